"""The collective kernels and their launchers (csrc/collectives/collectives.cu, direct_kernels.cuh) compiled for the host
- plus fused_adam.cu and the functor of the fused GEMM + all-reduce - and run without a GPU (csrc/collectives/emu/, tools/emu_build.py): every rank and every CUDA thread is a host thread, the
system-scope flags are release / acquire atomics, multicast windows are emulated.  Every collective entry point x dtype x
reduce op x size x strided view is checked against plain arithmetic at world sizes 2-8 (3, 5, 6, 7 exist on no box), each
call repeated back to back with the inputs rewritten in between, with thresholds lowered so that the large-message
kernels (pipelined all-reduce, chain, two-hop relay / pull, zero-copy direct) run on kilobytes.  One build runs under
ThreadSanitizer: a data access that the flag protocol does not order is reported as a race (SURVEY.md 5.2, section 4
"world_size in {1,2,4,8}, every op x dtype x reduce-op x size").  A negative control with weakened flags proves the
detector sees the protocol."""
import os
import re
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def _build(tmp_path_factory, name, *extra):
    out = tmp_path_factory.mktemp("emu") / name
    r = subprocess.run([sys.executable, str(ROOT / "tools" / "emu_build.py"), "--out", str(out), *extra], capture_output=True,
                       text=True, timeout=1200)
    assert r.returncode == 0, r.stderr[-3000:]
    return out


def _sanitizer_available(kind):
    probe = subprocess.run(["g++", f"-fsanitize={kind}", "-x", "c++", "-", "-o", "/dev/null"], input="int main(){}", text=True,
                           capture_output=True)
    return probe.returncode == 0


@pytest.fixture(scope="module")
def emu_builds(tmp_path_factory):
    """All variants of the emulation build, compiled side by side (one g++ each, about a minute)."""
    from concurrent.futures import ThreadPoolExecutor
    variants = {"plain": ()}
    if _sanitizer_available("thread"):
        variants["tsan"] = ("--sanitize", "thread")
        variants["weak"] = ("--sanitize", "thread", "--define", "EMU_WEAK_FLAGS")
    if _sanitizer_available("address"):
        variants["asan"] = ("--sanitize", "address")
    with ThreadPoolExecutor(max_workers=len(variants)) as pool:
        futs = {k: pool.submit(_build, tmp_path_factory, "emu_collectives_" + k, *extra) for k, extra in variants.items()}
        return {k: f.result() for k, f in futs.items()}


@pytest.fixture(scope="module")
def emu_plain(emu_builds):
    return emu_builds["plain"]


@pytest.fixture(scope="module")
def emu_tsan(emu_builds):
    if "tsan" not in emu_builds:
        pytest.skip("ThreadSanitizer runtime not available")
    return emu_builds["tsan"]


def _run(exe, world, mc, flt=None, env=None, timeout=1500):
    e = dict(os.environ)
    e.setdefault("EMU_REPS", "2")   # back-to-back repetitions per case (the program's own default is 4)
    e.update(env or {})
    cmd = [str(exe), str(world), "mc" if mc else "nomc"] + ([flt] if flt else [])
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=e)
    out = r.stdout + r.stderr
    assert r.returncode == 0 and "EMU COLLECTIVES OK" in out, out[-3000:]
    m = re.search(r"^kernels:(.*)$", out, re.M)
    kernels = dict((k, int(v)) for k, v in re.findall(r"(\S.*?)=(\d+)(?: |$)", m.group(1))) if m else {}
    return out, kernels


def _ran(kernels, prefix):
    return sum(n for k, n in kernels.items() if k.startswith(prefix))


@pytest.mark.parametrize("world,mc", [(1, False), (2, True), (3, True), (5, False), (7, False), (8, True)])   # (4 and 6: below)
def test_every_collective_at_every_world_size(emu_plain, world, mc):
    out, k = _run(emu_plain, world, mc, env={"EMU_REPS": "2", "EMU_MIXED_SEEDS": "2"})
    assert _ran(k, "fused_reduce_adam_kernel") and _ran(k, "emu_gemm_allreduce_kernel"), k
    assert _ran(k, "reduce_kernel") and _ran(k, "move_kernel"), k
    if world == 1:   # what the driver's one-GPU box runs: every collective degenerates to a local copy / reduction
        return
    assert _ran(k, "refresh_tile_flags_kernel") and _ran(k, "p2p_kernel") and _ran(k, "reduce_multi_kernel"), k
    assert _ran(k, "direct_reduce_kernel"), k                      # zero-copy in place
    if world > 2:
        assert _ran(k, "bcast_relay_kernel"), k
    if mc and world > 2:
        assert _ran(k, "allreduce_twoshot_kernel<T, OP, true>") and _ran(k, "broadcast_nvls_kernel"), k
        assert _ran(k, "direct_reduce_kernel<T, OP, true, SINK_MC"), k   # multimem.ld_reduce + multimem.st in place


@pytest.mark.parametrize("env,world,mc,flt,kernel", [
    ({"DLBB_PIPELINE_MIN_BYTES": "32768", "EMU_CTAS": "2", "EMU_STAGE_KB": "256"}, 3, True, "allreduce", "allreduce_pipelined_kernel"),
    ({"DLBB_CHAIN_MIN_BYTES": "16384"}, 4, True, "broadcast", "chain_broadcast_kernel"),
    ({"DLBB_CHAIN_MIN_BYTES": "16384"}, 3, False, "reduce ", "chain_reduce_kernel"),
    ({"DLBB_REDUCE_PULL": "1", "DLBB_REDUCE_PULL_MIN_BYTES": "8192", "DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "13"}, 7, True,
     "reduce ", "reduce_pull_kernel"),
    ({"DLBB_REDUCE_PULL": "1", "DLBB_REDUCE_PULL_MIN_BYTES": "8192", "DLBB_RELAY_CHUNK_KB": "1"}, 5, False, "reduce ", "reduce_pull_kernel"),
    ({"DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "13", "DLBB_BCAST_MC_MAX_BYTES": "0"}, 7, True, "broadcast", "bcast_relay_kernel<false>"),
    ({"DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "10"}, 6, False, "broadcast", "bcast_relay_kernel<false>"),
    # push form of the relay broadcast (opt-in), on a device that holds only P-1 of its 12 CTAs at a time
    ({"DLBB_BCAST_PUSH": "1", "DLBB_BCAST_MC_MAX_BYTES": "0", "DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "13", "EMU_RESIDENT_CTAS": "3"}, 4, False,
     "broadcast", "bcast_push_kernel"),
    ({"DLBB_BCAST_PUSH": "1", "DLBB_BCAST_MC_MAX_BYTES": "0", "DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "13"}, 7, True, "broadcast", "bcast_push_kernel"),
    # two CTA slots: the two-hop pull kernels only ever wait for lower-numbered CTAs, the pipelined all-reduce is clamped
    ({"DLBB_REDUCE_PULL": "1", "DLBB_REDUCE_PULL_MIN_BYTES": "8192", "DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "12", "EMU_RESIDENT_CTAS": "2",
      "EMU_REPS": "2"}, 4, True, "reduce ", "reduce_pull_kernel"),
    ({"DLBB_PIPELINE_MIN_BYTES": "32768", "EMU_CTAS": "4", "EMU_STAGE_KB": "256", "EMU_RESIDENT_CTAS": "2", "EMU_REPS": "2"}, 3, True,
     "allreduce", "allreduce_pipelined_kernel"),
    ({"DLBB_A2A_LIST_ZC_MIN_BYTES": "4096"}, 5, False, "ptrs", "move_kernel"),    # list all-to-all on separate tensors, zero-copy
    ({"DLBB_DIRECT_LEGACY": "1"}, 5, False, "allreduce", "allreduce_direct_kernel"),
    ({"DLBB_MC_REDUCE_SCATTER": "1"}, 3, True, "reduce_scatter", "direct_reduce_kernel<T, OP, true, SINK_ONE"),
    ({"DLBB_MC_WORLD2": "1"}, 2, True, "allreduce", "direct_reduce_kernel<T, OP, true, SINK_MC"),
    ({"DLBB_FUSION": "0"}, 3, True, "multi", "reduce_kernel"),
    ({"EMU_CTAS": "1", "EMU_STAGE_KB": "16", "EMU_ZC_MIN": "0", "EMU_REPS": "2"}, 4, True, None, "allreduce_twoshot_kernel"),   # everything staged, many slices
    ({"DLBB_DEBUG_POISON": "1", "EMU_REPS": "2"}, 4, True, None, "reduce_kernel"),                          # 0xFF-filled staging
    # DLBB_CTAS unset: the launchers' own grid choices (up to 294 CTAs for the two-hop kernels)
    ({"EMU_CTAS": "0"}, 3, True, "broadcast", "bcast_relay_kernel"),
    ({"EMU_CTAS": "0"}, 3, True, "allreduce", "direct_reduce_kernel"),
    # a device with 3 CTA slots and 7 CTAs asked for: grid-barrier kernels are clamped, the others run in waves
    ({"EMU_CTAS": "7", "EMU_RESIDENT_CTAS": "3", "EMU_REPS": "2", "EMU_MIXED_SEEDS": "2"}, 4, True, None, "move_kernel"),
    # the periodic rewrite of all flags of a bank (every 2^28 epochs in production, every 300 here), across the wrap
    ({"DLBB_FLAG_REFRESH_EVERY": "300", "EMU_EPOCH0": "0xFFFFFF00", "EMU_REPS": "2", "EMU_MIXED_SEEDS": "2"}, 3, True, None, "barrier_kernel"),
    ({"EMU_EPOCH0": "0xFFFFFFC0", "EMU_REPS": "3"}, 3, True, None, "reduce_kernel"),                       # every epoch counter wraps past 2^32
])
def test_threshold_selected_and_opt_in_kernels(emu_plain, env, world, mc, flt, kernel):
    out, k = _run(emu_plain, world, mc, flt, env)
    assert _ran(k, kernel), (kernel, k)


@pytest.mark.parametrize("env,world,mc,flt", [
    ({}, 3, True, "mixed"),            # different collectives back to back: what one kernel leaves behind, the next one meets
    ({}, 4, True, "mixed"),
    ({}, 4, False, "mixed"),
    ({"EMU_REPS": "3"}, 3, True, "views"),
    ({"EMU_REPS": "3"}, 3, False, "sendrecv"),
    ({"EMU_REPS": "3", "DLBB_A2A_LIST_ZC_MIN_BYTES": "4096"}, 3, False, "ptrs"),   # list collectives, zero-copy on separate tensors
    ({"EMU_REPS": "3"}, 4, True, "multi"),                                         # bucket fusion
    ({"EMU_REPS": "3"}, 3, True, "fused_adam"),
    ({"EMU_REPS": "2"}, 3, True, "gemm_allreduce"),     # reducer warps + per-tile flags of the fused GEMM + all-reduce
    ({"EMU_REPS": "2", "DLBB_REDUCE_PULL": "1", "DLBB_REDUCE_PULL_MIN_BYTES": "8192", "DLBB_RELAY_CHUNK_KB": "1"}, 4, True, "reduce "),
    ({"EMU_REPS": "2", "DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "6", "DLBB_BCAST_MC_MAX_BYTES": "0"}, 4, True, "broadcast"),
    # nothing zero-copy: the staged large-message kernels (pipelined, chain, multicast broadcast) share flag bank 4
    ({"EMU_ZC_MIN": "0", "DLBB_PIPELINE_MIN_BYTES": "16384", "DLBB_CHAIN_MIN_BYTES": "16384", "EMU_CTAS": "2"}, 3, True, "mixed"),
    ({"EMU_REPS": "3", "DLBB_BCAST_PUSH": "1", "DLBB_BCAST_MC_MAX_BYTES": "0", "DLBB_RELAY_CHUNK_KB": "1", "EMU_CTAS": "6"}, 4, True, "broadcast"),
    ({"DLBB_DEBUG_POISON": "1", "EMU_MIXED_SEEDS": "3"}, 3, False, "mixed"),   # the debug fill itself must not race with a slower reader
    ({"EMU_REPS": "2", "DLBB_PIPELINE_MIN_BYTES": "32768", "EMU_CTAS": "1", "EMU_STAGE_KB": "256"}, 2, True, "allreduce"),
])
def test_flag_protocols_are_race_free_under_thread_sanitizer(emu_tsan, env, world, mc, flt):
    out, _ = _run(emu_tsan, world, mc, flt, env, timeout=3000)
    assert "ThreadSanitizer" not in out, out[-4000:]


def test_device_watchdog_names_the_missing_rank(emu_plain):
    """a rank that never arrives ends the others' flag waits after the deadline, with its number in the message."""
    r = subprocess.run([str(emu_plain), "3", "nomc", "watchdog"], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, DLBB_WATCHDOG_SECONDS="2"))
    assert r.returncode != 0 and "watchdog did not fire" not in r.stdout
    assert re.search(r"device barrier timeout: block \d+ waited 2000 ms for rank 2", r.stderr), r.stderr[-2000:]


def test_grid_barrier_over_more_ctas_than_fit_is_detected(emu_plain):
    """Negative control for the residency model: with the library's clamp disabled, a grid-barrier kernel of 7 CTAs on a
    device with 3 slots stalls (CTA 0 waits for CTAs that cannot start) and the watchdog fires."""
    r = subprocess.run([str(emu_plain), "2", "nomc", "allgather"], capture_output=True, text=True, timeout=120,
                       env=dict(os.environ, DLBB_WATCHDOG_SECONDS="2", EMU_NO_CLAMP="1", EMU_CTAS="7", EMU_RESIDENT_CTAS="3", EMU_REPS="1"))
    assert r.returncode != 0 and "device barrier timeout" in r.stderr, r.stderr[-2000:]


def test_no_out_of_bounds_access_under_address_sanitizer(emu_builds):
    """Tensors are allocated to the byte: a kernel that rounds a tail up to a whole vector, or view arithmetic that leaves the
    tensor, is an AddressSanitizer report."""
    if "asan" not in emu_builds:
        pytest.skip("AddressSanitizer runtime not available")
    out, _ = _run(emu_builds["asan"], 3, True, env={"EMU_REPS": "1", "EMU_MIXED_SEEDS": "1", "ASAN_OPTIONS": "detect_leaks=0"})
    assert "AddressSanitizer" not in out, out[-4000:]


def test_thread_sanitizer_sees_the_protocol(emu_builds):
    """Negative control: with the flags demoted to relaxed atomics the same run must report data races."""
    if "weak" not in emu_builds:
        pytest.skip("ThreadSanitizer runtime not available")
    r = subprocess.run([str(emu_builds["weak"]), "2", "mc", "allreduce"], capture_output=True, text=True, timeout=1500,
                       env=dict(os.environ, EMU_REPS="2"))
    assert "WARNING: ThreadSanitizer: data race" in r.stderr
