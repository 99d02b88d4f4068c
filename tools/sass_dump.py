#!/usr/bin/env python3
"""Full SASS of ONE instantiation of every named kernel of libdlbb.so into profiles/sass/<kernel>.sass, plus the
mnemonic counts of all kernels (tools/sass_summary.py) into profiles/sass/mnemonic_counts.txt.

multimem.ld_reduce lowers to LDGMC.E.<op>.<type>.STRONG.SYS.  multimem.st has NO mnemonic of its own: ptxas emits the
same STG.E.128.STRONG.SYS for `multimem.st.relaxed.sys.global.v4.f32` as for `st.relaxed.sys.global.v4.f32` (checked
with a two-instruction kernel) - the multicast is a property of the address (the multicast mapping of the segment).
Plain stores of these kernels are weak (`STG.E.128`), so in bcast_relay_kernel<true>, direct_reduce_kernel<.., SINK_MC>,
fused_reduce_adam_kernel<true> and gemm_allreduce_kernel every STG.E.128.STRONG.SYS is a multimem.st site.

Listings above 12000 lines (kernels whose unaligned / strided fall-back paths are unrolled byte loops) are stored
condensed: only memory, synchronisation, multimem and tensor instructions, with a header that says so."""
import re, subprocess, sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
lib = ROOT / "distributed_llm_backend_benchmark_b200" / "lib" / "libdlbb.so"
out_dir = ROOT / "profiles" / "sass"
out_dir.mkdir(parents=True, exist_ok=True)
sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
blocks, cur, name = {}, [], None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        if name:
            blocks.setdefault(name, cur)
        name, cur = m.group(1), []
    if name:
        cur.append(line)
if name:
    blocks.setdefault(name, cur)
dem = subprocess.run(["c++filt"], input="\n".join(blocks), capture_output=True, text=True).stdout.splitlines()
# preferred instantiation of each kernel: the bf16 / SUM / multicast ones the model and the sweeps run
WANT = {
    "allreduce_direct_kernel": r"allreduce_direct_kernel<__nv_bfloat16, 0>",
    "direct_reduce_kernel_nvls_allreduce": r"direct_reduce_kernel<__nv_bfloat16, 0, true, 2, 8>",
    "direct_reduce_kernel_p2p_reduce_scatter": r"direct_reduce_kernel<__nv_bfloat16, 0, false, 0, 8>",
    "reduce_pull_kernel": r"reduce_pull_kernel<__nv_bfloat16, 0, true, 8>",
    "bcast_relay_kernel_p2p": r"bcast_relay_kernel<false>",
    "bcast_relay_kernel_mc": r"bcast_relay_kernel<true>",
    "bcast_push_kernel": r"bcast_push_kernel\(",
    "allreduce_pipelined_kernel": r"allreduce_pipelined_kernel<__nv_bfloat16, 0, true>",
    "allreduce_twoshot_kernel_nvls": r"allreduce_twoshot_kernel<__nv_bfloat16, 0, true>",
    "reduce_kernel": r"dlbb::reduce_kernel<__nv_bfloat16, 0>",
    "reduce_multi_kernel": r"reduce_multi_kernel<__nv_bfloat16, 0>",
    "move_kernel": r"move_kernel\(",
    "broadcast_nvls_kernel": r"broadcast_nvls_kernel\(",
    "chain_broadcast_kernel": r"chain_broadcast_kernel\(",
    "chain_reduce_kernel": r"chain_reduce_kernel<__nv_bfloat16, 0>",
    "p2p_kernel": r"p2p_kernel\(",
    "barrier_kernel": r"barrier_kernel\(",
    "fused_reduce_adam_kernel_nvls": r"fused_reduce_adam_kernel<true>",
    "fused_reduce_adam_kernel_p2p": r"fused_reduce_adam_kernel<false>",
    "gemm_allreduce_kernel": r"gemm_allreduce_kernel<256, false, true>",
    "adam_bf16_kernel": r"adam_bf16_kernel\(",
    "ln_fwd_vec_kernel": r"ln_fwd_vec_kernel<2>",
    "ln_bwd_vec_kernel": r"ln_bwd_vec_kernel<2>",
    "ln_bwd_reduce_kernel": r"ln_bwd_reduce_kernel\(",
    "mse_kernel": r"mse_kernel\(",
    "cast_kernel": r"cast_kernel<float, __nv_bfloat16>",
    "gemm_bf16_tcgen05_pair_kernel_kmajor_mnmajor": r"gemm_bf16_tcgen05_pair_lowreg_kernel<false, true>",
    "gemm_bf16_tcgen05_kernel_256_kmajor_mnmajor": r"gemm_bf16_tcgen05_kernel<256, false, true>",
}
found = {}
for (mangled, body), d in zip(blocks.items(), dem):
    for key, pat in WANT.items():
        if key not in found and re.search(pat, d):
            found[key] = (d, body)
KEEP = re.compile(r"Function|\b(LD|ST|LDG|STG|LDS|STS|LDGMC|RED|ATOM|MEMBAR|FENCE|BAR|SYNCS|UTC|UTMA|LDTM|STTM|CCTL|ERRBAR|EXIT|BPT|TRAP|WARPSYNC)")
for key, (d, body) in sorted(found.items()):
    head = f"// {d}\n"
    if len(body) > 12000:
        head += f"// condensed listing: {len(body)} SASS lines, only memory / synchronisation / multimem / tensor instructions kept\n"
        body = [l for l in body if KEEP.search(l)]
    # drop the hex encodings and the padding: addresses, predicates, mnemonics and operands stay
    body = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in body]
    body = [re.sub(r"^\s+(/\*[0-9a-f]+\*/)\s+", r"\1 ", l).rstrip() for l in body if l.strip()]
    (out_dir / f"{key}.sass").write_text(head + "\n".join(body) + "\n")
    print(f"{key}.sass  {len(body)} lines  ({d[:90]})")
for key in WANT:
    if key not in found:
        print("MISSING", key, file=sys.stderr)
counts = subprocess.run([sys.executable, str(ROOT / "tools" / "sass_summary.py"), str(lib)], capture_output=True, text=True).stdout
(out_dir / "mnemonic_counts.txt").write_text(counts)
