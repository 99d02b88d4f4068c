"""Host-side rendezvous of the zero-copy collectives (csrc/bootstrap/mailbox.h) without a GPU: forked ranks
connect over the bootstrap sockets, map each other's mailbox regions and run thousands of exchanges with
randomly delayed ranks; every record must carry the sender's rank and the current exchange number."""
import os
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


SRC = [ROOT / "csrc" / "bootstrap" / "tests" / "mailbox_test.cpp", ROOT / "csrc" / "bootstrap" / "bootstrap.cpp"]


def _build(out, extra=()):
    return subprocess.run(["g++", "-O2", "-g", "-std=c++17", "-Wall", *extra, "-o", str(out), *map(str, SRC), "-lpthread"],
                          capture_output=True, text=True)


@pytest.fixture(scope="module")
def mailbox_test_binary(tmp_path_factory):
    out = tmp_path_factory.mktemp("mailbox") / "mailbox_test"
    r = _build(out)
    assert r.returncode == 0, r.stderr
    return out


@pytest.mark.parametrize("world,exchanges", [(2, 20000), (4, 10000), (8, 3000)])
def test_exchanges_stay_in_lockstep(mailbox_test_binary, world, exchanges):
    r = subprocess.run([str(mailbox_test_binary), str(world), str(exchanges), f"mbx-{os.getpid()}-{world}"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "MAILBOX OK" in r.stdout, r.stdout + r.stderr


@pytest.mark.parametrize("name,flags", [("asan_ubsan", ["-fsanitize=address,undefined", "-fno-sanitize-recover=all"]),
                                        ("tsan", ["-fsanitize=thread"])])
def test_bootstrap_and_mailbox_under_sanitizers(tmp_path, name, flags):
    """the host side of the communicator (socket rendezvous, fd passing, shared-memory mailbox) built
    with AddressSanitizer + UBSan and with ThreadSanitizer; any report makes the rank exit non-zero."""
    out = tmp_path / f"mailbox_{name}"
    r = _build(out, flags)
    if r.returncode != 0:
        pytest.skip(f"sanitizer runtime not available: {r.stderr[-300:]}")
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1:halt_on_error=1", TSAN_OPTIONS="halt_on_error=1 exitcode=66")
    r = subprocess.run([str(out), "4", "2000", f"mbx-{name}-{os.getpid()}"], capture_output=True, text=True, timeout=300,
                       env=env)
    assert r.returncode == 0 and "MAILBOX OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
