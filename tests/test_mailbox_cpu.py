"""Host-side rendezvous of the zero-copy collectives (csrc/bootstrap/mailbox.h) without a GPU: forked ranks
connect over the bootstrap sockets, map each other's mailbox regions and run thousands of exchanges with
randomly delayed ranks; every record must carry the sender's rank and the current exchange number."""
import os
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def mailbox_test_binary(tmp_path_factory):
    out = tmp_path_factory.mktemp("mailbox") / "mailbox_test"
    src = [ROOT / "csrc" / "bootstrap" / "tests" / "mailbox_test.cpp", ROOT / "csrc" / "bootstrap" / "bootstrap.cpp"]
    subprocess.run(["g++", "-O2", "-std=c++17", "-Wall", "-o", str(out), *map(str, src), "-lpthread"], check=True)
    return out


@pytest.mark.parametrize("world,exchanges", [(2, 20000), (4, 10000), (8, 3000)])
def test_exchanges_stay_in_lockstep(mailbox_test_binary, world, exchanges):
    r = subprocess.run([str(mailbox_test_binary), str(world), str(exchanges), f"mbx-{os.getpid()}-{world}"],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and "MAILBOX OK" in r.stdout, r.stdout + r.stderr
