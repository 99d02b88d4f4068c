#!/usr/bin/env python3
"""Per-kernel count of the SASS mnemonics that prove the Blackwell paths (B200_PROFILING.md):
UTC*MMA = tcgen05.mma, LDTM = tcgen05.ld, UTMALDG/UTMASTG = TMA, LDGMC = multimem.ld_reduce, SYNCS = mbarrier.
multimem.st has no mnemonic of its own: it lowers to STG.E.128.STRONG.SYS on a multicast address (the same encoding as
st.relaxed.sys), counted here as STG.SYS; the kernels' ordinary stores are weak and count as STG."""
import collections, re, subprocess, sys
lib = sys.argv[1] if len(sys.argv) > 1 else "distributed_llm_backend_benchmark_b200/lib/libdlbb.so"
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()
pat = re.compile(r"\b(UTC[A-Z]*MMA|UTMALDG|UTMASTG|UTMAPF|LDTM|STTM|UTCBAR|UTCATOMSWS|SYNCS|UBLKCP|HMMA|LDG|STG|RED|ATOM|MULTIMEM|LDGMC|STGMC|REDGMC|LD\.E\.[A-Z0-9.]*STRONG\.SYS|ST\.E\.[A-Z0-9.]*STRONG\.SYS|CCTL)[A-Z0-9_.]*")
counts = collections.OrderedDict(); fn = None
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = m.group(1); counts[fn] = collections.Counter(); continue
    if fn is None: continue
    m = pat.search(line)
    if m:
        key = m.group(0)
        key = re.sub(r"^(LDG|STG)\..*", lambda mm: mm.group(1) + (".SYS" if "SYS" in key else ""), key)
        counts[fn][key.split(".")[0] if key.startswith(("SYNCS", "UTCATOMSWS", "CCTL")) else key] += 1
for fn, c in counts.items():
    name = re.sub(r"\(.*", "", demangle(fn))
    print(f"{name}\n    " + "  ".join(f"{k}:{v}" for k, v in sorted(c.items())))
