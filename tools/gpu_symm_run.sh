#!/bin/bash
# numerics of the symmetric / direct collectives + bandwidth of selected ops on N GPUs (compact output of tools/gpu_symm.py)
N=${1:-4}
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
SYMM_MIN_BYTES=${MINB:-33554432} SYMM_OPS=${OPS:-reduce,broadcast} timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/gpu_symm.py > gpurun_out/symm_run_$N.log 2>&1; echo rc=$?
grep -E "FAIL|CHECK|rror|timeout" gpurun_out/symm_run_$N.log | head -20
grep -h PERF gpurun_out/symm_run_$N.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l.split('PERF',1)[1]); print(d['op'],d['bytes']>>20,'MiB',{k:round(v) for k,v in d.items() if k.endswith('busbw')})"
