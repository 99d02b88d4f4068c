#!/bin/bash
N=${1:-4}
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/gpu_symm.py; }
export SYMM_SKIP_NUMERICS=1 SYMM_MIN_BYTES=134217728 SYMM_OPS=reduce
for m in ${MODES:-0 1 2}; do DLBB_REDUCE_EXP=$m run > gpurun_out/exp2_$m.log 2>&1; echo "mode $m rc=$?"; grep -h PERF gpurun_out/exp2_$m.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l.split('PERF',1)[1]); print(d['op'],d['bytes']>>20,'MiB',{k:round(v) for k,v in d.items() if k.endswith('busbw')})"; done
