#!/bin/bash
# reduce / broadcast exploration on N GPUs: CTA sweep of the direct + relay kernels, then the staged chains forced
N=${1:-4}
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
run() { timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/gpu_symm.py; }
export SYMM_SKIP_NUMERICS=1 SYMM_MIN_BYTES=33554432 SYMM_OPS=reduce,broadcast
SYMM_CTAS=0,64,192,256,300 run > gpurun_out/exp1_ctas_$N.log 2>&1; echo rc=$?
DLBB_RELAY_CHUNK_KB=512 SYMM_CTAS=0,300 run > gpurun_out/exp1_chunk512_$N.log 2>&1; echo rc=$?
DLBB_BCAST_RELAY=0 DLBB_REDUCE_DIRECT=0 DLBB_CHAIN_MIN_BYTES=1 SYMM_CTAS=0,300 run > gpurun_out/exp1_chain_$N.log 2>&1; echo rc=$?
grep -h PERF gpurun_out/exp1_ctas_$N.log | cut -c1-1500
echo ---- chunk512; grep -h PERF gpurun_out/exp1_chunk512_$N.log | cut -c1-900
echo ---- chain; grep -h PERF gpurun_out/exp1_chain_$N.log | cut -c1-900
