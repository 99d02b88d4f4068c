#!/bin/bash
# Round-end numbers on the 8-GPU box (charged 8x, so kept short): the DP bench at N=8 and the large-message
# collective bandwidth table (zero-copy / staged / NCCL), numerics skipped (validated on 2 and 4 GPUs).
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 \
  bench.py --gpus 8 --steps 8 --warmup 3 > gpurun_out/bench_n8_pair.json 2> gpurun_out/bench_n8_pair.err
echo "bench rc=$?"; cut -c1-300 gpurun_out/bench_n8_pair.json
ZC_SKIP_NUMERICS=1 timeout 200 python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 \
  tools/gpu_zero_copy.py > gpurun_out/zero_copy8.log 2>&1
echo "zero-copy rc=$?"; grep "PERF\|CHECK" gpurun_out/zero_copy8.log | cut -c1-330
