#!/bin/bash
# Round-end validation on a 2-GPU box: the whole GPU test tier, the smoke entry, and the DP bench for ours and for
# stock PyTorch DDP over NCCL (bench.py --impl torch-nccl) at N=2.
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2
for impl in ours torch-nccl; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 \
    bench.py --gpus 2 --steps 8 --warmup 3 --impl $impl > gpurun_out/bench_n2_$impl.json 2> gpurun_out/bench_n2_$impl.err
  echo "bench $impl rc=$?"; cut -c1-260 gpurun_out/bench_n2_$impl.json
done
