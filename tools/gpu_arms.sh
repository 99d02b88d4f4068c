#!/bin/bash
# The 2 x 2 context table at N GPUs: ours with the sparse / dense QKV backward, stock PyTorch (cuBLAS + autograd + DDP over
# NCCL) with bf16 / fp32 optimizer state.  One bench.py JSON line each under gpurun_out/arms_nN_*.json.
N=${1:-1}; STEPS=${2:-6}
cd "$(dirname "$0")/.." && mkdir -p gpurun_out
run() { # name, env..., -- bench args
  local name=$1; shift
  if [ "$N" = 1 ]; then env "$@" python bench.py --gpus 1 --steps $STEPS --warmup 3 --no-e2e --no-collectives $IMPL > gpurun_out/arms_n${N}_$name.json 2> gpurun_out/arms_n${N}_$name.err
  else env "$@" python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus $N --steps $STEPS --warmup 3 --no-e2e --no-collectives $IMPL > gpurun_out/arms_n${N}_$name.json 2> gpurun_out/arms_n${N}_$name.err; fi
  echo "$name rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/arms_n${N}_$name.json | head -1)"
}
IMPL="--impl ours";       run ours_sparse X=1
IMPL="--impl ours";       run ours_dense DLBB_DENSE_QKV_BWD=1
IMPL="--impl torch-nccl"; run torch_bf16state DLBB_TORCH_STATE=bf16
IMPL="--impl torch-nccl"; run torch_fp32state DLBB_TORCH_STATE=fp32
