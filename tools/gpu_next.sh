#!/bin/bash
# What to time first on the next box: the code paths that were built or changed after the
# last GPU run and are verified on the host emulation only.  One call, N GPUs (default 8), ~6 minutes.
#   gpurun --gpus 8 --timeout 900 -- 'bash tools/gpu_next.sh 8'
NP=${1:-8}
ROOT=$(cd "$(dirname "$0")/.." && pwd); cd "$ROOT"
export PATH="$ROOT/bin:$PATH"
OUT=gpurun_out/next_np$NP; mkdir -p $OUT
export DLBB_SWEEP_ITERS=20 DLBB_SWEEP_WARMUP=5
# 0. numerics first: everything below is meaningless if these fail
timeout 300 python -m pytest tests/test_gpu_distributed.py -x -q 2>&1 | tail -3
sweep1d() { local name=$1; shift   # 1-D sweep through the reference's entry point, one result directory per configuration
  ( cd collectives/1d && env "$@" DLBB_OUTPUT_DIR="$ROOT/$OUT/$name" timeout 150 deepspeed --num_gpus $NP --master_port 29721 dsccl.py \
      > "$ROOT/$OUT/$name.log" 2>&1; echo "$name rc=$?" ); }
# 1. small all-reduce: one-shot (all peers' loads in flight since reduce_wide) against the NVLS two-shot default
sweep1d ar_small_default DLBB_SWEEP_OPS=allreduce DLBB_SWEEP_SIZES=1KB,8KB,64KB,512KB DLBB_SWEEP_SYMMETRIC=0
sweep1d ar_small_oneshot64k DLBB_SWEEP_OPS=allreduce DLBB_SWEEP_SIZES=1KB,8KB,64KB,512KB DLBB_SWEEP_SYMMETRIC=0 DLBB_ONESHOT_MAX_BYTES=65536
# 2. broadcast: opening barrier (small sizes), pull relay against push relay (large sizes)
sweep1d bcast_pull DLBB_SWEEP_OPS=broadcast DLBB_SWEEP_SIZES=extended
sweep1d bcast_push DLBB_SWEEP_OPS=broadcast DLBB_SWEEP_SIZES=32MB,128MB,1GB DLBB_BCAST_PUSH=1 DLBB_BCAST_MC_MAX_BYTES=0
# 3. list all-to-all at 128 MB / 1 GB: staged (threshold out of reach) against the new zero-copy path
sweep1d a2a_list_zc DLBB_SWEEP_OPS=alltoall DLBB_SWEEP_SIZES=128MB,1GB
sweep1d a2a_list_staged DLBB_SWEEP_OPS=alltoall DLBB_SWEEP_SIZES=128MB,1GB DLBB_A2A_LIST_ZC_MIN_BYTES=1099511627776
# 4. bucket fusion of small tensors with the rewritten multi-tensor kernel
rm -f gpurun_out/fusion_study.jsonl
for f in 1 0; do DLBB_FUSION=$f timeout 100 mpirun -np $NP python tools/gpu_fusion_study.py > $OUT/fusion_$f.log 2>&1; echo "fusion $f rc=$?"; done
cp gpurun_out/fusion_study.jsonl $OUT/ 2>/dev/null
# 5. the bench line (DP training step + collective table)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus $NP --steps 10 --warmup 3 \
    > $OUT/bench_n$NP.json 2> $OUT/bench_n$NP.err; echo "bench rc=$?"; tail -c 600 $OUT/bench_n$NP.json
