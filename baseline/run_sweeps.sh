#!/bin/bash
# Usage: baseline/run_sweeps.sh NP [outdir]   - both arms of the collective sweeps on NP GPUs of this box.
NP=${1:-2}
OUT=$(realpath -m ${2:-gpurun_out/sweeps_np$NP})
ROOT=$(cd "$(dirname "$0")/.." && pwd)
export PATH="$ROOT/bin:$PATH"
INC=$(seq -s, 0 $((NP-1)))
export DLBB_SWEEP_ITERS=${DLBB_SWEEP_ITERS:-20} DLBB_SWEEP_WARMUP=${DLBB_SWEEP_WARMUP:-5}
mkdir -p "$OUT"
for arm in dlbb nccl; do
  [ $arm = nccl ] && export DLBB_SWEEP_BACKEND=nccl || unset DLBB_SWEEP_BACKEND
  ( cd "$ROOT/collectives/1d" && DLBB_SWEEP_SIZES=${SIZES_1D:-extended} DLBB_SWEEP_OPS=${OPS_1D:-allreduce,allgather,broadcast,reducescatter,alltoall,reduce} \
      DLBB_OUTPUT_DIR="$OUT/1d_$arm" timeout 600 deepspeed --master_port 29701 --include localhost:$INC dsccl.py > "$OUT/1d_$arm.log" 2>&1; echo "1d $arm rc=$?" )
  ( cd "$ROOT/collectives/3d" && DLBB_SWEEP_BATCH=${BATCH_3D:-1,8,32} DLBB_SWEEP_SEQ=${SEQ_3D:-1,2048,8192} DLBB_SWEEP_HIDDEN=4096 DLBB_SWEEP_OPS=${OPS_3D:-allreduce,allgather,broadcast,reduce,gather} \
      DLBB_SWEEP_MAX_MB=${MAX_MB_3D:-600} DLBB_OUTPUT_DIR="$OUT/3d_$arm" timeout 600 deepspeed --num_gpus $NP --master_port 29702 dsccl.py > "$OUT/3d_$arm.log" 2>&1; echo "3d $arm rc=$?" )
done
unset DLBB_SWEEP_BACKEND
( cd "$ROOT/collectives/3d" && DLBB_SWEEP_STRIDED=1 DLBB_SWEEP_BATCH=${BATCH_STRIDED:-1,8} DLBB_SWEEP_SEQ=${SEQ_STRIDED:-1,2048} DLBB_SWEEP_HIDDEN=4096 DLBB_SWEEP_OPS=${OPS_3D:-allreduce,allgather,broadcast,reduce,gather} \
    DLBB_OUTPUT_DIR="$OUT/3d_dlbb_strided" timeout 300 deepspeed --num_gpus $NP --master_port 29703 dsccl.py > "$OUT/3d_strided.log" 2>&1; echo "3d strided rc=$?" )
python "$ROOT/baseline/compare.py" "$OUT" | tee "$OUT/compare.md"
