#!/bin/bash
# only the NCCL arm (the dlbb arm results of this box are not comparable across boxes, so rerun a short dlbb arm too)
NP=$1; ROOT=$(pwd); OUT=$ROOT/gpurun_out/sweeps_np$NP; export PATH="$ROOT/bin:$PATH"; INC=$(seq -s, 0 $((NP-1)))
export DLBB_SWEEP_ITERS=20 DLBB_SWEEP_WARMUP=5
mkdir -p $OUT
for arm in dlbb nccl; do
  [ $arm = nccl ] && export DLBB_SWEEP_BACKEND=nccl || unset DLBB_SWEEP_BACKEND
  ( cd collectives/1d && DLBB_SWEEP_SIZES=extended DLBB_SWEEP_OPS=allreduce,allgather,broadcast,reducescatter,alltoall,reduce DLBB_OUTPUT_DIR=$OUT/1d_$arm timeout 400 deepspeed --master_port 29701 --include localhost:$INC dsccl.py > $OUT/1d_$arm.log 2>&1; echo "1d $arm rc=$?" )
  ( cd collectives/3d && DLBB_SWEEP_BATCH=1,8 DLBB_SWEEP_SEQ=1,2048 DLBB_SWEEP_HIDDEN=4096 DLBB_SWEEP_OPS=allreduce,allgather,broadcast,reduce,gather DLBB_OUTPUT_DIR=$OUT/3d_$arm timeout 300 deepspeed --num_gpus $NP --master_port 29702 dsccl.py > $OUT/3d_$arm.log 2>&1; echo "3d $arm rc=$?" )
done
python baseline/compare.py $OUT > $OUT/compare.md; grep -E "^\| (allreduce|allgather|broadcast|reducescatter|alltoall|reduce) " $OUT/compare.md | head -80
