#!/bin/bash
# Publish the collective sweeps at NP ranks through the entry points (both arms), the strided 3-D sweep, the stats, and at
# NP = 8 the tuning study (all-reduce algorithm x CTA count x fusion) in the reference's result layout
# (ref collectives/3d/launch_dsccl.sh:37-59, collectives/3d/results/dsccl_*).  Everything lands under gpurun_out/publish_npN.
NP=${1:-2}
ROOT=$(cd "$(dirname "$0")/.." && pwd)
cd "$ROOT"
OUT=gpurun_out/publish_np$NP
mkdir -p $OUT
export DLBB_SWEEP_ITERS=${DLBB_SWEEP_ITERS:-30} DLBB_SWEEP_WARMUP=${DLBB_SWEEP_WARMUP:-5}
bash baseline/run_sweeps.sh $NP $OUT/sweeps > $OUT/run_sweeps.log 2>&1
tail -3 $OUT/run_sweeps.log
export PATH="$ROOT/bin:$PATH"
if [ "$NP" -ge 4 ]; then
  # tuning study on the reference's two shapes: (8,2048,4096) = 128 MiB and (16,4096,4096) = 512 MiB, all-reduce
  study() {  # name, env...
    local name=$1; shift
    ( cd collectives/3d && env "$@" DLBB_SWEEP_OPS=allreduce DLBB_SWEEP_BATCH=8,16 DLBB_SWEEP_SEQ=2048,4096 DLBB_SWEEP_HIDDEN=4096 \
        DLBB_SWEEP_MAX_MB=600 DLBB_OUTPUT_DIR="$ROOT/$OUT/tuning/$name" timeout 200 deepspeed --num_gpus $NP --master_port 29711 dsccl.py \
        > "$ROOT/$OUT/tuning_$name.log" 2>&1; echo "tuning $name rc=$?" )
  }
  mkdir -p $OUT/tuning
  study dlbb_oneshot_allreduce DLBB_ALGO=oneshot DLBB_ONESHOT_MAX_BYTES=4000000000
  study dlbb_twoshot_allreduce DLBB_ALGO=twoshot
  study dlbb_nvls_allreduce DLBB_ALGO=nvls
  study dlbb_direct_allreduce DLBB_ALGO=direct DLBB_MC_ALLREDUCE=0
  study dlbb_directnvls_allreduce DLBB_ALGO=direct
  for c in 24 48 96 128; do study dlbb_ctas$c DLBB_CTAS=$c; done
  study dlbb_fusion0 DLBB_FUSION=0
  study dlbb_fusion DLBB_FUSION=1
fi
ls $OUT/sweeps | head; ls $OUT/tuning 2>/dev/null | head -20
