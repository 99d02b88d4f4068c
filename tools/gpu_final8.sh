#!/bin/bash
# One 8-GPU call, trimmed to the GPU budget: the sweeps through the entry points (both arms, 10 timed iterations), the
# strided sweep, the step breakdown of the DP step, a reduced tuning study, the TP forward and the fusion study.
NP=${1:-8}
ROOT=$(cd "$(dirname "$0")/.." && pwd); cd "$ROOT"
export PATH="$ROOT/bin:$PATH"
OUT=gpurun_out/publish_np$NP; mkdir -p $OUT/tuning
timeout 150 python -m pytest tests/test_gpu_distributed.py -x -q -k "dp_training" 2>&1 | tail -3
export DLBB_SWEEP_ITERS=10 DLBB_SWEEP_WARMUP=3
BATCH_3D=1,8,32 SEQ_3D=1,2048 MAX_MB_3D=600 BATCH_STRIDED=1,8 SEQ_STRIDED=1,2048 bash baseline/run_sweeps.sh $NP $OUT/sweeps > $OUT/run_sweeps.log 2>&1
grep -E "rc=" $OUT/run_sweeps.log
timeout 200 b200run -np $NP python tools/gpu_step_breakdown.py 2>&1 | grep -E "BREAKDOWN|rror" | cut -c1-600
study() { local name=$1; shift
  ( cd collectives/3d && env "$@" DLBB_SWEEP_OPS=allreduce DLBB_SWEEP_BATCH=8,16 DLBB_SWEEP_SEQ=2048,4096 DLBB_SWEEP_HIDDEN=4096 DLBB_SWEEP_MAX_MB=600 \
      DLBB_OUTPUT_DIR="$ROOT/$OUT/tuning/$name" timeout 120 deepspeed --num_gpus $NP --master_port 29711 dsccl.py > "$ROOT/$OUT/tuning_$name.log" 2>&1; echo "tuning $name rc=$?" ); }
study dlbb_directnvls_allreduce DLBB_ALGO=direct
study dlbb_direct_allreduce DLBB_ALGO=direct DLBB_MC_ALLREDUCE=0
study dlbb_nvls_allreduce DLBB_ALGO=nvls
study dlbb_twoshot_allreduce DLBB_ALGO=twoshot
study dlbb_ctas24 DLBB_CTAS=24
DLBB_TP_FUSED=1 DLBB_WORLD_SIZE=$NP DLBB_OUTPUT_DIR=$OUT/run_mpi timeout 200 mpirun -np $NP python run_mpi.py --config config/baseline_config.yaml --backend openmpi > $OUT/run_mpi_np$NP.log 2>&1
echo "run_mpi rc=$?"; grep -E "Forward pass mean|Forward \(device\)" $OUT/run_mpi_np$NP.log | head -3
rm -f gpurun_out/fusion_study.jsonl
for f in 1 0; do DLBB_FUSION=$f timeout 100 mpirun -np $NP python tools/gpu_fusion_study.py > $OUT/fusion_$f.log 2>&1; echo "fusion $f rc=$?"; done
cp gpurun_out/fusion_study.jsonl $OUT/ 2>/dev/null
DLBB_SKIP_PERF=1 SYMM_MAX_BYTES=67108864 timeout 120 mpirun -np $NP python tools/gpu_symm.py 2>&1 | grep -E "CHECK|FAIL" | head -5
