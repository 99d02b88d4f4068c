#!/bin/bash
# 7B tensor-parallel forward (run_mpi.py, the reference's headline model run) at the given rank count:
# fused GEMM+all-reduce kernel vs GEMM followed by the all-reduce kernel.
NP=${1:-2}
mkdir -p gpurun_out
for fused in 1 0; do
  DLBB_TP_FUSED=$fused DLBB_WORLD_SIZE=$NP DLBB_OUTPUT_DIR=gpurun_out/run_mpi_fused$fused timeout 300 bin/mpirun -np $NP python run_mpi.py --config config/baseline_config.yaml --backend openmpi > gpurun_out/run_mpi_np${NP}_fused$fused.log 2>&1
  echo "np=$NP fused=$fused rc=$?"; grep -E "Forward pass mean|Forward \(device\)|ERROR|Error" gpurun_out/run_mpi_np${NP}_fused$fused.log
done
