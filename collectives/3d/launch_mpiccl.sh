#!/bin/bash
# MPI-launched torch 'ccl' 3-D sweep (ref collectives/3d/launch_mpiccl.sh, whose target script
# mpitorchccl.py is missing from the reference; this repository provides it).
cd "$(dirname "$0")"
export PATH="$PWD/../../bin:$PATH"
RANKS=${RANKS:-8}
export MASTER_PORT=${MASTER_PORT:-29501}
# Intel MPI pinning variables are accepted and ignored by the dlbb launcher
export I_MPI_PIN=1
export I_MPI_PIN_DOMAIN=1
export I_MPI_PIN_ORDER=compact
export DLBB_LOG_LEVEL=${DLBB_LOG_LEVEL:-info}
export DLBB_CTAS=${DLBB_CTAS:-0}
echo "Launching MPI + dlbb('ccl') 3D Allreduce Benchmark..."
echo "Ranks: $RANKS"
mpirun -n ${RANKS} python mpitorchccl.py
