#!/bin/bash
# Tensor-parallel forward benchmark, Intel MPI spelling of the launch line (ref launch_intelmpi.sh).
# The I_MPI_* variables are kept for script compatibility; the dlbb launcher ignores them (there
# is no shared-memory fabric to configure: ranks talk over NVLink).

cd "$(dirname "$0")"
export PATH="$PWD/bin:$PATH"

CONFIG_FILE="config/baseline_config.yaml"
WORLD_SIZE=${WORLD_SIZE:-4}
CORES_PER_RANK=${CORES_PER_RANK:-14}

export OMP_NUM_THREADS=$CORES_PER_RANK
export MKL_NUM_THREADS=$CORES_PER_RANK
export OMP_NESTED=FALSE
export I_MPI_PIN_DOMAIN=omp
export I_MPI_FABRICS=shm
export I_MPI_SHM_LMT=shm

echo "Launching the IntelMPI-style run on $WORLD_SIZE rank(s)"
mpirun -np $WORLD_SIZE \
       -ppn $WORLD_SIZE \
       -genv OMP_NUM_THREADS=$CORES_PER_RANK \
       -genv MKL_NUM_THREADS=$CORES_PER_RANK \
       -genv DLBB_WORLD_SIZE=$WORLD_SIZE \
       python run_mpi.py \
       --config $CONFIG_FILE \
       --backend intelmpi || exit $?

echo ""
echo "IntelMPI experiment completed!"
