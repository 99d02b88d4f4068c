#!/bin/bash
# Tensor-parallel forward benchmark, Open MPI spelling of the launch line (ref launch_openmpi.sh).
# `mpirun` here is bin/mpirun -> the dlbb launcher: same flags, one rank per B200, NUMA-aware core
# binding.  WORLD_SIZE defaults to the YAML's 4; set it (1/2/4/8) to run on another GPU count.

cd "$(dirname "$0")"
export PATH="$PWD/bin:$PATH"

CONFIG_FILE="config/baseline_config.yaml"
WORLD_SIZE=${WORLD_SIZE:-4}
CORES_PER_RANK=${CORES_PER_RANK:-14}

export OMP_NUM_THREADS=$CORES_PER_RANK
export MKL_NUM_THREADS=$CORES_PER_RANK
export OMP_NESTED=FALSE
export DLBB_WORLD_SIZE=$WORLD_SIZE

echo "Launching the OpenMPI-style run on $WORLD_SIZE rank(s)"
mpirun -np $WORLD_SIZE \
       --bind-to core \
       --map-by socket:PE=$CORES_PER_RANK \
       -x OMP_NUM_THREADS \
       -x MKL_NUM_THREADS \
       -x DLBB_WORLD_SIZE \
       python run_mpi.py \
       --config $CONFIG_FILE \
       --backend openmpi || exit $?

echo ""
echo "OpenMPI experiment completed!"
