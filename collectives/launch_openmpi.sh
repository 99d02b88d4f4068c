#!/bin/bash
# 1-D sweep, Open MPI flavour, at every rank count the node offers (ref collectives/launch_openmpi.sh).
# RANKS="2 4 8" overrides the list; `mpirun` is the dlbb launcher (bin/mpirun).
cd "$(dirname "$0")/1d"
export PATH="$PWD/../../bin:$PATH"
echo "Running the OpenMPI-flavour 1-D sweep..."
export OMP_NUM_THREADS=2
for NP in ${RANKS:-2 4 8}; do
    mpirun -np $NP --report-bindings --map-by slot:PE=2 \
        python openmpi.py || exit $?
done
