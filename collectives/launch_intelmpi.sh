#!/bin/bash
# 1-D sweep, Intel MPI flavour (ref collectives/launch_intelmpi.sh).  RANKS="2 4 8" overrides the list.
cd "$(dirname "$0")/1d"
export PATH="$PWD/../../bin:$PATH"
echo "Running the IntelMPI-flavour 1-D sweep..."
export OMP_NUM_THREADS=2
for NP in ${RANKS:-2 4 8}; do
    mpirun -np $NP --map-by slot:PE=2 \
        python intelmpi.py || exit $?
done
