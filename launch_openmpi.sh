#!/bin/bash
# TP forward benchmark of models.py, Open MPI flavour of the command line (counterpart of the reference's
# launch_openmpi.sh).  bin/mpirun is the dlbb launcher: it takes Open MPI's flags, starts one rank per B200
# and binds each rank to host cores near its GPU.
#
#   WORLD_SIZE=8 ./launch_openmpi.sh        ranks = GPUs (default: the YAML's world_size, 4)
#   CORES_PER_RANK=8 ./launch_openmpi.sh    host cores per rank (default 14, as in the YAML)
set -u
here="$(cd "$(dirname "$0")" && pwd)"
cd "$here"
PATH="$here/bin:$PATH"

ranks="${WORLD_SIZE:-4}"
cores="${CORES_PER_RANK:-14}"
yaml="config/baseline_config.yaml"

# host threading for the little CPU work that is left (data generation, statistics)
export OMP_NUM_THREADS="$cores" MKL_NUM_THREADS="$cores" OMP_NESTED=FALSE
export DLBB_WORLD_SIZE="$ranks"

echo "run_mpi.py (backend label: openmpi) on $ranks rank(s), $cores core(s) each"
mpirun -np "$ranks" --bind-to core --map-by "socket:PE=$cores" \
       -x OMP_NUM_THREADS -x MKL_NUM_THREADS -x DLBB_WORLD_SIZE \
       python run_mpi.py --config "$yaml" --backend openmpi
rc=$?
[ $rc -eq 0 ] && echo "done: results under the YAML's output_dir" || echo "run failed with exit code $rc"
exit $rc
