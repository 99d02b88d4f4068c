#!/bin/bash
# TP forward benchmark of models.py, Intel MPI flavour of the command line (counterpart of the reference's
# launch_intelmpi.sh): `-ppn`, `-genv K=V` and the I_MPI_* pinning variables.  bin/mpirun is the dlbb launcher;
# it accepts the Hydra spelling and ignores I_MPI_FABRICS / I_MPI_SHM_LMT (ranks talk over NVLink, there is no
# shared-memory fabric to tune) - they are exported only so that scripts which read them keep working.
#
#   WORLD_SIZE=8 ./launch_intelmpi.sh
set -u
here="$(cd "$(dirname "$0")" && pwd)"
cd "$here"
PATH="$here/bin:$PATH"

ranks="${WORLD_SIZE:-4}"
cores="${CORES_PER_RANK:-14}"
yaml="config/baseline_config.yaml"

export OMP_NUM_THREADS="$cores" MKL_NUM_THREADS="$cores" OMP_NESTED=FALSE
export I_MPI_PIN_DOMAIN=omp I_MPI_FABRICS=shm I_MPI_SHM_LMT=shm

echo "run_mpi.py (backend label: intelmpi) on $ranks rank(s), $cores core(s) each"
mpirun -np "$ranks" -ppn "$ranks" \
       -genv OMP_NUM_THREADS="$cores" -genv MKL_NUM_THREADS="$cores" -genv DLBB_WORLD_SIZE="$ranks" \
       python run_mpi.py --config "$yaml" --backend intelmpi
rc=$?
[ $rc -eq 0 ] && echo "done: results under the YAML's output_dir" || echo "run failed with exit code $rc"
exit $rc
