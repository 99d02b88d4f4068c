#!/bin/bash
# 3-D tensor sweep on backend 'ccl' with explicit core binding and backend tuning knobs
# (ref collectives/3d/launch_dsccl.sh).  The `rank cores threads` table is parsed exactly as in the
# reference; the oneCCL knobs have dlbb analogues:
#
#   CCL_ALLREDUCE=<algo>          -> DLBB_ALGO=auto|oneshot|twoshot|nvls      all-reduce algorithm
#   CCL_WORKER_COUNT=<n>          -> DLBB_CTAS=<n>                            CTAs ("channels") per collective kernel
#   CCL_FUSION / *_THRESHOLD      -> DLBB_FUSION=0|1, DLBB_FUSION_BYTES=<n>   bucket fusion: all_reduce_coalesced / dlbb_allreduce_multi
#                                                                            pack up to 96 tensors into ONE launch (csrc/collectives: reduce_multi_kernel)
#   (message-size crossover)      -> DLBB_ONESHOT_MAX_BYTES=<n>
#   CCL_LOG_LEVEL=debug           -> DLBB_LOG_LEVEL=debug
# CCL_* variables that are set are translated below so old experiment scripts keep working.

cd "$(dirname "$0")"
export PATH="$PWD/../../bin:$PATH"

CONFIG_FILE=${CONFIG_FILE:-config_8.txt}
SCRIPT=${SCRIPT:-dsccl.py}
MASTER_PORT=${MASTER_PORT:-29502}

CORE_LIST=""
NUM_RANKS=0
while read -r line; do
    [[ "$line" =~ ^#.*$ || -z "$line" ]] && continue
    CORES=$(echo $line | awk '{print $2}')
    if [ -z "$CORES" ]; then
        echo "Invalid config line: $line"
        exit 1
    fi
    CORE_LIST="${CORE_LIST:+$CORE_LIST,}$CORES"
    NUM_RANKS=$((NUM_RANKS + 1))
done < "$CONFIG_FILE"

echo "Launching with $NUM_RANKS ranks"
echo "Core binding list: $CORE_LIST"

case "${CCL_ALLREDUCE:-}" in
    direct|nreduce)                 export DLBB_ALGO=${DLBB_ALGO:-oneshot} ;;
    ring|ring_rma|rabenseifner|2d)  export DLBB_ALGO=${DLBB_ALGO:-twoshot} ;;
    topo)                           export DLBB_ALGO=${DLBB_ALGO:-nvls} ;;
    double_tree|recursive_doubling) export DLBB_ALGO=${DLBB_ALGO:-twoshot} ;;
esac
[ -n "${CCL_WORKER_COUNT:-}" ] && export DLBB_CTAS=${DLBB_CTAS:-$((CCL_WORKER_COUNT * 24))}
[ -n "${CCL_FUSION:-}" ] && export DLBB_FUSION=${DLBB_FUSION:-$CCL_FUSION}
[ -n "${CCL_FUSION_BYTES_THRESHOLD:-}" ] && export DLBB_FUSION_BYTES=${DLBB_FUSION_BYTES:-$CCL_FUSION_BYTES_THRESHOLD}
[ "${CCL_LOG_LEVEL:-}" = "debug" ] && export DLBB_LOG_LEVEL=${DLBB_LOG_LEVEL:-debug}

mkdir -p logs
deepspeed \
    --num_gpus $NUM_RANKS \
    --bind_cores_to_rank \
    --bind_core_list $CORE_LIST \
    --master_port $MASTER_PORT \
    $SCRIPT > logs/${LOG_NAME:-dsccl_${NUM_RANKS}ranks}.log
echo "exit code $? - log in logs/${LOG_NAME:-dsccl_${NUM_RANKS}ranks}.log"
