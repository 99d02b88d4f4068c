#!/bin/bash
# Round-end style validation on all 8 GPUs: collectives numerics + bandwidth, stress, DP bench, TP forward.
NP=${1:-8}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29511 tools/gpu_check_comm.py > gpurun_out/comm$NP.log 2>&1; echo comm_rc=$?
grep -E "FAIL|PERF|FUSED|COMM|Error|error|Traceback|timeout" gpurun_out/comm$NP.log | cut -c1-400 | head -24
cp gpurun_out/comm_check_rank0.json gpurun_out/comm_check_${NP}gpu.json
( cd collectives && RANKS="$NP" DLBB_OUTPUT_DIR=$PWD/../gpurun_out/stress timeout 300 ./launch_stress.sh 2>&1 | tail -8 )
( cd collectives && DLBB_SWEEP_BACKEND=nccl RANKS="$NP" DLBB_OUTPUT_DIR=$PWD/../gpurun_out/stress timeout 300 ./launch_stress.sh 2>&1 | tail -8 )
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $NP --steps 8 --warmup 3 > gpurun_out/bench$NP.log 2>&1; echo bench_rc=$?; tail -1 gpurun_out/bench$NP.log | cut -c1-260
bash tools/gpu_tp_forward.sh $NP 2>&1 | tail -6
