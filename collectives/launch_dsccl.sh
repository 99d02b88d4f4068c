#!/bin/bash
# 1-D sweep through the DeepSpeed launcher on backend 'ccl' (ref collectives/launch_dsccl.sh).
# `deepspeed` is the dlbb launcher's DeepSpeed personality (bin/deepspeed); 'ccl' is the dlbb backend.
cd "$(dirname "$0")/1d"
export PATH="$PWD/../../bin:$PATH"
echo "Running the DeepSpeed/ccl 1-D sweep..."
export OMP_NUM_THREADS=4
deepspeed --master_port 29501 --include localhost:0,1 dsccl.py || exit $?
deepspeed --master_port 29501 --include localhost:0,1,2,3 dsccl.py || exit $?
deepspeed --master_port 29501 --include localhost:0,1,2,3,4,5,6,7 dsccl.py || exit $?
