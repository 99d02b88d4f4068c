#!/bin/bash
NP=${1:-2}
run() { env "$@" timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NP --master-addr 127.0.0.1 --master-port 29533 tools/gpu_probe_allreduce.py 2>&1 | grep -E "busbw|Error|error" ; }
run DLBB_PIPELINE_MIN_BYTES=99999999999 CTAS=64,96,128
run DLBB_PIPE_CHUNK_KB=64
run DLBB_PIPE_CHUNK_KB=256 CTAS=64,128
run DLBB_PIPE_CHUNK_KB=16 CTAS=64,128
run DLBB_PIPE_CHUNK_KB=64 STAGE_MB=256 CTAS=64,128
