#!/bin/bash
# All-to-all + mixed-collective stress at every rank count through the DeepSpeed-launcher entry point
# (BASELINE.json config 5).  RANKS="1 2 4 8" overrides the list.
cd "$(dirname "$0")"
export PATH="$PWD/../bin:$PATH"
for NP in ${RANKS:-1 2 4 8}; do
    INC=$(seq -s, 0 $((NP-1)))
    deepspeed --master_port 29503 --include localhost:$INC stress.py || exit $?
done
