#!/bin/bash
# round 3, fifth call: chunk length of the chunked accumulation and the precise SiLU: step time (interleaved) and end-to-end accuracy
mkdir -p gpurun_out/r3e
python tools/ab_interleaved.py 8 gpurun_out/r3e/ab_chunk.json head head_kb4 head_kb6 head_chunk_acc_off head_gn_precise > gpurun_out/r3e/ab_chunk.log 2>&1; echo "ab exit $?"; tail -5 gpurun_out/r3e/ab_chunk.log
timeout 1200 python tests/accuracy_study.py e2e --n 32 > gpurun_out/r3e/e2e.log 2>&1; echo "e2e exit $?"; cat gpurun_out/r3e/e2e.log | tail -20
