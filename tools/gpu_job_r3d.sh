#!/bin/bash
# round 3, fourth call: (1) properly powered upper bound for fusing the norms: HEAD vs all operand-feeding LN/GN launches skipped,
# 10 interleaved runs each; (2) launch list + ncu --set full of the non-GEMM kernels (attention, istft, gn_*, layernorm)
mkdir -p gpurun_out/r3d
python tools/ab_interleaved.py 10 gpurun_out/r3d/ab_norms.json head head_skip_norms > gpurun_out/r3d/ab_norms.log 2>&1; echo "ab exit $?"; tail -3 gpurun_out/r3d/ab_norms.log
bash tools/profile.sh r3d list attn istft gn ln > gpurun_out/r3d/profile.log 2>&1; tail -15 gpurun_out/r3d/profile.log
