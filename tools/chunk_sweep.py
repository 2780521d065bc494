"""A/B of the chunked tensor-core accumulation (gemm3.cu): chunk length vs C3 step time and GEMM accuracy, all variants
in ONE process with interleaved timing rounds (the board is power-capped, so separate runs drift by 2-3 %).
usage: python tools/chunk_sweep.py [rounds]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
import bench  # noqa: E402
from miocodec_b200 import binding, synth  # noqa: E402

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 6
gguf, _ = synth.ensure_assets()
wl = bench.make_workload("c3", 0, 1)
variants = [("off", None), ("prenet", 1), ("prenet", 2), ("prenet", 3), ("prenet", 4), ("prenet", 6), ("all", 2), ("all", 4)]
ctxs = []
for mode, kb in variants:
    os.environ["MIO_CHUNK_ACC"] = mode
    if kb:
        os.environ["MIO_CHUNK_KB"] = str(kb)
    c = binding.MioCodec(gguf, 0)
    c.batch_upload(wl["codes"], wl["cu"], wl["voice_table"], wl["voice_ids"])
    for _ in range(2):
        c.batch_run(); c.batch_sync()
    ctxs.append(c)
times = np.zeros((rounds, len(variants)))
for r in range(rounds):
    for i, c in enumerate(ctxs):
        t0 = time.perf_counter()
        for _ in range(3):
            c.batch_run()
        c.batch_sync()
        times[r, i] = (time.perf_counter() - t0) / 3 * 1e3
rng = np.random.default_rng(0)
print(f"C3 step time (ms), median of {rounds} interleaved rounds of 3 steps; GEMM rel. rms error vs fp64 (M=512, N=256)")
for i, ((mode, kb), c) in enumerate(zip(variants, ctxs)):
    errs = []
    for K in (768, 2048):
        A = rng.standard_normal((512, K)).astype(np.float32)
        B = (rng.standard_normal((256, K)) / np.sqrt(K)).astype(np.float32)
        out = c.debug_gemm(A, B, 3).astype(np.float64)
        true = A.astype(np.float64) @ B.astype(np.float64).T
        errs.append(np.sqrt(np.mean((out - true) ** 2) / np.mean(true ** 2)))
    print(f"  chunk_acc={mode:7s} kb={kb}: {np.median(times[:, i]):7.2f} ms (min {times[:, i].min():.2f})   K=768 {errs[0]:.2e}  K=2048 {errs[1]:.2e}", flush=True)
    c.close()
