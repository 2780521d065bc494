"""Where does the GPU path lose accuracy against the reference?  (attribute the ~1 dB.)

Runs on the GPU box.  Three parts, each printed as a table (copy under profiles/):
  gemm    bare GEMMs through the product tcgen05 kernel (mio_debug_gemm) against an fp64 product of the same split
          operands: rms error, signed bias toward zero, and two probes that tell round-to-nearest from truncation in the
          tensor core's fp32 accumulation
  stages  every tap of the decoder against the fp64 evaluation of the oracle port (oracle/port.py, dtype=float64),
          for the tensor-core path, the CUDA-core validation path, and the fp32 port itself (what fp32 CPU arithmetic
          reaches at the same tap)
  e2e     N utterances end to end against the live reference (oracle/_ref), one variant per process:
          tc | simt | attn_simt | gn_precise | all_precise; mean +- sd of the SNR and of max-abs

usage: python tools/accuracy_study.py gemm|stages|e2e|all [--n 32] [--variant tc]
"""
import argparse
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
OUT = os.path.join(ROOT, "gpurun_out")


def snr(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return 10 * np.log10(np.sum(b * b) / max(np.sum((a - b) ** 2), 1e-300))


def split16(x):
    hi = x.astype(np.float16)
    lo = (x - hi.astype(np.float32)).astype(np.float16)
    return hi.astype(np.float64), lo.astype(np.float64)


def part_gemm():
    from miocodec_b200 import binding, synth
    gguf, _ = synth.ensure_assets()
    rng = np.random.default_rng(0)
    print("== gemm: product kernel vs fp64 on the same operands (M=512, N=256; rel = rms(err)/rms(out); bias = mean(err*sign(out))/rms(out))")
    for mode in ("tc", "tc_chunked", "simt"):
        c = binding.MioCodec(gguf, 0, gemm=mode.split("_")[0])
        c.set_option("chunk_acc", "all" if mode == "tc_chunked" else "off")
        for K in (128, 512, 768, 2048):
            A = rng.standard_normal((512, K)).astype(np.float32)
            B = (rng.standard_normal((256, K)) / np.sqrt(K)).astype(np.float32)
            for npass in (3, 1):
                out = c.debug_gemm(A, B, npass).astype(np.float64)
                ah, al = split16(A); bh, bl = split16(B)
                if npass == 3:
                    kept = ah @ bh.T + ah @ bl.T + al @ bh.T          # what the three passes compute, exactly
                    full = (ah + al) @ (bh + bl).T
                    true = A.astype(np.float64) @ B.astype(np.float64).T
                else:
                    kept = full = true = ah @ bh.T
                rms = np.sqrt(np.mean(true ** 2))
                e_k, e_t = out - kept, out - true
                print(f"  {mode:10s} K={K:5d} npass={npass}: vs kept terms rel {np.sqrt(np.mean(e_k ** 2)) / rms:.3e} bias {np.mean(e_k * np.sign(kept)) / rms:+.3e}"
                      f" | vs fp32 operands rel {np.sqrt(np.mean(e_t ** 2)) / rms:.3e} | dropped lo*lo rel {np.sqrt(np.mean((kept - full) ** 2)) / rms:.3e}"
                      f" | operand split rel {np.sqrt(np.mean((full - true) ** 2)) / rms:.3e}", flush=True)
        if mode == "tc":
            # probe 1: one MMA instruction (k16 group): 1 + m * 2^-24, m = 3 -> RN gives 1 + 2 ulp, truncation 1 + 1 ulp
            K = 64
            A = np.zeros((128, K), np.float32); B = np.zeros((64, K), np.float32)
            A[:, 0] = 1.0; B[:, 0] = 1.0
            A[:, 1:4] = 2.0 ** -11; B[:, 1:4] = 2.0 ** -13
            r = c.debug_gemm(A, B, 1)[0, 0]
            print(f"  probe within one k16 step: 1 + 3*2^-24 -> (out-1)/ulp = {(float(r) - 1.0) / 2.0 ** -23:.2f}  (RN: 2, truncation: 1)")
            # probe 2: one 0.75-ulp increment per k16 step, 8 steps: RN climbs to ~1 + 6 ulp, truncation stays at 1
            K = 192
            A = np.zeros((128, K), np.float32); B = np.zeros((64, K), np.float32)
            A[:, 0] = 1.0; B[:, 0] = 1.0
            for s in range(1, 9):
                A[:, 16 * s] = 3.0 * 2.0 ** -12; B[:, 16 * s] = 2.0 ** -13
            r = c.debug_gemm(A, B, 1)[0, 0]
            print(f"  probe across k16 steps: 1 + 8 x 0.75 ulp -> (out-1)/ulp = {(float(r) - 1.0) / 2.0 ** -23:.2f}  (RN: 6, truncation: 0)")
            # probe 3: same with negative accumulator (direction of the truncation)
            A[:, 0] = -1.0
            r = c.debug_gemm(A, B, 1)[0, 0]
            print(f"  probe, accumulator -1 + 8 x 0.75 ulp -> (out+1)/ulp = {(float(r) + 1.0) / 2.0 ** -23:.2f}  (RN: 6, toward zero: 6 or 8, toward -inf: 0)")
        c.close()


STAGE_TAPS = [("embed", 1), ("pre0_qkv_rope", 1), ("pre0_attn", 1), ("pre0_x_attn", 1)] + [(f"pre{i}_out", 1) for i in range(6)] + \
             [("prenet_out", 1), ("upsample", 2), ("prior0_h1", 2), ("prior0_c1", 2), ("prior0_out", 2), ("prior", 2), ("dec0_out", 2), ("decoder_x", 2),
              ("post", 2), ("up0", 6), ("up1", 18), ("spec", 18)]


def part_stages(T=100):
    import torch
    from miocodec_b200 import binding, synth
    from oracle import port
    gguf, _ = synth.ensure_assets()
    hp, tensors = synth.synth_tensors(0)
    W = dict(tensors)
    codes, emb = synth.synth_codes(T, 5), synth.synth_voice(2)
    s64, i64 = port.decode(W, hp, codes, emb, want_inter=True, dtype=torch.float64)
    i64["spec"] = s64
    s32, i32 = port.decode(W, hp, codes, emb, want_inter=True)
    i32["spec"] = s32
    got = {}
    for mode in ("tc", "tc_nochunk", "simt"):
        c = binding.MioCodec(gguf, 0, gemm=mode.split("_")[0])
        if mode == "tc_nochunk":
            c.set_option("chunk_acc", "off")
        c.set_option("debug_snapshots", "1")
        c.decode_wave(codes, emb)
        got[mode] = {n: c.debug_read(n) for n, _ in STAGE_TAPS}
        c.close()
    print(f"== stages: SNR (dB) against the fp64 evaluation of the graph, T={T}")
    print(f"  {'tap':14s} {'tcgen05':>9s} {'unchunked':>9s} {'simt':>9s} {'fp32 port':>10s}")
    for name, r in STAGE_TAPS:
        ref = i64[name]
        row = []
        for mode in ("tc", "tc_nochunk", "simt"):
            g = got[mode][name].reshape(-1, ref.shape[1])[: T * r]
            row.append(snr(g, ref))
        print(f"  {name:14s} {row[0]:9.1f} {row[1]:9.1f} {row[2]:9.1f} {snr(i32[name], ref):10.1f}", flush=True)


VARIANTS = {
    "tc": {},
    "tc_unchunked": {"MIO_CHUNK_ACC": "off"},
    "tc_chunk_all": {"MIO_CHUNK_ACC": "all"},
    "tc_kb4": {"MIO_CHUNK_KB": "4"},
    "tc_kb6": {"MIO_CHUNK_KB": "6"},
    "simt": {"MIO_GEMM": "simt"},
    "attn_simt": {"MIO_ATTN": "simt"},
    "gn_precise": {"MIO_GN_PRECISE": "1"},
    "all_precise": {"MIO_GEMM": "simt", "MIO_ATTN": "simt", "MIO_GN_PRECISE": "1"},
}


def e2e_inputs(n):
    from miocodec_b200 import synth
    rng = np.random.default_rng(77)
    lens = rng.integers(40, 121, size=n)
    return [(synth.synth_codes(int(L), 1000 + i), i % 16) for i, L in enumerate(lens)]


def part_e2e_variant(variant, n):
    from miocodec_b200 import binding, synth
    gguf, _ = synth.ensure_assets()
    c = binding.MioCodec(gguf, 0)
    waves = [c.decode_wave(codes, synth.synth_voice(v)) for codes, v in e2e_inputs(n)]
    c.close()
    np.savez(os.path.join(OUT, f"e2e_{variant}.npz"), *waves)


def part_e2e(n):
    import torch
    from miocodec_b200 import synth
    from oracle import port, ref
    os.makedirs(OUT, exist_ok=True)
    gguf, _ = synth.ensure_assets()
    inputs = e2e_inputs(n)
    refs = {}
    for tag in ("v4", "v3"):
        os.environ["MIO_REF_VARIANT"] = tag
        try:
            R = ref.Reference(gguf)
        except (OSError, RuntimeError) as e:
            print(f"  reference {tag} unavailable here: {e}")
            continue
        refs[tag] = [R.decode_wave(codes, synth.synth_voice(v)) for codes, v in inputs]
        R.close()
    os.environ.pop("MIO_REF_VARIANT", None)
    main = "v4" if "v4" in refs else "v3"
    print(f"== e2e: {n} utterances of 40-120 codes against the live reference ({main}); SNR dB mean +- sd [min], max-abs mean [max]")

    def line(name, waves, against):
        s = np.array([snr(w, r) for w, r in zip(waves, against)])
        m = np.array([np.abs(w - r).max() for w, r in zip(waves, against)])
        print(f"  {name:28s} {s.mean():6.2f} +- {s.std():4.2f} [{s.min():5.2f}]   max-abs {m.mean():.2e} [{m.max():.2e}]", flush=True)

    if len(refs) == 2:
        line("reference v3 vs v4 (floor)", refs["v3"], refs["v4"])
    hp, tensors = synth.synth_tensors(0)
    W = dict(tensors)
    k = min(8, n)
    line(f"fp32 port (first {k})", [port.istft(port.decode(W, hp, c_, synth.synth_voice(v))) for c_, v in inputs[:k]], refs[main][:k])
    line(f"fp64 port (first {k})", [port.istft(port.decode(W, hp, c_, synth.synth_voice(v), dtype=torch.float64)) for c_, v in inputs[:k]], refs[main][:k])
    base = None
    for variant, env in VARIANTS.items():
        e = dict(os.environ); e.update(env)
        subprocess.check_call([sys.executable, os.path.abspath(__file__), "e2e_variant", "--variant", variant, "--n", str(n)], env=e)
        z = np.load(os.path.join(OUT, f"e2e_{variant}.npz"))
        waves = [z[f"arr_{i}"] for i in range(n)]
        line(f"GPU {variant}", waves, refs[main])
        if variant == "tc":
            base = waves
        os.remove(os.path.join(OUT, f"e2e_{variant}.npz"))
    # repeatability of the product path: two processes, same bits?
    subprocess.check_call([sys.executable, os.path.abspath(__file__), "e2e_variant", "--variant", "tc", "--n", str(n)])
    z = np.load(os.path.join(OUT, "e2e_tc.npz"))
    same = all(np.array_equal(z[f"arr_{i}"], base[i]) for i in range(n))
    print(f"  GPU tc, second process: bit-identical = {same}")
    os.remove(os.path.join(OUT, "e2e_tc.npz"))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("part")
    ap.add_argument("--n", type=int, default=32)
    ap.add_argument("--variant", default="tc")
    a = ap.parse_args()
    if a.part in ("gemm", "all"):
        part_gemm()
    if a.part in ("stages", "all"):
        part_stages()
    if a.part in ("e2e", "all"):
        part_e2e(a.n)
    if a.part == "e2e_variant":
        os.makedirs(OUT, exist_ok=True)
        part_e2e_variant(a.variant, a.n)
