"""Stage-by-stage parity report on the GPU box: CUDA snapshots vs the torch oracle port, for both GEMM paths.
usage: python tools/gpu_check.py [simt|tc] [T]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
from miocodec_b200 import binding, synth  # noqa: E402
from oracle import port  # noqa: E402


def snr(a, b):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return 10 * np.log10(np.sum(b * b) / max(np.sum((a - b) ** 2), 1e-300))


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "simt"
    T = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    gguf, voices = synth.ensure_assets()
    hp, tensors = synth.synth_tensors(0)
    W = dict(tensors)
    codes = synth.REFERENCE_TEST_CODES if T == 20 else synth.synth_codes(T, 99)
    emb = synth.synth_voice(0)
    t0 = time.time()
    spec_p, inter = port.decode(W, hp, codes, emb, want_inter=True)
    inter["spec"] = spec_p
    wave_p = port.istft(spec_p)
    print(f"[{mode}] port decode {time.time() - t0:.2f}s", flush=True)
    c = binding.MioCodec(gguf, 0, gemm=mode)
    c.set_option("debug_snapshots", "1")
    # ISTFT alone first (no GEMM involved)
    g = np.load(os.path.join(ROOT, "tests", "golden", "istft50.npz"))
    w = c.istft(g["spec"])
    print(f"[{mode}] istft-only vs reference golden: SNR {snr(w, g['wave']):.1f} dB maxabs {np.abs(w - g['wave']).max():.2e}", flush=True)
    t0 = time.time()
    wave = c.decode_wave(codes, emb)
    print(f"[{mode}] decode_wave T={T}: {time.time() - t0:.3f}s launches={c.kernel_launches()}", flush=True)
    res = {"embed": 1, "pre0_qkv_rope": 1, "pre0_attn": 1, "pre0_x_attn": 1, "pre0_out": 1, "prenet_x": 1, "prenet_out": 1,
           "upsample": 2, "prior": 2, "dec0_out": 2, "decoder_x": 2, "final_norm": 2, "post": 2, "up0_snake": 6, "up0": 6,
           "up1": 18, "spec": 18}
    for name, r in res.items():
        ref = inter[name]
        got = c.debug_read(name).reshape(-1, ref.shape[1])[: T * r]
        bad = int(np.sum(~np.isfinite(got)))
        print(f"[{mode}] {name:12s} SNR {snr(got, ref):7.1f} dB  maxabs {np.abs(got - ref).max():.3e}  ref_rms {np.sqrt(np.mean(ref ** 2)):.3f} nonfinite {bad}", flush=True)
    print(f"[{mode}] wave vs port: SNR {snr(wave, wave_p):.1f} dB maxabs {np.abs(wave - wave_p).max():.2e}")
    if T == 20:
        gg = np.load(os.path.join(ROOT, "tests", "golden", "ref20.npz"))
        print(f"[{mode}] wave vs reference golden: SNR {snr(wave, gg['wave']):.1f} dB maxabs {np.abs(wave - gg['wave']).max():.2e}")
        print(f"[{mode}] port vs reference golden: SNR {snr(wave_p, gg['wave']):.1f} dB")
    c.set_option("debug_snapshots", "0")
    # timing without snapshots
    for _ in range(3):
        t0 = time.time(); c.decode_wave(codes, emb); dt = time.time() - t0
    print(f"[{mode}] decode_wave T={T} warm: {dt * 1e3:.2f} ms")
    c.close()


if __name__ == "__main__":
    main()
