"""Which tap is not transparent under mio_debug_write of its own snapshot?  (debug helper)"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))
from miocodec_b200 import binding, synth
gguf, _ = synth.ensure_assets()
c = binding.MioCodec(gguf, 0)
c.set_option("debug_snapshots", "1")
codes, emb = synth.REFERENCE_TEST_CODES, synth.synth_voice(0)
base = c.decode_wave(codes, emb)
print("repeat equal:", np.array_equal(base, c.decode_wave(codes, emb)))
names = ["embed", "pre0_qkv_rope", "pre0_attn", "pre0_out", "prenet_out", "upsample", "prior0_h1", "prior0_c1", "prior0_h2", "prior0_out",
         "dec3_out", "final_norm", "up0_snake", "upres1_h2", "up1", "out_snake", "spec"]
snap = {n: c.debug_read(n) for n in names}
for n in names:
    c.debug_clear()
    c.debug_write(n, snap[n])
    w = c.decode_wave(codes, emb)
    back = c.debug_read(n)
    print(f"{n:14s} wave equal {np.array_equal(w, base)}  tap re-read equal {np.array_equal(back, snap[n])}  n={snap[n].size} "
          f"maxdiff {np.abs(w - base).max():.3e}", flush=True)
c.close()
