"""Synthetic assets in the reference's exact on-disk formats.

The real ``miocodec.gguf`` / ``*.emb.gguf`` are not available offline, so tests and the
benchmark use random-init weights in the exact tensor layout ``miocodec_load`` expects
(/root/reference/src/miocodec.cpp:426-504 for the KV keys, :599-740 for the tensor names;
layout table in SURVEY.md §8c) and voice embeddings in the layout written by
/root/reference/tools/create_voice_emb.py:125-133.

This module carries its own minimal GGUF v3 writer (format described in
third_party/llama.cpp/ggml/include/gguf.h:1-31) so nothing but numpy is needed.
"""
from __future__ import annotations

import os
import struct
import numpy as np

GGML_F32, GGML_F16, GGML_I32, GGML_BF16 = 0, 1, 26, 30
_KV_U32, _KV_F32, _KV_STR = 4, 6, 8
_ALIGN = 32


def _s(b: str) -> bytes:
    e = b.encode("utf-8")
    return struct.pack("<Q", len(e)) + e


class GGUFWriter:
    """Tiny GGUF v3 writer: u32 / f32 / string KVs, F32 / F16 / BF16 / I32 tensors."""

    def __init__(self, arch: str):
        self.kv: list[tuple[str, int, object]] = [("general.architecture", _KV_STR, arch)]
        self.tensors: list[tuple[str, np.ndarray, int]] = []

    def add_u32(self, k, v): self.kv.append((k, _KV_U32, int(v)))
    def add_f32(self, k, v): self.kv.append((k, _KV_F32, float(v)))
    def add_str(self, k, v): self.kv.append((k, _KV_STR, str(v)))

    def add_tensor(self, name: str, arr: np.ndarray, bf16: bool = False):
        if bf16:     # arr: float32 values -> BF16 bit patterns (round to nearest even), stored as uint16
            b = np.ascontiguousarray(arr, dtype=np.float32).view(np.uint32)
            b = ((b + 0x7FFF + ((b >> 16) & 1)) >> 16).astype(np.uint16)
            self.tensors.append((name, b, GGML_BF16))
            return
        if arr.dtype == np.float32: t = GGML_F32
        elif arr.dtype == np.float16: t = GGML_F16
        elif arr.dtype == np.int32: t = GGML_I32
        else: raise TypeError(arr.dtype)
        self.tensors.append((name, np.ascontiguousarray(arr), t))

    def write(self, path: str):
        tmp = path + ".tmp%d" % os.getpid()
        with open(tmp, "wb") as f:
            f.write(b"GGUF" + struct.pack("<IQQ", 3, len(self.tensors), len(self.kv)))
            for k, t, v in self.kv:
                f.write(_s(k) + struct.pack("<I", t))
                if t == _KV_U32: f.write(struct.pack("<I", v))
                elif t == _KV_F32: f.write(struct.pack("<f", v))
                else: f.write(_s(v))
            off = 0
            offs = []
            for name, arr, t in self.tensors:
                offs.append(off)
                # ggml ne order = reversed numpy shape (ne0 is the contiguous axis)
                ne = list(reversed(arr.shape))
                f.write(_s(name) + struct.pack("<I", len(ne)))
                f.write(struct.pack("<%dQ" % len(ne), *ne))
                f.write(struct.pack("<IQ", t, off))
                off += (arr.nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
            pad = (-f.tell()) % _ALIGN
            f.write(b"\0" * pad)
            for (name, arr, t), o in zip(self.tensors, offs):
                f.write(arr.tobytes())
                f.write(b"\0" * ((-arr.nbytes) % _ALIGN))
        os.replace(tmp, path)


# --------------------------------------------------------------------------------------
# hyper-parameters (defaults of /root/reference/src/miocodec.cpp:59-85)
# --------------------------------------------------------------------------------------
DEFAULT_HP = dict(
    sample_rate=44100, n_fft=392, hop_length=98, samples_per_token=1764, head_out=394,
    prenet_layers=6, prenet_dim=768, prenet_heads=12, prenet_ff=2048, prenet_window=65,
    decoder_layers=8, decoder_dim=512, decoder_heads=8, decoder_ff=1536, decoder_window=65,
    adaln_dim=128, resnet_blocks=2, resnet_groups=32, upsampler_stages=2,
    rope_theta=10000.0, norm_eps=1e-5, group_norm_eps=1e-6,
    vocab=12800, up_factors=(3, 3), up_kernels=(9, 9), up_channels=(256, 128), out_proj_dim=512,
)


def synth_tensors(seed: int = 0, hp: dict | None = None, scale_embd: float = 0.02):
    """Yield (name, ndarray in PyTorch-order shape) for all 227 tensors, deterministically."""
    hp = dict(DEFAULT_HP, **(hp or {}))
    rng = np.random.default_rng(seed)

    def W(*shape):  # linear / conv weight ~ N(0, 1/fan_in), fan_in = prod(shape[1:])
        fan_in = int(np.prod(shape[1:]))
        return (rng.standard_normal(shape, dtype=np.float32) / np.float32(np.sqrt(fan_in))).astype(np.float32)

    def G(n):   # norm gain
        return (1.0 + 0.05 * rng.standard_normal(n, dtype=np.float32)).astype(np.float32)

    def Bz(*shape):  # biases, snake params
        return (0.02 * rng.standard_normal(shape, dtype=np.float32)).astype(np.float32)

    out = []
    D, FF = hp["prenet_dim"], hp["prenet_ff"]
    out.append(("token_embd", (scale_embd * rng.standard_normal((hp["vocab"], D), dtype=np.float32)).astype(np.float32)))
    for i in range(hp["prenet_layers"]):
        p = f"wave_prenet.blk.{i}."
        out += [(p + "attn_norm.weight", G(D)), (p + "attn_norm.bias", Bz(D))]
        for n in ("attn_q", "attn_k", "attn_v", "attn_output"):
            out.append((p + n + ".weight", W(D, D)))
        out += [(p + "ffn_norm.weight", G(D)), (p + "ffn_norm.bias", Bz(D))]
        out += [(p + "ffn_gate.weight", W(FF, D)), (p + "ffn_up.weight", W(FF, D)), (p + "ffn_down.weight", W(D, FF))]
    C = hp["decoder_dim"]
    out += [("wave_prenet.norm.weight", G(D)), ("wave_prenet.norm.bias", Bz(D)),
            ("wave_prenet.output.weight", W(C, D)), ("wave_prenet.output.bias", Bz(C))]
    out += [("wave_upsample.weight", W(C, C, 2)), ("wave_upsample.bias", Bz(C))]   # ConvT: (Cin, Cout, K)

    def resblk(p, ch):
        r = []
        for j in (1, 2):
            r += [(p + f"norm{j}.weight", G(ch)), (p + f"norm{j}.bias", Bz(ch)),
                  (p + f"conv{j}.weight", W(ch, ch, 3)), (p + f"conv{j}.bias", Bz(ch))]  # Conv1d: (Cout, Cin, 3)
        return r

    for b in range(hp["resnet_blocks"]):
        out += resblk(f"wave_prior.{b}.", C)
    A, DF = hp["adaln_dim"], hp["decoder_ff"]
    for i in range(hp["decoder_layers"]):
        p = f"wave_decoder.blk.{i}."
        out += [(p + "attn_cond.weight", W(3 * C, A)), (p + "attn_cond.bias", Bz(3 * C)),
                (p + "ffn_cond.weight", W(3 * C, A)), (p + "ffn_cond.bias", Bz(3 * C))]
        for n in ("attn_q", "attn_k", "attn_v", "attn_output"):
            out.append((p + n + ".weight", W(C, C)))
        out += [(p + "ffn_gate.weight", W(DF, C)), (p + "ffn_up.weight", W(DF, C)), (p + "ffn_down.weight", W(C, DF))]
    out += [("wave_decoder.norm_cond.weight", W(2 * C, A)), ("wave_decoder.norm_cond.bias", Bz(2 * C))]
    for b in range(hp["resnet_blocks"]):
        out += resblk(f"wave_post.{b}.", C)
    out += [("miocodec.wave_upsampler.factors", np.asarray(hp["up_factors"], dtype=np.int32)),
            ("miocodec.wave_upsampler.kernel_sizes", np.asarray(hp["up_kernels"], dtype=np.int32))]
    cin = C
    for s in range(hp["upsampler_stages"]):
        co, k = hp["up_channels"][s], hp["up_kernels"][s]
        out += [(f"wave_upsampler.up.{s}.weight", W(cin, co, k)), (f"wave_upsampler.up.{s}.bias", Bz(co)),
                (f"wave_upsampler.snake.{s}.alpha", Bz(co)), (f"wave_upsampler.snake.{s}.beta", Bz(co))]
        out += resblk(f"wave_upsampler.resblk.{s}.", co)
        cin = co
    P = hp["out_proj_dim"]
    out += [("wave_upsampler.out_proj.weight", W(P, cin)), ("wave_upsampler.out_proj.bias", Bz(P)),
            ("wave_upsampler.out_snake.alpha", Bz(P)), ("wave_upsampler.out_snake.beta", Bz(P))]
    out += [("istft_head.out.weight", W(hp["head_out"], P)), ("istft_head.out.bias", Bz(hp["head_out"]))]
    return hp, out


def bf16_round(a: np.ndarray) -> np.ndarray:
    """float32 -> nearest BF16 value (ties to even), returned as float32."""
    b = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    return (((b + 0x7FFF + ((b >> 16) & 1)) >> 16) << 16).astype(np.uint32).view(np.float32)


def write_codec_gguf(path: str, seed: int = 0, hp: dict | None = None, f16_matrices: bool = False,
                     matrices: str = "f32", convs: str | None = None, pre_round: str | None = None) -> dict:
    """Write a random-init miocodec.gguf; returns the hyper-parameter dict used.

    matrices: storage type of the 2-D matrices (linear layers, conditioning linears, token table): "f32" | "f16" | "bf16".
    convs:    storage type of the conv / ConvTranspose kernels: "f32" | "f16" (default: "f16" when matrices == "f16").
    pre_round: round the VALUES of those tensors to "f16" / "bf16" but keep the storage chosen above -- two files that
              hold identical numbers in different storage types isolate the reference's activation rounding.
    f16_matrices: the round-1 switch (F16 for everything >= 2-D but the token table and the conditioning linears).
    """
    hp = dict(hp or {})
    if "up_factors" in hp and "samples_per_token" not in hp:
        fpc = 2 * int(np.prod(hp["up_factors"]))
        hp["samples_per_token"] = fpc * hp.get("hop_length", DEFAULT_HP["hop_length"])
    hp, tensors = synth_tensors(seed, hp)
    if convs is None:
        convs = "f16" if matrices == "f16" else "f32"
    w = GGUFWriter("miocodec")
    for k in ("sample_rate", "n_fft", "hop_length", "samples_per_token"):
        w.add_u32("miocodec." + k, hp[k])
    w.add_u32("embedding_length_out", hp["head_out"])
    for k in ("prenet_layers", "prenet_dim", "prenet_heads", "prenet_ff", "prenet_window",
              "decoder_layers", "decoder_dim", "decoder_heads", "decoder_ff", "decoder_window",
              "resnet_blocks", "resnet_groups"):
        w.add_u32("miocodec." + k, hp[k])
    w.add_u32("miocodec.decoder_adanorm_dim", hp["adaln_dim"])
    w.add_u32("miocodec.wave_upsampler_layers", hp["upsampler_stages"])
    w.add_f32("miocodec.rope_theta", hp["rope_theta"])
    w.add_f32("miocodec.norm_eps", hp["norm_eps"])
    w.add_f32("miocodec.group_norm_eps", hp["group_norm_eps"])
    for name, arr in tensors:
        if f16_matrices and arr.ndim >= 2 and name != "token_embd" and not name.endswith("_cond.weight"):
            arr = arr.astype(np.float16)
        elif arr.dtype == np.float32 and arr.ndim >= 2:
            kind = matrices if arr.ndim == 2 else convs
            rnd = pre_round if pre_round is not None else kind
            if rnd == "f16": arr = arr.astype(np.float16).astype(np.float32)
            elif rnd == "bf16": arr = bf16_round(arr)
            if kind == "f16": arr = arr.astype(np.float16)
            elif kind == "bf16":
                w.add_tensor(name, arr, bf16=True)
                continue
        w.add_tensor(name, arr)
    w.write(path)
    return hp


def synth_voice(voice_id: int, dim: int = 128) -> np.ndarray:
    return np.random.default_rng(1000 + voice_id).standard_normal(dim).astype(np.float32)


def write_voice_gguf(path: str, emb: np.ndarray):
    w = GGUFWriter("mio-embedding")
    w.add_u32("mio.embedding.dim", emb.shape[0])
    w.add_tensor("mio.global_embedding", emb.astype(np.float32))
    w.write(path)


def synth_codes(n: int, seed: int, vocab: int = 12800) -> np.ndarray:
    return np.random.default_rng(seed).integers(0, vocab, size=n, dtype=np.int64).astype(np.int32)


# the one golden INPUT vector the reference ships (/root/reference/tools/compare_codec.py:50-51)
REFERENCE_TEST_CODES = np.asarray(
    [12287, 11619, 11774, 12223, 2490, 826, 2257, 1668, 1219, 2319, 9994, 12683, 12745, 4215,
     12478, 8800, 8696, 375, 1406, 12396], dtype=np.int32)


def asset_dir() -> str:
    d = os.environ.get("MIO_ASSET_DIR") or os.path.join(os.environ.get("TMPDIR", "/tmp"), "mio_b200_assets")
    os.makedirs(d, exist_ok=True)
    return d


def ensure_assets(seed: int = 0, n_voices: int = 16) -> tuple[str, list[str]]:
    """Create (once) the synthetic codec GGUF and voice files; returns their paths."""
    d = asset_dir()
    g = os.path.join(d, f"miocodec_synth_seed{seed}.gguf")
    # several ranks may get here at once: write to a private name, then publish atomically
    if not os.path.exists(g):
        tmp = f"{g}.tmp{os.getpid()}"
        write_codec_gguf(tmp, seed)
        os.replace(tmp, g)
    voices = []
    for v in range(n_voices):
        p = os.path.join(d, f"voice_{v}.emb.gguf")
        if not os.path.exists(p):
            tmp = f"{p}.tmp{os.getpid()}"
            write_voice_gguf(tmp, synth_voice(v))
            os.replace(tmp, p)
        voices.append(p)
    return g, voices
