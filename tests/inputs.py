"""Deterministic input set shared by tests/golden/make_golden.py and the tests.

Everything is a pure function of seeds, so the GPU box regenerates byte-identical GGUFs / PCM (neither travels:
the 2-layer GGUFs are 108 MB / 256 MB). Files are cached under $NB200_TMP (default /tmp/nb200).
"""
from __future__ import annotations

import os
import sys
import zlib
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tools"))
TMP = Path(os.environ.get("NB200_TMP", "/tmp/nb200"))

SEED = 1234
TEST_LAYERS = 2  # parity fixtures use a 2-layer model of the real widths: every kernel and both cache paths run

# (tensor name, M) of the single-matmul fixtures
MATMUL_CASES = (
    ("encoder.layers.0.feed_forward1.linear1.weight", 1),
    ("encoder.layers.0.feed_forward1.linear1.weight", 6),
    ("encoder.layers.0.feed_forward1.linear2.weight", 6),
    ("encoder.layers.1.conv.pointwise_conv1.weight", 6),
    ("encoder.layers.1.self_attn.linear_out.weight", 6),
)

# (variant, right_context, pcm seed, seconds) of the stream fixtures. 8 s at rc=13 wraps the K/V ring (84 frames) and
# saturates cache_valid_len (70); every run ends with nemo_stream_finalize, so the zero-padded tail chunk is covered.
STREAM_CASES = (
    ("q8_0", 0, 11, 3.0),
    ("q8_0", 13, 11, 4.0),
    ("f32dq", 0, 11, 3.0),
    ("f32dq", 13, 11, 8.0),
    ("f32dq", 1, 12, 2.0),
    ("f32dq", 6, 12, 2.5),
)


# The full-depth model (24 layers, what the benchmark runs): (variant, rc, pcm seed, seconds). The F32 file pins the 24-layer
# stack against the reference (fp16-operand error through 24 layers); the Q8_0 file is stored from both ISA builds so the
# reference's own self-noise at full depth is known (tests/golden/make_golden_deep.py).
DEEP_LAYERS = 24
DEEP_CASES = (
    ("f32dq", 13, 71, 3.6),
    ("q8_0", 13, 71, 3.6),
)


# the converter's other on-disk formats: checked against the compiled reference on the GPU path only
FORMAT_CASES = (
    ("f16", 13, 11, 3.0),
    ("q4_0", 13, 11, 3.0),
    ("q4_0", 0, 12, 1.5),
)


# multilingual checkpoint (prompt fusion): (language code or None = the loader's default "auto", rc, seed, secs)
PROMPT_CASES = (
    (None, 13, 11, 3.0),
    ("de", 13, 11, 3.0),
    ("ja", 0, 12, 1.0),
)


# offline path (nemo_encode, full-sequence attention, timed tokens): (variant, seed, seconds, language);
# "ml" = the multilingual F32 file (prompt fusion; language None = the loader's default prompt). The 30 s case (376 frames,
# 751 relative positions) stores every 8th encoder frame.
OFFLINE_CASES = (
    ("f32dq", 41, 6.0, None),
    ("f32dq", 42, 0.9, None),
    ("q8_0", 41, 4.0, None),
    ("ml", 41, 3.0, None),
    ("ml", 41, 3.0, "de"),
    ("f32dq", 44, 30.0, None),
)
OFFLINE_ENC_STRIDE = 8  # for utterances longer than 10 s


def offline_gguf(variant: str) -> str:
    return gguf_path("f32dq", multilingual=True) if variant == "ml" else gguf_path(variant)


def offline_name(variant: str, seed: int, lang) -> str:
    return f"offline_{variant}{'_' + lang if lang else ''}_s{seed}.npz"


def gguf_path(variant: str, layers: int = TEST_LAYERS, blank_gain: float | None = None, multilingual: bool = False,
              decoder: str = "bistable", legacy_vocab: bool = False, res_scale: float = 1.0) -> str:
    import make_random_gguf as mg
    TMP.mkdir(parents=True, exist_ok=True)
    tag = ("" if blank_gain is None else f"_bg{blank_gain:g}") + ("_ml" if multilingual else "") + \
          ("" if decoder == "bistable" else f"_{decoder}") + ("_legacyvocab" if legacy_vocab else "") + \
          ("" if res_scale == 1.0 else f"_rs{res_scale:g}")
    p = TMP / f"L{layers}_{variant}_s{SEED}{tag}.gguf"
    if not p.exists():
        tmp = p.with_suffix(f".tmp{os.getpid()}")
        mg.write_gguf(str(tmp), SEED, layers, variant, blank_gain, multilingual, decoder=decoder, legacy_vocab=legacy_vocab, res_scale=res_scale)
        os.replace(tmp, p)
    return str(p)


def pcm(seed: int, seconds: float) -> np.ndarray:
    import make_pcm
    return make_pcm.make_pcm(seconds, seed)


def pcm_path(seed: int, seconds: float) -> str:
    TMP.mkdir(parents=True, exist_ok=True)
    p = TMP / f"pcm_s{seed}_{seconds:g}s.pcm"
    if not p.exists():
        tmp = p.with_suffix(f".tmp{os.getpid()}")
        pcm(seed, seconds).tofile(tmp)
        os.replace(tmp, p)
    return str(p)


def ensure_inputs() -> dict:
    return {"q8_0": gguf_path("q8_0"), "f32dq": gguf_path("f32dq")}


def matmul_input(name: str, M: int, K: int | None = None) -> np.ndarray:
    """Seeded activations with a realistic spread (a few large entries per 32-block, like LN outputs)."""
    if K is None:
        K = 4096 if name.endswith("linear2.weight") else 1024
    rng = np.random.default_rng([SEED, zlib.crc32(name.encode()), M])
    return (rng.standard_normal((M, K)) * rng.uniform(0.2, 3.0, (M, 1))).astype(np.float32)
