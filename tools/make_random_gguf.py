#!/usr/bin/env python3
"""Write a random-init nemotron-speech-streaming GGUF (same architecture and on-disk layout as the real
0.6B checkpoint) for parity tests and benchmarks. There is no network, so the real checkpoint is absent.

The on-disk format follows the reference's converter, which defines it:
  /root/reference/scripts/convert_to_gguf.py:491-536   header / KV / tensor-info / 32-byte aligned data
  /root/reference/scripts/convert_to_gguf.py:118-155   Q8_0 quantiser (fp16 scale = amax/127, rne(x/scale))
  /root/reference/scripts/convert_to_gguf.py:246-263   which tensors are quantised (11 matrices per layer)
  /root/reference/scripts/convert_to_gguf.py:221-243,399-413  conv reshapes (pointwise squeezed, depthwise (9,1024))
  /root/reference/src/nemo-ggml.cpp:108-182,296-398    keys and tensor names the loader reads

Everything is a pure function of (seed, tensor name), so the GPU box regenerates byte-identical files.
variants: q8_0 (11 matrices/layer as Q8_0), f32 (same weights unquantised), f32dq (F32 file holding the
dequantised Q8_0 weights), f16 and q4_0 (the converter's other two --quantize modes, convert_to_gguf.py:157-204,322-336).
"""
from __future__ import annotations

import argparse
import struct
import sys
import zlib

import numpy as np

GGML_F32, GGML_F16, GGML_Q4_0, GGML_Q8_0 = 0, 1, 2, 8
ALIGN = 32

N_MELS, D_MODEL, N_HEADS, D_HEAD, D_FF, KERNEL = 128, 1024, 8, 128, 4096, 9
VOCAB, DEC, JOINT, SUBS, LEFT = 1025, 640, 640, 8, 70
CONV_CH, FREQ_OUT = 256, 17


def _rng(seed: int, name: str) -> np.random.Generator:
    return np.random.default_rng([seed, zlib.crc32(name.encode())])


def normal(seed, name, shape, std):
    return (_rng(seed, name).standard_normal(shape, dtype=np.float32) * np.float32(std)).astype(np.float32)


def uniform(seed, name, shape, bound):
    return _rng(seed, name).uniform(-bound, bound, shape).astype(np.float32)


def quantize_q8_0(a: np.ndarray):
    """Returns (scales fp16 [nb], quants int8 [nb,32]) the way the converter does."""
    blocks = a.astype(np.float32).reshape(-1, 32)
    amax = np.max(np.abs(blocks), axis=1)
    scales = np.where(amax != 0, amax / np.float32(127.0), np.float32(0)).astype(np.float16)
    s32 = scales.astype(np.float32)[:, None]
    safe = np.where(s32 != 0, s32, np.float32(1))
    q = np.round(blocks / safe)
    q = np.where(s32 != 0, q, 0).astype(np.int8)
    return scales, q


def q8_0_bytes(a: np.ndarray) -> bytes:
    scales, q = quantize_q8_0(a)
    out = np.empty(scales.shape[0], dtype=np.dtype([("d", np.float16), ("q", np.int8, 32)]))
    out["d"] = scales
    out["q"] = q
    return out.tobytes()


def q4_0_bytes(a: np.ndarray) -> bytes:
    """Q4_0 blocks the way the converter writes them (convert_to_gguf.py:157-204): fp16 d = amax/7, q = clip(rne(x/d), -8, 7),
    byte j = (q[j] + 8) | (q[j + 16] + 8) << 4."""
    blocks = a.astype(np.float32).reshape(-1, 32)
    amax = np.max(np.abs(blocks), axis=1)
    scales = np.where(amax != 0, amax / 7.0, 0.0).astype(np.float16)
    s32 = scales.astype(np.float32)[:, None]
    safe = np.where(s32 != 0, s32, 1.0)
    q = np.clip(np.round(blocks / safe).astype(np.int8), -8, 7)
    q = np.where(s32 != 0, q, 0)
    u = (q + 8).astype(np.uint8)
    packed = ((u[:, :16] & 0x0F) | ((u[:, 16:] & 0x0F) << 4)).astype(np.uint8)
    out = np.empty(scales.shape[0], dtype=np.dtype([("d", np.float16), ("q", np.uint8, 16)]))
    out["d"] = scales
    out["q"] = packed
    return out.tobytes()


def dequantize_q8_0(a: np.ndarray) -> np.ndarray:
    scales, q = quantize_q8_0(a)
    return (q.astype(np.float32) * scales.astype(np.float32)[:, None]).reshape(a.shape)


def mel_filterbank() -> np.ndarray:
    """Slaney-style triangular filterbank [128, 257] (any positive triangular bank exercises the path)."""
    sr, n_fft = 16000, 512

    def hz_to_mel(f):
        f = np.asarray(f, dtype=np.float64)
        lin = f / (200.0 / 3)
        log = 15.0 + np.log(np.maximum(f, 1e-9) / 1000.0) / (np.log(6.4) / 27.0)
        return np.where(f >= 1000.0, log, lin)

    def mel_to_hz(m):
        m = np.asarray(m, dtype=np.float64)
        lin = m * (200.0 / 3)
        log = 1000.0 * np.exp((np.log(6.4) / 27.0) * (m - 15.0))
        return np.where(m >= 15.0, log, lin)

    freqs = np.linspace(0, sr / 2, n_fft // 2 + 1)
    pts = mel_to_hz(np.linspace(hz_to_mel(0.0), hz_to_mel(sr / 2), N_MELS + 2))
    fb = np.zeros((N_MELS, n_fft // 2 + 1))
    for m in range(N_MELS):
        lo, ce, hi = pts[m], pts[m + 1], pts[m + 2]
        up = (freqs - lo) / (ce - lo)
        dn = (hi - freqs) / (hi - ce)
        fb[m] = np.maximum(0, np.minimum(up, dn)) * (2.0 / (hi - lo))
    return fb.astype(np.float32)


NUM_PROMPTS, PROMPT_HIDDEN = 128, 2048
# language code -> prompt id, like the multilingual checkpoints' dictionary (src/nemo-ggml.cpp:171-181; 101 = "auto" is the
# default of nemo_init_with_backend, :459-462)
PROMPT_DICT = (("en", 0), ("de", 3), ("fr", 7), ("es", 12), ("ja", 64), ("auto", 101))


def prompt_tensors(seed: int):
    """prompt_kernel.{0,2}: Linear(1024 + 128 -> 2048) + ReLU + Linear(2048 -> 1024), F32 (src/nemo-ggml.cpp:390-394,1087-1105).
    The one-hot columns are large enough that the language visibly changes the encoder output."""
    s = seed
    w0 = normal(s, "prompt_kernel.0.weight", (PROMPT_HIDDEN, D_MODEL + NUM_PROMPTS), (2.0 / D_MODEL) ** 0.5)
    w0[:, D_MODEL:] = normal(s, "prompt_kernel.0.weight.onehot", (PROMPT_HIDDEN, NUM_PROMPTS), 0.7)
    yield "prompt_kernel.0.weight", w0, False
    yield "prompt_kernel.0.bias", normal(s, "prompt_kernel.0.bias", (PROMPT_HIDDEN,), 0.1), False
    yield "prompt_kernel.2.weight", normal(s, "prompt_kernel.2.weight", (D_MODEL, PROMPT_HIDDEN), (2.0 / PROMPT_HIDDEN) ** 0.5), False
    yield "prompt_kernel.2.bias", normal(s, "prompt_kernel.2.bias", (D_MODEL,), 0.1), False


def tensor_list(seed: int, n_layers: int, blank_gain: float | None = None, decoder: str = "bistable", res_scale: float = 1.0):
    """Yields (name, float32 array in PyTorch shape order, quantisable?). decoder: see DECODERS below. res_scale multiplies the
    output projection of every residual branch (FFN linear2, attention linear_out, pointwise_conv2), the usual depth-scaled
    initialisation: with 1.0 the output of a 24-layer random stack is an almost constant vector (96 random branch outputs swamp
    the frontend's features; adjacent frames correlate at 0.997), with ~0.1 it still follows the audio frame by frame."""
    s = seed
    rs = np.float32(res_scale)
    yield "preprocessor.featurizer.fb", mel_filterbank()[None], False
    n = np.arange(400, dtype=np.float64)
    yield "preprocessor.featurizer.window", (0.5 - 0.5 * np.cos(2 * np.pi * n / 399)).astype(np.float32), False

    pe = "encoder.pre_encode."
    yield pe + "conv.0.weight", normal(s, pe + "conv.0.weight", (CONV_CH, 1, 3, 3), (2.0 / 9) ** 0.5 * 0.25), False
    yield pe + "conv.0.bias", normal(s, pe + "conv.0.bias", (CONV_CH,), 0.5), False
    for i in (2, 5):
        yield pe + f"conv.{i}.weight", normal(s, pe + f"conv.{i}.weight", (CONV_CH, 1, 3, 3), (1.0 / 9) ** 0.5), False
        yield pe + f"conv.{i}.bias", normal(s, pe + f"conv.{i}.bias", (CONV_CH,), 0.1), False
    for i in (3, 6):
        yield pe + f"conv.{i}.weight", normal(s, pe + f"conv.{i}.weight", (CONV_CH, CONV_CH, 1, 1), (2.0 / CONV_CH) ** 0.5), False
        yield pe + f"conv.{i}.bias", normal(s, pe + f"conv.{i}.bias", (CONV_CH,), 0.1), False
    k_out = CONV_CH * FREQ_OUT
    yield pe + "out.weight", normal(s, pe + "out.weight", (D_MODEL, k_out), (1.0 / k_out) ** 0.5), False
    yield pe + "out.bias", normal(s, pe + "out.bias", (D_MODEL,), 0.1), False

    for l in range(n_layers):
        p = f"encoder.layers.{l}."
        for nm in ("norm_feed_forward1", "norm_self_att", "norm_conv", "norm_feed_forward2", "norm_out"):
            yield p + nm + ".weight", 1.0 + normal(s, p + nm + ".weight", (D_MODEL,), 0.05), False
            yield p + nm + ".bias", normal(s, p + nm + ".bias", (D_MODEL,), 0.02), False
        for ff in ("feed_forward1", "feed_forward2"):
            yield p + ff + ".linear1.weight", normal(s, p + ff + ".linear1.weight", (D_FF, D_MODEL), (1.0 / D_MODEL) ** 0.5), True
            yield p + ff + ".linear2.weight", normal(s, p + ff + ".linear2.weight", (D_MODEL, D_FF), (1.0 / D_FF) ** 0.5) * rs, True
        for nm in ("linear_q", "linear_k", "linear_v", "linear_pos", "linear_out"):
            yield p + f"self_attn.{nm}.weight", normal(s, p + f"self_attn.{nm}.weight", (D_MODEL, D_MODEL), (1.0 / D_MODEL) ** 0.5) * \
                (rs if nm == "linear_out" else np.float32(1.0)), True
        yield p + "self_attn.pos_bias_u", normal(s, p + "self_attn.pos_bias_u", (N_HEADS, D_HEAD), 0.3), False
        yield p + "self_attn.pos_bias_v", normal(s, p + "self_attn.pos_bias_v", (N_HEADS, D_HEAD), 0.3), False
        yield p + "conv.pointwise_conv1.weight", normal(s, p + "conv.pointwise_conv1.weight", (2 * D_MODEL, D_MODEL), (1.0 / D_MODEL) ** 0.5), True
        # stored (kernel, channels): /root/reference/scripts/convert_to_gguf.py:237-243
        yield p + "conv.depthwise_conv.weight", normal(s, p + "conv.depthwise_conv.weight", (KERNEL, D_MODEL), (1.0 / KERNEL) ** 0.5), False
        yield p + "conv.batch_norm.weight", 1.0 + normal(s, p + "conv.batch_norm.weight", (D_MODEL,), 0.05), False
        yield p + "conv.batch_norm.bias", normal(s, p + "conv.batch_norm.bias", (D_MODEL,), 0.02), False
        yield p + "conv.pointwise_conv2.weight", normal(s, p + "conv.pointwise_conv2.weight", (D_MODEL, D_MODEL), (1.0 / D_MODEL) ** 0.5) * rs, True

    d = "decoder.prediction."
    wo = normal(s, "joint.joint_net.2.weight", (VOCAB, JOINT), 1.0 / JOINT ** 0.5)
    emb = normal(s, d + "embed.weight", (VOCAB, DEC), 1.0)
    if decoder == "stationary":
        emb = wo.copy()  # the prediction network carries the previous token's OUTPUT row through to joint.pred (see DECODERS)
    emb[VOCAB - 1] = 0.0  # blank/pad row is zero in a trained model
    yield d + "embed.weight", emb, False
    b = 1.0 / DEC ** 0.5
    for l in (0, 1):
        for nm, shp in (("weight_ih", (4 * DEC, DEC)), ("weight_hh", (4 * DEC, DEC)), ("bias_ih", (4 * DEC,)), ("bias_hh", (4 * DEC,))):
            name = d + f"dec_rnn.lstm.{nm}_l{l}"
            w = uniform(s, name, shp, LSTM_BIAS if nm.startswith("bias") else b)
            if decoder == "stationary":  # gate order i, f, g, o (src/nemo-ggml.cpp:601-618)
                w = w * np.float32(STATIONARY_NOISE)
                if nm == "weight_ih":
                    w[2 * DEC:3 * DEC] += np.float32(STATIONARY_PASS) * np.eye(DEC, dtype=np.float32)  # g = tanh(pass * x): x passes through
                if nm == "bias_ih":
                    w[0:DEC] += 4.0            # input gates open
                    # layer 0 remembers: c = MEMORY c + tanh(PASS embed[prev]) accumulates the rows of the recently emitted tokens
                    # (state only moves on an emission); layer 1 passes its input through (forget gates shut)
                    w[DEC:2 * DEC] += np.float32(np.log(STATIONARY_MEMORY / (1.0 - STATIONARY_MEMORY)) if l == 0 else -4.0)
                    w[3 * DEC:4 * DEC] += 4.0  # output gates open
            yield name, w, False
    we = normal(s, "joint.enc.weight", (JOINT, D_MODEL), (1.0 / D_MODEL) ** 0.5)
    be = normal(s, "joint.enc.bias", (JOINT,), 0.1)
    if decoder == "stationary":
        # fold the encoder output's time-average (a measured constant of this random model, tests/golden/calibrate_bench_decoder.py) into the
        # bias and amplify what is left: the joint then sees the part of the encoder output that follows the audio
        m = encoder_mean(seed, n_layers, res_scale)
        if m is not None:
            we = we * np.float32(STATIONARY_ENC_GAIN)
            be = (be - we @ m).astype(np.float32)
    yield "joint.enc.weight", we, False
    yield "joint.enc.bias", be, False
    wp = normal(s, "joint.pred.weight", (JOINT, DEC), PRED_GAIN / DEC ** 0.5)
    if decoder == "stationary":  # pred = -SUPPRESS * (the previous token's joint_net row) + a small random part
        wp = wp * np.float32(STATIONARY_NOISE) - np.float32(STATIONARY_SUPPRESS / STATIONARY_PASS ** 2) * np.eye(JOINT, dtype=np.float32)
    yield "joint.pred.weight", wp, False
    yield "joint.pred.bias", normal(s, "joint.pred.bias", (JOINT,), 0.1), False
    # Scale-free blank: logit_blank = BLANK_GAIN * mean(relu(hidden)) grows with the hidden norm exactly like the
    # largest non-blank logit does, so the per-step blank probability is about the same in every decoder state
    # (a plain blank bias gave either all-blank or runaway 10-symbol loops once the LSTM state left zero).
    wo[VOCAB - 1] = (BLANK_GAIN if blank_gain is None else blank_gain) / JOINT
    yield "joint.joint_net.2.weight", wo, False
    ob = normal(s, "joint.joint_net.2.bias", (VOCAB,), 0.1)
    ob[VOCAB - 1] = 0.0
    yield "joint.joint_net.2.bias", ob, False


# The random decoder is bistable in this gain (the LSTM state feeds back into the joint): at 5.9 the reference emits
# 3-5 tokens per encoder frame (heavy multi-symbol decode: what the parity fixtures use); from about 6.0 (24 layers)
# it is blank-dominated, ~1.05 symbol steps per frame (real speech: ~1.3). The benchmark passes blank_gain=6.3 so that
# none of its concurrent streams falls into the runaway regime, and reports the measured steps per frame.
BLANK_GAIN = 5.9
PRED_GAIN = 4.0
LSTM_BIAS = 0.5
# DECODERS. "bistable": the plain random LSTM above (parity fixtures; token-heavy at gain 5.9, all-blank from ~6.0: once a frame
# emits, the hidden vector of its next symbol step is nearly the same, so it emits again up to the 10-symbol cap).
# "stationary" (bench.py): the architecture and every tensor shape are unchanged, but the prediction network is initialised so
# that it does what a trained RNNT predictor does after an emission - suppress the token just emitted. embed[v] = the
# joint_net row of v; both LSTM layers pass their input through (input / output gates open, forget gates shut, g = tanh(PASS x),
# recurrent weights ~0), joint.pred = -(SUPPRESS / PASS^2) I + noise: pred(prev) ~ -SUPPRESS * W_out[prev], which lowers the
# logit of prev by ~SUPPRESS |W_out[prev]|^2 / 2 and leaves the other tokens alone. Emissions then follow the encoder frame
# (through joint.enc), a second symbol on the same frame needs the runner-up to beat the blank as well, and the statistics are
# the same in every decoder state (no drift into all-blank or 10-symbols-per-frame regimes). The blank gain sets the rate:
# tests/golden/calibrate_bench_decoder.py. All other weights keep a random part (NOISE) so no product is trivially zero.
STATIONARY_NOISE = 0.1
STATIONARY_PASS = 0.5
STATIONARY_SUPPRESS = 1.0
STATIONARY_MEMORY = 0.8
STATIONARY_ENC_GAIN = 4.0


def calibration_key(seed: int, n_layers: int, res_scale: float) -> str:
    return f"enc_mean_s{seed}_L{n_layers}_rs{res_scale:g}"


def encoder_mean(seed: int, n_layers: int, res_scale: float):
    """Measured time-average of the encoder output of the model (seed, n_layers, res_scale), or None if it was never measured."""
    from pathlib import Path
    p = Path(__file__).resolve().parent / "decoder_calibration.npz"
    if not p.exists():
        return None
    d = np.load(p)
    k = calibration_key(seed, n_layers, res_scale)
    return d[k].astype(np.float32) if k in d.files else None


def _w_str(f, sbytes: bytes):
    f.write(struct.pack("<Q", len(sbytes)))
    f.write(sbytes)


def write_gguf(path: str, seed: int = 1234, n_layers: int = 24, variant: str = "q8_0", blank_gain: float | None = None,
               multilingual: bool = False, decoder: str = "bistable", legacy_vocab: bool = False, res_scale: float = 1.0) -> dict:
    """legacy_vocab: write the vocabulary as the old `tokenizer.vocab` blob of fixed 8-byte NUL-padded pieces instead of the
    `tokenizer.vocab_list` string array (the loader's fallback branch, /root/reference/src/nemo-ggml.cpp:156-165)."""
    assert variant in ("q8_0", "f32", "f32dq", "f16", "q4_0") and decoder in ("bistable", "stationary")
    vocab = [("▁" if i % 3 == 0 else "") + f"t{i}" for i in range(VOCAB - 1)]
    hparams = [("nemo.n_mels", N_MELS), ("nemo.d_model", D_MODEL), ("nemo.n_heads", N_HEADS), ("nemo.d_head", D_HEAD),
               ("nemo.d_ff", D_FF), ("nemo.n_layers", n_layers), ("nemo.kernel_size", KERNEL), ("nemo.vocab_size", VOCAB),
               ("nemo.decoder_dim", DEC), ("nemo.joint_dim", JOINT), ("nemo.subsampling_factor", SUBS),
               ("nemo.att_left_context", LEFT), ("nemo.num_prompts", NUM_PROMPTS if multilingual else 0)]
    infos, off = [], 0
    blobs = []
    import itertools
    for name, arr, quantisable in itertools.chain(tensor_list(seed, n_layers, blank_gain, decoder, res_scale), prompt_tensors(seed) if multilingual else ()):
        arr = np.ascontiguousarray(arr, dtype=np.float32)
        if quantisable and variant == "q8_0":
            data, typ = q8_0_bytes(arr), GGML_Q8_0
        elif quantisable and variant == "q4_0":
            data, typ = q4_0_bytes(arr), GGML_Q4_0
        elif quantisable and variant == "f16":
            data, typ = arr.astype(np.float16).tobytes(), GGML_F16
        elif quantisable and variant == "f32dq":
            data, typ = dequantize_q8_0(arr).astype(np.float32).tobytes(), GGML_F32
        else:
            data, typ = arr.tobytes(), GGML_F32
        off = (off + ALIGN - 1) // ALIGN * ALIGN
        infos.append((name, list(reversed(arr.shape)), typ, off, len(data)))
        blobs.append(data)
        off += len(data)
    with open(path, "wb") as f:
        f.write(b"GGUF")
        f.write(struct.pack("<I", 3))
        f.write(struct.pack("<q", len(infos)))
        f.write(struct.pack("<q", len(hparams) + 3 + (2 if multilingual else 0)))
        for k, v in (("general.architecture", "nemo"), ("general.name", f"random-init-{variant}-L{n_layers}-s{seed}")):
            _w_str(f, k.encode()); f.write(struct.pack("<I", 8)); _w_str(f, v.encode())
        if legacy_vocab:  # one string: vocab_size - 1 pieces of 8 bytes, NUL-padded (pieces longer than 8 bytes are cut)
            _w_str(f, b"tokenizer.vocab"); f.write(struct.pack("<I", 8))
            _w_str(f, b"".join(t.encode("utf-8")[:8].ljust(8, b"\0") for t in vocab))
        else:
            _w_str(f, b"tokenizer.vocab_list"); f.write(struct.pack("<I", 9)); f.write(struct.pack("<I", 8))
            f.write(struct.pack("<Q", len(vocab)))
            for t in vocab:
                _w_str(f, t.encode("utf-8"))
        for k, v in hparams:
            _w_str(f, k.encode()); f.write(struct.pack("<I", 4)); f.write(struct.pack("<I", v))
        if multilingual:  # parallel arrays: string[] codes, i32[] ids
            _w_str(f, b"nemo.prompt_langs"); f.write(struct.pack("<I", 9)); f.write(struct.pack("<I", 8))
            f.write(struct.pack("<Q", len(PROMPT_DICT)))
            for code, _ in PROMPT_DICT:
                _w_str(f, code.encode())
            _w_str(f, b"nemo.prompt_ids"); f.write(struct.pack("<I", 9)); f.write(struct.pack("<I", 5))
            f.write(struct.pack("<Q", len(PROMPT_DICT)))
            for _, pid in PROMPT_DICT:
                f.write(struct.pack("<i", pid))
        for name, shape, typ, o, _ in infos:
            _w_str(f, name.encode()); f.write(struct.pack("<I", len(shape)))
            for dim in shape:
                f.write(struct.pack("<q", dim))
            f.write(struct.pack("<i", typ)); f.write(struct.pack("<Q", o))
        pos = f.tell()
        f.write(b"\0" * ((pos + ALIGN - 1) // ALIGN * ALIGN - pos))
        start = f.tell()
        for (name, shape, typ, o, n), data in zip(infos, blobs):
            cur = f.tell() - start
            if o > cur:
                f.write(b"\0" * (o - cur))
            f.write(data)
        size = f.tell()
    return {"path": path, "bytes": size, "tensors": len(infos), "variant": variant, "n_layers": n_layers, "seed": seed}


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--seed", type=int, default=1234)
    ap.add_argument("--layers", type=int, default=24)
    ap.add_argument("--variant", default="q8_0", choices=("q8_0", "f32", "f32dq", "f16", "q4_0"))
    ap.add_argument("--blank-gain", type=float, default=None)
    ap.add_argument("--decoder", default="bistable", choices=("bistable", "stationary"))
    ap.add_argument("--legacy-vocab", action="store_true")
    ap.add_argument("--res-scale", type=float, default=1.0)
    a = ap.parse_args(argv)
    print(write_gguf(a.out, a.seed, a.layers, a.variant, a.blank_gain, decoder=a.decoder, legacy_vocab=a.legacy_vocab, res_scale=a.res_scale))


if __name__ == "__main__":
    sys.exit(main())
