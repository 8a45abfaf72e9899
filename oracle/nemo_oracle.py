"""oracle/nemo_oracle.py — TEST INFRASTRUCTURE, not product.

A numpy (CPU) restatement of the reference's streaming hot path
    s16le PCM -> log-mel -> cache-aware FastConformer encoder -> RNNT LSTM/joint -> greedy tokens
used ONLY by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg as the checker. Nothing under
nemotron-asr.cpp_b200/ imports it; the product path is the CUDA library and fails loudly without it.

Parity pin: this restatement is itself checked against outputs of the UNMODIFIED reference compiled from
/root/reference by oracle/Makefile (oracle/_ref/<isa>/ref_harness), committed as fixtures under tests/golden/
(generator: tests/golden/make_golden.py). See tests/test_oracle_vs_reference.py.

Every function cites the reference file:line it follows (paths relative to /root/reference).
All arithmetic is fp32 unless noted; Q8_0 weight matmuls emulate ggml-CPU's block-int8 arithmetic:
activations re-quantised to Q8_0 per 32 elements, integer dot per block, scaled by fp16(d_w)*fp16(d_x).
"""
from __future__ import annotations

import struct

import numpy as np

F32 = np.float32

# ---------------------------------------------------------------------------------------------
# GGUF v3 reader (format: scripts/convert_to_gguf.py:491-536; reader side src/nemo-ggml.cpp:94-283)
# ---------------------------------------------------------------------------------------------
_GGUF_SCALAR = {0: "<B", 1: "<b", 2: "<H", 3: "<h", 4: "<I", 5: "<i", 6: "<f", 7: "<?", 10: "<Q", 11: "<q", 12: "<d"}


class Q8Weight:
    """Q8_0 matrix [N, K]: d fp16 [N, K/32], q int8 [N, K] (ggml/src/ggml-common.h:251-255)."""

    def __init__(self, d: np.ndarray, q: np.ndarray):
        self.d, self.q = d, q
        self.shape = q.shape

    def dequant(self) -> np.ndarray:  # ggml/src/ggml-quants.c:553-567
        n, k = self.q.shape
        return (self.q.reshape(n, k // 32, 32).astype(F32) * self.d.astype(F32)[:, :, None]).reshape(n, k)


def load_gguf(path: str):
    """Returns (kv dict, tensors dict name -> np.ndarray fp32 in PyTorch shape order | Q8Weight)."""
    buf = np.fromfile(path, dtype=np.uint8)
    mv = memoryview(buf)
    pos = 0

    def rd(fmt):
        nonlocal pos
        v = struct.unpack_from(fmt, mv, pos)[0]
        pos += struct.calcsize(fmt)
        return v

    def rd_str():
        nonlocal pos
        n = rd("<Q")
        s = bytes(mv[pos:pos + n])
        pos += n
        return s

    def rd_val(t):
        if t == 8:
            return rd_str().decode("utf-8", errors="replace")
        if t == 9:
            et, cnt = rd("<I"), rd("<Q")
            return [rd_val(et) for _ in range(cnt)]
        return rd(_GGUF_SCALAR[t])

    assert bytes(mv[0:4]) == b"GGUF", "not a GGUF file"
    pos = 4
    version, n_tensors, n_kv = rd("<I"), rd("<q"), rd("<q")
    assert version == 3
    kv = {}
    for _ in range(n_kv):
        k = rd_str().decode()
        kv[k] = rd_val(rd("<I"))
    infos = []
    for _ in range(n_tensors):
        name = rd_str().decode()
        nd = rd("<I")
        dims = [rd("<q") for _ in range(nd)]
        typ, off = rd("<i"), rd("<Q")
        infos.append((name, dims, typ, off))
    align = kv.get("general.alignment", 32)
    start = (pos + align - 1) // align * align
    tensors = {}
    for name, dims, typ, off in infos:
        shape = tuple(reversed(dims))  # GGUF stores the fastest dim first
        n = int(np.prod(shape))
        o = start + off
        if typ == 0:
            tensors[name] = buf[o:o + 4 * n].view(F32).reshape(shape).copy()
        elif typ == 1:
            tensors[name] = buf[o:o + 2 * n].view(np.float16).astype(F32).reshape(shape)
        elif typ == 8:
            blk = buf[o:o + n // 32 * 34].view(np.dtype([("d", np.float16), ("q", np.int8, 32)]))
            rows, k = int(np.prod(shape[:-1])), shape[-1]
            tensors[name] = Q8Weight(blk["d"].reshape(rows, k // 32).copy(), blk["q"].reshape(rows, k).copy())
        else:
            raise ValueError(f"unsupported ggml type {typ} for {name}")
    return kv, tensors


# ---------------------------------------------------------------------------------------------
# a1. log-mel preprocessor (src/preprocessor.cpp)
# ---------------------------------------------------------------------------------------------
def _libm_f32(fn: str, x: np.ndarray) -> np.ndarray:
    """Apply the C library's float function (sinf/cosf/logf) elementwise: the reference's tables and logs come
    from the host libm, and a 1-ulp twiddle difference shows up at 1e-4 in quiet mel bins."""
    import ctypes
    import ctypes.util
    lib = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    f = getattr(lib, fn)
    f.restype, f.argtypes = ctypes.c_float, [ctypes.c_float]
    flat = np.asarray(x, F32).ravel()
    return np.array([f(float(v)) for v in flat], F32).reshape(np.shape(x))


class Preprocessor:
    """Streaming PCM -> log-mel. Mirrors the operation ORDER of the reference so fp32 results agree to an ulp:
    pre-emphasis with carried sample (:349-356), 256-zero centre pad and <512-sample tail (:212-225, :389-393),
    window multiply (:184-194), radix-2 DIT FFT with bit reversal and the sinf/cosf twiddle table
    (:80-110, :113-161), |X| via sqrtf then re-squared (:200-203, :363-368), sequential 257-term mel sum and
    log(x + 2^-24) (:374-383)."""

    N_FFT, HOP, N_MELS, N_BINS = 512, 160, 128, 257

    def __init__(self, fb: np.ndarray, window: np.ndarray):
        assert fb.size == 128 * 257 and window.size == 400  # :282-294
        self.fb = fb.reshape(128, 257).astype(F32)
        self.win = np.zeros(512, F32)
        self.win[56:456] = window.astype(F32)  # :297-299
        i = np.arange(512)
        theta = (F32(2.0) * F32(np.pi) * i.astype(F32)) / F32(512)  # :87 float arithmetic
        self.cos, self.sin = _libm_f32("cosf", theta), _libm_f32("sinf", theta)  # the host libm's sinf/cosf, :88-89
        rev = np.zeros(512, np.int64)
        for b in range(9):
            rev |= ((i >> b) & 1) << (8 - b)
        self.rev = rev  # :96-105
        self.buf = np.zeros(256, F32)  # :220-221
        self.last = F32(0.0)

    def _fft_pow(self, frames: np.ndarray) -> np.ndarray:
        """frames [F,512] windowed -> power [F,257]."""
        re = np.zeros_like(frames)
        re[:, self.rev] = frames  # real[bit_rev[i]] = frame[i]
        im = np.zeros_like(frames)
        m = 2
        while m <= 512:
            m2, step = m // 2, 512 // m
            j = np.arange(m2)
            wr = self.cos[j * step]
            wi = -self.sin[j * step]
            k = np.arange(0, 512, m)
            i1 = (k[:, None] + j[None, :]).ravel()
            i2 = i1 + m2
            wr_ = np.tile(wr, k.size)
            wi_ = np.tile(wi, k.size)
            r2, q2 = re[:, i2], im[:, i2]
            tr = wr_ * r2 - wi_ * q2  # two rounded products, one rounded subtract (no FMA: Makefile:7 has no -march)
            ti = wr_ * q2 + wi_ * r2
            r1, q1 = re[:, i1], im[:, i1]
            re[:, i2] = r1 - tr
            im[:, i2] = q1 - ti
            re[:, i1] = r1 + tr
            im[:, i1] = q1 + ti
            m <<= 1
        re, im = re[:, :257], im[:, :257]
        mag = np.sqrt(re * re + im * im)
        return mag * mag

    def process(self, pcm: np.ndarray) -> np.ndarray:
        """int16 [n] -> float32 [frames,128]."""
        n = int(pcm.size)
        if n == 0:
            return np.zeros((0, 128), F32)
        cur = pcm.astype(F32) * F32(1.0 / 32768.0)
        prev = np.concatenate([[self.last], cur[:-1]]).astype(F32)
        self.last = cur[-1]
        self.buf = np.concatenate([self.buf, cur - F32(0.97) * prev]).astype(F32)
        avail = self.buf.size
        nf = 0 if avail < 512 else (avail - 512 + 160) // 160  # :320-328
        if nf == 0:
            return np.zeros((0, 128), F32)
        idx = np.arange(nf)[:, None] * 160 + np.arange(512)[None, :]
        power = self._fft_pow(self.buf[idx] * self.win[None, :])
        mel = np.zeros((nf, 128), F32)
        for k in range(257):  # sequential fp32 accumulation over bins, :377-379
            mel += self.fb[None, :, k] * power[:, k:k + 1]
        self.buf = self.buf[nf * 160:]
        return np.log(mel + F32(5.960464477539063e-8)).astype(F32)


# ---------------------------------------------------------------------------------------------
# Matmuls the way ggml-CPU does them
# ---------------------------------------------------------------------------------------------
def quantize_act_q8_0(x: np.ndarray):
    """fp32 [M,K] -> (d fp32-of-fp16 [M,K/32], q float32-valued ints [M,K]).
    ggml/src/ggml-cpu/arch/x86/quants.c:331-346 (AVX2): d = amax/127 stored as fp16, q = rne(x * (127/amax))."""
    m, k = x.shape
    b = x.reshape(m, k // 32, 32).astype(F32)
    amax = np.max(np.abs(b), axis=2)
    d = (amax / F32(127.0)).astype(F32)
    inv = np.where(amax != 0, F32(127.0) / np.where(amax != 0, amax, F32(1)), F32(0)).astype(F32)
    q = np.rint(b * inv[:, :, None]).astype(F32)
    return d.astype(np.float16).astype(F32), q.reshape(m, k)


def mm(x: np.ndarray, w) -> np.ndarray:
    """x [M,K] @ w[N,K]^T -> [M,N]. Q8Weight: block-int8 arithmetic of ggml/src/ggml-cpu/quants.c:451-478
    (sum over blocks of (int dot) * fp16(d_w) * fp16(d_x)); ndarray: plain fp32."""
    if isinstance(w, Q8Weight):
        m, k = x.shape
        n = w.shape[0]
        dx, qx = quantize_act_q8_0(x)
        nb = k // 32
        qxb = qx.reshape(m, nb, 32).transpose(1, 0, 2)  # [nb,M,32]
        qwb = w.q.reshape(n, nb, 32).transpose(1, 2, 0).astype(F32)  # [nb,32,N]
        s = np.matmul(qxb, qwb)  # [nb,M,N]; exact: |sum| <= 32*127*127 < 2^24
        sc = dx.T[:, :, None] * w.d.astype(F32).T[:, None, :]  # [nb,M,N]
        return np.sum(s * sc, axis=0, dtype=F32)
    return x.astype(F32) @ w.astype(F32).T


def layer_norm(x, g, b, eps=F32(1e-5)):
    """src/nemo-stream.cpp:580-591 over ggml_norm (ggml/src/ggml-cpu/ops.cpp:3698-3741): mean, centred variance."""
    mean = x.mean(axis=-1, keepdims=True, dtype=F32)
    v = x - mean
    var = (v * v).mean(axis=-1, keepdims=True, dtype=F32)
    return (v * (F32(1.0) / np.sqrt(var + eps))) * g + b


def silu(x):
    return (x / (F32(1.0) + np.exp(-x))).astype(F32)


def sigmoid(x):
    return (F32(1.0) / (F32(1.0) + np.exp(-x))).astype(F32)


def f16r(x):
    return x.astype(np.float16).astype(F32)


# ---------------------------------------------------------------------------------------------
# a2. conv subsampling (src/nemo-ggml.cpp:897-1029), per chunk
# ---------------------------------------------------------------------------------------------
def _pad_causal(a):  # [C,H,W] -> pad 2 before / 1 after on H (time) and W (mel), :905-913
    return np.pad(a, ((0, 0), (2, 1), (2, 1)))


def _patches(a):
    """[C,H,W] padded -> [C, Ho, Wo, 3, 3] for a 3x3 stride-2 conv."""
    c, h, w = a.shape
    ho, wo = (h - 3) // 2 + 1, (w - 3) // 2 + 1
    out = np.empty((c, ho, wo, 3, 3), F32)
    for kh in range(3):
        for kw in range(3):
            out[:, :, :, kh, kw] = a[:, kh:kh + 2 * ho:2, kw:kw + 2 * wo:2]
    return out


def conv_subsampling(mel: np.ndarray, t: dict) -> np.ndarray:
    """mel [CM,128] -> [T3,1024] (before the drop of 2 frames). conv0/pw3/pw6 multiply fp16-rounded operands
    with fp32 accumulation (ggml_conv_2d = im2col(F16) + mul_mat, ggml/src/ggml.c:4658-4680); depthwise and the
    output Linear are fp32 (:944, :1020)."""
    pe = "encoder.pre_encode."
    a = _pad_causal(mel[None].astype(F32))  # [1,H,W], H=time
    p = _patches(f16r(a))[0]  # [Ho,Wo,3,3]
    w0 = f16r(t[pe + "conv.0.weight"]).reshape(256, 9)
    a = np.einsum("hwk,ck->chw", p.reshape(p.shape[0], p.shape[1], 9), w0).astype(F32) + t[pe + "conv.0.bias"][:, None, None]
    a = np.maximum(a, 0)
    for dw, pw in ((2, 3), (5, 6)):
        p = _patches(_pad_causal(a))  # [C,Ho,Wo,3,3]
        a = np.einsum("chwij,cij->chw", p, t[pe + f"conv.{dw}.weight"].reshape(256, 3, 3)).astype(F32) + t[pe + f"conv.{dw}.bias"][:, None, None]
        wp = f16r(t[pe + f"conv.{pw}.weight"].reshape(256, 256))
        a = np.einsum("ihw,oi->ohw", f16r(a), wp).astype(F32) + t[pe + f"conv.{pw}.bias"][:, None, None]
        a = np.maximum(a, 0)
    c, h, w = a.shape
    feat = a.transpose(1, 0, 2).reshape(h, c * w)  # feat[c*17+w], :1014-1017
    return mm(feat, t[pe + "out.weight"]) + t[pe + "out.bias"]


def pos_emb_row(p: np.ndarray, d_model=1024) -> np.ndarray:
    """Sinusoid rows for relative positions p (src/nemo-ggml.cpp:17-32): [2i]=sin(p*w_i), [2i+1]=cos(p*w_i)."""
    i = np.arange(0, d_model, 2).astype(F32)
    div = np.exp(-i * F32(np.log(F32(10000.0))) / F32(d_model)).astype(F32)
    ang = (p.astype(F32)[:, None] * div[None, :]).astype(F32)
    out = np.empty((p.size, d_model), F32)
    out[:, 0::2] = np.sin(ang)
    out[:, 1::2] = np.cos(ang)
    return out


# ---------------------------------------------------------------------------------------------
# a4-a12. streaming session (src/nemo-stream.cpp), one stream
# ---------------------------------------------------------------------------------------------
class StreamOracle:
    """Restates nemo_stream_context + process_mel_chunk_streaming + decode_one_step for one stream.
    State per SURVEY Appendix B: mel overlap buffer, K/V caches [L,70,1024], conv cache [L,8,1024],
    cache_valid_len, LSTM h/c [2,640], prev token."""

    MAX_SYMBOLS = 10  # src/nemo-stream.cpp:849

    def __init__(self, kv: dict, tensors: dict, right_context: int = 0):
        self.t = tensors
        self.L = int(kv["nemo.n_layers"])
        self.H, self.DH, self.D = int(kv["nemo.n_heads"]), int(kv["nemo.d_head"]), int(kv["nemo.d_model"])
        self.left = int(kv["nemo.att_left_context"])
        self.vocab = int(kv["nemo.vocab_size"])
        self.blank = self.vocab - 1
        self.rc = right_context
        # multilingual checkpoints: language prompt (src/nemo-ggml.cpp:459-462 default 101 = "auto"; set_language :735-749)
        self.num_prompts = int(kv.get("nemo.num_prompts", 0))
        self.prompt_index = 101 if self.num_prompts > 0 else -1
        self.prompt_dict = dict(zip(kv.get("nemo.prompt_langs", []), kv.get("nemo.prompt_ids", [])))
        self.chunk_mel = 9 + 8 * (1 + right_context)  # src/nemo-stream.h:65-72
        self.shift = 8 * (1 + right_context)  # :76-81
        self.valid_out = 1 + right_context  # :98-100
        self.pp = Preprocessor(tensors["preprocessor.featurizer.fb"], tensors["preprocessor.featurizer.window"])
        self.mel_buf = np.zeros((9, 128), F32)  # src/nemo-stream.cpp:73-74
        self.kc = np.zeros((self.L, self.left, self.D), F32)
        self.vc = np.zeros((self.L, self.left, self.D), F32)
        self.cc = np.zeros((self.L, 8, self.D), F32)
        self.valid = 0
        self.h = np.zeros((2, 640), F32)
        self.c = np.zeros((2, 640), F32)
        self.prev = self.blank
        self.tokens: list[int] = []
        self.token_frames: list[int] = []  # encoder frame (80 ms) each token was emitted at, counted over the decoded frames of the stream
        self.frames_done = 0
        self.enc_chunks: list[np.ndarray] = []  # all T frames of every chunk (before valid_out truncation)
        self.chunk_tokens: list[int] = []
        self.decode_iters = 0
        self._pos_cache = {}

    # -- encoder -------------------------------------------------------------------------------
    def _pos_proj(self, l: int, T: int) -> np.ndarray:
        """linear_pos(PE(p)) for p = 70+T-1 ... -(T-1)+... (src/nemo-stream.cpp:168-177,516): returns dict-like
        array indexed by p + (T-1), p in [-(T-1), left+T-1]."""
        key = (l, T)
        if key not in self._pos_cache:
            p = np.arange(-(T - 1), self.left + T)
            self._pos_cache[key] = mm(pos_emb_row(p, self.D), self.t[f"encoder.layers.{l}.self_attn.linear_pos.weight"])
        return self._pos_cache[key]

    def _layer(self, l: int, x: np.ndarray) -> np.ndarray:
        t, p = self.t, f"encoder.layers.{l}."
        T = x.shape[0]
        # FFN1 (:631-634, :593-603)
        y = layer_norm(x, t[p + "norm_feed_forward1.weight"], t[p + "norm_feed_forward1.bias"])
        y = mm(silu(mm(y, t[p + "feed_forward1.linear1.weight"])), t[p + "feed_forward1.linear2.weight"])
        x = x + y * F32(0.5)
        # cached rel-pos MHA (:463-573)
        y = layer_norm(x, t[p + "norm_self_att.weight"], t[p + "norm_self_att.bias"])
        q = mm(y, t[p + "self_attn.linear_q.weight"])
        k = mm(y, t[p + "self_attn.linear_k.weight"])
        v = mm(y, t[p + "self_attn.linear_v.weight"])
        K = np.concatenate([self.kc[l], k], axis=0)  # [70+T,1024]
        V = np.concatenate([self.vc[l], v], axis=0)
        if self.left > 0:
            self.kc[l], self.vc[l] = K[-self.left:], V[-self.left:]  # :504-512
        P = self._pos_proj(l, T)  # index p + T - 1
        kv_len = self.left + T
        qh = q.reshape(T, self.H, self.DH)
        qu = qh + t[p + "self_attn.pos_bias_u"][None]
        qv = qh + t[p + "self_attn.pos_bias_v"][None]
        Kh = K.reshape(kv_len, self.H, self.DH)
        Vh = V.reshape(kv_len, self.H, self.DH)
        Ph = P.reshape(-1, self.H, self.DH)
        content = np.einsum("ihd,jhd->hij", qu, Kh).astype(F32)
        # rel-shift (:419-461) reduces to position index p = left + i - j
        i_idx = np.arange(T)[:, None]
        j_idx = np.arange(kv_len)[None, :]
        pidx = (self.left + i_idx - j_idx) + (T - 1)
        posfull = np.einsum("ihd,phd->hip", qv, Ph).astype(F32)  # [H,T,npos]
        pos = np.take_along_axis(posfull, np.broadcast_to(pidx[None], (self.H, T, kv_len)), axis=2)
        s = (content + pos) * F32(1.0 / np.sqrt(F32(self.DH)))
        mask = np.where(np.arange(kv_len) < self.left - self.valid, F32(-1e9), F32(0))  # :1035-1045
        s = s + mask[None, None, :]
        s = s - s.max(axis=2, keepdims=True)
        e = np.exp(s)
        a = (e / e.sum(axis=2, keepdims=True)).astype(F32)
        ctx = np.einsum("hij,jhd->ihd", a, Vh).astype(F32).reshape(T, self.D)
        x = x + mm(ctx, t[p + "self_attn.linear_out.weight"])
        # conv module (:645-679, :336-412)
        y = layer_norm(x, t[p + "norm_conv.weight"], t[p + "norm_conv.bias"])
        z = mm(y, t[p + "conv.pointwise_conv1.weight"])
        g = z[:, :self.D] * sigmoid(z[:, self.D:])
        G = np.concatenate([self.cc[l], g], axis=0)  # [8+T,1024]
        wdw = t[p + "conv.depthwise_conv.weight"]  # [9,1024]
        d = np.zeros((T, self.D), F32)
        for kk in range(9):
            d = d + G[kk:kk + T] * wdw[kk][None, :]
        self.cc[l] = G[-8:]
        d = silu(layer_norm(d, t[p + "conv.batch_norm.weight"], t[p + "conv.batch_norm.bias"]))
        x = x + mm(d, t[p + "conv.pointwise_conv2.weight"])
        # FFN2, final LN (:682-687)
        y = layer_norm(x, t[p + "norm_feed_forward2.weight"], t[p + "norm_feed_forward2.bias"])
        y = mm(silu(mm(y, t[p + "feed_forward2.linear1.weight"])), t[p + "feed_forward2.linear2.weight"])
        x = x + y * F32(0.5)
        return layer_norm(x, t[p + "norm_out.weight"], t[p + "norm_out.bias"])

    def encode_chunk(self, mel: np.ndarray) -> np.ndarray:
        x = conv_subsampling(mel, self.t)[2:]  # drop_extra_pre_encoded = 2, src/nemo-stream.cpp:154-162
        for l in range(self.L):
            x = self._layer(l, x)
        self.valid = min(self.valid + x.shape[0], self.left)  # :1085
        if self.num_prompts > 0:  # build_prompt_fusion, src/nemo-ggml.cpp:1087-1105; one-hot fill src/nemo-stream.cpp:1047-1059
            idx = self.prompt_index if 0 <= self.prompt_index < self.num_prompts else 0
            onehot = np.zeros((x.shape[0], self.num_prompts), F32)
            onehot[:, idx] = 1.0
            cat = np.concatenate([x, onehot], axis=1)
            h = np.maximum(mm(cat, self.t["prompt_kernel.0.weight"]) + self.t["prompt_kernel.0.bias"], 0.0).astype(F32)
            x = (mm(h, self.t["prompt_kernel.2.weight"]) + self.t["prompt_kernel.2.bias"]).astype(F32)
        return x

    def set_language(self, lang: str) -> bool:
        """nemo_stream_set_language, src/nemo-stream.cpp:735-749."""
        if self.num_prompts <= 0 or lang not in self.prompt_dict:
            return False
        self.prompt_index = int(self.prompt_dict[lang])
        return True

    # -- decoder -------------------------------------------------------------------------------
    def _lstm(self, l: int, x, h, c):  # src/nemo-ggml.cpp:580-619, gate order i,f,g,o
        d = "decoder.prediction.dec_rnn.lstm."
        g = (mm(x[None], self.t[d + f"weight_ih_l{l}"])[0] + mm(h[None], self.t[d + f"weight_hh_l{l}"])[0]
             + self.t[d + f"bias_ih_l{l}"] + self.t[d + f"bias_hh_l{l}"])
        i, f, gg, o = g[:640], g[640:1280], g[1280:1920], g[1920:]
        c2 = sigmoid(f) * c + sigmoid(i) * np.tanh(gg)
        return (sigmoid(o) * np.tanh(c2)).astype(F32), c2.astype(F32)

    def decode_frame(self, enc: np.ndarray) -> list[int]:
        """src/nemo-stream.cpp:840-930: <=10 symbols; first-max argmax; blank leaves the state untouched."""
        t = self.t
        e = mm(enc[None], t["joint.enc.weight"])[0] + t["joint.enc.bias"]
        out = []
        for _ in range(self.MAX_SYMBOLS):
            self.decode_iters += 1
            emb = t["decoder.prediction.embed.weight"][self.prev]
            h0, c0 = self._lstm(0, emb, self.h[0], self.c[0])
            h1, c1 = self._lstm(1, h0, self.h[1], self.c[1])
            pr = mm(h1[None], t["joint.pred.weight"])[0] + t["joint.pred.bias"]
            logits = mm(np.maximum(e + pr, 0)[None], t["joint.joint_net.2.weight"])[0] + t["joint.joint_net.2.bias"]
            tok = int(np.argmax(logits))  # first maximum, like the strict '>' scan at :899-906
            self.last_logits = logits
            if tok == self.blank:
                break
            out.append(tok)
            self.prev = tok
            self.h = np.stack([h0, h1])
            self.c = np.stack([c0, c1])
        return out

    # -- driver --------------------------------------------------------------------------------
    def _step(self, mel: np.ndarray, valid_out: int):
        enc = self.encode_chunk(mel)
        self.enc_chunks.append(enc)
        n0 = len(self.tokens)
        nf = min(enc.shape[0], valid_out)
        for f in range(nf):
            new = self.decode_frame(enc[f])
            self.tokens.extend(new)
            self.token_frames.extend([self.frames_done + f] * len(new))
        self.frames_done += nf
        self.chunk_tokens.append(len(self.tokens) - n0)

    def feed(self, pcm: np.ndarray):
        """nemo_stream_process_incremental, src/nemo-stream.cpp:1145-1206."""
        mel = self.pp.process(pcm)
        self.mel_buf = np.concatenate([self.mel_buf, mel], axis=0)
        while self.mel_buf.shape[0] >= self.chunk_mel:
            self._step(self.mel_buf[:self.chunk_mel], self.valid_out)
            self.mel_buf = self.mel_buf[self.shift:]

    def finalize(self):
        """nemo_stream_finalize, src/nemo-stream.cpp:1217-1293: zero-pad the tail, keep floor(new/8) frames."""
        total = self.mel_buf.shape[0]
        if total > 9:
            n_valid = (total - 9) // 8
            if n_valid > 0:
                if total < self.chunk_mel:
                    self.mel_buf = np.concatenate([self.mel_buf, np.zeros((self.chunk_mel - total, 128), F32)])
                self._step(self.mel_buf[:self.chunk_mel], n_valid)


class OfflineOracle(StreamOracle):
    """The reference's offline path (nemo_encode, src/nemo-ggml.cpp:1592-1670): whole-utterance log-mel ->
    build_conv_subsampling with its causal zero padding, nothing dropped (:954-1029) -> 24 layers with FULL-sequence
    relative-position attention (build_rel_pos_mha :668-757: no cache, no mask, positions T-1 .. -(T-1); the rel-shift
    :625-662 picks position i - j for query i, key j) and the causal depthwise conv over zero left padding (:796-823) ->
    greedy_decode (:1233-1380) from a zero LSTM state, tokens stamped with their frame. That is the streaming layer
    with an empty left context, so it is restated as StreamOracle with left = 0 and zero conv cache."""

    def __init__(self, kv: dict, tensors: dict):
        super().__init__(kv, tensors, 0)
        self.left = 0
        self.kc = np.zeros((self.L, 0, self.D), F32)
        self.vc = np.zeros((self.L, 0, self.D), F32)

    def encode(self, mel: np.ndarray) -> np.ndarray:
        x = conv_subsampling(mel, self.t)
        for l in range(self.L):
            self.cc[l] = 0
            x = self._layer(l, x)
        if self.num_prompts > 0:
            idx = self.prompt_index if 0 <= self.prompt_index < self.num_prompts else 0
            onehot = np.zeros((x.shape[0], self.num_prompts), F32)
            onehot[:, idx] = 1.0
            cat = np.concatenate([x, onehot], axis=1)
            h = np.maximum(mm(cat, self.t["prompt_kernel.0.weight"]) + self.t["prompt_kernel.0.bias"], 0.0).astype(F32)
            x = (mm(h, self.t["prompt_kernel.2.weight"]) + self.t["prompt_kernel.2.bias"]).astype(F32)
        return x

    def transcribe(self, pcm: np.ndarray):
        """-> (encoder_out [T,1024], tokens, frames)"""
        mel = Preprocessor(self.t["preprocessor.featurizer.fb"], self.t["preprocessor.featurizer.window"]).process(pcm)
        enc = self.encode(mel)
        self.h[:] = 0
        self.c[:] = 0
        self.prev = self.blank
        tokens, frames = [], []
        for f in range(enc.shape[0]):
            new = self.decode_frame(enc[f])
            tokens += new
            frames += [f] * len(new)
        return enc, tokens, frames


def tokens_to_text(tokens, vocab):
    """src/nemo-ggml.cpp:1556-1583 (no timestamps): concatenate pieces, U+2581 -> space."""
    return "".join(vocab[t] for t in tokens if 0 <= t < len(vocab)).replace("▁", " ")


def tokens_to_text_timed(tokens, frames, vocab):
    """src/nemo-ggml.cpp:1556-1583 with timestamp_words: a piece that STARTS with U+2581 becomes ' {%.2f}' + rest, the
    time being timed_token::to_seconds (src/nemo-ggml.h:393-395): float(frame) * 1280 / 16000 in fp32. Only the leading
    U+2581 is replaced (the untimed helper above replaces all of them, which is the same for SentencePiece vocabularies)."""
    out = []
    for t, f in zip(tokens, frames):
        if not 0 <= t < len(vocab):
            continue
        piece = vocab[t]
        if piece.startswith("▁"):
            sec = F32(F32(F32(f) * F32(1280)) / F32(16000))
            out.append(" {%.2f}" % float(sec) + piece[1:])
        else:
            out.append(piece)
    return "".join(out)
