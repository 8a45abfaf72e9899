"""ctypes binding of libnemob200.so (include/nb200.h) plus a thin host-side mirror of the reference's stream API.

`Model` / `Stream` follow the reference's operator interface for the streaming path
(/root/reference/src/nemo-stream.h:271-326): `Stream.process_incremental(pcm)` == nemo_stream_process_incremental,
`Stream.finalize()` == nemo_stream_finalize, `Stream.tokens` == nemo_stream_get_tokens, same argument meaning and
error behaviour (empty string / exception with the library's message instead of nullptr + stderr).
The product path is the CUDA library only: importing this module without the built library, or loading a model
without a B200, raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "libnemob200.so"

FLAG_ACT_QUANT_AUTO, FLAG_ACT_QUANT_OFF, FLAG_ACT_QUANT_ON, FLAG_GEMM_SIMT = 0, 1, 2, 16
FLAG_NO_GRAPHS, FLAG_BF16 = 32, 64
TAP_MEL, TAP_ENC = 1, 2


class Info(C.Structure):
    _fields_ = [(n, C.c_int32) for n in (
        "n_mels", "d_model", "n_heads", "d_head", "d_ff", "n_layers", "kernel_size", "vocab_size", "blank_token",
        "decoder_dim", "joint_dim", "subsampling_factor", "att_left_context", "num_prompts", "weights_quantized",
        "act_quant", "max_streams", "device")]


class Stats(C.Structure):
    _fields_ = [(n, C.c_int64) for n in ("steps", "chunks", "frames", "tokens", "decode_rounds", "kernel_launches",
                                         "h2d_bytes", "d2h_bytes")] + \
               [(n, C.c_double) for n in ("audio_seconds", "gpu_seconds", "last_step_gpu_ms", "last_frontend_ms",
                                          "last_encoder_ms", "last_decoder_ms")]


# every symbol include/nb200.h declares: (restype, argtypes)
_I16P, _I32P, _F32P = C.POINTER(C.c_int16), C.POINTER(C.c_int32), C.POINTER(C.c_float)
_I64P = C.POINTER(C.c_int64)
SYMBOLS = {
    "nb200_last_error": (C.c_char_p, []),
    "nb200_version": (C.c_int, []),
    "nb200_model_load": (C.c_void_p, [C.c_char_p, C.c_int, C.c_int, C.c_int]),
    "nb200_gguf_check": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int]),
    "nb200_model_free": (None, [C.c_void_p]),
    "nb200_model_info": (C.c_int, [C.c_void_p, C.POINTER(Info)]),
    "nb200_model_token": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "nb200_mel_batch_noalloc": (C.c_int, [C.c_void_p, _I16P, C.c_int, C.c_int, _F32P]),
    "nb200_model_load_cached": (C.c_void_p, [C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int]),
    "nb200_model_cache_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "nb200_weight_cache_selftest": (C.c_int, [C.c_char_p]),
    "nb200_bench_mma_peak": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_double)]),
    "nb200_tokens_to_text": (C.c_int, [C.c_void_p, _I32P, C.c_int, C.c_char_p, C.c_int]),
    "nb200_tokens_to_text_timed": (C.c_int, [C.c_void_p, _I32P, _I64P, C.c_int, C.c_char_p, C.c_int]),
    "nb200_stream_create": (C.c_void_p, [C.c_void_p, C.c_int]),
    "nb200_stream_free": (None, [C.c_void_p]),
    "nb200_stream_reset": (C.c_int, [C.c_void_p]),
    "nb200_stream_set_language": (C.c_int, [C.c_void_p, C.c_char_p]),
    "nb200_model_set_language": (C.c_int, [C.c_void_p, C.c_char_p]),
    "nb200_stream_set_prompt_index": (C.c_int, [C.c_void_p, C.c_int]),
    "nb200_stream_feed": (C.c_int, [C.c_void_p, _I16P, C.c_int]),
    "nb200_stream_finalize": (C.c_int, [C.c_void_p]),
    "nb200_stream_ready_chunks": (C.c_int, [C.c_void_p]),
    "nb200_step": (C.c_int, [C.c_void_p]),
    "nb200_step_begin": (C.c_int, [C.c_void_p]),
    "nb200_step_end": (C.c_int, [C.c_void_p]),
    "nb200_feed_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p), _I32P, C.c_int]),
    "nb200_poll_batch": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.c_int, _I32P, C.c_int, _I32P]),
    "nb200_drain": (C.c_int, [C.c_void_p]),
    "nb200_encode_audio": (C.c_int, [C.c_void_p, _I16P, C.c_int, _I32P, _I64P, C.c_int, _F32P, C.c_int64, C.POINTER(C.c_int)]),
    "nb200_stream_poll_tokens": (C.c_int, [C.c_void_p, _I32P, C.c_int]),
    "nb200_stream_get_tokens": (C.c_int, [C.c_void_p, _I32P, C.c_int]),
    "nb200_stream_poll_timed": (C.c_int, [C.c_void_p, _I32P, _I64P, C.c_int]),
    "nb200_stream_get_token_frames": (C.c_int, [C.c_void_p, _I64P, C.c_int]),
    "nb200_stream_chunk_token_counts": (C.c_int, [C.c_void_p, _I32P, C.c_int]),
    "nb200_stream_chunk_margins": (C.c_int, [C.c_void_p, _F32P, C.c_int]),
    "nb200_stream_set_taps": (C.c_int, [C.c_void_p, C.c_int]),
    "nb200_stream_fetch_tap": (C.c_int64, [C.c_void_p, C.c_int, _F32P, C.c_int64]),
    "nb200_get_stats": (C.c_int, [C.c_void_p, C.POINTER(Stats)]),
    "nb200_timer_start": (C.c_int, [C.c_void_p]),
    "nb200_timer_stop": (C.c_int, [C.c_void_p, C.POINTER(C.c_double)]),
    "nb200_profile_enable": (C.c_int, [C.c_void_p, C.c_int]),
    "nb200_set_dup_family": (C.c_int, [C.c_void_p, C.c_char_p]),
    "nb200_profile_report": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int]),
    "nb200_mel_batch": (C.c_int, [C.c_void_p, _I16P, C.c_int, C.c_int, _F32P]),
    "nb200_debug_matmul": (C.c_int, [C.c_void_p, C.c_char_p, _F32P, C.c_int, _F32P]),
    "nb200_debug_gemm": (C.c_int, [C.c_void_p, C.c_char_p, _F32P, C.c_int, C.c_int, _F32P, _F32P, _I32P]),
    "nb200_bench_gemm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                   C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "nb200_trace_gemm": (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                   C.POINTER(C.c_uint64), C.c_int]),
}

_lib = None


def lib() -> C.CDLL:
    """Loads the CUDA library; raises if it has not been built (no fallback of any kind)."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}` (python __graft_entry__.py build)")
        _lib = C.CDLL(str(LIB_PATH))
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(_lib, name)
            fn.restype, fn.argtypes = res, args
    return _lib


class NB200Error(RuntimeError):
    pass


def _err() -> str:
    return (lib().nb200_last_error() or b"").decode(errors="replace")


def _ptr(a: np.ndarray, typ):
    return a.ctypes.data_as(typ)


def gguf_check(path: str) -> dict:
    """Host-only parse of a GGUF file the way nb200_model_load reads it (nb200_gguf_check); raises NB200Error on a malformed file."""
    import json
    buf = C.create_string_buffer(1 << 14)
    if lib().nb200_gguf_check(os.fsencode(path), buf, len(buf)) < 0:
        raise NB200Error(_err())
    return json.loads(buf.value.decode())


class Model:
    """nemo_context equivalent (src/nemo-ggml.h:236-261)."""

    def __init__(self, gguf_path: str, device: int = 0, max_streams: int = 64, flags: int = 0, weight_cache: "str | bool | None" = None):
        """weight_cache: None = build from the GGUF; True = through the native weight file next to the GGUF ("<gguf>.nb200w");
        a path = through that file (nb200_model_load_cached)."""
        if weight_cache is None or weight_cache is False:
            self._h = lib().nb200_model_load(os.fsencode(gguf_path), device, max_streams, flags)
        else:
            cp = None if weight_cache is True else os.fsencode(weight_cache)
            self._h = lib().nb200_model_load_cached(os.fsencode(gguf_path), cp, device, max_streams, flags)
        if not self._h:
            raise NB200Error(_err())
        self.info = Info()
        lib().nb200_model_info(self._h, C.byref(self.info))
        self._streams: list[Stream] = []

    def cache_stats(self) -> tuple[int, int]:
        """(big operands taken from the native weight file, built from the GGUF)"""
        a, b = C.c_int64(0), C.c_int64(0)
        lib().nb200_model_cache_stats(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def close(self):
        if self._h:
            lib().nb200_model_free(self._h)
            self._h = None
            for s in self._streams:
                s._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self, right_context: int = 0) -> "Stream":
        """nemo_stream_init(ctx, &cfg) with cfg.att_right_context = right_context."""
        s = Stream(self, right_context)
        self._streams.append(s)
        return s

    def step(self) -> int:
        n = lib().nb200_step(self._h)
        if n < 0:
            raise NB200Error(_err())
        return n

    def step_begin(self) -> int:
        """Launch one batched step without waiting (nb200_step_begin); feed() may go on until step_end()."""
        n = lib().nb200_step_begin(self._h)
        if n < 0:
            raise NB200Error(_err())
        return n

    def step_end(self) -> int:
        n = lib().nb200_step_end(self._h)
        if n < 0:
            raise NB200Error(_err())
        return n

    def batch_io(self, streams: "list[Stream]") -> "BatchIO":
        return BatchIO(self, streams)

    def drain(self) -> int:
        n = lib().nb200_drain(self._h)
        if n < 0:
            raise NB200Error(_err())
        return n

    def set_dup_family(self, family: "str | None"):
        """nb200_set_dup_family: launch every kernel of `family` twice inside the step (marginal in-graph cost); None = off."""
        if lib().nb200_set_dup_family(self._h, family.encode() if family else None) < 0:
            raise NB200Error(_err())

    def stats(self) -> Stats:
        st = Stats()
        lib().nb200_get_stats(self._h, C.byref(st))
        return st

    def timer_start(self):
        if lib().nb200_timer_start(self._h) < 0:
            raise NB200Error(_err())

    def timer_stop(self) -> float:
        """Elapsed device time (ms) on the library's stream since timer_start (CUDA events)."""
        ms = C.c_double()
        if lib().nb200_timer_stop(self._h, C.byref(ms)) < 0:
            raise NB200Error(_err())
        return ms.value

    def profile(self, on: bool):
        if lib().nb200_profile_enable(self._h, 1 if on else 0) < 0:
            raise NB200Error(_err())

    def profile_report(self) -> dict:
        import json
        buf = C.create_string_buffer(1 << 16)
        if lib().nb200_profile_report(self._h, buf, len(buf)) < 0:
            raise NB200Error(_err())
        return json.loads(buf.value.decode())

    def tokens_to_text(self, tokens) -> str:
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        buf = C.create_string_buffer(16 * max(1, t.size) + 16)
        n = lib().nb200_tokens_to_text(self._h, _ptr(t, _I32P), int(t.size), buf, len(buf))
        if n < 0:
            raise NB200Error(_err())
        return buf.raw[:n].decode("utf-8", errors="replace")

    def set_language(self, lang: str) -> bool:
        """nemo_set_language (src/nemo-ggml.cpp:497-511): False (and no change) for an unknown code / monolingual model"""
        return lib().nb200_model_set_language(self._h, lang.encode()) == 0

    def encode_audio(self, pcm: np.ndarray, want_enc: bool = False):
        """nemo_encode_audio (src/nemo-ggml.cpp:1690-1715), the offline path: -> (tokens, frames) or (tokens, frames, enc [T,1024])"""
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        max_frames = pcm.size // 1280 + 4
        cap = 10 * max_frames
        tok, fr = np.zeros(cap, np.int32), np.zeros(cap, np.int64)
        enc = np.zeros((max_frames, 1024), np.float32) if want_enc else None
        nfr = C.c_int(0)
        n = lib().nb200_encode_audio(self._h, _ptr(pcm, _I16P), int(pcm.size), _ptr(tok, _I32P), _ptr(fr, _I64P), cap,
                                     _ptr(enc, _F32P) if want_enc else None, int(enc.size) if want_enc else 0, C.byref(nfr))
        if n < 0:
            raise NB200Error(_err())
        if want_enc:
            return tok[:n].tolist(), fr[:n].tolist(), enc[: nfr.value].copy()
        return tok[:n].tolist(), fr[:n].tolist()

    def tokens_to_text_timed(self, tokens, frames) -> str:
        """tokens_to_text(timed tokens, vocab, timestamp_words=true), src/nemo-ggml.cpp:1556-1583"""
        t = np.ascontiguousarray(tokens, dtype=np.int32)
        f = np.ascontiguousarray(frames, dtype=np.int64)
        if t.size != f.size:
            raise ValueError("tokens and frames differ in length")
        buf = C.create_string_buffer(48 * max(1, t.size) + 16)
        n = lib().nb200_tokens_to_text_timed(self._h, _ptr(t, _I32P), _ptr(f, _I64P), int(t.size), buf, len(buf))
        if n < 0:
            raise NB200Error(_err())
        return buf.raw[:n].decode("utf-8", errors="replace")

    def mel_batch(self, pcm: np.ndarray) -> np.ndarray:
        """nemo_preprocessor_process for [n_streams, n_samples] int16 signals, each from a fresh preprocessor."""
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        ns, n = pcm.shape
        nf = 0 if n + 256 < 512 else (n + 256 - 512) // 160 + 1
        out = np.zeros((ns, nf, 128), np.float32)
        r = lib().nb200_mel_batch(self._h, _ptr(pcm, _I16P), ns, n, _ptr(out, _F32P))
        if r < 0:
            raise NB200Error(_err())
        return out

    def bench_gemm(self, M: int, N: int, K: int, epi_kind: int = 0, bn: int = 0, splits: int = 1, iters: int = 50) -> dict:
        """Device time of the tcgen05 GEMM alone (weights streamed from HBM); see include/nb200.h."""
        us, bn_used, sp_used = C.c_double(), C.c_int(), C.c_int()
        if lib().nb200_bench_gemm(self._h, M, N, K, epi_kind, bn, splits, iters, C.byref(us), C.byref(bn_used), C.byref(sp_used)) < 0:
            raise NB200Error(_err())
        return {"us": us.value, "bn": bn_used.value, "splits": sp_used.value, "tflops": 2.0 * M * N * K / us.value * 1e-6}

    def bench_mma_peak(self, kind: int, ctas: int = 148, iters: int = 4096, seed: int = 0):
        """Tensor-pipe peak of kind::f16 (0) or kind::i8 (1) with one resident 128 x 256 tile per CTA; also returns CTA 0's
        accumulator after one k-block and the operands, for an arithmetic check. -> (TFLOP/s or TOP/s, acc [128][256], a, b)"""
        rng = np.random.default_rng(seed)
        if kind == 0:
            a = rng.integers(-8, 9, size=(128, 64)).astype(np.float16)
            b = rng.integers(-8, 9, size=(256, 64)).astype(np.float16)
        else:
            a = rng.integers(-127, 128, size=(128, 128)).astype(np.int8)
            b = rng.integers(-127, 128, size=(256, 128)).astype(np.int8)
        acc = np.zeros((128, 256), np.uint32)
        tops = C.c_double(0.0)
        rc = lib().nb200_bench_mma_peak(self._h, kind, ctas, iters, a.ctypes.data, b.ctypes.data, acc.ctypes.data, C.byref(tops))
        if rc != 0:
            raise RuntimeError(_err())
        return tops.value, (acc.view(np.float32) if kind == 0 else acc.view(np.int32)), a, b

    def trace_gemm(self, M: int, N: int, K: int, epi_kind: int = 0, bn: int = 0, splits: int = 1, launches: int = 6) -> np.ndarray:
        """Per-CTA phase timestamps [launches, ctas, 8] (ns) of back-to-back launches; see include/nb200.h."""
        buf = np.zeros(launches * 4096 * 8, np.uint64)
        n = lib().nb200_trace_gemm(self._h, M, N, K, epi_kind, bn, splits, launches, _ptr(buf, C.POINTER(C.c_uint64)), buf.size)
        if n < 0:
            raise NB200Error(_err())
        return buf[: launches * n * 8].reshape(launches, n, 8)

    def debug_matmul(self, tensor_name: str, x: np.ndarray, n_out: int) -> np.ndarray:
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.zeros((x.shape[0], n_out), np.float32)
        if lib().nb200_debug_matmul(self._h, tensor_name.encode(), _ptr(x, _F32P), x.shape[0], _ptr(y, _F32P)) < 0:
            raise NB200Error(_err())
        return y


    EPI_F32, EPI_SILU_F16, EPI_GLU, EPI_SPLITK_LN, EPI_F16 = 0, 1, 4, 5, 6

    def debug_gemm(self, tensor_name: str, x: np.ndarray, n_out: int, epi_kind: int):
        """The step's GEMM + fused epilogue on one matrix (nb200_debug_gemm). -> (y [M, n_out], y_ln or None, cfg dict)."""
        x = np.ascontiguousarray(x, dtype=np.float32)
        y = np.zeros((x.shape[0], n_out), np.float32)
        y_ln = np.zeros_like(y) if epi_kind == self.EPI_SPLITK_LN else None
        cfg = np.zeros(4, np.int32)
        if lib().nb200_debug_gemm(self._h, tensor_name.encode(), _ptr(x, _F32P), x.shape[0], epi_kind, _ptr(y, _F32P),
                                  _ptr(y_ln, _F32P) if y_ln is not None else None, _ptr(cfg, _I32P)) < 0:
            raise NB200Error(_err())
        return y, y_ln, {"bn": int(cfg[0]), "splits": int(cfg[1]), "cta_group": int(cfg[2]), "ctas": int(cfg[3])}


class BatchIO:
    """Feeds / polls a fixed list of streams with one library call each (nb200_feed_batch / nb200_poll_batch)."""

    def __init__(self, model: Model, streams: "list[Stream]"):
        self.model, self.n = model, len(streams)
        self._handles = (C.c_void_p * self.n)(*[s._h for s in streams])
        self._ptrs = (C.c_void_p * self.n)()
        self._ptr_view = np.frombuffer(self._ptrs, dtype=np.uint64)
        self._lens = np.zeros(self.n, np.int32)
        self._counts = np.zeros(self.n, np.int32)
        self._toks = np.zeros(max(4096, 256 * self.n), np.int32)

    def feed_ptrs(self, addresses: np.ndarray, n_samples):
        """addresses: uint64 host addresses of int16 PCM, one per stream; n_samples: scalar or per-stream array."""
        self._ptr_view[:] = addresses
        self._lens[:] = n_samples
        if lib().nb200_feed_batch(self.model._h, self._handles, self._ptrs, _ptr(self._lens, _I32P), self.n) < 0:
            raise NB200Error(_err())

    def poll(self):
        """-> (tokens int32 [total], counts int32 [n_streams]) of everything emitted since the last poll."""
        for _ in range(2):
            n = lib().nb200_poll_batch(self.model._h, self._handles, self.n, _ptr(self._toks, _I32P), self._toks.size, _ptr(self._counts, _I32P))
            if n >= 0:
                return self._toks[:n].copy(), self._counts.copy()
            need = int(self._counts.sum())
            if need <= self._toks.size:
                break
            self._toks = np.zeros(need, np.int32)  # counts hold what is waiting: retry once with room for all of it
        raise NB200Error(_err())


class Stream:
    """nemo_stream_context equivalent (src/nemo-stream.h:177-326)."""

    def __init__(self, model: Model, right_context: int):
        self.model = model
        self.right_context = right_context
        self._h = lib().nb200_stream_create(model._h, right_context)
        if not self._h:
            raise NB200Error(_err())

    def free(self):
        if self._h and self.model._h:
            lib().nb200_stream_free(self._h)
        self._h = None

    def reset(self):
        if lib().nb200_stream_reset(self._h) < 0:
            raise NB200Error(_err())

    def set_language(self, lang: str):
        """nemo_stream_set_language (src/nemo-stream.cpp:735-749); raises where the reference returns false."""
        if lib().nb200_stream_set_language(self._h, lang.encode()) < 0:
            raise NB200Error(_err())

    def feed(self, pcm: np.ndarray):
        pcm = np.ascontiguousarray(pcm, dtype=np.int16)
        if lib().nb200_stream_feed(self._h, _ptr(pcm, _I16P), int(pcm.size)) < 0:
            raise NB200Error(_err())

    def feed_ptr(self, address: int, n_samples: int):
        """Feed from a raw host address (e.g. a pinned torch tensor's data_ptr())."""
        if lib().nb200_stream_feed(self._h, C.cast(C.c_void_p(address), _I16P), int(n_samples)) < 0:
            raise NB200Error(_err())

    def ready_chunks(self) -> int:
        return lib().nb200_stream_ready_chunks(self._h)

    def set_taps(self, taps: int):
        lib().nb200_stream_set_taps(self._h, taps)

    def fetch_tap(self, tap: int) -> np.ndarray:
        n = lib().nb200_stream_fetch_tap(self._h, tap, None, 0)
        out = np.zeros(max(n, 0), np.float32)
        if n > 0:
            lib().nb200_stream_fetch_tap(self._h, tap, _ptr(out, _F32P), n)
        return out

    def poll_tokens(self) -> list[int]:
        buf = np.zeros(4096, np.int32)
        out: list[int] = []
        while True:
            n = lib().nb200_stream_poll_tokens(self._h, _ptr(buf, _I32P), buf.size)
            if n <= 0:
                return out
            out.extend(buf[:n].tolist())

    def poll_timed(self) -> tuple[list[int], list[int]]:
        """new (tokens, encoder frames they were emitted at): the reference's timed_token pairs (src/nemo-ggml.h:383-395)"""
        buf, fr = np.zeros(4096, np.int32), np.zeros(4096, np.int64)
        toks: list[int] = []
        frames: list[int] = []
        while True:
            n = lib().nb200_stream_poll_timed(self._h, _ptr(buf, _I32P), _ptr(fr, _I64P), buf.size)
            if n <= 0:
                return toks, frames
            toks.extend(buf[:n].tolist())
            frames.extend(fr[:n].tolist())

    @property
    def token_frames(self) -> list[int]:
        n = lib().nb200_stream_get_token_frames(self._h, None, 0)
        buf = np.zeros(max(n, 1), np.int64)
        lib().nb200_stream_get_token_frames(self._h, _ptr(buf, _I64P), n)
        return buf[:n].tolist()

    @property
    def tokens(self) -> list[int]:
        n = lib().nb200_stream_get_tokens(self._h, None, 0)
        buf = np.zeros(max(n, 1), np.int32)
        lib().nb200_stream_get_tokens(self._h, _ptr(buf, _I32P), n)
        return buf[:n].tolist()

    @property
    def chunk_token_counts(self) -> list[int]:
        n = lib().nb200_stream_chunk_token_counts(self._h, None, 0)
        buf = np.zeros(max(n, 1), np.int32)
        lib().nb200_stream_chunk_token_counts(self._h, _ptr(buf, _I32P), n)
        return buf[:n].tolist()

    @property
    def chunk_margins(self) -> list[float]:
        n = lib().nb200_stream_chunk_margins(self._h, None, 0)
        buf = np.zeros(max(n, 1), np.float32)
        lib().nb200_stream_chunk_margins(self._h, _ptr(buf, _F32P), n)
        return buf[:n].tolist()

    # ---- the reference's call shapes ---------------------------------------------------------------
    def process_incremental(self, pcm: np.ndarray) -> str:
        """nemo_stream_process_incremental (src/nemo-stream.cpp:1145-1206): feed, run every complete chunk of THIS
        stream, return the newly decoded text."""
        if pcm is None or len(pcm) == 0:
            return ""
        self.feed(pcm)
        while lib().nb200_stream_ready_chunks(self._h) > 0:
            self.model.step()
        return self.model.tokens_to_text(self.poll_tokens())

    def finalize(self) -> str:
        """nemo_stream_finalize (src/nemo-stream.cpp:1217-1293): flush the zero-padded tail; text added by the flush."""
        lib().nb200_stream_finalize(self._h)
        while lib().nb200_stream_ready_chunks(self._h) > 0:
            self.model.step()
        self.model.step()  # consumes the finalize mark even when the tail yields no frame
        return self.model.tokens_to_text(self.poll_tokens())
