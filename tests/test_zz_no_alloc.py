"""The allocation-free entry point nb200_mel_batch_noalloc (include/nb200.h; no counterpart in the reference)
against the existing nb200_mel_batch, through the C ABI (ctypes). Collected last on purpose (file name):
a failure here hides no other test. Self-contained: passes alone in a fresh process and
after the other tests."""
import ctypes as C

import numpy as np
import pytest

import inputs

pytestmark = pytest.mark.gpu

MEL_TOL = 1e-4  # north_star's log-mel tolerance, the one tests/test_gpu_parity.py uses (same kernel: the difference is expected to be 0)


def _call(nb, fn, model, pcm):
    ns, n = pcm.shape
    nf = 0 if n + 256 < 512 else (n + 256 - 512) // 160 + 1
    out = np.full((ns, nf, 128), np.nan, np.float32)
    r = fn(model._h, pcm.ctypes.data_as(C.POINTER(C.c_int16)), ns, n, out.ctypes.data_as(C.POINTER(C.c_float)))
    return r, out


def test_mel_batch_noalloc_matches_mel_batch(lib_built, gguf_q8):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    nb = lib_built
    L = nb.lib()
    model = nb.Model(gguf_q8, device=0, max_streams=2)
    try:
        # two calls in a row on the shared scratch: a full-size batch first, then a smaller and shorter one whose results must not
        # see anything the first call left behind (stale PCM beyond n_samples, stale mel rows, stale metadata)
        first = np.stack([inputs.pcm(300 + i, 4.2)[:65536] for i in range(16)])
        second = np.stack([inputs.pcm(400 + i, 1.3)[:20000] for i in range(5)])
        for k, pcm in enumerate((first, second, first)):
            pcm = np.ascontiguousarray(pcm, np.int16)
            r_ref, ref = _call(nb, L.nb200_mel_batch, model, pcm)
            r_new, got = _call(nb, L.nb200_mel_batch_noalloc, model, pcm)
            assert r_ref > 0 and r_new == r_ref, (k, r_new, r_ref, nb._err())
            assert np.isfinite(got).all() and got.shape == ref.shape
            err = float(np.abs(got - ref).max())
            print(f"call {k}: {pcm.shape[0]} x {pcm.shape[1]} samples -> {r_new} frames, max |noalloc - mel_batch| = {err:.1e}")
            assert err <= MEL_TOL
        # beyond the fixed scratch: rejected with a clear error, nothing computed
        for ns, n in ((17, 16000), (2, 65537)):
            big = np.zeros((ns, n), np.int16)
            r, _ = _call(nb, L.nb200_mel_batch_noalloc, model, big)
            assert r < 0 and "fixed device scratch" in nb._err(), (ns, n, r, nb._err())
        # and the entry point still works after a rejected call
        r, got = _call(nb, L.nb200_mel_batch_noalloc, model, np.ascontiguousarray(second, np.int16))
        _, ref = _call(nb, L.nb200_mel_batch, model, np.ascontiguousarray(second, np.int16))
        assert r > 0 and float(np.abs(got - ref).max()) <= MEL_TOL
    finally:
        model.close()
