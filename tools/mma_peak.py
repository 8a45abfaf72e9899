"""Tensor-pipe peak of tcgen05.mma kind::f16 vs kind::i8 on this device, with an arithmetic check of each kind against numpy
('measure the int8 peak with the first int8 GEMM'). One JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("tests", "nemotron-asr.cpp_b200", "tools"):
    sys.path.insert(0, os.path.join(ROOT, p))
import inputs  # noqa: E402
import nb200 as nb  # noqa: E402

m = nb.Model(inputs.gguf_path("q8_0"), device=0, max_streams=2)
out = {}
for kind, name in ((0, "f16"), (1, "i8")):
    res = {}
    for ctas in (1, 148):
        for iters in (4096, 16384):
            tops, acc, a, b = m.bench_mma_peak(kind, ctas, iters)
            res[f"ctas{ctas}_iters{iters}"] = round(tops, 1)
    ref = a.astype(np.int64) @ b.astype(np.int64).T if kind == 1 else a.astype(np.float64) @ b.astype(np.float64).T
    res["max_abs_err_vs_numpy"] = float(np.abs(acc.astype(np.float64) - ref).max())
    res["unit"] = "TFLOP/s" if kind == 0 else "TOP/s"
    out[name] = res
# kind 2: kind::i8 with ggml's Q8_0 block scales applied per K = 32 block (the arithmetic of ggml-cpu/quants.c:451-478)
import ctypes as C
rng = np.random.default_rng(1)
a = rng.integers(-127, 128, size=(128, 128)).astype(np.int8)
b = rng.integers(-127, 128, size=(256, 128)).astype(np.int8)
sw = (0.5 + 0.001 * ((np.arange(8 * 256) * 7) % 31)).astype(np.float32).reshape(8, 256)
sx = (0.25 + 0.002 * ((np.arange(8 * 128) * 5) % 29)).astype(np.float32).reshape(8, 128)
ref = np.zeros((128, 256), np.float64)
for it in range(8):
    k = 32 * (it & 3)
    blk = a[:, k:k + 32].astype(np.int64) @ b[:, k:k + 32].astype(np.int64).T
    ref += blk * (sx[it][:, None].astype(np.float64) * sw[it][None, :].astype(np.float64))
res = {}
for ctas in (1, 148):
    for iters in (4096, 16384):
        acc = np.zeros((128, 256), np.uint32)
        tops = C.c_double(0.0)
        rc = nb.lib().nb200_bench_mma_peak(m._h, 2, ctas, iters, a.ctypes.data, b.ctypes.data, acc.ctypes.data, C.byref(tops))
        assert rc == 0, nb._err()
        res[f"ctas{ctas}_iters{iters}"] = round(tops.value, 1)
got = acc.view(np.float32).astype(np.float64)
res["max_rel_err_vs_numpy"] = float(np.abs(got - ref).max() / np.abs(ref).max())
res["unit"] = "TOP/s"
res["what"] = ("one K = 32 MMA per quantisation block into one of two TMEM accumulators; 16 epilogue warps read it out and add "
               "d_x[m][blk] * d_w[n][blk] * acc to fp32 running sums in registers")
out["i8_q8_0_block_scaled"] = res
m.close()
print(json.dumps(out))
