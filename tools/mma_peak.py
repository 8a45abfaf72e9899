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
m.close()
print(json.dumps(out))
