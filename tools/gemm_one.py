#!/usr/bin/env python3
"""python tools/gemm_one.py M N K kind bn splits [iters] — one GEMM configuration (for ncu)."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))
import inputs  # noqa: E402
import nb200  # noqa: E402
M, N, K, kind, bn, sp = map(int, sys.argv[1:7])
iters = int(sys.argv[7]) if len(sys.argv) > 7 else 5
m = nb200.Model(inputs.gguf_path("q8_0"), 0, 2)
print(m.bench_gemm(M, N, K, kind, bn, sp, iters))
m.close()
