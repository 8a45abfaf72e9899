#!/usr/bin/env python3
"""GEMM tile/split sweep on the GPU: python tools/tune_gemm.py [M ...] -> gpurun_out/tune_gemm.json"""
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))
import inputs  # noqa: E402
import nb200  # noqa: E402

SHAPES = [("ffn_w1", 4096, 1024, 1), ("ffn_w2", 1024, 4096, 5), ("qkv", 3072, 1024, 6), ("out_pw2", 1024, 1024, 5), ("pw1", 2048, 1024, 4)]


def main():
    Ms = [int(x) for x in sys.argv[1:]] or [896]
    m = nb200.Model(inputs.gguf_path("q8_0"), 0, 2)
    res = []
    for M in Ms:
        for name, N, K, kind in SHAPES:
            auto = m.bench_gemm(M, N, K, kind, 0, 1)
            row = {"M": M, "shape": name, "N": N, "K": K, "auto": auto, "sweep": []}
            for bn in (1256, 1192, 1128, 256, 128, 64, 32):
                for sp in ((1, 2, 3, 4, 5, 6, 8) if kind == 5 else (1,)):
                    if N % (bn % 1000) or (bn == 32 and kind in (1, 4, 6)) or (bn >= 1000 and M <= 128):
                        continue
                    r = m.bench_gemm(M, N, K, kind, bn, sp)
                    row["sweep"].append({"bn": bn, "splits": sp, "us": round(r["us"], 2), "tflops": round(r["tflops"], 1)})
            best = min(row["sweep"], key=lambda x: x["us"])
            print(f"M={M} {name:8s} N={N} K={K}: auto bn={auto['bn']} sp={auto['splits']} {auto['us']:.1f} us {auto['tflops']:.0f} TF | best {best}", flush=True)
            print("    " + " ".join(f"{x['bn']}x{x['splits']}:{x['us']}" for x in row["sweep"]), flush=True)
            res.append(row)
    m.close()
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "tune_gemm.json").write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
