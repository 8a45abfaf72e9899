#!/usr/bin/env python3
"""Phase breakdown of the tcgen05 GEMM from in-kernel globaltimer stamps: python tools/trace_gemm.py [M]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))
import inputs  # noqa: E402
import nb200  # noqa: E402

SHAPES = [("ffn_w1", 4096, 1024, 1), ("ffn_w2", 1024, 4096, 5), ("qkv", 3072, 1024, 6), ("out_pw2", 1024, 1024, 5), ("pw1", 2048, 1024, 4)]


def main():
    M = int(sys.argv[1]) if len(sys.argv) > 1 else 896
    force = [int(x) for x in sys.argv[2:4]] if len(sys.argv) > 3 else None  # bn (1000+ = CTA pair), splits
    m = nb200.Model(inputs.gguf_path("q8_0"), 0, 2)
    for name, N, K, kind in SHAPES:
        t = (m.trace_gemm(M, N, K, kind, force[0], force[1] if kind == 5 else 1, 8) if force else m.trace_gemm(M, N, K, kind, 0, 8 if kind == 5 else 1, 8)).astype(np.int64)
        L, C, _ = t.shape
        print(f"{name} M={M} N={N} K={K}: {C} CTAs")
        for l in range(2, L):
            x = t[l]
            t0 = x[:, 0].min()
            prev_end = t[l - 1][:, 5].max()
            ph = [(x[:, i] - t0) / 1e3 for i in range(7)]
            print(f"  launch {l}: entry-after-prev-end {(t0 - prev_end) / 1e3:6.2f} us | entry max {ph[0].max():5.2f} | prologue {np.median(ph[1]):5.2f} | "
                  f"dep-wait {np.median(ph[2]):5.2f} (max {ph[2].max():5.2f}) | first-stage {np.median(ph[3]):5.2f} | accum {np.median(ph[4]):5.2f} (max {ph[4].max():5.2f}) | "
                  f"epi-warp2 {np.median(ph[6]):5.2f} | ldtm/xform/sts/tma ns {int(np.median(x[:, 7] & 0xfff))}/{int(np.median((x[:, 7] >> 12) & 0xfff))}/{int(np.median((x[:, 7] >> 24) & 0xfff))}/{int(np.median((x[:, 7] >> 36) & 0xfff))} | end {np.median(ph[5]):5.2f} (max {ph[5].max():5.2f}) | launch-to-launch {(t0 - t[l - 1][:, 0].min()) / 1e3:6.2f}")
    m.close()


if __name__ == "__main__":
    main()
