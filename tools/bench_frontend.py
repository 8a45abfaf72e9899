#!/usr/bin/env python3
"""BASELINE.json config 4: mel preprocessor + conv-subsampling direct-conv kernels only, batch sweep over concurrent
streams. python tools/bench_frontend.py [--rc 0 13] [--streams 1 4 16 64 256 1024 4096] -> gpurun_out/frontend_sweep.json

Per batch size one chunk of every stream is stepped through a 2-layer model (the frontend does not depend on the layer
count) with the library's per-kernel profile on (CUDA events around every launch); reported per kernel: device time,
algorithmic GB/s against the measured HBM peak, audio-seconds per second of the frontend alone."""
import argparse
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))
import inputs  # noqa: E402
import nb200  # noqa: E402

FRONTEND = ("mel", "sub_conv0_dw2", "sub_dw5")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rc", type=int, nargs="+", default=[0, 13])
    ap.add_argument("--streams", type=int, nargs="+", default=[1, 4, 16, 64, 256, 1024, 4096])
    a = ap.parse_args()
    peak = 6462.4
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peak = json.loads(pk.read_text())["hbm_gbs"]
    gguf = inputs.gguf_path("q8_0")
    base = [inputs.pcm(100 + i, 4.0) for i in range(4)]
    rows = []
    for B in a.streams:
        m = nb200.Model(gguf, 0, B)
        for rc in a.rc:
            cs = 8 * (1 + rc) * 160
            n = 5 * cs
            streams = [m.stream(rc) for _ in range(B)]
            for i, s in enumerate(streams):
                sig = np.roll(base[i % 4], 997 * i)
                s.feed(np.tile(sig, n // sig.size + 1)[:n])
            m.step()
            m.step()  # shapes seen twice: lazy attribute calls done
            m.profile(True)
            reps = 2
            for _ in range(reps):
                assert m.step() == B
            prof = m.profile_report()
            m.profile(False)
            row = {"streams": B, "right_context": rc, "chunk_s": cs / 16000.0, "kernels": {}}
            tot = 0.0
            for k in FRONTEND:
                v = prof[k]
                us = v["ms"] / v["launches"] * 1e3
                tot += us
                row["kernels"][k] = {"us": round(us, 2), "gbs": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1),
                                     "hbm_frac": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9 / peak, 4)}
            row["frontend_us"] = round(tot, 2)
            row["frontend_audio_s_per_s"] = round(B * cs / 16000.0 / (tot * 1e-6), 1)
            rows.append(row)
            print(json.dumps(row), flush=True)
            for s in streams:
                s.free()
        m.close()
    out = ROOT / "gpurun_out"
    out.mkdir(exist_ok=True)
    (out / "frontend_sweep.json").write_text(json.dumps({"hbm_peak_gbs": peak, "rows": rows}, indent=1))


if __name__ == "__main__":
    main()
