#!/usr/bin/env python3
"""Offline path (nb200_encode_audio) on the 0.6B architecture: seconds of audio per second for single utterances of
several lengths, one after the other on one GPU (the offline path is not batched across utterances), next to the
unmodified reference's offline path on the host cores (`oracle/_ref/<isa>/ref_harness offline`, test infrastructure,
used here only as the timed CPU baseline).

    python tools/bench_offline.py [--secs 5 10 20 40] [--reps 5] [--no-cpu]"""
import argparse
import json
import subprocess
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--secs", type=float, nargs="+", default=[5.0, 10.0, 20.0, 40.0])
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--layers", type=int, default=24)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    import inputs
    import nb200
    gguf = inputs.gguf_path("q8_0", a.layers, 6.3)
    m = nb200.Model(gguf, 0, 40)
    rows = []
    for secs in a.secs:
        pcm = inputs.pcm(51, secs)
        m.encode_audio(pcm)  # warm-up: positional table of this length, lazy kernel attributes
        times = []
        for _ in range(a.reps):
            t0 = time.perf_counter()
            tok, fr = m.encode_audio(pcm)
            times.append(time.perf_counter() - t0)
        dt = float(np.median(times))
        row = {"seconds": secs, "frames": int(len(pcm) // 1280), "ms": dt * 1e3, "ms_min": min(times) * 1e3, "ms_max": max(times) * 1e3,
               "audio_s_per_s": secs / dt, "tokens": len(tok)}
        if not a.no_cpu:
            isa = "avx512" if "avx512f" in open("/proc/cpuinfo").read() else "avx2"
            h = ROOT / "oracle" / "_ref" / isa / "ref_harness"
            r = subprocess.run([str(h), "offline", gguf, inputs.pcm_path(51, secs), "0", str(inputs.TMP / "boff"), str(secs)],
                               stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            if r.returncode == 0:
                info = json.loads(r.stdout.strip().splitlines()[-1])
                row["cpu_reference_ms"] = info["proc_s"] * 1e3
                row["cpu_reference_audio_s_per_s"] = secs / info["proc_s"]
        rows.append(row)
        print(json.dumps(row), file=sys.stderr)
    m.close()
    print(json.dumps({"metric": "offline_rtfx_single_utterance", "unit": "audio-s/s", "model": f"0.6B architecture, {a.layers} layers, Q8_0, random init",
                      "timing": "wall clock around nb200_encode_audio (host PCM in, host tokens out), median of reps", "results": rows}))


if __name__ == "__main__":
    main()
