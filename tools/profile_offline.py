#!/usr/bin/env python3
"""Per-kernel times of one offline utterance (nb200_encode_audio) from the library's own event profiler:
python tools/profile_offline.py [--secs 40] [--layers 24]"""
import argparse
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--secs", type=float, default=40.0)
    ap.add_argument("--layers", type=int, default=24)
    a = ap.parse_args()
    import inputs
    import nb200
    m = nb200.Model(inputs.gguf_path("q8_0", a.layers, 6.3), 0, 40)
    pcm = inputs.pcm(51, a.secs)
    m.encode_audio(pcm)
    t0 = time.perf_counter()
    m.encode_audio(pcm)
    wall = (time.perf_counter() - t0) * 1e3
    m.profile(True)
    m.encode_audio(pcm)
    rep = m.profile_report()
    m.profile(False)
    print(json.dumps({"seconds": a.secs, "wall_ms": wall, "report": rep}))
    m.close()


if __name__ == "__main__":
    main()
