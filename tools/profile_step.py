#!/usr/bin/env python3
"""Runs a few batched steps of the bench workload (for ncu): python tools/profile_step.py [--streams 64] [--rc 13]
[--steps 2] [--layers 24]. Prints the library's own per-step stage times; meant to be wrapped by
  ncu --metrics gpu__time_duration.sum --clock-control none ...   (launch list)
  ncu --set full --clock-control none --import-source on -k regex:<kernel> ...   (one kernel)"""
import argparse
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=64)
    ap.add_argument("--rc", type=int, default=13)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--layers", type=int, default=24)
    a = ap.parse_args()
    import bench
    import inputs
    import nb200
    gguf = bench.bench_gguf("q8_0", a.layers)  # the benchmark's model (token-emitting decoder)
    m = nb200.Model(gguf, 0, a.streams)
    cs = 8 * (1 + a.rc) * 160
    base = [inputs.pcm(100 + i, 8.0) for i in range(8)]
    streams = []
    for i in range(a.streams):
        s = m.stream(a.rc)
        sig = np.roll(base[i % 8], 997 * i)
        s.feed(np.tile(sig, (a.steps + 1) * cs // sig.size + 1)[: (a.steps + 1) * cs])
        streams.append(s)
    for k in range(a.steps):
        n = m.step()
        st = m.stats()
        print(f"step {k}: items={n} gpu_ms={st.last_step_gpu_ms:.3f} frontend={st.last_frontend_ms:.3f} encoder={st.last_encoder_ms:.3f} "
              f"decoder={st.last_decoder_ms:.3f} launches={st.kernel_launches} rounds={st.decode_rounds} frames={st.frames} tokens={st.tokens}")
    m.close()


if __name__ == "__main__":
    main()
