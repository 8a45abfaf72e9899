#!/usr/bin/env python3
"""Where a batched step goes, measured IN the step's CUDA graph: for each kernel family the step is run with every launch of
that family issued twice (nb200_set_dup_family) and the difference to the plain step is the family's in-graph cost; the decode
kernel stamps its own duration. python tools/step_costs.py [--streams 64] [--rc 13] [--steps 20] -> one JSON line."""
import argparse
import json
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
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--layers", type=int, default=24)
    a = ap.parse_args()
    import bench
    import inputs
    import nb200
    m = nb200.Model(bench.bench_gguf("q8_0", a.layers), 0, a.streams)
    cs = 8 * (1 + a.rc) * 160
    fams = [None, "gemm", "layernorm", "attention", "dwconv", None]
    n_steps = len(fams) * (a.steps + 4) + 2
    base = [inputs.pcm(100 + i, 8.0) for i in range(8)]
    streams = []
    for i in range(a.streams):
        s = m.stream(a.rc)
        sig = np.roll(base[i % 8], 997 * i)
        s.feed(np.tile(sig, n_steps * cs // sig.size + 1)[: n_steps * cs])
        streams.append(s)
    res, dec = {}, []
    for fam in fams:
        m.set_dup_family(fam)
        for _ in range(4):  # eager pass + graph capture + warm replays
            m.step()
        t = []
        for _ in range(a.steps):
            m.step()
            st = m.stats()
            t.append(st.last_step_gpu_ms)  # device time of the step (CUDA events on the library's stream)
            dec.append(st.last_decoder_ms)
        res.setdefault(fam or "plain", []).append(float(np.median(t)))
    m.set_dup_family(None)
    plain = float(np.mean(res["plain"]))
    out = {"streams": a.streams, "right_context": a.rc, "ms_per_step": plain, "plain_runs_ms": res["plain"],
           "in_graph_cost_ms": {f: res[f][0] - plain for f in ("gemm", "layernorm", "attention", "dwconv")},
           "decode_kernel_ms": float(np.mean(dec)),
           "what": "in_graph_cost = device time of a step with every launch of the family issued twice minus the plain step (medians; CUDA events "
                   "on the library's stream); decode from the kernel's own globaltimer stamps; the rest (frontend, joint.enc, launch gaps) is the remainder"}
    out["remainder_ms"] = plain - sum(out["in_graph_cost_ms"].values()) - out["decode_kernel_ms"]
    print(json.dumps(out))
    m.close()


if __name__ == "__main__":
    main()
