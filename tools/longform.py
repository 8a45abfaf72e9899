#!/usr/bin/env python3
"""BASELINE.json config 5: long-form audio, many streams, cache roll-forward stress on one GPU (run once per GPU for
the 2/4/8-GPU variants: streams never interact). python tools/longform.py [--streams 256] [--minutes 60] [--rc 13]
-> gpurun_out/longform.json

Every stream is fed `minutes` of synthetic audio chunk by chunk in server-loop order (step_begin | feed next | step_end |
poll). Checked while running: device memory stays constant after the first steps, step time does not drift, and pairs
of streams that receive identical audio (stream i and i + streams/2) emit identical token sequences after the K/V
rings have wrapped thousands of times (their slots, ring phases and batch rows differ)."""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
for p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests", ROOT):
    sys.path.insert(0, str(p))
import inputs  # noqa: E402
import nb200  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=256)
    ap.add_argument("--minutes", type=float, default=60.0)
    ap.add_argument("--rc", type=int, default=13)
    ap.add_argument("--layers", type=int, default=24)
    a = ap.parse_args()
    import torch
    cs = 8 * (1 + a.rc) * 160
    steps = int(a.minutes * 60 * 16000 // cs)
    half = a.streams // 2
    gguf = inputs.gguf_path("q8_0", a.layers, 6.3)
    m = nb200.Model(gguf, 0, a.streams)
    streams = [m.stream(a.rc) for _ in range(a.streams)]
    io = m.batch_io(streams)
    # 64 s of distinct audio per stream pair, looped: stream i and i + half share a row
    period = 64 * 16000 // cs * cs
    base = [inputs.pcm(200 + i, 8.0) for i in range(8)]
    host = torch.empty((half, period + cs), dtype=torch.int16).pin_memory()
    arr = host.numpy()
    for i in range(half):
        sig = np.roll(base[i % 8], 1009 * i)
        arr[i, :period] = np.tile(sig, period // sig.size + 1)[:period]
        arr[i, period:] = arr[i, :cs]
    row = np.uint64(host.data_ptr()) + (np.arange(a.streams, dtype=np.uint64) % np.uint64(half)) * np.uint64(host.stride(0) * 2)
    counts = np.zeros(a.streams, np.int64)
    digest = np.zeros(a.streams, np.uint64)

    def feed(k):
        io.feed_ptrs(row + np.uint64(((k * cs) % period) * 2), cs)

    def collect():
        toks, cnt = io.poll()
        off = 0
        for i, c in enumerate(cnt):
            if c:
                t = toks[off:off + c].astype(np.uint64)
                for v in t:  # order-sensitive rolling hash of the stream's token sequence
                    digest[i] = (digest[i] * np.uint64(1099511628211) + v + np.uint64(1)) & np.uint64(0xFFFFFFFFFFFFFFFF)
                off += c
        counts[:] += cnt

    log = []
    feed(0)
    feed(1)  # one chunk ahead: a frame needs 256 samples beyond its hop
    free0 = None
    t0 = time.perf_counter()
    t_mark, k_mark = t0, 0
    for k in range(steps):
        n = m.step_begin()
        assert n == a.streams, (k, n)
        feed(k + 2)
        m.step_end()
        collect()
        if k == 20:
            free0 = torch.cuda.mem_get_info(0)[0]
        if (k + 1) % max(steps // 10, 1) == 0:
            now = time.perf_counter()
            st = m.stats()
            log.append({"step": k + 1, "ms_per_step_wall": (now - t_mark) / (k + 1 - k_mark) * 1e3, "last_step_gpu_ms": st.last_step_gpu_ms,
                        "free_bytes_delta": int(torch.cuda.mem_get_info(0)[0] - free0) if free0 else 0, "tokens": int(counts.sum())})
            print(json.dumps(log[-1]), flush=True)
            t_mark, k_mark = now, k + 1
    wall = time.perf_counter() - t0
    same = int(sum(digest[i] == digest[i + half] and counts[i] == counts[i + half] for i in range(half)))
    out = {"streams": a.streams, "right_context": a.rc, "minutes_per_stream": a.minutes, "steps": steps, "wall_s": wall,
           "audio_s_per_s": a.streams * steps * cs / 16000.0 / wall, "kv_ring_wraps": steps * (1 + a.rc) / 84.0,
           "identical_pairs": same, "pairs": half, "tokens_total": int(counts.sum()),
           "free_bytes_delta_end": log[-1]["free_bytes_delta"], "ms_per_step_first_last": [log[0]["ms_per_step_wall"], log[-1]["ms_per_step_wall"]],
           "log": log}
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / "longform.json").write_text(json.dumps(out, indent=1))
    print(json.dumps({k: v for k, v in out.items() if k != "log"}))
    assert same == half, "streams with identical audio diverged"
    assert abs(out["free_bytes_delta_end"]) <= (64 << 20), "device memory changed while stepping"
    m.close()


if __name__ == "__main__":
    main()
