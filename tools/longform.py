#!/usr/bin/env python3
"""BASELINE.json config 5: long-form audio, many streams, cache roll-forward stress on one GPU (run once per GPU for
the 2/4/8-GPU variants: streams never interact). python tools/longform.py [--streams 256] [--minutes 60] [--rc 13]
-> gpurun_out/longform.json

Every stream is fed `minutes` of synthetic audio chunk by chunk in server-loop order (step_begin | feed next | step_end |
poll). Checked while running: device memory stays constant after the first steps, step time does not drift, and pairs
of streams that receive identical audio (stream i and i + streams/2) emit identical token sequences after the K/V
rings have wrapped thousands of times (their slots, ring phases and batch rows differ). The model is the benchmark's
(token-emitting 'stationary' decoder, ~0.25 tokens per frame), so the pair check compares real token streams.

--replay-check ("bounded drift vs a fresh-cache replay of the last 6 s"): the encoder output of
8 streams is tapped at two ages — after ~66 s (the caches have just filled and wrapped once) and at the very end (tens of
thousands of roll-forwards) — and each is compared with a FRESH stream that is fed only the last `--replay-secs` of the same
audio. A fresh replay never equals the long-running stream exactly (a frame sees 70 cached frames per layer, i.e. up to
24 x 5.6 s of history through the stack), so the distance at age 66 s is the yardstick: the roll-forward is sound if the
distance at the end is no larger than that, whatever the number of ring wraps in between.

Under torchrun (one process per GPU, LOCAL_RANK) every rank runs its own copy: the 2/4/8-GPU variants of the config."""
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
    ap.add_argument("--replay-check", action="store_true")
    ap.add_argument("--replay-secs", type=float, default=6.0)
    a = ap.parse_args()
    import os

    import torch
    dev = int(os.environ.get("LOCAL_RANK", "0"))
    rank_tag = f"_rank{dev}" if "LOCAL_RANK" in os.environ else ""
    cs = 8 * (1 + a.rc) * 160
    steps = int(a.minutes * 60 * 16000 // cs)
    half = a.streams // 2
    import bench
    gguf = bench.bench_gguf("q8_0", a.layers)
    m = nb200.Model(gguf, dev, a.streams)
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

    # replay check: tap the encoder output of 8 streams at two ages (one eager step each)
    T = 1 + a.rc
    sample = sorted({(i * a.streams) // 8 for i in range(8)}) if a.replay_check else []
    n_rep = max(1, int(np.ceil(a.replay_secs * 16000 / cs)))
    k_early = min(steps - 1, max(n_rep, int(np.ceil(66.0 * 16000 / cs))))
    tapped = {}  # step -> {stream: enc [T][1024]}

    log = []
    feed(0)
    feed(1)  # one chunk ahead: a frame needs 256 samples beyond its hop
    free0 = None
    t0 = time.perf_counter()
    t_mark, k_mark = t0, 0
    for k in range(steps):
        tap_now = a.replay_check and k in (k_early, steps - 1)
        if tap_now:
            for i in sample:
                streams[i].set_taps(nb200.TAP_ENC)
        n = m.step_begin()
        assert n == a.streams, (k, n)
        feed(k + 2)
        m.step_end()
        collect()
        if tap_now:
            tapped[k] = {i: streams[i].fetch_tap(nb200.TAP_ENC).reshape(-1, T, 1024)[-1].copy() for i in sample}
            for i in sample:
                streams[i].set_taps(0)
        if k == 20:
            free0 = torch.cuda.mem_get_info(dev)[0]
        if (k + 1) % max(steps // 10, 1) == 0:
            now = time.perf_counter()
            st = m.stats()
            log.append({"step": k + 1, "ms_per_step_wall": (now - t_mark) / (k + 1 - k_mark) * 1e3, "last_step_gpu_ms": st.last_step_gpu_ms,
                        "free_bytes_delta": int(torch.cuda.mem_get_info(dev)[0] - free0) if free0 else 0, "tokens": int(counts.sum())})
            print(json.dumps(log[-1]), flush=True)
            t_mark, k_mark = now, k + 1
    wall = time.perf_counter() - t0
    same = int(sum(digest[i] == digest[i + half] and counts[i] == counts[i + half] for i in range(half)))
    out = {"streams": a.streams, "right_context": a.rc, "minutes_per_stream": a.minutes, "steps": steps, "wall_s": wall,
           "audio_s_per_s": a.streams * steps * cs / 16000.0 / wall, "kv_ring_wraps": steps * (1 + a.rc) / 84.0,
           "identical_pairs": same, "pairs": half, "tokens_total": int(counts.sum()),
           "free_bytes_delta_end": log[-1]["free_bytes_delta"], "ms_per_step_first_last": [log[0]["ms_per_step_wall"], log[-1]["ms_per_step_wall"]],
           "log": log}
    if a.replay_check:
        # fresh streams fed only the last n_rep chunks before each tapped step (same audio rows, same chunk alignment)
        for s_ in streams:
            s_.free()
        replay = {}
        for k_tap, encs in tapped.items():
            fresh = {i: m.stream(a.rc) for i in sample}
            j0 = k_tap - n_rep + 1
            for i, fs in fresh.items():
                r = arr[i % half]
                sig = np.concatenate([r[((j * cs) % period):((j * cs) % period) + cs] for j in range(j0, k_tap + 2)])  # + the look-ahead chunk
                fs.feed(sig)
            for step_i in range(n_rep):
                if step_i == n_rep - 1:
                    for fs in fresh.values():
                        fs.set_taps(nb200.TAP_ENC)
                assert m.step() == len(fresh)
            d = []
            for i, fs in fresh.items():
                e = fs.fetch_tap(nb200.TAP_ENC).reshape(-1, T, 1024)[-1]
                d.append(float(np.linalg.norm(e - encs[i]) / np.linalg.norm(encs[i])))
                fs.free()
            replay[k_tap] = d
        early, late = replay[k_early], replay[steps - 1]
        out["replay_check"] = {"replay_seconds": n_rep * cs / 16000.0, "streams": sample,
                               "age_early_s": (k_early + 1) * cs / 16000.0, "age_late_s": steps * cs / 16000.0,
                               "rel_l2_vs_fresh_replay_early": early, "rel_l2_vs_fresh_replay_late": late,
                               "median_early": float(np.median(early)), "median_late": float(np.median(late)),
                               "what": "relative L2 distance of the tapped chunk's encoder output between the long-running stream and a fresh "
                                       "stream fed only the last replay_seconds of the same audio; early = the yardstick (context truncation), "
                                       "late must not exceed it: no drift from the ring / conv-cache roll-forward"}
    (ROOT / "gpurun_out").mkdir(exist_ok=True)
    (ROOT / "gpurun_out" / f"longform{rank_tag}.json").write_text(json.dumps(out, indent=1))
    print(json.dumps({k: v for k, v in out.items() if k != "log"}))
    assert same == half, "streams with identical audio diverged"
    assert abs(out["free_bytes_delta_end"]) <= (64 << 20), "device memory changed while stepping"
    if a.replay_check:
        rc_ = out["replay_check"]
        assert rc_["median_late"] <= 1.5 * rc_["median_early"] + 1e-3 and max(rc_["rel_l2_vs_fresh_replay_late"]) <= 2.0 * max(rc_["rel_l2_vs_fresh_replay_early"]) + 1e-3, \
            "encoder output drifted against the fresh-cache replay"
    m.close()


if __name__ == "__main__":
    main()
