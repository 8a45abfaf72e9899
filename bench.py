#!/usr/bin/env python3
"""Benchmark of the streaming-ASR hot path (BASELINE.json): audio-seconds per wall-second (RTFx) over concurrent
streams, plus per-chunk latency, on N B200s of one node (one process per GPU; streams sharded by id, no collective).

    python bench.py --gpus 1 --steps 20 --warmup 3                         # own arm (CUDA library through its C ABI)
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference --steps 20 --warmup 3                 # the reference's ggml-CPU path (oracle/_ref)

Workload at N=1 = BASELINE.json configs[1]: nemotron-speech-streaming 0.6B architecture (24 layers, random-init Q8_0
GGUF, seed 1234), 64 concurrent streams, 1.12 s chunks (right_context 13). One "step" = one batched chunk of every
stream (64 x 1.12 s of audio). Weak scaling: 64 streams per GPU.

The model is random-init (no checkpoint offline) with a decoder calibrated to emit ~0.25 tokens per encoder frame like real
speech (tools/make_random_gguf.py), so the greedy loop's LSTM updates run inside every timed step.

Legs (all timed with CUDA events on the library's stream, max over ranks):
  value  K steps with the audio already resident in HBM (fed into the device PCM rings before the timed region)
  e2e    K steps through the public call sequence feed (pinned host PCM -> H2D) / step / poll tokens (D2H)
  roofline  the tcgen05 GEMMs' in-graph cost (K more steps with every GEMM launched twice, minus the K timed steps) against the
            sustained bf16 peak; the back-to-back microbenchmark against the burst peak and the eager per-launch events beside it
  latency_path  BASELINE config 3: 128 streams per GPU x 80 ms chunks (1024 streams at N = 8): ms per step, p50/p99 chunk latency,
            achieved fraction of the HBM roofline (weights + K/V caches streamed once per step)
  cpu_baseline  the unmodified reference (oracle/_ref ref_harness) on the host cores, bounded sample
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
for _p in (ROOT / "nemotron-asr.cpp_b200", ROOT / "tools", ROOT / "tests"):
    sys.path.insert(0, str(_p))

METRIC = "audio-sec/sec (RTFx) over concurrent streams"
UNIT = "audio-s/s"
HOP, SR = 160, 16000
# The benchmark model: random-init weights of the 0.6B architecture with the depth-scaled residual initialisation (the encoder
# output then follows the audio frame by frame) and the "stationary" decoder, calibrated to the profile of real speech: ~0.25
# tokens per encoder frame, hardly ever more than 3 symbols on a frame (tools/make_random_gguf.py DECODERS,
# tests/golden/calibrate_bench_decoder.py). Both arms load the same file.
BENCH_BLANK_GAIN = 6.0
BENCH_DECODER = "stationary"
BENCH_RES_SCALE = 0.1
LATENCY_STREAMS = 128  # BASELINE config 3: 1024 concurrent 80 ms-chunk streams sharded by stream over 8 GPUs = 128 per GPU
# fp16 operand bytes a step reads per conformer layer: two FFNs (2 x 2 x 4096 x 1024), q|k|v (3 x 1024^2), linear_out, pointwise_conv1
# (2 x 1024^2), pointwise_conv2 (linear_pos is folded into a per-mode table at stream creation)
WEIGHT_BYTES_PER_LAYER = 2 * (4 * 4096 * 1024 + 3 * 1024 * 1024 + 1024 * 1024 + 2 * 1024 * 1024 + 1024 * 1024)


def bench_gguf(variant: str, layers: int) -> str:
    import inputs
    return inputs.gguf_path(variant, layers, BENCH_BLANK_GAIN, decoder=BENCH_DECODER, res_scale=BENCH_RES_SCALE)


def chunk_samples(rc: int) -> int:
    return 8 * (1 + rc) * HOP


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm_gbs": d["hbm_gbs"], "tflops": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops": 1400.0, "source": "fallback"}


def ref_harness_bin():
    flags = open("/proc/cpuinfo").read()
    order = (["avx512"] if ("avx512_vnni" in flags or "avx512vnni" in flags) and "avx512bw" in flags else []) + ["avx2"]
    for v in order:
        p = ROOT / "oracle" / "_ref" / v / "ref_harness"
        if p.exists():
            return str(p), v
    return None, None


def host_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def run_reference(gguf: str, rc: int, chunks: int, threads: int, seed: int = 21) -> dict:
    """Times the unmodified reference on `chunks` chunks of one stream (it has no cross-stream batching)."""
    import inputs
    exe, isa = ref_harness_bin()
    if exe is None:
        raise RuntimeError("oracle/_ref/<isa>/ref_harness is missing (built by __graft_entry__.build() where /root/reference exists)")
    secs = chunks * chunk_samples(rc) / SR
    pcm = inputs.pcm_path(seed, round(secs + 0.05, 3))
    r = subprocess.run([exe, "stream", gguf, pcm, str(rc), str(threads), "-", f"{secs:.4f}"], check=True,
                       stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    d = json.loads(r.stdout.strip().splitlines()[-1])
    d["isa"] = isa
    return d


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed regions run."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        except Exception:
            pass

    def stop(self) -> dict:
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def build_pcm(n_streams: int, n_samples: int, first_id: int):
    """[n_streams, n_samples] int16 in pinned host memory; stream s = base signal (s mod 8) rolled by a stream-specific
    offset and tiled (seeded, see tools/make_pcm.py)."""
    import torch

    import inputs
    base = [inputs.pcm(100 + i, 8.0) for i in range(8)]
    out = torch.empty((n_streams, n_samples), dtype=torch.int16).pin_memory()
    arr = out.numpy()
    for s in range(n_streams):
        g = first_id + s
        sig = np.roll(base[g % 8], 997 * g)
        arr[s] = np.tile(sig, n_samples // sig.size + 1)[:n_samples]
    return out


def offline_main(a) -> None:
    """`bench.py --offline`: the offline path (nb200_encode_audio) on single utterances of several lengths, one after the other
    on one GPU (it is not batched across utterances), with the unmodified reference's offline path (nemo_encode through
    `oracle/_ref/<isa>/ref_harness offline`) on the host cores as the CPU baseline. Not the headline metric: one JSON line of
    its own."""
    import inputs
    lib_path = ROOT / "nemotron-asr.cpp_b200" / "libnemob200.so"
    if not lib_path.exists():
        import __graft_entry__ as ge
        ge.build_library()
    import nb200
    gguf = bench_gguf(a.variant, a.layers)
    m = nb200.Model(gguf, 0, 40)
    rows = []
    for secs in a.offline_secs:
        pcm = inputs.pcm(51, secs)
        m.encode_audio(pcm)  # warm-up: positional table of this length, lazy kernel attributes
        times = []
        for _ in range(a.offline_reps):
            t0 = time.perf_counter()
            tok, _fr = m.encode_audio(pcm)
            times.append(time.perf_counter() - t0)
        dt = float(np.median(times))
        row = {"seconds": secs, "frames": int(len(pcm) // 1280), "ms": dt * 1e3, "ms_min": min(times) * 1e3, "ms_max": max(times) * 1e3,
               "audio_s_per_s": secs / dt, "tokens": len(tok)}
        if not a.no_cpu_baseline:
            exe, isa = ref_harness_bin()
            if exe is not None:
                r = subprocess.run([exe, "offline", gguf, inputs.pcm_path(51, secs), "0", str(Path(inputs.TMP) / "boff"), str(secs)],
                                   stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
                if r.returncode == 0:
                    info = json.loads(r.stdout.strip().splitlines()[-1])
                    row["cpu_reference_ms"] = info["proc_s"] * 1e3
                    row["cpu_reference_audio_s_per_s"] = secs / info["proc_s"]
                    row["cpu_reference_isa"] = isa
        rows.append(row)
        print(json.dumps(row), file=sys.stderr)
    m.close()
    print(json.dumps({"metric": "offline_rtfx_single_utterance", "unit": "audio-s/s", "higher_is_better": True,
                      "model": f"0.6B architecture, {a.layers} layers, {a.variant}, random init", "data": "synthetic",
                      "timing": "wall clock around nb200_encode_audio (host PCM in, host tokens out), median of reps", "results": rows}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=("b200", "reference"))
    ap.add_argument("--streams", type=int, default=64, help="concurrent streams per GPU")
    ap.add_argument("--rc", type=int, default=13, choices=(0, 1, 6, 13))
    ap.add_argument("--layers", type=int, default=24)
    ap.add_argument("--variant", default="q8_0", choices=("q8_0", "f32dq", "f16"),
                    help="weight format of the GGUF: q8_0 (headline), f16 (the reference converter's 16-bit format; it has no BF16, scripts/convert_to_gguf.py:322-336), f32dq")
    ap.add_argument("--bf16", action="store_true", help="bf16 instead of fp16 operands in the layers' GEMMs (NB200_FLAG_BF16; needs --variant f16 or f32dq)")
    ap.add_argument("--no-latency-path", action="store_true", help="skip the config-3 leg (80 ms chunks, 128 streams per GPU)")
    ap.add_argument("--latency-steps", type=int, default=200)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--offline", action="store_true", help="side benchmark of the offline path (nb200_encode_audio), its own JSON line")
    ap.add_argument("--offline-secs", type=float, nargs="+", default=[5.0, 10.0, 20.0, 40.0])
    ap.add_argument("--offline-reps", type=int, default=10)
    a = ap.parse_args()
    if a.offline:
        return offline_main(a)
    if a.warmup < 3:
        a.warmup = 3
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    cs, T = chunk_samples(a.rc), 1 + a.rc
    workload = f"nemotron-speech-streaming-0.6B ({a.layers} layers, random-init {a.variant}), {a.streams} concurrent streams per GPU, " \
               f"{cs / SR:.2f} s chunks (right_context {a.rc})"
    config = {"workload": workload, "streams_per_gpu": a.streams, "right_context": a.rc, "chunk_seconds": cs / SR, "layers": a.layers,
              "weights": a.variant, "decoder": f"random-init '{BENCH_DECODER}' decoder, blank gain {BENCH_BLANK_GAIN:g}, residual scale {BENCH_RES_SCALE:g}: "
                                               "~0.25 tokens per encoder frame (tools/make_random_gguf.py)", "l2_policy": "inputs larger than L2: weights 1.2 GB + 0.9 GB of K/V caches are re-read every step"}

    import inputs
    gguf_done = Path(inputs.TMP) / f".gguf_ready_{a.layers}_{a.variant}_{BENCH_DECODER}_{BENCH_BLANK_GAIN:g}_{BENCH_RES_SCALE:g}"
    if local == 0:
        gguf = bench_gguf(a.variant, a.layers)
        gguf_done.touch()
    else:
        while not gguf_done.exists():
            time.sleep(0.5)
        gguf = bench_gguf(a.variant, a.layers)

    # ------------------------------------------------------------------------------------------------------
    if a.impl == "reference":
        if rank != 0:
            return
        threads = host_threads()
        run_reference(gguf, a.rc, a.warmup, threads)  # warm-up (page cache, thread pool)
        d = run_reference(gguf, a.rc, a.steps, threads)
        value = d["audio_s"] / d["proc_s"]
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": d["proc_s"] / max(d["chunks"], 1) * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "int8 x int8 blocks (ggml Q8_0) + f32", "data": "synthetic", "impl": "reference", "config": config,
                "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "reference",
                                 "sample": f"1 stream x {d['chunks']} chunks ({d['audio_s']:.1f} s); the reference has no cross-stream batching "
                                           f"(src/nemo-server.cpp:192-271), so its N-stream RTFx equals its 1-stream RTFx; ggml-CPU {d['isa']} build"},
                "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "decode_steps_per_frame": d["decode_iters"] / max(d["chunks"] * T, 1)}
        print(json.dumps(line))
        return

    # ------------------------------------------------------------------------------------------------------
    import torch
    import torch.distributed as dist

    import nb200
    import shard
    import __graft_entry__ as ge
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product path has no CPU fallback)")
    if not (ROOT / "nemotron-asr.cpp_b200" / "libnemob200.so").exists():  # normally prebuilt by __graft_entry__.build()
        if local == 0:
            ge.build_library()
        else:
            while not (ROOT / "nemotron-asr.cpp_b200" / "libnemob200.so").exists():
                time.sleep(1.0)
            time.sleep(2.0)
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local])
        torch.cuda.synchronize()

    n_total_steps = 3 * a.warmup + 4 * a.steps + 8
    need = (a.warmup + a.steps + 2) * cs + 1024
    os.environ["NB200_PCM_RING_LOG2"] = str(max(17, int(np.ceil(np.log2(need)))))
    my_ids = shard.shard_streams(a.streams * world, world, rank)
    assert len(my_ids) == a.streams
    if a.bf16 and a.variant == "q8_0":
        raise SystemExit("--bf16 is for --variant f16 / f32dq (the Q8_0 activation grid does not fit bf16)")
    model = nb200.Model(gguf, device=local, max_streams=max(a.streams, 0 if a.no_latency_path else LATENCY_STREAMS),
                        flags=(nb200.FLAG_ACT_QUANT_OFF | nb200.FLAG_BF16) if a.bf16 else 0)
    streams = [model.stream(a.rc) for _ in my_ids]
    pcm = build_pcm(a.streams, n_total_steps * cs, my_ids[0])
    base_ptr, row_bytes = pcm.data_ptr(), pcm.stride(0) * 2
    cursor = 0  # samples fed per stream so far

    io = model.batch_io(streams)
    row_addr = np.uint64(base_ptr) + np.arange(a.streams, dtype=np.uint64) * np.uint64(row_bytes)

    def feed_chunks(n_chunks: int):
        nonlocal cursor
        io.feed_ptrs(row_addr + np.uint64(cursor * 2), n_chunks * cs)  # nb200_feed_batch: one call for all streams
        cursor += n_chunks * cs

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()

    # ---- leg 1: device-resident inputs -------------------------------------------------------------------
    # one chunk more than the steps consume: a centred STFT frame needs 256 samples beyond its hop (preprocessor.cpp:320-328)
    feed_chunks(a.warmup + a.steps + 1)
    for _ in range(a.warmup):
        assert model.step() == a.streams
    st0 = model.stats()
    barrier()
    model.timer_start()
    t0 = time.perf_counter()
    for _ in range(a.steps):
        n = model.step()
        assert n == a.streams, n
    ms_dev = model.timer_stop()
    wall_ms = (time.perf_counter() - t0) * 1e3
    barrier()
    st1 = model.stats()
    audio_local = a.steps * a.streams * cs / SR
    ms_max, audio_all = shard.reduce_timing(max(ms_dev, wall_ms), audio_local, dev)
    value = audio_all / (ms_max * 1e-3)
    launches = st1.kernel_launches - st0.kernel_launches
    frames = st1.frames - st0.frames
    # symbol steps per frame: every decoded frame costs its blank step plus one step per emitted token (MAX_SYMBOLS cap aside);
    # a decode ROUND evaluates `decode window` consecutive frames of every stream at once, so rounds per chunk can be < T
    tokens = st1.tokens - st0.tokens
    steps_per_frame = (frames + tokens) / max(frames, 1)
    rounds_per_step = (st1.decode_rounds - st0.decode_rounds) / max(a.steps, 1)

    # ---- leg 2: end to end through feed / step / poll ------------------------------------------------------
    # The server-loop order: while the GPU runs step k the host stages the audio of step k+1 (nb200_step_begin /
    # nb200_feed_batch / nb200_step_end / nb200_poll_batch). Every step's PCM goes pinned host -> device and its
    # tokens device -> host inside the timed region.
    for _ in range(a.warmup):
        feed_chunks(1)
        model.step()
        io.poll()
    st2 = model.stats()
    barrier()
    model.timer_start()
    t0 = time.perf_counter()
    feed_chunks(1)
    for k in range(a.steps):
        n = model.step_begin()
        assert n == a.streams, n
        if k + 1 < a.steps:
            feed_chunks(1)
        model.step_end()
        io.poll()
    ms_dev2 = model.timer_stop()
    wall2 = (time.perf_counter() - t0) * 1e3
    barrier()
    st3 = model.stats()
    ms2_max, audio2 = shard.reduce_timing(max(ms_dev2, wall2), audio_local, dev)
    e2e = {"value": audio2 / (ms2_max * 1e-3), "unit": UNIT, "h2d_bytes_per_step": (st3.h2d_bytes - st2.h2d_bytes) // a.steps,
           "d2h_bytes_per_step": (st3.d2h_bytes - st2.d2h_bytes) // a.steps, "ms_per_step": ms2_max / a.steps,
           "order": "step_begin(k) | feed(k+1) | step_end(k) | poll(k)"}
    # chunk latency: the same calls strictly in sequence (feed -> step -> poll), wall clock per chunk
    lat = []
    for _ in range(a.steps):
        t1 = time.perf_counter()
        feed_chunks(1)
        n = model.step()
        assert n == a.streams, n
        io.poll()
        lat.append((time.perf_counter() - t1) * 1e3)

    # ---- leg 3: per-kernel device time (profiled pass, outside the timed regions) ---------------------------
    model.profile(True)
    for _ in range(2):
        feed_chunks(1)
        model.step()
    prof = model.profile_report()
    model.profile(False)
    tot_ms = sum(v["ms"] for v in prof.values()) or 1.0
    pk = peaks()
    gemm_fams = [k for k in prof if k.startswith("gemm_tc")]
    gemm_flops_step = sum(prof[k]["flops"] for k in gemm_fams) / 2
    gemm_launches_step = sum(prof[k]["launches"] for k in gemm_fams) // 2
    gemm_ms_eager = sum(prof[k]["ms"] for k in gemm_fams) / 2
    dom = max(prof, key=lambda k: prof[k]["ms"])
    M = a.streams * T

    # ---- leg 4: what the GEMMs cost INSIDE the step (the roofline figure) ------------------------------------------------------
    # Every tcgen05 GEMM of the step is launched twice inside the graph (nb200_set_dup_family("gemm"); their epilogues are pure
    # stores, so the results do not change); the step-time difference over the same K device-resident steps is the in-graph cost
    # of all GEMM launches: producer -> consumer dependencies, PDL overlap and graph replay included, weights streaming from HBM.
    model.set_dup_family("gemm")
    feed_chunks(a.warmup + a.steps)
    for _ in range(a.warmup):
        assert model.step() == a.streams
    barrier()
    model.timer_start()
    for _ in range(a.steps):
        assert model.step() == a.streams
    ms_dup = model.timer_stop()
    barrier()
    model.set_dup_family(None)
    gemm_ms_step = max((ms_dup - ms_dev) / a.steps, 1e-6)
    ach = gemm_flops_step / (gemm_ms_step * 1e-3) / 1e12
    # beside it: the FFN pair back to back on synthetic operands (a burst of 96 PDL launches per shape, weights cycling through
    # a 300 MB pool): held against the BURST peak
    runs = [model.bench_gemm(M, 4096, 1024, 1, 0, 1, 96), model.bench_gemm(M, 1024, 4096, 5, 0, 1, 96)]
    fl = [2.0 * M * 4096 * 1024, 2.0 * M * 1024 * 4096]
    ach_burst = sum(fl) / (sum(r["us"] for r in runs) * 1e-6) / 1e12
    burst_peak = pk["tflops"]
    try:
        burst_peak = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["bf16_tflops"]
    except Exception:
        pass
    traffic = None
    traffic_own = None
    try:  # DRAM bytes per launch of the dominant GEMM from this round's ncu --set full capture (profiles/ncu_traffic.json)
        if a.streams == 64 and a.rc == 13:
            tj = json.loads((ROOT / "profiles" / "ncu_traffic.json").read_text()).get("gemm_tc_128x256", {})
            traffic, traffic_own = tj.get("traffic"), tj.get("own_operand_bytes")
    except Exception:
        traffic = None
    traffic_note = ("dram bytes per launch of the 128x256 GEMMs (W1 / W2) under ncu --set full: the kernel's own operands (traffic_own_operands: weights + A, no "
                    "re-reads) plus the NEXT GEMM's weights and the layer's K / V slabs that its epilogue warps prefetch into L2; ncu serialises "
                    "kernels with cold caches, so it counts the prefetched bytes here and again in the next kernel, in a step they replace that kernel's reads")
    roofline = {"kernel": "gemm_tc_kernel (all tcgen05 GEMM launches of the step)", "bound": "tensor", "achieved": ach, "peak": pk["tflops"],
                "unit": "TFLOP/s", "frac": ach / pk["tflops"], "traffic": traffic, "traffic_own_operands": traffic_own, "traffic_note": traffic_note, "algorithmic": gemm_flops_step / max(gemm_launches_step, 1),
                "launches_per_step": gemm_launches_step, "avg_launch_us": gemm_ms_step / max(gemm_launches_step, 1) * 1e3,
                "gemm_ms_per_step": gemm_ms_step, "share_of_step": gemm_ms_step / (ms_dev / a.steps), "peak_source": pk["source"],
                "timing": "in-graph marginal cost: K steps with every GEMM launched twice minus the K timed steps (CUDA events on the library's "
                          "stream), divided by the GEMM FLOPs of a step; peak = sustained bf16",
                "burst_microbench": {"achieved": ach_burst, "peak_burst": burst_peak, "frac_of_burst_peak": ach_burst / burst_peak,
                                     "what": "FFN pair (W1 SiLU + re-quantise, W2 split-K), 96 back-to-back launches per shape on synthetic operands",
                                     "shapes": [{"N": 4096, "K": 1024, "us": runs[0]["us"], "bn": runs[0]["bn"], "splits": runs[0]["splits"]},
                                                {"N": 1024, "K": 4096, "us": runs[1]["us"], "bn": runs[1]["bn"], "splits": runs[1]["splits"]}]},
                "eager_events": {"achieved": gemm_flops_step / (gemm_ms_eager * 1e-3) / 1e12, "what": "event pair around every launch of a non-graph pass"},
                "step_tflops": gemm_flops_step / (ms_dev / a.steps * 1e-3) / 1e12}
    kernels = {k: {"share": round(v["ms"] / tot_ms, 4), "launches_per_step": v["launches"] // 2, "avg_us": round(v["ms"] / v["launches"] * 1e3, 2),
                   "tflops": round(v["flops"] / (v["ms"] * 1e-3) / 1e12, 2), "gbs": round(v["bytes"] / (v["ms"] * 1e-3) / 1e9, 1)}
               for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["ms"])}
    for s in streams:
        s.free()

    # ---- leg 5: the latency path (BASELINE config 3): 80 ms chunks, right_context 0, 128 streams per GPU -------------------------
    latency_path = None
    if not a.no_latency_path:
        L_S, cs0 = LATENCY_STREAMS, chunk_samples(0)
        warm0, k0 = 20, a.latency_steps
        ls = [model.stream(0) for _ in range(L_S)]
        pcm0 = build_pcm(L_S, (warm0 + 2 * k0 + 8) * cs0, my_ids[0] * 2)
        io0 = model.batch_io(ls)
        addr0 = np.uint64(pcm0.data_ptr()) + np.arange(L_S, dtype=np.uint64) * np.uint64(pcm0.stride(0) * 2)
        cur0 = 0

        def feed0(n):
            nonlocal cur0
            io0.feed_ptrs(addr0 + np.uint64(cur0 * 2), n * cs0)
            cur0 += n * cs0

        feed0(2)  # one chunk of look-ahead: a centred STFT frame needs 256 samples beyond its hop
        for _ in range(warm0):
            assert model.step() == L_S
            io0.poll()
            feed0(1)
        q0 = model.stats()
        barrier()
        # (a) chunk latency: feed -> step -> poll strictly in sequence, wall clock per 80 ms chunk of all 128 streams
        lat0 = []
        model.timer_start()
        for _ in range(k0):
            t1 = time.perf_counter()
            feed0(1)
            n = model.step()
            assert n == L_S, n
            io0.poll()
            lat0.append((time.perf_counter() - t1) * 1e3)
        ms_seq = model.timer_stop()
        barrier()
        q1 = model.stats()
        # (b) throughput: device-resident audio, K steps back to back
        feed0(k0)
        barrier()
        model.timer_start()
        for _ in range(k0):
            assert model.step() == L_S
        ms_res = model.timer_stop()
        barrier()
        q2 = model.stats()
        ms0_max, audio0 = shard.reduce_timing(ms_res, k0 * L_S * cs0 / SR, dev)
        seq_max, _ = shard.reduce_timing(ms_seq, 0.0, dev)
        kv_bytes = L_S * a.layers * 2 * 70 * 1024 * 2
        w_bytes = a.layers * WEIGHT_BYTES_PER_LAYER
        step_ms = ms0_max / k0
        latency_path = {"workload": f"{L_S} concurrent streams per GPU x 80 ms chunks (right_context 0), {L_S * world} streams on {world} GPU(s)",
                        "ms_per_step": step_ms, "value": audio0 / (ms0_max * 1e-3), "unit": UNIT, "steps": k0,
                        "realtime_duty": step_ms / 80.0, "streams_total": L_S * world,
                        "chunk_latency_ms": {"p50": float(np.percentile(lat0, 50)), "p99": float(np.percentile(lat0, 99)), "max": float(max(lat0)),
                                             "what": "feed (host PCM -> H2D) + step + poll (tokens D2H) of one 80 ms chunk of every stream, in sequence, wall clock, this rank"},
                        "ms_per_step_sequential": seq_max / k0,
                        "gpu_launches_per_step": int((q2.kernel_launches - q1.kernel_launches) // k0),
                        "tokens_per_frame": (q2.tokens - q0.tokens) / max(q2.frames - q0.frames, 1),
                        "hbm": {"bytes_per_step": w_bytes + kv_bytes, "weights": w_bytes, "kv_cache_read": kv_bytes, "achieved_gbs": (w_bytes + kv_bytes) / (step_ms * 1e-3) / 1e9,
                                "peak_gbs": pk["hbm_gbs"], "frac": (w_bytes + kv_bytes) / (step_ms * 1e-3) / 1e9 / pk["hbm_gbs"],
                                "what": "every step streams all fp16 weights and the 70 cached K and V rows of every stream and layer once: the floor of this path"}}
        for s in ls:
            s.free()
    clocks = sampler.stop() if rank == 0 else None
    model.close()
    if world > 1:
        dist.barrier(device_ids=[local])
        dist.destroy_process_group()
    if rank != 0:
        return

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_max / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ("bf16 operands, f32 accumulate" if a.bf16 else
                      "f16 operands, f32 accumulate" + (" (Q8_0 weights dequantised; ggml's int8 activation re-quantisation emulated)" if a.variant == "q8_0" else "")),
            "data": "synthetic", "config": config, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "latency_ms": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)), "what": "feed+step+poll of one batched chunk, in sequence (wall clock)"},
            "decode_steps_per_frame": steps_per_frame, "decode_rounds_per_step": rounds_per_step, "tokens_emitted": int(tokens),
            "frames_decoded": int(frames), "streams_total": a.streams * world,
            "tokens_per_frame": tokens / max(frames, 1), "realtime_streams_supported": value, "roofline": roofline, "kernels": kernels,
            "dominant_family_eager": dom, "latency_path": latency_path}

    if world == 1 and not a.no_cpu_baseline:
        try:
            threads = host_threads()
            run_reference(gguf, a.rc, 2, threads)
            r = run_reference(gguf, a.rc, 12, threads)
            line["cpu_baseline"] = {"value": r["audio_s"] / r["proc_s"], "unit": UNIT, "cores": threads, "kind": "reference",
                                    "sample": f"unmodified reference (ggml-CPU {r['isa']} build), 1 stream x {r['chunks']} chunks ({r['audio_s']:.1f} s) of the same "
                                              f"GGUF and latency mode; it has no cross-stream batching, so its N-stream RTFx equals this 1-stream RTFx",
                                    "ms_per_chunk": r["proc_s"] / max(r["chunks"], 1) * 1e3}
        except Exception as e:  # the checker binary did not travel / cannot run here
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_threads(), "kind": "reference", "sample": f"unavailable: {e}"}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
