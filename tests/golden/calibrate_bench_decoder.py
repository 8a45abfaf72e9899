#!/usr/bin/env python3
"""Calibration of the benchmark's random-init decoder (tools/make_random_gguf.py, decoder="stationary"). Test infrastructure
(it executes the compiled reference, oracle/_ref, and the numpy oracle's greedy decode): run in the build container; its
output tools/decoder_calibration.npz (4 KB: one 1024-vector per calibrated model) is committed and is plain data for the generator.

Two things are measured on the benchmark's model (24 layers, res_scale 0.1, seed 1234) and its kind of audio (bench.py
build_pcm: 8 s seeded signals, tiled):
  1. `--measure-mean`: the time-average of the encoder output (a model constant to within the frame-to-frame variation: the
     LayerNorm'd output of a random stack is a fixed vector plus a part that follows the audio). It is stored in
     tools/decoder_calibration.npz and folded into joint.enc.bias by the generator, so that the joint sees the varying part only
     and every frame draws new logits (otherwise the same few tokens stay above the blank and the greedy loop ping-pongs
     between them up to the 10-symbol cap).
  2. default: tokens per frame and symbols-per-frame histogram of the oracle's greedy decode for a list of blank gains, on
     audio the mean was NOT measured on.

    python tests/golden/calibrate_bench_decoder.py --measure-mean && python tests/golden/calibrate_bench_decoder.py --gains 5.6 6.0 6.4
"""
from __future__ import annotations

import argparse
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
for p in (ROOT / "tests", ROOT / "tools", ROOT / "oracle"):
    sys.path.insert(0, str(p))
CAL = ROOT / "tools" / "decoder_calibration.npz"


def bench_like_pcm(seed: int, seconds: float) -> str:
    import inputs
    p = inputs.TMP / f"cal_pcm_s{seed}_{seconds:g}.pcm"
    if not p.exists():
        sig = np.roll(inputs.pcm(seed, 8.0), 997 * seed)
        n = int(seconds * 16000)
        np.tile(sig, n // sig.size + 1)[:n].tofile(p)
    return str(p)


def encoder_output(gguf: str, seed: int, seconds: float) -> np.ndarray:
    import inputs
    pref = inputs.TMP / f"cal_enc_{Path(gguf).stem}_s{seed}"
    if not Path(str(pref) + ".enc.f32").exists():
        subprocess.run([str(ROOT / "oracle/_ref/avx2/ref_harness"), "stream", gguf, bench_like_pcm(seed, seconds), "13", "16", str(pref)],
                       check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return np.fromfile(str(pref) + ".enc.f32", dtype=np.float32).reshape(-1, 1024)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--measure-mean", action="store_true")
    ap.add_argument("--gains", type=float, nargs="+", default=[6.3])
    ap.add_argument("--layers", type=int, default=24)
    ap.add_argument("--variant", default="q8_0")
    ap.add_argument("--res-scale", type=float, default=0.1)
    ap.add_argument("--seconds", type=float, default=40.0)
    ap.add_argument("--mean-seeds", type=int, nargs="+", default=[100, 101, 102])
    ap.add_argument("--eval-seeds", type=int, nargs="+", default=[103, 104, 105, 106])
    a = ap.parse_args()
    import inputs
    import make_random_gguf as mg
    import nemo_oracle as no
    gguf = inputs.gguf_path(a.variant, a.layers, decoder="stationary", res_scale=a.res_scale)
    key = mg.calibration_key(inputs.SEED, a.layers, a.res_scale)
    if a.measure_mean:
        encs = [encoder_output(gguf, s, a.seconds) for s in a.mean_seeds]
        means = np.stack([e[28:].mean(0) for e in encs])  # skip the first two chunks (cache warm-up)
        m = means.mean(0).astype(np.float32)
        dyn = float(np.mean([np.sqrt(((e[28:] - m) ** 2).sum(1).mean()) for e in encs]))
        print(f"{key}: |mean| {np.linalg.norm(m):.2f}, spread between signals {[round(float(np.linalg.norm(x - m)), 2) for x in means]}, "
              f"rms distance of a frame from the mean {dyn:.2f}")
        d = dict(np.load(CAL)) if CAL.exists() else {}
        d[key] = m
        np.savez_compressed(CAL, **d)
        print(f"wrote {CAL}; regenerate the GGUF (delete {gguf}) before calibrating the gain")
        return
    kv, t = no.load_gguf(inputs.gguf_path("f32dq"))  # only the decoder / joint tensors are used, replaced below
    for gain in a.gains:
        dec = {n: arr for n, arr, _ in mg.tensor_list(inputs.SEED, a.layers, gain, "stationary", a.res_scale) if n.startswith(("decoder.", "joint."))}
        rates, hist = [], np.zeros(11, int)
        for s in a.eval_seeds:
            enc = encoder_output(gguf, s, a.seconds)
            o = no.StreamOracle(kv, dict(t, **dec), 13)
            n = 0
            for f in range(enc.shape[0]):
                k = len(o.decode_frame(enc[f]))
                hist[k] += 1
                n += k
            rates.append(round(n / enc.shape[0], 3))
        print(f"stationary gain {gain}: tokens/frame {np.mean(rates):.3f} (per signal {rates}); symbols per frame histogram 0..10: {hist.tolist()}")


if __name__ == "__main__":
    main()
