#!/usr/bin/env python3
"""Seeded synthetic 16 kHz s16le audio: a few AM/FM chirps in 100-4000 Hz under a syllable-rate envelope with
silence gaps, plus white noise at about -45 dBFS everywhere (keeps every mel bin above the fp32 FFT noise floor,
so a 1e-4 absolute log-mel check is meaningful). Peak about -6 dBFS. Pure function of (seed, seconds)."""
from __future__ import annotations

import argparse
import sys

import numpy as np

SR = 16000


def make_pcm(seconds: float, seed: int = 0) -> np.ndarray:
    n = int(round(seconds * SR))
    rng = np.random.default_rng([seed, 0xA5D10])
    t = np.arange(n, dtype=np.float64) / SR
    x = np.zeros(n)
    for _ in range(int(rng.integers(3, 6))):
        f0 = rng.uniform(100, 4000)
        f1 = rng.uniform(100, 4000)
        fm_rate = rng.uniform(0.3, 3.0)
        fm_depth = rng.uniform(10, 200)
        inst = f0 + (f1 - f0) * (t / max(seconds, 1e-3)) + fm_depth * np.sin(2 * np.pi * fm_rate * t)
        phase = 2 * np.pi * np.cumsum(inst) / SR + rng.uniform(0, 2 * np.pi)
        am = 0.6 + 0.4 * np.sin(2 * np.pi * rng.uniform(1.0, 8.0) * t + rng.uniform(0, 2 * np.pi))
        x += rng.uniform(0.3, 1.0) * am * np.sin(phase)
    syl = rng.uniform(3.0, 5.0)
    env = np.clip(0.5 + 0.8 * np.sin(2 * np.pi * syl * t + rng.uniform(0, 2 * np.pi)), 0.0, 1.0)
    gate = (np.sin(2 * np.pi * rng.uniform(0.2, 0.5) * t + rng.uniform(0, 2 * np.pi)) > -0.75).astype(np.float64)
    x = x * env * gate
    x = x / max(np.max(np.abs(x)), 1e-9) * 0.5
    x += rng.standard_normal(n) * 10 ** (-45 / 20)
    return np.clip(np.round(x * 32767.0), -32768, 32767).astype(np.int16)


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("out")
    ap.add_argument("--seconds", type=float, default=10.0)
    ap.add_argument("--seed", type=int, default=0)
    a = ap.parse_args(argv)
    make_pcm(a.seconds, a.seed).tofile(a.out)
    print({"path": a.out, "seconds": a.seconds, "seed": a.seed})


if __name__ == "__main__":
    sys.exit(main())
