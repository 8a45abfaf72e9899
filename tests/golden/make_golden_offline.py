#!/usr/bin/env python3
"""Generates tests/golden/offline_<variant>_s<seed>.npz: the reference's OFFLINE path (nemo_encode: whole-utterance
log-mel -> build_encoder with full-sequence relative-position attention -> greedy_decode with timed tokens,
src/nemo-ggml.cpp:1592-1670) on the deterministic test inputs, through `oracle/_ref/<isa>/ref_harness offline`.

    make -C oracle -j8 && python tests/golden/make_golden_offline.py

Per case: enc [T,1024] (build_encoder output), tokens, frames (timed_token::frame_idx). Q8_0 cases are stored from both
ISA builds (`enc`, `enc_avx512`, ...): their distance is the reference's own noise on the re-quantising path."""
from __future__ import annotations

import json
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT / "tools"))
sys.path.insert(0, str(ROOT / "tests"))
import inputs  # noqa: E402

GOLD = Path(__file__).resolve().parent


def run(isa, gguf, pcm, prefix, secs, lang=None):
    h = ROOT / "oracle" / "_ref" / isa / "ref_harness"
    if not h.exists():
        raise SystemExit(f"{h} missing: run `make -C oracle -j8` first")
    r = subprocess.run([str(h), "offline", gguf, pcm, "8", str(prefix), str(secs), lang or "-"], check=True, stdout=subprocess.PIPE,
                       stderr=subprocess.DEVNULL, text=True)
    info = json.loads(r.stdout.strip().splitlines()[-1])
    return info, {"enc": np.fromfile(f"{prefix}.enc.f32", np.float32).reshape(-1, 1024),
                  "tokens": np.fromfile(f"{prefix}.tokens.i32", np.int32), "frames": np.fromfile(f"{prefix}.frames.i64", np.int64)}


def main():
    manifest = {}
    for variant, seed, secs, lang in inputs.OFFLINE_CASES:
        gguf, pcm = inputs.offline_gguf(variant), inputs.pcm_path(seed, secs)
        info, out = run("avx2", gguf, pcm, inputs.TMP / "off_a", secs, lang)
        if variant == "q8_0":
            _, o2 = run("avx512", gguf, pcm, inputs.TMP / "off_b", secs, lang)
            out.update({k + "_avx512": v for k, v in o2.items()})
        if secs > 10.0:
            out["enc"] = out["enc"][:: inputs.OFFLINE_ENC_STRIDE]
        name = inputs.offline_name(variant, seed, lang)
        np.savez_compressed(GOLD / name, **out)
        manifest[name] = {"variant": variant, "seed": seed, "seconds": secs, "language": lang, "mel_frames": info["mel_frames"], "enc_T": info["enc_T"],
                          "tokens": info["tokens"]}
        print(name, manifest[name])
    (GOLD / "manifest_offline.json").write_text(json.dumps(manifest, indent=1) + "\n")


if __name__ == "__main__":
    main()
