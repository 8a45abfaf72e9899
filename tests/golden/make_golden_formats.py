#!/usr/bin/env python3
"""Golden fixtures for the converter's other on-disk formats (F16, Q4_0; scripts/convert_to_gguf.py:157-204,322-336 in the
reference), produced like make_golden.py by the UNMODIFIED reference (oracle/_ref/<isa>/ref_harness):

    make -C oracle -j8 && python tests/golden/make_golden_formats.py

  stream_ml_<lang>_rc<rc>.npz  multilingual GGUF (num_prompts = 128, prompt_kernel.*): nemo_stream_set_language + prompt fusion
  stream_<variant>_rc<rc>.npz  per-chunk encoder_out, tokens, tokens per chunk; both ISA builds for q4_0 (its activations
                               are re-quantised to Q8_0 like the Q8_0 path, so it has the same self-noise), avx2 only for f16.
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT / "tools"))
sys.path.insert(0, str(ROOT / "tests"))
sys.path.insert(0, str(ROOT / "tests" / "golden"))
import inputs  # noqa: E402
from make_golden import GOLD, ISAS, TMP, harness, run, sha  # noqa: E402


def main():
    manifest = {"reference": "m1el/nemotron-asr.cpp @ 4a863ff, ggml @ 9be3133", "inputs": {}, "fixtures": {}}
    for variant, rc, seed, secs in inputs.FORMAT_CASES:
        path = inputs.gguf_path(variant)
        manifest["inputs"][variant] = {"sha256": sha(path), "bytes": Path(path).stat().st_size}
        T = 1 + rc
        isas = ISAS if variant == "q4_0" else ("avx2",)
        d = {}
        for isa in isas:
            pref = TMP / f"gold_{variant}_rc{rc}_{isa}"
            info = run([harness(isa), "stream", path, inputs.pcm_path(seed, secs), str(rc), "4", str(pref)])
            sfx = "" if isa == isas[0] else "_alt"
            d["enc" + sfx] = np.fromfile(str(pref) + ".enc.f32", dtype=np.float32).reshape(-1, T, 1024)
            d["tokens" + sfx] = np.fromfile(str(pref) + ".tokens.i32", dtype=np.int32)
            d["chunk_tokens" + sfx] = np.fromfile(str(pref) + ".chunk_tokens.i32", dtype=np.int32)
            assert d["enc" + sfx].shape[0] == info["chunks"]
        name = f"stream_{variant}_rc{rc}.npz"
        np.savez_compressed(GOLD / name, **d)
        e = {"pcm": f"seed {seed}, {secs} s", "chunks": int(d["enc"].shape[0]), "tokens": int(d["tokens"].size), "isas": list(isas)}
        if "enc_alt" in d:
            e["self_noise_rel_l2"] = [round(float(np.linalg.norm(a - b) / np.linalg.norm(a)), 5) for a, b in zip(d["enc"], d["enc_alt"])]
        manifest["fixtures"][name] = e
        print(name, e)
    # ---- multilingual checkpoint: prompt fusion with the default prompt and with explicit languages ----
    path = inputs.gguf_path("f32dq", multilingual=True)
    manifest["inputs"]["f32dq_ml"] = {"sha256": sha(path), "bytes": Path(path).stat().st_size}
    for lang, rc, seed, secs in inputs.PROMPT_CASES:
        T = 1 + rc
        pref = TMP / f"gold_ml_{lang}_rc{rc}"
        info = run([harness("avx2"), "stream", path, inputs.pcm_path(seed, secs), str(rc), "4", str(pref), "0", lang or "-"])
        d = {"enc": np.fromfile(str(pref) + ".enc.f32", dtype=np.float32).reshape(-1, T, 1024),
             "tokens": np.fromfile(str(pref) + ".tokens.i32", dtype=np.int32),
             "chunk_tokens": np.fromfile(str(pref) + ".chunk_tokens.i32", dtype=np.int32)}
        assert d["enc"].shape[0] == info["chunks"]
        name = f"stream_ml_{lang or 'default'}_rc{rc}.npz"
        np.savez_compressed(GOLD / name, **d)
        manifest["fixtures"][name] = {"pcm": f"seed {seed}, {secs} s", "language": lang, "chunks": int(d["enc"].shape[0]), "tokens": int(d["tokens"].size)}
        print(name, manifest["fixtures"][name])
    (GOLD / "manifest_formats.json").write_text(json.dumps(manifest, indent=1) + "\n")


if __name__ == "__main__":
    main()
