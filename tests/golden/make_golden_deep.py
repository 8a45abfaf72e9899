#!/usr/bin/env python3
"""Golden fixtures of the FULL-DEPTH model (24 conformer layers: what bench.py times), produced like make_golden.py by the
UNMODIFIED reference (oracle/_ref/<isa>/ref_harness):

    make -C oracle -j8 && python tests/golden/make_golden_deep.py

  stream_deep_f32dq_rc13.npz  per-chunk encoder_out, tokens, tokens per chunk of a 3.6 s stream (3 chunks + finalize tail)
  stream_deep_q8_0_rc13.npz   the same from both ISA builds: their mutual distance is the reference's self-noise at 24 layers
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
    for variant, rc, seed, secs in inputs.DEEP_CASES:
        path = inputs.gguf_path(variant, inputs.DEEP_LAYERS)
        manifest["inputs"][f"{variant}_L{inputs.DEEP_LAYERS}"] = {"sha256": sha(path), "bytes": Path(path).stat().st_size}
        T = 1 + rc
        isas = ISAS if variant == "q8_0" else ("avx2",)
        d = {}
        for isa in isas:
            pref = TMP / f"gold_deep_{variant}_rc{rc}_{isa}"
            info = run([harness(isa), "stream", path, inputs.pcm_path(seed, secs), str(rc), "8", str(pref)])
            sfx = "" if isa == isas[0] else "_alt"
            d["enc" + sfx] = np.fromfile(str(pref) + ".enc.f32", dtype=np.float32).reshape(-1, T, 1024)
            d["tokens" + sfx] = np.fromfile(str(pref) + ".tokens.i32", dtype=np.int32)
            d["chunk_tokens" + sfx] = np.fromfile(str(pref) + ".chunk_tokens.i32", dtype=np.int32)
            assert d["enc" + sfx].shape[0] == info["chunks"]
        name = f"stream_deep_{variant}_rc{rc}.npz"
        np.savez_compressed(GOLD / name, **d)
        e = {"pcm": f"seed {seed}, {secs} s", "layers": inputs.DEEP_LAYERS, "chunks": int(d["enc"].shape[0]), "tokens": int(d["tokens"].size),
             "isas": list(isas)}
        if "enc_alt" in d:
            e["self_noise_rel_l2"] = [round(float(np.linalg.norm(a - b) / np.linalg.norm(a)), 5) for a, b in zip(d["enc"], d["enc_alt"])]
            e["tokens_equal_across_isas"] = bool(np.array_equal(d["tokens"], d["tokens_alt"]))
        manifest["fixtures"][name] = e
        print(name, e)
    (GOLD / "manifest_deep.json").write_text(json.dumps(manifest, indent=1) + "\n")


if __name__ == "__main__":
    main()
