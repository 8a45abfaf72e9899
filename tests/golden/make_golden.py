#!/usr/bin/env python3
"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/<isa>/ref_harness, built from
/root/reference by oracle/Makefile) on deterministic inputs. Run in the build container (needs /root/reference
only through the already-built oracle/_ref binaries):

    make -C oracle -j8 && python tests/golden/make_golden.py

Inputs are pure functions of seeds (tools/make_random_gguf.py, tools/make_pcm.py), so the tests regenerate them
byte-identically on any box; manifest.json records their sha256 so a drifted generator is caught.

Fixtures (all produced by the reference's own code paths):
  mel_s11.npz            nemo_preprocessor_process (src/preprocessor.cpp:330) on 4 s of PCM, fed in 1280- and
                         180-sample slices (the slice size must not matter; tests/test_preprocessor.cpp:100-105)
  matmul_<variant>.npz   ggml_mul_mat of encoder weight tensors (Q8_0 and F32) with seeded activations, from the avx2
                         and the avx512 build: pins the oracle's block-int8 restatement to ~1e-7
  stream_<variant>_rc<rc>.npz  nemo_stream_process_incremental + nemo_stream_finalize (src/nemo-stream.cpp:1145-1293):
                         per-chunk encoder_out (all T frames), tokens, tokens per chunk. Q8_0 runs are stored from
                         BOTH ISA builds: their mutual distance is the reference's own self-noise (activation
                         re-quantisation makes the Q8_0 path chaotic at the ~7e-3 level).
"""
from __future__ import annotations

import hashlib
import json
import subprocess
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT / "tools"))
sys.path.insert(0, str(ROOT / "tests"))
import inputs  # noqa: E402  (tests/inputs.py: the deterministic input set shared with the tests)

GOLD = Path(__file__).resolve().parent
TMP = inputs.TMP
ISAS = ("avx2", "avx512")


def harness(isa: str) -> str:
    p = ROOT / "oracle" / "_ref" / isa / "ref_harness"
    if not p.exists():
        raise SystemExit(f"{p} missing: run `make -C oracle -j8` first")
    return str(p)


def run(cmd) -> dict:
    r = subprocess.run(cmd, check=True, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    return json.loads(r.stdout.strip().splitlines()[-1])


def sha(path) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        for blk in iter(lambda: f.read(1 << 22), b""):
            h.update(blk)
    return h.hexdigest()


def main():
    paths = inputs.ensure_inputs()
    manifest = {"reference": "m1el/nemotron-asr.cpp @ 4a863ff, ggml @ 9be3133", "inputs": {}, "fixtures": {}}
    for k, p in paths.items():
        manifest["inputs"][k] = {"sha256": sha(p), "bytes": Path(p).stat().st_size}

    # ---- mel ------------------------------------------------------------------------------------------
    pcm = inputs.pcm_path(11, 4.0)
    mels = {}
    for sl in (1280, 180):
        out = TMP / f"gold_mel_{sl}.f32"
        run([harness("avx2"), "mel", paths["q8_0"], pcm, str(sl), str(out)])
        mels[sl] = np.fromfile(out, dtype=np.float32).reshape(-1, 128)
    assert mels[1280].shape == mels[180].shape and np.array_equal(mels[1280], mels[180]), "mel depends on slice size?"
    np.savez_compressed(GOLD / "mel_s11.npz", mel=mels[1280])
    manifest["fixtures"]["mel_s11.npz"] = {"frames": int(mels[1280].shape[0]), "pcm": "seed 11, 4.0 s"}

    # ---- single matmuls ------------------------------------------------------------------------------
    for variant in ("q8_0", "f32dq"):
        out = {}
        for name, M in inputs.MATMUL_CASES:
            x = inputs.matmul_input(name, M)
            xf = TMP / "gold_x.f32"
            x.tofile(xf)
            for isa in ISAS:
                yf = TMP / "gold_y.f32"
                info = run([harness(isa), "matmul", paths[variant], name, str(xf), str(M), str(yf), "1"])
                out[f"{name}|{M}|{isa}"] = np.fromfile(yf, dtype=np.float32).reshape(M, info["N"])
        np.savez_compressed(GOLD / f"matmul_{variant}.npz", **out)
        manifest["fixtures"][f"matmul_{variant}.npz"] = {"cases": [list(c) for c in inputs.MATMUL_CASES], "isas": list(ISAS)}

    # ---- streams --------------------------------------------------------------------------------------
    for variant, rc, seed, secs in inputs.STREAM_CASES:
        T = 1 + rc
        isas = ISAS if variant == "q8_0" else ("avx2",)
        d = {}
        for isa in isas:
            pref = TMP / f"gold_{variant}_rc{rc}_{isa}"
            info = run([harness(isa), "stream", paths[variant], inputs.pcm_path(seed, secs), str(rc), "4", str(pref)])
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
            e["tokens_equal_across_isas"] = bool(np.array_equal(d["tokens"], d["tokens_alt"]))
        manifest["fixtures"][name] = e
        print(name, e)
    (GOLD / "manifest.json").write_text(json.dumps(manifest, indent=1) + "\n")
    total = sum(p.stat().st_size for p in GOLD.glob("*.npz"))
    print(f"wrote {len(list(GOLD.glob('*.npz')))} fixtures, {total / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
