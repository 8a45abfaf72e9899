#!/usr/bin/env python3
"""Generates tests/golden/text_timed.json: the reference's tokens_to_text (src/nemo-ggml.cpp:1556-1583) over
timed_token{token_id, frame_idx} pairs (src/nemo-ggml.h:383-395), with and without timestamp_words, through
`oracle/_ref/<isa>/ref_harness text` and the vocabulary of the deterministic test GGUF (tools/make_random_gguf.py).

    make -C oracle -j8 && python tests/golden/make_golden_text.py

The token / frame lists cover: pieces with and without the leading U+2581, out-of-range ids (skipped), frame 0, frames
whose time is not exactly representable in fp32 (rounding of %.2f), large frame indices (an hour of audio)."""
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


def main():
    harness = ROOT / "oracle" / "_ref" / "avx2" / "ref_harness"
    if not harness.exists():
        raise SystemExit(f"{harness} missing: run `make -C oracle -j8` first")
    gguf = inputs.gguf_path("f32dq")
    rng = np.random.default_rng(77)
    tokens = [0, 1, 2, 3, 6, 1023, 1024, -1, 5000, 9, 12, 13]
    frames = [0, 0, 1, 3, 7, 13, 14, 15, 16, 17, 44999, 45000]
    tokens += rng.integers(0, 1024, 60).tolist()
    frames += np.sort(rng.integers(18, 4000, 60)).tolist()
    tok_p, frm_p = inputs.TMP / "text_tokens.i32", inputs.TMP / "text_frames.i64"
    np.asarray(tokens, np.int32).tofile(tok_p)
    np.asarray(frames, np.int64).tofile(frm_p)
    out = {"tokens": tokens, "frames": frames}
    for flag, key in ((0, "text_plain"), (1, "text_timed")):
        dst = inputs.TMP / f"text_{flag}.txt"
        subprocess.run([str(harness), "text", gguf, str(tok_p), str(frm_p), str(flag), str(dst)], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        out[key] = dst.read_bytes().decode("utf-8")
    (GOLD / "text_timed.json").write_text(json.dumps(out, ensure_ascii=False, indent=1) + "\n")
    print(out["text_timed"][:200])


if __name__ == "__main__":
    main()
