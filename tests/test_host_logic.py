"""CPU-side checks: the C ABI library loads and exports every symbol include/nb200.h declares, fails loudly without a
GPU, the latency-mode arithmetic, the GGUF writer/reader pair, and the stream->GPU sharding under a 2-rank gloo group."""
import ctypes
import os
import re
import socket
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

import inputs
import nemo_oracle as no

ROOT = Path(__file__).resolve().parents[1]


def test_header_symbols_exported(lib_built):
    nb200 = lib_built
    header = (ROOT / "include" / "nb200.h").read_text()
    declared = set(re.findall(r"\b(nb200_[a-z0-9_]+)\s*\(", header))
    assert len(declared) >= 25
    lib = ctypes.CDLL(str(nb200.LIB_PATH))
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} is declared in include/nb200.h but not exported by libnemob200.so"
    assert declared == set(nb200.SYMBOLS), "nb200.py must bind exactly the header's entry points"
    assert nb200.lib().nb200_version() >= 100


def test_no_cpu_fallback(lib_built, gguf_q8):
    """Without a CUDA device the product path must fail loudly (there is no CPU fallback of any kind)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    nb200 = lib_built
    with pytest.raises(nb200.NB200Error, match="no CUDA device"):
        nb200.Model(gguf_q8)


def test_product_does_not_import_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may touch oracle/: not the product, not tools/."""
    for p in (ROOT / "nemotron-asr.cpp_b200").rglob("*"):
        if p.suffix in (".py", ".cu", ".cuh", ".cpp", ".h") and p.is_file():
            txt = p.read_text(errors="ignore")
            assert "nemo_oracle" not in txt and "oracle/" not in txt, f"{p} references the test oracle"
    for p in (ROOT / "tools").glob("*.py"):
        txt = p.read_text(errors="ignore")
        assert "nemo_oracle" not in txt and "_ref" not in txt and '"oracle"' not in txt, f"{p} uses the test oracle"


def test_latency_modes():
    """src/nemo-stream.h:65-100 (the table in SURVEY §8): chunk_mel, shift, subsampled lengths, chunk_len."""
    table = {0: (17, 8, (9, 5, 3), 1), 1: (25, 16, (13, 7, 4), 2), 6: (65, 56, (33, 17, 9), 7), 13: (121, 112, (61, 31, 16), 14)}
    for rc, (chunk_mel, shift, ts, T) in table.items():
        s = no.StreamOracle({"nemo.n_layers": 0, "nemo.n_heads": 8, "nemo.d_head": 128, "nemo.d_model": 1024,
                             "nemo.att_left_context": 70, "nemo.vocab_size": 1025},
                            {"preprocessor.featurizer.fb": np.zeros((128, 257), np.float32),
                             "preprocessor.featurizer.window": np.zeros(400, np.float32)}, rc)
        assert (s.chunk_mel, s.shift, s.valid_out) == (chunk_mel, shift, T)
        t1 = chunk_mel // 2 + 1
        t2 = t1 // 2 + 1
        t3 = t2 // 2 + 1
        assert (t1, t2, t3) == ts and t3 - 2 == T


def test_gguf_roundtrip(oracle_q8, oracle_f32):
    """Writer (tools/make_random_gguf.py, the converter's format) -> oracle reader: Q8_0 blocks dequantise to exactly the
    values the F32-of-dequantised file stores."""
    kv, tq = oracle_q8
    _, tf = oracle_f32
    assert kv["nemo.n_layers"] == inputs.TEST_LAYERS and kv["nemo.vocab_size"] == 1025 and len(kv["tokenizer.vocab_list"]) == 1024
    name = "encoder.layers.1.feed_forward2.linear1.weight"
    assert isinstance(tq[name], no.Q8Weight) and tq[name].shape == (4096, 1024)
    assert np.array_equal(tq[name].dequant(), tf[name])
    assert tq["encoder.layers.0.conv.depthwise_conv.weight"].shape == (9, 1024)


def test_shard_partition():
    import shard
    for world in (1, 2, 4, 8):
        parts = [shard.shard_streams(1000, world, r) for r in range(world)]
        assert sorted(sum(parts, [])) == list(range(1000))
        assert max(map(len, parts)) - min(map(len, parts)) <= 1
    with pytest.raises(ValueError):
        shard.shard_streams(8, 2, 2)


_WORKER = r"""
import os, sys
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
import shard
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
mine = shard.shard_streams(129, w, r)
gathered = [None] * w
dist.all_gather_object(gathered, mine)
assert sorted(sum(gathered, [])) == list(range(129))
ms, units = shard.reduce_timing(10.0 + 5.0 * r, float(len(mine)))
assert ms == 10.0 + 5.0 * (w - 1) and units == 129.0, (ms, units)
dist.barrier()
dist.destroy_process_group()
print("rank", r, "ok")
"""


def test_two_rank_gloo(tmp_path):
    """The N>1 host path (sharding + max-over-ranks timing reduction) under torch.distributed/gloo, world_size 2."""
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    with socket.socket() as sk:
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), str(script), str(ROOT / "nemotron-asr.cpp_b200")],
                       env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert r.stdout.count("ok") == 2


def test_server_loop_header_compiles(lib_built, tmp_path):
    """nemotron-asr.cpp_b200/server_loop.h and its GPU test driver build against include/nb200.h and link with the library
    (no GPU needed to compile and link; the driver itself runs in the gpu suite)."""
    import subprocess
    root = Path(__file__).resolve().parents[1]
    pkg = root / "nemotron-asr.cpp_b200"
    exe = tmp_path / "server_loop_test"
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I", str(root / "include"), "-I", str(pkg),
                    str(root / "tests" / "server_loop_test.cpp"), "-o", str(exe), "-L", str(pkg), "-lnemob200", f"-Wl,-rpath,{pkg}", "-lpthread"],
                   check=True)
    assert exe.exists()


# ---- GGUF container: the host-only parser nb200_model_load uses (nb200_gguf_check needs no GPU) --------------------------
def _tiny_gguf(n_tensors=1, n_kv=1, nd=1, dims=(64,), ttype=0, data_bytes=256, str_len=None, vocab_key=b"tokenizer.vocab_list",
               legacy_blob=None):
    """A minimal GGUF v3 image: one vocabulary KV and one tensor "x"; every field of the header can be corrupted."""
    import struct

    def s(b, n=None):
        return struct.pack("<Q", len(b) if n is None else n) + b

    out = b"GGUF" + struct.pack("<I", 3) + struct.pack("<q", n_tensors) + struct.pack("<q", n_kv)
    if legacy_blob is not None:
        out += s(b"tokenizer.vocab") + struct.pack("<I", 8) + s(legacy_blob)
    else:
        out += s(vocab_key) + struct.pack("<I", 9) + struct.pack("<I", 8) + struct.pack("<Q", 2) + s(b"a", str_len) + s(b"bc")
    out += s(b"x") + struct.pack("<I", nd)
    for d in dims:
        out += struct.pack("<q", d)
    out += struct.pack("<i", ttype) + struct.pack("<Q", 0)
    out += b"\0" * ((len(out) + 31) // 32 * 32 - len(out))
    return out + b"\0" * data_bytes


def test_gguf_check_accepts_and_rejects(lib_built, tmp_path):
    """Well-formed files parse; truncated or crafted headers (wrapping 64-bit lengths, absurd counts, negative or overflowing
    dimensions, a quantised tensor that is not a whole number of blocks, more than 4 dimensions, data outside the file) are
    refused with an error instead of reading outside the mapping."""
    nb200 = lib_built

    def check(img):
        p = tmp_path / "t.gguf"
        p.write_bytes(img)
        return nb200.gguf_check(str(p))

    ok = check(_tiny_gguf())
    assert ok["tensors"] == 1 and ok["tensor_bytes"] == 256 and ok["vocab_source"] == "tokenizer.vocab_list" and ok["first_pieces"] == ["a", "bc"]
    assert check(_tiny_gguf(ttype=8, dims=(64,), data_bytes=68))["tensor_bytes"] == 68  # 2 Q8_0 blocks
    bad = {
        "truncated": _tiny_gguf()[:40],
        "string length wraps": _tiny_gguf(str_len=2 ** 64 - 8),
        "string past the end": _tiny_gguf(str_len=1 << 40),
        "absurd kv count": _tiny_gguf(n_kv=2 ** 62),
        "negative tensor count": _tiny_gguf(n_tensors=-5),
        "negative dimension": _tiny_gguf(dims=(-64,)),
        "dimension product overflows": _tiny_gguf(nd=3, dims=(2 ** 31, 2 ** 31, 2 ** 31)),
        "five dimensions": _tiny_gguf(nd=5, dims=(2, 2, 2, 2, 2)),
        "q8_0 with a partial block": _tiny_gguf(ttype=8, dims=(40,)),
        "data outside the file": _tiny_gguf(dims=(1 << 20,)),
        "unknown tensor type": _tiny_gguf(ttype=77),
        "no vocabulary": _tiny_gguf(vocab_key=b"tokenizer.something"),
        "bad magic": b"GGUX" + _tiny_gguf()[4:],
    }
    for what, img in bad.items():
        with pytest.raises(nb200.NB200Error):
            check(img)
            pytest.fail(f"accepted a malformed file: {what}")
    with pytest.raises(nb200.NB200Error, match="cannot open"):
        nb200.gguf_check(str(tmp_path / "missing.gguf"))


def test_legacy_vocab_blob(lib_built, tmp_path):
    """The loader's fallback vocabulary (src/nemo-ggml.cpp:156-165): `tokenizer.vocab` = vocab_size - 1 pieces of 8 bytes,
    NUL-padded. The generator's --legacy-vocab file must read back the pieces of the string-array file (cut to 8 bytes)."""
    nb200 = lib_built
    tiny = tmp_path / "l.gguf"
    tiny.write_bytes(_tiny_gguf(legacy_blob=b"ab\0\0\0\0\0\0" + "▁t3".encode() + b"\0\0\0" + b"12345678" + b"tail"))
    d = nb200.gguf_check(str(tiny))
    assert d["vocab_source"] == "tokenizer.vocab" and d["vocab_pieces"] == 3 and d["first_pieces"] == ["ab", "▁t3", "12345678"]
    import make_random_gguf as mg
    a, b = tmp_path / "list.gguf", tmp_path / "blob.gguf"
    mg.write_gguf(str(a), 7, 0, "f32")
    mg.write_gguf(str(b), 7, 0, "f32", legacy_vocab=True)
    da, db = nb200.gguf_check(str(a)), nb200.gguf_check(str(b))
    assert da["vocab_source"] == "tokenizer.vocab_list" and db["vocab_source"] == "tokenizer.vocab"
    assert da["vocab_pieces"] == db["vocab_pieces"] == 1024 and da["first_pieces"] == db["first_pieces"] == ["▁t0", "t1", "t2", "▁t3"]
    assert da["tensors"] == db["tensors"] and da["tensor_bytes"] == db["tensor_bytes"] and da["n_layers"] == 0


def test_bench_decoder_profile():
    """bench.py's random-init decoder (make_random_gguf decoder="stationary" + the committed encoder-mean calibration): greedy
    decode by the oracle over seeded stand-ins for the varying part of the encoder output must give the profile of real speech
    (a fraction of a token per frame, hardly ever more than 3 symbols on a frame), not the all-blank / 10-symbol-cap pair."""
    import make_random_gguf as mg
    m = mg.encoder_mean(inputs.SEED, 24, 0.1)
    assert m is not None and m.shape == (1024,), "tools/decoder_calibration.npz lost its entry for the benchmark model"
    dec = {n: a for n, a, _ in mg.tensor_list(inputs.SEED, 24, 6.0, "stationary", 0.1) if n.startswith(("decoder.", "joint."))}
    o = no.StreamOracle({"nemo.n_layers": 0, "nemo.n_heads": 8, "nemo.d_head": 128, "nemo.d_model": 1024, "nemo.att_left_context": 70,
                         "nemo.vocab_size": 1025}, dict({"preprocessor.featurizer.fb": np.zeros((128, 257), np.float32),
                                                         "preprocessor.featurizer.window": np.zeros(400, np.float32)}, **dec), 13)
    rng = np.random.default_rng(3)
    low = rng.standard_normal((16, 1024)).astype(np.float32)  # the varying part lives in a few directions and changes smoothly
    hist, n, frames = np.zeros(11, int), 0, 600
    state = np.zeros(16, np.float32)
    for f in range(frames):
        state = 0.6 * state + 0.8 * rng.standard_normal(16).astype(np.float32)
        enc = m + 0.03 * (state @ low) + 0.08 * rng.standard_normal(1024).astype(np.float32)  # ~5 from the mean, like the measured frames
        k = len(o.decode_frame(enc.astype(np.float32)))
        hist[k] += 1
        n += k
    print(f"stationary decoder on synthetic encoder frames: {n / frames:.3f} tokens/frame, histogram {hist.tolist()}")
    assert 0.03 < n / frames < 1.0 and hist[4:].sum() <= 0.02 * frames and hist[10] <= 2


def test_weight_cache_format_round_trip(lib_built, tmp_path):
    """The native weight file (csrc/weight_cache.cpp) without a GPU: blobs written by the writer come back
    bit-identical and 256-byte aligned through the mmap reader; a foreign source stamp, a wrong size request and files truncated
    at the header, inside a blob, inside the index and one byte short are all rejected."""
    import nb200
    rc = nb200.lib().nb200_weight_cache_selftest(os.fsencode(str(tmp_path / "selftest.nb200w")))
    assert rc == 0, nb200._err()
    assert not (tmp_path / "selftest.nb200w").exists()
