#!/usr/bin/env python3
"""Developer GPU check (run under gpurun): exercises every kernel family against the numpy oracle and the compiled
reference, each phase in its own subprocess with a timeout so a hung kernel cannot hang the box.

  python tools/gpu_check.py all            # runs the phases below, writes gpurun_out/gpu_check.json
  python tools/gpu_check.py phase <name>   # one phase in-process
"""
from __future__ import annotations

import json
import os
import subprocess
import sys
import time
import traceback
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "nemotron-asr.cpp_b200"))
sys.path.insert(0, str(ROOT / "oracle"))
sys.path.insert(0, str(ROOT / "tools"))
TMP = Path(os.environ.get("NB200_TMP", "/tmp/nb200"))
OUT = ROOT / "gpurun_out"


def ensure_inputs(layers=2):
    import make_pcm
    import make_random_gguf as mg
    TMP.mkdir(parents=True, exist_ok=True)
    paths = {}
    for variant in ("q8_0", "f32dq"):
        p = TMP / f"L{layers}_{variant}.gguf"
        if not p.exists():
            mg.write_gguf(str(p), 1234, layers, variant)
        paths[variant] = str(p)
    for seed, secs in ((11, 6.0),):
        p = TMP / f"pcm_s{seed}.pcm"
        if not p.exists():
            make_pcm.make_pcm(secs, seed).tofile(p)
        paths[f"pcm{seed}"] = str(p)
    return paths


def ref_harness_bin():
    flags = open("/proc/cpuinfo").read()
    for v in (("avx512",) if ("avx512vnni" in flags or "avx512_vnni" in flags) else ()) + ("avx2",):
        p = ROOT / "oracle" / "_ref" / v / "ref_harness"
        if p.exists():
            return str(p)
    return None


def rel(a, b):
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def phase_matmul(flags):
    import nb200
    import nemo_oracle as no
    paths = ensure_inputs()
    res = {}
    kv, t = no.load_gguf(paths["q8_0"])
    m = nb200.Model(paths["q8_0"], 0, 8, flags)
    rng = np.random.default_rng(0)
    for name, n_out in (("encoder.layers.0.feed_forward1.linear1.weight", 4096),
                        ("encoder.layers.0.feed_forward1.linear2.weight", 1024),
                        ("encoder.layers.1.self_attn.linear_out.weight", 1024)):
        w = t[name]
        for M in (1, 14, 130, 300):
            x = rng.standard_normal((M, w.shape[1])).astype(np.float32)
            y = m.debug_matmul(name, x, n_out)
            ref = no.mm(x, w)
            res[f"{name.split('.')[-3]}.{name.split('.')[-2]}:M{M}"] = rel(y, ref)
    m.close()
    return res


def phase_mel(flags):
    import nb200
    import nemo_oracle as no
    paths = ensure_inputs()
    kv, t = no.load_gguf(paths["q8_0"])
    m = nb200.Model(paths["q8_0"], 0, 8, flags)
    pcm = np.fromfile(paths["pcm11"], dtype=np.int16)[:16000 * 3]
    sig = np.stack([pcm, np.roll(pcm, 777)])
    got = m.mel_batch(sig)
    out = {}
    for i in range(2):
        pp = no.Preprocessor(t["preprocessor.featurizer.fb"], t["preprocessor.featurizer.window"])
        ref = pp.process(sig[i])
        d = np.abs(got[i] - ref)
        out[f"sig{i}"] = {"frames": int(ref.shape[0]), "max_abs": float(d.max()), "exact_frac": float((d == 0).mean())}
    m.close()
    return out


def run_stream(model, pcm, rc, taps=True):
    import nb200
    s = model.stream(rc)
    if taps:
        s.set_taps(nb200.TAP_MEL | nb200.TAP_ENC)
    feed = 8 * (1 + rc) * 160
    for i in range(0, len(pcm), feed):
        s.process_incremental(pcm[i:i + feed])
    s.finalize()
    return s


def phase_stream(flags, variant="f32dq", rcs=(13, 0), seconds=4.0):
    import nb200
    import nemo_oracle as no
    paths = ensure_inputs()
    pcm = np.fromfile(paths["pcm11"], dtype=np.int16)[:int(16000 * seconds)]
    pcm_path = TMP / "cut.pcm"
    pcm.tofile(pcm_path)
    out = {}
    m = nb200.Model(paths[variant], 0, 8, flags)
    harness = ref_harness_bin()
    for rc in rcs:
        s = run_stream(m, pcm, rc)
        T = 1 + rc
        enc = s.fetch_tap(nb200.TAP_ENC).reshape(-1, T, 1024)
        mel = s.fetch_tap(nb200.TAP_MEL).reshape(-1, 128)
        r = {"chunks": int(enc.shape[0]), "tokens": len(s.tokens), "chunk_tokens": s.chunk_token_counts[:12],
             "finite": bool(np.isfinite(enc).all())}
        if harness:
            pref = TMP / f"ref_{variant}_rc{rc}"
            subprocess.run([harness, "stream", paths[variant], str(pcm_path), str(rc), "8", str(pref)], check=True,
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE)
            renc = np.fromfile(str(pref) + ".enc.f32", dtype=np.float32).reshape(-1, T, 1024)
            rmel = np.fromfile(str(pref) + ".mel.f32", dtype=np.float32).reshape(-1, 128)
            rtok = np.fromfile(str(pref) + ".tokens.i32", dtype=np.int32)
            n = min(enc.shape[0], renc.shape[0])
            r["ref_chunks"] = int(renc.shape[0])
            r["enc_rel_vs_ref"] = [round(rel(enc[c], renc[c]), 6) for c in range(n)][:10]
            nm = min(mel.shape[0], rmel.shape[0])
            r["mel_frames"] = [int(mel.shape[0]), int(rmel.shape[0])]
            r["mel_max_abs_vs_ref"] = float(np.abs(mel[:nm] - rmel[:nm]).max()) if nm else None
            tk = np.array(s.tokens, dtype=np.int32)
            k = min(len(tk), len(rtok))
            r["ref_tokens"] = int(len(rtok))
            r["token_prefix_equal"] = int(np.argmax(np.concatenate([tk[:k] != rtok[:k], [True]])))
            r["ref_chunk_tokens"] = np.fromfile(str(pref) + ".chunk_tokens.i32", dtype=np.int32)[:12].tolist()
        out[f"rc{rc}"] = r
        s.free()
    st = m.stats()
    out["stats"] = {"steps": st.steps, "launches": st.kernel_launches, "last_step_ms": st.last_step_gpu_ms,
                    "frontend_ms": st.last_frontend_ms, "encoder_ms": st.last_encoder_ms, "decoder_ms": st.last_decoder_ms}
    m.close()
    return out


PHASES = {
    "mel": lambda: phase_mel(16),
    "matmul_simt": lambda: phase_matmul(16),
    "stream_simt_f32": lambda: phase_stream(16, "f32dq"),
    "stream_simt_q8": lambda: phase_stream(16, "q8_0"),
    "matmul_tc": lambda: phase_matmul(0),
    "stream_tc_f32": lambda: phase_stream(0, "f32dq"),
    "stream_tc_q8": lambda: phase_stream(0, "q8_0"),
}


def main():
    if len(sys.argv) >= 3 and sys.argv[1] == "phase":
        try:
            print("RESULT " + json.dumps(PHASES[sys.argv[2]]()))
        except Exception:
            traceback.print_exc()
            sys.exit(1)
        return
    names = sys.argv[2:] if len(sys.argv) > 2 else list(PHASES)
    OUT.mkdir(exist_ok=True)
    ensure_inputs()
    results = {}
    for name in names:
        t0 = time.time()
        try:
            p = subprocess.run([sys.executable, __file__, "phase", name], capture_output=True, text=True, timeout=240)
            line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")]
            results[name] = json.loads(line[-1][7:]) if line else {"error": (p.stdout + p.stderr)[-3000:]}
        except subprocess.TimeoutExpired:
            results[name] = {"error": "timeout"}
        results[name + "_secs"] = round(time.time() - t0, 1)
        print(name, json.dumps(results[name])[:1500], flush=True)
        (OUT / "gpu_check.json").write_text(json.dumps(results, indent=1))


if __name__ == "__main__":
    main()
