"""CPU arm: the reference's OWN code (the unmodified package vendored by oracle/make_ref.py into oracle/_ref/), run
as shipped — batch 1, `torch.set_num_threads(1)` (models/tts_model.py:49), generation thread + Mimi decoder thread
(tts_model.py:651-658,741-742) — through its public `TTSModel.generate_audio_stream`.

TEST / MEASUREMENT INFRASTRUCTURE ONLY: imported by bench.py's `--impl reference` / `cpu_baseline` legs, never by the
product.  Weights are the deterministic random-init set of `pocket_oracle.make_weights` (no network for checkpoints);
the sentencepiece download is replaced by a character tokenizer so that a 40-character text is 40 tokens; EOS is
disabled (threshold +inf) so every utterance runs its full max_gen_len, exactly like the GPU arm.

    python -m oracle.ref_runner --frames 24 --warmup 4 [--layers 6]      # prints one JSON line
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
REF_DIR = HERE / "_ref"


def reference_available() -> bool:
    return (REF_DIR / "pocket_tts" / "models" / "tts_model.py").exists()


class _CharTok:
    """One id per character (ids < n_bins); stands in for the sentencepiece model, which is a network download."""

    def __init__(self, nbins, tokenizer_path=None):
        self.nbins = nbins
        self.sp = self

    def decode(self, ids):
        return "".join(chr(int(i)) for i in ids)

    def encode(self, text):
        return [ord(c) % 4000 for c in text]

    def __call__(self, text):
        import torch
        from pocket_tts.conditioners.base import TokenizedText

        return TokenizedText(torch.tensor(self.encode(text))[None, :])


def build_reference_model(preset: str, n_layers: int, weights: dict, eos_threshold: float = float("inf")):
    """The reference's own constructor path (tts_model.py:130-230) on synthetic weights."""
    if str(REF_DIR) not in sys.path:
        sys.path.insert(0, str(REF_DIR))
    import torch
    import pocket_tts.conditioners.text as _text

    _text.SentencePieceTokenizer = _CharTok
    from pocket_tts.models.tts_model import TTSModel
    from pocket_tts.utils.config import CONFIGS_DIR, load_config

    cfg = load_config(CONFIGS_DIR / f"{preset}.yaml")
    cfg.weights_path = None
    cfg.weights_path_without_voice_cloning = None
    cfg.flow_lm.transformer.num_layers = n_layers
    torch.manual_seed(0)
    m = TTSModel._from_pydantic_config_with_weights(cfg, 0.7, 1, None, eos_threshold).eval()
    res = m.load_state_dict(dict(weights), strict=False)
    assert not res.unexpected_keys, res.unexpected_keys
    return m


def voice_state(ref, cond):
    """get_state_for_audio_prompt after the audio encoding (tts_model.py:893-899) on synthetic conditioning."""
    import torch
    from pocket_tts.modules.stateful_module import init_states

    with torch.no_grad():
        prompt = torch.cat([ref.flow_lm.bos_before_voice, cond], dim=1)
        st = init_states(ref.flow_lm, batch_size=1, sequence_length=prompt.shape[1])
        ref._run_flow_lm_and_increment_step(model_state=st, audio_conditioning=prompt)
    return st


def time_generate_stream(preset: str, n_layers: int, voice_p: int, text_t: int, warmup: int, frames: int, seed: int = 0,
                         handshake: bool = False) -> dict:
    """One utterance through the reference's public generate_audio_stream; the clock runs over `frames` yielded
    1920-sample chunks after `warmup` chunks (text prefill and thread start-up excluded, as on the GPU arm)."""
    import torch
    from oracle import pocket_oracle as po

    dims = po.PRESETS[preset] if hasattr(po, "PRESETS") and preset in po.PRESETS else po.ModelDims(n_layers=n_layers)
    W = po.make_weights(dims, seed=0)
    ref = build_reference_model(preset, dims.n_layers, W)
    g = torch.Generator().manual_seed(1 + seed)
    cond = torch.randn(1, voice_p - 1, dims.d_model, generator=g) * 0.1
    st = voice_state(ref, cond)
    text = ("a" * (text_t - 1)) + "."
    import logging

    logging.getLogger("pocket_tts").setLevel(logging.ERROR)
    torch.manual_seed(2 + seed)
    if handshake:           # multi-process run: everybody builds first, then the parent releases all at once
        print("READY", flush=True)
        sys.stdin.readline()
    n, t0, t1 = 0, None, None
    for chunk in ref.generate_audio_stream(st, text, max_tokens=max(50, text_t)):
        n += 1
        if n == warmup:
            t0 = time.perf_counter()
        if n == warmup + frames:
            t1 = time.perf_counter()
            break
    if t0 is None or t1 is None:
        raise RuntimeError(f"the utterance ended after {n} frames (< warmup + frames = {warmup + frames})")
    return {"frames": frames, "seconds": t1 - t0, "audio_s_per_s": frames * 0.08 / (t1 - t0),
            "torch_threads": torch.get_num_threads(), "t_start": time.time() - (time.perf_counter() - t0),
            "t_end": time.time() - (time.perf_counter() - t1)}


def run_processes(n_procs: int, preset: str, n_layers: int, voice_p: int, text_t: int, warmup: int, frames: int,
                  cores_per_proc: int = 2, timeout_s: float = 600.0) -> dict:
    """n_procs independent copies of the as-shipped reference (each: generation thread + Mimi thread), pinned to
    disjoint core pairs with taskset when it exists, released together.  Aggregate = total frames of audio over the
    span from the first start to the last end."""
    import shutil
    import subprocess

    root = HERE.parent
    procs = []
    ncpu = os.cpu_count() or 1
    for i in range(n_procs):
        cmd = [sys.executable, "-m", "oracle.ref_runner", "--preset", preset, "--layers", str(n_layers), "--voice", str(voice_p),
               "--text", str(text_t), "--warmup", str(warmup), "--frames", str(frames), "--seed", str(i), "--handshake"]
        if shutil.which("taskset") and n_procs > 1:
            c0 = (i * cores_per_proc) % ncpu
            cmd = ["taskset", "-c", ",".join(str((c0 + k) % ncpu) for k in range(cores_per_proc))] + cmd
        procs.append(subprocess.Popen(cmd, cwd=str(root), stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True))
    t_dead = time.time() + timeout_s
    try:
        for p in procs:
            line = p.stdout.readline()
            while line and line.strip() != "READY":
                if line.startswith("{") and "error" in line:
                    raise RuntimeError(line.strip())
                line = p.stdout.readline()
            if not line:
                raise RuntimeError("a reference process exited before it was ready")
            if time.time() > t_dead:
                raise RuntimeError("timeout while the reference processes were building their models")
        for p in procs:
            p.stdin.write("GO\n")
            p.stdin.flush()
        outs = []
        for p in procs:
            line = p.stdout.readline()
            while line and not line.startswith("{"):
                line = p.stdout.readline()
            outs.append(json.loads(line))
    finally:
        for p in procs:
            try:
                p.wait(timeout=5)
            except Exception:
                p.kill()
    span = max(o["t_end"] for o in outs) - min(o["t_start"] for o in outs)
    total_audio = sum(o["frames"] for o in outs) * 0.08
    return {"processes": n_procs, "aggregate_audio_s_per_s": total_audio / span, "span_s": span,
            "per_process_audio_s_per_s": [round(o["audio_s_per_s"], 3) for o in outs],
            "torch_threads_per_process": outs[0]["torch_threads"]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="english")
    ap.add_argument("--layers", type=int, default=6)
    ap.add_argument("--voice", type=int, default=126)
    ap.add_argument("--text", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--frames", type=int, default=24)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--handshake", action="store_true", help="print READY after the model is built, start on a line from stdin")
    a = ap.parse_args()
    if not reference_available():
        print(json.dumps({"error": "oracle/_ref/pocket_tts missing (run python oracle/make_ref.py in the build container)"}))
        return 2
    out = time_generate_stream(a.preset, a.layers, a.voice, a.text, a.warmup, a.frames, a.seed, a.handshake)
    print(json.dumps(out), flush=True)
    # the generator was abandoned mid-utterance: its worker threads are still running, leave without joining them
    os._exit(0)


if __name__ == "__main__":
    sys.exit(main())
