"""Shared helpers for parity tests: build engines from the oracle's deterministic weights and
drive engine and oracle with identical inputs."""
from __future__ import annotations

from pathlib import Path

import numpy as np
import torch

from oracle import pocket_oracle as po

GOLD = Path(__file__).parent / "golden"
_WEIGHTS: dict = {}


def weights(n_layers: int):
    if n_layers not in _WEIGHTS:
        _WEIGHTS[n_layers] = po.make_weights(po.ModelDims(n_layers=n_layers), seed=0)
    return _WEIGHTS[n_layers]


def make_engine(n_layers: int, *, max_streams=8, kv_capacity=256, precision="fp32", lsd_steps=1,
                max_prefill_rows=512):
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    cfg = ModelConfig(n_layers=n_layers)
    return GenerationEngine(cfg, weights(n_layers), max_streams=max_streams, kv_capacity=kv_capacity,
                            precision=precision, lsd_steps=lsd_steps, max_prefill_rows=max_prefill_rows)


def load_case(name: str):
    g = np.load(GOLD / f"{name}.npz")
    ns = int(g["n_streams"])
    case = dict(n_layers=int(g["n_layers"]), lsd=int(g["lsd_steps"]), G=int(g["G"]), ns=ns)
    for k in ("tokens", "noise", "latents", "eos_logits", "pcm"):
        case[k] = [torch.from_numpy(g[f"{k}_{i}"]) for i in range(ns)]
    if "cond_seed" in g:   # long voice prefix redrawn from its seed (make_golden.py: first draw of the generator)
        import hashlib

        gen = torch.Generator().manual_seed(int(g["cond_seed"]))
        cond = torch.randn(1, int(g["cond_P"]), 1024, generator=gen) * 0.3
        digest = hashlib.sha256(cond[0].numpy().tobytes()).hexdigest()
        assert digest == str(g["cond_sha256"]), "voice conditioning redrawn from the seed differs from the fixture's"
        case["cond"] = [cond[0]]
    else:
        case["cond"] = [torch.from_numpy(g[f"cond_{i}"]) for i in range(ns)]
    # long fixtures: PCM after the first frames is stored as float16 (SNR gate only)
    case["pcm_tail"] = [torch.from_numpy(g[f"pcm_tail_f16_{i}"].astype(np.float32)) if f"pcm_tail_f16_{i}" in g else None
                        for i in range(ns)]
    case["raw"] = g
    return case


def synth_audio(seed: int, samples: int) -> torch.Tensor:
    """[1, samples] synthetic voice prompt (noise + two tones, amplitude ~0.1) of the encoder fixtures."""
    g = torch.Generator().manual_seed(seed)
    t = torch.arange(samples, dtype=torch.float32) / 24000.0
    x = 0.05 * torch.randn(samples, generator=g)
    x = x + 0.08 * torch.sin(2 * torch.pi * 220.0 * t) + 0.04 * torch.sin(2 * torch.pi * 1333.0 * t + 0.5)
    return x[None].contiguous()


def load_encoder_case(name: str):
    """tests/golden/encode_*.npz (make_golden_encoder.py): the waveform is redrawn from its seed and checked."""
    import hashlib

    g = np.load(GOLD / f"{name}.npz")
    audio = synth_audio(int(g["seed"]), int(g["samples"]))
    assert hashlib.sha256(audio.numpy().tobytes()).hexdigest() == str(g["audio_sha256"]), "synthetic waveform differs from the fixture's"
    rows = torch.from_numpy(g["rows"]).long() if "rows" in g else torch.arange(int(g["frames"]))
    return dict(audio=audio, frames=int(g["frames"]), rows=rows, cond=torch.from_numpy(g["cond"]))


def encoder_weights():
    if "enc" not in _WEIGHTS:
        _WEIGHTS["enc"] = po.make_encoder_weights(po.ModelDims(n_layers=2), seed=0)
    return _WEIGHTS["enc"]


def oracle_for(n_layers: int, lsd=1):
    return po.Oracle(po.ModelDims(n_layers=n_layers), weights(n_layers), lsd_steps=lsd)


class CharTokenizer:
    """Reversible toy tokenizer for the API-level fixtures (one id per character, ids < n_bins):
    stands in for the sentencepiece model, which is a network download in the reference."""

    def encode(self, text: str) -> list:
        return [ord(c) % 4000 for c in text]

    def decode(self, ids) -> str:
        return "".join(chr(int(i)) for i in ids)
