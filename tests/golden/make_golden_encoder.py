"""Golden fixtures for the voice-prompt encoder, recorded from the REFERENCE itself.

Build container only (/root/reference importable).  The reference's own TTSModel is built by its own constructors,
loaded with `make_weights` + `make_encoder_weights` (every state_dict key present, strict), and its own
`_encode_audio` (models/tts_model.py:379-388 -> mimi.encode_to_latent, models/mimi.py:96-119) runs on a seeded
synthetic waveform.  The waveform is not stored: the loader redraws it from (seed, samples) and checks the digest.

    python tests/golden/make_golden_encoder.py
"""
import hashlib
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(Path(__file__).resolve().parent))

import make_golden as mg  # noqa: E402  (stubs the tokenizer, imports the reference)

from oracle import pocket_oracle as po  # noqa: E402


from tests.helpers import synth_audio  # noqa: E402


@torch.no_grad()
def case(name, seed, samples, keep_rows=None):
    dims = po.ModelDims(n_layers=2)
    W = dict(po.make_weights(dims, seed=0))
    W.update(po.make_encoder_weights(dims, seed=0))
    cfg = mg.load_config(mg.CONFIGS_DIR / "english.yaml")
    cfg.weights_path = None
    cfg.weights_path_without_voice_cloning = None
    cfg.flow_lm.transformer.num_layers = 2
    torch.manual_seed(0)
    m = mg.TTSModel._from_pydantic_config_with_weights(cfg, 0.7, 1, None, -4.0).eval()
    res = m.load_state_dict(W, strict=False)
    assert not res.unexpected_keys and not res.missing_keys, (res.unexpected_keys, res.missing_keys)
    audio = synth_audio(seed, samples)
    cond = m._encode_audio(audio.unsqueeze(0))[0]          # [frames, 1024]
    out = dict(seed=seed, samples=samples, audio_sha256=hashlib.sha256(audio.numpy().tobytes()).hexdigest(),
               frames=cond.shape[0], torch_version=str(torch.__version__), weights_seed=0)
    if keep_rows is None:
        out["cond"] = cond.numpy()
    else:
        rows = np.arange(0, cond.shape[0], keep_rows)
        out["rows"] = rows
        out["cond"] = cond[rows].numpy()
    np.savez_compressed(Path(__file__).resolve().parent / f"{name}.npz", **out)
    print(name, "frames", cond.shape[0], "rms", float(cond.pow(2).mean().sqrt()))


if __name__ == "__main__":
    case("encode_short", seed=11, samples=3 * 1920 + 777)          # ragged: right-padded to 4 frames
    case("encode_10s", seed=12, samples=240000, keep_rows=5)       # 125 frames = 2000 transformer positions (window 250 active)
