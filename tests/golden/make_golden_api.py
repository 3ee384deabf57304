"""API-level golden fixture from the REFERENCE itself (build container only): the reference's own
`TTSModel.generate_audio` (tts_model.py:477-543) — sentence splitting, per-chunk state copy, text
prefill, the autoregressive loop with its EOS rule and its own RNG draws, the Mimi decoder thread —
on the deterministic weights of `oracle.pocket_oracle.make_weights`, with a reversible character
tokenizer in place of the sentencepiece download.  The EOS threshold is chosen from a probe run so
that the first chunk stops by EOS a few frames in.

    python tests/golden/make_golden_api.py
"""
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference")

import pocket_tts.conditioners.text as _text  # noqa: E402
from pocket_tts.conditioners.base import TokenizedText  # noqa: E402

from tests.helpers import CharTokenizer  # noqa: E402


class _SP:
    def __init__(self, tok):
        self.tok = tok

    def decode(self, ids):
        return self.tok.decode(ids)


class _RefCharTokenizer:
    def __init__(self, nbins, tokenizer_path):
        self.tok = CharTokenizer()
        self.sp = _SP(self.tok)

    def __call__(self, text):
        return TokenizedText(torch.tensor(self.tok.encode(text))[None, :])


from pocket_tts.modules.stateful_module import init_states  # noqa: E402

from oracle import pocket_oracle as po  # noqa: E402
from tests.golden.make_golden import build_reference  # noqa: E402  (installs its own stub tokenizer)

_text.SentencePieceTokenizer = _RefCharTokenizer  # looked up when the conditioner is constructed

SEED = 1234
TEXT = "hello there. this is the b200 engine speaking, and it keeps going for a while! short one?"
MAX_TOKENS = 40
P = 12


@torch.no_grad()
def voice_state(ref, cond):
    """get_state_for_audio_prompt after the (host) audio encoding: tts_model.py:893-899."""
    prompt = torch.cat([ref.flow_lm.bos_before_voice, cond], dim=1)
    st = init_states(ref.flow_lm, batch_size=1, sequence_length=prompt.shape[1])
    ref._run_flow_lm_and_increment_step(model_state=st, audio_conditioning=prompt)
    return st


def run(ref, cond, thr, frames_after_eos, text=TEXT, copy_state=True):
    ref.eos_threshold = thr
    logits = []
    h = ref.flow_lm.out_eos.register_forward_hook(lambda m, i, o: logits.append(float(o.view(-1)[0])))
    torch.manual_seed(SEED)
    st = voice_state(ref, cond)
    chunks = [c.clone() for c in ref.generate_audio_stream(st, text, max_tokens=MAX_TOKENS,
                                                           frames_after_eos=frames_after_eos, copy_state=copy_state)]
    h.remove()
    return torch.cat(chunks), logits, st


def in_place_fixture(ref, cond):
    """copy_state=False with the EOS limit exactly at an engine block boundary (blocks of 1, 1, 2, 4 frames: EOS
    fires at frame 0, frames_after_eos = 4 -> 4 frames emitted, the reference still runs FlowLM for frame 4 and
    appends its K/V before breaking, tts_model.py:756-765): the state handed back holds voice + text + 5 positions."""
    text, fae = "hi there.", 4
    audio, logits, st = run(ref, cond, -1e9, fae, text=text, copy_state=False)
    off = int(st["transformer.layers.0.self_attn"]["offset"].view(-1)[0])
    tail = np.stack([st[f"transformer.layers.{i}.self_attn"]["cache"][:, 0, off - 3:off].numpy() for i in range(2)])
    after = np.stack([st[f"transformer.layers.{i}.self_attn"]["cache"][:, 0, off:off + 1].numpy() for i in range(2)]) \
        if st["transformer.layers.0.self_attn"]["cache"].shape[2] > off else np.zeros(0)
    print("in-place fixture: frames", audio.numel() // 1920, "offset", off, "flow calls", len(logits),
          "cache len", st["transformer.layers.0.self_attn"]["cache"].shape[2])
    np.savez_compressed(Path(__file__).parent / "api_inplace_2l.npz", cond=cond[0].numpy(), text=np.array(text),
                        frames_after_eos=fae, eos_threshold=-1e9, seed=SEED, audio=audio.numpy(), offset=off,
                        cache_tail=tail, next_draw=torch.randn(4).numpy(), torch_version=str(torch.__version__),
                        weights_seed=0)


TEXTS = [
    TEXT,
    "one",
    "  leading spaces and no punctuation  ",
    "a; b; c; d; e: f, g. h! i? j... k",
    "first sentence. second sentence is quite a bit longer than the first one, with a comma; and a semicolon: done. third!",
    "multi\nline\r\ntext  with   double spaces",
    "x" * 130 + ". " + "y" * 20 + ", " + "z" * 30 + "; tail",
    "ends with digit 7",
]


def text_fixture():
    """prepare_text_prompt / split_into_best_sentences of the reference on the character tokenizer."""
    import json

    from pocket_tts.models.tts_model import prepare_text_prompt, split_into_best_sentences

    tok = _RefCharTokenizer(4000, None)
    rows = []
    for text in TEXTS:
        for pad in (False, True):
            for semi in (False, True):
                for max_tokens in (20, 50):
                    prepared, guess = prepare_text_prompt(text, pad, semi)
                    chunks = split_into_best_sentences(tok, text, max_tokens, pad, semi)
                    rows.append(dict(text=text, pad=pad, remove_semicolons=semi, max_tokens=max_tokens,
                                     prepared=prepared, guess=guess, chunks=chunks))
    (Path(__file__).parent / "text_chunks.json").write_text(json.dumps(rows, indent=0))
    print("text fixture rows", len(rows))


if __name__ == "__main__":
    torch.set_num_threads(8)
    text_fixture()
    dims = po.ModelDims(n_layers=2)
    W = po.make_weights(dims, seed=0)
    ref = build_reference("english", 2, W, 1)
    g = torch.Generator().manual_seed(77)
    cond = torch.randn(1, P, dims.d_model, generator=g) * 0.3
    if "--inplace-only" in sys.argv:
        in_place_fixture(ref, cond)
        sys.exit(0)
    # probe: EOS never fires; logits[0] is the voice prefill, then per chunk: text prefill + frames
    _, logits, _ = run(ref, cond, 1e9, 2)
    first = np.array(logits[2:2 + 14])        # frames 0..13 of the first chunk
    k = max(i for i in range(1, len(first)) if first[i] > np.max(first[:i]))  # last running maximum: first EOS at frame k
    thr = float((first[k] + np.max(first[:k])) / 2)
    assert np.max(first[:k]) < thr < first[k], (first, k)
    audio, logits2, _ = run(ref, cond, thr, 2)
    print("threshold", thr, "first-chunk eos frame", k, "samples", audio.numel(), "frames", audio.numel() // 1920,
          "rms", float(audio.pow(2).mean().sqrt()))
    rng_after = torch.get_rng_state()
    np.savez_compressed(Path(__file__).parent / "api_generate_2l.npz", cond=cond[0].numpy(), text=np.array(TEXT),
                        max_tokens=MAX_TOKENS, frames_after_eos=2, eos_threshold=thr, seed=SEED,
                        audio=audio.numpy(), n_flow_calls=len(logits2), next_draw=torch.randn(4).numpy(),
                        torch_version=str(torch.__version__), weights_seed=0)
    in_place_fixture(ref, cond)
