"""Host-side text preparation that feeds the engine: prompt normalisation, sentence chunking and the
token -> conditioning-row gather.  Behaviour follows the reference so that the same text yields the
same token chunks (pocket_tts/models/tts_model.py:913-1044, conditioners/text.py:53-76); none of it
is on the per-frame path."""
from __future__ import annotations

import logging
from typing import Protocol, Sequence

logger = logging.getLogger(__name__)

MAX_TOKEN_PER_CHUNK = 50          # default_parameters.py:8
TOKENS_PER_SECOND_ESTIMATE = 3.0  # tts_model.py:907-910
GEN_SECONDS_PADDING = 2.0


class Tokenizer(Protocol):
    """What the host needs from a tokenizer: ids for a text, text for ids
    (SentencePieceTokenizer.__call__ / .sp.decode, conditioners/text.py:13-36)."""

    def encode(self, text: str) -> list[int]: ...
    def decode(self, ids: Sequence[int]) -> str: ...


class SentencePieceAdapter:
    """Tokenizer over a sentencepiece model file (conditioners/text.py:26-36)."""

    def __init__(self, model_path: str, n_bins: int | None = None):
        import sentencepiece

        self.sp = sentencepiece.SentencePieceProcessor(str(model_path))
        if n_bins is not None and self.sp.vocab_size() != n_bins:
            raise ValueError(f"sentencepiece model has {self.sp.vocab_size()} pieces, config says {n_bins}")

    def encode(self, text: str) -> list[int]:
        return list(self.sp.encode(text, out_type=int))

    def decode(self, ids: Sequence[int]) -> str:
        return self.sp.decode(list(ids))


class ReferenceTokenizerAdapter:
    """Wraps the reference's own tokenizer object (callable -> TokenizedText, `.sp.decode`)."""

    def __init__(self, ref_tokenizer):
        self.ref = ref_tokenizer

    def encode(self, text: str) -> list[int]:
        return self.ref(text).tokens[0].tolist()

    def decode(self, ids: Sequence[int]) -> str:
        return self.ref.sp.decode(list(ids))


def estimate_max_gen_len(token_count: int, frame_rate: float = 12.5) -> int:
    """tts_model.py:907-910: ceil((tokens / 3 + 2 s) * frame_rate)."""
    import math

    return math.ceil((token_count / TOKENS_PER_SECOND_ESTIMATE + GEN_SECONDS_PADDING) * frame_rate)


def prepare_text_prompt(text: str, pad_with_spaces_for_short_inputs: bool = False,
                        remove_semicolons: bool = False) -> tuple[str, int]:
    """tts_model.py:913-946 -> (normalised text, frames_after_eos guess before the +2)."""
    text = text.strip()
    if not text:
        raise ValueError("Text prompt cannot be empty")
    for ch in ("\n", "\r"):
        text = text.replace(ch, " ")
    text = text.replace("  ", " ")
    if remove_semicolons:
        text = text.replace(";", ",")
    guess = 3 if len(text.split()) <= 4 else 1
    if not text[0].isupper():
        text = text[0].upper() + text[1:]
    if text[-1].isalnum():
        text += "."
    if pad_with_spaces_for_short_inputs and len(text.split()) < 5:
        text = " " * 8 + text
    return text, guess


def _segment_starts(tokens: Sequence[int], boundary: set) -> list[int]:
    """Cut positions: a new segment starts at the first non-boundary token after a run of boundary
    tokens (tts_model.py:949-968)."""
    cuts, in_run = [0], False
    for i, t in enumerate(tokens):
        if t in boundary:
            in_run = True
        elif in_run:
            cuts.append(i)
            in_run = False
    cuts.append(len(tokens))
    return cuts


def _segments(tok: Tokenizer, tokens: Sequence[int], boundary: set) -> list[tuple[int, str]]:
    cuts = _segment_starts(tokens, boundary)
    return [(b - a, tok.decode(tokens[a:b])) for a, b in zip(cuts[:-1], cuts[1:])]


def split_into_best_sentences(tok: Tokenizer, text: str, max_tokens: int = MAX_TOKEN_PER_CHUNK,
                              pad_with_spaces_for_short_inputs: bool = False,
                              remove_semicolons: bool = False) -> list[str]:
    """Greedy packing of sentences into chunks of at most `max_tokens` tokens; over-long sentences are
    first cut at , ; : (tts_model.py:978-1044)."""
    text, _ = prepare_text_prompt(text, pad_with_spaces_for_short_inputs, remove_semicolons)
    text = text.strip()
    ids = tok.encode(text)
    enders = set(tok.encode(".!...?")[1:])
    soft = set(tok.encode(",;:")[1:])
    pieces: list[tuple[int, str]] = []
    for count, sentence in _segments(tok, ids, enders):
        if count > max_tokens:
            sub = _segments(tok, tok.encode(sentence.strip()), soft)
            if len(sub) > 1:
                pieces.extend(sub)
                continue
        pieces.append((count, sentence))
    chunks: list[str] = []
    cur, cur_tokens = "", 0
    for count, sentence in pieces:
        if cur and cur_tokens + count > max_tokens:
            chunks.append(cur.strip())
            cur, cur_tokens = "", 0
        cur = sentence if not cur else cur + " " + sentence
        cur_tokens += count
    if cur:
        chunks.append(cur.strip())
    for c in chunks:
        n = len(tok.encode(c.strip()))
        if n > max_tokens:
            logger.warning("Chunk has %d tokens (max %d), generation may skip words: '%.50s...'", n, max_tokens, c)
    return chunks
