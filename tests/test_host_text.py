"""Host-side text preparation against outputs of the reference's own functions
(tests/golden/text_chunks.json, written by tests/golden/make_golden_api.py)."""
import json

import pytest

from pocket_tts_b200.text import estimate_max_gen_len, prepare_text_prompt, split_into_best_sentences
from tests.helpers import GOLD, CharTokenizer

ROWS = json.loads((GOLD / "text_chunks.json").read_text())


def test_fixture_is_not_trivial():
    assert len(ROWS) >= 32 and any(len(r["chunks"]) > 2 for r in ROWS)


@pytest.mark.parametrize("i", range(len(ROWS)))
def test_prepare_and_split_match_reference(i):
    r = ROWS[i]
    prepared, guess = prepare_text_prompt(r["text"], r["pad"], r["remove_semicolons"])
    assert prepared == r["prepared"] and guess == r["guess"]
    chunks = split_into_best_sentences(CharTokenizer(), r["text"], r["max_tokens"], r["pad"], r["remove_semicolons"])
    assert chunks == r["chunks"]


def test_empty_text_raises():
    with pytest.raises(ValueError):
        prepare_text_prompt("   ")
    with pytest.raises(ValueError):
        split_into_best_sentences(CharTokenizer(), "\n", 50)


def test_max_gen_len_values():
    # tts_model.py:907-910: T=40 -> 192 frames, T=50 -> 234
    assert estimate_max_gen_len(40) == 192 and estimate_max_gen_len(50) == 234


def test_public_api_surface_matches_reference_pin():
    """reference tests/test_python_api.py:8-26: __all__ and the public methods / properties of TTSModel."""
    import pocket_tts_b200
    from pocket_tts_b200.tts_model import TTSModel

    assert pocket_tts_b200.__all__ == ["TTSModel", "export_model_state"]
    assert pocket_tts_b200.TTSModel is TTSModel
    for name in ("load_model", "generate_audio", "generate_audio_stream", "get_state_for_audio_prompt"):
        assert callable(getattr(TTSModel, name))
    for name in ("device", "sample_rate"):
        assert isinstance(getattr(TTSModel, name), property)
