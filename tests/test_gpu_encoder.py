"""Voice-prompt encoder on the device (ptts_encode_audio) against fixtures recorded from the
reference's own TTSModel._encode_audio (tests/golden/make_golden_encoder.py).  fp32 gate of the north_star: rel-L2 <= 1e-4."""
import pytest
import torch

from oracle import pocket_oracle as po
from tests.helpers import encoder_weights, load_encoder_case, weights

pytestmark = pytest.mark.gpu


def _engine(precision):
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    W = dict(weights(2))
    W.update(encoder_weights())
    return GenerationEngine(ModelConfig(n_layers=2), W, max_streams=2, kv_capacity=256, precision=precision,
                            max_prefill_rows=256), W


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["encode_short", "encode_10s"])
def test_encode_audio_matches_reference_encode_audio(name, precision):
    """Ragged right padding (short case), 2000 transformer positions with the 250-key window active (10 s case);
    the encoder runs in fp32 in both precision modes, so the fp32 gate applies to both."""
    c = load_encoder_case(name)
    eng, _ = _engine(precision)
    try:
        cond = eng.encode_audio(c["audio"])
        assert cond.shape == (c["frames"], 1024)
        err = po.rel_l2(cond[c["rows"]], c["cond"])
        print(f"{name} {precision}: {c['frames']} frames, rel-L2 {err:.2e}")
        assert err <= 1e-4
        # second call on the same engine reuses the scratch: bit-identical
        assert torch.equal(cond, eng.encode_audio(c["audio"]))
    finally:
        eng.close()


def test_encode_audio_rejects_bad_input_and_missing_weights():
    from pocket_tts_b200 import _native as N
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    eng, _ = _engine("fp32")
    try:
        with pytest.raises(N.NativeError):
            eng.encode_audio(torch.zeros(eng.ENCODE_MAX_SAMPLES + 1))
    finally:
        eng.close()
    eng = GenerationEngine(ModelConfig(n_layers=2), weights(2), max_streams=1, kv_capacity=64)   # no encoder weights
    try:
        assert not eng.has_encoder
        with pytest.raises(N.NativeError) as ei:
            eng.encode_audio(torch.zeros(1920))
        assert ei.value.code == N.PTTS_ESTATE
    finally:
        eng.close()


def test_voice_state_from_waveform_matches_oracle_prefill():
    """TTSModel.get_state_for_audio_prompt(waveform): device encoder -> bos_before_voice + conditioning -> prefill;
    the exported KV cache of layer 0 against the oracle's prefill of the fixture's conditioning."""
    from pocket_tts_b200.engine import ModelConfig
    from pocket_tts_b200.tts_model import TTSModel

    c = load_encoder_case("encode_short")
    W = dict(weights(2))
    W.update(encoder_weights())
    m = TTSModel.from_state_dict(ModelConfig(n_layers=2), W, None, max_streams=2, kv_capacity=128, precision="fp32")
    try:
        assert m.has_voice_cloning
        state = m.get_state_for_audio_prompt(c["audio"])
        key = "transformer.layers.0.self_attn"
        P = c["frames"] + 1
        assert int(state[key]["offset"][0]) == P
        orc = po.Oracle(po.ModelDims(n_layers=2), W)
        st = orc.new_flow_state(1, P)
        orc.prefill(st, orc.voice_conditioning(c["cond"][None]))
        got = state[key]["cache"][:, 0, :P]                     # [2, P, H, D]
        assert po.rel_l2(got[0], st.k[0][0, :P]) <= 1e-4 and po.rel_l2(got[1], st.v[0][0, :P]) <= 1e-4
    finally:
        m.engine.close()
