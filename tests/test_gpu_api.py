"""The drop-in Python API (pocket_tts_b200.TTSModel) on the GPU against (1) a fixture recorded from the
reference's own `generate_audio` and (2) the oracle, for the batched extension with ragged EOS stops."""
import numpy as np
import pytest
import torch

from oracle import pocket_oracle as po
from tests.helpers import GOLD, CharTokenizer, oracle_for, weights

pytestmark = pytest.mark.gpu


def make_model(**kw):
    from pocket_tts_b200 import ModelConfig, TTSModel

    args = dict(precision="fp32", max_streams=4, kv_capacity=512, max_prefill_rows=512)
    args.update(kw)
    return TTSModel.from_state_dict(ModelConfig(n_layers=2), weights(2), CharTokenizer(), **args)


def test_generate_audio_matches_reference_generate_audio():
    g = np.load(GOLD / "api_generate_2l.npz")
    m = make_model(eos_threshold=float(g["eos_threshold"]))
    try:
        torch.manual_seed(int(g["seed"]))
        state = m.get_state_for_audio_prompt(torch.from_numpy(g["cond"]))
        before = {k: v["cache"].clone() for k, v in state.items()}
        audio = m.generate_audio(state, str(g["text"]), max_tokens=int(g["max_tokens"]),
                                 frames_after_eos=int(g["frames_after_eos"]))
        nxt = torch.randn(4)
    finally:
        m.close()
    ref = torch.from_numpy(g["audio"])
    assert audio.shape == ref.shape, (audio.shape, ref.shape)       # same EOS decisions in every chunk
    assert po.snr_db(ref, audio) >= 40.0                             # north_star waveform gate (fp32 mode)
    assert torch.equal(nxt, torch.from_numpy(g["next_draw"]))        # RNG consumed exactly like the reference
    for k, v in state.items():                                       # copy_state=True: caller's state untouched
        assert torch.equal(v["cache"], before[k]) and int(v["offset"]) == g["cond"].shape[0] + 1


def test_stream_chunks_and_in_place_state():
    m = make_model(eos_threshold=1e9)  # EOS never fires: every chunk runs max_gen_len frames
    try:
        cond = torch.randn(9, 1024, generator=torch.Generator().manual_seed(3)) * 0.3
        state = m.get_state_for_audio_prompt(cond)
        assert set(state) == {f"transformer.layers.{i}.self_attn" for i in range(2)}
        assert state["transformer.layers.0.self_attn"]["cache"].shape == (2, 1, 10, 16, 64)
        torch.manual_seed(5)
        chunks = list(m.generate_audio_stream(state, "Hi.", copy_state=True))
        n = po.estimate_max_gen_len(3)
        assert len(chunks) == n and all(c.shape == (1920,) and c.dtype == torch.float32 for c in chunks)
        torch.manual_seed(5)
        again = m.generate_audio(state, "Hi.", copy_state=False)
        assert torch.equal(again, torch.cat(chunks))                 # bit-stable; same result either way
        off = int(state["transformer.layers.1.self_attn"]["offset"])
        assert off == 10 + 3 + n                                      # voice + text + every executed step
        assert state["transformer.layers.1.self_attn"]["cache"].shape[2] == off
    finally:
        m.close()


def test_in_place_state_with_eos_at_block_boundary_matches_reference():
    """copy_state=False when the EOS limit falls exactly on an engine block boundary (blocks of 1, 1, 2 frames, limit
    4): the reference runs FlowLM once more for frame 4 before breaking (tts_model.py:756-765), so the state handed
    back has voice + text + 5 positions, all of them written.  Fixture recorded from the reference itself."""
    g = np.load(GOLD / "api_inplace_2l.npz")
    m = make_model(eos_threshold=float(g["eos_threshold"]))
    try:
        torch.manual_seed(int(g["seed"]))
        state = m.get_state_for_audio_prompt(torch.from_numpy(g["cond"]))
        audio = m.generate_audio(state, str(g["text"]), frames_after_eos=int(g["frames_after_eos"]), copy_state=False)
        nxt = torch.randn(4)
    finally:
        m.close()
    ref = torch.from_numpy(g["audio"])
    off = int(g["offset"])
    assert audio.shape == ref.shape and po.snr_db(ref, audio) >= 40.0
    assert torch.equal(nxt, torch.from_numpy(g["next_draw"]))
    tail = torch.from_numpy(g["cache_tail"])                          # [layer, 2, 3, 16, 64]: the last three positions
    for i in range(2):
        st = state[f"transformer.layers.{i}.self_attn"]
        assert int(st["offset"].view(-1)[0]) == off
        got = st["cache"][:, 0, off - 3:off]
        assert po.rel_l2(got, tail[i]) <= 1e-4, (i, po.rel_l2(got, tail[i]))
        assert po.rel_l2(got[:, -1], tail[i][:, -1]) <= 1e-4         # the position of the extra FlowLM call


def test_generate_audio_batch_ragged_vs_oracle():
    dims = po.ModelDims(n_layers=2)
    orc = oracle_for(2)
    g = torch.Generator().manual_seed(21)
    voices = [torch.randn(P, dims.d_model, generator=g) * 0.3 for P in (7, 15)]
    reqs = [(0, 5), (1, 9), (0, 3), (1, 6), (0, 8)]                    # (voice, text tokens)
    toks = [torch.randint(0, dims.n_bins, (T,), generator=g) for _, T in reqs]
    G = [po.estimate_max_gen_len(T) for _, T in reqs]
    noise = [torch.randn(n, dims.ldim, generator=g) * 0.7 ** 0.5 for n in G]
    fae = 2
    # oracle, one stream at a time, EOS disabled -> logits; threshold between two sorted logits of the first 12 frames
    outs = []
    for (v, T), tk, nz in zip(reqs, toks, noise):
        st = orc.new_flow_state(1, 8)
        orc.prefill(st, orc.voice_conditioning(voices[v][None]))
        outs.append(orc.generate(st, orc.embed_tokens(tk)[None], nz[:16, None]))
    early = torch.sort(torch.cat([o["eos_logits"][:12, 0] for o in outs])).values
    thr = float((early[-4] + early[-5]) / 2)
    m = make_model(eos_threshold=thr)
    try:
        states = [m.get_state_for_audio_prompt(v) for v in voices]
        pcm = m.generate_audio_batch([states[v] for v, _ in reqs], toks, frames_after_eos=fae, noise=noise)
    finally:
        m.close()
    n_stopped = 0
    for i, o in enumerate(outs):
        hits = (o["eos_logits"][:, 0] > thr).nonzero()
        if hits.numel() and int(hits[0]) + fae <= 16:
            n_frames = int(hits[0]) + fae
            n_stopped += 1
            assert pcm[i].numel() == n_frames * 1920, (i, pcm[i].numel() // 1920, n_frames)
            assert po.snr_db(o["pcm"][0, : n_frames * 1920], pcm[i]) >= 40.0
        else:                                       # beyond the oracle's 16-frame horizon: compare the prefix
            assert pcm[i].numel() >= 16 * 1920
            assert po.snr_db(o["pcm"][0, : 16 * 1920], pcm[i][: 16 * 1920]) >= 40.0
    assert n_stopped >= 2


def test_api_error_conventions():
    from pocket_tts_b200 import TTSModel

    with pytest.raises(ValueError):
        TTSModel.load_model(language="english", config="english.yaml")
    m = make_model()
    try:
        state = m.get_state_for_audio_prompt(torch.zeros(4, 1024))
        with pytest.raises(ValueError):
            m.generate_audio(state, "   ")
        with pytest.raises(ValueError):
            m.get_state_for_audio_prompt(torch.zeros(1, 24000))      # raw audio without a host encoder
        with pytest.raises(ValueError):                               # predefined voice name: the reference's resolver is
            m.get_state_for_audio_prompt("alba")                      # absent on the GPU box / no language origin here
        with pytest.raises(ValueError):
            m.generate_audio_batch([state], ["a.", "b."])
        with pytest.raises(ValueError):                               # KV capacity: 5 + 400 + frames > 512
            m.generate_audio(state, torch.zeros(400, dtype=torch.long))
    finally:
        m.close()


@pytest.mark.gpu
def test_pcm_to_int16_matches_reference_wav_writer():
    """data/audio.py:77-78: (chunk.clamp(-1, 1) * 32767).short(), bit-exact, host and device buffers."""
    import torch
    from tests.helpers import make_engine

    eng = make_engine(1)
    g = torch.Generator().manual_seed(0)
    x = torch.randn(3, 1920, generator=g) * 0.8
    x[0, :4] = torch.tensor([1.0, -1.0, 1.5, -2.0])
    want = (x.clamp(-1, 1) * 32767).short()
    assert torch.equal(eng.pcm_to_int16(x), want)
    assert torch.equal(eng.pcm_to_int16(x.cuda()).cpu(), want)
    eng.close()
