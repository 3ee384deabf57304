"""Shared voice prefixes (ptts_stream_fork_shared): a stream forked WITHOUT a copy must
continue bit-identically to a deep-copied one, through prefill, single steps and the captured frame loop."""
import pytest
import torch

from oracle import pocket_oracle as po
from tests.helpers import make_engine, weights

pytestmark = pytest.mark.gpu


def _voice_and_text(seed, P=23, T=9, G=6):
    g = torch.Generator().manual_seed(seed)
    dims = po.ModelDims(n_layers=2)
    voice = torch.randn(P, dims.d_model, generator=g) * 0.3
    toks = torch.randint(0, dims.n_bins, (T,), generator=g).tolist()
    noise = torch.randn(G, 1, dims.ldim, generator=g) * 0.7 ** 0.5
    return voice, toks, noise


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_shared_fork_continues_bit_identically_to_deep_copy(precision):
    from pocket_tts_b200 import _native as N

    voice, toks, noise = _voice_and_text(5)
    eng = make_engine(2, max_streams=6, kv_capacity=96, precision=precision)
    try:
        v = eng.open_stream()
        eng.prefill([v], [voice])
        a, b, c = eng.open_stream(), eng.open_stream(), eng.open_stream()
        eng.fork(v, a)              # deep copy
        eng.fork_shared(v, b)       # shared prefix
        eng.fork_shared(b, c)       # sharing onward from a dependent that has not appended: shares the owner
        assert eng.offset(a) == eng.offset(b) == eng.offset(c) == voice.shape[0]
        outs = {}
        for name, s in (("copy", a), ("shared", b)):
            eng.prefill_tokens([s], [toks])
            lat, pcm = [], []
            for f in range(3):                                  # single steps
                l, _, p = eng.step([s], noise[f])
                lat.append(l); pcm.append(p)
            p2, l2, _ = eng.generate_frames([s], noise[3:])     # captured frame loop
            outs[name] = (torch.cat(lat), torch.cat(pcm, dim=1), l2.clone(), p2.clone())
        for x, y in zip(outs["copy"], outs["shared"]):
            assert torch.equal(x, y)
        # exported state reads through the sharing; a deep copy of a dependent continues like the dependent
        sa, sb = eng.export_state(a), eng.export_state(b)
        for k in sa:
            assert torch.equal(sa[k]["cache"], sb[k]["cache"]) and torch.equal(sa[k]["offset"], sb[k]["offset"])
        d = eng.open_stream()
        eng.fork(b, d)
        sd = eng.export_state(d)
        assert all(torch.equal(sd[k]["cache"], sb[k]["cache"]) for k in sb)
        # a mixed batch (copy + dependent + untouched dependent c after its own prefill) steps together
        eng.prefill_tokens([c], [toks])
        l_mixed, _, _ = eng.step([a, c], torch.cat([noise[0], noise[0]]))
        assert torch.isfinite(l_mixed).all()
        # ownership rules
        with pytest.raises(N.NativeError) as ei:
            eng.close_stream(v)
        assert ei.value.code == N.PTTS_ESTATE
        with pytest.raises(N.NativeError):
            eng.fork_shared(b, d)                               # b has appended: only a deep copy is possible
        with pytest.raises(N.NativeError):
            N.check(eng.lib.ptts_stream_set_offset(eng._h, c, 3))   # inside the shared prefix
        for s in (b, c):
            eng.close_stream(s)
        eng.close_stream(v)                                     # no dependents left
    finally:
        eng.close()


def test_batch_of_streams_sharing_one_voice_matches_private_copies_bf16():
    """16 streams of one voice, different texts: shared-prefix decode attention in the graph loop vs private copies."""
    dims = po.ModelDims(n_layers=2)
    g = torch.Generator().manual_seed(9)
    voice = torch.randn(40, dims.d_model, generator=g) * 0.3
    texts = [torch.randint(0, dims.n_bins, (5 + i % 7,), generator=g).tolist() for i in range(16)]
    noise = torch.randn(10, 16, dims.ldim, generator=g) * 0.7 ** 0.5
    res = {}
    for mode in ("copy", "shared"):
        eng = make_engine(2, max_streams=17, kv_capacity=96, precision="bf16")
        try:
            v = eng.open_stream()
            eng.prefill([v], [voice])
            slots = [eng.open_stream() for _ in range(16)]
            for s in slots:
                (eng.fork if mode == "copy" else eng.fork_shared)(v, s)
            eng.prefill_tokens(slots, texts)
            pcm, lat, eos = eng.generate_frames(slots, noise)
            res[mode] = (pcm.clone(), lat.clone(), eos.clone())
        finally:
            eng.close()
    for x, y in zip(res["copy"], res["shared"]):
        assert torch.equal(x, y)


def test_tts_model_keeps_voice_resident_and_forks_shared():
    """get_state_for_audio_prompt keeps the prefix on the device; generate_audio from it equals generation from a plain
    dict copy of the same state (the upload path), chunk for chunk, with the same seed."""
    from pocket_tts_b200.engine import ModelConfig
    from pocket_tts_b200.tts_model import TTSModel, VoiceState
    from tests.helpers import CharTokenizer

    W = weights(2)
    m = TTSModel.from_state_dict(ModelConfig(n_layers=2), W, CharTokenizer(), max_streams=8, kv_capacity=256, precision="fp32",
                                 eos_threshold=1e9)
    try:
        g = torch.Generator().manual_seed(3)
        cond = torch.randn(12, 1024, generator=g) * 0.3
        state = m.get_state_for_audio_prompt(cond)
        assert isinstance(state, VoiceState) and m._resident_slot(state) is not None
        plain = {k: {kk: vv.clone() for kk, vv in v.items()} for k, v in state.items()}   # no resident handle
        torch.manual_seed(11)
        a = m.generate_audio(state, "hi there")
        torch.manual_seed(11)
        b = m.generate_audio(plain, "hi there")
        assert a.numel() > 0 and torch.equal(a, b)
        out = m.generate_audio_batch([state, state, plain], ["abc", "defg", "abc"],
                                     noise=[torch.randn(64, 32, generator=torch.Generator().manual_seed(i % 2)) for i in (0, 1, 2)])
        assert torch.equal(out[0], out[2])
    finally:
        m.engine.close()
