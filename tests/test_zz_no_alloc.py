"""ptts_pcm_to_int16_static — the allocation-free variant of the PCM egress conversion (its staging scratch is a
file-scope device array, csrc/static_scratch.cu) — against the existing ptts_pcm_to_int16, through the C ABI.  Both
compute (clamp(x, -1, 1) * 32767) truncated to int16 (data/audio.py:77-78), so the tolerance is the one the existing
test uses: bit-exact.  Collected last on purpose."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_static_scratch_pcm_to_int16_matches_allocating_entry_point_twice_in_a_row():
    from pocket_tts_b200 import _native as N
    from tests.helpers import make_engine

    eng = make_engine(1)
    try:
        g = torch.Generator().manual_seed(0)
        # call 1: four frames; call 2: FEWER samples of different data — the array still holds call 1's contents,
        # none of which may leak into call 2's result
        a = torch.randn(4, 1920, generator=g) * 0.8
        a[0, :4] = torch.tensor([1.0, -1.0, 1.5, -2.0])
        b = torch.randn(1, 1000, generator=g) * 0.5
        for x in (a, b):
            want = eng.pcm_to_int16(x)                                    # existing entry point (cudaMalloc'd staging)
            assert torch.equal(want, (x.clamp(-1, 1) * 32767).short())
            assert torch.equal(eng.pcm_to_int16_static(x), want)          # host in, host out
            assert torch.equal(eng.pcm_to_int16_static(x.cuda()).cpu(), want)   # device in, device out (no staging)
            out = torch.empty(x.shape, dtype=torch.int16).pin_memory()
            assert torch.equal(eng.pcm_to_int16_static(x.pin_memory(), out=out), want)   # pinned in, caller's pinned out
        # the largest accepted size, then one sample too many
        big = torch.rand(eng.PCM_STATIC_MAX_SAMPLES, generator=g) * 2.2 - 1.1
        assert torch.equal(eng.pcm_to_int16_static(big), eng.pcm_to_int16(big))
        with pytest.raises(N.NativeError) as ei:
            eng.pcm_to_int16_static(torch.zeros(eng.PCM_STATIC_MAX_SAMPLES + 1))
        assert ei.value.code == N.PTTS_EINVAL and "PTTS_PCM_STATIC_MAX_SAMPLES" in str(ei.value)
        assert eng.pcm_to_int16_static(torch.zeros(0)).numel() == 0
    finally:
        eng.close()
