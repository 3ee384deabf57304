"""Batched serving loop: concurrent generate_audio_stream calls coalesced by StreamServer must
give every caller what a private call would have given it."""
import threading

import pytest
import torch

from oracle import pocket_oracle as po
from tests.helpers import CharTokenizer, weights

pytestmark = pytest.mark.gpu


def test_concurrent_requests_match_batch_api_and_fail_alone():
    from pocket_tts_b200.engine import ModelConfig
    from pocket_tts_b200.serving import StreamServer
    from pocket_tts_b200.tts_model import TTSModel

    m = TTSModel.from_state_dict(ModelConfig(n_layers=2), weights(2), CharTokenizer(), max_streams=8, kv_capacity=160,
                                 precision="fp32", eos_threshold=1e9)
    server = StreamServer([m], block=4)
    try:
        g = torch.Generator().manual_seed(21)
        voice_a = m.get_state_for_audio_prompt(torch.randn(10, 1024, generator=g) * 0.3)            # resident on the device
        voice_b = {k: {kk: vv.clone() for kk, vv in v.items()}
                   for k, v in m.get_state_for_audio_prompt(torch.randn(7, 1024, generator=g) * 0.3).items()}  # uploaded per use
        texts = ["ab", "hello", "xyz w", "q", "more text"]
        states = [voice_a, voice_b, voice_a, voice_b, voice_a]
        noise = [torch.randn(96, 32, generator=torch.Generator().manual_seed(100 + i)) * 0.7 ** 0.5 for i in range(5)]
        got = [None] * 5
        errs = []

        def work(i):
            try:
                got[i] = server.generate_audio(states[i], texts[i], noise=noise[i])
            except Exception as ex:   # pragma: no cover
                errs.append(ex)

        th = [threading.Thread(target=work, args=(i,)) for i in range(5)]
        for t in th:
            t.start()
        for t in th:
            t.join(timeout=120)
        assert not errs, errs
        want = m.generate_audio_batch(states, texts, noise=noise)
        for i in range(5):
            assert got[i].shape == want[i].shape and got[i].numel() > 0
            assert po.rel_l2(got[i], want[i]) <= 1e-5, (i, po.rel_l2(got[i], want[i]))
        # a request that cannot fit fails in its caller, the others keep running
        with pytest.raises(ValueError):
            server.generate_audio(voice_a, "x" * 48)      # needs more KV positions than kv_capacity=160
        again = server.generate_audio(voice_a, texts[0], noise=noise[0])
        assert po.rel_l2(again, want[0]) <= 1e-5
    finally:
        server.close()
        m.engine.close()
