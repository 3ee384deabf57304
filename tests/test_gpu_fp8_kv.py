"""FP8 (e4m3, per-row scale) FlowLM KV cache: the engine's low-precision mode (the analogue of
the reference's quantize=True, quantization.py:60-88).  Its own tolerance, stated here: teacher-forced latents within
rel-L2 <= 2e-2 of the reference's fp32 latents (the bf16 gate; measured 6e-3 .. 1.0e-2) and PCM SNR >= 25 dB (measured 40 dB), on the reference-generated
fixtures at the benched context lengths; exported caches within the e4m3 rounding of the bf16 engine's."""
import pytest
import torch

from oracle import pocket_oracle as po
from tests.helpers import load_case, make_engine, weights

pytestmark = pytest.mark.gpu

TOL_FP8_KV = 2e-2
SNR_FP8_KV = 25.0


def _engine(n_layers, **kw):
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    return GenerationEngine(ModelConfig(n_layers=n_layers), weights(n_layers), precision="bf16", kv_dtype="fp8", **kw)


@pytest.mark.parametrize("name,cap", [("gen_2l", 256), ("gen_english_6l", 256), ("gen_english_6l_long", 704)])
def test_fp8_kv_teacher_forced_latents_and_pcm_vs_reference(name, cap):
    case = load_case(name)
    orc_dims = po.ModelDims(n_layers=case["n_layers"])
    eng = _engine(case["n_layers"], max_streams=2, kv_capacity=cap, lsd_steps=case["lsd"], max_prefill_rows=512)
    try:
        bos = weights(case["n_layers"])["flow_lm.bos_before_voice"][0]
        worst, snr_min = 0.0, 1e9
        for i in range(case["ns"]):
            s = eng.open_stream()
            eng.prefill([s], [torch.cat([bos, case["cond"][i]])])
            eng.prefill_tokens([s], [case["tokens"][i].tolist()])
            pcm = []
            for f in range(case["G"]):
                if f > 0:
                    eng.set_prev_latent(s, case["latents"][i][f - 1])
                lat, _, p = eng.step([s], case["noise"][i][f][None])
                worst = max(worst, po.rel_l2(lat[0], case["latents"][i][f]))
                pcm.append(p[0])
            pcm = torch.cat(pcm)
            n_full = case["pcm"][i].numel()
            snr_min = min(snr_min, po.snr_db(case["pcm"][i], pcm[:n_full]))
            eng.close_stream(s)
        print(f"{name} fp8 KV: worst teacher-forced latent rel-L2 {worst:.3e}, min PCM SNR {snr_min:.1f} dB (dims {orc_dims.n_layers}L)")
        assert worst <= TOL_FP8_KV and snr_min >= SNR_FP8_KV
    finally:
        eng.close()


def test_fp8_kv_state_ops_round_trip():
    """import -> export is the e4m3 rounding of the input (relative error per element <= 2^-4 + scale granularity);
    fork, shared fork and export agree bit for bit; the frame loop runs on shared FP8 prefixes."""
    dims = po.ModelDims(n_layers=2)
    g = torch.Generator().manual_seed(4)
    eng = _engine(2, max_streams=5, kv_capacity=96, max_prefill_rows=256)
    try:
        voice = torch.randn(30, dims.d_model, generator=g) * 0.3
        a = eng.open_stream()
        eng.prefill([a], [voice])
        st = eng.export_state(a)
        b = eng.open_stream()
        eng.import_state(b, st)                         # values already on the e4m3 grid of their rows: the round trip
        st2 = eng.export_state(b)                       # only moves the row scale by an fp32 ulp
        for k in st:
            assert po.rel_l2(st2[k]["cache"], st[k]["cache"]) <= 1e-6
        ref = make_engine(2, max_streams=2, kv_capacity=96, precision="bf16")
        try:
            r = ref.open_stream()
            ref.prefill([r], [voice])
            stb = ref.export_state(r)
        finally:
            ref.close()
        key = "transformer.layers.0.self_attn"
        assert po.rel_l2(st[key]["cache"], stb[key]["cache"]) <= 4e-2     # e4m3: 3 mantissa bits
        c, d = eng.open_stream(), eng.open_stream()
        eng.fork(a, c)
        eng.fork_shared(a, d)
        toks = torch.randint(0, dims.n_bins, (7,), generator=g).tolist()
        noise = torch.randn(6, 1, dims.ldim, generator=g) * 0.7 ** 0.5
        outs = []
        for s in (c, d):
            eng.prefill_tokens([s], [toks])
            pcm, lat, _ = eng.generate_frames([s], noise)
            outs.append((pcm.clone(), lat.clone()))
        assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    finally:
        eng.close()
