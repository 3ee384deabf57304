"""Per-kernel cost on the FlowLM chain inside the captured loop (PTTS_TIMELINE=1 globaltimer stamps): median time from
a kernel's programmatic-dependent-launch wait to the next kernel's wait, over the last frames of a 28-frame run at the
bench shape.  Run once per loop form, e.g.
    python profiles/chain_timeline.py                      # default loop (head-gated Mimi passes)
    PTTS_HEAD_GATE=0 python profiles/chain_timeline.py     # ungated
    PTTS_PROFILE_PART=3 python profiles/chain_timeline.py  # FlowLM chain alone"""
import os
import statistics
import sys
from collections import defaultdict
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
os.environ["PTTS_TIMELINE"] = "1"

import torch  # noqa: E402

from oracle import pocket_oracle as po  # noqa: E402  (weights generator only)
from pocket_tts_b200.engine import GenerationEngine, ModelConfig  # noqa: E402

B, P, T, G = 64, 126, 40, 28
dims = po.PRESETS["english"]
W = po.make_weights(dims, seed=0)
emb, bos = W["flow_lm.conditioner.embed.weight"], W["flow_lm.bos_before_voice"][0]
eng = GenerationEngine(ModelConfig(n_layers=dims.n_layers), W, max_streams=B, kv_capacity=P + T + G + 8, precision="bf16",
                       max_prefill_rows=B * P)
g = torch.Generator().manual_seed(3)
voice = torch.randn(B, P - 1, dims.d_model, generator=g) * 0.1
tok = torch.randint(0, dims.n_bins, (B, T), generator=g)
sl = [eng.open_stream() for _ in range(B)]
eng.prefill(sl, [torch.cat([bos, voice[i]]) for i in range(B)])
eng.prefill(sl, [emb[tok[i]] for i in range(B)])
noise = (torch.randn(G, B, dims.ldim, generator=g) * 0.7 ** 0.5).cuda()
eng.run_frames(sl, noise)
eng.synchronize()
recs = eng.timeline()
cost = defaultdict(list)
half = len(recs) // 2
for i in range(half, len(recs) - 1):
    a, b = recs[i], recs[i + 1]
    if a[2] > 0 and b[2] > a[2] and b[2] - a[2] < 500:
        cost[a[0]].append(b[2] - a[2])
tot = 0.0
print(f"{'kernel':28s} {'n':>4s} {'median us':>10s} {'p90 us':>8s}")
for name, v in sorted(cost.items(), key=lambda kv: -statistics.median(kv[1]) * len(kv[1])):
    v.sort()
    print(f"{name:28s} {len(v):4d} {statistics.median(v):10.2f} {v[int(0.9 * (len(v) - 1))]:8.2f}")
    tot += sum(v)
frames = max(1, len(cost.get("attn_decode", [])) // dims.n_layers)
print(f"sum over {frames} frames: {tot / frames:.1f} us per frame")
eng.close()
