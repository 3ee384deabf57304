"""Developer probe: CUDA-event time of every tagged kernel of ptts_run_frames (graphs off, so that the
events bracket single launches).  Run with PTTS_PROFILE_PART=3 (FlowLM + head only) or =4 (Mimi only) so that
the two streams do not overlap inside the brackets.
    PTTS_PROFILE_PART=4 python profiles/kernel_times.py [--frames 32]"""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=32)
    ap.add_argument("--streams", type=int, default=64)
    a = ap.parse_args()
    import torch
    from oracle import pocket_oracle as po
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    dims = po.ModelDims(n_layers=6)
    W = po.make_weights(dims, seed=0)
    B = a.streams
    eng = GenerationEngine(ModelConfig(n_layers=6), W, max_streams=B, kv_capacity=166 + 2 * a.frames + 16,
                           precision="bf16", max_prefill_rows=B * 126)
    g = torch.Generator().manual_seed(1)
    slots = [eng.open_stream() for _ in range(B)]
    emb = W["flow_lm.conditioner.embed.weight"]
    eng.prefill(slots, [torch.randn(126, dims.d_model, generator=g) * 0.1 for _ in slots])
    eng.prefill(slots, [emb[torch.randint(0, dims.n_bins, (40,), generator=g)] for _ in slots])
    noise = (torch.randn(2 * a.frames, B, dims.ldim, generator=g) * 0.7 ** 0.5).cuda()
    eng.set_graphs(False)
    eng.run_frames(slots, noise[:a.frames].contiguous())
    eng.synchronize()
    eng.profile_begin()
    eng.run_frames(slots, noise[a.frames:].contiguous())
    eng.synchronize()
    prof = eng.profile_end()
    tot = sum(v["total_ms"] for v in prof.values())
    print(f"{tot * 1e3 / a.frames:.1f} us/frame summed over {len(prof)} tags, {a.frames} frames")
    for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["total_ms"]):
        print(f"{k:28s} n={v['count']:5d}  avg {v['total_ms'] * 1e3 / v['count']:8.1f} us   {v['total_ms'] * 1e3 / a.frames:7.1f} us/frame")


if __name__ == "__main__":
    main()
