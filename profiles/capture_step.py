"""Minimal driver for ncu captures: prefill B streams, then run a few frame steps.

    python profiles/capture_step.py [--streams 64 --layers 6 --frames 6 --precision bf16 --eager]

Prints the CUDA-event time per frame of the last `frames` steps (device-resident loop), so
the same command line can be run plain first and under ncu afterwards (B200_PROFILING.md).
The oracle module is imported only for its deterministic weight generator.
"""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--streams", type=int, default=64)
    ap.add_argument("--layers", type=int, default=6)
    ap.add_argument("--frames", type=int, default=6)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--voice", type=int, default=126)
    ap.add_argument("--text", type=int, default=40)
    ap.add_argument("--precision", default="bf16")
    ap.add_argument("--eager", action="store_true")
    ap.add_argument("--mimi-only", action="store_true")
    ap.add_argument("--kernel-profile", action="store_true", help="print per-kernel event times (eager)")
    a = ap.parse_args()

    import torch
    from oracle import pocket_oracle as po
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    dims = po.ModelDims(n_layers=a.layers)
    W = po.make_weights(dims, seed=0)
    B = a.streams
    eng = GenerationEngine(ModelConfig(n_layers=a.layers), W, max_streams=B,
                           kv_capacity=a.voice + a.text + a.warmup + a.frames + 8, precision=a.precision,
                           max_prefill_rows=min(B, 64) * a.voice)
    g = torch.Generator().manual_seed(1)
    slots = [eng.open_stream() for _ in range(B)]
    emb = W["flow_lm.conditioner.embed.weight"]
    for lo in range(0, B, 64):   # prefill in groups of 64 streams
        grp = slots[lo:lo + 64]
        eng.prefill(grp, [torch.randn(a.voice, dims.d_model, generator=g) * 0.1 for _ in grp])
        eng.prefill(grp, [emb[torch.randint(0, dims.n_bins, (a.text,), generator=g)] for _ in grp])
    noise = (torch.randn(a.warmup + a.frames, B, dims.ldim, generator=g) * 0.7 ** 0.5).cuda()
    if a.eager or a.kernel_profile:
        eng.set_graphs(False)
    stream = torch.cuda.ExternalStream(eng.cuda_stream)
    if a.mimi_only:
        lat = torch.randn(B, dims.ldim, generator=g)
        for _ in range(a.warmup):
            eng.mimi_decode(slots, lat)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(a.frames):
            eng.mimi_decode(slots, lat)
        e1.record(stream)
        eng.synchronize()
        print(f"mimi_decode (host latents/pcm): {e0.elapsed_time(e1) / a.frames * 1e3:.1f} us per frame-batch of {B}")
        return
    eng.run_frames(slots, noise[:a.warmup].contiguous())
    eng.synchronize()
    if a.kernel_profile:
        eng.profile_begin()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    eng.run_frames(slots, noise[a.warmup:].contiguous())
    e1.record(stream)
    eng.synchronize()
    us = e0.elapsed_time(e1) / a.frames * 1e3
    print(f"{a.layers}L {a.precision} B={B}: {us:.1f} us per frame step "
          f"({B * 0.08 / (us * 1e-6):.0f} audio-s/s), graphs={'off' if (a.eager or a.kernel_profile) else 'on'}")
    if a.kernel_profile:
        prof = eng.profile_end()
        tot = sum(v["total_ms"] for v in prof.values())
        for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["total_ms"]):
            print(f"  {k:24s} n={v['count']:4d}  avg {v['total_ms'] / v['count'] * 1e3:8.1f} us  share {v['total_ms'] / tot:.3f}")


if __name__ == "__main__":
    main()
