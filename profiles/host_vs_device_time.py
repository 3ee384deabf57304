"""Developer probe: host time spent inside ptts_run_frames (graph launches are asynchronous) against the
device time of the same frames — tells a launch-bound loop from a device-bound one.
    python profiles/host_vs_device_time.py [--frames 64]"""
import argparse
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=64)
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
    stream = torch.cuda.ExternalStream(eng.cuda_stream)
    eng.run_frames(slots, noise[:a.frames].contiguous())
    eng.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(stream)
    eng.run_frames(slots, noise[a.frames:].contiguous())
    e1.record(stream)
    t1 = time.perf_counter()
    eng.synchronize()
    t2 = time.perf_counter()
    print(f"host time in run_frames {1e6 * (t1 - t0) / a.frames:.1f} us/frame, until synchronized "
          f"{1e6 * (t2 - t0) / a.frames:.1f} us/frame, device {1e3 * e0.elapsed_time(e1) / a.frames:.1f} us/frame")


if __name__ == "__main__":
    main()
