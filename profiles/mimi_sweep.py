"""cfg4 (BASELINE.json configs[3]): Mimi decoder-only throughput sweep over N = frames x streams.

Runs the Mimi passes of the frame loop alone (PTTS_PROFILE_PART=4: latents -> upsample -> 2-layer window-attention
transformer -> SEANet ConvTranspose1d / Conv1d / ELU stack -> PCM), device resident, 4 frames of every stream per
pass, and reports per N: ms per pass, audio-s/s, dense TFLOP/s (0.525 GFLOP per frame-stream) against
the measured bf16 tensor peak, and GB/s of algorithmic bytes (weights once + KV window + conv state + PCM).
    PTTS_PROFILE_PART=4 python profiles/mimi_sweep.py [streams ...]"""
import json
import os
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
os.environ.setdefault("PTTS_PROFILE_PART", "4")

import torch  # noqa: E402

from oracle import pocket_oracle as po  # noqa: E402  (weights generator only)
from pocket_tts_b200.engine import GenerationEngine, ModelConfig  # noqa: E402

G, PASSES = 4, 8
peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
TF_PEAK = float(peaks.get("bf16_tflops_sustained", 1368.9))


def run(n):
    dims = po.PRESETS["english"]
    W = po.make_weights(dims, seed=0)
    eng = GenerationEngine(ModelConfig(n_layers=dims.n_layers), W, max_streams=n, kv_capacity=G * (PASSES + 4) + 8,
                           precision="bf16", max_prefill_rows=max(n, 64))
    slots = [eng.open_stream() for _ in range(n)]
    g = torch.Generator().manual_seed(0)
    noise = (torch.randn(G * (PASSES + 3), n, dims.ldim, generator=g)).cuda()
    eng.run_frames(slots, noise[:3 * G].contiguous())            # warm-up: group sizes 1, 1, 2, 4, 4 and graph capture
    eng.synchronize()
    st = torch.cuda.ExternalStream(eng.cuda_stream)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    eng.run_frames(slots, noise[3 * G:].contiguous())
    eng.synchronize()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3
    frames = G * PASSES
    items = frames * n
    per_pass = ms / (frames / G)
    tflops = items * 0.525e9 / (ms * 1e-3) / 1e12
    byts = (frames / G) * 20.6e6 + items * (2 * 2 * 512 * 2 * (265 + 16) + 2 * 14720 * 2 + 32 * 4 + 1920 * 4)
    print(json.dumps({"streams": n, "frames_x_streams_per_pass": G * n, "ms_per_pass": round(per_pass, 3),
                      "audio_s_per_s": round(items * 0.08 / (ms * 1e-3), 1), "dense_tflops": round(tflops, 1),
                      "frac_of_bf16_peak": round(tflops / TF_PEAK, 4), "algorithmic_GBps": round(byts / (ms * 1e-3) / 1e9, 1)}), flush=True)
    eng.close()


if __name__ == "__main__":
    for n in [int(a) for a in sys.argv[1:]] or [1, 4, 16, 64, 256, 1024]:
        run(n)
