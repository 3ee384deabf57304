"""Probe: is the frame step latency-bound enough that several independent sub-batches (each with its own
engine, graphs and CUDA streams) overlap?  Runs G engines x (64 / G) streams from G host threads and reports the
aggregate ms per 64-stream frame.  Usage: python profiles/probe_concurrent_chains.py [G ...]"""
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402

from oracle import pocket_oracle as po  # noqa: E402  (weights generator only)
from pocket_tts_b200.engine import GenerationEngine, ModelConfig  # noqa: E402

B, K, WM, P, T = 64, 64, 4, 126, 40


def run(G):
    dims = po.PRESETS["english"]
    W = po.make_weights(dims, seed=0)
    emb, bos = W["flow_lm.conditioner.embed.weight"], W["flow_lm.bos_before_voice"][0]
    n = B // G
    engs, slots, noises, outs = [], [], [], []
    for gi in range(G):
        eng = GenerationEngine(ModelConfig(n_layers=dims.n_layers), W, max_streams=n, kv_capacity=P + T + WM + K + 8,
                               precision="bf16", max_prefill_rows=n * P)
        g = torch.Generator().manual_seed(gi)
        voice = torch.randn(n, P - 1, dims.d_model, generator=g) * 0.1
        tok = torch.randint(0, dims.n_bins, (n, T), generator=g)
        sl = [eng.open_stream() for _ in range(n)]
        eng.prefill(sl, [torch.cat([bos, voice[i]]) for i in range(n)])
        eng.prefill(sl, [emb[tok[i]] for i in range(n)])
        nz = (torch.randn(WM + K, n, dims.ldim, generator=g) * 0.7 ** 0.5).pin_memory()
        eng.generate_frames(sl, nz[:WM])
        engs.append(eng); slots.append(sl); noises.append(nz); outs.append(eng.alloc_frame_outputs(n, K))
    torch.cuda.synchronize()
    bar = threading.Barrier(G + 1)

    def work(i):
        bar.wait()
        engs[i].generate_frames(slots[i], noises[i][WM:], out=outs[i])
        bar.wait()

    th = [threading.Thread(target=work, args=(i,)) for i in range(G)]
    for t in th:
        t.start()
    bar.wait()
    t0 = time.perf_counter()
    bar.wait()
    dt = time.perf_counter() - t0
    for t in th:
        t.join()
    print(f"G={G}: {n} streams per engine, {dt / K * 1e3:.3f} ms per 64-stream frame, "
          f"{B * K * 0.08 / dt:.0f} audio-s/s", flush=True)
    for e in engs:
        e.close()


if __name__ == "__main__":
    for G in [int(a) for a in sys.argv[1:]] or [1, 2, 4]:
        run(G)
