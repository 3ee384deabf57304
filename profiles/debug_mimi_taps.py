"""Developer tool: locate the first Mimi-decode buffer that diverges from the CPU oracle.

    python profiles/debug_mimi_taps.py [--precision fp32] [--streams 2] [--frames 3]

Runs `mimi_decode` frame by frame on the engine (eager, no graph) and on the oracle with the same
latents and prints rel-L2 of every debug-readable intermediate (channels-last layouts, history rows
skipped).  The oracle is used only as the checker."""
import argparse
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--precision", default="fp32")
    ap.add_argument("--streams", type=int, default=2)
    ap.add_argument("--frames", type=int, default=3)
    a = ap.parse_args()
    import torch
    from oracle import pocket_oracle as po
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    dims = po.ModelDims(n_layers=1)
    W = po.make_weights(dims, seed=0)
    orc = po.Oracle(dims, W)
    B = a.streams
    eng = GenerationEngine(ModelConfig(n_layers=1), W, max_streams=B, kv_capacity=64, precision=a.precision)
    eng.set_graphs(False)
    slots = [eng.open_stream() for _ in range(B)]
    mst = orc.new_mimi_state(B, a.frames + 1)
    g = torch.Generator().manual_seed(5)
    U, Md = dims.upsample_stride, dims.mimi_d
    for f in range(a.frames):
        lat = torch.randn(B, dims.ldim, generator=g)
        pcm = eng.mimi_decode(slots, lat)
        ref, taps = orc.mimi_step(mst, lat, return_taps=True)
        print(f"frame {f}: pcm SNR {min(po.snr_db(ref[i], pcm[i]) for i in range(B)):.1f} dB")
        xm = eng.debug_read("xm", B * U * Md).view(B, U, Md)
        print(f"   transformer out (xm)   rel-L2 {po.rel_l2(xm, taps['transformer']):.3e}")
        T, C = U, Md
        tr0 = eng.debug_read("b_tr0", B * (1 + T) * C).view(B, 1 + T, C)[:, 1:]
        print(f"   elu(conv0) (b_tr0)     rel-L2 {po.rel_l2(tr0, po.elu(taps['conv0'])):.3e}")
        for i, r in enumerate(dims.ratios):
            T, C = T * r, C // 2
            raw = eng.debug_read(f"b_raw{i}", B * T * C).view(B, T, C)
            print(f"   convtr{i} raw (b_raw{i})  rel-L2 {po.rel_l2(raw, taps[f'tr{i}']):.3e}")
    eng.close()


if __name__ == "__main__":
    main()
