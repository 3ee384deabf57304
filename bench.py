"""bench.py — throughput of the Pocket TTS per-frame generation path on B200.

Workload (BASELINE.json configs[1]): default english model (6-layer FlowLM), 64 concurrent
streams per GPU, voice prefix 126 positions + 40 text tokens, bf16 mode.  One step = one 80 ms
frame for every stream of the batch (FlowLM step + flow head + Mimi decode).
metric = audio-seconds generated per wall-second (RTFx), aggregate over all GPUs.

    python bench.py [--gpus N --steps K --warmup W] [--impl reference]

Under torchrun (N > 1) every rank drives its own GPU with its own 64 streams (streams shard,
weights replicate, no data-path collective): weak scaling.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

FRAME_S = 0.08
STREAMS = 64
VOICE_P, TEXT_T = 126, 40      # SURVEY.md 8(d) synthetic inputs
PRESET, N_LAYERS = "english", 6


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# --------------------------------------------------------------------------- clocks
class ClockSampler:
    """SM clock and clock-event reasons DURING a timed region.  NVML (the source nvidia-smi itself reads) polled every
    5 ms from a thread, because the default timed region is ~50 ms and an `nvidia-smi -lms` child takes longer than that
    to print its first row; `nvidia-smi -lms 100` is the fallback when the NVML binding is missing."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None
        self.nvml, self.handle, self.stop_flag, self.thread, self.max_mhz = None, None, False, None, None

    def _nvml_handle(self):
        import pynvml
        pynvml.nvmlInit()
        try:                                    # the PCI address survives any CUDA_VISIBLE_DEVICES remapping
            import torch
            pr = torch.cuda.get_device_properties(self.index)
            bus = f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
            return pynvml, pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        except Exception:
            return pynvml, pynvml.nvmlDeviceGetHandleByIndex(self.index)

    def start(self):
        try:
            self.nvml, self.handle = self._nvml_handle()
            nv = self.nvml
            self.max_mhz = float(nv.nvmlDeviceGetMaxClockInfo(self.handle, nv.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nvml = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv, h = self.nvml, self.handle
        bits = [(getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8), "hw_slowdown"),
                (getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40), "hw_thermal_slowdown"),
                (getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20), "sw_thermal_slowdown"),
                (getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4), "sw_power_cap")]
        while not self.stop_flag:
            try:
                mhz = float(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                mask = int(nv.nvmlDeviceGetCurrentClocksEventReasons(h))
                self.rows.append((mhz, self.max_mhz, [nm for b, nm in bits if mask & b]))
            except Exception:
                pass
            time.sleep(0.005)

    def _read(self):
        for line in self.proc.stdout:
            r = [x.strip() for x in line.split(",")]
            try:
                self.rows.append((float(r[0]), float(r[1]),
                                  [nm for nm, v in zip(self.NAMES, r[3:7]) if v.lower().startswith("active")]))
            except Exception:
                pass

    def stop(self) -> dict:
        if self.nvml is not None:
            self.stop_flag = True
            self.thread.join(timeout=1.0)
            source = "nvml, 5 ms period"
        elif self.proc is not None:
            time.sleep(0.15)
            self.proc.terminate()
            source = "nvidia-smi -lms 100"
        else:
            return {"sm_mhz": None, "sm_max_mhz": None, "samples": 0, "reasons": ["nvidia-smi unavailable"]}
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = sorted({nm for r in self.rows for nm in r[2]})
        busy = [s for s in sm if s > 0.5 * max(sm)] if sm else []
        return {"sm_mhz": statistics.median(busy) if busy else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": reasons, "source": source}


# --------------------------------------------------------------------------- workload
def synth_inputs(dims, B, G, seed_base=0):
    import torch
    g = torch.Generator().manual_seed(1 + seed_base)
    tokens = torch.randint(0, dims.n_bins, (B, TEXT_T), generator=g)
    voice = torch.randn(B, VOICE_P - 1, dims.d_model, generator=g) * 0.1
    g2 = torch.Generator().manual_seed(2 + seed_base)
    noise = torch.randn(G, B, dims.ldim, generator=g2) * (0.7 ** 0.5)
    return tokens, voice, noise


def algorithmic_bytes_per_step(dims, B, ctx_len, e, mimi_tokens, kv_e=None):
    """weights once + FlowLM KV read + Mimi KV window + conv state + I/O.  `mimi_tokens` is the
    Mimi decoder's own offset (16 tokens per generated frame), not the FlowLM context."""
    w_elems = 95_582_018 if dims.n_layers == 6 else 322_074_434
    kv_flow = B * dims.n_layers * 2 * dims.d_model * (kv_e if kv_e is not None else e) * (ctx_len + 1)
    kv_mimi = B * dims.mimi_layers * 2 * dims.mimi_d * e * (min(mimi_tokens, 265) + 16)
    state = B * 2 * 14_720 * e
    io = B * (dims.ldim * 4 + dims.frame_samples * 4)
    return w_elems * e + kv_flow + kv_mimi + state + io


def run_ours(args, rank, world, local_rank):
    import torch
    from oracle import pocket_oracle as po           # weights generator + CPU baseline only
    from pocket_tts_b200.engine import GenerationEngine, ModelConfig

    torch.cuda.set_device(local_rank)
    dims = po.PRESETS[PRESET]
    B, K, Wm = STREAMS, args.steps, args.warmup
    prec = args.precision
    e_bytes = 2 if prec == "bf16" else 4
    W = po.make_weights(dims, seed=0)
    kv_dtype = None if args.kv == "act" else args.kv
    kv_e = (1 + 4 / 64) if kv_dtype == "fp8" else e_bytes      # e4m3 row + its fp32 scale
    eng = GenerationEngine(ModelConfig(n_layers=dims.n_layers), W, max_streams=B,
                           kv_capacity=VOICE_P + TEXT_T + Wm + K + 8, precision=prec, device=local_rank,
                           max_prefill_rows=B * VOICE_P, kv_dtype=kv_dtype)
    emb = W["flow_lm.conditioner.embed.weight"]
    bos = W["flow_lm.bos_before_voice"][0]

    def fresh_streams(seed_base):
        tokens, voice, noise = synth_inputs(dims, B, Wm + K, seed_base + 1000 * rank)
        slots = [eng.open_stream() for _ in range(B)]
        eng.prefill(slots, [torch.cat([bos, voice[i]]) for i in range(B)])
        eng.prefill(slots, [emb[tokens[i]] for i in range(B)])
        return slots, noise

    def close(slots):
        for s in slots:
            eng.close_stream(s)

    dist = world > 1
    if dist:
        import torch.distributed as td
        td.init_process_group("gloo")   # control plane only (barrier + two scalar reductions): no NCCL on any path

    def barrier():
        if dist:
            td.barrier()
        torch.cuda.synchronize()

    # one-off set-up outside every timed region: a throw-away batch runs 12 frames so that every kernel module is
    # loaded and every graph variant of the loop (group sizes 1, 1, 2, 4, 4) is captured before anything is timed
    # (on a fresh box the first 4-frame Mimi pass otherwise pays lazy module loading inside the timed region)
    slots, noise = fresh_streams(99)
    eng.run_frames(slots, noise[:min(12, Wm + K)].cuda().contiguous())
    eng.synchronize()
    close(slots)

    # ---------------- device-resident timed region (value) ----------------
    slots, noise = fresh_streams(0)
    noise_dev = noise.cuda()
    stream = torch.cuda.ExternalStream(eng.cuda_stream)
    pcm_dev = torch.empty(B, K * dims.frame_samples, device="cuda")
    eng.run_frames(slots, noise_dev[:Wm].contiguous())          # warm-up (captures the graph)
    eng.synchronize()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    launches0 = eng.launch_count
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    eng.run_frames(slots, noise_dev[Wm:].contiguous(), out_pcm=pcm_dev)
    ev1.record(stream)
    eng.synchronize()
    barrier()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop()
    launches = eng.launch_count - launches0
    assert bool(torch.isfinite(pcm_dev).all()), "non-finite PCM"
    close(slots)

    # ---------------- end-to-end through the host-buffer C-ABI call (e2e) ----------------
    # ptts_generate_frames: per frame the noise comes from pinned host memory and latents, EOS logits
    # and PCM go back to pinned host memory, all inside the timed region; FlowLM(f) || Mimi(f-1).
    slots, noise = fresh_streams(1)
    noise_pin = noise.pin_memory()
    eng.generate_frames(slots, noise_pin[:Wm])
    out_bufs = eng.alloc_frame_outputs(B, K)                      # pinned result buffers, allocated once by the caller
    barrier()
    t0 = time.perf_counter()
    pcm_h, lat_h, eos_h = eng.generate_frames(slots, noise_pin[Wm:Wm + K], out=out_bufs)
    e2e_ms = (time.perf_counter() - t0) * 1e3
    assert bool(torch.isfinite(pcm_h).all()) and bool(torch.isfinite(lat_h).all())
    close(slots)
    # secondary: one blocking ptts_step per frame (host round trip every 80 ms frame, no overlap)
    slots, noise = fresh_streams(3)
    noise_pin = noise.pin_memory()
    for g in range(Wm):
        eng.step(slots, noise_pin[g])
    t0 = time.perf_counter()
    KS = min(K, 32)
    for g in range(Wm, Wm + KS):
        eng.step(slots, noise_pin[g])
    step_api_ms = (time.perf_counter() - t0) * 1e3 / KS
    close(slots)
    h2d = B * dims.ldim * 4
    d2h = B * (dims.ldim * 4 + 4 + dims.frame_samples * 4)

    # ---------------- first-chunk latency: text prefill + first FlowLM step + first Mimi frame, PCM on the host ------
    # (the voice prefix is already in the slot, as with the reference's built-in voice states; tts_model.py:718-779)
    first_chunk = {}
    if rank == 0:
        for label, nb in (("batch", B), ("single", 1)):
            times = []
            for trial in range(7):
                tokens, voice, noise = synth_inputs(dims, nb, 1, 50 + trial)
                sl = [eng.open_stream() for _ in range(nb)]
                eng.prefill(sl, [torch.cat([bos, voice[i]]) for i in range(nb)])
                text = [tokens[i].tolist() for i in range(nb)]
                nz = noise[0].pin_memory()
                eng.synchronize()
                t0 = time.perf_counter()
                eng.prefill_tokens(sl, text)                       # token ids in, embedding gather on the device
                eng.step(sl, nz)                                   # returns latents, EOS logits and PCM on the host
                times.append((time.perf_counter() - t0) * 1e3)
                close(sl)
            first_chunk[f"p50_ms_{label}_{nb}"] = round(statistics.median(times[2:]), 3)

    # the metric as SURVEY.md 8(d) defines it: from the generate_audio_stream CALL (state fork + prefill + first FlowLM
    # step + first Mimi frame) to the first [1920] chunk on the host, through the drop-in TTSModel; with the voice
    # prefix resident on the device (shared fork) and with a plain model_state dict (K/V upload, tts_model.py:637-638)
    if rank == 0:
        try:
            first_chunk.update(api_first_chunk(dims, W, prec, local_rank))
        except Exception as ex:   # secondary numbers must never break the bench line
            first_chunk["api_error"] = str(ex)[:200]

    # ---------------- per-kernel CUDA-event pass (roofline of the dominant kernel) ----------------
    prof = {}
    if rank == 0:
        slots, noise = fresh_streams(2)
        eng.set_graphs(False)
        for g in range(Wm):
            eng.step(slots, noise[g])
        eng.profile_begin()
        PK = min(K, 32)
        for g in range(Wm, Wm + PK):
            eng.step(slots, noise[g])
        prof = eng.profile_end()
        eng.set_graphs(True)
        close(slots)

    # ---------------- decode attention inside the captured frame loop (device timers, no profiler) ----------------
    # A second engine with the in-graph timeline on runs 24 frames of the real loop (FlowLM graph || Mimi graph);
    # the kernel's cost = time from the return of its programmatic-dependent-launch wait to the return of the next
    # kernel's wait (out_proj), i.e. until its whole grid has finished and flushed.  Secondary to `roofline` (CUDA
    # events around single eager launches), which cannot see the overlap the graph has.
    in_graph = None
    if rank == 0:
        try:
            os.environ["PTTS_TIMELINE"] = "1"
            eng2 = GenerationEngine(ModelConfig(n_layers=dims.n_layers), W, max_streams=B,
                                    kv_capacity=VOICE_P + TEXT_T + 40, precision=prec, device=local_rank,
                                    max_prefill_rows=B * VOICE_P, kv_dtype=kv_dtype)
            os.environ.pop("PTTS_TIMELINE", None)
            tokens, voice, noise2 = synth_inputs(dims, B, 24, 7)
            sl2 = [eng2.open_stream() for _ in range(B)]
            eng2.prefill(sl2, [torch.cat([bos, voice[i]]) for i in range(B)])
            eng2.prefill(sl2, [emb[tokens[i]] for i in range(B)])
            eng2.run_frames(sl2, noise2.cuda().contiguous())
            eng2.synchronize()
            recs = eng2.timeline()
            costs = []
            for i, r in enumerate(recs):
                if r[0] == "attn_decode" and r[2] > 0:
                    nxt = next((q for q in recs[i + 1:i + 40] if q[0] == "gemm N=1024 kb=16" and q[2] > r[2]), None)
                    if nxt:
                        costs.append(nxt[2] - r[2])
            if costs:
                costs = costs[len(costs) // 2:]                      # frames 12..23: context 178-189
                us = statistics.median(costs)
                per_launch = kernel_algorithmic_bytes("flow.attn_decode", dims, B, VOICE_P + TEXT_T + 18, e_bytes, kv_e)
                in_graph = {"kernel": "flow.attn_decode", "median_us_wait_to_next_wait": round(us, 2), "launches": len(costs),
                            "achieved_GBps": round(per_launch / (us * 1e-6) / 1e9, 1),
                            "method": "globaltimer stamps inside the CUDA-graph loop (PTTS_TIMELINE), FlowLM || Mimi running"}
            eng2.close()
        except Exception as ex:  # the secondary number must never break the bench line
            in_graph = {"error": str(ex)[:200]}
            os.environ.pop("PTTS_TIMELINE", None)

    if dist:   # max over ranks of the device-timed region and of the end-to-end region
        from pocket_tts_b200.sharding import aggregate_throughput
        _, ms = aggregate_throughput(B * K * FRAME_S, ms)
        _, e2e_ms = aggregate_throughput(B * K * FRAME_S, e2e_ms)
    if rank != 0:
        if dist:
            td.destroy_process_group()
        return

    audio_s = world * B * K * FRAME_S
    value = audio_s / (ms / 1e3)
    e2e_value = audio_s / (e2e_ms / 1e3)
    # roofline: dominant kernel by summed event time
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    total_prof = sum(v["total_ms"] for v in prof.values()) or 1.0
    # dominant kernel = the tag with the largest summed event time among those with a byte model (kernel_ms_share)
    ranked = sorted(prof.items(), key=lambda kv: -kv[1]["total_ms"])
    top = next(((k, v) for k, v in ranked if kernel_algorithmic_bytes(k, dims, B, 200, e_bytes, kv_e) > 0), (None, None))
    roofline = None
    if top[0]:
        ctx_mid = VOICE_P + TEXT_T + Wm + min(K, 32) // 2
        per_launch = kernel_algorithmic_bytes(top[0], dims, B, ctx_mid, e_bytes, kv_e)
        avg_ms = top[1]["total_ms"] / top[1]["count"]
        ach = per_launch / (avg_ms / 1e3) / 1e9
        roofline = {"kernel": top[0], "bound": "hbm", "achieved": round(ach, 1), "peak": hbm_peak,
                    "unit": "GB/s", "frac": round(ach / hbm_peak, 4), "traffic": ncu_traffic_bytes(top[0], args.workload),
                    "algorithmic_bytes_per_launch": int(per_launch),
                    "avg_launch_us": round(avg_ms * 1e3, 2), "share_of_step": round(top[1]["total_ms"] / total_prof, 3),
                    "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback"}
    if in_graph and "achieved_GBps" in in_graph:
        in_graph["frac"] = round(in_graph["achieved_GBps"] / hbm_peak, 4)
    ctx_avg = VOICE_P + TEXT_T + Wm + K / 2
    step_bytes = algorithmic_bytes_per_step(dims, B, ctx_avg, e_bytes, 16 * (Wm + K / 2), kv_e)
    step_frac = step_bytes / (ms / K / 1e3) / 1e9 / hbm_peak

    cpu = cpu_baseline(dims, W, sample_streams=8, frames=4)
    line = {
        "metric": "audio-sec generated per sec (RTFx), batched", "value": round(value, 1), "unit": "audio-s/s",
        "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": round(ms / K, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": prec, "data": "synthetic",
        "config": {"workload": f"{'default english' if PRESET == 'english' else PRESET} model ({dims.n_layers}L FlowLM), {B} concurrent streams per GPU, "
                               f"voice prefix {VOICE_P} + {TEXT_T} text tokens, lsd_steps 1, frame step = FlowLM + flow head + Mimi decode",
                   "streams_per_gpu": B, "precision": prec, "kv_cache": kv_dtype or prec, "parallelism": f"streams sharded x{world}, no collective",
                   "l2_policy": "working set per step (weights + KV, >= 0.6 GB) exceeds the 126 MB L2; no flush"},
        "e2e": {"value": round(e2e_value, 1), "unit": "audio-s/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": round(e2e_ms / K, 4),
                "api": "ptts_generate_frames (pinned host noise in; latents, EOS logits, PCM out per frame) via GenerationEngine.generate_frames",
                "blocking_step_api_ms_per_step": round(step_api_ms, 4)},
        "first_chunk": first_chunk,
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "roofline_in_graph": in_graph,
        "step_roofline": {"bound": "hbm", "algorithmic_bytes_per_step": int(step_bytes), "frac_of_hbm_peak": round(step_frac, 4)},
        "cpu_baseline": cpu,
        "kernel_ms_share": {k: round(v["total_ms"] / total_prof, 4) for k, v in sorted(prof.items(), key=lambda kv: -kv[1]["total_ms"])[:12]},
    }
    print(json.dumps(line), flush=True)
    if dist:
        td.destroy_process_group()


def api_first_chunk(dims, W, prec, device):
    import torch
    from pocket_tts_b200.engine import ModelConfig
    from pocket_tts_b200.tts_model import TTSModel

    m = TTSModel.from_state_dict(ModelConfig(n_layers=dims.n_layers), W, None, precision=prec, max_streams=8,
                                 kv_capacity=VOICE_P + TEXT_T + 200, max_prefill_rows=VOICE_P + 8, device=device, eos_threshold=1e9)
    try:
        g = torch.Generator().manual_seed(77)
        voice = m.get_state_for_audio_prompt(torch.randn(VOICE_P - 1, dims.d_model, generator=g) * 0.1)
        plain = {k: {kk: vv.clone() for kk, vv in v.items()} for k, v in voice.items()}
        out = {}
        for label, st in (("api_resident_voice", voice), ("api_uploaded_state", plain)):
            ts = []
            for trial in range(9):
                toks = torch.randint(0, dims.n_bins, (TEXT_T,), generator=g).tolist()
                it = m.generate_audio_stream(st, toks)
                t0 = time.perf_counter()
                next(it)
                ts.append((time.perf_counter() - t0) * 1e3)
                it.close()
            out[f"p50_ms_{label}"] = round(statistics.median(ts[2:]), 3)
        return out
    finally:
        m.engine.close()


def ncu_traffic_bytes(tag, workload):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed `ncu --set full` summary of THIS
    workload and kernel (profiles/), taken at the start of the generation; None when no such capture exists."""
    import csv
    f = {("cfg2", "flow.attn_decode"): ROOT / "profiles" / "r02_attn_decode_ncu_full_summary.csv"}.get((workload, tag))
    if not f or not f.exists():
        return None
    rows = list(csv.reader(open(f)))
    hdr, units = rows[0], rows[1]
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    ri, wi = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    vals = [float(r[ri]) * mult[units[ri]] + float(r[wi]) * mult[units[wi]] for r in rows[2:] if len(r) > max(ri, wi)]
    return int(sum(vals) / len(vals)) if vals else None


def kernel_algorithmic_bytes(tag, dims, B, ctx, e, kv_e=None):
    """Algorithmic bytes one launch of `tag` must move (DESIGN.md kernel table)."""
    d, ffn = dims.d_model, dims.ffn
    if tag == "flow.attn_decode":
        return B * 2 * d * (kv_e if kv_e is not None else e) * (ctx + 1) + B * 3 * d * 4 + B * d * e
    if tag == "head.cluster":     # 12 matrices of 512 x 512 + final 32 x 512 (bf16), input_proj fp32, modulation rows in, latents out
        F = dims.flow_dim
        return (2 * dims.flow_depth * F * F + dims.ldim * F) * e + F * dims.ldim * 4 + B * (3 * dims.flow_depth + 2) * F * 4 + 3 * B * dims.ldim * 4
    if tag == "head.adaln_mod":
        F = dims.flow_dim
        return (3 * dims.flow_depth + 2) * F * F * e + B * F * e + B * (3 * dims.flow_depth + 2) * F * 4
    gemm = {"flow.qkv": (3 * d, d), "flow.out_proj": (d, d), "flow.ffn1": (ffn, d), "flow.ffn2": (d, ffn),
            "mimi.qkv": (1536, 512), "mimi.out_proj": (512, 512), "mimi.ffn1": (2048, 512), "mimi.ffn2": (512, 2048)}
    if tag in gemm:
        N, K = gemm[tag]
        rows = B * (16 if tag.startswith("mimi") else 1)
        return N * K * e + rows * K * e + rows * N * 4
    conv = {  # (rows per stream, K, N, extra outputs)
        "seanet.conv0": (16, 3584, 512), "seanet.convtr0": (16, 1024, 1536), "seanet.res0.conv3": (96, 768, 128),
        "seanet.res0.conv1": (96, 128, 256), "seanet.convtr1": (96, 512, 640), "seanet.res1.conv3": (480, 384, 64),
        "seanet.res1.conv1": (480, 64, 128), "seanet.convtr2": (480, 256, 256), "seanet.res2.conv3": (1920, 192, 32),
        "seanet.res2.conv1": (1920, 32, 64)}
    if tag in conv:
        T, K, N = conv[tag]
        uniq_in = K // (7 if tag == "seanet.conv0" else (2 if "convtr" in tag else (3 if "conv3" in tag else 1)))
        out_mult = 2 if "convtr" in tag else 1
        res = N if "conv1" in tag else 0
        return N * K * e + B * T * (uniq_in + out_mult * N + res) * e
    return 0


# --------------------------------------------------------------------------- cfg5: long-text streaming through the API
def run_cfg5(args, rank, world, local_rank):
    """BASELINE.json configs[4] on one GPU of the box (every rank runs its own share, no collective): 256 concurrent
    streams per GPU, english 6L bf16, long texts as 8 chunks of 50 tokens, voices cloned from 10 s wav prompts by the
    device-side Mimi encoder and kept resident, every chunk forked from its voice without a copy (shared prefix),
    ragged completion with slot reuse, all through TTSModel.generate_audio_batch (host tensors in, host PCM out)."""
    import random

    import torch
    from oracle import pocket_oracle as po           # weights generator only
    from pocket_tts_b200.engine import ModelConfig
    from pocket_tts_b200.tts_model import TTSModel

    torch.cuda.set_device(local_rank)
    # the scheduler's host work (PCM block copies, noise draws) is torch CPU code: torchrun pins every rank to one OpenMP
    # thread by default, which made 8 ranks host-bound; give each rank its share of the box's cores
    torch.set_num_threads(max(1, (os.cpu_count() or 8) // max(1, world)))
    dims = po.PRESETS["english"]
    W = dict(po.make_weights(dims, seed=0))
    W.update(po.make_encoder_weights(dims, seed=0))
    S, CH, TOK, NV = 256, (8 if args.steps >= 8 else max(1, args.steps)), 50, 4
    R = S                                            # requests per GPU; each is a long text of CH chunks
    P = 126                                          # 10 s prompt -> 125 frames + bos_before_voice
    m = TTSModel.from_state_dict(ModelConfig(n_layers=dims.n_layers), W, None, precision=args.precision, max_streams=S + NV,
                                 kv_capacity=P + TOK + 234 + 8, max_prefill_rows=S * TOK, device=local_rank, eos_threshold=1e9)
    dist = world > 1
    if dist:
        import torch.distributed as td
        td.init_process_group("gloo")
    g = torch.Generator().manual_seed(100 + rank)
    t0 = time.perf_counter()
    voices = [m.get_state_for_audio_prompt(torch.randn(1, 240000, generator=g) * 0.1) for _ in range(NV)]
    m.engine.synchronize()
    voice_ms = (time.perf_counter() - t0) * 1e3 / NV
    rnd = random.Random(7 + rank)
    texts = [[torch.randint(0, dims.n_bins, (TOK,), generator=g).tolist() for _ in range(CH)] for _ in range(R)]
    limits = [rnd.randint(117, 234) for _ in range(R)]
    states = [voices[r % NV] for r in range(R)]
    # warm-up: captures the graphs of the full batch and touches every kernel module
    m.generate_audio_batch(states, [t[:1] for t in texts], frame_limits=[2 * args.block] * R, block=args.block)
    sampler = ClockSampler(local_rank)
    sampler.start()
    if dist:
        td.barrier()
    launches0 = m.engine.launch_count
    t0 = time.perf_counter()
    out = m.generate_audio_batch(states, texts, frame_limits=limits, block=args.block)
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    st = m.last_batch_stats
    frames = st["frames"]
    assert all(bool(torch.isfinite(o).all()) for o in out[:8])
    lat = st["first_block_ms"]
    # single-stream first chunk through generate_audio_stream, measured from the call: resident voice (shared fork)
    # against a plain dict of the same state (6 MB K/V upload + fork), tts_model.py:545-632
    first = {}
    if rank == 0:
        plain = {k: {kk: vv.clone() for kk, vv in v.items()} for k, v in voices[0].items()}
        for label, stt in (("resident_voice", voices[0]), ("uploaded_state", plain)):
            ts = []
            for trial in range(9):
                it = m.generate_audio_stream(stt, texts[trial][0])
                t1 = time.perf_counter()
                next(it)
                ts.append((time.perf_counter() - t1) * 1e3)
                it.close()
            first[f"p50_ms_{label}"] = round(statistics.median(ts[2:]), 3)
    audio_s = frames * FRAME_S
    if dist:
        from pocket_tts_b200.sharding import aggregate_throughput
        audio_s, wall_ms = aggregate_throughput(audio_s, wall * 1e3)
        wall = wall_ms / 1e3
    if rank != 0:
        if dist:
            td.destroy_process_group()
        return
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    steps = frames / S                                # engine steps if the batch stayed full
    ctx_avg = P + TOK + sum(limits) / len(limits) / 2
    step_bytes = algorithmic_bytes_per_step(dims, S, ctx_avg, 2 if args.precision == "bf16" else 4, 16 * sum(limits) / len(limits) / 2)
    ms_per_step = wall * 1e3 / steps
    value = audio_s / wall
    pct = lambda q: round(lat[min(len(lat) - 1, int(q * len(lat)))], 3) if lat else None
    line = {
        "metric": "audio-sec generated per sec (RTFx), batched", "value": round(value, 1), "unit": "audio-s/s",
        "n_gpus": world, "steps": int(steps), "warmup": 1, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": args.precision, "data": "synthetic",
        "config": {"workload": f"cfg5: default english model (6L FlowLM), {S} concurrent streams per GPU, {R} long texts of {CH} chunks x {TOK} tokens "
                               f"per GPU, {NV} voices cloned from 10 s wav prompts on the device (prefix {P}) and shared by their streams, "
                               "ragged completion (117-234 frames per chunk) with slot reuse, TTSModel.generate_audio_batch",
                   "streams_per_gpu": S, "precision": args.precision, "parallelism": f"streams sharded x{world}, no collective",
                   "l2_policy": "working set per step (weights + KV, > 1 GB) exceeds the 126 MB L2; no flush"},
        "e2e": {"value": round(value, 1), "unit": "audio-s/s", "h2d_bytes_per_step": S * dims.ldim * 4,
                "d2h_bytes_per_step": S * (4 + dims.frame_samples * 4), "ms_per_step": round(ms_per_step, 4),
                "api": "TTSModel.generate_audio_batch: token ids and noise from host tensors, PCM and EOS logits back to pinned host "
                       f"memory every block of {args.block} frames; the whole call is the timed region (value == e2e: there is no device-only form)"},
        "host_split": {"generate_frames_s": round(st["engine_s"], 3), "prefill_tokens_s": round(st["prefill_s"], 3), "call_s": round(st["total_s"], 3)},
        "first_chunk": {"batch_admission_to_first_block_p50_ms": pct(0.5), "batch_admission_to_first_block_p95_ms": pct(0.95),
                        "chunks": st["chunks"], "voice_prompt_encode_and_prefill_ms": round(voice_ms, 2), **first},
        "gpu_launches": int(m.engine.launch_count - launches0), "clocks": clocks,
        "roofline": {"kernel": "whole frame step (estimate: batch assumed full)", "bound": "hbm",
                     "achieved": round(step_bytes / (ms_per_step / 1e3) / 1e9, 1), "peak": hbm_peak, "unit": "GB/s",
                     "frac": round(step_bytes / (ms_per_step / 1e3) / 1e9 / hbm_peak, 4), "traffic": None,
                     "algorithmic_bytes_per_launch": int(step_bytes)},
        "cpu_baseline": cpu_baseline(dims, W, sample_streams=8, frames=4),
    }
    print(json.dumps(line), flush=True)
    if dist:
        td.destroy_process_group()


# --------------------------------------------------------------------------- CPU arm
def reference_cpu_arm(frames, warmup, n_procs=None):
    """The reference's OWN code (oracle/_ref, vendored by oracle/make_ref.py) on this box's host cores, as shipped:
    batch 1, torch.set_num_threads(1), generation thread + Mimi thread = 2 cores per utterance (tts_model.py:49,
    651-658,741-742), through TTSModel.generate_audio_stream; (i) one utterance alone, (ii) floor(cores / 2)
    utterances in as many processes pinned to disjoint core pairs.  Returns None when the copy is absent."""
    from oracle import ref_runner as rr
    if not rr.reference_available():
        return None
    cores = os.cpu_count() or 2
    P = n_procs or max(1, cores // 2)
    frames = max(4, min(frames, 180 - warmup))      # one utterance of 40 tokens is 192 frames long
    one = rr.run_processes(1, PRESET, N_LAYERS, VOICE_P, TEXT_T, warmup, frames)
    many = rr.run_processes(P, PRESET, N_LAYERS, VOICE_P, TEXT_T, warmup, frames) if P > 1 else one
    return {"as_shipped": {"value": round(one["aggregate_audio_s_per_s"], 3), "cores": 2, "processes": 1},
            "scaled": {"value": round(many["aggregate_audio_s_per_s"], 3), "cores": min(cores, 2 * P), "processes": P,
                       "per_process": many["per_process_audio_s_per_s"]},
            "frames": frames, "host_cores": cores}


def cpu_baseline(dims, W, sample_streams=8, frames=4, threads=None):
    """CPU baseline of the same workload on this box's host cores, bounded sample: the reference's own code when
    oracle/_ref exists (kind "reference"), else the oracle port (kind "port")."""
    try:
        ref = reference_cpu_arm(frames=24, warmup=4)
    except Exception as ex:   # the baseline must never break the bench line
        log("reference CPU arm failed, falling back to the oracle port:", repr(ex)[:300])
        ref = None
    if ref is not None:
        return {"value": ref["scaled"]["value"], "unit": "audio-s/s", "cores": ref["scaled"]["cores"], "kind": "reference",
                "sample": f"kyutai-labs/pocket-tts own code (oracle/_ref) as shipped, B=1 per process, 1 torch thread per stage, "
                          f"{ref['scaled']['processes']} processes x {ref['frames']} frames of one {TEXT_T}-token utterance at voice prefix {VOICE_P}",
                "as_shipped_2_cores": ref["as_shipped"]["value"], "host_cores": ref["host_cores"]}
    import torch
    from oracle import pocket_oracle as po
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    orc = po.Oracle(dims, W)
    tokens, voice, noise = synth_inputs(dims, sample_streams, frames + 1)
    st = orc.new_flow_state(sample_streams, VOICE_P + TEXT_T + frames + 2)
    with torch.no_grad():
        orc.prefill(st, orc.voice_conditioning(voice))
        orc.prefill(st, orc.embed_tokens(tokens))
        mst = orc.new_mimi_state(sample_streams, frames + 1)
        prev = torch.full((sample_streams, dims.ldim), float("nan"))
        lat, _ = orc.flow_step(st, prev, noise[0]); orc.mimi_step(mst, lat)        # warm-up frame
        t0 = time.perf_counter()
        for g in range(1, frames + 1):
            lat, _ = orc.flow_step(st, lat, noise[g])
            orc.mimi_step(mst, lat)
        dt = time.perf_counter() - t0
    return {"value": round(sample_streams * frames * FRAME_S / dt, 2), "unit": "audio-s/s", "cores": threads,
            "kind": "port", "sample": f"oracle (torch fp32 CPU restatement), {sample_streams} streams x {frames} frames "
            f"at context {VOICE_P + TEXT_T}, {threads} torch threads"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    K, Wm = args.steps, args.warmup
    try:
        ref = reference_cpu_arm(frames=K, warmup=Wm)
    except Exception as ex:
        log("reference CPU arm failed, falling back to the oracle port:", repr(ex)[:300])
        ref = None
    if ref is not None:
        value, P = ref["scaled"]["value"], ref["scaled"]["processes"]
        sample = (f"kyutai-labs/pocket-tts own code (oracle/_ref) as shipped: B=1, torch.set_num_threads(1), generation thread + Mimi "
                  f"thread per utterance; {P} concurrent utterances in {P} processes pinned to core pairs, {ref['frames']} timed frames "
                  f"each after {Wm} warm-up frames (a step = one 80 ms frame of every utterance)")
        line = {"impl": "reference", "metric": "audio-sec generated per sec (RTFx), batched", "value": value,
                "unit": "audio-s/s", "n_gpus": world, "steps": ref["frames"], "warmup": Wm,
                "ms_per_step": round(P * FRAME_S / value * 1e3, 3), "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{'default english' if PRESET == 'english' else PRESET} model ({N_LAYERS}L FlowLM), voice prefix {VOICE_P} + "
                                       f"{TEXT_T} text tokens; CPU arm: {P} concurrent utterances (all host cores) instead of {STREAMS} streams per GPU"},
                "cpu_baseline": {"value": value, "unit": "audio-s/s", "cores": ref["scaled"]["cores"], "kind": "reference", "sample": sample,
                                 "as_shipped_2_cores": ref["as_shipped"]["value"], "host_cores": ref["host_cores"],
                                 "per_process": ref["scaled"]["per_process"]},
                "e2e": {"value": value, "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return
    import torch
    from oracle import pocket_oracle as po
    dims = po.PRESETS[PRESET]
    W = po.make_weights(dims, seed=0)
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    Bs = 16   # bounded sample of the 64-stream workload per step
    orc = po.Oracle(dims, W)
    tokens, voice, noise = synth_inputs(dims, Bs, Wm + K)
    with torch.no_grad():
        st = orc.new_flow_state(Bs, VOICE_P + TEXT_T + Wm + K + 2)
        orc.prefill(st, orc.voice_conditioning(voice))
        orc.prefill(st, orc.embed_tokens(tokens))
        mst = orc.new_mimi_state(Bs, Wm + K)
        lat = torch.full((Bs, dims.ldim), float("nan"))
        for g in range(Wm):
            lat, _ = orc.flow_step(st, lat, noise[g]); orc.mimi_step(mst, lat)
        t0 = time.perf_counter()
        for g in range(Wm, Wm + K):
            lat, _ = orc.flow_step(st, lat, noise[g]); orc.mimi_step(mst, lat)
        dt = time.perf_counter() - t0
    value = Bs * K * FRAME_S / dt
    sample = f"oracle port of the reference CPU path, {Bs} of the {STREAMS} streams per step, {threads} torch threads"
    line = {"impl": "reference", "metric": "audio-sec generated per sec (RTFx), batched", "value": round(value, 2),
            "unit": "audio-s/s", "n_gpus": world, "steps": K, "warmup": Wm, "ms_per_step": round(dt / K * 1e3, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{'default english' if PRESET == 'english' else PRESET} model ({dims.n_layers}L FlowLM), {STREAMS} concurrent streams per GPU, "
                                   f"voice prefix {VOICE_P} + {TEXT_T} text tokens (CPU arm: {Bs}-stream sample per step)"},
            "cpu_baseline": {"value": round(value, 2), "unit": "audio-s/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": round(value, 2), "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=96)
    ap.add_argument("--warmup", type=int, default=4)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--precision", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--block", type=int, default=8, help="cfg5: frames per engine call of the batch scheduler")
    ap.add_argument("--kv", default="act", choices=["act", "fp8"],
                    help="FlowLM KV cache storage: the activation type (default) or FP8 e4m3 with per-row scales (bf16 mode)")
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg3", "cfg5"],
                    help="cfg2: english 6L, 64 streams per GPU (the headline line); cfg3: italian_24l, 256 streams per GPU; "
                         "cfg5: 256 streams per GPU, long texts in 50-token chunks from wav-prompt voices through generate_audio_batch")
    args = ap.parse_args()
    if args.workload == "cfg3":
        global STREAMS, PRESET, N_LAYERS
        STREAMS, PRESET, N_LAYERS = 256, "italian_24l", 24
    args.warmup = max(args.warmup, 3)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    elif args.workload == "cfg5":
        run_cfg5(args, rank, world, local_rank)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
