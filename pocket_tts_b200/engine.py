"""Python host side of the generation engine: owns a native engine handle, feeds it the
reference-format state_dict and exposes prefill / step / decode on torch tensors.

PyTorch is plumbing here (host/device buffers); all arithmetic of the generation path runs in
the CUDA library behind include/pocket_tts_b200.h."""

from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import torch

from . import _native as N


@dataclass(frozen=True)
class ModelConfig:
    """Decode-path architecture (pocket_tts/config/*.yaml, utils/config.py:11-118)."""
    n_layers: int = 6
    d_model: int = 1024
    n_heads: int = 16
    ffn: int = 4096
    ldim: int = 32
    flow_dim: int = 512
    flow_depth: int = 6
    max_period: float = 10000.0
    n_bins: int = 4000
    insert_bos_before_voice: bool = True
    mimi_d: int = 512
    mimi_heads: int = 8
    mimi_layers: int = 2
    mimi_ffn: int = 2048
    mimi_context: int = 250
    n_filters: int = 64
    ratios: tuple = (6, 5, 4)
    kernel: int = 7
    res_kernel: int = 3
    last_kernel: int = 3
    upsample_stride: int = 16
    frame_rate: float = 12.5
    sample_rate: int = 24000

    @property
    def frame_samples(self) -> int:
        return int(self.sample_rate / self.frame_rate)


def rope_freqs(head_dim: int, max_period: float) -> torch.Tensor:
    """rope.py:28-29, evaluated exactly as the reference does (torch fp32)."""
    ds = torch.arange(head_dim // 2, dtype=torch.float32)
    return torch.exp(ds * (-math.log(max_period) * 2 / head_dim))


def flow_time_table(sd: dict, lsd_steps: int) -> torch.Tensor:
    """Input-independent constant of the flow head: (te0(i/n) + te1((i+1)/n)) / 2 for each LSD
    step i (mlp.py:58-83,205-208 at the times of flow_lm.py:33-37).  Folded at load time."""
    def te(idx: int, tau: float) -> torch.Tensor:
        p = f"flow_lm.flow_net.time_embed.{idx}."
        args = torch.tensor(float(tau), dtype=torch.float32) * sd[p + "freqs"].float()
        emb = torch.cat([torch.cos(args), torch.sin(args)], dim=-1)
        hid = torch.nn.functional.silu(emb @ sd[p + "mlp.0.weight"].float().t() + sd[p + "mlp.0.bias"].float())
        hid = hid @ sd[p + "mlp.2.weight"].float().t() + sd[p + "mlp.2.bias"].float()
        var = 1e-5 + hid.var(dim=-1, keepdim=True)  # mlp.py:20-25 (unbiased variance)
        return hid * (sd[p + "mlp.3.alpha"].float() * torch.rsqrt(var))
    rows = [(te(0, i / lsd_steps) + te(1, (i + 1) / lsd_steps)) / 2 for i in range(lsd_steps)]
    return torch.stack(rows).contiguous()


def _ptr(t: torch.Tensor | None):
    return None if t is None else C.c_void_p(t.data_ptr())


def _i32(values) -> "C.Array":
    return (C.c_int32 * len(values))(*[int(v) for v in values])


class GenerationEngine:
    """One engine per GPU: weights, stream slots, frame step."""

    def __init__(self, cfg: ModelConfig, state_dict: dict, *, max_streams: int = 64, kv_capacity: int = 1024,
                 precision: str = "fp32", device: int = 0, lsd_steps: int = 1, max_prefill_rows: int = 4096,
                 kv_dtype: str | None = None):
        self.cfg = cfg
        self.lib = N.lib()
        self.device = torch.device("cuda", device)
        self.precision = precision
        self.lsd_steps = lsd_steps
        self.max_streams = max_streams
        self.kv_capacity = kv_capacity
        self.max_prefill_rows = max_prefill_rows
        pc = N.PttsConfig()
        for k in ("n_layers", "d_model", "n_heads", "ffn", "ldim", "flow_dim", "flow_depth", "mimi_d",
                  "mimi_heads", "mimi_layers", "mimi_ffn", "mimi_context", "n_filters", "kernel",
                  "res_kernel", "last_kernel", "upsample_stride"):
            setattr(pc, k, int(getattr(cfg, k)))
        pc.ratios = (C.c_int32 * 3)(*cfg.ratios)
        pc.max_streams, pc.kv_capacity = max_streams, kv_capacity
        pc.max_prefill_rows, pc.max_lsd_steps = max_prefill_rows, lsd_steps
        pc.precision = {"fp32": N.PRECISION_FP32, "bf16": N.PRECISION_BF16}[precision]
        pc.device = device
        if kv_dtype not in (None, "fp8"):
            raise ValueError("kv_dtype is None (cache in the activation type) or 'fp8'")
        if kv_dtype == "fp8" and precision != "bf16":
            raise ValueError("the FP8 KV cache belongs to the bf16 mode")
        pc.kv_fp8 = 1 if kv_dtype == "fp8" else 0
        self.kv_dtype = kv_dtype
        h = C.c_void_p()
        N.check(self.lib.ptts_engine_create(C.byref(pc), C.byref(h)))
        self._h = h
        try:
            self._load(state_dict)
        except Exception:
            self.close()
            raise

    # ---- weights
    def _load_one(self, name: str, t: torch.Tensor) -> None:
        t = t.detach().to("cpu", torch.float32).contiguous()
        shape = (C.c_int64 * t.dim())(*t.shape)
        N.check(self.lib.ptts_load_tensor(self._h, name.encode(), _ptr(t), shape, t.dim()))

    def _load(self, sd: dict) -> None:
        skip = ("flow_lm.conditioner.", "flow_lm.bos_before_voice",
                "flow_lm.flow_net.time_embed.")  # host-side only / folded into flow.tcomb
        keep = ("flow_lm.conditioner.embed.weight",)  # device-side embedding gather of prefill_tokens
        # the Mimi encoder side (mimi.encoder*, mimi.downsample*, speaker_proj_weight) is loaded when the state_dict
        # carries it: it enables encode_audio (voice-prompt encoding on the device)
        self.has_encoder = "mimi.encoder.model.0.conv.weight" in sd
        for name, t in sd.items():
            if name.startswith(skip) and name not in keep:
                continue
            self._load_one(name, t)
        self._load_one("rope.freqs", rope_freqs(self.cfg.d_model // self.cfg.n_heads, self.cfg.max_period))
        self._load_one("flow.tcomb", flow_time_table(sd, self.lsd_steps))
        N.check(self.lib.ptts_finalize(self._h))

    def close(self) -> None:
        if getattr(self, "_h", None):
            self.lib.ptts_engine_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- streams
    def open_stream(self) -> int:
        s = C.c_int32(-1)
        N.check(self.lib.ptts_stream_open(self._h, C.byref(s)))
        return s.value

    def close_stream(self, slot: int) -> None:
        N.check(self.lib.ptts_stream_close(self._h, slot))

    def reset_decoder(self, slot: int) -> None:
        N.check(self.lib.ptts_stream_reset_decoder(self._h, slot))

    def fork(self, src: int, dst: int) -> None:
        N.check(self.lib.ptts_stream_fork(self._h, src, dst))

    def fork_shared(self, src: int, dst: int) -> None:
        """Zero-copy fork: `dst` reads `src`'s positions from their owner's cache (one resident voice prefix for every
        stream of that voice) and appends behind them; `src` (or its owner) must outlive `dst`."""
        N.check(self.lib.ptts_stream_fork_shared(self._h, src, dst))

    def offset(self, slot: int) -> int:
        o = C.c_int32(0)
        N.check(self.lib.ptts_stream_get_offset(self._h, slot, C.byref(o)))
        return o.value

    def set_prev_latent(self, slot: int, latent: torch.Tensor) -> None:
        t = latent.detach().to(torch.float32).contiguous()
        N.check(self.lib.ptts_stream_set_prev_latent(self._h, slot, _ptr(t)))

    def import_state(self, slot: int, model_state: dict) -> None:
        """model_state in the reference format (transformer.py:46-57): per layer
        {"offset": int64[1], "cache": float[2,1,T,H,D]}; positions >= offset are ignored."""
        off = None
        for i in range(self.cfg.n_layers):
            st = model_state[f"transformer.layers.{i}.self_attn"]
            o = int(st["offset"].view(-1)[0])
            off = o if off is None else off
            if o != off:
                raise ValueError("model_state layers disagree on offset")
            cache = st["cache"][:, 0, :o].detach().to(torch.float32).contiguous()
            N.check(self.lib.ptts_stream_import_kv(self._h, slot, i, _ptr(cache), o))
        N.check(self.lib.ptts_stream_set_offset(self._h, slot, off))

    def export_state(self, slot: int) -> dict:
        o = self.offset(slot)
        out = {}
        for i in range(self.cfg.n_layers):
            cache = torch.empty(2, 1, o, self.cfg.n_heads, self.cfg.d_model // self.cfg.n_heads)
            N.check(self.lib.ptts_stream_export_kv(self._h, slot, i, _ptr(cache), o))
            out[f"transformer.layers.{i}.self_attn"] = {
                "offset": torch.full((1,), o, dtype=torch.long), "cache": cache}
        return out

    # ---- compute
    def prefill(self, slots, embeds: list) -> None:
        """embeds[i]: [len_i, d_model] conditioning rows for stream slots[i]."""
        lens = [int(e.shape[0]) for e in embeds]
        flat = torch.cat([e.detach().to(torch.float32).reshape(-1, self.cfg.d_model) for e in embeds]).contiguous()
        N.check(self.lib.ptts_prefill(self._h, len(slots), _i32(slots), _i32(lens), _ptr(flat)))

    def prefill_tokens(self, slots, tokens: list) -> None:
        """tokens[i]: 1-D integer ids of the text chunk of stream slots[i]; the embedding gather
        (conditioners/text.py:74-76) runs on the device."""
        lens = [int(len(t)) for t in tokens]
        flat = [int(v) for t in tokens for v in (t.tolist() if hasattr(t, "tolist") else t)]
        N.check(self.lib.ptts_prefill_tokens(self._h, len(slots), _i32(slots), _i32(lens), _i32(flat)))

    def step(self, slots, noise: torch.Tensor, *, lsd_steps: int | None = None, decode_audio: bool = True,
             out_device: str = "cpu"):
        """One frame.  Returns (latents [n,ldim], eos_logits [n], pcm [n,1920] | None)."""
        n = len(slots)
        noise = noise.detach().to(torch.float32).contiguous()
        assert noise.shape == (n, self.cfg.ldim), noise.shape
        dev = self.device if out_device == "cuda" else "cpu"
        lat = torch.empty(n, self.cfg.ldim, dtype=torch.float32, device=dev)
        eos = torch.empty(n, dtype=torch.float32, device=dev)
        pcm = torch.empty(n, self.cfg.frame_samples, dtype=torch.float32, device=dev) if decode_audio else None
        N.check(self.lib.ptts_step(self._h, n, _i32(slots), _ptr(noise), lsd_steps or self.lsd_steps,
                                   1 if decode_audio else 0, _ptr(lat), _ptr(eos), _ptr(pcm)))
        return lat, eos, pcm

    def mimi_decode(self, slots, latents: torch.Tensor) -> torch.Tensor:
        n = len(slots)
        latents = latents.detach().to(torch.float32).contiguous()
        pcm = torch.empty(n, self.cfg.frame_samples, dtype=torch.float32)
        N.check(self.lib.ptts_mimi_decode(self._h, n, _i32(slots), _ptr(latents), _ptr(pcm)))
        return pcm

    def mimi_decode_frames(self, slots, latents: torch.Tensor) -> torch.Tensor:
        """latents [F, n, ldim] (host or device) -> PCM [n, F*1920] where the latents live, decoded in grouped
        passes of up to 4 frames per stream."""
        F, n, _ = latents.shape
        assert n == len(slots)
        latents = latents.detach().to(torch.float32).contiguous()
        pcm = torch.empty(n, F * self.cfg.frame_samples, dtype=torch.float32, device=latents.device)
        N.check(self.lib.ptts_mimi_decode_frames(self._h, n, _i32(slots), F, _ptr(latents), _ptr(pcm)))
        return pcm

    def run_frames(self, slots, noise_dev: torch.Tensor, *, lsd_steps: int | None = None,
                   out_pcm: torch.Tensor | None = None, out_latents: torch.Tensor | None = None) -> None:
        """Device-resident loop: noise_dev [F, n, ldim] on the GPU; asynchronous."""
        assert noise_dev.is_cuda and noise_dev.dtype == torch.float32 and noise_dev.is_contiguous()
        F, n, _ = noise_dev.shape
        # the call is asynchronous on the engine's own streams: keep the tensors alive until synchronize(), or torch's
        # caching allocator may hand their memory to someone else while the engine still reads / writes it
        self._inflight = getattr(self, "_inflight", [])
        self._inflight.append((noise_dev, out_pcm, out_latents))
        N.check(self.lib.ptts_run_frames(self._h, n, _i32(slots), F, _ptr(noise_dev), lsd_steps or self.lsd_steps,
                                         _ptr(out_pcm), _ptr(out_latents)))

    def alloc_frame_outputs(self, n: int, F: int, *, want_latents: bool = True, pin: bool = True):
        """Host buffers for `generate_frames` (pinned by default): (pcm [n, F*1920], latents [F, n, ldim] | None,
        eos_logits [F, n]).  Allocate once and pass as `out=`: pinning memory costs milliseconds."""
        kw = dict(dtype=torch.float32, pin_memory=pin)
        pcm = torch.empty(n, F * self.cfg.frame_samples, **kw)
        lat = torch.empty(F, n, self.cfg.ldim, **kw) if want_latents else None
        eos = torch.empty(F, n, **kw)
        return pcm, lat, eos

    def generate_frames(self, slots, noise: torch.Tensor, *, lsd_steps: int | None = None, want_latents: bool = True,
                        pin: bool = True, out=None):
        """Streaming generation with HOST buffers: noise [F, n, ldim] (host) ->
        (pcm [n, F*1920], latents [F, n, ldim] | None, eos_logits [F, n]) on the host.  The FlowLM frames of a
        group overlap the Mimi pass of the previous group on the device; result copies ride a side stream."""
        F, n, _ = noise.shape
        assert n == len(slots)
        noise = noise.detach().to(torch.float32).contiguous()
        if pin and not noise.is_pinned() and not noise.is_cuda:
            noise = noise.pin_memory()
        pcm, lat, eos = out if out is not None else self.alloc_frame_outputs(n, F, want_latents=want_latents, pin=pin)
        assert pcm.shape == (n, F * self.cfg.frame_samples) and eos.shape == (F, n)
        N.check(self.lib.ptts_generate_frames(self._h, n, _i32(slots), F, _ptr(noise), lsd_steps or self.lsd_steps,
                                              _ptr(pcm), _ptr(lat), _ptr(eos)))
        return pcm, lat, eos

    def pcm_to_int16(self, pcm: torch.Tensor) -> torch.Tensor:
        """16-bit samples as the reference's WAV writer makes them (data/audio.py:77-78), converted on the device;
        the result lives where the input lives."""
        pcm = pcm.detach().to(torch.float32).contiguous()
        out = torch.empty(pcm.shape, dtype=torch.int16, device=pcm.device)
        N.check(self.lib.ptts_pcm_to_int16(self._h, _ptr(pcm), _ptr(out), pcm.numel()))
        return out

    ENCODE_MAX_SAMPLES = 1440000

    def encode_audio(self, audio: torch.Tensor) -> torch.Tensor:
        """TTSModel._encode_audio (tts_model.py:379-388) on the device: mono 24 kHz waveform [samples] or [1, samples]
        -> conditioning rows [ceil(samples / 1920), d_model] (fp32, on the host), to be prefilled after
        bos_before_voice.  Needs the Mimi encoder weights in the state_dict (models with voice cloning)."""
        a = audio.detach().to(torch.float32).reshape(-1).contiguous()
        nf = C.c_int32(0)
        N.check(self.lib.ptts_encode_audio(self._h, _ptr(a), a.numel(), None, 0, C.byref(nf)))
        out = torch.empty(nf.value, self.cfg.d_model, dtype=torch.float32)
        N.check(self.lib.ptts_encode_audio(self._h, _ptr(a), a.numel(), _ptr(out), nf.value, C.byref(nf)))
        return out

    PCM_STATIC_MAX_SAMPLES = 262144

    def pcm_to_int16_static(self, pcm: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
        """`pcm_to_int16` without any device allocation inside the library: host buffers are staged through a fixed
        device array (at most PCM_STATIC_MAX_SAMPLES samples per call).  Pass `out` (int16, same shape and place
        as `pcm`) to keep the call allocation-free on the Python side too."""
        assert pcm.dtype == torch.float32 and pcm.is_contiguous()
        if out is None:
            out = torch.empty(pcm.shape, dtype=torch.int16, device=pcm.device)
        assert out.dtype == torch.int16 and out.is_contiguous() and out.numel() == pcm.numel()
        N.check(self.lib.ptts_pcm_to_int16_static(self._h, _ptr(pcm), _ptr(out), pcm.numel()))
        return out

    def synchronize(self) -> None:
        N.check(self.lib.ptts_synchronize(self._h))
        self._inflight = []

    @property
    def launch_count(self) -> int:
        return int(self.lib.ptts_launch_count(self._h))

    @property
    def cuda_stream(self) -> int:
        return int(self.lib.ptts_stream_handle(self._h) or 0)

    def set_graphs(self, enabled: bool) -> None:
        N.check(self.lib.ptts_set_graphs(self._h, 1 if enabled else 0))

    def debug_gemm(self, backend: int, a: torch.Tensor, w: torch.Tensor, *, M: int, K: int, ld: int,
                   rows_per_batch: int = 0, valid_rows: int = 0) -> torch.Tensor:
        a = a.detach().to(torch.float32).contiguous()
        w = w.detach().to(torch.float32).contiguous()
        out = torch.empty(M, w.shape[0], dtype=torch.float32)
        N.check(self.lib.ptts_debug_gemm(self._h, backend, M, w.shape[0], K, ld, rows_per_batch, valid_rows,
                                         _ptr(a), a.numel(), _ptr(w), _ptr(out)))
        return out

    def profile_begin(self) -> None:
        N.check(self.lib.ptts_profile_begin(self._h))

    def profile_end(self) -> dict:
        import json

        buf = C.create_string_buffer(1 << 16)
        N.check(self.lib.ptts_profile_end(self._h, buf, len(buf)))
        return json.loads(buf.value.decode())

    def timeline(self, max_records: int = 2048) -> list:
        """In-graph device timeline (engine created with PTTS_TIMELINE=1): [[name, entry_us, wait_us, exit_us], ...]."""
        import json

        buf = C.create_string_buffer(1 << 20)
        N.check(self.lib.ptts_debug_timeline(self._h, buf, len(buf), max_records))
        return json.loads(buf.value.decode())

    def debug_read(self, what: str, numel: int) -> torch.Tensor:
        out = torch.empty(numel, dtype=torch.float32)
        got = C.c_int64(0)
        N.check(self.lib.ptts_debug_read(self._h, what.encode(), _ptr(out), numel, C.byref(got)))
        return out[: got.value]
