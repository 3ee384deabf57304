"""Drop-in host for the reference's public Python API over the B200 engine.

`TTSModel` keeps the names, argument meaning and error behaviour of
pocket_tts/models/tts_model.py (`load_model` :232-315, `get_state_for_audio_prompt` :788-905,
`generate_audio` :477-543, `generate_audio_stream` :545-632, `device`, `sample_rate`,
`export_model_state` :1047-1072) and adds the batched entry the reference lacks
(`generate_audio_batch`).  The frame loop of `_autoregressive_generation` (:745-779) and the decoder
thread (:434-474) run on the GPU through `GenerationEngine.generate_frames`; tokenisation, sentence
splitting, voice-prompt encoding and WAV I/O stay on the host.

State ownership follows the reference: `copy_state=True` never mutates the caller's model_state (the
voice prefix is imported into a slot once per call and forked on the device per text chunk);
`copy_state=False` advances and expands the caller's state in place."""
from __future__ import annotations

import os

import logging
import math
from dataclasses import dataclass, field
from pathlib import Path
from typing import Callable, Iterator, Sequence

import torch

from .engine import GenerationEngine, ModelConfig
from .text import (MAX_TOKEN_PER_CHUNK, ReferenceTokenizerAdapter, SentencePieceAdapter, Tokenizer,
                   estimate_max_gen_len, prepare_text_prompt, split_into_best_sentences)

logger = logging.getLogger(__name__)

VOICE_CLONING_UNSUPPORTED = (
    "Voice cloning from raw audio needs the Mimi encoder weights (mimi.encoder*, mimi.downsample*), which this "
    "state_dict does not carry (a model without voice cloning): pass a .safetensors model state or a "
    "pre-encoded conditioning tensor [P, d_model]."
)
_BLOCK_SCHEDULE = (1, 1, 2, 4)  # frames per engine call at the start of a chunk (first PCM after one frame)
_BLOCK_MAX = 8


def export_model_state(model_state: dict, dest) -> None:
    """tts_model.py:1047-1052: safetensors with keys "{module}/{key}"."""
    import safetensors.torch

    flat = {f"{mod}/{k}": v.contiguous() for mod, st in model_state.items() for k, v in st.items()}
    safetensors.torch.save_file(flat, str(dest))


def import_model_state(source, device="cpu") -> dict:
    """tts_model.py:1055-1072, including the legacy `current_end` key (offset = its length)."""
    import safetensors

    out: dict = {}
    with safetensors.safe_open(str(source), framework="pt") as f:
        for key in f.keys():
            mod, name = key.split("/")
            t = f.get_tensor(key)
            if name == "current_end":
                out.setdefault(mod, {})["offset"] = torch.full((1,), t.shape[0], dtype=torch.long)
            else:
                out.setdefault(mod, {})[name] = t.to(device)
    return out


class VoiceState(dict):
    """A model_state dict in the reference format (what get_state_for_audio_prompt returns and generate_audio takes)
    that also remembers the engine slot holding the same K/V on the device, so that every generation from this voice
    forks that resident prefix (ptts_stream_fork_shared) instead of uploading the cache again."""
    resident = None   # (engine token, slot) or None


def _state_offset(model_state: dict) -> int:
    """tts_model.py:423-431."""
    for st in model_state.values():
        if "offset" in st:
            return int(st["offset"].view(-1)[0])
    raise ValueError("Could not find offset in model state.")


@dataclass
class _Job:
    """One text chunk of one request on one stream slot."""
    request: int
    tokens: list
    frames_after_eos: int
    max_gen_len: int
    noise: torch.Tensor | None = None
    slot: int = -1
    done: int = 0            # frames decoded so far
    eos_step: int | None = None
    limit: int | None = None  # frames to emit once EOS is known
    pcm: list = field(default_factory=list)
    t_admit: float | None = None   # generate_audio_batch: slot opened / first PCM block on the host
    t_first: float | None = None

    def target(self) -> int:
        return self.max_gen_len if self.limit is None else self.limit


class TTSModel:
    """B200 engine behind the reference's `TTSModel` interface."""

    def __init__(self, cfg: ModelConfig, state_dict: dict, tokenizer: Tokenizer | None, *, temp: float = 0.7,
                 lsd_decode_steps: int = 1, noise_clamp: float | None = None, eos_threshold: float = -4.0,
                 precision: str = "bf16", max_streams: int = 64, kv_capacity: int = 1024, device: int = 0,
                 pad_with_spaces_for_short_inputs: bool = False, remove_semicolons: bool = False,
                 model_recommended_frames_after_eos: int | None = None,
                 audio_encoder: Callable[[torch.Tensor], torch.Tensor] | None = None,
                 max_prefill_rows: int = 4096):
        self.cfg = cfg
        self.temp = temp
        self.lsd_decode_steps = lsd_decode_steps
        self.noise_clamp = noise_clamp
        self.eos_threshold = eos_threshold
        self.tokenizer = tokenizer
        self.pad_with_spaces_for_short_inputs = pad_with_spaces_for_short_inputs
        self.remove_semicolons = remove_semicolons
        self.model_recommended_frames_after_eos = model_recommended_frames_after_eos
        self.audio_encoder = audio_encoder          # optional host override; default is the engine's own encoder
        self._embed = state_dict["flow_lm.conditioner.embed.weight"].detach().float().cpu()
        bos = state_dict.get("flow_lm.bos_before_voice")
        self._bos_before_voice = None if bos is None else bos.detach().float().cpu().reshape(1, cfg.d_model)
        self.engine = GenerationEngine(cfg, state_dict, max_streams=max_streams, kv_capacity=kv_capacity,
                                       precision=precision, device=device, lsd_steps=lsd_decode_steps,
                                       max_prefill_rows=max_prefill_rows)
        # voice cloning = the Mimi encoder side is in the state_dict (tts_model.py:871-872) -> ptts_encode_audio
        self.has_voice_cloning = self.engine.has_encoder or audio_encoder is not None
        self._outs: dict = {}
        # voice prefixes kept resident on the device: slot -> True, oldest first
        self._resident: dict = {}
        self.max_resident_voices = max(1, min(8, max_streams // 8))

    # ------------------------------------------------------------------ construction
    @classmethod
    def load_model(cls, language: str | None = None, config=None, temp: float = 0.7, lsd_decode_steps: int = 1,
                   noise_clamp: float | None = None, eos_threshold: float = -4.0, quantize: bool = False,
                   **engine_kwargs) -> "TTSModel":
        """Same signature and argument checks as the reference (tts_model.py:232-315).  Config parsing,
        weight download and the tokenizer are host-side pieces of the reference package: when it is
        importable they are reused as they are and the weights move into the engine."""
        if language is not None and config is not None:
            raise ValueError("Only one of `language` or `config` can be specified, not both.")
        if quantize:
            raise ValueError("quantize=True is the reference's CPU int8 path; the B200 engine offers "
                             "precision='bf16' or 'fp32'.")
        try:
            from pocket_tts import TTSModel as RefTTSModel
        except ImportError as e:
            raise RuntimeError("TTSModel.load_model resolves configs / weights / tokenizer through the reference "
                               "package `pocket_tts`, which is not importable here; build the model with "
                               "TTSModel.from_state_dict(cfg, state_dict, tokenizer) instead.") from e
        kw = dict(temp=temp, lsd_decode_steps=lsd_decode_steps, noise_clamp=noise_clamp, eos_threshold=eos_threshold)
        if language is not None:
            kw["language"] = language
        if config is not None:
            kw["config"] = config
        return cls.from_reference(RefTTSModel.load_model(**kw), **engine_kwargs)

    @classmethod
    def from_reference(cls, ref, **engine_kwargs) -> "TTSModel":
        """Take weights, config values, tokenizer and the (host) voice-prompt encoder of an instantiated
        reference model."""
        c = ref.config
        tr, fm = c.flow_lm.transformer, c.flow_lm.flow
        mt, sn = c.mimi.transformer, c.mimi.seanet
        cfg = ModelConfig(n_layers=tr.num_layers, d_model=tr.d_model, n_heads=tr.num_heads,
                          ffn=int(tr.d_model * tr.hidden_scale), ldim=c.mimi.quantizer.dimension,
                          flow_dim=fm.dim, flow_depth=fm.depth, max_period=float(tr.max_period),
                          n_bins=c.flow_lm.lookup_table.n_bins,
                          insert_bos_before_voice=bool(getattr(ref.flow_lm, "insert_bos_before_voice", False)),
                          mimi_d=mt.d_model, mimi_heads=mt.num_heads, mimi_layers=mt.num_layers,
                          mimi_ffn=mt.dim_feedforward, mimi_context=mt.context, n_filters=sn.n_filters,
                          ratios=tuple(sn.ratios), kernel=sn.kernel_size, res_kernel=sn.residual_kernel_size,
                          last_kernel=sn.last_kernel_size,
                          upsample_stride=int(ref.mimi.encoder_frame_rate / ref.mimi.frame_rate),
                          frame_rate=float(c.mimi.frame_rate), sample_rate=int(c.mimi.sample_rate))
        sd = ref.state_dict()
        if not getattr(ref, "has_voice_cloning", False):   # encoder weights of such a checkpoint are uninitialised
            sd = {k: v for k, v in sd.items() if not k.startswith(("mimi.encoder", "mimi.downsample"))}
        enc = None
        origin = getattr(ref, "origin", None)
        m = cls(cfg, sd, ReferenceTokenizerAdapter(ref.flow_lm.conditioner.tokenizer),
                   temp=ref.temp, lsd_decode_steps=ref.lsd_decode_steps, noise_clamp=ref.noise_clamp,
                   eos_threshold=ref.eos_threshold,
                   pad_with_spaces_for_short_inputs=getattr(ref, "pad_with_spaces_for_short_inputs", False),
                   remove_semicolons=getattr(ref, "remove_semicolons", False),
                   model_recommended_frames_after_eos=getattr(ref, "model_recommended_frames_after_eos", None),
                   audio_encoder=enc, **engine_kwargs)
        m.origin = origin          # config path of the language: predefined voices are resolved against it
        return m

    @classmethod
    def from_state_dict(cls, cfg: ModelConfig, state_dict: dict, tokenizer: Tokenizer | str | Path | None = None,
                        **kwargs) -> "TTSModel":
        """Build from a reference-format state_dict (TTSModel.state_dict() names) without the reference
        package; `tokenizer` may be a sentencepiece model path."""
        if isinstance(tokenizer, (str, Path)):
            tokenizer = SentencePieceAdapter(str(tokenizer), cfg.n_bins)
        return cls(cfg, state_dict, tokenizer, **kwargs)

    # ------------------------------------------------------------------ properties (tts_model.py:92-98)
    @property
    def device(self) -> str:
        return self.engine.device.type

    @property
    def sample_rate(self) -> int:
        return self.cfg.sample_rate

    # ------------------------------------------------------------------ helpers
    def _draw_noise(self, n: int) -> torch.Tensor:
        """flow_lm.py:131-137: one [n, ldim] draw from the global torch RNG."""
        noise = torch.empty(n, self.cfg.ldim, dtype=torch.float32)
        std = self.temp ** 0.5
        if self.noise_clamp is None:
            torch.nn.init.normal_(noise, mean=0.0, std=std)
        else:
            torch.nn.init.trunc_normal_(noise, mean=0.0, std=std, a=-self.noise_clamp, b=self.noise_clamp)
        return noise

    def _tokens_of(self, text: str) -> list:
        if self.tokenizer is None:
            raise ValueError("this model was built without a tokenizer; pass token ids instead of text")
        return self.tokenizer.encode(text)

    def _text_rows(self, tokens: Sequence[int]) -> torch.Tensor:
        """conditioners/text.py:74-76: embedding gather (host; feeds prefill only)."""
        return self._embed[torch.as_tensor(list(tokens), dtype=torch.long)]

    def _frames_after_eos(self, chunk: str, frames_after_eos: int | None) -> int:
        """tts_model.py:603-625."""
        if frames_after_eos is None:
            frames_after_eos = self.model_recommended_frames_after_eos
        _, guess = prepare_text_prompt(chunk, self.pad_with_spaces_for_short_inputs, self.remove_semicolons)
        return frames_after_eos if frames_after_eos is not None else guess + 2

    def _jobs_for_text(self, request: int, text, max_tokens: int, frames_after_eos: int | None) -> list:
        if isinstance(text, str):
            chunks = split_into_best_sentences(self.tokenizer_or_raise(), text, max_tokens,
                                               self.pad_with_spaces_for_short_inputs, self.remove_semicolons)
            jobs = []
            for chunk in chunks:
                toks = self._tokens_of(chunk)
                jobs.append(_Job(request, toks, self._frames_after_eos(chunk, frames_after_eos),
                                 estimate_max_gen_len(len(toks), self.cfg.frame_rate)))
            return jobs
        if isinstance(text, (list, tuple)) and text and isinstance(text[0], (list, tuple, torch.Tensor)):
            # extension: a long text already split into chunks of token ids (one stream per chunk, as the reference
            # generates sentence chunks one after the other from copies of the voice state, tts_model.py:618-631)
            jobs = []
            for chunk in text:
                jobs.extend(self._jobs_for_text(request, chunk, max_tokens, frames_after_eos))
            return jobs
        toks = [int(t) for t in torch.as_tensor(text).view(-1).tolist()]  # extension: token ids of one chunk
        if not toks:
            raise ValueError("Text prompt cannot be empty")
        fae = frames_after_eos if frames_after_eos is not None else (self.model_recommended_frames_after_eos or 3)
        return [_Job(request, toks, fae, estimate_max_gen_len(len(toks), self.cfg.frame_rate))]

    def tokenizer_or_raise(self) -> Tokenizer:
        if self.tokenizer is None:
            raise ValueError("this model was built without a tokenizer; pass token ids instead of text")
        return self.tokenizer

    def _buffers(self, n: int, F: int, rot: int = 0):
        """Pinned host buffers for one `generate_frames` call of n streams x F frames: views into one allocation per
        rotation index, grown when a larger call comes.  (Keyed by (n, F) the ragged tail of a batch - a different n
        almost every block - pinned fresh memory 68 times in the cfg5 run, 1.0 of its 2.8 s.)"""
        fs = self.cfg.frame_samples
        need_pcm, need_eos = n * F * fs, n * F
        pool = self._outs.get(rot)
        if pool is None or pool[0].numel() < need_pcm or pool[1].numel() < need_eos:
            cap_n = max(n, self.engine.max_streams)
            cap_f = max(F, _BLOCK_MAX)
            pool = (torch.empty(cap_n * cap_f * fs, dtype=torch.float32, pin_memory=True),
                    torch.empty(cap_n * cap_f, dtype=torch.float32, pin_memory=True))
            self._outs[rot] = pool
        return pool[0][:need_pcm].view(n, F * fs), None, pool[1][:need_eos].view(F, n)

    # ------------------------------------------------------------------ voice state
    @torch.no_grad()
    def get_state_for_audio_prompt(self, audio_conditioning, truncate: bool = False) -> dict:
        """tts_model.py:788-905.  Accepts a .safetensors model state, a waveform tensor
        [channels, samples] / audio file (encoded on the device by the engine's Mimi encoder, ptts_encode_audio =
        _encode_audio tts_model.py:379-388), or - extension - an already encoded conditioning tensor
        [P, d_model] / [1, P, d_model]."""
        if isinstance(audio_conditioning, str) and not Path(audio_conditioning).exists():
            audio_conditioning = self._resolve_voice_reference(audio_conditioning)   # predefined name / URL
        if isinstance(audio_conditioning, (str, Path)) and str(audio_conditioning).endswith(".safetensors"):
            return import_model_state(audio_conditioning)
        if isinstance(audio_conditioning, (str, Path)):
            if not self.has_voice_cloning:
                raise ValueError(VOICE_CLONING_UNSUPPORTED)
            path = Path(audio_conditioning)
            if not path.exists():
                raise FileNotFoundError(str(path))
            audio_conditioning = self._read_audio(path, truncate)
        if not isinstance(audio_conditioning, torch.Tensor):
            raise ValueError(f"unsupported audio conditioning of type {type(audio_conditioning).__name__}")
        t = audio_conditioning
        if t.numel() == 0:
            raise ValueError("audio conditioning is empty")
        if t.shape[-1] == self.cfg.d_model and t.dim() in (2, 3):
            prompt = t.reshape(-1, self.cfg.d_model).float().cpu()
        else:
            if not self.has_voice_cloning:
                raise ValueError(VOICE_CLONING_UNSUPPORTED)
            if self.audio_encoder is not None:
                prompt = self.audio_encoder(t.unsqueeze(0)).reshape(-1, self.cfg.d_model).float().cpu()
            else:
                if t.dim() == 2 and t.shape[0] != 1:
                    raise ValueError("audio conditioning must be mono [1, samples] (convert_audio, tts_model.py:885-887)")
                prompt = self.engine.encode_audio(t)
        if self.cfg.insert_bos_before_voice and self._bos_before_voice is not None:
            prompt = torch.cat([self._bos_before_voice, prompt], dim=0)
        slot = self.engine.open_stream()
        keep = False
        try:
            self.engine.prefill([slot], [prompt])
            self._draw_noise(1)  # the reference evaluates (and discards) a flow-head sample here: same RNG use
            state = VoiceState(self.engine.export_state(slot))
            keep = self._make_room_for_voice()
            if keep:
                self._resident[slot] = True
                state.resident = (id(self.engine), slot)
            return state
        finally:
            if not keep:
                self.engine.close_stream(slot)

    def _resolve_voice_reference(self, name: str):
        """Predefined voice names and http(s):// / hf:// URLs (tts_model.py:846-876) are resolved by the reference
        package's own resolver / downloader - host-side pieces that stay as they are; returns a local path."""
        is_url = name.startswith(("http://", "https://", "hf://"))
        try:
            from pocket_tts.utils.utils import _ORIGINS_OF_PREDEFINED_VOICES, download_if_necessary, get_predefined_voice
        except ImportError as e:
            if is_url or "/" not in name:
                raise ValueError(f"{name!r} is a predefined voice name or a URL: those are resolved by the reference package "
                                 "(pocket_tts.utils.utils), which is not importable here; pass a local .wav / .safetensors "
                                 "path or a conditioning tensor") from e
            raise FileNotFoundError(name) from e
        if name in _ORIGINS_OF_PREDEFINED_VOICES:
            origin = getattr(self, "origin", None)
            if origin is None:
                raise ValueError("Cannot use predefined voices when the model is not loaded from a config associated with "
                                 "a language (tts_model.py:858-863)")
            return download_if_necessary(get_predefined_voice(language=Path(origin).stem, name=name))
        if is_url:
            return download_if_necessary(name)
        raise FileNotFoundError(name)

    def _make_room_for_voice(self) -> bool:
        """Evict the oldest resident voice prefixes that no live stream reads until one more fits."""
        from . import _native as N

        for slot in list(self._resident):
            if len(self._resident) < self.max_resident_voices:
                break
            try:
                self.engine.close_stream(slot)
                del self._resident[slot]
            except N.NativeError:      # still the prefix of live streams
                continue
        return len(self._resident) < self.max_resident_voices

    def _resident_slot(self, model_state) -> int | None:
        r = getattr(model_state, "resident", None)
        if r is not None and r[0] == id(self.engine) and r[1] in self._resident:
            return r[1]
        return None

    def _read_audio(self, path: Path, truncate: bool) -> torch.Tensor:
        """data/audio.py:23-52 for PCM WAV + resampling to 24 kHz mono (host, scipy)."""
        import wave

        import numpy as np

        with wave.open(str(path), "rb") as w:
            sr, ch, width, n = w.getframerate(), w.getnchannels(), w.getsampwidth(), w.getnframes()
            raw = w.readframes(n)
        if width != 2:
            raise ValueError("only 16-bit PCM WAV prompts are read here")
        audio = torch.from_numpy(np.frombuffer(raw, dtype=np.int16).astype("float32") / 32768.0).reshape(-1, ch).t()
        if truncate:
            audio = audio[..., : int(30 * sr)]
        audio = audio.mean(dim=0, keepdim=True)
        if sr != self.sample_rate:
            from scipy.signal import resample_poly

            g = math.gcd(sr, self.sample_rate)
            audio = torch.from_numpy(resample_poly(audio.numpy(), self.sample_rate // g, sr // g, axis=-1).astype("float32"))
        return audio

    # ------------------------------------------------------------------ generation
    @torch.no_grad()
    def generate_audio(self, model_state: dict, text_to_generate, max_tokens: int = MAX_TOKEN_PER_CHUNK,
                       frames_after_eos: int | None = None, copy_state: bool = True) -> torch.Tensor:
        """tts_model.py:477-543: 1-D fp32 tensor [n_frames * 1920]."""
        chunks = list(self.generate_audio_stream(model_state, text_to_generate, max_tokens, frames_after_eos,
                                                 copy_state))
        return torch.cat(chunks, dim=0) if chunks else torch.zeros(0)

    @torch.no_grad()
    def generate_audio_stream(self, model_state: dict, text_to_generate, max_tokens: int = MAX_TOKEN_PER_CHUNK,
                              frames_after_eos: int | None = None, copy_state: bool = True) -> Iterator[torch.Tensor]:
        """tts_model.py:545-632: yields [1920] fp32 chunks as frames are decoded.  One stream; the RNG
        is consumed exactly as the reference consumes it (one [1, ldim] draw per FlowLM call, prefill
        included), so a seeded run reproduces the reference's noise."""
        jobs = self._jobs_for_text(0, text_to_generate, max_tokens, frames_after_eos)
        eng = self.engine
        resident = self._resident_slot(model_state) if copy_state else None
        base = resident if resident is not None else eng.open_stream()
        work = base
        try:
            if resident is None:
                eng.import_state(base, model_state)
            elif not copy_state and isinstance(model_state, VoiceState):
                model_state.resident = None      # about to be mutated in place
            if copy_state:
                work = eng.open_stream()
            for job in jobs:
                if copy_state:
                    # copy.deepcopy(model_state), tts_model.py:637-638 - without the copy: the chunk's stream reads
                    # the voice prefix from the slot that holds it and appends behind it
                    eng.fork_shared(base, work)
                eng.reset_decoder(work)          # fresh Mimi state + BOS, tts_model.py:444,748-753
                job.slot = work
                start = eng.offset(work)
                self._check_capacity(start, job)
                eng.prefill_tokens([work], [list(job.tokens)])     # embedding gather on the device
                self._draw_noise(1)              # sample of the prefill call, discarded (tts_model.py:723-725)
                yield from self._run_single(job, keep_state=not copy_state)
                if not copy_state:
                    steps = job.done if job.limit is None or job.limit >= job.max_gen_len else job.limit + 1
                    self._set_offset(work, start + len(job.tokens) + min(steps, job.max_gen_len))
            if not copy_state:
                self._write_back(model_state, eng.export_state(work))
                if isinstance(model_state, VoiceState):
                    model_state.resident = None
        finally:
            if work != base:
                eng.close_stream(work)
            if resident is None:
                eng.close_stream(base)

    def _set_offset(self, slot: int, offset: int) -> None:
        from . import _native as N

        N.check(self.engine.lib.ptts_stream_set_offset(self.engine._h, slot, offset))

    @staticmethod
    def _write_back(model_state: dict, new_state: dict) -> None:
        for mod, st in new_state.items():
            tgt = model_state.setdefault(mod, {})
            tgt["cache"] = st["cache"]
            tgt["offset"] = st["offset"]

    def _check_capacity(self, start: int, job: _Job) -> None:
        need = start + len(job.tokens) + job.max_gen_len
        if need > self.engine.kv_capacity:
            raise ValueError(f"stream needs {need} FlowLM positions (state {start} + text {len(job.tokens)} + "
                             f"{job.max_gen_len} frames) but the engine was built with kv_capacity={self.engine.kv_capacity}")

    def _update_eos(self, job: _Job, logits: torch.Tensor, first_frame: int) -> None:
        """EOS bookkeeping of tts_model.py:761-764 on the logits of frames first_frame.."""
        if job.eos_step is None:
            hit = (logits > self.eos_threshold).nonzero()
            if hit.numel():
                job.eos_step = first_frame + int(hit[0])
                job.limit = min(job.max_gen_len, job.eos_step + job.frames_after_eos)

    def _run_single(self, job: _Job, keep_state: bool = False) -> Iterator[torch.Tensor]:
        eng, fs = self.engine, self.cfg.frame_samples
        block_i = 0
        while job.done < job.target():
            F = _BLOCK_SCHEDULE[block_i] if block_i < len(_BLOCK_SCHEDULE) else _BLOCK_MAX
            F = min(F, job.max_gen_len - job.done)
            block_i += 1
            noise, rng = [], []
            for _ in range(F):
                noise.append(self._draw_noise(1))
                rng.append(torch.get_rng_state())
            pcm, _, eos = eng.generate_frames([job.slot], torch.stack(noise), out=self._buffers(1, F),
                                              want_latents=False)
            self._update_eos(job, eos[:, 0], job.done)
            upto = min(job.done + F, job.target())
            for f in range(job.done, upto):
                yield pcm[0, (f - job.done) * fs:(f - job.done + 1) * fs].clone()
            if job.limit is not None and job.limit < job.max_gen_len and job.limit < job.done + F:
                # the reference ran frame `limit` (one more draw) and then broke out of its loop
                torch.set_rng_state(rng[job.limit - job.done])
            elif job.limit is not None and job.limit < job.max_gen_len and job.limit == job.done + F:
                # the block ended exactly at the limit: the reference still runs FlowLM for frame `limit` (one more
                # draw, K/V of that position appended, tts_model.py:756-765) before it breaks.  With
                # copy_state=False that position is part of the state handed back, so it is really computed.
                extra = self._draw_noise(1)
                if keep_state:
                    eng.step([job.slot], extra, lsd_steps=self.lsd_decode_steps, decode_audio=False)
            job.done += F
        if job.limit is None or job.limit >= job.max_gen_len:
            logger.warning("Maximum generation length reached without EOS, this very often indicates an error.")

    @torch.no_grad()
    def generate_audio_batch(self, model_states: Sequence[dict], texts: Sequence, max_tokens: int = MAX_TOKEN_PER_CHUNK,
                             frames_after_eos: int | None = None, noise: Sequence[torch.Tensor] | None = None,
                             block: int = _BLOCK_MAX, frame_limits: Sequence[int] | None = None) -> list:
        """Extension: many independent utterances at once.  Every text chunk of every
        request is one stream; up to `max_streams` streams step together and a finished stream's slot is
        handed to the next waiting chunk after each block of frames (ragged completion).  The caller's
        states are never mutated.  Requests that share a model_state object share one imported voice
        slot.  `noise[i]` ([frames, ldim], single-chunk texts only) injects the draws of request i.
        `frame_limits[i]` caps the frames of every chunk of request i (synthetic workloads: random-init weights never
        raise the EOS logit, the cap stands in for it and makes completion ragged).  Per-chunk timings of the call
        (admission -> first PCM block on the host) are left in `self.last_batch_stats`.
        Returns one 1-D PCM tensor per request."""
        import time as _time

        if len(model_states) != len(texts):
            raise ValueError("model_states and texts differ in length")
        eng, fs = self.engine, self.cfg.frame_samples
        waiting: list = []
        per_request: list = []
        for r, text in enumerate(texts):
            jobs = self._jobs_for_text(r, text, max_tokens, frames_after_eos)
            if noise is not None:
                if len(jobs) != 1:
                    raise ValueError("injected noise needs single-chunk texts")
                if noise[r].shape[0] < jobs[0].max_gen_len:
                    raise ValueError(f"noise[{r}] holds {noise[r].shape[0]} frames, {jobs[0].max_gen_len} may be needed")
                jobs[0].noise = noise[r].float()
            if frame_limits is not None:
                for j in jobs:
                    j.max_gen_len = max(1, min(j.max_gen_len, int(frame_limits[r])))
            per_request.append(jobs)
            waiting.extend(jobs)
        voices: dict = {}
        temp_voices: list = []
        active: list = []
        admitted: list = []
        t_engine = t_prefill = 0.0
        t_call = _time.perf_counter()
        from concurrent.futures import ThreadPoolExecutor
        copier = ThreadPoolExecutor(max_workers=1)
        pending = [None, None, None]
        n_blocks = 0
        try:
            for st in model_states:          # one resident slot per distinct voice state
                if id(st) not in voices:
                    v = self._resident_slot(st)
                    if v is None:
                        v = eng.open_stream()
                        temp_voices.append(v)
                        voices[id(st)] = v
                        eng.import_state(v, st)
                    else:
                        voices[id(st)] = v
            capacity = eng.max_streams - len(temp_voices) - len(self._resident)
            if capacity < 1:
                raise ValueError("no stream slot left for generation: raise max_streams")
            while waiting or active:
                while waiting and len(active) + len(admitted) < capacity:
                    job = waiting[0]
                    src = voices[id(model_states[job.request])]
                    self._check_capacity(eng.offset(src), job)
                    job.slot = eng.open_stream()
                    job.t_admit = _time.perf_counter()
                    admitted.append(waiting.pop(0))      # from here on `finally` closes the slot
                    eng.fork_shared(src, job.slot)       # every chunk of a voice reads one resident prefix
                    eng.reset_decoder(job.slot)
                while admitted:      # text prefill of the newcomers, in groups that fit the engine's prefill buffers
                    grp, rows = [], 0
                    while admitted and (not grp or rows + len(admitted[0].tokens) <= eng.max_prefill_rows):
                        rows += len(admitted[0].tokens)
                        grp.append(admitted.pop(0))
                    try:
                        _t = _time.perf_counter()
                        eng.prefill_tokens([j.slot for j in grp], [list(j.tokens) for j in grp])
                        t_prefill += _time.perf_counter() - _t
                    finally:
                        active.extend(grp)   # `finally` of the call closes their slots on error
                F = max(1, min(block, min(j.max_gen_len - j.done for j in active)))
                n = len(active)
                if any(j.noise is not None for j in active):
                    nz = torch.empty(F, n, self.cfg.ldim)
                    for i, j in enumerate(active):
                        nz[:, i] = j.noise[j.done:j.done + F] if j.noise is not None else self._draw_noise(F)
                else:
                    nz = self._draw_noise(F * n).view(F, n, self.cfg.ldim)   # one draw for the whole block
                # three sets of pinned result buffers in rotation: the copy of block k out of its buffer runs on a helper
                # thread (memcpy releases the GIL) beside the GPU's block k+1; a buffer is reused only after its copy
                rot = n_blocks % 3
                n_blocks += 1
                if pending[rot] is not None:
                    pending[rot].result()
                _t = _time.perf_counter()
                pcm, _, eos = eng.generate_frames([j.slot for j in active], nz, out=self._buffers(n, F, rot),
                                                  want_latents=False)
                t_engine += _time.perf_counter() - _t
                eos = eos.clone()
                pcm = pending[rot] = copier.submit(pcm.clone)     # jobs keep (future, row, samples) and resolve it at the end
                hit = eos > self.eos_threshold         # [F, n]; EOS bookkeeping of tts_model.py:761-764 for the block
                any_hit, first_hit = hit.any(dim=0).tolist(), hit.to(torch.int8).argmax(dim=0).tolist()
                still = []
                for i, j in enumerate(active):
                    if j.eos_step is None and any_hit[i]:
                        j.eos_step = j.done + first_hit[i]
                        j.limit = min(j.max_gen_len, j.eos_step + j.frames_after_eos)
                    upto = min(j.done + F, j.target())
                    if upto > j.done:
                        j.pcm.append((pcm, i, (upto - j.done) * fs))
                        if j.t_first is None:
                            j.t_first = _time.perf_counter()
                    j.done += F
                    if j.done >= j.target():
                        eng.close_stream(j.slot)
                        j.slot = -1
                    else:
                        still.append(j)
                active = still
        finally:
            copier.shutdown(wait=True)
            for j in active + admitted:
                if j.slot >= 0:
                    eng.close_stream(j.slot)
            for v in temp_voices:
                eng.close_stream(v)
        for jobs in per_request:             # resolve the block copies into per-job views
            for j in jobs:
                j.pcm = [fut.result()[row, :ns] for fut, row, ns in j.pcm]
        lat = sorted((j.t_first - j.t_admit) * 1e3 for jobs in per_request for j in jobs if j.t_first is not None)
        self.last_batch_stats = {"chunks": len(lat), "first_block_ms": lat, "engine_s": t_engine, "prefill_s": t_prefill,
                                 "total_s": _time.perf_counter() - t_call,
                                 "frames": sum(min(j.done, j.target()) for jobs in per_request for j in jobs)}
        def join(jobs):
            return torch.cat([p for j in jobs for p in j.pcm]) if any(j.pcm for j in jobs) else torch.zeros(0)

        if len(per_request) < 8:
            return [join(jobs) for jobs in per_request]
        with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:   # memcpy-bound, releases the GIL
            return list(pool.map(join, per_request))

    def close(self) -> None:
        self.engine.close()
