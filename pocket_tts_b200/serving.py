"""Batched serving loop: concurrent `generate_audio_stream`-style requests coalesced into the
engine's frame step.

The reference's server handles one request at a time (main.py:121-181: a thread per request around
`generate_audio_stream`, one model, `torch.set_num_threads(1)`).  Here every text chunk of every request is a stream
slot; a scheduler thread per model (= per GPU) admits waiting chunks into free slots between blocks of frames, steps
all live streams together (`ptts_generate_frames`) and hands each request its PCM chunks through a queue, so callers
keep the reference's iterator interface while the GPU sees one batch.  With several models (one per GPU of the box)
new requests go to the GPU with the fewest live streams (`LeastLoadedPlacer`, SURVEY.md 8e: no collective).

    server = StreamServer([model])                     # TTSModel instances, one per GPU
    for chunk in server.generate_audio_stream(voice_state, "some text"):   # from any number of threads
        ...
    server.close()

`create_app(server, default_voice)` wraps it in the reference's HTTP surface (`/health`, `POST /tts` streaming a WAV).
"""
from __future__ import annotations

import queue
import threading
import time
from typing import Iterator, Sequence

import torch

from .sharding import LeastLoadedPlacer
from .tts_model import MAX_TOKEN_PER_CHUNK, TTSModel, _BLOCK_MAX

_END = object()


class _Request:
    def __init__(self, jobs, out: "queue.Queue", voice_state, noise):
        self.jobs, self.out, self.voice_state, self.noise = jobs, out, voice_state, noise
        self.next_emit = 0            # chunks are emitted in text order: job k streams only after jobs < k are done
        self.t_submit = time.perf_counter()
        self.t_first = None
        self.failed = False


class _ModelLoop(threading.Thread):
    """Scheduler thread of one model / GPU."""

    def __init__(self, model: TTSModel, block: int, on_release):
        super().__init__(daemon=True)
        self.m, self.block, self.on_release = model, block, on_release
        self.inbox: "queue.Queue" = queue.Queue()
        self.stop_flag = False
        self.error: BaseException | None = None
        self.steps = 0

    def submit(self, req: _Request) -> None:
        self.inbox.put(req)

    def run(self) -> None:
        m, eng, fs = self.m, self.m.engine, self.m.cfg.frame_samples
        waiting: list = []            # (request, job) in arrival order
        active: list = []
        temp_voices: dict = {}        # id(state) -> (slot, live jobs) for states without a resident slot
        try:
            with torch.no_grad():
                while not self.stop_flag:
                    # ---- take new requests (block only when idle)
                    try:
                        req = self.inbox.get(timeout=0.05) if not (waiting or active) else self.inbox.get_nowait()
                        while True:
                            if req is None:
                                self.stop_flag = True
                                break
                            waiting.extend((req, j) for j in req.jobs)
                            req = self.inbox.get_nowait()
                    except queue.Empty:
                        pass
                    if self.stop_flag:
                        break
                    if not (waiting or active):
                        continue
                    # ---- admit into free slots: fork the voice prefix (shared), fresh decoder, text prefill
                    capacity = eng.max_streams - len(m._resident) - len(temp_voices)
                    admitted = []
                    while waiting and len(active) + len(admitted) < capacity - 1:
                        req, job = waiting[0]
                        if req.failed:
                            waiting.pop(0)
                            self.on_release()
                            continue
                        try:   # a request that cannot be served fails alone (the reference raises in the caller's thread)
                            m._check_capacity(int(next(iter(req.voice_state.values()))["offset"].view(-1)[0]), job)
                        except (ValueError, KeyError, StopIteration) as ex:
                            req.failed = True
                            req.out.put(ValueError(str(ex)))
                            continue
                        src = m._resident_slot(req.voice_state)
                        if src is None:
                            key = id(req.voice_state)
                            if key not in temp_voices:
                                v = eng.open_stream()
                                eng.import_state(v, req.voice_state)
                                temp_voices[key] = [v, 0]
                            temp_voices[key][1] += 1
                            src = temp_voices[key][0]
                            job.temp_voice = key
                        job.slot = eng.open_stream()
                        eng.fork_shared(src, job.slot)
                        eng.reset_decoder(job.slot)
                        job.req = req
                        admitted.append(job)
                        waiting.pop(0)
                    while admitted:
                        grp, rows = [], 0
                        while admitted and (not grp or rows + len(admitted[0].tokens) <= eng.max_prefill_rows):
                            rows += len(admitted[0].tokens)
                            grp.append(admitted.pop(0))
                        eng.prefill_tokens([j.slot for j in grp], [list(j.tokens) for j in grp])
                        active.extend(grp)
                    if not active:
                        continue
                    # ---- one block of frames for every live stream
                    F = max(1, min(self.block, min(j.max_gen_len - j.done for j in active)))
                    n = len(active)
                    nz = torch.empty(F, n, m.cfg.ldim)
                    for i, j in enumerate(active):
                        nz[:, i] = j.noise[j.done:j.done + F] if j.noise is not None else m._draw_noise(F)
                    pcm, _, eos = eng.generate_frames([j.slot for j in active], nz, out=m._buffers(n, F), want_latents=False)
                    pcm = pcm.clone()
                    self.steps += F
                    still = []
                    for i, j in enumerate(active):
                        m._update_eos(j, eos[:, i], j.done)
                        upto = min(j.done + F, j.target())
                        for f in range(j.done, upto):
                            j.pcm.append(pcm[i, (f - j.done) * fs:(f - j.done + 1) * fs])
                        j.done += F
                        if j.done >= j.target():
                            eng.close_stream(j.slot)
                            j.slot = -1
                            j.finished = True
                            key = getattr(j, "temp_voice", None)
                            if key is not None:
                                temp_voices[key][1] -= 1
                                if temp_voices[key][1] == 0:
                                    eng.close_stream(temp_voices.pop(key)[0])
                            self.on_release()
                        else:
                            still.append(j)
                        self._emit(j.req)
                    active = still
        except BaseException as ex:     # surfaces in every waiting consumer, as the reference's worker errors do
            self.error = ex
            for req, _ in waiting:
                req.out.put(ex)
            for j in active:
                j.req.out.put(ex)
        finally:
            for j in active:
                if j.slot >= 0:
                    try:
                        eng.close_stream(j.slot)
                    except Exception:
                        pass
            for v, _ in temp_voices.values():
                try:
                    eng.close_stream(v)
                except Exception:
                    pass

    @staticmethod
    def _emit(req: _Request) -> None:
        """Hand out the chunks that are next in text order; close the request after its last job."""
        while req.next_emit < len(req.jobs):
            j = req.jobs[req.next_emit]
            while j.pcm:
                if req.t_first is None:
                    req.t_first = time.perf_counter()
                req.out.put(j.pcm.pop(0))
            if not getattr(j, "finished", False):
                return
            req.next_emit += 1
        req.out.put(_END)


class StreamServer:
    """Continuous batching of independent utterance streams over one model per GPU."""

    def __init__(self, models: Sequence[TTSModel], block: int = _BLOCK_MAX):
        self.models = list(models)
        cap = min(m.engine.max_streams for m in self.models)
        self._placer = LeastLoadedPlacer(world_size=len(self.models), capacity=1 << 30)
        self._lock = threading.Lock()
        self._loops = []
        for r, m in enumerate(self.models):
            loop = _ModelLoop(m, block, on_release=lambda r=r: self._release(r))
            loop.start()
            self._loops.append(loop)
        self.capacity_per_gpu = cap

    def _release(self, rank: int) -> None:
        with self._lock:
            self._placer.release(rank)

    def submit(self, model_state: dict, text, max_tokens: int = MAX_TOKEN_PER_CHUNK, frames_after_eos: int | None = None,
               noise: torch.Tensor | None = None) -> "queue.Queue":
        """Queue one utterance; returns the queue its [1920]-sample chunks arrive on (then the end marker)."""
        with self._lock:
            # a voice that is resident on one GPU stays there; otherwise the GPU with the fewest live streams
            ranks = [r for r, m in enumerate(self.models) if m._resident_slot(model_state) is not None]
            rank = ranks[0] if ranks else min(range(len(self.models)), key=lambda r: (self._placer.live[r], r))
            m = self.models[rank]
            jobs = m._jobs_for_text(0, text, max_tokens, frames_after_eos)
            if noise is not None:
                if len(jobs) != 1:
                    raise ValueError("injected noise needs single-chunk texts")
                if noise.shape[0] < jobs[0].max_gen_len:
                    raise ValueError(f"noise holds {noise.shape[0]} frames, {jobs[0].max_gen_len} may be needed")
                jobs[0].noise = noise.float()
            for _ in jobs:
                self._placer.live[rank] += 1
        if self._loops[rank].error is not None:
            raise RuntimeError("the scheduler of this GPU has stopped") from self._loops[rank].error
        req = _Request(jobs, queue.Queue(), model_state, noise)
        self._loops[rank].submit(req)
        return req.out

    def generate_audio_stream(self, model_state: dict, text, max_tokens: int = MAX_TOKEN_PER_CHUNK,
                              frames_after_eos: int | None = None, noise: torch.Tensor | None = None) -> Iterator[torch.Tensor]:
        """Same contract as TTSModel.generate_audio_stream (tts_model.py:545-632, copy_state=True): yields fp32 [1920]
        chunks in order; safe to call from many threads at once."""
        out = self.submit(model_state, text, max_tokens, frames_after_eos, noise)
        while True:
            item = out.get()
            if item is _END:
                return
            if isinstance(item, BaseException):
                raise item
            yield item

    def generate_audio(self, model_state: dict, text, **kw) -> torch.Tensor:
        chunks = list(self.generate_audio_stream(model_state, text, **kw))
        return torch.cat(chunks) if chunks else torch.zeros(0)

    def close(self) -> None:
        for loop in self._loops:
            loop.submit(None)
        for loop in self._loops:
            loop.join(timeout=30)


def create_app(server: StreamServer, default_voice: dict, voices: dict | None = None):
    """The reference's HTTP surface (main.py:60-181) over the batched loop: GET /health, POST /tts (form field `text`,
    optional `voice` naming an entry of `voices`) streaming a 16-bit mono WAV as chunks are decoded.  Many requests
    run at once; the reference serves them one after the other."""
    import struct

    from fastapi import FastAPI, Form, HTTPException
    from fastapi.responses import StreamingResponse

    app = FastAPI(title="pocket_tts_b200")
    model = server.models[0]

    @app.get("/health")
    def health():
        return {"status": "healthy"}

    @app.post("/tts")
    def tts(text: str = Form(...), voice: str | None = Form(None)):
        if not text.strip():
            raise HTTPException(status_code=400, detail="Text cannot be empty")
        if voice is not None and (voices is None or voice not in voices):
            raise HTTPException(status_code=400, detail=f"unknown voice {voice!r}")
        state = voices[voice] if voice is not None else default_voice

        def body():
            sr = model.sample_rate          # streaming WAV header: sizes unknown -> 0xFFFFFFFF (data/audio.py:55-75)
            yield (b"RIFF" + struct.pack("<I", 0xFFFFFFFF) + b"WAVEfmt " + struct.pack("<IHHIIHH", 16, 1, 1, sr, sr * 2, 2, 16) +
                   b"data" + struct.pack("<I", 0xFFFFFFFF))
            for chunk in server.generate_audio_stream(state, text):
                yield model.engine.pcm_to_int16(chunk).numpy().tobytes()   # data/audio.py:77-78 on the device

        return StreamingResponse(body(), media_type="audio/wav",
                                 headers={"Content-Disposition": "attachment; filename=generated_speech.wav"})

    return app
