/*
 * pocket_tts_b200 — C ABI of the B200-native per-frame generation engine.
 *
 * This header is the drop-in boundary for the hot path of kyutai-labs/pocket-tts
 * (SURVEY.md section 8b).  The reference has no FFI of its own: its boundary is the Python
 * API in pocket_tts/models/tts_model.py, so every entry point below cites the reference
 * Python function whose work it replaces.  Plain pointers and sizes only, no torch types;
 * every call returns 0 on success or a negative PTTS_E* code, with a thread-local message
 * available from ptts_last_error().  No C++ exception crosses this boundary.
 *
 * Pointers named *_any may be host (pageable or pinned) or device pointers; they are
 * copied with cudaMemcpyAsync(cudaMemcpyDefault) on the engine stream.  All floats are
 * IEEE fp32 on the ABI regardless of the engine's internal precision mode.
 */
#ifndef POCKET_TTS_B200_H
#define POCKET_TTS_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTTS_OK 0
#define PTTS_EINVAL (-1)   /* bad argument / unknown tensor name / shape mismatch  */
#define PTTS_ECUDA (-2)    /* CUDA runtime error (message holds cudaGetErrorString) */
#define PTTS_ESTATE (-3)   /* call order violated (e.g. step before finalize)       */
#define PTTS_ENOSLOT (-4)  /* no free stream slot / capacity exceeded               */

#define PTTS_PRECISION_FP32 0 /* fp32 storage, FFMA accumulate: parity mode (rel-L2 <= 1e-4) */
#define PTTS_PRECISION_BF16 1 /* bf16 weights/activations/KV, fp32 accumulate (tcgen05)     */

/* Architecture of the model, mirrors pocket_tts/utils/config.py:11-118 (the YAML schema)
 * for the fields the decode path uses (config/english.yaml:7-60). */
typedef struct ptts_config {
  int32_t n_layers;       /* flow_lm.transformer.num_layers: 6 or 24              */
  int32_t d_model;        /* 1024                                                  */
  int32_t n_heads;        /* 16                                                    */
  int32_t ffn;            /* d_model * hidden_scale = 4096                         */
  int32_t ldim;           /* mimi.quantizer.dimension = 32                         */
  int32_t flow_dim;       /* flow.dim = 512                                        */
  int32_t flow_depth;     /* flow.depth = 6                                        */
  int32_t mimi_d;         /* mimi.transformer.d_model = 512                        */
  int32_t mimi_heads;     /* 8                                                     */
  int32_t mimi_layers;    /* 2                                                     */
  int32_t mimi_ffn;       /* 2048                                                  */
  int32_t mimi_context;   /* 250                                                   */
  int32_t n_filters;      /* seanet.n_filters = 64                                 */
  int32_t ratios[3];      /* seanet.ratios = 6,5,4                                 */
  int32_t kernel;         /* seanet.kernel_size = 7                                */
  int32_t res_kernel;     /* seanet.residual_kernel_size = 3                       */
  int32_t last_kernel;    /* seanet.last_kernel_size = 3                           */
  int32_t upsample_stride;/* encoder_frame_rate / frame_rate = 16                  */
  /* engine sizing */
  int32_t max_streams;    /* stream slots resident on this GPU                     */
  int32_t kv_capacity;    /* FlowLM positions per slot (voice + text + frames)     */
  int32_t max_prefill_rows; /* max sum of rows in one ptts_prefill call            */
  int32_t max_lsd_steps;  /* rows accepted for the "flow.tcomb" table              */
  int32_t precision;      /* PTTS_PRECISION_*                                      */
  int32_t device;         /* CUDA ordinal                                          */
  int32_t kv_fp8;         /* 1: FlowLM KV cache stored as FP8 e4m3 with one fp32 scale per cached row (bf16 mode only):
                           * the low-precision mode of this engine, the analogue of the reference's quantize=True
                           * (quantization.py:60-88) on the term that bounds the step on a GPU - KV bytes - with its own
                           * tolerance; 0: cache in the activation type                                    */
} ptts_config;

typedef struct ptts_engine ptts_engine;

const char* ptts_last_error(void);
int ptts_version(void);

/* TTSModel._from_pydantic_config_with_weights (tts_model.py:130-230): build the model. */
int ptts_engine_create(const ptts_config* cfg, ptts_engine** out);
int ptts_engine_destroy(ptts_engine* e);

/* load_state_dict (tts_model.py:209-210): one call per state_dict entry, names exactly as in
 * TTSModel.state_dict() ("flow_lm.transformer.layers.0.self_attn.in_proj.weight", ...).
 * The Mimi encoder side (mimi.encoder.*, mimi.encoder_transformer.*, mimi.downsample.*,
 * flow_lm.speaker_proj_weight) is optional: loaded whole it enables ptts_encode_audio, absent (the
 * "without voice cloning" checkpoints, tts_model.py:871-872) that call returns PTTS_ESTATE.  Two derived,
 * input-independent tables are loaded the same way: "rope.freqs" [head_dim/2]
 * (rope.py:28-29) and "flow.tcomb" [lsd_steps, flow_dim] (mlp.py:205-209 evaluated at
 * s=i/n, t=(i+1)/n, flow_lm.py:33-37). */
int ptts_load_tensor(ptts_engine* e, const char* name, const float* host_data,
                     const int64_t* shape, int32_t ndim);
/* Checks that every required tensor arrived and freezes the weights. */
int ptts_finalize(ptts_engine* e);

/* init_states (stateful_module.py:7-16): a slot owns a FlowLM KV cache (offset 0) and a fresh
 * Mimi decoder state (tts_model.py:444). */
int ptts_stream_open(ptts_engine* e, int32_t* slot_out);
int ptts_stream_close(ptts_engine* e, int32_t slot);
/* Fresh Mimi state + BOS previous-latent for the slot; FlowLM KV is kept (new text chunk on a
 * forked voice state, tts_model.py:637-658). */
int ptts_stream_reset_decoder(ptts_engine* e, int32_t slot);
/* copy.deepcopy(model_state) (tts_model.py:637-638): dst gets src's FlowLM KV prefix+offset. */
int ptts_stream_fork(ptts_engine* e, int32_t src_slot, int32_t dst_slot);
/* The same copy.deepcopy(model_state) (tts_model.py:637-638) for many streams of ONE voice, without the copy: dst
 * reads positions [0, offset(src)) from the slot that owns them (one resident voice prefix per GPU, read by every
 * stream forked from it) and appends its own text / frame positions behind them.  The owner must stay open until every
 * dependent is closed (ptts_stream_close on it returns PTTS_ESTATE before that) and may itself keep appending.
 * A dependent that has not appended yet can be shared onward (the owner is shared instead); after it has,
 * ptts_stream_fork makes a deep copy.  Export / fork of a dependent read through the sharing; ptts_stream_import_kv
 * into a dependent returns PTTS_ESTATE.  (prefix sharing, zero-copy stream_fork.) */
int ptts_stream_fork_shared(ptts_engine* e, int32_t src_slot, int32_t dst_slot);
/* model_state import/export in the reference layout cache[2, 1, P, H, D] per layer
 * (transformer.py:46-57, tts_model.py:1047-1072). */
int ptts_stream_import_kv(ptts_engine* e, int32_t slot, int32_t layer, const float* cache_any,
                          int32_t n_positions);
int ptts_stream_export_kv(ptts_engine* e, int32_t slot, int32_t layer, float* cache_any,
                          int32_t n_positions);
int ptts_stream_set_offset(ptts_engine* e, int32_t slot, int32_t offset);
int ptts_stream_get_offset(ptts_engine* e, int32_t slot, int32_t* offset_out);
/* Teacher forcing for parity tests: overwrite the fed-back latent (tts_model.py:768). */
int ptts_stream_set_prev_latent(ptts_engine* e, int32_t slot, const float* latent_any);

/* _run_flow_lm_and_increment_step with text / audio conditioning (tts_model.py:317-346,
 * 722-725, 896-899): rows are already-embedded conditioning vectors [sum(lens), d_model],
 * stream i owning lens[i] consecutive rows; only the KV cache is kept (the reference discards
 * the sampled latent of a prefill call, tts_model.py:723-725). */
int ptts_prefill(ptts_engine* e, int32_t n, const int32_t* slots, const int32_t* lens,
                 const float* embeds_any);

/* Text prefill from token ids: the LUT conditioner's embedding gather (conditioners/text.py:74-76,
 * called from tts_model.py:718-725) runs on the device from the table loaded as
 * "flow_lm.conditioner.embed.weight", so a text chunk crosses the bus as 4 bytes per token instead
 * of 4 KB.  tokens [sum(lens)] are sentencepiece ids in [0, table rows); PTTS_EINVAL otherwise or
 * when the table was not loaded. */
int ptts_prefill_tokens(ptts_engine* e, int32_t n, const int32_t* slots, const int32_t* lens,
                        const int32_t* tokens);

/* One frame for n streams: FlowLM step + flow head (+ optional Mimi decode).
 * _autoregressive_generation body (tts_model.py:756-768) + _decode_audio_worker body
 * (tts_model.py:449-455).  noise_any [n, ldim] is the injected N(0, temp) draw
 * (flow_lm.py:131-137).  Outputs may be NULL.  out_eos_logit is out_eos(h) before the
 * threshold (flow_lm.py:129) so callers apply eos_threshold themselves. */
int ptts_step(ptts_engine* e, int32_t n, const int32_t* slots, const float* noise_any,
              int32_t lsd_steps, int32_t decode_audio, float* out_latents_any,
              float* out_eos_logit_any, float* out_pcm_any);

/* MimiModel.decode_from_latent on explicit latents (mimi.py:89-94; BASELINE config 4). */
int ptts_mimi_decode(ptts_engine* e, int32_t n, const int32_t* slots, const float* latents_any,
                     float* out_pcm_any);

/* The decoder thread's loop over a queue of latents (tts_model.py:434-474) for n streams at once: n_frames consecutive
 * latents per stream, latents_any [n_frames, n, ldim] -> out_pcm_any [n, n_frames*1920], decoded in passes of up to 4
 * frames x 16 tokens per stream (the pass form of ptts_generate_frames; arithmetically the same as n_frames calls of
 * ptts_mimi_decode).  BASELINE config 4 (Mimi decoder-only sweep) runs through this call. */
int ptts_mimi_decode_frames(ptts_engine* e, int32_t n, const int32_t* slots, int32_t n_frames,
                            const float* latents_any, float* out_pcm_any);

/* Device-resident multi-frame loop for throughput runs: n_frames frames back to back with
 * noise [n_frames, n, ldim] already on the device, PCM written to a device buffer
 * [n, n_frames*1920] (or NULL), no host round trip in between (CUDA-graph replay). */
int ptts_run_frames(ptts_engine* e, int32_t n, const int32_t* slots, int32_t n_frames,
                    const float* noise_dev, int32_t lsd_steps, float* out_pcm_dev,
                    float* out_latents_dev);

/* Streaming generation of n_frames frames for n streams with caller buffers that may live on the
 * host (pinned memory makes the copies asynchronous) or on the device: the body of
 * TTSModel._autoregressive_generation + the decoder thread (tts_model.py:651-658,745-779) as one
 * call.  Frames are issued in groups of 1, 1, 2, 4, 4, ...: each frame's FlowLM step + flow head is one
 * CUDA graph on the high-priority stream, the Mimi pass over a group's latents is one graph on a second
 * stream that runs beside the next group's FlowLM frames, and noise / latents / EOS logits / PCM copies
 * ride a third stream ordered by events.  Layouts:
 * noise [n_frames, n, ldim], out_pcm [n, n_frames*1920], out_latents [n_frames, n, ldim],
 * out_eos_logit [n_frames, n]; outputs may be NULL.  Returns after everything has landed. */
int ptts_generate_frames(ptts_engine* e, int32_t n, const int32_t* slots, int32_t n_frames,
                         const float* noise_any, int32_t lsd_steps, float* out_pcm_any,
                         float* out_latents_any, float* out_eos_logit_any);

/* Voice-prompt encoding on the device: TTSModel._encode_audio (tts_model.py:379-388) = MimiModel.encode_to_latent
 * (mimi.py:96-119: right padding to whole 1920-sample frames, SEANet encoder seanet.py:44-113, causal window-250
 * encoder transformer, ConvDownsample1d resample.py:7-29) followed by F.linear(latents, speaker_proj_weight).
 * audio_any: n_samples mono fp32 samples at 24 kHz (what convert_audio hands the reference, tts_model.py:885-887);
 * out_cond_any [n_frames, d_model] are the conditioning rows the caller prefills after bos_before_voice
 * (tts_model.py:893-899) with ptts_prefill.  n_frames = ceil(n_samples / 1920) is returned through n_frames_out;
 * out_cond_any == NULL only queries it.  Runs in fp32 in either precision mode.  At most
 * PTTS_ENCODE_MAX_SAMPLES samples (60 s; the reference's truncate=True keeps 30 s, tts_model.py:878-882). */
#define PTTS_ENCODE_MAX_SAMPLES 1440000
int ptts_encode_audio(ptts_engine* e, const float* audio_any, int64_t n_samples, float* out_cond_any,
                      int32_t max_frames, int32_t* n_frames_out);

/* PCM egress: fp32 -> 16-bit samples exactly as the reference's WAV writer does on the host
 * (data/audio.py:77-78: clamp(-1, 1) * 32767, truncated), on the device; n samples; both
 * pointers may be host or device memory.  Halves the bytes of the device-to-host PCM copy. */
int ptts_pcm_to_int16(ptts_engine* e, const float* pcm_any, int16_t* out_any, int64_t n);

/* The same conversion for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  NO counterpart in the reference.
 * ptts_pcm_to_int16 stages host buffers through a device buffer it grows with cudaMalloc; this entry point stages them
 * through a fixed file-scope device array instead (csrc/static_scratch.cu) and calls no cudaMalloc / cudaFree / other
 * allocating routine.  Limits and rules:
 *   - n <= PTTS_PCM_STATIC_MAX_SAMPLES (262144 samples = 10.9 s of 24 kHz audio); larger n returns PTTS_EINVAL.  Every
 *     GPU that loads the library holds the array (6 bytes per sample = 1.5 MiB) until the process exits.
 *   - The first call on a device resolves the array's device address (cudaGetSymbolAddress, once per device): make
 *     one call before entering graph capture.
 *   - All calls on a GPU share the array, which keeps the previous call's contents; a call only reads what it has
 *     just written, so nothing needs clearing.  The call returns after the engine stream has drained, so calls on
 *     one engine never overlap; calls that overlap on the same GPU (several engines / host threads) race on the
 *     array and must be serialised by the caller.
 * Results are bit-identical to ptts_pcm_to_int16. */
#define PTTS_PCM_STATIC_MAX_SAMPLES 262144
int ptts_pcm_to_int16_static(ptts_engine* e, const float* pcm_any, int16_t* out_any, int64_t n);

int ptts_synchronize(ptts_engine* e);
/* Number of engine kernels launched (or replayed through a graph) since creation. */
int64_t ptts_launch_count(ptts_engine* e);
/* CUDA stream the engine launches on (for event timing by the caller), as a void*. */
void* ptts_stream_handle(ptts_engine* e);
/* Enable / disable CUDA-graph replay of the frame step (default on). */
int ptts_set_graphs(ptts_engine* e, int32_t enabled);

/* Debug taps for per-kernel parity tests: copy a named internal activation buffer
 * (fp32-converted) to out_any; returns the element count through n_out. */
int ptts_debug_read(ptts_engine* e, const char* what, float* out_any, int64_t max_elems,
                    int64_t* n_out);

/* Test hook: D[M,N] = A W^T on bf16-rounded operands through one GEMM back-end (0 = SIMT FFMA,
 * 1 = tcgen05).  rows_per_batch <= 0: plain A [M, ld]; otherwise A is a stack of channels-last
 * stream buffers with overlapping rows (row t of a buffer = K elements starting at t*ld) and only
 * rows t < valid_rows of each buffer are written; other rows of out keep the 0xff fill. */
int ptts_debug_gemm(ptts_engine* e, int32_t backend, int32_t M, int32_t N, int32_t K, int32_t ld,
                    int32_t rows_per_batch, int32_t valid_rows, const float* a_host, int64_t a_elems,
                    const float* w_host, float* out_host);

/* Per-kernel timing with CUDA events on the engine stream (eager launches between begin and
 * end; graph replay is bypassed by the caller with ptts_set_graphs(e, 0)).  profile_end writes a
 * JSON object {"<kernel tag>": {"count": n, "total_ms": t}, ...} into json_out. */
int ptts_profile_begin(ptts_engine* e);
int ptts_profile_end(ptts_engine* e, char* json_out, int64_t capacity);
/* Developer: in-graph device timeline (engine created with PTTS_TIMELINE=1 in the environment): the last
 * max_records kernel records as a JSON array of [name, entry_us, pdl_wait_returned_us, exit_us]. */
int ptts_debug_timeline(ptts_engine* e, char* json_out, int64_t capacity, int32_t max_records);

#ifdef __cplusplus
}
#endif
#endif /* POCKET_TTS_B200_H */
