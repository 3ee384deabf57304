// extern "C" boundary (include/pocket_tts_b200.h).  No C++ exception leaves this file.
#include <new>

#include "engine.h"

using ptts::Engine;

struct ptts_engine {
  Engine* impl;
};

#define API_BEGIN try {
#define API_END                                              \
  }                                                          \
  catch (const std::bad_alloc&) {                            \
    ptts::set_error("host allocation failed");               \
    return PTTS_EINVAL;                                      \
  }                                                          \
  catch (const std::exception& ex) {                         \
    ptts::set_error(std::string("exception: ") + ex.what()); \
    return PTTS_EINVAL;                                      \
  }                                                          \
  catch (...) {                                              \
    ptts::set_error("unknown exception");                    \
    return PTTS_EINVAL;                                      \
  }
// Every entry point runs with the engine's CUDA device current and restores the caller's device on the way out
// (an engine on device 1 called from a fresh thread, or after torch switched devices, must not land on device 0).
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); }
    if (prev != dev) cudaSetDevice(dev); else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define NEED(e)                                 \
  if (!(e) || !(e)->impl) {                     \
    ptts::set_error("null engine handle");      \
    return PTTS_EINVAL;                         \
  }                                             \
  DeviceGuard _device_guard((e)->impl->cfg.device);

extern "C" {

const char* ptts_last_error(void) { return ptts::get_error(); }
int ptts_version(void) { return 100; }

int ptts_engine_create(const ptts_config* cfg, ptts_engine** out) {
  API_BEGIN
  if (!cfg || !out) { ptts::set_error("null argument"); return PTTS_EINVAL; }
  *out = nullptr;
  DeviceGuard guard(cfg->device);
  Engine* e = new Engine(*cfg);
  const int rc = e->init();
  if (rc != PTTS_OK) { delete e; return rc; }
  *out = new ptts_engine{e};
  return PTTS_OK;
  API_END
}

int ptts_engine_destroy(ptts_engine* e) {
  API_BEGIN
  if (!e) return PTTS_OK;
  {
    DeviceGuard guard(e->impl ? e->impl->cfg.device : 0);
    delete e->impl;
  }
  delete e;
  return PTTS_OK;
  API_END
}

int ptts_load_tensor(ptts_engine* e, const char* name, const float* host_data, const int64_t* shape, int32_t ndim) {
  API_BEGIN NEED(e) return e->impl->load_tensor(name, host_data, shape, ndim); API_END
}
int ptts_finalize(ptts_engine* e) { API_BEGIN NEED(e) return e->impl->finalize(); API_END }
int ptts_stream_open(ptts_engine* e, int32_t* slot_out) {
  API_BEGIN NEED(e)
  if (!slot_out) { ptts::set_error("null slot_out"); return PTTS_EINVAL; }
  int s = -1;
  const int rc = e->impl->stream_open(&s);
  *slot_out = s;
  return rc;
  API_END
}
int ptts_stream_close(ptts_engine* e, int32_t slot) { API_BEGIN NEED(e) return e->impl->stream_close(slot); API_END }
int ptts_stream_reset_decoder(ptts_engine* e, int32_t slot) {
  API_BEGIN NEED(e) return e->impl->stream_reset_decoder(slot); API_END
}
int ptts_stream_fork(ptts_engine* e, int32_t src, int32_t dst) {
  API_BEGIN NEED(e) return e->impl->stream_fork(src, dst); API_END
}
int ptts_stream_fork_shared(ptts_engine* e, int32_t src_slot, int32_t dst_slot) {
  API_BEGIN NEED(e) return e->impl->stream_fork_shared(src_slot, dst_slot); API_END
}
int ptts_stream_import_kv(ptts_engine* e, int32_t slot, int32_t layer, const float* cache_any, int32_t n_positions) {
  API_BEGIN NEED(e)
  if (!cache_any && n_positions > 0) { ptts::set_error("null cache"); return PTTS_EINVAL; }
  return e->impl->import_kv(slot, layer, cache_any, n_positions);
  API_END
}
int ptts_stream_export_kv(ptts_engine* e, int32_t slot, int32_t layer, float* cache_any, int32_t n_positions) {
  API_BEGIN NEED(e)
  if (!cache_any && n_positions > 0) { ptts::set_error("null cache"); return PTTS_EINVAL; }
  return e->impl->export_kv(slot, layer, cache_any, n_positions);
  API_END
}
int ptts_stream_set_offset(ptts_engine* e, int32_t slot, int32_t offset) {
  API_BEGIN NEED(e) return e->impl->set_offset(slot, offset); API_END
}
int ptts_stream_get_offset(ptts_engine* e, int32_t slot, int32_t* offset_out) {
  API_BEGIN NEED(e)
  if (!offset_out) { ptts::set_error("null offset_out"); return PTTS_EINVAL; }
  int o = 0;
  const int rc = e->impl->get_offset(slot, &o);
  *offset_out = o;
  return rc;
  API_END
}
int ptts_stream_set_prev_latent(ptts_engine* e, int32_t slot, const float* latent_any) {
  API_BEGIN NEED(e)
  if (!latent_any) { ptts::set_error("null latent"); return PTTS_EINVAL; }
  return e->impl->set_prev_latent(slot, latent_any);
  API_END
}
int ptts_prefill(ptts_engine* e, int32_t n, const int32_t* slots, const int32_t* lens, const float* embeds_any) {
  API_BEGIN NEED(e) return e->impl->prefill(n, slots, lens, embeds_any); API_END
}
int ptts_prefill_tokens(ptts_engine* e, int32_t n, const int32_t* slots, const int32_t* lens, const int32_t* tokens) {
  API_BEGIN NEED(e) return e->impl->prefill(n, slots, lens, nullptr, tokens); API_END
}
int ptts_step(ptts_engine* e, int32_t n, const int32_t* slots, const float* noise_any, int32_t lsd_steps,
              int32_t decode_audio, float* out_latents_any, float* out_eos_logit_any, float* out_pcm_any) {
  API_BEGIN NEED(e)
  return e->impl->step(n, slots, noise_any, lsd_steps, decode_audio, out_latents_any, out_eos_logit_any, out_pcm_any);
  API_END
}
int ptts_mimi_decode(ptts_engine* e, int32_t n, const int32_t* slots, const float* latents_any, float* out_pcm_any) {
  API_BEGIN NEED(e) return e->impl->mimi_decode(n, slots, latents_any, out_pcm_any); API_END
}
int ptts_mimi_decode_frames(ptts_engine* e, int32_t n, const int32_t* slots, int32_t n_frames, const float* latents_any,
                            float* out_pcm_any) {
  API_BEGIN NEED(e) return e->impl->mimi_decode_frames(n, slots, n_frames, latents_any, out_pcm_any); API_END
}
int ptts_run_frames(ptts_engine* e, int32_t n, const int32_t* slots, int32_t n_frames, const float* noise_dev,
                    int32_t lsd_steps, float* out_pcm_dev, float* out_latents_dev) {
  API_BEGIN NEED(e)
  return e->impl->run_frames(n, slots, n_frames, noise_dev, lsd_steps, out_pcm_dev, out_latents_dev);
  API_END
}

int ptts_generate_frames(ptts_engine* e, int32_t n, const int32_t* slots, int32_t n_frames, const float* noise_any,
                         int32_t lsd_steps, float* out_pcm_any, float* out_latents_any, float* out_eos_logit_any) {
  API_BEGIN NEED(e)
  return e->impl->run_frames(n, slots, n_frames, noise_any, lsd_steps, out_pcm_any, out_latents_any, out_eos_logit_any,
                             true);
  API_END
}
int ptts_pcm_to_int16(ptts_engine* e, const float* pcm_any, int16_t* out_any, int64_t n) {
  API_BEGIN NEED(e) return e->impl->pcm_to_int16(pcm_any, reinterpret_cast<short*>(out_any), (long long)n); API_END
}
int ptts_pcm_to_int16_static(ptts_engine* e, const float* pcm_any, int16_t* out_any, int64_t n) {
  API_BEGIN NEED(e)
  return ptts::pcm_to_int16_static(e->impl->stream, pcm_any, reinterpret_cast<short*>(out_any), (long long)n, &e->impl->launches);
  API_END
}
int ptts_encode_audio(ptts_engine* e, const float* audio_any, int64_t n_samples, float* out_cond_any, int32_t max_frames,
                      int32_t* n_frames_out) {
  API_BEGIN NEED(e)
  int nf = 0;
  const int rc = e->impl->encode_audio(audio_any, (long long)n_samples, out_cond_any, max_frames, &nf);
  if (n_frames_out) *n_frames_out = nf;
  return rc;
  API_END
}
int ptts_synchronize(ptts_engine* e) {
  API_BEGIN NEED(e)
  cudaError_t ce = cudaStreamSynchronize(e->impl->stream);
  if (ce != cudaSuccess) { ptts::set_error_cuda(ce, "cudaStreamSynchronize", __FILE__, __LINE__); return PTTS_ECUDA; }
  return PTTS_OK;
  API_END
}
int64_t ptts_launch_count(ptts_engine* e) { return (e && e->impl) ? e->impl->launches : -1; }
void* ptts_stream_handle(ptts_engine* e) { return (e && e->impl) ? (void*)e->impl->stream : nullptr; }
int ptts_set_graphs(ptts_engine* e, int32_t enabled) {
  API_BEGIN NEED(e) e->impl->use_graphs = enabled != 0; return PTTS_OK; API_END
}
int ptts_debug_read(ptts_engine* e, const char* what, float* out_any, int64_t max_elems, int64_t* n_out) {
  API_BEGIN NEED(e) return e->impl->debug_read(what, out_any, max_elems, n_out); API_END
}

int ptts_debug_gemm(ptts_engine* e, int32_t backend, int32_t M, int32_t N, int32_t K, int32_t ld,
                    int32_t rows_per_batch, int32_t valid_rows, const float* a_host, int64_t a_elems,
                    const float* w_host, float* out_host) {
  API_BEGIN NEED(e)
  return e->impl->debug_gemm(backend, M, N, K, ld, rows_per_batch, valid_rows, a_host, a_elems, w_host, out_host);
  API_END
}
int ptts_profile_begin(ptts_engine* e) { API_BEGIN NEED(e) return e->impl->profile_begin(); API_END }
int ptts_debug_timeline(ptts_engine* e, char* json_out, int64_t capacity, int32_t max_records) {
  API_BEGIN NEED(e)
  if (!json_out || capacity < 2) { ptts::set_error("null buffer"); return PTTS_EINVAL; }
  return e->impl->timeline_json(json_out, capacity, max_records);
  API_END
}
int ptts_profile_end(ptts_engine* e, char* json_out, int64_t capacity) {
  API_BEGIN NEED(e) return e->impl->profile_end(json_out, capacity); API_END
}

}  // extern "C"
