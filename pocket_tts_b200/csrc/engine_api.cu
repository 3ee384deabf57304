// Engine implementation, part 4: stream slots, prefill, frame step (eager / CUDA graph).
#include "engine.h"

#include <cmath>
#include <cstring>
#include <algorithm>
#include <cstdlib>
#include <functional>

namespace ptts {

#define NEED_READY() \
  if (!ready) PTTS_FAIL(PTTS_ESTATE, "engine not finalized")
#define CHECK_SLOT(s)                                                        \
  if ((s) < 0 || (s) >= cfg.max_streams || !slot_used[s])                    \
  PTTS_FAIL(PTTS_EINVAL, "invalid or closed stream slot %d", (int)(s))

int Engine::stream_open(int* slot) {
  NEED_READY();
  for (int s = 0; s < cfg.max_streams; ++s)
    if (!slot_used[s]) {
      slot_used[s] = 1;
      *slot = s;
      host_offsets[s] = 0;
      PTTS_CUDA_CHECK(cudaMemsetAsync(offsets + s, 0, sizeof(int), stream));
      return stream_reset_decoder(s);   // host_pfx_slot[s] is -1 here (stream_close / init)
    }
  PTTS_FAIL(PTTS_ENOSLOT, "all %d stream slots in use", cfg.max_streams);
}

int Engine::stream_close(int slot) {
  NEED_READY();
  CHECK_SLOT(slot);
  if (pfx_refs[slot] > 0)
    PTTS_FAIL(PTTS_ESTATE, "slot %d holds the shared voice prefix of %d live stream(s): close them first", slot, pfx_refs[slot]);
  int rc;
  if ((rc = unshare(slot)) != PTTS_OK) return rc;
  slot_used[slot] = 0;
  return PTTS_OK;
}

// drop the slot's reference to a shared prefix (its own rows [0, pfx_len) are NOT filled in: callers that keep the
// slot alive copy them first)
int Engine::unshare(int slot) {
  if (host_pfx_slot[slot] < 0) return PTTS_OK;
  --pfx_refs[host_pfx_slot[slot]];
  host_pfx_slot[slot] = -1;
  host_pfx_len[slot] = 0;
  PTTS_CUDA_CHECK(cudaMemsetAsync(d_pfx_slot + slot, 0xff, sizeof(int), stream));
  cur_slots.clear();   // batch_shared is recomputed by the next upload_slots
  return PTTS_OK;
}

// copy.deepcopy(model_state) without the copy (tts_model.py:637-638 for many streams of one voice): dst reads the
// positions [0, offset(src)) from the slot that owns them and appends its own positions behind; the owner must stay
// open (and may keep appending: dependents never read past their pfx_len) until every dependent is closed.
int Engine::stream_fork_shared(int src, int dst) {
  NEED_READY();
  CHECK_SLOT(src);
  CHECK_SLOT(dst);
  if (src == dst) PTTS_FAIL(PTTS_EINVAL, "fork_shared: source and destination are the same slot");
  if (pfx_refs[dst] > 0) PTTS_FAIL(PTTS_ESTATE, "fork_shared: slot %d is itself a shared prefix", dst);
  int owner = src;
  const int P = host_offsets[src];
  if (host_pfx_slot[src] >= 0) {
    // a dependent can be shared onward only while it has appended nothing of its own
    if (P != host_pfx_len[src]) PTTS_FAIL(PTTS_ESTATE, "fork_shared: slot %d already appended to its shared prefix; use ptts_stream_fork", src);
    owner = host_pfx_slot[src];
  }
  int rc;
  if ((rc = unshare(dst)) != PTTS_OK) return rc;
  host_pfx_slot[dst] = owner;
  host_pfx_len[dst] = P;
  ++pfx_refs[owner];
  host_offsets[dst] = P;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(offsets + dst, offsets + src, sizeof(int), cudaMemcpyDeviceToDevice, stream));
  int* hp = reinterpret_cast<int*>(h_pinned) + 3 * cfg.max_streams;
  hp[0] = owner; hp[1] = P;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_pfx_slot + dst, hp, sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_pfx_len + dst, hp + 1, sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));   // h_pinned is reused
  cur_slots.clear();
  return PTTS_OK;
}

// fresh Mimi state (tts_model.py:444) and BOS marker (tts_model.py:748-753)
int Engine::stream_reset_decoder(int slot) {
  NEED_READY();
  CHECK_SLOT(slot);
  host_mimi_offsets[slot] = 0;
  PTTS_CUDA_CHECK(cudaMemsetAsync(mimi_offsets + slot, 0, sizeof(int), stream));
  PTTS_CUDA_CHECK(cudaMemsetAsync(e_prev + (size_t)slot * cfg.mimi_d, 0, cfg.mimi_d * sizeof(float), stream));
  for (int k = 0; k < hist_tab.count; ++k) {
    const size_t bytes = (size_t)hist_tab.d[k].hist * hist_tab.d[k].C * act_size;
    PTTS_CUDA_CHECK(cudaMemsetAsync((char*)hist_tab.d[k].state + slot * bytes, 0, bytes, stream));
  }
  PTTS_CUDA_CHECK(launch_fill_f32(prev_latent + (size_t)slot * cfg.ldim, NAN, cfg.ldim, stream));
  return PTTS_OK;
}

int Engine::stream_fork(int src, int dst) {
  NEED_READY();
  CHECK_SLOT(src);
  CHECK_SLOT(dst);
  if (src == dst) return PTTS_OK;
  if (pfx_refs[dst] > 0) PTTS_FAIL(PTTS_ESTATE, "fork: slot %d holds a shared prefix in use", dst);
  const int P = host_offsets[src];
  int rc;
  if ((rc = unshare(dst)) != PTTS_OK) return rc;
  if (host_pfx_slot[src] >= 0) {   // deep copy of a sharing slot: the prefix rows from their owner, the rest from src
    const int pl = std::min(P, host_pfx_len[src]);
    PTTS_CUDA_CHECK(launch_kv_fork_range(kv_flow, cfg.n_layers, host_pfx_slot[src], dst, 0, pl, stream));
    PTTS_CUDA_CHECK(launch_kv_fork_range(kv_flow, cfg.n_layers, src, dst, pl, P, stream));
    ++launches;
  } else {
    PTTS_CUDA_CHECK(launch_kv_fork(kv_flow, cfg.n_layers, src, dst, P, stream));
  }
  ++launches;
  host_offsets[dst] = P;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(offsets + dst, offsets + src, sizeof(int), cudaMemcpyDeviceToDevice, stream));
  return PTTS_OK;
}

int Engine::import_kv(int slot, int layer, const float* cache_any, int P) {
  NEED_READY();
  CHECK_SLOT(slot);
  if (layer < 0 || layer >= cfg.n_layers || P < 0 || P > cfg.kv_capacity)
    PTTS_FAIL(PTTS_EINVAL, "import_kv: layer %d / positions %d out of range", layer, P);
  if (P == 0) return PTTS_OK;
  if (host_pfx_slot[slot] >= 0) PTTS_FAIL(PTTS_ESTATE, "import_kv: slot %d reads a shared prefix (fork it with ptts_stream_fork first)", slot);
  const size_t elems = 2ull * P * cfg.d_model;
  float* tmp = nullptr;
  PTTS_CUDA_CHECK(cudaMallocAsync((void**)&tmp, elems * sizeof(float), stream));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(tmp, cache_any, elems * sizeof(float), cudaMemcpyDefault, stream));
  PTTS_CUDA_CHECK(launch_kv_import(tmp, kv_flow, layer, slot, P, stream));
  PTTS_CUDA_CHECK(cudaFreeAsync(tmp, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

int Engine::export_kv(int slot, int layer, float* cache_any, int P) {
  NEED_READY();
  CHECK_SLOT(slot);
  if (layer < 0 || layer >= cfg.n_layers || P < 0 || P > cfg.kv_capacity)
    PTTS_FAIL(PTTS_EINVAL, "export_kv: layer %d / positions %d out of range", layer, P);
  if (P == 0) return PTTS_OK;
  const size_t elems = 2ull * P * cfg.d_model;
  float* tmp = nullptr;
  PTTS_CUDA_CHECK(cudaMallocAsync((void**)&tmp, elems * sizeof(float), stream));
  if (host_pfx_slot[slot] >= 0) {
    const int pl = std::min(P, host_pfx_len[slot]);
    PTTS_CUDA_CHECK(launch_kv_export_range(tmp, kv_flow, layer, host_pfx_slot[slot], P, 0, pl, stream));
    PTTS_CUDA_CHECK(launch_kv_export_range(tmp, kv_flow, layer, slot, P, pl, P, stream));
  } else {
    PTTS_CUDA_CHECK(launch_kv_export(tmp, kv_flow, layer, slot, P, stream));
  }
  PTTS_CUDA_CHECK(cudaMemcpyAsync(cache_any, tmp, elems * sizeof(float), cudaMemcpyDefault, stream));
  PTTS_CUDA_CHECK(cudaFreeAsync(tmp, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

int Engine::set_offset(int slot, int off) {
  NEED_READY();
  CHECK_SLOT(slot);
  if (off < 0 || off > cfg.kv_capacity) PTTS_FAIL(PTTS_EINVAL, "offset %d out of range", off);
  if (host_pfx_slot[slot] >= 0 && off < host_pfx_len[slot])
    PTTS_FAIL(PTTS_ESTATE, "offset %d lies inside the shared prefix of slot %d (%d positions)", off, slot, host_pfx_len[slot]);
  host_offsets[slot] = off;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(offsets + slot, &host_offsets[slot], sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

int Engine::get_offset(int slot, int* off) {
  NEED_READY();
  CHECK_SLOT(slot);
  int dev = -1;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(&dev, offsets + slot, sizeof(int), cudaMemcpyDeviceToHost, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  if (dev != host_offsets[slot]) PTTS_FAIL(PTTS_ESTATE, "offset mirror diverged (%d vs %d)", dev, host_offsets[slot]);
  *off = dev;
  return PTTS_OK;
}

int Engine::set_prev_latent(int slot, const float* latent_any) {
  NEED_READY();
  CHECK_SLOT(slot);
  PTTS_CUDA_CHECK(cudaMemcpyAsync(prev_latent + (size_t)slot * cfg.ldim, latent_any, cfg.ldim * sizeof(float),
                                  cudaMemcpyDefault, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

int Engine::check_slots(int n, const int* slots, bool need_decode_room) {
  if (n < 1 || n > cfg.max_streams || !slots) PTTS_FAIL(PTTS_EINVAL, "bad stream count %d", n);
  std::vector<char> seen(cfg.max_streams, 0);
  for (int i = 0; i < n; ++i) {
    CHECK_SLOT(slots[i]);
    if (seen[slots[i]]) PTTS_FAIL(PTTS_EINVAL, "slot %d listed twice", slots[i]);
    seen[slots[i]] = 1;
    if (need_decode_room && host_offsets[slots[i]] + 1 > cfg.kv_capacity)
      PTTS_FAIL(PTTS_ENOSLOT, "slot %d: KV capacity %d exhausted", slots[i], cfg.kv_capacity);
  }
  return PTTS_OK;
}

int Engine::upload_slots(int n, const int* slots) {
  if ((int)cur_slots.size() == n && std::memcmp(cur_slots.data(), slots, n * sizeof(int)) == 0) return PTTS_OK;
  int* hp = reinterpret_cast<int*>(h_pinned);
  const int S = cfg.max_streams;
  std::memcpy(hp, slots, n * sizeof(int));
  for (int i = 0; i < S; ++i) {
    hp[S + i] = i * cfg.upsample_stride;
    hp[2 * S + i] = cfg.upsample_stride;
  }
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_slot_ids, hp, n * sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_mimi_row_start, hp + S, S * sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_mimi_lens, hp + 2 * S, S * sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));  // h_pinned is reused
  cur_slots.assign(slots, slots + n);
  batch_shared = false;
  for (int i = 0; i < n; ++i) batch_shared = batch_shared || host_pfx_slot[slots[i]] >= 0;
  return PTTS_OK;
}

int Engine::prefill(int n, const int* slots, const int* lens, const float* embeds_any, const int* tokens) {
  NEED_READY();
  int rc;
  if ((rc = check_slots(n, slots, false)) != PTTS_OK) return rc;
  if (!lens || (!embeds_any && !tokens)) PTTS_FAIL(PTTS_EINVAL, "prefill: null argument");
  if (tokens && !text_embed) PTTS_FAIL(PTTS_EINVAL, "prefill_tokens: flow_lm.conditioner.embed.weight was not loaded");
  int rows = 0;
  std::vector<int> rs(n);
  for (int i = 0; i < n; ++i) {
    if (lens[i] < 1) PTTS_FAIL(PTTS_EINVAL, "prefill: stream %d has %d rows", i, lens[i]);
    if (host_offsets[slots[i]] + lens[i] > cfg.kv_capacity)
      PTTS_FAIL(PTTS_ENOSLOT, "slot %d: %d + %d positions exceed KV capacity %d", slots[i],
                host_offsets[slots[i]], lens[i], cfg.kv_capacity);
    rs[i] = rows;
    rows += lens[i];
  }
  if (rows > max_rows) PTTS_FAIL(PTTS_EINVAL, "prefill: %d rows exceed max_prefill_rows %d", rows, max_rows);
  prefill_max_len = 0;
  for (int i = 0; i < n; ++i) prefill_max_len = std::max(prefill_max_len, lens[i]);
  if ((rc = upload_slots(n, slots)) != PTTS_OK) return rc;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_row_start, rs.data(), n * sizeof(int), cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(d_lens, lens, n * sizeof(int), cudaMemcpyHostToDevice, stream));
  if (tokens) {
    for (int r = 0; r < rows; ++r)
      if (tokens[r] < 0 || tokens[r] >= text_embed_rows)
        PTTS_FAIL(PTTS_EINVAL, "prefill_tokens: token %d at row %d outside the table of %d rows", tokens[r], r, text_embed_rows);
    PTTS_CUDA_CHECK(cudaMemcpyAsync(d_tokens, tokens, (size_t)rows * sizeof(int), cudaMemcpyHostToDevice, stream));
    PTTS_CUDA_CHECK(launch_embed_gather(d_tokens, text_embed, h, rows, cfg.d_model, stream));
    ++launches;
  } else {
    PTTS_CUDA_CHECK(cudaMemcpyAsync(h, embeds_any, (size_t)rows * cfg.d_model * sizeof(float), cudaMemcpyDefault, stream));
  }
  for (int l = 0; l < cfg.n_layers; ++l) {
    // rows == n would select the decode kernel, which is equivalent for len == 1
    if ((rc = transformer_layer_common(flow_layers[l], h, rows, cfg.d_model, cfg.ffn, a_ln, qkv, attn_o, mid,
                                       false, l, n)) != PTTS_OK)
      return rc;
  }
  PTTS_CUDA_CHECK(launch_advance(offsets, d_slot_ids, d_lens, 0, n, stream));
  ++launches;
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));  // rs / lens host buffers
  for (int i = 0; i < n; ++i) host_offsets[slots[i]] += lens[i];
  return PTTS_OK;
}

static bool is_device_ptr(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// data/audio.py:77-78 on the device; host pointers are staged through one growing device buffer
int Engine::pcm_to_int16(const float* pcm_any, short* out_any, long long n) {
  NEED_READY();
  if (n < 0 || (n > 0 && (!pcm_any || !out_any))) PTTS_FAIL(PTTS_EINVAL, "pcm_to_int16: bad arguments");
  if (n == 0) return PTTS_OK;
  const bool din = is_device_ptr(pcm_any), dout = is_device_ptr(out_any);
  const size_t need = (din ? 0 : (size_t)n * 4) + (dout ? 0 : (size_t)n * 2 + 16);
  if (need > egress_cap) {
    if (egress_buf) cudaFree(egress_buf);
    egress_buf = nullptr; egress_cap = 0;
    PTTS_CUDA_CHECK(cudaMalloc(&egress_buf, need));
    egress_cap = need;
  }
  const float* src = pcm_any;
  char* scratch = reinterpret_cast<char*>(egress_buf);
  if (!din) {
    PTTS_CUDA_CHECK(cudaMemcpyAsync(scratch, pcm_any, (size_t)n * 4, cudaMemcpyHostToDevice, stream));
    src = reinterpret_cast<const float*>(scratch);
    scratch += ((size_t)n * 4 + 15) / 16 * 16;
  }
  short* dst = dout ? out_any : reinterpret_cast<short*>(scratch);
  PTTS_CUDA_CHECK(launch_pcm_to_int16(src, dst, n, stream));
  ++launches;
  if (!dout) PTTS_CUDA_CHECK(cudaMemcpyAsync(out_any, dst, (size_t)n * 2, cudaMemcpyDeviceToHost, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

// one frame for the streams in d_slot_ids with noise in noise_buf; graph-replayed when enabled
int Engine::frame(int n, int lsd, int audio) {
  int rc;
  if (mega_eligible(n)) {   // the phase program is built (cudaMalloc / cudaMemcpy) outside stream capture
    MegaPlan* plan = nullptr;
    if ((rc = mega_plan(lsd, &plan)) != PTTS_OK) return rc;
  }
  if (!use_graphs) {
    if ((rc = flow_sequence(n, lsd)) != PTTS_OK) return rc;
    if (audio && (rc = mimi_sequence(n, x_lat)) != PTTS_OK) return rc;
    return PTTS_OK;
  }
  GraphKey key{n, lsd, audio};
  key.shared = batch_shared ? 1 : 0;
  auto it = graphs.find(key);
  if (it == graphs.end()) {
    const long long before = launches;
    cudaGraph_t graph = nullptr;
    PTTS_CUDA_CHECK(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
    rc = flow_sequence(n, lsd);
    if (rc == PTTS_OK && audio) rc = mimi_sequence(n, x_lat);
    cudaError_t ce = cudaStreamEndCapture(stream, &graph);
    if (rc != PTTS_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
    PTTS_CUDA_CHECK(ce);
    GraphEntry ent;
    ent.kernels = launches - before;
    launches = before;
    PTTS_CUDA_CHECK(cudaGraphInstantiate(&ent.exec, graph, 0));
    cudaGraphDestroy(graph);
    it = graphs.emplace(key, ent).first;
  }
  PTTS_CUDA_CHECK(cudaGraphLaunch(it->second.exec, stream));
  launches += it->second.kernels;
  return PTTS_OK;
}

// Graphs of the pipelined loop.  Measured on B200: a CUDA graph's kernel nodes are issued at ~3.5 us per node by
// one front end shared by all streams, the issue of a graph overlaps the execution of the previous launch only when
// that is a DIFFERENT executable graph, and graphs of more than ~110 nodes lose that overlap.  So the loop launches
// one ~70-node FlowLM graph per frame (executables alternate by frame slot and group parity) and one Mimi graph per
// group of frames, ordered with events instead of intra-graph forks.
int Engine::launch_captured(const GraphKey& key_in, cudaStream_t on, const std::function<int()>& body) {
  if (!use_graphs) return body();
  GraphKey key = key_in;
  key.shared = batch_shared ? 1 : 0;
  auto it = graphs.find(key);
  if (it == graphs.end()) {
    const long long before = launches;
    cudaGraph_t graph = nullptr;
    PTTS_CUDA_CHECK(cudaStreamBeginCapture(on, cudaStreamCaptureModeThreadLocal));
    int rc = body();
    cudaError_t ce = cudaStreamEndCapture(on, &graph);
    if (rc != PTTS_OK) { if (graph) cudaGraphDestroy(graph); return rc; }
    PTTS_CUDA_CHECK(ce);
    GraphEntry ent;
    ent.kernels = launches - before;
    launches = before;
    PTTS_CUDA_CHECK(cudaGraphInstantiate(&ent.exec, graph, 0));
    cudaGraphDestroy(graph);
    it = graphs.emplace(key, ent).first;
  }
  PTTS_CUDA_CHECK(cudaGraphLaunch(it->second.exec, on));
  launches += it->second.kernels;
  return PTTS_OK;
}

// Alternative group form (PTTS_LOOP=fork): ONE graph with two branches per group — `g_flow` FlowLM frames in
// sequence on `stream` beside the Mimi pass over the previous group's `g_mimi` latents on `stream2`.
int Engine::group_graph(int n, int lsd, int parity, int g_flow, int g_mimi) {
  if (g_flow > 0 && mega_eligible(n)) {
    MegaPlan* plan = nullptr;
    const int prc = mega_plan(lsd, &plan);
    if (prc != PTTS_OK) return prc;
  }
  const size_t fl = (size_t)n * cfg.ldim;
  return launch_captured(GraphKey{n, lsd, 1, 1000 + parity * 100 + g_flow * 10 + g_mimi}, stream, [&]() -> int {
    int rc = PTTS_OK;
    const bool both = g_flow > 0 && g_mimi > 0;
    if (both) {
      PTTS_CUDA_CHECK(cudaEventRecord(ev_fork, stream));
      PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream2, ev_fork, 0));
    }
    for (int j = 0; j < g_flow; ++j) {
      commit_out = lat_ring[parity] + j * fl;
      rc = flow_sequence(n, lsd, noise_grp + j * fl, eos_grp[parity] + (size_t)j * n);
      commit_out = nullptr;
      if (rc != PTTS_OK) return rc;
    }
    if (g_mimi > 0) {
      cudaStream_t main_stream = stream;
      if (both) { stream = stream2; n_sms = n_sms_mimi; }   // the launchers read the members
      rc = mimi_sequence(n, lat_ring[parity ^ 1], g_mimi, pcm_ring[parity ^ 1]);
      stream = main_stream;
      n_sms = n_sms_flow;
      if (rc != PTTS_OK) return rc;
      if (both) {
        PTTS_CUDA_CHECK(cudaEventRecord(ev_join, stream2));
        PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_join, 0));
      }
    }
    return PTTS_OK;
  });
}

// FlowLM step + flow head of frame slot j of the group with parity `par`: noise noise_grp[j] ->
// latents lat_ring[par][j], EOS logits eos_grp[par][j]
int Engine::flow_graph(int n, int lsd, int par, int j) {
  if (mega_eligible(n)) {   // built outside capture
    MegaPlan* plan = nullptr;
    const int prc = mega_plan(lsd, &plan);
    if (prc != PTTS_OK) return prc;
  }
  const size_t fl = (size_t)n * cfg.ldim;
  return launch_captured(GraphKey{n, lsd, 0, 2000 + j * 2 + par + (gate_active ? 100 : 0)}, stream, [&]() -> int {
    commit_out = lat_ring[par] + j * fl;
    gate_record = gate_active && j < 8 ? ev_head[par][j] : nullptr;
    const int rc = flow_sequence(n, lsd, noise_grp + j * fl, eos_grp[par] + (size_t)j * n);
    gate_record = nullptr;
    commit_out = nullptr;
    return rc;
  });
}

// Mimi pass over the g latents of lat_ring[par] -> pcm_ring[par], on stream2
int Engine::mimi_graph(int n, int g, int par) {
  cudaStream_t main_stream = stream;
  stream = stream2;   // the launchers read the members
  n_sms = n_sms_mimi;
  const int rc = launch_captured(GraphKey{n, 0, 1, 3000 + g * 2 + par}, stream2,
                                 [&]() -> int { return mimi_sequence(n, lat_ring[par], g, pcm_ring[par]); });
  stream = main_stream;
  n_sms = n_sms_flow;
  return rc;
}

// Head-gated form of a full-size Mimi pass (group parity `par`, g frames): five graphs on stream2, the SEANet ones
// each behind "the flow head of frame slot j of the NEXT group (parity gate_par) has started".  The caller has
// already enqueued that group's FlowLM graphs, so the waits below refer to their event-record nodes.
int Engine::mimi_graph_gated(int n, int g, int par, int gate_par) {
  cudaStream_t main_stream = stream;
  stream = stream2;
  // the flow head keeps ceil(n / 8) SMs busy during the gated segments: persistent Mimi grids leave them alone
  static const int reserve = [] { const char* e = getenv("PTTS_GATE_RESERVE"); return e ? atoi(e) : -1; }();
  const int head_ctas = reserve >= 0 ? reserve : (n + 7) / 8;
  // PTTS_GATE_MAP="a,b,c,d": frame slot whose head start opens segments 1..4 (-1: not gated); default 0,1,2,3
  static int map[MIMI_SEGS] = {-1, 0, 1, 2, 3};
  static const bool map_read = [] {
    if (const char* e = getenv("PTTS_GATE_MAP")) sscanf(e, "%d,%d,%d,%d", &map[1], &map[2], &map[3], &map[4]);
    return true;
  }();
  (void)map_read;
  int rc = PTTS_OK;
  for (int seg = 0; seg < MIMI_SEGS && rc == PTTS_OK; ++seg) {
    static const int seg0_sms = [] { const char* e = getenv("PTTS_GATE_SEG0_SMS"); return e ? atoi(e) : 0; }();   // probe: persistent grids of the ungated segment
    n_sms = map[seg] < 0 ? (seg0_sms > 0 ? std::min(seg0_sms, n_sms_mimi) : n_sms_mimi) : std::max(8, n_sms_mimi - head_ctas);
    if (map[seg] >= 0) {
      const int j = std::min(map[seg], g - 1);
      cudaError_t e = cudaStreamWaitEvent(stream2, ev_head[gate_par][j], 0);
      if (e != cudaSuccess) { set_error(std::string("cudaStreamWaitEvent: ") + cudaGetErrorString(e)); rc = PTTS_ECUDA; break; }
    }
    rc = launch_captured(GraphKey{n, 0, 1, 5000 + g * 100 + seg * 2 + par}, stream2,
                         [&]() -> int { return mimi_sequence(n, lat_ring[par], g, pcm_ring[par], seg, seg); });
  }
  stream = main_stream;
  n_sms = n_sms_flow;
  return rc;
}

int Engine::step(int n, const int* slots, const float* noise_any, int lsd, int audio, float* out_lat,
                 float* out_eos, float* out_pcm) {
  NEED_READY();
  int rc;
  if ((rc = check_slots(n, slots, true)) != PTTS_OK) return rc;
  if (lsd < 1 || lsd > tcomb_rows) PTTS_FAIL(PTTS_EINVAL, "lsd_steps %d outside loaded flow.tcomb rows %d", lsd, tcomb_rows);
  if (!noise_any) PTTS_FAIL(PTTS_EINVAL, "step: noise is required");
  if ((rc = upload_slots(n, slots)) != PTTS_OK) return rc;
  const int Ld = cfg.ldim;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(noise_buf, noise_any, (size_t)n * Ld * sizeof(float), cudaMemcpyDefault, stream));
  if ((rc = frame(n, lsd, audio)) != PTTS_OK) return rc;
  if (out_lat) PTTS_CUDA_CHECK(cudaMemcpyAsync(out_lat, x_lat, (size_t)n * Ld * sizeof(float), cudaMemcpyDefault, stream));
  if (out_eos) PTTS_CUDA_CHECK(cudaMemcpyAsync(out_eos, eos_logit, (size_t)n * sizeof(float), cudaMemcpyDefault, stream));
  if (out_pcm && audio) {
    const size_t fs = (size_t)cfg.upsample_stride * cfg.ratios[0] * cfg.ratios[1] * cfg.ratios[2];
    PTTS_CUDA_CHECK(cudaMemcpyAsync(out_pcm, pcm, n * fs * sizeof(float), cudaMemcpyDefault, stream));
  }
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  for (int i = 0; i < n; ++i) {
    host_offsets[slots[i]] += 1;
    if (audio) host_mimi_offsets[slots[i]] += cfg.upsample_stride;
  }
  return PTTS_OK;
}

int Engine::mimi_decode(int n, const int* slots, const float* latents_any, float* out_pcm) {
  NEED_READY();
  int rc;
  if ((rc = check_slots(n, slots, false)) != PTTS_OK) return rc;
  if (!latents_any) PTTS_FAIL(PTTS_EINVAL, "mimi_decode: null latents");
  if ((rc = upload_slots(n, slots)) != PTTS_OK) return rc;
  PTTS_CUDA_CHECK(cudaMemcpyAsync(lat_in, latents_any, (size_t)n * cfg.ldim * sizeof(float), cudaMemcpyDefault, stream));
  if ((rc = mimi_sequence(n, lat_in)) != PTTS_OK) return rc;
  if (out_pcm) {
    const size_t fs = (size_t)cfg.upsample_stride * cfg.ratios[0] * cfg.ratios[1] * cfg.ratios[2];
    PTTS_CUDA_CHECK(cudaMemcpyAsync(out_pcm, pcm, n * fs * sizeof(float), cudaMemcpyDefault, stream));
  }
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  for (int i = 0; i < n; ++i) host_mimi_offsets[slots[i]] += cfg.upsample_stride;
  return PTTS_OK;
}

// Mimi decode of n_frames consecutive latents per stream (latents [n_frames][n][ldim] -> PCM [n][n_frames*1920]) in
// passes of up to mimi_gmax frames, the pass form the pipelined loop uses (BASELINE config 4: decoder-only sweep).
int Engine::mimi_decode_frames(int n, const int* slots, int n_frames, const float* latents_any, float* out_pcm_any) {
  NEED_READY();
  int rc;
  if ((rc = check_slots(n, slots, false)) != PTTS_OK) return rc;
  if (!latents_any || n_frames < 1) PTTS_FAIL(PTTS_EINVAL, "mimi_decode_frames: bad arguments");
  if ((rc = upload_slots(n, slots)) != PTTS_OK) return rc;
  const size_t fl = (size_t)n * cfg.ldim;
  const size_t fs = (size_t)cfg.upsample_stride * cfg.ratios[0] * cfg.ratios[1] * cfg.ratios[2];
  mimi_alone = true;
  struct Reset { bool& b; ~Reset() { b = false; } } reset_alone{mimi_alone};
  for (int f0 = 0; f0 < n_frames;) {
    const int g = std::min(mimi_gmax, n_frames - f0);
    PTTS_CUDA_CHECK(cudaMemcpyAsync(lat_ring[0], latents_any + (size_t)f0 * fl, g * fl * sizeof(float), cudaMemcpyDefault, stream));
    rc = launch_captured(GraphKey{n, 0, 1, 4000 + g}, stream, [&]() -> int { return mimi_sequence(n, lat_ring[0], g, pcm_ring[0]); });
    if (rc != PTTS_OK) return rc;
    if (out_pcm_any)
      PTTS_CUDA_CHECK(cudaMemcpy2DAsync(out_pcm_any + (size_t)f0 * fs, (size_t)n_frames * fs * sizeof(float), pcm_ring[0],
                                        g * fs * sizeof(float), g * fs * sizeof(float), n, cudaMemcpyDefault, stream));
    f0 += g;
  }
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  for (int i = 0; i < n; ++i) host_mimi_offsets[slots[i]] += n_frames * cfg.upsample_stride;
  return PTTS_OK;
}

// The frame loop (tts_model.py:745-779 + the decoder thread :434-474) for n streams and n_frames frames.
// Pipelined form: group k of the schedule 1, 1, 2, 4, ... (<= mimi_gmax frames) runs its FlowLM frames while the
// Mimi pass decodes group k-1; the result copies of a group ride `copy_stream` beside the next group's graph.
int Engine::run_frames(int n, const int* slots, int n_frames, const float* noise_any, int lsd,
                       float* out_pcm_any, float* out_lat_any, float* out_eos, bool sync) {
  NEED_READY();
  int rc;
  if ((rc = check_slots(n, slots, true)) != PTTS_OK) return rc;
  if (lsd < 1 || lsd > tcomb_rows) PTTS_FAIL(PTTS_EINVAL, "lsd_steps %d outside loaded flow.tcomb rows %d", lsd, tcomb_rows);
  if (n_frames < 1 || !noise_any) PTTS_FAIL(PTTS_EINVAL, "run_frames: bad arguments");
  for (int i = 0; i < n; ++i)
    if (host_offsets[slots[i]] + n_frames > cfg.kv_capacity)
      PTTS_FAIL(PTTS_ENOSLOT, "slot %d: %d frames exceed KV capacity", slots[i], n_frames);
  if ((rc = upload_slots(n, slots)) != PTTS_OK) return rc;
  const int Ld = cfg.ldim;
  const size_t fl = (size_t)n * Ld;
  const size_t fs = (size_t)cfg.upsample_stride * cfg.ratios[0] * cfg.ratios[1] * cfg.ratios[2];
  static const int part = [] { const char* e = getenv("PTTS_PROFILE_PART"); return e ? atoi(e) : 0; }();  // 1: FlowLM+head only, 2: Mimi only (profiling)
  const bool piped = use_pipeline && (part == 0 || part >= 3);   // 3: grouped loop without the Mimi branch (profiling)
  mimi_alone = part == 4;
  struct Reset { bool& b; ~Reset() { b = false; } } reset_alone{mimi_alone};
  if (!piped) {
    for (int f = 0; f < n_frames; ++f) {
      PTTS_CUDA_CHECK(cudaMemcpyAsync(noise_buf, noise_any + (size_t)f * fl, fl * sizeof(float), cudaMemcpyDefault, stream));
      if (part == 2) { if ((rc = mimi_sequence(n, x_lat)) != PTTS_OK) return rc; }
      else if ((rc = frame(n, lsd, part == 1 ? 0 : 1)) != PTTS_OK) return rc;
      if (out_pcm_any)
        PTTS_CUDA_CHECK(cudaMemcpy2DAsync(out_pcm_any + (size_t)f * fs, (size_t)n_frames * fs * sizeof(float), pcm,
                                          fs * sizeof(float), fs * sizeof(float), n, cudaMemcpyDefault, stream));
      if (out_lat_any)
        PTTS_CUDA_CHECK(cudaMemcpyAsync(out_lat_any + (size_t)f * fl, x_lat, fl * sizeof(float), cudaMemcpyDefault, stream));
      if (out_eos)
        PTTS_CUDA_CHECK(cudaMemcpyAsync(out_eos + (size_t)f * n, eos_logit, (size_t)n * sizeof(float), cudaMemcpyDefault, stream));
    }
  } else if (loop_fork) {
    const bool copies = out_pcm_any || out_lat_any || out_eos;
    int f0 = 0, k = 0, prev_f0 = 0, prev_g = 0;
    bool copied[2] = {false, false};
    while (f0 < n_frames || prev_g > 0) {
      const int par = k & 1;
      const int want = k < 2 ? 1 : std::min(mimi_gmax, 1 << std::min(k - 1, 4));
      const int g = std::min(want, n_frames - f0);   // 0 on the drain group
      if (copies && copied[par]) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_copied_lat[par], 0));   // buffers of this parity are free again
      if (g > 0)
        PTTS_CUDA_CHECK(cudaMemcpyAsync(noise_grp, noise_any + (size_t)f0 * fl, g * fl * sizeof(float), cudaMemcpyDefault, stream));
      if ((rc = group_graph(n, lsd, par, g, part == 3 ? 0 : prev_g)) != PTTS_OK) return rc;
      if (copies) {
        PTTS_CUDA_CHECK(cudaEventRecord(ev_lat[par], stream));
        PTTS_CUDA_CHECK(cudaStreamWaitEvent(copy_stream, ev_lat[par], 0));
        if (g > 0 && out_lat_any)
          PTTS_CUDA_CHECK(cudaMemcpyAsync(out_lat_any + (size_t)f0 * fl, lat_ring[par], g * fl * sizeof(float), cudaMemcpyDefault, copy_stream));
        if (g > 0 && out_eos)
          PTTS_CUDA_CHECK(cudaMemcpyAsync(out_eos + (size_t)f0 * n, eos_grp[par], (size_t)g * n * sizeof(float), cudaMemcpyDefault, copy_stream));
        if (prev_g > 0 && out_pcm_any && part != 3)
          PTTS_CUDA_CHECK(cudaMemcpy2DAsync(out_pcm_any + (size_t)prev_f0 * fs, (size_t)n_frames * fs * sizeof(float), pcm_ring[par ^ 1],
                                            prev_g * fs * sizeof(float), prev_g * fs * sizeof(float), n, cudaMemcpyDefault, copy_stream));
        PTTS_CUDA_CHECK(cudaEventRecord(ev_copied_lat[par], copy_stream));
        copied[par] = true;
      }
      prev_f0 = f0; prev_g = g;
      f0 += g;
      ++k;
    }
    for (int q = 0; q < 2; ++q)
      if (copied[q]) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_copied_lat[q], 0));
  } else {
    const bool copies = out_pcm_any || out_lat_any || out_eos;
    const bool do_mimi = part != 3;
    int f0 = 0;
    bool used[2] = {false, false}, used_pcm[2] = {false, false};
    // Gated only when the call holds at least two full-size groups (1 + 1 + 2 + gmax + gmax frames): graphs that carry
    // the event-record node launch more slowly from the host, which short calls (8-frame blocks of the batch
    // scheduler at 256 streams: 2.2 -> 3.4 s per run) cannot hide.
    const bool gate_loop = use_head_gate && n <= head_gate_max && n_frames >= 4 + 2 * mimi_gmax && do_mimi && part == 0 &&
                           use_graphs && use_fused_stage2 && act_type == 1 && use_tc && !mega_eligible(n) && head_mode != 0;
    gate_active = gate_loop;
    struct { bool valid = false; int g = 0, par = 0, f0 = 0; } pend;
    auto enqueue_mimi = [&](int g, int par, int f0m, int gate_par) -> int {
      int rc2;
      PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream2, ev_lat[par], 0));
      if (copies && used_pcm[par]) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream2, ev_copied_pcm[par], 0));
      if ((rc2 = gate_par >= 0 ? mimi_graph_gated(n, g, par, gate_par) : mimi_graph(n, g, par)) != PTTS_OK) return rc2;
      PTTS_CUDA_CHECK(cudaEventRecord(ev_mimi[par], stream2));
      if (copies) {
        PTTS_CUDA_CHECK(cudaStreamWaitEvent(copy_stream, ev_mimi[par], 0));
        if (out_pcm_any)
          PTTS_CUDA_CHECK(cudaMemcpy2DAsync(out_pcm_any + (size_t)f0m * fs, (size_t)n_frames * fs * sizeof(float), pcm_ring[par],
                                            g * fs * sizeof(float), g * fs * sizeof(float), n, cudaMemcpyDefault, copy_stream));
        PTTS_CUDA_CHECK(cudaEventRecord(ev_copied_pcm[par], copy_stream));
      }
      used_pcm[par] = true;
      return PTTS_OK;
    };
    // Group sizes of the call.  Nothing leaves a blocking call before it returns, so the schedule is a throughput
    // matter only: the call lasts as long as its FlowLM chain plus the Mimi pass of the LAST group, which nothing
    // overlaps, and only full-size passes can be head-gated.  Policy 1 (default): full-size groups only, the odd
    // remainder first.  0: the round-1 ramp-up 1, 1, 2, 4, ..; 2: full-size groups, then a ramp-down 2, 1, 1 (a shorter
    // drain, but three ungated passes beside the last frames); 3: ramp-up and ramp-down.  20-frame call at 64 streams:
    // 0.4824 / 0.4558 / 0.4791 / 0.4889 ms per step for 0 / 1 / 2 / 3 (PTTS_SCHED, profiles/r02_group_schedule.txt).
    // Calls shorter than two full groups keep the ramp-up.
    static const int sched_policy = [] { const char* e = getenv("PTTS_SCHED"); return e ? atoi(e) : 1; }();
    std::vector<int> sched;
    {
      const int G = mimi_gmax;
      int left = n_frames;
      const bool down = (sched_policy == 2 || sched_policy == 3) && G == 4 && n_frames >= 2 * G;
      const bool up = sched_policy == 0 || sched_policy == 3 || n_frames < 2 * G || G != 4;
      std::vector<int> tail;
      if (down) { tail = {2, 1, 1}; left -= 4; }
      if (up) {
        for (int k = 0; left > 0; ++k) {
          const int want = k < 2 ? 1 : std::min(G, 1 << std::min(k - 1, 4));
          const int g = std::min(want, left);
          sched.push_back(g);
          left -= g;
        }
      } else {
        if (left % G) { sched.push_back(left % G); left -= left % G; }   // the odd group first: it also starts the Mimi branch early
        for (; left > 0; left -= G) sched.push_back(G);
      }
      sched.insert(sched.end(), tail.begin(), tail.end());
    }
    for (int k = 0; k < (int)sched.size(); ++k) {
      const int par = k & 1;
      const int g = sched[k];
      if (used[par]) {   // buffers of this parity: the Mimi pass and the copies of group k-2 must have drained them
        if (do_mimi) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_mimi[par], 0));
        if (copies) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_copied_lat[par], 0));
      }
      static const bool per_frame_noise = [] { const char* e = getenv("PTTS_NOISE_PER_FRAME"); return !(e && e[0] == '0'); }();
      if (!per_frame_noise)
        PTTS_CUDA_CHECK(cudaMemcpyAsync(noise_grp, noise_any + (size_t)f0 * fl, g * fl * sizeof(float), cudaMemcpyDefault, stream));
      for (int j = 0; j < g; ++j) {
        if (per_frame_noise)
          PTTS_CUDA_CHECK(cudaMemcpyAsync(noise_grp + j * fl, noise_any + (size_t)(f0 + j) * fl, fl * sizeof(float), cudaMemcpyDefault, stream));
        if (part != 4 && (rc = flow_graph(n, lsd, par, j)) != PTTS_OK) return rc;   // 4: Mimi passes only (profiling)
      }
      PTTS_CUDA_CHECK(cudaEventRecord(ev_lat[par], stream));
      if (copies) {
        PTTS_CUDA_CHECK(cudaStreamWaitEvent(copy_stream, ev_lat[par], 0));
        if (out_lat_any)
          PTTS_CUDA_CHECK(cudaMemcpyAsync(out_lat_any + (size_t)f0 * fl, lat_ring[par], g * fl * sizeof(float), cudaMemcpyDefault, copy_stream));
        if (out_eos)
          PTTS_CUDA_CHECK(cudaMemcpyAsync(out_eos + (size_t)f0 * n, eos_grp[par], (size_t)g * n * sizeof(float), cudaMemcpyDefault, copy_stream));
        PTTS_CUDA_CHECK(cudaEventRecord(ev_copied_lat[par], copy_stream));
      }
      if (do_mimi) {
        // head-gated loop: the Mimi pass of a full-size group is enqueued one group late, behind the FlowLM graphs of
        // the next group, whose flow-head events gate its SEANet segments; small groups (first chunks, tail) and the
        // last group run ungated at once
        if (pend.valid) {
          const bool gated = g == pend.g;   // the next group has a head phase for every gated segment
          if ((rc = enqueue_mimi(pend.g, pend.par, pend.f0, gated ? par : -1)) != PTTS_OK) return rc;
          pend.valid = false;
        }
        if (gate_loop && g == mimi_gmax && g >= MIMI_SEGS - 1) {
          pend.valid = true; pend.g = g; pend.par = par; pend.f0 = f0;
        } else if ((rc = enqueue_mimi(g, par, f0, -1)) != PTTS_OK) {
          return rc;
        }
      }
      used[par] = true;
      f0 += g;
    }
    gate_active = false;
    if (pend.valid && (rc = enqueue_mimi(pend.g, pend.par, pend.f0, -1)) != PTTS_OK) return rc;   // last group: nothing to gate on
    // the caller synchronises `stream`: make it wait for the Mimi passes and copies still in flight
    for (int q = 0; q < 2; ++q)
      if (used[q]) {
        if (do_mimi) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_mimi[q], 0));
        if (copies) {
          PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_copied_lat[q], 0));
          if (do_mimi) PTTS_CUDA_CHECK(cudaStreamWaitEvent(stream, ev_copied_pcm[q], 0));
        }
      }
    // x_lat / eos_logit of the last frame stay readable through the single-frame buffers' owners
  }
  for (int i = 0; i < n; ++i) {
    host_offsets[slots[i]] += n_frames;
    host_mimi_offsets[slots[i]] += n_frames * cfg.upsample_stride;
  }
  if (sync) PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  return PTTS_OK;
}

// Debug/test entry: D = A W^T (+bias) on bf16-rounded operands through the selected back-end
// (0 = SIMT, 1 = tcgen05).  A is [n_batches, rows_per_batch(+taps-1 rows), ld] with overlapping
// rows of length K when ld < K (the convolution-as-GEMM addressing).
int Engine::debug_gemm(int backend, int M, int N, int K, int ld, int rows_per_batch, int valid_rows,
                       const float* a_host, long long a_elems, const float* w_host, float* out_host) {
  float *a32 = nullptr, *w32 = nullptr, *o = nullptr;
  bf16 *a16 = nullptr, *w16 = nullptr;
  PTTS_CUDA_CHECK(cudaMalloc(&a32, a_elems * 4 + 512));
  PTTS_CUDA_CHECK(cudaMalloc(&a16, a_elems * 2 + 512));
  PTTS_CUDA_CHECK(cudaMalloc(&w32, (size_t)N * K * 4));
  PTTS_CUDA_CHECK(cudaMalloc(&w16, (size_t)N * K * 2));
  PTTS_CUDA_CHECK(cudaMalloc(&o, (size_t)M * N * 4));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(a32, a_host, a_elems * 4, cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(cudaMemcpyAsync(w32, w_host, (size_t)N * K * 4, cudaMemcpyHostToDevice, stream));
  PTTS_CUDA_CHECK(launch_f32_to_bf16(a32, a16, a_elems, stream));
  PTTS_CUDA_CHECK(launch_f32_to_bf16(w32, w16, (long long)N * K, stream));
  PTTS_CUDA_CHECK(cudaMemsetAsync(o, 0xff, (size_t)M * N * 4, stream));
  GemmProblem p;
  p.A = a16; p.W = w16; p.M = M; p.N = N; p.K = K;
  const bool plain_a = rows_per_batch <= 0;
  p.a_map = plain_a ? RowMap{0, 1 << 30, 1 << 30, ld} : RowMap{(long long)rows_per_batch * ld, rows_per_batch, valid_rows, ld};
  p.ep.out = o;
  p.ep.out_map = plain_a ? RowMap{0, 1 << 30, 1 << 30, N} : RowMap{(long long)rows_per_batch * N, rows_per_batch, valid_rows, N};
  cudaError_t ce;
  if (backend == 1) {
    const long long a_rows = plain_a ? M : (long long)M - (K / ld - 1);
    ce = gemm_tc_launch(p, 0, a_rows, n_sms, pdl_enabled(), stream);
  } else {
    ce = gemm_simt_launch(p, 1, 0, stream);
  }
  PTTS_CUDA_CHECK(ce);
  if (backend == 1) { PTTS_CUDA_CHECK(cudaStreamSynchronize(stream)); tc_trace_dump("debug_gemm"); }
  PTTS_CUDA_CHECK(cudaMemcpyAsync(out_host, o, (size_t)M * N * 4, cudaMemcpyDeviceToHost, stream));
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  cudaFree(a32); cudaFree(a16); cudaFree(w32); cudaFree(w16); cudaFree(o);
  return PTTS_OK;
}

// ---- per-kernel CUDA-event profiling: events are recorded on the engine stream around every
// launch (eager mode), aggregated per tag at profile_end().
void Engine::prof_begin(const char* tag) {
  if (!prof_on) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(stream, &cs);
  if (cs != cudaStreamCaptureStatusNone) return;
  ProfRec r{tag, nullptr, nullptr};
  cudaEventCreate(&r.a);
  cudaEventCreate(&r.b);
  cudaEventRecord(r.a, stream);
  prof_recs.push_back(r);
}
void Engine::prof_end() {
  if (!prof_on || prof_recs.empty() || prof_recs.back().b == nullptr) return;
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(stream, &cs);
  if (cs != cudaStreamCaptureStatusNone) return;
  cudaEventRecord(prof_recs.back().b, stream);
}
int Engine::profile_begin() {
  for (auto& r : prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  prof_recs.clear();
  prof_on = true;
  return PTTS_OK;
}
int Engine::profile_end(char* json_out, int64_t cap) {
  prof_on = false;
  PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
  std::map<std::string, std::pair<long long, double>> agg;
  for (auto& r : prof_recs) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, r.a, r.b) == cudaSuccess) {
      auto& a = agg[r.tag];
      a.first += 1;
      a.second += ms;
    }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  prof_recs.clear();
  std::string js = "{";
  bool first = true;
  for (auto& kv : agg) {
    char b[256];
    snprintf(b, sizeof b, "%s\"%s\": {\"count\": %lld, \"total_ms\": %.6f}", first ? "" : ", ", kv.first.c_str(),
             kv.second.first, kv.second.second);
    js += b;
    first = false;
  }
  js += "}";
  if (!json_out || (int64_t)js.size() + 1 > cap) PTTS_FAIL(PTTS_EINVAL, "profile_end: buffer too small (%zu needed)", js.size() + 1);
  std::memcpy(json_out, js.c_str(), js.size() + 1);
  return PTTS_OK;
}

int Engine::debug_read(const char* what, float* out_any, int64_t max_elems, int64_t* n_out) {
  NEED_READY();
  const std::string w(what ? what : "");
  const int S = cfg.max_streams, U = cfg.upsample_stride;
  struct Ent { const char* name; const void* p; int type; long long elems; };
  const Ent tab[] = {
      {"h", h, 0, (long long)max_rows * cfg.d_model},
      {"qkv", qkv, 0, (long long)max_rows * 3 * cfg.d_model},
      {"attn_o", attn_o, act_type, (long long)max_rows * cfg.d_model},
      {"c_cond", c_cond, act_type, (long long)S * cfg.d_model},
      {"eos_logit", eos_logit, 0, S},
      {"x_lat", x_lat, 0, (long long)S * cfg.ldim},
      {"mod", mod, 0, (long long)S * (cfg.flow_depth * 3 + 2) * cfg.flow_dim},
      {"xm", xm, 0, (long long)S * U * cfg.mimi_d},
      {"b_conv0", b_conv0, act_type, (long long)S * (cfg.kernel - 1 + U) * cfg.mimi_d},
      {"b_tr0", b_tr[0], act_type, (long long)S * (1 + U) * cfg.mimi_d},
      {"b_raw0", b_raw[0], act_type, (long long)S * U * cfg.ratios[0] * cfg.mimi_d / 2},
      {"b_raw1", b_raw[1], act_type, (long long)S * U * cfg.ratios[0] * cfg.ratios[1] * cfg.mimi_d / 4},
      {"b_raw2", b_raw[2], act_type, (long long)S * U * cfg.ratios[0] * cfg.ratios[1] * cfg.ratios[2] * cfg.mimi_d / 8},
      {"pcm", pcm, 0, (long long)S * U * cfg.ratios[0] * cfg.ratios[1] * cfg.ratios[2]},
  };
  for (const Ent& e : tab) {
    if (w != e.name) continue;
    const long long n = std::min<long long>(e.elems, max_elems);
    if (e.type == 0) {
      PTTS_CUDA_CHECK(cudaMemcpyAsync(out_any, e.p, n * sizeof(float), cudaMemcpyDefault, stream));
    } else {
      float* tmp = nullptr;
      PTTS_CUDA_CHECK(cudaMallocAsync((void**)&tmp, n * sizeof(float), stream));
      PTTS_CUDA_CHECK(launch_bf16_to_f32((const bf16*)e.p, tmp, n, stream));
      PTTS_CUDA_CHECK(cudaMemcpyAsync(out_any, tmp, n * sizeof(float), cudaMemcpyDefault, stream));
      PTTS_CUDA_CHECK(cudaFreeAsync(tmp, stream));
    }
    PTTS_CUDA_CHECK(cudaStreamSynchronize(stream));
    if (n_out) *n_out = n;
    return PTTS_OK;
  }
  PTTS_FAIL(PTTS_EINVAL, "debug_read: unknown buffer '%s'", w.c_str());
}

}  // namespace ptts
