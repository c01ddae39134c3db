// extern "C" entry points of libsietill_b200.so (declared in include/sietill_b200.h).
// Host-side glue only: buffer management, table packing (K5), launch ordering, copies, NCCL.
#include <cuda_fp16.h>
#include <dlfcn.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <atomic>
#include <map>
#include <string>

#include "sgmm_internal.cuh"

namespace {
std::string g_create_error;

template <typename T>
int upload(sgmm_ctx* c, DevBuf<T>& buf, const T* src, size_t n) {
  SGMM_CHECK(c, buf.reserve(n ? n : 1));
  if (n) SGMM_CHECK(c, cudaMemcpyAsync(buf.p, src, n * sizeof(T), cudaMemcpyHostToDevice, c->stream));
  return 0;
}

void begin_call(sgmm_ctx* c) { sgmm_stage_reset(c); }

constexpr int kViterbiC[] = {2, 3, 4, 5, 6, 8, 10, 12, 16};

size_t viterbi_smem(uint32_t T, uint32_t S, uint32_t ncol, int C, int W = 1) {
  size_t b = 64;
  if (W > 1) b += 2 * (size_t)W * 4 * sizeof(double) + 16;     // k_viterbi_mw: boundary-cost exchange + end position
  b += (size_t)SGMM_STAGES * SGMM_TILE_T * ncol * sizeof(double);
  b += (size_t)((T + 15) / 16) * 32 * W * C * sizeof(uint32_t);
  b += 2 * (size_t)S * sizeof(uint16_t);
  return (b + 15) & ~(size_t)15;
}
}  // namespace

extern "C" {

const char* sgmm_version(void) { return "sietill-b200 0.1 (sm_100a)"; }

const char* sgmm_last_error(const sgmm_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int sgmm_create(int device, sgmm_ctx** out) {
  if (!out) return 1;
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) +
                     " (this library has no CPU fallback)";
    return 1;
  }
  if (device < 0 || device >= n) { g_create_error = "device index out of range"; return 1; }
  e = cudaSetDevice(device);
  if (e != cudaSuccess) { g_create_error = cudaGetErrorString(e); return 1; }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  if (prop.major != 10) {
    g_create_error = "device is sm_" + std::to_string(prop.major * 10 + prop.minor) +
                     "; this library is built for sm_100a (B200) only";
    return 1;
  }
  sgmm_ctx* c = new sgmm_ctx();
  c->device = device;
  { static std::atomic<uint64_t> next_serial{1}; c->serial = next_serial.fetch_add(1); }
  if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
    g_create_error = "cudaStreamCreate failed";
    delete c;
    return 1;
  }
  for (auto& ev : c->ev) cudaEventCreate(&ev);
  for (auto& ev : c->tev) cudaEventCreate(&ev);
  for (auto& st : c->aux) cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming);
  for (auto& ev : c->ev_join) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking);
  for (auto& ev : c->feat_ev) cudaEventCreateWithFlags(&ev, cudaEventDisableTiming);
  *out = c;
  return 0;
}

void sgmm_destroy(sgmm_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  c->release_all(); c->spare.release_all();
  c->d_x.release(); c->d_raw.release(); c->d_feat_off.release(); c->d_feat_norm.release();
  c->d_frame_off.release(); c->d_E_off.release(); c->d_aut.release(); c->d_aut_off.release();
  c->d_uniq.release(); c->d_uniq_off.release(); c->d_nuniq_pad.release(); c->d_pos2uniq.release();
  c->d_tasks.release(); c->d_tasks64.release(); c->d_tasks128.release(); c->d_tasks256.release();
  c->d_fallbacks.release(); c->d_Amin.release(); c->d_fcol.release(); c->d_Ahi.release(); c->d_Alo.release();
  c->d_xnorm.release(); c->d_ci1.release(); c->d_e32.release(); c->d_summask.release(); c->d_rescan.release(); c->d_E.release(); c->d_all_states.release(); c->d_ali.release();
  c->d_costs.release(); c->d_fdens.release(); c->d_fscore.release();
  c->d_sm_items.release(); c->d_sm_ent_a.release(); c->d_sm_ent_b.release(); c->d_sm_ivoff.release(); c->d_dsm_items.release(); c->d_dsm_ent_a.release(); c->d_dsm_ent_b.release();
  c->d_pack_desc.release(); c->d_pack_uniq.release(); c->d_pack_rows.release(); c->d_pack_colmap.release();
  for (int i = 0; i < 2; i++) { c->d_keys[i].release(); c->d_vals[i].release(); }
  c->d_hist.release(); c->d_cnt_u32.release(); c->d_seg_off.release(); c->d_item_off.release(); c->d_soft_keys.release();
  c->d_partial.release(); c->d_acc.release(); c->d_score_part.release();
  c->d_dense_scores.release(); c->d_dense_argmin.release(); c->d_model_flags.release(); c->d_mstep_tmp.release();
  c->d_lex_info.release(); c->d_lex_state.release(); c->d_lex_base.release(); c->d_rec_perm.release(); c->d_rec_woff.release(); c->d_rec_compact.release();
  c->d_rec_frame_off.release(); c->d_rec_tb_off.release(); c->d_rec_tb.release();
  c->d_rec_words.release(); c->d_rec_nwords.release(); c->d_rec_score.release();
  for (auto& vc : c->vclasses) vc.d_utts.release();
  for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
  for (auto& ev : c->tev) if (ev) cudaEventDestroy(ev);
  for (auto& st : c->aux) if (st) cudaStreamDestroy(st);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  for (auto& ev : c->ev_join) if (ev) cudaEventDestroy(ev);
  if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
  for (auto& ev : c->feat_ev) if (ev) cudaEventDestroy(ev);
  cudaStreamDestroy(c->stream);
  delete c;
}

int sgmm_set_features(sgmm_ctx* c, const float* x, uint64_t n_frames, uint32_t dim) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  begin_call(c);
  if (c->feat_pending) {   // a queued upload nobody consumed: let it finish before its target is reused
    SGMM_CHECK(c, cudaStreamSynchronize(c->copy_stream));
    c->feat_pending = 0;
  }
  SGMM_CHECK(c, c->d_x.reserve((size_t)n_frames * dim + 8));
  if (c->async_upload && n_frames >= 64 * (uint64_t)sgmm_ctx::kFeatChunks) {
    // every compute entry point ends with a stream synchronize, so nothing is reading d_x now.  The first chunk
    // is the smallest (the pipeline starts early), the others are equal.
    const int n = c->feat_chunks;
    uint64_t f0 = 0;
    for (int i = 0; i < n; i++) {
      const uint64_t f1 = (i + 1 == n) ? n_frames : (n_frames * (2 * (uint64_t)i + 1)) / (2 * (uint64_t)n - 1);
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_x.p + f0 * dim, x + f0 * dim, (f1 - f0) * dim * sizeof(float),
                                    cudaMemcpyHostToDevice, c->copy_stream));
      SGMM_CHECK(c, cudaEventRecord(c->feat_ev[i], c->copy_stream));
      c->feat_chunk_end[i] = f1;
      f0 = f1;
    }
    c->feat_pending = n;
  } else {
    {
      StageTimer st(c, SGMM_T_H2D);
      SGMM_CHECK(c, cudaMemcpyAsync(c->d_x.p, x, (size_t)n_frames * dim * sizeof(float), cudaMemcpyHostToDevice,
                                    c->stream));
    }
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  }
  c->n_frames = n_frames; c->feat_dim = dim;
  c->have_alignment = false;
  c->ali_has_cols = false;
  c->tc_A_valid = false;
  return 0;
}

int sgmm_set_tdp(sgmm_ctx* c, double loop, double forward, double skip, uint32_t silence_state) {
  if (!c) return 1;
  c->tdp_loop = loop; c->tdp_forward = forward; c->tdp_skip = skip; c->silence_state = silence_state;
  return 0;
}

int sgmm_set_beam(sgmm_ctx* c, double pruning_threshold) {
  if (!c) return 1;
  if (!(pruning_threshold >= 0.0)) SGMM_FAIL(c, "sgmm_set_beam: the threshold must be >= 0 (+inf disables pruning)");
  c->beam = pruning_threshold;
  return 0;
}

int sgmm_set_corpus(sgmm_ctx* c, uint32_t U, const uint64_t* frame_off, const uint32_t* automata,
                    const uint32_t* aut_off) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (frame_off[0] != 0 || aut_off[0] != 0) SGMM_FAIL(c, "set_corpus: offsets must start at 0");
  std::vector<uint32_t> uniq, uniq_off(U + 1, 0), ncol(U);
  std::vector<uint16_t> p2u(aut_off[U]);
  std::vector<uint64_t> E_off(U + 1, 0);
  std::vector<uint4> tasks, tasks64, tasks128, tasks256;
  std::vector<uint64_t> t128_end, t256_end;
  std::map<std::pair<int, size_t>, ViterbiClass> classes;   // key: (C + 100 W, shared-memory bucket)
  uint32_t max_state = 0, max_uniq = 0;
  for (uint32_t u = 0; u < U; u++) {
    if (frame_off[u + 1] <= frame_off[u]) SGMM_FAIL(c, "set_corpus: empty utterance");
    const uint64_t T64 = frame_off[u + 1] - frame_off[u];
    if (T64 > 65535) SGMM_FAIL(c, "set_corpus: utterance longer than 65535 frames");
    const uint32_t T = (uint32_t)T64;
    const uint32_t S = aut_off[u + 1] - aut_off[u];
    if (S == 0) SGMM_FAIL(c, "set_corpus: empty automaton");
    if (S > 512) SGMM_FAIL(c, "set_corpus: automaton longer than 512 positions");
    const uint32_t* a = automata + aut_off[u];
    for (uint32_t s = 0; s < S; s++) max_state = std::max(max_state, a[s]);
    std::vector<uint32_t> us(a, a + S);
    std::sort(us.begin(), us.end());
    us.erase(std::unique(us.begin(), us.end()), us.end());
    for (uint32_t s = 0; s < S; s++)
      p2u[aut_off[u] + s] = (uint16_t)(std::lower_bound(us.begin(), us.end(), a[s]) - us.begin());
    uniq.insert(uniq.end(), us.begin(), us.end());
    uniq_off[u + 1] = (uint32_t)uniq.size();
    ncol[u] = ((uint32_t)us.size() + 1u) & ~1u;
    max_uniq = std::max(max_uniq, (uint32_t)us.size());
    E_off[u + 1] = E_off[u] + sgmm_cells_of(T, ncol[u]);   // whole 16-frame tiles (sgmm_cell_rel)
    for (uint32_t t0 = 0; t0 < T; t0 += 32) tasks.push_back(make_uint4(u, t0, 0, 0));
    for (uint32_t t0 = 0; t0 < T; t0 += 64) tasks64.push_back(make_uint4(u, t0, 0, 0));
    for (uint32_t t0 = 0; t0 < T; t0 += 128) {
      tasks128.push_back(make_uint4(u, t0, 0, 0));
      t128_end.push_back(frame_off[u] + std::min(T, t0 + 128));
    }
    // long utterances: positions split over four warps (k_viterbi_mw), C <= 4 positions per lane
    const int W = (c->vit_mw && T >= (uint32_t)c->vit_mw) ? SGMM_VIT_MW_WARPS : 1;
    int C = 0;
    for (int cc : kViterbiC) if ((uint32_t)cc * 32 * W >= S) { C = cc; break; }
    size_t sm = viterbi_smem(T, S, ncol[u], C, W);
    if (sm > 227 * 1024) SGMM_FAIL(c, "set_corpus: utterance trellis exceeds shared memory (T*S too large)");
    size_t bucket = 8 * 1024;  // classes by shared-memory bucket so short utterances keep high occupancy
    while (bucket < sm) bucket += bucket < 64 * 1024 ? 8 * 1024 : 32 * 1024;
    bucket = std::min(bucket, (size_t)227 * 1024);
    ViterbiClass& vc = classes[{C + 100 * W, bucket}];
    vc.C = C; vc.W = W; vc.smem = bucket;
    vc.utts.push_back(u);
  }
  // ---- GEMM tasks of 256 frames.  The last task of an utterance often holds few frames (T = 279 leaves 23), yet a
  // 128-row MMA half costs the same however many rows are valid.  Short tails of consecutive utterances are
  // therefore PACKED into one 128-row half that is scored against the union of their state lists, when that
  // is cheaper than one half per tail.  A packed task is {0x80000000 | pack id}; per row it carries the frame,
  // the first emission cell of that frame's row, and a map union column -> column of the utterance's own list.
  std::vector<uint4> pack_desc;                  // {first entry in pack_uniq, states in the union, valid rows, 0}
  std::vector<uint32_t> pack_uniq;
  std::vector<sgmm_pack_row> pack_rows;          // [pack][128]
  std::vector<uint16_t> pack_colmap;
  {
    struct Seg { uint32_t u, t0, rows; };
    std::vector<Seg> open;
    uint32_t open_rows = 0;
    uint64_t last_end = 0;
    auto emit_plain = [&](uint32_t u, uint32_t t0, uint32_t T) {
      tasks256.push_back(make_uint4(u, t0, 0, 0));
      last_end = std::max(last_end, frame_off[u] + std::min(T, t0 + 256));
      t256_end.push_back(last_end);
    };
    auto close_pack = [&]() {
      if (open.empty()) return;
      std::vector<uint32_t> uni;
      size_t own = 0;
      for (const Seg& sg : open) {
        uni.insert(uni.end(), uniq.begin() + uniq_off[sg.u], uniq.begin() + uniq_off[sg.u + 1]);
        own += uniq_off[sg.u + 1] - uniq_off[sg.u];
      }
      std::sort(uni.begin(), uni.end());
      uni.erase(std::unique(uni.begin(), uni.end()), uni.end());
      if (open.size() == 1 || uni.size() * 10 > own * 9) {       // the union saves < 10 %: one half per tail after all
        for (const Seg& sg : open) emit_plain(sg.u, sg.t0, (uint32_t)(frame_off[sg.u + 1] - frame_off[sg.u]));
      } else {
        const uint32_t id = (uint32_t)pack_desc.size();
        pack_desc.push_back(make_uint4((uint32_t)pack_uniq.size(), (uint32_t)uni.size(), open_rows, 0));
        pack_uniq.insert(pack_uniq.end(), uni.begin(), uni.end());
        pack_rows.resize((size_t)(id + 1) * 128);
        sgmm_pack_row* rows = pack_rows.data() + (size_t)id * 128;
        uint32_t r = 0;
        uint64_t end = 0;
        for (const Seg& sg : open) {
          const uint32_t cm = (uint32_t)pack_colmap.size();
          const uint32_t* ub = uniq.data() + uniq_off[sg.u];
          const uint32_t nu_u = uniq_off[sg.u + 1] - uniq_off[sg.u];
          for (uint32_t s : uni) {
            const uint32_t* it = std::lower_bound(ub, ub + nu_u, s);
            pack_colmap.push_back((it != ub + nu_u && *it == s) ? (uint16_t)(it - ub) : (uint16_t)0xFFFF);
          }
          for (uint32_t i = 0; i < sg.rows; i++, r++) {
            rows[r].cell = E_off[sg.u] + (uint64_t)((sg.t0 + i) >> 4) * (16u * ncol[sg.u]) + ((sg.t0 + i) & 15u);
            rows[r].frame = (uint32_t)(frame_off[sg.u] + sg.t0 + i);
            rows[r].colmap = cm;
          }
          end = std::max(end, frame_off[sg.u] + sg.t0 + sg.rows);
        }
        for (; r < 128; r++) rows[r] = rows[0];                   // rows past the end: any valid frame, never written
        tasks256.push_back(make_uint4(0x80000000u | id, 0, 0, 0));
        last_end = std::max(last_end, end);
        t256_end.push_back(last_end);
      }
      open.clear(); open_rows = 0;
    };
    const uint32_t kPackMaxRows = 96;
    for (uint32_t u = 0; u < U; u++) {
      const uint32_t T = (uint32_t)(frame_off[u + 1] - frame_off[u]);
      const uint32_t tail = T % 256;
      const bool packable = c->tc_pack && frame_off[U] < 0xffffffffull && tail > 0 && tail <= kPackMaxRows;
      if (packable && (open_rows + tail > 128 || open.size() >= 8)) close_pack();
      for (uint32_t t0 = 0; t0 + 256 <= T; t0 += 256) emit_plain(u, t0, T);
      if (tail == 0) continue;
      if (packable) { open.push_back({u, T - tail, tail}); open_rows += tail; }
      else emit_plain(u, T - tail, T);
    }
    close_pack();
  }
  // ---- state-major refinement items.  Utterances are cut into blocks; an item is one state of one block and lists the
  // block's utterances that visit it (at most kSmMaxEntries of them, at most kSmMaxChunks chunks of 32 frames).  A block
  // ends at about kSmBlockFrames frames once its busiest state has kSmTargetChunks chunks: where every utterance visits
  // a quarter of the states (the digits) that is reached long before, and the frame bound keeps the items small enough to
  // balance over the SMs; where an utterance visits 150 of 4501 states (config 5) a 64 k-frame block would leave ~70
  // chunks per item, and the item's prologue (the state's table into shared memory, 1 CTA / SM) would cost more than
  // its chunks — measured 11.3 ms for 4.7 M chunks in 69 738 items.
  // Items are ordered block-major so that the CTAs in flight write into a few MB of the emission matrix.
  std::vector<uint4> sm_items, sm_ent_a;
  std::vector<uint2> sm_ent_b;
  {
    std::vector<std::vector<std::pair<uint32_t, uint32_t>>> by_state(max_state + 1);   // state -> (utt, column)
    std::vector<uint32_t> state_chunks(max_state + 1, 0u);
    std::vector<uint32_t> touched;
    uint32_t u0 = 0;
    while (u0 < U) {
      uint32_t u1 = u0, busiest = 0;
      touched.clear();
      while (u1 < U) {
        const uint32_t T = (uint32_t)(frame_off[u1 + 1] - frame_off[u1]), ch = (T + 31u) / 32u;
        if (u1 > u0) {
          if (frame_off[u1 + 1] - frame_off[u0] > kSmBlockFrames && busiest >= kSmTargetChunks) break;
          bool fits = true;
          for (uint32_t j = uniq_off[u1]; j < uniq_off[u1 + 1] && fits; j++)
            fits = by_state[uniq[j]].size() < kSmMaxEntries && state_chunks[uniq[j]] + ch <= kSmMaxChunks;
          if (!fits) break;
        }
        for (uint32_t j = uniq_off[u1]; j < uniq_off[u1 + 1]; j++) {
          const uint32_t st = uniq[j];
          if (by_state[st].empty()) touched.push_back(st);
          by_state[st].push_back({u1, j - uniq_off[u1]});
          state_chunks[st] += ch;
          busiest = std::max(busiest, state_chunks[st]);
        }
        u1++;
      }
      std::sort(touched.begin(), touched.end());
      for (uint32_t s : touched) {
        uint32_t chunks = 0;
        const uint32_t first = (uint32_t)sm_ent_a.size();
        for (auto& uc : by_state[s]) {
          const uint64_t e0 = E_off[uc.first], f0 = frame_off[uc.first];
          const uint32_t T = (uint32_t)(frame_off[uc.first + 1] - f0);
          sm_ent_a.push_back(make_uint4((uint32_t)e0, (uint32_t)(e0 >> 32), ncol[uc.first] | (uc.second << 16), chunks));
          sm_ent_b.push_back(make_uint2((uint32_t)f0, T));
          chunks += (T + 31u) / 32u;
        }
        sm_items.push_back(make_uint4(s, first, (uint32_t)by_state[s].size(), chunks));
        by_state[s].clear();
        state_chunks[s] = 0;
      }
      u0 = u1;
    }
  }
  for (auto& vc : c->vclasses) vc.d_utts.release();
  c->vclasses.clear();
  for (auto& kv : classes) {
    // longest utterances first inside a class: the tail of the launch is then made of short ones
    std::stable_sort(kv.second.utts.begin(), kv.second.utts.end(), [&](uint32_t a, uint32_t b) {
      return frame_off[a + 1] - frame_off[a] > frame_off[b + 1] - frame_off[b];
    });
    c->vclasses.push_back(std::move(kv.second));
  }
  // launch order (sgmm_launch_viterbi forks the classes over side streams): the class whose longest utterance is the
  // most work goes first — a frame costs about (1 + 2C) issue slots per lane — so the few very long utterances start at
  // once and the many short ones fill the machine around them
  {
    auto longest = [&](const ViterbiClass& vc) {
      const uint32_t u = vc.utts.front();
      return (uint64_t)(frame_off[u + 1] - frame_off[u]) * (uint64_t)(1 + 2 * vc.C);
    };
    std::stable_sort(c->vclasses.begin(), c->vclasses.end(),
                     [&](const ViterbiClass& a, const ViterbiClass& b) { return longest(a) > longest(b); });
  }
  for (auto& vc : c->vclasses)
    if (upload(c, vc.d_utts, vc.utts.data(), vc.utts.size())) return 1;
  if (upload(c, c->d_frame_off, frame_off, (size_t)U + 1) || upload(c, c->d_aut, automata, (size_t)aut_off[U]) ||
      upload(c, c->d_aut_off, aut_off, (size_t)U + 1) || upload(c, c->d_uniq, uniq.data(), uniq.size()) ||
      upload(c, c->d_uniq_off, uniq_off.data(), uniq_off.size()) || upload(c, c->d_nuniq_pad, ncol.data(), ncol.size()) ||
      upload(c, c->d_pos2uniq, p2u.data(), p2u.size()) || upload(c, c->d_E_off, E_off.data(), E_off.size()) ||
      upload(c, c->d_tasks, tasks.data(), tasks.size()) || upload(c, c->d_tasks64, tasks64.data(), tasks64.size()) ||
      upload(c, c->d_tasks128, tasks128.data(), tasks128.size()) ||
      upload(c, c->d_tasks256, tasks256.data(), tasks256.size()) ||
      upload(c, c->d_pack_desc, pack_desc.data(), pack_desc.size()) ||
      upload(c, c->d_pack_uniq, pack_uniq.data(), pack_uniq.size()) ||
      upload(c, c->d_pack_rows, pack_rows.data(), pack_rows.size()) ||
      upload(c, c->d_pack_colmap, pack_colmap.data(), pack_colmap.size()) ||
      upload(c, c->d_sm_items, sm_items.data(), sm_items.size()) ||
      upload(c, c->d_sm_ent_a, sm_ent_a.data(), sm_ent_a.size()) ||
      upload(c, c->d_sm_ent_b, sm_ent_b.data(), sm_ent_b.size()))
    return 1;
  c->n_packs = pack_desc.size();
  c->n_sm_items = sm_items.size();
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  c->max_aut_state = max_state; c->max_uniq = max_uniq;
  c->n_utts = U; c->n_tasks = tasks.size(); c->n_tasks64 = tasks64.size(); c->n_tasks128 = tasks128.size(); c->n_tasks256 = tasks256.size(); c->E_elems = E_off[U];
  c->h_t128_end.swap(t128_end); c->h_t256_end.swap(t256_end);
  c->h_frame_off.assign(frame_off, frame_off + U + 1);
  c->h_aut_off.assign(aut_off, aut_off + U + 1);
  // whatever the emission matrix / aligned columns held belongs to the previous corpus layout: the cached E-step
  // (k_rescore_cached) must not index it with the new offsets
  c->ali_has_cols = false;
  c->E_model_version = 0; c->tol_model_version = 0; c->tol_active = false;
  return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// NCCL, loaded lazily so that single-GPU users (and the CPU-only build check) never need libnccl.
// ---------------------------------------------------------------------------------------------------
namespace {
struct NcclApi {
  void* lib = nullptr;
  int (*GetUniqueId)(void*) = nullptr;
  int (*CommInitRank)(void**, int, /*ncclUniqueId by value*/ struct Id128, int) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
struct Id128 { char internal[128]; };
NcclApi g_nccl;

bool load_nccl(std::string& err) {
  if (g_nccl.lib) return true;
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names) {
    g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (g_nccl.lib) break;
  }
  if (!g_nccl.lib) { err = std::string("cannot load libnccl: ") + dlerror(); return false; }
  g_nccl.GetUniqueId = (int (*)(void*))dlsym(g_nccl.lib, "ncclGetUniqueId");
  g_nccl.CommInitRank = (int (*)(void**, int, Id128, int))dlsym(g_nccl.lib, "ncclCommInitRank");
  g_nccl.AllReduce = (int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t))dlsym(g_nccl.lib, "ncclAllReduce");
  g_nccl.CommDestroy = (int (*)(void*))dlsym(g_nccl.lib, "ncclCommDestroy");
  g_nccl.GetErrorString = (const char* (*)(int))dlsym(g_nccl.lib, "ncclGetErrorString");
  if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.AllReduce) { err = "libnccl lacks required symbols"; return false; }
  return true;
}
}  // namespace

int sgmm_nccl_allreduce(sgmm_ctx* c, double* d_buf, size_t n) {
  if (!c->nccl_comm || c->world <= 1 || !c->allreduce) return 0;
  StageTimer st(c, SGMM_T_ALLREDUCE);
  const int ncclFloat64 = 8, ncclSum = 0;
  int rc = g_nccl.AllReduce(d_buf, d_buf, n, ncclFloat64, ncclSum, c->nccl_comm, c->stream);
  if (rc != 0) SGMM_FAIL(c, std::string("ncclAllReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"));
  return 0;
}

extern "C" {

int sgmm_comm_unique_id(void* id_out, size_t id_bytes) {
  if (id_bytes < 128) return 1;
  if (!load_nccl(g_create_error)) return 1;
  return g_nccl.GetUniqueId(id_out) == 0 ? 0 : 1;
}

int sgmm_comm_init(sgmm_ctx* c, int rank, int world, const void* id, size_t id_bytes) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  if (world <= 1) { c->rank = 0; c->world = 1; return 0; }
  if (id_bytes < 128) SGMM_FAIL(c, "comm_init: unique id must be 128 bytes");
  if (!load_nccl(c->err)) return 1;
  Id128 uid;
  memcpy(uid.internal, id, 128);
  int rc = g_nccl.CommInitRank(&c->nccl_comm, world, uid, rank);
  if (rc != 0) SGMM_FAIL(c, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(rc) : "error"));
  c->rank = rank; c->world = world;
  return 0;
}

int sgmm_stage_ms(sgmm_ctx* c, float* ms) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
  for (int i = 0; i < SGMM_T_COUNT; i++) {
    sgmm_stage_fold(c, i);
    ms[i] = c->stage_ms[i];
  }
  return 0;
}

int sgmm_set_option(sgmm_ctx* c, const char* key, int64_t value) {
  if (!c || !key) return 1;
  if (!strcmp(key, "scorer")) {
    if (value < 0 || value > 3) SGMM_FAIL(c, "set_option: scorer must be 0 (auto), 1 (f64), 2 (ffma) or 3 (tensor)");
    c->scorer = (int)value;
    return 0;
  }
  if (!strcmp(key, "rescore_cache")) { c->use_cache = value != 0; return 0; }
  if (!strcmp(key, "tc_debug")) { c->tc_debug = (int)value; return 0; }
  if (!strcmp(key, "async_upload")) { c->async_upload = value != 0; return 0; }
  if (!strcmp(key, "tc_wide")) { c->tc_wide = (int)value; return 0; }   // takes effect at the next sgmm_set_model
  if (!strcmp(key, "max_approx")) { c->max_approx = value != 0; return 0; }
  if (!strcmp(key, "sum_prune")) { c->sum_prune = value != 0; return 0; }
  if (!strcmp(key, "tc_epi16")) { c->tc_epi16 = (int)value; return 0; }
  if (!strcmp(key, "tc_f16")) { c->tc_f16 = (int)value; return 0; }   // read by sgmm_set_model and at every launch
  if (!strcmp(key, "tc_rem")) { c->tc_rem = (int)value; return 0; }   // read by sgmm_set_model
  if (!strcmp(key, "rec_run_frames")) { c->rec_run_frames = value < 0 ? 0 : value; return 0; }
  if (!strcmp(key, "rec_reg")) { c->rec_reg = value != 0; return 0; }
  if (!strcmp(key, "vit_mw")) {   // read by sgmm_set_corpus: 0 = never, 1 = default threshold, n > 1 = utterances of >= n frames
    c->vit_mw = value <= 0 ? 0 : (value == 1 ? SGMM_VIT_MW_MIN_FRAMES : (int)value);
    return 0;
  }
  if (!strcmp(key, "tc_rescan")) { c->tc_rescan = (int)value; return 0; }
  if (!strcmp(key, "tc_refine")) { c->tc_refine = (int)value; return 0; }
  if (!strcmp(key, "tc_civ")) { c->tc_civ = value != 0; return 0; }
  if (!strcmp(key, "tc_pairs")) { c->tc_pairs = value != 0; return 0; }
  if (!strcmp(key, "tc_dense_sm")) { c->tc_dense_sm = value != 0; return 0; }
  if (!strcmp(key, "sum_tc")) { c->sum_tc = value != 0; return 0; }
  if (!strcmp(key, "tolerance")) { c->tol = value != 0; return 0; }
  if (!strcmp(key, "tc_lq")) { c->tc_lq = (int)value; return 0; }   // 0 none, 1 default size, n > 1 entries
  if (!strcmp(key, "host_tables")) { c->host_tables = value != 0; return 0; }   // read by sgmm_set_model
  if (!strcmp(key, "allreduce")) { c->allreduce = value != 0; return 0; }   // 0: sgmm_accumulate* return this rank's partial sums
  if (!strcmp(key, "tc_pack")) { c->tc_pack = value != 0; return 0; }   // takes effect at the next sgmm_set_corpus
  if (!strcmp(key, "upload_chunks")) {
    if (value < 1 || value > sgmm_ctx::kFeatChunks) SGMM_FAIL(c, "set_option: upload_chunks out of range");
    c->feat_chunks = (int)value;
    return 0;
  }
  if (!strcmp(key, "tc_halves")) {
    if (value < 0 || value > 2) SGMM_FAIL(c, "set_option: tc_halves must be 0 (auto), 1 or 2");
    c->tc_halves = (int)value;
    return 0;
  }
  SGMM_FAIL(c, std::string("set_option: unknown key ") + key);
}

int sgmm_get_counter(sgmm_ctx* c, const char* key, int64_t* value) {
  if (!c || !key || !value) return 1;
  cudaSetDevice(c->device);
  if (!strcmp(key, "fallbacks")) {
    unsigned long long v = 0;
    if (c->d_fallbacks.p) {
      SGMM_CHECK(c, cudaMemcpyAsync(&v, c->d_fallbacks.p, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
      SGMM_CHECK(c, cudaMemsetAsync(c->d_fallbacks.p, 0, sizeof(v), c->stream));
      SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
    }
    *value = (int64_t)v;
    return 0;
  }
  if (!strcmp(key, "scorer")) { *value = sgmm_effective_scorer(c); return 0; }
  if (!strcmp(key, "frames_global")) {   // frames behind the last (all-reduced) statistics / AM-score sum
    if (!c->d_acc.p || c->Dtot == 0) SGMM_FAIL(c, "get_counter: no statistics yet");
    double v = 0.0;
    SGMM_CHECK(c, cudaMemcpyAsync(&v, c->d_acc.p + (size_t)c->Dtot * (1 + 2 * (size_t)c->dim) + 1, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
    *value = (int64_t)v;
    return 0;
  }
  if (!strcmp(key, "tc_rem")) { *value = c->tc_f16_ok ? c->tc_rem16 : 0; return 0; }   // packed K remainder in use (terms)
  if (!strcmp(key, "packs")) { *value = (int64_t)c->n_packs; return 0; }   // packed GEMM tasks of the resident corpus
  // operand form the wide GEMM runs with (k_score_tc.cu): 0 tf32 hi/lo (3xTF32), 1 fp16 hi/lo (kind::f16);
  // k-steps per tile of that form
  if (!strcmp(key, "tc_form")) { *value = (c->tc_ok && c->tc_tile_rows == 256 && c->tc_f16_ok && c->tc_f16 != 0) ? 1 : 0; return 0; }
  if (!strcmp(key, "tc_ksteps")) {
    *value = (c->tc_ok && c->tc_tile_rows == 256 && c->tc_f16_ok && c->tc_f16 != 0) ? c->tc_ks16 : c->tc_ksteps;
    return 0;
  }
  if (!strncmp(key, "tc_prof", 7)) {   // cycle counters of the instrumented GEMM (option tc_debug & 32); read-and-clear
    const int i = atoi(key + 7);
    if (i < 1 || i > 15 || !c->d_fallbacks.p) SGMM_FAIL(c, "get_counter: tc_prof index out of range");
    unsigned long long v = 0;
    SGMM_CHECK(c, cudaMemcpyAsync(&v, c->d_fallbacks.p + i, sizeof(v), cudaMemcpyDeviceToHost, c->stream));
    SGMM_CHECK(c, cudaMemsetAsync(c->d_fallbacks.p + i, 0, sizeof(v), c->stream));
    SGMM_CHECK(c, cudaStreamSynchronize(c->stream));
    *value = (int64_t)v;
    return 0;
  }
  SGMM_FAIL(c, std::string("get_counter: unknown key ") + key);
}

int sgmm_timer_start(sgmm_ctx* c) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  SGMM_CHECK(c, cudaEventRecord(c->tev[0], c->stream));
  return 0;
}

int sgmm_timer_stop(sgmm_ctx* c, float* ms) {
  if (!c) return 1;
  cudaSetDevice(c->device);
  SGMM_CHECK(c, cudaEventRecord(c->tev[1], c->stream));
  SGMM_CHECK(c, cudaEventSynchronize(c->tev[1]));
  SGMM_CHECK(c, cudaEventElapsedTime(ms, c->tev[0], c->tev[1]));
  return 0;
}

void* sgmm_host_alloc(size_t bytes) {
  void* p = nullptr;
  if (cudaHostAlloc(&p, bytes ? bytes : 1, cudaHostAllocDefault) != cudaSuccess) return nullptr;
  return p;
}

void sgmm_host_free(void* p) { if (p) cudaFreeHost(p); }

uint64_t sgmm_launch_count(const sgmm_ctx* c) { return c ? c->launches : 0; }

}  // extern "C"
