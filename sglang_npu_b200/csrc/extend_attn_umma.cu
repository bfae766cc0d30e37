// Paged extend / prefill attention on tcgen05 (accumulators in TMEM).
//
// Replaces extend_attention_fwd (reference python/sglang/srt/layers/attention/triton_ops/extend_attention.py:
// 302-440, kernel :42-299: 128x64 tl.dot tiles, num_stages=1); semantics follow the oracle
// TorchNativeAttnBackend._run_sdpa_forward_extend (torch_native_backend.py:27-110): query i of request b
// attends KV positions 0 .. prefix_b + i (causal) of that request, prefix positions through the page table.
//
// One CTA = 128 query rows of one (request, q head); KV walked in tiles of 128 positions:
//   S = Q K^T   tcgen05.mma kind::f16, M=128 N=128 K=16 x D/16, A = Q and B = K both K-major SWIZZLE_128B in smem,
//               fp32 S in TMEM (two 128-column buffers: QK^T of tile j+1 is issued before the softmax of tile j)
//   P = softmax tile: one THREAD per query row (TMEM lane) — row max / sum need no shuffles; exp2 domain;
//               P is written as bf16/fp16 into a K-major SWIZZLE_128B smem tile (the A operand of the next MMA)
//   O += P V    tcgen05.mma M=128 N=D K=16 x 8, B = V tile read MN-major (the [kv][d] rows as they lie in the
//               cache — no transpose pass), O accumulates in TMEM across the whole KV loop
//   lazy rescale: O (and the running sum) is rescaled in TMEM only for warps where some row's max grew by more
//               than 2^8 since the max it is normalised to; otherwise the stale max keeps being used (exact after
//               the final 1/l), so the TMEM round-trip of O disappears from the steady state.
// Warp roles: 0-1 K gather, 2-3 V gather (16-byte cp.async of 256-byte cache rows through req_to_token — any
// page_size — placed directly in the UMMA swizzle; independent 2-stage rings, page ids fetched a tile ahead),
// 4 = MMA issuer (one lane) + TMEM allocator, 8-11 softmax / correction / epilogue (TMEM lane quarter = warp % 4).
// mbarriers: k_full/k_empty and v_full/v_empty per stage, s_full per S buffer, p_full, pv_done per P buffer.
// P is double-buffered in shared memory: the softmax of tile j+1 runs while P V of tile j is still on the tensor core;
// only the (rare) lazy rescale of O waits for the previous P V.
#include <type_traits>

#include "extend_attn.cuh"
#include "sm100.cuh"

namespace b200 {

using namespace sm100;

namespace {

constexpr int kUxBM = 128;        // query rows per CTA
constexpr int kUxBN = 128;        // KV positions per tile
constexpr int kUxThreads = 384;   // 12 warps (warps 5-7 idle: keeps softmax warps on lane quarters warp % 4)
constexpr int kUxProducers = 128;
constexpr float kRescaleThreshold = 8.f;  // log2 domain

template <int D>
struct UxSmem {
  static constexpr int kHalf = kUxBN * 128;                  // one 64-element (128-byte) column block of a 128-row tile
  static constexpr int kTile = kUxBN * D * 2;                // Q, K, V tile bytes
  static constexpr int kQ = 0;
  static constexpr int kK = kTile;                           // 2 stages
  static constexpr int kV = kK + 2 * kTile;                  // 2 stages
  static constexpr int kP = kV + 2 * kTile;                  // [128 x 128] bf16
  static constexpr int kPTile = kUxBM * kUxBN * 2;            // two P buffers: softmax of tile j+1 does not wait for P V of tile j
  static constexpr int kBar = kP + 2 * kPTile;
  static constexpr int kBytes = kBar + 16 * 8 + 1024;
};

// byte offset of (row r, 16-byte chunk c of the row) inside a [rows x D] tile stored as D/64 column blocks of
// [rows x 128 B] in the SWIZZLE_128B pattern (8-row groups of 1024 B, chunk index XOR row % 8)
__device__ __forceinline__ uint32_t sw128_off(int r, int c, int half_bytes) {
  return (uint32_t)((c >> 3) * half_bytes + (r >> 3) * 1024 + (r & 7) * 128 + (((c & 7) ^ (r & 7)) << 4));
}

// Gather one [128 x D] tile of K (or V) rows (KV positions kv0 .. kv0+127) with 16-byte cp.async.  A "stream" is
// two warps; warp w of the stream owns rows 64w .. 64w+63.  D/8 lanes cover one row, so a warp instruction reads
// 32 / (D/8) whole rows.  The rows' cache slots come from two coalesced page-table loads per warp (`slots`, loaded
// ahead of time by ux_load_slots) and are handed round with shuffles — no dependent global load per row.
struct UxSlots {
  int32_t lo, hi;  // slot of row 64w + lane, 64w + 32 + lane
};

__device__ __forceinline__ UxSlots ux_load_slots(const int32_t* __restrict__ r2t_row, int kv0, int kv_end, int w, int lane) {
  UxSlots s;
  const int p0 = kv0 + 64 * w + lane;
  s.lo = p0 < kv_end ? __ldg(r2t_row + p0) : 0;
  s.hi = p0 + 32 < kv_end ? __ldg(r2t_row + p0 + 32) : 0;
  return s;
}

template <typename T, int D>
__device__ __forceinline__ void ux_load_tile(const T* __restrict__ cache, const T* __restrict__ fresh, int64_t kv_token_stride,
                                             int64_t fresh_token_stride, const UxSlots& slots, int kv0, int kv_end, int prefix,
                                             int64_t new_base_tok, int kvh, uint32_t smem_tile, int w, int lane) {
  constexpr int kChunks = D / 8;             // 16-byte chunks per row
  constexpr int kRpp = 32 / kChunks;         // rows per warp pass (2 for D=128, 4 for D=64)
  constexpr int kPeriod = 8 / kRpp;          // passes per 8-row swizzle group
  const int c = lane % kChunks;
  const int rl = lane / kChunks;
  // Everything that depends on the lane is hoisted: inside the fully unrolled loop one copy costs a shuffle (slot),
  // a 64-bit multiply-add (address), a compare (tail) and the cp.async with an immediate shared-memory offset.
  uint32_t lane_off[kPeriod];
#pragma unroll
  for (int bq = 0; bq < kPeriod; ++bq) {
    const int r7 = bq * kRpp + rl;
    lane_off[bq] = smem_tile + (uint32_t)((c >> 3) * UxSmem<D>::kHalf + (8 * w) * 1024 + r7 * 128 + (((c & 7) ^ r7) << 4));
  }
  const T* cbase = cache + kvh * D + c * 8;
  const int pos0 = kv0 + 64 * w + rl;
  if (fresh == nullptr) {
#pragma unroll
    for (int p = 0; p < 64 / kRpp; ++p) {
      const int32_t slot = __shfl_sync(0xffffffffu, (p * kRpp < 32) ? slots.lo : slots.hi, ((p * kRpp) & 31) + rl);
      const bool valid = pos0 + p * kRpp < kv_end;
      cp_async16_pred(lane_off[p % kPeriod] + (p / kPeriod) * 1024, cbase + (int64_t)slot * kv_token_stride, valid);
    }
  } else {  // save_kv_cache = False callers: positions >= prefix come from the contiguous new-token activations
    const T* fbase = fresh + kvh * D + c * 8;
#pragma unroll 4
    for (int p = 0; p < 64 / kRpp; ++p) {
      const int32_t slot = __shfl_sync(0xffffffffu, (p * kRpp < 32) ? slots.lo : slots.hi, ((p * kRpp) & 31) + rl);
      const int pos = pos0 + p * kRpp;
      const T* src = pos >= prefix ? fbase + (new_base_tok + (pos - prefix)) * fresh_token_stride
                                   : cbase + (int64_t)slot * kv_token_stride;
      cp_async16_pred(lane_off[p % kPeriod] + (p / kPeriod) * 1024, pos < kv_end ? src : cbase, pos < kv_end);
    }
  }
}

// ---- softmax row helpers: the calling thread owns one query row = one TMEM lane ----------------------------------
// MODE 0: plain scale; 1: scale + causal / tail mask; 2: logit cap (tanh) + mask.  Scores are handled in the exp2
// domain: t = s * (sm_scale * log2 e).
struct SoftmaxCtx {
  float scale_log2, cap_in, cap_out;
  int lim, kv_end;
};

template <int MODE>
__device__ __forceinline__ float ux_score(float s, const SoftmaxCtx& cx, int pos) {
  float v = s;
  if (MODE == 2) v = cx.cap_out * tanh_approx(s * cx.cap_in);
  if (MODE >= 1 && (pos > cx.lim || pos >= cx.kv_end)) v = -INFINITY;
  return v;
}

// max over the row's 128 scores of this tile (already in the exp2 domain)
template <int MODE>
__device__ __forceinline__ float ux_row_max(uint32_t t_s, const SoftmaxCtx& cx, int kv0) {
  float mx = -INFINITY;
  uint32_t ra[32], rb[32];
  tmem_ld32(t_s, ra);
  tmem_ld_wait();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    tmem_ld32(t_s + 64 * h + 32, rb);  // in flight while the previous 32 columns are reduced
#pragma unroll
    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, ux_score<MODE>(__uint_as_float(ra[i]), cx, kv0 + 64 * h + i));
    tmem_ld_wait();
    if (h == 0) tmem_ld32(t_s + 64, ra);
#pragma unroll
    for (int i = 0; i < 32; ++i) mx = fmaxf(mx, ux_score<MODE>(__uint_as_float(rb[i]), cx, kv0 + 64 * h + 32 + i));
    if (h == 0) tmem_ld_wait();
  }
  return MODE == 2 ? mx : mx * cx.scale_log2;  // sm_scale > 0: max commutes with the scaling
}

// p = exp2(t - m) for the row's 128 scores, written as T into the K-major SWIZZLE_128B P tile; returns sum(p)
template <typename T, int MODE>
__device__ __forceinline__ void ux_exp_store32(const uint32_t (&r)[32], const SoftmaxCtx& cx, int pos0, float m_use, float& sum,
                                               float& amax, uint32_t sp, int row, int c0, int half_bytes) {
  float p[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    const float v = ux_score<MODE>(__uint_as_float(r[i]), cx, pos0 + i);
    const float arg = MODE == 2 ? v - m_use : fmaf(v, cx.scale_log2, -m_use);  // masked: fma(-inf, +scale, .) = -inf -> 0
    amax = fmaxf(amax, arg);
    p[i] = fast_exp2(arg);
    sum += p[i];
  }
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const uint32_t w0 = pack2<T>(p[8 * g + 0], p[8 * g + 1]), w1 = pack2<T>(p[8 * g + 2], p[8 * g + 3]);
    const uint32_t w2 = pack2<T>(p[8 * g + 4], p[8 * g + 5]), w3 = pack2<T>(p[8 * g + 6], p[8 * g + 7]);
    const uint32_t addr = sp + sw128_off(row, (c0 >> 3) + g, half_bytes);
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(addr), "r"(w0), "r"(w1), "r"(w2), "r"(w3) : "memory");
  }
}

// amax: largest exponent argument (score - m_use) seen in the row — how far the tile's max lies above the max the
// row is normalised to (the optimistic single-pass softmax below needs it to decide whether to redo the tile)
template <typename T, int MODE>
__device__ __forceinline__ float ux_row_exp(uint32_t t_s, const SoftmaxCtx& cx, int kv0, float m_use, uint32_t sp, int row,
                                            int half_bytes, float& amax) {
  float sum = 0.f;
  amax = -INFINITY;
  uint32_t ra[32], rb[32];
  tmem_ld32(t_s, ra);
  tmem_ld_wait();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    tmem_ld32(t_s + 64 * h + 32, rb);
    ux_exp_store32<T, MODE>(ra, cx, kv0 + 64 * h, m_use, sum, amax, sp, row, 64 * h, half_bytes);
    tmem_ld_wait();
    if (h == 0) tmem_ld32(t_s + 64, ra);
    ux_exp_store32<T, MODE>(rb, cx, kv0 + 64 * h + 32, m_use, sum, amax, sp, row, 64 * h + 32, half_bytes);
    if (h == 0) tmem_ld_wait();
  }
  return sum;
}

}  // namespace

template <typename T, int D>
__global__ void __launch_bounds__(kUxThreads, 1) extend_attn_umma_kernel(const ExtendArgs a) {
  using L = UxSmem<D>;
  constexpr int kFmt = std::is_same<T, __half>::value ? 0 : 1;  // UMMA a/b format: 0 f16, 1 bf16
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBar;
  auto k_full = [&](int s) { return bar + s * 8; };
  auto k_empty = [&](int s) { return bar + (2 + s) * 8; };
  auto v_full = [&](int s) { return bar + (4 + s) * 8; };
  auto v_empty = [&](int s) { return bar + (6 + s) * 8; };
  auto s_full = [&](int b) { return bar + (8 + b) * 8; };
  // per P buffer, like pv_done: with one barrier the softmax warps (up to two tiles ahead of the MMA warp when a V tile is
  // late) could complete two phases before the MMA warp tests the first — the parity wait would then never succeed
  auto p_full = [&](int j) { return bar + ((j & 1) ? 13 : 10) * 8; };
  auto pv_done = [&](int j) { return bar + (11 + (j & 1)) * 8; };  // P V of tile j commits to barrier j & 1, phase j >> 1

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int mblk = gridDim.x - 1 - blockIdx.x;  // heaviest (longest causal span) tiles first
  const int head = blockIdx.y;
  const int b = blockIdx.z;

  pdl_trigger();
  pdl_wait();
  const bool seg_mode = a.part_o != nullptr;
  const int rb = seg_mode ? 0 : b;  // the request whose q rows and page-table row this CTA uses
  const int q_start = seg_mode ? 0 : a.qo_indptr[rb];
  const int ext_len = seg_mode ? a.seg_rows : a.qo_indptr[rb + 1] - q_start;
  const int m0 = mblk * kUxBM;
  if (m0 >= ext_len) return;  // uniform for the CTA, before any barrier / TMEM use
  // segment mode: the segment is treated as a sequence of its own (positions relative to kv_begin, page-table row
  // pointer advanced by kv_begin) that consists of prefix tokens only
  const int kv_begin = seg_mode ? b * a.seg_len : 0;
  const int prefix = seg_mode ? min(a.seg_prefix - kv_begin, a.seg_len) : a.prefix_lens[b];
  if (prefix <= 0 && seg_mode) return;
  const int kv_total = seg_mode ? prefix : prefix + ext_len;
  const int kv_end = a.causal ? min(kv_total, prefix + m0 + kUxBM) : kv_total;
  const int n_tiles = (kv_end + kUxBN - 1) / kUxBN;
  const int group = a.num_q_heads / a.num_kv_heads;
  const int kvh = head / group;

  if (warp == 4) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) {
        mbar_init(k_full(s), 2);  // one arrive per warp of the stream
        mbar_init(k_empty(s), 1);
        mbar_init(v_full(s), 2);
        mbar_init(v_empty(s), 1);
        mbar_init(s_full(s), 1);
      }
      mbar_init(p_full(0), 4);  // one arrive per softmax warp
      mbar_init(p_full(1), 4);
      mbar_init(pv_done(0), 1);
      mbar_init(pv_done(1), 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc<512>(smem_u32(&s_tmem_base));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  constexpr uint32_t kOCol = 256;  // S buffers at columns 0 and 128, O at 256

  if (warp < 4) {
    // ---- producers: warps 0-1 stream K tiles (and Q once), warps 2-3 stream V tiles; each stream runs its own
    //      2-stage ring, so K of tile j+2 is fetched as soon as Q K^T of tile j has retired ---------------------------
    const bool is_v = warp >= 2;
    const int w = warp & 1;
    const int64_t req = load_index(a.req_pool_indices, rb, a.index_is_int64);
    const int32_t* r2t_row = a.req_to_token + req * a.req_to_token_stride + kv_begin;
    const T* cache = reinterpret_cast<const T*>(is_v ? a.v_cache : a.k_cache);
    const T* fresh = reinterpret_cast<const T*>(is_v ? a.v_new : a.k_new);
    const int64_t fresh_stride = is_v ? a.vnew_token_stride : a.knew_token_stride;
    const uint32_t ring = base + (is_v ? L::kV : L::kK);
    auto full = [&](int s) { return is_v ? v_full(s) : k_full(s); };
    auto empty = [&](int s) { return is_v ? v_empty(s) : k_empty(s); };
    if (!is_v) {
      constexpr int kChunks = D / 8;
      constexpr int kRowsPerPass = 64 / kChunks;
      const int t64 = w * 32 + lane;
      const int c = t64 % kChunks;
      const int r0 = t64 / kChunks;
      const T* qb = reinterpret_cast<const T*>(a.q) + (int64_t)head * D;
#pragma unroll 4
      for (int p = 0; p < kUxBM / kRowsPerPass; ++p) {
        const int r = r0 + p * kRowsPerPass;
        const bool valid = m0 + r < ext_len;
        const T* src = valid ? qb + (int64_t)(q_start + m0 + r) * a.q_token_stride + c * 8 : qb;
        cp_async16_pred(base + L::kQ + sw128_off(r, c, L::kHalf), src, valid);
      }
    }
    UxSlots slots = ux_load_slots(r2t_row, 0, kv_end, w, lane);
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j & 1;
      while (!mbar_try_wait(empty(s), ((j >> 1) & 1) ^ 1)) __nanosleep(64);  // do not steal issue slots from the softmax warp
      ux_load_tile<T, D>(cache, fresh, a.kv_token_stride, fresh_stride, slots, j * kUxBN, kv_end, prefix, q_start, kvh,
                         ring + s * L::kTile, w, lane);
      cp_async_commit();
      if (j + 1 < n_tiles) slots = ux_load_slots(r2t_row, (j + 1) * kUxBN, kv_end, w, lane);  // page ids of the next tile
      if (j > 0) {
        // tile j-1 has landed once at most one group is outstanding.  It is published only now, AFTER tile j was
        // issued (copies of consecutive tiles overlap); the stage tile j+1 needs is freed by an MMA that depends
        // on tile j-1, which is public by the time this loop waits for it — no cycle.
        cp_async_wait<1>();
        fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(full((j - 1) & 1));
      }
    }
    cp_async_wait<0>();
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(full((n_tiles - 1) & 1));
  } else if (warp == 4) {
    // ---- MMA issuer --------------------------------------------------------------------------------------------
    if (elect_one()) {
      constexpr uint32_t idesc_s = make_idesc(/*f32*/ 1, kFmt, kFmt, 128, kUxBN);
      constexpr uint32_t idesc_o = make_idesc(/*f32*/ 1, kFmt, kFmt, 128, D, /*A K-major*/ 0, /*B MN-major*/ 1);
      auto issue_qk = [&](int j) {
        const int s = j & 1;
        mbar_wait(k_full(s), (j >> 1) & 1);
        tc_fence_after();
        const uint32_t sq = base + L::kQ;
        const uint32_t sk = base + L::kK + s * L::kTile;
        const uint32_t d_tmem = tmem_base + (uint32_t)(j & 1) * kUxBN;
#pragma unroll
        for (int kk = 0; kk < D / 16; ++kk) {
          const uint64_t da = smem_desc_k_sw128(sq + (kk >> 2) * L::kHalf) + 2 * (kk & 3);
          const uint64_t db = smem_desc_k_sw128(sk + (kk >> 2) * L::kHalf) + 2 * (kk & 3);
          umma<kKindF16>(d_tmem, da, db, idesc_s, kk != 0);
        }
        tc_commit(k_empty(s));  // the K stage is free as soon as this Q K^T retires
        tc_commit(s_full(j & 1));
      };
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_qk(j + 1);
        mbar_wait(v_full(j & 1), (j >> 1) & 1);
        mbar_wait(p_full(j), (j >> 1) & 1);
        tc_fence_after();
        const uint32_t sp = base + L::kP + (j & 1) * L::kPTile;
        const uint32_t sv = base + L::kV + (j & 1) * L::kTile;
#pragma unroll
        for (int kk = 0; kk < kUxBN / 16; ++kk) {
          const uint64_t da = smem_desc_k_sw128(sp + (kk >> 2) * L::kHalf) + 2 * (kk & 3);
          // V tile: atoms of [8 kv x 64 d]; next 64 d = kHalf bytes (LBO), next 8 kv = 1024 bytes (SBO)
          const uint64_t db = smem_desc_mn_sw128(sv + kk * 2048, L::kHalf, 1024);
          umma<kKindF16>(tmem_base + kOCol, da, db, idesc_o, (j | kk) != 0);
        }
        tc_commit(v_empty(j & 1));
        tc_commit(pv_done(j));
      }
    }
  } else if (warp >= 8) {
    // ---- softmax / correction / epilogue: thread = query row = TMEM lane ---------------------------------------
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    const int lim = a.causal ? prefix + m0 + row : kv_total - 1;  // last KV position this row may see
    const bool use_cap = a.logit_cap > 0.f;
    float m_used = -INFINITY;  // the max (log2 domain) the row's P and O are currently normalised to
    float l_run = 0.f;
    const SoftmaxCtx cx{a.scale_log2, use_cap ? a.sm_scale / a.logit_cap : 0.f, a.logit_cap * kLog2eExt, lim, kv_end};

    for (int j = 0; j < n_tiles; ++j) {
      const int kv0 = j * kUxBN;
      const uint32_t t_s = t_lane + (uint32_t)(j & 1) * kUxBN;
      // tiles that touch the causal diagonal or the end of the sequence need per-element masks (CTA-uniform test)
      const bool need_mask = (kv0 + kUxBN > kv_end) || (a.causal && kv0 + kUxBN - 1 > prefix + m0);
      const int mode = use_cap ? 2 : (need_mask ? 1 : 0);  // warp-uniform: the hot loops carry no per-element branches
      mbar_wait(s_full(j & 1), (j >> 1) & 1);
      tc_fence_after();

      const uint32_t sp = base + L::kP + (j & 1) * L::kPTile;
      if (j >= 2) {  // P V of tile j-2 has retired: this P buffer is free again
        mbar_wait(pv_done(j), ((j >> 1) - 1) & 1);
        tc_fence_after();
      }
      auto row_exp = [&](float m_use, float& amax) {
        if (mode == 0) return ux_row_exp<T, 0>(t_s, cx, kv0, m_use, sp, row, L::kHalf, amax);
        if (mode == 1) return ux_row_exp<T, 1>(t_s, cx, kv0, m_use, sp, row, L::kHalf, amax);
        return ux_row_exp<T, 2>(t_s, cx, kv0, m_use, sp, row, L::kHalf, amax);
      };
      float l_add, amax;
      if (j == 0) {  // first tile: exact row max, then the exponentials
        if (mode == 0) m_used = ux_row_max<0>(t_s, cx, kv0);
        else if (mode == 1) m_used = ux_row_max<1>(t_s, cx, kv0);
        else m_used = ux_row_max<2>(t_s, cx, kv0);
        l_add = row_exp((m_used == -INFINITY) ? 0.f : m_used, amax);
      } else {
        // Optimistic single pass: exponentials against the max the row is already normalised to, tracking how far the
        // tile's own max lies above it.  Within the lazy-rescale threshold (2^8 of headroom in P and l) the stale max
        // simply stays in use — no separate max pass over S, half the TMEM reads.  Beyond it (rare after the first
        // tiles) the warp rescales O and l in TMEM and redoes the tile against the new max.
        const float m_use = (m_used == -INFINITY) ? 0.f : m_used;
        l_add = row_exp(m_use, amax);
        const bool grow = (m_used == -INFINITY) ? (amax > -INFINITY) : (amax > kRescaleThreshold);
        if (__any_sync(0xffffffffu, grow)) {
          mbar_wait(pv_done(j - 1), ((j - 1) >> 1) & 1);  // O is touched only here: P V of the previous tile must be done
          tc_fence_after();
          const float m_new = m_use + amax;
          const float alpha = grow ? fast_exp2(m_used - m_new) : 1.f;
          if (grow) m_used = m_new;
          l_run *= alpha;
#pragma unroll 1
          for (int c0 = 0; c0 < D; c0 += 16) {
            uint32_t r[16];
            tmem_ld16(t_lane + kOCol + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st16(t_lane + kOCol + c0, r);
          }
          tmem_st_wait();
          l_add = row_exp((m_used == -INFINITY) ? 0.f : m_used, amax);  // rows that did not grow rewrite the same P
        }
      }
      l_run += l_add;
      fence_proxy_async();  // P (generic-proxy stores) -> visible to the MMA's async-proxy reads
      tc_fence_before();    // orders this thread's TMEM loads / stores before the MMA warp's next tcgen05 ops
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full(j));
    }

    // ---- epilogue: O / l -> global ----------------------------------------------------------------------------------
    mbar_wait(pv_done(n_tiles - 1), ((n_tiles - 1) >> 1) & 1);
    tc_fence_after();
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    const bool row_ok = m0 + row < ext_len;
    if (seg_mode) {  // fp32 partial + log2-domain LSE into the decode merge's workspace
      const int64_t pidx = ((int64_t)(q_start + m0 + row) * a.num_q_heads + head) * a.part_stride + b;
      float* prow = a.part_o + pidx * D;
      if (row_ok) a.part_lse[pidx] = l_run > 0.f ? m_used + log2f(l_run) : -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < D; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(t_lane + kOCol + c0, r);
        tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int i = 0; i < 16; i += 4)
            *reinterpret_cast<float4*>(prow + c0 + i) =
                make_float4(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv,
                            __uint_as_float(r[i + 2]) * inv, __uint_as_float(r[i + 3]) * inv);
        }
      }
    } else {
    T* orow = reinterpret_cast<T*>(a.o) + (int64_t)(q_start + m0 + row) * a.o_token_stride + (int64_t)head * D;
#pragma unroll 1
    for (int c0 = 0; c0 < D; c0 += 16) {
      uint32_t r[16];
      tmem_ld16(t_lane + kOCol + c0, r);
      tmem_ld_wait();
      if (row_ok) {
        uint32_t w[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) w[i] = pack2<T>(__uint_as_float(r[2 * i]) * inv, __uint_as_float(r[2 * i + 1]) * inv);
        if ((reinterpret_cast<uintptr_t>(orow) & 15) == 0) {
          *reinterpret_cast<uint4*>(orow + c0) = make_uint4(w[0], w[1], w[2], w[3]);
          *reinterpret_cast<uint4*>(orow + c0 + 8) = make_uint4(w[4], w[5], w[6], w[7]);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) *reinterpret_cast<uint32_t*>(orow + c0 + 2 * i) = w[i];
        }
      }
    }
    }  // !seg_mode
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<512>(tmem_base);
}

template <typename T, int D>
static int launch_ux(const ExtendArgs& a, int max_extend_len, cudaStream_t stream) {
  using L = UxSmem<D>;
  auto kern = extend_attn_umma_kernel<T, D>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  dim3 grid((max_extend_len + kUxBM - 1) / kUxBM, a.num_q_heads, a.batch);  // segment mode: batch = segments
  return launch_k(kern, grid, dim3(kUxThreads), L::kBytes, stream, 1, true, a);
}

int launch_extend_umma(const ExtendArgs& a, int max_extend_len, int head_dim, int dtype, cudaStream_t stream) {
  if (dtype == B200_BF16) {
    if (head_dim == 128) return launch_ux<__nv_bfloat16, 128>(a, max_extend_len, stream);
    if (head_dim == 64) return launch_ux<__nv_bfloat16, 64>(a, max_extend_len, stream);
  } else if (dtype == B200_F16) {
    if (head_dim == 128) return launch_ux<__half, 128>(a, max_extend_len, stream);
    if (head_dim == 64) return launch_ux<__half, 64>(a, max_extend_len, stream);
  }
  return fail(B200_ERR_UNSUPPORTED, "extend_attn (tcgen05): head_dim %d / dtype %d not supported", head_dim, dtype);
}

}  // namespace b200
