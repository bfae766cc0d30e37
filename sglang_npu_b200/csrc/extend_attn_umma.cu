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
// Warp roles: 0-3 KV gather (16-byte cp.async of 256-byte cache rows through req_to_token — any page_size —
// placed directly in the UMMA swizzle), 4 = MMA issuer (one lane) + TMEM allocator, 8-11 softmax / correction /
// epilogue (TMEM lane quarter = warp % 4).  mbarriers: kv_full/kv_empty per stage, s_full per S buffer,
// p_full, pv_done.
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
  static constexpr int kBar = kP + kUxBM * kUxBN * 2;
  static constexpr int kBytes = kBar + 16 * 8 + 1024;
};

// byte offset of (row r, 16-byte chunk c of the row) inside a [rows x D] tile stored as D/64 column blocks of
// [rows x 128 B] in the SWIZZLE_128B pattern (8-row groups of 1024 B, chunk index XOR row % 8)
__device__ __forceinline__ uint32_t sw128_off(int r, int c, int half_bytes) {
  return (uint32_t)((c >> 3) * half_bytes + (r >> 3) * 1024 + (r & 7) * 128 + (((c & 7) ^ (r & 7)) << 4));
}

// Gather one K tile and one V tile (KV positions kv0 .. kv0+127) with 16-byte cp.async: 16 lanes cover one
// 256-byte row, so a warp instruction reads two whole rows.
template <typename T, int D>
__device__ __forceinline__ void ux_load_kv(const ExtendArgs& a, const int32_t* __restrict__ r2t_row, int kv0, int kv_end,
                                           int prefix, int64_t new_base_tok, int kvh, uint32_t smem_k, uint32_t smem_v,
                                           int tid) {
  constexpr int kChunks = D / 8;
  constexpr int kRowsPerPass = kUxProducers / kChunks;
  const int c = tid % kChunks;
  const int r0 = tid / kChunks;
  const T* k_cache = reinterpret_cast<const T*>(a.k_cache);
  const T* v_cache = reinterpret_cast<const T*>(a.v_cache);
  const T* k_new = reinterpret_cast<const T*>(a.k_new);
  const T* v_new = reinterpret_cast<const T*>(a.v_new);
#pragma unroll 4
  for (int p = 0; p < kUxBN / kRowsPerPass; ++p) {
    const int r = r0 + p * kRowsPerPass;
    const int pos = kv0 + r;
    const bool valid = pos < kv_end;
    const T* ksrc = k_cache;
    const T* vsrc = v_cache;
    if (valid) {
      if (k_new != nullptr && pos >= prefix) {
        const int64_t tok = new_base_tok + (pos - prefix);
        ksrc = k_new + tok * a.knew_token_stride + kvh * D + c * 8;
        vsrc = v_new + tok * a.vnew_token_stride + kvh * D + c * 8;
      } else {
        const int64_t off = (int64_t)__ldg(r2t_row + pos) * a.kv_token_stride + kvh * D + c * 8;
        ksrc = k_cache + off;
        vsrc = v_cache + off;
      }
    }
    const uint32_t so = sw128_off(r, c, UxSmem<D>::kHalf);
    cp_async16_pred(smem_k + so, ksrc, valid);  // rows past kv_end are zero-filled (and masked in the softmax)
    cp_async16_pred(smem_v + so, vsrc, valid);
  }
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
  auto kv_full = [&](int s) { return bar + s * 8; };
  auto kv_empty = [&](int s) { return bar + (2 + s) * 8; };
  auto s_full = [&](int b) { return bar + (4 + b) * 8; };
  const uint32_t p_full = bar + 6 * 8;
  const uint32_t pv_done = bar + 7 * 8;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int mblk = gridDim.x - 1 - blockIdx.x;  // heaviest (longest causal span) tiles first
  const int head = blockIdx.y;
  const int b = blockIdx.z;

  pdl_trigger();
  pdl_wait();
  const int q_start = a.qo_indptr[b];
  const int ext_len = a.qo_indptr[b + 1] - q_start;
  const int m0 = mblk * kUxBM;
  if (m0 >= ext_len) return;  // uniform for the CTA, before any barrier / TMEM use
  const int prefix = a.prefix_lens[b];
  const int kv_total = prefix + ext_len;
  const int kv_end = a.causal ? min(kv_total, prefix + m0 + kUxBM) : kv_total;
  const int n_tiles = (kv_end + kUxBN - 1) / kUxBN;
  const int group = a.num_q_heads / a.num_kv_heads;
  const int kvh = head / group;

  if (warp == 4) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) {
        mbar_init(kv_full(s), 4);  // one arrive per producer warp
        mbar_init(kv_empty(s), 1);
        mbar_init(s_full(s), 1);
      }
      mbar_init(p_full, 4);  // one arrive per softmax warp
      mbar_init(pv_done, 1);
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
    // ---- producers: Q once, then K/V tiles through a 2-stage ring -----------------------------------------
    const int tid = threadIdx.x;
    const int64_t req = load_index(a.req_pool_indices, b, a.index_is_int64);
    const int32_t* r2t_row = a.req_to_token + req * a.req_to_token_stride;
    {
      constexpr int kChunks = D / 8;
      constexpr int kRowsPerPass = kUxProducers / kChunks;
      const int c = tid % kChunks;
      const int r0 = tid / kChunks;
      const T* qb = reinterpret_cast<const T*>(a.q) + (int64_t)head * D;
#pragma unroll 4
      for (int p = 0; p < kUxBM / kRowsPerPass; ++p) {
        const int r = r0 + p * kRowsPerPass;
        const bool valid = m0 + r < ext_len;
        const T* src = valid ? qb + (int64_t)(q_start + m0 + r) * a.q_token_stride + c * 8 : qb;
        cp_async16_pred(base + L::kQ + sw128_off(r, c, L::kHalf), src, valid);
      }
    }
    for (int j = 0; j < n_tiles; ++j) {
      const int s = j & 1;
      if (j > 0) {
        // publish tile j-1 BEFORE waiting for a free stage: the MMA warp issues Q K^T of tile j-1 ahead of the
        // P V that frees the stage of tile j
        cp_async_wait<0>();
        fence_proxy_async();  // generic-proxy writes -> visible to the tensor core's async-proxy reads
        __syncwarp();
        if (lane == 0) mbar_arrive(kv_full((j - 1) & 1));
      }
      mbar_wait(kv_empty(s), ((j >> 1) & 1) ^ 1);
      ux_load_kv<T, D>(a, r2t_row, j * kUxBN, kv_end, prefix, q_start, kvh, base + L::kK + s * L::kTile,
                       base + L::kV + s * L::kTile, tid);
      cp_async_commit();
    }
    cp_async_wait<0>();
    fence_proxy_async();
    __syncwarp();
    if (lane == 0) mbar_arrive(kv_full((n_tiles - 1) & 1));
  } else if (warp == 4) {
    // ---- MMA issuer --------------------------------------------------------------------------------------------
    if (lane == 0) {
      constexpr uint32_t idesc_s = make_idesc(/*f32*/ 1, kFmt, kFmt, 128, kUxBN);
      constexpr uint32_t idesc_o = make_idesc(/*f32*/ 1, kFmt, kFmt, 128, D, /*A K-major*/ 0, /*B MN-major*/ 1);
      auto issue_qk = [&](int j) {
        const int s = j & 1;
        mbar_wait(kv_full(s), (j >> 1) & 1);
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
        tc_commit(s_full(j & 1));
      };
      issue_qk(0);
      for (int j = 0; j < n_tiles; ++j) {
        if (j + 1 < n_tiles) issue_qk(j + 1);
        mbar_wait(p_full, j & 1);
        tc_fence_after();
        const uint32_t sp = base + L::kP;
        const uint32_t sv = base + L::kV + (j & 1) * L::kTile;
#pragma unroll
        for (int kk = 0; kk < kUxBN / 16; ++kk) {
          const uint64_t da = smem_desc_k_sw128(sp + (kk >> 2) * L::kHalf) + 2 * (kk & 3);
          // V tile: atoms of [8 kv x 64 d]; next 64 d = kHalf bytes (LBO), next 8 kv = 1024 bytes (SBO)
          const uint64_t db = smem_desc_mn_sw128(sv + kk * 2048, L::kHalf, 1024);
          umma<kKindF16>(tmem_base + kOCol, da, db, idesc_o, (j | kk) != 0);
        }
        tc_commit(kv_empty(j & 1));  // K and V of this stage are free once these MMAs retire
        tc_commit(pv_done);
      }
    }
  } else if (warp >= 8) {
    // ---- softmax / correction / epilogue: thread = query row = TMEM lane ---------------------------------------
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    const int lim = a.causal ? prefix + m0 + row : kv_total - 1;  // last KV position this row may see
    const bool use_cap = a.logit_cap > 0.f;
    const float cap_in = use_cap ? a.sm_scale / a.logit_cap : 0.f;
    const float cap_out = a.logit_cap * kLog2eExt;
    const uint32_t sp = base + L::kP;
    float m_used = -INFINITY;  // the max (log2 domain) the row's P and O are currently normalised to
    float l_run = 0.f;

    for (int j = 0; j < n_tiles; ++j) {
      const int kv0 = j * kUxBN;
      const uint32_t t_s = t_lane + (uint32_t)(j & 1) * kUxBN;
      // tiles that touch the causal diagonal or the end of the sequence need per-element masks (CTA-uniform test)
      const bool need_mask = (kv0 + kUxBN > kv_end) || (a.causal && kv0 + kUxBN - 1 > prefix + m0);
      mbar_wait(s_full(j & 1), (j >> 1) & 1);
      tc_fence_after();

      float mx = -INFINITY;
#pragma unroll 1
      for (int c0 = 0; c0 < kUxBN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(t_s + c0, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float v = __uint_as_float(r[i]);
          v = use_cap ? cap_out * tanh_approx(v * cap_in) : v * a.scale_log2;
          if (need_mask) {
            const int pos = kv0 + c0 + i;
            if (pos > lim || pos >= kv_end) v = -INFINITY;
          }
          mx = fmaxf(mx, v);
        }
      }
      const float m_new = fmaxf(m_used, mx);
      if (j == 0) {
        m_used = m_new;
      } else {
        mbar_wait(pv_done, (j - 1) & 1);  // P V of the previous tile has retired: P smem and O are ours again
        tc_fence_after();
        const bool grow = m_new > m_used + kRescaleThreshold;
        if (__any_sync(0xffffffffu, grow)) {
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
        }
      }
      const float m_use = (m_used == -INFINITY) ? 0.f : m_used;

#pragma unroll 1
      for (int c0 = 0; c0 < kUxBN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(t_s + c0, r);
        tmem_ld_wait();
        float p[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float v = __uint_as_float(r[i]);
          v = use_cap ? cap_out * tanh_approx(v * cap_in) : v * a.scale_log2;
          if (need_mask) {
            const int pos = kv0 + c0 + i;
            if (pos > lim || pos >= kv_end) v = -INFINITY;
          }
          p[i] = fast_exp2(v - m_use);
          l_run += p[i];
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          w.x = pack2<T>(p[8 * g + 0], p[8 * g + 1]);
          w.y = pack2<T>(p[8 * g + 2], p[8 * g + 3]);
          w.z = pack2<T>(p[8 * g + 4], p[8 * g + 5]);
          w.w = pack2<T>(p[8 * g + 6], p[8 * g + 7]);
          const uint32_t addr = sp + sw128_off(row, (c0 >> 3) + g, L::kHalf);
          asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(addr), "r"(w.x), "r"(w.y), "r"(w.z), "r"(w.w) : "memory");
        }
      }
      fence_proxy_async();  // P (generic-proxy stores) -> visible to the MMA's async-proxy reads
      tc_fence_before();    // orders this thread's TMEM loads / stores before the MMA warp's next tcgen05 ops
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full);
    }

    // ---- epilogue: O / l -> global ----------------------------------------------------------------------------------
    mbar_wait(pv_done, (n_tiles - 1) & 1);
    tc_fence_after();
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
    const bool row_ok = m0 + row < ext_len;
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
  dim3 grid((max_extend_len + kUxBM - 1) / kUxBM, a.num_q_heads, a.batch);
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
