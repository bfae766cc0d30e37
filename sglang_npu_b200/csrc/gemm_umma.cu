// Quantised GEMMs on tcgen05 (accumulators in TMEM), TMA-fed: W8A8-int8, per-channel FP8, block-scaled FP8.
//
// Replaces (reference sgl-kernel/csrc/gemm/):
//   int8_scaled_mm          int8_gemm_kernel.cu:663-746        (NO sm_100 build in the reference, :742-743)
//   fp8_scaled_mm           fp8_gemm_kernel.cu:1110-1145
//   fp8_blockwise_scaled_mm fp8_blockwise_gemm_kernel.cu:345-423 (1-SM CUTLASS blockwise on sm100)
// Semantics follow the oracle restatements (oracle/quant_oracle.py; reference test/srt/cpu/utils.py:91-115,
// python/sglang/test/test_block_fp8.py:274-330, sgl-kernel/tests/test_fp8_blockwise_gemm.py:21-61).
//
// Both operands are K-major byte matrices (activations [M,K], weights [N,K]), so either can sit on the
// 128-row MMA "M" side:
//   * SWAP (decode, M <= 256 tokens): weights on MMA-M (128 output channels per CTA), tokens on MMA-N
//     (BN = 16..256).  Weight streaming: every weight byte is pulled from HBM exactly once by TMA into a
//     4-stage 128B-swizzled ring.  Split-K runs as a THREAD-BLOCK CLUSTER (1,1,S): the S CTAs of a
//     cluster own K-slices of the same 128-channel tile and reduce-scatter their partial accumulators
//     through distributed shared memory (rank r receives everyone's partials for token columns
//     [r*BN/S, (r+1)*BN/S), sums them in rank order and applies the scales) — no workspace, no atomics,
//     bit-deterministic.
//   * NORMAL (prefill): tokens on MMA-M (128 per CTA), weights on MMA-N (BN = 128/256 channels).
// Block-scaled FP8 accumulates each 128-wide K block into a double-buffered TMEM accumulator and the
// epilogue warps promote it into fp32 registers with sa[m,kb] * sb[n/128,kb] while the next block's MMAs run.
// Warp roles: warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one lane), warp 2 = TMEM allocator,
// warps 4-7 = epilogue (TMEM lane quarter = warp % 4).
#include <cstdlib>
#include <unordered_map>

#include "gemm_common.cuh"

namespace b200 {

using namespace sm100;

template <int BN, int STAGES>
struct GemmSmemLayout {
  static constexpr int kA = 128 * kBK;
  static constexpr int kB = BN * kBK;
  static constexpr int kStage = kA + kB;
  static constexpr int kBarOff = STAGES * kStage;
  static constexpr int kNumBars = 2 * STAGES + 4;
  static constexpr int kBytes = kBarOff + kNumBars * 8 + 1024;  // + alignment slack
  static_assert(BN * 128 * 4 <= STAGES * kStage, "split-K receive buffer must fit in the pipeline stages");
};

template <int MODE>
__device__ __forceinline__ float finalize(uint32_t bits, float sm, float sn, bool has_bias, float bias) {
  float v;
  if constexpr (MODE == kModeI8) v = __fmul_rn(__fmul_rn(sm, (float)(int32_t)bits), sn);  // (As * C) * Bs
  else if constexpr (MODE == kModeF8Chan) v = __fmul_rn(__fmul_rn(__uint_as_float(bits), sm), sn);
  else v = __uint_as_float(bits);  // block-FP8 (already promoted) and the un-scaled 16-bit modes
  return has_bias ? __fadd_rn(v, bias) : v;
}

// 16 consecutive token columns of one output channel: scales (+bias), one rounding to the output dtype, 2-byte
// stores N elements apart (a warp covers 64 contiguous bytes per token).  Branch-free; per-token scales come from
// shared memory.  p / ap point at (token c0, channel n).
template <int MODE>
__device__ __forceinline__ void finalize_store16(const uint32_t (&r)[16], const float* s_scale, float sn, bool has_bias,
                                                 float bias, bool bf, uint16_t* p, int32_t* ap, int64_t N, int m_left,
                                                 bool n_ok) {
  float sc[16];
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float4 t = *reinterpret_cast<const float4*>(s_scale + j);  // s_scale + c0 is 64-byte aligned
    sc[j] = t.x; sc[j + 1] = t.y; sc[j + 2] = t.z; sc[j + 3] = t.w;
  }
  uint16_t h[16];
  if (bf) {
#pragma unroll
    for (int j = 0; j < 16; ++j) h[j] = __bfloat16_as_ushort(__float2bfloat16_rn(finalize<MODE>(r[j], sc[j], sn, has_bias, bias)));
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) h[j] = __half_as_ushort(__float2half_rn(finalize<MODE>(r[j], sc[j], sn, has_bias, bias)));
  }
  if (!n_ok) return;
  if (m_left >= 16) {  // full group of tokens: straight-line stores
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      *p = h[j];
      p += N;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (j < m_left) *p = h[j];
      p += N;
    }
  }
  if (ap != nullptr) {
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      if (j < m_left) *ap = (int32_t)r[j];
      ap += N;
    }
  }
}

// CTAs per SM the shared-memory footprint allows (the decode tiles are small enough for 2: twice the bytes
// in flight per SM and no half-empty second wave when the tile count is not a multiple of the SM count).
template <int MODE, int BN, int STAGES>
constexpr int gemm_ctas_per_sm() {
  return (MODE != kModeF8Block && GemmSmemLayout<BN, STAGES>::kBytes <= 112 * 1024) ? 2 : 1;
}

template <int MODE, int BN, int STAGES, bool SWAP>
__global__ void __launch_bounds__(kGemmThreads, gemm_ctas_per_sm<MODE, BN, STAGES>())
gemm_umma_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const GemmArgs g) {
  using L = GemmSmemLayout<BN, STAGES>;
  constexpr bool kBlock = MODE == kModeF8Block;
  constexpr int kAccBufs = kBlock ? 2 : 1;
  constexpr int kTmemCols = (kAccBufs * BN) < 32 ? 32 : (kAccBufs * BN);
  constexpr int kKind = MODE == kModeI8 ? kKindI8 : kKindF8;
  static_assert(kTmemCols <= 512 && (kTmemCols & (kTmemCols - 1)) == 0, "TMEM columns must be a power of two <= 512");
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (STAGES + s) * 8; };
  auto tmem_full_bar = [&](int b) { return bar + (2 * STAGES + b) * 8; };
  auto tmem_empty_bar = [&](int b) { return bar + (2 * STAGES + 2 + b) * 8; };

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  unsigned long long* const tr = trace_slot(g);
  if (threadIdx.x == 0) {
    trace_mark(tr, 0);
    trace_mark(tr, 14);
  }

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar(b), 1);
      mbar_init(tmem_empty_bar(b), 4);  // one arrive per epilogue warp
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<kTmemCols>(smem_u32(&s_tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();  // barriers / TMEM are set up: the next kernel of the stream may start its own prologue
  if (threadIdx.x == 0) trace_mark(tr, 1);

  const int a_row0 = blockIdx.x * 128;
  const int b_row0 = blockIdx.y * BN;
  const int rank = SWAP ? (int)blockIdx.z : 0;  // cluster dims are (1,1,splits) and gridDim.z == splits
  const int kb0 = rank * g.kb_per_split;
  const int nkb = min(g.kb_per_split, g.num_kb - kb0);
  const int S = SWAP ? g.splits : 1;

  // epilogue state that lives across the phases below
  const int q = warp & 3;
  const int row = q * 32 + lane;  // accumulator row (TMEM lane) owned by an epilogue thread
  const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
  float facc[kBlock ? BN : 1];

  if (warp == 0) {
    if (elect_one()) {
      // weights are read once (evict-first); activations are re-read by every CTA (evict-last)
      const uint64_t pol_a = SWAP ? kEvictFirst : kEvictLast;
      const uint64_t pol_b = SWAP ? kEvictLast : kEvictFirst;
      // The weight operand does not depend on the previous kernel: fill the whole ring with weight tiles
      // BEFORE waiting on it, so weight streaming overlaps the predecessor's tail (PDL, common.cuh).
      const int pre = min(nkb, STAGES);
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(full_bar(i), L::kStage);
        const uint32_t sa = base + i * L::kStage;
        if (SWAP) tma_load_2d(sa, &tmap_a, full_bar(i), (kb0 + i) * kBK, a_row0, pol_a);
        else tma_load_2d(sa + L::kA, &tmap_b, full_bar(i), (kb0 + i) * kBK, b_row0, pol_b);
      }
      pdl_wait();  // activations (and their scales) are written by the previous kernel
      trace_mark(tr, 2);
      for (int i = 0; i < pre; ++i) {
        const uint32_t sa = base + i * L::kStage;
        if (SWAP) tma_load_2d(sa + L::kA, &tmap_b, full_bar(i), (kb0 + i) * kBK, b_row0, pol_b);
        else tma_load_2d(sa, &tmap_a, full_bar(i), (kb0 + i) * kBK, a_row0, pol_a);
      }
      for (int i = pre; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(empty_bar(s), ph ^ 1);
        mbar_expect_tx(full_bar(s), L::kStage);
        const uint32_t sa = base + s * L::kStage;
        tma_load_2d(sa, &tmap_a, full_bar(s), (kb0 + i) * kBK, a_row0, pol_a);
        tma_load_2d(sa + L::kA, &tmap_b, full_bar(s), (kb0 + i) * kBK, b_row0, pol_b);
      }
    } else {
      pdl_wait();
    }
  } else if (warp == 1) {
    pdl_wait();
    if (elect_one()) {
      constexpr uint32_t idesc = MODE == kModeI8 ? make_idesc(/*s32*/ 2, /*int8*/ 1, 1, 128, BN)
                                                 : make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 128, BN);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(full_bar(s), ph);
        if (i == 0) trace_mark(tr, 3);
        if (kBlock) mbar_wait(tmem_empty_bar(i & 1), ((i >> 1) & 1) ^ 1);  // epilogue drained this buffer
        tc_fence_after();
        const uint32_t sa = base + s * L::kStage;
        const uint64_t da = smem_desc_k_sw128(sa);
        const uint64_t db = smem_desc_k_sw128(sa + L::kA);
        const uint32_t d_tmem = tmem_base + (kBlock ? (i & 1) * BN : 0);
#pragma unroll
        for (int k = 0; k < kBK / 32; ++k)  // UMMA K = 32 bytes; +2 in the (addr >> 4) field per step
          umma<kKind>(d_tmem, da + 2 * k, db + 2 * k, idesc, kBlock ? (k != 0) : ((i | k) != 0));
        tc_commit(empty_bar(s));  // frees the smem stage when these MMAs retire
        if (kBlock) tc_commit(tmem_full_bar(i & 1));
      }
      if (!kBlock) tc_commit(tmem_full_bar(0));
      trace_mark(tr, 4);
    }
  } else if (warp >= 4) {
    pdl_wait();  // scales are read (and the output written) only after the previous kernels have completed
    if constexpr (kBlock) {
      // promote every 128-wide K block into fp32 registers while the next block's MMAs run
#pragma unroll
      for (int j = 0; j < BN; ++j) facc[j] = 0.f;
      const int nblk = SWAP ? blockIdx.x : (b_row0 >> 7);
      const int m_row = a_row0 + row;  // NORMAL: this thread's token
      for (int i = 0; i < nkb; ++i) {
        const int buf = i & 1;
        mbar_wait(tmem_full_bar(buf), (i >> 1) & 1);
        tc_fence_after();
        const int kb = kb0 + i;
        const float sb = g.scale_n[nblk * g.sn_rs + kb * g.sn_cs];
        float srow = 0.f;
        if (!SWAP) srow = (m_row < g.M ? g.scale_m[m_row * g.sm_rs + kb * g.sm_cs] : 0.f) * sb;
#pragma unroll
        for (int c0 = 0; c0 < BN; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + buf * BN + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float s = srow;
            if (SWAP) {
              const int m = b_row0 + c0 + j;
              s = (m < g.M ? g.scale_m[m * g.sm_rs + kb * g.sm_cs] : 0.f) * sb;
            }
            facc[c0 + j] = fmaf(__uint_as_float(r[j]), s, facc[c0 + j]);
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tmem_empty_bar(buf));
      }
    } else {
      mbar_wait(tmem_full_bar(0), 0);
      tc_fence_after();
    }
    if (threadIdx.x == 128) trace_mark(tr, 5);
  }

  if (warp == 2 || warp == 3) pdl_wait();

  // ---- split-K: reduce-scatter the partial accumulators across the cluster through DSMEM -----------------
  if constexpr (SWAP) {
    if (S > 1) {
      __syncwarp();
      cluster_sync();  // every CTA of the cluster has retired its MMAs: pipeline smem is free to reuse
      if (threadIdx.x == 128) trace_mark(tr, 6);
      if (warp >= 4) {
        const int C = BN / S;  // token columns owned by each rank
        const int logC = 31 - __clz(C);
        const uint32_t my_slot = base + (uint32_t)(rank * C * 128 + row) * 4;  // [src = rank][col][row]
        uint32_t dst[8];
#pragma unroll
        for (int d = 0; d < 8; ++d) dst[d] = d < S ? mapa(my_slot, d) : 0;
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          if (c0 >= g.M) break;  // no tokens in these columns
          uint32_t r[16];
          if constexpr (kBlock) {
#pragma unroll
            for (int j = 0; j < 16; ++j) r[j] = 0;  // filled below (register array needs static indexing)
          } else {
            tmem_ld16(taddr + c0, r);
            tmem_ld_wait();
          }
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int col = c0 + j;
            uint32_t v = r[j];
            if constexpr (kBlock) {
              // facc is statically indexed only when c0 is a compile-time constant: select by comparison
              float f = 0.f;
#pragma unroll
              for (int cc = 0; cc < BN; cc += 16)
                if (cc == c0) f = facc[cc + j];
              v = __float_as_uint(f);
            }
            if (col < g.M) {
              const int d = col >> logC;
              const int w = col & (C - 1);
              uint32_t a = dst[0];
#pragma unroll
              for (int dd = 1; dd < 8; ++dd)
                if (dd == d) a = dst[dd];
              st_cluster_u32(a + (uint32_t)(w * 128) * 4, v);
            }
          }
        }
      }
      __syncwarp();
      if (threadIdx.x == 128) trace_mark(tr, 7);
      cluster_sync();  // all partials have landed
      if (threadIdx.x == 128) trace_mark(tr, 8);
    }
  }

  // ---- final epilogue ------------------------------------------------------------------------------------------
  if (warp >= 4) {
    const bool has_bias = g.bias != nullptr;
    if constexpr (SWAP) {
      const int n = a_row0 + row;  // output channel owned by this thread
      const bool n_ok = n < g.N;
      float sn = 0.f, bias = 0.f;
      if (n_ok) {
        if (MODE != kModeF8Block) sn = g.scale_n[n];
        if (has_bias) bias = load_bias(g.bias, n, g.out_is_bf16);
      }
      if (S > 1) {
        const int C = BN / S;
        const uint32_t* recv = reinterpret_cast<const uint32_t*>(smem_raw + (base - smem_u32(smem_raw)));
        for (int w = 0; w < C; ++w) {
          const int m = b_row0 + rank * C + w;
          if (m >= g.M) break;
          uint32_t sum;
          if constexpr (MODE == kModeI8) {
            int32_t t = 0;
            for (int src = 0; src < S; ++src) t += (int32_t)recv[(src * C + w) * 128 + row];
            sum = (uint32_t)t;
          } else {
            float t = 0.f;
            for (int src = 0; src < S; ++src) t += __uint_as_float(recv[(src * C + w) * 128 + row]);
            sum = __float_as_uint(t);
          }
          if (n_ok) {
            const int64_t idx = (int64_t)m * g.N + n;
            const float smv = MODE == kModeF8Block ? 0.f : g.scale_m[m];
            store_out(g.out, idx, finalize<MODE>(sum, smv, sn, has_bias, bias), g.out_is_bf16);
            if (MODE == kModeI8 && g.acc_out) g.acc_out[idx] = (int32_t)sum;
          }
        }
      } else {
#pragma unroll 1
        for (int c0 = 0; c0 < BN; c0 += 16) {
          if (b_row0 + c0 >= g.M) break;  // warp-uniform
          uint32_t r[16];
          if constexpr (!kBlock) {
            tmem_ld16(taddr + c0, r);
            tmem_ld_wait();
          }
          if (n_ok) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = b_row0 + c0 + j;
              uint32_t v = 0;
              if constexpr (kBlock) {
                float f = 0.f;
#pragma unroll
                for (int cc = 0; cc < BN; cc += 16)
                  if (cc == c0) f = facc[cc + j];
                v = __float_as_uint(f);
              } else {
                v = r[j];
              }
              if (m < g.M) {
                const int64_t idx = (int64_t)m * g.N + n;
                const float smv = MODE == kModeF8Block ? 0.f : g.scale_m[m];
                store_out(g.out, idx, finalize<MODE>(v, smv, sn, has_bias, bias), g.out_is_bf16);
                if (MODE == kModeI8 && g.acc_out) g.acc_out[idx] = (int32_t)v;
              }
            }
          }
        }
      }
    } else {
      const int m = a_row0 + row;  // token row owned by this thread
      const bool m_ok = m < g.M;
      const float smv = (m_ok && MODE != kModeF8Block) ? g.scale_m[m] : 0.f;
#pragma unroll
      for (int c0 = 0; c0 < BN; c0 += 16) {
        const int n0 = b_row0 + c0;
        if (n0 < g.N) {  // warp-uniform
          uint32_t r[16];
          if constexpr (kBlock) {
#pragma unroll
            for (int j = 0; j < 16; ++j) r[j] = __float_as_uint(facc[c0 + j]);
          } else {
            tmem_ld16(taddr + c0, r);
            tmem_ld_wait();
          }
          if (m_ok) {
            // N % 8 == 0 is required, so a 16-column group is either full or exactly half full
            const int valid = min(16, g.N - n0);
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              float bias = 0.f, sn = 0.f;
              if (j < valid) {
                if (MODE != kModeF8Block) sn = g.scale_n[n0 + j];
                if (has_bias) bias = load_bias(g.bias, n0 + j, g.out_is_bf16);
              }
              v[j] = finalize<MODE>(r[j], smv, sn, has_bias, bias);
            }
            uint32_t p[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
              p[j] = g.out_is_bf16 ? pack2<__nv_bfloat16>(v[2 * j], v[2 * j + 1]) : pack2<__half>(v[2 * j], v[2 * j + 1]);
            uint8_t* dstp = reinterpret_cast<uint8_t*>(g.out) + ((int64_t)m * g.N + n0) * 2;
            *reinterpret_cast<uint4*>(dstp) = make_uint4(p[0], p[1], p[2], p[3]);
            if (valid > 8) *reinterpret_cast<uint4*>(dstp + 16) = make_uint4(p[4], p[5], p[6], p[7]);
            if (MODE == kModeI8 && g.acc_out) {
              int32_t* ad = g.acc_out + (int64_t)m * g.N + n0;
#pragma unroll
              for (int j = 0; j < 16; j += 4)
                if (j < valid) *reinterpret_cast<uint4*>(ad + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
            }
          }
        }
      }
    }
  }

  if (threadIdx.x == 128) trace_mark(tr, 9);
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<kTmemCols>(tmem_base);
  if (threadIdx.x == 0) trace_mark(tr, 15);
}

// ---- prefill: persistent CTAs, double-buffered TMEM accumulators -------------------------------------------------------
// Tensor-bound orientation for M > 256 (int8 / per-channel FP8): one CTA per SM walks output tiles of 128 tokens x
// 256 channels (tile t -> n_tile = t / tiles_m, m_tile = t % tiles_m: the CTAs running at any moment share two or
// three weight tiles and the whole activation matrix stays in L2).  The 512 TMEM columns hold TWO 128x256 fp32/s32
// accumulators: while the epilogue warps drain tile j (tcgen05.ld -> scales -> bf16 -> 32-byte row segments), the
// MMA thread already accumulates tile j+1 into the other buffer and the TMA thread runs ahead through the smem ring
// without draining it at tile boundaries — no per-tile prologue / epilogue bubbles.
constexpr int kPfBN = 256;
constexpr int kPfStages = 4;

// MC: the CTAs run as pairs (cluster 1x1x2) on two m-tiles of the SAME n-tile; each CTA loads its own A tile and HALF of
// the shared B tile, multicast into both CTAs' shared memory.  Per K block a CTA then pulls 32 KB over the L2 -> SM
// fabric instead of 48 KB (a 128x256 tile at 2.6 POP/s asks ~92 B/clk per SM from the fabric).  A stage is free when BOTH
// CTAs have consumed it: the MMA threads commit to the empty barrier of both CTAs (multicast commit, count 2).
template <int MODE, bool MC>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_prefill_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const GemmArgs g, const int tiles_m, const int num_tiles) {
  using L = GemmSmemLayout<kPfBN, kPfStages>;
  constexpr int kKind = MODE == kModeI8 ? kKindI8 : kKindF8;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (kPfStages + s) * 8; };
  auto tmem_full_bar = [&](int b) { return bar + (2 * kPfStages + b) * 8; };
  auto tmem_empty_bar = [&](int b) { return bar + (2 * kPfStages + 2 + b) * 8; };
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kPfStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), MC ? 2 : 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar(b), 1);
      mbar_init(tmem_empty_bar(b), 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(smem_u32(&s_tmem_base));
  tc_fence_before();
  if constexpr (MC) cluster_sync();  // the peer's barriers exist before anything is multicast at them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();
  pdl_wait();

  const int nkb = g.num_kb;
  // tile walk: without MC, tile t of this CTA; with MC, pair-tile t = (n_tile, pair of m-tiles), this CTA takes m-tile
  // 2 * pair + rank (an odd tiles_m leaves rank 1 a tile of zero rows: TMA fills zeros, the epilogue stores nothing)
  const int rank = MC ? (int)cluster_ctarank() : 0;
  const int tiles_mp = MC ? (tiles_m + 1) / 2 : tiles_m;
  auto tile_m = [&](int t) { return MC ? 2 * (t % tiles_mp) + rank : t % tiles_mp; };
  if (warp == 0) {
    if (elect_one()) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int n_tile = t / tiles_mp;
        const int a_row0 = tile_m(t) * 128;
        const int b_row0 = n_tile * kPfBN;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kPfStages;
          mbar_wait(empty_bar(s), ((it / kPfStages) & 1) ^ 1);
          mbar_expect_tx(full_bar(s), L::kStage);
          const uint32_t sa = base + s * L::kStage;
          tma_load_2d(sa, &tmap_a, full_bar(s), i * kBK, a_row0, kEvictLast);
          if constexpr (MC)  // tmap_b boxes are 128 rows here: this CTA's half, into both CTAs
            tma_load_2d_mc(sa + L::kA + rank * (128 * kBK), &tmap_b, full_bar(s), i * kBK, b_row0 + rank * 128, (uint16_t)3,
                           kEvictNormal);
          else
            tma_load_2d(sa + L::kA, &tmap_b, full_bar(s), i * kBK, b_row0, kEvictNormal);
        }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      constexpr uint32_t idesc = MODE == kModeI8 ? make_idesc(/*s32*/ 2, /*int8*/ 1, 1, 128, kPfBN)
                                                 : make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 128, kPfBN);
      uint32_t it = 0;
      int j = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++j) {
        const int buf = j & 1;
        mbar_wait(tmem_empty_bar(buf), ((j >> 1) & 1) ^ 1);  // epilogue has drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * kPfBN;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kPfStages;
          mbar_wait(full_bar(s), (it / kPfStages) & 1);
          tc_fence_after();
          const uint32_t sa = base + s * L::kStage;
          const uint64_t da = smem_desc_k_sw128(sa);
          const uint64_t db = smem_desc_k_sw128(sa + L::kA);
#pragma unroll
          for (int k = 0; k < kBK / 32; ++k) umma<kKind>(d_tmem, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
          if constexpr (MC) tc_commit_mc(empty_bar(s), (uint16_t)3);
          else tc_commit(empty_bar(s));
        }
        tc_commit(tmem_full_bar(buf));
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const bool has_bias = g.bias != nullptr;
    int j = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++j) {
      const int buf = j & 1;
      const int n_tile = t / tiles_mp;
      const int m = tile_m(t) * 128 + row;
      const int b_row0 = n_tile * kPfBN;
      const bool m_ok = m < g.M;
      const float smv = m_ok ? g.scale_m[m] : 0.f;
      mbar_wait(tmem_full_bar(buf), (j >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * kPfBN;
#pragma unroll 2
      for (int c0 = 0; c0 < kPfBN; c0 += 16) {
        const int n0 = b_row0 + c0;
        if (n0 >= g.N) break;  // warp-uniform
        uint32_t r[16];
        tmem_ld16(taddr + c0, r);
        tmem_ld_wait();
        if (m_ok) {
          const int valid = min(16, g.N - n0);  // N % 8 == 0: a 16-column group is full or exactly half full
          float v[16];
#pragma unroll
          for (int jj = 0; jj < 16; ++jj) {
            float bias = 0.f, sn = 0.f;
            if (jj < valid) {
              sn = __ldg(g.scale_n + n0 + jj);
              if (has_bias) bias = load_bias(g.bias, n0 + jj, g.out_is_bf16);
            }
            v[jj] = finalize<MODE>(r[jj], smv, sn, has_bias, bias);
          }
          uint32_t p[8];
#pragma unroll
          for (int jj = 0; jj < 8; ++jj)
            p[jj] = g.out_is_bf16 ? pack2<__nv_bfloat16>(v[2 * jj], v[2 * jj + 1]) : pack2<__half>(v[2 * jj], v[2 * jj + 1]);
          uint8_t* dstp = reinterpret_cast<uint8_t*>(g.out) + ((int64_t)m * g.N + n0) * 2;
          *reinterpret_cast<uint4*>(dstp) = make_uint4(p[0], p[1], p[2], p[3]);
          if (valid > 8) *reinterpret_cast<uint4*>(dstp + 16) = make_uint4(p[4], p[5], p[6], p[7]);
          if (MODE == kModeI8 && g.acc_out) {
            int32_t* ad = g.acc_out + (int64_t)m * g.N + n0;
#pragma unroll
            for (int jj = 0; jj < 16; jj += 4)
              if (jj < valid) *reinterpret_cast<uint4*>(ad + jj) = make_uint4(r[jj], r[jj + 1], r[jj + 2], r[jj + 3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty_bar(buf));
    }
  }
  tc_fence_before();
  if constexpr (MC) cluster_sync();  // no CTA leaves while its peer may still multicast into it / arrive at its barriers
  else __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// ---- prefill: CTA pairs on one 256 x 256 tile (tcgen05.mma.cta_group::2) ------------------------------------------------
// The 128x256 single-CTA tile above asks ~92 B/clk per SM from the L2 -> SM fabric (A 16 KB + B 32 KB per 128-byte K
// block) and measures 15.7 TB/s of fabric traffic with the tensor pipe 73 % active and the MMA thread waiting for
// stages (profiles/r02s_gemm_prefill_ncu.txt).  Here two CTAs of a cluster (2,1,1) — two SMs of one TPC — form ONE
// 256(M) x 256(N) MMA: each CTA holds its own 128 token rows of A and HALF of the weight tile (128 of the 256
// channels); the leader's thread issues tcgen05.mma.cta_group::2, the tensor cores of both SMs read A from their own
// shared memory and B from both.  Per K block a CTA ingests 32 KB instead of 48 KB for the same MACs, and the stage
// shrinks to 32 KB: six stages instead of four in the same shared memory.
//   full[s]       lives in the LEADER: both producers arrive.expect_tx on it, both CTAs' TMA loads complete_tx on it;
//   empty[s]      in each CTA (its producer waits on it): the leader's MMA thread commits to both (multicast);
//   tmem_full[b]  in each CTA (its epilogue warps wait on it): multicast commit;
//   tmem_empty[b] in the LEADER: the epilogue warps of BOTH CTAs arrive on it (count 8).
constexpr int kP2Stages = 6;
constexpr int kP2StageBytes = 2 * 128 * kBK;  // A rows of this CTA + this CTA's half of the weight tile
constexpr int kP2BarOff = kP2Stages * kP2StageBytes;
constexpr int kP2Bytes = kP2BarOff + (2 * kP2Stages + 4) * 8 + 1024;

// Remote arrives in their default (release at CTA scope) form: `.release.cluster` compiles to MEMBAR.ALL.GPU + ERRBAR
// per call, which throttled the non-leader's producer to one stage per fence (tensor pipe 37 % active).
__device__ __forceinline__ void mbar_arrive_remote(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];\n" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_remote(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;\n" ::"r"(cluster_addr), "r"(bytes)
               : "memory");
}
// TMA tile load of a CTA pair: lands in THIS CTA's shared memory, completes bytes on the barrier at `cluster_bar`
// (a shared::cluster address: the leader's)
__device__ __forceinline__ void tma_load_2d_pair(uint32_t smem_dst, const CUtensorMap* m, uint32_t cluster_bar, int c0, int c1,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;\n" ::"r"(smem_dst),
      "l"(reinterpret_cast<uint64_t>(m)), "r"(cluster_bar), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
template <int KIND>
__device__ __forceinline__ void umma_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  const uint32_t acc = accumulate ? 1u : 0u;
  if constexpr (KIND == kKindI8) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n}\n" ::"r"(d_tmem),
                 "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
                 : "memory");
  } else {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n}\n" ::"r"(
                     d_tmem),
                 "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc)
                 : "memory");
  }
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {  // arrives on the barrier at this offset in BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_prefill_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                         const GemmArgs g, const int tiles_m2, const int num_tiles) {
  constexpr int kKind = MODE == kModeI8 ? kKindI8 : kKindF8;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + kP2BarOff;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (kP2Stages + s) * 8; };
  auto tmem_full_bar = [&](int b) { return bar + (2 * kP2Stages + b) * 8; };
  auto tmem_empty_bar = [&](int b) { return bar + (2 * kP2Stages + 2 + b) * 8; };
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int rank = (int)cluster_ctarank();

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kP2Stages; ++s) {
      mbar_init(full_bar(s), 2);   // one arrive.expect_tx per producer of the pair (used in the leader)
      mbar_init(empty_bar(s), 1);  // one multicast commit of the leader's MMA thread
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar(b), 1);
      mbar_init(tmem_empty_bar(b), 8);  // four epilogue warps of each CTA (used in the leader)
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(smem_u32(&s_tmem_base)), "n"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
  }
  tc_fence_before();
  cluster_sync();  // both CTAs' barriers and TMEM exist before anything crosses the pair
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();
  pdl_wait();

  const int nkb = g.num_kb;
  if (warp == 0) {
    if (elect_one()) {
      uint32_t it = 0;
      for (int t = blockIdx.x >> 1; t < num_tiles; t += gridDim.x >> 1) {
        const int n_tile = t / tiles_m2;
        const int a_row0 = (2 * (t - n_tile * tiles_m2) + rank) * 128;
        const int b_row0 = n_tile * 256 + rank * 128;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kP2Stages;
          mbar_wait(empty_bar(s), ((it / kP2Stages) & 1) ^ 1);
          const uint32_t lead_full = mapa(full_bar(s), 0);
          if (rank == 0) mbar_expect_tx(full_bar(s), kP2StageBytes);
          else mbar_expect_tx_remote(lead_full, kP2StageBytes);
          const uint32_t sa = base + s * kP2StageBytes;
          tma_load_2d_pair(sa, &tmap_a, lead_full, i * kBK, a_row0, kEvictLast);
          tma_load_2d_pair(sa + 128 * kBK, &tmap_b, lead_full, i * kBK, b_row0, kEvictNormal);
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      constexpr uint32_t idesc = MODE == kModeI8 ? make_idesc(/*s32*/ 2, /*int8*/ 1, 1, 256, 256)
                                                 : make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 256, 256);
      uint32_t it = 0;
      int j = 0;
      for (int t = blockIdx.x >> 1; t < num_tiles; t += gridDim.x >> 1, ++j) {
        const int buf = j & 1;
        mbar_wait(tmem_empty_bar(buf), ((j >> 1) & 1) ^ 1);  // both CTAs' epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kP2Stages;
          mbar_wait(full_bar(s), (it / kP2Stages) & 1);
          tc_fence_after();
          const uint32_t sa = base + s * kP2StageBytes;
          const uint64_t da = smem_desc_k_sw128(sa);
          const uint64_t db = smem_desc_k_sw128(sa + 128 * kBK);
#pragma unroll
          for (int k = 0; k < kBK / 32; ++k) umma_pair<kKind>(d_tmem, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
          tc_commit_pair(empty_bar(s));
        }
        tc_commit_pair(tmem_full_bar(buf));
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const bool has_bias = g.bias != nullptr;
    int j = 0;
    for (int t = blockIdx.x >> 1; t < num_tiles; t += gridDim.x >> 1, ++j) {
      const int buf = j & 1;
      const int n_tile = t / tiles_m2;
      const int m = (2 * (t - n_tile * tiles_m2) + rank) * 128 + row;
      const int b_row0 = n_tile * 256;
      const bool m_ok = m < g.M;
      const float smv = m_ok ? g.scale_m[m] : 0.f;
      mbar_wait(tmem_full_bar(buf), (j >> 1) & 1);
      tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * 256;
#pragma unroll 2
      for (int c0 = 0; c0 < 256; c0 += 16) {
        const int n0 = b_row0 + c0;
        if (n0 >= g.N) break;  // warp-uniform
        uint32_t r[16];
        tmem_ld16(taddr + c0, r);
        tmem_ld_wait();
        if (m_ok) {
          const int valid = min(16, g.N - n0);  // N % 8 == 0: a 16-column group is full or exactly half full
          float v[16];
#pragma unroll
          for (int jj = 0; jj < 16; ++jj) {
            float bias = 0.f, sn = 0.f;
            if (jj < valid) {
              sn = __ldg(g.scale_n + n0 + jj);
              if (has_bias) bias = load_bias(g.bias, n0 + jj, g.out_is_bf16);
            }
            v[jj] = finalize<MODE>(r[jj], smv, sn, has_bias, bias);
          }
          uint32_t p[8];
#pragma unroll
          for (int jj = 0; jj < 8; ++jj)
            p[jj] = g.out_is_bf16 ? pack2<__nv_bfloat16>(v[2 * jj], v[2 * jj + 1]) : pack2<__half>(v[2 * jj], v[2 * jj + 1]);
          uint8_t* dstp = reinterpret_cast<uint8_t*>(g.out) + ((int64_t)m * g.N + n0) * 2;
          *reinterpret_cast<uint4*>(dstp) = make_uint4(p[0], p[1], p[2], p[3]);
          if (valid > 8) *reinterpret_cast<uint4*>(dstp + 16) = make_uint4(p[4], p[5], p[6], p[7]);
          if (MODE == kModeI8 && g.acc_out) {
            int32_t* ad = g.acc_out + (int64_t)m * g.N + n0;
#pragma unroll
            for (int jj = 0; jj < 16; jj += 4)
              if (jj < valid) *reinterpret_cast<uint4*>(ad + jj) = make_uint4(r[jj], r[jj + 1], r[jj + 2], r[jj + 3]);
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (rank == 0) mbar_arrive(tmem_empty_bar(buf));
        else mbar_arrive_remote(mapa(tmem_empty_bar(buf), 0));
      }
    }
  }
  tc_fence_before();
  cluster_sync();  // no CTA leaves (or frees TMEM) while the pair's MMAs / barrier traffic may still touch it
  if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem_base), "n"(512) : "memory");
}

// ---- decode: persistent stream-K weight streaming -----------------------------------------------------------------------
// Weight-streaming orientation (weights on MMA-M, tokens on MMA-N, M <= 256) for int8 / per-channel FP8, balanced
// to the byte: the work is U = (128-channel tiles) x (128-byte K blocks) "units" of one 16 KiB weight box each, and
// CTA c of a one-CTA-per-SM grid streams the contiguous unit range [c*U/G, (c+1)*U/G) through a deep TMA ring, so
// every SM pulls the same number of weight bytes whatever N and K are (the tile-per-CTA grid left 72 of 148 SMs with
// half the work on gate_up).  A range is cut into segments at tile boundaries; segment j accumulates in TMEM buffer
// j & 1 and is drained by the epilogue warps while the MMA thread is already on segment j+1.
//   * a segment that covers the whole K range is finalised straight from TMEM;
//   * a segment that does not start at K block 0 (always the FIRST of its CTA) is a contribution: the raw 32-bit
//     accumulators go to this CTA's slot of the workspace and a release-flag is raised;
//   * a segment that starts at K block 0 but stops early (always the LAST of its CTA) owns the tile: it waits for
//     the flags of the following CTAs — which wrote their contributions at the very start of their own ranges —
//     and adds their partials in CTA order.  int32 sums are exact and the fp32 order is fixed, so the result is
//     bit-deterministic; the flags are reset by their consumer (self-cleaning workspace).
// All G <= #SM CTAs are co-resident (one per SM by shared-memory footprint), so the flag wait cannot deadlock.
// Token blocks (g.m_blocks = MB > 1, M <= MB * BN <= 512): the "tile" of the scheme above becomes a (weight tile, token
// block) pair — unit u = ((tile * MB + mb) * num_kb + kb) — so a batch wider than BN multiplies the units instead of
// widening them.  On 1/TP-size weight shards (TP 4: M = 256, 12 .. 56 weight tiles) one BN = 256 tile was shared by
// ~12 CTAs whose 128 KB partials the owner summed one after the other (30 us per GEMM, profiles/r02e_timeline_*_n4*);
// with BN = 64 x 4 token blocks a CTA holds ~10 units, a tile has ~3 sharers and a partial is 32 KB.  The weight
// tile is re-read once per token block, from L2 (the blocks of a tile are adjacent in unit order).
constexpr int kSkMaxM = 512;
template <int BN>
struct StreamK {
  static constexpr int kStageBytes = 128 * kBK + BN * kBK;
  static constexpr int kStages = (200 * 1024) / kStageBytes > 10 ? 10 : (200 * 1024) / kStageBytes;
  static constexpr int kBarOff = kStages * kStageBytes;
  static constexpr int kBytes = kBarOff + (2 * kStages + 5) * 8 + 1024;
  static constexpr int kTmemCols = 2 * BN < 32 ? 32 : 2 * BN;
};
template <int MODE, int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_streamk_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                    const GemmArgs g, const int total_units, uint32_t* __restrict__ flags,
                    uint32_t* __restrict__ partials) {
  using L = StreamK<BN>;
  constexpr int STAGES = L::kStages;
  constexpr bool kDense16 = MODE == kModeBF16 || MODE == kModeF16;  // un-scaled bf16 / f16 operands (lm_head)
  constexpr int kKind = MODE == kModeI8 ? kKindI8 : kDense16 ? kKindF16 : kKindF8;
  constexpr int kWBytes = 128 * kBK;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  __shared__ __align__(16) float s_scale_m[kSkMaxM];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  const uint32_t fix_bar = bar + (2 * STAGES + 4) * 8;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (STAGES + s) * 8; };
  auto tmem_full_bar = [&](int b) { return bar + (2 * STAGES + b) * 8; };
  auto tmem_empty_bar = [&](int b) { return bar + (2 * STAGES + 2 + b) * 8; };
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  unsigned long long* const tr = trace_slot(g);
  if (threadIdx.x == 0) {
    trace_mark(tr, 0);
    trace_mark(tr, 14);
  }

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar(b), 1);
      mbar_init(tmem_empty_bar(b), 4);
    }
    mbar_init(fix_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<L::kTmemCols>(smem_u32(&s_tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();
  if (threadIdx.x == 0) trace_mark(tr, 1);

  const int G = gridDim.x;
  const int c = blockIdx.x;
  const int nkb = g.num_kb;
  const int MB = g.m_blocks;
  const int u0 = (int)((int64_t)c * total_units / G);
  const int u1 = (int)((int64_t)(c + 1) * total_units / G);

  if (warp == 0) {
    if (elect_one()) {
      const int n_units = u1 - u0;
      const int pre = min(n_units, STAGES);
      int vt = u0 / nkb, kb = u0 - vt * nkb;  // vt = tile * MB + mb
      // weights do not depend on the previous kernel: fill the ring with weight boxes before the PDL wait
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(full_bar(i), L::kStageBytes);
        tma_load_2d(base + i * L::kStageBytes, &tmap_w, full_bar(i), kb * kBK, (vt / MB) * 128, MB > 1 ? kEvictNormal : kEvictFirst);
        if (++kb == nkb) { kb = 0; ++vt; }
      }
      pdl_wait();
      trace_mark(tr, 2);
      vt = u0 / nkb, kb = u0 - vt * nkb;
      for (int i = 0; i < pre; ++i) {
        tma_load_2d(base + i * L::kStageBytes + kWBytes, &tmap_x, full_bar(i), kb * kBK, (vt % MB) * BN, kEvictLast);
        if (++kb == nkb) { kb = 0; ++vt; }
      }
      for (int i = pre; i < n_units; ++i) {
        const int s = i % STAGES;
        mbar_wait(empty_bar(s), ((i / STAGES) & 1) ^ 1);
        mbar_expect_tx(full_bar(s), L::kStageBytes);
        const uint32_t sa = base + s * L::kStageBytes;
        tma_load_2d(sa, &tmap_w, full_bar(s), kb * kBK, (vt / MB) * 128, MB > 1 ? kEvictNormal : kEvictFirst);
        tma_load_2d(sa + kWBytes, &tmap_x, full_bar(s), kb * kBK, (vt % MB) * BN, kEvictLast);
        if (++kb == nkb) { kb = 0; ++vt; }
      }
    } else {
      pdl_wait();
    }
  } else if (warp == 1) {
    pdl_wait();
    if (elect_one()) {
      // kind::f16 operand formats: 0 = f16, 1 = bf16; a 128-byte K block is 64 16-bit elements = 4 UMMAs of K = 16
      constexpr uint32_t idesc = MODE == kModeI8     ? make_idesc(/*s32*/ 2, /*int8*/ 1, 1, 128, BN)
                                 : MODE == kModeBF16 ? make_idesc(/*f32*/ 1, /*bf16*/ 1, 1, 128, BN)
                                 : MODE == kModeF16  ? make_idesc(/*f32*/ 1, /*f16*/ 0, 0, 128, BN)
                                                     : make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 128, BN);
      int it = 0, seg = 0;
      for (int u = u0; u < u1; ++seg) {
        const int tile = u / nkb;
        const int seg_end = min(u1, (tile + 1) * nkb);
        const int buf = seg & 1;
        mbar_wait(tmem_empty_bar(buf), ((seg >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * BN;
        for (int i = u; i < seg_end; ++i, ++it) {
          const int s = it % STAGES;
          mbar_wait(full_bar(s), (it / STAGES) & 1);
          if (it == 0) trace_mark(tr, 3);
          tc_fence_after();
          const uint32_t sa = base + s * L::kStageBytes;
          const uint64_t da = smem_desc_k_sw128(sa);
          const uint64_t db = smem_desc_k_sw128(sa + kWBytes);
#pragma unroll
          for (int k = 0; k < kBK / 32; ++k) umma<kKind>(d_tmem, da + 2 * k, db + 2 * k, idesc, (i > u) || (k != 0));
          tc_commit(empty_bar(s));
        }
        tc_commit(tmem_full_bar(buf));
        u = seg_end;
      }
      trace_mark(tr, 4);
    }
  } else if (warp >= 4) {
    pdl_wait();
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const bool has_bias = g.bias != nullptr;
    const bool bf = g.out_is_bf16 != 0;
    // per-token scales of this launch -> shared memory (the inner loops below touch no global memory but their stores)
    for (int i = threadIdx.x - 128; i < MB * BN; i += 128) s_scale_m[i] = (!kDense16 && i < g.M) ? g.scale_m[i] : 0.f;
    epi_bar_sync();
    uint32_t fix_phase = 0;
    int seg = 0;
    for (int u = u0; u < u1; ++seg) {
      const int vt = u / nkb;  // (weight tile, token block)
      const int tile = vt / MB;
      const int t0 = (vt - tile * MB) * BN;  // first token of the block
      const int m_cols = min(BN, (g.M - t0 + 15) & ~15);  // token columns that hold data, in groups of 16
      const int kb_begin = u - vt * nkb;
      const int seg_end = min(u1, (vt + 1) * nkb);
      const bool starts = kb_begin == 0;
      const bool ends = seg_end == (vt + 1) * nkb;
      const int buf = seg & 1;
      mbar_wait(tmem_full_bar(buf), (seg >> 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 128) trace_mark(tr, seg == 0 ? 5 : 7);  // 5: first segment accumulated, 7: latest one
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN;
      if (!starts) {
        // contribution: raw accumulators -> partials[c][col][row]; a warp store covers 128 contiguous bytes
        uint32_t* dst = partials + (size_t)c * BN * 128 + row;
#pragma unroll 1
        for (int c0 = 0; c0 < m_cols; c0 += 16) {
          uint32_t r[16];
          tmem_ld16(taddr + c0, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) __stcg(dst + (c0 + j) * 128, r[j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tmem_empty_bar(buf));
        epi_bar_sync();  // orders every epilogue thread's stores before the release below (cumulativity)
        if (threadIdx.x == 128) st_release_u32(flags + c, 1u);
        if (threadIdx.x == 128) trace_mark(tr, 6);  // contribution published
      } else {
        // owner of the tile: peers are the CTAs after this one whose ranges start inside the tile
        int n_peers = 0;
        if (!ends) {
          const int tile_end = (vt + 1) * nkb;
          while (c + 1 + n_peers < G && (int)((int64_t)(c + 1 + n_peers) * total_units / G) < tile_end) ++n_peers;
          for (int p = 0; p < n_peers; ++p)
            while (ld_acquire_u32(flags + c + 1 + p) == 0u) {
            }
          if (threadIdx.x == 128) trace_mark(tr, 8);  // peers' flags seen
          if (threadIdx.x == 128 && tr != nullptr) tr[10] = (unsigned long long)n_peers;
        }
        const int n = tile * 128 + row;
        const bool n_ok = n < g.N;
        float sn = 0.f, bias = 0.f;
        if (n_ok) {
          if constexpr (!kDense16) sn = g.scale_n[n];
          if (has_bias) bias = load_bias(g.bias, n, g.out_is_bf16);
        }
        // This is the CTA's last segment when it has peers: every MMA has retired, so the TMA ring is free and the
        // peers' partials are pulled into it with bulk copies (a batch = as many peers as fit) instead of
        // latency-bound per-thread loads; the running sum of all but the last batch goes back to TMEM.
        const uint32_t pbytes = (uint32_t)m_cols * 512u;
        const int per_batch = max(1, (int)((uint32_t)(STAGES * L::kStageBytes) / pbytes));
        int p0 = 0;
        do {
          const int nb = min(per_batch, n_peers - p0);
          const bool last = p0 + nb >= n_peers;
          if (nb > 0) {
            if (threadIdx.x == 128) {
              asm volatile("fence.proxy.async.global;\n" ::: "memory");  // peers wrote with generic-proxy stores
              mbar_expect_tx(fix_bar, (uint32_t)nb * pbytes);
              for (int i = 0; i < nb; ++i)
                bulk_load_1d(base + (uint32_t)i * pbytes, partials + (size_t)(c + 1 + p0 + i) * BN * 128, pbytes, fix_bar);
            }
            mbar_wait(fix_bar, fix_phase);
            fix_phase ^= 1u;
            if (threadIdx.x == 128) trace_mark(tr, 11);  // peers' partials in shared memory
          }
          const uint32_t* sp = reinterpret_cast<const uint32_t*>(smem_raw + (base - smem_u32(smem_raw))) + row;
          uint16_t* outp = reinterpret_cast<uint16_t*>(g.out) + n;
          int32_t* accp = (MODE == kModeI8 && g.acc_out != nullptr) ? g.acc_out + n : nullptr;
#pragma unroll 1
          for (int c0 = 0; c0 < m_cols; c0 += 16) {
            uint32_t r[16];
            tmem_ld16(taddr + c0, r);
            tmem_ld_wait();
            for (int i = 0; i < nb; ++i) {
              const uint32_t* src = sp + (size_t)i * (pbytes / 4) + c0 * 128;
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                const uint32_t v = src[j * 128];
                if constexpr (MODE == kModeI8) r[j] = (uint32_t)((int32_t)r[j] + (int32_t)v);
                else r[j] = __float_as_uint(__fadd_rn(__uint_as_float(r[j]), __uint_as_float(v)));
              }
            }
            if (last) {
              finalize_store16<MODE>(r, s_scale_m + t0 + c0, sn, has_bias, bias, bf, outp + (int64_t)(t0 + c0) * g.N,
                                     accp ? accp + (int64_t)(t0 + c0) * g.N : nullptr, g.N, g.M - t0 - c0, n_ok);
            } else {
              tmem_st16(taddr + c0, r);
            }
          }
          if (!last) {
            tmem_st_wait();
            epi_bar_sync();  // everyone is done reading this batch before the next one overwrites it
          }
          p0 += nb;
        } while (p0 < n_peers);
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(tmem_empty_bar(buf));
        if (n_peers > 0) {
          epi_bar_sync();  // every epilogue thread has consumed the peers' partials
          if (threadIdx.x == 128)
            for (int p = 0; p < n_peers; ++p) flags[c + 1 + p] = 0u;
        }
      }
      u = seg_end;
    }
    if (threadIdx.x == 128) trace_mark(tr, 9);
  } else {
    pdl_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<L::kTmemCols>(tmem_base);
  if (threadIdx.x == 0) trace_mark(tr, 15);
}

// ---- host side -----------------------------------------------------------------------------------
struct TmapKey {
  const void* p;
  uint64_t rows, kbytes;
  uint32_t box;  // box_rows | box_bytes << 12 | swizzle << 20
  bool operator==(const TmapKey& o) const { return p == o.p && rows == o.rows && kbytes == o.kbytes && box == o.box; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    return std::hash<const void*>()(k.p) ^ (k.rows * 0x9E3779B97F4A7C15ull) ^ (k.kbytes << 17) ^ k.box;
  }
};

int cached_tmap(CUtensorMap* out, const void* base, uint64_t rows, uint64_t kbytes, uint32_t box_rows, uint32_t box_bytes,
                uint32_t swizzle_bytes) {
  static thread_local std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  TmapKey key{base, rows, kbytes, box_rows | (box_bytes << 12) | (swizzle_bytes << 20)};
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return B200_OK;
  }
  int st = make_tmap_2d_bytes(out, base, rows, kbytes, kbytes, box_rows, box_bytes, swizzle_bytes);
  if (st != B200_OK) return st;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return B200_OK;
}

// Cluster size for the weight-streaming orientation: the smallest power of two that gives at least one CTA
// per SM, keeping >= 4 K-blocks per CTA and at most BN/ S >= 1 token columns per rank (S <= 8 portable).
static int pick_splits(int tiles, int num_kb, int bn, int sms) {
  int s = 1;
  while (s < 8 && tiles * s < sms && num_kb / (s * 2) >= 4 && bn / (s * 2) >= 1) s *= 2;
  return s;
}

template <int MODE, int BN, int STAGES, bool SWAP>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmArgs& g, dim3 grid, cudaStream_t st) {
  using L = GemmSmemLayout<BN, STAGES>;
  auto kern = gemm_umma_kernel<MODE, BN, STAGES, SWAP>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  return launch_k(kern, grid, dim3(kGemmThreads), L::kBytes, st, SWAP ? (unsigned)g.splits : 1u, true, ta, tb, g);
}

// workspace of the stream-K decode path: [flags, kSkFlagBytes][partials: G x BN x 128 x 4 B]
static int streamk_bn(int64_t M) { return M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256; }
static size_t streamk_workspace_bytes(int64_t M) {
  if (M > kSkMaxM) return 0;
  return (size_t)kSkFlagBytes + (size_t)sm_count() * streamk_bn(M) * 128 * 4;
}
// Token-block width of the stream-K kernel for the scaled (int8 / per-channel FP8) GEMMs: up to 64 tokens one block as
// wide as the batch; wider batches are cut into 128-token blocks when that still leaves every CTA >= 16 units of
// weight stream (few sharers per tile anyway) or the batch is compute-heavy (M > 256), else into 64-token blocks.
// Measured on the Llama-3.1-8B shard shapes, layer total of the four GEMMs (profiles/r02f_kbench_int8_tokenblocks.jsonl):
// TP 4 / M 256: 114.7 us with one 256-token block, 57.9 us with 128, 51.5 us with 64; TP 8 / M 512: 102 / 48.6 / 55.4;
// TP 1 / M 256: 112 / 99 / 133.  B200_SK_BN overrides the width for M > 64 (tuning runs).
static int streamk_block_bn(int64_t M, int64_t weight_units, int sms) {
  static const int forced = getenv("B200_SK_BN") ? atoi(getenv("B200_SK_BN")) : 0;
  if (M <= 64) return streamk_bn(M);
  if (forced == 64 || forced == 128 || forced == 256) return forced >= M ? streamk_bn(M) : forced;
  if (M > 256) return 128;
  return weight_units * ceil_div(M, 128) >= 16ll * sms ? 128 : 64;
}

template <int MODE, int BN>
static int launch_streamk(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& g, int total_units, int ctas,
                          void* ws, cudaStream_t st) {
  using L = StreamK<BN>;
  auto kern = gemm_streamk_kernel<MODE, BN>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  uint32_t* flags = reinterpret_cast<uint32_t*>(ws);
  uint32_t* partials = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(ws) + kSkFlagBytes);
  return launch_k(kern, dim3((unsigned)ctas), dim3(kGemmThreads), L::kBytes, st, 1, true, tw, tx, g, total_units, flags,
                  partials);
}

template <int MODE>
static int run_gemm(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, cudaStream_t st,
                    void* ws = nullptr, size_t ws_bytes = 0) {
  g.M = (int)M;
  g.N = (int)N;
  g.num_kb = (int)ceil_div(K, kBK);
  g.trace = g_trace;
  g.trace_slots = (int)g_trace_slots;
  g.m_blocks = 1;
  static const int debug = getenv("B200_GEMM_DEBUG") ? atoi(getenv("B200_GEMM_DEBUG")) : 0;
  g.debug = debug;
  CUtensorMap ta, tb;
  int rc;
  if constexpr (MODE != kModeF8Block) {
    static const bool streamk = getenv("B200_GEMM_STREAMK") == nullptr || getenv("B200_GEMM_STREAMK")[0] != '0';
    // measured (profiles/r01e_gemm_decode.jsonl): below 64 tokens the epilogue is negligible and the cluster kernel's
    // two CTAs per SM start streaming earlier; from 64 tokens up the balanced persistent kernel wins
    static const int streamk_min_m = getenv("B200_GEMM_STREAMK_MIN_M") ? atoi(getenv("B200_GEMM_STREAMK_MIN_M")) : 33;
    // 256 < M <= 512 (a TP-8 decode batch): weight streaming still wins while the 128x256 tiles of the prefill kernel
    // would not fill the SMs once (1/8-size shards: 12 .. 64 tiles)
    const bool sk_wide = M > 256 && M <= kSkMaxM && ceil_div(M, 128) * ceil_div(N, kPfBN) < sm_count();
    if ((M <= 256 || sk_wide) && M >= streamk_min_m && ws != nullptr && streamk) {
      const size_t need = streamk_workspace_bytes(M);
      if (ws_bytes < need)
        return fail(B200_ERR_WORKSPACE, "scaled_mm: workspace of %zu bytes, %zu needed", ws_bytes, need);
      B200_REQUIRE(((uintptr_t)ws % 16) == 0, "scaled_mm: workspace must be 16-byte aligned");
      const int bn = streamk_block_bn(M, ceil_div(N, 128) * g.num_kb, sm_count());
      g.m_blocks = (int)ceil_div(M, bn);
      const int64_t units = ceil_div(N, 128) * g.m_blocks * g.num_kb;
      B200_REQUIRE(units < (1ll << 30), "scaled_mm: too many work units");
      // Grid: one CTA per SM.  When there are fewer tiles than SMs, prefer P = #SM / tiles CTAs per tile with ranges
      // aligned to tile boundaries (one segment per CTA, P - 1 peers per owner) if that still fills >= 85% of the SMs.
      const int sms = sm_count();
      const int tiles = (int)ceil_div(N, 128) * g.m_blocks;
      int ctas = (int)(units < sms ? units : sms);
      if (tiles < sms) {
        const int per_tile = sms / tiles < g.num_kb ? sms / tiles : g.num_kb;
        if (tiles * per_tile * 100 >= sms * 85) ctas = tiles * per_tile;
        // Short K on a quarter of the SMs or more (decode qkv / o_proj of an 8B model: 32-48 tiles of <= 512 KB): ONE CTA
        // per tile.  A single SM streams its 512 KB in ~3 us, while the fix-up a K split brings (publish, flag, 32 KB
        // per peer back through L2) costs ~5 us after the last MMA: 11.2 -> 9.6 us (qkv) and 10.6 -> 9.3 us (o_proj) at
        // M = 64 (profiles/r02zj_int8_ctas_per_tile_sweep.jsonl).  Measured for 33 <= M <= 64 only, so limited to it.
        if (g.m_blocks == 1 && M <= 64 && tiles >= 32 && g.num_kb <= 32) ctas = tiles;
        // B200_SK_CTAS_PER_TILE = p (experiments): force p CTAs per tile (p - 1 peers per owner; 1 = no fix-up at all)
        static const int force = getenv("B200_SK_CTAS_PER_TILE") ? atoi(getenv("B200_SK_CTAS_PER_TILE")) : 0;
        if (force > 0 && force <= g.num_kb && tiles * force <= sms) ctas = tiles * force;
      }
      g.splits = 1;
      g.kb_per_split = g.num_kb;
      if ((rc = cached_tmap(&ta, b, (uint64_t)N, (uint64_t)K, 128)) != B200_OK) return rc;
      if ((rc = cached_tmap(&tb, a, (uint64_t)M, (uint64_t)K, (uint32_t)bn)) != B200_OK) return rc;
      switch (bn) {
        case 16: return launch_streamk<MODE, 16>(ta, tb, g, (int)units, ctas, ws, st);
        case 32: return launch_streamk<MODE, 32>(ta, tb, g, (int)units, ctas, ws, st);
        case 64: return launch_streamk<MODE, 64>(ta, tb, g, (int)units, ctas, ws, st);
        case 128: return launch_streamk<MODE, 128>(ta, tb, g, (int)units, ctas, ws, st);
        default: return launch_streamk<MODE, 256>(ta, tb, g, (int)units, ctas, ws, st);
      }
    }
  }
  if constexpr (MODE == kModeF8Block) {
    // gemm_fp8_block.cu: stream-K weight streaming with per-K-block promotion (M <= 128, needs the workspace) and the
    // persistent 128x256 prefill kernel; B200_FP8_BLOCK_LEGACY=1 keeps the round-1 kernels below for A/B runs
    static const bool legacy = getenv("B200_FP8_BLOCK_LEGACY") != nullptr && getenv("B200_FP8_BLOCK_LEGACY")[0] == '1';
    if (!legacy) {
      if (M <= 128 && ws != nullptr) {
        const size_t need = streamk_workspace_bytes(M);
        if (ws_bytes < need)
          return fail(B200_ERR_WORKSPACE, "fp8_blockwise_scaled_mm: workspace of %zu bytes, %zu needed", ws_bytes, need);
        B200_REQUIRE(((uintptr_t)ws % 16) == 0, "fp8_blockwise_scaled_mm: workspace must be 16-byte aligned");
        g.splits = 1;
        g.kb_per_split = g.num_kb;
        return launch_block_fp8_streamk(g, a, b, M, N, K, ws, st);
      }
      if (M > 128 && N >= 256) {
        g.splits = 1;
        g.kb_per_split = g.num_kb;
        return launch_block_fp8_prefill(g, a, b, M, N, K, st);
      }
    }
  }
  // block-scaled FP8 keeps BN fp32 accumulators per epilogue thread in registers, so its token tile stops at 128
  const int64_t swap_max_m = MODE == kModeF8Block ? 128 : 256;
  if (M <= swap_max_m) {
    // weight-streaming orientation: A = weights [N, K], B = tokens [M, K]
    const int bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
    const int tiles = (int)ceil_div(N, 128);
    g.splits = pick_splits(tiles, g.num_kb, bn, sm_count());
    g.kb_per_split = (int)ceil_div(g.num_kb, g.splits);
    while (g.splits > 1 && (int64_t)(g.splits - 1) * g.kb_per_split >= g.num_kb) {  // every rank needs >= 1 K block
      g.splits /= 2;
      g.kb_per_split = (int)ceil_div(g.num_kb, g.splits);
    }
    if ((rc = cached_tmap(&ta, b, (uint64_t)N, (uint64_t)K, 128)) != B200_OK) return rc;
    if ((rc = cached_tmap(&tb, a, (uint64_t)M, (uint64_t)K, (uint32_t)bn)) != B200_OK) return rc;
    dim3 grid(tiles, 1, g.splits);
    switch (bn) {
      case 16: return launch_gemm<MODE, 16, 4, true>(ta, tb, g, grid, st);
      case 32: return launch_gemm<MODE, 32, 4, true>(ta, tb, g, grid, st);
      case 64: return launch_gemm<MODE, 64, 4, true>(ta, tb, g, grid, st);
      case 128: return launch_gemm<MODE, 128, 4, true>(ta, tb, g, grid, st);
      default:
        if constexpr (MODE != kModeF8Block) return launch_gemm<MODE, 256, 4, true>(ta, tb, g, grid, st);
        return fail(B200_ERR_INVALID_ARG, "unreachable tile selection");
    }
  }
  // prefill orientation: A = tokens [M, K], B = weights [N, K]
  g.splits = 1;
  g.kb_per_split = g.num_kb;
  if constexpr (MODE != kModeF8Block) {
    static const bool persist = getenv("B200_GEMM_PERSIST") == nullptr || getenv("B200_GEMM_PERSIST")[0] != '0';
    if (N >= kPfBN && persist) {
      if ((rc = cached_tmap(&ta, a, (uint64_t)M, (uint64_t)K, 128)) != B200_OK) return rc;
      const int tiles_m = (int)ceil_div(M, 128);
      using L = GemmSmemLayout<kPfBN, kPfStages>;
      // CTA pairs with the weight tile multicast across the pair (see the kernel) from two m-tiles up;
      // B200_GEMM_MC=0: one CTA per tile (A/B runs)
      // B200_GEMM_PAIR=0 / B200_GEMM_MC=0 select the older kernels (A/B runs)
      static const bool pair = getenv("B200_GEMM_PAIR") == nullptr || getenv("B200_GEMM_PAIR")[0] != '0';
      if (pair && tiles_m >= 2) {
        if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, 128)) != B200_OK) return rc;
        const int tiles_m2 = (int)ceil_div(tiles_m, 2);
        const int num_pairs = tiles_m2 * (int)ceil_div(N, 256);
        auto kern = gemm_prefill_pair_kernel<MODE>;
        B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kP2Bytes));
        const int pairs = num_pairs < sm_count() / 2 ? num_pairs : sm_count() / 2;
        // a kernel that uses cta_group::2 must be launched with an even cluster x dimension (a (1,1,2) cluster is
        // rejected as "cluster misconfiguration")
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(2 * pairs));
        cfg.blockDim = dim3(kGemmThreads);
        cfg.dynamicSmemBytes = kP2Bytes;
        cfg.stream = st;
        cudaLaunchAttribute attr[2];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        int na = 1;
        if (pdl_enabled()) {
          attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
          attr[na].val.programmaticStreamSerializationAllowed = 1;
          ++na;
        }
        cfg.attrs = attr;
        cfg.numAttrs = na;
        B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, ta, tb, g, tiles_m2, num_pairs));
        return B200_OK;
      }
      static const bool mc = getenv("B200_GEMM_MC") == nullptr || getenv("B200_GEMM_MC")[0] != '0';
      if (mc && tiles_m >= 2) {
        if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, 128)) != B200_OK) return rc;
        const int num_pairs = (int)ceil_div(tiles_m, 2) * (int)ceil_div(N, kPfBN);
        auto kern = gemm_prefill_kernel<MODE, true>;
        B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
        const int pairs = num_pairs < sm_count() / 2 ? num_pairs : sm_count() / 2;
        return launch_k(kern, dim3((unsigned)pairs, 1, 2), dim3(kGemmThreads), L::kBytes, st, 2, true, ta, tb, g, tiles_m,
                        num_pairs);
      }
      if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, (uint32_t)kPfBN)) != B200_OK) return rc;
      const int num_tiles = tiles_m * (int)ceil_div(N, kPfBN);
      auto kern = gemm_prefill_kernel<MODE, false>;
      B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
      const int ctas = num_tiles < sm_count() ? num_tiles : sm_count();
      return launch_k(kern, dim3((unsigned)ctas), dim3(kGemmThreads), L::kBytes, st, 1, true, ta, tb, g, tiles_m, num_tiles);
    }
  }
  const int bn = (MODE != kModeF8Block && N >= 256) ? 256 : 128;
  if ((rc = cached_tmap(&ta, a, (uint64_t)M, (uint64_t)K, 128)) != B200_OK) return rc;
  if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, (uint32_t)bn)) != B200_OK) return rc;
  dim3 grid((unsigned)ceil_div(M, 128), (unsigned)ceil_div(N, bn), 1);
  if constexpr (MODE != kModeF8Block) {
    if (bn == 256) return launch_gemm<MODE, 256, 4, false>(ta, tb, g, grid, st);
  }
  return launch_gemm<MODE, 128, 4, false>(ta, tb, g, grid, st);
}

// Un-scaled bf16 / f16 weight-streaming GEMM for decode-size M (lm_head): the stream-K kernel with kind::f16 MMAs.
template <int MODE>
static int run_dense16(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, cudaStream_t st, void* ws,
                       size_t ws_bytes) {
  g.M = (int)M;
  g.N = (int)N;
  g.num_kb = (int)ceil_div(K * 2, kBK);
  g.trace = g_trace;
  g.trace_slots = (int)g_trace_slots;
  g.m_blocks = 1;
  g.splits = 1;
  g.kb_per_split = g.num_kb;
  const size_t need = streamk_workspace_bytes(M);
  if (ws == nullptr || ws_bytes < need)
    return fail(B200_ERR_WORKSPACE, "dense_gemm_nt: workspace of %zu bytes, %zu needed", ws_bytes, need);
  B200_REQUIRE(((uintptr_t)ws % 16) == 0, "dense_gemm_nt: workspace must be 16-byte aligned");
  const int bn = streamk_bn(M);
  const int tiles = (int)ceil_div(N, 128);
  const int64_t units = (int64_t)tiles * g.num_kb;
  B200_REQUIRE(units < (1ll << 30), "dense_gemm_nt: too many work units");
  const int sms = sm_count();
  int ctas = (int)(units < sms ? units : sms);
  if (tiles < sms) {
    const int per_tile = sms / tiles < g.num_kb ? sms / tiles : g.num_kb;
    if (tiles * per_tile * 100 >= sms * 85) ctas = tiles * per_tile;
  }
  CUtensorMap tw, tx;
  int rc;
  if ((rc = cached_tmap(&tw, b, (uint64_t)N, (uint64_t)K * 2, 128)) != B200_OK) return rc;
  if ((rc = cached_tmap(&tx, a, (uint64_t)M, (uint64_t)K * 2, (uint32_t)bn)) != B200_OK) return rc;
  switch (bn) {
    case 16: return launch_streamk<MODE, 16>(tw, tx, g, (int)units, ctas, ws, st);
    case 32: return launch_streamk<MODE, 32>(tw, tx, g, (int)units, ctas, ws, st);
    case 64: return launch_streamk<MODE, 64>(tw, tx, g, (int)units, ctas, ws, st);
    case 128: return launch_streamk<MODE, 128>(tw, tx, g, (int)units, ctas, ws, st);
    default: return launch_streamk<MODE, 256>(tw, tx, g, (int)units, ctas, ws, st);
  }
}

static int check_common(const char* what, const void* out, const void* a, const void* b, int64_t M, int64_t N,
                        int64_t K, int out_dtype) {
  B200_REQUIRE(M >= 0 && N > 0 && K > 0, "%s: bad sizes", what);
  B200_REQUIRE(out && a && b, "%s: null pointer", what);
  B200_REQUIRE(K % 16 == 0, "%s: K %lld must be a multiple of 16", what, (long long)K);
  B200_REQUIRE(N % 8 == 0, "%s: N %lld must be a multiple of 8", what, (long long)N);
  B200_REQUIRE(((uintptr_t)out % 16) == 0, "%s: output must be 16-byte aligned", what);
  B200_REQUIRE(M < (1ll << 30) && N < (1ll << 30), "%s: sizes exceed 32-bit tile math", what);
  if (out_dtype != B200_BF16 && out_dtype != B200_F16)
    return fail(B200_ERR_UNSUPPORTED, "%s: out dtype %d not supported (bf16/f16)", what, out_dtype);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

B200_API int b200_int8_scaled_mm_workspace_bytes(int64_t M, int64_t N, int64_t K, size_t* bytes) {
  B200_REQUIRE(bytes != nullptr, "int8_scaled_mm_workspace_bytes: null out pointer");
  // stream-K decode path (M <= 256): per-CTA partial slots + flags; must be zero-initialised ONCE by the caller and
  // is self-cleaning afterwards.  One workspace serves all scaled_mm calls of a stream (calls on different streams
  // need their own).  Prefill sizes need none.
  *bytes = streamk_workspace_bytes(M);
  return B200_OK;
}

B200_API int b200_int8_scaled_mm(void* out, const int8_t* a, const int8_t* b, const float* scale_a, const float* scale_b,
                                 const void* bias, int64_t M, int64_t N, int64_t K, int out_dtype, int32_t* acc_out,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  // same shape rules as the reference (int8_gemm_kernel.cu:677-679)
  int rc = check_common("int8_scaled_mm", out, a, b, M, N, K, out_dtype);
  if (rc != B200_OK) return rc;
  if (M == 0) return B200_OK;
  B200_REQUIRE(scale_a && scale_b, "int8_scaled_mm: null scale pointer");
  B200_REQUIRE(((uintptr_t)acc_out % 16) == 0, "int8_scaled_mm: acc_out must be 16-byte aligned");
  GemmArgs g = {};
  g.out = out; g.scale_m = scale_a; g.scale_n = scale_b; g.bias = bias; g.acc_out = acc_out;
  g.out_is_bf16 = out_dtype == B200_BF16;
  return run_gemm<kModeI8>(g, a, b, M, N, K, (cudaStream_t)stream, workspace, workspace_bytes);
}

B200_API int b200_fp8_scaled_mm(void* out, const void* a, const void* b, const float* scale_a, const float* scale_b,
                                const void* bias, int64_t M, int64_t N, int64_t K, int out_dtype, void* workspace,
                                size_t workspace_bytes, void* stream) {
  int rc = check_common("fp8_scaled_mm", out, a, b, M, N, K, out_dtype);
  if (rc != B200_OK) return rc;
  if (M == 0) return B200_OK;
  B200_REQUIRE(scale_a && scale_b, "fp8_scaled_mm: null scale pointer");
  GemmArgs g = {};
  g.out = out; g.scale_m = scale_a; g.scale_n = scale_b; g.bias = bias;
  g.out_is_bf16 = out_dtype == B200_BF16;
  return run_gemm<kModeF8Chan>(g, a, b, M, N, K, (cudaStream_t)stream, workspace, workspace_bytes);
}

B200_API int b200_fp8_blockwise_scaled_mm(void* out, const void* a, const void* b, const float* scale_a,
                                          int64_t sa_row_stride, int64_t sa_col_stride, const float* scale_b,
                                          int64_t sb_row_stride, int64_t sb_col_stride, const void* bias, int64_t M,
                                          int64_t N, int64_t K, int out_dtype, void* workspace, size_t workspace_bytes,
                                          void* stream) {
  int rc = check_common("fp8_blockwise_scaled_mm", out, a, b, M, N, K, out_dtype);
  if (rc != B200_OK) return rc;
  if (M == 0) return B200_OK;
  B200_REQUIRE(scale_a && scale_b, "fp8_blockwise_scaled_mm: null scale pointer");
  // the reference's CUTLASS path has the same shape rule (fp8_utils.py:190-197)
  B200_REQUIRE(K % 128 == 0 && N % 128 == 0, "fp8_blockwise_scaled_mm: N %lld and K %lld must be multiples of 128",
               (long long)N, (long long)K);
  GemmArgs g = {};
  g.out = out; g.scale_m = scale_a; g.scale_n = scale_b; g.bias = bias;
  g.sm_rs = sa_row_stride; g.sm_cs = sa_col_stride; g.sn_rs = sb_row_stride; g.sn_cs = sb_col_stride;
  g.out_is_bf16 = out_dtype == B200_BF16;
  return run_gemm<kModeF8Block>(g, a, b, M, N, K, (cudaStream_t)stream, workspace, workspace_bytes);
}

B200_API int b200_dense_gemm_nt(void* out, const void* a, const void* b, const void* bias, int64_t M, int64_t N, int64_t K,
                                int dtype, void* workspace, size_t workspace_bytes, void* stream) {
  // out[M, N] = a[M, K] . b[N, K]^T (+ bias), bf16 or f16 operands and output, fp32 accumulation in TMEM.
  // Decode-size M only (<= 256): the lm_head / logits GEMM of a decode step or of the last tokens of a prefill chunk.
  int rc = check_common("dense_gemm_nt", out, a, b, M, N, K, dtype);
  if (rc != B200_OK) return rc;
  if (M == 0) return B200_OK;
  B200_REQUIRE(K % 8 == 0, "dense_gemm_nt: K %lld must be a multiple of 8 (16-byte rows)", (long long)K);
  if (M > 256) return fail(B200_ERR_UNSUPPORTED, "dense_gemm_nt: M %lld > 256 (weight-streaming sizes only)", (long long)M);
  GemmArgs g = {};
  g.out = out; g.bias = bias;
  g.out_is_bf16 = dtype == B200_BF16;
  if (dtype == B200_BF16)
    return run_dense16<kModeBF16>(g, a, b, M, N, K, (cudaStream_t)stream, workspace, workspace_bytes);
  return run_dense16<kModeF16>(g, a, b, M, N, K, (cudaStream_t)stream, workspace, workspace_bytes);
}
