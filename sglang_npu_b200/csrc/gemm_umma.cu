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

#include "sm100.cuh"

namespace b200 {

using namespace sm100;

enum GemmMode { kModeI8 = 0, kModeF8Chan = 1, kModeF8Block = 2 };

struct GemmArgs {
  void* out;             // [M, N] bf16/f16
  const float* scale_m;  // per token: [M] (I8 / F8Chan) or [M, K/128] with strides (F8Block)
  const float* scale_n;  // per channel: [N] or [N/128, K/128] with strides (F8Block)
  const void* bias;      // [N] out dtype or null
  int32_t* acc_out;      // [M, N] raw int32 accumulators (I8 only) or null
  int M, N;
  int num_kb;        // K blocks of 128 bytes (ceil)
  int kb_per_split;  // K blocks per cluster rank
  int splits;        // cluster size along z (1, 2, 4, 8)
  int out_is_bf16;
  int64_t sm_rs, sm_cs;  // scale_m[m * sm_rs + kb * sm_cs]          (F8Block)
  int64_t sn_rs, sn_cs;  // scale_n[(n / 128) * sn_rs + kb * sn_cs]   (F8Block)
};

constexpr int kGemmThreads = 256;
constexpr int kBK = 128;  // bytes of K per pipeline stage = one 128B swizzle row

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
  else v = __uint_as_float(bits);
  return has_bias ? __fadd_rn(v, bias) : v;
}

__device__ __forceinline__ float load_bias(const void* bias, int n, int is_bf16) {
  return is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias)[n])
                 : __half2float(reinterpret_cast<const __half*>(bias)[n]);
}

__device__ __forceinline__ void store_out(void* out, int64_t idx, float v, int is_bf16) {
  if (is_bf16) reinterpret_cast<__nv_bfloat16*>(out)[idx] = __float2bfloat16_rn(v);
  else reinterpret_cast<__half*>(out)[idx] = __float2half_rn(v);
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
    if (lane == 0) {
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
    if (lane == 0) {
      constexpr uint32_t idesc = MODE == kModeI8 ? make_idesc(/*s32*/ 2, /*int8*/ 1, 1, 128, BN)
                                                 : make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 128, BN);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(full_bar(s), ph);
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
  }

  if (warp == 2 || warp == 3) pdl_wait();

  // ---- split-K: reduce-scatter the partial accumulators across the cluster through DSMEM -----------------
  if constexpr (SWAP) {
    if (S > 1) {
      __syncwarp();
      cluster_sync();  // every CTA of the cluster has retired its MMAs: pipeline smem is free to reuse
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
      cluster_sync();  // all partials have landed
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

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<kTmemCols>(tmem_base);
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

template <int MODE>
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
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(tmem_full_bar(b), 1);
      mbar_init(tmem_empty_bar(b), 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(smem_u32(&s_tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();
  pdl_wait();

  const int nkb = g.num_kb;
  if (warp == 0) {
    if (lane == 0) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int n_tile = t / tiles_m;
        const int a_row0 = (t - n_tile * tiles_m) * 128;
        const int b_row0 = n_tile * kPfBN;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kPfStages;
          mbar_wait(empty_bar(s), ((it / kPfStages) & 1) ^ 1);
          mbar_expect_tx(full_bar(s), L::kStage);
          const uint32_t sa = base + s * L::kStage;
          tma_load_2d(sa, &tmap_a, full_bar(s), i * kBK, a_row0, kEvictLast);
          tma_load_2d(sa + L::kA, &tmap_b, full_bar(s), i * kBK, b_row0, kEvictNormal);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
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
          tc_commit(empty_bar(s));
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
      const int n_tile = t / tiles_m;
      const int m = (t - n_tile * tiles_m) * 128 + row;
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
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
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

template <int MODE>
static int run_gemm(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, cudaStream_t st) {
  g.M = (int)M;
  g.N = (int)N;
  g.num_kb = (int)ceil_div(K, kBK);
  CUtensorMap ta, tb;
  int rc;
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
      if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, (uint32_t)kPfBN)) != B200_OK) return rc;
      const int tiles_m = (int)ceil_div(M, 128);
      const int num_tiles = tiles_m * (int)ceil_div(N, kPfBN);
      using L = GemmSmemLayout<kPfBN, kPfStages>;
      auto kern = gemm_prefill_kernel<MODE>;
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
  *bytes = 0;  // split-K reduces through distributed shared memory; no global workspace
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
  return run_gemm<kModeI8>(g, a, b, M, N, K, (cudaStream_t)stream);
}

B200_API int b200_fp8_scaled_mm(void* out, const void* a, const void* b, const float* scale_a, const float* scale_b,
                                const void* bias, int64_t M, int64_t N, int64_t K, int out_dtype, void* stream) {
  int rc = check_common("fp8_scaled_mm", out, a, b, M, N, K, out_dtype);
  if (rc != B200_OK) return rc;
  if (M == 0) return B200_OK;
  B200_REQUIRE(scale_a && scale_b, "fp8_scaled_mm: null scale pointer");
  GemmArgs g = {};
  g.out = out; g.scale_m = scale_a; g.scale_n = scale_b; g.bias = bias;
  g.out_is_bf16 = out_dtype == B200_BF16;
  return run_gemm<kModeF8Chan>(g, a, b, M, N, K, (cudaStream_t)stream);
}

B200_API int b200_fp8_blockwise_scaled_mm(void* out, const void* a, const void* b, const float* scale_a,
                                          int64_t sa_row_stride, int64_t sa_col_stride, const float* scale_b,
                                          int64_t sb_row_stride, int64_t sb_col_stride, const void* bias, int64_t M,
                                          int64_t N, int64_t K, int out_dtype, void* stream) {
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
  return run_gemm<kModeF8Block>(g, a, b, M, N, K, (cudaStream_t)stream);
}
