// Block-scaled FP8 GEMM (W8A8-FP8, 1x128 activation groups x 128x128 weight blocks) on tcgen05, TMA-fed.
//
// Replaces fp8_blockwise_scaled_mm (reference sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423; its sm100
// path is a 1-SM CUTLASS blockwise kernel, :165-343) behind Fp8LinearMethod.apply's block branch
// (python/sglang/srt/layers/quantization/fp8.py:444-487, fp8_utils.py:180-210).  Semantics = the reference test's
// native_w8a8_block_fp8_matmul (python/sglang/test/test_block_fp8.py:274-330; oracle/quant_oracle.py):
//     y[m, n] = sum_kb  (sum_{k in kb} a_q[m,k] * w_q[n,k]) * a_s[m, kb] * w_s[n / 128, kb]        (fp32)
// A single TMEM accumulator over all of K would be wrong: every 128-wide K block has its own fp32 scale pair.  Both
// kernels here give every K block its own TMEM accumulator (tcgen05.mma kind::f8f6f4, accumulate off at the block
// start) out of a ring of TMEM buffers, and the epilogue warps promote block after block into fp32 registers
// (acc * (a_s * w_s) + sum) while the MMA thread is already several K blocks ahead.
//
//  * gemm_block_streamk_kernel (decode, M <= 128): the weight-streaming stream-K schedule of gemm_umma.cu
//    (weights on MMA-M, tokens on MMA-N, one 16 KiB weight box per unit, CTA c streams units [c*U/G, (c+1)*U/G)
//    through a deep TMA ring).  HBM-bound: an SM has ~750 cycles per unit, the promotion of a 128 x BN block takes
//    ~250.  Combined scales of a unit are computed once (thread t -> token t) into a 2-deep shared-memory table, so the
//    inner loop is one broadcast LDS.128 per four FFMAs; up to 8 TMEM buffers decouple the MMA thread from the warps.
//  * gemm_block_prefill_kernel (M > 128): persistent 128 x 256 tiles, tokens on MMA-M.  Here a thread owns a token row,
//    so a block's scale is one scalar per 128 channels.  Eight epilogue warps (two per TMEM lane quarter, 128 columns
//    each = 128 fp32 accumulators per thread, registers re-balanced with setmaxnreg) keep up with one 128x256x128 MMA
//    group per ~540 cycles; the tile's bf16 store overlaps the next tile's first two K blocks.
#include <cstdlib>

#include "gemm_common.cuh"

namespace b200 {

// acc.{x,y} += bits(x0, x1) * s.{x,y}.  Two scalar FFMAs by default: the packed fp32x2 FMA (sm_100 FFMA2, one
// instruction for both, -DB200_PACKED_PROMOTION) halves the instruction count but measured SLOWER in the promotion
// loops — prefill at M=8192: 1.59-1.73 PFLOP/s packed vs 1.99-2.10 scalar, down_proj (K=14336) 1.44 vs 2.03
// (profiles/r02z_kbench_fp8_block_scalar_vs_packed.jsonl).  Same rounding either way.
__device__ __forceinline__ void ffma2(float& a0, float& a1, uint32_t x0, uint32_t x1, float s0, float s1) {
#ifndef B200_PACKED_PROMOTION
  a0 = fmaf(__uint_as_float(x0), s0, a0);
  a1 = fmaf(__uint_as_float(x1), s1, a1);
  return;
#endif
  asm("{\n.reg .b64 va, vb, vc;\n"
      "mov.b64 va, {%2, %3};\nmov.b64 vb, {%4, %5};\nmov.b64 vc, {%0, %1};\n"
      "fma.rn.f32x2 vc, va, vb, vc;\n"
      "mov.b64 {%0, %1}, vc;\n}\n"
      : "+f"(a0), "+f"(a1)
      : "r"(x0), "r"(x1), "f"(s0), "f"(s1));
}

// ---- decode: stream-K weight streaming with per-K-block promotion ------------------------------------------------------
constexpr int kScaleChunk = 8;  // units per scale prefetch

template <int BN>
struct BlkSk {
  static constexpr int kStageBytes = 128 * kBK + BN * kBK;
  static constexpr int kStages = (192 * 1024) / kStageBytes > 10 ? 10 : (192 * 1024) / kStageBytes;
  static constexpr int kBufs = 512 / BN > 8 ? 8 : 512 / BN;
  static constexpr int kCols = kBufs * BN;
  static constexpr int kTmemCols = kCols <= 32 ? 32 : kCols <= 64 ? 64 : kCols <= 128 ? 128 : kCols <= 256 ? 256 : 512;
  static constexpr int kBarOff = kStages * kStageBytes;
  static constexpr int kNumBars = 2 * kStages + kBufs;
  static constexpr int kBytes = kBarOff + kNumBars * 8 + 1024;
  // ONE tcgen05.commit per unit: the barrier that frees smem stage s for the TMA thread is also what tells the
  // promotion warps that the unit's accumulator is complete.  A waiter may trail the barrier by at most one phase:
  // the MMA thread runs at most kBufs units ahead of the warps, so kBufs <= kStages keeps it within that.
  static_assert(kBufs <= kStages, "accumulator ring deeper than the smem ring");
};

// EW promotion warps: 4 (one per TMEM lane quarter, all BN columns each) or 8 (two per quarter, BN/2 columns each — a
// warp's unit is a serial chain barrier -> tcgen05.ld -> FMAs, two warps per SM sub-partition overlap each other's
// latencies).  A column's K blocks are still summed in order by one thread: results do not depend on EW.
template <int BN, int EW>
__global__ void __launch_bounds__(128 + 32 * EW, 1)
gemm_block_streamk_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x,
                          const GemmArgs g, const int total_units, uint32_t* __restrict__ flags,
                          float* __restrict__ partials) {
  using L = BlkSk<BN>;
  constexpr int STAGES = L::kStages;
  constexpr int BUFS = L::kBufs;
  constexpr int kWBytes = 128 * kBK;
  constexpr int HB = BN / (EW / 4);        // accumulator columns (tokens) per promotion warp
  constexpr int CH = HB >= 32 ? 32 : 16;  // TMEM columns per tcgen05.ld
  static_assert(EW == 4 || EW == 8, "one or two promotion warps per TMEM lane quarter");
  static_assert(HB >= 16, "a promotion warp reads at least 16 columns");
  auto epi_sync = [] { asm volatile("bar.sync 1, %0;\n" ::"n"(32 * EW) : "memory"); };
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  __shared__ __align__(16) float s_tab[2][kScaleChunk][BN];  // a_s[token, kb] * w_s[tile, kb], two chunks of units
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (STAGES + s) * 8; };
  auto tempty_bar = [&](int b) { return bar + (2 * STAGES + b) * 8; };
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
    for (int b = 0; b < BUFS; ++b) mbar_init(tempty_bar(b), EW);  // one arrive per promotion warp
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<L::kTmemCols>(smem_u32(&s_tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();

  const int G = gridDim.x;
  const int c = blockIdx.x;
  const int nkb = g.num_kb;
  const int u0 = (int)((int64_t)c * total_units / G);
  const int u1 = (int)((int64_t)(c + 1) * total_units / G);

  if (warp == 0) {
    if (elect_one()) {
      const int n_units = u1 - u0;
      const int pre = min(n_units, STAGES);
      int tile = u0 / nkb, kb = u0 - tile * nkb;
      // weights do not depend on the previous kernel: fill the ring with weight boxes before the PDL wait
      for (int i = 0; i < pre; ++i) {
        mbar_expect_tx(full_bar(i), L::kStageBytes);
        tma_load_2d(base + i * L::kStageBytes, &tmap_w, full_bar(i), kb * kBK, tile * 128, kEvictFirst);
        if (++kb == nkb) { kb = 0; ++tile; }
      }
      pdl_wait();
      trace_mark(tr, 2);
      tile = u0 / nkb, kb = u0 - tile * nkb;
      for (int i = 0; i < pre; ++i) {
        tma_load_2d(base + i * L::kStageBytes + kWBytes, &tmap_x, full_bar(i), kb * kBK, 0, kEvictLast);
        if (++kb == nkb) { kb = 0; ++tile; }
      }
      for (int i = pre; i < n_units; ++i) {
        const int s = i % STAGES;
        mbar_wait(empty_bar(s), ((i / STAGES) & 1) ^ 1);
        mbar_expect_tx(full_bar(s), L::kStageBytes);
        const uint32_t sa = base + s * L::kStageBytes;
        tma_load_2d(sa, &tmap_w, full_bar(s), kb * kBK, tile * 128, kEvictFirst);
        tma_load_2d(sa + kWBytes, &tmap_x, full_bar(s), kb * kBK, 0, kEvictLast);
        if (++kb == nkb) { kb = 0; ++tile; }
      }
    } else {
      pdl_wait();
    }
  } else if (warp == 1) {
    pdl_wait();
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 128, BN);
      const int n_units = u1 - u0;
      for (int it = 0; it < n_units; ++it) {
        const int s = it % STAGES;
        const int buf = it % BUFS;
        mbar_wait(tempty_bar(buf), ((it / BUFS) & 1) ^ 1);  // the epilogue has promoted the block that lived here
        mbar_wait(full_bar(s), (it / STAGES) & 1);
        if (it == 0) trace_mark(tr, 3);
        tc_fence_after();
        const uint32_t sa = base + s * L::kStageBytes;
        const uint64_t da = smem_desc_k_sw128(sa);
        const uint64_t db = smem_desc_k_sw128(sa + kWBytes);
        const uint32_t d_tmem = tmem_base + buf * BN;
#pragma unroll
        for (int k = 0; k < kBK / 32; ++k) umma<kKindF8>(d_tmem, da + 2 * k, db + 2 * k, idesc, k != 0);
        tc_commit(empty_bar(s));  // frees the smem stage AND publishes the accumulator (see BlkSk)
      }
      trace_mark(tr, 4);
    }
  } else if (warp >= 4) {
    pdl_wait();
    const int q = warp & 3;
    const int row = q * 32 + lane;  // output channel within the tile = TMEM lane
    const int col0 = ((warp - 4) >> 2) * HB;  // first token column of this warp
    const int t = threadIdx.x - 128;  // token whose combined scale this thread computes per unit
    const bool has_bias = g.bias != nullptr;
    const int m_cols = min(BN, (g.M + 15) & ~15);
    float facc[HB];
#pragma unroll
    for (int j = 0; j < HB; ++j) facc[j] = 0.f;
    // Combined scales a_s[token t, kb] * w_s[tile, kb]: thread t fetches them a CHUNK of units ahead (global-load
    // latency is ~a unit's worth of cycles, so a one-unit-deep prefetch would pace the loop) into a double-buffered
    // shared table; the promotion loop reads them with broadcast LDS.128.
    const int n_units = u1 - u0;
    int lt = u0 / nkb, lkb = u0 - lt * nkb;  // (tile, kb) of the next unit whose scale is to be fetched
    // The loaded factors stay RAW in registers until store_chunk, a chunk of units later: a multiply placed next to the
    // loads makes the warp sit out the whole global-load latency (several microseconds while every SM's TMA ring is
    // pulling weights) once per chunk — that, not the FMAs, paced this kernel at 5.2 TB/s.
    float va[kScaleChunk], vb[kScaleChunk];
    auto fetch_chunk = [&]() {
      // addresses first, then ALL loads back to back (one memory latency per chunk, not one per unit)
      int64_t oa[kScaleChunk], ob[kScaleChunk];
      bool ok[kScaleChunk];
#pragma unroll
      for (int j = 0; j < kScaleChunk; ++j) {
        ok[j] = lt * nkb + lkb < u1;
        oa[j] = t * g.sm_rs + lkb * g.sm_cs;
        ob[j] = lt * g.sn_rs + lkb * g.sn_cs;
        if (ok[j] && ++lkb == nkb) { lkb = 0; ++lt; }
      }
      const bool tok = t < g.M && t < BN;  // only the threads that fill the table load anything
#pragma unroll
      for (int j = 0; j < kScaleChunk; ++j) {
        va[j] = (ok[j] && tok) ? __ldg(g.scale_m + oa[j]) : 0.f;
        vb[j] = (ok[j] && tok) ? __ldg(g.scale_n + ob[j]) : 0.f;
      }
    };
    auto store_chunk = [&](int b) {
      if (t < BN) {
#pragma unroll
        for (int j = 0; j < kScaleChunk; ++j) {
          asm volatile("" : "+f"(va[j]), "+f"(vb[j]));  // first use of the loaded values is HERE, not at the loads
          s_tab[b][j][t] = va[j] * vb[j];
        }
      }
    };
    if (n_units > 0) {
      fetch_chunk();
      store_chunk(0);
    }
    epi_sync();
    if (threadIdx.x == 128) trace_mark(tr, 12);  // first chunk of scales in shared memory
    int seg_kb0 = u0 - (u0 / nkb) * nkb;  // first K block of the running segment
    int tile = u0 / nkb, kb = seg_kb0;
    for (int it = 0; it < n_units; ++it) {
      const int u = u0 + it;
      const int cj = it % kScaleChunk;
      const int cb = (it / kScaleChunk) & 1;
      const bool more = (it / kScaleChunk + 1) * kScaleChunk < n_units;
      if (cj == 0 && more) fetch_chunk();  // the next chunk's loads fly under this chunk's promotions
      const int buf = it % BUFS;
      mbar_wait(empty_bar(it % STAGES), (it / STAGES) & 1);
      tc_fence_after();
      if (threadIdx.x == 128) trace_mark(tr, it == 0 ? 5 : 7);  // 5: first unit accumulated, 7: latest one
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * BN + col0;
      const float* sc = s_tab[cb][cj] + col0;
      const bool dbg_no_read = (g.debug & 0x20) != 0;  // timing experiments (B200_GEMM_DEBUG): wrong results
      const bool dbg_no_fma = (g.debug & 0x10) != 0;
      if (col0 >= m_cols || dbg_no_read) {  // nothing of this buffer is this warp's (M <= BN / 2, two warps per quarter)
        if (lane == 0) mbar_arrive(tempty_bar(buf));
      }
#pragma unroll
      for (int c0 = 0; c0 < HB; c0 += CH) {
        if (col0 + c0 < m_cols && !dbg_no_read) {  // warp-uniform
          uint32_t r[CH];
          if constexpr (CH == 32) tmem_ld32(taddr + c0, r);
          else tmem_ld16(taddr + c0, r);
          tmem_ld_wait();
          if (col0 + c0 + CH >= m_cols || c0 + CH >= HB) {  // this warp's last read of the buffer: hand it back
            tc_fence_before();                             // to the MMA thread before the FMAs
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(buf));
          }
          if (dbg_no_fma) {
            facc[c0] += __uint_as_float(r[0]);
            continue;
          }
#pragma unroll
          for (int j = 0; j < CH; j += 4) {
            const float4 s4 = *reinterpret_cast<const float4*>(sc + c0 + j);
            ffma2(facc[c0 + j], facc[c0 + j + 1], r[j], r[j + 1], s4.x, s4.y);
            ffma2(facc[c0 + j + 2], facc[c0 + j + 3], r[j + 2], r[j + 3], s4.z, s4.w);
          }
        }
      }
      if (cj == kScaleChunk - 1 && more) {
        store_chunk(cb ^ 1);
        epi_sync();
      }
      const int cur_tile = tile, cur_kb = kb;
      if (++kb == nkb) { kb = 0; ++tile; }
      {
        const int tile = cur_tile, kb = cur_kb;
      const bool ends = kb == nkb - 1;
      if (!ends && u != u1 - 1) continue;
      // ---- the segment [seg_kb0, kb] of `tile` is complete -------------------------------------------------------------
      const bool starts = seg_kb0 == 0;
      if (!starts) {
        // contribution: fp32 partial sums -> partials[c][col][row]; a warp store covers 128 contiguous bytes
        float* dst = partials + ((size_t)c * BN + col0) * 128 + row;
#pragma unroll
        for (int j = 0; j < HB; ++j)
          if (col0 + j < m_cols) __stcg(dst + j * 128, facc[j]);
        // The flag is released by warp 3, not here: st.release.gpu is a MEMBAR.ALL.GPU, which under the initial flood of
        // weight requests takes microseconds — long enough for the MMA thread to run out of TMEM buffers and for this
        // SM's weight stream to stop (CTAs with a short first segment finished 4-5 us after the others).  bar.arrive
        // orders this thread's stores before the barrier's completion and does not wait.
        asm volatile("bar.arrive 2, %0;\n" ::"n"(32 * EW + 32) : "memory");
      } else {
        int n_peers = 0;
        if (!ends) {  // owner of a tile that continues in the following CTAs: add their partials in CTA order
          const int tile_end = (tile + 1) * nkb;
          while (c + 1 + n_peers < G && (int)((int64_t)(c + 1 + n_peers) * total_units / G) < tile_end) ++n_peers;
          for (int p = 0; p < n_peers; ++p) {
            while (ld_acquire_u32(flags + c + 1 + p) == 0u) {
            }
            if (threadIdx.x == 128) trace_mark(tr, 8);
            const float* src = partials + ((size_t)(c + 1 + p) * BN + col0) * 128 + row;
#pragma unroll
            for (int j0 = 0; j0 < HB; j0 += 16) {
              if (col0 + j0 < m_cols) {  // 16 independent loads in flight, then the adds (fixed order: bit-deterministic)
                float tmp[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) tmp[j] = __ldcg(src + (j0 + j) * 128);
#pragma unroll
                for (int j = 0; j < 16; ++j) facc[j0 + j] = __fadd_rn(facc[j0 + j], tmp[j]);
              }
            }
          }
        }
        const int n = tile * 128 + row;
        if (n < g.N) {
          const float bias = has_bias ? load_bias(g.bias, n, g.out_is_bf16) : 0.f;
          uint16_t* p = reinterpret_cast<uint16_t*>(g.out) + n;
          const bool bf = g.out_is_bf16 != 0;
#pragma unroll
          for (int j = 0; j < HB; ++j) {
            if (col0 + j < g.M) {
              const float v = has_bias ? __fadd_rn(facc[j], bias) : facc[j];
              p[(int64_t)(col0 + j) * g.N] = bf ? __bfloat16_as_ushort(__float2bfloat16_rn(v)) : __half_as_ushort(__float2half_rn(v));
            }
          }
        }
        if (n_peers > 0) {
          epi_sync();  // every epilogue thread has consumed the peers' partials
          if (threadIdx.x == 128)
            for (int p = 0; p < n_peers; ++p) flags[c + 1 + p] = 0u;
        }
      }
#pragma unroll
      for (int j = 0; j < HB; ++j) facc[j] = 0.f;
      seg_kb0 = 0;  // the next unit (if any) opens the next tile at K block 0
      }
    }
    if (threadIdx.x == 128) trace_mark(tr, 9);
  } else if (warp == 3) {
    pdl_wait();
    // Publisher of this CTA's contribution (its first segment, when that starts inside a tile): completes the barrier
    // the promotion warps arrive on after their partial stores, then releases the flag (cumulativity covers their
    // stores: same pattern as a CTA barrier followed by one releasing thread, without stalling the promotion).
    if (u1 > u0 && u0 % nkb != 0) {
      asm volatile("bar.sync 2, %0;\n" ::"n"(32 * EW + 32) : "memory");
      if (lane == 0) {
        st_release_u32(flags + c, 1u);
        trace_mark(tr, 6);
      }
    }
  } else {
    pdl_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<L::kTmemCols>(tmem_base);
  if (threadIdx.x == 0) trace_mark(tr, 15);
}

// ---- prefill: persistent 128 x 256 tiles, eight promotion warps ----------------------------------------------------------
constexpr int kBpBN = 256;
constexpr int kBpStages = 4;
constexpr int kBpThreads = 384;
constexpr int kBpStage = 128 * kBK + kBpBN * kBK;
constexpr int kBpBarOff = kBpStages * kBpStage;
constexpr int kBpBytes = kBpBarOff + (2 * kBpStages + 2) * 8 + 1024;

template <int N>
__device__ __forceinline__ void reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(N)); }

__global__ void __launch_bounds__(kBpThreads, 1)
gemm_block_prefill_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                          const GemmArgs g, const int tiles_m, const int num_tiles) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + kBpBarOff;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (kBpStages + s) * 8; };
  auto tempty_bar = [&](int b) { return bar + (2 * kBpStages + b) * 8; };
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kBpStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int b = 0; b < 2; ++b) mbar_init(tempty_bar(b), 8);  // one arrive per epilogue warp
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
  if (warp < 4) {
    reg_dealloc<40>();
    if (warp == 0 && elect_one()) {
      uint32_t it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int n_tile = t / tiles_m;
        const int a_row0 = (t - n_tile * tiles_m) * 128;
        const int b_row0 = n_tile * kBpBN;
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kBpStages;
          mbar_wait(empty_bar(s), ((it / kBpStages) & 1) ^ 1);
          mbar_expect_tx(full_bar(s), kBpStage);
          const uint32_t sa = base + s * kBpStage;
          tma_load_2d(sa, &tmap_a, full_bar(s), i * kBK, a_row0, kEvictLast);
          tma_load_2d(sa + 128 * kBK, &tmap_b, full_bar(s), i * kBK, b_row0, kEvictNormal);
        }
      }
    } else if (warp == 1 && elect_one()) {
      constexpr uint32_t idesc = make_idesc(/*f32*/ 1, /*e4m3*/ 0, 0, 128, kBpBN);
      uint32_t it = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        for (int i = 0; i < nkb; ++i, ++it) {
          const int s = it % kBpStages;
          const int buf = it & 1;
          mbar_wait(tempty_bar(buf), ((it >> 1) & 1) ^ 1);  // both promotion warpgroups have read this buffer
          mbar_wait(full_bar(s), (it / kBpStages) & 1);
          tc_fence_after();
          const uint32_t sa = base + s * kBpStage;
          const uint64_t da = smem_desc_k_sw128(sa);
          const uint64_t db = smem_desc_k_sw128(sa + 128 * kBK);
          const uint32_t d_tmem = tmem_base + buf * kBpBN;
#pragma unroll
          for (int k = 0; k < kBK / 32; ++k) umma<kKindF8>(d_tmem, da + 2 * k, db + 2 * k, idesc, k != 0);
          tc_commit(empty_bar(s));  // one commit: frees the stage and publishes the accumulator (2 buffers <= 4 stages)
        }
      }
    }
  } else {
    reg_alloc<232>();
    const int q = warp & 3;           // TMEM lane quarter this warp may read
    const int half = (warp - 4) >> 2;  // which 128 of the tile's 256 channels
    const int row = q * 32 + lane;
    const bool has_bias = g.bias != nullptr;
    float facc[128];
    uint32_t it = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int n_tile = t / tiles_m;
      const int m = (t - n_tile * tiles_m) * 128 + row;
      const int n0 = n_tile * kBpBN + half * 128;
      const bool m_ok = m < g.M;
      const bool n_ok = n0 < g.N;  // N % 128 == 0: a half tile is all in or all out
      const float* sm_p = g.scale_m + (int64_t)(m_ok ? m : 0) * g.sm_rs;
      const float* sn_p = g.scale_n + (int64_t)(n_ok ? (n0 >> 7) : 0) * g.sn_rs;
#pragma unroll
      for (int j = 0; j < 128; ++j) facc[j] = 0.f;
      // Scales of K block i are LOADED two blocks ahead and MULTIPLIED only when block i is promoted: the activation
      // scale of a thread's row is its own cache line (32 lines per warp load, an L2 round trip), and a product
      // formed right behind the loads parked every promotion warp on that latency once per K block (17 % of all
      // warp samples on that one FMUL, tensor pipe 44 % active: profiles/r02z_fp8_block_prefill_ncu.txt).
      float a1 = __ldg(sm_p), w1 = __ldg(sn_p);
      float a2 = 0.f, w2 = 0.f;
      if (nkb > 1) {
        a2 = __ldg(sm_p + g.sm_cs);
        w2 = __ldg(sn_p + g.sn_cs);
      }
      for (int i = 0; i < nkb; ++i, ++it) {
        const float s = a1 * w1;
        a1 = a2;
        w1 = w2;
        if (i + 2 < nkb) {
          a2 = __ldg(sm_p + (i + 2) * g.sm_cs);
          w2 = __ldg(sn_p + (i + 2) * g.sn_cs);
        }
        const int buf = it & 1;
        mbar_wait(empty_bar(it % kBpStages), (it / kBpStages) & 1);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + buf * kBpBN + half * 128;
        const bool early = (g.debug & 4) != 0;
#pragma unroll
        for (int c0 = 0; c0 < 128; c0 += 64) {
          uint32_t r0[32], r1[32];
          tmem_ld32(taddr + c0, r0);
          tmem_ld32(taddr + c0 + 32, r1);
          tmem_ld_wait();
          if (c0 == 64 && early) {  // last read of this buffer: hand it back to the MMA thread before the FMAs
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tempty_bar(buf));
          }
#pragma unroll
          for (int j = 0; j < 32; j += 2) ffma2(facc[c0 + j], facc[c0 + j + 1], r0[j], r0[j + 1], s, s);
#pragma unroll
          for (int j = 0; j < 32; j += 2) ffma2(facc[c0 + 32 + j], facc[c0 + 33 + j], r1[j], r1[j + 1], s, s);
        }
        if (!early) {
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(tempty_bar(buf));
        }
      }
      if (m_ok && n_ok) {
        uint8_t* dstp = reinterpret_cast<uint8_t*>(g.out) + ((int64_t)m * g.N + n0) * 2;
#pragma unroll
        for (int c0 = 0; c0 < 128; c0 += 8) {
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) v[j] = has_bias ? __fadd_rn(facc[c0 + j], load_bias(g.bias, n0 + c0 + j, g.out_is_bf16)) : facc[c0 + j];
          uint32_t p[4];
#pragma unroll
          for (int j = 0; j < 4; ++j)
            p[j] = g.out_is_bf16 ? pack2<__nv_bfloat16>(v[2 * j], v[2 * j + 1]) : pack2<__half>(v[2 * j], v[2 * j + 1]);
          *reinterpret_cast<uint4*>(dstp + c0 * 2) = make_uint4(p[0], p[1], p[2], p[3]);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

// ---- host side -------------------------------------------------------------------------------------------------------------
template <int BN, int EW>
static int launch_blk_sk(const CUtensorMap& tw, const CUtensorMap& tx, const GemmArgs& g, int total_units, int ctas,
                         void* ws, cudaStream_t st) {
  using L = BlkSk<BN>;
  auto kern = gemm_block_streamk_kernel<BN, EW>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  uint32_t* flags = reinterpret_cast<uint32_t*>(ws);
  float* partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + kSkFlagBytes);
  return launch_k(kern, dim3((unsigned)ctas), dim3(128 + 32 * EW), L::kBytes, st, 1, true, tw, tx, g, total_units,
                  flags, partials);
}

int launch_block_fp8_streamk(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, void* ws,
                             cudaStream_t st) {
  const int bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  const int64_t units = ceil_div(N, 128) * g.num_kb;
  B200_REQUIRE(units < (1ll << 30), "fp8_blockwise_scaled_mm: too many work units");
  const int sms = sm_count();
  const int tiles = (int)ceil_div(N, 128);
  int ctas = (int)(units < sms ? units : sms);
  if (tiles < sms) {  // same grid rule as the int8 stream-K path (gemm_umma.cu)
    const int per_tile = sms / tiles < g.num_kb ? sms / tiles : g.num_kb;
    if (tiles * per_tile * 100 >= sms * 85) ctas = tiles * per_tile;
  }
  CUtensorMap tw, tx;
  int rc;
  if ((rc = cached_tmap(&tw, b, (uint64_t)N, (uint64_t)K, 128)) != B200_OK) return rc;
  if ((rc = cached_tmap(&tx, a, (uint64_t)M, (uint64_t)K, (uint32_t)bn)) != B200_OK) return rc;
  // two promotion warps per TMEM lane quarter from 33 tokens up (B200_BLK_SK_EW=4 keeps one, for A/B timing)
  static const int ew_env = [] {
    const char* e = std::getenv("B200_BLK_SK_EW");
    return e ? std::atoi(e) : 0;
  }();
  const bool wide = ew_env == 8 || (ew_env != 4 && bn >= 64);
  switch (bn) {
    case 16: return launch_blk_sk<16, 4>(tw, tx, g, (int)units, ctas, ws, st);
    case 32: return wide ? launch_blk_sk<32, 8>(tw, tx, g, (int)units, ctas, ws, st)
                         : launch_blk_sk<32, 4>(tw, tx, g, (int)units, ctas, ws, st);
    case 64: return wide ? launch_blk_sk<64, 8>(tw, tx, g, (int)units, ctas, ws, st)
                         : launch_blk_sk<64, 4>(tw, tx, g, (int)units, ctas, ws, st);
    default: return wide ? launch_blk_sk<128, 8>(tw, tx, g, (int)units, ctas, ws, st)
                         : launch_blk_sk<128, 4>(tw, tx, g, (int)units, ctas, ws, st);
  }
}

int launch_block_fp8_prefill(GemmArgs g, const void* a, const void* b, int64_t M, int64_t N, int64_t K, cudaStream_t st) {
  CUtensorMap ta, tb;
  int rc;
  if ((rc = cached_tmap(&ta, a, (uint64_t)M, (uint64_t)K, 128)) != B200_OK) return rc;
  if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, (uint32_t)kBpBN)) != B200_OK) return rc;
  const int tiles_m = (int)ceil_div(M, 128);
  const int num_tiles = tiles_m * (int)ceil_div(N, kBpBN);
  auto kern = gemm_block_prefill_kernel;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBpBytes));
  const int ctas = num_tiles < sm_count() ? num_tiles : sm_count();
  return launch_k(kern, dim3((unsigned)ctas), dim3(kBpThreads), kBpBytes, st, 1, true, ta, tb, g, tiles_m, num_tiles);
}

}  // namespace b200
