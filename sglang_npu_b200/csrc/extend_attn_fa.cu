// Paged extend / prefill attention, second generation: two query tiles per CTA, K/V and Q fed by TMA, P kept in
// tensor memory.
//
// Same op as extend_attn_umma.cu (reference triton_ops/extend_attention.py:302-440, oracle
// torch_native_backend.py:27-110); used when every KV position is read through the page table (k_new == NULL: the
// chunk's K/V were written to the cache first, which is what set_kv_buffer / the fused rope + KV-write launch do).
//
// Why a second kernel: extend_attn_umma.cu runs ONE softmax warpgroup next to the tensor pipe.  A 128x128 score tile
// costs the pipe ~1050 cycles (Q K^T + P V) and the warpgroup >= 1024 cycles of MUFU.EX2 alone (16 / clk / SM), plus
// its FMA / conversion / shared-memory work with a single warp per scheduler to hide latency — the pipe sat at 38 %
// (profiles/r01h_extend_attn_ncu_full.txt).  Here:
//   * one CTA = 256 query rows of one (request, q head) = two 128-row tiles that share every K/V tile, with one
//     softmax warpgroup per tile (two warps per scheduler);
//   * KV tiles of 64 positions and TWO S buffers per query tile: Q K^T of tile j+1 is issued before P V of tile j,
//     so a softmax warpgroup finds its next S tile waiting and the tensor pipe finds the next P waiting — the
//     barrier round trips (commit -> softmax -> arrive -> issue) are off the critical path.  (A first version with
//     128-position tiles and a single S buffer per query tile had each warpgroup wait for its own P V + Q K^T
//     block: tensor pipe 48 % active, MUFU 48 %, neither saturated — profiles/r02i_extend_fa_ncu.txt.)
//   * TMEM (512 columns): S0[2] | S1[2] | O0 | O1 (64 + 64 + 64 + 64 + 128 + 128).  P (bf16/f16) is written back over
//     the first 32 columns of its own S buffer with tcgen05.st and is the A operand of P V straight from tensor
//     memory — no shared-memory P tile, no bank conflicts, its shared memory goes to the K/V rings (4 K + 4 V stages);
//   * the row of S (64 fp32) is held in registers (setmaxnreg: 232 for the softmax warps, 40 for the rest), so the row
//     max and the exponentials read TMEM once; the max a row is normalised to only moves when the tile max exceeds
//     it by 2^8 (O is then rescaled in TMEM by the same thread; rare after the first tiles);
//   * a fixed fraction of the exponentials is evaluated on the FMA pipe (Cody-Waite + cubic, 7.5e-5 relative —
//     below the rounding of P to bf16/f16) because MUFU.EX2 throughput alone equals the tensor time of the tile;
//     packed fp32x2 FMA / ADD halve the issue slots of the scale and row-sum steps;
//   * Q and K/V arrive by cp.async.bulk.tensor (128-byte swizzle, the UMMA operand layout): each 16-token run of a
//     KV tile whose cache slots are consecutive (any page_size >= 16) is ONE [16 x 128 B] box per 64-element
//     column block, anything else (page_size 1, a tile's ragged tail) falls back to [1 x 128 B] boxes; rows past
//     the end of the sequence are zero-filled by the producer warp, so V never feeds stale NaNs into P V.  All 32
//     lanes of the producer warp issue copies; the page ids of a tile come from four coalesced loads.
// Warps: 0 = TMA producer, 1 / 2 = MMA issuer of query tile 0 / 1 (1 also allocates TMEM), 3 idle, 4-7 softmax / correction / epilogue of query
// tile 0, 8-11 of query tile 1 (TMEM lane quarter = warp % 4).
#include <cstdlib>
#include <mutex>
#include <type_traits>
#include <unordered_map>

#include "extend_attn.cuh"
#include "sm100.cuh"

namespace b200 {

using namespace sm100;

namespace {

constexpr int kFaBM = 128;       // query rows per tile (two tiles per CTA)
constexpr int kFaBN = 64;        // KV positions per tile
constexpr int kFaThreads = 384;
constexpr int kFaKStages = 4;
constexpr int kFaVStages = 4;
constexpr float kFaRescale = 8.f;  // log2 domain
constexpr int kFaPolyDefault = 2;   // of every 8 exponentials, this many run on the FMA pipe (B200_FA_POLY overrides)

template <int D>
struct FaSmem {
  static constexpr int kQBlk = kFaBM * 128;          // one [128 rows x 128 B] column block of a Q tile
  static constexpr int kQTile = kFaBM * D * 2;
  static constexpr int kBlk = kFaBN * 128;           // one [64 rows x 128 B] column block of a K / V tile
  static constexpr int kTile = kFaBN * D * 2;        // K / V tile bytes
  static constexpr int kQ = 0;                       // two query tiles
  static constexpr int kK = 2 * kQTile;
  static constexpr int kV = kK + kFaKStages * kTile;
  static constexpr int kBar = kV + kFaVStages * kTile;
  static constexpr int kBytes = kBar + 40 * 8 + 1024;
};

template <int N>
__device__ __forceinline__ void fa_reg_dealloc() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void fa_reg_alloc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(N)); }

// (d0, d1) = (a0, a1) * (b, b) + (c, c) as one packed FFMA2
__device__ __forceinline__ void fa_fma2(float& d0, float& d1, float a0, float a1, float b, float c) {
#ifdef B200_FA_SCALAR_MATH
  d0 = fmaf(a0, b, c);
  d1 = fmaf(a1, b, c);
  return;
#endif
  asm("{\n.reg .b64 va, vb, vc;\n"
      "mov.b64 va, {%2, %3};\nmov.b64 vb, {%4, %4};\nmov.b64 vc, {%5, %5};\n"
      "fma.rn.f32x2 vc, va, vb, vc;\n"
      "mov.b64 {%0, %1}, vc;\n}\n"
      : "=f"(d0), "=f"(d1)
      : "f"(a0), "f"(a1), "f"(b), "f"(c));
}
__device__ __forceinline__ void fa_add2(float& s0, float& s1, float a0, float a1) {
#ifdef B200_FA_SCALAR_MATH
  s0 += a0;
  s1 += a1;
  return;
#endif
  asm("{\n.reg .b64 va, vs;\n"
      "mov.b64 va, {%2, %3};\nmov.b64 vs, {%0, %1};\n"
      "add.rn.f32x2 vs, vs, va;\n"
      "mov.b64 {%0, %1}, vs;\n}\n"
      : "+f"(s0), "+f"(s1)
      : "f"(a0), "f"(a1));
}
__device__ __forceinline__ float fa_max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;\n" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}
// 2^x on the FMA pipe: x = n + f, n = round(x), f in [-0.5, 0.5]; 2^f by a cubic (max relative error 7.5e-5), n added
// to the exponent field.  x <= 127; x below -126 (masked scores are -inf) is clamped: the result is ~1e-38.
__device__ __forceinline__ float fa_exp2_fma(float x) {
  x = fmaxf(x, -126.f);
  const float xf = __fadd_rn(x, 12582912.f);   // 1.5 * 2^23: the low mantissa bits now hold n (two's complement)
  const float n = __fsub_rn(xf, 12582912.f);
  const float f = __fsub_rn(x, n);
  float p = fmaf(0.0551716685f, f, 0.2426111251f);
  p = fmaf(p, f, 0.6932609677f);
  p = fmaf(p, f, 0.9999280572f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(xf) << 23));
}

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
      "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};\n" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
      "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
      "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
      "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
      : "memory");
}

// tcgen05.mma, A operand in tensor memory (rows on the 128 lanes, K packed two 16-bit elements per column)
__device__ __forceinline__ void fa_umma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  const uint32_t acc = accumulate ? 1u : 0u;
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}

__device__ __forceinline__ void st_shared_zero16(uint32_t addr) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%1,%1,%1};\n" ::"r"(addr), "r"(0u) : "memory");
}

}  // namespace

template <typename T, int D, int POLY>
__global__ void __launch_bounds__(kFaThreads, 1)
extend_attn_fa_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k16,
                      const __grid_constant__ CUtensorMap tm_k1, const __grid_constant__ CUtensorMap tm_v16,
                      const __grid_constant__ CUtensorMap tm_v1, const ExtendArgs a) {
  using L = FaSmem<D>;
  constexpr int kFmt = std::is_same<T, __half>::value ? 0 : 1;
  constexpr int kBlocks = D / 64;  // 128-byte column blocks per row
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBar;
  auto k_full = [&](int s) { return bar + s * 8; };
  auto k_empty = [&](int s) { return bar + (kFaKStages + s) * 8; };
  auto v_full = [&](int s) { return bar + (2 * kFaKStages + s) * 8; };
  auto v_empty = [&](int s) { return bar + (2 * kFaKStages + kFaVStages + s) * 8; };
  constexpr int kB0 = 2 * kFaKStages + 2 * kFaVStages;
  // per (query tile, S buffer): tile j uses buffer j & 1, phase (j >> 1) & 1 — with one barrier per query tile a
  // warpgroup that is a tile ahead could complete two phases before the other side tests the first
  auto s_full = [&](int t, int j) { return bar + (kB0 + 2 * t + (j & 1)) * 8; };
  auto p_full = [&](int t, int j) { return bar + (kB0 + 4 + 2 * t + (j & 1)) * 8; };
  auto pv_done = [&](int t) { return bar + (kB0 + 8 + t) * 8; };  // P V of tile j: phase j (completions cannot run ahead)
  const uint32_t q_full = bar + (kB0 + 10) * 8;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int mblk = gridDim.x - 1 - blockIdx.x;  // heaviest (longest causal span) blocks first
  const int head = blockIdx.y;
  const int b = blockIdx.z;

  pdl_trigger();
  pdl_wait();
  // Segment mode (shared-prefix decode, decode_attn.cu): blockIdx.z = segment s of request 0's prefix; the query rows
  // are the decode batch (one row per request), the KV range is [s * seg_len, min(prefix, (s+1) * seg_len)) of request
  // 0's page-table row, no causal mask, and the result is a normalised fp32 partial + log2-LSE in the decode merge's
  // workspace instead of o.
  const bool seg_mode = a.part_o != nullptr;
  const int rb = seg_mode ? 0 : b;
  const int q_start = seg_mode ? 0 : a.qo_indptr[rb];
  const int ext_len = seg_mode ? a.seg_rows : a.qo_indptr[rb + 1] - q_start;
  const int m0 = mblk * (2 * kFaBM);
  if (m0 >= ext_len) return;  // CTA-uniform, before any barrier / TMEM use
  const int kv_begin = seg_mode ? b * a.seg_len : 0;
  const int prefix = seg_mode ? min(a.seg_prefix - kv_begin, a.seg_len) : a.prefix_lens[b];
  if (seg_mode && prefix <= 0) return;
  const int kv_total = seg_mode ? prefix : prefix + ext_len;
  int kv_end_t[2], n_t[2];
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    const int r0 = m0 + t * kFaBM;
    kv_end_t[t] = a.causal ? min(kv_total, prefix + r0 + kFaBM) : kv_total;
    n_t[t] = r0 < ext_len ? (kv_end_t[t] + kFaBN - 1) / kFaBN : 0;
  }
  const int n_max = max(n_t[0], n_t[1]);
  const int kv_end_max = n_t[1] > 0 ? max(kv_end_t[0], kv_end_t[1]) : kv_end_t[0];
  // Both query tiles walk the same n_max KV tiles (tile 0's last two lie entirely above its causal diagonal: every
  // score is masked, P = 0): each K / V stage then has exactly one reader per valid query tile, and the two MMA
  // warps need no per-tile bookkeeping about who reads what.
  const int n_valid = n_t[1] > 0 ? 2 : 1;
  n_t[0] = n_max;
  const int kvh = head / (a.num_q_heads / a.num_kv_heads);

  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < kFaKStages; ++s) {
        mbar_init(k_full(s), 1);
        mbar_init(k_empty(s), n_valid);  // one commit per MMA warp that reads the stage
      }
      for (int s = 0; s < kFaVStages; ++s) {
        mbar_init(v_full(s), 1);
        mbar_init(v_empty(s), n_valid);
      }
      for (int t = 0; t < 2; ++t) {
        for (int bf = 0; bf < 2; ++bf) {
          mbar_init(s_full(t, bf), 1);
          mbar_init(p_full(t, bf), 4);  // one arrive per softmax warp
        }
        mbar_init(pv_done(t), 1);
      }
      mbar_init(q_full, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc<512>(smem_u32(&s_tmem_base));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  constexpr uint32_t kOCol = 256;  // S_t[buf] / P_t[buf] at column 128 t + 64 buf, O0 at 256, O1 at 256 + D

  if (warp < 4) {
    fa_reg_dealloc<40>();  // 128 x (168 - 40) registers returned = 256 x (232 - 168) taken below
    if (warp == 0) {
      // ---- TMA producer ----------------------------------------------------------------------------------------
      const int64_t req = load_index(a.req_pool_indices, rb, a.index_is_int64);
      const int32_t* r2t_row = a.req_to_token + req * a.req_to_token_stride + kv_begin;
      if (lane == 0) {
        prefetch_tmap(&tm_q);
        prefetch_tmap(&tm_k16);
        prefetch_tmap(&tm_v16);
        mbar_expect_tx(q_full, (uint32_t)((n_t[1] > 0 ? 2 : 1) * L::kQTile));
        for (int t = 0; t < (n_t[1] > 0 ? 2 : 1); ++t)
#pragma unroll
          for (int blk = 0; blk < kBlocks; ++blk)
            tma_load_2d(base + L::kQ + t * L::kQTile + blk * L::kQBlk, &tm_q, q_full, head * D * 2 + blk * 128,
                        q_start + m0 + t * kFaBM, kEvictFirst);
      }
      const int col0 = kvh * D * 2;
      // page-table slice of a tile: lane l holds the cache slots of rows l and 32 + l; fetched one tile ahead so the
      // global-load latency is off the issue path
      auto load_slots = [&](int j, int32_t (&sl)[2]) {
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int pos = j * kFaBN + 32 * g + lane;
          sl[g] = pos < kv_end_max ? __ldg(r2t_row + pos) : 0;
        }
      };
      int32_t sl_next[2];
      load_slots(0, sl_next);
      for (int j = 0; j < n_max; ++j) {
        const int kv0 = j * kFaBN;
        int32_t sl[2] = {sl_next[0], sl_next[1]};
        if (j + 1 < n_max) load_slots(j + 1, sl_next);
        bool vld[2];
#pragma unroll
        for (int g = 0; g < 2; ++g) vld[g] = kv0 + 32 * g + lane < kv_end_max;
        // 16-row runs with consecutive slots become one box each
        uint32_t run16 = 0;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int32_t first = __shfl_sync(0xffffffffu, sl[g], lane & 16);
          const uint32_t bal = __ballot_sync(0xffffffffu, vld[g] && sl[g] == first + (lane & 15));
          if ((bal & 0xFFFFu) == 0xFFFFu) run16 |= 1u << (2 * g);
          if ((bal >> 16) == 0xFFFFu) run16 |= 1u << (2 * g + 1);
        }
        int32_t first_slot[4];  // slot of the first row of each run, in every lane
#pragma unroll
        for (int r = 0; r < 4; ++r) first_slot[r] = __shfl_sync(0xffffffffu, sl[r >> 1], (r & 1) * 16);
        const int rows_valid = min(kFaBN, kv_end_max - kv0);
        const uint32_t bytes = (uint32_t)rows_valid * D * 2;
#pragma unroll
        for (int which = 0; which < 2; ++which) {  // 0: K tile j, 1: V tile j
          const int st = j & 3;
          const uint32_t full = which ? v_full(st) : k_full(st);
          const uint32_t empty = which ? v_empty(st) : k_empty(st);
          const uint32_t dst = base + (which ? L::kV : L::kK) + st * L::kTile;
          const CUtensorMap* m16 = which ? &tm_v16 : &tm_k16;
          const CUtensorMap* m1 = which ? &tm_v1 : &tm_k1;
          mbar_wait(empty, ((j >> 2) & 1) ^ 1);
          if (run16 == 0xFu) {
            // the common case (page_size >= 16, full tile): eight boxes issued straight by one elected lane
            if (elect_one()) {
              mbar_expect_tx(full, (uint32_t)L::kTile);
#pragma unroll
              for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int blk = 0; blk < kBlocks; ++blk)
                  tma_load_2d(dst + blk * L::kBlk + r * 16 * 128, m16, full, col0 + blk * 128, first_slot[r], kEvictNormal);
            }
            __syncwarp();
            continue;
          }
          if (rows_valid < kFaBN) {  // rows past the end of the sequence: zeros (generic proxy), then publish
#pragma unroll
            for (int g = 0; g < 2; ++g)
              if (!vld[g])
#pragma unroll
                for (int blk = 0; blk < kBlocks; ++blk)
#pragma unroll
                  for (int c = 0; c < 8; ++c) st_shared_zero16(dst + blk * L::kBlk + (32 * g + lane) * 128 + c * 16);
            fence_proxy_async();
          }
          __syncwarp();
          if (lane == 0) mbar_expect_tx(full, bytes);
          __syncwarp();
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const int r = 2 * g + (lane >> 4);
            if ((run16 >> r) & 1u) {
              if ((lane & 15) == 0) {
#pragma unroll
                for (int blk = 0; blk < kBlocks; ++blk)
                  tma_load_2d(dst + blk * L::kBlk + (32 * g + lane) * 128, m16, full, col0 + blk * 128, sl[g], kEvictNormal);
              }
            } else if (vld[g]) {
#pragma unroll
              for (int blk = 0; blk < kBlocks; ++blk)
                tma_load_2d(dst + blk * L::kBlk + (32 * g + lane) * 128, m1, full, col0 + blk * 128, sl[g], kEvictNormal);
            }
          }
        }
      }
    } else if (warp - 1 < n_valid) {
      // ---- MMA issuers: warp 1 for query tile 0, warp 2 for query tile 1 (two schedulers: a single warp spent ~370
      //      instructions per KV tile on waits / descriptors / issue and could not keep the pipe fed).  Per tile j:
      //      Q_t K(j+1)^T -> P_t(j) V(j); tcgen05 ops of one thread execute in order, which is all the S / P buffer
      //      reuse needs.  The whole warp runs the loop and the waits, one elected lane issues.
      const int t = warp - 1;
      constexpr uint32_t idesc_s = make_idesc(/*f32*/ 1, kFmt, kFmt, 128, kFaBN);
      constexpr uint32_t idesc_o = make_idesc(/*f32*/ 1, kFmt, kFmt, 128, D, /*A K-major*/ 0, /*B MN-major*/ 1);
      static_assert(kFaKStages == 4 && kFaVStages == 4, "stage index / phase below use j & 3, j >> 2");
      const uint64_t dq = smem_desc_k_sw128(base + L::kQ + t * L::kQTile);
      const uint64_t dk = smem_desc_k_sw128(base + L::kK);                  // + stage * (kTile >> 4)
      // V tile: atoms of [8 kv x 64 d]; next 64 d = kBlk bytes (LBO), next 8 kv = 1024 bytes (SBO)
      const uint64_t dv = smem_desc_mn_sw128(base + L::kV, L::kBlk, 1024);
      const uint32_t s_tmem0 = tmem_base + (uint32_t)(t * 128);
      const uint32_t o_tmem = tmem_base + kOCol + (uint32_t)t * D;
      mbar_wait(q_full, 0);
      auto issue_qk = [&](int j) {  // S_t[j & 1] = Q_t K(j)^T; the buffer's previous reader P V(j-2) was issued earlier
        const int ks = j & 3;
        mbar_wait(k_full(ks), (j >> 2) & 1);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t db0 = dk + (uint64_t)(ks * (L::kTile >> 4));
          const uint32_t s_tmem = s_tmem0 + (uint32_t)((j & 1) * kFaBN);
#pragma unroll
          for (int kk = 0; kk < D / 16; ++kk)  // descriptor address field counts 16-byte units
            umma<kKindF16>(s_tmem, dq + (uint64_t)((kk >> 2) * (L::kQBlk >> 4) + 2 * (kk & 3)),
                           db0 + (uint64_t)((kk >> 2) * (L::kBlk >> 4) + 2 * (kk & 3)), idesc_s, kk != 0);
          tc_commit(s_full(t, j));
          tc_commit(k_empty(ks));
        }
        __syncwarp();
      };
      issue_qk(0);
      for (int j = 0; j < n_max; ++j) {
        if (j + 1 < n_max) issue_qk(j + 1);
        const int vs = j & 3;
        mbar_wait(v_full(vs), (j >> 2) & 1);
        mbar_wait(p_full(t, j), (j >> 1) & 1);
        tc_fence_after();
        if (elect_one()) {  // O_t += P_t(j) V(j)
          const uint64_t db0 = dv + (uint64_t)(vs * (L::kTile >> 4));
          const uint32_t p_tmem = s_tmem0 + (uint32_t)((j & 1) * kFaBN);
#pragma unroll
          for (int kk = 0; kk < kFaBN / 16; ++kk)
            fa_umma_ts(o_tmem, p_tmem + kk * 8, db0 + (uint64_t)(kk * (2048 >> 4)), idesc_o, (j | kk) != 0);
          tc_commit(pv_done(t));
          tc_commit(v_empty(vs));
        }
        __syncwarp();
      }
    }
  } else {
    fa_reg_alloc<232>();
    // ---- softmax / correction / epilogue: thread = query row = TMEM lane ------------------------------------------
    const int t = (warp - 4) >> 2;
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const int r0 = m0 + t * kFaBM;
    const int n_tiles = n_t[t];
    const int kv_end = kv_end_t[t];
    const uint32_t t_lane = tmem_base + ((uint32_t)(qd * 32) << 16);
    const uint32_t t_o = t_lane + kOCol + (uint32_t)t * D;
    const int lim = min(a.causal ? prefix + r0 + row : kv_total - 1, kv_end - 1);  // last KV position this row may see
    const bool use_cap = a.logit_cap > 0.f;
    const float cap_in = use_cap ? a.sm_scale / a.logit_cap : 0.f;
    const float cap_out = use_cap ? a.logit_cap / a.sm_scale : 0.f;
    const float sc = a.scale_log2;
    float m_used = -INFINITY;  // the max (log2 domain) the row's P and O are normalised to
    float l0 = 0.f, l1 = 0.f;  // running sum, two partial sums (packed adds)

    for (int j = 0; j < n_tiles; ++j) {
      const int kv0 = j * kFaBN;
      const bool need_mask = (kv0 + kFaBN > kv_end) || (a.causal && kv0 + kFaBN - 1 > prefix + r0);
      const uint32_t t_s = t_lane + (uint32_t)(t * 128 + (j & 1) * kFaBN);
      mbar_wait(s_full(t, j), (j >> 1) & 1);
      tc_fence_after();
      uint32_t s[kFaBN];
      {
        uint32_t(&s0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[0]);
        uint32_t(&s1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&s[32]);
        tmem_ld32(t_s, s0);
        tmem_ld32(t_s + 32, s1);
        tmem_ld_wait();
      }
      if (use_cap) {
#pragma unroll
        for (int i = 0; i < kFaBN; ++i) s[i] = __float_as_uint(cap_out * tanh_approx(__uint_as_float(s[i]) * cap_in));
      }
      if (need_mask) {
#pragma unroll
        for (int i = 0; i < kFaBN; ++i)
          if (kv0 + i > lim) s[i] = 0xff800000u;  // -inf
      }
      float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
      for (int i = 0; i < kFaBN; i += 4) {
        mx0 = fa_max3(mx0, __uint_as_float(s[i]), __uint_as_float(s[i + 1]));
        mx1 = fa_max3(mx1, __uint_as_float(s[i + 2]), __uint_as_float(s[i + 3]));
      }
      const float m_tile = fmaxf(mx0, mx1) * sc;  // sm_scale > 0: max commutes with the scaling
      const bool grow = m_tile > m_used + kFaRescale;  // also the first finite tile (m_used = -inf)
      if (__any_sync(0xffffffffu, grow)) {
        if (j > 0) {
          // O is touched only here: P V of tile j-1 must have retired (P V of tile j waits for p_full below)
          mbar_wait(pv_done(t), (j - 1) & 1);
          tc_fence_after();
          const float alpha = grow ? fast_exp2(m_used - m_tile) : 1.f;
          l0 *= alpha;
          l1 *= alpha;
#pragma unroll 1
          for (int c0 = 0; c0 < D; c0 += 16) {
            uint32_t r[16];
            tmem_ld16(t_o + c0, r);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 16; ++i) r[i] = __float_as_uint(__uint_as_float(r[i]) * alpha);
            tmem_st16(t_o + c0, r);
          }
        }
        if (grow) m_used = m_tile;
      }
      const float neg_m = (m_used == -INFINITY) ? 0.f : -m_used;
#pragma unroll
      for (int c = 0; c < kFaBN / 32; ++c) {
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float a0, a1;
          fa_fma2(a0, a1, __uint_as_float(s[32 * c + i]), __uint_as_float(s[32 * c + i + 1]), sc, neg_m);
          float p0, p1;
          if ((i & 7) < POLY) {
            p0 = fa_exp2_fma(a0);
            p1 = fa_exp2_fma(a1);
          } else {
            p0 = fast_exp2(a0);
            p1 = fast_exp2(a1);
          }
          fa_add2(l0, l1, p0, p1);
          w[i >> 1] = pack2<T>(p0, p1);
        }
        tmem_st16(t_s + 16 * c, w);  // P over the columns of S already consumed (16 c + 16 <= 32 c + 32)
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(p_full(t, j));
    }

    // ---- epilogue: O / l -> global ----------------------------------------------------------------------------------
    if (n_tiles > 0) {
      // a parity wait is only exact one phase ahead: S of the last tile proves P V(n-3) retired, not P V(n-2)
      if (n_tiles >= 2) mbar_wait(pv_done(t), (n_tiles - 2) & 1);
      mbar_wait(pv_done(t), (n_tiles - 1) & 1);
      tc_fence_after();
      const float l_run = l0 + l1;
      const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
      const bool row_ok = r0 + row < ext_len;
      if (seg_mode) {
        const int64_t pidx = ((int64_t)(r0 + row) * a.num_q_heads + head) * a.part_stride + b;
        float* prow = a.part_o + pidx * D;
        if (row_ok) a.part_lse[pidx] = l_run > 0.f ? m_used + log2f(l_run) : -INFINITY;
#pragma unroll 1
        for (int c0 = 0; c0 < D; c0 += 32) {
          uint32_t r[32];
          tmem_ld32(t_o + c0, r);
          tmem_ld_wait();
          if (row_ok) {
#pragma unroll
            for (int i = 0; i < 32; i += 4)
              *reinterpret_cast<float4*>(prow + c0 + i) =
                  make_float4(__uint_as_float(r[i]) * inv, __uint_as_float(r[i + 1]) * inv, __uint_as_float(r[i + 2]) * inv,
                              __uint_as_float(r[i + 3]) * inv);
          }
        }
      }
      T* orow = reinterpret_cast<T*>(a.o) + (int64_t)(q_start + r0 + row) * a.o_token_stride + (int64_t)head * D;
      const bool aligned = (reinterpret_cast<uintptr_t>(orow) & 15) == 0;
#pragma unroll 1
      for (int c0 = 0; c0 < D && !seg_mode; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(t_o + c0, r);
        tmem_ld_wait();
        if (row_ok) {
          uint32_t w[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) w[i] = pack2<T>(__uint_as_float(r[2 * i]) * inv, __uint_as_float(r[2 * i + 1]) * inv);
          if (aligned) {
#pragma unroll
            for (int i = 0; i < 4; ++i)
              *reinterpret_cast<uint4*>(orow + c0 + 8 * i) = make_uint4(w[4 * i], w[4 * i + 1], w[4 * i + 2], w[4 * i + 3]);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) *reinterpret_cast<uint32_t*>(orow + c0 + 2 * i) = w[i];
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ---- host ----------------------------------------------------------------------------------------------------------
namespace {

struct FaKey {
  const void* base;
  uint64_t rows, row_bytes, pitch;
  uint32_t box_rows;
  bool operator==(const FaKey& o) const {
    return base == o.base && rows == o.rows && row_bytes == o.row_bytes && pitch == o.pitch && box_rows == o.box_rows;
  }
};
struct FaKeyHash {
  size_t operator()(const FaKey& k) const {
    size_t h = std::hash<const void*>()(k.base);
    h ^= std::hash<uint64_t>()(k.rows * 1315423911ull + k.row_bytes * 2654435761ull + k.pitch * 40503ull + k.box_rows);
    return h;
  }
};

int fa_tmap(CUtensorMap* out, const void* base, uint64_t rows, uint64_t row_bytes, uint64_t pitch, uint32_t box_rows) {
  static thread_local std::unordered_map<FaKey, CUtensorMap, FaKeyHash> cache;
  FaKey key{base, rows, row_bytes, pitch, box_rows};
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return B200_OK;
  }
  int st = make_tmap_2d_bytes(out, base, rows, row_bytes, pitch, box_rows, 128, 128);
  if (st != B200_OK) return st;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return B200_OK;
}

template <typename T, int D, int POLY>
int launch_fa(const ExtendArgs& a, int max_extend_len, int64_t num_q_tokens, int64_t num_kv_slots, cudaStream_t stream) {
  using L = FaSmem<D>;
  CUtensorMap tq, tk16, tk1, tv16, tv1;
  const uint64_t q_row = (uint64_t)a.num_q_heads * D * 2, kv_row = (uint64_t)a.num_kv_heads * D * 2;
  int st = fa_tmap(&tq, a.q, (uint64_t)num_q_tokens, q_row, (uint64_t)a.q_token_stride * 2, 128);
  if (st == B200_OK) st = fa_tmap(&tk16, a.k_cache, (uint64_t)num_kv_slots, kv_row, (uint64_t)a.kv_token_stride * 2, 16);
  if (st == B200_OK) st = fa_tmap(&tk1, a.k_cache, (uint64_t)num_kv_slots, kv_row, (uint64_t)a.kv_token_stride * 2, 1);
  if (st == B200_OK) st = fa_tmap(&tv16, a.v_cache, (uint64_t)num_kv_slots, kv_row, (uint64_t)a.kv_token_stride * 2, 16);
  if (st == B200_OK) st = fa_tmap(&tv1, a.v_cache, (uint64_t)num_kv_slots, kv_row, (uint64_t)a.kv_token_stride * 2, 1);
  if (st != B200_OK) return st;
  auto kern = extend_attn_fa_kernel<T, D, POLY>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  dim3 grid((max_extend_len + 2 * kFaBM - 1) / (2 * kFaBM), a.num_q_heads, a.batch);  // segment mode: batch = segments
  return launch_k(kern, grid, dim3(kFaThreads), L::kBytes, stream, 1, true, tq, tk16, tk1, tv16, tv1, a);
}

}  // namespace

// Two-query-tile TMA kernel.  Covers head_dim 128 and every KV position read through the page table (k_new == NULL),
// incl. the segment mode of the shared-prefix decode; anything else returns B200_ERR_UNSUPPORTED and the caller takes
// extend_attn_umma.cu.
int launch_extend_fa(const ExtendArgs& a, int max_extend_len, int head_dim, int dtype, int64_t num_q_tokens,
                     int64_t num_kv_slots, cudaStream_t stream) {
  if (a.k_new != nullptr || head_dim != 128 || num_q_tokens <= 0 || num_kv_slots <= 0)
    return B200_ERR_UNSUPPORTED;
  static const int poly = getenv("B200_FA_POLY") ? atoi(getenv("B200_FA_POLY")) : kFaPolyDefault;
#define B200_FA_LAUNCH(TT)                                                                                        \
  switch (poly) {                                                                                                 \
    case 0: return launch_fa<TT, 128, 0>(a, max_extend_len, num_q_tokens, num_kv_slots, stream);                  \
    case 4: return launch_fa<TT, 128, 4>(a, max_extend_len, num_q_tokens, num_kv_slots, stream);                  \
    case 6: return launch_fa<TT, 128, 6>(a, max_extend_len, num_q_tokens, num_kv_slots, stream);                  \
    default: return launch_fa<TT, 128, kFaPolyDefault>(a, max_extend_len, num_q_tokens, num_kv_slots, stream);    \
  }
  if (dtype == B200_BF16) { B200_FA_LAUNCH(__nv_bfloat16) }
  if (dtype == B200_F16) { B200_FA_LAUNCH(__half) }
#undef B200_FA_LAUNCH
  return B200_ERR_UNSUPPORTED;
}

}  // namespace b200
