// AWQ-int4: standalone dequantisation and the fused dequant + tcgen05 GEMM.
//
// Replaces (reference):
//   awq_dequantize  sgl-kernel/csrc/gemm/awq_kernel.cu:126-221  (writes the full fp16 matrix)
//   AWQLinearMethod.apply  python/sglang/srt/layers/quantization/awq.py:401-418
//                   (awq_dequantize -> torch.matmul (cuBLAS fp16) -> bias add: 4.5 B of traffic per weight)
// Semantics follow the oracle awq_dequantize_torch (sgl-kernel/tests/test_awq_dequant.py:9-61):
//   W[k, 8c+i] = (nib(qweight[k,c], ORDER[i]) - nib(qzeros[k/g,c], ORDER[i])) * scales[k/g, 8c+i]
//   ORDER = [0,4,1,5,2,6,3,7]; the subtraction is exact and the product is ONE fp16 (bf16) rounding, so
//   the dequantised weights are bit-identical to the reference's.
//
// The nibble order is what makes the register dequant cheap: output pair (2p, 2p+1) = nibbles (p, p+4),
// i.e. `(word >> 4p) & 0x000f000f` already is the pair as two 16-bit lanes; OR-ing the magic exponent
// (fp16 0x6400 = 1024, bf16 0x4300 = 128) turns it into {magic + q} and the zero point is removed with a
// single packed subtract of {magic + z}.
//
// Fused GEMM (`awq_gemm_kernel`), weight-streaming at 0.5 B/weight, one orientation for decode and prefill:
//   D[n, m] (TMEM, 128 channels x BN tokens) += W^T[n, k] * x[m, k]^T   per 64-wide K block
//   * raw packed words: TMA (64-byte swizzle) into a RAW-deep ring — 4 KB per K block per CTA, evict-first;
//   * the group's scales / zero points of the tile ride in the same ring as 1-D bulk copies (no global-latency
//     stall in the dequant loop);
//   * 4 dequant groups x 2 warps own alternate K blocks (block i -> group i % 4), so four blocks are being
//     dequantised concurrently; each turns its block into the fp16 A operand in shared memory, MN-major
//     SWIZZLE_128B atoms (64 channels x 8 k), one 16-byte store per packed word, bank-conflict free;
//   * tokens x[m, k]: TMA, K-major SWIZZLE_128B (B operand), evict-last;
//   * one thread issues tcgen05.mma kind::f16 (M=128, N=BN, K=16) with fp32 accumulators in TMEM;
//   * split-K over a thread-block cluster (1,1,S): partial accumulators are reduce-scattered through
//     distributed shared memory and summed in rank order (deterministic, no workspace);
//   * epilogue: (+bias) -> fp16/bf16 -> out[m, n].
#include <cstdlib>

#include "sm100.cuh"

namespace b200 {

using namespace sm100;

constexpr int kAwqThreads = 384;   // warp 0 raw-W TMA, 1 MMA, 2 TMEM alloc, 3 x TMA, 4-11 dequant + epilogue
constexpr int kAwqDqThreads = 256;
constexpr int kAwqBK = 64;         // K elements per pipeline stage
constexpr int kAwqMaxGps = 2;      // quantisation groups per K block: group_size >= 32
constexpr int kAwqGroups = 4;      // dequant warp groups (2 warps each) working on different K blocks

template <typename T> struct AwqMagic;
template <> struct AwqMagic<__half> { static constexpr uint32_t kOr = 0x64006400u; static constexpr int kFmt = 0; };
template <> struct AwqMagic<__nv_bfloat16> { static constexpr uint32_t kOr = 0x43004300u; static constexpr int kFmt = 1; };

template <typename T>
__device__ __forceinline__ uint32_t sub2(uint32_t a, uint32_t b) {
  using T2 = typename DT<T>::T2;
  T2 r = __hsub2(*reinterpret_cast<T2*>(&a), *reinterpret_cast<T2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
template <typename T>
__device__ __forceinline__ uint32_t mul2(uint32_t a, uint32_t b) {
  using T2 = typename DT<T>::T2;
  T2 r = __hmul2(*reinterpret_cast<T2*>(&a), *reinterpret_cast<T2*>(&b));
  return *reinterpret_cast<uint32_t*>(&r);
}
// {magic + nibble p, magic + nibble p+4} as a packed pair
template <typename T>
__device__ __forceinline__ uint32_t nib_pair(uint32_t word, int p) {
  return ((word >> (4 * p)) & 0x000f000fu) | AwqMagic<T>::kOr;
}
// 8 dequantised weights (output columns 8c .. 8c+7) of one packed word
template <typename T>
__device__ __forceinline__ uint4 dequant_word(uint32_t q, const uint32_t (&zm)[4], const uint4& s) {
  uint4 w;
  w.x = mul2<T>(sub2<T>(nib_pair<T>(q, 0), zm[0]), s.x);
  w.y = mul2<T>(sub2<T>(nib_pair<T>(q, 1), zm[1]), s.y);
  w.z = mul2<T>(sub2<T>(nib_pair<T>(q, 2), zm[2]), s.z);
  w.w = mul2<T>(sub2<T>(nib_pair<T>(q, 3), zm[3]), s.w);
  return w;
}

// ---- standalone dequantisation (parity surface of sgl_kernel.awq_dequantize) ----------------------------
template <typename T>
__global__ void __launch_bounds__(256) awq_dequant_kernel(T* __restrict__ out, const uint32_t* __restrict__ qweight,
                                                          const T* __restrict__ scales,
                                                          const uint32_t* __restrict__ qzeros, int64_t K, int64_t n8,
                                                          int group_size) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // one packed word per thread
  if (idx >= K * n8) return;
  const int64_t k = idx / n8;
  const int64_t c = idx - k * n8;
  const int64_t grp = k / group_size;
  const uint32_t q = qweight[idx];
  const uint32_t z = qzeros[grp * n8 + c];
  const uint4 s = *reinterpret_cast<const uint4*>(scales + (grp * n8 + c) * 8);
  uint32_t zm[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) zm[p] = nib_pair<T>(z, p);
  *reinterpret_cast<uint4*>(out + idx * 8) = dequant_word<T>(q, zm, s);
}

// ---- fused GEMM ----------------------------------------------------------------------------------------------
struct AwqArgs {
  void* out;              // [M, N]
  const void* scales;     // [K/g, N]
  const uint32_t* qzeros; // [K/g, N/8]
  const void* bias;       // [N] or null
  int M, N, K, group_size;
  int num_kb, kb_per_split, splits;
  int cl;  // CTAs along n that share (multicast) the token tile: cluster dims (cl, 1, splits)
};

template <int BN, int STAGES, int RAW>
struct AwqSmem {
  static constexpr int kA = 128 * kAwqBK * 2;  // dequantised weights, MN-major: atom (k_atom * 2 + n_atom) * 1024
  static constexpr int kB = BN * 128;          // tokens, K-major
  static constexpr int kStage = kA + kB;
  static constexpr int kRawStage = kAwqBK * 64;  // 64 k rows x 16 packed words
  static constexpr int kRawOff = STAGES * kStage;
  // per raw stage: up to kAwqMaxGps groups x (128 scales (256 B) + 16 zero-point words (64 B))
  static constexpr int kConstStage = kAwqMaxGps * 320;
  static constexpr int kConstOff = kRawOff + RAW * kRawStage;
  static constexpr int kBarOff = kConstOff + RAW * kConstStage;
  static constexpr int kNumBars = 2 * STAGES + 2 * RAW + 1;
  static constexpr int kBytes = kBarOff + kNumBars * 8 + 1024;
  static_assert(BN * 128 * 4 <= STAGES * kStage, "split-K receive buffer must fit in the operand stages");
  static_assert(kBytes <= 232448, "exceeds the 227 KB shared-memory limit");
};

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts128(uint32_t addr, const uint4& v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};\n" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T, int BN, int STAGES, int RAW>
__global__ void __launch_bounds__(kAwqThreads, (AwqSmem<BN, STAGES, RAW>::kBytes <= 113 * 1024) ? 2 : 1)
awq_gemm_kernel(const __grid_constant__ CUtensorMap tmap_w, const __grid_constant__ CUtensorMap tmap_x, const AwqArgs g) {
  using L = AwqSmem<BN, STAGES, RAW>;
  constexpr int kTmemCols = BN < 32 ? 32 : BN;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  auto op_full = [&](int s) { return bar + s * 8; };
  auto op_empty = [&](int s) { return bar + (STAGES + s) * 8; };
  auto raw_full = [&](int s) { return bar + (2 * STAGES + s) * 8; };
  auto raw_empty = [&](int s) { return bar + (2 * STAGES + RAW + s) * 8; };
  const uint32_t tmem_full = bar + (2 * STAGES + 2 * RAW) * 8;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_w);
    prefetch_tmap(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(op_full(s), 1 + 2);  // x TMA (expect_tx arrive) + the two warps of the owning dequant group
      mbar_init(op_empty(s), g.cl);  // one tcgen05.commit per CTA sharing the token tile
    }
    for (int s = 0; s < RAW; ++s) {
      mbar_init(raw_full(s), 1);
      mbar_init(raw_empty(s), 2);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<kTmemCols>(smem_u32(&s_tmem_base));
  if (warp >= 4) {  // scales / zero points of channels past N (partial last tile) stay zero
    uint32_t* cz = reinterpret_cast<uint32_t*>(smem_raw + (base - smem_u32(smem_raw)) + L::kConstOff);
    for (int i = threadIdx.x - 128; i < RAW * L::kConstStage / 4; i += kAwqDqThreads) cz[i] = 0u;
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  const int CL = g.cl;
  if (CL * g.splits > 1) cluster_sync();  // peers' barriers are initialised before anything is multicast to them
  const int cx = (int)blockIdx.x % CL;                                        // position among the tile sharers
  const uint16_t mc_mask = (uint16_t)(((1u << CL) - 1u) << (CL * blockIdx.z));  // cluster ranks cx' + CL * z

  const int n0 = blockIdx.x * 128;   // first output channel of this CTA
  const int m0 = blockIdx.y * BN;    // first token
  const int rank = (int)blockIdx.z;  // cluster dims (1,1,splits), gridDim.z == splits
  const int S = g.splits;
  const int kb0 = rank * g.kb_per_split;
  const int nkb = min(g.kb_per_split, g.num_kb - kb0);

  // quantisation groups per K block (1 when group_size >= 64, else 64 / group_size) and valid channels of the tile
  const int gps = g.group_size >= kAwqBK ? 1 : kAwqBK / g.group_size;
  const int nv = max(0, min(128, g.N - n0));  // 0 for a surplus CTA that only feeds the multicast
  const int num_groups = g.K / g.group_size;

  if (warp == 0) {
    if (lane == 0) {
      const uint8_t* scales_b = reinterpret_cast<const uint8_t*>(g.scales);
      const uint8_t* zeros_b = reinterpret_cast<const uint8_t*>(g.qzeros);
      for (int i = 0; i < nkb; ++i) {
        const int rs = i % RAW;
        mbar_wait(raw_empty(rs), ((i / RAW) & 1) ^ 1);
        const int k0 = (kb0 + i) * kAwqBK;
        const int grp0 = k0 / g.group_size;
        const int ng = nv > 0 ? min(gps, num_groups - grp0) : 0;  // groups of this block that exist (K tail)
        mbar_expect_tx(raw_full(rs), L::kRawStage + ng * (nv * 2 + nv / 2));
        tma_load_2d(base + L::kRawOff + rs * L::kRawStage, &tmap_w, raw_full(rs), n0 / 2 /* bytes: n0/8 words */, k0,
                    kEvictFirst);
        for (int gi = 0; gi < ng; ++gi) {
          const uint32_t dst = base + L::kConstOff + rs * L::kConstStage + gi * 320;
          bulk_load_1d(dst, scales_b + ((int64_t)(grp0 + gi) * g.N + n0) * 2, nv * 2, raw_full(rs));
          bulk_load_1d(dst + 256, zeros_b + ((int64_t)(grp0 + gi) * g.N + n0) / 2, nv / 2, raw_full(rs));
        }
      }
    }
  } else if (warp == 3) {
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        mbar_wait(op_empty(s), ((i / STAGES) & 1) ^ 1);
        mbar_expect_tx(op_full(s), L::kB);
        if (CL == 1) {
          tma_load_2d(base + s * L::kStage + L::kA, &tmap_x, op_full(s), (kb0 + i) * kAwqBK * 2, m0, kEvictLast);
        } else {  // this CTA fetches rows [cx, cx+1) * BN/CL of the tile and multicasts them to all sharers
          const int rows = BN / CL;
          tma_load_2d_mc(base + s * L::kStage + L::kA + cx * rows * 128, &tmap_x, op_full(s), (kb0 + i) * kAwqBK * 2,
                         m0 + cx * rows, mc_mask, kEvictLast);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(/*f32*/ 1, AwqMagic<T>::kFmt, AwqMagic<T>::kFmt, 128, BN, /*A MN-major*/ 1, 0);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        mbar_wait(op_full(s), (i / STAGES) & 1);
        tc_fence_after();
        const uint32_t sa = base + s * L::kStage;
        const uint64_t da = smem_desc_mn_sw128(sa, /*LBO: next 64 channels*/ 1024, /*SBO: next 8 k*/ 2048);
        const uint64_t db = smem_desc_k_sw128(sa + L::kA);
#pragma unroll
        for (int k = 0; k < kAwqBK / 16; ++k)  // UMMA K = 16 elements: A advances 2 k-atoms (4096 B), B 32 bytes
          umma<kKindF16>(tmem_base, da + 256 * k, db + 2 * k, idesc, (i | k) != 0);
        if (CL == 1) tc_commit(op_empty(s));
        else tc_commit_mc(op_empty(s), mc_mask);  // the stage is free once EVERY sharer's MMAs have read it
      }
      tc_commit(tmem_full);
    }
  } else if (warp >= 4) {
    // ---- dequant producers: group = (warp - 4) / 2 owns K blocks i % 4 == group; within a group the two
    //      warps take alternate 8-row k-atoms; lane -> (row of the atom, 4 consecutive packed words) ------------
    const int dw = warp - 4;
    const int grp_id = dw >> 1;
    const int wsub = dw & 1;
    const int r = lane & 7;
    const int quad = lane >> 3;
    const int c_base = (quad & 1) * 4;
    for (int i = grp_id; i < nkb; i += kAwqGroups) {
      const int rs = i % RAW;
      const int s = i % STAGES;
      mbar_wait(raw_full(rs), (i / RAW) & 1);
      const uint32_t raw_base = base + L::kRawOff + rs * L::kRawStage;
      const uint32_t const_base = base + L::kConstOff + rs * L::kConstStage;
      uint4 q[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        // raw tile: row (64 B), 16-byte chunk = quad, 64-byte swizzle: chunk ^= (row >> 1) & 3
        const int krow = (wsub + 2 * j) * 8 + r;
        q[j] = lds128(raw_base + (uint32_t)(krow * 64 + ((quad ^ ((krow >> 1) & 3)) << 4)));
      }
      uint32_t zm[4][4];
      uint4 sc[4];
      int cur_slot = -1;
      mbar_wait(op_empty(s), ((i / STAGES) & 1) ^ 1);
      const uint32_t a_stage = base + s * L::kStage;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int katom = wsub + 2 * j;
        const int slot = gps == 1 ? 0 : (katom * 8 * gps) >> 6;
        if (slot != cur_slot) {
          cur_slot = slot;
          const uint32_t cb = const_base + slot * 320;
          const uint4 z = lds128(cb + 256 + quad * 16);
          const uint32_t zw[4] = {z.x, z.y, z.z, z.w};
#pragma unroll
          for (int w = 0; w < 4; ++w) {
            sc[w] = lds128(cb + quad * 64 + w * 16);
#pragma unroll
            for (int p = 0; p < 4; ++p) zm[w][p] = nib_pair<T>(zw[w], p);
          }
        }
        // A operand: atom = katom * 2 + (quad >> 1); row r; chunk c = (quad & 1) * 4 + w, swizzled by r
        const uint32_t sa = a_stage + (uint32_t)(((katom * 2 + (quad >> 1)) << 10) + (r << 7));
        sts128(sa + (((c_base + 0) ^ r) << 4), dequant_word<T>(q[j].x, zm[0], sc[0]));
        sts128(sa + (((c_base + 1) ^ r) << 4), dequant_word<T>(q[j].y, zm[1], sc[1]));
        sts128(sa + (((c_base + 2) ^ r) << 4), dequant_word<T>(q[j].z, zm[2], sc[2]));
        sts128(sa + (((c_base + 3) ^ r) << 4), dequant_word<T>(q[j].w, zm[3], sc[3]));
      }
      fence_proxy_async();  // generic-proxy stores -> visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(op_full(s));
        mbar_arrive(raw_empty(rs));
      }
    }
    mbar_wait(tmem_full, 0);
    tc_fence_after();
  }

  // ---- epilogue: warps 4-7 take token columns [0, BN/2), warps 8-11 [BN/2, BN) (BN >= 32) ----------------------
  constexpr int kHalves = BN >= 32 ? 2 : 1;
  constexpr int kColsPerHalf = BN / kHalves;
  const int q4 = warp & 3;
  const int row = q4 * 32 + lane;  // accumulator row (TMEM lane) = output channel n0 + row
  const int half = warp >= 8 ? 1 : 0;
  const bool epi = warp >= 4 && half < kHalves;
  const uint32_t taddr = tmem_base + ((uint32_t)(q4 * 32) << 16);
  const int n = n0 + row;
  const bool n_ok = n < g.N;
  const bool has_bias = g.bias != nullptr;
  float bias = 0.f;
  if (epi && n_ok && has_bias) bias = DT<T>::to_f(reinterpret_cast<const T*>(g.bias)[n]);
  T* out = reinterpret_cast<T*>(g.out);

  if (S > 1) {
    __syncwarp();
    cluster_sync();  // every CTA of the cluster has retired its MMAs: operand stages are free to reuse
    const int C = BN / S;  // token columns owned by each rank
    const int logC = 31 - __clz(C);
    if (epi) {
      const uint32_t my_slot = base + (uint32_t)(rank * C * 128 + row) * 4;  // receive buffer [src rank][col][row]
      uint32_t dst[8];
#pragma unroll
      for (int d = 0; d < 8; ++d) dst[d] = d < S ? mapa(my_slot, cx + CL * d) : 0;
#pragma unroll 1
      for (int c0 = half * kColsPerHalf; c0 < (half + 1) * kColsPerHalf; c0 += 16) {
        if (m0 + c0 >= g.M) break;
        uint32_t v[16];
        tmem_ld16(taddr + c0, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = c0 + j;
          if (m0 + col < g.M) {
            const int d = col >> logC;
            const int w = col & (C - 1);
            uint32_t a = dst[0];
#pragma unroll
            for (int dd = 1; dd < 8; ++dd)
              if (dd == d) a = dst[dd];
            st_cluster_u32(a + (uint32_t)(w * 128) * 4, v[j]);
          }
        }
      }
    }
    __syncwarp();
    cluster_sync();  // all partials have landed
    if (epi) {
      const float* recv = reinterpret_cast<const float*>(smem_raw + (base - smem_u32(smem_raw)));
      for (int w = half; w < C; w += kHalves) {
        const int m = m0 + rank * C + w;
        if (m >= g.M) break;
        float t = 0.f;
        for (int src = 0; src < S; ++src) t += recv[(src * C + w) * 128 + row];
        if (n_ok) out[(int64_t)m * g.N + n] = DT<T>::from_f(has_bias ? t + bias : t);
      }
    }
  } else if (epi) {
#pragma unroll 1
    for (int c0 = half * kColsPerHalf; c0 < (half + 1) * kColsPerHalf; c0 += 16) {
      if (m0 + c0 >= g.M) break;  // warp-uniform
      uint32_t v[16];
      tmem_ld16(taddr + c0, v);
      tmem_ld_wait();
      if (n_ok) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int m = m0 + c0 + j;
          if (m < g.M) {
            const float t = __uint_as_float(v[j]);
            out[(int64_t)m * g.N + n] = DT<T>::from_f(has_bias ? t + bias : t);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (CL > 1) cluster_sync();  // no CTA exits while a sharer may still signal its barriers
  if (warp == 2) tmem_dealloc<kTmemCols>(tmem_base);
}

template <typename T, int BN, int STAGES, int RAW>
static int launch_awq(const CUtensorMap& tw, const CUtensorMap& tx, const AwqArgs& g, dim3 grid, cudaStream_t st) {
  using L = AwqSmem<BN, STAGES, RAW>;
  auto kern = awq_gemm_kernel<T, BN, STAGES, RAW>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(kAwqThreads);
  cfg.dynamicSmemBytes = L::kBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = g.cl;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = g.splits;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_OK(cudaLaunchKernelEx(&cfg, kern, tw, tx, g));
  return B200_OK;
}

template <typename T>
static int run_awq_gemm(AwqArgs g, const void* x, const void* qweight, cudaStream_t st) {
  const int64_t M = g.M, N = g.N, K = g.K;
  g.num_kb = (int)ceil_div(K, kAwqBK);
  const int bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
  const int tiles = (int)ceil_div(N, 128);
  const int mtiles = (int)ceil_div(M, bn);
  int s = 1;
  if (mtiles == 1)  // weight streaming: split K over a cluster until every SM has a CTA
    while (s < 8 && tiles * s < sm_count() && g.num_kb / (s * 2) >= 4 && bn / (s * 2) >= 1) s *= 2;
  g.splits = s;
  g.kb_per_split = (int)ceil_div(g.num_kb, g.splits);
  while (g.splits > 1 && (int64_t)(g.splits - 1) * g.kb_per_split >= g.num_kb) {
    g.splits /= 2;
    g.kb_per_split = (int)ceil_div(g.num_kb, g.splits);
  }
  // token-tile sharing: the tile is BN * 128 B per K block against 4 KB of packed weights, so at BN >= 64 the L2 ->
  // SM traffic is dominated by tokens; cl CTAs along n fetch 1/cl of it each and multicast (cl * splits <= 8)
  // Measured on B200 (profiles/awq_r01.md): sharing does not pay at these sizes — the kernel is bound by the
  // packed-word stream, not by L2 -> SM token traffic — so it stays off unless B200_AWQ_CLUSTER_N asks for it.
  static const int cl_env = [] { const char* e = getenv("B200_AWQ_CLUSTER_N"); return e ? atoi(e) : 1; }();
  g.cl = (cl_env == 2 || cl_env == 4) && bn >= 64 ? cl_env : 1;
  while (g.cl > 1 && (g.cl * g.splits > 8 || tiles < g.cl)) g.cl /= 2;
  const int tiles_padded = (int)ceil_div(tiles, g.cl) * g.cl;  // surplus CTAs only feed the multicast
  CUtensorMap tw, tx;
  int rc;
  // packed weights [K rows, N/2 bytes]: box 64 rows x 64 bytes (16 words = 128 channels), 64-byte swizzle
  if ((rc = cached_tmap(&tw, qweight, (uint64_t)K, (uint64_t)(N / 2), kAwqBK, 64, 64)) != B200_OK) return rc;
  if ((rc = cached_tmap(&tx, x, (uint64_t)M, (uint64_t)(K * 2), (uint32_t)(bn / g.cl))) != B200_OK) return rc;
  dim3 grid(tiles_padded, mtiles, g.splits);
  switch (bn) {
    // operand stages >= dequant groups + 1 (one being consumed by the MMA); the raw ring holds the bytes in flight
    // BN <= 64: two CTAs per SM (<= 113 KB each) so one CTA's prologue / epilogue hides under the other's stream
    case 16: return launch_awq<T, 16, 3, 8>(tw, tx, g, grid, st);
    case 32: return launch_awq<T, 32, 3, 8>(tw, tx, g, grid, st);
    case 64: return launch_awq<T, 64, 3, 8>(tw, tx, g, grid, st);
    case 128: return launch_awq<T, 128, 4, 16>(tw, tx, g, grid, st);
    default: return launch_awq<T, 256, 4, 4>(tw, tx, g, grid, st);
  }
}

}  // namespace b200

using namespace b200;

B200_API int b200_awq_dequantize(void* w_out, const int32_t* qweight, const void* scales, const int32_t* qzeros,
                                 int64_t K, int64_t N, int group_size, int dtype, void* stream) {
  B200_REQUIRE(K >= 0 && N >= 0 && N % 8 == 0, "awq_dequantize: N %lld must be a multiple of 8", (long long)N);
  if (K == 0 || N == 0) return B200_OK;
  B200_REQUIRE(w_out && qweight && scales && qzeros, "awq_dequantize: null pointer");
  B200_REQUIRE(group_size > 0 && K % group_size == 0, "awq_dequantize: K %lld is not a multiple of group_size %d",
               (long long)K, group_size);
  B200_REQUIRE(((uintptr_t)w_out | (uintptr_t)scales) % 16 == 0, "awq_dequantize: out / scales must be 16-byte aligned");
  const int64_t words = K * (N / 8);
  const unsigned blocks = (unsigned)ceil_div(words, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_F16)
    awq_dequant_kernel<__half><<<blocks, 256, 0, st>>>((__half*)w_out, (const uint32_t*)qweight, (const __half*)scales,
                                                       (const uint32_t*)qzeros, K, N / 8, group_size);
  else if (dtype == B200_BF16)
    awq_dequant_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((__nv_bfloat16*)w_out, (const uint32_t*)qweight,
                                                              (const __nv_bfloat16*)scales, (const uint32_t*)qzeros, K,
                                                              N / 8, group_size);
  else
    return fail(B200_ERR_UNSUPPORTED, "awq_dequantize: dtype %d not supported (f16/bf16)", dtype);
  B200_LAUNCH_OK();
  return B200_OK;
}

B200_API int b200_awq_gemm(void* out, const void* x, const int32_t* qweight, const void* scales, const int32_t* qzeros,
                           const void* bias, int64_t M, int64_t N, int64_t K, int group_size, int dtype, void* stream) {
  B200_REQUIRE(M >= 0 && N > 0 && K > 0, "awq_gemm: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(out && x && qweight && scales && qzeros, "awq_gemm: null pointer");
  // same shape rules as AWQLinearMethod.create_weights (awq.py:339-352): K % group == 0, N % pack_factor == 0;
  // the TMA row pitch additionally needs N % 32 == 0 (16-byte rows of packed words) and K % 8 == 0
  B200_REQUIRE(group_size > 0 && K % group_size == 0, "awq_gemm: K %lld is not a multiple of group_size %d", (long long)K,
               group_size);
  if (group_size < 32 || (group_size < kAwqBK ? kAwqBK % group_size != 0 : group_size % kAwqBK != 0))
    return fail(B200_ERR_UNSUPPORTED, "awq_gemm: group_size %d not supported (32, or a multiple of 64)", group_size);
  B200_REQUIRE(N % 32 == 0, "awq_gemm: N %lld must be a multiple of 32", (long long)N);
  B200_REQUIRE(K % 8 == 0, "awq_gemm: K %lld must be a multiple of 8", (long long)K);
  B200_REQUIRE(((uintptr_t)out % 2) == 0 && ((uintptr_t)scales % 16) == 0, "awq_gemm: pointer alignment");
  B200_REQUIRE(M < (1ll << 30) && N < (1ll << 30) && K < (1ll << 30), "awq_gemm: sizes exceed 32-bit tile math");
  AwqArgs g = {};
  g.out = out; g.scales = scales; g.qzeros = (const uint32_t*)qzeros; g.bias = bias;
  g.M = (int)M; g.N = (int)N; g.K = (int)K; g.group_size = group_size;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_F16) return run_awq_gemm<__half>(g, x, qweight, st);
  if (dtype == B200_BF16) return run_awq_gemm<__nv_bfloat16>(g, x, qweight, st);
  return fail(B200_ERR_UNSUPPORTED, "awq_gemm: dtype %d not supported (f16/bf16)", dtype);
}
