// W8A8-int8 scaled GEMM on tcgen05 (kind::i8, int32 accumulators in TMEM), TMA-fed.
//
// Replaces int8_scaled_mm (reference sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-746, which has NO
// sm_100 path: TORCH_CHECK_NOT_IMPLEMENTED at :742-743).  out[m,n] = cast((sa[m] * acc[m,n]) * sb[n]
// + bias[n]) with acc[m,n] = sum_k a[m,k] * w[n,k] exact in int32 — the order of the fp32 epilogue
// follows the oracle restatement (test/srt/cpu/utils.py:91-115).
//
// One kernel, two operand assignments (both operands are K-major byte matrices, so either can sit on
// the 128-row MMA "M" side):
//   * SWAP (decode, M <= 256 tokens): weights are the MMA-M operand (128 output channels per CTA),
//     tokens the MMA-N operand (BN = 16..256).  Weight-streaming: every weight byte is pulled from
//     HBM exactly once by TMA into a 4-stage 128B-swizzled ring; split-K over gridDim.z spreads a
//     GEMM over >= ~1.5 x SM-count CTAs so HBM stays saturated.  Split partials are summed with
//     int32 `red.global.add` into a self-cleaning workspace — integer adds are associative, so the
//     result is exact and run-to-run deterministic; the last-arriving CTA of a tile applies the scales.
//   * NORMAL (prefill): activations on MMA-M (128 tokens), weights on MMA-N (BN = 128/256 channels).
// Warp roles: warp 0 = TMA producer (one lane), warp 1 = MMA issuer (one lane), warp 2 = TMEM
// allocator, warps 4-7 = epilogue (TMEM lane quarter = warp % 4).
#include <unordered_map>

#include "sm100.cuh"

namespace b200 {

using namespace sm100;

struct GemmI8Args {
  void* out;            // [M, N] bf16/f16
  const float* scale_m;  // [M] per-token
  const float* scale_n;  // [N] per-channel
  const void* bias;      // [N] out dtype or null
  int32_t* acc_out;      // [M, N] raw accumulators or null
  int M, N;
  int num_kb;        // K blocks of 128 bytes (ceil)
  int kb_per_split;  // K blocks per gridDim.z slice
  int splits;
  int32_t* ws_acc;      // [M, N] int32, zero between calls (split-K only)
  int32_t* ws_counter;  // [tiles] int32, zero between calls
  int out_is_bf16;
};

constexpr int kGemmThreads = 256;
constexpr int kBK = 128;  // bytes of K per pipeline stage = one 128B swizzle row

template <int BN, int STAGES>
struct GemmSmemLayout {
  static constexpr int kA = 128 * kBK;
  static constexpr int kB = BN * kBK;
  static constexpr int kStage = kA + kB;
  static constexpr int kBarOff = STAGES * kStage;
  static constexpr int kBytes = kBarOff + (2 * STAGES + 2) * 8 + 1024;  // + alignment slack
};

__device__ __forceinline__ float epilogue_value(int32_t acc, float sm, float sn, bool has_bias, float bias) {
  float v = __fmul_rn(__fmul_rn(sm, (float)acc), sn);
  return has_bias ? __fadd_rn(v, bias) : v;
}

template <int BN, int STAGES, bool SWAP>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_i8_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const GemmI8Args g) {
  using L = GemmSmemLayout<BN, STAGES>;
  constexpr int kTmemCols = BN < 32 ? 32 : BN;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  __shared__ int s_ticket;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  auto full_bar = [&](int s) { return bar + s * 8; };
  auto empty_bar = [&](int s) { return bar + (STAGES + s) * 8; };
  const uint32_t tmem_full_bar = bar + 2 * STAGES * 8;

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
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<kTmemCols>(smem_u32(&s_tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;

  const int a_row0 = blockIdx.x * 128;
  const int b_row0 = blockIdx.y * BN;
  const int kb0 = blockIdx.z * g.kb_per_split;
  const int nkb = min(g.kb_per_split, g.num_kb - kb0);

  if (warp == 0) {
    if (lane == 0) {
      // weights are read once (evict-first); activations are re-read by every CTA (evict-last)
      const uint64_t pol_a = SWAP ? kEvictFirst : kEvictLast;
      const uint64_t pol_b = SWAP ? kEvictLast : kEvictFirst;
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(empty_bar(s), ph ^ 1);
        mbar_expect_tx(full_bar(s), L::kStage);
        const uint32_t sa = base + s * L::kStage;
        tma_load_2d(sa, &tmap_a, full_bar(s), (kb0 + i) * kBK, a_row0, pol_a);
        tma_load_2d(sa + L::kA, &tmap_b, full_bar(s), (kb0 + i) * kBK, b_row0, pol_b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc(/*c=s32*/ 2, /*a=int8*/ 1, /*b=int8*/ 1, 128, BN);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % STAGES;
        const uint32_t ph = (i / STAGES) & 1;
        mbar_wait(full_bar(s), ph);
        tc_fence_after();
        const uint32_t sa = base + s * L::kStage;
        const uint64_t da = smem_desc_k_sw128(sa);
        const uint64_t db = smem_desc_k_sw128(sa + L::kA);
#pragma unroll
        for (int k = 0; k < kBK / 32; ++k)  // UMMA K = 32 bytes; +2 in the (addr >> 4) field per step
          umma<kKindI8>(tmem_base, da + 2 * k, db + 2 * k, idesc, (i | k) != 0);
        tc_commit(empty_bar(s));  // frees the smem stage when these MMAs retire
      }
      tc_commit(tmem_full_bar);
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
    const bool has_bias = g.bias != nullptr;

    if constexpr (SWAP) {
      const int n = a_row0 + row;  // output channel owned by this thread
      const bool n_ok = n < g.N;
      const float sn = n_ok ? g.scale_n[n] : 0.f;
      float bias = 0.f;
      if (has_bias && n_ok)
        bias = g.out_is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g.bias)[n])
                             : __half2float(reinterpret_cast<const __half*>(g.bias)[n]);
      const bool split = g.splits > 1;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        if (b_row0 + c0 >= g.M) break;  // warp-uniform
        uint32_t r[16];
        tmem_ld16(taddr + c0, r);
        tmem_ld_wait();
        if (n_ok) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int m = b_row0 + c0 + j;
            if (m < g.M) {
              const int64_t idx = (int64_t)m * g.N + n;
              const int32_t acc = (int32_t)r[j];
              if (split) {
                asm volatile("red.global.add.s32 [%0], %1;\n" ::"l"(g.ws_acc + idx), "r"(acc) : "memory");
              } else {
                const float v = epilogue_value(acc, g.scale_m[m], sn, has_bias, bias);
                if (g.out_is_bf16) reinterpret_cast<__nv_bfloat16*>(g.out)[idx] = __float2bfloat16_rn(v);
                else reinterpret_cast<__half*>(g.out)[idx] = __float2half_rn(v);
                if (g.acc_out) g.acc_out[idx] = acc;
              }
            }
          }
        }
      }
      if (split) {
        // publish partials, take a ticket; the last CTA of this tile finishes the epilogue
        __threadfence();
        asm volatile("bar.sync 1, 128;\n" ::: "memory");
        const int tile = blockIdx.y * gridDim.x + blockIdx.x;
        if (warp == 4 && lane == 0) {
          __threadfence();
          s_ticket = atomicAdd(g.ws_counter + tile, 1);
        }
        asm volatile("bar.sync 1, 128;\n" ::: "memory");
        if (s_ticket == g.splits - 1) {
          __threadfence();
          if (n_ok) {
            const int m_end = min(g.M, b_row0 + BN);
            for (int m = b_row0; m < m_end; ++m) {
              const int64_t idx = (int64_t)m * g.N + n;
              int32_t acc;
              asm volatile("ld.global.cg.s32 %0, [%1];\n" : "=r"(acc) : "l"(g.ws_acc + idx) : "memory");
              g.ws_acc[idx] = 0;  // leave the workspace clean for the next call
              const float v = epilogue_value(acc, g.scale_m[m], sn, has_bias, bias);
              if (g.out_is_bf16) reinterpret_cast<__nv_bfloat16*>(g.out)[idx] = __float2bfloat16_rn(v);
              else reinterpret_cast<__half*>(g.out)[idx] = __float2half_rn(v);
              if (g.acc_out) g.acc_out[idx] = acc;
            }
          }
          if (warp == 4 && lane == 0) g.ws_counter[tile] = 0;
        }
      }
    } else {
      const int m = a_row0 + row;  // token row owned by this thread
      const bool m_ok = m < g.M;
      const float smv = m_ok ? g.scale_m[m] : 0.f;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        const int n0 = b_row0 + c0;
        if (n0 >= g.N) break;  // warp-uniform
        uint32_t r[16];
        tmem_ld16(taddr + c0, r);
        tmem_ld_wait();
        if (m_ok) {
          // N % 8 == 0 is required, so a 16-column group is either full or exactly half full
          const int valid = min(16, g.N - n0);
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int n = n0 + j;
            float bias = 0.f;
            float sn = 0.f;
            if (j < valid) {
              sn = g.scale_n[n];
              if (has_bias)
                bias = g.out_is_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(g.bias)[n])
                                     : __half2float(reinterpret_cast<const __half*>(g.bias)[n]);
            }
            v[j] = epilogue_value((int32_t)r[j], smv, sn, has_bias, bias);
          }
          uint32_t p[8];
#pragma unroll
          for (int j = 0; j < 8; ++j)
            p[j] = g.out_is_bf16 ? pack2<__nv_bfloat16>(v[2 * j], v[2 * j + 1]) : pack2<__half>(v[2 * j], v[2 * j + 1]);
          uint8_t* dst = reinterpret_cast<uint8_t*>(g.out) + ((int64_t)m * g.N + n0) * 2;
          *reinterpret_cast<uint4*>(dst) = make_uint4(p[0], p[1], p[2], p[3]);
          if (valid > 8) *reinterpret_cast<uint4*>(dst + 16) = make_uint4(p[4], p[5], p[6], p[7]);
          if (g.acc_out) {
            int32_t* ad = g.acc_out + (int64_t)m * g.N + n0;
#pragma unroll
            for (int j = 0; j < 16; j += 4)
              if (j < valid) *reinterpret_cast<uint4*>(ad + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<kTmemCols>(tmem_base);
}

// ---- host side -----------------------------------------------------------------------------------
struct TmapKey {
  const void* p;
  uint64_t rows, kbytes;
  uint32_t box;
  bool operator==(const TmapKey& o) const { return p == o.p && rows == o.rows && kbytes == o.kbytes && box == o.box; }
};
struct TmapKeyHash {
  size_t operator()(const TmapKey& k) const {
    return std::hash<const void*>()(k.p) ^ (k.rows * 0x9E3779B97F4A7C15ull) ^ (k.kbytes << 17) ^ k.box;
  }
};

int cached_tmap(CUtensorMap* out, const void* base, uint64_t rows, uint64_t kbytes, uint32_t box_rows) {
  static thread_local std::unordered_map<TmapKey, CUtensorMap, TmapKeyHash> cache;
  TmapKey key{base, rows, kbytes, box_rows};
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return B200_OK;
  }
  int st = make_tmap_2d_bytes(out, base, rows, kbytes, kbytes, box_rows);
  if (st != B200_OK) return st;
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, *out);
  return B200_OK;
}

constexpr size_t kCounterBytes = 64 * 1024;  // 16384 tile tickets

static int pick_splits(int tiles, int num_kb, int sms) {
  // enough CTAs to keep every SM's TMA queue busy (2 CTAs/SM resident), >= 4 K-blocks per slice
  int want = (2 * sms + tiles - 1) / tiles;
  int max_by_k = num_kb / 4;
  if (max_by_k < 1) max_by_k = 1;
  int s = want < max_by_k ? want : max_by_k;
  if (s < 1) s = 1;
  if (s > 32) s = 32;
  return s;
}

template <int BN, int STAGES, bool SWAP>
static int launch_i8(const CUtensorMap& ta, const CUtensorMap& tb, const GemmI8Args& g, dim3 grid, cudaStream_t st) {
  using L = GemmSmemLayout<BN, STAGES>;
  auto kern = gemm_i8_kernel<BN, STAGES, SWAP>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  kern<<<grid, kGemmThreads, L::kBytes, st>>>(ta, tb, g);
  B200_LAUNCH_OK();
  return B200_OK;
}

}  // namespace b200

using namespace b200;

B200_API int b200_int8_scaled_mm_workspace_bytes(int64_t M, int64_t N, int64_t K, size_t* bytes) {
  B200_REQUIRE(bytes != nullptr, "int8_scaled_mm_workspace_bytes: null out pointer");
  *bytes = kCounterBytes + (M <= 256 ? (size_t)M * N * 4 : 0) + 256;
  return B200_OK;
}

B200_API int b200_int8_scaled_mm(void* out, const int8_t* a, const int8_t* b, const float* scale_a, const float* scale_b,
                                 const void* bias, int64_t M, int64_t N, int64_t K, int out_dtype, int32_t* acc_out,
                                 void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(M >= 0 && N > 0 && K > 0, "int8_scaled_mm: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(out && a && b && scale_a && scale_b, "int8_scaled_mm: null pointer");
  // same shape rules as the reference (int8_gemm_kernel.cu:677-679)
  B200_REQUIRE(K % 16 == 0, "int8_scaled_mm: K %lld must be a multiple of 16", (long long)K);
  B200_REQUIRE(N % 8 == 0, "int8_scaled_mm: N %lld must be a multiple of 8", (long long)N);
  B200_REQUIRE(((uintptr_t)out % 16) == 0 && ((uintptr_t)acc_out % 16) == 0, "int8_scaled_mm: output alignment");
  B200_REQUIRE(M < (1ll << 31) / 2 && N < (1 << 30), "int8_scaled_mm: sizes exceed 32-bit tile math");
  if (out_dtype != B200_BF16 && out_dtype != B200_F16)
    return fail(B200_ERR_UNSUPPORTED, "int8_scaled_mm: out dtype %d not supported (bf16/f16)", out_dtype);
  cudaStream_t st = (cudaStream_t)stream;
  GemmI8Args g;
  g.out = out; g.scale_m = scale_a; g.scale_n = scale_b; g.bias = bias; g.acc_out = acc_out;
  g.M = (int)M; g.N = (int)N;
  g.num_kb = (int)ceil_div(K, kBK);
  g.out_is_bf16 = out_dtype == B200_BF16;
  g.ws_counter = reinterpret_cast<int32_t*>(workspace);
  g.ws_acc = workspace ? reinterpret_cast<int32_t*>(reinterpret_cast<uint8_t*>(workspace) + kCounterBytes) : nullptr;
  CUtensorMap ta, tb;
  int rc;
  if (M <= 256) {
    // weight-streaming orientation: A = weights [N, K], B = tokens [M, K]
    const int bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
    const int tiles = (int)ceil_div(N, 128);
    g.splits = pick_splits(tiles, g.num_kb, sm_count());
    g.kb_per_split = (int)ceil_div(g.num_kb, g.splits);
    g.splits = (int)ceil_div(g.num_kb, g.kb_per_split);  // no empty slices
    if (g.splits > 1) {
      size_t need = 0;
      b200_int8_scaled_mm_workspace_bytes(M, N, K, &need);
      if (workspace == nullptr || workspace_bytes < need)
        return fail(B200_ERR_WORKSPACE, "int8_scaled_mm: workspace %zu B < required %zu B", workspace_bytes, need);
      B200_REQUIRE((size_t)tiles * 4 <= kCounterBytes, "int8_scaled_mm: too many tiles for the ticket table");
    }
    if ((rc = cached_tmap(&ta, b, (uint64_t)N, (uint64_t)K, 128)) != B200_OK) return rc;
    if ((rc = cached_tmap(&tb, a, (uint64_t)M, (uint64_t)K, (uint32_t)bn)) != B200_OK) return rc;
    dim3 grid(tiles, 1, g.splits);
    switch (bn) {
      case 16: return launch_i8<16, 4, true>(ta, tb, g, grid, st);
      case 32: return launch_i8<32, 4, true>(ta, tb, g, grid, st);
      case 64: return launch_i8<64, 4, true>(ta, tb, g, grid, st);
      case 128: return launch_i8<128, 4, true>(ta, tb, g, grid, st);
      default: return launch_i8<256, 4, true>(ta, tb, g, grid, st);
    }
  }
  // prefill orientation: A = tokens [M, K], B = weights [N, K]
  g.splits = 1;
  g.kb_per_split = g.num_kb;
  const int bn = N >= 256 ? 256 : 128;
  if ((rc = cached_tmap(&ta, a, (uint64_t)M, (uint64_t)K, 128)) != B200_OK) return rc;
  if ((rc = cached_tmap(&tb, b, (uint64_t)N, (uint64_t)K, (uint32_t)bn)) != B200_OK) return rc;
  dim3 grid((unsigned)ceil_div(M, 128), (unsigned)ceil_div(N, bn), 1);
  if (bn == 256) return launch_i8<256, 4, false>(ta, tb, g, grid, st);
  return launch_i8<128, 4, false>(ta, tb, g, grid, st);
}
