// Paged extend / prefill attention (prefix through the page table + causal new tokens).
//
// Replaces extend_attention_fwd (reference python/sglang/srt/layers/attention/triton_ops/
// extend_attention.py:302-440, kernel :42-299); semantics follow the oracle
// TorchNativeAttnBackend._run_sdpa_forward_extend (torch_native_backend.py:27-110): query i of
// request b (0-based among its extend tokens) attends KV positions 0 .. prefix_b + i of that
// request (all prefix_b + extend_b positions when causal = 0).
//
// Round-1 kernel: flash-attention-2 style on the legacy tensor path (mma.sync m16n8k16, fp32
// accumulate).  grid = (ceil(max_extend/64), Hq, batch); 4 warps x 16 query rows; K/V tiles of
// 64 tokens double-buffered with 16-byte cp.async gathers (one 256-byte row per 16 lanes),
// XOR-swizzled shared memory, ldmatrix operands, online softmax in the exp2 domain.  KV position
// p is fetched from the cache slot req_to_token[req, p]; when k_new/v_new are given, positions
// >= prefix come from those contiguous activations instead (save_kv_cache = False callers).
// The tcgen05/TMEM version of this tile loop is the next step for this row.
#include <cstdlib>

#include "extend_attn.cuh"

namespace b200 {

constexpr int kExtWarps = 4;
constexpr int kExtThreads = kExtWarps * 32;
constexpr int kExtBM = 64;
constexpr int kExtBN = 64;


// Load one [64 x D] K tile and V tile (KV positions kv0 .. kv0+63 of the request).
template <typename T, int D>
__device__ __forceinline__ void load_kv_tile(const ExtendArgs& a, const int32_t* __restrict__ r2t_row,
                                             int kv0, int kv_end, int prefix, int64_t new_base_tok,
                                             int kvh, uint32_t smem_k, uint32_t smem_v) {
  constexpr int kChunks = D / 8;
  constexpr int kRowsPerPass = kExtThreads / kChunks;
  const int c = threadIdx.x % kChunks;
  const int r0 = threadIdx.x / kChunks;
  const T* k_cache = reinterpret_cast<const T*>(a.k_cache);
  const T* v_cache = reinterpret_cast<const T*>(a.v_cache);
  const T* k_new = reinterpret_cast<const T*>(a.k_new);
  const T* v_new = reinterpret_cast<const T*>(a.v_new);
#pragma unroll
  for (int p = 0; p < kExtBN / kRowsPerPass; ++p) {
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
    const uint32_t so = (uint32_t)(r * (D * 2) + ((c ^ (r & 7)) << 4));
    cp_async16_pred(smem_k + so, ksrc, valid);
    cp_async16_pred(smem_v + so, vsrc, valid);
  }
}

template <typename T, int D>
__global__ void __launch_bounds__(kExtThreads) extend_attn_kernel(const ExtendArgs a) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  constexpr int kTileBytes = kExtBN * D * 2;
  const int mblk = blockIdx.x;
  const int head = blockIdx.y;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;
  const int t = lane & 3;

  const int q_start = a.qo_indptr[b];
  const int ext_len = a.qo_indptr[b + 1] - q_start;
  const int m0 = mblk * kExtBM;
  if (m0 >= ext_len) return;
  const int prefix = a.prefix_lens[b];
  const int kv_total = prefix + ext_len;
  const int kv_end = a.causal ? min(kv_total, prefix + m0 + kExtBM) : kv_total;
  const int group = a.num_q_heads / a.num_kv_heads;
  const int kvh = head / group;
  const int64_t req = load_index(a.req_pool_indices, b, a.index_is_int64);
  const int32_t* r2t_row = a.req_to_token + req * a.req_to_token_stride;

  const uint32_t sm_q = smem_u32(smem_raw);
  const uint32_t sm_k = sm_q + kExtBM * D * 2;
  const uint32_t sm_v = sm_k + 2 * kTileBytes;

  // ---- Q tile -> smem (zero-fill rows past the request's extend length) ----------------------
  {
    constexpr int kChunks = D / 8;
    constexpr int kRowsPerPass = kExtThreads / kChunks;
    const int c = threadIdx.x % kChunks;
    const int r0 = threadIdx.x / kChunks;
    const T* qb = reinterpret_cast<const T*>(a.q) + (int64_t)head * D;
#pragma unroll
    for (int p = 0; p < kExtBM / kRowsPerPass; ++p) {
      const int r = r0 + p * kRowsPerPass;
      const bool valid = m0 + r < ext_len;
      const T* src = valid ? qb + (int64_t)(q_start + m0 + r) * a.q_token_stride + c * 8 : qb;
      cp_async16_pred(sm_q + r * (D * 2) + ((c ^ (r & 7)) << 4), src, valid);
    }
  }
  const int n_tiles = (kv_end + kExtBN - 1) / kExtBN;
  load_kv_tile<T, D>(a, r2t_row, 0, kv_end, prefix, q_start, kvh, sm_k, sm_v);
  cp_async_commit();

  uint32_t qf[D / 16][4];
  float o_acc[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};
  const bool use_cap = a.logit_cap > 0.f;
  const float cap_in = use_cap ? a.sm_scale / a.logit_cap : 0.f;
  const float cap_out = a.logit_cap * kLog2eExt;
  const int lm_row = lane & 7;
  const int lm_mat = lane >> 3;
  // KV position limit per query row (inclusive) for the two rows this thread owns
  const int row_a = m0 + warp * 16 + g;
  const int lim0 = a.causal ? prefix + row_a : kv_total - 1;
  const int lim1 = a.causal ? prefix + row_a + 8 : kv_total - 1;

  for (int it = 0; it < n_tiles; ++it) {
    const int buf = it & 1;
    if (it + 1 < n_tiles)
      load_kv_tile<T, D>(a, r2t_row, (it + 1) * kExtBN, kv_end, prefix, q_start, kvh,
                         sm_k + (buf ^ 1) * kTileBytes, sm_v + (buf ^ 1) * kTileBytes);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();

    if (it == 0) {
      // A fragments of this warp's 16 query rows
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
        // matrices: 0:(rows 0-7, k lo) 1:(rows 8-15, k lo) 2:(rows 0-7, k hi) 3:(rows 8-15, k hi)
        const int r = warp * 16 + (lm_mat & 1) * 8 + lm_row;
        const int c = ks * 2 + (lm_mat >> 1);
        ldmatrix_x4(qf[ks][0], qf[ks][1], qf[ks][2], qf[ks][3], sm_q + r * (D * 2) + ((c ^ (r & 7)) << 4));
      }
    }

    const uint32_t sk = sm_k + buf * kTileBytes;
    const uint32_t sv = sm_v + buf * kTileBytes;
    const int kv0 = it * kExtBN;

    float s[kExtBN / 8][4];
#pragma unroll
    for (int j = 0; j < kExtBN / 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
      for (int jp = 0; jp < kExtBN / 16; ++jp) {
        const int r = jp * 16 + (lm_mat >> 1) * 8 + lm_row;
        const int c = ks * 2 + (lm_mat & 1);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(b0, b1, b2, b3, sk + r * (D * 2) + ((c ^ (r & 7)) << 4));
        mma_16816<T>(s[2 * jp], qf[ks], b0, b1);
        mma_16816<T>(s[2 * jp + 1], qf[ks], b2, b3);
      }
    }

    // scale, cap, mask
#pragma unroll
    for (int j = 0; j < kExtBN / 8; ++j) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = s[j][e];
        if (use_cap) v = cap_out * tanh_approx(v * cap_in);
        else v *= a.scale_log2;
        const int pos = kv0 + j * 8 + 2 * t + (e & 1);
        const int lim = (e < 2) ? lim0 : lim1;
        s[j][e] = (pos <= lim && pos < kv_end) ? v : -INFINITY;
      }
    }

    float alpha[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < kExtBN / 8; ++j) mx = fmaxf(mx, fmaxf(s[j][2 * h], s[j][2 * h + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[h], mx);
      // rows past the extend length can be fully masked in tile 0 only if kv_end == 0 (impossible);
      // guard anyway so that -inf - -inf never produces NaN
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      alpha[h] = fast_exp2(m_run[h] - m_use);
      m_run[h] = m_new;
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < kExtBN / 8; ++j) {
        s[j][2 * h] = fast_exp2(s[j][2 * h] - m_use);
        s[j][2 * h + 1] = fast_exp2(s[j][2 * h + 1] - m_use);
        sum += s[j][2 * h] + s[j][2 * h + 1];
      }
      l_run[h] = l_run[h] * alpha[h] + sum;
    }
#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) {
      o_acc[nt][0] *= alpha[0];
      o_acc[nt][1] *= alpha[0];
      o_acc[nt][2] *= alpha[1];
      o_acc[nt][3] *= alpha[1];
    }

#pragma unroll
    for (int kt = 0; kt < kExtBN / 16; ++kt) {
      uint32_t pf[4];
      pf[0] = pack2<T>(s[2 * kt][0], s[2 * kt][1]);
      pf[1] = pack2<T>(s[2 * kt][2], s[2 * kt][3]);
      pf[2] = pack2<T>(s[2 * kt + 1][0], s[2 * kt + 1][1]);
      pf[3] = pack2<T>(s[2 * kt + 1][2], s[2 * kt + 1][3]);
#pragma unroll
      for (int np = 0; np < D / 16; ++np) {
        const int r = kt * 16 + (lm_mat & 1) * 8 + lm_row;
        const int c = np * 2 + (lm_mat >> 1);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4_trans(b0, b1, b2, b3, sv + r * (D * 2) + ((c ^ (r & 7)) << 4));
        mma_16816<T>(o_acc[2 * np], pf, b0, b1);
        mma_16816<T>(o_acc[2 * np + 1], pf, b2, b3);
      }
    }
    __syncthreads();  // tile `buf` may be overwritten by the next iteration's prefetch
  }
  cp_async_wait<0>();

  // ---- epilogue ---------------------------------------------------------------------------------
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
  }
  const float inv0 = l_run[0] > 0.f ? 1.f / l_run[0] : 0.f;
  const float inv1 = l_run[1] > 0.f ? 1.f / l_run[1] : 0.f;
  T* ob = reinterpret_cast<T*>(a.o) + (int64_t)head * D;
  const int r_a = m0 + warp * 16 + g;
  const int r_b = r_a + 8;
#pragma unroll
  for (int nt = 0; nt < D / 8; ++nt) {
    const int col = nt * 8 + 2 * t;
    if (r_a < ext_len)
      *reinterpret_cast<uint32_t*>(ob + (int64_t)(q_start + r_a) * a.o_token_stride + col) =
          pack2<T>(o_acc[nt][0] * inv0, o_acc[nt][1] * inv0);
    if (r_b < ext_len)
      *reinterpret_cast<uint32_t*>(ob + (int64_t)(q_start + r_b) * a.o_token_stride + col) =
          pack2<T>(o_acc[nt][2] * inv1, o_acc[nt][3] * inv1);
  }
}

template <typename T, int D>
static int launch_extend(const ExtendArgs& a, int max_extend_len, cudaStream_t stream) {
  const int smem = kExtBM * D * 2 + 4 * kExtBN * D * 2;
  auto kern = extend_attn_kernel<T, D>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  dim3 grid((max_extend_len + kExtBM - 1) / kExtBM, a.num_q_heads, a.batch);
  kern<<<grid, kExtThreads, smem, stream>>>(a);
  B200_LAUNCH_OK();
  return B200_OK;
}

}  // namespace b200

B200_API int b200_extend_attn(
    void* o, const void* q, const void* k_new, const void* v_new, const void* k_cache, const void* v_cache,
    const int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices, int index_is_int64,
    const int32_t* qo_indptr, const int32_t* prefix_lens, int batch, int max_extend_len, int num_q_heads,
    int num_kv_heads, int head_dim, int64_t q_token_stride, int64_t o_token_stride, int64_t knew_token_stride,
    int64_t vnew_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap, int causal, int dtype,
    void* stream, int64_t num_q_tokens, int64_t num_kv_slots) {
  using namespace b200;
  B200_REQUIRE(batch >= 0 && max_extend_len >= 0, "extend_attn: negative sizes");
  if (batch == 0 || max_extend_len == 0) return B200_OK;
  B200_REQUIRE(o && q && k_cache && v_cache && req_to_token && req_pool_indices && qo_indptr && prefix_lens,
               "extend_attn: null pointer");
  B200_REQUIRE((k_new == nullptr) == (v_new == nullptr), "extend_attn: k_new and v_new must both be given or both NULL");
  B200_REQUIRE(num_kv_heads > 0 && num_q_heads % num_kv_heads == 0, "extend_attn: bad head counts");
  B200_REQUIRE(batch <= 65535 && num_q_heads <= 65535, "extend_attn: grid limits");
  B200_REQUIRE(q_token_stride % 8 == 0 && kv_token_stride % 8 == 0 && o_token_stride % 2 == 0 &&
                   knew_token_stride % 8 == 0 && vnew_token_stride % 8 == 0,
               "extend_attn: strides must keep 16-byte row alignment");
  B200_REQUIRE(((uintptr_t)q | (uintptr_t)k_cache | (uintptr_t)v_cache | (uintptr_t)k_new | (uintptr_t)v_new) % 16 == 0,
               "extend_attn: pointers must be 16-byte aligned");
  ExtendArgs a;
  a.o = o; a.q = q; a.k_new = k_new; a.v_new = v_new; a.k_cache = k_cache; a.v_cache = v_cache;
  a.req_to_token = req_to_token; a.req_to_token_stride = req_to_token_stride;
  a.req_pool_indices = req_pool_indices; a.index_is_int64 = index_is_int64;
  a.qo_indptr = qo_indptr; a.prefix_lens = prefix_lens;
  a.batch = batch; a.num_q_heads = num_q_heads; a.num_kv_heads = num_kv_heads;
  a.q_token_stride = q_token_stride; a.o_token_stride = o_token_stride;
  a.knew_token_stride = knew_token_stride; a.vnew_token_stride = vnew_token_stride;
  a.kv_token_stride = kv_token_stride;
  a.sm_scale = sm_scale; a.scale_log2 = sm_scale * kLog2eExt; a.logit_cap = logit_cap; a.causal = causal;
  cudaStream_t st = (cudaStream_t)stream;
  // tcgen05 / TMEM tiles (extend_attn_umma.cu) unless the whole batch is a handful of query rows, where 128-row
  // MMA tiles would be mostly padding; B200_EXTEND_LEGACY=1 forces the mma.sync kernel (A/B measurements)
  static const bool legacy = getenv("B200_EXTEND_LEGACY") != nullptr && getenv("B200_EXTEND_LEGACY")[0] == '1';
  if (!legacy && max_extend_len > 16 && (head_dim == 128 || head_dim == 64) && (dtype == B200_BF16 || dtype == B200_F16)) {
    // two query tiles per CTA + TMA (extend_attn_fa.cu) from 129 query rows per request up, when all K/V is read
    // through the page table and the caller gave the tensor extents TMA needs; B200_EXTEND_FA=0 keeps the
    // one-tile kernel (A/B measurements)
    static const bool fa = getenv("B200_EXTEND_FA") == nullptr || getenv("B200_EXTEND_FA")[0] != '0';
    if (fa && max_extend_len > 128 && k_new == nullptr && num_q_tokens > 0 && num_kv_slots > 0 &&
        q_token_stride % 8 == 0 && kv_token_stride % 8 == 0) {
      const int rc = launch_extend_fa(a, max_extend_len, head_dim, dtype, num_q_tokens, num_kv_slots, st);
      if (rc != B200_ERR_UNSUPPORTED) return rc;
    }
    return launch_extend_umma(a, max_extend_len, head_dim, dtype, st);
  }
  if (dtype == B200_BF16) {
    if (head_dim == 128) return launch_extend<__nv_bfloat16, 128>(a, max_extend_len, st);
    if (head_dim == 64) return launch_extend<__nv_bfloat16, 64>(a, max_extend_len, st);
  } else if (dtype == B200_F16) {
    if (head_dim == 128) return launch_extend<__half, 128>(a, max_extend_len, st);
    if (head_dim == 64) return launch_extend<__half, 64>(a, max_extend_len, st);
  } else {
    return fail(B200_ERR_UNSUPPORTED, "extend_attn: dtype %d not supported (bf16/f16 only)", dtype);
  }
  return fail(B200_ERR_UNSUPPORTED, "extend_attn: head_dim %d not supported (64, 128)", head_dim);
}
