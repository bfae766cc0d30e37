// Paged split-KV GQA flash-decode + LSE merge for sm_100a.
//
// Replaces decode_attention_fwd (reference python/sglang/srt/layers/attention/triton_ops/
// decode_attention.py:677-728; grouped stage 1 :241-401, stage 2 :492-552).  Semantics follow
// the oracle TorchNativeAttnBackend._run_sdpa_forward_decode (torch_native_backend.py:112-180):
//   o[b,h] = softmax(sm_scale * q[b,h] . K[idx_b, h/g]^T) . V[idx_b, h/g],
//   idx_b  = req_to_token[req_pool_indices[b], :seq_lens[b]],  no causal mask.
//
// Design (HBM-bound: 4 FLOP per KV byte):
//  * grid = (max_splits, Hkv * head_blocks, batch).  One CTA owns one (request, KV head,
//    KV chunk) and ALL q heads of that KV head's GQA group (<= 16 per head block), so every
//    KV byte is read from HBM exactly once per step.
//  * 4 warps per CTA; each warp streams its own 16-token blocks (block j of the chunk goes
//    to warp j % 4) through a private STAGES-deep cp.async ring: 16-byte vectorised copies,
//    a warp instruction covers 512 contiguous bytes of KV rows, XOR-swizzled in shared
//    memory so ldmatrix is bank-conflict free.  No CTA-wide barrier in the main loop.
//  * math on the legacy tensor path (mma.sync m16n8k16, fp32 accumulate): S = Q K^T with the
//    group's q heads on the M axis, online softmax in the exp2 domain with warp shuffles,
//    O += P V.  At 4 FLOP/B this is <10% of the tensor pipe; the kernel is judged on GB/s.
//  * the 4 warps' (m, l, O) are merged through shared memory; with one chunk per request the
//    CTA writes the final output, else fp32 partial (o, lse) go to the workspace and
//    decode_merge_kernel combines the chunks (log-sum-exp merge).
#include <mutex>

#include "common.cuh"
#include "extend_attn.cuh"
#include <type_traits>

namespace b200 {

constexpr int kDecWarps = 4;
constexpr int kDecThreads = kDecWarps * 32;
constexpr int kBlkTok = 16;  // tokens per warp block (one mma k-step of P.V, two n-tiles of Q.K^T)
constexpr float kLog2e = 1.4426950408889634f;

struct DecodeArgs {
  void* o;
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int32_t* req_to_token;
  int64_t req_to_token_stride;
  const void* req_pool_indices;
  const void* seq_lens;
  int index_is_int64;
  int batch, num_q_heads, num_kv_heads, group, head_blocks;
  int64_t q_token_stride, o_token_stride, kv_token_stride;
  float scale_log2;   // sm_scale * log2(e)
  float sm_scale;
  float logit_cap;
  float out_scale;    // fp8 KV only: v_scale, applied with the final 1/l
  int kv_chunk, max_splits;
  // shared-prefix (cascade) decode: tokens [0, kv_start) of every request were already reduced into partials
  // 0..n_seg-1 of the workspace by the tensor-core prefix kernel (extend_attn_umma.cu, segment mode); this kernel
  // covers tokens [kv_start, seq_len) and writes partials n_seg.. ; both are 0 for plain decode
  int kv_start, n_seg;
  int ws_splits;    // partial slots per (request, head) = n_seg + max_splits
  float* part_o;    // [batch, Hq, ws_splits, D]
  float* part_lse;  // [batch, Hq, ws_splits]   (log2 domain)
  int8_t* oq;       // optional fused output quant (b200_decode_attn_int8): [batch, Hq * D] int8 codes
  float* oq_scale;  // [batch] per-token scales
};

template <int D>
struct DecodeSmem {
  static constexpr int kChunks = D / 8;                       // 16-byte chunks per row
  static constexpr int kBlkBytes = kBlkTok * D * 2;           // one K (or V) block of a warp
  static constexpr int kStageBytes = 2 * kBlkBytes;           // K + V
  static constexpr int kOPad = D + 8;                         // fp32 row stride for the merge
  static constexpr int merge_bytes() { return kDecWarps * 16 * kOPad * 4 + 2 * kDecWarps * 16 * 4; }
  static constexpr int pipe_bytes(int stages) { return kDecWarps * stages * kStageBytes; }
};

// Tail of stage 1: the 4 warps' (m, l, O) sit in shared memory ([warps][16][kOPad] fp32 rows + m + l); merge them and
// write either the final output row (one chunk per request) or the fp32 partial (o, lse) of this chunk.
template <typename T, int D>
__device__ __forceinline__ void combine_warps_and_store(const DecodeArgs& a, const float* sm_o, const float* sm_m,
                                                        const float* sm_l, int n_heads, int head0, int b, int split,
                                                        int num_splits, float out_scale) {
  using S = DecodeSmem<D>;
  // thread -> (row = head within block, 8 threads per row, D/8 consecutive dims each)
  {
    const int row = threadIdx.x >> 3;
    const int seg = threadIdx.x & 7;
    if (row < n_heads) {
      float mw[kDecWarps], wgt[kDecWarps];
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < kDecWarps; ++w) {
        mw[w] = sm_m[w * 16 + row];
        M = fmaxf(M, mw[w]);
      }
      float L = 0.f;
#pragma unroll
      for (int w = 0; w < kDecWarps; ++w) {
        wgt[w] = (mw[w] == -INFINITY) ? 0.f : fast_exp2(mw[w] - M);
        L += sm_l[w * 16 + row] * wgt[w];
      }
      const float inv_l = out_scale / L;
      const int head = head0 + row;
      constexpr int kPer = D / 8;
      float acc[kPer];
#pragma unroll
      for (int d = 0; d < kPer; ++d) acc[d] = 0.f;
#pragma unroll
      for (int w = 0; w < kDecWarps; ++w) {
        const float* src = sm_o + (w * 16 + row) * S::kOPad + seg * kPer;
#pragma unroll
        for (int d = 0; d < kPer; ++d) acc[d] += src[d] * wgt[w];
      }
      if (num_splits == 1) {
        T* dst = reinterpret_cast<T*>(a.o) + (int64_t)b * a.o_token_stride + (int64_t)head * D + seg * kPer;
#pragma unroll
        for (int d = 0; d < kPer; d += 2)
          *reinterpret_cast<uint32_t*>(dst + d) = pack2<T>(acc[d] * inv_l, acc[d + 1] * inv_l);
      } else {
        const int64_t pidx = ((int64_t)b * a.num_q_heads + head) * a.ws_splits + a.n_seg + split;
        float* dst = a.part_o + pidx * D + seg * kPer;
#pragma unroll
        for (int d = 0; d < kPer; d += 4)
          *reinterpret_cast<float4*>(dst + d) =
              make_float4(acc[d] * inv_l, acc[d + 1] * inv_l, acc[d + 2] * inv_l, acc[d + 3] * inv_l);
        if (seg == 0) a.part_lse[pidx] = M + log2f(L);
      }
    }
  }
}

// Issue the cp.async copies of one 16-token block (K and V) owned by this warp.
template <typename T, int D>
__device__ __forceinline__ void issue_block(const T* __restrict__ k_cache, const T* __restrict__ v_cache,
                                            int64_t kv_token_stride, int kv_head_off,
                                            const int32_t* __restrict__ s_slots, int tok0, int n_tok,
                                            uint32_t smem_k, uint32_t smem_v, int lane) {
  constexpr int kChunks = D / 8;
  constexpr int kRowsPerIter = 32 / kChunks;
  const int c = lane % kChunks;
  const int r0 = lane / kChunks;
#pragma unroll
  for (int it = 0; it < kBlkTok / kRowsPerIter; ++it) {
    const int r = r0 + it * kRowsPerIter;
    const int tok = tok0 + r;
    const bool valid = tok < n_tok;
    const int slot = valid ? s_slots[tok] : 0;
    const int64_t off = (int64_t)slot * kv_token_stride + kv_head_off + c * 8;
    const uint32_t so = (uint32_t)(r * (D * 2) + ((c ^ (r & 7)) << 4));
    cp_async16_pred(smem_k + so, k_cache + off, valid);
    cp_async16_pred(smem_v + so, v_cache + off, valid);
  }
}

template <typename T, int D, int STAGES, bool UPPER>
__global__ void __launch_bounds__(kDecThreads, 2) decode_attn_kernel(const DecodeArgs a) {
  using S = DecodeSmem<D>;
  extern __shared__ __align__(128) uint8_t smem_raw[];

  const int split = blockIdx.x;
  const int kvh = blockIdx.y / a.head_blocks;
  const int hb = blockIdx.y - kvh * a.head_blocks;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;
  const int t = lane & 3;

  pdl_trigger();
  pdl_wait();
  const int seq_len = (int)load_index(a.seq_lens, b, a.index_is_int64) - a.kv_start;  // tokens this kernel covers
  const int start = split * a.kv_chunk;
  if (start >= seq_len) return;
  const int n_tok = min(seq_len - start, a.kv_chunk);
  const int num_splits = a.n_seg + (seq_len + a.kv_chunk - 1) / a.kv_chunk;

  // carve shared memory: [pipeline | slots]; the merge scratch aliases the pipeline.
  uint8_t* pipe = smem_raw;
  int32_t* s_slots = reinterpret_cast<int32_t*>(smem_raw + S::pipe_bytes(STAGES));

  // ---- 1. page-table slice of this chunk -> shared memory --------------------------------
  {
    const int64_t req = load_index(a.req_pool_indices, b, a.index_is_int64);
    const int32_t* row = a.req_to_token + req * a.req_to_token_stride + a.kv_start + start;
    for (int i = threadIdx.x; i < n_tok; i += kDecThreads) s_slots[i] = __ldg(row + i);
  }
  __syncthreads();

  const T* k_cache = reinterpret_cast<const T*>(a.k_cache);
  const T* v_cache = reinterpret_cast<const T*>(a.v_cache);
  const int kv_head_off = kvh * D;
  const uint32_t warp_pipe = smem_u32(pipe) + warp * STAGES * S::kStageBytes;

  const int n_blocks = (n_tok + kBlkTok - 1) / kBlkTok;            // blocks in the chunk
  const int my_blocks = (n_blocks - warp + kDecWarps - 1) / kDecWarps;  // blocks warp, warp+4, ...

  // ---- 2. pipeline prologue ---------------------------------------------------------------
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < my_blocks) {
      const int tok0 = (warp + s * kDecWarps) * kBlkTok;
      issue_block<T, D>(k_cache, v_cache, a.kv_token_stride, kv_head_off, s_slots, tok0, n_tok,
                        warp_pipe + s * S::kStageBytes, warp_pipe + s * S::kStageBytes + S::kBlkBytes,
                        lane);
    }
    cp_async_commit();
  }

  // ---- 3. Q fragments (rows = q heads of the group), overlapped with the first copies -----
  const int head0 = kvh * a.group + hb * 16;
  const int n_heads = min(16, a.group - hb * 16);
  uint32_t qf[D / 16][4];
  {
    const T* qb = reinterpret_cast<const T*>(a.q) + (int64_t)b * a.q_token_stride + (int64_t)head0 * D;
    const bool v0 = g < n_heads;
    const bool v1 = UPPER && (g + 8 < n_heads);
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      const int col = ks * 16 + 2 * t;
      qf[ks][0] = v0 ? *reinterpret_cast<const uint32_t*>(qb + g * D + col) : 0u;
      qf[ks][1] = v1 ? *reinterpret_cast<const uint32_t*>(qb + (g + 8) * D + col) : 0u;
      qf[ks][2] = v0 ? *reinterpret_cast<const uint32_t*>(qb + g * D + col + 8) : 0u;
      qf[ks][3] = v1 ? *reinterpret_cast<const uint32_t*>(qb + (g + 8) * D + col + 8) : 0u;
    }
  }

  float o_acc[D / 8][4];
#pragma unroll
  for (int i = 0; i < D / 8; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};  // running max (log2 domain), rows g and g+8
  float l_run[2] = {0.f, 0.f};              // per-thread partial row sums

  const bool use_cap = a.logit_cap > 0.f;
  const float cap_in = use_cap ? a.sm_scale / a.logit_cap : 0.f;
  const float cap_out = a.logit_cap * kLog2e;

  // ldmatrix lane addressing (see header comment of each use)
  const int lm_row = lane & 7;
  const int lm_mat = lane >> 3;

  // ---- 4. main loop over this warp's blocks ----------------------------------------------------
  for (int i = 0; i < my_blocks; ++i) {
    // prefetch block i + STAGES - 1
    {
      const int nb = i + STAGES - 1;
      if (nb < my_blocks) {
        const int st = nb % STAGES;
        const int tok0 = (warp + nb * kDecWarps) * kBlkTok;
        issue_block<T, D>(k_cache, v_cache, a.kv_token_stride, kv_head_off, s_slots, tok0, n_tok,
                          warp_pipe + st * S::kStageBytes,
                          warp_pipe + st * S::kStageBytes + S::kBlkBytes, lane);
      }
      cp_async_commit();
    }
    cp_async_wait<STAGES - 1>();
    __syncwarp();

    const int st = i % STAGES;
    const uint32_t sk = warp_pipe + st * S::kStageBytes;
    const uint32_t sv = sk + S::kBlkBytes;
    const int tok0 = (warp + i * kDecWarps) * kBlkTok;

    // S = Q K^T : [16 heads x 16 tokens] as two n-tiles of 8 tokens
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      // matrices: 0:(tok 0-7, dims lo) 1:(tok 0-7, dims hi) 2:(tok 8-15, lo) 3:(tok 8-15, hi)
      const int r = (lm_mat >> 1) * 8 + lm_row;
      const int c = ks * 2 + (lm_mat & 1);
      uint32_t b0, b1, b2, b3;
      ldmatrix_x4(b0, b1, b2, b3, sk + r * (D * 2) + ((c ^ (r & 7)) << 4));
      mma_16816<T>(s[0], qf[ks], b0, b1);
      mma_16816<T>(s[1], qf[ks], b2, b3);
    }

    // scale (+ optional logit cap) into the log2 domain, mask the tail of the chunk
#pragma unroll
    for (int j = 0; j < 2; ++j) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = s[j][e];
        if (use_cap) v = cap_out * tanh_approx(v * cap_in);
        else v *= a.scale_log2;
        const int tok = tok0 + j * 8 + 2 * t + (e & 1);
        s[j][e] = tok < n_tok ? v : -INFINITY;
      }
    }

    // online softmax, rows g (regs 0,1) and g+8 (regs 2,3)
    float alpha[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !UPPER) { alpha[1] = 1.f; break; }
      float mx = fmaxf(fmaxf(s[0][2 * h], s[0][2 * h + 1]), fmaxf(s[1][2 * h], s[1][2 * h + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[h], mx);  // finite: every block has >= 1 valid token
      alpha[h] = fast_exp2(m_run[h] - m_new);
      m_run[h] = m_new;
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        s[j][2 * h] = fast_exp2(s[j][2 * h] - m_new);
        s[j][2 * h + 1] = fast_exp2(s[j][2 * h + 1] - m_new);
        sum += s[j][2 * h] + s[j][2 * h + 1];
      }
      l_run[h] = l_run[h] * alpha[h] + sum;
    }
    if (!UPPER) { s[0][2] = s[0][3] = s[1][2] = s[1][3] = 0.f; }

    // P as the A operand of P.V
    uint32_t pf[4];
    pf[0] = pack2<T>(s[0][0], s[0][1]);
    pf[1] = pack2<T>(s[0][2], s[0][3]);
    pf[2] = pack2<T>(s[1][0], s[1][1]);
    pf[3] = pack2<T>(s[1][2], s[1][3]);

#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) {
      o_acc[nt][0] *= alpha[0];
      o_acc[nt][1] *= alpha[0];
      if (UPPER) {
        o_acc[nt][2] *= alpha[1];
        o_acc[nt][3] *= alpha[1];
      }
    }
#pragma unroll
    for (int np = 0; np < D / 16; ++np) {
      // matrices (transposed on load): 0:(tok 0-7, dims np*16..+7) 1:(tok 8-15, same dims)
      //                                2:(tok 0-7, dims +8..+15)   3:(tok 8-15, dims +8..+15)
      const int r = (lm_mat & 1) * 8 + lm_row;
      const int c = np * 2 + (lm_mat >> 1);
      uint32_t b0, b1, b2, b3;
      ldmatrix_x4_trans(b0, b1, b2, b3, sv + r * (D * 2) + ((c ^ (r & 7)) << 4));
      mma_16816<T>(o_acc[2 * np], pf, b0, b1);
      mma_16816<T>(o_acc[2 * np + 1], pf, b2, b3);
    }
    __syncwarp();  // all lanes done reading this stage before it is refilled
  }
  cp_async_wait<0>();

  // ---- 5. merge the 4 warps through shared memory ----------------------------------------------
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
  }
  __syncthreads();  // every warp is done with its pipeline buffers; reuse them
  float* sm_o = reinterpret_cast<float*>(pipe);                        // [warps][16][kOPad]
  float* sm_m = sm_o + kDecWarps * 16 * S::kOPad;                      // [warps][16]
  float* sm_l = sm_m + kDecWarps * 16;                                 // [warps][16]
  {
    float* wo = sm_o + warp * 16 * S::kOPad;
#pragma unroll
    for (int nt = 0; nt < D / 8; ++nt) {
      *reinterpret_cast<float2*>(wo + g * S::kOPad + nt * 8 + 2 * t) = make_float2(o_acc[nt][0], o_acc[nt][1]);
      if (UPPER)
        *reinterpret_cast<float2*>(wo + (g + 8) * S::kOPad + nt * 8 + 2 * t) =
            make_float2(o_acc[nt][2], o_acc[nt][3]);
    }
    if (t == 0) {
      sm_m[warp * 16 + g] = m_run[0];
      sm_l[warp * 16 + g] = l_run[0];
      sm_m[warp * 16 + g + 8] = UPPER ? m_run[1] : -INFINITY;
      sm_l[warp * 16 + g + 8] = UPPER ? l_run[1] : 0.f;
    }
  }
  __syncthreads();

  combine_warps_and_store<T, D>(a, sm_o, sm_m, sm_l, n_heads, head0, b, split, num_splits, 1.f);
}

// ---- FP8 (e4m3) KV cache ---------------------------------------------------------------------------------------
// --kv-cache-dtype fp8_e4m3 (reference memory_pool.py:114-118 stores it as uint8): the step's KV bytes
// halve.  Semantics follow the reference's Triton decode kernel, which up-casts K to the q dtype before the dot
// (decode_attention.py:336): K and V are taken at their exact e4m3 values; k_scale / v_scale (the factors
// set_kv_buffer divided by, memory_pool.py:385-391) are folded into the softmax scale and the final 1/l.
//
// Same CTA/warp decomposition as decode_attn_kernel.  Rows are D bytes, so there is no 16-bit ldmatrix path; instead
// the mma fragments are read straight from shared memory with 16-byte loads and converted in registers
// (cvt.rn.f16x2.e4m3x2, exact), using the freedom to permute the reduction axis of an MMA:
//  * S = Q K^T: thread (g, t) reads bytes [t*D/4, (t+1)*D/4) of token row g; k-step ks uses the 4 bytes at
//    t*D/4 + 4*ks as k-slots {2t, 2t+1, 2t+8, 2t+9}.  The Q fragments are loaded with the same dim permutation.
//  * O += P V: thread (g, t) reads D/8 consecutive bytes at g*D/8 of token rows {2t, 2t+1, 2t+8, 2t+9}; byte nt pairs
//    up (prmt) into the B fragment of n-tile nt, i.e. accumulator column j of n-tile nt is dim j*D/8 + nt.  The
//    permutation is undone when the accumulators are written to shared memory for the warp merge.
// Both MMAs run in f16 (e4m3 and P are exact / finer in f16 than in bf16; a bf16 q is converted with a clamp to the
// f16 range, exact for |q| in [2^-14, 65504]).
template <int D>
struct DecodeSmem8 {
  static constexpr int kChunks = D / 16;                      // 16-byte chunks per row
  static constexpr int kBlkBytes = kBlkTok * D;               // one K (or V) block of a warp
  static constexpr int kStageBytes = 2 * kBlkBytes;
  static constexpr int pipe_bytes(int stages) { return kDecWarps * stages * kStageBytes; }
  // physical 16-byte chunk of logical chunk c in row r (bank-conflict-free for both access patterns at D = 128)
  static __device__ __forceinline__ int swz(int c, int r) { return c ^ ((r / (8 / kChunks)) & (kChunks - 1)); }
};

template <int D>
__device__ __forceinline__ void issue_block8(const uint8_t* __restrict__ k_cache, const uint8_t* __restrict__ v_cache,
                                             int64_t kv_token_stride, int kv_head_off,
                                             const int32_t* __restrict__ s_slots, int tok0, int n_tok,
                                             uint32_t smem_k, uint32_t smem_v, int lane) {
  using S8 = DecodeSmem8<D>;
  constexpr int kChunks = S8::kChunks;
  constexpr int kRowsPerIter = 32 / kChunks;
  const int c = lane % kChunks;
  const int r0 = lane / kChunks;
#pragma unroll
  for (int it = 0; it < kBlkTok / kRowsPerIter; ++it) {
    const int r = r0 + it * kRowsPerIter;
    const int tok = tok0 + r;
    const bool valid = tok < n_tok;
    const int slot = valid ? s_slots[tok] : 0;
    const int64_t off = (int64_t)slot * kv_token_stride + kv_head_off + c * 16;
    const uint32_t so = (uint32_t)(r * D + (S8::swz(c, r) << 4));
    cp_async16_pred(smem_k + so, k_cache + off, valid);
    cp_async16_pred(smem_v + so, v_cache + off, valid);
  }
}

// two e4m3 bytes (low half of `pair` after the shift) -> f16x2, low byte in the low half
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t word, int hi) {
  uint32_t d;
  const uint16_t h = (uint16_t)(hi ? (word >> 16) : (word & 0xffffu));
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;\n" : "=r"(d) : "h"(h));
  return d;
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t d;
  asm("prmt.b32 %0, %1, %2, %3;\n" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
  return d;
}
__device__ __forceinline__ uint32_t lds32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.b32 %0, [%1];\n" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.b32 {%0,%1}, [%2];\n" : "=r"(v.x), "=r"(v.y) : "r"(addr));
  return v;
}
// 4 consecutive q values (dims d..d+3 of one head) -> two f16x2 registers
template <typename T>
__device__ __forceinline__ void load_q4_f16(const T* p, bool valid, uint32_t& lo, uint32_t& hi) {
  if (!valid) { lo = hi = 0u; return; }
  const uint2 raw = *reinterpret_cast<const uint2*>(p);
  if constexpr (std::is_same<T, __half>::value) {
    lo = raw.x; hi = raw.y;
  } else {
    const float2 a = unpack2<T>(raw.x), c = unpack2<T>(raw.y);
    lo = pack2<__half>(fminf(fmaxf(a.x, -65504.f), 65504.f), fminf(fmaxf(a.y, -65504.f), 65504.f));
    hi = pack2<__half>(fminf(fmaxf(c.x, -65504.f), 65504.f), fminf(fmaxf(c.y, -65504.f), 65504.f));
  }
}

template <typename T, int D, int STAGES, bool UPPER, int MINB>
__global__ void __launch_bounds__(kDecThreads, MINB) decode_attn_fp8kv_kernel(const DecodeArgs a) {
  using S = DecodeSmem<D>;
  using S8 = DecodeSmem8<D>;
  using H = __half;
  constexpr int NT = D / 8;       // n-tiles of O; also the bytes a thread reads per V row
  constexpr int KS = D / 16;      // k-steps of S = Q K^T; also the 4-byte words a thread reads per K row
  extern __shared__ __align__(128) uint8_t smem_raw[];

  const int split = blockIdx.x;
  const int kvh = blockIdx.y / a.head_blocks;
  const int hb = blockIdx.y - kvh * a.head_blocks;
  const int b = blockIdx.z;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int g = lane >> 2;
  const int t = lane & 3;

  pdl_trigger();
  pdl_wait();
  const int seq_len = (int)load_index(a.seq_lens, b, a.index_is_int64) - a.kv_start;  // tokens this kernel covers
  const int start = split * a.kv_chunk;
  if (start >= seq_len) return;
  const int n_tok = min(seq_len - start, a.kv_chunk);
  const int num_splits = a.n_seg + (seq_len + a.kv_chunk - 1) / a.kv_chunk;

  uint8_t* pipe = smem_raw;
  int32_t* s_slots = reinterpret_cast<int32_t*>(smem_raw + S8::pipe_bytes(STAGES));
  {
    const int64_t req = load_index(a.req_pool_indices, b, a.index_is_int64);
    const int32_t* row = a.req_to_token + req * a.req_to_token_stride + a.kv_start + start;
    for (int i = threadIdx.x; i < n_tok; i += kDecThreads) s_slots[i] = __ldg(row + i);
  }
  __syncthreads();

  const uint8_t* k_cache = reinterpret_cast<const uint8_t*>(a.k_cache);
  const uint8_t* v_cache = reinterpret_cast<const uint8_t*>(a.v_cache);
  const int kv_head_off = kvh * D;
  const uint32_t warp_pipe = smem_u32(pipe) + warp * STAGES * S8::kStageBytes;
  const int n_blocks = (n_tok + kBlkTok - 1) / kBlkTok;
  const int my_blocks = (n_blocks - warp + kDecWarps - 1) / kDecWarps;

#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (s < my_blocks) {
      const int tok0 = (warp + s * kDecWarps) * kBlkTok;
      issue_block8<D>(k_cache, v_cache, a.kv_token_stride, kv_head_off, s_slots, tok0, n_tok,
                      warp_pipe + s * S8::kStageBytes, warp_pipe + s * S8::kStageBytes + S8::kBlkBytes, lane);
    }
    cp_async_commit();
  }

  // Q fragments in f16 with the k-axis permutation of the header comment: k-step ks <-> dims t*D/4 + 4*ks + {0..3}
  const int head0 = kvh * a.group + hb * 16;
  const int n_heads = min(16, a.group - hb * 16);
  uint32_t qf[KS][4];
  {
    const T* qb = reinterpret_cast<const T*>(a.q) + (int64_t)b * a.q_token_stride + (int64_t)head0 * D;
    const bool v0 = g < n_heads;
    const bool v1 = UPPER && (g + 8 < n_heads);
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int col = t * (D / 4) + 4 * ks;
      load_q4_f16<T>(qb + g * D + col, v0, qf[ks][0], qf[ks][2]);
      load_q4_f16<T>(qb + (g + 8) * D + col, v1, qf[ks][1], qf[ks][3]);
    }
  }

  float o_acc[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) o_acc[i][0] = o_acc[i][1] = o_acc[i][2] = o_acc[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY};
  float l_run[2] = {0.f, 0.f};

  const bool use_cap = a.logit_cap > 0.f;
  const float cap_in = use_cap ? a.sm_scale / a.logit_cap : 0.f;
  const float cap_out = a.logit_cap * kLog2e;

  for (int i = 0; i < my_blocks; ++i) {
    {
      const int nb = i + STAGES - 1;
      if (nb < my_blocks) {
        const int st = nb % STAGES;
        const int tok0 = (warp + nb * kDecWarps) * kBlkTok;
        issue_block8<D>(k_cache, v_cache, a.kv_token_stride, kv_head_off, s_slots, tok0, n_tok,
                        warp_pipe + st * S8::kStageBytes, warp_pipe + st * S8::kStageBytes + S8::kBlkBytes, lane);
      }
      cp_async_commit();
    }
    cp_async_wait<STAGES - 1>();
    __syncwarp();

    const int st = i % STAGES;
    const uint32_t sk = warp_pipe + st * S8::kStageBytes;
    const uint32_t sv = sk + S8::kBlkBytes;
    const int tok0 = (warp + i * kDecWarps) * kBlkTok;

    // S = Q K^T
    float s[2][4];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int r = j * 8 + g;
      uint32_t w[KS];
#pragma unroll
      for (int c = 0; c < KS / 4; ++c) {
        const uint4 v = lds128(sk + r * D + (S8::swz(t * (KS / 4) + c, r) << 4));
        w[4 * c] = v.x; w[4 * c + 1] = v.y; w[4 * c + 2] = v.z; w[4 * c + 3] = v.w;
      }
      // even / odd k-steps accumulate separately: with two warps per scheduler the dependent-HMMA chain is what the
      // loop waits on, so keep four chains (2 n-tiles x 2) in flight instead of two
      float s2[4] = {0.f, 0.f, 0.f, 0.f};
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
#pragma unroll
      for (int ks = 0; ks < KS; ks += 2) {
        mma_16816<H>(s[j], qf[ks], e4m3x2_to_f16x2(w[ks], 0), e4m3x2_to_f16x2(w[ks], 1));
        mma_16816<H>(s2, qf[ks + 1], e4m3x2_to_f16x2(w[ks + 1], 0), e4m3x2_to_f16x2(w[ks + 1], 1));
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) s[j][e] += s2[e];
    }

#pragma unroll
    for (int j = 0; j < 2; ++j) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = s[j][e];
        if (use_cap) v = cap_out * tanh_approx(v * cap_in);
        else v *= a.scale_log2;
        const int tok = tok0 + j * 8 + 2 * t + (e & 1);
        s[j][e] = tok < n_tok ? v : -INFINITY;
      }
    }

    float alpha[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (h == 1 && !UPPER) { alpha[1] = 1.f; break; }
      float mx = fmaxf(fmaxf(s[0][2 * h], s[0][2 * h + 1]), fmaxf(s[1][2 * h], s[1][2 * h + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[h], mx);
      alpha[h] = fast_exp2(m_run[h] - m_new);
      m_run[h] = m_new;
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        s[j][2 * h] = fast_exp2(s[j][2 * h] - m_new);
        s[j][2 * h + 1] = fast_exp2(s[j][2 * h + 1] - m_new);
        sum += s[j][2 * h] + s[j][2 * h + 1];
      }
      l_run[h] = l_run[h] * alpha[h] + sum;
    }
    if (!UPPER) { s[0][2] = s[0][3] = s[1][2] = s[1][3] = 0.f; }

    uint32_t pf[4];
    pf[0] = pack2<H>(s[0][0], s[0][1]);
    pf[1] = pack2<H>(s[0][2], s[0][3]);
    pf[2] = pack2<H>(s[1][0], s[1][1]);
    pf[3] = pack2<H>(s[1][2], s[1][3]);

#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      o_acc[nt][0] *= alpha[0];
      o_acc[nt][1] *= alpha[0];
      if (UPPER) {
        o_acc[nt][2] *= alpha[1];
        o_acc[nt][3] *= alpha[1];
      }
    }

    // O += P V: token rows {2t, 2t+1, 2t+8, 2t+9}, bytes [g*NT, (g+1)*NT)
    uint32_t vw[4][NT / 4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int r = 2 * t + (q & 1) + (q >> 1) * 8;
      if constexpr (NT == 16) {
        const uint4 v = lds128(sv + r * D + (S8::swz(g, r) << 4));
        vw[q][0] = v.x; vw[q][1] = v.y; vw[q][2] = v.z; vw[q][3] = v.w;
      } else {
        const uint2 v = lds64(sv + r * D + (S8::swz(g >> 1, r) << 4) + (g & 1) * 8);
        vw[q][0] = v.x; vw[q][1] = v.y;
      }
    }
#pragma unroll
    for (int wi = 0; wi < NT / 4; ++wi) {
      const uint32_t lo01 = prmt(vw[0][wi], vw[1][wi], 0x5140u), lo23 = prmt(vw[0][wi], vw[1][wi], 0x7362u);
      const uint32_t hi01 = prmt(vw[2][wi], vw[3][wi], 0x5140u), hi23 = prmt(vw[2][wi], vw[3][wi], 0x7362u);
      mma_16816<H>(o_acc[4 * wi + 0], pf, e4m3x2_to_f16x2(lo01, 0), e4m3x2_to_f16x2(hi01, 0));
      mma_16816<H>(o_acc[4 * wi + 1], pf, e4m3x2_to_f16x2(lo01, 1), e4m3x2_to_f16x2(hi01, 1));
      mma_16816<H>(o_acc[4 * wi + 2], pf, e4m3x2_to_f16x2(lo23, 0), e4m3x2_to_f16x2(hi23, 0));
      mma_16816<H>(o_acc[4 * wi + 3], pf, e4m3x2_to_f16x2(lo23, 1), e4m3x2_to_f16x2(hi23, 1));
    }
    __syncwarp();
  }
  cp_async_wait<0>();

#pragma unroll
  for (int h = 0; h < 2; ++h) {
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 1);
    l_run[h] += __shfl_xor_sync(0xffffffffu, l_run[h], 2);
  }
  __syncthreads();
  float* sm_o = reinterpret_cast<float*>(pipe);
  float* sm_m = sm_o + kDecWarps * 16 * S::kOPad;
  float* sm_l = sm_m + kDecWarps * 16;
  {
    float* wo = sm_o + warp * 16 * S::kOPad;
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {  // accumulator column j of n-tile nt is dim j * NT + nt
      wo[g * S::kOPad + (2 * t) * NT + nt] = o_acc[nt][0];
      wo[g * S::kOPad + (2 * t + 1) * NT + nt] = o_acc[nt][1];
      if (UPPER) {
        wo[(g + 8) * S::kOPad + (2 * t) * NT + nt] = o_acc[nt][2];
        wo[(g + 8) * S::kOPad + (2 * t + 1) * NT + nt] = o_acc[nt][3];
      }
    }
    if (t == 0) {
      sm_m[warp * 16 + g] = m_run[0];
      sm_l[warp * 16 + g] = l_run[0];
      sm_m[warp * 16 + g + 8] = UPPER ? m_run[1] : -INFINITY;
      sm_l[warp * 16 + g + 8] = UPPER ? l_run[1] : 0.f;
    }
  }
  __syncthreads();
  combine_warps_and_store<T, D>(a, sm_o, sm_m, sm_l, n_heads, head0, b, split, num_splits, a.out_scale);
}

// Stage 2: combine the chunks of each (request, q head).  One warp per (b, h); each lane owns
// D/32 consecutive dims.  Requests with one chunk were finalised by stage 1; requests with
// seq_len 0 get zeros.
template <typename T, int D>
__global__ void __launch_bounds__(128) decode_merge_kernel(const DecodeArgs a) {
  pdl_trigger();
  pdl_wait();
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp_global >= a.batch * a.num_q_heads) return;
  const int b = warp_global / a.num_q_heads;
  const int h = warp_global - b * a.num_q_heads;
  const int seq_len = max((int)load_index(a.seq_lens, b, a.index_is_int64) - a.kv_start, 0);
  const int num_splits = a.n_seg + (seq_len + a.kv_chunk - 1) / a.kv_chunk;
  if (num_splits == 1) return;
  constexpr int kPer = D / 32;
  T* dst = reinterpret_cast<T*>(a.o) + (int64_t)b * a.o_token_stride + (int64_t)h * D + lane * kPer;
  float acc[kPer];
#pragma unroll
  for (int d = 0; d < kPer; ++d) acc[d] = 0.f;
  if (num_splits > 1) {
    const int64_t base = ((int64_t)b * a.num_q_heads + h) * a.ws_splits;
    float M = -INFINITY;
    for (int s = lane; s < num_splits; s += 32) M = fmaxf(M, a.part_lse[base + s]);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, off));
    float L = 0.f;
    for (int s = 0; s < num_splits; ++s) {
      const float w = fast_exp2(a.part_lse[base + s] - M);
      L += w;
      const float* src = a.part_o + (base + s) * D + lane * kPer;
      if (kPer == 4) {
        const float4 v = *reinterpret_cast<const float4*>(src);
        acc[0] += v.x * w; acc[1] += v.y * w; acc[2] += v.z * w; acc[3] += v.w * w;
      } else {
        const float2 v = *reinterpret_cast<const float2*>(src);
        acc[0] += v.x * w; acc[1] += v.y * w;
      }
    }
    const float inv = 1.f / L;
#pragma unroll
    for (int d = 0; d < kPer; ++d) acc[d] *= inv;
  }
#pragma unroll
  for (int d = 0; d < kPer; d += 2) *reinterpret_cast<uint32_t*>(dst + d) = pack2<T>(acc[d], acc[d + 1]);
}

// Stage 2 fused with the per-token int8 quantisation that feeds o_proj (b200_decode_attn_int8): one CTA per request
// merges every q head (or picks up the bf16 row stage 1 already finalised), rounds to T exactly like
// decode_merge_kernel, and then quantises the [Hq * D] row exactly like per_token_quant_int8_kernel (quant.cu) —
// bit-identical to the two kernels back to back, one launch and one activation round-trip less.
template <typename T, int D>
__global__ void __launch_bounds__(1024) decode_merge_quant_kernel(const DecodeArgs a) {
  extern __shared__ float row_f[];  // Hq * D
  __shared__ float red[32];
  pdl_trigger();
  pdl_wait();
  const int b = blockIdx.x;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int seq_len = max((int)load_index(a.seq_lens, b, a.index_is_int64) - a.kv_start, 0);
  const int num_splits = a.n_seg + (seq_len + a.kv_chunk - 1) / a.kv_chunk;
  constexpr int kPer = D / 32;
  float amax = 0.f;
  for (int h = warp; h < a.num_q_heads; h += nwarps) {  // one warp per q head: every head's loads are in flight at once
    T* dst = reinterpret_cast<T*>(a.o) + (int64_t)b * a.o_token_stride + (int64_t)h * D + lane * kPer;
    float acc[kPer];
    if (num_splits == 1) {
#pragma unroll
      for (int d = 0; d < kPer; d += 2) {
        const float2 v = unpack2<T>(*reinterpret_cast<const uint32_t*>(dst + d));
        acc[d] = v.x;
        acc[d + 1] = v.y;
      }
    } else {
#pragma unroll
      for (int d = 0; d < kPer; ++d) acc[d] = 0.f;
      if (num_splits > 1) {
        const int64_t base = ((int64_t)b * a.num_q_heads + h) * a.ws_splits;
        float M = -INFINITY;
        for (int s = lane; s < num_splits; s += 32) M = fmaxf(M, a.part_lse[base + s]);
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) M = fmaxf(M, __shfl_xor_sync(0xffffffffu, M, off));
        float L = 0.f;
        for (int s0 = 0; s0 < num_splits; s0 += 4) {  // 4 chunks' partial rows loaded before any is used
          float pv[4][kPer];
          float pl[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const bool ok = s0 + i < num_splits;
            const int s = ok ? s0 + i : s0;
            pl[i] = ok ? a.part_lse[base + s] : -INFINITY;
            const float* src = a.part_o + (base + s) * D + lane * kPer;
            if (kPer == 4) {
              const float4 v = *reinterpret_cast<const float4*>(src);
              pv[i][0] = v.x; pv[i][1] = v.y; pv[i][2] = v.z; pv[i][kPer - 1] = v.w;
            } else {
              const float2 v = *reinterpret_cast<const float2*>(src);
              pv[i][0] = v.x; pv[i][1] = v.y;
            }
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {  // same accumulation order as decode_merge_kernel (s ascending)
            if (s0 + i < num_splits) {
              const float w = fast_exp2(pl[i] - M);
              L += w;
#pragma unroll
              for (int d = 0; d < kPer; ++d) acc[d] += pv[i][d] * w;
            }
          }
        }
        const float inv = 1.f / L;
#pragma unroll
        for (int d = 0; d < kPer; ++d) acc[d] *= inv;
      }
#pragma unroll
      for (int d = 0; d < kPer; d += 2) {
        const uint32_t u = pack2<T>(acc[d], acc[d + 1]);
        *reinterpret_cast<uint32_t*>(dst + d) = u;
        const float2 v = unpack2<T>(u);  // the T-rounded value the un-fused quant kernel would read
        acc[d] = v.x;
        acc[d + 1] = v.y;
      }
    }
#pragma unroll
    for (int d = 0; d < kPer; ++d) {
      row_f[h * D + lane * kPer + d] = acc[d];
      amax = fmaxf(amax, fabsf(acc[d]));
    }
  }
#pragma unroll
  for (int off = 16; off > 0; off >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, off));
  if (lane == 0) red[warp] = amax;
  __syncthreads();
  amax = red[0];
  for (int i = 1; i < nwarps; ++i) amax = fmaxf(amax, red[i]);
  amax = fmaxf(amax, 1e-10f);
  const float inv = __fmul_rn(__frcp_rn(amax), 127.f);
  if (threadIdx.x == 0) a.oq_scale[b] = __fdiv_rn(amax, 127.f);
  const int H = a.num_q_heads * D;
  int8_t* qr = a.oq + (int64_t)b * H;
  for (int c = threadIdx.x * 8; c < H; c += (int)blockDim.x * 8) {
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int x0 = __float2int_rn(__fmul_rn(row_f[c + i], inv));
      const int x1 = __float2int_rn(__fmul_rn(row_f[c + 4 + i], inv));
      lo |= (uint32_t)(x0 & 0xff) << (8 * i);
      hi |= (uint32_t)(x1 & 0xff) << (8 * i);
    }
    *reinterpret_cast<uint2*>(qr + c) = make_uint2(lo, hi);
  }
}

template <typename T, int D, int STAGES, int MINB>
static int launch_decode_fp8kv_stage1(const DecodeArgs& a, cudaStream_t stream) {
  using S = DecodeSmem<D>;
  constexpr int pipe = DecodeSmem8<D>::pipe_bytes(STAGES);
  static_assert(S::merge_bytes() <= pipe, "merge scratch must fit in the pipeline buffers");
  const int smem = pipe + a.kv_chunk * 4;
  dim3 grid(a.max_splits, a.num_kv_heads * a.head_blocks, a.batch);
  if (a.group > 8) {
    auto kern = decode_attn_fp8kv_kernel<T, D, STAGES, true, MINB>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return launch_k(kern, grid, dim3(kDecThreads), smem, stream, 1, true, a);
  }
  auto kern = decode_attn_fp8kv_kernel<T, D, STAGES, false, MINB>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  return launch_k(kern, grid, dim3(kDecThreads), smem, stream, 1, true, a);
}

template <typename T, int D, bool KV8>
static int launch_decode(const DecodeArgs& a, cudaStream_t stream, cudaEvent_t join_before_merge = nullptr) {
  constexpr int STAGES = 3;
  using S = DecodeSmem<D>;
  constexpr int pipe = S::pipe_bytes(STAGES);
  static_assert(S::merge_bytes() <= pipe, "merge scratch must fit in the pipeline buffers");
  const int smem = pipe + a.kv_chunk * 4;
  dim3 grid(a.max_splits, a.num_kv_heads * a.head_blocks, a.batch);
  const bool upper = a.group > 8;
  int rc;
  if constexpr (KV8) {
    // 4 stages x 3 CTAs/SM (measured 0.118 ms at cfg 2, profiles/r01g_kbench_fp8kv.jsonl) vs 6 stages x 2 CTAs/SM
    // (0.128 ms, B200_FP8KV_OCC=2): the register-side conversions want more warps per scheduler
    static const int occ = [] { const char* e = getenv("B200_FP8KV_OCC"); return e ? atoi(e) : 3; }();
    rc = occ == 3 ? launch_decode_fp8kv_stage1<T, D, 4, 3>(a, stream) : launch_decode_fp8kv_stage1<T, D, 6, 2>(a, stream);
  } else if (upper) {
    auto kern = decode_attn_kernel<T, D, STAGES, true>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    rc = launch_k(kern, grid, dim3(kDecThreads), smem, stream, 1, true, a);
  } else {
    auto kern = decode_attn_kernel<T, D, STAGES, false>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    rc = launch_k(kern, grid, dim3(kDecThreads), smem, stream, 1, true, a);
  }
  if (rc != B200_OK) return rc;
  if (join_before_merge != nullptr) B200_CUDA_OK(cudaStreamWaitEvent(stream, join_before_merge, 0));
  if (a.oq != nullptr) {
    const int qsmem = a.num_q_heads * D * (int)sizeof(float);
    auto kern = decode_merge_quant_kernel<T, D>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, qsmem));
    int threads = 32 * a.num_q_heads;
    threads = threads > 1024 ? 1024 : threads < 128 ? 128 : threads;
    return launch_k(kern, dim3(a.batch), dim3(threads), qsmem, stream, 1, true, a);
  }
  const int warps = a.batch * a.num_q_heads;
  return launch_k(decode_merge_kernel<T, D>, dim3((warps + 3) / 4), dim3(128), 0, stream, 1, true, a);
}

}  // namespace b200

B200_API int b200_decode_attn_workspace_bytes(int batch, int num_q_heads, int head_dim,
                                                int max_splits, size_t* bytes) {
  using namespace b200;
  B200_REQUIRE(bytes != nullptr, "decode_attn_workspace_bytes: null out pointer");
  B200_REQUIRE(batch >= 0 && num_q_heads > 0 && head_dim > 0 && max_splits > 0,
               "decode_attn_workspace_bytes: bad sizes");
  const size_t n = (size_t)batch * num_q_heads * max_splits;
  *bytes = n * head_dim * sizeof(float) + n * sizeof(float) + 256;
  return B200_OK;
}


// Side stream + fork / join events of the shared-prefix overlap, one set per device ordinal (created under a mutex on
// first use on that device; a process that drives several GPUs gets a stream that belongs to each).
struct SideStream {
  cudaStream_t stream = nullptr;
  cudaEvent_t fork = nullptr, join = nullptr;
};
static int side_stream_for_current_device(SideStream* out) {
  static std::mutex mu;
  static SideStream per_device[64];
  int dev = 0;
  B200_CUDA_OK(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < 64, "decode_attn_shared_prefix: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  SideStream& s = per_device[dev];
  if (s.stream == nullptr) {
    B200_CUDA_OK(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking));
    B200_CUDA_OK(cudaEventCreateWithFlags(&s.fork, cudaEventDisableTiming));
    B200_CUDA_OK(cudaEventCreateWithFlags(&s.join, cudaEventDisableTiming));
  }
  *out = s;
  return B200_OK;
}

// prefix pass of the shared-prefix decode: the two-query-tile TMA kernel when the K/V extent is known and the batch
// fills more than one 128-row tile, else the one-tile cp.async kernel
static int launch_prefix_pass(const b200::ExtendArgs& e, int batch, int head_dim, int dtype, int64_t kv_rows,
                              cudaStream_t stream) {
  using namespace b200;
  static const bool fa = getenv("B200_EXTEND_FA") == nullptr || getenv("B200_EXTEND_FA")[0] != '0';
  if (fa && kv_rows > 0 && batch > 128 && e.q_token_stride % 8 == 0 && e.kv_token_stride % 8 == 0) {
    const int rc = launch_extend_fa(e, batch, head_dim, dtype, batch, kv_rows, stream);
    if (rc != B200_ERR_UNSUPPORTED) return rc;
  }
  return launch_extend_umma(e, batch, head_dim, dtype, stream);
}

static int decode_attn_impl(void* o, const void* q, const void* k_cache, const void* v_cache,
                            const int32_t* req_to_token, int64_t req_to_token_stride,
                            const void* req_pool_indices, const void* seq_lens,
                            int index_is_int64, int batch, int num_q_heads, int num_kv_heads,
                            int head_dim, int64_t q_token_stride, int64_t o_token_stride,
                            int64_t kv_token_stride, float sm_scale, float logit_cap,
                            int kv_chunk, int max_splits, void* workspace,
                            size_t workspace_bytes, int dtype, int8_t* oq, float* oq_scale, void* stream,
                            bool kv_fp8 = false, float k_scale = 1.f, float v_scale = 1.f, int shared_prefix = 0,
                            int seg_len = 0, const void* k_stage = nullptr, const void* v_stage = nullptr,
                            const int32_t* stage_table = nullptr, const void* stage_index = nullptr,
                            int64_t stage_token_stride = 0, int64_t num_kv_slots = 0) {
  using namespace b200;
  B200_REQUIRE(batch >= 0, "decode_attn: negative batch");
  if (batch == 0) return B200_OK;
  B200_REQUIRE(o && q && k_cache && v_cache && req_to_token && req_pool_indices && seq_lens,
               "decode_attn: null pointer");
  B200_REQUIRE(num_kv_heads > 0 && num_q_heads % num_kv_heads == 0,
               "decode_attn: num_q_heads %d not a multiple of num_kv_heads %d", num_q_heads, num_kv_heads);
  B200_REQUIRE(kv_chunk >= 16 && kv_chunk % 16 == 0 && kv_chunk <= 4096,
               "decode_attn: kv_chunk %d must be a multiple of 16 in [16, 4096]", kv_chunk);
  B200_REQUIRE(max_splits >= 1 && max_splits <= 65535, "decode_attn: max_splits %d", max_splits);
  B200_REQUIRE(batch <= 65535, "decode_attn: batch %d exceeds grid.z", batch);
  B200_REQUIRE(q_token_stride % 2 == 0 && o_token_stride % 2 == 0 && kv_token_stride % (kv_fp8 ? 16 : 8) == 0,
               "decode_attn: strides must keep 4-byte (q/o) and 16-byte (kv) alignment");
  B200_REQUIRE(!kv_fp8 || (q_token_stride % 4 == 0 && (uintptr_t)q % 8 == 0),
               "decode_attn_fp8kv: q rows must be 8-byte aligned");
  B200_REQUIRE(k_scale > 0.f && v_scale > 0.f, "decode_attn_fp8kv: k_scale / v_scale must be positive");
  B200_REQUIRE(((uintptr_t)k_cache | (uintptr_t)v_cache) % 16 == 0 && ((uintptr_t)q | (uintptr_t)o) % 4 == 0,
               "decode_attn: pointer alignment");
  int n_seg = 0;
  if (shared_prefix > 0) {
    // FP8 cache: the tensor-core prefix pass reads the shared tokens from a 16-bit staging copy (stage_table names
    // its rows for "request" *stage_index), the per-request pass reads the e4m3 cache
    B200_REQUIRE(!kv_fp8 || (k_stage && v_stage && stage_table && stage_index && stage_token_stride % 8 == 0 &&
                             ((uintptr_t)k_stage | (uintptr_t)v_stage) % 16 == 0),
                 "decode_attn_shared_prefix_fp8kv: the 16-bit staging copy of the shared prefix is required");
    B200_REQUIRE(seg_len >= 128 && seg_len % 128 == 0, "decode_attn_shared_prefix: seg_len %d must be a multiple of 128",
                 seg_len);
    n_seg = (shared_prefix + seg_len - 1) / seg_len;
    B200_REQUIRE(n_seg <= 65535 && (q_token_stride * 2) % 16 == 0 && (uintptr_t)q % 16 == 0,
                 "decode_attn_shared_prefix: %d segments / q alignment", n_seg);
  }
  const int ws_splits = n_seg + max_splits;
  size_t need = 0;
  b200_decode_attn_workspace_bytes(batch, num_q_heads, head_dim, ws_splits, &need);
  if (ws_splits > 1) {
    if (workspace == nullptr || workspace_bytes < need)
      return fail(B200_ERR_WORKSPACE, "decode_attn: workspace %zu B < required %zu B", workspace_bytes, need);
    B200_REQUIRE((uintptr_t)workspace % 16 == 0, "decode_attn: workspace must be 16-byte aligned");
  }
  DecodeArgs a;
  a.o = o; a.q = q; a.k_cache = k_cache; a.v_cache = v_cache;
  a.req_to_token = req_to_token; a.req_to_token_stride = req_to_token_stride;
  a.req_pool_indices = req_pool_indices; a.seq_lens = seq_lens; a.index_is_int64 = index_is_int64;
  a.batch = batch; a.num_q_heads = num_q_heads; a.num_kv_heads = num_kv_heads;
  a.group = num_q_heads / num_kv_heads;
  a.head_blocks = (a.group + 15) / 16;
  a.q_token_stride = q_token_stride; a.o_token_stride = o_token_stride; a.kv_token_stride = kv_token_stride;
  const float sm_scale_plain = sm_scale;  // for K values that already carry k_scale (the staged prefix)
  sm_scale *= k_scale;  // logits of the stored K/k_scale values, scaled back
  a.sm_scale = sm_scale; a.scale_log2 = sm_scale * kLog2e; a.logit_cap = logit_cap;
  a.out_scale = v_scale;
  a.kv_chunk = kv_chunk; a.max_splits = max_splits;
  a.kv_start = shared_prefix; a.n_seg = n_seg; a.ws_splits = ws_splits;
  const size_t n = (size_t)batch * num_q_heads * ws_splits;
  a.part_o = reinterpret_cast<float*>(workspace);
  a.part_lse = a.part_o ? a.part_o + n * head_dim : nullptr;
  a.oq = oq;
  a.oq_scale = oq_scale;
  B200_REQUIRE((oq == nullptr) == (oq_scale == nullptr), "decode_attn_int8: oq and oq_scale go together");
  B200_REQUIRE(oq == nullptr || ((uintptr_t)oq % 8 == 0 && (num_q_heads * head_dim) % 8 == 0 &&
                                 (size_t)num_q_heads * head_dim * 4 <= 200 * 1024),
               "decode_attn_int8: row of %d values not supported", num_q_heads * head_dim);
  cudaStream_t st = (cudaStream_t)stream;
  cudaEvent_t join = nullptr;
  if (n_seg > 0) {
    // prefix pass on the tensor cores: every request's q rows against the shared tokens, one segment per grid.z
    ExtendArgs e;
    e.o = nullptr; e.q = q; e.k_new = nullptr; e.v_new = nullptr; e.k_cache = k_cache; e.v_cache = v_cache;
    e.req_to_token = req_to_token; e.req_to_token_stride = req_to_token_stride;
    e.req_pool_indices = req_pool_indices; e.index_is_int64 = index_is_int64;
    e.qo_indptr = nullptr; e.prefix_lens = nullptr;
    e.batch = n_seg; e.num_q_heads = num_q_heads; e.num_kv_heads = num_kv_heads;
    e.q_token_stride = q_token_stride; e.o_token_stride = o_token_stride;
    e.knew_token_stride = 0; e.vnew_token_stride = 0; e.kv_token_stride = kv_token_stride;
    e.scale_log2 = a.scale_log2; e.sm_scale = a.sm_scale; e.logit_cap = logit_cap; e.causal = 0;
    if (kv_fp8) {  // staged prefix: up-cast and multiplied by k_scale / v_scale already
      e.k_cache = k_stage; e.v_cache = v_stage; e.req_to_token = stage_table; e.req_to_token_stride = 0;
      e.req_pool_indices = stage_index; e.kv_token_stride = stage_token_stride;
      e.sm_scale = sm_scale_plain; e.scale_log2 = sm_scale_plain * kLog2e;
    }
    e.part_o = a.part_o; e.part_lse = a.part_lse; e.part_stride = ws_splits; e.seg_len = seg_len;
    e.seg_rows = batch; e.seg_prefix = shared_prefix;
    // The prefix pass is tensor-bound (16 MB of HBM reads at cfg 5) and the per-request pass HBM-bound; they are
    // independent, so the prefix pass is forked onto a private stream and the merge joins both (plain event
    // record/wait: capturable in a CUDA graph).  Measured gain is small (0.1198 -> 0.1167 ms at cfg 5 per GPU): the
    // prefix grid fills the chip and the decode pass needs every SM to saturate HBM, so mostly the tails overlap.
    // B200_SHARED_PREFIX_OVERLAP=0: serial.
    // rows of the tensor the prefix pass reads K/V from (TMA extent): the staging copy, or the cache
    const int64_t prefix_rows = kv_fp8 ? (int64_t)shared_prefix : num_kv_slots;
    static const bool overlap = getenv("B200_SHARED_PREFIX_OVERLAP") == nullptr || getenv("B200_SHARED_PREFIX_OVERLAP")[0] != '0';
    if (overlap) {
      SideStream ss;
      int rc = side_stream_for_current_device(&ss);
      if (rc != B200_OK) return rc;
      B200_CUDA_OK(cudaEventRecord(ss.fork, st));
      B200_CUDA_OK(cudaStreamWaitEvent(ss.stream, ss.fork, 0));
      rc = launch_prefix_pass(e, batch, head_dim, dtype, prefix_rows, ss.stream);
      // join on every path after the fork: an un-joined side stream would break an enclosing graph capture
      B200_CUDA_OK(cudaEventRecord(ss.join, ss.stream));
      if (rc != B200_OK) {
        cudaStreamWaitEvent(st, ss.join, 0);
        return rc;
      }
      join = ss.join;
    } else {
      const int rc = launch_prefix_pass(e, batch, head_dim, dtype, prefix_rows, st);
      if (rc != B200_OK) return rc;
    }
  }
  if (dtype == B200_BF16) {
    if (head_dim == 128) return kv_fp8 ? launch_decode<__nv_bfloat16, 128, true>(a, st, join) : launch_decode<__nv_bfloat16, 128, false>(a, st, join);
    if (head_dim == 64) return kv_fp8 ? launch_decode<__nv_bfloat16, 64, true>(a, st, join) : launch_decode<__nv_bfloat16, 64, false>(a, st, join);
  } else if (dtype == B200_F16) {
    if (head_dim == 128) return kv_fp8 ? launch_decode<__half, 128, true>(a, st, join) : launch_decode<__half, 128, false>(a, st, join);
    if (head_dim == 64) return kv_fp8 ? launch_decode<__half, 64, true>(a, st, join) : launch_decode<__half, 64, false>(a, st, join);
  } else {
    return fail(B200_ERR_UNSUPPORTED, "decode_attn: dtype %d not supported (bf16/f16 only)", dtype);
  }
  return fail(B200_ERR_UNSUPPORTED, "decode_attn: head_dim %d not supported (64, 128)", head_dim);
}

B200_API int b200_decode_attn(void* o, const void* q, const void* k_cache, const void* v_cache,
                              const int32_t* req_to_token, int64_t req_to_token_stride,
                              const void* req_pool_indices, const void* seq_lens,
                              int index_is_int64, int batch, int num_q_heads, int num_kv_heads,
                              int head_dim, int64_t q_token_stride, int64_t o_token_stride,
                              int64_t kv_token_stride, float sm_scale, float logit_cap,
                              int kv_chunk, int max_splits, void* workspace,
                              size_t workspace_bytes, int dtype, void* stream) {
  return decode_attn_impl(o, q, k_cache, v_cache, req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                          index_is_int64, batch, num_q_heads, num_kv_heads, head_dim, q_token_stride, o_token_stride,
                          kv_token_stride, sm_scale, logit_cap, kv_chunk, max_splits, workspace, workspace_bytes, dtype,
                          nullptr, nullptr, stream);
}

B200_API int b200_decode_attn_int8(void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache,
                                   const void* v_cache, const int32_t* req_to_token, int64_t req_to_token_stride,
                                   const void* req_pool_indices, const void* seq_lens, int index_is_int64, int batch,
                                   int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
                                   int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap,
                                   int kv_chunk, int max_splits, void* workspace, size_t workspace_bytes, int dtype,
                                   void* stream) {
  B200_REQUIRE(oq != nullptr && oq_scale != nullptr, "decode_attn_int8: null quant output");
  return decode_attn_impl(o, q, k_cache, v_cache, req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                          index_is_int64, batch, num_q_heads, num_kv_heads, head_dim, q_token_stride, o_token_stride,
                          kv_token_stride, sm_scale, logit_cap, kv_chunk, max_splits, workspace, workspace_bytes, dtype,
                          oq, oq_scale, stream);
}

B200_API int b200_decode_attn_fp8kv(void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache,
                                    const void* v_cache, const int32_t* req_to_token, int64_t req_to_token_stride,
                                    const void* req_pool_indices, const void* seq_lens, int index_is_int64, int batch,
                                    int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
                                    int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap,
                                    float k_scale, float v_scale, int kv_chunk, int max_splits, void* workspace,
                                    size_t workspace_bytes, int dtype, void* stream) {
  return decode_attn_impl(o, q, k_cache, v_cache, req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                          index_is_int64, batch, num_q_heads, num_kv_heads, head_dim, q_token_stride, o_token_stride,
                          kv_token_stride, sm_scale, logit_cap, kv_chunk, max_splits, workspace, workspace_bytes, dtype,
                          oq, oq_scale, stream, true, k_scale, v_scale);
}

B200_API int b200_decode_attn_shared_prefix(void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache,
                                            const void* v_cache, const int32_t* req_to_token,
                                            int64_t req_to_token_stride, const void* req_pool_indices,
                                            const void* seq_lens, int index_is_int64, int batch, int num_q_heads,
                                            int num_kv_heads, int head_dim, int64_t q_token_stride,
                                            int64_t o_token_stride, int64_t kv_token_stride, float sm_scale,
                                            float logit_cap, int shared_prefix_len, int seg_len, int kv_chunk,
                                            int max_splits, void* workspace, size_t workspace_bytes, int dtype,
                                            void* stream, int64_t num_kv_slots) {
  B200_REQUIRE(shared_prefix_len > 0, "decode_attn_shared_prefix: shared_prefix_len must be positive");
  return decode_attn_impl(o, q, k_cache, v_cache, req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                          index_is_int64, batch, num_q_heads, num_kv_heads, head_dim, q_token_stride, o_token_stride,
                          kv_token_stride, sm_scale, logit_cap, kv_chunk, max_splits, workspace, workspace_bytes, dtype,
                          oq, oq_scale, stream, false, 1.f, 1.f, shared_prefix_len, seg_len, nullptr, nullptr, nullptr,
                          nullptr, 0, num_kv_slots);
}

B200_API int b200_decode_attn_shared_prefix_fp8kv(
    void* o, int8_t* oq, float* oq_scale, const void* q, const void* k_cache, const void* v_cache, const void* k_stage,
    const void* v_stage, const int32_t* stage_table, const void* stage_index, int64_t stage_token_stride,
    const int32_t* req_to_token, int64_t req_to_token_stride, const void* req_pool_indices, const void* seq_lens,
    int index_is_int64, int batch, int num_q_heads, int num_kv_heads, int head_dim, int64_t q_token_stride,
    int64_t o_token_stride, int64_t kv_token_stride, float sm_scale, float logit_cap, float k_scale, float v_scale,
    int shared_prefix_len, int seg_len, int kv_chunk, int max_splits, void* workspace, size_t workspace_bytes, int dtype,
    void* stream) {
  B200_REQUIRE(shared_prefix_len > 0, "decode_attn_shared_prefix_fp8kv: shared_prefix_len must be positive");
  return decode_attn_impl(o, q, k_cache, v_cache, req_to_token, req_to_token_stride, req_pool_indices, seq_lens,
                          index_is_int64, batch, num_q_heads, num_kv_heads, head_dim, q_token_stride, o_token_stride,
                          kv_token_stride, sm_scale, logit_cap, kv_chunk, max_splits, workspace, workspace_bytes, dtype,
                          oq, oq_scale, stream, true, k_scale, v_scale, shared_prefix_len, seg_len, k_stage, v_stage,
                          stage_table, stage_index, stage_token_stride);
}
