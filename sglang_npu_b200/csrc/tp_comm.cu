// Tensor-parallel exchange step of the row-parallel linears, fused with what follows it.
//
// Reference path being replaced (python/sglang/srt/): RowParallelLinear.forward -> tensor_model_parallel_all_reduce
// (layers/linear.py:1303-1304, distributed/communication_op.py:11-13: NCCL / custom one-shot all-reduce,
// sgl-kernel/csrc/allreduce/custom_all_reduce.cuh) followed by fused_add_rmsnorm and the per-token int8 quant of the
// next linear; the reference's own fusion of the first two is layers/flashinfer_comm_fusion.py (allreduce + rmsnorm).
//
// Here it is ONE kernel over NVLink peer memory, and a two-shot one: every rank's row-parallel GEMM writes its partial
// [M, H] output into a symmetric buffer (cudaIpc-mapped into every peer).  Rank r owns the token slice
// [r * ceil(M / tp), ...): its CTAs (one per token row)
//   1. wait until every peer's partial is complete (flag barrier, sys-scope release/acquire),
//   2. read that row from all tp partial buffers with 16-byte peer loads and add them in rank order in fp32
//      (one rounding to T: for tp = 2 this is bit-identical to NCCL's bf16 sum),
//   3. residual += sum (kept only on the owning rank: the residual stream is token-sharded between exchanges),
//      RMSNorm, per-token int8 quantisation — the arithmetic of rmsnorm_quant_int8_kernel (fused_steps.cu), bit for bit,
//   4. store the int8 row + its scale (and/or the normalised T row) straight into EVERY rank's activation buffer
//      (the all-gather half, at 1 byte per element instead of 2), then signal the peers and wait for theirs.
// Per exchange a rank moves (tp-1)/tp * M * H * (2 + 1) bytes over NVLink instead of 2 * (tp-1)/tp * M * H * 2 for a
// ring all-reduce of bf16, launches one kernel instead of three, and normalises M / tp rows instead of M.
#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace b200 {

constexpr int kMaxTp = 8;
constexpr int kCommThreads = 512;

struct TpArgs {
  const void* partial[kMaxTp];  // [M, H] T, one per rank (own included), rank order
  int8_t* xq[kMaxTp];           // [M, H] int8 destination on every rank (or null)
  float* xscale[kMaxTp];        // [M]
  void* y[kMaxTp];              // [M, H] T destination on every rank (or null)
  uint32_t* flags[kMaxTp];      // per rank: [2][kMaxTp][max_blocks] epochs (start, end)
  uint32_t* epoch;              // local: [max_blocks]
  void* residual;               // local [M, H] T or null
  const void* weight;           // [H] T
  int64_t M, H;
  int tp, rank, max_blocks;
  float eps;
  uint32_t* err;                  // local: first barrier failure of this rank (0 = none), read by b200_comm_check
  unsigned long long timeout_ns;  // how long a barrier waits for a peer before giving up (%globaltimer)
  int relaxed_start;              // start barrier with relaxed flags (default; 0 = sys-scope release / acquire, A/B)
  unsigned long long* trace;  // diagnostics (b200_debug_set_trace) or null
  int trace_slots;
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// Peer (and local) partial rows are read exactly once per launch, after the start barrier.  ld.volatile is the
// load that is guaranteed not to be served from this SM's L1 (peer lines may live there across launches — an
// ld.global.cg version read stale rows); as a volatile asm it also stays after the barrier, while consecutive
// loads still issue back to back: all ranks' rows are in flight together, one NVLink round trip per row.
__device__ __forceinline__ int4 ld_peer_v4(const void* p) {
  int4 r;
  asm volatile("ld.volatile.global.v4.s32 {%0,%1,%2,%3}, [%4];\n" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.relaxed.sys.global.u32 [%0], %1;\n" ::"l"(p), "r"(v) : "memory");
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t));
  return t;
}

// Block-level barrier across ranks: thread t < tp publishes `e` into peer t's slot [phase][rank][b] and waits for
// peer t's value in the local slot [phase][t][b].
//   END barrier (phase 1): a PTX release / acquire pair at system scope.  ONE releasing thread per peer: the CTA barrier
//   before it orders every thread's earlier accesses — the stores into the peers' buffers — before the release
//   (cumulativity), so the other ~500 threads do not fence.  The release costs 3-4 us (it waits for the NVLink acks of
//   those stores); that wait is the point of the barrier.
//   START barrier (phase 0): what has to be visible to the peers is this rank's GEMM output, written by the PREVIOUS
//   grid.  Default: relaxed sys-scope flags, the reference's own protocol (custom_all_reduce.cuh:168-171,
//   "Start barrier shouldn't need fence": st/ld.volatile).  It rests on three facts, each held by a test or a rule of
//   the programming model: (1) griddepcontrol.wait / the kernel boundary returns only when all memory operations of the
//   prerequisite grid are performed and visible to this grid, i.e. they sit in THIS GPU's L2 or HBM; (2) memory homed on
//   a GPU has one point of coherence, that GPU's L2 — a peer's NVLink load is served there and is never cached in the
//   peer's L2 — so a load that leaves the peer AFTER it saw the flag sees the data; (3) the peer's loads do leave the
//   peer: they are ld.volatile (L1 bypass, see ld_peer_v4) and sit behind the flag loop in program order as volatile
//   asm.  A sys-scope release here instead (B200_TP_RELEASE_START=1, kept for A/B runs) is a MEMBAR.SYS with nothing of
//   this grid to wait for and still costs 4.1 us per exchange (flag sent 4.77 us after kernel entry instead of 0.62 us:
//   profiles/r02zb_tp_exchange_trace_n2.jsonl) — 8 us per layer.  tools/tp_exchange_check.py holds the exchange bit-exact
//   against the rank-order sum of the gathered partials over repeated launches and graph replays at tp 2 / 4 / 8.
//   A peer that does not arrive within `timeout_ns` (wall clock, %globaltimer; default 20 s — host skew, a lazy module
//   load or a GC pause on a peer are legitimate reasons to be late) does NOT kill the context: the first failure is
//   recorded in `err` (b200_comm_check reports it on the host), the wait is abandoned and the kernel completes with
//   whatever data it has, so the process can tear down in an orderly way like it would after an NCCL timeout.
__device__ __forceinline__ void rank_barrier(const TpArgs& a, int phase, int b, uint32_t e) {
  __syncthreads();
  if ((int)threadIdx.x < a.tp) {
    const int t = threadIdx.x;
    const size_t slot = ((size_t)phase * kMaxTp) * a.max_blocks + b;
    uint32_t* theirs = a.flags[t] + slot + (size_t)a.rank * a.max_blocks;
    const uint32_t* mine = a.flags[a.rank] + slot + (size_t)t * a.max_blocks;
    const bool relaxed = phase == 0 && a.relaxed_start != 0;
    if (relaxed) st_relaxed_sys(theirs, e);
    else st_release_sys(theirs, e);
    if (t == 0 && a.trace != nullptr && b < a.trace_slots) trace_mark(a.trace + (size_t)b * 16, 7 + 2 * phase);  // flag sent
    unsigned long long t0 = 0;
    uint32_t spins = 0;
    while (ld_acquire_sys(mine) != e) {  // acquire on the waiting side in both modes: it measured free (1.5 us either way)
      if ((++spins & 0x3ffu) == 0) {  // look at the clock every 1024 polls
        const unsigned long long now = globaltimer_ns();
        if (t0 == 0) t0 = now;
        else if (now - t0 > a.timeout_ns) {
          // code: 1 | phase << 1 | peer << 4 | block << 8  (first failure wins)
          atomicCAS(a.err, 0u, 1u | ((uint32_t)phase << 1) | ((uint32_t)t << 4) | ((uint32_t)b << 8));
          break;
        }
      }
    }
  }
  __syncthreads();
}

template <typename T>
__device__ __forceinline__ void unpack8(const int4& v, float (&f)[8]) {
  const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = unpack2<T>(u[i]);
    f[2 * i] = x.x;
    f[2 * i + 1] = x.y;
  }
}

template <typename T>
__device__ __forceinline__ void tp_reduce_norm_quant_body(const TpArgs& a, float* row_f, float* red, float* red2) {
  constexpr int kWarps = kCommThreads / 32;
  const int b = blockIdx.x;
  unsigned long long* const tr = (a.trace != nullptr && b < a.trace_slots) ? a.trace + (size_t)b * 16 : nullptr;
  if (threadIdx.x == 0) {
    trace_mark(tr, 0);
    trace_mark(tr, 14);
  }
  pdl_trigger();
  pdl_wait();  // the local GEMM has written this rank's partial
  if (threadIdx.x == 0) trace_mark(tr, 1);
  const int64_t rows_per = (a.M + a.tp - 1) / a.tp;
  const uint32_t e = a.epoch[b] + 1u;
  rank_barrier(a, 0, b, e);  // every peer's GEMM output is complete and visible
  if (threadIdx.x == 0) trace_mark(tr, 2);
  const int64_t H = a.H;
  const int tp = a.tp;
  const bool plain = a.weight == nullptr;  // b200_tp_all_reduce: the summed rows go out as they are (no norm, no quant)
  // the grid is capped (every CTA of every rank must be resident for the barriers): a CTA walks rows b, b + grid, ...
  for (int64_t ri = b; ri < rows_per; ri += gridDim.x) {
    const int64_t row = (int64_t)a.rank * rows_per + ri;
    if (row >= a.M) break;
    __syncthreads();  // row_f / red are reused across rows
    T* rr = a.residual ? reinterpret_cast<T*>(a.residual) + row * H : nullptr;
    float ss = 0.f;
    for (int64_t c = threadIdx.x * 8; c < H; c += kCommThreads * 8) {
      int4 pv[kMaxTp];
#pragma unroll
      for (int r = 0; r < kMaxTp; ++r)
        if (r < tp) pv[r] = ld_peer_v4(reinterpret_cast<const T*>(a.partial[r]) + row * H + c);
      int4 rv = make_int4(0, 0, 0, 0);
      if (rr) rv = *reinterpret_cast<const int4*>(rr + c);
      float s[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) s[i] = 0.f;
#pragma unroll
      for (int r = 0; r < kMaxTp; ++r) {  // fixed rank order: deterministic, identical on every rank
        if (r < tp) {
          float f[8];
          unpack8<T>(pv[r], f);
#pragma unroll
          for (int i = 0; i < 8; ++i) s[i] += f[i];
        }
      }
      if (plain) {  // two-shot all-reduce: this rank's token slice, summed in rank order, to every rank
        int4 sv;
        uint32_t* su = reinterpret_cast<uint32_t*>(&sv);
#pragma unroll
        for (int i = 0; i < 4; ++i) su[i] = pack2<T>(s[2 * i], s[2 * i + 1]);
        for (int r = 0; r < tp; ++r) *reinterpret_cast<int4*>(reinterpret_cast<T*>(a.y[r]) + row * H + c) = sv;
        continue;
      }
      float f[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {  // the all-reduce result as a T tensor (what the un-fused path hands to the norm)
        const float2 x = unpack2<T>(pack2<T>(s[2 * i], s[2 * i + 1]));
        f[2 * i] = x.x;
        f[2 * i + 1] = x.y;
      }
      if (rr) {
        float r8[8];
        unpack8<T>(rv, r8);
        int4 nv;
        uint32_t* nu = reinterpret_cast<uint32_t*>(&nv);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          f[2 * i] += r8[2 * i];
          f[2 * i + 1] += r8[2 * i + 1];
          nu[i] = pack2<T>(f[2 * i], f[2 * i + 1]);
        }
        *reinterpret_cast<int4*>(rr + c) = nv;  // residual = T(x + residual), owner rank only
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) row_f[c + i] = f[i];
    }
    if (plain) continue;  // (uniform over the CTA)
    __syncthreads();
    // Sum of squares in EXACTLY the order of rmsnorm_quant_int8_kernel (256 threads, thread j runs over the 8-column
    // groups j, j + 256, ... sequentially, then a warp tree, then 8 warp sums in order): fp32 addition is not
    // associative and the fused exchange must give the bits of the un-fused path.
    if (threadIdx.x < 256) {
      for (int64_t c = threadIdx.x * 8; c < H; c += 256 * 8) {
#pragma unroll
        for (int i = 0; i < 8; ++i) ss += row_f[c + i] * row_f[c + i];
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
      if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
    }
    __syncthreads();
    if (threadIdx.x == 0 && ri == b) trace_mark(tr, 3);
    float tot = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) tot += red[i];
    const float rstd = rsqrtf(tot / (float)H + a.eps);
    float amax = 0.f;
    const T* w = reinterpret_cast<const T*>(a.weight);
    for (int64_t c = threadIdx.x * 8; c < H; c += kCommThreads * 8) {
      float wv[8];
      unpack8<T>(*reinterpret_cast<const int4*>(w + c), wv);
      int4 ov;
      uint32_t* ou = reinterpret_cast<uint32_t*>(&ov);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        ou[i] = pack2<T>(row_f[c + 2 * i] * rstd * wv[2 * i], row_f[c + 2 * i + 1] * rstd * wv[2 * i + 1]);
        const float2 y = unpack2<T>(ou[i]);
        row_f[c + 2 * i] = y.x;
        row_f[c + 2 * i + 1] = y.y;
        amax = fmaxf(amax, fmaxf(fabsf(y.x), fabsf(y.y)));
      }
      if (a.y[0] != nullptr) {
        for (int r = 0; r < tp; ++r) *reinterpret_cast<int4*>(reinterpret_cast<T*>(a.y[r]) + row * H + c) = ov;
      }
    }
    if (a.xq[0] != nullptr) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
      if ((threadIdx.x & 31) == 0) red2[threadIdx.x >> 5] = amax;
      __syncthreads();
      amax = red2[0];
#pragma unroll
      for (int i = 1; i < kWarps; ++i) amax = fmaxf(amax, red2[i]);
      amax = fmaxf(amax, 1e-10f);
      const float inv = __fmul_rn(__frcp_rn(amax), 127.f);
      if ((int)threadIdx.x < tp) a.xscale[threadIdx.x][row] = __fdiv_rn(amax, 127.f);
      for (int64_t c = threadIdx.x * 8; c < H; c += kCommThreads * 8) {
        uint32_t lo = 0, hi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int x0 = __float2int_rn(__fmul_rn(row_f[c + i], inv));
          const int x1 = __float2int_rn(__fmul_rn(row_f[c + 4 + i], inv));
          lo |= (uint32_t)(x0 & 0xff) << (8 * i);
          hi |= (uint32_t)(x1 & 0xff) << (8 * i);
        }
        const uint2 q = make_uint2(lo, hi);
        for (int r = 0; r < tp; ++r) *reinterpret_cast<uint2*>(a.xq[r] + row * H + c) = q;
      }
    }
  }
  if (threadIdx.x == 0) trace_mark(tr, 4);
  rank_barrier(a, 1, b, e);  // every peer has delivered its rows of block b; the grid's exit covers all blocks
  if (threadIdx.x == 0) {
    a.epoch[b] = e;
    trace_mark(tr, 6);
    trace_mark(tr, 15);
  }
}

template <typename T>
__global__ void __launch_bounds__(kCommThreads) tp_reduce_norm_quant_kernel(const TpArgs a) {
  extern __shared__ float row_f[];  // H floats
  __shared__ float red[kCommThreads / 32];
  __shared__ float red2[kCommThreads / 32];
  tp_reduce_norm_quant_body<T>(a, row_f, red, red2);
}

// Single-GPU emulation for the test suite: ALL ranks of the exchange in ONE cooperative launch, blockIdx.y = rank, every
// rank with its own argument block and buffer set on the same device.  Ranks wait on one another through the same flag
// protocol as over NVLink; the cooperative launch guarantees that all of them are resident (separate launches that spin
// on one another are not guaranteed to run together on one GPU).
template <typename T>
__global__ void __launch_bounds__(kCommThreads) tp_reduce_norm_quant_loopback_kernel(const TpArgs* __restrict__ rank_args) {
  extern __shared__ float row_f[];
  __shared__ float red[kCommThreads / 32];
  __shared__ float red2[kCommThreads / 32];
  __shared__ TpArgs a;
  if (threadIdx.x == 0) a = rank_args[blockIdx.y];
  __syncthreads();
  tp_reduce_norm_quant_body<T>(a, row_f, red, red2);
}

}  // namespace b200

using namespace b200;

B200_API int b200_comm_malloc(void** dev_ptr, size_t bytes) {
  B200_REQUIRE(dev_ptr != nullptr && bytes > 0, "comm_malloc: bad arguments");
  B200_CUDA_OK(cudaMalloc(dev_ptr, bytes));
  B200_CUDA_OK(cudaMemset(*dev_ptr, 0, bytes));
  B200_CUDA_OK(cudaDeviceSynchronize());
  return B200_OK;
}

B200_API int b200_comm_free(void* dev_ptr) {
  if (dev_ptr) B200_CUDA_OK(cudaFree(dev_ptr));
  return B200_OK;
}

B200_API int b200_comm_ipc_handle(void* dev_ptr, unsigned char* handle64) {
  B200_REQUIRE(dev_ptr && handle64, "comm_ipc_handle: null pointer");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  cudaIpcMemHandle_t h;
  B200_CUDA_OK(cudaIpcGetMemHandle(&h, dev_ptr));
  memcpy(handle64, &h, 64);
  return B200_OK;
}

B200_API int b200_comm_ipc_open(const unsigned char* handle64, void** peer_ptr) {
  B200_REQUIRE(handle64 && peer_ptr, "comm_ipc_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  B200_CUDA_OK(cudaIpcOpenMemHandle(peer_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}

B200_API int b200_comm_ipc_close(void* peer_ptr) {
  if (peer_ptr) B200_CUDA_OK(cudaIpcCloseMemHandle(peer_ptr));
  return B200_OK;
}

B200_API int b200_comm_check(const void* err_flag, void* stream) {
  // Synchronises `stream` and reports a barrier timeout recorded by an exchange kernel of this rank.
  B200_REQUIRE(err_flag != nullptr, "comm_check: null pointer");
  uint32_t code = 0;
  B200_CUDA_OK(cudaMemcpyAsync(&code, err_flag, sizeof(code), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  B200_CUDA_OK(cudaStreamSynchronize((cudaStream_t)stream));
  if (code != 0)
    return fail(B200_ERR_CUDA, "tp exchange: %s barrier of block %u timed out waiting for rank %u (B200_TP_TIMEOUT_MS)",
                ((code >> 1) & 1u) ? "end" : "start", code >> 8, (code >> 4) & 0xfu);
  return B200_OK;
}

static int build_tp_args(TpArgs& a, int64_t& grid, const void* const* partial_ptrs, void* const* xq_ptrs,
                         void* const* xscale_ptrs, void* const* y_ptrs, void* const* flag_ptrs, void* epoch, void* residual,
                         const void* weight, int64_t M, int64_t H, float eps, int tp_size, int tp_rank, int max_blocks,
                         void* err_flag) {
  B200_REQUIRE(tp_size >= 2 && tp_size <= kMaxTp && tp_rank >= 0 && tp_rank < tp_size, "tp exchange: bad tp %d / rank %d",
               tp_size, tp_rank);
  B200_REQUIRE(M > 0 && H > 0 && H % 8 == 0 && H * 4 <= 200 * 1024, "tp exchange: bad sizes M %lld H %lld", (long long)M,
               (long long)H);
  B200_REQUIRE(partial_ptrs && flag_ptrs && epoch && err_flag, "tp exchange: null pointer");
  // weight == null is the plain all-reduce (b200_tp_all_reduce): 16-bit output only, nothing to add or normalise
  B200_REQUIRE(weight != nullptr || (y_ptrs != nullptr && xq_ptrs == nullptr && residual == nullptr),
               "tp exchange: a norm weight is required unless only the summed rows are requested");
  B200_REQUIRE(xq_ptrs != nullptr || y_ptrs != nullptr, "tp exchange: no output requested");
  B200_REQUIRE((xq_ptrs == nullptr) == (xscale_ptrs == nullptr), "tp exchange: xq and xscale go together");
  const int64_t rows_per = ceil_div(M, tp_size);
  B200_REQUIRE(max_blocks >= 1, "tp exchange: no flag slots");
  grid = sm_count();  // all CTAs co-resident on every rank (one 512-thread CTA per SM)
  if (grid > max_blocks) grid = max_blocks;
  if (grid > rows_per) grid = rows_per;
  a = TpArgs{};
  for (int r = 0; r < tp_size; ++r) {
    B200_REQUIRE(partial_ptrs[r] && flag_ptrs[r], "tp exchange: null peer pointer");
    a.partial[r] = partial_ptrs[r];
    a.flags[r] = reinterpret_cast<uint32_t*>(flag_ptrs[r]);
    if (xq_ptrs) {
      B200_REQUIRE(xq_ptrs[r] && xscale_ptrs[r], "tp exchange: null peer pointer");
      a.xq[r] = reinterpret_cast<int8_t*>(xq_ptrs[r]);
      a.xscale[r] = reinterpret_cast<float*>(xscale_ptrs[r]);
    }
    if (y_ptrs) {
      B200_REQUIRE(y_ptrs[r], "tp exchange: null peer pointer");
      a.y[r] = y_ptrs[r];
    }
  }
  a.epoch = reinterpret_cast<uint32_t*>(epoch);
  a.residual = residual;
  a.weight = weight;
  a.trace = g_trace; a.trace_slots = (int)g_trace_slots;
  a.err = reinterpret_cast<uint32_t*>(err_flag);
  static const unsigned long long timeout_ms =
      getenv("B200_TP_TIMEOUT_MS") ? strtoull(getenv("B200_TP_TIMEOUT_MS"), nullptr, 10) : 20000ull;
  // start barrier: relaxed flags unless B200_TP_RELEASE_START=1 (or the older spelling B200_TP_RELAXED_START=0)
  static const int relaxed_start = [] {
    const char* rel = getenv("B200_TP_RELEASE_START");
    const char* old = getenv("B200_TP_RELAXED_START");
    if (rel && rel[0] == '1') return 0;
    if (old && old[0] == '0') return 0;
    return 1;
  }();
  a.timeout_ns = timeout_ms * 1000000ull;
  a.relaxed_start = relaxed_start;
  a.M = M; a.H = H; a.tp = tp_size; a.rank = tp_rank; a.max_blocks = max_blocks; a.eps = eps;
  return B200_OK;
}

B200_API int b200_tp_reduce_rmsnorm_quant_int8(const void* const* partial_ptrs, void* const* xq_ptrs,
                                               void* const* xscale_ptrs, void* const* y_ptrs, void* const* flag_ptrs,
                                               void* epoch, void* residual, const void* weight, int64_t M, int64_t H,
                                               float eps, int tp_size, int tp_rank, int max_blocks, int dtype,
                                               void* err_flag, void* stream) {
  if (M == 0) return B200_OK;
  TpArgs a;
  int64_t grid = 0;
  int rc = build_tp_args(a, grid, partial_ptrs, xq_ptrs, xscale_ptrs, y_ptrs, flag_ptrs, epoch, residual, weight, M, H, eps,
                         tp_size, tp_rank, max_blocks, err_flag);
  if (rc != B200_OK) return rc;
  const int smem = (int)(H * sizeof(float));
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_BF16) {
    auto kern = tp_reduce_norm_quant_kernel<__nv_bfloat16>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return launch_k(kern, dim3((unsigned)grid), dim3(kCommThreads), smem, st, 1, true, a);
  }
  if (dtype == B200_F16) {
    auto kern = tp_reduce_norm_quant_kernel<__half>;
    B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    return launch_k(kern, dim3((unsigned)grid), dim3(kCommThreads), smem, st, 1, true, a);
  }
  return fail(B200_ERR_UNSUPPORTED, "tp exchange: dtype %d not supported (bf16/f16)", dtype);
}

B200_API int b200_tp_all_reduce(const void* const* partial_ptrs, void* const* out_ptrs, void* const* flag_ptrs, void* epoch,
                                int64_t M, int64_t H, int tp_size, int tp_rank, int max_blocks, int dtype, void* err_flag,
                                void* stream) {
  // sum over ranks of the [M, H] partials, every rank receiving all rows: the exchange kernel without its norm / quant
  return b200_tp_reduce_rmsnorm_quant_int8(partial_ptrs, nullptr, nullptr, out_ptrs, flag_ptrs, epoch, nullptr, nullptr, M, H,
                                           0.f, tp_size, tp_rank, max_blocks, dtype, err_flag, stream);
}

B200_API int b200_tp_exchange_loopback(const void* const* partial_ptrs, void* const* xq_ptrs, void* const* xscale_ptrs,
                                       void* const* y_ptrs, void* const* flag_ptrs, void* const* epoch_ptrs,
                                       void* const* residual_ptrs, const void* weight, int64_t M, int64_t H, float eps,
                                       int tp_size, int max_blocks, int dtype, void* const* err_ptrs, void* args_scratch,
                                       size_t args_scratch_bytes, void* stream) {
  // Test entry: every rank of one exchange in a single cooperative launch on ONE device (see the loopback kernel).
  // args_scratch: device memory for tp_size argument blocks (b200_tp_exchange_loopback needs tp_size * 512 bytes).
  if (M == 0) return B200_OK;
  B200_REQUIRE(epoch_ptrs && err_ptrs && args_scratch, "tp loopback: null pointer");
  B200_REQUIRE(args_scratch_bytes >= (size_t)tp_size * sizeof(TpArgs), "tp loopback: argument scratch of %zu bytes, %zu needed",
               args_scratch_bytes, (size_t)tp_size * sizeof(TpArgs));
  TpArgs host_args[kMaxTp];
  int64_t grid = 0;
  for (int r = 0; r < tp_size; ++r) {
    int rc = build_tp_args(host_args[r], grid, partial_ptrs, xq_ptrs, xscale_ptrs, y_ptrs, flag_ptrs, epoch_ptrs[r],
                           residual_ptrs ? residual_ptrs[r] : nullptr, weight, M, H, eps, tp_size, r, max_blocks, err_ptrs[r]);
    if (rc != B200_OK) return rc;
  }
  cudaStream_t st = (cudaStream_t)stream;
  B200_CUDA_OK(cudaMemcpyAsync(args_scratch, host_args, (size_t)tp_size * sizeof(TpArgs), cudaMemcpyHostToDevice, st));
  const int smem = (int)(H * sizeof(float));
  const void* kern = dtype == B200_BF16 ? (const void*)tp_reduce_norm_quant_loopback_kernel<__nv_bfloat16>
                     : dtype == B200_F16 ? (const void*)tp_reduce_norm_quant_loopback_kernel<__half> : nullptr;
  if (kern == nullptr) return fail(B200_ERR_UNSUPPORTED, "tp loopback: dtype %d not supported (bf16/f16)", dtype);
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  int per_sm = 0;
  B200_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kCommThreads, smem));
  B200_REQUIRE(grid * tp_size <= (int64_t)per_sm * sm_count(), "tp loopback: %lld CTAs cannot be co-resident",
               (long long)(grid * tp_size));
  const TpArgs* dev_args = reinterpret_cast<const TpArgs*>(args_scratch);
  void* params[] = {(void*)&dev_args};
  B200_CUDA_OK(cudaLaunchCooperativeKernel(kern, dim3((unsigned)grid, (unsigned)tp_size), dim3(kCommThreads), params, smem, st));
  return B200_OK;
}
