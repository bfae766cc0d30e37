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
// The fused GEMM is described at awq_streamk_kernel below.
#include <cstdlib>

#include "gemm_common.cuh"

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
// {magic + nibble p, magic + nibble p+4} as a packed pair: one shift and ONE lop3 ((a & b) | c, LUT 0xEA)
template <typename T>
__device__ __forceinline__ uint32_t nib_pair(uint32_t word, int p) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;\n" : "=r"(r) : "r"(word >> (4 * p)), "r"(0x000f000fu), "r"(AwqMagic<T>::kOr));
  return r;
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

// ---- checkpoint layout -> streaming layout (process_weights_after_loading) ------------------------------------------
// qweight [K, N/8] (8 channels of one k per word) -> words that hold 8 consecutive k of ONE channel, in the same nibble
// order (k offset e sits in nibble ORDER[e] = (e >> 1) + 4 * (e & 1), so `nib_pair(word, p)` is the pair (k = 2p, 2p + 1)
// — one 32-bit TMEM cell of the MMA's A operand), stored UNIT-MAJOR: qweight_t [ceil(N/256) * ceil(K/128)][256][16] holds,
// per (256-channel tile, 128-k block), the exact 16 KB shared-memory image the GEMM wants (rows = channels, 64 B = 128 k
// per row, 16-byte chunks swizzled for conflict-free 16-byte reads; zero padding past N and K).  The stream-K ranges are
// contiguous in (tile, k block) order, so every CTA reads ONE contiguous byte range of the weights with 16 KB bulk copies;  scales [K/g, N] + qzeros [K/g, N/8] -> sz [K/g, N]
// uint32 = {scale (low half), magic + zero point (high half)}: one coalesced 4-byte load per (channel, group).
__device__ __forceinline__ int awq_order(int e) { return (e >> 1) + 4 * (e & 1); }

__global__ void __launch_bounds__(256) awq_repack_weight_kernel(uint32_t* __restrict__ qweight_t,
                                                                const uint32_t* __restrict__ qweight, int64_t K, int64_t N,
                                                                int64_t num_kb) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // one output word: (channel n, word j of its k run)
  const int64_t k8 = K / 8;
  if (idx >= N * k8) return;
  const int64_t n = idx / k8, j = idx - n * k8;
  const int sh = 4 * awq_order((int)(n & 7));
  uint32_t word = 0;
#pragma unroll
  for (int e = 0; e < 8; ++e) {
    const uint32_t q = (qweight[(j * 8 + e) * (N / 8) + (n >> 3)] >> sh) & 0xFu;
    word |= q << (4 * awq_order(e));
  }
  // unit-major image: unit (tile, kb) = 256 channels x 16 words, the 16-byte chunk index XOR-ed with (row >> 1) & 3
  const int64_t tile = n >> 8, kb = j >> 4;
  const int row = (int)(n & 255), jw = (int)(j & 15);
  const int chunk = (jw >> 2) ^ ((row >> 1) & 3);
  qweight_t[((tile * num_kb + kb) * 256 + row) * 16 + chunk * 4 + (jw & 3)] = word;
}

__global__ void __launch_bounds__(256) awq_repack_sz_kernel(uint32_t* __restrict__ sz, const uint16_t* __restrict__ scales,
                                                            const uint32_t* __restrict__ qzeros, int64_t groups, int64_t N,
                                                            uint32_t magic16) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // (group, n)
  if (idx >= groups * N) return;
  const int64_t gi = idx / N, n = idx - gi * N;
  const uint32_t z = (qzeros[gi * (N / 8) + (n >> 3)] >> (4 * awq_order((int)(n & 7)))) & 0xFu;
  sz[idx] = (uint32_t)scales[idx] | ((magic16 | z) << 16);  // 1024 + z (f16) / 128 + z (bf16): exact
}

// ---- fused dequant + GEMM: persistent stream-K weight streaming with the A operand in tensor memory ----------------------
// D[channel, token] (TMEM, fp32) += W^T[channel, k] * x[token, k]^T, weights on MMA-M, tokens on MMA-N (<= 128 per launch).
//   * work unit = 256 channels x 128 k: 16 KB of packed words — one contiguous bulk copy of the unit-major image that
//     awq_repack lays out (a [256 rows x 64 B] box of a row-major matrix touched half a cache line per row) — and two x chunks
//     [BN tokens x 64 k] (128-byte-swizzled K-major boxes).  One x chunk feeds TWO M=128 MMAs (two channel groups), which
//     halves the L2 -> SM token traffic per weight byte (at 128 tokens an x tile is 4x the bytes of a 128-channel box);
//   * the packed words and the x chunks ride SEPARATE rings: the words — the only bytes that come from HBM — get a deep
//     ring (8-10 units = 128-160 KB in flight per SM, freed as soon as the dequant warps hold them in registers), the x
//     chunks — L2 hits — a shallow one.  (One ring of [words | x] stages kept only 64 KB of HBM loads in flight: 2.7 TB/s
//     with the compute switched off.)
//   * CTA c of a one-CTA-per-SM grid streams the contiguous unit range [c*U/G, (c+1)*U/G) (tile-major, K inner) — every
//     SM pulls the same number of weight bytes whatever N and K are; ranges are cut into segments at tile boundaries;
//   * 8 dequant warps: warp -> (channel group c = 0/1, TMEM lane quadrant q), thread -> ONE channel.  A thread reads its
//     channel's 16 words (64 B, conflict-free through the swizzle), turns every word into four {k, k+1} cells with the
//     magic-exponent trick (shift, lop3, one packed subtract of {magic + z}, one packed multiply by {s, s} — the same
//     arithmetic as awq_dequant_kernel, so the weights are bit-identical to the reference's fp16 dequantisation) and
//     stores them with tcgen05.st into a ring of four 64-column TMEM buffers: the dequantised matrix NEVER passes
//     through shared memory (the round-1 kernel paid 16 KB of st.shared + 16 KB of tensor-core reads per 4 KB of weights);
//   * one thread issues tcgen05.mma kind::f16 with A from TMEM, B = x chunk from shared memory;
//   * the dequant warps double as the epilogue of their (group, quadrant).  A segment that covers a whole tile goes
//     (+bias) -> fp16/bf16 -> out[m, n] straight from TMEM.  A partial segment PUBLISHES its raw fp32 accumulators to the
//     workspace and moves on; when a CTA has streamed all its units it helps finish the (at most two) tiles it shares:
//     the P CTAs of a tile each sum 1/P of its token columns over all P partials in CTA order (bit-deterministic) — a
//     reduce-scatter, so no CTA ingests more than one tile's worth of partials however finely K was split.
constexpr int kSkThreads = 384;  // warp 0 TMA (packed words), 1 MMA, 2 TMEM alloc, 3 TMA (x chunks), 4-11 dequant + epilogue
constexpr int kUnitK = 128;
constexpr int kUnitN = 256;
constexpr int kARing = 4;     // A buffers of 64 TMEM columns (128 channels x 128 k): two units in flight
constexpr int kACol0 = 256;   // A ring = TMEM columns [256, 512); accumulators below
constexpr int kAwqStateBytes = 256;  // workspace header: [0] launch epoch, [1] CTAs done

struct AwqSkArgs {
  void* out;           // [M, N]
  const uint8_t* qweight_t;  // unit-major packed words: 16 KB per (256-channel tile, 128-k block)
  const uint32_t* sz;  // [K/g, N] {scale, magic + zero}
  const void* bias;    // [N] or null
  int M, N, K, group_size, num_groups;
  int group_shift;  // log2(group_size) when it is a power of two, else -1 (then group_size % 128 == 0)
  int num_kb, total_units;
  int debug;  // B200_AWQ_DEBUG bit mask — timing experiments only, results are wrong when set: 1 no dequant math,
              // 2 no tcgen05.st, 4 no MMA
  unsigned long long* trace;  // diagnostics (b200_debug_set_trace): CTA 0 stamps 16 clocks per unit, or null
  int trace_units;
  uint32_t* state;     // [0] epoch of the last completed launch, [1] CTAs of this launch that are done
  uint32_t* flags;     // [2 * G] epoch-valued: slot 0 = a CTA's leading partial tile, slot 1 = its trailing one
  uint32_t* partials;  // [2 * G][2 groups][BN cols][128 rows] fp32
};

template <int BN>
struct AwqSk {
  static constexpr int kW = kUnitN * (kUnitK / 2);  // packed words of a unit
  static constexpr int kXC = BN * 128;              // x chunk: BN tokens x 64 k
  static constexpr int kSX = 6;  // three units of x chunks in flight (a deeper ring does not help: at 128 tokens the kernel
                                 // is bound by shared-memory bandwidth — see the measurements in DESIGN.md)
  static constexpr int kSWmax = (208 * 1024 - kSX * kXC) / kW;
  static constexpr int kSW = kSWmax > 10 ? 10 : kSWmax;
  static constexpr int kAccBufs = 4 * BN <= kACol0 ? 2 : 1;  // a buffer = two channel groups x BN token columns
  static constexpr int kWOff = kSX * kXC;
  static constexpr int kBarOff = kWOff + kSW * kW;
  static constexpr int kNumBars = 2 * kSW + 2 * kSX + 2 * kARing + 4;
  static constexpr int kBytes = kBarOff + kNumBars * 8 + 1024;
  static_assert(2 * BN * kAccBufs <= kACol0, "accumulators overlap the A ring");
  static_assert(kXC % 1024 == 0 && kBytes <= 232448, "shared-memory layout");
};

__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];\n" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void dq_bar_sync() { asm volatile("bar.sync 2, 256;\n" ::: "memory"); }

// tcgen05.mma with the A operand in tensor memory (M rows on the 128 lanes, K packed two 16-bit elements per column)
__device__ __forceinline__ void umma_ts_f16(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  const uint32_t acc = accumulate ? 1u : 0u;
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc)
      : "memory");
}

// first unit of CTA c's range, and the CTA whose range holds unit u (ranges are [c*U/G, (c+1)*U/G))
__device__ __forceinline__ int sk_first_unit(int c, int U, int G) { return (int)((int64_t)c * U / G); }
__device__ __forceinline__ int sk_owner_of(int u, int U, int G) {
  int c = (int)(((int64_t)u * G) / U);
  while (c + 1 < G && sk_first_unit(c + 1, U, G) <= u) ++c;
  while (c > 0 && sk_first_unit(c, U, G) > u) --c;
  return c;
}

__device__ __forceinline__ void awq_stamp(const AwqSkArgs& g, int unit_idx, int k) {
  if (g.trace != nullptr && blockIdx.x == 0 && unit_idx < g.trace_units) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%clock64;\n" : "=l"(t));
    g.trace[(size_t)unit_idx * 16 + k] = t;
  }
}

template <typename T, int BN>
__global__ void __launch_bounds__(kSkThreads, 1)
awq_streamk_kernel(const __grid_constant__ CUtensorMap tmap_x, const AwqSkArgs g) {
  using L = AwqSk<BN>;
  constexpr int SW = L::kSW;
  constexpr int SX = L::kSX;
  constexpr int NACC = L::kAccBufs;
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint32_t s_tmem_base;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar = base + L::kBarOff;
  auto w_full = [&](int s) { return bar + s * 8; };
  auto w_empty = [&](int s) { return bar + (SW + s) * 8; };
  auto x_full = [&](int t) { return bar + (2 * SW + t) * 8; };
  auto x_empty = [&](int t) { return bar + (2 * SW + SX + t) * 8; };
  auto a_full = [&](int r) { return bar + (2 * SW + 2 * SX + r) * 8; };
  auto a_empty = [&](int r) { return bar + (2 * SW + 2 * SX + kARing + r) * 8; };
  auto d_full = [&](int b) { return bar + (2 * SW + 2 * SX + 2 * kARing + b) * 8; };
  auto d_empty = [&](int b) { return bar + (2 * SW + 2 * SX + 2 * kARing + 2 + b) * 8; };
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 3 && lane == 0) prefetch_tmap(&tmap_x);
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < SW; ++s) {
      mbar_init(w_full(s), 1);
      mbar_init(w_empty(s), 8);  // the eight dequant warps
    }
    for (int t = 0; t < SX; ++t) {
      mbar_init(x_full(t), 1);
      mbar_init(x_empty(t), 1);
    }
    for (int r = 0; r < kARing; ++r) {
      mbar_init(a_full(r), 4);  // the four warps (lane quadrants) of the owning channel group
      mbar_init(a_empty(r), 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(d_full(b), 1);
      mbar_init(d_empty(b), 8);
    }
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc<512>(smem_u32(&s_tmem_base));
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = s_tmem_base;
  pdl_trigger();

  const int G = gridDim.x;
  const int U = g.total_units;
  const int cta = blockIdx.x;
  const int nkb = g.num_kb;
  const int u0 = sk_first_unit(cta, U, G);
  const int u1 = sk_first_unit(cta + 1, U, G);

  if (warp == 0) {
    // ---- packed words: independent of the previous kernel, so the ring fills before the PDL wait --------------------
    if (elect_one()) {
      for (int i = 0; i < u1 - u0; ++i) {
        const int s = i % SW;
        mbar_wait(w_empty(s), ((i / SW) & 1) ^ 1);
        awq_stamp(g, i, 13);
        mbar_expect_tx(w_full(s), L::kW);
        bulk_load_1d_hint(base + L::kWOff + s * L::kW, g.qweight_t + (size_t)(u0 + i) * L::kW, L::kW, w_full(s), kEvictFirst);
      }
    }
    pdl_wait();
  } else if (warp == 3) {
    // ---- x chunks (64 k each, two per unit): written by the previous kernel --------------------------------------------
    pdl_wait();
    if (elect_one()) {
      int kb = u0 % nkb;
      for (int i = 0; i < u1 - u0; ++i) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int j = 2 * i + h;
          const int t = j % SX;
          mbar_wait(x_empty(t), ((j / SX) & 1) ^ 1);
          awq_stamp(g, i, 14 + h);
          mbar_expect_tx(x_full(t), L::kXC);
          tma_load_2d(base + t * L::kXC, &tmap_x, x_full(t), kb * kUnitK * 2 + h * 128, 0, kEvictLast);
        }
        if (++kb == nkb) kb = 0;
      }
    }
  } else if (warp == 1) {
    pdl_wait();
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(/*f32*/ 1, AwqMagic<T>::kFmt, AwqMagic<T>::kFmt, 128, BN);
      int it = 0, seg = 0;
      for (int u = u0; u < u1; ++seg) {
        const int tile = u / nkb;
        const int seg_end = min(u1, (tile + 1) * nkb);
        const int buf = seg % NACC;
        mbar_wait(d_empty(buf), ((seg / NACC) & 1) ^ 1);
        tc_fence_after();
        for (int i = u; i < seg_end; ++i, ++it) {
          const int r0 = 2 * (it & 1);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int j = 2 * it + h;
            const int t = j % SX;
            mbar_wait(x_full(t), (j / SX) & 1);
            awq_stamp(g, it, h == 0 ? 8 : 11);
            const uint64_t db = smem_desc_k_sw128(base + t * L::kXC);
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              if (h == 0) {
                mbar_wait(a_full(r0 + c), (it >> 1) & 1);
                awq_stamp(g, it, 9 + c);
              }
              tc_fence_after();
              const uint32_t d_tmem = tmem_base + buf * (2 * BN) + c * BN;
              const uint32_t a_tmem = tmem_base + kACol0 + (r0 + c) * 64 + h * 32;
#pragma unroll
              for (int k = 0; k < 4; ++k)  // UMMA K = 16: 8 TMEM columns of A, 32 bytes of the 128-byte-swizzled x row
                if (!(g.debug & 4)) umma_ts_f16(d_tmem, a_tmem + 8 * k, db + 2 * k, idesc, (i > u) || (h | k) != 0);
            }
            tc_commit(x_empty(t));
          }
          tc_commit(a_empty(r0));
          tc_commit(a_empty(r0 + 1));
          awq_stamp(g, it, 12);
        }
        tc_commit(d_full(buf));
        u = seg_end;
      }
    }
  } else if (warp >= 4) {
    const int c = (warp - 4) >> 2;   // channel group of the unit
    const int q = warp & 3;          // TMEM lane quadrant this warp may touch
    const int row = c * 128 + q * 32 + lane;  // channel inside the 256-channel unit
    const uint32_t sw = (uint32_t)((row >> 1) & 3);
    const uint32_t lane_base = tmem_base + ((uint32_t)(q * 32) << 16);
    const int g_size = g.group_size;
    const int last_group = g.num_groups - 1;
    const bool has_bias = g.bias != nullptr;
    const int m_cols = min(BN, (g.M + 15) & ~15);
    T* out = reinterpret_cast<T*>(g.out);
    constexpr size_t kSlot = (size_t)2 * BN * 128;  // words per partial slot

    // {scale, magic + zero} of the (channel, group)s a unit touches: four 32-k chunks; no division in the unit loop
    auto load_sz = [&](int tile, int kb, uint32_t (&dst)[4]) {
      const int n = tile * kUnitN + row;
      const int k0 = kb * kUnitK;
      const int gi0 = g.group_shift >= 0 ? 0 : min(k0 / g_size, last_group);  // group_size % 128 == 0: one group per unit
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int gi = g.group_shift >= 0 ? min((k0 + j * 32) >> g.group_shift, last_group) : gi0;
        dst[j] = n < g.N ? __ldg(g.sz + (int64_t)gi * g.N + n) : 0u;
      }
    };
    uint32_t sz_nxt[4] = {0u, 0u, 0u, 0u};
    int nx_tile = u0 / nkb, nx_kb = u0 - nx_tile * nkb;  // the unit whose constants are being prefetched
    if (u0 < u1) load_sz(nx_tile, nx_kb, sz_nxt);

    int it = 0, seg = 0;
    uint32_t epoch = 0;
    int pend0 = -1, pend1 = -1;  // tiles this CTA shares with others: its leading partial (slot 0), its trailing one (slot 1)
    for (int u = u0; u < u1; ++seg) {
      const int tile = u / nkb;
      const int kb_begin = u - tile * nkb;
      const int seg_end = min(u1, (tile + 1) * nkb);
      const bool starts = kb_begin == 0;
      const bool ends = seg_end == (tile + 1) * nkb;
      const int buf = seg % NACC;
      // ---- dequantise the segment's units into the TMEM A ring --------------------------------------------------------
      for (int i = u; i < seg_end; ++i, ++it) {
        uint32_t sz_cur[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) sz_cur[j] = sz_nxt[j];
        if (++nx_kb == nkb) { nx_kb = 0; ++nx_tile; }
        if (i + 1 < u1) load_sz(nx_tile, nx_kb, sz_nxt);
        const int s = it % SW;
        const int r = 2 * (it & 1) + c;
        mbar_wait(w_full(s), (it / SW) & 1);
        if (q == 0 && lane == 0) awq_stamp(g, it, 4 * c + 0);
        const uint32_t raw_row = base + L::kWOff + s * L::kW + (uint32_t)row * 64u;
        uint4 w[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) w[j] = lds128(raw_row + ((((uint32_t)j) ^ sw) << 4));
        mbar_wait(a_empty(r), ((it >> 1) & 1) ^ 1);
        if (q == 0 && lane == 0) awq_stamp(g, it, 4 * c + 1);
        tc_fence_after();
        const uint32_t a_addr = lane_base + kACol0 + r * 64;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t s2 = __byte_perm(sz_cur[j], sz_cur[j], 0x1010);
          const uint32_t zm = __byte_perm(sz_cur[j], sz_cur[j], 0x3232);
          const uint32_t ww[4] = {w[j].x, w[j].y, w[j].z, w[j].w};
          uint32_t cells[16];
#pragma unroll
          for (int t = 0; t < 4; ++t)
#pragma unroll
            for (int p = 0; p < 4; ++p)
              cells[t * 4 + p] = (g.debug & 1) ? ww[t] + p : mul2<T>(sub2<T>(nib_pair<T>(ww[t], p), zm), s2);
          if (!(g.debug & 2)) tmem_st16(a_addr + j * 16, cells);
          if (j == 3) {  // the unit's words are in registers (consumed above): hand the ring slot back to the TMA warp
            __syncwarp();
            if (lane == 0) mbar_arrive(w_empty(s));
          }
        }
        if (q == 0 && lane == 0) awq_stamp(g, it, 4 * c + 2);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(a_full(r));
        if (q == 0 && lane == 0) awq_stamp(g, it, 4 * c + 3);
      }
      // ---- epilogue of the segment: this warp drains its (group, quadrant) of the accumulator -------------------------
      if (epoch == 0) {  // the first global write of this thread comes after the previous kernel has completed
        pdl_wait();
        epoch = ld_acquire_u32(g.state) + 1u;
      }
      mbar_wait(d_full(buf), (seg / NACC) & 1);
      tc_fence_after();
      const uint32_t taddr = lane_base + buf * (2 * BN) + c * BN;
      const int n = tile * kUnitN + row;
      const bool n_ok = n < g.N;
      if (starts && ends) {
        float bias = 0.f;
        if (n_ok && has_bias) bias = DT<T>::to_f(reinterpret_cast<const T*>(g.bias)[n]);
#pragma unroll 1
        for (int c0 = 0; c0 < m_cols; c0 += 16) {
          uint32_t r16[16];
          tmem_ld16(taddr + c0, r16);
          tmem_ld_wait();
          if (n_ok) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int m = c0 + j;
              if (m < g.M) {
                const float v = __uint_as_float(r16[j]);
                out[(int64_t)m * g.N + n] = DT<T>::from_f(has_bias ? v + bias : v);
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(d_empty(buf));
      } else {
        // publish: raw fp32 accumulators -> partials[2 * cta + slot][group][col][128 rows]; a warp store covers 128 bytes
        const int slot = starts ? 1 : 0;
        if (slot) pend1 = tile; else pend0 = tile;
        uint32_t* dst = g.partials + (size_t)(2 * cta + slot) * kSlot + (size_t)c * BN * 128 + q * 32 + lane;
#pragma unroll 1
        for (int c0 = 0; c0 < m_cols; c0 += 16) {
          uint32_t r16[16];
          tmem_ld16(taddr + c0, r16);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) __stcg(dst + (c0 + j) * 128, r16[j]);
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(d_empty(buf));
        dq_bar_sync();  // orders every epilogue thread's stores before the release below (cumulativity)
        if (threadIdx.x == 128) st_release_u32(g.flags + 2 * cta + slot, epoch);
      }
      u = seg_end;
    }
    if (epoch == 0) {
      pdl_wait();
      epoch = ld_acquire_u32(g.state) + 1u;
    }
    // ---- shared tiles: reduce this CTA's share of the token columns over all contributors, in CTA order ---------------
#pragma unroll 1
    for (int slot = 0; slot < 2; ++slot) {
      const int tile = slot ? pend1 : pend0;
      if (tile < 0) continue;
      const int t_begin = tile * nkb, t_end = t_begin + nkb;
      const int c_lo = sk_owner_of(t_begin, U, G);
      const int c_hi = sk_owner_of(t_end - 1, U, G);
      const int P = c_hi - c_lo + 1;
      const int me = cta - c_lo;
      // contributor p used slot 1 iff the tile starts inside (or at the start of) its range: only the first one can
      for (int p = lane; p < P; p += 32) {
        const uint32_t* f = g.flags + 2 * (c_lo + p) + (p == 0 ? 1 : 0);
        while (ld_acquire_u32(f) != epoch) {
        }
      }
      __syncwarp();
      const int col0 = (int)((int64_t)me * m_cols / P), col1 = (int)((int64_t)(me + 1) * m_cols / P);
      const int n = tile * kUnitN + row;
      const bool n_ok = n < g.N;
      float bias = 0.f;
      if (n_ok && has_bias) bias = DT<T>::to_f(reinterpret_cast<const T*>(g.bias)[n]);
      const uint32_t* src0 = g.partials + (size_t)(2 * c_lo + 1) * kSlot + (size_t)c * BN * 128 + q * 32 + lane;
      const uint32_t* srcp = g.partials + (size_t)(2 * (c_lo + 1)) * kSlot + (size_t)c * BN * 128 + q * 32 + lane;
#pragma unroll 1
      for (int col = col0; col < col1; col += 4) {
        float acc[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] = col + j < col1 ? __uint_as_float(__ldcg(src0 + (col + j) * 128)) : 0.f;
        for (int p = 1; p < P; ++p) {
          const uint32_t* sp = srcp + (size_t)(p - 1) * 2 * kSlot;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            if (col + j < col1) acc[j] = __fadd_rn(acc[j], __uint_as_float(__ldcg(sp + (col + j) * 128)));
        }
        if (n_ok) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int m = col + j;
            if (m < col1 && m < g.M) out[(int64_t)m * g.N + n] = DT<T>::from_f(has_bias ? acc[j] + bias : acc[j]);
          }
        }
      }
    }
    // ---- launch epoch: the last CTA to finish publishes it (flags of this launch can never match a later one) ----------
    dq_bar_sync();
    if (threadIdx.x == 128) {
      __threadfence();
      if (atomicAdd(g.state + 1, 1u) == (uint32_t)G - 1u) {
        g.state[1] = 0u;
        st_release_u32(g.state, epoch);
      }
    }
  } else {
    pdl_wait();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc<512>(tmem_base);
}

template <typename T, int BN>
static int launch_awq_sk(const CUtensorMap& tx, const AwqSkArgs& g, int ctas, cudaStream_t st) {
  using L = AwqSk<BN>;
  auto kern = awq_streamk_kernel<T, BN>;
  B200_CUDA_OK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::kBytes));
  return launch_k(kern, dim3((unsigned)ctas), dim3(kSkThreads), L::kBytes, st, 1, true, tx, g);
}

// workspace: [state 256 B][flags 2 * #SM words, padded to 4 KB][partials 2 * #SM slots of 2 x 128 x 128 fp32]
static size_t awq_flag_bytes() { return ((size_t)2 * sm_count() * 4 + 4095) / 4096 * 4096; }
static size_t awq_workspace_bytes() { return kAwqStateBytes + awq_flag_bytes() + (size_t)2 * sm_count() * 2 * 128 * 128 * 4; }

template <typename T>
static int run_awq_gemm(AwqSkArgs g, const void* x, const void* qweight_t, int64_t M, void* ws, size_t ws_bytes,
                        cudaStream_t st) {
  const size_t need = awq_workspace_bytes();
  if (ws == nullptr || ws_bytes < need) return fail(B200_ERR_WORKSPACE, "awq_gemm: workspace of %zu bytes, %zu needed", ws_bytes, need);
  B200_REQUIRE(((uintptr_t)ws % 256) == 0, "awq_gemm: workspace must be 256-byte aligned");
  const int64_t N = g.N, K = g.K;
  g.num_kb = (int)ceil_div(K, kUnitK);
  const int tiles = (int)ceil_div(N, kUnitN);
  const int64_t units = (int64_t)tiles * g.num_kb;
  B200_REQUIRE(units < (1ll << 30), "awq_gemm: too many work units");
  g.total_units = (int)units;
  static const int dbg = getenv("B200_AWQ_DEBUG") ? atoi(getenv("B200_AWQ_DEBUG")) : 0;
  g.debug = dbg;
  g.trace = g_trace;
  g.trace_units = (int)g_trace_slots;
  uint8_t* wsb = reinterpret_cast<uint8_t*>(ws);
  g.state = reinterpret_cast<uint32_t*>(wsb);
  g.flags = reinterpret_cast<uint32_t*>(wsb + kAwqStateBytes);
  g.partials = reinterpret_cast<uint32_t*>(wsb + kAwqStateBytes + awq_flag_bytes());
  const int sms = sm_count();
  int ctas = (int)(units < sms ? units : sms);
  // Above 64 tokens a unit costs a CTA ~1.1 us whatever the grid does (the fp16 tensor pipe and the dequant chain, not
  // the weight stream, pace it), and what a launch pays on top is the fix-up.  With no more tiles than SMs the grid is
  // therefore tiles x P CTAs: ranges [c*U/G, (c+1)*U/G) then never straddle a tile, every tile is cut into exactly P
  // contiguous K ranges, and each CTA publishes and reduces at most once.  P minimises the measured cost model
  // ceil(nkb / P) * 1.1 + 0.025 * G + 0.1 * P + 3 us (P = 1: nkb * 1.1, no fix-up).  Per-shard GEMMs of cfg 4 at M = 128:
  // layer total 152 -> 128 us (TP 2), 125 -> 93 (TP 4), 88 -> 70 (TP 8): profiles/r02zd_awq_ctas_per_tile_sweep.jsonl.
  // B200_AWQ_CTAS_PER_TILE = p forces P (0 = one CTA per SM, the stream-K split of the larger shapes).
  static const int per_tile_env = getenv("B200_AWQ_CTAS_PER_TILE") ? atoi(getenv("B200_AWQ_CTAS_PER_TILE")) : -1;
  int per_tile = per_tile_env;
  if (per_tile < 0 && M > 64 && tiles <= sms) {
    const int pmax = sms / tiles < g.num_kb ? sms / tiles : g.num_kb;
    float best = 1.1f * (float)g.num_kb;
    per_tile = 1;
    for (int p = 2; p <= pmax; ++p) {
      const float c = 1.1f * (float)((g.num_kb + p - 1) / p) + 0.025f * (float)(tiles * p) + 0.1f * (float)p + 3.f;
      if (c < best) { best = c; per_tile = p; }
    }
  }
  if (per_tile > 0 && (int64_t)tiles * per_tile <= ctas) ctas = tiles * per_tile;
  CUtensorMap tx;
  int rc;
  g.qweight_t = reinterpret_cast<const uint8_t*>(qweight_t);
  uint8_t* outp = reinterpret_cast<uint8_t*>(g.out);
  const uint8_t* xp = reinterpret_cast<const uint8_t*>(x);
  // more than 128 tokens: one pass over the packed weights per 128-token chunk (at 128 tokens per pass the fp16 tensor
  // pipe, not the weight stream, is already the bound: 512 flop per weight byte)
  for (int64_t m0 = 0; m0 < M; m0 += 128) {
    const int64_t mc = M - m0 < 128 ? M - m0 : 128;
    g.M = (int)mc;
    g.out = outp + m0 * N * 2;
    const int bn = mc <= 16 ? 16 : mc <= 32 ? 32 : mc <= 64 ? 64 : 128;
    if ((rc = cached_tmap(&tx, xp + m0 * K * 2, (uint64_t)mc, (uint64_t)(K * 2), (uint32_t)bn)) != B200_OK) return rc;
    switch (bn) {
      case 16: rc = launch_awq_sk<T, 16>(tx, g, ctas, st); break;
      case 32: rc = launch_awq_sk<T, 32>(tx, g, ctas, st); break;
      case 64: rc = launch_awq_sk<T, 64>(tx, g, ctas, st); break;
      default: rc = launch_awq_sk<T, 128>(tx, g, ctas, st); break;
    }
    if (rc != B200_OK) return rc;
  }
  return B200_OK;
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

B200_API int b200_awq_repack(int32_t* qweight_t, uint32_t* sz, const int32_t* qweight, const void* scales,
                             const int32_t* qzeros, int64_t K, int64_t N, int group_size, int dtype, void* stream) {
  // one-time layout change for the streaming GEMM (process_weights_after_loading): see awq_repack_*_kernel
  B200_REQUIRE(K > 0 && N > 0 && N % 8 == 0 && K % 8 == 0, "awq_repack: K %lld and N %lld must be multiples of 8",
               (long long)K, (long long)N);
  B200_REQUIRE(qweight_t && sz && qweight && scales && qzeros, "awq_repack: null pointer");
  B200_REQUIRE(group_size > 0 && K % group_size == 0, "awq_repack: K %lld is not a multiple of group_size %d", (long long)K,
               group_size);
  if (dtype != B200_F16 && dtype != B200_BF16) return fail(B200_ERR_UNSUPPORTED, "awq_repack: dtype %d not supported", dtype);
  cudaStream_t st = (cudaStream_t)stream;
  const int64_t words = N * (K / 8), cells = (K / group_size) * N;
  const int64_t num_kb = ceil_div(K, 128);
  B200_CUDA_OK(cudaMemsetAsync(qweight_t, 0, (size_t)(ceil_div(N, 256) * num_kb) * 16384, st));  // padding past N and K
  awq_repack_weight_kernel<<<(unsigned)ceil_div(words, 256), 256, 0, st>>>((uint32_t*)qweight_t, (const uint32_t*)qweight, K, N,
                                                                           num_kb);
  B200_LAUNCH_OK();
  awq_repack_sz_kernel<<<(unsigned)ceil_div(cells, 256), 256, 0, st>>>(sz, (const uint16_t*)scales, (const uint32_t*)qzeros,
                                                                       K / group_size, N, dtype == B200_F16 ? 0x6400u : 0x4300u);
  B200_LAUNCH_OK();
  return B200_OK;
}

B200_API int b200_awq_gemm_workspace_bytes(size_t* bytes) {
  B200_REQUIRE(bytes != nullptr, "awq_gemm_workspace_bytes: null pointer");
  *bytes = awq_workspace_bytes();
  return B200_OK;
}

B200_API int b200_awq_gemm(void* out, const void* x, const int32_t* qweight_t, const uint32_t* sz, const void* bias, int64_t M,
                           int64_t N, int64_t K, int group_size, int dtype, void* workspace, size_t workspace_bytes,
                           void* stream) {
  B200_REQUIRE(M >= 0 && N > 0 && K > 0, "awq_gemm: bad sizes");
  if (M == 0) return B200_OK;
  B200_REQUIRE(out && x && qweight_t && sz, "awq_gemm: null pointer");
  // shape rules of AWQLinearMethod.create_weights (awq.py:339-352): K % group == 0, N % pack_factor == 0; the TMA row
  // pitches additionally need K % 32 == 0 (16-byte rows of packed words)
  B200_REQUIRE(group_size > 0 && K % group_size == 0, "awq_gemm: K %lld is not a multiple of group_size %d", (long long)K,
               group_size);
  if (!(group_size == 32 || group_size == 64 || group_size % 128 == 0))
    return fail(B200_ERR_UNSUPPORTED, "awq_gemm: group_size %d not supported (32, 64 or a multiple of 128)", group_size);
  B200_REQUIRE(N % 8 == 0, "awq_gemm: N %lld must be a multiple of 8", (long long)N);
  B200_REQUIRE(K % 32 == 0, "awq_gemm: K %lld must be a multiple of 32", (long long)K);
  B200_REQUIRE(((uintptr_t)out % 2) == 0 && ((uintptr_t)x % 16) == 0 && ((uintptr_t)qweight_t % 16) == 0, "awq_gemm: pointer alignment");
  B200_REQUIRE(M < (1ll << 30) && N < (1ll << 30) && K < (1ll << 30), "awq_gemm: sizes exceed 32-bit tile math");
  AwqSkArgs g = {};
  g.out = out; g.sz = sz; g.bias = bias;
  g.N = (int)N; g.K = (int)K; g.group_size = group_size; g.num_groups = (int)(K / group_size);
  g.group_shift = (group_size & (group_size - 1)) == 0 ? 31 - __builtin_clz((unsigned)group_size) : -1;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == B200_F16) return run_awq_gemm<__half>(g, x, qweight_t, M, workspace, workspace_bytes, st);
  if (dtype == B200_BF16) return run_awq_gemm<__nv_bfloat16>(g, x, qweight_t, M, workspace, workspace_bytes, st);
  return fail(B200_ERR_UNSUPPORTED, "awq_gemm: dtype %d not supported (f16/bf16)", dtype);
}
