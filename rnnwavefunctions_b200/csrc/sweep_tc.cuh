// Tensor-core regime of the lattice sweep (tcgen05 / TMEM), for the tensorized 2D GRU cell.
//
// Per lattice site and 128-row tile the cell needs  s, g = hc (W[:,:,sl] + W[:,:,2+su]) , m = hc Wmerge  with a weight
// slice that depends on the row's neighbouring spins (sl, su).  On the tensor cores the dependence is moved to
// the epilogue by linearity:
//        s = hc Wa + sl * (hc Wb) + su * (hc Wc),   Wa = W0 + W2,  Wb = W1 - W0,  Wc = W3 - W2   (same for g)
// so every row multiplies the SAME operand: one dense [128 x 208] x [208 x 700] contraction per site, issued as
// tcgen05.mma (kind::f16, FP16 in, FP32 accumulate in TMEM) in 7 column passes of 112 (16 hidden units x 7 blocks),
// double-buffered in TMEM against the epilogue.  Operands are FP16 rather than BF16: the hidden states are O(1) and
// the weights O(0.1), so the range is no issue and the 11-bit mantissa keeps E_loc within ~2e-4 of the FP32 path
// (measured on the oracle: BF16 activations give 2.4e-3, which misses the 1e-3 bound of the reduced-precision mode).
//
//   warps 0, 18  : MMA issue, one warp per accumulator buffer (even / odd column passes) -- tcgen05.mma with the A operand
//                  in TMEM and the weights in shared memory (UMMA K-major core-matrix layout), K loop fully unrolled,
//                  tcgen05.commit to the epilogue and to the loader.  Two issuers because a warp's waits after its own
//                  MMAs only return once they have executed.  Biases ride on a constant-1 operand position.
//   warp 19      : loader -- cp.async.bulk of the pre-packed weight images (mbarrier complete_tx) through 4 stages,
//                  up to 4 column passes ahead, across sites and tiles.
//   warps 1..16  : epilogue, four threads per row (row == TMEM lane; the four share the 16 units of every column pass).
//                  They read their 28 accumulators with tcgen05.ld, apply the gates, and write the new hidden state as
//                  FP16 straight into the next site's A operand in TMEM (tcgen05.st) and into the L2-resident
//                  history ring; the older neighbour half of the next operand is prefetched from the ring one site
//                  ahead.  No block barrier in the site loop.
//   warp 17      : heads, four rows per thread: softmax / U(1) mask / draw / log-probability / phase, one site behind
//                  the epilogue warps (partial head sums handed over in shared memory), off the critical path.
// Everything is synchronised by mbarriers; every wait is bounded and traps instead of hanging.
//
// Same reference behaviour as sweep.cuh for 2DHeisenberg/MDTensorizedGRUcell.py:49-61 and MDComplexRNN.py:69-286.
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace rnnwf {

constexpr int TC_ROWS = 128;
constexpr int TC_JP = 16;               // hidden units per column pass
constexpr int TC_NB = 7;                // column blocks per unit: Wa_s Wb_s Wc_s Wa_g Wb_g Wc_g Wmerge
constexpr int TC_N = TC_JP * TC_NB;     // 112 MMA columns per pass
constexpr int TC_SUB = 4;                // threads per row: each handles 4 of the 16 units of a pass
constexpr int TC_EPI_THREADS = TC_ROWS * TC_SUB;
constexpr int TC_HEAD_THREADS = 32;     // four rows per head thread
constexpr int TC_HEAD_ROWS = TC_ROWS / TC_HEAD_THREADS;
constexpr int TC_MMA2_WARP = 18;        // second MMA-issue warp
// warp 0: MMA issue (even passes), 1..16: epilogue, 17: heads, 18: MMA issue (odd passes), 19: loader
constexpr int TC_THREADS = 32 + TC_EPI_THREADS + TC_HEAD_THREADS + 32 + 32;
constexpr int TC_TPS = TC_NB * 4;       // accumulator columns one epilogue thread owns per pass (28, contiguous)
// one step's FP16 history image: [sub 4][quad <= 4][row 128][16 bytes]; quad i of thread (row, sub) holds its pieces 2i, 2i+1
// (a warp's 16-byte accesses are contiguous: fully coalesced)
constexpr uint32_t TC_RING_QUAD = TC_ROWS * 16;
// (only the quads that hold unit pieces are kept: 4 at 100 units, 2 at 50 -- the ring of C3 then fits the L2)
__host__ __device__ inline int tc_ring_quads(int nhp) { return (((nhp + TC_JP - 1) / TC_JP) + 1) / 2; }
__host__ __device__ inline uint32_t tc_ring_slot(int nhp) { return (uint32_t)(TC_SUB * tc_ring_quads(nhp)) * TC_RING_QUAD; }

// Thread-major order of the K axis inside one neighbour half (nhp units): epilogue thread `sub` owns units
// 16 p + 4 sub + e; its units occupy the contiguous positions tc_koff(sub) + 4 p + e.  Used identically by the operand
// in TMEM, the ring image and the packed weights.
__host__ __device__ inline int tc_kcnt(int nhp, int sub) { return (4 * sub < nhp) ? 4 * ((nhp - 4 * sub + 15) / 16) : 0; }
__host__ __device__ inline int tc_koff(int nhp, int sub) { int o = 0; for (int s = 0; s < sub; ++s) o += tc_kcnt(nhp, s); return o; }
// unit index of position kp (0 <= kp < nhp)
__host__ __device__ inline int tc_kunit(int nhp, int kp) {
  int sub = 0, off = 0;
  while (sub < TC_SUB - 1 && kp >= off + tc_kcnt(nhp, sub)) { off += tc_kcnt(nhp, sub); ++sub; }
  const int r = kp - off;
  return 16 * (r / 4) + 4 * sub + (r % 4);
}
constexpr int TC_TMEM_COLS = 512;         // 2 x 112 accumulator columns + 2 x (nhp) A-operand columns
constexpr int TC_STAGES = 4;              // weight stages in shared memory
constexpr uint32_t TC_SPIN_LIMIT = 1u << 26;

struct TcSmem {
  int off_B, off_headw, off_headb, off_bias, off_part, off_meta, off_step, off_sym, off_bar, off_tmem, total;
  int b_bytes, kc;
  __host__ __device__ TcSmem(int nhp, int Tsym, int Q = 0) {
    kc = 2 * (nhp / 8);                       // 16-byte k-cores of the concatenated [h_h ; h_v] operand
    b_bytes = kc * (TC_N / 8) * 128;
    int o = 0;
    off_B = o;   o += TC_STAGES * b_bytes;     // weight stages (the A operand lives in TMEM)
    off_headw = o; o += (nhp + TC_JP) * 16;     // head weights [unit][4], zero rows up to the last pass's end
    off_headb = o; o += 16;
    off_bias = o;  o += 6 * (nhp + TC_JP) * 4;    // tGRU: b, bg; vanilla: b' of the 4 neighbour variants, dUh, dUv
    off_part = o;  o += 2 * TC_ROWS * TC_SUB * 16;   // partial head sums [2][128][TC_SUB] float4
    off_meta = o;  o += ((Q > 0 ? Q : Tsym) * 8 + 15) & ~15;   // per step: nbr0, nbr1, site, pad (int16 x 4)
    off_step = o;  o += (Q > 0 ? Q : Tsym) * 16;                // per step: neighbours, carry half, ring slots (int4)
    off_bar = o;   o += 256;
    off_tmem = o;  o += 16;
    off_sym = o;   o += Tsym * 16;               // symbols, one bit per (site, row): 4 ballot words per site
    total = o + 1024;                          // slack for the 1024-byte alignment of the base
  }
};

// ---- PTX wrappers ------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok;
}
__device__ __forceinline__ uint32_t mbar_test_wait(uint64_t* bar, uint32_t parity) {   // non-blocking poll
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok;
}
__device__ __forceinline__ long long tc_clock() { long long c; asm volatile("mov.u64 %0, %%clock64;" : "=l"(c)); return c; }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t spin = 0; !mbar_try_wait(bar, parity); ++spin)
    if (spin > TC_SPIN_LIMIT) { asm volatile("trap;"); }
}
// the same on a precomputed shared-space address
__device__ __forceinline__ void mbar_arrive_a(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait_a(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  if (ok) return;
  for (uint32_t spin = 0;; ++spin) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    if (ok) return;
    if (spin > TC_SPIN_LIMIT) { asm volatile("trap;"); }
  }
}
// wait of a role that is off the critical path: let the hardware suspend the warp for up to ~4 us per poll instead of
// spinning on the issue slots the epilogue warps need
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  for (uint32_t spin = 0;; ++spin) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity), "r"(4000u) : "memory");
    if (ok) return;
    if (spin > TC_SPIN_LIMIT) { asm volatile("trap;"); }
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// one lane of a converged warp
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_alloc(uint32_t* smem_result, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tc_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T ; both operands K-major, no swizzle
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T ; A: lane = row, 32-bit column = two consecutive K halves; B K-major, no swizzle
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_st4(uint32_t taddr, const uint4& v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x4.b32 [%0], {%1, %2, %3, %4};" ::"r"(taddr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void tc_st2(uint32_t taddr, const uint2& v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x2.b32 [%0], {%1, %2};" ::"r"(taddr), "r"(v.x), "r"(v.y) : "memory");
}
__device__ __forceinline__ void tc_st8(uint32_t taddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d, uint32_t e, uint32_t f, uint32_t g, uint32_t h) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
               ::"r"(taddr), "r"(a), "r"(b), "r"(c), "r"(d), "r"(e), "r"(f), "r"(g), "r"(h) : "memory");
}
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t (&d)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(d[0]), "r"(d[1]), "r"(d[2]), "r"(d[3]), "r"(d[4]), "r"(d[5]), "r"(d[6]), "r"(d[7]), "r"(d[8]), "r"(d[9]),
        "r"(d[10]), "r"(d[11]), "r"(d[12]), "r"(d[13]), "r"(d[14]), "r"(d[15]) : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&d)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]),
        "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_ld8(uint32_t taddr, uint32_t* d) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tc_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// shared-memory matrix descriptor: K-major, SWIZZLE_NONE, core matrix = 8 rows x 16 bytes (128 contiguous bytes);
// LBO = byte distance between the two k-cores of one K=16 slice, SBO = byte distance between 8-row groups.
__device__ __forceinline__ uint64_t tc_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr >> 4) & 0x3FFF) | ((uint64_t)((lbo >> 4) & 0x3FFF) << 16) | ((uint64_t)((sbo >> 4) & 0x3FFF) << 32) |
         (1ull << 46);
}
// instruction descriptor: D=F32 (bit 4), A=B=F16 (format 0), both K-major, N>>3 at [17,23), M>>4 at [24,29)
__host__ __device__ constexpr uint32_t tc_idesc(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_f16x2(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// split-precision pair: hi = fp16(x), lo = fp16(x - hi)   (hi + lo carries 22 mantissa bits of x)
__device__ __forceinline__ void split_f16x2(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 h = __floats2half2_rn(a, b);
  const float2 hf = __half22float2(h);
  const __half2 l = __floats2half2_rn(a - hf.x, b - hf.y);
  hi = *reinterpret_cast<const uint32_t*>(&h);
  lo = *reinterpret_cast<const uint32_t*>(&l);
}

// ------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tc_ld4(uint32_t taddr, uint32_t (&d)[4]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]) : "r"(taddr) : "memory");
}
// the four warps that share a TMEM lane quarter (= the four threads of each of its 32 rows)
__device__ __forceinline__ void quarter_barrier(int quarter) { asm volatile("bar.sync %0, 128;" ::"r"(1 + quarter) : "memory"); }
// Gate non-linearities of the reduced-precision regime: one MUFU.TANH each (tanh.approx.f32, relative error 2^-11, the
// same size as the FP16 rounding of the operands; tools/tc_accuracy.py shows no measurable difference against exact
// gates).  sigmoid(g) = 1/2 + 1/2 tanh(g/2): the factor 1/2 is folded into the packed update-gate weights and biases
// (pack_tc_kernel; exact in FP16).  The special-function unit issues 16 ops/clk/SM, so the two MUFU per unit
// (instead of four with ex2 + rcp) matter: 512 threads x 4 units x 2 = 256 cycles per column pass.
constexpr float TC_SCALE_CAND = 1.0f, TC_SCALE_GATE = 0.5f;
// split-precision regime: hidden states are stored as hi + lo of TC_SPLIT_HS * h and weights as hi + lo of TC_SPLIT_WS * w
// (exact power-of-two scalings that keep the lo parts out of FP16's subnormal range: |h| >= 2e-3, |w| >= 1.2e-4);
// FP16 overflows beyond |h| = 1023 and |w| = 63.9
constexpr float TC_SPLIT_HS = 64.0f, TC_SPLIT_WS = 1024.0f;
// Sampling / evaluation / reverse sweeps over fewer 128-row tiles than SMs: every site is a latency chain per tile, so the
// rows are spread over all SMs -- live rows per tile (a multiple of 8, <= 128; the MMAs stay 128 rows high).
inline int tc_tile_rows(long long rows, int num_sms) {
  const long long tiles = (rows + TC_ROWS - 1) / TC_ROWS;
  if (tiles >= num_sms || std::getenv("RNNWF_NO_VTILES") != nullptr) return TC_ROWS;
  const long long per = (tiles * TC_ROWS + num_sms - 1) / num_sms;
  return (int)std::max<long long>(8, std::min<long long>(TC_ROWS, (per + 7) / 8 * 8));
}
__device__ __forceinline__ float mufu_tanh(float x) { float y; asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigmoid_prescaled(float x) { return fmaf(0.5f, mufu_tanh(x), 0.5f); }
__device__ __forceinline__ float tanh_prescaled(float x) { return mufu_tanh(x); }
// gates of the split-precision regime: ex2.approx (2 ulp) and rcp.approx (1 ulp) -- absolute error below 3e-7, the class of
// expf / tanhf at a third of their instructions (the epilogue's instruction issue bounds this regime, profiles/r2d_summary.md)
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rcp_approx(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sigmoid_split(float x) { return rcp_approx(1.0f + ex2_approx(-1.4426950408889634f * x)); }
__device__ __forceinline__ float tanh_split(float x) {
  const float e = ex2_approx(-2.8853900817779268f * fabsf(x));      // exp(-2 |x|)
  return copysignf((1.0f - e) * rcp_approx(1.0f + e), x);
}

// symbol of (site t, tile row) in the bit-packed table
__device__ __forceinline__ int tc_sym(const uint32_t* symS, int t, int row) { return (int)((symS[t * 4 + (row >> 5)] >> (row & 31)) & 1u); }

// K loop of one column pass.  HC > 0: fully unrolled, every operand address is base + immediate; HC == 0: generic.
// `acc0`: accumulate flag of the first K step (0 starts a fresh accumulator, 1 adds to it: the split-precision passes).
template <int HC>
__device__ __forceinline__ void tc_issue_pass(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, int hc = HC, uint32_t acc0 = 0u) {
  constexpr uint64_t B_KSTEP = (2ull * 128) >> 4;   // two k-cores of the weight image per K=16 step (n-core-major image: adjacent)
  if constexpr (HC > 0) {
#pragma unroll
    for (int ks = 0; ks < HC; ++ks) tc_mma_f16_ts(tmem_d, tmem_a + (uint32_t)ks * 8u, bdesc + (uint64_t)ks * B_KSTEP, idesc, ks > 0 ? 1u : acc0);
  } else {
#pragma unroll 1
    for (int ks = 0; ks < hc; ++ks) tc_mma_f16_ts(tmem_d, tmem_a + (uint32_t)ks * 8u, bdesc + (uint64_t)ks * B_KSTEP, idesc, ks > 0 ? 1u : acc0);
  }
}
// K steps [K0, K1) of a column pass, fully unrolled (see tc_issue_pass); `acc_first`: accumulate flag of step K0
template <int K0, int K1>
__device__ __forceinline__ void tc_issue_range(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t acc_first) {
  constexpr uint64_t B_KSTEP = (2ull * 128) >> 4;
#pragma unroll
  for (int ks = K0; ks < K1; ++ks) tc_mma_f16_ts(tmem_d, tmem_a + (uint32_t)ks * 8u, bdesc + (uint64_t)ks * B_KSTEP, idesc, ks > K0 ? 1u : acc_first);
}
// the 28 accumulator columns of one epilogue thread (contiguous in TMEM, see pack_tc_kernel): three loads
__device__ __forceinline__ void tc_ld28(uint32_t taddr, uint32_t (&d)[TC_TPS]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(d[0]), "=r"(d[1]), "=r"(d[2]), "=r"(d[3]), "=r"(d[4]), "=r"(d[5]), "=r"(d[6]), "=r"(d[7]), "=r"(d[8]),
        "=r"(d[9]), "=r"(d[10]), "=r"(d[11]), "=r"(d[12]), "=r"(d[13]), "=r"(d[14]), "=r"(d[15])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(d[16]), "=r"(d[17]), "=r"(d[18]), "=r"(d[19]), "=r"(d[20]), "=r"(d[21]), "=r"(d[22]), "=r"(d[23])
               : "r"(taddr + 16u) : "memory");
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
               : "=r"(d[24]), "=r"(d[25]), "=r"(d[26]), "=r"(d[27]) : "r"(taddr + 24u) : "memory");
}

}  // namespace rnnwf
