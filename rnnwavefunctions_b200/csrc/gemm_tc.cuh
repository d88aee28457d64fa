// Generic tcgen05 GEMM for the backward pass: C[M,N] (+)= A(M,K) * B(N,K)^T on the tensor cores, TF32 operands
// (FP32 in memory, 10-bit mantissa in the multiplier -- FP32 range, which gradients need), FP32 accumulation in TMEM.
//
// Operands stay FP32 row-major in global memory; the CTA's threads copy 128 x 32 tiles into shared memory in the
// UMMA core-matrix layout (8 rows x 16 bytes, K-major), whether the operand is contiguous along K or
// is contiguous along M / N (transposed on the fly by the staging copy), so none of the three dense-layer backward
// shapes needs a transposed copy in global memory:
//      dX = dY W^T   (A = dY K-contig, B = W   K-contig)          gemm "nt"
//      dX = dY W     (A = dY K-contig, B = W   N-contig)          gemm "nn"
//      dW += X^T dY  (A = X  M-contig, B = dY  N-contig, split over the batch, atomics)   gemm "tn"
// One elected thread issues tcgen05.mma (kind::tf32, M=128, N=128, K=8); a 3-stage cp.async pipeline keeps two tiles
// of loads in flight behind the MMAs; the epilogue reads TMEM with tcgen05.ld (one thread per accumulator row).
#pragma once
#include "common.cuh"
#include "sweep_tc.cuh"

namespace rnnwf {

constexpr int GT_BM = 128, GT_BN = 128, GT_BK = 32;
constexpr int GT_TILE_BYTES = GT_BM * GT_BK * 4;   // 16 KB per operand tile

struct GemmTcArgs {
  const float* A; long long lda; int a_kc;   // a_kc = 1: A[m*lda + k] ; 0: A[k*lda + m]
  const float* B; long long ldb; int b_kc;   // b_kc = 1: B[n*ldb + k] ; 0: B[k*ldb + n]
  float* C; long long ldc;
  long long M; int N; long long K;
  long long kchunk;                           // K range per blockIdx.z (multiple of GT_BK)
  int mode;                                   // 0: C = , 1: C += (plain), 2: atomicAdd
};

__device__ __forceinline__ void tc_mma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

// Stage a [128 (mn) x 32 (k)] FP32 tile into shared memory as K-major core matrices:
//   offset = (k/4)*2048 + (mn/8)*128 + (mn%8)*16 + (k%4)*4
//   kc = 1 (contiguous along k):  each 16-byte core row is one cp.async (zero-filled past the edges) -- asynchronous;
//   kc = 0 (contiguous along mn): coalesced scalar loads along mn, transposed in registers (MN-major TF32 operands
//                                 would need the special 128B/32B-base swizzle, so the copy does the transposition).
__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gmem_src, int src_bytes) {
  unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gmem_src), "r"(src_bytes));
}
__device__ __forceinline__ void gt_load_tile(unsigned char* smem, const float* __restrict__ G, long long ld, int kc,
                                             long long mn0, long long mn_lim, long long k0, long long k_lim, int tid) {
  const long long mn = mn0 + tid;
  const uint32_t base = (uint32_t)(tid >> 3) * 128 + (uint32_t)(tid & 7) * 16;
  const bool ok = mn < mn_lim;
  if (kc) {
    const float* src = G + (ok ? mn : 0) * ld + k0;
#pragma unroll
    for (int c = 0; c < GT_BK / 4; ++c) {
      const long long rem = k_lim - (k0 + 4 * c);
      const int bytes = (!ok || rem <= 0) ? 0 : (rem >= 4 ? 16 : (int)rem * 4);
      cp_async16_zfill(smem + (uint32_t)c * 2048 + base, bytes ? (const void*)(src + 4 * c) : (const void*)G, bytes);
    }
  } else {
    const float* src = G + k0 * ld + mn;
#pragma unroll
    for (int c = 0; c < GT_BK / 4; ++c) {
      const long long k = k0 + 4 * c;
      float4 v;
      v.x = (ok && k < k_lim) ? __ldg(src + (long long)(4 * c) * ld) : 0.f;
      v.y = (ok && k + 1 < k_lim) ? __ldg(src + (long long)(4 * c + 1) * ld) : 0.f;
      v.z = (ok && k + 2 < k_lim) ? __ldg(src + (long long)(4 * c + 2) * ld) : 0.f;
      v.w = (ok && k + 3 < k_lim) ? __ldg(src + (long long)(4 * c + 3) * ld) : 0.f;
      *reinterpret_cast<float4*>(smem + (uint32_t)c * 2048 + base) = v;
    }
  }
}

constexpr int GT_STAGES = 3;

__global__ void __launch_bounds__(128, 2) gemm_tc_kernel(const GemmTcArgs g) {
  extern __shared__ __align__(128) unsigned char gt_smem[];   // [stage][A tile | B tile]
  __shared__ __align__(8) uint64_t bars[GT_STAGES];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long m0 = (long long)blockIdx.y * GT_BM;
  const long long n0 = (long long)blockIdx.x * GT_BN;
  const long long kbeg = (long long)blockIdx.z * g.kchunk;
  const long long kend = (kbeg + g.kchunk < g.K) ? kbeg + g.kchunk : g.K;
  const int niter = (kend > kbeg) ? (int)((kend - kbeg + GT_BK - 1) / GT_BK) : 0;

  if (tid == 0) { for (int i = 0; i < GT_STAGES; ++i) mbar_init(bars + i, 1); fence_mbar_init(); }
  if (warp == 0) tc_alloc(&tmem_slot, GT_BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = tmem_slot;

  // instruction descriptor: D=F32, A=B=TF32 (format 2), both K-major
  const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(GT_BN >> 3) << 17) | ((uint32_t)(GT_BM >> 4) << 24);

  auto stage_A = [&](int s) { return gt_smem + (size_t)s * 2 * GT_TILE_BYTES; };
  auto stage_B = [&](int s) { return gt_smem + (size_t)s * 2 * GT_TILE_BYTES + GT_TILE_BYTES; };
  auto issue_loads = [&](int it) {       // tile `it` -> stage it % GT_STAGES ; always commits one cp.async group
    if (it < niter) {
      const int s = it % GT_STAGES;
      const long long k0 = kbeg + (long long)it * GT_BK;
      gt_load_tile(stage_A(s), g.A, g.lda, g.a_kc, m0, g.M, k0, kend, tid);
      gt_load_tile(stage_B(s), g.B, g.ldb, g.b_kc, n0, (long long)g.N, k0, kend, tid);
    }
    asm volatile("cp.async.commit_group;\n" ::);
  };

  uint32_t nwait[GT_STAGES];
#pragma unroll
  for (int i = 0; i < GT_STAGES; ++i) nwait[i] = 0;
#pragma unroll
  for (int it = 0; it < GT_STAGES - 1; ++it) issue_loads(it);

  for (int it = 0; it < niter; ++it) {
    const int s = it % GT_STAGES;
    asm volatile("cp.async.wait_group %0;\n" ::"n"(GT_STAGES - 2) : "memory");   // this thread's copies of tile `it` landed
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0) {
      tc_fence_after();
      if (elect_one()) {
        // per-MMA (K = 8) descriptor: two k-groups of 4 (LBO = 2048), 8-row groups SBO = 128, advance 4096 bytes per MMA
        const uint64_t ad = tc_desc(smem_u32(stage_A(s)), 2048u, 128u), bd = tc_desc(smem_u32(stage_B(s)), 2048u, 128u);
#pragma unroll
        for (int ks = 0; ks < GT_BK / 8; ++ks)
          tc_mma_tf32(tmem_d, ad + (uint64_t)(ks * 4096 >> 4), bd + (uint64_t)(ks * 4096 >> 4), idesc, (it > 0 || ks > 0) ? 1u : 0u);
        tc_commit(bars + s);
      }
      __syncwarp();
    }
    // refill the stage that tile it-1 used (its MMAs were committed one iteration ago) with tile it + STAGES - 1
    if (it >= 1) { const int sp = (it - 1) % GT_STAGES; mbar_wait(bars + sp, nwait[sp] & 1); ++nwait[sp]; }
    issue_loads(it + GT_STAGES - 1);
  }
  if (niter > 0) { const int sl = (niter - 1) % GT_STAGES; mbar_wait(bars + sl, nwait[sl] & 1); ++nwait[sl]; }
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
  __syncwarp();
  tc_fence_after();

  // epilogue: thread = accumulator row (TMEM lane)
  const long long m = m0 + warp * 32 + lane;
  if (niter > 0) {
#pragma unroll 1
    for (int c0 = 0; c0 < GT_BN; c0 += 16) {
      if (n0 + c0 >= g.N) break;
      uint32_t d[16];
      tc_ld16(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, d);
      tc_wait_ld();
      if (m < g.M) {
        float* crow = g.C + m * g.ldc + n0 + c0;
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          if (n0 + c0 + e < g.N) {
            const float v = __uint_as_float(d[e]);
            if (g.mode == 2) { if (v != 0.f) atomicAdd(crow + e, v); }
            else if (g.mode == 1) crow[e] += v;
            else crow[e] = v;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_d, GT_BN); }
}

// host-side launcher; returns false if the shapes do not meet the alignment the vector loads need
inline bool gemm_tc_launch(cudaStream_t st, const float* A, long long lda, int a_kc, const float* B, long long ldb, int b_kc,
                           float* C, long long ldc, long long M, int N, long long K, int mode) {
  if (M <= 0 || N <= 0 || K <= 0) return true;
  if (a_kc && ((lda & 3) || (reinterpret_cast<uintptr_t>(A) & 15))) return false;   // vector loads along K
  if (b_kc && ((ldb & 3) || (reinterpret_cast<uintptr_t>(B) & 15))) return false;
  GemmTcArgs g;
  g.A = A; g.lda = lda; g.a_kc = a_kc; g.B = B; g.ldb = ldb; g.b_kc = b_kc; g.C = C; g.ldc = ldc; g.M = M; g.N = N; g.K = K;
  g.mode = mode;
  const long long tiles = ((M + GT_BM - 1) / GT_BM) * ((N + GT_BN - 1) / GT_BN);
  long long split = 1;
  if (mode == 2) {               // split the reduction so that ~2 CTAs per SM are in flight
    split = (296 + tiles - 1) / tiles;
    const long long maxsplit = (K + 4 * GT_BK - 1) / (4 * GT_BK);
    if (split > maxsplit) split = maxsplit;
    if (split < 1) split = 1;
  }
  g.kchunk = ((K + split - 1) / split + GT_BK - 1) / GT_BK * GT_BK;
  split = (K + g.kchunk - 1) / g.kchunk;
  dim3 grid((unsigned)((N + GT_BN - 1) / GT_BN), (unsigned)((M + GT_BM - 1) / GT_BM), (unsigned)split);
  RNNWF_COUNT_LAUNCH(); gemm_tc_kernel<<<grid, 128, GT_STAGES * 2 * GT_TILE_BYTES, st>>>(g);
  return true;
}

}  // namespace rnnwf
