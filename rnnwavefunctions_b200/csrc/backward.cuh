// Backward of log psi: hand-written back-propagation through the lattice sweep.
//
// grad[k] = sum_b seed_logp[b] dlogP_b/dtheta_k + seed_phase[b] dphase_b/dtheta_k  -- what TF autodiff of
// the reference's surrogate costs yields (1DTFIM/TrainingRNN_1DTFIM.py:157-161, J1J2/TrainingRNN_J1J2.py:197-200,
// 2DHeisenberg/Run_RNNWF_2DHeis.py:199-217, RydbergKagome_2DGRU/Run_RNNWF_Rydberg.py:158-176).
//
// The sweep is walked in reverse generation order over the saved hidden-state history and gate
// activations of the samples (written by the forward sweep).  Per step: head backward (softmax / U(1)
// mask / phase), cell non-linearity backward (element-wise), then the three GEMM shapes of a dense layer
// backward, written here as plain FP32 FFMA tiles: dX = dY W^T (gemm_nt), dX = dY W (gemm_nn) and
// dW += X^T dY (gemm_tn_atomic, split over the batch).  The operands are the reference-form ones (the
// expanded outer product h (x) x of the tensorized cells, the concatenated [x, h] of the GRUs), so the
// weight gradients land directly in the TF parameter layout.  Backward runs on `numsamples` rows only,
// i.e. ~1-3% of the rows a VMC step pushes through the forward sweep for its local energies.
#pragma once
#include "gru1d_bwd.cuh"
#include <string>

#include "model.h"
#include "bwd_tc.cuh"
#include "dw_tc.cuh"
#include "bwd_tc4.cuh"

namespace rnnwf {
cudaError_t g1_bwd_launch(const G1BwdParams& B, int rg, int grid, cudaStream_t st);     // gru1d_inst.cu

// ------------------------------------------------------------------------------------------------
// generic FP32 GEMM tiles (64x64x16, 256 threads, 4x4 per thread)
// ------------------------------------------------------------------------------------------------
constexpr int GB = 64, GK = 16;

// C[M,N] (+)= A[M,K] * B[N,K]^T
__global__ void __launch_bounds__(256) gemm_nt_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                      int ldb, float* __restrict__ C, int ldc, int M, int N, int K, int beta) {
  __shared__ float As[GK][GB + 4], Bs[GK][GB + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const long long m0 = (long long)blockIdx.y * GB;
  const int n0 = blockIdx.x * GB;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += GK) {
    for (int i = threadIdx.x; i < GB * GK; i += 256) {
      const int r = i / GK, k = i % GK;
      As[k][r] = (m0 + r < M && k0 + k < K) ? A[(m0 + r) * lda + k0 + k] : 0.f;
      Bs[k][r] = (n0 + r < N && k0 + k < K) ? B[(long long)(n0 + r) * ldb + k0 + k] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      const long long m = m0 + ty * 4 + i;
      const int n = n0 + tx * 4 + j;
      if (m < M && n < N) { float* c = C + m * ldc + n; *c = beta ? (*c + acc[i][j]) : acc[i][j]; }
    }
}

// C[M,N] (+)= A[M,K] * B[K,N]
__global__ void __launch_bounds__(256) gemm_nn_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                      int ldb, float* __restrict__ C, int ldc, int M, int N, int K, int beta) {
  __shared__ float As[GK][GB + 4], Bs[GK][GB + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const long long m0 = (long long)blockIdx.y * GB;
  const int n0 = blockIdx.x * GB;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += GK) {
    for (int i = threadIdx.x; i < GB * GK; i += 256) {
      const int r = i / GK, k = i % GK;
      As[k][r] = (m0 + r < M && k0 + k < K) ? A[(m0 + r) * lda + k0 + k] : 0.f;
      const int kk = i / GB, c = i % GB;
      Bs[kk][c] = (k0 + kk < K && n0 + c < N) ? B[(long long)(k0 + kk) * ldb + n0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      const long long m = m0 + ty * 4 + i;
      const int n = n0 + tx * 4 + j;
      if (m < M && n < N) { float* c = C + m * ldc + n; *c = beta ? (*c + acc[i][j]) : acc[i][j]; }
    }
}

// C[K1,N] += A[M,K1]^T * B[M,N]; blockIdx.z splits the batch dimension M, partial tiles are added atomically
// (rows of A may come in blocks: row m lives at A + (m / ablk) * astride + (m % ablk) * lda -- the saved history of one
// layer of a stacked model is such a set; ablk >= M makes it an ordinary matrix)
__device__ __forceinline__ void gemm_tn_atomic_body(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                    int ldb, float* __restrict__ C, int ldc, long long M, int K1, int N,
                                                    long long mchunk, long long ablk, long long astride, int bx, int by, int bz) {
  __shared__ float As[GK][GB + 4], Bs[GK][GB + 4];
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int k0 = by * GB, n0 = bx * GB;
  const long long mlo = (long long)bz * mchunk;
  const long long mhi = (mlo + mchunk < M) ? mlo + mchunk : M;
  float acc[4][4] = {};
  for (long long m0 = mlo; m0 < mhi; m0 += GK) {
    for (int i = threadIdx.x; i < GB * GK; i += 256) {
      const int mm = i / GB, c = i % GB;
      const bool mok = m0 + mm < mhi;
      const long long ma = m0 + mm;
      As[mm][c] = (mok && k0 + c < K1) ? A[(ma / ablk) * astride + (ma % ablk) * lda + k0 + c] : 0.f;
      Bs[mm][c] = (mok && n0 + c < N) ? B[(m0 + mm) * ldb + n0 + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < GK; ++k) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { a[i] = As[k][ty * 4 + i]; b[i] = Bs[k][tx * 4 + i]; }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      const int kk = k0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (kk < K1 && n < N && acc[i][j] != 0.f) atomicAdd(C + (long long)kk * ldc + n, acc[i][j]);
    }
}
__global__ void __launch_bounds__(256) gemm_tn_atomic_kernel(const float* __restrict__ A, int lda, const float* __restrict__ B,
                                                             int ldb, float* __restrict__ C, int ldc, long long M, int K1, int N,
                                                             long long mchunk, long long ablk, long long astride) {
  gemm_tn_atomic_body(A, lda, B, ldb, C, ldc, M, K1, N, mchunk, ablk, astride, blockIdx.x, blockIdx.y, blockIdx.z);
}
// Several such products (and column sums) in ONE launch: the weight / bias gradients of the small 1D models are a dozen
// ~50 us GEMMs over the same 10^4 rows -- launch latency, not work.  A block finds its problem by a scan over the list.
constexpr int GEMM_BATCH_MAX = 24;
struct GemmTnDesc { const float* A; const float* B; float* C; long long M, mchunk, ablk, astride; int lda, ldb, ldc, K1, N, gx, gy, gz, first; };
struct GemmTnBatch { GemmTnDesc d[GEMM_BATCH_MAX]; int n; };
__global__ void __launch_bounds__(256) gemm_tn_batched_kernel(const GemmTnBatch T) {
  int i = 0;
  while (i + 1 < T.n && (int)blockIdx.x >= T.d[i + 1].first) ++i;
  const GemmTnDesc& D = T.d[i];
  const int b = (int)blockIdx.x - D.first;
  gemm_tn_atomic_body(D.A, D.lda, D.B, D.ldb, D.C, D.ldc, D.M, D.K1, D.N, D.mchunk, D.ablk, D.astride, b % D.gx, (b / D.gx) % D.gy, b / (D.gx * D.gy));
}

// out[n] += sum_m X[m, n]   (block = 32 columns x 8 row lanes, 4 loads in flight per thread)
__device__ __forceinline__ void colsum_atomic_body(const float* __restrict__ X, int ldx, long long M, int N, float* __restrict__ out,
                                                   int bx, int by, int gy) {
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int n = bx * 32 + tx;
  const long long stride = 8LL * gy;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (n < N) {
    long long m = (long long)by * 8 + ty;
    for (; m + 3 * stride < M; m += 4 * stride) {
      s0 += X[m * ldx + n]; s1 += X[(m + stride) * ldx + n]; s2 += X[(m + 2 * stride) * ldx + n]; s3 += X[(m + 3 * stride) * ldx + n];
    }
    for (; m < M; m += stride) s0 += X[m * ldx + n];
  }
  red[ty][tx] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (ty == 0 && n < N) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][tx];
    if (s != 0.f) atomicAdd(out + n, s);
  }
}
__global__ void __launch_bounds__(256) colsum_atomic_kernel(const float* __restrict__ X, int ldx, long long M, int N, float* __restrict__ out) {
  colsum_atomic_body(X, ldx, M, N, out, blockIdx.x, blockIdx.y, gridDim.y);
}
struct ColsumDesc { const float* X; float* out; long long M; int ldx, N, gx, gy, first; };
struct ColsumBatch { ColsumDesc d[GEMM_BATCH_MAX]; int n; };
__global__ void __launch_bounds__(256) colsum_batched_kernel(const ColsumBatch T) {
  int i = 0;
  while (i + 1 < T.n && (int)blockIdx.x >= T.d[i + 1].first) ++i;
  const ColsumDesc& D = T.d[i];
  const int b = (int)blockIdx.x - D.first;
  colsum_atomic_body(D.X, D.ldx, D.M, D.N, D.out, b % D.gx, b / D.gx, D.gy);
}

// ------------------------------------------------------------------------------------------------
// per-step element-wise kernels
// ------------------------------------------------------------------------------------------------
struct BwdStep {
  int q, t, cell, nh, nhp, Gp, D, ND, complex_head, u1, cap_up, cap_dn, bits, nsites, Tsym, L;
  int nbr[MAX_SLOTS], nsym_site[MAX_SLOTS];  // history index / configuration offset of each slot (-1 = none)
  int sym_site;                              // configuration offset of this step's own symbol
  long long R;
  const int8_t* cfg;
  const float* hist;      // [Q][R][nhp]
  const float4* gates;    // [Q][R][Gp]
  float* dh;              // [Q][R][nhp]
};

__device__ __forceinline__ int cfg_symbol(const int8_t* row, int site, int bits) {
  if (bits == 1) return row[site] & 1;
  return (row[site * 3] & 1) | ((row[site * 3 + 1] & 1) << 1) | ((row[site * 3 + 2] & 1) << 2);
}

// One warp per row: recompute the heads from the saved hidden state, form d(logits), push it into dh.
__global__ void head_bwd_kernel(const BwdStep S, const int16_t* __restrict__ site_of_t, const float* __restrict__ head_w,
                                const float* __restrict__ head_b, const double* __restrict__ seed_logp,
                                const double* __restrict__ seed_phase, float* __restrict__ dlog) {
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= S.R) return;
  const float* h = S.hist + ((long long)S.q * S.R + row) * S.nhp;
  float part[8];
#pragma unroll
  for (int d = 0; d < 8; ++d) part[d] = 0.f;
  for (int j = lane; j < S.nh; j += 32) {
    const float hv = h[j];
#pragma unroll
    for (int d = 0; d < 8; ++d) if (d < S.ND) part[d] = fmaf(hv, head_w[j * S.ND + d], part[d]);
  }
#pragma unroll
  for (int d = 0; d < 8; ++d)
    if (d < S.ND)
      for (int o = 16; o > 0; o >>= 1) part[d] += __shfl_xor_sync(0xffffffffu, part[d], o);
  const int8_t* crow = S.cfg + row * S.nsites;
  const int s = cfg_symbol(crow, S.sym_site, S.bits);
  float p[8];
  float mx = -INFINITY;
#pragma unroll
  for (int d = 0; d < 8; ++d) if (d < S.D) { p[d] = part[d] + head_b[d]; mx = fmaxf(mx, p[d]); } else p[d] = 0.f;
  float ssum = 0.f;
#pragma unroll
  for (int d = 0; d < 8; ++d) if (d < S.D) { p[d] = expf(p[d] - mx); ssum += p[d]; }
#pragma unroll
  for (int d = 0; d < 8; ++d) p[d] /= ssum;
  if (S.u1) {
    int nup = 0;
    for (int tt = lane; tt < S.t; tt += 32) nup += crow[site_of_t[tt]] & 1;
    for (int o = 16; o > 0; o >>= 1) nup += __shfl_xor_sync(0xffffffffu, nup, o);
    const int ndn = S.t - nup;
    p[0] *= (S.cap_dn - ndn >= 0) ? 1.f : 0.f;
    p[1] *= (S.cap_up - nup >= 0) ? 1.f : 0.f;
    const float nrm = p[0] + p[1];
    p[0] /= nrm;
    p[1] /= nrm;
  }
  const float sl = (float)seed_logp[row];
  float dl[8];
#pragma unroll
  for (int d = 0; d < 8; ++d) dl[d] = (d < S.D) ? sl * (((d == s) ? 1.f : 0.f) - p[d]) : 0.f;
  if (S.complex_head) {
    const float sp = (seed_phase != nullptr) ? (float)seed_phase[row] : 0.f;
    const float x = ((s == 0) ? part[2] : part[3]) + ((s == 0) ? head_b[2] : head_b[3]);
    const float den = 1.f + fabsf(x);
    const float g = sp * 3.14159265358979f / (den * den);
    dl[2] = (s == 0) ? g : 0.f;
    dl[3] = (s == 1) ? g : 0.f;
  }
  if (lane == 0)
    for (int d = 0; d < S.ND; ++d) dlog[row * S.ND + d] = dl[d];
  float* dh = S.dh + ((long long)S.q * S.R + row) * S.nhp;
  for (int j = lane; j < S.nh; j += 32) {
    float a = 0.f;
#pragma unroll
    for (int d = 0; d < 8; ++d) if (d < S.ND) a = fmaf(head_w[j * S.ND + d], dl[d], a);
    dh[j] += a;
  }
}

// d(pre-activations) of the cell from d(new state); blocks of `dpre`: each [R][nh] (GRU1D block 0 is [R][2nh])
__global__ void cell_bwd_kernel(const BwdStep S, float* __restrict__ dp0, float* __restrict__ dp1, float* __restrict__ dp2) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= S.R * S.nh) return;
  const long long row = idx / S.nh;
  const int j = (int)(idx - row * S.nh);
  const float dhp = S.dh[((long long)S.q * S.R + row) * S.nhp + j];
  if (S.cell == CELL_TENSGRU2D || S.cell == CELL_GRU4) {
    const float4 g = S.gates[((long long)S.q * S.R + row) * S.Gp + j];  // (u, ht, m, -)
    dp0[idx] = dhp * g.x * (1.f - g.y * g.y);
    dp1[idx] = dhp * (g.y - g.z) * g.x * (1.f - g.x);
    dp2[idx] = dhp * (1.f - g.x);
  } else if (S.cell == CELL_GRU1D) {
    const float4 g = S.gates[((long long)S.q * S.R + row) * S.Gp + j];  // (r, u, c, hh)
    float hown = 0.f;
    if (S.nbr[0] >= 0) {
      hown = S.hist[((long long)S.nbr[0] * S.R + row) * S.nhp + j];
      S.dh[((long long)S.nbr[0] * S.R + row) * S.nhp + j] += dhp * g.y;
    }
    const float du = dhp * (hown - g.z), dc = dhp * (1.f - g.y);
    const float dcp = dc * (1.f - g.z * g.z);
    dp0[row * 2 * S.nh + j] = dcp * g.w * g.x * (1.f - g.x);   // d r_pre
    dp0[row * 2 * S.nh + S.nh + j] = du * g.y * (1.f - g.y);   // d u_pre
    dp1[idx] = dcp;                                            // d (x Wci + bci)
    dp2[idx] = dcp * g.x;                                      // d (h Wch + bch)
  } else {
    const float hv = S.hist[((long long)S.q * S.R + row) * S.nhp + j];
    dp0[idx] = (S.cell == CELL_VANILLA2D) ? dhp * (hv > 0.f ? 1.f : hv + 1.f) : dhp * (1.f - hv * hv);
  }
}

// Tensorized cells, d(pre-activations) + the operand of the linearised input-gradient GEMM:
//   P[row] = [ds, cl*ds, cu*ds, dg, cl*dg, cu*dg, dm]   (cl, cu = left / up neighbour spin, 0 if absent)
// so that  d[h_h ; h_v] = P * Wstack  with  Wstack = [Wa; Wb; Wc; Wga; Wgb; Wgc; Wmerge^T]  (see sweep_tc.cuh).
__global__ void cell_bwd_tens_kernel(const BwdStep S, int gru, float* __restrict__ dp0, float* __restrict__ dp1,
                                     float* __restrict__ dp2, float* __restrict__ Pm) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= S.R * S.nh) return;
  const long long row = idx / S.nh;
  const int j = (int)(idx - row * S.nh);
  const int pw = (gru ? 7 : 3) * S.nh;
  const float dhp = S.dh[((long long)S.q * S.R + row) * S.nhp + j];
  const int8_t* crow = S.cfg + row * S.nsites;
  const float cl = (S.nsym_site[0] >= 0) ? (float)(crow[S.nsym_site[0]] & 1) : 0.f;
  const float cu = (S.nsym_site[1] >= 0) ? (float)(crow[S.nsym_site[1]] & 1) : 0.f;
  float* Pr = Pm + row * pw;
  if (gru) {
    const float4 g = S.gates[((long long)S.q * S.R + row) * S.Gp + j];  // (u, ht, m, -)
    const float ds = dhp * g.x * (1.f - g.y * g.y);
    const float dg = dhp * (g.y - g.z) * g.x * (1.f - g.x);
    const float dm = dhp * (1.f - g.x);
    dp0[idx] = ds; dp1[idx] = dg; dp2[idx] = dm;
    Pr[j] = ds; Pr[S.nh + j] = cl * ds; Pr[2 * S.nh + j] = cu * ds;
    Pr[3 * S.nh + j] = dg; Pr[4 * S.nh + j] = cl * dg; Pr[5 * S.nh + j] = cu * dg;
    Pr[6 * S.nh + j] = dm;
  } else {
    const float hv = S.hist[((long long)S.q * S.R + row) * S.nhp + j];
    const float ds = dhp * (1.f - hv * hv);
    dp0[idx] = ds;
    Pr[j] = ds; Pr[S.nh + j] = cl * ds; Pr[2 * S.nh + j] = cu * ds;
  }
}

// WstackT[variant][i (2nh)][c (nblk*nh)], K-contiguous for the gemm "nt" form.  variant: 0 both neighbours,
// 1 horizontal only, 2 vertical only (same definitions as pack_tc_kernel).
__global__ void pack_wstack_kernel(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                                   int nh, int gru, float* __restrict__ out) {
  const int nblk = gru ? 7 : 3, pw = nblk * nh;
  const long long total = 3LL * 2 * nh * pw;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(idx % pw);
    const int i = (int)((idx / pw) % (2 * nh));
    const int variant = (int)(idx / ((long long)pw * 2 * nh));
    const int blk = c / nh, j = c - blk * nh;
    float v;
    if (blk == 6) {
      v = Wmerge[(long long)i * nh + j];
    } else {
      const float* w = ((blk < 3) ? W : Wg) + ((long long)j * 2 * nh + i) * 4;
      const int kind = blk % 3;
      if (variant == 0) v = (kind == 0) ? (w[0] + w[2]) : (kind == 1 ? (w[1] - w[0]) : (w[3] - w[2]));
      else if (variant == 1) v = (kind == 0) ? w[0] : (kind == 1 ? (w[1] - w[0]) : 0.f);
      else v = (kind == 0) ? w[2] : (kind == 1 ? 0.f : (w[3] - w[2]));
    }
    out[idx] = v;
  }
}

// dh[neighbour s][row][k] += dhc[row][s*nh + k]
__global__ void scatter_hc_kernel(const BwdStep S, const float* __restrict__ dhc) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int nh2 = 2 * S.nh;
  if (idx >= S.R * nh2) return;
  const long long row = idx / nh2;
  const int i = (int)(idx - row * nh2);
  const int slot = i / S.nh, k = i - slot * S.nh;
  if (S.nbr[slot] < 0) return;
  S.dh[((long long)S.nbr[slot] * S.R + row) * S.nhp + k] += dhc[idx];
}

// tensorized cells: T[row][i*4+k] = hc_i * xc_k and hc[row][i]
__global__ void build_tens_kernel(const BwdStep S, float* __restrict__ T, float* __restrict__ hc) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int nh2 = 2 * S.nh;
  if (idx >= S.R * nh2) return;
  const long long row = idx / nh2;
  const int i = (int)(idx - row * nh2);
  const int slot = i / S.nh, k = i - slot * S.nh;
  const float hv = (S.nbr[slot] >= 0) ? S.hist[((long long)S.nbr[slot] * S.R + row) * S.nhp + k] : 0.f;
  const int8_t* crow = S.cfg + row * S.nsites;
  float xc[4] = {0.f, 0.f, 0.f, 0.f};
  if (S.nsym_site[0] >= 0) xc[crow[S.nsym_site[0]] & 1] = 1.f;
  if (S.nsym_site[1] >= 0) xc[2 + (crow[S.nsym_site[1]] & 1)] = 1.f;
  hc[idx] = hv;
  *reinterpret_cast<float4*>(T + idx * 4) = make_float4(hv * xc[0], hv * xc[1], hv * xc[2], hv * xc[3]);
}

__global__ void scatter_tens_kernel(const BwdStep S, const float* __restrict__ dT, const float* __restrict__ dhc) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int nh2 = 2 * S.nh;
  if (idx >= S.R * nh2) return;
  const long long row = idx / nh2;
  const int i = (int)(idx - row * nh2);
  const int slot = i / S.nh, k = i - slot * S.nh;
  if (S.nbr[slot] < 0) return;
  const int8_t* crow = S.cfg + row * S.nsites;
  const float4 d = *reinterpret_cast<const float4*>(dT + idx * 4);
  float v = (dhc != nullptr) ? dhc[idx] : 0.f;
  if (S.nsym_site[0] >= 0) v += (crow[S.nsym_site[0]] & 1) ? d.y : d.x;
  if (S.nsym_site[1] >= 0) v += (crow[S.nsym_site[1]] & 1) ? d.w : d.z;
  S.dh[((long long)S.nbr[slot] * S.R + row) * S.nhp + k] += v;
}

// Rydberg cell: z[row] = [x0..x3 (one-hots of 3 bits, 6 wide each), h0..h3]
__global__ void build_gru4_kernel(const BwdStep S, float* __restrict__ z) {
  const int Kz = 24 + 4 * S.nh;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= S.R * Kz) return;
  const long long row = idx / Kz;
  const int c = (int)(idx - row * Kz);
  float v = 0.f;
  if (c < 24) {
    const int slot = c / 6, b = (c % 6) >> 1, bit = c & 1;
    if (S.nsym_site[slot] >= 0) v = ((S.cfg[row * S.nsites + S.nsym_site[slot] * 3 + b] & 1) == bit) ? 1.f : 0.f;
  } else {
    const int slot = (c - 24) / S.nh, k = (c - 24) - slot * S.nh;
    if (S.nbr[slot] >= 0) v = S.hist[((long long)S.nbr[slot] * S.R + row) * S.nhp + k];
  }
  z[idx] = v;
}

__global__ void scatter_gru4_kernel(const BwdStep S, const float* __restrict__ dz) {
  const int Kz = 24 + 4 * S.nh, nh4 = 4 * S.nh;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= S.R * nh4) return;
  const long long row = idx / nh4;
  const int i = (int)(idx - row * nh4);
  const int slot = i / S.nh, k = i - slot * S.nh;
  if (S.nbr[slot] < 0) return;
  // atomic: on lattices with Lx <= 2 or Ly <= 2 two slots of a cell name the same neighbour (its periodic successor is its
  // predecessor), i.e. two threads add into the same element
  atomicAdd(&S.dh[((long long)S.nbr[slot] * S.R + row) * S.nhp + k], dz[row * Kz + 24 + i]);
}

// one-hot inputs of the two-neighbour / chain models: X[row][2*slot + v]
__global__ void build_onehot_kernel(const BwdStep S, int nslots, float* __restrict__ X) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= S.R * nslots) return;
  const long long row = idx / nslots;
  const int slot = (int)(idx - row * nslots);
  float a = 0.f, b = 0.f;
  if (S.nsym_site[slot] >= 0) { if (S.cfg[row * S.nsites + S.nsym_site[slot]] & 1) b = 1.f; else a = 1.f; }
  X[idx * 2] = a;
  X[idx * 2 + 1] = b;
}

// ------------------------------------------------------------------------------------------------
// orchestration
// ------------------------------------------------------------------------------------------------
inline cudaError_t backward_init() {
  return cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GT_STAGES * 2 * GT_TILE_BYTES);
}
inline int dw_smem_bytes(int nhp) { return 2 * tc_himg_floats(nhp) * 4 + 2 * nhp * DW_APITCH * 4 + 256; }
inline cudaError_t backward_init_tc(const rnnwf_model_impl& m) {
  cudaError_t e = cudaFuncSetAttribute(bwd_tc_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, BtSmem(m.nhp, m.Tsym, m.Q).total);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(bwd_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, BtSmem(m.nhp, m.Tsym, m.Q).total);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(dw_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, dw_smem_bytes(m.nhp));
}

struct BwdLayout { long long off_dh, off_dp, off_Z, off_dZ, off_hc, off_dhc, off_dlog, off_X, off_wstack, off_wimg, off_dpT, off_jobs, off_cls,
                   off_t4_wt, off_t4_dimg, off_t4_himg, off_t4_ximg, off_t4_dlimg, total; long long dpw; int Kz; bool batched; bool batched1d; };
inline cudaError_t backward_init_tc4(const rnnwf_model_impl& m) {
  cudaError_t e = cudaFuncSetAttribute(bwd4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B4Smem(m.tc_np, m.Q).total);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(dw4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DW4_SMEM);
}
// Work plan of dw4_kernel: every output tile's (cell, tile) items split into contiguous ranges, ~3 jobs per SM in total.
inline void dw4_plan(const rnnwf_model_impl& m, long long R, std::vector<Dw4Job>& jobs, std::vector<int>& sites, int sites_n[5]) {
  const int tiles = (int)((R + 127) / 128), N4 = 64 * m.tc_np;
  sites.assign((size_t)5 * m.Q, 0);
  for (int s = 0; s < 5; ++s) {
    sites_n[s] = 0;
    for (int q = 0; q < m.Q; ++q)
      if (s == 4 || m.steps[q].nbr[s] >= 0) sites[(size_t)s * m.Q + sites_n[s]++] = q;
  }
  struct Out { int kind, slot, m0, blk; long long items; };
  std::vector<Out> outs;
  long long total = 0;
  for (int s = 0; s < 4; ++s)
    for (int m0 = 0; m0 < N4 && m0 < m.nh; m0 += 128)
      for (int blk = 0; blk < 3; ++blk) { outs.push_back({0, s, m0, blk, (long long)sites_n[s] * tiles}); total += outs.back().items; }
  for (int blk = 0; blk < 2; ++blk) { outs.push_back({1, 0, 0, blk, (long long)m.Q * tiles}); total += outs.back().items; }
  for (int m0 = 0; m0 < N4 && m0 < m.nh; m0 += 128) { outs.push_back({2, 0, m0, 0, (long long)m.Q * tiles}); total += outs.back().items; }   // head weights
  const long long target = std::max<long long>(8, (total + 3LL * m.num_sms - 1) / (3LL * m.num_sms));   // items per job
  jobs.clear();
  for (auto& o : outs) {
    if (o.items <= 0) continue;
    const long long parts = (o.items + target - 1) / target;
    for (long long p = 0; p < parts && jobs.size() < 4096; ++p) {
      const long long b = o.items * p / parts, e = o.items * (p + 1) / parts;
      if (e > b) jobs.push_back({(int16_t)o.kind, (int16_t)o.slot, (int16_t)o.m0, (int16_t)o.blk, (int32_t)b, (int32_t)(e - b)});
    }
  }
}

inline BwdLayout backward_layout(const rnnwf_model_impl& m, long long R) {
  auto al = [](long long x) { return (x + 255) & ~255LL; };
  BwdLayout L;
  const long long nh = m.nh;
  L.Kz = (m.cell == CELL_TENS2D || m.cell == CELL_TENSGRU2D) ? 8 * m.nh : (m.cell == CELL_GRU4 ? 24 + 4 * m.nh : 0);
  L.dpw = (m.cell == CELL_GRU1D) ? 4 * nh : 3 * nh;
  // Tensorized cells: the operands of the weight-gradient GEMMs (d pre-activations, h (x) x, hc, d logits) are kept for
  // ALL steps, so that dW, dWg, dWmerge, the head kernels and every bias come out of one long-K GEMM / reduction each
  // after the reverse sweep instead of ~10 small launches per lattice site.
  L.batched = (m.cell == CELL_TENS2D || m.cell == CELL_TENSGRU2D);
  // Single-layer 1D GRU: the same idea -- d(pre-activations), d(logits) and the one-hot inputs of ALL sites are kept, so every
  // weight / bias gradient is one launch after the reverse recurrence (4 launches per site instead of 15).
  L.batched1d = (m.cell == CELL_GRU1D && (m.symL == 1 || m.g1_bwd));
  const long long nb = (L.batched || L.batched1d) ? m.Q : 1;
  // tensor-core regime of the tensorized GRU: the reverse sweep / weight-gradient kernels have their own layouts
  // (d h tile-blocked, feature-major ds/dg/dm); none of the per-site FFMA-regime operands below is needed
  const bool tcs = m.tc_enabled;   // (tensorized GRU and vanilla 2D RNN both have the tensor-core backward)
  long long o = 0;
  L.off_dh = o;   o += al(tcs ? tc_dh_bytes(m.Q, R, (m.nhp + TC_JP - 1) / TC_JP) : (long long)m.Q * R * m.nhp * 4);
  L.off_dp = o;   o += al(tcs ? 0 : nb * R * L.dpw * 4);
  L.off_Z = o;    o += al(tcs ? 0 : nb * R * (long long)L.Kz * 4);
  L.off_dZ = o;   o += al(tcs ? 0 : R * (long long)L.Kz * 4);   // tensorized cells: the per-step linearised operand P [R][7nh] lives here
  L.off_hc = o;   o += al(tcs ? 0 : nb * R * 2 * nh * 4);
  L.off_dhc = o;  o += al(tcs ? 0 : R * 2 * nh * 4);
  L.off_dlog = o; o += al(((L.batched || L.batched1d || tcs || m.tc4) ? (long long)m.Q : 1LL) * R * 8 * 4);   // d(logits) of all sites in the batched / tensor-core paths
  L.off_X = o;    o += al((L.batched1d ? (long long)m.Q : 1LL) * R * 8 * 4);
  L.off_wstack = o; o += al(L.batched ? 3LL * 2 * nh * 7 * nh * 4 : 0);
  L.off_wimg = o;   o += al(tcs ? 3LL * ((m.nhp + TC_JP - 1) / TC_JP) * BtSmem(m.nhp, m.Tsym, m.Q).b_bytes : 0);   // Wstack^T chunks of the reverse sweep
  L.off_dpT = o;    o += al(tcs ? tc_dpt_bytes(m.Q, R, m.nhp) : 0);
  L.off_jobs = o; L.off_cls = o;   // (the work plan lives in the handle)
  {
    // four-neighbour GRU, tensor-core regime (bwd_tc4.cuh): transposed weight image and the BF16 images of every (cell, tile)
    const long long t4 = m.tc4 ? 1 : 0, tiles = (R + 127) / 128, NP = m.tc_np;
    L.off_t4_wt = o;   o += al(t4 * 4 * 3 * NP * (64 * NP) * 128);
    L.off_t4_dimg = o; o += al(t4 * m.Q * tiles * 3 * NP * (long long)B4_A_BYTES);
    L.off_t4_himg = o; o += al(t4 * m.Q * tiles * NP * (long long)B4_A_BYTES);
    L.off_t4_ximg = o; o += al(t4 * m.Q * tiles * 4 * 128 * 16);
    L.off_t4_dlimg = o; o += al(t4 * m.Q * tiles * 4096);
  }
  L.total = o;
  return L;
}

// Work split of the weight-gradient kernel: one job per CTA, each a contiguous range of (site, tile) items of one
// (source, neighbour half, neighbour class); about one job per SM, sized by cost (Wmerge jobs skip the masked blocks).
inline void dw_plan(const rnnwf_model_impl& m, long long R, std::vector<DwJob>& jobs, std::vector<int>& cls_steps, int cls_count[3]) {
  const long long tiles = (R + TC_ROWS - 1) / TC_ROWS;
  cls_steps.assign(3 * (size_t)m.Q, 0);
  cls_count[0] = cls_count[1] = cls_count[2] = 0;
  for (int q = 0; q < m.Q; ++q) {
    const bool h = m.steps[q].nbr[0] >= 0, u = m.steps[q].nbr[1] >= 0;
    const int c = (h && u) ? 0 : (h ? 1 : (u ? 2 : -1));
    if (c >= 0) cls_steps[(size_t)c * m.Q + cls_count[c]++] = q;
  }
  // One job per CTA and at most one CTA per SM.  The jobs of a neighbour class -- the sources (ds, dg, dm) times the halves --
  // get the same item ranges; CTAs per class: small search for the best balance.  (Measured dead end: pacing the jobs that
  // read the same tiles so that they share them through the L2 -- ncu shows 9 % L2 hits -- made the kernel slower whatever
  // the window: it is bound by the per-item chain of each CTA, not by the 4.2 TB/s of DRAM traffic it generates.)
  struct Cls { int cls; long long items; int members; int n; };
  std::vector<Cls> cl;
  const int src_lo = (m.cell == CELL_VANILLA2D) ? 2 : 0;                // vanilla cell: one unmasked source (d pre -> Wh, Wv)
  const int src_hi = (m.cell == CELL_TENS2D) ? 1 : 3;                   // un-gated tensorized cell: ds -> W only
  const int nsrc = src_hi - src_lo;
  for (int cls = 0; cls < 3; ++cls) {
    const long long items = (long long)cls_count[cls] * tiles;
    if (items > 0) cl.push_back({cls, items, nsrc * (cls == 0 ? 2 : 1), 1});   // class 0 has both halves
  }
  jobs.clear();
  if (cl.empty()) return;
  std::sort(cl.begin(), cl.end(), [](const Cls& a, const Cls& b) { return a.items > b.items; });
  {
    // the largest class takes what the others leave; the others are enumerated (at most two of them, up to 32 groups each)
    double best = 1e300;
    std::vector<int> bestn(cl.size(), 1);
    const int lim1 = cl.size() > 1 ? (int)std::min<long long>(32, cl[1].items) : 1, lim2 = cl.size() > 2 ? (int)std::min<long long>(32, cl[2].items) : 1;
    for (int n1 = 1; n1 <= lim1; ++n1)
      for (int n2 = 1; n2 <= lim2; ++n2) {
        int rest = m.num_sms;
        if (cl.size() > 1) rest -= n1 * cl[1].members;
        if (cl.size() > 2) rest -= n2 * cl[2].members;
        const int n0 = (int)std::min<long long>(cl[0].items, std::max(1, rest / cl[0].members));
        double worst = (double)cl[0].items / n0;
        if (cl.size() > 1) worst = std::max(worst, (double)cl[1].items / n1);
        if (cl.size() > 2) worst = std::max(worst, (double)cl[2].items / n2);
        const int used = n0 * cl[0].members + (cl.size() > 1 ? n1 * cl[1].members : 0) + (cl.size() > 2 ? n2 * cl[2].members : 0);
        if (used > std::max(m.num_sms, (int)(cl[0].members + (cl.size() > 1 ? cl[1].members : 0) + (cl.size() > 2 ? cl[2].members : 0)))) continue;
        if (worst < best) { best = worst; bestn[0] = n0; if (cl.size() > 1) bestn[1] = n1; if (cl.size() > 2) bestn[2] = n2; }
      }
    for (size_t k = 0; k < cl.size(); ++k) cl[k].n = bestn[k];
  }
  for (const Cls& g : cl)
    for (int k = 0; k < g.n; ++k) {
      const long long b = g.items * k / g.n, e = g.items * (k + 1) / g.n;
      for (int half = 0; half < 2; ++half) {
        if ((half == 0 && g.cls == 2) || (half == 1 && g.cls == 1)) continue;   // that neighbour is absent in this class
        for (int src = src_lo; src < src_hi; ++src) jobs.push_back({src, half, g.cls, (int)b, (int)(e - b)});
      }
    }
}

inline long long backward_workspace_bytes(const rnnwf_model_impl& m, long long R) { return backward_layout(m, R).total; }

struct Gemm {
  cudaStream_t st;
  bool tc = false;   // reduced-precision regime: long-K weight-gradient GEMMs on the tensor cores (TF32, gemm_tc.cuh)
  // deferred mode: tn / tn_blocked / colsum calls are collected and go out as ONE launch each at flush() (they only add into
  // disjoint or atomically updated outputs, so their order does not matter)
  bool defer = false;
  mutable GemmTnBatch tb{};
  mutable ColsumBatch cb{};
  mutable int tb_blocks = 0, cb_blocks = 0;
  void flush() const {
    if (tb.n > 0) { RNNWF_COUNT_LAUNCH(); gemm_tn_batched_kernel<<<tb_blocks, 256, 0, st>>>(tb); }
    if (cb.n > 0) { RNNWF_COUNT_LAUNCH(); colsum_batched_kernel<<<cb_blocks, 256, 0, st>>>(cb); }
    tb.n = 0; cb.n = 0; tb_blocks = 0; cb_blocks = 0;
  }
  // C[K1,N] += A[M,K1]^T B[M,N] with a long reduction M
  void tn_long(const float* A, int lda, const float* B, int ldb, float* C, int ldc, long long M, int K1, int N) const {
    if (tc && N >= 16 && K1 >= 16 && gemm_tc_launch(st, A, lda, 0, B, ldb, 0, C, ldc, K1, N, M, 2)) return;
    tn(A, lda, B, ldb, C, ldc, M, K1, N);
  }
  void nt(const float* A, int lda, const float* B, int ldb, float* C, int ldc, long long M, int N, int K, int beta) const {
    if (M <= 0 || N <= 0 || K <= 0) return;
    dim3 g((N + GB - 1) / GB, (unsigned)((M + GB - 1) / GB));
    RNNWF_COUNT_LAUNCH(); gemm_nt_kernel<<<g, 256, 0, st>>>(A, lda, B, ldb, C, ldc, (int)M, N, K, beta);
  }
  // C[M,N] (+)= A[M,K] B[N,K]^T; reduced-precision regime: TF32 tensor cores when the operands are aligned for it
  void nt_fast(const float* A, int lda, const float* B, int ldb, float* C, int ldc, long long M, int N, int K, int beta) const {
    if (tc && N >= 16 && K >= 16 && gemm_tc_launch(st, A, lda, 1, B, ldb, 1, C, ldc, M, N, K, beta ? 1 : 0)) return;
    nt(A, lda, B, ldb, C, ldc, M, N, K, beta);
  }
  void nn(const float* A, int lda, const float* B, int ldb, float* C, int ldc, long long M, int N, int K, int beta) const {
    if (M <= 0 || N <= 0 || K <= 0) return;
    dim3 g((N + GB - 1) / GB, (unsigned)((M + GB - 1) / GB));
    RNNWF_COUNT_LAUNCH(); gemm_nn_kernel<<<g, 256, 0, st>>>(A, lda, B, ldb, C, ldc, (int)M, N, K, beta);
  }
  void tn(const float* A, int lda, const float* B, int ldb, float* C, int ldc, long long M, int K1, int N) const {
    tn_blocked(A, lda, M > 0 ? M : 1, 0, B, ldb, C, ldc, M, K1, N);
  }
  // the same with A given as blocks of `ablk` rows, `astride` floats apart
  void tn_blocked(const float* A, int lda, long long ablk, long long astride, const float* B, int ldb, float* C, int ldc, long long M, int K1, int N) const {
    if (M <= 0 || N <= 0 || K1 <= 0) return;
    const int tiles = ((N + GB - 1) / GB) * ((K1 + GB - 1) / GB);
    long long split = (592 + tiles - 1) / tiles;                 // ~4 CTAs per SM in flight
    const long long maxsplit = (M + 255) / 256;
    if (split > maxsplit) split = maxsplit;
    if (split < 1) split = 1;
    long long mchunk = ((M + split - 1) / split + GK - 1) / GK * GK;
    split = (M + mchunk - 1) / mchunk;
    dim3 g((N + GB - 1) / GB, (K1 + GB - 1) / GB, (unsigned)split);
    if (defer) {
      if (tb.n == GEMM_BATCH_MAX) flush();
      tb.d[tb.n++] = GemmTnDesc{A, B, C, M, mchunk, ablk, astride, lda, ldb, ldc, K1, N, (int)g.x, (int)g.y, (int)g.z, tb_blocks};
      tb_blocks += (int)(g.x * g.y * g.z);
      return;
    }
    RNNWF_COUNT_LAUNCH(); gemm_tn_atomic_kernel<<<g, 256, 0, st>>>(A, lda, B, ldb, C, ldc, M, K1, N, mchunk, ablk, astride);
  }
  void colsum(const float* X, int ldx, long long M, int N, float* out) const {
    if (M <= 0 || N <= 0) return;
    dim3 g((N + 31) / 32, (unsigned)std::max<long long>(1, std::min<long long>(1024, M / 64)));
    if (defer) {
      if (cb.n == GEMM_BATCH_MAX) flush();
      cb.d[cb.n++] = ColsumDesc{X, out, M, ldx, N, (int)g.x, (int)g.y, cb_blocks};
      cb_blocks += (int)(g.x * g.y);
      return;
    }
    RNNWF_COUNT_LAUNCH(); colsum_atomic_kernel<<<g, 256, 0, st>>>(X, ldx, M, N, out);
  }
};

inline int backward_run(const rnnwf_model_impl& m, long long R, const int8_t* cfg, const void* saved,
                        const double* seed_logp, const double* seed_phase, float* grad, void* workspace,
                        cudaStream_t st, std::string& err) {
  if (R > 0x7fffffffLL / std::max(1, m.nhp)) { err = "rnnwf_backward: too many rows"; return RNNWF_EINVAL; }
  const Saved sv = m.saved_layout(R);
  const char* sb = static_cast<const char*>(saved);
  const float* hist = reinterpret_cast<const float*>(sb + sv.off_hist);
  const float4* gates = reinterpret_cast<const float4*>(sb + sv.off_gates);
  const BwdLayout L = backward_layout(m, R);
  char* ws = static_cast<char*>(workspace);
  float* dh = reinterpret_cast<float*>(ws + L.off_dh);
  float* dp = reinterpret_cast<float*>(ws + L.off_dp);
  float* Z_base = reinterpret_cast<float*>(ws + L.off_Z);
  float* dZ = reinterpret_cast<float*>(ws + L.off_dZ);
  float* hc_base = reinterpret_cast<float*>(ws + L.off_hc);
  float* dhc = reinterpret_cast<float*>(ws + L.off_dhc);
  float* dlog_base = reinterpret_cast<float*>(ws + L.off_dlog);
  float* X = reinterpret_cast<float*>(ws + L.off_X);
  const float* params = m.d_params;
  const int nh = m.nh, nhp = m.nhp;
  Gemm G{st};
  G.tc = (m.precision == 1);
  float* wstack = reinterpret_cast<float*>(ws + L.off_wstack);
  if (L.batched && !m.tc_enabled) {
    const bool gru0 = m.cell == CELL_TENSGRU2D;
    RNNWF_COUNT_LAUNCH(); pack_wstack_kernel<<<512, 256, 0, st>>>(params + m.tensor("W")->offset, gru0 ? params + m.tensor("Wg")->offset : nullptr,
                                                     gru0 ? params + m.tensor("Wmerge")->offset : nullptr, nh, gru0 ? 1 : 0, wstack);
  }

  cudaError_t e = (m.tc_enabled || m.g1_bwd) ? cudaSuccess : cudaMemsetAsync(dh, 0, (size_t)m.Q * R * nhp * 4, st);
  if (e == cudaSuccess) e = cudaMemsetAsync(grad, 0, (size_t)m.num_params * 4, st);
  if (e != cudaSuccess) { err = std::string("rnnwf_backward memset: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }

  auto P = [&](const std::string& n) { return params + m.tensor(n)->offset; };
  auto Gd = [&](const std::string& n) { return grad + m.tensor(n)->offset; };
  const int tpb = 256;
  auto nblk = [&](long long n) { return (unsigned)((n + tpb - 1) / tpb); };
  auto hrow = [&](int q) { return hist + (long long)q * R * nhp; };
  auto dhrow = [&](int q) { return dh + (long long)q * R * nhp; };

  if (m.g1_bwd) {
    // Small 1D models: the whole reverse recurrence is one launch (gru1d_bwd.cuh); it leaves d(pre-activations), d(logits)
    // and the one-hot inputs of every (layer, site, row) layer-major in the workspace, and every weight / bias gradient is
    // one GEMM / reduction over all sites afterwards.
    const int Ls = m.symL, Ns = m.Tsym;
    const long long NR = (long long)Ns * R, N1R = (long long)(Ns - 1) * R;
    G1BwdParams B;
    std::memset(&B, 0, sizeof(B));
    B.nh = nh; B.nhp = nhp; B.Gp = m.Gp; B.L = Ls; B.N = Ns; B.D = m.D; B.ND = m.ND; B.complex_head = m.complex_head; B.u1 = m.u1;
    B.cap_up = m.cap_up; B.cap_dn = m.cap_dn; B.R = R; B.cfg = cfg; B.hist = hist; B.gates = gates; B.seed_logp = seed_logp;
    B.seed_phase = seed_phase; B.packT = m.d_g1packT;
    B.dpG = dp; B.dpCI = dp + (long long)m.Q * R * 2 * nh; B.dpCH = dp + (long long)m.Q * R * 3 * nh; B.dlog = dlog_base; B.X = X;
    for (int l = Ls - 1; l >= 0; --l) {       // one launch per layer, top first; d h^{l-1}_t of all sites handed down through `dh`
      B.layer = l;
      B.dx_in = (l < Ls - 1) ? dh + (long long)((l + 1) & 1) * NR * nhp : nullptr;
      B.dx_out = (l > 0) ? dh + (long long)(l & 1) * NR * nhp : nullptr;
      RNNWF_COUNT_LAUNCH();
      const int rg = std::getenv("RNNWF_NO_VTILES") == nullptr ? g1_pick_rg(R, m.num_sms) : G1_RG;    // narrow tiles for a training batch
      e = g1_bwd_launch(B, rg, (int)std::min<long long>((R + 8 * rg - 1) / (8 * rg), 2LL * m.num_sms), st);
      if (e != cudaSuccess) { err = std::string("rnnwf_backward 1D GRU reverse sweep: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
    }
    const long long hstride = (long long)Ls * R * nhp;       // between consecutive sites of one layer in the saved history
    G.defer = std::getenv("RNNWF_NO_GEMM_BATCH") == nullptr;   // all weight / bias gradients below: one GEMM launch + one column-sum launch
    for (int l = 0; l < Ls; ++l) {
      const std::string p = "gru" + std::to_string(l) + ".";
      const int din = (l == 0) ? 2 : nh;
      const float* dG = B.dpG + (long long)l * NR * 2 * nh;
      const float* dCI = B.dpCI + (long long)l * NR * nh;
      const float* dCH = B.dpCH + (long long)l * NR * nh;
      G.colsum(dG, 2 * nh, NR, 2 * nh, Gd(p + "bg"));
      G.colsum(dCI, nh, NR, nh, Gd(p + "bci"));
      G.colsum(dCH, nh, NR, nh, Gd(p + "bch"));
      if (Ns > 1) {      // rows fed by the own previous state h^l_{t-1}, t >= 1
        G.tn_blocked(hist + (long long)l * R * nhp, nhp, R, hstride, dG + R * 2LL * nh, 2 * nh, Gd(p + "Wg") + (long long)din * 2 * nh, 2 * nh, N1R, nh, 2 * nh);
        G.tn_blocked(hist + (long long)l * R * nhp, nhp, R, hstride, dCH + R * (long long)nh, nh, Gd(p + "Wch"), nh, N1R, nh, nh);
      }
      if (l == 0) {      // rows fed by the one-hot of the previous symbol (zero vector at the first site)
        if (Ns > 1) {
          G.tn(X + R * 2LL, 2, dG + R * 2LL * nh, 2 * nh, Gd(p + "Wg"), 2 * nh, N1R, 2, 2 * nh);
          G.tn(X + R * 2LL, 2, dCI + R * (long long)nh, nh, Gd(p + "Wci"), nh, N1R, 2, nh);
        }
      } else {           // rows fed by the layer below at the same site
        G.tn_blocked(hist + (long long)(l - 1) * R * nhp, nhp, R, hstride, dG, 2 * nh, Gd(p + "Wg"), 2 * nh, NR, nh, 2 * nh);
        G.tn_blocked(hist + (long long)(l - 1) * R * nhp, nhp, R, hstride, dCI, nh, Gd(p + "Wci"), nh, NR, nh, nh);
      }
    }
    const float* htop = hist + (long long)(Ls - 1) * R * nhp;
    G.tn_blocked(htop, nhp, R, hstride, dlog_base, m.ND, Gd("dense.W"), m.D, NR, nh, m.D);
    G.colsum(dlog_base, m.ND, NR, m.D, Gd("dense.b"));
    if (m.complex_head) {
      G.tn_blocked(htop, nhp, R, hstride, dlog_base + m.D, m.ND, Gd("dense_phase.W"), 2, NR, nh, 2);
      G.colsum(dlog_base + m.D, m.ND, NR, 2, Gd("dense_phase.b"));
    }
    G.flush();
    G.defer = false;
    e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("rnnwf_backward kernels: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
    return RNNWF_OK;
  }

  // Reduced-precision regime, tensorized GRU: the whole recurrence (heads' push into d h, gate derivatives, input
  // gradients, scatter to the neighbours) is one persistent tensor-core kernel (bwd_tc.cuh); d(logits) of every site is
  // formed inside it from the heads' outputs the forward sweep saved.
  const bool tc_sweep = m.tc_enabled;
  if (tc_sweep) {
    const bool van = m.cell == CELL_VANILLA2D;
    const int pieces = (nhp + TC_JP - 1) / TC_JP;      // unit pieces per thread layout (== column passes of the tensorized GRU)
    e = cudaMemsetAsync(dh, 0, (size_t)tc_dh_bytes(m.Q, R, pieces), st);
    if (e != cudaSuccess) { err = std::string("rnnwf_backward memset: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
    // transposed weight image of the reverse sweep: the tensorized cells keep it in the handle (packed by rnnwf_set_weights together
    // with the forward images); the vanilla cell packs its single chunk here
    float* wimg = van ? reinterpret_cast<float*>(ws + L.off_wimg) : m.d_wimgT;
    if (van) { RNNWF_COUNT_LAUNCH(); pack_bwd_tc_vanilla_kernel<<<256, 256, 0, st>>>(P("Wh"), P("Wv"), nh, nhp, wimg); }
    BwdTcParams B;
    B.steps = m.d_steps; B.Q = m.Q; B.Tsym = m.Tsym; B.nsites = m.nsites; B.nh = nh; B.nhp = nhp; B.np = pieces; B.ND = m.ND;
    B.rows = R; B.cfg = cfg; B.gates = gates; B.dlog = dlog_base; B.dh = reinterpret_cast<float4*>(dh);
    B.sv_head = reinterpret_cast<const float4*>(sb + sv.off_head); B.seed_logp = seed_logp; B.seed_phase = seed_phase;
    B.dpT = reinterpret_cast<float*>(ws + L.off_dpT);
    B.wimg = wimg; B.head_w = m.d_head_w; B.hist = hist;
    { const char* dbg = std::getenv("RNNWF_TC_DEBUG"); B.debug = dbg ? std::atoi(dbg) : 0; }
    {
      // horizontal contributions stay on the SM when the horizontal neighbour is always the previous site of the path and no
      // site is the vertical neighbour of two sites (every 2D zig-zag lattice)
      bool chain = true;
      std::vector<int> hits(m.Q, 0);
      for (int q = 0; q < m.Q; ++q) {
        if (m.steps[q].nbr[0] >= 0 && m.steps[q].nbr[0] != q - 1) chain = false;
        if (m.steps[q].nbr[1] >= 0 && ++hits[m.steps[q].nbr[1]] > 1) chain = false;
      }
      B.h_chain = (chain && std::getenv("RNNWF_NO_HCHAIN") == nullptr) ? 1 : 0;
    }
    const long long tiles = (R + TC_ROWS - 1) / TC_ROWS;
    const BtSmem bs(nhp, m.Tsym, m.Q);
    // (the reverse sweep is bound by its per-site chain and the 128-row MMAs: spreading pays once a tile would be at most half live)
    B.tile_rows = tc_tile_rows(R, m.num_sms) <= TC_ROWS / 2 ? tc_tile_rows(R, m.num_sms) : TC_ROWS;
    const long long vtiles = (tiles * TC_ROWS + B.tile_rows - 1) / B.tile_rows;
    RNNWF_COUNT_LAUNCH();
    if (van) bwd_tc_kernel<1><<<(unsigned)std::min<long long>(vtiles, m.num_sms), BT_THREADS, bs.total, st>>>(B);
    else bwd_tc_kernel<0><<<(unsigned)std::min<long long>(vtiles, m.num_sms), BT_THREADS, bs.total, st>>>(B);
    // ---- parameter gradients: weights on the tensor cores (dw_tc.cuh), biases and heads as reductions
    {
      DwParams D;
      if (m.dw_rows != R) {             // plan depends on the row count only: build and upload once, ordered on the caller's stream
        std::vector<DwJob> jobs;
        std::vector<int> cls_steps;
        dw_plan(m, R, jobs, cls_steps, m.dw_cls_count);
        if (jobs.size() > 1024) { err = "rnnwf_backward: too many weight-gradient jobs"; return RNNWF_EINVAL; }
        // the host copies live in the handle until the next plan, so the stream-ordered upload never outlives its source
        m.plan_host_a.assign(reinterpret_cast<const unsigned char*>(jobs.data()), reinterpret_cast<const unsigned char*>(jobs.data() + jobs.size()));
        m.plan_host_b.assign(reinterpret_cast<const unsigned char*>(cls_steps.data()), reinterpret_cast<const unsigned char*>(cls_steps.data() + cls_steps.size()));
        e = cudaMemcpyAsync(m.d_dw_jobs, m.plan_host_a.data(), m.plan_host_a.size(), cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(m.d_dw_cls, m.plan_host_b.data(), m.plan_host_b.size(), cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) { err = std::string("rnnwf_backward plan upload: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
        m.dw_njobs = (int)jobs.size();
        m.dw_rows = R;
      }
      for (int k = 0; k < 3; ++k) D.cls_count[k] = m.dw_cls_count[k];
      const DwJob* d_jobs = reinterpret_cast<const DwJob*>(m.d_dw_jobs);
      const int* d_cls = m.d_dw_cls;
      const int njobs = m.dw_njobs;
      const float* himg = reinterpret_cast<const float*>(sb + sv.off_himg);
      D.jobs = d_jobs; D.cls_steps = d_cls; D.steps = m.d_steps; D.Q = m.Q; D.nh = nh; D.nhp = nhp; D.nsites = m.nsites;
      D.rows = R; D.tiles = tiles; D.cfg = cfg; D.site_of_t = m.d_site_of_t; D.dpT = B.dpT; D.himg = himg;
      { const char* dbg = std::getenv("RNNWF_TC_DEBUG"); D.debug = dbg ? std::atoi(dbg) : 0; }
      const int nsplit = 16;
      int rowsum_nq = 0;
      bool rowsum_gbg = false;
      if (van) {
        D.gW = nullptr; D.gWg = nullptr; D.gM0 = Gd("Wh"); D.gM1 = Gd("Wv");
        D.slab[0] = D.slab[1] = D.slab[2] = 0; D.gB = nullptr; D.gBg = nullptr;
        if (njobs > 0) { RNNWF_COUNT_LAUNCH(); dw_tc_kernel<<<(unsigned)njobs, DW_THREADS, dw_smem_bytes(nhp), st>>>(D); }   // (a single-site lattice has no neighbour weights)
        RNNWF_COUNT_LAUNCH(); vanilla_input_grads_kernel<<<dim3((nh + VIG_UB - 1) / VIG_UB, (unsigned)std::max<long long>(1, std::min<long long>((long long)m.Q * tiles, 4LL * m.num_sms / ((nh + VIG_UB - 1) / VIG_UB) + 1))), 128, 0, st>>>(B.dpT, m.d_steps, m.d_site_of_t, cfg, m.Q, tiles, R, nh, nhp,
                                                                                          m.nsites, Gd("b"), Gd("Uh"), Gd("Uv"));
      } else {
        const bool gated = m.cell == CELL_TENSGRU2D;
        D.gW = Gd("W"); D.gWg = gated ? Gd("Wg") : nullptr;
        D.gM0 = gated ? Gd("Wmerge") : nullptr; D.gM1 = gated ? Gd("Wmerge") + (long long)nh * nh : nullptr;
        D.slab[0] = 0; D.slab[1] = 1; D.slab[2] = 2;
        // biases: a free padding position of the hidden-state image holds a constant 1 (sweep_tc_kernel.cuh), so the sums over
        // the rows come out of the weight-gradient MMAs for every site with a neighbour; the others (site 0) are summed directly
        const bool bias_mma = nh < tc_himg_ncores(nhp) * 8 && njobs > 0 && std::getenv("RNNWF_NO_BIAS_MMA") == nullptr;
        int nq = m.Q;
        if (bias_mma) { nq = 0; while (nq < m.Q && m.steps[nq].nbr[0] < 0 && m.steps[nq].nbr[1] < 0) ++nq; }
        for (int q = nq; bias_mma && q < m.Q; ++q)
          if (m.steps[q].nbr[0] < 0 && m.steps[q].nbr[1] < 0) { err = "rnnwf_backward: a site without neighbours inside the path"; return RNNWF_EINVAL; }
        D.gB = bias_mma ? Gd("b") : nullptr; D.gBg = (bias_mma && gated) ? Gd("bg") : nullptr;
        if (njobs > 0) { RNNWF_COUNT_LAUNCH(); dw_tc_kernel<<<(unsigned)njobs, DW_THREADS, dw_smem_bytes(nhp), st>>>(D); }   // (a single-site lattice has no neighbour weights)
        rowsum_nq = nq;                      // the sites without a neighbour: folded into head_grad_himg_kernel below ...
        if (nq > 0 && !bias_mma) {           // ... unless every site is summed directly (fallback path): its own launch
          const int ns2 = std::max(1, std::min<int>(nsplit, (int)std::min<long long>(1 << 20, (long long)nq * tiles)));
          RNNWF_COUNT_LAUNCH(); dpt_rowsum_kernel<<<dim3(nh, ns2, gated ? 2 : 1), 128, 0, st>>>(B.dpT, m.Q, nq, tiles, nhp, ns2, Gd("b"), gated ? Gd("bg") : nullptr);
          rowsum_nq = 0;
        }
        rowsum_gbg = gated;
      }
      RNNWF_COUNT_LAUNCH(); head_grad_himg_kernel<<<std::min<long long>((long long)m.Q * tiles, 4LL * m.num_sms), 128, 0, st>>>(
          himg, dlog_base, m.Q, tiles, R, nh, nhp, m.ND, Gd("dense.W"), m.complex_head ? Gd("dense_phase.W") : nullptr,
          (nh <= 124 && m.D == 2) ? Gd("dense.b") : nullptr, (nh <= 124 && m.D == 2 && m.complex_head) ? Gd("dense_phase.b") : nullptr,
          B.dpT, rowsum_nq, rowsum_nq > 0 ? Gd("b") : nullptr, (rowsum_nq > 0 && rowsum_gbg) ? Gd("bg") : nullptr);
      if (!(nh <= 124 && m.D == 2)) {      // (wide models: the heads' bias gradients as column sums of d(logits))
        const long long RQ = (long long)m.Q * R;
        G.colsum(dlog_base, m.ND, RQ, m.D, Gd("dense.b"));
        if (m.complex_head) G.colsum(dlog_base + m.D, m.ND, RQ, 2, Gd("dense_phase.b"));
      }
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("rnnwf_backward kernels: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
    return RNNWF_OK;
  }

  if (m.tc4) {
    // Four-neighbour GRU, tensor-core regime: the whole recurrence is one persistent kernel (bwd_tc4.cuh); the weight
    // gradients follow from the (ds, dg, dm) and d(logits) of all cells it leaves behind.
    const long long tiles = (R + 127) / 128, RQ = (long long)m.Q * R;
    unsigned char* wt = reinterpret_cast<unsigned char*>(ws + L.off_t4_wt);
    RNNWF_COUNT_LAUNCH(); pack_bwd4_kernel<<<512, 256, 0, st>>>(P("W"), P("Wg"), P("Wmerge"), nh, m.tc_np, reinterpret_cast<__nv_bfloat16*>(wt));
    Bwd4Params B;
    B.Q = m.Q; B.nh = nh; B.nhp = nhp; B.NP = m.tc_np; B.Gp = m.Gp; B.nsites = m.nsites; B.R = R;
    B.steps = m.d_steps; B.cfg = cfg; B.hist = hist; B.gates = gates; B.dh = dh; B.dp = nullptr; B.dlog = dlog_base;
    B.seed = seed_logp; B.head_w = m.d_head_w; B.head_b = m.d_head_b; B.Wt = wt;
    B.dimg = reinterpret_cast<unsigned char*>(ws + L.off_t4_dimg);
    B.himg = reinterpret_cast<unsigned char*>(ws + L.off_t4_himg);
    B.ximg = reinterpret_cast<unsigned char*>(ws + L.off_t4_ximg);
    B.dlimg = reinterpret_cast<unsigned char*>(ws + L.off_t4_dlimg);
    RNNWF_COUNT_LAUNCH(); bwd4_kernel<<<(unsigned)std::min<long long>(tiles, m.num_sms), B4_THREADS, B4Smem(m.tc_np, m.Q).total, st>>>(B);
    {
      // weight gradients: grouped GEMM over the images the reverse sweep left behind (work plan cached per row count)
      if (m.dw4_rows != R) {
        std::vector<Dw4Job> jobs;
        std::vector<int> sites;
        dw4_plan(m, R, jobs, sites, m.dw4_sites_n);
        m.plan_host_a.assign(reinterpret_cast<const unsigned char*>(jobs.data()), reinterpret_cast<const unsigned char*>(jobs.data() + jobs.size()));
        m.plan_host_b.assign(reinterpret_cast<const unsigned char*>(sites.data()), reinterpret_cast<const unsigned char*>(sites.data() + sites.size()));
        e = cudaMemcpyAsync(m.d_dw4_jobs, m.plan_host_a.data(), m.plan_host_a.size(), cudaMemcpyHostToDevice, st);
        if (e == cudaSuccess) e = cudaMemcpyAsync(m.d_dw4_sites, m.plan_host_b.data(), m.plan_host_b.size(), cudaMemcpyHostToDevice, st);
        if (e != cudaSuccess) { err = std::string("rnnwf_backward plan upload: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
        m.dw4_njobs = (int)jobs.size();
        m.dw4_rows = R;
      }
      Dw4Params D;
      D.Q = m.Q; D.nh = nh; D.NP = m.tc_np; D.tiles = (int)tiles; D.steps = m.d_steps;
      D.jobs = static_cast<const Dw4Job*>(m.d_dw4_jobs); D.sites = m.d_dw4_sites;
      for (int k = 0; k < 5; ++k) D.sites_n[k] = m.dw4_sites_n[k];
      D.dimg = B.dimg; D.himg = B.himg; D.ximg = B.ximg; D.dlimg = B.dlimg; D.gDense = Gd("dense.W");
      D.gW = Gd("W"); D.gWg = Gd("Wg"); D.gWmerge = Gd("Wmerge"); D.gb = Gd("b"); D.gbg = Gd("bg");
      if (m.dw4_njobs > 0) { RNNWF_COUNT_LAUNCH(); dw4_kernel<<<(unsigned)m.dw4_njobs, DW4_THREADS, DW4_SMEM, st>>>(D); }
    }
    G.colsum(dlog_base, m.ND, RQ, m.D, Gd("dense.b"));      // (dense.W is one more job type of dw4_kernel)
    e = cudaGetLastError();
    if (e != cudaSuccess) { err = std::string("rnnwf_backward kernels: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
    return RNNWF_OK;
  }

  for (int q = m.Q - 1; q >= 0; --q) {
    const StepMeta& sm = m.steps[q];
    BwdStep S;
    std::memset(&S, 0, sizeof(S));
    S.q = q; S.t = sm.sym; S.cell = m.cell; S.nh = nh; S.nhp = nhp; S.Gp = m.Gp; S.D = m.D; S.ND = m.ND;
    S.complex_head = m.complex_head; S.u1 = m.u1; S.cap_up = m.cap_up; S.cap_dn = m.cap_dn; S.bits = m.bits;
    S.nsites = m.nsites; S.Tsym = m.Tsym; S.L = m.symL; S.R = R; S.cfg = cfg; S.hist = hist; S.gates = gates; S.dh = dh;
    for (int k = 0; k < MAX_SLOTS; ++k) {
      S.nbr[k] = sm.nbr[k];
      S.nsym_site[k] = (sm.nsym[k] >= 0) ? m.site_of_t[sm.nsym[k]] : -1;
    }
    S.sym_site = sm.site;

    const bool batched = L.batched;
    const bool b1d = L.batched1d;
    float* dlog = (batched || b1d) ? dlog_base + (long long)q * R * m.ND : dlog_base;
    if (sm.head) {
      const int wpb = 8;
      RNNWF_COUNT_LAUNCH(); head_bwd_kernel<<<(unsigned)((R + wpb - 1) / wpb), wpb * 32, 0, st>>>(S, m.d_site_of_t, m.d_head_w, m.d_head_b, seed_logp, seed_phase, dlog);
      if (!batched && !b1d) {
        G.tn(hrow(q), nhp, dlog, m.ND, Gd("dense.W"), m.D, R, nh, m.D);
        G.colsum(dlog, m.ND, R, m.D, Gd("dense.b"));
        if (m.complex_head) {
          G.tn(hrow(q), nhp, dlog + m.D, m.ND, Gd("dense_phase.W"), 2, R, nh, 2);
          G.colsum(dlog + m.D, m.ND, R, 2, Gd("dense_phase.b"));
        }
      }
    }
    // batched layout: three arrays [Q][R*nh] (ds, dg, dm); otherwise one per-step block
    float* dp0 = batched ? dp + (long long)q * R * nh : dp;
    float* dp1 = batched ? dp + ((long long)m.Q + q) * R * nh : dp + ((m.cell == CELL_GRU1D) ? 2 * R * nh : R * nh);
    float* dp2 = batched ? dp + (2LL * m.Q + q) * R * nh : dp1 + R * nh;
    if (b1d) {   // three arrays over all sites: [Q][R][2nh] (d r_pre, d u_pre), [Q][R][nh] (d candidate input), [Q][R][nh] (d candidate hidden)
      dp0 = dp + (long long)q * R * 2 * nh;
      dp1 = dp + (long long)m.Q * R * 2 * nh + (long long)q * R * nh;
      dp2 = dp + (long long)m.Q * R * 3 * nh + (long long)q * R * nh;
    }
    float* Z = batched ? Z_base + (long long)q * R * L.Kz : Z_base;
    float* hc = batched ? hc_base + (long long)q * R * 2 * nh : hc_base;
    if (batched) { RNNWF_COUNT_LAUNCH(); cell_bwd_tens_kernel<<<nblk(R * nh), tpb, 0, st>>>(S, m.cell == CELL_TENSGRU2D ? 1 : 0, dp0, dp1, dp2, dZ); }
    else { RNNWF_COUNT_LAUNCH(); cell_bwd_kernel<<<nblk(R * nh), tpb, 0, st>>>(S, dp0, dp1, dp2); }

    switch (m.cell) {
      case CELL_TENS2D:
      case CELL_TENSGRU2D: {
        const bool gru = m.cell == CELL_TENSGRU2D;
        const int K8 = 8 * nh;
        if (sm.nbr[0] < 0 && sm.nbr[1] < 0) break;  // first site: no inputs (its bias gradient is in the batched sums)
        RNNWF_COUNT_LAUNCH(); build_tens_kernel<<<nblk(R * 2 * nh), tpb, 0, st>>>(S, Z, hc);   // operands of the batched dW GEMMs
        {
          // d[h_h ; h_v] = P * Wstack(variant): one GEMM instead of dT = ds W2 (+ dg Wg2) over the expanded h (x) x
          const int pw = (gru ? 7 : 3) * nh;
          const int variant = (sm.nbr[0] >= 0 && sm.nbr[1] >= 0) ? 0 : (sm.nbr[0] >= 0 ? 1 : 2);
          const float* WsT = wstack + (long long)variant * 2 * nh * pw;
          if (!(G.tc && gemm_tc_launch(st, dZ, pw, 1, WsT, pw, 1, dhc, 2 * nh, R, 2 * nh, pw, 0)))
            G.nt(dZ, pw, WsT, pw, dhc, 2 * nh, R, 2 * nh, pw, 0);
          RNNWF_COUNT_LAUNCH(); scatter_hc_kernel<<<nblk(R * 2 * nh), tpb, 0, st>>>(S, dhc);
        }
        (void)K8;
        break;
      }
      case CELL_GRU4: {
        const int Kz = 24 + 4 * nh;
        RNNWF_COUNT_LAUNCH(); build_gru4_kernel<<<nblk(R * Kz), tpb, 0, st>>>(S, Z);
        // (reduced-precision regime: these six GEMMs run on the tensor cores in TF32, gemm_tc.cuh)
        G.nt_fast(dp0, nh, P("W"), nh, dZ, Kz, R, Kz, nh, 0);                  // dz  = ds W^T
        G.nt_fast(dp1, nh, P("Wg"), nh, dZ, Kz, R, Kz, nh, 1);                 // dz += dg Wg^T
        G.nt_fast(dp2, nh, P("Wmerge"), nh, dZ + 24, Kz, R, 4 * nh, nh, 1);    // dz[:,24:] += dm Wmerge^T
        G.tn_long(Z, Kz, dp0, nh, Gd("W"), nh, R, Kz, nh);
        G.tn_long(Z, Kz, dp1, nh, Gd("Wg"), nh, R, Kz, nh);
        G.tn_long(Z + 24, Kz, dp2, nh, Gd("Wmerge"), nh, R, 4 * nh, nh);
        G.colsum(dp0, nh, R, nh, Gd("b"));
        G.colsum(dp1, nh, R, nh, Gd("bg"));
        RNNWF_COUNT_LAUNCH(); scatter_gru4_kernel<<<nblk(R * 4 * nh), tpb, 0, st>>>(S, dZ);
        break;
      }
      case CELL_VANILLA2D: {
        RNNWF_COUNT_LAUNCH(); build_onehot_kernel<<<nblk(R * 2), tpb, 0, st>>>(S, 2, X);       // X[row] = [x_h (2), x_v (2)]
        G.tn(X, 4, dp0, nh, Gd("Uh"), nh, R, 2, nh);
        G.tn(X + 2, 4, dp0, nh, Gd("Uv"), nh, R, 2, nh);
        G.colsum(dp0, nh, R, nh, Gd("b"));
        if (sm.nbr[0] >= 0) {
          G.tn(hrow(sm.nbr[0]), nhp, dp0, nh, Gd("Wh"), nh, R, nh, nh);
          G.nt(dp0, nh, P("Wh"), nh, dhrow(sm.nbr[0]), nhp, R, nh, nh, 1);
        }
        if (sm.nbr[1] >= 0) {
          G.tn(hrow(sm.nbr[1]), nhp, dp0, nh, Gd("Wv"), nh, R, nh, nh);
          G.nt(dp0, nh, P("Wv"), nh, dhrow(sm.nbr[1]), nhp, R, nh, nh, 1);
        }
        break;
      }
      case CELL_GRU1D: {
        const int l = sm.variant;
        const int din = (l == 0) ? 2 : nh;
        const std::string p = "gru" + std::to_string(l) + ".";
        const float* Wg = P(p + "Wg");
        float* dWg = Gd(p + "Wg");
        if (b1d) {
          // single layer: only the recurrence stays in the loop (d h_{q-1} += d gates Wg_h^T + d candidate Wch^T); the
          // one-hot inputs of this site are kept for the batched weight gradients below
          if (sm.nsym[0] >= 0) { RNNWF_COUNT_LAUNCH(); build_onehot_kernel<<<nblk(R), tpb, 0, st>>>(S, 1, X + (long long)q * R * 2); }
          if (sm.nbr[0] >= 0) {
            G.nt(dp0, 2 * nh, Wg + (long long)din * 2 * nh, 2 * nh, dhrow(sm.nbr[0]), nhp, R, nh, 2 * nh, 1);
            G.nt(dp2, nh, P(p + "Wch"), nh, dhrow(sm.nbr[0]), nhp, R, nh, nh, 1);
          }
          break;
        }
        G.colsum(dp0, 2 * nh, R, 2 * nh, Gd(p + "bg"));
        G.colsum(dp1, nh, R, nh, Gd(p + "bci"));
        G.colsum(dp2, nh, R, nh, Gd(p + "bch"));
        if (sm.nbr[0] >= 0) {  // own previous state h
          G.tn(hrow(sm.nbr[0]), nhp, dp0, 2 * nh, dWg + (long long)din * 2 * nh, 2 * nh, R, nh, 2 * nh);
          G.tn(hrow(sm.nbr[0]), nhp, dp2, nh, Gd(p + "Wch"), nh, R, nh, nh);
          G.nt(dp0, 2 * nh, Wg + (long long)din * 2 * nh, 2 * nh, dhrow(sm.nbr[0]), nhp, R, nh, 2 * nh, 1);
          G.nt(dp2, nh, P(p + "Wch"), nh, dhrow(sm.nbr[0]), nhp, R, nh, nh, 1);
        }
        if (l == 0) {          // one-hot input of the previous spin
          if (sm.nsym[0] >= 0) {
            RNNWF_COUNT_LAUNCH(); build_onehot_kernel<<<nblk(R), tpb, 0, st>>>(S, 1, X);
            G.tn(X, 2, dp0, 2 * nh, dWg, 2 * nh, R, 2, 2 * nh);
            G.tn(X, 2, dp1, nh, Gd(p + "Wci"), nh, R, 2, nh);
          }
        } else {               // dense input from the layer below (same lattice site)
          const int qb = sm.nbr[1];
          G.tn(hrow(qb), nhp, dp0, 2 * nh, dWg, 2 * nh, R, nh, 2 * nh);
          G.tn(hrow(qb), nhp, dp1, nh, Gd(p + "Wci"), nh, R, nh, nh);
          G.nt(dp0, 2 * nh, Wg, 2 * nh, dhrow(qb), nhp, R, nh, 2 * nh, 1);
          G.nt(dp1, nh, P(p + "Wci"), nh, dhrow(qb), nhp, R, nh, nh, 1);
        }
        break;
      }
    }
  }
  if (L.batched1d) {
    // Single-layer 1D GRU: weight and bias gradients over all sites at once.  Site 0 has neither a previous state nor a
    // one-hot input (the first input is the zero vector); sites 1..Q-1 are contiguous in every array.
    const long long RQ = (long long)m.Q * R, R1 = (long long)(m.Q - 1) * R;
    float* dgA = dp;                                          // [Q][R][2nh]
    float* dciA = dp + (long long)m.Q * R * 2 * nh;           // [Q][R][nh]
    float* dchA = dp + (long long)m.Q * R * 3 * nh;           // [Q][R][nh]
    float* dWg = Gd("gru0.Wg");
    G.colsum(dgA, 2 * nh, RQ, 2 * nh, Gd("gru0.bg"));
    G.colsum(dciA, nh, RQ, nh, Gd("gru0.bci"));
    G.colsum(dchA, nh, RQ, nh, Gd("gru0.bch"));
    if (m.Q > 1) {
      G.tn(hist, nhp, dgA + R * 2LL * nh, 2 * nh, dWg + 2LL * 2 * nh, 2 * nh, R1, nh, 2 * nh);     // rows of Wg fed by h_{q-1}
      G.tn(hist, nhp, dchA + R * (long long)nh, nh, Gd("gru0.Wch"), nh, R1, nh, nh);
      G.tn(X + R * 2LL, 2, dgA + R * 2LL * nh, 2 * nh, dWg, 2 * nh, R1, 2, 2 * nh);               // rows fed by the one-hot input
      G.tn(X + R * 2LL, 2, dciA + R * (long long)nh, nh, Gd("gru0.Wci"), nh, R1, 2, nh);
    }
    G.tn(hist, nhp, dlog_base, m.ND, Gd("dense.W"), m.D, RQ, nh, m.D);
    G.colsum(dlog_base, m.ND, RQ, m.D, Gd("dense.b"));
    if (m.complex_head) {
      G.tn(hist, nhp, dlog_base + m.D, m.ND, Gd("dense_phase.W"), 2, RQ, nh, 2);
      G.colsum(dlog_base + m.D, m.ND, RQ, 2, Gd("dense_phase.b"));
    }
  }
  if (L.batched) {
    // The step-0 site has no inputs; steps 1..Q-1 are contiguous in the batched operand arrays.
    const bool gru = m.cell == CELL_TENSGRU2D;
    const int K8 = 8 * nh;
    const long long RQ = (long long)m.Q * R, R1 = (long long)(m.Q - 1) * R;
    float* dsA = dp;
    float* dgA = dp + (long long)m.Q * R * nh;
    float* dmA = dp + 2LL * m.Q * R * nh;
    if (m.Q > 1) {
      G.tn_long(dsA + R * nh, nh, Z_base + R * (long long)K8, K8, Gd("W"), K8, R1, nh, K8);          // dW2 += ds^T T
      if (gru) {
        G.tn_long(dgA + R * nh, nh, Z_base + R * (long long)K8, K8, Gd("Wg"), K8, R1, nh, K8);
        G.tn_long(hc_base + R * 2LL * nh, 2 * nh, dmA + R * nh, nh, Gd("Wmerge"), nh, R1, 2 * nh, nh);  // dWmerge += hc^T dm
      }
    }
    G.colsum(dsA, nh, RQ, nh, Gd("b"));
    if (gru) G.colsum(dgA, nh, RQ, nh, Gd("bg"));
    G.tn(hist, nhp, dlog_base, m.ND, Gd("dense.W"), m.D, RQ, nh, m.D);
    G.colsum(dlog_base, m.ND, RQ, m.D, Gd("dense.b"));
    if (m.complex_head) {
      G.tn(hist, nhp, dlog_base + m.D, m.ND, Gd("dense_phase.W"), 2, RQ, nh, 2);
      G.colsum(dlog_base + m.D, m.ND, RQ, 2, Gd("dense_phase.b"));
    }
  }
  e = cudaGetLastError();
  if (e != cudaSuccess) { err = std::string("rnnwf_backward kernels: ") + cudaGetErrorString(e); return RNNWF_ECUDA; }
  return RNNWF_OK;
}

}  // namespace rnnwf
