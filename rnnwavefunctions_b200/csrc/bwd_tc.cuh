// Tensor-core reverse sweep of the tensorized 2D GRU (reduced-precision regime): the recurrence part of the backward
// pass of  sum_b seed_b * log psi(sigma_b)  as ONE persistent kernel instead of ~5 launches per lattice site.
//
// Per 128-row tile the sites are walked backwards.  For site q the tile's threads turn d h_q (accumulated from the
// sites that consumed h_q, plus the heads' contribution) into the pre-activation gradients (ds, dg, dm) with the gates
// saved by the sampling sweep, and form the linearised operand
//        P = [ds, cl ds, cu ds, dg, cl dg, cu dg, dm]            (cl, cu: the neighbours' spins, 0 if absent)
// so that  d[h_h ; h_v] = P * Wstack^T  with the same pre-summed weight blocks as the forward sweep (sweep_tc.cuh):
// one [128 x 784] x [784 x 208] contraction per site on the tensor cores (tcgen05.mma kind::tf32 -- gradients need the
// FP32 exponent range -- A operand written straight into TMEM by the threads that computed it, weights streamed from
// L2 in 7 chunks of 16 units, FP32 accumulator [128 x 208] in TMEM).  The K axis of a chunk is ordered exactly like the
// forward sweep's accumulator columns (thread-major: sub * 28 + block * 4 + e), the N axis like its operand (tc_kunit),
// so every thread only ever touches its own columns and its own pieces of d h: no atomics, no barriers on d h.
// d h of a site has two contributions.  The one of its vertical successor is written to global memory with plain stores
// (each site is the vertical neighbour of at most one site; sites without one read the zero-initialised buffer).  The one
// of its horizontal successor is consumed by the same thread at the very next site of the sweep, so it never leaves the
// SM: 20 of the thread's 28 values wait in the 80 TMEM columns left free by the accumulator and the operand buffers,
// 8 in registers (h_chain lattices: every 2D zig-zag path; otherwise both contributions are read-modify-write in global).
//
// Warp roles: warp 0 MMA issue, warps 1..16 four threads per row (thread (row, sub) owns units 16 p + 4 sub + e),
// warp 17 weight loader.  Everything is synchronised by mbarriers; waits are bounded and trap instead of hanging.
//
// Outputs: ds, dg, dm of every site ([Q][R][nh] each, for the batched weight-gradient GEMMs) -- the parameter
// gradients themselves are long-K GEMMs over these (backward.cuh).
// Reference behaviour: tf.gradients of the cost in 2DHeisenberg/Training2DRNN_Heisenberg.py through
// MDTensorizedGRUcell.py:49-61 / MDComplexRNN.py:69-286 (same math as cell_bwd_tens_kernel + gemm + scatter_hc_kernel).
#pragma once
#include "sweep_tc.cuh"

namespace rnnwf {

constexpr int BT_K = TC_TPS * TC_SUB;                 // 112 K elements per chunk (16 units x 7 blocks)
constexpr int BT_THREADS = 32 + TC_EPI_THREADS + 32;  // warp 0: MMA, warps 1..16: rows, warp 17: loader
constexpr int BT_NK8 = BT_K / 8;                      // tf32 MMAs (K = 8) per chunk
// feature-major (ds | dg | dm) tiles [site][tile][unit nhp][TC_DPT_PITCH]: 128 rows + 4 floats of padding per unit row, so that
// the weight-gradient kernel stages a tile with ONE bulk copy and reads it conflict-free at lane <-> unit (dw_tc.cuh)
constexpr int TC_DPT_PITCH = TC_ROWS + 4;

struct BtSmem {
  int off_B, off_headw, off_step, off_sym, off_bar, off_tmem, total;
  int b_bytes, ncores;
  __host__ __device__ BtSmem(int nhp, int Tsym, int Q) {
    ncores = 2 * nhp / 8;                             // 8-row groups of the N axis [d h_h ; d h_v]
    b_bytes = (BT_K / 4) * ncores * 128;              // one chunk of Wstack^T: [k-core 28][n-core][8 rows][4 tf32]
    int o = 0;
    off_B = o;     o += 2 * b_bytes;
    off_headw = o; o += (nhp + TC_JP) * 16;           // head weights [unit][4], zero rows past num_units
    off_step = o;  o += Q * 8;                        // per site: neighbours (int16 x 2), flags
    off_bar = o;   o += 128;
    off_tmem = o;  o += 16;
    off_sym = o;   o += Tsym * 16;                    // symbols, one bit per (site, row)
    total = o;
  }
};

struct BwdTcParams {
  const StepMeta* steps;
  int Q, Tsym, nsites, nh, nhp, np, ND;
  long long rows;
  const int8_t* cfg;            // [rows][nsites]
  const float4* gates;          // saved by the sampling sweep, tile-blocked (tc_gate_index)
  float* dlog;                  // [Q][rows][ND]  d(logits) of every site: written here (for the heads' weight gradients)
  const float4* sv_head;        // [Q][rows] heads' outputs saved by the forward sweep: (p0, p1, phase pre-activation of the symbol, 0)
  const double* seed_logp;      // [rows] d cost / d log P
  const double* seed_phase;     // [rows] d cost / d phase (complex head), may be null
  float4* dh;                   // [Q][tiles][sub 4][pass np][row 128] float4, zero-initialised
  float* dpT;                   // [src 3: ds, dg, dm][Q][tiles][unit nhp][row 128]  (feature-major tiles for dw_tc.cuh)
  const float* wimg;            // [variant 3][chunk np][b_bytes / 4]
  const float* head_w;          // [nh][ND]
  const float* hist;            // vanilla cell: saved hidden states [Q][rows][nhp] (elu' = h > 0 ? 1 : h + 1)
  int tile_rows;                // live rows per tile (<= 128): batches of fewer tiles than SMs are spread over all SMs (as in the forward sweep)
  int h_chain;                  // 1: every horizontal neighbour is the site generated right before (nbr[0] == q - 1 wherever present)
  int debug;                    // timing experiments (RNNWF_TC_DEBUG bits 19..21: no MMAs / no feature-major stores / no d h update), 0 in production
};

// tile-blocked layouts shared with the sampling sweep: everything a thread (row, sub) touches for column pass p is
// one float4 per quantity, and a warp's 32 rows are contiguous (fully coalesced)
__host__ __device__ inline size_t tc_gate_index(int q, long long tile, long long ntiles, int np, int sub, int p, int c, int row) {
  return ((((size_t)q * (size_t)ntiles + (size_t)tile) * TC_SUB + sub) * np + p) * 3 * TC_ROWS + (size_t)c * TC_ROWS + row;
}
__host__ __device__ inline size_t tc_dh_index(int q, long long tile, long long ntiles, int np, int sub, int p, int row) {
  return ((((size_t)q * (size_t)ntiles + (size_t)tile) * TC_SUB + sub) * np + p) * TC_ROWS + row;
}
__host__ __device__ inline long long tc_gate_bytes(int Q, long long rows, int np) {
  return (long long)Q * ((rows + TC_ROWS - 1) / TC_ROWS) * TC_SUB * np * 3 * TC_ROWS * 16;
}
__host__ __device__ inline long long tc_dh_bytes(int Q, long long rows, int np) {
  return (long long)Q * ((rows + TC_ROWS - 1) / TC_ROWS) * TC_SUB * np * TC_ROWS * 16;
}

// D[tmem] (+)= A[tmem] * B[smem]^T, tf32 operands (one 32-bit TMEM column per A element, K = 8 per instruction)
__device__ __forceinline__ void tc_mma_tf32_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// instruction descriptor: D = F32 (bit 4), A = B = TF32 (format 2 at bits 7 and 10), K-major, N >> 3 at 17, M >> 4 at 24
__host__ __device__ constexpr uint32_t tc_idesc_tf32(int M, int N) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tc_st28(uint32_t taddr, const uint32_t (&d)[TC_TPS]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(d[0]), "r"(d[1]), "r"(d[2]), "r"(d[3]), "r"(d[4]), "r"(d[5]), "r"(d[6]), "r"(d[7]), "r"(d[8]), "r"(d[9]),
        "r"(d[10]), "r"(d[11]), "r"(d[12]), "r"(d[13]), "r"(d[14]), "r"(d[15]) : "memory");
  tc_st8(taddr + 16u, d[16], d[17], d[18], d[19], d[20], d[21], d[22], d[23]);
  tc_st4(taddr + 24u, make_uint4(d[24], d[25], d[26], d[27]));
}

// ---- weight images: Wstack^T per variant and chunk ------------------------------------------------------------
// out[variant 3][chunk p][k-core 28][n-core][8 rows][4]:  element (n, k) of chunk p is the derivative of pre-activation
// block blk of unit j = 16 p + 4 sub + e  (k = sub * 28 + blk * 4 + e) with respect to input i of [h_h ; h_v]
// (n = half * nhp + position, unit of the half = tc_kunit(position)).  Same variants as pack_tc_kernel.
__device__ __forceinline__ void pack_bwd_tc_body(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                                                 int nh, int nhp, int NP, float* __restrict__ out, long long vb, long long vg) {
  const int ncores = 2 * nhp / 8;
  const long long per_chunk = (long long)(BT_K / 4) * ncores * 32;
  const long long total = 3LL * NP * per_chunk;
  for (long long idx = vb * (long long)blockDim.x + threadIdx.x; idx < total; idx += vg * blockDim.x) {
    long long r = idx;
    const int ek = (int)(r % 4); r /= 4;
    const int en = (int)(r % 8); r /= 8;
    const int nn = (int)(r % ncores); r /= ncores;
    const int kk = (int)(r % (BT_K / 4)); r /= (BT_K / 4);
    const int p = (int)(r % NP); r /= NP;
    const int variant = (int)r;
    const int k = kk * 4 + ek, n = nn * 8 + en;
    const int sub = k / TC_TPS, blk = (k % TC_TPS) / 4, j = p * TC_JP + sub * 4 + (k % 4);
    const int half = n / nhp, unit = tc_kunit(nhp, n - half * nhp);
    float v = 0.f;
    if (unit < nh && j < nh) {
      const int i = half * nh + unit;
      if (blk == 6) {
        v = (Wmerge != nullptr) ? Wmerge[(long long)i * nh + j] : 0.f;
      } else if (blk >= 3 && Wg == nullptr) {
        v = 0.f;                                   // un-gated tensorized cell
      } else {
        const float* w = ((blk < 3) ? W : Wg) + ((long long)j * 2 * nh + i) * 4;
        const int kind = blk % 3;
        if (variant == 0) v = (kind == 0) ? (w[0] + w[2]) : (kind == 1 ? (w[1] - w[0]) : (w[3] - w[2]));
        else if (variant == 1) v = (kind == 0) ? w[0] : (kind == 1 ? (w[1] - w[0]) : 0.f);
        else v = (kind == 0) ? w[2] : (kind == 1 ? 0.f : (w[3] - w[2]));
      }
    }
    out[idx] = v;
  }
}
__global__ void pack_bwd_tc_kernel(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                                   int nh, int nhp, int NP, float* __restrict__ out) {
  pack_bwd_tc_body(W, Wg, Wmerge, nh, nhp, NP, out, blockIdx.x, gridDim.x);
}

// ---- vanilla 2D RNN: one chunk, element (n, k) = d pre_j / d h_{half, i} = W_half[i][j]   (k <-> unit j, n <-> (half, unit i))
__global__ void pack_bwd_tc_vanilla_kernel(const float* __restrict__ Wh, const float* __restrict__ Wv, int nh, int nhp, float* __restrict__ out) {
  const int ncores = 2 * nhp / 8;
  const long long total = (long long)(BT_K / 4) * ncores * 32;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int ek = (int)(r % 4); r /= 4;
    const int en = (int)(r % 8); r /= 8;
    const int nn = (int)(r % ncores); r /= ncores;
    const int kk = (int)r;
    const int k = kk * 4 + ek, n = nn * 8 + en;
    const int sub = k / TC_TPS, blk = (k % TC_TPS) / 4, j = blk * TC_JP + sub * 4 + (k % 4);
    const int half = n / nhp, unit = tc_kunit(nhp, n - half * nhp);
    out[idx] = (unit < nh && j < nh) ? (half == 0 ? Wh : Wv)[(long long)unit * nh + j] : 0.f;
  }
}

// vanilla 2D RNN: gradients of the one-hot input kernels and the bias from the feature-major d(pre) tiles:
//   db[j] = sum d pre_j,   dUh[c][j] = sum [horizontal neighbour present and sigma_l == c] d pre_j,   dUv likewise.
// grid (ceil(nh / 8), nsplit), 128 threads = the rows of a tile.  A CTA sums a block of 8 units over its share of the
// (site, tile) items, so the two scattered symbol reads per row and item are shared by 8 coalesced 512-byte tile rows
// (one unit per CTA re-read them 50 times: 656 us at C3; now the tile reads dominate).
constexpr int VIG_UB = 8;
__global__ void __launch_bounds__(128) vanilla_input_grads_kernel(const float* __restrict__ dpT, const StepMeta* __restrict__ steps,
                                                                  const int16_t* __restrict__ site_of_t, const int8_t* __restrict__ cfg,
                                                                  int Q, long long tiles, long long rows, int nh, int nhp, int nsites,
                                                                  float* __restrict__ gb, float* __restrict__ gUh, float* __restrict__ gUv) {
  __shared__ float red[4][VIG_UB][5];
  const int j0 = blockIdx.x * VIG_UB, r = threadIdx.x;
  float acc[VIG_UB][5];
#pragma unroll
  for (int u = 0; u < VIG_UB; ++u)
#pragma unroll
    for (int k = 0; k < 5; ++k) acc[u][k] = 0.f;
  const long long nitems = (long long)Q * tiles;
  for (long long it = blockIdx.y; it < nitems; it += gridDim.y) {
    const int q = (int)(it / tiles);
    const long long row = (it % tiles) * TC_ROWS + r;
    const bool alive = row < rows;
    const float* src = dpT + ((size_t)it * nhp + j0) * TC_DPT_PITCH + r;
    float v[VIG_UB];
#pragma unroll
    for (int u = 0; u < VIG_UB; ++u) v[u] = (alive && j0 + u < nh) ? src[(size_t)u * TC_DPT_PITCH] : 0.f;
    const int n0 = steps[q].nbr[0], n1 = steps[q].nbr[1];
    const int8_t* cf = cfg + (alive ? row : 0) * nsites;
    const bool hl = n0 >= 0, hu = n1 >= 0;
    const bool l1 = hl && (cf[site_of_t[hl ? n0 : 0]] & 1), u1 = hu && (cf[site_of_t[hu ? n1 : 0]] & 1);
    const bool l0 = hl && !l1, u0 = hu && !u1;
#pragma unroll
    for (int u = 0; u < VIG_UB; ++u) {
      acc[u][0] += v[u];
      acc[u][1] += l0 ? v[u] : 0.f;
      acc[u][2] += l1 ? v[u] : 0.f;
      acc[u][3] += u0 ? v[u] : 0.f;
      acc[u][4] += u1 ? v[u] : 0.f;
    }
  }
#pragma unroll
  for (int u = 0; u < VIG_UB; ++u)
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      float x = acc[u][k];
      for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
      if ((r & 31) == 0) red[r >> 5][u][k] = x;
    }
  __syncthreads();
  if (r < VIG_UB * 5) {
    const int u = r / 5, k = r % 5, j = j0 + u;
    if (j < nh) {
      const float x = red[0][u][k] + red[1][u][k] + red[2][u][k] + red[3][u][k];
      float* dst = (k == 0) ? gb + j : (k <= 2 ? gUh + (k - 1) * nh + j : gUv + (k - 3) * nh + j);
      atomicAdd(dst, x);
    }
  }
}

// ---- the reverse sweep ------------------------------------------------------------------------------------------
// CELLK: 0 = tensorized GRU (np chunks of 16 units x 7 blocks per site); 1 = vanilla 2D RNN (one chunk per site: the
// K axis holds d(pre-activation) of all units, position sub * 28 + piece * 4 + e; P.np = unit pieces of the layouts)
template <int CELLK>
__global__ void __launch_bounds__(BT_THREADS, 1) bwd_tc_kernel(const BwdTcParams P) {
  extern __shared__ __align__(128) unsigned char bt_smem[];
  unsigned char* const smem = bt_smem;
  const BtSmem L(P.nhp, P.Tsym, P.Q);
  unsigned char* const Bst = smem + L.off_B;                              // [2][b_bytes] weight chunk stages
  float* const headW = reinterpret_cast<float*>(smem + L.off_headw);      // [nhp + 16][4], zero past num_units / ND
  short4* const stepS = reinterpret_cast<short4*>(smem + L.off_step);     // per site: (nbr0, nbr1, site, has_inputs)
  uint32_t* const symS = reinterpret_cast<uint32_t*>(smem + L.off_sym);   // [Tsym][4] ballot words (tc_sym)
  uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
  uint64_t* const bar_A = bars + 0;       // [2] operand chunk written to TMEM (512 arrivals)
  uint64_t* const bar_B = bars + 2;       // [2] weight chunk landed (complete_tx)
  uint64_t* const bar_free = bars + 4;    // [2] chunk consumed: operand buffer and weight stage reusable (tcgen05.commit)
  uint64_t* const bar_D = bars + 6;       // accumulator of the site complete (tcgen05.commit)
  uint64_t* const bar_Dfree = bars + 7;   // accumulator drained (512 arrivals)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nh = P.nh, nhp = P.nhp, NP = P.np;
  const int NCH = (CELLK == 1) ? 1 : NP;              // operand chunks (MMA K groups) per site
  const uint32_t ncols = 2u * (uint32_t)nhp;          // accumulator columns: [d h_h ; d h_v]
  // TMEM map: [0, ncols) accumulator; then two operand chunks of 112 columns (one tf32 per column)
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(bar_A + i, TC_EPI_THREADS); mbar_init(bar_B + i, 1); mbar_init(bar_free + i, 1); }
    mbar_init(bar_D, 1);
    mbar_init(bar_Dfree, TC_EPI_THREADS);
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, TC_TMEM_COLS);
  for (int i = tid; i < (nhp + TC_JP) * 4; i += BT_THREADS) {
    const int j = i >> 2, d = i & 3;
    headW[i] = (j < nh && d < P.ND) ? P.head_w[j * P.ND + d] : 0.f;
  }
  for (int i = tid; i < P.Q; i += BT_THREADS) {
    const StepMeta* m = P.steps + i;
    stepS[i] = make_short4(m->nbr[0], m->nbr[1], m->site, (m->nbr[0] >= 0 || m->nbr[1] >= 0) ? 1 : 0);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_A0 = tmem_base + ((ncols + 15u) & ~15u);
  const long long ntiles = (P.rows + TC_ROWS - 1) / TC_ROWS;        // 128-row tiles of the tile-blocked layouts (gates, d h, feature-major tiles)
  const int VR = P.tile_rows;                                          // rows this kernel processes per tile
  const long long vtiles = (ntiles * TC_ROWS + VR - 1) / VR;          // (they cover the rows up to a multiple of 128: zeros are written past the batch)

  if (warp == 0) {
    // ======================= MMA warp =======================
    const bool leader = elect_one();
    const uint32_t B_LBO = (uint32_t)L.ncores * 128u, SBO = 128u;
    const uint64_t bdesc0 = tc_desc(smem_u32(Bst), B_LBO, SBO);
    const uint64_t kstep = (uint64_t)((2u * B_LBO) >> 4);           // two k-cores (8 tf32) per MMA
    const uint32_t idesc = tc_idesc_tf32(TC_ROWS, (int)ncols);
    uint32_t phA = 0, phB = 0, c = 0, nsite = 0;
    for (long long tile = blockIdx.x; tile < vtiles; tile += gridDim.x) {
      for (int q = P.Q - 1; q >= 0; --q) {
        if (!stepS[q].w) continue;                                  // first site: no inputs to differentiate
        for (int p = 0; p < NCH; ++p, ++c) {
          const uint32_t s = c & 1u;
          mbar_wait(bar_B + s, (phB >> s) & 1u); phB ^= 1u << s;
          mbar_wait(bar_A + s, (phA >> s) & 1u); phA ^= 1u << s;
          if (p == 0 && nsite > 0) mbar_wait(bar_Dfree, (nsite - 1) & 1u);
          tc_fence_after();
          const uint32_t tmem_a = tmem_A0 + s * (uint32_t)BT_K;
          const uint64_t bd = bdesc0 + (uint64_t)((s * (uint32_t)L.b_bytes) >> 4);
          if (leader) {
#pragma unroll
            for (int ks = 0; ks < BT_NK8; ++ks)
              if (!(P.debug & (1 << 19))) tc_mma_tf32_ts(tmem_base, tmem_a + (uint32_t)ks * 8u, bd + (uint64_t)ks * kstep, idesc, (p > 0 || ks > 0) ? 1u : 0u);
            tc_commit(bar_free + s);
            if (p == NCH - 1) tc_commit(bar_D);
          }
          __syncwarp();
        }
        ++nsite;
      }
    }
  } else if (warp == BT_THREADS / 32 - 1) {
    // ======================= loader warp: weight chunks =======================
    const bool leader = elect_one();
    uint32_t phF = 0, c = 0;
    for (long long tile = blockIdx.x; tile < vtiles; tile += gridDim.x) {
      for (int q = P.Q - 1; q >= 0; --q) {
        const short4 sm = stepS[q];
        if (!sm.w) continue;
        const int v = (CELLK == 1) ? 0 : ((sm.x >= 0 && sm.y >= 0) ? 0 : (sm.x >= 0 ? 1 : 2));   // vanilla: one image
        const unsigned char* src = reinterpret_cast<const unsigned char*>(P.wimg) + (size_t)v * NCH * L.b_bytes;
        for (int p = 0; p < NCH; ++p, ++c) {
          const uint32_t s = c & 1u;
          if (c >= 2) { mbar_wait_relaxed(bar_free + s, (phF >> s) & 1u); phF ^= 1u << s; }
          if (leader) {
            mbar_expect_tx(bar_B + s, L.b_bytes);
            bulk_g2s(Bst + (size_t)s * L.b_bytes, src + (size_t)p * L.b_bytes, L.b_bytes, bar_B + s);
          }
          __syncwarp();
        }
      }
    }
  } else {
    // ======================= row warps: 4 threads per row =======================
    const int quarter = warp & 3;
    const int sub = (warp - 1) >> 2;
    const int row = (quarter << 5) + lane;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const int np_own = tc_kcnt(nhp, sub) / 4;                       // column passes in which this thread owns units
    const uint32_t kcol = (uint32_t)tc_koff(nhp, sub);              // its first column inside a neighbour half of D
    const uint32_t a_thr = tmem_A0 + lane_base + (uint32_t)(sub * TC_TPS);   // its 28 columns of operand buffer 0
    const float* const hW = headW + sub * 16;
    // the horizontal successor's contribution to d h of the next site: columns [432 + 20 sub, + 20) of this row's lane, and hreg
    const uint32_t stash_thr = tmem_A0 + lane_base + 2u * (uint32_t)BT_K + (uint32_t)(sub * 20);
    float hreg[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) hreg[i] = 0.f;
    bool have_h = false;                                            // warp-uniform: the site being processed has a stashed contribution
    auto h_piece = [&](int p, float (&hc)[4]) {                     // the stash's values of column pass p (units 16 p + 4 sub + e)
      hc[0] = hc[1] = hc[2] = hc[3] = 0.f;
      if (!have_h) return;
      if (p < 5) {
        uint32_t t[4];
        tc_ld4(stash_thr + (uint32_t)(4 * p), t);
        tc_wait_ld();
        hc[0] = __uint_as_float(t[0]); hc[1] = __uint_as_float(t[1]); hc[2] = __uint_as_float(t[2]); hc[3] = __uint_as_float(t[3]);
      } else if (p == 5) { hc[0] = hreg[0]; hc[1] = hreg[1]; hc[2] = hreg[2]; hc[3] = hreg[3]; }
      else { hc[0] = hreg[4]; hc[1] = hreg[5]; hc[2] = hreg[6]; hc[3] = hreg[7]; }
    };
    uint32_t phF = 0, c = 0, nsite = 0;
    for (long long tile = blockIdx.x; tile < vtiles; tile += gridDim.x) {
      const long long row0 = tile * VR;
      const bool live = row < VR && row0 + row < P.rows;
      const bool stores = row < VR && row0 + row < ntiles * TC_ROWS;   // owns a row of the feature-major tiles (zeros past the batch)
      const long long stile = (row0 + row) / TC_ROWS;                   // position of the row in the 128-row tile-blocked layouts
      const int srow = (int)((row0 + row) % TC_ROWS);
      have_h = false;
      const float seed_l = live ? (float)P.seed_logp[row0 + row] : 0.f;
      const float seed_p = (live && P.seed_phase != nullptr) ? (float)P.seed_phase[row0 + row] : 0.f;
      // ---- symbol table of the tile
      {
        const int8_t* cf = P.cfg + (row0 + row) * P.nsites;
        for (int t = sub; t < P.Tsym; t += TC_SUB) {
          const uint32_t w = __ballot_sync(0xffffffffu, live ? (cf[stepS[t].z] & 1) : 0);
          if (lane == 0) symS[t * 4 + quarter] = w;
        }
        quarter_barrier(quarter);
      }
      for (int q = P.Q - 1; q >= 0; --q) {
        const short4 sm = stepS[q];
        const int nb0 = sm.x, nb1 = sm.y;
        const bool has_in = sm.w != 0;
        const float cl = (nb0 >= 0) ? (float)tc_sym(symS, nb0, row) : 0.f;
        const float cu = (nb1 >= 0) ? (float)tc_sym(symS, nb1, row) : 0.f;
        // d(logits) of this (site, row) from the heads' saved outputs: d log p_s / d z_k = [k == s] - p_k (masked softmax),
        // d phase / d x_s = pi / (1 + |x_s|)^2 (softsign head); 2D cells: site q <-> symbol q
        float dl[4] = {0.f, 0.f, 0.f, 0.f};
        if (live) {
          const float4 hd = __ldg(P.sv_head + (size_t)q * (size_t)P.rows + (size_t)(row0 + row));
          const int sy = tc_sym(symS, q, row);
          dl[0] = seed_l * ((sy == 0 ? 1.f : 0.f) - hd.x);
          dl[1] = seed_l * ((sy == 1 ? 1.f : 0.f) - hd.y);
          if (P.ND == 4) {
            const float den = 1.f + fabsf(hd.z);
            const float g = seed_p * 3.14159265358979f / (den * den);
            dl[2] = (sy == 0) ? g : 0.f;
            dl[3] = (sy == 1) ? g : 0.f;
          }
          if (sub == 0) {
            float* dlp = P.dlog + ((size_t)q * (size_t)P.rows + (size_t)(row0 + row)) * P.ND;
            if (P.ND == 4) *reinterpret_cast<float4*>(dlp) = make_float4(dl[0], dl[1], dl[2], dl[3]);
            else *reinterpret_cast<float2*>(dlp) = make_float2(dl[0], dl[1]);
          }
        }
        if constexpr (CELLK == 1) {
          // ---- vanilla cell: d(pre) = (d h + heads' push) * elu'(pre), elu' = 1 for h > 0 else h + 1; one chunk per site
          const float4* dhq = P.dh + tc_dh_index(q, stile, ntiles, NP, sub, 0, srow);
          const float* hq = P.hist + ((size_t)q * (size_t)P.rows + (size_t)(row0 + row)) * nhp + sub * 4;
          uint32_t d[TC_TPS];
#pragma unroll
          for (int i = 0; i < TC_TPS; ++i) d[i] = 0u;
#pragma unroll
          for (int blk = 0; blk < TC_NB; ++blk) {
            if (blk < np_own) {                                       // warp-uniform
              float dp[4] = {0.f, 0.f, 0.f, 0.f};
              const int j0 = blk * TC_JP + sub * 4;
              float hc[4];
              h_piece(blk, hc);
              if (live) {
                const float4 dh4 = __ldcg(dhq + (size_t)blk * TC_ROWS);
                const float4 h4 = __ldg(reinterpret_cast<const float4*>(hq + blk * TC_JP));
                const float dhv[4] = {dh4.x, dh4.y, dh4.z, dh4.w}, hv[4] = {h4.x, h4.y, h4.z, h4.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float4 w = *reinterpret_cast<const float4*>(hW + blk * (TC_JP * 4) + e * 4);
                  const float g = dhv[e] + hc[e] + dl[0] * w.x + dl[1] * w.y + dl[2] * w.z + dl[3] * w.w;
                  dp[e] = (j0 + e < nh) ? g * (hv[e] > 0.f ? 1.f : hv[e] + 1.f) : 0.f;
                  d[blk * 4 + e] = __float_as_uint(dp[e]);
                }
              }
              float* o0 = P.dpT + (((size_t)q * (size_t)ntiles + (size_t)stile) * nhp + j0) * TC_DPT_PITCH + srow;
#pragma unroll
              for (int e = 0; e < 4; ++e)
                if (j0 + e < nhp && stores) o0[(size_t)e * TC_DPT_PITCH] = dp[e];
            }
          }
          if (has_in) {
            const uint32_t s = c & 1u;
            if (c >= 2) { mbar_wait(bar_free + s, (phF >> s) & 1u); phF ^= 1u << s; }
            __syncwarp();
            tc_st28(a_thr + s * (uint32_t)BT_K, d);
            tc_wait_st();
            tc_fence_before();
            mbar_arrive(bar_A + s);
            ++c;
          }
        } else {
          const float4* gq = P.gates + tc_gate_index(q, stile, ntiles, NP, sub, 0, 0, srow);
          const float4* dhq = P.dh + tc_dh_index(q, stile, ntiles, NP, sub, 0, srow);
          // The saved gates stream from DRAM and the loop below only looks one column pass ahead: pull the NEXT site's gates
          // of this warp (np_own passes x 3 quantities x 512 contiguous bytes) into the L2 a whole site ahead.
          if (q > 0 && (lane & 7) == 0 && !(P.debug & (1 << 22))) {
            const float4* gn = P.gates + tc_gate_index(q - 1, stile, ntiles, NP, sub, 0, 0, srow);
            for (int i = 0; i < 3 * np_own; ++i) asm volatile("prefetch.global.L2 [%0];" ::"l"(gn + (size_t)i * TC_ROWS));
            const float4* dn = P.dh + tc_dh_index(q - 1, stile, ntiles, NP, sub, 0, srow);       // (its vertical successor wrote it ~2 rows of sites ago)
            for (int i = 0; i < np_own; ++i) asm volatile("prefetch.global.L2 [%0];" ::"l"(dn + (size_t)i * TC_ROWS));
            if (sub == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(P.sv_head + (size_t)(q - 1) * (size_t)P.rows + (size_t)(row0 + row)));
          }
          float4 n_dh = make_float4(0.f, 0.f, 0.f, 0.f), n_u = n_dh, n_t = n_dh, n_m = n_dh;
          if (live && np_own > 0) { n_dh = __ldcg(dhq); n_u = __ldg(gq); n_t = __ldg(gq + TC_ROWS); n_m = __ldg(gq + 2 * TC_ROWS); }
  #pragma unroll 1
          for (int p = 0; p < NP; ++p) {
            uint32_t d[TC_TPS];
  #pragma unroll
            for (int i = 0; i < TC_TPS; ++i) d[i] = 0u;
            if (p < np_own) {
              float ds[4] = {0.f, 0.f, 0.f, 0.f}, dg[4] = {0.f, 0.f, 0.f, 0.f}, dm[4] = {0.f, 0.f, 0.f, 0.f};
              const int j0 = p * TC_JP + sub * 4;
              float hc[4];
              h_piece(p, hc);
              if (live) {
                const float4 dh4 = n_dh, u4 = n_u, t4 = n_t, m4 = n_m;      // loaded one pass ahead
                if (p + 1 < np_own) {
                  n_dh = __ldcg(dhq + (size_t)(p + 1) * TC_ROWS);
                  n_u = __ldg(gq + (size_t)(p + 1) * 3 * TC_ROWS);
                  n_t = __ldg(gq + (size_t)(p + 1) * 3 * TC_ROWS + TC_ROWS);
                  n_m = __ldg(gq + (size_t)(p + 1) * 3 * TC_ROWS + 2 * TC_ROWS);
                }
                const float dhv[4] = {dh4.x, dh4.y, dh4.z, dh4.w}, uv[4] = {u4.x, u4.y, u4.z, u4.w};
                const float tv[4] = {t4.x, t4.y, t4.z, t4.w}, mv[4] = {m4.x, m4.y, m4.z, m4.w};
  #pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float4 w = *reinterpret_cast<const float4*>(hW + p * (TC_JP * 4) + e * 4);
                  const float g = dhv[e] + hc[e] + dl[0] * w.x + dl[1] * w.y + dl[2] * w.z + dl[3] * w.w;   // + heads' contribution
                  const bool real = (j0 + e < nh);                     // padding units carry nothing (their saved gates are unset)
                  ds[e] = real ? g * uv[e] * (1.f - tv[e] * tv[e]) : 0.f;
                  dg[e] = real ? g * (tv[e] - mv[e]) * uv[e] * (1.f - uv[e]) : 0.f;
                  dm[e] = real ? g * (1.f - uv[e]) : 0.f;
                  d[e] = __float_as_uint(ds[e]); d[4 + e] = __float_as_uint(cl * ds[e]); d[8 + e] = __float_as_uint(cu * ds[e]);
                  d[12 + e] = __float_as_uint(dg[e]); d[16 + e] = __float_as_uint(cl * dg[e]); d[20 + e] = __float_as_uint(cu * dg[e]);
                  d[24 + e] = __float_as_uint(dm[e]);
                }
              }
              // feature-major tiles for the weight-gradient kernel (zeros for rows past the end of the batch)
              float* o0 = P.dpT + (((size_t)q * (size_t)ntiles + (size_t)stile) * nhp + j0) * TC_DPT_PITCH + srow;
              const size_t sstride = (size_t)P.Q * (size_t)ntiles * nhp * TC_DPT_PITCH;
  #pragma unroll
              for (int e = 0; e < 4; ++e) {
                if (j0 + e < nhp && stores && !(P.debug & (1 << 20))) { o0[(size_t)e * TC_DPT_PITCH] = ds[e]; o0[sstride + (size_t)e * TC_DPT_PITCH] = dg[e]; o0[2 * sstride + (size_t)e * TC_DPT_PITCH] = dm[e]; }
              }
            }
            if (has_in) {
              const uint32_t s = c & 1u;
              if (c >= 2) { mbar_wait(bar_free + s, (phF >> s) & 1u); phF ^= 1u << s; }
              __syncwarp();
              tc_st28(a_thr + s * (uint32_t)BT_K, d);
              tc_wait_st();
              tc_fence_before();
              mbar_arrive(bar_A + s);
              ++c;
            }
          }
        }
        if (has_in) {
          // ---- d[h_h ; h_v] of this row: add this thread's columns to the d h of the two neighbour sites
          mbar_wait(bar_D, nsite & 1u);
          ++nsite;
          __syncwarp();
          tc_fence_after();
          if (P.h_chain) {
            if (nb1 >= 0) {                                          // vertical neighbour: the only contribution it gets through memory
              uint32_t d[TC_TPS];
              tc_ld28(tmem_base + lane_base + (uint32_t)nhp + kcol, d);
              tc_wait_ld();
              if (live && !(P.debug & (1 << 21))) {
                float4* dst = P.dh + tc_dh_index(nb1, stile, ntiles, NP, sub, 0, srow);
#pragma unroll
                for (int p = 0; p < TC_TPS / 4; ++p)
                  if (p < np_own)
                    __stcg(dst + (size_t)p * TC_ROWS, make_float4(__uint_as_float(d[4 * p]), __uint_as_float(d[4 * p + 1]),
                                                                 __uint_as_float(d[4 * p + 2]), __uint_as_float(d[4 * p + 3])));
              }
            }
            if (nb0 >= 0) {                                          // horizontal neighbour == the next site of this sweep: stays on the SM
              uint32_t d[TC_TPS];
              tc_ld28(tmem_base + lane_base + kcol, d);
              tc_wait_ld();
              uint32_t lo[16];
#pragma unroll
              for (int i = 0; i < 16; ++i) lo[i] = d[i];
              tc_st16(stash_thr, lo);
              tc_st4(stash_thr + 16u, make_uint4(d[16], d[17], d[18], d[19]));
#pragma unroll
              for (int i = 0; i < 8; ++i) hreg[i] = __uint_as_float(d[20 + i]);
              tc_wait_st();
            }
            have_h = nb0 >= 0;
          } else
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int nb = half ? nb1 : nb0;                         // warp-uniform
            if (nb < 0) continue;
            uint32_t d[TC_TPS];
            tc_ld28(tmem_base + lane_base + (uint32_t)half * (uint32_t)nhp + kcol, d);
            tc_wait_ld();
            if (live && !(P.debug & (1 << 21))) {
              float4* dst = P.dh + tc_dh_index(nb, stile, ntiles, NP, sub, 0, srow);
#pragma unroll
              for (int p = 0; p < TC_TPS / 4; ++p) {
                if (p < np_own) {
                  float4 v = __ldcg(dst + (size_t)p * TC_ROWS);
                  v.x += __uint_as_float(d[4 * p]); v.y += __uint_as_float(d[4 * p + 1]);
                  v.z += __uint_as_float(d[4 * p + 2]); v.w += __uint_as_float(d[4 * p + 3]);
                  __stcg(dst + (size_t)p * TC_ROWS, v);
                }
              }
            }
          }
          tc_fence_before();
          mbar_arrive(bar_Dfree);
        }
      }  // sites
      quarter_barrier(quarter);   // symS is rewritten by the next tile's set-up
    }  // tiles
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_base, TC_TMEM_COLS); }
}

}  // namespace rnnwf
