// Sweeps of the vanilla 2D RNN (2DTFIM_2DRNN/MDRNNcell.py:51-66, RNNwavefunction.py:73-178) in the split-precision regime,
// TWO 128-row tiles per CTA: connected configurations (MODE_CHILD), sampling (MODE_SAMPLE) and evaluation (MODE_EVAL).
//
// This cell has ONE tensor-core pass per lattice site (all units at once), so inside a tile the MMAs of a site and the
// element-wise work on their result are strictly serial: the generic kernel (sweep_tc_kernel.cuh, CELLK = 1) keeps the tensor
// pipe 16 % busy and its warps wait a quarter of the time.  Here a CTA runs two independent tiles ("groups") side by side,
// each with its own 8 element-wise warps (2 threads per row: thread (row, half) owns half of the units), its own MMA warp,
// its own accumulator and operand in TMEM -- one group's element-wise work fills the other's MMA / hand-over latency.  The
// weights (hi + lo FP16 images of [Wh; Wv], 2 x 14 KB at 56 units) are resident in shared memory for the whole kernel: no
// weight stream, no loader warp.
//
//   h_q = elu(x_h Uh + h_h Wh + x_v Uv + h_v Wv + b):   acc = [h_h | h_v] (128 x 2 nhp, FP16 hi + lo pairs scaled by 64)
//                                                              x [Wh; Wv]^T (hi + lo scaled by 1024), three MMAs per K = 16 step;
//   one-hot inputs and the bias from one table of 16 rows, b'(neighbour variant) + sigma_l dUh + sigma_u dUv, built at
//   kernel start from pack_tc_vanilla_bias_kernel's three tables; elu branch-free through ex2.approx;
//   two-logit head (real wavefunction): only the difference of the logits enters log p = -softplus(logit_other - logit_sigma);
//   the two threads of a row exchange their partial sums through shared memory behind a 64-thread mbarrier (one per lane
//   quarter and site parity) and take turns in finishing a site; running log P as a compensated FP32 sum, FP64 at the tile end;
//   E_loc(parent) += coef exp((logP' - logP)/2).
// History ring in global memory, [slot][unit half][4-unit chunk][row] 16-byte words = the operand's hi / lo words (no second
// split on the way back), written and read back by the same thread, a warp's access 512 contiguous bytes; slots assigned at
// kernel start by live range (one per lattice column on the zigzag path, so the rings of all CTAs mostly stay in the L2);
// prefix reuse from the parents' saved history exactly as in the generic kernel.
// TMEM per group: accumulator 64 columns + operand hi 64 + lo 64 (2 x 192 of 512 columns).
// 18 warps: 0, 1 issue the MMAs of group 0, 1; 2..17 element-wise (group, lane quarter = warp & 3, unit half).
// Sampling / evaluation: tile_rows live rows per tile (small batches cover both groups of all SMs), the saved state of the
// samples (history, weight-gradient tile image, running log P, heads' outputs) written from the registers; SAMPLE: the finishing
// thread draws the symbol and signals the row's other thread through a second set of pair mbarriers.
// Every wait is bounded and traps instead of hanging.  Measured history and stall analysis: profiles/r2d_summary.md.
#pragma once
#include "sweep_tc.cuh"

namespace rnnwf {

constexpr int V2_ROWS = 128;
constexpr int V2_N = 64;                              // accumulator columns (units, natural order; nhp <= 64)
constexpr int V2_GROUPS = 2;
constexpr int V2_GWARPS = 8;                          // element-wise warps per group: (lane quarter 4) x (unit half 2)
constexpr int V2_THREADS = 32 * (2 + V2_GROUPS * V2_GWARPS);   // warps 0, 1: MMA issue of group 0, 1 (576 threads; registers are granted per 4 warps, so 96 each as with 640)
constexpr int V2_TABP = V2_N + 4;                     // pitch of the input / bias table (rows of one warp's lanes fall into distinct banks)
constexpr int V2_GCOLS = 192;                         // TMEM columns per group: [acc 64 | operand hi 64 | operand lo 64]

// weight image: [part 2: hi, lo][k-core 2 nhp / 8][n-core 8][8 rows (n)][8 halves (k)],  B[n][k] = 1024 * W_slot(k)[unit(k)][n]
__host__ __device__ inline int v2_image_bytes(int nhp) { return 2 * (2 * nhp / 8) * (V2_N / 8) * 128; }

struct V2Smem {
  int off_B, off_bias, off_headw, off_step, off_sym, off_part, off_lp, off_bar, off_tmem, total;
  __host__ __device__ V2Smem(int nhp, int Q) {
    int o = 0;
    off_B = o;     o += v2_image_bytes(nhp);
    off_bias = o;  o += 16 * V2_TABP * 4;             // b' (neighbour variant) + sym_l dUh + sym_u dUv: [variant 4][sym_l 2][sym_u 2][68]
    off_headw = o; o += V2_N * 4 + 16;                // head: w[unit][1] - w[unit][0], then b[1] - b[0]
    off_step = o;  o += Q * 8;                        // per site: (nbr0, nbr1, config offset, -)
    off_sym = o;   o += V2_GROUPS * ((Q * V2_ROWS + 15) & ~15);   // symbols of the group's tile, one byte per (site, row)
    off_part = o;  o += 2 * V2_GROUPS * V2_ROWS * 4;  // partial logit difference of the row's other unit half: [site parity][group][row]
    off_lp = o;    o += V2_GROUPS * V2_ROWS * 8;      // tile end: the upper half's share of the row's log P (double)
    off_bar = o;   o += 32 + 2 * (2 * V2_GROUPS * 4 * 8);   // operand / accumulator barriers, then [site parity][group][lane quarter] pair barriers: partial sums written; (sampling) symbol drawn
    off_tmem = o;  o += 16;
    total = o;
  }
};
inline bool v2_supported(int nh, int nhp, int Q, int complex_head, int u1) {
  return !complex_head && !u1 && nhp >= 8 && nhp <= V2_N && nhp % 8 == 0 && nh <= nhp && V2Smem(nhp, Q).total <= 200 * 1024;
}
// ring bytes per CTA: [group][slot ring_len][unit half 2][chunk][row 128] 16-byte words = 4 bytes per (row, unit)
__host__ __device__ inline long long v2_ring_bytes(int ring_len, int nhp) { return (long long)V2_GROUPS * ring_len * V2_ROWS * nhp * 4; }

template <int UNUSED = 0>
__global__ void pack_van2_kernel(const float* __restrict__ Wh, const float* __restrict__ Wv, int nh, int nhp, __half* __restrict__ out) {
  const int kc = 2 * nhp / 8;
  const long long per_part = (long long)kc * (V2_N / 8) * 64;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < 2 * per_part; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int e = (int)(r % 8); r /= 8;
    const int r8 = (int)(r % 8); r /= 8;
    const int ni = (int)(r % (V2_N / 8)); r /= (V2_N / 8);
    const int ki = (int)(r % kc); r /= kc;
    const int part = (int)r;
    const int n = ni * 8 + r8, k = ki * 8 + e;
    const int slot = k / nhp, unit = k - slot * nhp;
    float v = 0.f;
    if (unit < nh && n < nh) v = (slot == 0 ? Wh : Wv)[(long long)unit * nh + n] * TC_SPLIT_WS;
    const __half hi = __float2half_rn(v);
    out[idx] = (part == 0) ? hi : __float2half_rn(v - __half2float(hi));
  }
}

// NU4: 4-unit chunks per thread (units per thread = 4 NU4 = nhp / 2)
template <int NU4, int MODE>
__global__ void __launch_bounds__(V2_THREADS, 1) sweep_van2_kernel(const SweepParams P) {
  constexpr bool SAMPLE = (MODE == MODE_SAMPLE);        // symbols drawn site by site
  constexpr bool SV = (MODE != MODE_CHILD);             // sampling / evaluation: own rows, row spreading, saved state, log P out
  constexpr int U2 = 4 * NU4;                         // units per thread
  constexpr int NHP = 2 * U2;                         // padded units
  constexpr int KS = NHP / 8;                         // K = 16 steps over [h_h | h_v]
  constexpr float HS = TC_SPLIT_HS, DS = 1.0f / (TC_SPLIT_HS * TC_SPLIT_WS);
  extern __shared__ __align__(128) unsigned char v2_smem[];
  const V2Smem L(NHP, P.Q);
  unsigned char* const smem = v2_smem;
  float* const biasS = reinterpret_cast<float*>(smem + L.off_bias);
  float* const headW = reinterpret_cast<float*>(smem + L.off_headw);          // [64] logit-difference weights, then the bias difference
  short4* const stepS = reinterpret_cast<short4*>(smem + L.off_step);
  uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
  uint64_t* const bar_op = bars + 0;                  // [2] group's operand written (256 arrivals)
  uint64_t* const bar_acc = bars + 2;                 // [2] group's accumulator complete (tcgen05.commit)
  uint64_t* const bar_pair = bars + 4;                // [site parity 2][group 2][quarter 4]: the partner warp's partial logits are written (32 arrivals)
  uint64_t* const bar_symp = bars + 4 + 2 * V2_GROUPS * 4;   // same shape, sampling: the finishing warp has drawn the site's symbols (32 arrivals)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nh = P.nh;

  if (tid == 0) {
    for (int g = 0; g < V2_GROUPS; ++g) { mbar_init(bar_op + g, V2_GWARPS * 32); mbar_init(bar_acc + g, 1); }
    for (int i = 0; i < 2 * V2_GROUPS * 4; ++i) { mbar_init(bar_pair + i, 32); mbar_init(bar_symp + i, 32); }
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, 512);
  for (int i = tid; i < v2_image_bytes(NHP) / 16; i += V2_THREADS)
    reinterpret_cast<uint4*>(smem + L.off_B)[i] = __ldg(reinterpret_cast<const uint4*>(P.tc_B) + i);
  for (int i = tid; i < 16 * V2_TABP; i += V2_THREADS) {
    const int which = i / V2_TABP, j = i % V2_TABP;
    const int var = which >> 2, sl = (which >> 1) & 1, su = which & 1;
    biasS[i] = (j < NHP) ? P.tc_bias[var * NHP + j] + (sl ? P.tc_bias[4 * NHP + j] : 0.f) + (su ? P.tc_bias[5 * NHP + j] : 0.f) : 0.f;
  }
  // two-class softmax: only the difference of the logits enters log p
  for (int i = tid; i <= V2_N; i += V2_THREADS)
    headW[i] = (i == V2_N) ? P.head_b[1] - P.head_b[0] : ((i < nh) ? P.head_w[i * P.ND + 1] - P.head_w[i * P.ND] : 0.f);
  // Ring slots (assigned once on the host by live range, build_model: StepMeta::pad2).  A site's state is read back once,
  // as the older neighbour of a later site t (the other neighbour of t is site t - 1, handed over in registers); that read is
  // issued at the top of iteration t - 1, before iteration t - 1 stores its own state, so site q may take the slot of any
  // site whose reader t <= q + 1: on the zigzag path one slot per lattice column, half the look-back window.
  // w = own slot | slot of the older neighbour << 8.
  for (int i = tid; i < P.Q; i += V2_THREADS) {
    const StepMeta* m = P.steps + i;
    const int nbo = (i > 0) ? ((m->nbr[0] == i - 1) ? m->nbr[1] : m->nbr[0]) : -1;
    stepS[i] = make_short4(m->nbr[0], m->nbr[1], m->site, (short)((m->pad2 & 0xff) | ((nbo >= 0 ? (P.steps[nbo].pad2 & 0xff) : 0) << 8)));
  }
  fence_proxy_async_smem();                           // the weight image is read by the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const long long total_rows = (P.rows_dev != nullptr) ? (long long)(*P.rows_dev) : P.rows;
  // sampling: tile_rows live rows per 128-row MMA tile so that small batches cover all SMs, and the tiles cover the batch
  // rounded up to 128 rows (the saved tile images' rows past the batch are written as zeros)
  const int VR = SV ? P.tile_rows : V2_ROWS;
  const long long rows_cover = SV ? ((total_rows + V2_ROWS - 1) / V2_ROWS) * V2_ROWS : total_rows;
  const long long ntiles = (rows_cover + VR - 1) / VR;
  // tile sequence of group g of this CTA: tiles 2 (blockIdx.x + k gridDim.x) + g
  auto tile_of = [&](int g, long long k) -> long long { return 2 * ((long long)blockIdx.x + k * (long long)gridDim.x) + g; };
  auto tile_q0 = [&](long long row0) -> int { return (!SV && P.par_hist != nullptr) ? (int)P.children[row0].ta : 0; };   // single-flip lists: first changed site

  if (warp < V2_GROUPS) {
    // ======================= MMA warp of group `warp` =======================
    const int g = warp;
    const bool leader = elect_one();
    const uint32_t acc = tmem_base + (uint32_t)(g * V2_GCOLS), a_hi = acc + 64u, a_lo = acc + 128u;
    const uint32_t B_LBO = (V2_N / 8) * 128, SBO = 128;
    const uint64_t b_hi = tc_desc(smem_u32(smem + L.off_B), B_LBO, SBO);
    const uint64_t b_lo = tc_desc(smem_u32(smem + L.off_B) + (uint32_t)(v2_image_bytes(NHP) / 2), B_LBO, SBO);
    constexpr uint64_t B_KSTEP = (2ull * (V2_N / 8) * 128) >> 4;
    const uint32_t idesc = tc_idesc(V2_ROWS, V2_N);
    uint32_t n = 0;
    for (long long k = 0;; ++k) {
      const long long tile = tile_of(g, k);
      if (tile >= ntiles) break;
      const int q0 = tile_q0(tile * VR);
      for (int q = q0; q < P.Q; ++q, ++n) {
        mbar_wait(bar_op + g, n & 1u);
        tc_fence_after();
        if (leader) {
#pragma unroll
          for (int ks = 0; ks < KS; ++ks) {
            tc_mma_f16_ts(acc, a_hi + (uint32_t)ks * 8u, b_hi + (uint64_t)ks * B_KSTEP, idesc, ks > 0 ? 1u : 0u);
            tc_mma_f16_ts(acc, a_lo + (uint32_t)ks * 8u, b_hi + (uint64_t)ks * B_KSTEP, idesc, 1u);
            tc_mma_f16_ts(acc, a_hi + (uint32_t)ks * 8u, b_lo + (uint64_t)ks * B_KSTEP, idesc, 1u);
          }
          tc_commit(bar_acc + g);
        }
        __syncwarp();
      }
    }
  } else {
    // ======================= element-wise warps: group, lane quarter, unit half =======================
    const int g = (warp - 2) / V2_GWARPS;
    const int wg = (warp - 2) % V2_GWARPS;
    const int quarter = warp & 3;                       // TMEM lanes this warp may touch: 32 quarter .. + 31
    const int half = wg >> 2;                           // (four consecutive warps cover the four lane quarters: warp & 3)
    const int row = (quarter << 5) + lane;
    const int u0 = half * U2;                           // this thread's units u0 .. u0 + U2 - 1
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const uint32_t acc = tmem_base + lane_base + (uint32_t)(g * V2_GCOLS);
    const uint32_t a_hi = acc + 64u, a_lo = acc + 128u;
    uint8_t* const symS = smem + L.off_sym + (size_t)g * ((P.Q * V2_ROWS + 15) & ~15);
    float* const partS = reinterpret_cast<float*>(smem + L.off_part) + g * V2_ROWS;
    double* const lpS = reinterpret_cast<double*>(smem + L.off_lp) + g * V2_ROWS;
    // ring of this thread: [slot][unit half][chunk][row] 16-byte operand words (a warp's store covers 512 contiguous bytes)
    uint4* const ring = reinterpret_cast<uint4*>(P.ring) + (((size_t)blockIdx.x * V2_GROUPS + g) * (size_t)P.ring_len * 2 + half) * (NU4 * V2_ROWS) + row;
    constexpr int RING_SLOT = 2 * NU4 * V2_ROWS;        // uint4 per slot
    const int gbar = 1 + g;                             // named barrier of the group's 256 threads
    uint32_t n = 0;

    // the thread's 2 NU4 hi words and 2 NU4 lo words of neighbour slot `slot` into the operand, in the widest stores that fit
    auto put_words = [&](uint32_t t, const uint32_t (&w)[2 * NU4]) {
      constexpr int W = 2 * NU4;
      if constexpr (W >= 8) tc_st8(t, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
      if constexpr (W >= 16) tc_st8(t + 8u, w[8], w[9], w[10], w[11], w[12], w[13], w[14], w[15]);
      constexpr int O = W / 8 * 8;
      if constexpr (W - O >= 4) tc_st4(t + (uint32_t)O, make_uint4(w[O], w[O + 1], w[O + 2], w[O + 3]));
      constexpr int O2 = O + (W - O) / 4 * 4;
      if constexpr (W - O2 >= 2) tc_st2(t + (uint32_t)O2, make_uint2(w[O2], w[O2 + 1]));
    };
    auto put_operand = [&](int slot, const uint32_t (&hi)[2 * NU4], const uint32_t (&lo)[2 * NU4]) {
      const uint32_t c0 = (uint32_t)((slot * NHP + u0) >> 1);
      put_words(a_hi + c0, hi);
      put_words(a_lo + c0, lo);
    };
    for (long long k = 0;; ++k) {
      const long long tile = tile_of(g, k);
      if (tile >= ntiles) break;
      const long long row0 = tile * VR;
      const int nrows = (int)((total_rows - row0 < (long long)VR) ? (total_rows - row0 > 0 ? total_rows - row0 : 0) : (long long)VR);
      const bool live = row < nrows;
      const bool stores = SV && row < VR && row0 + row < rows_cover;   // owns a row of the saved tile images (zeros past the batch)
      const int q0 = tile_q0(row0);
      Child ch;
      ch.parent = 0; ch.ta = -1; ch.tb = -1; ch.coef = 0.f; ch.ka = ch.kb = 0; ch.pad = 0;
      if (!SV && live) ch = P.children[row0 + row];
      // ---- symbols of the tile (teacher forcing: the parent's configuration with the child's flip; sampling: drawn site by site)
      asm volatile("bar.sync %0, %1;" ::"r"(gbar), "n"(V2_GWARPS * 32) : "memory");      // the previous tile's symbols are no longer read
      if (!SAMPLE) {
        const int8_t* c = P.cfg_in + (SV ? (live ? row0 + row : 0) : (long long)ch.parent) * P.nsites;
        for (int t = half; t < P.Q; t += 2) {
          int sy = 0;
          if (live) { sy = c[stepS[t].z] & 1; if (!SV && (t == ch.ta || t == ch.tb)) sy ^= 1; }
          symS[t * V2_ROWS + row] = (uint8_t)sy;
        }
      }
      asm volatile("bar.sync %0, %1;" ::"r"(gbar), "n"(V2_GWARPS * 32) : "memory");
      // hidden state of an earlier site for this thread's units, as operand words: own ring (sites >= q0 of this tile; `rs` =
      // the site's ring slot) or the parent's saved FP32 history (split here)
      auto load_h = [&](int nb, int rs, uint32_t (&hi)[2 * NU4], uint32_t (&lo)[2 * NU4]) {
        if (nb < 0 || !live) {
#pragma unroll
          for (int i = 0; i < 2 * NU4; ++i) { hi[i] = 0u; lo[i] = 0u; }
        } else if (nb >= q0 || P.par_hist == nullptr) {
          const uint4* src = ring + (size_t)rs * RING_SLOT;
#pragma unroll
          for (int c = 0; c < NU4; ++c) {
            const uint4 w = __ldcg(src + c * V2_ROWS);
            hi[2 * c] = w.x; hi[2 * c + 1] = w.y; lo[2 * c] = w.z; lo[2 * c + 1] = w.w;
          }
        } else {
          const float* src = P.par_hist + ((size_t)nb * (size_t)P.par_rows + (size_t)ch.parent) * P.nhp + u0;
#pragma unroll
          for (int c = 0; c < NU4; ++c) {
            float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
            if (u0 + 4 * c < P.nhp) x = __ldg(reinterpret_cast<const float4*>(src) + c);
            if (u0 + 4 * c >= nh) x.x = 0.f;
            if (u0 + 4 * c + 1 >= nh) x.y = 0.f;
            if (u0 + 4 * c + 2 >= nh) x.z = 0.f;
            if (u0 + 4 * c + 3 >= nh) x.w = 0.f;
            split_f16x2(x.x * HS, x.y * HS, hi[2 * c], lo[2 * c]);
            split_f16x2(x.z * HS, x.w * HS, hi[2 * c + 1], lo[2 * c + 1]);
          }
        }
      };
      // ---- operand of the tile's first site: both neighbours from the history
      short4 eq = stepS[q0];
      {
        uint32_t hi[2 * NU4], lo[2 * NU4];
        load_h(eq.x, 0, hi, lo);                            // (before the tile's first site: never the ring)
        __syncwarp();
        put_operand(0, hi, lo);
        load_h(eq.y, 0, hi, lo);
        put_operand(1, hi, lo);
        tc_wait_st();
        tc_fence_before();
        mbar_arrive(bar_op + g);
      }
      float lps = 0.f, lpc = 0.f;                          // (compensated FP32 sum: the FP64 pipe is narrow)
      for (int q = q0; q < P.Q; ++q, ++n) {
        const bool has_next = q + 1 < P.Q;
        // the older neighbour of the next site (not this site's output): fetched before the wait for this site's accumulators
        short4 en = make_short4(-1, -1, 0, 0);
        int carry_slot = 0;
        uint32_t ohi[2 * NU4], olo[2 * NU4];
        if (has_next) {
          en = stepS[q + 1];
          carry_slot = (en.x == q) ? 0 : 1;                 // the next site's neighbour that is this site: horizontal or (row turn) vertical
          const int nbo = carry_slot == 0 ? en.y : en.x;
          load_h(nbo, (en.w >> 8) & 0xff, ohi, olo);
        }
        if (SAMPLE && q > q0 && half != (int)((n - 1u) & 1u))   // the previous site's symbol: drawn by the other thread of the row
          mbar_wait(bar_symp + (((n - 1u) & 1u) * V2_GROUPS + g) * 4 + quarter, ((n - 1u) >> 1) & 1u);
        const int sl = (eq.x >= 0) ? (int)symS[eq.x * V2_ROWS + row] : 0;
        const int su = (eq.y >= 0) ? (int)symS[eq.y * V2_ROWS + row] : 0;
        const float* bv = biasS + ((((eq.x >= 0) ? 1 : 0) + ((eq.y >= 0) ? 2 : 0)) * 4 + sl * 2 + su) * V2_TABP + u0;
        uint4* const rdst = ring + (size_t)(eq.w & 0xff) * RING_SLOT;
        mbar_wait(bar_acc + g, n & 1u);
        __syncwarp();
        tc_fence_after();
        uint32_t d[NU4][4];
#pragma unroll
        for (int c = 0; c < NU4; ++c) tc_ld4(acc + (uint32_t)(u0 + 4 * c), d[c]);
        tc_wait_ld();
        // ---- new hidden state in place of the accumulators (no ordered instruction in this loop: loads and MUFUs pipeline freely)
        float lg = 0.f;
        // sampling: this site's rows of the saved state (history FP32 [site][row][unit]; tile image, zeros past the batch)
        float hsv[4];
        float* const sv_h = (SV && live && P.sv_hist != nullptr) ? P.sv_hist + ((size_t)q * (size_t)P.sv_rows + (size_t)(row0 + row)) * NHP + u0 : nullptr;
        const int hnc = ((NHP + 15) / 16) * 2;
        float* const sv_i = (SV && stores && P.sv_himg != nullptr)
                                ? P.sv_himg + ((size_t)q * (size_t)((P.sv_rows + V2_ROWS - 1) / V2_ROWS) + (size_t)((row0 + row) / V2_ROWS)) * (size_t)(32 * hnc * 32) +
                                      (size_t)(((row0 + row) % V2_ROWS) >> 2) * hnc * 32 + ((row0 + row) & 3)
                                : nullptr;
#pragma unroll
        for (int c = 0; c < NU4; ++c) {
          const float4 b4 = *reinterpret_cast<const float4*>(bv + 4 * c);
          const float4 w4 = *reinterpret_cast<const float4*>(headW + u0 + 4 * c);
          const float bb[4] = {b4.x, b4.y, b4.z, b4.w}, ww[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float x = fmaf(__uint_as_float(d[c][e]), DS, bb[e]);
            // elu, branch-free: exp(min(x, 0)) - 1 by ex2.approx (2 ulp, the class of expf); padding units: zero weights and biases -> 0
            const float em = ex2_approx(fminf(x, 0.f) * 1.4426950408889634f) - 1.0f;
            const float hv = live ? ((x > 0.f) ? x : em) : 0.f;
            lg = fmaf(hv, ww[e], lg);
            d[c][e] = __float_as_uint(hv * HS);
            if (SV) {
              hsv[e] = hv;
              // hidden-state tile image for the weight-gradient kernel (dw_tc.cuh), straight from the registers
              if (sv_i != nullptr) { const int j = u0 + 4 * c + e; sv_i[(j >> 3) * 32 + (j & 7) * 4] = (P.sv_hist != nullptr) ? hv : 0.f; }
            }
          }
          if (SV && sv_h != nullptr) *reinterpret_cast<float4*>(sv_h + 4 * c) = make_float4(hsv[0], hsv[1], hsv[2], hsv[3]);
        }
        if (SV && sv_i != nullptr && half == 1) {       // the image covers the units rounded up to 16
          for (int j = NHP; j < hnc * 8; ++j) sv_i[(j >> 3) * 32 + (j & 7) * 4] = 0.f;
        }
        // ---- hi / lo words of the new state: the ring image of this site (the same thread reads it back as a vertical
        // neighbour ~2 Nx sites later) and, all MMAs of this site having completed, the next site's operand
        uint32_t nhi[2 * NU4], nlo[2 * NU4];
#pragma unroll
        for (int c = 0; c < NU4; ++c) {
          split_f16x2(__uint_as_float(d[c][0]), __uint_as_float(d[c][1]), nhi[2 * c], nlo[2 * c]);
          split_f16x2(__uint_as_float(d[c][2]), __uint_as_float(d[c][3]), nhi[2 * c + 1], nlo[2 * c + 1]);
        }
        if (live) {
#pragma unroll
          for (int c = 0; c < NU4; ++c) __stcg(rdst + c * V2_ROWS, make_uint4(nhi[2 * c], nhi[2 * c + 1], nlo[2 * c], nlo[2 * c + 1]));
        }
        // ---- operand of the next site
        if (has_next) {
          put_operand(carry_slot, nhi, nlo);
          put_operand(1 - carry_slot, ohi, olo);
          tc_wait_st();
          tc_fence_before();
          mbar_arrive(bar_op + g);
        }
        // ---- head: the two unit halves of a row meet in shared memory; the halves take turns in finishing a site
        {
          const uint32_t b = n & 1u;
          float* const ps = partS + b * (V2_GROUPS * V2_ROWS);
          uint64_t* const pb = bar_pair + (b * V2_GROUPS + g) * 4 + quarter;
          if (half != (int)b) {
            ps[row] = lg;
            mbar_arrive(pb);
          } else {
            mbar_wait(pb, (n >> 1) & 1u);
            if (!SV) {
              if (live) {
                // log p(sy) = -softplus(logit(other) - logit(sy))
                const float dl = lg + ps[row] + headW[V2_N];
                const float z = symS[q * V2_ROWS + row] ? -dl : dl;
                const float y = -(fmaxf(z, 0.f) + logf(1.0f + expf(-fabsf(z)))) - lpc;
                const float t = lps + y;
                lpc = (t - lps) - y;
                lps = t;
              }
            } else {
              // sampling: draw the site's symbol (inverse CDF on the supplied uniform or the Philox stream, as the generic
              // kernel), the running log P in FP64 through shared memory (the two threads of a row alternate), the heads'
              // outputs for the reverse sweep; then publish the symbol to the row's other thread
              int sy = 0;
              if (live) {
                const float dl = lg + ps[row] + headW[V2_N];                 // logit(1) - logit(0)
                const float ee = expf(-fabsf(dl)), rr = 1.0f / (1.0f + ee);
                const float p1 = (dl >= 0.f) ? rr : ee * rr, p0 = (dl >= 0.f) ? ee * rr : rr;
                float u = 0.f;
                if (SAMPLE) u = (P.uniforms != nullptr) ? P.uniforms[(size_t)(row0 + row) * P.Tsym + q]
                                                        : philox_uniform(P.seed, P.row_offset + (unsigned long long)(row0 + row), (uint32_t)q);
                sy = SAMPLE ? (((double)p0 <= (double)u * ((double)p0 + (double)p1)) ? 1 : 0) : (int)symS[q * V2_ROWS + row];
                const float z = sy ? -dl : dl;
                const double run = ((q > q0) ? lpS[row] : 0.0) - (double)(fmaxf(z, 0.f) + logf(1.0f + expf(-fabsf(z))));
                lpS[row] = run;
                if (P.sv_cum != nullptr) {
                  P.sv_cum[(size_t)q * (size_t)P.sv_rows + (size_t)(row0 + row)] = make_double2(run, 0.0);
                  P.sv_head[(size_t)q * (size_t)P.sv_rows + (size_t)(row0 + row)] = make_float4(p0, p1, 0.f, 0.f);
                }
              }
              if (SAMPLE) {
                symS[q * V2_ROWS + row] = (uint8_t)sy;
                mbar_arrive(bar_symp + (b * V2_GROUPS + g) * 4 + quarter);
              }
            }
          }
        }
        eq = en;
      }
      // ---- tile result: E_loc(parent) += <sigma'|H|sigma> psi(sigma') / psi(sigma)
      if (SV) {
        asm volatile("bar.sync %0, %1;" ::"r"(gbar), "n"(V2_GWARPS * 32) : "memory");      // the last site's symbol and running log P are written
        if (live) {
          if (half == 0) {
            if (P.out_logp != nullptr) P.out_logp[row0 + row] = lpS[row];
            if (P.out_phase != nullptr) P.out_phase[row0 + row] = 0.0;
          }
          if (SAMPLE && P.cfg_out != nullptr) {
            int8_t* c = P.cfg_out + (size_t)(row0 + row) * P.nsites;
            for (int t = half; t < P.Q; t += 2) c[stepS[t].z] = (int8_t)symS[t * V2_ROWS + row];
          }
        }
        continue;
      }
      double rlogp = (double)lps - (double)lpc;            // this thread's share of the row's log P (the sites it finished)
      if (half == 1) lpS[row] = rlogp;
      asm volatile("bar.sync %0, %1;" ::"r"(gbar), "n"(V2_GWARPS * 32) : "memory");
      if (half == 0 && live) {
        rlogp += lpS[row];
        double base_lp = P.par_logp[ch.parent];
        if (q0 > 0 && P.par_cum != nullptr) base_lp -= P.par_cum[(size_t)(q0 - 1) * (size_t)P.par_rows + ch.parent].x;
        atomicAdd(P.eloc + 2 * (size_t)ch.parent, exp(0.5 * (rlogp - base_lp)) * (double)ch.coef);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_base, 512); }
}

}  // namespace rnnwf
