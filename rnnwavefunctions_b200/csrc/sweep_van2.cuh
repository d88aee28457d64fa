// Connected-configuration sweep of the vanilla 2D RNN (2DTFIM_2DRNN/MDRNNcell.py:51-66, RNNwavefunction.py:73-178) in the
// split-precision regime, TWO 128-row tiles per CTA.
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
//   one-hot inputs and the bias from three small tables (b' of the neighbour variant, dUh, dUv: pack_tc_vanilla_bias_kernel);
//   two-logit head (real wavefunction): the two threads of a row exchange their partial logits through shared memory behind
//   a 64-thread mbarrier and take turns in finishing a site (softmax, log in FP64, running log P in FP64);
//   E_loc(parent) += coef exp((logP' - logP)/2).
// History ring in global memory ([slot][row][units of this thread] as the operand's hi / lo words -- written and read back
// by the same thread, no second split);
// prefix reuse from the parents' saved history exactly as in the generic kernel.
// TMEM per group: accumulator 64 columns + operand hi 64 + lo 64 (2 x 192 of 512 columns).
// Every wait is bounded and traps instead of hanging.
#pragma once
#include "sweep_tc.cuh"

namespace rnnwf {

constexpr int V2_ROWS = 128;
constexpr int V2_N = 64;                              // accumulator columns (units, natural order; nhp <= 64)
constexpr int V2_GROUPS = 2;
constexpr int V2_GWARPS = 8;                          // element-wise warps per group: (lane quarter 4) x (unit half 2)
constexpr int V2_THREADS = 32 * (4 + V2_GROUPS * V2_GWARPS);   // warps 0, 1: MMA issue of group 0, 1; warps 2, 3 spare
constexpr int V2_GCOLS = 192;                         // TMEM columns per group: [acc 64 | operand hi 64 | operand lo 64]

// weight image: [part 2: hi, lo][k-core 2 nhp / 8][n-core 8][8 rows (n)][8 halves (k)],  B[n][k] = 1024 * W_slot(k)[unit(k)][n]
__host__ __device__ inline int v2_image_bytes(int nhp) { return 2 * (2 * nhp / 8) * (V2_N / 8) * 128; }

struct V2Smem {
  int off_B, off_bias, off_headw, off_step, off_sym, off_part, off_lp, off_bar, off_tmem, total;
  __host__ __device__ V2Smem(int nhp, int Q) {
    int o = 0;
    off_B = o;     o += v2_image_bytes(nhp);
    off_bias = o;  o += 6 * V2_N * 4;                 // b' (4 variants), dUh, dUv: [6][64]
    off_headw = o; o += V2_N * 2 * 4 + 16;            // head weights [unit][2], biases [2]
    off_step = o;  o += Q * 8;                        // per site: (nbr0, nbr1, config offset, -)
    off_sym = o;   o += V2_GROUPS * ((Q * V2_ROWS + 15) & ~15);   // symbols of the group's tile, one byte per (site, row)
    off_part = o;  o += 2 * V2_GROUPS * V2_ROWS * 8;  // partial logits of the row's other unit half: [site parity][group][row] float2
    off_lp = o;    o += V2_GROUPS * V2_ROWS * 8;      // tile end: the upper half's share of the row's log P (double)
    off_bar = o;   o += 32 + 2 * V2_GROUPS * 4 * 8;   // operand / accumulator barriers, then [site parity][group][lane quarter] pair barriers
    off_tmem = o;  o += 16;
    total = o;
  }
};
inline bool v2_supported(int nh, int nhp, int Q, int complex_head, int u1) {
  return !complex_head && !u1 && nhp <= V2_N && (nhp == 32 || nhp == 48 || nhp == 56 || nhp == 64) && nh <= nhp && V2Smem(nhp, Q).total <= 200 * 1024;
}
// ring bytes per CTA: [group][slot ring_len][row 128][nhp] FP32
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

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// NU4: 4-unit chunks per thread (units per thread = 4 NU4 = nhp / 2)
template <int NU4>
__global__ void __launch_bounds__(V2_THREADS, 1) sweep_van2_kernel(const SweepParams P) {
  constexpr int U2 = 4 * NU4;                         // units per thread
  constexpr int NHP = 2 * U2;                         // padded units
  constexpr int KS = NHP / 8;                         // K = 16 steps over [h_h | h_v]
  constexpr float HS = TC_SPLIT_HS, DS = 1.0f / (TC_SPLIT_HS * TC_SPLIT_WS);
  extern __shared__ __align__(128) unsigned char v2_smem[];
  const V2Smem L(NHP, P.Q);
  unsigned char* const smem = v2_smem;
  float* const biasS = reinterpret_cast<float*>(smem + L.off_bias);
  float* const headW = reinterpret_cast<float*>(smem + L.off_headw);          // [64][2], then [2] biases at 128
  short4* const stepS = reinterpret_cast<short4*>(smem + L.off_step);
  uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
  uint64_t* const bar_op = bars + 0;                  // [2] group's operand written (256 arrivals)
  uint64_t* const bar_acc = bars + 2;                 // [2] group's accumulator complete (tcgen05.commit)
  uint64_t* const bar_pair = bars + 4;                // [site parity 2][group 2][quarter 4]: the partner warp's partial logits are written (32 arrivals)
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nh = P.nh;

  if (tid == 0) {
    for (int g = 0; g < V2_GROUPS; ++g) { mbar_init(bar_op + g, V2_GWARPS * 32); mbar_init(bar_acc + g, 1); }
    for (int i = 0; i < 2 * V2_GROUPS * 4; ++i) mbar_init(bar_pair + i, 32);
    fence_mbar_init();
  }
  if (warp == 2) tc_alloc(tmem_slot, 512);
  for (int i = tid; i < v2_image_bytes(NHP) / 16; i += V2_THREADS)
    reinterpret_cast<uint4*>(smem + L.off_B)[i] = __ldg(reinterpret_cast<const uint4*>(P.tc_B) + i);
  for (int i = tid; i < 6 * V2_N; i += V2_THREADS) {
    const int which = i / V2_N, j = i % V2_N;
    biasS[i] = (j < NHP) ? P.tc_bias[which * NHP + j] : 0.f;
  }
  for (int i = tid; i < V2_N * 2 + 2; i += V2_THREADS) {
    if (i < V2_N * 2) { const int j = i >> 1, d = i & 1; headW[i] = (j < nh) ? P.head_w[j * P.ND + d] : 0.f; }
    else headW[i] = P.head_b[i - V2_N * 2];
  }
  for (int i = tid; i < P.Q; i += V2_THREADS) stepS[i] = make_short4(P.steps[i].nbr[0], P.steps[i].nbr[1], P.steps[i].site, 0);
  fence_proxy_async_smem();                           // the weight image is read by the tensor core (async proxy)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const long long total_rows = (P.rows_dev != nullptr) ? (long long)(*P.rows_dev) : P.rows;
  const long long ntiles = (total_rows + V2_ROWS - 1) / V2_ROWS;
  // tile sequence of group g of this CTA: tiles 2 (blockIdx.x + k gridDim.x) + g
  auto tile_of = [&](int g, long long k) -> long long { return 2 * ((long long)blockIdx.x + k * (long long)gridDim.x) + g; };
  auto tile_q0 = [&](long long row0) -> int { return (P.par_hist != nullptr) ? (int)P.children[row0].ta : 0; };   // single-flip lists: first changed site

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
      const int q0 = tile_q0(tile * V2_ROWS);
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
  } else if (warp >= 4) {
    // ======================= element-wise warps: group, lane quarter, unit half =======================
    const int g = (warp - 4) / V2_GWARPS;
    const int wg = (warp - 4) % V2_GWARPS;
    const int quarter = warp & 3;                       // TMEM lanes this warp may touch: 32 quarter .. + 31
    const int half = wg >> 2;                           // (warps 4..7 / 12..15: half 0; 8..11 / 16..19: half 1 -- warp & 3 runs 0..3 in each)
    const int row = (quarter << 5) + lane;
    const int u0 = half * U2;                           // this thread's units u0 .. u0 + U2 - 1
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const uint32_t acc = tmem_base + lane_base + (uint32_t)(g * V2_GCOLS);
    const uint32_t a_hi = acc + 64u, a_lo = acc + 128u;
    uint8_t* const symS = smem + L.off_sym + (size_t)g * ((P.Q * V2_ROWS + 15) & ~15);
    float2* const partS = reinterpret_cast<float2*>(smem + L.off_part) + g * V2_ROWS;
    double* const lpS = reinterpret_cast<double*>(smem + L.off_lp) + g * V2_ROWS;
    float* const ring = P.ring + ((size_t)blockIdx.x * V2_GROUPS + g) * (size_t)P.ring_len * V2_ROWS * NHP + (size_t)row * NHP + u0;
    const int gbar = 1 + g;                             // named barrier of the group's 256 threads
    uint32_t n = 0;

    // operand words of 4-unit chunk c: (hi pair 0, hi pair 1, lo pair 0, lo pair 1) of 64 h -- also the ring's storage form
    auto put_operand = [&](int slot, const uint4 (&w)[NU4]) {
      const uint32_t c0 = (uint32_t)((slot * NHP + u0) >> 1);
#pragma unroll
      for (int c = 0; c < NU4; ++c) {
        tc_st2(a_hi + c0 + 2u * c, make_uint2(w[c].x, w[c].y));
        tc_st2(a_lo + c0 + 2u * c, make_uint2(w[c].z, w[c].w));
      }
    };
    for (long long k = 0;; ++k) {
      const long long tile = tile_of(g, k);
      if (tile >= ntiles) break;
      const long long row0 = tile * V2_ROWS;
      const int nrows = (int)((total_rows - row0 < (long long)V2_ROWS) ? (total_rows - row0) : (long long)V2_ROWS);
      const bool live = row < nrows;
      const int q0 = tile_q0(row0);
      Child ch;
      ch.parent = 0; ch.ta = -1; ch.tb = -1; ch.coef = 0.f; ch.ka = ch.kb = 0; ch.pad = 0;
      if (live) ch = P.children[row0 + row];
      // ---- symbols of the tile (teacher forcing: the parent's configuration with the child's flip)
      asm volatile("bar.sync %0, %1;" ::"r"(gbar), "n"(V2_GWARPS * 32) : "memory");      // the previous tile's symbols are no longer read
      {
        const int8_t* c = P.cfg_in + (long long)ch.parent * P.nsites;
        for (int t = half; t < P.Q; t += 2) {
          int sy = 0;
          if (live) { sy = c[stepS[t].z] & 1; if (t == ch.ta || t == ch.tb) sy ^= 1; }
          symS[t * V2_ROWS + row] = (uint8_t)sy;
        }
      }
      asm volatile("bar.sync %0, %1;" ::"r"(gbar), "n"(V2_GWARPS * 32) : "memory");
      // hidden state of an earlier site for this thread's units, as operand words: own ring (sites >= q0 of this tile) or
      // the parent's saved FP32 history (split here)
      auto load_h = [&](int nb, uint4 (&w)[NU4]) {
        if (nb < 0 || !live) {
#pragma unroll
          for (int c = 0; c < NU4; ++c) w[c] = make_uint4(0u, 0u, 0u, 0u);
        } else if (nb >= q0 || P.par_hist == nullptr) {
          const uint4* src = reinterpret_cast<const uint4*>(ring + (size_t)(nb % P.ring_len) * V2_ROWS * NHP);
#pragma unroll
          for (int c = 0; c < NU4; ++c) w[c] = __ldcg(src + c);
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
            split_f16x2(x.x * HS, x.y * HS, w[c].x, w[c].z);
            split_f16x2(x.z * HS, x.w * HS, w[c].y, w[c].w);
          }
        }
      };
      // ---- operand of the tile's first site: both neighbours from the history
      {
        const short4 e0 = stepS[q0];
        uint4 w[NU4];
        load_h(e0.x, w);
        __syncwarp();
        put_operand(0, w);
        load_h(e0.y, w);
        put_operand(1, w);
        tc_wait_st();
        tc_fence_before();
        mbar_arrive(bar_op + g);
      }
      double rlogp = 0.0;                                  // this thread's share of the row's log P (the sites it finished)
      for (int q = q0; q < P.Q; ++q, ++n) {
        const short4 eq = stepS[q];
        const bool has_next = q + 1 < P.Q;
        // the older neighbour of the next site (not this site's output): fetched before the wait for this site's accumulators
        int carry_slot = 0;
        uint4 vold[NU4];
        if (has_next) {
          const short4 en = stepS[q + 1];
          carry_slot = (en.x == q) ? 0 : 1;                 // the next site's neighbour that is this site: horizontal or (row turn) vertical
          load_h(carry_slot == 0 ? en.y : en.x, vold);
        }
        const float cl = (eq.x >= 0) ? (float)symS[eq.x * V2_ROWS + row] : 0.f;
        const float cu = (eq.y >= 0) ? (float)symS[eq.y * V2_ROWS + row] : 0.f;
        const float* bv = biasS + (((eq.x >= 0) ? 1 : 0) + ((eq.y >= 0) ? 2 : 0)) * V2_N + u0;
        const float* duh = biasS + 4 * V2_N + u0;
        const float* duv = biasS + 5 * V2_N + u0;
        uint4* const rdst = reinterpret_cast<uint4*>(ring + (size_t)(q % P.ring_len) * V2_ROWS * NHP);
        const uint32_t cnew = (uint32_t)((carry_slot * NHP + u0) >> 1);
        mbar_wait(bar_acc + g, n & 1u);
        __syncwarp();
        tc_fence_after();
        uint32_t d[NU4][4];
#pragma unroll
        for (int c = 0; c < NU4; ++c) tc_ld4(acc + (uint32_t)(u0 + 4 * c), d[c]);
        tc_wait_ld();
        // all MMAs of this site have completed: the operand may be rewritten chunk by chunk
        float lg0 = 0.f, lg1 = 0.f;
#pragma unroll
        for (int c = 0; c < NU4; ++c) {
          const float4 b4 = *reinterpret_cast<const float4*>(bv + 4 * c);
          const float4 h4 = *reinterpret_cast<const float4*>(duh + 4 * c);
          const float4 v4 = *reinterpret_cast<const float4*>(duv + 4 * c);
          const float bb[4] = {b4.x, b4.y, b4.z, b4.w}, hh[4] = {h4.x, h4.y, h4.z, h4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
          float hv[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float x = fmaf(cu, vv[e], fmaf(cl, hh[e], fmaf(__uint_as_float(d[c][e]), DS, bb[e])));
            // elu, branch-free: exp(min(x, 0)) - 1 by ex2.approx (2 ulp, the class of expf); padding units: zero weights and biases -> 0
            const float em = ex2_approx(fminf(x, 0.f) * 1.4426950408889634f) - 1.0f;
            hv[e] = live ? ((x > 0.f) ? x : em) : 0.f;
            const float2 w = *reinterpret_cast<const float2*>(headW + 2 * (u0 + 4 * c + e));
            lg0 = fmaf(hv[e], w.x, lg0); lg1 = fmaf(hv[e], w.y, lg1);
          }
          uint4 w;
          split_f16x2(hv[0] * HS, hv[1] * HS, w.x, w.z);
          split_f16x2(hv[2] * HS, hv[3] * HS, w.y, w.w);
          // ring image of this site (the same thread reads it back as a vertical neighbour ~2 Nx sites later)
          if (live) __stcg(rdst + c, w);
          if (has_next) {
            tc_st2(a_hi + cnew + 2u * c, make_uint2(w.x, w.y));
            tc_st2(a_lo + cnew + 2u * c, make_uint2(w.z, w.w));
          }
        }
        // ---- operand of the next site
        if (has_next) {
          put_operand(1 - carry_slot, vold);
          tc_wait_st();
          tc_fence_before();
          mbar_arrive(bar_op + g);
        }
        // ---- head: the two unit halves of a row meet in shared memory; the halves take turns in finishing a site
        {
          const uint32_t b = n & 1u;
          float2* const ps = partS + b * (V2_GROUPS * V2_ROWS);
          uint64_t* const pb = bar_pair + (b * V2_GROUPS + g) * 4 + quarter;
          if (half != (int)b) {
            ps[row] = make_float2(lg0, lg1);
            mbar_arrive(pb);
          } else {
            mbar_wait(pb, (n >> 1) & 1u);
            if (live) {
              const float2 o = ps[row];
              float p0 = lg0 + o.x + headW[2 * V2_N], p1 = lg1 + o.y + headW[2 * V2_N + 1];
              const float mx = fmaxf(p0, p1);
              p0 = expf(p0 - mx); p1 = expf(p1 - mx);
              const int sy = symS[q * V2_ROWS + row];
              rlogp += log((double)((sy ? p1 : p0) / (p0 + p1)));
            }
          }
        }
      }
      // ---- tile result: E_loc(parent) += <sigma'|H|sigma> psi(sigma') / psi(sigma)
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
  if (warp == 2) { __syncwarp(); tc_dealloc(tmem_base, 512); }
}

}  // namespace rnnwf
