// Dedicated kernels of the small 1D models (stacked CudnnCompatibleGRUCell chains: 1DTFIM/RNNwavefunction.py:35-118,
// J1J2/ComplexRNNwavefunction.py:45-169; BASELINE configs 1 and 2: N = 20 sites, 1-3 layers x 50 units, 500 samples).
//
// north_star (1): ONE persistent launch per direction, the whole weight set resident in shared memory for the whole sweep.
//   gru1d_sweep_kernel<MODE>  sampling / teacher-forced log psi / connected configurations (local energies, prefix reuse)
//   gru1d_bwd_kernel          reverse sweep: d(pre-activations) of every (site, layer, row) in one launch
// A CTA owns a tile of 64 rows and walks it through all sites and layers.  Weights (33.6 KB per layer at 50 units, plus the
// same again for the dense input of layers >= 1), biases, input tables and heads are copied to shared memory once; the hidden
// state of every layer lives in shared memory ([unit][row], so a thread's 8 rows are two 16-byte loads); nothing but the
// saved state (sampling / evaluation with `saved`) and the results touches global memory inside the sweep.
// Thread (rg, up) owns rows 8 rg .. 8 rg + 7 and units 2 up, 2 up + 1 and computes ALL gate pre-activations of those units
// (r, u, candidate-hidden [, candidate-input]) with FP32 FFMA: 48-64 accumulators, 48 FMA per 5 shared-memory loads, and the
// GRU non-linearity is thread-local (no exchange of pre-activations).  One warp = 8 row groups x 4 unit pairs: the row
// operand is broadcast over the unit pairs and the weight operand over the row groups (conflict-free).
// Heads: two lanes per row (softmax / U(1) mask / inverse-CDF draw or teacher forcing, FP64 running log P and phase).
#pragma once
#include "common.cuh"

namespace rnnwf {

// Rows per thread RG (template parameter of the sweeps) and rows per tile 8 RG.  RG = 8: the connected-configuration sweep and
// large batches (48 FMA per 5 shared-memory loads).  RG = 2: sampling / evaluation / reverse sweeps of a training batch
// (500 samples are 8 tiles of 64 rows but 32 of 16: four times the SMs, a quarter of the per-site chain per tile).
constexpr int G1_ROWS = 64;          // rows per tile at RG = 8
constexpr int G1_RG = 8;             // rows per thread (wide tiles)
constexpr int G1_RG_SMALL = 2;       // rows per thread (small batches)
constexpr int G1_MAXL = 4;
__host__ __device__ constexpr int g1_ldh(int rows) { return rows + 4; }   // row stride of the [unit][row] arrays (16-byte aligned, de-phased banks)
// rows per thread for a batch: the narrow tiles while they still fit one wave of CTAs
// (measured at 500 samples: RG = 2 -> 0.65 ms per C1 step, RG = 4 -> 0.76, RG = 8 -> 0.94; RG = 1 would leave half a warp of
//  head lanes, which the warp shuffles of the heads do not allow)
inline int g1_pick_rg(long long rows, int num_sms) { return (rows + 8 * G1_RG_SMALL - 1) / (8 * G1_RG_SMALL) <= 2LL * num_sms ? G1_RG_SMALL : G1_RG; }
// RG consecutive floats of a [unit][row] array (aligned: RG * 4 bytes)
template <int RG>
__device__ __forceinline__ void g1_load_rows(const float* p, float (&a)[RG]) {
  if constexpr (RG == 8) {
    const float4 a0 = *reinterpret_cast<const float4*>(p), a1 = *reinterpret_cast<const float4*>(p + 4);
    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
  } else if constexpr (RG == 4) {
    const float4 a0 = *reinterpret_cast<const float4*>(p);
    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w;
  } else if constexpr (RG == 2) {
    const float2 a0 = *reinterpret_cast<const float2*>(p);
    a[0] = a0.x; a[1] = a0.y;
  } else {
#pragma unroll
    for (int i = 0; i < RG; ++i) a[i] = p[i];
  }
}

// packed parameter image (floats), built by gru1d_pack_kernel, copied verbatim into shared memory:
//   per layer l: Wh [nh][3][njp] (r, u, candidate-hidden from h_{t-1});  l >= 1: Wx [nh][3][njp] (r, u, candidate-input from
//   the layer below);  l == 0: X0 [2][3][njp] (rows of Wg / Wci selected by the previous symbol);  bias [4][njp]: bg_r, bg_u, bci, bch
//   then headW [nh][4] (zero-padded to 4 outputs), headB [4]
struct G1Layout {
  int njp, off_wh[G1_MAXL], off_wx[G1_MAXL], off_b[G1_MAXL], off_headw, off_headb, pack_floats;
  // shared memory beyond the parameter image
  int off_H, off_sym, off_part, total_bytes;
  __host__ __device__ G1Layout(int nh, int L, int N, int rows = G1_ROWS) {
    njp = (nh + 7) & ~7;
    int o = 0;
    for (int l = 0; l < G1_MAXL; ++l) { off_wh[l] = off_wx[l] = off_b[l] = 0; }
    for (int l = 0; l < L; ++l) {
      off_wh[l] = o; o += nh * 3 * njp;
      off_wx[l] = o; o += (l == 0 ? 2 : nh) * 3 * njp;
      off_b[l] = o;  o += 4 * njp;
    }
    off_headw = o; o += nh * 4;
    off_headb = o; o += 4;
    pack_floats = (o + 3) & ~3;
    int b = pack_floats * 4;
    off_H = b;    b += L * nh * g1_ldh(rows) * 4;
    off_part = b; b += rows * 4 * 4;
    off_sym = b;  b += ((N * rows + 15) & ~15);
    total_bytes = b;
  }
};
inline int g1_threads(int nh) { const int t = 32 * ((((nh + 7) & ~7) / 2 + 3) / 4); return t < 2 * G1_ROWS ? 2 * G1_ROWS : t; }   // >= two head lanes per row
inline bool g1_supported(int nh, int L, int N) { return nh <= 64 && L <= G1_MAXL && N <= 1024 && G1Layout(nh, L, N).total_bytes <= 227 * 1024; }

// TF layout -> the image above.  Wg [din + nh, 2 nh] (rows [x; h], columns [r | u]), Wci [din, nh], Wch [nh, nh], bg [2 nh], bci, bch [nh]
struct G1PackSrc { const float* Wg[G1_MAXL]; const float* bg[G1_MAXL]; const float* Wci[G1_MAXL]; const float* Wch[G1_MAXL];
                   const float* bci[G1_MAXL]; const float* bch[G1_MAXL]; const float* head_w; const float* head_b; const float* phase_w; const float* phase_b; };
template <int UNUSED = 0>
__global__ void gru1d_pack_kernel(const G1PackSrc S, int nh, int L, int N, int D, int ND, float* __restrict__ out) {
  const G1Layout Y(nh, L, N);
  const int njp = Y.njp;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < Y.pack_floats; i += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (i >= Y.off_headb) {
      const int d = i - Y.off_headb;
      if (d < D) v = S.head_b[d]; else if (d < ND) v = S.phase_b[d - D];
    } else if (i >= Y.off_headw) {
      const int j = (i - Y.off_headw) >> 2, d = (i - Y.off_headw) & 3;
      if (d < D) v = S.head_w[j * D + d]; else if (d < ND) v = S.phase_w[j * 2 + (d - D)];
    } else {
      int l = L - 1;
      while (l > 0 && i < Y.off_wh[l]) --l;
      const int din = (l == 0) ? 2 : nh;
      if (i >= Y.off_b[l]) {
        const int which = (i - Y.off_b[l]) / njp, j = (i - Y.off_b[l]) % njp;
        if (j < nh) v = (which == 0) ? S.bg[l][j] : (which == 1) ? S.bg[l][nh + j] : (which == 2) ? S.bci[l][j] : S.bch[l][j];
      } else if (i >= Y.off_wx[l]) {
        const int r = i - Y.off_wx[l], k = r / (3 * njp), g = (r / njp) % 3, j = r % njp;      // k: input row (symbol or unit of the layer below)
        if (j < nh) v = (g == 0) ? S.Wg[l][(size_t)k * 2 * nh + j] : (g == 1) ? S.Wg[l][(size_t)k * 2 * nh + nh + j] : S.Wci[l][(size_t)k * nh + j];
      } else {
        const int r = i - Y.off_wh[l], k = r / (3 * njp), g = (r / njp) % 3, j = r % njp;
        if (j < nh) v = (g == 0) ? S.Wg[l][(size_t)(din + k) * 2 * nh + j] : (g == 1) ? S.Wg[l][(size_t)(din + k) * 2 * nh + nh + j] : S.Wch[l][(size_t)k * nh + j];
      }
    }
    out[i] = v;
  }
}

__device__ __forceinline__ float g1_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

template <int MODE, int RG>
__global__ void __launch_bounds__(256) gru1d_sweep_kernel(const SweepParams P) {
  constexpr int ROWS = 8 * RG, LDH = g1_ldh(ROWS);
  static_assert(2 * ROWS % 32 == 0, "the head lanes (two per row) must be whole warps: they exchange partial sums with warp shuffles");
  extern __shared__ __align__(16) unsigned char g1_smem[];
  const int nh = P.nh, L = P.symL, N = P.Tsym;
  const G1Layout Y(nh, L, N, ROWS);
  const int njp = Y.njp;
  float* const Wsm = reinterpret_cast<float*>(g1_smem);
  float* const Hs = reinterpret_cast<float*>(g1_smem + Y.off_H);            // [L][nh][LDH]
  uint8_t* const symS = g1_smem + Y.off_sym;                               // [N][64]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rg = lane >> 2, up = warp * 4 + (lane & 3);
  const int r0 = rg * RG, j0 = 2 * up;
  const bool jok0 = j0 < nh, jok1 = j0 + 1 < nh;
  const bool act = j0 < njp;           // warps beyond the last unit pair only serve as head lanes

  for (int i = tid; i < Y.pack_floats / 4; i += blockDim.x)
    reinterpret_cast<float4*>(Wsm)[i] = __ldg(reinterpret_cast<const float4*>(P.g1_pack) + i);

  const long long total_rows = (P.rows_dev != nullptr) ? (long long)(*P.rows_dev) : P.rows;
  const long long ntiles = (total_rows + ROWS - 1) / ROWS;
  // head threads: two lanes per row
  const int hrow = tid >> 1, hhalf = tid & 1;
  const bool head_thread = tid < 2 * ROWS;

  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long row0 = tile * ROWS;
    const int nrows = (int)((total_rows - row0 < (long long)ROWS) ? (total_rows - row0) : (long long)ROWS);
    int ts0 = 0;
    if (MODE == MODE_CHILD && P.par_hist != nullptr) {
      const Child c0 = P.children[row0];
      ts0 = (c0.tb >= 0 && c0.tb < c0.ta) ? c0.tb : c0.ta;
    }
    __syncthreads();     // previous tile done with shared memory
    // ---- symbols (teacher forcing) and the starting hidden state
    if (MODE != MODE_SAMPLE) {
      for (int idx = tid; idx < nrows * N; idx += blockDim.x) {
        const int r = idx / N, t = idx - r * N;
        const long long src = (MODE == MODE_CHILD) ? (long long)P.children[row0 + r].parent : row0 + r;
        symS[t * ROWS + r] = (uint8_t)(P.cfg_in[src * N + t] & 1);
      }
    }
    for (int idx = tid; idx < L * nh * LDH; idx += blockDim.x) Hs[idx] = 0.f;
    __syncthreads();
    if (MODE == MODE_CHILD) {
      if (tid < nrows) {
        const Child ch = P.children[row0 + tid];
        symS[ch.ta * ROWS + tid] ^= 1;
        if (ch.tb >= 0) symS[ch.tb * ROWS + tid] ^= 1;
      }
      if (ts0 > 0) {       // prefix reuse: hidden states of every layer after site ts0 - 1, from the parents' saved history
        for (int idx = tid; idx < nrows * L * nh; idx += blockDim.x) {
          const int r = idx / (L * nh), rem = idx - r * (L * nh), l = rem / nh, j = rem - l * nh;
          const int par = P.children[row0 + r].parent;
          Hs[(l * nh + j) * LDH + r] = __ldg(P.par_hist + ((size_t)((ts0 - 1) * L + l) * (size_t)P.par_rows + (size_t)par) * P.nhp + j);
        }
      }
      __syncthreads();
    }
    // per-row running sums live in the row's first head lane
    double rlogp = 0.0, rphase = 0.0;
    int rnup = 0, rparent = 0;
    float rcoef = 0.f;
    if (head_thread && hrow < nrows && MODE == MODE_CHILD) {
      const Child ch = P.children[row0 + hrow];
      rparent = ch.parent; rcoef = ch.coef;
      for (int t = 0; t < ts0; ++t) rnup += symS[t * ROWS + hrow];
    }

    for (int t = ts0; t < N; ++t) {
      for (int l = 0; l < L; ++l) {
        const int q = t * L + l;
        float* const Hl = Hs + (size_t)l * nh * LDH;
        const float* const Wh = Wsm + Y.off_wh[l];
        const float* const Wx = Wsm + Y.off_wx[l];
        const float* const Bl = Wsm + Y.off_b[l];
        // accumulators: [row][unit][r, u, ch, cx]
        float ar[RG][2], au[RG][2], ach[RG][2], acx[RG][2];
#pragma unroll
        for (int i = 0; i < RG; ++i)
#pragma unroll
          for (int e = 0; e < 2; ++e) { ar[i][e] = 0.f; au[i][e] = 0.f; ach[i][e] = 0.f; acx[i][e] = 0.f; }
        // ---- hidden part: [r, u, ch] += h_{t-1} Wh
        if (t > 0 && act) {
#pragma unroll 2
          for (int k = 0; k < nh; ++k) {
            float a[RG];
            g1_load_rows<RG>(Hl + k * LDH + r0, a);
            const float2 wr = *reinterpret_cast<const float2*>(Wh + (k * 3 + 0) * njp + j0);
            const float2 wu = *reinterpret_cast<const float2*>(Wh + (k * 3 + 1) * njp + j0);
            const float2 wc = *reinterpret_cast<const float2*>(Wh + (k * 3 + 2) * njp + j0);
#pragma unroll
            for (int i = 0; i < RG; ++i) {
              ar[i][0] = fmaf(a[i], wr.x, ar[i][0]); ar[i][1] = fmaf(a[i], wr.y, ar[i][1]);
              au[i][0] = fmaf(a[i], wu.x, au[i][0]); au[i][1] = fmaf(a[i], wu.y, au[i][1]);
              ach[i][0] = fmaf(a[i], wc.x, ach[i][0]); ach[i][1] = fmaf(a[i], wc.y, ach[i][1]);
            }
          }
        }
        // ---- input part of the layers above the first: [r, u, cx] += h^{l-1}_t Wx
        if (l > 0 && act) {
          const float* const Hb = Hs + (size_t)(l - 1) * nh * LDH;
#pragma unroll 2
          for (int k = 0; k < nh; ++k) {
            float a[RG];
            g1_load_rows<RG>(Hb + k * LDH + r0, a);
            const float2 wr = *reinterpret_cast<const float2*>(Wx + (k * 3 + 0) * njp + j0);
            const float2 wu = *reinterpret_cast<const float2*>(Wx + (k * 3 + 1) * njp + j0);
            const float2 wc = *reinterpret_cast<const float2*>(Wx + (k * 3 + 2) * njp + j0);
#pragma unroll
            for (int i = 0; i < RG; ++i) {
              ar[i][0] = fmaf(a[i], wr.x, ar[i][0]); ar[i][1] = fmaf(a[i], wr.y, ar[i][1]);
              au[i][0] = fmaf(a[i], wu.x, au[i][0]); au[i][1] = fmaf(a[i], wu.y, au[i][1]);
              acx[i][0] = fmaf(a[i], wc.x, acx[i][0]); acx[i][1] = fmaf(a[i], wc.y, acx[i][1]);
            }
          }
        }
        // own previous state of the thread's units (read before anybody overwrites it)
        float hold[RG][2];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const bool ok = (e == 0) ? jok0 : jok1;
          if (ok) {
            float hv[RG];
            g1_load_rows<RG>(Hl + (j0 + e) * LDH + r0, hv);
#pragma unroll
            for (int i = 0; i < RG; ++i) hold[i][e] = hv[i];
          } else {
#pragma unroll
            for (int i = 0; i < RG; ++i) hold[i][e] = 0.f;
          }
        }
        __syncthreads();      // everybody has read H^l_{t-1}
        // ---- gates, thread-local
        const float2 b_r = *reinterpret_cast<const float2*>(Bl + 0 * njp + j0), b_u = *reinterpret_cast<const float2*>(Bl + 1 * njp + j0);
        const float2 b_ci = *reinterpret_cast<const float2*>(Bl + 2 * njp + j0), b_ch = *reinterpret_cast<const float2*>(Bl + 3 * njp + j0);
        const bool save = (MODE != MODE_CHILD) && (P.sv_hist != nullptr);
        if (act) {
#pragma unroll
        for (int i = 0; i < RG; ++i) {
          const int r = r0 + i;
          float xr[2] = {0.f, 0.f}, xu[2] = {0.f, 0.f}, xc[2] = {acx[i][0], acx[i][1]};
          if (l == 0 && t > 0) {     // one-hot input: a row of the input table (the first input is the zero vector)
            const float* X0 = Wx + (int)(symS[(t - 1) * ROWS + r] & 1) * 3 * njp + j0;
            const float2 v0 = *reinterpret_cast<const float2*>(X0), v1 = *reinterpret_cast<const float2*>(X0 + njp), v2 = *reinterpret_cast<const float2*>(X0 + 2 * njp);
            xr[0] = v0.x; xr[1] = v0.y; xu[0] = v1.x; xu[1] = v1.y; xc[0] = v2.x; xc[1] = v2.y;
          }
          float hn[2];
          float4 gsv[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const float rgate = g1_sigmoid(ar[i][e] + (e ? b_r.y : b_r.x) + xr[e]);
            const float ugate = g1_sigmoid(au[i][e] + (e ? b_u.y : b_u.x) + xu[e]);
            const float hh = ach[i][e] + (e ? b_ch.y : b_ch.x);
            const float c = tanhf(xc[e] + (e ? b_ci.y : b_ci.x) + rgate * hh);
            hn[e] = (1.f - ugate) * c + ugate * hold[i][e];
            gsv[e] = make_float4(rgate, ugate, c, hh);
          }
          if (jok0) Hl[j0 * LDH + r] = hn[0];
          if (jok1) Hl[(j0 + 1) * LDH + r] = hn[1];
          if (save && r < nrows) {       // (padding units of the saved history are written as zeros)
            const size_t base = (size_t)q * (size_t)P.sv_rows + (size_t)(row0 + r);
            *reinterpret_cast<float2*>(P.sv_hist + base * P.nhp + j0) = make_float2(jok0 ? hn[0] : 0.f, jok1 ? hn[1] : 0.f);
            if (P.sv_gates != nullptr) {
              if (jok0) P.sv_gates[base * P.Gp + j0] = gsv[0];
              if (jok1) P.sv_gates[base * P.Gp + j0 + 1] = gsv[1];
            }
          }
        }
        }
        __syncthreads();      // H^l_t complete
      }  // layers

      // ---- heads on the top layer's state: two lanes per row
      if (head_thread) {
        const float* const Ht = Hs + (size_t)(L - 1) * nh * LDH;
        const float* const hw = Wsm + Y.off_headw;
        float lg0 = 0.f, lg1 = 0.f, lg2 = 0.f, lg3 = 0.f;
        for (int j = hhalf; j < nh; j += 2) {
          const float hv = Ht[j * LDH + hrow];
          const float4 w = *reinterpret_cast<const float4*>(hw + j * 4);
          lg0 = fmaf(hv, w.x, lg0); lg1 = fmaf(hv, w.y, lg1); lg2 = fmaf(hv, w.z, lg2); lg3 = fmaf(hv, w.w, lg3);
        }
        lg0 += __shfl_xor_sync(0xffffffffu, lg0, 1); lg1 += __shfl_xor_sync(0xffffffffu, lg1, 1);
        lg2 += __shfl_xor_sync(0xffffffffu, lg2, 1); lg3 += __shfl_xor_sync(0xffffffffu, lg3, 1);
        if (hhalf == 0 && hrow < nrows) {
          const float* hb = Wsm + Y.off_headb;
          float p0 = lg0 + hb[0], p1 = lg1 + hb[1];
          const float mx = fmaxf(p0, p1);
          p0 = expf(p0 - mx); p1 = expf(p1 - mx);
          const float ssum = p0 + p1;
          p0 = p0 / ssum; p1 = p1 / ssum;
          if (P.u1) {
            const int ndn = t - rnup;
            p0 *= (P.cap_dn - ndn >= 0) ? 1.f : 0.f;
            p1 *= (P.cap_up - rnup >= 0) ? 1.f : 0.f;
            const float nrm = p0 + p1;
            p0 /= nrm; p1 /= nrm;
          }
          int sy;
          if (MODE == MODE_SAMPLE) {
            const float u = (P.uniforms != nullptr) ? P.uniforms[(size_t)(row0 + hrow) * N + t]
                                                    : philox_uniform(P.seed, P.row_offset + (unsigned long long)(row0 + hrow), (uint32_t)t);
            const double target = (double)u * ((double)p0 + (double)p1);
            sy = ((double)p0 <= target) ? 1 : 0;
            symS[t * ROWS + hrow] = (uint8_t)sy;
          } else {
            sy = symS[t * ROWS + hrow];
          }
          rlogp += log((double)(sy ? p1 : p0));
          if (P.complex_head) {
            const float xs = (sy ? lg3 : lg2) + (sy ? hb[3] : hb[2]);
            rphase += 3.14159265358979323846 * ((double)xs / (1.0 + fabs((double)xs)));
          }
          rnup += sy;
          if (P.sv_cum != nullptr && MODE != MODE_CHILD)
            P.sv_cum[(size_t)t * (size_t)P.sv_rows + (size_t)(row0 + hrow)] = make_double2(rlogp, rphase);
        }
      }
      if (MODE == MODE_SAMPLE) __syncthreads();     // the drawn symbol selects the next site's input rows
    }  // sites

    // ---- tile results
    if (head_thread && hhalf == 0 && hrow < nrows) {
      if (MODE == MODE_CHILD) {
        double base_lp = P.par_logp[rparent], base_ph = (P.par_phase != nullptr) ? P.par_phase[rparent] : 0.0;
        if (ts0 > 0 && P.par_cum != nullptr) {
          const double2 c = P.par_cum[(size_t)(ts0 - 1) * (size_t)P.par_rows + rparent];
          base_lp -= c.x; base_ph -= c.y;
        }
        const double amp = exp(0.5 * (rlogp - base_lp)) * (double)rcoef;
        if (P.complex_head) {
          double sn, cs;
          sincos(rphase - base_ph, &sn, &cs);
          atomicAdd(P.eloc + 2 * (size_t)rparent, amp * cs);
          atomicAdd(P.eloc + 2 * (size_t)rparent + 1, amp * sn);
        } else {
          atomicAdd(P.eloc + 2 * (size_t)rparent, amp);
        }
      } else {
        if (P.out_logp != nullptr) P.out_logp[row0 + hrow] = rlogp;
        if (P.out_phase != nullptr) P.out_phase[row0 + hrow] = rphase;
        if (MODE == MODE_SAMPLE && P.cfg_out != nullptr)
          for (int t = 0; t < N; ++t) P.cfg_out[(size_t)(row0 + hrow) * N + t] = (int8_t)symS[t * ROWS + hrow];
      }
    }
  }  // tiles
}

}  // namespace rnnwf
