// Reverse sweep of the small 1D models (stacked CudnnCompatibleGRUCell chains) in ONE launch: the autodiff of
// 1DTFIM/TrainingRNN_1DTFIM.py:157-161 / J1J2/TrainingRNN_J1J2.py:197-200 through the N unrolled steps.
//
// One launch per layer, top layer first (a single launch for the one-layer models of the reference's entry scripts).  A CTA
// owns 64 rows and walks the sites N-1 .. 0 of its layer; the layer's transposed weights stay in shared memory for the whole
// sweep and the gradient with respect to the layer's own state never leaves shared memory; what a layer hands to the one
// below (d h^{l-1}_t of every site) goes through a small global buffer.  Per site:
//   A  thread (8 rows x 2 units) forms d(pre-activations) of its units from the saved gates (r, u, c, hh) and
//        d h = [from site t+1] + [from the layer above] + [head push, top layer], writes them to shared memory and, layer-major,
//        to global memory (operands of the weight-gradient GEMMs that follow the sweep: backward.cuh, one launch per tensor);
//   B  the same thread computes d h^l_{t-1} = u * d h + [dr, du] Wg_h^T + dch Wch^T  and, above the first layer,
//        d h^{l-1}_t = [dr, du] Wg_x^T + dci Wci^T   (FP32 FFMA, 96 FMA per 14 shared-memory loads).
// The heads' d(logits) (softmax recomputed from the saved top-layer state, U(1) mask, phase) are formed by two lanes per row.
#pragma once
#include "gru1d.cuh"

namespace rnnwf {

// transposed parameter image: per layer  WhT [nh (j)][3][njp (k)] = (Wg_h[k][r_j], Wg_h[k][u_j], Wch[k][j]);
// l >= 1: WxT [nh (j)][3][njp (k)] = (Wg_x[k][r_j], Wg_x[k][u_j], Wci[k][j]);  then headW [nh][4], headB [4]
struct G1BwdLayout {
  int njp, off_wh[G1_MAXL], off_wx[G1_MAXL], off_headw, off_headb, pack_floats;
  int off_dH, off_dX, off_DP, off_DL, total_bytes;
  __host__ __device__ G1BwdLayout(int nh, int L, int rows = G1_ROWS) {
    njp = (nh + 7) & ~7;
    int o = 0;
    for (int l = 0; l < G1_MAXL; ++l) { off_wh[l] = off_wx[l] = 0; }
    for (int l = 0; l < L; ++l) {
      off_wh[l] = o; o += nh * 3 * njp;
      off_wx[l] = o; o += (l == 0 ? 0 : nh * 3 * njp);
    }
    off_headw = o; o += nh * 4;
    off_headb = o; o += 4;
    pack_floats = (o + 3) & ~3;
    // shared memory of one launch (one layer): [WhT | WxT | headW | headB] then the work arrays
    int b = (2 * nh * 3 * njp + nh * 4 + 4) * 4;
    off_dH = b; b += njp * g1_ldh(rows) * 4;     // d h^l from site t+1
    off_dX = b;                                  // (unused: the hand-over between layers is in global memory)
    off_DP = b; b += 4 * nh * g1_ldh(rows) * 4;  // dr, du, dci, dch of the current site: [4][unit][row]
    off_DL = b; b += rows * 16;                  // d(logits) of the current site
    total_bytes = b;
  }
};
inline bool g1_bwd_supported(int nh, int L) { return L <= G1_MAXL && G1BwdLayout(nh, L).total_bytes <= 220 * 1024; }

template <int UNUSED = 0>
__global__ void gru1d_pack_bwd_kernel(const G1PackSrc S, int nh, int L, int D, int ND, float* __restrict__ out) {
  const G1BwdLayout Y(nh, L);
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
      const bool xpart = (l > 0) && i >= Y.off_wx[l];
      const int r = i - (xpart ? Y.off_wx[l] : Y.off_wh[l]);
      const int j = r / (3 * njp), g = (r / njp) % 3, k = r % njp;
      if (k < nh) {
        const int krow = xpart ? k : din + k;                 // row of Wg
        if (g == 0) v = S.Wg[l][(size_t)krow * 2 * nh + j];
        else if (g == 1) v = S.Wg[l][(size_t)krow * 2 * nh + nh + j];
        else v = xpart ? S.Wci[l][(size_t)k * nh + j] : S.Wch[l][(size_t)k * nh + j];
      }
    }
    out[i] = v;
  }
}

struct G1BwdParams {
  int nh, nhp, Gp, L, N, D, ND, complex_head, u1, cap_up, cap_dn;
  long long R;
  const int8_t* cfg;           // [R][N]
  const float* hist;           // [Q][R][nhp], q = t L + l
  const float4* gates;         // [Q][R][Gp]: (r, u, c, hh)
  const double* seed_logp;
  const double* seed_phase;    // or null
  const float* packT;
  // outputs, layer-major: index (l N + t) R + row
  float* dpG;                  // [L N R][2 nh]  (dr | du)
  float* dpCI;                 // [L N R][nh]
  float* dpCH;                 // [L N R][nh]
  float* dlog;                 // [N R][ND]
  float* X;                    // [N R][2]   one-hot of the previous symbol (zeros at t = 0)
  int layer;                   // the layer this launch walks
  const float* dx_in;          // [N][R][nhp]: d h^l_t handed down by layer l+1 (null for the top layer)
  float* dx_out;               // [N][R][nhp]: d h^{l-1}_t for the layer below (null for the first layer)
};

template <int RG>
__global__ void __launch_bounds__(256) gru1d_bwd_kernel(const G1BwdParams P) {
  constexpr int ROWS = 8 * RG, LDH = g1_ldh(ROWS);
  static_assert(2 * ROWS % 32 == 0, "the head lanes (two per row) must be whole warps");
  extern __shared__ __align__(16) unsigned char g1b_smem[];
  const int nh = P.nh, L = P.L, N = P.N, l = P.layer;
  const G1BwdLayout Y(nh, L, ROWS);
  const int njp = Y.njp;
  float* const Wsm = reinterpret_cast<float*>(g1b_smem);
  float* const dH = reinterpret_cast<float*>(g1b_smem + Y.off_dH);     // [njp][LDH]
  float* const DP = reinterpret_cast<float*>(g1b_smem + Y.off_DP);     // [4][nh][LDH]
  float4* const DL = reinterpret_cast<float4*>(g1b_smem + Y.off_DL);   // [64]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rg = lane >> 2, up = warp * 4 + (lane & 3);
  const int r0 = rg * RG, j0 = 2 * up;
  const bool jok0 = j0 < nh, jok1 = j0 + 1 < nh, act = j0 < njp;
  const int hrow = tid >> 1, hhalf = tid & 1;
  const bool head_thread = tid < 2 * ROWS;
  const long long R = P.R;

  const int wsz = nh * 3 * njp;
  float* const WhT = Wsm;
  float* const WxT = Wsm + wsz;
  float* const hw = Wsm + 2 * wsz;
  float* const hb = hw + nh * 4;
  for (int i = tid; i < wsz / 4; i += blockDim.x) {
    reinterpret_cast<float4*>(WhT)[i] = __ldg(reinterpret_cast<const float4*>(P.packT + Y.off_wh[l]) + i);
    if (l > 0) reinterpret_cast<float4*>(WxT)[i] = __ldg(reinterpret_cast<const float4*>(P.packT + Y.off_wx[l]) + i);
  }
  for (int i = tid; i < nh * 4 + 4; i += blockDim.x) hw[i] = __ldg(P.packT + Y.off_headw + i);
  float* const DPr = DP, *const DPu = DP + nh * LDH, *const DPci = DP + 2 * nh * LDH, *const DPch = DP + 3 * nh * LDH;

  const long long ntiles = (R + ROWS - 1) / ROWS;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long row0 = tile * ROWS;
    const int nrows = (int)((R - row0 < (long long)ROWS) ? (R - row0) : (long long)ROWS);
    __syncthreads();
    for (int i = tid; i < njp * LDH; i += blockDim.x) dH[i] = 0.f;
    // ups before the current site, per row (head lane 0 of the row)
    int nup = 0;
    if (head_thread && hhalf == 0 && hrow < nrows)
      for (int t = 0; t < N; ++t) nup += P.cfg[(row0 + hrow) * N + t] & 1;
    __syncthreads();

    for (int t = N - 1; t >= 0; --t) {
      // ---- heads: d(logits) of site t (top layer's launch)
      if (head_thread && l == L - 1) {
        float lg0 = 0.f, lg1 = 0.f, lg2 = 0.f, lg3 = 0.f;
        const bool live = hrow < nrows;
        if (live) {
          const float* h = P.hist + ((size_t)(t * L + L - 1) * (size_t)R + (size_t)(row0 + hrow)) * P.nhp;
          for (int j = hhalf; j < nh; j += 2) {
            const float hv = __ldg(h + j);
            const float4 w = *reinterpret_cast<const float4*>(hw + j * 4);
            lg0 = fmaf(hv, w.x, lg0); lg1 = fmaf(hv, w.y, lg1); lg2 = fmaf(hv, w.z, lg2); lg3 = fmaf(hv, w.w, lg3);
          }
        }
        lg0 += __shfl_xor_sync(0xffffffffu, lg0, 1); lg1 += __shfl_xor_sync(0xffffffffu, lg1, 1);
        lg2 += __shfl_xor_sync(0xffffffffu, lg2, 1); lg3 += __shfl_xor_sync(0xffffffffu, lg3, 1);
        if (hhalf == 0) {
          float4 dl = make_float4(0.f, 0.f, 0.f, 0.f);
          if (live) {
            const int s = P.cfg[(row0 + hrow) * N + t] & 1;
            nup -= s;                                   // ups among sites < t
            float p0 = lg0 + hb[0], p1 = lg1 + hb[1];
            const float mx = fmaxf(p0, p1);
            p0 = expf(p0 - mx); p1 = expf(p1 - mx);
            const float ssum = p0 + p1;
            p0 /= ssum; p1 /= ssum;
            if (P.u1) {
              const int ndn = t - nup;
              p0 *= (P.cap_dn - ndn >= 0) ? 1.f : 0.f;
              p1 *= (P.cap_up - nup >= 0) ? 1.f : 0.f;
              const float nrm = p0 + p1;
              p0 /= nrm; p1 /= nrm;
            }
            const float sl = (float)P.seed_logp[row0 + hrow];
            dl.x = sl * ((s == 0 ? 1.f : 0.f) - p0);
            dl.y = sl * ((s == 1 ? 1.f : 0.f) - p1);
            if (P.complex_head) {
              const float sp = (P.seed_phase != nullptr) ? (float)P.seed_phase[row0 + hrow] : 0.f;
              const float x = (s == 0) ? lg2 + hb[2] : lg3 + hb[3];
              const float den = 1.f + fabsf(x);
              const float g = sp * 3.14159265358979f / (den * den);
              dl.z = (s == 0) ? g : 0.f;
              dl.w = (s == 1) ? g : 0.f;
            }
            float* o = P.dlog + ((size_t)t * (size_t)R + (size_t)(row0 + hrow)) * P.ND;
            o[0] = dl.x; o[1] = dl.y;
            if (P.ND > 2) { o[2] = dl.z; o[3] = dl.w; }
            const int sprev = (t > 0) ? (P.cfg[(row0 + hrow) * N + t - 1] & 1) : -1;
            *reinterpret_cast<float2*>(P.X + ((size_t)t * (size_t)R + (size_t)(row0 + hrow)) * 2) =
                make_float2(sprev == 0 ? 1.f : 0.f, sprev == 1 ? 1.f : 0.f);
          }
          DL[hrow] = dl;
        }
      }
      __syncthreads();

      {
        const int q = t * L + l;
        float* const dHl = dH;
        float keep[RG][2];       // u * d h: the direct path into d h^l_{t-1}
        // ---- A: d(pre-activations) of this thread's units
        if (act) {
#pragma unroll
          for (int i = 0; i < RG; ++i) {
            const int r = r0 + i;
            const bool live = r < nrows;
            const size_t base = (size_t)q * (size_t)R + (size_t)(row0 + r);
            const size_t obase = ((size_t)(l * N + t) * (size_t)R + (size_t)(row0 + r));
            float dr[2], du[2], dci[2], dch[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int j = j0 + e;
              const bool ok = (e == 0 ? jok0 : jok1) && live;
              float dhp = dHl[j * LDH + r];
              if (l < L - 1) { if (live) dhp += __ldg(P.dx_in + ((size_t)t * (size_t)R + (size_t)(row0 + r)) * P.nhp + j); }
              else if (j < nh) {
                const float4 dl = DL[r];
                const float4 w = *reinterpret_cast<const float4*>(hw + j * 4);
                dhp += w.x * dl.x + w.y * dl.y + w.z * dl.z + w.w * dl.w;
              }
              float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
              float hown = 0.f;
              if (ok) {
                g = __ldg(P.gates + base * P.Gp + j);
                if (t > 0) hown = __ldg(P.hist + ((size_t)(q - L) * (size_t)R + (size_t)(row0 + r)) * P.nhp + j);
              } else dhp = 0.f;
              const float dug = dhp * (hown - g.z), dc = dhp * (1.f - g.y);
              const float dcp = dc * (1.f - g.z * g.z);
              dr[e] = dcp * g.w * g.x * (1.f - g.x);
              du[e] = dug * g.y * (1.f - g.y);
              dci[e] = dcp;
              dch[e] = dcp * g.x;
              keep[i][e] = dhp * g.y;
              if (e == 0 ? jok0 : jok1) {
                DPr[j * LDH + r] = dr[e]; DPu[j * LDH + r] = du[e]; DPci[j * LDH + r] = dci[e]; DPch[j * LDH + r] = dch[e];
              }
            }
            if (live) {
              if (jok1 && !(nh & 1)) {       // (8-byte stores need an even row length)
                *reinterpret_cast<float2*>(P.dpG + obase * 2 * nh + j0) = make_float2(dr[0], dr[1]);
                *reinterpret_cast<float2*>(P.dpG + obase * 2 * nh + nh + j0) = make_float2(du[0], du[1]);
                *reinterpret_cast<float2*>(P.dpCI + obase * nh + j0) = make_float2(dci[0], dci[1]);
                *reinterpret_cast<float2*>(P.dpCH + obase * nh + j0) = make_float2(dch[0], dch[1]);
              } else {
#pragma unroll
                for (int e = 0; e < 2; ++e)
                  if (e == 0 ? jok0 : jok1) {
                    P.dpG[obase * 2 * nh + j0 + e] = dr[e]; P.dpG[obase * 2 * nh + nh + j0 + e] = du[e];
                    P.dpCI[obase * nh + j0 + e] = dci[e]; P.dpCH[obase * nh + j0 + e] = dch[e];
                  }
              }
            }
          }
        }
        __syncthreads();
        // ---- B: gradients with respect to the inputs of this cell
        if (act && (t > 0 || l > 0)) {
          float at[RG][2], ab[RG][2];      // d h^l_{t-1} and d h^{l-1}_t of units (k) j0, j0 + 1
#pragma unroll
          for (int i = 0; i < RG; ++i) { at[i][0] = keep[i][0]; at[i][1] = keep[i][1]; ab[i][0] = 0.f; ab[i][1] = 0.f; }
          const bool below = l > 0;
#pragma unroll 2
          for (int j = 0; j < nh; ++j) {
            float a[4][RG];
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              if (g == 2 && !below) continue;
              g1_load_rows<RG>(DP + (g * nh + j) * LDH + r0, a[g]);
            }
            const float2 wr = *reinterpret_cast<const float2*>(WhT + (j * 3 + 0) * njp + j0);
            const float2 wu = *reinterpret_cast<const float2*>(WhT + (j * 3 + 1) * njp + j0);
            const float2 wc = *reinterpret_cast<const float2*>(WhT + (j * 3 + 2) * njp + j0);
#pragma unroll
            for (int i = 0; i < RG; ++i) {
              at[i][0] = fmaf(a[0][i], wr.x, fmaf(a[1][i], wu.x, fmaf(a[3][i], wc.x, at[i][0])));
              at[i][1] = fmaf(a[0][i], wr.y, fmaf(a[1][i], wu.y, fmaf(a[3][i], wc.y, at[i][1])));
            }
            if (below) {
              const float2 xr = *reinterpret_cast<const float2*>(WxT + (j * 3 + 0) * njp + j0);
              const float2 xu = *reinterpret_cast<const float2*>(WxT + (j * 3 + 1) * njp + j0);
              const float2 xc = *reinterpret_cast<const float2*>(WxT + (j * 3 + 2) * njp + j0);
#pragma unroll
              for (int i = 0; i < RG; ++i) {
                ab[i][0] = fmaf(a[0][i], xr.x, fmaf(a[1][i], xu.x, fmaf(a[2][i], xc.x, ab[i][0])));
                ab[i][1] = fmaf(a[0][i], xr.y, fmaf(a[1][i], xu.y, fmaf(a[2][i], xc.y, ab[i][1])));
              }
            }
          }
#pragma unroll
          for (int i = 0; i < RG; ++i) {
            const int r = r0 + i;
            dHl[j0 * LDH + r] = at[i][0]; dHl[(j0 + 1) * LDH + r] = at[i][1];
            if (below && r < nrows)
              *reinterpret_cast<float2*>(P.dx_out + ((size_t)t * (size_t)R + (size_t)(row0 + r)) * P.nhp + j0) = make_float2(ab[i][0], ab[i][1]);
          }
        }
        __syncthreads();
      }
    }  // sites
  }  // tiles
}

}  // namespace rnnwf
