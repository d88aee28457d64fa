// Weight images of the tensor-core sweep (sweep_tc.cuh): TF layout -> UMMA K-major core matrices, FP16 (hi [, lo]).
#pragma once
#include "sweep_tc.cuh"

namespace rnnwf {

// ---- weight image for the tensor-core sweep --------------------------------------------------------------
// [variant 3][pass NP][n-core 14][k-core kc][8 rows][8 halves]:  B[n][k] of the pass, K-major core matrices
// (n ordered thread-major, see below; n-core outermost so that the columns a narrow last pass needs are one contiguous run).
// variants: 0 both neighbours, 1 horizontal only, 2 vertical only.
// (`vb`, `vg`: the block's index and the number of blocks inside the part of the grid that works on this image)
__device__ __forceinline__ void pack_tc_body(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                                             const float* __restrict__ b, const float* __restrict__ bg,
                                             int nh, int nhp, int NP, int split, __half* __restrict__ out, long long vb, long long vg) {
  // split != 0 (split-precision regime): [variant][pass][part 2: hi, lo][image], unscaled gate weights
  const int kc = 2 * (nhp / 8);
  const int nparts = split ? 2 : 1;
  const long long per_pass = (long long)kc * (TC_N / 8) * 64;
  const long long total = 3LL * NP * nparts * per_pass;
  const float gate_scale = split ? 1.0f : TC_SCALE_GATE;
  for (long long idx = vb * (long long)blockDim.x + threadIdx.x; idx < total; idx += vg * blockDim.x) {
    long long r = idx;
    const int e = (int)(r % 8); r /= 8;
    const int r8 = (int)(r % 8); r /= 8;
    const int ki = (int)(r % kc); r /= kc;
    const int ni = (int)(r % (TC_N / 8)); r /= (TC_N / 8);
    const int part = (int)(r % nparts); r /= nparts;
    const int p = (int)(r % NP); r /= NP;
    const int variant = (int)r;
    const int n = ni * 8 + r8, k = ki * 8 + e;
    // MMA column n = sub * 28 + blk * 4 + e  <->  block blk of unit j = 16 p + 4 sub + e: the 28 accumulators an
    // epilogue thread needs are contiguous TMEM columns
    const int sub = n / TC_TPS, blk = (n % TC_TPS) / 4, j = p * TC_JP + sub * 4 + (n % 4);
    const int slot = k / nhp, unit = tc_kunit(nhp, k - slot * nhp);   // K axis in thread-major order
    float v = 0.f;
    if (nh < nhp && slot == 0 && unit == nh && j < nh) {
      // bias row: the operand carries a constant 1 at position num_units of the horizontal half (all variants)
      v = (blk == 0) ? b[j] : ((blk == 3 && bg != nullptr) ? bg[j] : 0.f);
    } else if (unit < nh && j < nh) {
      const int i = slot * nh + unit;
      if (blk == 6) {
        v = (Wmerge != nullptr) ? Wmerge[(long long)i * nh + j] : 0.f;
      } else if (blk >= 3 && Wg == nullptr) {
        v = 0.f;                                   // un-gated cell: no update gate
      } else {
        const float* T = (blk < 3) ? W : Wg;
        const float* w = T + ((long long)j * 2 * nh + i) * 4;
        const int kind = blk % 3;
        if (variant == 0) v = (kind == 0) ? (w[0] + w[2]) : (kind == 1 ? (w[1] - w[0]) : (w[3] - w[2]));
        else if (variant == 1) v = (kind == 0) ? w[0] : (kind == 1 ? (w[1] - w[0]) : 0.f);
        else v = (kind == 0) ? w[2] : (kind == 1 ? 0.f : (w[3] - w[2]));
      }
    }
    v *= (blk < 3 ? TC_SCALE_CAND : (blk < 6 ? gate_scale : 1.0f)) * (split ? TC_SPLIT_WS : 1.0f);
    const __half hi = __float2half_rn(v);
    out[idx] = (part == 0) ? hi : __float2half_rn(v - __half2float(hi));
  }
}

__global__ void pack_tc_kernel(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                               const float* __restrict__ b, const float* __restrict__ bg,
                               int nh, int nhp, int NP, int split, __half* __restrict__ out) {
  pack_tc_body(W, Wg, Wmerge, b, bg, nh, nhp, NP, split, out, blockIdx.x, gridDim.x);
}

__device__ __forceinline__ void pack_tc_bias_body(const float* __restrict__ b, const float* __restrict__ bg, int nh, int nhp, float* __restrict__ out,
                                                  int vb, int vg) {
  for (int i = vb * blockDim.x + threadIdx.x; i < 2 * nhp; i += vg * blockDim.x) {
    const int which = i / nhp, j = i - which * nhp;
    out[i] = (j < nh) ? (which == 0 ? b[j] : (bg != nullptr ? bg[j] : 0.f)) : 0.f;
  }
}
__global__ void pack_tc_bias_kernel(const float* __restrict__ b, const float* __restrict__ bg, int nh, int nhp, float* __restrict__ out) {
  pack_tc_bias_body(b, bg, nh, nhp, out, blockIdx.x, gridDim.x);
}

// ---- vanilla 2D RNN (CELLK = 1): one image [n-core 14][k-core kc][8 rows][8 halves], B[n][k] = W_half[unit(k)][unit(n)]
// with n = sub * 28 + blk * 4 + e  <->  output unit 16 blk + 4 sub + e, k in the thread-major order of the operand.
__global__ void pack_tc_vanilla_kernel(const float* __restrict__ Wh, const float* __restrict__ Wv, int nh, int nhp, int split, __half* __restrict__ out) {
  const int kc = 2 * (nhp / 8);
  const long long total = (long long)kc * (TC_N / 8) * 64 * (split ? 2 : 1);      // split: [part 2: hi, lo][image]
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int e = (int)(r % 8); r /= 8;
    const int r8 = (int)(r % 8); r /= 8;
    const int ki = (int)(r % kc); r /= kc;
    const int ni = (int)(r % (TC_N / 8)); r /= (TC_N / 8);
    const int part = (int)r;
    const int n = ni * 8 + r8, k = ki * 8 + e;
    const int sub = n / TC_TPS, blk = (n % TC_TPS) / 4, j = blk * TC_JP + sub * 4 + (n % 4);
    const int slot = k / nhp, unit = tc_kunit(nhp, k - slot * nhp);
    float v = 0.f;
    if (unit < nh && j < nh) v = (slot == 0 ? Wh : Wv)[(long long)unit * nh + j] * (split ? TC_SPLIT_WS : 1.0f);
    const __half hi = __float2half_rn(v);
    out[idx] = (part == 0) ? hi : __float2half_rn(v - __half2float(hi));
  }
}
// bias / input tables [6][nhp]: b' of the four neighbour variants (bit 0: horizontal present, bit 1: vertical present),
// then dUh = Uh[1] - Uh[0], dUv = Uv[1] - Uv[0]   (one-hot inputs: x U = U[0] + sigma (U[1] - U[0]); absent neighbour: 0)
__global__ void pack_tc_vanilla_bias_kernel(const float* __restrict__ Uh, const float* __restrict__ Uv, const float* __restrict__ b,
                                            int nh, int nhp, float* __restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 6 * nhp; i += gridDim.x * blockDim.x) {
    const int which = i / nhp, j = i - which * nhp;
    float v = 0.f;
    if (j < nh) {
      if (which < 4) v = b[j] + ((which & 1) ? Uh[j] : 0.f) + ((which & 2) ? Uv[j] : 0.f);
      else if (which == 4) v = Uh[nh + j] - Uh[j];
      else v = Uv[nh + j] - Uv[j];
    }
    out[i] = v;
  }
}

}  // namespace rnnwf
