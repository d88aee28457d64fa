// Weight packing: TF-layout parameter tensors -> the sweep engine's streaming layout, and back
// (gradient un-packing).  One thread per destination float4.
//
// Packed GEMM weights of a variant:  [column block][packed slot][k-chunk][class][kk (8)][Gblk] float4
// One-hot input table:               [slot][symbol][Gp] float4
// Bias table:                        [Gp] float4
// A float4 holds the (up to) four gate pre-activations of one hidden unit for the GRU-type cells
// (NT3/NT4 layout), or four consecutive hidden units for the un-gated cells (packed-4 layout).
#pragma once
#include "common.cuh"

namespace rnnwf {

struct PackCtx {
  int cell, nh, nhp, Gblk, ncb, Gp, Greal, cps, D;
  int ncls, nsl;           // classes and packed slots of this variant
  int slot_of_rank[MAX_SLOTS];
  int vkind;               // tensorized: 0 both neighbours, 1 horizontal only, 2 vertical only; GRU1D: layer
  int din;                 // GRU1D: input width of this layer
  // parameter tensors (device pointers into the flat parameter / gradient vector)
  const float* t[8];
  float* gt[8];            // same tensors inside the gradient vector (unpack only)
};

// Tensor slots in PackCtx::t
//  GRU1D:     0 Wg [din+nh,2nh]  1 bg [2nh]  2 Wci [din,nh]  3 Wch [nh,nh]  4 bci [nh]  5 bch [nh]
//  VANILLA2D: 0 Wh [nh,nh]  1 Uh [2,nh]  2 Wv [nh,nh]  3 Uv [2,nh]  4 b [nh]
//  TENS2D:    0 W [nh,2nh,4]  1 b [nh]
//  TENSGRU2D: 0 W  1 b  2 Wg [nh,2nh,4]  3 bg [nh]  4 Wmerge [2nh,nh]
//  GRU4:      0 W [4(nh+6),nh]  1 b  2 Wg  3 bg  4 Wmerge [4nh,nh]

// Linear combination describing one packed scalar: up to two (tensor, index) terms.
struct Src { int n; int ten[2]; long long idx[2]; };

// GEMM weight feeding output group g, component e, from hidden unit k of `slot`, weight class `cls`.
__device__ __forceinline__ Src w_src(const PackCtx& c, int slot, int k, int cls, int g, int e) {
  Src s; s.n = 0;
  const int nh = c.nh;
  if (k >= nh) return s;
  switch (c.cell) {
    case CELL_GRU1D: {
      const int j = g;
      if (j >= nh) return s;
      if (slot == 0) {  // own previous state
        if (e == 0) { s.n = 1; s.ten[0] = 0; s.idx[0] = (long long)(c.din + k) * 2 * nh + j; }
        else if (e == 1) { s.n = 1; s.ten[0] = 0; s.idx[0] = (long long)(c.din + k) * 2 * nh + nh + j; }
        else if (e == 3) { s.n = 1; s.ten[0] = 3; s.idx[0] = (long long)k * nh + j; }
      } else {          // output of the layer below
        if (e == 0) { s.n = 1; s.ten[0] = 0; s.idx[0] = (long long)k * 2 * nh + j; }
        else if (e == 1) { s.n = 1; s.ten[0] = 0; s.idx[0] = (long long)k * 2 * nh + nh + j; }
        else if (e == 2) { s.n = 1; s.ten[0] = 2; s.idx[0] = (long long)k * nh + j; }
      }
      return s;
    }
    case CELL_VANILLA2D: {
      const int j = 4 * g + e;
      if (j >= nh) return s;
      s.n = 1; s.ten[0] = (slot == 0) ? 0 : 2; s.idx[0] = (long long)k * nh + j;
      return s;
    }
    case CELL_TENS2D:
    case CELL_TENSGRU2D: {
      int j, ten;
      if (c.cell == CELL_TENS2D) { j = 4 * g + e; ten = 0; }
      else { j = g; if (e == 3) return s; ten = (e == 0) ? 0 : 2; }
      if (j >= nh) return s;
      const int i = slot * nh + k;  // index into hc = [h_h ; h_v]
      if (c.cell == CELL_TENSGRU2D && e == 2) { s.n = 1; s.ten[0] = 4; s.idx[0] = (long long)i * nh + j; return s; }
      const long long base = ((long long)j * 2 * nh + i) * 4;
      if (c.vkind == 0) { s.n = 2; s.ten[0] = s.ten[1] = ten; s.idx[0] = base + (cls & 1); s.idx[1] = base + 2 + (cls >> 1); }
      else if (c.vkind == 1) { s.n = 1; s.ten[0] = ten; s.idx[0] = base + cls; }
      else { s.n = 1; s.ten[0] = ten; s.idx[0] = base + 2 + cls; }
      return s;
    }
    case CELL_GRU4: {
      const int j = g;
      if (j >= nh || e == 3) return s;
      if (e == 2) { s.n = 1; s.ten[0] = 4; s.idx[0] = (long long)(slot * nh + k) * nh + j; return s; }
      s.n = 1; s.ten[0] = (e == 0) ? 0 : 2; s.idx[0] = (long long)(24 + slot * nh + k) * nh + j;
      return s;
    }
  }
  return s;
}

// One-hot input-table entry: contribution of symbol `sym` arriving through `slot`.
// (Rydberg: a symbol is three bits, i.e. three one-hot pairs -> three weight rows; handled by the caller.)
__device__ __forceinline__ Src x_src(const PackCtx& c, int slot, int row, int g, int e) {
  Src s; s.n = 0;
  const int nh = c.nh;
  switch (c.cell) {
    case CELL_GRU1D: {
      const int j = g;
      if (j >= nh || c.vkind != 0 || slot != 0) return s;
      if (e == 0) { s.n = 1; s.ten[0] = 0; s.idx[0] = (long long)row * 2 * nh + j; }
      else if (e == 1) { s.n = 1; s.ten[0] = 0; s.idx[0] = (long long)row * 2 * nh + nh + j; }
      else if (e == 2) { s.n = 1; s.ten[0] = 2; s.idx[0] = (long long)row * nh + j; }
      return s;
    }
    case CELL_VANILLA2D: {
      const int j = 4 * g + e;
      if (j >= nh) return s;
      s.n = 1; s.ten[0] = (slot == 0) ? 1 : 3; s.idx[0] = (long long)row * nh + j;
      return s;
    }
    case CELL_GRU4: {
      const int j = g;
      if (j >= nh || e >= 2) return s;
      s.n = 1; s.ten[0] = (e == 0) ? 0 : 2; s.idx[0] = (long long)(slot * 6 + row) * nh + j;
      return s;
    }
    default: return s;
  }
}

__device__ __forceinline__ Src b_src(const PackCtx& c, int g, int e) {
  Src s; s.n = 0;
  const int nh = c.nh;
  switch (c.cell) {
    case CELL_GRU1D: {
      if (g >= nh) return s;
      s.n = 1;
      if (e == 0) { s.ten[0] = 1; s.idx[0] = g; }
      else if (e == 1) { s.ten[0] = 1; s.idx[0] = nh + g; }
      else if (e == 2) { s.ten[0] = 4; s.idx[0] = g; }
      else { s.ten[0] = 5; s.idx[0] = g; }
      return s;
    }
    case CELL_VANILLA2D: { const int j = 4 * g + e; if (j >= nh) return s; s.n = 1; s.ten[0] = 4; s.idx[0] = j; return s; }
    case CELL_TENS2D:    { const int j = 4 * g + e; if (j >= nh) return s; s.n = 1; s.ten[0] = 1; s.idx[0] = j; return s; }
    case CELL_TENSGRU2D:
    case CELL_GRU4: {
      if (g >= nh || e >= 2) return s;
      s.n = 1; s.ten[0] = (e == 0) ? 1 : 3; s.idx[0] = g;
      return s;
    }
  }
  return s;
}

__device__ __forceinline__ float src_val(const PackCtx& c, const Src& s) {
  float v = 0.f;
  for (int i = 0; i < s.n; ++i) v += c.t[s.ten[i]][s.idx[i]];
  return v;
}

// number of weight rows one symbol selects per slot (Rydberg: 3 bits -> rows 2b + bit_b)
__device__ __forceinline__ int x_rows(const PackCtx& c, int sym, int rows[3]) {
  if (c.cell == CELL_GRU4) { rows[0] = sym & 1; rows[1] = 2 + ((sym >> 1) & 1); rows[2] = 4 + ((sym >> 2) & 1); return 3; }
  rows[0] = sym; return 1;
}

__global__ void pack_w_kernel(const PackCtx c, float4* __restrict__ dst) {
  const long long total = (long long)c.ncb * c.nsl * c.cps * c.ncls * KC * c.Gblk;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int gl = (int)(r % c.Gblk); r /= c.Gblk;
    const int kk = (int)(r % KC); r /= KC;
    const int cls = (int)(r % c.ncls); r /= c.ncls;
    const int ch = (int)(r % c.cps); r /= c.cps;
    const int sr = (int)(r % c.nsl); r /= c.nsl;
    const int cb = (int)r;
    const int slot = c.slot_of_rank[sr], k = ch * KC + kk, g = cb * c.Gblk + gl;
    float v[4];
    for (int e = 0; e < 4; ++e) v[e] = src_val(c, w_src(c, slot, k, cls, g, e));
    dst[idx] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

__global__ void pack_x_kernel(const PackCtx c, int nslots, float4* __restrict__ dst) {
  const long long total = (long long)nslots * c.D * c.Gp;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(idx % c.Gp);
    const int sym = (int)((idx / c.Gp) % c.D);
    const int slot = (int)(idx / ((long long)c.Gp * c.D));
    int rows[3];
    const int nr = x_rows(c, sym, rows);
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    for (int e = 0; e < 4; ++e)
      for (int i = 0; i < nr; ++i) v[e] += src_val(c, x_src(c, slot, rows[i], g, e));
    dst[idx] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

__global__ void pack_b_kernel(const PackCtx c, float4* __restrict__ dst) {
  for (int g = blockIdx.x * blockDim.x + threadIdx.x; g < c.Gp; g += gridDim.x * blockDim.x) {
    float v[4];
    for (int e = 0; e < 4; ++e) v[e] = src_val(c, b_src(c, g, e));
    dst[g] = make_float4(v[0], v[1], v[2], v[3]);
  }
}

// heads: dense.W [nh,D] (+ dense_phase.W [nh,2]) -> [nhp][ND]; biases -> [ND]
__device__ __forceinline__ void pack_head_body(const float* Wd, const float* bd, const float* Wp, const float* bp, int nh, int nhp,
                                               int D, int ND, float* head_w, float* head_b, int vb, int vg) {
  for (int idx = vb * blockDim.x + threadIdx.x; idx < nhp * ND; idx += vg * blockDim.x) {
    const int j = idx / ND, d = idx - j * ND;
    float v = 0.f;
    if (j < nh) {
      if (d < D) v = Wd[j * D + d];
      else if (Wp != nullptr && d < D + 2) v = Wp[j * 2 + (d - D)];
    }
    head_w[idx] = v;
    if (j == 0) head_b[d] = (d < D) ? bd[d] : ((bp != nullptr && d < D + 2) ? bp[d - D] : 0.f);
  }
}
__global__ void pack_head_kernel(const float* Wd, const float* bd, const float* Wp, const float* bp, int nh, int nhp,
                                 int D, int ND, float* head_w, float* head_b) {
  pack_head_body(Wd, bd, Wp, bp, nh, nhp, D, ND, head_w, head_b, blockIdx.x, gridDim.x);
}

}  // namespace rnnwf
