// Backward of log P for the four-neighbour Rydberg GRU in the tensor-core regime
// (gradient of RydbergKagome_2DGRU/MDGRUcell.py:48-59 + the softmax head of MDRNNWavefunction.py:157-238, i.e. what
// TF autodiff does for the cost of Run_RNNWF_Rydberg.py:158-159).
//
// bwd4_kernel -- persistent reverse sweep, one CTA per 128-row tile, unit cells walked backwards.  Per cell q the 512
// producer threads (4 per row; thread (row, sub) owns hidden units [64 sub, 64 sub + 64)) recompute the softmax from the
// saved hidden state, form d(logits), add the head's push to d h_q, turn it into (ds, dg, dm) with the saved gates
// (u, h~, m) and write them as BF16 UMMA K-major chunk images [chunk (blk, p)][k-core 8][row 128][16 B]; one
// tcgen05.mma pass per live neighbour slot, d h_nbr += [ds dg dm] * [W_s ; Wg_s ; Wmerge_s]^T  ([128 x 768] x [768 x 256],
// K streamed in 12 jobs of 64 from the image just written (A) and the transposed weight image (B)), and the same threads
// add their 64 accumulator columns to d h of that neighbour (a thread only ever touches its own (row, 64 units): no
// atomics, no barriers).  BF16 rather than FP16: gradients need the FP32 exponent range.
// The images it leaves behind -- d(pre-activations), hidden states and one-hot inputs per (cell, tile), all with the 128
// rows contiguous per 8-unit k-core -- are exactly MN-major UMMA operands of the weight-gradient contraction over rows,
// which dw4_kernel runs as a grouped GEMM:  dW_slot[unit k][(blk, j)] = sum_{cells, tiles} himg[nbr_slot]^T dimg.
#pragma once
#include <cuda_bf16.h>

#include "sweep_tc4.cuh"

namespace rnnwf {

constexpr int B4_THREADS = 32 * 18;       // warp 0: MMA issue, 1..16: producers / epilogue, 17: loader
constexpr int B4_STAGES = 3;
constexpr uint32_t B4_A_BYTES = 16384;    // [k-core 8][row 128][8 x bf16]
constexpr uint32_t B4_B_MAX = 32768;      // [k-core 8][n 64 NP][8 x bf16]
constexpr uint32_t B4_STAGE = B4_A_BYTES + B4_B_MAX;

struct Bwd4Params {
  int Q, nh, nhp, NP, Gp, nsites;
  long long R;
  const StepMeta* steps;
  const int8_t* cfg;
  const float* hist;        // [Q][R][nhp]
  const float4* gates;      // tile-blocked [Q][tiles][Gp][row 128]  (u, h~, m, -)
  float* dh;                // [Q][R][nhp], zero on entry
  float* dp;                // optional FP32 copy of (ds, dg, dm): [3][Q][R][nh]
  float* dlog;              // [Q][R][8]
  const double* seed;       // [R]
  const float* head_w;      // [nhp][8]
  const float* head_b;      // [8]
  const unsigned char* Wt;  // BF16 [slot 4][chunk 3 NP][k-core 8][n 64 NP][8]
  unsigned char* dimg;      // BF16 [Q][tiles][chunk 3 NP][k-core 8][row 128][8]
  unsigned char* himg;      // BF16 [Q][tiles][k-core 8 NP][row 128][8]
  unsigned char* ximg;      // BF16 [Q][tiles][k-core 4][row 128][8]
  unsigned char* dlimg;     // BF16 [Q][tiles][k-core 2][row 128][8]: d(logits) (second k-core zero: the MMA needs N >= 16)
};

struct B4Smem {
  int off_stage, off_headw, off_headb, off_part, off_step, off_bar, off_tmem, total;
  __host__ __device__ B4Smem(int NP, int Q) {
    int o = 0;
    off_stage = o; o += B4_STAGES * (int)B4_STAGE;
    off_headw = o; o += NP * 64 * 32;
    off_headb = o; o += 32;
    off_part = o;  o += 2 * 128 * 4 * 32;
    off_step = o;  o += Q * 16;
    off_bar = o;   o += 128;
    off_tmem = o;  o += 16;
    total = o + 128;
  }
};

__host__ __device__ constexpr uint32_t b4_idesc_bf16(int M, int N, int a_mn, int b_mn) {
  // D = F32 (bit 4), A = B = BF16 (format 1 at [7,10) and [10,13)), major-ness at bits 15 / 16, N >> 3 at [17,23), M >> 4 at [24,29)
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn << 15) | ((uint32_t)b_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

__global__ void __launch_bounds__(B4_THREADS, 1) bwd4_kernel(const Bwd4Params P) {
  extern __shared__ __align__(128) unsigned char b4_smem[];
  const int NP = P.NP, N4 = 64 * NP, NJ = 3 * NP;
  const B4Smem L(NP, P.Q);
  unsigned char* const stage0 = b4_smem + L.off_stage;
  float* const headW = reinterpret_cast<float*>(b4_smem + L.off_headw);
  float* const headB = reinterpret_cast<float*>(b4_smem + L.off_headb);
  float4* const partS = reinterpret_cast<float4*>(b4_smem + L.off_part);     // [2][row][sub][2]
  int4* const stepT = reinterpret_cast<int4*>(b4_smem + L.off_step);
  uint64_t* const bars = reinterpret_cast<uint64_t*>(b4_smem + L.off_bar);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(b4_smem + L.off_tmem);
  uint64_t* const bar_full = bars + 0;        // [3]
  uint64_t* const bar_empty = bars + 3;       // [3]
  uint64_t* const bar_acc_full = bars + 6;    // [2]
  uint64_t* const bar_acc_free = bars + 8;    // [2] (512 arrivals)
  uint64_t* const bar_ready = bars + 10;      // the cell's d(pre-activation) image is written (512 arrivals)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nh = P.nh, nhp = P.nhp;
  const uint32_t b_bytes = (uint32_t)N4 * 128u;                 // one weight chunk: 8 k-cores x N4 x 16 B
  const long long ntiles = (P.R + 127) / 128;
  const size_t dimg_tile = (size_t)NJ * B4_A_BYTES, himg_tile = (size_t)NP * B4_A_BYTES, ximg_tile = 4 * 128 * 16;

  if (tid == 0) {
    for (int i = 0; i < B4_STAGES; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_acc_full + i, 1); mbar_init(bar_acc_free + i, 512); }
    mbar_init(bar_ready, 512);
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, 512);
  for (int i = tid; i < N4 * 8; i += B4_THREADS) headW[i] = ((i >> 3) < nh) ? P.head_w[i] : 0.f;
  if (tid < 8) headB[tid] = P.head_b[tid];
  for (int i = tid; i < P.Q; i += B4_THREADS) {
    const StepMeta* m = P.steps + i;
    int order = 0, nact = 0;
    for (int s = 0; s < 4; ++s) if (m->nbr[s] >= 0) { order |= s << (4 * nact); ++nact; }
    stepT[i] = make_int4(order | (nact << 16), (int)(uint16_t)m->nbr[0] | ((int)(uint16_t)m->nbr[1] << 16),
                         (int)(uint16_t)m->nbr[2] | ((int)(uint16_t)m->nbr[3] << 16), (int)m->site);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  auto nbr_of = [](const int4& e, int s) -> int {
    const int w = (s < 2) ? e.y : e.z;
    return (int)(short)((s & 1) ? ((unsigned)w >> 16) : (w & 0xffff));
  };

  if (warp == 0) {
    // ======================= MMA issue =======================
    const bool leader = elect_one();
    const uint32_t idesc = b4_idesc_bf16(128, N4, 0, 0);
    const uint32_t A_LBO = 128 * 16, B_LBO = (uint32_t)N4 * 16, SBO = 128;
    uint32_t n = 0, it = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      for (int q = P.Q - 1; q >= 0; --q) {
        const int nact = (stepT[q].x >> 16) & 0xff;
        for (int a = 0; a < nact; ++a, ++it) {
          const uint32_t buf = it & 1u;
          mbar_wait(bar_acc_free + buf, ((it >> 1) & 1u) ^ 1u);
          const uint32_t tmem_d = tmem_base + buf * 256u;
          for (int c = 0; c < NJ; ++c, ++n) {
            const uint32_t s = n % B4_STAGES;
            mbar_wait(bar_full + s, (n / B4_STAGES) & 1u);
            tc_fence_after();
            if (leader) {
              const uint32_t sa = smem_u32(stage0 + (size_t)s * B4_STAGE);
              const uint64_t adesc = tc_desc(sa, A_LBO, SBO), bdesc = tc_desc(sa + B4_A_BYTES, B_LBO, SBO);
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                tc_mma_f16(tmem_d, adesc + (uint64_t)((ks * 2 * A_LBO) >> 4), bdesc + (uint64_t)((ks * 2 * B_LBO) >> 4), idesc, (c > 0 || ks > 0) ? 1u : 0u);
              tc_commit(bar_empty + s);
              if (c == NJ - 1) tc_commit(bar_acc_full + buf);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 17) {
    // ======================= loader =======================
    const bool leader = elect_one();
    uint32_t n = 0, hs = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      for (int q = P.Q - 1; q >= 0; --q, ++hs) {
        const int4 e = stepT[q];
        const int nact = (e.x >> 16) & 0xff;
        const unsigned char* asrc = P.dimg + ((size_t)q * (size_t)ntiles + (size_t)tile) * dimg_tile;
        // every cell's phase of bar_ready is observed, also when the cell has no neighbour to push into (a skipped phase
        // would alias the parity of the next one)
        if (nact == 0) mbar_wait(bar_ready, hs & 1u);
        for (int a = 0; a < nact; ++a) {
          const int slot = (e.x >> (4 * a)) & 15;
          for (int c = 0; c < NJ; ++c, ++n) {
            const uint32_t s = n % B4_STAGES;
            if (n >= B4_STAGES) mbar_wait_relaxed(bar_empty + s, ((n / B4_STAGES) - 1u) & 1u);
            unsigned char* dst = stage0 + (size_t)s * B4_STAGE;
            if (leader) {
              mbar_expect_tx(bar_full + s, B4_A_BYTES + b_bytes);
              bulk_g2s(dst + B4_A_BYTES, P.Wt + ((size_t)slot * NJ + c) * b_bytes, b_bytes, bar_full + s);
            }
            if (a == 0 && c == 0) { mbar_wait(bar_ready, hs & 1u); fence_proxy_async_all(); }
            if (leader) bulk_g2s(dst, asrc + (size_t)c * B4_A_BYTES, B4_A_BYTES, bar_full + s);
            __syncwarp();
          }
        }
      }
    }
  } else {
    // ======================= producers / epilogue: 4 threads per row =======================
    const int quarter = warp & 3, sub = (warp - 1) >> 2;
    const int row = (quarter << 5) + lane;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const bool active = sub < NP;                           // warp-uniform
    uint32_t it = 0, hs = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long long grow = tile * 128 + row;
      const bool live = grow < P.R;
      const float seed = live ? (float)P.seed[grow] : 0.f;
      const int8_t* crow = P.cfg + (live ? grow : 0) * P.nsites;
      for (int q = P.Q - 1; q >= 0; --q, ++hs) {
        const int4 e = stepT[q];
        const int nact = (e.x >> 16) & 0xff;
        const size_t qt = (size_t)q * (size_t)ntiles + (size_t)tile;
        const float* hq = P.hist + ((size_t)q * (size_t)P.R + (size_t)(live ? grow : 0)) * nhp;
        // ---- head: partial logits over this thread's 64 units; the hidden-state image for the weight gradients
        float lg[8];
#pragma unroll
        for (int d = 0; d < 8; ++d) lg[d] = 0.f;
        if (active) {
          unsigned char* hdst = P.himg + qt * himg_tile + (size_t)(8 * sub) * 2048 + (size_t)row * 16;
#pragma unroll 2
          for (int kc = 0; kc < 8; ++kc) {
            const int j0 = 64 * sub + 8 * kc;
            float hv[8];
#pragma unroll
            for (int i = 0; i < 8; i += 4) {
              float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
              if (live && j0 + i < nhp) v = __ldg(reinterpret_cast<const float4*>(hq + j0 + i));
              hv[i] = v.x; hv[i + 1] = v.y; hv[i + 2] = v.z; hv[i + 3] = v.w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float4 wa = *reinterpret_cast<const float4*>(headW + (j0 + i) * 8);
              const float4 wb = *reinterpret_cast<const float4*>(headW + (j0 + i) * 8 + 4);
              lg[0] = fmaf(hv[i], wa.x, lg[0]); lg[1] = fmaf(hv[i], wa.y, lg[1]); lg[2] = fmaf(hv[i], wa.z, lg[2]); lg[3] = fmaf(hv[i], wa.w, lg[3]);
              lg[4] = fmaf(hv[i], wb.x, lg[4]); lg[5] = fmaf(hv[i], wb.y, lg[5]); lg[6] = fmaf(hv[i], wb.z, lg[6]); lg[7] = fmaf(hv[i], wb.w, lg[7]);
            }
            *reinterpret_cast<uint4*>(hdst + (size_t)kc * 2048) =
                make_uint4(pack_bf16x2(hv[0], hv[1]), pack_bf16x2(hv[2], hv[3]), pack_bf16x2(hv[4], hv[5]), pack_bf16x2(hv[6], hv[7]));
          }
        }
        float4* pbuf = partS + (size_t)(hs & 1u) * (128 * 4 * 2);
        pbuf[(row * 4 + sub) * 2] = make_float4(lg[0], lg[1], lg[2], lg[3]);
        pbuf[(row * 4 + sub) * 2 + 1] = make_float4(lg[4], lg[5], lg[6], lg[7]);
        quarter_barrier(quarter);
        float dl[8];
        {
          float z[8];
#pragma unroll
          for (int d = 0; d < 8; ++d) z[d] = headB[d];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 a = pbuf[(row * 4 + i) * 2], b = pbuf[(row * 4 + i) * 2 + 1];
            z[0] += a.x; z[1] += a.y; z[2] += a.z; z[3] += a.w; z[4] += b.x; z[5] += b.y; z[6] += b.z; z[7] += b.w;
          }
          float mx = z[0];
#pragma unroll
          for (int d = 1; d < 8; ++d) mx = fmaxf(mx, z[d]);
          float ssum = 0.f;
#pragma unroll
          for (int d = 0; d < 8; ++d) { z[d] = expf(z[d] - mx); ssum += z[d]; }
          const int site = e.w;
          const int sy = (crow[site * 3] & 1) | ((crow[site * 3 + 1] & 1) << 1) | ((crow[site * 3 + 2] & 1) << 2);
#pragma unroll
          for (int d = 0; d < 8; ++d) dl[d] = seed * (((d == sy) ? 1.f : 0.f) - z[d] / ssum);
          if (sub == 0 && live) {
            float4* o = reinterpret_cast<float4*>(P.dlog + ((size_t)q * (size_t)P.R + (size_t)grow) * 8);
            o[0] = make_float4(dl[0], dl[1], dl[2], dl[3]);
            o[1] = make_float4(dl[4], dl[5], dl[6], dl[7]);
          }
          if (sub == 0) {     // image for the head weight gradient (dead rows: seed = 0 -> zeros)
            unsigned char* o = P.dlimg + ((size_t)q * (size_t)ntiles + (size_t)tile) * 4096 + (size_t)row * 16;
            *reinterpret_cast<uint4*>(o) = make_uint4(pack_bf16x2(dl[0], dl[1]), pack_bf16x2(dl[2], dl[3]), pack_bf16x2(dl[4], dl[5]), pack_bf16x2(dl[6], dl[7]));
            *reinterpret_cast<uint4*>(o + 2048) = make_uint4(0u, 0u, 0u, 0u);
          }
        }
        // ---- one-hot input image (k-core = neighbour slot = sub): pairs of the neighbour cell's three bits, constant 1 at
        //      position 6 of slot 0 (its weight-gradient row is the bias gradient)
        {
          const int nb = nbr_of(e, sub);
          uint32_t w0 = 0u, w1 = 0u, w2 = 0u, w3 = (sub == 0 && live) ? 0x3F80u : 0u;
          if (nb >= 0 && live) {
            const int ns = stepT[nb].w;
            w0 = (crow[ns * 3] & 1) ? 0x3F800000u : 0x3F80u;
            w1 = (crow[ns * 3 + 1] & 1) ? 0x3F800000u : 0x3F80u;
            w2 = (crow[ns * 3 + 2] & 1) ? 0x3F800000u : 0x3F80u;
          }
          *reinterpret_cast<uint4*>(P.ximg + qt * ximg_tile + (size_t)sub * 2048 + (size_t)row * 16) = make_uint4(w0, w1, w2, w3);
        }
        // ---- d h_q (+ the head's push) -> (ds, dg, dm): BF16 chunk images (+ optional FP32 copies)
        if (active) {
          const float* dhq = P.dh + ((size_t)q * (size_t)P.R + (size_t)(live ? grow : 0)) * nhp;
          const float4* gq = P.gates + qt * (size_t)P.Gp * 128 + row;         // unit j at gq[j * 128]
          unsigned char* ddst = P.dimg + qt * dimg_tile + (size_t)sub * B4_A_BYTES + (size_t)row * 16;
#pragma unroll 1
          for (int kc = 0; kc < 8; ++kc) {
            const int j0 = 64 * sub + 8 * kc;
            float dsv[8], dgv[8], dmv[8];
            float dh8[8];                                            // this thread's 8 d h values of the chunk: two 16-byte loads
#pragma unroll
            for (int i = 0; i < 8; i += 4) {
              float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
              if (live && j0 + i < nhp) v = *reinterpret_cast<const float4*>(dhq + j0 + i);
              dh8[i] = v.x; dh8[i + 1] = v.y; dh8[i + 2] = v.z; dh8[i + 3] = v.w;
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const int j = j0 + i;
              dsv[i] = dgv[i] = dmv[i] = 0.f;
              if (live && j < nh) {
                const float4 wa = *reinterpret_cast<const float4*>(headW + j * 8);
                const float4 wb = *reinterpret_cast<const float4*>(headW + j * 8 + 4);
                float d = dh8[i];
                d = fmaf(wa.x, dl[0], d); d = fmaf(wa.y, dl[1], d); d = fmaf(wa.z, dl[2], d); d = fmaf(wa.w, dl[3], d);
                d = fmaf(wb.x, dl[4], d); d = fmaf(wb.y, dl[5], d); d = fmaf(wb.z, dl[6], d); d = fmaf(wb.w, dl[7], d);
                const float4 g = __ldg(gq + (size_t)j * 128);        // (u, h~, m, -)
                dsv[i] = d * g.x * (1.f - g.y * g.y);
                dgv[i] = d * (g.y - g.z) * g.x * (1.f - g.x);
                dmv[i] = d * (1.f - g.x);
                if (P.dp != nullptr) {
                  const size_t o = ((size_t)q * (size_t)P.R + (size_t)grow) * nh + j, blk = (size_t)P.Q * (size_t)P.R * nh;
                  P.dp[o] = dsv[i]; P.dp[blk + o] = dgv[i]; P.dp[2 * blk + o] = dmv[i];
                }
              }
            }
            unsigned char* d8 = ddst + (size_t)kc * 2048;
            *reinterpret_cast<uint4*>(d8) = make_uint4(pack_bf16x2(dsv[0], dsv[1]), pack_bf16x2(dsv[2], dsv[3]), pack_bf16x2(dsv[4], dsv[5]), pack_bf16x2(dsv[6], dsv[7]));
            *reinterpret_cast<uint4*>(d8 + (size_t)NP * B4_A_BYTES) = make_uint4(pack_bf16x2(dgv[0], dgv[1]), pack_bf16x2(dgv[2], dgv[3]), pack_bf16x2(dgv[4], dgv[5]), pack_bf16x2(dgv[6], dgv[7]));
            *reinterpret_cast<uint4*>(d8 + (size_t)2 * NP * B4_A_BYTES) = make_uint4(pack_bf16x2(dmv[0], dmv[1]), pack_bf16x2(dmv[2], dmv[3]), pack_bf16x2(dmv[4], dmv[5]), pack_bf16x2(dmv[6], dmv[7]));
          }
        }
        fence_proxy_async_all();
        mbar_arrive(bar_ready);
        // ---- one pass per live neighbour: d h_nbr[row][64 sub ..] += accumulator columns
        for (int a = 0; a < nact; ++a, ++it) {
          const uint32_t buf = it & 1u;
          mbar_wait(bar_acc_full + buf, (it >> 1) & 1u);
          if (active) {
            const int nb = nbr_of(e, (e.x >> (4 * a)) & 15);
            float* dst = P.dh + ((size_t)nb * (size_t)P.R + (size_t)(live ? grow : 0)) * nhp + 64 * sub;
            __syncwarp();
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < 64; c0 += 16) {
              uint32_t d[16];
              tc_ld16(tmem_base + lane_base + buf * 256u + (uint32_t)(64 * sub + c0), d);
              tc_wait_ld();
              if (live) {
#pragma unroll
                for (int i = 0; i < 16; i += 4) {
                  if (64 * sub + c0 + i < nhp) {
                    float4 v = *reinterpret_cast<float4*>(dst + c0 + i);
                    v.x += __uint_as_float(d[i]); v.y += __uint_as_float(d[i + 1]); v.z += __uint_as_float(d[i + 2]); v.w += __uint_as_float(d[i + 3]);
                    *reinterpret_cast<float4*>(dst + c0 + i) = v;
                  }
                }
              }
            }
          }
          tc_fence_before();
          mbar_arrive(bar_acc_free + buf);
        }
      }  // cells
    }  // tiles
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_base, 512); }
}

// transposed weight image of the reverse sweep: Wt[slot][chunk (blk, p)][k-core][n][8]:
//   B[n = hidden unit n of neighbour `slot`][k = (blk, unit j = 64 p + 8 k-core + i)] = W_blk[row(slot, n)][j]
__global__ void pack_bwd4_kernel(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                                 int nh, int NP, __nv_bfloat16* __restrict__ out) {
  const int N4 = 64 * NP, NJ = 3 * NP;
  const long long total = 4LL * NJ * 8 * N4 * 8;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    long long r = idx;
    const int i = (int)(r % 8); r /= 8;
    const int n = (int)(r % N4); r /= N4;
    const int kc = (int)(r % 8); r /= 8;
    const int c = (int)(r % NJ); r /= NJ;
    const int slot = (int)r;
    const int blk = c / NP, p = c % NP, j = 64 * p + 8 * kc + i;
    float v = 0.f;
    if (n < nh && j < nh) {
      if (blk == 2) v = Wmerge[(long long)(slot * nh + n) * nh + j];
      else v = (blk == 0 ? W : Wg)[(long long)(24 + slot * nh + n) * nh + j];
    }
    out[idx] = __float2bfloat16_rn(v);
  }
}

// ---- weight gradients of the four-neighbour GRU: grouped GEMM over (cell, tile) items ------------------------------------
// One job = one [128 x N4] output tile -- (neighbour slot, 128 input units, block s|g|m) or (the 32 one-hot input positions,
// block s|g) -- accumulated in TMEM over a contiguous range of (cell, tile) items and flushed with atomics.  Per item the
// contraction runs over the tile's 128 rows: A = hidden-state (or one-hot) image, B = d(pre-activation) image of the block, both
// MN-major (8 units contiguous per row, rows 16 bytes apart, 8-row groups 128 bytes apart, 8-unit groups 2048 bytes apart).
struct Dw4Job { int16_t kind, slot, m0, blk; int32_t item_begin, item_count; };   // kind 0: hidden-state rows, 1: one-hot input rows, 2: head weights
static_assert(sizeof(Dw4Job) == 16, "Dw4Job layout");

struct Dw4Params {
  int Q, nh, NP, tiles;
  const StepMeta* steps;
  const Dw4Job* jobs;
  const int* sites;           // [5][Q]: cells in which neighbour slot s is live (s = 4: every cell); counts in sites_n
  int sites_n[5];
  const unsigned char* dimg;
  const unsigned char* himg;
  const unsigned char* ximg;
  const unsigned char* dlimg;
  float* gW; float* gWg; float* gWmerge; float* gb; float* gbg; float* gDense;
};

constexpr int DW4_THREADS = 128;
constexpr uint32_t DW4_A_MAX = 32768, DW4_B_MAX = 65536, DW4_STAGE = DW4_A_MAX + DW4_B_MAX;
constexpr int DW4_SMEM = 2 * (int)DW4_STAGE + 256;

__global__ void __launch_bounds__(DW4_THREADS, 1) dw4_kernel(const Dw4Params P) {
  extern __shared__ __align__(128) unsigned char d4_smem[];
  unsigned char* const stage0 = d4_smem;
  uint64_t* const bars = reinterpret_cast<uint64_t*>(d4_smem + 2 * DW4_STAGE);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(d4_smem + 2 * DW4_STAGE + 64);
  uint64_t* const bar_full = bars + 0;     // [2]
  uint64_t* const bar_empty = bars + 2;    // [2]
  uint64_t* const bar_done = bars + 4;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int NP = P.NP, N4 = 64 * NP, nh = P.nh;
  const Dw4Job job = P.jobs[blockIdx.x];
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { mbar_init(bar_full + i, 1); mbar_init(bar_empty + i, 1); }
    mbar_init(bar_done, 1);
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, 256);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_slot;
  const uint32_t a_bytes = (job.kind != 1) ? (uint32_t)((N4 - job.m0 < 128 ? N4 - job.m0 : 128) / 8) * 2048u : 4u * 2048u;
  const uint32_t b_bytes = (job.kind == 2) ? 4096u : (uint32_t)NP * B4_A_BYTES;
  const int Nk = (job.kind == 2) ? 16 : N4;          // MMA columns: head job = 8 logits padded to 16
  const size_t dimg_tile = (size_t)3 * NP * B4_A_BYTES, himg_tile = (size_t)NP * B4_A_BYTES, ximg_tile = 4 * 2048;
  const int* sites = P.sites + (job.kind == 0 ? job.slot : 4) * P.Q;     // (input and head jobs: every cell)

  if (warp == 0) {
    const bool leader = elect_one();
    for (int i = 0; i < job.item_count; ++i) {
      const int item = job.item_begin + i, q = sites[item / P.tiles], tile = item % P.tiles;
      const uint32_t s = (uint32_t)i & 1u;
      if (i >= 2) mbar_wait_relaxed(bar_empty + s, ((uint32_t)(i >> 1) - 1u) & 1u);
      if (leader) {
        unsigned char* dst = stage0 + (size_t)s * DW4_STAGE;
        const size_t qt = (size_t)q * (size_t)P.tiles + (size_t)tile;
        const unsigned char* asrc = (job.kind == 0)
            ? P.himg + ((size_t)P.steps[q].nbr[job.slot] * (size_t)P.tiles + (size_t)tile) * himg_tile + (size_t)(job.m0 / 8) * 2048
            : (job.kind == 2 ? P.himg + qt * himg_tile + (size_t)(job.m0 / 8) * 2048 : P.ximg + qt * ximg_tile);
        const unsigned char* bsrc = (job.kind == 2) ? P.dlimg + qt * 4096 : P.dimg + qt * dimg_tile + (size_t)job.blk * b_bytes;
        mbar_expect_tx(bar_full + s, a_bytes + b_bytes);
        bulk_g2s(dst, asrc, a_bytes, bar_full + s);
        bulk_g2s(dst + DW4_A_MAX, bsrc, b_bytes, bar_full + s);
      }
      __syncwarp();
    }
  } else if (warp == 1) {
    const bool leader = elect_one();
    const uint32_t idesc = b4_idesc_bf16(128, Nk, 1, 1);
    for (int i = 0; i < job.item_count; ++i) {
      const uint32_t s = (uint32_t)i & 1u;
      mbar_wait(bar_full + s, (uint32_t)(i >> 1) & 1u);
      tc_fence_after();
      if (leader) {
        const uint32_t sa = smem_u32(stage0 + (size_t)s * DW4_STAGE);
        // MN-major, no swizzle: "leading" offset = stride between 8-row (K) groups, "stride" offset = between 8-unit (MN) groups
        const uint64_t adesc = tc_desc(sa, 128u, 2048u), bdesc = tc_desc(sa + DW4_A_MAX, 128u, 2048u);
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          tc_mma_f16(tmem_d, adesc + (uint64_t)((ks * 256) >> 4), bdesc + (uint64_t)((ks * 256) >> 4), idesc, (i > 0 || ks > 0) ? 1u : 0u);
        tc_commit(bar_empty + s);
        if (i == job.item_count - 1) tc_commit(bar_done);
      }
      __syncwarp();
    }
  }
  __syncwarp();
  if (job.item_count > 0) {
    mbar_wait(bar_done, 0u);
    tc_fence_after();
    // flush: thread = accumulator row m (TMEM lane)
    const int m = warp * 32 + lane;
    float* dst = nullptr;          // row of the gradient tensor this accumulator row belongs to
    int ncol = nh;                 // real columns of this job's output rows
    if (job.kind == 0) {
      const int k = job.m0 + m;
      if (k < nh) {
        if (job.blk == 2) dst = P.gWmerge + (size_t)(job.slot * nh + k) * nh;
        else dst = (job.blk == 0 ? P.gW : P.gWg) + (size_t)(24 + job.slot * nh + k) * nh;
      }
    } else if (job.kind == 2) {
      const int k = job.m0 + m;
      if (k < nh) dst = P.gDense + (size_t)k * 8;
      ncol = 8;
    } else if (m < 32) {
      const int slot = m >> 3, i = m & 7;
      if (i < 6) dst = (job.blk == 0 ? P.gW : P.gWg) + (size_t)(slot * 6 + i) * nh;
      else if (i == 6 && slot == 0) dst = (job.blk == 0 ? P.gb : P.gbg);
    }
#pragma unroll 1
    for (int c0 = 0; c0 < Nk; c0 += 16) {
      uint32_t d[16];
      __syncwarp();
      tc_ld16(tmem_d + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, d);
      tc_wait_ld();
      if (dst != nullptr) {
#pragma unroll
        for (int e = 0; e < 16; ++e) {
          const float v = __uint_as_float(d[e]);
          if (c0 + e < ncol && v != 0.f) atomicAdd(dst + c0 + e, v);
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_d, 256); }
}

}  // namespace rnnwf
