// Weight gradients of the tensorized 2D GRU on the tensor cores (reduced-precision regime).
//
//   dW [j][i][c]   = sum_{sites q, rows b} ds_j  hc_i  x_c        x = (1 - sl, sl, 1 - su, su)   (0 for an absent neighbour)
//   dWg[j][i][c]   = same with dg,         dWmerge[i][j] = sum dm_j hc_i,        hc = [h_h ; h_v]
// The one-hot factors are {0,1} masks over the rows, so for each source s in {ds, dg, dm} and each neighbour half the
// kernel accumulates three [128 x 112] FP32 tiles in TMEM over all (site, 128-row tile) items it is given:
//        S = s^T h,    Cl = (sl . s)^T h,    Cu = (su . s)^T h              (contraction over the rows)
// with tcgen05.mma kind::tf32: the A operand (units x rows) is written into TMEM by threads (lane <-> unit) that read the
// feature-major (ds, dg, dm) tile of the reverse sweep from shared memory -- the loader warp stages it there with one
// bulk copy (the reverse sweep writes the tiles with unit rows padded to 528 bytes), two tiles deep -- and apply the row
// masks in registers; the B
// operand is the hidden-state tile image the sampling sweep saved in UMMA core-matrix order, one bulk copy per item.
// Items of one CTA all belong to one neighbour class (both / horizontal only / vertical only), so at the end
//        dW[..,0] += S - Cl, dW[..,1] += Cl   (horizontal neighbour present),   dW[..,2] += S - Cu, dW[..,3] += Cu   (vertical)
// go to the gradient with atomicAdd.  Replaces the expanded operand h (x) x (3.2 GB at the benchmark size) and three
// long-K GEMMs with transposing loads.
// Reference behaviour: tf.gradients w.r.t. MDTensorizedGRUcell.py:35-47 parameters (W, Wg, Wmerge).
#pragma once
#include "bwd_tc.cuh"

namespace rnnwf {

// The operand writers are a latency chain per chunk (loads -> masks -> three tcgen05.st -> wait -> arrive, ~1600 cycles against
// ~340 cycles of MMAs), so DW_GROUPS groups of four warps take the chunks round-robin (chunk c -> group c % DW_GROUPS,
// buffer c % DW_ABUF): three chains in flight per SM.
constexpr int DW_GROUPS = 3;
constexpr int DW_THREADS = 32 + 128 * DW_GROUPS + 32;   // warp 0: MMA, then DW_GROUPS x 4 operand-writer warps (lane <-> unit), last warp: loader
constexpr int DW_KC = 16;                   // rows per operand chunk
constexpr int DW_CHUNKS = TC_ROWS / DW_KC;  // 8 chunks per item
constexpr int DW_ABUF = 3;                  // operand chunk buffers in TMEM
static_assert(DW_GROUPS == DW_ABUF, "each writer group owns one operand buffer (it follows every phase of that buffer's barrier)");
constexpr int DW_APITCH = TC_DPT_PITCH;     // floats per unit row of the staged (ds | dg | dm) tile: 528 B, conflict-free 16-byte reads at lane <-> unit

// hidden-state tile image of one (site, 128-row tile): B operand [unit n][row k], K-major core matrices
//   float index = ((k / 4) * ncores + n / 8) * 32 + (n % 8) * 4 + (k % 4),     ncores = ceil(nhp / 16) * 2
__host__ __device__ inline int tc_himg_ncores(int nhp) { return ((nhp + 15) / 16) * 2; }
__host__ __device__ inline int tc_himg_floats(int nhp) { return (TC_ROWS / 4) * tc_himg_ncores(nhp) * 32; }
__host__ __device__ inline long long tc_himg_bytes(int Q, long long rows, int nhp) {
  return (long long)Q * ((rows + TC_ROWS - 1) / TC_ROWS) * tc_himg_floats(nhp) * 4;
}
// feature-major tiles of the reverse sweep: [src 3][site][tile][unit nhp][TC_DPT_PITCH (128 rows + padding)]
__host__ __device__ inline long long tc_dpt_bytes(int Q, long long rows, int nhp) {
  return 3LL * Q * ((rows + TC_ROWS - 1) / TC_ROWS) * nhp * TC_DPT_PITCH * 4;
}

struct DwJob {            // one CTA's share: a contiguous range of items of one (source, half, class)
  int src;                // 0 ds -> W, 1 dg -> Wg, 2 dm -> Wmerge
  int half;               // neighbour half of hc: 0 horizontal, 1 vertical
  int cls;                // 0 both neighbours present, 1 horizontal only, 2 vertical only
  int begin, count;       // items [begin, begin + count) of the class: item = step_index * tiles + tile
};

struct DwParams {
  const DwJob* jobs;
  const int* cls_steps;   // [3][Q] sites of each class, cls_count[3]
  int cls_count[3];
  const StepMeta* steps;
  int Q, nh, nhp, nsites;
  long long rows, tiles;
  const int8_t* cfg;
  const int16_t* site_of_t;
  const float* dpT;       // feature-major (ds, dg, dm)
  const float* himg;      // hidden-state tile images
  float* gW; float* gWg;
  float* gM0; float* gM1;   // unmasked source (src 2): [unit i][unit j] targets of the horizontal / vertical half
  float* gB; float* gBg;    // bias gradients of sources 0 / 1 from the constant-1 column (unit position nh) of the hidden-state image, or null
  int slab[3];              // which feature-major slab of dpT holds source 0, 1, 2
  int debug;                // timing experiments (RNNWF_TC_DEBUG bits 16..18: no MMAs / no operand loads / no masked copies), 0 in production
};

__global__ void __launch_bounds__(DW_THREADS, 1) dw_tc_kernel(const DwParams P) {
  extern __shared__ __align__(128) unsigned char dw_smem[];
  const int ncores = tc_himg_ncores(P.nhp);
  const uint32_t img_bytes = (uint32_t)tc_himg_floats(P.nhp) * 4u;
  unsigned char* const Bst = dw_smem;                                       // [2][img_bytes]
  const uint32_t at_bytes = (uint32_t)P.nhp * DW_APITCH * 4u;
  float* const Ast = reinterpret_cast<float*>(dw_smem + 2 * img_bytes);     // [2][nhp][DW_APITCH] staged (ds | dg | dm) tiles
  unsigned char* const ctl = dw_smem + 2 * img_bytes + 2 * at_bytes;
  uint64_t* const bars = reinterpret_cast<uint64_t*>(ctl);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(ctl + 120);
  uint32_t* const maskS = reinterpret_cast<uint32_t*>(ctl + 128);           // [2 items][2 masks][4 words]
  uint64_t* const bar_At = bars + 10;     // [2] staged tile landed (complete_tx)
  uint64_t* const bar_Atfree = bars + 12; // [2] staged tile read by every writer thread
  uint64_t* const bar_A = bars + 0;       // [DW_ABUF] operand chunk written (128 arrivals)
  uint64_t* const bar_Afree = bars + 3;   // [DW_ABUF] operand chunk consumed (commit)
  uint64_t* const bar_B = bars + 6;       // [2] hidden-state tile landed (complete_tx)
  uint64_t* const bar_Bfree = bars + 8;   // [2] hidden-state tile consumed (commit)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const DwJob job = P.jobs[blockIdx.x];
  const int N = ncores * 8;                                   // MMA N (units padded to a multiple of 16)
  const bool masks = job.src != 2;                            // Wmerge has no one-hot factor
  const int nacc = masks ? 3 : 1;
  // TMEM map: accumulators [0, 3 N); operand chunks at 3 N: [DW_ABUF buffers][3 blocks][16 columns]
  if (tid == 0) {
    for (int i = 0; i < DW_ABUF; ++i) { mbar_init(bar_A + i, 128); mbar_init(bar_Afree + i, 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(bar_B + i, 1); mbar_init(bar_Bfree + i, 1); mbar_init(bar_At + i, 1); mbar_init(bar_Atfree + i, 128 * DW_GROUPS); }
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, TC_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_A0 = tmem_base + 3u * (uint32_t)N;
  const int* steps_of_cls = P.cls_steps + job.cls * P.Q;

  if (job.count > 0) {
    if (warp == 0) {
      // ======================= MMA warp =======================
      const bool leader = elect_one();
      const uint32_t LBO = (uint32_t)ncores * 128u, SBO = 128u;
      const uint64_t bdesc0 = tc_desc(smem_u32(Bst), LBO, SBO);
      const uint32_t idesc = tc_idesc_tf32(TC_ROWS, N);
      uint32_t phB = 0, c = 0;
      for (int it = 0; it < job.count; ++it) {
        const uint32_t st = (uint32_t)it & 1u;
        mbar_wait(bar_B + st, (phB >> st) & 1u); phB ^= 1u << st;
        for (int kc = 0; kc < DW_CHUNKS; ++kc, ++c) {
          const uint32_t s = c % DW_ABUF;
          mbar_wait(bar_A + s, (c / DW_ABUF) & 1u);
          tc_fence_after();
          if (leader) {
            const uint32_t a0 = tmem_A0 + s * 48u;
            // rows 16 kc .. + 15 of the tile = k-cores 4 kc .. + 3 of the image
            const uint64_t bd = bdesc0 + (uint64_t)((st * img_bytes + (uint32_t)(kc * 4) * LBO) >> 4);
            for (int b = 0; b < ((P.debug & (1 << 16)) ? 0 : nacc); ++b) {
#pragma unroll
              for (int ks = 0; ks < 2; ++ks)
                tc_mma_tf32_ts(tmem_base + (uint32_t)(b * N), a0 + (uint32_t)(b * 16 + ks * 8), bd + (uint64_t)((ks * 2u * LBO) >> 4), idesc,
                               (it > 0 || kc > 0 || ks > 0) ? 1u : 0u);
            }
            tc_commit(bar_Afree + s);
            if (kc == DW_CHUNKS - 1) tc_commit(bar_Bfree + st);
          }
          __syncwarp();
        }
      }
    } else if (warp == DW_THREADS / 32 - 1) {
      // ======================= loader warp: per item one bulk copy of the hidden-state image, one per unit row of the tile =====
      const bool leader = elect_one();
      uint32_t phF = 0;
      for (int it = 0; it < job.count; ++it) {
        const uint32_t st = (uint32_t)it & 1u;
        const int item = job.begin + it;
        const int q = steps_of_cls[item / (int)P.tiles];
        const long long tile = item % (int)P.tiles;
        const int nb = P.steps[q].nbr[job.half];
        if (it >= 2) {
          mbar_wait_relaxed(bar_Bfree + st, (phF >> st) & 1u);
          mbar_wait_relaxed(bar_Atfree + st, (phF >> st) & 1u);
          phF ^= 1u << st;
        }
        if (leader) {
          const unsigned char* src = reinterpret_cast<const unsigned char*>(P.himg) + ((size_t)nb * (size_t)P.tiles + (size_t)tile) * img_bytes;
          mbar_expect_tx(bar_B + st, img_bytes);
          bulk_g2s(Bst + (size_t)st * img_bytes, src, img_bytes, bar_B + st);
          const float* tsrc = P.dpT + (((size_t)P.slab[job.src] * P.Q + q) * (size_t)P.tiles + (size_t)tile) * P.nhp * DW_APITCH;
          mbar_expect_tx(bar_At + st, (P.debug & (1 << 17)) ? 0u : at_bytes);
          if (!(P.debug & (1 << 17))) bulk_g2s(Ast + (size_t)st * P.nhp * DW_APITCH, tsrc, at_bytes, bar_At + st);
        }
        __syncwarp();
      }
    } else {
      // ======================= operand writers: lane <-> unit, DW_GROUPS groups of four warps =======================
      const int quarter = warp & 3;                                // TMEM lane quarter this warp may access
      const int grp = (warp - 1) >> 2;
      const int unit = (quarter << 5) + lane;                      // TMEM lane == M index == unit j
      const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
      const bool real = unit < P.nh;
      const uint32_t total = (uint32_t)job.count * DW_CHUNKS;      // chunks of this job; chunk c = item (c / 8), rows 16 (c % 8) ..
      int cur_it = -1;
      for (uint32_t c = (uint32_t)grp; c < total; c += DW_GROUPS) {
        const int it = (int)(c / DW_CHUNKS), kc = (int)(c % DW_CHUNKS);
        while (cur_it < it) {
          if (cur_it >= 0) mbar_arrive(bar_Atfree + (cur_it & 1));  // this thread has read its chunks of the previous staged tile
          // row masks of the next item: bit k of word w <-> row 32 w + k has its horizontal / vertical neighbour spin up.
          // Every group passes every item (8 chunks over DW_GROUPS <= 4 groups); group 0 computes, all writers meet.
          ++cur_it;
          if (masks) {
            if (grp == 0) {
              const int item = job.begin + cur_it;
              const int q = steps_of_cls[item / (int)P.tiles];
              const long long tile = item % (int)P.tiles;
              const StepMeta* sm = P.steps + q;
              const long long r = tile * TC_ROWS + quarter * 32 + lane;
              int sl = 0, su = 0;
              if (r < P.rows) {
                const int8_t* cf = P.cfg + r * P.nsites;
                if (sm->nbr[0] >= 0) sl = cf[P.site_of_t[sm->nbr[0]]] & 1;
                if (sm->nbr[1] >= 0) su = cf[P.site_of_t[sm->nbr[1]]] & 1;
              }
              const uint32_t bl = __ballot_sync(0xffffffffu, sl), bu = __ballot_sync(0xffffffffu, su);
              if (lane == 0) { maskS[((cur_it & 1) * 2 + 0) * 4 + quarter] = bl; maskS[((cur_it & 1) * 2 + 1) * 4 + quarter] = bu; }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(128 * DW_GROUPS) : "memory");     // all writer warps
          }
          mbar_wait(bar_At + (cur_it & 1), ((uint32_t)cur_it >> 1) & 1u);           // the item's staged tile has landed
        }
        float v[16];
        {
          const float4* s4 = reinterpret_cast<const float4*>(Ast + ((size_t)(it & 1) * P.nhp + (real ? unit : 0)) * DW_APITCH + kc * DW_KC);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 t = real ? s4[i] : make_float4(0.f, 0.f, 0.f, 0.f);
            v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
          }
        }
        const uint32_t s = c % DW_ABUF, use = c / DW_ABUF;         // buffer s is used for the (use)-th time
        if (use > 0) mbar_wait(bar_Afree + s, (use - 1u) & 1u);
        const uint32_t a0 = tmem_A0 + lane_base + s * 48u;
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i] = __float_as_uint(v[i]);
        __syncwarp();
        tc_st16(a0, w);
        if (masks && !(P.debug & (1 << 18))) {
          const uint32_t ml = (maskS[((it & 1) * 2 + 0) * 4 + (kc >> 1)] >> ((kc & 1) * 16)) & 0xffffu;
          const uint32_t mu = (maskS[((it & 1) * 2 + 1) * 4 + (kc >> 1)] >> ((kc & 1) * 16)) & 0xffffu;
#pragma unroll
          for (int i = 0; i < 16; ++i) w[i] = ((ml >> i) & 1u) ? __float_as_uint(v[i]) : 0u;
          tc_st16(a0 + 16u, w);
#pragma unroll
          for (int i = 0; i < 16; ++i) w[i] = ((mu >> i) & 1u) ? __float_as_uint(v[i]) : 0u;
          tc_st16(a0 + 32u, w);
        }
        tc_wait_st();
        tc_fence_before();
        mbar_arrive(bar_A + s);
      }
      // ---- flush: wait for the last MMAs, then add this unit's accumulator rows to the gradient (column blocks shared
      //      between the groups)
      {
        // commits complete in order: the completion of the last chunk's commit covers every MMA of the job
        // (a group that does not own the last chunk's buffer has not followed its phases: the phase before last may still
        //  be open, the one before that is closed by the group's own last wait -- follow the last two in order)
        const uint32_t lastc = total - 1u, lastuse = lastc / DW_ABUF;
        if (lastuse > 0) mbar_wait(bar_Afree + (lastc % DW_ABUF), (lastuse - 1u) & 1u);
        mbar_wait(bar_Afree + (lastc % DW_ABUF), lastuse & 1u);
        tc_fence_after();
        const bool has_h = job.cls != 2, has_u = job.cls != 1;
        const int nh = P.nh;
        for (int n0 = 16 * grp; n0 < N; n0 += 16 * DW_GROUPS) {
          uint32_t S[16], Cl[16], Cu[16];
          __syncwarp();
          tc_ld16(tmem_base + lane_base + (uint32_t)n0, S);
          if (masks) { tc_ld16(tmem_base + lane_base + (uint32_t)(N + n0), Cl); tc_ld16(tmem_base + lane_base + (uint32_t)(2 * N + n0), Cu); }
          tc_wait_ld();
          if (real) {
#pragma unroll
            for (int e = 0; e < 16; ++e) {
              const int i = n0 + e;
              if (i == nh && masks && (job.half == 0 || job.cls == 2)) {
                // constant-1 column: sum over the rows of this job's items (every site with a neighbour is counted once:
                // through its horizontal half, or through the vertical one if it has no horizontal neighbour)
                float* gb = (job.src == 0) ? P.gB : P.gBg;
                if (gb != nullptr) atomicAdd(gb + unit, __uint_as_float(S[e]));
              }
              if (i >= nh) continue;
              const float s = __uint_as_float(S[e]);
              if (!masks) {
                atomicAdd((job.half ? P.gM1 : P.gM0) + (size_t)i * nh + unit, s);
              } else {
                float* g = (job.src == 0 ? P.gW : P.gWg) + ((size_t)unit * 2 * nh + (size_t)(job.half * nh + i)) * 4;
                const float cl = __uint_as_float(Cl[e]), cu = __uint_as_float(Cu[e]);
                if (has_h) { atomicAdd(g + 0, s - cl); atomicAdd(g + 1, cl); }
                if (has_u) { atomicAdd(g + 2, s - cu); atomicAdd(g + 3, cu); }
              }
            }
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_base, TC_TMEM_COLS); }
}

// bias gradients from the feature-major tiles: out[j] += sum over all (site, tile, row) of src[j]
// (nq: only the first nq sites -- dw_tc_kernel's constant-1 column covers every site that has a neighbour)
// blockIdx.z selects the source (ds -> out0, dg -> out1)
__global__ void __launch_bounds__(128) dpt_rowsum_kernel(const float* __restrict__ dpT, int Q, int nq, long long tiles, int nhp,
                                                         int nsplit, float* __restrict__ out0, float* __restrict__ out1) {
  const int src = blockIdx.z;
  float* const out = src ? out1 : out0;
  __shared__ float red[4];
  const int j = blockIdx.x;
  const long long nitems = (long long)nq * tiles;
  const float* base = dpT + (size_t)src * Q * tiles * nhp * TC_DPT_PITCH + (size_t)j * TC_DPT_PITCH + threadIdx.x;
  float s = 0.f;
  for (long long it = blockIdx.y; it < nitems; it += nsplit) s += base[(size_t)it * nhp * TC_DPT_PITCH];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) atomicAdd(out + j, red[0] + red[1] + red[2] + red[3]);
}

// head weight gradients from the hidden-state tile images: gH[j][d] += sum_rows h_j dlog_d; the heads' bias gradients
// (column sums of d(logits)) are taken along by the last four threads when they own no unit (gB / gBp non-null)
__global__ void __launch_bounds__(128) head_grad_himg_kernel(const float* __restrict__ himg, const float* __restrict__ dlog, int Q,
                                                             long long tiles, long long rows, int nh, int nhp, int ND,
                                                             float* __restrict__ gW, float* __restrict__ gWp,
                                                             float* __restrict__ gB, float* __restrict__ gBp,
                                                             const float* __restrict__ dpT, int nq, float* __restrict__ gb, float* __restrict__ gbg) {
  // (dpT, nq, gb, gbg: the cell's bias gradients of the first nq sites -- the ones without a neighbour, which the
  //  weight-gradient MMAs do not cover -- as row sums of the feature-major (ds | dg) tiles; null / 0: nothing to do)
  __shared__ float dl[TC_ROWS * 4];
  const int ncores = tc_himg_ncores(nhp);
  const int j = threadIdx.x;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  float bsum = 0.f, cb0 = 0.f, cb1 = 0.f;
  const long long nitems = (long long)Q * tiles;
  for (long long it = blockIdx.x; it < nitems; it += gridDim.x) {
    const int q = (int)(it / tiles);
    const long long tile = it % tiles;
    __syncthreads();
    for (int i = threadIdx.x; i < TC_ROWS * 4; i += 128) {
      const long long r = tile * TC_ROWS + (i >> 2);
      const int d = i & 3;
      dl[i] = (r < rows && d < ND) ? dlog[((size_t)q * rows + r) * ND + d] : 0.f;
    }
    __syncthreads();
    if (gb != nullptr && q < nq && j < nh) {
      const float* d0 = dpT + ((size_t)it * nhp + j) * TC_DPT_PITCH;
      const float* d1 = d0 + (size_t)Q * tiles * nhp * TC_DPT_PITCH;
      for (int r = 0; r < TC_ROWS; r += 4) {
        const float4 a = *reinterpret_cast<const float4*>(d0 + r);
        cb0 += (a.x + a.y) + (a.z + a.w);
        if (gbg != nullptr) { const float4 c = *reinterpret_cast<const float4*>(d1 + r); cb1 += (c.x + c.y) + (c.z + c.w); }
      }
    }
    if (gB != nullptr && j >= 124) {
      float s = 0.f;
      for (int r = 0; r < TC_ROWS; ++r) s += dl[r * 4 + (j - 124)];
      bsum += s;
    }
    if (j < nh) {
      const float4* img = reinterpret_cast<const float4*>(himg + (size_t)it * tc_himg_floats(nhp)) + (size_t)(j >> 3) * 8 + (j & 7);
      for (int kc = 0; kc < TC_ROWS / 4; ++kc) {
        const float4 h = __ldg(img + (size_t)kc * ncores * 8);
        const float hv[4] = {h.x, h.y, h.z, h.w};
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int d = 0; d < 4; ++d) acc[d] = fmaf(hv[r], dl[(kc * 4 + r) * 4 + d], acc[d]);
      }
    }
  }
  if (j < nh) {
    // dense.W is [nh][D = 2], dense_phase.W is [nh][2]
    atomicAdd(gW + j * 2 + 0, acc[0]); atomicAdd(gW + j * 2 + 1, acc[1]);
    if (gWp != nullptr) { atomicAdd(gWp + j * 2 + 0, acc[2]); atomicAdd(gWp + j * 2 + 1, acc[3]); }
  }
  if (gb != nullptr && j < nh && nq > 0) {
    if (cb0 != 0.f) atomicAdd(gb + j, cb0);
    if (gbg != nullptr && cb1 != 0.f) atomicAdd(gbg + j, cb1);
  }
  if (gB != nullptr && j >= 124) {
    const int d = j - 124;                         // dense.b is [2], dense_phase.b is [2]
    if (d < 2) atomicAdd(gB + d, bsum);
    else if (gBp != nullptr && d < ND) atomicAdd(gBp + (d - 2), bsum);
  }
}

}  // namespace rnnwf
