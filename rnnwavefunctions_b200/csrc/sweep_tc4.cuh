// Tensor-core regime of the lattice sweep for the four-neighbour 2D GRU of the Rydberg Kagome wavefunction
// (RydbergKagome_2DGRU/MDGRUcell.py:48-59, MDRNNWavefunction.py:73-238).
//
// Per unit cell and 128-row tile the cell is one dense contraction  [128 x K] x [K x 3 nh]  with
// K = (active neighbours) * nh + 32:  candidate s, update gate g and the merge m of every unit, the neighbours being
// separable K blocks (z W = sum_i x_i Wx_i + h_i Wh_i), so a neighbour that is absent (lattice edge, not generated
// yet) is simply skipped.  With nh = 256 neither the operand (K = 1056 halves) nor the accumulators (768 columns) fit
// in TMEM or shared memory as a whole, so BOTH operands are streamed:
//   * column passes of 64 units x (s, g, m) = 192 MMA columns, accumulators double-buffered in TMEM (2 x 192 columns);
//   * per pass a K loop over "jobs" of K = 64: the A chunk (128 rows x 64 hidden units of one neighbour, FP16, 16 KB) is
//     bulk-copied from the L2-resident history ring, which the epilogue writes directly in the UMMA K-major core-matrix
//     layout; the B chunk (192 x 64 weights, 24 KB) from the pre-packed weight image; 4 stages of 40 KB;
//   * the one-hot inputs (4 neighbours x 3 bits) and the biases ride on a last K = 32 job whose A chunk the epilogue
//     threads write into shared memory from the row's symbols (one k-core per neighbour; a constant 1 for the biases).
// The in-row predecessor ("carry") is ordered last in the K loop and its chunk c is released as soon as column pass c
// of the previous site has been written, so the next site's MMAs start under the current site's epilogue.
//
//   warp 0       : MMA issue (tcgen05.mma kind::f16, both operands in shared memory), tcgen05.commit to the stages / epilogue
//   warps 1..16  : epilogue, four threads per row (row == TMEM lane); thread (row, sub) owns units 64 p + 16 sub + [0,16) of
//                  pass p: 48 accumulators -> gates -> FP16 ring chunk (two 16-byte k-core rows), FP32 saved state, head partials
//   warp 17      : heads, four rows per thread: 8-way softmax, inverse-CDF draw or teacher forcing, running log P (FP64)
//   warp 18, 19  : loaders: cp.async.bulk of the B chunk (weights; no dependencies) and of the A chunk (ring; gated by the
//                  epilogue's chunk counters) of every stage -- two issuers because one thread issues a bulk copy per ~300 cycles
// Ring writes (generic proxy) are published to the loader's bulk copies (async proxy) by fence.proxy.async on both
// sides of a shared-memory counter.  Every wait is bounded and traps instead of hanging.
#pragma once
#include "sweep_tc.cuh"

namespace rnnwf {

constexpr int T4_ROWS = 128;
constexpr int T4_UP = 64;                         // units per column pass == hidden units per K job
constexpr int T4_N = 3 * T4_UP;                   // 192 MMA columns per pass: (s, g, m) of 64 units
constexpr int T4_TPS = T4_N / 4;                  // 48 accumulator columns per epilogue thread
constexpr int T4_MAX_STAGES = 4;                  // stages of 40 KB (A chunk + B chunk)
constexpr uint32_t T4_A_BYTES = T4_ROWS * T4_UP * 2;      // 16 KB: [k-core 8][row 128][8 halves]
constexpr uint32_t T4_B_BYTES = T4_N * T4_UP * 2;         // 24 KB: [k-core 8][n 192][8 halves]
constexpr uint32_t T4_STAGE = T4_A_BYTES + T4_B_BYTES;
constexpr uint32_t T4_XA_BYTES = 4 * T4_ROWS * 16;        // input chunk: [k-core = neighbour 4][row 128][8 halves]
constexpr uint32_t T4_XB_BYTES = 4 * T4_N * 16;           // its weights: [k-core 4][n 192][8 halves]
constexpr int T4_EPI_THREADS = 512;
constexpr int T4_HEAD_WARP = 17, T4_LOADB_WARP = 18, T4_LOAD_WARP = 19;   // (T4_LOAD_WARP: the hidden-state loader)
constexpr int T4_THREADS = 32 * 20;
constexpr int T4_HEAD_ROWS = 4;
constexpr int T4_MAXNP = 4;                       // num_units <= 256
constexpr uint32_t T4_TMEM_COLS = 512;

// Split precision (precision tc3): every FP16 image exists twice, hi + lo of the scaled value (TC_SPLIT_HS h, TC_SPLIT_WS w),
// and a K job covers 32 hidden units instead of 64 so that a stage still is 40 KB:
//   stage  = [A hi 8 KB | A lo 8 KB | B hi 12 KB | B lo 12 KB],  six MMAs per job (hi hi + lo hi + hi lo per K = 16 step);
//   ring   = per site and 64-unit chunk [half 2][hi | lo][k-core 4][row 128][8 halves]   (one 16 KB bulk copy per job);
//   weights= per pass [job (slot, chunk, half)][hi | lo][k-core 4][n 192][8 halves], then the input / bias job [hi | lo][k-core 4].
__host__ __device__ inline long long t4_pass_bytes(int NP, bool split = false) { return (4LL * NP * T4_B_BYTES + T4_XB_BYTES) * (split ? 2 : 1); }
__host__ __device__ inline long long t4_step_bytes(int NP, bool split = false) { return (long long)NP * T4_A_BYTES * (split ? 2 : 1); }   // one site's ring image

struct T4Smem {
  int off_stage, off_xa, off_part, off_headw, off_headb, off_step, off_last, off_sym, off_bar, off_cnt, off_tmem, off_sched, total;
  __host__ __device__ T4Smem(int stages, int Q) {
    int o = 0;
    off_stage = o; o += stages * (int)T4_STAGE;
    off_xa = o;    o += (int)T4_XA_BYTES;
    off_part = o;  o += T4_ROWS * 4 * 32;           // partial head sums [row][sub][8]
    off_headw = o; o += T4_MAXNP * T4_UP * 32;      // head weights [unit][8], zero rows past num_units
    off_headb = o; o += 32;
    off_step = o;  o += Q * 16;
    off_last = o;  o += (Q * 2 + 15) & ~15;         // per step: last step that reads it as a neighbour
    off_sym = o;   o += Q * T4_ROWS;                // symbols (unit-cell classes 0..7), one byte per (cell, row)
    off_bar = o;   o += 256;
    off_cnt = o;   o += 32;                         // ring-chunk / prologue counters
    off_tmem = o;  o += 16;
    off_sched = o; o += 32;                         // this CTA's tile range under the parent-block-major schedule
    total = o + 128;
  }
};

// stages of the operand ring.  (Measured dead ends, tools/tc4_cycles.py: a fifth stage at the price of reading the head weights
// from global memory, streaming loads for the parent-history gather, an L2 prefetch of the next tile's parent history,
// separate A / B rings -- the stream is bound by
// bytes in flight per SM x ~3k cycles of L2 latency under load, tools/ubench/bulk_rate.cu.)
__host__ __device__ inline int t4_stages(int Q) { (void)Q; return T4_MAX_STAGES; }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
// generic-proxy writes to global memory -> later async-proxy (bulk copy) reads of them; executed by the WRITER before it
// signals.  The full fence.proxy.async costs ~1000 cycles here (event trace, tools/tc4_cycles.py), so the reader side
// relies on the writers' fence + the release/acquire chain of the counter, as TMA-store epilogues do.
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ uint32_t ld_volatile_shared(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
// poll a shared-memory counter (one warp, off the critical path of everything but itself)
__device__ __forceinline__ void t4_wait_count(const uint32_t* cnt, uint32_t target) {
  for (uint32_t spin = 0; (int32_t)(ld_volatile_shared(cnt) - target) < 0; ++spin) {
    if (spin > 64) __nanosleep(20);                 // the chunk it waits for is usually a few hundred cycles away
    if (spin > (1u << 25)) { asm volatile("trap;"); }
  }
  __threadfence_block();
}

template <int MODE, bool SPLIT>
__global__ void __launch_bounds__(T4_THREADS, 1) sweep_tc4_kernel(const SweepParams P) {
  constexpr int KH = SPLIT ? 2 : 1;                      // K jobs per 64-unit chunk
  constexpr float HS = SPLIT ? TC_SPLIT_HS : 1.0f, DS = SPLIT ? 1.0f / (TC_SPLIT_HS * TC_SPLIT_WS) : 1.0f;
  constexpr uint32_t A_CHUNK = T4_A_BYTES * KH;          // ring bytes of one 64-unit chunk
  extern __shared__ __align__(128) unsigned char t4_smem[];
  const int NP = P.tc_np;
  const uint32_t NS = (uint32_t)t4_stages(P.Q);
  const T4Smem L((int)NS, P.Q);
  unsigned char* const smem = t4_smem;
  unsigned char* const stage0 = smem + L.off_stage;
  unsigned char* const xa = smem + L.off_xa;
  float4* const partS = reinterpret_cast<float4*>(smem + L.off_part);      // [row][sub][2]
  float* const headW = reinterpret_cast<float*>(smem + L.off_headw);
  float* const headB = reinterpret_cast<float*>(smem + L.off_headb);
  int4* const stepT = reinterpret_cast<int4*>(smem + L.off_step);
  int16_t* const lastS = reinterpret_cast<int16_t*>(smem + L.off_last);
  uint8_t* const symS = smem + L.off_sym;                                   // [Q][128]
  uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint32_t* const cnt = reinterpret_cast<uint32_t*>(smem + L.off_cnt);      // [0..3] ring chunk c written (warps), [4] prologues (warps)
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
  uint64_t* const bar_full = bars + 0;                 // [<= 5] stage landed (complete_tx, two producers)
  uint64_t* const bar_empty = bars + 5;                // [<= 5] stage consumed (tcgen05.commit)
  uint64_t* const bar_acc_full = bars + 10;            // [2] accumulator buffer complete (tcgen05.commit)
  uint64_t* const bar_acc_free = bars + 12;            // [2] accumulator buffer drained (512 arrivals)
  uint64_t* const bar_x = bars + 14;                   // input chunk of a step written (512 arrivals)
  uint64_t* const bar_part = bars + 15;                // partial head sums written (512 arrivals)
  uint64_t* const bar_pfree = bars + 16;               // partial head sums consumed (32 arrivals)
  uint64_t* const bar_sym = bars + 17;                 // sampling: the step's symbols are drawn (32 arrivals)
  uint64_t* const bar_tile = bars + 18;                // head warp is done with the tile's symbol table (32 arrivals)

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nh = P.nh, nhp = P.nhp;
  const long long pass_bytes = t4_pass_bytes(NP, SPLIT);
  const long long step_bytes = t4_step_bytes(NP, SPLIT);

  if (tid == 0) {
    for (int i = 0; i < T4_MAX_STAGES; ++i) { mbar_init(bar_full + i, 2); mbar_init(bar_empty + i, 1); }   // two producers per stage
    for (int i = 0; i < 2; ++i) { mbar_init(bar_acc_full + i, 1); mbar_init(bar_acc_free + i, T4_EPI_THREADS); }
    mbar_init(bar_x, T4_EPI_THREADS);
    mbar_init(bar_part, T4_EPI_THREADS);
    mbar_init(bar_pfree, 32);
    mbar_init(bar_sym, 32);
    mbar_init(bar_tile, 32);
    for (int i = 0; i < 8; ++i) cnt[i] = 0u;
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, T4_TMEM_COLS);
  for (int i = tid; i < NP * T4_UP * 8; i += T4_THREADS) {
    const int j = i >> 3, d = i & 7;
    headW[i] = (j < nh && d < P.ND) ? P.head_w[j * P.ND + d] : 0.f;
  }
  if (tid < 8) headB[tid] = (tid < P.ND) ? P.head_b[tid] : 0.f;
  for (int i = tid; i < P.Q; i += T4_THREADS) {
    // job order of step i: the neighbours that are present, the one generated right before this step ("carry") last
    const StepMeta* m = P.steps + i;
    int order = 0, nact = 0, carry = -1;
    for (int s = 0; s < 4; ++s) {
      const int nb = m->nbr[s];
      if (nb < 0) continue;
      if (nb == i - 1 && carry < 0) { carry = s; continue; }
      order |= s << (4 * nact); ++nact;
    }
    if (carry >= 0) { order |= carry << (4 * nact); ++nact; }
    stepT[i] = make_int4(order | (nact << 16) | ((carry >= 0 ? 1 : 0) << 24),
                         (int)(uint16_t)m->nbr[0] | ((int)(uint16_t)m->nbr[1] << 16),
                         (int)(uint16_t)m->nbr[2] | ((int)(uint16_t)m->nbr[3] << 16), (int)m->site);
    int last = -1;
    for (int q = i + 1; q < P.Q; ++q) {
      const StepMeta* mq = P.steps + q;
      if (mq->nbr[0] == i || mq->nbr[1] == i || mq->nbr[2] == i || mq->nbr[3] == i) last = q;
    }
    lastS[i] = (int16_t)last;
  }
  // Parent-block-major schedule of the connected-configuration sweep (single-flip Hamiltonians: the child list is regular,
  // child index (t bits + k) ns + n).  With prefix reuse every tile first converts its parents' saved history of the steps it
  // looks back to into the CTA's ring -- on this periodic lattice nearly all earlier steps, ~q0 x 128 KB per tile.  Tiles of
  // the SAME 128 parents restarting at increasing steps share that prefix: walked in that order by one CTA, each tile only
  // has to restore the steps the previous tile overwrote.  The linear order L = block (T bits) + (t bits + k) is cut into
  // gridDim.x contiguous ranges of equal cost (a tile costs T - t sites).
  long long* const schedS = reinterpret_cast<long long*>(smem + L.off_sched);
  const bool block_major = (MODE == MODE_CHILD) && P.par_hist != nullptr && P.sched_ns > 0 && P.rows_dev == nullptr;
  if (tid == 0 && block_major) {
    const long long T = P.Q, bits = P.sched_bits, G = T * bits, nb = (P.sched_ns + T4_ROWS - 1) / T4_ROWS;
    const long long Cb = bits * T * (T + 1) / 2, totc = nb * Cb;
    auto cut = [&](long long x) -> long long {
      const long long b = x / Cb, rem = x % Cb;
      long long acc = 0, r = 0;
      while (acc < rem) { acc += T - r / bits; ++r; }
      return b * G + r;
    };
    schedS[0] = cut(totc * (long long)blockIdx.x / (long long)gridDim.x);
    schedS[1] = (blockIdx.x + 1 == gridDim.x) ? nb * G : cut(totc * (long long)(blockIdx.x + 1) / (long long)gridDim.x);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // optional cycle accounting (-DRNNWF_TC_PROF builds only; tools/tc4_cycles.py).  MMA warp: [0] accumulator-free wait
  // [1] stage-full wait [2] input-chunk wait [3] issue [5] total; loader: [4] stage-empty wait [6] ring-chunk wait [7] total;
  // epilogue thread 32: [8] accumulator-full wait [9] pass compute [10] step tail (head hand-off, next input chunk) [12] total
#ifdef RNNWF_TC_PROF
  long long prof4[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) prof4[i] = 0;
  const bool do_prof4 = (P.tc_prof != nullptr) && (tid == 0 || tid == 32 || tid == 32 * T4_LOAD_WARP);
  const long long kstart4 = tc_clock();
#define T4_T0() long long t0__ = do_prof4 ? tc_clock() : 0
#define T4_ACC(k) do { if (do_prof4) { const long long t1__ = tc_clock(); prof4[k] += t1__ - t0__; t0__ = t1__; } } while (0)
#define T4_TOTAL(k) do { if (do_prof4) prof4[k] = tc_clock() - kstart4; } while (0)
  // event trace of CTA 0: [role 3][T4_TRACE_LEN] words (tag << 40 | cycles since kernel start) after the counters
  constexpr int T4_TRACE_LEN = 4096;
  const int trole4 = (tid == 0) ? 0 : (tid == 32 * T4_LOAD_WARP ? 1 : (tid == 32 ? 2 : -1));
  const bool do_trace4 = (P.tc_prof != nullptr) && blockIdx.x == 0 && trole4 >= 0;
  unsigned long long* const tbuf4 = do_trace4 ? P.tc_prof + (size_t)gridDim.x * 16 + (size_t)trole4 * T4_TRACE_LEN : nullptr;
  int nev4 = 0;
#define T4_EV(tag) do { if (do_trace4 && nev4 < T4_TRACE_LEN) tbuf4[nev4++] = ((unsigned long long)(tag) << 40) | (unsigned long long)(tc_clock() - kstart4); } while (0)
#else
#define T4_EV(tag) do { } while (0)
#define T4_T0() do { } while (0)
#define T4_ACC(k) do { } while (0)
#define T4_TOTAL(k) do { } while (0)
#endif
  const long long total_rows = (P.rows_dev != nullptr) ? (long long)(*P.rows_dev) : P.rows;
  const long long ntiles = (total_rows + T4_ROWS - 1) / T4_ROWS;
  // tile iteration of this CTA: `ti` runs over the schedule above, or over the tiles blockIdx.x, + gridDim.x, ...
  const long long ti_begin = block_major ? schedS[0] : (long long)blockIdx.x;
  const long long ti_end = block_major ? schedS[1] : ntiles;
  const long long ti_step = block_major ? 1 : (long long)gridDim.x;
  const long long sched_G = (long long)P.Q * P.sched_bits;
  auto tile_row0 = [&](long long ti) -> long long {
    return block_major ? (ti % sched_G) * P.sched_ns + (ti / sched_G) * T4_ROWS : ti * T4_ROWS;
  };
  auto tile_rows = [&](long long ti, long long row0) -> int {
    const long long left = block_major ? P.sched_ns - (ti / sched_G) * T4_ROWS : total_rows - row0;
    return (int)(left < (long long)T4_ROWS ? left : (long long)T4_ROWS);
  };
  auto tile_q0 = [&](long long row0) -> int {
    int ts0 = 0;
    if (MODE == MODE_CHILD && P.par_hist != nullptr) {
      const Child c0 = P.children[row0];
      ts0 = (c0.tb >= 0 && c0.tb < c0.ta) ? c0.tb : c0.ta;
    }
    return ts0;
  };
  auto nbr_of = [](const int4& e, int s) -> int {
    const int w = (s < 2) ? e.y : e.z;
    return (int)(short)((s & 1) ? ((unsigned)w >> 16) : (w & 0xffff));
  };
  unsigned char* const ring_cta = reinterpret_cast<unsigned char*>(P.ring) + (size_t)blockIdx.x * (size_t)P.Q * (size_t)step_bytes;

  if (warp == 0) {
    // ======================= MMA issue =======================
    const bool leader = elect_one();
    const uint32_t idesc = tc_idesc(T4_ROWS, T4_N);
    const uint32_t A_LBO = T4_ROWS * 16, B_LBO = T4_N * 16, SBO = 128;
    const uint64_t xdesc = tc_desc(smem_u32(xa), A_LBO, SBO);
    uint32_t n = 0, it = 0, hs = 0;
    for (long long ti = ti_begin; ti < ti_end; ti += ti_step) {
      const int q0 = tile_q0(tile_row0(ti));
      for (int q = q0; q < P.Q; ++q, ++hs) {
        const int nact = (stepT[q].x >> 16) & 0xff;
        const int njobs = nact * NP * KH;
        for (int p = 0; p < NP; ++p, ++it) {
          const uint32_t buf = it & 1u;
          T4_T0();
          mbar_wait(bar_acc_free + buf, ((it >> 1) & 1u) ^ 1u);
          T4_ACC(0); T4_EV(1);
          const uint32_t tmem_d = tmem_base + buf * (uint32_t)T4_N;
          for (int j = 0; j <= njobs; ++j, ++n) {
            const uint32_t s = n % NS;
            mbar_wait(bar_full + s, (n / NS) & 1u);
            T4_ACC(1);
            const bool xjob = (j == njobs);
            if (xjob && p == 0) mbar_wait(bar_x, hs & 1u);
            T4_ACC(2);
            tc_fence_after();
            if (leader) {
              const uint32_t sa = smem_u32(stage0 + (size_t)s * T4_STAGE);
              const uint64_t bdesc = tc_desc(sa + T4_A_BYTES, B_LBO, SBO);
              if (SPLIT) {
                // hi hi + lo hi + hi lo per K = 16 step (the one-hot inputs are exact: hi only)
                const uint64_t blo = tc_desc(sa + T4_A_BYTES + (xjob ? T4_XB_BYTES : T4_B_BYTES / 2), B_LBO, SBO);
                const uint64_t ahi = xjob ? xdesc : tc_desc(sa, A_LBO, SBO);
                const uint64_t alo = tc_desc(sa + T4_A_BYTES / 2, A_LBO, SBO);
                uint32_t accf = (j > 0) ? 1u : 0u;
#pragma unroll
                for (int ks = 0; ks < 2; ++ks) {
                  const uint64_t oa = (uint64_t)((ks * 2 * A_LBO) >> 4), ob = (uint64_t)((ks * 2 * B_LBO) >> 4);
                  tc_mma_f16(tmem_d, ahi + oa, bdesc + ob, idesc, accf); accf = 1u;
                  if (!xjob) tc_mma_f16(tmem_d, alo + oa, bdesc + ob, idesc, 1u);
                  tc_mma_f16(tmem_d, ahi + oa, blo + ob, idesc, 1u);
                }
              } else if (!xjob) {
                const uint64_t adesc = tc_desc(sa, A_LBO, SBO);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                  tc_mma_f16(tmem_d, adesc + (uint64_t)((ks * 2 * A_LBO) >> 4), bdesc + (uint64_t)((ks * 2 * B_LBO) >> 4), idesc, (j > 0 || ks > 0) ? 1u : 0u);
              } else {
#pragma unroll
                for (int ks = 0; ks < 2; ++ks)
                  tc_mma_f16(tmem_d, xdesc + (uint64_t)((ks * 2 * A_LBO) >> 4), bdesc + (uint64_t)((ks * 2 * B_LBO) >> 4), idesc, (j > 0 || ks > 0) ? 1u : 0u);
              }
              tc_commit(bar_empty + s);
              if (xjob) tc_commit(bar_acc_full + buf);
            }
            __syncwarp();
            T4_ACC(3);
            if (xjob) T4_EV(2);
          }
        }
      }
    }
    T4_TOTAL(5);
  } else if (warp == T4_LOADB_WARP) {
    // ======================= weight loader (no dependencies): B chunk of every job =======================
    // A single thread issues one bulk copy per ~300 cycles whatever its size (measured, tools/ubench/bulk_rate.cu), which is
    // about the tensor-core time of a job: the two operands of a stage are therefore issued by two warps.
    const bool leader = elect_one();
    uint32_t n = 0;
    const uint32_t bB = (P.debug_flags & 1) ? T4_B_BYTES / 2 : T4_B_BYTES;   // (RNNWF_TC_DEBUG bit 0: timing experiment)
    for (long long ti = ti_begin; ti < ti_end; ti += ti_step) {
      const int q0 = tile_q0(tile_row0(ti));
      for (int q = q0; q < P.Q; ++q) {
        const int4 e = stepT[q];
        const int nact = (e.x >> 16) & 0xff;
        for (int p = 0; p < NP; ++p) {
          const unsigned char* bsrc = P.tc_B + (size_t)p * (size_t)pass_bytes;
          const int cj = NP * KH;                        // jobs per neighbour
          for (int j = 0; j <= nact * cj; ++j, ++n) {
            const bool xjob = (j == nact * cj);
            const uint32_t s = n % NS;
            if (n >= NS) mbar_wait_relaxed(bar_empty + s, ((n / NS) - 1u) & 1u);
            if (leader) {
              const uint32_t bytes = xjob ? T4_XB_BYTES * KH : bB;
              const size_t off = xjob ? (size_t)4 * cj * T4_B_BYTES : (size_t)(((e.x >> (4 * (j / cj))) & 15) * cj + (j % cj)) * T4_B_BYTES;
              mbar_expect_tx(bar_full + s, bytes);
              bulk_g2s(stage0 + (size_t)s * T4_STAGE + T4_A_BYTES, bsrc + off, bytes, bar_full + s);
            }
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == T4_LOAD_WARP) {
    // ======================= hidden-state loader: A chunk of every job, gated by the epilogue's chunk counters ===========
    const bool leader = elect_one();
    uint32_t n = 0, hs = 0, nt = 0;
    uint32_t seen[T4_MAXNP] = {0u, 0u, 0u, 0u};     // ring chunk c: writer warps already waited for
    const uint32_t bA = (P.debug_flags & 2) ? T4_A_BYTES / 2 : T4_A_BYTES;   // (RNNWF_TC_DEBUG bit 1: timing experiment)
    for (long long ti = ti_begin; ti < ti_end; ti += ti_step, ++nt) {
      const int q0 = tile_q0(tile_row0(ti));
      const uint32_t hs0 = hs;
      // the tile's prologue (symbols, parent history -> ring) must be complete before any of its A chunks is read;
      // it only starts once the previous tile is finished, so every earlier ring write is covered as well
      t4_wait_count(cnt + 4, 16u * (nt + 1u));
      fence_proxy_async_all();
      for (int q = q0; q < P.Q; ++q, ++hs) {
        const int4 e = stepT[q];
        const int nact = (e.x >> 16) & 0xff;
        for (int p = 0; p < NP; ++p) {
          for (int a = 0; a < nact; ++a) {
            const int slot = (e.x >> (4 * a)) & 15;
            const int nb = nbr_of(e, slot);
            for (int ck = 0; ck < NP * KH; ++ck, ++n) {
              const int c = ck / KH;
              const uint32_t s = n % NS;
              T4_T0();
              if (n >= NS) mbar_wait_relaxed(bar_empty + s, ((n / NS) - 1u) & 1u);
              T4_ACC(4);
              if (nb >= q0) {
                // written by this tile's epilogue at local step hs0 + (nb - q0): 16 warps per (step, chunk)
                const uint32_t target = 16u * (hs0 + (uint32_t)(nb - q0) + 1u);
#pragma unroll
                for (int cc = 0; cc < T4_MAXNP; ++cc)
                  if (cc == c && (int32_t)(seen[cc] - target) < 0) { T4_EV(10 + 4 * cc); t4_wait_count(cnt + cc, target); seen[cc] = target; T4_EV(11 + 4 * cc); }
              }
              T4_ACC(6);
              if (leader) {
                mbar_expect_tx(bar_full + s, bA);
                bulk_g2s(stage0 + (size_t)s * T4_STAGE, ring_cta + (size_t)nb * (size_t)step_bytes + (size_t)ck * T4_A_BYTES, bA, bar_full + s);
              }
              __syncwarp();
            }
          }
          {   // input / bias job: its A chunk is the shared-memory input chunk; this producer only arrives
            const uint32_t s = n % NS;
            if (n >= NS) mbar_wait_relaxed(bar_empty + s, ((n / NS) - 1u) & 1u);
            if (leader) mbar_arrive(bar_full + s);
            __syncwarp();
            ++n;
          }
        }
      }
    }
    T4_TOTAL(7);
  } else if (warp <= T4_EPI_THREADS / 32) {
    // ======================= epilogue warps: 4 threads per row =======================
    const int quarter = warp & 3;
    const int sub = (warp - 1) >> 2;
    const int row = (quarter << 5) + lane;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const uint32_t acc0 = tmem_base + lane_base + (uint32_t)(sub * T4_TPS);
    // this thread's two k-core rows (units 16 sub + [0,8), + [8,16)) inside a 64-unit ring chunk; split precision: the lo image
    // of the same half follows the hi image at + T4_A_BYTES / 2
    const size_t ring_thr = (SPLIT ? (size_t)(sub >> 1) * T4_A_BYTES + (size_t)(sub & 1) * 2 * (T4_ROWS * 16) : (size_t)(2 * sub) * (T4_ROWS * 16)) + (size_t)row * 16;
    uint32_t it = 0, hs = 0, ntile = 0;
    long long prev_block = -1;
    int prev_q0 = 0;
    for (long long ti = ti_begin; ti < ti_end; ti += ti_step, ++ntile) {
      const long long row0 = tile_row0(ti);
      const long long tile = row0 / T4_ROWS;            // (index into the tile-blocked saved gates: sampling / evaluation only)
      const int nrows = tile_rows(ti, row0);
      const int q0 = tile_q0(row0);
      // same 128 parents as the previous tile and a later restart: the ring still holds their history below prev_q0
      const int restore_from = (block_major && ti / sched_G == prev_block && prev_q0 <= q0) ? prev_q0 : 0;
      prev_block = block_major ? ti / sched_G : -1;
      prev_q0 = q0;
      const bool live = row < nrows;
      int rparent = 0;
      T4_T0();
      if (ntile > 0) mbar_wait(bar_tile, (ntile - 1) & 1);
      T4_ACC(11);
      // ---- symbol table of the tile (teacher forcing)
      {
        Child ch;
        ch.parent = 0; ch.ta = -1; ch.tb = -1; ch.coef = 0.f; ch.ka = ch.kb = 0;
        if (live && MODE == MODE_CHILD) { ch = P.children[row0 + row]; rparent = ch.parent; }
        const long long src = (MODE == MODE_CHILD) ? (long long)ch.parent : row0 + row;
        const int8_t* c = P.cfg_in + src * P.nsites;
        for (int t = sub; t < P.Q; t += 4) {
          int sy = 0;
          if (live && MODE != MODE_SAMPLE) {
            const int site = stepT[t].w;
            sy = (c[site * 3] & 1) | ((c[site * 3 + 1] & 1) << 1) | ((c[site * 3 + 2] & 1) << 2);
            if (MODE == MODE_CHILD) {
              if (t == ch.ta) sy ^= 1 << ch.ka;
              if (t == ch.tb) sy ^= 1 << ch.kb;
            }
          }
          symS[t * T4_ROWS + row] = (uint8_t)sy;
        }
      }
      T4_ACC(13);
      // ---- prefix reuse: the parents' saved FP32 history of every earlier step some step >= q0 looks back to,
      //      converted into this CTA's ring (FP16 chunk images)
      if (MODE == MODE_CHILD && q0 > 0) {
        // (all 16 loads of a step are issued before the first conversion: the gather is latency-bound)
        const float* pbase = P.par_hist + (size_t)rparent * nhp + sub * 16;
        unsigned char* dbase = ring_cta + ring_thr;
#pragma unroll 1
        for (int j = restore_from; j < q0; ++j) {
          if ((int)lastS[j] < q0) continue;           // no step of this tile looks back to it (uniform over the CTA)
          const float* psrc = pbase + (size_t)j * (size_t)P.par_rows * nhp;
          unsigned char* dst = dbase + (size_t)j * (size_t)step_bytes;
          float4 v[T4_MAXNP][4];
#pragma unroll
          for (int p = 0; p < T4_MAXNP; ++p)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              v[p][i] = make_float4(0.f, 0.f, 0.f, 0.f);
              if (p < NP && live && p * T4_UP + sub * 16 + 4 * i < nhp) v[p][i] = __ldg(reinterpret_cast<const float4*>(psrc + p * T4_UP + 4 * i));
            }
#pragma unroll
          for (int p = 0; p < T4_MAXNP; ++p) {
            if (p < NP && !SPLIT) {
              *reinterpret_cast<uint4*>(dst + (size_t)p * T4_A_BYTES) =
                  make_uint4(pack_f16x2(v[p][0].x, v[p][0].y), pack_f16x2(v[p][0].z, v[p][0].w), pack_f16x2(v[p][1].x, v[p][1].y), pack_f16x2(v[p][1].z, v[p][1].w));
              *reinterpret_cast<uint4*>(dst + (size_t)p * T4_A_BYTES + T4_ROWS * 16) =
                  make_uint4(pack_f16x2(v[p][2].x, v[p][2].y), pack_f16x2(v[p][2].z, v[p][2].w), pack_f16x2(v[p][3].x, v[p][3].y), pack_f16x2(v[p][3].z, v[p][3].w));
            }
            if (p < NP && SPLIT) {
#pragma unroll
              for (int kc = 0; kc < 2; ++kc) {
                uint4 hi, lo;
                split_f16x2(v[p][2 * kc].x * HS, v[p][2 * kc].y * HS, hi.x, lo.x); split_f16x2(v[p][2 * kc].z * HS, v[p][2 * kc].w * HS, hi.y, lo.y);
                split_f16x2(v[p][2 * kc + 1].x * HS, v[p][2 * kc + 1].y * HS, hi.z, lo.z); split_f16x2(v[p][2 * kc + 1].z * HS, v[p][2 * kc + 1].w * HS, hi.w, lo.w);
                *reinterpret_cast<uint4*>(dst + (size_t)p * A_CHUNK + (size_t)kc * (T4_ROWS * 16)) = hi;
                *reinterpret_cast<uint4*>(dst + (size_t)p * A_CHUNK + (size_t)kc * (T4_ROWS * 16) + T4_A_BYTES / 2) = lo;
              }
            }
          }
        }
      }
      T4_ACC(14);
      // symbols of the other sub-threads are needed below: the four warps of a lane quarter hold one row's four threads
      quarter_barrier(quarter);
      // input chunk of step q: k-core `sub` = neighbour slot `sub`: one-hot pairs of the neighbour cell's three bits; the
      // constant 1 of the bias row sits at position 6 of slot 0
      auto write_x = [&](int q) {
        const int nb = nbr_of(stepT[q], sub);
        constexpr uint32_t ONE = SPLIT ? 0x5400u : 0x3C00u;      // FP16 1.0 (split precision: TC_SPLIT_HS = 64.0)
        uint32_t w0 = 0u, w1 = 0u, w2 = 0u, w3 = (sub == 0) ? ONE : 0u;
        if (nb >= 0) {
          const int sy = symS[nb * T4_ROWS + row];
          w0 = (sy & 1) ? (ONE << 16) : ONE;
          w1 = (sy & 2) ? (ONE << 16) : ONE;
          w2 = (sy & 4) ? (ONE << 16) : ONE;
        }
        *reinterpret_cast<uint4*>(xa + (size_t)sub * (T4_ROWS * 16) + (size_t)row * 16) = make_uint4(w0, w1, w2, w3);
        fence_proxy_async_smem();
        mbar_arrive(bar_x);
      };
      write_x(q0);
      // publish the prologue (ring writes of this thread) to the loader
      fence_proxy_async_all();
      __syncwarp();
      if (lane == 0) { __threadfence_block(); atomicAdd(cnt + 4, 1u); }
      T4_ACC(15);

      for (int q = q0; q < P.Q; ++q, ++hs) {
        const bool save = (MODE != MODE_CHILD) && live && (P.sv_hist != nullptr);
        float lg[8];
#pragma unroll
        for (int d = 0; d < 8; ++d) lg[d] = 0.f;
        unsigned char* ring_q = ring_cta + (size_t)q * (size_t)step_bytes + ring_thr;
#pragma unroll 1
        for (int p = 0; p < NP; ++p, ++it) {
          const uint32_t buf = it & 1u;
          T4_T0();
          mbar_wait(bar_acc_full + buf, (it >> 1) & 1u);
          T4_ACC(8); T4_EV(20 + p);
          uint32_t ds[16], dg[16], dm[16];
          __syncwarp();
          tc_fence_after();
          const uint32_t acc = acc0 + buf * (uint32_t)T4_N;
          tc_ld16(acc, ds);
          tc_ld16(acc + 16u, dg);
          tc_ld16(acc + 32u, dm);
          tc_wait_ld();
          tc_fence_before();
          mbar_arrive(bar_acc_free + buf);
          const int j0 = p * T4_UP + sub * 16;
          const float* hW = headW + j0 * 8;
          // The new hidden states go out first (the next site's MMAs wait for this chunk); the head partial sums and the
          // saved state follow, off the site-to-site critical path.
          float hv[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float u = SPLIT ? sigmoid_split(__uint_as_float(dg[e]) * DS) : sigmoid_prescaled(__uint_as_float(dg[e]));
            const float ht = SPLIT ? tanh_split(__uint_as_float(ds[e]) * DS) : tanh_prescaled(__uint_as_float(ds[e]));
            const float mv = __uint_as_float(dm[e]) * DS;
            // u ht + (1 - u) m; padding units have zero weights and biases: u = 1/2, ht = m = 0 -> 0
            hv[e] = live ? fmaf(u, ht - mv, mv) : 0.f;
            if (save) {
              const int j = j0 + e;
              // tile-blocked [cell][tile][unit][row 128]: a warp's 32 rows are 512 contiguous bytes (read back the same way by bwd4_kernel)
              if (j < nh) P.sv_gates[(((size_t)q * (size_t)((P.sv_rows + T4_ROWS - 1) / T4_ROWS) + (size_t)tile) * P.Gp + j) * T4_ROWS + row] = make_float4(u, ht, mv, 0.f);
            }
          }
          if (save) {        // saved history (row-major, prefix reuse gathers whole rows): four 16-byte stores per pass
            float* hrow = P.sv_hist + ((size_t)q * (size_t)P.sv_rows + (size_t)(row0 + row)) * nhp + j0;
#pragma unroll
            for (int e = 0; e < 16; e += 4)
              if (j0 + e < nhp) *reinterpret_cast<float4*>(hrow + e) = make_float4(hv[e], hv[e + 1], hv[e + 2], hv[e + 3]);
          }
          if (!SPLIT) {
            *reinterpret_cast<uint4*>(ring_q + (size_t)p * T4_A_BYTES) =
                make_uint4(pack_f16x2(hv[0], hv[1]), pack_f16x2(hv[2], hv[3]), pack_f16x2(hv[4], hv[5]), pack_f16x2(hv[6], hv[7]));
            *reinterpret_cast<uint4*>(ring_q + (size_t)p * T4_A_BYTES + T4_ROWS * 16) =
                make_uint4(pack_f16x2(hv[8], hv[9]), pack_f16x2(hv[10], hv[11]), pack_f16x2(hv[12], hv[13]), pack_f16x2(hv[14], hv[15]));
          } else {
#pragma unroll
            for (int kc = 0; kc < 2; ++kc) {
              uint4 hi, lo;
              split_f16x2(hv[8 * kc] * HS, hv[8 * kc + 1] * HS, hi.x, lo.x); split_f16x2(hv[8 * kc + 2] * HS, hv[8 * kc + 3] * HS, hi.y, lo.y);
              split_f16x2(hv[8 * kc + 4] * HS, hv[8 * kc + 5] * HS, hi.z, lo.z); split_f16x2(hv[8 * kc + 6] * HS, hv[8 * kc + 7] * HS, hi.w, lo.w);
              *reinterpret_cast<uint4*>(ring_q + (size_t)p * A_CHUNK + (size_t)kc * (T4_ROWS * 16)) = hi;
              *reinterpret_cast<uint4*>(ring_q + (size_t)p * A_CHUNK + (size_t)kc * (T4_ROWS * 16) + T4_A_BYTES / 2) = lo;
            }
          }
          // chunk p of this step is written: release it to the loader
          fence_proxy_async_global();
          __syncwarp();
          if (lane == 0) { __threadfence_block(); atomicAdd(cnt + p, 1u); }
          T4_EV(30 + p);
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float4 wa = *reinterpret_cast<const float4*>(hW + e * 8);
            const float4 wb = *reinterpret_cast<const float4*>(hW + e * 8 + 4);
            lg[0] = fmaf(hv[e], wa.x, lg[0]); lg[1] = fmaf(hv[e], wa.y, lg[1]); lg[2] = fmaf(hv[e], wa.z, lg[2]); lg[3] = fmaf(hv[e], wa.w, lg[3]);
            lg[4] = fmaf(hv[e], wb.x, lg[4]); lg[5] = fmaf(hv[e], wb.y, lg[5]); lg[6] = fmaf(hv[e], wb.z, lg[6]); lg[7] = fmaf(hv[e], wb.w, lg[7]);
          }
          T4_ACC(9); T4_EV(40);
        }
        T4_T0();
        // partial head sums of this step -> head warp
        mbar_wait(bar_pfree, (hs & 1u) ^ 1u);
        partS[(row * 4 + sub) * 2] = make_float4(lg[0], lg[1], lg[2], lg[3]);
        partS[(row * 4 + sub) * 2 + 1] = make_float4(lg[4], lg[5], lg[6], lg[7]);
        mbar_arrive(bar_part);
        if (q + 1 < P.Q) {
          // every MMA of this step has completed (its last accumulators were just read): the input chunk can be rewritten
          if (MODE == MODE_SAMPLE) mbar_wait(bar_sym, hs & 1u);
          write_x(q + 1);
        }
        T4_ACC(10);
      }  // steps
    }  // tiles
    T4_TOTAL(12);
  } else if (warp == T4_HEAD_WARP) {
    // ======================= head warp: four rows per thread =======================
    const int hrow = lane;
    uint32_t hs = 0;
    for (long long ti = ti_begin; ti < ti_end; ti += ti_step) {
      const long long row0 = tile_row0(ti);
      const int nrows = tile_rows(ti, row0);
      const int ts0 = tile_q0(row0), q0 = ts0;
      double rlogp[T4_HEAD_ROWS];
      int rparent[T4_HEAD_ROWS];
      float rcoef[T4_HEAD_ROWS];
      bool live[T4_HEAD_ROWS];
#pragma unroll
      for (int k = 0; k < T4_HEAD_ROWS; ++k) {
        const int row = hrow + k * 32;
        rlogp[k] = 0.0; rparent[k] = 0; rcoef[k] = 0.f;
        live[k] = row < nrows;
        if (live[k] && MODE == MODE_CHILD) { const Child ch = P.children[row0 + row]; rparent[k] = ch.parent; rcoef[k] = ch.coef; }
      }
      for (int q = q0; q < P.Q; ++q, ++hs) {
        const int t = q;
        mbar_wait_relaxed(bar_part, hs & 1u);
        float l8[T4_HEAD_ROWS][8];
#pragma unroll
        for (int k = 0; k < T4_HEAD_ROWS; ++k) {
          const int row = hrow + k * 32;
#pragma unroll
          for (int d = 0; d < 8; ++d) l8[k][d] = headB[d];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 a = partS[(row * 4 + i) * 2], b = partS[(row * 4 + i) * 2 + 1];
            l8[k][0] += a.x; l8[k][1] += a.y; l8[k][2] += a.z; l8[k][3] += a.w;
            l8[k][4] += b.x; l8[k][5] += b.y; l8[k][6] += b.z; l8[k][7] += b.w;
          }
        }
        mbar_arrive(bar_pfree);
#pragma unroll
        for (int k = 0; k < T4_HEAD_ROWS; ++k) {
          const int row = hrow + k * 32;
          if (!live[k]) continue;
          float mx = l8[k][0];
#pragma unroll
          for (int d = 1; d < 8; ++d) mx = fmaxf(mx, l8[k][d]);
          float pr[8], ssum = 0.f;
#pragma unroll
          for (int d = 0; d < 8; ++d) { pr[d] = SPLIT ? expf(l8[k][d] - mx) : __expf(l8[k][d] - mx); ssum += pr[d]; }
#pragma unroll
          for (int d = 0; d < 8; ++d) pr[d] = SPLIT ? pr[d] / ssum : __fdividef(pr[d], ssum);
          int sy;
          if (MODE == MODE_SAMPLE) {
            const float u = (P.uniforms != nullptr)
                                ? P.uniforms[(size_t)(row0 + row) * P.Tsym + t]
                                : philox_uniform(P.seed, P.row_offset + (unsigned long long)(row0 + row), (uint32_t)t);
            double tot = 0.0;
#pragma unroll
            for (int d = 0; d < 8; ++d) tot += (double)pr[d];
            const double target = (double)u * tot;
            double c = 0.0;
            sy = 0;
#pragma unroll
            for (int d = 0; d < 8; ++d) { c += (double)pr[d]; if (c <= target) ++sy; }
            if (sy > 7) sy = 7;
            symS[t * T4_ROWS + row] = (uint8_t)sy;
          } else {
            sy = symS[t * T4_ROWS + row];
          }
          float ps = pr[0];
#pragma unroll
          for (int d = 1; d < 8; ++d) if (d == sy) ps = pr[d];
          rlogp[k] += SPLIT ? (double)logf(ps) : (double)__logf(ps);
          if (P.sv_cum != nullptr && MODE != MODE_CHILD)
            P.sv_cum[(size_t)t * (size_t)P.sv_rows + (size_t)(row0 + row)] = make_double2(rlogp[k], 0.0);
        }
        if (MODE == MODE_SAMPLE) mbar_arrive(bar_sym);
      }  // steps

#pragma unroll
      for (int k = 0; k < T4_HEAD_ROWS; ++k) {
        const int row = hrow + k * 32;
        if (!live[k]) continue;
        if (MODE == MODE_CHILD) {
          const int par = rparent[k];
          double base_lp = P.par_logp[par];
          if (ts0 > 0 && P.par_cum != nullptr) base_lp -= P.par_cum[(size_t)(ts0 - 1) * (size_t)P.par_rows + par].x;
          const double amp = exp(0.5 * (rlogp[k] - base_lp)) * (double)rcoef[k];
          atomicAdd(P.eloc + 2 * (size_t)par, amp);
        } else {
          if (P.out_logp != nullptr) P.out_logp[row0 + row] = rlogp[k];
          if (P.out_phase != nullptr) P.out_phase[row0 + row] = 0.0;
          if (MODE == MODE_SAMPLE && P.cfg_out != nullptr) {
            int8_t* c = P.cfg_out + (size_t)(row0 + row) * P.nsites;
            for (int tt = 0; tt < P.Q; ++tt) {
              const int site = stepT[tt].w, sy = symS[tt * T4_ROWS + row];
              c[site * 3] = (int8_t)(sy & 1); c[site * 3 + 1] = (int8_t)((sy >> 1) & 1); c[site * 3 + 2] = (int8_t)((sy >> 2) & 1);
            }
          }
        }
      }
      mbar_arrive(bar_tile);
    }  // tiles
  }

#ifdef RNNWF_TC_PROF
  if (do_prof4) {
    const int role = (tid == 0) ? 0 : (tid == 32 ? 2 : 1);
    for (int i = 0; i < 16; ++i) {
      const int owner = (i == 4 || i == 6 || i == 7) ? 1 : (i < 8 ? 0 : 2);
      if (owner == role) P.tc_prof[(size_t)blockIdx.x * 16 + i] = (unsigned long long)prof4[i];
    }
  }
#endif
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_base, T4_TMEM_COLS); }
}

// ---- weight image: [pass NP][job][k-core][n 192][8 halves], jobs = (neighbour slot 4) x (chunk NP) of 8 k-cores, then the
// input/bias job of 4 k-cores; split precision: jobs = (slot, chunk, half) of [hi | lo][k-core 4] (layout comment at the top).
// Column n = sub * 48 + blk * 16 + e  <->  block blk (s, g, m) of unit 64 p + 16 sub + e.
// Rows of W / Wg (RydbergKagome_2DGRU/MDGRUcell.py:34-46): 24 input rows (slot * 6 + 2 bit + value), then 4 nh hidden rows.
template <int UNUSED = 0>
__global__ void pack_tc4_kernel(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wmerge,
                                const float* __restrict__ b, const float* __restrict__ bg, int nh, int NP, int split, __half* __restrict__ out) {
  const int KH = split ? 2 : 1;
  const long long per_pass = t4_pass_bytes(NP, split != 0) / 2;
  const long long total = (long long)NP * per_pass;
  const long long hjob = T4_B_BYTES / 2;                  // halves per job: [part KH][k-core 8 / KH][n 192][8]
  const long long hpart = hjob / KH;
  const long long xpart = T4_XB_BYTES / 2;                // input / bias job: [part KH][k-core 4][n 192][8]
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int p = (int)(idx / per_pass);
    long long r = idx - (long long)p * per_pass;
    const int i = (int)(r % 8);
    float v = 0.f;
    int blk, part;
    if (r < 4LL * NP * KH * hjob) {
      const int job = (int)(r / hjob); r -= (long long)job * hjob;
      part = (int)(r / hpart); r -= (long long)part * hpart;
      const int n = (int)((r / 8) % T4_N), kc = (int)(r / (8 * T4_N));
      const int slot = job / (NP * KH), c = (job / KH) % NP, kh = job % KH;
      const int sub = n / T4_TPS, e = n % 16;
      blk = (n % T4_TPS) / 16;
      const int j = p * T4_UP + sub * 16 + e, k = c * T4_UP + kh * 32 + kc * 8 + i;
      if (j < nh && k < nh) {
        if (blk == 2) v = Wmerge[(long long)(slot * nh + k) * nh + j];
        else v = (blk == 0 ? W : Wg)[(long long)(24 + slot * nh + k) * nh + j];
      }
    } else {
      r -= 4LL * NP * KH * hjob;
      part = (int)(r / xpart); r -= (long long)part * xpart;
      const int n = (int)((r / 8) % T4_N), kc = (int)(r / (8 * T4_N));     // kc = neighbour slot
      const int sub = n / T4_TPS, e = n % 16;
      blk = (n % T4_TPS) / 16;
      const int j = p * T4_UP + sub * 16 + e;
      if (j < nh && blk < 2) {
        if (i < 6) v = (blk == 0 ? W : Wg)[(long long)(kc * 6 + i) * nh + j];
        else if (i == 6 && kc == 0) v = (blk == 0 ? b : bg)[j];
      }
    }
    if (split) {
      v *= TC_SPLIT_WS;
      const __half hi = __float2half_rn(v);
      out[idx] = part == 0 ? hi : __float2half_rn(v - __half2float(hi));
    } else {
      out[idx] = __float2half_rn(v * (blk == 1 ? TC_SCALE_GATE : 1.0f));
    }
  }
}

}  // namespace rnnwf
