// The tensor-core sweep kernel itself (see sweep_tc.cuh for the design notes, constants and PTX wrappers).  Compiled in its
// own translation unit (sweep_tc_inst.cu) so that its template instances build in parallel with the rest of the library.
#pragma once
#include "sweep_tc.cuh"

namespace rnnwf {

// BIASMMA: the biases ride on the tensor core (num_units < padded width): position `num_units` of the horizontal
// neighbour half of every operand is a constant 1 and that K row of the weight images holds the biases.
// CELLK: 0 = tensorized GRU (7 accumulator blocks per unit, 16 units per column pass); 2 = un-gated tensorized cell
// (2DHeisenberg/MDTensorizedcell.py: tanh(T:W + b)) = the same with u == 1 and zero gate / merge images; 1 = vanilla 2D RNN
// (2DTFIM_2DRNN/MDRNNcell.py:21-35: elu(x_h Uh + h_h Wh + x_v Uv + h_v Wv + b)): one pass per site whose 7 "blocks" are
// the 7 unit pieces a thread owns (unit = 16 blk + 4 sub + e); the neighbours are separable, so one weight image serves
// every site, and the one-hot inputs / bias are added in the epilogue from small shared-memory tables.
// SPLIT: split-precision regime (1e-5 parity on the tensor cores).  Hidden states and weights are each carried as a pair of
// FP16 numbers hi + lo (hi = fp16(x), lo = fp16(x - hi): 22 mantissa bits together) and every K=16 step is three MMAs into
// the same FP32 accumulator, A_hi B_hi + A_lo B_hi + A_hi B_lo (the lo x lo term is below FP32 rounding).  The operand is
// single-buffered in TMEM ([hi 8 hc | lo 8 hc] columns; two buffers of it would not fit beside the accumulators): the
// epilogue writes the next site's operand from the history ring once the last column pass's MMAs have completed.  Each
// column pass streams two weight images (hi, lo) through two stages.  Gates, softmax and logarithms use the accurate
// FP32 / FP64 functions of the FFMA regime (no MUFU approximations).
template <int MODE, bool BIASMMA, int CELLK, bool SPLIT>
__global__ void __launch_bounds__(TC_THREADS, 1) sweep_tc_kernel(const SweepParams P) {
  extern __shared__ __align__(128) unsigned char tc_smem[];   // no swizzle anywhere: 16-byte alignment is all the hardware needs
  unsigned char* const smem = tc_smem;
  const TcSmem L(P.nhp, P.Tsym, P.Q);
  unsigned char* const Bst = smem + L.off_B;                              // [TC_STAGES][b_bytes] weight stages
  float* const headW = reinterpret_cast<float*>(smem + L.off_headw);      // [nhp + 16][4], zero rows past num_units
  float* const headB = reinterpret_cast<float*>(smem + L.off_headb);
  float* const biasS = reinterpret_cast<float*>(smem + L.off_bias);       // [2][nhp + 16]: b, bg (zero past num_units)
  float4* const partS = reinterpret_cast<float4*>(smem + L.off_part);     // [2][128][TC_SUB] partial head sums
  short4* const metaS = reinterpret_cast<short4*>(smem + L.off_meta);     // per step: (nbr0, nbr1, site, -)
  int4* const stepT = reinterpret_cast<int4*>(smem + L.off_step);         // per step: see the set-up loop below
  uint32_t* const symS = reinterpret_cast<uint32_t*>(smem + L.off_sym);   // [Tsym][4]: bit (row & 31) of word row >> 5
  uint64_t* const bars = reinterpret_cast<uint64_t*>(smem + L.off_bar);
  uint32_t* const tmem_slot = reinterpret_cast<uint32_t*>(smem + L.off_tmem);
  uint64_t* const bar_A = bars + 0;                   // next operand written to TMEM (512 arrivals)
  uint64_t* const bar_mma = bars + 1;                 // [2] accumulator buffer full (tcgen05.commit)
  uint64_t* const bar_free = bars + 3;                // [2] accumulator buffer drained (512 arrivals)
  uint64_t* const bar_part = bars + 5;                // [2] partial head sums of a step written (512 arrivals)
  uint64_t* const bar_pfree = bars + 7;               // [2] partial head sums consumed (64 arrivals)
  uint64_t* const bar_sym = bars + 9;                 // sampling: the step's symbols are drawn (64 arrivals)
  uint64_t* const bar_tile = bars + 10;               // head warps are done with the tile's symbol table (64 arrivals)
  uint64_t* const bar_B = bars + 11;                  // [TC_STAGES] weight stage landed (complete_tx)
  uint64_t* const bar_Bfree = bars + 11 + TC_STAGES;  // [TC_STAGES] weight stage consumed (tcgen05.commit)
  // [2] older-neighbour half of a site's operand written to TMEM (512 arrivals), alternating with the CTA's running site
  // count: the arrival for site n + 1 comes at the START of site n's epilogue, i.e. possibly before an MMA warp has looked at
  // site n's phase -- on one barrier that warp could then find itself two phases behind; on two, a barrier's next arrivals
  // (site n + 2) only start after site n + 1's epilogue has begun, which needs every MMA of site n
  uint64_t* const bar_Aold = bars + 11 + 2 * TC_STAGES;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nhp = P.nhp, nh = P.nh, ldb = nhp + TC_JP;
  const int NP = P.tc_np, hc = L.kc / 2;              // column passes per site; k-cores (= K16 steps) per neighbour half
  const uint32_t B_LBO = 128, SBO = (uint32_t)L.kc * 128u;      // n-core-major weight image: k-cores adjacent, n-cores kc * 128 bytes apart
  const uint32_t idesc = tc_idesc(TC_ROWS, TC_N);
  // The last column pass holds only num_units - 16 (NP - 1) real units: issue it just wide enough for the epilogue
  // threads that own them (28 columns each, rounded up to the MMA's N granularity of 16).
  const int subs_last = (nh - (NP - 1) * TC_JP + 3) / 4;
  const uint32_t idesc_last = (CELLK == 1) ? idesc : tc_idesc(TC_ROWS, ((subs_last * TC_TPS + 15) / 16) * 16);
  // TMEM map: [0, 224) two accumulator buffers; then two A operands of 8*hc columns each (2 halves per 32-bit column)
  const uint32_t acols = 8u * (uint32_t)hc;           // columns of one A operand ([h_h ; h_v], 16*hc halves)
  const uint32_t NQ = (uint32_t)tc_ring_quads(P.nhp);              // quads (2 unit pieces each) a thread owns in a history image
  const uint32_t SLOT1 = tc_ring_slot(P.nhp);                      // one history image
  const uint32_t RSLOT = SPLIT ? 2u * SLOT1 : SLOT1;               // one step's history image(s): hi [, lo]
  constexpr uint32_t NPARTS = SPLIT ? 2u : 1u;        // weight images per column pass
  // split precision: operands are stored scaled by powers of two (TC_SPLIT_HS, TC_SPLIT_WS) so that their lo parts stay in
  // FP16's normal range; the accumulators are scaled back (exactly) as they are read
  constexpr float HS = SPLIT ? TC_SPLIT_HS : 1.0f, DS = SPLIT ? 1.0f / (TC_SPLIT_HS * TC_SPLIT_WS) : 1.0f;
  // Early K steps.  The operand of a site is [older-neighbour half | this sweep's previous output ("carry")] in either order;
  // the older half is in TMEM long before the carry half is complete, and with 13 K steps over 2 x 52 columns six of them
  // touch one half only.  The first pass of each MMA warp issues those six as soon as the older half has landed -- under the
  // tail of the previous site's epilogue -- and the other seven once the carry half is there: the site-to-site bubble of the
  // tensor pipe shrinks by ~2 x 6/13 of a pass.  (Double-buffered operand only: not in the split-precision regime.)
  const bool pre = !SPLIT && CELLK != 1 && hc == 13 && !(P.debug_flags & 32);

  if (tid == 0) {
    mbar_init(bar_A, TC_EPI_THREADS);
    mbar_init(bar_Aold, TC_EPI_THREADS); mbar_init(bar_Aold + 1, TC_EPI_THREADS);
    for (int i = 0; i < 2; ++i) {
      mbar_init(bar_mma + i, 1);
      mbar_init(bar_free + i, TC_EPI_THREADS);
      mbar_init(bar_part + i, TC_EPI_THREADS);
      mbar_init(bar_pfree + i, TC_HEAD_THREADS);
    }
    mbar_init(bar_sym, TC_HEAD_THREADS);
    mbar_init(bar_tile, TC_HEAD_THREADS);
    for (int i = 0; i < TC_STAGES; ++i) { mbar_init(bar_B + i, 1); mbar_init(bar_Bfree + i, 1); }
    fence_mbar_init();
  }
  if (warp == 0) tc_alloc(tmem_slot, TC_TMEM_COLS);
  for (int i = tid; i < (nhp + TC_JP) * 4; i += TC_THREADS) {     // packed heads are [unit][ND], ND = 2 (real) or 4 (complex)
    const int j = i >> 2, d = i & 3;
    headW[i] = (j < nh && d < P.ND) ? P.head_w[j * P.ND + d] : 0.f;
  }
  if (tid < 4) headB[tid] = (tid < P.ND) ? P.head_b[tid] : 0.f;
  for (int i = tid; i < (CELLK == 1 ? 6 : 2) * ldb; i += TC_THREADS) {
    const int which = i / ldb, j = i - which * ldb;
    const float sc = (CELLK == 1 || SPLIT) ? 1.0f : (which == 0 ? TC_SCALE_CAND : TC_SCALE_GATE);
    biasS[i] = (j < nh) ? P.tc_bias[which * nhp + j] * sc : 0.f;
  }
  for (int i = tid; i < P.Q; i += TC_THREADS) {
    const StepMeta* m = P.steps + i;
    metaS[i] = make_short4(m->nbr[0], m->nbr[1], m->site, 0);
    // operand of step i: the half that holds the step right before it ("carry", written as that step is computed) and
    // the other neighbour (copied from the ring); ring slots as byte offsets
    const int nb0 = m->nbr[0], nb1 = m->nbr[1];
    const int carry = (nb0 == i - 1) ? 0 : 1;
    const int other = (carry == 0) ? nb1 : nb0;
    // Ring slots by live range instead of step number (assigned once on the host, build_model: StepMeta::pad1): an image is
    // read back by the step below it in the same lattice column, and that load is issued before the step in between stores
    // (top of iteration t - 1 in the split regime, end of t - 2 otherwise) -- one slot per lattice column on the zigzag path,
    // half the ring footprint (C4: 157 -> 86 MB over all CTAs, inside the L2).  RNNWF_TC_DEBUG bit 64: slot = step mod ring_len.
    const bool modslots = (P.debug_flags & 64) != 0;
    const uint32_t sl_own = modslots ? (uint32_t)(i % P.ring_len) : (uint32_t)m->pad1;
    const uint32_t sl_other = other < 0 ? 0u : (modslots ? (uint32_t)(other % P.ring_len) : (uint32_t)P.steps[other].pad1);
    stepT[i] = make_int4((nb0 & 0xffff) | (nb1 << 16), (other & 0xffff) | (carry << 16), (int)(sl_own * RSLOT), (int)(sl_other * RSLOT));
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_A0 = tmem_base + 2u * TC_N;     // first A operand buffer

  // optional cycle accounting (-DRNNWF_TC_PROF builds only; tools/tc_cycles.py): MMA warp [0] bar_A [1] bar_B
  // [2] bar_free [3] mma issue [5] total; loader warp [4] stage wait [6] copy issue; epilogue thread 32: [8] bar_mma
  // [9] pass compute [10] next-operand half [11] step tail [12] total; head thread: [13] wait for partial sums [14] heads
#ifdef RNNWF_TC_PROF
  long long prof[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) prof[i] = 0;
  const bool do_prof = (P.tc_prof != nullptr) && (tid == 0 || tid == 32 || tid == 32 + TC_EPI_THREADS || tid == TC_THREADS - 32);
  const long long kstart = tc_clock();
#define TC_T0() long long t0__ = do_prof ? tc_clock() : 0
#define TC_ACC(k) do { if (do_prof) { const long long t1__ = tc_clock(); prof[k] += t1__ - t0__; t0__ = t1__; } } while (0)
#define TC_TOTAL(k) do { if (do_prof) prof[k] = tc_clock() - kstart; } while (0)
  // event trace of CTA 0: [role 8][TC_TRACE_LEN] words (tag << 40 | cycles since kernel start), after the counters
  constexpr int TC_TRACE_LEN = 1024;
  const int trole = (warp == 0) ? 0 : (warp == TC_THREADS / 32 - 1) ? 1 : (warp == 1) ? 2 : (warp == 16) ? 3 : (warp == 8) ? 4 : (warp == 17) ? 5 : -1;
  const bool do_trace = (P.tc_prof != nullptr) && blockIdx.x == 0 && lane == 0 && trole >= 0;
  unsigned long long* const tbuf = do_trace ? P.tc_prof + (size_t)gridDim.x * 16 + (size_t)trole * TC_TRACE_LEN : nullptr;
  int nev = 0;
#define TC_EV(tag) do { if (do_trace && nev < TC_TRACE_LEN) tbuf[nev++] = ((unsigned long long)(tag) << 40) | (unsigned long long)(tc_clock() - kstart); } while (0)
#else
#define TC_EV(tag) do { } while (0)
#define TC_T0() do { } while (0)
#define TC_ACC(k) do { } while (0)
#define TC_TOTAL(k) do { } while (0)
#endif

  const long long total_rows = (P.rows_dev != nullptr) ? (long long)(*P.rows_dev) : P.rows;
  // Rows per tile.  The connected-configuration sweep always runs full 128-row tiles (it has thousands of them).  Sampling
  // and evaluation batches are a few dozen tiles, fewer than SMs, and every site is a latency chain per tile: there the
  // host spreads the rows over all SMs (P.tile_rows <= 128 live rows per tile; the MMAs stay 128 rows high).  The saved
  // state keeps its 128-row tile-blocked layouts: a thread addresses them through its global row (srow_of / stile_of), and
  // the tiles cover the batch rounded up to 128 rows so that the images' rows past the batch are written (as zeros).
  const int VR = (MODE == MODE_CHILD) ? TC_ROWS : P.tile_rows;
  const long long rows_cover = (MODE == MODE_CHILD) ? total_rows : ((total_rows + TC_ROWS - 1) / TC_ROWS) * TC_ROWS;
  const long long ntiles = (rows_cover + VR - 1) / VR;
  // first sweep step of a tile: under prefix reuse the tile restarts where its first child departs from its parent
  auto tile_q0 = [&](long long row0) -> int {
    int ts0 = 0;
    if (MODE == MODE_CHILD && P.par_hist != nullptr) {
      const Child c0 = P.children[row0];
      ts0 = (c0.tb >= 0 && c0.tb < c0.ta) ? c0.tb : c0.ta;
    }
    return ts0;
  };
  // Every site goes through the tensor core, the first one included (its operand is all zeros): no special cases.

  if (warp == 0 || warp == TC_MMA2_WARP) {
    // ======================= two MMA warps: one per accumulator buffer =======================
    // A warp's instructions after its tcgen05.mma's (even a barrier poll) only return once those MMAs have executed,
    // so a single issuer cannot check the next pass's barriers while its MMAs are in flight and the tensor core idles
    // between passes.  Two issuers alternate: warp 0 takes the even column passes (accumulator buffer 0), warp 18 the
    // odd ones (buffer 1); each is only ever blocked behind its own pass, and the other one's MMAs queue up behind.
    // Warp-uniform code (descriptors stay in uniform registers); the single-issue instructions are predicated on one
    // elected lane; the K loop is fully unrolled (every descriptor is base + immediate).
    const int me = (warp == 0) ? 0 : 1;
    const bool leader = elect_one();
    const uint64_t bdesc0 = tc_desc(smem_u32(Bst), B_LBO, SBO);
    const uint32_t barA_a = smem_u32(bar_A), barB_a = smem_u32(bar_B), barF_a = smem_u32(bar_free + me);
    const uint32_t tmem_d = tmem_base + (uint32_t)me * (uint32_t)TC_N;
    uint32_t phA = 0, nsite = 0, nuse = 0, it0 = 0;   // operand parity; sites so far; uses of this warp's buffer; passes before this site
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int q0 = tile_q0(tile * VR) * P.symL;
      for (int q = q0; q < P.Q; ++q, it0 += (uint32_t)NP, ++nsite) {
        TC_T0();
        const uint32_t tmem_a = SPLIT ? tmem_A0 : tmem_A0 + (uint32_t)(q & 1) * acols;
        for (int p = me; p < NP; p += 2) {
          const uint32_t it = (it0 + (uint32_t)p) * NPARTS, s = it % TC_STAGES;   // first weight image of the pass
          // weights and accumulator buffer first (ready long before); the site's operand is the critical one: last
          mbar_wait_a(barB_a + s * 8u, (it / TC_STAGES) & 1u);   // the stage's (it / stages)-th use, by either warp
          if (SPLIT) mbar_wait_a(barB_a + (s + 1u) * 8u, (it / TC_STAGES) & 1u);   // (it is even: the lo image sits in the next stage)
          TC_ACC(1); TC_EV(2);
          mbar_wait_a(barF_a, (nuse & 1u) ^ 1u); ++nuse;
          TC_ACC(2); TC_EV(3);
          const uint64_t bd0 = bdesc0 + (uint64_t)((s * (uint32_t)L.b_bytes) >> 4);
          const uint32_t id = (p == NP - 1) ? idesc_last : idesc;
          const bool split_k = pre && p == me;        // this warp's first pass of the site: early K steps on the older half
          const int ch = stepT[q].y >> 16;             // which half of this site's operand is the carry
          if (split_k) {
            mbar_wait_a(smem_u32(bar_Aold + (nsite & 1u)), (nsite >> 1) & 1u);
            tc_fence_after();
            if (leader) {
              if (ch == 0) tc_issue_range<7, 13>(tmem_d, tmem_a, bd0, id, 0u);   // older half = columns 52..103: K steps 7..12
              else tc_issue_range<0, 6>(tmem_d, tmem_a, bd0, id, 0u);            // older half = columns 0..51: K steps 0..5
            }
            __syncwarp();
          }
          if (p == me) {
            mbar_wait_a(barA_a, phA); phA ^= 1u;
            TC_ACC(0); TC_EV(1);
          }
          tc_fence_after();
          if (leader) {
            if (split_k) {
              if (ch == 0) tc_issue_range<0, 7>(tmem_d, tmem_a, bd0, id, 1u);
              else tc_issue_range<6, 13>(tmem_d, tmem_a, bd0, id, 1u);
            }
            else if (hc == 13) tc_issue_pass<13>(tmem_d, tmem_a, bd0, id);
            else tc_issue_pass<0>(tmem_d, tmem_a, bd0, id, hc);
            if (SPLIT) {
              const uint64_t bd1 = bdesc0 + (uint64_t)(((s + 1u) * (uint32_t)L.b_bytes) >> 4);
              if (hc == 13) { tc_issue_pass<13>(tmem_d, tmem_a + acols, bd0, id, 13, 1u); tc_issue_pass<13>(tmem_d, tmem_a, bd1, id, 13, 1u); }
              else { tc_issue_pass<0>(tmem_d, tmem_a + acols, bd0, id, hc, 1u); tc_issue_pass<0>(tmem_d, tmem_a, bd1, id, hc, 1u); }
            }
            tc_commit(bar_mma + me);
            tc_commit(bar_Bfree + s);
            if (SPLIT) tc_commit(bar_Bfree + s + 1u);
          }
          __syncwarp();
          TC_ACC(3); TC_EV(4);
        }
      }
    }
    if (me == 0) { TC_TOTAL(5); }
  } else if (warp == TC_THREADS / 32 - 1) {
    // ======================= loader warp: weight stream =======================
    // One flat sequence of (site, pass) weight images through TC_STAGES shared-memory stages; runs ahead of the MMA
    // warp across sites and tiles, gated only by the stages' "consumed" barriers.
    const bool leader = elect_one();
    // debug_flags bit 0 (RNNWF_TC_DEBUG=1, timing experiments only, wrong results): copy half of every weight image
    const uint32_t load_bytes = (P.debug_flags & 1) ? (uint32_t)L.b_bytes / 2 : (uint32_t)L.b_bytes;
    // the last column pass is issued only as wide as its real units: its image is the first n-cores of the stage
    const uint32_t last_bytes = (CELLK == 1 || (P.debug_flags & 1)) ? load_bytes : (uint32_t)L.kc * 128u * (uint32_t)(((subs_last * TC_TPS + 15) / 16) * 2);
    uint32_t phBf = 0, n = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const int q0 = tile_q0(tile * VR) * P.symL;
      for (int q = q0; q < P.Q; ++q) {
        const short4 sm = metaS[q];
        const int v = (CELLK == 1) ? 0 : ((sm.x >= 0 && sm.y < 0) ? 1 : ((sm.x < 0 && sm.y >= 0) ? 2 : 0));   // first site: any image (operand is zero)
        const unsigned char* src = P.tc_B + (size_t)v * NP * NPARTS * L.b_bytes;
        for (int p = 0; p < NP * (int)NPARTS; ++p, ++n) {      // SPLIT: (pass, hi), (pass, lo), ...
          const uint32_t s = n % TC_STAGES;
          TC_T0();
          if (n >= TC_STAGES) { mbar_wait_relaxed(bar_Bfree + s, (phBf >> s) & 1u); phBf ^= 1u << s; }
          TC_ACC(4); TC_EV(5);
          if (leader) {
            const uint32_t nb = ((p / (int)NPARTS) == NP - 1) ? last_bytes : load_bytes;
            mbar_expect_tx(bar_B + s, nb);
            bulk_g2s(Bst + (size_t)s * L.b_bytes, src + (size_t)p * L.b_bytes, nb, bar_B + s);
          }
          __syncwarp();
          TC_ACC(6); TC_EV(6);
        }
      }
    }
  } else if (warp <= TC_EPI_THREADS / 32) {
    // ======================= epilogue warps: 4 threads per row =======================
    // Thread (row, sub) owns units 16 p + 4 sub .. + 3 of every column pass p: it reads those accumulators (28
    // contiguous TMEM columns), applies the gates, and writes the 4 new hidden values (FP16) to the next operand in
    // TMEM and to the history ring.  The K axis of the operand is ordered thread-major (tc_kpos), so everything a
    // thread owns of a neighbour half is one contiguous run of TMEM columns and one 64-byte chunk of the ring image:
    // only this thread ever reads those bytes back, so the ring needs no barrier.  The partial head sums go to the head
    // warps through shared memory; no block-wide barrier in the step loop.
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int sub = (warp - 1) >> 2;                    // which 4 of the 16 units of a pass this thread handles
    const int row = (quarter << 5) + lane;              // tile row == TMEM lane
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const int np_own = tc_kcnt(nhp, sub) / 4;           // column passes in which this thread owns units (7 or 6 at 104)
    const uint32_t hcols = (uint32_t)nhp / 2;           // TMEM columns of one neighbour half
    // constant-1 position (unit index num_units, a padding unit) of the horizontal half: owned by one sub of each row
    const bool one_mine = BIASMMA && sub == ((nh & 15) >> 2);
    const int one_p = nh >> 4;                          // its column pass == piece index
    const uint32_t one_bits = (SPLIT ? 0x5400u : 0x3C00u) << (16 * (nh & 1));   // FP16 1.0 (split precision: TC_SPLIT_HS = 64.0) in the low / high half of the word
    const int one_word = 2 * one_p + ((nh & 3) >> 1);                   // word inside the thread's run
    unsigned char* const ring_base = reinterpret_cast<unsigned char*>(P.ring);
    // this thread's first quad inside a step's ring image   (the whole ring is < 4 GB: 32-bit offsets)
    const uint32_t ring_off0 = (uint32_t)blockIdx.x * (uint32_t)P.ring_len * RSLOT + (uint32_t)sub * (NQ * TC_RING_QUAD) + (uint32_t)row * 16u;
    const uint32_t bar_mma_a = smem_u32(bar_mma), bar_free_a = smem_u32(bar_free);
    const uint32_t bar_part_a = smem_u32(bar_part), bar_pfree_a = smem_u32(bar_pfree);
    const uint32_t acc0 = tmem_base + lane_base + (uint32_t)(sub * TC_TPS);               // this thread's accumulators, buffer 0
    const uint32_t aop0 = tmem_A0 + lane_base + (uint32_t)tc_koff(nhp, sub) / 2;          // its operand columns, buffer 0, half 0
    const float* const bS = biasS + sub * 4;
    const float* const hW = headW + sub * 16;
    uint32_t phM = 0;                                   // parity bits of the two accumulator-full barriers
    uint32_t hs = 0, ntile = 0;                         // steps / tiles processed so far by this CTA
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++ntile) {
      const long long row0 = tile * VR;
      const int nrows = (int)((total_rows - row0 < (long long)VR) ? (total_rows - row0 > 0 ? total_rows - row0 : 0) : (long long)VR);
      const int q0 = tile_q0(row0) * P.symL;
      const bool live = row < nrows;
      const bool stores = row < VR && row0 + row < rows_cover;       // this thread owns a row of the saved images (zeros past the batch)
      const long long sv_tiles = ((long long)P.sv_rows + TC_ROWS - 1) / TC_ROWS;
      const long long stile = (row0 + row) / TC_ROWS;                 // position of the row in the 128-row tile-blocked saved layouts
      const int srow = (int)((row0 + row) % TC_ROWS);
      int rparent = 0;
      // ---- symbol table of the tile (teacher forcing); the head warps must be done with the previous tile's
      if (ntile > 0) mbar_wait(bar_tile, (ntile - 1) & 1);
      {
        Child ch;
        ch.parent = 0; ch.ta = 0; ch.tb = -1; ch.coef = 0.f; ch.ka = ch.kb = 0;
        if (live && MODE == MODE_CHILD) { ch = P.children[row0 + row]; rparent = ch.parent; }
        const long long src = (MODE == MODE_CHILD) ? (long long)ch.parent : row0 + row;
        const int8_t* c = P.cfg_in + src * P.nsites;
        for (int t = sub; t < P.Tsym; t += TC_SUB) {        // warp-uniform trip count: one ballot word per (site, 32 rows)
          int sy = 0;
          if (live && MODE != MODE_SAMPLE) {
            sy = c[metaS[t].z] & 1;
            if (MODE == MODE_CHILD && (t == ch.ta || t == ch.tb)) sy ^= 1;
          }
          const uint32_t w = __ballot_sync(0xffffffffu, sy);
          if (lane == 0) symS[t * 4 + quarter] = w;
        }
        quarter_barrier(quarter);
      }
      // Everything this thread owns of one neighbour half: the hidden state of step `other` (zeros if absent) as 14 (12)
      // packed FP16 words, from the ring image or, before the tile's first step under prefix reuse, from the parent's
      // saved FP32 history.  All branches are warp-uniform.
      // (SPLIT: `wl` receives the lo parts -- second ring image of the step, or the residual of the FP32 history)
      auto load_half = [&](uint32_t (&w)[16], uint32_t (&wl)[16], int other, uint32_t other_slot) {
        if (other < 0) {
#pragma unroll
          for (int i = 0; i < 16; ++i) { w[i] = 0u; if (SPLIT) wl[i] = 0u; }
        } else if (other >= q0 || P.par_hist == nullptr) {
          const unsigned char* src = ring_base + (ring_off0 + other_slot);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint4 v = make_uint4(0u, 0u, 0u, 0u), u = make_uint4(0u, 0u, 0u, 0u);
            if ((uint32_t)i < NQ) {          // (narrow models keep fewer quads per image)
              v = __ldcg(reinterpret_cast<const uint4*>(src + i * TC_RING_QUAD));
              if (SPLIT) u = __ldcg(reinterpret_cast<const uint4*>(src + SLOT1 + i * TC_RING_QUAD));
            }
            w[4 * i] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
            if (SPLIT) { wl[4 * i] = u.x; wl[4 * i + 1] = u.y; wl[4 * i + 2] = u.z; wl[4 * i + 3] = u.w; }
          }
        } else {
          const float* psrc = P.par_hist + ((size_t)other * (size_t)P.par_rows + (size_t)rparent) * nhp + sub * 4;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            w[2 * i] = 0u; w[2 * i + 1] = 0u;
            if (SPLIT) { wl[2 * i] = 0u; wl[2 * i + 1] = 0u; }
            if (i < np_own) {
              float4 v = __ldg(reinterpret_cast<const float4*>(psrc + i * TC_JP));
              const int j = i * TC_JP + sub * 4;                   // padding units must be exact zeros in the operand
              if (j + 0 >= nh) v.x = 0.f;
              if (j + 1 >= nh) v.y = 0.f;
              if (j + 2 >= nh) v.z = 0.f;
              if (j + 3 >= nh) v.w = 0.f;
              if (SPLIT) { split_f16x2(v.x * HS, v.y * HS, w[2 * i], wl[2 * i]); split_f16x2(v.z * HS, v.w * HS, w[2 * i + 1], wl[2 * i + 1]); }
              else { w[2 * i] = pack_f16x2(v.x, v.y); w[2 * i + 1] = pack_f16x2(v.z, v.w); }
            }
          }
        }
      };
      // TMEM lane == row; the thread's run of 2 np_own columns inside the half that starts at column a_half.  All 32
      // lanes execute the (warp-collective) tcgen05.st.
      auto store_half = [&](const uint32_t (&w)[16], uint32_t a_half) {
        __syncwarp();
        // 2 np_own columns, decomposed into x8 / x4 / x2 stores (register indices must be static)
        if (np_own & 4) tc_st8(a_half, w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7]);
        if (np_own & 2) {
          if (np_own & 4) tc_st4(a_half + 8u, make_uint4(w[8], w[9], w[10], w[11]));
          else tc_st4(a_half, make_uint4(w[0], w[1], w[2], w[3]));
        }
        if (np_own & 1) {
          switch (np_own & 6) {
            case 0: tc_st2(a_half, make_uint2(w[0], w[1])); break;
            case 2: tc_st2(a_half + 4u, make_uint2(w[4], w[5])); break;
            case 4: tc_st2(a_half + 8u, make_uint2(w[8], w[9])); break;
            default: tc_st2(a_half + 12u, make_uint2(w[12], w[13])); break;
          }
        }
      };

      auto set_one = [&](uint32_t (&w)[16]) {
        if (one_mine) {
#pragma unroll
          for (int i = 0; i < 14; ++i) if (i == one_word) w[i] |= one_bits;
        }
      };

      // operand of the tile's first step: both halves from the history (parent's saved history under prefix reuse;
      // zeros for the very first site), then prefetch the older-neighbour half of the second step's operand
      uint32_t onext[16], onlo[16];
      // both halves of a step's operand from the history (SPLIT: every step; otherwise the tile's first step only)
      auto operand_from_history = [&](int nb0, int nb1, uint32_t a0) {
        load_half(onext, onlo, nb0, nb0 >= 0 ? (uint32_t)stepT[nb0].z : 0u);
        set_one(onext);
        store_half(onext, a0);
        if (SPLIT) store_half(onlo, a0 + acols);
        load_half(onext, onlo, nb1, nb1 >= 0 ? (uint32_t)stepT[nb1].z : 0u);
        store_half(onext, a0 + hcols);
        if (SPLIT) store_half(onlo, a0 + acols + hcols);
        tc_wait_st();
        tc_fence_before();
        if (pre) mbar_arrive(bar_Aold + (hs & 1u));        // (hs: the running site count this tile's first site will have)
        mbar_arrive(bar_A);
      };
      {
        const int4 e0 = stepT[q0];
        const int nb0 = (int)(short)(e0.x & 0xffff), nb1 = (int)(short)((unsigned)e0.x >> 16);
        operand_from_history(nb0, nb1, SPLIT ? aop0 : aop0 + (uint32_t)(q0 & 1) * acols);
        if (!SPLIT && q0 + 1 < P.Q) { const int4 e1 = stepT[q0 + 1]; load_half(onext, onlo, (int)(short)(e1.y & 0xffff), (uint32_t)e1.w); }
      }

      for (int q = q0; q < P.Q; ++q, ++hs) {
        // The operand of step q+1 is produced here: its older-neighbour half was prefetched at the end of the previous
        // step and goes to TMEM first; the half that is this step's own output is written pass by pass.
        const int4 eq = stepT[q];
        const int nb0 = (int)(short)(eq.x & 0xffff), nb1 = (int)(short)((unsigned)eq.x >> 16);
        TC_T0();
        const bool has_next = (q + 1 < P.Q);
        const int carry_half = has_next ? (stepT[q + 1].y >> 16) : 0;
        const uint32_t a_next = SPLIT ? aop0 : aop0 + (uint32_t)((q + 1) & 1) * acols;
        if (has_next && !SPLIT) {
          if (carry_half == 1) set_one(onext);          // the prefetched half is the horizontal one
          store_half(onext, a_next + (uint32_t)(1 - carry_half) * hcols);
          if (pre) { tc_wait_st(); tc_fence_before(); mbar_arrive(bar_Aold + ((hs + 1u) & 1u)); }   // the MMA warps may start on this half
        }
        // split precision: the operand buffer is still being read by this site's MMAs; fetch the older-neighbour half of the
        // next operand now (its latency hides behind the column passes) and collect this site's pieces in `chi` / `clo`
        uint32_t chi[16], clo[16];
        if (SPLIT) {
#pragma unroll
          for (int i = 0; i < 16; ++i) { chi[i] = 0u; clo[i] = 0u; }
          if (has_next) { const int4 e1 = stepT[q + 1]; load_half(onext, onlo, (int)(short)(e1.y & 0xffff), (uint32_t)e1.w); }
        }
        TC_ACC(10); TC_EV(11);
        // sampling: the neighbours' symbols are drawn by the head warps (completion number hs covers steps < hs)
        if (MODE == MODE_SAMPLE && hs > 0) mbar_wait(bar_sym, (hs - 1) & 1);
        const float cl = (nb0 >= 0) ? (float)tc_sym(symS, nb0, row) : 0.f;
        const float cu = (nb1 >= 0) ? (float)tc_sym(symS, nb1, row) : 0.f;
        float lg[4] = {0.f, 0.f, 0.f, 0.f};
        uint2 o_one = make_uint2(0u, 0u);
        if constexpr (CELLK == 1) {
          // ---- vanilla 2D RNN: one 112-column pass, accumulator block blk = this thread's unit piece blk
          unsigned char* ring_p = ring_base + (ring_off0 + (uint32_t)eq.z);
          const bool save = (MODE != MODE_CHILD) && live && (P.sv_hist != nullptr);
          float* sv_h = save ? P.sv_hist + ((size_t)q * (size_t)P.sv_rows + (size_t)(row0 + row)) * nhp + sub * 4 : nullptr;
          uint32_t a_carry = a_next + (uint32_t)carry_half * hcols;
          // hidden-state tile image for the weight-gradient kernel (zeros for padding units / rows past the batch)
          const bool himg = MODE != MODE_CHILD && P.sv_himg != nullptr && stores;
          const int hnc = ((nhp + 15) / 16) * 2;
          float* const sv_i = himg ? P.sv_himg + ((size_t)q * (size_t)sv_tiles + (size_t)stile) * (size_t)(32 * hnc * 32) + (size_t)(srow >> 2) * hnc * 32 + (srow & 3) : nullptr;
          // b' = b + Uh[0] (horizontal neighbour present) + Uv[0] (vertical present); dUh = Uh[1] - Uh[0], dUv likewise
          const float* bv = biasS + (((nb0 >= 0) ? 1 : 0) + ((nb1 >= 0) ? 2 : 0)) * ldb + sub * 4;
          const float* duh = biasS + 4 * ldb + sub * 4;
          const float* duv = biasS + 5 * ldb + sub * 4;
          mbar_wait_a(bar_mma_a, phM & 1u); phM ^= 1u;
          TC_ACC(8); TC_EV(12);
          uint32_t d[TC_TPS];
          __syncwarp();
          tc_fence_after();
          tc_ld28(acc0, d);
          tc_wait_ld();
          tc_fence_before();
          mbar_arrive_a(bar_free_a);
          TC_EV(13);
#pragma unroll
          for (int blk = 0; blk < TC_NB; ++blk) {
            if (blk < np_own) {                                     // warp-uniform
              const float4 b4 = *reinterpret_cast<const float4*>(bv + blk * TC_JP);
              const float4 h4 = *reinterpret_cast<const float4*>(duh + blk * TC_JP);
              const float4 v4 = *reinterpret_cast<const float4*>(duv + blk * TC_JP);
              const float bb[4] = {b4.x, b4.y, b4.z, b4.w}, hh[4] = {h4.x, h4.y, h4.z, h4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
              float hnew[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float x = fmaf(cu, vv[e], fmaf(cl, hh[e], fmaf(__uint_as_float(d[blk * 4 + e]), DS, bb[e])));
                const float hv = (x > 0.f) ? x : (SPLIT ? (expf(x) - 1.0f) : (__expf(x) - 1.0f));   // elu (absolute error of exp(x) - 1 <= 1.2e-7: what FP32 h carries anyway); padding units: x = 0 -> 0
                hnew[e] = live ? hv : 0.f;
                const float4 w = *reinterpret_cast<const float4*>(hW + blk * (TC_JP * 4) + e * 4);
                lg[0] = fmaf(hv, w.x, lg[0]); lg[1] = fmaf(hv, w.y, lg[1]); lg[2] = fmaf(hv, w.z, lg[2]); lg[3] = fmaf(hv, w.w, lg[3]);
              }
              uint2 o, ol = make_uint2(0u, 0u);
              if (SPLIT) { split_f16x2(hnew[0] * HS, hnew[1] * HS, o.x, ol.x); split_f16x2(hnew[2] * HS, hnew[3] * HS, o.y, ol.y); }
              else { o.x = pack_f16x2(hnew[0], hnew[1]); o.y = pack_f16x2(hnew[2], hnew[3]); }
              if (save) *reinterpret_cast<float4*>(sv_h + blk * TC_JP) = make_float4(hnew[0], hnew[1], hnew[2], hnew[3]);
              if (himg) {   // hidden-state tile image for the weight-gradient kernel, straight from the registers
                const int j0 = blk * TC_JP + sub * 4;
                if (j0 < hnc * 8) {
#pragma unroll
                  for (int e = 0; e < 4; ++e)
                    sv_i[((j0 + e) >> 3) * 32 + ((j0 + e) & 7) * 4] = (P.sv_hist != nullptr && j0 + e < nhp) ? hnew[e] : 0.f;
                }
              }
              __syncwarp();
              if (has_next && !SPLIT) tc_st2(a_carry + (uint32_t)(2 * blk), o);
              *reinterpret_cast<uint2*>(ring_p) = o;
              if (SPLIT) { *reinterpret_cast<uint2*>(ring_p + SLOT1) = ol; chi[2 * blk] = o.x; chi[2 * blk + 1] = o.y; clo[2 * blk] = ol.x; clo[2 * blk + 1] = ol.y; }
              ring_p += (blk & 1) ? (TC_RING_QUAD - 8u) : 8u;
            }
          }
          if (himg) {
            // (zeros for this thread's unit pieces past the operand's K range that the image still covers)
#pragma unroll
            for (int blk = 0; blk < TC_NB; ++blk) {
              const int j0 = blk * TC_JP + sub * 4;
              if (blk >= np_own && j0 < hnc * 8) {
#pragma unroll
                for (int e = 0; e < 4; ++e) sv_i[((j0 + e) >> 3) * 32 + ((j0 + e) & 7) * 4] = 0.f;
              }
            }
          }
          TC_ACC(9);
        } else {
          unsigned char* ring_p = ring_base + (ring_off0 + (uint32_t)eq.z);   // this thread's pieces of this step's image, in pass order
          const uint32_t acc1 = acc0 + (uint32_t)TC_N;
          const int p_act = (sub < subs_last) ? NP : NP - 1;                  // passes in which this thread owns real units
          const bool save = (MODE != MODE_CHILD) && live && (P.sv_hist != nullptr);
          float* sv_h = save ? P.sv_hist + ((size_t)q * (size_t)P.sv_rows + (size_t)(row0 + row)) * nhp + sub * 4 : nullptr;
          // saved gates, tile-blocked for the reverse sweep (bwd_tc.cuh: tc_gate_index): one float4 (4 units) per quantity
          float4* sv_g = save ? P.sv_gates + ((((size_t)q * (size_t)sv_tiles + (size_t)stile) * TC_SUB + sub) * NP) * 3 * TC_ROWS + srow : nullptr;
          uint32_t a_carry = a_next + (uint32_t)carry_half * hcols;
          // hidden-state tile image for the weight-gradient kernel (dw_tc.cuh: [row / 4][unit / 8][unit % 8][row % 4]);
          // rows past the end of the batch and padding units are written as zeros
          const int hnc = ((nhp + 15) / 16) * 2;
          float* sv_i = (MODE != MODE_CHILD && P.sv_himg != nullptr && stores)
              ? P.sv_himg + ((size_t)q * (size_t)sv_tiles + (size_t)stile) * (size_t)(32 * hnc * 32) + (size_t)(srow >> 2) * hnc * 32 + (srow & 3) : nullptr;
          const float* bSp = bS;
          const float* hWp = hW;
  #pragma unroll 1
          for (int p = 0; p < NP; ++p) {
            const uint32_t s = (uint32_t)p & 1u;
            const bool active = p < p_act;                              // warp-uniform
            mbar_wait_a(bar_mma_a + s * 8u, (phM >> s) & 1u); phM ^= 1u << s;
            TC_ACC(8); TC_EV(12);
            uint2 o = make_uint2(0u, 0u), ol = make_uint2(0u, 0u);
            float hsave[4] = {0.f, 0.f, 0.f, 0.f};
            if (active) {
              uint32_t d[TC_TPS];
              __syncwarp();
              tc_fence_after();
              tc_ld28(s ? acc1 : acc0, d);
              tc_wait_ld();
              tc_fence_before();
              mbar_arrive_a(bar_free_a + s * 8u);
              TC_EV(13);
              float bsv[4] = {0.f, 0.f, 0.f, 0.f}, bgv[4] = {0.f, 0.f, 0.f, 0.f};
              if (!BIASMMA) {
                const float4 bs = *reinterpret_cast<const float4*>(bSp);
                const float4 bg = *reinterpret_cast<const float4*>(bSp + ldb);
                bsv[0] = bs.x; bsv[1] = bs.y; bsv[2] = bs.z; bsv[3] = bs.w;
                bgv[0] = bg.x; bgv[1] = bg.y; bgv[2] = bg.z; bgv[3] = bg.w;
              }
              float hnew[4], gu[4], gt[4], gm[4];
  #pragma unroll
              for (int e = 0; e < 4; ++e) {
                // blocks: 0..2 = Wa, Wb, Wc of the candidate; 3..5 of the gate; 6 = Wmerge
                float sv = fmaf(cu, __uint_as_float(d[8 + e]), fmaf(cl, __uint_as_float(d[4 + e]), __uint_as_float(d[e])));
                float gv = fmaf(cu, __uint_as_float(d[20 + e]), fmaf(cl, __uint_as_float(d[16 + e]), __uint_as_float(d[12 + e])));
                float mv = __uint_as_float(d[24 + e]);
                if (SPLIT) { sv *= DS; gv *= DS; mv *= DS; }
                if (!BIASMMA) { sv += bsv[e]; gv += bgv[e]; }
                const float u = (CELLK == 2) ? 1.0f : (SPLIT ? sigmoid_split(gv) : sigmoid_prescaled(gv));   // CELLK 2: un-gated tensorized cell, h = tanh(s)
                const float ht = SPLIT ? tanh_split(sv) : tanh_prescaled(sv);
                // u ht + (1 - u) m.  Padding units (>= num_units) have zero weights and biases: u = 1/2, ht = m = 0 -> 0.
                const float hv = fmaf(u, ht - mv, mv);
                hnew[e] = hv;
                const float4 w = *reinterpret_cast<const float4*>(hWp + e * 4);
                lg[0] = fmaf(hv, w.x, lg[0]); lg[1] = fmaf(hv, w.y, lg[1]); lg[2] = fmaf(hv, w.z, lg[2]); lg[3] = fmaf(hv, w.w, lg[3]);
                gu[e] = u; gt[e] = ht; gm[e] = mv;
              }
              if (live) {
                hsave[0] = hnew[0]; hsave[1] = hnew[1]; hsave[2] = hnew[2]; hsave[3] = hnew[3];
                if (SPLIT) { split_f16x2(hnew[0] * HS, hnew[1] * HS, o.x, ol.x); split_f16x2(hnew[2] * HS, hnew[3] * HS, o.y, ol.y); }
                else { o.x = pack_f16x2(hnew[0], hnew[1]); o.y = pack_f16x2(hnew[2], hnew[3]); }
                if (save && !(P.debug_flags & 8)) {
                  sv_g[0] = make_float4(gu[0], gu[1], gu[2], gu[3]);
                  sv_g[TC_ROWS] = make_float4(gt[0], gt[1], gt[2], gt[3]);
                  sv_g[2 * TC_ROWS] = make_float4(gm[0], gm[1], gm[2], gm[3]);
                }
              }
            } else {
              mbar_arrive_a(bar_free_a + s * 8u);   // nothing to read: its columns of the narrow last pass are padding
            }
            // saved history (FP32, row-major): padding units are written too (as zeros) -- prefix reuse reads whole pieces
            if (save && p < np_own && !(P.debug_flags & 4)) *reinterpret_cast<float4*>(sv_h) = make_float4(hsave[0], hsave[1], hsave[2], hsave[3]);
            if (p < np_own) {
              // 4 halves = 2 TMEM columns of the next operand's carry half; the same 8 bytes go to the ring image
              // (zeros for padding units and for rows past the end of the batch)
              __syncwarp();
              if (has_next && !SPLIT) tc_st2(a_carry, o);
              *reinterpret_cast<uint2*>(ring_p) = o;
              if (SPLIT) { *reinterpret_cast<uint2*>(ring_p + SLOT1) = ol; chi[2 * p] = o.x; chi[2 * p + 1] = o.y; clo[2 * p] = ol.x; clo[2 * p + 1] = ol.y; }
              if (p == one_p) o_one = o;                    // (kept for the constant-1 store after the loop)
            }
            ring_p += (p & 1) ? (TC_RING_QUAD - 8u) : 8u;
            if (sv_i != nullptr && !(P.debug_flags & 2)) {
              const int j0 = p * TC_JP + sub * 4;
              if (j0 < hnc * 8) {
  #pragma unroll
                for (int e = 0; e < 4; ++e)   // (unit position nh, the first padding unit: a constant 1 -> the bias gradients fall out of dw_tc_kernel's MMAs)
                  sv_i[((j0 + e) >> 3) * 32 + ((j0 + e) & 7) * 4] = (CELLK != 1 && j0 + e == nh && live) ? 1.0f : hsave[e];
              }
            }
            a_carry += 2u; bSp += TC_JP; hWp += TC_JP * 4; sv_h += TC_JP; sv_g += 3 * TC_ROWS;
            TC_ACC(9);
          }
        }
        TC_EV(14);
        __syncwarp();
        if (SPLIT) {
          // split-precision regime: all MMAs of this site have completed (the last pass's accumulators were read), so the
          // single operand buffer can be rewritten: this site's pieces (carry half) and the prefetched older-neighbour half
          if (has_next) {
            if (carry_half == 0) set_one(chi); else set_one(onext);       // the horizontal half carries the constant 1 of the bias row
            store_half(chi, aop0 + (uint32_t)carry_half * hcols);
            store_half(clo, aop0 + acols + (uint32_t)carry_half * hcols);
            store_half(onext, aop0 + (uint32_t)(1 - carry_half) * hcols);
            store_half(onlo, aop0 + acols + (uint32_t)(1 - carry_half) * hcols);
            tc_wait_st();
            tc_fence_before();
            mbar_arrive(bar_A);
          }
        } else if (one_mine && has_next && carry_half == 0) {
          // the TMEM copy of the horizontal half carries the constant 1 of the bias row: rewrite that piece (warp-uniform)
          if (nh & 2) o_one.y |= one_bits; else o_one.x |= one_bits;
          tc_st2(a_next + (uint32_t)(2 * one_p), o_one);
        }
        if (has_next && !SPLIT) {               // next step's operand complete: release the tensor core
          tc_wait_st();
          tc_fence_before();
          mbar_arrive(bar_A);
        }
        TC_EV(15);
        // partial head sums of this step -> head warps (buffer hs & 1; wait until the step before last was consumed)
        const uint32_t pb = hs & 1u;
        mbar_wait_a(bar_pfree_a + pb * 8u, ((hs >> 1) & 1u) ^ 1u);
        partS[(pb * TC_ROWS + row) * TC_SUB + sub] = make_float4(lg[0], lg[1], lg[2], lg[3]);
        mbar_arrive_a(bar_part_a + pb * 8u);
        // prefetch the older-neighbour half of the operand after next (its latency hides behind the wait for the next
        // step's first accumulators)
        if (!SPLIT && q + 2 < P.Q) { const int4 e2 = stepT[q + 2]; load_half(onext, onlo, (int)(short)(e2.y & 0xffff), (uint32_t)e2.w); }
        TC_ACC(11); TC_EV(16);
      }  // steps
    }  // tiles
    TC_TOTAL(12);
  } else if (warp == TC_EPI_THREADS / 32 + 1) {
    // ======================= head warp: four rows per thread =======================
    // Softmax / U(1) mask / draw or teacher forcing / running log-probability and phase of the rows, off the
    // tensor-core critical path: they consume the partial head sums one step behind the epilogue warps.  Transcendental
    // math in FP32 (this regime's tolerance is 1e-3), running sums in FP64.
    const int hrow = tid - 32 - TC_EPI_THREADS;          // rows hrow + 32 k
    uint32_t hs = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long long row0 = tile * VR;
      const int nrows = (int)((total_rows - row0 < (long long)VR) ? (total_rows - row0 > 0 ? total_rows - row0 : 0) : (long long)VR);
      const int ts0 = tile_q0(row0), q0 = ts0 * P.symL;
      double rlogp[TC_HEAD_ROWS], rphase[TC_HEAD_ROWS];
      int rnup[TC_HEAD_ROWS], rparent[TC_HEAD_ROWS];
      float rcoef[TC_HEAD_ROWS];
      bool live[TC_HEAD_ROWS];
#pragma unroll
      for (int k = 0; k < TC_HEAD_ROWS; ++k) {
        const int row = hrow + k * TC_HEAD_THREADS;
        rlogp[k] = 0.0; rphase[k] = 0.0; rnup[k] = 0; rparent[k] = 0; rcoef[k] = 0.f;
        live[k] = row < nrows;
        if (live[k] && MODE == MODE_CHILD) { const Child ch = P.children[row0 + row]; rparent[k] = ch.parent; rcoef[k] = ch.coef; }
      }
      for (int q = q0; q < P.Q; ++q, ++hs) {
        const int t = q;
        const uint32_t pb = hs & 1u;
        TC_T0();
        mbar_wait_relaxed(bar_part + pb, (hs >> 1) & 1u);
        TC_ACC(13); TC_EV(20);
        float l4[TC_HEAD_ROWS][4];
#pragma unroll
        for (int k = 0; k < TC_HEAD_ROWS; ++k) {
          const int row = hrow + k * TC_HEAD_THREADS;
          if (q == q0 && MODE == MODE_CHILD)            // the tile's symbol table is complete once its first step arrived
            for (int tt = 0; tt < ts0; ++tt) rnup[k] += tc_sym(symS, tt, row);
          l4[k][0] = headB[0]; l4[k][1] = headB[1]; l4[k][2] = headB[2]; l4[k][3] = headB[3];
#pragma unroll
          for (int i = 0; i < TC_SUB; ++i) {
            const float4 v = partS[(pb * TC_ROWS + row) * TC_SUB + i];
            l4[k][0] += v.x; l4[k][1] += v.y; l4[k][2] += v.z; l4[k][3] += v.w;
          }
        }
        mbar_arrive(bar_pfree + pb);
#pragma unroll
        for (int k = 0; k < TC_HEAD_ROWS; ++k) {
          const int row = hrow + k * TC_HEAD_THREADS;
          int sy = 0;
          if (live[k]) {
          float p0 = l4[k][0], p1 = l4[k][1];
          const float mx = fmaxf(p0, p1);
          if (SPLIT) { p0 = expf(p0 - mx); p1 = expf(p1 - mx); } else { p0 = __expf(p0 - mx); p1 = __expf(p1 - mx); }
          const float ssum = p0 + p1;
          if (SPLIT) { p0 = p0 / ssum; p1 = p1 / ssum; } else { p0 = __fdividef(p0, ssum); p1 = __fdividef(p1, ssum); }
          if (P.u1) {
            const int ndn = t - rnup[k];
            p0 *= (P.cap_dn - ndn >= 0) ? 1.f : 0.f;
            p1 *= (P.cap_up - rnup[k] >= 0) ? 1.f : 0.f;
            const float nrm = p0 + p1;
            if (SPLIT) { p0 = p0 / nrm; p1 = p1 / nrm; } else { p0 = __fdividef(p0, nrm); p1 = __fdividef(p1, nrm); }
          }
          if (MODE == MODE_SAMPLE) {
            const float u = (P.uniforms != nullptr)
                                ? P.uniforms[(size_t)(row0 + row) * P.Tsym + t]
                                : philox_uniform(P.seed, P.row_offset + (unsigned long long)(row0 + row), (uint32_t)t);
            const double tot = (double)p0 + (double)p1;
            const double target = (double)u * tot;
            sy = ((double)p0 <= target) ? 1 : 0;   // # of CDF entries <= target, clamped to the last class
          } else {
            sy = tc_sym(symS, t, row);
          }
          rlogp[k] += SPLIT ? log((double)(sy ? p1 : p0)) : (double)__logf(sy ? p1 : p0);
          if (P.complex_head) {
            const float xs = sy ? l4[k][3] : l4[k][2];
            if (SPLIT) rphase[k] += 3.14159265358979323846 * ((double)xs / (1.0 + fabs((double)xs)));
            else rphase[k] += (double)(3.14159265358979323846f * __fdividef(xs, 1.0f + fabsf(xs)));
          }
          rnup[k] += sy;
          if (P.sv_cum != nullptr && MODE != MODE_CHILD) {
            P.sv_cum[(size_t)t * (size_t)P.sv_rows + (size_t)(row0 + row)] = make_double2(rlogp[k], rphase[k]);
            // the heads' outputs, for the reverse sweep (bwd_tc.cuh forms d(logits) from them)
            P.sv_head[(size_t)t * (size_t)P.sv_rows + (size_t)(row0 + row)] = make_float4(p0, p1, P.complex_head ? (sy ? l4[k][3] : l4[k][2]) : 0.f, 0.f);
          }
          }
          if (MODE == MODE_SAMPLE) {      // drawn symbols of rows 32 k .. 32 k + 31: one ballot word (the head warp's lane == row & 31)
            const uint32_t w = __ballot_sync(0xffffffffu, sy);
            if (lane == 0) symS[t * 4 + k] = w;
          }
        }
        if (MODE == MODE_SAMPLE) mbar_arrive(bar_sym);
        TC_ACC(14); TC_EV(21);
      }  // steps

      // ---- tile results
#pragma unroll
      for (int k = 0; k < TC_HEAD_ROWS; ++k) {
        const int row = hrow + k * TC_HEAD_THREADS;
        if (!live[k]) continue;
        if (MODE == MODE_CHILD) {
          const int par = rparent[k];
          double base_lp = P.par_logp[par], base_ph = (P.par_phase != nullptr) ? P.par_phase[par] : 0.0;
          if (ts0 > 0 && P.par_cum != nullptr) {
            const double2 c = P.par_cum[(size_t)(ts0 - 1) * (size_t)P.par_rows + par];
            base_lp -= c.x;
            base_ph -= c.y;
          }
          const double amp = exp(0.5 * (rlogp[k] - base_lp)) * (double)rcoef[k];
          if (P.complex_head) {
            double sn, cs;
            sincos(rphase[k] - base_ph, &sn, &cs);
            atomicAdd(P.eloc + 2 * (size_t)par, amp * cs);
            atomicAdd(P.eloc + 2 * (size_t)par + 1, amp * sn);
          } else {
            atomicAdd(P.eloc + 2 * (size_t)par, amp);
          }
        } else {
          if (P.out_logp != nullptr) P.out_logp[row0 + row] = rlogp[k];
          if (P.out_phase != nullptr) P.out_phase[row0 + row] = rphase[k];
          if (MODE == MODE_SAMPLE && P.cfg_out != nullptr) {
            int8_t* c = P.cfg_out + (size_t)(row0 + row) * P.nsites;
            for (int tt = 0; tt < P.Tsym; ++tt) c[metaS[tt].z] = (int8_t)tc_sym(symS, tt, row);
          }
        }
      }
      mbar_arrive(bar_tile);          // the epilogue warps may rewrite the symbol table
    }  // tiles
  }

#ifdef RNNWF_TC_PROF
  if (do_prof) {
    const int role = (tid == 0) ? 0 : (tid == 32 ? 1 : (tid == TC_THREADS - 32 ? 3 : 2));
    for (int i = 0; i < 16; ++i) {
      const int owner = (i == 4 || i == 6) ? 3 : (i < 8 ? 0 : (i < 13 ? 1 : 2));
      if (owner == role) P.tc_prof[(size_t)blockIdx.x * 16 + i] = (unsigned long long)prof[i];
    }
  }
#endif
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { __syncwarp(); tc_dealloc(tmem_base, TC_TMEM_COLS); }
}

}  // namespace rnnwf
