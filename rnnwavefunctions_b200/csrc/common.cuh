// Shared device/host structures of the lattice-sweep engine.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rnnwf {

// number of kernels this library has launched (bench.py reports it as gpu_launches)
inline long long& launch_counter() { static long long c = 0; return c; }
#define RNNWF_COUNT_LAUNCH() (++::rnnwf::launch_counter())

constexpr int KC = 8;          // k-chunk depth of one pipeline stage
constexpr int TM = 4;          // rows per thread (one LDS.128 of the A operand)
constexpr int RL = 8;          // row lanes per warp
constexpr int CL = 4;          // column lanes per warp
constexpr int MAX_SLOTS = 4;   // neighbour slots per step
constexpr int HL = 8;          // lanes cooperating on one row in the head phase
constexpr int MAX_THREADS = 640;

enum Cell { CELL_GRU1D = 0, CELL_VANILLA2D = 1, CELL_TENS2D = 2, CELL_TENSGRU2D = 3, CELL_GRU4 = 4 };
enum Mode { MODE_SAMPLE = 0, MODE_EVAL = 1, MODE_CHILD = 2 };

// One (pseudo-)step of the sweep.  For the stacked 1D GRU a lattice site expands into num_layers
// pseudo-steps; for the 2D models one step = one lattice site / unit cell in zig-zag order.
struct __align__(16) StepMeta {
  int16_t nbr[MAX_SLOTS];   // history index of the hidden state feeding slot s, -1 = zeros
  int16_t nsym[MAX_SLOTS];  // symbol index whose one-hot feeds slot s, -1 = zeros
  int16_t sym;              // symbol produced (sampling) / consumed (evaluation) by this step's head, -1 = none
  int16_t site;             // flat offset of that symbol inside a configuration row
  uint8_t variant;          // index into the VariantMeta table
  uint8_t head;             // 1: the output heads are evaluated after this step
  uint8_t own;              // slot holding this unit's own previous state (GRU1D blend), 0xff = none
  uint8_t pad0;
  int32_t pad1, pad2;
};
static_assert(sizeof(StepMeta) == 32, "StepMeta layout");

struct __align__(16) VariantMeta {
  uint32_t w_off;        // float4 offset of the variant's packed GEMM weights (column block 0)
  uint32_t w_cb_stride;  // float4 stride between column blocks
  uint32_t x_off;        // float4 offset of the one-hot input table  [slot][D][Gp]
  uint32_t b_off;        // float4 offset of the bias table           [Gp]
  uint8_t ncls;          // weight classes (tensorized cells: one per (left, up) input pair)
  uint8_t slot_mask;     // slots that carry GEMM weights
  uint8_t x_mask;        // slots that carry one-hot table rows
  uint8_t pad;
  uint8_t cls_mult[MAX_SLOTS];   // class = sum_s cls_mult[s] * symbol(nsym[s])
  uint8_t slot_rank[MAX_SLOTS];  // position of slot s among the packed slots
  uint32_t pad2;
};
static_assert(sizeof(VariantMeta) == 32, "VariantMeta layout");

// One connected configuration behind a local energy: the parent sample with one or two symbols toggled.
struct __align__(16) Child {
  int32_t parent;   // row of the parent sample
  int16_t ta, tb;   // symbol indices to toggle (tb = -1: single flip)
  float coef;       // matrix element <sigma'|H|sigma>
  uint8_t ka, kb;   // bit inside the symbol (Rydberg: atom of the unit cell), 0 for spins
  uint16_t pad;
};
static_assert(sizeof(Child) == 16, "Child layout");

struct SweepParams {
  // ---- model (constant per handle)
  int cell, nh, nhp, ldH, nslots, D, ND, complex_head, u1, cap_up, cap_dn;
  int Q, Tsym, symL, nsites, bits, cps;
  int Gblk, ncb, Gp, Greal, ntu;
  int ring_len, WR, WC, Mp, Mr, ncls_max;
  const StepMeta* steps;
  const VariantMeta* variants;
  const float4* Wpack;
  const float4* Xtab;
  const float4* bias;
  const float* head_w;   // [nhp][ND]
  const float* head_b;   // [ND]
  // ---- batch
  int mode;
  long long rows;
  const int* rows_dev;      // optional: row count produced on the device (variable connected-config lists)
  const int8_t* cfg_in;     // EVAL: configurations; CHILD: parent samples
  int8_t* cfg_out;          // SAMPLE: generated configurations
  const float* uniforms;    // SAMPLE: [rows][Tsym] or null -> Philox
  unsigned long long seed, row_offset;
  double* out_logp;
  double* out_phase;
  // ---- CHILD mode
  const Child* children;
  const double* par_logp;
  const double* par_phase;
  double* eloc;             // [parents][2]
  const float* par_hist;    // saved history of the parents (prefix reuse) or null
  const double2* par_cum;   // saved running (logp, phase) of the parents
  long long par_rows;       // row stride of the parent's saved arrays
  // ---- saved state written in SAMPLE / EVAL mode (any may be null)
  float* sv_hist;           // [Q][sv_rows][nhp]
  float4* sv_gates;         // [Q][sv_rows][Gp]
  double2* sv_cum;          // [Tsym][sv_rows]
  float4* sv_head;          // tensor-core regime: heads' outputs (p0, p1, phase pre-activation of the symbol, 0) [Q][sv_rows]
  float* sv_himg;           // tensor-core regime: hidden states as UMMA tile images per (site, 128-row tile) (dw_tc.cuh)
  long long sv_rows;
  // ---- tensor-core regime (sweep_tc.cuh)
  const unsigned char* tc_B;   // [3 variants][tc_np passes][weight image]
  const float* tc_bias;        // [2][nhp]: b, bg
  int tc_np;
  unsigned long long* tc_prof;  // optional per-CTA cycle counters (debug)
  int tile_rows;                      // tensor-core sweeps, sampling / evaluation: live rows per tile (<= 128; rows spread over all SMs)
  int debug_flags;                    // tensor-core sweep timing experiments (RNNWF_TC_DEBUG), 0 in production
  long long sched_ns;          // CHILD mode of single-flip Hamiltonians: samples per (step, bit) group of the regular child list
  int sched_bits;              //   (child index (t bits + k) ns + n), 0 = no regular structure
  const float* g1_pack;        // dedicated 1D-GRU kernels (gru1d.cuh): parameter image copied to shared memory
  // ---- scratch
  float* ring;              // [gridDim.x][ring_len][Mp][nhp]
};

// Shared-memory carve-up, computed identically on host and device.
struct SmemLayout {
  int off_W, off_A, off_H, off_headw, off_headb, off_sym, off_logp, off_phase, off_nup, off_perm, off_permid,
      off_qcls, off_qzero, off_rcls, off_rpos, off_cnt, off_parent, off_coef, total;
  __host__ __device__ static int align16(int x) { return (x + 15) & ~15; }
  __host__ __device__ SmemLayout(int ncls_max, int Gblk, int Mp, int ldH, int nhp, int ND, int Tsym) {
    int o = 0;
    off_W = o;      o += 2 * ncls_max * KC * Gblk * 16;
    off_A = o;      o += align16(2 * KC * (Mp + 4) * 4);
    off_H = o;      o += align16(Mp * ldH * 4);
    off_headw = o;  o += align16(nhp * ND * 4);
    off_headb = o;  o += align16(ND * 4);
    off_logp = o;   o += align16(Mp * 8);
    off_phase = o;  o += align16(Mp * 8);
    off_nup = o;    o += align16(Mp * 4);
    off_parent = o; o += align16(Mp * 4);
    off_coef = o;   o += align16(Mp * 4);
    off_rpos = o;   o += align16(Mp * 4);
    off_cnt = o;    o += 64;
    off_perm = o;   o += align16(Mp * 2);
    off_permid = o; o += align16(Mp * 2);
    off_qcls = o;   o += align16(Mp / 4);
    off_qzero = o;  o += align16(Mp / 4);
    off_rcls = o;   o += align16(Mp);
    off_sym = o;    o += align16(Tsym * Mp);
    total = o;
  }
};

// Philox4x32-10 (Salmon et al. 2011), counter (c0,c1,c2,c3), key (k0,k1); returns word 0..3.
__host__ __device__ inline void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                                              uint32_t k1, uint32_t out[4]) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
    uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0, hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += W0; k1 += W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

// uniform in [0,1) with 24 random bits, float32-exact: u[b,t] = (philox(b_lo, b_hi, t, 0; seed)[0] >> 8) * 2^-24
__host__ __device__ inline float philox_uniform(unsigned long long seed, unsigned long long row, uint32_t t) {
  uint32_t o[4];
  philox4x32_10((uint32_t)row, (uint32_t)(row >> 32), t, 0u, (uint32_t)seed, (uint32_t)(seed >> 32), o);
  return (float)(o[0] >> 8) * (1.0f / 16777216.0f);
}

}  // namespace rnnwf
