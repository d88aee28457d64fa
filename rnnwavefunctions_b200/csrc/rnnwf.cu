// librnnwf.so -- C-ABI entry points (see include/rnnwf.h) and host-side orchestration.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <cstdlib>
#include <string>
#include <vector>

#include "model.h"
#include "sweep.cuh"
#include "sweep_tc.cuh"
#include "pack_tc.cuh"
#include "sweep_tc4.cuh"
#include "pack.cuh"
#include "eloc.cuh"
#include "gemm_tc.cuh"
#include "backward.cuh"
#include "symmetry.cuh"
#include "gru1d.cuh"
#include "sweep_van2.cuh"

using namespace rnnwf;

namespace rnnwf {   // gru1d_inst.cu
cudaError_t g1_set_smem(int nh, int L, int N);
int g1_occupancy(int threads, int smem_bytes);
cudaError_t g1_sweep_launch(const SweepParams& P, int rg, int grid, int threads, cudaStream_t st);
cudaError_t g1_pack_launch(const G1PackSrc& S, int nh, int L, int N, int D, int ND, float* out, cudaStream_t st);
cudaError_t g1_bwd_set_smem(int nh, int L);
cudaError_t g1_pack_bwd_launch(const G1PackSrc& S, int nh, int L, int D, int ND, float* out, cudaStream_t st);
}
namespace rnnwf {   // sweep_tc_inst.cu
cudaError_t tc_sweep_set_smem(int cellk, bool biasmma, bool split, int smem_bytes);
cudaError_t tc_sweep_launch(const SweepParams& P, int cellk, bool biasmma, bool split, int grid, int threads, int smem_bytes, cudaStream_t st);
// sweep_van2_inst.cu
cudaError_t van2_set_smem(int nhp, int Q);
cudaError_t van2_launch(const SweepParams& P, int grid, cudaStream_t st);
cudaError_t van2_pack_launch(const float* Wh, const float* Wv, int nh, int nhp, void* out, cudaStream_t st);
// sweep_tc4_inst.cu
cudaError_t tc4_sweep_set_smem(bool split, int smem_bytes);
cudaError_t tc4_sweep_launch(const SweepParams& P, bool split, int grid, int threads, int smem_bytes, cudaStream_t st);
cudaError_t tc4_pack_launch(const float* W, const float* Wg, const float* Wmerge, const float* b, const float* bg, int nh, int NP, bool split,
                            void* out, cudaStream_t st);
}

struct rnnwf_model : rnnwf::rnnwf_model_impl {};

namespace {

thread_local std::string g_err;
unsigned long long* g_tc_prof = nullptr;   // debug: per-CTA cycle counters of the tensor-core sweep

int fail(int code, const std::string& msg) { g_err = msg; return code; }

#define CUDA_TRY(expr)                                                                             \
  do {                                                                                             \
    cudaError_t e__ = (expr);                                                                      \
    if (e__ != cudaSuccess)                                                                        \
      return fail(RNNWF_ECUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));               \
  } while (0)

inline long long align256(long long x) { return (x + 255) & ~255LL; }
inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
// the Hamiltonian tables of the local-energy front end (first four regions of ElocLayout; they live in the handle)
inline long long hamtab_bytes(int T) {
  return align256((3LL * T + 16) * 8) + align256(3LL * T * (long long)sizeof(Bond) + 64) + align256((T + 2) * 4LL) + align256(256 * (long long)sizeof(RydNb));
}

// ------------------------------------------------------------------------------------------------
// model construction
// ------------------------------------------------------------------------------------------------
void add_tensor(rnnwf_model& m, const std::string& name, std::initializer_list<long long> shape) {
  TensorInfo t;
  t.name = name;
  t.offset = m.num_params;
  t.ndim = (int)shape.size();
  int i = 0;
  for (auto s : shape) t.shape[i++] = s;
  for (; i < 4; ++i) t.shape[i] = 1;
  m.num_params += t.numel();
  m.tensors.push_back(t);
}

int step_of_2d(int nx, int ny, int Nx) { return ny * Nx + ((ny % 2 == 0) ? nx : Nx - 1 - nx); }

int build_model(rnnwf_model& m) {
  const rnnwf_config& c = m.cfg;
  m.cell = c.cell;
  m.nh = c.num_units;
  m.nhp = (m.nh + 7) & ~7;
  m.ldH = (m.nhp % 16 == 0) ? m.nhp + 8 : m.nhp;
  m.complex_head = c.complex_head ? 1 : 0;
  m.u1 = c.u1 ? 1 : 0;
  m.cps = m.nhp / KC;
  m.num_params = 0;
  m.precision = c.reserved[0];
  if (m.precision < 0 || m.precision > 2)
    return fail(RNNWF_EINVAL, "precision must be 0 (FP32 FFMA sweep), 1 (tcgen05 sweep, FP16 operands) or 2 (tcgen05 sweep, split-precision FP16 hi+lo operands)");
  const int nh = m.nh, Nx = c.nx, Ny = c.ny;
  if (nh < 1 || nh > 1024 || Nx < 1 || Ny < 1) return fail(RNNWF_EINVAL, "bad lattice / num_units");
  if ((long long)Nx * Ny * std::max(1, c.num_layers) > 30000) return fail(RNNWF_EINVAL, "lattice too large for int16 step tables");

  auto blank_step = [] {
    StepMeta s;
    std::memset(&s, 0, sizeof(s));
    for (int k = 0; k < MAX_SLOTS; ++k) { s.nbr[k] = -1; s.nsym[k] = -1; }
    s.sym = -1; s.own = 0xff;
    return s;
  };
  auto blank_variant = [] {
    HostVariant v;
    std::memset(&v, 0, sizeof(v));
    v.meta.ncls = 1;
    return v;
  };

  switch (c.cell) {
    case RNNWF_CELL_GRU1D: {
      // ny > 1: the same chain raster-scanned over an Nx x Ny lattice, site (i, j) at flat index i * Ny + j
      // (2DTFIM_1DRNN/RNNwavefunction.py:74-76,123; Training1DRNN_2DTFIM.py:27); configurations stay flat [B, Nx * Ny]
      const int L = std::max(1, c.num_layers), N = Nx * Ny;
      m.nslots = 2; m.D = 2; m.bits = 1; m.Tsym = N; m.symL = L; m.Q = N * L; m.nsites = N;
      m.NTU = 4; m.Greal = nh; m.gru_type = true;
      for (int l = 0; l < L; ++l) {
        const long long din = (l == 0) ? 2 : nh;
        const std::string p = "gru" + std::to_string(l) + ".";
        add_tensor(m, p + "Wg", {din + nh, 2LL * nh});
        add_tensor(m, p + "bg", {2LL * nh});
        add_tensor(m, p + "Wci", {din, (long long)nh});
        add_tensor(m, p + "Wch", {(long long)nh, (long long)nh});
        add_tensor(m, p + "bci", {(long long)nh});
        add_tensor(m, p + "bch", {(long long)nh});
        HostVariant v = blank_variant();
        v.vkind = l;
        v.meta.slot_mask = (l == 0) ? 1 : 3;
        v.meta.x_mask = (l == 0) ? 1 : 0;
        v.nsl = (l == 0) ? 1 : 2;
        v.slot_of_rank[0] = 0; v.slot_of_rank[1] = 1;
        v.meta.slot_rank[0] = 0; v.meta.slot_rank[1] = 1;
        m.variants.push_back(v);
      }
      for (int t = 0; t < N; ++t)
        for (int l = 0; l < L; ++l) {
          StepMeta s = blank_step();
          const int q = t * L + l;
          if (t > 0) s.nbr[0] = (int16_t)(q - L);
          if (l > 0) s.nbr[1] = (int16_t)(q - 1);
          if (l == 0 && t > 0) s.nsym[0] = (int16_t)(t - 1);
          s.sym = (l == L - 1) ? (int16_t)t : (int16_t)-1;
          s.site = (int16_t)t;
          s.head = (l == L - 1);
          s.own = 0;
          s.variant = (uint8_t)l;
          m.steps.push_back(s);
        }
      m.site_of_t.resize(N);
      for (int t = 0; t < N; ++t) m.site_of_t[t] = (int16_t)t;
      break;
    }
    case RNNWF_CELL_VANILLA2D:
    case RNNWF_CELL_TENS2D:
    case RNNWF_CELL_TENSGRU2D: {
      m.nslots = 2; m.D = 2; m.bits = 1; m.Tsym = Nx * Ny; m.symL = 1; m.Q = Nx * Ny; m.nsites = Nx * Ny;
      const bool tens = c.cell != RNNWF_CELL_VANILLA2D;
      m.gru_type = (c.cell == RNNWF_CELL_TENSGRU2D);
      m.NTU = m.gru_type ? 3 : 4;
      m.Greal = m.gru_type ? nh : (nh + 3) / 4;
      if (!tens) {
        add_tensor(m, "Wh", {(long long)nh, (long long)nh});
        add_tensor(m, "Uh", {2, (long long)nh});
        add_tensor(m, "Wv", {(long long)nh, (long long)nh});
        add_tensor(m, "Uv", {2, (long long)nh});
        add_tensor(m, "b", {(long long)nh});
        HostVariant v = blank_variant();
        v.meta.slot_mask = 3; v.meta.x_mask = 3; v.nsl = 2;
        v.slot_of_rank[0] = 0; v.slot_of_rank[1] = 1; v.meta.slot_rank[0] = 0; v.meta.slot_rank[1] = 1;
        m.variants.push_back(v);
      } else {
        add_tensor(m, "W", {(long long)nh, 2LL * nh, 4});
        add_tensor(m, "b", {(long long)nh});
        if (m.gru_type) {
          add_tensor(m, "Wg", {(long long)nh, 2LL * nh, 4});
          add_tensor(m, "bg", {(long long)nh});
          add_tensor(m, "Wmerge", {2LL * nh, (long long)nh});
        }
        for (int k = 0; k < 4; ++k) {
          HostVariant v = blank_variant();
          v.vkind = k;
          if (k == 0) { v.meta.ncls = 4; v.meta.slot_mask = 3; v.nsl = 2; v.meta.cls_mult[0] = 1; v.meta.cls_mult[1] = 2;
                        v.slot_of_rank[0] = 0; v.slot_of_rank[1] = 1; v.meta.slot_rank[0] = 0; v.meta.slot_rank[1] = 1; }
          if (k == 1) { v.meta.ncls = 2; v.meta.slot_mask = 1; v.nsl = 1; v.meta.cls_mult[0] = 1; v.slot_of_rank[0] = 0; v.meta.slot_rank[0] = 0; }
          if (k == 2) { v.meta.ncls = 2; v.meta.slot_mask = 2; v.nsl = 1; v.meta.cls_mult[1] = 1; v.slot_of_rank[0] = 1; v.meta.slot_rank[1] = 0; }
          if (k == 3) { v.meta.ncls = 1; v.meta.slot_mask = 0; v.nsl = 0; }
          m.variants.push_back(v);
        }
      }
      m.site_of_t.resize(Nx * Ny);
      for (int ny = 0; ny < Ny; ++ny)
        for (int ix = 0; ix < Nx; ++ix) {
          const int nx = (ny % 2 == 0) ? ix : Nx - 1 - ix;
          StepMeta s = blank_step();
          const int t = ny * Nx + ix;
          const int px = (ny % 2 == 0) ? nx - 1 : nx + 1;
          const bool has_h = px >= 0 && px < Nx, has_v = ny > 0;
          if (has_h) s.nbr[0] = s.nsym[0] = (int16_t)step_of_2d(px, ny, Nx);
          if (has_v) s.nbr[1] = s.nsym[1] = (int16_t)step_of_2d(nx, ny - 1, Nx);
          s.sym = (int16_t)t;
          s.site = (int16_t)(nx * Ny + ny);
          s.head = 1;
          s.variant = tens ? (uint8_t)(has_h && has_v ? 0 : (has_h ? 1 : (has_v ? 2 : 3))) : 0;
          m.steps.push_back(s);
          m.site_of_t[t] = s.site;
        }
      break;
    }
    case RNNWF_CELL_GRU4: {
      m.nslots = 4; m.D = 8; m.bits = 3; m.Tsym = Nx * Ny; m.symL = 1; m.Q = Nx * Ny; m.nsites = Nx * Ny * 3;
      m.NTU = 3; m.Greal = nh; m.gru_type = true;
      add_tensor(m, "W", {4LL * (nh + 6), (long long)nh});
      add_tensor(m, "b", {(long long)nh});
      add_tensor(m, "Wg", {4LL * (nh + 6), (long long)nh});
      add_tensor(m, "bg", {(long long)nh});
      add_tensor(m, "Wmerge", {4LL * nh, (long long)nh});
      HostVariant v = blank_variant();
      v.meta.slot_mask = 15; v.meta.x_mask = 15; v.nsl = 4;
      for (int k = 0; k < 4; ++k) { v.slot_of_rank[k] = k; v.meta.slot_rank[k] = (uint8_t)k; }
      m.variants.push_back(v);
      m.site_of_t.resize(Nx * Ny);
      for (int ny = 0; ny < Ny; ++ny)
        for (int ix = 0; ix < Nx; ++ix) {
          const int nx = (ny % 2 == 0) ? ix : Nx - 1 - ix;
          const int t = ny * Nx + ix;
          const int dir = (ny % 2 == 0) ? 1 : -1;
          int keys[4][2] = {{nx - dir, ny}, {nx, ny - 1}, {((nx + dir) % Nx + Nx) % Nx, ny}, {nx, (ny + 1) % Ny}};
          StepMeta s = blank_step();
          for (int k = 0; k < 4; ++k) {
            const int kx = keys[k][0], ky = keys[k][1];
            if (kx < 0 || kx >= Nx || ky < 0 || ky >= Ny) continue;
            const int st = step_of_2d(kx, ky, Nx);
            if (st < t) s.nbr[k] = s.nsym[k] = (int16_t)st;
          }
          s.sym = (int16_t)t;
          s.site = (int16_t)(nx * Ny + ny);
          s.head = 1;
          s.variant = 0;
          m.steps.push_back(s);
          m.site_of_t[t] = s.site;
        }
      break;
    }
    default:
      return fail(RNNWF_EINVAL, "unknown cell type");
  }
  // heads
  add_tensor(m, "dense.W", {(long long)nh, (long long)m.D});
  add_tensor(m, "dense.b", {(long long)m.D});
  if (m.complex_head) {
    if (m.D != 2) return fail(RNNWF_EINVAL, "complex head needs a two-state model");
    add_tensor(m, "dense_phase.W", {(long long)nh, 2});
    add_tensor(m, "dense_phase.b", {2});
  }
  m.ND = m.complex_head ? 4 : m.D;
  // U(1) caps (2DHeisenberg/MDComplexRNN.py:71-72): H((N +- m)//2 - 1 - n)
  const int N = m.Tsym;
  auto floordiv2 = [](int a) { return (a >= 0) ? a / 2 : -((-a + 1) / 2); };
  m.cap_up = floordiv2(N + c.magnetization) - 1;
  m.cap_dn = floordiv2(N - c.magnetization) - 1;
  if (m.u1 && m.bits != 1) return fail(RNNWF_EINVAL, "U(1) projection needs a spin-1/2 model");

  m.t_of_site.assign(m.Tsym, 0);
  for (int t = 0; t < m.Tsym; ++t) m.t_of_site[m.site_of_t[t]] = (int16_t)t;

  // ring length = 1 + largest look-back
  int look = 1;
  for (int q = 0; q < m.Q; ++q)
    for (int k = 0; k < MAX_SLOTS; ++k)
      if (m.steps[q].nbr[k] >= 0) look = std::max(look, q - m.steps[q].nbr[k]);
  m.ring_len = std::min(m.Q, look + 1);
  // History-ring slots by live range (used by the tensor-core sweeps; sweep_tc_kernel.cuh, sweep_van2.cuh).  A step's image
  // is read back as a neighbour of later steps; every such load is issued no later than the top of iteration t - 1, before
  // iteration t - 1 stores its own image, so step q may take the slot of any step whose last reader is <= q + 1: on the
  // zigzag path one slot per lattice column instead of the whole look-back window.  pad1: readers = all neighbours (generic
  // kernel); pad2: readers = the neighbour that is not the step right before (two-tile kernel: that one stays in registers).
  for (int which = 0; which < 2; ++which) {
    std::vector<int> last(m.Q, -1), owner(m.ring_len, -1), slot(m.Q, 0);
    for (int t = 1; t < m.Q; ++t) {
      if (which == 0) {
        for (int k = 0; k < MAX_SLOTS; ++k) { const int nb = m.steps[t].nbr[k]; if (nb >= 0) last[nb] = std::max(last[nb], t); }
      } else {
        const int nb = (m.steps[t].nbr[0] == t - 1) ? m.steps[t].nbr[1] : m.steps[t].nbr[0];
        if (nb >= 0) last[nb] = std::max(last[nb], t);
      }
    }
    bool ok = true;
    for (int q = 0; q < m.Q && ok; ++q) {
      int sl = 0;
      while (sl < m.ring_len && owner[sl] >= 0 && last[owner[sl]] > q + 1) ++sl;
      if (sl >= m.ring_len) { ok = false; break; }
      owner[sl] = q; slot[q] = sl;
    }
    for (int q = 0; q < m.Q; ++q) (which == 0 ? m.steps[q].pad1 : m.steps[q].pad2) = ok ? slot[q] : q % m.ring_len;
  }

  // tile shape: TJ in {4,5} for the three-gate cells, 4 otherwise; WC <= 5 warp columns
  int best_tj = 4, best_wc = 1;
  long long best_cost = -1;
  for (int tj = 4; tj <= ((m.NTU == 3) ? 5 : 4); ++tj)
    for (int wc = 1; wc <= 5; ++wc) {
      const int gblk = tj * CL * wc;
      const int ncb = cdiv(m.Greal, gblk);
      const long long cost = (long long)ncb * gblk * 100 + ncb;  // padded columns first, then fewer passes
      const long long cost2 = cost + (wc < 2 ? 1 : 0);
      if (best_cost < 0 || cost2 < best_cost) { best_cost = cost2; best_tj = tj; best_wc = wc; }
    }
  m.TJ = best_tj; m.WC = best_wc;
  m.Gblk = m.TJ * CL * m.WC;
  m.ncb = cdiv(m.Greal, m.Gblk);
  m.Gp = m.ncb * m.Gblk;
  m.ncls_max = 1;
  for (auto& v : m.variants) m.ncls_max = std::max<int>(m.ncls_max, v.meta.ncls);

  // packed-buffer offsets
  long long woff = 0, xoff = 0, boff = 0;
  for (auto& v : m.variants) {
    const long long per_cb = (long long)v.nsl * m.cps * v.meta.ncls * KC * m.Gblk;
    v.meta.w_off = (uint32_t)woff;
    v.meta.w_cb_stride = (uint32_t)per_cb;
    v.w_float4 = per_cb * m.ncb;
    woff += v.w_float4;
    v.meta.x_off = (uint32_t)xoff;
    if (v.meta.x_mask) xoff += (long long)m.nslots * m.D * m.Gp;
    v.meta.b_off = (uint32_t)boff;
    boff += m.Gp;
  }
  m.tc_enabled = (m.precision >= 1 && (m.cell == CELL_TENSGRU2D || m.cell == CELL_TENS2D || m.cell == CELL_VANILLA2D));
  m.tc_split = m.tc_enabled && m.precision == 2;
  m.tc4 = (m.precision >= 1 && m.cell == CELL_GRU4);
  m.van2 = m.tc_split && m.cell == CELL_VANILLA2D && v2_supported(m.nh, m.nhp, m.Q, m.complex_head, m.u1) && std::getenv("RNNWF_NO_VAN2") == nullptr;
  m.tc4_split = m.tc4 && m.precision == 2;
  // small 1D models in the FFMA regime: dedicated kernels with the whole weight set resident in shared memory (gru1d.cuh)
  m.g1 = (m.precision == 0 && m.cell == CELL_GRU1D && g1_supported(m.nh, m.symL, m.Tsym) && std::getenv("RNNWF_NO_G1") == nullptr);
  m.g1_bwd = m.g1 && g1_bwd_supported(m.nh, m.symL) && std::getenv("RNNWF_NO_G1_BWD") == nullptr;
  if (m.precision == 1 && !m.tc_enabled && !m.tc4)
    return fail(RNNWF_EINVAL, "the tensor-core regime is implemented for the 2D cells (tensorized GRU, tensorized, vanilla RNN, four-neighbour GRU)");
  if (m.precision == 2 && !m.tc_enabled && !m.tc4)
    return fail(RNNWF_EINVAL, "the split-precision tensor-core regime is implemented for the 2D cells (tensorized GRU, tensorized, vanilla RNN, four-neighbour GRU)");
  // column passes per site: 16 units x 7 blocks each for the tensorized GRU; the vanilla cell needs one (7 unit pieces)
  m.tc_np = (m.cell == CELL_VANILLA2D) ? 1 : (m.nh + TC_JP - 1) / TC_JP;
  if (m.tc_enabled) {
    const TcSmem ts(m.nhp, m.Tsym, m.Q);
    if (ts.total > 227 * 1024 || 2 * TC_N + 2 * m.nhp > TC_TMEM_COLS || (m.nhp + TC_JP - 1) / TC_JP > 7)
      return fail(RNNWF_EINVAL, "the tensor-core regime needs num_units <= 104 (weight stages in shared memory, operand in TMEM)");
  }
  if (m.tc4) {
    m.tc_np = (m.nh + T4_UP - 1) / T4_UP;      // column passes of 64 units; the hidden axis is padded to 64 tc_np
    if (m.tc_np > T4_MAXNP) return fail(RNNWF_EINVAL, "the tensor-core regime of the four-neighbour GRU needs num_units <= 256");
    if (T4Smem(4, m.Q).total > 227 * 1024) return fail(RNNWF_EINVAL, "the tensor-core regime of the four-neighbour GRU: lattice too large for shared memory");
    m.ring_len = m.Q;
  }
  m.wpack_float4 = std::max(1LL, woff);
  m.xtab_float4 = std::max(1LL, xoff);
  m.bias_float4 = std::max(1LL, boff);
  return RNNWF_OK;
}

// ------------------------------------------------------------------------------------------------
// sweep launch
// ------------------------------------------------------------------------------------------------
struct LaunchShape { int WR, Mp, Mr, threads, smem, grid, tile_rows; };

const void* sweep_fn(const rnnwf_model& m) {
  if (m.NTU == 3 && m.TJ == 5) return reinterpret_cast<const void*>(&sweep_kernel<5, 3>);
  if (m.NTU == 3 && m.TJ == 4) return reinterpret_cast<const void*>(&sweep_kernel<4, 3>);
  return reinterpret_cast<const void*>(&sweep_kernel<4, 4>);
}
// persistent CTAs the scratch ring is sized for, in units of 128-row tiles: up to 4 per SM when shared memory allows
int max_ctas(const rnnwf_model& m) {
  if (m.tc_enabled || m.tc4) return m.num_sms;
  SmemLayout L(m.ncls_max, m.Gblk, 128, m.ldH, m.nhp, m.ND, m.Tsym);
  const int occ = std::max(1, std::min(4, (227 * 1024) / std::max(1, L.total)));
  return m.num_sms * occ;
}

// child: the connected-configuration sweep (always full tiles)
int pick_shape(const rnnwf_model& m, long long rows, LaunchShape& s, bool child = false) {
  s.tile_rows = 128;
  if (m.g1) {
    // sampling / evaluation of a training batch: narrow tiles (16 rows) so that the batch spreads over many SMs
    // the connected configurations of a training batch (10^4 rows at C1 / C2): 32-row tiles balance better over the SMs than
    // 157 tiles of 64 (measured: E_loc 0.19 -> 0.15 ms at C1, 0.69 -> 0.47 ms with three layers); wide tiles for larger sets
    int rg = G1_RG;
    if (std::getenv("RNNWF_NO_VTILES") == nullptr) rg = child ? ((rows + 31) / 32 <= 8LL * m.num_sms ? 4 : G1_RG) : g1_pick_rg(rows, m.num_sms);
    s.tile_rows = 8 * rg;
    const long long tiles = (rows + s.tile_rows - 1) / s.tile_rows;
    s.WR = 2; s.Mp = G1_ROWS; s.Mr = G1_ROWS; s.threads = g1_threads(m.nh); s.smem = G1Layout(m.nh, m.symL, m.Tsym, s.tile_rows).total_bytes;
    s.grid = (int)std::max(1LL, std::min<long long>(tiles, (long long)m.num_sms * (rg == G1_RG ? m.g1_occ : 4)));
    return RNNWF_OK;
  }
  if (m.tc_enabled) {
    const TcSmem ts(m.nhp, m.Tsym, m.Q);
    long long tiles = (rows + TC_ROWS - 1) / TC_ROWS;
    s.WR = 4; s.Mp = TC_ROWS; s.Mr = TC_ROWS; s.threads = TC_THREADS; s.smem = ts.total;
    if (!child) {     // sampling / evaluation: fewer tiles than SMs are spread over all SMs (sweep_tc.cuh: tc_tile_rows)
      s.tile_rows = tc_tile_rows(rows, m.num_sms);
      tiles = (tiles * TC_ROWS + s.tile_rows - 1) / s.tile_rows;
    }
    s.grid = (int)std::max(1LL, std::min<long long>(tiles, m.num_sms));
    return RNNWF_OK;
  }
  if (m.tc4) {
    const long long tiles = (rows + T4_ROWS - 1) / T4_ROWS;
    s.WR = 4; s.Mp = T4_ROWS; s.Mr = T4_ROWS; s.threads = T4_THREADS; s.smem = T4Smem(t4_stages(m.Q), m.Q).total;
    s.grid = (int)std::max(1LL, std::min<long long>(tiles, m.num_sms));
    return RNNWF_OK;
  }
  const int pad = (m.ncls_max > 1) ? 3 * m.ncls_max : 0;
  for (int wr = 4; wr >= 1; wr >>= 1) {
    const int Mp = 32 * wr, Mr = Mp - pad;
    if (Mr <= 0) continue;
    SmemLayout L(m.ncls_max, m.Gblk, Mp, m.ldH, m.nhp, m.ND, m.Tsym);
    const long long tiles = (rows + Mr - 1) / Mr;
    const bool fits = L.total <= 227 * 1024;
    if (!fits) continue;
    // prefer the tallest tile that still gives every SM a tile
    if (wr == 1 || tiles >= m.num_sms) {
      s.WR = wr; s.Mp = Mp; s.Mr = Mr; s.threads = 32 * wr * m.WC; s.smem = L.total;
      int occ = 1;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, sweep_fn(m), s.threads, s.smem) != cudaSuccess) { occ = 1; cudaGetLastError(); }
      occ = std::max(1, std::min(4, occ));
      const long long cap = (long long)max_ctas(m) * 128 / Mp;     // what the scratch ring was sized for
      s.grid = (int)std::max(1LL, std::min<long long>(tiles, std::min<long long>((long long)m.num_sms * occ, cap)));
      return RNNWF_OK;
    }
  }
  return fail(RNNWF_ENOMEM, "model does not fit in shared memory (lattice too long / too many units)");
}

void fill_model_params(const rnnwf_model& m, const LaunchShape& s, SweepParams& P) {
  std::memset(&P, 0, sizeof(P));
  P.cell = m.cell; P.nh = m.nh; P.nhp = m.nhp; P.ldH = m.ldH; P.nslots = m.nslots; P.D = m.D; P.ND = m.ND;
  P.complex_head = m.complex_head; P.u1 = m.u1; P.cap_up = m.cap_up; P.cap_dn = m.cap_dn;
  P.Q = m.Q; P.Tsym = m.Tsym; P.symL = m.symL; P.nsites = m.nsites; P.bits = m.bits; P.cps = m.cps;
  P.Gblk = m.Gblk; P.ncb = m.ncb; P.Gp = m.Gp; P.Greal = m.Greal; P.ntu = m.NTU;
  P.tile_rows = s.tile_rows;
  P.ring_len = m.ring_len; P.WR = s.WR; P.WC = m.WC; P.Mp = s.Mp; P.Mr = s.Mr; P.ncls_max = m.ncls_max;
  P.steps = m.d_steps; P.variants = m.d_variants; P.Wpack = m.d_wpack; P.Xtab = m.d_xtab; P.bias = m.d_bias;
  P.head_w = m.d_head_w; P.head_b = m.d_head_b;
  P.tc_B = m.d_tcB; P.tc_bias = m.d_tcbias; P.tc_np = m.tc_np; P.tc_prof = g_tc_prof; P.g1_pack = m.d_g1pack;
  { const char* dbg = std::getenv("RNNWF_TC_DEBUG"); P.debug_flags = dbg ? std::atoi(dbg) : 0; }
}

template <int TJ, int NTU>
cudaError_t launch_inst(const SweepParams& P, const LaunchShape& s, cudaStream_t st) {
  RNNWF_COUNT_LAUNCH(); sweep_kernel<TJ, NTU><<<s.grid, s.threads, s.smem, st>>>(P);
  return cudaGetLastError();
}

// the FP32 FFMA sweep (its weight packs are kept up to date in every regime)
int launch_sweep_ffma(const rnnwf_model& m, const SweepParams& P, const LaunchShape& s, cudaStream_t st) {
  cudaError_t e;
  if (m.NTU == 3 && m.TJ == 5) e = launch_inst<5, 3>(P, s, st);
  else if (m.NTU == 3 && m.TJ == 4) e = launch_inst<4, 3>(P, s, st);
  else e = launch_inst<4, 4>(P, s, st);
  if (e != cudaSuccess) return fail(RNNWF_ECUDA, std::string("sweep kernel launch: ") + cudaGetErrorString(e));
  return RNNWF_OK;
}

int launch_sweep(const rnnwf_model& m, const SweepParams& P, const LaunchShape& s, cudaStream_t st) {
  cudaError_t e;
  if (m.g1) {
    RNNWF_COUNT_LAUNCH();
    e = g1_sweep_launch(P, s.tile_rows / 8, s.grid, s.threads, st);
    if (e != cudaSuccess) return fail(RNNWF_ECUDA, std::string("1D GRU sweep launch: ") + cudaGetErrorString(e));
    return RNNWF_OK;
  }
  if (m.van2 && P.mode == MODE_CHILD && P.rows_dev == nullptr) {
    // vanilla 2D RNN, split precision, single-flip child list: two tiles per CTA (sweep_van2.cuh).  The parents' log P comes
    // from the same kernel (sampling / evaluation instances below): one accumulation order on both sides of the ratio.
    SweepParams P2 = P;
    P2.tc_B = m.d_van2B;
    const long long tiles = (P.rows + V2_ROWS - 1) / V2_ROWS;
    RNNWF_COUNT_LAUNCH();
    e = van2_launch(P2, (int)std::max(1LL, std::min<long long>((tiles + 1) / 2, m.num_sms)), st);
    if (e != cudaSuccess) return fail(RNNWF_ECUDA, std::string("two-tile vanilla sweep launch: ") + cudaGetErrorString(e));
    return RNNWF_OK;
  }
  if (m.van2 && (P.mode == MODE_SAMPLE || P.mode == MODE_EVAL) && P.rows_dev == nullptr && std::getenv("RNNWF_NO_VAN2_SAMPLE") == nullptr) {
    // vanilla 2D RNN, split precision, sampling / evaluation (with or without saved state): the same two-tile kernel, rows spread over
    // both tile groups of every SM.  (The ring workspace was checked against s.grid CTAs: never launch more.)
    SweepParams P2 = P;
    P2.tc_B = m.d_van2B;
    const long long tiles128 = (P.rows + V2_ROWS - 1) / V2_ROWS, groups = 2LL * std::max(1, s.grid);
    long long vr = V2_ROWS;
    if (tiles128 < groups) vr = std::max<long long>(8, std::min<long long>(V2_ROWS, ((tiles128 * V2_ROWS + groups - 1) / groups + 7) / 8 * 8));
    P2.tile_rows = (int)vr;
    const long long ntiles = (tiles128 * V2_ROWS + vr - 1) / vr;
    RNNWF_COUNT_LAUNCH();
    e = van2_launch(P2, (int)std::max(1LL, std::min<long long>((ntiles + 1) / 2, s.grid)), st);
    if (e != cudaSuccess) return fail(RNNWF_ECUDA, std::string("two-tile vanilla sampling sweep launch: ") + cudaGetErrorString(e));
    return RNNWF_OK;
  }
  if (m.tc_enabled) {
    RNNWF_COUNT_LAUNCH();
    const int cellk = (m.cell == CELL_VANILLA2D) ? 1 : (m.cell == CELL_TENS2D ? 2 : 0);
    const bool bm = cellk != 1 && m.nh < m.nhp;     // biases through the tensor core when the operand has a free padding position
    e = tc_sweep_launch(P, cellk, bm, m.tc_split, s.grid, s.threads, s.smem, st);
    if (e != cudaSuccess) return fail(RNNWF_ECUDA, std::string("tensor-core sweep launch: ") + cudaGetErrorString(e));
    return RNNWF_OK;
  }
  if (m.tc4) {
    RNNWF_COUNT_LAUNCH();
    e = tc4_sweep_launch(P, m.tc4_split, s.grid, s.threads, s.smem, st);
    if (e != cudaSuccess) return fail(RNNWF_ECUDA, std::string("tensor-core sweep (four-neighbour GRU) launch: ") + cudaGetErrorString(e));
    return RNNWF_OK;
  }
  return launch_sweep_ffma(m, P, s, st);
}

long long ring_bytes(const rnnwf_model& m, int grid, int Mp) {
  if (m.tc4) return (long long)grid * m.Q * t4_step_bytes(m.tc_np, m.tc4_split);   // FP16 chunk images of every site, per persistent CTA
  if (m.van2) {      // (the two-tile sweep keeps a ring of 32-bit operand words (hi + lo halves) per tile group; the generic kernel's ring serves sampling / evaluation)
    const long long slot = (long long)tc_ring_slot(m.nhp) * 2;
    return (long long)grid * std::max((long long)m.ring_len * slot, v2_ring_bytes(m.ring_len, m.nhp));
  }
  // tensor-core regime: one FP16 history image per slot in the fixed quad layout of sweep_tc.cuh (TC_RING_SLOT bytes)
  const long long slot = m.tc_enabled ? (long long)tc_ring_slot(m.nhp) * (m.tc_split ? 2 : 1) : (long long)Mp * m.nhp * 4;
  return (long long)grid * m.ring_len * slot;
}

int set_saved(const rnnwf_model& m, SweepParams& P, void* saved, long long rows) {
  if (saved == nullptr) return RNNWF_OK;
  const Saved sv = m.saved_layout(rows);
  char* b = static_cast<char*>(saved);
  P.sv_hist = reinterpret_cast<float*>(b + sv.off_hist);
  P.sv_gates = (m.gru_type || (m.tc_enabled && m.cell == CELL_TENS2D)) ? reinterpret_cast<float4*>(b + sv.off_gates) : nullptr;
  P.sv_cum = reinterpret_cast<double2*>(b + sv.off_cum);
  P.sv_himg = m.tc_enabled ? reinterpret_cast<float*>(b + sv.off_himg) : nullptr;
  P.sv_head = m.tc_enabled ? reinterpret_cast<float4*>(b + sv.off_head) : nullptr;
  P.sv_rows = rows;
  return RNNWF_OK;
}

// single-block exclusive scan of int32 (n <= ~1e8); out[n-1] + in[n-1] is not needed: callers append a 0.
__global__ void scan_kernel(const int* __restrict__ in, int* __restrict__ out, long long n) {
  __shared__ int tot[1024];
  const int tid = threadIdx.x;
  const long long chunk = (n + 1023) / 1024;
  long long lo = tid * chunk; if (lo > n) lo = n;
  long long hi = lo + chunk; if (hi > n) hi = n;
  int s = 0;
  for (long long i = lo; i < hi; ++i) s += in[i];
  tot[tid] = s;
  __syncthreads();
  for (int off = 1; off < 1024; off <<= 1) {
    const int v = (tid >= off) ? tot[tid - off] : 0;
    __syncthreads();
    tot[tid] += v;
    __syncthreads();
  }
  int run = tot[tid] - s;
  for (long long i = lo; i < hi; ++i) { const int v = in[i]; out[i] = run; run += v; }
}

// three-launch exclusive scan for long inputs: per-block sums, scan of the (<= 1024) sums, per-block scan with offset
constexpr int SCAN_TPB = 256;
__global__ void __launch_bounds__(SCAN_TPB) scan_reduce_kernel(const int* __restrict__ in, int* __restrict__ bsum, long long n, long long chunk) {
  __shared__ int red[SCAN_TPB / 32];
  const long long lo = blockIdx.x * chunk, hi = (lo + chunk < n) ? lo + chunk : n;
  int s = 0;
  for (long long i = lo + threadIdx.x; i < hi; i += SCAN_TPB) s += in[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < SCAN_TPB / 32; ++w) t += red[w];
    bsum[blockIdx.x] = t;
  }
}
__global__ void __launch_bounds__(SCAN_TPB) scan_apply_kernel(const int* __restrict__ in, const int* __restrict__ boff, int* __restrict__ out,
                                                             long long n, long long chunk) {
  __shared__ int wsum[SCAN_TPB / 32];
  __shared__ int carry_s;
  const long long lo = blockIdx.x * chunk, hi = (lo + chunk < n) ? lo + chunk : n;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry_s = boff[blockIdx.x];
  __syncthreads();
  for (long long base = lo; base < hi; base += SCAN_TPB) {
    const long long i = base + threadIdx.x;
    const int v = (i < hi) ? in[i] : 0;
    int x = v;                                   // inclusive scan inside the warp
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    int woff = 0;
    for (int k = 0; k < w; ++k) woff += wsum[k];
    const int carry = carry_s;
    if (i < hi) out[i] = carry + woff + x - v;
    __syncthreads();
    if (threadIdx.x == SCAN_TPB - 1) carry_s = carry + woff + x;
    __syncthreads();
  }
}
// The same in ONE launch (chained scan): a block takes a ticket (so that logical block b only ever waits for blocks that
// have already started), sums its chunk, publishes the sum, adds up the sums of the blocks before it as they appear, and
// scans its chunk with that offset.  state[0] = ticket, state[1 + b] = sum of block b, state[1 + nblk + b] = its flag; zeroed
// by the caller before the launch.
__global__ void __launch_bounds__(SCAN_TPB) scan_chained_kernel(const int* __restrict__ in, int* __restrict__ out, long long n, long long chunk,
                                                               int nblk, int* __restrict__ state) {
  __shared__ int red[SCAN_TPB / 32], wsum[SCAN_TPB / 32];
  __shared__ int bid_s, carry_s;
  if (threadIdx.x == 0) bid_s = atomicAdd(state, 1);
  __syncthreads();
  const int b = bid_s;
  volatile int* sums = state + 1;
  volatile int* flags = state + 1 + nblk;
  const long long lo = b * chunk, hi = (lo + chunk < n) ? lo + chunk : n;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  // ---- own sum, published
  int s = 0;
  for (long long i = lo + threadIdx.x; i < hi; i += SCAN_TPB) s += in[i];
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) red[w] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < SCAN_TPB / 32; ++k) t += red[k];
    sums[b] = t;
    __threadfence();
    flags[b] = 1;
  }
  // ---- sum of the blocks before this one
  int pre = 0;
  for (int k = threadIdx.x; k < b; k += SCAN_TPB) {
    for (unsigned spin = 0; flags[k] == 0; ++spin) { __nanosleep(40); if (spin > (1u << 26)) { asm volatile("trap;"); } }
    __threadfence();
    pre += sums[k];
  }
  for (int o = 16; o > 0; o >>= 1) pre += __shfl_xor_sync(0xffffffffu, pre, o);
  __syncthreads();
  if (lane == 0) red[w] = pre;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int k = 0; k < SCAN_TPB / 32; ++k) t += red[k];
    carry_s = t;
  }
  __syncthreads();
  // ---- scan of the chunk with that offset
  for (long long base = lo; base < hi; base += SCAN_TPB) {
    const long long i = base + threadIdx.x;
    const int v = (i < hi) ? in[i] : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    int woff = 0;
    for (int k = 0; k < w; ++k) woff += wsum[k];
    const int carry = carry_s;
    if (i < hi) out[i] = carry + woff + x - v;
    __syncthreads();
    if (threadIdx.x == SCAN_TPB - 1) carry_s = carry + woff + x;
    __syncthreads();
  }
}
// exclusive scan of n int32 counts on stream st; tmp holds 2048 ints
void exclusive_scan(const int* in, int* out, long long n, int* tmp, cudaStream_t st) {
  if (n <= 8192) { RNNWF_COUNT_LAUNCH(); scan_kernel<<<1, 1024, 0, st>>>(in, out, n); return; }
  const int nblk = (int)std::min<long long>(1000, (n + 4095) / 4096);
  const long long chunk = (n + nblk - 1) / nblk;
  if (std::getenv("RNNWF_SCAN3") == nullptr) {       // one launch
    cudaMemsetAsync(tmp, 0, (size_t)(1 + 2 * nblk) * sizeof(int), st);
    RNNWF_COUNT_LAUNCH(); scan_chained_kernel<<<nblk, SCAN_TPB, 0, st>>>(in, out, n, chunk, nblk, tmp);
    return;
  }
  RNNWF_COUNT_LAUNCH(); scan_reduce_kernel<<<nblk, SCAN_TPB, 0, st>>>(in, tmp, n, chunk);
  RNNWF_COUNT_LAUNCH(); scan_kernel<<<1, 1024, 0, st>>>(tmp, tmp + 1024, nblk);
  RNNWF_COUNT_LAUNCH(); scan_apply_kernel<<<nblk, SCAN_TPB, 0, st>>>(in, tmp + 1024, out, n, chunk);
}

__global__ void count_children_kernel(const int* rows_dev, long long fixed, long long* out) {
  if (threadIdx.x == 0 && blockIdx.x == 0) *out = (rows_dev != nullptr) ? (long long)(*rows_dev) : fixed;
}

int check_ready(rnnwf_handle h) {
  if (h == nullptr) return fail(RNNWF_EINVAL, "null handle");
  if (!h->weights_set) return fail(RNNWF_ESTATE, "rnnwf_set_weights has not been called");
  return RNNWF_OK;
}

}  // namespace

// ------------------------------------------------------------------------------------------------
// C-ABI
// ------------------------------------------------------------------------------------------------
extern "C" {

const char* rnnwf_last_error(void) { return g_err.c_str(); }
int rnnwf_version(void) { return 1; }

int rnnwf_create(const rnnwf_config* cfg, rnnwf_handle* out) {
  if (cfg == nullptr || out == nullptr) return fail(RNNWF_EINVAL, "null argument");
  *out = nullptr;
  int dev = 0;
  CUDA_TRY(cudaGetDevice(&dev));
  rnnwf_model* m = new rnnwf_model();
  m->cfg = *cfg;
  m->device = dev;
  int sms = 0;
  cudaError_t e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (e != cudaSuccess) { delete m; return fail(RNNWF_ECUDA, std::string("no CUDA device: ") + cudaGetErrorString(e)); }
  m->num_sms = sms;
  int rc = build_model(*m);
  if (rc != RNNWF_OK) { delete m; return rc; }
  std::vector<VariantMeta> vmeta;
  for (auto& v : m->variants) vmeta.push_back(v.meta);
#define CREATE_TRY(expr)                                                                        \
  do {                                                                                          \
    cudaError_t e__ = (expr);                                                                   \
    if (e__ != cudaSuccess) {                                                                   \
      std::string msg = std::string(#expr) + ": " + cudaGetErrorString(e__);                    \
      rnnwf_destroy(m);                                                                         \
      return fail(RNNWF_ECUDA, msg);                                                            \
    }                                                                                           \
  } while (0)
  CREATE_TRY(cudaMalloc(&m->d_steps, m->steps.size() * sizeof(StepMeta)));
  CREATE_TRY(cudaMalloc(&m->d_variants, vmeta.size() * sizeof(VariantMeta)));
  CREATE_TRY(cudaMalloc(&m->d_site_of_t, m->site_of_t.size() * sizeof(int16_t)));
  CREATE_TRY(cudaMalloc(&m->d_wpack, m->wpack_float4 * 16));
  CREATE_TRY(cudaMalloc(&m->d_xtab, m->xtab_float4 * 16));
  CREATE_TRY(cudaMalloc(&m->d_bias, m->bias_float4 * 16));
  {   // (rows padded to a multiple of 64 and zero-filled: the four-neighbour tensor-core sweep reads whole 64-unit passes)
    const size_t hw_bytes = (size_t)((m->nhp + 63) / 64 * 64) * m->ND * 4;
    CREATE_TRY(cudaMalloc(&m->d_head_w, hw_bytes));
    CREATE_TRY(cudaMemset(m->d_head_w, 0, hw_bytes));
  }
  CREATE_TRY(cudaMalloc(&m->d_head_b, 64));
  CREATE_TRY(cudaMalloc(&m->d_hamtab, (size_t)hamtab_bytes(m->Tsym)));     // couplings / bonds / neighbour table
  CREATE_TRY(cudaMemcpy(m->d_steps, m->steps.data(), m->steps.size() * sizeof(StepMeta), cudaMemcpyHostToDevice));
  CREATE_TRY(cudaMemcpy(m->d_variants, vmeta.data(), vmeta.size() * sizeof(VariantMeta), cudaMemcpyHostToDevice));
  CREATE_TRY(cudaMemcpy(m->d_site_of_t, m->site_of_t.data(), m->site_of_t.size() * sizeof(int16_t), cudaMemcpyHostToDevice));
  CREATE_TRY(cudaMemset(m->d_wpack, 0, m->wpack_float4 * 16));
  CREATE_TRY(cudaMemset(m->d_xtab, 0, m->xtab_float4 * 16));
  CREATE_TRY(cudaMemset(m->d_bias, 0, m->bias_float4 * 16));
  const int max_smem = 227 * 1024;
  CREATE_TRY(cudaFuncSetAttribute(sweep_kernel<5, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  CREATE_TRY(cudaFuncSetAttribute(sweep_kernel<4, 3>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  CREATE_TRY(cudaFuncSetAttribute(sweep_kernel<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem));
  if (m->tc_enabled) {
    const TcSmem ts(m->nhp, m->Tsym, m->Q);
    if (ts.total > 227 * 1024) { rnnwf_destroy(m); return fail(RNNWF_ENOMEM, "tensor-core sweep does not fit in shared memory for this lattice / width"); }
    CREATE_TRY(cudaMalloc(&m->d_tcB, (size_t)3 * m->tc_np * ts.b_bytes * (m->tc_split ? 2 : 1)));
    CREATE_TRY(cudaMalloc(&m->d_tcbias, (size_t)6 * m->nhp * 4));
    CREATE_TRY(cudaMalloc(&m->d_wimgT, (size_t)3 * ((m->nhp + TC_JP - 1) / TC_JP) * BtSmem(m->nhp, m->Tsym, m->Q).b_bytes));
    const int cellk = (m->cell == CELL_VANILLA2D) ? 1 : (m->cell == CELL_TENS2D ? 2 : 0);
    CREATE_TRY(tc_sweep_set_smem(cellk, cellk != 1 && m->nh < m->nhp, m->tc_split, ts.total));
  }
  if (m->van2) {
    CREATE_TRY(cudaMalloc(&m->d_van2B, (size_t)v2_image_bytes(m->nhp)));
    CREATE_TRY(van2_set_smem(m->nhp, m->Q));
  }
  if (m->tc4) {
    const int t4s = T4Smem(t4_stages(m->Q), m->Q).total;
    CREATE_TRY(cudaMalloc(&m->d_tcB, (size_t)m->tc_np * (size_t)t4_pass_bytes(m->tc_np, m->tc4_split)));
    CREATE_TRY(tc4_sweep_set_smem(m->tc4_split, t4s));
  }
  if (m->g1) {
    const G1Layout Y(m->nh, m->symL, m->Tsym);
    CREATE_TRY(cudaMalloc(&m->d_g1pack, (size_t)Y.pack_floats * 4));
    CREATE_TRY(g1_set_smem(m->nh, m->symL, m->Tsym));
    m->g1_occ = std::max(1, std::min(4, g1_occupancy(g1_threads(m->nh), Y.total_bytes)));
    if (m->g1_bwd) {
      const G1BwdLayout Yb(m->nh, m->symL);
      CREATE_TRY(cudaMalloc(&m->d_g1packT, (size_t)Yb.pack_floats * 4));
      CREATE_TRY(g1_bwd_set_smem(m->nh, m->symL));
    }
  }
  CREATE_TRY(backward_init());
  if (m->tc4) {
    CREATE_TRY(backward_init_tc4(*m));
    CREATE_TRY(cudaMalloc(&m->d_dw4_jobs, 4096 * sizeof(Dw4Job)));
    CREATE_TRY(cudaMalloc(&m->d_dw4_sites, (size_t)5 * m->Q * sizeof(int)));
  }
  if (m->tc_enabled) {
    CREATE_TRY(backward_init_tc(*m));
    CREATE_TRY(cudaMalloc(&m->d_dw_jobs, 1024 * sizeof(DwJob)));
    CREATE_TRY(cudaMalloc(&m->d_dw_cls, (size_t)3 * m->Q * sizeof(int)));
  }
#undef CREATE_TRY
  *out = m;
  return RNNWF_OK;
}

int rnnwf_destroy(rnnwf_handle h) {
  if (h == nullptr) return RNNWF_OK;
  cudaFree(h->d_steps); cudaFree(h->d_variants); cudaFree(h->d_site_of_t); cudaFree(h->d_wpack);
  cudaFree(h->d_xtab); cudaFree(h->d_bias); cudaFree(h->d_head_w); cudaFree(h->d_head_b);
  cudaFree(h->d_hamtab); cudaFree(h->d_g1pack); cudaFree(h->d_g1packT); cudaFree(h->d_tcB); cudaFree(h->d_tcbias); cudaFree(h->d_wimgT); cudaFree(h->d_van2B); cudaFree(h->d_dw_jobs); cudaFree(h->d_dw_cls); cudaFree(h->d_dw4_jobs); cudaFree(h->d_dw4_sites);
  delete h;
  return RNNWF_OK;
}

int64_t rnnwf_count_params(const rnnwf_config* cfg) {
  // host-only: parameter count of a configuration (no device needed)
  if (cfg == nullptr) return fail(RNNWF_EINVAL, "null argument");
  rnnwf_model m;
  m.cfg = *cfg;
  m.num_sms = 1;
  int rc = build_model(m);
  return rc == RNNWF_OK ? m.num_params : (int64_t)rc;
}

int64_t rnnwf_num_params(rnnwf_handle h) { return h ? h->num_params : 0; }
int rnnwf_num_tensors(rnnwf_handle h) { return h ? (int)h->tensors.size() : 0; }

int rnnwf_param_info(rnnwf_handle h, int index, char* name, int name_cap, int64_t* offset, int32_t* ndim, int64_t shape[4]) {
  if (h == nullptr || index < 0 || index >= (int)h->tensors.size()) return fail(RNNWF_EINVAL, "bad tensor index");
  const TensorInfo& t = h->tensors[index];
  if (name != nullptr && name_cap > 0) { std::strncpy(name, t.name.c_str(), name_cap - 1); name[name_cap - 1] = 0; }
  if (offset) *offset = t.offset;
  if (ndim) *ndim = t.ndim;
  if (shape) for (int i = 0; i < 4; ++i) shape[i] = t.shape[i];
  return RNNWF_OK;
}

static void make_pack_ctx(const rnnwf_model& m, const HostVariant& v, const float* params, float* grads, PackCtx& c) {
  std::memset(&c, 0, sizeof(c));
  c.cell = m.cell; c.nh = m.nh; c.nhp = m.nhp; c.Gblk = m.Gblk; c.ncb = m.ncb; c.Gp = m.Gp; c.Greal = m.Greal;
  c.cps = m.cps; c.D = m.D; c.ncls = v.meta.ncls; c.nsl = v.nsl; c.vkind = v.vkind;
  for (int k = 0; k < MAX_SLOTS; ++k) c.slot_of_rank[k] = v.slot_of_rank[k];
  auto bind = [&](int slot, const std::string& name) {
    const TensorInfo* t = m.tensor(name);
    c.t[slot] = (t && params) ? params + t->offset : nullptr;
    c.gt[slot] = (t && grads) ? grads + t->offset : nullptr;
  };
  switch (m.cell) {
    case CELL_GRU1D: {
      const std::string p = "gru" + std::to_string(v.vkind) + ".";
      c.din = (v.vkind == 0) ? 2 : m.nh;
      bind(0, p + "Wg"); bind(1, p + "bg"); bind(2, p + "Wci"); bind(3, p + "Wch"); bind(4, p + "bci"); bind(5, p + "bch");
      break;
    }
    case CELL_VANILLA2D: bind(0, "Wh"); bind(1, "Uh"); bind(2, "Wv"); bind(3, "Uv"); bind(4, "b"); break;
    case CELL_TENS2D: bind(0, "W"); bind(1, "b"); break;
    case CELL_TENSGRU2D:
    case CELL_GRU4: bind(0, "W"); bind(1, "b"); bind(2, "Wg"); bind(3, "bg"); bind(4, "Wmerge"); break;
  }
}

// weight packs of the generic FFMA sweep (sweep.cuh)
static void pack_ffma(rnnwf_handle h, const float* params_dev, cudaStream_t st) {
  for (auto& v : h->variants) {
    PackCtx c;
    make_pack_ctx(*h, v, params_dev, nullptr, c);
    if (v.w_float4 > 0) { RNNWF_COUNT_LAUNCH(); pack_w_kernel<<<std::min(1024, cdiv(v.w_float4, 256)), 256, 0, st>>>(c, h->d_wpack + v.meta.w_off); }
    if (v.meta.x_mask) { RNNWF_COUNT_LAUNCH(); pack_x_kernel<<<std::min(256, cdiv((long long)h->nslots * h->D * h->Gp, 256)), 256, 0, st>>>(c, h->nslots, h->d_xtab + v.meta.x_off); }
    RNNWF_COUNT_LAUNCH(); pack_b_kernel<<<cdiv(h->Gp, 256), 256, 0, st>>>(c, h->d_bias + v.meta.b_off);
  }
  h->ffma_stale = false;
}

// every image a tensorized-cell handle derives from the parameters, in ONE launch: forward FP16 images (pack_tc.cuh), the bias
// table, the reverse sweep's transposed TF32 image (bwd_tc.cuh) and the heads (pack.cuh); the grid is cut into four parts
struct PackTcAll {
  const float *W, *Wg, *Wmerge, *b, *bg, *Wd, *bd, *Wp, *bp;
  int nh, nhp, NP, split, D, ND, nb_tc, nb_bwd;
  __half* tcB; float* tcbias; float* wimgT; float* head_w; float* head_b;
};
__global__ void __launch_bounds__(256) pack_tc_all_kernel(const PackTcAll A) {
  int b = blockIdx.x;
  if (b < A.nb_tc) { pack_tc_body(A.W, A.Wg, A.Wmerge, A.b, A.bg, A.nh, A.nhp, A.NP, A.split, A.tcB, b, A.nb_tc); return; }
  b -= A.nb_tc;
  if (b < A.nb_bwd) { pack_bwd_tc_body(A.W, A.Wg, A.Wmerge, A.nh, A.nhp, A.NP, A.wimgT, b, A.nb_bwd); return; }
  b -= A.nb_bwd;
  if (b == 0) { pack_tc_bias_body(A.b, A.bg, A.nh, A.nhp, A.tcbias, 0, 1); return; }
  pack_head_body(A.Wd, A.bd, A.Wp, A.bp, A.nh, A.nhp, A.D, A.ND, A.head_w, A.head_b, b - 1, (int)gridDim.x - A.nb_tc - A.nb_bwd - 1);
}

int rnnwf_set_weights(rnnwf_handle h, const float* params_dev, void* stream) {
  if (h == nullptr || params_dev == nullptr) return fail(RNNWF_EINVAL, "null argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // handles whose sweeps run their own kernels (tensor-core regimes, dedicated 1D kernels) need the generic FFMA packs only
  // for rnnwf_logpsi_static: packed there on demand instead of after every optimiser step
  if (h->g1 || h->tc_enabled || h->tc4) h->ffma_stale = true;
  else pack_ffma(h, params_dev, st);
  const TensorInfo* wd = h->tensor("dense.W");
  const TensorInfo* bd = h->tensor("dense.b");
  const TensorInfo* wp = h->tensor("dense_phase.W");
  const TensorInfo* bp = h->tensor("dense_phase.b");
  const bool tc_all = h->tc_enabled && h->cell != CELL_VANILLA2D;     // tensorized cells: everything in one launch below
  if (!tc_all) {
    RNNWF_COUNT_LAUNCH(); pack_head_kernel<<<cdiv((long long)h->nhp * h->ND, 256), 256, 0, st>>>(
        params_dev + wd->offset, params_dev + bd->offset, wp ? params_dev + wp->offset : nullptr,
        bp ? params_dev + bp->offset : nullptr, h->nh, h->nhp, h->D, h->ND, h->d_head_w, h->d_head_b);
  }
  if (h->g1) {
    G1PackSrc S;
    std::memset(&S, 0, sizeof(S));
    for (int l = 0; l < h->symL; ++l) {
      const std::string p = "gru" + std::to_string(l) + ".";
      S.Wg[l] = params_dev + h->tensor(p + "Wg")->offset;   S.bg[l] = params_dev + h->tensor(p + "bg")->offset;
      S.Wci[l] = params_dev + h->tensor(p + "Wci")->offset; S.Wch[l] = params_dev + h->tensor(p + "Wch")->offset;
      S.bci[l] = params_dev + h->tensor(p + "bci")->offset; S.bch[l] = params_dev + h->tensor(p + "bch")->offset;
    }
    S.head_w = params_dev + wd->offset; S.head_b = params_dev + bd->offset;
    S.phase_w = wp ? params_dev + wp->offset : nullptr; S.phase_b = bp ? params_dev + bp->offset : nullptr;
    RNNWF_COUNT_LAUNCH();
    CUDA_TRY(g1_pack_launch(S, h->nh, h->symL, h->Tsym, h->D, h->ND, h->d_g1pack, st));
    if (h->g1_bwd) { RNNWF_COUNT_LAUNCH(); CUDA_TRY(g1_pack_bwd_launch(S, h->nh, h->symL, h->D, h->ND, h->d_g1packT, st)); }
  }
  if (h->tc_enabled) {
    const TcSmem ts(h->nhp, h->Tsym, h->Q);
    if (h->cell == CELL_VANILLA2D) {
      RNNWF_COUNT_LAUNCH(); pack_tc_vanilla_kernel<<<std::min(2048, cdiv((long long)ts.b_bytes / 2, 256)), 256, 0, st>>>(
          params_dev + h->tensor("Wh")->offset, params_dev + h->tensor("Wv")->offset, h->nh, h->nhp, h->tc_split ? 1 : 0, reinterpret_cast<__half*>(h->d_tcB));
      RNNWF_COUNT_LAUNCH(); pack_tc_vanilla_bias_kernel<<<4, 256, 0, st>>>(params_dev + h->tensor("Uh")->offset, params_dev + h->tensor("Uv")->offset,
                                                                           params_dev + h->tensor("b")->offset, h->nh, h->nhp, h->d_tcbias);
      if (h->van2) {
        RNNWF_COUNT_LAUNCH();
        CUDA_TRY(van2_pack_launch(params_dev + h->tensor("Wh")->offset, params_dev + h->tensor("Wv")->offset, h->nh, h->nhp, h->d_van2B, st));
      }
    } else {
      const long long n = 3LL * h->tc_np * (ts.b_bytes / 2) * (h->tc_split ? 2 : 1);
      auto opt = [&](const char* nm) -> const float* { const TensorInfo* t = h->tensor(nm); return t ? params_dev + t->offset : nullptr; };
      PackTcAll A;
      A.W = params_dev + h->tensor("W")->offset; A.Wg = opt("Wg"); A.Wmerge = opt("Wmerge"); A.b = params_dev + h->tensor("b")->offset; A.bg = opt("bg");
      A.Wd = params_dev + wd->offset; A.bd = params_dev + bd->offset; A.Wp = wp ? params_dev + wp->offset : nullptr; A.bp = bp ? params_dev + bp->offset : nullptr;
      A.nh = h->nh; A.nhp = h->nhp; A.NP = h->tc_np; A.split = h->tc_split ? 1 : 0; A.D = h->D; A.ND = h->ND;
      A.nb_tc = std::min(2048, cdiv(n, 256)); A.nb_bwd = 512;
      A.tcB = reinterpret_cast<__half*>(h->d_tcB); A.tcbias = h->d_tcbias; A.wimgT = h->d_wimgT; A.head_w = h->d_head_w; A.head_b = h->d_head_b;
      RNNWF_COUNT_LAUNCH();
      pack_tc_all_kernel<<<A.nb_tc + A.nb_bwd + 1 + cdiv((long long)h->nhp * h->ND, 256), 256, 0, st>>>(A);
    }
  }
  if (h->tc4) {
    RNNWF_COUNT_LAUNCH();
    CUDA_TRY(tc4_pack_launch(params_dev + h->tensor("W")->offset, params_dev + h->tensor("Wg")->offset, params_dev + h->tensor("Wmerge")->offset,
                             params_dev + h->tensor("b")->offset, params_dev + h->tensor("bg")->offset, h->nh, h->tc_np, h->tc4_split, h->d_tcB, st));
  }
  CUDA_TRY(cudaGetLastError());
  h->d_params = params_dev;
  h->weights_set = true;
  return RNNWF_OK;
}

int64_t rnnwf_workspace_bytes(rnnwf_handle h) {
  if (h == nullptr) return 0;
  return align256(ring_bytes(*h, max_ctas(*h), 128));
}

int64_t rnnwf_saved_bytes(rnnwf_handle h, int64_t rows) { return h ? h->saved_layout(rows).total : 0; }

int rnnwf_sample(rnnwf_handle h, int64_t rows, const float* uniforms_dev, uint64_t seed, uint64_t row_offset,
                 int8_t* samples_dev, double* logp_dev, double* phase_dev, void* saved_dev, void* workspace_dev,
                 int64_t workspace_bytes, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (rows <= 0 || samples_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_sample: rows must be positive and samples non-null");
  LaunchShape s;
  if ((rc = pick_shape(*h, rows, s))) return rc;
  if (workspace_dev == nullptr || workspace_bytes < ring_bytes(*h, s.grid, s.Mp)) return fail(RNNWF_ENOMEM, "rnnwf_sample: workspace too small");
  SweepParams P;
  fill_model_params(*h, s, P);
  P.mode = MODE_SAMPLE; P.rows = rows; P.cfg_out = samples_dev; P.uniforms = uniforms_dev; P.seed = seed; P.row_offset = row_offset;
  P.out_logp = logp_dev; P.out_phase = phase_dev; P.ring = static_cast<float*>(workspace_dev);
  set_saved(*h, P, saved_dev, rows);
  return launch_sweep(*h, P, s, static_cast<cudaStream_t>(stream));
}

int rnnwf_logpsi(rnnwf_handle h, int64_t rows, const int8_t* samples_dev, double* logp_dev, double* phase_dev,
                 void* saved_dev, void* workspace_dev, int64_t workspace_bytes, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (rows <= 0 || samples_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_logpsi: rows must be positive and samples non-null");
  LaunchShape s;
  if ((rc = pick_shape(*h, rows, s))) return rc;
  if (workspace_dev == nullptr || workspace_bytes < ring_bytes(*h, s.grid, s.Mp)) return fail(RNNWF_ENOMEM, "rnnwf_logpsi: workspace too small");
  SweepParams P;
  fill_model_params(*h, s, P);
  P.mode = MODE_EVAL; P.rows = rows; P.cfg_in = samples_dev; P.out_logp = logp_dev; P.out_phase = phase_dev;
  P.ring = static_cast<float*>(workspace_dev);
  set_saved(*h, P, saved_dev, rows);
  return launch_sweep(*h, P, s, static_cast<cudaStream_t>(stream));
}

// ---- local energies --------------------------------------------------------------------------------
namespace {
struct ElocLayout {
  long long off_ham, off_bonds, off_bstart, off_nb, off_counts, off_offsets, off_nchild, off_children, off_ring, total;
  long long max_children;
};

long long max_children_per_sample(const rnnwf_model& m, int ham) {
  switch (ham) {
    case RNNWF_HAM_TFIM1D:
    case RNNWF_HAM_TFIM2D: return m.Tsym;
    case RNNWF_HAM_RYDBERG: return 3LL * m.Tsym;
    case RNNWF_HAM_J1J2: return 2LL * m.Tsym;
    case RNNWF_HAM_HEIS2D: return 3LL * m.Tsym;
  }
  return 0;
}

ElocLayout eloc_layout(const rnnwf_model& m, int ham, long long rows) {
  ElocLayout L;
  long long o = 0;
  L.off_ham = o;      o += align256((3LL * m.Tsym + 16) * 8);
  L.off_bonds = o;    o += align256(3LL * m.Tsym * sizeof(Bond) + 64);
  L.off_bstart = o;   o += align256((m.Tsym + 2) * 4LL);
  L.off_nb = o;       o += align256(256 * sizeof(RydNb));
  const bool variable = (ham == RNNWF_HAM_J1J2 || ham == RNNWF_HAM_HEIS2D);
  const long long ncnt = variable ? (long long)m.Tsym * rows + 1 : 0;
  L.off_counts = o;   o += align256(ncnt * 4);
  L.off_offsets = o;  o += align256(ncnt * 4);
  L.off_nchild = o;   o += 2 * 1024 * 4;     // scratch of the multi-block scan
  L.max_children = max_children_per_sample(m, ham) * rows;
  L.off_children = o; o += align256(L.max_children * (long long)sizeof(Child));
  L.off_ring = o;     o += align256(ring_bytes(m, max_ctas(m), 128));
  L.total = o;
  return L;
}
}  // namespace

int64_t rnnwf_eloc_workspace_bytes(rnnwf_handle h, int hamiltonian, int64_t rows) {
  if (h == nullptr || rows <= 0) return 0;
  return eloc_layout(*h, hamiltonian, rows).total;
}

namespace {
struct EnumOut { long long nchild_fixed = 0; const int* rows_dev = nullptr; bool have_children = true; bool counted = false; };   // counted: *nconn written by the enumeration

// Diagonal matrix elements into eloc_dev and the list of connected configurations (Child records, ordered by first
// changed generation step) into the workspace -- shared by rnnwf_local_energies and rnnwf_local_energies_sym.
int enumerate_connected(const rnnwf_model& m, int hamiltonian, const double* ham, int ham_len, long long rows,
                        const int8_t* samples_dev, double* eloc_dev, char* ws, const ElocLayout& L, cudaStream_t st, EnumOut& out,
                        long long* nconn_dev = nullptr) {
  int rc;
  bool counted = false;          // the enumeration kernels wrote the number of connected configurations to *nconn_dev
  // Hamiltonian tables (couplings, bond list, neighbour table) live in the handle and are rebuilt / uploaded only when the
  // `ham` block changes: a training loop passes the same block every step, so the hot call stages nothing from the host.
  // The host images the uploads read from are kept in the handle, so the stream-ordered copies never outlive their source.
  char* const tab = static_cast<char*>(m.d_hamtab);
  double* d_ham = reinterpret_cast<double*>(tab + L.off_ham);
  Bond* d_bonds = reinterpret_cast<Bond*>(tab + L.off_bonds);
  int* d_bstart = reinterpret_cast<int*>(tab + L.off_bstart);
  RydNb* d_nb = reinterpret_cast<RydNb*>(tab + L.off_nb);
  const bool hit = (m.ham_kind == hamiltonian && (int)m.ham_block.size() == ham_len && std::equal(ham, ham + ham_len, m.ham_block.begin()));
  auto upload = [&](void* dst, const void* src, size_t bytes, std::vector<unsigned char>& keep) -> cudaError_t {
    keep.assign(static_cast<const unsigned char*>(src), static_cast<const unsigned char*>(src) + bytes);
    return cudaMemcpyAsync(dst, keep.data(), bytes, cudaMemcpyHostToDevice, st);
  };
  int* d_counts = reinterpret_cast<int*>(ws + L.off_counts);
  int* d_offsets = reinterpret_cast<int*>(ws + L.off_offsets);
  Child* d_children = reinterpret_cast<Child*>(ws + L.off_children);
  const int Nx = m.cfg.nx, Ny = m.cfg.ny, T = m.Tsym;
  const int tpb = 256;
  const int gsamp = std::min(4096, cdiv(rows, tpb));
  long long nchild_fixed = 0;
  const int* rows_dev = nullptr;
  bool have_children = true;

  auto need_cell = [&](bool ok, const char* what) { return ok ? RNNWF_OK : fail(RNNWF_EINVAL, std::string("rnnwf_local_energies: ") + what); };

  switch (hamiltonian) {
    case RNNWF_HAM_TFIM1D:
    case RNNWF_HAM_TFIM2D: {
      if ((rc = need_cell(m.bits == 1, "TFIM needs a spin-1/2 wavefunction"))) return rc;
      const int nJ = (hamiltonian == RNNWF_HAM_TFIM1D) ? std::max(1, Nx - 1) : Nx * Ny;
      if (hamiltonian == RNNWF_HAM_TFIM1D && Ny != 1) return fail(RNNWF_EINVAL, "TFIM1D needs a chain");
      if (ham_len < 1 + nJ) return fail(RNNWF_EINVAL, "TFIM: ham = {Bx, Jz...} too short");
      // 1D: Jz[i] couples (i,i+1) == 2D kernel with Ny = 1 and Jz[i*1+0]
      if (!hit) {
        std::vector<double> J(Nx * Ny, 0.0);
        for (int i = 0; i < nJ && i < Nx * Ny; ++i) J[i] = ham[1 + i];
        CUDA_TRY(upload(d_ham, J.data(), J.size() * 8, m.ham_host[0]));
      }
      RNNWF_COUNT_LAUNCH(); tfim_diag_kernel<<<gsamp, tpb, 0, st>>>(samples_dev, rows, Nx, Ny, d_ham, eloc_dev);
      const double Bx = ham[0];
      if (Bx == 0.0) have_children = false;
      else {
        nchild_fixed = (long long)T * rows;
        RNNWF_COUNT_LAUNCH(); all_flips_kernel<<<std::min(8192, cdiv(nchild_fixed, tpb)), tpb, 0, st>>>(rows, T, 1, (float)(-Bx), d_children, nconn_dev); counted = nconn_dev != nullptr;
      }
      break;
    }
    case RNNWF_HAM_RYDBERG: {
      if ((rc = need_cell(m.cell == CELL_GRU4, "Rydberg needs the Kagome unit-cell wavefunction"))) return rc;
      if (ham_len < 4) return fail(RNNWF_EINVAL, "Rydberg: ham = {V, delta, Omega, Rcutoff}");
      const double V = ham[0], delta = ham[1], Omega = ham[2], Rc = ham[3];
      // generate_kagome + neighbouring_atoms (RydbergKagome_2DGRU/Helping_Functions.py:33-96)
      auto pos = [&](int i, int j, int k, double& x, double& y) {
        const double lx = 2.0 * (i + j / 2.0), ly = 2.0 * (std::sqrt(3.0) / 2.0 * j);
        if (k == 1) { x = lx; y = ly; } else if (k == 2) { x = lx + 1.0; y = ly; } else { x = lx + 0.5; y = ly + std::sqrt(3.0) / 2.0; }
      };
      std::vector<RydNb> nb;
      const double a1x = 2.0, a1y = 0.0, a2x = 1.0, a2y = std::sqrt(3.0);
      if (!hit)
      for (int i = 0; i < Nx; ++i)
        for (int j = 0; j < Ny; ++j)
          for (int k = 0; k < 3; ++k)
            for (int kc = 0; kc < 3; ++kc) {
              double x, y, x0, y0;
              pos(i, j, k, x, y);
              pos(0, 0, kc, x0, y0);
              const double sx[4] = {0.0, Nx * a1x, Ny * a2x, Nx * a1x + Ny * a2x};
              const double sy[4] = {0.0, Nx * a1y, Ny * a2y, Nx * a1y + Ny * a2y};
              for (int im = 0; im < 4; ++im) {
                const double dx = x - sx[im] - x0, dy = y - sy[im] - y0;
                const double d = std::sqrt(dx * dx + dy * dy);
                if (d > 0 && d <= Rc) {
                  RydNb e; e.kc = (int16_t)kc; e.di = (int16_t)i; e.dj = (int16_t)j; e.k = (int16_t)k;
                  e.w = (V / 2.0) / std::pow(d, 6.0);
                  nb.push_back(e);
                }
              }
            }
      if (!hit) {
        if (nb.size() > 256) return fail(RNNWF_EINVAL, "Rydberg: too many neighbours within Rcutoff (max 256)");
        if (!nb.empty()) CUDA_TRY(upload(d_nb, nb.data(), nb.size() * sizeof(RydNb), m.ham_host[0]));
        m.ham_n[0] = (int)nb.size();
      }
      RNNWF_COUNT_LAUNCH(); rydberg_diag_kernel<<<std::min(8192, cdiv(rows * 32, tpb)), tpb, 0, st>>>(samples_dev, rows, Nx, Ny, d_nb, m.ham_n[0], delta, eloc_dev);
      if (Omega == 0.0) have_children = false;
      else {
        nchild_fixed = 3LL * T * rows;
        RNNWF_COUNT_LAUNCH(); all_flips_kernel<<<std::min(8192, cdiv(nchild_fixed, tpb)), tpb, 0, st>>>(rows, T, 3, (float)(-Omega / 2.0), d_children, nconn_dev); counted = nconn_dev != nullptr;
      }
      break;
    }
    case RNNWF_HAM_J1J2:
    case RNNWF_HAM_HEIS2D: {
      if ((rc = need_cell(m.bits == 1, "exchange Hamiltonians need a spin-1/2 wavefunction"))) return rc;
      struct HB { int ta, tb; float off, diag; };
      std::vector<HB> hb;
      std::vector<double> Bz;
      auto add_bond = [&](int sa, int sb, double off, double diag) {
        int ta = m.t_of_site[sa], tb = m.t_of_site[sb];
        if (ta > tb) std::swap(ta, tb);
        hb.push_back({ta, tb, (float)off, (float)diag});
      };
      if (hamiltonian == RNNWF_HAM_J1J2) {
        const int N = Nx;
        if (Ny != 1) return fail(RNNWF_EINVAL, "J1J2 needs a chain");
        if (ham_len < 2 + 3 * N) return fail(RNNWF_EINVAL, "J1J2: ham = {periodic, marshall, J1[N], J2[N], Bz[N]}");
        const bool periodic = ham[0] != 0.0, marshall = ham[1] != 0.0;
        const double* J1 = ham + 2; const double* J2 = ham + 2 + N; const double* B = ham + 2 + 2 * N;
        const int lim1 = periodic ? N : N - 1, lim2 = periodic ? N : N - 2;
        for (int i = 0; i < lim1; ++i) add_bond(i, (i + 1) % N, J1[i] != 0.0 ? (marshall ? -J1[i] / 2 : J1[i] / 2) : 0.0, 0.25 * J1[i]);
        for (int i = 0; i < lim2; ++i) if (J2[i] != 0.0) add_bond(i, (i + 2) % N, J2[i] / 2, 0.25 * J2[i]);
        Bz.assign(B, B + N);
      } else {
        if (ham_len < 1) return fail(RNNWF_EINVAL, "HEIS2D: ham = {J2}");
        if ((rc = need_cell(Ny >= 1 && m.cell != CELL_GRU1D, "HEIS2D needs a 2D wavefunction"))) return rc;
        const double J2 = ham[0];
        for (int nx = 0; nx < Nx; ++nx)
          for (int ny = 0; ny < Ny; ++ny) {
            if (nx + 1 < Nx) add_bond(nx * Ny + ny, (nx + 1) * Ny + ny, -0.5, 0.25);
            if (ny + 1 < Ny) add_bond(nx * Ny + ny, nx * Ny + ny + 1, -0.5, 0.25);
            if (nx + 1 < Nx && ny + 1 < Ny) add_bond(nx * Ny + ny, (nx + 1) * Ny + ny + 1, J2 / 2, 0.25 * J2);
          }
      }
      if ((long long)hb.size() > 3LL * T) return fail(RNNWF_EINVAL, "too many bonds");
      std::stable_sort(hb.begin(), hb.end(), [](const HB& a, const HB& b) { return a.ta < b.ta; });
      std::vector<Bond> bonds(hb.size());
      std::vector<int> bstart(T + 1, 0);
      for (size_t i = 0; i < hb.size(); ++i) {
        bonds[i].ta = (int16_t)hb[i].ta; bonds[i].tb = (int16_t)hb[i].tb; bonds[i].offdiag = hb[i].off; bonds[i].diag = hb[i].diag; bonds[i].pad = 0;
        bstart[hb[i].ta + 1]++;
      }
      for (int t = 0; t < T; ++t) bstart[t + 1] += bstart[t];
      if (!hit) {
        if (!bonds.empty()) CUDA_TRY(upload(d_bonds, bonds.data(), bonds.size() * sizeof(Bond), m.ham_host[0]));
        CUDA_TRY(upload(d_bstart, bstart.data(), bstart.size() * 4, m.ham_host[1]));
        if (!Bz.empty()) CUDA_TRY(upload(d_ham, Bz.data(), Bz.size() * 8, m.ham_host[2]));
      }
      // three launches: counts per (first changed symbol, sample) + the diagonal matrix elements; their exclusive scan; the
      // child records (+ the number of connected configurations)
      const long long ncnt = (long long)T * rows;
      RNNWF_COUNT_LAUNCH(); bond_count_kernel<<<std::min(8192, cdiv(ncnt + 1, tpb)), tpb, 0, st>>>(samples_dev, rows, m.nsites, T, m.d_site_of_t, d_bonds, d_bstart, d_counts,
                                                                                                (int)bonds.size(), d_ham, (int)Bz.size(), eloc_dev);
      exclusive_scan(d_counts, d_offsets, ncnt + 1, reinterpret_cast<int*>(ws + L.off_nchild), st);
      RNNWF_COUNT_LAUNCH(); bond_fill_kernel<<<std::min(8192, cdiv(ncnt, tpb)), tpb, 0, st>>>(samples_dev, rows, m.nsites, T, m.d_site_of_t, d_bonds, d_bstart, d_offsets, d_children, nconn_dev);
      counted = nconn_dev != nullptr;
      rows_dev = d_offsets + ncnt;
      nchild_fixed = L.max_children;
      break;
    }
    default:
      return fail(RNNWF_EINVAL, "unknown hamiltonian");
  }
  CUDA_TRY(cudaGetLastError());
  if (!hit) { m.ham_kind = hamiltonian; m.ham_block.assign(ham, ham + ham_len); }
  out.nchild_fixed = nchild_fixed; out.rows_dev = rows_dev; out.have_children = have_children; out.counted = counted;
  return RNNWF_OK;
}
}  // namespace

int rnnwf_local_energies(rnnwf_handle h, int hamiltonian, const double* ham, int ham_len, int64_t rows,
                         const int8_t* samples_dev, const double* logp_dev, const double* phase_dev,
                         const void* saved_dev, double* eloc_dev, int64_t* nconn_dev, void* workspace_dev,
                         int64_t workspace_bytes, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (rows <= 0 || samples_dev == nullptr || logp_dev == nullptr || eloc_dev == nullptr || ham == nullptr)
    return fail(RNNWF_EINVAL, "rnnwf_local_energies: null argument or rows <= 0");
  if (rows > 0x7fffffffLL) return fail(RNNWF_EINVAL, "rnnwf_local_energies: too many rows");
  const rnnwf_model& m = *h;
  const ElocLayout L = eloc_layout(m, hamiltonian, rows);
  if (workspace_dev == nullptr || workspace_bytes < L.total) return fail(RNNWF_ENOMEM, "rnnwf_local_energies: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* ws = static_cast<char*>(workspace_dev);
  EnumOut en;
  if ((rc = enumerate_connected(m, hamiltonian, ham, ham_len, rows, samples_dev, eloc_dev, ws, L, st, en, (long long*)nconn_dev))) return rc;
  Child* d_children = reinterpret_cast<Child*>(ws + L.off_children);
  if (nconn_dev != nullptr && !en.counted) {
    // connected configurations evaluated (+ the diagonal one per sample, already known from `logp`)
    RNNWF_COUNT_LAUNCH(); count_children_kernel<<<1, 32, 0, st>>>(en.have_children ? en.rows_dev : nullptr, en.have_children ? (en.rows_dev ? 0 : en.nchild_fixed) : 0, (long long*)nconn_dev);
  }
  if (!en.have_children) return RNNWF_OK;

  LaunchShape s;
  if ((rc = pick_shape(m, en.nchild_fixed, s, true))) return rc;
  SweepParams P;
  fill_model_params(m, s, P);
  P.mode = MODE_CHILD; P.rows = en.nchild_fixed; P.rows_dev = en.rows_dev; P.cfg_in = samples_dev; P.children = d_children;
  P.par_logp = logp_dev; P.par_phase = phase_dev; P.eloc = eloc_dev;
  P.ring = reinterpret_cast<float*>(ws + L.off_ring);
  if (en.rows_dev == nullptr) {      // single-flip Hamiltonians: the child list is regular (all_flips_kernel)
    P.sched_ns = rows;
    P.sched_bits = (hamiltonian == RNNWF_HAM_RYDBERG) ? 3 : 1;
  }
  if (saved_dev != nullptr) {
    const Saved sv = m.saved_layout(rows);
    const char* b = static_cast<const char*>(saved_dev);
    P.par_hist = reinterpret_cast<const float*>(b + sv.off_hist);
    P.par_cum = reinterpret_cast<const double2*>(b + sv.off_cum);
    P.par_rows = rows;
  }
  return launch_sweep(m, P, s, st);
}

// ---- backward ------------------------------------------------------------------------------------
int64_t rnnwf_backward_workspace_bytes(rnnwf_handle h, int64_t rows) {
  if (h == nullptr || rows <= 0) return 0;
  return backward_workspace_bytes(*h, rows);
}

int rnnwf_backward(rnnwf_handle h, int64_t rows, const int8_t* samples_dev, const void* saved_dev,
                   const double* seed_logp_dev, const double* seed_phase_dev, float* grad_dev, void* workspace_dev,
                   int64_t workspace_bytes, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (rows <= 0 || samples_dev == nullptr || saved_dev == nullptr || seed_logp_dev == nullptr || grad_dev == nullptr)
    return fail(RNNWF_EINVAL, "rnnwf_backward: null argument or rows <= 0");
  if (workspace_dev == nullptr || workspace_bytes < backward_workspace_bytes(*h, rows))
    return fail(RNNWF_ENOMEM, "rnnwf_backward: workspace too small");
  std::string err;
  rc = backward_run(*h, rows, samples_dev, saved_dev, seed_logp_dev, seed_phase_dev, grad_dev, workspace_dev,
                    static_cast<cudaStream_t>(stream), err);
  if (rc) return fail(rc, err);
  return RNNWF_OK;
}

long long rnnwf_launch_count(void) { return rnnwf::launch_counter(); }

// debug hook (not declared in the public header): run the tcgen05 TF32 GEMM on raw FP32 device matrices
int rnnwf_debug_gemm_tc(const float* A, long long lda, int a_kc, const float* B, long long ldb, int b_kc, float* C, long long ldc,
                        long long M, int N, long long K, int mode, void* stream) {
  cudaFuncSetAttribute(gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GT_STAGES * 2 * GT_TILE_BYTES);
  if (!gemm_tc_launch(static_cast<cudaStream_t>(stream), A, lda, a_kc, B, ldb, b_kc, C, ldc, M, N, K, mode)) return fail(RNNWF_EINVAL, "gemm_tc: unaligned operands");
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

// debug hook (not declared in the public header): device buffer of gridDim.x * 16 uint64 cycle counters
void rnnwf_debug_set_tc_prof(unsigned long long* dev) { g_tc_prof = dev; }

// FP32 FFMA peak probe (roofline denominator of the FFMA regime): every thread runs 16 independent
// FMA chains; flops = blocks * threads * iters * 16 * 2.
__global__ void __launch_bounds__(512) ffma_probe_kernel(float* out, int iters, float seedv) {
  float a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = seedv + (float)(threadIdx.x + i);
  const float m = 0.999f + seedv * 1e-9f, c = 1e-3f;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], m, c);
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  if (s == 123.456f) out[0] = s;
}

int rnnwf_ffma_probe(float* out_dev, int blocks, int iters, void* stream) {
  RNNWF_COUNT_LAUNCH(); ffma_probe_kernel<<<blocks, 512, 0, static_cast<cudaStream_t>(stream)>>>(out_dev, iters, 1.0f);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

// ---- fused optimiser step ---------------------------------------------------------------------------
// tf.train.AdamOptimizer.apply_gradients in one pass over the flat parameter vector (TF 1.13 semantics, SURVEY App. B:
// m <- b1 m + (1-b1) g, v <- b2 v + (1-b2) g^2, theta <- theta - lr_t m / (sqrt(v) + eps), lr_t = lr sqrt(1-b2^t)/(1-b1^t)),
// with the optional clip-by-value the Heisenberg / Rydberg scripts apply to the gradient first.
__global__ void __launch_bounds__(256) adam_step_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                        float* __restrict__ v, long long n, float lr_t, float b1, float b2, float eps,
                                                        float clip) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float gi = g[i];
    if (clip > 0.f) gi = fminf(fmaxf(gi, -clip), clip);
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi;
    v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

int rnnwf_adam_step(float* params_dev, const float* grad_dev, float* m_dev, float* v_dev, int64_t n, double lr, double beta1,
                    double beta2, double eps, int64_t t, double clip, void* stream) {
  if (params_dev == nullptr || grad_dev == nullptr || m_dev == nullptr || v_dev == nullptr || n <= 0 || t < 1)
    return fail(RNNWF_EINVAL, "rnnwf_adam_step: null argument, n <= 0 or t < 1");
  const double lr_t = lr * std::sqrt(1.0 - std::pow(beta2, (double)t)) / (1.0 - std::pow(beta1, (double)t));
  RNNWF_COUNT_LAUNCH();
  adam_step_kernel<<<std::min(1184, cdiv(n, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      params_dev, grad_dev, m_dev, v_dev, n, (float)lr_t, (float)beta1, (float)beta2, (float)eps, (float)clip);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

// theta <- theta - lr sign(g) u,  u = philox_uniform(seed, element, step): the `normalized_gradients` optimiser
__global__ void __launch_bounds__(256) sign_sgd_step_kernel(float* __restrict__ p, const float* __restrict__ g, long long n, float lr,
                                                            unsigned long long seed, uint32_t t) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float sg = (gi > 0.f) ? 1.f : (gi < 0.f ? -1.f : 0.f);
    p[i] -= lr * sg * philox_uniform(seed, (unsigned long long)i, t);
  }
}

int rnnwf_sign_sgd_step(float* params_dev, const float* grad_dev, int64_t n, double lr, uint64_t seed, int64_t t, void* stream) {
  if (params_dev == nullptr || grad_dev == nullptr || n <= 0 || t < 0) return fail(RNNWF_EINVAL, "rnnwf_sign_sgd_step: null argument, n <= 0 or t < 0");
  RNNWF_COUNT_LAUNCH();
  sign_sgd_step_kernel<<<std::min(1184, cdiv(n, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(params_dev, grad_dev, n, (float)lr,
                                                                                               (unsigned long long)seed, (uint32_t)t);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

// ---- energy moments and the seeds of the reference's surrogate costs ---------------------------------------
// moments[8] = {rows, sum Re E, sum Im E, sum (Re E)^2, sum (Im E)^2, sum logP, 0, 0}: one block, FP64, fixed order
__global__ void __launch_bounds__(1024) energy_moments_kernel(long long rows, const double* __restrict__ eloc, const double* __restrict__ logp,
                                                              double* __restrict__ out) {
  __shared__ double red[5][32];
  double a[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
  for (long long b = threadIdx.x; b < rows; b += blockDim.x) {
    const double re = eloc[2 * b], im = eloc[2 * b + 1];
    a[0] += re; a[1] += im; a[2] += re * re; a[3] += im * im;
    if (logp != nullptr) a[4] += logp[b];
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 5; ++k) {
    for (int o = 16; o > 0; o >>= 1) a[k] += __shfl_xor_sync(0xffffffffu, a[k], o);
    if (lane == 0) red[k][w] = a[k];
  }
  __syncthreads();
  if (w == 0) {
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      double v = (lane < (int)(blockDim.x >> 5)) ? red[k][lane] : 0.0;
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) out[1 + k] = v;
    }
    if (lane == 0) { out[0] = (double)rows; out[6] = 0.0; out[7] = 0.0; }
  }
}

// seeds from the (all-reduced) moments; stats[4] = {mean Re E, mean Im E, var E, mean logP}
__global__ void __launch_bounds__(256) cost_seeds_kernel(long long rows, const double* __restrict__ eloc, const double* __restrict__ logp,
                                                         const double* __restrict__ mom, int complex_head, double T, int free_energy,
                                                         double* __restrict__ seed_logp, double* __restrict__ seed_phase, double* __restrict__ stats) {
  const double n = mom[0], mre = mom[1] / n, mim = mom[2] / n, mlp = mom[5] / n;
  if (blockIdx.x == 0 && threadIdx.x == 0 && stats != nullptr) {
    stats[0] = mre; stats[1] = mim; stats[2] = mom[3] / n - mre * mre + mom[4] / n - mim * mim; stats[3] = mlp;
  }
  for (long long b = blockIdx.x * (long long)blockDim.x + threadIdx.x; b < rows; b += (long long)gridDim.x * blockDim.x) {
    const double dre = eloc[2 * b] - mre, dim_ = eloc[2 * b + 1] - mim;
    double s = dre;
    if (T != 0.0 && logp != nullptr && (complex_head || free_energy)) s += T * (logp[b] - mlp);
    seed_logp[b] = s / n;
    if (complex_head && seed_phase != nullptr) seed_phase[b] = 2.0 * dim_ / n;
  }
}

int rnnwf_energy_moments(int64_t rows, const double* eloc_dev, const double* logp_dev, double* moments_dev, void* stream) {
  if (rows <= 0 || eloc_dev == nullptr || moments_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_energy_moments: null argument or rows <= 0");
  RNNWF_COUNT_LAUNCH();
  energy_moments_kernel<<<1, 1024, 0, static_cast<cudaStream_t>(stream)>>>(rows, eloc_dev, logp_dev, moments_dev);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

int rnnwf_cost_seeds(int64_t rows, const double* eloc_dev, const double* logp_dev, const double* moments_dev, int complex_head, double T,
                     int free_energy, double* seed_logp_dev, double* seed_phase_dev, double* stats_dev, void* stream) {
  if (rows <= 0 || eloc_dev == nullptr || moments_dev == nullptr || seed_logp_dev == nullptr)
    return fail(RNNWF_EINVAL, "rnnwf_cost_seeds: null argument or rows <= 0");
  if (complex_head && seed_phase_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_cost_seeds: complex cost needs seed_phase");
  RNNWF_COUNT_LAUNCH();
  cost_seeds_kernel<<<std::min(592, cdiv(rows, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(rows, eloc_dev, logp_dev, moments_dev, complex_head, T,
                                                                                                  free_energy, seed_logp_dev, seed_phase_dev, stats_dev);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

// ---- symmetry-projected amplitudes --------------------------------------------------------------------
static int sym_chars(int G, const double* chi, SymChars& ch) {
  if (G < 1 || G > SYM_MAX_G) return fail(RNNWF_EINVAL, "symmetry epilogue: the group must have 1..64 elements");
  for (int g = 0; g < SYM_MAX_G; ++g) ch.inv_chi[g] = (g < G) ? ((chi != nullptr) ? 1.0 / chi[g] : 1.0) : 0.0;
  return RNNWF_OK;
}

int rnnwf_symmetry_epilogue(int G, int64_t rows, const double* logp_dev, const double* phase_dev, const double* chi,
                            double* out_logp_dev, double* out_phase_dev, void* stream) {
  if (rows <= 0 || logp_dev == nullptr || out_logp_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_symmetry_epilogue: null argument or rows <= 0");
  SymChars ch;
  int rc = sym_chars(G, chi, ch);
  if (rc) return rc;
  RNNWF_COUNT_LAUNCH();
  symmetry_epilogue_kernel<<<std::min(1184, cdiv(rows, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(G, rows, logp_dev, phase_dev, ch, out_logp_dev, out_phase_dev);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

int rnnwf_symmetry_seeds(int G, int64_t rows, const double* logp_dev, const double* phase_dev, const double* chi,
                         const double* s_logp_dev, const double* s_phase_dev, double* seed_logp_dev, double* seed_phase_dev, void* stream) {
  if (rows <= 0 || logp_dev == nullptr || s_logp_dev == nullptr || seed_logp_dev == nullptr)
    return fail(RNNWF_EINVAL, "rnnwf_symmetry_seeds: null argument or rows <= 0");
  SymChars ch;
  int rc = sym_chars(G, chi, ch);
  if (rc) return rc;
  RNNWF_COUNT_LAUNCH();
  symmetry_seeds_kernel<<<std::min(1184, cdiv(rows, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(G, rows, logp_dev, phase_dev, ch, s_logp_dev, s_phase_dev,
                                                                                                      seed_logp_dev, seed_phase_dev);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

// ---- symmetry groups as device-resident index-remap tables ------------------------------------------------
}  // extern "C" (reopened below)

struct rnnwf_symmetry {
  int G = 0, nsites = 0, bits = 1, device = 0;
  int32_t* d_perm = nullptr;
  int8_t* d_flip = nullptr;
  int16_t* d_site_of_t = nullptr;
  bool has_chi = false;
  SymChars ch;
};

namespace {
struct SymElocLayout { ElocLayout e; long long off_cfg, off_logp, off_phase, off_total, off_ring2, total, Rmax; };
SymElocLayout sym_eloc_layout(const rnnwf_model& m, int G, int ham, long long rows) {
  SymElocLayout S;
  S.e = eloc_layout(m, ham, rows);
  S.Rmax = rows + S.e.max_children;
  long long o = S.e.off_ring;                  // the plain E_loc sweep ring is not used: reuse its place and grow from there
  S.off_cfg = o;    o += align256((long long)G * S.Rmax * m.nsites);
  S.off_logp = o;   o += align256((long long)G * S.Rmax * 8);
  S.off_phase = o;  o += align256((long long)G * S.Rmax * 8);
  S.off_total = o;  o += 256;
  S.off_ring2 = o;  o += align256(ring_bytes(m, max_ctas(m), 128));
  S.total = o;
  return S;
}
}  // namespace

extern "C" {

int rnnwf_symmetry_create(rnnwf_handle h, int G, const int32_t* perm, const int8_t* flip, const double* chi, rnnwf_symmetry_handle* out) {
  if (h == nullptr || perm == nullptr || out == nullptr) return fail(RNNWF_EINVAL, "rnnwf_symmetry_create: null argument");
  *out = nullptr;
  if (G < 1 || G > SYM_MAX_G) return fail(RNNWF_EINVAL, "rnnwf_symmetry_create: the group must have 1..64 elements");
  const int ns = h->nsites;
  std::vector<char> seen(ns);
  for (int g = 0; g < G; ++g) {
    std::fill(seen.begin(), seen.end(), 0);
    for (int i = 0; i < ns; ++i) {
      const int j = perm[(size_t)g * ns + i];
      if (j < 0 || j >= ns || seen[j]) return fail(RNNWF_EINVAL, "rnnwf_symmetry_create: every group element must be a permutation of the sites");
      seen[j] = 1;
    }
  }
  rnnwf_symmetry* s = new rnnwf_symmetry();
  s->G = G; s->nsites = ns; s->bits = h->bits; s->device = h->device;
  s->has_chi = chi != nullptr;
  for (int g = 0; g < SYM_MAX_G; ++g) s->ch.inv_chi[g] = (g < G) ? ((chi != nullptr) ? 1.0 / chi[g] : 1.0) : 0.0;
  std::vector<int8_t> fl(G, 0);
  if (flip != nullptr) for (int g = 0; g < G; ++g) fl[g] = flip[g] ? 1 : 0;
  cudaError_t e = cudaMalloc(&s->d_perm, (size_t)G * ns * 4);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_flip, (size_t)G);
  if (e == cudaSuccess) e = cudaMalloc(&s->d_site_of_t, h->site_of_t.size() * sizeof(int16_t));
  if (e == cudaSuccess) e = cudaMemcpy(s->d_perm, perm, (size_t)G * ns * 4, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(s->d_flip, fl.data(), (size_t)G, cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMemcpy(s->d_site_of_t, h->site_of_t.data(), h->site_of_t.size() * sizeof(int16_t), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    cudaFree(s->d_perm); cudaFree(s->d_flip); cudaFree(s->d_site_of_t); delete s;
    return fail(RNNWF_ECUDA, std::string("rnnwf_symmetry_create: ") + cudaGetErrorString(e));
  }
  *out = s;
  return RNNWF_OK;
}

int rnnwf_symmetry_destroy(rnnwf_symmetry_handle s) {
  if (s == nullptr) return RNNWF_OK;
  cudaFree(s->d_perm); cudaFree(s->d_flip); cudaFree(s->d_site_of_t);
  delete s;
  return RNNWF_OK;
}

int rnnwf_symmetry_order(rnnwf_symmetry_handle s) { return s ? s->G : 0; }

int rnnwf_symmetry_apply(rnnwf_symmetry_handle s, int64_t rows, const int8_t* in_dev, int8_t* out_dev, void* stream) {
  if (s == nullptr || rows <= 0 || in_dev == nullptr || out_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_symmetry_apply: null argument or rows <= 0");
  SymExpandParams P;
  std::memset(&P, 0, sizeof(P));
  P.G = s->G; P.nsites = s->nsites; P.bits = s->bits; P.rows = rows; P.perm = s->d_perm; P.flip = s->d_flip;
  P.samples = in_dev; P.site_of_t = s->d_site_of_t; P.out = out_dev;
  const long long total = (long long)rows * s->G * s->nsites;
  RNNWF_COUNT_LAUNCH();
  sym_expand_kernel<<<std::min(148 * 16, cdiv(total, 256)), 256, 0, static_cast<cudaStream_t>(stream)>>>(P);
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

int64_t rnnwf_eloc_sym_workspace_bytes(rnnwf_handle h, rnnwf_symmetry_handle s, int hamiltonian, int64_t rows) {
  if (h == nullptr || s == nullptr || rows <= 0) return 0;
  return sym_eloc_layout(*h, s->G, hamiltonian, rows).total;
}

int rnnwf_local_energies_sym(rnnwf_handle h, rnnwf_symmetry_handle s, int hamiltonian, const double* ham, int ham_len, int64_t rows,
                             const int8_t* samples_dev, double* out_logp_dev, double* out_phase_dev, double* eloc_dev,
                             int64_t* nconn_dev, void* workspace_dev, int64_t workspace_bytes, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (s == nullptr || rows <= 0 || samples_dev == nullptr || out_logp_dev == nullptr || eloc_dev == nullptr || ham == nullptr)
    return fail(RNNWF_EINVAL, "rnnwf_local_energies_sym: null argument or rows <= 0");
  const rnnwf_model& m = *h;
  if (s->nsites != m.nsites || s->bits != m.bits) return fail(RNNWF_EINVAL, "rnnwf_local_energies_sym: the symmetry table was built for another lattice");
  if (m.complex_head && out_phase_dev == nullptr) return fail(RNNWF_EINVAL, "rnnwf_local_energies_sym: complex wavefunction needs out_phase");
  const int G = s->G;
  const SymElocLayout S = sym_eloc_layout(m, G, hamiltonian, rows);
  if ((long long)G * S.Rmax > 0x7fffffffLL) return fail(RNNWF_EINVAL, "rnnwf_local_energies_sym: too many rows");
  if (workspace_dev == nullptr || workspace_bytes < S.total) return fail(RNNWF_ENOMEM, "rnnwf_local_energies_sym: workspace too small");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* ws = static_cast<char*>(workspace_dev);
  EnumOut en;
  if ((rc = enumerate_connected(m, hamiltonian, ham, ham_len, rows, samples_dev, eloc_dev, ws, S.e, st, en, (long long*)nconn_dev))) return rc;
  Child* d_children = reinterpret_cast<Child*>(ws + S.e.off_children);
  if (nconn_dev != nullptr && !en.counted) {
    RNNWF_COUNT_LAUNCH(); count_children_kernel<<<1, 32, 0, st>>>(en.have_children ? en.rows_dev : nullptr, en.have_children ? (en.rows_dev ? 0 : en.nchild_fixed) : 0, (long long*)nconn_dev);
  }
  int8_t* d_cfg = reinterpret_cast<int8_t*>(ws + S.off_cfg);
  double* d_lp = reinterpret_cast<double*>(ws + S.off_logp);
  double* d_ph = m.complex_head ? reinterpret_cast<double*>(ws + S.off_phase) : nullptr;
  int* d_total = reinterpret_cast<int*>(ws + S.off_total);
  const long long nfixed = en.have_children ? en.nchild_fixed : 0;
  // 1. the G transformed copies of every sample and every connected configuration, formed through the remap tables
  SymExpandParams X;
  std::memset(&X, 0, sizeof(X));
  X.G = G; X.nsites = m.nsites; X.bits = m.bits; X.rows = rows; X.nchild_fixed = nfixed; X.nchild_dev = en.have_children ? en.rows_dev : nullptr;
  X.perm = s->d_perm; X.flip = s->d_flip; X.samples = samples_dev; X.children = en.have_children ? d_children : nullptr;
  X.site_of_t = s->d_site_of_t; X.out = d_cfg; X.total_rows_out = d_total;
  const long long upper = (long long)G * (rows + nfixed);
  RNNWF_COUNT_LAUNCH(); sym_expand_kernel<<<std::min(148 * 16, cdiv(upper * m.nsites, 256)), 256, 0, st>>>(X);
  // 2. one batched sweep over all of them
  LaunchShape sh;
  if ((rc = pick_shape(m, upper, sh))) return rc;
  SweepParams P;
  fill_model_params(m, sh, P);
  P.mode = MODE_EVAL; P.rows = upper; P.rows_dev = d_total; P.cfg_in = d_cfg; P.out_logp = d_lp; P.out_phase = d_ph;
  P.ring = reinterpret_cast<float*>(ws + S.off_ring2);
  if ((rc = launch_sweep(m, P, sh, st))) return rc;
  // 3. symmetrised log psi of the samples, then the ratios of the connected configurations reduced into E_loc
  RNNWF_COUNT_LAUNCH(); sym_parent_kernel<<<std::min(1184, cdiv(rows, 256)), 256, 0, st>>>(G, rows, nfixed, X.nchild_dev, d_lp, d_ph, s->ch, out_logp_dev, out_phase_dev);
  if (en.have_children) {
    RNNWF_COUNT_LAUNCH(); sym_eloc_reduce_kernel<<<std::min(148 * 8, cdiv(nfixed, 256)), 256, 0, st>>>(G, rows, nfixed, X.nchild_dev, d_children, d_lp, d_ph, s->ch,
                                                                                                       out_logp_dev, out_phase_dev, eloc_dev);
  }
  CUDA_TRY(cudaGetLastError());
  return RNNWF_OK;
}

int rnnwf_describe(rnnwf_handle h, char* buf, int cap) {
  if (h == nullptr || buf == nullptr || cap <= 0) return fail(RNNWF_EINVAL, "null argument");
  std::snprintf(buf, cap,
                "cell=%d nh=%d (padded %d) steps=%d symbols=%d slots=%d classes=%d | pipe=%s, tile=128 rows x %d col-groups "
                "(TJ=%d NTU=%d WC=%d passes=%d), k-chunk=%d, ring=%d steps, params=%lld",
                h->cell, h->nh, h->nhp, h->Q, h->Tsym, h->nslots, h->ncls_max,
                h->van2 ? "tcgen05 split precision: FP16 hi+lo operands, 3 MMAs per K step, FP32 accumulate in TMEM; sampling, evaluation, connected configurations: two-tile sweep (N = 64, weights resident in shared memory)"
                : h->tc_split ? "tcgen05 split precision: FP16 hi+lo operands, 3 MMAs per K step, FP32 accumulate in TMEM (column passes of 112, double-buffered)"
                : h->g1 ? "FFMA fp32, dedicated 1D-GRU kernels: one launch per sweep, weights resident in shared memory"
                : h->tc_enabled ? "tcgen05 FP16 operands, FP32 accumulate in TMEM (column passes of 112, double-buffered)"
                              : (h->tc4 ? (h->tc4_split ? "tcgen05 split precision: FP16 hi+lo operands streamed from the L2 ring (A) and the weight image (B), 3 MMAs per K step, FP32 accumulate in TMEM (column passes of 192, double-buffered)" : "tcgen05 FP16 operands streamed from the L2 ring (A) and the weight image (B), FP32 accumulate in TMEM (column passes of 192, double-buffered)") : "FFMA fp32"), h->Gblk, h->TJ, h->NTU, h->WC, h->ncb, KC,
                h->ring_len, (long long)h->num_params);
  return RNNWF_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------
// rnnwf_logpsi_static -- log-amplitudes for callers that cannot obtain device scratch memory.
//
// Some hosts of this library have no way to hand us a workspace: the process's deep-learning runtime
// has reserved the whole GPU up front (TensorFlow 1.x, the reference's own runtime, does so by default),
// or the binding is a bare ctypes/NumPy script with no device allocator of its own beyond the tensors
// it was given.  For them the history ring of the sweep lives in the fixed array below, which is part
// of the library image: every GPU that loads librnnwf.so carries these RNNWF_STATIC_SCRATCH_BYTES until
// the process exits, whether or not this entry point is ever called.
//
// The array's name is only meaningful in device code; on the host it is a handle for the symbol.  The
// device address is looked up once per device with cudaGetSymbolAddress (first call on that device,
// with that device current) and remembered.  Make that first call outside any stream capture.
// Nothing in the ring has to start at zero (every history slot is written before it is read), so it
// is never cleared; it simply keeps whatever the previous call left.  All calls on a device share the
// one array: two calls whose kernels overlap on the GPU (different streams) corrupt each other.
// ------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[RNNWF_STATIC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
float* g_static_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_static_scratch_mu;

int static_scratch_address(int device, float** out) {
  if (device < 0 || device >= kMaxDevices) return fail(RNNWF_EINVAL, "rnnwf_logpsi_static: device index out of range");
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_addr[device] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || p == nullptr)
      return fail(RNNWF_ECUDA, std::string("rnnwf_logpsi_static: cudaGetSymbolAddress(g_static_scratch): ") + cudaGetErrorString(e));
    g_static_scratch_addr[device] = static_cast<float*>(p);
  }
  *out = g_static_scratch_addr[device];
  return RNNWF_OK;
}
}  // namespace

extern "C" int rnnwf_logpsi_static(rnnwf_handle h, int64_t rows, const int8_t* samples_dev, double* logp_dev,
                                   double* phase_dev, void* stream) {
  int rc = check_ready(h);
  if (rc) return rc;
  if (rows <= 0 || samples_dev == nullptr || logp_dev == nullptr)
    return fail(RNNWF_EINVAL, "rnnwf_logpsi_static: rows must be positive, samples and logp non-null");
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev != h->device) return fail(RNNWF_EINVAL, "rnnwf_logpsi_static: the handle's device is not the current device");
  const rnnwf_model& m = *h;
  // tallest tile whose per-CTA ring fits; as many CTAs as the fixed array can serve
  const int pad = (m.ncls_max > 1) ? 3 * m.ncls_max : 0;
  LaunchShape s;
  bool found = false, fits_smem = false;
  long long smallest = 0;
  for (int wr = 4; wr >= 1 && !found; wr >>= 1) {
    const int Mp = 32 * wr, Mr = Mp - pad;
    if (Mr <= 0) continue;
    SmemLayout L(m.ncls_max, m.Gblk, Mp, m.ldH, m.nhp, m.ND, m.Tsym);
    if (L.total > 227 * 1024) continue;
    fits_smem = true;
    const long long per_cta = (long long)m.ring_len * Mp * m.nhp * 4;      // FP32 history ring of the FFMA sweep
    smallest = per_cta;
    const long long tiles = (rows + Mr - 1) / Mr;
    if (per_cta > (long long)RNNWF_STATIC_SCRATCH_BYTES) continue;
    if (wr == 1 || tiles >= m.num_sms) {
      s.WR = wr; s.Mp = Mp; s.Mr = Mr; s.threads = 32 * wr * m.WC; s.smem = L.total;
      const long long cap = (long long)RNNWF_STATIC_SCRATCH_BYTES / per_cta;
      s.grid = (int)std::max(1LL, std::min<long long>(std::min<long long>(tiles, m.num_sms), cap));
      found = true;
    }
  }
  if (!fits_smem) return fail(RNNWF_ENOMEM, "rnnwf_logpsi_static: model does not fit in shared memory (lattice too long / too many units)");
  if (!found) {
    char msg[256];
    std::snprintf(msg, sizeof(msg),
                  "rnnwf_logpsi_static: this model needs %lld bytes of history ring per thread block, the built-in scratch "
                  "holds %u (RNNWF_STATIC_SCRATCH_BYTES); use rnnwf_logpsi with a caller workspace",
                  smallest, (unsigned)RNNWF_STATIC_SCRATCH_BYTES);
    return fail(RNNWF_ENOMEM, msg);
  }
  float* ring = nullptr;
  if ((rc = static_scratch_address(dev, &ring))) return rc;
  SweepParams P;
  fill_model_params(m, s, P);
  P.mode = MODE_EVAL; P.rows = rows; P.cfg_in = samples_dev; P.out_logp = logp_dev; P.out_phase = phase_dev;
  P.ring = ring;
  // always the generic FP32 FFMA sweep, whatever the regime of the handle (the other sweeps size their scratch differently)
  if (h->ffma_stale) pack_ffma(h, h->d_params, static_cast<cudaStream_t>(stream));
  return launch_sweep_ffma(m, P, s, static_cast<cudaStream_t>(stream));
}
