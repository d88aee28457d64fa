// Host-side model description: lattice sweep tables, tile shape, parameter layout.
#pragma once
#include <string>
#include <vector>

#include "common.cuh"
#include "../../include/rnnwf.h"

namespace rnnwf {

struct TensorInfo {
  std::string name;
  long long offset;
  int ndim;
  long long shape[4];
  long long numel() const { long long n = 1; for (int i = 0; i < ndim; ++i) n *= shape[i]; return n; }
};

struct HostVariant {
  VariantMeta meta;
  int vkind;       // tensorized: 0 both, 1 horizontal only, 2 vertical only, 3 none; GRU1D: layer
  int nsl;         // packed slots
  int slot_of_rank[MAX_SLOTS];
  long long w_float4;  // packed size of this variant (all column blocks)
};

struct Saved {   // byte offsets inside a caller-provided saved-state buffer for `rows` rows
  long long off_hist, off_gates, off_cum, off_himg, off_head, total;
};

struct rnnwf_model_impl {
  rnnwf_config cfg;
  int device;
  int num_sms;
  // derived
  int cell, nh, nhp, ldH, nslots, D, ND, complex_head, u1, cap_up, cap_dn;
  int Q, Tsym, symL, nsites, bits, cps;
  int TJ, NTU, WC, Gblk, ncb, Gp, Greal, ncls_max, ring_len;
  bool gru_type;   // cells whose backward needs saved gate activations
  std::vector<StepMeta> steps;
  std::vector<HostVariant> variants;
  std::vector<int16_t> site_of_t, t_of_site;
  std::vector<TensorInfo> tensors;
  long long num_params;
  long long wpack_float4, xtab_float4, bias_float4;
  // device buffers owned by the handle
  StepMeta* d_steps = nullptr;
  VariantMeta* d_variants = nullptr;
  int16_t* d_site_of_t = nullptr;
  float4* d_wpack = nullptr;
  float4* d_xtab = nullptr;
  float4* d_bias = nullptr;
  float* d_head_w = nullptr;
  float* d_head_b = nullptr;
  // tensor-core regime
  int precision = 0;            // 0: FP32 FFMA sweep, 1: tcgen05 sweep (FP16 operands, FP32 accumulate in TMEM), 2: tcgen05 sweep, split precision
  bool tc_enabled = false;
  bool g1 = false;              // dedicated small-model 1D GRU kernels (gru1d.cuh)
  int g1_occ = 1;
  float* d_g1pack = nullptr;
  bool g1_bwd = false;          // ... and their one-launch reverse sweep (gru1d_bwd.cuh)
  float* d_g1packT = nullptr;
  bool tc_split = false;        // precision 2: split-precision operands (FP16 hi + lo, three MMAs per K step) -- the 1e-5 bound on the tensor cores
  bool tc4_split = false;       // ... in the split-precision regime (precision tc3)
  bool tc4 = false;             // tensor-core sweep of the four-neighbour GRU (sweep_tc4.cuh); saved state and backward as in the FFMA regime
  int tc_np = 0;
  unsigned char* d_tcB = nullptr;
  float* d_tcbias = nullptr;
  bool van2 = false;            // vanilla 2D RNN, split precision: two-tile connected-configuration sweep (sweep_van2.cuh)
  unsigned char* d_van2B = nullptr;
  float* d_wimgT = nullptr;     // tensorized cells: the reverse sweep's transposed TF32 weight image (bwd_tc.cuh), packed with the forward images
  // weight-gradient kernel work plan (dw_tc.cuh), cached on the device for the last row count it was built for
  mutable long long dw_rows = -1;
  mutable int dw_njobs = 0;
  mutable int dw_cls_count[3] = {0, 0, 0};
  void* d_dw_jobs = nullptr;        // [1024] DwJob
  int* d_dw_cls = nullptr;          // [3][Q]
  // work plan of the four-neighbour GRU's weight-gradient kernel (bwd_tc4.cuh), cached per row count
  mutable long long dw4_rows = -1;
  mutable int dw4_njobs = 0;
  mutable int dw4_sites_n[5] = {0, 0, 0, 0, 0};
  void* d_dw4_jobs = nullptr;       // [4096] Dw4Job
  int* d_dw4_sites = nullptr;       // [5][Q]
  mutable std::vector<unsigned char> plan_host_a, plan_host_b;   // host images of the last uploaded work plan (source of the stream-ordered copies)
  const float* d_params = nullptr;  // last parameter vector handed to set_weights (caller-owned)
  // Hamiltonian tables of the local-energy front end, cached per `ham` block (rnnwf.cu: enumerate_connected)
  void* d_hamtab = nullptr;
  mutable int ham_kind = -1;
  mutable std::vector<double> ham_block;
  mutable std::vector<unsigned char> ham_host[3];
  mutable int ham_n[2] = {0, 0};
  bool weights_set = false;
  bool ffma_stale = false;          // the generic FFMA packs lag behind d_params (repacked on demand by rnnwf_logpsi_static)

  const TensorInfo* tensor(const std::string& n) const {
    for (auto& t : tensors) if (t.name == n) return &t;
    return nullptr;
  }
  Saved saved_layout(long long rows) const {
    auto al = [](long long x) { return (x + 255) & ~255LL; };
    Saved s;
    s.off_hist = 0;
    s.off_gates = al((long long)Q * rows * nhp * 4);
    // tensor-core regime: gates are saved tile-blocked, three float4 per (row, 4 units) (bwd_tc.cuh: tc_gate_index)
    const bool tcs = tc_enabled && (cell == CELL_TENSGRU2D || cell == CELL_TENS2D);   // tile-blocked (u, h~, m); the vanilla cell saves none
    const long long gate_bytes = tcs ? (long long)Q * ((rows + 127) / 128) * 4 * tc_np * 3 * 128 * 16
                                     : (tc4 ? (long long)Q * ((rows + 127) / 128) * Gp * 128 * 16      // tile-blocked [cell][tile][unit][row 128]
                                            : (gru_type ? (long long)Q * rows * Gp * 16 : 0));
    s.off_cum = s.off_gates + al(gate_bytes);
    s.off_himg = s.off_cum + al((long long)Tsym * rows * 16);
    // tensor-core regime: hidden states once more as ready-made B operand tiles of the weight-gradient kernel
    const long long himg_bytes = tc_enabled ? (long long)Q * ((rows + 127) / 128) * (32LL * (((nhp + 15) / 16) * 2) * 32) * 4 : 0;
    // tensor-core regime: the heads' outputs of every (site, row) -- (p0, p1, phase pre-activation of the symbol, 0) -- so that
    // the reverse sweep forms d(logits) itself instead of recomputing the heads from the hidden states
    s.off_head = s.off_himg + al(himg_bytes);
    s.total = s.off_head + al(tc_enabled ? (long long)Q * rows * 16 : 0);
    return s;
  }
};

}  // namespace rnnwf
