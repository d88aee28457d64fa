// Symmetry-projected amplitudes: the epilogue that combines the G transformed copies of every configuration, and its
// backward (the seeds each copy's log P / phase receives).  Reference: 2DHeisenberg/MDComplexRNN.py:290-316
// (log_amplitudes_fromsymmetrygroup: amplitude sqrt(mean_g P_g), phase arg(reg(Re) + i Im) of sum_g e^{i phi_g} / chi_g with
// reg(x) = sign(x) sqrt(x^2 + 1e-8), sign under stop-gradient, :23-25) and RydbergKagome_2DGRU/MDRNNWavefunction.py:242-279
// (log mean_g P_g).  Inputs are copy-major: logp[g * rows + b].
#pragma once
#include "common.cuh"

namespace rnnwf {

constexpr int SYM_MAX_G = 64;
struct SymChars { double inv_chi[SYM_MAX_G]; };   // 1 / chi_g

// log(mean_g exp(logp_g)) and atan2(Im, reg(Re)) of sum_g e^{i phi_g} / chi_g over G entries `stride` apart
__device__ __forceinline__ void sym_combine(int G, const double* __restrict__ logp, const double* __restrict__ phase, long long stride,
                                            const SymChars& ch, double& out_logp, double& out_phase) {
  double mx = -INFINITY;
  for (int g = 0; g < G; ++g) mx = fmax(mx, logp[(long long)g * stride]);
  double s = 0.0, re = 0.0, im = 0.0;
  for (int g = 0; g < G; ++g) {
    s += exp(logp[(long long)g * stride] - mx);
    if (phase != nullptr) {
      double sn, cs;
      sincos(phase[(long long)g * stride], &sn, &cs);
      re += cs * ch.inv_chi[g];
      im += sn * ch.inv_chi[g];
    }
  }
  out_logp = mx + log(s / (double)G);
  out_phase = 0.0;
  if (phase != nullptr) {
    const double r = (re >= 0.0 ? 1.0 : -1.0) * sqrt(re * re + 1e-8);
    out_phase = atan2(im, r);
  }
}

// forward: out_logp[b] = log(mean_g exp(logp_g)),  out_phase[b] = atan2(Im, reg(Re))   (phase arrays may be null)
__global__ void __launch_bounds__(256) symmetry_epilogue_kernel(int G, long long rows, const double* __restrict__ logp,
                                                                const double* __restrict__ phase, const SymChars ch,
                                                                double* __restrict__ out_logp, double* __restrict__ out_phase) {
  for (long long b = blockIdx.x * (long long)blockDim.x + threadIdx.x; b < rows; b += (long long)gridDim.x * blockDim.x) {
    double lp, ph;
    sym_combine(G, logp + b, phase != nullptr ? phase + b : nullptr, rows, ch, lp, ph);
    out_logp[b] = lp;
    if (phase != nullptr && out_phase != nullptr) out_phase[b] = ph;
  }
}

// ---- index-remap tables: the group elements as site permutations (+ a spin flip), applied where the copies are formed ----
// A group element g maps a configuration s to s_g[i] = s[perm[g][i]] ^ flip[g]  (rot90 / reflections / transpose /
// 1 - s of 2DHeisenberg/MDComplexRNN.py:320-577, rotate60 and tf.roll of RydbergKagome_2DGRU/MDRNNWavefunction.py:246-291,
// the chain reversal of 1DTFIM/RNNwavefunction_paritysym.py:125).  Output is copy-major: out[(g * R + r) * nsites + i].
// Row r < rows is sample r; row rows + c is connected configuration c (its parent with one or two symbols toggled) when a
// child list is given.  R = rows + (number of children, read from the device when the list has a variable length).
struct SymExpandParams {
  int G, nsites, bits;
  long long rows, nchild_fixed;
  const int* nchild_dev;
  const int32_t* perm;        // [G][nsites]
  const int8_t* flip;         // [G]
  const int8_t* samples;      // [rows][nsites]
  const Child* children;      // or null
  const int16_t* site_of_t;
  int8_t* out;
  int* total_rows_out;        // optional: G * R for the sweep's device-side row count
};

__global__ void __launch_bounds__(256) sym_expand_kernel(const SymExpandParams P) {
  const long long nchild = (P.children == nullptr) ? 0 : (P.nchild_dev != nullptr ? (long long)(*P.nchild_dev) : P.nchild_fixed);
  const long long R = P.rows + nchild;
  if (P.total_rows_out != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *P.total_rows_out = (int)(R * P.G);
  const long long total = R * P.G * P.nsites;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int i = (int)(idx % P.nsites);
    const long long gr = idx / P.nsites;
    const long long r = gr % R;
    const int g = (int)(gr / R);
    const int j = P.perm[g * P.nsites + i];
    int v;
    if (r < P.rows) {
      v = P.samples[r * P.nsites + j] & 1;
    } else {
      const Child c = P.children[r - P.rows];
      v = P.samples[(long long)c.parent * P.nsites + j] & 1;
      const int oa = (P.bits == 1) ? P.site_of_t[c.ta] : P.site_of_t[c.ta] * 3 + c.ka;
      if (j == oa) v ^= 1;
      if (c.tb >= 0) {
        const int ob = (P.bits == 1) ? P.site_of_t[c.tb] : P.site_of_t[c.tb] * 3 + c.kb;
        if (j == ob) v ^= 1;
      }
    }
    P.out[idx] = (int8_t)(v ^ (P.flip[g] & 1));
  }
}

// symmetrised log psi of the samples themselves (rows [0, rows) of every copy)
__global__ void __launch_bounds__(256) sym_parent_kernel(int G, long long rows, long long nchild_fixed, const int* __restrict__ nchild_dev,
                                                         const double* __restrict__ logp, const double* __restrict__ phase, const SymChars ch,
                                                         double* __restrict__ out_logp, double* __restrict__ out_phase) {
  const long long R = rows + (nchild_dev != nullptr ? (long long)(*nchild_dev) : nchild_fixed);
  for (long long b = blockIdx.x * (long long)blockDim.x + threadIdx.x; b < rows; b += (long long)gridDim.x * blockDim.x) {
    double lp, ph;
    sym_combine(G, logp + b, phase != nullptr ? phase + b : nullptr, R, ch, lp, ph);
    out_logp[b] = lp;
    if (out_phase != nullptr) out_phase[b] = ph;
  }
}

// E_loc[parent] += coef * psi_sym(child) / psi_sym(parent), psi = exp(0.5 logP + i phase)
// (2DHeisenberg/Helper_Functions.py:146-148, 170-175 with the symmetrised log-amplitudes; 1DTFIM/TrainingRNN_1DTFIM.py:70-74)
__global__ void __launch_bounds__(256) sym_eloc_reduce_kernel(int G, long long rows, long long nchild_fixed, const int* __restrict__ nchild_dev,
                                                              const Child* __restrict__ children, const double* __restrict__ logp,
                                                              const double* __restrict__ phase, const SymChars ch,
                                                              const double* __restrict__ par_logp, const double* __restrict__ par_phase,
                                                              double* __restrict__ eloc) {
  const long long nchild = nchild_dev != nullptr ? (long long)(*nchild_dev) : nchild_fixed;
  const long long R = rows + nchild;
  for (long long c = blockIdx.x * (long long)blockDim.x + threadIdx.x; c < nchild; c += (long long)gridDim.x * blockDim.x) {
    const Child chd = children[c];
    double lp, ph;
    sym_combine(G, logp + rows + c, phase != nullptr ? phase + rows + c : nullptr, R, ch, lp, ph);
    const double mag = (double)chd.coef * exp(0.5 * (lp - par_logp[chd.parent]));
    if (phase != nullptr) {
      double sn, cs;
      sincos(ph - par_phase[chd.parent], &sn, &cs);
      atomicAdd(eloc + 2 * (long long)chd.parent, mag * cs);
      atomicAdd(eloc + 2 * (long long)chd.parent + 1, mag * sn);
    } else {
      atomicAdd(eloc + 2 * (long long)chd.parent, mag);
    }
  }
}

// backward: seed_logp_g[b] = s_logp[b] * P_g / sum_g' P_g',
//           seed_phase_g[b] = s_phase[b] * d atan2(Im, reg(Re)) / d phi_g      (sign of reg under stop-gradient)
__global__ void __launch_bounds__(256) symmetry_seeds_kernel(int G, long long rows, const double* __restrict__ logp,
                                                             const double* __restrict__ phase, const SymChars ch,
                                                             const double* __restrict__ s_logp, const double* __restrict__ s_phase,
                                                             double* __restrict__ seed_logp, double* __restrict__ seed_phase) {
  for (long long b = blockIdx.x * (long long)blockDim.x + threadIdx.x; b < rows; b += (long long)gridDim.x * blockDim.x) {
    double mx = -INFINITY;
    for (int g = 0; g < G; ++g) mx = fmax(mx, logp[(long long)g * rows + b]);
    double s = 0.0, re = 0.0, im = 0.0;
    for (int g = 0; g < G; ++g) {
      s += exp(logp[(long long)g * rows + b] - mx);
      if (phase != nullptr) {
        double sn, cs;
        sincos(phase[(long long)g * rows + b], &sn, &cs);
        re += cs * ch.inv_chi[g];
        im += sn * ch.inv_chi[g];
      }
    }
    const double sl = s_logp[b];
    double r = 0.0, dr = 0.0, den = 1.0, sp = 0.0;
    if (phase != nullptr && seed_phase != nullptr) {
      const double mag = sqrt(re * re + 1e-8);
      r = (re >= 0.0 ? 1.0 : -1.0) * mag;
      dr = fabs(re) / mag;                       // d reg / d Re
      den = r * r + im * im;
      sp = (s_phase != nullptr) ? s_phase[b] : 0.0;
    }
    for (int g = 0; g < G; ++g) {
      seed_logp[(long long)g * rows + b] = sl * exp(logp[(long long)g * rows + b] - mx) / s;
      if (phase != nullptr && seed_phase != nullptr) {
        double sn, cs;
        sincos(phase[(long long)g * rows + b], &sn, &cs);
        // d atan2(y, r) = (r dy - y dr) / (r^2 + y^2);  dy/dphi_g = cos / chi,  dRe/dphi_g = -sin / chi
        seed_phase[(long long)g * rows + b] = sp * (r * cs + im * dr * sn) * ch.inv_chi[g] / den;
      }
    }
  }
}

}  // namespace rnnwf
