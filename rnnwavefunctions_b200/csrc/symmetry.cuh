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

// forward: out_logp[b] = log(mean_g exp(logp_g)),  out_phase[b] = atan2(Im, reg(Re))   (phase arrays may be null)
__global__ void __launch_bounds__(256) symmetry_epilogue_kernel(int G, long long rows, const double* __restrict__ logp,
                                                                const double* __restrict__ phase, const SymChars ch,
                                                                double* __restrict__ out_logp, double* __restrict__ out_phase) {
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
    out_logp[b] = mx + log(s / (double)G);
    if (phase != nullptr && out_phase != nullptr) {
      const double r = (re >= 0.0 ? 1.0 : -1.0) * sqrt(re * re + 1e-8);
      out_phase[b] = atan2(im, r);
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
