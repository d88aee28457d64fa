// Local-energy front end: diagonal matrix elements and on-device enumeration of the connected
// configurations (never materialised: a child is its parent row plus one or two toggled symbols).
// Children are emitted ordered by the generation step of their first changed symbol so that a tile of
// the sweep can restart from the parent's saved hidden states at that step (prefix reuse).
//
// Reference behaviour restated (file:line under the reference root):
//   E1 1DTFIM/TrainingRNN_1DTFIM.py:13-75            E2 2DTFIM_2DRNN/Training2DRNN_2DTFIM.py:13-83
//   E3 J1J2/TrainingRNN_J1J2.py:12-127,277-279        E4 2DHeisenberg/Helper_Functions.py:19-177
//   E5 RydbergKagome_2DGRU/Helping_Functions.py:33-153
#pragma once
#include "common.cuh"

namespace rnnwf {

// An exchange bond between symbols ta < tb (generation order): S+S- element `offdiag` when the two
// spins are antiparallel, Sz Sz element +-`diag` (parallel +, antiparallel -).
struct __align__(16) Bond { int16_t ta, tb; float offdiag; float diag; int32_t pad; };

struct __align__(16) RydNb { int16_t kc, di, dj, k; double w; };  // (V/2)/d^6 pair term of atom kc of a cell

__device__ __forceinline__ int cfg_spin(const int8_t* row, int site) { return row[site] & 1; }

// ---- TFIM diagonal parts -------------------------------------------------------------------------
// Jz is indexed like the reference: 1D Jz[i] couples (i,i+1); 2D Jz[i*Ny+j] couples (i,j)-(i+1,j) and (i,j)-(i,j+1).
__global__ void tfim_diag_kernel(const int8_t* __restrict__ cfg, long long ns, int Nx, int Ny, const double* __restrict__ Jz,
                                 double* __restrict__ eloc) {
  for (long long n = blockIdx.x * (long long)blockDim.x + threadIdx.x; n < ns; n += (long long)gridDim.x * blockDim.x) {
    const int8_t* row = cfg + n * Nx * Ny;
    double e = 0.0;
    for (int i = 0; i < Nx; ++i)
      for (int j = 0; j < Ny; ++j) {
        const int s = 2 * cfg_spin(row, i * Ny + j) - 1;
        if (i + 1 < Nx) e -= Jz[i * Ny + j] * (double)(s * (2 * cfg_spin(row, (i + 1) * Ny + j) - 1));
        if (j + 1 < Ny) e -= Jz[i * Ny + j] * (double)(s * (2 * cfg_spin(row, i * Ny + j + 1) - 1));
      }
    eloc[2 * n] = e;
    eloc[2 * n + 1] = 0.0;
  }
}

// every symbol-bit of every sample flipped once: child index (t*bits + k)*ns + n
__global__ void all_flips_kernel(long long ns, int Tsym, int bits, float coef, Child* __restrict__ children, long long* __restrict__ nconn) {
  const long long total = (long long)Tsym * bits * ns;
  if (nconn != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *nconn = total;     // number of connected configurations
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const long long n = idx % ns;
    const int tk = (int)(idx / ns);
    Child c;
    c.parent = (int)n; c.ta = (int16_t)(tk / bits); c.tb = -1; c.coef = coef; c.ka = (uint8_t)(tk % bits); c.kb = 0; c.pad = 0;
    children[idx] = c;
  }
}

// ---- exchange Hamiltonians (J1-J2 chain, square / triangular Heisenberg) ---------------------------
// site_of_t maps a symbol index to its offset inside a configuration row.
// counts[ts*ns + n] = number of non-zero exchanges of sample n whose first changed symbol is ts; the threads of ts == 0 also
// take the diagonal matrix element of their sample along (bond_diag_kernel's arithmetic; eloc == nullptr: not wanted)
__global__ void bond_count_kernel(const int8_t* __restrict__ cfg, long long ns, int nsites, int Tsym,
                                  const int16_t* __restrict__ site_of_t, const Bond* __restrict__ bonds,
                                  const int* __restrict__ bond_start, int* __restrict__ counts,
                                  int nbonds, const double* __restrict__ Bz, int nBz, double* __restrict__ eloc) {
  const long long total = (long long)Tsym * ns;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx <= total; idx += (long long)gridDim.x * blockDim.x) {
    if (idx == total) { counts[idx] = 0; continue; }
    const long long n = idx % ns;
    const int ts = (int)(idx / ns);
    const int8_t* row = cfg + n * nsites;
    if (ts == 0 && eloc != nullptr) {
      double e = 0.0;
      for (int b = 0; b < nbonds; ++b) {
        const Bond bd = bonds[b];
        const bool anti = cfg_spin(row, site_of_t[bd.ta]) != cfg_spin(row, site_of_t[bd.tb]);
        e += anti ? -(double)bd.diag : (double)bd.diag;
      }
      for (int i = 0; i < nBz; ++i) e += ((double)cfg_spin(row, i) - 0.5) * Bz[i];
      eloc[2 * n] = e;
      eloc[2 * n + 1] = 0.0;
    }
    const int sa = cfg_spin(row, site_of_t[ts]);
    int c = 0;
    for (int b = bond_start[ts]; b < bond_start[ts + 1]; ++b)
      if (bonds[b].offdiag != 0.f && sa != cfg_spin(row, site_of_t[bonds[b].tb])) ++c;
    counts[idx] = c;
  }
}

__global__ void bond_fill_kernel(const int8_t* __restrict__ cfg, long long ns, int nsites, int Tsym,
                                 const int16_t* __restrict__ site_of_t, const Bond* __restrict__ bonds,
                                 const int* __restrict__ bond_start, const int* __restrict__ offsets,
                                 Child* __restrict__ children, long long* __restrict__ nconn) {
  const long long total = (long long)Tsym * ns;
  if (nconn != nullptr && blockIdx.x == 0 && threadIdx.x == 0) *nconn = (long long)offsets[total];   // number of connected configurations
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const long long n = idx % ns;
    const int ts = (int)(idx / ns);
    const int8_t* row = cfg + n * nsites;
    const int sa = cfg_spin(row, site_of_t[ts]);
    int o = offsets[idx];
    for (int b = bond_start[ts]; b < bond_start[ts + 1]; ++b) {
      const Bond bd = bonds[b];
      if (bd.offdiag != 0.f && sa != cfg_spin(row, site_of_t[bd.tb])) {
        Child c;
        c.parent = (int)n; c.ta = bd.ta; c.tb = bd.tb; c.coef = bd.offdiag; c.ka = 0; c.kb = 0; c.pad = 0;
        children[o++] = c;
      }
    }
  }
}

// ---- Rydberg diagonal part -------------------------------------------------------------------------
// one warp per sample: lanes over (unit cell, neighbour-table entry) pairs, float64 shuffle reduction
__global__ void rydberg_diag_kernel(const int8_t* __restrict__ cfg, long long ns, int Lx, int Ly, const RydNb* __restrict__ nb,
                                    int nnb, double delta, double* __restrict__ eloc) {
  const int lane = threadIdx.x & 31;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int cells = Lx * Ly;
  for (long long n = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; n < ns; n += nwarps) {
    const int8_t* row = cfg + n * cells * 3;
    double e = 0.0;
    for (int c = lane; c < cells; c += 32) {
      const int8_t* cell = row + c * 3;
      e -= delta * (double)((cell[0] & 1) + (cell[1] & 1) + (cell[2] & 1));
    }
    for (int idx = lane; idx < cells * nnb; idx += 32) {
      const int c = idx / nnb, m = idx - c * nnb;
      const int i = c / Ly, j = c - i * Ly;
      const RydNb q = nb[m];
      if (row[c * 3 + q.kc] & 1) {
        const int ii = ((i + q.di) % Lx + Lx) % Lx, jj = ((j + q.dj) % Ly + Ly) % Ly;
        if (row[(ii * Ly + jj) * 3 + q.k] & 1) e += (double)q.w;
      }
    }
    for (int o = 16; o > 0; o >>= 1) e += __shfl_xor_sync(0xffffffffu, e, o);
    if (lane == 0) {
      eloc[2 * n] = e;
      eloc[2 * n + 1] = 0.0;
    }
  }
}

}  // namespace rnnwf
