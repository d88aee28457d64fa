// Instances and launchers of the dedicated 1D-GRU kernels (gru1d.cuh), a translation unit of their own.
// `rg`: rows per thread (G1_RG = 8: tiles of 64 rows; G1_RG_SMALL = 2: tiles of 16 rows for small batches).
#include "gru1d.cuh"
#include "gru1d_bwd.cuh"

namespace rnnwf {

namespace {
using G1Kernel = void (*)(const SweepParams);
template <int RG>
G1Kernel g1_pick(int mode) {
  if (mode == MODE_SAMPLE) return gru1d_sweep_kernel<MODE_SAMPLE, RG>;
  if (mode == MODE_CHILD) return gru1d_sweep_kernel<MODE_CHILD, RG>;
  return gru1d_sweep_kernel<MODE_EVAL, RG>;
}
G1Kernel g1_kernel(int mode, int rg) { return rg == 4 ? g1_pick<4>(mode) : rg == G1_RG_SMALL ? g1_pick<G1_RG_SMALL>(mode) : g1_pick<G1_RG>(mode); }
}  // namespace

// (the wide instance needs the most shared memory; the attribute is set to each instance's own size)
cudaError_t g1_set_smem(int nh, int L, int N) {
  for (int rg : {G1_RG, G1_RG_SMALL, 4})
    for (int mode = 0; mode < 3; ++mode) {
      cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(g1_kernel(mode, rg)), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           G1Layout(nh, L, N, 8 * rg).total_bytes);
      if (e != cudaSuccess) return e;
    }
  return cudaSuccess;
}

int g1_occupancy(int threads, int smem_bytes) {
  int occ = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gru1d_sweep_kernel<MODE_CHILD, G1_RG>, threads, smem_bytes) != cudaSuccess) { occ = 1; cudaGetLastError(); }
  return occ < 1 ? 1 : occ;
}

cudaError_t g1_sweep_launch(const SweepParams& P, int rg, int grid, int threads, cudaStream_t st) {
  const int smem_bytes = G1Layout(P.nh, P.symL, P.Tsym, 8 * rg).total_bytes;
  g1_kernel(P.mode, rg)<<<grid, threads, smem_bytes, st>>>(P);
  return cudaGetLastError();
}

cudaError_t g1_pack_launch(const G1PackSrc& S, int nh, int L, int N, int D, int ND, float* out, cudaStream_t st) {
  const G1Layout Y(nh, L, N);
  gru1d_pack_kernel<0><<<(Y.pack_floats + 255) / 256, 256, 0, st>>>(S, nh, L, N, D, ND, out);
  return cudaGetLastError();
}

cudaError_t g1_bwd_set_smem(int nh, int L) {
  cudaError_t e = cudaFuncSetAttribute(gru1d_bwd_kernel<G1_RG>, cudaFuncAttributeMaxDynamicSharedMemorySize, G1BwdLayout(nh, L, 8 * G1_RG).total_bytes);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(gru1d_bwd_kernel<G1_RG_SMALL>, cudaFuncAttributeMaxDynamicSharedMemorySize, G1BwdLayout(nh, L, 8 * G1_RG_SMALL).total_bytes);
  return e;
}

cudaError_t g1_bwd_launch(const G1BwdParams& B, int rg, int grid, cudaStream_t st) {
  const G1BwdLayout Y(B.nh, B.L, 8 * rg);
  if (rg == G1_RG_SMALL) gru1d_bwd_kernel<G1_RG_SMALL><<<grid < 1 ? 1 : grid, g1_threads(B.nh), Y.total_bytes, st>>>(B);
  else gru1d_bwd_kernel<G1_RG><<<grid < 1 ? 1 : grid, g1_threads(B.nh), Y.total_bytes, st>>>(B);
  return cudaGetLastError();
}

cudaError_t g1_pack_bwd_launch(const G1PackSrc& S, int nh, int L, int D, int ND, float* out, cudaStream_t st) {
  const G1BwdLayout Y(nh, L);
  gru1d_pack_bwd_kernel<0><<<(Y.pack_floats + 255) / 256, 256, 0, st>>>(S, nh, L, D, ND, out);
  return cudaGetLastError();
}

}  // namespace rnnwf
