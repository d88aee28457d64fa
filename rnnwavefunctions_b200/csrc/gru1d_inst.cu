// Instances and launchers of the dedicated 1D-GRU kernels (gru1d.cuh), a translation unit of their own.
#include "gru1d.cuh"
#include "gru1d_bwd.cuh"

namespace rnnwf {

cudaError_t g1_set_smem(int smem_bytes) {
  cudaError_t e = cudaFuncSetAttribute(gru1d_sweep_kernel<MODE_SAMPLE>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gru1d_sweep_kernel<MODE_EVAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(gru1d_sweep_kernel<MODE_CHILD>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  return e;
}

int g1_occupancy(int threads, int smem_bytes) {
  int occ = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, gru1d_sweep_kernel<MODE_CHILD>, threads, smem_bytes) != cudaSuccess) { occ = 1; cudaGetLastError(); }
  return occ < 1 ? 1 : occ;
}

cudaError_t g1_sweep_launch(const SweepParams& P, int grid, int threads, int smem_bytes, cudaStream_t st) {
  if (P.mode == MODE_SAMPLE) gru1d_sweep_kernel<MODE_SAMPLE><<<grid, threads, smem_bytes, st>>>(P);
  else if (P.mode == MODE_CHILD) gru1d_sweep_kernel<MODE_CHILD><<<grid, threads, smem_bytes, st>>>(P);
  else gru1d_sweep_kernel<MODE_EVAL><<<grid, threads, smem_bytes, st>>>(P);
  return cudaGetLastError();
}

cudaError_t g1_pack_launch(const G1PackSrc& S, int nh, int L, int N, int D, int ND, float* out, cudaStream_t st) {
  const G1Layout Y(nh, L, N);
  gru1d_pack_kernel<0><<<(Y.pack_floats + 255) / 256, 256, 0, st>>>(S, nh, L, N, D, ND, out);
  return cudaGetLastError();
}

cudaError_t g1_bwd_set_smem(int smem_bytes) {
  return cudaFuncSetAttribute(gru1d_bwd_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
}

cudaError_t g1_bwd_launch(const G1BwdParams& B, int grid, cudaStream_t st) {
  const G1BwdLayout Y(B.nh, B.L);
  gru1d_bwd_kernel<0><<<grid < 1 ? 1 : grid, g1_threads(B.nh), Y.total_bytes, st>>>(B);
  return cudaGetLastError();
}

cudaError_t g1_pack_bwd_launch(const G1PackSrc& S, int nh, int L, int D, int ND, float* out, cudaStream_t st) {
  const G1BwdLayout Y(nh, L);
  gru1d_pack_bwd_kernel<0><<<(Y.pack_floats + 255) / 256, 256, 0, st>>>(S, nh, L, D, ND, out);
  return cudaGetLastError();
}

}  // namespace rnnwf
