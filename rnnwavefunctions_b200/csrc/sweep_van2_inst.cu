// Instances and launchers of the two-tile connected-configuration sweep of the vanilla 2D RNN (sweep_van2.cuh).
#include "sweep_van2.cuh"

namespace rnnwf {

namespace {
using V2Kernel = void (*)(const SweepParams);
V2Kernel v2_kernel(int nhp) {
  switch (nhp) {
    case 32: return sweep_van2_kernel<4>;
    case 48: return sweep_van2_kernel<6>;
    case 56: return sweep_van2_kernel<7>;
    case 64: return sweep_van2_kernel<8>;
  }
  return nullptr;
}
}  // namespace

cudaError_t van2_set_smem(int nhp, int Q) {
  V2Kernel k = v2_kernel(nhp);
  if (k == nullptr) return cudaErrorInvalidValue;
  return cudaFuncSetAttribute(reinterpret_cast<const void*>(k), cudaFuncAttributeMaxDynamicSharedMemorySize, V2Smem(nhp, Q).total);
}

cudaError_t van2_launch(const SweepParams& P, int grid, cudaStream_t st) {
  V2Kernel k = v2_kernel(P.nhp);
  if (k == nullptr) return cudaErrorInvalidValue;
  k<<<grid, V2_THREADS, V2Smem(P.nhp, P.Q).total, st>>>(P);
  return cudaGetLastError();
}

cudaError_t van2_pack_launch(const float* Wh, const float* Wv, int nh, int nhp, void* out, cudaStream_t st) {
  pack_van2_kernel<0><<<64, 256, 0, st>>>(Wh, Wv, nh, nhp, reinterpret_cast<__half*>(out));
  return cudaGetLastError();
}

}  // namespace rnnwf
