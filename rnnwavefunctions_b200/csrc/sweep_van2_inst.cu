// Instances and launchers of the two-tile connected-configuration sweep of the vanilla 2D RNN (sweep_van2.cuh).
#include "sweep_van2.cuh"

namespace rnnwf {

namespace {
using V2Kernel = void (*)(const SweepParams);
V2Kernel v2_kernel(int nhp, int mode) {
#define V2_PICK(N) (mode == MODE_SAMPLE ? sweep_van2_kernel<N, MODE_SAMPLE> : (mode == MODE_EVAL ? sweep_van2_kernel<N, MODE_EVAL> : sweep_van2_kernel<N, MODE_CHILD>))
  switch (nhp) {
    case 8: return V2_PICK(1);
    case 16: return V2_PICK(2);
    case 24: return V2_PICK(3);
    case 32: return V2_PICK(4);
    case 40: return V2_PICK(5);
    case 48: return V2_PICK(6);
    case 56: return V2_PICK(7);
    case 64: return V2_PICK(8);
  }
#undef V2_PICK
  return nullptr;
}
}  // namespace

cudaError_t van2_set_smem(int nhp, int Q) {
  for (int mode : {MODE_CHILD, MODE_SAMPLE, MODE_EVAL}) {
    V2Kernel k = v2_kernel(nhp, mode);
    if (k == nullptr) return cudaErrorInvalidValue;
    const cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(k), cudaFuncAttributeMaxDynamicSharedMemorySize, V2Smem(nhp, Q).total);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t van2_launch(const SweepParams& P, int grid, cudaStream_t st) {
  V2Kernel k = v2_kernel(P.nhp, P.mode);
  if (k == nullptr) return cudaErrorInvalidValue;
  k<<<grid, V2_THREADS, V2Smem(P.nhp, P.Q).total, st>>>(P);
  return cudaGetLastError();
}

cudaError_t van2_pack_launch(const float* Wh, const float* Wv, int nh, int nhp, void* out, cudaStream_t st) {
  pack_van2_kernel<0><<<64, 256, 0, st>>>(Wh, Wv, nh, nhp, reinterpret_cast<__half*>(out));
  return cudaGetLastError();
}

}  // namespace rnnwf
