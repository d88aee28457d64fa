// Instances of the four-neighbour GRU's tensor-core sweep (sweep_tc4.cuh) and its weight-image pack kernel, in a
// translation unit of their own so that they build in parallel with the rest of librnnwf.so.
#include "sweep_tc4.cuh"

namespace rnnwf {

namespace {
using T4Kernel = void (*)(const SweepParams);

T4Kernel t4_kernel(int mode, bool split) {
  if (mode == MODE_SAMPLE) return split ? sweep_tc4_kernel<MODE_SAMPLE, true> : sweep_tc4_kernel<MODE_SAMPLE, false>;
  if (mode == MODE_CHILD) return split ? sweep_tc4_kernel<MODE_CHILD, true> : sweep_tc4_kernel<MODE_CHILD, false>;
  return split ? sweep_tc4_kernel<MODE_EVAL, true> : sweep_tc4_kernel<MODE_EVAL, false>;
}
}  // namespace

cudaError_t tc4_sweep_set_smem(bool split, int smem_bytes) {
  for (int mode = 0; mode < 3; ++mode) {
    cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(t4_kernel(mode, split)), cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t tc4_sweep_launch(const SweepParams& P, bool split, int grid, int threads, int smem_bytes, cudaStream_t st) {
  t4_kernel(P.mode, split)<<<grid, threads, smem_bytes, st>>>(P);
  return cudaGetLastError();
}

cudaError_t tc4_pack_launch(const float* W, const float* Wg, const float* Wmerge, const float* b, const float* bg, int nh, int NP, bool split,
                            void* out, cudaStream_t st) {
  const long long n = (long long)NP * (t4_pass_bytes(NP, split) / 2);
  const int blocks = (int)std::min<long long>(2048, (n + 255) / 256);
  pack_tc4_kernel<0><<<blocks, 256, 0, st>>>(W, Wg, Wmerge, b, bg, nh, NP, split ? 1 : 0, reinterpret_cast<__half*>(out));
  return cudaGetLastError();
}

}  // namespace rnnwf
