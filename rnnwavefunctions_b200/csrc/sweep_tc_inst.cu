// Instances of the tensor-core sweep kernel (sweep_tc_kernel.cuh), in a translation unit of their own so that they build in
// parallel with the rest of librnnwf.so.  rnnwf.cu reaches them through the two functions below.
#include "sweep_tc_kernel.cuh"

namespace rnnwf {

namespace {
using TcKernel = void (*)(const SweepParams);

template <int MODE>
TcKernel tc_pick(int cellk, bool biasmma, bool split) {
  if (cellk == 1) return split ? sweep_tc_kernel<MODE, false, 1, true> : sweep_tc_kernel<MODE, false, 1, false>;
  if (cellk == 2) {
    if (split) return biasmma ? sweep_tc_kernel<MODE, true, 2, true> : sweep_tc_kernel<MODE, false, 2, true>;
    return biasmma ? sweep_tc_kernel<MODE, true, 2, false> : sweep_tc_kernel<MODE, false, 2, false>;
  }
  if (split) return biasmma ? sweep_tc_kernel<MODE, true, 0, true> : sweep_tc_kernel<MODE, false, 0, true>;
  return biasmma ? sweep_tc_kernel<MODE, true, 0, false> : sweep_tc_kernel<MODE, false, 0, false>;
}

TcKernel tc_kernel(int mode, int cellk, bool biasmma, bool split) {
  if (mode == MODE_SAMPLE) return tc_pick<MODE_SAMPLE>(cellk, biasmma, split);
  if (mode == MODE_CHILD) return tc_pick<MODE_CHILD>(cellk, biasmma, split);
  return tc_pick<MODE_EVAL>(cellk, biasmma, split);
}
}  // namespace

// cellk: 0 tensorized GRU, 1 vanilla 2D RNN, 2 un-gated tensorized cell; biasmma: biases ride on the operand's constant-1
// position; split: split-precision regime.
cudaError_t tc_sweep_set_smem(int cellk, bool biasmma, bool split, int smem_bytes) {
  for (int mode = 0; mode < 3; ++mode) {
    cudaError_t e = cudaFuncSetAttribute(reinterpret_cast<const void*>(tc_kernel(mode, cellk, biasmma, split)),
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t tc_sweep_launch(const SweepParams& P, int cellk, bool biasmma, bool split, int grid, int threads, int smem_bytes, cudaStream_t st) {
  tc_kernel(P.mode, cellk, biasmma, split)<<<grid, threads, smem_bytes, st>>>(P);
  return cudaGetLastError();
}

}  // namespace rnnwf
