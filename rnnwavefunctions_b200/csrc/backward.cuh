// Backward of log psi (BPTT over the lattice sweep) -- placeholder until the kernels land.
#pragma once
#include <string>
#include "model.h"

namespace rnnwf {
inline cudaError_t backward_init() { return cudaSuccess; }
inline long long backward_workspace_bytes(const rnnwf_model_impl&, long long) { return 256; }
inline int backward_run(const rnnwf_model_impl&, long long, const int8_t*, const void*, const double*, const double*,
                        float*, void*, cudaStream_t, std::string& err) {
  err = "rnnwf_backward: not implemented yet";
  return RNNWF_ESTATE;
}
}  // namespace rnnwf
