// Microbenchmark: per-SM throughput of cp.async.bulk (global -> shared, mbarrier complete_tx) from L2-resident data,
// as a function of request size and requests in flight.  One CTA per SM, one issuing thread.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/bulk_rate tools/ubench/bulk_rate.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1) k(const unsigned char* src, size_t src_bytes, int req_bytes, int depth, int iters, long long* out, int nissue) {
  extern __shared__ __align__(128) unsigned char sm[];
  __shared__ __align__(8) uint64_t bar[32];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 32; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(bar + i)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if ((threadIdx.x & 31) == 0 && (threadIdx.x >> 5) < nissue) {
    const int w = threadIdx.x >> 5;
    uint64_t* bar_w = bar + 8 * w;
    unsigned char* sm_w = sm + (size_t)w * depth * req_bytes;
    const unsigned char* base = src + ((size_t)blockIdx.x * 262144) % (src_bytes - 262144);
    long long t0 = clock64();
    for (int i = 0; i < iters + depth; ++i) {
      const int s = i % depth;
      if (i >= depth) {
        const uint32_t par = ((i / depth) - 1) & 1;
        uint32_t ok = 0;
        while (!ok) asm volatile("{.reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0,1,0,p;}" : "=r"(ok) : "r"(s32(bar_w + s)), "r"(par) : "memory");
      }
      if (i < iters) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar_w + s)), "r"(req_bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(s32(sm_w + (size_t)s * req_bytes)), "l"(base + ((size_t)i * req_bytes) % 196608), "r"(req_bytes), "r"(s32(bar_w + s)) : "memory");
      }
    }
    if (w == 0) out[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  const size_t src_bytes = 64u << 20;
  unsigned char* src; long long* out;
  cudaMalloc(&src, src_bytes); cudaMemset(src, 1, src_bytes); cudaMalloc(&out, 148 * 8);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  const int sizes[] = {4096, 8192, 16384, 24576, 49152};
  for (int grid : {1, 148}) for (int nissue : {1, 2, 4}) for (int sz : sizes) for (int depth : {2}) {
    if ((size_t)sz * depth * nissue > 196608) continue;
    const int iters = 2000;
    k<<<grid, 128, sz * depth * nissue>>>(src, src_bytes, sz, depth, iters, out, nissue);
    k<<<grid, 128, sz * depth * nissue>>>(src, src_bytes, sz, depth, iters, out, nissue);
    cudaError_t e = cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, out, grid * 8, cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < grid; ++i) avg += h[i]; avg /= grid;
    printf("grid %3d  issuers %d  req %6d B  in flight %d each : %.1f B/clk/SM  (%.0f cycles per request per issuer) %s\n", grid, nissue, sz, depth, (double)sz * iters * nissue / avg, avg / iters, cudaGetErrorString(e));
  }
  return 0;
}
