// How much does ONE bulk-copy request cost?  One CTA per SM streams a private slice of a large buffer through a ring of
// shared-memory stages; a stage of STAGE_BYTES is fetched as `nreq` requests of STAGE_BYTES / nreq bytes each (one request
// per lane, as the skinning kernels issue them).  Same bytes, same ring depth: only the request size changes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/tma_probe tools/tma_probe.cu && tools/tma_probe
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(n)); }
__device__ __forceinline__ void expect_tx(uint64_t* b, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void wait(uint64_t* b, uint32_t ph) {
  asm volatile("{ .reg .pred p; W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1; @!p bra W; }" ::"r"(s32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void g2s(void* d, const void* g, uint32_t bytes, uint64_t* b) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(s32(d)), "l"(g),
               "r"(bytes), "r"(s32(b)) : "memory");
}

constexpr int STAGE_BYTES = 24576;
template <int STAGES>
__global__ void __launch_bounds__(32, 1) probe(const char* src, size_t per_cta, int nreq, int iters) {
  extern __shared__ __align__(128) unsigned char sm[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + STAGES * STAGE_BYTES);
  const int lane = threadIdx.x;
  if (lane == 0) {
    for (int s = 0; s < STAGES; ++s) mbar_init(bar + s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  __syncwarp();
  const char* base = src + (size_t)blockIdx.x * per_cta;
  const uint32_t rb = STAGE_BYTES / nreq;
  for (int it = 0; it < iters + STAGES; ++it) {
    const int s = it % STAGES;
    if (it >= STAGES) {          // consume the stage fetched STAGES iterations ago
      wait(bar + s, ((it / STAGES) - 1) & 1);
    }
    __syncwarp();
    if (it < iters) {
      if (lane == 0) expect_tx(bar + s, STAGE_BYTES);
      __syncwarp();
      // requests of one stage come from rows 82 944 B apart (like the v_posed rows of different samples)
      if (lane < nreq) g2s(sm + s * STAGE_BYTES + lane * rb, base + ((size_t)it * nreq + lane) * 82944 % per_cta / 16 * 16, rb, bar + s);
    }
  }
}

int main() {
  const size_t per_cta = (size_t)32 << 20, total = per_cta * 148;
  char* src;
  cudaMalloc(&src, total + (1 << 20));
  cudaMemset(src, 1, total + (1 << 20));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto run = [&](auto kern, int stages, int nreq) {
    const size_t smem = (size_t)stages * STAGE_BYTES + 256;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 4000;
    kern<<<148, 32, smem>>>(src, per_cta, nreq, 200);
    cudaEventRecord(e0);
    kern<<<148, 32, smem>>>(src, per_cta, nreq, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    const double bytes = (double)iters * STAGE_BYTES * 148;
    printf("stages %d (%3d KB in flight)  requests/stage %2d (%5d B each): %7.1f GB/s  %6.0f cycles/stage @1.9GHz  err=%s\n", stages,
           stages * STAGE_BYTES / 1024, nreq, STAGE_BYTES / nreq, bytes / ms * 1e-6, ms * 1e-3 * 1.9e9 / iters,
           cudaGetErrorString(cudaGetLastError()));
  };
  for (int nreq : {1, 2, 4, 8, 16, 32}) run(probe<4>, 4, nreq);
  for (int nreq : {1, 4, 16, 32}) run(probe<8>, 8, nreq);
  for (int nreq : {1, 16}) run(probe<2>, 2, nreq);
  return 0;
}
