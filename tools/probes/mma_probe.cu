// Micro-probe (B200): cycles per tcgen05.mma (M=128, K=16, bf16, SS) as a function of N, of whether
// consecutive MMAs accumulate into the same TMEM columns, and of the shared-memory layout.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/probes/mma_probe.bin tools/probes/mma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../low_light_image_enhancement_b200/csrc/sm100_ptx.cuh"
using namespace llie::ptx;

__device__ __forceinline__ void mma_bf16_ss_a(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, int use) {
  if (use == 0)
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, 1, 0;\ntcgen05.mma.cta_group::1.kind::f16.collector::a::fill [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
  else
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, 1, 0;\ntcgen05.mma.cta_group::1.kind::f16.collector::a::lastuse [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc) : "memory");
}
// pairs of MMAs on the SAME A tile: N = n1 (plain or collector fill) then N = n2 (plain or collector lastuse)
__global__ void probe_pair(int n1, int n2, int iters, int reuse, long long* out) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tm;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&tm, 512); tmem_relinquish(); }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tm;
  if (threadIdx.x < 32) {
    long long t0 = 0, t1 = 0;
    if (elect_one()) {
      const uint64_t a0 = smem_desc(smem_u32(smem), 2112, 128);
      const uint64_t b0 = smem_desc(smem_u32(smem + 96 * 1024), 1536, 128);
      const uint32_t i1 = idesc_bf16_m128(n1), i2 = idesc_bf16_m128(n2);
      t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint64_t ad = a0 + (uint64_t)(i % 6);
        const uint64_t bd = b0 + (uint64_t)(((i % 6) * 3072) >> 4);
        if (reuse) { mma_bf16_ss_a(tmem, ad, bd, i1, 0); mma_bf16_ss_a(tmem + 256, ad, bd + 32, i2, 1); }
        else { mma_bf16_ss(tmem, ad, bd, i1, 1u); mma_bf16_ss(tmem + 256, ad, bd + 32, i2, 1u); }
      }
      mma_commit(&bar);
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

// mode bit0: alternate between two accumulators; bit1: 4 accumulators round-robin
__global__ void probe(int n, int iters, int mode, int a_stride16, int b_bytes_step, long long* out, int a_base = 0) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tm;
  for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&tm, 512); tmem_relinquish(); }
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  tc_fence_before(); __syncthreads(); tc_fence_after();
  const uint32_t tmem = tm;
  if (threadIdx.x < 32) {
    long long t0 = 0, t1 = 0;
    if (elect_one()) {
      const uint64_t a0 = smem_desc(smem_u32(smem) + a_base, 2112, 128);
      const uint64_t b0 = smem_desc(smem_u32(smem + 96 * 1024), 1536, 128);
      const uint32_t idesc = idesc_bf16_m128(n);
      const int nacc = (mode & 2) ? 4 : ((mode & 1) ? 2 : 1);
      t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint32_t d = tmem + (i % nacc) * 128;
        const uint64_t ad = a0 + (uint64_t)((i % 6) * a_stride16);
        const uint64_t bd = b0 + (uint64_t)(((i % 6) * b_bytes_step) >> 4);
        mma_bf16_ss(d, ad, bd, idesc, 1u);
      }
      mma_commit(&bar);
    }
    __syncwarp();
    mbar_wait(&bar, 0);
    t1 = clock64();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  }
  tc_fence_before(); __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 148 * 8);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  const int iters = 4096;
  struct Cfg { int n, mode, a16, bstep; const char* what; int abase = 0; };
  Cfg cfgs[] = {
      {96, 0, 1, 3072, "N=96 same D (as the conv kernel)"}, {96, 1, 1, 3072, "N=96 two D alternating"},
      {96, 2, 1, 3072, "N=96 four D round-robin"},          {32, 0, 1, 3072, "N=32 same D"},
      {32, 2, 1, 3072, "N=32 four D"},                      {128, 0, 1, 3072, "N=128 same D"},
      {128, 2, 1, 3072, "N=128 four D"},                    {256, 0, 1, 0, "N=256 same D"},
      {256, 1, 1, 0, "N=256 two D"},                        {96, 0, 0, 0, "N=96 same D, same operands"},
      {96, 2, 0, 0, "N=96 four D, same operands"},          {64, 0, 1, 3072, "N=64 same D"},
      {192, 0, 1, 3072, "N=192 same D"},                    {192, 1, 1, 3072, "N=192 two D"},
      {96, 0, 0, 3072, "N=96 A fixed, 128B-aligned"},        {96, 0, 8, 3072, "N=96 A steps of 128 B (aligned)"},
      {96, 0, 1, 3072, "N=96 A steps of 16 B (the conv)"},   {96, 0, 8, 3072, "N=96 A steps 128 B, base +64 B", 64},
      {96, 0, 8, 3072, "N=96 A steps 128 B, base +16 B", 16}, {96, 0, 0, 3072, "N=96 A fixed, base +16 B", 16},
      {96, 0, 0, 3072, "N=96 A fixed, base +64 B", 64},      {96, 0, 0, 3072, "N=96 A fixed, base +32 B", 32},
  };
  for (auto& c : cfgs) {
    for (int grid : {1, 148}) {
      probe<<<grid, 128, 160 * 1024>>>(c.n, iters, c.mode, c.a16, c.bstep, d, c.abase);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("%s: %s\n", c.what, cudaGetErrorString(e)); return 1; }
      long long h[148]; cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost);
      long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
      printf("%-40s grid %3d: %7.1f cycles / MMA  (floor 128*N/256 = %5.1f)\n", c.what, grid, (double)mx / iters, 128.0 * c.n / 256);
    }
  }
  cudaFuncSetAttribute(probe_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
  struct P { int n1, n2, reuse; const char* what; };
  P ps[] = {{64, 32, 0, "pair N=64 + N=32, same A, plain"}, {64, 32, 1, "pair N=64 + N=32, same A, collector fill/lastuse"},
            {96, 96, 0, "pair N=96 + N=96, same A, plain"}, {96, 96, 1, "pair N=96 + N=96, same A, collector fill/lastuse"},
            {32, 64, 1, "pair N=32 + N=64, same A, collector fill/lastuse"}};
  for (auto& c : ps) {
    probe_pair<<<148, 128, 160 * 1024>>>(c.n1, c.n2, iters, c.reuse, d);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: %s\n", c.what, cudaGetErrorString(e)); return 1; }
    long long h[148]; cudaMemcpy(h, d, 148 * 8, cudaMemcpyDeviceToHost);
    long long mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("%-52s: %7.1f cycles / pair\n", c.what, (double)mx / iters);
  }
  return 0;
}
