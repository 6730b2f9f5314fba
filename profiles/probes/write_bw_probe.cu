// Write-only HBM bandwidth of a B200 by store flavour: what bounds a kernel that only emits results
// (khop_emit_bits_kernel writes 650 MB per config-2 batch and reads 33 MB).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o write_bw_probe write_bw_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__global__ void st32(uint32_t* p, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = (uint32_t)i;
}
__global__ void st128(uint4* p, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    p[i] = make_uint4((uint32_t)i, 1u, 2u, 3u);
}
__global__ void st128_cs(uint4* p, size_t n) {   // streaming (evict-first) stores
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    __stcs(p + i, make_uint4((uint32_t)i, 1u, 2u, 3u));
}
// contiguous 16 KB per CTA step: bulk copy shared -> global (TMA 1D)
__global__ void bulk_s2g(unsigned char* p, size_t bytes) {
  extern __shared__ __align__(128) unsigned char sm[];
  const uint32_t CH = 16384;
  for (uint32_t i = threadIdx.x; i < CH / 16; i += blockDim.x) reinterpret_cast<uint4*>(sm)[i] = make_uint4(i, 1u, 2u, 3u);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(sm);
    for (size_t off = (size_t)blockIdx.x * CH; off + CH <= bytes; off += (size_t)gridDim.x * CH) {
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + off), "r"(s), "r"(CH) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
}
__global__ void copy128(const uint4* __restrict__ a, uint4* b, size_t n) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) b[i] = a[i];
}

int main() {
  const size_t bytes = 1ull << 30;
  unsigned char *a, *b;
  cudaMalloc(&a, bytes);
  cudaMalloc(&b, bytes);
  cudaMemset(a, 1, bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  auto run = [&](const char* name, auto fn, double moved) {
    float best = 1e9f;
    for (int it = 0; it < 8; ++it) {
      cudaEventRecord(e0);
      fn();
      cudaEventRecord(e1);
      cudaEventSynchronize(e1);
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      if (ms < best) best = ms;
    }
    printf("{\"probe\": \"%s\", \"ms\": %.4f, \"GBs\": %.1f, \"err\": \"%s\"}\n", name, best, moved / best / 1e6,
           cudaGetErrorString(cudaGetLastError()));
  };
  const int grid = 148 * 16;
  run("st.32 fill 1 GiB", [&] { st32<<<grid, 256>>>((uint32_t*)b, bytes / 4); }, (double)bytes);
  run("st.128 fill 1 GiB", [&] { st128<<<grid, 256>>>((uint4*)b, bytes / 16); }, (double)bytes);
  run("st.128.cs fill 1 GiB", [&] { st128_cs<<<grid, 256>>>((uint4*)b, bytes / 16); }, (double)bytes);
  cudaFuncSetAttribute(bulk_s2g, cudaFuncAttributeMaxDynamicSharedMemorySize, 16384);
  run("cp.async.bulk s2g fill 1 GiB", [&] { bulk_s2g<<<148 * 8, 128, 16384>>>(b, bytes); }, (double)bytes);
  run("cudaMemsetAsync 1 GiB", [&] { cudaMemsetAsync(b, 0, bytes); }, (double)bytes);
  run("ld/st.128 copy 1 GiB (read + write bytes)", [&] { copy128<<<grid, 256>>>((const uint4*)a, (uint4*)b, bytes / 16); }, 2.0 * bytes);
  return 0;
}
