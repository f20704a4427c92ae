// tools/micro/red_bench.cu — throughput of global reductions/atomics to (pseudo-)random addresses.
// Development microbenchmark behind DESIGN.md's statement about the L2 atomic rate.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t mix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}

template <int MODE>
__global__ void k(unsigned long long* tab, uint32_t mask, size_t n, int per_thread, unsigned long long* sink) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long acc = 0;
  for (int k = 0; k < per_thread; ++k) {
    const uint64_t h = mix(i * 131 + k);
    const uint32_t a = (uint32_t)h & mask;
    if (MODE == 0) asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(tab + a), "l"((unsigned long long)(h >> 40)) : "memory");
    else if (MODE == 1) asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"((uint32_t*)tab + a), "r"((uint32_t)(h >> 40)) : "memory");
    else if (MODE == 2) acc += atomicAdd(tab + a, (unsigned long long)(h >> 40));
    else if (MODE == 3) {  // three reds to one 64-byte row (the QuotaRow pattern)
      unsigned long long* row = tab + (size_t)(a & (mask >> 3)) * 8;
      asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(row), "l"(1ull) : "memory");
      asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(row + 2), "l"(2ull) : "memory");
      asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(row + 4), "l"(3ull) : "memory");
    } else if (MODE == 4) {  // three reds to three separate arrays (SoA)
      const size_t third = ((size_t)mask + 1) / 4;
      const uint32_t b = a & (uint32_t)(third - 1);
      asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(tab + b), "l"(1ull) : "memory");
      asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(tab + third + b), "l"(2ull) : "memory");
      asm volatile("red.relaxed.gpu.global.add.u64 [%0], %1;" ::"l"(tab + 2 * third + b), "l"(3ull) : "memory");
    }
  }
  if (acc == 0x1234567) *sink = acc;
}

template <int MODE>
void run(const char* name, unsigned long long* tab, size_t words, int ops_per_iter) {
  const int per_thread = 16;
  const size_t threads = (size_t)1 << 24;
  unsigned long long* sink;
  cudaMalloc(&sink, 8);
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEventRecord(a);
    k<MODE><<<(unsigned)(threads / 256), 256>>>(tab, (uint32_t)(words - 1), threads, per_thread, sink);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
  }
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  const double ops = (double)threads * per_thread * ops_per_iter;
  printf("{\"op\":\"%s\",\"table_MB\":%.0f,\"Gops_s\":%.1f,\"ms\":%.3f}\n", name, words * 8.0 / 1e6, ops / ms / 1e6, ms);
  cudaFree(sink);
}

int main() {
  for (size_t words : {(size_t)1 << 17, (size_t)1 << 20, (size_t)1 << 23}) {  // 1 MB, 8 MB, 64 MB
    unsigned long long* tab;
    cudaMalloc(&tab, words * 8);
    cudaMemset(tab, 0, words * 8);
    run<0>("red.u64 random", tab, words, 1);
    run<1>("red.u32 random", tab, words, 1);
    run<2>("atom.u64 random (returning)", tab, words, 1);
    run<3>("3 x red.u64 into one 64B row", tab, words, 3);
    run<4>("3 x red.u64 into three arrays", tab, words, 3);
    cudaFree(tab);
  }
  return 0;
}
