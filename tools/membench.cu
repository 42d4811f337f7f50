// tools/membench.cu -- measurement aid (not product): what the B200 memory system does with random 32-byte
// sector reads, as a function of footprint (TLB reach), load width and address order.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/membench tools/membench.cu && tools/membench
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x)                                                                       \
  do {                                                                              \
    cudaError_t e = (x);                                                            \
    if (e != cudaSuccess) {                                                         \
      printf("{\"error\": \"%s at %s\"}\n", cudaGetErrorString(e), #x);              \
      exit(1);                                                                      \
    }                                                                               \
  } while (0)

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}
__device__ __forceinline__ uint64_t mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdull; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ull; x ^= x >> 33;
  return x;
}

__device__ __forceinline__ uint32_t ld256(const uint32_t *p) {
  uint32_t a, b, c, d, e, f, g, h;
  asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(p));
  return a ^ b ^ c ^ d ^ e ^ f ^ g ^ h;
}
__device__ __forceinline__ uint32_t ld64(const uint32_t *p) {
  uint32_t a, b;
  asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p));
  return a ^ b;
}

// MODE 0: independent random sectors; 1: dependent chain; 2: sorted sparse sweep (thread t, step s reads a random
// sector inside its own window; windows ascend with the global access number)
template <int WIDE, int MODE, int MLP>
__global__ void __launch_bounds__(256) k_rand(const uint32_t *buf, uint64_t nsect, uint32_t steps, uint64_t nthreads,
                                              unsigned long long *sink) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nthreads) return;
  uint32_t acc = 0;
  uint64_t x[MLP];
#pragma unroll
  for (int m = 0; m < MLP; ++m) x[m] = mix64(t * MLP + m + 0x9E3779B97F4A7C15ull);
  for (uint32_t s = 0; s < steps; ++s) {
    uint32_t v[MLP];
#pragma unroll
    for (int m = 0; m < MLP; ++m) {
      uint64_t sect;
      if (MODE == 2) {
        // access number n = (s * nthreads + t) * MLP + m of N = steps * nthreads * MLP: window n of N equal windows
        const uint64_t n = ((uint64_t)s * nthreads + t) * MLP + m, N = (uint64_t)steps * nthreads * MLP;
        const uint64_t w = nsect / N ? nsect / N : 1;
        sect = (n * w + mix64(n) % w) % nsect;
      } else {
        x[m] = mix64(x[m] + s);
        sect = x[m] % nsect;
      }
      const uint32_t *p = buf + sect * 8;
      v[m] = WIDE ? ld256(p) : ld64(p);
    }
#pragma unroll
    for (int m = 0; m < MLP; ++m) {
      acc += v[m];
      if (MODE == 1) x[m] += v[m];
    }
  }
  if (acc == 0x7fffffffu) atomicAdd(sink, 1ull);
}

__global__ void k_fill(uint32_t *buf, uint64_t nwords) {
  for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nwords; i += (uint64_t)gridDim.x * blockDim.x)
    buf[i] = mix32((uint32_t)i * 2654435761u + (uint32_t)(i >> 32));
}

// unloaded latency: `nthreads` threads in all (one per SM block), each a dependent chain of random 8-byte loads
__global__ void k_latency(const uint32_t *buf, uint64_t nsect, uint32_t steps, unsigned long long *sink) {
  uint64_t x = mix64(blockIdx.x * 977 + threadIdx.x + 12345);
  uint32_t acc = 0;
  for (uint32_t s = 0; s < steps; ++s) {
    x = mix64(x + s);
    const uint32_t v = ld64(buf + (x % nsect) * 8);
    acc += v;
    x += v;
  }
  if (acc == 0x7fffffffu) atomicAdd(sink, 1ull);
}
double latency_ns(const uint32_t *buf, uint64_t nsect, unsigned blocks, unsigned threads, unsigned long long *sink) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  const uint32_t steps = 4000;
  k_latency<<<blocks, threads>>>(buf, nsect, 100, sink);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  k_latency<<<blocks, threads>>>(buf, nsect, steps, sink);
  CK(cudaEventRecord(b));
  CK(cudaEventSynchronize(b));
  float ms;
  CK(cudaEventElapsedTime(&ms, a, b));
  return ms * 1e6 / steps;
}

template <int WIDE, int MODE, int MLP>
double run(const uint32_t *buf, uint64_t nsect, uint32_t steps, uint64_t nthreads, unsigned long long *sink) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  const unsigned grid = (unsigned)((nthreads + 255) / 256);
  k_rand<WIDE, MODE, MLP><<<grid, 256>>>(buf, nsect, 2, nthreads, sink);
  CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int rep = 0; rep < 3; ++rep) {
    CK(cudaEventRecord(a));
    k_rand<WIDE, MODE, MLP><<<grid, 256>>>(buf, nsect, steps, nthreads, sink);
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    if (ms < best) best = ms;
  }
  cudaEventDestroy(a);
  cudaEventDestroy(b);
  return (double)nthreads * steps * MLP / (best * 1e-3) / 1e9;  // G loads / s
}

int main(int argc, char **argv) {
  const uint64_t max_bytes = argc > 1 ? strtoull(argv[1], nullptr, 10) << 30 : 64ull << 30;
  uint32_t *buf;
  unsigned long long *sink;
  CK(cudaMalloc(&buf, max_bytes));
  CK(cudaMalloc(&sink, 8));
  CK(cudaMemset(sink, 0, 8));
  k_fill<<<148 * 16, 256>>>(buf, max_bytes / 4);
  CK(cudaDeviceSynchronize());
  const uint64_t nthreads = 148ull * 2048 * 8;
  for (uint64_t bytes = 64ull << 20; bytes <= max_bytes; bytes <<= 1)
    printf("{\"mib\": %llu, \"latency_ns_1thread\": %.0f, \"latency_ns_148x32\": %.0f, \"latency_ns_148x256\": %.0f}\n",
           (unsigned long long)(bytes >> 20), latency_ns(buf, bytes / 32, 1, 1, sink), latency_ns(buf, bytes / 32, 148, 32, sink),
           latency_ns(buf, bytes / 32, 148, 256, sink));
  fflush(stdout);
  if (argc > 2) return 0;
  for (uint64_t bytes = 64ull << 20; bytes <= max_bytes; bytes <<= 1) {
    const uint64_t ns = bytes / 32;
    printf("{\"mib\": %llu, \"indep32_mlp1\": %.1f, \"indep32_mlp4\": %.1f, \"dep32_mlp1\": %.1f, \"dep32_mlp2\": %.1f, "
           "\"indep8_mlp1\": %.1f, \"dep8_mlp1\": %.1f, \"sorted32_mlp1\": %.1f, \"sorted8_mlp1\": %.1f}\n",
           (unsigned long long)(bytes >> 20), run<1, 0, 1>(buf, ns, 16, nthreads, sink), run<1, 0, 4>(buf, ns, 8, nthreads, sink),
           run<1, 1, 1>(buf, ns, 16, nthreads, sink), run<1, 1, 2>(buf, ns, 8, nthreads, sink),
           run<0, 0, 1>(buf, ns, 16, nthreads, sink), run<0, 1, 1>(buf, ns, 16, nthreads, sink),
           run<1, 2, 1>(buf, ns, 16, nthreads, sink), run<0, 2, 1>(buf, ns, 16, nthreads, sink));
    fflush(stdout);
  }
  return 0;
}
