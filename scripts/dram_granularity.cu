// DRAM fetch granularity probe (B200): every thread reads SZ bytes (one vector load) at a pseudo-random SZ-aligned
// offset of a buffer far larger than L2.  Run under `ncu --metrics dram__bytes_read.sum,gpu__time_duration.sum` and
// divide by the request count: the bytes DRAM moves per random SZ-byte request.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_bin/dram_granularity scripts/dram_granularity.cu
//   scripts/_bin/dram_granularity [GiB]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long mix(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

template <int SZ>
__global__ void k_random_read(const char *__restrict__ buf, unsigned long long slots, unsigned long long n, unsigned *__restrict__ sink) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const char *p = buf + (mix(i * 0x9e3779b97f4a7c15ULL + 12345) % slots) * SZ;
  unsigned acc = 0;
  if (SZ == 16) {
    uint4 v = __ldg(reinterpret_cast<const uint4 *>(p));
    acc = v.x ^ v.y ^ v.z ^ v.w;
  } else if (SZ == 32) {
    unsigned r[8];
    asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p));
    for (int k = 0; k < 8; k++) acc ^= r[k];
  } else {
    for (int c = 0; c < SZ / 32; c++) {
      unsigned r[8];
      asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "l"(p + 32 * c));
      for (int k = 0; k < 8; k++) acc ^= r[k];
    }
  }
  if (acc == 0xdeadbeefu) sink[0] = acc;
}

// 16-byte random reads with an explicit L2 prefetch-size qualifier (0 = none, 64 / 128 / 256 = ld.global.L2::<n>B)
template <int PF>
__global__ void k_random_read16_pf(const char *__restrict__ buf, unsigned long long slots, unsigned long long n, unsigned *__restrict__ sink) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const char *p = buf + (mix(i * 0x9e3779b97f4a7c15ULL + 999) % slots) * 16;
  unsigned a, b, c, d;
  if (PF == 64) asm volatile("ld.global.L2::64B.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
  else if (PF == 128) asm volatile("ld.global.L2::128B.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
  else if (PF == 256) asm volatile("ld.global.L2::256B.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
  else asm volatile("ld.global.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "l"(p));
  if ((a ^ b ^ c ^ d) == 0xdeadbeefu) sink[0] = a;
}
template <int PF>
static void run_pf(const char *buf, size_t bytes, unsigned long long n, unsigned *sink) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_random_read16_pf<PF><<<(unsigned)((n + 255) / 256), 256>>>(buf, bytes / 16, n, sink);
  cudaEventRecord(e0);
  k_random_read16_pf<PF><<<(unsigned)((n + 255) / 256), 256>>>(buf, bytes / 16, n, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("{\"kernel\": \"random_read16_pf\", \"l2_prefetch_qualifier\": %d, \"requests\": %llu, \"ms\": %.3f, \"Grequests_per_s\": %.2f}\n", PF, n, ms, n / ms / 1e6);
}

// 8-byte random writes (the result scatter of a binned batch): DRAM bytes read + written per store
__global__ void k_random_write8(double *__restrict__ buf, unsigned long long slots, unsigned long long n) {
  const unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  buf[mix(i * 0x9e3779b97f4a7c15ULL + 777) % slots] = (double)i;
}

template <int SZ>
static void run(const char *buf, size_t bytes, unsigned long long n, unsigned *sink) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_random_read<SZ><<<(unsigned)((n + 255) / 256), 256>>>(buf, bytes / SZ, n, sink);
  cudaEventRecord(e0);
  k_random_read<SZ><<<(unsigned)((n + 255) / 256), 256>>>(buf, bytes / SZ, n, sink);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("{\"kernel\": \"random_read\", \"request_bytes\": %d, \"requests\": %llu, \"ms\": %.3f, \"Grequests_per_s\": %.2f, \"requested_GBps\": %.1f}\n",
         SZ, n, ms, n / ms / 1e6, (double)n * SZ / ms / 1e6);
}

int main(int argc, char **argv) {
  const size_t gib = argc > 1 ? (size_t)atoi(argv[1]) : 4;
  const size_t bytes = gib << 30;
  if (argc > 2 && atoi(argv[2]) > 0) {   // cudaLimitMaxL2FetchGranularity, set before anything touches the device
    cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(argv[2]));
    size_t got = 0;
    cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
    printf("{\"set_limit_max_l2_fetch_granularity\": %d, \"rc\": \"%s\", \"readback\": %zu}\n", atoi(argv[2]), cudaGetErrorString(e), got);
  } else {
    size_t got = 0;
    cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
    printf("{\"default_max_l2_fetch_granularity\": %zu}\n", got);
  }
  char *buf;
  unsigned *sink;
  if (cudaMalloc(&buf, bytes) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMalloc(&sink, 4);
  cudaMemset(buf, 1, bytes);
  const unsigned long long n = 64ull << 20;
  run_pf<0>(buf, bytes, n, sink);
  run_pf<64>(buf, bytes, n, sink);
  run_pf<128>(buf, bytes, n, sink);
  run_pf<256>(buf, bytes, n, sink);
  run<16>(buf, bytes, n, sink);
  run<32>(buf, bytes, n, sink);
  run<64>(buf, bytes, n, sink);
  run<128>(buf, bytes, n, sink);
  {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_random_write8<<<(unsigned)((n + 255) / 256), 256>>>(reinterpret_cast<double *>(buf), bytes / 8, n);
    cudaEventRecord(e0);
    k_random_write8<<<(unsigned)((n + 255) / 256), 256>>>(reinterpret_cast<double *>(buf), bytes / 8, n);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    printf("{\"kernel\": \"random_write8\", \"requests\": %llu, \"ms\": %.3f, \"Grequests_per_s\": %.2f}\n", n, ms, n / ms / 1e6);
  }
  cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
