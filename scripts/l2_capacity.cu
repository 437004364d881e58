// Effective L2 capacity probe (B200): all SMs read random 64-byte records of a buffer of S MB, many passes; under
// `ncu --metrics dram__bytes_read.sum,lts__t_sector_hit_rate.pct` the DRAM bytes per request give the L2 hit rate as a
// function of S.  If data read by SMs of both dies is cached once per L2 partition, the hit rate of a uniformly random
// working set is ~ (L2 / 2) / S instead of L2 / S.
// Second part: the same reads issued only by the SMs of one "side" (smid < half or smid >= half, and even / odd smid) to
// see which split of the SMs, if any, doubles the capacity seen.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o scripts/_bin/l2_capacity scripts/l2_capacity.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ unsigned long long mix(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}
__device__ __forceinline__ unsigned smid() { unsigned r; asm volatile("mov.u32 %0, %%smid;" : "=r"(r)); return r; }

// mode 0: every SM; 1: smid < split only; 2: smid >= split only; 3: even smid; 4: odd smid
__global__ void k_read(const uint4 *__restrict__ buf, unsigned long long recs, int iters, int mode, unsigned split,
                       unsigned long long seed, unsigned *__restrict__ sink, unsigned long long *__restrict__ count) {
  const unsigned s = smid();
  const bool on = mode == 0 || (mode == 1 && s < split) || (mode == 2 && s >= split) || (mode == 3 && !(s & 1)) || (mode == 4 && (s & 1));
  if (!on) return;
  unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 0x9e3779b97f4a7c15ULL + seed;
  unsigned acc = 0;
  for (int it = 0; it < iters; it++) {
    i = mix(i + it);
    const uint4 *p = buf + (i % recs) * 4;   // 64-byte records
    uint4 a = __ldg(p), b = __ldg(p + 2);
    acc ^= a.x ^ b.y;
  }
  if (acc == 0xdeadbeefu) sink[0] = acc;
  if (threadIdx.x == 0) atomicAdd(count, (unsigned long long)blockDim.x * iters);
}

// Disjoint halves: SMs of group g = map(smid) read only half g of the buffer.  If the grouping matches the two L2
// partitions, each partition caches only its own half and the hit rate is that of a working set of S / 2.
__device__ __forceinline__ unsigned group_of(unsigned s, int map) {
  switch (map) {
    case 0: return s < 74 ? 0u : 1u;
    case 1: return s & 1u;
    case 2: return (s >> 1) & 1u;
    case 3: return (s >> 2) & 1u;
    case 4: return (s >> 3) & 1u;
    case 5: return (s >> 4) & 1u;
    case 6: return (s % 8) < 4 ? 0u : 1u;
    case 7: return (s % 16) < 8 ? 0u : 1u;
    case 8: return (s % 18) < 9 ? 0u : 1u;
    case 9: return (s / 18) & 1u;
    case 10: return (s / 20) & 1u;
    default: return (s / 37) & 1u;
  }
}
__global__ void k_read_split(const uint4 *__restrict__ buf, unsigned long long recs, int iters, int map, unsigned long long seed,
                             unsigned *__restrict__ sink) {
  const unsigned g = group_of(smid(), map);
  const unsigned long long half = recs / 2;
  unsigned long long i = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * 0x9e3779b97f4a7c15ULL + seed;
  unsigned acc = 0;
  for (int it = 0; it < iters; it++) {
    i = mix(i + it);
    const uint4 *p = buf + (g * half + i % half) * 4;
    uint4 a = __ldg(p), b = __ldg(p + 2);
    acc ^= a.x ^ b.y;
  }
  if (acc == 0xdeadbeefu) sink[0] = acc;
}

int main(int argc, char **argv) {
  if (argc > 1 && atoi(argv[1]) == 100) {   // split-map sweep on a 128 MB buffer
    uint4 *buf; unsigned *sink;
    cudaMalloc(&buf, 128ull << 20); cudaMalloc(&sink, 4);
    cudaMemset(buf, 1, 128ull << 20);
    const unsigned long long recs = (128ull << 20) / 64;
    for (int map = 0; map < 12; map++)
      for (int rep = 0; rep < 2; rep++) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        cudaEventRecord(e0);
        k_read_split<<<148 * 16, 256>>>(buf, recs, 64, map, 99 + rep, sink);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (rep == 1) printf("{\"split_map\": %d, \"MB\": 128, \"ms\": %.3f}\n", map, ms);
      }
    printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
  }
  const int mode = argc > 1 ? atoi(argv[1]) : 0;
  const unsigned split = argc > 2 ? (unsigned)atoi(argv[2]) : 74;
  uint4 *buf; unsigned *sink; unsigned long long *count;
  const size_t maxb = 512ull << 20;
  cudaMalloc(&buf, maxb); cudaMalloc(&sink, 4); cudaMalloc(&count, 8);
  cudaMemset(buf, 1, maxb);
  const int sizes[] = {32, 48, 64, 80, 96, 112, 128, 160, 192, 256, 384};
  for (int si = 0; si < 11; si++) {
    const unsigned long long recs = ((unsigned long long)sizes[si] << 20) / 64;
    cudaMemset(count, 0, 8);
    // warm pass then the measured pass (ncu sees both launches: use the second of each pair)
    for (int rep = 0; rep < 2; rep++) {
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaMemset(count, 0, 8);
      cudaEventRecord(e0);
      k_read<<<148 * 16, 256>>>(buf, recs, 64, mode, split, 1234 + rep * 77, sink, count);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      unsigned long long c; cudaMemcpy(&c, count, 8, cudaMemcpyDeviceToHost);
      if (rep == 1) printf("{\"mode\": %d, \"split\": %u, \"MB\": %d, \"requests\": %llu, \"ms\": %.3f, \"Greq_per_s\": %.2f}\n", mode, split, sizes[si], c, ms, c / ms / 1e6);
    }
  }
  printf("status: %s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
