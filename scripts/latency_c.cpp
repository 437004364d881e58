// Per-call latency of dli_interp seen by a C / C++ caller (no Python in the loop): cfg2-like mesh (1M random vertices),
// device-resident and pinned-host batches of 1 .. 1M queries, wall clock around call + stream synchronisation.
//   g++ -O2 -std=c++14 -I include -I /usr/local/cuda/include scripts/latency_c.cpp -o scripts/_bin/latency_c \
//       -L delaunay_linterp_b200 -ldli -L /usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/delaunay_linterp_b200
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <random>
#include <vector>
#include <cuda_runtime_api.h>
#include "dli.h"
#include "dli_host.h"

int main() {
  const int64_t nv = 1000000;
  std::mt19937_64 rng(1001);
  std::uniform_real_distribution<double> U(0.0, 1.0);
  std::vector<double> xy(2 * nv), f(nv);
  for (int64_t i = 0; i < nv; i++) { xy[2 * i] = U(rng); xy[2 * i + 1] = U(rng); f[i] = std::pow(xy[2 * i], 0.3) * std::pow(xy[2 * i + 1], 0.6); }
  dli_di2 *t = dli_di2_create(nullptr, nullptr);
  if (dli_di2_insert_many(t, nv, xy.data(), f.data(), 1, 0) != DLI_OK) { printf("insert_many failed\n"); return 1; }
  const int64_t nmax = 1 << 20;
  std::vector<double> q(2 * nmax);
  for (auto &v : q) v = 0.02 + 0.96 * U(rng);
  double *dq, *dy, *hq, *hy;
  cudaMalloc((void **)&dq, sizeof(double) * 2 * nmax); cudaMalloc((void **)&dy, sizeof(double) * nmax);
  cudaMallocHost((void **)&hq, sizeof(double) * 2 * nmax); cudaMallocHost((void **)&hy, sizeof(double) * nmax);
  cudaMemcpy(dq, q.data(), sizeof(double) * 2 * nmax, cudaMemcpyHostToDevice);
  std::copy(q.begin(), q.end(), hq);
  if (dli_di2_interp(t, nmax, dq, dy, nullptr, nullptr, nullptr) != DLI_OK) { printf("interp failed\n"); return 1; }   // flatten
  cudaDeviceSynchronize();
  const dli_mesh *m = dli_di2_device_mesh(t);
  printf("{\"what\": \"dli_interp per-call latency from C++, cfg2-like mesh (%lld simplices)\"}\n", (long long)dli_mesh_num_simplices(m));
  const int64_t sizes[] = {1024, 4096, 32768, 262144, 1048576};
  for (int64_t n : sizes) {
    for (int mode = 0; mode < 2; mode++) {
      const double *xs = mode ? hq : dq;
      double *ys = mode ? hy : dy;
      const int reps = n <= 32768 ? 400 : 50;
      std::vector<double> ts;
      for (int r = 0; r < reps + 10; r++) {
        auto t0 = std::chrono::steady_clock::now();
        dli_interp(m, n, xs, ys, nullptr, nullptr, nullptr);
        cudaStreamSynchronize(nullptr);
        auto t1 = std::chrono::steady_clock::now();
        if (r >= 10) ts.push_back(std::chrono::duration<double, std::micro>(t1 - t0).count());
      }
      std::sort(ts.begin(), ts.end());
      printf("{\"queries\": %lld, \"buffers\": \"%s\", \"median_us\": %.1f, \"p10_us\": %.1f}\n", (long long)n, mode ? "pinned host" : "device",
             ts[ts.size() / 2], ts[ts.size() / 10]);
    }
  }
  dli_di2_destroy(t);
  return 0;
}
