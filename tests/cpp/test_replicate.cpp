// dli_mesh_replicate called from C++ with real NCCL communicators: one process, one thread per GPU,
// communicators from ncclCommInitAll.  Rank 0 flattens a 2D and a 3D mesh; every other rank receives a replica through
// ncclBroadcast and must answer a batch bit-identically.  Needs >= 2 GPUs (exit code 77 = skipped).
// Build: g++ -std=c++11 test_replicate.cpp -I include -I /usr/local/cuda/include -L <libdir> -ldli -lnccl -lcudart -lpthread
#include <cmath>
#include <cstdio>
#include <cstring>
#include <thread>
#include <vector>

#include <cuda_runtime_api.h>
#include <nccl.h>

#include "dli.h"
#include "dli_host.h"

static double fn3(const double *x) { return pow(x[0] + 0.1, 0.3) * pow(x[1] + 0.1, 0.6) * pow(x[2] + 0.1, 0.2); }

int main() {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 2) { printf("skipped: needs 2 GPUs (found %d)\n", ndev); return 77; }
  const int nr = ndev < 4 ? ndev : 4;
  std::vector<ncclComm_t> comms((size_t)nr);
  std::vector<int> devs((size_t)nr);
  for (int i = 0; i < nr; i++) devs[(size_t)i] = i;
  if (ncclCommInitAll(comms.data(), nr, devs.data()) != ncclSuccess) { printf("ncclCommInitAll failed\n"); return 1; }

  // meshes on rank 0: 2D through the host triangulation class, 3D through dli_dd
  const int n2 = 20000, n3 = 5000, nq = 200000;
  unsigned long long s = 88172645463325252ull;
  auto rnd = [&]() { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (double)(s >> 11) / 9007199254740992.0; };
  std::vector<double> p2((size_t)n2 * 2), f2((size_t)n2), p3((size_t)n3 * 3), f3((size_t)n3);
  for (int i = 0; i < n2; i++) { p2[2 * i] = rnd(); p2[2 * i + 1] = rnd(); f2[i] = pow(p2[2 * i] + 0.1, 0.3) * pow(p2[2 * i + 1] + 0.1, 0.6); }
  for (int i = 0; i < n3; i++) { for (int k = 0; k < 3; k++) p3[3 * i + k] = rnd(); f3[i] = fn3(&p3[3 * i]); }
  cudaSetDevice(0);
  dli_di2 *h2 = dli_di2_create(nullptr, nullptr);
  dli_dd *h3 = dli_dd_create(3, nullptr, nullptr);
  if (dli_di2_insert_many(h2, n2, p2.data(), f2.data(), 1, 0) != DLI_OK || dli_dd_insert_many(h3, n3, p3.data(), f3.data(), 1, 0) != DLI_OK) {
    printf("triangulation failed\n");
    return 1;
  }
  const dli_mesh *src[2] = {dli_di2_device_mesh(h2), dli_dd_device_mesh(h3)};
  if (!src[0] || !src[1]) { printf("device mesh failed: %s\n", dli_last_error()); return 1; }
  const int dims[2] = {2, 3};
  std::vector<double> q[2], want[2];
  for (int t = 0; t < 2; t++) {
    q[t].resize((size_t)nq * dims[t]);
    for (double &v : q[t]) v = 0.05 + 0.9 * rnd();
    want[t].resize((size_t)nq);
    if (dli_interp(src[t], nq, q[t].data(), want[t].data(), nullptr, nullptr, nullptr) != DLI_OK) { printf("interp failed: %s\n", dli_last_error()); return 1; }
  }
  std::vector<int> fails((size_t)nr, 0);
  std::vector<std::thread> th;
  for (int rank = 0; rank < nr; rank++)
    th.emplace_back([&, rank]() {
      cudaSetDevice(rank);
      for (int t = 0; t < 2; t++) {
        dli_mesh *m = rank == 0 ? const_cast<dli_mesh *>(src[t]) : nullptr;
        if (dli_mesh_replicate(&m, comms[(size_t)rank], 0) != DLI_OK) { printf("rank %d: %s\n", rank, dli_last_error()); fails[(size_t)rank]++; return; }
        if (rank == 0) { if (m != src[t]) fails[(size_t)rank]++; continue; }
        std::vector<double> ys((size_t)nq);
        std::vector<unsigned char> st((size_t)nq);
        if (dli_interp(m, nq, q[t].data(), ys.data(), nullptr, st.data(), nullptr) != DLI_OK) { printf("rank %d: %s\n", rank, dli_last_error()); fails[(size_t)rank]++; }
        else if (memcmp(ys.data(), want[t].data(), sizeof(double) * (size_t)nq) != 0) { printf("rank %d dim %d: replica answers differ\n", rank, dims[t]); fails[(size_t)rank]++; }
        if (dli_mesh_dim(m) != dims[t] || dli_mesh_num_simplices(m) != dli_mesh_num_simplices(src[t])) fails[(size_t)rank]++;
        dli_mesh_destroy(m);
      }
    });
  for (auto &t : th) t.join();
  int bad = 0;
  for (int f : fails) bad += f;
  for (auto c : comms) ncclCommDestroy(c);
  dli_di2_destroy(h2);
  dli_dd_destroy(h3);
  if (bad) { printf("FAILED (%d)\n", bad); return 1; }
  printf("ok: %d ranks, 2D + 3D replicas bit-identical\n", nr);
  return 0;
}
