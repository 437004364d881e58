// dli_host_small.h -- when does a batch stay on the host?  A GPU call costs two launches, two copies and a
// synchronisation (~45 us from host buffers, profiles/r01_latency_by_batch.json) whatever its size; the reference's own
// usage pattern is one interp() per point (README.md:81-97, src/delaunay_2_test2.cpp:78-83).  Batches whose estimated
// host cost is below that floor are served by an exact walk on the host triangulation, started at the simplex the
// previous query ended in, with the reference-order epilogue of dli_refmath.h (the code the exact pass of the kernels
// runs, so host and device agree bit for bit in `reference` mode and within the 1e-13 bound of the default mode).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime_api.h>

namespace dli {

// estimated host cost: n queries x (un-hinted walk ~ n_simplices^(1/D) steps) x ~25 ns per step against ~40 us
inline bool host_batch_is_small(int64_t n, int64_t n_simplices, int dim) {
  static const double budget = [] {
    const char *e = getenv("DLI_HOST_SMALL_BATCH");   // 0 disables the host path; default 1600 step-equivalents
    return e ? atof(e) : 1600.0;
  }();
  if (n <= 0 || budget <= 0.0) return false;
  const double walk = n_simplices > 0 ? std::pow((double)n_simplices, 1.0 / dim) : 1.0;
  return (double)n * walk < budget;
}

// true for pointers the CUDA runtime knows nothing about or knows as host memory
inline bool is_plain_host_pointer(const void *p) {
  if (!p) return true;
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
  return a.type == cudaMemoryTypeUnregistered || a.type == cudaMemoryTypeHost;
}

}  // namespace dli
