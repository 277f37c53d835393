/* Roofline denominator for this path: random aligned loads of one granule (128 B = an L2 line, as
 * SURVEY.md 8(d) asks; 64 B = one block of rs_layout.h) over a buffer far larger than L2.  granule/16
 * lanes read one granule with one 128-bit load each, every thread keeps UNROLL independent granules in
 * flight, no dependence between loads. */
#include "rs_internal.h"

#define RG_UNROLL 8

__device__ __forceinline__ uint64_t rg_mix(uint64_t x) { /* splitmix64 finaliser */
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
  x ^= x >> 27; x *= 0x94D049BB133111EBull;
  x ^= x >> 31;
  return x;
}

template <int LANES>
__global__ void __launch_bounds__(256) k_random_gather(const uint4* __restrict__ buf, uint64_t nlines, uint64_t per_group,
                                                       uint64_t seed, unsigned long long* __restrict__ sink) {
  const uint32_t g = threadIdx.x & (LANES - 1);
  const uint64_t group = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES;
  uint32_t acc = 0;
  for (uint64_t it = 0; it < per_group; it += RG_UNROLL) {
    uint4 v[RG_UNROLL];
#pragma unroll
    for (int u = 0; u < RG_UNROLL; ++u) {
      const uint64_t line = rg_mix(seed + group * per_group + it + u) % nlines;
      v[u] = __ldg(buf + line * LANES + g);
    }
#pragma unroll
    for (int u = 0; u < RG_UNROLL; ++u) acc += v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) atomicAdd(sink, 1ull);
}

int rs_measure_random_access(int device, uint64_t buffer_bytes, uint64_t lines_to_read, double* gbytes_per_s) {
  return rs_measure_random_access_granule(device, buffer_bytes, lines_to_read, 128, gbytes_per_s);
}

int rs_measure_random_access_granule(int device, uint64_t buffer_bytes, uint64_t lines_to_read, int granule_bytes,
                                     double* gbytes_per_s) {
  if (!gbytes_per_s || buffer_bytes < (1u << 20) || (granule_bytes != 32 && granule_bytes != 64 && granule_bytes != 128))
    return fail(RS_STATUS_ARG, "bad argument");
  const int lanes = granule_bytes / 16;
  CU(cudaSetDevice(device));
  void* buf = nullptr;
  unsigned long long* sink = nullptr;
  CU(cudaMalloc(&buf, buffer_bytes));
  CU(cudaMalloc((void**)&sink, 8));
  CU(cudaMemset(buf, 1, buffer_bytes));
  CU(cudaMemset(sink, 0, 8));
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
  const uint64_t nlines = buffer_bytes / granule_bytes;
  const int grid = sms * 8;
  const uint64_t groups = (uint64_t)grid * 256 / lanes;
  uint64_t per_group = (lines_to_read + groups - 1) / groups;
  per_group = (per_group + RG_UNROLL - 1) / RG_UNROLL * RG_UNROLL;
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) { /* rep 0 warms up */
    CU(cudaEventRecord(e0));
    const uint64_t seed = 0x9E3779B97F4A7C15ull * (rep + 1);
    if (lanes == 8) k_random_gather<8><<<grid, 256>>>((const uint4*)buf, nlines, per_group, seed, sink);
    else if (lanes == 4) k_random_gather<4><<<grid, 256>>>((const uint4*)buf, nlines, per_group, seed, sink);
    else k_random_gather<2><<<grid, 256>>>((const uint4*)buf, nlines, per_group, seed, sink);
    CU(cudaEventRecord(e1));
    CU(cudaEventSynchronize(e1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, e0, e1));
    if (rep && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(buf);
  cudaFree(sink);
  *gbytes_per_s = (double)(groups * per_group) * (double)granule_bytes / (best * 1e-3) / 1e9;
  return RS_STATUS_OK;
}
