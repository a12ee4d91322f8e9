// Brute-force check of dfu_div_by (constant-divisor correctly rounded division, dfu_common.cuh) against __ddiv_rn:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -I include -o /tmp/divcheck tools/divcheck.cu && /tmp/divcheck
// Operands: (i) the back-projection's own domain, a = (col - cu) * depth with float32 cu, depth and focal lengths from
// 500 to 3000; (ii) random doubles over 40 binades for a and random float32-valued and full-double divisors.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "dfu3d.h"
#include "../dfu3d_b200/csrc/dfu_common.cuh"

__device__ __forceinline__ uint64_t mix(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

__global__ void k_check(uint64_t seed, int mode, unsigned long long* bad, unsigned long long* n) {
  const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  unsigned long long nb = 0;
  for (int it = 0; it < 256; it++) {
    const uint64_t h = mix(seed + t * 256 + it), g = mix(h);
    double a, d;
    if (mode == 0) {
      const float fu = 500.0f + (float)(h & 0xffffff) * (2500.0f / 16777216.0f);
      const float cu = 300.0f + (float)((h >> 24) & 0xffffff) * (1000.0f / 16777216.0f);
      const float depth = 0.001f + (float)(g & 0xffffff) * (120.0f / 16777216.0f);
      const int col = (int)((g >> 24) % 2048);
      a = __dmul_rn(__dsub_rn((double)col, (double)cu), (double)depth);
      d = (double)fu;
    } else {
      a = __longlong_as_double((long long)((h & 0x800fffffffffffffull) | ((uint64_t)(1003 + (h >> 52) % 40) << 52)));
      d = (mode == 1) ? (double)__uint_as_float((unsigned)((g & 0x007fffffu) | ((120u + (unsigned)((g >> 23) % 16)) << 23)))
                      : __longlong_as_double((long long)((g & 0x000fffffffffffffull) | ((uint64_t)(1015 + (g >> 52) % 16) << 52)));
    }
    const double y = __drcp_rn(d);
    if (__double_as_longlong(dfu_div_by(a, d, y)) != __double_as_longlong(__ddiv_rn(a, d))) nb++;
  }
  if (nb) atomicAdd(bad, nb);
  if (threadIdx.x == 0) atomicAdd(n, 256ull * blockDim.x);
}

int main() {
  unsigned long long *bad, *n, hb[3] = {0, 0, 0}, hn[3] = {0, 0, 0};
  cudaMalloc(&bad, 8); cudaMalloc(&n, 8);
  for (int mode = 0; mode < 3; mode++) {
    cudaMemset(bad, 0, 8); cudaMemset(n, 0, 8);
    k_check<<<148 * 128, 256>>>(0x1234567ull + mode, mode, bad, n);   // 148*128*256*256 = 1.24e9 pairs per mode
    cudaMemcpy(&hb[mode], bad, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&hn[mode], n, 8, cudaMemcpyDeviceToHost);
    printf("mode %d: %llu pairs, %llu differ from __ddiv_rn\n", mode, hn[mode], hb[mode]);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("%s\n", cudaGetErrorString(e));
  return (hb[0] | hb[1] | hb[2]) ? 1 : 0;
}
