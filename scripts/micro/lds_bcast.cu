// Shared-memory wavefronts of wide loads with broadcast: how many cycles does a warp-wide LDS.32 / .64 / .128 cost when
// every four lanes read the same address (the {offset, weight} loads of k_spmm_ring)?
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o lds_bcast lds_bcast.cu && ./lds_bcast
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ float ld32(unsigned a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a)); return v; }
__device__ __forceinline__ unsigned ld16(unsigned a) { unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
__device__ __forceinline__ float2 ld64(unsigned a) { float2 v; asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ float4 ld128(unsigned a) { float4 v; asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a)); return v; }

template <int MODE>
__global__ void k(long long* out, float* sink, int iters) {
  extern __shared__ __align__(16) unsigned char sm[];
  for (int i = threadIdx.x; i < 8192; i += blockDim.x) ((float*)sm)[i] = (float)i;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2;
  unsigned base = warp * 2048;
  const unsigned sbase = (unsigned)__cvta_generic_to_shared(sm);
  float acc = 0.f;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      const unsigned o = base + ((u + it) & 15) * 128;
      if (MODE == 0) acc += ld32(sbase + o + g * 4);                                // LDS.32, 8 unique words
      if (MODE == 1) { float2 v = ld64(sbase + o + g * 8); acc += v.x + v.y; }        // LDS.64, 8 unique
      if (MODE == 2) { float4 v = ld128(sbase + o + g * 16); acc += (v.x + v.y) + (v.z + v.w); }      // LDS.128, 8 unique (128 B)
      if (MODE == 3) { float4 v = ld128(sbase + (o & 1023u) * 4 + lane * 16); acc += (v.x + v.y) + (v.z + v.w); }  // LDS.128 all distinct
      if (MODE == 4) acc += (float)ld16(sbase + o + g * 2);                           // LDS.U16, 8 unique
      if (MODE == 5) { float4 v = ld128(sbase + o + (lane >> 3) * 16); acc += (v.x + v.y) + (v.z + v.w); }  // LDS.128, 4 unique: one per quarter warp
    }
  }
  long long t1 = clock64();
  if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

int main() {
  long long* d; float* s;
  cudaMalloc(&d, 8 * 148); cudaMalloc(&s, 4 * 148 * 512);
  const int iters = 2000, warps = 16;
  const char* names[] = {"LDS.32 x8 unique (4-lane broadcast)", "LDS.64 x8 unique", "LDS.128 x8 unique (128 B)", "LDS.128 all distinct (512 B)",
                         "LDS.U16 x8 unique", "LDS.128 x4 unique (one per quarter)"};
  for (int m = 0; m < 6; ++m) {
    for (int rep = 0; rep < 2; ++rep) {
      switch (m) {
        case 0: k<0><<<148, 32 * warps, 32768>>>(d, s, iters); break;
        case 1: k<1><<<148, 32 * warps, 32768>>>(d, s, iters); break;
        case 2: k<2><<<148, 32 * warps, 32768>>>(d, s, iters); break;
        case 3: k<3><<<148, 32 * warps, 32768>>>(d, s, iters); break;
        case 4: k<4><<<148, 32 * warps, 32768>>>(d, s, iters); break;
        case 5: k<5><<<148, 32 * warps, 32768>>>(d, s, iters); break;
      }
      cudaDeviceSynchronize();
    }
    long long h[148];
    cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
    printf("%-40s %.2f cycles per warp instruction per SM (16 warps)\n", names[m], (double)h[0] / ((double)iters * 16 * warps));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
