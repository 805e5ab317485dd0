// microbench.cu -- L2 / HBM bandwidth probes that bound the CSR gather kernel.
//   stream: coalesced float4 reads of a buffer (L2-resident or HBM-sized)
//   gather: warp-per-row reads of `row_bytes` rows at random indices (the SpMM B-row gather
//           without any arithmetic), U independent rows in flight per warp
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/build/microbench tools/microbench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__global__ void __launch_bounds__(256) stream_kernel(const float4* __restrict__ p, size_t n4, int reps, float* out) {
  float4 acc = make_float4(0, 0, 0, 0);
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (int r = 0; r < reps; ++r) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 3 * stride < n4; i += 4 * stride) {
      float4 a = __ldg(p + i), b = __ldg(p + i + stride), c = __ldg(p + i + 2 * stride), d = __ldg(p + i + 3 * stride);
      acc.x += a.x + b.x + c.x + d.x; acc.y += a.y + b.y + c.y + d.y;
      acc.z += a.z + b.z + c.z + d.z; acc.w += a.w + b.w + c.w + d.w;
    }
    for (; i < n4; i += stride) { float4 a = __ldg(p + i); acc.x += a.x; acc.y += a.y; acc.z += a.z; acc.w += a.w; }
  }
  if (acc.x + acc.y + acc.z + acc.w == 1234.5f) out[0] = acc.x;
}

template <int VEC, int U>
__global__ void __launch_bounds__(256) gather_kernel(const float4* __restrict__ tab, const int* __restrict__ idx,
                                                     size_t nidx, int row_vec, float* out) {
  const int lane = threadIdx.x & 31;
  const size_t warp = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const size_t nwarps = ((size_t)gridDim.x * blockDim.x) >> 5;
  float4 acc[VEC];
#pragma unroll
  for (int i = 0; i < VEC; ++i) acc[i] = make_float4(0, 0, 0, 0);
  // each warp owns contiguous chunks of 32 indices, like a CSR row stream
  for (size_t base = warp * 32; base < nidx; base += nwarps * 32) {
    const int my = (base + lane < nidx) ? __ldg(idx + base + lane) : 0;
#pragma unroll 1
    for (int kk = 0; kk < 32; kk += U) {
      float4 b[U][VEC];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int r = __shfl_sync(0xffffffffu, my, kk + u);
        const float4* __restrict__ row = tab + (size_t)r * row_vec + lane;
#pragma unroll
        for (int i = 0; i < VEC; ++i) b[u][i] = __ldg(row + 32 * i);
      }
#pragma unroll
      for (int u = 0; u < U; ++u)
#pragma unroll
        for (int i = 0; i < VEC; ++i) { acc[i].x += b[u][i].x; acc[i].y += b[u][i].y; acc[i].z += b[u][i].z; acc[i].w += b[u][i].w; }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < VEC; ++i) s += acc[i].x + acc[i].y + acc[i].z + acc[i].w;
  if (s == 1234.5f) out[0] = s;
}

static float time_ms(cudaEvent_t a, cudaEvent_t b) { float ms; CK(cudaEventElapsedTime(&ms, a, b)); return ms; }

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int nsm = prop.multiProcessorCount;
  printf("{\"device\": \"%s\", \"sms\": %d, \"l2_bytes\": %d}\n", prop.name, nsm, prop.l2CacheSize);
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  float* out; CK(cudaMalloc(&out, 64));
  const size_t big = (size_t)2560 << 20;  // 2.5 GiB table (products-sized B)
  float4* buf; CK(cudaMalloc(&buf, big)); CK(cudaMemset(buf, 0, big));
  // ---- stream reads
  for (size_t mb : {16, 32, 64, 96, 2560}) {
    const size_t n4 = (mb << 20) / 16;
    const int reps = mb > 1000 ? 1 : 64;
    for (int cps : {4, 8}) {
      stream_kernel<<<nsm * cps, 256>>>(buf, n4, 2, out);
      CK(cudaEventRecord(e0));
      stream_kernel<<<nsm * cps, 256>>>(buf, n4, reps, out);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      const float ms = time_ms(e0, e1);
      printf("{\"probe\": \"stream\", \"mib\": %zu, \"ctas_per_sm\": %d, \"ms\": %.4f, \"gbs\": %.1f}\n", mb, cps, ms,
             (double)n4 * 16 * reps / ms / 1e6);
    }
  }
  // ---- gathers
  const size_t nidx = (size_t)48 << 20;
  std::vector<int> h(nidx);
  int* idx; CK(cudaMalloc(&idx, nidx * 4));
  for (int row_bytes : {512, 1024}) {
    for (size_t tab_mb : {32, 64, 2560}) {
      const size_t rows = (tab_mb << 20) / row_bytes;
      uint64_t s = 88172645463325252ull;
      for (size_t i = 0; i < nidx; ++i) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; h[i] = (int)(s % rows); }
      CK(cudaMemcpy(idx, h.data(), nidx * 4, cudaMemcpyHostToDevice));
      const size_t use = tab_mb > 1000 ? nidx / 4 : nidx;
      for (int cps : {2, 4, 8}) {
        for (int variant = 0; variant < 2; ++variant) {
          auto launch = [&]() {
            if (row_bytes == 512) {
              if (variant == 0) gather_kernel<1, 8><<<nsm * cps, 256>>>(buf, idx, use, row_bytes / 16, out);
              else gather_kernel<1, 16><<<nsm * cps, 256>>>(buf, idx, use, row_bytes / 16, out);
            } else {
              if (variant == 0) gather_kernel<2, 4><<<nsm * cps, 256>>>(buf, idx, use, row_bytes / 16, out);
              else gather_kernel<2, 8><<<nsm * cps, 256>>>(buf, idx, use, row_bytes / 16, out);
            }
          };
          launch();
          CK(cudaEventRecord(e0));
          launch();
          CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
          const float ms = time_ms(e0, e1);
          printf("{\"probe\": \"gather\", \"row_bytes\": %d, \"table_mib\": %zu, \"ctas_per_sm\": %d, \"U\": %d, \"ms\": %.4f, \"gbs\": %.1f}\n",
                 row_bytes, tab_mb, cps, row_bytes == 512 ? (variant ? 16 : 8) : (variant ? 8 : 4), ms,
                 (double)use * row_bytes / ms / 1e6);
        }
      }
    }
  }
  CK(cudaGetLastError());
  return 0;
}
