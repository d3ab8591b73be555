// Box microbenchmarks that fix the second roofline of the scan (INT pipes) and the
// read-only streaming ceiling.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o
// tools/microbench tools/microbench.cu ; run on the GPU box.  Prints one JSON object.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>
#include <algorithm>

#define CHECK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

template <int MODE>
__global__ void __launch_bounds__(1024) int_kernel(uint32_t* out, uint32_t seed, int iters) {
  uint32_t a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed * (threadIdx.x + 1) + i * 0x9E3779B9u;
  uint32_t acc = 0, b = seed ^ 0x5555u, c = seed ^ 0x3333u;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) {            // POPC only (dependent through acc keeps it honest)
        a[i] = __popc(a[i] ^ acc) + a[i];
      } else if (MODE == 1) {     // LOP3 only
        a[i] = (a[i] & b) | (a[i] & c) | (b & c);
        b ^= a[i];
      } else if (MODE == 2) {     // 1 POPC : 4 LOP3
        uint32_t x = a[i] ^ b ^ c;
        uint32_t y = (a[i] & b) | (a[i] & c) | (b & c);
        uint32_t z = x ^ y ^ acc;
        uint32_t w = (x & y) | (x & acc) | (y & acc);
        a[i] = __popc(z) + w;
      } else {                    // 1 POPC : 1 XOR : 1 IADD (plain scan inner op)
        a[i] = a[i] + __popc(a[i] ^ b);
      }
    }
    acc += a[0];
  }
  uint32_t r = acc;
#pragma unroll
  for (int i = 0; i < 8; ++i) r ^= a[i];
  if (r == 0x12345678u) out[0] = r;
}

__global__ void __launch_bounds__(512) stream_read_kernel(const uint4* __restrict__ in, size_t n, uint32_t* out) {
  uint32_t acc = 0;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  for (; i + 3 * stride < n; i += 4 * stride) {
    uint4 v0, v1, v2, v3;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v0.x), "=r"(v0.y), "=r"(v0.z), "=r"(v0.w) : "l"(in + i));
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v1.x), "=r"(v1.y), "=r"(v1.z), "=r"(v1.w) : "l"(in + i + stride));
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v2.x), "=r"(v2.y), "=r"(v2.z), "=r"(v2.w) : "l"(in + i + 2 * stride));
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v3.x), "=r"(v3.y), "=r"(v3.z), "=r"(v3.w) : "l"(in + i + 3 * stride));
    acc ^= v0.x ^ v0.y ^ v0.z ^ v0.w ^ v1.x ^ v1.y ^ v1.z ^ v1.w ^ v2.x ^ v2.y ^ v2.z ^ v2.w ^ v3.x ^ v3.y ^ v3.z ^ v3.w;
  }
  for (; i < n; i += stride) { uint4 v = in[i]; acc ^= v.x ^ v.y ^ v.z ^ v.w; }
  if (acc == 0x12345678u) out[0] = acc;
}

template <int MODE>
static double run_int(uint32_t* d_out, int sms, int iters) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  int_kernel<MODE><<<sms * 2, 1024>>>(d_out, 12345u, 16);
  cudaDeviceSynchronize();
  float best = 1e30f;
  for (int r = 0; r < 5; ++r) {
    cudaEventRecord(a);
    int_kernel<MODE><<<sms * 2, 1024>>>(d_out, 12345u + r, iters);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); best = std::min(best, ms);
  }
  return best;
}

int main() {
  cudaDeviceProp p; CHECK(cudaGetDeviceProperties(&p, 0));
  int sms = p.multiProcessorCount;
  int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
  uint32_t* d_out; CHECK(cudaMalloc(&d_out, 64));
  const int iters = 20000;
  const double thread_iters = (double)sms * 2 * 1024 * iters * 8;
  double ms0 = run_int<0>(d_out, sms, iters), ms1 = run_int<1>(d_out, sms, iters);
  double ms2 = run_int<2>(d_out, sms, iters), ms3 = run_int<3>(d_out, sms, iters);
  // per-op rates (ops / s)
  double popc_only = thread_iters / (ms0 * 1e-3);
  double lop3_only = thread_iters * 2 / (ms1 * 1e-3);     // maj LOP3 + xor LOP3 per step
  double mix_popc = thread_iters / (ms2 * 1e-3);          // POPC rate while 4 LOP3 accompany each
  double plain_popc = thread_iters / (ms3 * 1e-3);
  // streaming read
  size_t bytes = (size_t)8 << 30;
  uint4* d_in; CHECK(cudaMalloc(&d_in, bytes)); CHECK(cudaMemset(d_in, 1, bytes));
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  double best_bw = 0;
  int grids[] = {sms * 2, sms * 4, sms * 8};
  int best_grid = 0;
  for (int g : grids) {
    stream_read_kernel<<<g, 512>>>(d_in, bytes / 16, d_out); cudaDeviceSynchronize();
    for (int r = 0; r < 5; ++r) {
      cudaEventRecord(a);
      stream_read_kernel<<<g, 512>>>(d_in, bytes / 16, d_out);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      double bw = bytes / (ms * 1e-3) / 1e9;
      if (bw > best_bw) { best_bw = bw; best_grid = g; }
    }
  }
  CHECK(cudaGetLastError());
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_rate_khz\": %d, \"popc_only_per_s\": %.4g, \"lop3_only_per_s\": %.4g, "
         "\"popc_with_4lop3_per_s\": %.4g, \"popc_xor_add_per_s\": %.4g, \"popc_only_per_clk_per_sm_at_max_clk\": %.3f, "
         "\"lop3_only_per_clk_per_sm_at_max_clk\": %.3f, \"stream_read_GBps\": %.1f, \"stream_read_grid\": %d}\n",
         p.name, sms, clk_khz, popc_only, lop3_only, mix_popc, plain_popc,
         popc_only / ((double)clk_khz * 1e3) / sms, lop3_only / ((double)clk_khz * 1e3) / sms, best_bw, best_grid);
  return 0;
}
