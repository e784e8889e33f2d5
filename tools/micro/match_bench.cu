// Microbenchmark (nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/micro/match_bench tools/micro/match_bench.cu): cycles per warp instruction of match.any / shared atomics / shuffles on sm_100a,
// with W warps per SM issuing independent (ILP) or dependent (chain) operations.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int OP, bool CHAIN>
__global__ void bench(uint32_t* out, int iters, long long* cycles) {
  __shared__ uint32_t s[1024 * 8];
  for (int i = threadIdx.x; i < 1024 * 8; i += blockDim.x) s[i] = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t v = threadIdx.x * 2654435761u + blockIdx.x, acc = 0;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      uint32_t d = ((v >> (k + 3)) + (CHAIN ? acc : 0u)) & 1023u;
      if (OP == 0) acc += __match_any_sync(0xffffffffu, d);
      if (OP == 1) acc += atomicAdd(&s[(warp & 7) * 1024 + d], 1u);
      if (OP == 2) acc += __shfl_sync(0xffffffffu, d, (lane + k) & 31);
      if (OP == 3) acc += s[(warp & 7) * 1024 + d];
      if (OP == 4) acc += __ballot_sync(0xffffffffu, d & 1);
      if (OP == 5) { uint32_t m = __match_any_sync(0xffffffffu, d & 7u); acc += m; }
    }
    v = v * 1664525u + 1013904223u;
  }
  const long long t1 = clock64();
  if (threadIdx.x == 0 && blockIdx.x == 0) *cycles = t1 - t0;
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int OP, bool CHAIN>
void run(const char* name, int threads) {
  uint32_t* out; long long* cyc; long long h;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 8);
  const int iters = 200;
  bench<OP, CHAIN><<<148, threads>>>(out, iters, cyc);
  bench<OP, CHAIN><<<148, threads>>>(out, iters, cyc);
  cudaDeviceSynchronize();
  cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  const double per_op_warp = (double)h / (iters * 8);                       // latency-ish per op for one warp
  const double per_op_sm = (double)h / (iters * 8.0 * (threads / 32));      // cycles per warp-op per SM
  printf("%-22s %-6s threads/SM %4d: %.1f cycles per op per warp, %.2f cycles per warp-op per SM (%.2f per SMSP)\n", name,
         CHAIN ? "chain" : "ilp", threads, per_op_warp, per_op_sm, per_op_sm * 4);
  cudaFree(out); cudaFree(cyc);
}

int main() {
  for (int threads : {32, 128, 512, 1024}) {
    run<0, false>("match.any 10-bit", threads);
    run<0, true>("match.any 10-bit", threads);
    run<5, false>("match.any 3-bit", threads);
    run<1, false>("atoms.add (ret)", threads);
    run<1, true>("atoms.add (ret)", threads);
    run<2, false>("shfl.idx", threads);
    run<3, false>("lds random", threads);
    run<4, false>("ballot", threads);
  }
  return 0;
}
