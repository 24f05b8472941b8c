// Issue-rate microbenchmark for the integer instructions the front end leans on (POPC, LOP3, IMAD,
// VIMNMX3.S16x2, IDP.4A, SHFL, LDS).  One CTA of 1024 threads per SM, 8 independent chains per thread,
// result = warp instructions per clock per SM.  DESIGN.md quotes these as the issue rooflines for
// K2 / K4 / K6 / K7.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_probe pipe_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int kIters = 4096, kChains = 8;

template <int OP>
__global__ void __launch_bounds__(1024) k_probe(uint32_t* out, uint32_t seed, long long* clocks) {
  __shared__ uint32_t s_buf[1024 + 64];
  uint32_t v[kChains];
#pragma unroll
  for (int c = 0; c < kChains; ++c) v[c] = seed * (threadIdx.x + 1) + c * 0x9E3779B9u;
  s_buf[threadIdx.x] = v[0];
  __syncthreads();
  const uint32_t k1 = seed | 1u, k2 = seed ^ 0x5bd1e995u;
  const long long t0 = clock64();
  for (int i = 0; i < kIters; ++i) {
#pragma unroll
    for (int c = 0; c < kChains; ++c) {
      if (OP == 0) v[c] = __popc(v[c]) + k1;                                   // POPC (+ IADD to keep the chain alive)
      else if (OP == 1) asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(v[c]) : "r"(k1), "r"(k2));
      else if (OP == 2) v[c] = v[c] * k1 + k2;                                 // IMAD
      else if (OP == 3) v[c] = __vimin3_s16x2(v[c], k1, k2);                   // VIMNMX3.S16x2
      else if (OP == 4) v[c] = __dp4a(v[c], k1, k2);                           // IDP.4A
      else if (OP == 5) v[c] = __shfl_xor_sync(0xffffffffu, v[c], 1);          // SHFL
      else if (OP == 6) v[c] = s_buf[(v[c] & 1023u)];                          // LDS (dependent address)
      else if (OP == 7) v[c] = v[c] + k1;                                      // IADD
      else if (OP == 8) { v[c] = __popc(v[c] ^ k1); }                          // LOP + POPC pair, no add
    }
  }
  const long long t1 = clock64();
  uint32_t acc = 0;
#pragma unroll
  for (int c = 0; c < kChains; ++c) acc ^= v[c];
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0) clocks[blockIdx.x] = t1 - t0;
}

template <int OP>
static void run(const char* name, int extra_per_op, uint32_t* d_out, long long* d_clk, int sms) {
  k_probe<OP><<<sms, 1024>>>(d_out, 12345u, d_clk);
  k_probe<OP><<<sms, 1024>>>(d_out, 12345u, d_clk);
  cudaDeviceSynchronize();
  long long h[1024];
  cudaMemcpy(h, d_clk, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
  double avg = 0;
  for (int i = 0; i < sms; ++i) avg += (double)h[i];
  avg /= sms;
  const double warp_instr = 32.0 * kIters * kChains;   // 32 warps per CTA
  printf("{\"op\": \"%s\", \"warp_instr_per_clk_per_sm\": %.3f, \"note\": \"%d instruction(s) per step counted as one\"}\n", name,
         warp_instr / avg, 1 + extra_per_op);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int sms = p.multiProcessorCount;
  uint32_t* d_out; long long* d_clk;
  cudaMalloc(&d_out, sizeof(uint32_t) * sms * 1024);
  cudaMalloc(&d_clk, sizeof(long long) * 1024);
  printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, sms, p.clockRate);
  run<0>("popc+iadd", 1, d_out, d_clk, sms);
  run<8>("lop+popc", 1, d_out, d_clk, sms);
  run<1>("lop3", 0, d_out, d_clk, sms);
  run<7>("iadd", 0, d_out, d_clk, sms);
  run<2>("imad", 0, d_out, d_clk, sms);
  run<3>("vimnmx3.s16x2", 0, d_out, d_clk, sms);
  run<4>("idp.4a", 0, d_out, d_clk, sms);
  run<5>("shfl", 0, d_out, d_clk, sms);
  run<6>("lds.dependent", 1, d_out, d_clk, sms);
  return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
