// Pipe-throughput microbenchmark for the softmax inner loop of tc_attn.cu (sm_100a): cycles per warp instruction and
// scheduler (SMSP) for MUFU.EX2, FFMA, FFMA2, FADD2, F2FP, FMNMX3 with 1, 2 and 4 warps per scheduler, 16 independent
// chains per thread.   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipes pipes.cu && ./pipes
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

#define CHAINS 16
#define ITERS 256

template <int OP>
__global__ void bench(float* out, long long* cyc, float seed) {
  float v[CHAINS];
  for (int i = 0; i < CHAINS; ++i) v[i] = seed + threadIdx.x * 0.001f + i;
  uint64_t w[CHAINS / 2];
  for (int i = 0; i < CHAINS / 2; ++i) asm("mov.b64 %0, {%1, %2};" : "=l"(w[i]) : "f"(v[2 * i]), "f"(v[2 * i + 1]));
  uint64_t c2;
  asm("mov.b64 %0, {%1, %1};" : "=l"(c2) : "f"(seed));
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; ++i) {
      if (OP == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
      if (OP == 1) asm volatile("fma.rn.f32 %0, %0, %1, %1;" : "+f"(v[i]) : "f"(seed));
      if (OP == 4) asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(v[i]) : "f"(v[(i + 1) % CHAINS]), "f"(seed));
      if (OP == 5) {
        uint32_t p;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p) : "f"(v[i]), "f"(v[(i + 1) % CHAINS]));
        v[i] = __uint_as_float(p);
      }
    }
#pragma unroll
    for (int i = 0; i < CHAINS / 2; ++i) {
      if (OP == 2) asm volatile("fma.rn.f32x2 %0, %0, %1, %1;" : "+l"(w[i]) : "l"(c2));
      if (OP == 3) asm volatile("add.rn.f32x2 %0, %0, %1;" : "+l"(w[i]) : "l"(c2));
    }
  }
  long long t1 = clock64();
  float acc = 0;
  for (int i = 0; i < CHAINS; ++i) acc += v[i];
  for (int i = 0; i < CHAINS / 2; ++i) acc += (float)(w[i] & 0xff);
  out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

// the attention mix: per element 1 FFMA + 1 MUFU + 1 FADD + 0.5 F2FP + 0.5 FMNMX3 (what the round-1 loop issues)
__global__ void mix(float* out, long long* cyc, float seed) {
  float v[CHAINS];
  for (int i = 0; i < CHAINS; ++i) v[i] = seed + threadIdx.x * 0.001f + i;
  float sum = 0, mx = -1e30f;
  uint32_t pk = 0;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < ITERS; ++it) {
#pragma unroll
    for (int i = 0; i < CHAINS; i += 2) {
      float a = fmaf(v[i], seed, -seed), b = fmaf(v[i + 1], seed, -seed);
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a));
      asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(b));
      sum += a;
      sum += b;
      uint32_t p;
      asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(p) : "f"(a), "f"(b));
      pk ^= p;
      asm volatile("max.f32 %0, %0, %1, %2;" : "+f"(mx) : "f"(v[i]), "f"(v[i + 1]));
      v[i] = a * 0.5f;
      v[i + 1] = b * 0.5f;
    }
  }
  long long t1 = clock64();
  out[blockIdx.x * blockDim.x + threadIdx.x] = sum + mx + __uint_as_float(pk);
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMalloc(&cyc, 8);
  const char* names[6] = {"MUFU.EX2", "FFMA", "FFMA2", "FADD2", "FMNMX3", "F2FP.BF16x2"};
  for (int warps_per_smsp = 1; warps_per_smsp <= 4; warps_per_smsp *= 2) {
    const int threads = 128 * warps_per_smsp;
    for (int op = 0; op < 6; ++op) {
      long long h = 0;
      for (int rep = 0; rep < 2; ++rep) {
        switch (op) {
          case 0: bench<0><<<1, threads>>>(out, cyc, 0.5f); break;
          case 1: bench<1><<<1, threads>>>(out, cyc, 0.5f); break;
          case 2: bench<2><<<1, threads>>>(out, cyc, 0.5f); break;
          case 3: bench<3><<<1, threads>>>(out, cyc, 0.5f); break;
          case 4: bench<4><<<1, threads>>>(out, cyc, 0.5f); break;
          case 5: bench<5><<<1, threads>>>(out, cyc, 0.5f); break;
        }
        cudaDeviceSynchronize();
        cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      }
      const double n = (op == 2 || op == 3) ? CHAINS / 2 : CHAINS;
      printf("%d warp(s)/SMSP %-12s %6.2f cycles per warp instruction per scheduler (%.2f per instr per warp)\n",
             warps_per_smsp, names[op], (double)h / (ITERS * n * warps_per_smsp), (double)h / (ITERS * n));
    }
    long long h = 0;
    for (int rep = 0; rep < 2; ++rep) {
      mix<<<1, threads>>>(out, cyc, 0.5f);
      cudaDeviceSynchronize();
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    }
    printf("%d warp(s)/SMSP softmax mix  %6.2f cycles per ELEMENT per scheduler (128 elements -> %.0f cycles per warp-block)\n",
           warps_per_smsp, (double)h / (ITERS * CHAINS * warps_per_smsp), (double)h / (ITERS * CHAINS) * 128);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
