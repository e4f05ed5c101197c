// Microbenchmarks behind the Kalman-filter kernel design (run on the B200):
//   throughput of DFMA and of mma.sync.m8n8k4.f64 (DMMA) at 4..16 warps per SM, and the latency of
//   the dependent chains the blocked factorisation is made of (DFMA, DMMA accumulate, DMMA
//   A-operand dependency, 64-bit shuffle, rsqrt + refinement, shared-memory load).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_probe fp64_probe.cu && ./fp64_probe
#include <cstdio>
#include <cuda_runtime.h>

#define DMMA(c0, c1, a, b)                                                                         \
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n" \
               : "+d"(c0), "+d"(c1)                                                                \
               : "d"(a), "d"(b))

__global__ void dfma_kernel(double* out, int iters) {
  double a[8], b = 1.000001, c = 0.5;
  for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 1e-3 + i;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = fma(a[i], b, c);
  }
  double s = 0;
  for (int i = 0; i < 8; i++) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void dmma_kernel(double* out, int iters) {
  double c[8][2];
  double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
  for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = 0.0;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) DMMA(c[i][0], c[i][1], a, b);
  }
  double s = 0;
  for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// mixed: per iteration 4 DMMA + 8 DFMA (independent)
__global__ void mix_kernel(double* out, int iters) {
  double c[4][2], f[8];
  double a = threadIdx.x * 1e-3, b = 1.0 + threadIdx.x * 1e-6;
  for (int i = 0; i < 4; i++) c[i][0] = c[i][1] = 0.0;
  for (int i = 0; i < 8; i++) f[i] = i + a;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 4; i++) DMMA(c[i][0], c[i][1], a, b);
#pragma unroll
    for (int i = 0; i < 8; i++) f[i] = fma(f[i], b, a);
  }
  double s = 0;
  for (int i = 0; i < 4; i++) s += c[i][0] + c[i][1];
  for (int i = 0; i < 8; i++) s += f[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__device__ __forceinline__ double rsqrt_pos(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x, y * y, 1.0);
  const double p = fma(0.375, e, 0.5);
  return fma(y * e, p, y);
}

// one warp, dependent chains; which selects the chain
__global__ void latency_kernel(double* out, long long* cyc, int iters, int which) {
  __shared__ double sh[64];
  const int lane = threadIdx.x;
  sh[lane] = 1.0 + lane * 1e-9;
  sh[lane + 32] = 0.5;
  __syncwarp();
  double x = 1.0 + lane * 1e-6, y = 0.999999, c0 = 0.0, c1 = 0.0;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
    if (which == 0) {          // DFMA chain
#pragma unroll
      for (int i = 0; i < 16; i++) x = fma(x, y, 1e-9);
    } else if (which == 1) {   // DMMA chain through the accumulator
#pragma unroll
      for (int i = 0; i < 16; i++) DMMA(c0, c1, x, y);
    } else if (which == 2) {   // DMMA chain through the A operand (C of one is A of the next)
#pragma unroll
      for (int i = 0; i < 16; i++) {
        double d0 = 0.0, d1 = 0.0;
        DMMA(d0, d1, x, y);
        x = d0 * 1e-3 + 1.0;
      }
    } else if (which == 3) {   // 64-bit shuffle chain
#pragma unroll
      for (int i = 0; i < 16; i++) x = __shfl_sync(0xffffffffu, x, (lane + 1) & 31);
    } else if (which == 4) {   // rsqrt + refinement chain
#pragma unroll
      for (int i = 0; i < 16; i++) x = rsqrt_pos(x) + 1.0;
    } else if (which == 5) {   // shared-memory load chain (index depends on the loaded value)
#pragma unroll
      for (int i = 0; i < 16; i++) x = sh[(int)(x) & 31];
    } else if (which == 6) {   // store -> syncwarp -> broadcast load chain (publishing a row)
#pragma unroll
      for (int i = 0; i < 16; i++) {
        sh[lane] = x;
        __syncwarp();
        x = sh[(lane + 1) & 31] + 1e-9;
        __syncwarp();
      }
    } else if (which == 7) {   // div chain
#pragma unroll
      for (int i = 0; i < 16; i++) x = 1.0 / x + 0.5;
    }
  }
  long long t1 = clock64();
  out[lane] = x + c0 + c1;
  if (lane == 0) cyc[0] = (t1 - t0);
}

int main() {
  double* out;
  long long* cyc;
  cudaMalloc(&out, sizeof(double) * 148 * 8 * 1024);
  cudaMalloc(&cyc, sizeof(long long));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int iters = 20000;
  for (int warps = 4; warps <= 16; warps += 4) {
    dim3 grid(148), block(warps * 32);
    for (int which = 0; which < 3; which++) {
      for (int rep = 0; rep < 2; rep++) {
        cudaEventRecord(e0);
        if (which == 0) dfma_kernel<<<grid, block>>>(out, iters);
        else if (which == 1) dmma_kernel<<<grid, block>>>(out, iters);
        else mix_kernel<<<grid, block>>>(out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
      }
      float ms;
      cudaEventElapsedTime(&ms, e0, e1);
      double fma = (which == 0) ? (double)grid.x * block.x * iters * 8
                   : (which == 1) ? (double)grid.x * (block.x / 32) * iters * 8 * 256
                                  : (double)grid.x * (block.x / 32) * iters * (4 * 256 + 8 * 32);
      printf("%s warps/SM=%2d: %8.3f ms  %6.2f TFLOP/s fp64\n", which == 0 ? "DFMA" : (which == 1 ? "DMMA" : "MIX "), warps, ms,
             2 * fma / ms / 1e9);
    }
  }
  const char* names[] = {"DFMA chain", "DMMA chain (accumulator)", "DMMA chain (A operand, +FMA)", "SHFL.64 chain",
                         "rsqrt+refine(+add) chain", "LDS chain (dependent index)", "STS+syncwarp+LDS chain", "DDIV(+add) chain"};
  for (int which = 0; which < 8; which++) {
    for (int rep = 0; rep < 2; rep++) latency_kernel<<<1, 32>>>(out, cyc, 1000, which);
    cudaDeviceSynchronize();
    long long h;
    cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-32s %7.1f cycles per op\n", names[which], (double)h / (1000.0 * 16));
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
