// Throughput of the epilogue's arithmetic on one SM: MUFU.TANH / MUFU.EX2 / MUFU.RCP / FFMA / F2FP pack, with
// 1..16 warps resident (1 CTA), 256 independent ops per thread in 8 chains.  Prints cycles per warp-instruction
// per scheduler.  Build + run (GPU box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/sfu_bench tools/sfu_bench.cu && /tmp/sfu_bench
#include <cstdio>
#include <cuda_bf16.h>

template <int OP>
__global__ void k(float* out, long long* cyc, int iters) {
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = 0.001f * (threadIdx.x + i);
  __syncthreads();
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (OP == 0) asm volatile("tanh.approx.f32 %0, %0;" : "+f"(v[i]));
      if (OP == 1) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
      if (OP == 2) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(v[i]));
      if (OP == 3) asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(v[i]));
      if (OP == 4) {  // the SiLU the conv epilogue uses: h + h * tanh(h)
        float h = 0.5f * v[i], t;
        asm volatile("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(h));
        v[i] = fmaf(h, t, h);
      }
      if (OP == 5) {  // SiLU through ex2 + rcp: x / (1 + 2^(-x*log2e))
        float e, r;
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * v[i]));
        asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.0f + e));
        v[i] = v[i] * r;
      }
      if (OP == 6) {  // bf16x2 pack
        unsigned u;
        asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(u) : "f"(v[i]), "f"(v[(i + 1) & 7]));
        v[i] = __uint_as_float(u);
      }
    }
  }
  const long long t1 = clock64();
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += v[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
  float* out;
  long long* cyc;
  cudaMalloc(&out, 1 << 20);
  cudaMalloc(&cyc, 1024);
  const char* names[] = {"tanh.approx", "ex2.approx", "rcp.approx", "fma", "silu = h + h*tanh(h)", "silu = x*rcp(1+ex2)", "cvt.bf16x2"};
  const int iters = 256;
  for (int op = 0; op < 7; ++op)
    for (int warps : {1, 4, 8, 16}) {
      void (*fn)(float*, long long*, int) = op == 0 ? k<0> : op == 1 ? k<1> : op == 2 ? k<2> : op == 3 ? k<3> : op == 4 ? k<4> : op == 5 ? k<5> : k<6>;
      fn<<<1, warps * 32>>>(out, cyc, iters);
      fn<<<1, warps * 32>>>(out, cyc, iters);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("error\n"); return 1; }
      long long h;
      cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
      // per scheduler: warps/4 warps each issuing iters*8 ops
      const double per_sched_ops = (warps < 4 ? 1 : warps / 4.0) * iters * 8;
      printf("%-24s warps=%2d  %8lld cycles  %.2f cycles per warp-op per scheduler\n", names[op], warps, h, h / per_sched_ops);
    }
  return 0;
}
