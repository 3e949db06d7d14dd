// Micro-benchmark of the pipeline primitives the conv kernel is built from, one CTA, clock64 around
// each group of instructions in ONE issuing thread.  Build + run (GPU box):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o /tmp/prim_bench tools/prim_bench.cu -lcuda
//   /tmp/prim_bench
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include "../telescope_cam_detection_b200/csrc/tcd_common.cuh"

using namespace tcd;

__global__ void __launch_bounds__(128) prim_kernel(long long* out, int n_mma_cols) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[16];
  __shared__ uint32_t tmem_base_s;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 16; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, 512);
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  if (warp == 0 && lane == 0) {
    long long t0, t1;
    int k = 0;
    // 0: clock64 back to back
    t0 = clock64();
    for (int i = 0; i < 64; ++i) asm volatile("" ::: "memory");
    t1 = clock64();
    out[k++] = t1 - t0;
    // 1: 64 x (arrive + try_wait on the phase just completed)
    uint32_t ph = 0;
    t0 = clock64();
    for (int i = 0; i < 64; ++i) {
      mbar_arrive(&bars[0]);
      mbar_wait(&bars[0], ph);
      ph ^= 1u;
    }
    t1 = clock64();
    out[k++] = (t1 - t0) / 64;
    // 2: 64 x try_wait on an already completed phase (parity of the previous phase)
    t0 = clock64();
    for (int i = 0; i < 64; ++i) mbar_wait(&bars[1], 1u);
    t1 = clock64();
    out[k++] = (t1 - t0) / 64;
    // 3: 64 x arrive.expect_tx(0)  (then drain)
    t0 = clock64();
    for (int i = 0; i < 64; ++i) mbar_expect_tx(&bars[2], 0u);
    t1 = clock64();
    out[k++] = (t1 - t0) / 64;
    // 4: 64 x (tcgen05.commit + wait) with nothing outstanding: commit -> arrival latency
    ph = 0;
    t0 = clock64();
    for (int i = 0; i < 64; ++i) {
      umma_commit(&bars[3]);
      mbar_wait(&bars[3], ph);
      ph ^= 1u;
    }
    t1 = clock64();
    out[k++] = (t1 - t0) / 64;
    // 5: issue cost of 64 x tcgen05.commit to distinct-phase-free barrier (count large): no waiting
    t0 = clock64();
    for (int i = 0; i < 16; ++i) umma_commit(&bars[4 + (i & 7)]);
    t1 = clock64();
    out[k++] = (t1 - t0) / 16;
    // let those land
    for (int i = 0; i < 2000; ++i) asm volatile("" ::: "memory");
    // 6: tcgen05.fence::after_thread_sync x 64
    t0 = clock64();
    for (int i = 0; i < 64; ++i) tc_fence_after();
    t1 = clock64();
    out[k++] = (t1 - t0) / 64;
    // 7/8: 64 MMAs (M128, N=n_mma_cols, K16, 128B swizzle) issue time, then commit+wait total
    {
      const uint32_t idesc = umma_idesc_bf16(128, n_mma_cols);
      const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 16384);
      mbar_init(&bars[12], 1);
      fence_mbar_init();
      t0 = clock64();
      for (int i = 0; i < 64; ++i) {
        const uint64_t ad = umma_smem_desc(a_addr + (i & 3) * 32, 1024, 2);
        const uint64_t bd = umma_smem_desc(b_addr + (i & 3) * 32, 1024, 2);
        umma_bf16(tmem, ad, bd, idesc, (uint32_t)(i != 0));
      }
      t1 = clock64();
      out[k++] = (t1 - t0) / 64;
      umma_commit(&bars[12]);
      mbar_wait(&bars[12], 0);
      t1 = clock64();
      out[k++] = (t1 - t0) / 64;
    }
    // 9: chain of (4 MMA + commit + wait) x 16: per-iteration latency when fully serialized
    {
      const uint32_t idesc = umma_idesc_bf16(128, n_mma_cols);
      const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 16384);
      mbar_init(&bars[13], 1);
      fence_mbar_init();
      ph = 0;
      t0 = clock64();
      for (int it = 0; it < 16; ++it) {
        for (int i = 0; i < 4; ++i) {
          const uint64_t ad = umma_smem_desc(a_addr + i * 32, 1024, 2);
          const uint64_t bd = umma_smem_desc(b_addr + i * 32, 1024, 2);
          umma_bf16(tmem, ad, bd, idesc, 1u);
        }
        umma_commit(&bars[13]);
        mbar_wait(&bars[13], ph);
        ph ^= 1u;
      }
      t1 = clock64();
      out[k++] = (t1 - t0) / 16;
    }
    out[k++] = -1;
  }
  __syncthreads();
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// two-thread ping-pong: producer (warp 0) arrives on full[s], consumer (warp 1) waits, commits to
// empty[s]; S stages, no data, no MMAs: the pure handshake rate of the operand ring.
__global__ void __launch_bounds__(64) ring_kernel(long long* out, int S, int iters, int use_commit) {
  __shared__ uint64_t full_bar[8], empty_bar[8];
  __shared__ uint32_t tmem_base_s;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) {
      mbar_init(&full_bar[i], 1);
      mbar_init(&empty_bar[i], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, 32);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  long long t0 = clock64();
  if (warp == 0 && lane == 0) {
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(&empty_bar[s], ph ^ 1u);
      mbar_expect_tx(&full_bar[s], 0u);
      if (++s == S) { s = 0; ph ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    int s = 0;
    uint32_t ph = 0;
    for (int it = 0; it < iters; ++it) {
      mbar_wait(&full_bar[s], ph);
      tc_fence_after();
      if (use_commit) umma_commit(&empty_bar[s]); else mbar_arrive(&empty_bar[s]);
      if (++s == S) { s = 0; ph ^= 1u; }
    }
    out[0] = (clock64() - t0) / iters;
  }
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base_s, 32);
}

// Can TWO threads (different warps) issue tcgen05.mma concurrently at twice the rate of one?  Each issuer sends `per`
// MMAs (M128 x N x K16) to its own accumulator columns, commits and waits; out[0] = cycles from the common start to the
// later of the two completions.  issuers = 1: one thread issues all 2 * per.
__global__ void __launch_bounds__(128) dual_issue_kernel(long long* out, int N, int per, int issuers) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ uint64_t bars[2];
  __shared__ uint32_t tmem_base_s;
  __shared__ long long t_end[2];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
    t_end[0] = t_end[1] = 0;
  }
  if (warp == 1) tmem_alloc(&tmem_base_s, 512);
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_base_s;
  const long long t0 = clock64();
  const int me = warp == 0 ? 0 : (warp == 2 ? 1 : -1);
  if (me >= 0 && lane == 0 && me < issuers) {
    const uint32_t idesc = umma_idesc_bf16(128, N);
    const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 16384);
    const int n = issuers == 1 ? 2 * per : per;
    for (int i = 0; i < n; ++i) {
      const uint64_t ad = umma_smem_desc(a_addr + (i & 3) * 32, 1024, 2);
      const uint64_t bd = umma_smem_desc(b_addr + (i & 3) * 32, 1024, 2);
      umma_bf16(tmem + me * 256, ad, bd, idesc, (uint32_t)(i != 0));
    }
    umma_commit(&bars[me]);
    mbar_wait(&bars[me], 0);
    t_end[me] = clock64();
  }
  __syncthreads();
  if (threadIdx.x == 0) out[0] = (t_end[0] > t_end[1] ? t_end[0] : t_end[1]) - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

int main() {
  long long* d;
  cudaMalloc(&d, 64 * sizeof(long long));
  cudaFuncSetAttribute(prim_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  const char* names[] = {"64 empty asm (clock pair overhead, total)", "arrive + wait same thread", "wait on completed phase",
                         "arrive.expect_tx(0) issue", "tcgen05.commit + wait (nothing outstanding)", "tcgen05.commit issue",
                         "tcgen05.fence::after", "MMA issue (per MMA, 64 back to back)", "MMA issue + drain (per MMA)",
                         "4 MMA + commit + wait (per iteration, serialized)"};
  for (int n : {32, 64, 128, 256}) {
    cudaMemset(d, 0, 64 * sizeof(long long));
    prim_kernel<<<1, 128, 80 * 1024>>>(d, n);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("prim_kernel N=%d: %s\n", n, cudaGetErrorString(e)); return 1; }
    long long h[64];
    cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    printf("--- UMMA N = %d (cycles)\n", n);
    for (int i = 0; i < 10; ++i) printf("  %-52s %lld\n", names[i], h[i]);
  }
  cudaFuncSetAttribute(dual_issue_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (int n : {32, 64, 128})
    for (int issuers : {1, 2}) {
      dual_issue_kernel<<<1, 128, 80 * 1024>>>(d, n, 64, issuers);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("dual_issue_kernel: %s\n", cudaGetErrorString(e)); return 1; }
      long long h;
      cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
      printf("128 MMAs N=%d from %d issuing thread(s): %lld cycles total = %.1f per MMA\n", n, issuers, h, h / 128.0);
    }
  for (int commit = 0; commit < 2; ++commit)
    for (int S : {1, 2, 4, 8}) {
      ring_kernel<<<1, 64>>>(d, S, 512, commit);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("ring_kernel: %s\n", cudaGetErrorString(e)); return 1; }
      long long h;
      cudaMemcpy(&h, d, sizeof(h), cudaMemcpyDeviceToHost);
      printf("ring handshake S=%d %s: %lld cycles/iteration\n", S, commit ? "tcgen05.commit" : "mbarrier.arrive", h);
    }
  return 0;
}
