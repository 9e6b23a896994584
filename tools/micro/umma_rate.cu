// Raw tcgen05.mma issue/execute rate on one SM (M = 128, K = 8, kind::tf32, A from TMEM) and the cost of
// what a pipelined GEMM loop puts between the MMAs.  Each variant is its own template instance: a
// runtime branch inside the elected region makes the compiler wrap every MMA in R2UR/ELECT loops.
// nvcc -gencode arch=compute_100a,code=sm_100a -I../../artspeech_b200/csrc -I../../include
#include <cstdio>
#include "umma.cuh"
using namespace asb;
using namespace asb::umma;

template <int MODE, int N>
__global__ void rate_kernel(long long* out, int iters, int stage_in) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar, bar2, bar3;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<float*>(smem)[i] = 0.f;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&bar2, 1);
    mbar_init(&bar3, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&slot, 512);
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = slot;
  if (warp == 0) {
    long long t0 = 0;
    uint32_t c = stage_in;
    for (int rep = 0; rep < 2; ++rep) {
      t0 = clock64();
      for (int i = 0; i < iters; i += 4) {
        if (MODE == 2 || MODE == 4 || MODE == 5) {
          mbar_wait(&bar3, 1);   // already complete: returns at once
          mbar_wait(&bar3, 1);
        }
        if (MODE == 3 || MODE == 4 || MODE == 5) fence_after_sync();
        const uint32_t st = (MODE == 5) ? (c % 3) : 0;
        const uint64_t bdesc = smem_desc_k_sw128(smem_u32(smem + st * 32768));
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < 4; ++k) mma_tf32_ts(tm, tm + 256 + k * 8, bdesc + uint64_t(2 * k), idesc_tf32_m128(N), (i + k) != 0);
          if (MODE == 1 || MODE == 4 || MODE == 5) commit(&bar2);
        }
        __syncwarp();
        ++c;
      }
      if (elect_one()) commit(&bar);
      __syncwarp();
      const long long t1 = clock64();
      mbar_wait(&bar, rep & 1);
      const long long t2 = clock64();
      if (lane == 0 && rep == 1) {
        out[0] = t1 - t0;
        out[1] = t2 - t0;
      }
    }
  }
  fence_before_sync();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tm, 512);
}

template <int MODE, int N>
void run(long long* d, const char* name) {
  const int iters = 512;
  cudaFuncSetAttribute(rate_kernel<MODE, N>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  rate_kernel<MODE, N><<<1, 128, 100 * 1024>>>(d, iters, 0);
  long long h[2];
  cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("%-34s N=%3d: %.1f clk per 4 MMAs (issue), %.1f (complete)  (%s)\n", name, N, 4.0 * double(h[0]) / iters,
         4.0 * double(h[1]) / iters, cudaGetErrorString(e));
}

template <int N>
void all(long long* d) {
  run<0, N>(d, "4 MMAs per elect");
  run<1, N>(d, "+ commit");
  run<2, N>(d, "+ 2 try_wait (complete)");
  run<3, N>(d, "+ fence::after_thread_sync");
  run<4, N>(d, "+ commit, waits, fence");
  run<5, N>(d, "+ all, runtime stage descriptor");
}

int main() {
  long long* d;
  cudaMalloc(&d, 16);
  all<256>(d);
  all<112>(d);
  all<16>(d);
  return 0;
}
