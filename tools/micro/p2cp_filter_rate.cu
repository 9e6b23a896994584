// What a tensor-core FILTER for P2CP / the tract variables would cost per contour pair, measured on one SM.
//
// The proposed formulation gets an approximate 50 x 50 d^2 tile per contour pair
// from tcgen05.mma, then needs (i) every value read back from tensor memory and (ii) a row minimum and a
// column minimum over it before any exact refinement can start.  This bench measures exactly those two
// floors, with nothing else in the loop:
//   MODE 0  tcgen05.ld only      (32 lanes x 64 columns per warp and iteration, values xor-folded)
//   MODE 1  tcgen05.ld + FMNMX3  (row minimum of the 64 values of the thread's lane: 32 min3)
//   MODE 2  FMNMX3 only          (the same 32 min3 on register values: ALU-pipe rate)
//   MODE 3  the present kernel's packed body on registers (2 sub2, mul2, fma2 + 2 min3 per 2 x 2 values),
//           same number of VALUES as the other modes: the fp32-pipe cost it would replace
// One "pair-equivalent" = 128 lanes x 50 columns = both orientations (u rows over v, v rows over u) of one
// 50 x 50 contour pair padded to the 128-lane tile; the table prints clk per pair-equivalent so that it
// reads against the budget: 1218 clk per frame = 122 clk per contour pair today (p2cp50_kernel, 2.31e8
// frames/s), 69 clk per pair for 50 % of the HBM roofline.
// nvcc -gencode arch=compute_100a,code=sm_100a -I../../artspeech_b200/csrc -I../../include
#include <cstdio>
#include "umma.cuh"
using namespace asb;
using namespace asb::umma;

template <int MODE>
__global__ void filter_kernel(long long* out, float* sink, int iters) {
  __shared__ uint32_t slot;
  __shared__ long long t_begin[16], t_end[16];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) tmem_alloc(&slot, 512);
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tm = slot + (uint32_t(warp & 3) << 21);   // lane quadrant of this warp: bits 31:16 = lane
  {  // defined contents (finite floats) in the columns that are read
    uint32_t z[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) z[i] = __float_as_uint(1.0f + float((lane * 37 + i * 11) & 63));
    if (warp < 4) {
      for (int c = 0; c < 512; c += 32) tmem_st32(tm + c, z);
      tmem_st_wait();
    }
  }
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  float acc = __int_as_float(0x7f800000);
  uint32_t fold = 0;
  float a0 = 1.f + lane, a1 = 2.f + lane;
  f32x2 ux = pack2(0.1f * lane, 0.2f * lane), uy = pack2(0.3f * lane, 0.4f);
  f32x2 vx = pack2(0.5f, 0.5f), vy = pack2(0.25f, 0.25f);
  float rlo = acc, rhi = acc, ca = acc;
  const long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
    const uint32_t col = (it & 3) * 64;
    if (MODE == 0 || MODE == 1) {
      uint32_t r0[32], r1[32];
      tmem_ld32(tm + col, r0);
      tmem_ld32(tm + col + 32, r1);
      tmem_ld_wait();
      if (MODE == 0) {
#pragma unroll
        for (int i = 0; i < 32; ++i) fold ^= r0[i] ^ r1[i];
      } else {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          acc = min3(acc, __uint_as_float(r0[i]), __uint_as_float(r0[i + 1]));
          acc = min3(acc, __uint_as_float(r1[i]), __uint_as_float(r1[i + 1]));
        }
      }
    } else if (MODE == 2) {
      // four independent chains, 32 min3 in all, operands live in registers
      float m0 = acc, m1 = acc, m2 = acc, m3 = acc;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        m0 = min3(m0, a0, a1);
        m1 = min3(m1, a1, a0 + float(i));
        m2 = min3(m2, a0, a1);
        m3 = min3(m3, a1, a0);
        a0 += 1.f;
      }
      acc = fminf(fminf(m0, m1), fminf(m2, m3));
    } else {
      // 64 values: 16 trips of the production body (2 rows x 2 columns each)
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const f32x2 dxa = sub2(ux, vx), dya = sub2(uy, vy);
        const f32x2 dxb = sub2(ux, vy), dyb = sub2(uy, vx);
        const f32x2 qa = fma2(dya, dya, mul2(dxa, dxa));
        const f32x2 qb = fma2(dyb, dyb, mul2(dxb, dxb));
        float p0, p1, q0, q1;
        unpack2(qa, p0, p1);
        unpack2(qb, q0, q1);
        rlo = min3(rlo, p0, q0);
        rhi = min3(rhi, p1, q1);
        ca = min3(ca, p0, p1);
        acc = min3(acc, q0, q1);
        vx = pack2(p0, q1);   // keeps the trips dependent on data (no hoisting), off the critical chain
      }
    }
  }
  const long long t1 = clock64();
  if (lane == 0) {
    t_begin[warp] = t0;
    t_end[warp] = t1;
  }
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc + __uint_as_float(fold) + rlo + rhi + ca;
  fence_before_sync();
  __syncthreads();
  if (threadIdx.x == 0) {
    long long b = t_begin[0], e = t_end[0];
    for (int w = 1; w < int(blockDim.x >> 5); ++w) {
      b = t_begin[w] < b ? t_begin[w] : b;
      e = t_end[w] > e ? t_end[w] : e;
    }
    out[blockIdx.x] = e - b;
  }
  if (warp == 0) tmem_dealloc(slot, 512);
}

template <int MODE>
void run(long long* d, float* sink, const char* name, int warps, int grid) {
  const int iters = 4096;
  filter_kernel<MODE><<<grid, warps * 32>>>(d, sink, iters);
  long long h[148];
  cudaError_t e = cudaMemcpy(h, d, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
  long long worst = 0;
  for (int i = 0; i < grid; ++i) worst = h[i] > worst ? h[i] : worst;
  // one iteration = warps x 32 lanes x 64 columns; pair-equivalent = 128 lanes x 50 columns
  const double clk_iter = double(worst) / iters;
  const double clk_pair = clk_iter * (128.0 * 50.0) / (warps * 32.0 * 64.0);
  const double bytes_clk = warps * 32.0 * 64.0 * 4.0 / clk_iter;
  printf("%-26s warps %2d grid %3d: %7.1f clk/iter  %6.1f clk per pair-equivalent  (%6.1f B/clk/SM of values)  %s\n",
         name, warps, grid, clk_iter, clk_pair, bytes_clk, cudaGetErrorString(e));
}

int main() {
  long long* d;
  float* sink;
  cudaMalloc(&d, sizeof(long long) * 148);
  cudaMalloc(&sink, sizeof(float) * 148 * 512);
  for (int grid : {1, 148}) {
    for (int warps : {4, 8, 16}) {
      run<0>(d, sink, "tcgen05.ld only", warps, grid);
      run<1>(d, sink, "tcgen05.ld + min3", warps, grid);
      run<2>(d, sink, "min3 only (registers)", warps, grid);
      run<3>(d, sink, "direct form (production)", warps, grid);
    }
  }
  return 0;
}
