// Rate of the P2CP inner loop (packed FADD2/FMUL2/FFMA2 + FMNMX3) on one SM as a function of the number
// of resident warps, and of a scalar-FFMA formulation, to see what bounds csrc/p2cp.cu: the fp32 pipe,
// the register file ports (64-bit operands) or the two warps per scheduler it runs with.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I../../artspeech_b200/csrc -I../../include fp32x2_rate.cu
#include <cstdio>
#include "asb_common.cuh"
using namespace asb;

constexpr int kD = 50;

// MODE 0: the production body (packed, two columns per trip); 1: scalar sub/mul/fma, one column per trip;
// MODE 2: packed without the minima (fma pipe only); 3: packed, four columns per trip
template <int MODE, int RP>
__global__ void __launch_bounds__(512, 1) body_kernel(const float* __restrict__ g, float* out, long long* clk, int reps) {
  __shared__ float sv[2 * kD + 8];
  for (int i = threadIdx.x; i < 2 * kD + 8; i += blockDim.x) sv[i] = g[i];
  __syncthreads();
  const float kInf = __int_as_float(0x7f800000);
  f32x2 ux[RP], uy[RP];
  float rlo[RP], rhi[RP];
#pragma unroll
  for (int p = 0; p < RP; ++p) {
    ux[p] = pack2(g[128 + threadIdx.x + 2 * p], g[129 + threadIdx.x + 2 * p]);
    uy[p] = pack2(g[256 + threadIdx.x + 2 * p], g[257 + threadIdx.x + 2 * p]);
    rlo[p] = kInf;
    rhi[p] = kInf;
  }
  float colsum = 0.f;
  const long long t0 = clock64();
  for (int r = 0; r < reps; ++r) {
    const float* pv = sv;
    if (MODE == 1) {
#pragma unroll 1
      for (int j = 0; j < kD; ++j) {
        const float vx = pv[j], vy = pv[kD + j];
        float c = kInf;
#pragma unroll
        for (int p = 0; p < RP; ++p) {
          float x0, x1, y0, y1;
          unpack2(ux[p], x0, x1);
          unpack2(uy[p], y0, y1);
          const float dx0 = x0 - vx, dy0 = y0 - vy, dx1 = x1 - vx, dy1 = y1 - vy;
          const float q0 = fmaf(dy0, dy0, dx0 * dx0), q1 = fmaf(dy1, dy1, dx1 * dx1);
          rlo[p] = fminf(rlo[p], q0);
          rhi[p] = fminf(rhi[p], q1);
          c = min3(c, q0, q1);
        }
        colsum += c;
      }
    } else if (MODE == 3) {
#pragma unroll 1
      for (int j = 0; j < kD - 2; j += 4) {
        const float4 vx = *reinterpret_cast<const float4*>(pv + j);
        const float2 vy0 = *reinterpret_cast<const float2*>(pv + kD + j);
        const float2 vy1 = *reinterpret_cast<const float2*>(pv + kD + j + 2);
        const f32x2 vxa = pack2(vx.x, vx.x), vxb = pack2(vx.y, vx.y), vxc = pack2(vx.z, vx.z), vxd = pack2(vx.w, vx.w);
        const f32x2 vya = pack2(vy0.x, vy0.x), vyb = pack2(vy0.y, vy0.y), vyc = pack2(vy1.x, vy1.x), vyd = pack2(vy1.y, vy1.y);
        float ca = kInf, cb = kInf, cc = kInf, cd = kInf;
#pragma unroll
        for (int p = 0; p < RP; ++p) {
          const f32x2 dxa = sub2(ux[p], vxa), dya = sub2(uy[p], vya);
          const f32x2 dxb = sub2(ux[p], vxb), dyb = sub2(uy[p], vyb);
          const f32x2 dxc = sub2(ux[p], vxc), dyc = sub2(uy[p], vyc);
          const f32x2 dxd = sub2(ux[p], vxd), dyd = sub2(uy[p], vyd);
          const f32x2 qa = fma2(dya, dya, mul2(dxa, dxa));
          const f32x2 qb = fma2(dyb, dyb, mul2(dxb, dxb));
          const f32x2 qc = fma2(dyc, dyc, mul2(dxc, dxc));
          const f32x2 qd = fma2(dyd, dyd, mul2(dxd, dxd));
          float a0, a1, b0, b1, c0, c1, d0, d1;
          unpack2(qa, a0, a1);
          unpack2(qb, b0, b1);
          unpack2(qc, c0, c1);
          unpack2(qd, d0, d1);
          rlo[p] = min3(min3(rlo[p], a0, b0), c0, d0);
          rhi[p] = min3(min3(rhi[p], a1, b1), c1, d1);
          ca = min3(ca, a0, a1);
          cb = min3(cb, b0, b1);
          cc = min3(cc, c0, c1);
          cd = min3(cd, d0, d1);
        }
        colsum += ca + cb + cc + cd;
      }
    } else {
      float2 vx = *reinterpret_cast<const float2*>(pv);
      float2 vy = *reinterpret_cast<const float2*>(pv + kD);
#pragma unroll 1
      for (int j = 0; j < kD; j += 2) {
        const f32x2 vxa = pack2(vx.x, vx.x), vxb = pack2(vx.y, vx.y);
        const f32x2 vya = pack2(vy.x, vy.x), vyb = pack2(vy.y, vy.y);
        const int jn = (j + 2 < kD) ? j + 2 : j;
        vx = *reinterpret_cast<const float2*>(pv + jn);
        vy = *reinterpret_cast<const float2*>(pv + kD + jn);
        float ca = kInf, cb = kInf;
#pragma unroll
        for (int p = 0; p < RP; ++p) {
          const f32x2 dxa = sub2(ux[p], vxa), dya = sub2(uy[p], vya);
          const f32x2 dxb = sub2(ux[p], vxb), dyb = sub2(uy[p], vyb);
          const f32x2 qa = fma2(dya, dya, mul2(dxa, dxa));
          const f32x2 qb = fma2(dyb, dyb, mul2(dxb, dxb));
          float a0, a1, b0, b1;
          unpack2(qa, a0, a1);
          unpack2(qb, b0, b1);
          if (MODE == 2) {
            rlo[p] += a0 * b1;
            rhi[p] += a1 * b0;
          } else {
            rlo[p] = min3(rlo[p], a0, b0);
            rhi[p] = min3(rhi[p], a1, b1);
            ca = min3(ca, a0, a1);
            cb = min3(cb, b0, b1);
          }
        }
        colsum += ca + cb;
      }
    }
  }
  const long long t1 = clock64();
  float s = colsum;
#pragma unroll
  for (int p = 0; p < RP; ++p) s += rlo[p] + rhi[p];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) clk[blockIdx.x] = t1 - t0;
}

template <int MODE, int RP>
void run(const char* name, const float* g, float* out, long long* clk) {
  const int reps = 200;
  for (int threads = 128; threads <= 512; threads += 128) {
    body_kernel<MODE, RP><<<1, threads>>>(g, out, clk, reps);
    body_kernel<MODE, RP><<<1, threads>>>(g, out, clk, reps);
    cudaError_t e = cudaDeviceSynchronize();
    long long c = 0;
    cudaMemcpy(&c, clk, 8, cudaMemcpyDeviceToHost);
    const int cols = (MODE == 3) ? 48 : kD;
    const double pairs = double(reps) * cols * 2 * RP * threads;
    printf("%-28s warps/SMSP %d  %8lld clk  %.2f pairs/clk/SM (fp32 ceiling 32)  %s\n", name, threads / 128, c,
           pairs / double(c), e == cudaSuccess ? "" : cudaGetErrorString(e));
  }
}

int main() {
  float *g, *out;
  long long* clk;
  cudaMalloc(&g, 4096 * 4);
  cudaMalloc(&out, 1024 * 4);
  cudaMalloc(&clk, 64);
  float h[4096];
  for (int i = 0; i < 4096; ++i) h[i] = float((i * 2654435761u) >> 8) * (1.f / 16777216.f);
  cudaMemcpy(g, h, sizeof(h), cudaMemcpyHostToDevice);
  run<0, 13>("packed, 2 cols, 13 row pairs", g, out, clk);
  run<0, 7>("packed, 2 cols, 7 row pairs", g, out, clk);
  run<3, 13>("packed, 4 cols, 13 row pairs", g, out, clk);
  run<3, 7>("packed, 4 cols, 7 row pairs", g, out, clk);
  run<2, 13>("packed, no minima", g, out, clk);
  run<1, 13>("scalar, 1 col, 13 row pairs", g, out, clk);
  run<1, 7>("scalar, 1 col, 7 row pairs", g, out, clk);
  return 0;
}
