// Training support: the forward-with-saved-activations and BACKWARD kernels of
// ArtSpeech / SimpleArtSpeech (train_phoneme_to_articulation.py:84-96 runs loss.backward() through the
// model).  The inference path is fused and saves nothing; training runs this unfused fp32 path instead,
// assembled from a few primitive kernels by artspeech_b200/training.py:
//
//   asb_gemm            C (+)= op(A) . op(B) (+ bias, ReLU / sigmoid epilogue) with arbitrary strides: the three
//                       forms of a Linear layer (y = x W^T + b, dx = dy W, dW = dy^T x), FFMA on 64 x 64 tiles
//   asb_colsum          bias gradients
//   asb_layernorm_*     LayerNorm forward (saves x-hat, rstd) and backward (dx, dgamma, dbeta; optional ReLU mask)
//   asb_gru_train_*     one direction-pair GRU layer forward saving the gates, and BPTT through it with W_hh
//                       resident in shared memory (packed-sequence semantics, models.py:136-138)
//   asb_sigmoid_bwd, asb_relu_mask, asb_scatter_add_rows (embedding gradient), asb_dropout
//
// Everything is fp32 (gradients match the reference's fp32 autograd to ~1e-5 relative); none of it is on
// the benchmarked evaluation path, and no cuDNN / cuBLAS kernel is involved.
#include "asb_common.cuh"

namespace {

using namespace asb;

constexpr int kH = 128;
constexpr int kGH = 3 * kH;

// ---------------------------------------------------------------- generic strided GEMM (FFMA)
// C[i, j] (+)= sum_l A(i, l) B(l, j) (+ bias[j]) with A(i, l) = A[i * a_i + l * a_l], B(l, j) = B[l * b_l + j * b_j]
struct GemmArgs {
  const float* A;
  const float* B;
  float* C;
  const float* bias;
  long long M;
  int N, K;
  long long a_i, a_l, b_l, b_j, ldc;
  int act;         // 0 none, 1 ReLU, 2 sigmoid
  int accumulate;  // C += ...
};

constexpr int kTM = 64, kTN = 64, kTK = 16;

__global__ void __launch_bounds__(256) gemm_kernel(const GemmArgs a) {
  __shared__ __align__(16) float As[kTK][kTM + 4];
  __shared__ __align__(16) float Bs[kTK][kTN + 4];
  const int tid = threadIdx.x;
  const long long i0 = (long long)blockIdx.x * kTM;
  const int j0 = blockIdx.y * kTN;
  const int ty = tid >> 4, tx = tid & 15;   // 16 x 16 threads, 4 x 4 outputs each
  float acc[4][4] = {};
  for (int l0 = 0; l0 < a.K; l0 += kTK) {
    // 64 x 16 elements of A and 16 x 64 of B: 4 each per thread
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int q = tid + 256 * e;
      {
        // A tile: walk the faster-varying memory direction with consecutive threads
        const int ii = a.a_l == 1 ? q / kTK : q % kTM;
        const int ll = a.a_l == 1 ? q % kTK : q / kTM;
        const long long i = i0 + ii;
        const int l = l0 + ll;
        As[ll][ii] = (i < a.M && l < a.K) ? __ldg(a.A + i * a.a_i + l * a.a_l) : 0.f;
      }
      {
        const int jj = a.b_j == 1 ? q % kTN : q / kTK;
        const int ll = a.b_j == 1 ? q / kTN : q % kTK;
        const int j = j0 + jj;
        const int l = l0 + ll;
        Bs[ll][jj] = (j < a.N && l < a.K) ? __ldg(a.B + l * a.b_l + j * a.b_j) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int l = 0; l < kTK; ++l) {
      const float4 av = *reinterpret_cast<const float4*>(&As[l][4 * ty]);
      const float4 bv = *reinterpret_cast<const float4*>(&Bs[l][4 * tx]);
      const float ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(ar[r], br[c], acc[r][c]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const long long i = i0 + 4 * ty + r;
    if (i >= a.M) continue;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int j = j0 + 4 * tx + c;
      if (j >= a.N) continue;
      float v = acc[r][c];
      if (a.bias) v += __ldg(a.bias + j);
      float* dst = a.C + i * a.ldc + j;
      if (a.accumulate) v += *dst;
      if (a.act == 1) v = fmaxf(v, 0.f);
      if (a.act == 2) v = 1.f / (1.f + expf(-v));
      *dst = v;
    }
  }
}

// out[j] += sum_i X[i * ldx + j]  (out zeroed by the launcher unless accumulating); grid = (column tiles, row chunks)
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ X, long long M, int N, long long ldx,
                                                     float* __restrict__ out) {
  __shared__ float red[8][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + lane;
  float s = 0.f;
  if (j < N)
    for (long long i = (long long)blockIdx.y * 8 + w; i < M; i += 8LL * gridDim.y) s += X[i * ldx + j];
  red[w][lane] = s;
  __syncthreads();
  if (w == 0 && j < N) {
    float t = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += red[k][lane];
    atomicAdd(out + j, t);
  }
}

// ---------------------------------------------------------------- LayerNorm (biased variance, eps 1e-5)
// warp per row; C <= 256, C % 32 == 0
__global__ void __launch_bounds__(256) layernorm_fwd_kernel(const float* __restrict__ x, long long M, int C,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float* __restrict__ xhat,
                                                            float* __restrict__ rstd_out, float* __restrict__ y) {
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31, per = C >> 5;
  float v[8], s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[i] = i < per ? x[row * C + lane + 32 * i] : 0.f;
    s += v[i];
  }
  const float mean = warp_sum(s) / float(C);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[i] = i < per ? v[i] - mean : 0.f;
    q = fmaf(v[i], v[i], q);
  }
  const float rstd = rsqrtf(warp_sum(q) / float(C) + 1e-5f);
  if (lane == 0) rstd_out[row] = rstd;
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (i < per) {
      const int c = lane + 32 * i;
      const float xh = v[i] * rstd;
      xhat[row * C + c] = xh;
      y[row * C + c] = fmaf(xh, gamma[c], beta[c]);
    }
}

// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)), g = dy * gamma; optionally masked by relu_src > 0 (the ReLU
// that produced this LayerNorm's input) and optionally accumulated into dx.
__global__ void __launch_bounds__(256) layernorm_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ xhat,
                                                            const float* __restrict__ rstd,
                                                            const float* __restrict__ gamma,
                                                            const float* __restrict__ relu_src, long long M, int C,
                                                            float* __restrict__ dx, int accumulate) {
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= M) return;
  const int lane = threadIdx.x & 31, per = C >> 5;
  float g[8], xh[8], s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane + 32 * i;
    g[i] = i < per ? dy[row * C + c] * gamma[c] : 0.f;
    xh[i] = i < per ? xhat[row * C + c] : 0.f;
    s1 += g[i];
    s2 = fmaf(g[i], xh[i], s2);
  }
  const float m1 = warp_sum(s1) / float(C), m2 = warp_sum(s2) / float(C), rs = rstd[row];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (i < per) {
      const int c = lane + 32 * i;
      float v = rs * (g[i] - m1 - xh[i] * m2);
      if (relu_src && !(relu_src[row * C + c] > 0.f)) v = 0.f;
      float* dst = dx + row * C + c;
      *dst = accumulate ? *dst + v : v;
    }
}

// dgamma[c] += sum_rows dy * xhat, dbeta[c] += sum_rows dy  (grid = (column tiles, row chunks), outputs zeroed
// by the launcher unless accumulating)
__global__ void __launch_bounds__(256) layernorm_param_grad_kernel(const float* __restrict__ dy,
                                                                   const float* __restrict__ xhat, long long M, int C,
                                                                   float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float rg[8][33], rb[8][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float sg = 0.f, sb = 0.f;
  if (c < C)
    for (long long i = (long long)blockIdx.y * 8 + w; i < M; i += 8LL * gridDim.y) {
      const float d = dy[i * C + c];
      sg = fmaf(d, xhat[i * C + c], sg);
      sb += d;
    }
  rg[w][lane] = sg;
  rb[w][lane] = sb;
  __syncthreads();
  if (w == 0 && c < C) {
    float tg = 0.f, tb = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      tg += rg[k][lane];
      tb += rb[k][lane];
    }
    atomicAdd(dgamma + c, tg);
    atomicAdd(dbeta + c, tb);
  }
}

// ---------------------------------------------------------------- elementwise
__global__ void sigmoid_bwd_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dz,
                                   long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dz[i] = dy[i] * y[i] * (1.f - y[i]);
}
__global__ void relu_mask_kernel(float* __restrict__ dx, const float* __restrict__ f, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && !(f[i] > 0.f)) dx[i] = 0.f;
}
// dst[tokens[r], :] += src[r, :]
__global__ void scatter_add_rows_kernel(const float* __restrict__ src, const long long* __restrict__ tokens,
                                        long long rows, int width, int vocab, float* __restrict__ dst) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * width) return;
  const long long r = i / width;
  const int c = int(i - r * width);
  const long long t = tokens[r];
  if (t >= 0 && t < vocab) atomicAdd(dst + t * width + c, src[i]);
}
// inverted dropout with a counter-based generator: keep iff hash(seed, i) >= p * 2^32; out = x * keep / (1 - p)
__device__ __forceinline__ uint32_t mix32(uint64_t z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return uint32_t((z ^ (z >> 31)) >> 16);
}
__global__ void dropout_kernel(const float* __restrict__ x, float* __restrict__ out, long long n, float p,
                               unsigned long long seed) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint32_t threshold = uint32_t(fminf(fmaxf(p, 0.f), 1.f) * 4294967295.f);
  const bool keep = mix32(seed * 0x100000001b3ull + uint64_t(i)) >= threshold;
  out[i] = keep ? x[i] / (1.f - p) : 0.f;
}

// ---------------------------------------------------------------- GRU layer, training forward
// Same cell and packed-sequence semantics as gru_fp32_kernel (gru.cu); W_hh^T resident in shared memory;
// additionally saves, per (row, step, direction, unit): r, z, n and hn = W_hn h + b_hn (what BPTT needs).
// gi holds W_ih x + b_ih only (b_hh is added here); thread = hidden unit x 4 batch rows.
__global__ void __launch_bounds__(256, 1)
gru_train_fwd_kernel(const float* __restrict__ whh_fwd, const float* __restrict__ whh_bwd,
                     const float* __restrict__ bhh_fwd, const float* __restrict__ bhh_bwd,
                     const float* __restrict__ gi, const int* __restrict__ lengths, float* __restrict__ out,
                     float* __restrict__ gates, int B, int T) {
  extern __shared__ __align__(16) float smem[];
  float* W = smem;                  // [k][384] = W_hh^T
  float* hsm = smem + kH * kGH;     // [2][8][128]
  const int tid = threadIdx.x, u = tid & (kH - 1), half = tid >> 7;
  const int dir = blockIdx.y, b0 = blockIdx.x * 8;
  const float* whh = dir ? whh_bwd : whh_fwd;   // (384, 128) row-major like nn.GRU.weight_hh
  const float* bhh = dir ? bhh_bwd : bhh_fwd;
  for (int i = tid; i < kGH * kH; i += 256) {
    const int g = i / kH, k = i - g * kH;
    W[k * kGH + g] = whh[i];
  }
  const float br = bhh[u], bz = bhh[kH + u], bn = bhh[2 * kH + u];
  for (int i = tid; i < 2 * 8 * kH; i += 256) hsm[i] = 0.f;
  int len[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int b = b0 + 4 * half + r;
    len[r] = b < B ? lengths[b] : 0;
  }
  const int Lmax = lengths[b0];
  float h[4] = {0.f, 0.f, 0.f, 0.f};
  __syncthreads();
  int cur = 0;
  for (int s = 0; s < Lmax; ++s) {
    const int t = dir ? Lmax - 1 - s : s;
    float ar[4] = {br, br, br, br}, az[4] = {bz, bz, bz, bz}, an[4] = {bn, bn, bn, bn};
    const float* hc = hsm + cur * 8 * kH + 4 * half * kH;
#pragma unroll 2
    for (int k = 0; k < kH; k += 4) {
      float4 hv[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) hv[r] = *reinterpret_cast<const float4*>(hc + r * kH + k);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float wr = W[(k + kk) * kGH + u], wz = W[(k + kk) * kGH + kH + u], wn = W[(k + kk) * kGH + 2 * kH + u];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float hk = kk == 0 ? hv[r].x : kk == 1 ? hv[r].y : kk == 2 ? hv[r].z : hv[r].w;
          ar[r] = fmaf(wr, hk, ar[r]);
          az[r] = fmaf(wz, hk, az[r]);
          an[r] = fmaf(wn, hk, an[r]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const bool valid = t < len[r];
      if (valid) {
        const long long bt = (long long)(b0 + 4 * half + r) * T + t;
        const float* p = gi + (bt * 2 + dir) * kGH + u;
        const float rg = 1.f / (1.f + expf(-(p[0] + ar[r])));
        const float zg = 1.f / (1.f + expf(-(p[kH] + az[r])));
        const float ng = tanhf(p[2 * kH] + rg * an[r]);
        h[r] = (1.f - zg) * ng + zg * h[r];
        out[bt * (2 * kH) + dir * kH + u] = h[r];
        float* gs = gates + (bt * 2 + dir) * 4 * kH + u;
        gs[0] = rg; gs[kH] = zg; gs[2 * kH] = ng; gs[3 * kH] = an[r];
      }
      hsm[(cur ^ 1) * 8 * kH + (4 * half + r) * kH + u] = h[r];
    }
    __syncthreads();
    cur ^= 1;
  }
}

// BPTT through one GRU layer (both directions).  W_hh (384 x 128, row-major) resident in shared memory.
// Per step, thread (unit u, 4 rows): the cell's local derivatives from the saved gates -> dgi (r, z, n) and
// dgh (r, z, hn) written to global (the weight gradients are GEMMs over all steps afterwards) and to shared
// memory; then dh_prev[k = u] = dh * z + sum_g dgh[g] W_hh[g][k].
// dout: gradient wrt the layer output (B, T, 2H); out: the saved forward output (for h_prev).
__global__ void __launch_bounds__(256, 1)
gru_train_bwd_kernel(const float* __restrict__ whh_fwd, const float* __restrict__ whh_bwd,
                     const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ gates,
                     const int* __restrict__ lengths, float* __restrict__ dgi, float* __restrict__ dgh,
                     float* __restrict__ hprev, int B, int T) {
  extern __shared__ __align__(16) float smem[];
  float* W = smem;                  // [g][128]
  float* gsm = smem + kGH * kH;     // [8][384] dgh of the current step
  const int tid = threadIdx.x, u = tid & (kH - 1), half = tid >> 7;
  const int dir = blockIdx.y, b0 = blockIdx.x * 8;
  const float* whh = dir ? whh_bwd : whh_fwd;
  for (int i = tid; i < kGH * kH / 4; i += 256)
    reinterpret_cast<float4*>(W)[i] = __ldg(reinterpret_cast<const float4*>(whh) + i);
  int len[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int b = b0 + 4 * half + r;
    len[r] = b < B ? lengths[b] : 0;
  }
  const int Lmax = lengths[b0];
  float dh[4] = {0.f, 0.f, 0.f, 0.f};
  __syncthreads();
  for (int s = 0; s < Lmax; ++s) {
    // reverse of the forward order: the forward direction ran t = 0 .. L-1, the reverse one t = L-1 .. 0
    const int t = dir ? s : Lmax - 1 - s;
    float zkeep[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const bool valid = t < len[r];
      float d_r = 0.f, d_z = 0.f, d_n = 0.f, d_hn = 0.f;
      zkeep[r] = 0.f;
      if (valid) {
        const long long row = b0 + 4 * half + r;
        const long long bt = row * T + t;
        const int tp = dir ? t + 1 : t - 1;                      // the step the forward pass came from
        const float hp = (tp >= 0 && tp < len[r]) ? out[(row * T + tp) * (2 * kH) + dir * kH + u] : 0.f;
        const float* gs = gates + (bt * 2 + dir) * 4 * kH + u;
        const float rg = gs[0], zg = gs[kH], ng = gs[2 * kH], hn = gs[3 * kH];
        const float dht = dh[r] + dout[bt * (2 * kH) + dir * kH + u];
        const float dn = dht * (1.f - zg);
        const float dz = dht * (hp - ng);
        d_n = dn * (1.f - ng * ng);
        d_hn = d_n * rg;
        d_r = d_n * hn * rg * (1.f - rg);
        d_z = dz * zg * (1.f - zg);
        zkeep[r] = dht * zg;
        float* o1 = dgi + (bt * 2 + dir) * kGH + u;
        o1[0] = d_r; o1[kH] = d_z; o1[2 * kH] = d_n;
        float* o2 = dgh + (bt * 2 + dir) * kGH + u;
        o2[0] = d_r; o2[kH] = d_z; o2[2 * kH] = d_hn;
        hprev[(bt * 2 + dir) * kH + u] = hp;
      }
      float* g = gsm + (4 * half + r) * kGH + u;
      g[0] = d_r; g[kH] = d_z; g[2 * kH] = d_hn;
    }
    __syncthreads();
    float acc[4] = {zkeep[0], zkeep[1], zkeep[2], zkeep[3]};
    const float* gr = gsm + 4 * half * kGH;
#pragma unroll 4
    for (int g = 0; g < kGH; ++g) {
      const float w = W[g * kH + u];
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = fmaf(gr[r * kGH + g], w, acc[r]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) dh[r] = (t < len[r]) ? acc[r] : dh[r];
    __syncthreads();
  }
}

// ---------------------------------------------------------------- LSTM layer, training forward / backward
// nn.LSTM cell (gate rows i, f, g, o): c' = f c + i g, h' = o tanh(c').  W_hh (512 x 128 fp32 = 256 KB) does not fit
// a CTA's shared memory: three gates' worth (192 KB) lives there and the remaining 128 values each thread needs sit
// in its REGISTERS for the whole sequence (the trick of lstm_fp32_kernel, gru.cu).  gi holds W_ih x + b_ih; b_hh is
// added here.  Saved per (row, step, direction, unit): i, f, g, o, and c' (5 values).
constexpr int kLGH = 4 * kH;

__global__ void __launch_bounds__(256, 1)
lstm_train_fwd_kernel(const float* __restrict__ whh_fwd, const float* __restrict__ whh_bwd,
                      const float* __restrict__ bhh_fwd, const float* __restrict__ bhh_bwd,
                      const float* __restrict__ gi, const int* __restrict__ lengths, float* __restrict__ out,
                      float* __restrict__ gates, int B, int T) {
  extern __shared__ __align__(16) float smem[];
  float* W = smem;                  // [k][384] = (W_hh^T) columns of gates i, f, g
  float* hsm = smem + kH * kGH;     // [2][8][128]
  const int tid = threadIdx.x, u = tid & (kH - 1), half = tid >> 7;
  const int dir = blockIdx.y, b0 = blockIdx.x * 8;
  const float* whh = dir ? whh_bwd : whh_fwd;   // (512, 128) row-major like nn.LSTM.weight_hh
  const float* bhh = dir ? bhh_bwd : bhh_fwd;
  for (int i = tid; i < kGH * kH; i += 256) {
    const int g = i / kH, k = i - g * kH;
    W[k * kGH + g] = whh[i];
  }
  float wo[kH];
#pragma unroll
  for (int k = 0; k < kH; ++k) wo[k] = __ldg(whh + (size_t)(3 * kH + u) * kH + k);
  const float bi = bhh[u], bf = bhh[kH + u], bg = bhh[2 * kH + u], bo = bhh[3 * kH + u];
  for (int i = tid; i < 2 * 8 * kH; i += 256) hsm[i] = 0.f;
  int len[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int b = b0 + 4 * half + r;
    len[r] = b < B ? lengths[b] : 0;
  }
  const int Lmax = lengths[b0];
  float h[4] = {0.f, 0.f, 0.f, 0.f}, c[4] = {0.f, 0.f, 0.f, 0.f};
  __syncthreads();
  int cur = 0;
  for (int s = 0; s < Lmax; ++s) {
    const int t = dir ? Lmax - 1 - s : s;
    float ai[4] = {bi, bi, bi, bi}, af[4] = {bf, bf, bf, bf}, ag[4] = {bg, bg, bg, bg}, ao[4] = {bo, bo, bo, bo};
    const float* hc = hsm + cur * 8 * kH + 4 * half * kH;
#pragma unroll
    for (int k = 0; k < kH; k += 4) {
      float4 hv[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) hv[r] = *reinterpret_cast<const float4*>(hc + r * kH + k);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float wi = W[(k + kk) * kGH + u], wf = W[(k + kk) * kGH + kH + u], wg = W[(k + kk) * kGH + 2 * kH + u];
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const float hk = kk == 0 ? hv[r].x : kk == 1 ? hv[r].y : kk == 2 ? hv[r].z : hv[r].w;
          ai[r] = fmaf(wi, hk, ai[r]);
          af[r] = fmaf(wf, hk, af[r]);
          ag[r] = fmaf(wg, hk, ag[r]);
          ao[r] = fmaf(wo[k + kk], hk, ao[r]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (t < len[r]) {
        const long long bt = (long long)(b0 + 4 * half + r) * T + t;
        const float* p = gi + (bt * 2 + dir) * kLGH + u;
        const float ig = 1.f / (1.f + expf(-(p[0] + ai[r])));
        const float fg = 1.f / (1.f + expf(-(p[kH] + af[r])));
        const float gg = tanhf(p[2 * kH] + ag[r]);
        const float og = 1.f / (1.f + expf(-(p[3 * kH] + ao[r])));
        c[r] = fg * c[r] + ig * gg;
        h[r] = og * tanhf(c[r]);
        out[bt * (2 * kH) + dir * kH + u] = h[r];
        float* gs = gates + (bt * 2 + dir) * 5 * kH + u;
        gs[0] = ig; gs[kH] = fg; gs[2 * kH] = gg; gs[3 * kH] = og; gs[4 * kH] = c[r];
      }
      hsm[(cur ^ 1) * 8 * kH + (4 * half + r) * kH + u] = h[r];
    }
    __syncthreads();
    cur ^= 1;
  }
}

// BPTT through one LSTM layer.  Rows i, f, g of W_hh (384 x 128) in shared memory, row block o (128 x 128): thread u
// keeps COLUMN u of it (W_hh[384 + g][u], g < 128) in registers.  Per step: local derivatives -> dgi (= dgh: every
// gate pre-activation is gi + W_hh h + b_hh) to global and to shared memory, then dh_prev = dpre . W_hh and
// dc_prev = dc f.  cprev of a step is the saved c' of the step the forward pass came from.
__global__ void __launch_bounds__(256, 1)
lstm_train_bwd_kernel(const float* __restrict__ whh_fwd, const float* __restrict__ whh_bwd,
                      const float* __restrict__ dout, const float* __restrict__ out, const float* __restrict__ gates,
                      const int* __restrict__ lengths, float* __restrict__ dgi, float* __restrict__ hprev, int B, int T) {
  extern __shared__ __align__(16) float smem[];
  float* W = smem;                  // [g < 384][128]
  float* gsm = smem + kGH * kH;     // [8][512] dpre of the current step
  const int tid = threadIdx.x, u = tid & (kH - 1), half = tid >> 7;
  const int dir = blockIdx.y, b0 = blockIdx.x * 8;
  const float* whh = dir ? whh_bwd : whh_fwd;
  for (int i = tid; i < kGH * kH / 4; i += 256)
    reinterpret_cast<float4*>(W)[i] = __ldg(reinterpret_cast<const float4*>(whh) + i);
  float wo[kH];
#pragma unroll
  for (int g = 0; g < kH; ++g) wo[g] = __ldg(whh + (size_t)(3 * kH + g) * kH + u);
  int len[4];
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int b = b0 + 4 * half + r;
    len[r] = b < B ? lengths[b] : 0;
  }
  const int Lmax = lengths[b0];
  float dh[4] = {0.f, 0.f, 0.f, 0.f}, dc[4] = {0.f, 0.f, 0.f, 0.f};
  __syncthreads();
  for (int s = 0; s < Lmax; ++s) {
    const int t = dir ? s : Lmax - 1 - s;
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      float d_i = 0.f, d_f = 0.f, d_g = 0.f, d_o = 0.f;
      if (t < len[r]) {
        const long long row = b0 + 4 * half + r;
        const long long bt = row * T + t;
        const int tp = dir ? t + 1 : t - 1;
        const bool has_prev = tp >= 0 && tp < len[r];
        const float hp = has_prev ? out[(row * T + tp) * (2 * kH) + dir * kH + u] : 0.f;
        const float cp = has_prev ? gates[((row * T + tp) * 2 + dir) * 5 * kH + 4 * kH + u] : 0.f;
        const float* gs = gates + (bt * 2 + dir) * 5 * kH + u;
        const float ig = gs[0], fg = gs[kH], gg = gs[2 * kH], og = gs[3 * kH], cn = gs[4 * kH];
        const float tc = tanhf(cn);
        const float dht = dh[r] + dout[bt * (2 * kH) + dir * kH + u];
        const float dct = dc[r] + dht * og * (1.f - tc * tc);
        d_o = dht * tc * og * (1.f - og);
        d_i = dct * gg * ig * (1.f - ig);
        d_f = dct * cp * fg * (1.f - fg);
        d_g = dct * ig * (1.f - gg * gg);
        dc[r] = dct * fg;
        float* o1 = dgi + (bt * 2 + dir) * kLGH + u;
        o1[0] = d_i; o1[kH] = d_f; o1[2 * kH] = d_g; o1[3 * kH] = d_o;
        hprev[(bt * 2 + dir) * kH + u] = hp;
      }
      float* g = gsm + (4 * half + r) * kLGH + u;
      g[0] = d_i; g[kH] = d_f; g[2 * kH] = d_g; g[3 * kH] = d_o;
    }
    __syncthreads();
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    const float* gr = gsm + 4 * half * kLGH;
#pragma unroll 4
    for (int g = 0; g < kGH; ++g) {
      const float w = W[g * kH + u];
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = fmaf(gr[r * kLGH + g], w, acc[r]);
    }
#pragma unroll
    for (int g = 0; g < kH; ++g) {
#pragma unroll
      for (int r = 0; r < 4; ++r) acc[r] = fmaf(gr[r * kLGH + 3 * kH + g], wo[g], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < 4; ++r) dh[r] = (t < len[r]) ? acc[r] : dh[r];
    __syncthreads();
  }
}

int grid_1d(long long n, int block, unsigned* out) {
  const long long blocks = (n + block - 1) / block;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "too many elements");
  *out = unsigned(blocks);
  return ASB_OK;
}

}  // namespace

extern "C" int asb_gemm(const float* A, const float* B, float* C, int64_t M, int N, int K, int64_t a_i, int64_t a_l,
                        int64_t b_l, int64_t b_j, int64_t ldc, const float* bias, int act, int accumulate,
                        void* stream) {
  if (M < 0 || N <= 0 || K <= 0 || act < 0 || act > 2) return asb::set_error(ASB_ERR_INVALID, "asb_gemm: bad shape");
  if (M == 0) return ASB_OK;
  if (!A || !B || !C) return asb::set_error(ASB_ERR_INVALID, "asb_gemm: null pointer");
  GemmArgs a{A, B, C, bias, M, N, K, a_i, a_l, b_l, b_j, ldc, act, accumulate};
  const long long gx = (M + kTM - 1) / kTM;
  if (gx > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_gemm: M too large");
  dim3 grid(unsigned(gx), unsigned((N + kTN - 1) / kTN));
  gemm_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(a);
  ASB_CHECK_LAUNCH("gemm_kernel");
  return ASB_OK;
}

extern "C" int asb_colsum(const float* X, int64_t M, int N, int64_t ldx, float* out, int accumulate, void* stream) {
  if (M < 0 || N <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_colsum: bad shape");
  if (!X || !out) return asb::set_error(ASB_ERR_INVALID, "asb_colsum: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (!accumulate)
    if (int rc = asb::check_cuda(cudaMemsetAsync(out, 0, sizeof(float) * N, st), "asb_colsum: memset")) return rc;
  const unsigned chunks = unsigned(M >= 8192 ? 64 : (M + 127) / 128);
  colsum_kernel<<<dim3((N + 31) / 32, chunks ? chunks : 1), 256, 0, st>>>(X, M, N, ldx, out);
  ASB_CHECK_LAUNCH("colsum_kernel");
  return ASB_OK;
}

extern "C" int asb_layernorm_fwd(const float* x, int64_t M, int C, const float* gamma, const float* beta, float* xhat,
                                 float* rstd, float* y, void* stream) {
  if (M < 0 || C <= 0 || C > 256 || (C & 31)) return asb::set_error(ASB_ERR_INVALID, "asb_layernorm_fwd: C must be a multiple of 32 up to 256");
  if (M == 0) return ASB_OK;
  if (!x || !gamma || !beta || !xhat || !rstd || !y) return asb::set_error(ASB_ERR_INVALID, "asb_layernorm_fwd: null pointer");
  unsigned g;
  if (int rc = grid_1d(M, 8, &g)) return rc;
  layernorm_fwd_kernel<<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, M, C, gamma, beta, xhat, rstd, y);
  ASB_CHECK_LAUNCH("layernorm_fwd_kernel");
  return ASB_OK;
}

extern "C" int asb_layernorm_bwd(const float* dy, const float* xhat, const float* rstd, const float* gamma,
                                 const float* relu_src, int64_t M, int C, float* dx, int accumulate_dx, float* dgamma,
                                 float* dbeta, int accumulate_params, void* stream) {
  if (M < 0 || C <= 0 || C > 256 || (C & 31)) return asb::set_error(ASB_ERR_INVALID, "asb_layernorm_bwd: C must be a multiple of 32 up to 256");
  if (M == 0) return ASB_OK;
  if (!dy || !xhat || !rstd || !gamma || !dx) return asb::set_error(ASB_ERR_INVALID, "asb_layernorm_bwd: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dgamma && dbeta) {   // before dx: dx may alias dy
    if (!accumulate_params) {
      if (int rc = asb::check_cuda(cudaMemsetAsync(dgamma, 0, sizeof(float) * C, st), "asb_layernorm_bwd: memset")) return rc;
      if (int rc = asb::check_cuda(cudaMemsetAsync(dbeta, 0, sizeof(float) * C, st), "asb_layernorm_bwd: memset")) return rc;
    }
    const unsigned chunks = unsigned(M >= 8192 ? 64 : (M + 127) / 128);
    layernorm_param_grad_kernel<<<dim3((C + 31) / 32, chunks ? chunks : 1), 256, 0, st>>>(dy, xhat, M, C, dgamma, dbeta);
    ASB_CHECK_LAUNCH("layernorm_param_grad_kernel");
  }
  unsigned g;
  if (int rc = grid_1d(M, 8, &g)) return rc;
  layernorm_bwd_kernel<<<g, 256, 0, st>>>(dy, xhat, rstd, gamma, relu_src, M, C, dx, accumulate_dx);
  ASB_CHECK_LAUNCH("layernorm_bwd_kernel");
  return ASB_OK;
}

extern "C" int asb_sigmoid_bwd(const float* dy, const float* y, float* dz, int64_t n, void* stream) {
  if (n < 0) return asb::set_error(ASB_ERR_INVALID, "asb_sigmoid_bwd: bad size");
  if (n == 0) return ASB_OK;
  unsigned g;
  if (int rc = grid_1d(n, 256, &g)) return rc;
  sigmoid_bwd_kernel<<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, y, dz, n);
  ASB_CHECK_LAUNCH("sigmoid_bwd_kernel");
  return ASB_OK;
}

extern "C" int asb_relu_mask(float* dx, const float* f, int64_t n, void* stream) {
  if (n < 0) return asb::set_error(ASB_ERR_INVALID, "asb_relu_mask: bad size");
  if (n == 0) return ASB_OK;
  unsigned g;
  if (int rc = grid_1d(n, 256, &g)) return rc;
  relu_mask_kernel<<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(dx, f, n);
  ASB_CHECK_LAUNCH("relu_mask_kernel");
  return ASB_OK;
}

extern "C" int asb_scatter_add_rows(const float* src, const int64_t* tokens, int64_t rows, int width, int vocab,
                                    float* dst, void* stream) {
  if (rows < 0 || width <= 0 || vocab <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_scatter_add_rows: bad shape");
  if (rows == 0) return ASB_OK;
  unsigned g;
  if (int rc = grid_1d(rows * width, 256, &g)) return rc;
  scatter_add_rows_kernel<<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, reinterpret_cast<const long long*>(tokens), rows, width, vocab, dst);
  ASB_CHECK_LAUNCH("scatter_add_rows_kernel");
  return ASB_OK;
}

extern "C" int asb_dropout(const float* x, float* out, int64_t n, float p, uint64_t seed, void* stream) {
  if (n < 0 || !(p >= 0.f) || !(p < 1.f)) return asb::set_error(ASB_ERR_INVALID, "asb_dropout: need 0 <= p < 1");
  if (n == 0) return ASB_OK;
  unsigned g;
  if (int rc = grid_1d(n, 256, &g)) return rc;
  dropout_kernel<<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(x, out, n, p, seed);
  ASB_CHECK_LAUNCH("dropout_kernel");
  return ASB_OK;
}

extern "C" int asb_gru_train_forward(const float* whh_fwd, const float* whh_bwd, const float* bhh_fwd,
                                     const float* bhh_bwd, const float* gi, const int32_t* lengths_dev, int B, int T,
                                     int hidden, float* out, float* gates, void* stream) {
  if (hidden != kH) return asb::set_error(ASB_ERR_UNSUPPORTED, "asb_gru_train_forward: hidden_size must be %d", kH);
  if (B <= 0 || T <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_gru_train_forward: empty batch");
  if (!whh_fwd || !whh_bwd || !bhh_fwd || !bhh_bwd || !gi || !lengths_dev || !out || !gates)
    return asb::set_error(ASB_ERR_INVALID, "asb_gru_train_forward: null pointer");
  const int smem = (kH * kGH + 2 * 8 * kH) * int(sizeof(float));
  static asb::DeviceOnce once;
  if (int rc = once.run([&] {
        return asb::check_cuda(cudaFuncSetAttribute(gru_train_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                               "gru_train_fwd_kernel smem attribute");
      }))
    return rc;
  dim3 grid((B + 7) / 8, 2);
  gru_train_fwd_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(whh_fwd, whh_bwd, bhh_fwd, bhh_bwd, gi,
                                                                              lengths_dev, out, gates, B, T);
  ASB_CHECK_LAUNCH("gru_train_fwd_kernel");
  return ASB_OK;
}

extern "C" int asb_gru_train_backward(const float* whh_fwd, const float* whh_bwd, const float* dout, const float* out,
                                      const float* gates, const int32_t* lengths_dev, int B, int T, int hidden,
                                      float* dgi, float* dgh, float* hprev, void* stream) {
  if (hidden != kH) return asb::set_error(ASB_ERR_UNSUPPORTED, "asb_gru_train_backward: hidden_size must be %d", kH);
  if (B <= 0 || T <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_gru_train_backward: empty batch");
  if (!whh_fwd || !whh_bwd || !dout || !out || !gates || !lengths_dev || !dgi || !dgh || !hprev)
    return asb::set_error(ASB_ERR_INVALID, "asb_gru_train_backward: null pointer");
  const int smem = (kGH * kH + 8 * kGH) * int(sizeof(float));
  static asb::DeviceOnce once;
  if (int rc = once.run([&] {
        return asb::check_cuda(cudaFuncSetAttribute(gru_train_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                               "gru_train_bwd_kernel smem attribute");
      }))
    return rc;
  dim3 grid((B + 7) / 8, 2);
  gru_train_bwd_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(whh_fwd, whh_bwd, dout, out, gates,
                                                                              lengths_dev, dgi, dgh, hprev, B, T);
  ASB_CHECK_LAUNCH("gru_train_bwd_kernel");
  return ASB_OK;
}

extern "C" int asb_lstm_train_forward(const float* whh_fwd, const float* whh_bwd, const float* bhh_fwd,
                                      const float* bhh_bwd, const float* gi, const int32_t* lengths_dev, int B, int T,
                                      int hidden, float* out, float* gates, void* stream) {
  if (hidden != kH) return asb::set_error(ASB_ERR_UNSUPPORTED, "asb_lstm_train_forward: hidden_size must be %d", kH);
  if (B <= 0 || T <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_lstm_train_forward: empty batch");
  if (!whh_fwd || !whh_bwd || !bhh_fwd || !bhh_bwd || !gi || !lengths_dev || !out || !gates)
    return asb::set_error(ASB_ERR_INVALID, "asb_lstm_train_forward: null pointer");
  const int smem = (kH * kGH + 2 * 8 * kH) * int(sizeof(float));
  static asb::DeviceOnce once;
  if (int rc = once.run([&] {
        return asb::check_cuda(cudaFuncSetAttribute(lstm_train_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                               "lstm_train_fwd_kernel smem attribute");
      }))
    return rc;
  dim3 grid((B + 7) / 8, 2);
  lstm_train_fwd_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(whh_fwd, whh_bwd, bhh_fwd, bhh_bwd, gi,
                                                                               lengths_dev, out, gates, B, T);
  ASB_CHECK_LAUNCH("lstm_train_fwd_kernel");
  return ASB_OK;
}

extern "C" int asb_lstm_train_backward(const float* whh_fwd, const float* whh_bwd, const float* dout, const float* out,
                                       const float* gates, const int32_t* lengths_dev, int B, int T, int hidden,
                                       float* dgi, float* hprev, void* stream) {
  if (hidden != kH) return asb::set_error(ASB_ERR_UNSUPPORTED, "asb_lstm_train_backward: hidden_size must be %d", kH);
  if (B <= 0 || T <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_lstm_train_backward: empty batch");
  if (!whh_fwd || !whh_bwd || !dout || !out || !gates || !lengths_dev || !dgi || !hprev)
    return asb::set_error(ASB_ERR_INVALID, "asb_lstm_train_backward: null pointer");
  const int smem = (kGH * kH + 8 * kLGH) * int(sizeof(float));
  static asb::DeviceOnce once;
  if (int rc = once.run([&] {
        return asb::check_cuda(cudaFuncSetAttribute(lstm_train_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                               "lstm_train_bwd_kernel smem attribute");
      }))
    return rc;
  dim3 grid((B + 7) / 8, 2);
  lstm_train_bwd_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(whh_fwd, whh_bwd, dout, out, gates,
                                                                               lengths_dev, dgi, hprev, B, T);
  ASB_CHECK_LAUNCH("lstm_train_bwd_kernel");
  return ASB_OK;
}
