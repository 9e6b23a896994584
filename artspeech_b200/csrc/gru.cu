// Bidirectional GRU layer with the recurrent weights resident on chip for all T steps.
//
// Replaces nn.GRU(64, 128, num_layers=2, bidirectional=True) + pack/pad_packed_sequence of the
// reference (phoneme_to_articulation/encoder_decoder/models.py:111,136-138).  Cell equations
// (gate rows r, z, n):  r = s(gi_r + gh_r), z = s(gi_z + gh_z), n = tanh(gi_n + r * gh_n),
// h' = (1 - z) * n + z * h, with gi = W_ih x + b_ih (computed ahead of time for all steps) and
// gh = W_hh h + b_hh.  Packed-sequence semantics: a row stops updating at its own length, the
// reverse direction starts at each row's last valid step, padded outputs are zero.
//
// One CTA owns 8*NT batch rows of one direction and walks the time steps without any inter-CTA
// communication.  Fast mode: gh^T = W_hh . h^T on mma.sync.m16n8k16 (bf16 in, fp32 accumulate)
// with W_hh held in REGISTERS as A fragments (96 registers / thread) for the whole sequence,
// h kept in fp32 registers and mirrored as bf16 in a double-buffered shared tile that feeds the
// B fragments; warp w owns hidden units [16w, 16w+16) of all three gates so the gate math is
// thread-local in the MMA epilogue.  fp32 mode: W_hh^T (192 KB) resident in shared memory, FFMA.
#include <cuda_bf16.h>

#include "model.cuh"

namespace {

using namespace asb;

constexpr int kG = 3;             // GRU gates
constexpr int kGH = kG * kH;      // 384
constexpr int kHPad = kH + 8;     // bf16 row stride of the h tile (conflict-free B-fragment loads)

__device__ __forceinline__ float sigmoidf_fast(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanhf_fast(float x) { return 2.f * sigmoidf_fast(2.f * x) - 1.f; }

__device__ __forceinline__ void mma_bf16_16816(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// ---------------------------------------------------------------- weight packing
// A-fragment order: frag[(((w*3 + gate)*8 + kt)*32 + lane)*4 + r]
__global__ void pack_whh_frag_kernel(const float* __restrict__ whh, uint32_t* __restrict__ frag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 8 * kG * 8 * 32 * 4) return;
  const int r = i & 3, lane = (i >> 2) & 31, kt = (i >> 7) & 7, wg = i >> 10;
  const int gate = wg % kG, w = wg / kG;
  const int g = lane >> 2, c = lane & 3;
  const int row = gate * kH + 16 * w + g + ((r & 1) ? 8 : 0);
  const int k = kt * 16 + 2 * c + ((r & 2) ? 8 : 0);
  const __nv_bfloat16 lo = __float2bfloat16_rn(whh[row * kH + k]);
  const __nv_bfloat16 hi = __float2bfloat16_rn(whh[row * kH + k + 1]);
  frag[i] = uint32_t(__bfloat16_as_ushort(lo)) | (uint32_t(__bfloat16_as_ushort(hi)) << 16);
}

__global__ void transpose_whh_kernel(const float* __restrict__ whh, float* __restrict__ wt, int GH) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;  // over [k][row]
  if (i >= kH * GH) return;
  const int k = i / GH, row = i - k * GH;
  wt[i] = whh[row * kH + k];
}

// tab0[dir][v][row] = sum_e E[v][e] * W_ih[row][e] + b_ih[row] (+ b_hh[row] for r, z)
// GH = gates * H; rows below `fold` also take b_hh (GRU: r, z; LSTM: every gate)
__global__ void build_tab0_kernel(const float* __restrict__ emb, const float* __restrict__ wih,
                                  const float* __restrict__ bih, const float* __restrict__ bhh,
                                  float* __restrict__ tab, int V, int E, int GH, int fold) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= V * GH) return;
  const int v = i / GH, row = i - v * GH;
  double acc = 0.0;
  for (int e = 0; e < E; ++e) acc += double(emb[v * E + e]) * double(wih[row * E + e]);
  float r = float(acc) + bih[row];
  if (row < fold) r += bhh[row];
  tab[i] = r;
}

__global__ void copy_slice_kernel(const float* __restrict__ src, float* __restrict__ dst, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dst[i] = src[i];
}

// gi[(b*T + t)][dir][:] = tab0[dir][x[b][t]][:]   (one float4 per thread)
__global__ void gather_gi0_kernel(const float* __restrict__ tab, const long long* __restrict__ x, int T_in,
                                  float* __restrict__ gi, int B, int T, int V, int GH, int* __restrict__ tok_flag) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)B * T * 2 * (GH / 4);
  if (i >= total) return;
  const int q = int(i % (GH / 4));
  const long long rest = i / (GH / 4);
  const int dir = int(rest & 1);
  const long long bt = rest >> 1;
  const int b = int(bt / T), t = int(bt - (long long)b * T);
  long long tok = x[(long long)b * T_in + t];
  if (tok < 0 || tok >= V) {   // nn.Embedding raises here; report it (model.cuh), then clamp
    if (q == 0 && dir == 0) {
      tok_flag[1] = int(tok);
      tok_flag[0] = 1;
    }
    tok = tok < 0 ? 0 : V - 1;
  }
  const float4 val = __ldg(reinterpret_cast<const float4*>(tab + ((long long)dir * V + tok) * GH) + q);
  reinterpret_cast<float4*>(gi)[i] = val;
}

// ---------------------------------------------------------------- fast mode (bf16 MMA)
template <int NT>
__global__ void __launch_bounds__(256, 1)
gru_fast_kernel(const uint32_t* __restrict__ frag_fwd, const uint32_t* __restrict__ frag_bwd,
                const float* __restrict__ bhhn_fwd, const float* __restrict__ bhhn_bwd,
                const float* __restrict__ gi, const int* __restrict__ lengths, float* __restrict__ out, int B,
                int T) {
  __shared__ __align__(16) __nv_bfloat16 hs[2][8 * NT][kHPad];
  const int tid = threadIdx.x, w = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int dir = blockIdx.y;
  const int b0 = blockIdx.x * 8 * NT;
  const uint32_t* frag = dir ? frag_bwd : frag_fwd;
  const float* bhhn = dir ? bhhn_bwd : bhhn_fwd;

  // W_hh as register-resident A fragments: [gate][k tile][4]
  uint32_t a[kG][8][4];
#pragma unroll
  for (int gate = 0; gate < kG; ++gate)
#pragma unroll
    for (int kt = 0; kt < 8; ++kt) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(frag) + ((w * kG + gate) * 8 + kt) * 32 + lane);
      a[gate][kt][0] = v.x; a[gate][kt][1] = v.y; a[gate][kt][2] = v.z; a[gate][kt][3] = v.w;
    }
  const int u0 = 16 * w + g;  // this thread's hidden units: u0 (e = 0, 1) and u0 + 8 (e = 2, 3)
  const float bn0 = bhhn[u0], bn1 = bhhn[u0 + 8];

  int len[NT][2];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int col = 0; col < 2; ++col) {
      const int b = b0 + nt * 8 + 2 * c + col;
      len[nt][col] = b < B ? lengths[b] : 0;
    }
  const int Lmax = lengths[b0];  // sorted descending: the tile's first row is its longest

  for (int i = tid; i < 2 * 8 * NT * kHPad; i += 256) (&hs[0][0][0])[i] = __float2bfloat16_rn(0.f);
  float h[NT][4];
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
#pragma unroll
    for (int e = 0; e < 4; ++e) h[nt][e] = 0.f;
  __syncthreads();

  float gin[NT][4][kG];
  auto load_gi = [&](int t) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int b = b0 + nt * 8 + 2 * c + (e & 1);
        const int u = u0 + ((e & 2) ? 8 : 0);
        if (t < len[nt][e & 1]) {
          const float* p = gi + (((long long)b * T + t) * 2 + dir) * kGH + u;
          gin[nt][e][0] = __ldg(p);
          gin[nt][e][1] = __ldg(p + kH);
          gin[nt][e][2] = __ldg(p + 2 * kH);
        } else {
          gin[nt][e][0] = gin[nt][e][1] = gin[nt][e][2] = 0.f;
        }
      }
  };
  load_gi(dir ? Lmax - 1 : 0);

  int cur = 0;
  for (int s = 0; s < Lmax; ++s) {
    const int t = dir ? Lmax - 1 - s : s;
    float gc[NT][4][kG];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e)
#pragma unroll
        for (int q = 0; q < kG; ++q) gc[nt][e][q] = gin[nt][e][q];
    if (s + 1 < Lmax) load_gi(dir ? t - 1 : t + 1);  // in flight during the MMA chain

    float acc[kG][NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[0][nt][e] = acc[1][nt][e] = 0.f;
      acc[2][nt][0] = acc[2][nt][1] = bn0;  // b_hn rides in the accumulator
      acc[2][nt][2] = acc[2][nt][3] = bn1;
    }
#pragma unroll
    for (int kt = 0; kt < 8; ++kt)
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const uint32_t f0 = *reinterpret_cast<const uint32_t*>(&hs[cur][nt * 8 + g][kt * 16 + 2 * c]);
        const uint32_t f1 = *reinterpret_cast<const uint32_t*>(&hs[cur][nt * 8 + g][kt * 16 + 2 * c + 8]);
#pragma unroll
        for (int gate = 0; gate < kG; ++gate) mma_bf16_16816(acc[gate][nt], a[gate][kt], f0, f1);
      }

    // gates, state update, outputs (all thread-local)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int n = 2 * c + (e & 1);
        const int u = u0 + ((e & 2) ? 8 : 0);
        const float r = sigmoidf_fast(gc[nt][e][0] + acc[0][nt][e]);
        const float z = sigmoidf_fast(gc[nt][e][1] + acc[1][nt][e]);
        const float nn = tanhf_fast(gc[nt][e][2] + r * acc[2][nt][e]);
        const float hn = nn + z * (h[nt][e] - nn);
        const bool valid = t < len[nt][e & 1];
        h[nt][e] = valid ? hn : h[nt][e];
        hs[cur ^ 1][nt * 8 + n][u] = __float2bfloat16_rn(h[nt][e]);
        if (valid) out[((long long)(b0 + nt * 8 + n) * T + t) * (2 * kH) + dir * kH + u] = h[nt][e];
      }
    __syncthreads();
    cur ^= 1;
  }
}

// ---------------------------------------------------------------- fp32 mode (FFMA, W_hh^T in smem)
// R = batch rows per thread (the CTA owns 2 R rows): the step is bound by the FFMA pipe (3 x 128 FMAs per row and
// thread), so the launcher takes R = 2 while that still fits one wave of the machine - twice the CTAs, half the step.
template <int R>
__global__ void __launch_bounds__(256, 1)
gru_fp32_kernel(const float* __restrict__ wt_fwd, const float* __restrict__ wt_bwd,
                const float* __restrict__ bhhn_fwd, const float* __restrict__ bhhn_bwd,
                const float* __restrict__ gi, const int* __restrict__ lengths, float* __restrict__ out, int B,
                int T) {
  extern __shared__ __align__(16) float smem[];
  float* W = smem;                  // [k][384]
  float* hsm = smem + kH * kGH;     // [2][2 R][128]
  const int tid = threadIdx.x, u = tid & (kH - 1), half = tid >> 7;
  const int dir = blockIdx.y, b0 = blockIdx.x * 2 * R;
  const float* wt = dir ? wt_bwd : wt_fwd;
  const float bn = (dir ? bhhn_bwd : bhhn_fwd)[u];
  for (int i = tid; i < kH * kGH / 4; i += 256)
    reinterpret_cast<float4*>(W)[i] = __ldg(reinterpret_cast<const float4*>(wt) + i);
  for (int i = tid; i < 2 * 2 * R * kH; i += 256) hsm[i] = 0.f;
  int len[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const int b = b0 + R * half + r;
    len[r] = b < B ? lengths[b] : 0;
  }
  const int Lmax = lengths[b0];
  float h[R];
#pragma unroll
  for (int r = 0; r < R; ++r) h[r] = 0.f;
  __syncthreads();
  int cur = 0;
  for (int s = 0; s < Lmax; ++s) {
    const int t = dir ? Lmax - 1 - s : s;
    float gr[R], gz[R], gn[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (t < len[r]) {
        const float* p = gi + (((long long)(b0 + R * half + r) * T + t) * 2 + dir) * kGH + u;
        gr[r] = __ldg(p); gz[r] = __ldg(p + kH); gn[r] = __ldg(p + 2 * kH);
      } else {
        gr[r] = gz[r] = gn[r] = 0.f;
      }
    }
    float ar[R], az[R], an[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      ar[r] = 0.f;
      az[r] = 0.f;
      an[r] = bn;
    }
    const float* hc = hsm + cur * 2 * R * kH + R * half * kH;
#pragma unroll 2
    for (int k = 0; k < kH; k += 4) {
      float4 hv[R];
#pragma unroll
      for (int r = 0; r < R; ++r) hv[r] = *reinterpret_cast<const float4*>(hc + r * kH + k);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
        const float wr = W[(k + kk) * kGH + u], wz = W[(k + kk) * kGH + kH + u], wn = W[(k + kk) * kGH + 2 * kH + u];
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const float hk = kk == 0 ? hv[r].x : kk == 1 ? hv[r].y : kk == 2 ? hv[r].z : hv[r].w;
          ar[r] = fmaf(wr, hk, ar[r]);
          az[r] = fmaf(wz, hk, az[r]);
          an[r] = fmaf(wn, hk, an[r]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float rg = 1.f / (1.f + expf(-(gr[r] + ar[r])));
      const float zg = 1.f / (1.f + expf(-(gz[r] + az[r])));
      const float ng = tanhf(gn[r] + rg * an[r]);
      const float hn = (1.f - zg) * ng + zg * h[r];
      const bool valid = t < len[r];
      h[r] = valid ? hn : h[r];
      hsm[(cur ^ 1) * 2 * R * kH + (R * half + r) * kH + u] = h[r];
      if (valid) out[((long long)(b0 + R * half + r) * T + t) * (2 * kH) + dir * kH + u] = h[r];
    }
    __syncthreads();
    cur ^= 1;
  }
}

// ---------------------------------------------------------------- LSTM in fp32 (RNNType.LSTM,
// phoneme_to_articulation/__init__.py:47-49; used by principal_components/models/rnn.py:58-68)
// nn.LSTM cell, gate rows i, f, g, o:  i = s(.), f = s(.), g = tanh(.), o = s(.), c' = f c + i g,
// h' = o tanh(c'); gi already holds W_ih x + b_ih + b_hh.  W_hh^T is 256 KB in fp32 - more than one
// CTA's shared memory - so the columns of gates i, f, g (192 KB) live in shared memory and the column of
// gate o of the thread's own hidden unit lives in 128 REGISTERS for the whole sequence.  Same
// packed-sequence semantics and thread mapping as gru_fp32_kernel (thread = hidden unit x 4 batch rows).
constexpr int kLG = 4, kLGH = kLG * kH;   // 512

__global__ void __launch_bounds__(256, 1)
lstm_fp32_kernel(const float* __restrict__ wt_fwd, const float* __restrict__ wt_bwd, const float* __restrict__ gi,
                 const int* __restrict__ lengths, float* __restrict__ out, int B, int T) {
  extern __shared__ __align__(16) float smem[];
  float* W = smem;                  // [k][384]: gates i, f, g
  float* hsm = smem + kH * kGH;     // [2][8][128]
  const int tid = threadIdx.x, u = tid & (kH - 1), half = tid >> 7;
  const int dir = blockIdx.y, b0 = blockIdx.x * 8;
  const float* wt = dir ? wt_bwd : wt_fwd;   // [k][512]
  for (int i = tid; i < kH * (kGH / 4); i += 256) {
    const int k = i / (kGH / 4), q = i - k * (kGH / 4);
    reinterpret_cast<float4*>(W)[i] = __ldg(reinterpret_cast<const float4*>(wt + (size_t)k * kLGH) + q);
  }
  float wo[kH];
#pragma unroll
  for (int k = 0; k < kH; ++k) wo[k] = __ldg(wt + (size_t)k * kLGH + 3 * kH + u);
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
    float ai[4], af[4], ag[4], ao[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      if (t < len[r]) {
        const float* p = gi + (((long long)(b0 + 4 * half + r) * T + t) * 2 + dir) * kLGH + u;
        ai[r] = __ldg(p); af[r] = __ldg(p + kH); ag[r] = __ldg(p + 2 * kH); ao[r] = __ldg(p + 3 * kH);
      } else {
        ai[r] = af[r] = ag[r] = ao[r] = 0.f;
      }
    }
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
      const float ig = 1.f / (1.f + expf(-ai[r]));
      const float fg = 1.f / (1.f + expf(-af[r]));
      const float gg = tanhf(ag[r]);
      const float og = 1.f / (1.f + expf(-ao[r]));
      const float cn = fg * c[r] + ig * gg;
      const float hn = og * tanhf(cn);
      const bool valid = t < len[r];
      c[r] = valid ? cn : c[r];
      h[r] = valid ? hn : h[r];
      hsm[(cur ^ 1) * 8 * kH + (4 * half + r) * kH + u] = h[r];
      if (valid) out[((long long)(b0 + 4 * half + r) * T + t) * (2 * kH) + dir * kH + u] = h[r];
    }
    __syncthreads();
    cur ^= 1;
  }
}

}  // namespace

namespace asb {

int launch_gather_gi0(const asb_model* m, const long long* x, int T_in, float* gi, int B, int T, cudaStream_t st) {
  const int GH = m->G * kH;
  const long long total = (long long)B * T * 2 * (GH / 4);
  const long long blocks = (total + 255) / 256;
  gather_gi0_kernel<<<unsigned(blocks), 256, 0, st>>>(m->w.tab0, x, T_in, gi, B, T, m->V, GH, m->tok_flag_dev);
  ASB_CHECK_LAUNCH("gather_gi0_kernel");
  return ASB_OK;
}

int launch_recurrent_layer(const asb_model* m, int layer, const float* gi, const int* lengths, float* out, int B,
                           int T, cudaStream_t st) {
  const PackedWeights& w = m->w;
  if (m->G == kLG && m->precision == ASB_PRECISION_FP32) {
    static DeviceOnce once;
    const int smem = (kH * kGH + 2 * 8 * kH) * int(sizeof(float));
    if (int rc = once.run([&] {
          return check_cuda(cudaFuncSetAttribute(lstm_fp32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem),
                            "lstm_fp32_kernel smem attribute");
        }))
      return rc;
    dim3 grid((B + 7) / 8, 2);
    lstm_fp32_kernel<<<grid, 256, smem, st>>>(w.whh_t[layer][0], w.whh_t[layer][1], gi, lengths, out, B, T);
    ASB_CHECK_LAUNCH("lstm_fp32_kernel");
    return ASB_OK;
  }
  if (m->G != kG)
    return set_error(ASB_ERR_UNSUPPORTED, "the mma.sync recurrent kernel implements the GRU cell only");
  if (m->precision == ASB_PRECISION_FP32) {
    static DeviceOnce once;
    const int smem = (kH * kGH + 2 * 8 * kH) * int(sizeof(float));
    if (int rc = once.run([&] {
          cudaError_t e = cudaFuncSetAttribute(gru_fp32_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
          if (e == cudaSuccess) e = cudaFuncSetAttribute(gru_fp32_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
          return check_cuda(e, "gru_fp32_kernel smem attribute");
        }))
      return rc;
    if (2 * ((B + 3) / 4) <= sm_count()) {
      gru_fp32_kernel<2><<<dim3((B + 3) / 4, 2), 256, smem, st>>>(w.whh_t[layer][0], w.whh_t[layer][1], w.bhh_n[layer][0],
                                                                 w.bhh_n[layer][1], gi, lengths, out, B, T);
    } else {
      gru_fp32_kernel<4><<<dim3((B + 7) / 8, 2), 256, smem, st>>>(w.whh_t[layer][0], w.whh_t[layer][1], w.bhh_n[layer][0],
                                                                 w.bhh_n[layer][1], gi, lengths, out, B, T);
    }
    ASB_CHECK_LAUNCH("gru_fp32_kernel");
    return ASB_OK;
  }
  // 8 rows per CTA while that still fits one wave of the machine, else 16
  const int sms = sm_count();
  if (2 * ((B + 7) / 8) <= 2 * sms) {
    dim3 grid((B + 7) / 8, 2);
    gru_fast_kernel<1><<<grid, 256, 0, st>>>(w.whh_frag[layer][0], w.whh_frag[layer][1], w.bhh_n[layer][0],
                                             w.bhh_n[layer][1], gi, lengths, out, B, T);
  } else {
    dim3 grid((B + 15) / 16, 2);
    gru_fast_kernel<2><<<grid, 256, 0, st>>>(w.whh_frag[layer][0], w.whh_frag[layer][1], w.bhh_n[layer][0],
                                             w.bhh_n[layer][1], gi, lengths, out, B, T);
  }
  ASB_CHECK_LAUNCH("gru_fast_kernel");
  return ASB_OK;
}

int pack_recurrent_weights(asb_model* m, const asb_weights* src, cudaStream_t st) {
  PackedWeights& w = m->w;
  for (int dir = 0; dir < 2; ++dir) {
    const int GH = m->G * kH;
    const int n = m->V * GH;
    build_tab0_kernel<<<(n + 255) / 256, 256, 0, st>>>(src->embedding, src->rnn_w_ih[0][dir], src->rnn_b_ih[0][dir],
                                                        src->rnn_b_hh[0][dir], w.tab0 + (size_t)dir * n, m->V, m->E,
                                                        GH, m->G == kG ? 2 * kH : GH);
    ASB_CHECK_LAUNCH("build_tab0_kernel");
  }
  if (m->G != kG) {
    // LSTM: b_hh is folded entirely; fast precision runs the tcgen05 kernel (gru_umma.cu), fp32 precision
    // lstm_fp32_kernel on W_hh^T
    for (int layer = 0; layer < 2; ++layer)
      for (int dir = 0; dir < 2; ++dir) {
        transpose_whh_kernel<<<(kH * kLGH + 255) / 256, 256, 0, st>>>(src->rnn_w_hh[layer][dir], w.whh_t[layer][dir], kLGH);
        ASB_CHECK_LAUNCH("transpose_whh_kernel");
      }
    return ASB_OK;
  }
  for (int layer = 0; layer < 2; ++layer)
    for (int dir = 0; dir < 2; ++dir) {
      const int nf = 8 * kG * 8 * 32 * 4;
      pack_whh_frag_kernel<<<(nf + 255) / 256, 256, 0, st>>>(src->rnn_w_hh[layer][dir], w.whh_frag[layer][dir]);
      ASB_CHECK_LAUNCH("pack_whh_frag_kernel");
      transpose_whh_kernel<<<(kH * kGH + 255) / 256, 256, 0, st>>>(src->rnn_w_hh[layer][dir], w.whh_t[layer][dir], kGH);
      ASB_CHECK_LAUNCH("transpose_whh_kernel");
      copy_slice_kernel<<<1, kH, 0, st>>>(src->rnn_b_hh[layer][dir] + 2 * kH, w.bhh_n[layer][dir], kH);
      ASB_CHECK_LAUNCH("copy_slice_kernel");
    }
  return ASB_OK;
}

}  // namespace asb
