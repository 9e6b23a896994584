// Dense layers of the model: layer-1 input projection and the fused contour decoder.
//
// Replaces, for a tile of 64 frames, the whole tail of ArtSpeech.forward
// (phoneme_to_articulation/encoder_decoder/models.py:140-145): linear.0 + ReLU, then for every
// articulator LayerNorm -> Linear(128,256) -> ReLU -> LayerNorm -> Linear(256,256) -> ReLU ->
// LayerNorm -> x/y heads (models.py:7-33), the two torch.stack copies and the sigmoid.  The
// reference runs ~9 library kernels per articulator and round-trips every intermediate through
// HBM; here activations never leave shared memory / registers and the contours are written once,
// directly in the final (frames, N_art, 2, D) layout.
//
// This file is the mma.sync (m16n8k8, tf32 in / fp32 accumulate) implementation: weights are
// streamed from L2 through a 2-stage cp.async ring as ONE continuous chunk sequence across all
// 31 GEMMs of a tile, so the pipeline never drains between layers.  fp32 mode runs the same code
// with 3xTF32 error compensation (a_lo*b_hi + a_hi*b_lo + a_hi*b_hi).
#include <new>

#include "model.cuh"

namespace {

using namespace asb;

constexpr int kBM = 64;            // frames per CTA tile
constexpr int kBK = 32;            // k extent of one streamed weight chunk
constexpr int kWStride = kBK + 8;  // 40: conflict-free 64-bit fragment loads (stride = 8 mod 32)
constexpr int kWStage = 256 * kWStride;          // floats per weight stage (up to 256 rows)
constexpr int kXStride = kH + 8;                 // 136
constexpr int kHStride = kFc + 8;                // 264
constexpr int kThreads = 256;
constexpr size_t kDecoderSmem = (size_t(kBM) * kXStride + size_t(kBM) * kHStride + 2 * size_t(kWStage)) * sizeof(float);

__device__ __forceinline__ uint32_t f2tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ float round_tf32(float x) { return __uint_as_float(f2tf32(x)); }

__device__ __forceinline__ void mma_tf32(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void cp_async16(void* dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst)), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

// One streamed chunk: `rows` weight rows x kBK floats starting at src (row stride ldw floats).
struct Chunk {
  const float* src;
  int rows;
  int ldw;
};

__device__ __forceinline__ void issue_chunk(const Chunk& ch, float* stage, int tid) {
  // rows * 8 sixteen-byte copies
  for (int i = tid; i < ch.rows * (kBK / 4); i += kThreads) {
    const int r = i >> 3, q = i & 7;
    cp_async16(stage + r * kWStride + 4 * q, ch.src + (size_t)r * ch.ldw + 4 * q);
  }
  cp_async_commit();
}

// acc += A[64][K] (smem, stride lda) . W^T for the chunk sequence [q0, q0 + K/kBK) of `next`.
// NTW = 8-column tiles per warp (warp grid 2 x 4: rows 32*wm, columns wn*8*NTW).
// The k index inside every 8-wide step is permuted (logical c <-> physical 2c, c+4 <-> 2c+1)
// identically for A and W, which turns all fragment loads into 64-bit loads.
template <int NTW, bool X3, typename NextChunk>
__device__ __forceinline__ void cta_gemm(float (&acc)[2][NTW][4], const float* As, int lda, int K, int& q,
                                         int q_end, NextChunk next, float* Ws, int tid) {
  const int w = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int wm = w >> 2, wn = w & 3;
  for (int k0 = 0; k0 < K; k0 += kBK, ++q) {
    cp_async_wait_all();
    __syncthreads();  // chunk q landed for everyone; everyone is done with chunk q-1's stage
    if (q + 1 < q_end) issue_chunk(next(q + 1), Ws + ((q + 1) & 1) * kWStage, tid);
    const float* Wc = Ws + (q & 1) * kWStage;
#pragma unroll
    for (int ks = 0; ks < kBK / 8; ++ks) {
      float2 af[2][2];
#pragma unroll
      for (int mt = 0; mt < 2; ++mt) {
        const float* ap = As + (32 * wm + 16 * mt + g) * lda + k0 + ks * 8 + 2 * c;
        af[mt][0] = *reinterpret_cast<const float2*>(ap);
        af[mt][1] = *reinterpret_cast<const float2*>(ap + 8 * lda);
      }
#pragma unroll
      for (int nt = 0; nt < NTW; ++nt) {
        const float2 bf = *reinterpret_cast<const float2*>(Wc + (wn * 8 * NTW + nt * 8 + g) * kWStride + ks * 8 + 2 * c);
        if (!X3) {
          const uint32_t b0 = __float_as_uint(bf.x), b1 = __float_as_uint(bf.y);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
            mma_tf32(acc[mt][nt], __float_as_uint(af[mt][0].x), __float_as_uint(af[mt][1].x),
                     __float_as_uint(af[mt][0].y), __float_as_uint(af[mt][1].y), b0, b1);
        } else {
          const uint32_t b0h = f2tf32(bf.x), b1h = f2tf32(bf.y);
          const uint32_t b0l = f2tf32(bf.x - __uint_as_float(b0h)), b1l = f2tf32(bf.y - __uint_as_float(b1h));
#pragma unroll
          for (int mt = 0; mt < 2; ++mt) {
            const float av[4] = {af[mt][0].x, af[mt][1].x, af[mt][0].y, af[mt][1].y};
            uint32_t ah[4], al[4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
              ah[r] = f2tf32(av[r]);
              al[r] = f2tf32(av[r] - __uint_as_float(ah[r]));
            }
            mma_tf32(acc[mt][nt], al[0], al[1], al[2], al[3], b0h, b1h);
            mma_tf32(acc[mt][nt], ah[0], ah[1], ah[2], ah[3], b0l, b1l);
            mma_tf32(acc[mt][nt], ah[0], ah[1], ah[2], ah[3], b0h, b1h);
          }
        }
      }
    }
  }
}

template <int NTW>
__device__ __forceinline__ void zero_acc(float (&acc)[2][NTW][4]) {
#pragma unroll
  for (int mt = 0; mt < 2; ++mt)
#pragma unroll
    for (int nt = 0; nt < NTW; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[mt][nt][e] = 0.f;
}

// acc + bias -> ReLU -> dst[64][ldd] (fp32, not rounded: LayerNorm follows or caller rounds)
template <int NTW, bool ROUND>
__device__ __forceinline__ void store_bias_relu(const float (&acc)[2][NTW][4], const float* __restrict__ bias,
                                                float* dst, int ldd, int tid) {
  const int w = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int wm = w >> 2, wn = w & 3;
#pragma unroll
  for (int nt = 0; nt < NTW; ++nt) {
    const int col = wn * 8 * NTW + nt * 8 + 2 * c;
    const float2 bb = __ldg(reinterpret_cast<const float2*>(bias + col));
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int row = 32 * wm + 16 * mt + g + 8 * hh;
        float2 v;
        v.x = fmaxf(acc[mt][nt][2 * hh] + bb.x, 0.f);
        v.y = fmaxf(acc[mt][nt][2 * hh + 1] + bb.y, 0.f);
        if (ROUND) {
          v.x = round_tf32(v.x);
          v.y = round_tf32(v.y);
        }
        *reinterpret_cast<float2*>(dst + row * ldd + col) = v;
      }
  }
}

// Row-wise LayerNorm of src[64][lds] (C columns) into dst[64][ldd]; biased variance, eps 1e-5
// (nn.LayerNorm defaults, models.py:11,14,17).  Each warp normalises 8 rows.
template <int C, bool ROUND>
__device__ __forceinline__ void layer_norm_rows(const float* src, int lds, float* dst, int ldd,
                                                const float* __restrict__ gamma, const float* __restrict__ beta,
                                                int tid) {
  const int w = tid >> 5, lane = tid & 31;
  constexpr int PER = C / 32;
  float gm[PER], bt[PER];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    gm[i] = __ldg(gamma + lane + 32 * i);
    bt[i] = __ldg(beta + lane + 32 * i);
  }
#pragma unroll 2
  for (int r = 0; r < 8; ++r) {
    const int row = 8 * w + r;
    float x[PER];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      x[i] = src[row * lds + lane + 32 * i];
      s += x[i];
    }
    const float mean = warp_sum(s) * (1.f / C);
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      x[i] -= mean;
      v = fmaf(x[i], x[i], v);
    }
    const float rstd = rsqrtf(warp_sum(v) * (1.f / C) + 1e-5f);
#pragma unroll
    for (int i = 0; i < PER; ++i) {
      float y = fmaf(x[i] * rstd, gm[i], bt[i]);
      if (ROUND) y = round_tf32(y);
      dst[row * ldd + lane + 32 * i] = y;
    }
  }
}

// Load a [64][K] fp32 tile (row stride K in global memory) into smem (stride lds); rows >= M are zero.
template <int K, bool ROUND>
__device__ __forceinline__ void load_rows(const float* __restrict__ src, long long row0, long long M, float* dst,
                                          int lds, int tid) {
  for (int i = tid; i < kBM * (K / 4); i += kThreads) {
    const int r = i / (K / 4), q = i - r * (K / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < M) v = __ldg(reinterpret_cast<const float4*>(src + (row0 + r) * K) + q);
    if (ROUND) {
      v.x = round_tf32(v.x); v.y = round_tf32(v.y); v.z = round_tf32(v.z); v.w = round_tf32(v.w);
    }
    *reinterpret_cast<float4*>(dst + r * lds + 4 * q) = v;
  }
}

struct DecoderArgs {
  const float* rnn_out;  // (M, 2H)
  float* out;            // (M, n_art, 2, D)
  long long M;
  int n_art, D;
  PackedWeights w;
};

template <bool X3>
__global__ void __launch_bounds__(kThreads, 1) decoder_kernel(const DecoderArgs a) {
  extern __shared__ __align__(16) float smem[];
  float* Xs = smem;                       // [64][136]  linear.0 output, kept for all articulators
  float* Hs = Xs + kBM * kXStride;        // [64][264]  current activation (A operand)
  float* Ws = Hs + kBM * kHStride;        // 2 weight stages
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * kBM;
  const PackedWeights& w = a.w;
  const int q_end = 8 + 20 * a.n_art;

  auto next = [&](int q) -> Chunk {
    if (q < 8) return Chunk{w.lin_w + q * kBK, kH, 2 * kH};
    int r = q - 8;
    const int art = r / 20;
    r -= art * 20;
    if (r < 4) return Chunk{w.fc1_w + (size_t)art * kFc * kH + r * kBK, kFc, kH};
    if (r < 12) return Chunk{w.fc2_w + (size_t)art * kFc * kFc + (r - 4) * kBK, kFc, kFc};
    return Chunk{w.head_w + (size_t)art * kHeadPad * kFc + (r - 12) * kBK, kHeadPad, kFc};
  };

  issue_chunk(next(0), Ws, tid);
  load_rows<2 * kH, !X3>(a.rnn_out, row0, a.M, Hs, kHStride, tid);
  int q = 0;
  {  // linear.0 + ReLU  (models.py:113-116,140)
    float acc[2][4][4];
    zero_acc<4>(acc);
    cta_gemm<4, X3>(acc, Hs, kHStride, 2 * kH, q, q_end, next, Ws, tid);
    store_bias_relu<4, false>(acc, w.lin_b, Xs, kXStride, tid);
  }
  __syncthreads();

  const int w_id = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int wm = w_id >> 2, wn = w_id & 3;
  const int n_out = 2 * a.D;

  for (int art = 0; art < a.n_art; ++art) {
    // LayerNorm(128) of the shared input with this articulator's affine   (models.py:11)
    layer_norm_rows<kH, !X3>(Xs, kXStride, Hs, kHStride, w.ln0_w + art * kH, w.ln0_b + art * kH, tid);
    {  // Linear(128,256) + ReLU + LayerNorm(256)   (models.py:12-14)
      float acc[2][8][4];
      zero_acc<8>(acc);
      cta_gemm<8, X3>(acc, Hs, kHStride, kH, q, q_end, next, Ws, tid);
      __syncthreads();  // all warps finished reading Hs before it is overwritten
      store_bias_relu<8, false>(acc, w.fc1_b + art * kFc, Hs, kHStride, tid);
    }
    __syncthreads();
    layer_norm_rows<kFc, !X3>(Hs, kHStride, Hs, kHStride, w.ln1_w + art * kFc, w.ln1_b + art * kFc, tid);
    {  // Linear(256,256) + ReLU + LayerNorm(256)   (models.py:15-17)
      float acc[2][8][4];
      zero_acc<8>(acc);
      cta_gemm<8, X3>(acc, Hs, kHStride, kFc, q, q_end, next, Ws, tid);
      __syncthreads();
      store_bias_relu<8, false>(acc, w.fc2_b + art * kFc, Hs, kHStride, tid);
    }
    __syncthreads();
    layer_norm_rows<kFc, !X3>(Hs, kHStride, Hs, kHStride, w.ln2_w + art * kFc, w.ln2_b + art * kFc, tid);
    {  // x/y heads + sigmoid, written in the final layout   (models.py:20-21,29-31,141-145)
      float acc[2][4][4];
      zero_acc<4>(acc);
      cta_gemm<4, X3>(acc, Hs, kHStride, kFc, q, q_end, next, Ws, tid);
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const int col = wn * 32 + nt * 8 + 2 * c;
        if (col < n_out) {
          const float2 bb = __ldg(reinterpret_cast<const float2*>(w.head_b + art * kHeadPad + col));
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const long long row = row0 + 32 * wm + 16 * mt + g + 8 * hh;
              if (row < a.M) {
                const float zx = acc[mt][nt][2 * hh] + bb.x, zy = acc[mt][nt][2 * hh + 1] + bb.y;
                float2 v;
                if (X3) {
                  v.x = 1.f / (1.f + expf(-zx));
                  v.y = 1.f / (1.f + expf(-zy));
                } else {
                  v.x = __fdividef(1.f, 1.f + __expf(-zx));
                  v.y = __fdividef(1.f, 1.f + __expf(-zy));
                }
                *reinterpret_cast<float2*>(a.out + (row * a.n_art + art) * n_out + col) = v;
              }
            }
        }
      }
    }
    __syncthreads();  // heads finished reading Hs before the next articulator's LayerNorm writes it
  }
}

// C[M][N] = A[M][256] . W[N][256]^T + bias   (layer-1 input projection, N = 768)
struct LinearArgs {
  const float* A;
  const float* W;
  const float* bias;
  float* C;
  long long M;
  int N;
};

template <bool X3>
__global__ void __launch_bounds__(kThreads, 1) linear256_kernel(const LinearArgs a) {
  extern __shared__ __align__(16) float smem[];
  float* Hs = smem;
  float* Ws = Hs + kBM * kHStride;
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * kBM;
  const int n0 = blockIdx.y * 256;
  auto next = [&](int q) -> Chunk { return Chunk{a.W + (size_t)n0 * 256 + q * kBK, 256, 256}; };
  issue_chunk(next(0), Ws, tid);
  load_rows<256, !X3>(a.A, row0, a.M, Hs, kHStride, tid);
  int q = 0;
  float acc[2][8][4];
  zero_acc<8>(acc);
  cta_gemm<8, X3>(acc, Hs, kHStride, 256, q, 8, next, Ws, tid);
  const int w_id = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
  const int wm = w_id >> 2, wn = w_id & 3;
#pragma unroll
  for (int nt = 0; nt < 8; ++nt) {
    const int col = n0 + wn * 64 + nt * 8 + 2 * c;
    const float2 bb = __ldg(reinterpret_cast<const float2*>(a.bias + col));
#pragma unroll
    for (int mt = 0; mt < 2; ++mt)
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const long long row = row0 + 32 * wm + 16 * mt + g + 8 * hh;
        if (row < a.M)
          *reinterpret_cast<float2*>(a.C + row * a.N + col) =
              make_float2(acc[mt][nt][2 * hh] + bb.x, acc[mt][nt][2 * hh + 1] + bb.y);
      }
  }
}

// ---------------------------------------------------------------- principal-components head
// PrincipalComponentsArtSpeech's tail (phoneme_to_articulation/principal_components/models/rnn.py:
// 11-33, 70-79, 106-108): linear (2H -> H) + ReLU, then PrincipalComponentsPredictor = LayerNorm(H) ->
// Linear(H,256) -> ReLU -> LayerNorm(256) -> Linear(256,128) -> ReLU -> LayerNorm(128) ->
// Linear(128, n_components), then tanh.  Same building blocks as the contour decoder: one 64-frame
// tile per CTA, activations in shared memory, weights streamed as one chunk sequence of 24 chunks.
constexpr int kPcMid = 128;    // hidden_features / 2
constexpr int kPcOutPad = 128; // n_components rows padded to one 128-row weight stage

struct PcWeights {
  float* lin_w; float* lin_b;    // [H][2H], [H]
  float* ln0_w; float* ln0_b;    // [H]
  float* fc1_w; float* fc1_b;    // [256][H], [256]
  float* ln1_w; float* ln1_b;    // [256]
  float* fc2_w; float* fc2_b;    // [128][256], [128]
  float* ln2_w; float* ln2_b;    // [128]
  float* fc3_w; float* fc3_b;    // [128 (n_components, zero padded)][128], [128]
};

struct PcArgs {
  const float* rnn_out;  // (M, 2H)
  float* out;            // (M, n_components)
  long long M;
  int n_components;
  PcWeights w;
};

template <bool X3>
__global__ void __launch_bounds__(kThreads, 1) pc_head_kernel(const PcArgs a) {
  extern __shared__ __align__(16) float smem[];
  float* Xs = smem;                       // [64][136]
  float* Hs = Xs + kBM * kXStride;        // [64][264]
  float* Ws = Hs + kBM * kHStride;        // 2 weight stages
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * kBM;
  const PcWeights& w = a.w;
  constexpr int q_end = 8 + 4 + 8 + 4;

  auto next = [&](int q) -> Chunk {
    if (q < 8) return Chunk{w.lin_w + q * kBK, kH, 2 * kH};
    if (q < 12) return Chunk{w.fc1_w + (q - 8) * kBK, kFc, kH};
    if (q < 20) return Chunk{w.fc2_w + (q - 12) * kBK, kPcMid, kFc};
    return Chunk{w.fc3_w + (q - 20) * kBK, kPcOutPad, kPcMid};
  };

  issue_chunk(next(0), Ws, tid);
  load_rows<2 * kH, !X3>(a.rnn_out, row0, a.M, Hs, kHStride, tid);
  int q = 0;
  {  // linear + ReLU   (rnn.py:70-73,106)
    float acc[2][4][4];
    zero_acc<4>(acc);
    cta_gemm<4, X3>(acc, Hs, kHStride, 2 * kH, q, q_end, next, Ws, tid);
    store_bias_relu<4, false>(acc, w.lin_b, Xs, kXStride, tid);
  }
  __syncthreads();
  layer_norm_rows<kH, !X3>(Xs, kXStride, Hs, kHStride, w.ln0_w, w.ln0_b, tid);       // rnn.py:21
  {  // Linear(H,256) + ReLU   (:22-23)
    float acc[2][8][4];
    zero_acc<8>(acc);
    cta_gemm<8, X3>(acc, Hs, kHStride, kH, q, q_end, next, Ws, tid);
    __syncthreads();
    store_bias_relu<8, false>(acc, w.fc1_b, Hs, kHStride, tid);
  }
  __syncthreads();
  layer_norm_rows<kFc, !X3>(Hs, kHStride, Hs, kHStride, w.ln1_w, w.ln1_b, tid);      // :24
  {  // Linear(256,128) + ReLU   (:25-26); written to Xs, Hs is the operand being read
    float acc[2][4][4];
    zero_acc<4>(acc);
    cta_gemm<4, X3>(acc, Hs, kHStride, kFc, q, q_end, next, Ws, tid);
    store_bias_relu<4, false>(acc, w.fc2_b, Xs, kXStride, tid);
  }
  __syncthreads();
  layer_norm_rows<kPcMid, !X3>(Xs, kXStride, Hs, kHStride, w.ln2_w, w.ln2_b, tid);   // :27
  {  // Linear(128, n_components) + tanh   (:28, :107)
    float acc[2][4][4];
    zero_acc<4>(acc);
    cta_gemm<4, X3>(acc, Hs, kHStride, kPcMid, q, q_end, next, Ws, tid);
    const int w_id = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
    const int wm = w_id >> 2, wn = w_id & 3;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int col = wn * 32 + nt * 8 + 2 * c + e;
        if (col < a.n_components) {
          const float bb = __ldg(w.fc3_b + col);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const long long row = row0 + 32 * wm + 16 * mt + g + 8 * hh;
              if (row < a.M) {
                const float z = acc[mt][nt][2 * hh + e] + bb;
                // tanh(z) = 1 - 2 / (1 + e^{2z}); fp32 mode uses the library tanhf
                a.out[row * a.n_components + col] =
                    X3 ? tanhf(z) : 1.f - __fdividef(2.f, 1.f + __expf(2.f * z));
              }
            }
        }
      }
    }
  }
}


// ---------------------------------------------------------------- standalone ArticulatorPredictor
// ArticulatorPredictor.forward (phoneme_to_articulation/encoder_decoder/models.py:23-33) on its own:
// the class is also instantiated outside ArtSpeech (transformer heads, transformer/models.py:337-339,
// with in_features = embed_dim).  LayerNorm(C) -> Linear(C,256) -> ReLU -> LayerNorm(256) ->
// Linear(256,256) -> ReLU -> LayerNorm(256) -> x/y heads, stacked to (rows, 2, D); NO sigmoid (the
// caller applies its own).  Same building blocks as decoder_kernel; C = in_features is a run-time
// multiple of 32 up to 256.
struct PredWeights {
  float* ln0_w; float* ln0_b;    // [C]
  float* fc1_w; float* fc1_b;    // [256][C], [256]
  float* ln1_w; float* ln1_b;    // [256]
  float* fc2_w; float* fc2_b;    // [256][256], [256]
  float* ln2_w; float* ln2_b;    // [256]
  float* head_w; float* head_b;  // [128 (2D rows, zero padded)][256], [128]
};

struct PredArgs {
  const float* features;  // (M, C)
  float* out;             // (M, 2, D)
  long long M;
  int C, D;
  PredWeights w;
};

// LayerNorm over C = 32 * per columns (run-time), each warp normalises 8 rows
template <bool ROUND>
__device__ __forceinline__ void layer_norm_rows_rt(const float* src, int lds, float* dst, int ldd, int C,
                                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                                   int tid) {
  const int w = tid >> 5, lane = tid & 31;
  const int per = C >> 5;
  const float inv = 1.f / float(C);
  for (int r = 0; r < 8; ++r) {
    const int row = 8 * w + r;
    float x[8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      x[i] = i < per ? src[row * lds + lane + 32 * i] : 0.f;
      s += x[i];
    }
    const float mean = warp_sum(s) * inv;
    float v = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      x[i] = i < per ? x[i] - mean : 0.f;
      v = fmaf(x[i], x[i], v);
    }
    const float rstd = rsqrtf(warp_sum(v) * inv + 1e-5f);
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (i < per) {
        float y = fmaf(x[i] * rstd, __ldg(gamma + lane + 32 * i), __ldg(beta + lane + 32 * i));
        if (ROUND) y = round_tf32(y);
        dst[row * ldd + lane + 32 * i] = y;
      }
  }
}

template <bool X3>
__global__ void __launch_bounds__(kThreads, 1) predictor_kernel(const PredArgs a) {
  extern __shared__ __align__(16) float smem[];
  float* Hs = smem + kBM * kXStride;      // [64][264] (same carve-up as decoder_kernel; Xs unused)
  float* Ws = Hs + kBM * kHStride;
  const int tid = threadIdx.x;
  const long long row0 = (long long)blockIdx.x * kBM;
  const PredWeights& w = a.w;
  const int C = a.C, q1 = C / kBK;
  const int q_end = q1 + 8 + 8;

  auto next = [&](int q) -> Chunk {
    if (q < q1) return Chunk{w.fc1_w + q * kBK, kFc, C};
    if (q < q1 + 8) return Chunk{w.fc2_w + (q - q1) * kBK, kFc, kFc};
    return Chunk{w.head_w + (q - q1 - 8) * kBK, kHeadPad, kFc};
  };

  issue_chunk(next(0), Ws, tid);
  for (int i = tid; i < kBM * (C / 4); i += kThreads) {   // raw fp32 rows: LayerNorm comes first
    const int r = i / (C / 4), qq = i - r * (C / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row0 + r < a.M) v = __ldg(reinterpret_cast<const float4*>(a.features + (row0 + r) * C) + qq);
    *reinterpret_cast<float4*>(Hs + r * kHStride + 4 * qq) = v;
  }
  __syncthreads();
  layer_norm_rows_rt<!X3>(Hs, kHStride, Hs, kHStride, C, w.ln0_w, w.ln0_b, tid);      // models.py:11
  int q = 0;
  {  // Linear(C,256) + ReLU   (:12-13)
    float acc[2][8][4];
    zero_acc<8>(acc);
    cta_gemm<8, X3>(acc, Hs, kHStride, C, q, q_end, next, Ws, tid);
    __syncthreads();
    store_bias_relu<8, false>(acc, w.fc1_b, Hs, kHStride, tid);
  }
  __syncthreads();
  layer_norm_rows<kFc, !X3>(Hs, kHStride, Hs, kHStride, w.ln1_w, w.ln1_b, tid);        // :14
  {  // Linear(256,256) + ReLU   (:15-16)
    float acc[2][8][4];
    zero_acc<8>(acc);
    cta_gemm<8, X3>(acc, Hs, kHStride, kFc, q, q_end, next, Ws, tid);
    __syncthreads();
    store_bias_relu<8, false>(acc, w.fc2_b, Hs, kHStride, tid);
  }
  __syncthreads();
  layer_norm_rows<kFc, !X3>(Hs, kHStride, Hs, kHStride, w.ln2_w, w.ln2_b, tid);        // :17
  {  // x_coords / y_coords, stacked on dim 2   (:20-21,29-31)
    float acc[2][4][4];
    zero_acc<4>(acc);
    cta_gemm<4, X3>(acc, Hs, kHStride, kFc, q, q_end, next, Ws, tid);
    const int w_id = tid >> 5, lane = tid & 31, g = lane >> 2, c = lane & 3;
    const int wm = w_id >> 2, wn = w_id & 3;
    const int n_out = 2 * a.D;
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int col = wn * 32 + nt * 8 + 2 * c + e;
        if (col < n_out) {
          const float bb = __ldg(w.head_b + col);
#pragma unroll
          for (int mt = 0; mt < 2; ++mt)
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              const long long row = row0 + 32 * wm + 16 * mt + g + 8 * hh;
              if (row < a.M) a.out[row * n_out + col] = acc[mt][nt][2 * hh + e] + bb;
            }
        }
      }
  }
}

// ---------------------------------------------------------------- weight packing
// dst[n][k] = src[n][k] (rows >= rows_src are zero), optionally rounded to tf32
__global__ void pack_matrix_kernel(const float* __restrict__ src, float* __restrict__ dst, int rows_src,
                                   int rows_dst, int K, int round) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows_dst * K) return;
  const int n = i / K;
  float v = n < rows_src ? src[i] : 0.f;
  if (round) v = round_tf32(v);
  dst[i] = v;
}
__global__ void pack_vector_kernel(const float* __restrict__ src, const float* __restrict__ add, int n_add,
                                   float* __restrict__ dst, int n_src, int n_dst) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_dst) return;
  float v = i < n_src ? src[i] : 0.f;
  if (add && i < n_add) v += add[i];
  dst[i] = v;
}

}  // namespace

namespace asb {

static int pack_matrix(const float* src, float* dst, int rows_src, int rows_dst, int K, bool round, cudaStream_t st) {
  const int n = rows_dst * K;
  pack_matrix_kernel<<<(n + 255) / 256, 256, 0, st>>>(src, dst, rows_src, rows_dst, K, round ? 1 : 0);
  ASB_CHECK_LAUNCH("pack_matrix_kernel");
  return ASB_OK;
}
static int pack_vector(const float* src, const float* add, int n_add, float* dst, int n_src, int n_dst,
                       cudaStream_t st) {
  pack_vector_kernel<<<(n_dst + 255) / 256, 256, 0, st>>>(src, add, n_add, dst, n_src, n_dst);
  ASB_CHECK_LAUNCH("pack_vector_kernel");
  return ASB_OK;
}

#define ASB_TRY(x)        \
  do {                    \
    int rc__ = (x);       \
    if (rc__) return rc__; \
  } while (0)

int pack_dense_weights(asb_model* m, const asb_weights* s, cudaStream_t st) {
  PackedWeights& w = m->w;
  const bool rnd = m->precision == ASB_PRECISION_FAST;
  const int GH = m->G * kH;
  for (int dir = 0; dir < 2; ++dir) {
    ASB_TRY(pack_matrix(s->rnn_w_ih[1][dir], w.wih1 + (size_t)dir * GH * 2 * kH, GH, GH, 2 * kH, rnd, st));
    // b_ih + the foldable part of b_hh (GRU: r and z rows; LSTM: everything)
    ASB_TRY(pack_vector(s->rnn_b_ih[1][dir], s->rnn_b_hh[1][dir], m->G == 3 ? 2 * kH : GH, w.bih1 + dir * GH, GH, GH, st));
  }
  ASB_TRY(pack_matrix(s->linear_w, w.lin_w, kH, kH, 2 * kH, rnd, st));
  ASB_TRY(pack_vector(s->linear_b, nullptr, 0, w.lin_b, kH, kH, st));
  const int D = m->D;
  for (int i = 0; i < m->N; ++i) {
    ASB_TRY(pack_vector(s->ln0_w[i], nullptr, 0, w.ln0_w + i * kH, kH, kH, st));
    ASB_TRY(pack_vector(s->ln0_b[i], nullptr, 0, w.ln0_b + i * kH, kH, kH, st));
    ASB_TRY(pack_matrix(s->fc1_w[i], w.fc1_w + (size_t)i * kFc * kH, kFc, kFc, kH, rnd, st));
    ASB_TRY(pack_vector(s->fc1_b[i], nullptr, 0, w.fc1_b + i * kFc, kFc, kFc, st));
    ASB_TRY(pack_vector(s->ln1_w[i], nullptr, 0, w.ln1_w + i * kFc, kFc, kFc, st));
    ASB_TRY(pack_vector(s->ln1_b[i], nullptr, 0, w.ln1_b + i * kFc, kFc, kFc, st));
    ASB_TRY(pack_matrix(s->fc2_w[i], w.fc2_w + (size_t)i * kFc * kFc, kFc, kFc, kFc, rnd, st));
    ASB_TRY(pack_vector(s->fc2_b[i], nullptr, 0, w.fc2_b + i * kFc, kFc, kFc, st));
    ASB_TRY(pack_vector(s->ln2_w[i], nullptr, 0, w.ln2_w + i * kFc, kFc, kFc, st));
    ASB_TRY(pack_vector(s->ln2_b[i], nullptr, 0, w.ln2_b + i * kFc, kFc, kFc, st));
    float* hw = w.head_w + (size_t)i * kHeadPad * kFc;
    ASB_TRY(pack_matrix(s->xh_w[i], hw, D, D, kFc, rnd, st));
    ASB_TRY(pack_matrix(s->yh_w[i], hw + (size_t)D * kFc, D, kHeadPad - D, kFc, rnd, st));
    ASB_TRY(pack_vector(s->xh_b[i], nullptr, 0, w.head_b + i * kHeadPad, D, D, st));
    ASB_TRY(pack_vector(s->yh_b[i], nullptr, 0, w.head_b + i * kHeadPad + D, D, kHeadPad - D, st));
  }
  return ASB_OK;
}

int launch_linear(const asb_model* m, const float* A, const float* W, const float* bias, float* C, int M, int N,
                  int K, cudaStream_t st) {
  if (K != 256 || N % 256 != 0) return set_error(ASB_ERR_UNSUPPORTED, "launch_linear: K must be 256 and N % 256 == 0");
  const size_t smem = (size_t(kBM) * kHStride + 2 * size_t(kWStage)) * sizeof(float);
  static DeviceOnce once;
  if (int rc = once.run([&] {
        cudaError_t e = cudaFuncSetAttribute(linear256_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(linear256_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
        return check_cuda(e, "linear256_kernel smem attribute");
      }))
    return rc;
  LinearArgs a{A, W, bias, C, M, N};
  dim3 grid((M + kBM - 1) / kBM, N / 256);
  if (m->precision == ASB_PRECISION_FP32)
    linear256_kernel<true><<<grid, kThreads, smem, st>>>(a);
  else
    linear256_kernel<false><<<grid, kThreads, smem, st>>>(a);
  ASB_CHECK_LAUNCH("linear256_kernel");
  return ASB_OK;
}

int launch_decoder(const asb_model* m, const float* rnn_out, float* out, int M, cudaStream_t st) {
  static DeviceOnce once;
  if (int rc = once.run([&] {
        cudaError_t e = cudaFuncSetAttribute(decoder_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kDecoderSmem));
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(decoder_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kDecoderSmem));
        return check_cuda(e, "decoder_kernel smem attribute");
      }))
    return rc;
  DecoderArgs a{rnn_out, out, M, m->N, m->D, m->w};
  const unsigned grid = unsigned((M + kBM - 1) / kBM);
  if (m->precision == ASB_PRECISION_FP32)
    decoder_kernel<true><<<grid, kThreads, kDecoderSmem, st>>>(a);
  else
    decoder_kernel<false><<<grid, kThreads, kDecoderSmem, st>>>(a);
  ASB_CHECK_LAUNCH("decoder_kernel");
  return ASB_OK;
}

}  // namespace asb

// ---- principal-components head: C ABI -------------------------------------------------------------
struct asb_pc_head {
  int n_components, precision;
  void* arena;
  PcWeights w;
};

extern "C" int asb_pc_head_create(const asb_pc_weights* src, int hidden, int n_components, int precision, void* stream,
                                  asb_pc_head** out) {
  using namespace asb;
  if (!src || !out) return set_error(ASB_ERR_INVALID, "asb_pc_head_create: null argument");
  *out = nullptr;
  if (hidden != kH) return set_error(ASB_ERR_UNSUPPORTED, "asb_pc_head_create: hidden_size must be %d (got %d)", kH, hidden);
  if (n_components < 1 || n_components > kPcOutPad)
    return set_error(ASB_ERR_INVALID, "asb_pc_head_create: need 1 <= n_components <= %d (got %d)", kPcOutPad, n_components);
  if (precision != ASB_PRECISION_FAST && precision != ASB_PRECISION_FP32)
    return set_error(ASB_ERR_INVALID, "asb_pc_head_create: bad precision %d", precision);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const size_t counts[14] = {size_t(kH) * 2 * kH, kH, kH, kH, size_t(kFc) * kH, kFc, kFc, kFc,
                             size_t(kPcMid) * kFc, kPcMid, kPcMid, kPcMid, size_t(kPcOutPad) * kPcMid, kPcOutPad};
  size_t off[15];
  off[0] = 0;
  for (int i = 0; i < 14; ++i) off[i + 1] = off[i] + (counts[i] + 63) / 64 * 64;
  asb_pc_head* h = new (std::nothrow) asb_pc_head();
  if (!h) return set_error(ASB_ERR_INVALID, "asb_pc_head_create: out of host memory");
  cudaError_t e = cudaMalloc(&h->arena, off[14] * sizeof(float));
  if (e != cudaSuccess) {
    delete h;
    return check_cuda(e, "asb_pc_head_create: cudaMalloc");
  }
  float* b = static_cast<float*>(h->arena);
  PcWeights& w = h->w;
  float** slots[14] = {&w.lin_w, &w.lin_b, &w.ln0_w, &w.ln0_b, &w.fc1_w, &w.fc1_b, &w.ln1_w, &w.ln1_b,
                       &w.fc2_w, &w.fc2_b, &w.ln2_w, &w.ln2_b, &w.fc3_w, &w.fc3_b};
  for (int i = 0; i < 14; ++i) *slots[i] = b + off[i];
  h->n_components = n_components;
  h->precision = precision;
  const bool rnd = precision == ASB_PRECISION_FAST;
  int rc = pack_matrix(src->lin_w, w.lin_w, kH, kH, 2 * kH, rnd, st);
  if (!rc) rc = pack_vector(src->lin_b, nullptr, 0, w.lin_b, kH, kH, st);
  if (!rc) rc = pack_vector(src->ln0_w, nullptr, 0, w.ln0_w, kH, kH, st);
  if (!rc) rc = pack_vector(src->ln0_b, nullptr, 0, w.ln0_b, kH, kH, st);
  if (!rc) rc = pack_matrix(src->fc1_w, w.fc1_w, kFc, kFc, kH, rnd, st);
  if (!rc) rc = pack_vector(src->fc1_b, nullptr, 0, w.fc1_b, kFc, kFc, st);
  if (!rc) rc = pack_vector(src->ln1_w, nullptr, 0, w.ln1_w, kFc, kFc, st);
  if (!rc) rc = pack_vector(src->ln1_b, nullptr, 0, w.ln1_b, kFc, kFc, st);
  if (!rc) rc = pack_matrix(src->fc2_w, w.fc2_w, kPcMid, kPcMid, kFc, rnd, st);
  if (!rc) rc = pack_vector(src->fc2_b, nullptr, 0, w.fc2_b, kPcMid, kPcMid, st);
  if (!rc) rc = pack_vector(src->ln2_w, nullptr, 0, w.ln2_w, kPcMid, kPcMid, st);
  if (!rc) rc = pack_vector(src->ln2_b, nullptr, 0, w.ln2_b, kPcMid, kPcMid, st);
  if (!rc) rc = pack_matrix(src->fc3_w, w.fc3_w, n_components, kPcOutPad, kPcMid, rnd, st);
  if (!rc) rc = pack_vector(src->fc3_b, nullptr, 0, w.fc3_b, n_components, kPcOutPad, st);
  if (!rc) rc = check_cuda(cudaStreamSynchronize(st), "asb_pc_head_create: packing kernels");
  if (rc) {
    cudaFree(h->arena);
    delete h;
    return rc;
  }
  *out = h;
  return ASB_OK;
}

extern "C" void asb_pc_head_destroy(asb_pc_head* h) {
  if (!h) return;
  cudaFree(h->arena);
  delete h;
}

extern "C" int asb_pc_head_forward(asb_pc_head* h, const float* features, int64_t rows, float* out, void* stream) {
  using namespace asb;
  if (!h) return set_error(ASB_ERR_INVALID, "asb_pc_head_forward: null head");
  if (rows < 0 || rows > 0x7fffffffLL) return set_error(ASB_ERR_INVALID, "asb_pc_head_forward: bad row count");
  if (rows == 0) return ASB_OK;
  if (!features || !out) return set_error(ASB_ERR_INVALID, "asb_pc_head_forward: null pointer");
  if (reinterpret_cast<uintptr_t>(features) & 15)
    return set_error(ASB_ERR_INVALID, "asb_pc_head_forward: features must be 16-byte aligned");
  static DeviceOnce once;
  if (int rc = once.run([&] {
        cudaError_t e = cudaFuncSetAttribute(pc_head_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kDecoderSmem));
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(pc_head_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kDecoderSmem));
        return check_cuda(e, "pc_head_kernel smem attribute");
      }))
    return rc;
  PcArgs a{features, out, rows, h->n_components, h->w};
  const unsigned grid = unsigned((rows + kBM - 1) / kBM);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (h->precision == ASB_PRECISION_FP32)
    pc_head_kernel<true><<<grid, kThreads, kDecoderSmem, st>>>(a);
  else
    pc_head_kernel<false><<<grid, kThreads, kDecoderSmem, st>>>(a);
  ASB_CHECK_LAUNCH("pc_head_kernel");
  return ASB_OK;
}

// ---- standalone ArticulatorPredictor: C ABI ---------------------------------------------------------
struct asb_predictor {
  int C, D, precision;
  void* arena;
  PredWeights w;
};

extern "C" int asb_predictor_create(const asb_predictor_weights* src, int in_features, int n_samples, int precision,
                                    void* stream, asb_predictor** out) {
  using namespace asb;
  if (!src || !out) return set_error(ASB_ERR_INVALID, "asb_predictor_create: null argument");
  *out = nullptr;
  if (in_features < 32 || in_features > 256 || (in_features & 31))
    return set_error(ASB_ERR_UNSUPPORTED, "asb_predictor_create: in_features must be a multiple of 32 in [32, 256] (got %d)",
                     in_features);
  if (n_samples < 1 || 2 * n_samples > kHeadPad)
    return set_error(ASB_ERR_INVALID, "asb_predictor_create: need 1 <= n_samples <= %d (got %d)", kHeadPad / 2, n_samples);
  if (precision != ASB_PRECISION_FAST && precision != ASB_PRECISION_FP32)
    return set_error(ASB_ERR_INVALID, "asb_predictor_create: bad precision %d", precision);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int C = in_features, D = n_samples;
  const size_t counts[12] = {size_t(C), size_t(C), size_t(kFc) * C, kFc, kFc, kFc, size_t(kFc) * kFc, kFc, kFc, kFc,
                             size_t(kHeadPad) * kFc, kHeadPad};
  size_t off[13];
  off[0] = 0;
  for (int i = 0; i < 12; ++i) off[i + 1] = off[i] + (counts[i] + 63) / 64 * 64;
  asb_predictor* h = new (std::nothrow) asb_predictor();
  if (!h) return set_error(ASB_ERR_INVALID, "asb_predictor_create: out of host memory");
  cudaError_t e = cudaMalloc(&h->arena, off[12] * sizeof(float));
  if (e != cudaSuccess) {
    delete h;
    return check_cuda(e, "asb_predictor_create: cudaMalloc");
  }
  float* b = static_cast<float*>(h->arena);
  PredWeights& w = h->w;
  float** slots[12] = {&w.ln0_w, &w.ln0_b, &w.fc1_w, &w.fc1_b, &w.ln1_w, &w.ln1_b, &w.fc2_w, &w.fc2_b,
                       &w.ln2_w, &w.ln2_b, &w.head_w, &w.head_b};
  for (int i = 0; i < 12; ++i) *slots[i] = b + off[i];
  h->C = C; h->D = D; h->precision = precision;
  const bool rnd = precision == ASB_PRECISION_FAST;
  int rc = pack_vector(src->ln0_w, nullptr, 0, w.ln0_w, C, C, st);
  if (!rc) rc = pack_vector(src->ln0_b, nullptr, 0, w.ln0_b, C, C, st);
  if (!rc) rc = pack_matrix(src->fc1_w, w.fc1_w, kFc, kFc, C, rnd, st);
  if (!rc) rc = pack_vector(src->fc1_b, nullptr, 0, w.fc1_b, kFc, kFc, st);
  if (!rc) rc = pack_vector(src->ln1_w, nullptr, 0, w.ln1_w, kFc, kFc, st);
  if (!rc) rc = pack_vector(src->ln1_b, nullptr, 0, w.ln1_b, kFc, kFc, st);
  if (!rc) rc = pack_matrix(src->fc2_w, w.fc2_w, kFc, kFc, kFc, rnd, st);
  if (!rc) rc = pack_vector(src->fc2_b, nullptr, 0, w.fc2_b, kFc, kFc, st);
  if (!rc) rc = pack_vector(src->ln2_w, nullptr, 0, w.ln2_w, kFc, kFc, st);
  if (!rc) rc = pack_vector(src->ln2_b, nullptr, 0, w.ln2_b, kFc, kFc, st);
  if (!rc) rc = pack_matrix(src->xh_w, w.head_w, D, D, kFc, rnd, st);
  if (!rc) rc = pack_matrix(src->yh_w, w.head_w + (size_t)D * kFc, D, kHeadPad - D, kFc, rnd, st);
  if (!rc) rc = pack_vector(src->xh_b, nullptr, 0, w.head_b, D, D, st);
  if (!rc) rc = pack_vector(src->yh_b, nullptr, 0, w.head_b + D, D, kHeadPad - D, st);
  if (!rc) rc = check_cuda(cudaStreamSynchronize(st), "asb_predictor_create: packing kernels");
  if (rc) {
    cudaFree(h->arena);
    delete h;
    return rc;
  }
  *out = h;
  return ASB_OK;
}

extern "C" void asb_predictor_destroy(asb_predictor* h) {
  if (!h) return;
  cudaFree(h->arena);
  delete h;
}

extern "C" int asb_predictor_forward(asb_predictor* h, const float* features, int64_t rows, float* out, void* stream) {
  using namespace asb;
  if (!h) return set_error(ASB_ERR_INVALID, "asb_predictor_forward: null predictor");
  if (rows < 0 || rows > 0x7fffffffLL) return set_error(ASB_ERR_INVALID, "asb_predictor_forward: bad row count");
  if (rows == 0) return ASB_OK;
  if (!features || !out) return set_error(ASB_ERR_INVALID, "asb_predictor_forward: null pointer");
  if (reinterpret_cast<uintptr_t>(features) & 15)
    return set_error(ASB_ERR_INVALID, "asb_predictor_forward: features must be 16-byte aligned");
  static DeviceOnce once;
  if (int rc = once.run([&] {
        cudaError_t e = cudaFuncSetAttribute(predictor_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kDecoderSmem));
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(predictor_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kDecoderSmem));
        return check_cuda(e, "predictor_kernel smem attribute");
      }))
    return rc;
  PredArgs a{features, out, rows, h->C, h->D, h->w};
  const unsigned grid = unsigned((rows + kBM - 1) / kBM);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (h->precision == ASB_PRECISION_FP32)
    predictor_kernel<true><<<grid, kThreads, kDecoderSmem, st>>>(a);
  else
    predictor_kernel<false><<<grid, kThreads, kDecoderSmem, st>>>(a);
  ASB_CHECK_LAUNCH("predictor_kernel");
  return ASB_OK;
}
