// Fused contour decoder on the 5th-generation tensor cores (tcgen05 + TMEM), fast mode.
//
// Same computation as decoder.cu (models.py:140-145 with ArticulatorPredictor :7-33) for a tile of
// 128 frames: linear.0 + ReLU, then per articulator LN -> FC(128,256) -> ReLU -> LN -> FC(256,256)
// -> ReLU -> LN -> x/y heads -> sigmoid, written once in the final (frames, N_art, 2, D) layout.
//
// Blackwell mapping
//   * every GEMM is D[128 x N] = A[128 x K] . W[N x K]^T with tcgen05.mma kind::tf32 (fp32 accumulate):
//     A (the activations) lives in TENSOR MEMORY (columns 256..511, one tf32 per column, lane = frame),
//     D accumulates in tensor memory (columns 0..255), W is streamed from L2 into shared memory by
//     1-D bulk async copies (UBLKCP) from a weight image that is pre-swizzled (128-byte swizzle,
//     K-major) at model-creation time, so no tensor map is needed.
//   * warp roles: warps 0-3 = epilogue (thread = frame = TMEM lane: bias, ReLU, LayerNorm, sigmoid are
//     all thread-local, no cross-thread reduction), warp 4 = weight/parameter producer, warp 5 = MMA
//     issuer (one elected thread).  Pipelines: weight ring full/empty mbarriers (tcgen05.commit frees
//     a stage), per-articulator parameter double buffer, A-ready (epilogue -> MMA) and D-ready
//     (tcgen05.commit -> epilogue) barriers.
//   * work unit = (128-frame tile, group of articulators); units are distributed round-robin over a
//     persistent grid of one CTA per SM.
#include "model.cuh"
#include "umma.cuh"

namespace {

using namespace asb;
using namespace asb::umma;

constexpr int kTileM = 128;
constexpr int kHeadN = 112;                 // 2 * 50 head rows padded to a multiple of 16
constexpr int kChunkK = 32;                 // tf32 per 128-byte swizzle row
constexpr int kStages = 4;
constexpr int kStageBytes = 256 * 128;      // 32 KB: up to 256 weight rows x 128 B
constexpr int kXStride = 129;               // x-hat row stride (floats), conflict-free per-thread rows
constexpr int kEpiWarps = 8;                // two warps per TMEM lane quarter, each owning half the columns
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kProducerWarp = 8, kMmaWarp = 9;
constexpr int kThreads = 320;

// weight image (floats).  LayerNorm affines are folded into the following layer at pack time:
// W' = W . diag(gamma), b' = b + W . beta, so the kernel only normalises.
constexpr int kLinParams = 128;
constexpr int kLinChunk = 128 * kChunkK;    // 4096
constexpr int kFcChunk = 256 * kChunkK;     // 8192
constexpr int kHeadChunk = kHeadN * kChunkK;  // 3584
constexpr int kArtParams = 640;             // b1' (256) | b2' (256) | bh' (128, zero padded)
constexpr int kLinBlock = kLinParams + 8 * kLinChunk;
constexpr int kArtBlock = kArtParams + 4 * kFcChunk + 8 * kFcChunk + 8 * kHeadChunk;
constexpr int P_B1 = 0, P_B2 = 256, P_BH = 512;

// dynamic shared memory layout (bytes, base 1024-aligned)
constexpr int kOffW = 0;
constexpr int kOffX = kOffW + kStages * kStageBytes;
constexpr int kOffP = kOffX + kTileM * kXStride * 4;
constexpr int kOffLinB = kOffP + 2 * kArtParams * 4;
constexpr int kOffStat = kOffLinB + kLinParams * 4;       // float2 [2 uses][2 halves][128 rows]
constexpr int kOffBar = kOffStat + 2 * 2 * kTileM * 8;
constexpr int kSmemBytes = kOffBar + 256 + 1024;  // + barriers + alignment slack

// dst image of (W . diag(scale))[N][K] as K/32 chunks of [Npad rows][32] with the 128-byte swizzle
__global__ void pack_umma_kernel(const float* __restrict__ src, const float* __restrict__ scale,
                                 float* __restrict__ dst, int n_src, int row0, int n_pad, int K) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_src * K) return;
  const int ns = i / K, k = i - ns * K;
  const int n = ns + row0;
  const int kc = k >> 5, kk = k & 31;
  const int off = kc * (n_pad * kChunkK) + (n >> 3) * 256 + (n & 7) * 32 + ((((kk >> 2) ^ (n & 7)) << 2) | (kk & 3));
  const float w = scale ? src[i] * scale[k] : src[i];
  dst[off] = __uint_as_float(to_tf32(w));
}
// dst[n] = b[n] + sum_k W[n][k] * beta[k]   (beta may be null)
__global__ void fold_bias_kernel(const float* __restrict__ W, const float* __restrict__ beta,
                                 const float* __restrict__ b, float* __restrict__ dst, int N, int K) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double acc = double(b[n]);
  if (beta)
    for (int k = 0; k < K; ++k) acc += double(W[(size_t)n * K + k]) * double(beta[k]);
  dst[n] = float(acc);
}

__device__ __forceinline__ void bar_sync_epilogue() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

__device__ __forceinline__ float sigmoid_fast(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

struct UmmaArgs {
  const float* image;    // weight image
  const float* rnn_out;  // (M, 256)
  float* out;            // (M, n_art, 2, D)
  long long M;
  int n_art, D, group, n_groups, n_units;
};

__global__ void __launch_bounds__(kThreads, 1) decoder_umma_kernel(const UmmaArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem =
      reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  unsigned char* Ws = smem + kOffW;
  float* Xs = reinterpret_cast<float*>(smem + kOffX);
  float* Ps = reinterpret_cast<float*>(smem + kOffP);
  float* LinB = reinterpret_cast<float*>(smem + kOffLinB);
  float2* Stat = reinterpret_cast<float2*>(smem + kOffStat);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* full = bars;            // [kStages]  weights landed
  uint64_t* empty = bars + 4;       // [kStages]  MMAs that read the stage completed
  uint64_t* pfull = bars + 8;       // [2]        articulator parameters landed
  uint64_t* pempty = bars + 10;     // [2]        epilogue finished with them
  uint64_t* a_ready = bars + 12;    //            A operand + bias-initialised accumulator are in TMEM
  uint64_t* d_ready = bars + 13;    //            accumulator complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&pfull[s], 1);
      mbar_init(&pempty[s], kEpiWarps);
    }
    mbar_init(a_ready, kEpiWarps);
    mbar_init(d_ready, 1);
    fence_mbar_init();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < kLinParams; i += kThreads) LinB[i] = a.image[i];
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_d = tmem_base;         // accumulator columns [0, 256)
  const uint32_t tmem_a = tmem_base + 256;   // A operand columns [256, 512)

  if (warp == kProducerWarp) {
    // ===================== producer: weight chunks and articulator parameters =====================
    if (lane == 0) {
      uint32_t c = 0, pa = 0;
      auto push = [&](const float* src, uint32_t bytes) {
        const uint32_t s = c % kStages;
        mbar_wait(&empty[s], ((c / kStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[s], bytes);
        bulk_g2s(Ws + s * kStageBytes, src, bytes, &full[s]);
        ++c;
      };
      for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
        const int grp = u % a.n_groups;
        const int g0 = grp * a.group, g1 = min(a.n_art, g0 + a.group);
        for (int j = 0; j < 8; ++j) push(a.image + kLinParams + j * kLinChunk, kLinChunk * 4);
        for (int art = g0; art < g1; ++art) {
          const float* blk = a.image + kLinBlock + (size_t)art * kArtBlock;
          const uint32_t pb = pa & 1;
          mbar_wait(&pempty[pb], ((pa >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&pfull[pb], kArtParams * 4);
          bulk_g2s(Ps + pb * kArtParams, blk, kArtParams * 4, &pfull[pb]);
          ++pa;
          const float* w1 = blk + kArtParams;
          const float* w2 = w1 + 4 * kFcChunk;
          const float* wh = w2 + 8 * kFcChunk;
          for (int j = 0; j < 4; ++j) push(w1 + j * kFcChunk, kFcChunk * 4);
          for (int j = 0; j < 8; ++j) push(w2 + j * kFcChunk, kFcChunk * 4);
          for (int j = 0; j < 8; ++j) push(wh + j * kHeadChunk, kHeadChunk * 4);
        }
      }
    }
    __syncwarp();
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      uint32_t c = 0, gcount = 0;
      auto gemm = [&](uint32_t n, int k_chunks) {
        mbar_wait(a_ready, gcount & 1);
        fence_after_sync();
        const uint32_t idesc = idesc_tf32_m128(n);
        for (int j = 0; j < k_chunks; ++j) {
          const uint32_t s = c % kStages;
          mbar_wait(&full[s], (c / kStages) & 1);
          fence_after_sync();
          const uint64_t bdesc = smem_desc_k_sw128(smem_u32(Ws + s * kStageBytes));
#pragma unroll
          for (int ks = 0; ks < 4; ++ks)  // the accumulator was pre-loaded with the bias: always accumulate
            mma_tf32_ts(tmem_d, tmem_a + j * kChunkK + ks * 8, bdesc + uint64_t(2 * ks), idesc, true);
          commit(&empty[s]);  // frees the stage once these MMAs have read it
          ++c;
        }
        commit(d_ready);
        ++gcount;
      };
      for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
        const int grp = u % a.n_groups;
        const int g0 = grp * a.group, g1 = min(a.n_art, g0 + a.group);
        gemm(128, 8);                       // linear.0          K = 256
        for (int art = g0; art < g1; ++art) {
          gemm(256, 4);                     // Linear(128, 256)
          gemm(256, 8);                     // Linear(256, 256)
          gemm(kHeadN, 8);                  // x/y heads
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== epilogue: thread = (frame = TMEM lane, half of the columns) ============
    const int quarter = warp & 3, half = warp >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = uint32_t(quarter * 32) << 16;
    const uint32_t td = tmem_d + lane_sel, ta = tmem_a + lane_sel;
    float* xrow = Xs + row * kXStride;
    uint32_t gcount = 0, pa = 0, use = 0;
    uint32_t r[32];

    auto signal_a = [&]() {
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_ready);
    };
    auto wait_d = [&]() {
      mbar_wait(d_ready, gcount & 1);
      ++gcount;
      fence_after_sync();
    };
    // row moments are split over the two column halves: exchange the partial sums through smem
    auto exchange = [&](float& s, float& ss) {
      Stat[((use & 1) * 2 + half) * kTileM + row] = make_float2(s, ss);
      bar_sync_epilogue();
      const float2 o = Stat[((use & 1) * 2 + (half ^ 1)) * kTileM + row];
      s += o.x;
      ss += o.y;
      ++use;
    };
    // pre-load this thread's accumulator columns [col0, col0 + 32 * chunks) with the next layer's bias
    auto init_d = [&](const float* bias, int col0, int chunks) {
      bar_sync_epilogue();  // the partner may still be reading columns that change owner
      for (int c = 0; c < chunks; ++c) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 b4 = *reinterpret_cast<const float4*>(bias + col0 + 32 * c + 4 * q);
          r[4 * q + 0] = __float_as_uint(b4.x);
          r[4 * q + 1] = __float_as_uint(b4.y);
          r[4 * q + 2] = __float_as_uint(b4.z);
          r[4 * q + 3] = __float_as_uint(b4.w);
        }
        tmem_st32(td + col0 + 32 * c, r);
      }
    };
    // relu(D) -> LayerNorm(256) without affine (folded into the next layer) -> tf32 -> A
    auto ln256_to_a = [&]() {
      const int col0 = half * 128;
      float s = 0.f, ss = 0.f;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        tmem_ld32(td + col0 + 32 * c, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float v = fmaxf(__uint_as_float(r[i]), 0.f);
          s += v;
          ss = fmaf(v, v, ss);
        }
      }
      exchange(s, ss);
      const float mean = s * (1.f / 256.f);
      const float var = fmaxf(ss * (1.f / 256.f) - mean * mean, 0.f);
      const float rstd = rsqrtf(var + 1e-5f);
      const float shift = -mean * rstd;
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        tmem_ld32(td + col0 + 32 * c, r);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) r[i] = to_tf32(fmaf(fmaxf(__uint_as_float(r[i]), 0.f), rstd, shift));
        tmem_st32(ta + col0 + 32 * c, r);
      }
    };

    for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
      const int tile = u / a.n_groups, grp = u % a.n_groups;
      const int g0 = grp * a.group, g1 = min(a.n_art, g0 + a.group);
      const long long grow = (long long)tile * kTileM + row;
      const bool live = grow < a.M;

      // ---- input tile -> A (tf32): this thread converts 128 of its frame's 256 features
      {
        const float4* src = reinterpret_cast<const float4*>(a.rnn_out + grow * 256 + half * 128);
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) v = __ldg(src + 8 * c + q);
            r[4 * q + 0] = to_tf32(v.x);
            r[4 * q + 1] = to_tf32(v.y);
            r[4 * q + 2] = to_tf32(v.z);
            r[4 * q + 3] = to_tf32(v.w);
          }
          tmem_st32(ta + half * 128 + 32 * c, r);
        }
        init_d(LinB, half * 64, 2);
        signal_a();
      }
      // ---- linear.0 epilogue: relu(D) -> LayerNorm(128) statistics -> x-hat kept in shared memory
      wait_d();
      {
        const int col0 = half * 64;
        float s = 0.f, ss = 0.f;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          tmem_ld32(td + col0 + 32 * c, r);
          tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float v = fmaxf(__uint_as_float(r[i]), 0.f);
            xrow[col0 + 32 * c + i] = v;
            s += v;
            ss = fmaf(v, v, ss);
          }
        }
        exchange(s, ss);
        const float mean = s * (1.f / 128.f);
        const float var = fmaxf(ss * (1.f / 128.f) - mean * mean, 0.f);
        const float rstd = rsqrtf(var + 1e-5f);
        const float shift = -mean * rstd;
#pragma unroll 8
        for (int k = 0; k < 64; ++k)
          xrow[col0 + k] = __uint_as_float(to_tf32(fmaf(xrow[col0 + k], rstd, shift)));
      }

      for (int art = g0; art < g1; ++art) {
        const uint32_t pb = pa & 1;
        const float* P = Ps + pb * kArtParams;
        mbar_wait(&pfull[pb], (pa >> 1) & 1);
        // ---- x-hat -> A (the LayerNorm(128) affine of this articulator lives in its folded W1', b1')
        {
          const int col0 = half * 64;
#pragma unroll 1
          for (int c = 0; c < 2; ++c) {
#pragma unroll
            for (int i = 0; i < 32; ++i) r[i] = __float_as_uint(xrow[col0 + 32 * c + i]);
            tmem_st32(ta + col0 + 32 * c, r);
          }
        }
        init_d(P + P_B1, half * 128, 4);
        signal_a();
        wait_d();
        ln256_to_a();
        init_d(P + P_B2, half * 128, 4);
        signal_a();
        wait_d();
        ln256_to_a();
        init_d(P + P_BH, half * 64, 2);
        signal_a();
        wait_d();
        // ---- heads: sigmoid(D) -> contours in their final layout (this thread: 64 of the 100 columns)
        {
          float* dst = a.out + (grow * a.n_art + art) * 100;
#pragma unroll 1
          for (int c = 0; c < 2; ++c) {
            const int cbase = half * 64 + 32 * c;
            if (cbase < 100) {  // warp-uniform
              tmem_ld32(td + cbase, r);
              tmem_ld_wait();
              if (live) {
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                  const int col = cbase + 4 * q;
                  if (col < 100) {
                    float4 o;
                    o.x = sigmoid_fast(__uint_as_float(r[4 * q + 0]));
                    o.y = sigmoid_fast(__uint_as_float(r[4 * q + 1]));
                    o.z = sigmoid_fast(__uint_as_float(r[4 * q + 2]));
                    o.w = sigmoid_fast(__uint_as_float(r[4 * q + 3]));
                    *reinterpret_cast<float4*>(dst + col) = o;
                  }
                }
              }
            }
          }
        }
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&pempty[pb]);
        ++pa;
      }
    }
  }

  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, 512);
}

}  // namespace

namespace asb {

size_t umma_image_floats(int n_art) { return (size_t)kLinBlock + (size_t)n_art * kArtBlock; }

bool umma_decoder_supported(const asb_model* m) {
  return m->precision == ASB_PRECISION_FAST && m->D == 50 && m->H == kH;
}

static int pack_mat(const float* src, const float* scale, float* dst, int n_src, int row0, int n_pad, int K,
                    cudaStream_t st) {
  const int n = n_src * K;
  pack_umma_kernel<<<(n + 255) / 256, 256, 0, st>>>(src, scale, dst, n_src, row0, n_pad, K);
  ASB_CHECK_LAUNCH("pack_umma_kernel");
  return ASB_OK;
}
static int fold_bias(const float* W, const float* beta, const float* b, float* dst, int N, int K, cudaStream_t st) {
  fold_bias_kernel<<<(N + 127) / 128, 128, 0, st>>>(W, beta, b, dst, N, K);
  ASB_CHECK_LAUNCH("fold_bias_kernel");
  return ASB_OK;
}

#define ASB_TRY(x)         \
  do {                     \
    int rc__ = (x);        \
    if (rc__) return rc__; \
  } while (0)

int pack_umma_image(asb_model* m, const asb_weights* s, cudaStream_t st) {
  float* img = m->w.umma_image;
  const size_t total = umma_image_floats(m->N);
  ASB_TRY(check_cuda(cudaMemsetAsync(img, 0, total * sizeof(float), st), "pack_umma_image: memset"));
  ASB_TRY(fold_bias(nullptr, nullptr, s->linear_b, img, kH, 0, st));
  ASB_TRY(pack_mat(s->linear_w, nullptr, img + kLinParams, kH, 0, kH, 2 * kH, st));
  const int D = m->D;
  for (int i = 0; i < m->N; ++i) {
    float* blk = img + kLinBlock + (size_t)i * kArtBlock;
    // LayerNorm affines folded forward: W' = W diag(gamma), b' = b + W beta   (models.py:11-21)
    ASB_TRY(fold_bias(s->fc1_w[i], s->ln0_b[i], s->fc1_b[i], blk + P_B1, kFc, kH, st));
    ASB_TRY(fold_bias(s->fc2_w[i], s->ln1_b[i], s->fc2_b[i], blk + P_B2, kFc, kFc, st));
    ASB_TRY(fold_bias(s->xh_w[i], s->ln2_b[i], s->xh_b[i], blk + P_BH, D, kFc, st));
    ASB_TRY(fold_bias(s->yh_w[i], s->ln2_b[i], s->yh_b[i], blk + P_BH + D, D, kFc, st));
    float* w1 = blk + kArtParams;
    float* w2 = w1 + 4 * kFcChunk;
    float* wh = w2 + 8 * kFcChunk;
    ASB_TRY(pack_mat(s->fc1_w[i], s->ln0_w[i], w1, kFc, 0, kFc, kH, st));
    ASB_TRY(pack_mat(s->fc2_w[i], s->ln1_w[i], w2, kFc, 0, kFc, kFc, st));
    ASB_TRY(pack_mat(s->xh_w[i], s->ln2_w[i], wh, D, 0, kHeadN, kFc, st));
    ASB_TRY(pack_mat(s->yh_w[i], s->ln2_w[i], wh, D, D, kHeadN, kFc, st));
  }
  return ASB_OK;
}

int launch_decoder_umma(const asb_model* m, const float* rnn_out, float* out, int M, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(decoder_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) return check_cuda(e, "decoder_umma_kernel smem attribute");
    configured = true;
  }
  const int sms = sm_count();
  const int n_tiles = (M + kTileM - 1) / kTileM;
  // articulators per work unit: minimise (waves) x (cycles per unit); linear.0 is redone per unit
  int best_g = m->N;
  double best_cost = 1e300;
  for (int g = 1; g <= m->N; ++g) {
    const int n_groups = (m->N + g - 1) / g;
    const long long units = (long long)n_tiles * n_groups;
    const double waves = double((units + sms - 1) / sms);
    const double cost = waves * (4500.0 + g * 12500.0);
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best_g = g;
    }
  }
  UmmaArgs a;
  a.image = m->w.umma_image;
  a.rnn_out = rnn_out;
  a.out = out;
  a.M = M;
  a.n_art = m->N;
  a.D = m->D;
  a.group = best_g;
  a.n_groups = (m->N + best_g - 1) / best_g;
  a.n_units = n_tiles * a.n_groups;
  const int grid = a.n_units < sms ? a.n_units : sms;
  decoder_umma_kernel<<<grid, kThreads, kSmemBytes, st>>>(a);
  ASB_CHECK_LAUNCH("decoder_umma_kernel");
  return ASB_OK;
}

}  // namespace asb
