// Layer-1 input projection of the recurrent stack on tcgen05 tensor cores, fast mode.
//
// gi[row, :] = out0[row, :] . W_ih_l1^T + b   for both directions at once (models.py:111: the
// second nn.GRU layer consumes [h_fwd | h_bwd] of the first), M = B*T rows, K = 2H = 256,
// N = 2*G*H (768 for the GRU).  Replaces the mma.sync linear256_kernel for this GEMM.
//
// Blackwell mapping (same building blocks as decoder_umma.cu)
//   * D[128 x 128] = A[128 x 256] . W[128 x 256]^T with tcgen05.mma kind::tf32, fp32 accumulate;
//     A (one 128-row tile of out0, rounded to tf32) lives in TENSOR MEMORY columns 256..511, the
//     accumulator is double buffered in columns 0..127 / 128..255, W streams from L2 through a
//     shared-memory ring of 16 KB stages by 1-D bulk async copies from a pre-swizzled image.
//   * work unit = (row tile, direction): A is loaded once and used for G column blocks of 128.
//   * warps 0-7: load A (thread = row, half of the features), then drain accumulators: bias add and
//     128-byte row segments straight to global memory while the next block's MMAs run;
//     warp 8: weight producer; warp 9: MMA issuer (one thread).
#include "model.cuh"
#include "umma.cuh"

namespace {

using namespace asb;
using namespace asb::umma;

constexpr int kTileM = 128;
constexpr int kBlockN = 128;
constexpr int kK = 256;
constexpr int kChunkK = 32;                       // tf32 per 128-byte swizzle row
constexpr int kChunks = kK / kChunkK;             // 8
constexpr int kStageFloats = kBlockN * kChunkK;   // 4096 floats = 16 KB
constexpr int kStages = 6;
constexpr int kEpiWarps = 8;
constexpr int kProducerWarp = 8, kMmaWarp = 9;
constexpr int kThreads = 320;
constexpr int kMaxN = 1024;                       // bias staged in shared memory

constexpr int kPitch = 36;                        // floats per staged row: 16-byte aligned, conflict-free
constexpr int kStageTile = 32 * kPitch;           // per-warp transposition tile [32 rows][32 + 4]

constexpr int kOffW = 0;
constexpr int kOffBias = kOffW + kStages * kStageFloats * 4;
constexpr int kOffTile = kOffBias + kMaxN * 4;
constexpr int kOffBar = kOffTile + kEpiWarps * kStageTile * 4;
constexpr int kSmemBytes = kOffBar + 256 + 1024;

// image block nb (128 output rows), chunk j: [128 rows][32] tf32 with the 128-byte swizzle
__global__ void pack_proj_kernel(const float* __restrict__ W, float* __restrict__ dst, int N) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * kK) return;
  const int n = i / kK, k = i - n * kK;
  const int nb = n / kBlockN, r = n % kBlockN;
  const int kc = k >> 5, kk = k & 31;
  const size_t off = (size_t)(nb * kChunks + kc) * kStageFloats + (r >> 3) * 256 + (r & 7) * 32 +
                     ((((kk >> 2) ^ (r & 7)) << 2) | (kk & 3));
  dst[off] = __uint_as_float(to_tf32(W[i]));
}

struct ProjArgs {
  const float* image;
  const float* bias;   // (N)
  const float* A;      // (M, 256)
  float* C;            // (M, N)
  long long M;
  int N, blocks_per_part, n_parts, n_units;
};

__global__ void __launch_bounds__(kThreads, 1) proj_umma_kernel(const ProjArgs a) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // keeps the shared address space
  float* Ws = reinterpret_cast<float*>(smem + kOffW);
  float* Bs = reinterpret_cast<float*>(smem + kOffBias);
  float* Ts = reinterpret_cast<float*>(smem + kOffTile);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* full = bars;                 // [kStages] weights landed
  uint64_t* empty = bars + kStages;      // [kStages] MMAs that read the stage completed
  uint64_t* d_full = bars + 2 * kStages;       // [2] accumulator complete
  uint64_t* d_empty = bars + 2 * kStages + 2;  // [2] accumulator drained
  uint64_t* a_ready = bars + 2 * kStages + 4;  //     A tile is in tensor memory
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 5);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&d_full[s], 1);
      mbar_init(&d_empty[s], kEpiWarps);
    }
    mbar_init(a_ready, kEpiWarps);
    fence_mbar_init();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < a.N; i += kThreads) Bs[i] = a.bias[i];
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_a = tmem_base + 256;

  if (warp == kProducerWarp) {
    if (lane == 0) {
      uint32_t c = 0;
      for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
        const int part = u % a.n_parts;
        for (int blk = 0; blk < a.blocks_per_part; ++blk) {
          const float* src = a.image + (size_t)(part * a.blocks_per_part + blk) * kChunks * kStageFloats;
          for (int j = 0; j < kChunks; ++j) {
            const uint32_t s = c % kStages;
            mbar_wait(&empty[s], ((c / kStages) & 1) ^ 1);
            mbar_arrive_expect_tx(&full[s], kStageFloats * 4);
            bulk_g2s(Ws + s * kStageFloats, src + (size_t)j * kStageFloats, kStageFloats * 4, &full[s]);
            ++c;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == kMmaWarp) {
    {   // the whole warp walks the loop, one elected lane issues (see decoder_umma.cu)
      const uint32_t idesc = idesc_tf32_m128(kBlockN);
      uint32_t c = 0, dcount = 0, ucount = 0;
      for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
        mbar_wait(a_ready, ucount & 1);
        ++ucount;
        fence_after_sync();
        for (int blk = 0; blk < a.blocks_per_part; ++blk) {
          const uint32_t b = dcount & 1;
          mbar_wait(&d_empty[b], ((dcount >> 1) & 1) ^ 1);
          fence_after_sync();
          const uint32_t tmem_d = tmem_base + b * kBlockN;
          for (int j = 0; j < kChunks; ++j) {
            const uint32_t s = c % kStages;
            mbar_wait(&full[s], (c / kStages) & 1);
            fence_after_sync();
            const uint64_t bdesc = smem_desc_k_sw128(smem_u32(Ws + s * kStageFloats));
            if (elect_one()) {
#pragma unroll
              for (int ks = 0; ks < 4; ++ks)
                mma_tf32_ts(tmem_d, tmem_a + j * kChunkK + ks * 8, bdesc + uint64_t(2 * ks), idesc, (j | ks) != 0);
              commit(&empty[s]);
              if (j == kChunks - 1) commit(&d_full[b]);
            }
            __syncwarp();
            ++c;
          }
          ++dcount;
        }
      }
    }
    __syncwarp();
  } else {
    // warp = (32 rows = one TMEM lane quarter, half of the columns).  Global memory is always
    // accessed as 4 rows x 128 B per instruction (lane -> row it*4 + lane/8, 16-byte column lane%8);
    // a per-warp shared-memory tile turns that into the thread-per-row form tcgen05.ld/st use.
    const int quarter = warp & 3, half = warp >> 2;
    const uint32_t lane_sel = uint32_t(quarter * 32) << 16;
    float* tile_s = Ts + warp * kStageTile;
    const int sub_row = lane >> 3, col4 = (lane & 7) * 4;
    uint32_t r[32];
    uint32_t dcount = 0;
    for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
      const int tile = u / a.n_parts, part = u % a.n_parts;
      const long long row0 = (long long)tile * kTileM + quarter * 32;
      // ---- A tile -> tensor memory (tf32).  The previous unit's MMAs have completed: its last
      // accumulator was waited for below.
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        float4 v[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const long long grow = row0 + it * 4 + sub_row;
          v[it] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (grow < a.M) v[it] = __ldg(reinterpret_cast<const float4*>(a.A + grow * kK + half * 128 + c * 32 + col4));
        }
#pragma unroll
        for (int it = 0; it < 8; ++it)
          *reinterpret_cast<float4*>(tile_s + (it * 4 + sub_row) * kPitch + col4) = v[it];
        __syncwarp();
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 t = *reinterpret_cast<const float4*>(tile_s + lane * kPitch + 4 * q);
          r[4 * q + 0] = to_tf32(t.x);
          r[4 * q + 1] = to_tf32(t.y);
          r[4 * q + 2] = to_tf32(t.z);
          r[4 * q + 3] = to_tf32(t.w);
        }
        tmem_st32(tmem_a + lane_sel + half * 128 + 32 * c, r);
        __syncwarp();
      }
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(a_ready);

      // ---- drain the accumulators of this unit
      for (int blk = 0; blk < a.blocks_per_part; ++blk) {
        const uint32_t b = dcount & 1;
        mbar_wait(&d_full[b], (dcount >> 1) & 1);
        ++dcount;
        fence_after_sync();
        const int n0 = (part * a.blocks_per_part + blk) * kBlockN + half * 64;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          tmem_ld32(tmem_base + lane_sel + b * kBlockN + half * 64 + 32 * c, r);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 8; ++q)
            *reinterpret_cast<float4*>(tile_s + lane * kPitch + 4 * q) =
                make_float4(__uint_as_float(r[4 * q + 0]), __uint_as_float(r[4 * q + 1]),
                            __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
          __syncwarp();
          const float4 bq = *reinterpret_cast<const float4*>(Bs + n0 + 32 * c + col4);
#pragma unroll
          for (int it = 0; it < 8; ++it) {
            const long long grow = row0 + it * 4 + sub_row;
            float4 o = *reinterpret_cast<const float4*>(tile_s + (it * 4 + sub_row) * kPitch + col4);
            o.x += bq.x; o.y += bq.y; o.z += bq.z; o.w += bq.w;
            if (grow < a.M) *reinterpret_cast<float4*>(a.C + grow * a.N + n0 + 32 * c + col4) = o;
          }
          __syncwarp();
        }
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(&d_empty[b]);
      }
    }
  }

  fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) {
    fence_after_sync();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace

namespace asb {

size_t proj_image_floats(const asb_model* m) { return (size_t)2 * m->G * kH * kK; }

bool proj_umma_supported(const asb_model* m) {
  return m->precision == ASB_PRECISION_FAST && 2 * m->G * kH <= kMaxN;
}

// from the packed row-major weights (already tf32-rounded in fast mode) to the swizzled image
int pack_proj_image(asb_model* m, cudaStream_t st) {
  const int N = 2 * m->G * kH;
  const int total = N * kK;
  pack_proj_kernel<<<(total + 255) / 256, 256, 0, st>>>(m->w.wih1, m->w.proj_image, N);
  ASB_CHECK_LAUNCH("pack_proj_kernel");
  return ASB_OK;
}

int launch_proj_umma(const asb_model* m, const float* A, float* C, int M, cudaStream_t st) {
  static DeviceOnce once;
  if (int rc = once.run([&] {
        return check_cuda(cudaFuncSetAttribute(proj_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes),
                          "proj_umma_kernel smem attribute");
      }))
    return rc;
  ProjArgs a;
  a.image = m->w.proj_image;
  a.bias = m->w.bih1;
  a.A = A;
  a.C = C;
  a.M = M;
  a.N = 2 * m->G * kH;
  a.n_parts = 2;
  a.blocks_per_part = m->G * kH / kBlockN;
  const int n_tiles = (M + kTileM - 1) / kTileM;
  a.n_units = n_tiles * a.n_parts;
  const int sms = sm_count();
  const int grid = a.n_units < sms ? a.n_units : sms;
  proj_umma_kernel<<<grid, kThreads, kSmemBytes, st>>>(a);
  ASB_CHECK_LAUNCH("proj_umma_kernel");
  return ASB_OK;
}

}  // namespace asb
