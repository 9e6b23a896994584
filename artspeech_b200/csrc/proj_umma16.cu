// Layer-1 input projection of the recurrent stack on tcgen05 tensor cores with FP16 OPERANDS (round 2).
//
// gi[row, :] = out0[row, :] . W_ih_l1^T + b   for both directions at once (models.py:111: the second
// nn.GRU layer consumes [h_fwd | h_bwd] of the first), M = B*T rows, K = 2H = 256, N = 2*G*H.
// Same job as proj_umma.cu (kind::tf32, kept as the "tcgen05-tf32" implementation); what changes:
//   * kind::f16 with fp16 operands (11-bit significand, tf32 has 10) and fp32 accumulation: twice the tensor
//     rate and half the weight stream (196 KB per unit instead of 393 KB); the weights are rounded to fp16
//     straight from the fp32 source (not from the tf32-rounded copy);
//   * the A tile (128 rows of out0) is 128 TMEM columns of fp16 pairs instead of 256, which leaves room for TWO
//     A buffers beside the double-buffered accumulator: the next unit's tile is loaded and converted while the
//     MMAs of the current one run (the tf32 kernel serialises load -> MMAs -> drain per unit);
//   * one MMA-warp iteration per 64-k weight block (4 MMAs, one commit) instead of per 32-k chunk;
//   * kSplit = true (fp32 precision): operands are hi + lo fp16 pairs, hi = fp16(v), lo = fp16(v - hi), and a
//     product is three MMAs (a_lo w_hi + a_hi w_lo + a_hi w_hi) - 22 significand bits, ~1e-6 on gi; A then takes
//     all 256 columns again (one buffer) and every weight piece is followed through the ring by its residual.
// Warps 0-7: A loader / accumulator drain (thread = row, half of the columns), warp 8: weight producer,
// warp 9: MMA issuer (converged warp, one elected lane).  Work unit = (row tile, direction).
#include <cuda_fp16.h>

#include "model.cuh"
#include "umma.cuh"

namespace {

using namespace asb;
using namespace asb::umma;

constexpr int kTileM = 128;
constexpr int kBlockN = 128;
constexpr int kK = 256;
constexpr int kBlockK = 64;                        // fp16 per 128-byte swizzle row
constexpr int kKBlocks = kK / kBlockK;             // 4
constexpr int kStageBytes = kBlockN * 128;         // 16 KB: [128 rows][64 fp16]
constexpr int kStageFloats = kStageBytes / 4;
constexpr int kStages = 8;
constexpr int kEpiWarps = 8;
constexpr int kProducerWarp = 8, kMmaWarp = 9;
constexpr int kThreads = 320;
constexpr int kMaxN = 1024;                        // bias staged in shared memory

constexpr int kPitch = 36;                         // floats per staged row: 16-byte aligned, conflict-free
constexpr int kStageTile = 32 * kPitch;            // per-warp transposition tile [32 rows][32 + 4]

constexpr int kOffW = 0;
constexpr int kOffBias = kOffW + kStages * kStageBytes;
constexpr int kOffTile = kOffBias + kMaxN * 4;
constexpr int kOffBar = kOffTile + kEpiWarps * kStageTile * 4;
constexpr int kSmemBytes = kOffBar + 256 + 1024;

// image piece (N-block nb, k-block kb): [128 rows][64 fp16] with the 128-byte swizzle; split: the hi piece is
// followed by its residual piece
__global__ void pack_proj16_kernel(const float* __restrict__ W0, const float* __restrict__ W1, __half* __restrict__ dst,
                                   int n_per_dir, int split) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2 * n_per_dir * kK) return;
  const int n = i / kK, k = i - n * kK;
  const float w = n < n_per_dir ? W0[(size_t)n * kK + k] : W1[(size_t)(n - n_per_dir) * kK + k];
  const int nb = n / kBlockN, r = n % kBlockN;
  const int kb = k >> 6, kk = k & 63;
  const int in_piece = (r >> 3) * 512 + (r & 7) * 64 + ((((kk >> 3) ^ (r & 7)) << 3) | (kk & 7));
  const int piece = kBlockN * kBlockK;
  const __half hi = __float2half_rn(w);
  if (!split) {
    dst[(size_t)(nb * kKBlocks + kb) * piece + in_piece] = hi;
  } else {
    const size_t base = (size_t)(nb * kKBlocks + kb) * 2 * piece + in_piece;
    dst[base] = hi;
    dst[base + piece] = __float2half_rn(w - __half2float(hi));
  }
}

struct ProjArgs {
  const float* image;
  const float* bias;   // (N)
  const float* A;      // (M, 256)
  float* C;            // (M, N)
  long long M;
  int N, blocks_per_part, n_parts, n_units;
};

template <bool kSplit>
__global__ void __launch_bounds__(kThreads, 1) proj_umma16_kernel(const ProjArgs a) {
  constexpr int kW = kSplit ? 2 : 1;              // ring stages per weight piece (hi, lo)
  constexpr int kABufs = kSplit ? 1 : 2;          // A tiles in tensor memory
  constexpr int kChunkCols = kSplit ? 32 : 16;    // TMEM columns of a 32-k chunk of A (split: 16 hi + 16 lo)
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // keeps the shared address space
  unsigned char* Ws = smem + kOffW;
  float* Bs = reinterpret_cast<float*>(smem + kOffBias);
  float* Ts = reinterpret_cast<float*>(smem + kOffTile);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* full = bars;                       // [kStages] weights landed
  uint64_t* empty = bars + kStages;            // [kStages] MMAs that read the stage completed
  uint64_t* d_full = bars + 2 * kStages;       // [2] accumulator complete
  uint64_t* d_empty = bars + 2 * kStages + 2;  // [2] accumulator drained
  uint64_t* a_ready = bars + 2 * kStages + 4;  // [2] A tile is in tensor memory
  uint64_t* a_free = bars + 2 * kStages + 6;   // [2] the MMAs that read the A tile completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&d_full[s], 1);
      mbar_init(&d_empty[s], kEpiWarps);
      mbar_init(&a_ready[s], kEpiWarps);
      mbar_init(&a_free[s], 1);
    }
    fence_mbar_init();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < a.N; i += kThreads) Bs[i] = a.bias[i];
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_a = tmem_base + 256;     // buffer b at + 128 b (fast); one 256-column buffer (split)

  if (warp == kProducerWarp) {
    if (lane == 0) {
      uint32_t c = 0;
      for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
        const int part = u % a.n_parts;
        for (int blk = 0; blk < a.blocks_per_part; ++blk) {
          const float* src = a.image + (size_t)(part * a.blocks_per_part + blk) * kKBlocks * kW * kStageFloats;
          for (int j = 0; j < kKBlocks * kW; ++j) {   // split: the pieces alternate hi, lo in the image
            const uint32_t s = c % kStages;
            mbar_wait(&empty[s], ((c / kStages) & 1) ^ 1);
            mbar_arrive_expect_tx(&full[s], kStageBytes);
            bulk_g2s(Ws + s * kStageBytes, src + (size_t)j * kStageFloats, kStageBytes, &full[s]);
            ++c;
          }
        }
      }
    }
    __syncwarp();
  } else if (warp == kMmaWarp) {
    {   // the whole warp walks the loop, one elected lane issues (see decoder_umma16.cu)
      const uint32_t idesc = idesc_f16_m128(kBlockN);
      uint32_t c = 0, dcount = 0, ucount = 0;
      for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
        const uint32_t buf = ucount % kABufs;
        mbar_wait(&a_ready[buf], (ucount / kABufs) & 1);
        fence_after_sync();
        const uint32_t abase = tmem_a + buf * 128;
        for (int blk = 0; blk < a.blocks_per_part; ++blk) {
          const uint32_t b = dcount & 1;
          mbar_wait(&d_empty[b], ((dcount >> 1) & 1) ^ 1);
          fence_after_sync();
          const uint32_t tmem_d = tmem_base + b * kBlockN;
          for (int kb = 0; kb < kKBlocks; ++kb) {
            const uint32_t s = c % kStages, s_lo = (c + 1) % kStages;
            mbar_wait(&full[s], (c / kStages) & 1);
            if (kSplit) mbar_wait(&full[s_lo], ((c + 1) / kStages) & 1);
            fence_after_sync();
            const uint64_t bdesc = smem_desc_k_sw128(smem_u32(Ws + s * kStageBytes));
            const uint64_t bdesc_lo = smem_desc_k_sw128(smem_u32(Ws + s_lo * kStageBytes));
            if (elect_one()) {
#pragma unroll
              for (int ks = 0; ks < 4; ++ks) {
                // 16 k = 32 bytes of the swizzled row; A: chunk 2 kb + ks / 2, fp16 pairs of its k-step ks % 2
                const uint32_t ac = abase + (2 * kb + (ks >> 1)) * kChunkCols + (ks & 1) * 8;
                const uint64_t bd = bdesc + uint64_t(2 * ks), bd_lo = bdesc_lo + uint64_t(2 * ks);
                if (kSplit) {   // small terms first
                  mma_bf16_ts(tmem_d, ac + 16, bd, idesc, (kb | ks) != 0);   // a_lo w_hi
                  mma_bf16_ts(tmem_d, ac, bd_lo, idesc, true);               // a_hi w_lo
                  mma_bf16_ts(tmem_d, ac, bd, idesc, true);                  // a_hi w_hi
                } else {
                  mma_bf16_ts(tmem_d, ac, bd, idesc, (kb | ks) != 0);
                }
              }
              commit(&empty[s]);
              if (kSplit) commit(&empty[s_lo]);
              if (kb == kKBlocks - 1) {
                commit(&d_full[b]);
                if (blk == a.blocks_per_part - 1) commit(&a_free[buf]);   // the A buffer may be refilled
              }
            }
            __syncwarp();
            c += kW;
          }
          ++dcount;
        }
        ++ucount;
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
    uint32_t dcount = 0, loads = 0;

    // A tile of unit u -> tensor memory as fp16 pairs (split: + residual pairs), this warp's 32 rows x 128 features
    auto load_a = [&](int u) {
      const int tile = u / a.n_parts;
      const long long row0 = (long long)tile * kTileM + quarter * 32;
      const uint32_t buf = loads % kABufs;
      mbar_wait(&a_free[buf], ((loads / kABufs) & 1) ^ 1);   // the MMAs of the buffer's previous tile completed
      fence_after_sync();
      const uint32_t abase = tmem_a + lane_sel + buf * 128;
      float4 v[8];
      auto fetch = [&](int c) {
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const long long grow = row0 + it * 4 + sub_row;
          v[it] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (grow < a.M) v[it] = __ldg(reinterpret_cast<const float4*>(a.A + grow * kK + half * 128 + c * 32 + col4));
        }
      };
      fetch(0);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
#pragma unroll
        for (int it = 0; it < 8; ++it)
          *reinterpret_cast<float4*>(tile_s + (it * 4 + sub_row) * kPitch + col4) = v[it];
        __syncwarp();
        if (c < 3) fetch(c + 1);   // the next chunk's loads fly under this chunk's conversion
        uint32_t h[16], l[kSplit ? 16 : 1];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const float4 t = *reinterpret_cast<const float4*>(tile_s + lane * kPitch + 4 * q);
          h[2 * q + 0] = pack_f16x2(t.x, t.y);
          h[2 * q + 1] = pack_f16x2(t.z, t.w);
          if constexpr (kSplit) {
            l[2 * q + 0] = residual_f16x2(t.x, t.y, h[2 * q + 0]);
            l[2 * q + 1] = residual_f16x2(t.z, t.w, h[2 * q + 1]);
          }
        }
        const uint32_t col = abase + (half * 4 + c) * kChunkCols;
        tmem_st16(col, h);
        if constexpr (kSplit) tmem_st16(col + 16, l);
        __syncwarp();   // the tile is rewritten by the next chunk
      }
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_ready[buf]);
      ++loads;
    };

    if (blockIdx.x < a.n_units) load_a(blockIdx.x);
    for (int u = blockIdx.x; u < a.n_units; u += gridDim.x) {
      const int tile = u / a.n_parts, part = u % a.n_parts;
      const long long row0 = (long long)tile * kTileM + quarter * 32;
      const int next = u + gridDim.x;
      if (!kSplit && next < a.n_units) load_a(next);   // second A buffer: under this unit's MMAs

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
      if (kSplit && next < a.n_units) load_a(next);    // one A buffer: after this unit's last accumulator
    }
  }

  fence_before_sync();
  __syncthreads();
  if (warp == kMmaWarp) {
    fence_after_sync();
    tmem_dealloc(tmem_base, 512);
  }
}

bool split_of(const asb_model* m) { return m->precision != ASB_PRECISION_FAST; }

}  // namespace

namespace asb {

size_t proj16_image_floats(const asb_model* m) {
  return (size_t)2 * m->G * kH * kK / 2 * (split_of(m) ? 2 : 1);   // fp16: two per float
}

bool proj_umma16_supported(const asb_model* m) { return m->H == kH && 2 * m->G * kH <= kMaxN; }

int pack_proj16_image(asb_model* m, const asb_weights* src, cudaStream_t st) {
  const int n_per_dir = m->G * kH;
  const int total = 2 * n_per_dir * kK;
  pack_proj16_kernel<<<(total + 255) / 256, 256, 0, st>>>(src->rnn_w_ih[1][0], src->rnn_w_ih[1][1],
                                                          reinterpret_cast<__half*>(m->w.proj16_image), n_per_dir,
                                                          split_of(m) ? 1 : 0);
  ASB_CHECK_LAUNCH("pack_proj16_kernel");
  return ASB_OK;
}

int launch_proj_umma16(const asb_model* m, const float* A, float* C, int M, cudaStream_t st) {
  static DeviceOnce once;
  if (int rc = once.run([&] {
        if (int r = check_cuda(cudaFuncSetAttribute(proj_umma16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                    kSmemBytes),
                               "proj_umma16_kernel smem attribute"))
          return r;
        return check_cuda(cudaFuncSetAttribute(proj_umma16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               kSmemBytes),
                          "proj_umma16_kernel (split) smem attribute");
      }))
    return rc;
  ProjArgs a;
  a.image = m->w.proj16_image;
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
  if (split_of(m))
    proj_umma16_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(a);
  else
    proj_umma16_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(a);
  ASB_CHECK_LAUNCH("proj_umma16_kernel");
  return ASB_OK;
}

}  // namespace asb
