// Fused contour decoder on the 5th-generation tensor cores (tcgen05 + TMEM), fast mode, FP16 OPERANDS.
//
// Round-2 variant of decoder_umma.cu (kind::tf32).  Same structure - warp roles, barriers, deferred
// LayerNorms, in-place accumulator -> operand conversion - but every GEMM is tcgen05.mma kind::f16 with
// fp16 operands and fp32 accumulation: fp16 carries an 11-bit significand (tf32: 10), so the contours are
// at least as accurate, while the tensor pipe runs kind::f16 at twice the tf32 rate, the streamed weight
// image is half the bytes (248 KB per articulator instead of 496 KB) and the epilogue stores half the
// tensor-memory columns.  What changes against the tf32 kernel:
//   * the A operand of a 32-wide K chunk is 16 TMEM columns (two fp16 per column) written into the first
//     half of the chunk's own 32 accumulator columns, so the in-place conversion still never touches a
//     column another warp has yet to read; two MMAs (K = 16) per chunk instead of four (K = 8);
//   * a weight block is [rows][64 fp16] = 128-byte rows (128-byte swizzle as before) and serves two K
//     chunks; FC1 / FC2 stream 32 KB blocks (one per ring stage), linear.0 and the heads two blocks per stage;
//   * values are rounded with cvt.rn.satfinite.f16x2.f32 (un-normalised ReLU outputs stay far below the
//     fp16 range for any sane checkpoint; satfinite keeps an outlier finite instead of producing NaN).
//
// FP32 PRECISION (kSplit = true): the same kernel with SPLIT operands.  Every activation and every weight is
// carried as hi + lo, hi = fp16(v), lo = fp16(v - hi) (22 significand bits together), and a product is three
// MMAs, a_lo w_hi + a_hi w_lo + a_hi w_hi, accumulated in fp32 in tensor memory (the dropped a_lo w_lo term is
// below 2^-22 of the product).  hi pairs occupy the first 16 columns of a chunk's 32 accumulator columns, the
// residual pairs the second 16 (all 32 were loaded into registers before either half is stored); a weight
// piece is followed through the ring by its residual piece, so a product costs two ring stages and the
// image is twice the fast one.  Contours agree with the fp32 FFMA decoder (decoder.cu) to ~1e-6.
//
// Same computation as decoder.cu (models.py:140-145 with ArticulatorPredictor :7-33) for a tile of
// 128 frames: linear.0 + ReLU, then per articulator LN -> FC(128,256) -> ReLU -> LN -> FC(256,256)
// -> ReLU -> LN -> x/y heads -> sigmoid, written once in the final (frames, N_art, 2, D) layout.
//
// Blackwell mapping
//   * every GEMM is D[128 x N] = A[128 x K] . W[N x K]^T with tcgen05.mma kind::f16 on fp16 operands (fp32 accumulate):
//     A (the activations) lives in TENSOR MEMORY (one tf32 per column, lane = frame), D accumulates
//     in tensor memory, W is streamed from L2 into shared memory by 1-D bulk async copies (UBLKCP)
//     from a weight image that is pre-swizzled (128-byte swizzle, K-major) at model-creation time,
//     so no tensor map is needed.
//   * the 512 TMEM columns are two 256-column regions that swap roles from layer to layer: a GEMM
//     reads A from one region and accumulates into the other; the epilogue then turns that
//     accumulator IN PLACE into the next GEMM's A operand, 32 columns at a time, and signals each
//     32-column K chunk on its own mbarrier - so the next GEMM starts on chunk 0 while the epilogue
//     is still converting chunks 1..7 (the tensor pipe and the epilogue warps overlap inside one
//     tile).  The in-place single pass is possible because the LayerNorms are DEFERRED: with
//     W' = W diag(gamma), b' = b + W beta folded at pack time,
//        W' ((x - mean) rstd) + b' = rstd (W' x) + (b' - rstd mean (W' 1)),
//     so the GEMM consumes the un-normalised relu output x (rounded to tf32 like before), the row
//     moments are accumulated while x is produced, and the following epilogue applies rstd and the
//     rank-1 mean term (W' 1 = row sums of the rounded weights, precomputed) as two FMAs.
//   * warp roles: warps 0-7 = epilogue (thread = frame = TMEM lane, the two warps of a lane quarter
//     own alternate 32-column chunks), warp 8 = weight/parameter producer, warp 9 = MMA issuer
//     (converged warp, one elected lane), warps 10-13 = output (heads accumulator -> sigmoid ->
//     global memory, off the epilogue's critical path).  Pipelines: 4-stage 32 KB weight ring with
//     full/empty mbarriers (tcgen05.commit frees a stage), per-articulator parameter double buffer,
//     per-K-chunk A-ready barriers (epilogue -> MMA), one D-complete barrier per layer
//     (tcgen05.commit -> epilogue / output warps) and a heads-accumulator-read barrier (output ->
//     MMA).  The A operand of the next articulator's first GEMM is written while the heads GEMM of
//     the current one runs.
//   * work = the (tile, articulator) pairs, tile-major, cut into one contiguous range per CTA of a
//     persistent grid (one CTA per SM); see UnitIter.
#include <cuda_fp16.h>

#include "model.cuh"
#include "umma.cuh"

// timing ablations for tools/decoder_variants.py (results are wrong when set): 1 = epilogue without
// arithmetic, 2 = epilogue without TMEM traffic either, 3 = no MMAs issued, 4 = no weight streaming
#ifndef ASB_DEC_ABLATE
#define ASB_DEC_ABLATE 0
#endif
// 1 = every lane stores 16-byte pieces of its own row (round-2 first form) instead of the transposed, coalesced stores
#ifndef ASB_DEC_OUT_DIRECT
#define ASB_DEC_OUT_DIRECT 0
#endif

#ifdef ASB_DEC_TRACE
// timeline of CTA 0 (tools/decoder_trace.py): [role][event] = (tag, clock64) for the MMA thread (0),
// lane 0 of epilogue warp 0 (1) and of epilogue warp 4 (2), the producer thread (3), lane 0 of the output warps (4..7)
__device__ long long g_dec16_trace[8 * 4096];
#define TR(role, tag)                                                     \
  do {                                                                    \
    if (blockIdx.x == 0 && lane == 0 && tr_n < 2047) {                    \
      g_dec16_trace[(role) * 4096 + 2 * tr_n] = (tag);                      \
      g_dec16_trace[(role) * 4096 + 2 * tr_n + 1] = clock64();              \
      ++tr_n;                                                             \
    }                                                                     \
  } while (0)
extern "C" __attribute__((visibility("default"))) int asb_debug_decoder16_trace(long long* host_out) {
  return int(cudaMemcpyFromSymbol(host_out, g_dec16_trace, sizeof(g_dec16_trace)));
}
#define TR_DECL int tr_n = 0
#else
#define TR(role, tag) do { } while (0)
#define TR_DECL do { } while (0)
#endif

namespace {

using namespace asb;
using namespace asb::umma;

constexpr int kTileM = 128;
constexpr int kHeadN = 112;                 // 2 * 50 head rows padded to a multiple of 16
constexpr int kChunkK = 32;                 // K extent of one A chunk (one A-ready barrier): 16 TMEM columns of fp16 pairs
constexpr int kBlockK = 64;                 // fp16 per 128-byte swizzle row: one weight block serves two A chunks
#ifndef ASB_DEC_STAGES
#define ASB_DEC_STAGES 4
#endif
// (measured with 2 / 3 / 4 stages: 0.098 / 0.100 / 0.100 ms fast precision - the ring is not what bounds the kernel;
// four keep two hi + lo piece pairs in flight in fp32 precision)
constexpr int kStages = ASB_DEC_STAGES;     // weight ring: 32 KB stages = one 256-row block or two 128- / 112-row ones
constexpr int kStageBytes = 256 * 128;
constexpr int kXStride = 129;               // x-hat row stride (floats), conflict-free per-thread rows
constexpr int kTilePitch = 36;              // per-warp 32 x 32 transposition tile of the input load (floats per row)
constexpr int kEpiWarps = 8;                // two warps per TMEM lane quarter, each owning half the columns
constexpr int kProducerWarp = 8, kMmaWarp = 9;
constexpr int kHeadWarp0 = 10, kHeadWarps = 4;   // warps 10..13: output warps, TMEM lane quarter = warp % 4
constexpr int kThreads = 32 * (kHeadWarp0 + kHeadWarps);

// weight image (floats).  LayerNorm affines are folded into the following layer at pack time:
// W' = W . diag(gamma), b' = b + W . beta, so the kernel only normalises.
// (sizes in FLOAT units of the image array; a weight block of R rows is R * 128 bytes = R * 32 floats)
constexpr int kLinParams = 128;
constexpr int kLinChunk = 128 * 32;         // one 64-k block of linear.0: 128 rows x 128 B
constexpr int kFcChunk = 256 * 32;          // one 64-k block of FC1 / FC2: 256 rows x 128 B = 32 KB
constexpr int kHeadChunk = kHeadN * 32;     // one 64-k block of the heads: 112 rows x 128 B = 14 KB
constexpr int kArtParams = 1024;            // b1' (256) | b2' (256) | W2'1 (256) | bh' (128, zero padded) | Wh'1 (128)
// split (fp32 precision) images hold every piece twice: [piece hi][piece lo], a piece = one ring stage
__host__ __device__ constexpr int lin_block(bool split) { return kLinParams + (split ? 2 : 1) * 4 * kLinChunk; }
__host__ __device__ constexpr int art_block(bool split) {
  return kArtParams + (split ? 2 : 1) * (2 * kFcChunk + 4 * kFcChunk + 4 * kHeadChunk);
}
constexpr int P_B1 = 0, P_B2 = 256, P_WS2 = 512, P_BH = 768, P_WSH = 896;

// dynamic shared memory layout (bytes, base 1024-aligned)
constexpr int kOffW = 0;
constexpr int kOffX = kOffW + kStages * kStageBytes;
constexpr int kOffP = kOffX + kTileM * kXStride * 4;
constexpr int kOffLinB = kOffP + 2 * kArtParams * 4;
constexpr int kOffStat = kOffLinB + kLinParams * 4;       // float2 [8][128 rows]: see Stat0/1/2 in the kernel
constexpr int kOffOut = kOffStat + 8 * kTileM * 8;          // [4 output warps][32 rows][32 floats]: store transposition
constexpr int kOutTile = 32 * 32 * 4;
constexpr int kOffBar = kOffOut + kHeadWarps * kOutTile;
constexpr int kSmemBytes = kOffBar + 512 + 1024;  // + barriers + alignment slack

__device__ __forceinline__ float round_f16(float x) { return __half2float(__float2half_rn(x)); }

// dst image of (W . diag(scale))[N][K] in fp16 as K/64 blocks of [Npad rows][64] with the 128-byte swizzle
// (16-byte unit u of a row lands at u ^ (row & 7))
// bpp = blocks per piece (ring stage).  split: piece p is stored as [hi blocks][lo blocks], lo = fp16(w - fp16(w)).
__global__ void pack_umma16_kernel(const float* __restrict__ src, const float* __restrict__ scale,
                                   __half* __restrict__ dst, int n_src, int row0, int n_pad, int K, int bpp,
                                   int split) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_src * K) return;
  const int ns = i / K, k = i - ns * K;
  const int n = ns + row0;
  const int kc = k >> 6, kk = k & 63;
  const int block = n_pad * kBlockK;
  const int in_block = (n >> 3) * 512 + (n & 7) * 64 + ((((kk >> 3) ^ (n & 7)) << 3) | (kk & 7));
  const int piece = kc / bpp, bi = kc - piece * bpp;
  const float w = scale ? src[i] * scale[k] : src[i];
  const __half hi = __float2half_rn(w);
  if (!split) {
    dst[kc * block + in_block] = hi;
  } else {
    const int base = piece * 2 * bpp * block + bi * block + in_block;
    dst[base] = hi;
    dst[base + bpp * block] = __float2half_rn(w - __half2float(hi));
  }
}
// dst[n] = sum_k fp16(W[n][k] * scale[k]): row sums of the folded weights exactly as the tensor cores
// see them (the mean term of the deferred LayerNorm)
__global__ void fold_wsum16_kernel(const float* __restrict__ W, const float* __restrict__ scale,
                                   float* __restrict__ dst, int N, int K, double out_scale, int split) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double acc = 0.0;
  for (int k = 0; k < K; ++k) {
    const float w = W[(size_t)n * K + k] * scale[k];
    const float hi = round_f16(w);
    acc += double(hi);
    if (split) acc += double(round_f16(w - hi));
  }
  dst[n] = float(acc * out_scale);
}
// dst[n] = b[n] + sum_k W[n][k] * beta[k]   (beta may be null)
__global__ void fold_bias16_kernel(const float* __restrict__ W, const float* __restrict__ beta,
                                 const float* __restrict__ b, float* __restrict__ dst, int N, int K,
                                 double out_scale) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  double acc = double(b[n]);
  if (beta)
    for (int k = 0; k < K; ++k) acc += double(W[(size_t)n * K + k]) * double(beta[k]);
  dst[n] = float(acc * out_scale);
}

// residual store of the split variant; the one-element overload exists only so that the fast instantiation compiles
__device__ __forceinline__ void st_lo16(uint32_t taddr, const uint32_t (&r)[16]) { tmem_st16(taddr, r); }
__device__ __forceinline__ void st_lo16(uint32_t, const uint32_t (&)[1]) {}

__device__ __forceinline__ void bar_sync_epilogue() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// 1 / (1 + 2^t): the sigmoid of z with t = -log2(e) z (the factor is folded into the head parameters)
__device__ __forceinline__ float sigmoid_exp2(float t) {
  float e, y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(t));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(1.f + e));
  return y;
}

// parameters are read with explicit shared-space loads (all threads read the same address: broadcast)
__device__ __forceinline__ float4 lds4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

struct UmmaArgs {
  const float* image;    // weight image
  const float* rnn_out;  // (M, 256)
  float* out;            // (M, n_art, 2, D)
  long long M;
  int n_art, D;
  long long n_items;    // (tile, articulator) pairs = tiles * n_art
};

// Work distribution: the (tile, articulator) pairs, tile-major, are cut into gridDim.x contiguous
// ranges of equal length; a range is walked as units = (tile, run of articulators of that tile), and
// linear.0 is (re)computed once per unit.  Every warp role iterates the same sequence.
struct UnitIter {
  long long item, hi;
  int n_art;
  __device__ UnitIter(const UmmaArgs& a)
      : item(a.n_items * blockIdx.x / gridDim.x), hi(a.n_items * (blockIdx.x + 1) / gridDim.x), n_art(a.n_art) {}
  __device__ bool next(int& tile, int& g0, int& g1) {
    if (item >= hi) return false;
    tile = int(item / n_art);
    g0 = int(item - (long long)tile * n_art);
    g1 = int(min((long long)n_art, g0 + (hi - item)));
    item += g1 - g0;
    return true;
  }
};

// TMEM regions.  With p = parity of the articulator counter, P = region p, Q = the other one:
//   linear.0 : A0 = Q[0,256)    -> D0 = P[0,128)                (epilogue: x-hat -> shared memory)
//   FC1      : A1 = P[128,256)  -> D1 = Q[0,256)   in place ->  A2 = Q
//   FC2      : A2 = Q           -> D2 = P[0,256)   in place ->  A3 = P
//   heads    : A3 = P           -> D3 = Q[0,112)                (read by the output warps)
// and the next articulator (parity flipped) puts its A1 into Q[128,256) while the heads GEMM runs;
// its FC1 and the first FC2 epilogue chunks overlap the output warps' sigmoid + stores.
template <bool kSplit>
__global__ void __launch_bounds__(kThreads, 1) decoder_umma16_kernel(const UmmaArgs a) {
  constexpr int kW = kSplit ? 2 : 1;   // ring stages per weight piece (hi, lo)
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // keeps the shared address space
  unsigned char* Ws = smem + kOffW;
  float* Xs = reinterpret_cast<float*>(smem + kOffX);
  float* Ps = reinterpret_cast<float*>(smem + kOffP);
  float* LinB = reinterpret_cast<float*>(smem + kOffLinB);
  float2* Stat0 = reinterpret_cast<float2*>(smem + kOffStat);   // [2 halves][128]  LayerNorm(128) partial moments
  float2* Stat1 = Stat0 + 2 * kTileM;                           // [2 halves][128]  after FC1
  float2* Stat2 = Stat1 + 2 * kTileM;                           // [2 articulator parities][2 halves][128]  after FC2
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kOffBar);
  uint64_t* full = bars;            // [kStages]  weights landed
  uint64_t* empty = bars + 8;       // [kStages]  MMAs that read the stage completed
  uint64_t* pfull = bars + 16;      // [2]        articulator parameters landed
  uint64_t* pempty = bars + 18;     // [2]        epilogue and output warps finished with them
  uint64_t* chunk = bars + 20;      // [4]        64 k (two 32-column chunks) of the A operand are in TMEM (K = 256 GEMMs)
  uint64_t* a1bar = bars + 28;      // [2]        the same for the A operand of FC1 (K = 128)
  uint64_t* dbar = bars + 32;       // [4]        accumulator of linear.0 / FC1 / FC2 / heads complete
  uint64_t* d3_free = bars + 36;    //            the output warps have read the heads accumulator
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 37);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&pfull[s], 1);
      mbar_init(&pempty[s], kEpiWarps + kHeadWarps);
    }
    for (int j = 0; j < 4; ++j) mbar_init(&chunk[j], 8);   // 64 k of A: one 32-k chunk of each half x four lane quarters
    for (int j = 0; j < 2; ++j) mbar_init(&a1bar[j], 8);
    for (int j = 0; j < 4; ++j) mbar_init(&dbar[j], 1);
    mbar_init(d3_free, kHeadWarps);
    fence_mbar_init();
  }
  if (warp == kMmaWarp) tmem_alloc(tmem_slot, 512);
  for (int i = tid; i < kLinParams; i += kThreads) LinB[i] = a.image[i];
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == kProducerWarp) {
    // ===================== producer: weight chunks and articulator parameters =====================
    if (lane == 0) {
      uint32_t c = 0, pa = 0;
      TR_DECL;
      auto push = [&](const float* src, uint32_t bytes) {   // one bulk copy per stage (split copies measured slower)
        if (ASB_DEC_ABLATE == 4) return;
        const uint32_t st = c % kStages;
        mbar_wait(&empty[st], ((c / kStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&full[st], bytes);
        bulk_g2s(Ws + st * kStageBytes, src, bytes, &full[st]);
        TR(3, 400 + st);
        ++c;
      };
      UnitIter units(a);
      int tile, g0, g1;
      while (units.next(tile, g0, g1)) {
        for (int j = 0; j < 2 * kW; ++j) push(a.image + kLinParams + 2 * j * kLinChunk, 2 * kLinChunk * 4);
        for (int k = 0; k < g1 - g0; ++k) {
          const int art = g0 + k;
          if (k == g1 - g0 - 1) {   // last articulator of this unit: pull the NEXT unit's input tile towards L2
            // (issued a whole unit ahead, the tile was evicted again by the output stream before its use)
            UnitIter peek = units;
            int t2, h0, h1;
            if (peek.next(t2, h0, h1) && t2 != tile) {
              const long long row0 = (long long)t2 * kTileM;
              const long long rows = min((long long)kTileM, a.M - row0);
              asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(a.rnn_out + row0 * 256),
                           "r"(uint32_t(rows * 256 * 4))
                           : "memory");
            }
          }
          const float* blk = a.image + lin_block(kSplit) + (size_t)art * art_block(kSplit);
          const uint32_t pb = pa & 1;
          mbar_wait(&pempty[pb], ((pa >> 1) & 1) ^ 1);
          mbar_arrive_expect_tx(&pfull[pb], kArtParams * 4);
          bulk_g2s(Ps + pb * kArtParams, blk, kArtParams * 4, &pfull[pb]);
          ++pa;
          const float* w1 = blk + kArtParams;
          const float* w2 = w1 + kW * 2 * kFcChunk;
          const float* wh = w2 + kW * 4 * kFcChunk;
          // split: the pieces alternate hi, lo in the image, so the same sequential walk streams both
          for (int j = 0; j < 2 * kW; ++j) push(w1 + j * kFcChunk, kFcChunk * 4);
          for (int j = 0; j < 4 * kW; ++j) push(w2 + j * kFcChunk, kFcChunk * 4);
          for (int j = 0; j < 2 * kW; ++j) push(wh + 2 * j * kHeadChunk, 2 * kHeadChunk * 4);   // 14 KB blocks stay 1 KB aligned
        }
      }
    }
    __syncwarp();
  } else if (warp == kMmaWarp) {
    // ===================== MMA issuer: the whole warp walks the loop, one elected lane issues =====
    {
      uint32_t c = 0, artc = 0, ph_chunk = 0, ph_a1 = 0;
      TR_DECL;
      // The whole warp walks the loop and one ELECTED lane issues (tools/micro/umma_rate.cu: behind
      // `lane == 0` every MMA is wrapped in an R2UR/ELECT loop and costs ~300 clk to issue).
      // One iteration = one 64-k WEIGHT BLOCK = two 32-k A chunks (one of each epilogue half, produced
      // concurrently and signalled on ONE barrier of 8 arrivals): 4 MMAs (12 split) + the commits per iteration.
      // With one iteration per 32-k chunk the issuing warp itself was the critical path of the kernel: ~450 clk
      // per iteration (waits, descriptor moves to uniform registers, UTCHMMA / UTCBAR issue) x 20 per articulator.
      // k_blocks blocks of 64 k (A columns at a_tmem + 64 b); a ring stage holds `blocks_per_stage` blocks.
      auto gemm = [&](uint32_t d_tmem, uint32_t a_tmem, uint32_t n, int k_blocks, int blocks_per_stage,
                      uint32_t block_bytes, uint64_t* abar, uint32_t& ph, uint64_t* done) {
        const uint32_t idesc = idesc_f16_m128(n);
        TR(0, 100 + n);
        for (int b = 0; b < k_blocks; ++b) {
          const int blk = b % blocks_per_stage;            // block inside the stage
          const bool opens = blk == 0;
          const bool closes = blk == blocks_per_stage - 1;
          const uint32_t st = c % kStages, st_lo = (c + 1) % kStages;   // split: the residual piece follows in the ring
          mbar_wait(&abar[b], (ph >> b) & 1);
          ph ^= 1u << b;
          TR(0, 30 + b);
          if (opens && ASB_DEC_ABLATE != 4) {
            mbar_wait(&full[st], (c / kStages) & 1);
            if (kSplit) mbar_wait(&full[st_lo], ((c + 1) / kStages) & 1);
          }
          fence_after_sync();
          TR(0, 10 + b);
          const uint64_t bdesc = smem_desc_k_sw128(smem_u32(Ws + st * kStageBytes + blk * block_bytes));
          const uint64_t bdesc_lo = smem_desc_k_sw128(smem_u32(Ws + st_lo * kStageBytes + blk * block_bytes));
          const uint32_t acol = a_tmem + b * 2 * kChunkK;
          if (elect_one()) {
#pragma unroll
            for (int ks = 0; ks < (ASB_DEC_ABLATE == 3 ? 0 : 4); ++ks) {
              // 16 k = 32 bytes into the 128-byte swizzled row: the descriptor's start address advances by 32 >> 4;
              // A: 32-column chunk ks / 2, fp16 pairs of its k-step ks % 2 at columns 8 (ks % 2)
              const uint32_t ac = acol + (ks >> 1) * kChunkK + (ks & 1) * 8;
              const uint64_t bd = bdesc + uint64_t(2 * ks), bd_lo = bdesc_lo + uint64_t(2 * ks);
              if (kSplit) {   // hi pairs in columns [0, 16) of the chunk, residual pairs in [16, 32); small terms first
                mma_bf16_ts(d_tmem, ac + 16, bd, idesc, (b | ks) != 0);   // a_lo w_hi
                mma_bf16_ts(d_tmem, ac, bd_lo, idesc, true);              // a_hi w_lo
                mma_bf16_ts(d_tmem, ac, bd, idesc, true);                 // a_hi w_hi
              } else {
                mma_bf16_ts(d_tmem, ac, bd, idesc, (b | ks) != 0);
              }
            }
            if (closes && ASB_DEC_ABLATE != 4) {   // frees the stage(s) once their MMAs have read them
              commit(&empty[st]);
              if (kSplit) commit(&empty[st_lo]);
            }
            if (b == k_blocks - 1) commit(done);
          }
          __syncwarp();
          TR(0, 50 + b);
          if (closes) c += kW;
        }
        TR(0, 99);
      };
      // the accumulator about to be overwritten still holds the previous heads output until the
      // output warps have loaded it
      auto wait_heads_read = [&]() {
        if (artc) {
          mbar_wait(d3_free, (artc - 1) & 1);
          fence_after_sync();
        }
      };
      UnitIter units(a);
      int tile, g0, g1;
      while (units.next(tile, g0, g1)) {
        {
          const uint32_t P = tmem_base + (artc & 1) * 256, Q = tmem_base + ((artc & 1) ^ 1) * 256;
          wait_heads_read();
          gemm(P, Q, 128, 4, 2, kLinChunk * 4, chunk, ph_chunk, &dbar[0]);     // linear.0          K = 256: 4 blocks of 16 KB
        }
        for (int k = 0; k < g1 - g0; ++k) {
          const uint32_t P = tmem_base + (artc & 1) * 256, Q = tmem_base + ((artc & 1) ^ 1) * 256;
          gemm(Q, P + 128, 256, 2, 1, 0, a1bar, ph_a1, &dbar[1]);              // Linear(128, 256): 2 blocks of 32 KB
          wait_heads_read();
          gemm(P, Q, 256, 4, 1, 0, chunk, ph_chunk, &dbar[2]);                 // Linear(256, 256): 4 blocks of 32 KB
          gemm(Q, P, kHeadN, 4, 2, kHeadChunk * 4, chunk, ph_chunk, &dbar[3]);  // x/y heads: 4 blocks of 14 KB
          ++artc;
        }
      }
    }
    __syncwarp();
  } else if (warp >= kHeadWarp0) {
    // ===================== output warps: thread = frame; sigmoid(rstd * D3 + ...) -> contours ======
    const int quarter = warp & 3;
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = uint32_t(quarter * 32) << 16;
    uint32_t pa = 0, artc = 0;
    uint32_t r[32], r2[32], r4[4];
    float4* otile = reinterpret_cast<float4*>(smem + kOffOut + (warp - kHeadWarp0) * kOutTile);
    TR_DECL;
    UnitIter units(a);
    int tile, g0, g1;
    while (units.next(tile, g0, g1)) {
      const long long grow0 = (long long)tile * kTileM + quarter * 32, grow = grow0 + lane;
      const bool live = grow < a.M;
      for (int k = 0; k < g1 - g0; ++k) {
        const int art = g0 + k;
        const uint32_t Q = tmem_base + lane_sel + ((artc & 1) ^ 1) * 256;
        const uint32_t pb = pa & 1;
        mbar_wait(&pfull[pb], (pa >> 1) & 1);
        mbar_wait(&dbar[3], artc & 1);
        fence_after_sync();
        TR(4 + (warp - kHeadWarp0), 600);
        const float2 m0 = Stat2[((artc & 1) * 2 + 0) * kTileM + row], m1 = Stat2[((artc & 1) * 2 + 1) * kTileM + row];
        const float mean = (m0.x + m1.x) * (1.f / 256.f);
        const float var = fmaxf((m0.y + m1.y) * (1.f / 256.f) - mean * mean, 0.f);
        const float rstd = rsqrtf(var + 1e-5f);
        const float nm = -mean * rstd;
        // the head parameters carry the factor -log2(e): t = -log2(e) (rstd D + b - rstd mean ws)
        const f32x2 rs2 = pack2(-1.4426950408889634f * rstd, -1.4426950408889634f * rstd), nm2 = pack2(nm, nm);
        float* dst = a.out + (grow * a.n_art + art) * 100;
        // plain shared-memory loads (not the volatile lds4): the compiler may batch them and interleave the 32
        // independent sigmoid chains of a column group (serialised, a group cost ~1.8 k clk of pure latency)
        const float4* prm4 = reinterpret_cast<const float4*>(Ps + pb * kArtParams);
        auto sig32 = [&](const uint32_t(&v)[32], int col0, float (&y)[32]) {
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 b4 = prm4[(P_BH + col0) / 4 + q];
            const float4 w4 = prm4[(P_WSH + col0) / 4 + q];
            unpack2(fma2(rs2, pack2(__uint_as_float(v[4 * q + 0]), __uint_as_float(v[4 * q + 1])),
                         fma2(nm2, pack2(w4.x, w4.y), pack2(b4.x, b4.y))), y[4 * q + 0], y[4 * q + 1]);
            unpack2(fma2(rs2, pack2(__uint_as_float(v[4 * q + 2]), __uint_as_float(v[4 * q + 3])),
                         fma2(nm2, pack2(w4.z, w4.w), pack2(b4.z, b4.w))), y[4 * q + 2], y[4 * q + 3]);
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y[i]) : "f"(y[i]));
#pragma unroll
          for (int i = 0; i < 32; ++i) asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y[i]) : "f"(1.f + y[i]));
        };
        // 32 columns of the thread's own row -> the warp's transposition tile (16-byte unit u of row r at
        // u ^ (r & 7): conflict-free both ways) -> COALESCED stores: one warp instruction writes 4 rows x 128
        // contiguous bytes (with every lane storing 16 bytes of its own row - 4000 bytes apart - an instruction
        // is 32 half-sector requests).
        auto emit32 = [&](const uint32_t(&v)[32], int chunk) {
          float y[32];
          sig32(v, 32 * chunk, y);
#if ASB_DEC_OUT_DIRECT
          if (live) {
#pragma unroll
            for (int q = 0; q < 8; ++q)
              __stcs(reinterpret_cast<float4*>(dst + 32 * chunk + 4 * q),
                     make_float4(y[4 * q], y[4 * q + 1], y[4 * q + 2], y[4 * q + 3]));
          }
#else
#pragma unroll
          for (int q = 0; q < 8; ++q)
            otile[lane * 8 + (q ^ (lane & 7))] = make_float4(y[4 * q], y[4 * q + 1], y[4 * q + 2], y[4 * q + 3]);
          __syncwarp();
          float4 val[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = 4 * i + (lane >> 3);
            val[i] = otile[rr * 8 + ((lane & 7) ^ (rr & 7))];
          }
          // rows 4 i + lane / 8 of this warp's 32: frame stride n_art * 100 floats
          float* o = a.out + ((grow0 + (lane >> 3)) * a.n_art + art) * 100 + 32 * chunk + 4 * (lane & 7);
          const long long rows_left = a.M - grow0 - (lane >> 3);
#pragma unroll
          for (int i = 0; i < 8; ++i)
            if (4 * i < rows_left) __stcs(reinterpret_cast<float4*>(o + (size_t)(4 * i) * a.n_art * 100), val[i]);
          __syncwarp();   // the tile is rewritten by the next chunk
#endif
        };
        // The accumulator columns are overwritten by FC2 of the NEXT articulator, which waits for the release
        // below: it comes after ONE column group (two groups stay in registers; with all 100 columns loaded first the
        // emission had no registers left to overlap its chains, with the release after two groups FC2 waited).
        tmem_ld32(Q, r);
        tmem_ld32(Q + 32, r2);
        tmem_ld_wait();
        emit32(r, 0);
        TR(4 + (warp - kHeadWarp0), 605);
        tmem_ld32(Q + 64, r);
        tmem_ld4(Q + 96, r4);
        tmem_ld_wait();
        fence_before_sync();
        __syncwarp();
        if (lane == 0) mbar_arrive(d3_free);       // the accumulator columns may be overwritten now
        TR(4 + (warp - kHeadWarp0), 603);
        emit32(r2, 1);
        TR(4 + (warp - kHeadWarp0), 606);
        emit32(r, 2);
        TR(4 + (warp - kHeadWarp0), 607);
        if (live) {   // the last 4 of the 100 columns
          const float4 b4 = prm4[(P_BH + 96) / 4], w4 = prm4[(P_WSH + 96) / 4];
          float t0, t1, t2, t3;
          unpack2(fma2(rs2, pack2(__uint_as_float(r4[0]), __uint_as_float(r4[1])),
                       fma2(nm2, pack2(w4.x, w4.y), pack2(b4.x, b4.y))), t0, t1);
          unpack2(fma2(rs2, pack2(__uint_as_float(r4[2]), __uint_as_float(r4[3])),
                       fma2(nm2, pack2(w4.z, w4.w), pack2(b4.z, b4.w))), t2, t3);
          __stcs(reinterpret_cast<float4*>(dst + 96),
                 make_float4(sigmoid_exp2(t0), sigmoid_exp2(t1), sigmoid_exp2(t2), sigmoid_exp2(t3)));
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&pempty[pb]);
        TR(4 + (warp - kHeadWarp0), 604);
        ++pa;
        ++artc;
      }
    }
  } else {
    // ===================== epilogue: thread = (frame = TMEM lane, alternate 32-column chunks) ======
    const int quarter = warp & 3, half = warp >> 2;
    const int row = quarter * 32 + lane;
    const uint32_t lane_sel = uint32_t(quarter * 32) << 16;
    float* xrow = Xs + row * kXStride;
    const uint32_t ps_addr = smem_u32(Ps), linb_addr = smem_u32(LinB);
    uint32_t pa = 0, artc = 0, unitc = 0;
    uint32_t r[32], r2[32];
    TR_DECL;
#ifdef ASB_DEC_TRACE
    const int tr_role = 1 + half;
#endif

    auto wait_d = [&](int which, uint32_t parity) {
      mbar_wait(&dbar[which], parity & 1);
      fence_after_sync();
      if (quarter == 0) TR(tr_role, 200);
    };
    // the A columns just stored are complete for this warp's 32 lanes: one arrival per warp
    auto signal = [&](uint64_t* bar) {
      tmem_st_wait();
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar);
      if (quarter == 0) TR(tr_role, 201);
    };
    // row moments are split over the two halves: each half publishes its partial sums after a layer;
    // the partner (or the output warp) reads them behind the chunk -> MMA -> D-complete chain
    float own_s = 0.f, own_ss = 0.f;
    auto publish = [&](float2* buf, float s, float ss) {
      own_s = s;
      own_ss = ss;
      buf[half * kTileM + row] = make_float2(s, ss);
    };
    auto moments = [&](const float2* buf, float inv_n, float& rstd, float& nm) {
      const float2 o = buf[(half ^ 1) * kTileM + row];
      const float mean = (own_s + o.x) * inv_n;
      const float var = fmaxf((own_ss + o.y) * inv_n - mean * mean, 0.f);
      rstd = rsqrtf(var + 1e-5f);
      nm = -mean * rstd;
    };
    // accumulator -> relu(rs * D + (b + nm * ws)) -> moments, fp16 pairs -> the chunk's own columns, chunk by chunk.
    // The TMEM load of the next chunk is in flight while this one is converted; the arithmetic is
    // packed fp32x2 (FFMA2) except the relu and the integer rounding.
    auto convert_in_place = [&](uint32_t region, uint32_t b_addr, uint32_t ws_addr, float rs, float nm,
                                bool deferred, float2* stat) {
      const f32x2 rs2 = pack2(rs, rs), nm2 = pack2(nm, nm), one2 = pack2(1.f, 1.f);
      f32x2 sa = pack2(0.f, 0.f), sb = sa, qa = sa, qb = sa;
      auto body = [&](uint32_t(&buf)[32], int ch, bool last) {
        uint32_t h[16], l[kSplit ? 16 : 1];
#pragma unroll
        for (int q = 0; q < (ASB_DEC_ABLATE == 1 || ASB_DEC_ABLATE == 2 ? 0 : 8); ++q) {
          const float4 b4 = lds4(b_addr + (32 * ch + 4 * q) * 4);
          f32x2 k01 = pack2(b4.x, b4.y), k23 = pack2(b4.z, b4.w);
          if (deferred) {
            const float4 w4 = lds4(ws_addr + (32 * ch + 4 * q) * 4);
            k01 = fma2(nm2, pack2(w4.x, w4.y), k01);
            k23 = fma2(nm2, pack2(w4.z, w4.w), k23);
          }
          float v0, v1, v2, v3;
          unpack2(fma2(rs2, pack2(__uint_as_float(buf[4 * q + 0]), __uint_as_float(buf[4 * q + 1])), k01), v0, v1);
          unpack2(fma2(rs2, pack2(__uint_as_float(buf[4 * q + 2]), __uint_as_float(buf[4 * q + 3])), k23), v2, v3);
          v0 = fmaxf(v0, 0.f);
          v1 = fmaxf(v1, 0.f);
          v2 = fmaxf(v2, 0.f);
          v3 = fmaxf(v3, 0.f);
          const f32x2 p01 = pack2(v0, v1), p23 = pack2(v2, v3);
          sa = fma2(p01, one2, sa);
          sb = fma2(p23, one2, sb);
          qa = fma2(p01, p01, qa);
          qb = fma2(p23, p23, qb);
          h[2 * q + 0] = pack_f16x2(v0, v1);
          h[2 * q + 1] = pack_f16x2(v2, v3);
          if constexpr (kSplit) {
            l[2 * q + 0] = residual_f16x2(v0, v1, h[2 * q + 0]);
            l[2 * q + 1] = residual_f16x2(v2, v3, h[2 * q + 1]);
          }
        }
        // 32 k as 16 columns of fp16 pairs, into the FIRST half of this chunk's own accumulator columns
        // (split: the residual pairs into the second half)
        if (ASB_DEC_ABLATE != 2) {
          tmem_st16(region + 32 * ch, h);
          if (kSplit) st_lo16(region + 32 * ch + 16, l);
        }
        if (last) {   // the moments are published BEFORE the last chunk is signalled (their readers sit behind it)
          float s0, s1, s2, s3, q0, q1, q2, q3;
          unpack2(sa, s0, s1);
          unpack2(sb, s2, s3);
          unpack2(qa, q0, q1);
          unpack2(qb, q2, q3);
          publish(stat, (s0 + s1) + (s2 + s3), (q0 + q1) + (q2 + q3));
        }
        signal(&chunk[ch >> 1]);
      };
      if (ASB_DEC_ABLATE != 2) tmem_ld32(region + 32 * half, r);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int ch = 2 * c + half;
        tmem_ld_wait();
        if (c & 1) {
          if (c < 3 && ASB_DEC_ABLATE != 2) tmem_ld32(region + 32 * (ch + 2), r);
          body(r2, ch, c == 3);
        } else {
          if (c < 3 && ASB_DEC_ABLATE != 2) tmem_ld32(region + 32 * (ch + 2), r2);
          body(r, ch, c == 3);
        }
      }
    };
    // x-hat (shared memory) -> A operand of FC1 in region[128, 256)
    auto write_a1 = [&](uint32_t region) {
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        const int ch = 2 * c + half;
        uint32_t h[16], l[kSplit ? 16 : 1];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float v0 = xrow[32 * ch + 2 * i], v1 = xrow[32 * ch + 2 * i + 1];
          h[i] = pack_f16x2(v0, v1);
          if constexpr (kSplit) l[i] = residual_f16x2(v0, v1, h[i]);
        }
        tmem_st16(region + 128 + 32 * ch, h);
        if (kSplit) st_lo16(region + 128 + 32 * ch + 16, l);
        signal(&a1bar[ch >> 1]);
      }
    };

    UnitIter units(a);
    int tile, g0, g1;
    while (units.next(tile, g0, g1)) {
      // ---- input tile -> A0 (fp16 pairs) in Q: this warp converts 32 rows x 128 of the 256 features
      {
        const uint32_t P = tmem_base + lane_sel + (artc & 1) * 256, Q = tmem_base + lane_sel + ((artc & 1) ^ 1) * 256;
        if (artc) wait_d(3, artc - 1);   // Q was the A operand of the previous heads GEMM
        // Coalesced: a warp instruction reads 4 rows x 128 B (one 16-byte piece per lane), the 32 x 32
        // block is turned to lane = row through a per-warp tile in the x-hat buffer, which is dead
        // between units.  (One row per lane costs an L1 tag look-up per lane and instruction: the
        // four chunks then took ~13 k clk per unit.)
        float* tile32 = Xs + warp * (32 * kTilePitch);
        const long long row0 = (long long)tile * kTileM + quarter * 32;
        auto load_block = [&](int ch, uint32_t(&buf)[32]) {   // 8 x float4 per lane, in flight
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const long long gr = row0 + 4 * i + (lane >> 3);
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gr < a.M) v = __ldg(reinterpret_cast<const float4*>(a.rnn_out + gr * 256 + 32 * ch) + (lane & 7));
            buf[4 * i + 0] = __float_as_uint(v.x);
            buf[4 * i + 1] = __float_as_uint(v.y);
            buf[4 * i + 2] = __float_as_uint(v.z);
            buf[4 * i + 3] = __float_as_uint(v.w);
          }
        };
        load_block(half, r2);
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          const int ch = 2 * c + half;
#pragma unroll
          for (int i = 0; i < 8; ++i)
            *reinterpret_cast<float4*>(tile32 + (4 * i + (lane >> 3)) * kTilePitch + 4 * (lane & 7)) =
                make_float4(__uint_as_float(r2[4 * i + 0]), __uint_as_float(r2[4 * i + 1]),
                            __uint_as_float(r2[4 * i + 2]), __uint_as_float(r2[4 * i + 3]));
          __syncwarp();
          if (c < 3) load_block(ch + 2, r2);   // the next block's loads fly under this block's conversion
          uint32_t h[16], l[kSplit ? 16 : 1];
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 v = *reinterpret_cast<const float4*>(tile32 + lane * kTilePitch + 4 * q);
            h[2 * q + 0] = pack_f16x2(v.x, v.y);
            h[2 * q + 1] = pack_f16x2(v.z, v.w);
            if constexpr (kSplit) {
              l[2 * q + 0] = residual_f16x2(v.x, v.y, h[2 * q + 0]);
              l[2 * q + 1] = residual_f16x2(v.z, v.w, h[2 * q + 1]);
            }
          }
          __syncwarp();   // the tile is rewritten by the next block
          tmem_st16(Q + 32 * ch, h);
          if (kSplit) st_lo16(Q + 32 * ch + 16, l);
          signal(&chunk[ch >> 1]);
        }
        // ---- linear.0 epilogue: relu(D0 + b) -> LayerNorm(128) -> x-hat (fp32) kept in shared memory
        wait_d(0, unitc);
        float s = 0.f, ss = 0.f;
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          const int ch = 2 * c + half;
          tmem_ld32(P + 32 * ch, r);
          tmem_ld_wait();
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const float4 b4 = lds4(linb_addr + (32 * ch + 4 * q) * 4);
            const float v0 = fmaxf(__uint_as_float(r[4 * q + 0]) + b4.x, 0.f);
            const float v1 = fmaxf(__uint_as_float(r[4 * q + 1]) + b4.y, 0.f);
            const float v2 = fmaxf(__uint_as_float(r[4 * q + 2]) + b4.z, 0.f);
            const float v3 = fmaxf(__uint_as_float(r[4 * q + 3]) + b4.w, 0.f);
            xrow[32 * ch + 4 * q + 0] = v0;
            xrow[32 * ch + 4 * q + 1] = v1;
            xrow[32 * ch + 4 * q + 2] = v2;
            xrow[32 * ch + 4 * q + 3] = v3;
            s += (v0 + v1) + (v2 + v3);
            ss = fmaf(v0, v0, fmaf(v1, v1, fmaf(v2, v2, fmaf(v3, v3, ss))));
          }
        }
        bar_sync_epilogue();   // the partner has read the previous unit's Stat0
        publish(Stat0, s, ss);
        bar_sync_epilogue();   // once per unit: the partner's partial sums are needed right away
        float rstd, nm;
        moments(Stat0, 1.f / 128.f, rstd, nm);
#pragma unroll 1
        for (int c = 0; c < 2; ++c) {
          const int ch = 2 * c + half;
#pragma unroll 8
          for (int k = 0; k < 32; ++k)
            xrow[32 * ch + k] = fmaf(xrow[32 * ch + k], rstd, nm);   // rounded to fp16 when it is packed (write_a1)
        }
        write_a1(P);
        ++unitc;
      }

      for (int k = 0; k < g1 - g0; ++k) {
        const uint32_t P = tmem_base + lane_sel + (artc & 1) * 256, Q = tmem_base + lane_sel + ((artc & 1) ^ 1) * 256;
        const uint32_t pb = pa & 1;
        const uint32_t prm = ps_addr + pb * kArtParams * 4;
        mbar_wait(&pfull[pb], (pa >> 1) & 1);
        float rstd, nm;
        wait_d(1, artc);                                      // FC1 done: D1 in Q
        convert_in_place(Q, prm + P_B1 * 4, 0u, 1.f, 0.f, false, Stat1);
        wait_d(2, artc);                                      // FC2 done: D2 in P
        moments(Stat1, 1.f / 256.f, rstd, nm);
        convert_in_place(P, prm + P_B2 * 4, prm + P_WS2 * 4, rstd, nm, true, Stat2 + (artc & 1) * 2 * kTileM);
        if (k + 1 < g1 - g0) write_a1(Q);                     // next articulator's FC1 operand, under the heads GEMM
        __syncwarp();
        if (lane == 0) mbar_arrive(&pempty[pb]);
        ++pa;
        ++artc;
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

size_t umma16_image_floats(int n_art, bool split) {
  return (size_t)lin_block(split) + (size_t)n_art * art_block(split);
}

// both precisions: fast = fp16 operands, fp32 = split fp16 operands (three MMAs per product)
bool umma16_decoder_supported(const asb_model* m) { return m->D == 50 && m->H == kH; }

static bool split_of(const asb_model* m) { return m->precision != ASB_PRECISION_FAST; }

static int pack_mat(const float* src, const float* scale, float* dst_floats, int n_src, int row0, int n_pad, int K,
                    int bpp, bool split, cudaStream_t st) {
  const int n = n_src * K;
  pack_umma16_kernel<<<(n + 255) / 256, 256, 0, st>>>(src, scale, reinterpret_cast<__half*>(dst_floats), n_src, row0,
                                                     n_pad, K, bpp, split ? 1 : 0);
  ASB_CHECK_LAUNCH("pack_umma16_kernel");
  return ASB_OK;
}
static int fold_wsum(const float* W, const float* scale, float* dst, int N, int K, bool split, cudaStream_t st,
                     double out_scale = 1.0) {
  fold_wsum16_kernel<<<(N + 127) / 128, 128, 0, st>>>(W, scale, dst, N, K, out_scale, split ? 1 : 0);
  ASB_CHECK_LAUNCH("fold_wsum16_kernel");
  return ASB_OK;
}
static int fold_bias(const float* W, const float* beta, const float* b, float* dst, int N, int K, cudaStream_t st,
                     double out_scale = 1.0) {
  fold_bias16_kernel<<<(N + 127) / 128, 128, 0, st>>>(W, beta, b, dst, N, K, out_scale);
  ASB_CHECK_LAUNCH("fold_bias16_kernel");
  return ASB_OK;
}

#define ASB_TRY(x)         \
  do {                     \
    int rc__ = (x);        \
    if (rc__) return rc__; \
  } while (0)

int pack_umma16_image(asb_model* m, const asb_weights* s, cudaStream_t st) {
  float* img = m->w.umma16_image;
  const bool split = split_of(m);
  const int W = split ? 2 : 1;
  const size_t total = umma16_image_floats(m->N, split);
  ASB_TRY(check_cuda(cudaMemsetAsync(img, 0, total * sizeof(float), st), "pack_umma16_image: memset"));
  ASB_TRY(fold_bias(nullptr, nullptr, s->linear_b, img, kH, 0, st));
  ASB_TRY(pack_mat(s->linear_w, nullptr, img + kLinParams, kH, 0, kH, 2 * kH, 2, split, st));
  const int D = m->D;
  for (int i = 0; i < m->N; ++i) {
    float* blk = img + lin_block(split) + (size_t)i * art_block(split);
    // LayerNorm affines folded forward: W' = W diag(gamma), b' = b + W beta   (models.py:11-21)
    ASB_TRY(fold_bias(s->fc1_w[i], s->ln0_b[i], s->fc1_b[i], blk + P_B1, kFc, kH, st));
    ASB_TRY(fold_bias(s->fc2_w[i], s->ln1_b[i], s->fc2_b[i], blk + P_B2, kFc, kFc, st));
    // the head parameters carry -log2(e): the output warps evaluate the sigmoid as 1 / (1 + 2^t)
    const double kNegLog2e = -1.4426950408889634;
    ASB_TRY(fold_bias(s->xh_w[i], s->ln2_b[i], s->xh_b[i], blk + P_BH, D, kFc, st, kNegLog2e));
    ASB_TRY(fold_bias(s->yh_w[i], s->ln2_b[i], s->yh_b[i], blk + P_BH + D, D, kFc, st, kNegLog2e));
    ASB_TRY(fold_wsum(s->fc2_w[i], s->ln1_w[i], blk + P_WS2, kFc, kFc, split, st));
    ASB_TRY(fold_wsum(s->xh_w[i], s->ln2_w[i], blk + P_WSH, D, kFc, split, st, kNegLog2e));
    ASB_TRY(fold_wsum(s->yh_w[i], s->ln2_w[i], blk + P_WSH + D, D, kFc, split, st, kNegLog2e));
    float* w1 = blk + kArtParams;
    float* w2 = w1 + W * 2 * kFcChunk;
    float* wh = w2 + W * 4 * kFcChunk;
    ASB_TRY(pack_mat(s->fc1_w[i], s->ln0_w[i], w1, kFc, 0, kFc, kH, 1, split, st));
    ASB_TRY(pack_mat(s->fc2_w[i], s->ln1_w[i], w2, kFc, 0, kFc, kFc, 1, split, st));
    ASB_TRY(pack_mat(s->xh_w[i], s->ln2_w[i], wh, D, 0, kHeadN, kFc, 2, split, st));
    ASB_TRY(pack_mat(s->yh_w[i], s->ln2_w[i], wh, D, D, kHeadN, kFc, 2, split, st));
  }
  return ASB_OK;
}

int launch_decoder_umma16(const asb_model* m, const float* rnn_out, float* out, int M, cudaStream_t st) {
  static DeviceOnce once;
  if (int rc = once.run([&] {
        if (int r = check_cuda(cudaFuncSetAttribute(decoder_umma16_kernel<false>,
                                                    cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes),
                               "decoder_umma16_kernel smem attribute"))
          return r;
        return check_cuda(cudaFuncSetAttribute(decoder_umma16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                               kSmemBytes),
                          "decoder_umma16_kernel (split) smem attribute");
      }))
    return rc;
  const int sms = sm_count();
  const int n_tiles = (M + kTileM - 1) / kTileM;
  UmmaArgs a;
  a.image = m->w.umma16_image;
  a.rnn_out = rnn_out;
  a.out = out;
  a.M = M;
  a.n_art = m->N;
  a.D = m->D;
  a.n_items = (long long)n_tiles * m->N;
  const int grid = int(a.n_items < sms ? a.n_items : sms);
  if (split_of(m))
    decoder_umma16_kernel<true><<<grid, kThreads, kSmemBytes, st>>>(a);
  else
    decoder_umma16_kernel<false><<<grid, kThreads, kSmemBytes, st>>>(a);
  ASB_CHECK_LAUNCH("decoder_umma16_kernel");
  return ASB_OK;
}

}  // namespace asb
