// Recurrent layer on tcgen05 tensor cores with W_hh resident in TENSOR MEMORY for all T steps.
//
// Same computation as gru.cu's fast kernel (nn.GRU cell of models.py:111 with packed-sequence
// semantics; G = 4: nn.LSTM): per step gh^T[G*H x batch] = W_hh[G*H x H] . h^T[H x batch], gates
// fused behind it.
//
// Blackwell mapping
//   * one CTA = 2*CPT (4, 8 or 16) batch rows of one direction - as few as still fill one wave of
//     SMs, because the gate math (6 MUFU per unit and row) is what bounds a step;
//   * W_hh (fp16; bf16 until round 2: 8 significand bits against 11, and the recurrent operands dominated the
//     fast-mode contour error) is the A operand and lives in tensor memory: lane = gate row, 64 columns per gate
//     (two consecutive-k fp16 per 32-bit column), written ONCE with tcgen05.st at kernel start -
//     G * 64 columns next to the G * 16 accumulator columns.  (The first version kept W_hh in shared
//     memory: every step then re-read 96 KB through the SS form, 32+ clk per MMA; the TS form
//     issues an N = 16 MMA every ~11 clk, tools/micro/umma_rate.cu.)
//   * per step ONE ELECTED lane of a converged warp issues G x 8 tcgen05.mma (kind::f16, fp16 in,
//     fp32 accumulate, M = 128 gate rows, N = 16 batch rows, K = 16): B = h as fp16 (shared
//     memory, 4 KB, K-major, 128-byte swizzle), D = G x 16 columns, lane = hidden unit;
//   * the 8 epilogue warps (thread = hidden unit, CPT batch columns) read D with tcgen05.ld - the
//     gate pre-activations of one (unit, row) sit in the same lane, so the gate math is
//     thread-local -, keep h (and the LSTM's c) in fp32 registers, write h as fp16 back into the B
//     operand and the fp32 output row, and hand over to the MMA warp through one mbarrier; the
//     input pre-activations of the next step are prefetched while the MMAs run.
//   * kSplit = true (fp32 precision, GRU): W_hh and h are carried as hi + lo fp16 pairs (hi = fp16(v),
//     lo = fp16(v - hi): 22 significand bits), a product is three MMAs (W_lo h_hi + W_hi h_lo + W_hi h_hi, fp32
//     accumulate), W_hh takes 2 x 192 columns of tensor memory, h two 4 KB operands, and the gates use expf /
//     tanhf.  72 MMAs per step instead of 24: ~0.9 us per step against 2.9 us of the FFMA kernel (gru.cu) with
//     the same 1e-5 agreement (the LSTM's 2 x 256 columns do not fit beside the accumulators).
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdlib>

#include "model.cuh"
#include "umma.cuh"

#ifndef ASB_RNN_FENCE_AT_MMA
#define ASB_RNN_FENCE_AT_MMA 0
#endif
#ifdef ASB_RNN_TRACE
// timeline of CTA (0, 0) (tools/rnn_trace.py): [role][event] = (tag, clock64); role 0 = MMA warp, 1 = epilogue warp 0
__device__ long long g_rnn_trace[2 * 4096];
#define TR(role, tag)                                                                   \
  do {                                                                                  \
    if (blockIdx.x == 0 && blockIdx.y == 0 && lane == 0 && tr_n < 2047) {               \
      g_rnn_trace[(role) * 4096 + 2 * tr_n] = (tag);                                    \
      g_rnn_trace[(role) * 4096 + 2 * tr_n + 1] = clock64();                            \
      ++tr_n;                                                                           \
    }                                                                                   \
  } while (0)
#define TR_DECL int tr_n = 0
extern "C" __attribute__((visibility("default"))) int asb_debug_rnn_trace(long long* host_out) {
  return int(cudaMemcpyFromSymbol(host_out, g_rnn_trace, sizeof(g_rnn_trace)));
}
#else
#define TR(role, tag) do { } while (0)
#define TR_DECL do { } while (0)
#endif

namespace {

using namespace asb;
using namespace asb::umma;

constexpr int kRows = 16;                    // the MMA's N (columns past the CTA's rows stay zero)
constexpr int kTileBytes = 128 * kH * 2;     // one gate tile of W_hh in fp16: 32 KB
constexpr int kAhead = 3, kGiSlots = 4;        // look-ahead of the input pre-activations (steps), ring slots
constexpr int kACol0 = 64;                   // TMEM: accumulators in columns [0, G*16), W_hh from column 64
constexpr int kHsBytes = 2 * kRows * 128;    // h as the B operand: 2 k-atoms x 16 rows x 128 B
constexpr int kEpiWarps = 8;
constexpr int kCtlWarp = 8;
constexpr int kThreads = 288;

__device__ __forceinline__ float sigmoid_f(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float tanh_f(float x) { return 2.f * sigmoid_f(2.f * x) - 1.f; }

// byte offset of element (row, k) inside a K-major, 128-byte-swizzled 16-bit operand whose k-atoms
// (64 elements) are `atom_bytes` apart
__host__ __device__ constexpr uint32_t sw128_bf16(uint32_t row, uint32_t k, uint32_t atom_bytes) {
  return (k >> 6) * atom_bytes + (row >> 3) * 1024u + (row & 7u) * 128u +
         ((((k & 63u) >> 3) ^ (row & 7u)) << 4) + (k & 7u) * 2u;
}

// whh (G*128, 128) fp32 -> fp16, same row-major layout (a row = the 64 TMEM columns of one lane);
// split: fp16 hi image followed by the fp16 residual image
__global__ void pack_whh_umma_kernel(const float* __restrict__ whh, uint16_t* __restrict__ img, int G, int split) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= G * kH * kH) return;
  if (!split) {
    img[i] = __half_as_ushort(__float2half_rn(whh[i]));
  } else {
    const __half hi = __float2half_rn(whh[i]);
    img[i] = __half_as_ushort(hi);
    img[G * kH * kH + i] = __half_as_ushort(__float2half_rn(whh[i] - __half2float(hi)));
  }
}

struct RnnArgs {
  const uint16_t* img[2];        // per direction: fp16 (fast) or fp16 hi | fp16 lo (split)
  const float* bhhn[2];          // GRU: b_hh of the n gate
  const float* gi;               // (B*T, 2, G*H) input pre-activations, or (layer 0) null:
  const long long* tokens;       //   (B, T_in) token ids and
  const float* tab;              //   tab0[dir][token][G*H] = embedding . W_ih_l0^T + folded biases (model.cuh)
  int T_in, V;
  int* tok_flag;                 // {1, id} is stored here when a token id is outside [0, V) (model.cuh)
  const int* lengths;            // (B) sorted descending
  float* out;                    // (B, T, 2H); every element is written (zeros at t >= length)
  int B, T;
};

// CPT = batch columns per epilogue thread; the CTA owns 2 * CPT batch rows
template <int G, int CPT, bool kSplit>
__global__ void __launch_bounds__(kThreads, 1) rnn_umma_kernel(const RnnArgs a) {
  static_assert(!kSplit || G == 3, "split operands: 2 x G x 64 columns of W_hh must fit beside the accumulators");
  constexpr int kW = kSplit ? 2 : 1;
  constexpr int kHs = kW * kHsBytes;                 // split: the hi operand followed by the residual operand
  extern __shared__ unsigned char smem_raw[];
  unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // keeps the shared address space
  unsigned char* Hs = smem;                          // 4 KB (8 KB split)
  uint64_t* bars = reinterpret_cast<uint64_t*>(Hs + kHs);
  float* Gs = reinterpret_cast<float*>(Hs + kHs + 64);   // [kGiSlots][CPT][G][256 threads] input pre-activations
  int* Toks = reinterpret_cast<int*>(Gs + kGiSlots * CPT * G * 256);   // [2 * CPT rows][T] token ids (layer 0)
  uint64_t* h_ready = bars + 1;  // h (fp16) of this step is in shared memory, D has been read
  uint64_t* d_ready = bars + 2;  // this step's MMAs completed
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int dir = blockIdx.y;
  const int b0 = blockIdx.x * 2 * CPT;
  const int Lmax = a.lengths[b0];   // sorted descending: the tile's first row is its longest
  constexpr int GH = G * kH;
  constexpr uint32_t kTmemCols = (kACol0 + kW * G * 64 <= 256) ? 256 : 512;
  static_assert(kACol0 + kW * G * 64 <= 512, "tensor memory");

  if (tid == 0) {
    mbar_init(h_ready, kEpiWarps);
    mbar_init(d_ready, 1);
    fence_mbar_init();
  }
  if (warp == kCtlWarp) tmem_alloc(tmem_slot, kTmemCols);
  for (int i = tid; i < kHs / 4; i += kThreads) reinterpret_cast<uint32_t*>(Hs)[i] = 0u;   // h0 = 0
  if (a.tokens) {   // layer 0: the embedding + input projection is a table look-up done by this kernel
    for (int i = tid; i < 2 * CPT * a.T; i += kThreads) {
      const int row = i / a.T, t = i - row * a.T;
      long long tok = (b0 + row < a.B) ? a.tokens[(long long)(b0 + row) * a.T_in + t] : 0;
      if (tok < 0 || tok >= a.V) {   // nn.Embedding raises here; report it, then clamp to stay in bounds
        a.tok_flag[1] = int(tok);
        a.tok_flag[0] = 1;
        tok = tok < 0 ? 0 : a.V - 1;
      }
      Toks[i] = int(tok);
    }
  }
  fence_proxy_async();
  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  const uint32_t tmem_d = *tmem_slot;

  if (warp == kCtlWarp) {
    // the whole warp walks the loop, one ELECTED lane issues (behind `lane == 0` the compiler wraps
    // every tcgen05.mma in an R2UR/ELECT loop that costs far more than the MMA itself)
    const uint32_t idesc = idesc_f16_m128(kRows);
    const uint64_t bdesc0 = smem_desc_k_sw128(smem_u32(Hs));
    const uint32_t tmem_a = tmem_d + kACol0;
    TR_DECL;
    for (int s = 0; s < Lmax; ++s) {
      mbar_wait(h_ready, s & 1);
#if ASB_RNN_FENCE_AT_MMA
      fence_proxy_async();     // the generic-proxy writes of h (ordered before the barrier) -> async proxy
#endif
      fence_after_sync();
      TR(0, 1);
      if (elect_one()) {
#pragma unroll
        for (int g = 0; g < G; ++g)
#pragma unroll
          for (int kt = 0; kt < 8; ++kt) {
            // k-step kt: 8 TMEM columns of A; B: atom kt / 4, 32 bytes (2 descriptor units) per step
            const uint64_t bd = bdesc0 + uint64_t(((kt >> 2) * kRows * 128) >> 4) + uint64_t(2 * (kt & 3));
            if (kSplit) {   // small terms first; W_lo from column (G + g) * 64, h_lo kHsBytes behind h_hi
              mma_bf16_ts(tmem_d + g * kRows, tmem_a + (G + g) * 64 + kt * 8, bd, idesc, kt != 0);
              mma_bf16_ts(tmem_d + g * kRows, tmem_a + g * 64 + kt * 8, bd + uint64_t(kHsBytes >> 4), idesc, true);
              mma_bf16_ts(tmem_d + g * kRows, tmem_a + g * 64 + kt * 8, bd, idesc, true);
            } else {
              mma_bf16_ts(tmem_d + g * kRows, tmem_a + g * 64 + kt * 8, bd, idesc, kt != 0);
            }
          }
        commit(d_ready);
      }
      __syncwarp();
      TR(0, 2);
    }
  } else {
    // thread = (hidden unit u = TMEM lane, CPT of the CTA's 2 * CPT batch rows)
    const int quarter = warp & 3, colhalf = warp >> 2;
    const int u = quarter * 32 + lane;
    const uint32_t taddr = tmem_d + (uint32_t(quarter * 32) << 16) + colhalf * CPT;
    const float bn = (G == 3) ? a.bhhn[dir][u] : 0.f;
    {   // W_hh -> tensor memory: this thread's gate rows, its half of the k range
      const uint32_t ta = tmem_d + (uint32_t(quarter * 32) << 16) + kACol0 + colhalf * 32;
      uint32_t w[32];
#pragma unroll 1
      for (int g = 0; g < kW * G; ++g) {   // split: the G residual tiles follow the G hi tiles, in the image and in TMEM
        const uint4* src = reinterpret_cast<const uint4*>(a.img[dir] + (size_t)(g * kH + u) * kH + colhalf * 64);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint4 v = __ldg(src + q);
          w[4 * q + 0] = v.x;
          w[4 * q + 1] = v.y;
          w[4 * q + 2] = v.z;
          w[4 * q + 3] = v.w;
        }
        tmem_st32(ta + g * 64, w);
      }
      tmem_st_wait();
      fence_before_sync();
    }
    int len[CPT];
    const float* gi_row[CPT];   // per-row bases: one multiply-add per step instead of a 64-bit chain
    float* out_row[CPT];
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      const int b = b0 + colhalf * CPT + j;
      len[j] = b < a.B ? a.lengths[b] : 0;
      gi_row[j] = a.gi + ((long long)b * a.T * 2 + dir) * GH + u;
      out_row[j] = a.out + (long long)b * a.T * (2 * kH) + dir * kH + u;
    }
    float h[CPT], cs[CPT];   // cs: LSTM cell state (unused by the GRU)
#pragma unroll
    for (int j = 0; j < CPT; ++j) h[j] = cs[j] = 0.f;

    // Input pre-activations are fetched kAhead steps before they are used, straight into shared
    // memory with cp.async: gi (78 MB per layer at batch 256 x 100) streams from L2 / HBM and one
    // step (~0.65 us) does not cover that latency - with one step of look-ahead in registers 31 % of
    // the epilogue's issue slots waited on these loads, and deeper look-ahead in registers does not
    // help because a scoreboard wait covers the youngest loads too; cp.async.wait_group counts groups.
    // Each thread reads back only what it copied itself.  Rows past their length are not fetched:
    // their gate values are computed from stale data and discarded by the `valid` selects.
    auto fetch_gi = [&](int step) {
      const int t = dir ? Lmax - 1 - step : step;
      float* dst = Gs + (size_t)(step & (kGiSlots - 1)) * (CPT * G * 256) + (tid & 255);
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        if (step < Lmax && t < len[j]) {
          const float* p = a.tokens
                               ? a.tab + ((long long)dir * a.V + Toks[(colhalf * CPT + j) * a.T + t]) * GH + u
                               : gi_row[j] + (long long)t * (2 * GH);
#pragma unroll
          for (int g = 0; g < G; ++g)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(dst + (j * G + g) * 256)),
                         "l"(p + g * kH)
                         : "memory");
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll 1
    for (int k = 0; k < kAhead; ++k) fetch_gi(k);
    // h0 (zeros) is in place: first hand-over to the MMA thread
    __syncwarp();
    if (lane == 0) mbar_arrive(h_ready);

    TR_DECL;
    for (int s = 0; s < Lmax; ++s) {
      const int t = dir ? Lmax - 1 - s : s;
      if (warp == 0) TR(1, 15);
      fetch_gi(s + kAhead);   // into the slot that was read a step ago
      asm volatile("cp.async.wait_group %0;" ::"n"(kAhead) : "memory");   // this step's group has landed
      float gc[CPT][G];
      {
        const float* src = Gs + (size_t)(s & (kGiSlots - 1)) * (CPT * G * 256) + (tid & 255);
#pragma unroll
        for (int j = 0; j < CPT; ++j)
#pragma unroll
          for (int g = 0; g < G; ++g) gc[j][g] = src[(j * G + g) * 256];
      }
      if (warp == 0) TR(1, 10);

      mbar_wait(d_ready, s & 1);
      fence_after_sync();
      if (warp == 0) TR(1, 11);
      uint32_t d[G][CPT];
#pragma unroll
      for (int g = 0; g < G; ++g) tmem_ldn(taddr + g * kRows, d[g]);
      tmem_ld_wait();
      if (warp == 0) TR(1, 12);

#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        const bool valid = t < len[j];
        float hn;
        if (G == 3 && kSplit) {   // fp32 precision: the gate functions of gru_fp32_kernel
          const float r = 1.f / (1.f + expf(-(gc[j][0] + __uint_as_float(d[0][j]))));
          const float z = 1.f / (1.f + expf(-(gc[j][1] + __uint_as_float(d[1][j]))));
          const float nn = tanhf(gc[j][2] + r * (__uint_as_float(d[2][j]) + bn));
          hn = nn + z * (h[j] - nn);
        } else if (G == 3) {   // GRU, gate rows r, z, n
          const float r = sigmoid_f(gc[j][0] + __uint_as_float(d[0][j]));
          const float z = sigmoid_f(gc[j][1] + __uint_as_float(d[1][j]));
          const float nn = tanh_f(gc[j][2] + r * (__uint_as_float(d[2][j]) + bn));
          hn = nn + z * (h[j] - nn);
        } else {        // LSTM, gate rows i, f, g, o (torch.nn.LSTM order): c' = f c + i g, h' = o tanh(c')
          const float ig = sigmoid_f(gc[j][0] + __uint_as_float(d[0][j]));
          const float fg = sigmoid_f(gc[j][1] + __uint_as_float(d[1][j]));
          const float gg = tanh_f(gc[j][2] + __uint_as_float(d[2][j]));
          const float og = sigmoid_f(gc[j][G - 1] + __uint_as_float(d[G - 1][j]));
          const float cn = fmaf(fg, cs[j], ig * gg);
          cs[j] = valid ? cn : cs[j];
          hn = og * tanh_f(cn);
        }
        h[j] = valid ? hn : h[j];
        const int n = colhalf * CPT + j;
        if (kSplit) {
          const __half hi = __float2half_rn(h[j]);
          *reinterpret_cast<__half*>(Hs + sw128_bf16(n, u, kRows * 128)) = hi;
          *reinterpret_cast<__half*>(Hs + kHsBytes + sw128_bf16(n, u, kRows * 128)) = __float2half_rn(h[j] - __half2float(hi));
        } else {
          *reinterpret_cast<__half*>(Hs + sw128_bf16(n, u, kRows * 128)) = __float2half_rn(h[j]);
        }
      }
      if (warp == 0) TR(1, 13);
#if !ASB_RNN_FENCE_AT_MMA
      fence_proxy_async();     // the generic-proxy writes of h must be visible to the tensor core
#endif
      fence_before_sync();
      __syncwarp();
      if (lane == 0) mbar_arrive(h_ready);
      if (warp == 0) TR(1, 14);
      // the fp32 output rows go out AFTER the hand-over: stores in flight ahead of the proxy fence
      // would be waited for on the step's critical path
#pragma unroll
      for (int j = 0; j < CPT; ++j) {
        if (t < len[j]) out_row[j][(long long)t * (2 * kH)] = h[j];
      }
    }
    // padded steps of the output are zero (pad_packed_sequence, models.py:138): written here, after
    // the last step, so that the caller does not have to clear the whole (B, T, 2H) buffer first
#pragma unroll
    for (int j = 0; j < CPT; ++j) {
      if (b0 + colhalf * CPT + j < a.B)
        for (int t = len[j]; t < a.T; ++t) out_row[j][(long long)t * (2 * kH)] = 0.f;
    }
  }

  fence_before_sync();
  __syncthreads();
  fence_after_sync();
  if (warp == kCtlWarp) tmem_dealloc(tmem_d, kTmemCols);
}

}  // namespace

namespace asb {

static bool split_of(const asb_model* m) { return m->precision != ASB_PRECISION_FAST; }

size_t whh_umma_bytes(const asb_model* m) { return (size_t)m->G * kTileBytes * (split_of(m) ? 2 : 1); }

// fast precision: GRU and LSTM (fp16 operands); fp32 precision: the GRU with split fp16 operands
bool rnn_umma_supported(const asb_model* m) {
  if (m->H != kH) return false;
  return split_of(m) ? m->G == 3 : (m->G == 3 || m->G == 4);
}

int pack_whh_umma(asb_model* m, const asb_weights* src, cudaStream_t st) {
  for (int layer = 0; layer < 2; ++layer)
    for (int dir = 0; dir < 2; ++dir) {
      const int n = m->G * kH * kH;
      pack_whh_umma_kernel<<<(n + 255) / 256, 256, 0, st>>>(
          src->rnn_w_hh[layer][dir], reinterpret_cast<uint16_t*>(m->w.whh_umma[layer][dir]), m->G, split_of(m) ? 1 : 0);
      ASB_CHECK_LAUNCH("pack_whh_umma_kernel");
    }
  return ASB_OK;
}

constexpr int kMaxTokenBytes = 64 * 1024;   // token ids staged per CTA (16 rows x 1024 steps)

// Batch rows per warp column: the fewest rows per CTA that still fit one wave of the machine (the
// gate math bounds a step).  -DASB_RNN_CPT=2|4|8 pins it for experiments (tools/rnn_variants.py).
static int rows_per_column(int B) {
#ifdef ASB_RNN_CPT
  return ASB_RNN_CPT;
#else
  const int sms = sm_count();
  return (2 * ((B + 3) / 4) <= sms) ? 2 : (2 * ((B + 7) / 8) <= sms) ? 4 : 8;
#endif
}

template <int G, bool kSplit>
static int launch_rnn_umma_g(const asb_model* m, RnnArgs& a, int B, bool use_tokens, cudaStream_t st) {
  const int cpt = rows_per_column(B);
  const int tok_bytes = 2 * cpt * a.T * 4;
  if (!use_tokens || tok_bytes > kMaxTokenBytes) a.tokens = nullptr;   // the caller materialised gi instead
  constexpr int kHs = (kSplit ? 2 : 1) * kHsBytes;
  const int smem = kHs + 64 + 1024 + kGiSlots * cpt * G * 1024 + (a.tokens ? tok_bytes : 0);
  static DeviceOnce once;   // one instance per (G, kSplit) (function template)
  if (int rc = once.run([&] {
        const int max_smem = kHs + 64 + 1024 + kGiSlots * 8 * G * 1024 + kMaxTokenBytes;
        cudaError_t e =
            cudaFuncSetAttribute(rnn_umma_kernel<G, 2, kSplit>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(rnn_umma_kernel<G, 4, kSplit>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess)
          e = cudaFuncSetAttribute(rnn_umma_kernel<G, 8, kSplit>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        return check_cuda(e, "rnn_umma_kernel smem attribute");
      }))
    return rc;
  if (cpt == 2) {
    rnn_umma_kernel<G, 2, kSplit><<<dim3((B + 3) / 4, 2), kThreads, smem, st>>>(a);
  } else if (cpt == 4) {
    rnn_umma_kernel<G, 4, kSplit><<<dim3((B + 7) / 8, 2), kThreads, smem, st>>>(a);
  } else {
    rnn_umma_kernel<G, 8, kSplit><<<dim3((B + 15) / 16, 2), kThreads, smem, st>>>(a);
  }
  ASB_CHECK_LAUNCH("rnn_umma_kernel");
  return ASB_OK;
}

// layer 0 may read its input pre-activations through the token table instead of a materialised gi
bool rnn_umma_reads_tokens(const asb_model* m, int B, int T) {
  const int cpt = rows_per_column(B);
  return rnn_umma_supported(m) && 2 * cpt * T * 4 <= kMaxTokenBytes;
}

int launch_rnn_umma(const asb_model* m, int layer, const float* gi, const long long* tokens, int T_in,
                    const int* lengths, float* out, int B, int T, cudaStream_t st) {
  RnnArgs a;
  for (int d = 0; d < 2; ++d) {
    a.img[d] = reinterpret_cast<const uint16_t*>(m->w.whh_umma[layer][d]);
    a.bhhn[d] = m->w.bhh_n[layer][d];
  }
  a.gi = gi; a.lengths = lengths; a.out = out; a.B = B; a.T = T;
  a.tokens = tokens; a.tab = m->w.tab0; a.T_in = T_in; a.V = m->V; a.tok_flag = m->tok_flag_dev;
  const bool use_tokens = tokens != nullptr && layer == 0;
  if (split_of(m)) return launch_rnn_umma_g<3, true>(m, a, B, use_tokens, st);
  return m->G == 3 ? launch_rnn_umma_g<3, false>(m, a, B, use_tokens, st)
                   : launch_rnn_umma_g<4, false>(m, a, B, use_tokens, st);
}

}  // namespace asb
