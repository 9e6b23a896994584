// P2CP (point-to-closest-point) distance and mean Euclidean distance between contours.
//
// Replaces MeanP2CPDistance.forward (phoneme_to_articulation/metrics.py:32-46) and the
// p2cp_distance / euclidean_distance wrappers (metrics.py:38-68) of the reference, which
// materialise a (…, n, m) cdist tensor and run on the CPU in the evaluation loop.
//
// Fast path (n = m = 50, planar (2,50) contours = the model's output layout):
//   persistent CTAs, one per SM; tiles of 128 contour pairs (2 x 51 200 B) are staged in
//   shared memory by 1-D bulk async copies (TMA, UBLKCP) into a 2-stage mbarrier ring; two
//   threads share a contour pair: each keeps 26 rows of u in registers as natural fp32x2
//   pairs and streams the columns of v two at a time, so that one distance evaluation feeds
//   both the row minimum and the column minimum.  Squared distances use the direct
//   (dx*dx + dy*dy) form on packed FADD2/FMUL2/FFMA2, minima use FMNMX3, sqrt is taken only
//   of the 100 minima.  The kernel is fp32-issue bound, not HBM bound (kernel K1).
#include "asb_common.cuh"

namespace {

using namespace asb;

constexpr int kD = 50;               // points per contour on the fast path
constexpr int kItemF = 2 * kD;       // floats per contour
constexpr int kTile = 128;           // contour pairs per tile
constexpr int kThreads = 2 * kTile;  // two threads per contour pair
constexpr int kRowPairs = 13;        // 26 rows per thread: half 0 rows 0..25, half 1 rows 24..49
constexpr int kStages = 2;
constexpr size_t kSmemBytes = size_t(kStages) * 2 * kTile * kItemF * sizeof(float) + kStages * sizeof(uint64_t);

__device__ __forceinline__ float sqrt_approx(float x) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

__global__ void __launch_bounds__(kThreads, 1)
p2cp50_kernel(const float* __restrict__ u, const float* __restrict__ v, float* __restrict__ out, long long items) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  float* su = reinterpret_cast<float*>(smem_raw);
  float* sv = su + kStages * kTile * kItemF;
  uint64_t* full = reinterpret_cast<uint64_t*>(sv + kStages * kTile * kItemF);

  const int tid = threadIdx.x;
  const long long n_tiles = (items + kTile - 1) / kTile;

  if (tid == 0) {
    mbar_init(&full[0], 1);
    mbar_init(&full[1], 1);
    fence_mbar_init();
  }
  __syncthreads();

  auto issue = [&](long long tile, int stage) {
    const long long first = tile * kTile;
    const long long left = items - first;
    const uint32_t cnt = left < kTile ? uint32_t(left) : uint32_t(kTile);
    const uint32_t bytes = cnt * kItemF * sizeof(float);
    mbar_arrive_expect_tx(&full[stage], 2 * bytes);
    bulk_g2s(su + stage * kTile * kItemF, u + first * kItemF, bytes, &full[stage]);
    bulk_g2s(sv + stage * kTile * kItemF, v + first * kItemF, bytes, &full[stage]);
  };

  long long tile = blockIdx.x;
  if (tid == 0 && tile < n_tiles) issue(tile, 0);

  const int local = tid >> 1;
  const int half = tid & 1;
  const int r0 = half * 24;
  const float kInf = __int_as_float(0x7f800000);

  for (int k = 0; tile < n_tiles; ++k, tile += gridDim.x) {
    const int stage = k & 1;
    const long long next = tile + gridDim.x;
    if (tid == 0 && next < n_tiles) issue(next, stage ^ 1);
    mbar_wait(&full[stage], (k >> 1) & 1);

    const float* pu = su + stage * kTile * kItemF + local * kItemF;
    const float* pv = sv + stage * kTile * kItemF + local * kItemF;

    f32x2 ux[kRowPairs], uy[kRowPairs];
    float rlo[kRowPairs], rhi[kRowPairs];
#pragma unroll
    for (int p = 0; p < kRowPairs; ++p) {
      ux[p] = *reinterpret_cast<const f32x2*>(pu + r0 + 2 * p);
      uy[p] = *reinterpret_cast<const f32x2*>(pu + kD + r0 + 2 * p);
      rlo[p] = kInf;
      rhi[p] = kInf;
    }

    // Software pipeline over column pairs: the operands of pair j+2 are fetched and the
    // partner exchange + sqrt of pair j-2 are retired while pair j is being evaluated, so the
    // LDS / SHFL / MUFU latencies overlap the FMA-pipe-bound body.
    float colsum = 0.f;
    float2 vx = *reinterpret_cast<const float2*>(pv);
    float2 vy = *reinterpret_cast<const float2*>(pv + kD);
    float prev_send = 0.f, prev_mine = 0.f;  // sqrt(min(0, 0)) = 0 retires harmlessly at j = 0
#pragma unroll 1
    for (int j = 0; j < kD; j += 2) {
      const f32x2 vxa = pack2(vx.x, vx.x), vxb = pack2(vx.y, vx.y);
      const f32x2 vya = pack2(vy.x, vy.x), vyb = pack2(vy.y, vy.y);
      const int jn = (j + 2 < kD) ? j + 2 : j;
      vx = *reinterpret_cast<const float2*>(pv + jn);
      vy = *reinterpret_cast<const float2*>(pv + kD + jn);
      const float recv = __shfl_xor_sync(0xffffffffu, prev_send, 1);
      float ca = kInf, cb = kInf;
#pragma unroll
      for (int p = 0; p < kRowPairs; ++p) {
        const f32x2 dxa = sub2(ux[p], vxa), dya = sub2(uy[p], vya);
        const f32x2 dxb = sub2(ux[p], vxb), dyb = sub2(uy[p], vyb);
        const f32x2 qa = fma2(dya, dya, mul2(dxa, dxa));
        const f32x2 qb = fma2(dyb, dyb, mul2(dxb, dxb));
        float a0, a1, b0, b1;
        unpack2(qa, a0, a1);
        unpack2(qb, b0, b1);
        rlo[p] = min3(rlo[p], a0, b0);
        rhi[p] = min3(rhi[p], a1, b1);
        ca = min3(ca, a0, a1);
        cb = min3(cb, b0, b1);
      }
      colsum += sqrt_approx(fminf(prev_mine, recv));
      // the partner thread holds the other 26 rows: exchange one column each way
      prev_send = half ? ca : cb;
      prev_mine = half ? cb : ca;
    }
    colsum += sqrt_approx(fminf(prev_mine, __shfl_xor_sync(0xffffffffu, prev_send, 1)));

    float rowsum = 0.f;
#pragma unroll
    for (int p = 0; p < kRowPairs; ++p) {
      // half 1 starts at row 24: its first pair (rows 24, 25) belongs to half 0
      const float s = sqrt_approx(rlo[p]) + sqrt_approx(rhi[p]);
      rowsum += (half && p == 0) ? 0.f : s;
    }
    rowsum += __shfl_xor_sync(0xffffffffu, rowsum, 1);
    colsum += __shfl_xor_sync(0xffffffffu, colsum, 1);
    const long long item = tile * kTile + local;
    if (half == 0 && item < items) out[item] = (rowsum / float(kD) + colsum / float(kD)) * 0.5f;

    __syncthreads();  // every thread is done with this stage before it is refilled
  }
}

// General contours (any n, m; planar or interleaved): one warp per contour pair, operands
// read through L1.  Used for ragged point counts and for MeanP2CPDistance on (*, n, 2) input.
__global__ void __launch_bounds__(256)
p2cp_generic_kernel(const float* __restrict__ u, const float* __restrict__ v, float* __restrict__ out,
                    long long items, int n, int m, int u_sp, int u_sc, int v_sp, int v_sc) {
  const long long item = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= items) return;
  const int lane = threadIdx.x & 31;
  const float* pu = u + item * 2 * n;
  const float* pv = v + item * 2 * m;
  const float kInf = __int_as_float(0x7f800000);
  float acc_u = 0.f, acc_v = 0.f;
  for (int i = lane; i < n; i += 32) {
    const float x = pu[i * u_sp], y = pu[i * u_sp + u_sc];
    float best = kInf;
    for (int j = 0; j < m; ++j) {
      const float dx = x - pv[j * v_sp], dy = y - pv[j * v_sp + v_sc];
      best = fminf(best, fmaf(dy, dy, dx * dx));
    }
    acc_u += sqrt_approx(best);
  }
  for (int j = lane; j < m; j += 32) {
    const float x = pv[j * v_sp], y = pv[j * v_sp + v_sc];
    float best = kInf;
    for (int i = 0; i < n; ++i) {
      const float dx = pu[i * u_sp] - x, dy = pu[i * u_sp + u_sc] - y;
      best = fminf(best, fmaf(dy, dy, dx * dx));
    }
    acc_v += sqrt_approx(best);
  }
  acc_u = warp_sum(acc_u);
  acc_v = warp_sum(acc_v);
  if (lane == 0) out[item] = (acc_u / float(n) + acc_v / float(m)) * 0.5f;
}

// mean over points of |u_k - v_k| for planar contours: one warp per contour pair.
__global__ void __launch_bounds__(256)
mean_euclidean_kernel(const float* __restrict__ u, const float* __restrict__ v, float* __restrict__ out,
                      long long items, int n) {
  const long long item = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= items) return;
  const int lane = threadIdx.x & 31;
  const float* pu = u + item * 2 * n;
  const float* pv = v + item * 2 * n;
  float acc = 0.f;
  for (int k = lane; k < n; k += 32) {
    const float dx = pu[k] - pv[k], dy = pu[n + k] - pv[n + k];
    acc += sqrtf(fmaf(dy, dy, dx * dx));
  }
  acc = warp_sum(acc);
  if (lane == 0) out[item] = acc / float(n);
}

// Closest pair between two planar contours u (2, n) and v (2, m): min over all (i, j) of the
// Euclidean distance, plus the pair that attains it (smallest flattened index i*m + j among
// ties).  This is the reduction of CriticalLoss (phoneme_to_articulation/principal_components/
// losses.py:86-96: cdist -> reshape(D*D) -> min) and _calculate_TV without slices; one warp per
// item, lane owns columns j = lane, lane + 32, ...; direct dx*dx + dy*dy form, sqrt of the minimum only.
__global__ void __launch_bounds__(256)
min_distance_kernel(const float* __restrict__ u, const float* __restrict__ v, float* __restrict__ out_d,
                    int* __restrict__ out_ij, long long items, int n, int m) {
  const long long item = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (item >= items) return;
  const int lane = threadIdx.x & 31;
  const float* pu = u + item * 2 * n;
  const float* pv = v + item * 2 * m;
  const float kInf = __int_as_float(0x7f800000);
  float best = kInf;
  int flat = 0x7fffffff;
  for (int j = lane; j < m; j += 32) {
    const float x = pv[j], y = pv[m + j];
    for (int i = 0; i < n; ++i) {
      const float dx = pu[i] - x, dy = pu[n + i] - y;
      const float d = fmaf(dy, dy, dx * dx);
      const int f = i * m + j;
      if (d < best || (d == best && f < flat)) {
        best = d;
        flat = f;
      }
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float od = __shfl_xor_sync(0xffffffffu, best, o);
    const int of = __shfl_xor_sync(0xffffffffu, flat, o);
    if (od < best || (od == best && of < flat)) {
      best = od;
      flat = of;
    }
  }
  if (lane == 0) {
    out_d[item] = sqrtf(best);
    if (out_ij) {
      if (flat == 0x7fffffff) flat = 0;   // only reachable with non-finite input
      out_ij[2 * item] = flat / m;
      out_ij[2 * item + 1] = flat - (flat / m) * m;
    }
  }
}

// EuclideanDistance (phoneme_to_articulation/metrics.py:5-24): per-point distance of planar contours, and its
// gradient wrt the first argument (the training criterion, train_phoneme_to_articulation.py:86).
__global__ void __launch_bounds__(256)
point_distance_kernel(const float* __restrict__ u, const float* __restrict__ v, float* __restrict__ out,
                      long long items, int n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= items * n) return;
  const long long item = i / n;
  const int k = int(i - item * n);
  const float dx = u[item * 2 * n + k] - v[item * 2 * n + k];
  const float dy = u[item * 2 * n + n + k] - v[item * 2 * n + n + k];
  out[i] = sqrtf(fmaf(dy, dy, dx * dx));
}
// grad_u[item][c][k] = g[item][k] * (u - v)[item][c][k] / d[item][k]  (0 where the two points coincide)
__global__ void __launch_bounds__(256)
point_distance_bwd_kernel(const float* __restrict__ u, const float* __restrict__ v, const float* __restrict__ g,
                          float* __restrict__ grad_u, long long items, int n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= items * n) return;
  const long long item = i / n;
  const int k = int(i - item * n);
  const float dx = u[item * 2 * n + k] - v[item * 2 * n + k];
  const float dy = u[item * 2 * n + n + k] - v[item * 2 * n + n + k];
  const float d = sqrtf(fmaf(dy, dy, dx * dx));
  const float s = d > 0.f ? g[i] / d : 0.f;
  grad_u[item * 2 * n + k] = s * dx;
  grad_u[item * 2 * n + n + k] = s * dy;
}

// pearsons_correlation (root metrics.py:9-35) with its quirk (the x TARGETS are centred with the x OUTPUT mean,
// :22): outputs / targets (B, T, N, 2, D) -> x_corr, y_corr (B, N, D); thread = (b, art, coordinate, point), two
// passes over T in fp64.
__global__ void __launch_bounds__(256)
pearson_kernel(const float* __restrict__ o, const float* __restrict__ t, int B, int T, int N, int D,
               float* __restrict__ x_corr, float* __restrict__ y_corr) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long total = (long long)B * N * 2 * D;
  if (i >= total) return;
  const int d = int(i % D);
  const int c = int((i / D) % 2);
  const int a = int((i / (2 * D)) % N);
  const int b = int(i / ((long long)2 * D * N));
  const size_t stride = (size_t)N * 2 * D;
  const size_t base = ((size_t)b * T * N + a) * 2 * D + (size_t)c * D + d;
  double so = 0.0, st = 0.0;
  for (int k = 0; k < T; ++k) {
    so += double(o[base + k * stride]);
    st += double(t[base + k * stride]);
  }
  const double mo = so / T, mt = c == 0 ? mo : st / T;   // :22 - the x targets use the OUTPUT mean
  double soo = 0.0, stt = 0.0, sot = 0.0;
  for (int k = 0; k < T; ++k) {
    const double vo = double(o[base + k * stride]) - mo, vt = double(t[base + k * stride]) - mt;
    soo += vo * vo;
    stt += vt * vt;
    sot += vo * vt;
  }
  const float r = float(sot / (sqrt(soo) * sqrt(stt) + 1e-5));
  (c == 0 ? x_corr : y_corr)[((size_t)b * N + a) * D + d] = r;
}

}  // namespace

extern "C" int asb_point_distance(const float* u, const float* v, float* out, int64_t items, int n, void* stream) {
  if (items < 0 || n <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_point_distance: bad shape");
  if (items == 0) return ASB_OK;
  if (!u || !v || !out) return asb::set_error(ASB_ERR_INVALID, "asb_point_distance: null pointer");
  const long long blocks = (items * n + 255) / 256;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_point_distance: too many items");
  point_distance_kernel<<<unsigned(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(u, v, out, items, n);
  ASB_CHECK_LAUNCH("point_distance_kernel");
  return ASB_OK;
}

extern "C" int asb_point_distance_backward(const float* u, const float* v, const float* grad_out, float* grad_u,
                                           int64_t items, int n, void* stream) {
  if (items < 0 || n <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_point_distance_backward: bad shape");
  if (items == 0) return ASB_OK;
  if (!u || !v || !grad_out || !grad_u) return asb::set_error(ASB_ERR_INVALID, "asb_point_distance_backward: null pointer");
  const long long blocks = (items * n + 255) / 256;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_point_distance_backward: too many items");
  point_distance_bwd_kernel<<<unsigned(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(u, v, grad_out, grad_u,
                                                                                            items, n);
  ASB_CHECK_LAUNCH("point_distance_bwd_kernel");
  return ASB_OK;
}

extern "C" int asb_pearson(const float* outputs, const float* targets, int B, int T, int n_art, int n_points,
                           float* x_corr, float* y_corr, void* stream) {
  if (B <= 0 || T <= 0 || n_art <= 0 || n_points <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_pearson: bad shape");
  if (!outputs || !targets || !x_corr || !y_corr) return asb::set_error(ASB_ERR_INVALID, "asb_pearson: null pointer");
  const long long total = (long long)B * n_art * 2 * n_points;
  const long long blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_pearson: too many items");
  pearson_kernel<<<unsigned(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(outputs, targets, B, T, n_art, n_points,
                                                                                 x_corr, y_corr);
  ASB_CHECK_LAUNCH("pearson_kernel");
  return ASB_OK;
}

extern "C" int asb_p2cp(const float* u, const float* v, float* out, int64_t items, int n, int m, int layout,
                        void* stream) {
  if (items < 0 || n <= 0 || m <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_p2cp: bad shape");
  if (layout != ASB_LAYOUT_PLANAR && layout != ASB_LAYOUT_INTERLEAVED)
    return asb::set_error(ASB_ERR_INVALID, "asb_p2cp: bad layout %d", layout);
  if (items == 0) return ASB_OK;
  if (!u || !v || !out) return asb::set_error(ASB_ERR_INVALID, "asb_p2cp: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool aligned = ((reinterpret_cast<uintptr_t>(u) | reinterpret_cast<uintptr_t>(v)) & 15) == 0;
  if (n == kD && m == kD && layout == ASB_LAYOUT_PLANAR && aligned) {
    static asb::DeviceOnce once;
    if (int rc = once.run([&] {
          return asb::check_cuda(cudaFuncSetAttribute(p2cp50_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                      int(kSmemBytes)),
                                 "asb_p2cp: smem attribute");
        }))
      return rc;
    const long long n_tiles = (items + kTile - 1) / kTile;
    const int grid = int(n_tiles < asb::sm_count() ? n_tiles : asb::sm_count());
    p2cp50_kernel<<<grid, kThreads, kSmemBytes, st>>>(u, v, out, items);
    ASB_CHECK_LAUNCH("p2cp50_kernel");
    return ASB_OK;
  }
  const int sp = layout == ASB_LAYOUT_PLANAR ? 1 : 2;
  const int warps = 8;
  const long long blocks = (items + warps - 1) / warps;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_p2cp: too many items");
  p2cp_generic_kernel<<<unsigned(blocks), warps * 32, 0, st>>>(u, v, out, items, n, m, sp,
                                                                 layout == ASB_LAYOUT_PLANAR ? n : 1, sp,
                                                                 layout == ASB_LAYOUT_PLANAR ? m : 1);
  ASB_CHECK_LAUNCH("p2cp_generic_kernel");
  return ASB_OK;
}

extern "C" int asb_mean_euclidean(const float* u, const float* v, float* out, int64_t items, int n, void* stream) {
  if (items < 0 || n <= 0) return asb::set_error(ASB_ERR_INVALID, "asb_mean_euclidean: bad shape");
  if (items == 0) return ASB_OK;
  if (!u || !v || !out) return asb::set_error(ASB_ERR_INVALID, "asb_mean_euclidean: null pointer");
  const int warps = 8;
  const long long blocks = (items + warps - 1) / warps;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_mean_euclidean: too many items");
  mean_euclidean_kernel<<<unsigned(blocks), warps * 32, 0, static_cast<cudaStream_t>(stream)>>>(u, v, out, items, n);
  ASB_CHECK_LAUNCH("mean_euclidean_kernel");
  return ASB_OK;
}

extern "C" int asb_min_distance(const float* u, const float* v, int64_t items, int n, int m, float* out_d,
                                int32_t* out_ij, void* stream) {
  if (items < 0 || n <= 0 || m <= 0 || (long long)n * m > 0x7ffffffeLL)
    return asb::set_error(ASB_ERR_INVALID, "asb_min_distance: bad shape");
  if (items == 0) return ASB_OK;
  if (!u || !v || !out_d) return asb::set_error(ASB_ERR_INVALID, "asb_min_distance: null pointer");
  const int warps = 8;
  const long long blocks = (items + warps - 1) / warps;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_min_distance: too many items");
  min_distance_kernel<<<unsigned(blocks), warps * 32, 0, static_cast<cudaStream_t>(stream)>>>(u, v, out_d, out_ij,
                                                                                             items, n, m);
  ASB_CHECK_LAUNCH("min_distance_kernel");
  return ASB_OK;
}
