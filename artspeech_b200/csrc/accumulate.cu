// Length-masked reduction of the per-frame metrics of one batch into the fp64 accumulator vector.
//
// Batched form of the aggregation in run_test (phoneme_to_articulation/encoder_decoder/
// evaluation.py:68-84,142-156): per sentence the mean P2CP over its valid frames, summed over
// sentences; plus sums of the tract variables over valid frames.  Two launches with a fixed
// summation order, so the result is bit-reproducible and a sharded run differs from a single-rank
// run only by the order of the final per-rank addition.
//   acc[0:N)      sum over utterances of the per-utterance mean P2CP per articulator
//   acc[N]        utterances
//   acc[N+1:N+5)  sum over valid frames of predicted TVs, [N+5:N+9) target TVs, [N+9:N+13) |pred-target|
//   acc[N+13]     valid frames
#include "asb_common.cuh"

namespace {

__device__ __forceinline__ double warp_sum_f64(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// one block per utterance, one warp per output column (strided)
__global__ void __launch_bounds__(256)
utterance_partials_kernel(const float* __restrict__ p2cp, const float* __restrict__ tv_p,
                          const float* __restrict__ tv_t, const int* __restrict__ lengths, int T, int n_art,
                          double* __restrict__ partial) {
  const int b = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int L = min(lengths[b], T);
  const int cols = n_art + 12;
  for (int j = warp; j < cols; j += 8) {
    double s = 0.0;
    if (j < n_art) {
      const float* p = p2cp + (size_t)b * T * n_art + j;
      for (int t = lane; t < L; t += 32) s += double(p[(size_t)t * n_art]);
      s = warp_sum_f64(s) / double(L);
    } else {
      const int k = (j - n_art) & 3, which = (j - n_art) >> 2;
      const float* a = tv_p + (size_t)b * T * 4 + k;
      const float* c = tv_t + (size_t)b * T * 4 + k;
      for (int t = lane; t < L; t += 32) {
        const float x = a[(size_t)t * 4], y = c[(size_t)t * 4];
        s += which == 0 ? double(x) : which == 1 ? double(y) : double(fabsf(x - y));
      }
      s = warp_sum_f64(s);
    }
    if (lane == 0) partial[(size_t)b * cols + j] = s;
  }
}

// one warp per accumulator column: lane l adds utterances l, l + 32, ... in order, then the fixed
// shuffle tree - a fixed summation order (a serial loop over the utterances took 24 us of
// dependent loads)
__global__ void __launch_bounds__(1024)
reduce_partials_kernel(const double* __restrict__ partial, const int* __restrict__ lengths, int B, int T, int n_art,
                       double* __restrict__ acc, double* __restrict__ total) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, cols = n_art + 12;
  for (int j = warp; j <= cols; j += 32) {
    double s = 0.0;
    if (j < cols) {
      for (int b = lane; b < B; b += 32) s += partial[(size_t)b * cols + j];
    } else {
      for (int b = lane; b < B; b += 32) s += double(min(lengths[b], T));   // exact: integers far below 2^53
    }
    s = warp_sum_f64(s);
    if (lane == 0) {
      // `total` (optional) is the running sum over the batches of an evaluation: one writer per element,
      // batches are stream-ordered, so its summation order is fixed as well
      if (j < cols) {
        const int o = j < n_art ? j : j + 1;
        acc[o] = s;
        if (total) total[o] += s;
      } else {
        acc[n_art] = double(B);
        acc[n_art + 13] = s;
        if (total) {
          total[n_art] += double(B);
          total[n_art + 13] += s;
        }
      }
    }
  }
}

// ---- per-sentence MED and Pearson correlations (run_test, evaluation.py:68-84) ------------------
// One block per (utterance, articulator); thread d < D owns point d: over the valid frames it sums
// the point distance (MED, metrics.py:54-68 over phoneme_to_articulation/metrics.py:10-24) and the
// five moments per coordinate from which pearsons_correlation (metrics.py:9-35) follows - including
// its quirk: the x TARGETS are centred with the mean of the x OUTPUTS (:22).  fp64 throughout.
// partial[(b * n_art + art) * 3 + {0, 1, 2}] = mean MED, mean_d x_corr, mean_d y_corr of the sentence.
__global__ void __launch_bounds__(64)
sentence_metrics_kernel(const float* __restrict__ outputs, const float* __restrict__ targets,
                        const int* __restrict__ lengths, int T, int n_art, int D, double* __restrict__ partial) {
  const int b = blockIdx.x, art = blockIdx.y, d = threadIdx.x;
  const int L = min(lengths[b], T);
  const size_t frame_stride = (size_t)n_art * 2 * D;
  const float* o = outputs + ((size_t)b * T * n_art + art) * 2 * D;
  const float* g = targets + ((size_t)b * T * n_art + art) * 2 * D;
  double med = 0.0, xc = 0.0, yc = 0.0;
  if (d < D) {
    double so[2] = {0, 0}, st[2] = {0, 0}, soo[2] = {0, 0}, stt[2] = {0, 0}, sot[2] = {0, 0};
    for (int t = 0; t < L; ++t) {
      const double xo = o[t * frame_stride + d], yo = o[t * frame_stride + D + d];
      const double xt = g[t * frame_stride + d], yt = g[t * frame_stride + D + d];
      med += sqrt((xo - xt) * (xo - xt) + (yo - yt) * (yo - yt));
      so[0] += xo; st[0] += xt; soo[0] += xo * xo; stt[0] += xt * xt; sot[0] += xo * xt;
      so[1] += yo; st[1] += yt; soo[1] += yo * yo; stt[1] += yt * yt; sot[1] += yo * yt;
    }
    const double n = double(L), eps = 1e-5;
    {   // x: both series centred with the OUTPUT mean
      const double m = so[0] / n;
      const double voo = fmax(soo[0] - n * m * m, 0.0), vtt = fmax(stt[0] - 2.0 * m * st[0] + n * m * m, 0.0);
      xc = (sot[0] - m * st[0]) / (sqrt(voo) * sqrt(vtt) + eps);
    }
    {   // y: the usual form
      const double mo = so[1] / n, mt = st[1] / n;
      const double voo = fmax(soo[1] - n * mo * mo, 0.0), vtt = fmax(stt[1] - n * mt * mt, 0.0);
      yc = (sot[1] - n * mo * mt) / (sqrt(voo) * sqrt(vtt) + eps);
    }
  }
  __shared__ double red[2][3];
  med = warp_sum_f64(med);
  xc = warp_sum_f64(xc);
  yc = warp_sum_f64(yc);
  if ((d & 31) == 0) {
    red[d >> 5][0] = med;
    red[d >> 5][1] = xc;
    red[d >> 5][2] = yc;
  }
  __syncthreads();
  if (d < 3) {
    const double v = red[0][d] + red[1][d];
    partial[((size_t)b * n_art + art) * 3 + d] = d == 0 ? v / (double(D) * double(L)) : v / double(D);
  }
}

// acc[k * n_art + art] = sum over utterances of partial[.][art][k]  (fixed order: lane-strided, then the tree)
__global__ void __launch_bounds__(1024)
reduce_sentence_metrics_kernel(const double* __restrict__ partial, int B, int n_art, double* __restrict__ acc) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int j = warp; j < 3 * n_art; j += 32) {
    const int k = j / n_art, art = j - k * n_art;
    double s = 0.0;
    for (int b = lane; b < B; b += 32) s += partial[((size_t)b * n_art + art) * 3 + k];
    s = warp_sum_f64(s);
    if (lane == 0) acc[j] = s;
  }
}

}  // namespace

extern "C" int asb_eval_accumulate(const float* p2cp, const float* tv_pred, const float* tv_target,
                                   const int32_t* lengths_dev, int B, int T, int n_art, double* scratch, double* acc,
                                   void* stream) {
  return asb_eval_accumulate_total(p2cp, tv_pred, tv_target, lengths_dev, B, T, n_art, scratch, acc, nullptr, stream);
}

extern "C" int asb_eval_accumulate_total(const float* p2cp, const float* tv_pred, const float* tv_target,
                                         const int32_t* lengths_dev, int B, int T, int n_art, double* scratch,
                                         double* acc, double* total, void* stream) {
  if (B <= 0 || T <= 0 || n_art <= 0 || n_art + 13 > 64)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_accumulate: bad shape (B %d, T %d, n_art %d)", B, T, n_art);
  if (!p2cp || !tv_pred || !tv_target || !lengths_dev || !scratch || !acc)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_accumulate: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  utterance_partials_kernel<<<B, 256, 0, st>>>(p2cp, tv_pred, tv_target, lengths_dev, T, n_art, scratch);
  ASB_CHECK_LAUNCH("utterance_partials_kernel");
  reduce_partials_kernel<<<1, 1024, 0, st>>>(scratch, lengths_dev, B, T, n_art, acc, total);
  ASB_CHECK_LAUNCH("reduce_partials_kernel");
  return ASB_OK;
}

// ---- allocation-free variant: the per-utterance partials live in a file-scope device array ------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (graph capture, serving loops).  Every GPU that loads the library holds the array until exit.
// Nothing in it has to start at zero: utterance_partials_kernel overwrites every element the reduction reads.
// All calls on a device share it, so two calls that overlap on the GPU (different streams) race.
constexpr int kStaticScratchDoubles = ASB_EVAL_STATIC_MAX_DOUBLES;
static __device__ __align__(256) double g_static_partials[kStaticScratchDoubles];

namespace {
// `g_static_partials` in host code is the host-side stand-in of the symbol, not a device address: the address
// is looked up once per device (with that device current) and cached.
int static_partials(double** out) {
  static std::atomic<double*> cache[64];
  int dev = -1;
  if (int rc = asb::check_cuda(cudaGetDevice(&dev), "asb_eval_accumulate_static: cudaGetDevice")) return rc;
  const bool cached = dev >= 0 && dev < 64;
  double* p = cached ? cache[dev].load(std::memory_order_acquire) : nullptr;
  if (!p) {
    if (int rc = asb::check_cuda(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_partials),
                                 "asb_eval_accumulate_static: cudaGetSymbolAddress"))
      return rc;
    if (!p) return asb::set_error(ASB_ERR_CUDA, "asb_eval_accumulate_static: null symbol address");
    if (cached) cache[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return ASB_OK;
}
}  // namespace

extern "C" int asb_eval_accumulate_static(const float* p2cp, const float* tv_pred, const float* tv_target,
                                          const int32_t* lengths_dev, int B, int T, int n_art, double* acc,
                                          void* stream) {
  if (B <= 0 || T <= 0 || n_art <= 0 || n_art + 13 > 64)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_accumulate_static: bad shape (B %d, T %d, n_art %d)", B, T, n_art);
  if ((long long)B * (n_art + 12) > kStaticScratchDoubles)
    return asb::set_error(ASB_ERR_WORKSPACE,
                          "asb_eval_accumulate_static: B * (n_art + 12) = %lld doubles exceeds the built-in scratch of "
                          "%d doubles; use asb_eval_accumulate with caller-provided scratch",
                          (long long)B * (n_art + 12), kStaticScratchDoubles);
  if (!p2cp || !tv_pred || !tv_target || !lengths_dev || !acc)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_accumulate_static: null pointer");
  double* scratch = nullptr;
  if (int rc = static_partials(&scratch)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  utterance_partials_kernel<<<B, 256, 0, st>>>(p2cp, tv_pred, tv_target, lengths_dev, T, n_art, scratch);
  ASB_CHECK_LAUNCH("utterance_partials_kernel (static scratch)");
  reduce_partials_kernel<<<1, 1024, 0, st>>>(scratch, lengths_dev, B, T, n_art, acc, nullptr);
  ASB_CHECK_LAUNCH("reduce_partials_kernel (static scratch)");
  return ASB_OK;
}

extern "C" int asb_eval_sentence_metrics(const float* outputs, const float* targets, const int32_t* lengths_dev, int B,
                                         int T, int n_art, int n_points, double* scratch, double* acc, void* stream) {
  if (B <= 0 || T <= 0 || n_art <= 0 || n_points <= 0 || n_points > 64)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_sentence_metrics: bad shape (B %d, T %d, n_art %d, points %d)", B,
                          T, n_art, n_points);
  if (!outputs || !targets || !lengths_dev || !scratch || !acc)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_sentence_metrics: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  sentence_metrics_kernel<<<dim3(B, n_art), 64, 0, st>>>(outputs, targets, lengths_dev, T, n_art, n_points, scratch);
  ASB_CHECK_LAUNCH("sentence_metrics_kernel");
  reduce_sentence_metrics_kernel<<<1, 1024, 0, st>>>(scratch, B, n_art, acc);
  ASB_CHECK_LAUNCH("reduce_sentence_metrics_kernel");
  return ASB_OK;
}
