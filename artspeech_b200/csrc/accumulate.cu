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
                       double* __restrict__ acc) {
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
      if (j < cols) {
        acc[j < n_art ? j : j + 1] = s;
      } else {
        acc[n_art] = double(B);
        acc[n_art + 13] = s;
      }
    }
  }
}

}  // namespace

extern "C" int asb_eval_accumulate(const float* p2cp, const float* tv_pred, const float* tv_target,
                                   const int32_t* lengths_dev, int B, int T, int n_art, double* scratch, double* acc,
                                   void* stream) {
  if (B <= 0 || T <= 0 || n_art <= 0 || n_art + 13 > 64)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_accumulate: bad shape (B %d, T %d, n_art %d)", B, T, n_art);
  if (!p2cp || !tv_pred || !tv_target || !lengths_dev || !scratch || !acc)
    return asb::set_error(ASB_ERR_INVALID, "asb_eval_accumulate: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  utterance_partials_kernel<<<B, 256, 0, st>>>(p2cp, tv_pred, tv_target, lengths_dev, T, n_art, scratch);
  ASB_CHECK_LAUNCH("utterance_partials_kernel");
  reduce_partials_kernel<<<1, 1024, 0, st>>>(scratch, lengths_dev, B, T, n_art, acc);
  ASB_CHECK_LAUNCH("reduce_partials_kernel");
  return ASB_OK;
}
