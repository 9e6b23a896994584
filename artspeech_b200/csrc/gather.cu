// Row gather: out[r, :] = table[tokens[r], :].
//
// Two users.  (1) nn.Embedding (models.py:63,135) for SimpleArtSpeech's padded feature rows.
// (2) SimpleArtSpeech.forward itself (models.py:84-96): that model has no recurrence, so the
// contours of a frame depend on its token only - the predictors are evaluated once per vocabulary
// entry (V x 4000 B, L2-resident) and the forward pass is this gather, bound by the 4000 B per
// frame it has to write.
//
// Flat float4 index space over rows x (row_floats / 4): consecutive threads write consecutive
// 16-byte pieces (fully coalesced streaming stores), the table reads hit L1/L2.  A token outside
// [0, vocab) produces a row of NaNs and, through asb_gather_rows_flagged, {1, id} in the caller's status
// word (nn.Embedding raises IndexError on the host; the Python wrapper raises it from that word).
#include "asb_common.cuh"

namespace {

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float4* __restrict__ table, const long long* __restrict__ tokens, long long rows, int vocab,
                   int q /* float4 per row */, float4* __restrict__ out, int* __restrict__ tok_flag) {
  const long long total = rows * q;
  const long long step = (long long)gridDim.x * blockDim.x;
  const float nan = __int_as_float(0x7fc00000);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += step) {
    const long long r = i / q;
    const int part = int(i - r * q);
    const long long t = __ldg(tokens + r);
    float4 v = make_float4(nan, nan, nan, nan);
    if (t >= 0 && t < vocab) {
      v = __ldg(table + t * q + part);
    } else if (tok_flag && part == 0) {   // {1, id}: the host raises IndexError like nn.Embedding
      tok_flag[1] = int(t);
      tok_flag[0] = 1;
    }
    __stcs(out + i, v);
  }
}

}  // namespace

extern "C" int asb_gather_rows(const float* table, const int64_t* tokens, int64_t rows, int vocab, int row_floats,
                               float* out, void* stream) {
  return asb_gather_rows_flagged(table, tokens, rows, vocab, row_floats, out, nullptr, stream);
}

extern "C" int asb_gather_rows_flagged(const float* table, const int64_t* tokens, int64_t rows, int vocab,
                                       int row_floats, float* out, int32_t* token_flag, void* stream) {
  if (rows < 0 || vocab <= 0 || row_floats <= 0 || (row_floats & 3))
    return asb::set_error(ASB_ERR_INVALID, "asb_gather_rows: bad sizes (rows %lld, vocab %d, row_floats %d, must be a multiple of 4)",
                          (long long)rows, vocab, row_floats);
  if (rows == 0) return ASB_OK;
  if (!table || !tokens || !out) return asb::set_error(ASB_ERR_INVALID, "asb_gather_rows: null pointer");
  if ((reinterpret_cast<uintptr_t>(table) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return asb::set_error(ASB_ERR_INVALID, "asb_gather_rows: table/out must be 16-byte aligned");
  const int q = row_floats / 4;
  const long long total = rows * q;
  long long blocks = (total + 255) / 256;
  const long long resident = (long long)asb::sm_count() * 8;   // one wave of 256-thread CTAs
  if (blocks > resident) blocks = resident;
  gather_rows_kernel<<<unsigned(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(table), reinterpret_cast<const long long*>(tokens), rows, vocab, q,
      reinterpret_cast<float4*>(out), token_flag);
  ASB_CHECK_LAUNCH("gather_rows_kernel");
  return ASB_OK;
}
