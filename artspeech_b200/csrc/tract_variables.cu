// Tract variables: closest-pair searches LA, TTCD, TBCD, VEL over contour sets.
//
// Replaces calculate_vocal_tract_variables (tract_variables.py:73-125) and its helper
// _calculate_TV (:23-35), which the reference calls once per frame from a Python loop with
// three .item() host syncs per search (phoneme_to_articulation/__init__.py:242-261).
//
// One warp owns one contour set.  The six required contours (2 400 B) are staged in shared
// memory with 128-bit loads; for each search a lane owns columns j = lane, lane + 32 of the
// second point set and scans the rows of the first two at a time on packed fp32x2 arithmetic
// (direct dx*dx + dy*dy form, fp32).  Per column only the running minimum (FMNMX3) and the
// row-pair step at which it last strictly improved are tracked; the winning row is resolved
// afterwards by re-evaluating that one pair.  The cross-lane reduction orders candidates by
// (distance, j), which together with the in-lane strict '<' reproduces the reference's
// tie-break: smallest j among the global minima, then smallest i for that j.
#include "asb_common.cuh"

namespace {

using namespace asb;

constexpr int kD = 50;
constexpr int kItemF = 2 * kD;
constexpr int kWarps = 8;
// shared-memory slots
enum { S_LLIP = 0, S_PHAR = 1, S_SPM = 2, S_TONGUE = 3, S_UINC = 4, S_ULIP = 5, S_COUNT = 6 };

struct ArtIndex {
  int v[6];
};

struct Best {
  float d;  // squared distance
  int i, j;
};

// Squared distance of row i of (ax, ay) to the point (cx, cy), with exactly the operations the
// scan uses (packed subtract, multiply, fused multiply-add), so that re-evaluating a pair gives
// the bit pattern the scan saw.
__device__ __forceinline__ float pair_d2(const float* __restrict__ ax, const float* __restrict__ ay, int i, float cx,
                                         float cy) {
  const f32x2 dx = sub2(pack2(ax[i], ax[i]), pack2(cx, cx));
  const f32x2 dy = sub2(pack2(ay[i], ay[i]), pack2(cy, cy));
  float d0, d1;
  unpack2(fma2(dy, dy, mul2(dx, dx)), d0, d1);
  return d0;
}

// Column scan: minimum over the rows [0, P) of (ax, ay) of the squared distance to each of NC
// column points.  The rows are read once (broadcast LDS.64) for all NC columns of the lane; only
// the minimum is tracked here, the winning row is resolved afterwards for the one winning column.
template <int P, int NC>
__device__ __forceinline__ void scan_min(const float* __restrict__ ax, const float* __restrict__ ay,
                                         const float (&cx)[NC], const float (&cy)[NC], float (&m)[NC]) {
  const float kInf = __int_as_float(0x7f800000);
  f32x2 vx[NC], vy[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    vx[c] = pack2(cx[c], cx[c]);
    vy[c] = pack2(cy[c], cy[c]);
    m[c] = kInf;
  }
#pragma unroll
  for (int s = 0; s < P / 2; ++s) {
    const f32x2 ux = *reinterpret_cast<const f32x2*>(ax + 2 * s);
    const f32x2 uy = *reinterpret_cast<const f32x2*>(ay + 2 * s);
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const f32x2 dx = sub2(ux, vx[c]), dy = sub2(uy, vy[c]);
      float d0, d1;
      unpack2(fma2(dy, dy, mul2(dx, dx)), d0, d1);
      m[c] = min3(m[c], d0, d1);
    }
  }
  if (P & 1) {
#pragma unroll
    for (int c = 0; c < NC; ++c) m[c] = fminf(m[c], pair_d2(ax, ay, P - 1, cx[c], cy[c]));
  }
}

// One search: rows = P points starting at (ax, ay); columns = R points, column j read through
// the functor col(j, cx, cy).  A lane owns the columns lane and lane + 32.  The reduction orders
// candidates by (distance, j) - the reference's argmin over the flattened (j, i) matrix picks the
// smallest j among the global minima, then the smallest i for that j - and the row is found by
// re-evaluating the winning column against every row, one row per lane.
template <int P, int R, typename ColFn>
__device__ __forceinline__ Best search(const float* ax, const float* ay, ColFn col, int lane) {
  const float kInf = __int_as_float(0x7f800000);
  constexpr int NC = (R + 31) / 32;
  float cx[NC], cy[NC], m[NC];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int j = lane + 32 * c;
    col(j < R ? j : R - 1, cx[c], cy[c]);  // clamp: idle lanes redo the last column, then are ignored
  }
  scan_min<P, NC>(ax, ay, cx, cy, m);
  float d = (lane < R) ? m[0] : kInf;
  int j = lane;
  if (NC > 1 && lane + 32 < R && m[NC - 1] < d) {  // strict: an equal later column never displaces an earlier one
    d = m[NC - 1];
    j = lane + 32;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float od = __shfl_xor_sync(0xffffffffu, d, o);
    const int oj = __shfl_xor_sync(0xffffffffu, j, o);
    const bool take = (od < d) || (od == d && oj < j);
    d = take ? od : d;
    j = take ? oj : j;
  }
  j = j < R ? j : R - 1;  // only reachable with non-finite input
  float wx, wy;
  col(j, wx, wy);
  int row = P - 1;
  bool found = false;
#pragma unroll
  for (int pass = 0; pass < (P + 31) / 32; ++pass) {
    const int i = lane + 32 * pass;
    const float di = pair_d2(ax, ay, i < P ? i : P - 1, wx, wy);
    const unsigned hit = __ballot_sync(0xffffffffu, i < P && di == d);
    if (!found && hit) {
      row = 32 * pass + __ffs(hit) - 1;
      found = true;
    }
  }
  return Best{d, row, j};
}

#ifndef ASB_TV_MINB
#define ASB_TV_MINB 2
#endif
__global__ void __launch_bounds__(kWarps * 32, ASB_TV_MINB)
tract_variables_kernel(const float* __restrict__ sets_ptr, long long sets, int n_art, ArtIndex ai,
                       const float* __restrict__ uinc_ext, long long uinc_stride,
                       float* __restrict__ val, int* __restrict__ idx, float* __restrict__ poc) {
  // persistent warps: the next set's six contours are copied global -> shared (cp.async, no
  // registers) while the current one is searched; without this 23 % of the issue slots waited on
  // the staging loads (profiles/r1c_geometry_ncu.txt)
  __shared__ __align__(16) float sm[kWarps][2][S_COUNT * kItemF];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long stride = (long long)gridDim.x * kWarps;
  auto stage = [&](long long set, float* dst) {
    const float* base = sets_ptr + set * (long long)n_art * kItemF;
    // 6 contours x 25 x 16 bytes
    for (int q = lane; q < S_COUNT * (kItemF / 4); q += 32) {
      const int slot = q / (kItemF / 4), part = q - slot * (kItemF / 4);
      // the upper incisor may live outside the set (evaluation.py:93-109 injects a reference contour)
      const float* src = (slot == S_UINC && uinc_ext) ? uinc_ext + set * uinc_stride
                                                      : base + (long long)ai.v[slot] * kItemF;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst + slot * kItemF + 4 * part)),
                   "l"(src + 4 * part)
                   : "memory");
    }
  };
  long long set = (long long)blockIdx.x * kWarps + warp;
  if (set < sets) stage(set, sm[warp][0]);
  asm volatile("cp.async.commit_group;" ::: "memory");
  for (int it = 0; set < sets; set += stride, ++it) {
  float* s = sm[warp][it & 1];
  if (set + stride < sets) stage(set + stride, sm[warp][(it + 1) & 1]);
  asm volatile("cp.async.commit_group;" ::: "memory");
  asm volatile("cp.async.wait_group 1;" ::: "memory");   // everything but the copy just issued has landed
  __syncwarp();

  const float* llip = s + S_LLIP * kItemF;
  const float* phar = s + S_PHAR * kItemF;
  const float* spm = s + S_SPM * kItemF;
  const float* tongue = s + S_TONGUE * kItemF;
  const float* uinc = s + S_UINC * kItemF;
  const float* ulip = s + S_ULIP * kItemF;

  Best r[4];
  // LA: lower lip [0:50) x upper lip [0:50)                         tract_variables.py:38-41
  r[0] = search<50, 50>(llip, llip + kD, [&](int j, float& x, float& y) { x = ulip[j]; y = ulip[kD + j]; }, lane);
  // TTCD: tongue [30:45) x upper incisor [25:50)                     :44-51
  r[1] = search<15, 25>(tongue + 30, tongue + kD + 30,
                        [&](int j, float& x, float& y) { x = uinc[25 + j]; y = uinc[kD + 25 + j]; }, lane);
  // TBCD: tongue [10:30) x cat(upper incisor [0:25), soft palate midline [35:50))   :54-62
  r[2] = search<20, 40>(tongue + 10, tongue + kD + 10,
                        [&](int j, float& x, float& y) {
                          const float* p = j < 25 ? uinc + j : spm + 10 + j;  // 35 + (j - 25)
                          x = p[0];
                          y = p[kD];
                        },
                        lane);
  // VEL: soft palate midline [0:15) x pharynx [0:50)                 :65-70
  r[3] = search<15, 50>(spm, spm + kD, [&](int j, float& x, float& y) { x = phar[j]; y = phar[kD + j]; }, lane);

  if (lane < 4) {
    // lanes 0..3 each publish one tract variable
    Best b = r[0];
    const float* a1 = llip;
    const float* a2 = ulip;
    if (lane == 1) { b = r[1]; a1 = tongue + 30; a2 = uinc + 25; }
    if (lane == 2) { b = r[2]; a1 = tongue + 10; a2 = (b.j < 25) ? uinc : spm + 10; }
    if (lane == 3) { b = r[3]; a1 = spm; a2 = phar; }
    const long long o = set * 4 + lane;
    val[o] = sqrtf(b.d);
    idx[2 * o] = b.i;
    idx[2 * o + 1] = b.j;
    float4 p;
    p.x = a1[b.i];
    p.y = a1[kD + b.i];
    p.z = a2[b.j];
    p.w = a2[kD + b.j];
    *reinterpret_cast<float4*>(poc + 4 * o) = p;
  }
  __syncwarp();   // every lane is done with this buffer before the copy after next overwrites it
  }
}

}  // namespace

static int tv_launch(const float* sets_ptr, int64_t sets, int n_art, const int32_t* art_index_host,
                     const float* uinc, int64_t uinc_stride, float* val, int32_t* idx, float* poc, void* stream,
                     int max_ctas = 0) {
  if (sets < 0 || n_art <= 0 || !art_index_host)
    return asb::set_error(ASB_ERR_INVALID, "asb_tract_variables: bad arguments");
  ArtIndex ai;
  for (int k = 0; k < 6; ++k) {
    ai.v[k] = art_index_host[k];
    const bool external = (k == S_UINC && uinc != nullptr);
    if (!external && (ai.v[k] < 0 || ai.v[k] >= n_art))
      return asb::set_error(ASB_ERR_INVALID, "asb_tract_variables: articulator row %d out of range [0,%d)", ai.v[k],
                            n_art);
  }
  if (sets == 0) return ASB_OK;
  if (!sets_ptr || !val || !idx || !poc) return asb::set_error(ASB_ERR_INVALID, "asb_tract_variables: null pointer");
  if ((reinterpret_cast<uintptr_t>(sets_ptr) & 15) || (reinterpret_cast<uintptr_t>(poc) & 15) ||
      (reinterpret_cast<uintptr_t>(uinc) & 15) || (uinc_stride & 3))
    return asb::set_error(ASB_ERR_INVALID, "asb_tract_variables: sets/poc/upper_incisor must be 16-byte aligned");
  long long blocks = (sets + kWarps - 1) / kWarps;
  const long long resident = (long long)asb::sm_count() * ASB_TV_MINB;   // persistent: one wave of CTAs
  if (blocks > resident) blocks = resident;
  if (max_ctas > 0 && blocks > max_ctas) blocks = max_ctas;   // the sets are strided over whatever grid there is
  tract_variables_kernel<<<unsigned(blocks), kWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      sets_ptr, sets, n_art, ai, uinc, uinc_stride, val, idx, poc);
  ASB_CHECK_LAUNCH("tract_variables_kernel");
  return ASB_OK;
}

extern "C" int asb_tract_variables(const float* sets_ptr, int64_t sets, int n_art, const int32_t* art_index_host,
                                   float* val, int32_t* idx, float* poc, void* stream) {
  return tv_launch(sets_ptr, sets, n_art, art_index_host, nullptr, 0, val, idx, poc, stream);
}

extern "C" int asb_tract_variables_ref(const float* sets_ptr, int64_t sets, int n_art,
                                       const int32_t* art_index_host, const float* upper_incisor,
                                       int64_t upper_incisor_stride, float* val, int32_t* idx, float* poc,
                                       void* stream) {
  if (!upper_incisor) return asb::set_error(ASB_ERR_INVALID, "asb_tract_variables_ref: null upper_incisor");
  return tv_launch(sets_ptr, sets, n_art, art_index_host, upper_incisor, upper_incisor_stride, val, idx, poc, stream);
}

extern "C" int asb_tract_variables_ex(const float* sets_ptr, int64_t sets, int n_art, const int32_t* art_index_host,
                                      const float* upper_incisor, int64_t upper_incisor_stride, int max_ctas,
                                      float* val, int32_t* idx, float* poc, void* stream) {
  if (max_ctas < 0) return asb::set_error(ASB_ERR_INVALID, "asb_tract_variables_ex: max_ctas < 0");
  return tv_launch(sets_ptr, sets, n_art, art_index_host, upper_incisor, upper_incisor_stride, val, idx, poc, stream,
                   max_ctas);
}
