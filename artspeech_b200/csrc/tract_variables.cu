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

// Column scan: rows [0, P) of (ax, ay) against the single point (cx, cy).
// Returns min squared distance and the smallest row index attaining it.
template <int P>
__device__ __forceinline__ void scan_rows(const float* __restrict__ ax, const float* __restrict__ ay, float cx,
                                          float cy, float& best, int& best_i) {
  const float kInf = __int_as_float(0x7f800000);
  const f32x2 vx = pack2(cx, cx), vy = pack2(cy, cy);
  float m = kInf;
  int step = 0;
#pragma unroll
  for (int s = 0; s < P / 2; ++s) {
    const f32x2 ux = *reinterpret_cast<const f32x2*>(ax + 2 * s);
    const f32x2 uy = *reinterpret_cast<const f32x2*>(ay + 2 * s);
    const f32x2 dx = sub2(ux, vx), dy = sub2(uy, vy);
    float d0, d1;
    unpack2(fma2(dy, dy, mul2(dx, dx)), d0, d1);
    const float mn = min3(m, d0, d1);
    step = (mn < m) ? s : step;
    m = mn;
  }
  if (P & 1) {
    const f32x2 ux = pack2(ax[P - 1], ax[P - 1]);
    const f32x2 uy = pack2(ay[P - 1], ay[P - 1]);
    const f32x2 dx = sub2(ux, vx), dy = sub2(uy, vy);
    float d0, d1;
    unpack2(fma2(dy, dy, mul2(dx, dx)), d0, d1);
    const float mn = fminf(m, d0);
    step = (mn < m) ? (P / 2) : step;
    m = mn;
  }
  // resolve the row inside the winning pair with the same arithmetic
  int row = 2 * step;
  if (2 * step + 1 < P) {
    const f32x2 ux = *reinterpret_cast<const f32x2*>(ax + 2 * step);
    const f32x2 uy = *reinterpret_cast<const f32x2*>(ay + 2 * step);
    const f32x2 dx = sub2(ux, vx), dy = sub2(uy, vy);
    float d0, d1;
    unpack2(fma2(dy, dy, mul2(dx, dx)), d0, d1);
    row = (d0 == m) ? 2 * step : 2 * step + 1;
  }
  best = m;
  best_i = row;
}

__device__ __forceinline__ Best warp_argmin(Best b) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float od = __shfl_xor_sync(0xffffffffu, b.d, o);
    const int oi = __shfl_xor_sync(0xffffffffu, b.i, o);
    const int oj = __shfl_xor_sync(0xffffffffu, b.j, o);
    const bool take = (od < b.d) || (od == b.d && oj < b.j);
    b.d = take ? od : b.d;
    b.i = take ? oi : b.i;
    b.j = take ? oj : b.j;
  }
  return b;
}

// One search: rows = P points starting at (ax, ay); columns = R points, column j read through
// the functor col(j, cx, cy).
template <int P, int R, typename ColFn>
__device__ __forceinline__ Best search(const float* ax, const float* ay, ColFn col, int lane) {
  const float kInf = __int_as_float(0x7f800000);
  Best b{kInf, 0, 0x7fffffff};
#pragma unroll
  for (int pass = 0; pass < (R + 31) / 32; ++pass) {
    const int j = lane + 32 * pass;
    const int jc = j < R ? j : R - 1;  // clamp: idle lanes redo the last column, then are ignored
    float cx, cy;
    col(jc, cx, cy);
    float d;
    int i;
    scan_rows<P>(ax, ay, cx, cy, d, i);
    if (j < R && d < b.d) {  // strict: an equal later column never displaces an earlier one
      b.d = d;
      b.i = i;
      b.j = j;
    }
  }
  return warp_argmin(b);
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
                     const float* uinc, int64_t uinc_stride, float* val, int32_t* idx, float* poc, void* stream) {
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
