// Area function: semipolar-grid / wall intersections -> tube areas -> evenly resampled profile.
//
// Replaces the chain intersect_semipolar_grid (area_function.py:175-223) -> area_function
// (:124-142, mid_point :113-121) -> evenly_spaced_fx (:145-159), which the reference runs per frame
// through shapely/GEOS calls in Python loops (2 * n_lines + 200 GEOS calls per frame).
//
// One warp owns one frame; lane owns grid lines l = lane, lane + 32, ... (4 rounds, <= 128 lines),
// whose fp32 line coefficients stay in registers across all frames of the persistent CTA.
//   1. both walls (2 * n_wall * 8 B) are staged in shared memory with coalesced 8-byte loads;
//   2. FILTER: for every (line, wall vertex) the signed area about the grid line is evaluated in
//      fp32 (2 FFMA); its sign bit is shifted into a per-line mask (SHF) and min |area| is tracked
//      (FMNMX).  A sign is trusted only if |area| exceeds a forward error bound 2^-21 * M; vertices
//      below it are re-evaluated with the fp64 expression the oracle uses, so every decision
//      equals the fp64 one;
//   3. sign changes in the masks are the candidate wall segments; only they get the fp64
//      intersection (orientation tests + one division), enumerated by ffs over the masks, with the
//      reference's closest-pair selection (row-major first minimum, both walls' views) and its
//      single-contact rule (pair with the other wall's first/last vertex, chosen by the distance
//      to the EXTERNAL wall's ends, :199-221);
//   4. touched lines are compacted in grid order (ballot/popc), mid-points, radii and
//      alpha * r^beta go to shared memory, the cumulative mid-point distance is a warp scan;
//   5. each lane evaluates the piecewise-linear profile at its share of the n_samples abscissae
//      (branch-free binary search in shared memory) and writes the (2, n_samples) fp32 result coalesced.
// All fp64 arithmetic uses the non-fusing intrinsics so that it rounds like NumPy / gcc
// -ffp-contract=off.
#include "asb_common.cuh"

#include <cfloat>

namespace {

using namespace asb;

#ifndef ASB_AREA_ABLATE
#define ASB_AREA_ABLATE 0   // timing experiments (tools/area_variants.py): 1 = no resampling; 2 = no crossings (then K = 0: staging + filter + NaN rows only); 6 = staging + NaN rows only
#endif
#ifndef ASB_AREA_UNROLL
#define ASB_AREA_UNROLL 1   // rounds of the crossing stage interleaved per trip
#endif
#ifndef ASB_AREA_MINB
#define ASB_AREA_MINB 2
#endif
constexpr int kRoundUnroll = ASB_AREA_UNROLL;
constexpr int kWarps = 8;
constexpr int kMaxWall = 64;    // vertices per wall (one 64-bit sign mask)
constexpr int kRounds = 4;      // grid lines per lane
constexpr int kMaxLines = 32 * kRounds;
typedef unsigned long long u64;

__device__ __forceinline__ double orient64(double ax, double ay, double bx, double by, double cx, double cy) {
  // (bx-ax)*(cy-ay) - (by-ay)*(cx-ax), every operation separately rounded
  return __dsub_rn(__dmul_rn(__dsub_rn(bx, ax), __dsub_rn(cy, ay)), __dmul_rn(__dsub_rn(by, ay), __dsub_rn(cx, ax)));
}
__device__ __forceinline__ double dist64(double ax, double ay, double bx, double by) {
  const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by);
  return __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
}

__device__ __forceinline__ double dist2_64(double ax, double ay, double bx, double by) {
  const double dx = __dsub_rn(ax, bx), dy = __dsub_rn(ay, by);
  return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}

// Shared point of grid line g0g1 and wall segment k (vertices k, k+1), if any.  Branch-free: all
// four orientations and the division are always evaluated (candidates come from sign changes, so
// the early-outs are rare) and the special cases are selects, which lets independent calls overlap.
__device__ __forceinline__ bool crossing(double g0x, double g0y, double g1x, double g1y, const float4* wall, int k,
                                         double& px, double& py) {
  const float4 a = wall[k], b = wall[k + 1];  // (x, x, y, y)
  const double ax = a.x, ay = a.z, bx = b.x, by = b.z;
  const double d1 = orient64(g0x, g0y, g1x, g1y, ax, ay);
  const double d2 = orient64(g0x, g0y, g1x, g1y, bx, by);
  const double d3 = orient64(ax, ay, bx, by, g0x, g0y);
  const double d4 = orient64(ax, ay, bx, by, g1x, g1y);
  const bool same12 = (d1 > 0 && d2 > 0) || (d1 < 0 && d2 < 0);
  const bool same34 = (d3 > 0 && d4 > 0) || (d3 < 0 && d4 < 0);
  const bool allz = d1 == 0 && d2 == 0 && d3 == 0 && d4 == 0;  // collinear overlap contributes nothing
  const double t = __ddiv_rn(d1, __dsub_rn(d1, d2));
  px = __dadd_rn(ax, __dmul_rn(t, __dsub_rn(bx, ax)));
  py = __dadd_rn(ay, __dmul_rn(t, __dsub_rn(by, ay)));
  // touching: the exact end point, priority d1, d2, d3, d4 (lowest priority applied first)
  if (d4 == 0) { px = g1x; py = g1y; }
  if (d3 == 0) { px = g0x; py = g0y; }
  if (d2 == 0) { px = bx; py = by; }
  if (d1 == 0) { px = ax; py = ay; }
  return !same12 && !same34 && !allz;
}

struct LineCoef {
  float ex, ey, c;   // area(w) = ex * w.y - ey * w.x - c
  float ma, mc;      // error-bound terms: |ex| + |ey| and |c|
};

// A wall vertex in shared memory, duplicated for the packed fp32x2 filter: (x, x, y, y).
typedef float4 WallVtx;

// Sign masks of the signed areas of every vertex of `wall` about the lane's kRounds lines.
// neg[r] bit k: area < 0; zer[r] bit k: area == 0 (exact fp64 decision for uncertain vertices).
// Lines are handled two per instruction (FFMA2), vertices two per FMNMX3.
__device__ __forceinline__ void wall_masks(const WallVtx* wall, int n_wall, const LineCoef (&lc)[kRounds], int lane,
                                           const double* sgrid, float wmax, u64 (&neg)[kRounds], u64 (&zer)[kRounds]) {
  static_assert(kRounds == 4, "the packed filter handles the lane's lines as two pairs");
  // area = ex * y + (-(ey * x + c)) = fma(ex, y, fma(-ey, x, -c))
  const f32x2 ex01 = pack2(lc[0].ex, lc[1].ex), ex23 = pack2(lc[2].ex, lc[3].ex);
  const f32x2 ney01 = pack2(-lc[0].ey, -lc[1].ey), ney23 = pack2(-lc[2].ey, -lc[3].ey);
  const f32x2 nc01 = pack2(-lc[0].c, -lc[1].c), nc23 = pack2(-lc[2].c, -lc[3].c);
  unsigned lo[kRounds], hi[kRounds];
  float m[kRounds];
#pragma unroll
  for (int r = 0; r < kRounds; ++r) { lo[r] = 0u; hi[r] = 0u; m[r] = FLT_MAX; }

  auto vertex_pair = [&](int k, unsigned (&acc)[kRounds]) {
    const ulonglong2 wa = *reinterpret_cast<const ulonglong2*>(wall + k);      // (x,x), (y,y)
    const ulonglong2 wb = *reinterpret_cast<const ulonglong2*>(wall + k + 1);
    float a[kRounds], b[kRounds];
    unpack2(fma2(ex01, wa.y, fma2(ney01, wa.x, nc01)), a[0], a[1]);
    unpack2(fma2(ex23, wa.y, fma2(ney23, wa.x, nc23)), a[2], a[3]);
    unpack2(fma2(ex01, wb.y, fma2(ney01, wb.x, nc01)), b[0], b[1]);
    unpack2(fma2(ex23, wb.y, fma2(ney23, wb.x, nc23)), b[2], b[3]);
#pragma unroll
    for (int r = 0; r < kRounds; ++r) {
      m[r] = fminf(fminf(m[r], fabsf(a[r])), fabsf(b[r]));
      acc[r] = __funnelshift_l(__float_as_uint(a[r]), acc[r], 1);
      acc[r] = __funnelshift_l(__float_as_uint(b[r]), acc[r], 1);
    }
  };
  auto vertex_one = [&](int k, unsigned (&acc)[kRounds]) {
    const ulonglong2 wa = *reinterpret_cast<const ulonglong2*>(wall + k);
    float a[kRounds];
    unpack2(fma2(ex01, wa.y, fma2(ney01, wa.x, nc01)), a[0], a[1]);
    unpack2(fma2(ex23, wa.y, fma2(ney23, wa.x, nc23)), a[2], a[3]);
#pragma unroll
    for (int r = 0; r < kRounds; ++r) {
      m[r] = fminf(m[r], fabsf(a[r]));
      acc[r] = __funnelshift_l(__float_as_uint(a[r]), acc[r], 1);
    }
  };

  const int n0 = n_wall < 32 ? n_wall : 32;
  int k = 0;
#pragma unroll 4
  for (; k + 1 < n0; k += 2) vertex_pair(k, lo);
  if (k < n0) vertex_one(k, lo);
  k = 32;
#pragma unroll 3
  for (; k + 1 < n_wall; k += 2) vertex_pair(k, hi);
  if (k < n_wall) vertex_one(k, hi);

  const int n1 = n_wall - n0;
#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    // first vertex shifted in first -> reverse so that vertex k sits at bit k
    const unsigned l32 = __brev(lo[r]) >> (32 - n0);
    const unsigned h32 = n1 > 0 ? (__brev(hi[r]) >> (32 - n1)) : 0u;
    neg[r] = (u64)l32 | ((u64)h32 << 32);
    zer[r] = 0ull;
    const float bound = fmaxf(0x1p-21f * __fmaf_rn(lc[r].ma, wmax, lc[r].mc), FLT_MIN);
    if (m[r] < bound) {
      // rare: some vertex is too close to this line for fp32; decide those in fp64
      const double* g = sgrid + 4 * (r * 32 + lane);
      const double g0x = g[0], g0y = g[1], g1x = g[2], g1y = g[3];
      for (int q = 0; q < n_wall; ++q) {
        const WallVtx w = wall[q];
        const float det = __fmaf_rn(lc[r].ex, w.z, __fmaf_rn(-lc[r].ey, w.x, -lc[r].c));
        if (fabsf(det) < bound) {
          const double d = orient64(g0x, g0y, g1x, g1y, (double)w.x, (double)w.z);
          const u64 bit = 1ull << q;
          neg[r] = d < 0 ? (neg[r] | bit) : (neg[r] & ~bit);
          zer[r] = d == 0 ? (zer[r] | bit) : zer[r];
        }
      }
    }
  }
}

__device__ __forceinline__ u64 candidates(u64 neg, u64 zer, int n_wall) {
  const u64 segs = (n_wall >= 65) ? ~0ull : ((1ull << (n_wall - 1)) - 1ull);
  return ((neg ^ (neg >> 1)) | zer | (zer >> 1)) & segs;
}

// evenly_spaced_fx (:145-159): the polyline (x, fx) of K >= 2 points evaluated at
// np.linspace(x[0], x[K-1], n_samples); out (2, n_samples) fp32.  One warp, one sample per lane and
// pass, so the global writes are coalesced as they are.  The segment of an abscissa - the first k
// with x[k] <= xv <= x[k+1], i.e. the number of x[1..K-1] below xv (x is non-decreasing), capped at
// K-2 - comes from a branch-free binary search with a warp-uniform trip count; the special cases
// are selects.  (The earlier version gave each lane 8 consecutive abscissae and walked the segment
// index forward: its divergent loops took 25 % of the kernel's stall samples for 7 % of its work.)
template <bool kSlope>
__device__ __forceinline__ void resample(const double* x, const double* fx, const double* slope, int K,
                                         int n_samples, float* out, int lane) {
  const double x0 = x[0], x1 = x[K - 1];
  const double step = __ddiv_rn(__dsub_rn(x1, x0), (double)(n_samples - 1));
  const double qnan = __longlong_as_double(0x7ff8000000000000LL);
  const int n = K - 1;                       // candidates x[1..K-1]
  const int top = 1 << (31 - __clz(n));      // n >= 1
  for (int s = lane; s < n_samples; s += 32) {
    const double xv = (s == n_samples - 1) ? x1 : __dadd_rn(x0, __dmul_rn((double)s, step));
    int pos = 0;                             // number of x[1..K-1] that are < xv
    for (int half = top; half > 0; half >>= 1) {
      const int t = pos + half;
      const double probe = x[t <= n ? t : n];
      pos = (t <= n && probe < xv) ? t : pos;
    }
    const int seg = pos < K - 2 ? pos : K - 2;
    const double da = x[seg], db = x[seg + 1], f0 = fx[seg], f1 = fx[seg + 1];
    double yv;
    if (kSlope)
      yv = __dadd_rn(f0, __dmul_rn(__dsub_rn(xv, da), slope[seg]));
    else
      yv = __dadd_rn(f0, __dmul_rn(__ddiv_rn(__dsub_rn(xv, da), __dsub_rn(db, da)), __dsub_rn(f1, f0)));
    yv = (xv == db) ? f1 : yv;
    yv = (db == da || xv == da) ? f0 : yv;
    yv = (da <= xv && xv <= db) ? yv : qnan;  // outside the polyline
    out[s] = (float)xv;
    out[n_samples + s] = (float)yv;
  }
}

struct AreaArgs {
  const float2* wi;
  const float2* we;
  long long frames;
  int n_wall;
  const double* grid;  // (n_lines, 2, 2)
  int n_lines;
  double alpha, beta;
  int n_samples;
  float* xfx;
  int* K;
  double* pts;
  int* lines;
};

__global__ void __launch_bounds__(kWarps * 32, ASB_AREA_MINB) area_function_kernel(const AreaArgs p) {
  __shared__ __align__(16) double sgrid[kMaxLines * 4];
  __shared__ __align__(16) WallVtx swall[kWarps][2][kMaxWall];
  __shared__ __align__(16) double smid[kWarps][2][kMaxLines];
  __shared__ __align__(16) double sfx[kWarps][kMaxLines];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int q = threadIdx.x; q < kMaxLines * 4; q += blockDim.x) {
    sgrid[q] = (q < p.n_lines * 4) ? p.grid[q] : 0.0;
  }
  __syncthreads();

  LineCoef lc[kRounds];
#pragma unroll
  for (int r = 0; r < kRounds; ++r) {
    const int l = r * 32 + lane;
    if (l < p.n_lines) {
      const double* g = sgrid + 4 * l;
      const double ex = g[2] - g[0], ey = g[3] - g[1];
      const double c = ex * g[1] - ey * g[0];
      lc[r].ex = (float)ex; lc[r].ey = (float)ey; lc[r].c = (float)c;
      lc[r].ma = fabsf(lc[r].ex) + fabsf(lc[r].ey);
      lc[r].mc = fabsf(lc[r].c);
    } else {
      lc[r].ex = 0.f; lc[r].ey = 0.f; lc[r].c = 1.f; lc[r].ma = 0.f; lc[r].mc = 1.f;  // area == -1 everywhere
    }
  }

  WallVtx* wI = swall[warp][0];
  WallVtx* wE = swall[warp][1];
  double* mx = smid[warp][0];
  double* my = smid[warp][1];
  double* dist = mx;  // the cumulative distances overwrite the mid-point abscissae once those are consumed
  double* fxs = sfx[warp];
  const int nw = p.n_wall;
  const unsigned lt_mask = (1u << lane) - 1u;

  // The walls of the NEXT frame are copied global -> shared (cp.async, no registers) while the
  // current one is processed: the staging loads of a frame used to sit un-overlapped at its start
  // (12 % of the stall samples on their first consumer, profiles/r1d_geometry_ncu.txt).
  extern __shared__ __align__(16) float2 sraw_all[];                // [kWarps][2 walls][kMaxWall]
  float2* sraw = sraw_all + warp * 2 * kMaxWall;
  const long long fstride = (long long)gridDim.x * kWarps;
  auto prefetch = [&](long long f) {
    for (int q = lane; q < 2 * nw; q += 32) {
      const int which = q >= nw;
      const int k = q - which * nw;
      asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(smem_u32(sraw + which * kMaxWall + k)),
                   "l"((which ? p.we : p.wi) + f * nw + k)
                   : "memory");
    }
  };
  {
    const long long f0 = (long long)blockIdx.x * kWarps + warp;
    if (f0 < p.frames) prefetch(f0);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (long long f = (long long)blockIdx.x * kWarps + warp; f < p.frames; f += fstride) {
    // ---- 1. expand the staged walls to the filter's (x, x, y, y) layout, then refill the staging buffer
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    float wmax = 0.f;
    for (int q = lane; q < 2 * nw; q += 32) {
      const int which = q >= nw;
      const int k = q - which * nw;
      const float2 v = sraw[which * kMaxWall + k];
      swall[warp][which][k] = make_float4(v.x, v.x, v.y, v.y);
      wmax = fmaxf(wmax, fmaxf(fabsf(v.x), fabsf(v.y)));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wmax = fmaxf(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
    __syncwarp();
    if (f + fstride < p.frames) prefetch(f + fstride);
    asm volatile("cp.async.commit_group;" ::: "memory");

    // ---- 2. sign masks
    u64 negI[kRounds], zerI[kRounds], negE[kRounds], zerE[kRounds];
#if ASB_AREA_ABLATE & 4
    for (int r = 0; r < kRounds; ++r) { negI[r] = __float_as_uint(wI[lane].x); zerI[r] = 0; negE[r] = __float_as_uint(wE[lane].x); zerE[r] = 0; }
#else
    wall_masks(wI, nw, lc, lane, sgrid, wmax, negI, zerI);
    wall_masks(wE, nw, lc, lane, sgrid, wmax, negE, zerE);
#endif

    const double eFx = wE[0].x, eFy = wE[0].z, eLx = wE[nw - 1].x, eLy = wE[nw - 1].z;

    // ---- 3./4. per line: intersections, selection, compaction
    int base = 0;
#pragma unroll(kRoundUnroll)
    for (int r = 0; r < kRounds; ++r) {
      const int l = r * 32 + lane;
#if ASB_AREA_ABLATE & 2
      const u64 cI = (negI[r] ^ negE[r]) == 0x123456789ull, cE = 0;   // never a candidate, masks stay live
#else
      const u64 cI = candidates(negI[r], zerI[r], nw);
      const u64 cE = candidates(negE[r], zerE[r], nw);
#endif
      bool anyI = false, anyE = false;
      double ix = 0, iy = 0, ox = 0, oy = 0;  // selected internal / external point
      // Distances are only compared, so the selection works on squared distances.
      const int nI = __popcll(cI), nE = __popcll(cE);
      if (__all_sync(0xffffffffu, nI <= 1 && nE <= 1)) {
        // fast path (the usual case): at most one candidate segment per wall on every lane's line;
        // both intersections are evaluated side by side, everything else is selects
        if (nI | nE) {
          const double* g = sgrid + 4 * l;
          const double g0x = g[0], g0y = g[1], g1x = g[2], g1y = g[3];
          double px, py, qx, qy;
          const bool okP = crossing(g0x, g0y, g1x, g1y, wI, nI ? __ffsll((long long)cI) - 1 : 0, px, py);
          const bool okQ = crossing(g0x, g0y, g1x, g1y, wE, nE ? __ffsll((long long)cE) - 1 : 0, qx, qy);
          anyI = nI && okP;
          anyE = nE && okQ;
          ix = px; iy = py; ox = qx; oy = qy;
          if (anyI != anyE) {
            // one wall only (:199-221): the met point against the ends of the EXTERNAL wall
            const double mx_ = anyI ? px : qx, my_ = anyI ? py : qy;
            const bool first = dist2_64(mx_, my_, eFx, eFy) <= dist2_64(mx_, my_, eLx, eLy);
            if (anyI) {
              ox = first ? eFx : eLx; oy = first ? eFy : eLy;
            } else {
              const WallVtx v = first ? wI[0] : wI[nw - 1];
              ix = v.x; iy = v.z;
            }
          }
        }
      } else if (cI | cE) {
        const double* g = sgrid + 4 * l;
        const double g0x = g[0], g0y = g[1], g1x = g[2], g1y = g[3];
        double b1 = INFINITY, b2 = INFINITY;
        int j2 = 0x7fffffff;
        for (u64 ci = cI; ci; ci &= ci - 1) {
          double px, py;
          if (!crossing(g0x, g0y, g1x, g1y, wI, __ffsll((long long)ci) - 1, px, py)) continue;
          anyI = true;
          for (u64 ce = cE; ce; ce &= ce - 1) {
            const int kj = __ffsll((long long)ce) - 1;
            double qx, qy;
            if (!crossing(g0x, g0y, g1x, g1y, wE, kj, qx, qy)) continue;
            anyE = true;
            const double d = dist2_64(px, py, qx, qy);
            if (d < b1) { b1 = d; ix = px; iy = py; }                          // rows = internal points
            if (d < b2 || (d == b2 && kj < j2)) { b2 = d; j2 = kj; ox = qx; oy = qy; }  // rows = external
          }
        }
        if (anyI && !anyE) {
          // only the internal wall is met (:204-212): closest to an end of the external wall
          double best = INFINITY;
          for (u64 ci = cI; ci; ci &= ci - 1) {
            double px, py;
            if (!crossing(g0x, g0y, g1x, g1y, wI, __ffsll((long long)ci) - 1, px, py)) continue;
            const double dF = dist2_64(px, py, eFx, eFy), dL = dist2_64(px, py, eLx, eLy);
            if (dF < best) { best = dF; ix = px; iy = py; ox = eFx; oy = eFy; }
            if (dL < best) { best = dL; ix = px; iy = py; ox = eLx; oy = eLy; }
          }
        } else if (!anyI) {
          // only the external wall may be met (:214-221): distances still to the external ends
          double best = INFINITY;
          for (u64 ce = cE; ce; ce &= ce - 1) {
            double qx, qy;
            if (!crossing(g0x, g0y, g1x, g1y, wE, __ffsll((long long)ce) - 1, qx, qy)) continue;
            anyE = true;
            const double dF = dist2_64(qx, qy, eFx, eFy), dL = dist2_64(qx, qy, eLx, eLy);
            if (dF < best) { best = dF; ox = qx; oy = qy; ix = wI[0].x; iy = wI[0].z; }
            if (dL < best) { best = dL; ox = qx; oy = qy; ix = wI[nw - 1].x; iy = wI[nw - 1].z; }
          }
        }
      }
      const bool valid = anyI || anyE;
      const unsigned ball = __ballot_sync(0xffffffffu, valid);
      if (valid) {
        const int k = base + __popc(ball & lt_mask);
        // mid_point :113-121 and radius / alpha * r^beta :126-133
        mx[k] = __dadd_rn(fmin(ix, ox), __dmul_rn(fabs(__dsub_rn(ix, ox)), 0.5));
        my[k] = __dadd_rn(fmin(iy, oy), __dmul_rn(fabs(__dsub_rn(iy, oy)), 0.5));
        const double d2 = dist2_64(ix, iy, ox, oy);   // (|int - ext| / 2)^2 = d2 / 4: no square root for beta = 2
        fxs[k] = __dmul_rn(p.alpha, p.beta == 2.0 ? __dmul_rn(d2, 0.25)
                                                  : pow(__dmul_rn(__dsqrt_rn(d2), 0.5), p.beta));
        if (p.pts) {
          double* o = p.pts + (f * p.n_lines + k) * 4;
          reinterpret_cast<double2*>(o)[0] = make_double2(ix, iy);
          reinterpret_cast<double2*>(o)[1] = make_double2(ox, oy);
        }
        if (p.lines) p.lines[f * p.n_lines + k] = l;
      }
      base += __popc(ball);
    }
    const int K = base;
    __syncwarp();
    if (lane == 0) p.K[f] = K;
    if (p.pts)
      for (int k = K * 4 + lane; k < p.n_lines * 4; k += 32) p.pts[f * p.n_lines * 4 + k] = 0.0;
    if (p.lines)
      for (int k = K + lane; k < p.n_lines; k += 32) p.lines[f * p.n_lines + k] = -1;

    float* out = p.xfx + f * 2 * (long long)p.n_samples;
    if (K < 2) {
      const float nanv = __int_as_float(0x7fc00000);
      for (int s = lane; s < 2 * p.n_samples; s += 32) out[s] = nanv;
      __syncwarp();
      continue;
    }

    // ---- cumulative distance between successive mid-points (:135-140): lane owns kRounds entries
    double seg[kRounds];
    double run = 0.0;
#pragma unroll
    for (int q = 0; q < kRounds; ++q) {
      const int k = lane * kRounds + q;
      seg[q] = (k > 0 && k < K) ? dist64(mx[k - 1], my[k - 1], mx[k], my[k]) : 0.0;
      run = __dadd_rn(run, seg[q]);
      seg[q] = run;
    }
    double incl = run;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl = __dadd_rn(incl, t);
    }
    double excl = __shfl_up_sync(0xffffffffu, incl, 1);
    if (lane == 0) excl = 0.0;
    __syncwarp();  // every lane has read its mid-points; dist aliases mx
#pragma unroll
    for (int q = 0; q < kRounds; ++q) {
      const int k = lane * kRounds + q;
      if (k < K) dist[k] = __dadd_rn(excl, seg[q]);
    }
    __syncwarp();

    // ---- 5. evenly_spaced_fx (:145-159); the walls are consumed, their shared memory holds the
    // per-segment slopes (one division per segment instead of one per sample)
    double* slope = reinterpret_cast<double*>(wI);
#pragma unroll
    for (int q = 0; q < kRounds; ++q) {
      const int k = lane * kRounds + q;
      if (k < K - 1) {
        const double w = __dsub_rn(dist[k + 1], dist[k]);
        slope[k] = (w == 0.0) ? 0.0 : __ddiv_rn(__dsub_rn(fxs[k + 1], fxs[k]), w);
      }
    }
    __syncwarp();
#if !(ASB_AREA_ABLATE & 1)
    resample<true>(dist, fxs, slope, K, p.n_samples, out, lane);
#endif
    __syncwarp();
  }
}

// area_function (:124-142) on already paired wall points: (frames, K, 2) fp64 x 2 -> dists, fx (frames, K).
__global__ void __launch_bounds__(kWarps * 32)
area_from_pairs_kernel(const double2* __restrict__ pi, const double2* __restrict__ pe, long long frames, int K,
                       double alpha, double beta, double* __restrict__ dists, double* __restrict__ fx) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long f = (long long)blockIdx.x * kWarps + warp;
  if (f >= frames) return;
  pi += f * K; pe += f * K; dists += f * K; fx += f * K;
  double carry = 0.0;
  for (int k0 = 0; k0 < K; k0 += 32) {
    const int k = k0 + lane;
    double seg = 0.0;
    if (k < K) {
      const double2 a = pi[k], b = pe[k];
      const double mxk = __dadd_rn(fmin(a.x, b.x), __dmul_rn(fabs(__dsub_rn(a.x, b.x)), 0.5));
      const double myk = __dadd_rn(fmin(a.y, b.y), __dmul_rn(fabs(__dsub_rn(a.y, b.y)), 0.5));
      const double radius = __dmul_rn(dist64(a.x, a.y, b.x, b.y), 0.5);
      fx[k] = __dmul_rn(alpha, beta == 2.0 ? __dmul_rn(radius, radius) : pow(radius, beta));
      if (k > 0) {
        const double2 a0 = pi[k - 1], b0 = pe[k - 1];
        const double mx0 = __dadd_rn(fmin(a0.x, b0.x), __dmul_rn(fabs(__dsub_rn(a0.x, b0.x)), 0.5));
        const double my0 = __dadd_rn(fmin(a0.y, b0.y), __dmul_rn(fabs(__dsub_rn(a0.y, b0.y)), 0.5));
        seg = dist64(mx0, my0, mxk, myk);
      }
    }
    double incl = seg;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const double t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl = __dadd_rn(incl, t);
    }
    if (k < K) dists[k] = __dadd_rn(carry, incl);
    carry = __dadd_rn(carry, __shfl_sync(0xffffffffu, incl, 31));
  }
}

__global__ void __launch_bounds__(kWarps * 32)
evenly_spaced_kernel(const double* __restrict__ x, const double* __restrict__ fx, long long frames, int K,
                     int n_samples, float* __restrict__ out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long f = (long long)blockIdx.x * kWarps + warp;
  if (f >= frames) return;
  resample<false>(x + f * K, fx + f * K, nullptr, K, n_samples, out + f * 2 * (long long)n_samples, lane);
}

}  // namespace

extern "C" int asb_area_function(const float* walls_int, const float* walls_ext, int64_t frames, int n_wall,
                                 const double* grid_ends, int n_lines, double alpha, double beta, int n_samples,
                                 float* xfx, int32_t* K, double* pts, int32_t* lines, void* stream) {
  if (frames < 0 || n_wall < 2 || n_wall > kMaxWall || n_lines < 1 || n_lines > kMaxLines || n_samples < 2)
    return asb::set_error(ASB_ERR_INVALID,
                          "asb_area_function: need 2 <= n_wall <= %d, 1 <= n_lines <= %d, n_samples >= 2 (got %d, %d, %d)",
                          kMaxWall, kMaxLines, n_wall, n_lines, n_samples);
  if (frames == 0) return ASB_OK;
  if (!walls_int || !walls_ext || !grid_ends || !xfx || !K)
    return asb::set_error(ASB_ERR_INVALID, "asb_area_function: null pointer");
  if ((reinterpret_cast<uintptr_t>(walls_int) & 7) || (reinterpret_cast<uintptr_t>(walls_ext) & 7) ||
      (reinterpret_cast<uintptr_t>(grid_ends) & 7) || (reinterpret_cast<uintptr_t>(pts) & 15))
    return asb::set_error(ASB_ERR_INVALID, "asb_area_function: walls/grid must be 8-byte, pts 16-byte aligned");
  AreaArgs a;
  a.wi = reinterpret_cast<const float2*>(walls_int);
  a.we = reinterpret_cast<const float2*>(walls_ext);
  a.frames = frames; a.n_wall = n_wall; a.grid = grid_ends; a.n_lines = n_lines;
  a.alpha = alpha; a.beta = beta; a.n_samples = n_samples;
  a.xfx = xfx; a.K = K; a.pts = pts; a.lines = lines;
  const long long want = (frames + kWarps - 1) / kWarps;
  const long long cap = (long long)asb::sm_count() * 4;
  const unsigned blocks = (unsigned)(want < cap ? want : cap);
  const int raw_bytes = kWarps * 2 * kMaxWall * int(sizeof(float2));   // wall staging buffer (dynamic part)
  static asb::DeviceOnce once;
  if (int rc = once.run([&] {
        return asb::check_cuda(cudaFuncSetAttribute(area_function_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                    raw_bytes),
                               "asb_area_function: smem attribute");
      }))
    return rc;
  area_function_kernel<<<blocks, kWarps * 32, raw_bytes, static_cast<cudaStream_t>(stream)>>>(a);
  ASB_CHECK_LAUNCH("area_function_kernel");
  return ASB_OK;
}

extern "C" int asb_area_from_pairs(const double* int_pts, const double* ext_pts, int64_t frames, int K, double alpha,
                                   double beta, double* dists, double* fx, void* stream) {
  if (frames < 0 || K < 1) return asb::set_error(ASB_ERR_INVALID, "asb_area_from_pairs: need K >= 1 (got %d)", K);
  if (frames == 0) return ASB_OK;
  if (!int_pts || !ext_pts || !dists || !fx) return asb::set_error(ASB_ERR_INVALID, "asb_area_from_pairs: null pointer");
  if ((reinterpret_cast<uintptr_t>(int_pts) & 15) || (reinterpret_cast<uintptr_t>(ext_pts) & 15))
    return asb::set_error(ASB_ERR_INVALID, "asb_area_from_pairs: points must be 16-byte aligned");
  const long long blocks = (frames + kWarps - 1) / kWarps;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_area_from_pairs: too many frames");
  area_from_pairs_kernel<<<unsigned(blocks), kWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const double2*>(int_pts), reinterpret_cast<const double2*>(ext_pts), frames, K, alpha, beta,
      dists, fx);
  ASB_CHECK_LAUNCH("area_from_pairs_kernel");
  return ASB_OK;
}

extern "C" int asb_evenly_spaced_fx(const double* x, const double* fx, int64_t frames, int K, int n_samples,
                                    float* out, void* stream) {
  if (frames < 0 || K < 2 || n_samples < 2)
    return asb::set_error(ASB_ERR_INVALID, "asb_evenly_spaced_fx: need K >= 2 and n_samples >= 2 (got %d, %d)", K,
                          n_samples);
  if (frames == 0) return ASB_OK;
  if (!x || !fx || !out) return asb::set_error(ASB_ERR_INVALID, "asb_evenly_spaced_fx: null pointer");
  const long long blocks = (frames + kWarps - 1) / kWarps;
  if (blocks > 0x7fffffffLL) return asb::set_error(ASB_ERR_INVALID, "asb_evenly_spaced_fx: too many frames");
  evenly_spaced_kernel<<<unsigned(blocks), kWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(x, fx, frames, K,
                                                                                               n_samples, out);
  ASB_CHECK_LAUNCH("evenly_spaced_kernel");
  return ASB_OK;
}
