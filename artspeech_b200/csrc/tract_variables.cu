// Tract variables: closest-pair searches LA, TTCD, TBCD, VEL over contour sets.
//
// Replaces calculate_vocal_tract_variables (tract_variables.py:73-125) and its helper
// _calculate_TV (:23-35), which the reference calls once per frame from a Python loop with
// three .item() host syncs per search (phoneme_to_articulation/__init__.py:242-261).
//
// Round-2 kernel.  A TEAM of 8 lanes owns one contour set (4 sets per warp).  Persistent CTAs, one per SM, of
// NW <= 14 compute warps + one PRODUCER warp: the contours of the next group of 4 NW sets are staged in
// shared memory by 1-D bulk async copies (TMA) into a 2-stage full/empty mbarrier ring while the current group
// is searched; adjacent contours of a set are merged into one copy (two copies per set in the thesis layout
// instead of five: a 400-byte bulk copy costs as much to issue as a large one, and with five the kernel was
// bound by issuing them).  Inside a team a search is split HS x (8 / HS): lane (h, q) keeps the rows of part h
// of the first point set in registers as natural fp32x2 pairs and streams the columns of part q of the second,
// two columns at a time with the next pair's operands prefetched; per (row pair, column) the body is
// 2 FADD2 + FMUL2 + FFMA2 + one FMNMX3 - all 32 lanes busy, <= 12 % padding (the round-1 kernel gave a lane
// whole columns: 62-78 % of the lanes had work, and every row was an LDS broadcast per two columns).
// Distances use the direct (dx*dx + dy*dy) form in fp32; only the minimum per column is tracked.  The four
// searches of a set are reduced TOGETHER by (distance, j) shuffles with the smaller j winning ties, then the
// rows are re-evaluated against the winning column with the scan's roundings and the smallest i that
// reproduces the minimum bit for bit is taken: the reference's tie-break (smallest j among the global minima,
// then smallest i for that j, tract_variables.py:26-29).
// Shared-memory banks: a set is 500 (or 600) floats, so the four teams of a warp start 20 (24) banks apart
// and the row parts / column parts of a team fall into distinct banks for the 50 x 50 search (a few 2-way
// conflicts remain in the 15-row searches); lanes that share a row part or a column part read the same word.
#include <utility>

#include "asb_common.cuh"

namespace {

using namespace asb;

constexpr int kD = 50;
constexpr int kItemF = 2 * kD;
constexpr int kTeam = 8;                          // lanes per contour set
constexpr int kStages = 2;
// NW compute warps (4 teams each) + one producer warp; NW <= 14 when a set needs 5 staged contours (the
// upper incisor shared by all sets, evaluation.py:93-109), <= 8 when it needs 6 (shared memory bound)
constexpr int kTeamsPerWarp = 32 / kTeam;
// shared-memory slots
enum { S_LLIP = 0, S_PHAR = 1, S_SPM = 2, S_TONGUE = 3, S_UINC = 4, S_ULIP = 5, S_COUNT = 6 };
__host__ __device__ constexpr size_t tv_smem_bytes(int nw, int slots) {
  return size_t(kStages) * nw * kTeamsPerWarp * slots * kItemF * sizeof(float) + kItemF * sizeof(float) +
         2 * kStages * sizeof(uint64_t);
}

struct ArtIndex {
  int v[6];
};

// How a set is staged: the needed contours in SOURCE order, adjacent ones merged into one bulk copy
// (the thesis layout needs lower-lip, pharynx, soft-palate-midline = rows 3..5 and tongue, upper-lip =
// rows 7..8: two copies per set instead of five - a 400-byte bulk copy costs as much to issue as a large one).
struct Staging {
  int n_runs;
  int run_src[S_COUNT];   // first source row of the run
  int run_len[S_COUNT];   // contours in the run
  int run_dst[S_COUNT];   // first destination slot of the run
  int slot_pos[S_COUNT];  // destination slot of every logical contour (S_LLIP ...)
  int n_slots;            // contours staged per set from the set itself
};

struct Best {
  float d;  // squared distance
  int i, j;
};

// One search of a team: rows = P points at (ax, ay) (8-byte aligned, P <= 50), columns = R points read
// through col(j, x, y).  The 8 lanes split the search HS x (8 / HS): lane (h, q) keeps the rows of part h
// (of HS) in registers and streams the columns of part q.  HS = 2 for the 50-row search; the 15- and
// 20-row searches keep ALL rows in every lane (HS = 1) and split the columns 8 ways, which halves the
// per-column bookkeeping per evaluated pair.
template <int P, int R, int HS, typename ColFn>
__device__ __forceinline__ Best search(const float* __restrict__ ax, const float* __restrict__ ay, ColFn col, int tl) {
  const float kInf = __int_as_float(0x7f800000);
  constexpr int QS = kTeam / HS;                  // column parts
  constexpr int PH = (P + HS - 1) / HS;           // rows per part
  constexpr int NP = HS == 1 ? (P + 1) / 2 : (PH + 2) / 2;   // row pairs per lane (a later part starts on an even row)
  constexpr int RQ = (R + QS - 1) / QS;           // columns per part
  const int h = tl / QS, q = tl % QS;
  const int r0 = (h * PH) & ~1;                   // first row of this lane (even: LDS.64 pairs)
  f32x2 ux[NP], uy[NP];
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    int i0 = r0 + 2 * p;
    i0 = i0 < P - 1 ? i0 : ((P - 1) & ~1);        // pairs past the end repeat the last pair
    float2 x = *reinterpret_cast<const float2*>(ax + i0);
    float2 y = *reinterpret_cast<const float2*>(ay + i0);
    if (i0 + 1 >= P) {                            // odd P: the last pair's second row is not part of the set
      x.y = x.x;
      y.y = y.x;
    }
    ux[p] = pack2(x.x, x.y);
    uy[p] = pack2(y.x, y.y);
  }
  float best = kInf;
  int bj = 0;
  // Columns two at a time (two independent minimum chains per column, four in flight), the operands of
  // the NEXT pair fetched before the current pair is evaluated so that the LDS latency is hidden.
  auto column = [&](int c) { const int j = q * RQ + c; return j < R ? j : R - 1; };   // a repeated column never wins a tie
  float cxa, cya, cxb, cyb;
  col(column(0), cxa, cya);
  col(column(1), cxb, cyb);
#pragma unroll (RQ <= 8 ? 4 : 1)
  for (int c = 0; c + 1 < RQ; c += 2) {
    const int ja = column(c), jb = column(c + 1);
    const f32x2 vxa = pack2(cxa, cxa), vya = pack2(cya, cya);
    const f32x2 vxb = pack2(cxb, cxb), vyb = pack2(cyb, cyb);
    col(column(c + 2), cxa, cya);
    col(column(c + 3), cxb, cyb);
    float a0 = kInf, a1 = kInf, b0 = kInf, b1 = kInf;
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const f32x2 dxa = sub2(ux[p], vxa), dya = sub2(uy[p], vya);
      const f32x2 dxb = sub2(ux[p], vxb), dyb = sub2(uy[p], vyb);
      float d0, d1, e0, e1;
      unpack2(fma2(dya, dya, mul2(dxa, dxa)), d0, d1);
      unpack2(fma2(dyb, dyb, mul2(dxb, dxb)), e0, e1);
      if (p & 1) {
        a1 = min3(a1, d0, d1);
        b1 = min3(b1, e0, e1);
      } else {
        a0 = min3(a0, d0, d1);
        b0 = min3(b0, e0, e1);
      }
    }
    const float ma = fminf(a0, a1), mb = fminf(b0, b1);
    if (ma < best) {
      best = ma;
      bj = ja;
    }
    if (mb < best && jb > ja) {   // jb == ja only for the repeated last column
      best = mb;
      bj = jb;
    }
  }
  if (RQ & 1) {   // odd part: the last column on its own (its operands are in cxa, cya)
    const int ja = column(RQ - 1);
    const f32x2 vxa = pack2(cxa, cxa), vya = pack2(cya, cya);
    float a0 = kInf, a1 = kInf;
#pragma unroll
    for (int p = 0; p < NP; ++p) {
      const f32x2 dxa = sub2(ux[p], vxa), dya = sub2(uy[p], vya);
      float d0, d1;
      unpack2(fma2(dya, dya, mul2(dxa, dxa)), d0, d1);
      if (p & 1) a1 = min3(a1, d0, d1); else a0 = min3(a0, d0, d1);
    }
    const float ma = fminf(a0, a1);
    if (ma < best) {
      best = ma;
      bj = ja;
    }
  }
  return Best{best, 0, bj};   // lane-local: the team reduction and the row follow in finish()
}

// Team reduction of the four searches TOGETHER (four independent shuffle chains hide each other's
// latency), ordered by (distance, j); then the row of each: the smallest i whose distance to column j
// reproduces the minimum - lane tl checks the rows tl, tl + 8, ... with the scan's roundings (sub.rn,
// mul.rn, fma.rn: each half of the packed instructions is exactly the scalar operation).
struct Rows {
  const float* ax;
  const float* ay;
  int P;
};

__device__ __forceinline__ void finish(Best (&r)[4], const Rows (&rows)[4], const float (&wx)[4], const float (&wy)[4],
                                       int tl) {
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    int bi = 0x7fffffff;
    for (int i = tl; i < rows[k].P; i += kTeam) {
      const float dx = __fsub_rn(rows[k].ax[i], wx[k]), dy = __fsub_rn(rows[k].ay[i], wy[k]);
      const float d = __fmaf_rn(dy, dy, __fmul_rn(dx, dx));
      bi = (d == r[k].d && i < bi) ? i : bi;
    }
    r[k].i = bi;
  }
#pragma unroll
  for (int o = 1; o < kTeam; o <<= 1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int oi = __shfl_xor_sync(0xffffffffu, r[k].i, o);
      r[k].i = oi < r[k].i ? oi : r[k].i;
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) r[k].i = r[k].i < rows[k].P ? r[k].i : rows[k].P - 1;   // only reachable with non-finite input
}

__device__ __forceinline__ void reduce_dj(Best (&r)[4]) {
#pragma unroll
  for (int o = 1; o < kTeam; o <<= 1) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float od = __shfl_xor_sync(0xffffffffu, r[k].d, o);
      const int oj = __shfl_xor_sync(0xffffffffu, r[k].j, o);
      const bool take = (od < r[k].d) || (od == r[k].d && oj < r[k].j);
      r[k].d = take ? od : r[k].d;
      r[k].j = take ? oj : r[k].j;
    }
  }
}

__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int NW>
__global__ void __launch_bounds__((NW + 1) * 32, 1)
tract_variables_kernel(const float* __restrict__ sets_ptr, long long sets, int n_art, Staging sg,
                       const float* __restrict__ uinc_ext, long long uinc_stride,
                       float* __restrict__ val, int* __restrict__ idx, float* __restrict__ poc) {
  constexpr int kSets = NW * kTeamsPerWarp;        // contour sets per group (one per team)
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const bool shared_uinc = uinc_ext != nullptr && uinc_stride == 0;
  const bool own_uinc = uinc_ext != nullptr && !shared_uinc;
  const int slots = sg.n_slots + (own_uinc ? 1 : 0);   // contours staged per set
  const int set_f = slots * kItemF;
  float* sbuf = reinterpret_cast<float*>(smem_raw);                    // [stage][set][slots * 100]
  float* suinc = sbuf + kStages * kSets * set_f;                       // shared upper incisor (stride 0)
  uint64_t* full = reinterpret_cast<uint64_t*>(suinc + kItemF);        // [stage] contours landed
  uint64_t* empty = full + kStages;                                    // [stage] every compute warp is done with it

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long n_groups = (sets + kSets - 1) / kSets;

  if (tid == 0) {
    for (int st = 0; st < kStages; ++st) {
      mbar_init(&full[st], 1);
      mbar_init(&empty[st], NW);
    }
    fence_mbar_init();
  }
  if (shared_uinc && tid < kItemF / 4)   // evaluation.py:93-109 injects one fixed reference contour
    reinterpret_cast<float4*>(suinc)[tid] = __ldg(reinterpret_cast<const float4*>(uinc_ext) + tid);
  __syncthreads();

  if (warp == NW) {
    // producer warp: one bulk copy per run of adjacent contours (and per external upper incisor) of every
    // set of the group, spread over the lanes; a stage is refilled when all compute warps released it
    int k = 0;
    for (long long group = blockIdx.x; group < n_groups; group += gridDim.x, ++k) {
      const int stage = k & 1;
      if (k >= kStages) mbar_wait(&empty[stage], ((k >> 1) - 1) & 1);
      const long long first = group * kSets;
      const long long left = sets - first;
      const int cnt = left < kSets ? int(left) : kSets;
      if (lane == 0) mbar_arrive_expect_tx(&full[stage], uint32_t(cnt * slots * kItemF * sizeof(float)));
      __syncwarp();
      const int copies = sg.n_runs + (own_uinc ? 1 : 0);
      for (int w = lane; w < cnt * copies; w += 32) {
        const int s_local = w / copies, r = w - s_local * copies;
        const long long set = first + s_local;
        float* dst = sbuf + (stage * kSets + s_local) * set_f;
        if (r < sg.n_runs) {
          bulk_g2s(dst + sg.run_dst[r] * kItemF, sets_ptr + (set * n_art + sg.run_src[r]) * kItemF,
                   uint32_t(sg.run_len[r] * kItemF * sizeof(float)), &full[stage]);
        } else {
          bulk_g2s(dst + sg.slot_pos[S_UINC] * kItemF, uinc_ext + set * uinc_stride, kItemF * sizeof(float),
                   &full[stage]);
        }
      }
    }
    return;
  }

  const int team = lane >> 3, tl = lane & 7;
  const int local = warp * kTeamsPerWarp + team;
  int k = 0;
  for (long long group = blockIdx.x; group < n_groups; group += gridDim.x, ++k) {
    const int stage = k & 1;
    mbar_wait(&full[stage], (k >> 1) & 1);

    const long long set = group * kSets + local;
    const bool valid = set < sets;
    // teams past the end of the input redo set 0 of the group (no divergence around the shuffles), no stores
    const float* s = sbuf + (stage * kSets + (valid ? local : 0)) * set_f;
    const float* llip = s + sg.slot_pos[S_LLIP] * kItemF;
    const float* phar = s + sg.slot_pos[S_PHAR] * kItemF;
    const float* spm = s + sg.slot_pos[S_SPM] * kItemF;
    const float* tongue = s + sg.slot_pos[S_TONGUE] * kItemF;
    const float* uinc = shared_uinc ? suinc : s + sg.slot_pos[S_UINC] * kItemF;
    const float* ulip = s + sg.slot_pos[S_ULIP] * kItemF;

    auto c_la = [&](int j, float& x, float& y) { x = ulip[j]; y = ulip[kD + j]; };
    auto c_ttcd = [&](int j, float& x, float& y) { x = uinc[25 + j]; y = uinc[kD + 25 + j]; };
    auto c_tbcd = [&](int j, float& x, float& y) {
      const float* p = j < 25 ? uinc + j : spm + 10 + j;  // 35 + (j - 25)
      x = p[0];
      y = p[kD];
    };
    auto c_vel = [&](int j, float& x, float& y) { x = phar[j]; y = phar[kD + j]; };
    Best r[4];
    // LA: lower lip [0:50) x upper lip [0:50)                         tract_variables.py:38-41
    r[0] = search<50, 50, 2>(llip, llip + kD, c_la, tl);
    // TTCD: tongue [30:45) x upper incisor [25:50)                     :44-51
    r[1] = search<15, 25, 1>(tongue + 30, tongue + kD + 30, c_ttcd, tl);
    // TBCD: tongue [10:30) x cat(upper incisor [0:25), soft palate midline [35:50))   :54-62
    r[2] = search<20, 40, 1>(tongue + 10, tongue + kD + 10, c_tbcd, tl);
    // VEL: soft palate midline [0:15) x pharynx [0:50)                 :65-70
    r[3] = search<15, 50, 1>(spm, spm + kD, c_vel, tl);
    reduce_dj(r);
    const Rows rows[4] = {{llip, llip + kD, 50}, {tongue + 30, tongue + kD + 30, 15}, {tongue + 10, tongue + kD + 10, 20},
                          {spm, spm + kD, 15}};
    float wx[4], wy[4];
    c_la(r[0].j, wx[0], wy[0]);
    c_ttcd(r[1].j, wx[1], wy[1]);
    c_tbcd(r[2].j, wx[2], wy[2]);
    c_vel(r[3].j, wx[3], wy[3]);
    finish(r, rows, wx, wy, tl);

    if (valid && tl < 4) {
      // lanes 0..3 of the team each publish one tract variable
      Best b = r[0];
      const float* a1 = llip;
      const float* a2 = ulip;
      if (tl == 1) { b = r[1]; a1 = tongue + 30; a2 = uinc + 25; }
      if (tl == 2) { b = r[2]; a1 = tongue + 10; a2 = (b.j < 25) ? uinc : spm + 10; }
      if (tl == 3) { b = r[3]; a1 = spm; a2 = phar; }
      const long long o = set * 4 + tl;
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
    __syncwarp();
    if (lane == 0) mbar_arrive_cta(&empty[stage]);   // this warp is done with the stage
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
  // staging plan: the contours taken from the set itself in source order, adjacent rows merged
  Staging sg{};
  {
    int order[S_COUNT], n = 0;
    for (int k = 0; k < S_COUNT; ++k)
      if (!(k == S_UINC && uinc != nullptr)) order[n++] = k;
    for (int a = 1; a < n; ++a)          // insertion sort by source row
      for (int b = a; b > 0 && ai.v[order[b]] < ai.v[order[b - 1]]; --b) std::swap(order[b], order[b - 1]);
    for (int a = 0; a < n; ++a) {
      // the same source row may serve two logical contours only if the caller repeats an index: give each its own slot
      sg.slot_pos[order[a]] = a;
      const bool extend = sg.n_runs > 0 && ai.v[order[a]] == sg.run_src[sg.n_runs - 1] + sg.run_len[sg.n_runs - 1];
      if (extend) {
        sg.run_len[sg.n_runs - 1] += 1;
      } else {
        sg.run_src[sg.n_runs] = ai.v[order[a]];
        sg.run_len[sg.n_runs] = 1;
        sg.run_dst[sg.n_runs] = a;
        sg.n_runs += 1;
      }
    }
    sg.n_slots = n;
    if (uinc != nullptr) sg.slot_pos[S_UINC] = n;   // its own slot after the set's contours (unused when shared)
  }
  // Compute warps per CTA (4 sets each): up to 14 when five contours are staged per set (the shared upper
  // incisor), up to 8 when six (shared memory).  A CTA works in rounds of one set per team, so a mid-sized
  // input is cut into ceil(groups / SMs) rounds: the warp count is chosen to minimise rounds x warps (e.g.
  // 25 600 sets on 148 SMs: 11 warps = 4 rounds of 44 sets instead of 4 rounds of 56), larger counts winning ties.
  const int slots = sg.n_slots + ((uinc != nullptr && uinc_stride != 0) ? 1 : 0);
  const int max_nw = slots <= 5 ? 14 : 8;
  const long long sms = (max_ctas > 0 && max_ctas < asb::sm_count()) ? max_ctas : asb::sm_count();
  int nw = max_nw;
  double best = 1e300;
  for (int cand = max_nw; cand >= 8; --cand) {
    const long long groups = (sets + cand * kTeamsPerWarp - 1) / (cand * kTeamsPerWarp);
    const long long rounds = (groups + sms - 1) / sms;
    const double cost = double(rounds) * cand * (1.0 + 0.015 * (max_nw - cand));
    if (cost < best * 0.999) {
      best = cost;
      nw = cand;
    }
  }
#ifdef ASB_TV_FORCE_NW   // experiments (tools/build_variant.py)
  nw = ASB_TV_FORCE_NW < max_nw ? ASB_TV_FORCE_NW : max_nw;
#endif
  const size_t smem = tv_smem_bytes(nw, slots);
  static asb::DeviceOnce once;
  if (int rc = once.run([&] {
        cudaError_t e = cudaSuccess;
        auto opt_in = [&](auto kernel, int w) {
          if (e == cudaSuccess)
            e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(tv_smem_bytes(w, w > 8 ? 5 : 6)));
        };
        opt_in(tract_variables_kernel<8>, 8);
        opt_in(tract_variables_kernel<9>, 9);
        opt_in(tract_variables_kernel<10>, 10);
        opt_in(tract_variables_kernel<11>, 11);
        opt_in(tract_variables_kernel<12>, 12);
        opt_in(tract_variables_kernel<13>, 13);
        opt_in(tract_variables_kernel<14>, 14);
        return asb::check_cuda(e, "asb_tract_variables: smem attribute");
      }))
    return rc;
  const int per_group = nw * kTeamsPerWarp;
  long long blocks = (sets + per_group - 1) / per_group;
  if (blocks > sms) blocks = sms;   // persistent: one CTA per SM, the groups strided over whatever grid there is
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define ASB_TV_LAUNCH(NW)                                                                                       \
  case NW:                                                                                                      \
    tract_variables_kernel<NW><<<unsigned(blocks), (NW + 1) * 32, smem, st>>>(sets_ptr, sets, n_art, sg, uinc,  \
                                                                               uinc_stride, val, idx, poc);     \
    break;
  switch (nw) {
    ASB_TV_LAUNCH(8)
    ASB_TV_LAUNCH(9)
    ASB_TV_LAUNCH(10)
    ASB_TV_LAUNCH(11)
    ASB_TV_LAUNCH(12)
    ASB_TV_LAUNCH(13)
    ASB_TV_LAUNCH(14)
  }
#undef ASB_TV_LAUNCH
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
