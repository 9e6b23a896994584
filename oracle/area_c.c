/* Plain-C fp64 restatement of the reference's area-function chain for a batch of frames
 * (area_function.py:124-223: intersect_semipolar_grid -> area_function -> evenly_spaced_fx).
 *
 * TEST INFRASTRUCTURE ONLY: compiled by oracle/build_c.py (gcc -O2 -ffp-contract=off) into
 * oracle/_build/libarea_oracle.so and called from tests/, smoke() and bench.py's CPU-baseline
 * legs.  Same definitions as oracle/area.py (see its header for the parity status: the GEOS /
 * vt_tools parts are restated from their published behaviour, PARITY UNPINNED); the two are
 * checked against each other in tests/test_oracle_area.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static double orient(double ax, double ay, double bx, double by, double cx, double cy) {
  return (bx - ax) * (cy - ay) - (by - ay) * (cx - ax);
}

static double euclid(double ax, double ay, double bx, double by) {
  const double dx = ax - bx, dy = ay - by;
  return sqrt(dx * dx + dy * dy);
}

/* segment p0p1 (grid line) against q0q1 (wall segment); 1 = one shared point written to out */
static int seg_x(const double* p0, const double* p1, const double* q0, const double* q1, double* out) {
  const double d1 = orient(p0[0], p0[1], p1[0], p1[1], q0[0], q0[1]);
  const double d2 = orient(p0[0], p0[1], p1[0], p1[1], q1[0], q1[1]);
  if ((d1 > 0 && d2 > 0) || (d1 < 0 && d2 < 0)) return 0;
  const double d3 = orient(q0[0], q0[1], q1[0], q1[1], p0[0], p0[1]);
  const double d4 = orient(q0[0], q0[1], q1[0], q1[1], p1[0], p1[1]);
  if ((d3 > 0 && d4 > 0) || (d3 < 0 && d4 < 0)) return 0;
  if (d1 == 0 && d2 == 0 && d3 == 0 && d4 == 0) return 0; /* collinear overlap: contributes nothing */
  if (d1 == 0) { out[0] = q0[0]; out[1] = q0[1]; return 1; }
  if (d2 == 0) { out[0] = q1[0]; out[1] = q1[1]; return 1; }
  if (d3 == 0) { out[0] = p0[0]; out[1] = p0[1]; return 1; }
  if (d4 == 0) { out[0] = p1[0]; out[1] = p1[1]; return 1; }
  const double t = d1 / (d1 - d2);
  out[0] = q0[0] + t * (q1[0] - q0[0]);
  out[1] = q0[1] + t * (q1[1] - q0[1]);
  return 1;
}

/* all intersections of one grid line with one wall, in wall-segment order, exact duplicates removed */
static int line_wall(const double* g0, const double* g1, const double* wall, int nw, double* pts) {
  int n = 0;
  for (int k = 0; k + 1 < nw; ++k) {
    double p[2];
    if (!seg_x(g0, g1, wall + 2 * k, wall + 2 * k + 2, p)) continue;
    int dup = 0;
    for (int q = 0; q < n; ++q) dup |= (pts[2 * q] == p[0] && pts[2 * q + 1] == p[1]);
    if (!dup) { pts[2 * n] = p[0]; pts[2 * n + 1] = p[1]; ++n; }
  }
  return n;
}

/* argmatrix(distance_matrix(A, B), "min"): first minimum in row-major order */
static void argmin_rows(const double* A, int na, const double* B, int nb, int* i_min, int* j_min) {
  double best = INFINITY;
  *i_min = 0; *j_min = 0;
  for (int i = 0; i < na; ++i)
    for (int j = 0; j < nb; ++j) {
      const double d = euclid(A[2 * i], A[2 * i + 1], B[2 * j], B[2 * j + 1]);
      if (d < best) { best = d; *i_min = i; *j_min = j; }
    }
}

/* One frame.  wi / we: (nw, 2) fp32 walls; ends: (n_lines, 2, 2) fp64 [line][int|ext][x|y].
 * pts: (n_lines, 2, 2) [pair][int|ext][x|y]; lines: (n_lines) grid-line index of each pair;
 * xfx: (2, n_samples).  Returns K. */
static int one_frame(const float* wi32, const float* we32, int nw, const double* ends, int n_lines, double alpha,
                     double beta, int n_samples, double* xfx, double* pts, int32_t* lines, double* scratch) {
  double* wi = scratch;
  double* we = wi + 2 * nw;
  double* li = we + 2 * nw;           /* up to nw-1 intersections each */
  double* le = li + 2 * nw;
  double* dist = le + 2 * nw;         /* n_lines */
  double* fx = dist + n_lines;        /* n_lines */
  for (int k = 0; k < 2 * nw; ++k) { wi[k] = wi32[k]; we[k] = we32[k]; }
  int K = 0;
  for (int l = 0; l < n_lines; ++l) {
    const double* g0 = ends + 4 * l;
    const double* g1 = g0 + 2;
    const int ni = line_wall(g0, g1, wi, nw, li);
    const int ne = line_wall(g0, g1, we, nw, le);
    if (!ni && !ne) continue;                                     /* area_function.py:185-186 */
    const double dflt[4] = {we[0], we[1], we[2 * (nw - 1)], we[2 * (nw - 1) + 1]}; /* :199-202 */
    double* pi = pts + 4 * K;
    double* pe = pi + 2;
    int i_min, j_min;
    if (ni) {                                                     /* :204-212 */
      argmin_rows(li, ni, ne ? le : dflt, ne ? ne : 2, &i_min, &j_min);
      pi[0] = li[2 * i_min]; pi[1] = li[2 * i_min + 1];
      if (!ne) { const int v = j_min ? nw - 1 : 0; pe[0] = we[2 * v]; pe[1] = we[2 * v + 1]; }
    }
    if (ne) {                                                     /* :214-221 */
      argmin_rows(le, ne, ni ? li : dflt, ni ? ni : 2, &i_min, &j_min);
      pe[0] = le[2 * i_min]; pe[1] = le[2 * i_min + 1];
      if (!ni) { const int v = j_min ? nw - 1 : 0; pi[0] = wi[2 * v]; pi[1] = wi[2 * v + 1]; }
    }
    lines[K++] = l;
  }
  for (int k = K; k < n_lines; ++k) { lines[k] = -1; memset(pts + 4 * k, 0, 4 * sizeof(double)); }
  if (K < 2) {
    for (int s = 0; s < 2 * n_samples; ++s) xfx[s] = NAN;
    return K;
  }
  /* area_function, :124-142 */
  double mx_prev = 0, my_prev = 0, d = 0;
  for (int k = 0; k < K; ++k) {
    const double* pi = pts + 4 * k;
    const double* pe = pi + 2;
    const double mx = fmin(pi[0], pe[0]) + fabs(pi[0] - pe[0]) / 2;   /* mid_point :113-121 */
    const double my = fmin(pi[1], pe[1]) + fabs(pi[1] - pe[1]) / 2;
    const double radius = euclid(pi[0], pi[1], pe[0], pe[1]) / 2;
    fx[k] = alpha * (beta == 2.0 ? radius * radius : pow(radius, beta));
    if (k) d = euclid(mx_prev, my_prev, mx, my) + d;
    dist[k] = d;
    mx_prev = mx; my_prev = my;
  }
  /* evenly_spaced_fx, :145-159: np.linspace then the piecewise-linear graph at each abscissa */
  const double x0 = dist[0], x1 = dist[K - 1];
  const double step = (x1 - x0) / (n_samples - 1);
  for (int s = 0; s < n_samples; ++s) {
    const double xv = (s == n_samples - 1) ? x1 : x0 + s * step;
    int hit = -1;
    for (int k = 0; k + 1 < K; ++k)
      if (dist[k] <= xv && xv <= dist[k + 1]) { hit = k; break; }
    double yv;
    if (hit < 0) yv = NAN;
    else if (dist[hit + 1] == dist[hit] || xv == dist[hit]) yv = fx[hit];
    else if (xv == dist[hit + 1]) yv = fx[hit + 1];
    else yv = fx[hit] + (xv - dist[hit]) / (dist[hit + 1] - dist[hit]) * (fx[hit + 1] - fx[hit]);
    xfx[s] = xv;
    xfx[n_samples + s] = yv;
  }
  return K;
}

/* Batch entry.  wi / we: (frames, nw, 2) fp32; outputs: xfx (frames, 2, n_samples) fp64,
 * K (frames) int32, pts (frames, n_lines, 2, 2) fp64 or NULL, lines (frames, n_lines) int32 or NULL.
 * threads > 1 splits the frames over OpenMP threads (frames are independent). */
int area_oracle_batch(const float* wi, const float* we, int64_t frames, int nw, const double* ends, int n_lines,
                      double alpha, double beta, int n_samples, double* xfx, int32_t* K, double* pts,
                      int32_t* lines, int threads) {
  if (nw < 2 || n_lines < 1 || n_samples < 2) return 1;
  if (threads < 1) threads = 1;
#pragma omp parallel num_threads(threads)
  {
    double* scratch = (double*)malloc(sizeof(double) * (8 * (size_t)nw + 2 * (size_t)n_lines));
    double* my_pts = (double*)malloc(sizeof(double) * 4 * (size_t)n_lines);
    int32_t* my_lines = (int32_t*)malloc(sizeof(int32_t) * (size_t)n_lines);
#pragma omp for schedule(static)
    for (int64_t f = 0; f < frames; ++f) {
      K[f] = one_frame(wi + f * 2 * nw, we + f * 2 * nw, nw, ends, n_lines, alpha, beta, n_samples,
                       xfx + f * 2 * n_samples, my_pts, my_lines, scratch);
      if (pts) memcpy(pts + f * 4 * n_lines, my_pts, sizeof(double) * 4 * (size_t)n_lines);
      if (lines) memcpy(lines + f * n_lines, my_lines, sizeof(int32_t) * (size_t)n_lines);
    }
    free(scratch); free(my_pts); free(my_lines);
  }
  return 0;
}
