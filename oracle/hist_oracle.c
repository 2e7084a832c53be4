/*
 * hist_oracle.c — CPU ORACLE (test infrastructure, not product; see hist_oracle.h for the rules and the
 * "PARITY UNPINNED" statement).  Plain C11, built with -O2 -ffp-contract=off so that no fused multiply-add
 * changes an fp64/fp32 result relative to the device code (which uses explicit *_rn intrinsics).
 *
 * Citations of the form [H2O-3 Class.method] name the unit restated, recalled from H2O-3's public source
 * (hex.tree, 3.30-era); that source is not in /root/reference (only src/k8s/templates.rs:29-30 names the jar)
 * so the statements are unverified and this file is the normative definition for parity.
 */
#include "hist_oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------------------------------------
 * Synthetic generator — SURVEY.md §8(d).  Integer hash + a handful of individually rounded fp32 ops;
 * no libm, so host and device agree bit for bit.
 * ---------------------------------------------------------------------------------------------- */
uint64_t h2ok_oracle_splitmix64(uint64_t z) {
  z += 0x9E3779B97F4A7C15ULL;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
  return z ^ (z >> 31);
}
static inline uint64_t cell_hash(uint64_t seed, uint64_t col, uint64_t row) {
  return h2ok_oracle_splitmix64(seed ^ (col * 0x9E3779B97F4A7C15ULL) ^ row);
}
static inline float u24(uint64_t h) { return (float)(h >> 40) * 0x1p-24f; }

/* 28-column HIGGS-shaped recipe: b-tag-like 3-valued columns at 8/12/16/20, high-level columns 21..27,
 * the rest cycle pT-like / eta-like / phi-like. */
static float base28(int j, float u) {
  if (j == 8 || j == 12 || j == 16 || j == 20) return u < 0.5f ? 0.0f : (u < 0.8f ? 1.0865f : 2.173f);
  if (j >= 21) {
    float u2 = u * u;
    float u3 = u2 * u;
    float t = 3.0f * u3;
    return 0.3f + t;
  }
  switch (j % 3) {
    case 0: {
      float u2 = u * u;
      float t = 4.0f * u2;
      return 0.25f + t;
    }
    case 1: {
      float t = 5.0f * u;
      return t - 2.5f;
    }
    default: {
      float t = 6.28318548f * u;
      return t - 3.14159274f;
    }
  }
}

float h2ok_oracle_gen_x(int32_t shape, uint64_t seed, int64_t row, int32_t col, int32_t na_permille) {
  uint64_t h = cell_hash(seed, (uint64_t)col, (uint64_t)row);
  if (na_permille > 0) {
    uint64_t h2 = cell_hash(seed ^ 0xA5A5A5A5A5A5A5A5ULL, (uint64_t)col, (uint64_t)row);
    if ((int32_t)(h2 % 1000ULL) < na_permille) return NAN;
  }
  float u = u24(h);
  if (shape == H2OK_SHAPE_SPARSE) return ((h & 0xFFFFULL) < 58982ULL) ? 0.0f : u;
  float b = base28(col % 28, u);
  if (shape == H2OK_SHAPE_WIDE64) {
    float s = 1.0f + (float)col / 64.0f;
    float t = b * s;
    return t + (float)col;
  }
  return b;
}

/* y depends on the base uniforms of a few X columns so that real splits exist at every level. */
float h2ok_oracle_gen_y(uint64_t seed, int64_t row, int32_t cols) {
  float uy = u24(cell_hash(seed + 1, 0, (uint64_t)row));
  float u0 = u24(cell_hash(seed, (uint64_t)(0 % cols), (uint64_t)row));
  float u1 = u24(cell_hash(seed, (uint64_t)(1 % cols), (uint64_t)row));
  float u3 = u24(cell_hash(seed, (uint64_t)(3 % cols), (uint64_t)row));
  float u5 = u24(cell_hash(seed, (uint64_t)(5 % cols), (uint64_t)row));
  float t = 2.0f * uy;
  float noise = 0.5f * (t - 1.0f);
  float a = u0 - 0.5f;
  float b = u3 - 0.5f;
  if (u5 < 0.5f) b = -b;
  float c = u1 > 0.7f ? 0.5f : 0.0f;
  float s = noise + a;
  s = s + b;
  return s + c;
}

float h2ok_oracle_gen_w(uint64_t seed, int64_t row, int32_t weight_mode) {
  if (weight_mode == 0) return 1.0f;
  uint64_t h = cell_hash(seed + 2, 0, (uint64_t)row);
  if ((h & 0xFFFFULL) < 1311ULL) return 0.0f; /* 2 % zero weights */
  return 0.5f + u24(h);
}

void h2ok_oracle_synth_fill(int32_t shape, uint64_t seed, int64_t row0, int64_t rows, int32_t cols, int64_t ld,
                            int32_t na_permille, int32_t weight_mode, float* X, float* y, float* w) {
  for (int32_t c = 0; c < cols; c++)
    for (int64_t r = 0; r < rows; r++) X[(int64_t)c * ld + r] = h2ok_oracle_gen_x(shape, seed, row0 + r, c, na_permille);
  for (int64_t r = 0; r < rows; r++) {
    if (y) y[r] = h2ok_oracle_gen_y(seed, row0 + r, cols);
    if (w) w[r] = h2ok_oracle_gen_w(seed, row0 + r, weight_mode);
  }
}

/* ------------------------------------------------------------------------------------------------
 * a1 — DHistogram constructor.  [H2O-3 DHistogram.<init>, DHistogram.find_maxEx, DHistogram.initialHist]
 * ---------------------------------------------------------------------------------------------- */
static double ulp_d(double x) { /* java.lang.Math.ulp(double) */
  x = fabs(x);
  if (isinf(x)) return INFINITY;
  if (x == DBL_MAX) return x - nextafter(x, 0.0);
  return nextafter(x, INFINITY) - x;
}

double h2ok_oracle_find_maxex(double maxin, int32_t is_int) {
  double ulp = ulp_d(maxin);
  if (is_int > 0 && 1 > ulp) ulp = 1;
  double res = maxin + ulp;
  return isinf(res) ? maxin : res;
}

/* returns 0 and sets nbin=0 when no histogram can be made for the range */
int32_t h2ok_oracle_make_desc(double min, double maxex, int32_t is_int, int32_t nbins, h2ok_col_desc* out) {
  out->min = min;
  out->maxex = maxex;
  out->is_int = is_int;
  out->nbin = 0;
  out->step = 0.0;
  if (!(maxex > min)) return 0;
  int32_t xbins = nbins;
  double step;
  if (is_int > 0 && maxex - min <= (double)xbins) { /* fewer unique values than bins: one bin per integer */
    xbins = (int32_t)((long long)maxex - (long long)min);
    step = 1.0;
  } else {
    step = (double)xbins / (maxex - min);
    if (!(step > 0.0) || isinf(step)) return 0;
  }
  if (xbins < 1) return 0;
  out->step = step;
  out->nbin = xbins;
  return 1;
}

void h2ok_oracle_col_ranges(const float* X, int64_t ld, int64_t rows, int32_t cols, float* mn, float* mx) {
  for (int32_t c = 0; c < cols; c++) {
    float lo = INFINITY, hi = -INFINITY;
    const float* x = X + (int64_t)c * ld;
    for (int64_t r = 0; r < rows; r++) {
      float v = x[r];
      if (v < lo) lo = v;
      if (v > hi) hi = v;
    }
    mn[c] = lo + 0.0f;
    mx[c] = hi + 0.0f;
  }
}

/* [H2O-3 DHistogram.initialHist]: constant / all-NA columns get no histogram. */
static void root_desc_from_range(float lo, float hi, int32_t is_int, int32_t nbins_top, h2ok_col_desc* out) {
  memset(out, 0, sizeof(*out));
  out->is_int = is_int;
  if (!(lo < hi) || isinf(lo) || isinf(hi)) return;
  h2ok_oracle_make_desc((double)lo, h2ok_oracle_find_maxex((double)hi, is_int), is_int, nbins_top, out);
}

int32_t h2ok_oracle_root_descs(const float* X, int64_t ld, int64_t rows, int32_t cols, const h2ok_params* p,
                               const int32_t* is_int, h2ok_col_desc* out) {
  float* mn = (float*)malloc(sizeof(float) * (size_t)cols * 2);
  if (!mn) return H2OK_ERR_ARG;
  float* mx = mn + cols;
  h2ok_oracle_col_ranges(X, ld, rows, cols, mn, mx);
  for (int32_t c = 0; c < cols; c++) root_desc_from_range(mn[c], mx[c], is_int ? is_int[c] : 0, p->nbins_top_level, &out[c]);
  free(mn);
  return H2OK_OK;
}

/* ------------------------------------------------------------------------------------------------
 * a2 — [H2O-3 DHistogram.bin]: NaN → NA bucket (slot nbin); ±Inf → end bins; else
 * idx = (int)((x - min) * step) in fp64, with the round-off clamp `if (idx == nbin) idx--`.  H2O asserts
 * 0 <= idx < nbin; here out-of-range values are clamped to the end bins, which equals H2O wherever its
 * assertion holds.
 * ---------------------------------------------------------------------------------------------- */
int32_t h2ok_oracle_bin(float x, const h2ok_col_desc* d) {
  if (isnan(x)) return d->nbin;
  double xd = (double)x;
  if (isinf(xd)) return xd < 0 ? 0 : d->nbin - 1;
  double diff = xd - d->min;
  double pos = diff * d->step;
  if (pos < 0.0) return 0;
  if (!(pos < (double)d->nbin)) return d->nbin - 1;
  return (int32_t)pos;
}

void h2ok_oracle_bin_many(const float* x, int64_t n, const h2ok_col_desc* d, int32_t* bins) {
  for (int64_t i = 0; i < n; i++) bins[i] = h2ok_oracle_bin(x[i], d);
}

/* ------------------------------------------------------------------------------------------------
 * Fixed-point targets (no H2O counterpart: H2O adds doubles; the device adds integers so that sums are
 * order-independent and the cross-GPU reduce is exact).
 * ---------------------------------------------------------------------------------------------- */
void h2ok_oracle_target_absmax(const float* y, const float* w, int64_t rows, double* m) {
  m[0] = m[1] = m[2] = 0.0;
  for (int64_t r = 0; r < rows; r++) {
    double wd = w ? (double)w[r] : 1.0, yd = (double)y[r];
    double wy = wd * yd;
    double wyy = wy * yd;
    if (fabs(wd) > m[0]) m[0] = fabs(wd);
    if (fabs(wy) > m[1]) m[1] = fabs(wy);
    if (fabs(wyy) > m[2]) m[2] = fabs(wyy);
  }
}

static double pow2_scale(double absmax, int32_t qbits) {
  if (!(absmax > 0.0) || isinf(absmax)) return 1.0;
  int ex;
  frexp(absmax, &ex); /* absmax = m * 2^ex, m in [0.5,1) → absmax < 2^ex */
  return ldexp(1.0, qbits - ex);
}

void h2ok_oracle_choose_quant(const double* m, int32_t weighted, int32_t qbits, h2ok_quant* q) {
  q->weighted = weighted;
  q->qbits = qbits;
  q->scale_w = weighted ? pow2_scale(m[0], qbits) : 1.0;
  q->scale_wy = pow2_scale(m[1], qbits);
  q->scale_wyy = pow2_scale(m[2], qbits);
}

void h2ok_oracle_quantize_rows(const float* y, const float* w, int64_t rows, const h2ok_quant* q, int32_t* wq,
                               int32_t* wyq, int32_t* wyyq) {
  for (int64_t r = 0; r < rows; r++) {
    double wd = (q->weighted && w) ? (double)w[r] : 1.0, yd = (double)y[r];
    double wy = wd * yd;
    double wyy = wy * yd;
    if (q->weighted && wq) wq[r] = (int32_t)nearbyint(wd * q->scale_w);
    wyq[r] = (int32_t)nearbyint(wy * q->scale_wy);
    wyyq[r] = (int32_t)nearbyint(wyy * q->scale_wyy);
  }
}

/* ------------------------------------------------------------------------------------------------
 * a3 / a4 — [H2O-3 DHistogram.updateHisto via ScoreBuildHistogram2 local phase]: for every active row of a
 * node and every tracked column: skip zero weight; track observed min/max; b = bin(x);
 * vals[3b] += w; vals[3b+1] += w*y; vals[3b+2] += (w*y)*y, all in fp64.
 * ---------------------------------------------------------------------------------------------- */
static void build_one_col(const float* x, int64_t rows, const float* y, const float* w, const int32_t* nid,
                          int32_t n_nodes, const h2ok_col_desc* descs, int32_t cols, int32_t c, int32_t nb1,
                          double* hist, double* abs_wy, float* min2, float* maxin) {
  for (int64_t r = 0; r < rows; r++) {
    int32_t g = nid[r];
    if (g < 0 || g >= n_nodes) continue;
    const h2ok_col_desc* d = &descs[(int64_t)g * cols + c];
    if (d->nbin <= 0) continue;
    double wd = w ? (double)w[r] : 1.0;
    if (wd == 0.0) continue;
    float xv = x[r];
    int64_t nc = (int64_t)g * cols + c;
    if (xv < min2[nc]) min2[nc] = xv;
    if (xv > maxin[nc]) maxin[nc] = xv;
    double yd = (double)y[r];
    double wy = wd * yd;
    double wyy = wy * yd;
    int32_t b = h2ok_oracle_bin(xv, d);
    double* v = hist + (nc * nb1 + b) * 3;
    v[0] += wd;
    v[1] += wy;
    v[2] += wyy;
    if (abs_wy) abs_wy[nc * nb1 + b] += fabs(wy);
  }
}

void h2ok_oracle_build_level(const float* X, int64_t ld, int64_t rows, int32_t cols, const float* y, const float* w,
                             const int32_t* nid, int32_t n_nodes, const h2ok_col_desc* descs, int32_t max_bins,
                             double* hist, double* abs_wy, float* min2, float* maxin, int32_t threads) {
  int32_t nb1 = max_bins + 1;
  int64_t ncell = (int64_t)n_nodes * cols;
  memset(hist, 0, sizeof(double) * (size_t)(ncell * nb1 * 3));
  if (abs_wy) memset(abs_wy, 0, sizeof(double) * (size_t)(ncell * nb1));
  for (int64_t i = 0; i < ncell; i++) {
    min2[i] = INFINITY;
    maxin[i] = -INFINITY;
  }
  /* LocalMR over columns is H2O's own parallel shape (ScoreBuildHistogram2); columns are independent. */
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) num_threads(threads > 1 ? threads : 1)
#endif
  for (int32_t c = 0; c < cols; c++)
    build_one_col(X + (int64_t)c * ld, rows, y, w, nid, n_nodes, descs, cols, c, nb1, hist, abs_wy, min2, maxin);
  (void)threads;
  for (int64_t i = 0; i < ncell; i++) {
    min2[i] += 0.0f; /* canonical +0 */
    maxin[i] += 0.0f;
  }
}

void h2ok_oracle_build_level_q(const float* X, int64_t ld, int64_t rows, int32_t cols, const int32_t* nid,
                               const int32_t* wq, const int32_t* wyq, const int32_t* wyyq, int32_t n_nodes,
                               const h2ok_col_desc* descs, int32_t max_bins, int64_t* acc, float* min2, float* maxin) {
  int32_t nb1 = max_bins + 1;
  int64_t ncell = (int64_t)n_nodes * cols;
  memset(acc, 0, sizeof(int64_t) * (size_t)(ncell * nb1 * 3));
  for (int64_t i = 0; i < ncell; i++) {
    min2[i] = INFINITY;
    maxin[i] = -INFINITY;
  }
  for (int32_t c = 0; c < cols; c++) {
    const float* x = X + (int64_t)c * ld;
    for (int64_t r = 0; r < rows; r++) {
      int32_t g = nid[r];
      if (g < 0 || g >= n_nodes) continue;
      const h2ok_col_desc* d = &descs[(int64_t)g * cols + c];
      if (d->nbin <= 0) continue;
      if (wq && wq[r] == 0) continue;
      float xv = x[r];
      int64_t nc = (int64_t)g * cols + c;
      if (xv < min2[nc]) min2[nc] = xv;
      if (xv > maxin[nc]) maxin[nc] = xv;
      int64_t* v = acc + (nc * nb1 + h2ok_oracle_bin(xv, d)) * 3;
      v[0] += wq ? wq[r] : 1;
      v[1] += wyq[r];
      v[2] += wyyq[r];
    }
  }
  for (int64_t i = 0; i < ncell; i++) {
    min2[i] += 0.0f;
    maxin[i] += 0.0f;
  }
}

void h2ok_oracle_finalize(const int64_t* acc, int64_t n, int32_t max_bins, const h2ok_quant* q, double* hist) {
  (void)max_bins;
  double inv[3] = {1.0 / q->scale_w, 1.0 / q->scale_wy, 1.0 / q->scale_wyy};
  for (int64_t i = 0; i < n; i++) hist[i] = (double)acc[i] * inv[i % 3];
}

/* ------------------------------------------------------------------------------------------------
 * a7 — [H2O-3 DTree.findBestSplitPoint] for numeric columns, then [UndecidedNode.bestCol].
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t found, bin, na_dir;
  double se_before, se_l, se_r, n_l, n_r, p_l, p_r;
} col_split;

static int iabs(int a) { return a < 0 ? -a : a; }

static col_split best_split_point(const double* vals, int32_t nbins, double min_rows, double msi, double* scratch) {
  col_split out;
  memset(&out, 0, sizeof(out));
  if (nbins <= 1) return out; /* bestCol skips hs[i].nbins() <= 1 */
  double* wlo = scratch;
  double* wYlo = wlo + (nbins + 1);
  double* wYYlo = wYlo + (nbins + 1);
  double* whi = wYYlo + (nbins + 1);
  double* wYhi = whi + (nbins + 1);
  double* wYYhi = wYhi + (nbins + 1);
  wlo[0] = wYlo[0] = wYYlo[0] = 0.0;
  for (int b = 0; b < nbins; b++) {
    wlo[b + 1] = wlo[b] + vals[3 * b];
    wYlo[b + 1] = wYlo[b] + vals[3 * b + 1];
    wYYlo[b + 1] = wYYlo[b] + vals[3 * b + 2];
  }
  double wNA = vals[3 * nbins], wYNA = vals[3 * nbins + 1], wYYNA = vals[3 * nbins + 2];
  double tot = wlo[nbins] + wNA;
  if (tot < 2 * min_rows) return out;
  double sy = wYlo[nbins] + wYNA;
  double var = (wYYlo[nbins] + wYYNA) * tot - sy * sy;
  if ((float)var == 0.0f) return out;
  whi[nbins] = wYhi[nbins] = wYYhi[nbins] = 0.0;
  for (int b = nbins - 1; b >= 0; b--) {
    whi[b] = whi[b + 1] + vals[3 * b];
    wYhi[b] = wYhi[b + 1] + vals[3 * b + 1];
    wYYhi[b] = wYYhi[b + 1] + vals[3 * b + 2];
  }
  double best_seL = DBL_MAX, best_seR = DBL_MAX;
  int nasplit = H2OK_NA_NONE;
  double seNonNA = wYYhi[0] - wYhi[0] * wYhi[0] / whi[0];
  if (seNonNA < 0) seNonNA = 0;
  double seBefore = seNonNA;
  double nLeft = 0, nRight = 0, predLeft = 0, predRight = 0;
  if (wNA >= min_rows) { /* try NA vs rest */
    double sall = wYhi[0] + wYNA;
    double seAll = (wYYhi[0] + wYYNA) - sall * sall / (whi[0] + wNA);
    double seNA = wYYNA - wYNA * wYNA / wNA;
    if (seNA < 0) seNA = 0;
    best_seL = seNonNA;
    best_seR = seNA;
    nasplit = H2OK_NA_VS_REST;
    seBefore = seAll;
    nLeft = whi[0];
    predLeft = wYhi[0];
    nRight = wNA;
    predRight = wYNA;
  }
  int best = 0;
  int mid = nbins >> 1;
  for (int b = 1; b <= nbins - 1; b++) {
    if (vals[3 * b] == 0) continue;
    if (wlo[b] + wNA < min_rows) continue;
    if (whi[b] + wNA < min_rows) break;
    if (wNA == 0) {
      double selo = wYYlo[b] - wYlo[b] * wYlo[b] / wlo[b];
      double sehi = wYYhi[b] - wYhi[b] * wYhi[b] / whi[b];
      if (selo < 0) selo = 0;
      if (sehi < 0) sehi = 0;
      if ((selo + sehi < best_seL + best_seR) ||
          (selo + sehi == best_seL + best_seR && iabs(b - mid) < iabs(best - mid))) {
        if (wlo[b] >= min_rows && whi[b] >= min_rows) {
          best_seL = selo;
          best_seR = sehi;
          best = b;
          nLeft = wlo[b];
          nRight = whi[b];
          predLeft = wYlo[b];
          predRight = wYhi[b];
          nasplit = H2OK_NA_NONE;
        }
      }
    } else {
      { /* NAs to the left */
        double sl = wYlo[b] + wYNA;
        double selo = wYYlo[b] + wYYNA - sl * sl / (wlo[b] + wNA);
        double sehi = wYYhi[b] - wYhi[b] * wYhi[b] / whi[b];
        if (selo < 0) selo = 0;
        if (sehi < 0) sehi = 0;
        if ((selo + sehi < best_seL + best_seR) ||
            (selo + sehi == best_seL + best_seR && iabs(b - mid) < iabs(best - mid))) {
          if ((wlo[b] + wNA) >= min_rows && whi[b] >= min_rows) {
            best_seL = selo;
            best_seR = sehi;
            best = b;
            nLeft = wlo[b] + wNA;
            nRight = whi[b];
            predLeft = wYlo[b] + wYNA;
            predRight = wYhi[b];
            nasplit = H2OK_NA_LEFT;
          }
        }
      }
      { /* NAs to the right */
        double selo = wYYlo[b] - wYlo[b] * wYlo[b] / wlo[b];
        double sr = wYhi[b] + wYNA;
        double sehi = wYYhi[b] + wYYNA - sr * sr / (whi[b] + wNA);
        if (selo < 0) selo = 0;
        if (sehi < 0) sehi = 0;
        if ((selo + sehi < best_seL + best_seR) ||
            (selo + sehi == best_seL + best_seR && iabs(b - mid) < iabs(best - mid))) {
          if (wlo[b] >= min_rows && (whi[b] + wNA) >= min_rows) {
            best_seL = selo;
            best_seR = sehi;
            best = b;
            nLeft = wlo[b];
            nRight = whi[b] + wNA;
            predLeft = wYlo[b];
            predRight = wYhi[b] + wYNA;
            nasplit = H2OK_NA_RIGHT;
          }
        }
      }
    }
  }
  if (best == 0 && nasplit == H2OK_NA_NONE) return out;
  if (!(best_seL + best_seR < seBefore * (1 - msi))) return out;
  if (nasplit == H2OK_NA_NONE) nasplit = nLeft > nRight ? H2OK_NA_LEFT : H2OK_NA_RIGHT; /* test-time default */
  out.found = 1;
  out.bin = best;
  out.na_dir = nasplit;
  out.se_before = seBefore;
  out.se_l = best_seL;
  out.se_r = best_seR;
  out.n_l = nLeft;
  out.n_r = nRight;
  out.p_l = predLeft / nLeft;
  out.p_r = predRight / nRight;
  return out;
}

/* [H2O-3 DTree.Split.splat]: midpoint between the nearest non-empty bins either side of the split, as a float. */
static float split_splat(const double* vals, const h2ok_col_desc* d, int32_t bin, int32_t na_dir) {
  if (na_dir == H2OK_NA_VS_REST) return -1.0f;
  int x = bin - 1;
  while (x >= 0 && vals[3 * x] == 0) x--;
  int n = bin;
  while (n < d->nbin && vals[3 * n] == 0) n++;
  double lo = d->min + (double)(x + 1) / d->step; /* DHistogram.binAt */
  double hi = d->min + (double)n / d->step;
  if (d->is_int > 0) lo = d->step == 1 ? lo - 1 : floor(lo);
  if (d->is_int > 0) hi = d->step == 1 ? hi : ceil(hi);
  return (float)((lo + hi) / 2.0);
}

/* [H2O-3 DTree.findBestSplitPoint, categorical part, recalled]: a categorical column with one category per bin and
 * at least four bins is scanned in the order of its bins' mean response (idxs sorted by avgs; empty bins carry
 * -Double.MAX_VALUE and come first; the sort is stable), which finds the best two-set partition for a squared-error
 * criterion.  The split then carries a bit set instead of a threshold. */
static int cat_sorted(const h2ok_col_desc* d) { return d->is_int == 2 && d->step == 1.0 && d->nbin >= 4; }

static void sort_bins_by_mean(const double* vals, int32_t nb, int32_t* idx) {
  double* avg = (double*)malloc(sizeof(double) * (size_t)nb);
  for (int32_t b = 0; b < nb; b++) avg[b] = vals[3 * b] == 0 ? -DBL_MAX : vals[3 * b + 1] / vals[3 * b];
  for (int32_t i = 0; i < nb; i++) { /* rank by counting: (mean, bin) ascending */
    int32_t r = 0;
    for (int32_t j = 0; j < nb; j++) r += (avg[j] < avg[i] || (avg[j] == avg[i] && j < i)) ? 1 : 0;
    idx[r < nb ? r : nb - 1] = i;
  }
  free(avg);
}

/* [IcedBitSet of DTree.Split, recalled]: the categories at the sorted positions [bin, nb) go right; a category the node
 * has not seen goes where most of the seen ones went (left on a tie). */
static void cat_bits(const double* vals, const h2ok_col_desc* d, const int32_t* idx, int32_t bin, uint32_t* bits) {
  int nl = 0, nr = 0;
  for (int32_t i = 0; i < d->nbin; i++)
    if (vals[3 * idx[i]] != 0) {
      if (i >= bin) nr++;
      else nl++;
    }
  int empties_left = nl >= nr;
  int off = (int)d->min;
  for (int w = 0; w < H2OK_CAT_WORDS; w++) bits[w] = 0;
  for (int32_t i = 0; i < d->nbin; i++) {
    int right = vals[3 * idx[i]] != 0 ? i >= bin : !empties_left;
    int c = idx[i] + off;
    if (right && c >= 0 && c < 32 * H2OK_CAT_WORDS) bits[c >> 5] |= 1u << (c & 31);
  }
}

void h2ok_oracle_find_splits(const double* hist, const h2ok_col_desc* descs, int32_t n_nodes, int32_t cols,
                             int32_t max_bins, double min_rows, double msi, h2ok_split* splits) {
  int32_t nb1 = max_bins + 1;
  double* scratch = (double*)malloc(sizeof(double) * 6 * (size_t)(nb1 + 1));
  double* sorted = (double*)malloc(sizeof(double) * 3 * (size_t)nb1);
  int32_t* idx = (int32_t*)malloc(sizeof(int32_t) * (size_t)nb1);
  for (int32_t g = 0; g < n_nodes; g++) {
    h2ok_split* s = &splits[g];
    memset(s, 0, sizeof(*s));
    s->col = -1;
    double best_se = 0;
    for (int32_t c = 0; c < cols; c++) {
      const h2ok_col_desc* d = &descs[(int64_t)g * cols + c];
      if (d->nbin <= 1) continue;
      const double* vals = hist + ((int64_t)g * cols + c) * nb1 * 3;
      const int cat = cat_sorted(d);
      const double* scan = vals;
      if (cat) { /* a sorted copy; the NA bucket stays behind the last bin */
        sort_bins_by_mean(vals, d->nbin, idx);
        for (int32_t b = 0; b < d->nbin; b++)
          for (int ch = 0; ch < 3; ch++) sorted[3 * b + ch] = vals[3 * idx[b] + ch];
        for (int ch = 0; ch < 3; ch++) sorted[3 * d->nbin + ch] = vals[3 * d->nbin + ch];
        scan = sorted;
      }
      col_split cs = best_split_point(scan, d->nbin, min_rows, msi, scratch);
      if (!cs.found) continue;
      double se = cs.se_l + cs.se_r;
      if (s->col < 0 || se < best_se) { /* bestCol: strictly smaller se() wins, first column on ties */
        best_se = se;
        s->col = c;
        s->bin = cs.bin;
        s->na_dir = cs.na_dir;
        s->equal = 0;
        memset(s->cat_bits, 0, sizeof(s->cat_bits));
        if (cat && cs.na_dir != H2OK_NA_VS_REST) {
          s->equal = 2;
          s->splat = 0.0f;
          cat_bits(vals, d, idx, cs.bin, s->cat_bits);
        } else {
          s->splat = cat ? -1.0f : split_splat(vals, d, cs.bin, cs.na_dir);
        }
        s->se_before = cs.se_before;
        s->se_left = cs.se_l;
        s->se_right = cs.se_r;
        s->n_left = cs.n_l;
        s->n_right = cs.n_r;
        s->pred_left = cs.p_l;
        s->pred_right = cs.p_r;
      }
    }
  }
  free(idx);
  free(sorted);
  free(scratch);
}

/* ------------------------------------------------------------------------------------------------
 * a8 — [H2O-3 DTree.Split.nextLevelHistos + DecidedNode.<init>]: children ranges from the parent's observed
 * ranges; the split column is cut at splat; columns that became constant are dropped.
 * ---------------------------------------------------------------------------------------------- */
static int equals_within_one_small_ulp(double a, double b) { /* water.util.MathUtils */
  if (isinf(a) || (isinf(b) && (a < b || b < a))) return 0;
  double ua = ulp_d(a), ub = ulp_d(b);
  double su = ua < ub ? ua : ub;
  return fabs(a - b) <= su;
}

void h2ok_oracle_next_level(const h2ok_split* splits, const h2ok_col_desc* descs, const float* min2,
                            const float* maxin, int32_t n_nodes, int32_t cols, const h2ok_params* p, int32_t depth,
                            h2ok_col_desc* child_descs, int32_t* child_ids, int32_t* n_children) {
  (void)depth;
  int32_t nc = 0;
  for (int32_t g = 0; g < n_nodes; g++) {
    const h2ok_split* s = &splits[g];
    for (int way = 0; way < 2; way++) {
      child_ids[2 * g + way] = -1;
      if (s->col < 0) continue;
      double n = way == 0 ? s->n_left : s->n_right;
      if (n < p->min_rows) continue;
      double se = way == 0 ? s->se_left : s->se_right;
      if (se <= 1e-30) continue;
      h2ok_col_desc* out = child_descs + (int64_t)nc * cols;
      int cnt = 0;
      for (int32_t j = 0; j < cols; j++) {
        const h2ok_col_desc* h = &descs[(int64_t)g * cols + j];
        h2ok_col_desc* o = &out[j];
        memset(o, 0, sizeof(*o));
        o->is_int = h->is_int;
        if (h->nbin <= 0) continue;
        int32_t adj_nbins = (h->nbin >> 1) > p->nbins ? (h->nbin >> 1) : p->nbins;
        double mn = (double)min2[(int64_t)g * cols + j];
        double mxin = (double)maxin[(int64_t)g * cols + j];
        if (mxin == mn) continue;
        if (!(mn <= mxin)) continue; /* nothing observed */
        double maxex = h2ok_oracle_find_maxex(mxin, h->is_int);
        if (s->col == j) {
          if (s->na_dir == H2OK_NA_VS_REST) {
            if (way == 1) continue; /* the NA child has nothing but NAs in this column */
          } else if (s->equal >= 2) {
            /* [Split.nextLevelHistos case 2 / 3]: a set of categories each side, no tighter bounds */
          } else {
            double split = (double)s->splat;
            if (h->is_int > 0) split = (double)(float)ceil(split);
            if (way == 0) maxex = split;
            else mn = split;
          }
        }
        if (mn > maxex) continue;
        if (equals_within_one_small_ulp(mn, maxex)) continue;
        if (isinf((double)adj_nbins / (maxex - mn))) continue;
        if (h->is_int > 0 && !(mn + 1 < maxex)) continue;
        if (h2ok_oracle_make_desc(mn, maxex, h->is_int, adj_nbins, o)) cnt++;
      }
      if (cnt == 0) continue;
      child_ids[2 * g + way] = nc++;
    }
  }
  *n_children = nc;
}

/* ------------------------------------------------------------------------------------------------
 * a6 — [H2O-3 ScoreBuildHistogram.score_decide / DTree.DecidedNode.getChildNodeID]: NaN follows na_dir,
 * everything else compares against the float threshold splat.
 * ---------------------------------------------------------------------------------------------- */
void h2ok_oracle_apply_splits(const float* X, int64_t ld, int64_t rows, const h2ok_split* splits,
                              const int32_t* child_ids, int32_t gid_base, int32_t n_nodes, int32_t* nid, int32_t* leaf) {
  for (int64_t r = 0; r < rows; r++) {
    int32_t g = nid[r];
    if (g < 0 || g >= n_nodes) continue;
    const h2ok_split* s = &splits[g];
    if (s->col < 0) {
      nid[r] = -1;
      if (leaf) leaf[r] = (gid_base + g) * 3;
      continue;
    }
    float d = X[(int64_t)s->col * ld + r];
    int way;
    if (isnan(d)) way = (s->na_dir == H2OK_NA_LEFT) ? 0 : 1;
    else if (s->na_dir == H2OK_NA_VS_REST) way = 0;
    else if (s->equal >= 2) { /* [DecidedNode with a bit set]: the row's category; one outside the set's range follows the NAs */
      int c = (int)d;
      if (c < 0 || c >= 32 * H2OK_CAT_WORDS) way = (s->na_dir == H2OK_NA_LEFT) ? 0 : 1;
      else way = (int)((s->cat_bits[c >> 5] >> (c & 31)) & 1u);
    } else way = d >= s->splat ? 1 : 0;
    int32_t ch = child_ids[2 * g + way];
    nid[r] = ch;
    if (ch < 0 && leaf) leaf[r] = (gid_base + g) * 3 + 1 + way;
  }
}

/* ------------------------------------------------------------------------------------------------
 * a9 — the tree-level caller.  [H2O-3 GBM.ComputePredAndRes / GammaPass / AddTreeContributions, recalled]
 * Residual r = y - mu(f); leaf value gamma = sum(w r) / sum(w mu (1 - mu)) (bernoulli) or sum(w r) / sum(w)
 * (gaussian); f += learn_rate * gamma.  The leaf sums run over the same fixed-point integers the device uses.
 * ---------------------------------------------------------------------------------------------- */
void h2ok_oracle_gbm_residual(int32_t dist, const float* label, const float* f, const float* w, int64_t rows,
                              double den_scale, float* res, int32_t* denq) {
  for (int64_t r = 0; r < rows; r++) {
    float fi = f[r];
    if (dist == H2OK_DIST_BERNOULLI) {
      float e = expf(-fi); /* libm: agrees with the device's expf to a few ulp, not bit for bit */
      float den = 1.0f + e;
      float mu = 1.0f / den;
      res[r] = label[r] - mu;
      double wd = w ? (double)w[r] : 1.0;
      float om = 1.0f - mu;
      double v = (wd * (double)mu) * (double)om;
      if (denq) denq[r] = (int32_t)nearbyint(v * den_scale);
    } else {
      res[r] = label[r] - fi;
      if (denq) denq[r] = w ? (int32_t)nearbyint((double)w[r] * den_scale) : 1;
    }
  }
}

/* a9, K classes: GBM.java ComputePredAndRes for multinomial — softmax of the K scores (fp32, max-subtracted, summed in
 * class order), residual (label == k) - p_k, GammaPass denominator w * |res| * (1 - |res|) quantised. */
void h2ok_oracle_gbm_residual_multi(int32_t K, const float* label, const float* f, int64_t ld, const float* w, int64_t rows,
                                    double den_scale, float* res, int32_t* denq) {
  for (int64_t r = 0; r < rows; r++) {
    float m = f[r];
    for (int k = 1; k < K; k++) m = fmaxf(m, f[(int64_t)k * ld + r]);
    float s = 0.0f;
    for (int k = 0; k < K; k++) {
      float e = expf(f[(int64_t)k * ld + r] - m); /* libm: a few ulp from the device's expf */
      s = s + e;
    }
    int cls = (int)label[r];
    double wd = w ? (double)w[r] : 1.0;
    for (int k = 0; k < K; k++) {
      float e = expf(f[(int64_t)k * ld + r] - m);
      float p = e / s;
      float rk = (k == cls ? 1.0f : 0.0f) - p;
      res[(int64_t)k * ld + r] = rk;
      float a = fabsf(rk);
      float om = 1.0f - a;
      double v = (wd * (double)a) * (double)om;
      denq[(int64_t)k * ld + r] = (int32_t)nearbyint(v * den_scale);
    }
  }
}

void h2ok_oracle_gbm_gamma_pass(const int32_t* leaf, const int32_t* numq, const int32_t* denq, int64_t rows,
                                int32_t n_codes, int64_t* num_acc, int64_t* den_acc) {
  memset(num_acc, 0, sizeof(int64_t) * (size_t)n_codes);
  memset(den_acc, 0, sizeof(int64_t) * (size_t)n_codes);
  for (int64_t r = 0; r < rows; r++) {
    int32_t c = leaf[r];
    if (c < 0 || c >= n_codes) continue;
    num_acc[c] += numq[r];
    den_acc[c] += denq ? denq[r] : 1;
  }
}

void h2ok_oracle_gbm_add_gamma(const int32_t* leaf, int64_t rows, const double* gamma, int32_t n_codes, double learn_rate,
                               float* f) {
  for (int64_t r = 0; r < rows; r++) {
    int32_t c = leaf[r];
    if (c < 0 || c >= n_codes) continue;
    double step = learn_rate * gamma[c];
    f[r] = f[r] + (float)step;
  }
}

int32_t h2ok_oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
