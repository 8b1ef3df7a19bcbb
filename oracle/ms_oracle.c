/*
 * oracle/ms_oracle.c  --  TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * Plain-C, CPU-only restatement of magicsoup's per-step world math
 * (reference v1.0.5 under /root/reference):
 *
 *   Kinetics.integrate_signals and helpers   python/magicsoup/kinetics.py:725-918
 *   World.enzymatic_activity                 python/magicsoup/world.py:610-625
 *   World.diffuse_molecules                  python/magicsoup/world.py:627-665
 *   World._get_diffuse / _get_permeate       python/magicsoup/world.py:935-984 (host side, in oracle.py)
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; the product path (magicsoup_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py replays every call the
 * reference's own hot-path tests make (tests/fast/test_kinetics.py:955-2234,
 * tests/fast/test_world.py:8-51), recorded from the real reference by
 * tests/golden/make_golden.py, plus seeded synthetic populations run through the
 * real reference, against this file.
 *
 * The reference evaluates with dense aten ops whose float32 reduction order and
 * powf are implementation details of torch (not under /root/reference; pinned
 * torch>=2.0,<2.2 in pyproject.toml:14).  This restatement fixes the arithmetic
 * the CUDA kernels are held to bit-for-bit ("spec order"):
 *   - every sum over proteins / signals runs sequentially in ascending index
 *     order in `real` precision, starting from +0;
 *   - x^n for integer n is binary exponentiation by repeated multiplication
 *     (powi below), x^-n = 1 / x^n;
 *   - no fused multiply-add (compile with -ffp-contract=off);
 *   - the per-plane mass sums of diffuse_molecules are accumulated exactly as
 *     2^-32 fixed point integers (order independent) and rounded once to float32,
 *     standing in for torch.sum's float32 tree.
 * Compile with -DREAL=double for the float64 conditioning reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef REAL
#define REAL float
#endif
typedef REAL real;
typedef int64_t i64;
typedef int32_t i32;

#define IS_F32 (sizeof(real) == 4)
/* kinetics.py:11-13 : _EPS, _MAX, _MIN; python floats are cast to the tensor dtype */
#define V_EPS (IS_F32 ? (real)1e-36f : (real)1e-36)
#define V_MAX (IS_F32 ? (real)1e36f : (real)1e36)
#define V_MIN (-V_MAX)

int ms_oracle_real_bytes(void) { return (int)sizeof(real); }

static int n_threads_or_default(int n) {
#ifdef _OPENMP
  return n > 0 ? n : omp_get_max_threads();
#else
  (void)n;
  return 1;
#endif
}

/* x^n, n > 0, by right-to-left binary exponentiation (spec order) */
static real powi(real x, i32 n) {
  real r = (real)1, b = x;
  while (n) {
    if (n & 1) r = r * b;
    n >>= 1;
    if (n) b = b * b;
  }
  return r;
}

/* kinetics.py:894-918 _multiply_signals for one protein row */
static real mulsig(const real *X, const i32 *K, i64 s, int *involved) {
  real xx = (real)1;
  int any = 0;
  for (i64 i = 0; i < s; i++) {
    i32 k = K[i];
    if (k > 0) { /* M = N > 0 ; x = X ; pow(x, k) */
      any = 1;
      xx = xx * powi(X[i], k);
    } else if (k < 0) { /* x = 0 ; pow(0, k<0) = +inf */
      xx = xx * (real)INFINITY;
    } /* k == 0: pow(0, 0) = 1 */
  }
  if (isnan(xx)) xx = (real)0;  /* :913 */
  if (xx < (real)0) xx = (real)0; /* :914 */
  if (isinf(xx)) xx = V_MAX;    /* :915 */
  *involved = any;              /* :918 */
  return xx;
}

/* kinetics.py:881-892 _get_quotient for one protein */
static real quotient(const real *X, const i32 *Nf, const i32 *Nb, i64 s) {
  int inv_b, inv_f;
  real xb = mulsig(X, Nb, s, &inv_b);
  if (!inv_b) xb = (real)0;
  real xf = mulsig(X, Nf, s, &inv_f);
  if (!inv_f) xf = (real)0;
  real q = xb / xf; /* x/0=inf, 0/0=nan */
  if (q < V_EPS) q = V_EPS; /* clamp passes NaN */
  if (q > V_MAX) q = V_MAX;
  if (isnan(q)) q = (real)1; /* nan_to_num(1.0) */
  return q;
}

/* kinetics.py:771-806 _get_velocities for one protein */
static real velocity(const real *X, const i32 *Nf, const i32 *Nb, const i32 *A,
                     const real *Kmr, real Kmf, real Kmb, real vmax, i64 s) {
  int inv;
  real kf = mulsig(X, Nf, s, &inv) / Kmf;
  if (!inv) kf = (real)0;
  if (isinf(kf)) kf = V_MAX;
  real kb = mulsig(X, Nb, s, &inv) / Kmb;
  if (!inv) kb = (real)0;
  if (isinf(kb)) kb = V_MAX;
  real a_cat = (kf - kb) / (((real)1 + kf) + kb); /* :786 */

  real a_reg = (real)1;
  for (i64 i = 0; i < s; i++) { /* :793-799 */
    i32 a = A[i];
    if (a == 0) continue;
    real t = a > 0 ? powi(X[i], a) : (real)1 / powi(X[i], -a);
    t = t / (t + Kmr[i]);
    if (isnan(t)) t = (real)1;
    a_reg = a_reg * t;
  }
  if (isinf(a_reg)) a_reg = V_MAX; /* :800 */
  real v = (a_cat * vmax) * a_reg;  /* :803 */
  if (v < V_MIN) v = V_MIN;         /* :806 clamp passes NaN */
  if (v > V_MAX) v = V_MAX;
  return v;
}

/* min with torch semantics (NaN propagates) */
static real nanmin(real a, real b) {
  if (isnan(a)) return a;
  if (isnan(b)) return b;
  return b < a ? b : a;
}

/* ------------------------------------------------------------------ */
/* staged entry points, dense tensors exactly as the reference's       */
/* ------------------------------------------------------------------ */

/* _multiply_signals(X, N) -> xx (c,p), involved (c,p) as bytes */
void ms_oracle_multiply_signals(const real *X, const i32 *K, i64 c, i64 p, i64 s,
                                real *xx, uint8_t *involved) {
  for (i64 ci = 0; ci < c; ci++)
    for (i64 j = 0; j < p; j++) {
      int inv;
      xx[ci * p + j] = mulsig(X + ci * s, K + (ci * p + j) * s, s, &inv);
      involved[ci * p + j] = (uint8_t)inv;
    }
}

void ms_oracle_quotient(const real *X, const i32 *Nf, const i32 *Nb, i64 c, i64 p,
                        i64 s, real *Q) {
  for (i64 ci = 0; ci < c; ci++)
    for (i64 j = 0; j < p; j++)
      Q[ci * p + j] = quotient(X + ci * s, Nf + (ci * p + j) * s, Nb + (ci * p + j) * s, s);
}

void ms_oracle_velocities(const real *X, const real *vmax, const i32 *Nf, const i32 *Nb,
                          const i32 *A, const real *Kmr, const real *Kmf,
                          const real *Kmb, i64 c, i64 p, i64 s, real *V) {
  for (i64 ci = 0; ci < c; ci++)
    for (i64 j = 0; j < p; j++) {
      i64 r = ci * p + j;
      V[r] = velocity(X + ci * s, Nf + r * s, Nb + r * s, A + r * s, Kmr + r * s, Kmf[r],
                      Kmb[r], vmax[r], s);
    }
}

/* kinetics.py:861-879, one cell; NV (p,s) dense -> Fmin (p) */
static void neg_adjust_cell(const real *NV, const real *X, i64 p, i64 s, real *Fmin,
                            real *Fsig /* scratch (s) */) {
  for (i64 i = 0; i < s; i++) {
    real d = (real)0;
    for (i64 j = 0; j < p; j++) {
      real t = -NV[j * s + i];
      if (t < (real)0) t = (real)0; /* clamp(min=0) passes NaN */
      d = d + t;
    }
    real f = X[i] / d;
    if (f > (real)1) f = (real)1;
    Fsig[i] = f;
  }
  for (i64 j = 0; j < p; j++) {
    real fm = (real)1;
    int first = 1;
    for (i64 i = 0; i < s; i++) {
      real f = NV[j * s + i] < (real)0 ? Fsig[i] : (real)1;
      fm = first ? f : nanmin(fm, f);
      first = 0;
    }
    Fmin[j] = fm;
  }
}

void ms_oracle_negative_adjusted_nv(const real *NV, const real *X, i64 c, i64 p, i64 s,
                                    real *NV_adj) {
  real *Fmin = (real *)malloc(sizeof(real) * (size_t)(p + s + 1));
  real *Fsig = Fmin + p;
  for (i64 ci = 0; ci < c; ci++) {
    neg_adjust_cell(NV + ci * p * s, X + ci * s, p, s, Fmin, Fsig);
    for (i64 j = 0; j < p; j++)
      for (i64 i = 0; i < s; i++)
        NV_adj[(ci * p + j) * s + i] = NV[(ci * p + j) * s + i] * Fmin[j];
  }
  free(Fmin);
}

/* lower/upper thresholds of kinetics.py:822-823 cast to the tensor dtype */
#define UPPER ((real)1.5)
#define LOWER ((real)(1.0 / 1.5))
#define IMPACT (IS_F32 ? (real)0.1f : (real)0.1)
static const double INCS[4] = {0.5, 0.25, 0.125, 0.0625};

typedef struct {
  const i32 *N, *Nf, *Nb, *A;
  const real *Kmr, *Ke, *Kmf, *Kmb;
  i64 c, p, s;
} params_t;

/*
 * kinetics.py:808-859.  The early return at :846-847 is a GLOBAL any() over all
 * cells and proteins, so each iteration is a batch-wide pass.  The per-protein,
 * per-signal rate NV'_ji is either read from a dense (c,p,s) tensor (staged call,
 * NV_dense != NULL) or recomputed bit-identically as (float(N_ji)*V_j)*Fmin_j.
 * X1 holds the unadjusted X1 on entry and the adjusted one on return; F (c,p) and
 * dir (c,p) are scratch.  Returns the number of F updates applied (0..max_iters).
 */
static int eq_adjust(const params_t *P, const real *NV_dense, const real *V,
                     const real *Fmin, const real *X0, real *X1, real *F, int8_t *dir,
                     int max_iters, int nt, int forced, uint32_t *flag_imp, uint32_t *flag_any) {
  const i64 c = P->c, p = P->p, s = P->s;
  (void)nt;
  for (i64 r = 0; r < c * p; r++) F[r] = (real)1; /* :829 */
  int done = 0;
  for (int it = 0; it < max_iters && it < 4; it++) {
    int any = 0, anyflag = 0;
#pragma omp parallel for num_threads(nt) reduction(| : any, anyflag) schedule(static)
    for (i64 ci = 0; ci < c; ci++) {
      for (i64 j = 0; j < p; j++) {
        i64 r = ci * p + j;
        real v = V[r];
        int has_impact = (v < (real)0 ? -v : v) > IMPACT; /* :825 */
        int is_fwd = v > (real)0;                          /* :826 */
        real q = quotient(X1 + ci * s, P->Nf + r * s, P->Nb + r * s, s);
        real qk = q / P->Ke[r]; /* :833 */
        int lo = is_fwd ? (qk < LOWER) : (qk > UPPER); /* :839 */
        if (is_fwd && F[r] == (real)1) lo = 0;         /* :840 */
        int hi = is_fwd ? (qk > UPPER) : (qk < LOWER); /* :842 */
        if (!is_fwd && F[r] == (real)0) hi = 0;        /* :843 */
        dir[r] = (int8_t)((hi ? 1 : 0) | (lo ? 2 : 0));
        if ((lo || hi) && has_impact) any = 1;
        if (lo || hi) anyflag = 1;
      }
    }
    if (flag_imp && any) flag_imp[it] = 1;
    if (flag_any && anyflag) flag_any[it] = 1;
    /* :846-847; `forced` mirrors the CUDA entry points, which apply a given number of
       iterations and report the flags instead of deciding themselves */
    if (!any && !forced) break;
    done++;
    real inc = (real)INCS[it];
#pragma omp parallel for num_threads(nt) schedule(static)
    for (i64 ci = 0; ci < c; ci++) {
      for (i64 j = 0; j < p; j++) { /* :850-853 */
        i64 r = ci * p + j;
        real f = F[r];
        if (dir[r] & 1) f = f - inc;
        if (dir[r] & 2) f = f + inc;
        if (f > (real)1) f = (real)1;
        if (f < (real)0) f = (real)0;
        F[r] = f;
      }
      for (i64 i = 0; i < s; i++) { /* :856-857 */
        real acc = (real)0;
        for (i64 j = 0; j < p; j++) {
          i64 r = ci * p + j;
          real nv = NV_dense ? NV_dense[r * s + i]
                             : ((real)P->N[r * s + i] * V[r]) * Fmin[r];
          acc = acc + nv * F[r];
        }
        real x = X0[ci * s + i] + acc;
        if (x < (real)0) x = (real)0;
        X1[ci * s + i] = x;
      }
    }
  }
  return done;
}

/* _get_equilibrium_adjusted_x(X0, X1, NV, V) on dense NV (c,p,s) */
int ms_oracle_equilibrium_adjusted_x(const real *X0, const real *X1_in, const real *NV,
                                     const real *V, const i32 *Nf, const i32 *Nb,
                                     const real *Ke, i64 c, i64 p, i64 s, real *X1,
                                     int *n_done, int n_threads) {
  params_t P = {0, Nf, Nb, 0, 0, Ke, 0, 0, c, p, s};
  real *F = (real *)malloc(sizeof(real) * (size_t)(c * p + 1));
  int8_t *dir = (int8_t *)malloc((size_t)(c * p + 1));
  if (!F || !dir) return -1;
  memcpy(X1, X1_in, sizeof(real) * (size_t)(c * s));
  int done = eq_adjust(&P, NV, V, 0, X0, X1, F, dir, 4, n_threads_or_default(n_threads), 0, 0, 0);
  if (n_done) *n_done = done;
  free(F);
  free(dir);
  return 0;
}

/* ------------------------------------------------------------------ */
/* integrate_signals: memory-light batch version (no (c,p,s) floats)    */
/* ------------------------------------------------------------------ */

/*
 * kinetics.py:753-769 _integrate_signals_part for the whole batch.
 * max_eq_iters: 4 = reference behaviour, 0 = the reference tests' _MockedKinetics
 * (tests/fast/test_kinetics.py:87-96) which disables the equilibrium adjustment.
 * eq_done receives the number of F updates applied (global early exit).
 */
static void integrate_part(const params_t *P, const real *adj_vmax, const real *X0,
                           real *X1, int max_eq_iters, int *eq_done, real *V, real *Fmin,
                           real *F, int8_t *dir, int nt, int forced, uint32_t *flag_imp,
                           uint32_t *flag_any) {
  const i64 c = P->c, p = P->p, s = P->s;
#pragma omp parallel num_threads(nt)
  {
    real *NV = (real *)malloc(sizeof(real) * (size_t)(p * s + s + 1));
    real *Fsig = NV + p * s;
#pragma omp for schedule(static)
    for (i64 ci = 0; ci < c; ci++) {
      const real *x0 = X0 + ci * s;
      for (i64 j = 0; j < p; j++) {
        i64 r = ci * p + j;
        V[r] = velocity(x0, P->Nf + r * s, P->Nb + r * s, P->A + r * s, P->Kmr + r * s,
                        P->Kmf[r], P->Kmb[r], adj_vmax[r], s);
        for (i64 i = 0; i < s; i++) NV[j * s + i] = (real)P->N[r * s + i] * V[r]; /* :759 */
      }
      neg_adjust_cell(NV, x0, p, s, Fmin + ci * p, Fsig);
      for (i64 i = 0; i < s; i++) { /* :763-764 */
        real acc = (real)0;
        for (i64 j = 0; j < p; j++) acc = acc + NV[j * s + i] * Fmin[ci * p + j];
        real x = x0[i] + acc;
        if (x < (real)0) x = (real)0;
        X1[ci * s + i] = x;
      }
    }
    free(NV);
  }
  int done = eq_adjust(P, 0, V, Fmin, X0, X1, F, dir, max_eq_iters, nt, forced, flag_imp, flag_any);
  if (eq_done) *eq_done = done;
}

/*
 * One _integrate_signals_part call (adj_vmax given by the caller).
 * V_out (c,p) optional.
 */
int ms_oracle_integrate_part(const i32 *N, const i32 *Nf, const i32 *Nb, const i32 *A,
                             const real *Kmr, const real *Ke, const real *Kmf,
                             const real *Kmb, const real *adj_vmax, i64 c, i64 p, i64 s,
                             const real *X0, real *X1, int max_eq_iters, int *eq_done,
                             real *V_out, int n_threads) {
  params_t P = {N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, c, p, s};
  size_t cp = (size_t)(c * p + 1);
  real *V = (real *)malloc(sizeof(real) * cp * 3);
  int8_t *dir = (int8_t *)malloc(cp);
  if (!V || !dir) return -1;
  integrate_part(&P, adj_vmax, X0, X1, max_eq_iters, eq_done, V, V + cp, V + 2 * cp, dir,
                 n_threads_or_default(n_threads), 0, 0, 0);
  if (V_out) memcpy(V_out, V, sizeof(real) * (size_t)(c * p));
  free(V);
  free(dir);
  return 0;
}

/*
 * kinetics.py:725-751 integrate_signals: three parts with trim 0.7, 0.2, 0.1.
 * eq_done (optional, int[3]) receives the number of F updates per part.
 */
static int integrate_signals_impl(const i32 *N, const i32 *Nf, const i32 *Nb, const i32 *A,
                                  const real *Kmr, const real *Ke, const real *Kmf,
                                  const real *Kmb, const real *Vmax, i64 c, i64 p, i64 s,
                                  const real *X_in, real *X_out, int *eq_done, int n_threads,
                                  const int *eq_iters, uint32_t *flags) {
  params_t P = {N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, c, p, s};
  static const double TRIMS[3] = {0.7, 0.2, 0.1};
  size_t cp = (size_t)(c * p + 1), cs = (size_t)(c * s + 1);
  real *V = (real *)malloc(sizeof(real) * (cp * 4 + cs * 2));
  int8_t *dir = (int8_t *)malloc(cp);
  if (!V || !dir) return -1;
  real *Fmin = V + cp, *F = V + 2 * cp, *adj = V + 3 * cp;
  real *Xa = V + 4 * cp, *Xb = Xa + cs;
  memcpy(Xa, X_in, sizeof(real) * (size_t)(c * s));
  int nt = n_threads_or_default(n_threads);
  for (int q = 0; q < 3; q++) {
    real trim = (real)TRIMS[q];
    for (i64 r = 0; r < c * p; r++) { /* :749 (Vmax * trim).clamp(0.0) */
      real v = Vmax[r] * trim;
      if (v < (real)0) v = (real)0;
      adj[r] = v;
    }
    integrate_part(&P, adj, Xa, Xb, eq_iters ? eq_iters[q] : 4, eq_done ? eq_done + q : 0, V,
                   Fmin, F, dir, nt, eq_iters != 0, flags ? flags + 4 * q : 0,
                   flags ? flags + 12 + 4 * q : 0);
    real *t = Xa;
    Xa = Xb;
    Xb = t;
  }
  memcpy(X_out, Xa, sizeof(real) * (size_t)(c * s));
  free(V);
  free(dir);
  return 0;
}

int ms_oracle_integrate_signals(const i32 *N, const i32 *Nf, const i32 *Nb, const i32 *A,
                                const real *Kmr, const real *Ke, const real *Kmf,
                                const real *Kmb, const real *Vmax, i64 c, i64 p, i64 s,
                                const real *X_in, real *X_out, int *eq_done,
                                int n_threads) {
  return integrate_signals_impl(N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, Vmax, c, p, s, X_in, X_out,
                                eq_done, n_threads, 0, 0);
}

/*
 * CPU mirror of the CONTRACT of the CUDA entry point ms_integrate_signals
 * (include/magicsoup_b200.h): apply eq_iters[q] adjustment iterations in part q without
 * deciding the global early exit, and report flags[q*4+k] (some cell had an adjustable
 * protein with |V| > 0.1) and flags[12+q*4+k] (some cell had any adjustable protein).
 * Lets the multi-rank host protocol (flags all-reduced over ranks) be tested on CPU.
 */
int ms_oracle_integrate_signals_iters(const i32 *N, const i32 *Nf, const i32 *Nb,
                                      const i32 *A, const real *Kmr, const real *Ke,
                                      const real *Kmf, const real *Kmb, const real *Vmax,
                                      i64 c, i64 p, i64 s, const real *X_in, real *X_out,
                                      const int *eq_iters, uint32_t *flags, int n_threads) {
  memset(flags, 0, 24 * sizeof(uint32_t));
  if (c == 0) return 0;
  return integrate_signals_impl(N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, Vmax, c, p, s, X_in, X_out, 0,
                                n_threads, eq_iters, flags);
}

/* ------------------------------------------------------------------ */
/* world.py:610-625 enzymatic_activity: gather, integrate, scatter      */
/* ------------------------------------------------------------------ */
int ms_oracle_enzymatic_activity(const i32 *N, const i32 *Nf, const i32 *Nb, const i32 *A,
                                 const real *Kmr, const real *Ke, const real *Kmf,
                                 const real *Kmb, const real *Vmax, i64 c, i64 p, i64 m,
                                 const i32 *pos /* (c,2) */, real *cell_molecules /* (c,m) */,
                                 real *molecule_map /* (m,H,W) */, i64 H, i64 W,
                                 int *eq_done, int n_threads) {
  if (c == 0) return 0;
  i64 s = 2 * m;
  real *X0 = (real *)malloc(sizeof(real) * (size_t)(c * s) * 2);
  if (!X0) return -1;
  real *X1 = X0 + c * s;
  for (i64 ci = 0; ci < c; ci++) {
    i64 x = pos[2 * ci], y = pos[2 * ci + 1];
    for (i64 k = 0; k < m; k++) {
      X0[ci * s + k] = cell_molecules[ci * m + k];
      X0[ci * s + m + k] = molecule_map[(k * H + x) * W + y];
    }
  }
  int rc = ms_oracle_integrate_signals(N, Nf, Nb, A, Kmr, Ke, Kmf, Kmb, Vmax, c, p, s, X0,
                                       X1, eq_done, n_threads);
  for (i64 ci = 0; ci < c; ci++) {
    i64 x = pos[2 * ci], y = pos[2 * ci + 1];
    for (i64 k = 0; k < m; k++) {
      cell_molecules[ci * m + k] = X1[ci * s + k];
      molecule_map[(k * H + x) * W + y] = X1[ci * s + m + k];
    }
  }
  free(X0);
  return rc;
}

/* ------------------------------------------------------------------ */
/* world.py:638-649 diffusion; :651-665 permeation                      */
/* ------------------------------------------------------------------ */

/* exact, order independent plane sum: integers of 2^-32 (see file header) */
static void fixsum_add(real x, i64 *hi, i64 *lo) {
  long long q = llrint((double)x * 4294967296.0);
  *hi += q >> 32;
  *lo += q & 0xffffffffLL;
}

/*
 * 3x3 circular stencil of a ROW SLAB (m,H,W) with weights [[a,a,a],[a,b,a],[a,a,a]]
 * (world.py:966-972) into out (m,H,W).  The rows above row 0 / below row H-1 come from
 * halo_top / halo_bot (m,W) (NULL: wrap around inside the slab = the whole map).
 * sums (4 per plane) receive the exact fixed-point plane sums {hi,lo} before and after.
 * Spec order: top=(nw+n)+ne ; mid=w+e ; bot=(sw+s)+se ; nb=(top+mid)+bot ;
 * out = b*c + a*nb (two roundings for the products, one for the sum).
 */
int ms_oracle_diffuse_stencil(const real *map, const real *halo_top, const real *halo_bot,
                              const real *a, const real *b, i64 m, i64 H, i64 W, real *out,
                              i64 *sums) {
  for (i64 k = 0; k < m; k++) {
    const real *in = map + k * H * W;
    real *o_ = out + k * H * W;
    i64 hi0 = 0, lo0 = 0, hi1 = 0, lo1 = 0;
    for (i64 x = 0; x < H; x++) {
      const real *rm = x > 0 ? in + (x - 1) * W : (halo_top ? halo_top + k * W : in + (H - 1) * W);
      const real *r0 = in + x * W;
      const real *rp = x < H - 1 ? in + (x + 1) * W : (halo_bot ? halo_bot + k * W : in);
      for (i64 y = 0; y < W; y++) {
        i64 ym = (y + W - 1) % W, yp = (y + 1) % W;
        real top = (rm[ym] + rm[y]) + rm[yp];
        real mid = r0[ym] + r0[yp];
        real bot = (rp[ym] + rp[y]) + rp[yp];
        real nb = (top + mid) + bot;
        real c0 = r0[y];
        real o = b[k] * c0 + a[k] * nb;
        o_[x * W + y] = o;
        fixsum_add(c0, &hi0, &lo0);
        fixsum_add(o, &hi1, &lo1);
      }
    }
    sums[4 * k] = hi0;
    sums[4 * k + 1] = lo0;
    sums[4 * k + 2] = hi1;
    sums[4 * k + 3] = lo1;
  }
  return 0;
}

/*
 * world.py:640-649: total_before / total_after are float32 scalars (torch.sum), their
 * difference and the division by n_pxls are float32 too.  The exact sum rounded once to
 * `real` stands in for torch.sum's float32 tree.  map = clamp(out + corr, 0).
 */
int ms_oracle_diffuse_correct(const real *out, real *map, const i64 *sums, i64 m, i64 H, i64 W,
                              i64 n_pxls_total) {
  for (i64 k = 0; k < m; k++) {
    real s0 = (real)((double)sums[4 * k] + (double)sums[4 * k + 1] * 2.3283064365386963e-10);
    real s1 = (real)((double)sums[4 * k + 2] + (double)sums[4 * k + 3] * 2.3283064365386963e-10);
    real corr = (s0 - s1) / (real)n_pxls_total;
    for (i64 t = 0; t < H * W; t++) {
      real v = out[k * H * W + t] + corr;
      if (v < (real)0) v = (real)0;
      map[k * H * W + t] = v;
    }
  }
  return 0;
}

/* world.py:638-649 on a whole map, in place */
int ms_oracle_diffuse(real *molecule_map, const real *a, const real *b, i64 m, i64 H,
                      i64 W, int n_threads) {
  (void)n_threads;
  real *out = (real *)malloc(sizeof(real) * (size_t)(m * H * W + 1));
  i64 *sums = (i64 *)malloc(sizeof(i64) * (size_t)(4 * m + 1));
  if (!out || !sums) return -1;
  ms_oracle_diffuse_stencil(molecule_map, 0, 0, a, b, m, H, W, out, sums);
  ms_oracle_diffuse_correct(out, molecule_map, sums, m, H, W, H * W);
  free(out);
  free(sums);
  return 0;
}

/* world.py:654-665, perm = float32 factors from _get_permeate */
int ms_oracle_permeate(real *cell_molecules, real *molecule_map, const i32 *pos,
                       const real *perm, i64 c, i64 m, i64 H, i64 W) {
  for (i64 ci = 0; ci < c; ci++) {
    i64 x = pos[2 * ci], y = pos[2 * ci + 1];
    for (i64 k = 0; k < m; k++) {
      real xi = cell_molecules[ci * m + k];
      real xe = molecule_map[(k * H + x) * W + y];
      real di = xi * perm[k], de = xe * perm[k];
      cell_molecules[ci * m + k] = xi + (de - di);
      molecule_map[(k * H + x) * W + y] = xe + (di - de);
    }
  }
  return 0;
}
