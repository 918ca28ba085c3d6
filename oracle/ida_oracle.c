/*
 * ida_oracle.c -- CPU ORACLE (test infrastructure, NOT the product path).
 *
 * A plain-C, single-system-at-a-time restatement of the algorithm behind
 * pybamm.IDAKLUSolver for the batched-solve hot path:
 *
 *   driver   : IDAKLUSolverOpenMP::solve and helpers
 *              (packages/pybammsolvers/src/pybammsolvers/idaklu_source/IDAKLUSolverOpenMP.inl:327-444
 *               setup :514-563, breakpoints :618-645, consistent init :787-873, t_interp :900-928)
 *   IC Newton: NonlinearSolver::RunNewtonLoop (idaklu_source/NonlinearSolver.inl:100-177),
 *              sub-block system (idaklu_source/AlgebraicICBuilder.inl:52-119)
 *   residual : F = f(t,y,p) - M y'   (idaklu_source/sundials_functions.inl:9-39)
 *   Jacobian : J - cj M              (idaklu_source/sundials_functions.inl:150-259)
 *   time grid: IDAKLUSolverGroup::solve (idaklu_source/IDAKLUSolverGroup.cpp:25-43), perturb_time (common.hpp:112-118)
 *   options  : idaklu_solver.py:286-325
 *
 * The integrator proper is SUNDIALS IDAS (pinned submodule ddf5daba..., v7 API), whose
 * source is NOT in the reference tree.  It is restated here from its published
 * algorithm (variable-order variable-coefficient BDF in fixed-leading-coefficient
 * form; Brenan, Campbell & Petzold 1996; Hindmarsh et al. 2005, "SUNDIALS"):
 * IDASetCoeffs / IDANls + Newton / IDATestError / IDAHandleNFlag / IDACompleteStep /
 * IDAGetSolution / IDARcheck3 + IDARootfind, with PyBaMM's option values.  The linear
 * solver is a banded LU with partial pivoting on an RCM ordering (KLU is not in the
 * tree either); any accurate direct solve gives the same Newton iterates to round-off.
 *
 * PARITY STATUS: pinned against the reference's analytic known-answer tests and
 * against pybamm's own expression evaluation (tests/golden); NOT pinned against a
 * running IDAKLU (it cannot be built or installed in this environment).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <float.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "pbo_model.h"

#define MXORDP1 6
#define UROUND DBL_EPSILON

/* IDA return codes used by the python side (sundials_error_handler.hpp:11-61) */
#define IDA_SUCCESS 0
#define IDA_TSTOP_RETURN 1
#define IDA_ROOT_RETURN 2
#define IDA_TOO_MUCH_WORK (-1)
#define IDA_TOO_MUCH_ACC (-2)
#define IDA_ERR_FAIL (-3)
#define IDA_CONV_FAIL (-4)
#define IDA_LSETUP_FAIL (-6)

typedef struct {
  double rtol;
  int maxord;            /* max_order_bdf 5 */
  long max_num_steps;    /* 100000 */
  int maxnef;            /* max_error_test_failures 10 */
  int maxcor;            /* max_nonlinear_iterations 40 */
  int maxncf;            /* max_convergence_failures 100 */
  double epcon;          /* nonlinear_convergence_coefficient 0.33 */
  double epsNewtIC;      /* nonlinear_convergence_coefficient_ic 0.0033 */
  double step_tol_ic;    /* newton_step_tol 1e-4 */
  int max_iter_ic;       /* max_num_iterations_ic 100 */
  int max_backtracks_ic; /* max_linesearch_backtracks_ic 5 */
  int scalesol;          /* linear_solution_scaling 1 */
  double dt_init;        /* 0 -> IDA heuristic */
  double dt_max;         /* 0 -> unbounded */
} pbo_options;

void pbo_default_options(pbo_options* o) {
  o->rtol = 1e-4; o->maxord = 5; o->max_num_steps = 100000; o->maxnef = 10; o->maxcor = 40;
  o->maxncf = 100; o->epcon = 0.33; o->epsNewtIC = 0.0033; o->step_tol_ic = 1e-4;
  o->max_iter_ic = 100; o->max_backtracks_ic = 5; o->scalesol = 1; o->dt_init = 0.0; o->dt_max = 0.0;
}

/* ------------------------------------------------------------------ banded LU */
typedef struct {
  int n, kl, ku, W;
  double* a;   /* n * W, row i holds columns i-kl .. i+kl+ku */
  int* ipiv;
} band_t;

#define BA(B, i, j) ((B)->a[(size_t)(i) * (B)->W + ((j) - (i) + (B)->kl)])

static int band_factor(band_t* B) {
  const int n = B->n, kl = B->kl, kw = B->kl + B->ku;
  for (int k = 0; k < n; k++) {
    int rmax = k + kl < n - 1 ? k + kl : n - 1;
    int jmax = k + kw < n - 1 ? k + kw : n - 1;
    int piv = k;
    double best = fabs(BA(B, k, k));
    for (int r = k + 1; r <= rmax; r++) {
      double v = fabs(BA(B, r, k));
      if (v > best) { best = v; piv = r; }
    }
    B->ipiv[k] = piv;
    if (best == 0.0 || !(best == best)) return k + 1;
    if (piv != k)
      for (int j = k; j <= jmax; j++) { double tmp = BA(B, k, j); BA(B, k, j) = BA(B, piv, j); BA(B, piv, j) = tmp; }
    const double inv = 1.0 / BA(B, k, k);
    for (int r = k + 1; r <= rmax; r++) {
      double l = BA(B, r, k) * inv;
      BA(B, r, k) = l;
      if (l != 0.0)
        for (int j = k + 1; j <= jmax; j++) BA(B, r, j) -= l * BA(B, k, j);
    }
  }
  return 0;
}

static void band_solve(const band_t* B, double* b) {
  const int n = B->n, kl = B->kl, kw = B->kl + B->ku;
  for (int k = 0; k < n; k++) {
    int piv = B->ipiv[k];
    if (piv != k) { double tmp = b[k]; b[k] = b[piv]; b[piv] = tmp; }
    int rmax = k + kl < n - 1 ? k + kl : n - 1;
    const double bk = b[k];
    if (bk != 0.0)
      for (int r = k + 1; r <= rmax; r++) b[r] -= BA(B, r, k) * bk;
  }
  for (int i = n - 1; i >= 0; i--) {
    int jmax = i + kw < n - 1 ? i + kw : n - 1;
    double s = b[i];
    for (int j = i + 1; j <= jmax; j++) s -= BA(B, i, j) * b[j];
    b[i] = s / BA(B, i, i);
  }
}

/* ------------------------------------------------------------------ dense LU (IC sub-block) */
static int dense_factor(double* a, int n, int* ipiv) {
  for (int k = 0; k < n; k++) {
    int piv = k; double best = fabs(a[(size_t)k * n + k]);
    for (int r = k + 1; r < n; r++) { double v = fabs(a[(size_t)r * n + k]); if (v > best) { best = v; piv = r; } }
    ipiv[k] = piv;
    if (best == 0.0 || !(best == best)) return k + 1;
    if (piv != k) for (int j = 0; j < n; j++) { double t = a[(size_t)k * n + j]; a[(size_t)k * n + j] = a[(size_t)piv * n + j]; a[(size_t)piv * n + j] = t; }
    double inv = 1.0 / a[(size_t)k * n + k];
    for (int r = k + 1; r < n; r++) {
      double l = a[(size_t)r * n + k] * inv; a[(size_t)r * n + k] = l;
      if (l != 0.0) for (int j = k + 1; j < n; j++) a[(size_t)r * n + j] -= l * a[(size_t)k * n + j];
    }
  }
  return 0;
}
static void dense_solve(const double* a, int n, const int* ipiv, double* b) {
  for (int k = 0; k < n; k++) {
    int piv = ipiv[k]; if (piv != k) { double t = b[k]; b[k] = b[piv]; b[piv] = t; }
    for (int r = k + 1; r < n; r++) b[r] -= a[(size_t)r * n + k] * b[k];
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = b[i];
    for (int j = i + 1; j < n; j++) s -= a[(size_t)i * n + j] * b[j];
    b[i] = s / a[(size_t)i * n + i];
  }
}

/* ------------------------------------------------------------------ static-pattern sparse LU (CPU baseline leg)
 * What SUNLinSol_KLU does in steady state (IDAKLUSolverOpenMP_solvers.hpp:29-33): the fill-reducing ordering and the
 * pivot sequence are fixed once (klu_analyze + the first klu_factor), every later set-up is a klu_refactor -- a numeric
 * factorisation on the frozen pattern with no pivot search.  Here the ordering (row / column permutations) comes from
 * pybamm_b200.symbolic_lu, the same one the device uses; the symbolic fill is recomputed below and the numeric phase
 * is the row-wise (IKJ) elimination on CSR.  lu_mode 1 of pbo_solve; the banded LU with partial pivoting above stays
 * the checker's linear solver (lu_mode 0). */
typedef struct {
  int n, nnz;
  int *rp, *ci, *dg;     /* CSR of L+U in permuted numbering, columns ascending; dg[i] = position of the diagonal */
  double *v, *w;         /* values; dense work row */
  int* a_pos;            /* position in v of every structural entry of J - cj M */
  int *ipr, *ipc;        /* old row -> new row, old column -> new column */
  const int *pr, *pc;    /* new -> old */
} splu_t;

static void splu_init(splu_t* L, int n, int nnz_a, const int* rows, const int* cols, const int* perm_r, const int* perm_c) {
  L->n = n; L->pr = perm_r; L->pc = perm_c;
  L->ipr = (int*)malloc(sizeof(int) * n); L->ipc = (int*)malloc(sizeof(int) * n);
  for (int i = 0; i < n; i++) { L->ipr[perm_r[i]] = i; L->ipc[perm_c[i]] = i; }
  /* pattern of the permuted matrix, row by row, then symbolic fill: row i gains the U pattern of every row k < i it holds */
  char* in = (char*)calloc((size_t)n * n, 1);     /* n <= a few thousand: a dense bitmap is the simplest exact symbolic phase */
  for (int k = 0; k < nnz_a; k++) in[(size_t)L->ipr[rows[k]] * n + L->ipc[cols[k]]] = 1;
  for (int i = 0; i < n; i++) in[(size_t)i * n + i] = 1;
  for (int i = 0; i < n; i++) {
    char* ri = in + (size_t)i * n;
    for (int k = 0; k < i; k++)
      if (ri[k]) {
        const char* rk = in + (size_t)k * n;
        for (int j = k + 1; j < n; j++) ri[j] |= rk[j];
      }
  }
  L->rp = (int*)malloc(sizeof(int) * (n + 1)); L->dg = (int*)malloc(sizeof(int) * n);
  int nnz = 0;
  for (size_t q = 0; q < (size_t)n * n; q++) nnz += in[q];
  L->nnz = nnz;
  L->ci = (int*)malloc(sizeof(int) * nnz); L->v = (double*)calloc(nnz, sizeof(double)); L->w = (double*)calloc(n, sizeof(double));
  int p = 0;
  for (int i = 0; i < n; i++) {
    L->rp[i] = p;
    for (int j = 0; j < n; j++)
      if (in[(size_t)i * n + j]) { if (j == i) L->dg[i] = p; L->ci[p++] = j; }
  }
  L->rp[n] = p;
  L->a_pos = (int*)malloc(sizeof(int) * (nnz_a > 0 ? nnz_a : 1));
  for (int k = 0; k < nnz_a; k++) {
    const int i = L->ipr[rows[k]], j = L->ipc[cols[k]];
    int lo = L->rp[i], hi = L->rp[i + 1] - 1;
    while (lo < hi) { int mid = (lo + hi) >> 1; if (L->ci[mid] < j) lo = mid + 1; else hi = mid; }
    L->a_pos[k] = lo;
  }
  free(in);
}
static void splu_free(splu_t* L) {
  free(L->rp); free(L->ci); free(L->dg); free(L->v); free(L->w); free(L->a_pos); free(L->ipr); free(L->ipc);
}
/* numeric factorisation of the values scattered into L->v; the diagonal of U is stored inverted */
static int splu_factor(splu_t* L) {
  const int n = L->n;
  double* w = L->w;
  for (int i = 0; i < n; i++) {
    const int p0 = L->rp[i], p1 = L->rp[i + 1], pd = L->dg[i];
    for (int p = p0; p < p1; p++) w[L->ci[p]] = L->v[p];
    for (int p = p0; p < pd; p++) {
      const int k = L->ci[p];
      const double l = w[k] * L->v[L->dg[k]];
      w[k] = l;
      if (l != 0.0)
        for (int q = L->dg[k] + 1; q < L->rp[k + 1]; q++) w[L->ci[q]] -= l * L->v[q];
    }
    const double d = w[i];
    if (d == 0.0 || !(d == d)) return i + 1;
    for (int p = p0; p < p1; p++) L->v[p] = w[L->ci[p]];
    L->v[pd] = 1.0 / d;
  }
  return 0;
}
static void splu_solve(const splu_t* L, const double* b, double* x, double* work) {
  const int n = L->n;
  for (int i = 0; i < n; i++) {
    double s = b[L->pr[i]];
    for (int p = L->rp[i]; p < L->dg[i]; p++) s -= L->v[p] * work[L->ci[p]];
    work[i] = s;
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = work[i];
    for (int p = L->dg[i] + 1; p < L->rp[i + 1]; p++) s -= L->v[p] * work[L->ci[p]];
    work[i] = s * L->v[L->dg[i]];
  }
  for (int j = 0; j < n; j++) x[L->pc[j]] = work[j];
}

/* ------------------------------------------------------------------ solver state */
typedef struct {
  const pbo_model* m;
  pbo_options opt;
  const double* atol;
  const int* perm;   /* new -> old (RCM), NULL = identity */
  int* iperm;        /* old -> new */
  int n;
  /* IDA memory */
  double* phi[MXORDP1];
  double psi[MXORDP1], alpha[MXORDP1], beta[MXORDP1], sigma[MXORDP1], gamma[MXORDP1];
  double *ewt, *ee, *delta, *yypredict, *yppredict, *yy, *yp, *tempv, *tempv2, *savres, *jacv, *tmp, *bw;
  double *p;
  double tn, hh, hused, cj, cjlast, cjold, cjratio, ss, rr, epsNewt, toldel, hmax_inv;
  int kk, kused, knew, phase, ns;
  long nst;
  double tstop; int tstopset;
  double tretlast;
  /* roots */
  double *glo, *ghi, *grout; int* gactive; int* iroots; double tlo, thi, trout, ttol;
  int irfnd;
  /* linear solver: banded + partial pivoting (checker) or static-pattern sparse (baseline leg) */
  band_t B;
  splu_t SP; int lu_mode;
  /* IC */
  double *Jaa; int* ipiv_a; double *xa, *xsave, *resa, *dela, *ewta;
  /* stats: nsteps nrevals nlinsetups netfails nniters nncfails njevals nge */
  long st[8];
} ida_t;

enum { ST_NST, ST_NRE, ST_NSETUPS, ST_NETF, ST_NNI, ST_NCFN, ST_NJE, ST_NGE };

static double wrms(const ida_t* S, const double* x) {
  double s = 0.0;
  for (int i = 0; i < S->n; i++) { double w = x[i] * S->ewt[i]; s += w * w; }
  return sqrt(s / S->n);
}
static int ewt_set(ida_t* S, const double* y) {
  for (int i = 0; i < S->n; i++) {
    double w = S->opt.rtol * fabs(y[i]) + S->atol[i];
    if (!(w > 0.0)) return -1;
    S->ewt[i] = 1.0 / w;
  }
  return 0;
}

/* F(t, y, y') = f(t,y,p) - M y'   (sundials_functions.inl:9-39; M = I on the first n_diff rows) */
static void residual(ida_t* S, double t, const double* y, const double* yp, double* r) {
  S->m->res(y, S->p, t, 0.0, r, S->tmp);
  for (int i = 0; i < S->m->n_diff; i++) r[i] -= yp[i];
  S->st[ST_NRE]++;
}

static int lsetup(ida_t* S, double t, const double* y) {
  const pbo_model* m = S->m;
  m->jac(y, S->p, t, S->cj, S->jacv, S->tmp);
  S->st[ST_NJE]++;
  S->st[ST_NSETUPS]++;
  if (S->lu_mode) {
    memset(S->SP.v, 0, sizeof(double) * (size_t)S->SP.nnz);
    for (int k = 0; k < m->nnz; k++) S->SP.v[S->SP.a_pos[k]] += S->jacv[k];
    return splu_factor(&S->SP);
  }
  memset(S->B.a, 0, sizeof(double) * (size_t)S->n * S->B.W);
  for (int k = 0; k < m->nnz; k++) {
    int i = S->iperm[m->jac_rows[k]], j = S->iperm[m->jac_cols[k]];
    BA(&S->B, i, j) += S->jacv[k];
  }
  return band_factor(&S->B);
}
static void lsolve(ida_t* S, double* b) {
  if (S->lu_mode) {
    splu_solve(&S->SP, b, S->tempv2, S->bw);
    memcpy(b, S->tempv2, sizeof(double) * S->n);
  } else {
    /* permute, solve, un-permute */
    for (int i = 0; i < S->n; i++) S->bw[S->iperm[i]] = b[i];
    band_solve(&S->B, S->bw);
    for (int i = 0; i < S->n; i++) b[i] = S->bw[S->iperm[i]];
  }
  if (S->opt.scalesol && S->cjratio != 1.0) {
    double f = 2.0 / (1.0 + S->cjratio);
    for (int i = 0; i < S->n; i++) b[i] *= f;
  }
}

/* IDAGetSolution: interpolating polynomial of order kused at t (and derivative) */
static void get_solution(const ida_t* S, double t, double* yret, double* ypret) {
  double cvals[MXORDP1], dvals[MXORDP1];
  int kord = S->kused; if (kord == 0) kord = 1;
  double delt = t - S->tn, c = 1.0, d = 0.0, gam = delt / S->psi[0];
  cvals[0] = c;
  for (int j = 1; j <= kord; j++) {
    d = d * gam + c / S->psi[j - 1];
    c = c * gam;
    gam = (delt + S->psi[j - 1]) / S->psi[j];
    cvals[j] = c; dvals[j - 1] = d;
  }
  for (int i = 0; i < S->n; i++) {
    double s = 0.0;
    for (int j = kord; j >= 0; j--) s += cvals[j] * S->phi[j][i];
    yret[i] = s;
  }
  if (ypret)
    for (int i = 0; i < S->n; i++) {
      double s = 0.0;
      for (int j = kord; j >= 1; j--) s += dvals[j - 1] * S->phi[j][i];
      ypret[i] = s;
    }
}

static void reinit(ida_t* S, double t0, const double* y0, const double* yp0) {
  S->tn = t0;
  memcpy(S->phi[0], y0, sizeof(double) * S->n);
  memcpy(S->phi[1], yp0, sizeof(double) * S->n);
  S->nst = 0; S->kk = 0; S->kused = 0; S->hused = 0.0; S->hh = 0.0;
  S->irfnd = 0; S->tstopset = 0; S->tretlast = t0;
  S->epsNewt = S->opt.epcon; S->toldel = 1e-4 * S->epsNewt;
}

/* ------------------------------------------------------------------ one BDF step (IDAStep) */
static void set_coeffs(ida_t* S, double* ck) {
  int kk = S->kk;
  if (S->hh != S->hused || kk != S->kused) S->ns = 0;
  S->ns = (S->ns + 1 < S->kused + 2) ? S->ns + 1 : S->kused + 2;
  if (kk + 1 >= S->ns) {
    S->beta[0] = 1.0; S->alpha[0] = 1.0; S->gamma[0] = 0.0; S->sigma[0] = 1.0;
    double temp1 = S->hh;
    for (int i = 1; i <= kk; i++) {
      double temp2 = S->psi[i - 1];
      S->psi[i - 1] = temp1;
      S->beta[i] = S->beta[i - 1] * S->psi[i - 1] / temp2;
      temp1 = temp2 + S->hh;
      S->alpha[i] = S->hh / temp1;
      S->sigma[i] = i * S->sigma[i - 1] * S->alpha[i];
      S->gamma[i] = S->gamma[i - 1] + S->alpha[i - 1] / S->hh;
    }
    S->psi[kk] = temp1;
  }
  double alphas = 0.0, alpha0 = 0.0;
  for (int i = 0; i < kk; i++) { alphas -= 1.0 / (i + 1); alpha0 -= S->alpha[i]; }
  S->cjlast = S->cj;
  S->cj = -alphas / S->hh;
  *ck = fabs(S->alpha[kk] + alphas - alpha0);
  if (S->alpha[kk] > *ck) *ck = S->alpha[kk];
  for (int i = S->ns; i <= kk; i++)
    for (int q = 0; q < S->n; q++) S->phi[i][q] *= S->beta[i];
  S->tn += S->hh;
}

static void restore(ida_t* S, double saved_t) {
  S->tn = saved_t;
  for (int j = 1; j <= S->kk; j++) S->psi[j - 1] = S->psi[j] - S->hh;
  for (int j = S->ns; j <= S->kk; j++) {
    double r = 1.0 / S->beta[j];
    for (int q = 0; q < S->n; q++) S->phi[j][q] *= r;
  }
}

#define NLS_SUCCESS 0
#define NLS_CONV_RECVR 1
#define NLS_LSETUP_RECVR 2

/* IDANls + SUNNonlinSol_Newton + idaNlsConvTest */
static int nls(ida_t* S) {
  const int n = S->n;
  int callLSetup = 0;
  if (S->nst == 0) { S->cjold = S->cj; S->ss = 20.0; callLSetup = 1; }
  S->cjratio = S->cj / S->cjold;
  {
    const double temp1 = (1.0 - 0.25) / (1.0 + 0.25), temp2 = 1.0 / temp1;
    if (S->cjratio < temp1 || S->cjratio > temp2) callLSetup = 1;
    if (S->cj != S->cjlast) S->ss = 100.0;
  }
  /* predictor */
  for (int i = 0; i < n; i++) {
    double a = 0.0, b = 0.0;
    for (int j = S->kk; j >= 0; j--) a += S->phi[j][i];
    for (int j = S->kk; j >= 1; j--) b += S->gamma[j] * S->phi[j][i];
    S->yypredict[i] = a; S->yppredict[i] = b;
    S->ee[i] = 0.0;
  }
  int jcur = 0, retval;
  double oldnrm = 0.0;
  for (;;) {
    for (int i = 0; i < n; i++) { S->yy[i] = S->yypredict[i] + S->ee[i]; S->yp[i] = S->yppredict[i] + S->cj * S->ee[i]; }
    residual(S, S->tn, S->yy, S->yp, S->delta);
    if (callLSetup) {
      int ier = lsetup(S, S->tn, S->yy);
      S->cjold = S->cj; S->cjratio = 1.0; S->ss = 20.0; jcur = 1;
      if (ier != 0) { retval = NLS_LSETUP_RECVR; goto failed; }
    }
    int mnewt = 0;
    for (;;) {
      S->st[ST_NNI]++;
      for (int i = 0; i < n; i++) S->delta[i] = -S->delta[i];
      lsolve(S, S->delta);
      for (int i = 0; i < n; i++) S->ee[i] += S->delta[i];
      /* convergence test */
      double delnrm = wrms(S, S->delta);
      retval = -99;
      if (!(delnrm == delnrm)) { retval = NLS_CONV_RECVR; break; }
      if (mnewt == 0) {
        oldnrm = delnrm;
        if (delnrm <= 1e-4 * S->toldel) retval = NLS_SUCCESS;
      } else {
        double rate = pow(delnrm / oldnrm, 1.0 / mnewt);
        if (rate > 0.9) { retval = NLS_CONV_RECVR; break; }
        S->ss = rate / (1.0 - rate);
      }
      if (retval != NLS_SUCCESS && S->ss * delnrm <= S->epsNewt) retval = NLS_SUCCESS;
      if (retval == NLS_SUCCESS) {
        for (int i = 0; i < n; i++) { S->yy[i] = S->yypredict[i] + S->ee[i]; S->yp[i] = S->yppredict[i] + S->cj * S->ee[i]; }
        return NLS_SUCCESS;
      }
      mnewt++;
      if (mnewt >= S->opt.maxcor) { retval = NLS_CONV_RECVR; break; }
      for (int i = 0; i < n; i++) { S->yy[i] = S->yypredict[i] + S->ee[i]; S->yp[i] = S->yppredict[i] + S->cj * S->ee[i]; }
      residual(S, S->tn, S->yy, S->yp, S->delta);
    }
  failed:
    if (retval > 0 && !jcur) {
      callLSetup = 1;
      for (int i = 0; i < n; i++) S->ee[i] = 0.0;
      continue;
    }
    return retval;
  }
}

#define ERROR_TEST_FAIL 7
#define PREDICT_AGAIN 20

static int test_error(ida_t* S, double ck, double* err_k, double* err_km1) {
  const int n = S->n, kk = S->kk;
  double enorm_k = wrms(S, S->ee);
  *err_k = S->sigma[kk] * enorm_k;
  double terr_k = (kk + 1) * (*err_k);
  S->knew = kk;
  if (kk > 1) {
    for (int i = 0; i < n; i++) S->delta[i] = S->phi[kk][i] + S->ee[i];
    *err_km1 = S->sigma[kk - 1] * wrms(S, S->delta);
    double terr_km1 = kk * (*err_km1);
    if (kk > 2) {
      for (int i = 0; i < n; i++) S->delta[i] += S->phi[kk - 1][i];
      double err_km2 = S->sigma[kk - 2] * wrms(S, S->delta);
      double terr_km2 = (kk - 1) * err_km2;
      if (fmax(terr_km1, terr_km2) <= terr_k) S->knew = kk - 1;
    } else if (terr_km1 <= 0.5 * terr_k) S->knew = kk - 1;
  }
  return (ck * enorm_k > 1.0) ? ERROR_TEST_FAIL : 0;
}

static int handle_nflag(ida_t* S, int nflag, double err_k, double err_km1, int* ncf, int* nef) {
  S->phase = 1;
  if (nflag != ERROR_TEST_FAIL) {
    (*ncf)++; S->st[ST_NCFN]++;
    if (nflag < 0) return nflag;
    S->rr = 0.25;
    S->hh *= S->rr;
    if (*ncf < S->opt.maxncf) return PREDICT_AGAIN;
    return (nflag == NLS_LSETUP_RECVR) ? IDA_LSETUP_FAIL : IDA_CONV_FAIL;
  }
  (*nef)++; S->st[ST_NETF]++;
  if (*nef == 1) {
    double err_knew = (S->kk == S->knew) ? err_k : err_km1;
    S->kk = S->knew;
    S->rr = 0.9 * pow(2.0 * err_knew + 0.0001, -1.0 / (S->kk + 1));
    S->rr = fmax(0.25, fmin(0.9, S->rr));
    S->hh *= S->rr;
    return PREDICT_AGAIN;
  } else if (*nef == 2) {
    S->kk = S->knew; S->rr = 0.25; S->hh *= S->rr; return PREDICT_AGAIN;
  } else if (*nef < S->opt.maxnef) {
    S->kk = 1; S->rr = 0.25; S->hh *= S->rr; return PREDICT_AGAIN;
  }
  return IDA_ERR_FAIL;
}

static void complete_step(ida_t* S, double err_k, double err_km1) {
  const int n = S->n;
  S->nst++; S->st[ST_NST]++;
  int kdiff = S->kk - S->kused;
  S->kused = S->kk; S->hused = S->hh;
  if (S->knew == S->kk - 1 || S->kk == S->opt.maxord) S->phase = 1;
  if (S->phase == 0) {
    /* IDACompleteStep: "after the first step, the order is not increased, as not all of the necessary
     * information is available yet" -- the raise-and-double of phase 0 starts with the second step */
    if (S->nst > 1) {
      S->kk++;
      double hnew = 2.0 * S->hh;
      const double tmp = fabs(hnew) * S->hmax_inv;
      if (tmp > 1.0) hnew /= tmp;
      S->hh = hnew;
    }
  } else {
    enum { UNSET, LOWER, MAINTAIN, RAISE } action = UNSET;
    double err_kp1 = 0.0, err_knew;
    if (S->knew == S->kk - 1) action = LOWER;
    else if (S->kk == S->opt.maxord) action = MAINTAIN;
    else if (S->kk + 1 >= S->ns || kdiff == 1) action = MAINTAIN;
    if (action == UNSET) {
      for (int i = 0; i < n; i++) S->tempv[i] = S->ee[i] - S->phi[S->kk + 1][i];
      double enorm = wrms(S, S->tempv);
      err_kp1 = enorm / (S->kk + 2);
      double terr_k = (S->kk + 1) * err_k, terr_kp1 = (S->kk + 2) * err_kp1;
      if (S->kk == 1) action = (terr_kp1 >= 0.5 * terr_k) ? MAINTAIN : RAISE;
      else {
        double terr_km1 = S->kk * err_km1;
        if (terr_km1 <= fmin(terr_k, terr_kp1)) action = LOWER;
        else if (terr_kp1 >= terr_k) action = MAINTAIN;
        else action = RAISE;
      }
    }
    if (action == RAISE) { S->kk++; err_knew = err_kp1; }
    else if (action == LOWER) { S->kk--; err_knew = err_km1; }
    else err_knew = err_k;
    double hnew = S->hh;
    S->rr = pow(2.0 * err_knew + 0.0001, -1.0 / (S->kk + 1));
    if (S->rr >= 2.0) {
      hnew = 2.0 * S->hh;
      const double tmp = fabs(hnew) * S->hmax_inv;
      if (tmp > 1.0) hnew /= tmp;
    }
    else if (S->rr <= 1.0) { S->rr = fmax(0.5, fmin(0.9, S->rr)); hnew = S->hh * S->rr; }
    S->hh = hnew;
  }
  if (S->kused < S->opt.maxord) memcpy(S->phi[S->kused + 1], S->ee, sizeof(double) * n);
  for (int i = 0; i < n; i++) S->phi[S->kused][i] += S->ee[i];
  for (int j = S->kused - 1; j >= 0; j--)
    for (int i = 0; i < n; i++) S->phi[j][i] += S->phi[j + 1][i];
}

static int ida_step(ida_t* S) {
  double saved_t = S->tn, ck, err_k = 0.0, err_km1 = 0.0;
  int ncf = 0, nef = 0;
  if (S->nst == 0) {
    S->kk = 1; S->kused = 0; S->hused = 0.0; S->psi[0] = S->hh; S->cj = 1.0 / S->hh; S->phase = 0; S->ns = 0;
  }
  for (;;) {
    set_coeffs(S, &ck);
    int nflag = nls(S);
    if (nflag == NLS_SUCCESS) nflag = test_error(S, ck, &err_k, &err_km1);
    if (nflag != 0) {
      restore(S, saved_t);
      int kflag = handle_nflag(S, nflag, err_k, err_km1, &ncf, &nef);
      if (kflag != PREDICT_AGAIN) return kflag;
      if (S->nst == 0) {
        S->psi[0] = S->hh;
        for (int i = 0; i < S->n; i++) S->phi[1][i] *= S->rr;
      }
      if (!(fabs(S->hh) > 0.0) || S->tn + S->hh == S->tn) return IDA_ERR_FAIL;
      continue;
    }
    break;
  }
  complete_step(S, err_k, err_km1);
  return IDA_SUCCESS;
}

/* ------------------------------------------------------------------ root finding */
static void gfun(ida_t* S, double t, const double* y, double* g) {
  S->m->events(y, S->p, t, 0.0, g, S->tmp);
  S->st[ST_NGE]++;
}

static int rootfind(ida_t* S) {
  const int ne = S->m->n_ev;
  int imax = 0, zroot = 0, sgnchg = 0, side = 0, sideprev = -1;
  double maxfrac = 0.0, alph = 1.0;
  for (int i = 0; i < ne; i++) {
    if (!S->gactive[i]) continue;
    if (fabs(S->ghi[i]) == 0.0) zroot = 1;
    else if (S->glo[i] * S->ghi[i] < 0.0) {
      double gfrac = fabs(S->ghi[i] / (S->ghi[i] - S->glo[i]));
      if (gfrac > maxfrac) { sgnchg = 1; maxfrac = gfrac; imax = i; }
    }
  }
  if (!sgnchg) {
    S->trout = S->thi;
    for (int i = 0; i < ne; i++) S->grout[i] = S->ghi[i];
    if (!zroot) return 0;
    for (int i = 0; i < ne; i++) {
      S->iroots[i] = 0;
      if (S->gactive[i] && fabs(S->ghi[i]) == 0.0) S->iroots[i] = S->glo[i] > 0 ? -1 : 1;
    }
    return 1;
  }
  for (;;) {
    if (fabs(S->thi - S->tlo) <= S->ttol) break;
    if (sideprev == side) alph = (side == 2) ? alph * 2.0 : alph * 0.5;
    else alph = 1.0;
    double tmid = S->thi - (S->thi - S->tlo) * S->ghi[imax] / (S->ghi[imax] - alph * S->glo[imax]);
    if (fabs(tmid - S->tlo) < 0.5 * S->ttol) {
      double fracint = fabs(S->thi - S->tlo) / S->ttol;
      double fracsub = (fracint > 5.0) ? 0.1 : 0.5 / fracint;
      tmid = S->tlo + fracsub * (S->thi - S->tlo);
    }
    if (fabs(S->thi - tmid) < 0.5 * S->ttol) {
      double fracint = fabs(S->thi - S->tlo) / S->ttol;
      double fracsub = (fracint > 5.0) ? 0.1 : 0.5 / fracint;
      tmid = S->thi - fracsub * (S->thi - S->tlo);
    }
    get_solution(S, tmid, S->yy, S->yp);
    gfun(S, tmid, S->yy, S->grout);
    maxfrac = 0.0; zroot = 0; sgnchg = 0; sideprev = side;
    for (int i = 0; i < ne; i++) {
      if (!S->gactive[i]) continue;
      if (fabs(S->grout[i]) == 0.0) zroot = 1;
      else if (S->glo[i] * S->grout[i] < 0.0) {
        double gfrac = fabs(S->grout[i] / (S->grout[i] - S->glo[i]));
        if (gfrac > maxfrac) { sgnchg = 1; maxfrac = gfrac; imax = i; }
      }
    }
    if (sgnchg) {
      S->thi = tmid; for (int i = 0; i < ne; i++) S->ghi[i] = S->grout[i];
      side = 1;
      if (fabs(S->thi - S->tlo) <= S->ttol) break;
      continue;
    }
    if (zroot) {
      S->thi = tmid; for (int i = 0; i < ne; i++) S->ghi[i] = S->grout[i];
      break;
    }
    S->tlo = tmid; for (int i = 0; i < ne; i++) S->glo[i] = S->grout[i];
    side = 2;
    if (fabs(S->thi - S->tlo) <= S->ttol) break;
  }
  S->trout = S->thi;
  for (int i = 0; i < ne; i++) {
    S->grout[i] = S->ghi[i];
    S->iroots[i] = 0;
    if (!S->gactive[i]) continue;
    if (fabs(S->ghi[i]) == 0.0) S->iroots[i] = S->glo[i] > 0 ? -1 : 1;
    if (S->glo[i] * S->ghi[i] < 0.0) S->iroots[i] = S->glo[i] > 0 ? -1 : 1;
  }
  return 1;
}

/* IDARcheck3 (ONE_STEP task: thi = tn) */
static int rcheck3(ida_t* S) {
  const int ne = S->m->n_ev;
  S->thi = S->tn;
  get_solution(S, S->thi, S->yy, S->yp);
  gfun(S, S->thi, S->yy, S->ghi);
  S->ttol = (fabs(S->tn) + fabs(S->hh)) * UROUND * 100.0;
  int ier = rootfind(S);
  for (int i = 0; i < ne; i++) if (!S->gactive[i] && S->grout[i] != 0.0) S->gactive[i] = 1;
  S->tlo = S->trout;
  for (int i = 0; i < ne; i++) S->glo[i] = S->grout[i];
  if (ier == 0) return 0;
  get_solution(S, S->trout, S->yy, S->yp);
  return 1;
}

/* IDASolve(tout, ONE_STEP): returns IDA flag; *tret, yy (and yp) hold the returned point */
static int ida_solve_one_step(ida_t* S, double tout, double* tret) {
  const int ne = S->m->n_ev;
  if (S->nst == 0) {
    if (ewt_set(S, S->phi[0])) return -22;
    double tdist = fabs(tout - S->tn);
    if (S->opt.dt_init != 0.0) {
      S->hh = S->opt.dt_init;
    } else {
      S->hh = 0.001 * tdist;
      double ypnorm = wrms(S, S->phi[1]);
      if (ypnorm > 0.5 / S->hh) S->hh = 0.5 / ypnorm;
      if (tout < S->tn) S->hh = -S->hh;
    }
    { const double rh = fabs(S->hh) * S->hmax_inv; if (rh > 1.0) S->hh /= rh; }
    if (S->tstopset && (S->tn + S->hh - S->tstop) * S->hh > 0.0) S->hh = (S->tstop - S->tn) * (1.0 - 4.0 * UROUND);
    S->kk = 0; S->kused = 0;
    for (int i = 0; i < S->n; i++) S->phi[1][i] *= S->hh;
    if (ne > 0) { /* IDARcheck1 */
      S->tlo = S->tn; S->ttol = (fabs(S->tn) + fabs(S->hh)) * UROUND * 100.0;
      gfun(S, S->tlo, S->phi[0], S->glo);
      int zroot = 0;
      for (int i = 0; i < ne; i++) { S->gactive[i] = 1; S->iroots[i] = 0; if (fabs(S->glo[i]) == 0.0) { zroot = 1; S->gactive[i] = 0; } }
      if (zroot) {
        double hratio = fmax(S->ttol / fabs(S->hh), 0.1), smallh = hratio * S->hh;
        for (int i = 0; i < S->n; i++) S->yy[i] = S->phi[0][i] + hratio * S->phi[1][i];
        gfun(S, S->tlo + smallh, S->yy, S->ghi);
        for (int i = 0; i < ne; i++) if (!S->gactive[i] && fabs(S->ghi[i]) != 0.0) { S->gactive[i] = 1; S->glo[i] = S->ghi[i]; }
      }
    }
  } else {
    /* IDAStopTest1, ONE_STEP */
    if ((S->tn - S->tretlast) * S->hh > 0.0) {
      get_solution(S, S->tn, S->yy, S->yp); *tret = S->tretlast = S->tn; return IDA_SUCCESS;
    }
    if (S->tstopset) {
      double troundoff = 100.0 * UROUND * (fabs(S->tn) + fabs(S->hh));
      if (fabs(S->tn - S->tstop) <= troundoff) {
        get_solution(S, S->tstop, S->yy, S->yp); *tret = S->tretlast = S->tstop; S->tstopset = 0; return IDA_TSTOP_RETURN;
      }
      if ((S->tn + S->hh - S->tstop) * S->hh > 0.0) S->hh = (S->tstop - S->tn) * (1.0 - 4.0 * UROUND);
    }
  }
  /* one pass of the main loop */
  if (S->nst > 0 && ewt_set(S, S->phi[0])) return -22;
  {
    double nrm = wrms(S, S->phi[0]);
    if (UROUND * nrm > 1.0) return IDA_TOO_MUCH_ACC;
  }
  int sflag = ida_step(S);
  if (sflag != IDA_SUCCESS) {
    *tret = S->tretlast = S->tn;
    get_solution(S, S->tn, S->yy, S->yp);
    return sflag;
  }
  if (S->tstopset) {
    double troundoff = 100.0 * UROUND * (fabs(S->tn) + fabs(S->hh));
    if (fabs(S->tn - S->tstop) <= troundoff) S->tn = S->tstop;
  }
  if (ne > 0) {
    if (rcheck3(S)) { S->irfnd = 1; *tret = S->tretlast = S->tlo; return IDA_ROOT_RETURN; }
  }
  /* IDAStopTest2, ONE_STEP */
  if (S->tstopset) {
    double troundoff = 100.0 * UROUND * (fabs(S->tn) + fabs(S->hh));
    if (fabs(S->tn - S->tstop) <= troundoff) {
      get_solution(S, S->tstop, S->yy, S->yp); *tret = S->tretlast = S->tstop; S->tstopset = 0; return IDA_TSTOP_RETURN;
    }
    if ((S->tn + S->hh - S->tstop) * S->hh > 0.0) S->hh = (S->tstop - S->tn) * (1.0 - 4.0 * UROUND);
  }
  get_solution(S, S->tn, S->yy, S->yp); *tret = S->tretlast = S->tn;
  return IDA_SUCCESS;
}

/* ------------------------------------------------------------------ consistent initialisation */
/* y' = f(t, y) on differential rows, 0 on algebraic rows (ConsistentInitializationODE / RecoverYp) */
static void recover_yp(ida_t* S, double t, const double* y, double* yp) {
  S->m->res(y, S->p, t, 0.0, yp, S->tmp);
  for (int i = S->m->n_diff; i < S->n; i++) yp[i] = 0.0;
}

static double alg_res_norm(ida_t* S, double t, double* y) {
  const int nd = S->m->n_diff, na = S->n - nd;
  memcpy(y + nd, S->xa, sizeof(double) * na);
  S->m->res(y, S->p, t, 0.0, S->tempv, S->tmp);
  double mx = 0.0;
  for (int i = 0; i < na; i++) { S->resa[i] = S->tempv[nd + i]; double a = fabs(S->resa[i]); if (a > mx || !(a == a)) mx = a; }
  return mx;
}

/* NonlinearSolver::RunNewtonLoop on the algebraic sub-block; returns 1 on success.  y is updated in place. */
static int newton_ic(ida_t* S, double t, double* y, double* init_norm, double* final_norm) {
  const pbo_model* m = S->m;
  const int nd = m->n_diff, na = S->n - nd;
  double delnorm = INFINITY, prev_res_norm = INFINITY;
  int converged = 0;
  memcpy(S->xa, y + nd, sizeof(double) * na);
  for (int i = 0; i < na; i++) S->ewta[i] = 1.0 / (S->opt.rtol * fabs(S->xa[i]) + S->atol[nd + i]);
  for (int iter = 0; iter < S->opt.max_iter_ic; iter++) {
    double res_norm = alg_res_norm(S, t, y);
    if (iter == 0) *init_norm = res_norm;
    *final_norm = res_norm;
    /* Jacobian sub-block d g / d y_a at cj = 0 */
    m->jac(y, S->p, t, 0.0, S->jacv, S->tmp);
    memset(S->Jaa, 0, sizeof(double) * (size_t)na * na);
    for (int k = 0; k < m->nnz; k++) {
      int r = m->jac_rows[k], c = m->jac_cols[k];
      if (r >= nd && c >= nd) S->Jaa[(size_t)(r - nd) * na + (c - nd)] += S->jacv[k];
    }
    if (dense_factor(S->Jaa, na, S->ipiv_a)) return 0;
    memcpy(S->dela, S->resa, sizeof(double) * na);
    dense_solve(S->Jaa, na, S->ipiv_a, S->dela);
    double s = 0.0;
    for (int i = 0; i < na; i++) { double w = S->dela[i] * S->ewta[i]; s += w * w; }
    delnorm = sqrt(s / na);
    if (delnorm <= S->opt.epsNewtIC) {
      converged = 1;
      if (delnorm <= S->opt.step_tol_ic) {
        for (int i = 0; i < na; i++) S->xa[i] -= S->dela[i];
        memcpy(y + nd, S->xa, sizeof(double) * na);
        return 1;
      }
      if (iter > 0 && res_norm >= prev_res_norm) { memcpy(S->xa, S->xsave, sizeof(double) * na); memcpy(y + nd, S->xa, sizeof(double) * na); return 1; }
    }
    prev_res_norm = res_norm;
    memcpy(S->xsave, S->xa, sizeof(double) * na);
    double alpha = 1.0;
    for (int ls = 0; ls < S->opt.max_backtracks_ic; ls++) {
      for (int i = 0; i < na; i++) S->xa[i] = S->xsave[i] - alpha * S->dela[i];
      double trial = alg_res_norm(S, t, y);
      if (trial <= (1.0 - alpha * 0.5) * res_norm) break;
      if (alpha * delnorm <= S->opt.step_tol_ic) break;
      alpha *= 0.5;
    }
  }
  memcpy(y + nd, S->xa, sizeof(double) * na);
  return converged;
}

/* ConsistentInitialization (IDAKLUSolverOpenMP.inl:787-873) + IDAReInit */
static void consistent_init(ida_t* S, double t, double* y, double* yp) {
  if (S->n > S->m->n_diff) {
    double in = 0, fn = 0;
    memcpy(S->savres, y, sizeof(double) * S->n);
    int ok = newton_ic(S, t, y, &in, &fn);
    if (!(ok || fn <= in)) memcpy(y, S->savres, sizeof(double) * S->n);
    else recover_yp(S, t, y, yp);
  } else {
    recover_yp(S, t, y, yp);
  }
  reinit(S, t, y, yp);
}

/* ------------------------------------------------------------------ the PyBaMM driver loop */
static void store_point(ida_t* S, int save_vars, int n_ret, double t, const double* y, int* i_save, int n_max,
                        double* t_out, double* y_out) {
  if (*i_save >= n_max) return;
  t_out[*i_save] = t;
  if (save_vars) S->m->vars(y, S->p, t, 0.0, y_out + (size_t)(*i_save) * n_ret, S->tmp);
  else memcpy(y_out + (size_t)(*i_save) * n_ret, y, sizeof(double) * S->n);
  (*i_save)++;
}

static double perturb_time(double t) {
  const double delta = sqrt(DBL_EPSILON);
  return (1.0 + delta) * t + delta;
}

static int solve_one(ida_t* S, const double* t_eval, int n_eval, const double* t_interp, int n_interp,
                     const double* y0_in, int save_vars, int n_max, double* t_out, double* y_out, int* n_out,
                     double* y_term) {
  const pbo_model* m = S->m;
  const int n = S->n, ne = m->n_ev;
  const int n_ret = save_vars ? m->n_var : n;
  memset(S->st, 0, sizeof(S->st));
  double* y = S->yy; double* yp = S->yp;
  if (y0_in) memcpy(y, y0_in, sizeof(double) * n);
  else m->ic(NULL, S->p, t_eval[0], 0.0, y, S->tmp);
  const double t0 = t_eval[0], tf = t_eval[n_eval - 1];
  const double tf_perturbed = perturb_time(tf);
  int i_save = 0, i_eval = 1, i_interp = 0;
  double t_eval_next = t_eval[1];
  recover_yp(S, t0, y, yp);      /* python side: ydot0 = rhs(t0, y0) (idaklu_solver.py:967-1013) */
  reinit(S, t0, y, yp);
  consistent_init(S, t0, y, yp);
  S->tstop = t_eval_next; S->tstopset = 1;
  int retval = IDA_SUCCESS;
  double t_val = t0, t_prev = t0;
  if (ne > 0) {
    double g[16];
    m->events(y, S->p, t0, 0.0, g, S->tmp);
    int trig = 0;
    for (int i = 0; i < ne; i++) trig |= (g[i] <= 0.0);
    if (trig) {
      store_point(S, save_vars, n_ret, t0, y, &i_save, n_max, t_out, y_out);
      *n_out = i_save; if (y_term) memcpy(y_term, y, sizeof(double) * n);
      return IDA_ROOT_RETURN;
    }
  }
  retval = ida_solve_one_step(S, tf_perturbed, &t_val);
  if (retval < 0) { *n_out = 0; return retval; }
  /* initial point from the interpolant at t0 (GetSolutionFull(t0)) */
  get_solution(S, t0, S->tempv, NULL);
  store_point(S, save_vars, n_ret, t0, S->tempv, &i_save, n_max, t_out, y_out);
  long guard = 0;
  for (;;) {
    if (retval < 0) break;
    if (t_prev == t_val) { retval = IDA_ERR_FAIL; break; }
    int hit_teval = retval == IDA_TSTOP_RETURN;
    int hit_final = t_val >= tf || (hit_teval && i_eval == n_eval);
    int hit_event = retval == IDA_ROOT_RETURN;
    int hit_adaptive = (n_interp == 0) && retval == IDA_SUCCESS;
    int final_step = hit_final || hit_event;
    if (n_interp > 0 && i_interp < n_interp && t_interp[i_interp] >= t_prev) {
      while (i_interp < n_interp && t_interp[i_interp] <= t_val) {
        get_solution(S, t_interp[i_interp], S->tempv, NULL);
        store_point(S, save_vars, n_ret, t_interp[i_interp], S->tempv, &i_save, n_max, t_out, y_out);
        i_interp++;
      }
    }
    if (hit_adaptive || hit_teval || hit_event || hit_final) {
      if (i_save == 0 || t_val != t_out[i_save - 1]) {
        get_solution(S, t_val, S->tempv, NULL);
        store_point(S, save_vars, n_ret, t_val, S->tempv, &i_save, n_max, t_out, y_out);
      }
    }
    if (final_step) break;
    if (hit_teval) {
      i_eval++;
      t_eval_next = t_eval[i_eval];
      get_solution(S, t_val, y, yp);
      reinit(S, t_val, y, yp);
      consistent_init(S, t_val, y, yp);
      S->tstop = t_eval_next; S->tstopset = 1;
    }
    t_prev = t_val;
    if (++guard > S->opt.max_num_steps) { retval = IDA_TOO_MUCH_WORK; break; }
    retval = ida_solve_one_step(S, tf_perturbed, &t_val);
  }
  *n_out = i_save;
  if (y_term) get_solution(S, t_val, y_term, NULL);
  return retval;
}

/* ------------------------------------------------------------------ allocation + batch entry point */
static ida_t* ida_alloc(const pbo_model* m, const pbo_options* opt, const double* atol, const int* perm, const int* perm_c,
                        int lu_mode) {
  ida_t* S = (ida_t*)calloc(1, sizeof(ida_t));
  const int n = m->n;
  S->m = m; S->opt = *opt; S->atol = atol; S->perm = perm; S->n = n;
  S->hmax_inv = (opt->dt_max > 0.0) ? 1.0 / opt->dt_max : 0.0;
  S->iperm = (int*)malloc(sizeof(int) * n);
  for (int i = 0; i < n; i++) S->iperm[perm ? perm[i] : i] = i;
  for (int j = 0; j < MXORDP1; j++) S->phi[j] = (double*)calloc(n, sizeof(double));
  double** vecs[] = {&S->ewt, &S->ee, &S->delta, &S->yypredict, &S->yppredict, &S->yy, &S->yp, &S->tempv, &S->tempv2, &S->savres, &S->bw};
  for (unsigned k = 0; k < sizeof(vecs) / sizeof(vecs[0]); k++) *vecs[k] = (double*)calloc(n, sizeof(double));
  S->jacv = (double*)calloc(m->nnz > 0 ? m->nnz : 1, sizeof(double));
  S->tmp = (double*)calloc(m->n_tmp > 0 ? m->n_tmp : 1, sizeof(double));
  S->p = (double*)calloc(m->n_p > 0 ? m->n_p : 1, sizeof(double));
  int ne = m->n_ev > 0 ? m->n_ev : 1;
  S->glo = (double*)calloc(ne, sizeof(double)); S->ghi = (double*)calloc(ne, sizeof(double)); S->grout = (double*)calloc(ne, sizeof(double));
  S->gactive = (int*)calloc(ne, sizeof(int)); S->iroots = (int*)calloc(ne, sizeof(int));
  S->lu_mode = (lu_mode && perm && perm_c) ? 1 : 0;
  if (S->lu_mode) splu_init(&S->SP, n, m->nnz, m->jac_rows, m->jac_cols, perm, perm_c);
  /* band widths of the permuted pattern */
  int kl = 0, ku = 0;
  for (int k = 0; k < m->nnz; k++) {
    int i = S->iperm[m->jac_rows[k]], j = S->iperm[m->jac_cols[k]];
    if (i - j > kl) kl = i - j;
    if (j - i > ku) ku = j - i;
  }
  if (S->lu_mode) kl = ku = 0;   /* the band store is not used */
  S->B.n = n; S->B.kl = kl; S->B.ku = ku; S->B.W = 2 * kl + ku + 1;
  S->B.a = (double*)calloc((size_t)n * S->B.W, sizeof(double));
  S->B.ipiv = (int*)calloc(n, sizeof(int));
  int na = n - m->n_diff;
  if (na > 0) {
    S->Jaa = (double*)calloc((size_t)na * na, sizeof(double)); S->ipiv_a = (int*)calloc(na, sizeof(int));
    S->xa = (double*)calloc(na, sizeof(double)); S->xsave = (double*)calloc(na, sizeof(double));
    S->resa = (double*)calloc(na, sizeof(double)); S->dela = (double*)calloc(na, sizeof(double)); S->ewta = (double*)calloc(na, sizeof(double));
  }
  return S;
}
static void ida_free(ida_t* S) {
  for (int j = 0; j < MXORDP1; j++) free(S->phi[j]);
  free(S->ewt); free(S->ee); free(S->delta); free(S->yypredict); free(S->yppredict); free(S->yy); free(S->yp);
  free(S->tempv); free(S->savres); free(S->bw); free(S->jacv); free(S->tmp); free(S->p);
  free(S->glo); free(S->ghi); free(S->grout); free(S->gactive); free(S->iroots);
  free(S->B.a); free(S->B.ipiv); free(S->iperm); free(S->tempv2);
  if (S->lu_mode) splu_free(&S->SP);
  free(S->Jaa); free(S->ipiv_a); free(S->xa); free(S->xsave); free(S->resa); free(S->dela); free(S->ewta);
  free(S);
}

/*
 * Batch entry (mirrors IDAKLUSolverGroup::solve's data contract, IDAKLUSolverGroup.cpp:6-170):
 *  t_eval sorted unique (n_eval >= 2), t_interp sorted unique with t_eval removed (may be empty -> adaptive steps saved),
 *  inputs [n_sys][n_p], y0 [n_sys][n] or NULL (-> model IC), atol [n],
 *  outputs t_out [n_sys][n_max], y_out [n_sys][n_max][n_ret], n_out [n_sys], flag [n_sys], y_term [n_sys][n] (or NULL),
 *  stats [n_sys][8].
 */
int pbo_solve(int n_sys, const double* t_eval, int n_eval, const double* t_interp, int n_interp,
              const double* inputs, const double* y0, const double* atol, const pbo_options* opt, const int* perm,
              int save_vars, int n_max, double* t_out, double* y_out, int* n_out, int* flag, double* y_term,
              long* stats, int n_threads, const int* perm_c, int lu_mode) {
  const pbo_model* m = pbo_get_model();
  if (n_eval < 2) return -1;
  const int n_ret = save_vars ? m->n_var : m->n;
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel
  {
    ida_t* S = ida_alloc(m, opt, atol, perm, perm_c, lu_mode);
#pragma omp for schedule(dynamic, 1)
    for (int s = 0; s < n_sys; s++) {
      for (int k = 0; k < m->n_p; k++) S->p[k] = inputs[(size_t)s * m->n_p + k];
      flag[s] = solve_one(S, t_eval, n_eval, t_interp, n_interp, y0 ? y0 + (size_t)s * m->n : NULL, save_vars, n_max,
                          t_out + (size_t)s * n_max, y_out + (size_t)s * n_max * n_ret, n_out + s,
                          y_term ? y_term + (size_t)s * m->n : NULL);
      if (stats) for (int k = 0; k < 8; k++) stats[(size_t)s * 8 + k] = S->st[k];
    }
    ida_free(S);
  }
  return 0;
}

/* direct access to the model functions (tests: generated C == pybamm's own evaluation) */
void pbo_eval(int which, const double* y, const double* p, double t, double cj, double* out) {
  const pbo_model* m = pbo_get_model();
  double* tmp = (double*)calloc(m->n_tmp > 0 ? m->n_tmp : 1, sizeof(double));
  switch (which) {
    case 0: m->res(y, p, t, cj, out, tmp); break;
    case 1: m->jac(y, p, t, cj, out, tmp); break;
    case 2: m->events(y, p, t, cj, out, tmp); break;
    case 3: m->ic(y, p, t, cj, out, tmp); break;
    case 4: m->vars(y, p, t, cj, out, tmp); break;
  }
  free(tmp);
}
void pbo_sizes(int* out) {
  const pbo_model* m = pbo_get_model();
  out[0] = m->n; out[1] = m->n_diff; out[2] = m->n_p; out[3] = m->n_ev; out[4] = m->nnz; out[5] = m->n_var;
}
