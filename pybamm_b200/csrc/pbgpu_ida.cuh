// pbgpu_ida.cuh -- batched variable-order BDF (IDA-style) DAE integrator, one team per system.
//
// Device-side replacement of the per-system solver object of the reference:
//   IDAKLUSolverOpenMP::solve + helpers  (idaklu_source/IDAKLUSolverOpenMP.inl:327-699, 787-928)
//   the SUNDIALS IDAS calls it makes     (IDASolve/IDAReInit/IDAGetDky/IDASetStopTime; call sites
//                                         IDAKLUSolverOpenMP.inl:383,431,573,632,780)
//   SUNLinSol_KLU                        (IDAKLUSolverOpenMP_solvers.hpp:29-33) -> static-pattern sparse LU
//   NonlinearSolver (IC Newton)          (idaklu_source/NonlinearSolver.inl:100-177, AlgebraicICBuilder.inl:52-119)
//
// Design (B200): every per-system vector (BDF history phi[0..5], weights, Newton work vectors,
// the L+U values, code-generator temporaries) lives in the team's shared memory for the whole
// integration; HBM is touched only for the inputs and the requested outputs.  Vector work is
// strided over the team's threads, scalar control state is replicated in registers (every thread
// takes the same decisions from bitwise-identical reductions), the BDF coefficient arrays are
// written by thread 0.  Teams pull systems from an atomic queue, so a team that finishes (event,
// final time, failure) immediately starts the next system: no lane idles on a terminated system.
//
// Included after pbgpu_team.cuh and the generated model header (PBM_* macros, pbm_* functions).
#pragma once

#define PB_MXORDP1 6
#define PB_UROUND 2.220446049250313e-16

#define PB_IDA_SUCCESS 0
#define PB_IDA_TSTOP_RETURN 1
#define PB_IDA_ROOT_RETURN 2
#define PB_IDA_TOO_MUCH_WORK (-1)
#define PB_IDA_TOO_MUCH_ACC (-2)
#define PB_IDA_ERR_FAIL (-3)
#define PB_IDA_CONV_FAIL (-4)
#define PB_IDA_LSETUP_FAIL (-6)
#define PB_IDA_ILL_INPUT (-22)

#define PB_FAC_PAD ((unsigned int)PBM_NNZ_LU | ((unsigned int)PBM_NNZ_LU << 16))   /* zero slot times zero slot */
#define PB_AUX_U 0xFFFF
#define PB_AUX_DIAG 0xFFFE
#define PB_AUX_NEG 0xFFFD   /* out = -sum(pairs): the old value of the slot is not part of the result */

#define PB_NEV_PAD ((PBM_NEV > 0) ? PBM_NEV : 1)

// Scalar divisions and square roots of the step / order control: the shared out-of-line copies (PB_DIV, PB_SQRT) instead of an
// inline expansion of twenty-odd instructions at every site (-DPB_SCALAR_INLINE: the plain operators).
#ifdef PB_VECTOR_DIV_INLINE   /* the divisions of the error-weight pass and of the factorisation: out of line as well (DFN 5 562 -> 5 609) */
#define PB_VDIV(a, b) ((a) / (b))
#else
#define PB_VDIV(a, b) PB_DIV(a, b)
#endif
#ifdef PB_SCALAR_INLINE
#define PB_IDIV(a, b) ((a) / (b))
#define PB_RMS(s) sqrt((s) / PBM_N)
#else
#define PB_IDIV(a, b) PB_DIV(a, b)
#define PB_RMS(s) PB_SQRT(PB_DIV(s, (double)PBM_N))
#endif


// ---- memory layout (in doubles) ---------------------------------------------------------------------
// Shared memory always holds what is accessed at random inside the latency-critical loops: y (stencil reads of the
// generated functions), delta / w (triangular solves), the L+U values, small temporaries, the packed solve stream.
// The vectors that are only streamed (BDF history phi[0..5], ewt, ee, yypred, yppred) and the temporaries of the
// Jacobian program live in shared memory for small models (PB_GLOBAL_VEC == 0: SPM, SPMe) and in a per-team global
// scratch block (L2-resident: teams x ~100 KB) for large ones (PB_GLOBAL_VEC == 1: DFN), so that two teams fit per SM.
#ifndef PB_GLOBAL_VEC
#define PB_GLOBAL_VEC 0
#endif
// Output variables of one-warp models are evaluated PB_VARS_BATCH output points at a time (lanes = points): the terminal
// voltage of the SPM / SPMe is one long scalar expression, a serial FP64 chain on ONE lane (8.6k cycles for the SPM against
// 2.5k for its residual), and output points were a third of the SPM integration.  The interpolated states of the pending
// points, their times and the temporaries of the batched evaluation live in a per-team block of global scratch (L2):
// [K][N] states, [K] times, [K][PBM_NTMP_VARS] temporaries.
#if defined(PBM_VARS_MULTI) && !defined(PB_EMU) && !PB_GLOBAL_VEC && !defined(PB_NO_VARS_BATCH)
#ifndef PB_VARS_BATCH
#define PB_VARS_BATCH 32   /* measured at 10 000 sets: SPM 125.0k (off) 130.1k (8) 138.4k (16) 145.6k (32) sims/s, SPMe 19.6k 20.2k 20.9k 21.3k */
#endif
#else
#undef PB_VARS_BATCH
#define PB_VARS_BATCH 1
#endif
#if PB_VARS_BATCH > 1
#define PB_VB_DOUBLES (PB_VARS_BATCH * (PBM_N + 1 + PBM_NTMP_VARS))
#else
#define PB_VB_DOUBLES 0
#endif
#define PB_GOFF_PHI 0
#define PB_GOFF_EWT (PB_GOFF_PHI + PB_MXORDP1 * PBM_N)
#define PB_GOFF_EE (PB_GOFF_EWT + PBM_N)
#define PB_GOFF_YYPRED (PB_GOFF_EE + PBM_N)
#define PB_GOFF_YPPRED (PB_GOFF_YYPRED + PBM_N)
#define PB_GOFF_TMPJ (PB_GOFF_YPPRED + PBM_N)
#ifndef PBM_DENSE_N
#define PBM_DENSE_N 0   /* rows of the dense top block of the elimination tree (symbolic_lu._dense_top) */
#endif
#define PB_GOFF_WD (PB_GOFF_TMPJ + PBM_NTMP_JAC)   /* W = (L_TT U_TT)^-1, row-major */
// PB_GLOBAL_VEC == 2 (models of several thousand states: the DFN with 40-point meshes and a lumped thermal state, 3 523
// states, 140 KB of L+U values): one 512-thread team per SM, and shared memory keeps only what the triangular solves and
// the stencil reads touch -- y, w, the L+U values.  The residual vector, the correction, the error weights and the solve
// stream join the streamed vectors in L2.
#if PB_GLOBAL_VEC == 2
#define PB_GOFF_DELTA (PB_GOFF_WD + PBM_DENSE_N * PBM_DENSE_N)
#define PB_GVEC_DOUBLES (PB_GOFF_DELTA + PBM_N)
#define PB_OFF_YY 0
#define PB_SCRATCH_DOUBLES PB_GVEC_DOUBLES
#else
#define PB_GVEC_DOUBLES (PB_GOFF_WD + PBM_DENSE_N * PBM_DENSE_N)
#if PB_GLOBAL_VEC
#define PB_OFF_DELTA 0
#define PB_SCRATCH_DOUBLES PB_GVEC_DOUBLES
#else
#define PB_OFF_DELTA PB_GVEC_DOUBLES
#define PB_SCRATCH_DOUBLES PB_VB_DOUBLES
#endif
#define PB_OFF_YY (PB_OFF_DELTA + PBM_N)
#endif
// The temporaries of the generated functions (residual, events, output variables) ALIAS the solve's w scratch: no
// generated function runs while w is live (w holds a right-hand side only between the Newton loop's rhs pass and its
// update pass, and inside pb_lu_solve / pb_lu_factor).  tmp[0, PB_TMP_EXTRA) is private, the rest is w[0, N); w[N], the
// zero operand of the padded solve records, is never touched.
#define PB_TMP_EXTRA ((PBM_NTMP_SMALL > PBM_N) ? (PBM_NTMP_SMALL - PBM_N) : 0)
#define PB_OFF_TMP (PB_OFF_YY + PBM_N)
#define PB_OFF_W (PB_OFF_TMP + PB_TMP_EXTRA)
#define PB_OFF_LU (PB_OFF_W + PBM_N + 1)
#define PB_OFF_P (PB_OFF_LU + PBM_NNZ_LU + 3)
#define PB_OFF_G (PB_OFF_P + PBM_NP)
#define PB_OFF_COEF (PB_OFF_G + 3 * PB_NEV_PAD)
#define PB_OFF_RED (PB_OFF_COEF + 32)
#define PB_OFF_STATS (PB_OFF_RED + 64)
#define PB_OFF_NPG (PB_OFF_STATS + 8)                          /* NoProgressGuard window: 32 step sizes + verdict */
#define PB_NPG_MAX 32
#define PB_OFF_SLV (PB_OFF_NPG + PB_NPG_MAX + 2)               /* packed solve stream (u32 words) */
#if PB_GLOBAL_VEC == 2
#define PB_OFF_FAC PB_OFF_SLV                                  /* the solve stream is read from the global table */
#else
#define PB_OFF_FAC (PB_OFF_SLV + (PBM_SLV_ALLOC + 1) / 2)      /* packed factor stream when it fits (PBM_FAC_SMEM) */
#endif
#if PBM_FAC_SMEM
#define PB_OFF_END0 (PB_OFF_FAC + (PBM_FAC_WORDS + 1) / 2)
#else
#define PB_OFF_END0 PB_OFF_FAC
#endif
// Large models: the correction ee and the error weights ewt, read by every Newton iteration and every norm, stay in
// shared memory as well; the history, the predictors and the Jacobian temporaries are the ones streamed from L2.
// (DFN: 113.9 KB per team; two teams and their 1 KB reservations still fit the 228 KB of an SM.)
#if PB_GLOBAL_VEC == 1
#define PB_OFF_EE PB_OFF_END0
#define PB_OFF_EWT_S (PB_OFF_EE + PBM_N)
#define PB_SMEM_DOUBLES (PB_OFF_EWT_S + PBM_N)
#else
#define PB_SMEM_DOUBLES PB_OFF_END0
#endif
#define PB_SMEM_BYTES (PB_SMEM_DOUBLES * 8)

// Strided vector loops, written so that every load of a pass is in flight before the first use: the trip count is a
// compile-time constant (elements per thread), the body is fully unrolled and guarded.  The streamed vectors of the
// large models live in L2 (PB_GLOBAL_VEC): a pass then costs about one L2 round trip instead of one per element.
#define PB_VPT ((PBM_N + PB_TEAM - 1) / PB_TEAM)
#ifdef PB_VEC_ROLLED   /* A/B switch: rolled loops (a quarter of the code, one L2 round trip per trip) */
#define PB_VEC_FOR(k, i) _Pragma("unroll 1") for (int k = 0, i = PB_TID; k < PB_VPT; k++, i += PB_TEAM)
#else
#define PB_VEC_FOR(k, i) _Pragma("unroll") for (int k = 0, i = PB_TID; k < PB_VPT; k++, i += PB_TEAM)
#endif

enum { PB_ST_NST, PB_ST_NRE, PB_ST_NSETUPS, PB_ST_NETF, PB_ST_NNI, PB_ST_NCFN, PB_ST_NJE, PB_ST_NGE };

struct pb_sys {
  double* sm;          // base of the team's shared memory
  double* gv;          // base of the streamed vectors: == sm (PB_GLOBAL_VEC 0) or the team's global scratch block
  double* vb;          // pending output points of the batched variable evaluation (global scratch), PB_VARS_BATCH > 1
  double* vb_row0;     // output row of the first pending point (the rows of a system are consecutive)
  int vb_n;            // pending points
  const unsigned int* slv;  // solve stream (shared memory)
  const unsigned int* fac;  // factor stream (shared memory when PBM_FAC_SMEM, else the global table)
  pb_team_ctx team;
  volatile unsigned int* al_me;     // phase-alignment mailbox of this team / of the other team on the SM (pb_align), or null
  volatile unsigned int* al_other;
  unsigned int al_count, al_seen;   // meaningful in thread 0 only
  const double* atol;  // global
  // options (copied from pbgpu_options)
  double rtol, epcon, epsNewtIC, step_tol_ic, dt_init, hmax_inv;
  int maxord, maxnef, maxcor, maxncf, max_iter_ic, max_backtracks_ic, scalesol;
  long long max_num_steps;
  // IDA scalar memory (replicated in every thread)
  double tn, hh, hused, cj, cjlast, cjold, cjratio, ss, rr, epsNewt, toldel;
  double tstop, tretlast, tlo, thi, trout, ttol;
  int kk, kused, knew, phase, ns, tstopset;
  long long nst;
  unsigned gactive;
};

#define S_PHI(S, j) ((S).gv + PB_GOFF_PHI + (j) * PBM_N)
#if PB_GLOBAL_VEC == 1
#define S_EE(S) ((S).sm + PB_OFF_EE)
#define S_EWT(S) ((S).sm + PB_OFF_EWT_S)
#else
#define S_EE(S) ((S).gv + PB_GOFF_EE)
#define S_EWT(S) ((S).gv + PB_GOFF_EWT)
#endif
#define S_YYPRED(S) ((S).gv + PB_GOFF_YYPRED)
#define S_YPPRED(S) ((S).gv + PB_GOFF_YPPRED)
#define S_TMPJ(S) ((S).gv + PB_GOFF_TMPJ)
#define S_WD(S) ((S).gv + PB_GOFF_WD)
#if PB_GLOBAL_VEC == 2
#define S_DELTA(S) ((S).gv + PB_GOFF_DELTA)
#else
#define S_DELTA(S) ((S).sm + PB_OFF_DELTA)
#endif
#define S_YY(S) ((S).sm + PB_OFF_YY)
#define S_W(S) ((S).sm + PB_OFF_W)
#define S_TMP(S) ((S).sm + PB_OFF_TMP)
#define S_LU(S) ((S).sm + PB_OFF_LU)
#define S_P(S) ((S).sm + PB_OFF_P)
#define S_GLO(S) ((S).sm + PB_OFF_G)
#define S_GHI(S) ((S).sm + PB_OFF_G + PB_NEV_PAD)
#define S_GROUT(S) ((S).sm + PB_OFF_G + 2 * PB_NEV_PAD)
#define S_PSI(S) ((S).sm + PB_OFF_COEF)
#define S_ALPHA(S) ((S).sm + PB_OFF_COEF + 6)
#define S_BETA(S) ((S).sm + PB_OFF_COEF + 12)
#define S_SIGMA(S) ((S).sm + PB_OFF_COEF + 18)
#define S_GAMMA(S) ((S).sm + PB_OFF_COEF + 24)
#define S_STATS(S) ((S).sm + PB_OFF_STATS)
#define S_NPG(S) ((S).sm + PB_OFF_NPG)

#ifdef PB_PAIR_DEBUG
#define PB_STAT_KEEP(k) ((k) < 4)   /* slots 4..7 carry the join statistics of tools/gpu_pair_debug.py */
#else
#define PB_STAT_KEEP(k) 1
#endif
#ifndef PB_PROFILE
#define PB_STAT(S, k)                        \
  do {                                       \
    if (PB_TID == 0 && PB_STAT_KEEP(k)) S_STATS(S)[k] += 1.0;   \
  } while (0)
#define PB_T0(v) ((void)0)
#define PB_T1(S, v, k) ((void)0)
#else
// Profiling build (tools/gpu_phases.py): the 8 stats slots hold clock64() cycles per phase instead of the IDA
// counters: 0 residual, 1 triangular solves, 2 Jacobian evaluation, 3 LU factorisation, 4 Newton vector work,
// 5 step control (coefficients, error test, history update), 6 events + root finding, 7 output points.
#define PB_STAT(S, k) ((void)0)
#define PB_T0(v) const long long v = clock64()
#define PB_T1(S, v, k)                                                   \
  do {                                                                   \
    if (PB_TID == 0) S_STATS(S)[k] += (double)(clock64() - (v));         \
  } while (0)
#endif
#if defined(PB_PROFILE) && PB_PROFILE == 2
#define PB_T1X(S, v, k) ((void)0)
#else
#define PB_T1X(S, v, k) PB_T1(S, v, k)
#endif

// ---- static-pattern sparse LU (packed streams from pybamm_b200/symbolic_lu.py: build_streams) -------
// The factor stream is a sequence of records [meta][task x nt][operand x nt x ml] walked with one running offset:
//   meta = nt | ml << 16 | flags << 28;  flags: 1 END of a dependency level, 4 team-wide level, 8 GROUP mode
//   task = out_slot | aux << 16,  operand = lu_slot | lu_slot << 16
// Padding operands point at lu[PBM_NNZ_LU] = 0 / w[PBM_N] = 0; lu[PBM_NNZ_LU + 1] = 1.
// The triangular solves (1000 per simulation against 40 factorisations) use the SEQ stream further down, which
// lives in shared memory.
#define PB_F_END 1u
#define PB_F_BWD 2u
#define PB_F_CHAINS 2u   /* factor stream: the chain record (no direction there) */
#define PB_F_TEAM 4u
#define PB_F_GROUP 8u

// Where the factor stream is read from:
//   PBM_FAC_SMEM   the whole stream sits in the team's shared memory (small models);
//   PB_FAC_STAGE   large models: the stream (DFN: 95 KB) stays in global memory and is *staged* through the solve's w
//                  scratch (free during a factorisation) in chunks of whole records (table pbm_fac_chunk, built by
//                  codegen).  The next chunk travels from L2 into registers while the current one is being processed,
//                  so a record costs shared-memory round trips only -- read straight from global memory every record
//                  paid three dependent L2 round trips (meta, task words, operand words; ~1 800 cycles, 197 records);
//   otherwise      direct global reads (host emulation).
#if PBM_FAC_SMEM
#define PB_FAC_LD(x) (x)
#define PB_FAC_STAGE 0
#elif !defined(PB_EMU) && defined(PBM_FAC_NCHUNK) && !defined(PB_NO_FAC_STAGE)
#define PB_FAC_LD(x) (x)
#define PB_FAC_STAGE 1
#else
#define PB_FAC_LD(x) PB_LDU(x)
#define PB_FAC_STAGE 0
#endif

// The bidiagonal chains, one lane per chain, before every level (symbolic_lu._build): words r0 | K << 16 and L0 | U0 << 16;
// the diagonal of row i is slot i, the L and U entries of a chain sit in consecutive slots.  Out of line and without
// register arrays on purpose: it runs once per factorisation, and a larger frame of the solve kernel costs everywhere (a
// first version with batched loads doubled the kernel's stack and slowed the whole DFN integration by 11 %).
#ifdef PB_EMU
static void pb_lu_factor_chains(double* lu, const unsigned int* body, const int nt, double& bad) {
#else
static __device__ __noinline__ void pb_lu_factor_chains(double* PB_RESTRICT lu, const unsigned int* PB_RESTRICT body, const int nt,
                                                        double& bad) {
#endif
  for (int q = PB_TID; q < nt; q += PB_TEAM) {
    const unsigned int h = PB_FAC_LD(body[q]), sl = PB_FAC_LD(body[nt + q]);
    const int r0 = (int)(h & 0xFFFFu), K = (int)(h >> 16);
    double* l = lu + (sl & 0xFFFFu) - 1;       // l[k] = L(r0 + k, r0 + k - 1)
    const double* u = lu + (sl >> 16) - 1;     // u[k] = U(r0 + k - 1, r0 + k)
    double* d = lu + r0;
    double inv = 1.0 / d[0];
    if (!(fabs(inv) <= 1.7e308) || d[0] == 0.0) bad = 1.0;
    d[0] = inv;
    for (int k = 1; k < K; k++) {
      const double lk = l[k] * inv;
      const double acc = d[k] - lk * u[k];
      inv = 1.0 / acc;
      if (!(fabs(inv) <= 1.7e308) || acc == 0.0) bad = 1.0;
      l[k] = lk;
      d[k] = inv;
    }
  }
}

// One record of the factor stream at st[off] (meta already fetched). Returns the offset of the next record.
static PB_DEV int pb_lu_factor_record(double* PB_RESTRICT lu, const unsigned int* PB_RESTRICT st, const int off,
                                      const unsigned int meta, double& bad) {
  const int nt = (int)(meta & 0xFFFFu), ml = (int)((meta >> 16) & 0xFFFu);
  const unsigned int fl = meta >> 28;
  const int base = off + 1;
  const int wide = (fl & PB_F_TEAM) != 0;
  if (fl & PB_F_CHAINS) {
    pb_lu_factor_chains(lu, st + base, nt, bad);
    PB_SYNC();
    return base + nt * (1 + ml);
  }
  if (fl & PB_F_GROUP) {   // narrow level, few long tasks: warp 0, 32/nt lanes per task
    if (PB_WARP == 0) {
#ifndef PB_EMU
      const unsigned int tw = PB_FAC_LD(st[base + (PB_LANE & (nt - 1))]);
      const unsigned int* op = st + base + nt + PB_LANE;
      const int k_n = (ml * nt) >> 5;
      double a0 = 0.0, a1 = 0.0;
      for (int k = 0; k < k_n; k += 8) {   // operand words in batches of eight
        unsigned int o[8];
#pragma unroll
        for (int j = 0; j < 8; j++) o[j] = (k + j < k_n) ? PB_FAC_LD(op[32 * (k + j)]) : PB_FAC_PAD;
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
          a0 += lu[o[j] & 0xFFFFu] * lu[o[j] >> 16];
          a1 += lu[o[j + 1] & 0xFFFFu] * lu[o[j + 1] >> 16];
        }
      }
      a0 += a1;
      for (int o = nt; o < 32; o <<= 1) a0 += __shfl_xor_sync(0xffffffffu, a0, o);
      if (PB_LANE < nt) {
        const int out = (int)(tw & 0xFFFFu), aux = (int)(tw >> 16);
        const double acc = (aux == PB_AUX_NEG ? 0.0 : lu[out]) - a0;
        if (aux == PB_AUX_U || aux == PB_AUX_NEG) {
          lu[out] = acc;
        } else if (aux == PB_AUX_DIAG) {
          const double inv = PB_VDIV(1.0, acc);
          if (!(fabs(inv) <= 1.7e308) || acc == 0.0) bad = 1.0;
          lu[out] = inv;
        } else {
          lu[out] = acc * lu[aux];
        }
      }
#else
      for (int q = 0; q < nt; q++) {
        const unsigned int tw = st[base + q];
        const int out = (int)(tw & 0xFFFFu), aux = (int)(tw >> 16);
        double dot = 0.0;
        for (int p = 0; p < ml; p++) {
          const unsigned int o0 = st[base + nt + p * nt + q];
          dot += lu[o0 & 0xFFFFu] * lu[o0 >> 16];
        }
        const double acc = (aux == PB_AUX_NEG ? 0.0 : lu[out]) - dot;
        if (aux == PB_AUX_U || aux == PB_AUX_NEG) {
          lu[out] = acc;
        } else if (aux == PB_AUX_DIAG) {
          const double inv = PB_VDIV(1.0, acc);
          if (!(fabs(inv) <= 1.7e308) || acc == 0.0) bad = 1.0;
          lu[out] = inv;
        } else {
          lu[out] = acc * lu[aux];
        }
      }
#endif
    }
    if (fl & PB_F_END) PB_SYNCWARP();
    return base + nt * (1 + ml);
  }
  if (wide || PB_WARP == 0) {
    const int stride = wide ? PB_TEAM : PB_LANES;
    for (int q = wide ? PB_TID : PB_LANE; q < nt; q += stride) {
      const unsigned int tw = PB_FAC_LD(st[base + q]);
      const int out = (int)(tw & 0xFFFFu), aux = (int)(tw >> 16);
      const unsigned int* op = st + base + nt + q;
      double acc0 = (aux == PB_AUX_NEG) ? 0.0 : lu[out], acc1 = 0.0;
#ifndef PB_EMU
      for (int p = 0; p < ml; p += 8) {   // operand words in batches of eight
        unsigned int o[8];
#pragma unroll
        for (int j = 0; j < 8; j++) o[j] = (p + j < ml) ? PB_FAC_LD(op[(p + j) * nt]) : PB_FAC_PAD;
#pragma unroll
        for (int j = 0; j < 8; j += 2) {
          acc0 -= lu[o[j] & 0xFFFFu] * lu[o[j] >> 16];
          acc1 -= lu[o[j + 1] & 0xFFFFu] * lu[o[j + 1] >> 16];
        }
      }
#else
      for (int p = 0; p < ml; p++) {
        const unsigned int o0 = op[p * nt];
        if (p & 1) acc1 -= lu[o0 & 0xFFFFu] * lu[o0 >> 16];
        else acc0 -= lu[o0 & 0xFFFFu] * lu[o0 >> 16];
      }
#endif
      const double acc = acc0 + acc1;
      if (aux == PB_AUX_U || aux == PB_AUX_NEG) {
        lu[out] = acc;
      } else if (aux == PB_AUX_DIAG) {
        const double inv = PB_VDIV(1.0, acc);
        if (!(fabs(inv) <= 1.7e308) || acc == 0.0) bad = 1.0;
        lu[out] = inv;
      } else {
        lu[out] = acc * lu[aux];
      }
    }
  }
  if (fl & PB_F_END) {
    if (wide) PB_SYNC();
    else PB_SYNCWARP();
  }
  return base + nt * (1 + ml);
}

// Entry-parallel left-looking factorisation of the values already scattered into lu[] by pbm_jac.
// Returns non-zero (team-uniform) when a pivot is zero / not finite.
static PB_DEV int pb_lu_factor(pb_sys& S) {
  double* lu = S_LU(S);
  double bad = 0.0;
#if PB_FAC_STAGE
  {
    constexpr int WPT = (PBM_FAC_CHUNK_WORDS + PB_TEAM - 1) / PB_TEAM;
    static_assert(PBM_FAC_CHUNK_WORDS <= 2 * PBM_N, "factor chunk larger than the w scratch (w[N] is the solve's zero slot)");
    unsigned int* buf = (unsigned int*)S_W(S);
    unsigned int pre[WPT];
    int c_off = (int)pbm_fac_chunk[0], c_len = (int)pbm_fac_chunk[2] - c_off;
#pragma unroll
    for (int k = 0; k < WPT; k++) {
      const int i = PB_TID + k * PB_TEAM;
      pre[k] = (i < c_len) ? PB_LDU(pbm_fac_tab[c_off + i]) : 0u;
    }
    for (int c = 0; c < PBM_FAC_NCHUNK; c++) {
      const int nrec = (int)pbm_fac_chunk[2 * c + 1];
      PB_SYNC();   // everybody is done with the previous chunk (and with w before the first one)
#pragma unroll
      for (int k = 0; k < WPT; k++) {
        const int i = PB_TID + k * PB_TEAM;
        if (i < c_len) buf[i] = pre[k];
      }
      PB_SYNC();
      if (c + 1 < PBM_FAC_NCHUNK) {   // next chunk: L2 -> registers while this one is processed
        c_off = (int)pbm_fac_chunk[2 * c + 2];
        c_len = (int)pbm_fac_chunk[2 * c + 4] - c_off;
#pragma unroll
        for (int k = 0; k < WPT; k++) {
          const int i = PB_TID + k * PB_TEAM;
          pre[k] = (i < c_len) ? PB_LDU(pbm_fac_tab[c_off + i]) : 0u;
        }
      }
      int off = 0;
      unsigned int meta = buf[0];
      for (int rec = 0; rec < nrec; rec++) {
        const int nxt = pb_lu_factor_record(lu, buf, off, meta, bad);
        off = nxt;
        if (rec + 1 < nrec) meta = buf[off];
      }
    }
  }
#else
  {
    const unsigned int* st = S.fac;
    int off = 0;
    unsigned int meta = PB_FAC_LD(st[0]);
    for (int rec = 0; rec < PBM_FAC_NREC; rec++) {
      off = pb_lu_factor_record(lu, st, off, meta, bad);
      if (rec + 1 < PBM_FAC_NREC) meta = PB_FAC_LD(st[off]);
    }
  }
#endif
  PB_SYNC();
#if PBM_DENSE_N > 0
  {
    // Dense top block: thread j forms column j of W = U_TT^-1 L_TT^-1 in the solve's w scratch (column-major, odd
    // leading dimension: conflict-free), reading L_TT / U_TT through the slot table (zero slot where the pattern has
    // no entry), and writes it row-major to the team's global scratch.
    constexpr int NT = PBM_DENSE_N, LD = NT | 1;
    static_assert(NT * LD <= PBM_N + 1, "dense block scratch");
    double* cw = S_W(S);
    double* wd = S_WD(S);
    for (int j = PB_TID; j < NT; j += PB_TEAM) {
      double* c = cw + j * LD;
      for (int i = 0; i < NT; i++) {
        double sa = (i == j) ? 1.0 : 0.0, sb = 0.0;
        int k = 0;
        for (; k + 1 < i; k += 2) {
          sa = fma(-lu[(int)PB_LD(pbm_dense_slot[i * NT + k])], c[k], sa);
          sb = fma(-lu[(int)PB_LD(pbm_dense_slot[i * NT + k + 1])], c[k + 1], sb);
        }
        if (k < i) sa = fma(-lu[(int)PB_LD(pbm_dense_slot[i * NT + k])], c[k], sa);
        c[i] = sa + sb;
      }
      for (int i = NT - 1; i >= 0; i--) {
        double sa = c[i], sb = 0.0;
        int k = i + 1;
        for (; k + 1 < NT; k += 2) {
          sa = fma(-lu[(int)PB_LD(pbm_dense_slot[i * NT + k])], c[k], sa);
          sb = fma(-lu[(int)PB_LD(pbm_dense_slot[i * NT + k + 1])], c[k + 1], sb);
        }
        if (k < NT) sa = fma(-lu[(int)PB_LD(pbm_dense_slot[i * NT + k])], c[k], sa);
        c[i] = (sa + sb) * lu[(int)PB_LD(pbm_dense_slot[i * NT + i])];
      }
      for (int i = 0; i < NT; i++) PB_STCG(wd + i * NT + j, c[i]);
    }
    PB_SYNC();
  }
#endif
  return pb_team_max(S.team, bad) != 0.0;
}

// ---- SEQ triangular solves (stream layout: pybamm_b200/symbolic_lu.py, build_seq_solve_stream) ------------------
// One thread walks one cluster of the elimination tree (a particle chain, a separator) row by row; the only
// loop-carried dependency is the w[] value it has just stored.  Everything that does not depend on it -- the next
// task's header, its first operand word and that operand's L/U value -- is fetched one task ahead, so a chain row
// costs one LDS + one DFMA (+ one DMUL backward) instead of a dependency level of the interpreter.
#define PB_SEQ_HDR 4
#define PB_PAR_ADD 2u
#define PB_PAR_DIAG 4u
#ifndef PB_PAR_CH
#define PB_PAR_CH 6   /* operands fetched together per lane */
#endif
#define PB_PAR_PAD ((unsigned int)PBM_NNZ_LU | ((unsigned int)PBM_N << 16))   /* zero L/U slot times w[N] = 0 */

template <int BWD>
static PB_DEV void pb_seq_thread(const unsigned int* PB_RESTRICT p, const int nt, const double* PB_RESTRICT lu,
                                 double* PB_RESTRICT w) {
  const int ntask = (int)p[0];
  p += nt;
  unsigned int hdr = p[0], o0 = p[nt];
  double l0 = lu[o0 & 0xFFFFu];
  double ld = BWD ? lu[hdr & 0xFFFFu] : 1.0;
  for (int k = 0; k < ntask; k++) {
    const int row = (int)(hdr & 0xFFFFu), ml = (int)(hdr >> 16);
    const unsigned int* pn = p + (1 + ml) * nt;
    const double wv0 = w[o0 >> 16], wr = w[row];       // the chain-critical loads first
    const unsigned int hdr_n = pn[0], o0n = pn[nt];    // prefetch of the next task (independent of the chain)
    double acc = wr - l0 * wv0;
    if (ml > 1) {
      // operands 1..ml-1 in chunks of four: the words of chunk c+1 are in flight while chunk c multiplies
      const unsigned int* q = p + 2 * nt;
      unsigned int a0 = q[0], a1 = q[nt], a2 = q[2 * nt], a3 = q[3 * nt];
      double acc1 = 0.0;
      for (int j = 1; j < ml; j += 4) {
        q += 4 * nt;
        const double x0 = lu[a0 & 0xFFFFu], y0 = w[a0 >> 16], x1 = lu[a1 & 0xFFFFu], y1 = w[a1 >> 16];
        const double x2 = lu[a2 & 0xFFFFu], y2 = w[a2 >> 16], x3 = lu[a3 & 0xFFFFu], y3 = w[a3 >> 16];
        a0 = q[0];
        a1 = q[nt];
        a2 = q[2 * nt];
        a3 = q[3 * nt];
        acc -= x0 * y0;
        acc1 -= x1 * y1;
        acc -= x2 * y2;
        acc1 -= x3 * y3;
      }
      acc += acc1;
    }
    const double l0n = lu[o0n & 0xFFFFu];
    double ldn = 1.0;
    if (BWD) {
      ldn = lu[hdr_n & 0xFFFFu];
      acc *= ld;
    }
    w[row] = acc;
    hdr = hdr_n;
    o0 = o0n;
    l0 = l0n;
    ld = ldn;
    p = pn;
  }
}

// ---- specialised records (device): shapes are template parameters, calls come from PBM_SOLVE_PROGRAM ------------
#ifndef PB_EMU
// PAR record: NT lanes, G lanes per row, MLG operands per lane.  Partial dot product, xor-butterfly over the G lanes,
// w[row] = base +/- sum with base = w[row] or w[row] / u_ii.  Every lane reads before any lane writes (rows of one
// cluster read each other's old values through the explicit inverse block).
// Out of line and shared by every record of the same shape (G, MLG, MODE): the solve is one call per record.
template <int G, int MLG, unsigned int MODE>
static __device__ __noinline__ void pb_par_rec(const unsigned int* PB_RESTRICT blk, const int nt, const double* PB_RESTRICT lu,
                                               double* PB_RESTRICT w) {
  {   // whole warp, constant mask: see pb_par_generic
    const unsigned int* p = blk + (PB_LANE < nt ? PB_LANE : nt - 1);
    const int row = (int)p[0];
    unsigned int a[MLG > 0 ? MLG : 1];
#pragma unroll
    for (int j = 0; j < MLG; j++) a[j] = p[(1 + j) * nt];
    const double wr = w[row];
    const double ld = (MODE & PB_PAR_DIAG) ? lu[row] : 1.0;
    double x[MLG > 0 ? MLG : 1], y[MLG > 0 ? MLG : 1];
#pragma unroll
    for (int j = 0; j < MLG; j++) {
      x[j] = lu[a[j] & 0xFFFFu];
      y[j] = w[a[j] >> 16];
    }
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int j = 0; j < MLG; j += 2) {
      s0 = fma(x[j], y[j], s0);
      if (j + 1 < MLG) s1 = fma(x[j + 1], y[j + 1], s1);
    }
    double sum = MLG > 1 ? s0 + s1 : s0;
#pragma unroll
    for (int o = 1; o < G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const double base = (MODE & PB_PAR_DIAG) ? wr * ld : wr;
    const double acc = (MODE & PB_PAR_ADD) ? base + sum : base - sum;
    __syncwarp();
    if ((G == 1 || (PB_LANE & (G - 1)) == 0) && PB_LANE < nt) w[row] = acc;
  }
}

// The same record with run-time shape: ONE copy of the code for every PAR record (about 1.5 KB of instructions
// instead of one function per shape).  Operands in chunks of four, each chunk's loads issued together.
static __device__ __noinline__ void pb_par_generic(const unsigned int* PB_RESTRICT blk, const int nt, const int G, const int mlg,
                                                   const unsigned int mode, const double* PB_RESTRICT lu,
                                                   double* PB_RESTRICT w) {
  // The whole warp runs the record with the full constant mask (a run-time mask costs a MATCH/REDUX sequence per
  // shuffle): lanes past nt shadow lane nt - 1 -- whole groups, since nt is a multiple of G -- and never write.
  {
    const unsigned int* p = blk + (PB_LANE < nt ? PB_LANE : nt - 1);
    const int row = (int)p[0];
    const double wr = w[row];
    const double ld = (mode & PB_PAR_DIAG) ? lu[row] : 1.0;
    double s0 = 0.0, s1 = 0.0;
    p += nt;
    // operand words of the next chunk are fetched while the current chunk's values are in flight
    unsigned int a[PB_PAR_CH];
#pragma unroll
    for (int k = 0; k < PB_PAR_CH; k++) a[k] = (k < mlg) ? p[k * nt] : PB_PAR_PAD;
    for (int j = 0; j < mlg; j += PB_PAR_CH) {
      double x[PB_PAR_CH], y[PB_PAR_CH];
#pragma unroll
      for (int k = 0; k < PB_PAR_CH; k++) {
        x[k] = lu[a[k] & 0xFFFFu];
        y[k] = w[a[k] >> 16];
      }
      p += PB_PAR_CH * nt;
      if (j + PB_PAR_CH < mlg) {
#pragma unroll
        for (int k = 0; k < PB_PAR_CH; k++) a[k] = (j + PB_PAR_CH + k < mlg) ? p[k * nt] : PB_PAR_PAD;
      }
#pragma unroll
      for (int k = 0; k < PB_PAR_CH; k += 2) {
        s0 = fma(x[k], y[k], s0);
        if (k + 1 < PB_PAR_CH) s1 = fma(x[k + 1], y[k + 1], s1);
      }
    }
    double sum = s0 + s1;
    for (int o = 1; o < G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const double base = wr * ld;
    const double acc = (mode & PB_PAR_ADD) ? base + sum : base - sum;
    __syncwarp();
    if ((PB_LANE & (G - 1)) == 0 && PB_LANE < nt) w[row] = acc;
  }
}

#if defined(PBM_PAR_RUN_N) && !defined(PB_EMU)
// A run of consecutive single-warp PAR records in one loop, software-pipelined over the records: everything that
// does not depend on w (descriptor, rows, operand words, L/U values) is fetched two / one records ahead, so the
// critical path of a record is w loads -> DFMAs -> butterfly -> store.  Descriptors: pbm_par_run[2 r] =
// body offset | nt << 16, pbm_par_run[2 r + 1] = G | mlg << 8 | mode << 16 (constant memory, emitted by codegen).
#define PB_RUN_ROWS(D0, D1, P, ROW, A)                                              \
  {                                                                                 \
    const int nt_ = (int)((D0) >> 16), mlg_ = (int)(((D1) >> 8) & 0xFFu);           \
    P = st + ((D0) & 0xFFFFu) + (PB_LANE < nt_ ? PB_LANE : nt_ - 1);                \
    ROW = (int)P[0];                                                                \
    _Pragma("unroll") for (int k = 0; k < CH; k++) A[k] = (k < mlg_) ? P[(1 + k) * nt_] : PB_PAR_PAD; \
  }
static __device__ __noinline__ void pb_par_run(const unsigned int* PB_RESTRICT st, const int r0, const int nrec,
                                               const double* PB_RESTRICT lu, double* PB_RESTRICT w) {
  constexpr int CH = PB_PAR_CH;
  const int rlast = r0 + nrec - 1;
  unsigned int c0 = pbm_par_run[2 * r0], c1 = pbm_par_run[2 * r0 + 1];
  int q = r0 + 1 < rlast ? r0 + 1 : rlast;
  unsigned int n0 = pbm_par_run[2 * q], n1 = pbm_par_run[2 * q + 1];
  q = r0 + 2 < rlast ? r0 + 2 : rlast;
  unsigned int m0 = pbm_par_run[2 * q], m1 = pbm_par_run[2 * q + 1];
  const unsigned int *pc, *pn;
  int rowc, rown;
  unsigned int ac[CH], an[CH];
  double xc[CH], ldc;
  PB_RUN_ROWS(c0, c1, pc, rowc, ac)
  PB_RUN_ROWS(n0, n1, pn, rown, an)
  ldc = lu[rowc];
#pragma unroll
  for (int k = 0; k < CH; k++) xc[k] = lu[ac[k] & 0xFFFFu];
  for (int r = 0; r < nrec; r++) {
    // --- fetched ahead, consumed in later iterations
    double xn[CH];
    const double ldn = lu[rown];
#pragma unroll
    for (int k = 0; k < CH; k++) xn[k] = lu[an[k] & 0xFFFFu];
    const unsigned int* pm;
    int rowm;
    unsigned int am[CH];
    PB_RUN_ROWS(m0, m1, pm, rowm, am)
    q = r0 + r + 3 < rlast ? r0 + r + 3 : rlast;
    const unsigned int l0 = pbm_par_run[2 * q], l1 = pbm_par_run[2 * q + 1];
    // --- this record
    const int nt = (int)(c0 >> 16), G = (int)(c1 & 0xFFu), mlg = (int)((c1 >> 8) & 0xFFu);
    const unsigned int mode = c1 >> 16;
    const double wr = w[rowc];
    double y[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) y[k] = w[ac[k] >> 16];
    double s0 = 0.0, s1 = 0.0;
#pragma unroll
    for (int k = 0; k < CH; k += 2) {
      s0 = fma(xc[k], y[k], s0);
      if (k + 1 < CH) s1 = fma(xc[k + 1], y[k + 1], s1);
    }
    for (int j = CH; j < mlg; j += CH) {   // the few long records: remaining operands, not pipelined
      unsigned int b[CH];
#pragma unroll
      for (int k = 0; k < CH; k++) b[k] = (j + k < mlg) ? pc[(1 + j + k) * nt] : PB_PAR_PAD;
#pragma unroll
      for (int k = 0; k < CH; k += 2) {
        s0 = fma(lu[b[k] & 0xFFFFu], w[b[k] >> 16], s0);
        if (k + 1 < CH) s1 = fma(lu[b[k + 1] & 0xFFFFu], w[b[k + 1] >> 16], s1);
      }
    }
    double sum = s0 + s1;
    for (int o = 1; o < G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const double base = (mode & PB_PAR_DIAG) ? wr * ldc : wr;
    const double acc = (mode & PB_PAR_ADD) ? base + sum : base - sum;
    __syncwarp();
    if ((PB_LANE & (G - 1)) == 0 && PB_LANE < nt) w[rowc] = acc;
    __syncwarp();
    // --- rotate the pipeline
    c0 = n0; c1 = n1; pc = pn; rowc = rown; ldc = ldn;
    n0 = m0; n1 = m1; pn = pm; rown = rowm;
    m0 = l0; m1 = l1;
#pragma unroll
    for (int k = 0; k < CH; k++) {
      ac[k] = an[k];
      xc[k] = xn[k];
      an[k] = am[k];
    }
  }
}
#endif

// CHAIN record: one lane per bidiagonal chain of K consecutive rows.  All operands are independent of the recurrence
// and are loaded in batches ahead of it: the chain costs one DFMA per row.
template <int BWD>
static __device__ __noinline__ void pb_chain_rec(const unsigned int* PB_RESTRICT blk, const int nt, const double* PB_RESTRICT lu,
                                                 double* PB_RESTRICT w) {
  if (PB_LANE >= nt) return;
  const unsigned int h = blk[PB_LANE], sl = blk[nt + PB_LANE];
  const int r0 = (int)(h & 0xFFFFu), K = (int)(h >> 16);
  constexpr int B = 6;
  if (!BWD) {
    const double* l = lu + (sl & 0xFFFFu) - 1;   // l[k] = L(r0 + k, r0 + k - 1)
    double* z = w + r0;
    double prev = z[0];
    for (int k = 1; k < K; k += B) {
      double wv[B], lv[B];
#pragma unroll
      for (int j = 0; j < B; j++)
        if (k + j < K) {
          wv[j] = z[k + j];
          lv[j] = l[k + j];
        }
#pragma unroll
      for (int j = 0; j < B; j++)
        if (k + j < K) {
          prev = fma(-lv[j], prev, wv[j]);
          z[k + j] = prev;
        }
    }
  } else {
    const double* u = lu + (sl >> 16);           // u[k] = U(r0 + k, next row)
    const double* d = lu + r0;                   // 1 / u_kk
    double* z = w + r0;
    double nxt = w[blk[2 * nt + PB_LANE]];
    for (int k = K - 1; k >= 0; k -= B) {
      double zd[B], ud[B];
#pragma unroll
      for (int j = 0; j < B; j++)
        if (k - j >= 0) {
          const double dd = d[k - j];
          zd[j] = z[k - j] * dd;
          ud[j] = u[k - j] * dd;
        }
#pragma unroll
      for (int j = 0; j < B; j++)
        if (k - j >= 0) {
          nxt = fma(-ud[j], nxt, zd[j]);
          z[k - j] = nxt;
        }
    }
  }
}

template <int BWD>
static __device__ __noinline__ void pb_seq_rec(const unsigned int* PB_RESTRICT blk, const int nt, const double* PB_RESTRICT lu,
                                               double* PB_RESTRICT w) {
  if (PB_LANE < nt) pb_seq_thread<BWD>(blk + PB_LANE, nt, lu, w);
}
#endif

#ifndef PB_EMU
#if PBM_DENSE_N > 0
// Dense top block x_T = W r_T: PBM_DENSE_G lanes per row, every lane a few columns.  The thread's share of W comes
// from L2 and does not depend on the right-hand side: it is fetched before the forward sweep starts.
#define PB_DENSE_CPL ((PBM_DENSE_N + PBM_DENSE_G - 1) / PBM_DENSE_G)
#define PB_DENSE_PROLOGUE(wd)                                                                            \
  double wdv[PB_DENSE_CPL];                                                                              \
  const int drow = PB_TID / PBM_DENSE_G, dpart = PB_TID % PBM_DENSE_G;                                   \
  _Pragma("unroll") for (int k = 0; k < PB_DENSE_CPL; k++) {                                             \
    const int col = dpart + k * PBM_DENSE_G;                                                             \
    wdv[k] = (drow < PBM_DENSE_N && col < PBM_DENSE_N) ? __ldcg((wd) + drow * PBM_DENSE_N + col) : 0.0;     \
  }
#define PB_DENSE_REC(off)                                                                                \
  {                                                                                                      \
    const unsigned int* rows = st + (off);                                                               \
    double sum = 0.0;                                                                                    \
    _Pragma("unroll") for (int k = 0; k < PB_DENSE_CPL; k++) {                                           \
      const int col = dpart + k * PBM_DENSE_G;                                                           \
      sum = fma(wdv[k], w[rows[col < PBM_DENSE_N ? col : 0]], sum);                                      \
    }                                                                                                    \
    _Pragma("unroll") for (int o = 1; o < PBM_DENSE_G; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o); \
    PB_SYNC();                                                                                           \
    if (dpart == 0 && drow < PBM_DENSE_N) w[rows[drow]] = sum;                                           \
    PB_SYNC();                                                                                           \
  }
static_assert(PBM_DENSE_N * PBM_DENSE_G <= PB_TEAM, "dense block wider than the team");
#else
#define PB_DENSE_PROLOGUE(wd)
#endif
// Models whose solve stream is read from the global table (PB_GLOBAL_VEC == 2): the whole team can touch the next record's
// lines while the current record runs.  Measured on the 3 523-state model and left off: 61.2k against 54.5k cycles per
// solve, 1 154 against 1 206 sims/s -- the stream words are not what the records wait for.
#if PB_GLOBAL_VEC == 2 && defined(PB_SLV_PREFETCH_ON)
#define PB_SLV_PREFETCH(off, n)                                                     \
  for (int i_ = PB_TID * 32; i_ < (n); i_ += PB_TEAM * 32)                          \
    asm volatile("prefetch.global.L1 [%0];" ::"l"(st + (off) + i_));
#else
#define PB_SLV_PREFETCH(off, n)
#endif
// one copy of the call sequence for every call site of pb_lu_solve
static __device__ __noinline__ void pb_solve_program(const unsigned int* PB_RESTRICT st, const double* PB_RESTRICT lu,
                                                     double* PB_RESTRICT w, const double* PB_RESTRICT wd) {
  PB_DENSE_PROLOGUE(wd)
#ifdef PBM_SOLVE_NREC
  // table-driven: one loop over the records, this warp's descriptor from constant memory (codegen.emit_solve_table)
  for (int r = 0; r < PBM_SOLVE_NREC; r++) {
    const unsigned int sy = pbm_solve_sync[r];
#if PBM_DENSE_N > 0
    if (sy == 3u) {
      PB_DENSE_REC(PBM_SOLVE_DENSE_OFF)
      continue;
    }
#endif
    if (PB_WARP < PBM_SOLVE_NW) {
      const unsigned int d0 = pbm_solve_desc[(r * PBM_SOLVE_NW + PB_WARP) * 2], d1 = pbm_solve_desc[(r * PBM_SOLVE_NW + PB_WARP) * 2 + 1];
      const unsigned int kind = d1 >> 24;
      const unsigned int* blk = st + (d0 & 0xFFFFu);
      const int nt = (int)(d0 >> 16);
      if (kind == 1u) pb_par_generic(blk, nt, (int)(d1 & 0xFFu), (int)((d1 >> 8) & 0xFFu), (d1 >> 16) & 0xFFu, lu, w);
      else if (kind == 2u) pb_chain_rec<0>(blk, nt, lu, w);
      else if (kind == 3u) pb_chain_rec<1>(blk, nt, lu, w);
    }
    if (sy == 2u) PB_SYNC();
    else if (sy == 1u) PB_SYNCWARP();
  }
#else
#define PB_REC_MARK
  PBM_SOLVE_PROGRAM
#undef PB_REC_MARK
#endif
}
#endif

// Solve (J - cj M) x = b in place (b: shared vector), w = S_W scratch. Ends with a team barrier.
// INW: the caller has already put the permuted right-hand side into w (w[pbm_iperm_r[i]] = b_i, barrier done) and
// reads the solution from it (x_j = w[pbm_iperm_c[j]]): no gather / scatter passes (the Newton loop, pb_nls_impl).
// TRACE (tools/gpu_cycles.py only): thread 0 stores the clock64() cycles of gather, every record, scatter in trace[].
template <int TRACE, int INW = 0>
static PB_DEV void pb_lu_solve_t(pb_sys& S, double* b, double* trace) {
  const double* lu = S_LU(S);
  long long tr0 = 0;
  int trk = 0;
  if (TRACE) tr0 = clock64();
  double* w = S_W(S);
#if defined(PB_PROFILE) && PB_PROFILE == 2
  PB_T0(tga_);
#endif
  if (!INW) {
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) w[i] = b[(int)PB_LD(pbm_perm_r[i])];
    PB_SYNC();
  }
#if defined(PB_PROFILE) && PB_PROFILE == 2
  PB_T1(S, tga_, 5);   // PB_PROFILE=2: slots 5/6/7 = gather / records / scatter of the triangular solves
  PB_T0(tlv_);
#endif
  if (TRACE) {
    const long long t1 = clock64();
    if (PB_TID == 0) trace[trk] = (double)(t1 - tr0);
    trk++;
    tr0 = t1;
  }
  {
#ifndef PB_EMU
    if (TRACE) {   // an inline copy with a clock sample after every record (eval kernel only)
      const unsigned int* st = S.slv;
      PB_DENSE_PROLOGUE(S_WD(S))
#define PB_REC_MARK                                              \
  {                                                              \
    const long long t1 = clock64();                              \
    if (PB_TID == 0 && trk < 60) trace[trk] = (double)(t1 - tr0); \
    trk++;                                                       \
    tr0 = t1;                                                    \
  }
      PBM_SOLVE_PROGRAM
#undef PB_REC_MARK
    } else {
      pb_solve_program(S.slv, lu, w, S_WD(S));
    }
#else
    const unsigned int* st = S.slv;
    int off = 0;
    while (off < PBM_SLV_WORDS) {   // host emulation: the generic interpreter of the same stream
      const unsigned int h0 = st[off];
      const int total = (int)st[off + 1];
      const unsigned int kind = h0 & 0xFu, mode = (h0 >> 4) & 0xFu;
      const int nb = (int)((h0 >> 16) & 0xFFu);
      const unsigned int* body = st + off + PB_SEQ_HDR + nb;
      if (kind == 0) {
        for (int v = 0; v < nb; v++) {
          const unsigned int d = st[off + PB_SEQ_HDR + v];
          const int nt = (int)(d & 0xFFFFu), len = (int)(d >> 16);
          for (int q = 0; q < nt; q++) {
            if (mode & 1u) pb_seq_thread<1>(body + q, nt, lu, w);
            else pb_seq_thread<0>(body + q, nt, lu, w);
          }
          body += nt * len;
        }
      } else if (kind == 3) {   // dense top block: x_T = W r_T
        const int nT = (int)(st[off + 3] >> 8);
        const double* wd = S_WD(S);
        double rt[64];
        for (int q = 0; q < nT; q++) rt[q] = w[body[q]];
        for (int i = 0; i < nT; i++) {
          double sum = 0.0;
          for (int q = 0; q < nT; q++) sum += wd[i * nT + q] * rt[q];
          w[body[i]] = sum;
        }
      } else if (kind == 2) {
        for (int v = 0; v < nb; v++) {
          const int nt = (int)st[off + PB_SEQ_HDR + v];
          for (int q = 0; q < nt; q++) {
            const int r0 = (int)(body[q] & 0xFFFFu), K = (int)(body[q] >> 16);
            const int L0 = (int)(body[nt + q] & 0xFFFFu), U0 = (int)(body[nt + q] >> 16);
            if (mode & 1u) {
              double nxt = w[body[2 * nt + q]];
              for (int k = K - 1; k >= 0; k--) {
                nxt = (w[r0 + k] - lu[U0 + k] * nxt) * lu[r0 + k];
                w[r0 + k] = nxt;
              }
            } else {
              for (int k = 1; k < K; k++) w[r0 + k] -= lu[L0 + k - 1] * w[r0 + k - 1];
            }
          }
          body += 3 * nt;
        }
      } else {
        const unsigned int prm = st[off + 3];
        const int G = (int)(prm & 0xFFu), mlg = (int)(prm >> 8);
        for (int v = 0; v < nb; v++) {   // a block = a warp: all its lanes read before any of them writes
          double accs[32];
          int rws[32], na = 0;
          const int nt = (int)st[off + PB_SEQ_HDR + v];
          for (int g = 0; g < nt / G; g++) {
            const int row = (int)body[g * G];
            double sum = 0.0;
            for (int l = g * G; l < (g + 1) * G; l++)
              for (int j = 0; j < mlg; j++) {
                const unsigned int a = body[(1 + j) * nt + l];
                sum += lu[a & 0xFFFFu] * w[a >> 16];
              }
            const double base = (mode & PB_PAR_DIAG) ? w[row] * lu[row] : w[row];
            rws[na] = row;
            accs[na++] = (mode & PB_PAR_ADD) ? base + sum : base - sum;
          }
          body += (1 + mlg) * nt;
          for (int k = 0; k < na; k++) w[rws[k]] = accs[k];
        }
      }
      off += total;
    }
#endif
  }
#if defined(PB_PROFILE) && PB_PROFILE == 2
  PB_T1(S, tlv_, 6);
  PB_T0(tsc_);
#endif
  if (!INW) {
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) b[(int)PB_LD(pbm_perm_c[i])] = w[i];
    PB_SYNC();
  }
#if defined(PB_PROFILE) && PB_PROFILE == 2
  PB_T1(S, tsc_, 7);
#endif
  if (TRACE && PB_TID == 0) trace[trk] = (double)(clock64() - tr0);
}
static PB_DEV void pb_lu_solve(pb_sys& S, double* b) { pb_lu_solve_t<0>(S, b, nullptr); }
static PB_DEV void pb_lu_solve_in_w(pb_sys& S) { pb_lu_solve_t<0, 1>(S, nullptr, nullptr); }

// Evaluate J - cj*M at (t, y) into the LU value array and factor it.  ic_mode: replace the
// differential rows by identity rows (the algebraic sub-block system of the IC Newton,
// AlgebraicICBuilder.inl:52-119, solved through the same static LU).
static PB_DEV int pb_lsetup(pb_sys& S, double t, const double* y, double cj, int ic_mode) {
  double* lu = S_LU(S);
  PB_T0(tj_);
  for (int i = PB_TID; i < PBM_NFILL; i += PB_TEAM) lu[(int)PB_LD(pbm_fill[i])] = 0.0;
  pbm_jac(y, S_P(S), t, cj, lu, S_TMPJ(S));
  PB_SYNC();
  PB_T1(S, tj_, 2);
  if (ic_mode) {
    for (int k = PB_TID; k < PBM_NNZ; k += PB_TEAM) {
      const int kind = (int)PB_LD(pbm_ent_kind[k]);
      if (kind) lu[(int)PB_LD(pbm_ent_slot[k])] = (kind == 2) ? 1.0 : 0.0;
    }
    PB_SYNC();
  }
  PB_STAT(S, PB_ST_NJE);
  PB_STAT(S, PB_ST_NSETUPS);
  PB_T0(tf_);
  const int ier = pb_lu_factor(S);
  PB_T1(S, tf_, 3);
  return ier;
}

static PB_DEV double pb_wrms(pb_sys& S, const double* PB_RESTRICT x) {
  const double* PB_RESTRICT ewt = S_EWT(S);
  double xv[PB_VPT], ev[PB_VPT];
  PB_VEC_FOR(k, i) {
    xv[k] = (i < PBM_N) ? x[i] : 0.0;
    ev[k] = (i < PBM_N) ? ewt[i] : 0.0;
  }
  double acc = 0.0;
  PB_VEC_FOR(k, i) {
    const double v = xv[k] * ev[k];
    acc += v * v;
  }
  return PB_RMS(pb_team_sum(S.team, acc));
}

// ewt_i = 1 / (rtol |y_i| + atol_i); returns non-zero if a weight is not positive. Ends with a barrier.
static PB_DEV int pb_ewt_set(pb_sys& S, const double* y) {
  double* PB_RESTRICT ewt = S_EWT(S);
  double bad = 0.0;
  double yv[PB_VPT], av[PB_VPT];
  PB_VEC_FOR(k, i) {
    yv[k] = (i < PBM_N) ? y[i] : 1.0;
    av[k] = (i < PBM_N) ? PB_LD(S.atol[i]) : 1.0;
  }
  PB_VEC_FOR(k, i) {
    const double w = S.rtol * fabs(yv[k]) + av[k];
    if (!(w > 0.0)) bad = 1.0;
    if (i < PBM_N) ewt[i] = PB_VDIV(1.0, w);
  }
  PB_SYNC();
  return pb_team_max(S.team, bad) != 0.0;
}

// ewt_set(y) and wrms(y) in one pass (IDASolve computes both from phi[0] before every step): returns the norm, NaN when
// a weight is not positive.  Ends with a barrier.
static PB_DEV double pb_ewt_set_norm(pb_sys& S, const double* y) {
  double* PB_RESTRICT ewt = S_EWT(S);
  double yv[PB_VPT], av[PB_VPT];
  PB_VEC_FOR(k, i) {
    yv[k] = (i < PBM_N) ? y[i] : 0.0;
    av[k] = (i < PBM_N) ? PB_LD(S.atol[i]) : 1.0;
  }
  double acc = 0.0;
  PB_VEC_FOR(k, i) {
    const double w = S.rtol * fabs(yv[k]) + av[k];
    if (!(w > 0.0)) acc = PB_NAN;
    const double e = PB_VDIV(1.0, w);
    if (i < PBM_N) ewt[i] = e;
    const double v = yv[k] * e;
    acc += v * v;
  }
  PB_SYNC();
  return PB_RMS(pb_team_sum(S.team, acc));
}

// ---- interpolation (IDAGetSolution / IDAGetDky k=0,1) ------------------------------------------
struct pb_interp {
  double c[PB_MXORDP1], d[PB_MXORDP1];
  int kord;
};
// One out-of-line copy: eleven FP64 divisions (twenty instructions each, inline) at a dozen call sites of
// pb_get_solution were 40 KB of the kernel; the integration is bound by instruction fetch.
#ifndef PB_EMU
static __device__ __noinline__
#else
static
#endif
void pb_interp_coeffs_impl(const double* psi, int kord, const double delt, pb_interp* I) {
  if (kord == 0) kord = 1;
  double c = 1.0, d = 0.0, gam = delt / psi[0];
  I->c[0] = c;
  I->d[0] = 0.0;
#pragma unroll 1
  for (int j = 1; j < PB_MXORDP1; j++) {
    if (j <= kord) {
      d = d * gam + c / psi[j - 1];
      c = c * gam;
      gam = (delt + psi[j - 1]) / psi[j < PB_MXORDP1 - 1 ? j : PB_MXORDP1 - 1];
      I->c[j] = c;
      I->d[j] = d;   // multiplies phi[j] in y'
    } else {
      I->c[j] = 0.0;
      I->d[j] = 0.0;
    }
  }
  I->kord = kord;
}
static PB_DEV void pb_interp_coeffs(const pb_sys& S, double t, pb_interp& I) {
  pb_interp_coeffs_impl(S_PSI(S), S.kused, t - S.tn, &I);
}
// y(t) -> dst (shared or global). No barrier.
static PB_DEV void pb_get_solution(const pb_sys& S, double t, double* dst) {
  pb_interp I;
  pb_interp_coeffs(S, t, I);
  const double* PB_RESTRICT phi = S_PHI(S, 0);
  double pv[PB_VPT][PB_MXORDP1];
  PB_VEC_FOR(k, i) {
#pragma unroll
    for (int j = 0; j < PB_MXORDP1; j++) pv[k][j] = (j <= I.kord && i < PBM_N) ? phi[j * PBM_N + i] : 0.0;
  }
  PB_VEC_FOR(k, i) {
    if (i < PBM_N) {
      double s = 0.0;
#pragma unroll
      for (int j = PB_MXORDP1 - 1; j >= 0; j--)
        if (j <= I.kord) s += I.c[j] * pv[k][j];
      dst[i] = s;
    }
  }
}
// y(t) -> ydst and y'(t) -> ypdst. No barrier.
static PB_DEV void pb_get_solution2(const pb_sys& S, double t, double* ydst, double* ypdst) {
  pb_interp I;
  pb_interp_coeffs(S, t, I);
  const double* phi = S_PHI(S, 0);
  for (int i = PB_TID; i < PBM_N; i += PB_TEAM) {
    double s = 0.0, sp = 0.0;
#pragma unroll
    for (int j = PB_MXORDP1 - 1; j >= 0; j--)
      if (j <= I.kord) {
        const double v = phi[j * PBM_N + i];
        s += I.c[j] * v;
        if (j >= 1) sp += I.d[j] * v;
      }
    ydst[i] = s;
    ypdst[i] = sp;
  }
}

// IDAReInit: phi[0] = y, phi[1] = y' must already be in place.
static PB_DEV void pb_reinit(pb_sys& S, double t0) {
  S.tn = t0;
  S.nst = 0;
  S.kk = 0;
  S.kused = 0;
  S.hused = 0.0;
  S.hh = 0.0;
  S.tstopset = 0;
  S.tretlast = t0;
  S.epsNewt = S.epcon;
  S.toldel = 1e-4 * S.epsNewt;
}

// IDASetCoeffs.  Thread 0 updates the coefficient arrays (one out-of-line copy: five divisions per order) and leaves
// cj and ck in the two spare slots of the coefficient block; everybody reads them after the barrier.
// The phi -> phi* scaling (phi[i] *= beta[i], i >= ns) is folded into the predictor pass of pb_nls.
#define S_CJCK(S) ((S).sm + PB_OFF_COEF + 30)
#ifndef PB_EMU
static __device__ __noinline__
#else
static
#endif
void pb_set_coeffs_t0(double* coef, const int kk, const int ns, const double hh) {
  double* psi = coef;
  double* alpha = coef + 6;
  double* beta = coef + 12;
  double* sigma = coef + 18;
  double* gamma = coef + 24;
  if (kk + 1 >= ns) {
    beta[0] = 1.0;
    alpha[0] = 1.0;
    gamma[0] = 0.0;
    sigma[0] = 1.0;
    double temp1 = hh;
#pragma unroll 1
    for (int i = 1; i <= kk; i++) {
      const double temp2 = psi[i - 1];
      psi[i - 1] = temp1;
      beta[i] = PB_IDIV(beta[i - 1] * psi[i - 1], temp2);
      temp1 = temp2 + hh;
      alpha[i] = PB_IDIV(hh, temp1);
      sigma[i] = i * sigma[i - 1] * alpha[i];
      gamma[i] = gamma[i - 1] + PB_IDIV(alpha[i - 1], hh);
    }
    psi[kk] = temp1;
  }
  double alphas = 0.0, alpha0 = 0.0;
#pragma unroll 1
  for (int i = 0; i < kk; i++) {
    alphas -= PB_IDIV(1.0, (double)(i + 1));
    alpha0 -= alpha[i];
  }
  double c = fabs(alpha[kk] + alphas - alpha0);
  if (alpha[kk] > c) c = alpha[kk];
  coef[30] = PB_IDIV(-alphas, hh);
  coef[31] = c;
}
static PB_DEV void pb_set_coeffs(pb_sys& S, double* ck) {
  const int kk = S.kk;
  if (S.hh != S.hused || kk != S.kused) S.ns = 0;
  S.ns = (S.ns + 1 < S.kused + 2) ? S.ns + 1 : S.kused + 2;
  if (PB_TID == 0) pb_set_coeffs_t0(S_PSI(S), kk, S.ns, S.hh);
  PB_SYNC();
  S.cjlast = S.cj;
  S.cj = S_CJCK(S)[0];
  *ck = S_CJCK(S)[1];
  S.tn += S.hh;
}

// IDARestore. phi_scaled: whether the predictor pass (and so the phi* scaling) ran. Ends with a barrier.
static PB_DEV void pb_restore(pb_sys& S, double saved_t, int phi_scaled) {
  S.tn = saved_t;
  double* psi = S_PSI(S);
  const double* beta = S_BETA(S);
  PB_SYNC();
  if (PB_TID == 0)
    for (int j = 1; j <= S.kk; j++) psi[j - 1] = psi[j] - S.hh;
  if (phi_scaled)
    for (int j = S.ns; j <= S.kk; j++) {
      const double r = 1.0 / beta[j];
      double* ph = S_PHI(S, j);
      for (int i = PB_TID; i < PBM_N; i += PB_TEAM) ph[i] *= r;
    }
  PB_SYNC();
}

#define PB_NLS_SUCCESS 0
#define PB_NLS_CONV_RECVR 1
#define PB_NLS_LSETUP_RECVR 2
#define PB_ERROR_TEST_FAIL 7
#define PB_PREDICT_AGAIN 20

// IDANls + Newton iteration + idaNlsConvTest (modified Newton, J held between lsetups).
// ---- keeping the two teams of an SM in phase -----------------------------------------------------------------
// Two teams share an SM and its instruction cache.  Measured on the DFN: 296 copies of one system (teams in phase by
// construction) run 11 % slower than one team alone, 296 different systems of the same length 62 % slower -- a team
// that executes other code than its neighbour evicts the neighbour's instructions (icc hit rate 67 %).  So the teams
// meet before every Newton residual: each publishes an arrival count in a per-SM mailbox (global memory, L2) and
// waits, for a bounded time, until the neighbour has arrived at its own next meeting point.  The wait is skipped while
// the neighbour is absent, between systems, or inside a Jacobian set-up (state != 1).
#ifndef PB_ALIGN_TEAMS
#define PB_ALIGN_TEAMS 0
#endif
#ifndef PB_ALIGN_LIMIT
#define PB_ALIGN_LIMIT 8000   /* cycles */
#endif
#if PB_ALIGN_TEAMS && !defined(PB_EMU)
static PB_DEV void pb_align_state(pb_sys& S, unsigned int st) {
  if (S.al_me != nullptr && PB_TID == 0) S.al_me[1] = st;
}
static PB_DEV void pb_align(pb_sys& S) {
  if (S.al_me == nullptr) return;
  if (PB_TID == 0) {
    S.al_me[0] = ++S.al_count;
    const long long t0 = clock64();
    while (S.al_other[1] == 1u && S.al_other[0] == S.al_seen && clock64() - t0 < PB_ALIGN_LIMIT) {
    }
    S.al_seen = S.al_other[0];
  }
  PB_SYNC();
}
#else
#define pb_align_state(S, st) ((void)0)
#define pb_align(S) ((void)0)
#endif

// x^y of the step-size and convergence-rate formulas (x > 0): through the out-of-line exp / log the generated functions
// already keep resident, instead of another 150 instructions of libdevice pow() on the path of every step (the kernels are
// bound by instruction fetch).  SPM x 1 000: 135.4k -> 147.7k, x 10 000: 170k -> 185.7k, DFN 5 297 -> 5 385 sims/s.  The last
// bits of the step sizes differ from pow()'s; -DPB_POW_LIBM restores it.
#if !defined(PB_POW_LIBM) && !defined(PB_EMU)
#define PB_IPOW(x, y) PB_EXP((y) * PB_LOG(x))
#else
#define PB_IPOW(x, y) pow(x, y)
#endif

// ---- two teams of one CTA in lock-step through the large functions (PB_PAIR) -------------------------------------------
// The integration is bound by instruction fetch: the generated residual alone is 45 KB of code against a 32 KB instruction
// cache, and a team runs it 1.7x slower beside a neighbour that executes other code.  Two separate CTAs cannot be held
// close enough (tried in round 1: a mailbox in L2, iteration granularity, +-1 %).  Inside one CTA they can: a team about
// to enter function f posts a ticket; if the other team arrives for the same f within `limit` cycles it claims the
// ticket and both run f on the CTA-wide barrier, stage by stage side by side; otherwise the first team withdraws and
// runs alone on its own barrier.  Both functions execute a fixed sequence of barriers, whatever the data.
// MEASURED, and therefore off (build with -DPB_PAIR=1 to repeat; tools/gpu_pair_debug.py prints the join statistics): DFN,
// 2 960 systems, 4 669 sims/s without against 4 119 with.  With a 6 000-cycle wait 168 of a system's 1 030 residuals run
// jointly (386 with 20 000 cycles, 53 with 300), and a joint residual takes 23.1k cycles against 26.1k alone: stage-level
// lock-step of two DIFFERENT systems buys 10 %, not the 1.7x of two identical ones, and the waiting costs more than that.
#define PB_FN_RES 1u
#define PB_FN_SOLVE 2u
#define PB_PAIR_JOINED 0x100u
#ifndef PB_PAIR_LIMIT_RES
#define PB_PAIR_LIMIT_RES 6000
#endif
#ifndef PB_PAIR_LIMIT_SOLVE
#define PB_PAIR_LIMIT_SOLVE 1500
#endif
#if PB_PAIR && !defined(PB_EMU)
static PB_DEV int pb_pair_enter(const unsigned int f, const long long limit) {
  volatile unsigned int* ctl = PB_PAIR_CTL;
  if (blockDim.x <= PB_TEAM_SIZE) return 0;        // kernels launched with one team per CTA
  if (PB_TID == 0) {
    unsigned int joined = 0u;
    const unsigned int old = atomicCAS((unsigned int*)&ctl[0], 0u, f);
    if (old == 0u) {                               // first to arrive: wait for the other team, for a bounded time
      const long long t0 = clock64();
      for (;;) {
        if (ctl[0] == (f | PB_PAIR_JOINED)) {
          joined = 1u;
          break;
        }
        if (clock64() - t0 > limit) {
          if (atomicCAS((unsigned int*)&ctl[0], f, 0u) != f) joined = 1u;   // claimed while giving up
          break;
        }
      }
      if (joined) ctl[0] = 0u;                     // the claimer is held at the first joint barrier until we arrive
    } else if (old == f) {                         // the other team is waiting for the same function
      if (atomicCAS((unsigned int*)&ctl[0], f, f | PB_PAIR_JOINED) == f) joined = 1u;
    }
    ctl[2 + PB_TEAM_ID] = joined;
  }
  pb_bar_solo();
  return (int)ctl[2 + PB_TEAM_ID];
}
static PB_DEV void pb_pair_leave(const int joined) {
  if (joined) {      // every thread of the team is past the last joint barrier
    if (PB_TID == 0) PB_PAIR_CTL[2 + PB_TEAM_ID] = 0u;
    pb_bar_solo();
  }
}
#else
#define pb_pair_enter(f, limit) 0
#define pb_pair_leave(j) ((void)0)
#endif

static PB_DEV int pb_nls_impl(pb_sys& S) {
  int callLSetup = 0;
  if (S.nst == 0) {
    S.cjold = S.cj;
    S.ss = 20.0;
    callLSetup = 1;
  }
  S.cjratio = PB_IDIV(S.cj, S.cjold);
  {
    const double temp1 = (1.0 - 0.25) / (1.0 + 0.25), temp2 = 1.0 / temp1;
    if (S.cjratio < temp1 || S.cjratio > temp2) callLSetup = 1;
    if (S.cj != S.cjlast) S.ss = 100.0;
  }
  double* ee = S_EE(S);
  double* delta = S_DELTA(S);
  double* yypred = S_YYPRED(S);
  double* yy = S_YY(S);
  const double* ewt = S_EWT(S);
  // What a thread needs of the streamed vectors inside the Newton loop stays in its registers: the predicted
  // derivative of its rows and their places in the solve's permuted vector.  An iteration then touches global memory
  // only inside the generated residual (tables): every L2 round trip here used to sit on the critical path.
  double ypp[PB_VPT];
  int ipr[PB_VPT], ipc[PB_VPT];
  PB_VEC_FOR(k, i) {
    ipr[k] = (i < PBM_N) ? (int)PB_LD(pbm_iperm_r[i]) : 0;
    ipc[k] = (i < PBM_N) ? (int)PB_LD(pbm_iperm_c[i]) : 0;
  }
  {  // phi -> phi* and predictor
    const double* beta = S_BETA(S);
    const double* gamma = S_GAMMA(S);
    double* phi = S_PHI(S, 0);
    const int kk = S.kk, ns = S.ns;
    double bj[PB_MXORDP1], gj[PB_MXORDP1];
#pragma unroll
    for (int j = 0; j < PB_MXORDP1; j++) {
      bj[j] = (j <= kk) ? beta[j] : 0.0;
      gj[j] = (j <= kk) ? gamma[j] : 0.0;
    }
    // orders above kk contribute nothing: their coefficients are zero and the loads are skipped (team-uniform branch)
    double pv[PB_VPT][PB_MXORDP1];
    PB_VEC_FOR(k, i) {
#pragma unroll
      for (int j = 0; j < PB_MXORDP1; j++) pv[k][j] = (j <= kk && i < PBM_N) ? phi[j * PBM_N + i] : 0.0;
    }
    PB_VEC_FOR(k, i) {
      if (i < PBM_N) {
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int j = PB_MXORDP1 - 1; j >= 0; j--) {
          if (j <= kk) {
            double v = pv[k][j];
            if (j >= ns) {
              v *= bj[j];
              phi[j * PBM_N + i] = v;
            }
            a += v;
            if (j >= 1) b += gj[j] * v;
          }
        }
        yypred[i] = a;
        ee[i] = 0.0;
        yy[i] = a;
        ypp[k] = b;
      } else {
        ypp[k] = 0.0;
      }
    }
    PB_SYNC();
  }
  int jcur = 0, retval = 0;
  double oldnrm = 0.0;
  for (;;) {
    {
      pb_align(S);
      PB_T0(tr_);
      const int jn_ = pb_pair_enter(PB_FN_RES, PB_PAIR_LIMIT_RES);
#ifdef PB_PAIR_DEBUG
      const long long dbg0_ = clock64();
#endif
      pbm_res(yy, S_P(S), S.tn, 0.0, delta, S_TMP(S));
      PB_STAT(S, PB_ST_NRE);
      PB_SYNC();
#ifdef PB_PAIR_DEBUG
      if (PB_TID == 0) {   // debug build: stats slots 4..7 = cycles of solo / joint residuals and their counts
        S_STATS(S)[jn_ ? 5 : 4] += (double)(clock64() - dbg0_);
        S_STATS(S)[jn_ ? 7 : 6] += 1.0;
      }
#endif
      pb_pair_leave(jn_);
      PB_T1(S, tr_, 0);
    }
    int fail = 0;
    if (callLSetup) {
      pb_align_state(S, 2u);
      const int ier = pb_lsetup(S, S.tn, yy, S.cj, 0);
      pb_align_state(S, 1u);
      S.cjold = S.cj;
      S.cjratio = 1.0;
      S.ss = 20.0;
      jcur = 1;
      if (ier != 0) {
        retval = PB_NLS_LSETUP_RECVR;
        fail = 1;
      }
    }
    if (!fail) {
      int mnewt = 0;
      for (;;) {
        PB_STAT(S, PB_ST_NNI);
        // rhs = -F = -(f - M y'),  y' = yppred + cj * ee on the differential rows
        {
          double* PB_RESTRICT w = S_W(S);
          PB_VEC_FOR(k, i) {
            if (i < PBM_N) {
              double r = delta[i];
              if (i < PBM_NDIFF) r -= ypp[k] + S.cj * ee[i];
              w[ipr[k]] = -r;     // straight into the solve's permuted vector
            }
          }
        }
        PB_SYNC();
        {
          PB_T0(ts_);
          const int jn_ = pb_pair_enter(PB_FN_SOLVE, PB_PAIR_LIMIT_SOLVE);
          pb_lu_solve_in_w(S);
          pb_pair_leave(jn_);
          PB_T1(S, ts_, 1);
        }
        const double scale = (S.scalesol && S.cjratio != 1.0) ? PB_IDIV(2.0, 1.0 + S.cjratio) : 1.0;
        double acc = 0.0;
        {
#if PB_GLOBAL_VEC
          // everything in shared memory: yy = yypred + ee is advanced by the same increment as ee
          const double* PB_RESTRICT w = S_W(S);
          PB_VEC_FOR(k, i) {
            if (i < PBM_N) {
              const double d = w[ipc[k]] * scale;
              ee[i] += d;
              yy[i] += d;
              const double v = d * ewt[i];
              acc += v * v;
            }
          }
#else
          double ev[PB_VPT], pv[PB_VPT], wv[PB_VPT];
          PB_VEC_FOR(k, i) {
            ev[k] = (i < PBM_N) ? ee[i] : 0.0;
            pv[k] = (i < PBM_N) ? yypred[i] : 0.0;
            wv[k] = (i < PBM_N) ? ewt[i] : 0.0;
          }
          const double* PB_RESTRICT w = S_W(S);
          PB_VEC_FOR(k, i) {
            if (i < PBM_N) {
              const double d = w[ipc[k]] * scale;
              const double e = ev[k] + d;
              ee[i] = e;
              yy[i] = pv[k] + e;
              const double v = d * wv[k];
              acc += v * v;
            }
          }
#endif
        }
        PB_SYNC();
        const double delnrm = PB_RMS(pb_team_sum(S.team, acc));
        retval = -99;
        if (!(delnrm == delnrm)) {
          retval = PB_NLS_CONV_RECVR;
          break;
        }
        if (mnewt == 0) {
          oldnrm = delnrm;
          if (delnrm <= 1e-4 * S.toldel) retval = PB_NLS_SUCCESS;
        } else {
          const double rate = PB_IPOW(PB_IDIV(delnrm, oldnrm), PB_IDIV(1.0, (double)mnewt));
          if (rate > 0.9) {
            retval = PB_NLS_CONV_RECVR;
            break;
          }
          S.ss = PB_IDIV(rate, 1.0 - rate);
        }
        if (retval != PB_NLS_SUCCESS && S.ss * delnrm <= S.epsNewt) retval = PB_NLS_SUCCESS;
        if (retval == PB_NLS_SUCCESS) return PB_NLS_SUCCESS;
        mnewt++;
        if (mnewt >= S.maxcor) {
          retval = PB_NLS_CONV_RECVR;
          break;
        }
        {
          pb_align(S);
          PB_T0(tr_);
          const int jn_ = pb_pair_enter(PB_FN_RES, PB_PAIR_LIMIT_RES);
#ifdef PB_PAIR_DEBUG
          const long long dbg0_ = clock64();
#endif
          pbm_res(yy, S_P(S), S.tn, 0.0, delta, S_TMP(S));
          PB_STAT(S, PB_ST_NRE);
          PB_SYNC();
#ifdef PB_PAIR_DEBUG
          if (PB_TID == 0) {
            S_STATS(S)[jn_ ? 5 : 4] += (double)(clock64() - dbg0_);
            S_STATS(S)[jn_ ? 7 : 6] += 1.0;
          }
#endif
          pb_pair_leave(jn_);
          PB_T1(S, tr_, 0);
        }
      }
    }
    if (retval > 0 && !jcur) {
      callLSetup = 1;
      for (int i = PB_TID; i < PBM_N; i += PB_TEAM) {
        ee[i] = 0.0;
        yy[i] = yypred[i];
      }
      PB_SYNC();
      continue;
    }
    return retval;
  }
}

static PB_DEV int pb_nls(pb_sys& S) {
  PB_T0(tn_);
  const int r = pb_nls_impl(S);
  PB_T1(S, tn_, 4);   // inclusive of phases 0-3 spent inside the Newton solve
  return r;
}

// IDATestError
static PB_DEV int pb_test_error(pb_sys& S, double ck, double* err_k, double* err_km1) {
  const int kk = S.kk;
  const double* ee = S_EE(S);
  const double* ewt = S_EWT(S);
  const double* sigma = S_SIGMA(S);
  const double* phik = S_PHI(S, kk);
  const double* phikm1 = S_PHI(S, kk > 0 ? kk - 1 : 0);
  double a0 = 0.0, a1 = 0.0, a2 = 0.0;
  {
    double wv[PB_VPT], ev[PB_VPT], p1[PB_VPT], p2[PB_VPT];
    PB_VEC_FOR(k, i) {
      const bool in = i < PBM_N;
      wv[k] = in ? ewt[i] : 0.0;
      ev[k] = in ? ee[i] : 0.0;
      p1[k] = (in && kk > 1) ? phik[i] : 0.0;
      p2[k] = (in && kk > 2) ? phikm1[i] : 0.0;
    }
    PB_VEC_FOR(k, i) {
      const double w = wv[k];
      double v = ev[k];
      a0 += (v * w) * (v * w);
      if (kk > 1) {
        v += p1[k];
        a1 += (v * w) * (v * w);
        if (kk > 2) {
          v += p2[k];
          a2 += (v * w) * (v * w);
        }
      }
    }
  }
  const double enorm_k = PB_RMS(pb_team_sum(S.team, a0));
  *err_k = sigma[kk] * enorm_k;
  const double terr_k = (kk + 1) * (*err_k);
  S.knew = kk;
  if (kk > 1) {
    *err_km1 = sigma[kk - 1] * PB_RMS(pb_team_sum(S.team, a1));
    const double terr_km1 = kk * (*err_km1);
    if (kk > 2) {
      const double err_km2 = sigma[kk - 2] * PB_RMS(pb_team_sum(S.team, a2));
      const double terr_km2 = (kk - 1) * err_km2;
      if (fmax(terr_km1, terr_km2) <= terr_k) S.knew = kk - 1;
    } else if (terr_km1 <= 0.5 * terr_k) {
      S.knew = kk - 1;
    }
  }
  return (ck * enorm_k > 1.0) ? PB_ERROR_TEST_FAIL : 0;
}

// IDAHandleNFlag
static PB_DEV int pb_handle_nflag(pb_sys& S, int nflag, double err_k, double err_km1, int* ncf, int* nef) {
  S.phase = 1;
  if (nflag != PB_ERROR_TEST_FAIL) {
    (*ncf)++;
    PB_STAT(S, PB_ST_NCFN);
    if (nflag < 0) return nflag;
    S.rr = 0.25;
    S.hh *= S.rr;
    if (*ncf < S.maxncf) return PB_PREDICT_AGAIN;
    return (nflag == PB_NLS_LSETUP_RECVR) ? PB_IDA_LSETUP_FAIL : PB_IDA_CONV_FAIL;
  }
  (*nef)++;
  PB_STAT(S, PB_ST_NETF);
  if (*nef == 1) {
    const double err_knew = (S.kk == S.knew) ? err_k : err_km1;
    S.kk = S.knew;
    S.rr = 0.9 * PB_IPOW(2.0 * err_knew + 0.0001, PB_IDIV(-1.0, (double)(S.kk + 1)));
    S.rr = fmax(0.25, fmin(0.9, S.rr));
    S.hh *= S.rr;
    return PB_PREDICT_AGAIN;
  } else if (*nef == 2) {
    S.kk = S.knew;
    S.rr = 0.25;
    S.hh *= S.rr;
    return PB_PREDICT_AGAIN;
  } else if (*nef < S.maxnef) {
    S.kk = 1;
    S.rr = 0.25;
    S.hh *= S.rr;
    return PB_PREDICT_AGAIN;
  }
  return PB_IDA_ERR_FAIL;
}

// IDACompleteStep. Ends with a barrier.
static PB_DEV void pb_complete_step(pb_sys& S, double err_k, double err_km1) {
  const double* ee = S_EE(S);
  S.nst++;
  PB_STAT(S, PB_ST_NST);
  const int kdiff = S.kk - S.kused;
  S.kused = S.kk;
  S.hused = S.hh;
  if (S.knew == S.kk - 1 || S.kk == S.maxord) S.phase = 1;
  if (S.phase == 0) {
    // "after the first step, the order is not increased, as not all of the necessary information is available yet"
    // (IDACompleteStep): the raise-and-double of phase 0 starts with the second step
    if (S.nst > 1) {
      S.kk++;
      double hnew = 2.0 * S.hh;
      const double tmp = fabs(hnew) * S.hmax_inv;
      if (tmp > 1.0) hnew /= tmp;
      S.hh = hnew;
    }
  } else {
    int action = 0;  // 0 unset, 1 lower, 2 maintain, 3 raise
    double err_kp1 = 0.0, err_knew;
    if (S.knew == S.kk - 1) action = 1;
    else if (S.kk == S.maxord) action = 2;
    else if (S.kk + 1 >= S.ns || kdiff == 1) action = 2;
    if (action == 0) {
      const double* phikp1 = S_PHI(S, S.kk + 1);
      const double* ewt = S_EWT(S);
      double acc = 0.0;
      {
        double ev[PB_VPT], pv[PB_VPT], wv[PB_VPT];
        PB_VEC_FOR(k, i) {
          const bool in = i < PBM_N;
          ev[k] = in ? ee[i] : 0.0;
          pv[k] = in ? phikp1[i] : 0.0;
          wv[k] = in ? ewt[i] : 0.0;
        }
        PB_VEC_FOR(k, i) {
          const double v = (ev[k] - pv[k]) * wv[k];
          acc += v * v;
        }
      }
      const double enorm = PB_RMS(pb_team_sum(S.team, acc));
      err_kp1 = PB_IDIV(enorm, (double)(S.kk + 2));
      const double terr_k = (S.kk + 1) * err_k, terr_kp1 = (S.kk + 2) * err_kp1;
      if (S.kk == 1) {
        action = (terr_kp1 >= 0.5 * terr_k) ? 2 : 3;
      } else {
        const double terr_km1 = S.kk * err_km1;
        if (terr_km1 <= fmin(terr_k, terr_kp1)) action = 1;
        else if (terr_kp1 >= terr_k) action = 2;
        else action = 3;
      }
    }
    if (action == 3) {
      S.kk++;
      err_knew = err_kp1;
    } else if (action == 1) {
      S.kk--;
      err_knew = err_km1;
    } else {
      err_knew = err_k;
    }
    double hnew = S.hh;
    S.rr = PB_IPOW(2.0 * err_knew + 0.0001, PB_IDIV(-1.0, (double)(S.kk + 1)));
    if (S.rr >= 2.0) {
      hnew = 2.0 * S.hh;
      const double tmp = fabs(hnew) * S.hmax_inv;
      if (tmp > 1.0) hnew /= tmp;
    } else if (S.rr <= 1.0) {
      S.rr = fmax(0.5, fmin(0.9, S.rr));
      hnew = S.hh * S.rr;
    }
    S.hh = hnew;
  }
  // history update: phi[kused+1] = ee (if kused < maxord); phi[kused] += ee; phi[j] += phi[j+1]
  {
    double* phi = S_PHI(S, 0);
    const int ku = S.kused, save = (ku < S.maxord);
    double pv[PB_VPT][PB_MXORDP1], ev[PB_VPT];
    PB_VEC_FOR(k, i) {
      ev[k] = (i < PBM_N) ? ee[i] : 0.0;
#pragma unroll
      for (int j = 0; j < PB_MXORDP1; j++) pv[k][j] = (j <= ku && i < PBM_N) ? phi[j * PBM_N + i] : 0.0;
    }
    PB_VEC_FOR(k, i) {
      if (i < PBM_N) {
        const double e = ev[k];
        if (save) phi[(ku + 1) * PBM_N + i] = e;
        double run = e;
#pragma unroll
        for (int j = PB_MXORDP1 - 1; j >= 0; j--) {
          if (j <= ku) {
            run += pv[k][j];
            phi[j * PBM_N + i] = run;
          }
        }
      }
    }
  }
  PB_SYNC();
}

// IDAStep
static PB_DEV int pb_ida_step_impl(pb_sys& S) {
  const double saved_t = S.tn;
  double ck = 0.0, err_k = 0.0, err_km1 = 0.0;
  int ncf = 0, nef = 0;
  if (S.nst == 0) {
    S.kk = 1;
    S.kused = 0;
    S.hused = 0.0;
    if (PB_TID == 0) S_PSI(S)[0] = S.hh;
    S.cj = 1.0 / S.hh;
    S.phase = 0;
    S.ns = 0;
    PB_SYNC();
  }
  for (;;) {
    pb_set_coeffs(S, &ck);
    int nflag = pb_nls(S);
    if (nflag == PB_NLS_SUCCESS) nflag = pb_test_error(S, ck, &err_k, &err_km1);
    if (nflag != 0) {
      pb_restore(S, saved_t, 1);
      const int kflag = pb_handle_nflag(S, nflag, err_k, err_km1, &ncf, &nef);
      if (kflag != PB_PREDICT_AGAIN) return kflag;
      if (S.nst == 0) {
        if (PB_TID == 0) S_PSI(S)[0] = S.hh;
        double* phi1 = S_PHI(S, 1);
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi1[i] *= S.rr;
        PB_SYNC();
      }
      if (!(fabs(S.hh) > 0.0) || S.tn + S.hh == S.tn) return PB_IDA_ERR_FAIL;
      continue;
    }
    break;
  }
  pb_complete_step(S, err_k, err_km1);
  return PB_IDA_SUCCESS;
}

static PB_DEV int pb_ida_step(pb_sys& S) {
  PB_T0(tt_);
  const int r = pb_ida_step_impl(S);
  PB_T1X(S, tt_, 5);   // inclusive of phase 4
  return r;
}

// ---- termination events: IDARcheck1 / IDARcheck3 / IDARootfind (Illinois) on the BDF interpolant ----
static PB_DEV void pb_gfun(pb_sys& S, double t, const double* y, double* g) {
#if PBM_NEV > 0
  pbm_events(y, S_P(S), t, 0.0, g, S_TMP(S));
  PB_STAT(S, PB_ST_NGE);
  PB_SYNC();
#endif
}

#if PBM_NEV > 0
static PB_DEV int pb_rootfind(pb_sys& S) {
  double* glo = S_GLO(S);
  double* ghi = S_GHI(S);
  double* grout = S_GROUT(S);
  int imax = 0, zroot = 0, sgnchg = 0, side = 0, sideprev = -1;
  double maxfrac = 0.0, alph = 1.0;
  for (int i = 0; i < PBM_NEV; i++) {
    if (!((S.gactive >> i) & 1u)) continue;
    if (fabs(ghi[i]) == 0.0) zroot = 1;
    else if (glo[i] * ghi[i] < 0.0) {
      const double gfrac = fabs(ghi[i] / (ghi[i] - glo[i]));
      if (gfrac > maxfrac) {
        sgnchg = 1;
        maxfrac = gfrac;
        imax = i;
      }
    }
  }
  if (!sgnchg) {
    S.trout = S.thi;
    PB_SYNC();
    if (PB_TID == 0)
      for (int i = 0; i < PBM_NEV; i++) grout[i] = ghi[i];
    PB_SYNC();
    return zroot;
  }
  for (;;) {
    if (fabs(S.thi - S.tlo) <= S.ttol) break;
    if (sideprev == side) alph = (side == 2) ? alph * 2.0 : alph * 0.5;
    else alph = 1.0;
    double tmid = S.thi - (S.thi - S.tlo) * ghi[imax] / (ghi[imax] - alph * glo[imax]);
    if (fabs(tmid - S.tlo) < 0.5 * S.ttol) {
      const double fracint = fabs(S.thi - S.tlo) / S.ttol;
      const double fracsub = (fracint > 5.0) ? 0.1 : 0.5 / fracint;
      tmid = S.tlo + fracsub * (S.thi - S.tlo);
    }
    if (fabs(S.thi - tmid) < 0.5 * S.ttol) {
      const double fracint = fabs(S.thi - S.tlo) / S.ttol;
      const double fracsub = (fracint > 5.0) ? 0.1 : 0.5 / fracint;
      tmid = S.thi - fracsub * (S.thi - S.tlo);
    }
    pb_get_solution(S, tmid, S_YY(S));
    PB_SYNC();
    pb_gfun(S, tmid, S_YY(S), grout);
    maxfrac = 0.0;
    zroot = 0;
    sgnchg = 0;
    sideprev = side;
    for (int i = 0; i < PBM_NEV; i++) {
      if (!((S.gactive >> i) & 1u)) continue;
      if (fabs(grout[i]) == 0.0) zroot = 1;
      else if (glo[i] * grout[i] < 0.0) {
        const double gfrac = fabs(grout[i] / (grout[i] - glo[i]));
        if (gfrac > maxfrac) {
          sgnchg = 1;
          maxfrac = gfrac;
          imax = i;
        }
      }
    }
    PB_SYNC();  // everybody has read grout/glo/ghi before thread 0 shuffles them
    if (sgnchg) {
      S.thi = tmid;
      if (PB_TID == 0)
        for (int i = 0; i < PBM_NEV; i++) ghi[i] = grout[i];
      PB_SYNC();
      side = 1;
      if (fabs(S.thi - S.tlo) <= S.ttol) break;
      continue;
    }
    if (zroot) {
      S.thi = tmid;
      if (PB_TID == 0)
        for (int i = 0; i < PBM_NEV; i++) ghi[i] = grout[i];
      PB_SYNC();
      break;
    }
    S.tlo = tmid;
    if (PB_TID == 0)
      for (int i = 0; i < PBM_NEV; i++) glo[i] = grout[i];
    PB_SYNC();
    side = 2;
    if (fabs(S.thi - S.tlo) <= S.ttol) break;
  }
  S.trout = S.thi;
  PB_SYNC();
  if (PB_TID == 0)
    for (int i = 0; i < PBM_NEV; i++) grout[i] = ghi[i];
  PB_SYNC();
  return 1;
}

// IDARcheck3 for the ONE_STEP task (thi = tn). Returns 1 when a root was found; y(trout) is then in S_YY.
static PB_DEV int pb_rcheck3(pb_sys& S) {
  double* glo = S_GLO(S);
  double* grout = S_GROUT(S);
  S.thi = S.tn;
  // y(tn) is phi[0] itself: the interpolant at delt = 0 has c[0] = 1 and c[j >= 1] = 0 (IDAGetSolution), so the
  // event functions read the history array directly instead of a six-array interpolation pass
  pb_gfun(S, S.thi, S_PHI(S, 0), S_GHI(S));
  S.ttol = (fabs(S.tn) + fabs(S.hh)) * PB_UROUND * 100.0;
  const int ier = pb_rootfind(S);
  for (int i = 0; i < PBM_NEV; i++)
    if (!((S.gactive >> i) & 1u) && grout[i] != 0.0) S.gactive |= (1u << i);
  S.tlo = S.trout;
  PB_SYNC();
  if (PB_TID == 0)
    for (int i = 0; i < PBM_NEV; i++) glo[i] = grout[i];
  PB_SYNC();
  if (!ier) return 0;
  pb_get_solution(S, S.trout, S_YY(S));
  PB_SYNC();
  return 1;
}
#endif

// IDASolve(tout, ONE_STEP). On return *tret is the time reached; for ROOT returns y(tret) is in S_YY.
static PB_DEV int pb_ida_solve_one_step(pb_sys& S, double tout, double* tret) {
  if (S.nst == 0) {
    if (pb_ewt_set(S, S_PHI(S, 0))) return PB_IDA_ILL_INPUT;
    const double tdist = fabs(tout - S.tn);
    if (S.dt_init != 0.0) {
      S.hh = S.dt_init;
    } else {
      S.hh = 0.001 * tdist;
      const double ypnorm = pb_wrms(S, S_PHI(S, 1));
      if (ypnorm > 0.5 / S.hh) S.hh = 0.5 / ypnorm;
      if (tout < S.tn) S.hh = -S.hh;
    }
    {
      const double rh = fabs(S.hh) * S.hmax_inv;
      if (rh > 1.0) S.hh /= rh;
    }
    if (S.tstopset && (S.tn + S.hh - S.tstop) * S.hh > 0.0) S.hh = (S.tstop - S.tn) * (1.0 - 4.0 * PB_UROUND);
    S.kk = 0;
    S.kused = 0;
    {
      double* phi1 = S_PHI(S, 1);
      for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi1[i] *= S.hh;
      PB_SYNC();
    }
#if PBM_NEV > 0
    {  // IDARcheck1
      double* glo = S_GLO(S);
      double* ghi = S_GHI(S);
      S.tlo = S.tn;
      S.ttol = (fabs(S.tn) + fabs(S.hh)) * PB_UROUND * 100.0;
      pb_gfun(S, S.tlo, S_PHI(S, 0), glo);
      int zroot = 0;
      S.gactive = 0u;
      for (int i = 0; i < PBM_NEV; i++) {
        if (fabs(glo[i]) == 0.0) zroot = 1;
        else S.gactive |= (1u << i);
      }
      if (zroot) {
        const double hratio = fmax(S.ttol / fabs(S.hh), 0.1), smallh = hratio * S.hh;
        const double* phi0 = S_PHI(S, 0);
        const double* phi1 = S_PHI(S, 1);
        double* yy = S_YY(S);
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) yy[i] = phi0[i] + hratio * phi1[i];
        PB_SYNC();
        pb_gfun(S, S.tlo + smallh, yy, ghi);
        unsigned act = S.gactive;
        for (int i = 0; i < PBM_NEV; i++)
          if (!((act >> i) & 1u) && fabs(ghi[i]) != 0.0) act |= (1u << i);
        PB_SYNC();
        if (PB_TID == 0)
          for (int i = 0; i < PBM_NEV; i++)
            if (!((S.gactive >> i) & 1u) && ((act >> i) & 1u)) glo[i] = ghi[i];
        PB_SYNC();
        S.gactive = act;
      }
    }
#endif
  } else {
    // IDAStopTest1, ONE_STEP
    if ((S.tn - S.tretlast) * S.hh > 0.0) {
      *tret = S.tretlast = S.tn;
      return PB_IDA_SUCCESS;
    }
    if (S.tstopset) {
      const double troundoff = 100.0 * PB_UROUND * (fabs(S.tn) + fabs(S.hh));
      if (fabs(S.tn - S.tstop) <= troundoff) {
        *tret = S.tretlast = S.tstop;
        S.tstopset = 0;
        return PB_IDA_TSTOP_RETURN;
      }
      if ((S.tn + S.hh - S.tstop) * S.hh > 0.0) S.hh = (S.tstop - S.tn) * (1.0 - 4.0 * PB_UROUND);
    }
  }
  {
    const double nrm = (S.nst > 0) ? pb_ewt_set_norm(S, S_PHI(S, 0)) : pb_wrms(S, S_PHI(S, 0));
    if (!(nrm == nrm)) return PB_IDA_ILL_INPUT;
    if (PB_UROUND * nrm > 1.0) return PB_IDA_TOO_MUCH_ACC;
  }
  const int sflag = pb_ida_step(S);
  if (sflag != PB_IDA_SUCCESS) {
    *tret = S.tretlast = S.tn;
    return sflag;
  }
  if (S.tstopset) {
    const double troundoff = 100.0 * PB_UROUND * (fabs(S.tn) + fabs(S.hh));
    if (fabs(S.tn - S.tstop) <= troundoff) S.tn = S.tstop;
  }
#if PBM_NEV > 0
  {
    PB_T0(tg_);
    const int root = pb_rcheck3(S);
    PB_T1X(S, tg_, 6);
    if (root) {
      *tret = S.tretlast = S.tlo;
      return PB_IDA_ROOT_RETURN;
    }
  }
#endif
  // IDAStopTest2, ONE_STEP
  if (S.tstopset) {
    const double troundoff = 100.0 * PB_UROUND * (fabs(S.tn) + fabs(S.hh));
    if (fabs(S.tn - S.tstop) <= troundoff) {
      *tret = S.tretlast = S.tstop;
      S.tstopset = 0;
      return PB_IDA_TSTOP_RETURN;
    }
    if ((S.tn + S.hh - S.tstop) * S.hh > 0.0) S.hh = (S.tstop - S.tn) * (1.0 - 4.0 * PB_UROUND);
  }
  *tret = S.tretlast = S.tn;
  return PB_IDA_SUCCESS;
}

// ---- consistent initialisation (IDAKLUSolverOpenMP.inl:787-873) ---------------------------------
// y' = f(t, y) on the differential rows, 0 on the algebraic rows: phi[1] <- y'(phi[0]). Ends with a barrier.
static PB_DEV void pb_recover_yp(pb_sys& S, double t) {
  double* phi1 = S_PHI(S, 1);
  pbm_res(S_PHI(S, 0), S_P(S), t, 0.0, phi1, S_TMP(S));
  PB_SYNC();
  for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) phi1[i] = 0.0;
  PB_SYNC();
}

#if PBM_N > PBM_NDIFF
#define PB_NALG (PBM_N - PBM_NDIFF)
// max-norm of the algebraic residual at y (result also left in S_DELTA). Ends with a barrier.
static PB_DEV double pb_alg_res_norm(pb_sys& S, double t, const double* y) {
  double* res = S_DELTA(S);
  pbm_res(y, S_P(S), t, 0.0, res, S_TMP(S));
  PB_SYNC();
  double mx = 0.0;
  for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) {
    const double a = fabs(res[i]);
    mx = (a != a || mx != mx) ? PB_NAN : fmax(mx, a);
  }
  mx = pb_team_max(S.team, mx);
  PB_SYNC();
  return mx;
}

// NonlinearSolver::RunNewtonLoop (NonlinearSolver.inl:100-177) on the algebraic sub-block; the
// sub-block system J_aa d = g is solved through the full static LU with identity differential rows.
// y (= S_YY) is updated in place.  Returns 1 on success.
static PB_DEV int pb_newton_ic(pb_sys& S, double t, double* init_norm, double* final_norm, const int max_iter,
                               const int max_backtracks) {
  double* y = S_YY(S);
  double* xsave = S_EE(S);
  double* res = S_DELTA(S);
  double* del = S_YYPRED(S);
  double* ewt = S_EWT(S);
  double delnorm = PB_INF, prev_res_norm = PB_INF;
  int converged = 0;
  for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) {
    ewt[i] = 1.0 / (S.rtol * fabs(y[i]) + PB_LD(S.atol[i]));
    xsave[i] = y[i];
  }
  PB_SYNC();
  for (int iter = 0; iter < max_iter; iter++) {
    const double res_norm = pb_alg_res_norm(S, t, y);
    if (iter == 0) *init_norm = res_norm;
    *final_norm = res_norm;
    if (pb_lsetup(S, t, y, 0.0, 1)) return 0;
    for (int i = PB_TID; i < PBM_NDIFF; i += PB_TEAM) res[i] = 0.0;
    PB_SYNC();
    pb_lu_solve(S, res);
    double acc = 0.0;
    for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) {
      const double d = res[i];
      del[i] = d;
      const double v = d * ewt[i];
      acc += v * v;
    }
    PB_SYNC();
    delnorm = sqrt(pb_team_sum(S.team, acc) / PB_NALG);
    if (delnorm <= S.epsNewtIC) {
      converged = 1;
      if (delnorm <= S.step_tol_ic) {
        for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) y[i] -= del[i];
        PB_SYNC();
        return 1;
      }
      if (iter > 0 && res_norm >= prev_res_norm) {
        for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) y[i] = xsave[i];
        PB_SYNC();
        return 1;
      }
    }
    if (!(delnorm == delnorm)) return 0;
    prev_res_norm = res_norm;
    for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) xsave[i] = y[i];
    PB_SYNC();
    double alpha = 1.0;
    for (int ls = 0; ls < max_backtracks; ls++) {
      for (int i = PBM_NDIFF + PB_TID; i < PBM_N; i += PB_TEAM) y[i] = xsave[i] - alpha * del[i];
      PB_SYNC();
      const double trial = pb_alg_res_norm(S, t, y);
      if (trial <= (1.0 - alpha * 0.5) * res_norm) break;
      if (alpha * delnorm <= S.step_tol_ic) break;
      alpha *= 0.5;
    }
  }
  return converged;
}
#endif

// ConsistentInitialization + IDAReInit at time t; on entry phi[0] holds y.
static PB_DEV void pb_consistent_init(pb_sys& S, double t, int calc_ic) {
#if PBM_N > PBM_NDIFF
  if (calc_ic) {
    double* phi0 = S_PHI(S, 0);
    double* y = S_YY(S);
    double* backup = S_YPPRED(S);
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) {
      y[i] = phi0[i];
      backup[i] = phi0[i];
    }
    PB_SYNC();
    // First the Newton loop of NonlinearSolver (five halvings per iteration); when it fails AND leaves a larger residual
    // than it started from, the reference restores the state and hands over to IDACalcIC (IDAKLUSolverOpenMP.inl:838-850),
    // whose Newton carries a proper backtracking line search.  The second pass below plays that role: the same iteration
    // from the restored state with up to thirty halvings per iteration (a step is always found unless the Jacobian is singular).
    for (int attempt = 0; attempt < 2; attempt++) {
      double in = 0.0, fn = 0.0;
      const int ok = pb_newton_ic(S, t, &in, &fn, S.max_iter_ic, attempt ? 30 : S.max_backtracks_ic);
      if (ok || fn <= in) {
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi0[i] = y[i];
        PB_SYNC();
        break;
      }
      for (int i = PB_TID; i < PBM_N; i += PB_TEAM) y[i] = backup[i];
      PB_SYNC();
    }
  }
#endif
  pb_recover_yp(S, t);
  pb_reinit(S, t);
}

// ---- the PyBaMM driver loop (IDAKLUSolverOpenMP::solve, IDAKLUSolverOpenMP.inl:327-444) ----------
static PB_DEV double pb_perturb_time(double t) {
  const double delta = 1.4901161193847656e-08;  // sqrt(DBL_EPSILON)
  return (1.0 + delta) * t + delta;
}

// Store the solution at time t (a point inside the last step) as output point *i_save.  Points beyond the block's
// capacity are counted but not stored: n_t > n_max tells the host that the block was too small (the reference grows
// its arrays without bound, ExtendAdaptiveArrays, IDAKLUSolverOpenMP.inl:608-612; the host re-solves with n_max = n_t).
// Evaluate the output variables of the pending points (all lanes busy) and write their rows.  Called when the batch is
// full, before the inputs change (experiment steps) and when the system is done.
static PB_DEV void pb_vars_flush(pb_sys& S) {
#if PB_VARS_BATCH > 1
  if (S.vb_n > 0) {
    PB_SYNC();
    pbm_vars_multi(S.vb, S_P(S), S.vb + PB_VARS_BATCH * PBM_N, S.vb_n, S.vb_row0,
                   S.vb + PB_VARS_BATCH * (PBM_N + 1));
    PB_SYNC();
    S.vb_n = 0;
  }
#else
  (void)S;
#endif
}

static PB_DEV void pb_store_point(pb_sys& S, const pbgpu_launch& L, int sys, double t, int* i_save, double* t_last) {
  *t_last = t;
  if (*i_save >= L.n_max) {
    (*i_save)++;
    return;
  }
  PB_T0(to_);
  const size_t slot = (size_t)sys * L.n_max + *i_save;
  double* row = L.y_out + slot * L.n_ret;
  if (L.save_vars) {
#if PBM_NVAR > 0
#if PB_VARS_BATCH > 1
    pb_get_solution(S, t, S.vb + S.vb_n * PBM_N);
    if (PB_TID == 0) S.vb[PB_VARS_BATCH * PBM_N + S.vb_n] = t;
    if (S.vb_n == 0) S.vb_row0 = row;
    S.vb_n++;
    if (S.vb_n == PB_VARS_BATCH) pb_vars_flush(S);
#else
    pb_get_solution(S, t, S_YY(S));
    PB_SYNC();
    pbm_vars(S_YY(S), S_P(S), t, 0.0, row, S_TMP(S));
    PB_SYNC();
#endif
#endif
  } else if (L.yp_out) {
    // Hermite data: y and dy/dt from the same interpolating polynomial (IDAGetDky k = 0, 1; save_hermite,
    // IDAKLUSolverOpenMP.inl:432-435, 447-457)
    pb_get_solution2(S, t, row, L.yp_out + slot * PBM_N);
  } else {
    pb_get_solution(S, t, row);
  }
  if (PB_TID == 0) L.t_out[slot] = t;
  (*i_save)++;
  PB_T1X(S, to_, 7);
}

static PB_DEV void pb_solve_system(pb_sys& S, const pbgpu_launch& L, int sys) {
  const double t0 = PB_LDU(L.t_eval[0]), tf = PB_LDU(L.t_eval[L.n_eval - 1]);
  const double tf_perturbed = pb_perturb_time(tf);
  PB_SYNC();
  for (int k = PB_TID; k < PBM_NP_REAL; k += PB_TEAM) S_P(S)[k] = L.inputs[(size_t)sys * PBM_NP_REAL + k];
  for (int k = PB_TID; k < 8; k += PB_TEAM) S_STATS(S)[k] = 0.0;
  PB_SYNC();
  if (L.y0) {
    double* phi0 = S_PHI(S, 0);
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi0[i] = L.y0[(size_t)sys * PBM_N + i];
  } else {
    pbm_ic(S_YY(S), S_P(S), t0, 0.0, S_PHI(S, 0), S_TMP(S));
  }
  PB_SYNC();
  pb_consistent_init(S, t0, L.opt.calc_ic);
  if (L.yp0 && !L.opt.calc_ic) {
    // calc_ic off: the caller's derivatives are used as given (SetupInitialState, IDAKLUSolverOpenMP.inl:514-563)
    double* phi1 = S_PHI(S, 1);
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi1[i] = L.yp0[(size_t)sys * PBM_N + i];
    PB_SYNC();
  }
  // NoProgressGuard (NoProgressGuard.hpp:12-66): thread 0 keeps a window of the last step sizes, initialised with the
  // threshold so that it cannot trip before the window is full; the verdict travels through shared memory
  const int npg_n = (L.opt.t_no_progress != 0.0) ? L.opt.num_steps_no_progress : 0;
  int npg_idx = 0;
  if (npg_n > 0) {
    if (PB_TID == 0) {
      for (int k = 0; k < npg_n; k++) S_NPG(S)[k] = L.opt.t_no_progress;
      S_NPG(S)[PB_NPG_MAX] = 0.0;
    }
    PB_SYNC();
  }
  int ev_idx = -1;
  int i_save = 0, i_eval = 1, i_interp = 0;
  double t_eval_next = PB_LDU(L.t_eval[1]);
  S.tstop = t_eval_next;
  S.tstopset = 1;
  int retval = PB_IDA_SUCCESS;
  double t_val = t0, t_prev = t0, t_last = t0;
  int done = 0;
#if PBM_NEV > 0
  {
    double* g = S_GHI(S);
    pb_gfun(S, t0, S_PHI(S, 0), g);
    int trig = 0;
    {
      double best = PB_INF;
      for (int i = 0; i < PBM_NEV; i++) {
        trig |= (g[i] <= 0.0);
        if (fabs(g[i]) < best) {
          best = fabs(g[i]);
          ev_idx = i;
        }
      }
      if (!trig) ev_idx = -1;
    }
    PB_SYNC();
    if (trig) {
      // IDA has not stepped: psi is not set, store phi[0] directly
      if (i_save < L.n_max) {
        const size_t slot = (size_t)sys * L.n_max;
        double* row = L.y_out + slot * L.n_ret;
        if (L.save_vars) {
#if PBM_NVAR > 0
          pbm_vars(S_PHI(S, 0), S_P(S), t0, 0.0, row, S_TMP(S));
          PB_SYNC();
#endif
        } else {
          const double* phi0 = S_PHI(S, 0);
          const double* phi1 = S_PHI(S, 1);
          for (int i = PB_TID; i < PBM_N; i += PB_TEAM) {
            row[i] = phi0[i];
            if (L.yp_out) L.yp_out[slot * PBM_N + i] = phi1[i];
          }
        }
        if (PB_TID == 0) L.t_out[slot] = t0;
      }
      i_save = 1;
      retval = PB_IDA_ROOT_RETURN;
      done = 1;
      if (L.y_term) {
        const double* phi0 = S_PHI(S, 0);
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) L.y_term[(size_t)sys * PBM_N + i] = phi0[i];
      }
    }
  }
#endif
  if (!done) {
    retval = pb_ida_solve_one_step(S, tf_perturbed, &t_val);
    if (retval >= 0) pb_store_point(S, L, sys, t0, &i_save, &t_last);
    long long guard = 0;
    if (npg_n > 0) {
      if (PB_TID == 0) {
        S_NPG(S)[0] = t_val - t0;
        npg_idx = 1 % npg_n;
      }
    }
    for (;;) {
      if (retval < 0) break;
      int stalled = 0;
      if (npg_n > 0) {
        PB_SYNC();
        if (PB_TID == 0) {
          double sum = 0.0;
          int viol = 1;
          for (int k = 0; k < npg_n; k++) {
            sum += S_NPG(S)[k];
            if (sum >= L.opt.t_no_progress) {
              viol = 0;
              break;
            }
          }
          S_NPG(S)[PB_NPG_MAX] = viol ? 1.0 : 0.0;
        }
        PB_SYNC();
        stalled = S_NPG(S)[PB_NPG_MAX] != 0.0;
      }
      if (t_prev == t_val || stalled) {
        retval = PB_IDA_ERR_FAIL;
        break;
      }
      const int hit_teval = retval == PB_IDA_TSTOP_RETURN;
      const int hit_final = t_val >= tf || (hit_teval && i_eval == L.n_eval);
      const int hit_event = retval == PB_IDA_ROOT_RETURN;
      const int hit_adaptive = (L.n_interp == 0) && retval == PB_IDA_SUCCESS;
      const int final_step = hit_final || hit_event;
      if (L.n_interp > 0 && i_interp < L.n_interp && PB_LDU(L.t_interp[i_interp]) >= t_prev) {
        while (i_interp < L.n_interp) {
          const double ti = PB_LDU(L.t_interp[i_interp]);
          if (!(ti <= t_val)) break;
          pb_store_point(S, L, sys, ti, &i_save, &t_last);
          i_interp++;
        }
      }
      if (hit_adaptive || hit_teval || hit_event || hit_final) {
        // duplicate check against the last stored time (SavePoint, IDAKLUSolverOpenMP.inl:608-616)
        if (i_save == 0 || t_val != t_last) pb_store_point(S, L, sys, t_val, &i_save, &t_last);
      }
      if (final_step) break;
      if (hit_teval) {
        i_eval++;
        t_eval_next = PB_LDU(L.t_eval[i_eval]);
        PB_SYNC();
        pb_get_solution2(S, t_val, S_YY(S), S_YPPRED(S));
        PB_SYNC();
        {
          double* phi0 = S_PHI(S, 0);
          const double* yy = S_YY(S);
          for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi0[i] = yy[i];
          PB_SYNC();
        }
        pb_consistent_init(S, t_val, 1);
        S.tstop = t_eval_next;
        S.tstopset = 1;
        if (npg_n > 0) {   // HandleBreakpoint re-initialises the guard
          PB_SYNC();
          if (PB_TID == 0) {
            for (int k = 0; k < npg_n; k++) S_NPG(S)[k] = L.opt.t_no_progress;
            npg_idx = 0;
          }
        }
      }
      t_prev = t_val;
      if (++guard > S.max_num_steps) {
        retval = PB_IDA_TOO_MUCH_WORK;
        break;
      }
      retval = pb_ida_solve_one_step(S, tf_perturbed, &t_val);
      if (npg_n > 0 && PB_TID == 0) {
        S_NPG(S)[npg_idx] = t_val - t_prev;
        npg_idx = (npg_idx + 1) % npg_n;
      }
    }
#if PBM_NEV > 0
    if (retval == PB_IDA_ROOT_RETURN) {
      // which event ended the run: the one closest to zero at the root (closest_event_idx, idaklu_solver.py:826-834);
      // S_GROUT holds g(t_root, y(t_root)) of the root finder
      const double* gr = S_GROUT(S);
      double best = PB_INF;
      for (int i = 0; i < PBM_NEV; i++) {
        const double a = fabs(gr[i]);
        if (a < best) {
          best = a;
          ev_idx = i;
        }
      }
    }
#endif
    if (L.y_term) {
      PB_SYNC();
      if (S.nst > 0 || S.kused > 0) pb_get_solution(S, t_val, L.y_term + (size_t)sys * PBM_N);
      else {
        const double* phi0 = S_PHI(S, 0);
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) L.y_term[(size_t)sys * PBM_N + i] = phi0[i];
      }
    }
  }
  pb_vars_flush(S);
  PB_SYNC();
  if (PB_TID == 0) {
    L.n_out[sys] = i_save;
    L.flag[sys] = retval;
    if (L.event_idx) L.event_idx[sys] = (retval == PB_IDA_ROOT_RETURN) ? ev_idx : -1;
    if (L.t_event) L.t_event[sys] = t_last;
    if (L.stats)
      for (int k = 0; k < 8; k++) L.stats[(size_t)sys * 8 + k] = (long long)S_STATS(S)[k];
  }
}

// ---- experiment schedule on the device (Simulation.solve's step loop, simulation.py:991-1167) ----------------------------
// One call integrates every step of the schedule for one system: the step's inputs are switched, the state is carried over,
// the algebraic states are re-initialised, and the step runs to its duration or to a root of the experiment's termination
// event.  See pbgpu_schedule in include/pbgpu.h for the per-step rules and the citations.
static PB_DEV double pb_next_after(double t) {
  if (t == 0.0) return t;   // base_solver.py:1380-1390: the first step starts at 0 exactly
#ifndef PB_EMU
  return __longlong_as_double(__double_as_longlong(t) + (t > 0.0 ? 1LL : -1LL));
#else
  return std::nextafter(t, INFINITY);
#endif
}

static PB_DEV void pb_solve_experiment(pb_sys& S, const pbgpu_launch& L, int sys) {
  const pbgpu_schedule& Q = L.sched;
  PB_SYNC();
  for (int k = PB_TID; k < PBM_NP_REAL; k += PB_TEAM) S_P(S)[k] = L.inputs[(size_t)sys * PBM_NP_REAL + k];
  for (int k = PB_TID; k < 8; k += PB_TEAM) S_STATS(S)[k] = 0.0;
  for (int k = PB_TID; k < Q.n_steps; k += PB_TEAM) {
    L.step_end[(size_t)sys * Q.n_steps + k] = 0;
    L.step_term[(size_t)sys * Q.n_steps + k] = PBGPU_STEP_NOT_RUN;
  }
  PB_SYNC();
  for (int j = PB_TID; j < Q.n_over; j += PB_TEAM) S_P(S)[Q.over_slot[j]] = Q.over_val[j];   // step 0: y0(p) may depend on them
  PB_SYNC();
  double t_cur = L.t_start;
  if (L.y0) {
    double* phi0 = S_PHI(S, 0);
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi0[i] = L.y0[(size_t)sys * PBM_N + i];
  } else {
    pbm_ic(S_YY(S), S_P(S), t_cur, 0.0, S_PHI(S, 0), S_TMP(S));
  }
  PB_SYNC();
  int i_save = 0, overall = PB_IDA_TSTOP_RETURN, ev_idx = -1;
  double t_last = t_cur;
  int have_history = 0;   // phi holds a stepped history of the previous step (so y(t_cur) comes from the interpolant)
  for (int k = 0; k < Q.n_steps; k++) {
    pb_vars_flush(S);   // pending output points belong to the inputs of the step that produced them
    PB_SYNC();
    for (int j = PB_TID; j < Q.n_over; j += PB_TEAM) S_P(S)[Q.over_slot[j]] = Q.over_val[(size_t)k * Q.n_over + j];
    PB_SYNC();
    if (have_history) {   // carry the state over: y(t_cur) of the finished step becomes the new start state
      pb_get_solution(S, t_cur, S_YY(S));
      PB_SYNC();
      double* phi0 = S_PHI(S, 0);
      const double* yy = S_YY(S);
      for (int i = PB_TID; i < PBM_N; i += PB_TEAM) phi0[i] = yy[i];
      PB_SYNC();
      have_history = 0;
    }
    const double t0 = pb_next_after(t_cur);   // base_solver.py:1380-1390 (0 stays 0)
    const double dur = PB_LDU(Q.duration[k]);
    const double tf = t_cur + dur;
    const int npts = Q.n_pts[k] < 2 ? 2 : Q.n_pts[k];
    const double dt_out = dur / (npts - 1);
    pb_consistent_init(S, t0, 1);
    int term = PBGPU_STEP_NOT_RUN;
#if PBM_NEV > 0
    {   // termination events at the start of the step (IDAKLUSolverOpenMP.inl:365-378, simulation.py:1062-1070)
      double* g = S_GHI(S);
      pb_gfun(S, t0, S_PHI(S, 0), g);
      int trig_exp = 0, trig_other = 0;
      double best = PB_INF;
      for (int i = 0; i < PBM_NEV; i++) {
        if (g[i] <= 0.0) {
          if (i == Q.experiment_event) trig_exp = 1;
          else trig_other = 1;
          if (fabs(g[i]) < best) {
            best = fabs(g[i]);
            ev_idx = i;
          }
        }
      }
      PB_SYNC();
      if (trig_exp && !trig_other) term = PBGPU_STEP_SKIPPED;
      else if (trig_other) term = PBGPU_STEP_OTHER_EVENT;
    }
#endif
    if (term == PBGPU_STEP_SKIPPED) {
      if (PB_TID == 0) {
        L.step_end[(size_t)sys * Q.n_steps + k] = i_save;
        L.step_term[(size_t)sys * Q.n_steps + k] = term;
      }
      continue;   // t_cur and the state stay; the next step re-initialises from phi[0]
    }
    if (term == PBGPU_STEP_OTHER_EVENT) {
      overall = PB_IDA_ROOT_RETURN;
      if (PB_TID == 0) {
        L.step_end[(size_t)sys * Q.n_steps + k] = i_save;
        L.step_term[(size_t)sys * Q.n_steps + k] = term;
      }
      break;
    }
    S.tstop = tf;
    S.tstopset = 1;
    double t_val = t0, t_prev = t0;
    int retval = pb_ida_solve_one_step(S, pb_perturb_time(tf), &t_val);
    if (retval >= 0) {
      pb_store_point(S, L, sys, t0, &i_save, &t_last);
      have_history = 1;
    }
    int j_out = 1;
    long long guard = 0;
    for (;;) {
      if (retval < 0) break;
      if (t_prev == t_val) {
        retval = PB_IDA_ERR_FAIL;
        break;
      }
      while (j_out < npts - 1) {
        const double ti = t_cur + j_out * dt_out;
        if (!(ti <= t_val)) break;
        if (ti > t0) pb_store_point(S, L, sys, ti, &i_save, &t_last);
        j_out++;
      }
      const int hit_final = retval == PB_IDA_TSTOP_RETURN || t_val >= tf;
      const int hit_event = retval == PB_IDA_ROOT_RETURN;
      if (hit_final || hit_event) {
        if (t_val != t_last) pb_store_point(S, L, sys, t_val, &i_save, &t_last);
        break;
      }
      t_prev = t_val;
      if (++guard > S.max_num_steps) {
        retval = PB_IDA_TOO_MUCH_WORK;
        break;
      }
      retval = pb_ida_solve_one_step(S, pb_perturb_time(tf), &t_val);
    }
    if (retval < 0) {
      term = retval;
      overall = retval;
    } else if (retval == PB_IDA_ROOT_RETURN) {
#if PBM_NEV > 0
      const double* gr = S_GROUT(S);
      double best = PB_INF;
      for (int i = 0; i < PBM_NEV; i++) {
        const double a = fabs(gr[i]);
        if (a < best) {
          best = a;
          ev_idx = i;
        }
      }
#endif
      term = (ev_idx == Q.experiment_event) ? PBGPU_STEP_EXPERIMENT_EVENT : PBGPU_STEP_OTHER_EVENT;
      if (term == PBGPU_STEP_OTHER_EVENT) overall = PB_IDA_ROOT_RETURN;
    } else {
      term = PBGPU_STEP_FINAL_TIME;
      if (Q.stop_on_final_time && Q.stop_on_final_time[k]) overall = PB_IDA_SUCCESS;   // infeasible: the default duration ran out
    }
    PB_SYNC();
    if (PB_TID == 0) {
      L.step_end[(size_t)sys * Q.n_steps + k] = i_save;
      L.step_term[(size_t)sys * Q.n_steps + k] = term;
    }
    t_cur = t_val;
    if (term != PBGPU_STEP_EXPERIMENT_EVENT && !(term == PBGPU_STEP_FINAL_TIME && overall == PB_IDA_TSTOP_RETURN)) break;
  }
  if (L.y_term) {
    PB_SYNC();
    if (have_history) pb_get_solution(S, t_cur, L.y_term + (size_t)sys * PBM_N);
    else {
      const double* phi0 = S_PHI(S, 0);
      for (int i = PB_TID; i < PBM_N; i += PB_TEAM) L.y_term[(size_t)sys * PBM_N + i] = phi0[i];
    }
  }
  pb_vars_flush(S);
  PB_SYNC();
  if (PB_TID == 0) {
    L.n_out[sys] = i_save;
    L.flag[sys] = overall;
    if (L.event_idx) L.event_idx[sys] = (overall == PB_IDA_ROOT_RETURN) ? ev_idx : -1;
    if (L.t_event) L.t_event[sys] = t_last;
    if (L.stats)
      for (int q = 0; q < 8; q++) L.stats[(size_t)sys * 8 + q] = (long long)S_STATS(S)[q];
  }
}

// ---- variable observation (observe.cpp:18-354) -------------------------------------------------------------------------
// One team per solution: for every requested time, the state (stored, or cubic-Hermite interpolated from y and y' of the
// enclosing stored interval exactly as HermiteInterpolator::compute_knots / interpolate do) goes to shared memory and the
// generated output-variable function writes the row.  S.sm and S_P(S) must be set; uses S_YY and S_TMP.
#if PBM_NVAR > 0
static PB_DEV void pb_observe_system(pb_sys& S, const pbgpu_observe_launch& O, int sys) {
  const int nt = O.n_t[sys];
  const double* t = O.t + (size_t)sys * O.n_max;
  const double* y = O.y + (size_t)sys * O.n_max * PBM_N;
  double* yy = S_YY(S);
  PB_SYNC();
  for (int k = PB_TID; k < PBM_NP_REAL; k += PB_TEAM) S_P(S)[k] = O.inputs[(size_t)sys * PBM_NP_REAL + k];
  PB_SYNC();
  if (O.n_obs == 0) {
    for (int i = 0; i < nt; i++) {
      for (int q = PB_TID; q < PBM_N; q += PB_TEAM) yy[q] = y[(size_t)i * PBM_N + q];
      PB_SYNC();
      pbm_vars(yy, S_P(S), PB_LDU(t[i]), 0.0, O.out + ((size_t)sys * O.n_max + i) * PBM_NVAR, S_TMP(S));
      PB_SYNC();
    }
    return;
  }
  const double* yp = O.yp + (size_t)sys * O.n_max * PBM_N;
  int j = 0;
  for (int k = 0; k < O.n_obs; k++) {
    double* row = O.out + ((size_t)sys * O.n_obs + k) * PBM_NVAR;
    const double ti = PB_LDU(O.t_obs[k]);
    if (nt < 2) {   // a single stored point: nothing to interpolate between
      for (int q = PB_TID; q < PBM_N; q += PB_TEAM) yy[q] = (nt == 1) ? y[q] : PB_NAN;
    } else {
      // interval search as in TimeSeriesInterpolator::process_within_bounds (t_obs ascending; the last interval also serves
      // every time beyond the final stored point: extrapolate_remaining)
      while (j < nt - 2 && !(PB_LDU(t[j]) <= ti && ti <= PB_LDU(t[j + 1]))) j++;
      const double t0 = PB_LDU(t[j]), hf = PB_LDU(t[j + 1]) - t0;
      const double inv_h = 1.0 / hf, inv_h2 = inv_h * inv_h, inv_h3 = inv_h2 * inv_h;
      const double h = ti - t0, h2 = h * h, h3 = h2 * h;
      for (int q = PB_TID; q < PBM_N; q += PB_TEAM) {
        const double y0 = y[(size_t)j * PBM_N + q], y1 = y[(size_t)(j + 1) * PBM_N + q];
        const double p0 = yp[(size_t)j * PBM_N + q], p1 = yp[(size_t)(j + 1) * PBM_N + q];
        const double c = 3.0 * (y1 - y0) * inv_h2 - (2.0 * p0 + p1) * inv_h;
        const double d = 2.0 * (y0 - y1) * inv_h3 + (p0 + p1) * inv_h2;
        yy[q] = y0 + p0 * h + c * h2 + d * h3;
      }
    }
    PB_SYNC();
    pbm_vars(yy, S_P(S), ti, 0.0, row, S_TMP(S));
    PB_SYNC();
  }
}
#endif

// Constant slots of the LU value array and the packed streams (copied to shared memory once per team). S.sm must be set.
static PB_DEV void pb_sys_tables(pb_sys& S) {
  if (PB_TID == 0) {
    S_LU(S)[PBM_NNZ_LU] = 0.0;      // padding operands
    S_LU(S)[PBM_NNZ_LU + 1] = 1.0;
    S_W(S)[PBM_N] = 0.0;
  }
#if PB_GLOBAL_VEC == 2
  S.slv = pbm_slv_tab;
#else
  unsigned int* slv = (unsigned int*)(S.sm + PB_OFF_SLV);
  for (int i = PB_TID; i < PBM_SLV_ALLOC; i += PB_TEAM) slv[i] = PB_LD(pbm_slv_tab[i]);
  S.slv = slv;
#endif
#if PBM_FAC_SMEM
  unsigned int* fac = (unsigned int*)(S.sm + PB_OFF_FAC);
  for (int i = PB_TID; i < PBM_FAC_WORDS; i += PB_TEAM) fac[i] = PB_LD(pbm_fac_tab[i]);
  S.fac = fac;
#else
  S.fac = pbm_fac_tab;
#endif
  PB_SYNC();
}

static PB_DEV void pb_sys_init(pb_sys& S, double* sm, double* gv, const pbgpu_launch& L) {
  S.sm = sm;
  S.gv = gv;
  S.team.red = sm + PB_OFF_RED;
  S.al_me = nullptr;
  S.al_other = nullptr;
  S.al_count = 0u;
  S.al_seen = 0u;
  S.vb = nullptr;
  S.vb_row0 = nullptr;
  S.vb_n = 0;
  pb_sys_tables(S);
#ifndef PB_EMU
  S.team.red_phase = 0;
#endif
  S.atol = L.atol;
  S.rtol = L.opt.rtol;
  S.epcon = L.opt.nonlinear_convergence_coefficient;
  S.epsNewtIC = L.opt.nonlinear_convergence_coefficient_ic;
  S.step_tol_ic = L.opt.newton_step_tol;
  S.dt_init = L.opt.dt_init;
  S.hmax_inv = (L.opt.dt_max > 0.0) ? 1.0 / L.opt.dt_max : 0.0;
  S.maxord = L.opt.max_order_bdf < 5 ? (L.opt.max_order_bdf < 1 ? 1 : L.opt.max_order_bdf) : 5;
  S.maxnef = L.opt.max_error_test_failures;
  S.maxcor = L.opt.max_nonlinear_iterations;
  S.maxncf = L.opt.max_convergence_failures;
  S.max_iter_ic = L.opt.max_num_iterations_ic;
  S.max_backtracks_ic = L.opt.max_linesearch_backtracks_ic;
  S.scalesol = L.opt.linear_solution_scaling;
  S.max_num_steps = L.opt.max_num_steps;
  S.cj = 0.0;
  S.cjlast = 0.0;
  S.cjold = 1.0;
  S.cjratio = 1.0;
  S.ss = 20.0;
  S.rr = 0.0;
  S.gactive = 0u;
  S.ns = 0;
  S.knew = 0;
  S.phase = 0;
  S.tlo = S.thi = S.trout = S.ttol = 0.0;
}
