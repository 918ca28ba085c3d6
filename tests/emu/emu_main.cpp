// tests/emu/emu_main.cpp -- HOST EMULATION of the device integrator, for tests only.
//
// Compiles the very same sources the GPU runs (pbgpu_ida.cuh + the CUDA-dialect generated model
// header) with PB_EMU: a team of one "thread", barriers are no-ops.  It lets the CPU-only test tier
// check the integrator's control flow, the generated LU tables and the driver loop against the
// independent C oracle before any GPU time is spent.  It cannot find missing barriers or races
// (those are checked on the GPU with compute-sanitizer) and it is NOT part of the product: nothing
// in pybamm_b200/ or libpbgpu loads it.
#define PB_EMU 1
#include <cstdlib>
#include <cstring>

#include "pbgpu.h"
#include "pbgpu_team.cuh"
#include PBM_HEADER
#include "pbgpu_ida.cuh"

extern "C" int emu_solve(const pbgpu_launch* L) {
  double* sm = (double*)calloc(PB_SMEM_DOUBLES + PB_SCRATCH_DOUBLES + 1, sizeof(double));
  pb_sys S;
  pb_sys_init(S, sm, PB_GLOBAL_VEC ? sm + PB_SMEM_DOUBLES : sm, *L);
  for (int sys = 0; sys < L->n_sys; sys++) {
    if (L->sched.n_steps > 0) pb_solve_experiment(S, *L, sys);
    else pb_solve_system(S, *L, sys);
  }
  free(sm);
  return 0;
}

extern "C" void emu_sizes(int* out) {
  out[0] = PBM_N; out[1] = PBM_NDIFF; out[2] = PBM_NP_REAL; out[3] = PBM_NEV; out[4] = PBM_NNZ; out[5] = PBM_NVAR;
  out[6] = PBM_NNZ_LU; out[7] = (int)PB_SMEM_BYTES;
}

// which: 0 res, 1 jac (CSC order), 2 events, 3 ic, 4 vars, 5 (J - cj M) x = y2
extern "C" int emu_eval(int which, const double* y, const double* p, double t, double cj, const double* y2, double* out) {
  double* sm = (double*)calloc(PB_SMEM_DOUBLES + PB_SCRATCH_DOUBLES + 1, sizeof(double));
  pb_sys S;
  S.sm = sm;
  S.gv = PB_GLOBAL_VEC ? sm + PB_SMEM_DOUBLES : sm;
  S.team.red = sm + PB_OFF_RED;
  pb_sys_tables(S);
  if (y) memcpy(S_YY(S), y, sizeof(double) * PBM_N);
  memcpy(S_P(S), p, sizeof(double) * PBM_NP_REAL);
  int rc = 0;
  switch (which) {
    case 0: pbm_res(S_YY(S), S_P(S), t, cj, out, S_TMP(S)); break;
    case 1:
      memset(S_LU(S), 0, sizeof(double) * PBM_NNZ_LU);
      pbm_jac(S_YY(S), S_P(S), t, cj, S_LU(S), S_TMPJ(S));
      for (int k = 0; k < PBM_NNZ; k++) out[k] = S_LU(S)[pbm_ent_slot[k]];
      break;
    case 2:
#if PBM_NEV > 0
      pbm_events(S_YY(S), S_P(S), t, cj, out, S_TMP(S));
#endif
      break;
    case 3: pbm_ic(S_YY(S), S_P(S), t, cj, out, S_TMP(S)); break;
    case 4:
#if PBM_NVAR > 0
      pbm_vars(S_YY(S), S_P(S), t, cj, out, S_TMP(S));
#endif
      break;
    case 5:
      rc = pb_lsetup(S, t, S_YY(S), cj, 0);
      memcpy(S_DELTA(S), y2, sizeof(double) * PBM_N);
      pb_lu_solve(S, S_DELTA(S));
      memcpy(out, S_DELTA(S), sizeof(double) * PBM_N);
      break;
  }
  free(sm);
  return rc;
}

extern "C" int emu_observe(const pbgpu_observe_launch* O) {
#if PBM_NVAR > 0
  double* sm = (double*)calloc(PB_SMEM_DOUBLES + PB_SCRATCH_DOUBLES + 1, sizeof(double));
  pb_sys S;
  S.sm = sm;
  S.gv = PB_GLOBAL_VEC ? sm + PB_SMEM_DOUBLES : sm;
  for (int sys = 0; sys < O->n_sys; sys++) pb_observe_system(S, *O, sys);
  free(sm);
  return 0;
#else
  return 1;
#endif
}
