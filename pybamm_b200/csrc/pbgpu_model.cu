// pbgpu_model.cu -- translation unit of ONE compiled model image (sm_100a).
//
// nvcc -gencode arch=compute_100a,code=sm_100a -DPBM_HEADER="<generated header>" -DPB_TEAM_SIZE=<32|64|128|256>
//
// Contains the generated model functions (residual, J - cj*M into the static LU slots, events, initial
// conditions, output variables), the sparse-LU tables and the batched BDF integrator specialised for them,
// plus the three host entry points libpbgpu.so binds with dlsym.  Role in the reference: the per-solver
// CasadiFunctions + IDAKLUSolverOpenMP_KLU objects built by create_casadi_solver_group
// (packages/pybammsolvers/src/pybammsolvers/idaklu.cpp:30-124).
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "pbgpu.h"
#include "pbgpu_team.cuh"
#include PBM_HEADER
#include "pbgpu_ida.cuh"

static_assert(PB_TEAM_SIZE >= PBM_LU_TEAM_MIN, "team smaller than the widest record of the SEQ solve stream");

#ifndef PB_MIN_TEAMS
#define PB_MIN_TEAMS 1
#endif
// Kernels with one team per CTA.  In a PB_PAIR image every barrier looks at the control words in front of the team's
// shared memory: they are there (and zero) for these kernels as well.
#if PB_PAIR
#define PB_SM0 (pb_smem + PB_PAIR_CTRL_DOUBLES)
#define PB_SMEM_LAUNCH_BYTES (PB_SMEM_BYTES + 8 * PB_PAIR_CTRL_DOUBLES)
#define PB_ONE_TEAM_PROLOGUE                                                                   \
  if (threadIdx.x < 2 * PB_PAIR_CTRL_DOUBLES) ((unsigned int*)pb_smem)[threadIdx.x] = 0u;      \
  __syncthreads();
#else
#define PB_SM0 pb_smem
#define PB_SMEM_LAUNCH_BYTES PB_SMEM_BYTES
#define PB_ONE_TEAM_PROLOGUE
#endif
static __device__ __forceinline__ void pb_solve_body(const pbgpu_launch& L) {
  extern __shared__ double pb_smem[];
  __shared__ unsigned int next_sys;
  pb_sys S;
  PB_ONE_TEAM_PROLOGUE
#if PB_GLOBAL_VEC
  pb_sys_init(S, PB_SM0, L.scratch + (size_t)blockIdx.x * PB_SCRATCH_DOUBLES, L);
#else
  pb_sys_init(S, PB_SM0, PB_SM0, L);
#if PB_VARS_BATCH > 1
  S.vb = L.scratch + (size_t)blockIdx.x * PB_SCRATCH_DOUBLES;
#endif
#endif
#if PB_ALIGN_TEAMS
  {   // the first two teams to arrive on an SM share its mailbox (persistent kernel: a team never leaves its SM)
    unsigned int smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    unsigned int* box = L.queue + 32 + 8 * (smid & 255u);
    if (threadIdx.x == 0) next_sys = atomicAdd(box, 1u);
    __syncthreads();
    const unsigned int me = next_sys;
    __syncthreads();
    if (me < 2u) {
      S.al_me = box + 2 + 2 * me;
      S.al_other = box + 2 + 2 * (1u - me);
    }
  }
#endif
  for (;;) {
    if (threadIdx.x == 0) next_sys = atomicAdd(L.queue, 1u);
    __syncthreads();
    const unsigned int slot = next_sys;
    __syncthreads();
    if (slot >= (unsigned int)L.n_sys) break;
    pb_align_state(S, 1u);
    pb_solve_system(S, L, L.order ? L.order[slot] : (int)slot);
    pb_align_state(S, 2u);
  }
}

extern "C" __global__ void __launch_bounds__(PB_TEAM_SIZE, PB_MIN_TEAMS) pbgpu_solve_kernel(const pbgpu_launch L) { pb_solve_body(L); }

#if PB_PAIR
// Two teams per CTA (pbgpu_team.cuh, pb_pair_enter): team k of CTA b is global team 2 b + k, with its own slice of the
// dynamic shared memory (after the control words) and of the global scratch, and pulls systems from the same queue.
extern "C" __global__ void __launch_bounds__(2 * PB_TEAM_SIZE, 1) pbgpu_solve_kernel_pair(const pbgpu_launch L) {
  extern __shared__ double pb_smem[];
  __shared__ unsigned int next_sys2[2];
  if (threadIdx.x < 2 * PB_PAIR_CTRL_DOUBLES) ((unsigned int*)pb_smem)[threadIdx.x] = 0u;
  __syncthreads();
  const int team = PB_TEAM_ID;
  pb_sys S;
  pb_sys_init(S, pb_smem + PB_PAIR_CTRL_DOUBLES + (size_t)team * PB_SMEM_DOUBLES,
              L.scratch + ((size_t)blockIdx.x * 2 + team) * PB_SCRATCH_DOUBLES, L);
  for (;;) {
    if (PB_TID == 0) next_sys2[team] = atomicAdd(L.queue, 1u);
    pb_bar_solo();
    const unsigned int slot = next_sys2[team];
    pb_bar_solo();
    if (slot >= (unsigned int)L.n_sys) break;
    pb_solve_system(S, L, L.order ? L.order[slot] : (int)slot);
  }
}
#endif

// The same integration under a tighter register cap (PB_MIN_TEAMS_WIDE teams per SM), for one-warp models whose uncapped
// kernel takes 255 registers: more resident teams win once a batch has more systems than the uncapped kernel keeps
// resident (SPM, 10 000 sets: +25 %), and lose below that (1 000 sets: -9 %, spills without anything to hide them).
// pb_use_wide() picks per launch.
#ifdef PB_MIN_TEAMS_WIDE
extern "C" __global__ void __launch_bounds__(PB_TEAM_SIZE, PB_MIN_TEAMS_WIDE) pbgpu_solve_kernel_wide(const pbgpu_launch L) {
  pb_solve_body(L);
}
#endif

// Evaluate generated functions / the LU for parity tests of the code generator (one team per system).
extern "C" __global__ void __launch_bounds__(PB_TEAM_SIZE, PB_MIN_TEAMS) pbgpu_eval_kernel(const pbgpu_eval_launch E) {
  extern __shared__ double pb_smem[];
  pb_sys S;
  PB_ONE_TEAM_PROLOGUE
  S.sm = PB_SM0;
#if PB_GLOBAL_VEC
  S.gv = E.scratch + (size_t)blockIdx.x * PB_SCRATCH_DOUBLES;
#else
  S.gv = PB_SM0;
#endif
  S.team.red = PB_SM0 + PB_OFF_RED;
  S.team.red_phase = 0;
  pb_sys_tables(S);
  for (int sys = blockIdx.x; sys < E.n_sys; sys += gridDim.x) {
    __syncthreads();
    double* y = S_YY(S);
    double* out = S_DELTA(S);
    for (int i = PB_TID; i < PBM_N; i += PB_TEAM) y[i] = E.y ? E.y[(size_t)sys * PBM_N + i] : 0.0;
    for (int k = PB_TID; k < PBM_NP_REAL; k += PB_TEAM) S_P(S)[k] = E.p[(size_t)sys * PBM_NP_REAL + k];
    for (int k = PB_TID; k < 8; k += PB_TEAM) S_STATS(S)[k] = 0.0;
    __syncthreads();
    double* gout = E.out + (size_t)sys * E.out_len;
    switch (E.which) {
      case 0:
        pbm_res(y, S_P(S), E.t, E.cj, out, S_TMP(S));
        PB_SYNC();
        for (int i = PB_TID; i < PBM_N && i < E.out_len; i += PB_TEAM) gout[i] = out[i];
        break;
      case 1: {
        double* lu = S_LU(S);
        for (int i = PB_TID; i < PBM_NNZ_LU; i += PB_TEAM) lu[i] = 0.0;
        PB_SYNC();
        pbm_jac(y, S_P(S), E.t, E.cj, lu, S_TMPJ(S));
        PB_SYNC();
        for (int k = PB_TID; k < PBM_NNZ && k < E.out_len; k += PB_TEAM) gout[k] = lu[(int)pbm_ent_slot[k]];
      } break;
      case 2:
#if PBM_NEV > 0
        pbm_events(y, S_P(S), E.t, E.cj, S_GHI(S), S_TMP(S));
        PB_SYNC();
        for (int i = PB_TID; i < PBM_NEV && i < E.out_len; i += PB_TEAM) gout[i] = S_GHI(S)[i];
#endif
        break;
      case 3:
        pbm_ic(y, S_P(S), E.t, E.cj, out, S_TMP(S));
        PB_SYNC();
        for (int i = PB_TID; i < PBM_N && i < E.out_len; i += PB_TEAM) gout[i] = out[i];
        break;
      case 4:
#if PBM_NVAR > 0
        pbm_vars(y, S_P(S), E.t, E.cj, gout, S_TMP(S));
#endif
        break;
      case 5: {
        const int bad = pb_lsetup(S, E.t, y, E.cj, 0);
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) out[i] = E.y2[(size_t)sys * PBM_N + i];
        PB_SYNC();
        pb_lu_solve(S, out);
        for (int i = PB_TID; i < PBM_N && i < E.out_len; i += PB_TEAM) gout[i] = bad ? CUDART_NAN : out[i];
      } break;
      case 8:   // events at per-system times: t of system sys is y2[sys][0] (event identification of a whole batch)
#if PBM_NEV > 0
        pbm_events(y, S_P(S), E.y2[(size_t)sys * PBM_N], E.cj, S_GHI(S), S_TMP(S));
        PB_SYNC();
        for (int i = PB_TID; i < PBM_NEV && i < E.out_len; i += PB_TEAM) gout[i] = S_GHI(S)[i];
#endif
        break;
      case 6: {   // clock64() cycles per call of the building blocks, measured in place (tools/gpu_cycles.py)
        const int reps = 16;
        long long c[6];
        pb_lsetup(S, E.t, y, E.cj, 0);
        long long t0 = clock64();
        for (int r = 0; r < reps; r++) {
          pbm_res(y, S_P(S), E.t, 0.0, out, S_TMP(S));
          PB_SYNC();
        }
        c[0] = clock64() - t0;
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) out[i] = E.y2[(size_t)sys * PBM_N + i];
        PB_SYNC();
        t0 = clock64();
        for (int r = 0; r < reps; r++) pb_lu_solve(S, out);
        c[1] = clock64() - t0;
        t0 = clock64();
        for (int r = 0; r < reps; r++) {
          double* lu = S_LU(S);
          for (int i = PB_TID; i < PBM_NFILL; i += PB_TEAM) lu[(int)PB_LD(pbm_fill[i])] = 0.0;
          pbm_jac(y, S_P(S), E.t, E.cj, lu, S_TMPJ(S));
          PB_SYNC();
        }
        c[2] = clock64() - t0;
        t0 = clock64();
        for (int r = 0; r < reps; r++) pb_lu_factor(S);   // refactoring a factor: same instruction stream, values irrelevant
        c[3] = clock64() - t0;
        S.rtol = 1e-8;
        S.atol = E.y2;   // any readable vector of N doubles
        t0 = clock64();
        for (int r = 0; r < reps; r++) pb_ewt_set(S, y);
        c[4] = clock64() - t0;
        double acc = 0.0;
        t0 = clock64();
        for (int r = 0; r < reps; r++) acc += pb_wrms(S, S_PHI(S, r % 6));
        c[5] = clock64() - t0;
        long long c_ev = 0, c_var = 0;
#if PBM_NEV > 0
        t0 = clock64();
        for (int r = 0; r < reps; r++) {
          pbm_events(y, S_P(S), E.t, 0.0, S_GHI(S), S_TMP(S));
          PB_SYNC();
        }
        c_ev = clock64() - t0;
#endif
#if PBM_NVAR > 0
        t0 = clock64();
        for (int r = 0; r < reps; r++) {
          pbm_vars(y, S_P(S), E.t, 0.0, out, S_TMP(S));
          PB_SYNC();
        }
        c_var = clock64() - t0;
#endif
        if (PB_TID == 0) {
          for (int k = 0; k < 6 && k < E.out_len; k++) gout[k] = (double)c[k] / reps;
          if (E.out_len > 7) {
            gout[6] = (double)c_ev / reps + (acc == 12345.678 ? 1.0 : 0.0);
            gout[7] = (double)c_var / reps;
          }
        }
        if (E.out_len >= 8 + 3) {
          pb_lu_solve_t<1>(S, out, gout + 8);   // warm
          pb_lu_solve_t<1>(S, out, gout + 8);
        }
      } break;
      case 7: {   // profiling target: nothing but triangular solves (E.cj = repetitions), for ncu source-level sampling
        pb_lsetup(S, E.t, y, 1.0, 0);
        for (int i = PB_TID; i < PBM_N; i += PB_TEAM) out[i] = E.y2[(size_t)sys * PBM_N + i];
        PB_SYNC();
        const long long t0 = clock64();
        const int reps = (int)E.cj;
        for (int r = 0; r < reps; r++) pb_lu_solve(S, out);
        if (PB_TID == 0 && E.out_len > 0) gout[0] = (double)(clock64() - t0) / reps;
      } break;
      default:
        break;
    }
  }
}

// The experiment schedule (pbgpu_solve_experiment): its own kernel, so that the driver of the plain solve stays as small as
// it is (the integration is bound by instruction fetch).
extern "C" __global__ void __launch_bounds__(PB_TEAM_SIZE, PB_MIN_TEAMS) pbgpu_experiment_kernel(const pbgpu_launch L) {
  extern __shared__ double pb_smem[];
  __shared__ unsigned int next_sys;
  pb_sys S;
  PB_ONE_TEAM_PROLOGUE
#if PB_GLOBAL_VEC
  pb_sys_init(S, PB_SM0, L.scratch + (size_t)blockIdx.x * PB_SCRATCH_DOUBLES, L);
#else
  pb_sys_init(S, PB_SM0, PB_SM0, L);
#if PB_VARS_BATCH > 1
  S.vb = L.scratch + (size_t)blockIdx.x * PB_SCRATCH_DOUBLES;
#endif
#endif
  for (;;) {
    if (threadIdx.x == 0) next_sys = atomicAdd(L.queue, 1u);
    __syncthreads();
    const unsigned int slot = next_sys;
    __syncthreads();
    if (slot >= (unsigned int)L.n_sys) break;
    pb_solve_experiment(S, L, L.order ? L.order[slot] : (int)slot);
  }
}

// Batched observation of the compiled output variables (pbgpu_observe): one team per solution, grid-stride.
extern "C" __global__ void __launch_bounds__(PB_TEAM_SIZE, PB_MIN_TEAMS) pbgpu_observe_kernel(const pbgpu_observe_launch O) {
#if PBM_NVAR > 0
  extern __shared__ double pb_smem[];
  pb_sys S;
  PB_ONE_TEAM_PROLOGUE
  S.sm = PB_SM0;
#if PB_GLOBAL_VEC
  S.gv = O.scratch + (size_t)blockIdx.x * PB_SCRATCH_DOUBLES;
#else
  S.gv = PB_SM0;
#endif
  for (int sys = blockIdx.x; sys < O.n_sys; sys += gridDim.x) pb_observe_system(S, O, sys);
#endif
}

// Launch configuration per device ordinal: the multi-device path runs one host thread per device through the same
// image, so nothing here may be a single file-scope value.
#include <mutex>
struct pb_dev_cfg {
  int ready, blocks_per_sm, num_sms;
  int blocks_per_sm_wide;   // resident teams per SM of pbgpu_solve_kernel_wide, 0 when the image has none
};

static inline bool pb_use_wide(const pb_dev_cfg& c, int n_sys) {
  return c.blocks_per_sm_wide > c.blocks_per_sm && n_sys > c.blocks_per_sm * c.num_sms;
}
static inline int pb_grid_for(const pb_dev_cfg& c, int n_sys) {
  int grid = (pb_use_wide(c, n_sys) ? c.blocks_per_sm_wide : c.blocks_per_sm) * c.num_sms;
  if (n_sys > 0 && grid > n_sys) grid = n_sys;
  return grid < 1 ? 1 : grid;
}
static pb_dev_cfg g_cfg[64];
static std::mutex g_cfg_mutex;

static int pb_configure(pb_dev_cfg* out, char* err, size_t errlen) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    snprintf(err, errlen, "cudaGetDevice: %s", e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range");
    return 1;
  }
  std::lock_guard<std::mutex> lock(g_cfg_mutex);
  if (g_cfg[dev].ready) {
    *out = g_cfg[dev];
    return 0;
  }
  e = cudaFuncSetAttribute(pbgpu_solve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PB_SMEM_LAUNCH_BYTES);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(pbgpu_eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PB_SMEM_LAUNCH_BYTES);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(pbgpu_observe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PB_SMEM_LAUNCH_BYTES);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(pbgpu_experiment_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PB_SMEM_LAUNCH_BYTES);
  if (e != cudaSuccess) {
    snprintf(err, errlen, "model '%s' needs %d bytes of shared memory per team: %s", PBM_NAME, (int)PB_SMEM_BYTES,
             cudaGetErrorString(e));
    return 1;
  }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, dev);
  pb_dev_cfg c;
  c.num_sms = prop.multiProcessorCount;
  // Shared-memory carve-out.  The integration is latency-bound (dependent shared-memory round trips in the
  // triangular solves), so resident teams per SM is what buys throughput: every model takes all the shared memory the
  // SM has (the index tables of the generated functions live in L2 either way).  PBGPU_SMEM_CARVEOUT_KB /
  // PBGPU_TEAMS_PER_SM override.
  const int per_team = (int)PB_SMEM_BYTES + 1024;  // + the per-CTA reservation
  int carve_kb = 227;   // measured on SPMe (22 KB per team): 9 resident teams at 227 KB against 7 at 164 KB, +3.5 %
  if (const char* env = getenv("PBGPU_SMEM_CARVEOUT_KB")) carve_kb = atoi(env);
  if (carve_kb * 1024 < per_team) carve_kb = (per_team + 1023) / 1024;
  if (carve_kb > 227) carve_kb = 227;
  const int total_kb = (int)(prop.sharedMemPerMultiprocessor / 1024);
  e = cudaFuncSetAttribute(pbgpu_solve_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c.blocks_per_sm, pbgpu_solve_kernel, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES);
  if (e != cudaSuccess || c.blocks_per_sm < 1) {
    snprintf(err, errlen, "occupancy query failed for model '%s': %s", PBM_NAME, cudaGetErrorString(e));
    return 1;
  }
  int by_carve = carve_kb * 1024 / per_team;
  if (by_carve < 1) by_carve = 1;
  if (c.blocks_per_sm > by_carve) c.blocks_per_sm = by_carve;
  if (const char* env = getenv("PBGPU_TEAMS_PER_SM")) {
    const int v = atoi(env);
    if (v >= 1 && v < c.blocks_per_sm) c.blocks_per_sm = v;
  }
  // give the rest back to L1: carve out exactly what the resident teams use
  int use_kb = (c.blocks_per_sm * per_team + 1023) / 1024;
  if (use_kb > carve_kb) use_kb = carve_kb;
  int pct = (use_kb * 100 + total_kb - 1) / total_kb;
  if (pct > 100) pct = 100;
  cudaFuncSetAttribute(pbgpu_solve_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  cudaFuncSetAttribute(pbgpu_eval_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  cudaFuncSetAttribute(pbgpu_observe_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
  cudaFuncSetAttribute(pbgpu_experiment_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct);
#if PB_PAIR
  {
    const int pair_bytes = 2 * (int)PB_SMEM_BYTES + 8 * PB_PAIR_CTRL_DOUBLES;
    e = cudaFuncSetAttribute(pbgpu_solve_kernel_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, pair_bytes);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(pbgpu_solve_kernel_pair, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    int bp = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bp, pbgpu_solve_kernel_pair, 2 * PB_TEAM_SIZE, pair_bytes);
    if (e != cudaSuccess || bp < 1) {
      snprintf(err, errlen, "pair kernel of model '%s' does not fit an SM: %s", PBM_NAME, cudaGetErrorString(e));
      return 1;
    }
  }
#endif
  c.blocks_per_sm_wide = 0;
#ifdef PB_MIN_TEAMS_WIDE
  if (!getenv("PBGPU_NO_WIDE")) {
    e = cudaFuncSetAttribute(pbgpu_solve_kernel_wide, cudaFuncAttributeMaxDynamicSharedMemorySize, PB_SMEM_LAUNCH_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(pbgpu_solve_kernel_wide, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    int bw = 0;
    if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bw, pbgpu_solve_kernel_wide, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES);
    if (e == cudaSuccess && bw > by_carve) bw = by_carve;
    if (e == cudaSuccess && bw > c.blocks_per_sm && !getenv("PBGPU_TEAMS_PER_SM")) {
      c.blocks_per_sm_wide = bw;
      int wkb = (bw * per_team + 1023) / 1024;
      if (wkb > carve_kb) wkb = carve_kb;
      int wpct = (wkb * 100 + total_kb - 1) / total_kb;
      cudaFuncSetAttribute(pbgpu_solve_kernel_wide, cudaFuncAttributePreferredSharedMemoryCarveout, wpct > 100 ? 100 : wpct);
    }
    (void)cudaGetLastError();
  }
#endif
  c.ready = 1;
  g_cfg[dev] = c;
  *out = c;
  return 0;
}

extern "C" int pbgpu_image_describe(pbgpu_model_desc* d) {
  memset(d, 0, sizeof(*d));
  d->abi_version = PBGPU_ABI_VERSION;
  d->n_states = PBM_N;
  d->n_diff = PBM_NDIFF;
  d->n_inputs = PBM_NP_REAL;
  d->n_events = PBM_NEV;
  d->nnz_jac = PBM_NNZ;
  d->nnz_lu = PBM_NNZ_LU;
  d->n_vars = PBM_NVAR;
  d->team_size = PB_TEAM_SIZE;
  d->smem_bytes = PB_SMEM_BYTES;
  d->n_tmp = PBM_NTMP;
  d->lu_levels[0] = PBM_FAC_NLEV;
  d->lu_levels[1] = PBM_FWD_NLEV;
  d->lu_levels[2] = PBM_BWD_NLEV;
  d->flops[0] = PBM_FLOPS_RES;
  d->flops[1] = PBM_FLOPS_JAC;
  d->flops[2] = PBM_FLOPS_FACTOR;
  d->flops[3] = PBM_FLOPS_SOLVE;
  d->flops[4] = PBM_FLOPS_EVENTS;
  d->flops[5] = PBM_FLOPS_VARS;
  d->scratch_doubles_per_team = PB_SCRATCH_DOUBLES;
  strncpy(d->name, PBM_NAME, sizeof(d->name) - 1);
  return 0;
}

// Teams a launch for n_sys systems will use on the current device (the solver handle sizes its scratch block with it
// and passes the same number back to pbgpu_image_launch / pbgpu_image_eval).
extern "C" int pbgpu_image_plan(int n_sys, int* teams, char* err, size_t errlen) {
  pb_dev_cfg c;
  if (pb_configure(&c, err, errlen)) return 1;
  *teams = pb_grid_for(c, n_sys);
#if PB_PAIR
  *teams = (*teams + 1) / 2 * 2;     // whole CTAs of two teams (the scratch block has a slot per team)
#endif
  return 0;
}

// All pointers in L are device pointers on the current device.  grid: the number of teams pbgpu_image_plan returned
// for this n_sys (the caller's scratch block holds exactly that many team slots).
extern "C" int pbgpu_image_launch(const pbgpu_launch* L, void* stream, int grid, char* err, size_t errlen) {
  pb_dev_cfg c;
  if (pb_configure(&c, err, errlen)) return 1;
  const bool wide = pb_use_wide(c, L->n_sys) && L->sched.n_steps == 0;
  if (grid < 1 || grid > (wide ? c.blocks_per_sm_wide : c.blocks_per_sm) * c.num_sms) {
    snprintf(err, errlen, "pbgpu_image_launch: grid %d was not planned for this device", grid);
    return 1;
  }
#if PB_PAIR
  static const bool no_pair = getenv("PBGPU_NO_PAIR") != nullptr;
  if (L->sched.n_steps == 0 && !no_pair && grid % 2 == 0) {
    pbgpu_solve_kernel_pair<<<grid / 2, 2 * PB_TEAM_SIZE, 2 * PB_SMEM_BYTES + 8 * PB_PAIR_CTRL_DOUBLES, (cudaStream_t)stream>>>(*L);
  } else
#endif
  if (L->sched.n_steps > 0) {
    pbgpu_experiment_kernel<<<grid, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES, (cudaStream_t)stream>>>(*L);
  } else if (wide) {
#ifdef PB_MIN_TEAMS_WIDE
    pbgpu_solve_kernel_wide<<<grid, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES, (cudaStream_t)stream>>>(*L);
#endif
  } else {
    pbgpu_solve_kernel<<<grid, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES, (cudaStream_t)stream>>>(*L);
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(err, errlen, "pbgpu_solve_kernel launch: %s", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

// ---- launch without any allocation (pbgpu_solve_device_noalloc) -------------------------------------------------
// The work queue and the per-team scratch of the solve kernel in a file-scope device array instead of the solver
// handle's cudaMalloc'ed blocks: for callers that may not allocate (CUDA graph capture, serving loops).  Every device
// that loads the image holds the array until the process exits; every call on a device shares it, so calls that overlap
// on the GPU race.  Layout: [queue words, padded to 256 bytes][PB_STATIC_TEAMS team slots of PB_SCRATCH_DOUBLES].
#define PB_STATIC_QUEUE_DOUBLES (((PBGPU_QUEUE_WORDS * 4 + 255) / 256) * 32)
#if PB_SCRATCH_DOUBLES > 0
#define PB_STATIC_FIT ((int)((PBGPU_STATIC_MAX_BYTES / 8 - PB_STATIC_QUEUE_DOUBLES) / PB_SCRATCH_DOUBLES))
#define PB_STATIC_TEAMS (PB_STATIC_FIT < PBGPU_STATIC_MAX_TEAMS ? PB_STATIC_FIT : PBGPU_STATIC_MAX_TEAMS)
#else
#define PB_STATIC_TEAMS PBGPU_STATIC_MAX_TEAMS   /* no team scratch: only the queue lives in the array */
#endif
#define PB_STATIC_DOUBLES (PB_STATIC_QUEUE_DOUBLES + (size_t)(PB_STATIC_TEAMS > 0 ? PB_STATIC_TEAMS : 0) * PB_SCRATCH_DOUBLES)
static_assert(PB_STATIC_DOUBLES * 8 <= PBGPU_STATIC_MAX_BYTES, "static scratch larger than its stated maximum");
static __device__ __align__(256) double pb_static_scratch[PB_STATIC_DOUBLES];

// In host code pb_static_scratch is only a handle for the runtime, not a device address: the address is asked for once
// per device, with that device current, and cached.
static double* g_static_ptr[64];
static std::mutex g_static_mutex;

static int pb_static_setup(double** ptr, char* err, size_t errlen) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= 64) {
    snprintf(err, errlen, "cudaGetDevice: %s", e != cudaSuccess ? cudaGetErrorString(e) : "ordinal out of range");
    return 1;
  }
  std::lock_guard<std::mutex> lock(g_static_mutex);
  if (!g_static_ptr[dev]) {
    double* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, pb_static_scratch);
    if (e != cudaSuccess || !p) {
      snprintf(err, errlen, "cudaGetSymbolAddress(pb_static_scratch) on device %d: %s", dev, cudaGetErrorString(e));
      return 1;
    }
    g_static_ptr[dev] = p;
  }
  *ptr = g_static_ptr[dev];
  return 0;
}

// L: as for pbgpu_image_launch, its queue and scratch fields are ignored.  L->n_sys == 0: set-up only (kernel
// attributes, the array's address on the current device).  *grid_out: teams launched.
extern "C" int pbgpu_image_launch_static(const pbgpu_launch* L, void* stream, int* grid_out, char* err, size_t errlen) {
  pb_dev_cfg c;
  if (pb_configure(&c, err, errlen)) return 1;
  double* base = nullptr;
  if (pb_static_setup(&base, err, errlen)) return 1;
  if (grid_out) *grid_out = 0;
  if (L->n_sys == 0) return 0;
  if (L->n_sys < 0 || L->n_sys > PBGPU_STATIC_MAX_SYSTEMS) {
    snprintf(err, errlen, "pbgpu_solve_device_noalloc: n_sys = %d is beyond the static limit of %d systems per call", L->n_sys,
             (int)PBGPU_STATIC_MAX_SYSTEMS);
    return 1;
  }
  if (PB_STATIC_TEAMS < 1) {
    snprintf(err, errlen, "model '%s' needs %lld bytes of scratch per team: more than the static array of %lld bytes holds",
             PBM_NAME, (long long)PB_SCRATCH_DOUBLES * 8, (long long)PBGPU_STATIC_MAX_BYTES);
    return 1;
  }
  if (L->sched.n_steps > 0) {
    snprintf(err, errlen, "pbgpu_solve_device_noalloc: experiment schedules are not supported on this entry point");
    return 1;
  }
  int grid = c.blocks_per_sm * c.num_sms;
  if (grid > L->n_sys) grid = L->n_sys;
  if (PB_SCRATCH_DOUBLES > 0 && grid > PB_STATIC_TEAMS) grid = PB_STATIC_TEAMS;
  pbgpu_launch K = *L;
  K.queue = (unsigned int*)base;
  K.scratch = base + PB_STATIC_QUEUE_DOUBLES;
  // the queue must start at zero; the team slots keep the last call's contents, which no team reads before writing
  cudaError_t e = cudaMemsetAsync(base, 0, sizeof(double) * PB_STATIC_QUEUE_DOUBLES, (cudaStream_t)stream);
  if (e == cudaSuccess) {
    pbgpu_solve_kernel<<<grid, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES, (cudaStream_t)stream>>>(K);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    snprintf(err, errlen, "pbgpu_solve_kernel launch (static scratch): %s", cudaGetErrorString(e));
    return 1;
  }
  if (grid_out) *grid_out = grid;
  return 0;
}

extern "C" int pbgpu_image_eval(const pbgpu_eval_launch* E, void* stream, int grid, char* err, size_t errlen) {
  pb_dev_cfg c;
  if (pb_configure(&c, err, errlen)) return 1;
  if (grid > c.blocks_per_sm * c.num_sms && grid <= c.blocks_per_sm_wide * c.num_sms)
    grid = c.blocks_per_sm * c.num_sms;   // planned for the wide solve kernel: this kernel is grid-stride, fewer teams do
  if (grid < 1 || grid > c.blocks_per_sm * c.num_sms) {
    snprintf(err, errlen, "pbgpu_image_eval: grid %d was not planned for this device", grid);
    return 1;
  }
  pbgpu_eval_kernel<<<grid, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES, (cudaStream_t)stream>>>(*E);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(err, errlen, "pbgpu_eval_kernel launch: %s", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}

extern "C" int pbgpu_image_observe(const pbgpu_observe_launch* O, void* stream, int grid, char* err, size_t errlen) {
  pb_dev_cfg c;
  if (pb_configure(&c, err, errlen)) return 1;
  if (PBM_NVAR < 1) {
    snprintf(err, errlen, "the model image has no compiled output variables");
    return 1;
  }
  if (grid > c.blocks_per_sm * c.num_sms && grid <= c.blocks_per_sm_wide * c.num_sms)
    grid = c.blocks_per_sm * c.num_sms;   // planned for the wide solve kernel: this kernel is grid-stride, fewer teams do
  if (grid < 1 || grid > c.blocks_per_sm * c.num_sms) {
    snprintf(err, errlen, "pbgpu_image_observe: grid %d was not planned for this device", grid);
    return 1;
  }
  pbgpu_observe_kernel<<<grid, PB_TEAM_SIZE, PB_SMEM_LAUNCH_BYTES, (cudaStream_t)stream>>>(*O);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    snprintf(err, errlen, "pbgpu_observe_kernel launch: %s", cudaGetErrorString(e));
    return 1;
  }
  return 0;
}
