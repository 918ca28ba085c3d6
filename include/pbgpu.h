/*
 * pbgpu.h -- C ABI of the B200-native batched DAE solver behind pybamm.IDAKLUSolver.
 *
 * Drop-in boundary.  These entry points replace, for the batched
 * solve path, what the reference binds through pybind11:
 *
 *   pbgpu_model_load / pbgpu_solver_create   <->  idaklu.create_casadi_solver_group(...)
 *        packages/pybammsolvers/src/pybammsolvers/idaklu.cpp:30-124,149-175
 *        (CasADi Function arguments cannot cross a C ABI: a *compiled model image* --
 *         residual, J - cj*M, events, initial conditions, output variables and the static
 *         sparse-LU tables lowered to sm_100a by pybamm_b200.codegen -- takes their place)
 *   pbgpu_solve                              <->  IDAKLUSolverGroup.solve(t_eval, t_interp, y0, yp0, inputs)
 *        packages/pybammsolvers/src/pybammsolvers/idaklu.cpp:138-147,
 *        idaklu_source/IDAKLUSolverGroup.cpp:6-170
 *   pbgpu_result                             <->  struct solution {flag, t, y, yp, yS, ypS, y_term}
 *        idaklu_source/Solution.hpp:9-39, SolutionData.hpp:75-86
 *   pbgpu_options                            <->  SolverOptions (idaklu_source/Options.cpp:9-206,
 *        defaults packages/pybamm/src/pybamm/solvers/idaklu_solver.py:286-325)
 *   per-system flag values                   <->  IDA return codes (idaklu_source/sundials_error_handler.hpp:11-61):
 *        0 success, 1 tstop (final t_eval reached), 2 root (termination event), <0 failure
 *
 * All arrays are C-contiguous float64 / int32 / int64 exactly as the pybind11 side requires
 * (np_array is c_style|forcecast, idaklu_source/common.hpp:35).  Output buffers are caller-owned
 * (no ownership crosses the ABI).  Every function returns 0 on success, non-zero on error with a
 * message available from pbgpu_last_error() (argument/shape errors mirror the std::domain_error /
 * invalid_argument texts of IDAKLUSolverGroup.cpp:46-97).  Integration failures are NOT errors:
 * they are reported per system in flag[] with the partial solution, as the reference does
 * (IDAKLUSolverOpenMP.inl:401-406,443).
 *
 * There is no CPU fallback: pbgpu_model_load fails when no CUDA device is usable.
 */
#ifndef PBGPU_H
#define PBGPU_H

#ifdef __cplusplus
extern "C" {
#endif

#define PBGPU_QUEUE_WORDS (32 + 8 * 256)
#define PBGPU_ABI_VERSION 6

typedef struct pbgpu_model_s* pbgpu_model_t;
typedef struct pbgpu_solver_s* pbgpu_solver_t;

typedef struct {
  int abi_version;
  int n_states;       /* number_of_states */
  int n_diff;         /* len_rhs: states [0, n_diff) are differential, the rest algebraic */
  int n_inputs;       /* number_of_parameters (flattened inputs) */
  int n_events;       /* number_of_events (TERMINATION events) */
  int nnz_jac;        /* structural nnz of J - cj*M */
  int nnz_lu;         /* nnz of the static L+U pattern */
  int n_vars;         /* total length of the compiled output variables */
  int team_size;      /* threads cooperating on one system */
  int smem_bytes;     /* dynamic shared memory per team */
  int n_tmp;          /* scratch temporaries of the generated functions */
  int lu_levels[3];   /* dependency levels: factor, forward, backward */
  long long flops[6]; /* static work model: res, jac, factor, solve, events, vars (DFMA-equivalents / flops) */
  long long scratch_doubles_per_team; /* global scratch a team needs (streamed vectors of large models), 0 if none */
  char name[64];
} pbgpu_model_desc;

/* Solver options: same names and defaults as IDAKLUSolver's options dict (idaklu_solver.py:286-325). */
typedef struct {
  double rtol;
  int max_order_bdf;                            /* 5 */
  long long max_num_steps;                      /* 100000 */
  int max_error_test_failures;                  /* 10 */
  int max_nonlinear_iterations;                 /* 40 */
  int max_convergence_failures;                 /* 100 */
  double nonlinear_convergence_coefficient;     /* 0.33 */
  double nonlinear_convergence_coefficient_ic;  /* 0.0033 */
  double newton_step_tol;                       /* 1e-4 */
  int max_num_iterations_ic;                    /* 100 */
  int max_linesearch_backtracks_ic;             /* 5 */
  int linear_solution_scaling;                  /* 1 */
  int calc_ic;                                  /* 1 */
  double dt_init;                               /* 0 -> IDA heuristic */
  double dt_max;                                /* 0 -> unbounded */
  int num_steps_no_progress;                    /* 0 (off); window of the NoProgressGuard (NoProgressGuard.hpp:12-66), <= 32 */
  double t_no_progress;                         /* 0.0 (off); the window's steps must advance t by at least this much */
} pbgpu_options;

/* Caller-owned output block of one batch (host pointers for pbgpu_solve, device for pbgpu_solve_device). */
typedef struct {
  int n_max;            /* capacity (time points) per system */
  int n_ret;            /* row length: n_states (output_mode 0) or n_vars (output_mode 1) */
  double* t;            /* [n_sys][n_max]  */
  double* y;            /* [n_sys][n_max][n_ret], time-major rows like solution.y */
  int* n_t;             /* [n_sys] number of time points the system produced; only the first n_max are stored, so
                           n_t > n_max means the block was too small (adaptive-step mode): re-solve with n_max >= n_t */
  int* flag;            /* [n_sys] IDA-style return code */
  double* y_term;       /* [n_sys][n_states] final state (y_event), may be NULL */
  long long* stats;     /* [n_sys][8]: nsteps nrevals nlinsetups netfails nniters nncfails njevals nge; may be NULL */
  double* yp;           /* [n_sys][n_max][n_states] dy/dt at the stored points (solution.yp, Solution.hpp:9-39: the Hermite
                           data of IDAKLUSolverOpenMP.inl:447-457), output_mode 0 only; may be NULL */
  int* event_idx;       /* [n_sys] index of the termination event closest to zero at the final state when flag == 2
                           (closest_event_idx, idaklu_solver.py:826-834), else -1; may be NULL */
  double* t_event;      /* [n_sys] time of the last stored point (t_event of idaklu_solver.py:814); may be NULL */
  float kernel_ms;      /* out: device time of the integration kernel (CUDA events) */
  float total_ms;       /* out: device time including copies (pbgpu_solve only) */
} pbgpu_result;

#define PBGPU_OUTPUT_STATES 0
#define PBGPU_OUTPUT_VARS 1

void pbgpu_default_options(pbgpu_options* opt);

/* Load a compiled model image (pbgpu_model_<name>.so built by pybamm_b200.build.build_model). */
int pbgpu_model_load(const char* path, pbgpu_model_t* model);
int pbgpu_model_describe(pbgpu_model_t model, pbgpu_model_desc* desc);
int pbgpu_model_unload(pbgpu_model_t model);

/* atol: [n_states].  device: CUDA ordinal.  One solver <-> one model, as in the reference
 * (base_solver.py:970-977). */
int pbgpu_solver_create(pbgpu_model_t model, const double* atol, const pbgpu_options* opt, int device,
                        pbgpu_solver_t* solver);
int pbgpu_solver_destroy(pbgpu_solver_t solver);

/* Launch order of the next solves with n systems (host array of n distinct row indices, or NULL / n = 0 to clear):
 * teams pull systems from a queue, and a batch ends when its slowest system does, so the expensive rows should be
 * handed out first.  Results stay in row order.  The reference has no equivalent: it splits the batch in static
 * contiguous chunks (IDAKLUSolverGroup.cpp:99-126). */
int pbgpu_solver_set_launch_order(pbgpu_solver_t solver, const int* order, int n);

/* Number of time points a result block must hold per system for these grids. */
int pbgpu_result_capacity(int n_teval, int n_tinterp);

/*
 * Integrate n_sys independent systems.  Host buffers in, host buffers out (H2D, kernel, D2H inside).
 *   t_eval   [n_teval]   sorted, unique, >= 2 entries (breakpoints; IDA stop times)
 *   t_interp [n_tinterp] sorted, unique, t_eval values removed (may be empty: n_tinterp = 0 -> adaptive steps stored)
 *   y0       [n_sys][n_states] or NULL (-> the model's initial conditions y0(p) evaluated on device)
 *   yp0      [n_sys][n_states] or NULL: ydot0full of IDAKLUSolverGroup.solve.  Used as given only when calc_ic is off
 *            (SetupInitialState, IDAKLUSolverOpenMP.inl:514-563); with calc_ic the derivatives are recomputed from the
 *            consistent state (RecoverYp, :868-873), so NULL is the normal call
 *   inputs   [n_sys][n_inputs]
 */
int pbgpu_solve(pbgpu_solver_t solver, int n_sys, const double* t_eval, int n_teval, const double* t_interp,
                int n_tinterp, const double* y0, const double* yp0, const double* inputs, int output_mode,
                pbgpu_result* out);

/* Same, but every pointer (inputs and the result block's arrays) is device memory on the solver's
 * device; nothing is copied.  stream: a cudaStream_t (NULL = default stream).  Asynchronous unless
 * sync != 0; kernel_ms is only filled when sync != 0. */
int pbgpu_solve_device(pbgpu_solver_t solver, int n_sys, const double* t_eval, int n_teval, const double* t_interp,
                       int n_tinterp, const double* y0, const double* yp0, const double* inputs, int output_mode,
                       pbgpu_result* out, void* stream, int sync);

/* pbgpu_solve_device without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it).  Same arguments and results as pbgpu_solve_device, but the work queue and the per-team scratch vectors of
 * the solve kernel live in a file-scope `static __device__` array of the model image instead of the solver handle's
 * cudaMalloc'ed blocks, and the call makes no cudaMalloc / cudaFree / allocating library call.
 *   limits   n_sys <= PBGPU_STATIC_MAX_SYSTEMS per call; the array holds at most PBGPU_STATIC_MAX_TEAMS team slots and
 *            PBGPU_STATIC_MAX_BYTES bytes, so a large model runs on fewer teams than pbgpu_solve_device would use (same
 *            results, lower throughput).  A larger n_sys is rejected with an error, nothing is launched.
 *   memory   every GPU that loads the model image holds the array (<= PBGPU_STATIC_MAX_BYTES) until the process exits.
 *   sharing  every call on a device uses the same array: the queue words are cleared on `stream` before the launch, the
 *            team slots keep the previous call's contents (never read before written).  Calls that overlap on the GPU
 *            (different streams, different solver handles of the same image) race: serialise them.
 *   set-up   n_sys == 0 launches nothing and only performs the one-time set-up for the solver's device (kernel attributes,
 *            address of the array): call it once before a region that must not touch the runtime's module loader.
 * The launch order of pbgpu_solver_set_launch_order (stored in the handle beforehand) is honoured; experiment schedules are
 * not available here. */
#define PBGPU_STATIC_MAX_SYSTEMS 1024
#define PBGPU_STATIC_MAX_TEAMS 64
#define PBGPU_STATIC_MAX_BYTES (8u << 20)
int pbgpu_solve_device_noalloc(pbgpu_solver_t solver, int n_sys, const double* t_eval, int n_teval, const double* t_interp,
                               int n_tinterp, const double* y0, const double* yp0, const double* inputs, int output_mode,
                               pbgpu_result* out, void* stream, int sync);

/* Evaluate generated model functions / the sparse LU on the device for n_sys (y, p) pairs (parity tests
 * of the code generator).  which: 0 res, 1 jac (bundle CSC order), 2 events, 3 ic, 4 vars,
 * 5 solve (J - cj*M) x = rhs with rhs = y2; 8 events at per-system times t_sys = y2[sys][0] (which event ended each
 * system of a batch: idaklu_solver.py:826-834); 6 / 7 cycle probes of the building blocks (tools/gpu_cycles.py,
 * tools/gpu_solve_loop.py).  Host pointers. */
int pbgpu_model_eval(pbgpu_solver_t solver, int which, int n_sys, const double* y, const double* p, double t,
                     double cj, const double* y2, double* out, int out_len);

/*
 * Experiment schedule (BASELINE config 5): the per-step loop of Simulation.solve
 * (packages/pybamm/src/pybamm/simulation/simulation.py:991-1167) for the *unified* experiment model (:269-416), whose
 * control law and step termination are selected by input parameters ("Experiment step index", "Experiment step value",
 * "Ambient temperature [K]").  The reference runs it as one Python solver.step() per step and per input set and refuses a
 * list of input sets (:832-835); here the whole schedule runs inside the kernel for every system of the batch.
 * Per step k (base_solver.py:1334-1588, BaseStep.setup_timestepping base_step.py:355-421):
 *   the over_slot inputs take over_val[k][:]; t_start = end of the previous step (moved to the next floating-point value when
 *   non-zero); consistent initialisation; termination events at t_start: the experiment event <= 0 skips the step
 *   ("Event exceeded in initial conditions" -> EmptySolution, simulation.py:1062-1070), any other one ends the run; integrate to
 *   t_start + duration[k] storing t_start, t_start + j * duration[k] / (n_pts[k] - 1) (j = 1 .. n_pts[k] - 2) and the end point;
 *   a root of the experiment event ends the step, a root of any other termination event ends the run (infeasible experiment),
 *   reaching the duration of a step with stop_on_final_time[k] set ends the run as well (default durations, :1140-1147).
 * All pointers are host memory for pbgpu_solve_experiment.
 */
typedef struct {
  int n_steps;
  int n_over;                            /* inputs overridden by the schedule */
  const int* over_slot;                  /* [n_over] columns of `inputs` */
  const double* over_val;                /* [n_steps][n_over] */
  const double* duration;                /* [n_steps] seconds */
  const int* n_pts;                      /* [n_steps] output points of the step's linspace(0, duration, n_pts), >= 2 */
  const unsigned char* stop_on_final_time; /* [n_steps] or NULL */
  int experiment_event;                  /* index of the termination event that ends a step normally, -1: none */
} pbgpu_schedule;

/* step_term codes */
#define PBGPU_STEP_NOT_RUN 0
#define PBGPU_STEP_FINAL_TIME 1
#define PBGPU_STEP_EXPERIMENT_EVENT 2
#define PBGPU_STEP_SKIPPED 3
#define PBGPU_STEP_OTHER_EVENT 4

/* Points a result block must hold per system for a schedule: sum over steps of n_pts[k] + 1. */
int pbgpu_schedule_capacity(const pbgpu_schedule* sched);

/*
 * Run the schedule for n_sys input sets starting at time t0.  out as in pbgpu_solve (t / y concatenated over the steps; flag:
 * 1 all steps done, 2 ended on a termination event other than the experiment's, < 0 IDA failure); step_end [n_sys][n_steps]:
 * number of stored points up to and including step k; step_term [n_sys][n_steps]: PBGPU_STEP_* or a negative IDA flag.
 */
int pbgpu_solve_experiment(pbgpu_solver_t solver, int n_sys, double t0, const pbgpu_schedule* sched, const double* y0,
                           const double* inputs, int output_mode, pbgpu_result* out, int* step_end, int* step_term);

/*
 * Observe the compiled output variables of a batch of solutions on the device: the batched counterpart of
 * idaklu.observe / idaklu.observe_hermite_interp (idaklu_source/observe.cpp:319-354), which the reference calls once per
 * variable and per solution from ProcessedVariable (processed_variable.py:90-297) through CasADi.
 *   t [n_sys][n_max], y [n_sys][n_max][n_states], n_t [n_sys]: a result block of pbgpu_solve (output_mode 0)
 *   yp [n_sys][n_max][n_states] or NULL, t_obs [n_obs] or NULL:
 *     t_obs == NULL: out[s][i][:] = vars(t[s][i], y[s][i], p[s]) at the stored points i < n_t[s]   (observe; out rows: n_max)
 *     t_obs != NULL: cubic Hermite interpolation of y between the stored points from (y, yp), then vars(t_obs[k], y(t_obs[k]), p[s])
 *                    (observe_hermite_interp: interval j with t[j] <= t_obs[k] <= t[j+1], the last interval extrapolates; out
 *                    rows: n_obs); yp is required
 *   out [n_sys][n_rows][n_vars]
 * pbgpu_observe takes host pointers, pbgpu_observe_device device pointers on the solver's device (no copies; stream as in
 * pbgpu_solve_device).
 */
int pbgpu_observe(pbgpu_solver_t solver, int n_sys, int n_max, const int* n_t, const double* t, const double* y,
                  const double* yp, const double* inputs, const double* t_obs, int n_obs, double* out);
int pbgpu_observe_device(pbgpu_solver_t solver, int n_sys, int n_max, const int* n_t, const double* t, const double* y,
                         const double* yp, const double* inputs, const double* t_obs, int n_obs, double* out, void* stream,
                         int sync);

const char* pbgpu_last_error(void);

/* ---- interface between libpbgpu and a compiled model image (not for end users) ---- */
typedef struct {
  int n_sys, n_eval, n_interp, n_max, save_vars, n_ret;
  const double* t_eval;
  const double* t_interp;
  const double* inputs;
  const double* y0;
  const double* yp0;
  const double* atol;
  double* t_out;
  double* y_out;
  int* n_out;
  int* flag;
  double* y_term;
  long long* stats;
  double* yp_out;
  int* event_idx;
  double* t_event;
  unsigned int* queue;   /* [PBGPU_QUEUE_WORDS], zeroed before every launch: word 0 = next system to hand out; from word
                            32 on, 8 words per SM: the mailbox through which the two teams resident on one SM keep their
                            Newton iterations in phase (pbgpu_ida.cuh, pb_align) */
  double* scratch;      /* [teams][scratch_doubles_per_team] device memory owned by the solver handle, may be NULL if 0 */
  const int* order;     /* [n_sys] or NULL: the k-th system handed out by the queue is order[k] */
  pbgpu_options opt;
  /* experiment schedule (device pointers), sched.n_steps == 0: plain solve */
  pbgpu_schedule sched;
  double t_start;       /* start time of the schedule */
  int* step_end;
  int* step_term;
} pbgpu_launch;

typedef struct {
  int n_sys, n_max, n_obs;   /* n_obs == 0: at the stored points */
  const int* n_t;
  const double* t;
  const double* y;
  const double* yp;
  const double* inputs;
  const double* t_obs;
  double* out;
  double* scratch;      /* as in pbgpu_launch */
} pbgpu_observe_launch;

typedef struct {
  int which, n_sys, out_len;
  const double* y;
  const double* p;
  const double* y2;
  double t, cj;
  double* out;
  double* scratch;      /* as in pbgpu_launch */
} pbgpu_eval_launch;

#ifdef __cplusplus
}
#endif
#endif
