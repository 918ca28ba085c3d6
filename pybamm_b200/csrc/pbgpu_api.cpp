// pbgpu_api.cpp -- libpbgpu.so: the C-ABI host library (include/pbgpu.h).
//
// Generic host side of the batched solver: loads compiled model images, owns device buffers and the
// stream, validates arguments the way IDAKLUSolverGroup::solve does
// (packages/pybammsolvers/src/pybammsolvers/idaklu_source/IDAKLUSolverGroup.cpp:46-97), moves the batch
// to the device, launches the image's persistent integration kernel and brings the results back.
// It contains no numerical code and no CPU fallback: every path ends in a kernel of the model image.
#include <cuda_runtime.h>
#include <dlfcn.h>

#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "pbgpu.h"

namespace {
thread_local std::string g_err;

int fail(const std::string& msg) {
  g_err = msg;
  return 1;
}
int cuda_fail(const char* what, cudaError_t e) { return fail(std::string(what) + ": " + cudaGetErrorString(e)); }

#define CK(call)                                    \
  do {                                              \
    cudaError_t e__ = (call);                       \
    if (e__ != cudaSuccess) return cuda_fail(#call, e__); \
  } while (0)

typedef int (*describe_fn)(pbgpu_model_desc*);
typedef int (*launch_fn)(const pbgpu_launch*, void*, int, char*, size_t);
typedef int (*eval_fn)(const pbgpu_eval_launch*, void*, int, char*, size_t);
typedef int (*plan_fn)(int, int*, char*, size_t);
typedef int (*observe_fn)(const pbgpu_observe_launch*, void*, int, char*, size_t);
typedef int (*launch_static_fn)(const pbgpu_launch*, void*, int*, char*, size_t);

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return cuda_fail("cudaMalloc", e);
    cap = bytes;
    return 0;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};
}  // namespace

struct pbgpu_model_s {
  void* handle = nullptr;
  describe_fn describe = nullptr;
  launch_fn launch = nullptr;
  eval_fn eval = nullptr;
  plan_fn plan = nullptr;
  observe_fn observe = nullptr;
  launch_static_fn launch_static = nullptr;   // optional: images built before the entry point existed lack it
  pbgpu_model_desc desc;
  std::string path;
};

struct pbgpu_solver_s {
  pbgpu_model_s* model = nullptr;
  int device = 0;
  pbgpu_options opt;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  DevBuf atol, t_eval, t_interp, inputs, y0, yp0, t_out, y_out, yp_out, n_out, flag, y_term, stats, event_idx, t_event, queue,
      scratch_a, scratch_b, team_scratch, order, sc_slot, sc_val, sc_dur, sc_npts, sc_stop, sc_end, sc_term, ob_nt, ob_t, ob_y, ob_yp, ob_tobs, ob_out;
  int last_grid = 0;
  int order_n = 0;
};

extern "C" {

const char* pbgpu_last_error(void) { return g_err.c_str(); }

void pbgpu_default_options(pbgpu_options* o) {
  o->rtol = 1e-4;
  o->max_order_bdf = 5;
  o->max_num_steps = 100000;
  o->max_error_test_failures = 10;
  o->max_nonlinear_iterations = 40;
  o->max_convergence_failures = 100;
  o->nonlinear_convergence_coefficient = 0.33;
  o->nonlinear_convergence_coefficient_ic = 0.0033;
  o->newton_step_tol = 1e-4;
  o->max_num_iterations_ic = 100;
  o->max_linesearch_backtracks_ic = 5;
  o->linear_solution_scaling = 1;
  o->calc_ic = 1;
  o->dt_init = 0.0;
  o->dt_max = 0.0;
  o->num_steps_no_progress = 0;
  o->t_no_progress = 0.0;
}

int pbgpu_model_load(const char* path, pbgpu_model_t* model) {
  if (!path || !model) return fail("pbgpu_model_load: null argument");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev < 1)
    return fail(std::string("pbgpu_model_load: no usable CUDA device (") + (e != cudaSuccess ? cudaGetErrorString(e) : "0 devices") +
                "); this solver has no CPU fallback");
  void* h = dlopen(path, RTLD_NOW | RTLD_LOCAL);
  if (!h) return fail(std::string("pbgpu_model_load: ") + dlerror());
  pbgpu_model_s* m = new pbgpu_model_s();
  m->handle = h;
  m->path = path;
  m->describe = (describe_fn)dlsym(h, "pbgpu_image_describe");
  m->launch = (launch_fn)dlsym(h, "pbgpu_image_launch");
  m->eval = (eval_fn)dlsym(h, "pbgpu_image_eval");
  m->plan = (plan_fn)dlsym(h, "pbgpu_image_plan");
  m->observe = (observe_fn)dlsym(h, "pbgpu_image_observe");
  m->launch_static = (launch_static_fn)dlsym(h, "pbgpu_image_launch_static");
  if (!m->describe || !m->launch || !m->eval || !m->plan || !m->observe) {
    dlclose(h);
    delete m;
    return fail(std::string("pbgpu_model_load: '") + path + "' is not a pbgpu model image");
  }
  m->describe(&m->desc);
  if (m->desc.abi_version != PBGPU_ABI_VERSION) {
    dlclose(h);
    delete m;
    return fail("pbgpu_model_load: model image was built against a different ABI version");
  }
  *model = m;
  return 0;
}

int pbgpu_model_describe(pbgpu_model_t model, pbgpu_model_desc* desc) {
  if (!model || !desc) return fail("pbgpu_model_describe: null argument");
  *desc = model->desc;
  return 0;
}

int pbgpu_model_unload(pbgpu_model_t model) {
  if (!model) return 0;
  // The image registers its kernels with the CUDA runtime when it is mapped and unregisters them from a static
  // destructor, so it can be unmapped once no solver uses it (the caller destroys its solvers first).
  if (model->handle) dlclose(model->handle);
  delete model;
  return 0;
}

int pbgpu_solver_destroy(pbgpu_solver_t s);

int pbgpu_solver_create(pbgpu_model_t model, const double* atol, const pbgpu_options* opt, int device,
                        pbgpu_solver_t* solver) {
  if (!model || !atol || !solver) return fail("pbgpu_solver_create: null argument");
  int ndev = 0;
  CK(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return fail("pbgpu_solver_create: invalid device ordinal " + std::to_string(device));
  CK(cudaSetDevice(device));
  pbgpu_solver_s* s = new pbgpu_solver_s();
  s->model = model;
  s->device = device;
  if (opt) s->opt = *opt;
  else pbgpu_default_options(&s->opt);
  for (int i = 0; i < model->desc.n_states; i++)
    if (!(atol[i] >= 0.0)) {
      delete s;
      return fail("pbgpu_solver_create: absolute tolerances must be non-negative");
    }
  if (!(s->opt.rtol >= 0.0)) {
    delete s;
    return fail("pbgpu_solver_create: rtol must be non-negative");
  }
  if (s->opt.num_steps_no_progress > 32) {
    delete s;
    return fail("pbgpu_solver_create: num_steps_no_progress is limited to 32 on this path");
  }
  cudaError_t e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    delete s;
    return cuda_fail("cudaStreamCreate", e);
  }
  for (int i = 0; i < 4; i++) cudaEventCreate(&s->ev[i]);
  int rc = s->atol.ensure(sizeof(double) * model->desc.n_states);
  if (!rc) {
    e = cudaMemcpy(s->atol.p, atol, sizeof(double) * model->desc.n_states, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) rc = cuda_fail("cudaMemcpy(atol)", e);
  }
  if (!rc) rc = s->queue.ensure(sizeof(unsigned int) * PBGPU_QUEUE_WORDS);
  if (rc) {
    const std::string msg = g_err;
    pbgpu_solver_destroy(s);   // releases the stream, the events and whatever was allocated
    g_err = msg;
    return 1;
  }
  *solver = s;
  return 0;
}

int pbgpu_solver_destroy(pbgpu_solver_t s) {
  if (!s) return 0;
  cudaSetDevice(s->device);
  if (s->stream) cudaStreamSynchronize(s->stream);
  DevBuf* bufs[] = {&s->atol, &s->t_eval, &s->t_interp, &s->inputs, &s->y0, &s->yp0, &s->t_out, &s->y_out, &s->yp_out,
                    &s->n_out, &s->flag, &s->y_term, &s->stats, &s->event_idx, &s->t_event, &s->queue, &s->scratch_a,
                    &s->scratch_b, &s->team_scratch, &s->order, &s->sc_slot, &s->sc_val, &s->sc_dur, &s->sc_npts, &s->sc_stop,
                    &s->sc_end, &s->sc_term, &s->ob_nt, &s->ob_t, &s->ob_y, &s->ob_yp, &s->ob_tobs, &s->ob_out};
  for (DevBuf* b : bufs) b->release();
  for (int i = 0; i < 4; i++)
    if (s->ev[i]) cudaEventDestroy(s->ev[i]);
  if (s->stream) cudaStreamDestroy(s->stream);
  delete s;
  return 0;
}

int pbgpu_result_capacity(int n_teval, int n_tinterp) { return n_teval + n_tinterp + 2; }

static int check_grids(const double* t_eval, int n_teval, const double* t_interp, int n_tinterp) {
  if (!t_eval || n_teval < 2) return fail("t_eval must have at least 2 entries");
  for (int i = 1; i < n_teval; i++)
    if (!(t_eval[i] > t_eval[i - 1])) return fail("t_eval must be sorted and unique");
  for (int i = 1; i < n_tinterp; i++)
    if (!(t_interp[i] > t_interp[i - 1])) return fail("t_interp must be sorted and unique");
  if (n_tinterp > 0) {
    if (t_interp[0] < t_eval[0])
      return fail("t_interp values must be greater than the smallest t_eval value: " + std::to_string(t_eval[0]));
    if (t_interp[n_tinterp - 1] > t_eval[n_teval - 1])
      return fail("t_interp values must be less than the greatest t_eval value: " + std::to_string(t_eval[n_teval - 1]));
  }
  return 0;
}

// Plan the launch (teams on this device for n_sys systems) and size the per-team global scratch (streamed vectors of
// large models) for exactly that grid.
static int plan_launch(pbgpu_solver_s* s, int n_sys, int* teams) {
  char err[512];
  err[0] = 0;
  if (s->model->plan(n_sys, teams, err, sizeof(err))) return fail(err);
  const long long per_team = s->model->desc.scratch_doubles_per_team;
  if (per_team <= 0) return 0;
  return s->team_scratch.ensure(sizeof(double) * (size_t)per_team * (size_t)*teams);
}

struct DevResult {
  double* t;
  double* y;
  double* yp;
  int* n;
  int* flag;
  double* yterm;
  long long* stats;
  int* event_idx;
  double* t_event;
};

static int launch_common(pbgpu_solver_s* s, int n_sys, const double* d_t_eval, int n_teval, const double* d_t_interp,
                         int n_tinterp, const double* d_y0, const double* d_yp0, const double* d_inputs, int output_mode,
                         int n_max, const DevResult& R, cudaStream_t stream, const pbgpu_schedule* d_sched = nullptr,
                         double t_start = 0.0, int* d_step_end = nullptr, int* d_step_term = nullptr) {
  const pbgpu_model_desc& D = s->model->desc;
  pbgpu_launch L;
  memset(&L, 0, sizeof(L));
  L.n_sys = n_sys;
  L.n_eval = n_teval;
  L.n_interp = n_tinterp;
  L.n_max = n_max;
  L.save_vars = (output_mode == PBGPU_OUTPUT_VARS);
  L.n_ret = L.save_vars ? D.n_vars : D.n_states;
  L.t_eval = d_t_eval;
  L.t_interp = d_t_interp;
  L.inputs = d_inputs;
  L.y0 = d_y0;
  L.yp0 = d_yp0;
  L.atol = (const double*)s->atol.p;
  L.t_out = R.t;
  L.y_out = R.y;
  L.yp_out = L.save_vars ? nullptr : R.yp;
  L.n_out = R.n;
  L.flag = R.flag;
  L.y_term = R.yterm;
  L.stats = R.stats;
  L.event_idx = R.event_idx;
  L.t_event = R.t_event;
  L.queue = (unsigned int*)s->queue.p;
  int teams = 0;
  if (plan_launch(s, n_sys, &teams)) return 1;
  L.scratch = (double*)s->team_scratch.p;
  L.order = (s->order_n == n_sys && n_sys > 0) ? (const int*)s->order.p : nullptr;
  L.opt = s->opt;
  if (d_sched) {
    L.sched = *d_sched;
    L.t_start = t_start;
    L.step_end = d_step_end;
    L.step_term = d_step_term;
  }
  CK(cudaMemsetAsync(s->queue.p, 0, sizeof(unsigned int) * PBGPU_QUEUE_WORDS, stream));
  char err[512];
  err[0] = 0;
  if (s->model->launch(&L, (void*)stream, teams, err, sizeof(err))) return fail(err);
  s->last_grid = teams;
  return 0;
}

int pbgpu_solve_device(pbgpu_solver_t s, int n_sys, const double* t_eval, int n_teval, const double* t_interp,
                       int n_tinterp, const double* y0, const double* yp0, const double* inputs, int output_mode,
                       pbgpu_result* out, void* stream, int sync) {
  if (!s || !out) return fail("pbgpu_solve_device: null argument");
  if (n_sys < 1) return fail("pbgpu_solve_device: n_sys must be positive");
  const pbgpu_model_desc& D = s->model->desc;
  if (output_mode == PBGPU_OUTPUT_VARS && D.n_vars < 1) return fail("the model image has no compiled output variables");
  if (n_teval < 2) return fail("t_eval must have at least 2 entries");
  if (D.n_inputs > 0 && !inputs) return fail("inputs has wrong number of dimensions. Expected 2 but got 0");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  if (sync) CK(cudaEventRecord(s->ev[0], st));
  const DevResult R = {out->t, out->y, out->yp, out->n_t, out->flag, out->y_term, out->stats, out->event_idx, out->t_event};
  if (launch_common(s, n_sys, t_eval, n_teval, t_interp, n_tinterp, y0, yp0, inputs, output_mode, out->n_max, R, st))
    return 1;
  if (sync) {
    CK(cudaEventRecord(s->ev[1], st));
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&out->kernel_ms, s->ev[0], s->ev[1]));
    out->total_ms = out->kernel_ms;
  }
  return 0;
}

// pbgpu_solve_device on the model image's static scratch array: nothing here (or below it) allocates.  Not routed through
// launch_common, whose plan_launch sizes the handle's cudaMalloc'ed team scratch.
int pbgpu_solve_device_noalloc(pbgpu_solver_t s, int n_sys, const double* t_eval, int n_teval, const double* t_interp,
                               int n_tinterp, const double* y0, const double* yp0, const double* inputs, int output_mode,
                               pbgpu_result* out, void* stream, int sync) {
  if (!s) return fail("pbgpu_solve_device_noalloc: null argument");
  if (!s->model->launch_static)
    return fail("pbgpu_solve_device_noalloc: the model image has no static-scratch launch (rebuild it)");
  if (n_sys < 0) return fail("pbgpu_solve_device_noalloc: n_sys must not be negative");
  if (n_sys > PBGPU_STATIC_MAX_SYSTEMS)
    return fail("pbgpu_solve_device_noalloc: n_sys = " + std::to_string(n_sys) + " is beyond the static limit of " +
                std::to_string(PBGPU_STATIC_MAX_SYSTEMS) + " systems per call (use pbgpu_solve_device, or split the batch)");
  const pbgpu_model_desc& D = s->model->desc;
  CK(cudaSetDevice(s->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  pbgpu_launch L;
  memset(&L, 0, sizeof(L));
  char err[512];
  err[0] = 0;
  int teams = 0;
  if (n_sys == 0) {   // set-up only
    if (s->model->launch_static(&L, (void*)st, &teams, err, sizeof(err))) return fail(err);
    return 0;
  }
  if (!out) return fail("pbgpu_solve_device_noalloc: null argument");
  if (output_mode == PBGPU_OUTPUT_VARS && D.n_vars < 1) return fail("the model image has no compiled output variables");
  if (n_teval < 2) return fail("t_eval must have at least 2 entries");
  if (D.n_inputs > 0 && !inputs) return fail("inputs has wrong number of dimensions. Expected 2 but got 0");
  L.n_sys = n_sys;
  L.n_eval = n_teval;
  L.n_interp = n_tinterp;
  L.n_max = out->n_max;
  L.save_vars = (output_mode == PBGPU_OUTPUT_VARS);
  L.n_ret = L.save_vars ? D.n_vars : D.n_states;
  L.t_eval = t_eval;
  L.t_interp = t_interp;
  L.inputs = inputs;
  L.y0 = y0;
  L.yp0 = yp0;
  L.atol = (const double*)s->atol.p;   // uploaded by pbgpu_solver_create
  L.t_out = out->t;
  L.y_out = out->y;
  L.yp_out = L.save_vars ? nullptr : out->yp;
  L.n_out = out->n_t;
  L.flag = out->flag;
  L.y_term = out->y_term;
  L.stats = out->stats;
  L.event_idx = out->event_idx;
  L.t_event = out->t_event;
  L.order = (s->order_n == n_sys) ? (const int*)s->order.p : nullptr;
  L.opt = s->opt;
  if (sync) CK(cudaEventRecord(s->ev[0], st));
  if (s->model->launch_static(&L, (void*)st, &teams, err, sizeof(err))) return fail(err);
  s->last_grid = teams;
  if (sync) {
    CK(cudaEventRecord(s->ev[1], st));
    CK(cudaStreamSynchronize(st));
    CK(cudaEventElapsedTime(&out->kernel_ms, s->ev[0], s->ev[1]));
    out->total_ms = out->kernel_ms;
  }
  return 0;
}

int pbgpu_solve(pbgpu_solver_t s, int n_sys, const double* t_eval, int n_teval, const double* t_interp, int n_tinterp,
                const double* y0, const double* yp0, const double* inputs, int output_mode, pbgpu_result* out) {
  if (!s || !out) return fail("pbgpu_solve: null argument");
  if (n_sys < 1) return fail("pbgpu_solve: n_sys must be positive");
  const pbgpu_model_desc& D = s->model->desc;
  if (check_grids(t_eval, n_teval, t_interp, n_tinterp)) return 1;
  if (output_mode != PBGPU_OUTPUT_STATES && output_mode != PBGPU_OUTPUT_VARS) return fail("invalid output_mode");
  if (output_mode == PBGPU_OUTPUT_VARS && D.n_vars < 1) return fail("the model image has no compiled output variables");
  if (output_mode == PBGPU_OUTPUT_VARS && out->yp) return fail("yp is only stored with the full state (output_mode 0)");
  const int n_ret = output_mode == PBGPU_OUTPUT_VARS ? D.n_vars : D.n_states;
  if (out->n_ret != n_ret)
    return fail("result block has wrong row length. Expected " + std::to_string(n_ret) + " but got " + std::to_string(out->n_ret));
  if (out->n_max < 1 || !out->t || !out->y || !out->n_t || !out->flag) return fail("result block is incomplete");
  if (n_tinterp > 0 && out->n_max < pbgpu_result_capacity(n_teval, n_tinterp))
    return fail("result block is too small for the requested t_eval/t_interp grids");
  if (D.n_inputs > 0 && !inputs) return fail("inputs has wrong number of dimensions. Expected 2 but got 0");
  CK(cudaSetDevice(s->device));
  const size_t N = D.n_states, nm = out->n_max;
  if (s->t_eval.ensure(sizeof(double) * n_teval)) return 1;
  if (s->t_interp.ensure(sizeof(double) * (n_tinterp > 0 ? n_tinterp : 1))) return 1;
  if (s->inputs.ensure(sizeof(double) * (size_t)n_sys * (D.n_inputs > 0 ? D.n_inputs : 1))) return 1;
  if (y0 && s->y0.ensure(sizeof(double) * n_sys * N)) return 1;
  if (yp0 && s->yp0.ensure(sizeof(double) * n_sys * N)) return 1;
  if (s->t_out.ensure(sizeof(double) * n_sys * nm)) return 1;
  if (s->y_out.ensure(sizeof(double) * n_sys * nm * n_ret)) return 1;
  if (out->yp && s->yp_out.ensure(sizeof(double) * n_sys * nm * N)) return 1;
  if (s->n_out.ensure(sizeof(int) * n_sys)) return 1;
  if (s->flag.ensure(sizeof(int) * n_sys)) return 1;
  if (out->y_term && s->y_term.ensure(sizeof(double) * n_sys * N)) return 1;
  if (out->stats && s->stats.ensure(sizeof(long long) * n_sys * 8)) return 1;
  if (out->event_idx && s->event_idx.ensure(sizeof(int) * n_sys)) return 1;
  if (out->t_event && s->t_event.ensure(sizeof(double) * n_sys)) return 1;
  cudaStream_t st = s->stream;
  CK(cudaEventRecord(s->ev[0], st));
  CK(cudaMemcpyAsync(s->t_eval.p, t_eval, sizeof(double) * n_teval, cudaMemcpyHostToDevice, st));
  if (n_tinterp > 0) CK(cudaMemcpyAsync(s->t_interp.p, t_interp, sizeof(double) * n_tinterp, cudaMemcpyHostToDevice, st));
  if (D.n_inputs > 0)
    CK(cudaMemcpyAsync(s->inputs.p, inputs, sizeof(double) * (size_t)n_sys * D.n_inputs, cudaMemcpyHostToDevice, st));
  if (y0) CK(cudaMemcpyAsync(s->y0.p, y0, sizeof(double) * n_sys * N, cudaMemcpyHostToDevice, st));
  if (yp0) CK(cudaMemcpyAsync(s->yp0.p, yp0, sizeof(double) * n_sys * N, cudaMemcpyHostToDevice, st));
  CK(cudaEventRecord(s->ev[1], st));
  const DevResult R = {(double*)s->t_out.p, (double*)s->y_out.p, out->yp ? (double*)s->yp_out.p : nullptr, (int*)s->n_out.p,
                       (int*)s->flag.p, out->y_term ? (double*)s->y_term.p : nullptr,
                       out->stats ? (long long*)s->stats.p : nullptr, out->event_idx ? (int*)s->event_idx.p : nullptr,
                       out->t_event ? (double*)s->t_event.p : nullptr};
  if (launch_common(s, n_sys, (const double*)s->t_eval.p, n_teval, n_tinterp > 0 ? (const double*)s->t_interp.p : nullptr,
                    n_tinterp, y0 ? (const double*)s->y0.p : nullptr, yp0 ? (const double*)s->yp0.p : nullptr,
                    (const double*)s->inputs.p, output_mode, out->n_max, R, st))
    return 1;
  CK(cudaEventRecord(s->ev[2], st));
  CK(cudaMemcpyAsync(out->t, s->t_out.p, sizeof(double) * n_sys * nm, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out->y, s->y_out.p, sizeof(double) * n_sys * nm * n_ret, cudaMemcpyDeviceToHost, st));
  if (out->yp) CK(cudaMemcpyAsync(out->yp, s->yp_out.p, sizeof(double) * n_sys * nm * N, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out->n_t, s->n_out.p, sizeof(int) * n_sys, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out->flag, s->flag.p, sizeof(int) * n_sys, cudaMemcpyDeviceToHost, st));
  if (out->y_term) CK(cudaMemcpyAsync(out->y_term, s->y_term.p, sizeof(double) * n_sys * N, cudaMemcpyDeviceToHost, st));
  if (out->stats) CK(cudaMemcpyAsync(out->stats, s->stats.p, sizeof(long long) * n_sys * 8, cudaMemcpyDeviceToHost, st));
  if (out->event_idx) CK(cudaMemcpyAsync(out->event_idx, s->event_idx.p, sizeof(int) * n_sys, cudaMemcpyDeviceToHost, st));
  if (out->t_event) CK(cudaMemcpyAsync(out->t_event, s->t_event.p, sizeof(double) * n_sys, cudaMemcpyDeviceToHost, st));
  CK(cudaEventRecord(s->ev[3], st));
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return cuda_fail("integration kernel", e);
  CK(cudaEventElapsedTime(&out->kernel_ms, s->ev[1], s->ev[2]));
  CK(cudaEventElapsedTime(&out->total_ms, s->ev[0], s->ev[3]));
  return 0;
}

int pbgpu_schedule_capacity(const pbgpu_schedule* q) {
  if (!q || q->n_steps < 1 || !q->n_pts) return 0;
  long long n = 0;
  for (int k = 0; k < q->n_steps; k++) n += (q->n_pts[k] < 2 ? 2 : q->n_pts[k]) + 1;
  return n > 2000000000LL ? 0 : (int)n;
}

int pbgpu_solve_experiment(pbgpu_solver_t s, int n_sys, double t0, const pbgpu_schedule* q, const double* y0,
                           const double* inputs, int output_mode, pbgpu_result* out, int* step_end, int* step_term) {
  if (!s || !out || !q || !step_end || !step_term) return fail("pbgpu_solve_experiment: null argument");
  if (n_sys < 1) return fail("pbgpu_solve_experiment: n_sys must be positive");
  const pbgpu_model_desc& D = s->model->desc;
  if (q->n_steps < 1 || !q->duration || !q->n_pts) return fail("pbgpu_solve_experiment: the schedule has no steps");
  if (q->n_over < 0 || (q->n_over > 0 && (!q->over_slot || !q->over_val))) return fail("pbgpu_solve_experiment: incomplete input overrides");
  for (int j = 0; j < q->n_over; j++)
    if (q->over_slot[j] < 0 || q->over_slot[j] >= D.n_inputs) return fail("pbgpu_solve_experiment: override slot out of range");
  for (int k = 0; k < q->n_steps; k++)
    if (!(q->duration[k] > 0.0)) return fail("pbgpu_solve_experiment: step durations must be positive");
  if (q->experiment_event >= D.n_events) return fail("pbgpu_solve_experiment: experiment_event out of range");
  if (output_mode != PBGPU_OUTPUT_STATES && output_mode != PBGPU_OUTPUT_VARS) return fail("invalid output_mode");
  if (output_mode == PBGPU_OUTPUT_VARS && D.n_vars < 1) return fail("the model image has no compiled output variables");
  const int n_ret = output_mode == PBGPU_OUTPUT_VARS ? D.n_vars : D.n_states;
  if (out->n_ret != n_ret)
    return fail("result block has wrong row length. Expected " + std::to_string(n_ret) + " but got " + std::to_string(out->n_ret));
  if (out->n_max < pbgpu_schedule_capacity(q) || !out->t || !out->y || !out->n_t || !out->flag)
    return fail("result block is too small for the schedule (pbgpu_schedule_capacity) or incomplete");
  if (D.n_inputs > 0 && !inputs) return fail("inputs has wrong number of dimensions. Expected 2 but got 0");
  CK(cudaSetDevice(s->device));
  const size_t N = D.n_states, nm = out->n_max, ns = q->n_steps;
  if (s->inputs.ensure(sizeof(double) * (size_t)n_sys * (D.n_inputs > 0 ? D.n_inputs : 1))) return 1;
  if (y0 && s->y0.ensure(sizeof(double) * n_sys * N)) return 1;
  if (s->t_out.ensure(sizeof(double) * n_sys * nm)) return 1;
  if (s->y_out.ensure(sizeof(double) * n_sys * nm * n_ret)) return 1;
  if (out->yp && s->yp_out.ensure(sizeof(double) * n_sys * nm * N)) return 1;
  if (s->n_out.ensure(sizeof(int) * n_sys)) return 1;
  if (s->flag.ensure(sizeof(int) * n_sys)) return 1;
  if (out->y_term && s->y_term.ensure(sizeof(double) * n_sys * N)) return 1;
  if (out->stats && s->stats.ensure(sizeof(long long) * n_sys * 8)) return 1;
  if (out->event_idx && s->event_idx.ensure(sizeof(int) * n_sys)) return 1;
  if (out->t_event && s->t_event.ensure(sizeof(double) * n_sys)) return 1;
  if (s->sc_slot.ensure(sizeof(int) * (q->n_over > 0 ? q->n_over : 1))) return 1;
  if (s->sc_val.ensure(sizeof(double) * (ns * (q->n_over > 0 ? q->n_over : 1)))) return 1;
  if (s->sc_dur.ensure(sizeof(double) * ns)) return 1;
  if (s->sc_npts.ensure(sizeof(int) * ns)) return 1;
  if (s->sc_stop.ensure(ns)) return 1;
  if (s->sc_end.ensure(sizeof(int) * n_sys * ns)) return 1;
  if (s->sc_term.ensure(sizeof(int) * n_sys * ns)) return 1;
  cudaStream_t st = s->stream;
  CK(cudaEventRecord(s->ev[0], st));
  if (D.n_inputs > 0)
    CK(cudaMemcpyAsync(s->inputs.p, inputs, sizeof(double) * (size_t)n_sys * D.n_inputs, cudaMemcpyHostToDevice, st));
  if (y0) CK(cudaMemcpyAsync(s->y0.p, y0, sizeof(double) * n_sys * N, cudaMemcpyHostToDevice, st));
  if (q->n_over > 0) {
    CK(cudaMemcpyAsync(s->sc_slot.p, q->over_slot, sizeof(int) * q->n_over, cudaMemcpyHostToDevice, st));
    CK(cudaMemcpyAsync(s->sc_val.p, q->over_val, sizeof(double) * ns * q->n_over, cudaMemcpyHostToDevice, st));
  }
  CK(cudaMemcpyAsync(s->sc_dur.p, q->duration, sizeof(double) * ns, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->sc_npts.p, q->n_pts, sizeof(int) * ns, cudaMemcpyHostToDevice, st));
  if (q->stop_on_final_time) CK(cudaMemcpyAsync(s->sc_stop.p, q->stop_on_final_time, ns, cudaMemcpyHostToDevice, st));
  else CK(cudaMemsetAsync(s->sc_stop.p, 0, ns, st));
  pbgpu_schedule dq = *q;
  dq.over_slot = (const int*)s->sc_slot.p;
  dq.over_val = (const double*)s->sc_val.p;
  dq.duration = (const double*)s->sc_dur.p;
  dq.n_pts = (const int*)s->sc_npts.p;
  dq.stop_on_final_time = (const unsigned char*)s->sc_stop.p;
  CK(cudaEventRecord(s->ev[1], st));
  const DevResult R = {(double*)s->t_out.p, (double*)s->y_out.p, out->yp ? (double*)s->yp_out.p : nullptr, (int*)s->n_out.p,
                       (int*)s->flag.p, out->y_term ? (double*)s->y_term.p : nullptr,
                       out->stats ? (long long*)s->stats.p : nullptr, out->event_idx ? (int*)s->event_idx.p : nullptr,
                       out->t_event ? (double*)s->t_event.p : nullptr};
  if (launch_common(s, n_sys, nullptr, 2, nullptr, 0, y0 ? (const double*)s->y0.p : nullptr, nullptr, (const double*)s->inputs.p,
                    output_mode, out->n_max, R, st, &dq, t0, (int*)s->sc_end.p, (int*)s->sc_term.p))
    return 1;
  CK(cudaEventRecord(s->ev[2], st));
  CK(cudaMemcpyAsync(out->t, s->t_out.p, sizeof(double) * n_sys * nm, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out->y, s->y_out.p, sizeof(double) * n_sys * nm * n_ret, cudaMemcpyDeviceToHost, st));
  if (out->yp) CK(cudaMemcpyAsync(out->yp, s->yp_out.p, sizeof(double) * n_sys * nm * N, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out->n_t, s->n_out.p, sizeof(int) * n_sys, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(out->flag, s->flag.p, sizeof(int) * n_sys, cudaMemcpyDeviceToHost, st));
  if (out->y_term) CK(cudaMemcpyAsync(out->y_term, s->y_term.p, sizeof(double) * n_sys * N, cudaMemcpyDeviceToHost, st));
  if (out->stats) CK(cudaMemcpyAsync(out->stats, s->stats.p, sizeof(long long) * n_sys * 8, cudaMemcpyDeviceToHost, st));
  if (out->event_idx) CK(cudaMemcpyAsync(out->event_idx, s->event_idx.p, sizeof(int) * n_sys, cudaMemcpyDeviceToHost, st));
  if (out->t_event) CK(cudaMemcpyAsync(out->t_event, s->t_event.p, sizeof(double) * n_sys, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(step_end, s->sc_end.p, sizeof(int) * n_sys * ns, cudaMemcpyDeviceToHost, st));
  CK(cudaMemcpyAsync(step_term, s->sc_term.p, sizeof(int) * n_sys * ns, cudaMemcpyDeviceToHost, st));
  CK(cudaEventRecord(s->ev[3], st));
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return cuda_fail("experiment kernel", e);
  CK(cudaEventElapsedTime(&out->kernel_ms, s->ev[1], s->ev[2]));
  CK(cudaEventElapsedTime(&out->total_ms, s->ev[0], s->ev[3]));
  return 0;
}

int pbgpu_observe_device(pbgpu_solver_t s, int n_sys, int n_max, const int* n_t, const double* t, const double* y,
                         const double* yp, const double* inputs, const double* t_obs, int n_obs, double* out, void* stream,
                         int sync) {
  if (!s || !n_t || !t || !y || !out) return fail("pbgpu_observe: null argument");
  if (n_sys < 1 || n_max < 1) return fail("pbgpu_observe: n_sys and n_max must be positive");
  const pbgpu_model_desc& D = s->model->desc;
  if (D.n_vars < 1) return fail("the model image has no compiled output variables");
  if (D.n_inputs > 0 && !inputs) return fail("inputs has wrong number of dimensions. Expected 2 but got 0");
  if (t_obs && n_obs < 1) return fail("pbgpu_observe: t_obs given but n_obs < 1");
  if (t_obs && !yp) return fail("pbgpu_observe: Hermite interpolation needs yp (solve with the full state and adaptive steps)");
  CK(cudaSetDevice(s->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : s->stream;
  pbgpu_observe_launch O;
  memset(&O, 0, sizeof(O));
  O.n_sys = n_sys;
  O.n_max = n_max;
  O.n_obs = t_obs ? n_obs : 0;
  O.n_t = n_t;
  O.t = t;
  O.y = y;
  O.yp = yp;
  O.inputs = inputs;
  O.t_obs = t_obs;
  O.out = out;
  int teams = 0;
  if (plan_launch(s, n_sys, &teams)) return 1;
  O.scratch = (double*)s->team_scratch.p;
  char err[512];
  err[0] = 0;
  if (s->model->observe(&O, (void*)st, teams, err, sizeof(err))) return fail(err);
  if (sync) {
    cudaError_t e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) return cuda_fail("observation kernel", e);
  }
  return 0;
}

int pbgpu_observe(pbgpu_solver_t s, int n_sys, int n_max, const int* n_t, const double* t, const double* y, const double* yp,
                  const double* inputs, const double* t_obs, int n_obs, double* out) {
  if (!s || !n_t || !t || !y || !out) return fail("pbgpu_observe: null argument");
  if (n_sys < 1 || n_max < 1) return fail("pbgpu_observe: n_sys and n_max must be positive");
  const pbgpu_model_desc& D = s->model->desc;
  if (D.n_vars < 1) return fail("the model image has no compiled output variables");
  if (t_obs && !yp) return fail("pbgpu_observe: Hermite interpolation needs yp (solve with the full state and adaptive steps)");
  for (int i = 1; t_obs && i < n_obs; i++)
    if (!(t_obs[i] >= t_obs[i - 1])) return fail("pbgpu_observe: t_obs must be sorted");
  CK(cudaSetDevice(s->device));
  const size_t N = D.n_states, rows = t_obs ? (size_t)n_obs : (size_t)n_max;
  const size_t np = D.n_inputs > 0 ? D.n_inputs : 1;
  if (s->ob_nt.ensure(sizeof(int) * n_sys)) return 1;
  if (s->ob_t.ensure(sizeof(double) * (size_t)n_sys * n_max)) return 1;
  if (s->ob_y.ensure(sizeof(double) * (size_t)n_sys * n_max * N)) return 1;
  if (yp && s->ob_yp.ensure(sizeof(double) * (size_t)n_sys * n_max * N)) return 1;
  if (t_obs && s->ob_tobs.ensure(sizeof(double) * n_obs)) return 1;
  if (s->inputs.ensure(sizeof(double) * (size_t)n_sys * np)) return 1;
  if (s->ob_out.ensure(sizeof(double) * (size_t)n_sys * rows * D.n_vars)) return 1;
  cudaStream_t st = s->stream;
  CK(cudaMemcpyAsync(s->ob_nt.p, n_t, sizeof(int) * n_sys, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->ob_t.p, t, sizeof(double) * (size_t)n_sys * n_max, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(s->ob_y.p, y, sizeof(double) * (size_t)n_sys * n_max * N, cudaMemcpyHostToDevice, st));
  if (yp) CK(cudaMemcpyAsync(s->ob_yp.p, yp, sizeof(double) * (size_t)n_sys * n_max * N, cudaMemcpyHostToDevice, st));
  if (t_obs) CK(cudaMemcpyAsync(s->ob_tobs.p, t_obs, sizeof(double) * n_obs, cudaMemcpyHostToDevice, st));
  if (D.n_inputs > 0) CK(cudaMemcpyAsync(s->inputs.p, inputs, sizeof(double) * (size_t)n_sys * D.n_inputs, cudaMemcpyHostToDevice, st));
  CK(cudaMemsetAsync(s->ob_out.p, 0, sizeof(double) * (size_t)n_sys * rows * D.n_vars, st));
  if (pbgpu_observe_device(s, n_sys, n_max, (const int*)s->ob_nt.p, (const double*)s->ob_t.p, (const double*)s->ob_y.p,
                           yp ? (const double*)s->ob_yp.p : nullptr, (const double*)s->inputs.p,
                           t_obs ? (const double*)s->ob_tobs.p : nullptr, n_obs, (double*)s->ob_out.p, (void*)st, 0))
    return 1;
  CK(cudaMemcpyAsync(out, s->ob_out.p, sizeof(double) * (size_t)n_sys * rows * D.n_vars, cudaMemcpyDeviceToHost, st));
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return cuda_fail("observation kernel", e);
  return 0;
}

int pbgpu_solver_set_launch_order(pbgpu_solver_t s, const int* order, int n) {
  if (!s) return fail("pbgpu_solver_set_launch_order: bad argument");
  if (!order || n <= 0) {
    s->order_n = 0;
    return 0;
  }
  std::vector<char> seen((size_t)n, 0);
  for (int i = 0; i < n; i++) {
    if (order[i] < 0 || order[i] >= n || seen[(size_t)order[i]]) return fail("pbgpu_solver_set_launch_order: not a permutation of 0..n-1");
    seen[(size_t)order[i]] = 1;
  }
  CK(cudaSetDevice(s->device));
  if (s->order.ensure(sizeof(int) * (size_t)n)) return 1;
  CK(cudaMemcpyAsync(s->order.p, order, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, s->stream));
  CK(cudaStreamSynchronize(s->stream));
  s->order_n = n;
  return 0;
}

int pbgpu_model_eval(pbgpu_solver_t s, int which, int n_sys, const double* y, const double* p, double t, double cj,
                     const double* y2, double* out, int out_len) {
  if (!s || !out || n_sys < 1) return fail("pbgpu_model_eval: bad argument");
  const pbgpu_model_desc& D = s->model->desc;
  CK(cudaSetDevice(s->device));
  const size_t N = D.n_states;
  const int np = D.n_inputs > 0 ? D.n_inputs : 1;
  if (s->y0.ensure(sizeof(double) * n_sys * N)) return 1;
  if (s->inputs.ensure(sizeof(double) * (size_t)n_sys * np)) return 1;
  if (s->scratch_a.ensure(sizeof(double) * n_sys * N)) return 1;
  if (s->scratch_b.ensure(sizeof(double) * (size_t)n_sys * out_len)) return 1;
  cudaStream_t st = s->stream;
  if (y) CK(cudaMemcpyAsync(s->y0.p, y, sizeof(double) * n_sys * N, cudaMemcpyHostToDevice, st));
  else CK(cudaMemsetAsync(s->y0.p, 0, sizeof(double) * n_sys * N, st));
  if (D.n_inputs > 0) CK(cudaMemcpyAsync(s->inputs.p, p, sizeof(double) * (size_t)n_sys * D.n_inputs, cudaMemcpyHostToDevice, st));
  if (y2) CK(cudaMemcpyAsync(s->scratch_a.p, y2, sizeof(double) * n_sys * N, cudaMemcpyHostToDevice, st));
  else CK(cudaMemsetAsync(s->scratch_a.p, 0, sizeof(double) * n_sys * N, st));
  CK(cudaMemsetAsync(s->scratch_b.p, 0, sizeof(double) * (size_t)n_sys * out_len, st));
  pbgpu_eval_launch E;
  memset(&E, 0, sizeof(E));
  E.which = which;
  E.n_sys = n_sys;
  E.out_len = out_len;
  E.y = (const double*)s->y0.p;
  E.p = (const double*)s->inputs.p;
  E.y2 = (const double*)s->scratch_a.p;
  E.t = t;
  E.cj = cj;
  E.out = (double*)s->scratch_b.p;
  int teams = 0;
  if (plan_launch(s, n_sys, &teams)) return 1;
  E.scratch = (double*)s->team_scratch.p;
  char err[512];
  err[0] = 0;
  if (s->model->eval(&E, (void*)st, teams, err, sizeof(err))) return fail(err);
  CK(cudaMemcpyAsync(out, s->scratch_b.p, sizeof(double) * (size_t)n_sys * out_len, cudaMemcpyDeviceToHost, st));
  cudaError_t e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return cuda_fail("evaluation kernel", e);
  return 0;
}

}  // extern "C"
