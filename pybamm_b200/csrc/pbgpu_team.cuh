// pbgpu_team.cuh -- team (one warp or one CTA per system) primitives for sm_100a.
//
// A *team* of PB_TEAM_SIZE threads integrates one system; all per-system vectors live in the
// team's shared memory.  The generated model code (pybamm_b200/codegen.py, CUDA dialect) and
// the integrator (pbgpu_ida.cuh) are written against the macros below.
//
// PB_EMU: single-"thread" host build of the very same sources (TEAM = 1, barriers are no-ops).
// It exists so that tests/ can exercise the device integrator's control flow on a CPU-only
// machine; it is never compiled into the product libraries.
#pragma once

#ifdef PB_EMU
#include <cmath>
#include <cstdint>
#include <cstring>
#define PB_DEV
#define PB_ALIGN16
#define PB_TABLE
#define PB_FN static inline
#define PB_RESTRICT __restrict__
#define PB_TEAM 1
#define PB_TID 0
#define PB_WARP 0
#define PB_NWARPS 1
#define PB_LANE 0
#define PB_LANES 1
#define PB_SYNC() ((void)0)
#define PB_SYNCWARP() ((void)0)
#define PB_STCG(p, v) (*(p) = (v))
#define PB_LD(x) (x)
#define PB_LDU(x) (x)
#define PB_NAN NAN
#define PB_INF INFINITY
struct pb_team_ctx {
  double* red;  // unused
};
static inline long long clock64() { return 0; }   // cycle traces are device-only
PB_FN double pb_team_sum(pb_team_ctx&, double x) { return x; }
PB_FN double pb_team_max(pb_team_ctx&, double x) { return x; }
#else
#include <cuda_runtime.h>
#include <math_constants.h>
#ifndef PB_TEAM_SIZE
#define PB_TEAM_SIZE 32
#endif
#define PB_DEV __device__ __forceinline__
#define PB_ALIGN16 __align__(16)
#define PB_TABLE __device__
#define PB_CTABLE __constant__
#define PB_FN static __device__ __noinline__
#define PB_RESTRICT __restrict__
#define PB_TEAM PB_TEAM_SIZE
#define PB_NWARPS (PB_TEAM_SIZE / 32)
#define PB_LANE ((int)(threadIdx.x & 31))
#define PB_LANES 32
#ifndef PB_PAIR
#define PB_PAIR 0
#endif
#if PB_PAIR
// Two teams in ONE CTA (pbgpu_solve_kernel_pair): each keeps its own named barrier, and when both are about to run the
// same large function (generated residual, triangular solves) they run it on the CTA-wide barrier instead -- lock-step
// stage by stage, so that the second team finds the instructions the first one fetched (see pb_pair_enter).
// Control words at the start of dynamic shared memory: [0] ticket, [2 + team] 1 while the team is on the joint barrier.
#define PB_PAIR_CTRL_DOUBLES 8
extern __shared__ double pb_smem[];
#define PB_TEAM_ID ((int)(threadIdx.x / PB_TEAM_SIZE))
#define PB_TID ((int)(threadIdx.x % PB_TEAM_SIZE))
#define PB_WARP ((int)((threadIdx.x % PB_TEAM_SIZE) >> 5))
#define PB_PAIR_CTL ((volatile unsigned int*)pb_smem)
static __device__ __forceinline__ void pb_bar_solo() {
  asm volatile("bar.sync %0, %1;" ::"r"(1 + PB_TEAM_ID), "n"(PB_TEAM_SIZE) : "memory");
}
static __device__ __forceinline__ void pb_bar_team() {
  if (PB_PAIR_CTL[2 + PB_TEAM_ID]) asm volatile("bar.sync 0, %0;" ::"n"(2 * PB_TEAM_SIZE) : "memory");
  else pb_bar_solo();
}
#define PB_SYNC() pb_bar_team()
#else
#define PB_TID ((int)threadIdx.x)
#define PB_WARP ((int)(threadIdx.x >> 5))
#if PB_TEAM_SIZE == 32
#define PB_SYNC() __syncwarp()
#else
#define PB_SYNC() __syncthreads()
#endif
#endif
#define PB_SYNCWARP() __syncwarp()
#define PB_STCG(p, v) __stcg(p, v)   /* L2 only: keep the small L1 for the tables */
#define PB_LD(x) __ldg(&(x))
#define PB_LDU(x) __ldg(&(x))
#define PB_NAN CUDART_NAN
#define PB_INF CUDART_INF

// Out-of-line copies of the long libm sequences: the generated model functions call them from dozens of sites, and
// the integration is bound by instruction fetch (1 MB of straight-line code against a 128 KB instruction cache).
#define PB_NOINLINE_FN(name, expr) \
  static __device__ __noinline__ double name(double x) { return expr; }
PB_NOINLINE_FN(pb_f_exp, exp(x))
PB_NOINLINE_FN(pb_f_log, log(x))
PB_NOINLINE_FN(pb_f_tanh, tanh(x))
PB_NOINLINE_FN(pb_f_sinh, sinh(x))
PB_NOINLINE_FN(pb_f_cosh, cosh(x))
PB_NOINLINE_FN(pb_f_asinh, asinh(x))
static __device__ __noinline__ double pb_f_pow(double x, double y) { return pow(x, y); }
#define PB_EXP(x) pb_f_exp(x)
#define PB_LOG(x) pb_f_log(x)
#define PB_TANH(x) pb_f_tanh(x)
#define PB_SINH(x) pb_f_sinh(x)
#define PB_COSH(x) pb_f_cosh(x)
#define PB_ASINH(x) pb_f_asinh(x)
#define PB_POW(x, y) pb_f_pow(x, y)
// reg_power (functions.py:412-431) is x / (x^2 + delta^2)^(1/4): its only powers are x^-1/4 and, in the Jacobian, x^-5/4.
// Two square roots (correctly rounded each) replace libdevice's 150-instruction pow() on the critical path of the
// electrode-kinetics templates.  PBGPU_LIBM_POW=1 at build time keeps pow() (A/B switch).
#ifndef PB_LIBM_POW
static __device__ __noinline__ double pb_f_powm025(double x) { return rsqrt(sqrt(x)); }
static __device__ __noinline__ double pb_f_powm125(double x) { return rsqrt(sqrt(x)) / x; }
#define PB_POWM025(x) pb_f_powm025(x)
#define PB_POWM125(x) pb_f_powm125(x)
#endif
// FP64 division is a 20-instruction sequence with a slow path: one shared copy (PB_INLINE_DIV=1 inlines it again).
#ifdef PB_PAD_NOPS   /* experiment: how much does the integration slow down per KB of (executed) code?  16 NOPs = 256 B per template */
#define PB_PAD1 (void)pad_[0];
#define PB_PAD { volatile const double* pad_ = p; PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 \
                 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 PB_PAD1 }
#endif
#ifndef PB_INLINE_SQRT
static __device__ __noinline__ double pb_f_sqrt(double x) { return sqrt(x); }
#define PB_SQRT(x) pb_f_sqrt(x)
#endif
#ifndef PB_INLINE_DIV
static __device__ __noinline__ double pb_f_div(double a, double b) { return a / b; }
#define PB_DIV(a, b) pb_f_div(a, b)
#endif

// 2- and 4-wide copies of the out-of-line libm routines (PB_B2 / PB_B4 in the generated code): one call evaluates
// independent arguments, and the compiler interleaves their dependency chains.  A lane's template body is a serial FP64
// chain at ~8 cycles per instruction; a terminal-voltage expression is sixteen such routines in a row.
#ifndef PB_NO_BATCH
#define PB_BATCH_FN(name, expr)                                                                       \
  static __device__ __noinline__ double2 pb_f2_##name(double x0, double x1) {                         \
    double2 r;                                                                                        \
    { const double x = x0; r.x = expr; }                                                              \
    { const double x = x1; r.y = expr; }                                                              \
    return r;                                                                                         \
  }                                                                                                   \
  static __device__ __noinline__ double4 pb_f4_##name(double x0, double x1, double x2, double x3) {   \
    double4 r;                                                                                        \
    { const double x = x0; r.x = expr; }                                                              \
    { const double x = x1; r.y = expr; }                                                              \
    { const double x = x2; r.z = expr; }                                                              \
    { const double x = x3; r.w = expr; }                                                              \
    return r;                                                                                         \
  }
PB_BATCH_FN(exp, exp(x))
PB_BATCH_FN(log, log(x))
PB_BATCH_FN(tanh, tanh(x))
PB_BATCH_FN(sinh, sinh(x))
PB_BATCH_FN(cosh, cosh(x))
PB_BATCH_FN(asinh, asinh(x))
PB_BATCH_FN(sqrt, sqrt(x))
PB_BATCH_FN(powm025, rsqrt(sqrt(x)))
PB_BATCH_FN(powm125, rsqrt(sqrt(x)) / x)
#define PB_B2(f, r0, r1, x0, x1)          \
  const double2 r0##_b = pb_f2_##f(x0, x1); \
  const double r0 = r0##_b.x, r1 = r0##_b.y;
#define PB_B4(f, r0, r1, r2, r3, x0, x1, x2, x3)    \
  const double4 r0##_b = pb_f4_##f(x0, x1, x2, x3);   \
  const double r0 = r0##_b.x, r1 = r0##_b.y, r2 = r0##_b.z, r3 = r0##_b.w;
#endif

struct pb_team_ctx {
  double* red;   // 2 * 32 doubles of shared memory for cross-warp reductions (double buffered)
  int red_phase;
};

// Team-wide sum; every thread returns the bitwise-identical value (butterfly inside the warp,
// fixed-order sum of the per-warp partials across warps).
// Inlined at about twenty call sites of ~110 instructions each.  One out-of-line copy (PB_TEAM_INLINE_RED = 0) makes
// the DFN image 43 KB smaller and the integration 2 % slower (4 550 against 4 636 sims/s): every reduction sits on the
// critical path of a step, and the call costs more than the instruction fetches it saves.
#ifndef PB_TEAM_INLINE_RED
#define PB_TEAM_INLINE_RED 1
#endif
#if PB_TEAM_INLINE_RED
#define PB_RED_FN static __device__ __forceinline__
#else
#define PB_RED_FN static __device__ __noinline__
#endif
PB_RED_FN double pb_team_sum_buf(double* buf, double x) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
#if PB_TEAM_SIZE > 32
  if (PB_LANE == 0) buf[PB_WARP] = x;
  PB_SYNC();
  x = buf[0];
#pragma unroll
  for (int w = 1; w < PB_NWARPS; w++) x += buf[w];
#endif
  return x;
}
PB_RED_FN double pb_team_max_buf(double* buf, double x) {
  // NaN-propagating max (a NaN anywhere must poison the result so that failures are seen)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    double y = __shfl_xor_sync(0xffffffffu, x, o);
    x = (x != x || y != y) ? CUDART_NAN : fmax(x, y);
  }
#if PB_TEAM_SIZE > 32
  if (PB_LANE == 0) buf[PB_WARP] = x;
  PB_SYNC();
  x = buf[0];
#pragma unroll
  for (int w = 1; w < PB_NWARPS; w++) {
    double y = buf[w];
    x = (x != x || y != y) ? CUDART_NAN : fmax(x, y);
  }
#endif
  return x;
}
static __device__ __forceinline__ double pb_team_sum(pb_team_ctx& c, double x) {
  double* buf = c.red + 32 * (c.red_phase & 1);   // double buffered: no barrier needed after the read-out
  c.red_phase++;
  return pb_team_sum_buf(buf, x);
}
static __device__ __forceinline__ double pb_team_max(pb_team_ctx& c, double x) {
  double* buf = c.red + 32 * (c.red_phase & 1);
  c.red_phase++;
  return pb_team_max_buf(buf, x);
}
#endif
