/* pbgpu_peak.h -- measurement helper of bench.py (not part of the drop-in boundary): the FP64 DFMA
 * rate of one device, the roofline denominator of the batched DAE path. */
#ifndef PBGPU_PEAK_H
#define PBGPU_PEAK_H
#ifdef __cplusplus
extern "C" {
#endif
/* Runs a DFMA-chain kernel over all SMs; best of 5. Returns 0 on success. */
int pbgpu_measure_fp64_peak(int device, int iters, double* tflops, float* ms);
#ifdef __cplusplus
}
#endif
#endif
