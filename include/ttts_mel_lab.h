/*
 * ttts_mel_lab.h -- development / measurement entry points of libttts_mel.  NOT part of the drop-in boundary
 * (include/ttts_mel.h is): nothing in the product path calls them.
 *
 *   measurement hooks   ttts_mel_profile_*: exported by both libraries; per plan, they never change a result
 *   lab build only      (libttts_mel_lab.so, compiled with -DTTTS_MEL_LAB; used by tests/ and tools/):
 *                       the cross-check kernels, the DFT-stage dump, the leg-removal switch and the TTTS_MEL_ALGO /
 *                       TTTS_MEL_PAIR / TTTS_MEL_NB_MAX / TTTS_RESAMPLE_* / TTTS_VQ_EXP environment knobs
 */
#ifndef TTTS_MEL_LAB_H_
#define TTTS_MEL_LAB_H_

#include "ttts_mel.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Measurement hooks: when enabled for a plan, every forward of THAT plan records a plan-owned CUDA event pair around
 * the fused log-mel kernel on the launch stream (up to max_calls calls; 0 disables and frees).
 * ttts_mel_profile_collect synchronises the device, writes the per-call kernel durations in ms and returns how many
 * were written (then restarts the ring). */
int ttts_mel_profile_enable(ttts_mel_plan* plan, int32_t max_calls);
int ttts_mel_profile_collect(ttts_mel_plan* plan, float* ms_out, int32_t capacity);

#ifdef TTTS_MEL_LAB
/* Same contract, computed by a plain CUDA-core fp32 kernel (no tensor cores, no pipeline).
 * A second, independent device implementation used by the GPU tests to cross-check the tcgen05
 * kernel at sizes the CPU oracle cannot reach.  Not a fallback: nothing dispatches to it. */
int ttts_mel_forward_simt(const ttts_mel_plan* plan, const float* wav_dev, int64_t row_stride,
                          const int64_t* lengths_dev_or_null, int32_t B, int64_t Lmax, int32_t Tmax,
                          float* mel_out_dev, int32_t* n_frames_dev_or_null, uint8_t* mask_dev_or_null,
                          void* workspace_dev, size_t workspace_bytes, void* stream);

/* ttts_mel_forward plus a dump of the DFT stage for stage-level parity tests:
 *   reim_out_dev [total_frames, 2*n_bins] fp32: per frame (clips concatenated, valid frames only)
 *   n_bins real parts then n_bins imaginary-part magnitudes, bins ttts_mel_first_bin(plan) .. +n_bins. */
int ttts_mel_forward_debug(const ttts_mel_plan* plan, const float* wav_dev, int64_t row_stride,
                           const int64_t* lengths_dev_or_null, int32_t B, int64_t Lmax, int32_t Tmax,
                           float* mel_out_dev, int32_t* n_frames_dev_or_null, uint8_t* mask_dev_or_null,
                           void* workspace_dev, size_t workspace_bytes, void* stream,
                           float* reim_out_dev);

/* Timing experiments for kernel development (RESULTS ARE INVALID while a flag is set; tools/gpu_experiments.py):
 * 1 removes the converter -> MMA operand handoff, 2 the basis loads, 4 the epilogue, 8 the drain arithmetic (the
 * accumulators are still read), 16 the log pass, 32 the MMAs themselves.  0 restores normal operation. */
void ttts_mel_set_experiment(int32_t flags);
#endif /* TTTS_MEL_LAB */

#ifdef __cplusplus
}
#endif
#endif /* TTTS_MEL_LAB_H_ */
