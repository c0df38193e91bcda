/*
 * ttts_mel.h -- C ABI of the B200 (sm_100a) batched log-mel front-end.
 *
 * The reference (boostpapa/xtts, package `ttts`) has no FFI layer for this path: the boundary
 * is a Python nn.Module, `TorchMelSpectrogram.forward(wav) -> mel`
 * (ttts/utils/utils.py:486-522), whose arithmetic is torchaudio.transforms.MelSpectrogram ->
 * torch.stft -> matmul, plus log(clamp) and / mel_norms.  This header is what a host binding
 * (ctypes here) calls instead of those library calls.  Plain pointers and
 * sizes only; no torch types.  All device pointers are CUDA device memory of the current
 * device; `stream` is a cudaStream_t passed as void*.
 *
 * Ownership: the caller allocates inputs, outputs and the workspace (so the caching allocator
 * and stream semantics stay the host framework's); the library owns only the plan constants.
 * Nothing in ttts_mel_forward allocates or synchronises.
 *
 * Threading: a plan is immutable after create/set_norms; forward is re-entrant across
 * streams and threads as long as each in-flight call has its own workspace.  The library keeps no
 * process-global state that can change a result and never reads the environment (kernel-selection
 * and leg-removal knobs exist only in the lab build, include/ttts_mel_lab.h).
 *
 * Errors: every int function returns 0 on success or a negative TTTS_MEL_E* code;
 * ttts_mel_last_error() returns a thread-local message for the last failure.
 */
#ifndef TTTS_MEL_H_
#define TTTS_MEL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TTTS_MEL_OK        0
#define TTTS_MEL_EINVAL   (-1)   /* bad shape / unsupported configuration / length <= pad      */
#define TTTS_MEL_ECUDA    (-2)   /* CUDA runtime error (text in ttts_mel_last_error)           */
#define TTTS_MEL_ENOMEM   (-3)   /* workspace too small / host allocation failed               */
#define TTTS_MEL_EDEVICE  (-4)   /* device-side check failed (pipeline timeout, bad length)    */

/* Constructor arguments of the reference modules, already resolved to numbers.
 *   V1  TorchMelSpectrogram.__init__      ttts/utils/utils.py:487-502      (n_mels 80, pad 512, power 2, clamp 1e-5)
 *   V2  MelSpectrogramFeatures.__init__   ttts/vocoder/feature_extractors.py:74-91 + safe_log modules.py:194-205
 *                                          (n_mels 100, pad 512 | 384 for padding="same", power 1, clamp 1e-7)
 *   V3  mel_spectrogram() / MelSpectrogramFeatures1   ttts/vocoder/feature_extractors.py:27-53,102-121
 *                                          (librosa Slaney filterbank, pad 384, power 1 with eps 1e-9, clamp 1e-5) */
typedef struct ttts_mel_config {
  int32_t n_fft;       /* 1024 (the only size built this round)                                  */
  int32_t hop;         /* 256                                                                    */
  int32_t n_mels;      /* columns of the filterbank, <= 128                                      */
  int32_t pad;         /* reflect padding on each side before framing: n_fft/2 for center=True   */
  float   power;       /* 2.0 (|X|^2) or 1.0 (|X|)                                                */
  float   clamp;       /* floor applied before the natural log                                   */
  float   spec_scale;  /* scales the complex spectrum: 1, or 1/||window||_2 for normalized=True  */
  float   power_eps;   /* power == 1 only: magnitude = sqrt(re^2 + im^2 + power_eps); 1e-9 for the   */
                       /* HiFi-GAN style mel_spectrogram(), feature_extractors.py:46; else 0        */
} ttts_mel_config;

typedef struct ttts_mel_plan ttts_mel_plan;

/* Replaces the buffers torchaudio builds in MelSpectrogram.__init__ (window, fb) -- S0 in
 * SURVEY.md 8a; reference call site ttts/utils/utils.py:498-502.
 *   fb_host     [n_fft/2+1, n_mels] row-major fp32, the exact filterbank of the reference
 *   window_host [n_fft] fp32 analysis window as torch.stft applies it (win_length taps centred)
 * Builds the folded, windowed, bin-pruned cos/sin basis (TF32 hi/lo split) and the sparse mel
 * table on the host and uploads them to the current device. */
int ttts_mel_plan_create(const ttts_mel_config* cfg, const float* fb_host,
                         const float* window_host, ttts_mel_plan** out_plan);

/* Replaces `self.mel_norms = torch.load(...)` + the broadcast divide, utils.py:503-507,519-521.
 * norms_host: [n_mels] fp32 host pointer, or NULL to skip the division.  Copied. */
int ttts_mel_plan_set_norms(ttts_mel_plan* plan, const float* norms_host_or_null);

void ttts_mel_plan_destroy(ttts_mel_plan* plan);

/* Frame count torch.stft produces for a clip of `length` samples (torch/functional.py:648):
 * 1 + (length + 2*pad - n_fft) / hop, or TTTS_MEL_EINVAL when length <= pad (reflect needs it). */
int64_t ttts_mel_frames(const ttts_mel_plan* plan, int64_t length);

/* Plan queries: the bin range the plan computes (rows of the filterbank with a non-zero weight: bins outside it are
 * never evaluated). */
int32_t ttts_mel_first_bin(const ttts_mel_plan* plan);
int32_t ttts_mel_num_bins(const ttts_mel_plan* plan);   /* padded bin count                                 */
int32_t ttts_mel_live_bins(const ttts_mel_plan* plan);  /* bins with a non-zero filterbank row span          */

/* Bytes of device workspace ttts_mel_forward needs for a batch of B clips of at most Lmax samples. */
size_t ttts_mel_workspace_bytes(const ttts_mel_plan* plan, int32_t B, int64_t Lmax);

/* The same for one particular input: when its rows are 16-byte addressable (base and row_stride * 4 multiples of 16)
 * the kernel reads the waveform in place and the padded copy -- as large as the batch -- is not part of the workspace.
 * ttts_mel_forward accepts either size for such an input. */
size_t ttts_mel_workspace_bytes_for(const ttts_mel_plan* plan, int32_t B, int64_t Lmax, const float* wav_dev,
                                    int64_t row_stride);

/* Replaces TorchMelSpectrogram.forward / MelSpectrogramFeatures.forward
 * (ttts/utils/utils.py:509-522, ttts/vocoder/feature_extractors.py:93-99) on a batch.
 *   wav_dev      [B, row_stride] fp32, clip b occupies wav_dev[b*row_stride .. +length_b)
 *   lengths_dev  [B] int64 device pointer or NULL (all clips are Lmax long).  With lengths, each
 *                clip is reflect-padded at its own end (what the reference does: one clip per
 *                call, ttts/prepare/mel_extract.py:15-27) and frames t >= n_frames[b] are 0.0
 *                (collate contract, ttts/gpt/dataset.py:116-117).
 *   Tmax         time extent of the output (>= frames of the longest clip)
 *   mel_out_dev  [B, Tmax, n_mels] fp32 -- the PHYSICAL layout of the reference's result, which
 *                is a transposed view [B, n_mels, Tmax] with strides (Tmax*n_mels, 1, n_mels)
 *   n_frames_dev [B] int32 out (may be NULL)
 *   mask_dev     [B, Tmax] uint8 out or NULL: make_pad_mask, utils.py:77-103 (1 = padding)
 * Asynchronous on `stream`. */
int ttts_mel_forward(const ttts_mel_plan* plan, const float* wav_dev, int64_t row_stride,
                     const int64_t* lengths_dev_or_null, int32_t B, int64_t Lmax, int32_t Tmax,
                     float* mel_out_dev, int32_t* n_frames_dev_or_null, uint8_t* mask_dev_or_null,
                     void* workspace_dev, size_t workspace_bytes, void* stream);

/* ttts_mel_forward with stage (4) fused into the kernel's log pass: the per-channel sufficient statistics of the
 * UN-normalised log-mel of this batch's valid frames are added to acc_dev[2*n_mels+1] fp64 = [sum_m | sumsq_m | frames]
 * (same meaning as ttts_mel_stats_accumulate, without the extra pass over HBM).  Inside the launch the sums are kept
 * in 64-bit fixed point (steps of 2^-32 for the sums and 2^-30 for the sums of squares), so they are bit-identical from run to run and independent of
 * how a corpus is cut into batches or sharded over GPUs up to the final fp64 additions.  acc_dev is NOT zeroed. */
int ttts_mel_forward_stats(const ttts_mel_plan* plan, const float* wav_dev, int64_t row_stride,
                           const int64_t* lengths_dev_or_null, int32_t B, int64_t Lmax, int32_t Tmax,
                           float* mel_out_dev, int32_t* n_frames_dev_or_null, uint8_t* mask_dev_or_null,
                           void* workspace_dev, size_t workspace_bytes, void* stream, double* acc_dev);

/* ttts_mel_forward without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The workspace of the call (status word, frame / group offsets, frame counts,
 * the reflect-padded edge samples and, for inputs the kernel cannot read in place, the padded copy) lives in ONE
 * file-scope `static __device__` array of TTTS_MEL_STATIC_WS_BYTES that every GPU loading the library holds until the
 * process exits.  The call makes no cudaMalloc / cudaFree / allocating library call; the plan must exist already.
 *   - a batch whose workspace (ttts_mel_workspace_bytes_for) exceeds TTTS_MEL_STATIC_WS_BYTES is rejected with
 *     TTTS_MEL_ENOMEM and a message naming both sizes; use ttts_mel_forward with a caller workspace for those
 *   - every call on a device shares the array and finds the previous call's contents in it; k_prep rewrites every word
 *     the kernels read (status, offsets, counts, edges) at the start of each call, so nothing has to be cleared by the
 *     caller -- but two calls that OVERLAP on the GPU (different streams, or graph branches) race on it and are not
 *     supported: serialise them on one stream
 *   - workspace_out_or_null receives the device address of the array (after the library resolved it with
 *     cudaGetSymbolAddress on the current device), which is what ttts_mel_status() takes
 * All other arguments and the results are those of ttts_mel_forward, bit for bit. */
#define TTTS_MEL_STATIC_WS_BYTES (8u << 20)
int ttts_mel_forward_static(const ttts_mel_plan* plan, const float* wav_dev, int64_t row_stride,
                            const int64_t* lengths_dev_or_null, int32_t B, int64_t Lmax, int32_t Tmax,
                            float* mel_out_dev, int32_t* n_frames_dev_or_null, uint8_t* mask_dev_or_null,
                            void* stream, const void** workspace_out_or_null);

/* sizeof(ttts_mel_config) as the library was built: a binding asserts its own struct against it. */
size_t ttts_mel_config_size(void);

/* Synchronises `stream` and returns the device-side status word of the workspace's last call:
 * 0, or TTTS_MEL_EDEVICE with the reason in ttts_mel_last_error(). */
int ttts_mel_status(const void* workspace_dev, void* stream);

/* Stage (4): per-channel sufficient statistics of UN-normalised log-mel over valid frames,
 * accumulated into acc_dev[2*n_mels+1] fp64 = [sum_m | sumsq_m | frame count].  The reference
 * only loads a finished mel_norms tensor (utils.py:503-507); the structural template is the EMA
 * codebook all_reduce in ttts/vqvae/xtts_dvae.py:104-110 (sum local statistics, allreduce, then
 * normalise).  acc_dev is NOT zeroed by the call (so batches accumulate).
 *   logmel_dev [B, Tmax, n_mels] fp32 as written by ttts_mel_forward with norms == NULL */
int ttts_mel_stats_accumulate(const ttts_mel_plan* plan, const float* logmel_dev,
                              const int32_t* n_frames_dev, int32_t B, int32_t Tmax,
                              double* acc_dev, void* stream);

/* Kernel launches the last ttts_mel_forward* call on this thread enqueued (for bench accounting). */
int32_t ttts_mel_last_launch_count(void);

const char* ttts_mel_last_error(void);
const char* ttts_mel_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TTTS_MEL_H_ */
