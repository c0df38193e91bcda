/*
 * ttts_ingest.h -- C ABI of the ingestion step in front of the log-mel path, same library
 * (libttts_mel.so), same conventions as ttts_mel.h: plain pointers and sizes, caller-owned device buffers,
 * `stream` is a cudaStream_t passed as void*, 0 / negative TTTS_MEL_E* return codes, ttts_mel_last_error().
 *
 * Reference: `load_audio` (ttts/utils/utils.py:32-47) and the loop body of ttts/prepare/mel_extract.py:15-21:
 *     audio, sr = torchaudio.load(path)            -> int16 PCM / 32768 as float32 [C, L]      (decode: host I/O, not here)
 *     audio = audio[0].unsqueeze(0)                -> channel 0                                  ttts_pcm16_to_f32
 *     torchaudio.functional.resample(audio, sr, sampling_rate)   (sinc_interp_hann, width 6, rolloff 0.99)   ttts_resample_forward
 *     audio.clip_(-1, 1)                                                                          ttts_resample_forward(clip = 1)
 */
#ifndef TTTS_INGEST_H_
#define TTTS_INGEST_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Interleaved int16 PCM (the sample order of a WAV file) -> float32 channel 0, x / 32768 (what torchaudio.load
 * returns for 16-bit PCM with normalize=True), utils.py:33-37.
 *   pcm_dev   [B, in_row_stride] int16, clip b holds frames_b * channels samples (frame-major, channel-minor)
 *   frames    frames of the longest clip; out[b, frames_b .. frames) is zero-filled when frames_dev is given
 *   out_dev   [B, out_row_stride] float32 */
int ttts_pcm16_to_f32(const int16_t* pcm_dev, int64_t in_row_stride, int32_t channels,
                      const int64_t* frames_dev_or_null, int32_t B, int64_t frames,
                      float* out_dev, int64_t out_row_stride, void* stream);

typedef struct ttts_resample_plan ttts_resample_plan;

/* Replaces torchaudio.functional.resample's kernel construction + conv1d (reference call site utils.py:41;
 * torchaudio/functional/functional.py `_get_sinc_resample_kernel`, `_apply_sinc_resample_kernel`).
 *   taps_host [new_freq, 2*width + orig_freq] float32: the polyphase FIR exactly as torchaudio builds it
 *             (the host mirror restates those torch ops; passed in like the filterbank of ttts_mel_plan_create)
 *   orig_freq, new_freq: the two rates divided by their gcd */
int ttts_resample_plan_create(const float* taps_host, int32_t orig_freq, int32_t new_freq, int32_t width,
                              ttts_resample_plan** out_plan);
void ttts_resample_plan_destroy(ttts_resample_plan* plan);

/* ceil(new_freq * length / orig_freq): the length torchaudio keeps (`target_length`). */
int64_t ttts_resample_out_len(const ttts_resample_plan* plan, int64_t length);

/* out[b, j] = sum_k xpad_b[(j / new) * orig + k] * taps[j % new][k], xpad_b = clip b zero-padded by `width` on the
 * left (and by zeros on the right), for j < out_len(length_b); zeros beyond.  clip != 0 clamps to [-1, 1]
 * (utils.py:46).  Asynchronous on `stream`. */
int ttts_resample_forward(const ttts_resample_plan* plan, const float* x_dev, int64_t row_stride,
                          const int64_t* lengths_dev_or_null, int32_t B, int64_t L,
                          float* out_dev, int64_t out_row_stride, int32_t clip, void* stream);

/* The two steps fused: the FIR reads channel 0 of the interleaved int16 PCM directly (x / 32768 while staging), so
 * the float32 waveform at the source rate never exists in HBM.  Same result as ttts_pcm16_to_f32 followed by
 * ttts_resample_forward, bit for bit.  pcm_dev [B, in_row_stride] int16, `frames` per-clip valid frames or NULL. */
int ttts_resample_pcm16_forward(const ttts_resample_plan* plan, const int16_t* pcm_dev, int64_t in_row_stride,
                                int32_t channels, const int64_t* frames_dev_or_null, int32_t B, int64_t F,
                                float* out_dev, int64_t out_row_stride, int32_t clip, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TTTS_INGEST_H_ */
