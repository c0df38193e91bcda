/*
 * ttts_stft.h -- C ABI of the STFT-family step behind the log-mel path (first slice; forward only),
 * same library and conventions as ttts_mel.h (plain pointers and sizes, caller-owned device buffers, `stream` is a
 * cudaStream_t passed as void*, 0 / negative TTTS_MEL_E* codes, ttts_mel_last_error()).
 *
 * Reference:
 *   TorchSTFT.__call__          ttts/hifigan/losses.py:107-148  (STFTLoss :156-172, MultiScaleSTFTLoss :196-218: the shapes
 *                               (n_fft, hop, win) = (1024, 120, 600), (2048, 240, 1200), (512, 50, 240) at :201)
 *   ISTFT.forward               ttts/vocoder/spectral_ops.py:32-60 (padding "same"; "center" = torch.istft, same sums)
 *
 * The dense contraction itself (frames x windowed DFT basis, spectrum x windowed inverse basis) is
 * ttts_resample_forward (include/ttts_ingest.h): [strides x taps] x [taps x phases] on tcgen05 with the 3xTF32 split.
 * For the forward transform stride = hop, taps = win_length, phases = (re, im) of every bin; for the inverse
 * stride = taps = 2 * bins, phases = n_fft.  This header holds the steps around it.
 */
#ifndef TTTS_STFT_H_
#define TTTS_STFT_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* torch.nn.functional.pad(x, (pad, pad), mode="reflect") per clip at its OWN length (torch/functional.py:675-680 inside
 * torch.stft(center=True); losses.py:124-126 for pad_wav): xp[b, 0 .. L_b + 2 pad) = padded clip b, zeros beyond.
 *   wav_dev [B, row_stride], lengths_dev [B] int64 or NULL, xp_dev [B, Lp] with Lp % 4 == 0, Lp >= Lmax + 2 pad, pad % 4 == 0 */
int ttts_reflect_pad(const float* wav_dev, int64_t row_stride, const int64_t* lengths_dev_or_null, int32_t B,
                     int64_t Lmax, int32_t pad, float* xp_dev, int64_t Lp, void* stream);

/* losses.py:141-148: S = sqrt(clamp(re^2 + im^2, min=clamp)) [** power] [log(clamp(S, 1e-5) * db_gain)], written as the
 * reference lays it out, [B, bins, T].
 *   y_dev   [B, y_stride] rows of n_cols floats as the contraction wrote them: re of bin k at column 2k, im at 2k + 1;
 *           frame t of clip b is row row0 + t
 *   power   <= 0: none (the reference's power=None) */
int ttts_stft_magnitude(const float* y_dev, int64_t y_stride, int32_t row0, int32_t n_cols, int32_t B, int32_t T,
                        int32_t nbins, float clamp, float power, int32_t to_db, float db_gain, float* out_dev,
                        void* stream);

/* The rest of TorchMelSpectrogram.forward (ttts/utils/utils.py:509-522) for transform sizes the fused kernel of
 * ttts_mel.h is not built for (any filter_length / hop_length / win_length the reference's constructor takes,
 * utils.py:487-507): power -> matmul(spec^T, fb) (torchaudio _transforms.py:407-418) -> log(clamp(., clamp)) [/ mel_norms],
 * written [B, Tmax, n_mels] like ttts_mel_forward.  Frames t >= n_frames[b] (or >= T) are zeros.
 *   y_dev       rows as ttts_stft_magnitude reads them (re of bin k at column 2k, im at 2k + 1; frame t = row row0 + t)
 *   fbT_dev     [n_mels, nbins] float32: the reference's filterbank, transposed
 *   krange_dev  [n_mels, 2] int32: first / last bin with a non-zero weight of each filter (first > last: empty filter)
 *   power       1 (sqrt(re^2 + im^2 + power_eps)) or 2 */
int ttts_stft_logmel(const float* y_dev, int64_t y_stride, int32_t row0, int32_t n_cols, int32_t B, int32_t T,
                     const int32_t* n_frames_dev_or_null, int32_t Tmax, int32_t nbins, const float* fbT_dev,
                     const int32_t* krange_dev, int32_t n_mels, float power, float power_eps, float clamp,
                     const float* norms_dev_or_null, float* out_dev, void* stream);

/* spectral_ops.py:49-52 input side: complex spectrogram [B, N, T] (complex64 = float pairs) -> [B, T, 2N] rows
 * (re of bin k at column k, im at column N + k), the row layout the inverse contraction strides over. */
int ttts_spec_pack(const float* spec_ri_dev, int32_t B, int32_t N, int32_t T, float* x_dev, void* stream);

/* spectral_ops.py:53-69: overlap-add of the windowed inverse-DFT frames, division by the window envelope, trim:
 *   out[b, i] = sum_t v[b, t, i + pad - t hop] / sum_t w^2[i + pad - t hop],  i < (T - 1) hop + n_fft - 2 pad
 * (`fold`, `fold` of the squared window and the divide in one pass).  status_dev[0] = 3 if the envelope is <= 1e-11
 * anywhere (the reference asserts).  window_dev [n_fft] float32. */
int ttts_istft_overlap_add(const float* v_dev, int32_t B, int32_t T, int32_t n_fft, int32_t hop, int32_t pad,
                           const float* window_dev, float* out_dev, int64_t out_stride, int32_t* status_dev,
                           void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TTTS_STFT_H_ */
