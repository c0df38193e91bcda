// k_stft.cuh -- SURVEY 8f N4, first slice: the STFT family on the library's dense-contraction machinery
// (included by ttts_mel.cu after k_ingest.cuh; C ABI in include/ttts_stft.h).
//
//   forward   TorchSTFT.__call__ (ttts/hifigan/losses.py:107-148; the three MultiScaleSTFTLoss shapes at :201 have hop !=
//             n_fft / 4 and win < n_fft, which k_logmel2's slot sharing cannot express): frames x windowed-DFT basis is the
//             [strides x taps] x [taps x phases] product k_resample_tc already computes on tcgen05 with 3xTF32 -- stride =
//             hop, taps = win_length, phases = (re, im) of every bin -- behind a per-clip reflect pad (k_pad) and in front
//             of k_stft_mag: sqrt(clamp(re^2 + im^2, 1e-8)) [^power] [log(clamp(., 1e-5) * gain)], transposed to [B, bins, T]
//   inverse   Vocos ISTFT (ttts/vocoder/spectral_ops.py:7-60): k_spec_pack ([B, bins, T] complex -> [B, T, re | im]),
//             the same product with stride = taps = 2 * bins and phases = n_fft (windowed irfft basis), then k_istft_ola:
//             overlap-add, division by the window envelope, trim -- fold / fold / divide of the reference in one pass
#pragma once

namespace stft {

// y [B][y_stride] holds rows of n_cols floats (re of bin k at column 2k, im at 2k+1); frame t of clip b is row row0 + t.
// out[b][k][t] = magnitude; 32 x 32 tiles through shared memory: coalesced reads along columns, coalesced writes along t.
__global__ void __launch_bounds__(256) k_stft_mag(const float* __restrict__ y, long long y_stride, int row0, int n_cols,
                                                  int T, int nbins, float clamp, float power, int to_db, float db_gain,
                                                  float* __restrict__ out) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;       // 32 x 8
  const float* yb = y + (long long)b * y_stride;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int t = t0 + ty + 8 * r, k = k0 + tx;
    float m = 0.f;
    if (t < T && k < nbins) {
      const float2 v = *reinterpret_cast<const float2*>(yb + (long long)(row0 + t) * n_cols + 2 * k);
      m = sqrtf(fmaxf(fmaf(v.x, v.x, v.y * v.y), clamp));
      if (power > 0.f && power != 1.f) m = (power == 2.f) ? m * m : powf(m, power);
      if (to_db) m = logf(fmaxf(m, 1e-5f) * db_gain);
    }
    tile[ty + 8 * r][tx] = m;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + 8 * r, t = t0 + tx;
    if (k < nbins && t < T) out[((long long)b * nbins + k) * T + t] = tile[tx][ty + 8 * r];
  }
}

// General-shape log-mel behind the same contraction (TorchMelSpectrogram with filter_length / hop_length other than
// 1024 / 256, ttts/utils/utils.py:487-507: torchaudio MelSpectrogram takes any n_fft / hop / win; the fused kernel
// k_logmel2 is built for hop = n_fft / 4 = 256 only).  y rows as k_stft_mag reads them; per block kFR frames:
//   phase 1  power (|X|^2, or sqrt(|X|^2 + eps) for the magnitude plans) of every bin -> shared memory, coalesced float2 loads
//   phase 2  one thread per (frame, mel): sum over the filter's own bin range [k_lo, k_hi] of fbT[m][k] * P[k] (ascending k),
//            log(max(., clamp)) [/ norms[m]] -> out[b][t][m]; frames at or past n_frames[b] are zeros (the collate contract)
template <int kFR>
__global__ void __launch_bounds__(256) k_stft_logmel(const float* __restrict__ y, long long y_stride, int row0, int n_cols,
                                                     const int32_t* __restrict__ n_frames, int T, int Tmax, int nbins,
                                                     const float* __restrict__ fbT, const int2* __restrict__ krange,
                                                     int n_mels, int power1, float eps, float clamp,
                                                     const float* __restrict__ norms, float* __restrict__ out) {
  extern __shared__ float s_pw[];                               // [kFR][nbins + 1]
  const int pitch = nbins + 1;
  const int b = blockIdx.y, t0 = blockIdx.x * kFR;
  const int nf = n_frames ? min(n_frames[b], T) : T;
  const float* yb = y + (long long)b * y_stride;
  for (int r = 0; r < kFR; ++r) {
    const int t = t0 + r;
    if (t >= nf) break;                                         // block-uniform
    const float* row = yb + (long long)(row0 + t) * n_cols;
    for (int k = threadIdx.x; k < nbins; k += 256) {
      const float2 v = *reinterpret_cast<const float2*>(row + 2 * k);
      float pw = fmaf(v.x, v.x, v.y * v.y);
      if (power1) pw = sqrtf(pw + eps);
      s_pw[r * pitch + k] = pw;
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < kFR * n_mels; idx += 256) {
    const int r = idx / n_mels, m = idx - r * n_mels, t = t0 + r;
    if (t >= Tmax) continue;
    float v = 0.f;
    if (t < nf) {
      const int2 kr = krange[m];
      const float* w = fbT + (long long)m * nbins;
      const float* pw = s_pw + r * pitch;
      float acc = 0.f;
      for (int k = kr.x; k <= kr.y; ++k) acc = fmaf(__ldg(w + k), pw[k], acc);
      v = logf(acc < clamp ? clamp : acc);               // torch.clamp keeps NaN (fmaxf would not)
      if (norms) v = v / norms[m];
    }
    out[((long long)b * Tmax + t) * n_mels + m] = v;
  }
}

// spec [B][N][T] complex64 (float pairs) -> x [B][T][2N]: re of bin k at column k, im at column N + k
__global__ void __launch_bounds__(256) k_spec_pack(const float2* __restrict__ spec, int N, int T, float* __restrict__ x) {
  __shared__ float2 tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int k = k0 + ty + 8 * r, t = t0 + tx;
    tile[ty + 8 * r][tx] = (k < N && t < T) ? spec[((long long)b * N + k) * T + t] : make_float2(0.f, 0.f);
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int t = t0 + ty + 8 * r, k = k0 + tx;
    if (t < T && k < N) {
      float* row = x + ((long long)b * T + t) * (2 * N);
      const float2 v = tile[tx][ty + 8 * r];
      row[k] = v.x;
      row[N + k] = v.y;
    }
  }
}

// v [B][T][n_fft] windowed inverse-DFT frames -> out[b][i] = sum_t v[b][t][i + pad - t hop] / sum_t w^2[i + pad - t hop]
__global__ void __launch_bounds__(256) k_istft_ola(const float* __restrict__ v, int T, int n_fft, int hop, int pad,
                                                   const float* __restrict__ win, long long out_len,
                                                   float* __restrict__ out, long long out_stride, int32_t* status) {
  const int b = blockIdx.y;
  const float* vb = v + (long long)b * T * n_fft;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < out_len; i += (long long)gridDim.x * blockDim.x) {
    const long long j = i + pad;
    long long t_hi = j / hop, t_lo = (j - n_fft + hop) / hop;      // frames with 0 <= j - t hop < n_fft
    if (t_lo < 0) t_lo = 0;
    if (t_hi > T - 1) t_hi = T - 1;
    float acc = 0.f, env = 0.f;
    for (long long t = t_lo; t <= t_hi; ++t) {
      const int n = (int)(j - t * hop);
      const float w = __ldg(win + n);
      acc += vb[t * n_fft + n];
      env = fmaf(w, w, env);
    }
    if (!(env > 1e-11f)) atomicExch(status, 3);                     // the reference asserts window_envelope > 1e-11
    out[(long long)b * out_stride + i] = acc / env;
  }
}

}  // namespace stft

extern "C" int ttts_reflect_pad(const float* wav, int64_t row_stride, const int64_t* lengths, int32_t B, int64_t Lmax,
                                int32_t pad, float* xp, int64_t Lp, void* stream) {
  if (!wav || !xp) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || Lmax <= pad || pad < 0 || pad % 4 != 0 || Lp < Lmax + 2 * (int64_t)pad || Lp % 4 != 0)
    return fail(TTTS_MEL_EINVAL, "reflect pad: need B > 0, Lmax > pad, pad %% 4 == 0, Lp %% 4 == 0 and Lp >= Lmax + 2 pad");
  const int vec_ok = (row_stride % 4 == 0) && (((uintptr_t)wav & 15) == 0);
  const int64_t vecs = (Lp + 3) / 4;
  const int64_t per_clip = std::min<int64_t>((vecs + 255) / 256, B >= 64 ? 48 : (B >= 8 ? 256 : 4096));
  // k_pad writes each clip's own L + 2 pad samples; the tail of the row (what frames past the clip's end read) is zeroed first
  CUDA_TRY(cudaMemsetAsync(xp, 0, (size_t)B * Lp * sizeof(float), (cudaStream_t)stream));
  k_pad<<<dim3((unsigned)per_clip, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(wav, row_stride, lengths, Lmax, pad, xp, Lp, vec_ok);
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}

extern "C" int ttts_stft_magnitude(const float* y, int64_t y_stride, int32_t row0, int32_t n_cols, int32_t B, int32_t T,
                                   int32_t nbins, float clamp, float power, int32_t to_db, float db_gain, float* out,
                                   void* stream) {
  if (!y || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || T <= 0 || nbins <= 0 || n_cols < 2 * nbins || (n_cols & 1) || row0 < 0 || (((uintptr_t)y | (uintptr_t)(y_stride * 4)) & 7))
    return fail(TTTS_MEL_EINVAL, "stft magnitude: bad geometry");
  dim3 grid((unsigned)((T + 31) / 32), (unsigned)((nbins + 31) / 32), (unsigned)B);
  stft::k_stft_mag<<<grid, 256, 0, (cudaStream_t)stream>>>(y, y_stride, row0, n_cols, T, nbins, clamp, power, to_db, db_gain, out);
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}

extern "C" int ttts_stft_logmel(const float* y, int64_t y_stride, int32_t row0, int32_t n_cols, int32_t B, int32_t T,
                                const int32_t* n_frames, int32_t Tmax, int32_t nbins, const float* fbT,
                                const int32_t* krange, int32_t n_mels, float power, float power_eps, float clamp,
                                const float* norms, float* out, void* stream) {
  if (!y || !fbT || !krange || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || T <= 0 || Tmax < T || nbins <= 0 || n_mels <= 0 || n_cols < 2 * nbins || (n_cols & 1) || row0 < 0 ||
      (((uintptr_t)y | (uintptr_t)(y_stride * 4) | (uintptr_t)krange) & 7))
    return fail(TTTS_MEL_EINVAL, "stft log-mel: bad geometry");
  if (!(power == 1.0f || power == 2.0f) || !(clamp > 0.f) || !(power_eps >= 0.f))
    return fail(TTTS_MEL_EINVAL, "stft log-mel: power must be 1 or 2, clamp > 0, power_eps >= 0");
  constexpr int kFR = 8;
  const size_t smem = (size_t)kFR * (nbins + 1) * sizeof(float);
  if (smem > 200 * 1024) return fail(TTTS_MEL_EINVAL, "stft log-mel: %d bins do not fit in shared memory", nbins);
  if (smem > 48 * 1024) {
    static std::atomic<size_t> granted[64];
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || granted[dev].load() < smem) {
      CUDA_TRY(cudaFuncSetAttribute(stft::k_stft_logmel<kFR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      if (dev >= 0 && dev < 64) granted[dev].store(smem);
    }
  }
  dim3 grid((unsigned)((Tmax + kFR - 1) / kFR), (unsigned)B);
  stft::k_stft_logmel<kFR><<<grid, 256, smem, (cudaStream_t)stream>>>(y, y_stride, row0, n_cols, n_frames, T, Tmax, nbins, fbT,
                                                                    reinterpret_cast<const int2*>(krange), n_mels,
                                                                    power == 1.0f ? 1 : 0, power_eps, clamp, norms, out);
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}

extern "C" int ttts_spec_pack(const float* spec_ri, int32_t B, int32_t N, int32_t T, float* x, void* stream) {
  if (!spec_ri || !x) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || N <= 0 || T <= 0 || ((uintptr_t)spec_ri & 7)) return fail(TTTS_MEL_EINVAL, "spec pack: bad geometry");
  dim3 grid((unsigned)((T + 31) / 32), (unsigned)((N + 31) / 32), (unsigned)B);
  stft::k_spec_pack<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float2*>(spec_ri), N, T, x);
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}

extern "C" int ttts_istft_overlap_add(const float* v, int32_t B, int32_t T, int32_t n_fft, int32_t hop, int32_t pad,
                                      const float* window_dev, float* out, int64_t out_stride, int32_t* status_dev,
                                      void* stream) {
  if (!v || !window_dev || !out || !status_dev) return fail(TTTS_MEL_EINVAL, "null argument");
  const int64_t out_len = (int64_t)(T - 1) * hop + n_fft - 2 * (int64_t)pad;
  if (B <= 0 || T <= 0 || hop <= 0 || n_fft < hop || pad < 0 || out_len <= 0 || out_stride < out_len)
    return fail(TTTS_MEL_EINVAL, "istft overlap-add: bad geometry");
  CUDA_TRY(cudaMemsetAsync(status_dev, 0, sizeof(int32_t), (cudaStream_t)stream));
  const unsigned gx = (unsigned)std::min<int64_t>((out_len + 255) / 256, 4096);
  stft::k_istft_ola<<<dim3(gx, (unsigned)B), 256, 0, (cudaStream_t)stream>>>(v, T, n_fft, hop, pad, window_dev, out_len, out,
                                                                           out_stride, status_dev);
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}
