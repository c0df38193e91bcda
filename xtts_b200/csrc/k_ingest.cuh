// k_ingest.cuh -- ingestion in front of the log-mel path (C ABI in include/ttts_ingest.h):
//   k_pcm16      interleaved int16 PCM -> float32 channel 0 (x / 32768)        HBM-bound: 2 C + 4 bytes per frame
//   k_decimate   integer-ratio (2:1, 3:1, 4:1) windowed-sinc FIR, one phase     HBM-bound: 4 (or 2 C) ORIG + 4 bytes per output
//   k_resample_tc general polyphase windowed-sinc FIR (orig -> new, >= 16 phases) on tcgen05 (k_resample_tc.cuh)
//   k_resample_few  fewer than 16 phases (2:3, 1:3, 4:3, 6:1 ...): flat staged segment, one output per thread
//   k_resample   many phases on the FMA pipe (fallback of k_resample_tc)        FMA-bound: 2 (2 width + orig) flop per output
// The two FIR kernels read float32 or (fused decode) interleaved int16 PCM.
// Reference: load_audio (ttts/utils/utils.py:32-47), mel_extract.py:15-21.  Included at the end of ttts_mel.cu
// (one shared library); uses its fail() / CUDA_TRY / g_launches.
#pragma once

namespace ingest {

constexpr int kThreads = 256;
constexpr float kPcmScale = 1.0f / 32768.0f;

// sample s of a clip: float32 as is, int16 -> channel 0 of frame s, scaled
__device__ __forceinline__ float load_sample(const float* src, long long s, int) { return __ldg(src + s); }
// int16 -> x / 32768 without the conversion pipe (I2F runs at a quarter of the ALU rate): the sample in offset
// binary is dropped into the mantissa of 2^23, and one FMA scales and removes the offset -- exact:
//   (2^23 + x + 32768) * 2^-15 - 257 = x / 32768
__device__ __forceinline__ float pcm16_to_float(uint32_t low16) {
  return fmaf(__uint_as_float(((low16 ^ 0x8000u) & 0xFFFFu) | 0x4B000000u), kPcmScale, -257.0f);
}
__device__ __forceinline__ float load_sample(const int16_t* src, long long s, int channels) {
  return pcm16_to_float(__ldg(reinterpret_cast<const unsigned short*>(src) + s * channels));
}

// ---------------------------------------------------------------------------------------------- k_pcm16
// One thread = 8 consecutive frames of one clip.  kC = 1: one 16-byte load; kC = 2: two; kC = 0: scalar loads.
template <int kC>
__global__ void __launch_bounds__(kThreads) k_pcm16(const int16_t* __restrict__ pcm, long long in_stride, int channels,
                                                    const long long* __restrict__ frames_b, long long frames,
                                                    float* __restrict__ out, long long out_stride, int vec_ok) {
  const int b = blockIdx.y;
  const long long nf = frames_b ? min(max(frames_b[b], 0LL), frames) : frames;
  const int16_t* src = pcm + (long long)b * in_stride;
  float* dst = out + (long long)b * out_stride;
  for (long long f0 = 8LL * (blockIdx.x * (long long)kThreads + threadIdx.x); f0 < frames;
       f0 += 8LL * kThreads * gridDim.x) {
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = 0.f;
    if (f0 + 8 <= nf && vec_ok && kC == 1) {
      const int4 s = __ldg(reinterpret_cast<const int4*>(src + f0));
      const int w[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        v[2 * j] = pcm16_to_float((uint32_t)w[j]);
        v[2 * j + 1] = pcm16_to_float((uint32_t)w[j] >> 16);
      }
    } else if (f0 + 8 <= nf && vec_ok && kC == 2) {
      const int4 s0 = __ldg(reinterpret_cast<const int4*>(src + 2 * f0));       // 4 frames x (ch0, ch1)
      const int4 s1 = __ldg(reinterpret_cast<const int4*>(src + 2 * f0 + 8));
      const int w[8] = {s0.x, s0.y, s0.z, s0.w, s1.x, s1.y, s1.z, s1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = pcm16_to_float((uint32_t)w[j]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f0 + j < nf) v[j] = load_sample(src, f0 + j, channels);
    }
    if (f0 + 8 <= frames && vec_ok) {
      reinterpret_cast<float4*>(dst + f0)[0] = make_float4(v[0], v[1], v[2], v[3]);
      reinterpret_cast<float4*>(dst + f0)[1] = make_float4(v[4], v[5], v[6], v[7]);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (f0 + j < frames) dst[f0 + j] = v[j];
    }
  }
}

// ---------------------------------------------------------------------------------------------- k_decimate
struct TapsParam { float t[64]; };     // by value: the FMAs read the taps straight from the constant bank

constexpr int kDecR = 8;               // consecutive outputs per thread
constexpr int kDecTile = kThreads * kDecR;
// shared-memory index with 4 floats of padding per 16: a thread's window starts 8 ORIG floats after its
// neighbour's, and the padding spreads the 16-byte chunks of 8 neighbouring lanes over distinct banks
__device__ __forceinline__ int dec_pad(int i) { return i + ((i >> 4) << 2); }

// four consecutive samples s .. s+3 (s % 4 == 0, all inside the clip, row base suitably aligned), split in a raw
// 16-byte load and a conversion so that a thread can issue all its loads before it touches the first result
__device__ __forceinline__ int4 raw_quad(const float* src, long long s, int) {
  return __ldg(reinterpret_cast<const int4*>(src + s));
}
__device__ __forceinline__ int4 raw_quad(const int16_t* src, long long s, int channels) {
  if (channels == 1) {
    const int2 w = __ldg(reinterpret_cast<const int2*>(src + s));
    return make_int4(w.x, w.y, 0, 0);
  }
  return __ldg(reinterpret_cast<const int4*>(src + 2 * s));          // channels == 2: 4 frames x (ch0, ch1)
}
__device__ __forceinline__ float4 quad_to_float(const float*, int4 w, int) {
  return make_float4(__int_as_float(w.x), __int_as_float(w.y), __int_as_float(w.z), __int_as_float(w.w));
}
__device__ __forceinline__ float4 quad_to_float(const int16_t*, int4 w, int channels) {
  if (channels == 1)
    return make_float4(pcm16_to_float((uint32_t)w.x), pcm16_to_float((uint32_t)w.x >> 16),
                       pcm16_to_float((uint32_t)w.y), pcm16_to_float((uint32_t)w.y >> 16));
  return make_float4(pcm16_to_float((uint32_t)w.x), pcm16_to_float((uint32_t)w.y),
                     pcm16_to_float((uint32_t)w.z), pcm16_to_float((uint32_t)w.w));
}

// (Round 2, tried and dropped -- profiles/r2p_decimator_variants.txt, 64 x 10 s 44.1 -> 22.05 kHz: this kernel 46 us fp32 /
//  53 us fused stereo PCM = 3.7 / 3.2 TB/s; a persistent CTA with a two-deep cp.async ring 51 / 58 us (fewer resident warps
//  overlap less than eight independent CTAs per SM do); packed fma.rn.f32x2 taps 52 / 66 us (FFMA2 issues at half rate).)
// new == 1: out[j] = sum_k xpad[j * ORIG + k] * taps[k].  One CTA = 2048 consecutive outputs of one clip: the
// segment is staged once in shared memory (16-byte loads and stores: it starts at the multiple of 4 at or below
// the first sample the tile needs, kOff samples early), each thread pulls the samples its 8 outputs share into
// registers with 16-byte loads, and the R x KW FMAs take their taps from the kernel-parameter constant bank.
template <int ORIG, int KW, typename TIn, bool kHoist>
__global__ void __launch_bounds__(kThreads) k_decimate(const TIn* __restrict__ x, long long row_stride, int channels,
                                                       const long long* __restrict__ lengths, long long L,
                                                       const TapsParam taps, float* __restrict__ out,
                                                       long long out_stride, long long out_cols, int clip, int vec) {
  constexpr int kWidth = (KW - ORIG) / 2;
  constexpr int kOff = (4 - (kWidth & 3)) & 3;
  constexpr int kWin = kOff + (kDecR - 1) * ORIG + KW;
  constexpr int kWinQ = (kWin + 3) / 4;
  constexpr int kSeg = (kThreads - 1) * kDecR * ORIG + kWinQ * 4;
  static_assert((kDecTile * ORIG) % 4 == 0 && kSeg % 4 == 0, "tile origin must stay 4-aligned");
  __shared__ __align__(16) float xs[kSeg + (kSeg >> 4) * 4 + 4];
  const int b = blockIdx.y;
  const long long len = lengths ? min(max(lengths[b], 0LL), L) : L;
  const long long olen = (len + ORIG - 1) / ORIG;
  const long long j0 = (long long)blockIdx.x * kDecTile;
  float* dst = out + (long long)b * out_stride;
  const long long jend = min(j0 + kDecTile, out_cols);
  if (j0 >= olen) {
    for (long long j = j0 + threadIdx.x; j < jend; j += kThreads) dst[j] = 0.f;
    return;
  }
  const TIn* src = x + (long long)b * row_stride;
  const long long s0 = j0 * ORIG - kWidth - kOff;          // multiple of 4 (negative for the first tile)
  constexpr int kQuads = kSeg / 4, kIt = (kQuads + kThreads - 1) / kThreads;

  int4 raw[kIt];
#pragma unroll
  for (int it = 0; it < kIt; ++it) {                        // all the 16-byte loads first ...
    const int q = threadIdx.x + it * kThreads;
    const long long s = s0 + 4 * q;
    raw[it] = make_int4(0, 0, 0, 0);
    if (kHoist && vec && q < kQuads && s >= 0 && s + 4 <= len) raw[it] = raw_quad(src, s, channels);
  }
#pragma unroll
  for (int it = 0; it < kIt; ++it) {                        // ... then conversion and the padded stores
    const int q = threadIdx.x + it * kThreads;
    const long long s = s0 + 4 * q;
    if (q >= kQuads) break;
    float4 v;
    if (vec && s >= 0 && s + 4 <= len) {
      v = quad_to_float(src, kHoist ? raw[it] : raw_quad(src, s, channels), channels);
    } else {
      v.x = (s >= 0 && s < len) ? load_sample(src, s, channels) : 0.f;
      v.y = (s + 1 >= 0 && s + 1 < len) ? load_sample(src, s + 1, channels) : 0.f;
      v.z = (s + 2 >= 0 && s + 2 < len) ? load_sample(src, s + 2, channels) : 0.f;
      v.w = (s + 3 >= 0 && s + 3 < len) ? load_sample(src, s + 3, channels) : 0.f;
    }
    *reinterpret_cast<float4*>(&xs[dec_pad(4 * q)]) = v;
  }
  __syncthreads();
  float w[kWinQ * 4];
  const int start = threadIdx.x * (kDecR * ORIG);
#pragma unroll
  for (int c = 0; c < kWinQ; ++c) {
    const float4 v = *reinterpret_cast<const float4*>(&xs[dec_pad(start + 4 * c)]);
    w[4 * c] = v.x; w[4 * c + 1] = v.y; w[4 * c + 2] = v.z; w[4 * c + 3] = v.w;
  }
  float acc[kDecR];
#pragma unroll
  for (int r = 0; r < kDecR; ++r) acc[r] = 0.f;
#pragma unroll
  for (int k = 0; k < KW; ++k) {
    const float t = taps.t[k];
#pragma unroll
    for (int r = 0; r < kDecR; ++r) acc[r] = fmaf(w[kOff + r * ORIG + k], t, acc[r]);
  }
  const long long j = j0 + (long long)threadIdx.x * kDecR;
#pragma unroll
  for (int r = 0; r < kDecR; ++r) {
    float v = (j + r < olen) ? acc[r] : 0.f;
    if (clip) v = fminf(fmaxf(v, -1.f), 1.f);
    acc[r] = v;
  }
  if (j + kDecR <= out_cols && (out_stride & 3) == 0 && ((uintptr_t)out & 15) == 0) {
    reinterpret_cast<float4*>(dst + j)[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    reinterpret_cast<float4*>(dst + j)[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  } else {
#pragma unroll
    for (int r = 0; r < kDecR; ++r)
      if (j + r < out_cols) dst[j + r] = acc[r];
  }
}

// ---------------------------------------------------------------------------------------------- k_resample
constexpr int kRI = 8;                 // rows (input strides) per thread item
constexpr int kRP = 4;                 // consecutive phases per thread item
constexpr int kMaxSmemFloats = 24 * 1024;   // staged rows per CTA (96 KB)

// The polyphase FIR as a tiled [strides x taps] x [taps x phases] product on the FMA pipe.  One CTA = `rows` input
// strides of one clip x a block of up to 32 phase quads (blockIdx.z), so that the taps a CTA streams (kw x 4 nqb)
// are shared by all its rows -- through L1: the 8 warps differ in row slot and read the same tap lines.
// Row i of the tile needs the kw samples x[(row0 + i) * orig - width ...]; each row is staged on its own in shared
// memory (pitch kwp = kw rounded up to 4, zeros outside the clip and in the pad) so that a thread reads 4 consecutive
// taps' worth of samples with one 16-byte load.  A thread item is (row slot g, phase quad q): rows g, g + ng, ...,
// g + 7 ng x phases 4 q .. 4 q + 3; per 4 taps it issues 4 16-byte tap loads (tapsT is tap-major, phases padded to a
// multiple of 4 with zero taps, so the lanes of a warp -- consecutive q -- read it coalesced) and 8 16-byte sample
// loads (the same address across the lanes that share g: a broadcast) for 128 FMAs.
// (Tried and dropped, B200 timings for 64 x 10 s at 320->441 / 147->80: register prefetch of the next taps 0.36 /
//  0.29 ms; 8 x 8 register tile with a 4-stage cp.async tap ring in shared memory 0.44 / 0.40 ms; this form 0.42 /
//  0.23 ms.)
template <typename TIn>
__global__ void __launch_bounds__(kThreads) k_resample(const TIn* __restrict__ x, long long row_stride, int channels,
                                                       const long long* __restrict__ lengths, long long L,
                                                       const float* __restrict__ tapsT, int orig, int new_f, int new_pad,
                                                       int width, int kw, int kwp, int rows, int nqb,
                                                       float* __restrict__ out, long long out_stride,
                                                       long long out_cols, int clip) {
  extern __shared__ __align__(16) float xs[];
  const int b = blockIdx.y;
  const long long len = lengths ? min(max(lengths[b], 0LL), L) : L;
  const long long olen = (len * new_f + orig - 1) / orig;          // ceil(new * len / orig)
  const long long row0 = (long long)blockIdx.x * rows;
  float* dst = out + (long long)b * out_stride;
  const long long j0 = row0 * new_f;
  if (j0 >= out_cols) return;
  const long long jend = min(j0 + (long long)rows * new_f, out_cols);
  if (j0 >= olen) {                                                // tile entirely past this clip: zero fill
    if (blockIdx.z == 0)
      for (long long j = j0 + threadIdx.x; j < jend; j += kThreads) dst[j] = 0.f;
    return;
  }
  const TIn* src = x + (long long)b * row_stride;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < rows; i += kThreads / 32) {
    const long long s0 = (row0 + i) * orig - width;
    float* xrow = xs + i * kwp;
    for (int k = lane; k < kwp; k += 32) {
      const long long s = s0 + k;
      xrow[k] = (k < kw && s >= 0 && s < len) ? load_sample(src, s, channels) : 0.f;
    }
  }
  __syncthreads();
  const int ng = rows / kRI;
  const int nq = new_pad / kRP;
  const int q0 = blockIdx.z * nqb;                  // this CTA's block of phase quads
  const int nql = min(nqb, nq - q0);
  const int items = ng * nql;
  const int k4n = kwp >> 2;
  for (int it = threadIdx.x; it < items; it += kThreads) {
    const int g = it / nql, q = q0 + it - g * nql;
    float acc[kRI][kRP];
#pragma unroll
    for (int r = 0; r < kRI; ++r)
#pragma unroll
      for (int c = 0; c < kRP; ++c) acc[r][c] = 0.f;
    const float4* tp = reinterpret_cast<const float4*>(tapsT) + q;
    const float4* xp = reinterpret_cast<const float4*>(xs + g * kwp);
    const int xstep = ng * k4n;                     // float4s between two of this item's rows
#pragma unroll 1
    for (int k4 = 0; k4 < k4n; ++k4) {
      float4 t[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) t[u] = __ldg(tp + (long long)(4 * k4 + u) * nq);
#pragma unroll
      for (int r = 0; r < kRI; ++r) {
        const float4 v = xp[r * xstep + k4];
        const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          acc[r][0] = fmaf(vv[u], t[u].x, acc[r][0]);
          acc[r][1] = fmaf(vv[u], t[u].y, acc[r][1]);
          acc[r][2] = fmaf(vv[u], t[u].z, acc[r][2]);
          acc[r][3] = fmaf(vv[u], t[u].w, acc[r][3]);
        }
      }
    }
#pragma unroll
    for (int r = 0; r < kRI; ++r) {
      const long long jr = (row0 + g + r * ng) * new_f;
#pragma unroll
      for (int c = 0; c < kRP; ++c) {
        const int p = q * kRP + c;
        const long long j = jr + p;
        if (p < new_f && j < out_cols) {
          float v = j < olen ? acc[r][c] : 0.f;
          if (clip) v = fminf(fmaxf(v, -1.f), 1.f);
          dst[j] = v;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------- k_resample_few
// Few phases (new < 16: 2:3, 1:3, 4:3, 3:2, 6:1 ...): the rows overlap almost completely (kw >> orig), so the tile of
// k_resample -- one staged copy of the taps' worth of samples per row -- would stage every sample kw / orig times.
// Here one CTA stages the contiguous segment its `rows` strides need once (coalesced) plus the whole tap table, and
// a thread computes one output (consecutive threads = consecutive outputs: coalesced stores; the lanes of a warp
// read shared-memory words at most `orig` apart).
template <typename TIn>
__global__ void __launch_bounds__(kThreads) k_resample_few(const TIn* __restrict__ x, long long row_stride, int channels,
                                                           const long long* __restrict__ lengths, long long L,
                                                           const float* __restrict__ taps, int orig, int new_f,
                                                           int width, int kw, int rows, float* __restrict__ out,
                                                           long long out_stride, long long out_cols, int clip) {
  extern __shared__ __align__(16) float fs[];
  const int b = blockIdx.y;
  const long long len = lengths ? min(max(lengths[b], 0LL), L) : L;
  const long long olen = (len * new_f + orig - 1) / orig;
  const long long row0 = (long long)blockIdx.x * rows;
  const long long j0 = row0 * new_f;
  if (j0 >= out_cols) return;
  float* dst = out + (long long)b * out_stride + j0;
  const int nout = (int)min((long long)rows * new_f, out_cols - j0);
  if (j0 >= olen) {
    for (int j = threadIdx.x; j < nout; j += kThreads) dst[j] = 0.f;
    return;
  }
  const int seg = (rows - 1) * orig + kw;
  float* xs = fs;
  float* tp = fs + ((seg + 3) & ~3);
  const TIn* src = x + (long long)b * row_stride;
  const long long s0 = row0 * orig - width;
  for (int i = threadIdx.x; i < seg; i += kThreads) {
    const long long s = s0 + i;
    xs[i] = (s >= 0 && s < len) ? load_sample(src, s, channels) : 0.f;
  }
  for (int i = threadIdx.x; i < new_f * kw; i += kThreads) tp[i] = __ldg(taps + i);
  __syncthreads();
  const long long keep = olen - j0;
  for (int jl = threadIdx.x; jl < nout; jl += kThreads) {
    const int i = jl / new_f, p = jl - i * new_f;
    const float* xr = xs + i * orig;
    const float* t = tp + p * kw;
    float a0 = 0.f, a1 = 0.f;
    int k = 0;
#pragma unroll 4
    for (; k + 1 < kw; k += 2) {
      a0 = fmaf(xr[k], t[k], a0);
      a1 = fmaf(xr[k + 1], t[k + 1], a1);
    }
    if (k < kw) a0 = fmaf(xr[k], t[k], a0);
    float v = jl < keep ? a0 + a1 : 0.f;
    if (clip) v = fminf(fmaxf(v, -1.f), 1.f);
    dst[jl] = v;
  }
}

#include "k_resample_tc.cuh"

}  // namespace ingest

// ------------------------------------------------------------------------------------------------ host / C ABI

struct ttts_resample_plan {
  int orig = 0, new_f = 0, new_pad = 0, width = 0, kw = 0;
  int kwp = 0;                // kw rounded up to a multiple of 4 (row pitch of the staged tile, rows of d_tapsT)
  int rows = 0;               // k_resample: input strides per CTA (multiple of ingest::kRI)
  int nqb = 0;                // ... and phase quads per CTA (grid.z = ceil(new_pad / 4 / nqb))
  int decimate = 0;           // k_decimate<orig, kw> applies (new == 1, orig in 2..4, torchaudio's default width)
  int device = 0;
  float* d_tapsT = nullptr;   // [kwp][new_pad], taps and phases zero-padded to multiples of 4
  ingest::TapsParam taps1;    // the single phase of a decimator
  int few = 0;                // k_resample_few applies (< 16 phases, not a templated decimator)
  int few_rows = 0;           // ... input strides per CTA
  float* d_taps = nullptr;    // ... [new_f][kw], the host layout
  int tc = 0;                 // k_resample_tc applies (>= 16 phases; TTTS_RESAMPLE_SIMT=1 forces k_resample)
  int ksteps = 0, npass = 0, nbp = 0, num_sms = 0;
  float* d_tapsB = nullptr;   // [npass][ksteps][hi | lo][2 k-halves][nbp][4] TF32-exact hi, remainder lo
  int32_t* d_status = nullptr;
};

extern "C" int ttts_pcm16_to_f32(const int16_t* pcm, int64_t in_stride, int32_t channels, const int64_t* frames_b,
                                 int32_t B, int64_t frames, float* out, int64_t out_stride, void* stream) {
  if (!pcm || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || frames <= 0 || channels <= 0) return fail(TTTS_MEL_EINVAL, "empty batch (B=%d frames=%lld channels=%d)", B, (long long)frames, channels);
  if (in_stride < frames * channels) return fail(TTTS_MEL_EINVAL, "in_row_stride %lld < frames * channels", (long long)in_stride);
  if (out_stride < frames) return fail(TTTS_MEL_EINVAL, "out_row_stride %lld < frames %lld", (long long)out_stride, (long long)frames);
  if (B > 65535) return fail(TTTS_MEL_EINVAL, "B %d > 65535: split the batch", B);
  const int vec_ok = (in_stride % 8 == 0) && (out_stride % 4 == 0) && ((uintptr_t)pcm % 16 == 0) && ((uintptr_t)out % 16 == 0);
  const long long octs = (frames + 7) / 8;
  long long gx = (octs + ingest::kThreads - 1) / ingest::kThreads;
  if (gx > 148 * 16) gx = 148 * 16;
  dim3 grid((unsigned)gx, (unsigned)B);
  cudaStream_t st = (cudaStream_t)stream;
  const long long* fb = (const long long*)frames_b;
  if (channels == 1) ingest::k_pcm16<1><<<grid, ingest::kThreads, 0, st>>>(pcm, in_stride, channels, fb, frames, out, out_stride, vec_ok);
  else if (channels == 2) ingest::k_pcm16<2><<<grid, ingest::kThreads, 0, st>>>(pcm, in_stride, channels, fb, frames, out, out_stride, vec_ok);
  else ingest::k_pcm16<0><<<grid, ingest::kThreads, 0, st>>>(pcm, in_stride, channels, fb, frames, out, out_stride, 0);
  CUDA_TRY(cudaGetLastError());
  g_launches = 1;
  return 0;
}

extern "C" int ttts_resample_plan_create(const float* taps, int32_t orig, int32_t new_f, int32_t width,
                                         ttts_resample_plan** out) {
  if (!taps || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (orig < 1 || new_f < 1 || width < 0) return fail   /* width 0: the rows do not overlap (inverse STFT, ttts_stft.h) */
     (TTTS_MEL_EINVAL, "bad resample geometry %d -> %d, width %d", orig, new_f, width);
  const int kw = 2 * width + orig;
  const int new_pad = (new_f + ingest::kRP - 1) / ingest::kRP * ingest::kRP;
  const int kwp = (kw + 3) / 4 * 4;
  // CTA tile (k_resample): up to 32 phase quads (one warp's lanes) x as many row slots of 8 rows as fill the 256
  // threads, bounded by the staged tile (rows x kwp floats)
  constexpr size_t kMaxTableBytes = (size_t)256 << 20;      // per tap table; beyond it the rates are not reduced sensibly
  if ((size_t)kwp * new_pad * sizeof(float) > kMaxTableBytes)
    return fail(TTTS_MEL_EINVAL, "resample %d -> %d: %d x %d taps do not fit a table (are the rates divided by their gcd?)", orig, new_f, new_f, kw);
  long long max_ng = ingest::kMaxSmemFloats / kwp / ingest::kRI;
  const bool simt_ok = max_ng >= 1;
  if (!simt_ok) max_ng = 1;
  const int nq = new_pad / ingest::kRP;
  const int npb = (nq + 31) / 32;
  const int nqb = (nq + npb - 1) / npb;
  long long best_ng = ingest::kThreads / nqb;
  if (best_ng < 1) best_ng = 1;
  if (best_ng > max_ng) best_ng = max_ng;
  auto* p = new (std::nothrow) ttts_resample_plan();
  if (!p) return fail(TTTS_MEL_ENOMEM, "host allocation failed");
  p->orig = orig; p->new_f = new_f; p->new_pad = new_pad; p->width = width; p->kw = kw; p->kwp = kwp;
  p->rows = (int)(best_ng * ingest::kRI);
  p->nqb = nqb;
  p->decimate = (new_f == 1 && ((orig == 2 && kw == 28) || (orig == 3 && kw == 41) || (orig == 4 && kw == 54))) ? 1 : 0;
  if (lab_env("TTTS_RESAMPLE_GENERIC")) p->decimate = 0;
  memset(&p->taps1, 0, sizeof(p->taps1));
  if (p->decimate) memcpy(p->taps1.t, taps, kw * sizeof(float));
  std::vector<float> t((size_t)kwp * new_pad, 0.f);
  for (int ph = 0; ph < new_f; ++ph)
    for (int k = 0; k < kw; ++k) t[(size_t)k * new_pad + ph] = taps[(size_t)ph * kw + k];
  cudaError_t e = cudaGetDevice(&p->device);
  // tensor-core path: phases in blocks of <= 192 (multiples of 16), taps in k-steps of 8, split hi / lo
  p->tc = (new_f >= 16 && !p->decimate && !lab_env("TTTS_RESAMPLE_SIMT")) ? 1 : 0;
  p->npass = (new_f + ingest::tc::kMaxNb - 1) / ingest::tc::kMaxNb;
  p->nbp = ((new_f + p->npass - 1) / p->npass + 15) / 16 * 16;
  p->ksteps = (kw + 31) / 32 * 2;                     // pairs of k-steps, rounded up to whole quads
  const size_t arr = (size_t)2 * p->nbp * 4;          // one operand array: [2 k-halves][nbp][4]
  if ((size_t)p->npass * p->ksteps * 4 * arr * sizeof(float) > kMaxTableBytes) p->tc = 0;
  p->few = (!p->tc && !p->decimate && new_f < 16 && !lab_env("TTTS_RESAMPLE_SIMT") &&
            (long long)new_f * kw + kw + orig <= ingest::kMaxSmemFloats) ? 1 : 0;
  if (p->few) {
    long long fr = 2048 / new_f, cap = (ingest::kMaxSmemFloats - (long long)new_f * kw - kw - 4) / orig + 1;
    if (fr > cap) fr = cap;
    if (fr < 1) fr = 1;
    p->few_rows = (int)fr;
  }
  if (!p->tc && !p->decimate && !p->few && !simt_ok) {
    delete p;
    return fail(TTTS_MEL_EINVAL, "resample %d -> %d: filter of %d taps too long for any kernel", orig, new_f, kw);
  }
  std::vector<float> tb;
  if (p->tc) {
    tb.assign((size_t)p->npass * p->ksteps * 4 * arr, 0.f);
    for (int pb = 0; pb < p->npass; ++pb)
      for (int pr = 0; pr < p->ksteps; ++pr) {
        float* stage = tb.data() + ((size_t)pb * p->ksteps + pr) * 4 * arr;       // [k-step of the pair][hi | lo]
        for (int n = 0; n < p->nbp; ++n)
          for (int kk = 0; kk < 16; ++kk) {
            const int ph = pb * p->nbp + n, k = 16 * pr + kk;
            if (ph >= new_f || k >= kw) continue;
            const float v = taps[(size_t)ph * kw + k];
            uint32_t u;
            memcpy(&u, &v, 4);
            u &= 0xFFFFE000u;
            float hi;
            memcpy(&hi, &u, 4);
            const int h = kk >> 3, k8 = kk & 7;
            const size_t at = ((size_t)(k8 >> 2) * p->nbp + n) * 4 + (k8 & 3);
            stage[(2 * h) * arr + at] = hi;
            stage[(2 * h + 1) * arr + at] = v - hi;
          }
      }
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, p->device);
    if (e == cudaSuccess) p->num_sms = prop.multiProcessorCount;
    if (e == cudaSuccess) e = cudaMalloc(&p->d_tapsB, tb.size() * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(p->d_tapsB, tb.data(), tb.size() * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc(&p->d_status, sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMemset(p->d_status, 0, sizeof(int32_t));
    // the attribute belongs to the function, not to the plan: always the largest tile, or a later plan with fewer
    // phases would make an earlier plan's launches fail
    const int dyn = (int)ingest::tc::smem_bytes(ingest::tc::kMaxNb);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ingest::tc::k_resample_tc<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ingest::tc::k_resample_tc<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, dyn);
  }
  if (p->few) {
    if (e == cudaSuccess) e = cudaMalloc(&p->d_taps, (size_t)new_f * kw * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(p->d_taps, taps, (size_t)new_f * kw * sizeof(float), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ingest::k_resample_few<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (ingest::kMaxSmemFloats + 8) * 4);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(ingest::k_resample_few<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (ingest::kMaxSmemFloats + 8) * 4);
  }
  if (e == cudaSuccess) e = cudaMalloc(&p->d_tapsT, t.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->d_tapsT, t.data(), t.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(ingest::k_resample<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, ingest::kMaxSmemFloats * 4);
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(ingest::k_resample<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, ingest::kMaxSmemFloats * 4);
  if (e != cudaSuccess) {
    int rc = fail(TTTS_MEL_ECUDA, "resample plan: %s", cudaGetErrorString(e));
    ttts_resample_plan_destroy(p);
    return rc;
  }
  *out = p;
  return 0;
}

extern "C" void ttts_resample_plan_destroy(ttts_resample_plan* p) {
  if (!p) return;
  cudaFree(p->d_tapsT);
  cudaFree(p->d_tapsB);
  cudaFree(p->d_taps);
  cudaFree(p->d_status);
  delete p;
}

extern "C" int64_t ttts_resample_out_len(const ttts_resample_plan* p, int64_t length) {
  if (!p || length < 0) return fail(TTTS_MEL_EINVAL, "null plan or negative length");
  return (length * p->new_f + p->orig - 1) / p->orig;
}

namespace {

template <typename TIn>
int resample_launch(const ttts_resample_plan* p, const TIn* x, int64_t row_stride, int channels, const int64_t* lengths,
                    int32_t B, int64_t L, float* out, int64_t out_stride, int32_t clip, void* stream) {
  if (!p || !x || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || L <= 0 || channels <= 0) return fail(TTTS_MEL_EINVAL, "empty batch (B=%d L=%lld channels=%d)", B, (long long)L, channels);
  if (B > 65535) return fail(TTTS_MEL_EINVAL, "B %d > 65535: split the batch", B);
  if (row_stride < L * channels) return fail(TTTS_MEL_EINVAL, "row_stride %lld < L %lld x channels %d", (long long)row_stride, (long long)L, channels);
  int cur_dev = -1;
  CUDA_TRY(cudaGetDevice(&cur_dev));
  if (cur_dev != p->device) return fail(TTTS_MEL_EINVAL, "plan lives on device %d, the current device is %d", p->device, cur_dev);
  const long long out_cols = ttts_resample_out_len(p, L);
  if (out_stride < out_cols) return fail(TTTS_MEL_EINVAL, "out_row_stride %lld < %lld output samples", (long long)out_stride, out_cols);
  cudaStream_t st = (cudaStream_t)stream;
  const long long* len = (const long long*)lengths;
  if (p->decimate) {
    const long long gx = (out_cols + ingest::kDecTile - 1) / ingest::kDecTile;
    if (gx > 0x7FFFFFFFLL) return fail(TTTS_MEL_EINVAL, "clip too long");
    dim3 grid((unsigned)gx, (unsigned)B);
    // 16-byte (float / stereo int16) or 8-byte (mono int16) staging loads need aligned row starts
    const int64_t unit = sizeof(TIn) == 4 ? 4 : (channels == 2 ? 8 : 4);
    const uintptr_t align = sizeof(TIn) == 4 ? 16 : (channels == 2 ? 16 : 8);
    const int vec = (sizeof(TIn) == 4 || channels <= 2) && row_stride % unit == 0 && (uintptr_t)x % align == 0;
    // all staging loads before the first conversion pays for fp32 and mono PCM; for stereo PCM (two 16-byte loads'
    // worth of raw words per quad of samples) the registers it takes cost more occupancy than the overlap wins
    const bool hoist = sizeof(TIn) == 4 || channels == 1;
#define TTTS_DEC(O, K) do { \
      if (hoist) ingest::k_decimate<O, K, TIn, true><<<grid, ingest::kThreads, 0, st>>>( \
          x, row_stride, channels, len, L, p->taps1, out, out_stride, out_cols, clip, vec); \
      else ingest::k_decimate<O, K, TIn, false><<<grid, ingest::kThreads, 0, st>>>( \
          x, row_stride, channels, len, L, p->taps1, out, out_stride, out_cols, clip, vec); } while (0)
    if (p->orig == 2) TTTS_DEC(2, 28);
    else if (p->orig == 3) TTTS_DEC(3, 41);
    else TTTS_DEC(4, 54);
#undef TTTS_DEC
  } else if (p->few) {
    const long long tile = (long long)p->few_rows * p->new_f;
    const long long gx = (out_cols + tile - 1) / tile;
    if (gx > 0x7FFFFFFFLL) return fail(TTTS_MEL_EINVAL, "clip too long");
    const int seg = (p->few_rows - 1) * p->orig + p->kw;
    const size_t smem = ((size_t)((seg + 3) & ~3) + (size_t)p->new_f * p->kw) * sizeof(float);
    dim3 grid((unsigned)gx, (unsigned)B);
    ingest::k_resample_few<TIn><<<grid, ingest::kThreads, smem, st>>>(
        x, row_stride, channels, len, L, p->d_taps, p->orig, p->new_f, p->width, p->kw, p->few_rows, out, out_stride,
        out_cols, clip);
  } else if (p->tc) {
    ingest::tc::Params q;
    q.x = x; q.row_stride = row_stride; q.channels = channels; q.lengths = len; q.L = L; q.B = B;
    q.orig = p->orig; q.new_f = p->new_f; q.width = p->width; q.kw = p->kw; q.ksteps = p->ksteps;
    q.npass = p->npass; q.nbp = p->nbp; q.tapsB = p->d_tapsB; q.out = out; q.out_stride = out_stride;
    q.out_cols = out_cols;
    const long long rows_total = (out_cols + p->new_f - 1) / p->new_f;
    q.tiles_per_clip = (int)((rows_total + ingest::tc::kTileM - 1) / ingest::tc::kTileM);
    q.units = (long long)B * q.tiles_per_clip * p->npass;
    q.clip = clip; q.status = p->d_status; q.b_stages = ingest::tc::b_stages(p->nbp);
    q.exp = lab_env("TTTS_RESAMPLE_EXP") ? atoi(lab_env("TTTS_RESAMPLE_EXP")) : 0;
    const int grid = (int)std::min<long long>(p->num_sms, q.units);
    ingest::tc::k_resample_tc<TIn><<<grid, ingest::tc::kThreadsTc, ingest::tc::smem_bytes(p->nbp), st>>>(q);
  } else {
    const long long tile = (long long)p->rows * p->new_f;
    const long long gx = (out_cols + tile - 1) / tile;
    if (gx > 0x7FFFFFFFLL) return fail(TTTS_MEL_EINVAL, "clip too long");
    const size_t smem = (size_t)p->rows * p->kwp * sizeof(float);
    const int nq = p->new_pad / ingest::kRP;
    dim3 grid((unsigned)gx, (unsigned)B, (unsigned)((nq + p->nqb - 1) / p->nqb));
    ingest::k_resample<TIn><<<grid, ingest::kThreads, smem, st>>>(
        x, row_stride, channels, len, L, p->d_tapsT, p->orig, p->new_f, p->new_pad, p->width, p->kw, p->kwp, p->rows,
        p->nqb, out, out_stride, out_cols, clip);
  }
  CUDA_TRY(cudaGetLastError());
  g_launches = 1;
  return 0;
}

}  // namespace

extern "C" int ttts_resample_forward(const ttts_resample_plan* p, const float* x, int64_t row_stride,
                                     const int64_t* lengths, int32_t B, int64_t L, float* out, int64_t out_stride,
                                     int32_t clip, void* stream) {
  return resample_launch<float>(p, x, row_stride, 1, lengths, B, L, out, out_stride, clip, stream);
}

extern "C" int ttts_resample_pcm16_forward(const ttts_resample_plan* p, const int16_t* pcm, int64_t in_row_stride,
                                           int32_t channels, const int64_t* frames, int32_t B, int64_t F, float* out,
                                           int64_t out_stride, int32_t clip, void* stream) {
  return resample_launch<int16_t>(p, pcm, in_row_stride, channels, frames, B, F, out, out_stride, clip, stream);
}
