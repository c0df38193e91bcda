// ttts_mel.cu -- B200 (sm_100a) batched log-mel front-end behind the C ABI of include/ttts_mel.h.
//
// Replaces, for the path TorchMelSpectrogram.forward (ttts/utils/utils.py:509-522) takes through
// torchaudio/torch.stft/matmul, the whole chain reflect-pad -> frame -> Hann -> rDFT -> |.|^p ->
// mel filterbank -> log(clamp) -> / mel_norms with these kernels:
//
//   k_prep    block 0: lengths -> n_frames (torch.stft frame count) + exclusive prefix (frame -> clip map);
//             blocks 1..B: per-clip edge buffers (the few frames that touch the reflect padding), zeros for padded
//             output frames, padding mask
//   k_pad     full per-clip reflect padding into the workspace (only for inputs that are not 16 B addressable)
//   k_logmel2 the fused hot kernel (k_logmel2.cuh; default): persistent, warp-specialised, CTA pairs (cta_group::2),
//             double-folded DFT-GEMM (K = 256 + centre tap, even / odd bin passes), 3xTF32, A operand in TMEM,
//             16 epilogue warps; spectrum magnitudes never leave the SM
//   k_logmel  the first-generation fused kernel (single fold, K = 512), kept as a second tcgen05 implementation
//             (TTTS_MEL_ALGO=1):
//               converter warps cp.async 16 B gathers of the overlapping hop windows, even/odd fold
//                               E[n]=f[n]+f[N-n], O[n]=f[n]-f[N-n]  (halves K), TF32 hi/lo split, tcgen05.st into TMEM
//               TMA warp        cp.async.bulk of pre-tiled windowed cos/sin basis stages -> smem ring
//               MMA warp        tcgen05.mma kind::tf32, A from TMEM, B from smem, fp32 accumulators in
//                               TMEM; 3-term error-compensated product hi*hi + lo*hi + hi*lo
//               epilogue warps  tcgen05.ld -> re^2+im^2 -> sparse mel contraction -> log(clamp) -> /norm
//   k_simt    the same maths on CUDA cores in plain fp32 (independent cross-check, not a fallback)
//   k_stats   per-channel sum / sum-of-squares / count of log-mel for mel_norms
//
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -shared -Xcompiler -fPIC
#include "ttts_mel.h"
#include "ttts_ingest.h"
#include "ttts_stft.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <type_traits>
#include <vector>

// ------------------------------------------------------------------------------------------
// constants of the tiling
// ------------------------------------------------------------------------------------------
namespace {

#ifndef TTTS_PREP_PDL
#define TTTS_PREP_PDL 1      /* k_prep as a programmatic dependent of the previous kernel in the stream: -1 % per step (profiles/r3a) */
#endif
constexpr int kNfft = 1024;            // the only transform size built this round
constexpr int kHop = 256;
constexpr int kKfold = kNfft / 2;      // folded contraction length (kappa = n-1, n = 1..512)
constexpr int kTileM = 128;            // frames per tile == TMEM lanes
constexpr int kKC = 16;                // kappa per A (TMEM) stage and per raw-audio stage
constexpr int kKStep = 8;              // K of one tcgen05.mma kind::tf32
constexpr int kChunks = kKfold / kKC;  // 32 chunks per pass
constexpr int kKSteps = kKfold / kKStep;  // 64 basis stages per pass
constexpr int kNBMax = 192;            // bins per pass (accumulator columns: re | im)
constexpr int kBStages = 4;            // basis ring depth (one k-step each)
constexpr int kConvWarps = 8;           // converter warps (2 sets x 4 TMEM lane quarters)
constexpr int kConvDepth = 2;           // raw-audio stages per converter warp (= 4 chunks ahead overall)
constexpr int kRawPitch = 20;          // floats per staged row (80 B: conflict-free row-per-thread LDS.128)
constexpr int kConvStageBytes = 2 * 32 * kRawPitch * 4;      // 32 rows x (forward window + mirrored window)
constexpr int kRawBytes = kConvWarps * kConvDepth * kConvStageBytes;
constexpr int kEdgeA = 2880;            // per-clip edge buffer, region A: padded samples of frames 0..7
constexpr int kEdgeB = 1792;            //                       region B: padded samples of the last 4 frames
constexpr int kEdgeClip = kEdgeA + kEdgeB;
constexpr int kTmemCols = 512;
constexpr int kAccStride = kNBMax;     // accumulator buffers: one (re [0,nb) | im [192,192+nb)) when nb > 96,
                                       // two (buffer b at 192 b: re [0,nb) | im [96,96+nb)) when nb <= 96
constexpr int kColA = 2 * kNBMax;      // A ring: 4 stages x 32 columns (E_hi,E_lo,O_hi,O_lo x 8 = one k-step)
constexpr int kAStages = 4;
constexpr int kThreads = 512;
constexpr int kMaxMels = 128;
constexpr long long kTimeoutCycles = 4000000000LL;
#ifndef F2_CHAINS
#define F2_CHAINS 4
#endif
constexpr int f2_kChains = F2_CHAINS;  // mel-scan chains per epilogue thread of k_logmel2 (== f2::kChains)

// warp roles of k_logmel
constexpr int kWarpTma = 0, kWarpMma = 1, kWarpTmem = 2;   // warp 3 is spare
constexpr int kWarpConv0 = 4;          // 8 converter warps: 4..11   (warp % 4 == TMEM lane quarter)
constexpr int kWarpEpi0 = 12;          // 4 epilogue warps: 12..15

thread_local char g_err[512] = "";
thread_local int g_launches = 0;

// optional per-call timing of the hot kernel (plan-owned CUDA event pairs, see ttts_mel_profile_*)
struct ProfileRing {
  cudaEvent_t* start = nullptr;
  cudaEvent_t* stop = nullptr;
  int capacity = 0;
  std::atomic<int> count{0};
};
#ifdef TTTS_MEL_LAB
std::atomic<int> g_exp_flags{0};     // lab build only (include/ttts_mel_lab.h): leg-removal timing experiments
#define TTTS_EXP_FLAGS() g_exp_flags.load(std::memory_order_relaxed)
#else
#define TTTS_EXP_FLAGS() 0
#endif

int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

#define CUDA_TRY(expr)                                                                          \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess) return fail(TTTS_MEL_ECUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// Environment knobs (kernel selection for cross-checks, leg-removal experiments) exist in the lab build only; the
// product library never reads the environment.
inline const char* lab_env(const char* name) {
#ifdef TTTS_MEL_LAB
  return getenv(name);
#else
  (void)name;
  return nullptr;
#endif
}

}  // namespace

struct ttts_mel_plan {
  ttts_mel_config cfg;
  int device = 0;
  int num_sms = 0;
  int kmin = 0;        // first live bin
  int n_live = 0;      // live bins (kmin .. kmin+n_live-1)
  int passes = 0;      // accumulator passes over the bin axis
  int nb = 0;          // bins per pass (multiple of 16, <= 192)
  int nbt = 0;         // passes * nb
  float* d_basis = nullptr;    // [passes][64 k-steps][4 arrays][2 k-halves][nb][4] fp32 (TF32-exact values)
  float2* d_melw = nullptr;    // [nbt] weights of bin i into the running filter pair (cur, cur+1)
  uint32_t* d_meladv = nullptr;  // [nbt/16] bit j: one filter closes before bin 16*i+j
  float* d_norms = nullptr;    // [n_mels] or nullptr
  float* d_cosf = nullptr;     // [512][nbt] plain fp32 folded basis for k_simt
  float* d_sinf = nullptr;
  float* d_fb = nullptr;       // [nbt][n_mels] dense filterbank rows of the live bins for k_simt
  // double-fold kernel (k_logmel2.cuh): even / odd bins are separate passes over K = 256 (+ centre tap)
  int algo = 2;        // 2 = k_logmel2 (default), 1 = k_logmel (TTTS_MEL_ALGO=1)
  int kmax = 0;
  int first_k[2] = {0, 0};   // first live bin of each parity
  int sp = 0;          // passes per parity
  int nb2 = 0;         // bins per pass
  int nbt2 = 0;        // 2 * sp * nb2
  float wc = 0.f;      // centre tap weight w[512] * spec_scale
  int cstart[16] = {0};  // [2 sp][4 chains] filter index of each mel-scan chain's first bin
  int cend[16] = {0};    // ... and of the first of the two filters still open at its end
  float* d_basis2 = nullptr;   // [2 sp][33 k-steps][4 arrays][2 k-halves][nb2][4] fp32 (TF32-exact values), no window
  float2* d_melw2 = nullptr;   // [nbt2]
  uint32_t* d_meladv2 = nullptr;  // [nbt2/16], 2 bits per bin
  float* d_win2 = nullptr;     // [512]
  int pair = 0;                // 1: CTA pairs (cta_group::2), each CTA stages half of every basis stage
  float* d_basis2p = nullptr;  // [2 sp][33][2 halves][4 arrays][2 k-halves][nb2/2][4]: the per-CTA half stages
  ProfileRing prof;            // ttts_mel_profile_enable / _collect (include/ttts_mel_lab.h): measurement only
  std::mutex prof_mu;
};

// ------------------------------------------------------------------------------------------
// host: plan construction
// ------------------------------------------------------------------------------------------
namespace {

// round-to-nearest to TF32 (10 explicit mantissa bits); result is exactly representable in TF32
float tf32_rn(double v) {
  float f = (float)v;
  uint32_t u;
  memcpy(&u, &f, 4);
  u = (u + 0x1000u) & 0xFFFFE000u;
  memcpy(&f, &u, 4);
  return f;
}

}  // namespace

extern "C" int ttts_mel_plan_create(const ttts_mel_config* cfg, const float* fb, const float* window,
                                    ttts_mel_plan** out) {
  if (!cfg || !fb || !window || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (cfg->n_fft != kNfft || cfg->hop != kHop)
    return fail(TTTS_MEL_EINVAL, "unsupported n_fft/hop %d/%d (built: %d/%d)", cfg->n_fft, cfg->hop, kNfft, kHop);
  if (cfg->n_mels < 1 || cfg->n_mels > kMaxMels) return fail(TTTS_MEL_EINVAL, "n_mels %d not in [1,%d]", cfg->n_mels, kMaxMels);
  if (cfg->pad < 0 || cfg->pad > kNfft / 2 || cfg->pad % 4 != 0)
    return fail(TTTS_MEL_EINVAL, "pad %d must be a multiple of 4 in [0,%d]", cfg->pad, kNfft / 2);
  if (!(cfg->power == 1.0f || cfg->power == 2.0f)) return fail(TTTS_MEL_EINVAL, "power must be 1 or 2");
  if (!(cfg->clamp > 0.f)) return fail(TTTS_MEL_EINVAL, "clamp must be > 0");
  if (!(cfg->power_eps >= 0.f) || (cfg->power == 2.0f && cfg->power_eps != 0.f))
    return fail(TTTS_MEL_EINVAL, "power_eps must be >= 0 and is only defined for power == 1");
  const int N = kNfft, nfreq = N / 2 + 1, nm = cfg->n_mels;

  // The even/odd fold needs w[n] == w[N-n] and contributes nothing at n = 0 only if w[0] == 0.
  double wmax = 0;
  for (int n = 0; n < N; ++n) wmax = fmax(wmax, fabs((double)window[n]));
  if (fabs((double)window[0]) > 1e-7 * wmax) return fail(TTTS_MEL_EINVAL, "window[0] must be 0 (periodic window)");
  for (int n = 1; n < N; ++n)
    if (fabs((double)window[n] - (double)window[N - n]) > 1e-6 * wmax)
      return fail(TTTS_MEL_EINVAL, "window must satisfy w[n] == w[n_fft-n] (tap %d)", n);

  // live bin range = rows of fb with any non-zero weight
  int kmin = nfreq, kmax = -1;
  for (int k = 0; k < nfreq; ++k)
    for (int m = 0; m < nm; ++m)
      if (fb[(size_t)k * nm + m] != 0.f) { kmin = k < kmin ? k : kmin; kmax = k; break; }
  if (kmax < 0) return fail(TTTS_MEL_EINVAL, "filterbank is all zero");

  ttts_mel_plan* p = new ttts_mel_plan();
  p->cfg = *cfg;
  p->kmin = kmin;
  p->n_live = kmax - kmin + 1;
  int nb_cap = kNBMax;
#ifdef TTTS_MEL_LAB
  if (const char* e = lab_env("TTTS_MEL_NB_MAX")) {          // tuning knob: 96 selects double-buffered accumulators
    const int v = atoi(e);
    if (v >= 16 && v <= kNBMax && v % 16 == 0) nb_cap = v;
  }
#endif
  p->passes = (p->n_live + nb_cap - 1) / nb_cap;
  p->nb = (int)align_up((size_t)((p->n_live + p->passes - 1) / p->passes), 16);
  p->nbt = p->passes * p->nb;
  const int nb = p->nb, nbt = p->nbt;

  // sparse mel table: every bin feeds at most two ADJACENT filters (triangular filterbanks)
  std::vector<float2> melw(nbt);
  std::vector<uint32_t> meladv(nbt / 16, 0u);
  std::vector<float> fbl((size_t)nbt * nm, 0.f);
  int cur = 0;
  for (int i = 0; i < nbt; ++i) {
    float wlo = 0.f, whi = 0.f;
    int adv = 0;
    if (i < p->n_live) {
      const float* row = fb + (size_t)(kmin + i) * nm;
      int first = -1, last = -1;
      for (int m = 0; m < nm; ++m) {
        fbl[(size_t)i * nm + m] = row[m];
        if (row[m] != 0.f) { if (first < 0) first = m; last = m; }
      }
      if (first >= 0) {
        if (last - first > 1 || first < cur || first - cur > 1) {
          delete p;
          return fail(TTTS_MEL_EINVAL, "filterbank row %d: need two adjacent, non-decreasing filters per bin and no filter without bins", kmin + i);
        }
        adv = first - cur;
        cur = first;
        wlo = row[first];
        whi = (last > first) ? row[last] : 0.f;
      }
    }
    melw[i] = make_float2(wlo, whi);
    meladv[i >> 4] |= (uint32_t)adv << (i & 15);
  }

  // folded, windowed basis.  kappa = n-1 for n = 1..512;  the n = 512 tap appears once, so it
  // carries the factor 1/2 (E[512] = 2 f[512]).  Trig argument reduced as an integer first.
  std::vector<double> cb((size_t)kKfold * nbt, 0.0), sb((size_t)kKfold * nbt, 0.0);
  const double scale = (double)cfg->spec_scale;
  for (int kap = 0; kap < kKfold; ++kap) {
    const int n = kap + 1;
    const double w = 0.5 * ((double)window[n] + (double)window[N - n]) * (n == N / 2 ? 0.5 : 1.0) * scale;
    for (int i = 0; i < p->n_live; ++i) {
      const int k = kmin + i;
      const int r = (int)(((long long)k * n) % N);
      const double ang = 2.0 * M_PI * (double)r / (double)N;
      cb[(size_t)kap * nbt + i] = w * cos(ang);
      sb[(size_t)kap * nbt + i] = w * sin(ang);
    }
  }
  std::vector<float> cosf((size_t)kKfold * nbt), sinf_((size_t)kKfold * nbt);
  for (size_t j = 0; j < cosf.size(); ++j) { cosf[j] = (float)cb[j]; sinf_[j] = (float)sb[j]; }

  // tiled stage images: [pass][kstep][array: C_hi, C_lo, S_hi, S_lo][k-half j][row][4 floats]
  // = the UMMA K-major SWIZZLE_NONE canonical layout (8 rows x 16 B core matrices, LBO = nb*16 B
  // between the two k-halves, SBO = 128 B between 8-row groups), so one bulk copy lands a stage.
  const size_t stage_floats = (size_t)4 * 2 * nb * 4;
  std::vector<float> img((size_t)p->passes * kKSteps * stage_floats);
  for (int ps = 0; ps < p->passes; ++ps)
    for (int ks = 0; ks < kKSteps; ++ks) {
      float* st = img.data() + ((size_t)ps * kKSteps + ks) * stage_floats;
      for (int arr = 0; arr < 4; ++arr)
        for (int j = 0; j < 2; ++j)
          for (int r = 0; r < nb; ++r)
            for (int e = 0; e < 4; ++e) {
              const int kap = ks * kKStep + j * 4 + e;
              const double v = (arr < 2 ? cb : sb)[(size_t)kap * nbt + ps * nb + r];
              const float hi = tf32_rn(v);
              const float val = (arr & 1) ? tf32_rn(v - (double)hi) : hi;
              st[(((size_t)arr * 2 + j) * nb + r) * 4 + e] = val;
            }
    }

  // ---- double-fold tables (k_logmel2.cuh)
  p->kmax = kmax;
  p->first_k[kmin & 1] = kmin;
  p->first_k[(kmin & 1) ^ 1] = kmin + 1;
  int n_par[2];
  for (int q = 0; q < 2; ++q) n_par[q] = p->first_k[q] <= kmax ? (kmax - p->first_k[q]) / 2 + 1 : 0;
  const int n_par_max = n_par[0] > n_par[1] ? n_par[0] : n_par[1];
  p->sp = (n_par_max + kNBMax - 1) / kNBMax;
  p->nb2 = (int)align_up((size_t)((n_par_max + p->sp - 1) / p->sp), 16 * f2_kChains);   // scan chains of whole 16-bin blocks
  if (p->nb2 > kNBMax) { p->sp += 1; p->nb2 = (int)align_up((size_t)((n_par_max + p->sp - 1) / p->sp), 16 * f2_kChains); }
  p->nbt2 = 2 * p->sp * p->nb2;
#ifdef TTTS_MEL_LAB
  if (const char* e2 = lab_env("TTTS_MEL_ALGO")) p->algo = atoi(e2) == 1 ? 1 : 2;
#endif
  const int nb2 = p->nb2, nbt2 = p->nbt2, per_par = p->sp * p->nb2;
  std::vector<float2> melw2(nbt2, make_float2(0.f, 0.f));
  std::vector<uint32_t> meladv2(2 * (nbt2 / 16), 0u);   // [0]: single / multi close masks, [1]: 2-bit counts
  const int chain_bins = nb2 / f2_kChains;
  for (int par = 0; par < 2; ++par) {
    int cur2 = 0;
    for (int idx = 0; idx < per_par; ++idx) {
      const int k = p->first_k[par] + 2 * idx, j = par * per_par + idx;
      const bool chain_start = idx % chain_bins == 0;
      int first = -1, last = -1;
      if (k <= kmax) {
        const float* row = fb + (size_t)k * nm;
        for (int m = 0; m < nm; ++m)
          if (row[m] != 0.f) { if (first < 0) first = m; last = m; }
      }
      if (first >= 0 && (last - first > 1 || first < cur2 || first - cur2 > 3)) {
        delete p;
        return fail(TTTS_MEL_EINVAL, "filterbank row %d: need <= two adjacent, non-decreasing filters per bin", k);
      }
      if (chain_start) {                      // a scan chain starts on this bin's own first filter: nothing closes before it
        if (idx > 0) p->cend[j / chain_bins - 1] = cur2;   // the previous chain ends with (cur2, cur2 + 1) open
        if (first >= 0) cur2 = first;
        p->cstart[j / chain_bins] = cur2;
      }
      if (first < 0) continue;
      const float* row = fb + (size_t)k * nm;
      const uint32_t cnt = (uint32_t)(first - cur2);
      meladv2[nbt2 / 16 + (j >> 4)] |= cnt << (2 * (j & 15));
      if (cnt == 1) meladv2[j >> 4] |= 1u << (j & 15);
      if (cnt >= 2) meladv2[j >> 4] |= 0x10000u << (j & 15);
      cur2 = first;
      melw2[j] = make_float2(row[first], last > first ? row[last] : 0.f);
    }
    p->cend[(par + 1) * per_par / chain_bins - 1] = cur2;
  }
  std::vector<float> win2(512, 0.f);
  for (int n = 1; n <= 256; ++n) {
    win2[n - 1] = (float)(0.5 * ((double)window[n] + (double)window[N - n]) * scale);
    if (n < 256) win2[256 + n - 1] = (float)(0.5 * ((double)window[512 - n] + (double)window[512 + n]) * scale);
  }
  p->wc = (float)((double)window[N / 2] * scale);
  const size_t stage2_floats = (size_t)4 * 2 * nb2 * 4;
  std::vector<float> img2((size_t)2 * p->sp * 33 * stage2_floats, 0.f);
  {
    static const double kQuarter[4] = {1.0, 0.0, -1.0, 0.0};
    for (int ps = 0; ps < 2 * p->sp; ++ps) {
      const int par = ps / p->sp, sub = ps % p->sp;
      for (int ks = 0; ks < 33; ++ks) {
        float* st = img2.data() + ((size_t)ps * 33 + ks) * stage2_floats;
        for (int j = 0; j < 2; ++j)
          for (int r = 0; r < nb2; ++r) {
            const int k = p->first_k[par] + 2 * (sub * nb2 + r);
            if (k > kmax) continue;
            for (int e4 = 0; e4 < 4; ++e4) {
              double c, sn;
              if (ks < 32) {
                const int n = ks * kKStep + j * 4 + e4 + 1;
                const int rr = (int)(((long long)k * n) % N);
                if (rr % (N / 4) == 0) { c = kQuarter[rr / (N / 4)]; sn = kQuarter[(rr / (N / 4) + 3) & 3]; }
                else { const double ang = 2.0 * M_PI * (double)rr / (double)N; c = cos(ang); sn = sin(ang); }
              } else {                       // centre tap n = 512: cos(pi k) = (-1)^k, sin = 0; one live slot
                c = (j == 0 && e4 == 0) ? ((k & 1) ? -1.0 : 1.0) : 0.0;
                sn = 0.0;
              }
              for (int arr = 0; arr < 4; ++arr) {
                const double v = arr < 2 ? c : sn;
                const float hi = tf32_rn(v);
                st[(((size_t)arr * 2 + j) * nb2 + r) * 4 + e4] = (arr & 1) ? tf32_rn(v - (double)hi) : hi;
              }
            }
          }
      }
    }
  }

  p->pair = p->nb2 % 32 == 0;                 // CTA pairs by default (TTTS_MEL_PAIR=0: one CTA per tile stream)
#ifdef TTTS_MEL_LAB
  if (const char* e3 = lab_env("TTTS_MEL_PAIR")) p->pair = p->pair && atoi(e3) != 0;
#endif
  std::vector<float> img2p;
  if (p->pair) {                              // same values, each stage split into the two CTAs' bin halves
    const int nh = nb2 / 2;
    img2p.resize(img2.size());
    for (size_t stg = 0; stg < (size_t)2 * p->sp * 33; ++stg)
      for (int half = 0; half < 2; ++half)
        for (int arr = 0; arr < 4; ++arr)
          for (int j = 0; j < 2; ++j)
            for (int r = 0; r < nh; ++r)
              for (int e4 = 0; e4 < 4; ++e4)
                img2p[stg * stage2_floats + ((((size_t)half * 4 + arr) * 2 + j) * nh + r) * 4 + e4] =
                    img2[stg * stage2_floats + (((size_t)arr * 2 + j) * nb2 + half * nh + r) * 4 + e4];
  }

  cudaError_t e;
  int dev = 0;
  e = cudaGetDevice(&dev);
  if (e == cudaSuccess) e = cudaDeviceGetAttribute(&p->num_sms, cudaDevAttrMultiProcessorCount, dev);
  p->device = dev;
  auto up = [&](void** dptr, const void* src, size_t bytes) {
    if (e != cudaSuccess) return;
    e = cudaMalloc(dptr, bytes);
    if (e == cudaSuccess) e = cudaMemcpy(*dptr, src, bytes, cudaMemcpyHostToDevice);
  };
  up((void**)&p->d_basis, img.data(), img.size() * 4);
  up((void**)&p->d_melw, melw.data(), melw.size() * sizeof(float2));
  up((void**)&p->d_meladv, meladv.data(), meladv.size() * sizeof(uint32_t));
  up((void**)&p->d_cosf, cosf.data(), cosf.size() * 4);
  up((void**)&p->d_sinf, sinf_.data(), sinf_.size() * 4);
  up((void**)&p->d_fb, fbl.data(), fbl.size() * 4);
  up((void**)&p->d_basis2, img2.data(), img2.size() * 4);
  up((void**)&p->d_melw2, melw2.data(), melw2.size() * sizeof(float2));
  up((void**)&p->d_meladv2, meladv2.data(), meladv2.size() * sizeof(uint32_t));
  up((void**)&p->d_win2, win2.data(), win2.size() * 4);
  if (p->pair) up((void**)&p->d_basis2p, img2p.data(), img2p.size() * 4);
  if (e != cudaSuccess) {
    int rc = fail(TTTS_MEL_ECUDA, "plan upload: %s", cudaGetErrorString(e));
    ttts_mel_plan_destroy(p);
    return rc;
  }
  *out = p;
  return TTTS_MEL_OK;
}

extern "C" int ttts_mel_plan_set_norms(ttts_mel_plan* p, const float* norms) {
  if (!p) return fail(TTTS_MEL_EINVAL, "null plan");
  if (!norms) {
    if (p->d_norms) cudaFree(p->d_norms);
    p->d_norms = nullptr;
    return TTTS_MEL_OK;
  }
  if (!p->d_norms) CUDA_TRY(cudaMalloc((void**)&p->d_norms, p->cfg.n_mels * sizeof(float)));
  CUDA_TRY(cudaMemcpy(p->d_norms, norms, p->cfg.n_mels * sizeof(float), cudaMemcpyHostToDevice));
  return TTTS_MEL_OK;
}

static void profile_free(ttts_mel_plan* p);
extern "C" void ttts_mel_plan_destroy(ttts_mel_plan* p) {
  if (!p) return;
  profile_free(p);
  cudaFree(p->d_basis); cudaFree(p->d_melw); cudaFree(p->d_meladv); cudaFree(p->d_norms);
  cudaFree(p->d_cosf); cudaFree(p->d_sinf); cudaFree(p->d_fb);
  cudaFree(p->d_basis2); cudaFree(p->d_melw2); cudaFree(p->d_meladv2); cudaFree(p->d_win2); cudaFree(p->d_basis2p);
  delete p;
}

extern "C" int64_t ttts_mel_frames(const ttts_mel_plan* p, int64_t length) {
  if (!p) return fail(TTTS_MEL_EINVAL, "null plan");
  if (length <= p->cfg.pad || length + 2 * p->cfg.pad < kNfft)
    return fail(TTTS_MEL_EINVAL, "length %lld too short for reflect pad %d / n_fft %d", (long long)length, p->cfg.pad, kNfft);
  return 1 + (length + 2 * p->cfg.pad - kNfft) / kHop;
}

extern "C" int32_t ttts_mel_first_bin(const ttts_mel_plan* p) { return p ? p->kmin : -1; }
extern "C" int32_t ttts_mel_num_bins(const ttts_mel_plan* p) { return p ? p->nbt : -1; }
extern "C" int32_t ttts_mel_live_bins(const ttts_mel_plan* p) { return p ? p->n_live : -1; }
extern "C" const char* ttts_mel_last_error(void) { return g_err; }
extern "C" const char* ttts_mel_version(void) { return "ttts_mel 0.1 (sm_100a, tcgen05 tf32x3)"; }
extern "C" int32_t ttts_mel_last_launch_count(void) { return g_launches; }

// ------------------------------------------------------------------------------------------
// workspace
// ------------------------------------------------------------------------------------------
namespace {

struct Workspace {
  int32_t* status;      // [64] word 0 = device error code
  int64_t* frame_off;   // [B+1]
  int64_t* group_off;   // [B+1] exclusive prefix of ceil(n_frames / 32) (k_logmel2's row layout)
  int32_t* n_frames;    // [B]
  float* xp;            // [B][Lp]   (k_simt and unaligned inputs only)
  float* edge;          // [B][kEdgeClip] reflect-padded samples of each clip's first and last frames
  unsigned long long* stat;  // [2 * kMaxMels + 1] fixed-point statistics of one launch + CTA ticket (ttts_mel_forward_stats)
  int64_t Lp;
  size_t bytes;
};

// `direct`: the fused kernel reads the caller's waveform in place (16-byte addressable rows), so the fully padded copy
// `xp` -- as large as the batch itself -- is not needed and not reserved.
Workspace carve(const ttts_mel_plan* p, void* base, int B, int64_t Lmax, bool direct = false) {
  Workspace w{};
  const int64_t T = 1 + (Lmax + 2 * p->cfg.pad - kNfft) / kHop;
  w.Lp = (int64_t)align_up((size_t)(kHop * (T > 0 ? T - 1 : 0) + kNfft + 64), 64);
  size_t off = 0;
  char* b = (char*)base;
  w.status = (int32_t*)(b + off); off += 256;
  w.stat = (unsigned long long*)(b + off); off += align_up((size_t)(2 * kMaxMels + 1) * 8, 256);
  w.frame_off = (int64_t*)(b + off); off += align_up((size_t)(B + 1) * 8, 256);
  w.group_off = (int64_t*)(b + off); off += align_up((size_t)(B + 1) * 8, 256);
  w.n_frames = (int32_t*)(b + off); off += align_up((size_t)B * 4, 256);
  w.edge = (float*)(b + off); off += align_up((size_t)B * kEdgeClip * 4, 256);
  w.xp = (float*)(b + off); off += direct ? 0 : align_up((size_t)B * (size_t)w.Lp * 4, 256);
  w.bytes = off;
  return w;
}

}  // namespace

extern "C" size_t ttts_mel_workspace_bytes(const ttts_mel_plan* p, int32_t B, int64_t Lmax) {
  if (!p || B <= 0 || Lmax <= 0) return 0;
  return carve(p, nullptr, B, Lmax).bytes;
}

static bool reads_in_place(const void* wav, int64_t row_stride, int32_t B) {
  return (row_stride % 4 == 0) && (((uintptr_t)wav & 15) == 0) && (int64_t)B * row_stride < ((int64_t)1 << 31);
}

extern "C" size_t ttts_mel_workspace_bytes_for(const ttts_mel_plan* p, int32_t B, int64_t Lmax, const float* wav,
                                               int64_t row_stride) {
  if (!p || B <= 0 || Lmax <= 0) return 0;
  return carve(p, nullptr, B, Lmax, reads_in_place(wav, row_stride, B)).bytes;
}

// ------------------------------------------------------------------------------------------
// device helpers (PTX)
// ------------------------------------------------------------------------------------------
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ uint32_t mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok;
}
// Bounded wait: a pipeline bug must surface as an error code, never as a hung GPU.  The slow path is out of line (the
// wait sites stay two instructions long); kRole gives every register-budget region of k_logmel2 (setmaxnreg) its own
// copy, because ptxas compiles a function shared by several regions for the smallest of their budgets -- and then the
// callers too.
__device__ __forceinline__ void mbar_timeout(int32_t* status, int code) {
  atomicExch(status, code);
  __threadfence_system();
  __trap();
}
// Slow path: try_wait with a suspend-time hint, so that a waiting warp sleeps in hardware (it wakes as soon as the phase
// completes) instead of re-issuing the probe every ~80 cycles -- ncu of round 1 showed 58 % of all issued warp
// instructions in these loops, competing with the converter warps for issue slots.
#ifndef TTTS_WAIT_HINT_NS
#define TTTS_WAIT_HINT_NS 0      /* measured (profiles/r2a_variants.json): no gain on V1, 2 % slower on V2 */
#endif
__device__ __forceinline__ uint32_t mbar_try_wait_hint(uint32_t bar, uint32_t parity) {
#if TTTS_WAIT_HINT_NS > 0
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity), "r"((uint32_t)TTTS_WAIT_HINT_NS) : "memory");
  return ok;
#else
  return mbar_try_wait(bar, parity);
#endif
}
template <int kRole>
__device__ __noinline__ void mbar_wait_slow(uint32_t bar, uint32_t parity, int32_t* status, int code) {
  // (a suspend-time hint on the probe, or a nanosleep between probes of the epilogue's pass-long waits, changes nothing
  // measurable: profiles/r2a_variants.json, r2l_variants_sleep_fastlog.txt -- the spinning warps do not starve the others)
  const long long t0 = clock64();
  while (!mbar_try_wait_hint(bar, parity)) {
    if (clock64() - t0 > kTimeoutCycles) mbar_timeout(status, code);
  }
}
template <int kRole = 0>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int32_t* status, int code) {
  if (!mbar_try_wait(bar, parity)) mbar_wait_slow<kRole>(bar, parity, status, code);
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[tmem] * B[smem]^T, kind::tf32, M = 128, K = 8
__device__ __forceinline__ void mma_tf32_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
      ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld4(uint32_t taddr, uint32_t* v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

__device__ __forceinline__ uint32_t elect_one_sync() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
      "elect.sync rx|px, %1;\n\t"
      "@px mov.s32 %0, 1;\n\t}"
      : "+r"(pred) : "r"(0xffffffffu));
  return pred;
}

// frame index g (clips concatenated) -> clip b; frame_off is the exclusive prefix of n_frames
__device__ __forceinline__ int find_clip(const int64_t* __restrict__ frame_off, int B, int64_t g) {
  int lo = 0, hi = B - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (frame_off[mid] <= g) lo = mid; else hi = mid - 1;
  }
  return lo;
}

}  // namespace

// ------------------------------------------------------------------------------------------
// k_prep: lengths -> n_frames, frame_off (exclusive prefix), status
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ long long clip_frames(long long L, long long Lmax, int pad, int Tmax, int* err) {
  long long T = 0;
  if (L > pad && L <= Lmax && L + 2 * pad >= kNfft) T = 1 + (L + 2 * pad - kNfft) / kHop;
  else *err = 1;                                // bad length
  if (T > Tmax) { T = Tmax; *err = 2; }
  return T;
}

constexpr int kPrepSplit = 4;            // blocks per clip for the side work
__device__ void clip_body(const float* __restrict__ wav, int64_t row_stride, const int64_t* __restrict__ lengths,
                          int64_t Lmax, int pad, int Tmax, float* __restrict__ edge, float* __restrict__ out, int n_mels,
                          uint8_t* __restrict__ mask, int b, int part);

// Block 0: frame counts + prefix.  Blocks 1..B: per-clip side work of clip blockIdx.x-1 -- its edge buffer (when `edge`
// is given), zeros for its padded output frames (ragged batches) and its row of the padding mask (when `mask` is given).
__global__ void __launch_bounds__(1024) k_prep(const float* __restrict__ wav, int64_t row_stride, float* __restrict__ edge,
                                               float* __restrict__ out, int n_mels, uint8_t* __restrict__ mask,
                                               const int64_t* __restrict__ lengths, int B, int64_t Lmax, int pad,
                                               int Tmax, int32_t* __restrict__ n_frames,
                                               int32_t* __restrict__ n_frames_user, int64_t* __restrict__ frame_off,
                                               int64_t* __restrict__ group_off, int32_t* __restrict__ status,
                                               unsigned long long* __restrict__ stat_ws) {
#if TTTS_PREP_PDL
  // k_prep is itself a programmatic dependent of whatever precedes it in the stream (the previous call's k_logmel2, which
  // triggers early): its launch is staged while that kernel still runs; nothing is read or written before the wait
  asm volatile("griddepcontrol.wait;" ::: "memory");
#endif
  asm volatile("griddepcontrol.launch_dependents;");   // the fused kernel may start its prologue (it waits for this grid's end)
  if (blockIdx.x > 0) {
    clip_body(wav, row_stride, lengths, Lmax, pad, Tmax, edge, out, n_mels, mask, (blockIdx.x - 1) / kPrepSplit,
              (blockIdx.x - 1) % kPrepSplit);
    return;
  }
  __shared__ long long part[1024];
  const int tid = threadIdx.x;
  if (tid == 0) *status = 0;
  if (stat_ws && tid <= 2 * kMaxMels) stat_ws[tid] = 0ull;
  __syncthreads();
  if (!lengths) {                           // dense batch: closed forms, no scan
    int err = 0;
    const long long T = clip_frames(Lmax, Lmax, pad, Tmax, &err);
    if (err && tid == 0) atomicExch(status, err);
    const long long G = (T + 31) >> 5;
    for (int b = tid; b <= B; b += 1024) {
      if (b < B) { n_frames[b] = (int32_t)T; if (n_frames_user) n_frames_user[b] = (int32_t)T; }
      frame_off[b] = T * b;
      group_off[b] = G * b;
    }
    return;
  }
  const int per = (B + 1023) / 1024;
  const int b0 = tid * per, b1 = min(B, b0 + per);
  long long sum = 0;
  for (int b = b0; b < b1; ++b) {
    const long long L = lengths ? lengths[b] : Lmax;
    int err = 0;
    const long long T = clip_frames(L, Lmax, pad, Tmax, &err);
    if (err) atomicExch(status, err);
    n_frames[b] = (int32_t)T;
    if (n_frames_user) n_frames_user[b] = (int32_t)T;
    sum += T;
  }
  part[tid] = sum;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    const long long v = tid >= d ? part[tid - d] : 0;
    __syncthreads();
    part[tid] += v;
    __syncthreads();
  }
  long long run = part[tid] - sum;
  for (int b = b0; b < b1; ++b) { frame_off[b] = run; run += n_frames[b]; }
  if (tid == 1023) frame_off[B] = part[1023];
  // second prefix: groups of 32 frames per clip
  __syncthreads();
  long long gsum = 0;
  for (int b = b0; b < b1; ++b) gsum += (n_frames[b] + 31) >> 5;
  part[tid] = gsum;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {
    const long long v = tid >= d ? part[tid - d] : 0;
    __syncthreads();
    part[tid] += v;
    __syncthreads();
  }
  run = part[tid] - gsum;
  for (int b = b0; b < b1; ++b) { group_off[b] = run; run += (n_frames[b] + 31) >> 5; }
  if (tid == 1023) group_off[B] = part[1023];
}

// ------------------------------------------------------------------------------------------
// k_pad: xp[b][i] = reflect-padded clip b (pad on each side at the clip's OWN length), 0 beyond
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_pad(const float* __restrict__ wav, int64_t row_stride,
                                             const int64_t* __restrict__ lengths, int64_t Lmax, int pad,
                                             float* __restrict__ xp, int64_t Lp, int vec_ok) {
  const int b = blockIdx.y;
  const int64_t L = lengths ? lengths[b] : Lmax;
  const float* __restrict__ src = wav + (int64_t)b * row_stride;
  float* __restrict__ dst = xp + (int64_t)b * Lp;
  const bool okL = (L > pad) && (L <= Lmax);
  // only what this clip's frames read: its own L + 2 pad samples (a ragged batch of 1..30 s clips is 2/3 padding;
  // writing zeros over all of it made this kernel 0.29 ms for 256 clips, 40 % of the log-mel kernel's time)
  const int64_t Lb = okL ? min(Lp, (L + 2 * (int64_t)pad + 3) & ~(int64_t)3) : 0;
  for (int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i4 < Lb;
       i4 += (int64_t)gridDim.x * blockDim.x * 4) {
    float4 v;
    const int64_t s0 = i4 - pad;
    if (okL && vec_ok && s0 >= 0 && s0 + 3 < L) {
      v = __ldg(reinterpret_cast<const float4*>(src + s0));
    } else {
      float t[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        int64_t s = s0 + e;
        float val = 0.f;
        if (okL && s >= -(int64_t)pad && s < L + pad) {
          if (s < 0) s = -s;
          if (s >= L) s = 2 * (L - 1) - s;
          val = __ldg(src + s);
        }
        t[e] = val;
      }
      v = make_float4(t[0], t[1], t[2], t[3]);
    }
    *reinterpret_cast<float4*>(dst + i4) = v;
  }
}

// ------------------------------------------------------------------------------------------
// edge buffers (k_prep blocks 1..B): only the frames that touch the reflect padding are materialised (first frames and last
// <= 3 frames of every clip, or the whole clip when it has <= 8 frames); every other frame is read
// straight from the caller's waveform by k_logmel.  edge[b] = A: padded[0, 2880) | B: padded[256 (T-4), +1792)
// ------------------------------------------------------------------------------------------
__device__ void clip_body(const float* __restrict__ wav, int64_t row_stride, const int64_t* __restrict__ lengths,
                          int64_t Lmax, int pad, int Tmax, float* __restrict__ edge, float* __restrict__ out, int n_mels,
                          uint8_t* __restrict__ mask, int b, int part) {
  const int64_t L = lengths ? lengths[b] : Lmax;
  const int tid = part * blockDim.x + threadIdx.x, nthr = kPrepSplit * blockDim.x;
  int err = 0;
  const int T = (int)clip_frames(L, Lmax, pad, Tmax, &err);
  if (edge) {
    const float* __restrict__ src = wav + (int64_t)b * row_stride;
    float* __restrict__ dst = edge + (int64_t)b * kEdgeClip;
    const bool okL = T > 0 && err != 1;
    const int64_t startB = (int64_t)kHop * (T - 4);
    for (int i = tid; i < kEdgeClip; i += nthr) {
      const bool inB = i >= kEdgeA;
      int64_t s = (inB ? startB + (i - kEdgeA) : (int64_t)i) - pad;
      float v = 0.f;
      if (okL && (!inB || T > 8) && s >= -(int64_t)pad && s < L + pad) {
        if (s < 0) s = -s;
        if (s >= L) s = 2 * (L - 1) - s;
        v = __ldg(src + s);
      }
      dst[i] = v;
    }
  }
  if (out && lengths) {          // collate contract (ttts/gpt/dataset.py:116-117): frames t >= n_frames[b] are 0.0
    float* __restrict__ z = out + ((int64_t)b * Tmax + T) * n_mels;
    const int64_t n = (int64_t)(Tmax - T) * n_mels;
    for (int64_t i = tid; i < n; i += nthr) z[i] = 0.f;
  }
  if (mask) {                    // make_pad_mask (ttts/utils/utils.py:77-103)
    uint8_t* __restrict__ mrow = mask + (int64_t)b * Tmax;
    for (int t = tid; t < Tmax; t += nthr) mrow[t] = t >= T ? 1 : 0;
  }
}

// ------------------------------------------------------------------------------------------
// k_logmel: the fused tcgen05 kernel
// ------------------------------------------------------------------------------------------
struct LogmelParams {
  const float* src;           // direct: the caller's waveform [B][src_stride]; else workspace xp [B][Lp]
  int64_t src_stride;
  const float* edge;          // [B][kEdgeClip] (direct mode)
  const int64_t* lengths;     // [B] or null
  int64_t Lmax;
  int direct, pad;
  int T_uniform;              // frames per clip when every clip is Lmax long (lengths == null), else 0
  const int64_t* frame_off;   // [B+1]
  const int32_t* n_frames;    // [B]
  int B;
  int Tmax;
  float* out;                 // [B][Tmax][n_mels]
  const float* basis;         // tiled stage images
  const float2* melw;         // [nbt]
  const uint32_t* meladv;     // [nbt/16]
  const float* norms;         // [n_mels] or null
  int passes, nb, nbt, n_mels;
  float clamp, power_eps;
  int power_is_1;
  int32_t* status;
  float* dbg_reim;            // optional [total_frames][2*nbt]
  int exp_flags;              // timing experiments (results invalid): 1 no A handoff, 2 no basis loads, 4 no epilogue, 32 no MMAs (commits only)
};

struct __align__(8) LogmelBarriers {
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t a_full[kAStages], a_empty[kAStages];
  uint64_t acc_full[2], acc_empty[2];
};

template <bool kPower1, bool kDebug>
__global__ void __launch_bounds__(kThreads, 1) k_logmel(const LogmelParams P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  // ---- carve shared memory
  const int nb = P.nb;
  const uint32_t b_stage_bytes = (uint32_t)nb * 128u;       // 4 arrays x 2 k-halves x nb rows x 16 B
  uint8_t* s_b = smem;                                       // basis ring
  uint8_t* s_raw = s_b + (size_t)kBStages * b_stage_bytes;   // raw audio ring
  float2* s_melw = reinterpret_cast<float2*>(s_raw + (size_t)kRawBytes);
  uint32_t* s_adv = reinterpret_cast<uint32_t*>(s_melw + P.nbt);
  float* s_norm = reinterpret_cast<float*>(s_adv + ((P.nbt / 16 + 3) & ~3));      // [kMaxMels] mel_norms (or 1)
  float* s_park = s_norm + kMaxMels;                                 // [128][n_mels+1] raw mel sums of the tile
  const int park_pitch = P.n_mels + 1;
  LogmelBarriers* bars = reinterpret_cast<LogmelBarriers*>(s_park + ((kTileM * park_pitch + 1) & ~1));
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t total_frames = P.frame_off[P.B];
  const int num_tiles = (int)((total_frames + kTileM - 1) / kTileM);

  // ---- one-time setup
  for (int i = threadIdx.x; i < P.nbt; i += kThreads) s_melw[i] = P.melw[i];
  for (int i = threadIdx.x; i < P.nbt / 16; i += kThreads) s_adv[i] = P.meladv[i];
  for (int i = threadIdx.x; i < P.n_mels; i += kThreads) s_norm[i] = P.norms ? P.norms[i] : 1.0f;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kBStages; ++i) { mbar_init(smem_u32(&bars->b_full[i]), 1); mbar_init(smem_u32(&bars->b_empty[i]), 1); }
    for (int i = 0; i < kAStages; ++i) { mbar_init(smem_u32(&bars->a_full[i]), 4); mbar_init(smem_u32(&bars->a_empty[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&bars->acc_full[i]), 1); mbar_init(smem_u32(&bars->acc_empty[i]), 4); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpTmem) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"((uint32_t)kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s_tmem;
  const uint32_t acc_bufs = nb <= 96 ? 2u : 1u;              // double-buffered accumulators when they fit
  const uint32_t col_im = nb <= 96 ? 96u : (uint32_t)kNBMax;

  if (warp == kWarpTma) {
    // ===== basis producer: one bulk copy per k-step stage =====
    if (lane == 0 && !(P.exp_flags & 2)) {
      uint32_t seq = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x)
        for (int ps = 0; ps < P.passes; ++ps)
          for (int ks = 0; ks < kKSteps; ++ks, ++seq) {
            const uint32_t st = seq % kBStages, use = seq / kBStages;
            mbar_wait(smem_u32(&bars->b_empty[st]), (use & 1) ^ 1, P.status, 101);
            const uint32_t full = smem_u32(&bars->b_full[st]);
            mbar_expect_tx(full, b_stage_bytes);
            bulk_g2s(smem_u32(s_b + (size_t)st * b_stage_bytes),
                     P.basis + ((size_t)ps * kKSteps + ks) * (b_stage_bytes / 4), b_stage_bytes, full);
          }
    }
    __syncwarp();
  } else if (warp == kWarpMma) {
    // ===== MMA issuer: the whole warp stays converged (waits are warp-uniform), one elected lane issues =====
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(nb >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
    // K-major SWIZZLE_NONE descriptors: high word is constant (SBO = 128 B, version 1), low word = addr>>4 | LBO>>4 << 16
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);
    const uint32_t desc_lo0 = (((uint32_t)nb * 16u) >> 4) << 16;
    const uint32_t arr16 = ((uint32_t)nb * 32u) >> 4;           // one operand array, in 16 B units
    const uint32_t sb0 = smem_u32(s_b) >> 4, stage16 = b_stage_bytes >> 4;
    uint32_t bseq = 0, cseq = 0, pseq = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x)
      for (int ps = 0; ps < P.passes; ++ps, ++pseq) {
        const uint32_t abuf = pseq % acc_bufs, ause_acc = pseq / acc_bufs;
        const uint32_t d_re = tmem + abuf * kAccStride, d_im = d_re + col_im;
        if (!(P.exp_flags & 4)) mbar_wait(smem_u32(&bars->acc_empty[abuf]), (ause_acc & 1) ^ 1, P.status, 201);
        for (int c = 0; c < kChunks; ++c, ++cseq) {
#pragma unroll
          for (int kk = 0; kk < kKC / kKStep; ++kk, ++bseq) {
            // A stage: chunks of parity p use stages 2p (first k-step) and 2p+1 (second k-step)
            const uint32_t as = 2 * (cseq & 1) + kk, ause = cseq >> 1;
            const uint32_t st = bseq % kBStages, use = bseq / kBStages;
            if (!(P.exp_flags & 1)) mbar_wait(smem_u32(&bars->a_full[as]), ause & 1, P.status, 202);
            if (!(P.exp_flags & 2)) mbar_wait(smem_u32(&bars->b_full[st]), use & 1, P.status, 203);
            tc_fence_after();
            if (elect_one_sync()) {
              const uint32_t a_base = tmem + kColA + as * 32;
              const uint32_t lo = desc_lo0 | (sb0 + st * stage16);
              const uint64_t c_hi = ((uint64_t)desc_hi << 32) | lo;
              const uint64_t c_lo = ((uint64_t)desc_hi << 32) | (lo + arr16);
              const uint64_t s_hi = ((uint64_t)desc_hi << 32) | (lo + 2 * arr16);
              const uint64_t s_lo = ((uint64_t)desc_hi << 32) | (lo + 3 * arr16);
              const uint32_t e_hi = a_base, e_lo = a_base + 8, o_hi = a_base + 16, o_lo = a_base + 24;
              const uint32_t first = (c == 0 && kk == 0) ? 0u : 1u;
              if (!(P.exp_flags & 32)) {
              mma_tf32_ts(d_re, e_lo, c_hi, idesc, first);   // small terms first
              mma_tf32_ts(d_re, e_hi, c_lo, idesc, 1u);
              mma_tf32_ts(d_re, e_hi, c_hi, idesc, 1u);
              mma_tf32_ts(d_im, o_lo, s_hi, idesc, first);
              mma_tf32_ts(d_im, o_hi, s_lo, idesc, 1u);
              mma_tf32_ts(d_im, o_hi, s_hi, idesc, 1u);
              }
              tc_commit(smem_u32(&bars->b_empty[st]));       // basis stage free once these MMAs retire
              tc_commit(smem_u32(&bars->a_empty[as]));       // A stage free
              if (c == kChunks - 1 && kk == kKC / kKStep - 1)
                tc_commit(smem_u32(&bars->acc_full[abuf]));  // accumulators of this pass complete
            }
            __syncwarp();
          }
        }
      }
  } else if (warp >= kWarpConv0 && warp < kWarpEpi0 && !(P.exp_flags & 1)) {
    // ===== converters: gather own rows (cp.async) -> fold + TF32 split -> TMEM A operand =====
    // Warp (q, set) owns frame rows 32q..32q+31 of the tile for the chunks c == set (mod 2).  It stages
    // exactly the rows it later reads, so the only synchronisation of the raw-audio ring is
    // cp.async.wait_group + __syncwarp: no loader warp, no mbarriers on this leg.
    const int cw = warp - kWarpConv0;
    const int set = cw >> 2;                           // which A stage (and chunk parity) this warp feeds
    const int q = warp & 3;                            // TMEM lane quarter
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    uint8_t* my_raw = s_raw + (size_t)cw * kConvDepth * kConvStageBytes;
    const int fj = lane & 7, fr = lane >> 3;           // forward window: 8 lanes per row (5 active), rows fr + 4k
    const int mj = lane & 3, mr = lane >> 2;           // mirrored window: 4 lanes per row, rows mr + 8k

    // issue cursor (runs kConvDepth chunks ahead of the consume cursor)
    int i_tile = blockIdx.x, i_ps = 0, i_c = set;
    int32_t off_f[8], off_m[4];
    uint32_t rows_valid = 0, rows_edge = 0;            // per row of this warp: has a frame / lives in the edge buffer
    auto load_row_offsets = [&](int tile) {
      int32_t base = 0;
      bool valid = false, edge = false;
      if (tile < num_tiles) {
        const int64_t g = (int64_t)tile * kTileM + q * 32 + lane;
        if (g < total_frames) {
          valid = true;
          int b, t;
          if (P.T_uniform) { b = (int)(g / P.T_uniform); t = (int)(g - (int64_t)b * P.T_uniform); }   // no memory traffic
          else { b = find_clip(P.frame_off, P.B, g); t = (int)(g - P.frame_off[b]); }
          if (!P.direct) {
            base = (int32_t)((int64_t)b * P.src_stride + (int64_t)t * kHop);
          } else {
            const int T = P.T_uniform ? P.T_uniform : P.n_frames[b];
            const int64_t L = P.lengths ? P.lengths[b] : P.Lmax;
            const int64_t s0 = (int64_t)t * kHop - P.pad;
            if (T <= 8 || s0 < 0) { edge = true; base = b * kEdgeClip + t * kHop; }
            else if (s0 + kNfft > L) { edge = true; base = b * kEdgeClip + kEdgeA + (t - (T - 4)) * kHop; }
            else base = (int32_t)((int64_t)b * P.src_stride + s0);
          }
        }
      }
      rows_valid = __ballot_sync(0xffffffffu, valid);
      rows_edge = __ballot_sync(0xffffffffu, edge);
#pragma unroll
      for (int k = 0; k < 8; ++k) off_f[k] = __shfl_sync(0xffffffffu, base, fr + 4 * k) + 4 * fj;
#pragma unroll
      for (int k = 0; k < 4; ++k) off_m[k] = __shfl_sync(0xffffffffu, base, mr + 8 * k) + (kNfft - kKC) + 4 * mj;
    };
    auto issue = [&](uint32_t stage) {
      if (i_tile < num_tiles) {
        const uint32_t fw = smem_u32(my_raw + (size_t)stage * kConvStageBytes);
        const uint32_t mw = fw + 32 * kRawPitch * 4;
        const int adv = kKC * i_c;
        // forward window: xp[row + 16c + i], i = 0..19 (taps n = 16c+1 .. 16c+16 are i = 1..16)
        if (fj < 5) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int row = fr + 4 * k;
            const bool ok = (rows_valid >> row) & 1u;
            const float* bp = ((rows_edge >> row) & 1u) ? P.edge : P.src;
            cp_async16(fw + (uint32_t)(row * kRawPitch + 4 * fj) * 4u, ok ? bp + (off_f[k] + adv) : P.src, ok ? 16u : 0u);
          }
        }
        // mirrored window: xp[row + 1008 - 16c + i], i = 0..15 (tap N-n is i = 15-j)
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int row = mr + 8 * k;
          const bool ok = (rows_valid >> row) & 1u;
          const float* bp = ((rows_edge >> row) & 1u) ? P.edge : P.src;
          cp_async16(mw + (uint32_t)(row * kRawPitch + 4 * mj) * 4u, ok ? bp + (off_m[k] - adv) : P.src, ok ? 16u : 0u);
        }
        // advance the cursor over (tile, pass, chunk of my parity)
        i_c += 2;
        if (i_c >= kChunks) {
          i_c = set;
          if (++i_ps >= P.passes) { i_ps = 0; i_tile += gridDim.x; load_row_offsets(i_tile); }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");   // empty groups keep the wait_group count uniform
    };

    load_row_offsets(i_tile);
#pragma unroll
    for (int d = 0; d < kConvDepth; ++d) issue(d);

    uint32_t cseq = set, mine = 0;                     // cseq: global chunk counter (A-stage phases); mine: my chunk counter
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x)
      for (int ps = 0; ps < P.passes; ++ps)
        for (int c = set; c < kChunks; c += 2, cseq += 2, ++mine) {
          const uint32_t stage = mine % kConvDepth;
          asm volatile("cp.async.wait_group %0;" ::"n"(kConvDepth - 1) : "memory");
          __syncwarp();
          const float4* fw = reinterpret_cast<const float4*>(my_raw + (size_t)stage * kConvStageBytes + (size_t)lane * kRawPitch * 4);
          const float4* mw = fw + 32 * kRawPitch / 4;
          float f[20], m[16];
#pragma unroll
          for (int i = 0; i < 5; ++i) { const float4 v = fw[i]; f[4 * i] = v.x; f[4 * i + 1] = v.y; f[4 * i + 2] = v.z; f[4 * i + 3] = v.w; }
#pragma unroll
          for (int i = 0; i < 4; ++i) { const float4 v = mw[i]; m[4 * i] = v.x; m[4 * i + 1] = v.y; m[4 * i + 2] = v.z; m[4 * i + 3] = v.w; }
          // TF32 split: hi = x with the 13 low mantissa bits cleared (exactly representable, so it does not matter
          // how the tensor core would round it), lo = x - hi exactly; the core drops lo's bits below 2^-11 lo.
          uint32_t ehi[16], elo[16], ohi[16], olo[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float e = f[1 + j] + m[15 - j];          // E = f[n] + f[N-n]
            const float o = f[1 + j] - m[15 - j];          // O = f[n] - f[N-n]
            const uint32_t he = __float_as_uint(e) & 0xFFFFE000u, ho = __float_as_uint(o) & 0xFFFFE000u;
            ehi[j] = he; elo[j] = __float_as_uint(e - __uint_as_float(he));
            ohi[j] = ho; olo[j] = __float_as_uint(o - __uint_as_float(ho));
          }
          __syncwarp();                                  // every lane has its row in registers: the stage can be refilled
          issue(stage);
          const uint32_t ause = cseq >> 1;               // == mine: how often this set's A stages were used
#pragma unroll
          for (int h = 0; h < 2; ++h) {                  // one TMEM stage per k-step (8 kappa)
            const uint32_t as = 2 * set + h;
            const uint32_t a_base = tmem + lane_addr + kColA + as * 32;
            mbar_wait(smem_u32(&bars->a_empty[as]), (ause & 1) ^ 1, P.status, 402);
            tc_fence_after();
            tmem_st8(a_base + 0, ehi + 8 * h);
            tmem_st8(a_base + 8, elo + 8 * h);
            tmem_st8(a_base + 16, ohi + 8 * h);
            tmem_st8(a_base + 24, olo + 8 * h);
            tmem_wait_st();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(smem_u32(&bars->a_full[as]));
          }
        }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp >= kWarpEpi0 && !(P.exp_flags & 4)) {
    // ===== epilogue: power -> sparse mel -> log(clamp) -> / norm =====
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    uint32_t pseq = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int64_t g = (int64_t)tile * kTileM + r;
      bool valid = g < total_frames;
      float* orow = nullptr;
      if (valid) {
        int b, t;
        if (P.T_uniform) { b = (int)(g / P.T_uniform); t = (int)(g - (int64_t)b * P.T_uniform); }
        else { b = find_clip(P.frame_off, P.B, g); t = (int)(g - P.frame_off[b]); }
        valid = t < P.Tmax;
        orow = P.out + ((int64_t)b * P.Tmax + t) * P.n_mels;
      }
      // Sparse mel contraction: every bin feeds the running pair of adjacent filters (cur, cur+1); the
      // host table says how many filters close before each bin.  Closed filters are parked as raw sums
      // in shared memory and turned into log-mel in a coalesced pass once the tile is finished, so the
      // accumulator drain below is branch-light and free of transcendental code.
      float run_lo = 0.f, run_hi = 0.f;            // sums of the filter pair (cur, cur+1)
      float* pp = s_park + r * park_pitch;         // &park[row][cur]
      float* const pp_end = pp + P.n_mels;
      // one 16-bin block: power -> two FMAs per bin; closed filters are parked with a predicated store
      auto drain16 = [&](const uint32_t (&re)[16], const uint32_t (&im)[16], int bin0) {
        const uint32_t advw = s_adv[bin0 >> 4];
        const float4* wp = reinterpret_cast<const float4*>(s_melw + bin0);
#pragma unroll
        for (int j2 = 0; j2 < 8; ++j2) {
          const float4 w2 = wp[j2];
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int j = 2 * j2 + h;
            if (advw & (1u << j)) {                   // warp-uniform: one filter closes before this bin
              *pp++ = run_lo;
              run_lo = run_hi; run_hi = 0.f;
            }
            const float a = __uint_as_float(re[j]), bq = __uint_as_float(im[j]);
            float pw = fmaf(a, a, bq * bq);
            if (kPower1) pw = sqrtf(pw + P.power_eps);
            run_lo = fmaf(h ? w2.z : w2.x, pw, run_lo);
            run_hi = fmaf(h ? w2.w : w2.y, pw, run_hi);
          }
        }
      };
      auto dump16 = [&](const uint32_t (&re)[16], const uint32_t (&im)[16], int bin0) {
        if (kDebug) {
          if (valid) {
            float* d = P.dbg_reim + g * (2 * P.nbt) + bin0;
#pragma unroll
            for (int j = 0; j < 16; ++j) { d[j] = __uint_as_float(re[j]); d[P.nbt + j] = __uint_as_float(im[j]); }
          }
        }
      };
      uint32_t abuf = 0;
      auto release_acc = [&]() {        // accumulators of this pass are in registers: hand TMEM back to the MMA warp
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->acc_empty[abuf]));
      };
      const int nblk = nb >> 4;
      for (int ps = 0; ps < P.passes; ++ps, ++pseq) {
        abuf = pseq % acc_bufs;
        mbar_wait(smem_u32(&bars->acc_full[abuf]), (pseq / acc_bufs) & 1, P.status, 501);
        tc_fence_after();
        const uint32_t t_re = tmem + lane_addr + abuf * kAccStride, t_im = t_re + col_im;
        // two register sets: the TMEM load of block i+1 is in flight while block i is reduced
        uint32_t reA[16], imA[16], reB[16], imB[16];
        tmem_ld16(t_re, reA);
        tmem_ld16(t_im, imA);
        for (int blk = 0; blk < nblk; blk += 2) {
          tmem_wait_ld();                                          // set A = block blk
          if (blk + 1 < nblk) {
            tmem_ld16(t_re + (blk + 1) * 16, reB);
            tmem_ld16(t_im + (blk + 1) * 16, imB);
          } else {
            release_acc();
          }
          dump16(reA, imA, ps * nb + blk * 16);
          if (!(P.exp_flags & 8)) drain16(reA, imA, ps * nb + blk * 16);
          if (blk + 1 < nblk) {
            tmem_wait_ld();                                        // set B = block blk+1
            if (blk + 2 < nblk) {
              tmem_ld16(t_re + (blk + 2) * 16, reA);
              tmem_ld16(t_im + (blk + 2) * 16, imA);
            } else {
              release_acc();
            }
            dump16(reB, imB, ps * nb + (blk + 1) * 16);
            if (!(P.exp_flags & 8)) drain16(reB, imB, ps * nb + (blk + 1) * 16);
          }
        }
      }
      // close the running filters and the (empty) ones above them
      for (; pp < pp_end; ++pp) { *pp = run_lo; run_lo = run_hi; run_hi = 0.f; }
      // raw mel sums -> log(max(., clamp)) / norm, one coalesced output row at a time
      __syncwarp();
#pragma unroll 4
      for (int rr = 0; rr < ((P.exp_flags & 16) ? 0 : 32); ++rr) {
        const unsigned long long pr = __shfl_sync(0xffffffffu, (unsigned long long)(uintptr_t)orow, rr);
        const int ok = __shfl_sync(0xffffffffu, (int)valid, rr);
        if (!ok) continue;
        float* rowp = reinterpret_cast<float*>((uintptr_t)pr);
        const float* parked = s_park + (q * 32 + rr) * park_pitch;
        for (int mth = lane; mth < P.n_mels; mth += 32)
          rowp[mth] = logf(parked[mth] < P.clamp ? P.clamp : parked[mth]) / s_norm[mth];   // true division, as the reference
      }
      __syncwarp();
    }
  }

  // ---- teardown
  tc_fence_before();
  __syncthreads();
  if (warp == kWarpTmem) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)kTmemCols) : "memory");
  }
}

#include "k_logmel2.cuh"

// ------------------------------------------------------------------------------------------
// k_simt: plain CUDA-core fp32 implementation of the same maths (cross-check kernel)
// ------------------------------------------------------------------------------------------
constexpr int kSimtFrames = 8;
constexpr int kSimtThreads = 256;

__global__ void __launch_bounds__(kSimtThreads) k_simt(const float* __restrict__ xp, int64_t Lp,
                                                       const int64_t* __restrict__ frame_off,
                                                       int B, int Tmax, float* __restrict__ out,
                                                       const float* __restrict__ cosf_, const float* __restrict__ sinf__,
                                                       const float* __restrict__ fb, const float* __restrict__ norms,
                                                       int nbt, int n_mels, float clamp, int power_is_1, float power_eps) {
  extern __shared__ float sm[];
  float* sE = sm;                                  // [F][512]
  float* sO = sE + kSimtFrames * kKfold;           // [F][512]
  float* sP = sO + kSimtFrames * kKfold;           // [F][nbt]
  __shared__ int64_t s_out[kSimtFrames];
  const int64_t total = frame_off[B];
  const int64_t g0 = (int64_t)blockIdx.x * kSimtFrames;
  for (int f = 0; f < kSimtFrames; ++f) {
    const int64_t g = g0 + f;
    int64_t base = -1, oidx = -1;
    if (g < total) {
      const int b = find_clip(frame_off, B, g);
      const int64_t t = g - frame_off[b];
      base = (int64_t)b * Lp + t * kHop;
      if (t < Tmax) oidx = ((int64_t)b * Tmax + t) * n_mels;
    }
    if (threadIdx.x == 0) s_out[f] = oidx;
    for (int kap = threadIdx.x; kap < kKfold; kap += kSimtThreads) {
      const int n = kap + 1;
      const float a = base >= 0 ? xp[base + n] : 0.f, c = base >= 0 ? xp[base + kNfft - n] : 0.f;
      sE[f * kKfold + kap] = a + c;
      sO[f * kKfold + kap] = a - c;
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nbt; i += kSimtThreads) {
    float re[kSimtFrames], im[kSimtFrames];
#pragma unroll
    for (int f = 0; f < kSimtFrames; ++f) re[f] = im[f] = 0.f;
    for (int kap = 0; kap < kKfold; ++kap) {
      const float c = cosf_[(size_t)kap * nbt + i], s = sinf__[(size_t)kap * nbt + i];
#pragma unroll
      for (int f = 0; f < kSimtFrames; ++f) {
        re[f] = fmaf(sE[f * kKfold + kap], c, re[f]);
        im[f] = fmaf(sO[f * kKfold + kap], s, im[f]);
      }
    }
#pragma unroll
    for (int f = 0; f < kSimtFrames; ++f) {
      float pw = fmaf(re[f], re[f], im[f] * im[f]);
      if (power_is_1) pw = sqrtf(pw + power_eps);
      sP[f * nbt + i] = pw;
    }
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < kSimtFrames * n_mels; idx += kSimtThreads) {
    const int f = idx / n_mels, m = idx - f * n_mels;
    if (s_out[f] < 0) continue;
    float acc = 0.f;
    for (int i = 0; i < nbt; ++i) acc = fmaf(sP[f * nbt + i], fb[(size_t)i * n_mels + m], acc);
    float y = logf(acc < clamp ? clamp : acc);
    if (norms) y = y / norms[m];
    out[s_out[f] + m] = y;
  }
}

// ------------------------------------------------------------------------------------------
// k_stats: per-channel sum / sumsq / count of log-mel over valid frames (fp64 accumulation)
// ------------------------------------------------------------------------------------------
constexpr int kStatRows = 8;

__global__ void __launch_bounds__(kMaxMels* kStatRows) k_stats(const float* __restrict__ logmel,
                                                               const int32_t* __restrict__ n_frames, int B, int Tmax,
                                                               int n_mels, double* __restrict__ acc) {
  __shared__ double s_sum[kStatRows][kMaxMels], s_sq[kStatRows][kMaxMels];
  const int c = threadIdx.x, ry = threadIdx.y;
  double sum = 0.0, sq = 0.0;
  long long cnt = 0;
  // segments = clips; rows beyond n_frames[b] are padding and are skipped
  for (int b = blockIdx.y; b < B; b += gridDim.y) {
    const int T = min(n_frames[b], Tmax);
    const float* base = logmel + (int64_t)b * Tmax * n_mels;
    for (int t = blockIdx.x * kStatRows + ry; t < T; t += gridDim.x * kStatRows) {
      if (c < n_mels) {
        const double v = (double)base[(int64_t)t * n_mels + c];
        sum += v;
        sq += v * v;
      }
      if (c == 0) ++cnt;
    }
  }
  s_sum[ry][c] = sum;
  s_sq[ry][c] = sq;
  // frame count: warp-shuffle reduction over the lanes that counted, one atomic per warp
  unsigned long long wc = (unsigned long long)cnt;
  for (int d = 16; d > 0; d >>= 1) wc += __shfl_down_sync(0xffffffffu, wc, d);
  __syncthreads();
  if (ry == 0 && c < n_mels) {
    double a = 0.0, q = 0.0;
#pragma unroll
    for (int y = 0; y < kStatRows; ++y) { a += s_sum[y][c]; q += s_sq[y][c]; }
    atomicAdd(acc + c, a);
    atomicAdd(acc + n_mels + c, q);
  }
  if (((threadIdx.y * blockDim.x + threadIdx.x) & 31) == 0 && wc) atomicAdd(acc + 2 * n_mels, (double)wc);
}

// ------------------------------------------------------------------------------------------
// host: launches
// ------------------------------------------------------------------------------------------
namespace {

size_t logmel_smem_bytes(const ttts_mel_plan* p) {
  size_t s = (size_t)kBStages * p->nb * 128 + (size_t)kRawBytes + (size_t)p->nbt * sizeof(float2) + (size_t)(p->nbt / 16 + 4) * 4 +
             kMaxMels * 4 + (size_t)(kTileM * (p->cfg.n_mels + 1) + 2) * 4 + sizeof(LogmelBarriers) + 16;
  return align_up(s, 16) + 1024;   // slack for the 1024 B alignment of the dynamic segment
}

size_t logmel2_smem_bytes(const ttts_mel_plan* p) {
  size_t s = (size_t)(p->pair ? f2::kBStagesPair * (p->nb2 / 2) : f2::kBStages2 * p->nb2) * 128 + (size_t)f2::kConvWarps2 * f2::kDepth * f2::kStageBytes +
             (size_t)p->nbt2 * sizeof(float2) + (size_t)(2 * (p->nbt2 / 16) + 4) * 4 + 512 * 4 + kMaxMels * 4 +
             (size_t)(kTileM * (p->cfg.n_mels + 1) + 2) * 4 + (size_t)2 * f2::kChains * 2 * kTileM * 4 + sizeof(f2::Barriers) + 16;
  s += (size_t)2 * kMaxMels * 8;   // fused statistics
  return align_up(s, 16) + 1024;   // slack for the 1024 B alignment of the dynamic segment
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per function and device: raise it once (it never shrinks), not per call
template <class Kern>
int ensure_dynamic_smem(Kern kern, int slot, int device, size_t smem) {
  static std::atomic<size_t> granted[8][64];
  if (slot < 0 || slot >= 8 || device < 0 || device >= 64) {
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    return TTTS_MEL_OK;
  }
  if (granted[slot][device].load(std::memory_order_acquire) >= smem) return TTTS_MEL_OK;
  CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  size_t cur = granted[slot][device].load();
  while (cur < smem && !granted[slot][device].compare_exchange_weak(cur, smem)) {}
  return TTTS_MEL_OK;
}

enum class Impl { Tcgen05, Simt };

int forward_impl(const ttts_mel_plan* p, const float* wav, int64_t row_stride, const int64_t* lengths, int32_t B,
                 int64_t Lmax, int32_t Tmax, float* out, int32_t* n_frames_user, uint8_t* mask, void* ws_ptr,
                 size_t ws_bytes, void* stream_, Impl impl, float* dbg_reim, double* stat_acc = nullptr) {
  g_launches = 0;
  if (!p || !wav || !out || !ws_ptr) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || Lmax <= 0 || Tmax <= 0) return fail(TTTS_MEL_EINVAL, "empty batch (B=%d Lmax=%lld Tmax=%d)", B, (long long)Lmax, Tmax);
  if (row_stride < Lmax) return fail(TTTS_MEL_EINVAL, "row_stride %lld < Lmax %lld", (long long)row_stride, (long long)Lmax);
  int cur_dev = -1;
  CUDA_TRY(cudaGetDevice(&cur_dev));
  if (cur_dev != p->device) return fail(TTTS_MEL_EINVAL, "plan lives on device %d, the current device is %d", p->device, cur_dev);
  if (ttts_mel_frames(p, Lmax) < 0) return TTTS_MEL_EINVAL;
  const bool in_place = impl == Impl::Tcgen05 && reads_in_place(wav, row_stride, B);
  Workspace w = carve(p, ws_ptr, B, Lmax, in_place);
  if (ws_bytes < w.bytes) return fail(TTTS_MEL_ENOMEM, "workspace %zu < %zu bytes", ws_bytes, w.bytes);
  if ((int64_t)B * w.Lp >= (int64_t)1 << 31) return fail(TTTS_MEL_EINVAL, "batch too large for 32-bit row offsets: split it");
  cudaStream_t stream = (cudaStream_t)stream_;

  const int vec_ok = (row_stride % 4 == 0) && (((uintptr_t)wav & 15) == 0);
  // k_logmel reads the caller's waveform directly when it is 16 B-addressable; only the frames that touch
  // the reflect padding go through a small edge buffer.  Otherwise (and for k_simt) the clips are padded
  // into the workspace first.
  const bool direct = in_place;
#if TTTS_PREP_PDL
  {
    cudaLaunchConfig_t lc{};
    lc.gridDim = dim3((unsigned)(kPrepSplit * B + 1)); lc.blockDim = dim3(1024); lc.dynamicSmemBytes = 0; lc.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = 1;
    CUDA_TRY(cudaLaunchKernelEx(&lc, k_prep, wav, row_stride, direct ? w.edge : (float*)nullptr, out, (int)p->cfg.n_mels, mask, lengths,
                                (int)B, Lmax, (int)p->cfg.pad, (int)Tmax, w.n_frames, n_frames_user, w.frame_off, w.group_off,
                                w.status, stat_acc ? w.stat : (unsigned long long*)nullptr));
  }
#else
  k_prep<<<kPrepSplit * B + 1, 1024, 0, stream>>>(wav, row_stride, direct ? w.edge : nullptr, out, p->cfg.n_mels, mask, lengths, B, Lmax,
                                     p->cfg.pad, Tmax, w.n_frames, n_frames_user, w.frame_off, w.group_off, w.status,
                                     stat_acc ? w.stat : nullptr);
#endif
  ++g_launches;
  if (!direct) {
    const int64_t vecs = (w.Lp + 3) / 4;
    // grid-stride over each clip's own extent: with ragged lengths most blocks of a (Lp / 1024)-wide grid had nothing to do
    const int64_t per_clip = std::min<int64_t>((vecs + 255) / 256, B >= 64 ? 48 : (B >= 8 ? 256 : 4096));
    dim3 grid((unsigned)per_clip, (unsigned)B);
    k_pad<<<grid, 256, 0, stream>>>(wav, row_stride, lengths, Lmax, p->cfg.pad, w.xp, w.Lp, vec_ok);
    ++g_launches;
  }
  if (impl == Impl::Tcgen05 && p->algo == 2) {
    f2::Params P{};
    P.src = direct ? wav : w.xp; P.src_stride = direct ? row_stride : w.Lp; P.edge = w.edge;
    P.lengths = lengths; P.Lmax = Lmax; P.direct = direct; P.pad = p->cfg.pad;
    {
      const int64_t Tu = ttts_mel_frames(p, Lmax);
      P.T_uniform = lengths ? 0 : (int)(Tu < Tmax ? Tu : Tmax);
      P.G_uniform = (P.T_uniform + f2::kGroup - 1) / f2::kGroup;
    }
    P.group_off = w.group_off; P.frame_off = w.frame_off; P.n_frames = w.n_frames; P.B = B; P.Tmax = Tmax;
    P.out = out; P.basis = p->d_basis2; P.melw = p->d_melw2; P.meladv = p->d_meladv2; P.norms = p->d_norms;
    P.win = p->d_win2; P.wc = p->wc;
    P.passes = 2 * p->sp; P.sp = p->sp; P.nb = p->nb2; P.nbt = p->nbt2; P.n_mels = p->cfg.n_mels;
    P.clamp = p->cfg.clamp; P.power_eps = p->cfg.power_eps; P.status = w.status;
    P.dbg_reim = dbg_reim; P.dbg_nbt = p->nbt; P.dbg_kmin = p->kmin; P.kmax = p->kmax;
    P.first_k[0] = p->first_k[0]; P.first_k[1] = p->first_k[1];
    for (int i = 0; i < 16; ++i) { P.cstart[i] = p->cstart[i]; P.cend[i] = p->cend[i]; }
    static_assert(f2_kChains == f2::kChains, "chain count");
    P.exp_flags = TTTS_EXP_FLAGS();
    P.stat_ws = stat_acc ? w.stat : nullptr; P.stat_acc = stat_acc;
    const size_t smem = logmel2_smem_bytes(p);
    const bool pw1 = p->cfg.power == 1.0f;
    const bool pair = p->pair && !dbg_reim && p->num_sms >= 2;
    if (pair) P.basis = p->d_basis2p;
    void (*kern)(const f2::Params) =
        dbg_reim ? (pw1 ? k_logmel2<true, true, false> : k_logmel2<false, true, false>)
        : pair   ? (pw1 ? k_logmel2<true, false, true> : k_logmel2<false, false, true>)
                 : (pw1 ? k_logmel2<true, false, false> : k_logmel2<false, false, false>);
    {
      const int rc = ensure_dynamic_smem(kern, (dbg_reim ? 4 : pair ? 2 : 0) + (pw1 ? 1 : 0), p->device, smem);
      if (rc != TTTS_MEL_OK) return rc;
    }
    ProfileRing& pr = const_cast<ttts_mel_plan*>(p)->prof;
    const int pslot = pr.capacity > 0 ? pr.count.fetch_add(1) : -1;
    const bool prof = pslot >= 0 && pslot < pr.capacity;
    if (prof) cudaEventRecord(pr.start[pslot], stream);
    {
      // programmatic dependent launch behind k_prep: the prologue of the persistent kernel overlaps it
      cudaLaunchConfig_t lc{};
      lc.gridDim = dim3((unsigned)(pair ? p->num_sms & ~1 : p->num_sms)); lc.blockDim = dim3(f2::kThreads2);
      lc.dynamicSmemBytes = smem; lc.stream = stream;
      cudaLaunchAttribute at[2];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      at[1].id = cudaLaunchAttributeClusterDimension;           // CTA pair = the two SMs of a TPC
      at[1].val.clusterDim.x = 2; at[1].val.clusterDim.y = 1; at[1].val.clusterDim.z = 1;
      lc.attrs = at; lc.numAttrs = pair ? 2 : 1;
      CUDA_TRY(cudaLaunchKernelEx(&lc, kern, P));
    }
    if (prof) cudaEventRecord(pr.stop[pslot], stream);
    ++g_launches;
  } else if (impl == Impl::Tcgen05) {
    if (stat_acc) return fail(TTTS_MEL_EINVAL, "fused statistics need the k_logmel2 kernel");
    LogmelParams P{};
    P.src = direct ? wav : w.xp; P.src_stride = direct ? row_stride : w.Lp; P.edge = w.edge;
    P.lengths = lengths; P.Lmax = Lmax; P.direct = direct; P.pad = p->cfg.pad;
    {
      const int64_t Tu = ttts_mel_frames(p, Lmax);
      P.T_uniform = lengths ? 0 : (int)(Tu < Tmax ? Tu : Tmax);
    }
    P.frame_off = w.frame_off; P.n_frames = w.n_frames; P.B = B; P.Tmax = Tmax;
    P.out = out; P.basis = p->d_basis; P.melw = p->d_melw; P.meladv = p->d_meladv; P.norms = p->d_norms;
    P.passes = p->passes; P.nb = p->nb; P.nbt = p->nbt; P.n_mels = p->cfg.n_mels;
    P.clamp = p->cfg.clamp; P.power_eps = p->cfg.power_eps; P.power_is_1 = p->cfg.power == 1.0f; P.status = w.status; P.dbg_reim = dbg_reim; P.exp_flags = TTTS_EXP_FLAGS();
    const size_t smem = logmel_smem_bytes(p);
    void (*kern)(const LogmelParams) =
        dbg_reim ? (P.power_is_1 ? k_logmel<true, true> : k_logmel<false, true>)
                 : (P.power_is_1 ? k_logmel<true, false> : k_logmel<false, false>);
    CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));   // lab path only
    ProfileRing& pr = const_cast<ttts_mel_plan*>(p)->prof;
    const int pslot = pr.capacity > 0 ? pr.count.fetch_add(1) : -1;
    const bool prof = pslot >= 0 && pslot < pr.capacity;
    if (prof) cudaEventRecord(pr.start[pslot], stream);
    kern<<<p->num_sms, kThreads, smem, stream>>>(P);
    if (prof) cudaEventRecord(pr.stop[pslot], stream);
    ++g_launches;
  } else {
    if (stat_acc) return fail(TTTS_MEL_EINVAL, "fused statistics need the k_logmel2 kernel");
    const int64_t Tall = ttts_mel_frames(p, Lmax);
    const int64_t max_frames = (int64_t)B * Tall;
    const unsigned blocks = (unsigned)((max_frames + kSimtFrames - 1) / kSimtFrames);
    const size_t smem = (size_t)(2 * kSimtFrames * kKfold + kSimtFrames * p->nbt) * 4;
    static thread_local size_t configured = 0;
    if (configured < smem) {
      CUDA_TRY(cudaFuncSetAttribute(k_simt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      configured = smem;
    }
    k_simt<<<blocks, kSimtThreads, smem, stream>>>(w.xp, w.Lp, w.frame_off, B, Tmax, out, p->d_cosf, p->d_sinf, p->d_fb,
                                                  p->d_norms, p->nbt, p->cfg.n_mels, p->cfg.clamp, p->cfg.power == 1.0f, p->cfg.power_eps);
    ++g_launches;
  }
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}

}  // namespace

extern "C" int ttts_mel_forward(const ttts_mel_plan* p, const float* wav, int64_t row_stride, const int64_t* lengths,
                                int32_t B, int64_t Lmax, int32_t Tmax, float* out, int32_t* n_frames, uint8_t* mask,
                                void* ws, size_t ws_bytes, void* stream) {
  return forward_impl(p, wav, row_stride, lengths, B, Lmax, Tmax, out, n_frames, mask, ws, ws_bytes, stream,
                      Impl::Tcgen05, nullptr);
}

extern "C" int ttts_mel_forward_stats(const ttts_mel_plan* p, const float* wav, int64_t row_stride,
                                      const int64_t* lengths, int32_t B, int64_t Lmax, int32_t Tmax, float* out,
                                      int32_t* n_frames, uint8_t* mask, void* ws, size_t ws_bytes, void* stream,
                                      double* acc) {
  if (!acc) return fail(TTTS_MEL_EINVAL, "acc_dev is null");
  return forward_impl(p, wav, row_stride, lengths, B, Lmax, Tmax, out, n_frames, mask, ws, ws_bytes, stream,
                      Impl::Tcgen05, nullptr, acc);
}

// ------------------------------------------------------------------------------------------
// ttts_mel_forward_static: the workspace in a file-scope device array (no allocation anywhere on the call path).
// No counterpart in the reference (include/ttts_mel.h).
// In HOST code `g_static_ws` is only a handle for cudaGetSymbolAddress -- never a device address: the address is resolved
// once per device, with that device current, and it is that pointer the kernels get.
// ------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_ws[TTTS_MEL_STATIC_WS_BYTES];

namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
void* g_static_ptr[kStaticMaxDevices] = {};

int static_workspace(int device, void** out) {
  if (device < 0 || device >= kStaticMaxDevices) return fail(TTTS_MEL_EINVAL, "device %d out of range", device);
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_ws);     // the current device is `device` (checked by the caller)
    if (e != cudaSuccess || !ptr)
      return fail(TTTS_MEL_ECUDA, "cudaGetSymbolAddress(static workspace) on device %d: %s", device, cudaGetErrorString(e));
    g_static_ptr[device] = ptr;
  }
  *out = g_static_ptr[device];
  return TTTS_MEL_OK;
}
}  // namespace

extern "C" int ttts_mel_forward_static(const ttts_mel_plan* p, const float* wav, int64_t row_stride,
                                       const int64_t* lengths, int32_t B, int64_t Lmax, int32_t Tmax, float* out,
                                       int32_t* n_frames, uint8_t* mask, void* stream, const void** ws_out) {
  if (ws_out) *ws_out = nullptr;
  if (!p || !wav || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || Lmax <= 0 || Tmax <= 0) return fail(TTTS_MEL_EINVAL, "empty batch (B=%d Lmax=%lld Tmax=%d)", B, (long long)Lmax, Tmax);
  const size_t need = carve(p, nullptr, B, Lmax, reads_in_place(wav, row_stride, B)).bytes;
  if (need > (size_t)TTTS_MEL_STATIC_WS_BYTES)
    return fail(TTTS_MEL_ENOMEM, "this batch needs a %zu-byte workspace, the static one holds %zu: call ttts_mel_forward "
                "with a caller workspace", need, (size_t)TTTS_MEL_STATIC_WS_BYTES);
  int cur_dev = -1;
  CUDA_TRY(cudaGetDevice(&cur_dev));
  if (cur_dev != p->device) return fail(TTTS_MEL_EINVAL, "plan lives on device %d, the current device is %d", p->device, cur_dev);
  void* ws = nullptr;
  const int rc = static_workspace(cur_dev, &ws);
  if (rc != TTTS_MEL_OK) return rc;
  if (ws_out) *ws_out = ws;
  return forward_impl(p, wav, row_stride, lengths, B, Lmax, Tmax, out, n_frames, mask, ws, (size_t)TTTS_MEL_STATIC_WS_BYTES,
                      stream, Impl::Tcgen05, nullptr);
}

extern "C" size_t ttts_mel_config_size(void) { return sizeof(ttts_mel_config); }

#ifdef TTTS_MEL_LAB
extern "C" int ttts_mel_forward_simt(const ttts_mel_plan* p, const float* wav, int64_t row_stride,
                                     const int64_t* lengths, int32_t B, int64_t Lmax, int32_t Tmax, float* out,
                                     int32_t* n_frames, uint8_t* mask, void* ws, size_t ws_bytes, void* stream) {
  return forward_impl(p, wav, row_stride, lengths, B, Lmax, Tmax, out, n_frames, mask, ws, ws_bytes, stream,
                      Impl::Simt, nullptr);
}

extern "C" int ttts_mel_forward_debug(const ttts_mel_plan* p, const float* wav, int64_t row_stride,
                                      const int64_t* lengths, int32_t B, int64_t Lmax, int32_t Tmax, float* out,
                                      int32_t* n_frames, uint8_t* mask, void* ws, size_t ws_bytes, void* stream,
                                      float* reim_out) {
  if (!reim_out) return fail(TTTS_MEL_EINVAL, "reim_out_dev is null");
  return forward_impl(p, wav, row_stride, lengths, B, Lmax, Tmax, out, n_frames, mask, ws, ws_bytes, stream,
                      Impl::Tcgen05, reim_out);
}

#endif  // TTTS_MEL_LAB

extern "C" int ttts_mel_status(const void* ws, void* stream) {
  if (!ws) return fail(TTTS_MEL_EINVAL, "null workspace");
  cudaError_t e = cudaStreamSynchronize((cudaStream_t)stream);
  if (e != cudaSuccess) return fail(TTTS_MEL_ECUDA, "stream sync: %s", cudaGetErrorString(e));
  int32_t code = 0;
  CUDA_TRY(cudaMemcpy(&code, ws, sizeof(code), cudaMemcpyDeviceToHost));
  if (code == 0) return TTTS_MEL_OK;
  if (code == 1) return fail(TTTS_MEL_EDEVICE, "a clip length is <= pad, shorter than the window, or > Lmax");
  if (code == 2) return fail(TTTS_MEL_EDEVICE, "Tmax is smaller than a clip's frame count");
  return fail(TTTS_MEL_EDEVICE, "pipeline wait timed out (code %d)", code);
}

extern "C" int ttts_mel_stats_accumulate(const ttts_mel_plan* p, const float* logmel, const int32_t* n_frames,
                                         int32_t B, int32_t Tmax, double* acc, void* stream) {
  if (!p || !logmel || !n_frames || !acc) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || Tmax <= 0) return fail(TTTS_MEL_EINVAL, "empty batch");
  dim3 block(kMaxMels, kStatRows);
  int gx = (Tmax + kStatRows * 8 - 1) / (kStatRows * 8);
  if (gx < 1) gx = 1;
  if (gx > 64) gx = 64;
  int gy = B < 64 ? B : 64;
  k_stats<<<dim3(gx, gy), block, 0, (cudaStream_t)stream>>>(logmel, n_frames, B, Tmax, p->cfg.n_mels, acc);
  g_launches = 1;
  CUDA_TRY(cudaGetLastError());
  return TTTS_MEL_OK;
}

// ------------------------------------------------------------------------------------------
// measurement hooks (include/ttts_mel_lab.h): CUDA-event duration of the fused kernel alone, per forward call of ONE plan,
// on the launch stream.  They never change what a forward computes.
// ------------------------------------------------------------------------------------------
static void profile_free(ttts_mel_plan* p) {
  for (int i = 0; i < p->prof.capacity; ++i) { cudaEventDestroy(p->prof.start[i]); cudaEventDestroy(p->prof.stop[i]); }
  delete[] p->prof.start; delete[] p->prof.stop;
  p->prof.start = p->prof.stop = nullptr;
  p->prof.capacity = 0;
  p->prof.count.store(0);
}

extern "C" int ttts_mel_profile_enable(ttts_mel_plan* p, int32_t max_calls) {
  if (!p) return fail(TTTS_MEL_EINVAL, "null plan");
  std::lock_guard<std::mutex> lk(p->prof_mu);
  profile_free(p);
  if (max_calls <= 0) return TTTS_MEL_OK;
  p->prof.start = new cudaEvent_t[max_calls];
  p->prof.stop = new cudaEvent_t[max_calls];
  for (int i = 0; i < max_calls; ++i) {
    CUDA_TRY(cudaEventCreate(&p->prof.start[i]));
    CUDA_TRY(cudaEventCreate(&p->prof.stop[i]));
  }
  p->prof.capacity = max_calls;
  return TTTS_MEL_OK;
}

extern "C" int ttts_mel_profile_collect(ttts_mel_plan* p, float* ms_out, int32_t capacity) {
  if (!p || !ms_out) return fail(TTTS_MEL_EINVAL, "null argument");
  std::lock_guard<std::mutex> lk(p->prof_mu);
  CUDA_TRY(cudaDeviceSynchronize());
  int n = p->prof.count.load();
  if (n > p->prof.capacity) n = p->prof.capacity;
  if (n > capacity) n = capacity;
  for (int i = 0; i < n; ++i) CUDA_TRY(cudaEventElapsedTime(&ms_out[i], p->prof.start[i], p->prof.stop[i]));
  p->prof.count.store(0);
  return n;
}

#ifdef TTTS_MEL_LAB
extern "C" void ttts_mel_set_experiment(int32_t flags) { g_exp_flags.store(flags); }
#endif

// ------------------------------------------------------------------------------------------
// ingestion in front of the path (int16 PCM -> float32, sinc resample): include/ttts_ingest.h
// ------------------------------------------------------------------------------------------
#include "k_ingest.cuh"
#include "k_stft.cuh"
#include "k_vq.cuh"
