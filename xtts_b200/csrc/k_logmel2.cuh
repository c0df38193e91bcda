// k_logmel2.cuh -- second-generation fused log-mel kernel ("double fold"), included by ttts_mel.cu.
//
// Same contract as k_logmel (ttts_mel.cu) and the same reference chain (ttts/utils/utils.py:509-522 via
// torchaudio MelSpectrogram / torch.stft), but the DFT-GEMM contracts over K = 256 (+1) instead of 512:
//
//   first fold (as k_logmel)    E'[n] = w[n] (f[n] + f[N-n]),  O'[n] = w[n] (f[n] - f[N-n]),   n = 1..512
//   second fold (new)           cos(2 pi k (512-n)/N) = (-1)^k cos(2 pi k n/N),  sin(..(512-n)..) = -(-1)^k sin(..n..)
//        even k:  Re[k] = sum_{n=1..256} (E'[n] + E'[512-n]) cos(th_kn) + E'[512]
//                 Im[k] = sum_{n=1..256} (O'[n] - O'[512-n]) sin(th_kn)
//        odd  k:  Re[k] = sum_{n=1..256} (E'[n] - E'[512-n]) cos(th_kn) - E'[512]
//                 Im[k] = sum_{n=1..256} (O'[n] + O'[512-n]) sin(th_kn)
//   (the n = 256 slot pairs with itself: its partner weight is 0).  Even and odd bins are separate accumulator
//   passes with different A operands, so tensor work, basis bytes and basis L2 traffic all halve; the window moves
//   from the basis into the converter (two FMULs per tap), the basis is the pure TF32 hi/lo split of cos/sin.
//   The centre tap E'[512] = w[512] f[512] is a 33rd k-step that carries one live column (basis +-1).
//
// Pipeline differences from k_logmel:
//   * converters work at k-step granularity: kSets sets x 4 lane quarters; set s owns A stage s and every k-step with
//     sequence number == s (mod kSets).  Per k-step a warp stages, for its 32 rows, 34 "slots" of 20 floats:
//     slot v = [ frame v's taps 8ks..8ks+11 | frame v-2's taps 1016-8ks..1023-8ks ].  Because hop = N/4, row r's
//     four windows are slot r (a = f[n], c = f[512-n]) and slot r+2 (d = f[512+n], b = f[N-n]): every sample is
//     staged once per pass instead of twice.  This needs the 32 rows of a warp to be consecutive frames of ONE clip,
//     so clips are laid out in groups of 32 frames (the last group of a clip is partly empty).
//   * the epilogue restarts its running-filter scan per parity and adds the second parity onto the parked sums.
#pragma once

namespace f2 {

#ifndef F2_SETS
#define F2_SETS 3
#endif
#ifndef F2_CHAINS
#define F2_CHAINS 4
#endif
constexpr int kSets = F2_SETS;                   // converter sets: set s serves the k-steps with sequence number == s (mod kSets)
#ifndef F2_ASTAGES
#define F2_ASTAGES 4
#endif
#ifndef F2_ACC_STRIDE
#define F2_ACC_STRIDE kNBMax
#endif
constexpr int kAStages2 = F2_ASTAGES;            // A-operand TMEM stages (k-step sequence number mod 4): 128 columns
constexpr int kAccStride2 = F2_ACC_STRIDE;        // columns between the Re and Im accumulators; the A ring starts at twice that
                                                 // (8 stages in V2's spare columns: no gain, and the run-time depth cost V1 3 %: profiles/r2g_variants.txt)
#ifndef F2_CENTRE_EPI
#define F2_CENTRE_EPI 2
#endif
// The centre tap E'[512] = w[512] f[512] (basis +-1: a rank-1 term) is added by the epilogue instead of costing a 33rd
// k-step (handshake, conversion, 3 MMAs): -5.7 % on the magnitude plans (V2), -1.7 % on V1 (profiles/r3b, r3c).  Magnitude
// plans add it with one FADD.  Power plans (V1) repeat the two accumulations the 33rd k-step performed -- D += tf32(e_lo) * (+-1),
// then D += e_hi * (+-1), the operand truncated as the tensor core truncates it: a plain FADD changed the rounding enough to
// move ONE element of the committed tone golden from 0.959 to 1.006 of its dual-oracle bound (a mel value of 2e-5 next to
// the clamp floor, cancellation regime: tools/gpu_tone_stage_diag.py); with the two-step form that element is
// bit-identical to the 33-k-step kernel's.  2: every plan; 1: magnitude plans only; 0: none (33 k-steps).
constexpr int kCentreEpiMode = F2_CENTRE_EPI;
constexpr int kKStab = 33;                       // basis stages per pass in the plan's table: 32 x 8 folded taps + the centre tap
constexpr int kSlots = 34;                       // staged slots per warp and k-step (32 rows + 2)
constexpr int kSlotFloats = 20;                  // 12 forward taps + 8 mirrored taps; 80 B pitch: conflict-free LDS.128
constexpr int kStageBytes = kSlots * kSlotFloats * 4;
#ifndef F2_DEPTH
#define F2_DEPTH 2
#endif
constexpr int kDepth = F2_DEPTH;                 // raw-audio stages per converter warp
constexpr int kPieces = kSlots * 5;              // 16 B pieces per stage
constexpr int kPieceIters = (kPieces + 31) / 32;
constexpr int kConvWarps2 = 4 * kSets;
constexpr int kWarpConv02 = 4;
constexpr int kWarpEpi02 = kWarpConv02 + kConvWarps2;
constexpr int kEpiWarps = 4 * F2_CHAINS;           // 4 lane quarters x kChains scan chains
constexpr int kThreads2 = (kWarpEpi02 + kEpiWarps) * 32;
// registers per thread by role (setmaxnreg, per warpgroup): 128*40 + 384*96 + 512*40 = 62464 <= 65536; launch value 64
#ifndef F2_REGS_MISC
#define F2_REGS_MISC 40
#endif
#ifndef F2_REGS_CONV
#define F2_REGS_CONV 96
#endif
#ifndef F2_REGS_EPI
#define F2_REGS_EPI 40
#endif
constexpr int kRegsMisc = F2_REGS_MISC, kRegsConv = F2_REGS_CONV, kRegsEpi = F2_REGS_EPI;
static_assert(128 * kRegsMisc + 128 * kSets * kRegsConv + 128 * F2_CHAINS * kRegsEpi <= 65536, "register budgets exceed the file");
// magnitude (power 1) instantiations: the epilogue holds its chain's last 32 bins as power values in registers and hands
// the accumulators back before scanning them (measured, profiles/r2e_variants_hold_sqrt.txt: V2 -3 %, V1 +1.6 %: V1 keeps
// the streaming scan and round 1's budgets)
#ifndef F2_REGS_CONV_P1
#define F2_REGS_CONV_P1 80
#endif
#ifndef F2_REGS_EPI_P1
#define F2_REGS_EPI_P1 56
#endif
#ifndef F2_HOLD_P1
#define F2_HOLD_P1 32
#endif
static_assert(128 * kRegsMisc + 128 * kSets * F2_REGS_CONV_P1 + 128 * F2_CHAINS * F2_REGS_EPI_P1 <= 65536, "register budgets exceed the file");
constexpr int kBStages2 = 4;                      // basis ring depth (single CTA: 4 x 24 KB)
#ifndef F2_BSTAGES_PAIR
#define F2_BSTAGES_PAIR 8
#endif
constexpr int kBStagesPair = F2_BSTAGES_PAIR;     // ... of a CTA pair (8 x 12 KB): the peer's half arrives through a relay
constexpr int kColA2 = 2 * kAccStride2;            // A ring behind the accumulators
static_assert(kColA2 + 32 * kAStages2 <= 512, "TMEM columns");
constexpr int kChains = F2_CHAINS;               // mel-scan chains per pass, one epilogue warp each (per lane quarter)
static_assert(kThreads2 <= 1024 && kAStages2 > kSets, "at most 32 warps; the operand ring is deeper than the number of converter sets");
#ifndef F2_QR_UNROLL
#define F2_QR_UNROLL 1
#endif
constexpr int kQrUnroll = F2_QR_UNROLL;          // unrolling of the epilogue's 4-bin rounds
constexpr int kGroup = 32;                       // frames per group (one converter / epilogue warp)
#ifndef F2_HOLD
#define F2_HOLD 0
#endif

#ifndef F2_FAST_LOG
#define F2_FAST_LOG 1
#endif
#ifndef F2_SQRT_APPROX
#define F2_SQRT_APPROX 1
#endif
static_assert(F2_HOLD % 16 == 0 && F2_HOLD >= 0 && F2_HOLD <= 48 && F2_HOLD_P1 % 16 == 0 && F2_HOLD_P1 <= 48, "F2_HOLD");

struct Params {
  const float* src;           // direct: the caller's waveform [B][src_stride]; else workspace xp [B][Lp]
  int64_t src_stride;
  const float* edge;          // [B][kEdgeClip] (direct mode)
  const int64_t* lengths;     // [B] or null
  int64_t Lmax;
  int direct, pad;
  int T_uniform, G_uniform;   // frames / groups per clip when lengths == null, else 0
  const int64_t* group_off;   // [B+1] exclusive prefix of ceil(n_frames / 32)
  const int64_t* frame_off;   // [B+1] (debug dump only)
  const int32_t* n_frames;    // [B]
  int B, Tmax;
  float* out;                 // [B][Tmax][n_mels]
  const float* basis;         // [passes][33][C_hi,C_lo,S_hi,S_lo][2][nb][4]
  const float2* melw;         // [nbt]
  const uint32_t* meladv;     // [2][nbt/16]: (bit j: one filter closes before bin j | bit 16+j: 2..3 close), 2-bit counts
  const float* norms;         // [n_mels] or null
  const float* win;           // [512]: w[n] (n = 1..256) | w[512-n] (n = 1..255), 0
  float wc;                   // w[512]
  int passes, sp, nb, nbt, n_mels;
  float clamp, power_eps;
  int32_t* status;
  unsigned long long* stat_ws;  // fused statistics: [2 * kMaxMels] fixed-point sums of this launch + [1] CTA ticket (zeroed by k_prep), or null
  double* stat_acc;           // [2 * n_mels + 1] fp64, += at the end of the launch by the last CTA
  float* dbg_reim;            // optional [total_frames][2*dbg_nbt], bins dbg_kmin..
  int dbg_nbt, dbg_kmin, kmax;
  int first_k[2];             // first live bin of each parity
  int cstart[16];             // [passes][kChains] filter index of each scan chain's first bin
  int cend[16];               // [passes][kChains] first of the two filters still open when the chain ends
  int exp_flags;
};

struct __align__(8) Barriers {
  uint64_t b_full[kBStagesPair], b_empty[kBStagesPair];
  uint64_t b_peer[kBStagesPair];                      // CTA pair: the peer CTA's half of the basis stage has landed
  uint64_t a_full[kAStages2], a_empty[kAStages2];
  uint64_t acc_full, acc_empty;
};

struct GroupInfo {
  bool valid;
  int b, t0, T;
};

__device__ __forceinline__ GroupInfo locate_group(const Params& P, int64_t G, int64_t total_groups) {
  GroupInfo gi{false, 0, 0, 0};
  if (G >= total_groups) return gi;
  gi.valid = true;
  if (P.T_uniform) {
    gi.b = (int)(G / P.G_uniform);
    gi.t0 = (int)(G - (int64_t)gi.b * P.G_uniform) * kGroup;
    gi.T = P.T_uniform;
  } else {
    gi.b = find_clip(P.group_off, P.B, G);
    gi.t0 = (int)(G - P.group_off[gi.b]) * kGroup;
    gi.T = P.n_frames[gi.b];
  }
  return gi;
}

// Offset (in floats, relative to P.edge when *edge else P.src) of padded sample 256*tv + woff of clip b.
__device__ __forceinline__ int32_t sample_offset(const Params& P, int b, int T, int tv, int woff, bool* edge) {
  *edge = false;
  if (!P.direct) return (int32_t)((int64_t)b * P.src_stride + (int64_t)kHop * tv + woff);
  if (T <= 8 || tv < 8) { *edge = true; return b * kEdgeClip + kHop * tv + woff; }
  if (tv >= T - 4) { *edge = true; return b * kEdgeClip + kEdgeA + kHop * (tv - (T - 4)) + woff; }
  return (int32_t)((int64_t)b * P.src_stride + (int64_t)kHop * tv - P.pad + woff);
}

__device__ __forceinline__ void st_s(uint32_t addr, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory"); }
__device__ __forceinline__ float ld_s(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}

// Rare event of the mel scan (two or three filters close before a bin: narrow low-frequency filters) and the close
// of a chain: add the running pair to the parked sums with fresh loads (a neighbouring chain may have stored to the same
// two slots).  Out of line on purpose, the accumulator drain stays short.
__device__ __noinline__ void park_multi(uint32_t pa, float lo, float hi) {
  st_s(pa, ld_s(pa) + lo);
  st_s(pa + 4, ld_s(pa + 4) + hi);
}

// ---- the mel scan of four consecutive bins (one tcgen05.ld round), specialised by the block's close pattern.
// Which filters close before which bin is warp-uniform and known to the plan, so instead of testing two mask bits and
// stepping over seven predicated-off instructions at EVERY bin (round 2's ncu: ~25 issued instructions per bin, the
// accumulator drain issue-bound at ~3.7 us per pass), the block's 4-bit pattern selects one of 16 straight-line bodies:
// 8 FMAs plus 4 instructions per close (the shifts run_lo = run_hi, run_hi = 0 fold into the FMAs).
struct ScanState { float run_lo, run_hi, old; uint32_t pa; };

template <int NIB>
__device__ __forceinline__ void scan4_body(ScanState& s, const float (&pw)[4], const float4& wa, const float4& wb) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (NIB & (1 << j)) {                      // one filter closes before this bin
      st_s(s.pa, s.old + s.run_lo);
      s.run_lo = s.run_hi; s.run_hi = 0.f;
      s.pa += 4;
      s.old = ld_s(s.pa);                      // consumed at the next close
    }
    const float wl = j == 0 ? wa.x : j == 1 ? wa.z : j == 2 ? wb.x : wb.z;
    const float wh = j == 0 ? wa.y : j == 1 ? wa.w : j == 2 ? wb.y : wb.w;
    s.run_lo = fmaf(wl, pw[j], s.run_lo);
    s.run_hi = fmaf(wh, pw[j], s.run_hi);
  }
}
__device__ __forceinline__ void scan4_switch(ScanState& s, const float (&pw)[4], const float4& wa, const float4& wb, uint32_t nib) {
  switch (nib) {
    case 0: scan4_body<0>(s, pw, wa, wb); break;
    case 1: scan4_body<1>(s, pw, wa, wb); break;
    case 2: scan4_body<2>(s, pw, wa, wb); break;
    case 3: scan4_body<3>(s, pw, wa, wb); break;
    case 4: scan4_body<4>(s, pw, wa, wb); break;
    case 5: scan4_body<5>(s, pw, wa, wb); break;
    case 6: scan4_body<6>(s, pw, wa, wb); break;
    case 7: scan4_body<7>(s, pw, wa, wb); break;
    case 8: scan4_body<8>(s, pw, wa, wb); break;
    case 9: scan4_body<9>(s, pw, wa, wb); break;
    case 10: scan4_body<10>(s, pw, wa, wb); break;
    case 11: scan4_body<11>(s, pw, wa, wb); break;
    case 12: scan4_body<12>(s, pw, wa, wb); break;
    case 13: scan4_body<13>(s, pw, wa, wb); break;
    case 14: scan4_body<14>(s, pw, wa, wb); break;
    default: scan4_body<15>(s, pw, wa, wb); break;
  }
}
// ---- CTA-pair (cta_group::2) helpers
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the mbarrier at the same shared-memory offset in CTA `cta` of the cluster
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t local_bar, uint32_t cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_bar), "r"(cta));
  // default semantics (release at CTA scope), as CUTLASS ClusterBarrier::arrive: what is handed over lives in TMEM / was
  // written by TMA and is ordered by the tcgen05 fences and the transaction barrier; a cluster-scope release would cost
  // MEMBAR.ALL.GPU + CCTL.IVALL per arrive
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}
__device__ __forceinline__ uint32_t mbar_try_wait_cluster(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok;
}
// bounded wait with cluster-scope acquire (barriers the peer CTA arrives on)
__device__ __noinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity, int32_t* status, int code) {
  const long long t0 = clock64();
  while (!mbar_try_wait_hint(bar, parity)) {
    if (clock64() - t0 > kTimeoutCycles) mbar_timeout(status, code);
  }
}
// D[tmem, both CTAs] (+)= A[tmem, both CTAs] * B[smem halves of both CTAs]^T, kind::tf32, M = 256, K = 8; leader CTA only
__device__ __forceinline__ void mma_tf32_ts2(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::tf32 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// completion of the pair's MMAs -> one arrival on the barrier at this offset in BOTH CTAs
__device__ __forceinline__ void tc_commit2(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(bar), "h"((uint16_t)3) : "memory");
}

}  // namespace f2

// kPair: two CTAs of a cluster (the two SMs of a TPC) work on adjacent tiles with ONE tcgen05.mma.cta_group::2 stream:
// M = 256 (128 rows per CTA, each CTA's A operand and accumulators in its own TMEM), and each CTA stages only HALF of
// every basis stage (96 of the 192 bins) -- the tensor cores of both SMs read both halves.  Per SM that halves the
// basis L2 -> smem traffic, the smem space of the basis ring and the B-operand smem reads.  CTA 0 issues the MMAs; the
// peer's converters / epilogue arrive on CTA 0's barriers, tcgen05.commit multicasts to both CTAs' barriers.
template <bool kPower1, bool kDebug, bool kPair>
__global__ void __launch_bounds__(f2::kThreads2, 1) k_logmel2(const f2::Params P) {
  using namespace f2;
  constexpr bool kCentreEpi = kCentreEpiMode == 2 || (kCentreEpiMode == 1 && kPower1);
  constexpr int kKS = kCentreEpi ? 32 : 33;                // k-steps per pass the pipeline runs
  constexpr int kBS = kPair ? kBStagesPair : kBStages2;    // basis ring depth
  constexpr int kRegsConvT = kPower1 ? F2_REGS_CONV_P1 : kRegsConv, kRegsEpiT = kPower1 ? F2_REGS_EPI_P1 : kRegsEpi;
  constexpr int kHold = kPower1 ? F2_HOLD_P1 : F2_HOLD;    // trailing bins of an epilogue chain held in registers
  extern __shared__ __align__(1024) uint8_t smem[];
  // ---- carve shared memory
  const int nb = P.nb;
  const uint32_t cta_rank = kPair ? cluster_ctarank() : 0u;
  const int nbh = kPair ? nb / 2 : nb;                      // basis rows (bins) staged by this CTA
  const uint32_t b_stage_bytes = (uint32_t)nbh * 128u;      // 4 arrays x 2 k-halves x nbh rows x 16 B
  uint8_t* s_b = smem;                                       // basis ring
  uint8_t* s_raw = s_b + (size_t)kBS * b_stage_bytes;   // raw audio ring
  float2* s_melw = reinterpret_cast<float2*>(s_raw + (size_t)kConvWarps2 * kDepth * kStageBytes);
  uint32_t* s_adv = reinterpret_cast<uint32_t*>(s_melw + P.nbt);                  // [nbt/16] single / multi close masks
  uint32_t* s_adv2 = s_adv + P.nbt / 16;                                           // [nbt/16] 2-bit close counts
  float* s_win = reinterpret_cast<float*>(s_adv + ((2 * (P.nbt / 16) + 3) & ~3));  // [512]
  float* s_norm = s_win + 512;                                       // [kMaxMels] mel_norms (or 1)
  float* s_park = s_norm + kMaxMels;                                 // [128][n_mels+1] raw mel sums of the tile
  const int park_pitch = P.n_mels + 1;
  float* s_fl = s_park + ((kTileM * park_pitch + 1) & ~1);          // [2][kChains][lo,hi][128] open pairs at chain ends
  Barriers* bars = reinterpret_cast<Barriers*>(s_fl + 2 * kChains * 2 * kTileM);
  unsigned long long* s_stat = reinterpret_cast<unsigned long long*>(bars + 1);   // [2][kMaxMels] fixed-point sum | sum of squares
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(s_stat + 2 * kMaxMels);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // ---- one-time setup
  for (int i = threadIdx.x; i < 2 * kMaxMels; i += kThreads2) s_stat[i] = 0ull;
  for (int i = threadIdx.x; i < P.nbt; i += kThreads2) s_melw[i] = P.melw[i];
  for (int i = threadIdx.x; i < 2 * (P.nbt / 16); i += kThreads2) s_adv[i] = P.meladv[i];
  for (int i = threadIdx.x; i < 512; i += kThreads2) s_win[i] = P.win[i];
  for (int i = threadIdx.x; i < P.n_mels; i += kThreads2) s_norm[i] = P.norms ? P.norms[i] : 1.0f;
  for (int i = threadIdx.x; i < kTileM * park_pitch; i += kThreads2) s_park[i] = 0.f;
  if (threadIdx.x == 0) {
    for (int i = 0; i < kBS; ++i) {
      mbar_init(smem_u32(&bars->b_full[i]), 1); mbar_init(smem_u32(&bars->b_empty[i]), 1); mbar_init(smem_u32(&bars->b_peer[i]), 1);
    }
    for (int i = 0; i < kAStages2; ++i) { mbar_init(smem_u32(&bars->a_full[i]), kPair ? 8 : 4); mbar_init(smem_u32(&bars->a_empty[i]), 1); }
    mbar_init(smem_u32(&bars->acc_full), 1);
    mbar_init(smem_u32(&bars->acc_empty), kPair ? 2 * kEpiWarps : kEpiWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == kWarpTmem) {
    if (kPair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"((uint32_t)kTmemCols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"((uint32_t)kTmemCols) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  // Programmatic dependent launch: everything above (plan constants, barriers, TMEM) overlaps k_prep, which produces
  // the frame counts, group prefix and edge buffers read below.
  asm volatile("griddepcontrol.wait;" ::: "memory");
#if TTTS_PREP_PDL
  asm volatile("griddepcontrol.launch_dependents;");   // the next call's k_prep may be staged now (it waits for this grid's end)
#endif
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();       // pair: the peer's barriers must exist before anyone arrives on them
  tc_fence_after();
  const uint32_t tmem = *s_tmem;
  const int64_t total_groups = P.group_off[P.B];
  const int num_tiles = (int)((total_groups + 3) / 4);
  // a pair walks tiles (2j, 2j+1) in lock step: both CTAs run the same number of iterations, a missing odd tile is empty
  auto tile_live = [&](int tile) { return (kPair ? (tile & ~1) : tile) < num_tiles; };

  const int wgroup = warp >> 2;                        // setmaxnreg works on warpgroups: misc | converters | epilogue
  if (wgroup == 0) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsMisc));
  if (warp == kWarpTma) {
    // ===== basis producer: one bulk copy per k-step stage =====
    if (lane == 0 && !(P.exp_flags & 2)) {
      uint32_t seq = 0;
      for (int tile = blockIdx.x; tile_live(tile); tile += gridDim.x)
        for (int ps = 0; ps < P.passes; ++ps)
          for (int ks = 0; ks < kKS; ++ks, ++seq) {
            const uint32_t st = seq % kBS, use = seq / kBS;
            mbar_wait<1>(smem_u32(&bars->b_empty[st]), (use & 1) ^ 1, P.status, 101);
            const uint32_t full = smem_u32(&bars->b_full[st]);
            mbar_expect_tx(full, b_stage_bytes);
            bulk_g2s(smem_u32(s_b + (size_t)st * b_stage_bytes),
                     P.basis + (((size_t)ps * kKStab + ks) * (kPair ? 2 : 1) + cta_rank) * (b_stage_bytes / 4), b_stage_bytes, full);
          }
    }
    __syncwarp();
  } else if (warp == kWarpMma) {
    // ===== MMA issuer: warp-uniform waits, one elected lane issues =====
    if (kPair && cta_rank != 0) {
      // peer CTA: no MMAs to issue; tell the leader when my half of each basis stage has landed
      if (lane == 0 && !(P.exp_flags & 2)) {
        uint32_t seq = 0;
        for (int tile = blockIdx.x; tile_live(tile); tile += gridDim.x)
          for (int ps = 0; ps < P.passes; ++ps)
            for (int ks = 0; ks < kKS; ++ks, ++seq) {
              const uint32_t st = seq % kBS, use = seq / kBS;
              mbar_wait<1>(smem_u32(&bars->b_full[st]), use & 1, P.status, 204);
              mbar_arrive_cluster(smem_u32(&bars->b_peer[st]), 0);
            }
      }
      __syncwarp();
    } else {
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(nb >> 3) << 17) |
                           ((uint32_t)((kPair ? 2 * kTileM : kTileM) >> 4) << 24);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);          // SBO = 128 B, descriptor version 1
    const uint32_t desc_lo0 = (((uint32_t)nbh * 16u) >> 4) << 16;  // LBO = nbh*16 B between the two k-halves
    const uint32_t arr16 = ((uint32_t)nbh * 32u) >> 4;          // one operand array, in 16 B units
    const uint32_t sb0 = smem_u32(s_b) >> 4, stage16 = b_stage_bytes >> 4;
    const uint32_t d_re = tmem, d_im = tmem + kAccStride2;
    uint32_t kseq = 0, pseq = 0;
    uint32_t as = 0, ause = 0;                                   // A stage = kseq % kAStages2, its use count
    for (int tile = blockIdx.x; tile_live(tile); tile += gridDim.x)
      for (int ps = 0; ps < P.passes; ++ps, ++pseq) {
        if (!(P.exp_flags & 4)) {
          if (kPair) { if (!mbar_try_wait_cluster(smem_u32(&bars->acc_empty), (pseq & 1) ^ 1)) mbar_wait_cluster(smem_u32(&bars->acc_empty), (pseq & 1) ^ 1, P.status, 201); }
          else mbar_wait<1>(smem_u32(&bars->acc_empty), (pseq & 1) ^ 1, P.status, 201);
        }
        for (int ks = 0; ks < kKS; ++ks, ++kseq) {
          const uint32_t st = kseq % kBS, use = kseq / kBS;
          if (!(P.exp_flags & 1)) {
            if (kPair) { if (!mbar_try_wait_cluster(smem_u32(&bars->a_full[as]), ause & 1)) mbar_wait_cluster(smem_u32(&bars->a_full[as]), ause & 1, P.status, 202); }
            else mbar_wait<1>(smem_u32(&bars->a_full[as]), ause & 1, P.status, 202);
          }
          if (!(P.exp_flags & 2)) {
            mbar_wait<1>(smem_u32(&bars->b_full[st]), use & 1, P.status, 203);
            if (kPair && !mbar_try_wait_cluster(smem_u32(&bars->b_peer[st]), use & 1)) mbar_wait_cluster(smem_u32(&bars->b_peer[st]), use & 1, P.status, 205);
          }
          tc_fence_after();
          if (elect_one_sync()) {
            const uint32_t a_base = tmem + kColA2 + as * 32;
            const uint32_t lo = desc_lo0 | (sb0 + st * stage16);
            const uint64_t c_hi = ((uint64_t)desc_hi << 32) | lo;
            const uint64_t c_lo = ((uint64_t)desc_hi << 32) | (lo + arr16);
            const uint64_t s_hi = ((uint64_t)desc_hi << 32) | (lo + 2 * arr16);
            const uint64_t s_lo = ((uint64_t)desc_hi << 32) | (lo + 3 * arr16);
            const uint32_t e_hi = a_base, e_lo = a_base + 8, o_hi = a_base + 16, o_lo = a_base + 24;
            const uint32_t first = ks == 0 ? 0u : 1u;
            auto mma = [&](uint32_t d, uint32_t a, uint64_t b, uint32_t acc) {
              if (kPair) mma_tf32_ts2(d, a, b, idesc, acc); else mma_tf32_ts(d, a, b, idesc, acc);
            };
            auto commit = [&](uint32_t bar) { if (kPair) tc_commit2(bar); else tc_commit(bar); };
            if (!(P.exp_flags & 32)) {
              mma(d_re, e_lo, c_hi, first);                  // small terms first
              mma(d_re, e_hi, c_lo, 1u);
              mma(d_re, e_hi, c_hi, 1u);
              if (kCentreEpi || ks < kKS - 1) {              // the centre-tap k-step has no sine part
                mma(d_im, o_lo, s_hi, first);
                mma(d_im, o_hi, s_lo, 1u);
                mma(d_im, o_hi, s_hi, 1u);
              }
            }
            if (!(P.exp_flags & 64)) {                       // (64: timing experiment, only valid together with 1 | 2)
            commit(smem_u32(&bars->b_empty[st]));            // basis stage free once these MMAs retire
            commit(smem_u32(&bars->a_empty[as]));            // A stage free
            }
            if (ks == kKS - 1) commit(smem_u32(&bars->acc_full));
          }
          __syncwarp();
          if (++as == kAStages2) { as = 0; ++ause; }
        }
      }
    }
  }
  } else if (wgroup <= kSets) {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsConvT));
  if (!(P.exp_flags & 1)) {
    // ===== converters: stage own slots (cp.async) -> window + double fold + TF32 split -> TMEM A stage `set` =====
    const float *src_p = P.src, *edge_p = P.edge;
    const int cw = warp - kWarpConv02;
    const int set = cw >> 2;                           // A stage and k-step residue this warp serves
    const int q = warp & 3;                            // TMEM lane quarter == group within the tile
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    uint8_t* my_raw = s_raw + (size_t)cw * kDepth * kStageBytes;

    // piece table: piece idx = 32 i + lane of the stage = (slot sigma, 16 B piece 0..4 of the slot); pieces 0..2 are
    // the forward window of frame sigma, pieces 3..4 the mirrored window of frame sigma - 2
    uint32_t pc_dst[kPieceIters];                      // byte offset in the stage | piece << 16 | live << 31
#pragma unroll
    for (int i = 0; i < kPieceIters; ++i) {
      const int idx = i * 32 + lane, sigma = idx / 5, piece = idx - sigma * 5;
      pc_dst[i] = (uint32_t)(sigma * kSlotFloats * 4 + piece * 16) | ((uint32_t)piece << 16) | (idx < kPieces ? 0x80000000u : 0u);
    }

    struct Cursor { int tile, ps, ks; };
    auto advance = [&](Cursor& c) -> bool {           // to this set's next k-step; true when it enters a new tile
      c.ks += kSets;
      if (c.ks >= kKS) {
        c.ks -= kKS;
        if (++c.ps >= P.passes) { c.ps = 0; c.tile += gridDim.x; return true; }
      }
      return false;
    };

    int32_t off[kPieceIters];
    uint32_t okm = 0, edm = 0;
    bool plain = false;                                // every live piece of the warp is valid and read from P.src
    float xc_next = 0.f;                               // centre tap f[512] of this lane's row in the issue tile
    auto load_tile = [&](int tile) {
      okm = edm = 0;
      xc_next = 0.f;
      plain = false;
      if (tile >= num_tiles) return;
      const GroupInfo gi = locate_group(P, (int64_t)tile * 4 + q, total_groups);
      if (!gi.valid) return;
#pragma unroll
      for (int i = 0; i < kPieceIters; ++i) {
        const int piece = (pc_dst[i] >> 16) & 7, sigma = (int)(pc_dst[i] & 0xffffu) / (kSlotFloats * 4);
        const bool fwd = piece < 3;
        const int tv = gi.t0 + (fwd ? sigma : sigma - 2);
        const int woff = fwd ? 4 * piece : (kNfft - 8) + 4 * (piece - 3);
        const bool ok = (pc_dst[i] >> 31) && (fwd ? tv <= gi.T + 1 : tv <= gi.T - 1);
        bool edge;
        off[i] = sample_offset(P, gi.b, gi.T, tv, woff, &edge);
        okm |= (uint32_t)ok << i;
        edm |= (uint32_t)(ok && edge) << i;
      }
      uint32_t live = 0;
#pragma unroll
      for (int i = 0; i < kPieceIters; ++i) live |= (pc_dst[i] >> 31) << i;
      plain = __all_sync(0xffffffffu, okm == live && edm == 0);
      const int t = gi.t0 + lane;
      if (!kCentreEpi && t < gi.T) {
        bool edge;
        const int32_t o = sample_offset(P, gi.b, gi.T, t, kNfft / 2, &edge);
        xc_next = __ldg((edge ? edge_p : src_p) + o);
      }
    };

    Cursor ic{(int)blockIdx.x, 0, set};                // issue cursor, kDepth own k-steps ahead of the consume cursor
    auto issue = [&](uint32_t stage) {
      if (ic.tile < num_tiles && (kCentreEpi || ic.ks < kKS - 1) && !(P.exp_flags & 128)) {   // 128: timing experiment, no raw-audio loads
        const uint32_t base = smem_u32(my_raw + (size_t)stage * kStageBytes);
        const int adv = kKStep * ic.ks;
        if (plain) {                                   // warp-uniform: interior groups (all but the first / last of a clip)
#pragma unroll
          for (int i = 0; i < kPieceIters; ++i) {
            if (i < kPieces / 32 || (pc_dst[i] >> 31)) {
              const bool fwd = ((pc_dst[i] >> 16) & 7) < 3;
              cp_async16(base + (pc_dst[i] & 0xffffu), src_p + (off[i] + (fwd ? adv : -adv)), 16u);
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < kPieceIters; ++i) {
            if (pc_dst[i] >> 31) {
              const bool ok = (okm >> i) & 1u;
              const bool fwd = ((pc_dst[i] >> 16) & 7) < 3;
              const float* bp = ((edm >> i) & 1u) ? edge_p : src_p;
              cp_async16(base + (pc_dst[i] & 0xffffu), ok ? bp + (off[i] + (fwd ? adv : -adv)) : src_p, ok ? 16u : 0u);
            }
          }
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");   // empty groups keep the wait_group count uniform
      if (tile_live(ic.tile) && advance(ic)) load_tile(ic.tile);
    };

    load_tile(ic.tile);
    float xc = xc_next;
#pragma unroll
    for (int d = 0; d < kDepth; ++d) issue(d);

    Cursor cc{(int)blockIdx.x, 0, set};
    uint32_t mine = 0, kseq = set;                     // kseq: sequence number of my current k-step (A stage = kseq % 4)
    while (tile_live(cc.tile)) {
      const uint32_t stage = mine % kDepth;
      asm volatile("cp.async.wait_group %0;" ::"n"(kDepth - 1) : "memory");
      __syncwarp();
      uint32_t ehi[8], elo[8], ohi[8], olo[8];
      const bool centre = !kCentreEpi && cc.ks == kKS - 1;
      if (!centre) {
        const float4* s0 = reinterpret_cast<const float4*>(my_raw + (size_t)stage * kStageBytes) + lane * 5;
        const float4* s2 = s0 + 10;                    // slot lane + 2
        float a[12], c[8], d[12], b[8];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const float4 v = s0[i], u = s2[i];
          a[4 * i] = v.x; a[4 * i + 1] = v.y; a[4 * i + 2] = v.z; a[4 * i + 3] = v.w;
          d[4 * i] = u.x; d[4 * i + 1] = u.y; d[4 * i + 2] = u.z; d[4 * i + 3] = u.w;
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          const float4 v = s0[3 + i], u = s2[3 + i];
          c[4 * i] = v.x; c[4 * i + 1] = v.y; c[4 * i + 2] = v.z; c[4 * i + 3] = v.w;
          b[4 * i] = u.x; b[4 * i + 1] = u.y; b[4 * i + 2] = u.z; b[4 * i + 3] = u.w;
        }
        float w1[8], w2[8];
        {
          const float4* wp = reinterpret_cast<const float4*>(s_win + kKStep * cc.ks);
          const float4 p0 = wp[0], p1 = wp[1], r0 = wp[64], r1 = wp[65];
          w1[0] = p0.x; w1[1] = p0.y; w1[2] = p0.z; w1[3] = p0.w; w1[4] = p1.x; w1[5] = p1.y; w1[6] = p1.z; w1[7] = p1.w;
          w2[0] = r0.x; w2[1] = r0.y; w2[2] = r0.z; w2[3] = r0.w; w2[4] = r1.x; w2[5] = r1.y; w2[6] = r1.z; w2[7] = r1.w;
        }
        const float sgn = (cc.ps >= P.sp) ? -1.f : 1.f;   // passes [0, sp) are the even bins, [sp, 2 sp) the odd bins
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float fa = a[1 + j], fb = b[7 - j], fc = c[7 - j], fd = d[1 + j];   // f[n], f[N-n], f[512-n], f[512+n]
          const float w2s = sgn * w2[j];
          const float e = fmaf(w2s, fc + fd, w1[j] * (fa + fb));
          const float o = fmaf(-w2s, fc - fd, w1[j] * (fa - fb));
          const uint32_t he = __float_as_uint(e) & 0xFFFFE000u, ho = __float_as_uint(o) & 0xFFFFE000u;
          ehi[j] = he; elo[j] = __float_as_uint(e - __uint_as_float(he));
          ohi[j] = ho; olo[j] = __float_as_uint(o - __uint_as_float(ho));
        }
      } else {
        const float e = P.wc * xc;
        const uint32_t he = __float_as_uint(e) & 0xFFFFE000u;
#pragma unroll
        for (int j = 0; j < 8; ++j) { ehi[j] = 0u; elo[j] = 0u; ohi[j] = 0u; olo[j] = 0u; }
        ehi[0] = he; elo[0] = __float_as_uint(e - __uint_as_float(he));
      }
      __syncwarp();                                    // every lane has its taps in registers: the stage can be refilled
      issue(stage);
      const uint32_t as = kseq % kAStages2, ause = kseq / kAStages2;
      const uint32_t a_base = tmem + lane_addr + kColA2 + as * 32;
      mbar_wait<2>(smem_u32(&bars->a_empty[as]), (ause & 1) ^ 1, P.status, 402);
      tc_fence_after();
      tmem_st8(a_base + 0, ehi);
      tmem_st8(a_base + 8, elo);
      if (!centre) {
        tmem_st8(a_base + 16, ohi);
        tmem_st8(a_base + 24, olo);
      }
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (kPair) mbar_arrive_cluster(smem_u32(&bars->a_full[as]), 0); else mbar_arrive(smem_u32(&bars->a_full[as])); }
      ++mine;
      kseq += kSets;
      if (advance(cc)) xc = xc_next;
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  }
  } else {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsEpiT));
  if (!(P.exp_flags & 4)) {
    // ===== epilogue: power -> sparse mel -> log(clamp) -> / norm.  4 warps per TMEM lane quarter, one scan chain each =====
    const int q = warp & 3;
    const int ch = (warp - kWarpEpi02) >> 2;           // my chain: bins [ch*cb, (ch+1)*cb) of every pass
    const int r = q * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const int cb = nb / kChains;                       // bins per chain (multiple of 16)
    const uint32_t t_re = tmem + lane_addr + ch * cb, t_im = t_re + kAccStride2;
    const uint32_t pa0 = smem_u32(s_park + r * park_pitch);
    const uint32_t bar_done = 1 + q, bar_flushed = 5 + q, bar_clean = 9 + q;   // named barriers of this quarter's four warps
    uint32_t pseq = 0;
    for (int tile = blockIdx.x; tile_live(tile); tile += gridDim.x) {
      // Sparse mel contraction.  Every bin feeds the running pair of adjacent filters (cur, cur+1) of its chain; the host
      // table says how many filters close before each bin (bit j: one, predicated inline; bit 16+j: two or three, rare,
      // out of line).  A closed filter's partial sum is added to its parked sum (`old`, prefetched one close ahead; the
      // park rows start a tile zeroed).  The filters still open at the end of a chain are shared with the next chain, so
      // they go through s_fl and chain 0's warp adds them once all four chains of the quarter are done.
      float ec = 0.f;                                    // w[512] f[512] of my row: the centre tap, added to Re with the parity's sign
      if (kCentreEpi) {
        const GroupInfo gc = locate_group(P, (int64_t)tile * 4 + q, total_groups);
        const int tc = gc.t0 + lane;
        if (gc.valid && tc < gc.T) {
          bool edge;
          const int32_t o = sample_offset(P, gc.b, gc.T, tc, kNfft / 2, &edge);
          ec = P.wc * __ldg((edge ? P.edge : P.src) + o);
        }
      }
      for (int ps = 0; ps < P.passes; ++ps, ++pseq) {
        const float ecs = (ps >= P.sp) ? -ec : ec;       // passes [0, sp): even bins, cos(pi k) = +1; [sp, 2 sp): odd bins, -1
        // power plans: the 33rd k-step's two accumulations, D += tf32(e_lo) * (+-1), D += e_hi * (+-1)
        const float ecs_hi = __uint_as_float(__float_as_uint(ecs) & 0xFFFFE000u);
        const float ecs_lo = __uint_as_float(__float_as_uint(ecs - ecs_hi) & 0xFFFFE000u);
        ScanState sc;
        sc.run_lo = 0.f; sc.run_hi = 0.f;
        sc.pa = pa0 + 4u * (uint32_t)P.cstart[ps * kChains + ch];
        sc.old = ld_s(sc.pa);
        const int bin_base = ps * nb + ch * cb;
        mbar_wait<3>(smem_u32(&bars->acc_full), pseq & 1, P.status, 501);
        tc_fence_after();
        // scan of four consecutive bins whose power is in registers.  Power plans (V1): the block's 4-bit close pattern
        // selects one of 16 straight-line bodies (-2 % on the kernel, short runs and at the power cap alike:
        // profiles/r2v, r2w); magnitude plans (V2) scan from registers at eight unrolled call sites, where the 16 bodies
        // would have to go out of line (+1-2 %): they keep the predicated per-bin form.
        auto scan4 = [&](const float (&pw4)[4], uint32_t advq, int bin0) {
          if (kPower1) {
            const float4 wa = *reinterpret_cast<const float4*>(s_melw + bin0);
            const float4 wb = *reinterpret_cast<const float4*>(s_melw + bin0 + 2);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (advq & (0x10000u << j)) {            // warp-uniform, rare: two or three filters close
                const uint32_t cnt = (s_adv2[bin0 >> 4] >> (2 * ((bin0 & 15) + j))) & 3u;
                park_multi(sc.pa, sc.run_lo, sc.run_hi);
                sc.run_lo = 0.f; sc.run_hi = 0.f;
                sc.pa += 4 * cnt;
                sc.old = ld_s(sc.pa);
              }
              if (advq & (1u << j)) {                  // warp-uniform: one filter closes before this bin
                st_s(sc.pa, sc.old + sc.run_lo);
                sc.run_lo = sc.run_hi; sc.run_hi = 0.f;
                sc.pa += 4;
                sc.old = ld_s(sc.pa);                     // consumed at the next close
              }
              const float wl = j == 0 ? wa.x : j == 1 ? wa.z : j == 2 ? wb.x : wb.z;
              const float wh = j == 0 ? wa.y : j == 1 ? wa.w : j == 2 ? wb.y : wb.w;
              sc.run_lo = fmaf(wl, pw4[j], sc.run_lo);
              sc.run_hi = fmaf(wh, pw4[j], sc.run_hi);
            }
          } else if (advq & 0xf0000u) {                     // warp-uniform: some bin of the block has two or three filters
                                                     // closing before it (the narrow low-frequency filters: all of
                                                     // chain 0's first blocks) -- one four-way switch per bin
            const float4 wa = *reinterpret_cast<const float4*>(s_melw + bin0);
            const float4 wb = *reinterpret_cast<const float4*>(s_melw + bin0 + 2);
            const uint32_t cnt8 = (s_adv2[bin0 >> 4] >> (2 * (bin0 & 15))) & 0xffu;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              switch ((cnt8 >> (2 * j)) & 3u) {
                case 0: break;
                case 1:
                  st_s(sc.pa, sc.old + sc.run_lo);
                  sc.run_lo = sc.run_hi; sc.run_hi = 0.f;
                  sc.pa += 4;
                  sc.old = ld_s(sc.pa);
                  break;
                case 2: {
                  const float o2 = ld_s(sc.pa + 4);
                  st_s(sc.pa, sc.old + sc.run_lo);
                  st_s(sc.pa + 4, o2 + sc.run_hi);
                  sc.run_lo = 0.f; sc.run_hi = 0.f;
                  sc.pa += 8;
                  sc.old = ld_s(sc.pa);
                  break;
                }
                default: {
                  const float o2 = ld_s(sc.pa + 4);
                  st_s(sc.pa, sc.old + sc.run_lo);
                  st_s(sc.pa + 4, o2 + sc.run_hi);
                  sc.run_lo = 0.f; sc.run_hi = 0.f;
                  sc.pa += 12;
                  sc.old = ld_s(sc.pa);
                  break;
                }
              }
              const float wl = j == 0 ? wa.x : j == 1 ? wa.z : j == 2 ? wb.x : wb.z;
              const float wh = j == 0 ? wa.y : j == 1 ? wa.w : j == 2 ? wb.y : wb.w;
              sc.run_lo = fmaf(wl, pw4[j], sc.run_lo);
              sc.run_hi = fmaf(wh, pw4[j], sc.run_hi);
            }
          } else {
            const float4 wa = *reinterpret_cast<const float4*>(s_melw + bin0);
            const float4 wb = *reinterpret_cast<const float4*>(s_melw + bin0 + 2);
            scan4_switch(sc, pw4, wa, wb, advq & 0xfu);
          }
        };
        auto power_of = [&](uint32_t re, uint32_t im) {
          const float r0 = __uint_as_float(re), bq = __uint_as_float(im);
          const float a = !kCentreEpi ? r0 : kPower1 ? r0 + ecs : __fadd_rn(__fadd_rn(r0, ecs_lo), ecs_hi);
          float pw = fmaf(a, a, bq * bq);
          if (kPower1) {
#if F2_SQRT_APPROX
            // |X| = sqrt(re^2 + im^2 + eps): the approximate square root (MUFU, <= 1 ulp class error, denormals flushed:
            // powers below 1e-38 are far under the clamp) instead of the IEEE sequence, 10 instructions per bin shorter
            pw += P.power_eps;
            asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(pw) : "f"(pw));
#else
            pw = sqrtf(pw + P.power_eps);
#endif
          }
          return pw;
        };
        auto release_acc = [&]() {                     // my accumulator columns are in registers: hand them back
          tc_fence_before();
          __syncwarp();
          if (lane == 0) { if (kPair) mbar_arrive_cluster(smem_u32(&bars->acc_empty), 0); else mbar_arrive(smem_u32(&bars->acc_empty)); }
        };
        // The last n_hold bins of the chain are pulled into registers as power values in one go, so that the accumulators
        // go back to the MMA warp BEFORE their scan runs (the scan is issue-bound while all 16 epilogue warps burst); the
        // first n_stream bins are scanned straight out of TMEM as before.
        const int n_hold = kDebug ? 0 : (kHold < cb ? kHold : cb);
        const int n_stream = cb - n_hold;
        for (int r16 = 0; r16 < n_stream; r16 += 16) {
          const uint32_t advw = s_adv[(bin_base + r16) >> 4];
#pragma unroll(kQrUnroll)
          for (int qr = 0; qr < 4; ++qr) {             // barely unrolled: the scan stays resident in the instruction cache
            uint32_t re[4], im[4];
            tmem_ld4(t_re + r16 + 4 * qr, re);
            tmem_ld4(t_im + r16 + 4 * qr, im);
            tmem_wait_ld();
            if (n_hold == 0 && r16 + 16 >= cb && qr == 3) release_acc();
            if (kDebug) {
              const GroupInfo gi = locate_group(P, (int64_t)tile * 4 + q, total_groups);
              const int t = gi.t0 + lane;
              if (gi.valid && t < gi.T && t < P.Tmax) {
                const int64_t gdbg = P.frame_off[gi.b] + t;
                const int par = ps / P.sp, sub = ps - par * P.sp;
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const int k = P.first_k[par] + 2 * (sub * nb + ch * cb + r16 + 4 * qr + j);
                  if (k <= P.kmax) {
                    float* d = P.dbg_reim + gdbg * (2 * P.dbg_nbt) + (k - P.dbg_kmin);
                    d[0] = !kCentreEpi ? __uint_as_float(re[j]) : kPower1 ? __uint_as_float(re[j]) + ecs
                                                                            : __fadd_rn(__fadd_rn(__uint_as_float(re[j]), ecs_lo), ecs_hi);
                    d[P.dbg_nbt] = __uint_as_float(im[j]);
                  }
                }
              }
            }
            if (P.exp_flags & 8) continue;
            float pw4[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) pw4[j] = power_of(re[j], im[j]);
            scan4(pw4, advw >> (4 * qr), bin_base + r16 + 4 * qr);
          }
        }
        if (kHold > 0 && n_hold > 0) {
          float pw[kHold > 0 ? kHold : 4];
#pragma unroll
          for (int k8 = 0; k8 < kHold / 8; ++k8) {
            if (8 * k8 < n_hold) {
              uint32_t re[8], im[8];
              tmem_ld8(t_re + n_stream + 8 * k8, re);
              tmem_ld8(t_im + n_stream + 8 * k8, im);
              tmem_wait_ld();
#pragma unroll
              for (int j = 0; j < 8; ++j) pw[8 * k8 + j] = power_of(re[j], im[j]);
            }
          }
          release_acc();
          if (!(P.exp_flags & 8)) {
#pragma unroll
            for (int k16 = 0; k16 < kHold / 16; ++k16) {
              if (16 * k16 < n_hold) {
                const uint32_t advw = s_adv[(bin_base + n_stream + 16 * k16) >> 4];
#pragma unroll
                for (int qr = 0; qr < 4; ++qr) {
                  const float pw4[4] = {pw[16 * k16 + 4 * qr], pw[16 * k16 + 4 * qr + 1], pw[16 * k16 + 4 * qr + 2],
                                        pw[16 * k16 + 4 * qr + 3]};
                  scan4(pw4, advw >> (4 * qr), bin_base + n_stream + 16 * k16 + 4 * qr);
                }
              }
            }
          }
        }
        // the pair still open at the end of my chain
        float* fl = s_fl + ((pseq & 1) * kChains + ch) * 2 * kTileM + r;
        fl[0] = sc.run_lo;
        fl[kTileM] = sc.run_hi;
        if (ch != 0) {
          asm volatile("bar.arrive %0, %1;" ::"r"(bar_done), "n"(32 * kChains) : "memory");
        } else {
          asm volatile("bar.sync %0, %1;" ::"r"(bar_done), "n"(32 * kChains) : "memory");
#pragma unroll
          for (int c = 0; c < kChains; ++c) {
            const float* f = s_fl + ((pseq & 1) * kChains + c) * 2 * kTileM + r;
            park_multi(pa0 + 4u * (uint32_t)P.cend[ps * kChains + c], f[0], f[kTileM]);
          }
        }
        // nobody starts the next scan (or the log pass) before the open pairs are in the park rows
        asm volatile("bar.sync %0, %1;" ::"r"(bar_flushed), "n"(32 * kChains) : "memory");
        if (ps == P.passes - 1) {
          // raw mel sums -> log(max(., clamp)) / norm.  Each of the quarter's four warps takes 8 of its 32 rows: coalesced
          // output rows, 2 rows x up to 4 columns per lane in flight (independent logf / division chains); the park rows
          // are handed back zeroed for the next tile
          const GroupInfo gi = locate_group(P, (int64_t)tile * 4 + q, total_groups);
          const int t = gi.t0 + lane;
          float* orow = (gi.valid && t < gi.T && t < P.Tmax) ? P.out + ((int64_t)gi.b * P.Tmax + t) * P.n_mels : nullptr;
#if F2_FAST_LOG
          const float log_clamp = logf(P.clamp);
#endif
          // the quarter's 32 rows are split over its kChains warps (8 each for four chains; 10 / 11 / 11 for three)
          const int row_lo = (32 * ch) / kChains, row_hi = (32 * (ch + 1)) / kChains;
          for (int rr0 = row_lo; rr0 < ((P.exp_flags & 16) ? 0 : row_hi); rr0 += 2) {
            float* rowp[2];
            float v[2][4];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
              const bool mine = (32 % kChains == 0) || rr0 + i < row_hi;      // an odd share ends on a single row
              rowp[i] = reinterpret_cast<float*>((uintptr_t)__shfl_sync(0xffffffffu, (unsigned long long)(uintptr_t)orow, (rr0 + i) & 31));
              if (!mine) rowp[i] = nullptr;
              float* parked = s_park + (q * 32 + rr0 + i) * park_pitch + lane;
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                v[i][k] = 1.f;
                if (mine && lane + 32 * k < P.n_mels) { v[i][k] = parked[32 * k]; parked[32 * k] = 0.f; }
              }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (32 * k < P.n_mels) {
                const float nrm = (lane + 32 * k < P.n_mels) ? s_norm[lane + 32 * k] : 1.f, rn = 1.0f / nrm;
                long long st_s = 0, st_q = 0;
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                  // x / norm as reciprocal + one Newton correction: branch-free (the IEEE division's slow-path call
                  // would serialise the chains) and equal to the correctly rounded quotient except in rare last-bit ties
#if F2_FAST_LOG
                  // lg2.approx * ln 2 (<= 3 ulp of the result away from 1, 2^-21.4 absolute near it) instead of logf's ~20
                  // instructions; values at or below the clamp take the exactly rounded constant, so silence stays bit-exact
                  const float x = !(v[i][k] <= P.clamp) ? __logf(v[i][k]) : log_clamp;   // (NaN goes through the log, as torch.clamp keeps it)
#else
                  const float x = logf(v[i][k] < P.clamp ? P.clamp : v[i][k]);
#endif
                  if (P.stat_ws && rowp[i]) {
                    // stage (4), fused: per-channel sum / sum of squares of the UN-normalised log-mel in 64-bit fixed point
                    // (steps of 2^-32 for x, 2^-30 for x^2; |x| < 2^5 leaves room for 2^24 frames per launch) -- integer addition is
                    // associative, so the statistics are bit-identical from run to run and for any sharding of the corpus
                    const double xd = (double)x;
                    st_s += __double2ll_rn(xd * 4294967296.0);
                    st_q += __double2ll_rn(xd * xd * 1073741824.0);
                  }
                  const float y = x * rn;
                  v[i][k] = fabsf(y) < 3.0e38f ? fmaf(fmaf(-y, nrm, x), rn, y) : y;   // norm 0 / inf: keep x * (1/norm)
                }
                if (P.stat_ws && lane + 32 * k < P.n_mels) {
                  atomicAdd(s_stat + lane + 32 * k, (unsigned long long)st_s);
                  atomicAdd(s_stat + kMaxMels + lane + 32 * k, (unsigned long long)st_q);
                }
              }
            }
#pragma unroll
            for (int i = 0; i < 2; ++i)
#pragma unroll
              for (int k = 0; k < 4; ++k)
                if (rowp[i] && lane + 32 * k < P.n_mels) rowp[i][lane + 32 * k] = v[i][k];
          }
          asm volatile("bar.sync %0, %1;" ::"r"(bar_clean), "n"(32 * kChains) : "memory");   // all 32 park rows of the quarter are zero again
        }
      }
    }
  }

  }

  // ---- teardown
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();       // pair: the leader's MMAs read the peer's TMEM and smem
  if (P.stat_ws) {
    // this CTA's fixed-point sums -> the launch's (integer atomics: order-independent); the last CTA to arrive converts
    // the launch totals to fp64 and adds them, with the frame count, to the caller's accumulator
    __shared__ int s_last;
    for (int i = threadIdx.x; i < 2 * kMaxMels; i += kThreads2)
      if (s_stat[i]) atomicAdd(P.stat_ws + i, s_stat[i]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(P.stat_ws + 2 * kMaxMels, 1ull) == (unsigned long long)gridDim.x - 1;
    __syncthreads();
    if (s_last) {
      __threadfence();
      for (int i = threadIdx.x; i < 2 * P.n_mels; i += kThreads2) {
        const bool sq = i >= P.n_mels;
        const long long tot = (long long)atomicExch(P.stat_ws + (sq ? kMaxMels + i - P.n_mels : i), 0ull);
        P.stat_acc[i] += (double)tot * (sq ? 1.0 / 1073741824.0 : 1.0 / 4294967296.0);
      }
      if (threadIdx.x == 0) {
        P.stat_acc[2 * P.n_mels] += (double)P.frame_off[P.B];
        P.stat_ws[2 * kMaxMels] = 0ull;
      }
    }
  }
  if (warp == kWarpTmem) {
    tc_fence_after();
    if (kPair) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)kTmemCols) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"((uint32_t)kTmemCols) : "memory");
  }
}
