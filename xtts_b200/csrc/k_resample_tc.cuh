// k_resample_tc.cuh -- the general-ratio polyphase FIR on the tensor cores (included by k_ingest.cuh, namespace ingest).
//
//   out[(i, p)] = sum_k X[i, k] * T[k, p],   X[i, k] = xpad[i * orig + k]  (rows overlap by kw - orig samples)
//
// is a [strides x taps] x [taps x phases] product: M = 128 strides per tile, K = taps (8 per k-step), N = a block of
// <= 192 phases.  Same machinery as k_logmel (ttts_mel.cu): tcgen05.mma kind::tf32 with the A operand in TMEM, written by
// converter warps (tcgen05.st), the B operand (pre-tiled taps) in shared memory by cp.async.bulk, accumulators in TMEM
// (two buffers: the epilogue of unit u overlaps the MMAs of unit u + 1), and the 3xTF32 split (hi = 13 low mantissa bits
// cleared, lo = x - hi; products lo*hi, hi*lo, hi*hi) that keeps fp32-grade accuracy.
//
//   warp 0      B producer: one bulk copy per quad of k-steps (taps hi | lo, [2 k-halves][nbp][4] each, K-major
//               SWIZZLE_NONE)
//   warp 1      MMA issuer (one elected lane), 12 MMAs per quad (a stage = four k-steps: the barrier and commit
//               traffic of a single k-step costs more than its three short MMAs)
//   warp 2      TMEM allocator
//   warps 4-15  converters: warp (set, q) owns rows 32 q .. 32 q + 31 for every third pair of k-steps: samples straight
//               from the waveform (fp32, or int16 PCM scaled while loading; coalesced, transposed through shared
//               memory), split, 2 x tcgen05.st.x8 per k-step
//   warps 16-23 epilogue (two per TMEM lane quarter, alternating 32-column blocks): tcgen05.ld 32 columns -> per-warp shared-memory transpose -> one 128-byte row segment per
//               store instruction (clip, zeros past the clip's end)
// Work unit = (clip, 128-stride tile, phase block), dealt round-robin to one persistent CTA per SM.
#pragma once

namespace tc {

constexpr int kTileM = 128;
#ifndef RTC_CONV_SETS
#define RTC_CONV_SETS 3
#endif
#ifndef RTC_EPI_SETS
#define RTC_EPI_SETS 2
#endif
constexpr int kConvSets = RTC_CONV_SETS, kEpiSets = RTC_EPI_SETS;
constexpr int kThreadsTc = 32 * (4 + 4 * kConvSets + 4 * kEpiSets);   // 4 misc + 12 converter + 8 epilogue warps
constexpr int kBStagesMax = 4;
constexpr int kAStagesTc = 3;          // x 64 TMEM columns: one quad of k-steps, (hi 8 | lo 8) each
constexpr int kMaxNb = 160;            // phases per block; two accumulator buffers = 320 columns
constexpr int kColATc = 2 * kMaxNb;    // A ring: columns 320..511
constexpr int kWarpConvTc = 4, kWarpEpiTc = 4 + 4 * kConvSets;
constexpr int kTPitch = 33;            // epilogue transpose buffer pitch (floats)

struct Params {
  const void* x;              // float32 [B][row_stride] or int16 [B][row_stride] (frame-major, channel-minor)
  long long row_stride;
  int channels;
  const long long* lengths;   // [B] or null
  long long L;
  int B;
  int orig, new_f, width, kw;
  int ksteps;                 // pairs of k-steps (16 taps each), rounded up to even: a stage is a quad (32 taps)
  int npass, nbp;             // phase blocks, phases per block (multiple of 16)
  const float* tapsB;         // [npass][quads][4 k-steps][hi | lo][2 k-halves][nbp][4]
  float* out;
  long long out_stride, out_cols;
  int tiles_per_clip;
  long long units;            // B * tiles_per_clip * npass
  int clip;
  int b_stages;               // taps ring depth
  int exp;                    // timing experiments (RESULTS INVALID): 1 no converters, 2 no taps loads, 4 no epilogue, 32 no MMAs, 128 converters without global loads, 256 without TMEM stores
  int32_t* status;
};

struct __align__(8) Barriers {
  uint64_t b_full[kBStagesMax], b_empty[kBStagesMax];
  uint64_t a_full[kAStagesTc], a_empty[kAStagesTc];
  uint64_t acc_full[2], acc_empty[2];
};

struct Unit {
  int b, pass;
  long long row0;             // first stride of the tile
  long long len, olen;        // this clip's input length and output length
  bool live;                  // false: the tile lies past the clip's end (zero fill only)
};

__device__ __forceinline__ Unit locate(const Params& P, long long u) {
  Unit r;
  const long long t = u / P.npass;
  r.pass = (int)(u - t * P.npass);
  r.b = (int)(t / P.tiles_per_clip);
  r.row0 = (t - (long long)r.b * P.tiles_per_clip) * kTileM;
  r.len = P.lengths ? min(max(P.lengths[r.b], 0LL), P.L) : P.L;
  r.olen = (r.len * P.new_f + P.orig - 1) / P.orig;
  r.live = r.row0 * P.new_f < r.olen;
  return r;
}

template <typename TIn>
__global__ void __launch_bounds__(kThreadsTc, 1) k_resample_tc(const Params P) {
  extern __shared__ __align__(1024) uint8_t smem_tc[];
  const uint32_t b_stage_bytes = (uint32_t)P.nbp * 256u;     // 4 k-steps x 2 arrays x 2 k-halves x nbp rows x 16 B
  const int nquads = P.ksteps >> 1;
  uint8_t* s_b = smem_tc;
  const uint32_t nbs = (uint32_t)P.b_stages;
  float* s_t = reinterpret_cast<float*>(s_b + (size_t)nbs * b_stage_bytes);     // [8 epilogue warps][32][kTPitch]
  float* s_c = s_t + 4 * kEpiSets * 32 * kTPitch;                                // [12 converter warps][32][20]
  Barriers* bars = reinterpret_cast<Barriers*>(s_c + 4 * kConvSets * 32 * 20);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kBStagesMax; ++i) { mbar_init(smem_u32(&bars->b_full[i]), 1); mbar_init(smem_u32(&bars->b_empty[i]), 1); }
    for (int i = 0; i < kAStagesTc; ++i) { mbar_init(smem_u32(&bars->a_full[i]), 8); mbar_init(smem_u32(&bars->a_empty[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&bars->acc_full[i]), 1); mbar_init(smem_u32(&bars->acc_empty[i]), 4 * kEpiSets); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(s_tmem)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *s_tmem;

  if (warp == 0) {
    // ===== taps producer =====
    if (lane == 0 && !(P.exp & 2)) {
      uint32_t seq = 0;
      for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
        const Unit un = locate(P, u);
        if (!un.live) continue;
        for (int ks = 0; ks < nquads; ++ks, ++seq) {
          const uint32_t st = seq % nbs, use = seq / nbs;
          mbar_wait(smem_u32(&bars->a_empty[st]), (use & 1) ^ 1, P.status, 101);   // one "stage free" barrier for A and B
          const uint32_t full = smem_u32(&bars->b_full[st]);
          mbar_expect_tx(full, b_stage_bytes);
          bulk_g2s(smem_u32(s_b + (size_t)st * b_stage_bytes),
                   P.tapsB + ((size_t)un.pass * nquads + ks) * (b_stage_bytes / 4), b_stage_bytes, full);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(P.nbp >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);             // SBO = 128 B, descriptor version 1
    const uint32_t desc_lo0 = (((uint32_t)P.nbp * 16u) >> 4) << 16;   // LBO = distance between the two k-halves
    const uint32_t arr16 = ((uint32_t)P.nbp * 32u) >> 4;             // one operand array, in 16 B units
    const uint32_t sb0 = smem_u32(s_b) >> 4, stage16 = b_stage_bytes >> 4;
    uint32_t kseq = 0, useq = 0;
    for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
      const Unit un = locate(P, u);
      if (!un.live) continue;
      const uint32_t abuf = useq & 1, ause_acc = useq >> 1;
      const uint32_t d = tmem + abuf * kMaxNb;
      if (!(P.exp & 4)) mbar_wait(smem_u32(&bars->acc_empty[abuf]), (ause_acc & 1) ^ 1, P.status, 201);
      for (int ks = 0; ks < nquads; ++ks, ++kseq) {
        const uint32_t as = kseq % kAStagesTc, ause = kseq / kAStagesTc;
        const uint32_t st = kseq % nbs, use = kseq / nbs;
        if (!(P.exp & 1)) mbar_wait(smem_u32(&bars->a_full[as]), ause & 1, P.status, 202);
        if (!(P.exp & 2)) mbar_wait(smem_u32(&bars->b_full[st]), use & 1, P.status, 203);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t a0 = tmem + kColATc + as * 64;
          const uint32_t lo = desc_lo0 | (sb0 + st * stage16);
          if (!(P.exp & 32)) {
#pragma unroll
            for (uint32_t h = 0; h < 4; ++h) {               // the four k-steps of the quad; small terms first
              const uint32_t a_hi = a0 + 16 * h, a_lo = a_hi + 8;
              const uint64_t b_hi = ((uint64_t)desc_hi << 32) | (lo + 2 * h * arr16);
              const uint64_t b_lo = ((uint64_t)desc_hi << 32) | (lo + (2 * h + 1) * arr16);
              mma_tf32_ts(d, a_lo, b_hi, idesc, (ks == 0 && h == 0) ? 0u : 1u);
              mma_tf32_ts(d, a_hi, b_lo, idesc, 1u);
              mma_tf32_ts(d, a_hi, b_hi, idesc, 1u);
            }
          }
          tc_commit(smem_u32(&bars->a_empty[as]));       // frees the A stage and the B stage (same index, same depth)
          if (ks == nquads - 1) tc_commit(smem_u32(&bars->acc_full[abuf]));
        }
        __syncwarp();
      }
      ++useq;
    }
  } else if (warp >= kWarpConvTc && warp < kWarpEpiTc && !(P.exp & 1)) {
    // ===== converters =====
    // Iteration `it` of set `set` (of 3) covers the pair of k-steps (16 taps) 3 it + set.  The 32 rows x
    // 16 samples a warp needs per iteration are 32 segments of 64 bytes, `orig` samples apart: they are loaded
    // coalesced (a lane takes sample lane & 15 of rows 2 i + (lane >> 4): two segments per load instruction instead of
    // 32), one iteration ahead, transposed through a per-warp shared-memory tile, and each lane then reads its own
    // row back, splits it and stores it to TMEM.  Tiles whose whole sample range lies inside the clip (all but the
    // first and last of a clip) load unmasked.
    constexpr int kPer = 2, kS = 8 * kPer, kCP = 20;               // kCP: tile pitch (floats), conflict-free 16 B reads
    const int cw = warp - kWarpConvTc, set = cw >> 2, q = warp & 3;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    float* ct = s_c + cw * 32 * kCP;
    const TIn* xin = reinterpret_cast<const TIn*>(P.x);
    const int iters = (P.ksteps + kConvSets - 1) / kConvSets;      // pairs per set, rounded up
    const int lj = lane & 15, lr = lane >> 4;
    const long long rstep = 2LL * P.orig * P.channels;             // elements between the rows of consecutive loads
    const bool stereo32 = P.channels == 2 && ((uintptr_t)P.x & 3) == 0 && (P.row_stride & 1) == 0;
    // fetch cursor: the (unit, iteration) whose samples are loaded next; it runs one iteration ahead of the consume
    // cursor and crosses unit boundaries, so the load pipeline never drains between units
    long long f_u = blockIdx.x, f_sl = 0, f_len = 0;
    const TIn* f_lp = xin;
    int f_it = 0;
    bool f_valid = false, f_interior = false;
    auto enter = [&](long long u) {                          // first live unit at or after u
      f_valid = false;
      for (; u < P.units; u += gridDim.x) {
        const Unit un = locate(P, u);
        if (!un.live) continue;
        f_u = u; f_it = 0; f_valid = true; f_len = un.len;
        f_sl = (un.row0 + q * 32 + lr) * P.orig - P.width + lj;          // my sample of row lr at tap 0
        f_lp = xin + (long long)un.b * P.row_stride + f_sl * P.channels;
        f_interior = un.row0 * P.orig - P.width >= 0 &&
                     (un.row0 + kTileM - 1) * P.orig - P.width + (long long)iters * kConvSets * kS <= un.len;
        return;
      }
    };
    auto advance = [&]() { if (++f_it >= iters) enter(f_u + gridDim.x); };
    // the prefetch registers hold the raw words (float bits, or the 16-bit sample zero-extended); int16 samples are
    // converted when they are written to the tile, so that the 16 loads of an iteration are issued back to back
    auto raw_load = [&](const TIn* p) -> uint32_t {
      if (sizeof(TIn) == 4) return __float_as_uint(__ldg(reinterpret_cast<const float*>(p)));
      if (stereo32) return (uint32_t)__ldg(reinterpret_cast<const int*>(p));         // frame = (ch0 | ch1 << 16)
      return (uint32_t)__ldg(reinterpret_cast<const unsigned short*>(p));
    };
    auto to_float = [&](uint32_t w) -> float {
      return sizeof(TIn) == 4 ? __uint_as_float(w) : pcm16_to_float(w);
    };
    auto fetch = [&](uint32_t (&v)[kS]) {
      const int k0 = kS * (kConvSets * f_it + set);
      const TIn* pk = f_lp + (long long)k0 * P.channels;
      if (P.exp & 128) {
#pragma unroll
        for (int i = 0; i < kS; ++i) v[i] = 0u;
      } else if (f_interior) {
#pragma unroll
        for (int i = 0; i < kS; ++i) v[i] = raw_load(pk + i * rstep);
      } else {
#pragma unroll
        for (int i = 0; i < kS; ++i) {
          const long long sj = f_sl + k0 + 2LL * i * P.orig;
          v[i] = (sj >= 0 && sj < f_len) ? raw_load(pk + i * rstep) : 0u;          // raw 0 is 0.0 in both encodings
        }
      }
    };
    enter(blockIdx.x);
    if (f_valid) {
      uint32_t cur[kS], nxt[kS];
      fetch(cur);
      long long c_u = f_u;
      int c_it = f_it;
      advance();
      uint32_t qseq = 0;                                     // quads of all earlier units
      for (;;) {
        if (f_valid) fetch(nxt);
        __syncwarp();                                        // the previous iteration's reads of the tile are done
#pragma unroll
        for (int i = 0; i < kS; ++i) ct[(2 * i + lr) * kCP + lj] = to_float(cur[i]);
        __syncwarp();
        const int pr = kConvSets * c_it + set;               // my pair of k-steps (phantom pairs of the padding included)
        if (pr < P.ksteps) {
          uint32_t hi[kS], lo[kS];
#pragma unroll
          for (int c = 0; c < kS / 4; ++c) {
            const float4 v4 = *reinterpret_cast<const float4*>(ct + lane * kCP + 4 * c);
            const float rv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t hb = __float_as_uint(rv[j]) & 0xFFFFE000u;
              hi[4 * c + j] = hb;
              lo[4 * c + j] = __float_as_uint(rv[j] - __uint_as_float(hb));
            }
          }
          const uint32_t qs = qseq + (uint32_t)(pr >> 1);
          const uint32_t as = qs % kAStagesTc, ause = qs / kAStagesTc;
          const uint32_t a_base = tmem + lane_addr + kColATc + as * 64 + (uint32_t)(pr & 1) * 32;
          mbar_wait(smem_u32(&bars->a_empty[as]), (ause & 1) ^ 1, P.status, 402);
          tc_fence_after();
          if (!(P.exp & 256)) {
            tmem_st8(a_base, hi);
            tmem_st8(a_base + 8, lo);
            tmem_st8(a_base + 16, hi + 8);
            tmem_st8(a_base + 24, lo + 8);
            tmem_wait_st();
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bars->a_full[as]));
        }
        if (!(f_valid && f_u == c_u)) qseq += (uint32_t)nquads;      // that was this unit's last iteration
        if (!f_valid) break;
        c_u = f_u;
        c_it = f_it;
        advance();
#pragma unroll
        for (int j = 0; j < kS; ++j) cur[j] = nxt[j];
      }
    }
  } else if (warp >= kWarpEpiTc && !(P.exp & 4)) {
    // ===== epilogue =====
    const int q = warp & 3, es = (warp - kWarpEpiTc) >> 2;       // TMEM lane quarter, column-block parity
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    float* tb = s_t + (warp - kWarpEpiTc) * 32 * kTPitch;
    uint32_t useq = 0;
    for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
      const Unit un = locate(P, u);
      const int p0 = un.pass * P.nbp;
      const int pn = min(P.nbp, P.new_f - p0);               // valid phases of this block
      // my 32 rows start at output sample j0 of the clip; everything below is relative to it (32-bit)
      const long long j0 = (un.row0 + q * 32) * P.new_f + p0;
      float* dst = P.out + (long long)un.b * P.out_stride + j0;
      const long long room = P.out_cols - j0, keep = un.olen - j0;     // samples that exist / that are not zero
      if (!un.live) {
        // tile past the clip's end: zeros, row segments of this phase block
        for (int rr = es; rr < 32; rr += kEpiSets)
          for (int c = lane; c < pn; c += 32)
            if ((long long)rr * P.new_f + c < room) dst[rr * P.new_f + c] = 0.f;
        continue;
      }
      const uint32_t abuf = useq & 1;
      mbar_wait(smem_u32(&bars->acc_full[abuf]), (useq >> 1) & 1, P.status, 501);
      tc_fence_after();
      const uint32_t t_acc = tmem + lane_addr + abuf * kMaxNb;
      const int nblk = (P.nbp + 31) >> 5;
      for (int blk = es; blk < nblk; blk += kEpiSets) {
        const int c0 = blk * 32;
        uint32_t va[16], vb[16];
        tmem_ld16(t_acc + c0, va);
        if (c0 + 16 < P.nbp) tmem_ld16(t_acc + c0 + 16, vb);
        tmem_wait_ld();
        if (blk + kEpiSets >= nblk) {                        // my last accumulator columns are in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bars->acc_empty[abuf]));
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) tb[lane * kTPitch + j] = __uint_as_float(va[j]);
        if (c0 + 16 < P.nbp) {
#pragma unroll
          for (int j = 0; j < 16; ++j) tb[lane * kTPitch + 16 + j] = __uint_as_float(vb[j]);
        }
        __syncwarp();
        const int c = c0 + lane;
        if (c < pn) {
#pragma unroll 8
          for (int rr = 0; rr < 32; ++rr) {
            const long long jr = (long long)rr * P.new_f + c;
            float v = tb[rr * kTPitch + lane];
            if (jr >= keep) v = 0.f;
            if (P.clip) v = fminf(fmaxf(v, -1.f), 1.f);
            if (jr < room) dst[jr] = v;
          }
        }
        __syncwarp();
      }
      if (es >= nblk) {                                      // fewer column blocks than epilogue sets: nothing to read
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->acc_empty[abuf]));
      }
      ++useq;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

inline int b_stages(int) { return kAStagesTc; }       // the taps ring shares the A ring's depth and "free" barrier
inline size_t smem_bytes(int nbp) {
  return (size_t)b_stages(nbp) * nbp * 256 + (4 * kEpiSets * 32 * kTPitch + 4 * kConvSets * 32 * 20) * sizeof(float) +
         sizeof(Barriers) + 16;
}

}  // namespace tc
