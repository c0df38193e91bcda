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
//   warp 0      B producer: one bulk copy per k-step (taps hi | lo, [2 k-halves][nbp][4] each, K-major SWIZZLE_NONE)
//   warp 1      MMA issuer (one elected lane), 3 MMAs per k-step
//   warp 2      TMEM allocator
//   warps 4-11  converters: warp (set, q) owns rows 32 q .. 32 q + 31 for the k-steps == set (mod 2): 8 samples of its
//               row straight from the waveform (fp32, or int16 PCM scaled while loading), split, 2 x tcgen05.st.x8
//   warps 12-15 epilogue: tcgen05.ld 32 columns -> per-warp shared-memory transpose -> one 128-byte row segment per
//               store instruction (clip, zeros past the clip's end)
// Work unit = (clip, 128-stride tile, phase block), dealt round-robin to one persistent CTA per SM.
#pragma once

namespace tc {

constexpr int kTileM = 128;
constexpr int kThreadsTc = 512;
constexpr int kBStagesTc = 6;
constexpr int kAStagesTc = 8;          // x 16 TMEM columns (hi 8 | lo 8)
constexpr int kMaxNb = 192;            // phases per block; two accumulator buffers = 384 columns
constexpr int kColATc = 2 * kMaxNb;    // A ring: columns 384..511
constexpr int kWarpConvTc = 4, kWarpEpiTc = 12;
constexpr int kTPitch = 33;            // epilogue transpose buffer pitch (floats)

struct Params {
  const void* x;              // float32 [B][row_stride] or int16 [B][row_stride] (frame-major, channel-minor)
  long long row_stride;
  int channels;
  const long long* lengths;   // [B] or null
  long long L;
  int B;
  int orig, new_f, width, kw;
  int ksteps;                 // ceil(kw / 8)
  int npass, nbp;             // phase blocks, phases per block (multiple of 16)
  const float* tapsB;         // [npass][ksteps][hi | lo][2 k-halves][nbp][4]
  float* out;
  long long out_stride, out_cols;
  int tiles_per_clip;
  long long units;            // B * tiles_per_clip * npass
  int clip;
  int32_t* status;
};

struct __align__(8) Barriers {
  uint64_t b_full[kBStagesTc], b_empty[kBStagesTc];
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
  const uint32_t b_stage_bytes = (uint32_t)P.nbp * 64u;      // 2 arrays x 2 k-halves x nbp rows x 16 B
  uint8_t* s_b = smem_tc;
  float* s_t = reinterpret_cast<float*>(s_b + (size_t)kBStagesTc * b_stage_bytes);     // [4 warps][32][kTPitch]
  Barriers* bars = reinterpret_cast<Barriers*>(s_t + 4 * 32 * kTPitch);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kBStagesTc; ++i) { mbar_init(smem_u32(&bars->b_full[i]), 1); mbar_init(smem_u32(&bars->b_empty[i]), 1); }
    for (int i = 0; i < kAStagesTc; ++i) { mbar_init(smem_u32(&bars->a_full[i]), 4); mbar_init(smem_u32(&bars->a_empty[i]), 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(smem_u32(&bars->acc_full[i]), 1); mbar_init(smem_u32(&bars->acc_empty[i]), 4); }
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
    if (lane == 0) {
      uint32_t seq = 0;
      for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
        const Unit un = locate(P, u);
        if (!un.live) continue;
        for (int ks = 0; ks < P.ksteps; ++ks, ++seq) {
          const uint32_t st = seq % kBStagesTc, use = seq / kBStagesTc;
          mbar_wait(smem_u32(&bars->b_empty[st]), (use & 1) ^ 1, P.status, 101);
          const uint32_t full = smem_u32(&bars->b_full[st]);
          mbar_expect_tx(full, b_stage_bytes);
          bulk_g2s(smem_u32(s_b + (size_t)st * b_stage_bytes),
                   P.tapsB + ((size_t)un.pass * P.ksteps + ks) * (b_stage_bytes / 4), b_stage_bytes, full);
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
      mbar_wait(smem_u32(&bars->acc_empty[abuf]), (ause_acc & 1) ^ 1, P.status, 201);
      for (int ks = 0; ks < P.ksteps; ++ks, ++kseq) {
        const uint32_t as = kseq % kAStagesTc, ause = kseq / kAStagesTc;
        const uint32_t st = kseq % kBStagesTc, use = kseq / kBStagesTc;
        mbar_wait(smem_u32(&bars->a_full[as]), ause & 1, P.status, 202);
        mbar_wait(smem_u32(&bars->b_full[st]), use & 1, P.status, 203);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t a_hi = tmem + kColATc + as * 16, a_lo = a_hi + 8;
          const uint32_t lo = desc_lo0 | (sb0 + st * stage16);
          const uint64_t b_hi = ((uint64_t)desc_hi << 32) | lo;
          const uint64_t b_lo = ((uint64_t)desc_hi << 32) | (lo + arr16);
          mma_tf32_ts(d, a_lo, b_hi, idesc, ks == 0 ? 0u : 1u);     // small terms first
          mma_tf32_ts(d, a_hi, b_lo, idesc, 1u);
          mma_tf32_ts(d, a_hi, b_hi, idesc, 1u);
          tc_commit(smem_u32(&bars->b_empty[st]));
          tc_commit(smem_u32(&bars->a_empty[as]));
          if (ks == P.ksteps - 1) tc_commit(smem_u32(&bars->acc_full[abuf]));
        }
        __syncwarp();
      }
      ++useq;
    }
  } else if (warp >= kWarpConvTc && warp < kWarpEpiTc) {
    // ===== converters =====
    const int set = (warp - kWarpConvTc) >> 2, q = warp & 3;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const TIn* xin = reinterpret_cast<const TIn*>(P.x);
    uint32_t kseq = 0;
    for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
      const Unit un = locate(P, u);
      if (!un.live) continue;
      const TIn* src = xin + (long long)un.b * P.row_stride;
      const long long s0 = (un.row0 + q * 32 + lane) * P.orig - P.width;
      auto fetch = [&](int ks, float (&v)[8]) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const long long s = s0 + 8 * ks + j;
          v[j] = (ks < P.ksteps && s >= 0 && s < un.len) ? load_sample(src, s, P.channels) : 0.f;
        }
      };
      float cur[8], nxt[8];
      fetch(set, cur);
      for (int ks = set; ks < P.ksteps; ks += 2) {
        fetch(ks + 2, nxt);                                 // in flight while this k-step is split and stored
        uint32_t hi[8], lo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t h = __float_as_uint(cur[j]) & 0xFFFFE000u;
          hi[j] = h;
          lo[j] = __float_as_uint(cur[j] - __uint_as_float(h));
        }
        const uint32_t ksq = kseq + ks;
        const uint32_t as = ksq % kAStagesTc, ause = ksq / kAStagesTc;
        const uint32_t a_base = tmem + lane_addr + kColATc + as * 16;
        mbar_wait(smem_u32(&bars->a_empty[as]), (ause & 1) ^ 1, P.status, 402);
        tc_fence_after();
        tmem_st8(a_base, hi);
        tmem_st8(a_base + 8, lo);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->a_full[as]));
#pragma unroll
        for (int j = 0; j < 8; ++j) cur[j] = nxt[j];
      }
      kseq += P.ksteps;
    }
  } else if (warp >= kWarpEpiTc) {
    // ===== epilogue =====
    const int q = warp & 3;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    float* tb = s_t + q * 32 * kTPitch;
    uint32_t useq = 0;
    for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
      const Unit un = locate(P, u);
      float* dst = P.out + (long long)un.b * P.out_stride;
      const int p0 = un.pass * P.nbp;
      const int pn = min(P.nbp, P.new_f - p0);               // valid phases of this block
      const long long rbase = un.row0 + q * 32;               // first stride of my 32 rows
      if (!un.live) {
        // tile past the clip's end: zeros, row segments of this phase block
        for (int rr = 0; rr < 32; ++rr) {
          const long long j0 = (rbase + rr) * P.new_f + p0;
          for (int c = lane; c < pn; c += 32)
            if (j0 + c < P.out_cols) dst[j0 + c] = 0.f;
        }
        continue;
      }
      const uint32_t abuf = useq & 1;
      mbar_wait(smem_u32(&bars->acc_full[abuf]), (useq >> 1) & 1, P.status, 501);
      tc_fence_after();
      const uint32_t t_acc = tmem + lane_addr + abuf * kMaxNb;
      for (int c0 = 0; c0 < P.nbp; c0 += 32) {
        uint32_t va[16], vb[16];
        tmem_ld16(t_acc + c0, va);
        if (c0 + 16 < P.nbp) tmem_ld16(t_acc + c0 + 16, vb);
        tmem_wait_ld();
        if (c0 + 32 >= P.nbp) {                              // my accumulator columns are in registers
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
          for (int rr = 0; rr < 32; ++rr) {
            const long long j = (rbase + rr) * P.new_f + p0 + c;
            if (j < P.out_cols) {
              float v = j < un.olen ? tb[rr * kTPitch + lane] : 0.f;
              if (P.clip) v = fminf(fmaxf(v, -1.f), 1.f);
              dst[j] = v;
            }
          }
        }
        __syncwarp();
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

inline size_t smem_bytes(int nbp) {
  return (size_t)kBStagesTc * nbp * 64 + 4 * 32 * kTPitch * sizeof(float) + sizeof(Barriers) + 16;
}

}  // namespace tc
