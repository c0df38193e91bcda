// k_vq.cuh -- the DVAE codebook search behind the log-mel path (C ABI in include/ttts_vq.h).
//
// Reference: Quantize.forward in eval mode, ttts/vqvae/xtts_dvae.py:87-93,119-120,129-130 (called by
// DiscreteVAE.get_codebook_indices :327-334 from ttts/prepare/extract_vq.py:24-29):
//     dist = |z|^2 - 2 z @ embed + |e|^2 ;  embed_ind = argmax(-dist) ;  quantize = embed[:, embed_ind]
// |z|^2 is the same for every code of a row, so the search is  argmin_j  |e_j|^2 - 2 z . e_j : a
// [rows x dim] x [dim x n_embed] product with an arg-min epilogue (config.json: dim 512, n_embed 8192:
// 8.4 MFLOP per row).  The distances never exist in HBM.
//
//   k_vq_tile    z [rows, stride] -> zt [tiles][pairs of k-steps][4][128 rows][4]: each of the four 16-byte loads a
//                converter warp of k_vq_tc issues per iteration covers 512 contiguous bytes; also resets the keys
//   k_vq_tile_cm the same from the encoder's channel-major output [B, dim, T'] (ttts_vq_nearest_bct)
//   k_vq_tc      the search on tcgen05, same machinery as k_resample_tc / k_logmel2: kind::tf32 MMAs with the A operand
//                (128 rows of z, split hi / lo by converter warps) in TMEM, the B operand (a block of <= 192 codes,
//                pre-tiled and pre-split at plan creation) in shared memory by cp.async.bulk, 3xTF32 products
//                (lo*hi + hi*lo + hi*hi: fp32-grade dot products), two accumulator buffers; the epilogue warps keep a
//                running (score, index) per row and merge blocks with one 64-bit atomicMin per row and unit on the
//                key (orderable score bits << 32 | index): equal scores resolve to the lowest index.
//                Work unit = (tile of 128 rows, block of codes), dealt round-robin to one persistent CTA per SM.
//   k_vq_finish  key -> int64 index; optional gather z + (e - z) and the per-row squared error
//   k_vq_mean    fixed-order fp64 sum of the per-row errors -> diff (deterministic)
//   k_vq_gather  embed_code
// Included at the end of ttts_mel.cu (one shared library); uses its fail() / CUDA_TRY / g_launches and PTX wrappers.
#pragma once

#include "ttts_vq.h"

namespace vq {

constexpr int kTileM = 128;
constexpr int kConvSets = 3, kEpiSets = 2;
constexpr int kThreadsTc = 32 * (4 + 4 * kConvSets + 4 * kEpiSets);   // 4 misc + 12 converter + 8 epilogue warps
#ifndef VQ_NB
#define VQ_NB 192                      // measured on one box: 160 -> 0.508 ms, 176 -> 0.498, 192 -> 0.487 (64 x 216 rows, 8192 codes):
#endif                                 // fewer, longer MMAs (96 cycles) and 43 instead of 52 conversions of every row tile
constexpr int kMaxNb = VQ_NB;          // codes per block; two accumulator buffers = 2 x VQ_NB TMEM columns
constexpr int kStages = (512 - 2 * kMaxNb) / 64;   // A ring: TMEM, 64 columns (a quad of k-steps) per stage
// a converter set's consecutive items are up to two quads apart; a parity wait must never be two phases ahead of its
// barrier, i.e. that distance must not exceed the ring depth (see the converter loop)
static_assert(kStages >= 2 && kMaxNb % 16 == 0 && kMaxNb <= 256, "VQ_NB: multiple of 16, at most 192 for a 2-deep A ring");
#ifndef VQ_B_STAGES
#define VQ_B_STAGES ((220 * 1024) / (VQ_NB * 256) < 5 ? (220 * 1024) / (VQ_NB * 256) : 5)
#endif
constexpr int kBStages = VQ_B_STAGES;  // B ring: shared memory, 48 KB per stage at 192 codes, 4 deep (own "free" barriers)
constexpr int kColA = 2 * kMaxNb;      // A ring: the columns behind the accumulators
constexpr int kWarpConv = 4, kWarpEpi = 4 + 4 * kConvSets;

struct Params {
  const float* zt;            // [tiles][pairs][4 quads of columns][128 rows][4]
  int pairs;                  // pairs of k-steps (16 columns of z each), even: a stage is a quad (32 columns)
  int npass, nbp;             // code blocks, codes per block (multiple of 16)
  const float* codeB;         // [npass][quads][4 k-steps][hi | lo][2 k-halves][nbp][4]
  const float* ee;            // [npass * nbp] |e_j|^2, +inf for the padding codes of the last block
  unsigned long long* keys;   // [tiles * 128]
  long long units;            // tiles * npass
  int exp;                    // timing experiments (RESULTS INVALID): 1 no converters, 2 no codebook loads, 4 no epilogue,
                              // 8 one product per k-step instead of three, 32 no MMAs
  int32_t* status;
};

struct __align__(8) Barriers {
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t a_full[kStages], a_empty[kStages];
  uint64_t acc_full[2], acc_empty[2];
};

__device__ __forceinline__ unsigned long long make_key(float score, uint32_t idx) {
  uint32_t b = __float_as_uint(score);
  b = (b & 0x80000000u) ? ~b : (b | 0x80000000u);          // unsigned order = float order
  return ((unsigned long long)b << 32) | idx;
}

// ------------------------------------------------------------------------------------------------ k_vq_tile
// One thread = four consecutive columns of one row.
__global__ void __launch_bounds__(256) k_vq_tile(const float* __restrict__ z, long long row_stride, long long n_rows,
                                                 int dim, int pairs, long long rows_pad, float* __restrict__ zt,
                                                 unsigned long long* __restrict__ keys, int vec_ok) {
  const unsigned cpr = (unsigned)pairs * 4u;                // float4 chunks per row
  const long long total = rows_pad * cpr;
  for (long long i = blockIdx.x * 256LL + threadIdx.x; i < total; i += 256LL * gridDim.x) {
    const long long row = total <= 0xFFFFFFFFLL ? (long long)((unsigned)i / cpr) : i / cpr;
    const int c = (int)(i - row * cpr), k = 4 * c;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row < n_rows) {
      const float* src = z + row * row_stride + k;
      if (vec_ok && k + 4 <= dim) {
        v = __ldg(reinterpret_cast<const float4*>(src));
      } else {
        if (k < dim) v.x = __ldg(src);
        if (k + 1 < dim) v.y = __ldg(src + 1);
        if (k + 2 < dim) v.z = __ldg(src + 2);
        if (k + 3 < dim) v.w = __ldg(src + 3);
      }
    }
    const long long tile = row / kTileM;
    const int r = (int)(row - tile * kTileM);
    *reinterpret_cast<float4*>(zt + ((((tile * pairs + (c >> 2)) * 4 + (c & 3)) * kTileM + r) * 4)) = v;
    if (c == 0) keys[row] = ~0ull;
  }
}

// Channel-major input z[b, k, t] (the DVAE encoder's own output layout, before the permute of xtts_dvae.py:331): row
// n = b T + t.  One thread = one row x four consecutive channels; consecutive threads take consecutive rows, so both
// the four loads (along t) and the 16-byte store (along the tile's rows) are coalesced.
__global__ void __launch_bounds__(256) k_vq_tile_cm(const float* __restrict__ z, long long batch_stride,
                                                    long long chan_stride, unsigned T, unsigned n_rows, int dim,
                                                    int pairs, unsigned rows_pad, float* __restrict__ zt,
                                                    unsigned long long* __restrict__ keys) {
  const int c = blockIdx.y, k = 4 * c;                      // grid.y = float4 chunks per row: no 64-bit divisions
  for (unsigned row = blockIdx.x * 256u + threadIdx.x; row < rows_pad; row += 256u * gridDim.x) {
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (row < n_rows) {
      const unsigned b = row / T, t = row - b * T;
      const float* src = z + b * batch_stride + (long long)k * chan_stride + t;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (k + j < dim) v[j] = __ldg(src + j * chan_stride);
    }
    const unsigned tile = row / kTileM, r = row % kTileM;
    *reinterpret_cast<float4*>(zt + ((((long long)tile * pairs + (c >> 2)) * 4 + (c & 3)) * kTileM + r) * 4) = make_float4(v[0], v[1], v[2], v[3]);
    if (c == 0) keys[row] = ~0ull;
  }
}

// ------------------------------------------------------------------------------------------------ k_vq_tc
__global__ void __launch_bounds__(kThreadsTc, 1) k_vq_tc(const Params P) {
  extern __shared__ __align__(1024) uint8_t smem_vq[];
  const uint32_t b_stage_bytes = (uint32_t)P.nbp * 256u;     // 4 k-steps x 2 arrays x 2 k-halves x nbp rows x 16 B
  const int nquads = P.pairs >> 1;
  uint8_t* s_b = smem_vq;
  Barriers* bars = reinterpret_cast<Barriers*>(s_b + (size_t)kBStages * b_stage_bytes);
  uint32_t* s_tmem = reinterpret_cast<uint32_t*>(bars + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kBStages; ++i) { mbar_init(smem_u32(&bars->b_full[i]), 1); mbar_init(smem_u32(&bars->b_empty[i]), 1); }
    for (int i = 0; i < kStages; ++i) {
      mbar_init(smem_u32(&bars->a_full[i]), 8);              // 2 pairs x 4 lane quarters per quad
      mbar_init(smem_u32(&bars->a_empty[i]), 1);
    }
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
    // ===== codebook producer: one bulk copy per quad of k-steps =====
    if (lane == 0 && !(P.exp & 2)) {
      uint32_t seq = 0;
      for (long long u = blockIdx.x; u < P.units; u += gridDim.x) {
        const int pass = (int)(u % P.npass);
        for (int ks = 0; ks < nquads; ++ks, ++seq) {
          const uint32_t st = seq % kBStages, use = seq / kBStages;
          mbar_wait(smem_u32(&bars->b_empty[st]), (use & 1) ^ 1, P.status, 1101);
          const uint32_t full = smem_u32(&bars->b_full[st]);
          mbar_expect_tx(full, b_stage_bytes);
          bulk_g2s(smem_u32(s_b + (size_t)st * b_stage_bytes),
                   P.codeB + ((size_t)pass * nquads + ks) * (b_stage_bytes / 4), b_stage_bytes, full);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(P.nbp >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);                // SBO = 128 B, descriptor version 1
    const uint32_t desc_lo0 = (((uint32_t)P.nbp * 16u) >> 4) << 16;   // LBO = distance between the two k-halves
    const uint32_t arr16 = ((uint32_t)P.nbp * 32u) >> 4;              // one operand array, in 16 B units
    const uint32_t sb0 = smem_u32(s_b) >> 4, stage16 = b_stage_bytes >> 4;
    uint32_t kseq = 0, useq = 0;
    for (long long u = blockIdx.x; u < P.units; u += gridDim.x, ++useq) {
      const uint32_t abuf = useq & 1, ause_acc = useq >> 1;
      const uint32_t d = tmem + abuf * kMaxNb;
      if (!(P.exp & 4)) mbar_wait(smem_u32(&bars->acc_empty[abuf]), (ause_acc & 1) ^ 1, P.status, 1201);
      for (int ks = 0; ks < nquads; ++ks, ++kseq) {
        const uint32_t as = kseq % kStages, ause = kseq / kStages;
        const uint32_t st = kseq % kBStages, use = kseq / kBStages;
        if (!(P.exp & 1)) mbar_wait(smem_u32(&bars->a_full[as]), ause & 1, P.status, 1202);
        if (!(P.exp & 2)) mbar_wait(smem_u32(&bars->b_full[st]), use & 1, P.status, 1203);
        tc_fence_after();
        if (elect_one_sync()) {
          const uint32_t a0 = tmem + kColA + as * 64;
          const uint32_t lo = desc_lo0 | (sb0 + st * stage16);
          if (!(P.exp & 32)) {
#pragma unroll
            for (uint32_t h = 0; h < 4; ++h) {               // the four k-steps of the quad; small terms first
              const uint32_t a_hi = a0 + 16 * h, a_lo = a_hi + 8;
              const uint64_t b_hi = ((uint64_t)desc_hi << 32) | (lo + 2 * h * arr16);
              const uint64_t b_lo = ((uint64_t)desc_hi << 32) | (lo + (2 * h + 1) * arr16);
              if (P.exp & 8) {                                 // timing experiment: one product instead of three
                mma_tf32_ts(d, a_hi, b_hi, idesc, (ks == 0 && h == 0) ? 0u : 1u);
                continue;
              }
              mma_tf32_ts(d, a_lo, b_hi, idesc, (ks == 0 && h == 0) ? 0u : 1u);
              mma_tf32_ts(d, a_hi, b_lo, idesc, 1u);
              mma_tf32_ts(d, a_hi, b_hi, idesc, 1u);
            }
          }
          tc_commit(smem_u32(&bars->a_empty[as]));
          tc_commit(smem_u32(&bars->b_empty[st]));
          if (ks == nquads - 1) tc_commit(smem_u32(&bars->acc_full[abuf]));
        }
        __syncwarp();
      }
    }
  } else if (warp >= kWarpConv && warp < kWarpEpi && !(P.exp & 1)) {
    // ===== converters =====
    // Per pair of k-steps: lane = row, 16 columns of z = four 16-byte loads, fetched one item ahead, split into
    // TF32-exact hi and the remainder lo, 4 x tcgen05.st.x8 into the quad's TMEM stage.
    const int cw = warp - kWarpConv, set = cw >> 2, q = warp & 3;
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    // A set walks the CTA's pairs of k-steps as one sequence across units (g = unit * pairs + pair, g = set mod 3): its
    // consecutive items are never more than two quads apart, so the parity wait on a stage's "free" barrier cannot be
    // two phases ahead of the barrier (with a per-unit restart the gap at a unit boundary is three quads: a 2-deep A
    // ring then aliases and deadlocks).
    const long long my_units = (P.units - blockIdx.x + gridDim.x - 1) / gridDim.x;
    struct Cur { long long ui; int pr; };
    auto fetch = [&](float4 (&v)[4], const Cur& c) {
      const long long tile = (blockIdx.x + c.ui * gridDim.x) / P.npass;
      const float4* src = reinterpret_cast<const float4*>(P.zt + (tile * P.pairs + c.pr) * (kTileM * 16)) + q * 32 + lane;
#pragma unroll
      for (int i = 0; i < 4; ++i) v[i] = __ldg(src + i * kTileM);         // 512 contiguous bytes per warp and load
    };
    auto step = [&](Cur& c) {
      c.pr += kConvSets;
      while (c.pr >= P.pairs) { c.pr -= P.pairs; ++c.ui; }
    };
    Cur c{0, set};
    while (c.pr >= P.pairs) { c.pr -= P.pairs; ++c.ui; }
    if (c.ui < my_units) {
      float4 cur[4], nxt[4];
      fetch(cur, c);
      for (;;) {
        Cur n = c;
        step(n);
        const bool n_valid = n.ui < my_units;
        if (n_valid) fetch(nxt, n);
        uint32_t hi[16], lo[16];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float rv[4] = {cur[k].x, cur[k].y, cur[k].z, cur[k].w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t hb = __float_as_uint(rv[j]) & 0xFFFFE000u;
            hi[4 * k + j] = hb;
            lo[4 * k + j] = __float_as_uint(rv[j] - __uint_as_float(hb));
          }
        }
        const uint32_t qs = (uint32_t)c.ui * (uint32_t)nquads + (uint32_t)(c.pr >> 1);     // quad sequence number
        const uint32_t as = qs % kStages, ause = qs / kStages;
        const uint32_t a_base = tmem + lane_addr + kColA + as * 64 + (uint32_t)(c.pr & 1) * 32;
        mbar_wait(smem_u32(&bars->a_empty[as]), (ause & 1) ^ 1, P.status, 1402);
        tc_fence_after();
        tmem_st8(a_base, hi);
        tmem_st8(a_base + 8, lo);
        tmem_st8(a_base + 16, hi + 8);
        tmem_st8(a_base + 24, lo + 8);
        tmem_wait_st();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->a_full[as]));
        if (!n_valid) break;
        c = n;
#pragma unroll
        for (int j = 0; j < 4; ++j) cur[j] = nxt[j];
      }
    }
  } else if (warp >= kWarpEpi && !(P.exp & 4)) {
    // ===== epilogue: running arg-min per row =====
    const int q = warp & 3, es = (warp - kWarpEpi) >> 2;         // TMEM lane quarter, column-block parity
    const uint32_t lane_addr = (uint32_t)(q * 32) << 16;
    const int nblk = (P.nbp + 31) >> 5;
    uint32_t useq = 0;
    for (long long u = blockIdx.x; u < P.units; u += gridDim.x, ++useq) {
      const long long tile = u / P.npass;
      const int p0 = (int)(u - tile * P.npass) * P.nbp;
      const uint32_t abuf = useq & 1;
      mbar_wait(smem_u32(&bars->acc_full[abuf]), (useq >> 1) & 1, P.status, 1501);
      tc_fence_after();
      const uint32_t t_acc = tmem + lane_addr + abuf * kMaxNb;
      float best = __int_as_float(0x7f800000);
      uint32_t bidx = 0xFFFFFFFFu;
      for (int blk = es; blk < nblk; blk += kEpiSets) {
        const int c0 = blk * 32;
        const bool two = c0 + 16 < P.nbp;
        uint32_t va[16], vb[16];
        tmem_ld16(t_acc + c0, va);
        if (two) tmem_ld16(t_acc + c0 + 16, vb);
        tmem_wait_ld();
        if (blk + kEpiSets >= nblk) {                        // my last accumulator columns are in registers
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(smem_u32(&bars->acc_empty[abuf]));
        }
        const float4* e4 = reinterpret_cast<const float4*>(P.ee + p0 + c0);      // warp-uniform addresses
#pragma unroll
        for (int j4 = 0; j4 < 4; ++j4) {
          const float4 e = __ldg(e4 + j4);
          const float ev[4] = {e.x, e.y, e.z, e.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float s = fmaf(-2.f, __uint_as_float(va[4 * j4 + j]), ev[j]);
            if (s < best) { best = s; bidx = (uint32_t)(p0 + c0 + 4 * j4 + j); }
          }
        }
        if (two) {
#pragma unroll
          for (int j4 = 0; j4 < 4; ++j4) {
            const float4 e = __ldg(e4 + 4 + j4);
            const float ev[4] = {e.x, e.y, e.z, e.w};
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float s = fmaf(-2.f, __uint_as_float(vb[4 * j4 + j]), ev[j]);
              if (s < best) { best = s; bidx = (uint32_t)(p0 + c0 + 16 + 4 * j4 + j); }
            }
          }
        }
      }
      if (es >= nblk) {                                      // fewer column blocks than epilogue sets: nothing to read
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(smem_u32(&bars->acc_empty[abuf]));
      }
      if (bidx != 0xFFFFFFFFu) atomicMin(P.keys + tile * kTileM + q * 32 + lane, make_key(best, bidx));
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

inline size_t smem_bytes(int nbp) { return (size_t)kBStages * nbp * 256 + sizeof(Barriers) + 16; }

// ------------------------------------------------------------------------------------------------ k_vq_finish
// One warp per row.
__global__ void __launch_bounds__(256) k_vq_finish(const unsigned long long* __restrict__ keys, long long n_rows,
                                                   const float* __restrict__ z, long long row_stride, int dim,
                                                   const float* __restrict__ embT, long long* __restrict__ ind,
                                                   float* __restrict__ quant, float* __restrict__ row_err) {
  const int lane = threadIdx.x & 31;
  for (long long row = blockIdx.x * 8LL + (threadIdx.x >> 5); row < n_rows; row += 8LL * gridDim.x) {
    uint32_t idx = (uint32_t)keys[row];
    if (idx == 0xFFFFFFFFu) idx = 0;                         // every score NaN: torch.max returns the first
    if (lane == 0) ind[row] = (long long)idx;
    if (!quant && !row_err) continue;
    const float* e = embT + (long long)idx * dim;
    const float* zr = z + row * row_stride;
    float acc = 0.f;
    for (int k = lane; k < dim; k += 32) {
      const float zv = __ldg(zr + k);
      const float d = __fsub_rn(__ldg(e + k), zv);
      if (quant) quant[row * dim + k] = __fadd_rn(zv, d);    // input + (quantize - input), xtts_dvae.py:120
      acc = fmaf(d, d, acc);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (row_err && lane == 0) row_err[row] = acc;
  }
}

// One block: fixed-order fp64 sum of the per-row errors -> mean over rows * dim.
__global__ void __launch_bounds__(1024) k_vq_mean(const float* __restrict__ row_err, long long n_rows, double denom,
                                                  float* __restrict__ out) {
  __shared__ double s[1024];
  double a = 0.0;
  for (long long i = threadIdx.x; i < n_rows; i += 1024) a += (double)row_err[i];
  s[threadIdx.x] = a;
  __syncthreads();
  for (int o = 512; o > 0; o >>= 1) {
    if ((int)threadIdx.x < o) s[threadIdx.x] += s[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) *out = (float)(s[0] / denom);
}

// embed_code: one warp per row.
__global__ void __launch_bounds__(256) k_vq_gather(const long long* __restrict__ ind, long long n_rows, int dim,
                                                   int n_embed, const float* __restrict__ embT,
                                                   float* __restrict__ out, long long out_stride, int32_t* status) {
  const int lane = threadIdx.x & 31;
  for (long long row = blockIdx.x * 8LL + (threadIdx.x >> 5); row < n_rows; row += 8LL * gridDim.x) {
    const long long idx = ind[row];
    const bool ok = idx >= 0 && idx < n_embed;
    if (!ok && lane == 0) atomicExch(status, 1601);
    const float* e = embT + (ok ? idx : 0) * dim;
    for (int k = lane; k < dim; k += 32) out[row * out_stride + k] = ok ? __ldg(e + k) : 0.f;
  }
}

}  // namespace vq

struct ttts_vq_plan {
  int dim = 0, n_embed = 0, pairs = 0, npass = 0, nbp = 0, num_sms = 0, device = 0;
  float* d_codeB = nullptr;   // [npass][pairs][2 k-steps][hi | lo][2 k-halves][nbp][4]
  float* d_ee = nullptr;      // [npass * nbp]
  float* d_embT = nullptr;    // [n_embed][dim]
  int32_t* d_status = nullptr;
};

namespace {
struct VqLayout { size_t zt, keys, row_err, total; long long tiles; };
VqLayout vq_layout(const ttts_vq_plan* p, long long n_rows) {
  VqLayout l;
  l.tiles = (n_rows + vq::kTileM - 1) / vq::kTileM;
  auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
  l.zt = 0;
  l.keys = up((size_t)l.tiles * p->pairs * vq::kTileM * 16 * sizeof(float));
  l.row_err = l.keys + up((size_t)l.tiles * vq::kTileM * sizeof(unsigned long long));
  l.total = l.row_err + up((size_t)l.tiles * vq::kTileM * sizeof(float));
  return l;
}
}  // namespace

extern "C" void ttts_vq_plan_destroy(ttts_vq_plan* p) {
  if (!p) return;
  cudaFree(p->d_codeB);
  cudaFree(p->d_ee);
  cudaFree(p->d_embT);
  cudaFree(p->d_status);
  delete p;
}

extern "C" int ttts_vq_plan_create(const float* embed, int32_t dim, int32_t n_embed, ttts_vq_plan** out) {
  if (!embed || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (dim < 1 || dim > 4096) return fail(TTTS_MEL_EINVAL, "codebook dim %d outside 1..4096", dim);
  if (n_embed < 16 || n_embed > (1 << 24)) return fail(TTTS_MEL_EINVAL, "n_embed %d outside 16..2^24", n_embed);
  auto* p = new (std::nothrow) ttts_vq_plan();
  if (!p) return fail(TTTS_MEL_ENOMEM, "host allocation failed");
  p->dim = dim; p->n_embed = n_embed;
  p->pairs = (dim + 31) / 32 * 2;
  p->npass = (n_embed + vq::kMaxNb - 1) / vq::kMaxNb;
  p->nbp = ((n_embed + p->npass - 1) / p->npass + 15) / 16 * 16;
  const size_t arr = (size_t)2 * p->nbp * 4;          // one operand array: [2 k-halves][nbp][4]
  std::vector<float> tb((size_t)p->npass * p->pairs * 4 * arr, 0.f), embT((size_t)n_embed * dim);
  std::vector<float> ee((size_t)p->npass * p->nbp, INFINITY);
  for (int j = 0; j < n_embed; ++j) {
    double s = 0.0;
    for (int k = 0; k < dim; ++k) {
      const float v = embed[(size_t)k * n_embed + j];
      embT[(size_t)j * dim + k] = v;
      s += (double)v * v;
    }
    ee[j] = (float)s;
  }
  for (int pb = 0; pb < p->npass; ++pb)
    for (int pr = 0; pr < p->pairs; ++pr) {
      float* stage = tb.data() + ((size_t)pb * p->pairs + pr) * 4 * arr;       // [k-step of the pair][hi | lo]
      for (int n = 0; n < p->nbp; ++n) {
        const int j = pb * p->nbp + n;
        if (j >= n_embed) continue;
        for (int kk = 0; kk < 16; ++kk) {
          const int k = 16 * pr + kk;
          if (k >= dim) continue;
          const float v = embT[(size_t)j * dim + k];
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
    }
  cudaError_t e = cudaGetDevice(&p->device);
  cudaDeviceProp prop;
  if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, p->device);
  if (e == cudaSuccess) p->num_sms = prop.multiProcessorCount;
  if (e == cudaSuccess) e = cudaMalloc(&p->d_codeB, tb.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->d_codeB, tb.data(), tb.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&p->d_ee, ee.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->d_ee, ee.data(), ee.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&p->d_embT, embT.size() * sizeof(float));
  if (e == cudaSuccess) e = cudaMemcpy(p->d_embT, embT.data(), embT.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = cudaMalloc(&p->d_status, sizeof(int32_t));
  if (e == cudaSuccess) e = cudaMemset(p->d_status, 0, sizeof(int32_t));
  if (e == cudaSuccess)
    e = cudaFuncSetAttribute(vq::k_vq_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)vq::smem_bytes(vq::kMaxNb));
  if (e != cudaSuccess) {
    int rc = fail(TTTS_MEL_ECUDA, "vq plan: %s", cudaGetErrorString(e));
    ttts_vq_plan_destroy(p);
    return rc;
  }
  *out = p;
  return 0;
}

extern "C" size_t ttts_vq_workspace_bytes(const ttts_vq_plan* p, int64_t n_rows) {
  if (!p || n_rows <= 0) return 0;
  return vq_layout(p, n_rows).total;
}

namespace {
// z_rm != null: rows [n_rows, row_stride]; else channel-major z_cm[b, k, t] with n_rows = B T
int vq_nearest_impl(const ttts_vq_plan* p, const float* z_rm, int64_t row_stride, const float* z_cm, int64_t batch_stride,
                    int64_t chan_stride, int64_t T, int64_t n_rows, int64_t* ind, float* quant, float* diff, void* ws,
                    size_t ws_bytes, void* stream) {
  int cur_dev = -1;
  CUDA_TRY(cudaGetDevice(&cur_dev));
  if (cur_dev != p->device) return fail(TTTS_MEL_EINVAL, "plan lives on device %d, the current device is %d", p->device, cur_dev);
  if (n_rows > 0x7FFFFF00LL) return fail(TTTS_MEL_EINVAL, "n_rows %lld: split the batch (limit 2^31 - 256)", (long long)n_rows);
  const VqLayout l = vq_layout(p, n_rows);
  if (ws_bytes < l.total) return fail(TTTS_MEL_EINVAL, "workspace %zu < %zu bytes", ws_bytes, l.total);
  if ((uintptr_t)ws % 256) return fail(TTTS_MEL_EINVAL, "workspace not 256-byte aligned");
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* w = (uint8_t*)ws;
  float* zt = (float*)(w + l.zt);
  auto* keys = (unsigned long long*)(w + l.keys);
  float* row_err = (float*)(w + l.row_err);
  const long long rows_pad = l.tiles * vq::kTileM;
  {
    const long long total = rows_pad * p->pairs * 4;
    const int grid = (int)std::min<long long>((total + 255) / 256, (long long)p->num_sms * 16);
    if (z_rm) {
      const int vec_ok = (row_stride % 4 == 0) && ((uintptr_t)z_rm % 16 == 0);
      vq::k_vq_tile<<<grid, 256, 0, st>>>(z_rm, row_stride, n_rows, p->dim, p->pairs, rows_pad, zt, keys, vec_ok);
    } else {
      const unsigned gx = (unsigned)std::min<long long>((rows_pad + 255) / 256, (long long)p->num_sms * 4);
      vq::k_vq_tile_cm<<<dim3(gx, (unsigned)(p->pairs * 4)), 256, 0, st>>>(z_cm, batch_stride, chan_stride, (unsigned)T, (unsigned)n_rows,
                                                                         p->dim, p->pairs, (unsigned)rows_pad, zt, keys);
    }
  }
  vq::Params q;
  q.zt = zt; q.pairs = p->pairs; q.npass = p->npass; q.nbp = p->nbp; q.codeB = p->d_codeB; q.ee = p->d_ee;
  q.keys = keys; q.units = l.tiles * p->npass; q.status = p->d_status;
  q.exp = lab_env("TTTS_VQ_EXP") ? atoi(lab_env("TTTS_VQ_EXP")) : 0;
  {
    const int grid = (int)std::min<long long>(p->num_sms, q.units);
    vq::k_vq_tc<<<grid, vq::kThreadsTc, vq::smem_bytes(p->nbp), st>>>(q);
  }
  {
    const int grid = (int)std::min<long long>((n_rows + 7) / 8, (long long)p->num_sms * 16);
    vq::k_vq_finish<<<grid, 256, 0, st>>>(keys, n_rows, z_rm, row_stride, p->dim, p->d_embT, (long long*)ind, quant,
                                          diff ? row_err : nullptr);
  }
  int launches = 3;
  if (diff) {
    vq::k_vq_mean<<<1, 1024, 0, st>>>(row_err, n_rows, (double)n_rows * p->dim, diff);
    ++launches;
  }
  CUDA_TRY(cudaGetLastError());
  g_launches = launches;
  return 0;
}
}  // namespace

extern "C" int ttts_vq_nearest(const ttts_vq_plan* p, const float* z, int64_t row_stride, int64_t n_rows,
                               int64_t* ind, float* quant, float* diff, void* ws, size_t ws_bytes, void* stream) {
  if (!p || !z || !ind || !ws) return fail(TTTS_MEL_EINVAL, "null argument");
  if (n_rows <= 0) return fail(TTTS_MEL_EINVAL, "empty input (n_rows=%lld)", (long long)n_rows);
  if (row_stride < p->dim) return fail(TTTS_MEL_EINVAL, "row_stride %lld < dim %d", (long long)row_stride, p->dim);
  return vq_nearest_impl(p, z, row_stride, nullptr, 0, 0, 0, n_rows, ind, quant, diff, ws, ws_bytes, stream);
}

extern "C" int ttts_vq_nearest_bct(const ttts_vq_plan* p, const float* z, int64_t batch_stride, int64_t chan_stride,
                                   int64_t B, int64_t T, int64_t* ind, void* ws, size_t ws_bytes, void* stream) {
  if (!p || !z || !ind || !ws) return fail(TTTS_MEL_EINVAL, "null argument");
  if (B <= 0 || T <= 0) return fail(TTTS_MEL_EINVAL, "empty input (B=%lld T=%lld)", (long long)B, (long long)T);
  if (chan_stride < T || batch_stride < chan_stride * p->dim)
    return fail(TTTS_MEL_EINVAL, "strides (%lld, %lld) too small for [B, %d, %lld]", (long long)batch_stride, (long long)chan_stride, p->dim, (long long)T);
  return vq_nearest_impl(p, nullptr, 0, z, batch_stride, chan_stride, T, B * T, ind, nullptr, nullptr, ws, ws_bytes, stream);
}

extern "C" int ttts_vq_embed_code(const ttts_vq_plan* p, const int64_t* ind, int64_t n_rows, float* out,
                                  int64_t out_stride, void* stream) {
  if (!p || !ind || !out) return fail(TTTS_MEL_EINVAL, "null argument");
  if (n_rows <= 0) return fail(TTTS_MEL_EINVAL, "empty input (n_rows=%lld)", (long long)n_rows);
  if (out_stride < p->dim) return fail(TTTS_MEL_EINVAL, "out_row_stride %lld < dim %d", (long long)out_stride, p->dim);
  const int grid = (int)std::min<long long>((n_rows + 7) / 8, (long long)p->num_sms * 16);
  vq::k_vq_gather<<<grid, 256, 0, (cudaStream_t)stream>>>((const long long*)ind, n_rows, p->dim, p->n_embed, p->d_embT,
                                                          out, out_stride, p->d_status);
  CUDA_TRY(cudaGetLastError());
  g_launches = 1;
  return 0;
}

extern "C" int ttts_vq_status(const ttts_vq_plan* p, void* stream) {
  if (!p) return fail(TTTS_MEL_EINVAL, "null plan");
  CUDA_TRY(cudaStreamSynchronize((cudaStream_t)stream));
  int32_t s = 0;
  CUDA_TRY(cudaMemcpy(&s, p->d_status, sizeof(s), cudaMemcpyDeviceToHost));
  if (s != 0) {
    cudaMemset(p->d_status, 0, sizeof(int32_t));
    return fail(TTTS_MEL_EDEVICE, "vq kernels flagged device status %d (%s)", s,
                s == 1601 ? "embed_code index out of range" : "pipeline wait timed out");
  }
  return 0;
}
