// Probe (GPU box): does cuTensorMapEncodeTiled accept OVERLAPPING rows (row stride 1024 B < row extent 5120 B), does a
// box land as expected, and how fast are 34-row x 112-byte boxes?   nvcc -arch=sm_100a -o tma_overlap tma_overlap.cu
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

constexpr int kRows = 34, kCols = 28;

__global__ void k_one(const __grid_constant__ CUtensorMap map, int c0, int r0, int b, float* out) {
  __shared__ __align__(128) float tile[kRows * kCols];
  __shared__ __align__(8) uint64_t bar;
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(kRows * kCols * 4) : "memory");
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(smem_u32(tile)), "l"(&map), "r"(c0), "r"(r0), "r"(b), "r"(smem_u32(&bar)) : "memory");
  }
  uint32_t ok = 0;
  while (!ok) {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
  }
  for (int i = threadIdx.x; i < kRows * kCols; i += blockDim.x) out[i] = tile[i];
}

// throughput: every CTA streams `iters` boxes through a 4-deep ring, one issuing thread, consumers just wait
template <int DEPTH>
__global__ void k_rate(const __grid_constant__ CUtensorMap map, int iters, int rows_avail, long long* cycles) {
  __shared__ __align__(128) float tile[DEPTH][960];   // TMA destinations must be 128-byte aligned: 3808 -> 3840 B
  __shared__ __align__(8) uint64_t bar[DEPTH];
  if (threadIdx.x == 0) {
    for (int i = 0; i < DEPTH; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar[i])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    for (int it = 0; it < iters + DEPTH; ++it) {
      const int st = it % DEPTH;
      if (it >= DEPTH) {
        uint32_t ok = 0; const uint32_t par = ((it - DEPTH) / DEPTH) & 1;
        while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                                 : "=r"(ok) : "r"(smem_u32(&bar[st])), "r"(par) : "memory");
      }
      if (it < iters) {
        const int r0 = (int)((blockIdx.x * 977u + it * 32u) % (unsigned)(rows_avail - 40));
        const int c0 = 24 * (it % 11);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar[st])), "r"(kRows * kCols * 4) : "memory");
        asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                     ::"r"(smem_u32(tile[st])), "l"(&map), "r"(c0), "r"(r0), "r"(0), "r"(smem_u32(&bar[st])) : "memory");
      }
    }
    cycles[blockIdx.x] = clock64() - t0;
  }
}

int main() {
  EncodeFn encode = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&encode, cudaEnableDefault, &q);
  printf("entry point: %s q=%d fn=%p\n", cudaGetErrorString(e), (int)q, (void*)encode);
  if (!encode) return 1;
  const int B = 4; const long long stride = 2646000 / 4 * 4;          // floats per clip row, multiple of 4
  std::vector<float> h((size_t)B * stride);
  for (size_t i = 0; i < h.size(); ++i) h[i] = (float)(i % 1000003);
  float* d; cudaMalloc(&d, h.size() * 4); cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
  const int rows = (int)(stride / 256);
  CUtensorMap map;
  cuuint64_t gdim[3] = {1280, (cuuint64_t)rows, (cuuint64_t)B};
  cuuint64_t gstr[2] = {1024, (cuuint64_t)stride * 4};
  cuuint32_t box[3] = {kCols, kRows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = encode(&map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, d, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  printf("encode overlapping rows (stride 1024 B, extent 5120 B): CUresult %d\n", (int)r);
  if (r != CUDA_SUCCESS) return 2;
  float* out; cudaMalloc(&out, kRows * kCols * 4);
  std::vector<float> ho(kRows * kCols);
  int cases[4][3] = {{24, 100, 1}, {996, 7, 2}, {128 + 240, 300, 3}, {0, -2, 0}};
  for (auto& c : cases) {
    k_one<<<1, 128>>>(map, c[0], c[1], c[2], out);
    cudaError_t ee = cudaDeviceSynchronize();
    cudaMemcpy(ho.data(), out, ho.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int rr = 0; rr < kRows; ++rr)
      for (int cc = 0; cc < kCols; ++cc) {
        const long long row = c[1] + rr;
        const float want = row < 0 ? 0.f : h[(size_t)c[2] * stride + row * 256 + c[0] + cc];
        if (ho[rr * kCols + cc] != want) ++bad;
      }
    printf("box c0=%d r0=%d b=%d: %s, mismatches %d\n", c[0], c[1], c[2], cudaGetErrorString(ee), bad);
  }
  long long* cyc; cudaMalloc(&cyc, 148 * 8);
  for (int iters : {2000, 8000}) {  // depth 8, depth 12
    if (iters == 2000) k_rate<8><<<148, 32>>>(map, iters, rows, cyc); else k_rate<12><<<148, 32>>>(map, iters, rows, cyc);
    cudaError_t e2 = cudaDeviceSynchronize();
    printf("k_rate: %s / %s\n", cudaGetErrorString(e2), cudaGetErrorString(cudaGetLastError()));
    std::vector<long long> hc(148); cudaMemcpy(hc.data(), cyc, 148 * 8, cudaMemcpyDeviceToHost);
    long long mx = 0; for (auto v : hc) mx = v > mx ? v : mx;
    printf("rate: %d boxes of 34x112 B per CTA, 148 CTAs: max %lld cycles = %.1f cycles/box (%.1f B/cycle/SM)\n", iters, mx,
           (double)mx / iters, (double)kRows * kCols * 4 * iters / mx);
  }
  return 0;
}
