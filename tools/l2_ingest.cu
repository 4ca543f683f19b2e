// How many bytes per clock one SM can take from L2 through the TMA unit, and what all SMs take together:
// the denominator of the pair kernel's K loop ("the third roof").
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/l2_ingest tools/l2_ingest.cu && tools/l2_ingest
// One CTA per SM, one thread issues 1-D bulk copies (cp.async.bulk, CH bytes each) of an L2-resident region into a
// ring of NS shared-memory slots, NS copies in flight; bytes / clock64 cycles per CTA, averaged over the CTAs.
// The second half repeats it with 2-D TILED tensor-map loads ([rows x 64] 16-bit boxes, SWIZZLE_128B), issued by one or two
// threads: the form the GEMM kernels' operand tiles take.
// Regions: "own" = every CTA reads its own 256 KB (distinct lines), "shared" = all CTAs read the same 256 KB (the
// operand tile several CTAs of one row block want at once).
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

template <int CH, int NS>
__global__ void __launch_bounds__(128) ingest(const char* __restrict__ src, size_t region, size_t cta_stride, int iters,
                                              unsigned long long* cycles) {
  extern __shared__ __align__(128) unsigned char ring[];
  __shared__ __align__(8) uint64_t bar[NS];
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    const char* base = src + (size_t)blockIdx.x * cta_stride;
    const long long t0 = clock64();
    for (int i = 0; i < iters + NS; ++i) {
      const int s = i % NS;
      if (i >= NS) {
        const uint32_t parity = ((i / NS) - 1) & 1;
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                       : "=r"(ok) : "r"(smem_u32(&bar[s])), "r"(parity) : "memory");
      }
      if (i < iters) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(&bar[s])), "r"(CH) : "memory");
        const char* g = base + ((size_t)i * CH) % region;
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(
                         smem_u32(ring + (size_t)s * CH)), "l"(g), "r"(CH), "r"(smem_u32(&bar[s])) : "memory");
      }
    }
    cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

// the same ring fed by 2-D TILED tensor-map loads: [ROWS x 64] 16-bit boxes (128-byte rows, SWIZZLE_128B) - the
// operand tiles of the GEMM kernels.  `nthr` threads (one per warp) share the issuing, slot s belongs to thread s % nthr.
template <int ROWS, int NS>
__global__ void __launch_bounds__(128) ingest_tiled(const __grid_constant__ CUtensorMap tm, int rows_per_cta, int cta_stride_rows,
                                                    int kcols, int iters, int nthr, unsigned long long* cycles) {
  extern __shared__ __align__(128) unsigned char ring_raw[];
  unsigned char* ring = ring_raw + ((1024u - (smem_u32(ring_raw) & 1023u)) & 1023u);     // swizzle atoms are 1 KB
  __shared__ __align__(8) uint64_t bar[NS];
  constexpr int CH = ROWS * 128;
  if (threadIdx.x == 0) {
    for (int s = 0; s < NS; ++s)
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bar[s])));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  long long t0 = 0;
  if ((threadIdx.x & 31) == 0 && w < nthr) {
    const int row0 = blockIdx.x * cta_stride_rows;
    t0 = clock64();
    for (int i = w; i < iters + NS; i += nthr) {        // NS % nthr == 0: thread w owns slots w, w + nthr, ...
      const int s = i % NS;
      if (i >= NS) {
        const uint32_t parity = ((i / NS) - 1) & 1;
        uint32_t ok = 0;
        while (!ok)
          asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
                       : "=r"(ok) : "r"(smem_u32(&bar[s])), "r"(parity) : "memory");
      }
      if (i < iters) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_u32(&bar[s])), "r"(CH) : "memory");
        const int kb = i % (kcols / 64);
        const int r = row0 + ((i / (kcols / 64)) * ROWS) % rows_per_cta;
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
                     ::"r"(smem_u32(ring + (size_t)s * CH)), "l"(reinterpret_cast<uint64_t>(&tm)), "r"(smem_u32(&bar[s])),
                       "r"(kb * 64), "r"(r) : "memory");
      }
    }
    if (w == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
  }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int ROWS, int NS>
static void run_tiled(const char* name, void* src, int total_rows, int kcols, int rows_per_cta, int stride_rows, int ctas,
                      int iters, int nthr) {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  CUtensorMap tm;
  cuuint64_t gdim[2] = {(cuuint64_t)kcols, (cuuint64_t)total_rows};
  cuuint64_t gstr[1] = {(cuuint64_t)kcols * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)ROWS};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = ((EncodeFn)fp)(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, src, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) { printf("encode failed %d\n", (int)rc); exit(1); }
  unsigned long long* d_cyc;
  CK(cudaMalloc(&d_cyc, ctas * sizeof(unsigned long long)));
  const int smem = ROWS * 128 * NS + 1024;
  CK(cudaFuncSetAttribute(ingest_tiled<ROWS, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  ingest_tiled<ROWS, NS><<<ctas, 128, smem>>>(tm, rows_per_cta, stride_rows, kcols, iters / 8, nthr, d_cyc);
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  ingest_tiled<ROWS, NS><<<ctas, 128, smem>>>(tm, rows_per_cta, stride_rows, kcols, iters, nthr, d_cyc);
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  std::vector<unsigned long long> h(ctas);
  CK(cudaMemcpy(h.data(), d_cyc, ctas * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  double sum = 0;
  for (auto c : h) sum += (double)c;
  const double bytes = (double)ROWS * 128 * iters;
  printf("{\"case\": \"%s\", \"ctas\": %d, \"box_rows\": %d, \"in_flight\": %d, \"issuing_threads\": %d, "
         "\"bytes_per_clk_per_sm\": %.1f, \"chip_tb_s\": %.2f, \"ms\": %.3f}\n",
         name, ctas, ROWS, NS, nthr, bytes / (sum / ctas), bytes * ctas / (ms * 1e-3) / 1e12, ms);
  CK(cudaFree(d_cyc));
  CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
}

template <int CH, int NS>
static void run(const char* name, const char* src, size_t region, size_t stride, int ctas, int iters) {
  unsigned long long* d_cyc;
  CK(cudaMalloc(&d_cyc, ctas * sizeof(unsigned long long)));
  const int smem = CH * NS;
  CK(cudaFuncSetAttribute(ingest<CH, NS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b));
  ingest<CH, NS><<<ctas, 128, smem>>>(src, region, stride, iters / 8, d_cyc);      // warm L2
  CK(cudaDeviceSynchronize());
  CK(cudaEventRecord(a));
  ingest<CH, NS><<<ctas, 128, smem>>>(src, region, stride, iters, d_cyc);
  CK(cudaEventRecord(b));
  CK(cudaDeviceSynchronize());
  float ms = 0;
  CK(cudaEventElapsedTime(&ms, a, b));
  std::vector<unsigned long long> h(ctas);
  CK(cudaMemcpy(h.data(), d_cyc, ctas * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  double sum = 0, mx = 0;
  for (auto c : h) { sum += (double)c; if ((double)c > mx) mx = (double)c; }
  const double bytes = (double)CH * iters;
  printf("{\"case\": \"%s\", \"ctas\": %d, \"chunk\": %d, \"in_flight\": %d, \"bytes_per_clk_per_sm\": %.1f, "
         "\"bytes_per_clk_slowest_sm\": %.1f, \"chip_tb_s\": %.2f, \"ms\": %.3f}\n",
         name, ctas, CH, NS, bytes / (sum / ctas), bytes / mx, bytes * ctas / (ms * 1e-3) / 1e12, ms);
  CK(cudaFree(d_cyc));
  CK(cudaEventDestroy(a)); CK(cudaEventDestroy(b));
}

int main() {
  int sms = 0;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const size_t region = 256 << 10;
  char* src;
  CK(cudaMalloc(&src, region * sms));
  CK(cudaMemset(src, 1, region * sms));
  const int iters = 4000;
  const int counts[] = {sms, sms / 2, 32, 8, 1};
  for (int c : counts) {
    run<16384, 12>("own region, 16 KB copies", src, region, region, c, iters);
    run<16384, 12>("shared region, 16 KB copies", src, region, 0, c, iters);
  }
  run<4096, 32>("own region, 4 KB copies", src, region, region, sms, iters * 4);
  run<32768, 6>("own region, 32 KB copies", src, region, region, sms, iters / 2);
  // tiled loads: a [148 * 256 rows, 512 columns] bf16 matrix (256 KB per CTA, as above); "shared" = every CTA walks
  // the first 256 rows
  const int kcols = 512, rpc = 256;
  for (int c : {sms, 32, 1}) {
    run_tiled<128, 12>("tiled 2-D boxes 128x64, own rows", src, rpc * sms, kcols, rpc, rpc, c, iters, 1);
    run_tiled<128, 12>("tiled 2-D boxes 128x64, own rows", src, rpc * sms, kcols, rpc, rpc, c, iters, 2);
    run_tiled<128, 12>("tiled 2-D boxes 128x64, shared rows", src, rpc * sms, kcols, rpc, 0, c, iters, 2);
  }
  // per-THREAD issue latency or per-SM operation rate?  the same boxes from 3 and 4 issuing threads
  run_tiled<128, 12>("tiled 2-D boxes 128x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters, 3);
  run_tiled<128, 12>("tiled 2-D boxes 128x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters, 4);
  run_tiled<64, 24>("tiled 2-D boxes 64x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters * 2, 1);
  run_tiled<64, 24>("tiled 2-D boxes 64x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters * 2, 3);
  run_tiled<64, 24>("tiled 2-D boxes 64x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters * 2, 4);
  run_tiled<16, 48>("tiled 2-D boxes 16x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters * 4, 1);
  run_tiled<16, 48>("tiled 2-D boxes 16x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters * 4, 4);
  run_tiled<64, 24>("tiled 2-D boxes 64x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters * 2, 2);
  run_tiled<256, 6>("tiled 2-D boxes 256x64, own rows", src, rpc * sms, kcols, rpc, rpc, sms, iters / 2, 2);
  CK(cudaFree(src));
  return 0;
}
