// Token reductions for the LoRA gradients (HBM-bound):
//   dB[n,j] += s · Σ_m dY[m,n]·T[m,j]       dA[j,k] += s · Σ_m dT[m,j]·X[m,k]
// Both are the same shape of work:  G[c,j] = s · Σ_m Z[m,c]·P[m,j]  with Z = dY or X
// ([M,C], row-major) and P = T or dT ([M,r] compact), so ONE launch covers both: the first
// `blocks0` CTAs work on problem 0, the rest on problem 1.  Each CTA owns a strip of CT columns
// and a slice of rows; 8 lanes x 16-byte loads cover one 128-byte row segment, 4 rows per warp
// instruction, and 4 such loads are kept in flight per thread.  Partial sums go lane -> warp
// (shuffle) -> CTA (smem) -> either fp32 atomics into the gradient accumulator (default) or a
// [splits, C, RP] workspace reduced by a second kernel (deterministic mode).
//
// 16-bit io takes the tensor-pipe variant below (lora_grad_mma_kernel): the SIMT kernel needs
// r/2 FMA per byte plus conversions, which at r = 4 is ~70 % of the SM's issue slots at full HBM
// rate -- it measured 1-2 TB/s.  The variant streams [128 tokens x 64 columns] boxes with TMA
// (128-byte swizzle) through a 4-stage mbarrier ring and contracts over tokens with
// ldmatrix.trans + mma.sync m16n8k16 (2 instructions per 512 bytes), so the kernel is bound by
// HBM, not by issue.  fp32 io (the correctness mode) keeps the SIMT kernel.
#include <cuda.h>
#include <mutex>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

bool pdl_enabled();   // linear_tc2.cu

namespace red {
constexpr int NT = 256;
constexpr int UNROLL = 4;
}

struct RedProblem {
  const void* Z; long long ldz;
  const void* P;
  float* out; long long out_cs, out_js;
  float* partial;
  int C, col_blocks, splits;
  long long rows_per_block;
};

template <typename T, int RP>
__device__ __forceinline__ void load_p(const T* __restrict__ P, long long m, int r, int ldp, float (&pv)[RP]) {
  if (r == RP && ldp == RP) {  // compact row of exactly RP elements: vector loads (row stride = RP*sizeof(T))
    constexpr int BYTES = RP * (int)sizeof(T);
    if constexpr (BYTES % 16 == 0) {
      const uint4* src = reinterpret_cast<const uint4*>(P + m * RP);
#pragma unroll
      for (int i = 0; i < BYTES / 16; ++i) {
        const uint4 raw = __ldg(src + i);
        const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
        for (int j = 0; j < 16 / (int)sizeof(T); ++j) pv[i * (16 / (int)sizeof(T)) + j] = to_f32(e[j]);
      }
    } else {  // 8 bytes (r = 4, 16-bit)
      const uint2 raw = __ldg(reinterpret_cast<const uint2*>(P + m * RP));
      const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int j = 0; j < RP; ++j) pv[j] = to_f32(e[j]);
    }
  } else {
#pragma unroll
    for (int j = 0; j < RP; ++j) pv[j] = (j < r) ? to_f32(P[m * ldp + j]) : 0.f;
  }
}

template <typename T, int RP>
__global__ void __launch_bounds__(red::NT)
lora_grad_reduce_kernel(const RedProblem p0, const RedProblem p1, int blocks0, int r, int ldp, float scale,
                        long long M) {   // r ranks of P starting at the problem's P pointer, row stride ldp
  constexpr int VEC = 16 / sizeof(T);   // columns per lane
  constexpr int CT = 8 * VEC;           // columns per block
  __shared__ float red_s[8][CT][RP + 1];

  const bool second = (int)blockIdx.x >= blocks0;
  const RedProblem& p = second ? p1 : p0;
  const int bid = second ? blockIdx.x - blocks0 : blockIdx.x;
  const int cb = bid % p.col_blocks;
  const int sp = bid / p.col_blocks;
  const T* __restrict__ Z = reinterpret_cast<const T*>(p.Z);
  const T* __restrict__ P = reinterpret_cast<const T*>(p.P);

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int cl = lane & 7;                   // column lane
  const int rl = (warp << 2) | (lane >> 3);  // row lane 0..31
  const int c0 = cb * CT + cl * VEC;
  const long long m_begin = (long long)sp * p.rows_per_block;
  long long m_end = m_begin + p.rows_per_block;
  if (m_end > M) m_end = M;

  float acc[VEC][RP];
#pragma unroll
  for (int v = 0; v < VEC; ++v)
#pragma unroll
    for (int j = 0; j < RP; ++j) acc[v][j] = 0.f;

  if (c0 < p.C) {  // C % VEC == 0 is guaranteed by the host
    long long m = m_begin + rl;
    // main loop: UNROLL independent row loads in flight
    for (; m + 32 * (red::UNROLL - 1) < m_end; m += 32 * red::UNROLL) {
      uint4 raw[red::UNROLL];
      float pv[red::UNROLL][RP];
#pragma unroll
      for (int u = 0; u < red::UNROLL; ++u)
        raw[u] = __ldg(reinterpret_cast<const uint4*>(Z + (m + 32 * u) * p.ldz + c0));
#pragma unroll
      for (int u = 0; u < red::UNROLL; ++u) load_p<T, RP>(P, m + 32 * u, r, ldp, pv[u]);
#pragma unroll
      for (int u = 0; u < red::UNROLL; ++u) {
        const T* zv = reinterpret_cast<const T*>(&raw[u]);
#pragma unroll
        for (int v = 0; v < VEC; ++v) {
          const float z = to_f32(zv[v]);
#pragma unroll
          for (int j = 0; j < RP; ++j) acc[v][j] = fmaf(z, pv[u][j], acc[v][j]);
        }
      }
    }
    for (; m < m_end; m += 32) {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(Z + m * p.ldz + c0));
      float pv[RP];
      load_p<T, RP>(P, m, r, ldp, pv);
      const T* zv = reinterpret_cast<const T*>(&raw);
#pragma unroll
      for (int v = 0; v < VEC; ++v) {
        const float z = to_f32(zv[v]);
#pragma unroll
        for (int j = 0; j < RP; ++j) acc[v][j] = fmaf(z, pv[j], acc[v][j]);
      }
    }
  }
  // lanes cl, cl+8, cl+16, cl+24 hold the same columns
#pragma unroll
  for (int v = 0; v < VEC; ++v)
#pragma unroll
    for (int j = 0; j < RP; ++j) {
      float a = acc[v][j];
      a += __shfl_xor_sync(0xffffffffu, a, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16);
      acc[v][j] = a;
    }
  if (lane < 8) {
#pragma unroll
    for (int v = 0; v < VEC; ++v)
#pragma unroll
      for (int j = 0; j < RP; ++j) red_s[warp][cl * VEC + v][j] = acc[v][j];
  }
  __syncthreads();
  for (int idx = tid; idx < CT * RP; idx += red::NT) {
    const int c = idx / RP, j = idx % RP;
    const int gc = cb * CT + c;
    if (gc >= p.C || j >= r) continue;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red_s[w][c][j];
    if (p.partial != nullptr) {
      p.partial[((long long)sp * p.C + gc) * RP + j] = s;
    } else {
      atomicAdd(p.out + gc * p.out_cs + j * p.out_js, s * scale);
    }
  }
}

template <int RP>
__global__ void lora_grad_finalize_kernel(const RedProblem p0, const RedProblem p1, int n0, int r,
                                          float scale) {
  int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const bool second = idx >= n0;
  const RedProblem& p = second ? p1 : p0;
  if (second) idx -= n0;
  if (p.partial == nullptr || idx >= p.C * RP) return;
  const int c = idx / RP, j = idx % RP;
  if (j >= r) return;
  float s = 0.f;
  for (int y = 0; y < p.splits; ++y) s += p.partial[((long long)y * p.C + c) * RP + j];
  p.out[c * p.out_cs + j * p.out_js] += s * scale;
}

// ------------------------------------------------------------------------------------------
// tensor-pipe variant (16-bit io)
// ------------------------------------------------------------------------------------------
namespace rtc {
constexpr int TOK = 128;                     // tokens per stage: 16 per consumer warp
constexpr int COLS = 64;                     // columns per CTA strip = one 128-byte swizzle row
constexpr int STAGE_BYTES = TOK * COLS * 2;  // 16 KB
constexpr int NS = 4;                        // stages in flight per CTA (3 CTAs per SM)
constexpr int CONSUMERS = 8;
constexpr int NT = (CONSUMERS + 1) * 32;
constexpr int SMEM_BYTES = NS * STAGE_BYTES + 1024;
constexpr int CTAS_PER_SM = 3;
}

// One reduction problem of a launch: G[c, j] = scale * sum_m Z[m, c] * P[m, j0 + j], j < r.
// Ranks [r_off[d], r_off[d+1]) go to out[d] (several destinations = the dA rows of fused siblings).
struct RedTcProb {
  const void* P;
  int ldp, j0, r;
  float scale;
  float* out[3];
  int r_off[4];
  long long out_cs, out_js;
  float* partial;          // deterministic mode: [splits, C, rp] workspace instead of atomics
  int C, col_blocks, block_start;
};
struct RedTcGroup {
  int n;
  RedTcProb p[4];
};

struct RedTcGeom {
  long long M;
  int rows_per_block;   // multiple of rtc::TOK
  int rp;               // rank stride of the deterministic workspace (4 / 8 / 16)
  // problem 0 as a convolution's dA: Z = im2col(X) [M, taps * cpb * 64], never materialised
  int conv, cv_S, cv_cpb, cv_C, cv_Ho, cv_Wo, cv_pad_h, cv_pad_w, cv_str_h, cv_str_w, cv_dil_h, cv_dil_w;
};

template <typename T, int NJ>
__global__ void __launch_bounds__(rtc::NT)
lora_grad_mma_kernel(const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1,
                     const __grid_constant__ CUtensorMap tm2, const __grid_constant__ CUtensorMap tm3,
                     const __grid_constant__ RedTcGroup grp, const RedTcGeom g) {
  extern __shared__ uint8_t red_smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[rtc::NS];
  __shared__ __align__(8) uint64_t empty_bar[rtc::NS];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(red_smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));

  int pid = 0;
#pragma unroll
  for (int i = 1; i < 4; ++i)
    if (i < grp.n && (int)blockIdx.x >= grp.p[i].block_start) pid = i;
  const RedTcProb& p = grp.p[pid];
  const CUtensorMap* tm = pid == 0 ? &tm0 : (pid == 1 ? &tm1 : (pid == 2 ? &tm2 : &tm3));
  const int bid = blockIdx.x - p.block_start;
  const int cb = bid % p.col_blocks;
  const int sp = bid / p.col_blocks;
  const long long m_begin = (long long)sp * g.rows_per_block;
  long long m_end = m_begin + g.rows_per_block;
  if (m_end > g.M) m_end = g.M;
  const int nstage = (int)((m_end - m_begin + rtc::TOK - 1) / rtc::TOK);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  pdl_launch_dependents();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < rtc::NS; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], rtc::CONSUMERS); }
    fence_mbar_init();
    tma_prefetch_desc(tm);
  }
  __syncthreads();
  pdl_wait();   // dY / X / T / dT come from the kernels before this one

  if (warp == rtc::CONSUMERS) {
    // ------------------------------ producer: one box per stage ------------------------------
    const bool im2col = pid == 0 && g.conv;
    const uint64_t pol = im2col ? policy_evict_normal() : policy_evict_first();   // taps re-read X
    const int tap = im2col ? cb / g.cv_cpb : 0;
    const int c0 = im2col ? (cb % g.cv_cpb) * rtc::COLS : cb * rtc::COLS;
    const uint16_t off_w = (uint16_t)((tap % (im2col ? g.cv_S : 1)) * g.cv_dil_w);
    const uint16_t off_h = (uint16_t)((tap / (im2col ? g.cv_S : 1)) * g.cv_dil_h);
    for (int it = 0; it < nstage; ++it) {
      const int s = it % rtc::NS;
      if (it >= rtc::NS) mbar_wait(&empty_bar[s], ((it / rtc::NS) - 1) & 1, 40);
      if (elect_one()) {
        const long long m0 = m_begin + (long long)it * rtc::TOK;
        mbar_arrive_expect_tx(&full_bar[s], rtc::STAGE_BYTES);
        if (im2col) {
          const int hw = g.cv_Ho * g.cv_Wo;
          const int n = (int)(m0 / hw);
          const int rem = (int)(m0 - (long long)n * hw);
          tma_load_im2col(smem + s * rtc::STAGE_BYTES, tm, &full_bar[s], c0,
                          (rem % g.cv_Wo) * g.cv_str_w - g.cv_pad_w,
                          (rem / g.cv_Wo) * g.cv_str_h - g.cv_pad_h, n, off_w, off_h, pol);
        } else {
          tma_load_2d(smem + s * rtc::STAGE_BYTES, tm, &full_bar[s], c0, (int)m0, pol);
        }
      }
      __syncwarp();
    }
    return;
  }

  // ------------------------------ consumers: 16 tokens per warp ------------------------------
  const int gq = lane >> 2, tq = lane & 3;
  const unsigned short* __restrict__ P = reinterpret_cast<const unsigned short*>(p.P) + p.j0;
  float acc[4][NJ][4];
#pragma unroll
  for (int ct = 0; ct < 4; ++ct)
#pragma unroll
    for (int jh = 0; jh < NJ; ++jh)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[ct][jh][i] = 0.f;

  // ldmatrix.trans row address of this lane: matrix mi = lane/8 -> (token half, column chunk parity)
  const int mi = lane >> 3, rr = lane & 7;
  const uint32_t row_off = (uint32_t)((warp * 16 + rr + ((mi >> 1) << 3)) * 128);
  const uint32_t smem_base = smem_u32(smem);

  uint32_t bnext[NJ][2];
#pragma unroll
  for (int jh = 0; jh < NJ; ++jh) load_p_frag(P, m_begin + warp * 16, gq + 8 * jh, p.r, p.ldp, g.M, tq, bnext[jh]);

  for (int it = 0; it < nstage; ++it) {
    const int s = it % rtc::NS;
    uint32_t bcur[NJ][2];
#pragma unroll
    for (int jh = 0; jh < NJ; ++jh) { bcur[jh][0] = bnext[jh][0]; bcur[jh][1] = bnext[jh][1]; }
    if (it + 1 < nstage) {
#pragma unroll
      for (int jh = 0; jh < NJ; ++jh)
        load_p_frag(P, m_begin + (long long)(it + 1) * rtc::TOK + warp * 16, gq + 8 * jh, p.r, p.ldp, g.M, tq,
                    bnext[jh]);
    }
    mbar_wait(&full_bar[s], (it / rtc::NS) & 1, 41);
    const uint32_t st_addr = smem_base + s * rtc::STAGE_BYTES + row_off;
#pragma unroll
    for (int ct = 0; ct < 4; ++ct) {
      uint32_t a[4];
      const uint32_t addr = st_addr + ((uint32_t)((2 * ct + (mi & 1)) ^ rr) << 4);
      asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
                   : "=r"(a[0]), "=r"(a[1]), "=r"(a[2]), "=r"(a[3])
                   : "r"(addr));
#pragma unroll
      for (int jh = 0; jh < NJ; ++jh) mma_16816<T>(acc[ct][jh], a, bcur[jh][0], bcur[jh][1]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[s]);
  }

  // ------------------------------ CTA reduction over the 8 token groups ----------------------
  constexpr int JW = 8 * NJ;
  float* red = reinterpret_cast<float*>(smem);   // [8 warps][64 cols][JW]; the ring is drained
  named_bar_sync(1, rtc::CONSUMERS * 32);
#pragma unroll
  for (int ct = 0; ct < 4; ++ct)
#pragma unroll
    for (int jh = 0; jh < NJ; ++jh) {
      float* dst = red + ((warp * rtc::COLS + ct * 16 + gq) * JW + jh * 8 + 2 * tq);
      *reinterpret_cast<float2*>(dst) = make_float2(acc[ct][jh][0], acc[ct][jh][1]);
      *reinterpret_cast<float2*>(dst + 8 * JW) = make_float2(acc[ct][jh][2], acc[ct][jh][3]);
    }
  named_bar_sync(1, rtc::CONSUMERS * 32);
  for (int idx = threadIdx.x; idx < rtc::COLS * JW; idx += rtc::CONSUMERS * 32) {
    const int c = idx / JW, j = idx % JW;
    int gc = cb * rtc::COLS + c;
    long long tap_off = 0;
    if (pid == 0 && g.conv) {          // column block = (tap, 64-channel block); dA is OIHW
      gc = (cb % g.cv_cpb) * rtc::COLS + c;
      tap_off = cb / g.cv_cpb;
    }
    if (gc >= p.C || j >= p.r) continue;
    float sum = 0.f;
#pragma unroll
    for (int w = 0; w < rtc::CONSUMERS; ++w) sum += red[(w * rtc::COLS + c) * JW + j];
    if (p.partial != nullptr) {
      p.partial[((long long)sp * p.C + gc) * g.rp + j] = sum;
    } else {
      const int d = (j >= p.r_off[1]) + (j >= p.r_off[2]);
      atomicAdd(p.out[d] + tap_off + gc * p.out_cs + (j - p.r_off[d]) * p.out_js, sum * p.scale);
    }
  }
}

// rows per CTA (multiple of 128) so that both problems together fill ~3 CTAs per SM
static int tc_rows_per_block(long long M, int C0, int C1, int sms) {
  const long long stages = (M + rtc::TOK - 1) / rtc::TOK;
  const long long cbs = (C0 + rtc::COLS - 1) / rtc::COLS + (C1 + rtc::COLS - 1) / rtc::COLS;
  const long long slots = (long long)sms * rtc::CTAS_PER_SM;
  long long spb = (cbs * stages + slots - 1) / slots;
  if (spb < 1) spb = 1;
  if (spb > stages) spb = stages;
  return (int)(spb * rtc::TOK);
}

static int pick_splits(long long M, int col_blocks) {
  long long target = (148 * 4 + col_blocks - 1) / col_blocks;
  long long max_splits = (M + 127) / 128;
  if (target > max_splits) target = max_splits;
  if (target < 1) target = 1;
  if (target > 1024) target = 1024;
  return (int)target;
}

// Row splits of the problem over C columns when it runs together with a problem over C_other
// columns (0 = alone).  Sizes the deterministic workspace; must match launch_lora_grads.
int grad_reduce_splits(long long M, int C, int C_other, int esize, int sms) {
  if (esize == 2) {
    const int rows = tc_rows_per_block(M, C, C_other, sms);
    return (int)((M + rows - 1) / rows);
  }
  const int CT = 8 * (16 / esize);
  return pick_splits(M, (C + CT - 1) / CT);
}

static void fill_problem(RedProblem& p, const void* Z, long long ldz, const void* P, float* out,
                         long long out_cs, long long out_js, float* partial, long long M, int C,
                         int esize) {
  const int CT = 8 * (16 / esize);
  p.Z = Z; p.ldz = ldz; p.P = P; p.out = out; p.out_cs = out_cs; p.out_js = out_js;
  p.partial = partial; p.C = C;
  p.col_blocks = (C + CT - 1) / CT;
  const int splits = pick_splits(M, p.col_blocks);
  long long rpb = (M + splits - 1) / splits;
  rpb = (rpb + 31) / 32 * 32;
  p.rows_per_block = rpb;
  p.splits = (int)((M + rpb - 1) / rpb);
}

// launch helper shared by the three entry points below
template <typename T>
static cudaError_t launch_mma_group(const CUtensorMap* const* tms, const RedTcGroup& grp, const RedTcGeom& g,
                                    int nblocks, int max_r, cudaStream_t stream) {
  const CUtensorMap& m0 = *tms[0];
  const CUtensorMap& m1 = *tms[grp.n > 1 ? 1 : 0];
  const CUtensorMap& m2 = *tms[grp.n > 2 ? 2 : 0];
  const CUtensorMap& m3 = *tms[grp.n > 3 ? 3 : 0];
  cudaError_t e = cudaSuccess;
#define LAUNCH_TC(NJ_)                                                                           \
  do {                                                                                           \
    auto kern = lora_grad_mma_kernel<T, NJ_>;                                                    \
    static std::once_flag once;                                                                  \
    static cudaError_t cap = cudaSuccess;                                                        \
    std::call_once(once, [&] {                                                                   \
      cap = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, rtc::SMEM_BYTES); \
    });                                                                                          \
    if (cap != cudaSuccess) return cap;                                                          \
    cudaLaunchConfig_t cfg = {};                                                                 \
    cudaLaunchAttribute attr[1];                                                                 \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                             \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                      \
    cfg.gridDim = dim3(nblocks, 1, 1);                                                           \
    cfg.blockDim = dim3(rtc::NT, 1, 1);                                                          \
    cfg.dynamicSmemBytes = rtc::SMEM_BYTES;                                                      \
    cfg.stream = stream;                                                                         \
    cfg.attrs = attr;                                                                            \
    cfg.numAttrs = pdl_enabled() ? 1 : 0;                                                        \
    e = cudaLaunchKernelEx(&cfg, kern, m0, m1, m2, m3, grp, g);                                  \
  } while (0)
  if (max_r <= 8) LAUNCH_TC(1); else LAUNCH_TC(2);
#undef LAUNCH_TC
  return e;
}

static void fill_tc_prob(RedTcProb& q, const void* P, int ldp, int j0, int r, float scale, float* out,
                         long long out_cs, long long out_js, float* partial, int C) {
  q.P = P; q.ldp = ldp; q.j0 = j0; q.r = r; q.scale = scale;
  q.out[0] = q.out[1] = q.out[2] = out;
  q.r_off[0] = 0; q.r_off[1] = q.r_off[2] = q.r_off[3] = r;
  q.out_cs = out_cs; q.out_js = out_js; q.partial = partial; q.C = C;
  q.col_blocks = (C + rtc::COLS - 1) / rtc::COLS;
}

template <typename T>
static cudaError_t launch_grads_tc(const CUtensorMap* tmDY, const CUtensorMap* tmX, const void* Tm,
                                   const void* dT, float scale, float* dA, float* dB, float* partA,
                                   float* partB, long long M, int K, int N, int r, int sms,
                                   cudaStream_t stream, int ldp = 0, int j0 = 0) {
  // ranks [j0, j0 + r) of T / dT whose rows hold ldp ranks (ldp = 0: compact, ldp = r); dA / dB point at the
  // FULL gradients [ldp, K] / [N, ldp]
  if (ldp == 0) ldp = r;
  if (dB) dB += j0;
  if (dA) dA += (long long)j0 * K;
  RedProblem pb = {}, pa = {};   // for the finalize kernel of the deterministic mode
  RedTcGroup grp = {};
  RedTcGeom g = {};
  g.M = M;
  g.rp = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
  g.rows_per_block = tc_rows_per_block(M, dB ? N : 0, dA ? K : 0, sms);
  const int splits = (int)((M + g.rows_per_block - 1) / g.rows_per_block);
  const CUtensorMap* tms[4] = {tmDY, tmX, tmDY, tmDY};
  int nblk = 0;
  if (dB) {
    RedTcProb& q = grp.p[grp.n];
    fill_tc_prob(q, Tm, ldp, j0, r, scale, dB, ldp, 1, partB, N);
    q.block_start = nblk; nblk += q.col_blocks * splits;
    tms[grp.n++] = tmDY;
    pb.out = dB; pb.out_cs = ldp; pb.out_js = 1; pb.partial = partB; pb.C = N; pb.splits = splits;
  }
  if (dA) {
    RedTcProb& q = grp.p[grp.n];
    fill_tc_prob(q, dT, ldp, j0, r, scale, dA, 1, K, partA, K);
    q.block_start = nblk; nblk += q.col_blocks * splits;
    tms[grp.n++] = tmX;
    pa.out = dA; pa.out_cs = 1; pa.out_js = K; pa.partial = partA; pa.C = K; pa.splits = splits;
  }
  cudaError_t e = launch_mma_group<T>(tms, grp, g, nblk, r, stream);
  if (e != cudaSuccess) return e;
  if ((partA && dA) || (partB && dB)) {
    const int n0 = dB ? N * g.rp : 0, n1 = dA ? K * g.rp : 0;
    const int blocks = (n0 + n1 + 255) / 256;
    if (g.rp == 4) lora_grad_finalize_kernel<4><<<blocks, 256, 0, stream>>>(pb, pa, n0, r, scale);
    else if (g.rp == 8) lora_grad_finalize_kernel<8><<<blocks, 256, 0, stream>>>(pb, pa, n0, r, scale);
    else lora_grad_finalize_kernel<16><<<blocks, 256, 0, stream>>>(pb, pa, n0, r, scale);
    e = cudaGetLastError();
  }
  return e;
}

// Fused sibling layers (same input X, outputs side by side): nseg problems dB_i <- (dY_i, T[:, off_i ..])
// and one problem dA_cat <- (X, dT) whose rank rows land in nseg destinations.  T / dT are [M, r_total].
template <typename T>
cudaError_t launch_grads_fused_tc(const CUtensorMap* const* tmDY, const CUtensorMap* tmX, const void* Tm,
                                  const void* dT, const float* scales, float* const* dA, float* const* dB,
                                  long long M, int K, int Nseg, int nseg, const int* r_each, int sms,
                                  cudaStream_t stream) {
  RedTcGroup grp = {};
  RedTcGeom g = {};
  g.M = M;
  int r_total = 0, r_max = 0;
  for (int i = 0; i < nseg; ++i) { r_total += r_each[i]; if (r_each[i] > r_max) r_max = r_each[i]; }
  g.rp = 16;
  g.rows_per_block = tc_rows_per_block(M, Nseg * nseg, K, sms);
  const int splits = (int)((M + g.rows_per_block - 1) / g.rows_per_block);
  const CUtensorMap* tms[4] = {tmX, tmX, tmX, tmX};
  int nblk = 0, off = 0;
  for (int i = 0; i < nseg; ++i) {       // dB_i: scale s_i, ranks [off, off + r_i) of T
    RedTcProb& q = grp.p[grp.n];
    fill_tc_prob(q, Tm, r_total, off, r_each[i], scales[i], dB[i], r_each[i], 1, nullptr, Nseg);
    q.block_start = nblk; nblk += q.col_blocks * splits;
    tms[grp.n++] = tmDY[i];
    off += r_each[i];
  }
  {                                      // dA_cat: dT already carries s_i (B was folded with it)
    RedTcProb& q = grp.p[grp.n];
    fill_tc_prob(q, dT, r_total, 0, r_total, 1.0f, dA[0], 1, K, nullptr, K);
    int o = 0;
    for (int i = 0; i < 3; ++i) {
      q.out[i] = dA[i < nseg ? i : nseg - 1];
      q.r_off[i] = o;
      o += i < nseg ? r_each[i] : 0;
    }
    q.r_off[3] = r_total;
    for (int i = nseg; i < 3; ++i) q.r_off[i] = r_total;
    q.block_start = nblk; nblk += q.col_blocks * splits;
    tms[grp.n++] = tmX;
  }
  return launch_mma_group<T>(tms, grp, g, nblk, r_total, stream);
}
template cudaError_t launch_grads_fused_tc<__nv_bfloat16>(const CUtensorMap* const*, const CUtensorMap*,
                                                          const void*, const void*, const float*, float* const*,
                                                          float* const*, long long, int, int, int, const int*,
                                                          int, cudaStream_t);
template cudaError_t launch_grads_fused_tc<__half>(const CUtensorMap* const*, const CUtensorMap*, const void*,
                                                   const void*, const float*, float* const*, float* const*,
                                                   long long, int, int, int, const int*, int, cudaStream_t);

// dA of a LoRA convolution: dA[j, c, kh, kw] += s * sum_m dT[m, j] * X[pixel(m) + tap, c], with the
// im2col view of X streamed by TMA (tmX: the forward's im2col map).  geo = {N,H,W,C,Ho,Wo,R,S,
// str_h,str_w,pad_h,pad_w,dil_h,dil_w}; dA is OIHW fp32 (accumulated with atomics).
template <typename T>
cudaError_t launch_conv_grad_a_tc(const CUtensorMap* tmX, const void* dT, float scale, float* dA,
                                  const int* geo, int r, int sms, cudaStream_t stream) {
  const int N = geo[0], C = geo[3], Ho = geo[4], Wo = geo[5], R = geo[6], S = geo[7];
  const long long M = (long long)N * Ho * Wo;
  RedTcGroup grp = {};
  RedTcGeom g = {};
  g.M = M; g.rp = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
  g.conv = 1; g.cv_S = S; g.cv_cpb = (C + rtc::COLS - 1) / rtc::COLS; g.cv_C = C; g.cv_Ho = Ho; g.cv_Wo = Wo;
  g.cv_str_h = geo[8]; g.cv_str_w = geo[9]; g.cv_pad_h = geo[10]; g.cv_pad_w = geo[11];
  g.cv_dil_h = geo[12]; g.cv_dil_w = geo[13];
  const int col_blocks = R * S * g.cv_cpb;
  g.rows_per_block = tc_rows_per_block(M, col_blocks * rtc::COLS, 0, sms);
  const int splits = (int)((M + g.rows_per_block - 1) / g.rows_per_block);
  RedTcProb& q = grp.p[0];
  fill_tc_prob(q, dT, r, 0, r, scale, dA, (long long)R * S, (long long)C * R * S, nullptr, C);
  q.col_blocks = col_blocks;
  q.block_start = 0;
  grp.n = 1;
  const CUtensorMap* tms[4] = {tmX, tmX, tmX, tmX};
  return launch_mma_group<T>(tms, grp, g, col_blocks * splits, r, stream);
}
template cudaError_t launch_conv_grad_a_tc<__nv_bfloat16>(const CUtensorMap*, const void*, float, float*,
                                                          const int*, int, int, cudaStream_t);
template cudaError_t launch_conv_grad_a_tc<__half>(const CUtensorMap*, const void*, float, float*, const int*,
                                                   int, int, cudaStream_t);

// Both gradient reductions in one launch.  dB <- (dY, T), dA <- (X, dT); either may be skipped.
// tmDY / tmX: [M,N] / [M,K] tensor maps with a [128 x 64] box (16-bit io); null selects the SIMT
// kernel (fp32 io).
template <typename T>
cudaError_t launch_lora_grads(const void* dY, const void* X, const void* Tm, const void* dT,
                              float scale, float* dA, float* dB, float* partA, float* partB,
                              long long M, int K, int N, int r, const CUtensorMap* tmDY,
                              const CUtensorMap* tmX, int sms, cudaStream_t stream) {
  // Ranks above 16 (the reference accepts any r <= min(in, out); --lora_rank 32 / 64 are common) run as
  // windows of 16 ranks: one launch per window over the same Z, each adding into its own rows of dA /
  // columns of dB.  (The deterministic workspace is sized for 16 ranks and reused: the launches are
  // stream-ordered.)
  for (int j0 = 0; j0 < r; j0 += 16) {
    const int rc = (r - j0) < 16 ? (r - j0) : 16;
    if constexpr (sizeof(T) == 2) {
      if (tmDY != nullptr && tmX != nullptr) {
        cudaError_t e = launch_grads_tc<T>(tmDY, tmX, Tm, dT, scale, dA, dB, partA, partB, M, K, N, rc, sms, stream,
                                           r, j0);
        if (e != cudaSuccess) return e;
        continue;
      }
    }
    const T* Tw = Tm ? reinterpret_cast<const T*>(Tm) + j0 : nullptr;
    const T* dTw = dT ? reinterpret_cast<const T*>(dT) + j0 : nullptr;
    float* dBw = dB ? dB + j0 : nullptr;
    float* dAw = dA ? dA + (long long)j0 * K : nullptr;
    RedProblem pb = {}, pa = {};
    int nb = 0, na = 0;
    if (dB) { fill_problem(pb, dY, N, Tw, dBw, /*cs=*/r, /*js=*/1, partB, M, N, sizeof(T)); nb = pb.col_blocks * pb.splits; }
    if (dA) { fill_problem(pa, X, K, dTw, dAw, /*cs=*/1, /*js=*/K, partA, M, K, sizeof(T)); na = pa.col_blocks * pa.splits; }
    if (nb + na == 0) return cudaSuccess;
    const int RPv = rc <= 4 ? 4 : (rc <= 8 ? 8 : 16);
#define LAUNCH(RP_)                                                                              \
  do {                                                                                           \
    lora_grad_reduce_kernel<T, RP_><<<nb + na, red::NT, 0, stream>>>(pb, pa, nb, rc, r, scale, M);  \
    if ((partA && dA) || (partB && dB)) {                                                        \
      const int n0 = dB ? N * RP_ : 0, n1 = dA ? K * RP_ : 0;                                     \
      lora_grad_finalize_kernel<RP_><<<(n0 + n1 + 255) / 256, 256, 0, stream>>>(pb, pa, n0, rc, scale); \
    }                                                                                            \
  } while (0)
    if (RPv == 4) LAUNCH(4); else if (RPv == 8) LAUNCH(8); else LAUNCH(16);
#undef LAUNCH
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

template cudaError_t launch_lora_grads<float>(const void*, const void*, const void*, const void*,
                                              float, float*, float*, float*, float*, long long,
                                              int, int, int, const CUtensorMap*, const CUtensorMap*,
                                              int, cudaStream_t);
template cudaError_t launch_lora_grads<__half>(const void*, const void*, const void*, const void*,
                                               float, float*, float*, float*, float*, long long,
                                               int, int, int, const CUtensorMap*, const CUtensorMap*,
                                               int, cudaStream_t);
template cudaError_t launch_lora_grads<__nv_bfloat16>(const void*, const void*, const void*,
                                                      const void*, float, float*, float*, float*,
                                                      float*, long long, int, int, int,
                                                      const CUtensorMap*, const CUtensorMap*, int,
                                                      cudaStream_t);

}  // namespace b200
