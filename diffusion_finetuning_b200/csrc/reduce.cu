// Token reductions for the LoRA gradients (HBM-bound):
//   dB[n,j] += s · Σ_m dY[m,n]·T[m,j]       dA[j,k] += s · Σ_m dT[m,j]·X[m,k]
// Both are the same shape of work:  G[c,j] = s · Σ_m Z[m,c]·P[m,j]  with Z = dY or X
// ([M,C], row-major) and P = T or dT ([M,r] compact), so ONE launch covers both: the first
// `blocks0` CTAs work on problem 0, the rest on problem 1.  Each CTA owns a strip of CT columns
// and a slice of rows; 8 lanes x 16-byte loads cover one 128-byte row segment, 4 rows per warp
// instruction, and 4 such loads are kept in flight per thread.  Partial sums go lane -> warp
// (shuffle) -> CTA (smem) -> either fp32 atomics into the gradient accumulator (default) or a
// [splits, C, RP] workspace reduced by a second kernel (deterministic mode).
#include "common.cuh"

namespace b200 {

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
__device__ __forceinline__ void load_p(const T* __restrict__ P, long long m, int r, float (&pv)[RP]) {
  if (r == RP) {  // compact row of exactly RP elements: vector loads (row stride = RP*sizeof(T))
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
    for (int j = 0; j < RP; ++j) pv[j] = (j < r) ? to_f32(P[m * r + j]) : 0.f;
  }
}

template <typename T, int RP>
__global__ void __launch_bounds__(red::NT)
lora_grad_reduce_kernel(const RedProblem p0, const RedProblem p1, int blocks0, int r, float scale,
                        long long M) {
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
      for (int u = 0; u < red::UNROLL; ++u) load_p<T, RP>(P, m + 32 * u, r, pv[u]);
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
      load_p<T, RP>(P, m, r, pv);
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

static int pick_splits(long long M, int col_blocks) {
  long long target = (148 * 4 + col_blocks - 1) / col_blocks;
  long long max_splits = (M + 127) / 128;
  if (target > max_splits) target = max_splits;
  if (target < 1) target = 1;
  if (target > 1024) target = 1024;
  return (int)target;
}

int grad_reduce_splits(long long M, int C, int esize) {
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

// Both gradient reductions in one launch.  dB <- (dY, T), dA <- (X, dT); either may be skipped.
template <typename T>
cudaError_t launch_lora_grads(const void* dY, const void* X, const void* Tm, const void* dT,
                              float scale, float* dA, float* dB, float* partA, float* partB,
                              long long M, int K, int N, int r, cudaStream_t stream) {
  RedProblem pb = {}, pa = {};
  int nb = 0, na = 0;
  if (dB) { fill_problem(pb, dY, N, Tm, dB, /*cs=*/r, /*js=*/1, partB, M, N, sizeof(T)); nb = pb.col_blocks * pb.splits; }
  if (dA) { fill_problem(pa, X, K, dT, dA, /*cs=*/1, /*js=*/K, partA, M, K, sizeof(T)); na = pa.col_blocks * pa.splits; }
  if (nb + na == 0) return cudaSuccess;
  const int RPv = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
#define LAUNCH(RP_)                                                                              \
  do {                                                                                           \
    lora_grad_reduce_kernel<T, RP_><<<nb + na, red::NT, 0, stream>>>(pb, pa, nb, r, scale, M);  \
    if ((partA && dA) || (partB && dB)) {                                                        \
      const int n0 = dB ? N * RP_ : 0, n1 = dA ? K * RP_ : 0;                                     \
      lora_grad_finalize_kernel<RP_><<<(n0 + n1 + 255) / 256, 256, 0, stream>>>(pb, pa, n0, r, scale); \
    }                                                                                            \
  } while (0)
  if (RPv == 4) LAUNCH(4); else if (RPv == 8) LAUNCH(8); else LAUNCH(16);
#undef LAUNCH
  return cudaGetLastError();
}

template cudaError_t launch_lora_grads<float>(const void*, const void*, const void*, const void*,
                                              float, float*, float*, float*, float*, long long,
                                              int, int, int, cudaStream_t);
template cudaError_t launch_lora_grads<__half>(const void*, const void*, const void*, const void*,
                                               float, float*, float*, float*, float*, long long,
                                               int, int, int, cudaStream_t);
template cudaError_t launch_lora_grads<__nv_bfloat16>(const void*, const void*, const void*,
                                                      const void*, float, float*, float*, float*,
                                                      float*, long long, int, int, int,
                                                      cudaStream_t);

}  // namespace b200
