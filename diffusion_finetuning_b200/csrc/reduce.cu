// Token reductions for the LoRA gradients (HBM-bound):
//   dB[n,j] += s · Σ_m dY[m,n]·T[m,j]       dA[j,k] += s · Σ_m dT[m,j]·X[m,k]
// Both are the same shape of work:  G[c,j] = s · Σ_m Z[m,c]·P[m,j]  with Z = dY or X
// ([M,C], row-major) and P = T or dT ([M,r] compact).  Each block owns a strip of CT
// columns and a slice of rows; 8 lanes x 16-byte loads cover one 128-byte row segment and
// 4 rows are read per warp instruction, so every global load is a full 128 B line.  Partial
// sums go lane -> warp (shuffle) -> block (smem) -> either fp32 atomics into the gradient
// accumulator (default) or a [splits, C, RP] workspace reduced by a second kernel
// (deterministic mode).
#include "common.cuh"

namespace b200 {

namespace red {
constexpr int NT = 256;
}

template <typename T, int RP>
__global__ void __launch_bounds__(red::NT)
lora_grad_reduce_kernel(const T* __restrict__ Z, long long ldz, const T* __restrict__ P, int r,
                        float scale, float* __restrict__ out, long long out_cs, long long out_js,
                        float* __restrict__ partial, long long M, int C, long long rows_per_block) {
  constexpr int VEC = 16 / sizeof(T);   // columns per lane
  constexpr int CT = 8 * VEC;           // columns per block
  __shared__ float red_s[8][CT][RP + 1];

  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const int cl = lane & 7;              // column lane
  const int rl = (warp << 2) | (lane >> 3);  // row lane 0..31
  const int c0 = blockIdx.x * CT + cl * VEC;
  const long long m_begin = (long long)blockIdx.y * rows_per_block;
  long long m_end = m_begin + rows_per_block;
  if (m_end > M) m_end = M;

  float acc[VEC][RP];
#pragma unroll
  for (int v = 0; v < VEC; ++v)
#pragma unroll
    for (int j = 0; j < RP; ++j) acc[v][j] = 0.f;

  const bool col_ok = c0 < C;  // C % VEC == 0 is guaranteed by the host
  if (col_ok) {
    for (long long m = m_begin + rl; m < m_end; m += 32) {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(Z + m * ldz + c0));
      const T* zv = reinterpret_cast<const T*>(&raw);
      float pv[RP];
#pragma unroll
      for (int j = 0; j < RP; ++j) pv[j] = (j < r) ? to_f32(P[m * r + j]) : 0.f;
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
    const int gc = blockIdx.x * CT + c;
    if (gc >= C || j >= r) continue;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red_s[w][c][j];
    if (partial != nullptr) {
      partial[((long long)blockIdx.y * C + gc) * RP + j] = s;
    } else {
      atomicAdd(out + gc * out_cs + j * out_js, s * scale);
    }
  }
}

template <int RP>
__global__ void lora_grad_finalize_kernel(const float* __restrict__ partial, int splits, int C,
                                          int r, float scale, float* __restrict__ out,
                                          long long out_cs, long long out_js) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= C * RP) return;
  const int c = idx / RP, j = idx % RP;
  if (j >= r) return;
  float s = 0.f;
  for (int y = 0; y < splits; ++y) s += partial[((long long)y * C + c) * RP + j];
  out[c * out_cs + j * out_js] += s * scale;
}

static int pick_splits(long long M, int col_blocks) {
  long long target = (148 * 4 + col_blocks - 1) / col_blocks;
  long long max_splits = (M + 63) / 64;
  if (target > max_splits) target = max_splits;
  if (target < 1) target = 1;
  if (target > 1024) target = 1024;
  return (int)target;
}

int grad_reduce_splits(long long M, int C, int esize) {
  const int CT = 8 * (16 / esize);
  return pick_splits(M, (C + CT - 1) / CT);
}

template <typename T, int RP>
static cudaError_t launch_one(const T* Z, long long ldz, const T* P, int r, float scale,
                              float* out, long long out_cs, long long out_js, float* partial,
                              long long M, int C, cudaStream_t stream) {
  constexpr int VEC = 16 / sizeof(T);
  constexpr int CT = 8 * VEC;
  const int col_blocks = (C + CT - 1) / CT;
  const int splits = pick_splits(M, col_blocks);
  long long rpb = (M + splits - 1) / splits;
  rpb = (rpb + 31) / 32 * 32;
  dim3 grid(col_blocks, (unsigned)((M + rpb - 1) / rpb));
  lora_grad_reduce_kernel<T, RP><<<grid, red::NT, 0, stream>>>(Z, ldz, P, r, scale, out, out_cs,
                                                             out_js, partial, M, C, rpb);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  if (partial != nullptr) {
    const int n = C * RP;
    lora_grad_finalize_kernel<RP><<<(n + 255) / 256, 256, 0, stream>>>(partial, (int)grid.y, C, r,
                                                                       scale, out, out_cs, out_js);
    e = cudaGetLastError();
  }
  return e;
}

// G[c,j] (+)= s·Σ_m Z[m,c]·P[m,j]; out element (c,j) at out[c*out_cs + j*out_js]
template <typename T>
cudaError_t launch_grad_reduce(const void* Z, long long ldz, const void* P, int r, float scale,
                               float* out, long long out_cs, long long out_js, float* partial,
                               long long M, int C, cudaStream_t stream) {
  const T* z = (const T*)Z;
  const T* pp = (const T*)P;
  if (r <= 4) return launch_one<T, 4>(z, ldz, pp, r, scale, out, out_cs, out_js, partial, M, C, stream);
  if (r <= 8) return launch_one<T, 8>(z, ldz, pp, r, scale, out, out_cs, out_js, partial, M, C, stream);
  return launch_one<T, 16>(z, ldz, pp, r, scale, out, out_cs, out_js, partial, M, C, stream);
}

template cudaError_t launch_grad_reduce<float>(const void*, long long, const void*, int, float,
                                               float*, long long, long long, float*, long long,
                                               int, cudaStream_t);
template cudaError_t launch_grad_reduce<__half>(const void*, long long, const void*, int, float,
                                                float*, long long, long long, float*, long long,
                                                int, cudaStream_t);
template cudaError_t launch_grad_reduce<__nv_bfloat16>(const void*, long long, const void*, int,
                                                       float, float*, long long, long long, float*,
                                                       long long, int, cudaStream_t);

}  // namespace b200
