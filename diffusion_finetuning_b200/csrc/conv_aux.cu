// Helpers of the LoRA-Conv2d implicit GEMM (no counterpart in the reference
// snapshot: upstream cloneofsimo/lora `LoraInjectedConv2d`).  The GEMM itself is the CTA-pair kernel
// of linear_tc2.cu with TMA-im2col loads; this file holds the bandwidth-bound side kernels:
//   repack   OIHW weights -> [rows, taps x Cp] GEMM operand (forward: KRSC; backward: flipped/transposed)
//   pack     lora_down / lora_up / bias -> rank rows, tail block, and the A-columns of the backward operand
//   zcat     backward: Z = [dY | s·dT | 0] along channels and dT = dY·B (one pass over dY)
//   dA       backward: dA[j,c,kh,kw] += s · Σ_pixels dT[m,j] · X[pixel shifted by (kh,kw), c]
#include <string.h>

#include "common.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------------
// weights: W [Cout, Cin, R, S] (OIHW, any dtype) -> GEMM B operand in the io dtype
//   mode 0 (forward):   Wk [Cout, (kh,kw), Cp]   Wk[o][tap][c] = W[o][c][kh][kw]          (c >= Cin: 0)
//   mode 1 (backward):  Wb [Cin, (a,b), Cz]      Wb[c][tap'][o] = W[o][c][R-1-a][S-1-b]   (o >= Cout: 0)
// ---------------------------------------------------------------------------------------------
template <typename TIO>
__global__ void conv_repack_kernel(const void* __restrict__ W, int w_dtype, TIO* __restrict__ out,
                                   int Cout, int Cin, int R, int S, int Cp, int mode) {
  const int taps = R * S;
  const long long rows = mode == 0 ? Cout : Cin;
  const long long total = rows * taps * Cp;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int ch = (int)(i % Cp);
    const int tap = (int)((i / Cp) % taps);
    const int row = (int)(i / ((long long)Cp * taps));
    float v = 0.f;
    if (mode == 0) {
      if (ch < Cin)
        v = load_as_float(W, w_dtype, (((long long)row * Cin + ch) * R + tap / S) * S + tap % S);
    } else {
      const int a = tap / S, b = tap % S;
      if (ch < Cout)
        v = load_as_float(W, w_dtype, (((long long)ch * Cin + row) * R + (R - 1 - a)) * S + (S - 1 - b));
    }
    out[i] = from_f32<TIO>(v);
  }
}

cudaError_t launch_conv_repack(const void* W, int w_dtype, void* out, int Cout, int Cin, int R, int S,
                               int Cp, int mode, int io_dtype, cudaStream_t stream) {
  const long long total = (long long)(mode == 0 ? Cout : Cin) * R * S * Cp;
  int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  if (blocks < 1) blocks = 1;
  if (io_dtype == B200LORA_BF16)
    conv_repack_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(W, w_dtype, (__nv_bfloat16*)out, Cout,
                                                                  Cin, R, S, Cp, mode);
  else
    conv_repack_kernel<__half><<<blocks, 256, 0, stream>>>(W, w_dtype, (__half*)out, Cout, Cin, R, S,
                                                           Cp, mode);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// LoRA pack for the conv:  A [r, Cin, R, S], B [Cout, r], bias [Cout]
//   A_pack  [16, taps x Cp]    A_pack[j][tap][c] = A[j][c][kh][kw]                 (forward rank rows)
//   Bn_pack [Cout, 64]         B | bias in column r | 0                            (forward tail block)
//   Wb A-columns (optional)    Wb[c][tap'][Coutp + j] = A[j][c][R-1-a][S-1-b], j < 64 (0 for j >= r)
// ---------------------------------------------------------------------------------------------
template <typename TIO>
__global__ void conv_pack_kernel(const void* __restrict__ A, const void* __restrict__ B, int ab_dtype,
                                 const void* __restrict__ bias, int bias_dtype, TIO* __restrict__ A_pack,
                                 TIO* __restrict__ Bn_pack, TIO* __restrict__ Wb, int Cout, int Cin,
                                 int R, int S, int Cp, int Coutp, int Cz, int r) {
  const int taps = R * S;
  const long long n0 = 16ll * taps * Cp, n1 = n0 + 64ll * Cout;
  const long long n2 = n1 + (Wb ? (long long)Cin * taps * 64 : 0);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2;
       i += (long long)gridDim.x * blockDim.x) {
    if (i < n0) {
      const int c = (int)(i % Cp), tap = (int)((i / Cp) % taps), j = (int)(i / ((long long)Cp * taps));
      float v = 0.f;
      if (j < r && c < Cin)
        v = load_as_float(A, ab_dtype, (((long long)j * Cin + c) * R + tap / S) * S + tap % S);
      A_pack[i] = from_f32<TIO>(v);
    } else if (i < n1) {
      const long long t = i - n0;
      const int o = (int)(t >> 6), j = (int)(t & 63);
      float v = 0.f;
      if (j < r) v = load_as_float(B, ab_dtype, (long long)o * r + j);
      else if (j == r && bias != nullptr) v = load_as_float(bias, bias_dtype, o);
      Bn_pack[t] = from_f32<TIO>(v);
    } else {
      const long long t = i - n1;
      const int j = (int)(t & 63), tap = (int)((t >> 6) % taps), c = (int)(t / (64ll * taps));
      const int a = tap / S, b = tap % S;
      float v = 0.f;
      if (j < r)
        v = load_as_float(A, ab_dtype, (((long long)j * Cin + c) * R + (R - 1 - a)) * S + (S - 1 - b));
      Wb[((long long)c * taps + tap) * Cz + Coutp + j] = from_f32<TIO>(v);
    }
  }
}

cudaError_t launch_conv_pack(const void* A, const void* B, int ab_dtype, const void* bias,
                             int bias_dtype, void* A_pack, void* Bn_pack, void* Wb, int Cout, int Cin,
                             int R, int S, int Cp, int Coutp, int Cz, int r, int io_dtype,
                             cudaStream_t stream) {
  const long long total = 16ll * R * S * Cp + 64ll * Cout + (Wb ? (long long)Cin * R * S * 64 : 0);
  int blocks = (int)((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
  if (io_dtype == B200LORA_BF16)
    conv_pack_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(
        A, B, ab_dtype, bias, bias_dtype, (__nv_bfloat16*)A_pack, (__nv_bfloat16*)Bn_pack,
        (__nv_bfloat16*)Wb, Cout, Cin, R, S, Cp, Coutp, Cz, r);
  else
    conv_pack_kernel<__half><<<blocks, 256, 0, stream>>>(A, B, ab_dtype, bias, bias_dtype,
                                                         (__half*)A_pack, (__half*)Bn_pack, (__half*)Wb,
                                                         Cout, Cin, R, S, Cp, Coutp, Cz, r);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// backward, one pass over dY [M, Cout]:  dT = dY·B  (written compact [M, r]) and
// Z [M, Cz] = [dY | 0 .. Coutp | s·dT | 0 .. 64]   — the channel-concatenated input of the dX conv
// ---------------------------------------------------------------------------------------------
template <typename TIO, int RP>
__global__ void __launch_bounds__(256)
conv_zcat_kernel(const TIO* __restrict__ dY, const void* __restrict__ B, int ab_dtype, float scale,
                 TIO* __restrict__ Z, TIO* __restrict__ dT, long long M, int Cout, int Coutp, int Cz,
                 int r) {
  extern __shared__ float Bs[];   // [Cout][RP], rounded through the io dtype as autocast would
  for (int i = threadIdx.x; i < Cout * RP; i += blockDim.x) {
    const int c = i / RP, j = i % RP;
    Bs[i] = j < r ? to_f32(from_f32<TIO>(load_as_float(B, ab_dtype, (long long)c * r + j))) : 0.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = Cout >> 3;           // Cout % 8 == 0 (host-checked)
  for (long long m = (long long)blockIdx.x * 8 + warp; m < M; m += (long long)gridDim.x * 8) {
    float acc[RP];
#pragma unroll
    for (int j = 0; j < RP; ++j) acc[j] = 0.f;
    const uint4* src = reinterpret_cast<const uint4*>(dY + m * Cout);
    uint4* dst = reinterpret_cast<uint4*>(Z + m * Cz);
    for (int v = lane; v < nvec; v += 32) {
      const uint4 raw = __ldg(src + v);
      dst[v] = raw;
      const TIO* e = reinterpret_cast<const TIO*>(&raw);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float y = to_f32(e[k]);
        const float* brow = Bs + (v * 8 + k) * RP;
#pragma unroll
        for (int j = 0; j < RP; ++j) acc[j] = fmaf(y, brow[j], acc[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < RP; ++j) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
    }
    // zero padding between Cout and Coutp, then the 64-channel LoRA block
    TIO* zrow = Z + m * Cz;
    for (int c = Cout + lane; c < Coutp; c += 32) zrow[c] = from_f32<TIO>(0.f);
    for (int j = lane; j < 64; j += 32) {
      float v = 0.f;
#pragma unroll
      for (int q = 0; q < RP; ++q)
        if (q == j) v = acc[q];
      const TIO t = from_f32<TIO>(v);                 // dT rounded to the io dtype (reference: bf16 matmul)
      if (j < r) dT[m * r + j] = t;
      zrow[Coutp + j] = from_f32<TIO>(j < r ? to_f32(t) * scale : 0.f);
    }
  }
}

template <typename TIO>
static cudaError_t zcat_dispatch(const void* dY, const void* B, int ab_dtype, float scale, void* Z,
                                 void* dT, long long M, int Cout, int Coutp, int Cz, int r,
                                 cudaStream_t stream) {
  const int RP = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
  const size_t smem = (size_t)Cout * RP * sizeof(float);
  long long want = (M + 7) / 8;
  const int blocks = (int)(want < 148 * 8 ? want : 148 * 8);
#define ZC(RP_)                                                                                    \
  do {                                                                                             \
    auto k = conv_zcat_kernel<TIO, RP_>;                                                           \
    if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    k<<<blocks, 256, smem, stream>>>((const TIO*)dY, B, ab_dtype, scale, (TIO*)Z, (TIO*)dT, M, Cout, \
                                     Coutp, Cz, r);                                                \
  } while (0)
  if (RP == 4) ZC(4); else if (RP == 8) ZC(8); else ZC(16);
#undef ZC
  return cudaGetLastError();
}

cudaError_t launch_conv_zcat(const void* dY, const void* B, int ab_dtype, float scale, void* Z, void* dT,
                             long long M, int Cout, int Coutp, int Cz, int r, int io_dtype,
                             cudaStream_t stream) {
  return io_dtype == B200LORA_BF16
             ? zcat_dispatch<__nv_bfloat16>(dY, B, ab_dtype, scale, Z, dT, M, Cout, Coutp, Cz, r, stream)
             : zcat_dispatch<__half>(dY, B, ab_dtype, scale, Z, dT, M, Cout, Coutp, Cz, r, stream);
}

// ---------------------------------------------------------------------------------------------
// backward: dA[j, c, kh, kw] += s · Σ_m dT[m, j] · X[n, oh·st + kh·d − p, ow·st + kw·d − p, c]
// grid = (64-channel blocks, row splits, taps); fp32 atomics into dA (OIHW, e.g. an arena slice)
// ---------------------------------------------------------------------------------------------
struct ConvGeo {
  int N, H, W, C, Ho, Wo, R, S, str_h, str_w, pad_h, pad_w, dil_h, dil_w;
};

template <typename TIO, int RP>
__global__ void __launch_bounds__(256)
conv_dA_kernel(const TIO* __restrict__ X, const TIO* __restrict__ dT, float scale, float* __restrict__ dA,
               const ConvGeo g, int r, long long M, long long rows_per_block) {
  __shared__ float red[8][64][RP + 1];
  const int tap = blockIdx.z, kh = tap / g.S, kw = tap % g.S;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cl = lane & 7, rl = (warp << 2) | (lane >> 3);     // 8 lanes x 8 channels, 32 row lanes
  const int c0 = blockIdx.x * 64 + cl * 8;
  long long m_end = ((long long)blockIdx.y + 1) * rows_per_block;
  if (m_end > M) m_end = M;
  float acc[8][RP];
#pragma unroll
  for (int v = 0; v < 8; ++v)
#pragma unroll
    for (int j = 0; j < RP; ++j) acc[v][j] = 0.f;
  if (c0 < g.C) {   // C % 8 == 0
    const int hw = g.Ho * g.Wo;
    for (long long m = (long long)blockIdx.y * rows_per_block + rl; m < m_end; m += 32) {
      const int n = (int)(m / hw);
      const int rem = (int)(m - (long long)n * hw);
      const int ih = (rem / g.Wo) * g.str_h + kh * g.dil_h - g.pad_h;
      const int iw = (rem % g.Wo) * g.str_w + kw * g.dil_w - g.pad_w;
      if (ih < 0 || ih >= g.H || iw < 0 || iw >= g.W) continue;
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(
          X + (((long long)n * g.H + ih) * g.W + iw) * g.C + c0));
      float pv[RP];
#pragma unroll
      for (int j = 0; j < RP; ++j) pv[j] = j < r ? to_f32(dT[m * r + j]) : 0.f;
      const TIO* xv = reinterpret_cast<const TIO*>(&raw);
#pragma unroll
      for (int v = 0; v < 8; ++v) {
        const float x = to_f32(xv[v]);
#pragma unroll
        for (int j = 0; j < RP; ++j) acc[v][j] = fmaf(x, pv[j], acc[v][j]);
      }
    }
  }
#pragma unroll
  for (int v = 0; v < 8; ++v)
#pragma unroll
    for (int j = 0; j < RP; ++j) {
      float a = acc[v][j];
      a += __shfl_xor_sync(0xffffffffu, a, 8);
      a += __shfl_xor_sync(0xffffffffu, a, 16);
      acc[v][j] = a;
    }
  if (lane < 8) {
#pragma unroll
    for (int v = 0; v < 8; ++v)
#pragma unroll
      for (int j = 0; j < RP; ++j) red[warp][cl * 8 + v][j] = acc[v][j];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 64 * RP; idx += 256) {
    const int c = idx / RP, j = idx % RP;
    const int gc = blockIdx.x * 64 + c;
    if (gc >= g.C || j >= r) continue;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += red[w][c][j];
    atomicAdd(dA + (((long long)j * g.C + gc) * g.R + kh) * g.S + kw, s * scale);
  }
}

cudaError_t launch_conv_dA(const void* X, const void* dT, float scale, float* dA, const int* geo14, int r,
                           int io_dtype, cudaStream_t stream) {
  ConvGeo g;
  memcpy(&g, geo14, sizeof(g));
  const long long M = (long long)g.N * g.Ho * g.Wo;
  const int cblocks = (g.C + 63) / 64, taps = g.R * g.S;
  long long splits = (148 * 4 + cblocks * taps - 1) / (cblocks * taps);
  const long long maxs = (M + 255) / 256;
  if (splits > maxs) splits = maxs;
  if (splits < 1) splits = 1;
  long long rpb = (M + splits - 1) / splits;
  rpb = (rpb + 31) / 32 * 32;
  splits = (M + rpb - 1) / rpb;
  dim3 grid(cblocks, (unsigned)splits, taps);
  const int RP = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
#define DA(T_, RP_) conv_dA_kernel<T_, RP_><<<grid, 256, 0, stream>>>((const T_*)X, (const T_*)dT, scale, dA, g, r, M, rpb)
  if (io_dtype == B200LORA_BF16) { if (RP == 4) DA(__nv_bfloat16, 4); else if (RP == 8) DA(__nv_bfloat16, 8); else DA(__nv_bfloat16, 16); }
  else { if (RP == 4) DA(__half, 4); else if (RP == 8) DA(__half, 8); else DA(__half, 16); }
#undef DA
  return cudaGetLastError();
}

}  // namespace b200
