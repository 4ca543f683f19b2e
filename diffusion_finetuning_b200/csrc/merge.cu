// Bandwidth-bound utility kernels around the LoRA weights:
//   merge        W <- W + alpha * (up @ down)            (lora.py:423, cli_lora_add.py:78)
//   absmean      Σ |up @ down|                            (lora.py:744-746, inspect_lora)
//   svd epilogue U <- clamp(U·diag(S)), Vh <- clamp(Vh)   (cli_svd.py:73-84)
// Each thread owns one 16-byte vector of a W row; the r x VEC slice of `down` it needs stays
// in registers while it walks down a strip of rows, so per 16 B of W moved it issues one
// 16 B load, one 16 B store and one broadcast load of up[n, 0:r].
#include "ptx.cuh"
#include "common.cuh"

namespace b200 {

namespace mrg {
constexpr int TX = 32;        // 16-byte vectors along K per block (512 contiguous bytes per row)
constexpr int TY = 8;         // row lanes
constexpr int NT = TX * TY;
constexpr int UNROLL = 4;     // rows in flight per thread
}

// rounding of the rank-r product follows the reference: (up @ down) is produced in the LoRA
// dtype, cast to the weight dtype, scaled by alpha (rounded), then added (rounded).
template <typename TW, typename TL>
__device__ __forceinline__ float merged_value(float w, float p, float alpha) {
  const float p_l = to_f32(from_f32<TL>(p));          // product rounded to LoRA dtype
  const float p_w = to_f32(from_f32<TW>(p_l));        // .type(weight.dtype)
  const float ap = to_f32(from_f32<TW>(__fmul_rn(alpha, p_w)));
  return __fadd_rn(w, ap);
}

// Block = 32 vectors x 8 row lanes; every thread keeps its RP x VEC slice of `down` in registers and
// walks rows ty, ty+8, ... of the block's row panel with UNROLL independent 16-byte loads in flight.
// r > RP (only RP = 16) makes several passes, accumulating through Wout in fp32.
template <typename TW, typename TL, int MODE, int RP>  // MODE 0 = merge, 1 = abs-sum
__global__ void __launch_bounds__(mrg::NT)
lora_merge_kernel(const TW* __restrict__ Win, TW* __restrict__ Wout, const TL* __restrict__ up,
                  const TL* __restrict__ down, float alpha, int N, int K, int r, int rows_per_block,
                  float* __restrict__ abs_sum) {
  constexpr int VEC = 16 / sizeof(TW);
  const int tx = threadIdx.x & (mrg::TX - 1), ty = threadIdx.x / mrg::TX;
  const int k0 = (blockIdx.x * mrg::TX + tx) * VEC;
  const int n_begin = blockIdx.y * rows_per_block;
  int n_end = n_begin + rows_per_block;
  if (n_end > N) n_end = N;
  float local = 0.f;
  if (k0 < K) {   // K % VEC == 0 guaranteed by the host
    for (int j0 = 0; j0 < r; j0 += RP) {
      float d[RP][VEC];
      const int jn = (r - j0) < RP ? (r - j0) : RP;
#pragma unroll
      for (int j = 0; j < RP; ++j)
#pragma unroll
        for (int v = 0; v < VEC; ++v)
          d[j][v] = (j < jn) ? to_f32(down[(long long)(j0 + j) * K + k0 + v]) : 0.f;
      const bool first = (j0 == 0), last = (j0 + RP >= r);
      const TW* src = first ? Win : Wout;
      for (int nb = n_begin + ty; nb < n_end; nb += mrg::TY * mrg::UNROLL) {
        constexpr bool BATCH_U = RP <= 8;     // RP = 16 would not fit the register file batched
        uint4 raw[mrg::UNROLL];
        float u[BATCH_U ? mrg::UNROLL : 1][RP];
        if (MODE == 0) {
#pragma unroll
          for (int q = 0; q < mrg::UNROLL; ++q) {
            const int n = nb + q * mrg::TY;
            if (n < n_end) raw[q] = *reinterpret_cast<const uint4*>(src + (long long)n * K + k0);
          }
        }
        if (BATCH_U) {
#pragma unroll
          for (int q = 0; q < mrg::UNROLL; ++q) {
            const int n = nb + q * mrg::TY;
#pragma unroll
            for (int j = 0; j < RP; ++j)
              u[q][j] = (n < n_end && j < jn) ? to_f32(up[(long long)n * r + j0 + j]) : 0.f;
          }
        }
#pragma unroll
        for (int q = 0; q < mrg::UNROLL; ++q) {
          const int n = nb + q * mrg::TY;
          if (n >= n_end) continue;
          const int uq = BATCH_U ? q : 0;
          if (!BATCH_U) {
#pragma unroll
            for (int j = 0; j < RP; ++j) u[0][j] = (j < jn) ? to_f32(up[(long long)n * r + j0 + j]) : 0.f;
          }
          float pacc[VEC];
#pragma unroll
          for (int v = 0; v < VEC; ++v) pacc[v] = 0.f;
#pragma unroll
          for (int j = 0; j < RP; ++j)
#pragma unroll
            for (int v = 0; v < VEC; ++v) pacc[v] = fmaf(u[uq][j], d[j][v], pacc[v]);
          if (MODE == 0) {
            TW* wv = reinterpret_cast<TW*>(&raw[q]);
            if (first && last) {
#pragma unroll
              for (int v = 0; v < VEC; ++v)
                wv[v] = from_f32<TW>(merged_value<TW, TL>(to_f32(wv[v]), pacc[v], alpha));
            } else {  // r > 16: plain fp32 accumulation of alpha * partial product
#pragma unroll
              for (int v = 0; v < VEC; ++v)
                wv[v] = from_f32<TW>(fmaf(alpha, pacc[v], to_f32(wv[v])));
            }
            *reinterpret_cast<uint4*>(Wout + (long long)n * K + k0) = raw[q];
          } else {
#pragma unroll
            for (int v = 0; v < VEC; ++v) local += fabsf(pacc[v]);
          }
        }
      }
    }
  }
  if (MODE == 1) {
    // block reduce -> one atomic per block (MODE 1 requires r <= 16, checked on the host)
    __shared__ float ws[mrg::NT / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = local;
    __syncthreads();
    if (threadIdx.x == 0) {
      float s = 0.f;
      for (int w = 0; w < mrg::NT / 32; ++w) s += ws[w];
      atomicAdd(abs_sum, s);
    }
  }
}

template <typename TW, typename TL, int MODE>
cudaError_t launch_merge(const void* Win, void* Wout, const void* up, const void* down,
                         float alpha, int N, int K, int r, float* abs_sum, cudaStream_t stream) {
  constexpr int VEC = 16 / sizeof(TW);
  const int kvecs = K / VEC;
  const int gx = (kvecs + mrg::TX - 1) / mrg::TX;
  // row panel per block: as tall as possible (amortises the `down` slice) while >= ~4 blocks per SM
  int rows = mrg::TY * mrg::UNROLL * 4;
  while (rows > mrg::TY && (long long)gx * ((N + rows - 1) / rows) < 148 * 4) rows >>= 1;
  dim3 grid(gx, (N + rows - 1) / rows);
#define MERGE_LAUNCH(RP_)                                                                        \
  lora_merge_kernel<TW, TL, MODE, RP_><<<grid, mrg::NT, 0, stream>>>(                            \
      (const TW*)Win, (TW*)Wout, (const TL*)up, (const TL*)down, alpha, N, K, r, rows, abs_sum)
  if (r <= 4) MERGE_LAUNCH(4);
  else if (r <= 8) MERGE_LAUNCH(8);
  else MERGE_LAUNCH(16);
#undef MERGE_LAUNCH
  return cudaGetLastError();
}

#define INST(TW, TL)                                                                           \
  template cudaError_t launch_merge<TW, TL, 0>(const void*, void*, const void*, const void*,   \
                                               float, int, int, int, float*, cudaStream_t);
INST(float, float) INST(float, __half) INST(float, __nv_bfloat16)
INST(__half, float) INST(__half, __half) INST(__half, __nv_bfloat16)
INST(__nv_bfloat16, float) INST(__nv_bfloat16, __half) INST(__nv_bfloat16, __nv_bfloat16)
#undef INST
// abs-sum: the "weight" type only sets the vector width; instantiate with TW == TL
template cudaError_t launch_merge<float, float, 1>(const void*, void*, const void*, const void*,
                                                   float, int, int, int, float*, cudaStream_t);
template cudaError_t launch_merge<__half, __half, 1>(const void*, void*, const void*, const void*,
                                                     float, int, int, int, float*, cudaStream_t);
template cudaError_t launch_merge<__nv_bfloat16, __nv_bfloat16, 1>(const void*, void*, const void*,
                                                                   const void*, float, int, int,
                                                                   int, float*, cudaStream_t);

// ---- abs-sum for any rank / any K (inspect_lora on r > 16 layers; diagnostic, not bandwidth-tuned):
// one output element per thread per step, the r-term dot product in fp32, rounded to the LoRA dtype like up @ down ----
__global__ void absmean_any_kernel(const void* __restrict__ up, const void* __restrict__ down, int dt,
                                   float* __restrict__ abs_sum, int N, int K, int r) {
  float local = 0.f;
  const long long total = (long long)N * K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int n = (int)(i / K), k = (int)(i - (long long)n * K);
    float p = 0.f;
    for (int j = 0; j < r; ++j)
      p = fmaf(load_as_float(up, dt, (long long)n * r + j), load_as_float(down, dt, (long long)j * K + k), p);
    if (dt == B200LORA_F16) p = __half2float(__float2half_rn(p));
    else if (dt == B200LORA_BF16) p = __bfloat162float(__float2bfloat16_rn(p));
    local += fabsf(p);
  }
  __shared__ float ws[8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += ws[w];
    atomicAdd(abs_sum, s);
  }
}

cudaError_t launch_absmean_any(const void* up, const void* down, float* out_sum, int N, int K, int r, int dt,
                               cudaStream_t stream) {
  long long total = (long long)N * K;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  absmean_any_kernel<<<blocks, 256, 0, stream>>>(up, down, dt, out_sum, N, K, r);
  return cudaGetLastError();
}

// ---- SVD epilogue ----
__global__ void svd_epilogue_kernel(float* __restrict__ U, const float* __restrict__ S,
                                    float* __restrict__ Vh, float hi, int do_scale, int N, int K,
                                    int r) {
  const long long nu = (long long)N * r, nv = (long long)r * K;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nu + nv;
       i += (long long)gridDim.x * blockDim.x) {
    if (i < nu) {
      float v = U[i];
      if (do_scale) v *= S[i % r];
      if (hi >= 0.f) v = fminf(fmaxf(v, -hi), hi);
      U[i] = v;
    } else {
      float v = Vh[i - nu];
      if (hi >= 0.f) v = fminf(fmaxf(v, -hi), hi);
      Vh[i - nu] = v;
    }
  }
}

cudaError_t launch_svd_epilogue(float* U, const float* S, float* Vh, float hi, int do_scale,
                                int N, int K, int r, cudaStream_t stream) {
  const long long total = (long long)N * r + (long long)r * K;
  int blocks = (int)((total + 255) / 256);
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  svd_epilogue_kernel<<<blocks, 256, 0, stream>>>(U, S, Vh, hi, do_scale, N, K, r);
  return cudaGetLastError();
}

// ---- LoRA pack: the TMA-loadable operands of the fused tcgen05 GEMM, all in the io dtype ----
//   A_pack  [16, K]  = A  rows (lora_down.weight), rows r..15 zero      forward  rank rows  (S1)
//   Bt_pack [16, N]  = Bᵀ rows (lora_up.weight),   rows r..15 zero      backward rank rows  (S1)
//   Bn_pack [N, 64]  = B  rows, column r = bias[n] (if any), rest zero  forward  tail block (S2)
//   At_pack [K, 64]  = Aᵀ rows, columns r..63 zero                      backward tail block (S2)
// One launch per (A, B) version; O(80·(K+N)) elements.
// fp32 -> TIO conversion of the layer's input riding in the pack launch (training under autocast: the input of a layer
// behind a LayerNorm is fp32; one launch instead of an ATen cast + the pack, and the consuming GEMM's prologue overlaps
// this grid through the programmatic launch).  8 elements per thread and pass: two 16-byte loads, one 16-byte store.
template <typename TIO>
__device__ __forceinline__ void cast_rows(const float* __restrict__ src, TIO* __restrict__ dst, long long n) {
  const long long nv = n >> 3;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
    const float4 a = __ldcs(reinterpret_cast<const float4*>(src) + 2 * i);
    const float4 b = __ldcs(reinterpret_cast<const float4*>(src) + 2 * i + 1);
    TIO o[8];
    o[0] = from_f32<TIO>(a.x); o[1] = from_f32<TIO>(a.y); o[2] = from_f32<TIO>(a.z); o[3] = from_f32<TIO>(a.w);
    o[4] = from_f32<TIO>(b.x); o[5] = from_f32<TIO>(b.y); o[6] = from_f32<TIO>(b.z); o[7] = from_f32<TIO>(b.w);
    *(reinterpret_cast<uint4*>(dst) + i) = *reinterpret_cast<const uint4*>(o);
  }
  for (long long i = (nv << 3) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    dst[i] = from_f32<TIO>(src[i]);
}

template <typename TIO>
__global__ void lora_pack_kernel(const void* __restrict__ A, const void* __restrict__ B, int ab_dtype,
                                 const void* __restrict__ bias, int bias_dtype,
                                 TIO* __restrict__ A_pack, TIO* __restrict__ Bt_pack,
                                 TIO* __restrict__ Bn_pack, TIO* __restrict__ At_pack, int K, int N,
                                 int r, const float* __restrict__ xsrc, TIO* __restrict__ xdst, long long xn) {
  const int n0 = 16 * K, n1 = n0 + 16 * N, n2 = n1 + 64 * N, total = n2 + 64 * K;
  // the GEMM that consumes this pack is the next launch on the stream: let its prologue (TMEM allocation,
  // barrier init, descriptor prefetch) run under this grid; its griddepcontrol.wait still sees a finished pack
  pdl_launch_dependents();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < n0) {
      const int j = i / K, k = i - j * K;
      A_pack[i] = from_f32<TIO>(j < r ? load_as_float(A, ab_dtype, (long long)j * K + k) : 0.f);
    } else if (i < n1) {
      const int t = i - n0;
      const int j = t / N, n = t - j * N;
      Bt_pack[t] = from_f32<TIO>(j < r ? load_as_float(B, ab_dtype, (long long)n * r + j) : 0.f);
    } else if (i < n2) {
      const int t = i - n1;
      const int n = t >> 6, j = t & 63;
      float v = 0.f;
      if (j < r) v = load_as_float(B, ab_dtype, (long long)n * r + j);
      else if (j == r && bias != nullptr) v = load_as_float(bias, bias_dtype, n);
      Bn_pack[t] = from_f32<TIO>(v);
    } else {
      const int t = i - n2;
      const int k = t >> 6, j = t & 63;
      At_pack[t] = from_f32<TIO>(j < r ? load_as_float(A, ab_dtype, (long long)j * K + k) : 0.f);
    }
  }
  if (xn > 0) cast_rows<TIO>(xsrc, xdst, xn);
}

cudaError_t launch_pack(const void* A, const void* B, int ab_dtype, const void* bias, int bias_dtype,
                        void* A_pack, void* Bt_pack, void* Bn_pack, void* At_pack, int K, int N,
                        int r, int io_dtype, cudaStream_t stream, const float* xsrc, void* xdst, long long xn) {
  const int total = 80 * (K + N);
  long long work = total;
  if (xn / 8 > work) work = xn / 8;
  int blocks = (int)((work + 255) / 256 < 148 * 8 ? (work + 255) / 256 : 148 * 8);
  if (blocks < 1) blocks = 1;
  if (io_dtype == B200LORA_BF16)
    lora_pack_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(
        A, B, ab_dtype, bias, bias_dtype, (__nv_bfloat16*)A_pack, (__nv_bfloat16*)Bt_pack,
        (__nv_bfloat16*)Bn_pack, (__nv_bfloat16*)At_pack, K, N, r, xsrc, (__nv_bfloat16*)xdst, xn);
  else
    lora_pack_kernel<__half><<<blocks, 256, 0, stream>>>(
        A, B, ab_dtype, bias, bias_dtype, (__half*)A_pack, (__half*)Bt_pack, (__half*)Bn_pack,
        (__half*)At_pack, K, N, r, xsrc, (__half*)xdst, xn);
  return cudaGetLastError();
}

// ---- segmented pack: the pack of up to 4 sibling layers' CONCATENATED problem (fused q/k/v), written
// straight from the layers' own factors: A_cat = [A_0; A_1; ...] (r_total rows), B block-diagonal with each
// layer's scale folded in, bias concatenated.  Replaces cat / zeros / slice-assign / mul launches per group.
struct PackSegs {
  const void* A[4];
  const void* B[4];
  const void* bias[4];
  float scale[4];
  int r[4];
  int off[4];
  int nseg;
};

template <typename TIO>
__global__ void lora_pack_seg_kernel(const PackSegs sg, int ab_dtype, int bias_dtype, TIO* __restrict__ A_pack,
                                     TIO* __restrict__ Bt_pack, TIO* __restrict__ Bn_pack,
                                     TIO* __restrict__ At_pack, int K, int Nseg, int r_total,
                                     const float* __restrict__ xsrc, TIO* __restrict__ xdst, long long xn) {
  const int N = sg.nseg * Nseg;
  const int n0 = 16 * K, n1 = n0 + 16 * N, n2 = n1 + 64 * N, total = n2 + 64 * K;
  pdl_launch_dependents();      // as in lora_pack_kernel
  auto seg_of_rank = [&](int j) {
    int s = 0;
#pragma unroll
    for (int q = 1; q < 4; ++q)
      if (q < sg.nseg && j >= sg.off[q]) s = q;
    return s;
  };
  auto a_val = [&](int j, int k) -> float {
    if (j >= r_total) return 0.f;
    const int s = seg_of_rank(j);
    return load_as_float(sg.A[s], ab_dtype, (long long)(j - sg.off[s]) * K + k);
  };
  auto b_val = [&](int n, int j) -> float {      // block-diagonal, scale folded (rounded once, to TIO)
    if (j >= r_total) return 0.f;
    const int s = n / Nseg;
    const int jl = j - sg.off[s];
    if (jl < 0 || jl >= sg.r[s]) return 0.f;
    return load_as_float(sg.B[s], ab_dtype, (long long)(n - s * Nseg) * sg.r[s] + jl) * sg.scale[s];
  };
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    if (i < n0) {
      const int j = i / K, k = i - j * K;
      A_pack[i] = from_f32<TIO>(a_val(j, k));
    } else if (i < n1) {
      const int t = i - n0;
      const int j = t / N, n = t - j * N;
      Bt_pack[t] = from_f32<TIO>(b_val(n, j));
    } else if (i < n2) {
      const int t = i - n1;
      const int n = t >> 6, j = t & 63;
      float v = b_val(n, j);
      if (j == r_total) {
        const int s = n / Nseg;
        if (sg.bias[s] != nullptr) v = load_as_float(sg.bias[s], bias_dtype, n - s * Nseg);
      }
      Bn_pack[t] = from_f32<TIO>(v);
    } else {
      const int t = i - n2;
      const int k = t >> 6, j = t & 63;
      At_pack[t] = from_f32<TIO>(a_val(j, k));
    }
  }
  if (xn > 0) cast_rows<TIO>(xsrc, xdst, xn);
}

cudaError_t launch_pack_seg(const void* const* A, const void* const* B, const void* const* bias,
                            const float* scales, const int* r_each, int nseg, int ab_dtype, int bias_dtype,
                            void* A_pack, void* Bt_pack, void* Bn_pack, void* At_pack, int K, int Nseg,
                            int io_dtype, cudaStream_t stream, const float* xsrc, void* xdst, long long xn) {
  PackSegs sg = {};
  sg.nseg = nseg;
  int off = 0;
  for (int s = 0; s < nseg; ++s) {
    sg.A[s] = A[s];
    sg.B[s] = B[s];
    sg.bias[s] = bias ? bias[s] : nullptr;
    sg.scale[s] = scales[s];
    sg.r[s] = r_each[s];
    sg.off[s] = off;
    off += r_each[s];
  }
  const int total = 80 * (K + nseg * Nseg);
  long long work = total;
  if (xn / 8 > work) work = xn / 8;
  int blocks = (int)((work + 255) / 256 < 148 * 8 ? (work + 255) / 256 : 148 * 8);
  if (blocks < 1) blocks = 1;
  if (io_dtype == B200LORA_BF16)
    lora_pack_seg_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>(
        sg, ab_dtype, bias_dtype, (__nv_bfloat16*)A_pack, (__nv_bfloat16*)Bt_pack, (__nv_bfloat16*)Bn_pack,
        (__nv_bfloat16*)At_pack, K, Nseg, off, xsrc, (__nv_bfloat16*)xdst, xn);
  else
    lora_pack_seg_kernel<__half><<<blocks, 256, 0, stream>>>(
        sg, ab_dtype, bias_dtype, (__half*)A_pack, (__half*)Bt_pack, (__half*)Bn_pack, (__half*)At_pack, K,
        Nseg, off, xsrc, (__half*)xdst, xn);
  return cudaGetLastError();
}

}  // namespace b200
