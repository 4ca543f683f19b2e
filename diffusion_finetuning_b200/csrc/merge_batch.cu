// Whole-model weight merge in ONE launch:  W_i <- W_i + alpha * (up_i @ down_i)  for every injected layer i
// (lora.py:410-424 `weight_apply_lora` loops over the layers; cli_lora_add.py:72-110 does it for a UNet's 144 and a
// text encoder's 48 matrices).  One launch per matrix was launch-bound (144 x 7.6 us against 0.24 ms of HBM time for the
// whole fp32 UNet); here a device-side table describes every matrix and a persistent grid walks a flat list of
// [32 rows x 512 bytes] tiles across all of them, so the kernel is one long HBM stream.
//
// Rounding follows the reference expression exactly (same as merge.cu): the rank-r product is accumulated in fp32 and
// rounded ONCE to the LoRA dtype (what `up @ down` yields), cast to the weight dtype, multiplied by alpha in fp32 and
// rounded to the weight dtype, added in fp32 and rounded to the weight dtype.  All 16-bit roundings are PACKED
// conversions of two values (F2FP on the ALU pipe): the scalar cvt.rn.f16.f32 form issues on the quarter-rate
// conversion pipe and capped the 16-bit merge at ~2.4 TB/s.
#include <type_traits>

#include "common.cuh"

namespace b200 {

struct MergeItem {          // device-side record, 48 bytes
  const void* Win;
  void* Wout;
  const void* up;
  const void* down;
  int N, K, r;
  int tile0;                // index of this matrix's first tile in the flat tile list
  int tiles_k;              // tiles along K
  int pad;
};

namespace mb {
constexpr int TX = 32;      // 16-byte vectors along K per tile (512 contiguous bytes per row)
constexpr int TY = 4;       // row lanes
constexpr int NT = TX * TY; // 128 threads: FOUR CTAs per SM, so the fill / load / compute / store phases of a tile
                            // (serial inside one CTA) overlap across four independent streams per SM
constexpr int UNROLL = 4;   // rows in flight per thread (multi-pass path)
constexpr int ROWS = 32;    // rows per tile: 8 per thread, all 8 loads in flight at once (single-pass path)
}  // namespace mb

// round two fp32 values to T (round-to-nearest-even) and widen them back
template <typename T> __device__ __forceinline__ void round2(float& a, float& b);
template <> __device__ __forceinline__ void round2<float>(float&, float&) {}
template <> __device__ __forceinline__ void round2<__half>(float& a, float& b) {
  const float2 f = __half22float2(__floats2half2_rn(a, b));
  a = f.x; b = f.y;
}
template <> __device__ __forceinline__ void round2<__nv_bfloat16>(float& a, float& b) {
  const float2 f = __bfloat1622float2(__floats2bfloat162_rn(a, b));
  a = f.x; b = f.y;
}
template <typename T> __device__ __forceinline__ float2 unpack2(uint32_t v);
template <> __device__ __forceinline__ float2 unpack2<__half>(uint32_t v) {
  return __half22float2(*reinterpret_cast<const __half2*>(&v));
}
template <> __device__ __forceinline__ float2 unpack2<__nv_bfloat16>(uint32_t v) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&v));
}
template <typename T> __device__ __forceinline__ uint32_t pack2(float a, float b);
template <> __device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  const __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  const __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<const uint32_t*>(&h);
}

template <typename TW> struct WVec;       // one 16-byte vector of W as fp32 values
template <> struct WVec<float> {
  static constexpr int VEC = 4;
  static __device__ __forceinline__ void load(const uint4& raw, float (&w)[4]) {
    w[0] = __uint_as_float(raw.x); w[1] = __uint_as_float(raw.y);
    w[2] = __uint_as_float(raw.z); w[3] = __uint_as_float(raw.w);
  }
  static __device__ __forceinline__ uint4 store(const float (&w)[4]) {
    return make_uint4(__float_as_uint(w[0]), __float_as_uint(w[1]), __float_as_uint(w[2]), __float_as_uint(w[3]));
  }
};
template <typename T16> struct WVec16 {
  static constexpr int VEC = 8;
  static __device__ __forceinline__ void load(const uint4& raw, float (&w)[8]) {
    float2 f;
    f = unpack2<T16>(raw.x); w[0] = f.x; w[1] = f.y;
    f = unpack2<T16>(raw.y); w[2] = f.x; w[3] = f.y;
    f = unpack2<T16>(raw.z); w[4] = f.x; w[5] = f.y;
    f = unpack2<T16>(raw.w); w[6] = f.x; w[7] = f.y;
  }
  static __device__ __forceinline__ uint4 store(const float (&w)[8]) {
    return make_uint4(pack2<T16>(w[0], w[1]), pack2<T16>(w[2], w[3]), pack2<T16>(w[4], w[5]), pack2<T16>(w[6], w[7]));
  }
};
template <> struct WVec<__half> : WVec16<__half> {};
template <> struct WVec<__nv_bfloat16> : WVec16<__nv_bfloat16> {};

template <typename TL> __device__ __forceinline__ float ld_l(const TL* p) { return to_f32(__ldg(p)); }

// VEC consecutive LoRA elements as fp32: one 8 / 16 / 32-byte load when the address allows it
template <typename TL, int VEC>
__device__ __forceinline__ void ld_row(const TL* __restrict__ p, bool aligned, float (&out)[VEC]) {
  constexpr int BYTES = VEC * (int)sizeof(TL);
  if (aligned) {
    if constexpr (BYTES == 8) {
      const uint2 raw = __ldg(reinterpret_cast<const uint2*>(p));
      const TL* e = reinterpret_cast<const TL*>(&raw);
#pragma unroll
      for (int v = 0; v < VEC; ++v) out[v] = to_f32(e[v]);
    } else {
#pragma unroll
      for (int i = 0; i < BYTES / 16; ++i) {
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>(p) + i);
        const TL* e = reinterpret_cast<const TL*>(&raw);
#pragma unroll
        for (int v = 0; v < 16 / (int)sizeof(TL); ++v) out[i * (16 / (int)sizeof(TL)) + v] = to_f32(e[v]);
      }
    }
  } else {
#pragma unroll
    for (int v = 0; v < VEC; ++v) out[v] = ld_l(p + v);
  }
}

// JR: rank rows of `down` resident in registers per pass (r <= 4 -> 4, r <= 8 -> 8, larger r -> passes of 8)
//
// Per tile (r <= JR, the common case): the tile's rows of `up` (ROWS x r contiguous elements) go to shared memory with
// one coalesced pass, the thread's r x VEC slice of `down` to registers with vector loads, then EIGHT independent
// 16-byte loads of W per thread are in flight before the first is consumed (4 CTAs x 128 threads x 128 B = 64 KB per
// SM: what HBM latency x bandwidth asks for; with 4 loads in flight the fp32 merge sat at 4.7 TB/s = 32 KB per SM).
template <typename TW, typename TL, int JR>
__global__ void __launch_bounds__(mb::NT, JR <= 4 ? 4 : 2)
lora_merge_batch_kernel(const MergeItem* __restrict__ items, int n_items, int total_tiles, float alpha) {
  constexpr int VEC = WVec<TW>::VEC;
  constexpr int RPT = mb::ROWS / mb::TY;            // rows per thread per tile (8)
  __shared__ float s_up[2][mb::ROWS * JR];
  const int tx = threadIdx.x & (mb::TX - 1), ty = threadIdx.x / mb::TX;
  int flip = 0;
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    // which matrix: last item with tile0 <= tile (uniform across the block: broadcast loads)
    int lo = 0, hi = n_items - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (__ldg(&items[mid].tile0) <= tile) lo = mid; else hi = mid - 1;
    }
    const MergeItem it = items[lo];
    const int local = tile - it.tile0;
    const int kc = local % it.tiles_k, panel = local / it.tiles_k;
    const int k0 = (kc * mb::TX + tx) * VEC;
    const int n_begin = panel * mb::ROWS;
    const int n_end = min(n_begin + mb::ROWS, it.N);
    const TW* Win = static_cast<const TW*>(it.Win);
    TW* Wout = static_cast<TW*>(it.Wout);
    const TL* up = static_cast<const TL*>(it.up);
    const TL* down = static_cast<const TL*>(it.down);
    const int r = it.r, K = it.K;
    if (r <= JR) {
      // ---- the tile's rows of `up`: (n_end - n_begin) * r contiguous elements -> shared memory, as fp32 ----
      float* su = s_up[flip];
      flip ^= 1;
      const int cnt = (n_end - n_begin) * r;
      for (int i = threadIdx.x; i < cnt; i += mb::NT) su[i] = ld_l(up + (long long)n_begin * r + i);
      __syncthreads();      // (two buffers: the next tile's fill cannot overtake this tile's readers)
      if (k0 < K) {                                   // K % VEC == 0 (host-checked): whole vectors only
        float d[JR][VEC];
        const bool d_al = ((reinterpret_cast<uintptr_t>(down) | ((uintptr_t)K * sizeof(TL))) & (VEC * sizeof(TL) - 1)) == 0;
#pragma unroll
        for (int j = 0; j < JR; ++j) {
          if (j < r) ld_row<TL, VEC>(down + (long long)j * K + k0, d_al, d[j]);
          else {
#pragma unroll
            for (int v = 0; v < VEC; ++v) d[j][v] = 0.f;
          }
        }
        uint4 raw[RPT];
#pragma unroll
        for (int q = 0; q < RPT; ++q) {
          const int n = n_begin + ty + q * mb::TY;
          if (n < n_end) raw[q] = __ldcs(reinterpret_cast<const uint4*>(Win + (long long)n * K + k0));
        }
#pragma unroll
        for (int q = 0; q < RPT; ++q) {
          const int ln = ty + q * mb::TY;
          const int n = n_begin + ln;
          if (n >= n_end) continue;
          float p[VEC], w[VEC];
#pragma unroll
          for (int v = 0; v < VEC; ++v) p[v] = 0.f;
#pragma unroll
          for (int j = 0; j < JR; ++j) {
            const float u = (j < r) ? su[ln * r + j] : 0.f;
#pragma unroll
            for (int v = 0; v < VEC; ++v) p[v] = fmaf(u, d[j][v], p[v]);
          }
          WVec<TW>::load(raw[q], w);
#pragma unroll
          for (int v = 0; v < VEC; v += 2) {
            float a = p[v], b = p[v + 1];
            round2<TL>(a, b);                 // up @ down in the LoRA dtype
            if constexpr (!std::is_same<TL, TW>::value) round2<TW>(a, b);     // .type(weight.dtype)
            a = __fmul_rn(alpha, a); b = __fmul_rn(alpha, b);
            round2<TW>(a, b);                 // alpha * (...) is a tensor of the weight dtype
            w[v] = __fadd_rn(w[v], a); w[v + 1] = __fadd_rn(w[v + 1], b);
          }
          __stcs(reinterpret_cast<uint4*>(Wout + (long long)n * K + k0), WVec<TW>::store(w));
        }
      }
    } else if constexpr (JR == 8) {     // (the JR = 4 kernel is only launched when every rank is <= 4)
      // ---- r > JR (JR == 8 here): passes of 8 rank rows, product accumulated in registers per row ----
      for (int nb = n_begin + ty; nb < n_end && k0 < K; nb += mb::TY * mb::UNROLL) {
        uint4 raw[mb::UNROLL];
        float p[mb::UNROLL][VEC];
#pragma unroll
        for (int q = 0; q < mb::UNROLL; ++q) {
          const int n = nb + q * mb::TY;
          if (n < n_end) raw[q] = __ldcs(reinterpret_cast<const uint4*>(Win + (long long)n * K + k0));
#pragma unroll
          for (int v = 0; v < VEC; ++v) p[q][v] = 0.f;
        }
        for (int j0 = 0; j0 < r; j0 += JR) {
          float d[JR][VEC];
#pragma unroll
          for (int j = 0; j < JR; ++j)
#pragma unroll
            for (int v = 0; v < VEC; ++v)
              d[j][v] = (j0 + j < r) ? ld_l(down + (long long)(j0 + j) * K + k0 + v) : 0.f;
#pragma unroll
          for (int q = 0; q < mb::UNROLL; ++q) {
            const int n = nb + q * mb::TY;
#pragma unroll
            for (int j = 0; j < JR; ++j) {
              const float u = (n < n_end && j0 + j < r) ? ld_l(up + (long long)n * r + j0 + j) : 0.f;
#pragma unroll
              for (int v = 0; v < VEC; ++v) p[q][v] = fmaf(u, d[j][v], p[q][v]);
            }
          }
        }
#pragma unroll
        for (int q = 0; q < mb::UNROLL; ++q) {
          const int n = nb + q * mb::TY;
          if (n >= n_end) continue;
          float w[VEC];
          WVec<TW>::load(raw[q], w);
#pragma unroll
          for (int v = 0; v < VEC; v += 2) {
            float a = p[q][v], b = p[q][v + 1];
            round2<TL>(a, b);
            round2<TW>(a, b);
            a = __fmul_rn(alpha, a); b = __fmul_rn(alpha, b);
            round2<TW>(a, b);
            w[v] = __fadd_rn(w[v], a); w[v + 1] = __fadd_rn(w[v + 1], b);
          }
          __stcs(reinterpret_cast<uint4*>(Wout + (long long)n * K + k0), WVec<TW>::store(w));
        }
      }
    }
  }
}

// host: fill tile0 / tiles_k of `items` (host copy), return the total tile count
long long merge_batch_plan(MergeItem* items, int n, int w_dtype) {
  const int vec = 16 / dtype_size(w_dtype);
  long long t = 0;
  for (int i = 0; i < n; ++i) {
    const int kvecs = items[i].K / vec;
    items[i].tiles_k = (kvecs + mb::TX - 1) / mb::TX;
    items[i].tile0 = (int)t;
    t += (long long)items[i].tiles_k * ((items[i].N + mb::ROWS - 1) / mb::ROWS);
  }
  return t;
}

template <typename TW, typename TL>
static cudaError_t launch_merge_batch_t(const MergeItem* items_dev, int n, int total_tiles, int max_r, float alpha,
                                        int sms, cudaStream_t stream) {
  int grid = sms * 16;      // 4 resident CTAs per SM, 4 waves: short tails, tiles handed out round-robin
  if (grid > total_tiles) grid = total_tiles;
  if (grid < 1) grid = 1;
  if (max_r <= 4)
    lora_merge_batch_kernel<TW, TL, 4><<<grid, mb::NT, 0, stream>>>(items_dev, n, total_tiles, alpha);
  else
    lora_merge_batch_kernel<TW, TL, 8><<<grid, mb::NT, 0, stream>>>(items_dev, n, total_tiles, alpha);
  return cudaGetLastError();
}

cudaError_t launch_merge_batch(const MergeItem* items_dev, int n, int total_tiles, int max_r, float alpha,
                               int w_dtype, int lora_dtype, int sms, cudaStream_t stream) {
#define M2(TW, TL) launch_merge_batch_t<TW, TL>(items_dev, n, total_tiles, max_r, alpha, sms, stream)
#define M1(TW)                                                          \
  (lora_dtype == B200LORA_F32 ? M2(TW, float)                           \
   : lora_dtype == B200LORA_F16 ? M2(TW, __half) : M2(TW, __nv_bfloat16))
  if (w_dtype == B200LORA_F32) return M1(float);
  if (w_dtype == B200LORA_F16) return M1(__half);
  return M1(__nv_bfloat16);
#undef M1
#undef M2
}

}  // namespace b200
