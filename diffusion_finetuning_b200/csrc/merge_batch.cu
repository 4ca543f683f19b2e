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
//
// 16-bit weights stay PACKED from the product's rounding to the store (r02: the unpack / fp32 op / re-pack form issued
// 130 instructions per 16-byte vector and the kernel sat at 54 % issue-active, 3.5 TB/s).  Both replacements are
// bit-identical to ATen's "compute in fp32, round to the tensor dtype":
//   * W + x  ->  one add.rn.{f16x2,bf16x2}: rounding the exact sum to 24 bits and then to p bits equals rounding it once
//     to p bits whenever 24 >= 2p + 2 (p = 11 for fp16, 8 for bf16: innocuous double rounding);
//   * alpha * x, alpha representable in the weight dtype  ->  one mul.rn.{f16x2,bf16x2}: the fp32 product of two p-bit
//     values is exact, so both forms round the exact product once.  alpha == 1 skips the multiply; any other alpha
//     takes the unpack / __fmul_rn / re-pack form.
// The _rn intrinsics are the ones the compiler may not contract into an fma (which would round once instead of twice).
#include <type_traits>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

struct MergeItem {          // device-side record, 56 bytes
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

template <typename T> __device__ __forceinline__ uint32_t hadd2_rn(uint32_t a, uint32_t b);
template <> __device__ __forceinline__ uint32_t hadd2_rn<__half>(uint32_t a, uint32_t b) {
  const __half2 c = __hadd2_rn(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
  return *reinterpret_cast<const uint32_t*>(&c);
}
template <> __device__ __forceinline__ uint32_t hadd2_rn<__nv_bfloat16>(uint32_t a, uint32_t b) {
  const __nv_bfloat162 c = __hadd2_rn(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&c);
}
template <typename T> __device__ __forceinline__ uint32_t hmul2_rn(uint32_t a, uint32_t b);
template <> __device__ __forceinline__ uint32_t hmul2_rn<__half>(uint32_t a, uint32_t b) {
  const __half2 c = __hmul2_rn(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
  return *reinterpret_cast<const uint32_t*>(&c);
}
template <> __device__ __forceinline__ uint32_t hmul2_rn<__nv_bfloat16>(uint32_t a, uint32_t b) {
  const __nv_bfloat162 c = __hmul2_rn(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b));
  return *reinterpret_cast<const uint32_t*>(&c);
}

// alpha as the kernel applies it: 0 = alpha == 1 (no multiply), 1 = representable in the weight dtype (packed multiply),
// 2 = any other value (fp32 multiply of the unpacked values)
enum { ALPHA_ONE = 0, ALPHA_PACKED = 1, ALPHA_F32 = 2, ALPHA_RUNTIME = 3 };

// W (+)= alpha * x for one 16-byte vector, p = the fp32 rank-r products of its VEC elements
template <typename TW, typename TL, int MODE> struct Apply;
template <typename TL, int MODE> struct Apply<float, TL, MODE> {
  static __device__ __forceinline__ uint4 run(const uint4& raw, float (&p)[4], float alpha, uint32_t) {
    round2<TL>(p[0], p[1]); round2<TL>(p[2], p[3]);                 // up @ down in the LoRA dtype
    return make_uint4(__float_as_uint(__fadd_rn(__uint_as_float(raw.x), __fmul_rn(alpha, p[0]))),
                      __float_as_uint(__fadd_rn(__uint_as_float(raw.y), __fmul_rn(alpha, p[1]))),
                      __float_as_uint(__fadd_rn(__uint_as_float(raw.z), __fmul_rn(alpha, p[2]))),
                      __float_as_uint(__fadd_rn(__uint_as_float(raw.w), __fmul_rn(alpha, p[3]))));
  }
};
template <typename T16, typename TL, int MODE> struct Apply16 {
  static __device__ __forceinline__ uint32_t one(uint32_t w, float a, float b, float alpha, uint32_t alpha2) {
    if constexpr (!std::is_same<TL, T16>::value) round2<TL>(a, b);  // up @ down in the LoRA dtype ...
    uint32_t x = pack2<T16>(a, b);                                  // ... .type(weight.dtype) (TL == T16: the same rounding)
    if constexpr (MODE == ALPHA_PACKED) x = hmul2_rn<T16>(alpha2, x);
    if constexpr (MODE == ALPHA_F32) {
      const float2 f = unpack2<T16>(x);
      x = pack2<T16>(__fmul_rn(alpha, f.x), __fmul_rn(alpha, f.y));
    }
    if constexpr (MODE == ALPHA_RUNTIME) {        // alpha2 = 0: the fp32 form; otherwise the packed multiply (1.0 included)
      if (alpha2 != 0u) x = hmul2_rn<T16>(alpha2, x);
      else {
        const float2 f = unpack2<T16>(x);
        x = pack2<T16>(__fmul_rn(alpha, f.x), __fmul_rn(alpha, f.y));
      }
    }
    return hadd2_rn<T16>(w, x);
  }
  static __device__ __forceinline__ uint4 run(const uint4& raw, float (&p)[8], float alpha, uint32_t alpha2) {
    return make_uint4(one(raw.x, p[0], p[1], alpha, alpha2), one(raw.y, p[2], p[3], alpha, alpha2),
                      one(raw.z, p[4], p[5], alpha, alpha2), one(raw.w, p[6], p[7], alpha, alpha2));
  }
};
template <typename TL, int MODE> struct Apply<__half, TL, MODE> : Apply16<__half, TL, MODE> {};
template <typename TL, int MODE> struct Apply<__nv_bfloat16, TL, MODE> : Apply16<__nv_bfloat16, TL, MODE> {};

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
  int mode = ALPHA_F32;       // fp32 weights: one form
  uint32_t alpha2 = 0;
  if constexpr (!std::is_same<TW, float>::value) {
    if (alpha == 1.0f) mode = ALPHA_ONE;
    else if (to_f32(from_f32<TW>(alpha)) == alpha) { mode = ALPHA_PACKED; alpha2 = pack2<TW>(alpha, alpha); }
  }
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
      // ---- the tile's rows of `up` -> shared memory as fp32, JR per row, zero beyond r and beyond the last row ----
      float* su = s_up[flip];
      flip ^= 1;
#pragma unroll
      for (int i = threadIdx.x; i < mb::ROWS * JR; i += mb::NT) {
        const int row = i / JR, j = i % JR;
        su[i] = (j < r && n_begin + row < n_end) ? ld_l(up + (long long)(n_begin + row) * r + j) : 0.f;
      }
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
        const TW* src = Win + (long long)(n_begin + ty) * K + k0;
        TW* dst = Wout + (long long)(n_begin + ty) * K + k0;
        const long long row_step = (long long)mb::TY * K;
        const int n_rows = n_end - n_begin;
        uint4 raw[RPT];
#pragma unroll
        for (int q = 0; q < RPT; ++q)
          if (ty + q * mb::TY < n_rows) raw[q] = __ldcs(reinterpret_cast<const uint4*>(src + q * row_step));
        auto rows = [&](auto mode_c) {
          constexpr int MODE = decltype(mode_c)::value;
#pragma unroll
          for (int q = 0; q < RPT; ++q) {
            const int ln = ty + q * mb::TY;
            if (ln >= n_rows) continue;
            float u[JR];
#pragma unroll
            for (int j4 = 0; j4 < JR; j4 += 4) {
              const float4 t = *reinterpret_cast<const float4*>(su + ln * JR + j4);
              u[j4] = t.x; u[j4 + 1] = t.y; u[j4 + 2] = t.z; u[j4 + 3] = t.w;
            }
            float pr[VEC];
#pragma unroll
            for (int v = 0; v < VEC; ++v) pr[v] = 0.f;
#pragma unroll
            for (int j = 0; j < JR; ++j)
#pragma unroll
              for (int v = 0; v < VEC; ++v) pr[v] = fmaf(u[j], d[j][v], pr[v]);
            __stcs(reinterpret_cast<uint4*>(dst + q * row_step), Apply<TW, TL, MODE>::run(raw[q], pr, alpha, alpha2));
          }
        };
        if (mode == ALPHA_ONE) rows(std::integral_constant<int, ALPHA_ONE>{});
        else if (mode == ALPHA_PACKED) rows(std::integral_constant<int, ALPHA_PACKED>{});
        else rows(std::integral_constant<int, ALPHA_F32>{});
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
          __stcs(reinterpret_cast<uint4*>(Wout + (long long)n * K + k0),
                 Apply<TW, TL, ALPHA_F32>::run(raw[q], p[q], alpha, 0u));
        }
      }
    }
  }
}

// -----------------------------------------------------------------------------------------------------------------
// Streamed variant (every rank <= 8): the loads leave the consumers' instruction stream altogether.
//
// A tile is R WHOLE ROWS of one matrix (R = what fits 32 KB, at least 8), i.e. ONE contiguous span of W, so one
// `cp.async.bulk` lands it in shared memory and a second one lands the tile's rows of `up` (also contiguous); rows longer
// than 4 KB are cut into equal column blocks of <= 4 KB, 8 rows per tile, one copy per row.  The first version of this
// kernel kept the [32 rows x 512 bytes] tiles of the kernel above and issued one 512-byte copy per row: 37 copies per
// 16 KB, and the TMA unit retires a bulk copy every ~80 cycles per SM whatever its size (measured: 2.8-3.4 TB/s,
// consumers 70 % of the time in the full-barrier wait, no gain from four producer warps) — few large copies is the rule.
// One producer warp per CTA walks the CTA's tiles ahead of the math through a ring of NS stages, completion counted in
// bytes on the stage's `full` mbarrier; G groups of four consumer warps (3 for fp32 weights, 6 for 16-bit ones: twice the
// arithmetic per byte) take alternate tiles.  Within a tile a thread owns one 16-byte column per pass (passes = power-of-
// two column chunks, so all 128 threads of a group work in every pass): its r x VEC slice of `down` sits in registers
// (fetched through L1 BEFORE the wait on `full`), it walks the rows with LDS.128, FFMA2 for the rank-r product, the
// rounding chain, and a coalesced 16-byte store to global.  Bytes in flight are set by the ring (6 x 32 KB per SM), not
// by registers; the table is mirrored in shared memory and walked by a forward-only pointer (no search); there is no
// __syncthreads in the loop.  Tiles -> CTAs: round-robin for fp32 weights, contiguous byte-balanced ranges for 16-bit
// ones (`down` then stays in L1 across a CTA's tiles); both measured, see MS_CONTIG*.
// -----------------------------------------------------------------------------------------------------------------
#ifndef MS_G32
#define MS_G32 3           // consumer groups, fp32 weights (more registers per thread: 4 rows in flight)
#endif
#ifndef MS_G16
#define MS_G16 6           // ... 16-bit weights (twice the arithmetic per byte: more warps to hide its latency)
#endif
#ifndef MS_CONTIG32
#define MS_CONTIG32 0      // tiles -> CTAs: 1 = contiguous byte-balanced ranges, 0 = round-robin (measured: fp32 6.29 vs 5.73 TB/s, 16-bit 5.24 vs 5.33)
#endif
#ifndef MS_CONTIG16
#define MS_CONTIG16 1
#endif
#ifndef MS_RU16
#define MS_RU16 1          // rows in flight per thread, 16-bit weights (72 registers per thread at 6 groups)
#endif
#ifndef MS_RU32
#define MS_RU32 4          // ... fp32 weights
#endif
#ifndef MS_MODE16
#define MS_MODE16 1        // 0: warp-uniform branch on alpha's form per packed word, 1: one copy of the row loop per form
#endif
namespace ms {
constexpr int GT = 128;                         // threads of a consumer group; group g takes the CTA's tiles it % G == g
constexpr int W_BYTES = 32 * 1024;              // W span per stage
constexpr int MAX_ROWS = 128;                   // rows per tile (bounds the `up` span)
constexpr int BLOCK_ROWS = 8, BLOCK_BYTES = 4096;   // tiles of rows longer than BLOCK_BYTES
constexpr int MAX_NS = 8;
constexpr int TABLE_ITEMS = 192;                // table entries mirrored in shared memory (larger tables are read from global)

// geometry of tile `local` of an item.  pad = rows per tile R, tiles_k = column blocks per row:
//   tiles_k == 1: R whole rows = one contiguous span (one bulk copy);
//   tiles_k  > 1: rows longer than 4 KB are cut into tiles_k equal column blocks of <= 4 KB, R = 8 rows per tile
//                 (one bulk copy per row), so a thread still walks >= 8 rows per slice of `down` it fetches.
struct TileGeo {
  int n0, rows;               // first row, number of rows
  int col0_bytes, row_bytes;  // first byte of the rows covered, bytes per row covered (multiple of 16)
};
__host__ __device__ inline int block_bytes(int full_row, int blocks) {      // width of a column block (last may be narrower)
  return ((full_row + blocks - 1) / blocks + 15) & ~15;
}
__device__ __forceinline__ TileGeo tile_geo(const MergeItem& im, int local, int esize) {
  TileGeo g;
  const int full_row = im.K * esize;
  if (im.tiles_k == 1) {
    g.n0 = local * im.pad; g.col0_bytes = 0; g.row_bytes = full_row;
  } else {
    const int cb = block_bytes(full_row, im.tiles_k);
    g.n0 = (local / im.tiles_k) * im.pad;
    g.col0_bytes = (local % im.tiles_k) * cb; g.row_bytes = min(cb, full_row - g.col0_bytes);
  }
  g.rows = min(im.pad, im.N - g.n0);
  return g;
}
// the tile's rows of `up` travel by bulk copy when source address and size are multiples of 16 (both sides evaluate this)
template <typename TL>
__device__ __forceinline__ bool up_by_copy(const MergeItem& im, const TileGeo& g) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(im.up) + (uintptr_t)g.n0 * im.r * sizeof(TL);
  return ((a | ((uintptr_t)g.rows * im.r * sizeof(TL))) & 15) == 0;
}
}  // namespace ms

// two fp32 fused multiply-adds in one issue slot (FFMA2, sm_100): each half rounds exactly as fmaf does
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)),
        "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}

__device__ __forceinline__ void bulk_load_1d(uint32_t dst_smem, const void* src, uint32_t bytes, uint64_t* bar,
                                             uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;\n" ::"r"(
          dst_smem),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

template <typename TW, typename TL, int JR, int G>
__global__ void __launch_bounds__(G * ms::GT + 32, 1)
lora_merge_stream_kernel(const MergeItem* __restrict__ items, int n_items, int total_tiles, float alpha, int ns, int contig) {
  constexpr int NT = G * ms::GT + 32;               // G consumer groups + the producer warp
  constexpr int VEC = WVec<TW>::VEC;
  constexpr int UP_BYTES = ms::MAX_ROWS * JR * (int)sizeof(TL);
  constexpr int STAGE = ms::W_BYTES + UP_BYTES;
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t full_bar[ms::MAX_NS], empty_bar[ms::MAX_NS];
  __shared__ __align__(16) MergeItem s_items[ms::TABLE_ITEMS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // ns is a multiple of G (host): a stage is always consumed by the SAME group.  Otherwise a group that runs ahead could
  // test a stage's `full` barrier for the parity of its NEXT use while the current use is still loading — a parity
  // wait only tells "not the phase in flight" — and read the other group's tile (seen with G = 4, ns = 6).
  if (threadIdx.x == 0) {
    for (int s = 0; s < ns; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], ms::GT / 32); }
    fence_mbar_init();
  }
  // the table walk is on every tile's critical path (round-robin tiles change matrix almost every time): mirror it
  const bool tab_smem = n_items <= ms::TABLE_ITEMS;
  if (tab_smem) {
    static_assert(sizeof(MergeItem) % 8 == 0, "MergeItem is copied in 8-byte words");
    const uint2* src = reinterpret_cast<const uint2*>(items);
    uint2* dst = reinterpret_cast<uint2*>(s_items);
    for (int i = threadIdx.x; i < n_items * (int)(sizeof(MergeItem) / 8); i += NT) dst[i] = __ldg(src + i);
  }
  __syncthreads();
  const MergeItem* tab = tab_smem ? s_items : items;      // generic pointer
  // This CTA's tiles: a CONTIGUOUS range, cut so that every CTA gets the same number of BYTES (tiles differ in size
  // between matrices).  Consecutive tiles then belong to the same matrix: its `down` stays in L1 (the per-pass
  // slices are re-read from there, ~40 cycles instead of an L2 round trip per tile) and the table pointer rarely moves.
  __shared__ int s_range[3];                              // first tile, one past the last, item of the first tile
  if (threadIdx.x == 0) {
    long long total = 0;
    for (int i = 0; i < n_items; ++i) total += (long long)tab[i].N * tab[i].K;
    const long long lo = total / gridDim.x * blockIdx.x + total % gridDim.x * blockIdx.x / gridDim.x;
    const long long hi = total / gridDim.x * (blockIdx.x + 1) + total % gridDim.x * (blockIdx.x + 1) / gridDim.x;
    long long acc = 0;
    int first = total_tiles, last = total_tiles, first_item = n_items - 1;
    bool got_lo = false;
    for (int i = 0; i < n_items; ++i) {
      const long long sz = (long long)tab[i].N * tab[i].K;
      const int nt = (i + 1 < n_items ? tab[i + 1].tile0 : total_tiles) - tab[i].tile0;
      // tile boundary at or after element offset x of this item: ceil(x * nt / sz)
      if (!got_lo && lo < acc + sz) { first = tab[i].tile0 + (int)(((lo - acc) * nt + sz - 1) / sz); first_item = i; got_lo = true; }
      if (hi < acc + sz) { last = tab[i].tile0 + (int)(((hi - acc) * nt + sz - 1) / sz); break; }
      acc += sz;
    }
    if (blockIdx.x == gridDim.x - 1) last = total_tiles;
    s_range[0] = first; s_range[1] = last; s_range[2] = first_item;
  }
  __syncthreads();
  const int t_begin = contig ? s_range[0] : (int)blockIdx.x;
  const int t_step = contig ? 1 : (int)gridDim.x;           // (round-robin kept as the A/B arm: MS_CONTIG*)
  const int my_tiles = contig ? max(0, s_range[1] - t_begin)
                              : ((int)blockIdx.x < total_tiles ? (total_tiles - (int)blockIdx.x + t_step - 1) / t_step : 0);
  // forward-only walk of the table: `cur` = the item of the current tile
  int cur = contig ? s_range[2] : 0;
  while (cur > 0 && tab[cur].tile0 > t_begin) --cur;      // (a boundary that rounded up into the next item)
  int next_t0 = cur + 1 < n_items ? tab[cur + 1].tile0 : 0x7fffffff;
  auto seek = [&](int tile) {
    while (tile >= next_t0) {
      ++cur;
      next_t0 = cur + 1 < n_items ? tab[cur + 1].tile0 : 0x7fffffff;
    }
  };

  if (warp == G * (ms::GT / 32)) {
    // ------------------------------- producer warp -------------------------------
    const uint64_t pol_stream = policy_evict_first(), pol_keep = policy_evict_last();
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = t_begin + it * t_step;
      const int st = it % ns;
      seek(tile);
      const MergeItem im = tab[cur];
      const ms::TileGeo g = ms::tile_geo(im, tile - im.tile0, (int)sizeof(TW));
      const bool up_ok = ms::up_by_copy<TL>(im, g);
      const uint32_t w_bytes = (uint32_t)g.rows * (uint32_t)g.row_bytes;
      const uint32_t u_bytes = (uint32_t)(g.rows * im.r) * (uint32_t)sizeof(TL);
      if (it >= ns) mbar_wait(&empty_bar[st], ((it / ns) & 1) ^ 1, 40);
      const uint32_t sbase = smem_u32(smem + (size_t)st * STAGE);
      const long long ld_bytes = (long long)im.K * (long long)sizeof(TW);
      const unsigned char* wsrc = static_cast<const unsigned char*>(im.Win) + g.n0 * ld_bytes + g.col0_bytes;
      if (lane == 0) mbar_arrive_expect_tx(&full_bar[st], w_bytes + (up_ok ? u_bytes : 0u));
      __syncwarp();
      if (im.tiles_k == 1) {
        if (lane == 0) bulk_load_1d(sbase, wsrc, w_bytes, &full_bar[st], pol_stream);
      } else if (lane < g.rows) {       // column block: one copy per row (<= 8), rows dense in the stage
        bulk_load_1d(sbase + lane * g.row_bytes, wsrc + lane * ld_bytes, g.row_bytes, &full_bar[st], pol_stream);
      }
      if (up_ok && lane == 31)
        bulk_load_1d(sbase + ms::W_BYTES, static_cast<const TL*>(im.up) + (long long)g.n0 * im.r, u_bytes, &full_bar[st], pol_keep);
    }
    return;
  }

  // --------------------------------- consumer groups ---------------------------------
  const int grp = warp / (ms::GT / 32), t = threadIdx.x % ms::GT;
  uint32_t alpha2 = 0;        // != 0: alpha is representable in the weight dtype, packed multiply (one code path: a
                              // warp-uniform branch per packed word instead of three copies of the row loop)
  if constexpr (!std::is_same<TW, float>::value) {
    if (to_f32(from_f32<TW>(alpha)) == alpha) alpha2 = pack2<TW>(alpha, alpha);
  }
  for (int it = grp; it < my_tiles; it += G) {
    const int tile = t_begin + it * t_step;
    const int st = it % ns;
    seek(tile);
    const MergeItem im = tab[cur];
    const ms::TileGeo g = ms::tile_geo(im, tile - im.tile0, (int)sizeof(TW));
    const bool up_ok = ms::up_by_copy<TL>(im, g);
    const int r = im.r, K = im.K;
    const int kv = g.row_bytes / 16;                  // 16-byte columns of the tile (<= 256)
    const unsigned char* base = smem + (size_t)st * STAGE;
    const TL* down = static_cast<const TL*>(im.down) + g.col0_bytes / (int)sizeof(TW);
    const bool d_al = ((reinterpret_cast<uintptr_t>(down) | ((uintptr_t)K * sizeof(TL))) & (VEC * sizeof(TL) - 1)) == 0;
    const TL* us_s = reinterpret_cast<const TL*>(base + ms::W_BYTES);
    const TL* us_g = static_cast<const TL*>(im.up) + (long long)g.n0 * r;
    unsigned char* out = static_cast<unsigned char*>(im.Wout) + (long long)g.n0 * K * (long long)sizeof(TW) + g.col0_bytes;
    const long long ld_bytes = (long long)K * (long long)sizeof(TW);
    bool waited = false;
    // Column passes: the tile's kv 16-byte columns are cut into chunks whose width w is a power of two <= 128 (80 =
    // 64 + 16, 160 = 128 + 32, ...); in each pass ALL 128 threads work: thread -> (column c0 + t % w, row lane t / w),
    // 128 / w row lanes side by side.  A thread's r x VEC slice of `down` is fixed for the pass (registers).
    for (int c0 = 0; c0 < kv;) {
      const int w = 1 << (31 - __clz(min(kv - c0, ms::GT)));
      const int tyc = ms::GT / w, c = c0 + (t & (w - 1)), ty = t / w;
      c0 += w;
      float d[JR][VEC];
#pragma unroll
      for (int j = 0; j < JR; ++j) {
        if (j < r) ld_row<TL, VEC>(down + (long long)j * K + c * VEC, d_al, d[j]);
        else {
#pragma unroll
          for (int v = 0; v < VEC; ++v) d[j][v] = 0.f;
        }
      }
      if (!waited) { mbar_wait(&full_bar[st], (it / ns) & 1, 41); waited = true; }
      const unsigned char* wcol = base + c * 16;
      unsigned char* ocol = out + c * 16;
      auto rows = [&](auto mode_c, auto up_c) {
        constexpr int MODE = decltype(mode_c)::value;
        constexpr int UP = decltype(up_c)::value;      // 0: JR-wide rows of `up` in the stage (vector LDS), 1: any r, stage or global
        constexpr int UB = JR * (int)sizeof(TL);       // bytes of one JR-wide row of `up`
        constexpr int UW = UB / 4;                     // ... in 32-bit words
        constexpr int RU = JR > 4 ? 2 : (std::is_same<TW, float>::value ? MS_RU32 : MS_RU16);   // rows in flight: all their shared-memory loads are issued first
        for (int l0 = ty; l0 < g.rows; l0 += RU * tyc) {
          uint4 raw[RU];
          uint32_t uw[RU][UW];
#pragma unroll
          for (int q = 0; q < RU; ++q) {
            const int ln = l0 + q * tyc;
            if (ln < g.rows) {
              raw[q] = *reinterpret_cast<const uint4*>(wcol + ln * g.row_bytes);
              if constexpr (UP == 0) {
                const unsigned char* src = reinterpret_cast<const unsigned char*>(us_s) + ln * UB;
                if constexpr (UB == 8) {
                  const uint2 v2 = *reinterpret_cast<const uint2*>(src);
                  uw[q][0] = v2.x; uw[q][1] = v2.y;
                } else {
#pragma unroll
                  for (int i = 0; i < UB / 16; ++i) {
                    const uint4 v4 = *(reinterpret_cast<const uint4*>(src) + i);
                    uw[q][4 * i] = v4.x; uw[q][4 * i + 1] = v4.y; uw[q][4 * i + 2] = v4.z; uw[q][4 * i + 3] = v4.w;
                  }
                }
              }
            }
          }
#pragma unroll
          for (int q = 0; q < RU; ++q) {
            const int ln = l0 + q * tyc;
            if (ln >= g.rows) continue;
            float u[JR];
            if constexpr (UP == 0) {
              if constexpr (std::is_same<TL, float>::value) {
#pragma unroll
                for (int j = 0; j < JR; ++j) u[j] = __uint_as_float(uw[q][j]);
              } else {
#pragma unroll
                for (int j = 0; j < JR; j += 2) {
                  const float2 f = unpack2<TL>(uw[q][j / 2]);
                  u[j] = f.x; u[j + 1] = f.y;
                }
              }
            } else {
              const TL* us = up_ok ? us_s : us_g;       // generic pointer: the stage or global memory
#pragma unroll
              for (int j = 0; j < JR; ++j) u[j] = (j < r) ? to_f32(us[ln * r + j]) : 0.f;
            }
            float2 p2[VEC / 2];
#pragma unroll
            for (int v = 0; v < VEC / 2; ++v) p2[v] = make_float2(0.f, 0.f);
#pragma unroll
            for (int j = 0; j < JR; ++j) {
              const float2 uj = make_float2(u[j], u[j]);
#pragma unroll
              for (int v = 0; v < VEC / 2; ++v) p2[v] = ffma2(uj, make_float2(d[j][2 * v], d[j][2 * v + 1]), p2[v]);
            }
            float pr[VEC];
#pragma unroll
            for (int v = 0; v < VEC / 2; ++v) { pr[2 * v] = p2[v].x; pr[2 * v + 1] = p2[v].y; }
            __stcs(reinterpret_cast<uint4*>(ocol + ln * ld_bytes), Apply<TW, TL, MODE>::run(raw[q], pr, alpha, alpha2));
          }
        }
      };
      auto by_up = [&](auto mode_c) {
        if (up_ok && r == JR) rows(mode_c, std::integral_constant<int, 0>{});
        else rows(mode_c, std::integral_constant<int, 1>{});
      };
#if MS_MODE16
      if constexpr (std::is_same<TW, float>::value) by_up(std::integral_constant<int, ALPHA_RUNTIME>{});
      else if (alpha == 1.0f) by_up(std::integral_constant<int, ALPHA_ONE>{});
      else if (alpha2 != 0u) by_up(std::integral_constant<int, ALPHA_PACKED>{});
      else by_up(std::integral_constant<int, ALPHA_F32>{});
#else
      by_up(std::integral_constant<int, ALPHA_RUNTIME>{});
#endif
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty_bar[st]);     // this warp has read everything it needs from the stage
  }
}

template <typename TW, typename TL, int JR>
static int stream_stage_bytes() { return ms::W_BYTES + ms::MAX_ROWS * JR * (int)sizeof(TL); }

static bool merge_stream_enabled() {
  static const int on = [] { const char* e = getenv("B200LORA_MERGE_STREAM"); return e ? atoi(e) : 1; }();
  return on != 0;
}
// which kernel a table is laid out for; plan and run evaluate the same predicate
static bool merge_uses_stream(int max_r) { return max_r <= 8 && merge_stream_enabled(); }

// host: fill tile0 / tiles_k / pad of `items` (host copy), return the total tile count.
//   register-staged kernel: [32 rows x 512 bytes] tiles, tiles_k = tiles along K;
//   streamed kernel: pad = rows per tile, tiles_k = column blocks per row (1 = whole rows; ms::tile_geo).
long long merge_batch_plan(MergeItem* items, int n, int w_dtype) {
  const int es = dtype_size(w_dtype);
  const int vec = 16 / es;
  int max_r = 0;
  for (int i = 0; i < n; ++i) if (items[i].r > max_r) max_r = items[i].r;
  const bool stream = merge_uses_stream(max_r);
  long long t = 0;
  for (int i = 0; i < n; ++i) {
    items[i].tile0 = (int)t;
    if (!stream) {
      const int kvecs = items[i].K / vec;
      items[i].tiles_k = (kvecs + mb::TX - 1) / mb::TX;
      items[i].pad = 0;
      t += (long long)items[i].tiles_k * ((items[i].N + mb::ROWS - 1) / mb::ROWS);
      continue;
    }
    const long long row_bytes = (long long)items[i].K * es;
    if (row_bytes <= ms::BLOCK_BYTES) {           // whole rows: R >= 8 of them, one contiguous span
      int R = (int)(ms::W_BYTES / row_bytes);
      if (R > ms::MAX_ROWS) R = ms::MAX_ROWS;
      R &= ~3;
      items[i].tiles_k = 1;
      items[i].pad = R;
      t += (items[i].N + R - 1) / R;
    } else {                                      // equal column blocks of <= 4 KB, 8 rows each
      const int blocks = (int)((row_bytes + ms::BLOCK_BYTES - 1) / ms::BLOCK_BYTES);
      items[i].tiles_k = blocks;
      items[i].pad = ms::BLOCK_ROWS;
      t += (long long)((items[i].N + ms::BLOCK_ROWS - 1) / ms::BLOCK_ROWS) * blocks;
    }
  }
  return t;
}

template <typename TW, typename TL, int JR>
static cudaError_t launch_stream(const MergeItem* items_dev, int n, int total_tiles, float alpha, int sms,
                                 cudaStream_t stream) {
  constexpr bool W32 = std::is_same<TW, float>::value;
  constexpr int G = W32 ? MS_G32 : (JR > 4 ? 3 : MS_G16);       // (ranks 5..8: 64 registers of `down` per thread, 3 groups)
  const int stage = stream_stage_bytes<TW, TL, JR>();
  int ns = (216 * 1024) / stage;                  // one CTA per SM: the ring fills what the table mirror leaves of the 227 KB
  if (ns > ms::MAX_NS) ns = ms::MAX_NS;
  ns -= ns % G;                                   // a stage belongs to one group (see the kernel)
  if (ns < G) return cudaErrorInvalidConfiguration;
  static const int contig = [] {
    const char* e = getenv("B200LORA_MERGE_CONTIG");
    return e ? atoi(e) : (W32 ? MS_CONTIG32 : MS_CONTIG16);
  }();
  auto* kern = lora_merge_stream_kernel<TW, TL, JR, G>;
  static int attr_dev = -1;                       // per instantiation
  int dev = 0;
  cudaGetDevice(&dev);
  if (attr_dev != dev) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, ns * stage);
    if (e != cudaSuccess) return e;
    attr_dev = dev;
  }
  int grid = sms;
  if (grid > total_tiles) grid = total_tiles;
  if (grid < 1) grid = 1;
  kern<<<grid, G * ms::GT + 32, ns * stage, stream>>>(items_dev, n, total_tiles, alpha, ns, contig);
  return cudaGetLastError();
}

template <typename TW, typename TL>
static cudaError_t launch_merge_batch_t(const MergeItem* items_dev, int n, int total_tiles, int max_r, float alpha,
                                        int sms, cudaStream_t stream) {
  if (merge_uses_stream(max_r)) {     // the bulk-copy ring; B200LORA_MERGE_STREAM=0 keeps the register-staged kernel (A/B)
    if (max_r <= 4) return launch_stream<TW, TL, 4>(items_dev, n, total_tiles, alpha, sms, stream);
    return launch_stream<TW, TL, 8>(items_dev, n, total_tiles, alpha, sms, stream);
  }
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
