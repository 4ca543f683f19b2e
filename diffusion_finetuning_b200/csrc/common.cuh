// Shared helpers for all b200lora translation units.
#pragma once
#include <cuda.h>  // CUtensorMap (types only; the driver entry point is resolved at run time)
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/b200lora.h"

namespace b200 {

__host__ __device__ inline int dtype_size(int dt) { return dt == B200LORA_F32 ? 4 : 2; }

// generic scalar load of element `idx` of an array of runtime dtype, widened to fp32
__device__ __forceinline__ float load_as_float(const void* p, int dt, long long idx) {
  if (dt == B200LORA_F32) return __ldg(reinterpret_cast<const float*>(p) + idx);
  if (dt == B200LORA_BF16)
    return __bfloat162float(__ldg(reinterpret_cast<const __nv_bfloat16*>(p) + idx));
  return __half2float(__ldg(reinterpret_cast<const __half*>(p) + idx));
}

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) {
  return __float2half_rn(v);
}
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// parameters of the tcgen05 fused GEMM (linear_tc.cu), filled by api.cu
struct TcParams {
  long long M;
  int Kc, Nout, r, BN;
  int num_m_blk, num_n_blk, num_k_blk, stages;
  int nst, acc_stride;        // TMEM accumulator stages and their column stride
  int has_bias, tail_ksteps;  // bias rides on a ones column of the tail MMA; K=16 steps in the tail
  float scale;
  void* T_out;
  unsigned idesc_main, idesc_main_t, idesc_tail;
  // implicit-GEMM convolution (conv != 0): X is NHWC, the k-block index walks (tap, 64-channel
  // block); rows are output pixels in (n, oh, ow) order
  int conv, cv_S, cv_cpb, cv_Ho, cv_Wo;
  int cv_pad_h, cv_pad_w, cv_str_h, cv_str_w, cv_dil_h, cv_dil_w;
  int kseg;                   // != 0: X is split along K into tensors of kseg k-blocks each (<= 3)
  int cluster_scope;          // pair kernel: release.cluster / acquire.cluster on the cross-CTA hand-offs (default 1)
  int w_prefetch;             // pair kernel: request the first stages' W half-tiles before griddepcontrol.wait
  int geglu;                  // pair kernel, != 0: W is [2*geglu, K] = [hidden ; gate] (GEGLU projection); a tile covers BN/2
                              // hidden columns and the BN/2 gate columns of the same outputs, the epilogue also writes
                              // hidden * gelu(gate) [M, geglu] (tmOut); Nout = 2*geglu, num_n_blk counts BN/2-wide blocks
  int geglu_store_y;          // ... and still stores the pre-activation Y [M, 2*geglu] (needed by the backward)
  int out_f32;                // pair kernel: Y is an fp32 tensor (tmY: 16-column boxes): the accumulator is stored unrounded
                              // (backward dX of a layer whose input is fp32 under autocast: no bf16 -> fp32 cast launch)
  int split_b;                // pair kernel: a CTA whose B half is two TMA boxes per k-block issues the second one from
                              // warp 2 (third producer thread) instead of back to back on warp 3
  long long* dbg;             // optional clock64 timeline of CTA `dbg_cta` (tools/timeline.py)
  int dbg_cta;
};

// ---- error plumbing (api.cu owns the storage) ----
void set_error(const char* fmt, ...);

}  // namespace b200
