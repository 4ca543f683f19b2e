// Flat element-wise pass over LoRA artefacts: out = round(alpha·a) (+ round(beta·b)).
//   * b == nullptr, alpha == 1: dtype conversion of a whole parameter arena in one launch — the `.pt` writer
//     (lora.py:201-213 casts 2·L tensors to fp16 and copies them to the host one by one);
//   * b != nullptr: the `lpl` mix of two LoRA lists, cli_lora_add.py:49-58:  alpha·x1 + (1-alpha)·x2,
//     evaluated as ATen does on tensors of the file's dtype — each product is rounded to that dtype, then the
//     sum is rounded (no fused multiply-add), so the result is bit-identical to the reference's CPU result.
// HBM-bound, 16-byte accesses when the pointers allow it; the payloads are a few MB.
#include "common.cuh"

namespace b200 {

template <typename TI, typename TO>
__global__ void __launch_bounds__(256) flat_axpby_kernel(const TI* __restrict__ a, const TI* __restrict__ b,
                                                         TO* __restrict__ out, long long n, float alpha,
                                                         float beta) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n;
       i += (long long)gridDim.x * blockDim.x) {
    // products in fp32 (ATen's opmath for 16-bit tensors), rounded to the INPUT dtype like the reference's
    // temporaries; __fmul_rn / __fadd_rn keep the compiler from contracting them into an FMA
    float v = to_f32<TI>(from_f32<TI>(__fmul_rn(to_f32<TI>(__ldg(a + i)), alpha)));
    if (b != nullptr) {
      const float w = to_f32<TI>(from_f32<TI>(__fmul_rn(to_f32<TI>(__ldg(b + i)), beta)));
      v = to_f32<TI>(from_f32<TI>(__fadd_rn(v, w)));
    }
    out[i] = from_f32<TO>(v);
  }
}

template <typename TI>
static cudaError_t flat_dispatch_out(const void* a, const void* b, void* out, long long n, float alpha, float beta,
                                     int out_dtype, int blocks, cudaStream_t stream) {
  const TI* pa = static_cast<const TI*>(a);
  const TI* pb = static_cast<const TI*>(b);
  if (out_dtype == B200LORA_F32)
    flat_axpby_kernel<TI, float><<<blocks, 256, 0, stream>>>(pa, pb, static_cast<float*>(out), n, alpha, beta);
  else if (out_dtype == B200LORA_F16)
    flat_axpby_kernel<TI, __half><<<blocks, 256, 0, stream>>>(pa, pb, static_cast<__half*>(out), n, alpha, beta);
  else
    flat_axpby_kernel<TI, __nv_bfloat16><<<blocks, 256, 0, stream>>>(pa, pb, static_cast<__nv_bfloat16*>(out), n, alpha, beta);
  return cudaGetLastError();
}

cudaError_t launch_flat_axpby(const void* a, const void* b, void* out, long long n, float alpha, float beta,
                              int in_dtype, int out_dtype, cudaStream_t stream) {
  long long want = (n + 1023) / 1024;
  const int blocks = (int)(want < 148 * 8 ? (want < 1 ? 1 : want) : 148 * 8);
  if (in_dtype == B200LORA_F32) return flat_dispatch_out<float>(a, b, out, n, alpha, beta, out_dtype, blocks, stream);
  if (in_dtype == B200LORA_F16) return flat_dispatch_out<__half>(a, b, out, n, alpha, beta, out_dtype, blocks, stream);
  return flat_dispatch_out<__nv_bfloat16>(a, b, out, n, alpha, beta, out_dtype, blocks, stream);
}

}  // namespace b200
