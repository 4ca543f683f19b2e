// Backward of the GEGLU activation that b200lora_linear_fwd_geglu applies in the GEMM epilogue:
//   out = hidden * gelu(gate),  Y = [hidden | gate] [M, 2*d_ff]  ->  dY = [dout * gelu(gate) | dout * hidden * gelu'(gate)]
// One pass: reads dout and Y once, writes dY once (stock PyTorch runs mul-backward x2 and gelu-backward: three launches
// and two [M, d_ff] temporaries).  Roundings follow that composition: gelu(gate) and dout*hidden are tensors of the io
// dtype there, so they are rounded here too.  HBM-bound elementwise work, 16-byte vectors.
#include "common.cuh"

namespace b200 {

template <typename T>
__global__ void __launch_bounds__(256) geglu_bwd_kernel(const T* __restrict__ dout, const T* __restrict__ Y,
                                                        T* __restrict__ dY, long long M, int d_ff) {
  constexpr int VEC = 16 / (int)sizeof(T);
  const int vpr = d_ff / VEC;                                    // vectors per row (d_ff % VEC == 0, host-checked)
  const long long total = M * vpr;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / vpr;
    const int j = (int)(i - m * vpr) * VEC;
    const uint4 qd = __ldcs(reinterpret_cast<const uint4*>(dout + m * d_ff + j));
    const uint4 qh = __ldcs(reinterpret_cast<const uint4*>(Y + m * 2 * d_ff + j));
    const uint4 qg = __ldcs(reinterpret_cast<const uint4*>(Y + m * 2 * d_ff + d_ff + j));
    const T* pd = reinterpret_cast<const T*>(&qd);
    const T* ph = reinterpret_cast<const T*>(&qh);
    const T* pg = reinterpret_cast<const T*>(&qg);
    uint4 oh, og;
    T* rh = reinterpret_cast<T*>(&oh);
    T* rg = reinterpret_cast<T*>(&og);
#pragma unroll
    for (int v = 0; v < VEC; ++v) {
      const float d = to_f32(pd[v]), h = to_f32(ph[v]), g = to_f32(pg[v]);
      const float cdf = 0.5f * (1.0f + erff(g * 0.70710678118654752f));
      const float pdf = 0.3989422804014327f * __expf(-0.5f * g * g);
      const float ge = to_f32(from_f32<T>(g * cdf));             // gelu(gate), a tensor of the io dtype in the reference
      const float dge = to_f32(from_f32<T>(d * h));              // grad of gelu(gate): dout * hidden, likewise
      rh[v] = from_f32<T>(d * ge);
      rg[v] = from_f32<T>(dge * (cdf + g * pdf));
    }
    __stcs(reinterpret_cast<uint4*>(dY + m * 2 * d_ff + j), oh);
    __stcs(reinterpret_cast<uint4*>(dY + m * 2 * d_ff + d_ff + j), og);
  }
}

cudaError_t launch_geglu_bwd(const void* dout, const void* Y, void* dY, long long M, int d_ff, int io_dtype, int sms,
                             cudaStream_t stream) {
  const int vec = 16 / dtype_size(io_dtype);
  const long long total = M * (d_ff / vec);
  long long blocks = (total + 255) / 256;
  const long long cap = (long long)sms * 16;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (io_dtype == B200LORA_F32)
    geglu_bwd_kernel<float><<<(int)blocks, 256, 0, stream>>>(static_cast<const float*>(dout), static_cast<const float*>(Y),
                                                              static_cast<float*>(dY), M, d_ff);
  else if (io_dtype == B200LORA_F16)
    geglu_bwd_kernel<__half><<<(int)blocks, 256, 0, stream>>>(static_cast<const __half*>(dout), static_cast<const __half*>(Y),
                                                               static_cast<__half*>(dY), M, d_ff);
  else
    geglu_bwd_kernel<__nv_bfloat16><<<(int)blocks, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(dout),
                                                                      static_cast<const __nv_bfloat16*>(Y),
                                                                      static_cast<__nv_bfloat16*>(dY), M, d_ff);
  return cudaGetLastError();
}

}  // namespace b200
