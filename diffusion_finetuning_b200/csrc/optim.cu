// Optimizer step over the flat LoRA arenas: global-norm clip + AdamW for ALL LoRA
// parameters in two launches over contiguous fp32 memory, instead of torch's clip_grad_norm_
// (norm per tensor, stack, norm, clamp, scale per tensor) + multi-tensor AdamW over ~250 small tensors.
// Semantics follow torch.nn.utils.clip_grad_norm_ (train_lora_dreambooth.py:878-884, cli_lora_pti.py:448)
// and torch.optim.AdamW (train_lora_dreambooth.py:676): decoupled weight decay, bias correction by
// the step count, eps added to sqrt(v)/sqrt(1-b2^t).  The step count lives on the device so that the
// whole update is CUDA-graph capturable (kernel arguments are frozen at capture).
// HBM-bound: reads g twice, p/m/v once, writes p/m/v/g once — 28 bytes per parameter.
#include "common.cuh"

namespace b200 {

__global__ void __launch_bounds__(256) arena_sqnorm_kernel(const float* __restrict__ g, long long n,
                                                           float* __restrict__ scratch) {
  // scratch[0] += sum g^2 (zeroed by the caller); scratch[1] = step count, advanced here once
  __shared__ float ws[8];
  float acc = 0.f;
  const long long n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = __ldg(g4 + i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (int)(n & 3)) {
    const float v = g[(n4 << 2) + threadIdx.x];
    acc += v * v;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += ws[w];
    atomicAdd(scratch, s);
    if (blockIdx.x == 0) scratch[1] += 1.0f;
  }
}

__global__ void __launch_bounds__(256) arena_adamw_kernel(float* __restrict__ p, float* __restrict__ g,
                                                          float* __restrict__ m, float* __restrict__ v,
                                                          long long n, float lr, float b1, float b2, float eps,
                                                          float wd, float max_norm,
                                                          const float* __restrict__ scratch) {
  const float step = scratch[1];
  float coef = 1.f;
  if (max_norm > 0.f) {
    const float total = sqrtf(scratch[0]);
    coef = fminf(max_norm / (total + 1e-6f), 1.f);
  }
  const float bc1 = 1.f - powf(b1, step);
  const float bc2_sqrt = sqrtf(1.f - powf(b2, step));
  const float step_size = lr / bc1;
  const float decay = 1.f - lr * wd;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gi = g[i] * coef;
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    p[i] = p[i] * decay - step_size * (mi / denom);
    m[i] = mi;
    v[i] = vi;
    g[i] = gi;          // clip_grad_norm_ scales the gradients in place
  }
}

cudaError_t launch_arena_adamw(float* p, float* g, float* m, float* v, long long n, float lr, float b1,
                               float b2, float eps, float wd, float max_norm, float* scratch,
                               cudaStream_t stream) {
  long long want = (n + 1023) / 1024;
  const int blocks = (int)(want < 148 * 8 ? (want < 1 ? 1 : want) : 148 * 8);
  cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(float), stream);
  if (e != cudaSuccess) return e;
  arena_sqnorm_kernel<<<blocks, 256, 0, stream>>>(g, n, scratch);
  arena_adamw_kernel<<<blocks, 256, 0, stream>>>(p, g, m, v, n, lr, b1, b2, eps, wd, max_norm, scratch);
  return cudaGetLastError();
}

}  // namespace b200
