// Textual-inversion step beside the LoRA path: the placeholder-token rows of the CLIP
// token embedding are the only rows that train (cli_lora_pti.py:280, 309-346).  The reference runs a
// DENSE AdamW over the whole [49410, 768] table (152 MB of gradient, 3 x 152 MB of optimizer traffic),
// re-normalises the placeholder rows and then copies every other row back from a 152 MB clone.  Here
// the work is what actually changes: P rows of D elements.
//   ti_grad_rows_kernel : g[p, :] += sum over tokens t with ids[t] == rows[p] of dY[t, :]
//                         (the only rows of the embedding gradient that survive cli_lora_pti.py:344-346)
//   ti_step_kernel      : AdamW on the fp32 master rows (torch.optim.AdamW semantics, cli_lora_pti.py:651-657),
//                         norm decay  w <- normalize(w) * (|w| + lambda * (target - |w|))  (cli_lora_pti.py:317-335),
//                         write-through of the rows into the embedding table in the table's dtype.
// Rows that are not placeholders are never touched, which is what the reference's restore step amounts to.
// Both kernels are latency-bound (a few KB); one CTA per (row, column chunk) / per row.
#include "common.cuh"

namespace b200 {

template <typename TIO>
__global__ void __launch_bounds__(256) ti_grad_rows_kernel(const TIO* __restrict__ dY,
                                                           const long long* __restrict__ ids, long long n_tok,
                                                           int D, const long long* __restrict__ rows,
                                                           float* __restrict__ g) {
  // blockIdx.y = placeholder row, blockIdx.x = chunk of 256 columns.  Tokens are visited in index order
  // by every thread, so the sum is reproducible run to run (no atomics).
  __shared__ int hits[256];
  __shared__ int n_hits;
  const int p = blockIdx.y;
  const long long row = rows[p];
  const int d = blockIdx.x * 256 + threadIdx.x;
  float acc = 0.f;
  for (long long base = 0; base < n_tok; base += 256) {
    // one pass of 256 token ids: matching positions are compacted in order through a ballot scan
    if (threadIdx.x == 0) n_hits = 0;
    __syncthreads();
    const long long t = base + threadIdx.x;
    const bool hit = t < n_tok && ids[t] == row;
    const unsigned m = __ballot_sync(0xffffffffu, hit);
    __shared__ int warp_cnt[8];
    if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = __popc(m);
    __syncthreads();
    int off = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); ++w) off += warp_cnt[w];
    if (hit) hits[off + __popc(m & ((1u << (threadIdx.x & 31)) - 1u))] = threadIdx.x;
    if (threadIdx.x == 255) n_hits = off + __popc(m);
    __syncthreads();
    const int nh = n_hits;
    if (d < D)
      for (int i = 0; i < nh; ++i) acc += to_f32<TIO>(dY[(base + hits[i]) * D + d]);
    __syncthreads();
  }
  if (d < D) g[(long long)p * D + d] += acc;
}

__device__ __forceinline__ float block_sum_256(float v, float* ws) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();                      // ws may still be read from a previous call
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) s += ws[w];
  return s;
}

template <typename TW>
__global__ void __launch_bounds__(256) ti_step_kernel(TW* __restrict__ table, int D,
                                                      const long long* __restrict__ rows,
                                                      float* __restrict__ master, const float* __restrict__ g,
                                                      float* __restrict__ m, float* __restrict__ v,
                                                      float step, float lr, float b1, float b2, float eps,
                                                      float wd, float lambda_, float target_norm,
                                                      float* __restrict__ norms_out, int P,
                                                      const float* __restrict__ lr_dev,
                                                      float* __restrict__ step_dev) {
  __shared__ float ws[8];
  const int p = blockIdx.x;
  const long long row = rows[p];
  // Replayable form (the step lives in a captured CUDA graph): the learning rate and the per-row step count
  // are read from device memory, so an LR schedule and the bias correction keep moving under replay.
  if (lr_dev != nullptr) {
    lr = lr_dev[0];
    if (lambda_ >= 0.f) lambda_ = fminf(1.f, 100.f * lr);      // cli_lora_pti.py:318
  }
  if (step_dev != nullptr) {
    step = step_dev[p] + 1.f;
    __syncthreads();
    if (threadIdx.x == 0) step_dev[p] = step;
  }
  const float bc1 = 1.f - powf(b1, step);
  const float bc2_sqrt = sqrtf(1.f - powf(b2, step));
  const float step_size = lr / bc1;
  const float decay = 1.f - lr * wd;
  float* wp = master + (long long)p * D;
  float sq = 0.f;
  for (int d = threadIdx.x; d < D; d += 256) {
    const long long i = (long long)p * D + d;
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    const float w = wp[d] * decay - step_size * (mi / denom);
    m[i] = mi;
    v[i] = vi;
    wp[d] = w;
    sq += w * w;
  }
  const float pre = sqrtf(block_sum_256(sq, ws));
  float mul = 1.f;
  if (lambda_ >= 0.f)     // F.normalize(w) * (pre + lambda * (target - pre)); F.normalize clamps the norm at 1e-12
    mul = (pre + lambda_ * (target_norm - pre)) / fmaxf(pre, 1e-12f);
  float sq2 = 0.f;
  TW* trow = table + row * D;
  for (int d = threadIdx.x; d < D; d += 256) {
    const float w = wp[d] * mul;
    wp[d] = w;
    trow[d] = from_f32<TW>(w);
    sq2 += w * w;
  }
  const float post = sqrtf(block_sum_256(sq2, ws));
  if (norms_out != nullptr && threadIdx.x == 0) {
    norms_out[p] = pre;
    norms_out[P + p] = post;
  }
}

cudaError_t launch_ti_grad_rows(const void* dY, const long long* ids, long long n_tok, int D,
                                const long long* rows, int P, float* g, int io_dtype, cudaStream_t stream) {
  dim3 grid((D + 255) / 256, P, 1);
  if (io_dtype == B200LORA_F32)
    ti_grad_rows_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(dY), ids, n_tok, D, rows, g);
  else if (io_dtype == B200LORA_F16)
    ti_grad_rows_kernel<__half><<<grid, 256, 0, stream>>>(static_cast<const __half*>(dY), ids, n_tok, D, rows, g);
  else
    ti_grad_rows_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(static_cast<const __nv_bfloat16*>(dY), ids, n_tok, D, rows, g);
  return cudaGetLastError();
}

cudaError_t launch_ti_step(void* table, int w_dtype, int D, const long long* rows, int P, float* master,
                           const float* g, float* m, float* v, float step, float lr, float b1, float b2,
                           float eps, float wd, float lambda_, float target_norm, float* norms_out,
                           const float* lr_dev, float* step_dev, cudaStream_t stream) {
#define RUN(TW) ti_step_kernel<TW><<<P, 256, 0, stream>>>(static_cast<TW*>(table), D, rows, master, g, m, v, step, \
                                                          lr, b1, b2, eps, wd, lambda_, target_norm, norms_out, P, \
                                                          lr_dev, step_dev)
  if (w_dtype == B200LORA_F32) RUN(float);
  else if (w_dtype == B200LORA_F16) RUN(__half);
  else RUN(__nv_bfloat16);
#undef RUN
  return cudaGetLastError();
}

}  // namespace b200
