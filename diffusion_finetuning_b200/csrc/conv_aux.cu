// Helpers of the LoRA-Conv2d implicit GEMM (no counterpart in the reference
// snapshot: upstream cloneofsimo/lora `LoraInjectedConv2d`).  The GEMM itself is the CTA-pair kernel
// of linear_tc2.cu with TMA-im2col loads; this file holds the bandwidth-bound side kernels:
//   repack   OIHW weights -> [rows, taps x Cp] GEMM operand (forward: KRSC; backward: flipped/transposed)
//   pack     lora_down / lora_up / bias -> rank rows, tail block, and the A-columns of the backward operand
//   zcat     backward: Z = [dY | s·dT | 0] along channels and dT = dY·B (one pass over dY)
//   (dA      backward: the im2col mode of the token-reduction kernel, reduce.cu)
#include <string.h>

#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

bool pdl_enabled();   // linear_tc2.cu

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
// Tensor-pipe variant (the one that runs): a [128 pixels x 64 channels] box of dY lands in smem by
// TMA, goes straight back out to Z by a TMA store (the copy costs no thread instructions), and is
// contracted against Bᵀ (smem, bf16) with ldmatrix + mma.sync into dT; one CTA owns whole pixel rows,
// so dT needs no atomics.  HBM-bound: reads dY once, writes Z once.
namespace zc {
constexpr int TOK = 128, COLS = 64, STAGE_BYTES = TOK * COLS * 2, NS = 6, CONSUMERS = 8;
constexpr int LAG = 2;   // warp 0 releases a stage LAG iterations late: its TMA store must have read it
constexpr int NT = (CONSUMERS + 1) * 32;
}

template <typename TIO>
__global__ void __launch_bounds__(zc::NT)
conv_zcat_tc_kernel(const __grid_constant__ CUtensorMap tmDY, const __grid_constant__ CUtensorMap tmZ,
                    const void* __restrict__ B, int ab_dtype, float scale, TIO* __restrict__ Z,
                    TIO* __restrict__ dT, long long M, int Cout, int Coutp, int Cz, int r) {
  extern __shared__ uint8_t zc_smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[zc::NS];
  __shared__ __align__(8) uint64_t empty_bar[zc::NS];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(zc_smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int ldb = Coutp + 8;                                       // +16 bytes: conflict-free B fragments
  TIO* Bt = reinterpret_cast<TIO*>(smem + zc::NS * zc::STAGE_BYTES);   // [16][ldb]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ncb = Coutp / zc::COLS;
  const int nrb = (int)((M + zc::TOK - 1) / zc::TOK);

  pdl_launch_dependents();
  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < zc::NS; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], zc::CONSUMERS); }
    fence_mbar_init();
    tma_prefetch_desc(&tmDY);
    tma_prefetch_desc(&tmZ);
  }
  pdl_wait();
  // Bᵀ [16][ldb] in the io dtype (rounded as autocast would round B): zero fill, then a coalesced
  // 16-byte sweep over B [Cout, r] scattered into place
  for (int i = threadIdx.x; i < 16 * ldb / 8; i += zc::NT)
    reinterpret_cast<uint4*>(Bt)[i] = make_uint4(0, 0, 0, 0);
  __syncthreads();
  {
    const int esz = ab_dtype == B200LORA_F32 ? 4 : 2, per = 16 / esz;
    const int total = Cout * r;
    if ((reinterpret_cast<uintptr_t>(B) & 15) == 0) {
      for (int v = threadIdx.x; v * per < total; v += zc::NT) {
        const uint4 raw = __ldg(reinterpret_cast<const uint4*>(B) + v);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int e = v * per + q;
          if (q < per && e < total) {
            float f;
            if (ab_dtype == B200LORA_F32) f = reinterpret_cast<const float*>(&raw)[q & 3];
            else if (ab_dtype == B200LORA_BF16) f = __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(&raw)[q]);
            else f = __half2float(reinterpret_cast<const __half*>(&raw)[q]);
            Bt[(e % r) * ldb + e / r] = from_f32<TIO>(f);
          }
        }
      }
    } else {
      for (int e = threadIdx.x; e < total; e += zc::NT)
        Bt[(e % r) * ldb + e / r] = from_f32<TIO>(load_as_float(B, ab_dtype, e));
    }
  }
  __syncthreads();

  if (warp == zc::CONSUMERS) {
    const uint64_t pol = policy_evict_first();
    int it = 0;
    for (int rb = blockIdx.x; rb < nrb; rb += gridDim.x)
      for (int cb = 0; cb < ncb; ++cb, ++it) {
        const int s = it % zc::NS;
        if (it >= zc::NS) mbar_wait(&empty_bar[s], ((it / zc::NS) - 1) & 1, 50);
        if (elect_one()) {
          mbar_arrive_expect_tx(&full_bar[s], zc::STAGE_BYTES);
          tma_load_2d(smem + s * zc::STAGE_BYTES, &tmDY, &full_bar[s], cb * zc::COLS, rb * zc::TOK, pol);
        }
        __syncwarp();
      }
    return;
  }

  const int gq = lane >> 2, tq = lane & 3;
  const int mi = lane >> 3, rr = lane & 7;
  const uint32_t row_off = (uint32_t)((warp * 16 + rr + ((mi & 1) << 3)) * 128);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t* Bt32 = reinterpret_cast<const uint32_t*>(Bt);
  int it = 0;
  for (int rb = blockIdx.x; rb < nrb; rb += gridDim.x) {
    float acc[2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[nt][i] = 0.f;
    for (int cb = 0; cb < ncb; ++cb, ++it) {
      const int s = it % zc::NS;
      mbar_wait(&full_bar[s], (it / zc::NS) & 1, 51);
      if (warp == 0 && lane == 0) {       // the copy: same tile, straight back out to Z[:, cb*64 ..]
        fence_proxy_async_smem();
        tma_store_2d(&tmZ, smem + s * zc::STAGE_BYTES, cb * zc::COLS, rb * zc::TOK);
        tma_store_commit();
      }
      const uint32_t st_addr = smem_base + s * zc::STAGE_BYTES + row_off;
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        uint32_t a[4];
        ldmatrix_x4(a, st_addr + ((uint32_t)((2 * ks + (mi >> 1)) ^ rr) << 4));
        const int kc = (cb * zc::COLS + ks * 16 + 2 * tq) >> 1;      // uint32 index of the pair
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const uint32_t* brow = Bt32 + ((gq + 8 * nt) * ldb >> 1);
          mma_16816<TIO>(acc[nt], a, brow[kc], brow[kc + 4]);
        }
      }
      __syncwarp();
      if (lane == 0) {
        if (warp != 0) {
          mbar_arrive(&empty_bar[s]);
        } else if (it >= zc::LAG) {      // stores up to iteration it - LAG have finished reading smem
          tma_store_wait_read<zc::LAG>();
          mbar_arrive(&empty_bar[(it - zc::LAG) % zc::NS]);
        }
      }
    }
    // dT (compact, io dtype) and the 64-channel LoRA block of Z: [s·dT | 0]
    const long long m0 = (long long)rb * zc::TOK + warp * 16;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const long long m = m0 + gq + 8 * h;
      if (m < M) {
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          const int j = 8 * nt + 2 * tq;
          const TIO t0 = from_f32<TIO>(acc[nt][2 * h]), t1 = from_f32<TIO>(acc[nt][2 * h + 1]);
          if (j < r) dT[m * r + j] = t0;
          if (j + 1 < r) dT[m * r + j + 1] = t1;
          TIO zz[2] = {from_f32<TIO>(to_f32(t0) * scale), from_f32<TIO>(to_f32(t1) * scale)};
          *reinterpret_cast<uint32_t*>(Z + m * Cz + Coutp + j) = *reinterpret_cast<uint32_t*>(zz);
        }
      }
    }
    for (int idx = lane; idx < 16 * 6; idx += 32) {
      const long long m = m0 + idx / 6;
      if (m < M) *reinterpret_cast<uint4*>(Z + m * Cz + Coutp + 16 + (idx % 6) * 8) = make_uint4(0, 0, 0, 0);
    }
  }
  if (warp == 0 && lane == 0) tma_store_wait_all<0>();   // (the last LAG stages are never reloaded)
}

template <typename TIO>
cudaError_t launch_conv_zcat_tc(const CUtensorMap* tmDY, const CUtensorMap* tmZ, const void* B, int ab_dtype,
                                float scale, void* Z, void* dT, long long M, int Cout, int Coutp, int Cz, int r,
                                int sms, cudaStream_t stream) {
  const int smem = zc::NS * zc::STAGE_BYTES + 16 * (Coutp + 8) * 2 + 1024;
  auto kern = conv_zcat_tc_kernel<TIO>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  if (e != cudaSuccess) return e;
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  const long long nrb = (M + zc::TOK - 1) / zc::TOK;
  const int per_sm = smem <= 110 * 1024 ? 2 : 1;   // blocks that fit an SM's shared memory
  const int grid = (int)(nrb < (long long)sms * per_sm ? nrb : (long long)sms * per_sm);
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(zc::NT, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, *tmDY, *tmZ, B, ab_dtype, scale, (TIO*)Z, (TIO*)dT, M, Cout, Coutp, Cz, r);
}
template cudaError_t launch_conv_zcat_tc<__nv_bfloat16>(const CUtensorMap*, const CUtensorMap*, const void*, int,
                                                        float, void*, void*, long long, int, int, int, int, int,
                                                        cudaStream_t);
template cudaError_t launch_conv_zcat_tc<__half>(const CUtensorMap*, const CUtensorMap*, const void*, int, float,
                                                 void*, void*, long long, int, int, int, int, int, cudaStream_t);

// backward dA[j, c, kh, kw] += s · Σ_m dT[m, j] · X[pixel(m) + tap, c] lives in reduce.cu: the token
// reduction kernel with TMA im2col loads (launch_conv_grad_a_tc).  A one-pass variant (halo tile in
// smem, nine taps per load) was measured and dropped: at r = 16 both are bound by the mma.sync issue
// rate (9·M·C·r MACs), not by the 9x re-read of X out of L2 (63 vs 57 us at 320 channels, 64x64).

}  // namespace b200
