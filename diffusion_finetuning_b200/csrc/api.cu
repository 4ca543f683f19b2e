// extern "C" entry points of libb200lora.so (declared in include/b200lora.h).
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------
// error plumbing
// ---------------------------------------------------------------------------------------
static thread_local char g_err[512] = {0};
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

// forward declarations of the per-file launchers
int tc_pick_bn(long long M, int Kc, int Nout, int rows, int units);
int tc_smem_bytes(int BN, int* stages_out);
void tc_tmem_plan(int BN, int* nst, int* stride);
unsigned int tc_read_watchdog();
template <typename TIO>
cudaError_t launch_tc(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                      const CUtensorMap&, const TcParams&, int, int, cudaStream_t);
int tc2_smem_bytes(int BN, int* stages_out);
template <typename TIO>
cudaError_t launch_tc2(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                       const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                       const CUtensorMap&, const TcParams&, int, cudaStream_t);
cudaError_t launch_conv_repack(const void*, int, void*, int, int, int, int, int, int, int, cudaStream_t);
cudaError_t launch_conv_pack(const void*, const void*, int, const void*, int, void*, void*, void*, int,
                             int, int, int, int, int, int, int, int, cudaStream_t);
cudaError_t launch_conv_zcat(const void*, const void*, int, float, void*, void*, long long, int, int,
                             int, int, int, cudaStream_t);

template <typename TIO>
cudaError_t launch_conv_zcat_tc(const CUtensorMap*, const CUtensorMap*, const void*, int, float, void*, void*,
                                long long, int, int, int, int, int, cudaStream_t);
cudaError_t launch_pack(const void*, const void*, int, const void*, int, void*, void*, void*, void*,
                        int, int, int, int, cudaStream_t, const float*, void*, long long);
cudaError_t launch_pack_seg(const void* const*, const void* const*, const void* const*, const float*, const int*,
                            int, int, int, void*, void*, void*, void*, int, int, int, cudaStream_t, const float*, void*,
                            long long);
template <typename TIO>
cudaError_t launch_simt(const void*, long long, const void*, long long, int, const void*, int,
                        const void*, long long, long long, const void*, long long, long long, int,
                        float, void*, long long, void*, long long, int, int, int, cudaStream_t);
template <typename T>
cudaError_t launch_lora_grads(const void*, const void*, const void*, const void*, float, float*,
                              float*, float*, float*, long long, int, int, int, const CUtensorMap*,
                              const CUtensorMap*, int, cudaStream_t);
int grad_reduce_splits(long long M, int C, int C_other, int esize, int sms);
template <typename T>
cudaError_t launch_conv_grad_a_tc(const CUtensorMap*, const void*, float, float*, const int*, int, int,
                                  cudaStream_t);
template <typename T>
cudaError_t launch_grads_fused_tc(const CUtensorMap* const*, const CUtensorMap*, const void*, const void*,
                                  const float*, float* const*, float* const*, long long, int, int, int,
                                  const int*, int, cudaStream_t);
template <typename TW, typename TL, int MODE>
cudaError_t launch_merge(const void*, void*, const void*, const void*, float, int, int, int,
                         float*, cudaStream_t);
cudaError_t launch_svd_epilogue(float*, const float*, float*, float, int, int, int, int,
                                cudaStream_t);
cudaError_t launch_absmean_any(const void*, const void*, float*, int, int, int, int, cudaStream_t);
struct MergeItem { const void* Win; void* Wout; const void* up; const void* down; int N, K, r, tile0, tiles_k, pad; };
long long merge_batch_plan(MergeItem*, int, int);
cudaError_t launch_merge_batch(const MergeItem*, int, int, int, float, int, int, int, cudaStream_t);
cudaError_t launch_arena_adamw(float*, float*, float*, float*, long long, float, float, float, float, float,
                               float, float*, cudaStream_t);
cudaError_t launch_geglu_bwd(const void*, const void*, void*, long long, int, int, int, cudaStream_t);
cudaError_t launch_flat_axpby(const void*, const void*, void*, long long, float, float, int, int, cudaStream_t);
cudaError_t launch_ti_grad_rows(const void*, const long long*, long long, int, const long long*, int, float*, int,
                                cudaStream_t);
cudaError_t launch_ti_step(void*, int, int, const long long*, int, float*, const float*, float*, float*, float,
                           float, float, float, float, float, float, float, float*, const float*, float*,
                           cudaStream_t);

// ---------------------------------------------------------------------------------------
// device info
// ---------------------------------------------------------------------------------------
struct DevInfo { int ok, major, minor, sms; };
static DevInfo dev_info() {
  static thread_local DevInfo cached[16] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return DevInfo{0, 0, 0, 0};
  if (!cached[dev].ok) {
    cudaDeviceProp pr;
    if (cudaGetDeviceProperties(&pr, dev) != cudaSuccess) return DevInfo{0, 0, 0, 0};
    cached[dev] = DevInfo{1, pr.major, pr.minor, pr.multiProcessorCount};
  }
  return cached[dev];
}

// ---------------------------------------------------------------------------------------
// TMA descriptors: cuTensorMapEncodeTiled resolved through the runtime (no -lcuda, so the
// library also loads on a box without a driver), results cached per (ptr, shape, box).
// ---------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct TmKey {
  const void* ptr; uint64_t d0, d1, ld; uint32_t b0, b1; int dt, sw, dev;
  bool operator==(const TmKey& o) const { return memcmp(this, &o, sizeof(TmKey)) == 0; }
};
struct TmKeyHash {
  size_t operator()(const TmKey& k) const {
    const uint64_t* w = reinterpret_cast<const uint64_t*>(&k);
    size_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(TmKey) / 8; ++i) h = (h ^ w[i]) * 1099511628211ull;
    return h;
  }
};
static std::mutex g_tm_mu;
static std::unordered_map<TmKey, CUtensorMap, TmKeyHash> g_tm_cache;

// The tensor-map encoders are driver entry points and need the primary context CURRENT on the calling
// thread; a thread whose first CUDA work is one of our calls (the autograd backward thread) has only
// had cudaSetDevice, which binds lazily.  One runtime no-op per thread and device does the binding.
static thread_local int g_bound_dev = -1;   // file scope: b200lora_lora_grads_static binds without cudaFree and records it here
static void bind_context() {
  int dev = -1;
  cudaGetDevice(&dev);
  if (dev != g_bound_dev) {
    cudaFree(0);
    g_bound_dev = dev;
  }
}

// 2-D row-major [rows, cols] tensor of dtype `dt` (16-bit operands; fp32 for the fp32-output store map), box [box_rows, box_cols]
static int get_tmap(CUtensorMap* out, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld,
                    uint32_t box_rows, uint32_t box_cols, int dt, CUtensorMapSwizzle sw) {
  TmKey key;
  memset(&key, 0, sizeof(key));
  key.ptr = ptr; key.d0 = cols; key.d1 = rows; key.ld = ld; key.b0 = box_cols; key.b1 = box_rows;
  key.dt = dt; key.sw = (int)sw;
  cudaGetDevice(&key.dev);
  {
    std::lock_guard<std::mutex> g(g_tm_mu);
    auto it = g_tm_cache.find(key);
    if (it != g_tm_cache.end()) { *out = it->second; return 0; }
  }
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return B200LORA_E_CUDA; }
  bind_context();
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstr[1] = {ld * (dt == B200LORA_F32 ? 4ull : 2ull)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType cdt = dt == B200LORA_F32    ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                            : dt == B200LORA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16
                                                  : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUresult rc = fn(out, cdt, 2, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed rc=%d (ptr=%p rows=%llu cols=%llu ld=%llu box=%ux%u)",
              (int)rc, ptr, (unsigned long long)rows, (unsigned long long)cols,
              (unsigned long long)ld, box_rows, box_cols);
    return B200LORA_E_CUDA;
  }
  std::lock_guard<std::mutex> g(g_tm_mu);
  if (g_tm_cache.size() > 8192) g_tm_cache.clear();
  g_tm_cache.emplace(key, *out);
  return 0;
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                   cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeIm2colFn get_im2col_fn() {
  static EncodeIm2colFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &p, cudaEnableDefault, &q) ==
            cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeIm2colFn>(p);
  });
  return fn;
}

// NHWC activations [N,H,W,C] (16-bit) as an im2col-mode map: 128 pixels x 64 channels per load,
// base pixels restricted to [-pad, dim + pad - (filter-1)*dil) so that base + tap offset covers
// exactly the padded window; out-of-image elements are zero-filled by the TMA unit.
static int get_tmap_im2col(CUtensorMap* out, const void* ptr, int N, int H, int W, int C, int R, int S,
                           int str_h, int str_w, int pad_h, int pad_w, int dil_h, int dil_w, int dt) {
  EncodeIm2colFn fn = get_im2col_fn();
  if (!fn) { set_error("cuTensorMapEncodeIm2col entry point not available"); return B200LORA_E_CUDA; }
  bind_context();
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  int lower[2] = {-pad_w, -pad_h};
  int upper[2] = {pad_w - (S - 1) * dil_w, pad_h - (R - 1) * dil_h};
  cuuint32_t estr[4] = {1, (cuuint32_t)str_w, (cuuint32_t)str_h, 1};
  CUtensorMapDataType cdt =
      dt == B200LORA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUresult rc = fn(out, cdt, 4, const_cast<void*>(ptr), gdim, gstr, lower, upper, 64, 128, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeIm2col failed rc=%d (N=%d H=%d W=%d C=%d R=%d S=%d)", (int)rc, N, H, W,
              C, R, S);
    return B200LORA_E_CUDA;
  }
  return 0;
}

// Stride-1 convolutions whose 128-pixel operand tiles are rectangles of the output image (whole rows of one
// image, or whole images): the same tile as a tiled-mode 4-D box [64 channels, Wo, rows, images] of the NHWC
// input, loaded at (c0, tap_w*dil - pad_w, oh0 + tap_h*dil - pad_h, n0) with the halo zero-filled.  Measured
// equal to the im2col mode (1-3 % faster forward on the config-3 shapes): kept as the default
// because it is the plain tiled form the linear layers' X operand uses.
static bool conv_tiles_are_boxes(int Ho, int Wo, int str_h, int str_w, int* box_h, int* box_n) {
  if (str_h != 1 || str_w != 1 || Wo < 1 || Wo > 128 || 128 % Wo) return false;
  const int rows = 128 / Wo;
  if (rows <= Ho) {
    if (Ho % rows) return false;
    *box_h = rows; *box_n = 1;
  } else {
    if (rows % Ho) return false;
    *box_h = Ho; *box_n = rows / Ho;
  }
  return true;
}
static bool conv_tiled_enabled() {   // B200LORA_CONV_TILED=0 keeps the im2col-mode loads (A/B measurements)
  static int v = -1;
  if (v < 0) { const char* e = getenv("B200LORA_CONV_TILED"); v = (e && e[0] == '0') ? 0 : 1; }
  return v != 0;
}
static int get_tmap_conv_box(CUtensorMap* out, const void* ptr, int N, int H, int W, int C, int box_w, int box_h,
                             int box_n, int dt) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return B200LORA_E_CUDA; }
  bind_context();
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_w, (cuuint32_t)box_h, (cuuint32_t)box_n};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMapDataType cdt =
      dt == B200LORA_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  CUresult rc = fn(out, cdt, 4, const_cast<void*>(ptr), gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled (conv box) failed rc=%d (N=%d H=%d W=%d C=%d box=%dx%dx%d)", (int)rc, N, H,
              W, C, box_w, box_h, box_n);
    return B200LORA_E_CUDA;
  }
  return 0;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// can the tcgen05 path take this problem?  (writes the reason to the error slot if not)
static bool tc_applicable(const void* X, long long ldx, const void* W, long long ldw, int w_layout,
                          const void* Y, long long ldy, int r, int io_dtype, int Nout = 0) {
  if (Nout % 8) { set_error("tc path needs Nout %% 8 == 0"); return false; }
  if (io_dtype == B200LORA_F32) { set_error("tc path needs 16-bit io"); return false; }
  if (w_layout != B200LORA_W_NK) { set_error("tc path needs W as [Nout,Kc]"); return false; }
  if (r < 1 || r > 16) { set_error("tc path needs 1 <= r <= 16"); return false; }
  if (!aligned16(X) || !aligned16(W) || !aligned16(Y) || (ldx % 8) || (ldw % 8) || (ldy % 8)) {
    set_error("tc path needs 16-byte aligned pointers and leading dimensions");
    return false;
  }
  DevInfo di = dev_info();
  if (!di.ok || di.major != 10) { set_error("tc path needs an sm_100 device"); return false; }
  return true;
}

// debug timeline target (tools/timeline.py): device buffer of 6*256 int64 and the CTA to trace
static long long* g_dbg_buf = nullptr;
static int g_dbg_cta = 0;

static int run_tc(const void* X, long long ldx, const void* W, long long ldw, bool has_bias,
                  const void* S1_pack, const void* S2_pack, float scale, void* Y, long long ldy,
                  void* T_out, long long M, int Kc, int Nout, int r, int io_dtype,
                  cudaStream_t stream) {
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.Kc = Kc; p.Nout = Nout; p.r = r;
  p.BN = tc_pick_bn(M, Kc, Nout, 128, dev_info().sms);
  p.num_m_blk = (int)((M + 127) / 128);
  p.num_n_blk = (Nout + p.BN - 1) / p.BN;
  p.num_k_blk = (Kc + 63) / 64;
  const int smem = tc_smem_bytes(p.BN, &p.stages);
  if (p.stages < 2) { set_error("tc: not enough shared memory for BN=%d", p.BN); return B200LORA_E_SHAPE; }
  tc_tmem_plan(p.BN, &p.nst, &p.acc_stride);
  p.scale = scale;
  p.has_bias = has_bias ? 1 : 0;
  p.tail_ksteps = (r + p.has_bias > 16) ? 2 : 1;
  p.T_out = T_out;
  p.dbg = g_dbg_buf;
  p.dbg_cta = g_dbg_cta;
  const unsigned fmt = io_dtype == B200LORA_BF16 ? 1u : 0u;
  // kind::f16 idesc: D=f32 (1<<4), A/B format, N>>3 at bit 17, M>>4 at bit 24
  const unsigned base = (1u << 4) | (fmt << 7) | (fmt << 10) | ((128u >> 4) << 24);
  p.idesc_main = base | (((unsigned)p.BN >> 3) << 17);
  p.idesc_main_t = base | (((unsigned)(p.BN + 16) >> 3) << 17);
  p.idesc_tail = p.idesc_main;

  CUtensorMap tmX, tmW, tmS1, tmS2, tmY;
  int rc;
  if ((rc = get_tmap(&tmX, X, (uint64_t)M, (uint64_t)Kc, (uint64_t)ldx, 128, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmW, W, (uint64_t)Nout, (uint64_t)Kc, (uint64_t)ldw, (uint32_t)p.BN, 64,
                     io_dtype, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmS1, S1_pack, 16, (uint64_t)Kc, (uint64_t)Kc, 16, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmS2, S2_pack, (uint64_t)Nout, 64, 64, (uint32_t)p.BN, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmY, Y, (uint64_t)M, (uint64_t)Nout, (uint64_t)ldy, 32, 32, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_64B))) return rc;

  DevInfo di = dev_info();
  cudaError_t e = io_dtype == B200LORA_BF16
                      ? launch_tc<__nv_bfloat16>(tmX, tmW, tmS1, tmS2, tmY, p, smem, di.sms, stream)
                      : launch_tc<__half>(tmX, tmW, tmS1, tmS2, tmY, p, smem, di.sms, stream);
  if (e != cudaSuccess) {
    set_error("tc launch failed: %s", cudaGetErrorString(e));
    return B200LORA_E_CUDA;
  }
  return 0;
}

#ifndef B200LORA_SPLIT_B_DEFAULT
#define B200LORA_SPLIT_B_DEFAULT 0
#endif
// CTA-pair variant (linear_tc2.cu): same math, tcgen05 cta_group::2, 256-row blocks per pair
struct ConvGeoH { int N, H, W, C, Ho, Wo, R, S, str_h, str_w, pad_h, pad_w, dil_h, dil_w; };
// X given as `n` tensors [M, cols] side by side along K (cols % 64 == 0, n <= 3)
struct XSegs { int n; int cols; const void* ptr[3]; long long ld[3]; };
// GEGLU projection (W = [hidden ; gate], Nout = 2*d_ff): also write hidden * gelu(gate) to out [M, d_ff]
struct GegluH { int d_ff; void* out; long long ldo; int store_y; };

static int run_tc2(const void* X, long long ldx, const void* W, long long ldw, bool has_bias,
                   const void* S1_pack, const void* S2_pack, float scale, void* Y, long long ldy,
                   void* T_out, long long M, int Kc, int Nout, int r, int io_dtype,
                   cudaStream_t stream, const ConvGeoH* cg = nullptr, const XSegs* xs = nullptr,
                   const GegluH* gg = nullptr, bool out_f32 = false) {
  TcParams p;
  memset(&p, 0, sizeof(p));
  p.out_f32 = out_f32 ? 1 : 0;
  if (xs) p.kseg = xs->cols / 64;
  if (gg) { p.geglu = gg->d_ff; p.geglu_store_y = gg->store_y; }
  if (cg) {
    p.conv = 1; p.cv_S = cg->S; p.cv_cpb = (cg->C + 63) / 64; p.cv_Ho = cg->Ho; p.cv_Wo = cg->Wo;
    p.cv_pad_h = cg->pad_h; p.cv_pad_w = cg->pad_w; p.cv_str_h = cg->str_h; p.cv_str_w = cg->str_w;
    p.cv_dil_h = cg->dil_h; p.cv_dil_w = cg->dil_w;
  }
  p.M = M; p.Kc = Kc; p.Nout = Nout; p.r = r;
  {   // B200LORA_WPREFETCH=0 switches the pre-wait W loads off (A/B measurements)
    static int wp = -1;
    if (wp < 0) { const char* e = getenv("B200LORA_WPREFETCH"); wp = (e && e[0] == '0') ? 0 : 1; }
    p.w_prefetch = wp;
    // How the PEER CTA's two hand-offs (staged [sT | 1] tile, drained accumulator stage) reach the leader's barriers:
    //   2 (default) relay: the peer's warps arrive on peer-LOCAL barriers (.cta scope, arrivers and observer in one
    //               CTA); a relay warp forwards each phase as an async-proxy transaction (16-byte DSMEM bulk copy with
    //               complete_tx on the leader's barrier) — no generic-proxy release crosses the pair, no GPU-scope
    //               fence in the MMA's path;
    //   1           direct remote arrive with release.cluster, acquire.cluster wait on the leader (MEMBAR.GPU +
    //               CCTL.IVALL per hand-off: +2..6 us per launch, measured);
    //   0           direct remote arrive with the default .cta scope (what round 1 shipped; CUTLASS's 2-SM pipelines
    //               do the same, but the PTX memory model does not order the peer's writes for the leader).
    // B200LORA_SPLIT_B: the second TMA box of a CTA's B half on a third producer thread (warp 2)
    static int sb = -1;
    if (sb < 0) { const char* e = getenv("B200LORA_SPLIT_B"); sb = e ? (e[0] != '0' ? 1 : 0) : B200LORA_SPLIT_B_DEFAULT; }
    p.split_b = sb;
    static int cs = -1;
    if (cs < 0) { const char* e = getenv("B200LORA_CLUSTER_SCOPE"); cs = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2; }
    p.cluster_scope = cs;
  }
  p.BN = tc_pick_bn(M, Kc, Nout, 256, dev_info().sms / 2);
  p.num_m_blk = (int)((M + 255) / 256);
  p.num_n_blk = (Nout + p.BN - 1) / p.BN;
  if (gg) {     // a tile = BN/2 hidden + BN/2 gate columns (chunk pairs of the epilogue are 32 wide: BN % 64 == 0); the
                // last tile may be ragged (the epilogue masks chunk pairs beyond d_ff); no pre-wait W loads
    static int gbn = -1;
    if (gbn < 0) { const char* e = getenv("B200LORA_GEGLU_BN"); gbn = e ? atoi(e) : 128; if (gbn != 64 && gbn != 128 && gbn != 192) gbn = 128; }
    p.BN = gbn;
    p.num_n_blk = (gg->d_ff + p.BN / 2 - 1) / (p.BN / 2);
    p.w_prefetch = 0;
  }
  p.num_k_blk = (Kc + 63) / 64;
  const int smem = tc2_smem_bytes(p.BN, &p.stages);
  if (p.stages < 2) { set_error("tc2: not enough shared memory for BN=%d", p.BN); return B200LORA_E_SHAPE; }
  tc_tmem_plan(p.BN, &p.nst, &p.acc_stride);
  p.scale = scale;
  p.has_bias = has_bias ? 1 : 0;
  p.tail_ksteps = (r + p.has_bias > 16) ? 2 : 1;
  p.T_out = T_out;
  p.dbg = g_dbg_buf;
  p.dbg_cta = g_dbg_cta;
  const unsigned fmt = io_dtype == B200LORA_BF16 ? 1u : 0u;
  const unsigned base = (1u << 4) | (fmt << 7) | (fmt << 10) | ((256u >> 4) << 24);   // M = 256
  p.idesc_main = base | (((unsigned)p.BN >> 3) << 17);
  p.idesc_main_t = base | (((unsigned)(p.BN + 16) >> 3) << 17);
  p.idesc_tail = p.idesc_main;
  const int H = (p.BN + 16) / 2;

  CUtensorMap tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY;
  int rc;
  int box_h = 0, box_n = 0;
  if (cg && conv_tiled_enabled() && conv_tiles_are_boxes(cg->Ho, cg->Wo, cg->str_h, cg->str_w, &box_h, &box_n)) {
    p.conv = 2;
    if ((rc = get_tmap_conv_box(&tmX, X, cg->N, cg->H, cg->W, cg->C, cg->Wo, box_h, box_n, io_dtype))) return rc;
  } else if (cg) {
    if ((rc = get_tmap_im2col(&tmX, X, cg->N, cg->H, cg->W, cg->C, cg->R, cg->S, cg->str_h, cg->str_w,
                              cg->pad_h, cg->pad_w, cg->dil_h, cg->dil_w, io_dtype))) return rc;
  } else if (xs) {
    CUtensorMap* dst[3] = {&tmX, &tmX1, &tmX2};
    for (int i = 0; i < xs->n; ++i)
      if ((rc = get_tmap(dst[i], xs->ptr[i], (uint64_t)M, (uint64_t)xs->cols, (uint64_t)xs->ld[i], 128, 64,
                         io_dtype, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    for (int i = xs->n; i < 3; ++i) *dst[i] = tmX;
  } else if ((rc = get_tmap(&tmX, X, (uint64_t)M, (uint64_t)Kc, (uint64_t)ldx, 128, 64, io_dtype,
                            CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if (!xs) { tmX1 = tmX; tmX2 = tmX; }
  CUtensorMap tmOut;
  if (gg) {     // hidden rows: [BN/2 x 64] boxes; the leader's first 8 gate rows: [8 x 64] boxes
    if ((rc = get_tmap(&tmX1, W, (uint64_t)Nout, (uint64_t)Kc, (uint64_t)ldw, (uint32_t)(p.BN / 2), 64, io_dtype,
                       CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if ((rc = get_tmap(&tmX2, W, (uint64_t)Nout, (uint64_t)Kc, (uint64_t)ldw, 8, 64, io_dtype,
                       CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if ((rc = get_tmap(&tmOut, gg->out, (uint64_t)M, (uint64_t)gg->d_ff, (uint64_t)gg->ldo, 32, 32, io_dtype,
                       CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  }
  if ((rc = get_tmap(&tmWa, W, (uint64_t)Nout, (uint64_t)Kc, (uint64_t)ldw, (uint32_t)(H - 16), 64,
                     io_dtype, CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmWb, W, (uint64_t)Nout, (uint64_t)Kc, (uint64_t)ldw, (uint32_t)H, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;   // leader: H rows in one box
  if ((rc = get_tmap(&tmS1, S1_pack, 16, (uint64_t)Kc, (uint64_t)Kc, 16, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmS2, S2_pack, (uint64_t)Nout, 64, 64, (uint32_t)(p.BN / 2), 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if (gg && !gg->store_y) tmY = tmOut;      // (no pre-activation wanted: the map is never used)
  else if (out_f32) {                       // fp32 Y: [32 rows x 16 columns] boxes = the 64-byte staging rows
    if ((rc = get_tmap(&tmY, Y, (uint64_t)M, (uint64_t)Nout, (uint64_t)ldy, 32, 16, B200LORA_F32,
                       CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  } else if ((rc = get_tmap(&tmY, Y, (uint64_t)M, (uint64_t)Nout, (uint64_t)ldy, 32, 32, io_dtype,
                            CU_TENSOR_MAP_SWIZZLE_64B))) return rc;
  if (!gg) tmOut = tmY;
  cudaError_t e = io_dtype == B200LORA_BF16
                      ? launch_tc2<__nv_bfloat16>(tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p, smem, stream)
                      : launch_tc2<__half>(tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p, smem, stream);
  if (e != cudaSuccess) {
    if (!g_err[0]) set_error("tc2 launch failed: %s", cudaGetErrorString(e));
    return B200LORA_E_CUDA;
  }
  return 0;
}

// B200LORA_PAIR=0 selects the single-CTA kernel (linear_tc.cu), anything else the CTA-pair kernel
static bool use_pair() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200LORA_PAIR");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

static int run_simt(const void* X, long long ldx, const void* W, long long ldw, int w_layout,
                    const void* bias, int bias_dtype, const void* S1, long long s1_rs,
                    long long s1_cs, const void* S2, long long s2_rs, long long s2_cs, int ab_dtype,
                    float scale, void* Y, long long ldy, void* T_out, long long M, int Kc, int Nout,
                    int r, int io_dtype, cudaStream_t stream) {
  if (r > 64) { set_error("simt path supports r <= 64 (got %d)", r); return B200LORA_E_SHAPE; }
  cudaError_t e;
#define CALL(T)                                                                                   \
  launch_simt<T>(X, ldx, W, ldw, w_layout, bias, bias_dtype, S1, s1_rs, s1_cs, S2, s2_rs, s2_cs,  \
                 ab_dtype, scale, Y, ldy, T_out, M, Kc, Nout, r, stream)
  if (io_dtype == B200LORA_F32) e = CALL(float);
  else if (io_dtype == B200LORA_F16) e = CALL(__half);
  else e = CALL(__nv_bfloat16);
#undef CALL
  if (e != cudaSuccess) { set_error("simt launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

}  // namespace b200

// =========================================================================================
// extern "C" surface
// =========================================================================================
using namespace b200;

static bool valid_dtype(int dt) { return dt == B200LORA_F32 || dt == B200LORA_F16 || dt == B200LORA_BF16; }

extern "C" int b200lora_version(void) { return 100; }
extern "C" const char* b200lora_last_error(void) { return g_err; }
extern "C" int b200lora_tc_supported(void) {
  DevInfo di = dev_info();
  return di.ok && di.major == 10 ? 1 : 0;
}
// diagnostic: non-zero if the tensor-core kernel's watchdog fired (site<<16 | block)
extern "C" unsigned int b200lora_watchdog_code(void) { return tc_read_watchdog(); }

// diagnostic: the CUDA runtime's pending (sticky-until-read) error code, without clearing it
extern "C" int b200lora_peek_cuda_error(void) { return (int)cudaPeekAtLastError(); }

// diagnostic, not part of the drop-in surface: trace CTA `cta` of the next tcgen05 launches
extern "C" void b200lora_debug_timeline(void* buf, int cta) {
  g_dbg_buf = reinterpret_cast<long long*>(buf);
  g_dbg_cta = cta;
}

extern "C" int64_t b200lora_pack_elems(int K, int N) { return 80ll * ((int64_t)K + N); }

static bool cast_args_ok(const float* x_src, const void* x_dst, int64_t x_elems) {
  if (x_elems == 0) return true;
  if (x_elems < 0 || !x_src || !x_dst) { set_error("pack + cast: null source / destination"); return false; }
  if (!aligned16(x_src) || !aligned16(x_dst)) { set_error("pack + cast: source and destination must be 16-byte aligned"); return false; }
  return true;
}

static int pack_lora_impl(const void* A, const void* B, int ab_dtype, const void* bias,
                          int bias_dtype, void* pack, int K, int N, int r, int io_dtype,
                          b200lora_stream_t stream_, const float* x_src, void* x_dst, int64_t x_elems) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  if (!cast_args_ok(x_src, x_dst, x_elems)) return B200LORA_E_ARG;
  if (!A || !B || !pack) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(ab_dtype) || (bias && !valid_dtype(bias_dtype)) ||
      (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16)) {
    set_error("pack_lora: io dtype must be bf16/fp16"); return B200LORA_E_DTYPE;
  }
  if (K < 1 || N < 1 || r < 1 || r > 16) { set_error("pack_lora: needs 1 <= r <= 16"); return B200LORA_E_SHAPE; }
  if (!aligned16(pack)) { set_error("pack_lora: pack buffer must be 16-byte aligned"); return B200LORA_E_ALIGN; }
  char* base = reinterpret_cast<char*>(pack);
  void* A_pack = base;
  void* Bt_pack = base + 2ll * 16 * K;
  void* Bn_pack = base + 2ll * 16 * (K + N);
  void* At_pack = base + 2ll * (16 * (K + N) + 64ll * N);
  cudaError_t e = launch_pack(A, B, ab_dtype, bias, bias_dtype, A_pack, Bt_pack, Bn_pack, At_pack, K,
                              N, r, io_dtype, stream, x_src, x_dst, x_elems);
  if (e != cudaSuccess) { set_error("pack launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}
extern "C" int b200lora_pack_lora(const void* A, const void* B, int ab_dtype, const void* bias,
                                  int bias_dtype, void* pack, int K, int N, int r, int io_dtype,
                                  b200lora_stream_t stream_) {
  return pack_lora_impl(A, B, ab_dtype, bias, bias_dtype, pack, K, N, r, io_dtype, stream_, nullptr, nullptr, 0);
}
extern "C" int b200lora_pack_lora_cast(const void* A, const void* B, int ab_dtype, const void* bias,
                                       int bias_dtype, void* pack, int K, int N, int r, int io_dtype,
                                       const float* x_src, void* x_dst, int64_t x_elems, b200lora_stream_t stream_) {
  return pack_lora_impl(A, B, ab_dtype, bias, bias_dtype, pack, K, N, r, io_dtype, stream_, x_src, x_dst, x_elems);
}

static int pack_lora_fused_impl(const void* const* A, const void* const* B, const void* const* bias,
                                const float* scales, const int* r_each, int nseg, int ab_dtype,
                                int bias_dtype, void* pack, int K, int Nseg, int io_dtype,
                                b200lora_stream_t stream_, const float* x_src, void* x_dst, int64_t x_elems) {
  g_err[0] = 0;
  if (!cast_args_ok(x_src, x_dst, x_elems)) return B200LORA_E_ARG;
  if (!A || !B || !scales || !r_each || !pack) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (nseg < 1 || nseg > 4) { set_error("pack_lora_fused: 1 <= nseg <= 4"); return B200LORA_E_SHAPE; }
  if (!valid_dtype(ab_dtype) || (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16)) {
    set_error("pack_lora_fused: io dtype must be bf16/fp16"); return B200LORA_E_DTYPE;
  }
  int r_total = 0;
  bool any_bias = false;
  for (int s = 0; s < nseg; ++s) {
    if (!A[s] || !B[s] || r_each[s] < 1) { set_error("pack_lora_fused: segment %d has a null factor or r < 1", s); return B200LORA_E_ARG; }
    r_total += r_each[s];
    any_bias |= bias && bias[s];
  }
  if (any_bias && !valid_dtype(bias_dtype)) { set_error("bad bias dtype enum"); return B200LORA_E_DTYPE; }
  if (K < 1 || Nseg < 1 || r_total > 16) {
    set_error("pack_lora_fused: needs K, Nseg >= 1 and sum of ranks <= 16, got %d", r_total); return B200LORA_E_SHAPE;
  }
  if (!aligned16(pack)) { set_error("pack_lora_fused: pack buffer must be 16-byte aligned"); return B200LORA_E_ALIGN; }
  const int64_t N = (int64_t)nseg * Nseg;
  char* base = reinterpret_cast<char*>(pack);
  cudaError_t e = launch_pack_seg(A, B, bias, scales, r_each, nseg, ab_dtype, bias_dtype, base,
                                  base + 2ll * 16 * K, base + 2ll * 16 * (K + N),
                                  base + 2ll * (16 * (K + N) + 64ll * N), K, Nseg, io_dtype,
                                  reinterpret_cast<cudaStream_t>(stream_), x_src, x_dst, x_elems);
  if (e != cudaSuccess) { set_error("pack_lora_fused launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}
extern "C" int b200lora_pack_lora_fused(const void* const* A, const void* const* B, const void* const* bias,
                                        const float* scales, const int* r_each, int nseg, int ab_dtype,
                                        int bias_dtype, void* pack, int K, int Nseg, int io_dtype,
                                        b200lora_stream_t stream_) {
  return pack_lora_fused_impl(A, B, bias, scales, r_each, nseg, ab_dtype, bias_dtype, pack, K, Nseg, io_dtype, stream_,
                              nullptr, nullptr, 0);
}
extern "C" int b200lora_pack_lora_fused_cast(const void* const* A, const void* const* B, const void* const* bias,
                                             const float* scales, const int* r_each, int nseg, int ab_dtype,
                                             int bias_dtype, void* pack, int K, int Nseg, int io_dtype,
                                             const float* x_src, void* x_dst, int64_t x_elems,
                                             b200lora_stream_t stream_) {
  return pack_lora_fused_impl(A, B, bias, scales, r_each, nseg, ab_dtype, bias_dtype, pack, K, Nseg, io_dtype, stream_,
                              x_src, x_dst, x_elems);
}

extern "C" int b200lora_lora_gemm(const void* X, int64_t ldx, const void* W, int64_t ldw,
                                  int w_layout, const void* bias, int bias_dtype, const void* S1,
                                  int64_t s1_rs, int64_t s1_cs, const void* S1_pack, const void* S2,
                                  int64_t s2_rs, int64_t s2_cs, int ab_dtype, const void* S2_pack,
                                  float scale, void* Y, int64_t ldy, void* T_out, int64_t M, int Kc,
                                  int Nout, int r, int io_dtype, int impl,
                                  b200lora_stream_t stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  const bool have_packs = S1_pack && S2_pack;
  if (!X || !W || !Y || (!(S1 && S2) && !have_packs)) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(io_dtype) || !valid_dtype(ab_dtype) || (bias && !valid_dtype(bias_dtype))) {
    set_error("bad dtype enum"); return B200LORA_E_DTYPE;
  }
  if (w_layout != B200LORA_W_NK && w_layout != B200LORA_W_KN) { set_error("bad w_layout"); return B200LORA_E_ARG; }
  if (M < 0 || Kc < 1 || Nout < 1 || r < 1) { set_error("bad shape M=%lld Kc=%d Nout=%d r=%d", (long long)M, Kc, Nout, r); return B200LORA_E_SHAPE; }
  if (M == 0) return 0;
  if (ldx < Kc || ldy < Nout || ldw < (w_layout == B200LORA_W_NK ? Kc : Nout)) {
    set_error("leading dimension smaller than the row length"); return B200LORA_E_SHAPE;
  }
  if (impl == B200LORA_IMPL_TC || impl == B200LORA_IMPL_AUTO) {
    bool ok = tc_applicable(X, ldx, W, ldw, w_layout, Y, ldy, r, io_dtype, Nout);
    if (ok && (!have_packs || !aligned16(S1_pack) || !aligned16(S2_pack) || (Kc % 8))) {
      set_error("tc path needs the packed LoRA operands (b200lora_pack_lora), 16-byte aligned");
      ok = false;
      if (impl == B200LORA_IMPL_TC) return B200LORA_E_ARG;
    }
    if (ok)
      return use_pair() ? run_tc2(X, ldx, W, ldw, bias != nullptr, S1_pack, S2_pack, scale, Y, ldy,
                                  T_out, M, Kc, Nout, r, io_dtype, stream)
                        : run_tc(X, ldx, W, ldw, bias != nullptr, S1_pack, S2_pack, scale, Y, ldy,
                                 T_out, M, Kc, Nout, r, io_dtype, stream);
    if (impl == B200LORA_IMPL_TC) {
      if (io_dtype == B200LORA_F32) return B200LORA_E_DTYPE;
      DevInfo di = dev_info();
      if (!di.ok || di.major != 10) return B200LORA_E_ARCH;
      return (r > 16 || w_layout != B200LORA_W_NK) ? B200LORA_E_SHAPE : B200LORA_E_ALIGN;
    }
    g_err[0] = 0;
  }
  if (!S1 || !S2) { set_error("simt path needs the unpacked LoRA factors S1/S2"); return B200LORA_E_ARG; }
  return run_simt(X, ldx, W, ldw, w_layout, bias, bias_dtype, S1, s1_rs, s1_cs, S2, s2_rs, s2_cs,
                  ab_dtype, scale, Y, ldy, T_out, M, Kc, Nout, r, io_dtype, stream);
}

extern "C" int b200lora_linear_fwd(const void* X, const void* W0, const void* bias, int bias_dtype,
                                   const void* A, const void* B, int ab_dtype, const void* pack,
                                   float scale, void* Y, void* T_out, int64_t M, int K, int N,
                                   int r, int io_dtype, int impl, b200lora_stream_t stream) {
  // S1 = A [r,K] (row stride K), S2 = B [N,r] (row stride r); packs: A_pack and Bn_pack
  const char* pk = reinterpret_cast<const char*>(pack);
  const void* s1p = pk;
  const void* s2p = pk ? pk + 2ll * 16 * ((int64_t)K + N) : nullptr;
  return b200lora_lora_gemm(X, K, W0, K, B200LORA_W_NK, bias, bias_dtype, A, K, 1, s1p, B, r, 1,
                            ab_dtype, s2p, scale, Y, N, T_out, M, K, N, r, io_dtype, impl, stream);
}

extern "C" int b200lora_linear_bwd_dx(const void* dY, const void* W, int w_layout, const void* A,
                                      const void* B, int ab_dtype, const void* pack, float scale,
                                      void* dX, void* dT_out, int64_t M, int K, int N, int r,
                                      int io_dtype, int impl, b200lora_stream_t stream) {
  // dX = dY·W0 + (s·dY·B)·A : contraction over N.  S1 = Bᵀ (logical [r,N]: (j,n) -> B[n*r+j]),
  // S2 = Aᵀ (logical [K,r]: (k,j) -> A[j*K+k]).  W as [K,N] (NK layout of the transposed
  // problem) or W0 itself [N,K] (KN layout).  packs: Bt_pack and At_pack
  const int64_t ldw = (w_layout == B200LORA_W_NK) ? N : K;
  const char* pk = reinterpret_cast<const char*>(pack);
  const void* s1p = pk ? pk + 2ll * 16 * K : nullptr;
  const void* s2p = pk ? pk + 2ll * (16 * ((int64_t)K + N) + 64ll * N) : nullptr;
  return b200lora_lora_gemm(dY, N, W, ldw, w_layout, nullptr, 0, B, 1, r, s1p, A, 1, K, ab_dtype,
                            s2p, scale, dX, K, dT_out, M, N, K, r, io_dtype, impl, stream);
}

// Backward dX with an fp32 result (the layer's input was fp32 under autocast: autograd wants an fp32 gradient, and the
// accumulator is fp32 already — the bf16 -> fp32 cast launch behind the kernel disappears).  tcgen05 pair kernel only.
extern "C" int b200lora_linear_bwd_dx_f32out(const void* dY, const void* Wt, const void* pack, float scale, float* dX,
                                             void* dT_out, int64_t M, int K, int N, int r, int io_dtype,
                                             b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!dY || !Wt || !pack || !dX) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16) { set_error("fp32 dX output needs 16-bit io"); return B200LORA_E_DTYPE; }
  if (M <= 0) return 0;
  if (!use_pair()) { set_error("fp32 dX output lives in the CTA-pair kernel (B200LORA_PAIR=0 is set)"); return B200LORA_E_ARG; }
  if (!tc_applicable(dY, N, Wt, N, B200LORA_W_NK, dX, K, r, io_dtype, K) || (N % 8)) return B200LORA_E_ALIGN;
  const char* pk = reinterpret_cast<const char*>(pack);
  const void* s1p = pk + 2ll * 16 * K;                                   // Bt_pack [16, N]
  const void* s2p = pk + 2ll * (16 * ((int64_t)K + N) + 64ll * N);       // At_pack [K, 64]
  return run_tc2(dY, N, Wt, N, false, s1p, s2p, scale, dX, K, dT_out, M, N, K, r, io_dtype,
                 reinterpret_cast<cudaStream_t>(stream_), nullptr, nullptr, nullptr, true);
}

// GEGLU projection with the activation in the epilogue: W0 = [hidden ; gate] rows [2*d_ff, K];
// out = hidden * gelu(gate) [M, d_ff]; Y [M, 2*d_ff] (the pre-activation, what the backward needs) is optional.
extern "C" int b200lora_linear_fwd_geglu(const void* X, const void* W0, const void* pack, int has_bias, float scale,
                                         void* Y, void* out, void* T_out, int64_t M, int K, int d_ff, int r,
                                         int io_dtype, b200lora_stream_t stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  if (!X || !W0 || !pack || !out) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16) { set_error("the GEGLU epilogue is 16-bit io only"); return B200LORA_E_DTYPE; }
  if (d_ff < 64 || d_ff % 32) { set_error("the GEGLU epilogue needs d_ff %% 32 == 0 and d_ff >= 64 (got %d)", d_ff); return B200LORA_E_SHAPE; }
  if (r < 1 || r > 16) { set_error("the GEGLU epilogue supports 1 <= r <= 16 (got %d)", r); return B200LORA_E_SHAPE; }
  if (M <= 0) return 0;
  const int N = 2 * d_ff;
  if (!use_pair()) { set_error("the GEGLU epilogue lives in the CTA-pair kernel (B200LORA_PAIR=0 is set)"); return B200LORA_E_ARG; }
  if (!tc_applicable(X, K, W0, K, B200LORA_W_NK, Y ? Y : out, Y ? N : d_ff, r, io_dtype, K) || !aligned16(out) || (d_ff % 8)) return B200LORA_E_ALIGN;
  const char* pk = reinterpret_cast<const char*>(pack);
  const void* s1p = pk;
  const void* s2p = pk + 2ll * 16 * ((int64_t)K + N);
  GegluH gg{d_ff, out, d_ff, Y ? 1 : 0};
  return run_tc2(X, K, W0, K, has_bias != 0, s1p, s2p, scale, Y, N, T_out, M, K, N, r, io_dtype, stream, nullptr, nullptr, &gg);
}

extern "C" int b200lora_geglu_bwd(const void* dout, const void* Y, void* dY, int64_t M, int d_ff, int io_dtype,
                                  b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!dout || !Y || !dY) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(io_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  const int vec = 16 / dtype_size(io_dtype);
  if (d_ff < 1 || d_ff % vec) { set_error("geglu_bwd needs d_ff %% %d == 0", vec); return B200LORA_E_SHAPE; }
  if (!aligned16(dout) || !aligned16(Y) || !aligned16(dY)) { set_error("geglu_bwd needs 16-byte aligned pointers"); return B200LORA_E_ALIGN; }
  if (M <= 0) return 0;
  DevInfo di = dev_info();
  cudaError_t e = launch_geglu_bwd(dout, Y, dY, M, d_ff, io_dtype, di.sms > 0 ? di.sms : 148, reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("geglu_bwd launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

// =========================================================================================
// Fused sibling projections: layers that read the SAME input (attention to_q/to_k/to_v)
// run as one GEMM over the concatenated outputs.  Forward is b200lora_lora_gemm on the concatenated
// operands (W_cat [nseg*Nseg, K], pack of A_cat [r_total, K] and the block-diagonal, scale-folded B);
// these two entry points are the backward: dY arrives as nseg separate tensors.
// =========================================================================================
static int bwd_dx_fused_impl(const void* const* dY, const int64_t* ld_dy, int nseg, const void* Wt, const void* pack,
                             void* dX, void* dT_out, int64_t M, int K, int Nseg, int r_total, int io_dtype,
                             b200lora_stream_t stream_, bool out_f32) {
  // dX = sum_i dY_i·W0_i + dT·A_cat with dT = [dY_0 | dY_1 | ..]·B_bd riding as rank rows; Wt is
  // W_catᵀ stored [K, nseg*Nseg] (the W_NK operand of the transposed problem)
  g_err[0] = 0;
  if (!dY || !ld_dy || !Wt || !pack || !dX) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (nseg < 1 || nseg > 3) { set_error("fused backward takes 1..3 segments (got %d)", nseg); return B200LORA_E_SHAPE; }
  if (Nseg % 64) { set_error("fused backward needs segment width %% 64 == 0 (got %d)", Nseg); return B200LORA_E_SHAPE; }
  if (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16) { set_error("fused path is 16-bit io only"); return B200LORA_E_DTYPE; }
  if (M <= 0) return 0;
  const int Ncat = nseg * Nseg;
  XSegs xs;
  xs.n = nseg; xs.cols = Nseg;
  for (int i = 0; i < nseg; ++i) {
    if (!dY[i] || !aligned16(dY[i]) || (ld_dy[i] % 8)) { set_error("fused backward: dY[%d] null or misaligned", i); return B200LORA_E_ALIGN; }
    xs.ptr[i] = dY[i]; xs.ld[i] = ld_dy[i];
  }
  if (!tc_applicable(dY[0], ld_dy[0], Wt, Ncat, B200LORA_W_NK, dX, K, r_total, io_dtype, K)) return B200LORA_E_ALIGN;
  const char* pk = reinterpret_cast<const char*>(pack);
  const void* s1p = pk + 2ll * 16 * K;                                       // Bt_pack [16, Ncat]
  const void* s2p = pk + 2ll * (16 * ((int64_t)K + Ncat) + 64ll * Ncat);     // At_pack [K, 64]
  if (out_f32 && !use_pair()) { set_error("fp32 dX output lives in the CTA-pair kernel (B200LORA_PAIR=0 is set)"); return B200LORA_E_ARG; }
  return run_tc2(dY[0], ld_dy[0], Wt, Ncat, false, s1p, s2p, 1.0f, dX, K, dT_out, M, Ncat, K, r_total, io_dtype,
                 reinterpret_cast<cudaStream_t>(stream_), nullptr, &xs, nullptr, out_f32);
}
extern "C" int b200lora_linear_bwd_dx_fused(const void* const* dY, const int64_t* ld_dy, int nseg,
                                            const void* Wt, const void* pack, void* dX, void* dT_out,
                                            int64_t M, int K, int Nseg, int r_total, int io_dtype,
                                            b200lora_stream_t stream_) {
  return bwd_dx_fused_impl(dY, ld_dy, nseg, Wt, pack, dX, dT_out, M, K, Nseg, r_total, io_dtype, stream_, false);
}
extern "C" int b200lora_linear_bwd_dx_fused_f32out(const void* const* dY, const int64_t* ld_dy, int nseg,
                                                   const void* Wt, const void* pack, float* dX, void* dT_out,
                                                   int64_t M, int K, int Nseg, int r_total, int io_dtype,
                                                   b200lora_stream_t stream_) {
  return bwd_dx_fused_impl(dY, ld_dy, nseg, Wt, pack, dX, dT_out, M, K, Nseg, r_total, io_dtype, stream_, true);
}

extern "C" int b200lora_lora_grads_fused(const void* const* dY, const int64_t* ld_dy, int nseg, const void* X,
                                         const void* T, const void* dT, const float* scales,
                                         float* const* dA, float* const* dB, int64_t M, int K, int Nseg,
                                         const int* r_each, int io_dtype, b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!dY || !ld_dy || !X || !T || !dT || !scales || !dA || !dB || !r_each) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (nseg < 1 || nseg > 3) { set_error("fused grads take 1..3 segments (got %d)", nseg); return B200LORA_E_SHAPE; }
  if (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16) { set_error("fused path is 16-bit io only"); return B200LORA_E_DTYPE; }
  if ((K % 8) || (Nseg % 8)) { set_error("fused grads need K and N multiples of 8"); return B200LORA_E_ALIGN; }
  int r_total = 0;
  for (int i = 0; i < nseg; ++i) {
    if (r_each[i] < 1 || !dA[i] || !dB[i] || !dY[i]) { set_error("fused grads: bad segment %d", i); return B200LORA_E_ARG; }
    r_total += r_each[i];
  }
  if (r_total > 16) { set_error("fused grads need sum of ranks <= 16 (got %d)", r_total); return B200LORA_E_SHAPE; }
  if (M <= 0) return 0;
  DevInfo di = dev_info();
  if (!di.ok || di.major != 10) { set_error("fused grads need an sm_100 device"); return B200LORA_E_CUDA; }
  CUtensorMap tmDY[3], tmX;
  const CUtensorMap* pDY[3];
  int rc;
  for (int i = 0; i < nseg; ++i) {
    if (!aligned16(dY[i]) || (ld_dy[i] % 8)) { set_error("fused grads: dY[%d] misaligned", i); return B200LORA_E_ALIGN; }
    if ((rc = get_tmap(&tmDY[i], dY[i], (uint64_t)M, (uint64_t)Nseg, (uint64_t)ld_dy[i], 128, 64, io_dtype,
                       CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    pDY[i] = &tmDY[i];
  }
  if (!aligned16(X)) { set_error("fused grads: X misaligned"); return B200LORA_E_ALIGN; }
  if ((rc = get_tmap(&tmX, X, (uint64_t)M, (uint64_t)K, (uint64_t)K, 128, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  cudaError_t e = io_dtype == B200LORA_BF16
                      ? launch_grads_fused_tc<__nv_bfloat16>(pDY, &tmX, T, dT, scales, dA, dB, M, K, Nseg, nseg,
                                                             r_each, di.sms, stream)
                      : launch_grads_fused_tc<__half>(pDY, &tmX, T, dT, scales, dA, dB, M, K, Nseg, nseg, r_each,
                                                      di.sms, stream);
  if (e != cudaSuccess) { set_error("fused grads launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

// =========================================================================================
// LoRA-Conv2d (implicit GEMM over NHWC activations)
// =========================================================================================
static int conv_out(int in, int k, int st, int pad, int dil) { return (in + 2 * pad - dil * (k - 1) - 1) / st + 1; }

extern "C" int b200lora_conv2d_repack_weight(const void* W_oihw, int w_dtype, void* out, int Cout,
                                             int Cin, int R, int S, int mode, int io_dtype,
                                             b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!W_oihw || !out) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(w_dtype) || (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16)) { set_error("bad dtype"); return B200LORA_E_DTYPE; }
  if (mode != 0 && mode != 1) { set_error("bad mode"); return B200LORA_E_ARG; }
  // forward rows are padded to 64 input channels; backward rows to 64 output channels + the 64 LoRA columns
  const int Cp = mode == 0 ? (Cin + 63) / 64 * 64 : (Cout + 63) / 64 * 64 + 64;
  cudaError_t e = launch_conv_repack(W_oihw, w_dtype, out, Cout, Cin, R, S, Cp, mode, io_dtype,
                                     reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("conv repack launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_conv2d_pack_lora(const void* A, const void* B, int ab_dtype, const void* bias,
                                         int bias_dtype, void* A_pack, void* Bn_pack, void* Wb,
                                         int Cout, int Cin, int R, int S, int r, int io_dtype,
                                         b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!A || !B || !A_pack || !Bn_pack) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(ab_dtype) || (bias && !valid_dtype(bias_dtype)) ||
      (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16)) { set_error("bad dtype"); return B200LORA_E_DTYPE; }
  if (r < 1 || r > 16) { set_error("conv pack: needs 1 <= r <= 16"); return B200LORA_E_SHAPE; }
  const int Cp = (Cin + 63) / 64 * 64, Coutp = (Cout + 63) / 64 * 64, Cz = Coutp + 64;
  cudaError_t e = launch_conv_pack(A, B, ab_dtype, bias, bias_dtype, A_pack, Bn_pack, Wb, Cout, Cin, R, S,
                                   Cp, Coutp, Cz, r, io_dtype, reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("conv pack launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_conv2d_gemm(const void* X_nhwc, const void* Wk, int has_bias, const void* S1_pack,
                                    const void* S2_pack, float scale, void* Y_nhwc, void* T_out, int N,
                                    int H, int W, int C, int Cout, int R, int S, int str_h, int str_w,
                                    int pad_h, int pad_w, int dil_h, int dil_w, int r, int io_dtype,
                                    b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!X_nhwc || !Wk || !S1_pack || !S2_pack || !Y_nhwc) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16) { set_error("conv2d: 16-bit io only"); return B200LORA_E_DTYPE; }
  if (N < 1 || H < 1 || W < 1 || C < 1 || Cout < 1 || R < 1 || S < 1 || str_h < 1 || str_w < 1 ||
      dil_h < 1 || dil_w < 1 || pad_h < 0 || pad_w < 0 || r < 1 || r > 16) { set_error("conv2d: bad shape"); return B200LORA_E_SHAPE; }
  if ((C % 8) || (Cout % 8)) { set_error("conv2d: channel counts must be multiples of 8"); return B200LORA_E_ALIGN; }
  if (!aligned16(X_nhwc) || !aligned16(Wk) || !aligned16(Y_nhwc) || !aligned16(S1_pack) || !aligned16(S2_pack)) {
    set_error("conv2d: pointers must be 16-byte aligned"); return B200LORA_E_ALIGN;
  }
  DevInfo di = dev_info();
  if (!di.ok || di.major != 10) { set_error("conv2d needs an sm_100 device"); return B200LORA_E_ARCH; }
  ConvGeoH g;
  g.N = N; g.H = H; g.W = W; g.C = C; g.R = R; g.S = S; g.str_h = str_h; g.str_w = str_w;
  g.pad_h = pad_h; g.pad_w = pad_w; g.dil_h = dil_h; g.dil_w = dil_w;
  g.Ho = conv_out(H, R, str_h, pad_h, dil_h);
  g.Wo = conv_out(W, S, str_w, pad_w, dil_w);
  if (g.Ho < 1 || g.Wo < 1) { set_error("conv2d: empty output"); return B200LORA_E_SHAPE; }
  const long long M = (long long)N * g.Ho * g.Wo;
  const int Kc = R * S * ((C + 63) / 64 * 64);
  return run_tc2(X_nhwc, 0, Wk, Kc, has_bias != 0, S1_pack, S2_pack, scale, Y_nhwc, Cout, T_out, M, Kc,
                 Cout, r, io_dtype, reinterpret_cast<cudaStream_t>(stream_), &g);
}

extern "C" int b200lora_conv2d_zcat(const void* dY, const void* B, int ab_dtype, float scale, void* Z,
                                    void* dT, int64_t M, int Cout, int r, int io_dtype,
                                    b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!dY || !B || !Z || !dT) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(ab_dtype) || (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16)) { set_error("bad dtype"); return B200LORA_E_DTYPE; }
  if (Cout % 8 || r < 1 || r > 16) { set_error("zcat: Cout %% 8 == 0 and 1 <= r <= 16"); return B200LORA_E_SHAPE; }
  if (M <= 0) return 0;
  const int Coutp = (Cout + 63) / 64 * 64;
  DevInfo di = dev_info();
  if (!di.ok || di.major != 10) { set_error("zcat needs an sm_100 device"); return B200LORA_E_CUDA; }
  if (!aligned16(dY) || !aligned16(Z)) { set_error("zcat needs 16-byte aligned dY / Z"); return B200LORA_E_ALIGN; }
  const int Cz = Coutp + 64;
  CUtensorMap tmDY, tmZ;
  int rc;
  if ((rc = get_tmap(&tmDY, dY, (uint64_t)M, (uint64_t)Cout, (uint64_t)Cout, 128, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  if ((rc = get_tmap(&tmZ, Z, (uint64_t)M, (uint64_t)Cz, (uint64_t)Cz, 128, 64, io_dtype,
                     CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  cudaError_t e = io_dtype == B200LORA_BF16
                      ? launch_conv_zcat_tc<__nv_bfloat16>(&tmDY, &tmZ, B, ab_dtype, scale, Z, dT, M, Cout, Coutp, Cz,
                                                           r, di.sms, stream)
                      : launch_conv_zcat_tc<__half>(&tmDY, &tmZ, B, ab_dtype, scale, Z, dT, M, Cout, Coutp, Cz, r,
                                                    di.sms, stream);
  if (e != cudaSuccess) { set_error("zcat launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_conv2d_grad_a(const void* X_nhwc, const void* dT, float scale, float* dA, int N, int H,
                                  int W, int C, int R, int S, int str_h, int str_w, int pad_h, int pad_w,
                                  int dil_h, int dil_w, int r, int io_dtype, b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!X_nhwc || !dT || !dA) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (io_dtype != B200LORA_BF16 && io_dtype != B200LORA_F16) { set_error("conv2d_dA: 16-bit io only"); return B200LORA_E_DTYPE; }
  if (C % 8 || r < 1 || r > 16) { set_error("conv2d_dA: C %% 8 == 0 and 1 <= r <= 16"); return B200LORA_E_SHAPE; }
  int geo[14] = {N, H, W, C, conv_out(H, R, str_h, pad_h, dil_h), conv_out(W, S, str_w, pad_w, dil_w), R, S,
                 str_h, str_w, pad_h, pad_w, dil_h, dil_w};
  // tensor-pipe reduction over the im2col view of X (same TMA map as the forward's X operand)
  DevInfo di = dev_info();
  if (!di.ok || di.major != 10) { set_error("conv2d_grad_a needs an sm_100 device"); return B200LORA_E_CUDA; }
  if (!aligned16(X_nhwc)) { set_error("conv2d_grad_a needs a 16-byte aligned X"); return B200LORA_E_ALIGN; }
  CUtensorMap tmX;
  int rc;
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  if ((rc = get_tmap_im2col(&tmX, X_nhwc, N, H, W, C, R, S, str_h, str_w, pad_h, pad_w, dil_h, dil_w, io_dtype))) return rc;
  cudaError_t e = io_dtype == B200LORA_BF16
                      ? launch_conv_grad_a_tc<__nv_bfloat16>(&tmX, dT, scale, dA, geo, r, di.sms, stream)
                      : launch_conv_grad_a_tc<__half>(&tmX, dT, scale, dA, geo, r, di.sms, stream);
  if (e != cudaSuccess) { set_error("conv2d_grad_a launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int64_t b200lora_lora_grads_workspace(int64_t M, int K, int N, int r) {
  const int RP = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
  int sms = dev_info().sms;
  if (sms <= 0) sms = 148;
  // worst case over element sizes and over which of dA / dB the call asks for
  int64_t best = 0;
  for (int es = 2; es <= 4; es += 2)
    for (int both = 0; both < 2; ++both) {
      const int64_t a = (int64_t)grad_reduce_splits(M, K, both ? N : 0, es, sms) * K * RP * 4;
      const int64_t b = (int64_t)grad_reduce_splits(M, N, both ? K : 0, es, sms) * N * RP * 4;
      if (a + b > best) best = a + b;
    }
  return best + 256;
}

extern "C" int b200lora_lora_grads(const void* dY, const void* X, const void* T, const void* dT,
                                   float scale, float* dA, float* dB, int64_t M, int K, int N,
                                   int r, int io_dtype, int deterministic, void* workspace,
                                   int64_t workspace_bytes, b200lora_stream_t stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  if (!valid_dtype(io_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (r < 1 || r > 64) { set_error("lora_grads supports 1 <= r <= 64 (got %d)", r); return B200LORA_E_SHAPE; }
  if (M <= 0) return 0;
  if (!dA && !dB) return 0;
  const int vec = 16 / dtype_size(io_dtype);
  if ((dB && (N % vec)) || (dA && (K % vec))) { set_error("lora_grads needs N and K multiples of %d", vec); return B200LORA_E_ALIGN; }
  if ((dB && (!dY || !T)) || (dA && (!X || !dT))) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  DevInfo di = dev_info();
  const int sms = di.sms > 0 ? di.sms : 148;
  float* partA = nullptr;
  float* partB = nullptr;
  if (deterministic) {
    if (!workspace || workspace_bytes < b200lora_lora_grads_workspace(M, K, N, r)) {
      set_error("deterministic mode needs a workspace of b200lora_lora_grads_workspace bytes");
      return B200LORA_E_ARG;
    }
    const int RP = r <= 4 ? 4 : (r <= 8 ? 8 : 16);
    partA = reinterpret_cast<float*>(workspace);
    partB = partA + (dA ? (int64_t)grad_reduce_splits(M, K, dB ? N : 0, dtype_size(io_dtype), sms) * K * RP : 0);
  }
  // 16-bit io: tensor-pipe reduction fed by TMA ([128 tokens x 64 columns] boxes)
  CUtensorMap tmDY, tmX;
  const CUtensorMap* pDY = nullptr;
  const CUtensorMap* pX = nullptr;
  if (io_dtype != B200LORA_F32) {
    if (!di.ok || di.major != 10) { set_error("lora_grads (16-bit io) needs an sm_100 device"); return B200LORA_E_CUDA; }
    if ((dB && !aligned16(dY)) || (dA && !aligned16(X))) { set_error("lora_grads needs 16-byte aligned dY / X"); return B200LORA_E_ALIGN; }
    int rc;
    if (dB && (rc = get_tmap(&tmDY, dY, (uint64_t)M, (uint64_t)N, (uint64_t)N, 128, 64, io_dtype,
                             CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    if (dA && (rc = get_tmap(&tmX, X, (uint64_t)M, (uint64_t)K, (uint64_t)K, 128, 64, io_dtype,
                             CU_TENSOR_MAP_SWIZZLE_128B))) return rc;
    pDY = dB ? &tmDY : &tmX;
    pX = dA ? &tmX : &tmDY;
  }
  cudaError_t e = cudaSuccess;
#define RUN(T_) e = launch_lora_grads<T_>(dY, X, T, dT, scale, dA, dB, partA, partB, M, K, N, r, pDY, pX, sms, stream)
  if (io_dtype == B200LORA_F32) RUN(float);
  else if (io_dtype == B200LORA_F16) RUN(__half);
  else RUN(__nv_bfloat16);
#undef RUN
  if (e != cudaSuccess) { set_error("lora_grads launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// No-allocation variant of the deterministic gradient reductions (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The [splits, C, rp] partial sums live in a file-scope device array that every GPU
// loading this library holds until exit.  `g_static_scratch` is a host-side STAND-IN in host code, not a device
// address: the pointer the kernels get comes from cudaGetSymbolAddress, once per thread and device, with that
// device current (the call also binds the primary context to the thread, so the tensor-map encoders need no
// cudaFree(0) on this path).  Every slot the finalize kernel reads is written by the same call's first stage
// (reduce.cu: partial[(split * C + c) * rp + j] for every split, c < C, j < r), so nothing has to start at zero;
// the array keeps the last call's contents and calls that overlap on the GPU (two streams, two threads) race.
// ---------------------------------------------------------------------------------------------------------
static __device__ __align__(256) unsigned char g_static_scratch[B200LORA_STATIC_SCRATCH_BYTES];

static int static_scratch_ptr(void** out) {
  constexpr int MAX_DEV = 64;
  static thread_local void* tl_ptr[MAX_DEV] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
    set_error("lora_grads_static: no current device (%s)", cudaGetErrorString(e));
    return B200LORA_E_CUDA;
  }
  if (!tl_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      set_error("lora_grads_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return B200LORA_E_CUDA;
    }
    tl_ptr[dev] = p;
    g_bound_dev = dev;      // the runtime call above made the primary context current on this thread
  }
  *out = tl_ptr[dev];
  return 0;
}

extern "C" int b200lora_lora_grads_static(const void* dY, const void* X, const void* T, const void* dT, float scale,
                                          float* dA, float* dB, int64_t M, int K, int N, int r, int io_dtype,
                                          b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!valid_dtype(io_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (r < 1 || r > 64 || M < 0 || K < 1 || N < 1) { set_error("lora_grads_static: bad shape M=%lld K=%d N=%d r=%d", (long long)M, K, N, r); return B200LORA_E_SHAPE; }
  const int64_t need = b200lora_lora_grads_workspace(M, K, N, r);
  if (need > (int64_t)B200LORA_STATIC_SCRATCH_BYTES) {
    set_error("lora_grads_static: M=%lld K=%d N=%d r=%d needs %lld bytes of scratch, above the built-in %lld "
              "(use b200lora_lora_grads with a caller workspace)", (long long)M, K, N, r, (long long)need,
              (long long)B200LORA_STATIC_SCRATCH_BYTES);
    return B200LORA_E_SHAPE;
  }
  void* ws = nullptr;
  int rc = static_scratch_ptr(&ws);
  if (rc) return rc;
  return b200lora_lora_grads(dY, X, T, dT, scale, dA, dB, M, K, N, r, io_dtype, /*deterministic=*/1, ws,
                             (int64_t)B200LORA_STATIC_SCRATCH_BYTES, stream_);
}

extern "C" int b200lora_arena_adamw(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                                    float lr, float beta1, float beta2, float eps, float weight_decay,
                                    float max_grad_norm, float* scratch, b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!params || !grads || !exp_avg || !exp_avg_sq || !scratch) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (n <= 0) return 0;
  if ((reinterpret_cast<uintptr_t>(grads) & 15) != 0) { set_error("arena_adamw needs a 16-byte aligned gradient arena"); return B200LORA_E_ALIGN; }
  cudaError_t e = launch_arena_adamw(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, weight_decay,
                                     max_grad_norm, scratch, reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("arena_adamw launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_flat_axpby(const void* a, const void* b, void* out, int64_t n, float alpha, float beta,
                                   int in_dtype, int out_dtype, b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!a || !out) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(in_dtype) || !valid_dtype(out_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (n < 0) { set_error("flat_axpby: negative length"); return B200LORA_E_SHAPE; }
  if (n == 0) return 0;
  cudaError_t e = launch_flat_axpby(a, b, out, n, alpha, beta, in_dtype, out_dtype, reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("flat_axpby launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_ti_grad_rows(const void* dY, const int64_t* token_ids, int64_t n_tokens, int D,
                                     const int64_t* rows, int P, float* grad_rows, int io_dtype,
                                     b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!dY || !token_ids || !rows || !grad_rows) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(io_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (D < 1 || P < 1 || P > 65535 || n_tokens < 0) { set_error("ti_grad_rows: bad shape (D=%d, P=%d)", D, P); return B200LORA_E_SHAPE; }
  if (n_tokens == 0) return 0;
  cudaError_t e = launch_ti_grad_rows(dY, reinterpret_cast<const long long*>(token_ids), n_tokens, D,
                                      reinterpret_cast<const long long*>(rows), P, grad_rows, io_dtype,
                                      reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("ti_grad_rows launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_ti_step(void* table, int table_dtype, int D, const int64_t* rows, int P, float* master_rows,
                                const float* grad_rows, float* exp_avg, float* exp_avg_sq, int step, float lr,
                                float beta1, float beta2, float eps, float weight_decay, float lambda_,
                                float target_norm, float* norms_out, const float* lr_dev, float* step_dev,
                                b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (!table || !rows || !master_rows || !grad_rows || !exp_avg || !exp_avg_sq) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(table_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (D < 1 || P < 1 || (!step_dev && step < 1)) { set_error("ti_step: bad shape or step (D=%d, P=%d, step=%d)", D, P, step); return B200LORA_E_SHAPE; }
  cudaError_t e = launch_ti_step(table, table_dtype, D, reinterpret_cast<const long long*>(rows), P, master_rows,
                                 grad_rows, exp_avg, exp_avg_sq, (float)step, lr, beta1, beta2, eps, weight_decay,
                                 lambda_, target_norm, norms_out, lr_dev, step_dev,
                                 reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("ti_step launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_merge(const void* Win, void* Wout, const void* up, const void* down,
                              float alpha, int N, int K, int r, int w_dtype, int lora_dtype,
                              b200lora_stream_t stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  if (!Win || !Wout || !up || !down) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(w_dtype) || !valid_dtype(lora_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (N < 1 || K < 1 || r < 1) { set_error("bad shape"); return B200LORA_E_SHAPE; }
  const int vec = 16 / dtype_size(w_dtype);
  if (K % vec || !aligned16(Win) || !aligned16(Wout)) { set_error("merge needs K %% %d == 0 and 16-byte aligned W", vec); return B200LORA_E_ALIGN; }
  cudaError_t e;
#define M2(TW, TL) launch_merge<TW, TL, 0>(Win, Wout, up, down, alpha, N, K, r, nullptr, stream)
#define M1(TW)                                                          \
  (lora_dtype == B200LORA_F32 ? M2(TW, float)                           \
   : lora_dtype == B200LORA_F16 ? M2(TW, __half) : M2(TW, __nv_bfloat16))
  if (w_dtype == B200LORA_F32) e = M1(float);
  else if (w_dtype == B200LORA_F16) e = M1(__half);
  else e = M1(__nv_bfloat16);
#undef M1
#undef M2
  if (e != cudaSuccess) { set_error("merge launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int64_t b200lora_merge_batch_workspace(int n_items) {
  return n_items > 0 ? (int64_t)n_items * (int64_t)sizeof(MergeItem) : 0;
}

// host-side half: validate the items and write the device-format problem table into `table_host`
extern "C" int b200lora_merge_batch_plan(const b200lora_merge_item* items, int n_items, int w_dtype,
                                         void* table_host, int64_t table_bytes, int64_t* total_tiles,
                                         int* max_rank) {
  g_err[0] = 0;
  if (!items || n_items < 1 || !table_host || !total_tiles || !max_rank) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(w_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (table_bytes < b200lora_merge_batch_workspace(n_items)) { set_error("merge_batch_plan: table buffer too small"); return B200LORA_E_ARG; }
  const int vec = 16 / dtype_size(w_dtype);
  MergeItem* host = static_cast<MergeItem*>(table_host);
  int max_r = 0;
  for (int i = 0; i < n_items; ++i) {
    const b200lora_merge_item& s = items[i];
    if (!s.Win || !s.Wout || !s.up || !s.down) { set_error("merge_batch: item %d has a null pointer", i); return B200LORA_E_ARG; }
    if (s.N < 1 || s.K < 1 || s.r < 1) { set_error("merge_batch: item %d has a bad shape", i); return B200LORA_E_SHAPE; }
    if (s.K % vec || !aligned16(s.Win) || !aligned16(s.Wout)) {
      set_error("merge_batch: item %d needs K %% %d == 0 and 16-byte aligned W", i, vec); return B200LORA_E_ALIGN;
    }
    host[i] = MergeItem{s.Win, s.Wout, s.up, s.down, s.N, s.K, s.r, 0, 0, 0};
    if (s.r > max_r) max_r = s.r;
  }
  const long long tiles = merge_batch_plan(host, n_items, w_dtype);
  if (tiles > 0x7fffffffll) { set_error("merge_batch: too many tiles"); return B200LORA_E_SHAPE; }
  *total_tiles = tiles;
  *max_rank = max_r;
  return 0;
}

// device-side half: ONE launch over the table (already resident in device memory)
extern "C" int b200lora_merge_batch_run(const void* table_dev, int n_items, int64_t total_tiles, int max_rank,
                                        float alpha, int w_dtype, int lora_dtype, b200lora_stream_t stream_) {
  g_err[0] = 0;
  if (n_items == 0 || total_tiles == 0) return 0;
  if (!table_dev || n_items < 0 || total_tiles < 0 || total_tiles > 0x7fffffffll || max_rank < 1) { set_error("merge_batch_run: bad argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(w_dtype) || !valid_dtype(lora_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (!aligned16(table_dev)) { set_error("merge_batch_run: table must be 16-byte aligned"); return B200LORA_E_ALIGN; }
  DevInfo di = dev_info();
  cudaError_t e = launch_merge_batch(static_cast<const MergeItem*>(table_dev), n_items, (int)total_tiles, max_rank,
                                     alpha, w_dtype, lora_dtype, di.sms > 0 ? di.sms : 148,
                                     reinterpret_cast<cudaStream_t>(stream_));
  if (e != cudaSuccess) { set_error("merge_batch launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_absmean_accum(const void* up, const void* down, float* out_sum, int N,
                                      int K, int r, int lora_dtype, b200lora_stream_t stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  if (!up || !down || !out_sum) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  if (!valid_dtype(lora_dtype)) { set_error("bad dtype enum"); return B200LORA_E_DTYPE; }
  if (r < 1) { set_error("absmean needs r >= 1"); return B200LORA_E_SHAPE; }
  const int vec = 16 / dtype_size(lora_dtype);
  cudaError_t e;
  if (r > 16 || (K % vec)) e = launch_absmean_any(up, down, out_sum, N, K, r, lora_dtype, stream);   // any rank / any K
  else if (lora_dtype == B200LORA_F32) e = launch_merge<float, float, 1>(nullptr, nullptr, up, down, 0.f, N, K, r, out_sum, stream);
  else if (lora_dtype == B200LORA_F16) e = launch_merge<__half, __half, 1>(nullptr, nullptr, up, down, 0.f, N, K, r, out_sum, stream);
  else e = launch_merge<__nv_bfloat16, __nv_bfloat16, 1>(nullptr, nullptr, up, down, 0.f, N, K, r, out_sum, stream);
  if (e != cudaSuccess) { set_error("absmean launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}

extern "C" int b200lora_svd_epilogue(float* U, const float* S, float* Vh, float hi, int do_scale,
                                     int N, int K, int r, b200lora_stream_t stream_) {
  cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_);
  g_err[0] = 0;
  if (!U || !Vh || (do_scale && !S)) { set_error("null pointer argument"); return B200LORA_E_ARG; }
  cudaError_t e = launch_svd_epilogue(U, S, Vh, hi, do_scale, N, K, r, stream);
  if (e != cudaSuccess) { set_error("svd_epilogue launch failed: %s", cudaGetErrorString(e)); return B200LORA_E_CUDA; }
  return 0;
}
