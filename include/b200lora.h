/*
 * b200lora.h — C-ABI of the B200-native LoRA-adapted-layer hot path.
 *
 * This is the drop-in boundary one layer below the reference's Python API
 * (the reference, lora_diffusion/lora.py, has no FFI of its own: its "operator
 * interface" is the nn.Module LoraInjectedLinear, lora.py:32-50).  Every entry
 * point below names the reference code it replaces.
 *
 * Conventions
 *   - all functions are extern "C", return 0 on success or a negative
 *     B200LORA_E_* code, never throw, never allocate or free caller memory;
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - matrices are row-major; "ld" arguments are leading dimensions in ELEMENTS;
 *   - every launch goes to the caller's stream (pass the torch current stream);
 *   - functions are re-entrant; the only internal mutable state is a mutex-guarded
 *     cache of TMA descriptors keyed on (pointer, shape, box).
 *
 * Notation: X [M,K], W0 [N,K], bias [N], A = lora_down.weight [r,K],
 * B = lora_up.weight [N,r], s = scale, T = X·Aᵀ [M,r], Y [M,N].
 */
#ifndef B200LORA_H_
#define B200LORA_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* opaque: a cudaStream_t passed through as void* so this header needs no CUDA includes */
typedef void* b200lora_stream_t;

enum {
  B200LORA_OK = 0,
  B200LORA_E_ARCH = -1,   /* device is not sm_100 (tensor-core path) / no device     */
  B200LORA_E_SHAPE = -2,  /* unsupported / inconsistent shape                         */
  B200LORA_E_ALIGN = -3,  /* pointer or leading dimension not aligned for this path   */
  B200LORA_E_DTYPE = -4,  /* unsupported dtype combination                            */
  B200LORA_E_CUDA = -5,   /* a CUDA runtime/driver call failed (see last_error)       */
  B200LORA_E_ARG = -6     /* null pointer / bad enum                                  */
};

/* element types */
enum { B200LORA_F32 = 0, B200LORA_F16 = 1, B200LORA_BF16 = 2 };

/* kernel family selection */
enum {
  B200LORA_IMPL_AUTO = 0, /* tcgen05 for 16-bit io when shape/alignment allow, else SIMT */
  B200LORA_IMPL_SIMT = 1, /* CUDA-core FFMA kernels (true-fp32 accumulate; fp32 mode)  */
  B200LORA_IMPL_TC = 2    /* tcgen05/TMEM/TMA kernel; error if not applicable          */
};

/* layout of the big frozen operand handed to b200lora_lora_gemm */
enum {
  B200LORA_W_NK = 0, /* W given as [Nout, Kc] (contraction index contiguous)  */
  B200LORA_W_KN = 1  /* W given as [Kc, Nout] (output index contiguous)       */
};

int b200lora_version(void);
/* human-readable text of the last failure on the calling thread ("" if none) */
const char* b200lora_last_error(void);
/* 1 if the current device can run the tcgen05 path (compute capability 10.x) */
int b200lora_tc_supported(void);

/*
 * Diagnostics (not part of the drop-in surface; used by tools/ and by B200LORA_DEBUG_ERR=1):
 *   b200lora_watchdog_code    non-zero after a tensor-core kernel's bounded barrier wait gave up:
 *                             0x80000000 | wait-site << 16 | block index (every mbarrier wait in the tcgen05
 *                             kernels is bounded so a protocol bug traps instead of wedging the GPU).
 *   b200lora_peek_cuda_error  the CUDA runtime's pending error code, without clearing it.
 *   b200lora_debug_timeline   `buf` = device buffer of 6*256 int64: the next tcgen05 launches record a clock64
 *                             timeline of CTA `cta` into it (tools/timeline.py); NULL switches it off.
 */
unsigned int b200lora_watchdog_code(void);
int b200lora_peek_cuda_error(void);
void b200lora_debug_timeline(void* buf, int cta);

/*
 * The generic fused kernel behind forward and backward-dX:
 *
 *     T = X · S1ᵀ                       [M, r]   (optionally written to T_out)
 *     Y = X · Wᵀ + bias + (s·T) · S2ᵀ   [M, Nout]
 *
 * X [M,Kc] (ldx), W per w_layout (ldw), S1 logical [r,Kc] with element (j,k) at
 * S1[j*s1_rs + k*s1_cs], S2 logical [Nout,r] with element (n,j) at S2[n*s2_rs + j*s2_cs].
 * io_dtype is the type of X, W, Y and T_out; ab_dtype the type of S1/S2 (fp32 masters
 * under autocast are read directly); bias may be NULL.
 * The tcgen05 path fetches the rank rows and the tail block by TMA from S1_pack / S2_pack (see
 * b200lora_pack_lora) and ignores S1/S2; the bias it adds is the one PACKED into S2_pack (`bias`
 * only says whether there is one).  The SIMT path reads S1/S2/bias with their strides and ignores
 * the packs.  Either set may be NULL if the selected path does not need it.
 * On the tcgen05 path this is ONE kernel: the rank rows ride along as 16 extra accumulator
 * columns of the same MMA and the up-projection AND the bias are one more K=16 MMA step
 * ([X | sT | 1]·[Wᵀ ; S2ᵀ ; bias]), so no separate low-rank GEMM or elementwise add runs.
 */
int b200lora_lora_gemm(const void* X, int64_t ldx,
                       const void* W, int64_t ldw, int w_layout,
                       const void* bias, int bias_dtype,
                       const void* S1, int64_t s1_rs, int64_t s1_cs,
                       const void* S1_pack /* [16,Kc] io dtype, rows >= r zero; tcgen05 path */,
                       const void* S2, int64_t s2_rs, int64_t s2_cs, int ab_dtype,
                       const void* S2_pack /* [Nout,64] io dtype: S2 | bias | 0; tcgen05 path */,
                       float scale,
                       void* Y, int64_t ldy,
                       void* T_out /* nullable, [M,r] compact */,
                       int64_t M, int Kc, int Nout, int r,
                       int io_dtype, int impl, b200lora_stream_t stream);

/*
 * Pack the LoRA factors for the tcgen05 path (one tiny launch per weight version) into ONE
 * buffer of b200lora_pack_elems(K, N) = 80·(K+N) elements of io_dtype, laid out as
 *   A_pack  [16,K] @ 0                 A  rows zero-padded to 16 (forward rank rows)
 *   Bt_pack [16,N] @ 16·K              Bᵀ rows zero-padded to 16 (backward rank rows)
 *   Bn_pack [N,64] @ 16·(K+N)          B rows | bias[n] in column r | zeros (forward tail block)
 *   At_pack [K,64] @ 16·(K+N) + 64·N   Aᵀ rows | zeros                     (backward tail block)
 * fp32 masters (autocast) are rounded to io_dtype here, as autocast would round them; `bias`
 * (nullable) is the frozen bias of the base layer (lora.py:34), added by the forward tail MMA.
 */
int64_t b200lora_pack_elems(int K, int N);
int b200lora_pack_lora(const void* A, const void* B, int ab_dtype,
                       const void* bias, int bias_dtype, void* pack,
                       int K, int N, int r, int io_dtype, b200lora_stream_t stream);

/*
 * The same pack for the CONCATENATED problem of up to 4 sibling layers that read one input (fused
 * to_q/to_k/to_v), written straight from the layers' own factors: A_cat = [A_0; A_1; ...]
 * (sum of r_each rows), B block-diagonal [nseg*Nseg, r_total] with scales[i] folded into block i, biases
 * concatenated (bias, or any bias[i], may be NULL).  A[i] [r_each[i], K], B[i] [Nseg, r_each[i]] in ab_dtype,
 * bias[i] [Nseg] in bias_dtype; the pointer ARRAYS are host memory (read during the call).
 * pack: b200lora_pack_elems(K, nseg*Nseg) elements of io_dtype.
 */
int b200lora_pack_lora_fused(const void* const* A, const void* const* B, const void* const* bias,
                             const float* scales, const int* r_each, int nseg, int ab_dtype,
                             int bias_dtype, void* pack, int K, int Nseg, int io_dtype,
                             b200lora_stream_t stream);

/*
 * The two pack launches with the fp32 -> io_dtype conversion of the layer's INPUT riding along: x_dst[i] = rnd(x_src[i])
 * for x_elems elements (both 16-byte aligned) in the same launch.  Replaces the input cast autocast performs in front
 * of lora.py:49 (`F.linear` under autocast casts its fp32 input) when the factors are trained and the pack is rebuilt
 * on every call anyway: one launch instead of two, and the consuming GEMM's prologue overlaps it (programmatic launch).
 */
int b200lora_pack_lora_cast(const void* A, const void* B, int ab_dtype,
                            const void* bias, int bias_dtype, void* pack,
                            int K, int N, int r, int io_dtype,
                            const float* x_src, void* x_dst, int64_t x_elems, b200lora_stream_t stream);
int b200lora_pack_lora_fused_cast(const void* const* A, const void* const* B, const void* const* bias,
                                  const float* scales, const int* r_each, int nseg, int ab_dtype,
                                  int bias_dtype, void* pack, int K, int Nseg, int io_dtype,
                                  const float* x_src, void* x_dst, int64_t x_elems, b200lora_stream_t stream);

/*
 * Forward of LoraInjectedLinear — replaces lora.py:49-50
 *   `self.linear(input) + self.lora_up(self.lora_down(input)) * self.scale`
 * (3 F.linear + mul + add in the reference; one kernel here).
 * T_out (nullable) receives T = X·Aᵀ for the backward; pass NULL under no_grad.
 */
int b200lora_linear_fwd(const void* X, const void* W0, const void* bias, int bias_dtype,
                        const void* A, const void* B, int ab_dtype,
                        const void* pack /* from b200lora_pack_lora; NULL on the SIMT path */,
                        float scale,
                        void* Y, void* T_out,
                        int64_t M, int K, int N, int r,
                        int io_dtype, int impl, b200lora_stream_t stream);

/*
 * Backward w.r.t. the input — replaces the autograd of lora.py:50:
 *   dX = dY·W0 + s·(dY·B)·A ;  dT_out = dY·B  [M,r]
 * W0 is frozen, so no dW0 is produced.  w_layout says how W is laid out:
 *   B200LORA_W_NK: W is W0ᵀ stored [K,N] (a cached transpose; required by the tcgen05 path)
 *   B200LORA_W_KN: W is W0 itself [N,K] (accepted by the SIMT path only)
 */
int b200lora_linear_bwd_dx(const void* dY, const void* W, int w_layout,
                           const void* A, const void* B, int ab_dtype,
                           const void* pack /* from b200lora_pack_lora; NULL on the SIMT path */,
                           float scale,
                           void* dX, void* dT_out,
                           int64_t M, int K, int N, int r,
                           int io_dtype, int impl, b200lora_stream_t stream);

/*
 * LoRA-Conv2d — `conv(x) + lora_up(lora_down(x)) * scale` with lora_down a conv of the same
 * geometry to r channels and lora_up a 1x1 conv.  NOT in the reference snapshot (lora.py has only
 * LoraInjectedLinear); upstream cloneofsimo/lora `LoraInjectedConv2d` semantics, SURVEY §8(c).
 * Implicit GEMM over NHWC activations: rows = output pixels (n, oh, ow), K = (kh, kw, 64-channel
 * blocks) fetched by TMA im2col loads, same tcgen05 CTA-pair kernel as the linear layer.
 *   Cp = ceil(Cin/64)*64, Coutp = ceil(Cout/64)*64, Cz = Coutp + 64.
 *
 * repack_weight  W [Cout,Cin,R,S] (OIHW) -> mode 0: Wk [Cout, R*S*Cp] (forward operand)
 *                                          mode 1: Wb [Cin,  R*S*Cz] (backward operand: flipped,
 *                                                  transposed; LoRA columns filled by pack_lora)
 * pack_lora      A [r,Cin,R,S], B [Cout,r], bias -> A_pack [16, R*S*Cp], Bn_pack [Cout,64] and
 *                (Wb != NULL) the A-columns of Wb
 * gemm           Y[N,Ho,Wo,Cout] = conv(X; Wk) + bias + (s·T)·S2ᵀ,  T = conv(X; S1) -> T_out [M,r]
 *                (backward dX: X = Z from zcat, Wk = Wb, zero packs, pad' = (R-1)*dil - pad)
 * zcat           backward: dT = dY·B [M,r] and Z [M,Cz] = [dY | 0 | s·dT | 0]
 * dA             backward: dA[r,Cin,R,S] (fp32, ADDED into) += s·Σ_m dT[m,:]ᵀ · im2col(X)[m,:]
 * dB uses b200lora_lora_grads(dY, NULL, T, NULL, ..., dA = NULL, dB).
 */
int b200lora_conv2d_repack_weight(const void* W_oihw, int w_dtype, void* out, int Cout, int Cin,
                                  int R, int S, int mode, int io_dtype, b200lora_stream_t stream);
int b200lora_conv2d_pack_lora(const void* A, const void* B, int ab_dtype, const void* bias,
                              int bias_dtype, void* A_pack, void* Bn_pack, void* Wb /*nullable*/,
                              int Cout, int Cin, int R, int S, int r, int io_dtype,
                              b200lora_stream_t stream);
int b200lora_conv2d_gemm(const void* X_nhwc, const void* Wk, int has_bias, const void* S1_pack,
                         const void* S2_pack, float scale, void* Y_nhwc, void* T_out /*nullable*/,
                         int N, int H, int W, int C, int Cout, int R, int S, int stride_h,
                         int stride_w, int pad_h, int pad_w, int dil_h, int dil_w, int r,
                         int io_dtype, b200lora_stream_t stream);
int b200lora_conv2d_zcat(const void* dY, const void* B, int ab_dtype, float scale, void* Z, void* dT,
                         int64_t M, int Cout, int r, int io_dtype, b200lora_stream_t stream);
int b200lora_conv2d_grad_a(const void* X_nhwc, const void* dT, float scale, float* dA, int N, int H,
                       int W, int C, int R, int S, int stride_h, int stride_w, int pad_h, int pad_w,
                       int dil_h, int dil_w, int r, int io_dtype, b200lora_stream_t stream);

/*
 * Token reductions for the LoRA gradients — replaces the autograd of lora.py:50:
 *   dB[n,j] (+)= s · Σ_m dY[m,n]·T[m,j]      [N,r]
 *   dA[j,k] (+)= s · Σ_m dT[m,j]·X[m,k]      [r,K]
 * dA/dB are fp32 accumulators (e.g. slices of the flat gradient arena).  The kernel ADDS
 * into them (zero them first for a plain gradient).  Either of dA/dB may be NULL to skip.
 * deterministic=1 uses a two-stage reduction through `workspace` (bytes from
 * b200lora_lora_grads_workspace) instead of fp32 atomics.  1 <= r <= 64; ranks above 16 run as windows of
 * 16 ranks (one launch each) over the same dY / X.
 */
int b200lora_lora_grads(const void* dY, const void* X, const void* T, const void* dT,
                        float scale, float* dA, float* dB,
                        int64_t M, int K, int N, int r, int io_dtype,
                        int deterministic, void* workspace, int64_t workspace_bytes,
                        b200lora_stream_t stream);
int64_t b200lora_lora_grads_workspace(int64_t M, int K, int N, int r);

/*
 * No-allocation variant of the deterministic reductions above.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  Same arithmetic and launches as b200lora_lora_grads(...,
 * deterministic = 1, workspace) — bit-identical results — but the two-stage partial sums live in a device array
 * built into the library (B200LORA_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until exit)
 * instead of a caller workspace.  The call itself issues no cudaMalloc / cudaFree and calls no allocating
 * library routine; the caller still owns every input and output.
 *   - shapes whose b200lora_lora_grads_workspace(M, K, N, r) exceeds B200LORA_STATIC_SCRATCH_BYTES are
 *     rejected with B200LORA_E_SHAPE and a message naming both sizes (nothing is launched);
 *   - all calls on a device share the one array, which keeps the previous call's contents (every slot a call
 *     reads it has written first, so nothing needs clearing): calls that OVERLAP on the GPU — different
 *     streams, different threads — race on it.  Serialise them on one stream.
 */
#define B200LORA_STATIC_SCRATCH_BYTES (4 << 20)
int b200lora_lora_grads_static(const void* dY, const void* X, const void* T, const void* dT,
                               float scale, float* dA, float* dB,
                               int64_t M, int K, int N, int r, int io_dtype,
                               b200lora_stream_t stream);

/*
 * b200lora_linear_bwd_dx / _fused with an FP32 dX: the same kernels, the accumulator stored unrounded.  Replaces the
 * bf16 -> fp32 cast autograd needs behind the backward of lora.py:50 when the layer's input was fp32 under autocast
 * (LayerNorm outputs are): one elementwise launch and a [M, K] round trip less per layer.  Wt = W0ᵀ [K, N] (the W_NK
 * operand of the transposed problem); 16-bit io, tcgen05 CTA-pair kernel only (other cases: the plain entry + a cast).
 */
int b200lora_linear_bwd_dx_f32out(const void* dY, const void* Wt, const void* pack, float scale, float* dX,
                                  void* dT_out, int64_t M, int K, int N, int r, int io_dtype,
                                  b200lora_stream_t stream);
int b200lora_linear_bwd_dx_fused_f32out(const void* const* dY, const int64_t* ld_dy, int nseg,
                                        const void* Wt, const void* pack, float* dX, void* dT_out,
                                        int64_t M, int K, int Nseg, int r_total, int io_dtype,
                                        b200lora_stream_t stream);

/*
 * GEGLU projection with the activation in the GEMM epilogue — replaces, for a LoRA-injected
 * `GEGLU.proj` (lora.py:53 lists GEGLU among the default targets), the caller's
 *     hidden, gate = proj(x).chunk(2, dim=-1); out = hidden * gelu(gate)
 * around lora.py:49-50.  W0 is proj's weight [2*d_ff, K] (hidden rows first), pack as for b200lora_linear_fwd with
 * N = 2*d_ff.  out [M, d_ff] = hidden * gelu(gate) with the reference's intermediate roundings; Y [M, 2*d_ff] (the
 * pre-activation, needed by the backward) is written when non-NULL; T [M, r] as in b200lora_linear_fwd.
 * 16-bit io, d_ff % 64 == 0, 1 <= r <= 16, CTA-pair kernel only (everything else: run the unfused composition).
 * b200lora_geglu_bwd: dY [M, 2*d_ff] = [dout * gelu(gate) | dout * hidden * gelu'(gate)] in one pass over dout and Y
 * (the autograd of the two caller-side ops); dY then feeds b200lora_linear_bwd_dx / b200lora_lora_grads unchanged.
 */
int b200lora_linear_fwd_geglu(const void* X, const void* W0, const void* pack, int has_bias, float scale,
                              void* Y, void* out, void* T_out, int64_t M, int K, int d_ff, int r,
                              int io_dtype, b200lora_stream_t stream);
int b200lora_geglu_bwd(const void* dout, const void* Y, void* dY, int64_t M, int d_ff, int io_dtype,
                       b200lora_stream_t stream);

/*
 * Fused sibling projections — layers that evaluate lora.py:49-50 on the SAME input (the to_q / to_k /
 * to_v of one attention block, all called on `hidden_states` by the caller of LoraInjectedLinear)
 * run as ONE problem over their concatenated outputs.
 *   forward : b200lora_lora_gemm on W_cat [nseg*Nseg, K], b200lora_pack_lora(A_cat [r_total, K],
 *             B_bd [nseg*Nseg, r_total] block diagonal with each layer's scale folded in), scale = 1;
 *             Y_cat [M, nseg*Nseg] holds the layers' outputs side by side, T_out is [M, r_total].
 *   backward: the callers' gradients arrive as nseg separate tensors dY[i] [M, Nseg] (row pitch
 *             ld_dy[i] elements):
 *     b200lora_linear_bwd_dx_fused  dX = sum_i dY_i·W0_i + (sum_i s_i·dY_i·B_i)·A_cat, dT_out [M, r_total];
 *                                   Wt = W_catᵀ stored [K, nseg*Nseg]; `pack` as in the forward.
 *     b200lora_lora_grads_fused     dB[i] += s_i·dY_iᵀ·T[:, ranks of i]   (fp32 [Nseg, r_each[i]])
 *                                   dA[i] += (dTᵀ·X)[ranks of i]          (fp32 [r_each[i], K])
 *                                   in ONE launch (nseg + 1 reductions).
 * 16-bit io only, nseg <= 3, Nseg % 64 == 0, sum of ranks <= 16.
 */
int b200lora_linear_bwd_dx_fused(const void* const* dY, const int64_t* ld_dy, int nseg,
                                 const void* Wt, const void* pack, void* dX, void* dT_out,
                                 int64_t M, int K, int Nseg, int r_total, int io_dtype,
                                 b200lora_stream_t stream);
int b200lora_lora_grads_fused(const void* const* dY, const int64_t* ld_dy, int nseg, const void* X,
                              const void* T, const void* dT, const float* scales,
                              float* const* dA, float* const* dB, int64_t M, int K, int Nseg,
                              const int* r_each, int io_dtype, b200lora_stream_t stream);

/*
 * Optimizer step over the flat arenas — replaces clip_grad_norm_ + AdamW.step() over the
 * LoRA parameters (train_lora_dreambooth.py:676, 878-888; cli_lora_pti.py:448-451) when parameters,
 * gradients and both moments are contiguous fp32 arrays of n elements in the same order
 * (dp.LoraGradArena + dp.ArenaAdamW).  max_grad_norm <= 0 disables the global-norm clip; otherwise
 * grads are scaled in place by min(1, max_grad_norm / (||g|| + 1e-6)) as clip_grad_norm_ does.
 * scratch: 2 floats on the device owned by the caller — [0] norm accumulator (reset here every call),
 * [1] the step count t (initialise to 0; advanced here), kept on the device so the step can be captured
 * in a CUDA graph.  Three stream operations (memset, norm, update); counts as one C-ABI call.
 */
int b200lora_arena_adamw(float* params, float* grads, float* exp_avg, float* exp_avg_sq, int64_t n,
                         float lr, float beta1, float beta2, float eps, float weight_decay,
                         float max_grad_norm, float* scratch, b200lora_stream_t stream);

/*
 * Flat pass over LoRA artefacts: out[i] = rnd(alpha·a[i]) (+ rnd(beta·b[i]), then rnd of the sum),
 * every rounding to in_dtype as ATen's temporaries are, then converted to out_dtype.
 *   b != NULL  replaces the `lpl` mix of two LoRA lists, cli_lora_add.py:49-58
 *              (`alpha * x1.data + (1 - alpha) * x2.data`, bit-identical to the CPU result);
 *   b == NULL  with alpha = 1 is the dtype conversion of a whole parameter arena for the `.pt` writer
 *              (lora.py:201-213: `.to(torch.float16)` per tensor).
 * a, b [n] in in_dtype, out [n] in out_dtype; out may alias a when the dtypes are equal.
 */
int b200lora_flat_axpby(const void* a, const void* b, void* out, int64_t n, float alpha, float beta,
                        int in_dtype, int out_dtype, b200lora_stream_t stream);

/*
 * Textual-inversion step — replaces, for the placeholder-token rows of the CLIP token
 * embedding, cli_lora_pti.py:309-346: the dense embedding gradient + dense AdamW over [vocab, D]
 * (cli_lora_pti.py:651-657), the norm decay of the placeholder rows (:317-335) and the restore of every
 * other row from a clone of the table (:280, :344-346).  Only P rows ever change, so only P rows are touched.
 *   b200lora_ti_grad_rows  grad_rows[p, :] += sum over tokens t with token_ids[t] == rows[p] of dY[t, :]
 *                          dY [n_tokens, D] in io_dtype (gradient of the embedding OUTPUT), token_ids [n_tokens]
 *                          int64, rows [P] int64 (placeholder token ids), grad_rows [P, D] fp32 (ADDED into;
 *                          tokens are summed in index order: bit-reproducible).
 *   b200lora_ti_step       AdamW (torch.optim.AdamW semantics: decoupled decay, bias correction with `step`
 *                          >= 1, eps outside the corrected sqrt) on master_rows [P, D] fp32 with moments
 *                          exp_avg / exp_avg_sq [P, D]; then, when lambda_ >= 0,
 *                          w <- w / max(|w|, 1e-12) * (|w| + lambda_ * (target_norm - |w|))
 *                          (lambda_ = min(1, 100 * lr), target_norm = 0.4 in the reference; lambda_ < 0 skips
 *                          the decay = clip_ti_decay False); then table[rows[p], :] = master_rows[p, :] cast to
 *                          table_dtype.  norms_out (nullable) [2, P] fp32: row norms before and after the decay
 *                          (the reference prints both).  Replayable form (the call sits in a captured CUDA
 *                          graph): lr_dev (nullable) points at ONE device float that overrides `lr` (and, when
 *                          lambda_ >= 0, lambda_ becomes min(1, 100 * lr_dev[0])); step_dev (nullable) [P] device
 *                          floats hold the per-row step count, read as step = step_dev[p] + 1 and written back.
 */
int b200lora_ti_grad_rows(const void* dY, const int64_t* token_ids, int64_t n_tokens, int D,
                          const int64_t* rows, int P, float* grad_rows, int io_dtype,
                          b200lora_stream_t stream);
int b200lora_ti_step(void* table, int table_dtype, int D, const int64_t* rows, int P, float* master_rows,
                     const float* grad_rows, float* exp_avg, float* exp_avg_sq, int step, float lr,
                     float beta1, float beta2, float eps, float weight_decay, float lambda_,
                     float target_norm, float* norms_out, const float* lr_dev, float* step_dev,
                     b200lora_stream_t stream);

/*
 * Weight merge — replaces lora.py:423  `weight + alpha * (up @ down).type(weight.dtype)`
 * (cli_lora_add.py:78 runs it on the CPU).  Wout may alias Win (in-place).
 * up [N,r], down [r,K] in lora_dtype; W in w_dtype.  Rounding follows the reference:
 * the product is rounded to lora_dtype, cast to w_dtype, multiplied by alpha, then added.
 */
int b200lora_merge(const void* Win, void* Wout, const void* up, const void* down,
                   float alpha, int N, int K, int r, int w_dtype, int lora_dtype,
                   b200lora_stream_t stream);

/*
 * The same merge for MANY matrices in ONE launch — replaces the per-layer loop of lora.py:410-424
 * (`weight_apply_lora`: 144 UNet + 48 text-encoder matrices in cli_lora_add.py:72-110).  All weights share w_dtype,
 * all factors share lora_dtype, ranks may differ per item; same rounding as b200lora_merge; Wout may alias Win.
 *   b200lora_merge_batch_plan  host only: validates `items` (a HOST array) and writes the device-format problem
 *                              table (b200lora_merge_batch_workspace(n_items) bytes) into HOST memory `table_host`;
 *                              returns the tile count and the largest rank the launch has to cover.
 *   b200lora_merge_batch_run   one launch over the table, which the caller has copied to DEVICE memory (any way it
 *                              likes; the table can be kept and re-run, e.g. for a sweep over alpha).
 */
typedef struct b200lora_merge_item {
  const void* Win;
  void* Wout;
  const void* up;      /* [N, r] */
  const void* down;    /* [r, K] */
  int32_t N, K, r, reserved;
} b200lora_merge_item;
int64_t b200lora_merge_batch_workspace(int n_items);
int b200lora_merge_batch_plan(const b200lora_merge_item* items, int n_items, int w_dtype, void* table_host,
                              int64_t table_bytes, int64_t* total_tiles, int* max_rank);
int b200lora_merge_batch_run(const void* table_dev, int n_items, int64_t total_tiles, int max_rank, float alpha,
                             int w_dtype, int lora_dtype, b200lora_stream_t stream);

/*
 * mean(|up @ down|) for one layer — replaces lora.py:744-746 (inspect_lora "moved" metric).
 * out_sum receives Σ|up@down| (fp32, ADDED into; divide by N*K on the host).
 */
int b200lora_absmean_accum(const void* up, const void* down, float* out_sum,
                           int N, int K, int r, int lora_dtype, b200lora_stream_t stream);

/*
 * Epilogue of the SVD distillation — replaces cli_svd.py:73-84: U ← clamp(U·diag(S), ±hi),
 * Vh ← clamp(Vh, ±hi).  The SVD itself stays on cuSOLVER and the quantile `hi` is computed
 * by the caller (cli_svd.py:79-80); pass hi<0 to skip clamping (used to build the quantile
 * input).  U [N,r] fp32 in/out, S [r], Vh [r,K] fp32 in/out.
 */
int b200lora_svd_epilogue(float* U, const float* S, float* Vh, float hi, int do_scale,
                          int N, int K, int r, b200lora_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200LORA_H_ */
