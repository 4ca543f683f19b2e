// Fused LoRA GEMM for sm_100a: TMA -> smem (128B swizzle) -> tcgen05.mma -> TMEM -> epilogue.
//
//   T = X · S1ᵀ                       [M, r]
//   Y = X · Wᵀ + bias + (s·T) · S2ᵀ   [M, Nout]
//
// One persistent, warp-specialised kernel (1 CTA per SM, 512 threads):
//   warp 0       TMA producer: per k-block X tile [128 x 64], W tile [BN x 64] and the 16 packed
//                rank rows S1 [16 x 64] (b200lora_pack_lora) right behind the W rows
//   warp 1       MMA issuer (one thread): per k-block 4 x tcgen05.mma (K=16 each) with
//                N = BN+16 — the rank rows ride along as 16 extra accumulator columns (= T);
//                then ONE more K=16 MMA, [X | sT]·[Wᵀ ; S2ᵀ], adds the low-rank update.
//                While the T step of tile i is in flight it already issues the first k-blocks
//                of tile i+1 into the other accumulator stage.
//   warp 2       TMEM allocator, then S2 loader: writes the [BN x 16] S2 tile of every output
//                tile into swizzled smem (any dtype / strides: forward reads B, backward Aᵀ)
//   warps 4-7    T step: T columns TMEM -> regs -> T_out (global) and s·T -> smem as the
//                A-operand block of the tail MMA
//   warps 8-15   epilogue: accumulator TMEM -> regs -> +bias -> round -> smem transpose ->
//                coalesced 16-byte global stores (two warps per TMEM lane quarter)
// Accumulators are double-buffered in TMEM (2 x 256 columns): the epilogue of tile i overlaps
// the main loop of tile i+1.
#include "ptx.cuh"
#include "common.cuh"

namespace b200 {

namespace tc {
constexpr int BM = 128;          // rows per tile == TMEM lanes == UMMA M
constexpr int BK = 64;           // k-block: 64 16-bit elements = one 128-byte swizzle row
constexpr int RP = 16;           // padded rank (extra MMA N columns / one UMMA K step)
constexpr int UK = 16;           // UMMA K for 16-bit inputs
constexpr int ACC_STRIDE = 256;  // TMEM columns per accumulator stage
constexpr int EPI_COLS = 32;     // columns per epilogue unit
constexpr int MAX_STAGES = 8;
constexpr int MAX_CHUNKS = 7;    // BN <= 224
constexpr int NUM_THREADS = 512;
constexpr int NUM_EPI_WARPS = 8;
constexpr int LOOKAHEAD_KB = 2;  // k-blocks of the next tile issued before the tail MMA

constexpr int A_TILE_BYTES = BM * 128;                            // 16 KB
constexpr int T_TILE_BYTES = BM * 128;                            // 16 KB (32 B of each row used)
constexpr int STAGING_BYTES = NUM_EPI_WARPS * 32 * EPI_COLS * 2;  // 8 warps x 2 KB
constexpr int BAR_BYTES = 1024;

__host__ __device__ inline int b_tile_bytes(int BN) { return (BN + RP) * 128; }
__host__ __device__ inline int stage_bytes(int BN) { return A_TILE_BYTES + b_tile_bytes(BN); }
__host__ __device__ inline int fixed_bytes(int BN) {
  return T_TILE_BYTES + BN * 128 + STAGING_BYTES + BAR_BYTES;
}
}  // namespace tc

template <typename TIO>
__device__ __forceinline__ uint32_t pack2(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// watchdog site ids
enum : uint32_t {
  SITE_PROD_EMPTY = 1, SITE_MMA_ACC_EMPTY, SITE_MMA_FULL, SITE_MMA_BS, SITE_MMA_TSM,
  SITE_LD_BS_EMPTY, SITE_T_TFULL, SITE_T_PREV_TAIL, SITE_EPI_ACC
};

template <typename TIO>
__global__ void __launch_bounds__(tc::NUM_THREADS, 1)
lora_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                    const __grid_constant__ CUtensorMap tmS, const TcParams p) {
  using namespace tc;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem base is only guaranteed 16B aligned by the ABI -> align by hand
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int BN = p.BN;
  const int stages = p.stages;
  const int stg_bytes = stage_bytes(BN);
  uint8_t* stage_base = smem;                                  // stages x [A tile | B tile]
  uint8_t* t_tile = stage_base + (size_t)stages * stg_bytes;   // 16 KB
  uint8_t* bs_tile = t_tile + T_TILE_BYTES;                    // BN x 128 B
  uint8_t* staging = bs_tile + (size_t)BN * 128;               // 16 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + STAGING_BYTES);
  uint64_t* full_bar = bars;                      // [MAX_STAGES] TMA -> MMA
  uint64_t* empty_bar = bars + MAX_STAGES;        // [MAX_STAGES] MMA -> TMA
  uint64_t* tfull_bar = bars + 2 * MAX_STAGES;    // [2] main loop done: T columns valid
  uint64_t* accfull_bar = tfull_bar + 2;          // [2] tail MMA done: accumulator valid
  uint64_t* accempty_bar = accfull_bar + 2;       // [2] epilogue drained the accumulator
  uint64_t* tsm_full_bar = accempty_bar + 2;      // [1] T warps wrote the sT smem tile
  uint64_t* bs_full_bar = tsm_full_bar + 1;       // [1] loader wrote the S2 smem tile
  uint64_t* bs_empty_bar = bs_full_bar + 1;       // [1] tail MMA consumed the S2 / sT tiles
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bs_empty_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_m_blk * p.num_n_blk;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmS);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);   // TMA thread (arrive.expect_tx)
      mbar_init(&empty_bar[s], 1);  // tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull_bar[a], 1);
      mbar_init(&accfull_bar[a], 1);
      mbar_init(&accempty_bar[a], NUM_EPI_WARPS * 32);
    }
    mbar_init(tsm_full_bar, 128);
    mbar_init(bs_full_bar, 1);
    mbar_init(bs_empty_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // =============================== TMA producer =====================================
    if (lane == 0) {
      const uint64_t pol_x = policy_evict_normal();
      const uint64_t pol_w = policy_evict_last();
      int s = 0;
      uint32_t ph = 0;
      const uint32_t tx_bytes = A_TILE_BYTES + (BN + RP) * 128;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int n_blk = tile % p.num_n_blk;
        const int m_blk = tile / p.num_n_blk;
        for (int kb = 0; kb < p.num_k_blk; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1, SITE_PROD_EMPTY);
          uint8_t* a_dst = stage_base + (size_t)s * stg_bytes;
          uint8_t* b_dst = a_dst + A_TILE_BYTES;
          mbar_arrive_expect_tx(&full_bar[s], tx_bytes);
          tma_load_2d(a_dst, &tmX, &full_bar[s], kb * BK, m_blk * BM, pol_x);
          tma_load_2d(b_dst, &tmW, &full_bar[s], kb * BK, n_blk * BN, pol_w);
          // the 16 rank rows (packed [16,Kc], zero padded) land right behind the W rows
          tma_load_2d(b_dst + (size_t)BN * 128, &tmS, &full_bar[s], kb * BK, 0, pol_w);
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer =======================================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      // issue k-blocks [kb0, kb1) of the tile with local index `it`
      auto issue = [&](int it, int kb0, int kb1) {
        const uint32_t d_tmem = tmem_base + (it & 1) * ACC_STRIDE;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[s], ph, SITE_MMA_FULL);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(stage_base + (size_t)s * stg_bytes);
          const uint64_t a_desc = make_desc_sw128(a_addr);
          const uint64_t b_desc = make_desc_sw128(a_addr + A_TILE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            // advance both descriptors by 32 bytes (16 elements) inside the swizzle row
            umma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, p.idesc_main, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);  // smem slot reusable once these MMAs have read it
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      };
      const int nkb = p.num_k_blk;
      const int look = nkb < LOOKAHEAD_KB ? nkb : LOOKAHEAD_KB;
      int it = 0;
      int tile = blockIdx.x;
      if (tile < num_tiles) {
        mbar_wait(&accempty_bar[0], 1, SITE_MMA_ACC_EMPTY);
        tc_fence_after_sync();
        issue(0, 0, nkb);
      }
      for (; tile < num_tiles; tile += gridDim.x, ++it) {
        const int a = it & 1;
        umma_commit(&tfull_bar[a]);  // T columns (and the partial accumulator) are complete
        const bool has_next = tile + (int)gridDim.x < num_tiles;
        if (has_next) {
          // hide the T step behind the first k-blocks of the next tile (other TMEM stage)
          const int nit = it + 1;
          mbar_wait(&accempty_bar[nit & 1], ((nit >> 1) & 1) ^ 1, SITE_MMA_ACC_EMPTY);
          tc_fence_after_sync();
          issue(nit, 0, look);
        }
        // tail: one K=16 step  acc[:, 0:BN] += (sT) · S2ᵀ
        mbar_wait(bs_full_bar, it & 1, SITE_MMA_BS);
        mbar_wait(tsm_full_bar, it & 1, SITE_MMA_TSM);
        tc_fence_after_sync();
        umma_f16(tmem_base + a * ACC_STRIDE, make_desc_sw128(smem_u32(t_tile)),
                 make_desc_sw128(smem_u32(bs_tile)), p.idesc_tail, 1u);
        umma_commit(&accfull_bar[a]);
        umma_commit(bs_empty_bar);
        if (has_next) issue(it + 1, look, nkb);
      }
    }
  } else if (warp == 2) {
    // =============================== S2 loader ========================================
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int n_blk = tile % p.num_n_blk;
      // S2 tile for this output tile: BN rows x 16 k (k >= r zero), rows >= Nout zero
      mbar_wait(bs_empty_bar, (it & 1) ^ 1, SITE_LD_BS_EMPTY);
      for (int n = lane; n < BN; n += 32) {
        const long long gn = (long long)n_blk * BN + n;
        float v[RP];
#pragma unroll
        for (int j = 0; j < RP; ++j)
          v[j] = (j < p.r && gn < p.Nout)
                     ? load_as_float(p.S2, p.ab_dtype, gn * p.s2_rs + j * p.s2_cs)
                     : 0.f;
        uint4 q0, q1;
        q0.x = pack2<TIO>(v[0], v[1]);   q0.y = pack2<TIO>(v[2], v[3]);
        q0.z = pack2<TIO>(v[4], v[5]);   q0.w = pack2<TIO>(v[6], v[7]);
        q1.x = pack2<TIO>(v[8], v[9]);   q1.y = pack2<TIO>(v[10], v[11]);
        q1.z = pack2<TIO>(v[12], v[13]); q1.w = pack2<TIO>(v[14], v[15]);
        uint8_t* row = bs_tile + n * 128;
        *reinterpret_cast<uint4*>(row + ((0 ^ (n & 7)) << 4)) = q0;
        *reinterpret_cast<uint4*>(row + ((1 ^ (n & 7)) << 4)) = q1;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(bs_full_bar);
    }
  } else if (warp >= 4 && warp < 8) {
    // =============================== T step ===========================================
    const int qd = warp & 3;                 // TMEM lane quarter
    const int row = qd * 32 + lane;          // row inside the 128-row tile
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    TIO* t_out = reinterpret_cast<TIO*>(p.T_out);
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int n_blk = tile % p.num_n_blk;
      const int m_blk = tile / p.num_n_blk;
      const int a = it & 1;
      const long long gm = (long long)m_blk * BM + row;
      mbar_wait(&tfull_bar[a], (it >> 1) & 1, SITE_T_TFULL);
      tc_fence_after_sync();
      uint32_t tv[16];
      tmem_ld_x16(tmem_base + lane_addr + a * ACC_STRIDE + BN, tv);
      tmem_ld_wait();
      if (t_out != nullptr && n_blk == 0 && gm < p.M) {
        TIO* trow_out = t_out + gm * p.r;
#pragma unroll
        for (int j = 0; j < 16; ++j)   // static register indexing; r is warp-uniform
          if (j < p.r) trow_out[j] = TIO(__uint_as_float(tv[j]));
      }
      float tf[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) tf[j] = __uint_as_float(tv[j]) * p.scale;
      uint4 q0, q1;
      q0.x = pack2<TIO>(tf[0], tf[1]);   q0.y = pack2<TIO>(tf[2], tf[3]);
      q0.z = pack2<TIO>(tf[4], tf[5]);   q0.w = pack2<TIO>(tf[6], tf[7]);
      q1.x = pack2<TIO>(tf[8], tf[9]);   q1.y = pack2<TIO>(tf[10], tf[11]);
      q1.z = pack2<TIO>(tf[12], tf[13]); q1.w = pack2<TIO>(tf[14], tf[15]);
      // the sT tile is single-buffered: the previous tile's tail MMA must have read it
      if (it > 0) mbar_wait(&accfull_bar[(it - 1) & 1], ((it - 1) >> 1) & 1, SITE_T_PREV_TAIL);
      uint8_t* trow = t_tile + row * 128;
      *reinterpret_cast<uint4*>(trow + ((0 ^ (row & 7)) << 4)) = q0;
      *reinterpret_cast<uint4*>(trow + ((1 ^ (row & 7)) << 4)) = q1;
      fence_proxy_async_smem();
      tc_fence_before_sync();
      mbar_arrive(tsm_full_bar);
    }
  } else if (warp >= 8) {
    // =============================== epilogue =========================================
    const int ew = warp - 8;                 // 0..7
    const int qd = warp & 3;                 // TMEM lane quarter
    const int half = ew >> 2;                // this warp takes chunks c with (c & 1) == half
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    uint8_t* my_staging = staging + ew * (32 * EPI_COLS * 2);
    TIO* Y = reinterpret_cast<TIO*>(p.Y);
    // write phase: lane owns row `lane`, 4 x 16 B; 64B-swizzle keeps 8 lanes on distinct banks
    uint8_t* wr_row = my_staging + lane * 64;
    const int wr_sw = (lane >> 1) & 3;
    // read phase: instruction i covers rows 8i..8i+7; lane -> row 8i + lane/4, 16B piece lane%4
    const int rd_q = lane & 3;
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int n_blk = tile % p.num_n_blk;
      const int m_blk = tile / p.num_n_blk;
      const int a = it & 1;
      const uint32_t acc_addr = tmem_base + lane_addr + a * ACC_STRIDE;
      const int n_chunks = BN / EPI_COLS;
      // bias of my chunks: lane e holds bias[n0 + 32c + e]; fetched before the accumulator
      // wait so the latency hides behind the MMAs, broadcast by shuffle below
      float bias_reg[(MAX_CHUNKS + 1) / 2];
#pragma unroll
      for (int ci = 0; ci < (MAX_CHUNKS + 1) / 2; ++ci) {
        const int c = 2 * ci + half;
        const int gn = n_blk * BN + c * EPI_COLS + lane;
        bias_reg[ci] = (p.bias != nullptr && c < n_chunks && gn < p.Nout)
                           ? load_as_float(p.bias, p.bias_dtype, gn) : 0.f;
      }
      mbar_wait(&accfull_bar[a], (it >> 1) & 1, SITE_EPI_ACC);
      tc_fence_after_sync();
#pragma unroll
      for (int ci = 0; ci < (MAX_CHUNKS + 1) / 2; ++ci) {
        const int c = 2 * ci + half;
        const int gn0 = n_blk * BN + c * EPI_COLS;
        if (c < n_chunks && gn0 < p.Nout) {   // warp-uniform
          uint32_t v[32];
          tmem_ld_x32(acc_addr + c * EPI_COLS, v);
          tmem_ld_wait();
          if (p.bias != nullptr) {
#pragma unroll
            for (int e = 0; e < 32; ++e)
              v[e] = __float_as_uint(__uint_as_float(v[e]) +
                                     __shfl_sync(0xffffffffu, bias_reg[ci], e));
          }
          __syncwarp();   // every lane has finished reading the previous chunk from staging
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 o;
            o.x = pack2<TIO>(__uint_as_float(v[q * 8 + 0]), __uint_as_float(v[q * 8 + 1]));
            o.y = pack2<TIO>(__uint_as_float(v[q * 8 + 2]), __uint_as_float(v[q * 8 + 3]));
            o.z = pack2<TIO>(__uint_as_float(v[q * 8 + 4]), __uint_as_float(v[q * 8 + 5]));
            o.w = pack2<TIO>(__uint_as_float(v[q * 8 + 6]), __uint_as_float(v[q * 8 + 7]));
            *reinterpret_cast<uint4*>(wr_row + ((q ^ wr_sw) << 4)) = o;
          }
          __syncwarp();
          const int gcol = gn0 + rd_q * 8;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int rr = i * 8 + (lane >> 2);
            const uint4 o = *reinterpret_cast<const uint4*>(
                my_staging + rr * 64 + ((rd_q ^ ((rr >> 1) & 3)) << 4));
            const long long gm = (long long)m_blk * BM + qd * 32 + rr;
            if (gm < p.M && gcol < p.Nout)
              *reinterpret_cast<uint4*>(Y + gm * p.ldy + gcol) = o;
          }
        }
      }
      // all TMEM reads of this accumulator stage are complete (tmem_ld_wait above)
      tc_fence_before_sync();
      mbar_arrive(&accempty_bar[a]);
    }
  }

  // =============================== teardown ===========================================
  tc_fence_before_sync();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
int tc_pick_bn(int Nout) {
  const char* env = getenv("B200LORA_BN");
  if (env) {
    int v = atoi(env);
    if (v >= 32 && v <= 224 && v % 32 == 0) return v;
  }
  // prefer the largest tile <= 192 that divides Nout; fall back to 128 (ragged last tile)
  const int cands[] = {192, 160, 128, 96, 64};
  for (int c : cands)
    if (Nout % c == 0) return c;
  if (Nout <= 32) return 32;
  if (Nout <= 64) return 64;
  if (Nout <= 96) return 96;
  return 128;
}

int tc_smem_bytes(int BN, int* stages_out) {
  const int budget = 232448 - 1024;  // 227 KB minus manual alignment slack
  int stages = (budget - tc::fixed_bytes(BN)) / tc::stage_bytes(BN);
  if (stages > tc::MAX_STAGES) stages = tc::MAX_STAGES;
  if (stages_out) *stages_out = stages;
  return stages * tc::stage_bytes(BN) + tc::fixed_bytes(BN) + 1024;
}

template <typename TIO>
cudaError_t launch_tc(const CUtensorMap& tmX, const CUtensorMap& tmW, const CUtensorMap& tmS,
                      const TcParams& p, int smem_bytes, int num_sms, cudaStream_t stream) {
  auto kern = lora_gemm_tc_kernel<TIO>;
  static thread_local int configured_smem = 0;
  if (configured_smem < smem_bytes) {
    cudaError_t e =
        cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    if (e != cudaSuccess) return e;
    configured_smem = smem_bytes;
  }
  const int num_tiles = p.num_m_blk * p.num_n_blk;
  const int grid = num_tiles < num_sms ? num_tiles : num_sms;
  kern<<<grid, tc::NUM_THREADS, smem_bytes, stream>>>(tmX, tmW, tmS, p);
  return cudaGetLastError();
}

template cudaError_t launch_tc<__nv_bfloat16>(const CUtensorMap&, const CUtensorMap&,
                                              const CUtensorMap&, const TcParams&, int, int,
                                              cudaStream_t);
template cudaError_t launch_tc<__half>(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                       const TcParams&, int, int, cudaStream_t);

unsigned int tc_read_watchdog() {
  unsigned int v = 0;
  cudaMemcpyFromSymbol(&v, g_watchdog_code, sizeof(v));
  return v;
}

}  // namespace b200
