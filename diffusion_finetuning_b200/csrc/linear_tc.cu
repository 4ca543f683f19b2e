// Fused LoRA GEMM for sm_100a: TMA -> smem (128B swizzle) -> tcgen05.mma -> TMEM -> TMA store.
//
//   T = X · S1ᵀ                       [M, r]
//   Y = X · Wᵀ + bias + (s·T) · S2ᵀ   [M, Nout]        ( = [X | sT | 1] · [Wᵀ ; S2ᵀ ; bias] )
//
// One persistent, warp-specialised kernel (1 CTA per SM, 512 threads).  Every CTA owns a
// CONTIGUOUS range of output tiles in n-fastest order, so consecutive tiles share their 128-row
// block of X: T depends on the row block only, so it is produced once per row block ("new_m"
// tile) and reused by the following tiles of that block.
//
//   warp 0       TMA producer.  Per k-block: X tile [128 x 64], W tile [BN x 64] and, on new_m
//                tiles, the 16 packed rank rows S1 [16 x 64] behind the W rows.  Per tile one more
//                "k-block": the [BN x 64] slice of the packed S2 (+ bias column) for the tail MMA.
//   warp 1       MMA issuer (one thread).  Main loop: 4 x tcgen05.mma (K=16) per k-block with
//                N = BN (+16 on new_m tiles: the rank rows ride along as 16 extra accumulator
//                columns = T).  Tail: one (two if r + bias > 16) K=16 MMA of the [sT | 1] smem tile
//                against the S2 slice: low-rank update AND bias land in the accumulator.
//                On a new_m tile the main loop of the NEXT tile is issued before the tail, which
//                hides the TMEM -> registers -> smem round trip of T.
//   warp 2       TMEM allocator
//   warps 4-7    T step (new_m tiles only): T columns TMEM -> regs -> T_out (global, for the
//                backward) and [s·T | 1] -> smem as the A operand of the tail MMAs
//   warps 8-15   epilogue: accumulator TMEM -> regs (software-pipelined tcgen05.ld) -> round ->
//                swizzled smem staging -> TMA store (bounds clipped by the tensor map)
// Accumulators are multi-buffered in TMEM (2-4 stages): the epilogue of tile i overlaps the main
// loops of the following tiles.
#include <mutex>

#include "ptx.cuh"
#include "common.cuh"

namespace b200 {

namespace tc {
constexpr int BM = 128;          // rows per tile == TMEM lanes == UMMA M
constexpr int BK = 64;           // k-block: 64 16-bit elements = one 128-byte swizzle row
constexpr int RP = 16;           // padded rank (extra MMA N columns / one UMMA K step)
constexpr int UK = 16;           // UMMA K for 16-bit inputs
constexpr int EPI_COLS = 32;     // columns per epilogue unit
constexpr int MAX_STAGES = 8;
constexpr int MAX_ACC = 4;       // TMEM accumulator stages
constexpr int MAX_CI = 4;        // epilogue units per warp (BN <= 256)
constexpr int NUM_THREADS = 512;
constexpr int NUM_EPI_WARPS = 8;

constexpr int A_TILE_BYTES = BM * 128;                            // 16 KB
constexpr int T_TILE_BYTES = BM * 128;                            // 16 KB (32/64 B of each row used)
constexpr int STAGING_BYTES = NUM_EPI_WARPS * 2 * 32 * EPI_COLS * 2;  // 8 warps x 2 buffers x 2 KB
constexpr int BAR_BYTES = 1024;

__host__ __device__ inline int b_tile_bytes(int BN) { return (BN + RP) * 128; }
__host__ __device__ inline int stage_bytes(int BN) { return A_TILE_BYTES + b_tile_bytes(BN); }
__host__ __device__ inline int fixed_bytes() { return T_TILE_BYTES + STAGING_BYTES + BAR_BYTES; }
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
  SITE_PROD_EMPTY = 1, SITE_MMA_ACC_EMPTY, SITE_MMA_FULL, SITE_MMA_TAIL_FULL, SITE_MMA_TSM,
  SITE_T_TFULL, SITE_T_FREE, SITE_EPI_ACC
};

// which tiles start a new 128-row block for this CTA / get the next tile's main loop issued
// before their own tail (the same predicate drives the producer and the MMA issuer)
struct Sched {
  int t0, t1, nnb;
  __device__ __forceinline__ bool new_m(int tile) const { return tile == t0 || (tile % nnb) == 0; }
  __device__ __forceinline__ bool look(int tile) const {
    return new_m(tile) && tile + 1 < t1 && ((tile + 1) % nnb) != 0;
  }
};

// timeline slots (DBG builds): [role*256 + it*8 + k]
#define TS(role, it_, k)                                                              \
  do {                                                                                \
    if (DBG && (int)blockIdx.x == p.dbg_cta && (it_) < 32)                            \
      p.dbg[(role) * 256 + (it_) * 8 + (k)] = clock64();                              \
  } while (0)

template <typename TIO, bool DBG>
__global__ void __launch_bounds__(tc::NUM_THREADS, 1)
lora_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmW,
                    const __grid_constant__ CUtensorMap tmS1, const __grid_constant__ CUtensorMap tmS2,
                    const __grid_constant__ CUtensorMap tmY, const TcParams p) {
  using namespace tc;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // dynamic smem base is only guaranteed 16B aligned by the ABI -> align by hand
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int BN = p.BN;
  const int stages = p.stages;
  const int stg_bytes = stage_bytes(BN);
  uint8_t* stage_base = smem;                                  // stages x [A tile | B tile]
  uint8_t* t_tile = stage_base + (size_t)stages * stg_bytes;   // 16 KB  [sT | 1] tile
  uint8_t* staging = t_tile + T_TILE_BYTES;                    // 32 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + STAGING_BYTES);
  uint64_t* full_bar = bars;                      // [MAX_STAGES] TMA -> MMA
  uint64_t* empty_bar = bars + MAX_STAGES;        // [MAX_STAGES] MMA -> TMA
  uint64_t* accfull_bar = bars + 2 * MAX_STAGES;  // [MAX_ACC] tail MMA done: accumulator valid
  uint64_t* accempty_bar = accfull_bar + MAX_ACC; // [MAX_ACC] epilogue drained the accumulator
  uint64_t* tfull_bar = accempty_bar + MAX_ACC;   // [1] main loop of a new_m tile done: T valid
  uint64_t* tsm_full_bar = tfull_bar + 1;         // [1] T warps wrote the [sT | 1] smem tile
  uint64_t* st_free_bar = tsm_full_bar + 1;       // [1] all tails of the previous row block done
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(st_free_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = p.num_m_blk * p.num_n_blk;
  Sched sc;
  sc.t0 = (int)((long long)blockIdx.x * num_tiles / gridDim.x);
  sc.t1 = (int)((long long)(blockIdx.x + 1) * num_tiles / gridDim.x);
  sc.nnb = p.num_n_blk;
  const int NST = p.nst;
  const int ACC_STRIDE = p.acc_stride;
  if (threadIdx.x == 0) TS(0, 0, 0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    tma_prefetch_desc(&tmW);
    tma_prefetch_desc(&tmS1);
    tma_prefetch_desc(&tmS2);
    tma_prefetch_desc(&tmY);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 1);   // TMA thread (arrive.expect_tx)
      mbar_init(&empty_bar[s], 1);  // tcgen05.commit
    }
    for (int a = 0; a < MAX_ACC; ++a) {
      mbar_init(&accfull_bar[a], 1);
      mbar_init(&accempty_bar[a], NUM_EPI_WARPS * 32);
    }
    mbar_init(tfull_bar, 1);
    mbar_init(tsm_full_bar, 128);
    mbar_init(st_free_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) TS(0, 0, 1);

  if (warp == 0) {
    // =============================== TMA producer =====================================
    if (lane == 0) {
      const uint64_t pol_x = policy_evict_normal();
      const uint64_t pol_w = policy_evict_last();
      int s = 0;
      uint32_t ph = 0;
      auto load_main = [&](int tile, bool nm) {
        const int n_blk = tile % p.num_n_blk;
        const int m_blk = tile / p.num_n_blk;
        const uint32_t tx_bytes = A_TILE_BYTES + BN * 128 + (nm ? RP * 128 : 0);
        for (int kb = 0; kb < p.num_k_blk; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1, SITE_PROD_EMPTY);
          uint8_t* a_dst = stage_base + (size_t)s * stg_bytes;
          uint8_t* b_dst = a_dst + A_TILE_BYTES;
          mbar_arrive_expect_tx(&full_bar[s], tx_bytes);
          tma_load_2d(a_dst, &tmX, &full_bar[s], kb * BK, m_blk * BM, pol_x);
          tma_load_2d(b_dst, &tmW, &full_bar[s], kb * BK, n_blk * BN, pol_w);
          // the 16 rank rows (packed [16,Kc], zero padded) land right behind the W rows
          if (nm) tma_load_2d(b_dst + (size_t)BN * 128, &tmS1, &full_bar[s], kb * BK, 0, pol_w);
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      };
      auto load_tail = [&](int tile) {
        const int n_blk = tile % p.num_n_blk;
        mbar_wait(&empty_bar[s], ph ^ 1, SITE_PROD_EMPTY);
        uint8_t* b_dst = stage_base + (size_t)s * stg_bytes + A_TILE_BYTES;
        mbar_arrive_expect_tx(&full_bar[s], BN * 128);
        tma_load_2d(b_dst, &tmS2, &full_bar[s], 0, n_blk * BN, pol_w);
        if (++s == stages) { s = 0; ph ^= 1; }
      };
      int pit = 0;
      for (int tile = sc.t0; tile < sc.t1;) {
        const bool lk = sc.look(tile);
        TS(1, pit, 0);
        load_main(tile, sc.new_m(tile));
        TS(1, pit, 1);
        if (lk) load_main(tile + 1, false);
        load_tail(tile);
        if (lk) load_tail(tile + 1);
        TS(1, pit, 2);
        tile += lk ? 2 : 1;
        pit += lk ? 2 : 1;
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer =======================================
    if (lane == 0) {
      int s = 0;
      uint32_t ph = 0;
      const uint64_t t_desc = make_desc_sw128(smem_u32(t_tile));
      auto main_loop = [&](int st, bool nm) {
        const uint32_t d_tmem = tmem_base + st * ACC_STRIDE;
        const uint32_t idesc = nm ? p.idesc_main_t : p.idesc_main;
        for (int kb = 0; kb < p.num_k_blk; ++kb) {
          mbar_wait(&full_bar[s], ph, SITE_MMA_FULL);
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(stage_base + (size_t)s * stg_bytes);
          const uint64_t a_desc = make_desc_sw128(a_addr);
          const uint64_t b_desc = make_desc_sw128(a_addr + A_TILE_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k) {
            // advance both descriptors by 32 bytes (16 elements) inside the swizzle row
            umma_f16(d_tmem, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);  // smem slot reusable once these MMAs have read it
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      };
      auto tail = [&](int st) {
        mbar_wait(&full_bar[s], ph, SITE_MMA_TAIL_FULL);
        tc_fence_after_sync();
        const uint64_t b_desc =
            make_desc_sw128(smem_u32(stage_base + (size_t)s * stg_bytes + A_TILE_BYTES));
        for (int k = 0; k < p.tail_ksteps; ++k)
          umma_f16(tmem_base + st * ACC_STRIDE, t_desc + 2 * k, b_desc + 2 * k, p.idesc_tail, 1u);
        umma_commit(&empty_bar[s]);
        umma_commit(&accfull_bar[st]);
        if (++s == stages) { s = 0; ph ^= 1; }
      };
      int it = 0;
      int e = 0;   // index of the current row-block event (new_m tiles seen so far)
      for (int tile = sc.t0; tile < sc.t1;) {
        const bool nm = sc.new_m(tile);
        const bool lk = sc.look(tile);
        const int st = it % NST;
        TS(2, it, 0);
        mbar_wait(&accempty_bar[st], ((it / NST) & 1) ^ 1, SITE_MMA_ACC_EMPTY);
        tc_fence_after_sync();
        TS(2, it, 1);
        // every tail of the previous row block has been issued: once they complete the
        // [sT | 1] tile may be overwritten by the T warps
        if (nm && e > 0) umma_commit(st_free_bar);
        main_loop(st, nm);
        TS(2, it, 2);
        if (nm) umma_commit(tfull_bar);   // T columns complete
        int st2 = 0;
        if (lk) {
          // hide the T round trip behind the next tile's main loop (another TMEM stage)
          st2 = (it + 1) % NST;
          mbar_wait(&accempty_bar[st2], (((it + 1) / NST) & 1) ^ 1, SITE_MMA_ACC_EMPTY);
          tc_fence_after_sync();
          main_loop(st2, false);
          TS(2, it, 3);
        }
        if (nm) {
          mbar_wait(tsm_full_bar, e & 1, SITE_MMA_TSM);
          tc_fence_after_sync();
          ++e;
        }
        TS(2, it, 4);
        tail(st);
        if (lk) tail(st2);
        TS(2, it, 5);
        it += lk ? 2 : 1;
        tile += lk ? 2 : 1;
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // =============================== T step ===========================================
    const int qd = warp & 3;                 // TMEM lane quarter
    const int row = qd * 32 + lane;          // row inside the 128-row tile
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    const uint32_t trow = smem_u32(t_tile) + row * 128;
    const int sw = row & 7;
    TIO* t_out = reinterpret_cast<TIO*>(p.T_out);
    int it = 0, e = 0;
    for (int tile = sc.t0; tile < sc.t1; ++tile, ++it) {
      if (!sc.new_m(tile)) continue;
      const int n_blk = tile % p.num_n_blk;
      const int m_blk = tile / p.num_n_blk;
      const int st = it % NST;
      const long long gm = (long long)m_blk * BM + row;
      if (threadIdx.x == 128) TS(3, it, 0);
      mbar_wait(tfull_bar, e & 1, SITE_T_TFULL);
      tc_fence_after_sync();
      if (threadIdx.x == 128) TS(3, it, 1);
      uint32_t tv[16];
      tmem_ld_x16(tmem_base + lane_addr + st * ACC_STRIDE + BN, tv);
      tmem_ld_wait();
      if (t_out != nullptr && n_blk == 0 && gm < p.M) {
        TIO* dst = t_out + gm * p.r;
        if (p.r == 4) {
          uint2 q;
          q.x = pack2<TIO>(__uint_as_float(tv[0]), __uint_as_float(tv[1]));
          q.y = pack2<TIO>(__uint_as_float(tv[2]), __uint_as_float(tv[3]));
          *reinterpret_cast<uint2*>(dst) = q;
        } else if (p.r == 8 || p.r == 16) {
          uint4 q;
          q.x = pack2<TIO>(__uint_as_float(tv[0]), __uint_as_float(tv[1]));
          q.y = pack2<TIO>(__uint_as_float(tv[2]), __uint_as_float(tv[3]));
          q.z = pack2<TIO>(__uint_as_float(tv[4]), __uint_as_float(tv[5]));
          q.w = pack2<TIO>(__uint_as_float(tv[6]), __uint_as_float(tv[7]));
          *reinterpret_cast<uint4*>(dst) = q;
          if (p.r == 16) {
            q.x = pack2<TIO>(__uint_as_float(tv[8]), __uint_as_float(tv[9]));
            q.y = pack2<TIO>(__uint_as_float(tv[10]), __uint_as_float(tv[11]));
            q.z = pack2<TIO>(__uint_as_float(tv[12]), __uint_as_float(tv[13]));
            q.w = pack2<TIO>(__uint_as_float(tv[14]), __uint_as_float(tv[15]));
            *reinterpret_cast<uint4*>(dst + 8) = q;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)   // static register indexing; r is warp-uniform
            if (j < p.r) dst[j] = TIO(__uint_as_float(tv[j]));
        }
      }
      float tf[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        tf[j] = __uint_as_float(tv[j]) * p.scale;     // columns >= r are exact zeros
        if (p.has_bias && j == p.r) tf[j] = 1.0f;     // the bias row of S2 rides on a ones column
      }
      // the [sT | 1] tile is single-buffered: the previous row block's tails must have read it
      if (e > 0) mbar_wait(st_free_bar, (e - 1) & 1, SITE_T_FREE);
      st_shared_v4(trow + ((0 ^ sw) << 4), pack2<TIO>(tf[0], tf[1]), pack2<TIO>(tf[2], tf[3]),
                   pack2<TIO>(tf[4], tf[5]), pack2<TIO>(tf[6], tf[7]));
      st_shared_v4(trow + ((1 ^ sw) << 4), pack2<TIO>(tf[8], tf[9]), pack2<TIO>(tf[10], tf[11]),
                   pack2<TIO>(tf[12], tf[13]), pack2<TIO>(tf[14], tf[15]));
      if (p.tail_ksteps > 1) {   // r == 16 with bias: the ones column is element 16
        st_shared_v4(trow + ((2 ^ sw) << 4), pack2<TIO>(1.0f, 0.f), 0u, 0u, 0u);
        st_shared_v4(trow + ((3 ^ sw) << 4), 0u, 0u, 0u, 0u);
      }
      fence_proxy_async_smem();
      tc_fence_before_sync();
      mbar_arrive(tsm_full_bar);
      if (threadIdx.x == 128) TS(3, it, 2);
      ++e;
    }
  } else if (warp >= 8) {
    // =============================== epilogue =========================================
    const int ew = warp - 8;                 // 0..7
    const int qd = warp & 3;                 // TMEM lane quarter
    const int half = ew >> 2;                // this warp takes units c with (c & 1) == half
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    // two 2 KB staging buffers per warp: [32 rows][64 B], 64B-swizzled as the Y tensor map expects
    uint8_t* my_staging = staging + ew * (2 * 32 * EPI_COLS * 2);
    const uint32_t stg_row = smem_u32(my_staging) + lane * 64;
    const int wr_sw = (lane >> 1) & 3;
    const int n_units = BN / EPI_COLS;
    int it = 0;
    uint32_t buf = 0;
    for (int tile = sc.t0; tile < sc.t1; ++tile, ++it) {
      const int n_blk = tile % p.num_n_blk;
      const int m_blk = tile / p.num_n_blk;
      const int st = it % NST;
      const uint32_t acc_addr = tmem_base + lane_addr + st * ACC_STRIDE;
      if (lane == 0 && qd == 0) TS(4 + half, it, 0);
      mbar_wait(&accfull_bar[st], (it / NST) & 1, SITE_EPI_ACC);
      tc_fence_after_sync();
      if (lane == 0 && qd == 0) TS(4 + half, it, 1);
      uint32_t v[2][32];
      // unit c is live if it exists in the tile and starts inside the matrix (warp-uniform)
      auto live = [&](int c) { return c < n_units && n_blk * BN + c * EPI_COLS < p.Nout; };
      if (live(half)) tmem_ld_x32(acc_addr + half * EPI_COLS, v[0]);
#pragma unroll
      for (int ci = 0; ci < MAX_CI; ++ci) {
        const int c = 2 * ci + half;
        if (live(c)) {
          tmem_ld_wait_dep(v[ci & 1]);
          if (ci + 1 < MAX_CI && live(c + 2))
            tmem_ld_x32(acc_addr + (c + 2) * EPI_COLS, v[(ci + 1) & 1]);
          // the TMA store that last read this staging buffer (two units ago) must be done with it
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          const uint32_t dst = stg_row + buf * (32 * EPI_COLS * 2);
          const uint32_t* w = v[ci & 1];
#pragma unroll
          for (int q = 0; q < 4; ++q)
            st_shared_v4(dst + ((q ^ wr_sw) << 4),
                         pack2<TIO>(__uint_as_float(w[q * 8 + 0]), __uint_as_float(w[q * 8 + 1])),
                         pack2<TIO>(__uint_as_float(w[q * 8 + 2]), __uint_as_float(w[q * 8 + 3])),
                         pack2<TIO>(__uint_as_float(w[q * 8 + 4]), __uint_as_float(w[q * 8 + 5])),
                         pack2<TIO>(__uint_as_float(w[q * 8 + 6]), __uint_as_float(w[q * 8 + 7])));
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmY, my_staging + buf * (32 * EPI_COLS * 2), n_blk * BN + c * EPI_COLS,
                         m_blk * BM + qd * 32);
            tma_store_commit();
          }
          buf ^= 1;
        }
      }
      // all TMEM reads of this accumulator stage are complete (wait::ld above)
      tc_fence_before_sync();
      mbar_arrive(&accempty_bar[st]);
      if (lane == 0 && qd == 0) TS(4 + half, it, 2);
    }
    if (lane == 0) tma_store_wait_all<0>();
    if (lane == 0 && qd == 0) TS(4 + half, 31, 7);
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
// Tile width: minimise  waves x (MMA columns x k-blocks + fixed per-tile cost)  over the candidate
// widths.  `rows` = output rows per tile (128 single CTA, 256 CTA pair), `units` = CTAs or pairs
// that run concurrently.  The fixed cost (tail MMA hand-off, accumulator switch, epilogue drain;
// ~0.45 us measured against 1.03 us for 176 columns x 5 k-blocks) is what makes wide tiles win on
// wide layers even when they leave a ragged last tile.
int tc_pick_bn(long long M, int Kc, int Nout, int rows, int units) {
  const char* env = getenv("B200LORA_BN");
  if (env) {
    int v = atoi(env);
    if (v >= 32 && v <= 224 && v % 32 == 0) return v;
  }
  if (Nout <= 32) return 32;
  if (Nout <= 64) return 64;
  const int cands[] = {160, 128, 192, 224, 96, 64};   // ties go to the earlier entry
  const double kb = (double)((Kc + 63) / 64) + 0.25;
  const double fixed = 385.0;
  const long long mb = (M + rows - 1) / rows;
  int best = 128;
  double best_cost = 1e300;
  for (int c : cands) {
    const long long tiles = mb * ((Nout + c - 1) / c);
    const long long waves = (tiles + units - 1) / units;
    const double cost = (double)waves * ((c + 16) * kb + fixed);
    if (cost < best_cost * 0.999) { best_cost = cost; best = c; }
  }
  return best;
}

// TMEM accumulator stages for a tile of BN (+16 rank) columns, and their column stride
void tc_tmem_plan(int BN, int* nst, int* stride) {
  int n = 512 / (BN + tc::RP);
  if (n > tc::MAX_ACC) n = tc::MAX_ACC;
  if (n < 1) n = 1;
  *nst = n;
  *stride = (512 / n) & ~31;
}

int tc_smem_bytes(int BN, int* stages_out) {
  const int budget = 232448 - 1024;  // 227 KB minus manual alignment slack
  int stages = (budget - tc::fixed_bytes()) / tc::stage_bytes(BN);
  if (stages > tc::MAX_STAGES) stages = tc::MAX_STAGES;
  if (stages_out) *stages_out = stages;
  return stages * tc::stage_bytes(BN) + tc::fixed_bytes() + 1024;
}

template <typename TIO>
cudaError_t launch_tc(const CUtensorMap& tmX, const CUtensorMap& tmW, const CUtensorMap& tmS1,
                      const CUtensorMap& tmS2, const CUtensorMap& tmY, const TcParams& p,
                      int smem_bytes, int num_sms, cudaStream_t stream) {
  const int num_tiles = p.num_m_blk * p.num_n_blk;
  const int grid = num_tiles < num_sms ? num_tiles : num_sms;
  if (p.dbg != nullptr) {   // timeline build (tools/timeline.py)
    auto dk = lora_gemm_tc_kernel<TIO, true>;
    cudaError_t e = cudaFuncSetAttribute(dk, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    if (e != cudaSuccess) return e;
    dk<<<grid, tc::NUM_THREADS, smem_bytes, stream>>>(tmX, tmW, tmS1, tmS2, tmY, p);
    return cudaGetLastError();
  }
  auto kern = lora_gemm_tc_kernel<TIO, false>;
  // process-wide attribute: raise it once to the device maximum (a per-thread cache would let the
  // autograd thread lower the cap under the main thread)
  static std::once_flag cap_once;
  static cudaError_t cap_err = cudaSuccess;
  std::call_once(cap_once, [&] {
    cap_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
  });
  if (cap_err != cudaSuccess) return cap_err;
  kern<<<grid, tc::NUM_THREADS, smem_bytes, stream>>>(tmX, tmW, tmS1, tmS2, tmY, p);
  return cudaGetLastError();
}

template cudaError_t launch_tc<__nv_bfloat16>(const CUtensorMap&, const CUtensorMap&,
                                              const CUtensorMap&, const CUtensorMap&,
                                              const CUtensorMap&, const TcParams&, int, int,
                                              cudaStream_t);
template cudaError_t launch_tc<__half>(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                       const CUtensorMap&, const CUtensorMap&, const TcParams&, int,
                                       int, cudaStream_t);

unsigned int tc_read_watchdog() {
  unsigned int v = 0;
  cudaMemcpyFromSymbol(&v, g_watchdog_code, sizeof(v));
  return v;
}

}  // namespace b200
