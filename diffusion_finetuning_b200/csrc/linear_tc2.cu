// Fused LoRA GEMM for sm_100a, CTA-PAIR version: tcgen05.mma.cta_group::2 (UMMA M = 256 over two
// SMs of one TPC), TMA -> smem (128B swizzle) -> TMEM -> TMA store.
//
//   T = X · S1ᵀ                       [M, r]
//   Y = X · Wᵀ + bias + (s·T) · S2ᵀ   [M, Nout]        ( = [X | sT | 1] · [Wᵀ ; S2ᵀ ; bias] )
//
// Why a pair: the single-CTA kernel (linear_tc.cu) moves (128 + BN + 16)·128 B from L2 per k-block
// per 128 x BN tile and was measured AT the L2->SM throughput cap (~12 TB/s) on every large shape.
// In a pair each CTA loads its own 128 rows of X but only HALF of the W tile; the MMA reads the
// two halves from both CTAs' shared memory.  Per CTA and k-block: 16 KB + (BN+16)/2·128 B, i.e.
// ~30 % less L2 traffic at BN = 160, and half the smem footprint for W (6 pipeline stages).
//
// Per CTA (512 threads), roles as in linear_tc.cu:
//   warp 0       TMA producer (BOTH CTAs): own X rows + own half of [W rows ; 16 rank rows]; all
//                transaction bytes of a stage are signalled on the LEADER's full barrier
//   warp 1       MMA issuer (leader CTA only, one thread): main loop N = BN+16 (the rank rows ride
//                along as 16 extra accumulator columns = T), tail K=16 step(s) [sT | 1]·[S2 | bias]ᵀ;
//                tcgen05.commit multicasts to the barriers of both CTAs
//   warp 2       TMEM allocator (cta_group::2, both CTAs)
//   warps 4-7    T step for the CTA's own 128 rows (first tile of a row block only); arrives on
//                the leader's barrier through shared::cluster
//   warps 8-15   epilogue for the CTA's own 128 rows: pipelined tcgen05.ld -> staging -> TMA store
// A pair owns a contiguous range of (256-row block, n-tile) pairs in n-fastest order.
#include <map>
#include <mutex>

#include "ptx.cuh"
#include "common.cuh"

namespace b200 {

namespace tc2 {
constexpr int BM = 128;          // rows per CTA == TMEM lanes
constexpr int BK = 64;
constexpr int RP = 16;
constexpr int UK = 16;
constexpr int EPI_COLS = 32;
constexpr int MAX_STAGES = 8;
constexpr int MAX_ACC = 4;
constexpr int MAX_CI = 4;
constexpr int NUM_THREADS = 512;
constexpr int NUM_EPI_WARPS = 8;
constexpr int A_TILE_BYTES = BM * 128;
constexpr int T_TILE_BYTES = BM * 128;
constexpr int STAGING_BYTES = NUM_EPI_WARPS * 2 * 32 * EPI_COLS * 2;
constexpr int BAR_BYTES = 1024;

__host__ __device__ inline int half_rows(int BN) { return (BN + RP) / 2; }   // B rows per CTA
__host__ __device__ inline int stage_bytes(int BN) { return A_TILE_BYTES + half_rows(BN) * 128; }
__host__ __device__ inline int fixed_bytes() { return T_TILE_BYTES + STAGING_BYTES + BAR_BYTES; }
}  // namespace tc2

template <typename TIO>
__device__ __forceinline__ uint32_t pack2b(float a, float b);
template <>
__device__ __forceinline__ uint32_t pack2b<__nv_bfloat16>(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
template <>
__device__ __forceinline__ uint32_t pack2b<__half>(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

enum : uint32_t {
  S2_PROD_EMPTY = 17, S2_MMA_ACC_EMPTY, S2_MMA_FULL, S2_MMA_TAIL_FULL, S2_MMA_TSM, S2_T_TFULL,
  S2_T_FREE, S2_EPI_ACC
};

struct Sched2 {
  int t0, t1, nnb;
  __device__ __forceinline__ bool new_m(int tile) const { return tile == t0 || (tile % nnb) == 0; }
  __device__ __forceinline__ bool look(int tile) const {
    return new_m(tile) && tile + 1 < t1 && ((tile + 1) % nnb) != 0;
  }
};

#define TS2(role, it_, k)                                                             \
  do {                                                                                \
    if (DBG && (int)blockIdx.x == p.dbg_cta && (it_) < 32)                            \
      p.dbg[(role) * 256 + (it_) * 8 + (k)] = clock64();                              \
  } while (0)

// MODE: 0 = plain (Y in the io dtype), 1 = GEGLU projection (hidden * gelu(gate) in the epilogue), 2 = fp32 Y.
// Compile-time, not a TcParams flag: as run-time branches the two extra epilogues cost the PLAIN launches of the training
// step ~0.5 us each (register ceiling, longer instruction stream).
template <typename TIO, bool DBG, int MODE>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(tc2::NUM_THREADS, 1)
lora_gemm_tc2_kernel(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmX1,
                     const __grid_constant__ CUtensorMap tmX2, const __grid_constant__ CUtensorMap tmWa,
                     const __grid_constant__ CUtensorMap tmWh, const __grid_constant__ CUtensorMap tmS1,
                     const __grid_constant__ CUtensorMap tmS2, const __grid_constant__ CUtensorMap tmY,
                     const __grid_constant__ CUtensorMap tmOut, const TcParams p) {
  using namespace tc2;
  constexpr bool GEGLU = MODE == 1;
  constexpr bool F32OUT = MODE == 2;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                             ~static_cast<uintptr_t>(1023));
  const int BN = p.BN;
  const int H = half_rows(BN);
  const int stages = p.stages;
  const int stg_bytes = stage_bytes(BN);
  uint8_t* stage_base = smem;
  uint8_t* t_tile = stage_base + (size_t)stages * stg_bytes;
  uint8_t* staging = t_tile + T_TILE_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(staging + STAGING_BYTES);
  uint64_t* full_bar = bars;                      // [MAX_STAGES] leader's is the live one
  uint64_t* empty_bar = bars + MAX_STAGES;        // [MAX_STAGES] per CTA (multicast commit)
  uint64_t* accfull_bar = bars + 2 * MAX_STAGES;  // [MAX_ACC]    per CTA (multicast commit)
  uint64_t* accempty_bar = accfull_bar + MAX_ACC; // [MAX_ACC]    leader's: 16 warp arrivals
  uint64_t* tfull_bar = accempty_bar + MAX_ACC;   // [1] per CTA (multicast commit)
  uint64_t* tsm_full_bar = tfull_bar + 1;         // [1] leader's: 256 thread arrivals
  uint64_t* st_free_bar = tsm_full_bar + 1;       // [1] per CTA (multicast commit)
  // relay mode (p.cluster_scope == 2): the peer's hand-off arrivals stay INSIDE the peer CTA (these two barriers),
  // a relay warp forwards each completed phase to the leader as an async-proxy transaction (bulk_copy_to_peer)
  uint64_t* accempty_loc = st_free_bar + 1;       // [MAX_ACC] peer's: 8 epilogue-warp arrivals
  uint64_t* tsm_loc = accempty_loc + MAX_ACC;     // [1] peer's: 128 thread arrivals
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tsm_loc + 1);
  uint8_t* relay_box = reinterpret_cast<uint8_t*>(bars) + 512;   // [MAX_ACC + 1][16 B] source (peer) / sink (leader)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int num_tiles = p.num_m_blk * p.num_n_blk;     // num_m_blk counts 256-row blocks here
  const int ncl = gridDim.x >> 1;
  const int cl = blockIdx.x >> 1;
  Sched2 sc;
  sc.t0 = (int)((long long)cl * num_tiles / ncl);
  sc.t1 = (int)((long long)(cl + 1) * num_tiles / ncl);
  sc.nnb = p.num_n_blk;
  const int NST = p.nst;
  const int ACC_STRIDE = p.acc_stride;
  if (threadIdx.x == 0) TS2(0, 0, 0);
  pdl_launch_dependents();   // the next launch may run its prologue (below) under this grid's tail

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmX);
    if (p.kseg) {
      tma_prefetch_desc(&tmX1);
      tma_prefetch_desc(&tmX2);
    }
    tma_prefetch_desc(&tmWa);
    tma_prefetch_desc(&tmWh);
    tma_prefetch_desc(&tmS1);
    tma_prefetch_desc(&tmS2);
    tma_prefetch_desc(&tmY);
    if (GEGLU) {
      tma_prefetch_desc(&tmX1);      // GEGLU mode: W with a [BN/2 x 64] box (hidden rows)
      tma_prefetch_desc(&tmX2);      //             W with an [8 x 64] box (first gate rows of the leader's half)
      tma_prefetch_desc(&tmOut);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full_bar[s], 2);   // leader: arrive.expect_tx; peer: remote arrive
      mbar_init(&empty_bar[s], 1);  // tcgen05.commit (multicast)
    }
    const bool relay = p.cluster_scope == 2;
    for (int a = 0; a < MAX_ACC; ++a) {
      mbar_init(&accfull_bar[a], 1);
      // direct modes: 16 warp arrivals (8 per CTA); relay mode: 8 local warps + the MMA warp's arrive.expect_tx
      mbar_init(&accempty_bar[a], relay ? NUM_EPI_WARPS + 1 : 2 * NUM_EPI_WARPS);
      mbar_init(&accempty_loc[a], NUM_EPI_WARPS);
    }
    mbar_init(tfull_bar, 1);
    mbar_init(tsm_full_bar, relay ? 129 : 256);
    mbar_init(tsm_loc, 128);
    mbar_init(st_free_bar, 1);
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2sm(tmem_slot, 512);
  }
  tc_fence_before_sync();
  cluster_sync_all();       // peer barriers initialised before any remote arrive / multicast commit
  tc_fence_after_sync();
  const uint32_t tmem_base = *tmem_slot;
  // W producer: the base-weight half-tiles of the first tile's first stages are requested BEFORE
  // griddepcontrol.wait.  W is the frozen weight (or its cached cast / transpose / conv repack): no kernel that can
  // still be running under programmatic launch writes it (only pack / GEMM / reduction kernels trigger early, and
  // they write packs, activations and gradients), so these loads overlap the tail of the previous kernel and the
  // first MMA finds its B operand already in shared memory.  X, the rank rows and the tail blocks come from earlier
  // kernels and stay behind the wait; the stage's expect_tx is posted by the X producer as before (a transaction
  // count may go negative until it is).
  int w_pre = 0;
  if (warp == 3 && p.w_prefetch && sc.t0 < sc.t1) {
    w_pre = p.num_k_blk < stages ? p.num_k_blk : stages;
    const uint64_t pol_w0 = policy_evict_last();
    const uint32_t full0_b = mapa_shared(smem_u32(&full_bar[0]), 0);
    const int n0 = (sc.t0 % p.num_n_blk) * BN;
    if (elect_one()) {
      for (int kb = 0; kb < w_pre; ++kb) {
        uint8_t* b_dst = stage_base + (size_t)kb * stg_bytes + A_TILE_BYTES;
        if (leader) tma_load_2d_2sm(b_dst, &tmWh, full0_b + kb * 8, kb * BK, n0, pol_w0);
        else tma_load_2d_2sm(b_dst, &tmWa, full0_b + kb * 8, kb * BK, n0 + H, pol_w0);
      }
    }
    __syncwarp();
  }
  pdl_wait();               // nothing above reads or writes global memory WRITTEN BY earlier kernels
  if (threadIdx.x == 0) TS2(0, 0, 1);

  // third producer (p.split_b): a CTA whose B half is TWO boxes per k-block (the peer's [W rows ; rank rows], the
  // GEGLU leader's [hidden rows ; first gate rows]) issues the second box from warp 2, idle otherwise between the
  // TMEM allocation and the teardown.  One thread sustains one {wait, expect_tx, tensor load} per ~420 cycles
  // (tools/l2_ingest.cu); off by default: measured equal on the conv shapes.
  const bool second_box = p.split_b && (GEGLU || !leader);
  if (warp == 0 || warp == 3 || (warp == 2 && second_box)) {
    // =============================== TMA producers (both CTAs) ========================
    // TMA issue is serial per thread: the loads of a stage are split over two or three threads.
    // warp 0 : the CTA's X rows + the arrive / expect_tx of the stage;
    // warp 3 : the CTA's half of [W rows ; rank rows] and the tail blocks;
    // warp 2 : the second box of that half, when there is one and p.split_b says so.
    {
      // the whole warp walks the schedule (warp-uniform control flow); one elected lane issues.
      // With `if (lane == 0)` around the loop ptxas wraps every uniform-datapath instruction
      // (UTMALDG / UTCHMMA / UTCBAR) in an ELECT + BRA.U.ANY convergence loop.
      const bool xrole = warp == 0;
      const bool box1 = warp == 3;                                   // first (or only) box of the B half
      const bool box2 = second_box ? warp == 2 : warp == 3;          // second box
      const uint64_t pol_x = policy_evict_normal();
      const uint64_t pol_w = policy_evict_last();
      int s = 0;
      uint32_t ph = 0;
      const uint32_t main_tx = 2u * (A_TILE_BYTES + H * 128);   // both CTAs' bytes of one stage
      const uint32_t tail_tx = 2u * (BN / 2) * 128;
      const uint32_t full0_base = mapa_shared(smem_u32(&full_bar[0]), 0);   // leader's full_bar[0]
      auto load_main = [&](int n_blk, int m_blk, int w_done) {     // w_done: k-blocks whose W part is in flight
        const int n0 = n_blk * BN;
        // convolution: first output pixel of this CTA's 128 rows -> (image, oh, ow)
        int cv_n = 0, cv_h = 0, cv_w = 0;
        if (p.conv) {
          const long long m0 = (long long)m_blk * 256 + (long long)rank * BM;
          const int hw = p.cv_Ho * p.cv_Wo;
          cv_n = (int)(m0 / hw);
          const int rem = (int)(m0 - (long long)cv_n * hw);
          cv_h = (rem / p.cv_Wo) * p.cv_str_h - p.cv_pad_h;
          cv_w = (rem % p.cv_Wo) * p.cv_str_w - p.cv_pad_w;
        }
        int tap = 0, cb = 0;
        int sg = 0, sk = 0;                      // K segment and k-block inside it (p.kseg != 0)
        const CUtensorMap* xm = &tmX;
        for (int kb = 0; kb < p.num_k_blk; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1, S2_PROD_EMPTY);
          uint8_t* a_dst = stage_base + (size_t)s * stg_bytes;
          uint8_t* b_dst = a_dst + A_TILE_BYTES;
          const uint32_t full0 = full0_base + s * 8;
          if (elect_one()) {
            if (xrole) {
              if (leader) mbar_arrive_expect_tx(&full_bar[s], main_tx);
              else mbar_arrive_cluster(full0);
              if (p.conv == 2)      // whole-row tiles of a stride-1 convolution: one tiled 4-D box per tap
                tma_load_4d_2sm(a_dst, &tmX, full0, cb * BK, cv_w + (tap % p.cv_S) * p.cv_dil_w,
                                cv_h + (tap / p.cv_S) * p.cv_dil_h, cv_n, pol_x);
              else if (p.conv)
                tma_load_im2col_2sm(a_dst, &tmX, full0, cb * BK, cv_w, cv_h, cv_n,
                                    (uint16_t)((tap % p.cv_S) * p.cv_dil_w),
                                    (uint16_t)((tap / p.cv_S) * p.cv_dil_h), pol_x);
              else if (p.kseg == 0)
                tma_load_2d_2sm(a_dst, &tmX, full0, kb * BK, m_blk * 256 + (int)rank * BM, pol_x);
              else     // X split along K into equal segments, one tensor each (fused sibling backward)
                tma_load_2d_2sm(a_dst, xm, full0, sk * BK, m_blk * 256 + (int)rank * BM, pol_x);
            } else if (GEGLU) {
              // GEGLU projection: B rows of the pair = [hidden j0 .. j0+BN/2) ; gate j0 .. j0+BN/2) ; 16 rank rows].
              // The leader's half (H = BN/2 + 8 rows) is the hidden rows plus the first 8 gate rows (two boxes), the
              // peer's the other BN/2 - 8 gate rows (= H - 16: the usual box) and the rank rows.
              const int j0 = n_blk * (BN / 2);
              if (leader) {
                if (box1) tma_load_2d_2sm(b_dst, &tmX1, full0, kb * BK, j0, pol_w);
                if (box2) tma_load_2d_2sm(b_dst + (size_t)(BN / 2) * 128, &tmX2, full0, kb * BK, p.geglu + j0, pol_w);
              } else {
                if (box1) tma_load_2d_2sm(b_dst, &tmWa, full0, kb * BK, p.geglu + j0 + 8, pol_w);
                if (box2) tma_load_2d_2sm(b_dst + (size_t)(H - RP) * 128, &tmS1, full0, kb * BK, 0, pol_w);
              }
            } else if (leader) {
              // B rows of the pair: [W n0 .. n0+BN) ; 16 rank rows], split in halves of H rows
              if (box1 && kb >= w_done) tma_load_2d_2sm(b_dst, &tmWh, full0, kb * BK, n0, pol_w);   // H rows
            } else {
              if (box1 && kb >= w_done) tma_load_2d_2sm(b_dst, &tmWa, full0, kb * BK, n0 + H, pol_w);   // H-16 rows
              if (box2) tma_load_2d_2sm(b_dst + (size_t)(H - RP) * 128, &tmS1, full0, kb * BK, 0, pol_w);
            }
          }
          __syncwarp();
          if (++cb == p.cv_cpb) { cb = 0; ++tap; }
          if (p.kseg && ++sk == p.kseg) { sk = 0; ++sg; xm = sg == 1 ? &tmX1 : &tmX2; }
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      };
      auto load_tail = [&](int n_blk) {
        if (warp != 2) mbar_wait(&empty_bar[s], ph ^ 1, S2_PROD_EMPTY);    // warp 2 has no part in a tail block: it
                                                                           // only keeps its stage counter in step
        uint8_t* b_dst = stage_base + (size_t)s * stg_bytes + A_TILE_BYTES;
        const uint32_t full0 = full0_base + s * 8;
        if (elect_one()) {
          if (xrole) {
            if (leader) mbar_arrive_expect_tx(&full_bar[s], tail_tx);
            else mbar_arrive_cluster(full0);
          } else if (box1) {
            const int s2_row = GEGLU ? n_blk * (BN / 2) + (int)rank * p.geglu      // leader: hidden rows, peer: gate rows
                                       : n_blk * BN + (int)rank * (BN / 2);
            tma_load_2d_2sm(b_dst, &tmS2, full0, 0, s2_row, pol_w);
          }
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1; }
      };
      int pit = 0;
      int n_blk = sc.t0 % p.num_n_blk, m_blk = sc.t0 / p.num_n_blk;
      for (int tile = sc.t0; tile < sc.t1;) {
        const bool lk = sc.look(tile);
        int n1 = n_blk + 1, m1 = m_blk;
        if (n1 == p.num_n_blk) { n1 = 0; ++m1; }
        if (xrole && lane == 0) TS2(1, pit, 0);
        load_main(n_blk, m_blk, tile == sc.t0 ? w_pre : 0);
        if (xrole && lane == 0) TS2(1, pit, 1);
        if (lk) load_main(n1, m1, 0);
        load_tail(n_blk);
        if (lk) load_tail(n1);
        if (xrole && lane == 0) TS2(1, pit, 2);
        if (lk) {
          n_blk = n1 + 1; m_blk = m1;
          if (n_blk == p.num_n_blk) { n_blk = 0; ++m_blk; }
        } else {
          n_blk = n1; m_blk = m1;
        }
        tile += lk ? 2 : 1;
        pit += lk ? 2 : 1;
      }
    }
  } else if (warp == 1) {
    // =============================== MMA issuer (leader only) =========================
    if (leader) {       // whole warp, warp-uniform; the elected lane issues MMAs and commits
      int s = 0;
      uint32_t ph = 0;
      const uint64_t t_desc = make_desc_sw128(smem_u32(t_tile));
      long long wsum = 0;   // DBG: cycles the issuer spent waiting for operands
      auto main_loop = [&](int st) {
        const uint32_t d_tmem = tmem_base + st * ACC_STRIDE;
        for (int kb = 0; kb < p.num_k_blk; ++kb) {
          long long w0 = 0;
          if (DBG) w0 = clock64();
          mbar_wait(&full_bar[s], ph, S2_MMA_FULL);
          if (DBG) wsum += clock64() - w0;
          tc_fence_after_sync();
          const uint32_t a_addr = smem_u32(stage_base + (size_t)s * stg_bytes);
          const uint64_t a_desc = make_desc_sw128(a_addr);
          const uint64_t b_desc = make_desc_sw128(a_addr + A_TILE_BYTES);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UK; ++k)
              umma_f16_2sm(d_tmem, a_desc + 2 * k, b_desc + 2 * k, p.idesc_main_t,
                           (kb | k) != 0 ? 1u : 0u);
            umma_commit_2sm(&empty_bar[s], 3);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      };
      auto tail = [&](int st) {
        mbar_wait(&full_bar[s], ph, S2_MMA_TAIL_FULL);
        tc_fence_after_sync();
        const uint64_t b_desc =
            make_desc_sw128(smem_u32(stage_base + (size_t)s * stg_bytes + A_TILE_BYTES));
        if (elect_one()) {
          for (int k = 0; k < p.tail_ksteps; ++k)
            umma_f16_2sm(tmem_base + st * ACC_STRIDE, t_desc + 2 * k, b_desc + 2 * k, p.idesc_tail, 1u);
          umma_commit_2sm(&empty_bar[s], 3);
          umma_commit_2sm(&accfull_bar[st], 3);
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1; }
      };
      int it = 0;
      int e = 0;
      int n_blk = sc.t0 % p.num_n_blk;          // tracked incrementally: no div/mod per tile
      int st = 0;                               // TMEM stage of tile `it` and its use parity
      uint32_t st_par = 0;
      const bool relay = p.cluster_scope == 2;
      const int nt = sc.t1 - sc.t0;
      // relay mode: a phase of accempty / tsm_full completes on 8 (128) LOCAL arrivals + this warp's arrive.expect_tx
      // + the 16 bytes the peer's relay warp sends when the peer's half of the hand-off is done.  The post for the
      // phase that tile `i` on stage s will complete goes out when the tile takes the stage — and only if a later tile
      // will wait for it (the relay sends under the same condition: nothing may be in flight towards a CTA that exits).
      auto post_acc = [&](int s_, int i_) {
        if (relay && i_ + NST < nt && elect_one()) mbar_arrive_expect_tx(&accempty_bar[s_], 16);
        __syncwarp();
      };
      auto wait_acc = [&](int s_, uint32_t par_) {
        if (p.cluster_scope == 1) mbar_wait_cluster(&accempty_bar[s_], par_, S2_MMA_ACC_EMPTY);   // acquire.cluster
        else mbar_wait(&accempty_bar[s_], par_, S2_MMA_ACC_EMPTY);
      };
      for (int tile = sc.t0; tile < sc.t1;) {
        const bool nm = tile == sc.t0 || n_blk == 0;
        const bool lk = nm && tile + 1 < sc.t1 && n_blk + 1 != p.num_n_blk;
        if (lane == 0) TS2(2, it, 0);
        wait_acc(st, st_par ^ 1);
        tc_fence_after_sync();
        post_acc(st, it);
        if (lane == 0) TS2(2, it, 1);
        if (nm && relay && elect_one()) mbar_arrive_expect_tx(tsm_full_bar, 16);    // this row block's T hand-off
        if (nm && e > 0 && elect_one()) umma_commit_2sm(st_free_bar, 3);
        __syncwarp();
        main_loop(st);
        if (lane == 0) TS2(2, it, 2);
        if (nm && elect_one()) umma_commit_2sm(tfull_bar, 3);
        __syncwarp();
        int st2 = st + 1;
        uint32_t par2 = st_par;
        if (st2 == NST) { st2 = 0; par2 ^= 1; }
        if (lk) {
          wait_acc(st2, par2 ^ 1);
          tc_fence_after_sync();
          post_acc(st2, it + 1);
          main_loop(st2);
          if (lane == 0) TS2(2, it, 3);
        }
        if (nm) {
          if (p.cluster_scope == 1) mbar_wait_cluster(tsm_full_bar, e & 1, S2_MMA_TSM);     // peer's [sT | 1] st.shared: acquire.cluster
          else mbar_wait(tsm_full_bar, e & 1, S2_MMA_TSM);
          tc_fence_after_sync();
          ++e;
        }
        if (lane == 0) TS2(2, it, 4);
        tail(st);
        if (lk) tail(st2);
        if (lane == 0) TS2(2, it, 5);
        if (DBG && lane == 0 && (int)blockIdx.x == p.dbg_cta && it < 32) { p.dbg[2 * 256 + it * 8 + 6] = wsum + 1; wsum = 0; }
        const int adv = lk ? 2 : 1;
        for (int a = 0; a < adv; ++a) {
          if (++st == NST) { st = 0; st_par ^= 1; }
          if (++n_blk == p.num_n_blk) n_blk = 0;
        }
        it += adv;
        tile += adv;
      }
    }
    else if (p.cluster_scope == 2) {
      // =============================== relay (peer CTA only) ==========================
      // The peer's T warps and epilogue warps arrive on PEER-LOCAL barriers (.cta scope: arrivers and observer share a
      // CTA).  This warp observes each completed phase and forwards it to the leader as an async-proxy transaction on
      // the leader's barrier — a 16-byte distributed-shared-memory bulk copy with complete_tx — so no generic-proxy
      // release has to cross the CTA pair and no GPU-scope fence sits in the MMA's path.
      const int nt = sc.t1 - sc.t0;
      const int n_rb = nt > 0 ? 1 + ((sc.t1 - 1) / sc.nnb - sc.t0 / sc.nnb) : 0;
      int ia = 0, sa = 0, et = 0;
      uint32_t pa = 0;
      const int na = nt > NST ? nt - NST : 0;      // accumulator releases a later tile waits for
      const long long t0c = clock64();
      // lane 0 observes (acquire) and forwards; the other lanes only follow its control flow
      auto seen = [&](uint64_t* bar, uint32_t parity) {
        int ok = 0;
        if (lane == 0) ok = mbar_try_wait(bar, parity) ? 1 : 0;
        return __shfl_sync(0xffffffffu, ok, 0) != 0;
      };
      while (ia < nt || et < n_rb) {
        if (ia < nt && seen(&accempty_loc[sa], pa)) {
          if (ia < na && lane == 0) {
            tc_fence_after_sync();
            tc_fence_before_sync();
            bulk_copy_to_peer(mapa_shared(smem_u32(relay_box + 16 * sa), 0), relay_box + 16 * sa, 16,
                              mapa_shared(smem_u32(&accempty_bar[sa]), 0));
          }
          __syncwarp();
          ++ia;
          if (++sa == NST) { sa = 0; pa ^= 1; }
        }
        if (et < n_rb && seen(tsm_loc, et & 1)) {
          if (lane == 0) {
            tc_fence_after_sync();
            tc_fence_before_sync();
            bulk_copy_to_peer(mapa_shared(smem_u32(relay_box + 16 * MAX_ACC), 0), relay_box + 16 * MAX_ACC, 16,
                              mapa_shared(smem_u32(tsm_full_bar), 0));
          }
          __syncwarp();
          ++et;
        }
        if (clock64() - t0c > B200LORA_WATCHDOG_CYCLES) {
          atomicCAS(&g_watchdog_code, 0u, 0x80000000u | (30u << 16) | (blockIdx.x & 0xffffu));
          __threadfence_system();
          __trap();
        }
      }
    }
  } else if (warp >= 4 && warp < 8) {
    // =============================== T step (both CTAs, own rows) =====================
    const int qd = warp & 3;
    const int row = qd * 32 + lane;
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
    const uint32_t trow = smem_u32(t_tile) + row * 128;
    const int sw = row & 7;
    const uint32_t tsm0 = mapa_shared(smem_u32(tsm_full_bar), 0);
    TIO* t_out = reinterpret_cast<TIO*>(p.T_out);
    int it = 0, e = 0;
    for (int tile = sc.t0; tile < sc.t1; ++tile, ++it) {
      if (!sc.new_m(tile)) continue;
      const int n_blk = tile % p.num_n_blk;
      const int m_blk = tile / p.num_n_blk;
      const int st = it % NST;
      const long long gm = (long long)m_blk * 256 + rank * BM + row;
      if (threadIdx.x == 128) TS2(3, it, 0);
      mbar_wait(tfull_bar, e & 1, S2_T_TFULL);
      tc_fence_after_sync();
      if (threadIdx.x == 128) TS2(3, it, 1);
      uint32_t tv[16];
      tmem_ld_x16(tmem_base + lane_addr + st * ACC_STRIDE + BN, tv);
      tmem_ld_wait();
      float tf[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        tf[j] = __uint_as_float(tv[j]) * p.scale;
        if (p.has_bias && j == p.r) tf[j] = 1.0f;
      }
      if (e > 0) mbar_wait(st_free_bar, (e - 1) & 1, S2_T_FREE);
      st_shared_v4(trow + ((0 ^ sw) << 4), pack2b<TIO>(tf[0], tf[1]), pack2b<TIO>(tf[2], tf[3]),
                   pack2b<TIO>(tf[4], tf[5]), pack2b<TIO>(tf[6], tf[7]));
      st_shared_v4(trow + ((1 ^ sw) << 4), pack2b<TIO>(tf[8], tf[9]), pack2b<TIO>(tf[10], tf[11]),
                   pack2b<TIO>(tf[12], tf[13]), pack2b<TIO>(tf[14], tf[15]));
      if (p.tail_ksteps > 1) {
        st_shared_v4(trow + ((2 ^ sw) << 4), pack2b<TIO>(1.0f, 0.f), 0u, 0u, 0u);
        st_shared_v4(trow + ((3 ^ sw) << 4), 0u, 0u, 0u, 0u);
      }
      fence_proxy_async_smem();
      tc_fence_before_sync();
      // the leader's tcgen05.mma reads THIS CTA's t_tile: the peer publishes with release.cluster (PTX memory
      // model: a .cta-scope release does not order the st.shared above for an observer in another CTA); the
      // leader's own T warps are in the observer's CTA, where the default .cta scope is the right one
      if (leader) mbar_arrive(tsm_full_bar);
      else if (p.cluster_scope == 2) mbar_arrive(tsm_loc);                 // relay mode: stays in this CTA
      else if (p.cluster_scope == 1) mbar_arrive_cluster_release(tsm0);
      else mbar_arrive_cluster(tsm0);
      // T_out (saved for the backward) goes to global memory AFTER the hand-off: the cluster-scope release above is
      // a MEMBAR.GPU, which would otherwise sit on the MMA's critical path until these stores are acknowledged
      // (measured +2.5 us per launch at 128 row blocks when the stores came first)
      if (t_out != nullptr && n_blk == 0 && gm < p.M) {
        TIO* dst = t_out + gm * p.r;
        if (p.r == 4) {
          uint2 q;
          q.x = pack2b<TIO>(__uint_as_float(tv[0]), __uint_as_float(tv[1]));
          q.y = pack2b<TIO>(__uint_as_float(tv[2]), __uint_as_float(tv[3]));
          *reinterpret_cast<uint2*>(dst) = q;
        } else if (p.r == 8 || p.r == 16) {
          uint4 q;
          q.x = pack2b<TIO>(__uint_as_float(tv[0]), __uint_as_float(tv[1]));
          q.y = pack2b<TIO>(__uint_as_float(tv[2]), __uint_as_float(tv[3]));
          q.z = pack2b<TIO>(__uint_as_float(tv[4]), __uint_as_float(tv[5]));
          q.w = pack2b<TIO>(__uint_as_float(tv[6]), __uint_as_float(tv[7]));
          *reinterpret_cast<uint4*>(dst) = q;
          if (p.r == 16) {
            q.x = pack2b<TIO>(__uint_as_float(tv[8]), __uint_as_float(tv[9]));
            q.y = pack2b<TIO>(__uint_as_float(tv[10]), __uint_as_float(tv[11]));
            q.z = pack2b<TIO>(__uint_as_float(tv[12]), __uint_as_float(tv[13]));
            q.w = pack2b<TIO>(__uint_as_float(tv[14]), __uint_as_float(tv[15]));
            *reinterpret_cast<uint4*>(dst + 8) = q;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j)
            if (j < p.r) dst[j] = TIO(__uint_as_float(tv[j]));
        }
      }
      if (threadIdx.x == 128) TS2(3, it, 2);
      ++e;
    }
  } else if (warp >= 8) {
    // =============================== epilogue (both CTAs, own rows) ===================
    const int ew = warp - 8;
    const int qd = warp & 3;
    const int half = ew >> 2;
    const uint32_t lane_addr = static_cast<uint32_t>(qd * 32) << 16;
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
      if (lane == 0 && qd == 0) TS2(4 + half, it, 0);
      mbar_wait(&accfull_bar[st], (it / NST) & 1, S2_EPI_ACC);
      tc_fence_after_sync();
      if (lane == 0 && qd == 0) TS2(4 + half, it, 1);
      uint32_t v[2][32];
      auto live = [&](int c) { return c < n_units && n_blk * BN + c * EPI_COLS < p.Nout; };
      if (live(half)) tmem_ld_x32(acc_addr + half * EPI_COLS, v[0]);
      bool released = false;
      // this warp's TMEM reads of the stage are complete once the LAST chunk's tcgen05.wait::ld has returned: the
      // arrival goes out right there, before that chunk is converted and stored, so the (cluster-scope) release
      // overlaps the rest of the epilogue instead of extending the MMA's wait for the accumulator stage
      auto release_stage = [&]() {
        tc_fence_before_sync();
        __syncwarp();
        if (lane == 0) {     // peer: release.cluster (its TMEM reads are ordered before the leader's next MMA)
          if (leader) mbar_arrive(&accempty_bar[st]);
          else if (p.cluster_scope == 2) mbar_arrive(&accempty_loc[st]);   // relay mode: stays in this CTA
          else if (p.cluster_scope == 1) mbar_arrive_cluster_release(mapa_shared(smem_u32(&accempty_bar[st]), 0));
          else mbar_arrive_cluster(mapa_shared(smem_u32(&accempty_bar[st]), 0));
        }
        released = true;
      };
      if (GEGLU) {
        // ---- GEGLU projection: accumulator columns [0, BN/2) = hidden, [BN/2, BN) = gate of the same BN/2 outputs.
        // This warp takes the 32-column chunk pairs pi = half, half + 2, ...: out = hidden * gelu(gate), rounded as the
        // reference composition rounds (projection output, gelu, product: each a tensor of the io dtype), stored
        // to tmOut [M, geglu]; the pre-activation chunks go to Y [M, 2*geglu] when the backward needs them.
        const int np = BN / (2 * EPI_COLS);
        const int j0 = n_blk * (BN / 2);
        const int row0 = m_blk * 256 + (int)rank * BM + qd * 32;
        int last_pi = -1;
        for (int pi = half; pi < np; pi += 2)
          if (j0 + pi * EPI_COLS < p.geglu) last_pi = pi;
        auto emit = [&](const uint32_t (&w)[32], const CUtensorMap* tm, int col) {
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          const uint32_t dst = stg_row + buf * (32 * EPI_COLS * 2);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            st_shared_v4(dst + ((q ^ wr_sw) << 4),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 0]), __uint_as_float(w[q * 8 + 1])),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 2]), __uint_as_float(w[q * 8 + 3])),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 4]), __uint_as_float(w[q * 8 + 5])),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 6]), __uint_as_float(w[q * 8 + 7])));
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(tm, my_staging + buf * (32 * EPI_COLS * 2), col, row0);
            tma_store_commit();
          }
          buf ^= 1;
        };
        for (int pi = half; pi < np; pi += 2) {
          if (j0 + pi * EPI_COLS >= p.geglu) continue;
          tmem_ld_x32(acc_addr + pi * EPI_COLS, v[0]);
          tmem_ld_x32(acc_addr + (np + pi) * EPI_COLS, v[1]);
          tmem_ld_wait_dep(v[0]);
          tmem_ld_wait_dep(v[1]);
          if (pi == last_pi) release_stage();
          if (p.geglu_store_y) {
            emit(v[0], &tmY, j0 + pi * EPI_COLS);
            emit(v[1], &tmY, p.geglu + j0 + pi * EPI_COLS);
          }
#pragma unroll
          for (int i = 0; i < 32; ++i) {
            const float h = to_f32(from_f32<TIO>(__uint_as_float(v[0][i])));
            const float g = to_f32(from_f32<TIO>(__uint_as_float(v[1][i])));
            const float ge = to_f32(from_f32<TIO>(0.5f * g * (1.0f + erff(g * 0.70710678118654752f))));
            v[0][i] = __float_as_uint(h * ge);
          }
          emit(v[0], &tmOut, j0 + pi * EPI_COLS);
        }
      } else
#pragma unroll
      for (int ci = 0; ci < MAX_CI; ++ci) {
        const int c = 2 * ci + half;
        if (live(c)) {
          tmem_ld_wait_dep(v[ci & 1]);
          const bool more = ci + 1 < MAX_CI && live(c + 2);
          if (more) tmem_ld_x32(acc_addr + (c + 2) * EPI_COLS, v[(ci + 1) & 1]);
          else release_stage();
          const uint32_t* w = v[ci & 1];
          if (F32OUT) {
            // fp32 output: the chunk's 32 accumulator columns leave as two [32 rows x 16 columns] boxes (64-byte rows,
            // the staging buffers' size and swizzle unchanged), unrounded
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
              if (lane == 0) tma_store_wait_read<1>();
              __syncwarp();
              const uint32_t dst = stg_row + buf * (32 * EPI_COLS * 2);
#pragma unroll
              for (int q = 0; q < 4; ++q)
                st_shared_v4(dst + ((q ^ wr_sw) << 4), w[hh * 16 + q * 4 + 0], w[hh * 16 + q * 4 + 1],
                             w[hh * 16 + q * 4 + 2], w[hh * 16 + q * 4 + 3]);
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_2d(&tmY, my_staging + buf * (32 * EPI_COLS * 2), n_blk * BN + c * EPI_COLS + hh * 16,
                             m_blk * 256 + (int)rank * BM + qd * 32);
                tma_store_commit();
              }
              buf ^= 1;
            }
          } else {
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          const uint32_t dst = stg_row + buf * (32 * EPI_COLS * 2);
#pragma unroll
          for (int q = 0; q < 4; ++q)
            st_shared_v4(dst + ((q ^ wr_sw) << 4),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 0]), __uint_as_float(w[q * 8 + 1])),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 2]), __uint_as_float(w[q * 8 + 3])),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 4]), __uint_as_float(w[q * 8 + 5])),
                         pack2b<TIO>(__uint_as_float(w[q * 8 + 6]), __uint_as_float(w[q * 8 + 7])));
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&tmY, my_staging + buf * (32 * EPI_COLS * 2), n_blk * BN + c * EPI_COLS,
                         m_blk * 256 + (int)rank * BM + qd * 32);
            tma_store_commit();
          }
          buf ^= 1;
          }
        }
      }
      if (!released) release_stage();      // a warp with no live chunk in this tile (ragged N) still owes its arrival
      if (lane == 0 && qd == 0) TS2(4 + half, it, 2);
    }
    if (lane == 0) tma_store_wait_all<0>();
    if (lane == 0 && qd == 0) TS2(4 + half, 31, 7);
  }

  // =============================== teardown ===========================================
  tc_fence_before_sync();
  cluster_sync_all();      // no CTA leaves while its pair may still touch its smem / barriers
  if (warp == 2) {
    tc_fence_after_sync();
    tmem_dealloc_2sm(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
int tc2_smem_bytes(int BN, int* stages_out) {
  const int budget = 232448 - 1024;
  int stages = (budget - tc2::fixed_bytes()) / tc2::stage_bytes(BN);
  if (stages > tc2::MAX_STAGES) stages = tc2::MAX_STAGES;
  if (stages_out) *stages_out = stages;
  return stages * tc2::stage_bytes(BN) + tc2::fixed_bytes() + 1024;
}

// B200LORA_PDL=0 turns programmatic dependent launch off (A/B measurements)
bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200LORA_PDL");
    v = (e && e[0] == '0') ? 0 : 1;
  }
  return v == 1;
}

template <typename TIO, bool DBG, int MODE>
static cudaError_t launch_tc2_impl(const CUtensorMap& tmX, const CUtensorMap& tmX1, const CUtensorMap& tmX2,
                                   const CUtensorMap& tmWa,
                                   const CUtensorMap& tmWb, const CUtensorMap& tmS1,
                                   const CUtensorMap& tmS2, const CUtensorMap& tmY, const CUtensorMap& tmOut,
                                   const TcParams& p, int smem_bytes, cudaStream_t stream) {
  auto kern = lora_gemm_tc2_kernel<TIO, DBG, MODE>;
  // The opt-in smem cap is a process-wide attribute of the function: raise it ONCE to the device
  // maximum (autograd runs the backward on another thread — a per-thread cache of "what I last
  // set" let one thread lower the cap under the other).  Co-resident pair count is cached per smem.
  static std::mutex mu;
  static bool cap_raised = false;
  static std::map<int, int> clusters_for_smem;
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.blockDim = dim3(tc2::NUM_THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  int max_clusters = 0;
  {
    std::lock_guard<std::mutex> g(mu);
    if (!cap_raised) {
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
      if (e != cudaSuccess) { set_error("tc2: cudaFuncSetAttribute: %s", cudaGetErrorString(e)); return e; }
      cap_raised = true;
    }
    auto it = clusters_for_smem.find(smem_bytes);
    if (it == clusters_for_smem.end()) {
      cfg.gridDim = dim3(2, 1, 1);
      int n = 0;
      cudaError_t e = cudaOccupancyMaxActiveClusters(&n, kern, &cfg);
      if (e != cudaSuccess) { set_error("tc2: cudaOccupancyMaxActiveClusters(smem=%d): %s", smem_bytes, cudaGetErrorString(e)); return e; }
      if (n < 1) return cudaErrorInvalidConfiguration;
      it = clusters_for_smem.emplace(smem_bytes, n).first;
    }
    max_clusters = it->second;
  }
  const int num_tiles = p.num_m_blk * p.num_n_blk;
  const int ncl = num_tiles < max_clusters ? num_tiles : max_clusters;
  cfg.gridDim = dim3(2 * ncl, 1, 1);
  cudaError_t e = cudaLaunchKernelEx(&cfg, kern, tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p);
  if (e != cudaSuccess)
    set_error("tc2: cudaLaunchKernelEx(grid=%d, smem=%d, BN=%d, stages=%d): %s", 2 * ncl, smem_bytes, p.BN,
              p.stages, cudaGetErrorString(e));
  return e;
}

template <typename TIO>
cudaError_t launch_tc2(const CUtensorMap& tmX, const CUtensorMap& tmX1, const CUtensorMap& tmX2,
                       const CUtensorMap& tmWa, const CUtensorMap& tmWb, const CUtensorMap& tmS1,
                       const CUtensorMap& tmS2, const CUtensorMap& tmY, const CUtensorMap& tmOut, const TcParams& p,
                       int smem_bytes, cudaStream_t stream) {
  if (p.geglu)
    return launch_tc2_impl<TIO, false, 1>(tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p, smem_bytes, stream);
  if (p.out_f32)
    return launch_tc2_impl<TIO, false, 2>(tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p, smem_bytes, stream);
  if (p.dbg != nullptr)      // (the clock64 timeline exists for the plain kernel only)
    return launch_tc2_impl<TIO, true, 0>(tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p, smem_bytes, stream);
  return launch_tc2_impl<TIO, false, 0>(tmX, tmX1, tmX2, tmWa, tmWb, tmS1, tmS2, tmY, tmOut, p, smem_bytes, stream);
}

template cudaError_t launch_tc2<__nv_bfloat16>(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                               const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                               const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                               const TcParams&, int, cudaStream_t);
template cudaError_t launch_tc2<__half>(const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                        const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                        const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const TcParams&,
                                        int, cudaStream_t);

}  // namespace b200
