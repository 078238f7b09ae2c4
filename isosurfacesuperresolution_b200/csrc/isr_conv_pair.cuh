// isr_conv_pair.cuh -- the 64-wide conv3x3 layers as a CTA-PAIR implicit GEMM (tcgen05.mma.cta_group::2, M = 256).
//
// Why: a tcgen05.mma with both operands in shared memory costs ~54 cycles per instruction for every N <= 64 when issued
// for one CTA (M = 128), but 43 cycles for a CTA pair (M = 256): 21.5 cycles per 128-pixel tile instead of 54, and 64
// instead of 2 x 67.7 at N = 128 (profiles/r01_probe_mma_cycles.json, r01_probe_mma2.json).  The pair also halves the
// weight footprint: each CTA keeps only its half of the B operand (N/2 rows per tap).
//
// Structure (both CTAs run the same code; rank 0 = leader):
//   * each CTA owns one 8x16 tile per iteration (tiles 2j and 2j+1), loads ITS halo into ITS shared memory, and drains
//     ITS half of the accumulator (TMEM lanes 0..127 of its own SM) with the same epilogue as the single-CTA kernel;
//   * only the leader's MMA warp issues instructions; descriptors are shared-memory OFFSETS valid in both CTAs, so the
//     two CTAs use identical layouts and advance their halo stages in lock step;
//   * barriers: full[s] and the weight barrier live in the LEADER (2 arrivals + both CTAs' TMA bytes, the peer's TMA
//     signals it through .cta_group::2); empty[s] and tmem_full[a] are per CTA and are arrived by a multicast
//     tcgen05.commit; tmem_empty[a] lives in the leader and collects 4 epilogue warps x 2 CTAs.
//   * STACK (split precision): pass 0 = A_hi x [W_hi ; W_lo] as one N=128 operand (leader holds the 64 W_hi rows of a
//     tap, the peer the 64 W_lo rows), pass 1 = A_lo x W_hi with N=64 (each CTA holds 32 rows of W_hi).
#pragma once
#include "isr_conv.cuh"

namespace isr {

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx_cluster(uint32_t cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_addr), "r"(bytes)
               : "memory");
}
// TMA loads whose completion bytes are credited to a barrier that may live in the pair's other CTA
__device__ __forceinline__ void tma2_load_2d(uint32_t dst, const void *tmap, uint32_t bar_cluster, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma2_load_4d(uint32_t dst, const void *tmap, uint32_t bar_cluster, int c0, int c1, int c2,
                                             int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, "
      "%6}], [%2];" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(tmap)), "r"(bar_cluster), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc2(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// four M256 x N x K16 MMAs covering K = 64 of one tap (see umma_f16_k64)
__device__ __forceinline__ void umma2_f16_k64(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                              uint32_t accumulate_first) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\t.reg .b64 a1, a2, a3, b1, b2, b3;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "setp.eq.b32 q, %3, %3;\n\t"
      "add.s64 a1, %1, 2;\n\t"
      "add.s64 b1, %2, 2;\n\t"
      "add.s64 a2, %1, 4;\n\t"
      "add.s64 b2, %2, 4;\n\t"
      "add.s64 a3, %1, 6;\n\t"
      "add.s64 b3, %2, 6;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a1, b1, %3, q;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a2, b2, %3, q;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], a3, b3, %3, q;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate_first)
      : "memory");
}
// two M256 x N x K32 8-bit MMAs (kind::f8f6f4) covering K = 64 of one tap; always accumulating
__device__ __forceinline__ void umma2_f8_k64(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc) {
  asm volatile(
      "{\n\t.reg .pred q;\n\t.reg .b64 a1, b1;\n\t"
      "setp.eq.b32 q, %3, %3;\n\t"
      "add.s64 a1, %1, 2;\n\t"
      "add.s64 b1, %2, 2;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, q;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], a1, b1, %3, q;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc)
      : "memory");
}
// arrive on the barrier at this offset in BOTH CTAs when all previously issued MMAs of the pair have completed
__device__ __forceinline__ void umma2_commit_multicast(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"(static_cast<uint16_t>(3))
      : "memory");
}

// (tx, ty, b) of tile t, advanced by a constant step without divisions (one division set per role and launch)
struct PairTileCursor {
  int tx, ty, b, dx, dy, db, tiles_x, tiles_y, rev;
  // rev: the cursor starts at tile t and walks DOWN by `step` (ConvParams::reverse)
  __device__ __forceinline__ void init(int t, int step, int txs, int tys, int reverse = 0) {
    tiles_x = txs; tiles_y = tys; rev = reverse;
    tx = t % txs; ty = (t / txs) % tys; b = t / (txs * tys);
    dx = step % txs; dy = (step / txs) % tys; db = step / (txs * tys);
  }
  __device__ __forceinline__ void advance() {
    if (rev) {
      tx -= dx; ty -= dy; b -= db;
      if (tx < 0) { tx += tiles_x; --ty; }
      if (ty < 0) { ty += tiles_y; --b; }
    } else {
      tx += dx; ty += dy; b += db;
      if (tx >= tiles_x) { tx -= tiles_x; ++ty; }
      if (ty >= tiles_y) { ty -= tiles_y; ++b; }
    }
  }
};

// Two epilogue groups of four warps and FOUR accumulator buffers in TMEM: tile i uses buffer i % 4 and is drained by
// group i % 2 (so a group only ever waits on its own two buffers and observes every phase of their barriers).  With two
// buffers the epilogue of tile i had to finish within the MMA time of tile i+1 (~1500 cycles) or the issuer stalled;
// with four it has three tiles' worth of slack while each group still has two tile times per tile.  (Three groups of
// warps instead -- 448 threads, 128 registers -- spill in the epilogue and run 30 % slower: measured.)
constexpr int kPairGroups = 2;
constexpr int kPairBufs = 4;
constexpr int kPairIssuers = 2;   // MMA-issuing warps (leader CTA), see the issuer section of the kernel
// Warps per epilogue group: 4 (a warp owns TMEM lanes 32q..32q+31 and all 64 columns: 32-byte stores) or 8 (EPW = 8: warps q
// and q + 4 of a group own the same lanes and split the accumulator columns, epilogue_store16_half in isr_conv.cuh).  The
// 8-warp form is used for the split-precision convs WITH a residual addend, whose per-tile chain in one warp (addend loads ->
// accumulator -> ~1000 dependent instructions -> stores, ~9000 cycles) exceeded the two tiles of MMAs a group has per tile
// (6200 cycles): 151-164 -> 134-141 us.  The store-heavy 4x layers lose with it (16-byte instead of 32-byte stores: fused
// upsample + conv 743 -> 942 us, 64->64 at 4x 1040 -> 1110 us) and keep four warps.
constexpr int pair_threads(int epw) { return (epw * kPairGroups + 1 + kPairIssuers) * 32; }
constexpr int kPairThreads = pair_threads(4);

// N = 64 output channels.  Weight rows of one CTA per tap: STACK: 64 (X region: W_hi for the leader, W_lo for the peer)
// + 32 (Y region: its half of W_hi); otherwise 32 per weight part.
// N = 16 with EPI_FINAL: the network's output conv (Cout <= 16) with its fused skip / state epilogue; an N = 16 pair-MMA
// costs 39 cycles, and the epilogue is a plain TMEM row load (no shared-memory gather as in isr_conv_final.cuh).
// PH = 4: fused bilinear-2x-upsample + conv (isr_up2.cu): every low-res tile is convolved four times, once per output
// phase (py, px), with that phase's composite weights; the four results are four "virtual tiles" v = 4*i + phase that
// share the tile's halo stage, each with its own accumulator buffer (8 buffers) and its own strided output pixels.
// PH = 2 (the default form of that layer): the same, with the two x-phases of an output row stacked along N -- the A
// window of a tap is the same for all four phases, so [W'(py,0) ; W'(py,1)] is ONE N = 128 operand (leader holds the
// px = 0 rows, the peer the px = 1 rows, as the split-precision stack does): 64 cycles per k16-step for two phases, the
// tensor floor, instead of 2 x 43 (an N = 64 pair-MMA is bound by its shared-memory operand reads).  Two virtual tiles
// v = 2*i + py per low-res tile, four 128-column accumulator buffers; the epilogue drains columns 0..63 to output pixel
// (2y+py, 2x) and 64..127 to (2y+py, 2x+1).
// ADD = false: an instance for layers without a residual addend (its prefetch registers are not allocated).
template <int STACK, int N = 64, int EPI = EPI_STORE16, int PH = 1, bool ADD = true, int EPW = 4>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(pair_threads(EPW), 1)
conv3x3_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                    const __grid_constant__ CUtensorMap tmA8, const __grid_constant__ CUtensorMap tmW8,
                    const ConvParams p) {
  // tmA8 / tmW8: byte-typed maps of the 8-bit low activation plane and the 8-bit weight plane (SWIZZLE_64B), used only
  // when p.a_lo8 is set (STACK): pass 1 is then A_lo8 x W8 as kind::f8f6f4 -- two K = 32 MMAs per tap instead of four
  // K = 16 ones, half the shared-memory operand bytes of the fp16 correction pass (profiles/r02_probe_mma4_fp8_lo.json).
  static_assert(N == 64 || N == 16, "N = 64 (hidden layers) or 16 (output conv)");
  static_assert(EPW == 4 || (EPW == 8 && N == 64 && EPI == EPI_STORE16 && PH == 1), "EPW = 8: plain 64-wide layers only");
  static_assert(PH == 1 || ((PH == 4 || PH == 2) && STACK == 0 && N == 64 && EPI == EPI_STORE16), "PH = 4 / 2: single-pass 64-wide layers only");
  constexpr bool kUp = PH != 1;     // fused upsample + conv
  constexpr bool kWide = PH == 2;   // ... with two phases per MMA (N = 128)
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();

  constexpr uint32_t kHalfTap = (N / 2) * 128;   // 32 weight rows = 4 KB
  constexpr uint32_t kHalfTap8 = (N / 2) * 64;   // 32 rows of 64 bytes = 2 KB
  const bool lo8 = STACK && p.a_lo8 != 0;
  const uint32_t w_bytes = STACK ? (lo8 ? 9 * 2 * kHalfTap + 9 * kHalfTap8 : 9 * 3 * kHalfTap)
                                 : (kUp ? 36 : p.n_w_parts * 9) * kHalfTap;   // per CTA
  const uint32_t s_w = smem_base;
  const uint32_t s_wy = s_w + 9 * 2 * kHalfTap;   // STACK: Y region after the X region
  const uint32_t s_stage0 = s_w + ((w_bytes + 1023u) & ~1023u);
  const uint32_t s_bar = s_stage0 + p.num_stages * p.stage_bytes;
  constexpr int G = kPairGroups, NB = (PH == 4 ? 8 : kPairBufs), kTmaWarp = EPW * G, kMmaWarp = EPW * G + 1;
  const uint32_t b_full = s_bar, b_empty = s_bar + 64, b_w = s_bar + 128, b_tfull = s_bar + 136,
                 b_tempty = b_tfull + 8 * NB, s_tptr = b_tempty + 8 * NB;
  static_assert(136 + 16 * NB + 4 <= (PH == 4 ? 512 : 256), "barrier block overflows its reservation (256 B; 512 B for PH = 4)");
  __shared__ float s_bias[64];
  if (threadIdx.x < 64) s_bias[threadIdx.x] = (p.bias != nullptr && threadIdx.x < N) ? p.bias[threadIdx.x] : 0.f;

  if (warp == kTmaWarp && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(b_full + 8 * s, 2);    // used in the leader: one arrive.expect_tx per CTA
      mbar_init(b_empty + 8 * s, kUp ? 2 : 1);   // per CTA: multicast commit (fused upsample: one per issuing warp)
    }
    mbar_init(b_w, 2);
    for (int a = 0; a < NB; ++a) {
      mbar_init(b_tfull + 8 * a, 1);   // per CTA: multicast commit
      mbar_init(b_tempty + 8 * a, 2 * EPW);  // leader: the epilogue warps of one group x 2 CTAs
    }
    fence_barrier_init();
  }
  constexpr uint32_t kAccCols = (STACK || kWide) ? 2 * N : N;
  constexpr uint32_t kTmemCols = (NB * kAccCols <= 64) ? 64 : (NB * kAccCols <= 128) ? 128 : ((NB * kAccCols <= 256) ? 256 : 512);
  static_assert(NB * kAccCols <= 512 && NB % G == 0, "accumulator buffers exceed TMEM");
  if (warp == kMmaWarp) tmem_alloc2(s_tptr, kTmemCols);
  tc_fence_before();
  cluster_sync_all();   // barriers of both CTAs initialised, TMEM allocated
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(s_tptr));

  // Programmatic dependent launch: everything above (barriers, TMEM, tensor-map prefetch) and the weight load below do
  // not depend on the previous kernel; the warps that touch activations call pdl_wait() first.
  pdl_launch_dependents();
  const int num_pairs = static_cast<int>(gridDim.x) >> 1;
  const int pair_id = static_cast<int>(blockIdx.x) >> 1;
  const int pair_tiles = (p.num_tiles + 1) >> 1;
  const int my_iters = (pair_tiles - pair_id + num_pairs - 1) / num_pairs;

  if (warp == kTmaWarp) {
    // ===================== TMA producer (one per CTA) =====================
    const uint32_t lead_w = mapa_shared(b_w, 0);
    if (elect_one()) {
      mbar_arrive_expect_tx_cluster(lead_w, w_bytes);
      if constexpr (STACK) {
        for (int tap = 0; tap < 9; ++tap) {
          const int row_x = p.w_row0 + (static_cast<int>(rank) * 9 + tap) * p.w_row_stride;   // part = rank
          tma2_load_2d(s_w + tap * 2 * kHalfTap, &tmW, lead_w, 0, row_x);
          tma2_load_2d(s_w + tap * 2 * kHalfTap + kHalfTap, &tmW, lead_w, 0, row_x + N / 2);
          if (lo8)
            tma2_load_2d(s_wy + tap * kHalfTap8, &tmW8, lead_w, 0, tap * N + static_cast<int>(rank) * (N / 2));
          else
            tma2_load_2d(s_wy + tap * kHalfTap, &tmW, lead_w, 0, p.w_row0 + tap * p.w_row_stride + static_cast<int>(rank) * (N / 2));
        }
      } else if constexpr (kWide) {
        // this CTA keeps all 64 rows of the phases with px = rank: block (py, tap) = 8 KB at (py * 9 + tap) * 8 KB
        for (int j = 0; j < 36; ++j) {
          const int phase = j / 9, tap = j - 9 * phase;
          if ((phase & 1) != static_cast<int>(rank)) continue;
          const uint32_t dst = s_w + static_cast<uint32_t>(((phase >> 1) * 9 + tap) * 2) * kHalfTap;
          const int row = p.w_row0 + j * p.w_row_stride;
          tma2_load_2d(dst, &tmW, lead_w, 0, row);
          tma2_load_2d(dst + kHalfTap, &tmW, lead_w, 0, row + N / 2);
        }
      } else {
        const int nblk = PH == 4 ? 36 : p.n_w_parts * 9;   // (part | phase, tap) blocks of N rows
        for (int j = 0; j < nblk; ++j)
          tma2_load_2d(s_w + j * kHalfTap, &tmW, lead_w, 0, p.w_row0 + j * p.w_row_stride + static_cast<int>(rank) * (N / 2));
      }
    }
    __syncwarp();
    pdl_wait();   // the activations are the previous kernel's output
    // Per-tile bookkeeping without integer divisions (a division by a run-time value is ~30 dependent instructions in a
    // single-thread role whose per-tile chain is on the critical path: the output conv lost 7 % to three of them per tile,
    // tools/final_experiments2.py): the tile coordinates and the stage ring are counters advanced by constant steps.
    uint32_t it = 0, s = 0, ph = 0;
    PairTileCursor cur;
    cur.init(2 * (p.reverse ? pair_tiles - 1 - pair_id : pair_id) + static_cast<int>(rank), 2 * num_pairs, p.tiles_x,
             p.tiles_y, p.reverse);
    for (int i = 0; i < my_iters; ++i, cur.advance()) {
      bool valid = cur.b < p.B;
      int tx = cur.tx, ty = cur.ty;
      int b = valid ? cur.b : p.B;   // invalid tile: fully out of bounds = zero fill
      if (p.dbg & 1024) {   // bench experiment: the divisions of the first form
        const int t = 2 * (pair_id + i * num_pairs) + static_cast<int>(rank);
        valid = t < p.num_tiles; tx = t % p.tiles_x; ty = (t / p.tiles_x) % p.tiles_y;
        b = valid ? t / (p.tiles_x * p.tiles_y) : p.B;
      }
      for (int ap = 0; ap < p.n_act_parts; ++ap, ++it) {
        if (it != 0 && ++s == static_cast<uint32_t>(p.num_stages)) { s = 0; ph ^= 1; }
        if (p.dbg & 1024) { s = it % p.num_stages; ph = (it / p.num_stages) & 1; }
        mbar_wait_parked(b_empty + 8 * s, ph ^ 1);
        if (elect_one()) {
          const uint32_t lead_full = mapa_shared(b_full + 8 * s, 0);
          if ((p.dbg & 16) && it >= static_cast<uint32_t>(p.num_stages)) {   // bench experiment: stages are filled once
            mbar_arrive_cluster(lead_full);
          } else if (lo8 && ap == 1) {   // the 8-bit plane: 64-byte rows
            mbar_arrive_expect_tx_cluster(lead_full, kPitch * kHaloH * 64);
            tma2_load_4d(s_stage0 + s * p.stage_bytes, &tmA8, lead_full, 0, tx * kTileW - 1, ty * kTileH - 1, b);
          } else {
            mbar_arrive_expect_tx_cluster(lead_full, kPitch * kHaloH * 128);
            tma2_load_4d(s_stage0 + s * p.stage_bytes, &tmA, lead_full, ap * 64, tx * kTileW - 1, ty * kTileH - 1, b);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp >= kMmaWarp) {
    // ===================== MMA issuers (leader CTA only): TWO warps, warp m issues tiles m, m+2, ... =================
    // The tcgen05.mma queue is shallow: whatever the issuing thread does between the last MMA of a tile and the first of
    // the next (barrier waits, commits, index arithmetic) is a tensor-pipe bubble -- measured 49.8 cycles per MMA in this
    // loop against 43.0 for back-to-back issue (tools/probe_mma3.cu), and moving the waits into the middle of a tile made
    // it worse (58.8).  With two issuing warps one of them always has MMAs to feed while the other does its per-tile
    // bookkeeping; tiles alternate, accumulators and halo stages are barrier-protected per tile, and a tcgen05.commit
    // tracks the MMAs of the issuing thread, i.e. of that warp's own tile.
    if (rank == 0) {
      const int mw = warp - kMmaWarp;
      const uint32_t idesc = make_idesc_f16(256, N, p.is_bf16);
      const uint32_t idesc2 = make_idesc_f16(256, 2 * N, p.is_bf16);
      const uint64_t b_desc_base = make_smem_desc_sw128(s_w, 1024, 0);
      const uint64_t by_desc_base = make_smem_desc_sw128(s_wy, 1024, 0);
      const uint64_t by8_desc_base = make_smem_desc_sw64(s_wy, 512);
      const uint32_t idesc8 = make_idesc_f8(256, N, 0, 1);   // A e4m3, B e5m2
      mbar_wait(b_w, 0);
      const long long c_start = clock64();
      for (int v = mw; v < my_iters * PH; v += kPairIssuers) {   // virtual tile v = PH * i + phase
        const int i = v / PH, phase = v % PH;
        const uint32_t acc = v % NB;
        const uint32_t acc_ph = (v / NB) & 1;
        if (!(p.dbg & 8)) mbar_wait(b_tempty + 8 * acc, acc_ph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * kAccCols;
        uint32_t first = 1;
        for (int ap = 0; ap < p.n_act_parts; ++ap) {
          const uint32_t it = static_cast<uint32_t>(i) * p.n_act_parts + ap;
          const uint32_t s = it % p.num_stages;
          const uint32_t ph = (it / p.num_stages) & 1;
          mbar_wait(b_full + 8 * s, ph);
          tc_fence_after();
          const uint32_t a_base = s_stage0 + s * p.stage_bytes;
          const uint64_t a_desc0 = make_smem_desc_sw128(a_base, kPitch * 128, 0);
          if (elect_one()) {
            if (p.dbg & 2) {
            } else if (STACK && lo8 && ap == 1) {
              // correction pass on the 8-bit planes: the same nine tap windows, 64-byte rows (SWIZZLE_64B)
              const uint64_t a8_desc0 = make_smem_desc_sw64(a_base, kPitch * 64);
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                const uint64_t adesc = a8_desc0 + static_cast<uint64_t>((((tap / 3) * kPitch + (tap % 3)) * 64) >> 4);
                const uint64_t bdesc = by8_desc_base + static_cast<uint64_t>((tap * kHalfTap8) >> 4);
                umma2_f8_k64(d_tmem, adesc, bdesc, idesc8);
              }
            } else if constexpr (STACK) {
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                const uint64_t adesc = a_desc0 + static_cast<uint64_t>((((tap / 3) * kPitch + (tap % 3)) * 128) >> 4);
                const uint64_t bdesc = ap == 0 ? b_desc_base + static_cast<uint64_t>((tap * 2 * kHalfTap) >> 4)
                                               : by_desc_base + static_cast<uint64_t>((tap * kHalfTap) >> 4);
                umma2_f16_k64(d_tmem, adesc, bdesc, ap == 0 ? idesc2 : idesc, (tap == 0) ? (first ? 0u : 1u) : 1u);
              }
            } else {
              const int nwp = (ap == 0) ? p.n_w_parts : 1;  // (hi,hi) (hi,lo) | (lo,hi)
              for (int wp = 0; wp < nwp; ++wp) {
#pragma unroll
                for (int tap = 0; tap < 9; ++tap) {
                  const uint64_t adesc = a_desc0 + static_cast<uint64_t>((((tap / 3) * kPitch + (tap % 3)) * 128) >> 4);
                  const uint64_t bdesc = b_desc_base + static_cast<uint64_t>((((kUp ? phase : wp) * 9 + tap) * (kWide ? 2 : 1) * kHalfTap) >> 4);
                  umma2_f16_k64(d_tmem, adesc, bdesc, kWide ? idesc2 : idesc, (tap == 0) ? (first ? 0u : 1u) : 1u);
                }
                first = 0;
              }
            }
            // frees this halo stage in both CTAs (PH = 4: after this warp's last phase of the tile; both warps commit)
            if (PH == 1 || phase >= PH - kPairIssuers) umma2_commit_multicast(b_empty + 8 * s);
            if (ap == p.n_act_parts - 1) umma2_commit_multicast(b_tfull + 8 * acc);
          }
          __syncwarp();
          first = 0;
        }
      }
      if (p.dbg_cycles != nullptr && lane == 0 && my_iters > 0) {
        // wait for this warp's last accumulator so the count covers the execution of its MMAs
        const int nv = my_iters * PH;
        const int last = ((nv - 1 - mw) / kPairIssuers) * kPairIssuers + mw;
        if (last >= 0 && last < nv) mbar_wait(b_tfull + 8 * (last % NB), (last / NB) & 1);
        p.dbg_cycles[pair_id + 128 * mw] = static_cast<unsigned long long>(clock64() - c_start);
      }
    }
  } else {
    // ===================== epilogue warps: group g = warp/4 drains tiles g, g+G, ... =====================
    const int grp = warp / EPW, q = warp & 3, chalf = (warp >> 2) & 1;   // warp % 4 = its TMEM lane quarter
    pdl_wait();   // addend reads / output writes must not overtake the previous kernel
    PairTileCursor cur;
    int cur_i = -1;
    for (int v = grp; v < my_iters * PH; v += G) {
      const int i = v / PH, phase = v % PH;
      if (cur_i < 0) {
        const int j0 = pair_id + i * num_pairs;   // pair index in walking order
        cur.init(2 * (p.reverse ? pair_tiles - 1 - j0 : j0) + static_cast<int>(rank), 2 * num_pairs, p.tiles_x, p.tiles_y,
                 p.reverse);
        cur_i = i;
      }
      for (; cur_i < i; ++cur_i) cur.advance();   // G / PH tiles (PH = 1) or none (the next phase of the same tile)
      bool valid = cur.b < p.B;
      int tx = cur.tx, ty = cur.ty;
      int b = valid ? cur.b : 0;
      if (p.dbg & 1024) {   // bench experiment: the divisions of the first form
        const int t = 2 * (pair_id + i * num_pairs) + static_cast<int>(rank);
        valid = t < p.num_tiles; tx = t % p.tiles_x; ty = (t / p.tiles_x) % p.tiles_y;
        b = valid ? t / (p.tiles_x * p.tiles_y) : 0;
      }
      const uint32_t acc = v % NB;
      const uint32_t acc_ph = (v / NB) & 1;
      const uint32_t taddr = tmem_base + acc * kAccCols + (static_cast<uint32_t>(q * 32) << 16);
      if constexpr (kWide) {
#pragma unroll 1   // one copy of the epilogue body (two inlined copies push its arrays into local memory)
        for (int u = 0; u < 2; ++u)
          epilogue_store16<N, 0, false>(p, s_bias, taddr + u * N, q, lane, b, ty, tx, valid, b_tfull + 8 * acc, acc_ph, 2, phase, u);
      } else if constexpr (EPI == EPI_STORE16 && EPW == 8)
        epilogue_store16_half<N, STACK, ADD>(p, s_bias, taddr, q, chalf, lane, b, ty, tx, valid, b_tfull + 8 * acc, acc_ph);
      else if constexpr (EPI == EPI_STORE16)
        epilogue_store16<N, STACK, ADD>(p, s_bias, taddr, q, lane, b, ty, tx, valid, b_tfull + 8 * acc, acc_ph,
                                        PH == 4 ? 2 : 1, PH == 4 ? (phase >> 1) : 0, PH == 4 ? (phase & 1) : 0);
      else
        epilogue_final<N, STACK>(p, s_bias, taddr, q, lane, b, ty, tx, valid, b_tfull + 8 * acc, acc_ph);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_shared(b_tempty + 8 * acc, 0));
    }
  }

  tc_fence_before();
  cluster_sync_all();   // the leader's MMAs read the peer's shared memory: nobody leaves before everything retired
  if (warp == kMmaWarp) tmem_dealloc2(tmem_base, kTmemCols);
}

}  // namespace isr
