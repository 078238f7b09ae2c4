// isr_conv_final.cuh -- the network's output convolution (64 -> Cout <= 8, 3x3, pad 1) with its fused epilogue
// (/root/reference/SuperresolutionNetwork/models/enhancenet.py:124 postblock[8], :51-90 _recon_image;
//  mainVideoUnshaded.py:460-465 clamp / normalise), as a TAP-STACKED implicit GEMM.
//
// Why not the generic kernel: a tcgen05.mma with operands in shared memory costs ~54 cycles for every N <= 64
// (profiles/r01_probe_mma_cycles.json), so 9 taps x 4 k-steps of N=16 MMAs cost as much as a 64-channel layer while using
// 7.6 % of the tensor pipe.  Here the nine taps are stacked along N instead of being nine shifted A windows:
//     P[halo pixel r, tap*CP + co] = sum_ci X[r, ci] * W[co, ci, tap]          (one un-shifted pass over the halo tile)
//     out[p, co] = sum_tap P[p + shift(tap), tap*CP + co]                      (gather in the epilogue, via shared memory)
// The 10x18 halo (180 pixel rows of 128 B, contiguous in shared memory) is covered by two M=128 row groups (rows 0..127
// and 52..179): 2 x 4 = 8 MMAs per pass instead of 36.  CP = Cout padded to {4, 6, 8}; N = 9*CP rounded up to 16.
//
// Warp roles as in isr_conv.cuh: warps 0-3 / 4-7 = epilogue group 0 / 1 (kFinalGroups groups; group g drains accumulator
// buffer g and owns gather buffer g), then the TMA producer warp and the MMA issuer warp (TMEM owner).
#pragma once
#include "isr_conv.cuh"

namespace isr {

// Tile of the output conv: 16 px wide x 8 rows (the hidden layers use 8 x 16): its outputs are fp32 NCHW planes, and a
// 16-px row segment is 64 B = one HBM burst, where the 8-px segment (32 B) cost 0.3 ms of the launch in partial bursts
// (tools/final_experiments.py).  Same 180 halo rows (18 x 10), same two M=128 row groups.
constexpr int kFTileW = 16, kFTileH = 8, kFPitch = kFTileW + 2, kFHaloH = kFTileH + 2;
static_assert(kFPitch * kFHaloH == 180, "the two MMA row groups (0..127, 52..179) assume 180 halo rows");
constexpr int kFinalGroups = 3;                                   // epilogue groups of four warps
constexpr int kFinalIssuers = kFinalGroups;                       // MMA-issuing warps: warp m issues tiles m, m+G, ... into buffer m
constexpr int kFinalThreads = (4 * kFinalGroups + 1 + kFinalIssuers) * 32;   // + TMA producer warp + MMA issuer warps

template <int CP>
struct FinalCfg {
  static constexpr int kCols = 9 * CP;                       // useful accumulator columns
  static constexpr int kN = (kCols + 15) / 16 * 16;          // UMMA N
  static constexpr int kLd = (kCols + 7) / 8 * 8;            // columns fetched from TMEM (x8 loads)
  // floats per halo row in the gather buffer: an ODD number of 8-byte units, so that both the 64-bit stores of phase A
  // (a half-warp = 16 consecutive halo rows) and the 64-bit tap loads of phase B (a half-warp = the 16 pixels of one
  // tile row) fall on 16 distinct bank pairs.  (An odd number of 16-byte units with 128-bit stores left phase B with
  // 2-way conflicts: ncu counted 63 M conflicts in 149 M shared wavefronts, LSU data pipe 68 % busy.)
  static constexpr int kPPitch = ((kLd / 2) % 2 == 1) ? kLd : kLd + 2;
  static constexpr int kSkipFloats = 6 * 4 * 8;              // low-res skip patch of a tile: 6 rows x 4 cols x <= 8 channels
  static constexpr int kPFloats = kFPitch * kFHaloH * kPPitch;  // gather buffer proper; the skip patch follows it
  static constexpr int kPBytes = ((kPFloats + kSkipFloats) * 4 + 1023) / 1024 * 1024;
  static constexpr uint32_t kTmemCols = (2 * kFinalGroups * kN <= 256) ? 256 : 512;   // one accumulator pair per group
};

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t *v) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <int CP>
__global__ void __launch_bounds__(kFinalThreads, 1)
conv3x3_final_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                     const ConvParams p) {
  using C = FinalCfg<CP>;
  constexpr int N = C::kN;
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr uint32_t kWPartBytes = N * 128;                  // one weight part: N rows of 64 16-bit
  const uint32_t w_bytes = p.n_w_parts * kWPartBytes;
  const uint32_t s_w = smem_base;
  const uint32_t s_p0 = s_w + ((w_bytes + 1023u) & ~1023u);  // two gather buffers
  const uint32_t s_stage0 = s_p0 + kFinalGroups * C::kPBytes;
  const uint32_t s_bar = s_stage0 + p.num_stages * p.stage_bytes;
  // barrier block: full[8], empty[8], wbar, tmem_full[G], tmem_empty[G], tmem_ptr.  One accumulator buffer PER GROUP: a
  // group must observe every phase of the barriers it waits on (parity waits alias after two phases), so buffers are
  // not shared between groups.
  const uint32_t b_full = s_bar, b_empty = s_bar + 64, b_w = s_bar + 128, b_tfull = s_bar + 136,
                 b_tempty = b_tfull + 8 * kFinalGroups, s_tptr = b_tempty + 8 * kFinalGroups;
  static_assert(136 + 16 * kFinalGroups + 4 <= 256, "barrier block overflows its 256-byte reservation");
  __shared__ float s_bias[8];
  if (threadIdx.x < 8) s_bias[threadIdx.x] = (p.bias != nullptr && threadIdx.x < p.cout) ? p.bias[threadIdx.x] : 0.f;

  constexpr int kTmaWarp = 4 * kFinalGroups, kMmaWarp = 4 * kFinalGroups + 1;
  if (warp == kTmaWarp && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(b_full + 8 * s, 1);
      mbar_init(b_empty + 8 * s, 1);
    }
    mbar_init(b_w, 1);
    for (int a = 0; a < kFinalGroups; ++a) {
      mbar_init(b_tfull + 8 * a, 1);
      mbar_init(b_tempty + 8 * a, 4);
    }
    fence_barrier_init();
  }
  if (warp == kMmaWarp) tmem_alloc(s_tptr, C::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(s_tptr));

  const int my_tiles = (p.num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                       static_cast<int>(gridDim.x);

  if (warp == kTmaWarp) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(b_w, w_bytes);
      for (int wp = 0; wp < p.n_w_parts; ++wp)
        tma_load_2d(s_w + wp * kWPartBytes, &tmW, b_w, 0, p.w_row0 + wp * p.w_row_stride);
    }
    __syncwarp();
    uint32_t it = 0;
    for (int i = 0; i < my_tiles; ++i) {
      const int t = blockIdx.x + i * gridDim.x;
      const int tx = t % p.tiles_x;
      const int ty = (t / p.tiles_x) % p.tiles_y;
      const int b = t / (p.tiles_x * p.tiles_y);
      for (int ap = 0; ap < p.n_act_parts; ++ap, ++it) {
        const uint32_t s = it % p.num_stages;
        const uint32_t ph = (it / p.num_stages) & 1;
        mbar_wait(b_empty + 8 * s, ph ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(b_full + 8 * s, kFPitch * kFHaloH * 128);
          tma_load_4d(s_stage0 + s * p.stage_bytes, &tmA, b_full + 8 * s, ap * 64, tx * kFTileW - 1, ty * kFTileH - 1, b);
        }
        __syncwarp();
      }
    }
  } else if (warp >= kMmaWarp) {
    // ===================== MMA issuers: one warp per accumulator buffer (p.issuers = 3), so the per-tile bookkeeping of
    // one warp (barrier waits, commits) hides behind the MMAs of the others -- see isr_conv_pair.cuh.  p.issuers = 1 when
    // the stage count is not a multiple of 3 * n_act_parts (a stage must always be consumed by the same warp).
    const int mw = warp - kMmaWarp;
    const uint32_t idesc = make_idesc_f16(128, N, p.is_bf16);
    const uint64_t b_desc_base = make_smem_desc_sw128(s_w, 1024, 0);
    mbar_wait(b_w, 0);
    for (int i = mw; i < my_tiles && mw < p.issuers; i += p.issuers) {
      const uint32_t acc = i % kFinalGroups;
      const uint32_t acc_ph = (i / kFinalGroups) & 1;
      mbar_wait(b_tempty + 8 * acc, acc_ph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * (2 * N);
      uint32_t first = 1;
      for (int ap = 0; ap < p.n_act_parts; ++ap) {
        const uint32_t it = static_cast<uint32_t>(i) * p.n_act_parts + ap;
        const uint32_t s = it % p.num_stages;
        const uint32_t ph = (it / p.num_stages) & 1;
        mbar_wait(b_full + 8 * s, ph);
        tc_fence_after();
        const uint32_t a_base = s_stage0 + s * p.stage_bytes;
        // halo rows are contiguous 128-byte rows: canonical K-major SW128 tiles, 8-row groups 1024 B apart
        const uint64_t a_desc_g0 = make_smem_desc_sw128(a_base, 1024, 0);               // halo rows 0..127
        const uint64_t a_desc_g1 = make_smem_desc_sw128(a_base + 52 * 128, 1024, 0);    // halo rows 52..179
        const int nwp = (ap == 0) ? p.n_w_parts : 1;  // (hi,hi) (hi,lo) | (lo,hi)
        if (elect_one()) {
          for (int wp = 0; wp < nwp; ++wp) {
            const uint64_t bdesc = b_desc_base + static_cast<uint64_t>((wp * kWPartBytes) >> 4);
            umma_f16_k64(d_tmem, a_desc_g0, bdesc, idesc, first ? 0u : 1u);
            umma_f16_k64(d_tmem + N, a_desc_g1, bdesc, idesc, first ? 0u : 1u);
            first = 0;
          }
          umma_commit(b_empty + 8 * s);
          if (ap == p.n_act_parts - 1) umma_commit(b_tfull + 8 * acc);
        }
        __syncwarp();
        first = 0;
      }
    }
  } else {
    // ===================== epilogue warps: group g handles tiles g, g+G, ... (accumulator buffer = tile & 1) ======
    const int grp = warp >> 2, q = warp & 3;
    const uint32_t s_p = s_p0 + grp * C::kPBytes;
    float *P = reinterpret_cast<float *>(smem_raw + (s_p - smem_u32(smem_raw)));
    for (int i = grp; i < my_tiles; i += kFinalGroups) {
      const int t = blockIdx.x + i * gridDim.x;
      const int tx = t % p.tiles_x;
      const int ty = (t / p.tiles_x) % p.tiles_y;
      const int b = t / (p.tiles_x * p.tiles_y);
      const uint32_t acc = grp;
      const uint32_t acc_ph = (i / kFinalGroups) & 1;
      const uint32_t taddr = tmem_base + acc * (2 * N) + (static_cast<uint32_t>(q * 32) << 16);
      const int row = q * 32 + lane;   // output pixel of this thread inside the tile
      const int py = row >> 4, px = row & 15;   // a warp = two 16-px rows = two 64-byte segments per plane
      const int y = ty * kFTileH + py, x = tx * kFTileW + px;
      const bool inb = (y < p.H) && (x < p.W);
      // Skip connection (_recon_image, enhancenet.py:67-71): the 4x bilinear / nearest upsample of the low-res input.
      // A tile's 16 x 8 pixels read a 6 x 4 (cols x rows) low-res patch per channel: the group stages it in shared memory (one or two
      // loads per thread, issued before the accumulator wait; visible after the barrier between phases A and B) instead
      // of 24 scattered global loads per thread.  Patch origin: row 2 ty - 1, column 4 tx - 1, indices clamped like bil_src.
      float lx0 = 1.f, lx1 = 0.f, ly0 = 1.f, ly1 = 0.f;
      int pr0 = 0, pr1 = 0, pc0 = 0, pc1 = 0;   // patch rows / columns of this pixel's taps
      float *S = P + C::kPFloats;
      float sv[2] = {0.f, 0.f};
      int n_skip = 0;
      const bool want_skip = p.recon_residual && !(p.dbg & 128);   // dbg bit7: bench experiment without the skip
      if (want_skip) {
        const int by = 2 * ty - 1, bx = 4 * tx - 1;
        const int nm = p.n_mask > 0 ? (p.n_mask < 8 ? p.n_mask : 8) : 1;
        n_skip = nm * 24;
        // the loads go to REGISTERS here and to shared memory only after phase A: a store to S right away would make the
        // warp wait for the load (queued behind the previous tile's output stores) before it can start draining TMEM
#pragma unroll
        for (int k = 0; k < 2; ++k) {
          const int e = row + 128 * k;
          if (e < nm * 24) {
            const int c = e / 24, r = (e % 24) / 6, cc = (e % 24) % 6;
            const int yy = min(max(by + r, 0), p.h_low - 1), xx = min(max(bx + cc, 0), p.w_low - 1);
            sv[k] = __ldg(p.skip_src + b * p.skip_bstride + (static_cast<size_t>(c) * p.h_low + yy) * p.w_low + xx);
          }
        }
        int x0, x1, y0, y1;
        if (p.upsample == 1) {
          bil_src(x, p.w_low, 0.25f, x0, x1, lx0, lx1);
          bil_src(y, p.h_low, 0.25f, y0, y1, ly0, ly1);
        } else {
          x0 = x1 = min(x >> 2, p.w_low - 1);
          y0 = y1 = min(y >> 2, p.h_low - 1);
        }
        pr0 = min(max(y0 - by, 0), 3); pr1 = min(max(y1 - by, 0), 3);
        pc0 = min(max(x0 - bx, 0), 5); pc1 = min(max(x1 - bx, 0), 5);
      }
      mbar_wait(b_tfull + 8 * acc, acc_ph);
      tc_fence_after();
      // ---- phase A: per-tap partial sums TMEM -> registers -> gather buffer (row = halo pixel) ----
      {
        uint32_t v[C::kLd];
#pragma unroll
        for (int c8 = 0; c8 < C::kLd / 8; ++c8) tmem_ld8(taddr + 8 * c8, v + 8 * c8);
        tmem_ld_wait();
        float2 *dst = reinterpret_cast<float2 *>(P + static_cast<size_t>(row) * C::kPPitch);
#pragma unroll
        for (int c2 = 0; c2 < C::kLd / 2; ++c2) dst[c2] = make_float2(__uint_as_float(v[2 * c2]), __uint_as_float(v[2 * c2 + 1]));
        if (q >= 2) {   // second row group: accumulator lane L holds halo row 52 + L; rows 128..179 are new
#pragma unroll
          for (int c8 = 0; c8 < C::kLd / 8; ++c8) tmem_ld8(taddr + N + 8 * c8, v + 8 * c8);
          tmem_ld_wait();
          const int hr = 52 + row;
          if (hr >= 128) {
            float2 *d2 = reinterpret_cast<float2 *>(P + static_cast<size_t>(hr) * C::kPPitch);
#pragma unroll
            for (int c2 = 0; c2 < C::kLd / 2; ++c2) d2[c2] = make_float2(__uint_as_float(v[2 * c2]), __uint_as_float(v[2 * c2 + 1]));
          }
        }
      }
      if (row < n_skip) S[row] = sv[0];
      if (row + 128 < n_skip) S[row + 128] = sv[1];
      // the accumulator buffer is free as soon as it is in shared memory
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(b_tempty + 8 * acc);
      named_bar_sync(1 + grp, 128);
      // ---- phase B: out[p, co] = sum over taps of P[p + shift(tap)][tap*CP + co] ----
      float rec[CP];
#pragma unroll
      for (int c = 0; c < CP; ++c) rec[c] = s_bias[c];
#pragma unroll
      for (int tap = 0; tap < 9; ++tap) {
        const float *src = P + static_cast<size_t>((py + tap / 3) * kFPitch + px + tap % 3) * C::kPPitch + tap * CP;
#pragma unroll
        for (int c2 = 0; c2 < CP / 2; ++c2) {
          const float2 f = *reinterpret_cast<const float2 *>(src + 2 * c2);
          rec[2 * c2] += f.x;
          rec[2 * c2 + 1] += f.y;
        }
      }
      if (inb && !((p.dbg & 64) && rec[0] != 12345.678f)) {   // dbg bit6: bench experiment without the global stores
        const size_t plane = static_cast<size_t>(p.H) * p.W;
        const size_t o0 = static_cast<size_t>(b) * p.cout * plane + static_cast<size_t>(y) * p.W + x;
#pragma unroll
        for (int c = 0; c < CP; ++c) {
          if (c < p.cout) {
            const float raw = rec[c];
            if (p.out_raw != nullptr) __stcs(p.out_raw + o0 + c * plane, raw);
            if (want_skip && c < p.n_mask) {
              // _recon_image (enhancenet.py:67-71): bilinear (x first, then y) or nearest 4x skip + raw
              const float *sc = S + c * 24;
              const float t00 = sc[pr0 * 6 + pc0], t01 = sc[pr0 * 6 + pc1], t10 = sc[pr1 * 6 + pc0], t11 = sc[pr1 * 6 + pc1];
              const float r0 = fmaf(lx0, t00, lx1 * t01);
              const float r1 = fmaf(lx0, t10, lx1 * t11);
              rec[c] = (p.upsample == 1 ? fmaf(ly0, r0, ly1 * r1) : t00) + raw;
            }
            if (p.out_recon != nullptr) __stcs(p.out_recon + o0 + c * plane, rec[c]);
          }
        }
        if (p.out_state != nullptr) {
          if (p.state_unshaded && p.cout >= 5) {
            if constexpr (CP >= 6) {
              // mainVideoUnshaded.py:460-465: clamp mask, normalise normal (shading.py:193-207), clamp depth / ao
              __stcs(p.out_state + o0, fminf(fmaxf(rec[0], -1.f), 1.f));
              float len = sqrtf(rec[1] * rec[1] + rec[2] * rec[2] + rec[3] * rec[3]);
              len = fmaxf(len, 1e-7f);
              __stcs(p.out_state + o0 + plane, rec[1] / len);
              __stcs(p.out_state + o0 + 2 * plane, rec[2] / len);
              __stcs(p.out_state + o0 + 3 * plane, rec[3] / len);
#pragma unroll
              for (int c = 4; c < CP; ++c)
                if (c < p.cout) __stcs(p.out_state + o0 + c * plane, fminf(fmaxf(rec[c], 0.f), 1.f));
            }
          } else {
#pragma unroll
            for (int c = 0; c < CP; ++c)
              if (c < p.cout) __stcs(p.out_state + o0 + c * plane, fminf(fmaxf(rec[c], 0.f), 1.f));
          }
        }
      }
      named_bar_sync(1 + grp, 128);   // gather buffer free for this group's next tile
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kMmaWarp) tmem_dealloc(tmem_base, C::kTmemCols);
}

}  // namespace isr
