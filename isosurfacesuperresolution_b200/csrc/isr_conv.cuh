// isr_conv.cuh -- conv3x3 (pad 1, Cin = 64 per launch) as an implicit GEMM on tcgen05 tensor cores.
//
// Replaces the cuDNN convolutions behind nn.Conv2d in the reference's EnhanceNet
// (/root/reference/SuperresolutionNetwork/models/enhancenet.py:93-125): preblock, 20 residual-block convs, 4 postblock convs.
//
// GEMM view per launch: D[M = pixels, N = Cout] = sum over 9 taps of A_tap[M, K = 64] * W_tap[N, K]^T.
//   * Activations are NHWC 16-bit (bf16 or fp16), channel dim = 64 * n_parts (hi | lo planes for split precision).
//   * One CTA tile = 8 px wide x 16 rows = 128 GEMM rows.  Its (8+2) x (16+2) halo is fetched ONCE by a 4-D TMA box
//     (out-of-bounds = zero fill = the conv's zero padding) into 128B-swizzled shared memory (one pixel = one 128-byte row);
//     the nine taps are nine UMMA shared-memory descriptors into that same halo tile (start address shifted by
//     (ky*pitch + kx) rows, 8-row-group stride = pitch*128 B) -- no im2col copy, no re-read.
//   * All nine taps' weights ([part][tap][Cout][64], K-major, 128B swizzle) stay resident in shared memory for the whole
//     persistent CTA; accumulators live in TMEM (double buffered: epilogue of tile i overlaps the MMAs of tile i+1).
//   * Warp roles: warps 0-7 epilogue in two groups of four (group g drains accumulator buffer g; warp w reads TMEM lanes
//     32*(w%4)..+31), warp 8 TMA producer, warp 9 MMA issuer + TMEM owner.
//   * 64-wide epilogue: tcgen05.ld 16x256b + a permuted GEMM-column order (pack_weights_kernel) leave 16 consecutive
//     channels of a pixel in each thread, so a warp's 16-byte stores cover whole 128-byte lines (no smem staging).
//   * Split precision ("x3"): passes (Ahi,Whi), (Ahi,Wlo), (Alo,Whi) accumulate into the same TMEM tile.
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include "isr_ptx.cuh"

namespace isr {

constexpr int kTileW = 8;
constexpr int kTileH = 16;
constexpr int kHaloH = kTileH + 2;
constexpr int kPitch = kTileW + 2;  // halo pitch in pixels; 8-row-group stride of the A descriptors = kPitch * 128 B
constexpr int kConvThreads = 320;  // warps 0-7 epilogue (two groups of 4), warp 8 TMA producer, warp 9 MMA issuer

enum { EPI_STORE16 = 0, EPI_FINAL = 1 };

struct ConvParams {
  int B, H, W;
  int tiles_x, tiles_y, num_tiles;
  int n_act_parts;       // 1 | 2 (input channel dim = 64 * n_act_parts)
  int n_w_parts;         // 1 | 2
  int is_bf16;
  int w_row0;            // first row of this launch's weights inside the weight tensor map
  int w_row_stride;      // rows between consecutive (part, tap) blocks
  int num_stages, stage_bytes;
  int issuers;           // conv3x3_final_kernel: MMA-issuing warps in use (1 | 3)
  int dbg;               // bench experiments only: bit0 = epilogue does not store, bit1 = no MMAs are issued,
                         // bit2 = epilogue only does the barrier handshake, bit3 = MMA warp does not wait for the epilogue,
                         // bit4 = the TMA producer fills each halo stage once and only signals afterwards (pair kernel)
  unsigned long long *dbg_cycles;   // bench experiment (dbg bit5): cycles spent in the MMA issue loop, per CTA pair
  int relu;
  const float *bias;     // N floats (zero padded), or nullptr
  const uint16_t *addend;  // [B,H,W,64*add_parts] or nullptr
  int add_parts;
  uint16_t *out16;       // [B,H,W,64*out_parts]
  int out_parts;
  // EPI_FINAL
  float *out_raw, *out_recon, *out_state;  // B x cout x H x W fp32 NCHW (any may be null)
  int cout, n_mask, upsample, recon_residual, state_unshaded;
  const float *skip_src;  // low-res fp32 NCHW input, first n_mask channels used
  long long skip_bstride;  // floats between batches of skip_src
  int h_low, w_low;
  // 8-bit low plane of the split-precision trunk (pair kernel, STACK; 0 everywhere = the 16-bit forms):
  // Tensors of that form are PLANAR: a hi plane (128 B per pixel), a lo fp16 plane (128 B) and a lo8 plane (64 B), each
  // contiguous (interleaving them in 320-byte pixels made the TMA / L2 fetch the unused plane too: 604 MB of DRAM reads
  // where 280 MB were needed, and a DRAM-bound 190 us launch).
  int a_lo8;              // the second activation part is the e4m3(lo * 2^6) plane (tmA8): the A_lo x W correction pass runs
                          // as kind::f8f6f4 (K = 32 per MMA) against e5m2(W * 2^-6) (tmW8)
  int out_pitch;          // bytes per output pixel of the hi (+ interleaved lo) plane (0 = 128 * out_parts)
  int add_pitch;          // bytes per addend pixel (0 = 128 * add_parts)
  uint16_t *out_lo16;     // planar lo fp16 output plane (nullptr: interleaved at +128 B of the output pixel when out_parts == 2)
  uint8_t *out_lo8;       // planar e4m3(lo * 2^6) output plane, 64 B per pixel (nullptr: none)
  const uint16_t *addend_lo;   // planar lo fp16 plane of the addend (nullptr: interleaved at +128 B when add_parts == 2)
  int reverse;            // 1: the persistent CTAs walk the tiles from the last to the first.  Consecutive launches alternate,
                          // so that a launch starts on the part of its input the previous launch wrote (or read) last --
                          // the part that is still in the 126 MB L2 (the tensors of a chunk are 100 MB .. 2 GB each)
  const uint8_t *addend_lo8;   // planar e4m3(lo * 2^6) plane of the addend (add_parts == 1): the residual stream kept as
                               // hi + 8-bit lo, the very planes the convs read (192 instead of 320 B per pixel and tensor)
};
constexpr int kLo8Bytes = 320;   // bytes per pixel of a tensor with all three planes (128 + 128 + 64)

__device__ __forceinline__ uint32_t pack2(float a, float b, int is_bf16) {
  if (is_bf16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&v);
  }
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t *>(&v);
}
__device__ __forceinline__ float2 unpack2(uint32_t u, int is_bf16) {
  if (is_bf16) return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162 *>(&u));
  return __half22float2(*reinterpret_cast<__half2 *>(&u));
}

// torch upsample_bilinear2d(align_corners=False) source index / weights for an integer scale
__device__ __forceinline__ void bil_src(int dst, int n_in, float inv_scale, int &i0, int &i1, float &l0, float &l1) {
  float src = inv_scale * (static_cast<float>(dst) + 0.5f) - 0.5f;
  src = src < 0.f ? 0.f : src;
  int a = static_cast<int>(src);
  a = a > n_in - 1 ? n_in - 1 : a;
  i0 = a;
  i1 = a < n_in - 1 ? a + 1 : a;
  float t = src - static_cast<float>(a);
  t = fminf(fmaxf(t, 0.f), 1.f);
  l1 = t;
  l0 = 1.f - t;
}

// Epilogue of a 64-wide layer for one 8x16 tile (one warp = TMEM lanes 32q..32q+31): waits for the accumulator, then
// bias (+ residual addend) (+ ReLU) -> 16-bit NHWC hi (and lo) planes.  Shared by the single-CTA and the CTA-pair kernel.
// thread (m = lane%4, rr = lane/4): pixel column rr of image rows 4q..4q+3 of the tile, channels 16m..16m+15
template <int N, int STACK, bool ADD = true>   // ADD = false: the layer never has a residual addend (frees its 64 registers)
__device__ __forceinline__ void epilogue_store16(const ConvParams &p, const float *s_bias, uint32_t taddr, int q, int lane,
                                                 int b, int ty, int tx, bool tile_valid, uint32_t bar_tfull,
                                                 uint32_t acc_ph, int omul = 1, int oyadd = 0, int oxadd = 0) {
  // omul / oyadd / oxadd: tile pixel (y, x) is stored at output pixel (omul * y + oyadd, omul * x + oxadd) of an
  // (omul * H) x (omul * W) tensor -- the phase interleave of the fused upsample + conv layer (omul = 2)
  const uint32_t b_tfull = bar_tfull;
  constexpr uint32_t acc = 0;   // bar_tfull already addresses this tile's accumulator barrier
    // thread (m = lane%4, rr = lane/4): pixel column rr of image rows 4q..4q+3 of the tile, channels 16m..16m+15
    const int m = lane & 3, rr = lane >> 2;
    const int x = tx * kTileW + rr;
    float bias_r[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) bias_r[j] = s_bias[16 * m + j];
    // residual addend (the block's skip): fetched BEFORE waiting for the accumulator, so its HBM latency hides
    // behind the tile's MMAs.  ad[pixel j = 2*half+hh][part][2 x 16 B]
    uint32_t ad[4][2][8];
    const bool has_add = ADD && p.addend != nullptr;
    const size_t add_pitch = p.add_pitch > 0 ? p.add_pitch : 128 * p.add_parts;
    const size_t out_pitch = p.out_pitch > 0 ? p.out_pitch : 128 * p.out_parts;
    if (has_add) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int y = ty * kTileH + 4 * q + j;
        const bool ok = tile_valid && (y < p.H) && (x < p.W);
        const size_t pix = (static_cast<size_t>(b) * p.H + (ok ? y : 0)) * p.W + (ok ? x : 0);
        const uint8_t *ab = reinterpret_cast<const uint8_t *>(p.addend) + pix * add_pitch + 32 * m;
        ldg_v8(ab, ad[j][0]);
        if (p.add_parts == 2)
          ldg_v8(p.addend_lo != nullptr ? reinterpret_cast<const uint8_t *>(p.addend_lo) + pix * 128 + 32 * m : ab + 128, ad[j][1]);
        else if (p.addend_lo8 != nullptr)
          asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(ad[j][1][0]), "=r"(ad[j][1][1]), "=r"(ad[j][1][2]), "=r"(ad[j][1][3])
                       : "l"(p.addend_lo8 + pix * 64 + 16 * m));
      }
    }
    mbar_wait_parked(b_tfull + 8 * acc, acc_ph);
    tc_fence_after();
    if (p.dbg & 4) return;   // bench experiment: barrier handshake only
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      uint32_t v[32];
      tmem_ld_16x256b_x8(taddr + (static_cast<uint32_t>(half * 16) << 16), v);
      if constexpr (STACK) {
        uint32_t v2[32];
        tmem_ld_16x256b_x8(taddr + N + (static_cast<uint32_t>(half * 16) << 16), v2);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]));
      } else {
        tmem_ld_wait();
      }
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int y = ty * kTileH + 4 * q + 2 * half + hh;
        if (tile_valid && y < p.H && x < p.W) {
          const size_t pix = (static_cast<size_t>(b) * (p.H * omul) + (y * omul + oyadd)) * (p.W * omul) + (x * omul + oxadd);
          float f[16];
#pragma unroll
          for (int rep = 0; rep < 8; ++rep) {
            f[2 * rep] = __uint_as_float(v[4 * rep + 2 * hh]) + bias_r[2 * rep];
            f[2 * rep + 1] = __uint_as_float(v[4 * rep + 2 * hh + 1]) + bias_r[2 * rep + 1];
          }
          if (has_add) {
#pragma unroll
            for (int ap = 0; ap < 2; ++ap) {
              if (ap < p.add_parts) {
                const uint32_t *w = ad[2 * half + hh][ap];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const float2 a = unpack2(w[j], p.is_bf16);
                  f[2 * j] += a.x;
                  f[2 * j + 1] += a.y;
                }
              }
            }
          }
          if (has_add && p.add_parts == 1 && p.addend_lo8 != nullptr) {
            const uint32_t *w = ad[2 * half + hh][1];   // 16 e4m3 bytes = lo * 2^6 of channels 16 m .. 16 m + 15
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const __half2_raw h01 = __nv_cvt_fp8x2_to_halfraw2(static_cast<__nv_fp8x2_storage_t>(w[j] & 0xffffu), __NV_E4M3);
              const __half2_raw h23 = __nv_cvt_fp8x2_to_halfraw2(static_cast<__nv_fp8x2_storage_t>(w[j] >> 16), __NV_E4M3);
              const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&h01));
              const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&h23));
              f[4 * j] += a.x * 0.015625f;
              f[4 * j + 1] += a.y * 0.015625f;
              f[4 * j + 2] += c.x * 0.015625f;
              f[4 * j + 3] += c.y * 0.015625f;
            }
          }
          if (p.relu) {
#pragma unroll
            for (int j = 0; j < 16; ++j) f[j] = fmaxf(f[j], 0.f);
          }
          uint32_t hi[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) hi[j] = pack2(f[2 * j], f[2 * j + 1], p.is_bf16);
          uint8_t *ob = reinterpret_cast<uint8_t *>(p.out16) + pix * out_pitch + 32 * m;
          st_global_v8(ob, hi);
          if (p.out_parts == 2 || p.out_lo8 != nullptr) {
            float2 lf[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float2 h = unpack2(hi[j], p.is_bf16);
              lf[j] = make_float2(f[2 * j] - h.x, f[2 * j + 1] - h.y);
            }
            if (p.out_parts == 2) {
              uint32_t lo[8];
#pragma unroll
              for (int j = 0; j < 8; ++j) lo[j] = pack2(lf[j].x, lf[j].y, p.is_bf16);
              st_global_v8(p.out_lo16 != nullptr ? reinterpret_cast<uint8_t *>(p.out_lo16) + pix * 128 + 32 * m : ob + 128, lo);
            }
            if (p.out_lo8 != nullptr) {   // e4m3(lo * 2^6): the operand of the 8-bit correction pass of the next conv
              uint32_t l8[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const uint32_t a = __nv_cvt_float2_to_fp8x2(make_float2(lf[2 * j].x * 64.f, lf[2 * j].y * 64.f), __NV_SATFINITE, __NV_E4M3);
                const uint32_t c = __nv_cvt_float2_to_fp8x2(make_float2(lf[2 * j + 1].x * 64.f, lf[2 * j + 1].y * 64.f), __NV_SATFINITE, __NV_E4M3);
                l8[j] = a | (c << 16);
              }
              // channels 16 m .. 16 m + 15 = 16 bytes of the pixel's 64-byte row in the lo8 plane
              asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(p.out_lo8 + pix * 64 + 16 * m), "r"(l8[0]),
                           "r"(l8[1]), "r"(l8[2]), "r"(l8[3])
                           : "memory");
            }
          }
        }
      }
    }
}

// The same epilogue for HALF of the channels: the CTA-pair kernel runs EIGHT warps per epilogue group -- warps q and q + 4
// both own TMEM lanes 32q..32q+31 and split the accumulator columns (`chalf` = 0: GEMM columns 0..31, 1: columns 32..63;
// with the column permutation of pack_weights_kernel a thread then holds channels 16 m + 8 chalf .. + 7 of its four pixels).
// Why: one warp executes the whole chain of a tile serially (addend loads -> accumulator -> ~1000 dependent instructions
// -> stores); with an addend that chain took ~9000 cycles per tile and group against a budget of two tiles' MMAs (6200), so
// the convs with a residual addend were epilogue-bound (151 us against 119 without addend).  Half the work per warp, twice
// the warps; 16-byte instead of 32-byte accesses (two lanes per sector).
template <int N, int STACK, bool ADD = true>
__device__ __forceinline__ void epilogue_store16_half(const ConvParams &p, const float *s_bias, uint32_t taddr, int q, int chalf,
                                                      int lane, int b, int ty, int tx, bool tile_valid, uint32_t bar_tfull,
                                                      uint32_t acc_ph, int omul = 1, int oyadd = 0, int oxadd = 0) {
  const int m = lane & 3, rr = lane >> 2;
  const int x = tx * kTileW + rr;
  const int ch0 = 16 * m + 8 * chalf;   // first of this thread's 8 channels
  float bias_r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) bias_r[j] = s_bias[ch0 + j];
  uint32_t ad[4][2][4];   // [pixel][hi | lo][16 B]  (lo8: 8 B in [1][0..1])
  const bool has_add = ADD && p.addend != nullptr;
  const size_t add_pitch = p.add_pitch > 0 ? p.add_pitch : 128 * p.add_parts;
  const size_t out_pitch = p.out_pitch > 0 ? p.out_pitch : 128 * p.out_parts;
  if (has_add) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int y = ty * kTileH + 4 * q + j;
      const bool ok = tile_valid && (y < p.H) && (x < p.W);
      const size_t pix = (static_cast<size_t>(b) * p.H + (ok ? y : 0)) * p.W + (ok ? x : 0);
      const uint8_t *ab = reinterpret_cast<const uint8_t *>(p.addend) + pix * add_pitch + 2 * ch0;
      ldg_v4(ab, ad[j][0]);
      if (p.add_parts == 2)
        ldg_v4(p.addend_lo != nullptr ? reinterpret_cast<const uint8_t *>(p.addend_lo) + pix * 128 + 2 * ch0 : ab + 128, ad[j][1]);
      else if (p.addend_lo8 != nullptr)
        ldg_v2(p.addend_lo8 + pix * 64 + ch0, ad[j][1]);
    }
  }
  mbar_wait_parked(bar_tfull, acc_ph);
  tc_fence_after();
  if (p.dbg & 4) return;   // bench experiment: barrier handshake only
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    uint32_t v[16];
    tmem_ld_16x256b_x4(taddr + 32 * chalf + (static_cast<uint32_t>(half * 16) << 16), v);
    if constexpr (STACK) {
      uint32_t v2[16];
      tmem_ld_16x256b_x4(taddr + N + 32 * chalf + (static_cast<uint32_t>(half * 16) << 16), v2);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(v2[j]));
    } else {
      tmem_ld_wait();
    }
#pragma unroll
    for (int hh = 0; hh < 2; ++hh) {
      const int y = ty * kTileH + 4 * q + 2 * half + hh;
      if (tile_valid && y < p.H && x < p.W) {
        const size_t pix = (static_cast<size_t>(b) * (p.H * omul) + (y * omul + oyadd)) * (p.W * omul) + (x * omul + oxadd);
        float f[8];
#pragma unroll
        for (int rep = 0; rep < 4; ++rep) {
          f[2 * rep] = __uint_as_float(v[4 * rep + 2 * hh]) + bias_r[2 * rep];
          f[2 * rep + 1] = __uint_as_float(v[4 * rep + 2 * hh + 1]) + bias_r[2 * rep + 1];
        }
        if (has_add) {
#pragma unroll
          for (int ap = 0; ap < 2; ++ap) {
            if (ap < p.add_parts) {
              const uint32_t *w = ad[2 * half + hh][ap];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float2 a = unpack2(w[j], p.is_bf16);
                f[2 * j] += a.x;
                f[2 * j + 1] += a.y;
              }
            }
          }
          if (p.add_parts == 1 && p.addend_lo8 != nullptr) {
            const uint32_t *w = ad[2 * half + hh][1];   // 8 e4m3 bytes = lo * 2^6 of channels ch0 .. ch0 + 7
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const __half2_raw h01 = __nv_cvt_fp8x2_to_halfraw2(static_cast<__nv_fp8x2_storage_t>(w[j] & 0xffffu), __NV_E4M3);
              const __half2_raw h23 = __nv_cvt_fp8x2_to_halfraw2(static_cast<__nv_fp8x2_storage_t>(w[j] >> 16), __NV_E4M3);
              const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&h01));
              const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&h23));
              f[4 * j] += a.x * 0.015625f;
              f[4 * j + 1] += a.y * 0.015625f;
              f[4 * j + 2] += c.x * 0.015625f;
              f[4 * j + 3] += c.y * 0.015625f;
            }
          }
        }
        if (p.relu) {
#pragma unroll
          for (int j = 0; j < 8; ++j) f[j] = fmaxf(f[j], 0.f);
        }
        uint32_t hi[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) hi[j] = pack2(f[2 * j], f[2 * j + 1], p.is_bf16);
        uint8_t *ob = reinterpret_cast<uint8_t *>(p.out16) + pix * out_pitch + 2 * ch0;
        st_global_v4(ob, hi);
        if (p.out_parts == 2 || p.out_lo8 != nullptr) {
          float2 lf[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float2 h = unpack2(hi[j], p.is_bf16);
            lf[j] = make_float2(f[2 * j] - h.x, f[2 * j + 1] - h.y);
          }
          if (p.out_parts == 2) {
            uint32_t lo[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) lo[j] = pack2(lf[j].x, lf[j].y, p.is_bf16);
            st_global_v4(p.out_lo16 != nullptr ? reinterpret_cast<uint8_t *>(p.out_lo16) + pix * 128 + 2 * ch0 : ob + 128, lo);
          }
          if (p.out_lo8 != nullptr) {
            uint32_t l8[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              const uint32_t a = __nv_cvt_float2_to_fp8x2(make_float2(lf[2 * j].x * 64.f, lf[2 * j].y * 64.f), __NV_SATFINITE, __NV_E4M3);
              const uint32_t c = __nv_cvt_float2_to_fp8x2(make_float2(lf[2 * j + 1].x * 64.f, lf[2 * j + 1].y * 64.f), __NV_SATFINITE, __NV_E4M3);
              l8[j] = a | (c << 16);
            }
            st_global_v2(p.out_lo8 + pix * 64 + ch0, l8[0], l8[1]);
          }
        }
      }
    }
  }
}

// Epilogue of the output conv (Cout <= 16) for one 8x16 tile: raw = acc + bias; recon = raw + upsampled skip
// (enhancenet.py:51-90); optional recurrent state (mainVideoUnshaded.py:460-465).  thread = GEMM row (32x32b load):
// pixel (py, px) = (row/8, row%8); NCHW fp32 planes, 8 px = one 32 B sector.  Shared by the single-CTA and the CTA-pair
// kernel.
template <int N, int STACK>
__device__ __forceinline__ void epilogue_final(const ConvParams &p, const float *s_bias, uint32_t taddr, int q, int lane,
                                               int b, int ty, int tx, bool tile_valid, uint32_t bar_tfull,
                                               uint32_t acc_ph) {
  const int row = q * 32 + lane;
  const int y = ty * kTileH + (row >> 3), x = tx * kTileW + (row & 7);
  const bool inb = tile_valid && (y < p.H) && (x < p.W);
  // the skip taps do not depend on the accumulator: fetch them before waiting for the MMAs
  constexpr int kSkipPrefetch = 8;
  float sk[kSkipPrefetch];
#pragma unroll
  for (int c = 0; c < kSkipPrefetch; ++c) sk[c] = 0.f;
  if (inb && p.recon_residual) {
    int x0, x1, y0, y1;
    float lx0, lx1, ly0, ly1;
    if (p.upsample == 1) {
      bil_src(x, p.w_low, 0.25f, x0, x1, lx0, lx1);
      bil_src(y, p.h_low, 0.25f, y0, y1, ly0, ly1);
    } else {
      x0 = x1 = x >> 2;
      y0 = y1 = y >> 2;
      lx0 = ly0 = 1.f;
      lx1 = ly1 = 0.f;
    }
    // all taps of all channels are issued before the first use (no per-channel branch)
    const int o00 = y0 * p.w_low + x0, o01 = y0 * p.w_low + x1, o10 = y1 * p.w_low + x0, o11 = y1 * p.w_low + x1;
    const int nm = p.n_mask > 0 ? p.n_mask : 1;
    float t00[kSkipPrefetch], t01[kSkipPrefetch], t10[kSkipPrefetch], t11[kSkipPrefetch];
#pragma unroll
    for (int c = 0; c < kSkipPrefetch; ++c) {
      const int cc = c < nm ? c : nm - 1;
      const float *sp = p.skip_src + b * p.skip_bstride + static_cast<size_t>(cc) * p.h_low * p.w_low;
      t00[c] = __ldg(sp + o00);
      t01[c] = __ldg(sp + o01);
      t10[c] = __ldg(sp + o10);
      t11[c] = __ldg(sp + o11);
    }
#pragma unroll
    for (int c = 0; c < kSkipPrefetch; ++c) {
      const float r0 = fmaf(lx0, t00[c], lx1 * t01[c]);
      const float r1 = fmaf(lx0, t10[c], lx1 * t11[c]);
      const float v = p.upsample == 1 ? fmaf(ly0, r0, ly1 * r1) : t00[c];
      sk[c] = c < p.n_mask ? v : 0.f;
    }
  }
  mbar_wait_parked(bar_tfull, acc_ph);
  tc_fence_after();
  uint32_t v[16];
  tmem_ld16(taddr, v);
  if constexpr (STACK) {   // columns 16..31 hold A x W_lo^T
    uint32_t v2[16];
    tmem_ld16(taddr + N, v2);
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < 16; ++c) v[c] = __float_as_uint(__uint_as_float(v[c]) + __uint_as_float(v2[c]));
  } else {
    tmem_ld_wait();
  }
  if (inb) {
    const size_t plane = static_cast<size_t>(p.H) * p.W;
    const size_t o0 = static_cast<size_t>(b) * p.cout * plane + static_cast<size_t>(y) * p.W + x;
    float rec[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) {
      rec[c] = 0.f;
      if (c < p.cout) {
        const float raw = __uint_as_float(v[c]) + s_bias[c];
        if (p.out_raw != nullptr) p.out_raw[o0 + c * plane] = raw;
        float r = raw;
        if (p.recon_residual && c < p.n_mask && c < kSkipPrefetch) r = sk[c < kSkipPrefetch ? c : 0] + raw;
        rec[c] = r;
        if (p.out_recon != nullptr) p.out_recon[o0 + c * plane] = r;
      }
    }
    if (p.out_state != nullptr) {
      if (p.state_unshaded && p.cout >= 5) {
        // mainVideoUnshaded.py:460-465: clamp mask, normalise normal (shading.py:193-207), clamp depth / ao
        p.out_state[o0] = fminf(fmaxf(rec[0], -1.f), 1.f);
        float len = sqrtf(rec[1] * rec[1] + rec[2] * rec[2] + rec[3] * rec[3]);
        len = fmaxf(len, 1e-7f);
        p.out_state[o0 + plane] = rec[1] / len;
        p.out_state[o0 + 2 * plane] = rec[2] / len;
        p.out_state[o0 + 3 * plane] = rec[3] / len;
#pragma unroll
        for (int c = 4; c < 16; ++c)
          if (c < p.cout) p.out_state[o0 + c * plane] = fminf(fmaxf(rec[c], 0.f), 1.f);
      } else {
#pragma unroll
        for (int c = 0; c < 16; ++c)
          if (c < p.cout) p.out_state[o0 + c * plane] = fminf(fmaxf(rec[c], 0.f), 1.f);
      }
    }
  }
}

// STACK = 1 (split precision, N = 64 only): the hi and lo weight planes of a tap sit back to back in shared memory and
// are ONE N=128 B operand, so A_hi*W_hi^T and A_hi*W_lo^T are a single M128xN128 MMA (67.7 cycles measured) instead of two
// N=64 MMAs (2 x 53.9, profiles/r01_probe_mma_cycles.json); the epilogue adds accumulator columns c and 64+c.
template <int N, int EPI, int STACK>
__global__ void __launch_bounds__(kConvThreads, 1)
conv3x3_umma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                    const ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr uint32_t kWTapBytes = N * 128;               // one (part, tap) weight tile
  const uint32_t w_bytes = p.n_w_parts * 9 * kWTapBytes;
  const uint32_t s_w = smem_base;
  const uint32_t s_stage0 = s_w + ((w_bytes + 1023u) & ~1023u);
  const uint32_t s_bar = s_stage0 + p.num_stages * p.stage_bytes;
  // barrier block: full[8], empty[8], wbar, tmem_full[2], tmem_empty[2], tmem_ptr
  const uint32_t b_full = s_bar, b_empty = s_bar + 64, b_w = s_bar + 128, b_tfull = s_bar + 136,
                 b_tempty = s_bar + 152, s_tptr = s_bar + 168;
  __shared__ float s_bias[64];

  if (threadIdx.x < 64) s_bias[threadIdx.x] = (p.bias != nullptr && threadIdx.x < N) ? p.bias[threadIdx.x] : 0.f;

  if (warp == 8 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmW);
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(b_full + 8 * s, 1);
      mbar_init(b_empty + 8 * s, 1);
    }
    mbar_init(b_w, 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(b_tfull + 8 * a, 1);
      mbar_init(b_tempty + 8 * a, 4);
    }
    fence_barrier_init();
  }
  constexpr uint32_t kAccCols = STACK ? 2 * N : N;   // TMEM columns of one accumulator buffer
  constexpr uint32_t kTmemCols = (2 * kAccCols < 32) ? 32 : 2 * kAccCols;
  if (warp == 9) tmem_alloc(s_tptr, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(s_tptr));

  const int my_tiles = (p.num_tiles - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) /
                       static_cast<int>(gridDim.x);

  if (warp == 8) {
    // ===================== TMA producer (whole warp runs the loop; one elected lane issues) =====================
    if (elect_one()) {
      mbar_arrive_expect_tx(b_w, w_bytes);
      for (int wp = 0; wp < p.n_w_parts; ++wp)
        for (int tap = 0; tap < 9; ++tap)
          tma_load_2d(s_w + (STACK ? (tap * 2 + wp) : (wp * 9 + tap)) * kWTapBytes, &tmW, b_w, 0,
                      p.w_row0 + (wp * 9 + tap) * p.w_row_stride);
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
          mbar_arrive_expect_tx(b_full + 8 * s, kPitch * kHaloH * 128);
          tma_load_4d(s_stage0 + s * p.stage_bytes, &tmA, b_full + 8 * s, ap * 64, tx * kTileW - 1, ty * kTileH - 1, b);
        }
        __syncwarp();
      }
    }
  } else if (warp == 9) {
    // ===================== MMA issuer (whole warp runs the loop; one elected lane issues) =====================
    const uint32_t idesc = make_idesc_f16(128, N, p.is_bf16);
    const uint32_t idesc2 = make_idesc_f16(128, 2 * N, p.is_bf16);   // STACK: [W_hi ; W_lo] as one N=128 operand
    const uint64_t b_desc_base = make_smem_desc_sw128(s_w, 1024, 0);
    mbar_wait(b_w, 0);
    uint32_t it = 0;
    for (int i = 0; i < my_tiles; ++i) {
      const uint32_t acc = i & 1;
      const uint32_t acc_ph = (i >> 1) & 1;
      mbar_wait(b_tempty + 8 * acc, acc_ph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * kAccCols;
      uint32_t first = 1;
      for (int ap = 0; ap < p.n_act_parts; ++ap, ++it) {
        const uint32_t s = it % p.num_stages;
        const uint32_t ph = (it / p.num_stages) & 1;
        mbar_wait(b_full + 8 * s, ph);
        tc_fence_after();
        const uint32_t a_base = s_stage0 + s * p.stage_bytes;
        const uint64_t a_desc0 = make_smem_desc_sw128(a_base, kPitch * 128, 0);
        if (elect_one()) {
          if constexpr (STACK) {
            // pass 0: A_hi x [W_hi ; W_lo] (N=128, zero-initialises all 128 columns); pass 1: A_lo x W_hi (N=64)
#pragma unroll
            for (int tap = 0; tap < 9; ++tap) {
              const uint64_t adesc = a_desc0 + static_cast<uint64_t>((((tap / 3) * kPitch + (tap % 3)) * 128) >> 4);
              const uint64_t bdesc = b_desc_base + static_cast<uint64_t>((tap * 2 * kWTapBytes) >> 4);
              umma_f16_k64(d_tmem, adesc, bdesc, ap == 0 ? idesc2 : idesc, (tap == 0) ? (first ? 0u : 1u) : 1u);
            }
          } else {
            const int nwp = (ap == 0) ? p.n_w_parts : 1;  // (hi,hi) (hi,lo) | (lo,hi)
            for (int wp = 0; wp < nwp; ++wp) {
              const uint64_t b_desc0 = b_desc_base + static_cast<uint64_t>((wp * 9 * kWTapBytes) >> 4);
#pragma unroll
              for (int tap = 0; tap < 9; ++tap) {
                // tap (ky,kx): the same halo tile, start address shifted by (ky*pitch + kx) pixel rows of 128 B
                const uint64_t adesc = a_desc0 + static_cast<uint64_t>((((tap / 3) * kPitch + (tap % 3)) * 128) >> 4);
                const uint64_t bdesc = b_desc0 + static_cast<uint64_t>((tap * kWTapBytes) >> 4);
                umma_f16_k64(d_tmem, adesc, bdesc, idesc, (tap == 0) ? (first ? 0u : 1u) : 1u);
              }
              first = 0;
            }
          }
          umma_commit(b_empty + 8 * s);  // frees this halo stage when its MMAs retire
          if (ap == p.n_act_parts - 1) umma_commit(b_tfull + 8 * acc);  // accumulator complete -> epilogue
        }
        __syncwarp();
        first = 0;
      }
    }
  } else {
    // ===================== epilogue warps 0..7: group g = warp/4 drains accumulator buffer g =====================
    const int grp = warp >> 2, q = warp & 3;
    for (int i = grp; i < my_tiles; i += 2) {
      const int t = blockIdx.x + i * gridDim.x;
      const int tx = t % p.tiles_x;
      const int ty = (t / p.tiles_x) % p.tiles_y;
      const int b = t / (p.tiles_x * p.tiles_y);
      const uint32_t acc = grp;
      const uint32_t acc_ph = (i >> 1) & 1;
      const uint32_t taddr = tmem_base + acc * kAccCols + (static_cast<uint32_t>(q * 32) << 16);
      if constexpr (EPI == EPI_STORE16) {
        epilogue_store16<N, STACK>(p, s_bias, taddr, q, lane, b, ty, tx, true, b_tfull + 8 * acc, acc_ph);
      } else {
        epilogue_final<N, STACK>(p, s_bias, taddr, q, lane, b, ty, tx, true, b_tfull + 8 * acc, acc_ph);
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(b_tempty + 8 * acc);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 9) tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace isr
