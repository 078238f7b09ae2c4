// isr_up2.cu -- support kernels of the fused "bilinear 2x upsample + conv3x3" layer (enhancenet.py:116-121: nn.Upsample(2,
// 'bilinear') followed by Conv2d), which conv3x3_pair_kernel<.., PH = 4> computes WITHOUT materialising the upsampled
// tensor: for output phase (py, px) = (Y % 2, X % 2) the conv of the upsampled image is a 3x3 conv of the LOW-res image
// with composite weights  W'[py,px][a,b] = sum_{ky,kx} Ry[py][a][ky] * Rx[px][b][kx] * W[ky,kx]  (same FLOPs, no 4x tensor).
//   py = 0:  rows Y-1, Y, Y+1 of the upsampled image are 0.75 x[i-1] + 0.25 x[i],  0.25 x[i-1] + 0.75 x[i],  0.75 x[i] + 0.25 x[i+1]
//   py = 1:                                               0.25 x[i-1] + 0.75 x[i],  0.75 x[i] + 0.25 x[i+1],  0.25 x[i] + 0.75 x[i+1]
// This identity holds away from the image border.  Within 2 pixels of it the upsample clamps its source index and the
// conv zero-pads the UPSAMPLED image, so those pixels are recomputed the plain way on four thin bands (3 rows / columns of
// the upsampled image -> 2 correct output rows / columns each): band_upsample_kernel builds the bands, the ordinary conv
// kernel runs on them as small images, band_scatter_kernel copies the two valid rows / columns over the fused result.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "isr_internal.h"

namespace isr {

// R[p][a][k]: coefficient of low-res tap a (offset a-1) in upsampled row (offset k-1) for output phase p
__constant__ float kUpR[2][3][3] = {{{0.75f, 0.25f, 0.f}, {0.25f, 0.75f, 0.75f}, {0.f, 0.f, 0.25f}},
                                    {{0.25f, 0.f, 0.f}, {0.75f, 0.75f, 0.25f}, {0.f, 0.25f, 0.75f}}};

// fp32 OIHW [64][64][3][3] -> [fmt 2 (bf16, fp16)][phase 4 = py*2+px][tap 9 = a*3+b][n 64 (permuted as pack_weights)][k 64]
__global__ void pack_weights_up2_kernel(const float *__restrict__ w, uint16_t *__restrict__ out) {
  const int per_fmt = 4 * 9 * 64 * 64;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < per_fmt; i += gridDim.x * blockDim.x) {
    const int k = i & 63, n = (i >> 6) & 63, tap = (i >> 12) % 9, ph = i / (9 * 4096);
    const int a = tap / 3, b = tap % 3, py = ph >> 1, px = ph & 1;
    const int co = 16 * ((n & 7) >> 1) + 2 * (n >> 3) + (n & 1);
    const float *wp = w + (static_cast<size_t>(co) * 64 + k) * 9;
    float acc = 0.f;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky)
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) acc += (kUpR[py][a][ky] * kUpR[px][b][kx]) * wp[ky * 3 + kx];
    const __nv_bfloat16 bh = __float2bfloat16_rn(acc);
    const __half hh = __float2half_rn(acc);
    out[i] = *reinterpret_cast<const uint16_t *>(&bh);
    out[per_fmt + i] = *reinterpret_cast<const uint16_t *>(&hh);
  }
}

__device__ __forceinline__ float ld16(const uint16_t *p, int is_bf16) {
  const uint16_t u = *p;
  return is_bf16 ? __bfloat162float(*reinterpret_cast<const __nv_bfloat16 *>(&u))
                 : __half2float(*reinterpret_cast<const __half *>(&u));
}
__device__ __forceinline__ uint16_t st16(float v, int is_bf16) {
  if (is_bf16) {
    const __nv_bfloat16 b = __float2bfloat16_rn(v);
    return *reinterpret_cast<const uint16_t *>(&b);
  }
  const __half h = __float2half_rn(v);
  return *reinterpret_cast<const uint16_t *>(&h);
}
__device__ __forceinline__ void up_src(int dst, int n_in, int &i0, int &i1, float &l0, float &l1) {
  float src = 0.5f * (static_cast<float>(dst) + 0.5f) - 0.5f;   // torch upsample_bilinear2d, align_corners=False
  src = src < 0.f ? 0.f : src;
  int a = static_cast<int>(src);
  a = a > n_in - 1 ? n_in - 1 : a;
  i0 = a;
  i1 = a < n_in - 1 ? a + 1 : a;
  l1 = src - static_cast<float>(a);
  l0 = 1.f - l1;
}

// Bands of the 2x-upsampled image (x first, then y, rounded once to 16 bit like upsample2x_nhwc_kernel):
//   rows: [B][2][3][W2][64]  (side 0: rows 0..2, side 1: rows H2-3..H2-1);  cols: [B][2][H2][3][64] (cols 0..2 / W2-3..W2-1)
// in: [B][hl][wl][64 * in_parts] 16-bit (first plane used)
__global__ void band_upsample_kernel(const uint16_t *__restrict__ in, uint16_t *__restrict__ rows, uint16_t *__restrict__ cols,
                                     int B, int hl, int wl, int in_parts, int is_bf16) {
  pdl_prologue();
  const int H2 = 2 * hl, W2 = 2 * wl;
  const size_t n_row = static_cast<size_t>(B) * 2 * 3 * W2, n_col = static_cast<size_t>(B) * 2 * H2 * 3;
  const size_t total = (n_row + n_col) * 8;   // one thread = 8 channels of one band pixel
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int c8 = static_cast<int>(i & 7);
    size_t p = i >> 3;
    int b, Y, X;
    uint16_t *dst;
    if (p < n_row) {
      X = static_cast<int>(p % W2);
      const int r = static_cast<int>((p / W2) % 3), side = static_cast<int>((p / (static_cast<size_t>(3) * W2)) % 2);
      b = static_cast<int>(p / (static_cast<size_t>(6) * W2));
      Y = side == 0 ? r : H2 - 3 + r;
      dst = rows + p * 64 + c8 * 8;
    } else {
      p -= n_row;
      const int c = static_cast<int>(p % 3);
      Y = static_cast<int>((p / 3) % H2);
      const int side = static_cast<int>((p / (static_cast<size_t>(3) * H2)) % 2);
      b = static_cast<int>(p / (static_cast<size_t>(6) * H2));
      X = side == 0 ? c : W2 - 3 + c;
      dst = cols + p * 64 + c8 * 8;
    }
    int x0, x1, y0, y1;
    float lx0, lx1, ly0, ly1;
    up_src(X, wl, x0, x1, lx0, lx1);
    up_src(Y, hl, y0, y1, ly0, ly1);
    const size_t cs = static_cast<size_t>(64) * in_parts;
    const uint16_t *base = in + static_cast<size_t>(b) * hl * wl * cs + c8 * 8;
    const uint16_t *p00 = base + (static_cast<size_t>(y0) * wl + x0) * cs, *p01 = base + (static_cast<size_t>(y0) * wl + x1) * cs;
    const uint16_t *p10 = base + (static_cast<size_t>(y1) * wl + x0) * cs, *p11 = base + (static_cast<size_t>(y1) * wl + x1) * cs;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float r0 = lx0 * ld16(p00 + j, is_bf16) + lx1 * ld16(p01 + j, is_bf16);
      const float r1 = lx0 * ld16(p10 + j, is_bf16) + lx1 * ld16(p11 + j, is_bf16);
      dst[j] = st16(ly0 * r0 + ly1 * r1, is_bf16);
    }
  }
}

// copies the two valid rows / columns of each convolved band over the fused result out[B][H2][W2][64 * out_parts]
__global__ void band_scatter_kernel(const uint16_t *__restrict__ rows, const uint16_t *__restrict__ cols,
                                    uint16_t *__restrict__ out, int B, int H2, int W2, int parts) {
  pdl_prologue();
  const size_t cs = static_cast<size_t>(64) * parts;     // 16-bit elements per pixel, same in bands and output
  const size_t n_row = static_cast<size_t>(B) * 2 * 2 * W2, n_col = static_cast<size_t>(B) * 2 * H2 * 2;
  const size_t chunks = cs / 8;
  const size_t total = (n_row + n_col) * chunks;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t ch = i % chunks;
    size_t p = i / chunks;
    const uint16_t *src;
    int b, Y, X;
    if (p < n_row) {
      X = static_cast<int>(p % W2);
      const int r = static_cast<int>((p / W2) % 2), side = static_cast<int>((p / (static_cast<size_t>(2) * W2)) % 2);
      b = static_cast<int>(p / (static_cast<size_t>(4) * W2));
      const int br = side == 0 ? r : 1 + r;              // valid band rows: 0,1 (top) / 1,2 (bottom)
      Y = side == 0 ? r : H2 - 2 + r;
      src = rows + (((static_cast<size_t>(b) * 2 + side) * 3 + br) * W2 + X) * cs;
    } else {
      p -= n_row;
      const int c = static_cast<int>(p % 2);
      Y = static_cast<int>((p / 2) % H2);
      const int side = static_cast<int>((p / (static_cast<size_t>(2) * H2)) % 2);
      b = static_cast<int>(p / (static_cast<size_t>(4) * H2));
      const int bc = side == 0 ? c : 1 + c;
      X = side == 0 ? c : W2 - 2 + c;
      src = cols + (((static_cast<size_t>(b) * 2 + side) * H2 + Y) * 3 + bc) * cs;
    }
    uint16_t *dst = out + ((static_cast<size_t>(b) * H2 + Y) * W2 + X) * cs;
    reinterpret_cast<uint4 *>(dst)[ch] = reinterpret_cast<const uint4 *>(src)[ch];
  }
}

static inline int grid_up(size_t total) {
  size_t g = (total + 255) / 256;
  const size_t cap = 148 * 8;
  return static_cast<int>(g < cap ? (g == 0 ? 1 : g) : cap);
}

cudaError_t launch_pack_weights_up2(const float *w, uint16_t *out, cudaStream_t s) {
  pack_weights_up2_kernel<<<(4 * 9 * 4096 + 255) / 256, 256, 0, s>>>(w, out);
  count_launch();
  return cudaGetLastError();
}
cudaError_t launch_band_upsample(const uint16_t *in, uint16_t *rows, uint16_t *cols, int B, int hl, int wl, int in_parts,
                                 int is_bf16, cudaStream_t s) {
  const size_t total = (static_cast<size_t>(B) * 6 * 2 * wl + static_cast<size_t>(B) * 6 * 2 * hl) * 8;
  launch_pdl(band_upsample_kernel, dim3(grid_up(total)), dim3(256), 0, s, in, rows, cols, B, hl, wl, in_parts, is_bf16);
  count_launch();
  return cudaGetLastError();
}
cudaError_t launch_band_scatter(const uint16_t *rows, const uint16_t *cols, uint16_t *out, int B, int H2, int W2, int parts,
                                cudaStream_t s) {
  const size_t total = (static_cast<size_t>(B) * 4 * W2 + static_cast<size_t>(B) * 4 * H2) * (8 * parts);
  launch_pdl(band_scatter_kernel, dim3(grid_up(total)), dim3(256), 0, s, rows, cols, out, B, H2, W2, parts);
  count_launch();
  return cudaGetLastError();
}

}  // namespace isr
