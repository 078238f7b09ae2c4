// probe_tma.cu -- per-SM throughput of TMA tensor loads as a function of the box shape, for an NHWC 16-bit activation
// tensor [B][H][W][64] (one pixel = 128 B).  Persistent kernel, one CTA per SM, 8-stage ring, a consumer warp that only
// recycles the stages: what is measured is the load path alone (the conv kernel's "loads_only" experiment showed 13 B/clk/SM).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I isosurfacesuperresolution_b200/csrc tools/probe_tma.cu -o tools/probe_tma
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>

#include "isr_ptx.cuh"

using namespace isr;

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
  void *p = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  return reinterpret_cast<EncodeTiledFn>(p);
}

struct Shape {
  int rank;           // 2 | 4
  int bx, by;         // box: bx pixels x by rows (rank 4) or bx*by rows (rank 2)
  int step_x, step_y; // tile stride in pixels
  int halo;           // start offset -halo
  int split;          // number of TMA instructions the box is split into (by rows)
};

__global__ void __launch_bounds__(64, 1)
tma_probe(const __grid_constant__ CUtensorMap tm, Shape sh, int B, int H, int W, int tiles_x, int tiles_y, int num_tiles,
          int stage_bytes, int num_stages, unsigned long long *cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_bar = base + num_stages * stage_bytes;
  const uint32_t b_full = s_bar, b_empty = s_bar + 64;
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    for (int s = 0; s < num_stages; ++s) {
      mbar_init(b_full + 8 * s, 1);
      mbar_init(b_empty + 8 * s, 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  const int my_tiles = (num_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const uint32_t box_bytes = sh.bx * sh.by * 128;
  const long long t0 = clock64();
  if (warp == 0) {
    for (int i = 0; i < my_tiles; ++i) {
      const int t = blockIdx.x + i * gridDim.x;
      const int tx = t % tiles_x, ty = (t / tiles_x) % tiles_y, b = t / (tiles_x * tiles_y);
      const uint32_t s = i % num_stages, ph = (i / num_stages) & 1;
      mbar_wait(b_empty + 8 * s, ph ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(b_full + 8 * s, box_bytes);
        const uint32_t dst = base + s * stage_bytes;
        if (sh.rank == 4) {
          const int rows = sh.by / sh.split;
          for (int k = 0; k < sh.split; ++k)
            tma_load_4d(dst + k * rows * sh.bx * 128, &tm, b_full + 8 * s, 0, tx * sh.step_x - sh.halo,
                        ty * sh.step_y - sh.halo + k * rows, b);
        } else {
          tma_load_2d(dst, &tm, b_full + 8 * s, 0, t * sh.bx * sh.by);
        }
      }
      __syncwarp();
    }
  } else {
    for (int i = 0; i < my_tiles; ++i) {
      const uint32_t s = i % num_stages, ph = (i / num_stages) & 1;
      mbar_wait(b_full + 8 * s, ph);
      if (elect_one()) mbar_arrive(b_empty + 8 * s);
      __syncwarp();
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
}

static void run(const char *name, void *buf, int B, int H, int W, Shape sh, CUtensorMapL2promotion promo, bool last) {
  EncodeTiledFn enc = get_encode();
  CUtensorMap tm;
  CUresult r;
  if (sh.rank == 4) {
    cuuint64_t dims[4] = {64, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
    cuuint64_t strides[3] = {128, (cuuint64_t)W * 128, (cuuint64_t)H * W * 128};
    cuuint32_t box[4] = {64, (cuuint32_t)sh.bx, (cuuint32_t)(sh.by / sh.split), 1};
    cuuint32_t es[4] = {1, 1, 1, 1};
    r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    cuuint64_t dims[2] = {64, (cuuint64_t)B * H * W};
    cuuint64_t strides[1] = {128};
    cuuint32_t box[2] = {64, (cuuint32_t)(sh.bx * sh.by)};
    cuuint32_t es[2] = {1, 1};
    r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) {
    printf("  {\"case\": \"%s\", \"err\": \"encode %d\"}%s\n", name, (int)r, last ? "" : ",");
    return;
  }
  int tiles_x, tiles_y, num_tiles;
  if (sh.rank == 4) {
    tiles_x = W / sh.step_x; tiles_y = H / sh.step_y; num_tiles = tiles_x * tiles_y * B;
  } else {
    num_tiles = (int)((size_t)B * H * W / (sh.bx * sh.by)); tiles_x = num_tiles; tiles_y = 1;
  }
  const int stage_bytes = ((sh.bx * sh.by * 128) + 1023) / 1024 * 1024;
  int num_stages = (200 * 1024) / stage_bytes;
  if (num_stages > 8) num_stages = 8;
  const size_t smem = 1024 + (size_t)num_stages * stage_bytes + 256;
  cudaFuncSetAttribute(tma_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  unsigned long long *d;
  cudaMalloc(&d, 148 * sizeof(unsigned long long));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  tma_probe<<<148, 64, smem>>>(tm, sh, B, H, W, tiles_x, tiles_y, num_tiles, stage_bytes, num_stages, d);
  cudaEventRecord(e0);
  tma_probe<<<148, 64, smem>>>(tm, sh, B, H, W, tiles_x, tiles_y, num_tiles, stage_bytes, num_stages, d);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  unsigned long long h[148], mx = 0;
  cudaMemcpy(h, d, sizeof h, cudaMemcpyDeviceToHost);
  for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  const double bytes = (double)num_tiles * sh.bx * sh.by * 128;
  printf("  {\"case\": \"%s\", \"tensor_MB\": %.0f, \"box_px\": [%d, %d], \"split\": %d, \"stages\": %d, \"ms\": %.4f, "
         "\"loaded_GBps\": %.0f, \"bytes_per_clk_per_sm\": %.2f, \"cycles_per_128B_line\": %.2f, \"err\": \"%s\"}%s\n",
         name, (double)B * H * W * 128 / 1e6, sh.bx, sh.by, sh.split, num_stages, ms, bytes / ms / 1e6,
         bytes / 148.0 / (double)mx, (double)mx / ((double)num_tiles / 148.0 * sh.bx * sh.by), cudaGetErrorString(e),
         last ? "" : ",");
  cudaFree(d);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
}

int main() {
  void *big, *small_;
  const int B = 16, H = 1024, W = 1024;             // 2.1 GB (the 4x stage of config 2)
  cudaMalloc(&big, (size_t)B * H * W * 128);
  cudaMemset(big, 0, (size_t)B * H * W * 128);
  cudaMalloc(&small_, (size_t)2 * 512 * 512 * 128);  // 67 MB: L2-resident
  cudaMemset(small_, 0, (size_t)2 * 512 * 512 * 128);
  const CUtensorMapL2promotion P256 = CU_TENSOR_MAP_L2_PROMOTION_L2_256B, P128 = CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                               PN = CU_TENSOR_MAP_L2_PROMOTION_NONE;
  printf("[\n");
  run("hbm_4d_halo_10x18 (conv kernel)", big, B, H, W, {4, 10, 18, 8, 16, 1, 1}, P256, false);
  run("hbm_4d_halo_10x18_split2", big, B, H, W, {4, 10, 18, 8, 16, 1, 2}, P256, false);
  run("hbm_4d_halo_10x18_split6", big, B, H, W, {4, 10, 18, 8, 16, 1, 6}, P256, false);
  run("hbm_4d_halo_10x18_promo128", big, B, H, W, {4, 10, 18, 8, 16, 1, 1}, P128, false);
  run("hbm_4d_halo_10x18_promoNone", big, B, H, W, {4, 10, 18, 8, 16, 1, 1}, PN, false);
  run("hbm_4d_nohalo_8x16", big, B, H, W, {4, 8, 16, 8, 16, 0, 1}, P256, false);
  run("hbm_4d_wide_18x10", big, B, H, W, {4, 18, 10, 16, 8, 1, 1}, P256, false);
  run("hbm_4d_wide_34x6", big, B, H, W, {4, 34, 6, 32, 4, 1, 1}, P256, false);
  run("hbm_4d_row_128x1", big, B, H, W, {4, 128, 1, 128, 1, 0, 1}, P256, false);
  run("hbm_2d_180rows", big, B, H, W, {2, 180, 1, 0, 0, 0, 1}, P256, false);
  run("hbm_2d_256rows", big, B, H, W, {2, 256, 1, 0, 0, 0, 1}, P256, false);
  run("l2_4d_halo_10x18", small_, 2, 512, 512, {4, 10, 18, 8, 16, 1, 1}, P256, false);
  run("l2_4d_halo_10x18_split2", small_, 2, 512, 512, {4, 10, 18, 8, 16, 1, 2}, P256, false);
  run("l2_4d_wide_18x10", small_, 2, 512, 512, {4, 18, 10, 16, 8, 1, 1}, P256, false);
  run("l2_4d_row_128x1", small_, 2, 512, 512, {4, 128, 1, 128, 1, 0, 1}, P256, false);
  run("l2_2d_180rows", small_, 2, 512, 512, {2, 180, 1, 0, 0, 0, 1}, P256, false);
  run("l2_2d_256rows", small_, 2, 512, 512, {2, 256, 1, 0, 0, 0, 1}, P256, true);
  printf("]\n");
  return 0;
}
