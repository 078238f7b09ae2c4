// probe_store.cu -- how fast can 148 persistent CTAs write fp32 NCHW planes in the output conv's pattern?
// Each "tile" = SEG px x ROWS rows of each of NPL planes (2048 x 2048 x B images), one thread per pixel, 4-byte streaming
// stores, tiles strided over the CTAs like the conv kernels do.  Reports ms and GB/s for several segment widths.
//   nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o tools/probe_store tools/probe_store.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int SEG, int ROWS, int VEC>
__global__ void __launch_bounds__(512, 1) store_kernel(float *out, int B, int H, int W, int npl) {
  const int tiles_x = W / SEG, tiles_y = (H + ROWS - 1) / ROWS;
  const int num_tiles = tiles_x * tiles_y * B;
  const size_t plane = (size_t)H * W;
  constexpr int PX = SEG * ROWS / VEC;   // threads with work per tile
  for (int t = blockIdx.x * 4 + (threadIdx.x >> 7); t < num_tiles; t += gridDim.x * 4) {   // 4 groups of 128 threads
    const int r = threadIdx.x & 127;
    if (r >= PX) continue;
    const int tx = t % tiles_x, ty = (t / tiles_x) % tiles_y, b = t / (tiles_x * tiles_y);
    const int py = (r * VEC) / SEG, px = (r * VEC) % SEG;
    const int y = ty * ROWS + py, x = tx * SEG + px;
    if (y >= H) continue;
    float *o = out + (size_t)b * npl * plane + (size_t)y * W + x;
    for (int c = 0; c < npl; ++c) {
      if (VEC == 1) __stcs(o + c * plane, (float)c);
      else if (VEC == 4) __stcs(reinterpret_cast<float4 *>(o + c * plane), make_float4(c, c, c, c));
    }
  }
}
template <int SEG, int ROWS, int VEC>
void run(const char *name, float *buf, int B, int H, int W, int npl) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  store_kernel<SEG, ROWS, VEC><<<148, 512>>>(buf, B, H, W, npl);
  cudaEventRecord(e0);
  for (int i = 0; i < 10; ++i) store_kernel<SEG, ROWS, VEC><<<148, 512>>>(buf, B, H, W, npl);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 10;
  printf("%-40s %.3f ms  %.0f GB/s  (%s)\n", name, ms, (double)B * npl * H * W * 4 / ms / 1e6, cudaGetErrorString(cudaGetLastError()));
}
int main() {
  const int B = 4, H = 2048, W = 2048, npl = 12;
  float *buf; cudaMalloc(&buf, (size_t)B * npl * H * W * 4);
  run<16, 8, 1>("16 px x 8 rows, 4-byte stores (64 B)", buf, B, H, W, npl);
  run<16, 7, 1>("16 px x 7 rows, 4-byte stores (64 B)", buf, B, H, W, npl);
  run<32, 4, 1>("32 px x 4 rows, 4-byte stores (128 B)", buf, B, H, W, npl);
  run<64, 2, 1>("64 px x 2 rows, 4-byte stores (256 B)", buf, B, H, W, npl);
  run<128, 1, 1>("128 px x 1 row, 4-byte stores (512 B)", buf, B, H, W, npl);
  run<16, 8, 4>("16 px x 8 rows, 16-byte stores (64 B)", buf, B, H, W, npl);
  run<64, 8, 4>("64 px x 8 rows, 16-byte stores (256 B)", buf, B, H, W, npl);
  run<128, 4, 4>("128 px x 4 rows, 16-byte stores (512 B)", buf, B, H, W, npl);
  run<16, 8, 1>("16 px x 8 rows, 6 planes", buf, B, H, W, 6);
  return 0;
}
