// probe_mma.cu -- micro-benchmark: sustained cost of one tcgen05.mma (M=128, K=16, kind::f16, operands in shared
// memory) as a function of N and of the A-operand addressing used by the conv kernel (pitch-10 halo windows).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I isosurfacesuperresolution_b200/csrc \
//        tools/probe_mma.cu -o tools/probe_mma     Run on a B200: tools/probe_mma > gpurun_out/probe_mma.json
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "isr_ptx.cuh"

using namespace isr;

// a_mode 0: one canonical 128-row tile (SBO 1024), same for all MMAs
// a_mode 1: conv-style: 9 tap windows into a pitch-10 halo tile (SBO 1280, start shifted by (ky*10+kx)*128 B)
// a_mode 2: A operand in TMEM (columns 256..), B in shared memory
// b_mode 0: one B tile reused; 1: nine B tiles (tap * N * 128 B apart)
template <int N, int M = 128, int ISSUERS = 1>
__global__ void __launch_bounds__(128, 1) mma_probe(int iters, int a_mode, int b_mode, int two_acc,
                                                    unsigned long long *cycles) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_a = base;                       // 24 KB halo
  const uint32_t s_b = base + 24 * 1024;           // up to 9 * N * 128 B
  const uint32_t nb = b_mode == 1 ? 9 : 1;
  const uint32_t s_bar = s_b + nb * N * 128;
  const uint32_t s_tptr = s_bar + 32;
  for (uint32_t i = threadIdx.x * 16; i < 24 * 1024 + nb * N * 128; i += blockDim.x * 16)
    asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + i), "r"(0));
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(s_bar, 1);
    mbar_init(s_bar + 8, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(s_tptr, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(s_tptr));
  if (warp < ISSUERS) {
    const uint32_t idesc = make_idesc_f16(M, N, 1);
    const uint64_t a0 = make_smem_desc_sw128(s_a, a_mode == 1 ? 1280 : 1024, 0);
    const uint64_t b0 = make_smem_desc_sw128(s_b, 1024, 0);
    {
      const long long t0 = clock64();
      for (int it = 0; it < iters; ++it) {
       if (elect_one()) {
        const uint32_t d = tmem + ((two_acc && (it & 1)) ? N : 0) + (warp ? 256 : 0);
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const uint64_t ad = a_mode == 1 ? a0 + (uint64_t)((((tap / 3) * 10 + tap % 3) * 128) >> 4) : a0;
          const uint64_t bd = b_mode == 1 ? b0 + (uint64_t)((tap * N * 128) >> 4) : b0;
          if (a_mode == 2) {
            const uint32_t at = tmem + 256;
#pragma unroll
            for (int k = 0; k < 4; ++k)
              asm volatile(
                  "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                  "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
                  "r"(at + 8 * k), "l"(bd + 2 * k), "r"(idesc), "r"(1)
                  : "memory");
          } else if (two_acc == 2) {   // alternate accumulators per tap (4 dependent MMAs in a row)
            umma_f16_k64(tmem + (tap & 1) * N, ad, bd, idesc, 1);
          } else if (two_acc == 3) {   // alternate accumulators per MMA
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16(tmem + (k & 1) * N, ad + 2 * k, bd + 2 * k, idesc, 1);
          } else if (two_acc == 4) {   // four accumulators round robin
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_f16(tmem + k * N, ad + 2 * k, bd + 2 * k, idesc, 1);
          } else {
            umma_f16_k64(d, ad, bd, idesc, 1);
          }
        }
       }
       __syncwarp();
      }
      if (elect_one()) umma_commit(s_bar + 8 * warp);
      __syncwarp();
      mbar_wait(s_bar + 8 * warp, 0);
      if (threadIdx.x == 0) cycles[blockIdx.x] = (unsigned long long)(clock64() - t0);
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 512);
}

template <int N, int M = 128, int ISSUERS = 1>
static void run(const char *name, int a_mode, int b_mode, int two_acc, bool last) {
  const int iters = 2000;
  unsigned long long *d;
  cudaMalloc(&d, 148 * sizeof(unsigned long long));
  const size_t smem = 1024 + 24 * 1024 + (b_mode == 1 ? 9 : 1) * N * 128 + 128;
  cudaFuncSetAttribute(mma_probe<N, M, ISSUERS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int grid : {1, 148}) {
    mma_probe<N, M, ISSUERS><<<grid, 128, smem>>>(iters, a_mode, b_mode, two_acc, d);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[148];
    cudaMemcpy(h, d, grid * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    unsigned long long mx = 0;
    for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("  {\"case\": \"%s\", \"N\": %d, \"grid\": %d, \"cycles_per_mma\": %.2f, \"floor_N_over_2\": %d, \"err\": \"%s\"}%s\n",
           name, N, grid, (double)mx / (iters * 36.0), N / 2, cudaGetErrorString(e), (last && grid == 148) ? "" : ",");
  }
  cudaFree(d);
}

int main() {
  printf("[\n");
  run<64>("ss_canonicalA_oneB", 0, 0, 0, false);
  run<64>("ss_haloA_nineB (conv kernel)", 1, 1, 1, false);
  run<64>("ss_haloA_oneB", 1, 0, 0, false);
  run<128>("ss_canonicalA_oneB", 0, 0, 0, false);
  run<128>("ss_haloA_nineB", 1, 1, 1, false);
  run<256>("ss_canonicalA_oneB", 0, 0, 0, false);
  run<32>("ss_haloA_nineB", 1, 1, 1, false);
  run<16>("ss_haloA_nineB", 1, 1, 1, false);
  run<64>("tmemA_nineB", 2, 1, 1, false);
  run<128>("tmemA_nineB", 2, 1, 1, false);
  run<64, 64, 1>("M64_ss_haloA_nineB", 1, 1, 1, false);
  run<128, 64, 1>("M64_ss_haloA_nineB", 1, 1, 1, false);
  run<64, 128, 2>("two_issuer_warps_ss_haloA_nineB (cycles per MMA of ONE warp)", 1, 1, 1, false);
  run<16, 128, 2>("two_issuer_warps_ss_haloA_nineB (cycles per MMA of ONE warp)", 1, 1, 1, false);
  run<64>("alt_acc_per_tap", 1, 1, 2, false);
  run<64>("alt_acc_per_mma", 1, 1, 3, false);
  run<64>("four_acc_round_robin", 1, 1, 4, false);
  run<16>("alt_acc_per_mma", 1, 1, 3, false);
  run<32>("four_acc_round_robin", 1, 1, 4, false);
  run<128>("alt_acc_per_mma", 1, 1, 3, true);
  printf("]\n");
  return 0;
}
