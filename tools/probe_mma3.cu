// probe_mma3.cu -- variant of probe_mma2.cu: what the per-tile tcgen05.commit pair and non-zero operands add to the
// cost of tcgen05.mma.cta_group::2 (M=256 across a CTA pair, K=16, kind::f16, operands in shared
// memory) per instruction, for N = 64 / 128 / 256.  Question: is the ~54-cycle floor of an N<=64 shared-memory MMA
// (probe_mma.cu) per INSTRUCTION (then a CTA pair does twice the work per instruction) or per SM?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I isosurfacesuperresolution_b200/csrc tools/probe_mma2.cu -o tools/probe_mma2
#include <cstdio>
#include <cuda_runtime.h>
#include <cooperative_groups.h>

#include "isr_ptx.cuh"

using namespace isr;
namespace cg = cooperative_groups;

template <int N>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128, 1)
mma2_probe(int iters, unsigned long long *cycles, int flags) {   // flags: 1 = two multicast commits per 36 MMAs, 2 = non-zero operands, 4 = one commit
  extern __shared__ uint8_t smem_raw[];
  cg::cluster_group cluster = cg::this_cluster();
  const uint32_t rank = cluster.block_rank();
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t s_a = base;                         // 24 KB halo-like A region
  const uint32_t s_b = base + 24 * 1024;             // nine taps of N/2 rows (this CTA's half of B)
  const uint32_t b_bytes = 9 * (N / 2) * 128;
  const uint32_t s_bar = s_b + b_bytes;
  const uint32_t s_tptr = s_bar + 40;
  for (uint32_t i = threadIdx.x * 16; i < 24 * 1024 + b_bytes; i += blockDim.x * 16)
    asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};" ::"r"(base + i), "r"((flags & 2) ? (0x3c003800u + (i * 2654435761u & 0x03ff03ffu)) : 0u));
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(s_bar, 1);
    mbar_init(s_bar + 8, 1);
    mbar_init(s_bar + 16, 1);
    mbar_init(s_bar + 24, 1);
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_bar + 24) : "memory");   // phase 0 complete: waits return at once
    fence_barrier_init();
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_tptr), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  fence_proxy_async();
  tc_fence_before();
  cluster.sync();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(s_tptr));
  if (warp == 0 && rank == 0) {
    const uint32_t idesc = make_idesc_f16(256, N, 1);
    const uint64_t a0 = make_smem_desc_sw128(s_a, 1280, 0);
    const uint64_t b0 = make_smem_desc_sw128(s_b, 1024, 0);
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
      if (flags & 8) tc_fence_after();
      if (flags & 64) { mbar_wait(s_bar + 24, 0); tc_fence_after(); }
      if (elect_one()) {
        const uint32_t d = tmem + ((flags & 32) ? (it & 3) * N : ((it & 1) ? N : 0));
#pragma unroll
        for (int tap = 0; tap < 9; ++tap) {
          const uint64_t ad = a0 + (uint64_t)((((tap / 3) * 10 + tap % 3) * 128) >> 4);
          const uint64_t bd = b0 + (uint64_t)((tap * (N / 2) * 128) >> 4);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            asm volatile(
                "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
                "l"(ad + 2 * k), "l"(bd + 2 * k), "r"(idesc), "r"(((flags & 16) && tap == 0 && k == 0) ? 0 : 1)
                : "memory");
        }
        if (flags & 5)
          asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(s_bar + 8),
                       "h"(static_cast<uint16_t>(3)) : "memory");
        if (flags & 1)
          asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(s_bar + 16),
                       "h"(static_cast<uint16_t>(3)) : "memory");
      }
      __syncwarp();
    }
    if (elect_one())
      asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s_bar) : "memory");
    __syncwarp();
    mbar_wait(s_bar, 0);
    if (threadIdx.x == 0) cycles[blockIdx.x / 2] = (unsigned long long)(clock64() - t0);
  }
  tc_fence_before();
  cluster.sync();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

template <int N>
static void run(bool last, int flags) {
  const int iters = 2000;
  unsigned long long *d;
  cudaMalloc(&d, 74 * sizeof(unsigned long long));
  const size_t smem = 1024 + 24 * 1024 + 9 * (N / 2) * 128 + 128;
  cudaFuncSetAttribute(mma2_probe<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int grid : {148}) {
    mma2_probe<N><<<grid, 128, smem>>>(iters, d, flags);
    cudaError_t e = cudaDeviceSynchronize();
    unsigned long long h[74];
    cudaMemcpy(h, d, (grid / 2) * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    unsigned long long mx = 0;
    for (int i = 0; i < grid / 2; ++i) mx = h[i] > mx ? h[i] : mx;
    printf("  {\"case\": \"cta_group2_ss_haloA_nineB\", \"flags\": %d, \"M\": 256, \"N\": %d, \"grid\": %d, \"cycles_per_mma\": %.2f, "
           "\"floor_per_sm_N_over_2\": %d, \"err\": \"%s\"}%s\n",
           flags, N, grid, (double)mx / (iters * 36.0), N / 2, cudaGetErrorString(e), (last && grid == 148) ? "" : ",");
  }
  cudaFree(d);
}

int main() {
  printf("[\n");
  for (int f : {0, 8, 16, 32, 64, 1 + 16 + 32 + 64}) { run<64>(false, f); }
  run<64>(true, 0);
  printf("]\n");
  return 0;
}
