// probe_mma4.cu -- can the LOW activation plane of the split-precision trunk run as an 8-bit MMA?
//   (1) kinds mix: D = A16 x B16^T (tcgen05.mma kind::f16, K = 64) and then D += A8 x B8^T (kind::f8f6f4, A e4m3, B e5m2,
//       K = 64) into the SAME fp32 TMEM accumulator, against a host reference of the same quantised operands;
//   (2) addressing: the 8-bit A operand read as a tap-shifted window of a halo tile (pixel pitch 10, 64-byte rows,
//       SWIZZLE_64B, start address + (ky*10 + kx) * 64 B, 8-row-group stride 640 B), the way isr_conv_pair.cuh reads its
//       16-bit halo with SWIZZLE_128B;
//   (3) cost: cycles per back-to-back f8f6f4 K32 MMA (M = 128, N = 64) against the f16 K16 MMA.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I isosurfacesuperresolution_b200/csrc tools/probe_mma4.cu -o tools/probe_mma4
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "isr_ptx.cuh"

using namespace isr;

__device__ __forceinline__ uint64_t make_smem_desc_sw64(uint32_t addr, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>(1) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(4) << 61;   // SWIZZLE_64B
  return d;
}
// kind::f8f6f4 instruction descriptor: c F32 (bit 4), a_format at bit 7, b_format at bit 10 (E4M3 = 0, E5M2 = 1)
__host__ __device__ constexpr uint32_t make_idesc_f8(int m, int n, int a_fmt, int b_fmt) {
  return (1u << 4) | (static_cast<uint32_t>(a_fmt) << 7) | (static_cast<uint32_t>(b_fmt) << 10) |
         (static_cast<uint32_t>(n >> 3) << 17) | (static_cast<uint32_t>(m >> 4) << 24);
}
__device__ __forceinline__ void umma_f8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}

constexpr int kPitch = 10, kRows = 18, kHalo = kPitch * kRows;   // 180 halo pixels

// a16: 128 x 64 fp16 (row-major), b16: 64 x 64 fp16, a8: kHalo x 64 bytes (e4m3), b8: 64 x 64 bytes (e5m2)
// out: 128 x 64 fp32.  mode 0: f16 only; 1: f16 then f8 accumulate; 2: f8 only.  shift = ky*10+kx window start.
__global__ void __launch_bounds__(128, 1)
probe(const uint16_t *a16, const uint16_t *b16, const uint8_t *a8, const uint8_t *b8, float *out, int mode, int shift,
      int iters, unsigned long long *cycles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *sm = smem_raw + ((1024 - (smem_u32(smem_raw) & 1023)) & 1023);
  const uint32_t base = smem_u32(sm);
  const uint32_t s_a16 = base, s_b16 = base + 16384, s_a8 = base + 24576, s_b8 = s_a8 + 12288, s_bar = s_b8 + 4096,
                 s_tptr = s_bar + 16;
  // fill shared memory with the hardware swizzles applied to the ABSOLUTE address (what TMA does)
  for (int i = threadIdx.x; i < 128 * 8; i += 128) {   // A16: row r, 16-byte chunk c
    const int r = i >> 3, c = i & 7;
    const uint32_t a = s_a16 + r * 128 + c * 16, pa = a ^ (((a >> 7) & 7) << 4);
    *reinterpret_cast<uint4 *>(sm + (pa - base)) = *reinterpret_cast<const uint4 *>(a16 + r * 64 + c * 8);
  }
  for (int i = threadIdx.x; i < 64 * 8; i += 128) {
    const int r = i >> 3, c = i & 7;
    const uint32_t a = s_b16 + r * 128 + c * 16, pa = a ^ (((a >> 7) & 7) << 4);
    *reinterpret_cast<uint4 *>(sm + (pa - base)) = *reinterpret_cast<const uint4 *>(b16 + r * 64 + c * 8);
  }
  for (int i = threadIdx.x; i < kHalo * 4; i += 128) {   // A8 halo: pixel q, 16-byte chunk c of its 64-byte row
    const int q = i >> 2, c = i & 3;
    const uint32_t a = s_a8 + q * 64 + c * 16, pa = a ^ (((a >> 7) & 3) << 4);
    *reinterpret_cast<uint4 *>(sm + (pa - base)) = *reinterpret_cast<const uint4 *>(a8 + q * 64 + c * 16);
  }
  for (int i = threadIdx.x; i < 64 * 4; i += 128) {
    const int r = i >> 2, c = i & 3;
    const uint32_t a = s_b8 + r * 64 + c * 16, pa = a ^ (((a >> 7) & 3) << 4);
    *reinterpret_cast<uint4 *>(sm + (pa - base)) = *reinterpret_cast<const uint4 *>(b8 + r * 64 + c * 16);
  }
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    mbar_init(s_bar, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc(s_tptr, 64);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem;
  asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem) : "r"(s_tptr));
  if (warp == 0) {
    const uint32_t idesc16 = make_idesc_f16(128, 64, 0);
    const uint32_t idesc8 = make_idesc_f8(128, 64, 0, 1);
    const uint64_t a16d = make_smem_desc_sw128(s_a16, 1024, 0), b16d = make_smem_desc_sw128(s_b16, 1024, 0);
    const uint64_t a8d = make_smem_desc_sw64(s_a8 + shift * 64, kPitch * 64), b8d = make_smem_desc_sw64(s_b8, 512);
    if (elect_one()) {
      const long long t0 = clock64();
      for (int it = 0; it < iters; ++it) {
        uint32_t acc = 0;
        if (mode != 2) {
          umma_f16_k64(tmem, a16d, b16d, idesc16, 0);
          acc = 1;
        }
        if (mode != 0) {
          umma_f8(tmem, a8d, b8d, idesc8, acc);            // K = 32: bytes 0..31 of every row
          umma_f8(tmem, a8d + 2, b8d + 2, idesc8, 1);      // bytes 32..63 (+32 B = +2 in the start-address field)
        }
      }
      umma_commit(s_bar);
      if (cycles != nullptr) {
        mbar_wait(s_bar, 0);
        cycles[0] = static_cast<unsigned long long>(clock64() - t0);
      }
    }
    __syncwarp();
  }
  mbar_wait(s_bar, 0);
  tc_fence_after();
  for (int c0 = 0; c0 < 64; c0 += 16) {
    uint32_t v[16];
    tmem_ld16(tmem + c0 + (static_cast<uint32_t>(warp * 32) << 16), v);
    tmem_ld_wait();
    for (int j = 0; j < 16; ++j) out[(warp * 32 + (threadIdx.x & 31)) * 64 + c0 + j] = __uint_as_float(v[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, 64);
}

static float dec_e4m3(uint8_t b) {
  const int s = b >> 7, e = (b >> 3) & 15, m = b & 7;
  float v = e == 0 ? std::ldexp(static_cast<float>(m), -9) : std::ldexp(1.f + m / 8.f, e - 7);
  return s ? -v : v;
}
static float dec_e5m2(uint8_t b) {
  const int s = b >> 7, e = (b >> 2) & 31, m = b & 3;
  float v = e == 0 ? std::ldexp(static_cast<float>(m), -16) : std::ldexp(1.f + m / 4.f, e - 15);
  return s ? -v : v;
}

int main() {
  std::vector<uint16_t> a16(128 * 64), b16(64 * 64);
  std::vector<uint8_t> a8(kHalo * 64), b8(64 * 64);
  std::vector<float> fa16(128 * 64), fb16(64 * 64);
  srand(7);
  auto rnd = [] { return (rand() / (float)RAND_MAX) * 2.f - 1.f; };
  for (size_t i = 0; i < a16.size(); ++i) { __half h = __float2half(rnd()); a16[i] = *reinterpret_cast<uint16_t *>(&h); fa16[i] = __half2float(h); }
  for (size_t i = 0; i < b16.size(); ++i) { __half h = __float2half(rnd() * 0.1f); b16[i] = *reinterpret_cast<uint16_t *>(&h); fb16[i] = __half2float(h); }
  for (auto &v : a8) { do { v = rand() & 0xFF; } while ((v & 0x7F) == 0x7F || ((v >> 3) & 15) > 9); }     // e4m3, |x| <= 7.5, no NaN
  for (auto &v : b8) { do { v = rand() & 0xFF; } while (((v >> 2) & 31) > 16 || ((v >> 2) & 31) < 6); }       // e5m2, 2^-9 .. 3.5
  uint16_t *da16, *db16; uint8_t *da8, *db8; float *dout; unsigned long long *dcyc;
  cudaMalloc(&da16, a16.size() * 2); cudaMalloc(&db16, b16.size() * 2); cudaMalloc(&da8, a8.size()); cudaMalloc(&db8, b8.size());
  cudaMalloc(&dout, 128 * 64 * 4); cudaMalloc(&dcyc, 8);
  cudaMemcpy(da16, a16.data(), a16.size() * 2, cudaMemcpyHostToDevice); cudaMemcpy(db16, b16.data(), b16.size() * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(da8, a8.data(), a8.size(), cudaMemcpyHostToDevice); cudaMemcpy(db8, b8.data(), b8.size(), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  std::vector<float> out(128 * 64);
  int bad_total = 0;
  for (int mode = 0; mode < 3; ++mode)
    for (int shift : {0, 1, 2, 10, 11, 12, 20, 21, 22}) {
      if (mode == 0 && shift != 0) continue;
      probe<<<1, 128, 64 * 1024>>>(da16, db16, da8, db8, dout, mode, shift, 1, nullptr);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("mode %d shift %d: CUDA error %s\n", mode, shift, cudaGetErrorString(e)); return 1; }
      cudaMemcpy(out.data(), dout, out.size() * 4, cudaMemcpyDeviceToHost);
      double maxerr = 0, maxref = 0;
      for (int m = 0; m < 128; ++m)
        for (int n = 0; n < 64; ++n) {
          double ref = 0;
          if (mode != 2) for (int k = 0; k < 64; ++k) ref += (double)fa16[m * 64 + k] * fb16[n * 64 + k];
          if (mode != 0) {
            const int q = shift + (m >> 3) * kPitch + (m & 7);
            for (int k = 0; k < 64; ++k) ref += (double)dec_e4m3(a8[q * 64 + k]) * dec_e5m2(b8[n * 64 + k]);
          }
          maxerr = fmax(maxerr, fabs(ref - out[m * 64 + n])); maxref = fmax(maxref, fabs(ref));
        }
      const bool ok = maxerr <= 1e-3 * fmax(maxref, 1.0);
      bad_total += !ok;
      printf("mode %d (%s) window shift %2d: max |err| %.3e (max |ref| %.2f) %s\n", mode,
             mode == 0 ? "f16" : mode == 1 ? "f16 + f8f6f4 into one accumulator" : "f8f6f4", shift, maxerr, maxref, ok ? "ok" : "MISMATCH");
    }
  for (int mode : {0, 2, 1}) {
    unsigned long long cyc = 0;
    probe<<<1, 128, 64 * 1024>>>(da16, db16, da8, db8, dout, mode, 11, 2000, dcyc);
    cudaDeviceSynchronize();
    cudaMemcpy(&cyc, dcyc, 8, cudaMemcpyDeviceToHost);
    const int per_it = mode == 0 ? 4 : mode == 2 ? 2 : 6;
    printf("mode %d: %.1f cycles per K=64 step (%d MMAs: %.1f cycles per MMA)  [M=128 N=64 cta_group::1]\n", mode, cyc / 2000.0, per_it,
           cyc / 2000.0 / per_it);
  }
  printf(bad_total ? "RESULT: %d mismatches\n" : "RESULT: all ok\n", bad_total);
  return bad_total != 0;
}
