// isr_ssao.cu -- screen-space ambient occlusion of the reference's renderer post-pass as a compute kernel
// (SURVEY.md section 8f rank 4): ShadowFX::screenSpaceAmbientOcclusion, /root/reference/GPURenderer/ShadowFX.cpp:23-134
// (called on the 12-channel frame at GPURenderer.cpp:716-723: normal = channels 4..6, depth = 7, AO written to 10).
// One thread per pixel, 64 hemisphere samples rotated by a 4x4 noise tile, depth range check; the blur of :114-133 is
// commented out in the reference (the result is copied), so it is a copy here too.
// The sample kernel / noise tables come from std::default_random_engine(42) on the host (isr_api.cu), exactly the
// reference's call sequence; float / double usage per expression follows the reference (0.5 and 2.0 literals are double).
// Compiled with -fmad=false.  Arithmetic restated on the CPU in oracle/ssao_oracle.cpp.
#include <cuda_runtime.h>
#include <stdint.h>

#include "isr_internal.h"

namespace isr {

struct V3 {
  float x, y, z;
};
__device__ __forceinline__ V3 v3(float x, float y, float z) { return V3{x, y, z}; }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// openvdb Vec3::normalize(): d = sqrt(x*x + y*y + z*z); if d ~ 0 (1e-7) leave unchanged; else *= 1/d
__device__ __forceinline__ V3 normalize3(V3 a) {
  const float d = sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
  if (fabsf(d) <= 1.0e-7f) return a;
  const float inv = 1.0f / d;
  return v3(a.x * inv, a.y * inv, a.z * inv);
}
__device__ __forceinline__ float smoothstep01(float t) {   // ShadowFX.cpp:17-21 with a = 0, b = 1
  t = fmaxf(0.0f, fminf(1.0f, (t - 0.0f) / (1.0f - 0.0f)));
  return t * t * (3.0f - 2.0f * t);
}

__global__ void __launch_bounds__(128)
ssao_kernel(const float *__restrict__ input, int es, int rs, int inx, int iny, int inz, int idepth, float *output,
            int oes, int ors, int oidx, int width, int height, int samples, float radius, float bias, int ns,
            SsaoTables tab) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 4 + (threadIdx.x >> 5);
  if (x >= width || y >= height) return;
  const size_t in_o = static_cast<size_t>(es) * (x + static_cast<size_t>(rs) * y);
  const size_t out_o = oidx + static_cast<size_t>(oes) * (x + static_cast<size_t>(ors) * y);
  const float depth = input[idepth + in_o];
  if (depth == 0.f || depth == 1.f) {   // :74-78
    output[out_o] = 0.f;
    return;
  }
  const V3 normal = normalize3(v3(input[inx + in_o], input[iny + in_o], input[inz + in_o]));
  const V3 frag = v3(static_cast<float>(x) / static_cast<float>(width) * 2.f - 1.f,
                     static_cast<float>(y) / static_cast<float>(height) * 2.f - 1.f, depth * 2.f - 1.f);   // :88
  const int ni = ((x % ns) * ns + (y % ns)) * 3;
  const V3 rnd = v3(tab.noise[ni], tab.noise[ni + 1], tab.noise[ni + 2]);
  const float rn = dot3(rnd, normal);
  const V3 tangent = normalize3(v3(rnd.x - normal.x * rn, rnd.y - normal.y * rn, rnd.z - normal.z * rn));   // :90
  const V3 bit = v3(normal.y * tangent.z - normal.z * tangent.y, normal.z * tangent.x - normal.x * tangent.z,
                    normal.x * tangent.y - normal.y * tangent.x);                                         // normal x tangent
  float ssao = 0.f;
  for (int i = 0; i < samples; ++i) {
    const V3 k = v3(tab.kernel[3 * i], tab.kernel[3 * i + 1], tab.kernel[3 * i + 2]);
    // TBN has tangent, bitangent, normal as COLUMNS (Mat3(v1, v2, v3, rows=false)); Mat3 * Vec3 sums along the row
    const V3 t = v3(tangent.x * k.x + bit.x * k.y + normal.x * k.z, tangent.y * k.x + bit.y * k.y + normal.y * k.z,
                    tangent.z * k.x + bit.z * k.y + normal.z * k.z);
    const V3 s = v3(frag.x + t.x * radius, frag.y + t.y * radius, frag.z + t.z * radius);                 // :98
    int sx = static_cast<int>(round(static_cast<double>(s.x + 1.0f) * 0.5 * width));   // :99 (float add, double product)
    int sy = static_cast<int>(round(static_cast<double>(s.y + 1.0f) * 0.5 * height));
    sx = min(width - 1, max(0, sx));
    sy = min(height - 1, max(0, sy));
    const float sample_depth = static_cast<float>(static_cast<double>(s.z) * 0.5 + 0.5);                  // :103
    const float screen_depth = input[idepth + static_cast<size_t>(es) * (sx + static_cast<size_t>(rs) * sy)];
    const float range = smoothstep01(static_cast<float>(0.5 / static_cast<double>(fabsf(depth - screen_depth))));   // :105
    ssao += (screen_depth > sample_depth + bias ? 1.0f : 0.0f) * range;
  }
  output[out_o] = 1.f - (ssao / static_cast<float>(samples));   // :108; the blur of :114-133 is a copy
}

cudaError_t launch_ssao(const float *input, int es, int rs, int inx, int iny, int inz, int idepth, float *output, int oes,
                        int ors, int oidx, int width, int height, int samples, float radius, float bias, int ns,
                        const SsaoTables &tab, cudaStream_t s) {
  dim3 grid((width + 31) / 32, (height + 3) / 4);
  ssao_kernel<<<grid, 128, 0, s>>>(input, es, rs, inx, iny, inz, idepth, output, oes, ors, oidx, width, height, samples,
                                  radius, bias, ns, tab);
  count_launch();
  return cudaGetLastError();
}

}  // namespace isr
