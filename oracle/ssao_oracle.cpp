// ssao_oracle.cpp -- TEST INFRASTRUCTURE ONLY (checker for isr_ssao).
// CPU restatement of ShadowFX::screenSpaceAmbientOcclusion (/root/reference/GPURenderer/ShadowFX.cpp:23-134) with a plain
// 3-float struct instead of openvdb::math::Vec3 / Mat3 (OpenVDB is not vendored in the reference tree; the members used
// are restated from its published headers: normalize = multiply by 1/length unless length <= 1e-7; Mat3(v1,v2,v3,false)
// takes the vectors as columns).  Pinned against the reference's own ShadowFX.cpp compiled in place
// (oracle/_ref/libshadowfx_ref.so, oracle/Makefile target _ref) by tests/test_ssao_oracle.py.
// Build: g++ -O2 -ffp-contract=off (no FMA contraction: every multiply and add rounds separately).
#include <algorithm>
#include <cmath>
#include <random>
#include <vector>

namespace {
struct V3 {
  float x, y, z;
};
inline float dot3(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 normalize3(V3 a) {
  const float d = std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z);
  if (!(std::fabs(d) > 1.0e-7f)) return a;
  const float inv = 1.0f / d;
  return V3{a.x * inv, a.y * inv, a.z * inv};
}
inline float smoothstep(float a, float b, float t) {   // ShadowFX.cpp:17-21
  t = std::max(0.0f, std::min(1.0f, (t - a) / (b - a)));
  return t * t * (3.0f - 2.0f * t);
}
}  // namespace

// ShadowFX.cpp:35-66; kernel: 3*samples floats, noise: 3*ns*ns floats indexed [x][y]
extern "C" void orc_ssao_tables(int samples, int ns, float *kernel, float *noise) {
  std::uniform_real_distribution<float> randomFloats(0.0, 1.0);
  std::default_random_engine generator(42);
  for (int i = 0; i < samples; ++i) {
    V3 s;
    // The three draws are constructor ARGUMENTS in the reference (:40-44): their evaluation order is unspecified in C++;
    // g++ and MSVC (the reference's compiler) both evaluate right to left -- z, y, x -- which oracle/_ref confirms.
    s.z = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);
    s.y = randomFloats(generator);
    s.x = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);
    s = normalize3(s);
    const float r = randomFloats(generator);
    s = V3{s.x * r, s.y * r, s.z * r};
    float scale = static_cast<float>(static_cast<float>(i) / 64.0);
    scale = 0.1f + scale * scale * (1.0f - 0.1f);
    kernel[3 * i] = s.x * scale;
    kernel[3 * i + 1] = s.y * scale;
    kernel[3 * i + 2] = s.z * scale;
  }
  for (int x = 0; x < ns; ++x)
    for (int y = 0; y < ns; ++y) {
      noise[(x * ns + y) * 3 + 2] = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);   // right to left, as above
      noise[(x * ns + y) * 3 + 1] = 0.0f;
      noise[(x * ns + y) * 3] = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);
    }
}

extern "C" void orc_ssao(int width, int height, const float *input, int es, int rs, int inx, int iny, int inz, int idepth,
                         float *output, int oes, int ors, int oidx, int samples, float radius, float bias, int ns) {
  std::vector<float> kernel(3 * samples), noise(3 * ns * ns);
  orc_ssao_tables(samples, ns, kernel.data(), noise.data());
  std::vector<float> buf(static_cast<size_t>(width) * height);   // :69 (the result is copied out after the loop, :114-133)
  std::vector<char> skip(static_cast<size_t>(width) * height, 0);
  for (int y = 0; y < height; ++y)
    for (int x = 0; x < width; ++x) {
      const size_t in_o = static_cast<size_t>(es) * (x + static_cast<size_t>(rs) * y);
      const float depth = input[idepth + in_o];
      if (depth == 0 || depth == 1) {   // :74-78: written immediately (and overwritten by the copy loop with buf = 0)
        output[oidx + static_cast<size_t>(oes) * (x + static_cast<size_t>(ors) * y)] = 0;
        skip[x + static_cast<size_t>(width) * y] = 1;
        continue;
      }
      const V3 normal = normalize3(V3{input[inx + in_o], input[iny + in_o], input[inz + in_o]});
      const V3 frag{x / float(width) * 2 - 1, y / float(height) * 2 - 1, depth * 2 - 1};
      const int ni = ((x % ns) * ns + (y % ns)) * 3;
      const V3 rnd{noise[ni], noise[ni + 1], noise[ni + 2]};
      const float rn = dot3(rnd, normal);
      const V3 tangent = normalize3(V3{rnd.x - normal.x * rn, rnd.y - normal.y * rn, rnd.z - normal.z * rn});
      const V3 bit{normal.y * tangent.z - normal.z * tangent.y, normal.z * tangent.x - normal.x * tangent.z,
                   normal.x * tangent.y - normal.y * tangent.x};
      float ssao = 0;
      for (int i = 0; i < samples; ++i) {
        const V3 k{kernel[3 * i], kernel[3 * i + 1], kernel[3 * i + 2]};
        const V3 t{tangent.x * k.x + bit.x * k.y + normal.x * k.z, tangent.y * k.x + bit.y * k.y + normal.y * k.z,
                   tangent.z * k.x + bit.z * k.y + normal.z * k.z};
        const V3 s{frag.x + t.x * radius, frag.y + t.y * radius, frag.z + t.z * radius};
        int sx = int(std::round((s.x + 1) * 0.5 * width));
        int sy = int(std::round((s.y + 1) * 0.5 * height));
        sx = std::min(width - 1, std::max(0, sx));
        sy = std::min(height - 1, std::max(0, sy));
        const float sampleDepth = s.z * 0.5 + 0.5;
        const float screenDepth = input[idepth + static_cast<size_t>(es) * (sx + static_cast<size_t>(rs) * sy)];
        const float rangeCheck = smoothstep(0.0, 1.0, 0.5 / std::fabs(depth - screenDepth));
        ssao += (screenDepth > sampleDepth + bias ? 1.0f : 0.0f) * rangeCheck;
      }
      buf[x + static_cast<size_t>(width) * y] = 1 - (ssao / samples);
    }
  for (int y = 0; y < height; ++y)
    for (int x = 0; x < width; ++x)
      output[oidx + static_cast<size_t>(oes) * (x + static_cast<size_t>(ors) * y)] = buf[x + static_cast<size_t>(width) * y];
}

// Ambient-occlusion tables of the ray-caster, GPURendererDirect.cpp:146-190: a default-seeded engine and ONE uniform(0,1)
// stream; 512 cosine-distributed hemisphere directions scaled into [0.1, 1], then 16 unit rotation vectors in the plane.
extern "C" void orc_raycast_ao_tables(float *hemisphere /* 4 * 512 */, float *rotations /* 4 * 16 */) {
  std::default_random_engine engine;
  std::uniform_real_distribution<float> uni(0.0f, 1.0f);
  for (int i = 0; i < 512; ++i) {
    float u1 = uni(engine);
    float u2 = uni(engine);
    float r = std::sqrt(u1);
    float theta = 2 * M_PI * u2;
    float s = uni(engine);
    s = 0.1 + 0.9 * s * s;
    hemisphere[4 * i + 0] = r * std::cos(theta) * s;
    hemisphere[4 * i + 1] = r * std::sin(theta) * s;
    hemisphere[4 * i + 2] = std::sqrt(1 - u1) * s;
    hemisphere[4 * i + 3] = 0;
  }
  for (int i = 0; i < 16; ++i) {
    float x = uni(engine) * 2 - 1;
    float y = uni(engine) * 2 - 1;
    float linv = 1.0f / std::sqrt(x * x + y * y);
    rotations[4 * i + 0] = x * linv;
    rotations[4 * i + 1] = y * linv;
    rotations[4 * i + 2] = 0;
    rotations[4 * i + 3] = 0;
  }
}
