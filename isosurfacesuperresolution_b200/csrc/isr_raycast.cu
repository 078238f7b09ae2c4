// isr_raycast.cu -- dense-grid isosurface ray-caster emitting the renderer's 12-channel G-buffer.
//
// What it restates: custom_iso_kernel of GPURendererDirect/render_kernel.cu:212-266 (per-pixel ray, first sample >= iso
// on a fixed-step march, 10 bisection steps back to the last point OUTSIDE the surface :175-190, gradient normal, Phong
// colour :234-239, flow from the current and next view-projection matrices :241-247, screen-space depth and normal
// :249-251, ray-cast ambient occlusion :109-148) and the stock surface tracer used for the AO rays
// (third-party/include/gvdb/cuda_gvdb_raycast.cuh:257-276, gradient :132-141).  The reference walks a sparse GVDB brick
// atlas through hardware-filtered 3-D textures; GVDB is a closed Windows binary and no .vbx volume ships with the
// repository, so this kernel works on ONE dense fp32 brick [nz][ny][nx] (x fastest, voxel centres at i + 0.5 as in a
// texture, clamp-to-edge) with trilinear interpolation in fp32 arithmetic (written out, compiled -fmad=false) so that
// the CPU restatement in oracle/raycast_oracle.c reproduces every hit decision bit for bit.
//
// Thread mapping: one thread per pixel, 8 x 4 pixel tiles per warp (neighbouring rays take the same number of steps and
// touch the same voxels; the volume is read through the read-only path and stays L1 / L2 resident).
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/isr.h"
#include "isr_internal.h"

namespace isr {

namespace {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 v3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 cross3(V3 a, V3 b) {
  return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float len3(V3 a) { return sqrtf(dot3(a, a)); }
__device__ __forceinline__ V3 normalize3(V3 a) {
  const float inv = 1.0f / len3(a);
  return v3(a.x * inv, a.y * inv, a.z * inv);
}
__device__ __forceinline__ V3 ld3(const float *p) { return v3(p[0], p[1], p[2]); }

constexpr int kSkipCell = 8;   // coarse macro-cell edge in voxels
constexpr int kSkipFine = 2;   // fine macro-cell edge

struct Grid {
  const float *__restrict__ v;
  int nx, ny, nz;
  const float *__restrict__ skip;   // max per coarse macro-cell, or nullptr
  int cx, cy, cz;
  const float *__restrict__ fine;   // max per fine macro-cell (follows the coarse grid in the same buffer)
  int fx, fy, fz;
  const float *__restrict__ dual;   // max of the 8 voxels of each interpolation cell, [nz][ny][nx] (follows the fine grid)
};

__device__ __forceinline__ void cell(float p, int n, int &i, float &f) {
  float q = p - 0.5f;
  q = q < 0.f ? 0.f : q;
  const float hi = static_cast<float>(n - 1);
  q = q > hi ? hi : q;
  int k = static_cast<int>(floorf(q));
  k = k > n - 2 ? n - 2 : k;
  i = k;
  f = q - static_cast<float>(k);
}

// trilinear sample at index coordinates p (x then y then z, each lerp as a + f * (b - a))
__device__ __forceinline__ float sample(const Grid &g, V3 p) {
  int ix, iy, iz;
  float fx, fy, fz;
  cell(p.x, g.nx, ix, fx);
  cell(p.y, g.ny, iy, fy);
  cell(p.z, g.nz, iz, fz);
  const float *b = g.v + (static_cast<size_t>(iz) * g.ny + iy) * g.nx + ix;
  const size_t sy = g.nx, sz = static_cast<size_t>(g.nx) * g.ny;
  const float v000 = __ldg(b), v100 = __ldg(b + 1), v010 = __ldg(b + sy), v110 = __ldg(b + sy + 1);
  const float v001 = __ldg(b + sz), v101 = __ldg(b + sz + 1), v011 = __ldg(b + sz + sy), v111 = __ldg(b + sz + sy + 1);
  const float c00 = v000 + fx * (v100 - v000), c10 = v010 + fx * (v110 - v010);
  const float c01 = v001 + fx * (v101 - v001), c11 = v011 + fx * (v111 - v011);
  const float c0 = c00 + fy * (c10 - c00), c1 = c01 + fy * (c11 - c01);
  return c0 + fz * (c1 - c0);
}

// slab test against the brick [0, n] in index coordinates; returns false on a miss
__device__ __forceinline__ bool clip(const Grid &g, V3 p0, V3 d, float &tn, float &tf) {
  tn = 0.f;
  tf = 3.0e38f;
  const float o[3] = {p0.x, p0.y, p0.z}, dd[3] = {d.x, d.y, d.z};
  const float n[3] = {static_cast<float>(g.nx), static_cast<float>(g.ny), static_cast<float>(g.nz)};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    if (fabsf(dd[a]) > 1e-12f) {
      float t1 = (0.f - o[a]) / dd[a], t2 = (n[a] - o[a]) / dd[a];
      if (t1 > t2) { const float t = t1; t1 = t2; t2 = t; }
      tn = t1 > tn ? t1 : tn;
      tf = t2 < tf ? t2 : tf;
    } else if (o[a] < 0.f || o[a] > n[a]) {
      return false;
    }
  }
  return tn <= tf;
}

// Number of samples the march may step over from p inside the macro-cell of edge `cell` that contains it, or 0 when the
// cell's max reaches iso.  A trilinear sample is a convex combination of its 8 voxels, so inside a cell whose max (over
// the cell plus a 1-voxel apron, the trilinear footprint) is below iso every sample is below iso.
__device__ __forceinline__ int hop(const float *__restrict__ grid, int gx, int gy, int gz, int cell, V3 p, V3 d, float iso,
                                   float inv_step) {
  const float inv_cell = 1.0f / static_cast<float>(cell);
  int cx = static_cast<int>(floorf(p.x * inv_cell)), cy = static_cast<int>(floorf(p.y * inv_cell)),
      cz = static_cast<int>(floorf(p.z * inv_cell));
  cx = cx < 0 ? 0 : (cx > gx - 1 ? gx - 1 : cx);
  cy = cy < 0 ? 0 : (cy > gy - 1 ? gy - 1 : cy);
  cz = cz < 0 ? 0 : (cz > gz - 1 ? gz - 1 : cz);
  if (!(__ldg(grid + (static_cast<size_t>(cz) * gy + cy) * gx + cx) < iso)) return 0;
  // distance to the cell's exit face (a point outside the brick sits in a clamped edge cell: its exit distance is then
  // over-estimated towards the brick only, where the clamped samples read the same edge voxels)
  float te = 3.0e38f;
  const float pc[3] = {p.x, p.y, p.z}, dc[3] = {d.x, d.y, d.z};
  const int cc[3] = {cx, cy, cz};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    if (fabsf(dc[a]) > 1e-12f) {
      const float face = static_cast<float>(dc[a] > 0.f ? (cc[a] + 1) * cell : cc[a] * cell);
      const float ta = (face - pc[a]) / dc[a];
      te = ta < te ? ta : te;
    }
  }
  int k = te > 0.f ? static_cast<int>(floorf((te - 0.25f) * inv_step)) : 1;   // stay 1/4 voxel short of the face
  return k < 1 ? 1 : (k > (1 << 20) ? (1 << 20) : k);
}

// Third level: the interpolation cell itself.  A sample at p interpolates the 8 voxels (i0, i0 + 1) per axis with
// i0 = the index cell() computes; dual[i0] is their max, so the bound is exact for the sample at p, and for the
// following samples as long as they stay inside [i0 + 0.5, i0 + 1.5) per axis (open-ended at the clamped brick faces).
// The samples stepped over keep eps = 1/512 voxel of distance to the faces along every axis (rounding of p is ~1e-4).
__device__ __forceinline__ int hop_dual(const Grid &g, V3 p, V3 d, float iso, float inv_step) {
  int i0[3];
  float f;
  cell(p.x, g.nx, i0[0], f);
  cell(p.y, g.ny, i0[1], f);
  cell(p.z, g.nz, i0[2], f);
  if (!(__ldg(g.dual + (static_cast<size_t>(i0[2]) * g.ny + i0[1]) * g.nx + i0[0]) < iso)) return 0;
  const float eps = 1.0f / 512.0f;
  float te = 3.0e38f;
  const float pc[3] = {p.x, p.y, p.z}, dc[3] = {d.x, d.y, d.z};
  const int n[3] = {g.nx, g.ny, g.nz};
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    if (dc[a] > 1e-12f) {
      if (i0[a] < n[a] - 2) {                                // otherwise the cell extends to +infinity (clamped)
        const float ta = ((static_cast<float>(i0[a]) + 1.5f - eps) - pc[a]) / dc[a];
        te = ta < te ? ta : te;
      }
    } else if (dc[a] < -1e-12f) {
      if (i0[a] > 0) {
        const float ta = ((static_cast<float>(i0[a]) + 0.5f + eps) - pc[a]) / dc[a];
        te = ta < te ? ta : te;
      }
    }
  }
  int k = te > 0.f ? static_cast<int>(floorf(te * inv_step)) : 1;
  return k < 1 ? 1 : (k > (1 << 20) ? (1 << 20) : k);
}

// fixed-step march: first sample >= iso at t = t0 + i * step; returns false when the ray leaves the brick.
// With the skip grids (8^3 macro-cells, then 2^3, then the interpolation cell itself) whole runs of samples that cannot reach iso are
// stepped over.  Only the index i advances -- the samples that ARE evaluated sit at the same t = t0 + i * step as
// without the grids, so the result is identical to the plain march (tests compare the two bit for bit).
__device__ __forceinline__ bool march(const Grid &g, V3 p0, V3 d, float iso, float step, float &t_hit) {
  float tn, tf;
  if (!clip(g, p0, d, tn, tf)) return false;
  const float t0 = step * ceilf(tn / step);
  const float inv_step = 1.0f / step;
  for (int i = 0; i < (1 << 22);) {
    const float t = t0 + static_cast<float>(i) * step;
    if (t > tf) return false;
    const V3 p = p0 + t * d;
    if (g.skip != nullptr) {
      int k = hop(g.skip, g.cx, g.cy, g.cz, kSkipCell, p, d, iso, inv_step);
      if (k == 0) k = hop(g.fine, g.fx, g.fy, g.fz, kSkipFine, p, d, iso, inv_step);
      if (k == 0) k = hop_dual(g, p, d, iso, inv_step);
      if (k > 0) { i += k; continue; }
    }
    if (sample(g, p) >= iso) { t_hit = t; return true; }
    ++i;
  }
  return false;
}

__device__ __forceinline__ float powi(float b, int e) {   // b^e by squaring (e >= 0), the same sequence as the oracle
  float r = 1.f;
  while (e > 0) {
    if (e & 1) r = r * b;
    b = b * b;
    e >>= 1;
  }
  return r;
}

__device__ __forceinline__ void matvec4(const float *m, V3 v, float w, float out[4]) {   // row-major 4x4
#pragma unroll
  for (int r = 0; r < 4; ++r) out[r] = ((m[4 * r] * v.x + m[4 * r + 1] * v.y) + m[4 * r + 2] * v.z) + m[4 * r + 3] * w;
}

}  // namespace

// max over the voxels a sample inside macro-cell c can touch: [cell * c - 1, cell * c + cell] per axis (the trilinear
// footprint), clamped to the brick.  One block per coarse cell, one thread per fine cell.
__global__ void skip_cells_kernel(const float *__restrict__ vol, float *__restrict__ skip, int nx, int ny, int nz, int cx,
                                  int cy, int cz) {
  const int c = blockIdx.x;
  const int ix = c % cx, iy = (c / cx) % cy, iz = c / (cx * cy);
  const int e = kSkipCell + 2;
  float m = -3.0e38f;
  for (int j = threadIdx.x; j < e * e * e; j += blockDim.x) {
    int x = ix * kSkipCell - 1 + j % e, y = iy * kSkipCell - 1 + (j / e) % e, z = iz * kSkipCell - 1 + j / (e * e);
    x = x < 0 ? 0 : (x > nx - 1 ? nx - 1 : x);
    y = y < 0 ? 0 : (y > ny - 1 ? ny - 1 : y);
    z = z < 0 ? 0 : (z > nz - 1 ? nz - 1 : z);
    m = fmaxf(m, __ldg(vol + (static_cast<size_t>(z) * ny + y) * nx + x));
  }
  __shared__ float red[4];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) skip[c] = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
}

__global__ void skip_fine_kernel(const float *__restrict__ vol, float *__restrict__ fine, int nx, int ny, int nz, int fx,
                                 int fy, int fz) {
  const size_t total = static_cast<size_t>(fx) * fy * fz;
  for (size_t c = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; c < total;
       c += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int ix = static_cast<int>(c % fx), iy = static_cast<int>((c / fx) % fy), iz = static_cast<int>(c / (static_cast<size_t>(fx) * fy));
    const int e = kSkipFine + 2;
    float m = -3.0e38f;
    for (int j = 0; j < e * e * e; ++j) {
      int x = ix * kSkipFine - 1 + j % e, y = iy * kSkipFine - 1 + (j / e) % e, z = iz * kSkipFine - 1 + j / (e * e);
      x = x < 0 ? 0 : (x > nx - 1 ? nx - 1 : x);
      y = y < 0 ? 0 : (y > ny - 1 ? ny - 1 : y);
      z = z < 0 ? 0 : (z > nz - 1 ? nz - 1 : z);
      m = fmaxf(m, __ldg(vol + (static_cast<size_t>(z) * ny + y) * nx + x));
    }
    fine[c] = m;
  }
}

// dual[z][y][x] = max of the 8 voxels (x..x+1, y..y+1, z..z+1), indices clamped: the interpolation cell of cell()
__global__ void skip_dual_kernel(const float *__restrict__ vol, float *__restrict__ dual, int nx, int ny, int nz) {
  const size_t total = static_cast<size_t>(nx) * ny * nz;
  for (size_t c = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; c < total;
       c += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int x = static_cast<int>(c % nx), y = static_cast<int>((c / nx) % ny), z = static_cast<int>(c / (static_cast<size_t>(nx) * ny));
    const int x1 = x + 1 < nx ? x + 1 : x, y1 = y + 1 < ny ? y + 1 : y, z1 = z + 1 < nz ? z + 1 : z;
    float m = -3.0e38f;
    const int xs[2] = {x, x1}, ys[2] = {y, y1}, zs[2] = {z, z1};
#pragma unroll
    for (int j = 0; j < 8; ++j)
      m = fmaxf(m, __ldg(vol + (static_cast<size_t>(zs[j >> 2]) * ny + ys[(j >> 1) & 1]) * nx + xs[j & 1]));
    dual[c] = m;
  }
}

__global__ void __launch_bounds__(128) raycast_iso_kernel(const float *__restrict__ vol, const float *__restrict__ skip,
                                                          float *__restrict__ out, const __grid_constant__ RaycastArgs a) {
  // 8 x 4 pixels per warp, 2 x 2 warps per block (16 x 8 pixels)
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int x = blockIdx.x * 16 + (wid & 1) * 8 + (lane & 7);
  const int y = blockIdx.y * 8 + (wid >> 1) * 4 + (lane >> 3);
  if (x >= a.width || y >= a.height) return;
  const float *P = a.p;
  Grid g;
  g.v = vol; g.nx = a.nx; g.ny = a.ny; g.nz = a.nz;
  g.skip = skip;
  g.cx = (a.nx + kSkipCell - 1) / kSkipCell; g.cy = (a.ny + kSkipCell - 1) / kSkipCell; g.cz = (a.nz + kSkipCell - 1) / kSkipCell;
  g.fine = skip + static_cast<size_t>(g.cx) * g.cy * g.cz;
  g.fx = (a.nx + kSkipFine - 1) / kSkipFine; g.fy = (a.ny + kSkipFine - 1) / kSkipFine; g.fz = (a.nz + kSkipFine - 1) / kSkipFine;
  g.dual = g.fine + static_cast<size_t>(g.fx) * g.fy * g.fz;

  float color[3] = {0.f, 0.f, 0.f}, mask = 0.f, normal[3] = {0.f, 0.f, 0.f}, depth = 0.f, flow[2] = {0.f, 0.f}, ao = 1.f;
  const float shadow = 1.f;

  if (x >= a.viewport[0] && y >= a.viewport[1] && x < a.viewport[2] && y < a.viewport[3]) {
    const float sx = (static_cast<float>(x) + 0.5f) / static_cast<float>(a.width);
    const float sy = (static_cast<float>(y) + 0.5f) / static_cast<float>(a.height);
    const V3 rpos = ld3(P + ISR_RC_CAM_POS);
    const V3 rdir = normalize3((sx * ld3(P + ISR_RC_CAM_U) + sy * ld3(P + ISR_RC_CAM_V)) + ld3(P + ISR_RC_CAM_S));
    const V3 bmin = ld3(P + ISR_RC_BOX_MIN);
    const float vdel = P[ISR_RC_VOXEL], iso = P[ISR_RC_ISO], step = P[ISR_RC_STEP];
    const float inv_vdel = 1.0f / vdel;
    const V3 p0 = inv_vdel * (rpos - bmin);
    float t_in;
    if (march(g, p0, rdir, iso, step, t_in)) {
      // bisection towards the last point outside the surface (render_kernel.cu:175-190)
      float lo = t_in - step, up = t_in;
      for (int i = 0; i < a.binary_steps; ++i) {
        const float mid = 0.5f * (lo + up);
        if (sample(g, p0 + mid * rdir) >= iso) up = mid; else lo = mid;
      }
      const V3 p = p0 + lo * rdir;
      const V3 hit = vdel * p + bmin;
      V3 grad = v3(sample(g, v3(p.x - 0.5f, p.y, p.z)) - sample(g, v3(p.x + 0.5f, p.y, p.z)),
                   sample(g, v3(p.x, p.y - 0.5f, p.z)) - sample(g, v3(p.x, p.y + 0.5f, p.z)),
                   sample(g, v3(p.x, p.y, p.z - 0.5f)) - sample(g, v3(p.x, p.y, p.z + 0.5f)));
      const float gl = len3(grad);
      V3 norm = gl > 1e-6f ? v3(grad.x / gl, grad.y / gl, grad.z / gl) : v3(0.f, 0.f, 0.f);   // safe_normalize :152-157
      mask = 1.f;
      // Phong colour
      const V3 light = ld3(P + ISR_RC_LIGHT);
      const V3 eyedir = normalize3(rpos - hit);
      const float ndl = fabsf(dot3(norm, light));
      const V3 R = normalize3(light - (2.0f * dot3(norm, light)) * norm);
      float rde = dot3(R, eyedir);
      rde = rde > 0.f ? rde : 0.f;
      const int e = static_cast<int>(P[ISR_RC_SPEC_EXP]);
      const float sfac = (static_cast<float>(e) + 2.0f) / 6.82f * powi(rde, e);
#pragma unroll
      for (int c = 0; c < 3; ++c)
        color[c] = (P[ISR_RC_AMBIENT + c] + P[ISR_RC_DIFFUSE + c] * ndl) + P[ISR_RC_SPECULAR + c] * sfac;
      // flow, depth, screen-space normal
      float cur[4], nxt[4], nm[4];
      matvec4(P + ISR_RC_VIEW_CUR, hit, 1.f, cur);
      matvec4(P + ISR_RC_VIEW_NEXT, hit, 1.f, nxt);
      matvec4(P + ISR_RC_NORMAL_MAT, norm, 0.f, nm);
      flow[0] = 0.5f * (cur[0] / cur[3] - nxt[0] / nxt[3]);
      flow[1] = 0.5f * (cur[1] / cur[3] - nxt[1] / nxt[3]);
      depth = cur[2] / cur[3];
      normal[0] = nm[0]; normal[1] = nm[1]; normal[2] = nm[2];
      // ambient occlusion: hemisphere rays through the plain surface tracer (render_kernel.cu:109-148)
      if (a.ao_samples > 0 && a.ao_two_pass) ao = lo;   // the hit parameter, for raycast_ao_kernel (which overwrites it)
      if (a.ao_samples > 0 && !a.ao_two_pass) {
        const float ao_radius = P[ISR_RC_AO_RADIUS];
        const V3 apos = hit - P[ISR_RC_AO_BIAS] * rdir;
        const V3 ap0 = inv_vdel * (apos - bmin);
        const float *nz4 = a.rotations + 4 * ((x % 4) + 4 * (y % 4));
        const V3 noise = ld3(nz4);
        const V3 tangent = normalize3(noise - dot3(noise, norm) * norm);
        const V3 bitangent = cross3(norm, tangent);
        float acc = 0.f;
        for (int s = 0; s < a.ao_samples; ++s) {
          const V3 st = normalize3(ld3(a.hemisphere + 4 * s));
          const V3 sw = v3(dot3(v3(tangent.x, bitangent.x, norm.x), st), dot3(v3(tangent.y, bitangent.y, norm.y), st),
                           dot3(v3(tangent.z, bitangent.z, norm.z), st));
          float value = 1.f, th;
          if (march(g, ap0, sw, iso, step, th)) {
            const V3 h2 = vdel * (ap0 + th * sw) + bmin;
            const float dist = len3(apos - h2);
            float yy = (ao_radius / dist - 1.0f) / (0.0f - 1.0f);    // smoothstep(1, 0, r / d)
            yy = yy < 0.f ? 0.f : (yy > 1.f ? 1.f : yy);
            value = (yy * yy) * (3.0f - 2.0f * yy);
          }
          acc = acc + value;
        }
        ao = acc / static_cast<float>(a.ao_samples);
      }
    }
  }
  // one 48-byte record per pixel: three 16-byte stores
  float4 *o = reinterpret_cast<float4 *>(out + 12 * (static_cast<size_t>(y) * a.width + x));
  o[0] = make_float4(color[0], color[1], color[2], mask);
  o[1] = make_float4(normal[0], normal[1], normal[2], depth);
  o[2] = make_float4(flow[0], flow[1], ao, shadow);
}

// Second pass of the two-pass form (RaycastArgs::ao_two_pass): ambient occlusion with ONE THREAD PER (pixel, AO ray).
// In the one-pass kernel a thread traces all AO rays of its pixel: ncu showed 13 of 32 lanes active and 16 % of the warp
// slots occupied (pixels that miss idle, rays differ in length).  Here a block owns 8 x 2 pixels: 16 threads rebuild the
// hit point and normal from the ray parameter the first pass left in the AO channel (same expressions, no march), all
// 128 threads then take (pixel, ray) pairs, and the per-ray values
// are summed per pixel IN SAMPLE ORDER from shared memory, so the frame equals the one-pass frame bit for bit.
constexpr int kAoPix = 16;
__global__ void __launch_bounds__(128) raycast_ao_kernel(const float *__restrict__ vol, const float *__restrict__ skip,
                                                         float *__restrict__ out, const __grid_constant__ RaycastArgs a) {
  __shared__ float s_val[kAoPix][512];
  __shared__ float s_rec[kAoPix][12];     // apos (index space) 0..2, apos (world) 3..5, normal 6..8, hit flag 9
  const float *P = a.p;
  Grid g;
  g.v = vol; g.nx = a.nx; g.ny = a.ny; g.nz = a.nz;
  g.skip = skip;
  g.cx = (a.nx + kSkipCell - 1) / kSkipCell; g.cy = (a.ny + kSkipCell - 1) / kSkipCell; g.cz = (a.nz + kSkipCell - 1) / kSkipCell;
  g.fine = skip + static_cast<size_t>(g.cx) * g.cy * g.cz;
  g.fx = (a.nx + kSkipFine - 1) / kSkipFine; g.fy = (a.ny + kSkipFine - 1) / kSkipFine; g.fz = (a.nz + kSkipFine - 1) / kSkipFine;
  g.dual = g.fine + static_cast<size_t>(g.fx) * g.fy * g.fz;
  const V3 bmin = ld3(P + ISR_RC_BOX_MIN);
  const float vdel = P[ISR_RC_VOXEL], iso = P[ISR_RC_ISO], step = P[ISR_RC_STEP];
  const float inv_vdel = 1.0f / vdel;
  const int tid = threadIdx.x;
  if (tid < kAoPix) {
    const int x = blockIdx.x * 8 + (tid & 7), y = blockIdx.y * 2 + (tid >> 3);
    float flag = 0.f;
    if (x < a.width && y < a.height && x >= a.viewport[0] && y >= a.viewport[1] && x < a.viewport[2] && y < a.viewport[3]) {
      const float sx = (static_cast<float>(x) + 0.5f) / static_cast<float>(a.width);
      const float sy = (static_cast<float>(y) + 0.5f) / static_cast<float>(a.height);
      const V3 rpos = ld3(P + ISR_RC_CAM_POS);
      const V3 rdir = normalize3((sx * ld3(P + ISR_RC_CAM_U) + sy * ld3(P + ISR_RC_CAM_V)) + ld3(P + ISR_RC_CAM_S));
      const V3 p0 = inv_vdel * (rpos - bmin);
      const float *rec = out + 12 * (static_cast<size_t>(y) * a.width + x);
      if (rec[3] == 1.f) {                                   // hit: the first pass left the ray parameter in the AO channel
        const float lo = rec[10];
        const V3 p = p0 + lo * rdir;
        const V3 hit = vdel * p + bmin;
        V3 grad = v3(sample(g, v3(p.x - 0.5f, p.y, p.z)) - sample(g, v3(p.x + 0.5f, p.y, p.z)),
                     sample(g, v3(p.x, p.y - 0.5f, p.z)) - sample(g, v3(p.x, p.y + 0.5f, p.z)),
                     sample(g, v3(p.x, p.y, p.z - 0.5f)) - sample(g, v3(p.x, p.y, p.z + 0.5f)));
        const float gl = len3(grad);
        const V3 norm = gl > 1e-6f ? v3(grad.x / gl, grad.y / gl, grad.z / gl) : v3(0.f, 0.f, 0.f);
        const V3 apos = hit - P[ISR_RC_AO_BIAS] * rdir;
        const V3 ap0 = inv_vdel * (apos - bmin);
        s_rec[tid][0] = ap0.x; s_rec[tid][1] = ap0.y; s_rec[tid][2] = ap0.z;
        s_rec[tid][3] = apos.x; s_rec[tid][4] = apos.y; s_rec[tid][5] = apos.z;
        s_rec[tid][6] = norm.x; s_rec[tid][7] = norm.y; s_rec[tid][8] = norm.z;
        flag = 1.f;
      }
    }
    s_rec[tid][9] = flag;
  }
  __syncthreads();
  const float ao_radius = P[ISR_RC_AO_RADIUS];
  for (int pair = tid; pair < kAoPix * a.ao_samples; pair += 128) {
    const int pix = pair % kAoPix, sidx = pair / kAoPix;
    if (s_rec[pix][9] == 0.f) continue;
    const int x = blockIdx.x * 8 + (pix & 7), y = blockIdx.y * 2 + (pix >> 3);
    const V3 ap0 = v3(s_rec[pix][0], s_rec[pix][1], s_rec[pix][2]);
    const V3 apos = v3(s_rec[pix][3], s_rec[pix][4], s_rec[pix][5]);
    const V3 norm = v3(s_rec[pix][6], s_rec[pix][7], s_rec[pix][8]);
    const V3 noise = ld3(a.rotations + 4 * ((x % 4) + 4 * (y % 4)));
    const V3 tangent = normalize3(noise - dot3(noise, norm) * norm);
    const V3 bitangent = cross3(norm, tangent);
    const V3 st = normalize3(ld3(a.hemisphere + 4 * sidx));
    const V3 sw = v3(dot3(v3(tangent.x, bitangent.x, norm.x), st), dot3(v3(tangent.y, bitangent.y, norm.y), st),
                     dot3(v3(tangent.z, bitangent.z, norm.z), st));
    float value = 1.f, th;
    if (march(g, ap0, sw, iso, step, th)) {
      const V3 h2 = vdel * (ap0 + th * sw) + bmin;
      const float dist = len3(apos - h2);
      float yy = (ao_radius / dist - 1.0f) / (0.0f - 1.0f);
      yy = yy < 0.f ? 0.f : (yy > 1.f ? 1.f : yy);
      value = (yy * yy) * (3.0f - 2.0f * yy);
    }
    s_val[pix][sidx] = value;
  }
  __syncthreads();
  if (tid < kAoPix && s_rec[tid][9] != 0.f) {
    const int x = blockIdx.x * 8 + (tid & 7), y = blockIdx.y * 2 + (tid >> 3);
    float acc = 0.f;
    for (int sidx = 0; sidx < a.ao_samples; ++sidx) acc = acc + s_val[tid][sidx];
    out[12 * (static_cast<size_t>(y) * a.width + x) + 10] = acc / static_cast<float>(a.ao_samples);
  }
}

cudaError_t launch_raycast_iso(const float *vol, const float *skip, float *out, const RaycastArgs &args, cudaStream_t s) {
  const dim3 grid((args.width + 15) / 16, (args.height + 7) / 8);
  raycast_iso_kernel<<<grid, 128, 0, s>>>(vol, skip, out, args);
  count_launch();
  if (args.ao_samples > 0 && args.ao_two_pass) {
    const dim3 grid2((args.width + 7) / 8, (args.height + 1) / 2);
    raycast_ao_kernel<<<grid2, 128, 0, s>>>(vol, skip, out, args);
    count_launch();
  }
  return cudaGetLastError();
}

size_t raycast_skip_floats(int nx, int ny, int nz) {
  const size_t c = static_cast<size_t>((nx + kSkipCell - 1) / kSkipCell) * ((ny + kSkipCell - 1) / kSkipCell) *
                   ((nz + kSkipCell - 1) / kSkipCell);
  const size_t f = static_cast<size_t>((nx + kSkipFine - 1) / kSkipFine) * ((ny + kSkipFine - 1) / kSkipFine) *
                   ((nz + kSkipFine - 1) / kSkipFine);
  return c + f + static_cast<size_t>(nx) * ny * nz;
}

cudaError_t launch_raycast_skip(const float *vol, float *skip, int nx, int ny, int nz, cudaStream_t s) {
  const int cx = (nx + kSkipCell - 1) / kSkipCell, cy = (ny + kSkipCell - 1) / kSkipCell, cz = (nz + kSkipCell - 1) / kSkipCell;
  const int fx = (nx + kSkipFine - 1) / kSkipFine, fy = (ny + kSkipFine - 1) / kSkipFine, fz = (nz + kSkipFine - 1) / kSkipFine;
  skip_cells_kernel<<<cx * cy * cz, 128, 0, s>>>(vol, skip, nx, ny, nz, cx, cy, cz);
  count_launch();
  const size_t nf = static_cast<size_t>(fx) * fy * fz;
  const size_t blocks = (nf + 255) / 256;
  skip_fine_kernel<<<static_cast<unsigned>(blocks < 148 * 16 ? blocks : 148 * 16), 256, 0, s>>>(
      vol, skip + static_cast<size_t>(cx) * cy * cz, nx, ny, nz, fx, fy, fz);
  count_launch();
  const size_t nv = static_cast<size_t>(nx) * ny * nz;
  const size_t vb = (nv + 255) / 256;
  skip_dual_kernel<<<static_cast<unsigned>(vb < 148 * 16 ? vb : 148 * 16), 256, 0, s>>>(
      vol, skip + static_cast<size_t>(cx) * cy * cz + nf, nx, ny, nz);
  count_launch();
  return cudaGetLastError();
}

}  // namespace isr
