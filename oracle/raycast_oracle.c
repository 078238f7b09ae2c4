/* raycast_oracle.c -- CPU restatement of the dense-grid isosurface ray-caster.  TEST INFRASTRUCTURE ONLY: used by tests/,
 * never by the product path.
 *
 * PARITY UNPINNED: the reference's renderer (GPURendererDirect/render_kernel.cu:109-266) runs only inside NVIDIA GVDB (a
 * closed Windows binary, third-party/dll-release/libgvdb.dll) on .vbx volumes the repository does not ship, and the
 * reference holds no golden frames.  This file restates the per-pixel arithmetic of that kernel on one dense brick:
 *   ray set-up        getViewRay, third-party/include/gvdb/cuda_gvdb_geom.cuh:66-74; render_kernel.cu:221-222
 *   march + bisection raySurfaceTrilinearBrickCustom, render_kernel.cu:161-201
 *   gradient          getGradient, third-party/include/gvdb/cuda_gvdb_raycast.cuh:132-141
 *   colour, flow, depth, normal   render_kernel.cu:231-251
 *   ambient occlusion computeAmbientOcclusion (ray mode), render_kernel.cu:109-148, with the stock tracer
 *                     raySurfaceTrilinearBrick, cuda_gvdb_raycast.cuh:257-276
 * Compiled with -ffp-contract=off; every expression is written in the order the CUDA kernel (compiled -fmad=false) uses,
 * so results are expected to agree bit for bit.  Analytic checks (sphere) live in tests/test_raycast_oracle.py. */
#include <math.h>
#include <stddef.h>
#include <string.h>

enum { P_POS = 0, P_U = 3, P_V = 6, P_S = 9, P_BMIN = 12, P_VOXEL = 15, P_ISO = 16, P_STEP = 17, P_CUR = 18, P_NEXT = 34,
       P_NMAT = 50, P_LIGHT = 66, P_AMB = 69, P_DIF = 72, P_SPEC = 75, P_EXP = 78, P_AOR = 79, P_AOB = 80 };

typedef struct { const float *v; int nx, ny, nz; } brick;

static float dot(const float *a, const float *b) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }
static float norm(const float *a) { return sqrtf(dot(a, a)); }
static void unit(const float *a, float *o) {
  float inv = 1.0f / norm(a);
  o[0] = a[0] * inv; o[1] = a[1] * inv; o[2] = a[2] * inv;
}
static void at(const float *p0, float t, const float *d, float *o) {
  for (int k = 0; k < 3; ++k) o[k] = p0[k] + t * d[k];
}

static void split(float p, int n, int *i, float *f) {
  float q = p - 0.5f;
  if (q < 0.f) q = 0.f;
  if (q > (float)(n - 1)) q = (float)(n - 1);
  int k = (int)floorf(q);
  if (k > n - 2) k = n - 2;
  *i = k;
  *f = q - (float)k;
}

static float trilinear(const brick *b, const float *p) {
  int i[3];
  float f[3];
  split(p[0], b->nx, &i[0], &f[0]);
  split(p[1], b->ny, &i[1], &f[1]);
  split(p[2], b->nz, &i[2], &f[2]);
  float plane[2];
  for (int dz = 0; dz < 2; ++dz) {
    float row[2];
    for (int dy = 0; dy < 2; ++dy) {
      const float *r = b->v + ((size_t)(i[2] + dz) * b->ny + (i[1] + dy)) * b->nx + i[0];
      row[dy] = r[0] + f[0] * (r[1] - r[0]);
    }
    plane[dz] = row[0] + f[1] * (row[1] - row[0]);
  }
  return plane[0] + f[2] * (plane[1] - plane[0]);
}

static int slab(const brick *b, const float *o, const float *d, float *tn, float *tf) {
  const float n[3] = {(float)b->nx, (float)b->ny, (float)b->nz};
  *tn = 0.f;
  *tf = 3.0e38f;
  for (int k = 0; k < 3; ++k) {
    if (fabsf(d[k]) > 1e-12f) {
      float t1 = (0.f - o[k]) / d[k], t2 = (n[k] - o[k]) / d[k];
      if (t1 > t2) { float t = t1; t1 = t2; t2 = t; }
      if (t1 > *tn) *tn = t1;
      if (t2 < *tf) *tf = t2;
    } else if (o[k] < 0.f || o[k] > n[k]) {
      return 0;
    }
  }
  return *tn <= *tf;
}

static int trace(const brick *b, const float *p0, const float *d, float iso, float step, float *t_hit) {
  float tn, tf, p[3];
  if (!slab(b, p0, d, &tn, &tf)) return 0;
  const float t0 = step * ceilf(tn / step);
  for (int i = 0; i < (1 << 22); ++i) {
    const float t = t0 + (float)i * step;
    if (t > tf) return 0;
    at(p0, t, d, p);
    if (trilinear(b, p) >= iso) { *t_hit = t; return 1; }
  }
  return 0;
}

static float ipow(float base, int e) {
  float r = 1.f;
  while (e > 0) {
    if (e & 1) r = r * base;
    base = base * base;
    e >>= 1;
  }
  return r;
}

static void mat4(const float *m, const float *v, float w, float *o) {
  for (int r = 0; r < 4; ++r) o[r] = ((m[4 * r] * v[0] + m[4 * r + 1] * v[1]) + m[4 * r + 2] * v[2]) + m[4 * r + 3] * w;
}

static void pixel(const brick *b, const float *P, const float *hemi, const float *rot, int ao_samples, int binary_steps,
                  int x, int y, int width, int height, float *rec) {
  const float sx = ((float)x + 0.5f) / (float)width, sy = ((float)y + 0.5f) / (float)height;
  float raw[3], rdir[3], p0[3];
  for (int k = 0; k < 3; ++k) raw[k] = (sx * P[P_U + k] + sy * P[P_V + k]) + P[P_S + k];
  unit(raw, rdir);
  const float vdel = P[P_VOXEL], iso = P[P_ISO], step = P[P_STEP], inv_vdel = 1.0f / vdel;
  for (int k = 0; k < 3; ++k) p0[k] = inv_vdel * (P[P_POS + k] - P[P_BMIN + k]);
  float t_in;
  if (!trace(b, p0, rdir, iso, step, &t_in)) return;
  float lo = t_in - step, up = t_in, p[3], hit[3], q[3];
  for (int i = 0; i < binary_steps; ++i) {
    const float mid = 0.5f * (lo + up);
    at(p0, mid, rdir, q);
    if (trilinear(b, q) >= iso) up = mid; else lo = mid;
  }
  at(p0, lo, rdir, p);
  for (int k = 0; k < 3; ++k) hit[k] = vdel * p[k] + P[P_BMIN + k];
  float grad[3], n[3];
  for (int k = 0; k < 3; ++k) {
    float m[3] = {p[0], p[1], p[2]}, pl[3] = {p[0], p[1], p[2]};
    m[k] = p[k] - 0.5f;
    pl[k] = p[k] + 0.5f;
    grad[k] = trilinear(b, m) - trilinear(b, pl);
  }
  const float gl = norm(grad);
  for (int k = 0; k < 3; ++k) n[k] = gl > 1e-6f ? grad[k] / gl : 0.f;
  rec[3] = 1.f;
  /* colour */
  const float *L = P + P_LIGHT;
  float eye_raw[3], eye[3], refl_raw[3], refl[3];
  for (int k = 0; k < 3; ++k) eye_raw[k] = P[P_POS + k] - hit[k];
  unit(eye_raw, eye);
  const float nl = dot(n, L), ndl = fabsf(nl);
  for (int k = 0; k < 3; ++k) refl_raw[k] = L[k] - (2.0f * nl) * n[k];
  unit(refl_raw, refl);
  float rde = dot(refl, eye);
  if (!(rde > 0.f)) rde = 0.f;
  const int e = (int)P[P_EXP];
  const float sfac = ((float)e + 2.0f) / 6.82f * ipow(rde, e);
  for (int c = 0; c < 3; ++c) rec[c] = (P[P_AMB + c] + P[P_DIF + c] * ndl) + P[P_SPEC + c] * sfac;
  /* flow, depth, normal */
  float cur[4], nxt[4], nm[4];
  mat4(P + P_CUR, hit, 1.f, cur);
  mat4(P + P_NEXT, hit, 1.f, nxt);
  mat4(P + P_NMAT, n, 0.f, nm);
  rec[8] = 0.5f * (cur[0] / cur[3] - nxt[0] / nxt[3]);
  rec[9] = 0.5f * (cur[1] / cur[3] - nxt[1] / nxt[3]);
  rec[7] = cur[2] / cur[3];
  rec[4] = nm[0]; rec[5] = nm[1]; rec[6] = nm[2];
  /* ambient occlusion */
  if (ao_samples > 0) {
    const float radius = P[P_AOR];
    float apos[3], ap0[3], tan_raw[3], tg[3], bt[3];
    for (int k = 0; k < 3; ++k) apos[k] = hit[k] - P[P_AOB] * rdir[k];
    for (int k = 0; k < 3; ++k) ap0[k] = inv_vdel * (apos[k] - P[P_BMIN + k]);
    const float *noise = rot + 4 * ((x % 4) + 4 * (y % 4));
    const float nn = dot(noise, n);
    for (int k = 0; k < 3; ++k) tan_raw[k] = noise[k] - nn * n[k];
    unit(tan_raw, tg);
    bt[0] = n[1] * tg[2] - n[2] * tg[1];
    bt[1] = n[2] * tg[0] - n[0] * tg[2];
    bt[2] = n[0] * tg[1] - n[1] * tg[0];
    float acc = 0.f;
    for (int s = 0; s < ao_samples; ++s) {
      float st[3], sw[3], th;
      unit(hemi + 4 * s, st);
      for (int k = 0; k < 3; ++k) {
        const float col[3] = {tg[k], bt[k], n[k]};
        sw[k] = dot(col, st);
      }
      float value = 1.f;
      if (trace(b, ap0, sw, iso, step, &th)) {
        float h2[3], dv[3];
        at(ap0, th, sw, h2);
        for (int k = 0; k < 3; ++k) dv[k] = apos[k] - (vdel * h2[k] + P[P_BMIN + k]);
        const float dist = norm(dv);
        float yy = (radius / dist - 1.0f) / (0.0f - 1.0f);
        yy = yy < 0.f ? 0.f : (yy > 1.f ? 1.f : yy);
        value = (yy * yy) * (3.0f - 2.0f * yy);
      }
      acc = acc + value;
    }
    rec[10] = acc / (float)ao_samples;
  }
}

/* out: height x width x 12 */
void isr_oracle_raycast_iso(const float *volume, int nx, int ny, int nz, const float *params, const float *hemisphere,
                            const float *rotations, int ao_samples, int binary_steps, const int *viewport, float *out,
                            int width, int height) {
  brick b = {volume, nx, ny, nz};
#pragma omp parallel for schedule(dynamic, 1)
  for (int y = 0; y < height; ++y)
    for (int x = 0; x < width; ++x) {
      float *rec = out + 12 * ((size_t)y * width + x);
      memset(rec, 0, 12 * sizeof(float));
      rec[10] = 1.f;
      rec[11] = 1.f;
      if (viewport == NULL || (x >= viewport[0] && y >= viewport[1] && x < viewport[2] && y < viewport[3]))
        pixel(&b, params, hemisphere, rotations, ao_samples, binary_steps, x, y, width, height, rec);
    }
}
