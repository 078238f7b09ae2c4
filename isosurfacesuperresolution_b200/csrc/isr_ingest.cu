// isr_ingest.cu -- the step in front of the network on the interactive path:
//   * G-buffer ingest: the renderer's H x W x 12 interleaved fp32 frame (GPURendererDirect/render_kernel.cu:254-265:
//     RGB, mask, nx, ny, nz, depth, flow_x, flow_y, AO, shadow) -> the network's planar low-res input with the channel
//     slicing / mask remap / clamps of inference/loadedmodel.py:84-105 and mainGUI.py:713-718, the raw flow, and the
//     in-painting hole mask (mask == 0), in ONE pass;
//   * flow in-painting on the GPU: replaces the per-frame GPU->CPU->GPU round trip through cv.inpaint(INPAINT_NS, radius 3)
//     of loadedmodel.py:77-82 by a push-pull pyramid fill.  A different algorithm from OpenCV's fast-marching
//     Navier-Stokes fill (parity with OpenCV is "visually equivalent", unpinned -- SURVEY.md 8f-1); the OpenCV path stays
//     selectable in LoadedModel.  The exact arithmetic is restated in oracle/oracle.py::inpaint_flow_pushpull.
// Compiled with -fmad=false: every multiply and add rounds separately, as the numpy oracle does.
#include <cuda_runtime.h>
#include <stdint.h>

#include "isr_internal.h"

namespace isr {

// ---- ingest ---------------------------------------------------------------------------------------------------
// variant: 0 unshaded (mask*2-1, normal, depth); 1 shaded RGB+mask; 2 + depth; 3 + normal; 4 + normal + depth (clamped)
__device__ __forceinline__ int ingest_channels(int variant) {
  return variant == 0 ? 5 : (variant == 1 ? 4 : (variant == 2 ? 5 : (variant == 3 ? 7 : 8)));
}

__global__ void __launch_bounds__(256)
ingest_kernel(const float *__restrict__ render, int interleaved, int c_render, float *__restrict__ low,
              float *__restrict__ flow, unsigned char *__restrict__ hole, float *__restrict__ image12, int B, int h,
              int w, int variant) {
  const size_t plane = static_cast<size_t>(h) * w;
  const size_t total = plane * B;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / plane, pix = i - b * plane;
    float v[12];
    if (interleaved) {
      const float *src = render + i * c_render;
      if (c_render == 12) {   // 48-byte pixels: three aligned 16-byte loads
        const float4 a = __ldg(reinterpret_cast<const float4 *>(src));
        const float4 c = __ldg(reinterpret_cast<const float4 *>(src) + 1);
        const float4 d = __ldg(reinterpret_cast<const float4 *>(src) + 2);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = c.x; v[5] = c.y; v[6] = c.z; v[7] = c.w;
        v[8] = d.x; v[9] = d.y; v[10] = d.z; v[11] = d.w;
      } else {
#pragma unroll
        for (int c = 0; c < 12; ++c) v[c] = c < c_render ? __ldg(src + c) : 0.f;
      }
    } else {
#pragma unroll
      for (int c = 0; c < 12; ++c) v[c] = c < c_render ? __ldg(render + (b * c_render + c) * plane + pix) : 0.f;
    }
    if (low != nullptr) {
      const int cl = ingest_channels(variant);
      float *dst = low + b * cl * plane + pix;
      if (variant == 0) {   // loadedmodel.py:85: cat(mask*2-1, normal, depth)
        dst[0] = v[3] * 2.f - 1.f;
        dst[plane] = v[4]; dst[2 * plane] = v[5]; dst[3 * plane] = v[6]; dst[4 * plane] = v[7];
      } else if (variant == 4) {   // loadedmodel.py:98: clamp(current_low[:, 0:8], 0, 1)
#pragma unroll
        for (int c = 0; c < 8; ++c) dst[c * plane] = fminf(fmaxf(v[c], 0.f), 1.f);
      } else if (variant == 3) {   // :100
#pragma unroll
        for (int c = 0; c < 7; ++c) dst[c * plane] = v[c];
      } else {                     // :102 / :104
#pragma unroll
        for (int c = 0; c < 4; ++c) dst[c * plane] = v[c];
        if (variant == 2) dst[4 * plane] = v[7];
      }
    }
    if (flow != nullptr) {
      flow[(b * 2) * plane + pix] = v[8];
      flow[(b * 2 + 1) * plane + pix] = v[9];
    }
    if (hole != nullptr) hole[i] = v[3] == 0.f ? 1 : 0;
    if (image12 != nullptr) {   // mainGUI.py:713-716: planar frame with the mask moved to [-1, 1]
      float *dst = image12 + b * 12 * plane + pix;
#pragma unroll
      for (int c = 0; c < 12; ++c) dst[c * plane] = c == 3 ? v[3] * 2.f - 1.f : v[c];
    }
  }
}

// ---- push-pull flow in-painting -----------------------------------------------------------------------------
// level 0: out = flow, weight = 1 - hole
__global__ void inpaint_init_kernel(const float *__restrict__ flow, const unsigned char *__restrict__ hole,
                                    float *__restrict__ out, float *__restrict__ w0, size_t plane, int B) {
  const size_t total = plane * B;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / plane, pix = i - b * plane;
    const bool valid = hole[i] == 0;
    w0[i] = valid ? 1.f : 0.f;
    out[(b * 2) * plane + pix] = valid ? flow[(b * 2) * plane + pix] : 0.f;
    out[(b * 2 + 1) * plane + pix] = valid ? flow[(b * 2 + 1) * plane + pix] : 0.f;
  }
}

// pull: coarse (Y, X) = weighted mean of its (up to) 2x2 valid children; weight 1 if any child was valid
__global__ void inpaint_pull_kernel(const float *__restrict__ sv, const float *__restrict__ sw, int sh, int swd,
                                    float *__restrict__ dv, float *__restrict__ dw, int dh, int dwd, int B) {
  const size_t dplane = static_cast<size_t>(dh) * dwd, splane = static_cast<size_t>(sh) * swd;
  const size_t total = dplane * B;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t b = i / dplane, pix = i - b * dplane;
    const int Y = static_cast<int>(pix / dwd), X = static_cast<int>(pix - static_cast<size_t>(Y) * dwd);
    float W = 0.f, V0 = 0.f, V1 = 0.f;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int y = 2 * Y + dy, x = 2 * X + dx;
        if (y < sh && x < swd) {
          const size_t o = static_cast<size_t>(y) * swd + x;
          const float wt = sw[b * splane + o];
          W = W + wt;
          V0 = V0 + wt * sv[(b * 2) * splane + o];
          V1 = V1 + wt * sv[(b * 2 + 1) * splane + o];
        }
      }
    const bool any = W > 0.f;
    dw[i] = any ? 1.f : 0.f;
    dv[(b * 2) * dplane + pix] = any ? V0 / W : 0.f;
    dv[(b * 2 + 1) * dplane + pix] = any ? V1 / W : 0.f;
  }
}

__device__ __forceinline__ void half_src(int dst, int n_in, int &i0, int &i1, float &l0, float &l1) {
  float src = 0.5f * (static_cast<float>(dst) + 0.5f) - 0.5f;
  src = src < 0.f ? 0.f : src;
  int a = static_cast<int>(src);
  a = a > n_in - 1 ? n_in - 1 : a;
  i0 = a;
  i1 = a < n_in - 1 ? a + 1 : a;
  l1 = src - static_cast<float>(a);
  l1 = fminf(fmaxf(l1, 0.f), 1.f);
  l0 = 1.f - l1;
}

// push: every fine pixel without a value takes the bilinear sample (x first, then y) of the already filled coarse level
__global__ void inpaint_push_kernel(const float *__restrict__ cv, int ch, int cw, float *__restrict__ fv,
                                    float *__restrict__ fw, int fh, int fwd, int B) {
  const size_t fplane = static_cast<size_t>(fh) * fwd, cplane = static_cast<size_t>(ch) * cw;
  const size_t total = fplane * B;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    if (fw[i] != 0.f) continue;
    const size_t b = i / fplane, pix = i - b * fplane;
    const int y = static_cast<int>(pix / fwd), x = static_cast<int>(pix - static_cast<size_t>(y) * fwd);
    int x0, x1, y0, y1;
    float lx0, lx1, ly0, ly1;
    half_src(x, cw, x0, x1, lx0, lx1);
    half_src(y, ch, y0, y1, ly0, ly1);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const float *p = cv + (b * 2 + c) * cplane;
      const float r0 = lx0 * p[static_cast<size_t>(y0) * cw + x0] + lx1 * p[static_cast<size_t>(y0) * cw + x1];
      const float r1 = lx0 * p[static_cast<size_t>(y1) * cw + x0] + lx1 * p[static_cast<size_t>(y1) * cw + x1];
      fv[(b * 2 + c) * fplane + pix] = ly0 * r0 + ly1 * r1;
    }
    fw[i] = 1.f;
  }
}

static inline int grid_for_n(size_t total) {
  size_t g = (total + 255) / 256;
  const size_t cap = 148 * 8;
  return static_cast<int>(g < cap ? (g == 0 ? 1 : g) : cap);
}

int ingest_low_channels(int variant) { return variant == 0 ? 5 : (variant == 1 ? 4 : (variant == 2 ? 5 : (variant == 3 ? 7 : 8))); }

cudaError_t launch_ingest(const float *render, int interleaved, int c_render, float *low, float *flow,
                          unsigned char *hole, float *image12, int B, int h, int w, int variant, cudaStream_t s) {
  ingest_kernel<<<grid_for_n(static_cast<size_t>(B) * h * w), 256, 0, s>>>(render, interleaved, c_render, low, flow, hole,
                                                                          image12, B, h, w, variant);
  count_launch();
  return cudaGetLastError();
}

// pyramid levels 1..L (level l has ceil(h / 2^l) x ceil(w / 2^l) pixels, down to 1 x 1): per level 2 value planes + 1
// weight plane per image; plus the level-0 weight plane
size_t inpaint_workspace_floats(int B, int h, int w) {
  size_t n = static_cast<size_t>(B) * h * w;   // level-0 weights
  int hh = h, ww = w;
  while (hh > 1 || ww > 1) {
    hh = (hh + 1) / 2;
    ww = (ww + 1) / 2;
    n += static_cast<size_t>(3) * B * hh * ww;
  }
  return n;
}

cudaError_t launch_inpaint_flow(const float *flow, const unsigned char *hole, float *out, int B, int h, int w,
                                float *ws, cudaStream_t s) {
  float *w0 = ws;
  inpaint_init_kernel<<<grid_for_n(static_cast<size_t>(B) * h * w), 256, 0, s>>>(flow, hole, out, w0,
                                                                                static_cast<size_t>(h) * w, B);
  count_launch();
  constexpr int kMaxLevels = 32;
  float *lv[kMaxLevels], *lw[kMaxLevels];
  int lh[kMaxLevels], lwd[kMaxLevels];
  lv[0] = out; lw[0] = w0; lh[0] = h; lwd[0] = w;
  int L = 0;
  float *cur = ws + static_cast<size_t>(B) * h * w;
  while ((lh[L] > 1 || lwd[L] > 1) && L + 1 < kMaxLevels) {
    const int nh = (lh[L] + 1) / 2, nw = (lwd[L] + 1) / 2;
    lv[L + 1] = cur;
    lw[L + 1] = cur + static_cast<size_t>(2) * B * nh * nw;
    cur += static_cast<size_t>(3) * B * nh * nw;
    lh[L + 1] = nh; lwd[L + 1] = nw;
    inpaint_pull_kernel<<<grid_for_n(static_cast<size_t>(B) * nh * nw), 256, 0, s>>>(lv[L], lw[L], lh[L], lwd[L], lv[L + 1],
                                                                                    lw[L + 1], nh, nw, B);
    count_launch();
    ++L;
  }
  for (int l = L - 1; l >= 0; --l) {
    inpaint_push_kernel<<<grid_for_n(static_cast<size_t>(B) * lh[l] * lwd[l]), 256, 0, s>>>(lv[l + 1], lh[l + 1], lwd[l + 1],
                                                                                           lv[l], lw[l], lh[l], lwd[l], B);
    count_launch();
  }
  return cudaGetLastError();
}

}  // namespace isr
