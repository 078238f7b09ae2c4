/*
 * isr_oracle.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement (plain C, fp32) of the
 * reference's frame-recurrent 4x EnhanceNet inference path.  It is the checker for the
 * CUDA product path; nothing under isosurfacesuperresolution_b200/ may call it.
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/SuperresolutionNetwork).  Third-party arithmetic on the path lives in
 * PyTorch ATen (conv2d, upsample_bilinear2d / nearest2d, grid_sampler_2d; torch 2.11.0, not
 * vendored in the reference): their published fp32 CPU algorithms are restated here in the
 * exact operation order pinned by SURVEY.md 7.4, and the restatement is pinned against the
 * golden vectors in tests/golden/ (generated from the live reference by oracle/gen_golden.py).
 *
 * Compile WITHOUT fast-math and with -ffp-contract=off: the fused multiply-adds that matter
 * are written explicitly as fmaf().
 */
#include <math.h>
#include <stddef.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))

/* ------------------------------------------------------------------------------------------
 * models/videotools.py:8-25  VideoTools.flatten_high  (inverse PixelShuffle)
 * out[b, c*r*r + dy*r + dx, y, x] = in[b, c, y*r+dy, x*r+dx]
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_flatten_high(const float *in, float *out, int B, int C, int H, int W, int r)
{
    int h = H / r, w = W / r;
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < C; ++c)
            for (int dy = 0; dy < r; ++dy)
                for (int dx = 0; dx < r; ++dx) {
                    int oc = c * r * r + dy * r + dx;
                    for (int y = 0; y < h; ++y)
                        for (int x = 0; x < w; ++x)
                            out[(((size_t)b * C * r * r + oc) * h + y) * w + x] =
                                in[(((size_t)b * C + c) * H + (y * r + dy)) * W + (x * r + dx)];
                }
}

/* torch upsample_bilinear2d, align_corners=False: src = max(0, scale*(dst+0.5)-0.5) with
 * scale = 1/s; i0 = floor(src); i1 = min(i0+1, n-1); l1 = src-i0; l0 = 1-l1.
 * Used by videotools.py:70, enhancenet.py:67-70,116,119, initial_image.py:41. */
static inline void bil_src(int dst, int n_in, float scale, int *i0, int *i1, float *l0, float *l1)
{
    float src = scale * ((float)dst + 0.5f) - 0.5f;
    if (src < 0.f) src = 0.f;
    int a = (int)src;
    if (a > n_in - 1) a = n_in - 1;
    *i0 = a;
    *i1 = (a < n_in - 1) ? a + 1 : a;
    float t = src - (float)a;
    if (t < 0.f) t = 0.f;
    if (t > 1.f) t = 1.f;
    *l1 = t;
    *l0 = 1.f - t;
}

static inline float bil_sample(const float *p, int w, int y0, int y1, int x0, int x1,
                               float ly0, float ly1, float lx0, float lx1)
{
    /* x first, then y; FMA placement per SURVEY 7.4 */
    float r0 = fmaf(lx0, p[(size_t)y0 * w + x0], lx1 * p[(size_t)y0 * w + x1]);
    float r1 = fmaf(lx0, p[(size_t)y1 * w + x0], lx1 * p[(size_t)y1 * w + x1]);
    return fmaf(ly0, r0, ly1 * r1);
}

/* F.interpolate(x, scale_factor=s, mode='bilinear') / nn.Upsample(scale_factor=s, 'bilinear') */
ORC_API void orc_upsample_bilinear(const float *in, float *out, int B, int C, int h, int w, int s)
{
    int H = h * s, W = w * s;
    float scale = 1.0f / (float)s;
#pragma omp parallel for collapse(2) schedule(static)
    for (int bc = 0; bc < B * C; ++bc)
        for (int y = 0; y < H; ++y) {
            const float *p = in + (size_t)bc * h * w;
            float *o = out + ((size_t)bc * H + y) * W;
            int y0, y1;
            float ly0, ly1;
            bil_src(y, h, scale, &y0, &y1, &ly0, &ly1);
            for (int x = 0; x < W; ++x) {
                int x0, x1;
                float lx0, lx1;
                bil_src(x, w, scale, &x0, &x1, &lx0, &lx1);
                o[x] = bil_sample(p, w, y0, y1, x0, x1, ly0, ly1, lx0, lx1);
            }
        }
}

/* nn.Upsample(scale_factor=s, mode='nearest') and F.interpolate(size=..., 'nearest'): src = dst / s */
ORC_API void orc_upsample_nearest(const float *in, float *out, int B, int C, int h, int w, int s)
{
    int H = h * s, W = w * s;
#pragma omp parallel for collapse(2) schedule(static)
    for (int bc = 0; bc < B * C; ++bc)
        for (int y = 0; y < H; ++y) {
            const float *p = in + (size_t)bc * h * w + (size_t)(y / s) * w;
            float *o = out + ((size_t)bc * H + y) * W;
            for (int x = 0; x < W; ++x) o[x] = p[x / s];
        }
}

/* torch.linspace(-1, 1, n) in fp32: symmetric evaluation from both ends */
static inline float lin_m1p1(int i, int n)
{
    if (n == 1) return -1.f;
    float step = 2.0f / (float)(n - 1);
    if (i < n / 2) return fmaf(step, (float)i, -1.f);
    return fmaf(-step, (float)(n - 1 - i), 1.f);
}

/* ------------------------------------------------------------------------------------------
 * models/videotools.py:49-87  VideoTools.warp_upscale(image_high, flow_low, r, special_mask)
 * grid_sample: bilinear, zeros padding, align_corners=False (installed-torch default).
 * img: B x C x H x W (H=h*r), flow: B x 2 x h x w, out: B x C x H x W
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_warp_upscale(const float *img, const float *flow, float *out, int B, int C, int h,
                              int w, int r, int special_mask)
{
    int H = h * r, W = w * r;
    float scale = 1.0f / (float)r;
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b)
        for (int y = 0; y < H; ++y) {
            const float *fx = flow + ((size_t)b * 2 + 0) * h * w;
            const float *fy = flow + ((size_t)b * 2 + 1) * h * w;
            int fy0, fy1;
            float ly0, ly1;
            bil_src(y, h, scale, &fy0, &fy1, &ly0, &ly1);
            for (int x = 0; x < W; ++x) {
                int fx0, fx1;
                float lx0, lx1;
                bil_src(x, w, scale, &fx0, &fx1, &lx0, &lx1);
                /* videotools.py:65-70: scale by (-2,+2) (exact), then bilinear x r */
                float r0 = fmaf(lx0, -2.f * fx[(size_t)fy0 * w + fx0], lx1 * (-2.f * fx[(size_t)fy0 * w + fx1]));
                float r1 = fmaf(lx0, -2.f * fx[(size_t)fy1 * w + fx0], lx1 * (-2.f * fx[(size_t)fy1 * w + fx1]));
                float Fx = fmaf(ly0, r0, ly1 * r1);
                r0 = fmaf(lx0, 2.f * fy[(size_t)fy0 * w + fx0], lx1 * (2.f * fy[(size_t)fy0 * w + fx1]));
                r1 = fmaf(lx0, 2.f * fy[(size_t)fy1 * w + fx0], lx1 * (2.f * fy[(size_t)fy1 * w + fx1]));
                float Fy = fmaf(ly0, r0, ly1 * r1);
                /* videotools.py:71-73 grid = offsets + flow_high */
                float gx = lin_m1p1(x, W) + Fx;
                float gy = lin_m1p1(y, H) + Fy;
                /* grid_sampler_unnormalize, align_corners=False: ((g+1)*size-1)/2 */
                float ix = fmaf(gx + 1.f, (float)W, -1.f) / 2.f;
                float iy = fmaf(gy + 1.f, (float)H, -1.f) / 2.f;
                float x0f = floorf(ix), y0f = floorf(iy);
                float x1f = x0f + 1.f, y1f = y0f + 1.f;
                float nw = (x1f - ix) * (y1f - iy);
                float ne = (ix - x0f) * (y1f - iy);
                float sw = (x1f - ix) * (iy - y0f);
                float se = (ix - x0f) * (iy - y0f);
                /* guard the int conversion for wild flows */
                int inx0 = (x0f >= -2.f && x0f <= (float)W + 1.f);
                int iny0 = (y0f >= -2.f && y0f <= (float)H + 1.f);
                int x0 = inx0 ? (int)x0f : -5, y0 = iny0 ? (int)y0f : -5;
                int x1 = x0 + 1, y1 = y0 + 1;
                int vx0 = (x0 >= 0 && x0 < W), vx1 = (x1 >= 0 && x1 < W);
                int vy0 = (y0 >= 0 && y0 < H), vy1 = (y1 >= 0 && y1 < H);
                for (int c = 0; c < C; ++c) {
                    const float *p = img + ((size_t)b * C + c) * H * W;
                    int sm = (special_mask && c == 0);
                    float acc = 0.f;
                    if (vy0 && vx0) { float v = p[(size_t)y0 * W + x0]; if (sm) v = v * 0.5f + 0.5f; acc += v * nw; }
                    if (vy0 && vx1) { float v = p[(size_t)y0 * W + x1]; if (sm) v = v * 0.5f + 0.5f; acc += v * ne; }
                    if (vy1 && vx0) { float v = p[(size_t)y1 * W + x0]; if (sm) v = v * 0.5f + 0.5f; acc += v * sw; }
                    if (vy1 && vx1) { float v = p[(size_t)y1 * W + x1]; if (sm) v = v * 0.5f + 0.5f; acc += v * se; }
                    if (sm) acc = acc * 2.f - 1.f;
                    out[(((size_t)b * C + c) * H + y) * W + x] = acc;
                }
            }
        }
}

/* ------------------------------------------------------------------------------------------
 * utils/initial_image.py:5-54  initialImage(current_input, channels, mode, aoInverted, upscaling)
 * mode: 0 = "zero", 1 = "unshaded", 2 = "input".  Returns 0, or -1 for an invalid request.
 * ---------------------------------------------------------------------------------------- */
ORC_API int orc_initial_image(const float *in, float *out, int B, int Cin, int h, int w, int channels,
                              int mode, int ao_inverted, int r)
{
    int H = h * r, W = w * r;
    size_t plane = (size_t)H * W;
    if (mode == 0) {
        memset(out, 0, sizeof(float) * B * channels * plane);
        return 0;
    }
    if (mode == 1) {
        if (channels != 5 && channels != 6) return -1;
        float d[6] = {-1.f, 0.f, 0.f, 1.f, 0.5f, ao_inverted ? 0.f : 1.f};
        for (int b = 0; b < B; ++b)
            for (int c = 0; c < channels; ++c)
                for (size_t i = 0; i < plane; ++i) out[((size_t)b * channels + c) * plane + i] = d[c];
        return 0;
    }
    if (mode == 2) {
        float *tmp = (float *)malloc(sizeof(float) * B * Cin * plane);
        orc_upsample_bilinear(in, tmp, B, Cin, h, w, r);
        for (int b = 0; b < B; ++b)
            for (int c = 0; c < channels; ++c) {
                float *o = out + ((size_t)b * channels + c) * plane;
                if (c < Cin)
                    memcpy(o, tmp + ((size_t)b * Cin + c) * plane, sizeof(float) * plane);
                else
                    for (size_t i = 0; i < plane; ++i) o[i] = 1.f;
            }
        free(tmp);
        return 0;
    }
    return -1;
}

/* nn.Conv2d(Cin, Cout, 3, padding=1) (+ optional ReLU), fp32, NCHW.
 * enhancenet.py:93-95,108-112,115-125.  Accumulation order: bias, then ci-major, ky, kx. */
ORC_API void orc_conv3x3(const float *in, const float *wgt, const float *bias, float *out, int B, int Cin,
                         int Cout, int H, int W, int relu)
{
#pragma omp parallel for collapse(2) schedule(static)
    for (int b = 0; b < B; ++b)
        for (int co = 0; co < Cout; ++co) {
            float *o = out + ((size_t)b * Cout + co) * H * W;
            for (size_t i = 0; i < (size_t)H * W; ++i) o[i] = bias[co];
            for (int ci = 0; ci < Cin; ++ci) {
                const float *p = in + ((size_t)b * Cin + ci) * H * W;
                const float *k = wgt + ((size_t)co * Cin + ci) * 9;
                for (int ky = 0; ky < 3; ++ky)
                    for (int kx = 0; kx < 3; ++kx) {
                        float wv = k[ky * 3 + kx];
                        int ylo = ky == 0 ? 1 : 0, yhi = ky == 2 ? H - 1 : H;
                        int xlo = kx == 0 ? 1 : 0, xhi = kx == 2 ? W - 1 : W;
                        for (int y = ylo; y < yhi; ++y) {
                            const float *src = p + (size_t)(y + ky - 1) * W + (kx - 1);
                            float *dst = o + (size_t)y * W;
                            for (int x = xlo; x < xhi; ++x) dst[x] += wv * src[x];
                        }
                    }
            }
            if (relu)
                for (size_t i = 0; i < (size_t)H * W; ++i) o[i] = o[i] > 0.f ? o[i] : 0.f;
        }
}

/* ------------------------------------------------------------------------------------------
 * models/enhancenet.py:136-145 EnhanceNet.forward (+ _recon_image :51-90), useBN=False.
 * params: 50 pointers in state_dict order (preblock.0.{weight,bias}, blocks.i.{0,2}.{weight,bias},
 * postblock.{1,4,6,8}.{weight,bias}).  upsample: 0 nearest, 1 bilinear.  recon_residual: 1/0.
 * n_mask = len(channel_mask).  out_recon / out_raw: B x Cout x 4h x 4w.
 * ---------------------------------------------------------------------------------------- */
ORC_API int orc_enhancenet_forward(const float *inputs, const float *const *params, float *out_recon,
                                   float *out_raw, int B, int Cin, int Cout, int h, int w, int upsample,
                                   int recon_residual, int n_mask)
{
    if (n_mask > Cout) return -1;
    size_t lo = (size_t)B * 64 * h * w;
    float *f = (float *)malloc(sizeof(float) * lo);
    float *t = (float *)malloc(sizeof(float) * lo);
    float *u = (float *)malloc(sizeof(float) * lo);
    int pi = 0;
    orc_conv3x3(inputs, params[0], params[1], f, B, Cin, 64, h, w, 1);
    pi = 2;
    for (int blk = 0; blk < 10; ++blk) {
        orc_conv3x3(f, params[pi], params[pi + 1], t, B, 64, 64, h, w, 1);
        orc_conv3x3(t, params[pi + 2], params[pi + 3], u, B, 64, 64, h, w, 0);
        for (size_t i = 0; i < lo; ++i) f[i] = f[i] + u[i];
        pi += 4;
    }
    free(t);
    free(u);
    int h2 = 2 * h, w2 = 2 * w, H = 4 * h, W = 4 * w;
    float *a2 = (float *)malloc(sizeof(float) * lo * 4);
    float *b2 = (float *)malloc(sizeof(float) * lo * 4);
    if (upsample) orc_upsample_bilinear(f, a2, B, 64, h, w, 2); else orc_upsample_nearest(f, a2, B, 64, h, w, 2);
    free(f);
    orc_conv3x3(a2, params[pi], params[pi + 1], b2, B, 64, 64, h2, w2, 1);
    pi += 2;
    free(a2);
    float *a4 = (float *)malloc(sizeof(float) * lo * 16);
    float *b4 = (float *)malloc(sizeof(float) * lo * 16);
    if (upsample) orc_upsample_bilinear(b2, a4, B, 64, h2, w2, 2); else orc_upsample_nearest(b2, a4, B, 64, h2, w2, 2);
    free(b2);
    orc_conv3x3(a4, params[pi], params[pi + 1], b4, B, 64, 64, H, W, 1);
    pi += 2;
    orc_conv3x3(b4, params[pi], params[pi + 1], a4, B, 64, 64, H, W, 1);
    pi += 2;
    orc_conv3x3(a4, params[pi], params[pi + 1], out_raw, B, 64, Cout, H, W, 0);
    free(a4);
    free(b4);
    /* _recon_image: skip = interpolate(inputs[:, 0:n_mask]) added to the first n_mask channels */
    size_t plane = (size_t)H * W;
    memcpy(out_recon, out_raw, sizeof(float) * B * Cout * plane);
    if (recon_residual && n_mask > 0) {
        float *sk = (float *)malloc(sizeof(float) * plane);
        for (int b = 0; b < B; ++b)
            for (int c = 0; c < n_mask; ++c) {
                const float *src = inputs + ((size_t)b * Cin + c) * h * w;
                if (upsample) orc_upsample_bilinear(src, sk, 1, 1, h, w, 4); else orc_upsample_nearest(src, sk, 1, 1, h, w, 4);
                float *o = out_recon + ((size_t)b * Cout + c) * plane;
                for (size_t i = 0; i < plane; ++i) o[i] = sk[i] + o[i];
            }
        free(sk);
    }
    return 0;
}

/* ------------------------------------------------------------------------------------------
 * Inter-frame post-processing: mainVideoUnshaded.py:460-465, mainGUI.py:594-598, with
 * ScreenSpaceShading.normalize (utils/shading.py:193-207): v / max(||v||, 1e-7).
 * unshaded (C==6): clamp(mask,-1,1), normalize(normal), clamp(depth,0,1), clamp(ao,0,1)
 * shaded: clamp(pred, 0, 1)  (mainVideo.py:416)
 * ---------------------------------------------------------------------------------------- */
static inline float clampf(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }

ORC_API void orc_postprocess(const float *in, float *out, int B, int C, int H, int W, int unshaded)
{
    size_t plane = (size_t)H * W;
    for (int b = 0; b < B; ++b) {
        const float *p = in + (size_t)b * C * plane;
        float *o = out + (size_t)b * C * plane;
        if (!unshaded) {
            for (size_t i = 0; i < (size_t)C * plane; ++i) o[i] = clampf(p[i], 0.f, 1.f);
            continue;
        }
        for (size_t i = 0; i < plane; ++i) {
            o[i] = clampf(p[i], -1.f, 1.f);
            float nx = p[plane + i], ny = p[2 * plane + i], nz = p[3 * plane + i];
            float len = sqrtf(nx * nx + ny * ny + nz * nz);
            if (len < 1e-7f) len = 1e-7f;
            o[plane + i] = nx / len;
            o[2 * plane + i] = ny / len;
            o[3 * plane + i] = nz / len;
            for (int c = 4; c < C; ++c) o[c * plane + i] = clampf(p[c * plane + i], 0.f, 1.f);
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * utils/shading.py:148-191  ScreenSpaceShading.forward.
 * prm[0..2] ambient, [3..5] diffuse, [6..8] specular, [9..11] light dir (normalised),
 * [12..14] material, [15..17] background, [18] ao strength, [19] specular exponent (integer),
 * [20] enable_specular, [21] inverse_ao.
 * in: B x C x H x W (mask, nx, ny, nz, depth[, ao]); out: B x 3 x H x W
 * ---------------------------------------------------------------------------------------- */
ORC_API void orc_shade(const float *in, const float *prm, float *out, int B, int C, int H, int W)
{
    size_t plane = (size_t)H * W;
    float aos = prm[18];
    int expo = (int)prm[19];
    int spec = prm[20] != 0.f, inv = prm[21] != 0.f;
    float sf = (float)(((double)expo + 2.0) / (2.0 * 3.14159265358979323846));
    for (int b = 0; b < B; ++b) {
        const float *p = in + (size_t)b * C * plane;
        float *o = out + (size_t)b * 3 * plane;
        for (size_t i = 0; i < plane; ++i) {
            float m = p[i], nx = p[plane + i], ny = p[2 * plane + i], nz = p[3 * plane + i];
            float ao = 1.f;
            if (C >= 6) {
                float a = p[5 * plane + i];
                if (inv) a = 1.f - a;
                ao = aos * clampf(a, 0.f, 1.f) + (1.f - aos) * 1.f;
            }
            float ldn = prm[9] * nx + prm[10] * ny + prm[11] * nz;
            float diff = fabsf(ldn);
            float sfac = 0.f;
            if (spec) {
                float rz = 2.f * ldn * nz - prm[11]; /* reflect . (0,0,1) */
                float cz = clampf(rz, 0.f, 1.f);
                float pw = 1.f;
                for (int k = 0; k < expo; ++k) pw *= cz;
                sfac = sf * pw;
            }
            float mk = clampf(m * 0.5f + 0.5f, 0.f, 1.f);
            for (int c = 0; c < 3; ++c) {
                float col = 0.f;
                col += prm[c] * prm[12 + c];
                col += (prm[3 + c] * prm[12 + c]) * diff;
                if (spec) col += sfac * prm[6 + c];
                col *= ao;
                float bg = prm[15 + c];
                float v = bg + mk * (col - bg);
                o[c * plane + i] = clampf(v, 0.f, 1.f);
            }
        }
    }
}

ORC_API int orc_version(void) { return 1; }
