/*
 * isr.h -- C ABI of libisr.so: the B200 (sm_100a) implementation of the frame-recurrent 4x
 * EnhanceNet inference path of shamanDevel/IsosurfaceSuperresolution.
 *
 * Conventions (modelled on the reference's own FFI, SuperresolutionNetwork/inference/renderer.py:78-117
 * and GPURendererDirect/GPURendererDirect.cpp:228-246,393-446: plain C functions, int status, the CALLER
 * allocates every tensor and passes raw device pointers, like `render(CUdeviceptr)` there):
 *   - every function returns 0 on success and a negative value on error; isr_last_error() returns the
 *     message of the last error on the calling thread;
 *   - all tensor pointers are CUDA device pointers unless the name ends in _host; fp32 tensors are NCHW
 *     contiguous exactly as the reference's torch tensors;
 *   - launches are asynchronous on `stream` (a cudaStream_t passed as void*; NULL = legacy default stream);
 *     nothing synchronises unless documented; no function allocates device memory except isr_net_create /
 *     isr_net_load_weights (packed weights, a few MB) -- activations live in the caller's workspace;
 *   - there is no CPU path: with no CUDA device every compute entry point fails with ISR_ERR_CUDA.
 *
 * Reference paths are relative to /root/reference/SuperresolutionNetwork.
 */
#ifndef ISR_H_
#define ISR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ISR_OK 0
#define ISR_ERR_ARG (-1)      /* invalid argument (the reference raises AssertionError / ValueError) */
#define ISR_ERR_CUDA (-2)     /* CUDA runtime / driver error, or no device */
#define ISR_ERR_WORKSPACE (-3)/* workspace too small */
#define ISR_ERR_UNSUPPORTED (-4)

/* upsample modes of EnhanceNet (opt.upsample, models/enhancenet.py:40-48) */
#define ISR_UPSAMPLE_NEAREST 0
#define ISR_UPSAMPLE_BILINEAR 1

/* initialImage modes (utils/initial_image.py:24-54) */
#define ISR_INITIAL_ZERO 0
#define ISR_INITIAL_UNSHADED 1
#define ISR_INITIAL_INPUT 2

/* Tensor-core operand schemes of the convolution stack (accumulation is always fp32 in TMEM).
 * "x3" = split operands: A=Ahi+Alo, W=Whi+Wlo, three MMAs (hi*hi + hi*lo + lo*hi) per k-step. */
#define ISR_PREC_BF16 0        /* single-pass bf16 everywhere (fast; misses 1e-2 with random-init weights) */
#define ISR_PREC_BF16X3 1      /* split-bf16 everywhere (safe) */
#define ISR_PREC_MIXED 2       /* split-bf16 in the low-res trunk, single-pass fp16 in the high-res tail (the output
                                * conv uses split fp16 weights stacked along N, which costs nothing: N=32) */
#define ISR_PREC_FP16 3        /* single-pass fp16 everywhere */
#define ISR_PREC_FP16X3 4      /* split-fp16 everywhere */
#define ISR_PREC_MIXED2 5      /* split-bf16 trunk, fp16 tail with split activations (2 MMAs per k-step) */
#define ISR_PREC_MIXED3 6      /* split-bf16 trunk, fp16 tail with split WEIGHTS stacked along N (one N=128 MMA per k-step) */
#define ISR_PREC_MIXED4 7      /* fp16 trunk with split weights stacked along N (one N=128 MMA per k-step), fp16 tail */
#define ISR_PREC_MIXED5 8      /* MIXED4 + the residual stream stored as hi + lo planes (MMAs read hi, the addend is exact) */
#define ISR_PREC_HYBRID 10     /* fp16: preblock + first K residual blocks fully split (x3), remaining blocks hi activations x split
                                * weights (one N=128 MMA per k-step), single-pass fp16 tail; K = tunable trunk_full_blocks (10 = the
                                * whole trunk).  In the fully split residual convs the A_lo x W correction pass runs on 8-bit planes
                                * (e4m3 x e5m2, kind::f8f6f4; tunable trunk_lo8).  DEFAULT of the Python front end and of bench.py */
#define ISR_PREC_MIXED6 9      /* single-pass fp16 everywhere + hi/lo residual stream */

int isr_version(void);
const char *isr_last_error(void);
/* number of kernels launched by this library on the calling process since load (bench.py's gpu_launches) */
uint64_t isr_launch_count(void);

/* ---- memory-bound passes (fp32 NCHW) ------------------------------------------------------------------ */

/* VideoTools.flatten_high (models/videotools.py:8-25): out[b, c*r*r+dy*r+dx, y, x] = in[b, c, y*r+dy, x*r+dx].
 * in: B x C x H x W, out: B x C*r*r x H/r x W/r.  Bit-exact. */
int isr_flatten_high(const float *in, float *out, int B, int C, int H, int W, int r, void *stream);

/* VideoTools.warp_upscale (models/videotools.py:49-87): flow*(-2,+2), bilinear x r, + linspace grid,
 * grid_sample(bilinear, zeros).  align_corners=0 reproduces the installed-torch default (the parity oracle);
 * 1 = the torch-1.0.1 behaviour the reference was written for.  prev/out: B x C x h*r x w*r, flow: B x 2 x h x w. */
int isr_warp_upscale(const float *prev, const float *flow, float *out, int B, int C, int h, int w, int r,
                     int special_mask, int align_corners, void *stream);

/* initialImage (utils/initial_image.py:5-54).  in: B x Cin x h x w, out: B x channels x h*r x w*r. */
int isr_initial_image(const float *in, float *out, int B, int Cin, int h, int w, int channels, int mode,
                      int ao_inverted, int r, void *stream);

/* inter-frame post-processing (mainVideoUnshaded.py:460-465, mainGUI.py:594-598; mainVideo.py:416):
 * unshaded: clamp(mask,-1,1), normal/max(|normal|,1e-7), clamp(depth,0,1), clamp(ao,0,1); shaded: clamp(x,0,1). */
int isr_postprocess(const float *in, float *out, int B, int C, int H, int W, int unshaded, void *stream);

/* ScreenSpaceShading.forward (utils/shading.py:148-191).  params_host: 22 floats on the HOST:
 * [0..2] ambient, [3..5] diffuse, [6..8] specular, [9..11] light direction (normalised), [12..14] material,
 * [15..17] background, [18] ao strength, [19] specular exponent, [20] enable_specular, [21] inverse_ao.
 * gbuf: B x C x H x W (C>=5: mask,nx,ny,nz,depth[,ao]), rgb: B x 3 x H x W. */
int isr_shade(const float *gbuf, const float *params_host, float *rgb, int B, int C, int H, int W, void *stream);

/* isr_shade straight to the display format: rgb8 = B x H x W x 3 interleaved bytes, uint8(clip(rgb * 255, 0, 255)), the
 * conversion mainGUI.py:507 / mainComparisonVideo3.py:589 do on the host after copying the fp32 frame back (12 B/px
 * instead of 3).  H * W must be a multiple of 4; gbuf 16-byte aligned. */
int isr_shade_rgb8(const float *gbuf, const float *params_host, unsigned char *rgb8, int B, int C, int H, int W,
                   void *stream);

/* Temporal post-smoothing of the displayed frame (mainGUI.py:834-849): out = factor * warp_upscale(prev_rgb, flow, r,
 * special_mask = False) + (1 - factor) * cur_rgb in one pass (the GUI's second warp of the previous RGB frame + exponential
 * blend; factor = slider / 100, a double as in the reference: it multiplies by fp32(factor) and fp32(1 - factor)).  prev_rgb / cur_rgb / out: B x C x h*r x w*r, flow: B x 2 x h x w (in-painted). */
int isr_temporal_blend(const float *prev_rgb, const float *flow, const float *cur_rgb, float *out, int B, int C, int h,
                       int w, int r, double factor, void *stream);

/* GUI / video composition (mainGUI.py:603-608,615-636; mainComparisonVideo3.py:540-556): bilinear x4 of the c_low-channel
 * low-res render (RGB, mask, normal, depth, flow, AO, shadow: c_low >= 12 for the unshaded form, >= 4 otherwise);
 * unshaded != 0: channels 0:3 = shading(pred), 3:8 = pred[0:5], 10 = pred[5] (pred = post-processed state, c_pred 5|6);
 * unshaded == 0: channels 0:3 = clamp(pred, 0, 1) (c_pred = 3).  masking != 0: every channel becomes
 * background0 + (up(low[3]) * 0.5 + 0.5) * (value - background0).  low: B x c_low x h x w; pred: B x c_pred x 4h x 4w;
 * out: B x c_low x 4h x 4w.  params_host: as isr_shade (ignored when unshaded == 0, may then be NULL). */
int isr_compose(const float *low, const float *pred, const float *params_host, float *out, int B, int c_low, int c_pred,
                int h, int w, int unshaded, int masking, float background0, void *stream);

/* ---- in front of the network on the interactive path ---------------------------- */

/* G-buffer ingest.  render: the renderer's frame, interleaved != 0: B x h x w x c_render (the wire format of
 * GPURendererDirect/render_kernel.cu:254-265: RGB, mask, nx, ny, nz, depth, flow_x, flow_y, AO, shadow; c_render = 12),
 * interleaved == 0: planar B x c_render x h x w (the tensor inference/loadedmodel.py:70 receives).  One pass writes
 *   low     B x c_low x h x w: the network input slice of loadedmodel.py:84-105.  variant 0 (unshaded): mask*2-1,
 *           normal, depth (5); 1: RGB, mask (4); 2: + depth (5); 3: + normal (7); 4: clamp(channels 0:8, 0, 1) (8);
 *   flow    B x 2 x h x w: channels 8:10 (raw, holes not filled);
 *   hole    B x h x w bytes: 1 where mask == 0 (the in-painting mask of loadedmodel.py:80);
 *   image12 B x 12 x h x w planar copy with the mask moved to [-1, 1] (mainGUI.py:713-716).
 * Any output may be NULL. */
int isr_ingest_gbuffer(const float *render, int interleaved, int B, int c_render, int h, int w, int variant, float *low,
                       float *flow, unsigned char *hole, float *image12, void *stream);
/* channels of `low` for a variant (5, 4, 5, 7, 8), or ISR_ERR_ARG */
int isr_ingest_low_channels(int variant);

/* Flow in-painting on the device, replacing cv.inpaint(flow, mask == 0, 3, INPAINT_NS) of loadedmodel.py:77-82 (no
 * GPU -> CPU -> GPU round trip).  Push-pull pyramid: valid pixels keep their value exactly; a hole takes the bilinear
 * sample of the next coarser level, whose pixels are the means of their valid children.  Not OpenCV's algorithm: the
 * fill is an average of valid neighbours at the scale of the hole: identical to OpenCV's on constant fields, inside
 * the range of the valid flow always, and within 10 % of the flow range of OpenCV's for silhouette-sized holes
 * (tests/test_ingest_oracle.py); arithmetic restated in oracle/oracle.py::inpaint_flow_pushpull.
 * flow, out: B x 2 x h x w; hole: B x h x w bytes; workspace: isr_inpaint_workspace_bytes, 16-byte aligned. */
int isr_inpaint_workspace_bytes(int B, int h, int w, size_t *bytes);
int isr_inpaint_flow(const float *flow, const unsigned char *hole, float *out, int B, int h, int w, void *workspace,
                     size_t workspace_bytes, void *stream);

/* isr_inpaint_flow without a workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid allocation).  Same arguments, kernels and results as isr_inpaint_flow; the pyramid scratch is a
 * static device array inside libisr.so of ISR_INPAINT_STATIC_MAX_BYTES (8 MiB, which every GPU that loads the library
 * holds until the process exits): enough for B * h * w up to about one million low-res pixels (4 x 512 x 512,
 * 1 x 1080 x 960; need = isr_inpaint_workspace_bytes).  Larger problems are refused with ISR_ERR_WORKSPACE.  The call
 * itself calls no cudaMalloc / cudaFree or other allocating routine.  ALL calls share the one array (per device): calls
 * that overlap on the GPU -- different streams, different threads -- race; serialise them or use isr_inpaint_flow.
 * Nothing in the scratch needs clearing between calls (every element is written before it is read within a call). */
#define ISR_INPAINT_STATIC_MAX_BYTES (8u << 20)
int isr_inpaint_flow_static(const float *flow, const unsigned char *hole, float *out, int B, int h, int w, void *stream);

/* Screen-space ambient occlusion of the renderer's post-pass on the device:
 * ShadowFX::screenSpaceAmbientOcclusion, GPURenderer/ShadowFX.cpp:23-134, same argument list and meaning (strided
 * addressing: value(channel, x, y) = buf[channel + entry_stride * (x + row_stride * y)]; the reference calls it in place
 * on the 12-channel frame with normal = 4..6, depth = 7, output = 10, 64 samples, GPURenderer.cpp:716-723).  input and
 * output are DEVICE pointers and may alias when the channels differ.  samples <= 128, noise_samples <= 8.
 * The hemisphere kernel and rotation noise are drawn on the host from std::default_random_engine(42) with the reference's
 * call sequence (ShadowFX.cpp:35-66); isr_ssao_tables returns them (kernel: 3*samples floats, noise: 3*noise_samples^2
 * floats, x-major) for inspection. */
int isr_ssao(int width, int height, const float *input, int input_entry_stride, int input_row_stride, int normal_x_index,
             int normal_y_index, int normal_z_index, int depth_index, float *output, int output_entry_stride,
             int output_row_stride, int output_index, int samples, float sample_radius, float bias, int noise_samples,
             void *stream);
int isr_ssao_tables(int samples, int noise_samples, float *kernel_host, float *noise_host);

/* Dense-grid isosurface ray-caster writing the renderer's 12-channel frame: the arithmetic of
 * custom_iso_kernel, GPURendererDirect/render_kernel.cu:212-266 (fixed-step march, 10-step bisection :175-190, gradient
 * normal, Phong colour, flow from the current / next view-projection matrices, screen-space depth and normal, ray-cast
 * ambient occlusion :109-148) on ONE dense fp32 brick instead of GVDB's sparse atlas (GVDB is a closed binary and the
 * repository ships no volume).  volume: DEVICE nz x ny x nx floats (x fastest), voxel centres at index + 0.5, clamp to
 * edge, trilinear in fp32; the surface is {value >= iso}.  out: DEVICE height x width x 12 interleaved floats
 * (RGB, mask, nx, ny, nz, depth, flow_x, flow_y, AO, shadow; render_kernel.cu:254-265), 16-byte aligned.
 * params: HOST block of ISR_RC_PARAMS floats (indices below; matrices row-major as uploaded by
 * GPURendererDirect.cpp:325-338).  ao_hemisphere: HOST 4 * ao_samples floats, ao_rotations: HOST 64 floats (4 x 4
 * float4, index x%4 + 4*(y%4)); isr_raycast_ao_tables draws both as GPURendererDirect.cpp:146-190 does (512 samples).
 * viewport: HOST minX, minY, maxX, maxY or NULL for the whole frame.  ao_samples in [0, 512]; nx, ny, nz >= 2. */
enum {
  ISR_RC_CAM_POS = 0,     /* 3: ray origin (world) */
  ISR_RC_CAM_U = 3,       /* 3: ray(sx, sy) = normalize(sx * U + sy * V + S), sx = (x + 0.5) / width, sy likewise */
  ISR_RC_CAM_V = 6,
  ISR_RC_CAM_S = 9,
  ISR_RC_BOX_MIN = 12,    /* 3: world position of the brick's index-space origin */
  ISR_RC_VOXEL = 15,      /* world size of one voxel */
  ISR_RC_ISO = 16,        /* isovalue */
  ISR_RC_STEP = 17,       /* march step in voxels (the reference: 0.05, GPURendererDirect.cpp:365) */
  ISR_RC_VIEW_CUR = 18,   /* 16: projection * view of the current camera */
  ISR_RC_VIEW_NEXT = 34,  /* 16: projection * view of the next camera */
  ISR_RC_NORMAL_MAT = 50, /* 16: view matrix (applied to (normal, 0)) */
  ISR_RC_LIGHT = 66,      /* 3: light direction, normalised */
  ISR_RC_AMBIENT = 69,
  ISR_RC_DIFFUSE = 72,
  ISR_RC_SPECULAR = 75,
  ISR_RC_SPEC_EXP = 78,   /* integer-valued */
  ISR_RC_AO_RADIUS = 79,  /* world units */
  ISR_RC_AO_BIAS = 80,    /* world units the AO origin steps back along the primary ray (1e-3) */
  ISR_RC_PARAMS = 81
};
int isr_raycast_iso(const float *volume, int nx, int ny, int nz, const float *params_host, const float *ao_hemisphere_host,
                    const float *ao_rotations_host, int ao_samples, int binary_steps, const int *viewport_host,
                    const float *skip, float *out, int width, int height, void *stream);
/* Optional empty-space grids for isr_raycast_iso (`skip`, DEVICE, may be NULL): the max of the brick over each 8^3 and
 * each 2^3 macro-cell plus the trilinear footprint, and over the 8 voxels of every interpolation cell, built once per volume (GVDB's brick topology plays this role in the
 * reference).  A cell whose max is below iso cannot contain a sample >= iso, so the march steps over it; the samples
 * that are evaluated and the frame are IDENTICAL to the plain march. */
int isr_raycast_skip_bytes(int nx, int ny, int nz, size_t *bytes);
int isr_raycast_build_skip(const float *volume, int nx, int ny, int nz, float *skip, size_t skip_bytes, void *stream);
int isr_raycast_ao_tables(float *hemisphere_host /* 4 * 512 */, float *rotations_host /* 64 */);

/* ---- EnhanceNet generator (models/enhancenet.py:8-26,92-145) ------------------------------------------- */

typedef struct isr_net isr_net;

/* Mirrors EnhanceNet(upscale_factor=4, input_channels, channel_mask, output_channels, opt): n_mask =
 * len(channel_mask) (the skip uses inputs[:, 0:n_mask], enhancenet.py:64-66), upsample = opt.upsample,
 * recon_residual = (opt.reconType == 'residual').  useBN is not supported (fold it first). */
int isr_net_create(isr_net **net, int input_channels, int output_channels, int n_mask, int upsample,
                   int recon_residual);
void isr_net_destroy(isr_net *net);

/* params: 50 DEVICE pointers to fp32 tensors in state_dict order: preblock.0.{weight,bias},
 * blocks.{0..9}.{0,2}.{weight,bias}, postblock.{1,4,6,8}.{weight,bias} (OIHW).  Packs them into the tensor-core
 * layouts ([part][tap][Cout][Cin] bf16/fp16 hi+lo).  Synchronises `stream` before returning. */
int isr_net_load_weights(isr_net *net, const float *const *params, int n_params, void *stream);

/* bytes of caller-provided device workspace needed by isr_net_forward* for a batch of B low-res frames h x w */
int isr_net_workspace_bytes(const isr_net *net, int B, int h, int w, int precision, size_t *bytes);

/* EnhanceNet.forward(inputs) -> (recon, raw)  (models/enhancenet.py:136-145).
 * inputs: B x input_channels x h x w.  out_recon / out_raw: B x output_channels x 4h x 4w (either may be NULL).
 * out_state (optional): post-processed recon (isr_postprocess; unshaded iff output_channels==6 semantics flag). */
int isr_net_forward(isr_net *net, const float *inputs, float *out_recon, float *out_raw, float *out_state,
                    int state_unshaded, int B, int h, int w, int precision, void *workspace, size_t workspace_bytes,
                    void *stream);

/* One frame of the recurrence (mainVideoUnshaded.py:419-465 / inference/loadedmodel.py:70-121) with the previous-frame
 * path fused: network input = cat(low, flatten_high(warp_upscale(prev, flow, 4, special_mask), 4)) is produced
 * directly in the convolution's NHWC 16-bit layout, never materialised in fp32.
 * low: B x c_low x h x w; prev: B x c_prev x 4h x 4w (already the warped/initial image if flow==NULL);
 * flow: B x 2 x h x w or NULL (= prev is used as is).  c_low + 16*c_prev must equal input_channels. */
int isr_net_forward_recurrent(isr_net *net, const float *low, const float *prev, const float *flow, int c_low,
                              int c_prev, int special_mask, int align_corners, float *out_recon, float *out_raw,
                              float *out_state, int state_unshaded, int B, int h, int w, int precision,
                              void *workspace, size_t workspace_bytes, void *stream);

/* Same as isr_net_forward_recurrent but low_host / flow_host are (pinned) HOST buffers copied to the device inside
 * the call, and recon_host (B x Cout x 4h x 4w, HOST) receives the prediction; the recurrent state stays on the
 * device in state_dev (read as `prev` when has_prev != 0, then overwritten).  Synchronises `stream` before returning.
 * staging_dev must hold B*(c_low+2)*h*w + B*Cout*16*h*w floats. */
int isr_net_forward_recurrent_host(isr_net *net, const float *low_host, const float *flow_host, float *state_dev,
                                   int has_prev, int initial_mode, int c_low, int c_prev, int special_mask,
                                   float *recon_host, int B, int h, int w, int precision, void *workspace,
                                   size_t workspace_bytes, float *staging_dev, void *stream);

/* ---- pipelined host-buffer sequence API ------------------------------------------------------------------------
 * The frame recurrence of one batch of sequences with HOST input / output buffers (what a renderer process that owns host
 * frames calls; inference/loadedmodel.py:70-121 per frame), software-pipelined over three internal streams: the H2D copy of
 * frame j+1 and the D2H copy of prediction j overlap the kernels of frame j+1.  The recurrent state never leaves the
 * device.  Host buffers must be pinned for the copies to overlap.  No device memory is allocated: the caller passes the
 * network workspace, the state tensor (B x c_prev x 4h x 4w) and `depth` staging slots. */
typedef struct isr_pipe isr_pipe;
/* floats of device staging needed for `depth` slots of (low + flow + prediction) */
int isr_pipe_staging_floats(const isr_net *net, int c_low, int B, int h, int w, int depth, size_t *floats);
int isr_pipe_create(isr_pipe **pipe, isr_net *net, int c_low, int c_prev, int special_mask, int initial_mode, int B, int h,
                    int w, int precision, void *workspace, size_t workspace_bytes, float *state_dev, float *staging_dev,
                    size_t staging_floats, int depth);
/* Enqueues one frame and returns without waiting for it (it blocks only until the slot it reuses has been drained, i.e.
 * until frame j-depth is in host memory).  flow_host == NULL starts a new sequence: the state is initialised with
 * initialImage(low, c_prev, initial_mode).  recon_host (B x Cout x 4h x 4w) is complete after isr_pipe_wait(). */
int isr_pipe_submit(isr_pipe *pipe, const float *low_host, const float *flow_host, float *recon_host);
/* Blocks until every submitted frame's prediction is in host memory (and the state is final on the device). */
/* Display form of the pipeline: after isr_pipe_enable_display (shading parameters as isr_shade; rgb8_staging_dev: depth x
 * B x 4h x 4w x 3 bytes on the device) isr_pipe_submit_display enqueues a frame whose host-side result is the shaded
 * 8-bit RGB frame (B x 4h x 4w x 3, what the GUI / video writers of the reference consume) instead of the fp32
 * prediction: 3 instead of 24 bytes per output pixel cross PCIe; the prediction itself is not written at all. */
int isr_pipe_enable_display(isr_pipe *pipe, const float *params_host, unsigned char *rgb8_staging_dev, size_t bytes);
int isr_pipe_submit_display(isr_pipe *pipe, const float *low_host, const float *flow_host, unsigned char *rgb8_host);
/* Switches the recurrent-state tensor (B x c_prev x 4h x 4w, device) used by the frames submitted from now on: one pipeline
 * (one workspace, one set of staging slots) can serve several chunks of independent sequences in turn -- BASELINE.json
 * configs[4] runs 64 sequences per GPU as 16 chunks of 4.  Stream-ordered: frames already submitted keep their state. */
int isr_pipe_set_state(isr_pipe *pipe, float *state_dev);
/* 16-bit form of the full prediction: after isr_pipe_enable_f16 (f16_staging_dev: depth x B x Cout x 4h x 4w x 2 bytes on
 * the device) isr_pipe_submit_f16 enqueues a frame whose host-side result is the prediction rounded to IEEE fp16 (round to
 * nearest even; |prediction| < 16 for this network, so the rounding error is <= 2^-8 * 2^-3 relative to the 1e-2 parity
 * bar): 12 instead of 24 bytes per output pixel cross the host link.  The recurrent state stays fp32 on the device. */
int isr_pipe_enable_f16(isr_pipe *pipe, void *f16_staging_dev, size_t bytes);
int isr_pipe_submit_f16(isr_pipe *pipe, const float *low_host, const float *flow_host, void *recon_f16_host);
int isr_pipe_wait(isr_pipe *pipe);
void isr_pipe_destroy(isr_pipe *pipe);

/* ---- test / bench hooks --------------------------------------------------------------------------------- */

/* One conv3x3(pad 1) 64->cout (cout in {64, <=16}) through the tensor-core kernel, fp32 NCHW in and out, for unit tests:
 * x: B x 64 x H x W, weight: cout x 64 x 3 x 3, bias: cout, addend: B x 64 x H x W or NULL (only cout==64),
 * out: B x cout x H x W.  fmt: 0 bf16, 1 fp16; n_act_parts / n_w_parts in {1,2}.  workspace: isr_debug_conv_workspace_bytes. */
int isr_debug_conv3x3(const float *x, const float *weight, const float *bias, const float *addend, float *out, int B,
                      int H, int W, int cout, int relu, int fmt, int n_act_parts, int n_w_parts, void *workspace,
                      size_t workspace_bytes, void *stream);
int isr_debug_conv_workspace_bytes(int B, int H, int W, size_t *bytes);

/* Tunables for tests, experiments and bench ablations (defaults in brackets).  isr_set_tunable changes the process-wide
 * DEFAULTS, of which every isr_net takes a snapshot at isr_net_create; isr_net_set_tunable changes ONE handle (its cached
 * plans are dropped), so handles in different threads can run different settings.  Both return the previous value, or a
 * negative ISR_ERR_* for an unknown key / value ("raycast_two_pass" is process-wide only):
 *   "halo_pitch" (10 | [16]), "base_offset_mode" ([0] | 1), "max_ctas" ([0] = #SMs)   operand addressing, DESIGN.md 4.1
 *   "cta_pair" (0 | [1])             CTA-pair (cta_group::2) form of the 64->64 conv
 *   "stack_w" (0 | [1])              hi/lo weight planes as one N=128 B operand
 *   "pdl" (0 | [1])                  programmatic dependent launch between consecutive convs
 *   "fuse_up" (0 | [1])              bilinear 2x upsample fused into the following conv (four phase convolutions)
 *   "up2_wide" (0 | [1])             ... with the two x-phases of an output row stacked as one N = 128 MMA operand
 *                                    (tensor floor: 64 cycles per two phases instead of 2 x 43); 0 = four N = 64 phases
 *   "trunk_lo8" (0 | [1])            ISR_PREC_HYBRID: 8-bit low activation plane + 8-bit (kind::f8f6f4) correction pass in the
 *                                    fully split trunk convs (0 = fp16 low plane, 25 % more trunk time, same parity)
 *   "alt_order" (0 | [1])            consecutive convs walk their tiles in alternating directions, so that each starts on the
 *                                    part of its input that the previous launch left in L2
 *   "input_lo8" (0 | [1])            with trunk_lo8: the packed network input is planar fp16 hi + 8-bit low plane as well (the
 *                                    preblock runs the 8-bit correction pass); 0 = interleaved fp16 hi | lo
 *   "stream_lo8" (0 | [1])           with trunk_lo8: the residual stream is stored as hi + that 8-bit low plane only (0 = it
 *                                    also carries an fp16 low plane for the residual addend: +67 % stream traffic)
 *   "trunk_full_blocks" (0..10, [10]) residual blocks of ISR_PREC_HYBRID that run fully split (4 is 11 % faster per
 *                                    step and misses the 1e-2 bar at full frame sizes: profiles/r02_parity_margin.json)
 *   "final_stacked" (0 | 1 | 2 | [3]) output conv: generic / tap-stacked (shared-memory gather) / CTA-pair N=16 /
 *                                    column-stacked with a register gather (isr_conv_final2.cuh)
 *   "final_issuers" (1 | [3])        MMA-issuing warps of the tap-stacked output conv
 *   "final_groups" ([3] | 4)         epilogue groups (= accumulator buffers = issuing warps) of the column-stacked output conv
 *                                    (4 spills in the epilogue: 0.65 against 0.61 ms)
 *   "final_w_parts" ([0] = scheme | 1 | 2)   weight planes of the output conv
 *   "raycast_two_pass" (0 | [1] = from 8 AO rays on | 2)   ambient occlusion of the ray-caster as a second pass
 *   "pdl_small" ([0] | 1)            process-wide: packing / band / shading / composition kernels launched programmatically
 *                                    too (measured 8 % slower per step: tools/pdl_experiments.py)
 *   "compose_variant" (1 | [2])      process-wide: GUI composition kernel (2 = 4 pixels per thread, bit-identical to 1)
 *   "pack_variant" (1 | [2] | 3)     process-wide: network-input packing kernel (2 = per-column form, bit-identical to 1;
 *                                    3 = the same with 16-byte tap loads, measured slower)
 *   "dbg"                            bit mask of bench experiments (tools/), 0 in production */
int isr_set_tunable(const char *key, int value);
int isr_net_set_tunable(isr_net *net, const char *key, int value);

/* Times `iters` back-to-back launches of ONE conv of the network's plan with CUDA events on `stream`; returns the average
 * milliseconds per launch in *ms.  scale 4: the dominant kernel, the 64->64 conv at 4h x 4w (bench.py's roofline leg);
 * 1 / 2: first / second conv of the first residual block (h x w; the second carries the residual addend); 3: the fused
 * upsample + conv that produces the 4h x 4w tensor (bilinear nets); 5: the output conv (writes into scratch).  The handle's
 * "dbg" tunable is applied to the timed launches only. */
int isr_net_time_conv(isr_net *net, int B, int h, int w, int precision, int scale, void *workspace,
                      size_t workspace_bytes, int iters, float *ms, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* ISR_H_ */
