// isr_internal.h -- declarations shared between the translation units of libisr.so (not part of the C ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace isr {

void count_launch();

// Programmatic dependent launch for the small kernels between the convolutions (packing, band kernels, output conv,
// shading, composition): launched with programmatic stream serialization, they start while the previous kernel drains and
// call pdl_prologue() first -- griddepcontrol.launch_dependents (let the NEXT kernel start launching) and
// griddepcontrol.wait (block until everything before this kernel has completed and is visible).  Without it every such
// boundary of the 36-launch frame pays ~3.6 us of launch latency (tools/launch_floor.py: 10.2 vs 6.6 us per launch).
extern int g_pdl_small;   // process-wide switch (tunable "pdl_small"): 1 = the small kernels are launched programmatically
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_prologue() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = g_pdl_small ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

cudaError_t launch_flatten_high(const float *in, float *out, int B, int C, int H, int W, int r, cudaStream_t s);
cudaError_t launch_warp_upscale(const float *prev, const float *flow, float *out, int B, int C, int h, int w, int r,
                                int special_mask, int align_corners, cudaStream_t s);
cudaError_t launch_temporal_blend(const float *prev, const float *flow, const float *cur, float *out, int B, int C, int h,
                                  int w, int r, float fa, float fb, cudaStream_t s);
cudaError_t launch_f32_to_f16(const float *in, void *out, size_t n, cudaStream_t s);
cudaError_t launch_initial_image(const float *in, float *out, int B, int Cin, int h, int w, int channels, int mode,
                                 int ao_inverted, int r, cudaStream_t s);
cudaError_t launch_postprocess(const float *in, float *out, int B, int C, int H, int W, int unshaded, cudaStream_t s);
cudaError_t launch_shade(const float *in, const float *params22_host, float *out, int B, int C, int H, int W,
                         cudaStream_t s);
cudaError_t launch_compose(const float *low, const float *pred, const float *params22_host, float *out, int B, int c_low,
                           int c_pred, int h, int w, int unshaded, int masking, float bg0, cudaStream_t s);
cudaError_t launch_pack_recurrent(const float *low, const float *prev, const float *flow, uint16_t *g0, uint16_t *g1,
                                  int B, int c_low, int c_prev, int h, int w, int special_mask, int align_corners,
                                  int nparts, int is_bf16, cudaStream_t s);
int set_compose_variant(int v);   // 1 | 2: scalar / 4-pixels-per-thread composition kernel; returns the old value
int set_pack_variant(int v);   // 1 | 2: which network-input packing kernel launch_pack_recurrent uses; returns the old value
cudaError_t launch_pack_nchw(const float *in, uint16_t *g0, uint16_t *g1, int B, int cin, int h, int w, int nparts,
                             int is_bf16, cudaStream_t s);
cudaError_t launch_unpack_nhwc(const uint16_t *in, float *out, int B, int H, int W, int nparts, int is_bf16,
                               cudaStream_t s);
cudaError_t launch_upsample2x(const uint16_t *in, uint16_t *out, int B, int h, int w, int in_parts, int out_parts,
                              int in_bf16, int out_bf16, int bilinear, cudaStream_t s);
// out layout: [fmt 2 (bf16, fp16)][part 2 (hi, lo)][tap 9][group G][npad][64]
cudaError_t launch_pack_weights(const float *w, uint16_t *out, int cout, int cin, int npad, int G, cudaStream_t s);
// tap-stacked output-conv weights: [fmt 2][part 2][nrows (n = tap*cp + co)][64]
cudaError_t launch_pack_weights_final(const float *w, uint16_t *out, int cout, int cp, int nrows, cudaStream_t s);

// 8-bit (e5m2, scaled 2^-6) weight plane of a 64 -> 64 layer: [tap 9][n 64 (permuted like pack_weights)][k 64] bytes
cudaError_t launch_pack_weights_lo8(const float *w, uint8_t *out, int cin, int G, cudaStream_t s);
// column-stacked output-conv weights (isr_conv_final2.cuh): [fmt 2][ky 3][part 2][nk (n = kx*cp + co)][64]
cudaError_t launch_pack_weights_final2(const float *w, uint16_t *out, int cout, int cp, int nk, cudaStream_t s);

cudaError_t launch_shade_rgb8(const float *in, const float *params22, unsigned char *out, int B, int C, int H, int W,
                              cudaStream_t s);
// isr_ingest.cu
int ingest_low_channels(int variant);
cudaError_t launch_ingest(const float *render, int interleaved, int c_render, float *low, float *flow,
                          unsigned char *hole, float *image12, int B, int h, int w, int variant, cudaStream_t s);
size_t inpaint_workspace_floats(int B, int h, int w);
cudaError_t launch_inpaint_flow(const float *flow, const unsigned char *hole, float *out, int B, int h, int w, float *ws,
                                cudaStream_t s);

// isr_up2.cu: fused bilinear-2x-upsample + conv support
// out: [fmt 2 (bf16, fp16)][phase 4][tap 9][n 64][k 64] composite weights of a 64 -> 64 layer
cudaError_t launch_pack_weights_up2(const float *w, uint16_t *out, cudaStream_t s);
cudaError_t launch_band_upsample(const uint16_t *in, uint16_t *rows, uint16_t *cols, int B, int hl, int wl, int in_parts,
                                 int is_bf16, cudaStream_t s);
cudaError_t launch_band_scatter(const uint16_t *rows, const uint16_t *cols, uint16_t *out, int B, int H2, int W2, int parts,
                                cudaStream_t s);

// isr_ssao.cu: hemisphere kernel (<= 128 samples) and rotation noise (<= 8 x 8), passed by value to the kernel
struct SsaoTables {
  float kernel[3 * 128];
  float noise[3 * 64];
};
cudaError_t launch_ssao(const float *input, int es, int rs, int inx, int iny, int inz, int idepth, float *output, int oes,
                        int ors, int oidx, int width, int height, int samples, float radius, float bias, int ns,
                        const SsaoTables &tab, cudaStream_t s);

// isr_raycast.cu: everything the ray-caster needs, passed by value (about 8.6 KB of kernel parameters)
struct RaycastArgs {
  int width, height, nx, ny, nz, binary_steps, ao_samples, ao_two_pass;
  int viewport[4];
  float p[81];                // include/isr.h ISR_RC_*
  float hemisphere[4 * 512];
  float rotations[4 * 16];
};
cudaError_t launch_raycast_iso(const float *vol, const float *skip, float *out, const RaycastArgs &args, cudaStream_t s);
size_t raycast_skip_floats(int nx, int ny, int nz);   // coarse (8^3) + fine (2^3) macro-cell grids
cudaError_t launch_raycast_skip(const float *vol, float *skip, int nx, int ny, int nz, cudaStream_t s);

}  // namespace isr
