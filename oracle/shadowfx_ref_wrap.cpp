// TEST INFRASTRUCTURE ONLY.  C entry point around the reference's own ShadowFX::screenSpaceAmbientOcclusion
// (/root/reference/GPURenderer/ShadowFX.cpp:23-134), which oracle/Makefile (target _ref) compiles in place from
// /root/reference -- no reference source is copied into this repository.
#include "ShadowFX.h"

extern "C" void ref_ssao(int width, int height, const float *input, int inputEntryStride, int inputRowStride, int nx,
                         int ny, int nz, int depthIndex, float *output, int outputEntryStride, int outputRowStride,
                         int outputIndex, int samples, float sampleRadius, float bias, int noiseSamples) {
  ShadowFX::screenSpaceAmbientOcclusion(width, height, input, inputEntryStride, inputRowStride, nx, ny, nz, depthIndex,
                                        output, outputEntryStride, outputRowStride, outputIndex, samples, sampleRadius,
                                        bias, noiseSamples);
}
