// isr_api.cu -- host side of libisr.so: the C ABI of include/isr.h, the EnhanceNet execution plan (workspace carving,
// TMA tensor maps, per-layer conv launches) and the unit-test hooks.
//
// Reference paths are relative to /root/reference/SuperresolutionNetwork.
#include <cuda.h>
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/isr.h"
#include "isr_conv.cuh"
#include "isr_internal.h"

namespace isr {

static thread_local std::string g_err;
static std::atomic<uint64_t> g_launches{0};
static int g_halo_pitch = 10;
static int g_base_offset_mode = 0;
static int g_max_ctas = 0;

void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

static int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}
#define ISR_CUDA(expr)                                                                          \
  do {                                                                                          \
    cudaError_t e_ = (expr);                                                                    \
    if (e_ != cudaSuccess) return fail(ISR_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
  } while (0)

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- TMA tensor maps -------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// NHWC 16-bit activation tensor [B][H][W][ctot]; box = 64 channels x box_w x box_h pixels of one image
static int make_act_map(CUtensorMap *m, const void *ptr, int B, int H, int W, int ctot, int box_w, int box_h,
                        int is_bf16) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(ctot), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(B)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(ctot) * 2, static_cast<cuuint64_t>(W) * ctot * 2,
                           static_cast<cuuint64_t>(H) * W * ctot * 2};
  cuuint32_t box[4] = {64, static_cast<cuuint32_t>(box_w), static_cast<cuuint32_t>(box_h), 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                   const_cast<void *>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled(activations) failed: %d", (int)r);
  return ISR_OK;
}
// packed weights [rows][64] 16-bit; box = 64 x n rows
static int make_w_map(CUtensorMap *m, const void *ptr, int rows, int n, int is_bf16) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {64, static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {128};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(n)};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                   const_cast<void *>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled(weights) failed: %d", (int)r);
  return ISR_OK;
}

// ---- conv launch -----------------------------------------------------------------------------------------------
struct ConvLaunch {
  CUtensorMap tmA, tmW;
  ConvParams p;
  int n;  // 64 (EPI_STORE16) or 16 (EPI_FINAL)
  size_t smem;
  int grid;
};

constexpr size_t kSmemLimit = 232448;  // 227 KB per CTA
constexpr size_t kStaticSmem = 512;    // s_bias + slack

static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
      n = 148;
  }
  return n;
}

static int setup_kernel_attrs() {
  static bool done = false;
  if (done) return ISR_OK;
  ISR_CUDA(cudaFuncSetAttribute(conv3x3_umma_kernel<64, EPI_STORE16>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(kSmemLimit - kStaticSmem)));
  ISR_CUDA(cudaFuncSetAttribute(conv3x3_umma_kernel<16, EPI_FINAL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(kSmemLimit - kStaticSmem)));
  done = true;
  return ISR_OK;
}

// fills geometry, smem plan and tensor maps of one conv launch; epilogue fields are set by the caller
static int init_conv(ConvLaunch *L, int n, const void *act, int B, int H, int W, int n_act_parts, const void *wpacked,
                     int w_rows_total, int w_row0, int w_row_stride, int n_w_parts, int is_bf16) {
  memset(&L->p, 0, sizeof L->p);
  L->n = n;
  ConvParams &p = L->p;
  p.B = B; p.H = H; p.W = W;
  p.tiles_x = (W + kTileW - 1) / kTileW;
  p.tiles_y = (H + kTileH - 1) / kTileH;
  p.num_tiles = p.tiles_x * p.tiles_y * B;
  p.pitch = g_halo_pitch;
  p.base_offset_mode = g_base_offset_mode;
  p.n_act_parts = n_act_parts;
  p.n_w_parts = n_w_parts;
  p.is_bf16 = is_bf16;
  p.w_row0 = w_row0;
  p.w_row_stride = w_row_stride;
  const size_t w_bytes = align_up(static_cast<size_t>(n_w_parts) * 9 * n * 128, 1024);
  p.stage_bytes = static_cast<int>(align_up(static_cast<size_t>(p.pitch) * kHaloH * 128, 1024));
  const size_t avail = kSmemLimit - kStaticSmem - 1024 - w_bytes - 256;
  int stages = static_cast<int>(avail / p.stage_bytes);
  if (stages > 8) stages = 8;
  if (stages < 2) return fail(ISR_ERR_UNSUPPORTED, "conv smem plan: fewer than 2 halo stages fit");
  p.num_stages = stages;
  L->smem = 1024 + w_bytes + static_cast<size_t>(stages) * p.stage_bytes + 256;
  int cap = g_max_ctas > 0 ? g_max_ctas : num_sms();
  L->grid = p.num_tiles < cap ? p.num_tiles : cap;
  int rc = make_act_map(&L->tmA, act, B, H, W, 64 * n_act_parts, p.pitch, kHaloH, is_bf16);
  if (rc != ISR_OK) return rc;
  return make_w_map(&L->tmW, wpacked, w_rows_total, n, is_bf16);
}

static int run_conv(const ConvLaunch &L, cudaStream_t s) {
  if (L.n == 64)
    conv3x3_umma_kernel<64, EPI_STORE16><<<L.grid, kConvThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
  else
    conv3x3_umma_kernel<16, EPI_FINAL><<<L.grid, kConvThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
  count_launch();
  ISR_CUDA(cudaGetLastError());
  return ISR_OK;
}

// ---- precision schemes -----------------------------------------------------------------------------------------
struct Scheme {
  int bf16, na, nw;  // operand format, activation parts, weight parts
};
static bool get_schemes(int precision, Scheme *trunk, Scheme *tail) {
  switch (precision) {
    case ISR_PREC_BF16: *trunk = {1, 1, 1}; *tail = {1, 1, 1}; return true;
    case ISR_PREC_BF16X3: *trunk = {1, 2, 2}; *tail = {1, 2, 2}; return true;
    case ISR_PREC_MIXED: *trunk = {1, 2, 2}; *tail = {0, 1, 1}; return true;
    case ISR_PREC_FP16: *trunk = {0, 1, 1}; *tail = {0, 1, 1}; return true;
    case ISR_PREC_FP16X3: *trunk = {0, 2, 2}; *tail = {0, 2, 2}; return true;
    case ISR_PREC_MIXED2: *trunk = {1, 2, 2}; *tail = {0, 2, 1}; return true;
  }
  return false;
}

}  // namespace isr
