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
#include <mutex>
#include <random>
#include <string>
#include <vector>

#include "../../include/isr.h"
#include "isr_conv.cuh"
#include "isr_conv_final.cuh"
#include "isr_conv_final2.cuh"
#include "isr_conv_pair.cuh"
#include "isr_internal.h"

namespace isr {

static thread_local std::string g_err;
static std::atomic<uint64_t> g_launches{0};
// Tunables of the network plans.  `g_tun` holds the process-wide DEFAULTS (isr_set_tunable); every isr_net takes a snapshot
// at isr_net_create and is from then on independent of them (isr_net_set_tunable changes one handle), so two handles in two
// threads can run different settings and a test that dies between two isr_set_tunable calls cannot alter existing handles.
struct Tunables {
  int halo_pitch = 10;
  int base_offset_mode = 0;
  int max_ctas = 0;
  int dbg = 0;             // ConvParams::dbg of isr_net_time_conv launches (bench experiments)
  int cta_pair = 1;        // 64-wide convs: CTA-pair kernel (tcgen05 cta_group::2, isr_conv_pair.cuh)
  int final_stacked = 3;   // output conv (Cout <= 8): 3 = column-stacked kernel with the register gather (isr_conv_final2.cuh),
                           // 1 = tap-stacked kernel with the shared-memory gather (isr_conv_final.cuh), 2 = CTA-pair N=16
                           // kernel (measured slower: 1.24 vs 0.92 ms), 0 = generic single-CTA N=16 kernel
  int stack_w = 1;         // split-precision 64-wide convs: [W_hi ; W_lo] as one N=128 MMA operand
  int fuse_up = 1;         // bilinear nets: upsample + conv run as one phase-decomposed kernel (isr_up2.cu)
  int up2_wide = 1;        // ... with the two x-phases of an output row as one N = 128 MMA operand (0: four N = 64 phases)
  int final_w_parts = 0;   // output conv weight planes: 0 = the scheme's choice, 1 = single, 2 = hi + lo
  int final_issuers = 3;   // MMA-issuing warps of the tap-stacked output conv (1 | 3)
  int final_groups = 3;    // column-stacked output conv: epilogue groups (3 | 4)
  int pdl = 1;             // CTA-pair convs are launched with programmatic stream serialization
  int trunk_lo8 = 1;       // ISR_PREC_HYBRID: the low activation plane of the fully split trunk convs is 8-bit (e4m3 * 2^6)
                           // and the A_lo x W correction pass runs as kind::f8f6f4 against e5m2 weights (half the MMA
                           // time of the fp16 pass); 0 = the fp16 low plane
  int alt_order = 1;       // consecutive convs walk their tiles in alternating directions (ConvParams::reverse): each starts on
                           // the part of its input that is still in L2
  int input_lo8 = 1;       // with trunk_lo8: the packed network input is hi + 8-bit low plane too (the preblock runs the 8-bit
                           // correction pass like the residual blocks); 0 = interleaved fp16 hi | lo, fp16 correction pass
  int stream_lo8 = 1;      // with trunk_lo8: the residual stream itself is kept as hi + the 8-bit low plane (the planes the convs
                           // read anyway); 0 = it also carries an fp16 low plane for the residual addend (320 instead of
                           // 192 B per pixel and tensor: the convs with an addend were HBM-bound, 156 vs 119 us)
  int trunk_full_blocks = 10;  // ISR_PREC_HYBRID: residual blocks (and the preblock) that run fully split (x3).  10 = the whole
                               // trunk: at realistic frame sizes K = 4 does NOT keep the 1e-2 bar (1.04e-2 on frame 0 of a
                               // 2 x 256^2 crop, 7.0e-3 worst over 6 weight seeds; K = 10: 6.7e-3 / 4.6e-3 --
                               // profiles/r02_recurrence_drift.json, r02_parity_margin.json); K < 10 is an opt-in
};
static Tunables g_tun;
static std::mutex g_tun_mutex;
static int g_raycast_two_pass = 1;   // (process-wide, not a network tunable) AO of the ray-caster as a second pass: 0 never, 1 auto, 2 always

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

// the 8-bit low plane of an activation tensor with `pitch` bytes per pixel: byte-typed, box = 64 bytes x box_w x box_h
static int make_act8_map(CUtensorMap *m, const void *ptr, int B, int H, int W, int pitch, int box_w, int box_h) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[4] = {static_cast<cuuint64_t>(pitch), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(B)};
  cuuint64_t strides[3] = {static_cast<cuuint64_t>(pitch), static_cast<cuuint64_t>(W) * pitch,
                           static_cast<cuuint64_t>(H) * W * pitch};
  cuuint32_t box[4] = {64, static_cast<cuuint32_t>(box_w), static_cast<cuuint32_t>(box_h), 1};
  cuuint32_t es[4] = {1, 1, 1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void *>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled(8-bit activations) failed: %d", (int)r);
  return ISR_OK;
}
// 8-bit weight plane [rows][64] bytes; box = 64 x n rows
static int make_w8_map(CUtensorMap *m, const void *ptr, int rows, int n) {
  EncodeTiledFn enc = get_encode();
  if (enc == nullptr) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[2] = {64, static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {64};
  cuuint32_t box[2] = {64, static_cast<cuuint32_t>(n)};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(ptr), dims, strides, box, es,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(ISR_ERR_CUDA, "cuTensorMapEncodeTiled(8-bit weights) failed: %d", (int)r);
  return ISR_OK;
}

// ---- conv launch -----------------------------------------------------------------------------------------------
struct ConvLaunch {
  CUtensorMap tmA, tmW;
  CUtensorMap tmA8, tmW8;   // 8-bit low activation plane / 8-bit weight plane (copies of tmA / tmW when unused)
  ConvParams p;
  int n;  // 64 (EPI_STORE16) or 16 (EPI_FINAL)
  int cp = 0;  // > 0: tap-stacked output conv (conv3x3_final_kernel<cp>)
  int final2 = 0;  // 1: column-stacked output conv (conv3x3_final2_kernel<cp>)
  int groups = 3;  // conv3x3_final2_kernel: epilogue groups (= accumulator buffers = MMA-issuing warps), 3 | 4
  int pair = 0;  // 1: CTA-pair kernel (conv3x3_pair_kernel)
  int ph = 1;    // fused upsample + conv: 4 = four N=64 output phases per low-res tile, 2 = two N=128 phase pairs
  int stack = 0; // split weights stacked as one N=128 operand
  int pdl = 1;   // programmatic dependent launch
  size_t smem;
  int grid;
};

constexpr size_t kSmemLimit = 232448;  // 227 KB per CTA
constexpr size_t kStaticSmem = 512;    // s_bias + slack

constexpr int kMaxDevices = 64;

static int num_sms() {
  static std::atomic<int> cache[kMaxDevices];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return 148;
  int n = cache[dev].load(std::memory_order_relaxed);
  if (n == 0) {
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
    cache[dev].store(n, std::memory_order_relaxed);
  }
  return n;
}

// cudaFuncSetAttribute applies to the device that is current when it is called: the opt-in to 227 KB of dynamic shared
// memory is made once PER DEVICE (a process may drive several GPUs through several handles), under a mutex
template <typename K>
static cudaError_t set_max_smem(K kernel) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kSmemLimit - kStaticSmem));
}
static int setup_kernel_attrs() {
  static std::mutex mu;
  static bool done[kMaxDevices] = {false};
  int dev = 0;
  ISR_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(ISR_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (done[dev]) return ISR_OK;
  ISR_CUDA(set_max_smem(conv3x3_umma_kernel<64, EPI_STORE16, 0>));
  ISR_CUDA(set_max_smem(conv3x3_umma_kernel<64, EPI_STORE16, 1>));
  ISR_CUDA(set_max_smem(conv3x3_umma_kernel<16, EPI_FINAL, 0>));
  ISR_CUDA(set_max_smem(conv3x3_umma_kernel<16, EPI_FINAL, 1>));
  ISR_CUDA(set_max_smem(conv3x3_pair_kernel<0>));
  ISR_CUDA(set_max_smem((conv3x3_pair_kernel<0, 64, EPI_STORE16, 1, false>)));
  ISR_CUDA(set_max_smem((conv3x3_pair_kernel<1, 64, EPI_STORE16, 1, false>)));
  ISR_CUDA(set_max_smem(conv3x3_pair_kernel<1>));
  ISR_CUDA(set_max_smem((conv3x3_pair_kernel<1, 64, EPI_STORE16, 1, true, 8>)));
  ISR_CUDA(set_max_smem(conv3x3_pair_kernel<0, 64, EPI_STORE16, 4>));
  ISR_CUDA(set_max_smem(conv3x3_pair_kernel<0, 64, EPI_STORE16, 2>));
  ISR_CUDA(set_max_smem(conv3x3_pair_kernel<0, 16, EPI_FINAL>));
  ISR_CUDA(set_max_smem(conv3x3_pair_kernel<1, 16, EPI_FINAL>));
  ISR_CUDA(set_max_smem(conv3x3_final_kernel<4>));
  ISR_CUDA(set_max_smem(conv3x3_final_kernel<6>));
  ISR_CUDA(set_max_smem(conv3x3_final_kernel<8>));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<4, 0, 3>)));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<6, 0, 3>)));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<8, 0, 3>)));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<4, 1, 3>)));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<6, 1, 3>)));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<4, 1, 4>)));
  ISR_CUDA(set_max_smem((conv3x3_final2_kernel<6, 1, 4>)));
  done[dev] = true;
  return ISR_OK;
}

// fills geometry, smem plan and tensor maps of one conv launch; epilogue fields are set by the caller
static int init_conv(const Tunables &tun, ConvLaunch *L, int n, const void *act, int B, int H, int W, int n_act_parts,
                     const void *wpacked, int w_rows_total, int w_row0, int w_row_stride, int n_w_parts, int is_bf16,
                     int tensor_parts = 0, const void *act8 = nullptr, const void *w8 = nullptr) {
  // act8 / w8 != nullptr: the second activation part is the planar 8-bit plane act8 (64 B per pixel) and the correction pass
  // runs against the 8-bit weight plane w8; `act` is then the planar hi plane (tensor_parts = 1)
  if (tensor_parts < n_act_parts) tensor_parts = n_act_parts;   // planes stored per pixel (>= planes the MMAs read)
  memset(&L->p, 0, sizeof L->p);
  L->n = n;
  ConvParams &p = L->p;
  p.B = B; p.H = H; p.W = W;
  p.tiles_x = (W + kTileW - 1) / kTileW;
  p.tiles_y = (H + kTileH - 1) / kTileH;
  p.num_tiles = p.tiles_x * p.tiles_y * B;
  p.n_act_parts = n_act_parts;
  p.n_w_parts = n_w_parts;
  p.is_bf16 = is_bf16;
  p.w_row0 = w_row0;
  p.w_row_stride = w_row_stride;
  L->final2 = 0;
  L->pair = ((n == 64 && tun.cta_pair) || (n == 16 && tun.final_stacked == 2)) ? 1 : 0;
  const bool stack = n_w_parts == 2 && tun.stack_w;
  L->stack = stack ? 1 : 0;
  L->pdl = tun.pdl;
  L->cp = 0; L->ph = 1;
  // weight bytes per CTA: the pair kernel keeps half of the B operand per CTA (+ its half of W_hi when stacked)
  size_t w_bytes = align_up(static_cast<size_t>(n_w_parts) * 9 * n * 128, 1024);
  if (L->pair) w_bytes = align_up(static_cast<size_t>(stack ? 3 : n_w_parts) * 9 * (n / 2) * 128, 1024);
  if (L->pair && stack && w8 != nullptr) w_bytes = align_up(static_cast<size_t>(9) * 2 * (n / 2) * 128 + 9 * (n / 2) * 64, 1024);
  p.stage_bytes = static_cast<int>(align_up(static_cast<size_t>(kPitch) * kHaloH * 128, 1024));
  const size_t avail = kSmemLimit - kStaticSmem - 1024 - w_bytes - 256;
  int stages = static_cast<int>(avail / p.stage_bytes);
  if (stages > 8) stages = 8;
  // two issuing warps alternate tiles: a stage (and its barriers) must always be consumed by the same warp, or that
  // warp would skip barrier phases and its parity waits would alias
  if (L->pair) stages -= stages % (kPairIssuers * n_act_parts);
  if (stages < 2) return fail(ISR_ERR_UNSUPPORTED, "conv smem plan: fewer than 2 halo stages fit");
  p.num_stages = stages;
  L->smem = 1024 + w_bytes + static_cast<size_t>(stages) * p.stage_bytes + 256;
  int cap = tun.max_ctas > 0 ? tun.max_ctas : num_sms();
  L->grid = p.num_tiles < cap ? p.num_tiles : cap;
  if (L->pair) {
    const int want = 2 * ((p.num_tiles + 1) / 2);
    cap &= ~1;
    if (cap < 2) cap = 2;
    L->grid = want < cap ? want : cap;
  }
  int rc = make_act_map(&L->tmA, act, B, H, W, w8 != nullptr ? 64 : 64 * tensor_parts, kPitch, kHaloH, is_bf16);
  if (rc != ISR_OK) return rc;
  if ((rc = make_w_map(&L->tmW, wpacked, w_rows_total, L->pair ? n / 2 : n, is_bf16)) != ISR_OK) return rc;
  L->tmA8 = L->tmA;
  L->tmW8 = L->tmW;
  if (w8 != nullptr) {
    if (!(L->pair && stack && n == 64 && n_act_parts == 2 && act8 != nullptr))
      return fail(ISR_ERR_UNSUPPORTED, "the 8-bit low plane needs the stacked CTA-pair kernel");
    p.a_lo8 = 1;
    if ((rc = make_act8_map(&L->tmA8, act8, B, H, W, 64, kPitch, kHaloH)) != ISR_OK) return rc;
    if ((rc = make_w8_map(&L->tmW8, w8, 9 * 64, 32)) != ISR_OK) return rc;
  }
  return ISR_OK;
}

// fused bilinear-2x-upsample + conv (64 -> 64): act = LOW-res input [B][hl][wl][64 * tensor_parts], wpacked = composite
// weights of the chosen format [phase 4][tap 9][64][64]; the launch writes the (2 hl) x (2 wl) output
static int init_conv_up2(const Tunables &tun, ConvLaunch *L, const void *act, int B, int hl, int wl, int tensor_parts,
                         const void *wpacked, int is_bf16) {
  memset(&L->p, 0, sizeof L->p);
  L->n = 64; L->cp = 0; L->final2 = 0; L->pair = 1; L->ph = tun.up2_wide ? 2 : 4; L->stack = 0; L->pdl = tun.pdl;
  ConvParams &p = L->p;
  p.B = B; p.H = hl; p.W = wl;
  p.tiles_x = (wl + kTileW - 1) / kTileW;
  p.tiles_y = (hl + kTileH - 1) / kTileH;
  p.num_tiles = p.tiles_x * p.tiles_y * B;
  p.n_act_parts = 1; p.n_w_parts = 1; p.is_bf16 = is_bf16;
  p.w_row0 = 0; p.w_row_stride = 64;
  const size_t w_bytes = static_cast<size_t>(36) * 32 * 128;   // per CTA: 4 phases x 9 taps x 32 rows
  p.stage_bytes = static_cast<int>(align_up(static_cast<size_t>(kPitch) * kHaloH * 128, 1024));
  int stages = static_cast<int>((kSmemLimit - kStaticSmem - 1024 - w_bytes - 512) / p.stage_bytes);
  if (stages > 4) stages = 4;
  if (stages < 2) return fail(ISR_ERR_UNSUPPORTED, "up2 conv smem plan: fewer than 2 halo stages fit");
  p.num_stages = stages;
  L->smem = 1024 + w_bytes + static_cast<size_t>(stages) * p.stage_bytes + 512;   // 8 accumulator barriers: 512 B block
  int cap = (tun.max_ctas > 0 ? tun.max_ctas : num_sms()) & ~1;
  if (cap < 2) cap = 2;
  const int want = 2 * ((p.num_tiles + 1) / 2);
  L->grid = want < cap ? want : cap;
  int rc = make_act_map(&L->tmA, act, B, hl, wl, 64 * tensor_parts, kPitch, kHaloH, is_bf16);
  if (rc != ISR_OK) return rc;
  if ((rc = make_w_map(&L->tmW, wpacked, 36 * 64, 32, is_bf16)) != ISR_OK) return rc;
  L->tmA8 = L->tmA;
  L->tmW8 = L->tmW;
  return ISR_OK;
}

static int final_cp(int cout) { return cout <= 4 ? 4 : (cout <= 6 ? 6 : 8); }
static int final_rows(int cp) { return (9 * cp + 15) / 16 * 16; }
static size_t final_pbytes(int cp) {
  const int ld = (9 * cp + 7) / 8 * 8;
  const int pitch4 = ((ld / 4) % 2 == 0) ? ld / 4 + 1 : ld / 4;
  return align_up(static_cast<size_t>(kPitch) * kHaloH * pitch4 * 16 + 6 * 4 * 8 * 4, 1024);   // gather buffer + skip patch
}

// tap-stacked output conv: wpacked = [part 2][nrows][64] of the chosen format
static int init_conv_final(const Tunables &tun, ConvLaunch *L, int cp, const void *act, int B, int H, int W,
                           int n_act_parts, const void *wpacked, int n_w_parts, int is_bf16) {
  memset(&L->p, 0, sizeof L->p);
  const int nrows = final_rows(cp);
  L->n = nrows;
  L->cp = cp;
  L->final2 = 0;
  L->pair = 0; L->ph = 1; L->stack = 0; L->pdl = tun.pdl;
  ConvParams &p = L->p;
  p.B = B; p.H = H; p.W = W;
  p.tiles_x = (W + kFTileW - 1) / kFTileW;
  p.tiles_y = (H + kFTileH - 1) / kFTileH;
  p.num_tiles = p.tiles_x * p.tiles_y * B;
  p.n_act_parts = n_act_parts;
  p.n_w_parts = n_w_parts;
  p.is_bf16 = is_bf16;
  p.w_row0 = 0;
  p.w_row_stride = nrows;
  const size_t w_bytes = align_up(static_cast<size_t>(n_w_parts) * nrows * 128, 1024);
  p.stage_bytes = static_cast<int>(align_up(static_cast<size_t>(kFPitch) * kFHaloH * 128, 1024));
  const size_t fixed = 1024 + w_bytes + kFinalGroups * final_pbytes(cp) + 256;
  int stages = static_cast<int>((kSmemLimit - kStaticSmem - fixed) / p.stage_bytes);
  if (stages > 8) stages = 8;
  if (stages < 2) return fail(ISR_ERR_UNSUPPORTED, "final conv smem plan: fewer than 2 halo stages fit");
  p.issuers = 1;
  if (tun.final_issuers == 3 && stages >= kFinalIssuers * n_act_parts) {   // three issuing warps need every stage to belong to one of them
    stages -= stages % (kFinalIssuers * n_act_parts);
    p.issuers = kFinalIssuers;
  }
  p.num_stages = stages;
  L->smem = fixed + static_cast<size_t>(stages) * p.stage_bytes;
  int cap = tun.max_ctas > 0 ? tun.max_ctas : num_sms();
  L->grid = p.num_tiles < cap ? p.num_tiles : cap;
  int rc = make_act_map(&L->tmA, act, B, H, W, 64 * n_act_parts, kFPitch, kFHaloH, is_bf16);
  if (rc != ISR_OK) return rc;
  return make_w_map(&L->tmW, wpacked, 2 * nrows, nrows, is_bf16);
}

static int final2_nk(int cp) { return (3 * cp + 15) / 16 * 16; }

// column-stacked output conv (isr_conv_final2.cuh): wpacked = [ky 3][part 2][nk][64] of the chosen format
static int init_conv_final2(const Tunables &tun, ConvLaunch *L, int cp, const void *act, int B, int H, int W,
                            int n_act_parts, const void *wpacked, int n_w_parts, int is_bf16) {
  memset(&L->p, 0, sizeof L->p);
  const int nk = final2_nk(cp);
  L->n = nk;
  L->cp = cp;
  L->final2 = 1;
  // four groups only exist for the specialised epilogues (the generic one needs more than the 96 registers 21 warps leave)
  L->groups = (tun.final_groups == 4 && (cp == 6 || cp == 4)) ? 4 : 3;
  L->pair = 0; L->ph = 1; L->stack = 0; L->pdl = tun.pdl;
  ConvParams &p = L->p;
  p.B = B; p.H = H; p.W = W;
  p.tiles_x = (W + kGTileW - 1) / kGTileW;
  p.tiles_y = (H + kGTileH - 1) / kGTileH;
  p.num_tiles = p.tiles_x * p.tiles_y * B;
  p.n_act_parts = n_act_parts;
  p.n_w_parts = n_w_parts;
  p.is_bf16 = is_bf16;
  p.w_row0 = 0;
  p.w_row_stride = nk;
  const size_t w_bytes = align_up(static_cast<size_t>(3) * n_w_parts * nk * 128, 1024);
  p.stage_bytes = static_cast<int>(align_up(static_cast<size_t>(kGStageRows) * 128, 1024));
  const size_t fixed = 1024 + w_bytes + kG2MaxGroups * 2 * kGScratch + 256;
  int stages = static_cast<int>((kSmemLimit - kStaticSmem - fixed) / p.stage_bytes);
  if (stages > 8) stages = 8;
  if (stages < 2) return fail(ISR_ERR_UNSUPPORTED, "final2 conv smem plan: fewer than 2 halo stages fit");
  p.issuers = 1;
  if (tun.final_issuers == 3 && stages >= L->groups * n_act_parts) {   // one issuing warp per accumulator buffer
    stages -= stages % (L->groups * n_act_parts);
    p.issuers = L->groups;
  }
  p.num_stages = stages;
  L->smem = fixed + static_cast<size_t>(stages) * p.stage_bytes;
  int cap = tun.max_ctas > 0 ? tun.max_ctas : num_sms();
  L->grid = p.num_tiles < cap ? p.num_tiles : cap;
  int rc = make_act_map(&L->tmA, act, B, H, W, 64 * n_act_parts, kGPitch, kGHaloH, is_bf16);
  if (rc != ISR_OK) return rc;
  return make_w_map(&L->tmW, wpacked, 3 * 2 * nk, nk, is_bf16);
}

static int run_conv(const ConvLaunch &L, cudaStream_t s) {
  if (L.final2) {
    // the compile-time specialised epilogues: the unshaded (6 / 5 / normalise state) and the shaded (3 / 3 / clamp) network
    const ConvParams &q = L.p;
    const bool fast6 = L.cp == 6 && q.cout == 6 && q.n_mask == 5 && q.recon_residual && q.out_raw == nullptr &&
                       (q.out_state == nullptr || q.state_unshaded);
    const bool fast4 = L.cp == 4 && q.cout == 3 && q.n_mask == 3 && q.recon_residual && q.out_raw == nullptr &&
                       (q.out_state == nullptr || !q.state_unshaded);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(L.grid);
    cfg.blockDim = dim3(g2_threads(L.groups));
    cfg.dynamicSmemBytes = L.smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = L.pdl ? 1 : 0;   // follows a CTA-pair conv: its CTAs cannot become resident before that kernel's exit
    if (fast6 && L.groups == 4) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<6, 1, 4>, L.tmA, L.tmW, L.p));
    else if (fast4 && L.groups == 4) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<4, 1, 4>, L.tmA, L.tmW, L.p));
    else if (fast6) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<6, 1, 3>, L.tmA, L.tmW, L.p));
    else if (fast4) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<4, 1, 3>, L.tmA, L.tmW, L.p));
    else if (L.cp == 4) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<4, 0, 3>, L.tmA, L.tmW, L.p));
    else if (L.cp == 6) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<6, 0, 3>, L.tmA, L.tmW, L.p));
    else ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_final2_kernel<8, 0, 3>, L.tmA, L.tmW, L.p));
    count_launch();
    ISR_CUDA(cudaGetLastError());
    return ISR_OK;
  }
  if (L.cp > 0) {
    if (L.cp == 4) conv3x3_final_kernel<4><<<L.grid, kFinalThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
    else if (L.cp == 6) conv3x3_final_kernel<6><<<L.grid, kFinalThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
    else conv3x3_final_kernel<8><<<L.grid, kFinalThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
    count_launch();
    ISR_CUDA(cudaGetLastError());
    return ISR_OK;
  }
  const bool stack = L.stack != 0;
  if (L.pair) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(L.grid);
    cfg.blockDim = dim3(kPairThreads);
    cfg.dynamicSmemBytes = L.smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = L.pdl ? 1 : 0;
    if (L.ph == 2) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<0, 64, EPI_STORE16, 2>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else if (L.ph == 4) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<0, 64, EPI_STORE16, 4>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else if (L.n == 16 && stack) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<1, 16, EPI_FINAL>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else if (L.n == 16) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<0, 16, EPI_FINAL>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else if (stack && L.p.addend == nullptr && !(L.p.dbg & 2048)) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<1, 64, EPI_STORE16, 1, false>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else if (stack && L.p.addend != nullptr && !(L.p.dbg & 2048)) {
      // with a residual addend: eight warps per epilogue group (isr_conv_pair.cuh)
      cfg.blockDim = dim3(pair_threads(8));
      ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<1, 64, EPI_STORE16, 1, true, 8>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    }
    else if (stack) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<1, 64, EPI_STORE16>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else if (L.p.addend == nullptr && !(L.p.dbg & 2048)) ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<0, 64, EPI_STORE16, 1, false>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    else ISR_CUDA(cudaLaunchKernelEx(&cfg, conv3x3_pair_kernel<0, 64, EPI_STORE16>, L.tmA, L.tmW, L.tmA8, L.tmW8, L.p));
    count_launch();
    return ISR_OK;
  }
  if (L.n == 64 && stack)
    conv3x3_umma_kernel<64, EPI_STORE16, 1><<<L.grid, kConvThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
  else if (L.n == 64)
    conv3x3_umma_kernel<64, EPI_STORE16, 0><<<L.grid, kConvThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
  else if (stack)
    conv3x3_umma_kernel<16, EPI_FINAL, 1><<<L.grid, kConvThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
  else
    conv3x3_umma_kernel<16, EPI_FINAL, 0><<<L.grid, kConvThreads, L.smem, s>>>(L.tmA, L.tmW, L.p);
  count_launch();
  ISR_CUDA(cudaGetLastError());
  return ISR_OK;
}

// ---- precision schemes -----------------------------------------------------------------------------------------
struct Scheme {
  int bf16, na, nw;  // operand format, activation parts read by the MMAs, weight parts
  int fs = 0;        // parts STORED for the residual stream f (0 = na).  fs = 2 with na = 1: the stream keeps hi + lo planes
                     // (the block's addend and the next block's output are exact to ~22 bits) while the MMAs read hi only
  int stream_parts() const { return fs > 0 ? fs : na; }
};
static bool get_schemes(int precision, Scheme *trunk, Scheme *tail, int *fin_nw = nullptr) {
  int dummy;
  if (fin_nw == nullptr) fin_nw = &dummy;
  *fin_nw = 0;  // 0 = same as the tail
  switch (precision) {
    case ISR_PREC_BF16: *trunk = {1, 1, 1}; *tail = {1, 1, 1}; return true;
    case ISR_PREC_BF16X3: *trunk = {1, 2, 2}; *tail = {1, 2, 2}; return true;
    case ISR_PREC_MIXED: *trunk = {1, 2, 2}; *tail = {0, 1, 1}; *fin_nw = 2; return true;
    case ISR_PREC_FP16: *trunk = {0, 1, 1}; *tail = {0, 1, 1}; return true;
    case ISR_PREC_FP16X3: *trunk = {0, 2, 2}; *tail = {0, 2, 2}; return true;
    case ISR_PREC_MIXED2: *trunk = {1, 2, 2}; *tail = {0, 2, 1}; return true;
    case ISR_PREC_MIXED3: *trunk = {1, 2, 2}; *tail = {0, 1, 2}; return true;
    case ISR_PREC_MIXED4: *trunk = {0, 1, 2, 1}; *tail = {0, 1, 1}; *fin_nw = 2; return true;
    case ISR_PREC_MIXED5: *trunk = {0, 1, 2, 2}; *tail = {0, 1, 1}; *fin_nw = 2; return true;
    case ISR_PREC_MIXED6: *trunk = {0, 1, 1, 2}; *tail = {0, 1, 1}; *fin_nw = 2; return true;
    case ISR_PREC_HYBRID: *trunk = {0, 2, 2}; *tail = {0, 1, 1}; *fin_nw = 2; return true;   // per block, see build_plan
  }
  return false;
}

// ---- the network ------------------------------------------------------------------------------------------------
struct Layer {
  int cin = 0, cout = 0, npad = 0, G = 0;
  uint16_t *packed = nullptr;  // [fmt 2][part 2][tap 9][G][npad][64]
  float *bias = nullptr;       // npad floats, zero padded
  uint16_t *packed_fin = nullptr;  // output conv only: tap-stacked [fmt 2][part 2][nrows][64]
  uint16_t *packed_fin2 = nullptr; // output conv only: column-stacked [fmt 2][ky 3][part 2][nk][64] (isr_conv_final2.cuh)
  uint8_t *packed_w8 = nullptr;    // 64 -> 64 trunk layers (1..20): e5m2(W * 2^-6) plane [tap 9][n 64][k 64] for the 8-bit correction pass
  uint16_t *packed_up = nullptr;   // layers 21, 22: composite weights of the fused upsample + conv [fmt 2][phase 4][tap 9][64][64]
  size_t per_fmt() const { return static_cast<size_t>(2) * 9 * G * npad * 64; }
  int rows() const { return 2 * 9 * G * npad; }
};

struct Plan {
  int B = 0, h = 0, w = 0, precision = -1, f_last_parts = 1;
  int in_parts = 1;     // what the packing kernels write: 1 | 2 interleaved 16-bit planes, 3 = planar fp16 hi + e4m3 lo8
  void *ws = nullptr;
  size_t ws_need = 0;   // bytes of workspace this plan uses (checked on every call, also on a cache hit)
  Scheme trunk, tail;
  int fin_nw = 0;
  uint16_t *in_g[2] = {nullptr, nullptr};
  std::vector<ConvLaunch> convs;  // in execution order
  // pointwise steps interleaved with convs: up2 after conv index up_after[0], up4 after up_after[1]
  uint16_t *f_last = nullptr, *u2 = nullptr, *c2 = nullptr, *u4 = nullptr, *c4 = nullptr;
  int idx_post1 = 0, idx_post4 = 0, idx_final = 0;
  // fused 4x upsample + conv (isr_up2.cu): the phase kernel, the two band convs and their buffers
  struct Fused {
    int on = 0, hl = 0, wl = 0, in_parts = 1;   // low-res size and stored planes of the input
    const uint16_t *src = nullptr;
    uint16_t *dst = nullptr;
    ConvLaunch up, band_r, band_c;
  } fuse2, fuse4;   // (upsample -> layers[21]) at 2x, (upsample -> layers[22]) at 4x
  uint16_t *rows_in = nullptr, *cols_in = nullptr, *rows_out = nullptr, *cols_out = nullptr;   // shared by both
};

}  // namespace isr

struct isr_net {
  isr::Tunables tun;   // snapshot of the process defaults at isr_net_create; isr_net_set_tunable changes it (and drops the plans)
  int cin, cout, n_mask, upsample, recon_residual;
  int G;  // 64-channel input groups of the first conv
  isr::Layer layers[25];
  bool loaded = false;
  std::vector<isr::Plan *> plans;
};

namespace isr {

// does this plan keep the 8-bit low plane (320-byte pixels) in its trunk tensors?
static bool plan_lo8(const isr_net *net, int precision) {
  return precision == ISR_PREC_HYBRID && net->tun.trunk_lo8 && net->tun.cta_pair && net->tun.stack_w &&
         net->tun.trunk_full_blocks > 0;
}

static size_t plan_bytes(const isr_net *net, int B, int h, int w, const Scheme &trunk, const Scheme &tail,
                         size_t *offs /* 13 offsets */, bool lo8 = false) {
  const size_t P = static_cast<size_t>(B) * h * w;
  size_t lo = align_up(P * 128 * (trunk.stream_parts() > trunk.na ? trunk.stream_parts() : trunk.na), 1024);
  if (lo8) lo = align_up(P * kLo8Bytes, 1024);   // planar: hi (P x 128 B), lo fp16 (P x 128 B), lo8 (P x 64 B)
  const size_t t2 = align_up(P * 4 * 128 * tail.na, 1024);
  const size_t t4 = align_up(P * 16 * 128 * tail.na, 1024);
  size_t o = 0;
  size_t tmp[13];
  if (offs == nullptr) offs = tmp;
  offs[0] = o; o += lo;                       // in_g0
  offs[1] = o; o += (net->G == 2 ? lo : 0);   // in_g1
  offs[2] = o; o += (net->G == 2 ? lo : 0);   // partial
  offs[3] = o; o += lo;                       // f0
  offs[4] = o; o += lo;                       // f1
  offs[5] = o; o += lo;                       // t
  offs[6] = o; o += 2 * t2;                   // u2, c2
  offs[7] = o; o += t4;                       // u4
  offs[8] = o; o += t4;                       // c4
  // bands of the fused 4x upsample + conv (isr_up2.cu): 3 rows / 3 columns on each side, input and output
  const size_t rb = align_up(static_cast<size_t>(B) * 6 * 4 * w * 128 * tail.na, 1024);
  const size_t cb = align_up(static_cast<size_t>(B) * 6 * 4 * h * 128 * tail.na, 1024);
  offs[9] = o; o += rb;                       // rows_in
  offs[10] = o; o += cb;                      // cols_in
  offs[11] = o; o += rb;                      // rows_out
  offs[12] = o; o += cb;                      // cols_out
  return o;
}

static void set_store(ConvLaunch *L, const float *bias, int relu, const uint16_t *addend, int add_parts, uint16_t *out,
                      int out_parts) {
  L->p.bias = bias;
  L->p.relu = relu;
  L->p.addend = addend;
  L->p.add_parts = add_parts;
  L->p.out16 = out;
  L->p.out_parts = out_parts;
}

static int conv_for_layer(const Tunables &tun, ConvLaunch *L, const Layer &ly, int g, const uint16_t *act, int B, int H,
                          int W, const Scheme &sc, int tensor_parts = 0, const void *act8 = nullptr) {
  const uint16_t *wbase = ly.packed + (sc.bf16 ? 0 : ly.per_fmt());
  return init_conv(tun, L, ly.npad, act, B, H, W, sc.na, wbase, ly.rows(), g * ly.npad, ly.G * ly.npad, sc.nw, sc.bf16,
                   tensor_parts, act8, act8 != nullptr ? ly.packed_w8 + static_cast<size_t>(g) * 9 * 64 * 64 : nullptr);
}

static int build_plan(isr_net *net, Plan *pl, int B, int h, int w, int precision, void *ws, size_t ws_bytes) {
  if (!get_schemes(precision, &pl->trunk, &pl->tail, &pl->fin_nw))
    return fail(ISR_ERR_ARG, "unknown precision %d", precision);
  size_t offs[13];
  const bool lo8 = plan_lo8(net, precision);
  const size_t need = plan_bytes(net, B, h, w, pl->trunk, pl->tail, offs, lo8);
  if (ws == nullptr || ws_bytes < need)
    return fail(ISR_ERR_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, need);
  if (reinterpret_cast<uintptr_t>(ws) % 1024 != 0) return fail(ISR_ERR_ARG, "workspace must be 1024-byte aligned");
  pl->B = B; pl->h = h; pl->w = w; pl->precision = precision; pl->ws = ws; pl->ws_need = need;
  const Tunables &tun = net->tun;
  uint8_t *base = static_cast<uint8_t *>(ws);
  auto at = [&](int i) { return reinterpret_cast<uint16_t *>(base + offs[i]); };
  const size_t t2 = align_up(static_cast<size_t>(B) * h * w * 4 * 128 * pl->tail.na, 1024);
  pl->in_g[0] = at(0);
  pl->in_g[1] = net->G == 2 ? at(1) : nullptr;
  uint16_t *partial = at(2), *f0 = at(3), *f1 = at(4), *t = at(5);
  pl->u2 = at(6);
  pl->c2 = reinterpret_cast<uint16_t *>(base + offs[6] + t2);
  pl->u4 = at(7);
  pl->c4 = at(8);
  const Scheme &T = pl->trunk, &Hs = pl->tail;
  const int FS = T.stream_parts();   // planes of the residual-stream tensors (partial, f0, f1)
  pl->convs.clear();
  ConvLaunch L;
  int rc;
  // Trunk tensors (partial, f0, f1, t) of the 8-bit-low-plane form are PLANAR inside their workspace block: hi plane
  // (P x 128 B), lo fp16 plane (P x 128 B), lo8 plane (P x 64 B).  The fully split convs read hi + lo8 (the correction pass
  // is an 8-bit MMA), the residual addend reads hi + lo fp16 (the stream itself keeps its 22 bits), the epilogues write the
  // planes their consumers need.  The network input (in_g0 / in_g1, written by the packing kernels) keeps the interleaved
  // 16-bit low plane: the preblock runs the fp16 correction pass.
  const size_t PX = static_cast<size_t>(B) * h * w;
  auto lo16_of = [&](uint16_t *x) { return reinterpret_cast<uint16_t *>(reinterpret_cast<uint8_t *>(x) + PX * 128); };
  auto lo8_of = [&](uint16_t *x) { return reinterpret_cast<uint8_t *>(x) + PX * 256; };
  // planar output: hi always, lo fp16 when `parts` == 2, lo8 when `out8`; planar addend
  const bool s8 = lo8 && tun.stream_lo8;   // two-plane stream tensors are hi + lo8 (no fp16 low plane)
  auto planar = [&](ConvLaunch *c, uint16_t *out, int parts, bool out8, const uint16_t *add, int aparts) {
    if (!lo8) return;
    c->p.out_pitch = 128;
    if (s8 && parts == 2) {
      c->p.out_parts = 1;
      c->p.out_lo16 = nullptr;
      c->p.out_lo8 = lo8_of(out);
    } else {
      c->p.out_lo16 = parts == 2 ? lo16_of(out) : nullptr;
      c->p.out_lo8 = out8 ? lo8_of(out) : nullptr;
    }
    if (add != nullptr) {
      c->p.add_pitch = 128;
      if (s8 && aparts == 2) {
        c->p.add_parts = 1;
        c->p.addend_lo = nullptr;
        c->p.addend_lo8 = lo8_of(const_cast<uint16_t *>(add));
      } else {
        c->p.addend_lo = aparts == 2 ? lo16_of(const_cast<uint16_t *>(add)) : nullptr;
      }
    }
  };
  // preblock (enhancenet.py:93-95): one launch per 64-channel input group; the second adds the first's partial sums
  const bool i8 = lo8 && tun.input_lo8 && net->layers[0].packed_w8 != nullptr;   // packed input = planar hi + lo8
  pl->in_parts = i8 ? 3 : T.na;
  auto in8 = [&](int g) -> const void * { return i8 ? lo8_of(pl->in_g[g]) : nullptr; };
  if (net->G == 1) {
    if ((rc = conv_for_layer(tun, &L, net->layers[0], 0, pl->in_g[0], B, h, w, T, i8 ? 1 : 0, in8(0))) != ISR_OK) return rc;
    set_store(&L, net->layers[0].bias, 1, nullptr, 0, f0, FS);
    planar(&L, f0, FS, true, nullptr, 0);
    pl->convs.push_back(L);
  } else {
    if ((rc = conv_for_layer(tun, &L, net->layers[0], 0, pl->in_g[0], B, h, w, T, i8 ? 1 : 0, in8(0))) != ISR_OK) return rc;
    set_store(&L, nullptr, 0, nullptr, 0, partial, FS);
    planar(&L, partial, FS, false, nullptr, 0);
    pl->convs.push_back(L);
    if ((rc = conv_for_layer(tun, &L, net->layers[0], 1, pl->in_g[1], B, h, w, T, i8 ? 1 : 0, in8(1))) != ISR_OK) return rc;
    set_store(&L, net->layers[0].bias, 1, partial, FS, f0, FS);
    planar(&L, f0, FS, true, partial, FS);
    pl->convs.push_back(L);
  }
  // 10 residual blocks (enhancenet.py:108-112,140-141): f = f + conv(relu(conv(f)))
  // ISR_PREC_HYBRID: the first K blocks run fully split (hi*hi + hi*lo + lo*hi), the rest read the hi plane only against
  // split weights (one N=128 MMA per k-step); rounding errors injected early are amplified by every later block, so the
  // expensive scheme is spent where it matters.
  const bool hybrid = precision == ISR_PREC_HYBRID;
  const int K = hybrid ? tun.trunk_full_blocks : 10;
  const Scheme cheap = {0, 1, 2, 1};
  uint16_t *fc = f0, *fn = f1;
  int pf = FS;   // planes stored in the current stream tensor
  for (int blk = 0; blk < 10; ++blk) {
    const bool full = !(hybrid && blk >= K);
    const Scheme &Sb = full ? T : cheap;
    const int pn = hybrid ? ((blk + 1 < K) ? 2 : 1) : FS;
    const bool a8 = lo8 && full;              // this block's convs read the 8-bit low plane
    const bool next8 = lo8 && blk + 1 < K;    // ... and so do the next block's: its input needs one
    // planar tensors are read through their hi plane (tensor_parts 1); interleaved ones carry pf planes per pixel
    if ((rc = conv_for_layer(tun, &L, net->layers[1 + 2 * blk], 0, fc, B, h, w, Sb, lo8 ? 1 : pf, a8 ? lo8_of(fc) : nullptr)) != ISR_OK) return rc;
    set_store(&L, net->layers[1 + 2 * blk].bias, 1, nullptr, 0, t, a8 ? 1 : Sb.na);   // t: hi (+ lo8) suffices with a8
    planar(&L, t, a8 ? 1 : Sb.na, a8, nullptr, 0);
    pl->convs.push_back(L);
    if ((rc = conv_for_layer(tun, &L, net->layers[2 + 2 * blk], 0, t, B, h, w, Sb, lo8 ? 1 : 0, a8 ? lo8_of(t) : nullptr)) != ISR_OK) return rc;
    set_store(&L, net->layers[2 + 2 * blk].bias, 0, fc, pf, fn, pn);
    planar(&L, fn, pn, next8, fc, pf);
    pl->convs.push_back(L);
    uint16_t *sw = fc; fc = fn; fn = sw;
    pf = pn;
  }
  pl->f_last_parts = pf;
  pl->f_last = fc;
  // postblock (enhancenet.py:115-125): up2 -> conv -> relu -> up2 -> conv -> relu -> conv -> relu -> conv
  pl->idx_post1 = static_cast<int>(pl->convs.size());
  if ((rc = conv_for_layer(tun, &L, net->layers[21], 0, pl->u2, B, 2 * h, 2 * w, Hs)) != ISR_OK) return rc;
  set_store(&L, net->layers[21].bias, 1, nullptr, 0, pl->c2, Hs.na);
  pl->convs.push_back(L);
  pl->idx_post4 = static_cast<int>(pl->convs.size());
  if ((rc = conv_for_layer(tun, &L, net->layers[22], 0, pl->u4, B, 4 * h, 4 * w, Hs)) != ISR_OK) return rc;
  set_store(&L, net->layers[22].bias, 1, nullptr, 0, pl->c4, Hs.na);
  pl->convs.push_back(L);
  // fused form of (upsample 2x -> conv): phase-decomposed kernel on the low-res tensor + the plain conv on four border
  // bands.  At 2x the input is the trunk's last stream tensor: fused only when it is a single plane in the tail's format.
  pl->fuse2.on = pl->fuse4.on = 0;
  pl->rows_in = at(9); pl->cols_in = at(10); pl->rows_out = at(11); pl->cols_out = at(12);
  auto build_fused = [&](Plan::Fused *f, const Layer &ly, const uint16_t *src, int in_parts, int hl, int wl,
                         uint16_t *dst) -> int {
    const uint16_t *wu = ly.packed_up + (Hs.bf16 ? 0 : static_cast<size_t>(36) * 4096);
    int r;
    if ((r = init_conv_up2(tun, &f->up, src, B, hl, wl, in_parts, wu, Hs.bf16)) != ISR_OK) return r;
    set_store(&f->up, ly.bias, 1, nullptr, 0, dst, Hs.na);
    if ((r = conv_for_layer(tun, &f->band_r, ly, 0, pl->rows_in, 2 * B, 3, 2 * wl, Hs)) != ISR_OK) return r;
    set_store(&f->band_r, ly.bias, 1, nullptr, 0, pl->rows_out, Hs.na);
    if ((r = conv_for_layer(tun, &f->band_c, ly, 0, pl->cols_in, 2 * B, 2 * hl, 3, Hs)) != ISR_OK) return r;
    set_store(&f->band_c, ly.bias, 1, nullptr, 0, pl->cols_out, Hs.na);
    f->src = src; f->dst = dst; f->hl = hl; f->wl = wl; f->in_parts = in_parts; f->on = 1;
    return ISR_OK;
  };
  const bool can_fuse = tun.fuse_up && tun.cta_pair && net->upsample == ISR_UPSAMPLE_BILINEAR && Hs.na == 1 && Hs.nw == 1;
  if (can_fuse && net->layers[22].packed_up != nullptr && 2 * h >= 4 && 2 * w >= 4)
    if ((rc = build_fused(&pl->fuse4, net->layers[22], pl->c2, Hs.na, 2 * h, 2 * w, pl->c4)) != ISR_OK) return rc;
  if (can_fuse && net->layers[21].packed_up != nullptr && h >= 4 && w >= 4 && pl->f_last_parts == 1 && T.bf16 == Hs.bf16)
    if ((rc = build_fused(&pl->fuse2, net->layers[21], pl->f_last, 1, h, w, pl->c2)) != ISR_OK) return rc;
  if ((rc = conv_for_layer(tun, &L, net->layers[23], 0, pl->c4, B, 4 * h, 4 * w, Hs)) != ISR_OK) return rc;
  set_store(&L, net->layers[23].bias, 1, nullptr, 0, pl->u4, Hs.na);
  pl->convs.push_back(L);
  pl->idx_final = static_cast<int>(pl->convs.size());
  Scheme Fs = Hs;
  if (pl->fin_nw > 0) Fs.nw = pl->fin_nw;
  if (tun.final_w_parts > 0) Fs.nw = tun.final_w_parts;
  if (net->layers[24].packed_fin2 != nullptr && tun.final_stacked == 3) {
    const int cp = final_cp(net->cout), nk = final2_nk(cp);
    const uint16_t *wb = net->layers[24].packed_fin2 + (Fs.bf16 ? 0 : static_cast<size_t>(3) * 2 * nk * 64);
    if ((rc = init_conv_final2(tun, &L, cp, pl->u4, B, 4 * h, 4 * w, Fs.na, wb, Fs.nw, Fs.bf16)) != ISR_OK) return rc;
  } else if (net->layers[24].packed_fin != nullptr && (tun.final_stacked == 1 || tun.final_stacked == 3)) {
    const int cp = final_cp(net->cout), nrows = final_rows(cp);
    const uint16_t *wb = net->layers[24].packed_fin + (Fs.bf16 ? 0 : static_cast<size_t>(2) * nrows * 64);
    if ((rc = init_conv_final(tun, &L, cp, pl->u4, B, 4 * h, 4 * w, Fs.na, wb, Fs.nw, Fs.bf16)) != ISR_OK) return rc;
  } else if ((rc = conv_for_layer(tun, &L, net->layers[24], 0, pl->u4, B, 4 * h, 4 * w, Fs)) != ISR_OK) {
    return rc;
  }
  L.p.bias = net->layers[24].bias;
  L.p.cout = net->cout;
  L.p.n_mask = net->n_mask;
  L.p.upsample = net->upsample;
  L.p.recon_residual = net->recon_residual;
  L.p.h_low = h;
  L.p.w_low = w;
  pl->convs.push_back(L);
  if (tun.alt_order) {
    // the packing kernels finish at the end of their tensors, so the first conv walks down, the next one up, ...
    int k = 1;
    for (size_t i = 0; i < pl->convs.size(); ++i, ++k) {
      ConvLaunch *c = &pl->convs[i];
      if (static_cast<int>(i) == pl->idx_post1 && pl->fuse2.on) c = &pl->fuse2.up;
      if (static_cast<int>(i) == pl->idx_post4 && pl->fuse4.on) c = &pl->fuse4.up;
      c->p.reverse = k & 1;
    }
  }
  return ISR_OK;
}

static int get_plan(isr_net *net, Plan **out, int B, int h, int w, int precision, void *ws, size_t ws_bytes) {
  for (Plan *pl : net->plans)
    if (pl->B == B && pl->h == h && pl->w == w && pl->precision == precision && pl->ws == ws) {
      if (ws_bytes < pl->ws_need)
        return fail(ISR_ERR_WORKSPACE, "workspace too small: %zu < %zu bytes", ws_bytes, pl->ws_need);
      *out = pl;
      return ISR_OK;
    }
  Plan *pl = new Plan();
  int rc = build_plan(net, pl, B, h, w, precision, ws, ws_bytes);
  if (rc != ISR_OK) {
    delete pl;
    return rc;
  }
  if (net->plans.size() >= 16) {  // bounded cache
    delete net->plans.front();
    net->plans.erase(net->plans.begin());
  }
  net->plans.push_back(pl);
  *out = pl;
  return ISR_OK;
}

// runs everything after the network-input packing
static int run_plan(isr_net *net, Plan *pl, const float *skip_src, long long skip_bstride, float *out_recon,
                    float *out_raw, float *out_state, int state_unshaded, cudaStream_t s) {
  const Scheme &T = pl->trunk, &Hs = pl->tail;
  int rc;
  for (int i = 0; i < static_cast<int>(pl->convs.size()); ++i) {
    if (i == pl->idx_post1 && !pl->fuse2.on)
      ISR_CUDA(launch_upsample2x(pl->f_last, pl->u2, pl->B, pl->h, pl->w, pl->f_last_parts, Hs.na, T.bf16, Hs.bf16,
                                 net->upsample == ISR_UPSAMPLE_BILINEAR, s));
    if ((i == pl->idx_post1 && pl->fuse2.on) || (i == pl->idx_post4 && pl->fuse4.on)) {
      // conv3x3(bilinear_up2(x)) without the upsampled tensor; then the 2-px border ring from the four bands
      const Plan::Fused &f = i == pl->idx_post1 ? pl->fuse2 : pl->fuse4;
      if ((rc = run_conv(f.up, s)) != ISR_OK) return rc;
      ISR_CUDA(launch_band_upsample(f.src, pl->rows_in, pl->cols_in, pl->B, f.hl, f.wl, f.in_parts, Hs.bf16, s));
      if ((rc = run_conv(f.band_r, s)) != ISR_OK) return rc;
      if ((rc = run_conv(f.band_c, s)) != ISR_OK) return rc;
      ISR_CUDA(launch_band_scatter(pl->rows_out, pl->cols_out, f.dst, pl->B, 2 * f.hl, 2 * f.wl, Hs.na, s));
      continue;
    }
    if (i == pl->idx_post4)
      ISR_CUDA(launch_upsample2x(pl->c2, pl->u4, pl->B, 2 * pl->h, 2 * pl->w, Hs.na, Hs.na, Hs.bf16, Hs.bf16,
                                 net->upsample == ISR_UPSAMPLE_BILINEAR, s));
    if (i == pl->idx_final) {
      ConvParams &p = pl->convs[i].p;
      p.out_raw = out_raw;
      p.out_recon = out_recon;
      p.out_state = out_state;
      p.state_unshaded = state_unshaded;
      p.skip_src = skip_src;
      p.skip_bstride = skip_bstride;
    }
    if ((rc = run_conv(pl->convs[i], s)) != ISR_OK) return rc;
  }
  return ISR_OK;
}

}  // namespace isr

// ================================================================================================================
// C ABI
// ================================================================================================================
using namespace isr;

static cudaStream_t S(void *stream) { return static_cast<cudaStream_t>(stream); }

static int check_device() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    cudaGetLastError();
    return fail(ISR_ERR_CUDA, "no CUDA device: libisr has no CPU path");
  }
  return ISR_OK;
}

// ---- allocation-free flow in-painting (isr_inpaint_flow_static) -----------------------------------------------------
// The pyramid scratch of the in-painting kernels in a file-scope device array instead of a caller workspace: for callers
// that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  Every device that loads the
// library holds these ISR_INPAINT_STATIC_MAX_BYTES until the process exits.  In host code the array's NAME is only a
// host-side stand-in; the device address comes from cudaGetSymbolAddress, asked once per device with that device current.
// Nothing in the scratch has to start at zero: inpaint_init_kernel writes the whole level-0 weight plane and every pull
// step writes every value / weight element of its level before any kernel of the same call reads it, so the contents the
// previous call left behind are never observed.  All calls share the one array: calls that overlap on the GPU (two
// streams, two threads) race.
static_assert(ISR_INPAINT_STATIC_MAX_BYTES % 256 == 0, "scratch size");
static __device__ __align__(256) float g_inpaint_static_scratch[ISR_INPAINT_STATIC_MAX_BYTES / sizeof(float)];

static int inpaint_static_scratch(float **out) {
  static std::mutex mu;
  static float *ptr[kMaxDevices] = {nullptr};
  int dev = 0;
  ISR_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return fail(ISR_ERR_UNSUPPORTED, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (ptr[dev] == nullptr) {
    void *p = nullptr;
    ISR_CUDA(cudaGetSymbolAddress(&p, g_inpaint_static_scratch));   // the address on the CURRENT device
    if (p == nullptr) return fail(ISR_ERR_CUDA, "inpaint_flow_static: cudaGetSymbolAddress returned a null address");
    ptr[dev] = static_cast<float *>(p);
  }
  *out = ptr[dev];
  return ISR_OK;
}

extern "C" {

int isr_version(void) { return 100; }
const char *isr_last_error(void) { return g_err.c_str(); }
uint64_t isr_launch_count(void) { return g_launches.load(); }

// key -> slot of a Tunables block, with the range check; nullptr (and the error set) when the key / value is refused
static int *tunable_slot(Tunables *t, const char *key, int value) {
  auto bad = [](const char *msg) -> int * { fail(ISR_ERR_ARG, "%s", msg); return nullptr; };
  if (key == nullptr) return bad("unknown tunable");
  if (strcmp(key, "halo_pitch") == 0)
    return value == 10 ? &t->halo_pitch : bad("halo_pitch is fixed at 10 (validated on B200, profiles/r01_probe_umma_addressing.json)");
  if (strcmp(key, "base_offset_mode") == 0) return value == 0 ? &t->base_offset_mode : bad("base_offset_mode is fixed at 0 (validated on B200)");
  if (strcmp(key, "max_ctas") == 0) return value >= 0 ? &t->max_ctas : bad("max_ctas must be >= 0");
  if (strcmp(key, "dbg") == 0) return &t->dbg;
  if (strcmp(key, "cta_pair") == 0) return (value == 0 || value == 1) ? &t->cta_pair : bad("cta_pair must be 0 or 1");
  if (strcmp(key, "final_stacked") == 0)
    return (value >= 0 && value <= 3) ? &t->final_stacked
                                      : bad("final_stacked must be 0 (generic), 1 (tap-stacked), 2 (CTA pair) or 3 (column-stacked)");
  if (strcmp(key, "trunk_full_blocks") == 0)
    return (value >= 0 && value <= 10) ? &t->trunk_full_blocks : bad("trunk_full_blocks must be in [0, 10]");
  if (strcmp(key, "trunk_lo8") == 0) return (value == 0 || value == 1) ? &t->trunk_lo8 : bad("trunk_lo8 must be 0 or 1");
  if (strcmp(key, "fuse_up") == 0) return (value == 0 || value == 1) ? &t->fuse_up : bad("fuse_up must be 0 or 1");
  if (strcmp(key, "alt_order") == 0) return (value == 0 || value == 1) ? &t->alt_order : bad("alt_order must be 0 or 1");
  if (strcmp(key, "input_lo8") == 0) return (value == 0 || value == 1) ? &t->input_lo8 : bad("input_lo8 must be 0 or 1");
  if (strcmp(key, "stream_lo8") == 0) return (value == 0 || value == 1) ? &t->stream_lo8 : bad("stream_lo8 must be 0 or 1");
  if (strcmp(key, "up2_wide") == 0) return (value == 0 || value == 1) ? &t->up2_wide : bad("up2_wide must be 0 or 1");
  if (strcmp(key, "final_issuers") == 0) return (value == 1 || value == 3) ? &t->final_issuers : bad("final_issuers must be 1 or 3");
  if (strcmp(key, "final_groups") == 0) return (value == 3 || value == 4) ? &t->final_groups : bad("final_groups must be 3 or 4");
  if (strcmp(key, "final_w_parts") == 0)
    return (value >= 0 && value <= 2) ? &t->final_w_parts : bad("final_w_parts must be 0 (scheme), 1 or 2");
  if (strcmp(key, "pdl") == 0) return (value == 0 || value == 1) ? &t->pdl : bad("pdl must be 0 or 1");
  if (strcmp(key, "stack_w") == 0) return (value == 0 || value == 1) ? &t->stack_w : bad("stack_w must be 0 or 1");
  return bad("unknown tunable");
}

int isr_set_tunable(const char *key, int value) {
  std::lock_guard<std::mutex> lock(g_tun_mutex);
  if (key != nullptr && strcmp(key, "raycast_two_pass") == 0) {   // the one process-wide switch that is not a network tunable
    if (value < 0 || value > 2) return fail(ISR_ERR_ARG, "raycast_two_pass must be 0 (never), 1 (auto) or 2 (always)");
    const int old = g_raycast_two_pass;
    g_raycast_two_pass = value;
    return old;
  }
  if (key != nullptr && strcmp(key, "pack_variant") == 0) {       // process-wide: which network-input packing kernel
    if (value < 1 || value > 3) return fail(ISR_ERR_ARG, "pack_variant must be 1, 2 or 3");
    return set_pack_variant(value);
  }
  if (key != nullptr && strcmp(key, "pdl_small") == 0) {          // process-wide: programmatic launch of the small kernels
    if (value != 0 && value != 1) return fail(ISR_ERR_ARG, "pdl_small must be 0 or 1");
    const int old = g_pdl_small;
    g_pdl_small = value;
    return old;
  }
  if (key != nullptr && strcmp(key, "compose_variant") == 0) {    // process-wide: which composition kernel
    if (value != 1 && value != 2) return fail(ISR_ERR_ARG, "compose_variant must be 1 or 2");
    return set_compose_variant(value);
  }
  int *slot = tunable_slot(&g_tun, key, value);
  if (slot == nullptr) return ISR_ERR_ARG;
  const int old = *slot;
  *slot = value;
  return old;
}

int isr_net_set_tunable(isr_net *net, const char *key, int value) {
  if (net == nullptr) return fail(ISR_ERR_ARG, "net_set_tunable: null handle");
  int *slot = tunable_slot(&net->tun, key, value);
  if (slot == nullptr) return ISR_ERR_ARG;
  const int old = *slot;
  if (old != value) {
    *slot = value;
    for (Plan *pl : net->plans) delete pl;   // cached plans were built under the old settings
    net->plans.clear();
  }
  return old;
}

int isr_flatten_high(const float *in, float *out, int B, int C, int H, int W, int r, void *stream) {
  if (in == nullptr || out == nullptr || B <= 0 || C <= 0 || r <= 0 || H <= 0 || W <= 0 || H % r != 0 || W % r != 0)
    return fail(ISR_ERR_ARG, "flatten_high: bad shape (H, W must be multiples of the upscale factor)");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_flatten_high(in, out, B, C, H, W, r, S(stream)));
  return ISR_OK;
}

int isr_warp_upscale(const float *prev, const float *flow, float *out, int B, int C, int h, int w, int r,
                     int special_mask, int align_corners, void *stream) {
  if (prev == nullptr || flow == nullptr || out == nullptr || B <= 0 || C <= 0 || h <= 0 || w <= 0 || r <= 0)
    return fail(ISR_ERR_ARG, "warp_upscale: bad arguments");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_warp_upscale(prev, flow, out, B, C, h, w, r, special_mask, align_corners, S(stream)));
  return ISR_OK;
}

int isr_temporal_blend(const float *prev_rgb, const float *flow, const float *cur_rgb, float *out, int B, int C, int h,
                       int w, int r, double factor, void *stream) {
  if (prev_rgb == nullptr || flow == nullptr || cur_rgb == nullptr || out == nullptr || B <= 0 || C <= 0 || h <= 0 ||
      w <= 0 || r <= 0)
    return fail(ISR_ERR_ARG, "temporal_blend: bad arguments");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  // the reference multiplies fp32 tensors by the Python floats `factor` and `1 - factor` (taken in double, mainGUI.py:846-847)
  const float fa = static_cast<float>(factor), fb = static_cast<float>(1.0 - factor);
  ISR_CUDA(launch_temporal_blend(prev_rgb, flow, cur_rgb, out, B, C, h, w, r, fa, fb, S(stream)));
  return ISR_OK;
}

int isr_initial_image(const float *in, float *out, int B, int Cin, int h, int w, int channels, int mode,
                      int ao_inverted, int r, void *stream) {
  if (mode < 0 || mode > 2) return fail(ISR_ERR_ARG, "unknown input mode");
  if (mode == ISR_INITIAL_UNSHADED && channels != 5 && channels != 6)
    return fail(ISR_ERR_ARG, "for mode='unshaded', channels is expected to be 5 or 6");
  if (in == nullptr || out == nullptr || B <= 0 || Cin <= 0 || h <= 0 || w <= 0 || channels <= 0 || r <= 0)
    return fail(ISR_ERR_ARG, "initial_image: bad arguments");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_initial_image(in, out, B, Cin, h, w, channels, mode, ao_inverted, r, S(stream)));
  return ISR_OK;
}

int isr_postprocess(const float *in, float *out, int B, int C, int H, int W, int unshaded, void *stream) {
  if (in == nullptr || out == nullptr || B <= 0 || C <= 0 || H <= 0 || W <= 0 || (unshaded && C < 5))
    return fail(ISR_ERR_ARG, "postprocess: bad arguments");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_postprocess(in, out, B, C, H, W, unshaded, S(stream)));
  return ISR_OK;
}

int isr_shade(const float *gbuf, const float *params_host, float *rgb, int B, int C, int H, int W, void *stream) {
  if (gbuf == nullptr || params_host == nullptr || rgb == nullptr || B <= 0 || H <= 0 || W <= 0)
    return fail(ISR_ERR_ARG, "shade: bad arguments");
  if (C < 5) return fail(ISR_ERR_ARG, "shade: input must have at least 5 channels (shading.py:150)");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_shade(gbuf, params_host, rgb, B, C, H, W, S(stream)));
  return ISR_OK;
}

int isr_shade_rgb8(const float *gbuf, const float *params_host, unsigned char *rgb8, int B, int C, int H, int W,
                   void *stream) {
  if (gbuf == nullptr || params_host == nullptr || rgb8 == nullptr || B <= 0 || H <= 0 || W <= 0)
    return fail(ISR_ERR_ARG, "shade_rgb8: bad arguments");
  if (C < 5) return fail(ISR_ERR_ARG, "shade_rgb8: input must have at least 5 channels (shading.py:150)");
  if ((static_cast<size_t>(H) * W) % 4 != 0 || reinterpret_cast<uintptr_t>(gbuf) % 16 != 0 || reinterpret_cast<uintptr_t>(rgb8) % 4 != 0)
    return fail(ISR_ERR_UNSUPPORTED, "shade_rgb8: H*W must be a multiple of 4 and the buffers 16 / 4 byte aligned");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_shade_rgb8(gbuf, params_host, rgb8, B, C, H, W, S(stream)));
  return ISR_OK;
}

int isr_compose(const float *low, const float *pred, const float *params_host, float *out, int B, int c_low, int c_pred,
                int h, int w, int unshaded, int masking, float background0, void *stream) {
  if (low == nullptr || pred == nullptr || out == nullptr || B <= 0 || h <= 0 || w <= 0)
    return fail(ISR_ERR_ARG, "compose: bad arguments");
  if (unshaded && (c_low < 12 || c_pred < 5 || params_host == nullptr))
    return fail(ISR_ERR_ARG, "compose: the unshaded form needs a 12-channel render, a 5|6-channel state and shading parameters");
  if (!unshaded && (c_low < 4 || c_pred != 3)) return fail(ISR_ERR_ARG, "compose: the shaded form needs c_low >= 4 and c_pred == 3");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  float zeros[22] = {0};
  ISR_CUDA(launch_compose(low, pred, params_host != nullptr ? params_host : zeros, out, B, c_low, c_pred, h, w, unshaded,
                          masking, background0, S(stream)));
  return ISR_OK;
}

// ShadowFX.cpp:35-66: sample hemisphere and kernel rotations from std::default_random_engine(42); expressions keep the
// reference's float / double mix (2.0, 1.0 and 64.0 are double literals, Vec3<float> stores the rounded result)
static void ssao_fill_tables(int samples, int ns, SsaoTables *t) {
  std::uniform_real_distribution<float> randomFloats(0.0, 1.0);
  std::default_random_engine generator(42);
  for (int i = 0; i < samples; ++i) {
    // constructor arguments in the reference: evaluated right to left by MSVC and g++ (z, y, x)
    float sz = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);
    float sy = randomFloats(generator);
    float sx = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);
    const float d = sqrtf(sx * sx + sy * sy + sz * sz);   // Vec3::normalize()
    if (!(fabsf(d) <= 1.0e-7f)) {
      const float inv = 1.0f / d;
      sx *= inv; sy *= inv; sz *= inv;
    }
    const float r = randomFloats(generator);
    sx *= r; sy *= r; sz *= r;
    float scale = static_cast<float>(static_cast<float>(i) / 64.0);
    scale = 0.1f + scale * scale * (1.0f - 0.1f);
    t->kernel[3 * i] = sx * scale; t->kernel[3 * i + 1] = sy * scale; t->kernel[3 * i + 2] = sz * scale;
  }
  for (int x = 0; x < ns; ++x)
    for (int y = 0; y < ns; ++y) {
      const float nz = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);   // right to left
      const float nx = static_cast<float>(randomFloats(generator) * 2.0 - 1.0);
      t->noise[(x * ns + y) * 3] = nx; t->noise[(x * ns + y) * 3 + 1] = 0.0f; t->noise[(x * ns + y) * 3 + 2] = nz;
    }
}

int isr_ssao_tables(int samples, int noise_samples, float *kernel_host, float *noise_host) {
  if (samples <= 0 || samples > 128 || noise_samples <= 0 || noise_samples > 8)
    return fail(ISR_ERR_UNSUPPORTED, "ssao: samples must be in [1,128] and noise_samples in [1,8]");
  SsaoTables t;
  memset(&t, 0, sizeof t);
  ssao_fill_tables(samples, noise_samples, &t);
  if (kernel_host != nullptr) memcpy(kernel_host, t.kernel, sizeof(float) * 3 * samples);
  if (noise_host != nullptr) memcpy(noise_host, t.noise, sizeof(float) * 3 * noise_samples * noise_samples);
  return ISR_OK;
}

int isr_ssao(int width, int height, const float *input, int input_entry_stride, int input_row_stride, int normal_x_index,
             int normal_y_index, int normal_z_index, int depth_index, float *output, int output_entry_stride,
             int output_row_stride, int output_index, int samples, float sample_radius, float bias, int noise_samples,
             void *stream) {
  if (input == nullptr || output == nullptr || width <= 0 || height <= 0 || input_entry_stride <= 0 ||
      output_entry_stride <= 0 || input_row_stride < width || output_row_stride < width)
    return fail(ISR_ERR_ARG, "ssao: bad arguments");
  if (samples <= 0 || samples > 128 || noise_samples <= 0 || noise_samples > 8)
    return fail(ISR_ERR_UNSUPPORTED, "ssao: samples must be in [1,128] and noise_samples in [1,8]");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  static thread_local SsaoTables tab;
  static thread_local int tab_samples = -1, tab_ns = -1;
  if (tab_samples != samples || tab_ns != noise_samples) {
    memset(&tab, 0, sizeof tab);
    ssao_fill_tables(samples, noise_samples, &tab);
    tab_samples = samples; tab_ns = noise_samples;
  }
  ISR_CUDA(launch_ssao(input, input_entry_stride, input_row_stride, normal_x_index, normal_y_index, normal_z_index,
                       depth_index, output, output_entry_stride, output_row_stride, output_index, width, height, samples,
                       sample_radius, bias, noise_samples, tab, S(stream)));
  return ISR_OK;
}

int isr_raycast_ao_tables(float *hemisphere_host, float *rotations_host) {
  // GPURendererDirect.cpp:146-190: default-seeded engine, one shared uniform(0,1) stream, hemisphere first
  std::default_random_engine rnd;
  std::uniform_real_distribution<float> distr(0.0f, 1.0f);
  for (int i = 0; i < 512; ++i) {
    const float u1 = distr(rnd);
    const float u2 = distr(rnd);
    const float r = std::sqrt(u1);
    const float theta = static_cast<float>(2 * M_PI * u2);
    const float x = r * std::cos(theta), y = r * std::sin(theta);
    float scale = distr(rnd);
    scale = static_cast<float>(0.1 + 0.9 * scale * scale);
    if (hemisphere_host != nullptr) {
      hemisphere_host[4 * i] = x * scale;
      hemisphere_host[4 * i + 1] = y * scale;
      hemisphere_host[4 * i + 2] = std::sqrt(1 - u1) * scale;
      hemisphere_host[4 * i + 3] = 0.f;
    }
  }
  for (int i = 0; i < 16; ++i) {
    const float x = distr(rnd) * 2 - 1;
    const float y = distr(rnd) * 2 - 1;
    const float linv = 1.0f / std::sqrt(x * x + y * y);
    if (rotations_host != nullptr) {
      rotations_host[4 * i] = x * linv;
      rotations_host[4 * i + 1] = y * linv;
      rotations_host[4 * i + 2] = 0.f;
      rotations_host[4 * i + 3] = 0.f;
    }
  }
  return ISR_OK;
}

int isr_raycast_skip_bytes(int nx, int ny, int nz, size_t *bytes) {
  if (bytes == nullptr || nx < 2 || ny < 2 || nz < 2) return fail(ISR_ERR_ARG, "raycast: bad brick size");
  *bytes = sizeof(float) * raycast_skip_floats(nx, ny, nz);
  return ISR_OK;
}

int isr_raycast_build_skip(const float *volume, int nx, int ny, int nz, float *skip, size_t skip_bytes, void *stream) {
  size_t need = 0;
  int rc = isr_raycast_skip_bytes(nx, ny, nz, &need);
  if (rc != ISR_OK) return rc;
  if (volume == nullptr || skip == nullptr || skip_bytes < need) return fail(ISR_ERR_ARG, "raycast: skip grid too small");
  rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_raycast_skip(volume, skip, nx, ny, nz, S(stream)));
  return ISR_OK;
}

int isr_raycast_iso(const float *volume, int nx, int ny, int nz, const float *params_host, const float *ao_hemisphere_host,
                    const float *ao_rotations_host, int ao_samples, int binary_steps, const int *viewport_host,
                    const float *skip, float *out, int width, int height, void *stream) {
  if (volume == nullptr || params_host == nullptr || out == nullptr || width <= 0 || height <= 0)
    return fail(ISR_ERR_ARG, "raycast: bad arguments");
  if (nx < 2 || ny < 2 || nz < 2) return fail(ISR_ERR_ARG, "raycast: the brick needs at least 2 voxels per axis");
  if (ao_samples < 0 || ao_samples > 512 || binary_steps < 0 || binary_steps > 64)
    return fail(ISR_ERR_UNSUPPORTED, "raycast: ao_samples must be in [0,512] and binary_steps in [0,64]");
  if (ao_samples > 0 && (ao_hemisphere_host == nullptr || ao_rotations_host == nullptr))
    return fail(ISR_ERR_ARG, "raycast: ambient-occlusion tables missing");
  if (!(params_host[ISR_RC_STEP] > 0.f) || !(params_host[ISR_RC_VOXEL] > 0.f))
    return fail(ISR_ERR_ARG, "raycast: step and voxel size must be positive");
  if ((reinterpret_cast<uintptr_t>(out) & 15) != 0) return fail(ISR_ERR_ARG, "raycast: out must be 16-byte aligned");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  static thread_local RaycastArgs a;
  a.width = width; a.height = height; a.nx = nx; a.ny = ny; a.nz = nz;
  a.binary_steps = binary_steps; a.ao_samples = ao_samples; 
  // measured on B200 (profiles/r01_raycast_one_vs_two_pass.txt): from 8 rays on the second pass wins at every frame size
  // (480x270, 32 rays: 2.3-2.5x; 1080p: 1.2x); with 4 rays its fixed cost is not recovered
  a.ao_two_pass = g_raycast_two_pass == 2 || (g_raycast_two_pass == 1 && ao_samples >= 8);
  if (viewport_host != nullptr) {
    memcpy(a.viewport, viewport_host, sizeof a.viewport);
  } else {
    a.viewport[0] = 0; a.viewport[1] = 0; a.viewport[2] = width; a.viewport[3] = height;
  }
  memcpy(a.p, params_host, sizeof a.p);
  memset(a.hemisphere, 0, sizeof a.hemisphere);
  memset(a.rotations, 0, sizeof a.rotations);
  if (ao_samples > 0) {
    memcpy(a.hemisphere, ao_hemisphere_host, sizeof(float) * 4 * ao_samples);
    memcpy(a.rotations, ao_rotations_host, sizeof a.rotations);
  }
  ISR_CUDA(launch_raycast_iso(volume, skip, out, a, S(stream)));
  return ISR_OK;
}

int isr_ingest_low_channels(int variant) {
  if (variant < 0 || variant > 4) return fail(ISR_ERR_ARG, "ingest: unknown variant %d", variant);
  return ingest_low_channels(variant);
}

int isr_ingest_gbuffer(const float *render, int interleaved, int B, int c_render, int h, int w, int variant, float *low,
                       float *flow, unsigned char *hole, float *image12, void *stream) {
  if (render == nullptr || B <= 0 || h <= 0 || w <= 0) return fail(ISR_ERR_ARG, "ingest: bad arguments");
  if (variant < 0 || variant > 4) return fail(ISR_ERR_ARG, "ingest: unknown variant %d", variant);
  if (c_render < 10 || (image12 != nullptr && c_render < 12))
    return fail(ISR_ERR_ARG, "ingest: the render needs >= 10 channels (RGB, mask, normal, depth, flow), 12 for image12");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_ingest(render, interleaved, c_render, low, flow, hole, image12, B, h, w, variant, S(stream)));
  return ISR_OK;
}

int isr_inpaint_workspace_bytes(int B, int h, int w, size_t *bytes) {
  if (bytes == nullptr || B <= 0 || h <= 0 || w <= 0) return fail(ISR_ERR_ARG, "inpaint workspace: bad arguments");
  *bytes = inpaint_workspace_floats(B, h, w) * sizeof(float);
  return ISR_OK;
}

int isr_inpaint_flow(const float *flow, const unsigned char *hole, float *out, int B, int h, int w, void *workspace,
                     size_t workspace_bytes, void *stream) {
  if (flow == nullptr || hole == nullptr || out == nullptr || workspace == nullptr || B <= 0 || h <= 0 || w <= 0)
    return fail(ISR_ERR_ARG, "inpaint_flow: bad arguments");
  const size_t need = inpaint_workspace_floats(B, h, w) * sizeof(float);
  if (workspace_bytes < need) return fail(ISR_ERR_WORKSPACE, "inpaint_flow: workspace too small (%zu < %zu)", workspace_bytes, need);
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_inpaint_flow(flow, hole, out, B, h, w, static_cast<float *>(workspace), S(stream)));
  return ISR_OK;
}

int isr_inpaint_flow_static(const float *flow, const unsigned char *hole, float *out, int B, int h, int w, void *stream) {
  if (flow == nullptr || hole == nullptr || out == nullptr || B <= 0 || h <= 0 || w <= 0)
    return fail(ISR_ERR_ARG, "inpaint_flow_static: bad arguments");
  const size_t need = inpaint_workspace_floats(B, h, w) * sizeof(float);
  if (need > static_cast<size_t>(ISR_INPAINT_STATIC_MAX_BYTES))
    return fail(ISR_ERR_WORKSPACE,
                "inpaint_flow_static: %d x %d x %d needs %zu bytes of scratch, the built-in array holds %zu "
                "(ISR_INPAINT_STATIC_MAX_BYTES); use isr_inpaint_flow with a caller workspace",
                B, h, w, need, static_cast<size_t>(ISR_INPAINT_STATIC_MAX_BYTES));
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  float *scratch = nullptr;
  if ((rc = inpaint_static_scratch(&scratch)) != ISR_OK) return rc;
  ISR_CUDA(launch_inpaint_flow(flow, hole, out, B, h, w, scratch, S(stream)));
  return ISR_OK;
}

int isr_net_create(isr_net **out, int input_channels, int output_channels, int n_mask, int upsample,
                   int recon_residual) {
  if (out == nullptr) return fail(ISR_ERR_ARG, "net_create: null output");
  if (input_channels <= 0 || input_channels > 128)
    return fail(ISR_ERR_UNSUPPORTED, "input_channels must be in [1,128] (got %d)", input_channels);
  if (output_channels <= 0 || output_channels > 16)
    return fail(ISR_ERR_UNSUPPORTED, "output_channels must be in [1,16] (got %d)", output_channels);
  if (upsample != ISR_UPSAMPLE_NEAREST && upsample != ISR_UPSAMPLE_BILINEAR)
    return fail(ISR_ERR_UNSUPPORTED, "upsample must be nearest or bilinear");
  if (recon_residual && n_mask > output_channels)
    return fail(ISR_ERR_ARG, "number of output channels must be at least the number of masked input channels");
  if (n_mask < 0 || n_mask > input_channels) return fail(ISR_ERR_ARG, "bad channel mask length");
  if (recon_residual && n_mask > 8) return fail(ISR_ERR_UNSUPPORTED, "channel masks longer than 8 are not supported");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  isr_net *net = new isr_net();
  {
    std::lock_guard<std::mutex> lock(g_tun_mutex);
    net->tun = g_tun;
  }
  net->cin = input_channels;
  net->cout = output_channels;
  net->n_mask = n_mask;
  net->upsample = upsample;
  net->recon_residual = recon_residual;
  net->G = (input_channels + 63) / 64;
  for (int i = 0; i < 25; ++i) {
    Layer &ly = net->layers[i];
    ly.cin = i == 0 ? input_channels : 64;
    ly.cout = i == 24 ? output_channels : 64;
    ly.npad = i == 24 ? 16 : 64;
    ly.G = i == 0 ? net->G : 1;
    if (cudaMalloc(&ly.packed, 2 * ly.per_fmt() * sizeof(uint16_t)) != cudaSuccess ||
        cudaMalloc(&ly.bias, 64 * sizeof(float)) != cudaSuccess) {
      isr_net_destroy(net);
      return fail(ISR_ERR_CUDA, "net_create: cudaMalloc failed");
    }
  }
  for (int i = 0; i <= 20; ++i)
    if (cudaMalloc(&net->layers[i].packed_w8, static_cast<size_t>(net->layers[i].G) * 9 * 64 * 64) != cudaSuccess) {
      isr_net_destroy(net);
      return fail(ISR_ERR_CUDA, "net_create: cudaMalloc failed");
    }
  for (int i = 21; i <= 22; ++i)
    if (cudaMalloc(&net->layers[i].packed_up, static_cast<size_t>(2) * 36 * 4096 * sizeof(uint16_t)) != cudaSuccess) {
      isr_net_destroy(net);
      return fail(ISR_ERR_CUDA, "net_create: cudaMalloc failed");
    }
  if (output_channels <= 8) {
    const int nrows = final_rows(final_cp(output_channels));
    const int nk = final2_nk(final_cp(output_channels));
    if (cudaMalloc(&net->layers[24].packed_fin, static_cast<size_t>(4) * nrows * 64 * sizeof(uint16_t)) != cudaSuccess ||
        cudaMalloc(&net->layers[24].packed_fin2, static_cast<size_t>(2) * 3 * 2 * nk * 64 * sizeof(uint16_t)) != cudaSuccess) {
      isr_net_destroy(net);
      return fail(ISR_ERR_CUDA, "net_create: cudaMalloc failed");
    }
  }
  *out = net;
  return ISR_OK;
}

void isr_net_destroy(isr_net *net) {
  if (net == nullptr) return;
  for (Plan *pl : net->plans) delete pl;
  for (int i = 0; i < 25; ++i) {
    if (net->layers[i].packed) cudaFree(net->layers[i].packed);
    if (net->layers[i].bias) cudaFree(net->layers[i].bias);
    if (net->layers[i].packed_fin) cudaFree(net->layers[i].packed_fin);
    if (net->layers[i].packed_fin2) cudaFree(net->layers[i].packed_fin2);
    if (net->layers[i].packed_up) cudaFree(net->layers[i].packed_up);
    if (net->layers[i].packed_w8) cudaFree(net->layers[i].packed_w8);
  }
  delete net;
}

int isr_net_load_weights(isr_net *net, const float *const *params, int n_params, void *stream) {
  if (net == nullptr || params == nullptr || n_params != 50)
    return fail(ISR_ERR_ARG, "load_weights: expected 50 parameter tensors (useBN=False state_dict)");
  for (int i = 0; i < 25; ++i) {
    Layer &ly = net->layers[i];
    if (params[2 * i] == nullptr || params[2 * i + 1] == nullptr) return fail(ISR_ERR_ARG, "load_weights: null tensor");
    ISR_CUDA(launch_pack_weights(params[2 * i], ly.packed, ly.cout, ly.cin, ly.npad, ly.G, S(stream)));
    if (ly.packed_up != nullptr) ISR_CUDA(launch_pack_weights_up2(params[2 * i], ly.packed_up, S(stream)));
    if (ly.packed_w8 != nullptr) ISR_CUDA(launch_pack_weights_lo8(params[2 * i], ly.packed_w8, ly.cin, ly.G, S(stream)));
    if (ly.packed_fin != nullptr)
      ISR_CUDA(launch_pack_weights_final(params[2 * i], ly.packed_fin, ly.cout, final_cp(ly.cout),
                                         final_rows(final_cp(ly.cout)), S(stream)));
    if (ly.packed_fin2 != nullptr)
      ISR_CUDA(launch_pack_weights_final2(params[2 * i], ly.packed_fin2, ly.cout, final_cp(ly.cout),
                                          final2_nk(final_cp(ly.cout)), S(stream)));
    ISR_CUDA(cudaMemsetAsync(ly.bias, 0, 64 * sizeof(float), S(stream)));
    ISR_CUDA(cudaMemcpyAsync(ly.bias, params[2 * i + 1], ly.cout * sizeof(float), cudaMemcpyDeviceToDevice, S(stream)));
  }
  ISR_CUDA(cudaStreamSynchronize(S(stream)));
  net->loaded = true;
  return ISR_OK;
}

int isr_net_workspace_bytes(const isr_net *net, int B, int h, int w, int precision, size_t *bytes) {
  Scheme t, hs;
  if (net == nullptr || bytes == nullptr || B <= 0 || h <= 0 || w <= 0) return fail(ISR_ERR_ARG, "workspace: bad arguments");
  if (!get_schemes(precision, &t, &hs)) return fail(ISR_ERR_ARG, "unknown precision %d", precision);
  *bytes = plan_bytes(net, B, h, w, t, hs, nullptr, plan_lo8(net, precision));
  return ISR_OK;
}

static int forward_common(isr_net *net, int B, int h, int w, int precision, void *ws, size_t ws_bytes, Plan **pl) {
  if (net == nullptr || B <= 0 || h <= 0 || w <= 0) return fail(ISR_ERR_ARG, "forward: bad arguments");
  if (!net->loaded) return fail(ISR_ERR_ARG, "forward: weights not loaded");
  int rc = setup_kernel_attrs();
  if (rc != ISR_OK) return rc;
  return get_plan(net, pl, B, h, w, precision, ws, ws_bytes);
}

int isr_net_forward(isr_net *net, const float *inputs, float *out_recon, float *out_raw, float *out_state,
                    int state_unshaded, int B, int h, int w, int precision, void *workspace, size_t workspace_bytes,
                    void *stream) {
  if (inputs == nullptr) return fail(ISR_ERR_ARG, "forward: null input");
  Plan *pl = nullptr;
  int rc = forward_common(net, B, h, w, precision, workspace, workspace_bytes, &pl);
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_pack_nchw(inputs, pl->in_g[0], pl->in_g[1], B, net->cin, h, w, pl->in_parts, pl->trunk.bf16, S(stream)));
  return run_plan(net, pl, inputs, static_cast<long long>(net->cin) * h * w, out_recon, out_raw, out_state,
                  state_unshaded, S(stream));
}

int isr_net_forward_recurrent(isr_net *net, const float *low, const float *prev, const float *flow, int c_low,
                              int c_prev, int special_mask, int align_corners, float *out_recon, float *out_raw,
                              float *out_state, int state_unshaded, int B, int h, int w, int precision,
                              void *workspace, size_t workspace_bytes, void *stream) {
  if (low == nullptr || prev == nullptr) return fail(ISR_ERR_ARG, "forward_recurrent: null input");
  if (net != nullptr && c_low + 16 * c_prev != net->cin)
    return fail(ISR_ERR_ARG, "forward_recurrent: c_low + 16*c_prev (%d) != input_channels (%d)", c_low + 16 * c_prev,
                net->cin);
  if (net != nullptr && net->n_mask > c_low) return fail(ISR_ERR_ARG, "forward_recurrent: channel mask longer than c_low");
  Plan *pl = nullptr;
  int rc = forward_common(net, B, h, w, precision, workspace, workspace_bytes, &pl);
  if (rc != ISR_OK) return rc;
  ISR_CUDA(launch_pack_recurrent(low, prev, flow, pl->in_g[0], pl->in_g[1], B, c_low, c_prev, h, w, special_mask,
                                 align_corners, pl->in_parts, pl->trunk.bf16, S(stream)));
  return run_plan(net, pl, low, static_cast<long long>(c_low) * h * w, out_recon, out_raw, out_state, state_unshaded,
                  S(stream));
}

int isr_net_forward_recurrent_host(isr_net *net, const float *low_host, const float *flow_host, float *state_dev,
                                   int has_prev, int initial_mode, int c_low, int c_prev, int special_mask,
                                   float *recon_host, int B, int h, int w, int precision, void *workspace,
                                   size_t workspace_bytes, float *staging_dev, void *stream) {
  if (net == nullptr || low_host == nullptr || state_dev == nullptr || recon_host == nullptr || staging_dev == nullptr)
    return fail(ISR_ERR_ARG, "forward_recurrent_host: null argument");
  if (has_prev && flow_host == nullptr) return fail(ISR_ERR_ARG, "forward_recurrent_host: flow required");
  const size_t n_low = static_cast<size_t>(B) * c_low * h * w, n_flow = static_cast<size_t>(B) * 2 * h * w;
  const size_t n_out = static_cast<size_t>(B) * net->cout * 16 * h * w;
  float *low_d = staging_dev, *flow_d = staging_dev + n_low, *recon_d = staging_dev + n_low + n_flow;
  cudaStream_t s = S(stream);
  ISR_CUDA(cudaMemcpyAsync(low_d, low_host, n_low * sizeof(float), cudaMemcpyHostToDevice, s));
  if (has_prev) ISR_CUDA(cudaMemcpyAsync(flow_d, flow_host, n_flow * sizeof(float), cudaMemcpyHostToDevice, s));
  int rc;
  if (!has_prev) {
    rc = isr_initial_image(low_d, state_dev, B, c_low, h, w, c_prev, initial_mode, 0, 4, stream);
    if (rc != ISR_OK) return rc;
  }
  rc = isr_net_forward_recurrent(net, low_d, state_dev, has_prev ? flow_d : nullptr, c_low, c_prev, special_mask, 0,
                                 recon_d, nullptr, state_dev, net->cout == 6, B, h, w, precision, workspace,
                                 workspace_bytes, stream);
  if (rc != ISR_OK) return rc;
  ISR_CUDA(cudaMemcpyAsync(recon_host, recon_d, n_out * sizeof(float), cudaMemcpyDeviceToHost, s));
  ISR_CUDA(cudaStreamSynchronize(s));
  return ISR_OK;
}

}  // extern "C"

struct isr_pipe {
  isr_net *net = nullptr;
  int c_low = 0, c_prev = 0, special_mask = 0, initial_mode = 0, B = 0, h = 0, w = 0, precision = 0, depth = 0;
  void *ws = nullptr;
  size_t ws_bytes = 0;
  float *state = nullptr, *staging = nullptr;
  size_t slot_floats = 0, n_low = 0, n_flow = 0, n_out = 0;
  cudaStream_t s_in = nullptr, s_comp = nullptr, s_out = nullptr;
  std::vector<cudaEvent_t> ev_in, ev_comp, ev_out;
  uint64_t submitted = 0;
  // display form: shaded 8-bit RGB frames leave the device instead of the fp32 prediction
  bool display = false;
  float shade_params[22] = {0};
  unsigned char *rgb8 = nullptr;
  size_t n_rgb8 = 0;
  // 16-bit form of the prediction
  uint16_t *f16 = nullptr;
};

extern "C" {

int isr_pipe_staging_floats(const isr_net *net, int c_low, int B, int h, int w, int depth, size_t *floats) {
  if (net == nullptr || floats == nullptr || c_low <= 0 || B <= 0 || h <= 0 || w <= 0 || depth < 1)
    return fail(ISR_ERR_ARG, "pipe_staging_floats: bad arguments");
  const size_t P = static_cast<size_t>(B) * h * w;
  *floats = static_cast<size_t>(depth) * (P * (c_low + 2) + P * 16 * net->cout);
  return ISR_OK;
}

int isr_pipe_create(isr_pipe **out, isr_net *net, int c_low, int c_prev, int special_mask, int initial_mode, int B, int h,
                    int w, int precision, void *workspace, size_t workspace_bytes, float *state_dev, float *staging_dev,
                    size_t staging_floats, int depth) {
  if (out == nullptr || net == nullptr || workspace == nullptr || state_dev == nullptr || staging_dev == nullptr)
    return fail(ISR_ERR_ARG, "pipe_create: null argument");
  if (depth < 1 || depth > 8) return fail(ISR_ERR_ARG, "pipe_create: depth must be in [1, 8]");
  if (c_low + 16 * c_prev != net->cin)
    return fail(ISR_ERR_ARG, "pipe_create: c_low + 16*c_prev (%d) != input_channels (%d)", c_low + 16 * c_prev, net->cin);
  if (initial_mode < 0 || initial_mode > 2) return fail(ISR_ERR_ARG, "unknown input mode");
  size_t need = 0;
  int rc = isr_pipe_staging_floats(net, c_low, B, h, w, depth, &need);
  if (rc != ISR_OK) return rc;
  if (staging_floats < need) return fail(ISR_ERR_WORKSPACE, "pipe_create: staging too small (%zu < %zu floats)", staging_floats, need);
  Plan *pl = nullptr;
  if ((rc = forward_common(net, B, h, w, precision, workspace, workspace_bytes, &pl)) != ISR_OK) return rc;
  isr_pipe *p = new isr_pipe();
  p->net = net; p->c_low = c_low; p->c_prev = c_prev; p->special_mask = special_mask; p->initial_mode = initial_mode;
  p->B = B; p->h = h; p->w = w; p->precision = precision; p->depth = depth;
  p->ws = workspace; p->ws_bytes = workspace_bytes; p->state = state_dev; p->staging = staging_dev;
  const size_t P = static_cast<size_t>(B) * h * w;
  p->n_low = P * c_low; p->n_flow = P * 2; p->n_out = P * 16 * net->cout;
  p->slot_floats = p->n_low + p->n_flow + p->n_out;
  bool ok = cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&p->s_comp, cudaStreamNonBlocking) == cudaSuccess &&
            cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking) == cudaSuccess;
  p->ev_in.resize(depth); p->ev_comp.resize(depth); p->ev_out.resize(depth);
  for (int i = 0; i < depth && ok; ++i)
    ok = cudaEventCreateWithFlags(&p->ev_in[i], cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&p->ev_comp[i], cudaEventDisableTiming) == cudaSuccess &&
         cudaEventCreateWithFlags(&p->ev_out[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) {
    isr_pipe_destroy(p);
    return fail(ISR_ERR_CUDA, "pipe_create: stream / event creation failed");
  }
  *out = p;
  return ISR_OK;
}

static int pipe_submit(isr_pipe *p, const float *low_host, const float *flow_host, float *recon_host,
                       unsigned char *rgb8_host, void *f16_host = nullptr) {
  const int k = static_cast<int>(p->submitted % p->depth);
  // the slot is reused: frame (j - depth) must have left it (its D2H finished => its kernels finished too)
  if (p->submitted >= static_cast<uint64_t>(p->depth)) ISR_CUDA(cudaEventSynchronize(p->ev_out[k]));
  float *low_d = p->staging + static_cast<size_t>(k) * p->slot_floats, *flow_d = low_d + p->n_low,
        *recon_d = flow_d + p->n_flow;
  ISR_CUDA(cudaMemcpyAsync(low_d, low_host, p->n_low * sizeof(float), cudaMemcpyHostToDevice, p->s_in));
  if (flow_host != nullptr)
    ISR_CUDA(cudaMemcpyAsync(flow_d, flow_host, p->n_flow * sizeof(float), cudaMemcpyHostToDevice, p->s_in));
  ISR_CUDA(cudaEventRecord(p->ev_in[k], p->s_in));
  ISR_CUDA(cudaStreamWaitEvent(p->s_comp, p->ev_in[k], 0));
  int rc;
  if (flow_host == nullptr) {
    rc = isr_initial_image(low_d, p->state, p->B, p->c_low, p->h, p->w, p->c_prev, p->initial_mode, 0, 4, p->s_comp);
    if (rc != ISR_OK) return rc;
  }
  // display form: the prediction is not written at all, only the recurrent state (which the shading reads)
  rc = isr_net_forward_recurrent(p->net, low_d, p->state, flow_host != nullptr ? flow_d : nullptr, p->c_low, p->c_prev,
                                 p->special_mask, 0, rgb8_host != nullptr ? nullptr : recon_d, nullptr, p->state,
                                 p->net->cout == 6, p->B, p->h, p->w, p->precision, p->ws, p->ws_bytes, p->s_comp);
  if (rc != ISR_OK) return rc;
  unsigned char *rgb8_d = nullptr;
  if (rgb8_host != nullptr) {
    rgb8_d = p->rgb8 + static_cast<size_t>(k) * p->n_rgb8;
    rc = isr_shade_rgb8(p->state, p->shade_params, rgb8_d, p->B, p->c_prev, 4 * p->h, 4 * p->w, p->s_comp);
    if (rc != ISR_OK) return rc;
  }
  uint16_t *f16_d = nullptr;
  if (f16_host != nullptr) {
    f16_d = p->f16 + static_cast<size_t>(k) * p->n_out;
    ISR_CUDA(launch_f32_to_f16(recon_d, f16_d, p->n_out, p->s_comp));
  }
  ISR_CUDA(cudaEventRecord(p->ev_comp[k], p->s_comp));
  ISR_CUDA(cudaStreamWaitEvent(p->s_out, p->ev_comp[k], 0));
  if (rgb8_host != nullptr)
    ISR_CUDA(cudaMemcpyAsync(rgb8_host, rgb8_d, p->n_rgb8, cudaMemcpyDeviceToHost, p->s_out));
  else if (f16_host != nullptr)
    ISR_CUDA(cudaMemcpyAsync(f16_host, f16_d, p->n_out * sizeof(uint16_t), cudaMemcpyDeviceToHost, p->s_out));
  else
    ISR_CUDA(cudaMemcpyAsync(recon_host, recon_d, p->n_out * sizeof(float), cudaMemcpyDeviceToHost, p->s_out));
  ISR_CUDA(cudaEventRecord(p->ev_out[k], p->s_out));
  ++p->submitted;
  return ISR_OK;
}

int isr_pipe_submit(isr_pipe *p, const float *low_host, const float *flow_host, float *recon_host) {
  if (p == nullptr || low_host == nullptr || recon_host == nullptr) return fail(ISR_ERR_ARG, "pipe_submit: null argument");
  return pipe_submit(p, low_host, flow_host, recon_host, nullptr);
}

int isr_pipe_enable_display(isr_pipe *p, const float *params_host, unsigned char *rgb8_staging_dev, size_t bytes) {
  if (p == nullptr || params_host == nullptr || rgb8_staging_dev == nullptr)
    return fail(ISR_ERR_ARG, "pipe_enable_display: null argument");
  if (p->c_prev < 5) return fail(ISR_ERR_ARG, "pipe_enable_display: shading needs the unshaded network's 5|6-channel state");
  const size_t per = static_cast<size_t>(p->B) * 16 * p->h * p->w * 3;
  if (bytes < per * p->depth) return fail(ISR_ERR_WORKSPACE, "pipe_enable_display: staging too small (%zu < %zu bytes)", bytes, per * p->depth);
  memcpy(p->shade_params, params_host, sizeof p->shade_params);
  p->rgb8 = rgb8_staging_dev;
  p->n_rgb8 = per;
  p->display = true;
  return ISR_OK;
}

int isr_pipe_submit_display(isr_pipe *p, const float *low_host, const float *flow_host, unsigned char *rgb8_host) {
  if (p == nullptr || low_host == nullptr || rgb8_host == nullptr) return fail(ISR_ERR_ARG, "pipe_submit_display: null argument");
  if (!p->display) return fail(ISR_ERR_ARG, "pipe_submit_display: call isr_pipe_enable_display first");
  return pipe_submit(p, low_host, flow_host, nullptr, rgb8_host);
}

int isr_pipe_set_state(isr_pipe *p, float *state_dev) {
  if (p == nullptr || state_dev == nullptr) return fail(ISR_ERR_ARG, "pipe_set_state: null argument");
  p->state = state_dev;   // read at submit time; launches already enqueued captured the old pointer
  return ISR_OK;
}

int isr_pipe_enable_f16(isr_pipe *p, void *f16_staging_dev, size_t bytes) {
  if (p == nullptr || f16_staging_dev == nullptr) return fail(ISR_ERR_ARG, "pipe_enable_f16: null argument");
  const size_t need = p->n_out * sizeof(uint16_t) * p->depth;
  if (bytes < need) return fail(ISR_ERR_WORKSPACE, "pipe_enable_f16: staging too small (%zu < %zu bytes)", bytes, need);
  if (reinterpret_cast<uintptr_t>(f16_staging_dev) % 16 != 0 || (p->n_out * sizeof(uint16_t)) % 16 != 0)
    return fail(ISR_ERR_ARG, "pipe_enable_f16: staging must be 16-byte aligned (and B*Cout*16*h*w a multiple of 8)");
  p->f16 = static_cast<uint16_t *>(f16_staging_dev);
  return ISR_OK;
}

int isr_pipe_submit_f16(isr_pipe *p, const float *low_host, const float *flow_host, void *recon_f16_host) {
  if (p == nullptr || low_host == nullptr || recon_f16_host == nullptr) return fail(ISR_ERR_ARG, "pipe_submit_f16: null argument");
  if (p->f16 == nullptr) return fail(ISR_ERR_ARG, "pipe_submit_f16: call isr_pipe_enable_f16 first");
  return pipe_submit(p, low_host, flow_host, nullptr, nullptr, recon_f16_host);
}

int isr_pipe_wait(isr_pipe *p) {
  if (p == nullptr) return fail(ISR_ERR_ARG, "pipe_wait: null argument");
  ISR_CUDA(cudaStreamSynchronize(p->s_comp));
  ISR_CUDA(cudaStreamSynchronize(p->s_out));
  return ISR_OK;
}

void isr_pipe_destroy(isr_pipe *p) {
  if (p == nullptr) return;
  if (p->s_comp) cudaStreamSynchronize(p->s_comp);
  if (p->s_out) cudaStreamSynchronize(p->s_out);
  for (cudaEvent_t e : p->ev_in) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : p->ev_comp) if (e) cudaEventDestroy(e);
  for (cudaEvent_t e : p->ev_out) if (e) cudaEventDestroy(e);
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_comp) cudaStreamDestroy(p->s_comp);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  delete p;
}

int isr_net_time_conv(isr_net *net, int B, int h, int w, int precision, int scale, void *workspace,
                      size_t workspace_bytes, int iters, float *ms, void *stream) {
  if (ms == nullptr || iters <= 0 || (scale < 1 || scale > 5)) return fail(ISR_ERR_ARG, "time_conv: bad arguments");
  Plan *pl = nullptr;
  int rc = forward_common(net, B, h, w, precision, workspace, workspace_bytes, &pl);
  if (rc != ISR_OK) return rc;
  // scale 4: postblock.4 (64->64 at 4h x 4w, the dominant kernel); scale 1: first residual-block conv (h x w)
  // scale 5: the output conv, writing into scratch (c4) with the packed-input buffer standing in for the skip source
  // scale 3: the fused upsample + conv launch that produces the 4h x 4w tensor (needs a bilinear net)
  if (scale == 3 && !pl->fuse4.on) return fail(ISR_ERR_UNSUPPORTED, "time_conv: this plan has no fused upsample + conv");
  // scale 2: the second conv of the first residual block (h x w, with the residual addend)
  ConvLaunch L = scale == 3 ? pl->fuse4.up : pl->convs[scale == 5 ? pl->idx_final : (scale == 4 ? pl->idx_post4 : (net->G == 2 ? 2 : 1) + (scale == 2 ? 1 : 0))];
  if (scale == 5) {
    const size_t n_out = static_cast<size_t>(B) * net->cout * 16 * h * w;
    L.p.out_recon = reinterpret_cast<float *>(pl->c4);
    L.p.out_state = reinterpret_cast<float *>(pl->c4) + n_out;
    L.p.state_unshaded = net->cout == 6;
    L.p.skip_src = reinterpret_cast<const float *>(pl->in_g[0]);
    L.p.skip_bstride = static_cast<long long>(net->n_mask > 0 ? net->n_mask : 1) * h * w;
  }
  L.p.dbg = net->tun.dbg;
  unsigned long long *cyc = nullptr;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  float t = 0.f;
  auto body = [&]() -> int {
    if (net->tun.dbg & 32) {   // bench experiment: cycles of the MMA issue loop (clock64 in the issuing warp)
      ISR_CUDA(cudaMalloc(&cyc, 256 * sizeof(unsigned long long)));
      ISR_CUDA(cudaMemset(cyc, 0, 256 * sizeof(unsigned long long)));
      L.p.dbg_cycles = cyc;
    }
    ISR_CUDA(cudaEventCreate(&e0));
    ISR_CUDA(cudaEventCreate(&e1));
    int r;
    if ((r = run_conv(L, S(stream))) != ISR_OK) return r;  // warm-up
    ISR_CUDA(cudaEventRecord(e0, S(stream)));
    for (int i = 0; i < iters; ++i)
      if ((r = run_conv(L, S(stream))) != ISR_OK) return r;
    ISR_CUDA(cudaEventRecord(e1, S(stream)));
    ISR_CUDA(cudaEventSynchronize(e1));
    ISR_CUDA(cudaEventElapsedTime(&t, e0, e1));
    return ISR_OK;
  };
  rc = body();
  if (e0 != nullptr) cudaEventDestroy(e0);   // also on the error paths
  if (e1 != nullptr) cudaEventDestroy(e1);
  if (rc != ISR_OK) {
    if (cyc != nullptr) cudaFree(cyc);
    return rc;
  }
  *ms = t / iters;
  if (cyc != nullptr) {
    unsigned long long hc[256];
    const cudaError_t ce = cudaMemcpy(hc, cyc, sizeof hc, cudaMemcpyDeviceToHost);
    cudaFree(cyc);
    if (ce != cudaSuccess) return fail(ISR_ERR_CUDA, "time_conv: %s", cudaGetErrorString(ce));
    unsigned long long mx = 0, sum = 0; int n = 0;
    for (int i = 0; i < 128; ++i) if (hc[i]) { mx = hc[i] > mx ? hc[i] : mx; sum += hc[i]; ++n; }
    const double mmas = static_cast<double>((L.p.num_tiles + 1) / 2) / (L.grid / 2) * 36.0 * (L.p.n_act_parts == 2 ? 2.0 : 1.0);
    fprintf(stderr, "[isr dbg] MMA loop cycles: max %llu avg %.0f over %d pairs; %.1f cycles per k64-tap MMA group/4 (=per MMA) ; kernel %.1f us => %.0f MHz\n",
            mx, n ? static_cast<double>(sum) / n : 0.0, n, n ? static_cast<double>(sum) / n / mmas : 0.0, t / iters * 1e3,
            mx / (t / iters * 1e3));
  }
  return ISR_OK;
}

int isr_debug_conv_workspace_bytes(int B, int H, int W, size_t *bytes) {
  if (bytes == nullptr || B <= 0 || H <= 0 || W <= 0) return fail(ISR_ERR_ARG, "bad arguments");
  const size_t act = align_up(static_cast<size_t>(B) * H * W * 256, 1024);
  *bytes = 3 * act + align_up(2 * 2 * 9 * 64 * 64 * sizeof(uint16_t), 1024) + 1024;
  return ISR_OK;
}

int isr_debug_conv3x3(const float *x, const float *weight, const float *bias, const float *addend, float *out, int B,
                      int H, int W, int cout, int relu, int fmt, int n_act_parts, int n_w_parts, void *workspace,
                      size_t workspace_bytes, void *stream) {
  if (x == nullptr || weight == nullptr || out == nullptr || workspace == nullptr)
    return fail(ISR_ERR_ARG, "debug_conv: null argument");
  if (!(cout == 64 || (cout >= 1 && cout <= 16)) || (fmt != 0 && fmt != 1) || n_act_parts < 1 || n_act_parts > 2 ||
      n_w_parts < 1 || n_w_parts > 2 || (addend != nullptr && cout != 64))
    return fail(ISR_ERR_ARG, "debug_conv: unsupported configuration");
  size_t need = 0;
  isr_debug_conv_workspace_bytes(B, H, W, &need);
  if (workspace_bytes < need) return fail(ISR_ERR_WORKSPACE, "debug_conv: workspace too small (%zu < %zu)", workspace_bytes, need);
  if (reinterpret_cast<uintptr_t>(workspace) % 1024 != 0) return fail(ISR_ERR_ARG, "workspace must be 1024-byte aligned");
  int rc = check_device();
  if (rc != ISR_OK) return rc;
  if ((rc = setup_kernel_attrs()) != ISR_OK) return rc;
  cudaStream_t s = S(stream);
  const int is_bf16 = fmt == 0;
  const int npad = cout == 64 ? 64 : 16;
  const size_t act = align_up(static_cast<size_t>(B) * H * W * 256, 1024);
  uint8_t *base = static_cast<uint8_t *>(workspace);
  uint16_t *in16 = reinterpret_cast<uint16_t *>(base), *add16 = reinterpret_cast<uint16_t *>(base + act),
           *out16 = reinterpret_cast<uint16_t *>(base + 2 * act), *wp = reinterpret_cast<uint16_t *>(base + 3 * act);
  float *bias_d = reinterpret_cast<float *>(base + 3 * act + align_up(2 * 2 * 9 * 64 * 64 * sizeof(uint16_t), 1024));
  ISR_CUDA(launch_pack_nchw(x, in16, nullptr, B, 64, H, W, n_act_parts, is_bf16, s));
  if (addend != nullptr) ISR_CUDA(launch_pack_nchw(addend, add16, nullptr, B, 64, H, W, n_act_parts, is_bf16, s));
  ISR_CUDA(launch_pack_weights(weight, wp, cout, 64, npad, 1, s));
  ISR_CUDA(cudaMemsetAsync(bias_d, 0, 64 * sizeof(float), s));
  if (bias != nullptr) ISR_CUDA(cudaMemcpyAsync(bias_d, bias, cout * sizeof(float), cudaMemcpyDeviceToDevice, s));
  const size_t per_fmt = static_cast<size_t>(2) * 9 * npad * 64;
  ConvLaunch L;
  Tunables tun;
  {
    std::lock_guard<std::mutex> lock(g_tun_mutex);
    tun = g_tun;
  }
  rc = init_conv(tun, &L, npad, in16, B, H, W, n_act_parts, wp + (is_bf16 ? 0 : per_fmt), 2 * 9 * npad, 0, npad, n_w_parts,
                 is_bf16);
  if (rc != ISR_OK) return rc;
  if (cout == 64) {
    set_store(&L, bias_d, relu, addend != nullptr ? add16 : nullptr, n_act_parts, out16, 2);
    if ((rc = run_conv(L, s)) != ISR_OK) return rc;
    ISR_CUDA(launch_unpack_nhwc(out16, out, B, H, W, 2, is_bf16, s));
  } else {
    L.p.bias = bias_d;
    L.p.cout = cout;
    L.p.out_raw = out;
    L.p.h_low = H / 4 > 0 ? H / 4 : 1;
    L.p.w_low = W / 4 > 0 ? W / 4 : 1;
    if ((rc = run_conv(L, s)) != ISR_OK) return rc;
  }
  return ISR_OK;
}

}  // extern "C"
