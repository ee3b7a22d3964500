// C-ABI entry points (include/dsdf.h) of the B200-native DeepSDF hot path.
// Host side only orchestrates launches on the caller's stream; all arithmetic is in the
// sm_100a kernels (fwd3_kernel.cuh, bwd2_kernel.cuh, wgrad2_kernel.cuh, small_kernels.cuh,
// fold_kernels.cuh, recon_kernels.cuh, dense_kernels.cuh).  There is deliberately no CPU path: if no CUDA device / kernel image is
// available every entry fails with DSDF_ECUDA.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dsdf.h"
#include "bwd2_kernel.cuh"
#include "common.cuh"
#include "dense_kernels.cuh"
#include "fold_kernels.cuh"
#include "fwd3_kernel.cuh"
#include "layout.cuh"
#include "mc_kernels.cuh"
#include "recon_kernels.cuh"
#include "sampler_kernels.cuh"
#include "small_kernels.cuh"
#include "wgrad2_kernel.cuh"
#ifdef DSDF_DIAG_BUILD   // diagnostics variant only (build.py DSDF_BUILD_VARIANT): descriptor self-tests, MMA-rate probe
#include "selftest.cuh"
#endif

using namespace dsdf;

// partial-sum slots of the wgrad schedule (one per (CTA pair, work segment))
static inline int wg2_max_slots(int num_sms) { return 3 * (num_sms / 2) + 32; }
// segments per launch of the fold-mode inference entries (per-segment E steps: 32 KB each, handle-owned)
constexpr long long kInferSegs = 256;

struct dsdf_handle_s {
  dsdf_config cfg;
  int device = 0;
  int num_sms = 0;
  std::string err;
  // handle-owned device buffers (bf16 shadows + small fp32 derived state)
  uint8_t* w2_img = nullptr;     // forward K-step stream
  uint8_t* dg2_img = nullptr;    // dgrad K-step stream
  uint8_t* scene_e = nullptr;    // fold-mode per-segment E steps for the inference entries (kInferSegs, fixed)
  // tensor maps over the three handle-owned streams (encoded once at dsdf_create) + a 2-entry cache for the
  // caller-workspace scene_e buffers (training / reconstruct), re-encoded on the host only when base or size change
  // side stream of the fused step: the reductions that only need the dgrad chain's outputs (latent gradients, loss
  // scalars, bias / xyz / latent-column gradients) run beside the wgrad GEMMs; forked and joined with events on the
  // caller's stream, so the call is still ordered on that stream alone
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  CUtensorMap tmap_w{}, tmap_dg{}, tmap_e{};
  struct MapCache { const void* base = nullptr; size_t bytes = 0; CUtensorMap map{}; } ws_map[3];
  // A/B knobs: fixed in the product build; the diagnostics variant (-DDSDF_DIAG_BUILD) reads DSDF_FWD_CG,
  // DSDF_WG_QUAD, DSDF_DW8_WGRAD from the environment at dsdf_create
  int fwd_cg = 2;                // 1: never pair CTAs (cta_group::1 only)
  int dw8_wgrad = 0;             // 1: dW8 as a wgrad item instead of the forward epilogue
  int wg_quad = 1;               // 0: item-major wgrad schedule (every out-chunk pair re-reads X from HBM)
  unsigned long long* dbg = nullptr;   // diagnostics variant + DSDF_DBG=1: per-CTA cycle counters (dsdf_debug_read)
  float* weff = nullptr;
  float* scale = nullptr;
  float* inv_norm = nullptr;
  float4* wx = nullptr;          // [2][512] fp32 xyz columns of lin0 / lin4 (Eikonal backward)
  // wgrad schedule (rebuilt on the host when the tile count changes; fixed-capacity device buffers + pinned
  // staging, uploaded asynchronously on the caller's stream: no allocation, no synchronisation per call)
  WgSeg* wg2_segs = nullptr;
  int* wg2_seg_begin = nullptr;
  uint8_t* wg2_stage = nullptr;        // pinned host staging for both arrays
  cudaEvent_t wg2_uploaded = nullptr;  // last upload has been consumed by the copy engine
  long long wg2_tiles = -1;
  int wg2_extra_pairs = 0;       // CTA pairs that fit on the SMs left over by wg2_max_quads clusters (0: unknown / none)
  int wg2_quads = 0;             // clusters of the current multicast schedule
  int wg2_max_quads = 0;         // 4-CTA clusters of k_wgrad2<.., QUAD> the device can hold at once (GPC granularity)
  bool wg2_mc = false;           // schedule is for k_wgrad2<.., QUAD>: 4-CTA clusters, multicast X loads
  int wg2_with_l8 = -1;          // schedule currently holds a lin8 item (autograd entries) or not (fused step)
  int wg2_pairs = 0;
  int wg2_slots = 0;
  Wg2Layout wg2_layout{};
  ParamLayout pl{};
  const float* params_ptr = nullptr;   // caller's flat fp32 parameters at the last dsdf_set_weights
  bool weights_set = false;
  bool attrs_set = false;
  DropoutArgs drop{};                  // dsdf_set_dropout; layers == 0: no dropout
  // optional per-kernel timing (CUDA events on the caller's stream), see dsdf_profile_*
  bool profile = false;
  bool pdl = false;          // chain launches carry the programmatic-serialisation attribute (dsdf_reconstruct's loop)
  std::vector<cudaEvent_t> prof_events;   // one per section boundary of dsdf_train_fwd_bwd
  std::vector<int> prof_ids;              // section that ENDS at the event (-1: start of a call)
};
// sections of dsdf_train_fwd_bwd timed by dsdf_profile_*: 0 scene fold + forward chain, 1 dgrad chain,
// 2 Eikonal recompute pass, 3 Eikonal row-sum corrections, 4 wgrad GEMMs, 5 split-K reduction,
// 6 decoder-gradient small kernels, 7 latent-gradient small kernels + loss
constexpr int kProfSections = 8;
constexpr int kProfMaxEvents = 512 * 10;

static void prof_mark(dsdf_handle h, cudaStream_t st, int id) {
  if (!h->profile || (int)h->prof_events.size() >= kProfMaxEvents) return;
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, st);
  h->prof_events.push_back(e);
  h->prof_ids.push_back(id);
}

static thread_local std::string g_err;

#define DSDF_FAIL(h, code, ...)                         \
  do {                                                  \
    char _b[512];                                       \
    snprintf(_b, sizeof(_b), __VA_ARGS__);              \
    if (h) (h)->err = _b;                               \
    g_err = _b;                                         \
    return (code);                                      \
  } while (0)

#define CU(h, expr)                                                                    \
  do {                                                                                 \
    cudaError_t _e = (expr);                                                           \
    if (_e != cudaSuccess)                                                             \
      DSDF_FAIL(h, DSDF_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                __FILE__, __LINE__);                                                   \
  } while (0)

// ---- tensor maps for the CTA-pair weight loads (UTMALDG.2D.2CTA, ptx.cuh tma2_load_2d) ----------------------
// A linear byte stream seen as [rows of 1 KB][256 x u32], box {256, 4} = one 4 KB CTA slice of a K-step, no
// swizzle / interleave: the box lands in shared memory as the same 4 KB byte for byte.  The encoder is a host-side
// driver function without a runtime-API twin; it is fetched through the runtime (no libcuda link dependency).
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled encode_tiled_fn() {
  static PFN_encodeTiled fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (PFN_encodeTiled)p;
  }();
  return fn;
}
static int make_stream_map(dsdf_handle h, CUtensorMap* out, const void* base, size_t bytes) {
  PFN_encodeTiled enc = encode_tiled_fn();
  if (!enc) DSDF_FAIL(h, DSDF_ECUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if (((uintptr_t)base & 15) != 0 || bytes % 4096 != 0 || bytes == 0)
    DSDF_FAIL(h, DSDF_EINVAL, "stream map: base %p / size %zu must be 16-byte aligned / a multiple of 4 KB", base, bytes);
  const cuuint64_t dims[2] = {kMapRowBytes / 4, (cuuint64_t)(bytes / kMapRowBytes)};
  const cuuint64_t strides[1] = {kMapRowBytes};
  const cuuint32_t box[2] = {kMapRowBytes / 4, 4};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) DSDF_FAIL(h, DSDF_ECUDA, "cuTensorMapEncodeTiled failed (%d) for %p + %zu", (int)r, base, bytes);
  return DSDF_OK;
}
// map over a caller-workspace scene_e buffer (slot 0: training step, 1: segments autograd entry, 2: reconstruct)
static int ws_scene_map(dsdf_handle h, int slot, const void* base, size_t bytes, CUtensorMap* out) {
  auto& c = h->ws_map[slot];
  if (c.base != base || c.bytes != bytes) {
    int rc = make_stream_map(h, &c.map, base, bytes);
    if (rc != DSDF_OK) return rc;
    c.base = base;
    c.bytes = bytes;
  }
  *out = c.map;
  return DSDF_OK;
}

static int set_kernel_attrs(dsdf_handle h) {
  if (h->attrs_set) return DSDF_OK;
#define F3ATTR(MODE, FOLD, TRAIN)                                                                                   \
  CU(h, cudaFuncSetAttribute(k_forward3<MODE, FOLD, TRAIN, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3<1>::kSmemBytes)); \
  CU(h, cudaFuncSetAttribute(k_forward3<MODE, FOLD, TRAIN, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3<2>::kSmemBytes));
  F3ATTR(MODE_DENSE, false, false)
  F3ATTR(MODE_TABLE, false, false)
  F3ATTR(MODE_TABLE, true, false)
  F3ATTR(MODE_GRID, true, false)
  F3ATTR(MODE_TABLE, true, true)
  F3ATTR(MODE_DENSE, false, true)
#undef F3ATTR
  CU(h, cudaFuncSetAttribute(k_forward3<MODE_TABLE, true, true, 1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3<1>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_forward3<MODE_TABLE, true, true, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3<2>::kSmemBytes));
#ifndef DSDF_ARCH_SIREN
  CU(h, cudaFuncSetAttribute(k_forward3<MODE_TABLE, true, true, 1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3<1>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_forward3<MODE_TABLE, true, true, 2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, F3<2>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_backward2<1, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, B2<1>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_backward2<2, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, B2<2>::kSmemBytes));
#endif
  CU(h, cudaFuncSetAttribute(k_backward2<1, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, B2<1>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_backward2<2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, B2<2>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_backward2<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, B2<1>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_backward2<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, B2<2>::kSmemBytes));
  CU(h, cudaFuncSetAttribute(k_wgrad2<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kW2SmemBytes));
  CU(h, cudaFuncSetAttribute(k_wgrad2<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kW2SmemBytes));
  CU(h, (cudaFuncSetAttribute(k_wgrad2<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kW2SmemBytes)));
  CU(h, (cudaFuncSetAttribute(k_wgrad2<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kW2SmemBytes)));
  {
    // a 4-CTA cluster must sit inside one GPC, and the GPCs' SM counts are not all multiples of 4: fewer than
    // num_sms / 4 clusters fit at once.  The static wgrad schedule must not have more work ranges than that, or the
    // surplus clusters run as a second wave behind the first.
    cudaLaunchConfig_t oc;
    memset(&oc, 0, sizeof(oc));
    oc.gridDim = dim3((unsigned)(h->num_sms / 4 * 4));
    oc.blockDim = dim3(kW2Threads);
    oc.dynamicSmemBytes = kW2SmemBytes;
    cudaLaunchAttribute oa[1];
    oa[0].id = cudaLaunchAttributeClusterDimension;
    oa[0].val.clusterDim.x = 4;
    oa[0].val.clusterDim.y = 1;
    oa[0].val.clusterDim.z = 1;
    oc.attrs = oa;
    oc.numAttrs = 1;
    int nq = 0;
    if (cudaOccupancyMaxActiveClusters(&nq, k_wgrad2<true, true>, &oc) != cudaSuccess) { nq = 0; cudaGetLastError(); }
    h->wg2_max_quads = nq < h->num_sms / 4 ? nq : h->num_sms / 4;
    // the SMs those clusters leave over (GPC size mod 4) can still hold CTA pairs -- all of them if every GPC has an
    // even number of SMs, which is what "num_sms / 2 two-CTA clusters fit" says
    oc.gridDim = dim3((unsigned)(h->num_sms / 2 * 2));
    oa[0].val.clusterDim.x = 2;
    int np = 0;
    if (cudaOccupancyMaxActiveClusters(&np, k_wgrad2<true, false>, &oc) != cudaSuccess) { np = 0; cudaGetLastError(); }
    h->wg2_extra_pairs = (np == h->num_sms / 2 && h->wg2_max_quads >= 1) ? (h->num_sms - 4 * h->wg2_max_quads) / 2 : 0;
    h->wg2_extra_pairs &= ~1;                                  // whole (old-style) quads only
  }
  h->attrs_set = true;
  return DSDF_OK;
}

extern "C" int dsdf_create(const dsdf_config* cfg, dsdf_handle* out) {
  if (!cfg || !out) DSDF_FAIL((dsdf_handle) nullptr, DSDF_EINVAL, "null argument");
  if (cfg->latent_size != kLatent || cfg->hidden != kHidden || cfg->num_hidden != 8 ||
      cfg->latent_in != kSkipLayer)
    DSDF_FAIL((dsdf_handle) nullptr, DSDF_EUNSUPPORTED,
              "unsupported NetworkSpecs: fused path needs latent 256, dims [512]*8, latent_in [4] "
              "(got latent %d hidden %d layers %d latent_in %d)",
              cfg->latent_size, cfg->hidden, cfg->num_hidden, cfg->latent_in);
  if (cfg->weight_norm_mask & ~0xFFu)
    DSDF_FAIL((dsdf_handle) nullptr, DSDF_EUNSUPPORTED, "weight_norm on lin8 is not a reference configuration");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    DSDF_FAIL((dsdf_handle) nullptr, DSDF_ECUDA, "no CUDA device: %s (there is no CPU fallback)",
              cudaGetErrorString(e));
  dsdf_handle h = new dsdf_handle_s();
  h->cfg = *cfg;
  h->device = cfg->device;
  cudaDeviceProp prop;
  if ((e = cudaSetDevice(h->device)) != cudaSuccess ||
      (e = cudaGetDeviceProperties(&prop, h->device)) != cudaSuccess) {
    delete h;
    DSDF_FAIL((dsdf_handle) nullptr, DSDF_ECUDA, "device %d: %s", cfg->device, cudaGetErrorString(e));
  }
  if (prop.major != 10) {
    delete h;
    DSDF_FAIL((dsdf_handle) nullptr, DSDF_ECUDA,
              "device %d is sm_%d%d; this library is built for sm_100a (B200) only",
              cfg->device, prop.major, prop.minor);
  }
  h->num_sms = prop.multiProcessorCount;
#ifdef DSDF_DIAG_BUILD
  if (const char* e = getenv("DSDF_FWD_CG")) h->fwd_cg = atoi(e) == 1 ? 1 : 2;
  if (const char* e = getenv("DSDF_WG_QUAD")) h->wg_quad = atoi(e) != 0;
  if (const char* e = getenv("DSDF_DW8_WGRAD")) h->dw8_wgrad = atoi(e) != 0;
  if (const char* e = getenv("DSDF_DBG")) {
    if (atoi(e) != 0 && cudaMalloc(&h->dbg, (size_t)h->num_sms * kDbgSlots * 8) == cudaSuccess)
      cudaMemset(h->dbg, 0, (size_t)h->num_sms * kDbgSlots * 8);
  }
#endif
  const size_t sched_bytes = (size_t)wg2_max_slots(h->num_sms) * sizeof(WgSeg) + (size_t)(h->num_sms + 2) * sizeof(int);
  bool ok = cudaMalloc(&h->w2_img, (size_t)kW2Steps * kStepBytes) == cudaSuccess &&
            cudaMalloc(&h->dg2_img, (size_t)kDg2Steps * kStepBytes) == cudaSuccess &&
            cudaMalloc(&h->scene_e, (size_t)kInferSegs * kSceneESteps * kStepBytes) == cudaSuccess &&
            cudaMalloc(&h->wg2_segs, (size_t)wg2_max_slots(h->num_sms) * sizeof(WgSeg)) == cudaSuccess &&
            cudaMalloc(&h->wg2_seg_begin, (size_t)(h->num_sms + 2) * sizeof(int)) == cudaSuccess &&
            cudaHostAlloc((void**)&h->wg2_stage, sched_bytes, cudaHostAllocDefault) == cudaSuccess &&
            cudaEventCreateWithFlags(&h->wg2_uploaded, cudaEventDisableTiming) == cudaSuccess &&
            cudaStreamCreateWithFlags(&h->side, cudaStreamNonBlocking) == cudaSuccess &&
            cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming) == cudaSuccess &&
            cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming) == cudaSuccess &&
            cudaMalloc(&h->weff, (size_t)(kWeffElems + 16) * 4) == cudaSuccess &&
            cudaMalloc(&h->scale, (size_t)kScaleElems * 4) == cudaSuccess &&
            cudaMalloc(&h->inv_norm, (size_t)kScaleElems * 4) == cudaSuccess &&
            cudaMalloc(&h->wx, (size_t)2 * kHidden * sizeof(float4)) == cudaSuccess;
  if (!ok) {
    dsdf_destroy(h);
    DSDF_FAIL((dsdf_handle) nullptr, DSDF_ENOMEM, "cudaMalloc of weight shadows failed");
  }
  int rc = set_kernel_attrs(h);
  if (rc == DSDF_OK) rc = make_stream_map(h, &h->tmap_w, h->w2_img, (size_t)kW2Steps * kStepBytes);
  if (rc == DSDF_OK) rc = make_stream_map(h, &h->tmap_dg, h->dg2_img, (size_t)kDg2Steps * kStepBytes);
  if (rc == DSDF_OK) rc = make_stream_map(h, &h->tmap_e, h->scene_e, (size_t)kInferSegs * kSceneESteps * kStepBytes);
  if (rc != DSDF_OK) {
    std::string m = h->err;
    dsdf_destroy(h);
    g_err = m;
    return rc;
  }
  *out = h;
  return DSDF_OK;
}

extern "C" int dsdf_destroy(dsdf_handle h) {
  if (!h) return DSDF_OK;
  if (h->side) cudaStreamDestroy(h->side);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  cudaFree(h->w2_img);
  cudaFree(h->dg2_img);
  cudaFree(h->scene_e);
  cudaFree(h->weff);
  cudaFree(h->scale);
  cudaFree(h->inv_norm);
  cudaFree(h->wx);
  cudaFree(h->wg2_segs);
  cudaFree(h->wg2_seg_begin);
  if (h->wg2_stage) cudaFreeHost(h->wg2_stage);
  if (h->wg2_uploaded) cudaEventDestroy(h->wg2_uploaded);
  cudaFree(h->dbg);
  for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
  delete h;
  return DSDF_OK;
}

// which network family this build of the library implements (one source tree, two builds)
extern "C" const char* dsdf_arch(void) {
#ifdef DSDF_ARCH_SIREN
  return "siren_decoder";
#else
  return "deep_sdf_decoder";
#endif
}

extern "C" const char* dsdf_last_error(dsdf_handle h) { return h ? h->err.c_str() : g_err.c_str(); }
extern "C" int dsdf_num_sms(dsdf_handle h) { return h ? h->num_sms : 0; }

extern "C" int dsdf_set_weights(dsdf_handle h, const float* params, const dsdf_param_layout* lay,
                                void* stream) {
  if (!h || !params || !lay) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  cudaStream_t st = (cudaStream_t)stream;
  for (int l = 0; l < kNumLin; ++l) {
    const bool wn = (h->cfg.weight_norm_mask >> l) & 1u;
    if (wn != (lay->g_off[l] >= 0))
      DSDF_FAIL(h, DSDF_EINVAL, "layout/weight_norm_mask mismatch at lin%d", l);
    h->pl.w[l] = lay->w_off[l];
    h->pl.g[l] = lay->g_off[l];
    h->pl.b[l] = lay->b_off[l];
  }
  k_rowscale_weff<<<(kScaleElems * 32 + 255) / 256, 256, 0, st>>>(params, h->pl, h->scale, h->inv_norm, h->weff);
#ifndef DSDF_ARCH_SIREN
  const unsigned fold_blocks = kFoldFwdBlocks + kFoldDgBlocks + kFoldWxBlocks;
#else
  const unsigned fold_blocks = kFoldFwdBlocks + kFoldDgBlocks;
#endif
  k_fold_images<<<fold_blocks, 256, 0, st>>>(h->weff, params, h->pl, h->w2_img, h->dg2_img, h->wx);
  CU(h, cudaGetLastError());
  h->params_ptr = params;
  h->weights_set = true;
  return DSDF_OK;
}

// ---- v3 forward launcher: persistent grid, CTA pairs (cta_group::2) when `pair_ok` -------------
template <int MODE, bool FOLD, bool TRAIN, int CG, bool DROP = false, bool TAN = false>
static cudaError_t launch_fwd3_cg(dsdf_handle h, const FwdArgs& a, cudaStream_t st) {
  const long long tiles = (a.n_rows + kTileM - 1) / kTileM;
  long long g = tiles < h->num_sms ? tiles : h->num_sms;
  g = (g / CG) * CG;
  if (g < CG) g = CG;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)g);
  cfg.blockDim = dim3(kF3Threads);
  cfg.dynamicSmemBytes = F3<CG>::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CG;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = h->pdl ? 2 : 1;
  return cudaLaunchKernelEx(&cfg, k_forward3<MODE, FOLD, TRAIN, CG, DROP, TAN>, a);
}
// pair_ok: both tiles of a pair may share the B operand (always true unless FOLD segments are
// shorter than 256 rows-aligned, see callers)
template <int MODE, bool FOLD, bool TRAIN, bool DROP = false, bool TAN = false>
static cudaError_t launch_fwd(dsdf_handle h, const FwdArgs& a, cudaStream_t st, bool pair_ok) {
  const long long tiles = (a.n_rows + kTileM - 1) / kTileM;
  if (pair_ok && h->fwd_cg == 2 && tiles >= 2) return launch_fwd3_cg<MODE, FOLD, TRAIN, 2, DROP, TAN>(h, a, st);
  return launch_fwd3_cg<MODE, FOLD, TRAIN, 1, DROP, TAN>(h, a, st);
}
// common fields of every forward launch
static void fwd_args_init(dsdf_handle h, FwdArgs& a, long long n) {
  memset(&a, 0, sizeof(a));
  a.w8 = h->weff + weff_off(8);
  a.b8 = h->weff + kWeffElems;   // lin8 bias rides behind the effective weights (k_rowscale_weff)
  a.n_rows = n;
  a.w2img = h->w2_img;
  a.tmap_w = h->tmap_w;
  a.tmap_e = h->tmap_e;          // handle-owned scene_e (inference entries); workspace callers overwrite it
  a.dbg = h->dbg;
}

extern "C" int dsdf_forward_dense(dsdf_handle h, const float* input, int64_t n, float* out,
                                  void* stream) {
  if (!h || !input || !out) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (n <= 0) return DSDF_OK;
  FwdArgs a;
  fwd_args_init(h, a, n);
  a.input = input;
  a.out = out;
  CU(h, (launch_fwd<MODE_DENSE, false, false>(h, a, (cudaStream_t)stream, true)));
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

extern "C" int dsdf_forward_segments(dsdf_handle h, const float* table, const int64_t* seg_ids,
                                     int64_t n_seg, int64_t rows_per_seg, const float* xyz,
                                     int64_t xyz_stride, int64_t n, float* out, void* stream) {
  if (!h || !table || !xyz || !out) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (xyz_stride < 3) DSDF_FAIL(h, DSDF_EINVAL, "xyz_stride must be >= 3");
  if (seg_ids && (n_seg <= 0 || rows_per_seg <= 0 || n_seg * rows_per_seg < n))
    DSDF_FAIL(h, DSDF_EINVAL, "segments do not cover n rows");
  if (n <= 0) return DSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  FwdArgs a;
  fwd_args_init(h, a, n);
  a.table = table;
  a.seg_ids = (const long long*)seg_ids;
  a.rows_per_seg = seg_ids ? rows_per_seg : n;
  a.xyzs = xyz;
  a.xyz_stride = xyz_stride;
  a.out = out;
  if (!seg_ids || rows_per_seg % kTileM == 0) {
    // every 128-row tile lies inside one segment: fold W[:, latent] z into a per-segment bias.  The
    // per-segment E steps live in the handle's fixed kInferSegs-segment buffer: longer id lists go in chunks.
    const long long ns = seg_ids ? n_seg : 1;
    const bool pair_ok = !seg_ids || rows_per_seg % (2 * kTileM) == 0;
    a.scene_e = h->scene_e;
    for (long long s0 = 0; s0 < ns; s0 += kInferSegs) {
      const long long cs = ns - s0 < kInferSegs ? ns - s0 : kInferSegs;
      const long long r0 = s0 * a.rows_per_seg;
      if (r0 >= n) break;
      long long cn = cs * a.rows_per_seg;
      if (cn > n - r0) cn = n - r0;
      a.seg_ids = seg_ids ? (const long long*)seg_ids + s0 : nullptr;
      a.xyzs = xyz + r0 * xyz_stride;
      a.out = out + r0;
      a.n_rows = cn;
      k_scene_fold<<<dim3((unsigned)cs, 2, kFoldSplit), 512, 0, st>>>(h->weff, h->params_ptr, h->pl, table,
                                                                  a.seg_ids, h->scene_e);
      CU(h, (launch_fwd<MODE_TABLE, true, false>(h, a, st, pair_ok)));
    }
  } else {
    CU(h, (launch_fwd<MODE_TABLE, false, false>(h, a, st, true)));
  }
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

extern "C" int dsdf_decode_grid(dsdf_handle h, const float* latent, int32_t grid_n, int64_t lo,
                                int64_t hi, float* out, void* stream) {
  if (!h || !latent || !out) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  const long long n3 = (long long)grid_n * grid_n * grid_n;
  if (grid_n < 2 || lo < 0 || hi > n3 || lo > hi) DSDF_FAIL(h, DSDF_EINVAL, "bad grid range");
  if (hi == lo) return DSDF_OK;
  FwdArgs a;
  fwd_args_init(h, a, hi - lo);
  a.table = latent;
  a.grid_n = grid_n;
  a.grid_lo = lo;
  a.voxel_size = (float)(2.0 / (grid_n - 1));
  a.out = out;
  a.scene_e = h->scene_e;
  k_scene_fold<<<dim3(1, 2, kFoldSplit), 512, 0, (cudaStream_t)stream>>>(h->weff, h->params_ptr, h->pl, latent,
                                                             nullptr, h->scene_e);
  CU(h, (launch_fwd<MODE_GRID, true, false>(h, a, (cudaStream_t)stream, true)));
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

__global__ void k_grid_coords(int n, long long lo, long long cnt, float vs, float* xyz) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cnt) return;
  float x, y, z;
  grid_xyz(lo + i, n, vs, x, y, z);
  xyz[i * 3] = x;
  xyz[i * 3 + 1] = y;
  xyz[i * 3 + 2] = z;
}

extern "C" int dsdf_grid_coords(dsdf_handle h, int32_t grid_n, int64_t lo, int64_t hi, float* xyz,
                                void* stream) {
  if (!h || !xyz || grid_n < 2 || lo > hi) DSDF_FAIL(h, DSDF_EINVAL, "bad argument");
  const long long cnt = hi - lo;
  if (cnt == 0) return DSDF_OK;
  k_grid_coords<<<(unsigned)((cnt + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      grid_n, lo, cnt, (float)(2.0 / (grid_n - 1)), xyz);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// ------------------------------------------------------------------------------------ training

static inline int eik_corr_blocks(int num_sms) { return 4 * num_sms; }

namespace {
struct Ws {
  uint8_t* xstash;
  uint8_t* dystash;
  float* g;
  float* tile_partials;
  float* tile_colsum;
  float* cta_partials;
  float* wg_partials;
  float* seg_sum;
  float* dlat_seg;
  float* seg_norm;
  long long* ids;        // identity segment ids (segments autograd entry)
  float* dweff;
  uint8_t* scene_e;
  size_t scene_e_bytes;
  uint32_t* relu_mask;
  uint8_t* g_img;
  float4* eik_row;       // Eikonal: per-row (x', y', z', eps)
  float* eik_tile;       // Eikonal: per-tile {sum (1-||G||)^2, sum eps g}
  float* eik_corr;       // Eikonal: k_eik_corr block partials
  size_t total;
};
inline size_t al(size_t x) { return (x + 1023) & ~(size_t)1023; }


Ws carve(uint8_t* base, long long max_rows, long long max_seg, int num_sms) {
  const size_t tiles = (size_t)((max_rows + kTileM - 1) / kTileM);
  Ws w;
  size_t o = 0;
  auto take = [&](size_t bytes) { uint8_t* p = base ? base + o : nullptr; o += al(bytes); return p; };
  w.xstash = take(tiles * kStashTileBytes);
  w.dystash = take(tiles * kStashTileBytes);
  w.g = (float*)take(tiles * kTileM * 4);
  w.tile_partials = (float*)take(tiles * 2 * 4);
  w.tile_colsum = (float*)take(tiles * 2 * kHidden * 4);
  w.cta_partials = (float*)take((size_t)num_sms * kCtaPartialFloats * 4);
  {
    w.wg_partials = (float*)take((size_t)wg2_max_slots(num_sms) * 256 * kHidden * 4);
  }
  w.seg_sum = (float*)take((size_t)max_seg * 2 * kHidden * 4);
  w.dlat_seg = (float*)take((size_t)max_seg * kLatent * 4);
  w.seg_norm = (float*)take((size_t)max_seg * 4);
  w.ids = (long long*)take((size_t)max_seg * 8);
  w.dweff = (float*)take((size_t)kWeffElems * 4);
  w.scene_e_bytes = (size_t)max_seg * kSceneESteps * kStepBytes;
  w.scene_e = take(w.scene_e_bytes);
  w.relu_mask = (uint32_t*)take(tiles * kActAuxTileBytes);
  w.g_img = take(tiles * kSubBytes);
  w.eik_row = (float4*)take(tiles * kTileM * sizeof(float4));
  w.eik_tile = (float*)take(tiles * 2 * 4);
  w.eik_corr = (float*)take((size_t)eik_corr_blocks(num_sms) * kCorrFloats * 4);
  w.total = o;
  return w;
}
}  // namespace

extern "C" size_t dsdf_train_workspace_bytes(dsdf_handle h, int64_t max_rows, int64_t max_segments) {
  if (!h) return 0;
  return carve(nullptr, max_rows, max_segments, h->num_sms).total + 1024;
}

// v2 wgrad schedule: items (layer 1..8, 256-row out chunk) x tiles, cut into equal-cost contiguous
// ranges, one per CTA pair.  Cost per tile: 2 (N = 512) or 1 (lin4's x3 part, N = 256).
static int build_wg2_schedule(dsdf_handle h, long long tiles, cudaStream_t st, bool with_l8 = true) {
  if (h->wg2_tiles == tiles && h->wg2_with_l8 == (int)with_l8) return DSDF_OK;
  const int l_last = with_l8 ? 8 : 7;    // fused step: the forward epilogue accumulates dW8 itself
  struct Item { int layer, ocp, n_sub, cost; };
  std::vector<Item> items;
  for (int l = 1; l <= l_last; ++l) {
    h->wg2_layout.item_base[l] = (int)items.size();
    const int n_ocp = (l == 3 || l == 8) ? 1 : 2;
    for (int ocp = 0; ocp < n_ocp; ++ocp) items.push_back({l, ocp, l == 4 ? 4 : 8, l == 4 ? 1 : 2});
  }
  h->wg2_layout.item_base[0] = 0;
  int n_pairs = h->num_sms / 2;
  if (n_pairs < 1) n_pairs = 1;
  std::vector<WgSeg> segs;
  std::vector<int> seg_begin;
  // multicast variant (fused step: no lin8 item): the two pairs of a quad run as one 4-CTA cluster in lock step, so
  // every segment of a quad must give both pairs the same tile range
  const bool mc = !with_l8 && h->wg_quad && h->wg2_max_quads >= 1;
  // every cluster resident from the start; the left-over SMs run ordinary pairs (old-style quads: separate loads)
  const int mc_pairs = mc ? 2 * h->wg2_max_quads : 0;
  if (mc) n_pairs = mc_pairs + h->wg2_extra_pairs;
  if (h->wg_quad && n_pairs >= 2) {
    // "quad" schedule: two CTA pairs walk the SAME tile range of a layer side by side, one per
    // 256-row out chunk, so the X tiles the second pair needs are L2 hits behind the first pair's
    // HBM reads (the item-major schedule read X twice from HBM: 21.0 GB instead of 14.7 GB per step at
    // 1 M rows, ncu profiles/r01_ncu_full_v4_benchsize.txt, which made the kernel HBM-bound).
    // Layers with a single out chunk (lin3, lin8) split the tile range between the two pairs.
    const int n_quads = n_pairs / 2;
    auto quad_cost = [](int l) { return (l == 4 || l == 3 || l == 8) ? 1 : 2; };   // quad-time per tile
    long long total = 0;
    for (int l = 1; l <= l_last; ++l) total += (long long)quad_cost(l) * tiles;
    const long long target = (total + n_quads - 1) / n_quads;
    std::vector<std::vector<WgSeg>> per_pair(n_pairs);
    int l = 1;
    long long t0 = 0;
    for (int q = 0; q < n_quads && l <= l_last; ++q) {
      long long budget = target;
      while (l <= l_last && budget > 0) {
        const int c = quad_cost(l);
        long long take = (budget + c - 1) / c;
        if (take > tiles - t0) take = tiles - t0;
        const int n_sub = l == 4 ? 4 : 8;
        const bool two = !(l == 3 || l == 8);
        WgSeg a;
        a.layer = l; a.n_sub = n_sub; a.slot = 0;
        if (two) {
          a.ocp = 0; a.tile_lo = t0; a.tile_hi = t0 + take; per_pair[2 * q].push_back(a);
          a.ocp = 1; per_pair[2 * q + 1].push_back(a);
        } else if (mc && 2 * q < mc_pairs) {
          // lin3 (one out chunk): both pairs walk the whole range; pair 0 contracts X columns 0..255, pair 1 256..511
          // (n_sub 4 each) and they fill the two column halves of ONE partial slot
          a.ocp = 0; a.n_sub = 4; a.tile_lo = t0; a.tile_hi = t0 + take;
          per_pair[2 * q].push_back(a);
          per_pair[2 * q + 1].push_back(a);
        } else {
          const long long mid = t0 + (take + 1) / 2;
          a.ocp = 0; a.tile_lo = t0; a.tile_hi = mid; per_pair[2 * q].push_back(a);
          if (mid < t0 + take) { a.tile_lo = mid; a.tile_hi = t0 + take; per_pair[2 * q + 1].push_back(a); }
        }
        budget -= take * c;
        t0 += take;
        if (t0 == tiles) { ++l; t0 = 0; }
      }
    }
    if (l <= l_last) DSDF_FAIL(h, DSDF_EINVAL, "internal: wgrad schedule did not cover all layers");
    // slots: contiguous per item (layer, ocp), in item order
    std::vector<int> next_slot(items.size(), 0), base_slot(items.size(), 0), cnt(items.size(), 0);
    auto shares_slot = [&](int p, const WgSeg& sg) { return p < mc_pairs && sg.layer == 3 && (p & 1); };   // slot of pair p - 1
    for (int p = 0; p < n_pairs; ++p)
      for (WgSeg& sg : per_pair[p])
        if (!shares_slot(p, sg)) ++cnt[h->wg2_layout.item_base[sg.layer] + sg.ocp];
    for (size_t i = 1; i < items.size(); ++i) base_slot[i] = base_slot[i - 1] + cnt[i - 1];
    for (int p = 0; p < n_pairs; ++p) {
      seg_begin.push_back((int)segs.size());
      for (size_t k = 0; k < per_pair[p].size(); ++k) {
        WgSeg& sg = per_pair[p][k];
        const int item = h->wg2_layout.item_base[sg.layer] + sg.ocp;
        if (shares_slot(p, sg)) sg.slot = per_pair[p - 1][k].slot;   // mc: both lists have the same shape
        else sg.slot = base_slot[item] + next_slot[item]++;
        segs.push_back(sg);
      }
    }
    seg_begin.push_back((int)segs.size());
  } else {
  long long total = 0;
  for (const Item& it : items) total += (long long)it.cost * tiles;
  const long long target = (total + n_pairs - 1) / n_pairs;
  size_t ii = 0;
  long long t0 = 0;   // next unassigned tile of item ii
  for (int p = 0; p < n_pairs; ++p) {
    seg_begin.push_back((int)segs.size());
    long long budget = target;
    while (ii < items.size() && budget > 0) {
      const Item& it = items[ii];
      long long take = (budget + it.cost - 1) / it.cost;
      if (take > tiles - t0) take = tiles - t0;
      WgSeg sg;
      sg.layer = it.layer; sg.ocp = it.ocp; sg.n_sub = it.n_sub; sg.slot = (int)segs.size();
      sg.tile_lo = t0; sg.tile_hi = t0 + take;
      segs.push_back(sg);
      budget -= take * it.cost;
      t0 += take;
      if (t0 == tiles) { ++ii; t0 = 0; }
    }
  }
  seg_begin.push_back((int)segs.size());
  if (ii != items.size()) DSDF_FAIL(h, DSDF_EINVAL, "internal: wgrad schedule did not cover all items");
  }
  if ((int)segs.size() > wg2_max_slots(h->num_sms))
    DSDF_FAIL(h, DSDF_EINVAL, "internal: wgrad schedule has %d segments, partial buffer holds %d",
              (int)segs.size(), wg2_max_slots(h->num_sms));
  for (size_t i = 0; i < items.size(); ++i) { h->wg2_layout.slot_lo[i] = 1 << 30; h->wg2_layout.slot_hi[i] = 0; }
  for (const WgSeg& sg : segs) {
    const int item = h->wg2_layout.item_base[sg.layer] + sg.ocp;
    if (sg.slot < h->wg2_layout.slot_lo[item]) h->wg2_layout.slot_lo[item] = sg.slot;
    if (sg.slot + 1 > h->wg2_layout.slot_hi[item]) h->wg2_layout.slot_hi[item] = sg.slot + 1;
  }
  if ((int)seg_begin.size() > h->num_sms + 2) DSDF_FAIL(h, DSDF_EINVAL, "internal: wgrad schedule has too many CTA pairs");
  // upload through the pinned staging buffer, asynchronously on the caller's stream; the staging bytes may be
  // rewritten only after the previous upload has been consumed (an event wait that happens on a shape change only)
  CU(h, cudaEventSynchronize(h->wg2_uploaded));
  WgSeg* stage_segs = reinterpret_cast<WgSeg*>(h->wg2_stage);
  int* stage_begin = reinterpret_cast<int*>(h->wg2_stage + (size_t)wg2_max_slots(h->num_sms) * sizeof(WgSeg));
  memcpy(stage_segs, segs.data(), segs.size() * sizeof(WgSeg));
  memcpy(stage_begin, seg_begin.data(), seg_begin.size() * sizeof(int));
  CU(h, cudaMemcpyAsync(h->wg2_segs, stage_segs, segs.size() * sizeof(WgSeg), cudaMemcpyHostToDevice, st));
  CU(h, cudaMemcpyAsync(h->wg2_seg_begin, stage_begin, seg_begin.size() * sizeof(int), cudaMemcpyHostToDevice, st));
  CU(h, cudaEventRecord(h->wg2_uploaded, st));
  h->wg2_pairs = n_pairs;
  h->wg2_slots = (int)segs.size();
  h->wg2_tiles = tiles;
  h->wg2_with_l8 = (int)with_l8;
  h->wg2_mc = mc;
  h->wg2_quads = mc_pairs / 2;
  return DSDF_OK;
}

// wgrad GEMMs (lin1..lin8) over the stashed images + ordered split-K reduction into dweff
static int launch_wgrad2(dsdf_handle h, const uint8_t* xstash, const uint8_t* dystash, const uint8_t* g_img,
                         float* partials, float* dweff, cudaStream_t st, bool prof = false,
                         const float* g = nullptr, const uint32_t* relu_mask = nullptr) {
  Wg2Args w2;
  memset(&w2, 0, sizeof(w2));
  w2.segs = h->wg2_segs;
  w2.seg_begin = h->wg2_seg_begin;
  w2.xstash = xstash;
  w2.dystash = dystash;
  w2.g_img = g_img;
  w2.partials = partials;
  const bool rebuild = g != nullptr && relu_mask != nullptr;     // dY7 rebuilt in the kernel instead of read from the stash
  w2.g = g;
  w2.w8 = h->weff + weff_off(8);
  w2.relu_mask = relu_mask;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)(h->wg2_mc ? 4 * h->wg2_quads : 2 * h->wg2_pairs));
  cfg.blockDim = dim3(kW2Threads);
  cfg.dynamicSmemBytes = kW2SmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = h->wg2_mc ? 4 : 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (h->wg2_mc) {
    if (rebuild) CU(h, (cudaLaunchKernelEx(&cfg, k_wgrad2<true, true>, w2)));
    else CU(h, (cudaLaunchKernelEx(&cfg, k_wgrad2<false, true>, w2)));
    // The pairs for the SMs the clusters leave over: same stream, launched with the programmatic-serialisation
    // attribute and never waiting on the grid dependency.  They become eligible when every CTA of the cluster launch
    // has executed griddepcontrol.launch_dependents, i.e. when all clusters are resident in their GPC slots (had the
    // pairs been placed first -- a second stream cannot exclude that -- a GPC could come up one cluster short and
    // that cluster would run as a second wave), and then run beside them.  Everything they read was complete before
    // the cluster launch started; the next ordinary launch in the stream waits for both.
    const int extra = h->wg2_pairs - 2 * h->wg2_quads;
    if (extra > 0) {
      w2.pair_base = 2 * h->wg2_quads;
      cfg.gridDim = dim3((unsigned)(2 * extra));
      attr[0].val.clusterDim.x = 2;
      cfg.numAttrs = 2;
      if (rebuild) CU(h, cudaLaunchKernelEx(&cfg, k_wgrad2<true>, w2));
      else CU(h, cudaLaunchKernelEx(&cfg, k_wgrad2<false>, w2));
    }
  } else if (rebuild) CU(h, cudaLaunchKernelEx(&cfg, k_wgrad2<true>, w2));
  else CU(h, cudaLaunchKernelEx(&cfg, k_wgrad2<false>, w2));
  if (prof) prof_mark(h, st, 4);
  k_reduce_wgrad2<<<dim3(kHidden, h->wg2_with_l8 ? 8 : 7), kHidden, 0, st>>>(partials, h->wg2_layout, dweff);
  if (prof) prof_mark(h, st, 5);
  return DSDF_OK;
}

// persistent dgrad chain launch (CTA pairs when there are at least two tiles)
static int launch_backward2(dsdf_handle h, const Bwd2Args& b2, long long tiles, cudaStream_t st,
                            unsigned* grid_out) {
  const int cg = (h->fwd_cg == 2 && tiles >= 2) ? 2 : 1;
  long long g2 = tiles < h->num_sms ? tiles : h->num_sms;
  g2 = (g2 / cg) * cg;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3((unsigned)g2);
  cfg.blockDim = dim3(kB2Threads);
  cfg.dynamicSmemBytes = cg == 2 ? B2<2>::kSmemBytes : B2<1>::kSmemBytes;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cg;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = h->pdl ? 2 : 1;
  if (b2.eik_row) {
#ifndef DSDF_ARCH_SIREN
    if (cg == 2) CU(h, (cudaLaunchKernelEx(&cfg, k_backward2<2, false, true>, b2)));
    else CU(h, (cudaLaunchKernelEx(&cfg, k_backward2<1, false, true>, b2)));
#else
    DSDF_FAIL(h, DSDF_EUNSUPPORTED, "the Eikonal term is built for the ReLU decoder family only");
#endif
  } else if (b2.drop.layers != 0) {
    if (cg == 2) CU(h, (cudaLaunchKernelEx(&cfg, k_backward2<2, true>, b2)));
    else CU(h, (cudaLaunchKernelEx(&cfg, k_backward2<1, true>, b2)));
  } else {
    if (cg == 2) CU(h, (cudaLaunchKernelEx(&cfg, k_backward2<2>, b2)));
    else CU(h, (cudaLaunchKernelEx(&cfg, k_backward2<1>, b2)));
  }
  if (grid_out) *grid_out = (unsigned)g2;
  return DSDF_OK;
}

extern "C" int dsdf_train_fwd_bwd(dsdf_handle h, const float* params, const float* table,
                                  int64_t num_scenes, const int64_t* seg_ids, int64_t n_seg,
                                  int64_t rows_per_seg, const float* sdf_data, int64_t n_total,
                                  float clamp_dist, float reg_lambda_ramp, float eikonal_lambda,
                                  void* workspace, size_t workspace_bytes, float* grads, float* lat_grad,
                                  float* pred, float* loss_out, void* grads_ready_event, void* stream) {
  if (!h || !params || !table || !seg_ids || !sdf_data || !workspace || !grads || !lat_grad ||
      !pred || !loss_out)
    DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  const bool eik = eikonal_lambda != 0.f;
#ifdef DSDF_ARCH_SIREN
  if (eik) DSDF_FAIL(h, DSDF_EUNSUPPORTED, "UseEikonal is built for the ReLU decoder family only (sine layers have a second derivative)");
#endif
  if (eik && h->drop.layers != 0)
    DSDF_FAIL(h, DSDF_EUNSUPPORTED, "UseEikonal together with NetworkSpecs.dropout is not built");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (n_seg <= 0 || rows_per_seg <= 0 || rows_per_seg % kTileM != 0)
    DSDF_FAIL(h, DSDF_EUNSUPPORTED,
              "fused step needs scene-contiguous segments with SamplesPerScene %% 128 == 0 (got %lld)",
              (long long)rows_per_seg);
  if (n_total <= 0 || num_scenes <= 0) DSDF_FAIL(h, DSDF_EINVAL, "n_total and num_scenes must be positive");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = n_seg * rows_per_seg;
  const long long tiles = n / kTileM;
  const int tiles_per_seg = (int)(rows_per_seg / kTileM);
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  Ws w = carve(base, n, n_seg, h->num_sms);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes)
    DSDF_FAIL(h, DSDF_ENOMEM, "workspace too small: need %zu bytes, have %zu", w.total + 1024,
              workspace_bytes);
  // the forward epilogue accumulates dW8 = g^T X8 itself (no X8 stash, no lin8 wgrad item); the diagnostics
  // variant can keep the previous split (dW8 as a wgrad item) for A/B timing
  const bool dw8_fwd = !h->dw8_wgrad || eik;
  int rc = build_wg2_schedule(h, tiles, st, !dw8_fwd);
  if (rc != DSDF_OK) return rc;

  const float inv_n = 1.0f / (float)n_total;

  // ---- forward (gather + 9 layers + tanh + clamped L1 + dL/ds), stashing activations
  FwdArgs fa;
  fwd_args_init(h, fa, n);
  fa.table = table;
  fa.seg_ids = (const long long*)seg_ids;
  fa.rows_per_seg = rows_per_seg;
  fa.xyzs = sdf_data;
  fa.xyz_stride = 4;
  fa.out = pred;
  fa.xstash = w.xstash;
#ifdef DSDF_DBG_BUILD
  // diagnostics build only: time the chain kernels without their stash stores (results are then meaningless)
  if (const char* e = getenv("DSDF_DBG_NOSTASH")) if (atoi(e) & 1) fa.xstash = nullptr;
#endif
  fa.gout = w.g;
  fa.tile_partials = w.tile_partials;
  fa.fused_loss = 1;
  fa.clamp_dist = clamp_dist;
  fa.inv_n_total = inv_n;
  const bool prof = h->profile && (int)h->prof_events.size() + 10 <= kProfMaxEvents;
  if (prof) prof_mark(h, st, -1);
  fa.scene_e = w.scene_e;
  if (int mrc = ws_scene_map(h, 0, w.scene_e, w.scene_e_bytes, &fa.tmap_e)) return mrc;
  fa.relu_mask = w.relu_mask;
  fa.g_img = dw8_fwd ? nullptr : w.g_img;
  if (dw8_fwd && !eik) {
    fa.dw8_partials = w.cta_partials + kPartW8;
    fa.dw8_stride = kCtaPartialFloats;
  }
  // Eikonal (train_deep_sdf.py:724-729; tests/eikonal_algebra.py): this pass only decides the ReLU masks and the
  // L1 upstream; the activations the wgrad contracts are written by the recompute pass after the dgrad chain
  if (eik) fa.xstash = nullptr;
  const bool pair_ok = rows_per_seg % (2 * kTileM) == 0;
  // CTAs of the forward launch (launch_fwd3_cg): one dW8 partial each
  int fwd_ctas = 0;
  {
    const int cg = (pair_ok && h->fwd_cg == 2 && tiles >= 2) ? 2 : 1;
    long long g = tiles < h->num_sms ? tiles : h->num_sms;
    g = (g / cg) * cg;
    if (g < cg) g = cg;
    fwd_ctas = (int)g;
  }
  // seg_ids outside [0, num_scenes) would index the latent table out of bounds (nn.Embedding raises): the step
  // runs on a sanitised copy and reports the first bad position in loss_out[3]
  k_sanitize_ids<<<1, 256, 0, st>>>((const long long*)seg_ids, (int)n_seg, num_scenes, w.ids, loss_out + 3);
  seg_ids = (const int64_t*)w.ids;
  fa.seg_ids = w.ids;
  k_scene_fold<<<dim3((unsigned)n_seg, 2, kFoldSplit), 512, 0, st>>>(h->weff, params, h->pl, table,
                                                        (const long long*)seg_ids, w.scene_e);
  fa.drop = h->drop;
  if (h->drop.layers != 0)
    CU(h, (launch_fwd<MODE_TABLE, true, true, true>(h, fa, st, pair_ok)));
  else
    CU(h, (launch_fwd<MODE_TABLE, true, true>(h, fa, st, pair_ok)));
  if (prof) prof_mark(h, st, 0);

  // ---- backward dgrad chain + reducers
  Bwd2Args b2;
  memset(&b2, 0, sizeof(b2));
  b2.n_rows = n;
  b2.dg2img = h->dg2_img;
  b2.tmap_dg = h->tmap_dg;
  b2.w8 = h->weff + weff_off(8);
  b2.g = w.g;
  b2.relu_mask = w.relu_mask;
  b2.dystash = w.dystash;
#ifdef DSDF_DBG_BUILD
  if (const char* e = getenv("DSDF_DBG_NOSTASH")) if (atoi(e) & 2) b2.dystash = nullptr;
#endif
  b2.xyz = sdf_data;
  b2.xyz_stride = 4;
  b2.tile_colsum = w.tile_colsum;
  b2.cta_partials = w.cta_partials;
  b2.drop = h->drop;
  // dY7 = g w8^T . mask7 is rebuilt inside the wgrad kernel (ReLU family, no dropout): the dgrad chain then does not
  // stash it; with the Eikonal term k_eik_corr takes lin7's share of the corrections from the masks as well
#ifndef DSDF_ARCH_SIREN
  const bool rebuild_dy7 = h->drop.layers == 0;     // (the Eikonal correction pass takes lin7's share from the masks)
#else
  const bool rebuild_dy7 = false;
#endif
  b2.skip_dy7_stash = rebuild_dy7 ? 1 : 0;
  if (eik) {
    b2.pred = pred;
    b2.wx = h->wx;
    b2.clamp_dist = clamp_dist;
    b2.eik_c2 = 2.0f * eikonal_lambda * inv_n;
    b2.eik_row = w.eik_row;
    b2.eik_tile = w.eik_tile;
  }
  unsigned bwd_grid = 0;
  rc = launch_backward2(h, b2, tiles, st, &bwd_grid);
  if (rc != DSDF_OK) return rc;
  if (prof) prof_mark(h, st, 1);

  int n_corr = 0;
  if (eik) {
#ifndef DSDF_ARCH_SIREN
    // ---- recompute pass: the chain again over (xyz', 1 + eps) with the stored masks -> Xt stash, dW8
    FwdArgs ta = fa;
    ta.xstash = w.xstash;
    ta.eik_row = w.eik_row;
    ta.g_tan = w.g;
    ta.out = nullptr;
    ta.gout = nullptr;
    ta.tile_partials = nullptr;
    ta.fused_loss = 0;
    ta.dw8_partials = w.cta_partials + kPartW8;
    ta.dw8_stride = kCtaPartialFloats;
    CU(h, (launch_fwd<MODE_TABLE, true, true, false, true>(h, ta, st, pair_ok)));
    if (prof) prof_mark(h, st, 2);
    // ---- eps-weighted corrections of the row sums that are not GEMMs (bias grads, colsum(dY4), dW4[:, xyz])
    n_corr = eik_corr_blocks(h->num_sms);
    if (n_corr > tiles) n_corr = (int)tiles;
    k_eik_corr<<<n_corr, 256, 0, st>>>(w.dystash, w.eik_row, sdf_data, tiles, w.tile_colsum, w.eik_corr,
                                       rebuild_dy7 ? w.g : nullptr, w.relu_mask, h->weff + weff_off(8));
    if (prof) prof_mark(h, st, 3);
#endif
  }

  // ---- segmented (scene-contiguous) reductions that need only what the dgrad chain (and the Eikonal passes) left
  // behind: on the side stream, beside the wgrad GEMMs (they co-reside with the persistent wgrad CTAs: no shared
  // memory, few registers).  Every kernel writes its own elements with plain stores, fixed summation orders.
  const float reg_coef = reg_lambda_ramp * (float)rows_per_seg * inv_n;
  cudaStream_t sd = h->side;
  CU(h, cudaEventRecord(h->ev_fork, st));
  CU(h, cudaStreamWaitEvent(sd, h->ev_fork, 0));
  k_seg_colsum<<<dim3((unsigned)n_seg, 2), kHidden, 0, sd>>>(w.tile_colsum, tiles_per_seg, w.seg_sum);
  k_dw_latent<<<dim3(kHidden, 2), kLatent, 0, sd>>>(w.seg_sum, table, (const long long*)seg_ids,
                                                    (int)n_seg, w.dweff);
  k_reduce_cta<<<16, kHidden, 0, sd>>>(w.cta_partials, (int)bwd_grid, w.seg_sum, (int)n_seg,
                                      w.tile_partials, tiles, h->pl, grads, w.dweff, dw8_fwd ? fwd_ctas : 0,
                                      eik ? w.eik_corr : nullptr, n_corr, eik ? w.eik_tile : nullptr);
  k_lat_grad_seg<<<(unsigned)n_seg, 1024, 0, sd>>>(w.seg_sum, h->weff, table,
                                                      (const long long*)seg_ids, reg_coef,
                                                      w.dlat_seg, w.seg_norm);
  CU(h, cudaMemsetAsync(lat_grad, 0, (size_t)num_scenes * kLatent * 4, sd));
  k_lat_grad_scatter<<<(unsigned)n_seg, kLatent, 0, sd>>>(w.dlat_seg, (const long long*)seg_ids, (int)n_seg, lat_grad);
  k_loss<<<1, 1024, 0, sd>>>(w.tile_partials, tiles, inv_n, w.seg_norm, (int)n_seg, reg_coef, loss_out,
                             eik ? w.eik_tile : nullptr, eikonal_lambda);
  CU(h, cudaEventRecord(h->ev_join, sd));

  // ---- wgrad GEMMs (lin1..lin7) + split-K reduction on the caller's stream
  rc = launch_wgrad2(h, w.xstash, w.dystash, w.g_img, w.wg_partials, w.dweff, st, prof, rebuild_dy7 ? w.g : nullptr,
                     rebuild_dy7 ? w.relu_mask : nullptr);
  if (rc != DSDF_OK) return rc;
  CU(h, cudaStreamWaitEvent(st, h->ev_join, 0));
  // weight-norm backward over the complete effective-weight gradient; `grads` is final after it (a multi-GPU caller
  // all-reduces it behind `grads_ready_event`)
  k_param_grads<<<(kScaleElems * 32 + 255) / 256, 256, 0, st>>>(w.dweff, params, h->pl, h->scale,
                                                                h->inv_norm, grads);
  if (grads_ready_event) CU(h, cudaEventRecord((cudaEvent_t)grads_ready_event, st));
  if (prof) { prof_mark(h, st, 6); prof_mark(h, st, 7); }
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// ------------------------------------------------------------------------------- dense autograd
extern "C" size_t dsdf_dense_workspace_bytes(dsdf_handle h, int64_t max_rows) {
  if (!h || max_rows <= 0) return 0;
  return carve(nullptr, max_rows, 64, h->num_sms).total + 1024;
}

// Decoder.forward(input[N,259]) in a grad-enabled context: same forward chain, additionally keeping
// the activation stash + ReLU masks in `workspace` for dsdf_dense_backward.
extern "C" int dsdf_dense_forward_train(dsdf_handle h, const float* input, int64_t n, float* out,
                                        void* workspace, size_t workspace_bytes, void* stream) {
  if (!h || !input || !out || !workspace) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (n <= 0) return DSDF_OK;
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  Ws w = carve(base, n, 64, h->num_sms);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes)
    DSDF_FAIL(h, DSDF_ENOMEM, "workspace too small: need %zu bytes, have %zu", w.total + 1024, workspace_bytes);
  FwdArgs a;
  fwd_args_init(h, a, n);
  a.input = input;
  a.out = out;
  a.xstash = w.xstash;
  a.relu_mask = w.relu_mask;
  const long long tiles = (n + kTileM - 1) / kTileM;
  if (h->fwd_cg == 2 && tiles >= 2) CU(h, (launch_fwd3_cg<MODE_DENSE, false, true, 2>(h, a, (cudaStream_t)stream)));
  else CU(h, (launch_fwd3_cg<MODE_DENSE, false, true, 1>(h, a, (cudaStream_t)stream)));
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// autograd backward of Decoder.forward: given dOut[N] returns the parameter gradients (flat,
// dsdf_param_layout addressing, OVERWRITTEN) and, when dinput != NULL, dInput[N,259].
extern "C" int dsdf_dense_backward(dsdf_handle h, const float* params, const float* input, int64_t n,
                                   const float* out, const float* dout, void* workspace,
                                   size_t workspace_bytes, float* grads, float* dinput, void* stream) {
  if (!h || !params || !input || !out || !dout || !workspace || !grads) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (n <= 0) return DSDF_OK;
  cudaStream_t st = (cudaStream_t)stream;
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  Ws w = carve(base, n, 64, h->num_sms);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes)
    DSDF_FAIL(h, DSDF_ENOMEM, "workspace too small: need %zu bytes, have %zu", w.total + 1024, workspace_bytes);
  const long long tiles = (n + kTileM - 1) / kTileM;
  int rc = build_wg2_schedule(h, tiles, st);
  if (rc != DSDF_OK) return rc;
  k_dense_g<<<(unsigned)tiles, kTileM, 0, st>>>(dout, out, n, w.g, w.g_img, w.tile_partials);
  Bwd2Args b2;
  memset(&b2, 0, sizeof(b2));
  b2.n_rows = n;
  b2.dg2img = h->dg2_img;
  b2.tmap_dg = h->tmap_dg;
  b2.w8 = h->weff + weff_off(8);
  b2.g = w.g;
  b2.relu_mask = w.relu_mask;
  b2.dystash = w.dystash;
  b2.xyz = input + kLatent;
  b2.xyz_stride = kIn0;
  b2.tile_colsum = w.tile_colsum;
  b2.cta_partials = w.cta_partials;
  b2.stash_dy0 = 1;
  unsigned bwd_grid = 0;
  rc = launch_backward2(h, b2, tiles, st, &bwd_grid);
  if (rc != DSDF_OK) return rc;
  {
    Wg2Args w2;
    w2.segs = h->wg2_segs;
    w2.seg_begin = h->wg2_seg_begin;
    w2.xstash = w.xstash;
    w2.dystash = w.dystash;
    w2.g_img = w.g_img;
    w2.partials = w.wg_partials;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)(2 * h->wg2_pairs));
    cfg.blockDim = dim3(kW2Threads);
    cfg.dynamicSmemBytes = kW2SmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    CU(h, cudaLaunchKernelEx(&cfg, k_wgrad2<false>, w2));
    k_reduce_wgrad2<<<dim3(kHidden, 8), kHidden, 0, st>>>(w.wg_partials, h->wg2_layout, w.dweff);
  }
  // latent columns of dW0 / dW4 (wg_partials is free again: stream order)
  int n_split = (int)(tiles < 16 ? tiles : 16);
  k_dense_dw_latent<<<dim3(16, 2, (unsigned)n_split), 256, 0, st>>>(w.dystash, input, n, tiles, w.wg_partials);
  k_dense_dw_reduce<<<dim3(kHidden, 2), kLatent, 0, st>>>(w.wg_partials, n_split, w.dweff);
  if (dinput) k_dense_dinput<<<dim3((unsigned)tiles, 5), 256, 0, st>>>(w.dystash, h->weff, n, dinput);
  const int n_chunks = (int)(tiles < 64 ? tiles : 64);
  k_colsum_chunks<<<dim3((unsigned)n_chunks, 2), kHidden, 0, st>>>(w.tile_colsum, tiles, w.seg_sum);
  k_reduce_cta<<<16, kHidden, 0, st>>>(w.cta_partials, (int)bwd_grid, w.seg_sum, n_chunks, w.tile_partials, tiles,
                                      h->pl, grads, w.dweff, 0);
  k_param_grads<<<(kScaleElems * 32 + 255) / 256, 256, 0, st>>>(w.dweff, params, h->pl, h->scale, h->inv_norm, grads);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// ------------------------------------------------------------------------------- segments autograd
__global__ void k_iota(long long* ids, long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) ids[i] = i;
}

// decoder(cat([latents[b].expand(P), xyz[b]])) for b < n_seg in a grad-enabled context: the fold-mode
// forward (per-scene latent product in the bias) keeping stash + masks in `workspace`.
extern "C" int dsdf_segments_forward_train(dsdf_handle h, const float* params, const float* latents,
                                           int64_t n_seg, int64_t rows_per_seg, const float* xyz,
                                           int64_t xyz_stride, float* out, void* workspace,
                                           size_t workspace_bytes, void* stream) {
  if (!h || !params || !latents || !xyz || !out || !workspace) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (n_seg <= 0 || rows_per_seg <= 0 || rows_per_seg % kTileM != 0 || xyz_stride < 3)
    DSDF_FAIL(h, DSDF_EUNSUPPORTED, "segments need rows_per_seg %% 128 == 0 (got %lld) and xyz_stride >= 3",
              (long long)rows_per_seg);
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = n_seg * rows_per_seg;
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  Ws w = carve(base, n, n_seg, h->num_sms);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes)
    DSDF_FAIL(h, DSDF_ENOMEM, "workspace too small: need %zu bytes, have %zu", w.total + 1024, workspace_bytes);
  FwdArgs fa;
  fwd_args_init(h, fa, n);
  fa.table = latents;
  fa.seg_ids = w.ids;
  fa.rows_per_seg = rows_per_seg;
  fa.xyzs = xyz;
  fa.xyz_stride = xyz_stride;
  fa.out = out;
  fa.xstash = w.xstash;
  fa.relu_mask = w.relu_mask;
  fa.scene_e = w.scene_e;
  if (int mrc = ws_scene_map(h, 1, w.scene_e, w.scene_e_bytes, &fa.tmap_e)) return mrc;
  k_iota<<<(unsigned)((n_seg + 255) / 256), 256, 0, st>>>(w.ids, n_seg);
  k_scene_fold<<<dim3((unsigned)n_seg, 2, kFoldSplit), 512, 0, st>>>(h->weff, params, h->pl, latents, w.ids, w.scene_e);
  CU(h, (launch_fwd<MODE_TABLE, true, true>(h, fa, st, rows_per_seg % (2 * kTileM) == 0)));
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// backward of the above: dOut[n] -> grads (flat, OVERWRITTEN) and dlatents[n_seg, 256] (OVERWRITTEN).
extern "C" int dsdf_segments_backward(dsdf_handle h, const float* params, const float* latents,
                                      int64_t n_seg, int64_t rows_per_seg, const float* xyz,
                                      int64_t xyz_stride, const float* out, const float* dout,
                                      void* workspace, size_t workspace_bytes, float* grads,
                                      float* dlatents, void* stream) {
  if (!h || !params || !latents || !xyz || !out || !dout || !workspace || !grads || !dlatents)
    DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (n_seg <= 0 || rows_per_seg <= 0 || rows_per_seg % kTileM != 0)
    DSDF_FAIL(h, DSDF_EUNSUPPORTED, "segments need rows_per_seg %% 128 == 0");
  cudaStream_t st = (cudaStream_t)stream;
  const long long n = n_seg * rows_per_seg;
  const long long tiles = n / kTileM;
  const int tiles_per_seg = (int)(rows_per_seg / kTileM);
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  Ws w = carve(base, n, n_seg, h->num_sms);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes)
    DSDF_FAIL(h, DSDF_ENOMEM, "workspace too small: need %zu bytes, have %zu", w.total + 1024, workspace_bytes);
  int rc = build_wg2_schedule(h, tiles, st);
  if (rc != DSDF_OK) return rc;
  k_dense_g<<<(unsigned)tiles, kTileM, 0, st>>>(dout, out, n, w.g, w.g_img, w.tile_partials);
  Bwd2Args b2;
  memset(&b2, 0, sizeof(b2));
  b2.n_rows = n;
  b2.dg2img = h->dg2_img;
  b2.tmap_dg = h->tmap_dg;
  b2.w8 = h->weff + weff_off(8);
  b2.g = w.g;
  b2.relu_mask = w.relu_mask;
  b2.dystash = w.dystash;
  b2.xyz = xyz;
  b2.xyz_stride = xyz_stride;
  b2.tile_colsum = w.tile_colsum;
  b2.cta_partials = w.cta_partials;
  unsigned bwd_grid = 0;
  rc = launch_backward2(h, b2, tiles, st, &bwd_grid);
  if (rc != DSDF_OK) return rc;
  rc = launch_wgrad2(h, w.xstash, w.dystash, w.g_img, w.wg_partials, w.dweff, st);
  if (rc != DSDF_OK) return rc;
  k_seg_colsum<<<dim3((unsigned)n_seg, 2), kHidden, 0, st>>>(w.tile_colsum, tiles_per_seg, w.seg_sum);
  k_lat_grad_seg<<<(unsigned)n_seg, 1024, 0, st>>>(w.seg_sum, h->weff, latents, w.ids, 0.f, dlatents, w.seg_norm);
  k_dw_latent<<<dim3(kHidden, 2), kLatent, 0, st>>>(w.seg_sum, latents, w.ids, (int)n_seg, w.dweff);
  k_reduce_cta<<<16, kHidden, 0, st>>>(w.cta_partials, (int)bwd_grid, w.seg_sum, (int)n_seg, w.tile_partials, tiles,
                                       h->pl, grads, w.dweff, 0);
  k_param_grads<<<(kScaleElems * 32 + 255) / 256, 256, 0, st>>>(w.dweff, params, h->pl, h->scale, h->inv_norm, grads);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// ------------------------------------------------------------------------------- dropout
extern "C" int dsdf_set_dropout(dsdf_handle h, uint32_t layer_mask, float prob, uint64_t seed,
                                uint64_t offset, int64_t row_base) {
  if (!h) return DSDF_EINVAL;
  if (layer_mask & ~0xFFu) DSDF_FAIL(h, DSDF_EINVAL, "dropout layers must be hidden layers 0..7");
#ifdef DSDF_ARCH_SIREN
  if (layer_mask != 0 && prob > 0.f) DSDF_FAIL(h, DSDF_EUNSUPPORTED, "dropout is not built for the SIREN decoder");
#endif
  if (!(prob >= 0.f) || prob >= 1.f) DSDF_FAIL(h, DSDF_EINVAL, "dropout_prob must be in [0, 1)");
  memset(&h->drop, 0, sizeof(h->drop));
  if (layer_mask == 0 || prob == 0.f) return DSDF_OK;
  h->drop.layers = layer_mask;
  h->drop.thresh16 = (uint32_t)lrint((double)prob * 65536.0);
  h->drop.scale = 1.0f / (1.0f - prob);
  h->drop.seed_lo = (uint32_t)seed;
  h->drop.seed_hi = (uint32_t)(seed >> 32);
  h->drop.offset = (uint32_t)offset;
  h->drop.row_base = row_base;
  return DSDF_OK;
}

__global__ void k_dropout_keep(DropoutArgs d, int layer, long long row_lo, long long n_rows, uint8_t* out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // (row, chunk32)
  if (i >= n_rows * 16) return;
  const long long r = i >> 4;
  const int ch = (int)(i & 15);
  const uint32_t m = dropout_keep32(d, d.row_base + row_lo + r, layer, ch);
  for (int k = 0; k < 32; ++k) out[r * kHidden + ch * 32 + k] = (m >> (31 - k)) & 1u;
}

extern "C" int dsdf_dropout_keep(dsdf_handle h, int32_t layer, int64_t row_lo, int64_t n_rows,
                                 uint8_t* keep_out, void* stream) {
  if (!h || !keep_out || layer < 0 || layer > 7 || n_rows < 0) DSDF_FAIL(h, DSDF_EINVAL, "bad argument");
  if (n_rows == 0) return DSDF_OK;
  DropoutArgs d = h->drop;
  if (!((d.layers >> layer) & 1u)) d.thresh16 = 0;   // layer without dropout: everything kept
  k_dropout_keep<<<(unsigned)((n_rows * 16 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(d, layer, row_lo, n_rows, keep_out);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// ------------------------------------------------------------------------------- reconstruct
namespace {
struct RWs {
  float* sdf;            // [n_shapes][rows_pad][4]
  long long* ids;        // [n_shapes] identity segment ids
  float* g;
  float* pred;
  float* tile_partials;
  float* tile_colsum;
  float* cta_partials;
  uint8_t* scene_e;
  size_t scene_e_bytes;
  uint32_t* relu_mask;
  float* loss_sink;      // [n_shapes] when the caller wants no loss history
  size_t total;
};
inline long long recon_rows_pad(long long num_samples) {
  return (num_samples + 2 * kTileM - 1) / (2 * kTileM) * (2 * kTileM);   // whole CTA pairs
}
RWs recon_carve(uint8_t* base, long long n_shapes, long long num_samples, int num_sms) {
  const long long rows_pad = recon_rows_pad(num_samples);
  const size_t rows = (size_t)(n_shapes * rows_pad);
  const size_t tiles = rows / kTileM;
  RWs w;
  size_t o = 0;
  auto take = [&](size_t bytes) { uint8_t* p = base ? base + o : nullptr; o += al(bytes); return p; };
  w.sdf = (float*)take(rows * 16);
  w.ids = (long long*)take((size_t)n_shapes * 8);
  w.g = (float*)take(rows * 4);
  w.pred = (float*)take(rows * 4);
  w.tile_partials = (float*)take(tiles * 2 * 4);
  w.tile_colsum = (float*)take(tiles * 2 * kHidden * 4);
  w.cta_partials = (float*)take((size_t)num_sms * kCtaPartialFloats * 4);
  w.scene_e_bytes = (size_t)n_shapes * kSceneESteps * kStepBytes;
  w.scene_e = take(w.scene_e_bytes);
  w.relu_mask = (uint32_t*)take(tiles * kActAuxTileBytes);
  w.loss_sink = (float*)take((size_t)n_shapes * 4);
  w.total = o;
  return w;
}
}  // namespace

extern "C" size_t dsdf_reconstruct_workspace_bytes(dsdf_handle h, int64_t n_shapes, int64_t num_samples) {
  if (!h || n_shapes <= 0 || num_samples <= 0) return 0;
  return recon_carve(nullptr, n_shapes, num_samples, h->num_sms).total + 1024;
}

extern "C" int dsdf_reconstruct(dsdf_handle h, const float* params, const dsdf_reconstruct_cfg* cfg,
                                int64_t n_shapes, const float* samples, const int64_t* shape_base,
                                const int64_t* windows, int64_t first_step, int64_t n_steps,
                                float* latents, float* exp_avg, float* exp_avg_sq, void* workspace,
                                size_t workspace_bytes, float* loss_hist, void* stream) {
  if (!h || !params || !cfg || !samples || !shape_base || !windows || !latents || !exp_avg ||
      !exp_avg_sq || !workspace)
    DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->weights_set) DSDF_FAIL(h, DSDF_EINVAL, "dsdf_set_weights has not been called");
  if (n_shapes <= 0 || cfg->num_samples < 2 || (cfg->num_samples & 1))
    DSDF_FAIL(h, DSDF_EINVAL, "num_samples must be even and >= 2 (half pos, half neg; got %lld)",
              (long long)cfg->num_samples);
  if (first_step < 1 || n_steps < 0 || first_step - 1 + n_steps > cfg->num_iterations)
    DSDF_FAIL(h, DSDF_EINVAL, "iterations [%lld, %lld) outside 1..num_iterations",
              (long long)first_step, (long long)(first_step + n_steps));
  if (cfg->reg_kind < 0 || cfg->reg_kind > 2 || cfg->dist_kind < 0 || cfg->dist_kind > 2)
    DSDF_FAIL(h, DSDF_EINVAL, "unknown regulariser kind");
  cudaStream_t st = (cudaStream_t)stream;
  const long long num = cfg->num_samples;
  const long long rows_pad = recon_rows_pad(num);
  const long long n = n_shapes * rows_pad;
  const long long tiles = n / kTileM;
  const int tiles_per_seg = (int)(rows_pad / kTileM);
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 1023) & ~(uintptr_t)1023);
  RWs w = recon_carve(base, n_shapes, num, h->num_sms);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes)
    DSDF_FAIL(h, DSDF_ENOMEM, "workspace too small: need %zu bytes, have %zu", w.total + 1024,
              workspace_bytes);

  FwdArgs fa;
  fwd_args_init(h, fa, n);
  fa.table = latents;
  fa.seg_ids = w.ids;
  fa.rows_per_seg = rows_pad;
  fa.valid_per_seg = num;
  fa.xyzs = w.sdf;
  fa.xyz_stride = 4;
  fa.out = w.pred;
  fa.gout = w.g;
  fa.tile_partials = w.tile_partials;
  fa.fused_loss = 1;
  fa.clamp_dist = cfg->clamp_dist;
  fa.inv_n_total = 1.0f / (float)num;          // L1Loss() = mean over num_samples (reconstruct.py:102)
  fa.scene_e = w.scene_e;
  if (int mrc = ws_scene_map(h, 2, w.scene_e, w.scene_e_bytes, &fa.tmap_e)) return mrc;
  fa.relu_mask = w.relu_mask;                   // xstash / g_img stay null: frozen decoder, no wgrad

  Bwd2Args b2;
  memset(&b2, 0, sizeof(b2));
  b2.n_rows = n;
  b2.dg2img = h->dg2_img;
  b2.tmap_dg = h->tmap_dg;
  b2.w8 = h->weff + weff_off(8);
  b2.g = w.g;
  b2.relu_mask = w.relu_mask;
  b2.xyz = w.sdf;
  b2.xyz_stride = 4;
  b2.tile_colsum = w.tile_colsum;
  b2.cta_partials = w.cta_partials;

  ReconTailArgs ta;
  memset(&ta, 0, sizeof(ta));
  ta.tile_colsum = w.tile_colsum;
  ta.tile_partials = w.tile_partials;
  ta.tiles_per_seg = tiles_per_seg;
  ta.inv_n = 1.0f / (float)num;
  ta.reg_kind = cfg->reg_lambda > 0.f ? cfg->reg_kind : 0;
  ta.reg_lambda = cfg->reg_lambda;
  ta.dist_kind = (cfg->dist_weight > 0.f && cfg->dist_mean) ? cfg->dist_kind : 0;
  ta.dist_weight = cfg->dist_weight;
  ta.dist_mean = cfg->dist_mean;
  ta.dist_std = cfg->dist_std;
  ta.b1 = 0.9f;
  ta.b2 = 0.999f;
  ta.eps = 1e-8f;
  ta.code_bound = cfg->code_bound;
  ta.latents = latents;
  ta.exp_avg = exp_avg;
  ta.exp_avg_sq = exp_avg_sq;
  ta.loss_out = w.loss_sink;
  ta.weff = h->weff;
  ta.params = params;
  ta.pl = h->pl;
  ta.scene_e = w.scene_e;
  ta.samples = (const float4*)samples;
  ta.shape_base = (const long long*)shape_base;
  ta.half = num / 2;
  ta.rows_pad = rows_pad;
  ta.sdf = (float4*)w.sdf;
  ta.ids = w.ids;

  // k_recon_tail: one cluster of kReconParts CTAs per shape.  Inside the loop every launch carries the
  // programmatic-serialisation attribute: the next grid's prologue (barrier init, TMEM allocation, cluster
  // handshake) runs on idle SMs under the current grid, and its griddepcontrol.wait orders the data.
  auto launch_tail = [&](bool do_update, bool do_next, long long e_next, bool pdl) -> cudaError_t {
    ta.do_update = do_update;
    ta.do_next = do_next;
    ta.win = (const long long*)windows + (do_next ? e_next : 0) * n_shapes * 2;
    cudaLaunchConfig_t lc;
    memset(&lc, 0, sizeof(lc));
    lc.gridDim = dim3(kReconParts, (unsigned)n_shapes);
    lc.blockDim = dim3(kReconThreads);
    lc.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kReconParts;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = attr;
    lc.numAttrs = pdl ? 2 : 1;
    return cudaLaunchKernelEx(&lc, k_recon_tail, ta);
  };
  if (n_shapes > 65535) DSDF_FAIL(h, DSDF_EINVAL, "more than 65535 shapes in one call");
  if (n_steps == 0) return DSDF_OK;
  long long adjust_every = cfg->num_iterations / 2;
  if (adjust_every < 1) adjust_every = 1;
  // the first launch of the call is ordered the ordinary way behind whatever the caller enqueued before
  CU(h, launch_tail(false, true, first_step - 1, false));
#ifndef DSDF_NO_PDL
  h->pdl = true;
#endif
  int rc = DSDF_OK;
  cudaError_t ce = cudaSuccess;
  for (long long i = 0; i < n_steps && rc == DSDF_OK && ce == cudaSuccess; ++i) {
    const long long e = first_step - 1 + i;     // 0-based iteration (reconstruct.py:76)
    ce = launch_fwd<MODE_TABLE, true, true>(h, fa, st, true);
    if (ce != cudaSuccess) break;
    rc = launch_backward2(h, b2, tiles, st, nullptr);
    if (rc != DSDF_OK) break;
    const double lr = (double)cfg->lr * pow(0.1, (double)(e / adjust_every));
    const double bc1 = 1.0 - pow(0.9, (double)(e + 1));
    const double bc2 = 1.0 - pow(0.999, (double)(e + 1));
    ta.step_size = (float)(lr / bc1);
    ta.bc2_sqrt = (float)sqrt(bc2);
    ta.loss_out = loss_hist ? loss_hist + e * n_shapes : w.loss_sink;
    ce = launch_tail(true, i + 1 < n_steps, e + 1, h->pdl);
  }
  h->pdl = false;
  if (rc != DSDF_OK) return rc;
  CU(h, ce);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

extern "C" int dsdf_profile_enable(dsdf_handle h, int32_t enable) {
  if (!h) return DSDF_EINVAL;
  for (cudaEvent_t e : h->prof_events) cudaEventDestroy(e);
  h->prof_events.clear();
  h->prof_ids.clear();
  h->profile = enable != 0;
  return DSDF_OK;
}

extern "C" int dsdf_profile_read(dsdf_handle h, float* ms_sum, int32_t* n_calls) {
  if (!h || !ms_sum || !n_calls) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  for (int k = 0; k < kProfSections; ++k) ms_sum[k] = 0.f;
  int calls = 0;
  if (!h->prof_events.empty()) CU(h, cudaEventSynchronize(h->prof_events.back()));
  for (size_t i = 0; i < h->prof_events.size(); ++i) {
    const int id = h->prof_ids[i];
    if (id < 0) { ++calls; continue; }
    if (i == 0 || id >= kProfSections) continue;
    float ms = 0.f;
    CU(h, cudaEventElapsedTime(&ms, h->prof_events[i - 1], h->prof_events[i]));
    ms_sum[id] += ms;
  }
  *n_calls = calls;
  return DSDF_OK;
}

extern "C" int dsdf_debug_read(dsdf_handle h, uint64_t* out, int32_t max_words) {
  if (!h || !out) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (!h->dbg) DSDF_FAIL(h, DSDF_EINVAL, "diagnostic counters are off (set DSDF_DBG=1 before dsdf_create)");
  int n = h->num_sms * kDbgSlots;
  if (n > max_words) n = max_words;
  CU(h, cudaDeviceSynchronize());
  CU(h, cudaMemcpy(out, h->dbg, (size_t)n * 8, cudaMemcpyDeviceToHost));
  CU(h, cudaMemset(h->dbg, 0, (size_t)h->num_sms * kDbgSlots * 8));
  return n;
}

extern "C" int dsdf_renorm_rows(dsdf_handle h, float* table, int64_t num_scenes, const int64_t* seg_ids,
                                int64_t n_seg, float max_norm, void* stream) {
  if (!h || !table || !seg_ids) DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (n_seg <= 0) return DSDF_OK;
  k_renorm_rows<<<(unsigned)n_seg, kLatent, 0, (cudaStream_t)stream>>>(table, (const long long*)seg_ids, (int)n_seg,
                                                                   num_scenes, max_norm);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

extern "C" int dsdf_clip_adam(dsdf_handle h, float* params, float* grads, float* exp_avg,
                              float* exp_avg_sq, int64_t n, float max_norm, float lr, float beta1,
                              float beta2, float eps, int64_t step, float* scratch, void* stream) {
  if (!h || !params || !grads || !exp_avg || !exp_avg_sq || !scratch)
    DSDF_FAIL(h, DSDF_EINVAL, "null argument");
  if (step < 1) DSDF_FAIL(h, DSDF_EINVAL, "Adam step count is 1-based");
  cudaStream_t st = (cudaStream_t)stream;
  const int nb = 1024;
  float* part = scratch + 16;
  if (max_norm > 0.f)     // the reference computes the norm only when it clips (train_deep_sdf.py:780-782)
    k_sqnorm_partial<<<nb, 256, 0, st>>>(grads, n, part);
  const double bc1 = 1.0 - pow((double)beta1, (double)step);
  const double bc2 = 1.0 - pow((double)beta2, (double)step);
  const float step_size = (float)((double)lr / bc1);
  const float bc2_sqrt = (float)sqrt(bc2);
  const unsigned blocks = (unsigned)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  k_adam<<<blocks, 256, 0, st>>>(params, grads, exp_avg, exp_avg_sq, n, max_norm > 0.f ? part : nullptr, nb, max_norm,
                                 scratch, step_size, bc2_sqrt, beta1, beta2, eps);
  CU(h, cudaGetLastError());
  return DSDF_OK;
}

// ------------------------------------------------------------------------------- iso-surface extraction
static int mc_check(const float* sdf, int32_t nx, int32_t ny, int32_t nz) {
  if (!sdf) { g_err = "null argument"; return DSDF_EINVAL; }
  if (nx < 2 || ny < 2 || nz < 2) { g_err = "marching cubes needs at least 2 grid points per axis"; return DSDF_EINVAL; }
  if ((long long)nx * ny * nz > (1ll << 31)) { g_err = "grid too large (more than 2^31 points)"; return DSDF_EINVAL; }
  return DSDF_OK;
}
extern "C" int dsdf_mc_classify(const float* sdf, int32_t nx, int32_t ny, int32_t nz, float level, uint8_t* eflags,
                                int64_t* counts, void* stream) {
  int rc = mc_check(sdf, nx, ny, nz);
  if (rc != DSDF_OK) return rc;
  if (!eflags || !counts) { g_err = "null argument"; return DSDF_EINVAL; }
  const long long n = (long long)nx * ny * nz;
  McGrid g{nx, ny, nz, level};
  k_mc_classify<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(sdf, g, eflags, (long long*)counts);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  return DSDF_OK;
}
extern "C" int dsdf_mc_emit(const float* sdf, int32_t nx, int32_t ny, int32_t nz, float level, float spacing_x,
                            float spacing_y, float spacing_z, const uint8_t* eflags, const int64_t* scan, float* verts,
                            int32_t* faces, void* stream) {
  int rc = mc_check(sdf, nx, ny, nz);
  if (rc != DSDF_OK) return rc;
  if (!eflags || !scan) { g_err = "null argument"; return DSDF_EINVAL; }
  const long long n = (long long)nx * ny * nz;
  McGrid g{nx, ny, nz, level};
  k_mc_emit<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(sdf, g, spacing_x, spacing_y, spacing_z, eflags,
                                                                          (const long long*)scan, verts, faces);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  return DSDF_OK;
}

// ------------------------------------------------------------------------------- SDF-sample generation
#define MC_CU(expr)                                                                   \
  do {                                                                                \
    cudaError_t _e = (expr);                                                          \
    if (_e != cudaSuccess) { g_err = cudaGetErrorString(_e); return DSDF_ECUDA; }     \
  } while (0)
extern "C" int dsdf_rasterize_views(const float* verts, int64_t n_verts, const int32_t* faces, int64_t n_faces,
                                    const float* views, int32_t n_views, uint64_t* zbuf, float* out_vert,
                                    float* out_norm, uint8_t* valid, void* stream) {
  if (!verts || !faces || !views || !zbuf || !out_vert || !out_norm || !valid) { g_err = "null argument"; return DSDF_EINVAL; }
  if (n_verts <= 0 || n_faces <= 0 || n_views <= 0 || n_faces > (1ll << 31) - 2) { g_err = "empty mesh / bad sizes"; return DSDF_EINVAL; }
  cudaStream_t st = (cudaStream_t)stream;
  const size_t px = (size_t)n_views * kRastW * kRastH;
  MC_CU(cudaMemsetAsync(zbuf, 0xFF, px * 8, st));
  k_rast_draw<<<dim3((unsigned)((n_faces + 127) / 128), (unsigned)n_views), 128, 0, st>>>(verts, faces, (int)n_faces, views,
                                                                                       (unsigned long long*)zbuf);
  k_rast_resolve<<<dim3((kRastW * kRastH + 255) / 256, (unsigned)n_views), 256, 0, st>>>(
      verts, faces, views, (const unsigned long long*)zbuf, (float4*)out_vert, (float4*)out_norm, valid);
  MC_CU(cudaGetLastError());
  return DSDF_OK;
}
// Allocation-free variant of dsdf_rasterize_views (no counterpart in the reference; added
// for callers that may not call cudaMalloc: graph capture, serving loops).  The depth buffer lives
// in a file-scope device array instead of a caller buffer: kRastStaticMaxViews views x 400 x 400 x 8 B = 10.24 MB,
// held by every GPU that loads the library until process exit.  In HOST code `g_rast_zbuf` is only the symbol's
// host-side stand-in, never a device address: the pointer handed to the kernels comes from cudaGetSymbolAddress,
// taken once per device with that device current.  All calls on a device share the array (it keeps the last
// call's contents, so it is reset to "empty" = all ones at the start of every call); two calls whose kernels
// overlap on the GPU (different streams, no ordering between them) race on it.
constexpr int kRastStaticMaxViews = DSDF_RAST_STATIC_MAX_VIEWS;
static __device__ __align__(256) unsigned long long g_rast_zbuf[(size_t)kRastStaticMaxViews * kRastW * kRastH];
constexpr int kMaxDevices = 64;
static unsigned long long* g_rast_zbuf_dev[kMaxDevices];      // per-device address of g_rast_zbuf (null = not taken yet)
static std::mutex g_rast_zbuf_mu;

static int rast_static_scratch(unsigned long long** out) {
  int dev = -1;
  MC_CU(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) { g_err = "device ordinal beyond the static scratch table"; return DSDF_EINVAL; }
  std::lock_guard<std::mutex> lock(g_rast_zbuf_mu);
  if (!g_rast_zbuf_dev[dev]) {
    unsigned long long* p = nullptr;
    MC_CU(cudaGetSymbolAddress((void**)&p, g_rast_zbuf));     // the address on the CURRENT device
    if (!p) { g_err = "cudaGetSymbolAddress returned a null scratch address"; return DSDF_ECUDA; }
    g_rast_zbuf_dev[dev] = p;
  }
  *out = g_rast_zbuf_dev[dev];
  return DSDF_OK;
}
extern "C" int dsdf_rasterize_views_static(const float* verts, int64_t n_verts, const int32_t* faces, int64_t n_faces,
                                           const float* views, int32_t n_views, float* out_vert, float* out_norm,
                                           uint8_t* valid, void* stream) {
  if (!verts || !faces || !views || !out_vert || !out_norm || !valid) { g_err = "null argument"; return DSDF_EINVAL; }
  if (n_verts <= 0 || n_faces <= 0 || n_views <= 0 || n_faces > (1ll << 31) - 2) { g_err = "empty mesh / bad sizes"; return DSDF_EINVAL; }
  if (n_views > kRastStaticMaxViews) {
    g_err = "dsdf_rasterize_views_static: n_views = " + std::to_string(n_views) + " exceeds the static depth buffer (" +
            std::to_string(kRastStaticMaxViews) + " views per call); split the views or use dsdf_rasterize_views";
    return DSDF_EINVAL;
  }
  unsigned long long* zbuf = nullptr;
  int rc = rast_static_scratch(&zbuf);
  if (rc != DSDF_OK) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t px = (size_t)n_views * kRastW * kRastH;
  MC_CU(cudaMemsetAsync(zbuf, 0xFF, px * 8, st));               // the array keeps the previous call's depths
  k_rast_draw<<<dim3((unsigned)((n_faces + 127) / 128), (unsigned)n_views), 128, 0, st>>>(verts, faces, (int)n_faces, views, zbuf);
  k_rast_resolve<<<dim3((kRastW * kRastH + 255) / 256, (unsigned)n_views), 256, 0, st>>>(
      verts, faces, views, zbuf, (float4*)out_vert, (float4*)out_norm, valid);
  MC_CU(cudaGetLastError());
  return DSDF_OK;
}
extern "C" int dsdf_sample_points(const float* verts, const int32_t* faces, const float* cdf, int64_t n_faces,
                                  int64_t n_surf, int64_t n_rand, float stdv1, float stdv2, float cube_dim, uint64_t seed,
                                  float* xyz, float* surf_out, void* stream) {
  if (!verts || !faces || !cdf || !xyz) { g_err = "null argument"; return DSDF_EINVAL; }
  if (n_faces <= 0 || n_surf < 0 || n_rand < 0 || n_surf + n_rand == 0) { g_err = "bad sizes"; return DSDF_EINVAL; }
  const long long n = n_surf + n_rand;
  k_sample_points<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      verts, faces, cdf, (int)n_faces, n_surf, n_rand, stdv1, stdv2, cube_dim, (uint32_t)seed, (uint32_t)(seed >> 32), xyz,
      surf_out);
  MC_CU(cudaGetLastError());
  return DSDF_OK;
}
extern "C" int dsdf_knn_sdf(const float* xyz, int64_t n, const float* pts, const float* normals, int64_t n_pts, float stdv,
                            float* sdf_out, int32_t* nn_out, void* stream) {
  if (!xyz || !pts || !normals || !sdf_out) { g_err = "null argument"; return DSDF_EINVAL; }
  if (n <= 0 || n_pts <= 0 || n_pts > (1ll << 31) - 2) { g_err = "bad sizes"; return DSDF_EINVAL; }
  k_knn_sdf<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(xyz, n, (const float4*)pts, (const float4*)normals,
                                                                         (int)n_pts, stdv, sdf_out, nn_out);
  MC_CU(cudaGetLastError());
  return DSDF_OK;
}
// ---- the same query through a uniform grid (exact; bit-identical results).  Workspace: header | counts / cursors |
// cell starts | block sums | cell-sorted points
namespace {
struct KnnWs { KnnGridHdr* hdr; int* cnt; int* start; int* block_sum; float4* sorted; size_t total; };
KnnWs knn_carve(uint8_t* base, long long n_pts) {
  KnnWs w;
  size_t o = 0;
  auto take = [&](size_t bytes) { uint8_t* p = base ? base + o : nullptr; o += (bytes + 255) & ~(size_t)255; return p; };
  w.hdr = (KnnGridHdr*)take(sizeof(KnnGridHdr));
  w.cnt = (int*)take((size_t)(kKnnCells + 1) * 4);
  w.start = (int*)take((size_t)(kKnnCells + 1) * 4);
  w.block_sum = (int*)take((size_t)kKnnScanBlocks * 4);
  w.sorted = (float4*)take((size_t)n_pts * 16);
  w.total = o;
  return w;
}
}  // namespace
extern "C" size_t dsdf_knn_grid_workspace_bytes(int64_t n_pts) {
  if (n_pts <= 0) return 0;
  return knn_carve(nullptr, n_pts).total + 256;
}
extern "C" int dsdf_knn_sdf_grid(const float* xyz, int64_t n, const float* pts, const float* normals, int64_t n_pts,
                                 float stdv, void* workspace, size_t workspace_bytes, float* sdf_out, int32_t* nn_out,
                                 void* stream) {
  if (!xyz || !pts || !normals || !sdf_out || !workspace) { g_err = "null argument"; return DSDF_EINVAL; }
  if (n <= 0 || n_pts <= 0 || n_pts > (1ll << 31) - 2) { g_err = "bad sizes"; return DSDF_EINVAL; }
  uint8_t* base = (uint8_t*)(((uintptr_t)workspace + 255) & ~(uintptr_t)255);
  KnnWs w = knn_carve(base, n_pts);
  if ((size_t)(base - (uint8_t*)workspace) + w.total > workspace_bytes) {
    g_err = "workspace too small: need " + std::to_string(w.total + 256) + " bytes, have " + std::to_string(workspace_bytes);
    return DSDF_ENOMEM;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int np = (int)n_pts;
  const float4* p4 = (const float4*)pts;
  MC_CU(cudaMemsetAsync(w.cnt, 0, (size_t)(kKnnCells + 1) * 4, st));
  k_knn_grid_init<<<1, 32, 0, st>>>(w.hdr);
  int bb = (np + 255) / 256;
  if (bb > 1184) bb = 1184;
  k_knn_grid_bbox<<<bb, 256, 0, st>>>(p4, np, w.hdr);
  k_knn_grid_setup<<<1, 32, 0, st>>>(w.hdr);
  k_knn_grid_count<<<(np + 255) / 256, 256, 0, st>>>(p4, np, w.hdr, w.cnt);
  k_knn_scan_local<<<kKnnScanBlocks, kKnnScanBlock, 0, st>>>(w.cnt, w.start, w.block_sum);
  k_knn_scan_blocks<<<1, kKnnScanBlock, 0, st>>>(w.block_sum);
  k_knn_scan_add<<<kKnnScanBlocks, kKnnScanBlock, 0, st>>>(w.start, w.block_sum, w.cnt);   // counts become the cursors
  k_knn_grid_fill<<<(np + 255) / 256, 256, 0, st>>>(p4, np, w.hdr, w.start, w.cnt, w.sorted);
  k_knn_sdf_grid<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(xyz, n, w.sorted, w.start, w.hdr, p4, (const float4*)normals,
                                                             stdv, sdf_out, nn_out);
  MC_CU(cudaGetLastError());
  return DSDF_OK;
}
extern "C" int dsdf_gather_windows(const float* samples, const int64_t* base, const int64_t* win, int64_t n_scenes,
                                   int64_t half, float* out, void* stream) {
  if (!samples || !base || !win || !out) { g_err = "null argument"; return DSDF_EINVAL; }
  if (n_scenes <= 0 || half <= 0 || n_scenes > 65535) { g_err = "bad sizes"; return DSDF_EINVAL; }
  k_gather_windows<<<dim3((unsigned)((2 * half + 255) / 256), (unsigned)n_scenes), 256, 0, (cudaStream_t)stream>>>(
      (const float4*)samples, (const long long*)base, (const long long*)win, half, (float4*)out);
  MC_CU(cudaGetLastError());
  return DSDF_OK;
}

#ifdef DSDF_DIAG_BUILD
extern "C" int dsdf_selftest_gemm(int32_t kind, const void* a_img, const void* b_img, float* d_out,
                                  void* stream) {
  if (!a_img || !b_img || !d_out || kind < 0 || kind > 1) {
    g_err = "bad argument";
    return DSDF_EINVAL;
  }
  cudaError_t e = cudaFuncSetAttribute(k_selftest_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  k_selftest_gemm<<<1, 128, 65536, (cudaStream_t)stream>>>(kind, (const uint8_t*)a_img,
                                                           (const uint8_t*)b_img, d_out);
  e = cudaGetLastError();
  if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  return DSDF_OK;
}

extern "C" int dsdf_selftest_mma_rate(int32_t kind, int32_t iters, int32_t n_cta, uint64_t* cycles_host) {
  if (!cycles_host || iters <= 0 || n_cta <= 0 || n_cta > 1024) { g_err = "bad argument"; return DSDF_EINVAL; }
  const int smem = 196608 + 64;
  cudaError_t e = cudaFuncSetAttribute(k_mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  unsigned long long* d = nullptr;
  if ((e = cudaMalloc(&d, (size_t)n_cta * 8)) != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  if (kind & 64) {
    cudaFuncSetAttribute(k_mma_rate2, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3((unsigned)((n_cta + 1) / 2 * 2));
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, k_mma_rate2, (int)(kind & 63), (int)iters, d);
  } else {
    k_mma_rate<<<n_cta, 128, smem>>>(kind, iters, d);
  }
  e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(cycles_host, d, (size_t)n_cta * 8, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) { g_err = cudaGetErrorString(e); return DSDF_ECUDA; }
  return DSDF_OK;
}

#else
// product build: the self-test / micro-benchmark kernels are not compiled in (diagnostics variant only)
extern "C" int dsdf_selftest_gemm(int32_t, const void*, const void*, float*, void*) {
  g_err = "diagnostics entry: build the variant with -DDSDF_DIAG_BUILD (build.py DSDF_BUILD_VARIANT=diag)";
  return DSDF_EUNSUPPORTED;
}
extern "C" int dsdf_selftest_mma_rate(int32_t, int32_t, int32_t, uint64_t*) {
  g_err = "diagnostics entry: build the variant with -DDSDF_DIAG_BUILD (build.py DSDF_BUILD_VARIANT=diag)";
  return DSDF_EUNSUPPORTED;
}
#endif

// diagnostics builds (-DDSDF_HANG_TRAP): mapped host buffer the device writes deadlock reports to
extern "C" int dsdf_debug_progress_buffer(uint32_t** host_ptr) {
#ifdef DSDF_PROGRESS
  static uint32_t* buf = nullptr;
  if (!buf) {
    if (cudaHostAlloc((void**)&buf, 16 * 64 * 4, cudaHostAllocMapped) != cudaSuccess) return DSDF_ECUDA;
    memset(buf, 0, 16 * 64 * 4);
    unsigned int* dptr = nullptr;
    if (cudaHostGetDevicePointer((void**)&dptr, buf, 0) != cudaSuccess) return DSDF_ECUDA;
    volatile unsigned int* v = dptr;
    if (cudaMemcpyToSymbol(g_prog, &v, sizeof(v)) != cudaSuccess) return DSDF_ECUDA;
  }
  if (host_ptr) *host_ptr = buf;
  return DSDF_OK;
#else
  (void)host_ptr;
  g_err = "library was not built with -DDSDF_PROGRESS";
  return DSDF_EUNSUPPORTED;
#endif
}

extern "C" int dsdf_debug_hang_buffer(uint32_t** host_ptr) {
#ifdef DSDF_HANG_TRAP
  static uint32_t* buf = nullptr;
  if (!buf) {
    if (cudaHostAlloc((void**)&buf, 4096, cudaHostAllocMapped) != cudaSuccess) return DSDF_ECUDA;
    memset(buf, 0, 4096);
    unsigned int* dptr = nullptr;
    if (cudaHostGetDevicePointer((void**)&dptr, buf, 0) != cudaSuccess) return DSDF_ECUDA;
    volatile unsigned int* v = dptr;
    if (cudaMemcpyToSymbol(g_hang, &v, sizeof(v)) != cudaSuccess) return DSDF_ECUDA;
  }
  if (host_ptr) *host_ptr = buf;
  return DSDF_OK;
#else
  (void)host_ptr;
  g_err = "library was not built with -DDSDF_HANG_TRAP";
  return DSDF_EUNSUPPORTED;
#endif
}
