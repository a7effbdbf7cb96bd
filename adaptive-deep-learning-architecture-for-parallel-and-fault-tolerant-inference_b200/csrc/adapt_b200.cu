// C-ABI implementation: stage executor (op program -> kernel launches -> CUDA graph),
// single-layer entry point, and hand-off plumbing.  See include/adapt_b200.h.
#include "../../include/adapt_b200.h"

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <atomic>
#include <chrono>
#include <thread>
#include <time.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "conv_gemm_sm100.cuh"
#include "conv_chain_sm100.cuh"
#include "stem_pool_sm100.cuh"
#include "simt_kernels.cuh"

namespace {

thread_local std::string g_last_error;

// Configuration-time CUDA work -- allocation, kernel attribute set-up, weight upload, stream capture and graph
// instantiation -- is serialised per process.  Several stage workers can live in one process (one thread per GPU,
// or several stages sharing a GPU); measured on B200: a stage's capture was occasionally invalidated ("operation
// failed due to a previous error during capture") while a sibling thread was creating its stage (cudaMalloc,
// cudaFuncSetAttribute, synchronous copies) even in thread-local capture mode.  Steady-state calls (enqueue, flag
// writes, async copies) never take this lock.
std::recursive_mutex g_config_mu;
#define CONFIG_LOCK() std::lock_guard<std::recursive_mutex> _config_lock(g_config_mu)

int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                                     \
    do {                                                                                                   \
        cudaError_t _e = (expr);                                                                           \
        if (_e != cudaSuccess)                                                                             \
            return fail(_e == cudaErrorMemoryAllocation ? ADAPT_ENOMEM : ADAPT_ECUDA, "%s failed: %s (%s:%d)", \
                        #expr, cudaGetErrorString(_e), __FILE__, __LINE__);                                \
    } while (0)

// ---------------------------------------------------------------- driver entry points
// libcuda is resolved at run time through the runtime so that the library still dlopen()s
// (and exports its symbols) on a machine without a driver.
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*PFN_encodeIm2col)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                     const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                     const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
typedef CUresult (*PFN_streamWriteValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);
typedef CUresult (*PFN_streamWaitValue32)(CUstream, CUdeviceptr, cuuint32_t, unsigned int);

struct DriverApi {
    PFN_encodeTiled encodeTiled = nullptr;
    PFN_encodeIm2col encodeIm2col = nullptr;
    PFN_streamWriteValue32 writeValue32 = nullptr;
    PFN_streamWaitValue32 waitValue32 = nullptr;
    int driver_version = 0;
    bool ok = false;
};

int get_driver(DriverApi** out) {
    static DriverApi api;
    static bool tried = false;
    if (!tried) {
        tried = true;
        cudaDriverEntryPointQueryResult q;
        void* fn = nullptr;
        bool good = true;
        auto load = [&](const char* name, void** dst) {
            fn = nullptr;
            cudaError_t e = cudaGetDriverEntryPoint(name, &fn, cudaEnableDefault, &q);
            if (e != cudaSuccess || fn == nullptr || q != cudaDriverEntryPointSuccess) {
                good = false;
                cudaGetLastError();
            } else {
                *dst = fn;
            }
        };
        load("cuTensorMapEncodeTiled", reinterpret_cast<void**>(&api.encodeTiled));
        load("cuTensorMapEncodeIm2col", reinterpret_cast<void**>(&api.encodeIm2col));
        load("cuStreamWriteValue32", reinterpret_cast<void**>(&api.writeValue32));
        load("cuStreamWaitValue32", reinterpret_cast<void**>(&api.waitValue32));
        cudaDriverGetVersion(&api.driver_version);
        api.ok = good;
    }
    if (!api.ok) return fail(ADAPT_ENODEV, "CUDA driver entry points unavailable (no GPU / no driver)");
    *out = &api;
    return ADAPT_OK;
}

int require_device(int device) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        return fail(ADAPT_ENODEV, "no CUDA device visible (this library has no CPU fallback)");
    }
    if (device < 0 || device >= n) return fail(ADAPT_EINVAL, "device %d out of range (count %d)", device, n);
    CUDA_TRY(cudaSetDevice(device));
    return ADAPT_OK;
}

// ---------------------------------------------------------------- tensor maps
int encode_tiled_2d_bf16(CUtensorMap* m, const void* base, uint64_t inner, uint64_t rows, uint32_t box_inner,
                         uint32_t box_rows) {
    DriverApi* d;
    int rc = get_driver(&d);
    if (rc) return rc;
    cuuint64_t dims[2] = {inner, rows};
    cuuint64_t strides[1] = {inner * 2};
    cuuint32_t box[2] = {box_inner, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = d->encodeTiled(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box,
                                estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(ADAPT_ECUDA, "cuTensorMapEncodeTiled failed (%d) inner=%llu rows=%llu box=%ux%u base=%p", (int)r,
                    (unsigned long long)inner, (unsigned long long)rows, box_inner, box_rows, base);
    return ADAPT_OK;
}

// Output of the stem-by-phase kernel as a 5-D tensor (channel, column phase, unit column, row, image): pixel
// q = 4 * unit + phase.  One store = 32 unit columns of one phase of one row; whatever lies beyond Wo / Ho is clipped.
int encode_phase_out_bf16(CUtensorMap* m, const void* base, int C, int Wo, int Ho, int N) {
    DriverApi* d;
    int rc = get_driver(&d);
    if (rc) return rc;
    cuuint64_t dims[5] = {(cuuint64_t)C, 4, (cuuint64_t)(Wo / 4), (cuuint64_t)Ho, (cuuint64_t)N};
    cuuint64_t strides[4] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * 4, (cuuint64_t)Wo * C * 2, (cuuint64_t)Ho * Wo * C * 2};
    cuuint32_t box[5] = {(cuuint32_t)C, 1, 32, 1, 1};
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    CUtensorMap tmp;
    CUresult r = d->encodeTiled(&tmp, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides, box, estr,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return 1;  // the caller keeps the st.global write-out (and its descriptor untouched)
    *m = tmp;
    return ADAPT_OK;
}

// NHWC activation tensor described as (C, W, H, N) for im2col-mode loads of a
// kh x kw / stride / symmetric-pad convolution; one load = 128 output pixels x 64 channels.
int encode_im2col_bf16(CUtensorMap* m, const void* base, int C, int W, int H, int N, int kh, int kw, int stride,
                       int pad) {
    DriverApi* d;
    int rc = get_driver(&d);
    if (rc) return rc;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    int lower[2] = {-pad, -pad};
    int upper[2] = {pad - (kw - 1), pad - (kh - 1)};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    CUresult r = d->encodeIm2col(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides,
                                 lower, upper, /*channelsPerPixel=*/adapt::kBlockK, /*pixelsPerColumn=*/adapt::kBlockM,
                                 estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(ADAPT_ECUDA, "cuTensorMapEncodeIm2col failed (%d) C=%d W=%d H=%d N=%d k=%dx%d s=%d p=%d", (int)r, C,
                    W, H, N, kh, kw, stride, pad);
    // Known driver issue (also worked around by CUTLASS): for tensors smaller than 128 KiB
    // older drivers set a descriptor bit that makes im2col loads fault; clear it.
    if (d->driver_version <= 13010) {
        const uint64_t bytes = (uint64_t)C * W * H * N * 2;
        if (bytes < 131072) reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
    }
    return ADAPT_OK;
}

// Row-reuse view of a 3x3 / stride 1 / pad 1 input (A_ROW3): base positions run over W + 2 columns
// (-1 .. W) and H rows (-1 .. H-2); the only filter offset used is the vertical one.  One load brings
// kRow3Rows consecutive positions of that padded-flat walk.
int encode_row3_bf16(CUtensorMap* m, const void* base, int C, int W, int H, int N) {
    DriverApi* d;
    int rc = get_driver(&d);
    if (rc) return rc;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    int lower[2] = {-1, -1};
    int upper[2] = {+1, -1};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = d->encodeIm2col(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper,
                                 /*channelsPerPixel=*/adapt::kBlockK, /*pixelsPerColumn=*/adapt::kRow3Rows, estr,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                 CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return -1;  // the caller falls back to the per-tap im2col path
    if (d->driver_version <= 13010) {
        const uint64_t bytes = (uint64_t)C * W * H * N * 2;
        if (bytes < 131072) reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
    }
    return ADAPT_OK;
}

// True only when `ptr` is plain device memory allocated on `device` itself (not a peer GPU's allocation, not a
// CUDA-IPC mapping of one, not host or managed memory).  Used to decide whether a caller-provided weight tensor
// may be the source of a cluster-multicast TMA load.
bool pointer_is_local_device_memory(const void* ptr, int device) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, ptr) != cudaSuccess) {
        (void)cudaGetLastError();
        return false;
    }
    return a.type == cudaMemoryTypeDevice && a.device == device;
}

// Residual / output positions of the chain kernel: the NHWC tensor (C, W, H, N) walked in padded-flat order -- rows
// of W + 2 positions starting at pixel 0, the last two of each row outside the tensor (zero-filled on load).  One
// load = 32 consecutive positions x 64 channels, the epilogue's slab.
int encode_padded_flat_slab_bf16(CUtensorMap* m, const void* base, int C, int W, int H, int N) {
    DriverApi* d;
    int rc = get_driver(&d);
    if (rc) return rc;
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    int lower[2] = {0, 0};
    int upper[2] = {+2, 0};
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = d->encodeIm2col(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, lower, upper,
                                 /*channelsPerPixel=*/adapt::kEpiCols, /*pixelsPerColumn=*/32, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                 CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(ADAPT_ECUDA, "cuTensorMapEncodeIm2col (padded-flat slab) failed (%d) C=%d W=%d H=%d N=%d", (int)r, C, W, H, N);
    if (d->driver_version <= 13010) {
        const uint64_t bytes = (uint64_t)C * W * H * N * 2;
        if (bytes < 131072) reinterpret_cast<uint64_t*>(m)[1] &= ~(1ull << 21);
    }
    return ADAPT_OK;
}

// ---------------------------------------------------------------- conv launch records
struct ConvGeom {
    int batch, H, W, Cin, Ho, Wo, Cout, kh, kw, stride, pad;
};

struct ConvLaunch {
    CUtensorMap tmap_a;
    CUtensorMap tmap_b;
    CUtensorMap tmap_out;  // [M, N] bf16 output, 64-column x 32-row boxes (staged epilogue)
    CUtensorMap tmap_res;  // [M, N] bf16 residual, same boxes
    int smem = 0;          // dynamic shared memory of this launch (operand stages + epilogue staging)
    int cluster = 1;       // thread-block cluster size (B operand multicast) or 1
    bool pair = false;     // cluster of 2 running as one CTA pair (tcgen05 cta_group::2, 256-row tiles)
    adapt::ConvGemmParams p;
    int block_n = 0;
    int a_mode = 0;
    int grid = 0;
};

// Launch record of the chained `_2_conv` (3x3) -> `_3_conv` (1x1 + shortcut + ReLU) kernel (conv_chain_sm100.cuh).
struct ChainLaunch {
    CUtensorMap tmap_a, tmap_w2, tmap_w3, tmap_res;
    adapt::ConvChainParams p;
    int C = 0;
    int smem = 0;
    int grid = 0;
};

// Launch record of the fused stem + max-pool kernel (stem_pool_sm100.cuh).
struct StemPoolLaunch {
    CUtensorMap tmap_a, tmap_b;
    adapt::StemPoolParams p;
    int smem = 0;
    int grid = 0;
};

int g_force_a_mode = -2;  // -2 = read env lazily
int g_force_block_n = -2;

int env_int(const char* name, int dflt) {
    const char* v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

// Tile width for the tensor-bound class.  `pair` = the layer will run on CTA pairs (256-row tiles, one
// unit per pair of SMs): the wave count is then taken over pairs.
int pick_block_n(int m_tiles, int N, int num_sms, bool pair) {
    const int cands[3] = {256, 128, 64};
    const float penalty[3] = {1.0f, 1.1f, 1.5f};  // narrower tiles re-read A from smem more often
    int best = 64;
    float best_cost = 1e30f;
    for (int i = 0; i < 3; ++i) {
        const int bn = cands[i];
        if (bn > 64 && N < bn) continue;
        const int n_tiles = (N + bn - 1) / bn;
        // do not waste more than ~12% of the MMA columns on padding
        if ((float)(n_tiles * bn - N) > 0.125f * (float)N && bn > 64) continue;
        const bool as_pair = pair && bn >= 128 && N % bn == 0;
        const int units = (as_pair ? (m_tiles + 1) / 2 : m_tiles) * n_tiles;
        const int slots = as_pair ? num_sms / 2 : num_sms;
        const int waves = (units + slots - 1) / slots;
        // a pair halves the B bytes each SM ingests, which is what paces these layers (profiles/r1)
        const float cost = (float)waves * (float)bn * penalty[i] * (as_pair ? 0.85f : 1.0f);
        if (cost < best_cost) {
            best_cost = cost;
            best = bn;
        }
    }
    return best;
}

// Number of clusters of `cluster` CTAs (one CTA per SM: each takes the whole shared memory) that can be
// resident at once; cached per cluster size.  Returns 0 if the query fails (the caller keeps its bound).
int max_resident_clusters(int cluster);

template <int BN, int MODE>
int set_smem_attr() {
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_gemm_kernel<BN, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  adapt::kMaxSmemBytes));
    return ADAPT_OK;
}

int init_conv_kernels_for_device() {
    int rc;
#define SET(BN, MODE)                        \
    rc = set_smem_attr<BN, adapt::MODE>();   \
    if (rc) return rc;
    SET(64, A_TILED) SET(128, A_TILED) SET(256, A_TILED)
    SET(64, A_IM2COL) SET(128, A_IM2COL) SET(256, A_IM2COL)
    SET(64, A_GATHER) SET(128, A_GATHER) SET(256, A_GATHER)
#undef SET
#define SETC(BN, MODE, CL)                                                                                       \
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_gemm_kernel<BN, adapt::MODE, CL>,                                  \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, adapt::kMaxSmemBytes));
    SETC(128, A_TILED, 2) SETC(256, A_TILED, 2) SETC(128, A_IM2COL, 2) SETC(256, A_IM2COL, 2)
    SETC(128, A_TILED, 4) SETC(256, A_TILED, 4) SETC(128, A_IM2COL, 4) SETC(256, A_IM2COL, 4)
#undef SETC
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_gemm_kernel<256, adapt::A_S2DP, 1, false>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, adapt::kMaxSmemBytes));
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_gemm_kernel<64, adapt::A_ROW3, 1, false>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, adapt::kMaxSmemBytes));
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_gemm_kernel<128, adapt::A_ROW3, 2, true>,
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, adapt::kMaxSmemBytes));
#define SETP(BN, MODE)                                                                                           \
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_gemm_kernel<BN, adapt::MODE, 2, true>,                             \
                                  cudaFuncAttributeMaxDynamicSharedMemorySize, adapt::kMaxSmemBytes));
    SETP(128, A_TILED) SETP(256, A_TILED) SETP(128, A_IM2COL) SETP(256, A_IM2COL)
#undef SETP
    CUDA_TRY(cudaFuncSetAttribute(adapt::stem_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, adapt::kMaxSmemBytes));
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_chain_kernel<64, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  adapt::kMaxSmemBytes));
    CUDA_TRY(cudaFuncSetAttribute(adapt::conv_chain_kernel<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  adapt::kMaxSmemBytes));
    return ADAPT_OK;
}

// Per-device one-time set-up of the conv kernels' attributes, for the entry points that have no stage handle;
// callers on different threads may race to be first, so the bookkeeping is locked.
int init_conv_kernels_once(int device) {
    static std::mutex mu;
    static bool done[64] = {false};
    if (device < 0 || device >= 64) return init_conv_kernels_for_device();
    std::lock_guard<std::mutex> lock(mu);
    if (done[device]) return ADAPT_OK;
    const int rc = init_conv_kernels_for_device();
    if (rc == ADAPT_OK) done[device] = true;
    return rc;
}

// Build the launch record (tensor maps + params) for one convolution.
// Space-to-depth stem by column phase (A_S2DP, conv_gemm_sm100.cuh): 4x4 / stride 1 / VALID over 16-channel pixels
// stored in rows of whole 128-byte units, 64 output channels.  Returns 1 if the geometry is not this case.
int build_s2dp_launch(ConvLaunch* L, const ConvGeom& g, const void* in, const void* w_bf16, const float* scale,
                      const float* bias, void* out, int relu, int num_sms) {
    if (!(g.kh == 4 && g.kw == 4 && g.Cin == 16 && g.Cout == 64 && g.stride == 1 && g.pad == 0 && g.W % 4 == 0 &&
          g.Ho <= g.H - 3 && g.Wo <= g.W - 3))
        return 1;
    const int units_per_row = g.W / 4;
    const long long M_walk = (long long)g.batch * g.H * units_per_row;  // one GEMM row per unit of every image row
    if (M_walk <= 0 || M_walk > 0x7fffffffLL) return 1;
    adapt::ConvGemmParams& p = L->p;
    memset(&p, 0, sizeof(p));
    p.M = (int)M_walk;
    p.N = g.Cout;
    p.num_k_blocks = 4;  // one per filter row: 4 taps x 16 channels
    p.m_tiles = (int)((M_walk + adapt::kBlockM - 1) / adapt::kBlockM);
    p.n_tiles = 1;
    p.H = g.H; p.W = g.W; p.Cin = g.Cin; p.Ho = g.Ho; p.Wo = g.Wo;
    p.kh = 4; p.kw = 4; p.stride = 1; p.pad = 0;
    p.c_blocks = 1;
    p.K_real = 256;
    p.scale = scale;
    p.bias = bias;
    p.out_bf16 = reinterpret_cast<__nv_bfloat16*>(out);
    p.relu = relu;
    p.in = reinterpret_cast<const __nv_bfloat16*>(in);
    p.row_pitch = units_per_row;
    p.row_valid = (g.Wo + 3) / 4;
    p.epi_tma = 2;  // staged epilogue, st.global write-out maps (unit, phase) to pixels
    p.b_resident = 1;
    p.d_out = 1;
    p.d_res = 0;
    L->block_n = 256;  // four 64-column accumulators
    L->a_mode = adapt::A_S2DP;
    L->cluster = 1;
    L->pair = false;
    L->grid = p.m_tiles < num_sms ? p.m_tiles : num_sms;
    int rc = encode_tiled_2d_bf16(&L->tmap_b, w_bf16, 256, (uint64_t)g.Cout, adapt::kBlockK, 64);
    if (rc) return rc;
    rc = encode_tiled_2d_bf16(&L->tmap_a, in, adapt::kBlockK, (uint64_t)M_walk, adapt::kBlockK, adapt::kRow3Rows);
    if (rc) return rc;
    L->tmap_out = L->tmap_b;
    // rows of exactly 32 units (what the ingest writes for 224-pixel images): a warp's 32-row slab is one image row
    // of one phase, so it can leave through one 5-D TMA store instead
    if (units_per_row == 32 && g.Wo % 4 == 0 && g.Cout == 64 && env_int("ADAPT_S2DP_TMA_STORE", 1) != 0 &&
        encode_phase_out_bf16(&L->tmap_out, out, g.Cout, g.Wo, g.Ho, g.batch) == ADAPT_OK) {
        p.epi_tma = 1;
        p.d_out = 2;
    }
    L->tmap_res = L->tmap_b;
    const int fixed = 4 * 64 * adapt::kBlockK * 2 + adapt::conv_gemm_epi_bytes(256, p.d_out, 0) + adapt::kBarrierBytes + 1024;
    int stages = (adapt::kMaxSmemBytes - fixed) / adapt::kRow3ABytes;
    if (stages > adapt::kMaxStages) stages = adapt::kMaxStages;
    if (stages < 2) return 1;
    p.stages = stages;
    L->smem = fixed + stages * adapt::kRow3ABytes;
    return ADAPT_OK;
}

// `w_local`: the weight tensor is known to be an allocation of the GPU the kernel runs on.  The optional
// cluster variant multicasts the B operand (ADAPT_CLUSTER=2/4); TMA multicast loads here read only this GPU's
// memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.  Without that knowledge
// the launch never multicasts.
int build_conv_launch(ConvLaunch* L, const ConvGeom& g, const void* in, const void* w_bf16, const float* scale,
                      const float* bias, const void* residual, void* out, bool out_f32, int relu, int a_mode_req,
                      int block_n_req, int num_sms, bool w_local) {
    if (g.Cin % 8 != 0) return fail(ADAPT_EINVAL, "conv: stored Cin=%d must be a multiple of 8", g.Cin);
    if (g.Cout % 8 != 0) return fail(ADAPT_EINVAL, "conv: Cout=%d must be a multiple of 8", g.Cout);
    if (g_force_a_mode == -2) g_force_a_mode = env_int("ADAPT_FORCE_A_MODE", -1);
    if (g_force_block_n == -2) g_force_block_n = env_int("ADAPT_FORCE_BLOCK_N", 0);
    const int K = g.kh * g.kw * g.Cin;
    const long long M = (long long)g.batch * g.Ho * g.Wo;
    if (M <= 0 || M > 0x7fffffffLL) return fail(ADAPT_EINVAL, "conv: bad M=%lld", M);

    if ((a_mode_req == adapt::A_S2DP || (a_mode_req < 0 && g_force_a_mode < 0 && env_int("ADAPT_S2DP", 1) != 0)) &&
        residual == nullptr && !out_f32) {
        const int rc = build_s2dp_launch(L, g, in, w_bf16, scale, bias, out, relu, num_sms);
        if (rc <= 0) return rc;  // built (0) or failed (< 0); 1 = not the stem geometry, fall through
    }
    int a_mode = a_mode_req >= 0 ? a_mode_req : g_force_a_mode;
    if (a_mode == adapt::A_S2DP) a_mode = -1;
    const bool plain = g.kh == 1 && g.kw == 1 && g.stride == 1 && g.pad == 0;
    if (a_mode < 0) {
        if (g.Cin % 64 != 0)
            a_mode = adapt::A_GATHER;
        else if (plain)
            a_mode = adapt::A_TILED;
        else
            a_mode = adapt::A_IM2COL;
    }
    if (a_mode == adapt::A_TILED && !plain) a_mode = adapt::A_IM2COL;
    if (a_mode != adapt::A_GATHER && g.Cin % 64 != 0) a_mode = adapt::A_GATHER;
    // Row reuse for 3x3 / stride 1 / pad 1 (ADAPT_ROW3: 0 off, 1 rows of at least 28 pixels, 2 any width): the
    // padded-flat walk spends (W + 2) / W of the MMA work and adds tiles; measured at batch 128 it wins on the
    // 56- and 28-wide maps (-27 %, -12 %), and loses a wave on the 14-wide ones (224 instead of 196 units).
    const int row3_env = env_int("ADAPT_ROW3", 1);
    const bool row3_shape = g.kh == 3 && g.kw == 3 && g.stride == 1 && g.pad == 1 && g.Cin % 64 == 0 && g.Cout % 64 == 0 &&
                            residual == nullptr && !out_f32 && env_int("ADAPT_EPI_MODE", 1) != 0;
    bool row3 = row3_shape && ((a_mode_req == adapt::A_ROW3) ||
                               (a_mode_req < 0 && g_force_a_mode < 0 && row3_env != 0 && (row3_env == 2 || g.W >= 28)));
    if (a_mode == adapt::A_ROW3 && !row3) a_mode = adapt::A_IM2COL;
    long long M_walk = M;  // rows of the GEMM the kernel walks (padded-flat positions in row-reuse mode)
    if (row3) {
        M_walk = (long long)g.batch * g.H * (g.W + 2);
        if (M_walk > 0x7fffffffLL) row3 = false;
    }
    if (row3 && encode_row3_bf16(&L->tmap_a, in, g.Cin, g.W, g.H, g.batch) != ADAPT_OK) row3 = false;
    if (row3) a_mode = adapt::A_ROW3;
    if (!row3) M_walk = M;

    adapt::ConvGemmParams& p = L->p;
    memset(&p, 0, sizeof(p));
    p.M = (int)M_walk;
    p.N = g.Cout;
    p.num_k_blocks = (K + adapt::kBlockK - 1) / adapt::kBlockK;
    p.m_tiles = (int)((M_walk + adapt::kBlockM - 1) / adapt::kBlockM);
    if (row3) {
        p.row_pitch = g.W + 2;
        p.row_valid = g.W;
    }
    // Roofline class of the layer: FLOP per byte of unavoidable HBM traffic (input read once, output
    // written once, residual read once, weights once).  Below the machine balance (~210-250 FLOP/B) the
    // layer is HBM-bound and the launch is configured for bytes in flight, not for MMA efficiency.
    const double hbm_bytes = 2.0 * ((double)g.batch * g.H * g.W * g.Cin / (g.kh == 1 ? g.stride * g.stride : 1) +
                                    (double)M * g.Cout * (residual ? 2 : 1) + (double)g.Cout * K);
    const bool memory_bound = (2.0 * (double)M * g.Cout * K) / hbm_bytes < 300.0;
    int bn = block_n_req > 0 ? block_n_req : (g_force_block_n > 0 ? g_force_block_n : 0);
    // 0: never, 1 (default): tensor-bound layers run on CTA pairs, 2: every eligible layer (tests)
    const int pair_env = env_int("ADAPT_PAIR", 1);
    // (memory-bound class: pairs only pay where the operand ring is the shallow part -- residual layers with
    // K >= 256, whose 24 KiB pair stages fit 5 deep next to the residual ring instead of 3: -5 % measured)
    const bool pair_ok = a_mode != adapt::A_GATHER && !out_f32 && env_int("ADAPT_EPI_MODE", 1) != 0 && p.m_tiles >= 2 &&
                         (pair_env == 2 || (pair_env == 1 && (!memory_bound || (residual != nullptr && K >= 256))));
    if (bn == 0) bn = memory_bound ? (g.Cout >= 128 ? 128 : 64) : pick_block_n(p.m_tiles, g.Cout, num_sms, pair_ok);
    // a layer of a handful of tiles (the Dense layer: 1 x 8; the deep layers at micro-batch 1) is paced by how fast ONE
    // SM ingests its operand stream: narrower tiles put twice as many SMs on it
    if (block_n_req <= 0 && g_force_block_n <= 0 && bn == 128 && memory_bound &&
        (long long)p.m_tiles * ((g.Cout + 127) / 128) < num_sms / 4)
        bn = 64;
    if (row3) {
        // a row-reuse stage holds three B tiles: 128-wide tiles on CTA pairs (8 KiB per tile and CTA), or
        // 64-wide tiles on single CTAs
        if (bn > 128) bn = 128;
        if (bn == 128 && (g.Cout % 128 != 0 || p.m_tiles < 2)) bn = 64;
    }
    if (bn != 64 && bn != 128 && bn != 256) return fail(ADAPT_EINVAL, "conv: block_n=%d unsupported", bn);
    p.n_tiles = (g.Cout + bn - 1) / bn;
    p.H = g.H; p.W = g.W; p.Cin = g.Cin; p.Ho = g.Ho; p.Wo = g.Wo;
    p.kh = g.kh; p.kw = g.kw; p.stride = g.stride; p.pad = g.pad;
    p.c_blocks = g.Cin / adapt::kBlockK;
    p.K_real = K;
    p.scale = scale;
    p.bias = bias;
    p.residual = reinterpret_cast<const __nv_bfloat16*>(residual);
    p.out_bf16 = out_f32 ? nullptr : reinterpret_cast<__nv_bfloat16*>(out);
    p.out_f32 = out_f32 ? reinterpret_cast<float*>(out) : nullptr;
    p.relu = relu;
    p.in = reinterpret_cast<const __nv_bfloat16*>(in);
    L->block_n = bn;
    L->a_mode = a_mode;
    // Thread-block clusters with a multicast B operand for the layers that are bound by L2 -> SM traffic
    // (tensor-bound class, TMA A path, bf16 output whose width the tile divides).
    int cluster = 1;
    if (w_local && !memory_bound && a_mode != adapt::A_GATHER && bn >= 128 && !out_f32 && g.Cout % bn == 0 &&
        env_int("ADAPT_EPI_MODE", 1) != 0) {
        // Measured on B200 (profiles/): multicasting B does not speed these layers up -- they are bound by
        // shared-memory bandwidth (TMA fills + MMA operand reads), not by L2 -> SM traffic -- so it is off by
        // default and kept as a tested switch.
        cluster = env_int("ADAPT_CLUSTER", 1);
        if (cluster != 2 && cluster != 4) cluster = 1;
        while (cluster > 1 && p.m_tiles < 2 * cluster) cluster /= 2;
    }
    // CTA pairs (cta_group::2): same eligibility; the pair replaces the multicast cluster
    bool pair = false;
    if (pair_ok && bn >= 128 && g.Cout % bn == 0) {
        pair = true;
        cluster = 2;
    }
    if (row3) {
        pair = bn == 128;
        cluster = pair ? 2 : 1;
    }
    L->pair = pair;
    L->cluster = cluster;
    const long long units = (long long)((p.m_tiles + cluster - 1) / cluster) * p.n_tiles;  // per cluster
    long long max_clusters = num_sms / cluster;
    if (cluster > 1) {
        // a persistent grid must be co-resident: ask how many clusters of this size the device can hold
        // (GPCs with an odd number of usable SMs leave some SMs without a partner)
        const int resident = max_resident_clusters(cluster);
        if (resident > 0 && resident < max_clusters) max_clusters = resident;
    }
    L->grid = (int)((units < max_clusters ? units : max_clusters) * cluster);

    int rc = encode_tiled_2d_bf16(&L->tmap_b, w_bf16, (uint64_t)K, (uint64_t)g.Cout, adapt::kBlockK,
                                  (uint32_t)(bn / cluster));  // each CTA of a cluster loads (and multicasts) 1/cluster of B
    if (rc) return rc;
    if (a_mode == adapt::A_TILED) {
        rc = encode_tiled_2d_bf16(&L->tmap_a, in, (uint64_t)g.Cin, (uint64_t)M, adapt::kBlockK, adapt::kBlockM);
        if (rc) return rc;
    } else if (a_mode == adapt::A_IM2COL) {
        rc = encode_im2col_bf16(&L->tmap_a, in, g.Cin, g.W, g.H, g.batch, g.kh, g.kw, g.stride, g.pad);
        if (rc) return rc;
    } else if (a_mode == adapt::A_ROW3) {
        // (tmap_a was encoded when the mode was chosen)
    } else {
        L->tmap_a = L->tmap_b;  // unused by the gather kernel
        // "Wide pixel" view for thin-channel stride-1 VALID convolutions whose filter row is exactly one
        // 64-element k-block (the space-to-depth stem: 4 taps x 16 channels): kw horizontally adjacent
        // pixels are contiguous in NHWC, so the tensor is described to the TMA as (C'=kw*Cin, W'=Wo, H, N)
        // with a W stride of ONE pixel (overlapping windows) and the layer runs on the im2col path as a
        // kh x 1 convolution over 64 channels.  Falls back to the software gather if the driver refuses.
        static int wide_ok = env_int("ADAPT_WIDE_PIXEL", 1);
        if (wide_ok && a_mode_req < 0 && g.Cin * g.kw == adapt::kBlockK && g.stride == 1 && g.pad == 0) {
            DriverApi* d;
            rc = get_driver(&d);
            if (rc) return rc;
            cuuint64_t dims[4] = {(cuuint64_t)adapt::kBlockK, (cuuint64_t)g.Wo, (cuuint64_t)g.H, (cuuint64_t)g.batch};
            cuuint64_t strides[3] = {(cuuint64_t)g.Cin * 2, (cuuint64_t)g.W * g.Cin * 2, (cuuint64_t)g.H * g.W * g.Cin * 2};
            int lower[2] = {0, 0};
            int upper[2] = {0, -(g.kh - 1)};
            cuuint32_t estr[4] = {1, 1, 1, 1};
            CUresult r = d->encodeIm2col(&L->tmap_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(in), dims, strides,
                                         lower, upper, adapt::kBlockK, adapt::kBlockM, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                         CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r == CUDA_SUCCESS) {
                a_mode = adapt::A_IM2COL;
                L->a_mode = a_mode;
                p.kw = 1;
                p.c_blocks = 1;
            } else {
                L->tmap_a = L->tmap_b;
            }
        }
    }
    // staged epilogue: bf16 output whose width the tile divides; the direct epilogue remains for fp32
    // logits (Dense) and as a switchable fall-back (ADAPT_EPI_DIRECT=1)
    const int epi_mode = env_int("ADAPT_EPI_MODE", 1);  // 0 direct, 1 staged + TMA store, 2 staged + coalesced st.global
    L->tmap_out = L->tmap_b;
    L->tmap_res = L->tmap_b;
    p.epi_tma = (!out_f32 && g.Cout % bn == 0) ? epi_mode : 0;
    if (row3) p.epi_tma = 2;  // the write-out maps padded-flat positions to pixels row by row
    if (p.epi_tma) {
        rc = encode_tiled_2d_bf16(&L->tmap_out, out, (uint64_t)g.Cout, (uint64_t)M, adapt::kEpiCols, 32);
        if (rc) return rc;
        if (residual) {
            rc = encode_tiled_2d_bf16(&L->tmap_res, residual, (uint64_t)g.Cout, (uint64_t)M, adapt::kEpiCols, 32);
            if (rc) return rc;
        }
    }
    // Split the 227 KiB of shared memory between operand stages and epilogue staging.  Each of the 8 epilogue
    // warps owns a ring of 4 KiB slabs: `d_out` deep without a residual; with one, `d_res` deep and used in
    // place (residual in, sum out), so d_out is 0.  d_res >= 3 keeps at least one residual chunk in flight
    // while another is being worked on.
    const bool has_res = residual != nullptr && p.epi_tma != 0;
    int stages = adapt::kMaxStages;  // as many operand stages as fit beside the epilogue rings (loop below)
    int d_out = has_res ? 0 : 2;
    int d_res = has_res ? (memory_bound ? 4 : 3) : 0;
    if (!p.epi_tma) d_out = 2;  // slabs unused by the direct epilogue; keep the layout minimal
    stages = env_int("ADAPT_STAGES", stages);
    if (!has_res) {
        d_out = env_int("ADAPT_DOUT", d_out);
        d_out = d_out >= 4 ? 4 : (d_out >= 2 ? 2 : 1);
    } else {
        d_res = env_int("ADAPT_DRES", d_res);
        if (d_res > adapt::kMaxRes) d_res = adapt::kMaxRes;
        if (d_res < 2) d_res = 2;
    }
    if (stages > adapt::kMaxStages) stages = adapt::kMaxStages;
    if (stages < 2) stages = 2;
    // Resident B: one n-tile, several m-tiles per CTA and a weight matrix of at most 96 KiB -> keep it in shared
    // memory for the whole launch instead of re-fetching it with every tile (stem, conv2 3x3, 1x1 -> 64).
    int b_res_blocks = 0;
    if (env_int("ADAPT_B_RESIDENT", 1) && p.n_tiles == 1 && L->cluster == 1 && a_mode != adapt::A_GATHER &&
        p.num_k_blocks * bn * adapt::kBlockK * 2 <= 96 * 1024 && p.m_tiles > L->grid) {
        b_res_blocks = p.num_k_blocks;
        stages = adapt::kMaxStages;  // the ring carries 16 KiB A tiles only: take what fits
    }
    p.b_resident = b_res_blocks > 0 ? 1 : 0;
    if (row3) {
        stages = adapt::kMaxStages;  // a stage is three k-blocks deep: take what fits
        d_out = 1;                   // the st.global write-out reads the slab synchronously: one slab per warp
    }
    while (adapt::conv_gemm_smem_bytes(bn, stages, d_out, d_res, L->pair, b_res_blocks, row3) > adapt::kMaxSmemBytes) {
        if (stages > 4) --stages;
        else if (d_res > 3) --d_res;
        else if (stages > 3) --stages;
        else if (d_out > 2) d_out = 2;
        else if (stages > 2) --stages;
        else if (d_res > 2) --d_res;
        else if (d_out > 1) d_out = 1;
        else return fail(ADAPT_EINVAL, "conv: no shared-memory configuration fits (bn=%d)", bn);
    }
    p.stages = stages;
    p.d_out = d_out;
    p.d_res = d_res;
    L->smem = adapt::conv_gemm_smem_bytes(bn, stages, d_out, d_res, L->pair, b_res_blocks, row3);
    return ADAPT_OK;
}

int max_resident_clusters(int cluster) {
    static int cached[5] = {-1, -1, -1, -1, -1};
    if (cluster < 2 || cluster > 4) return 0;
    if (cached[cluster] >= 0) return cached[cluster];
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(148 * 4);
    cfg.blockDim = dim3((unsigned)adapt::conv_gemm_threads<adapt::A_TILED>());
    cfg.dynamicSmemBytes = (size_t)adapt::kMaxSmemBytes;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)cluster;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    cudaError_t e = cluster == 2
                        ? cudaOccupancyMaxActiveClusters(&n, adapt::conv_gemm_kernel<128, adapt::A_TILED, 2, true>, &cfg)
                        : cudaOccupancyMaxActiveClusters(&n, adapt::conv_gemm_kernel<128, adapt::A_TILED, 4>, &cfg);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        n = 0;
    }
    if (env_int("ADAPT_DEBUG", 0)) fprintf(stderr, "[adapt] resident clusters of %d: %d\n", cluster, n);
    cached[cluster] = n;
    return n;
}

// Every conv_gemm launch goes through cudaLaunchKernelEx: cluster dimension for the multicast / CTA-pair
// variants, and (unless ADAPT_PDL=0) programmatic stream serialization, so that a layer's CTAs may start
// their prologue -- barrier init, TMEM allocation, descriptor and weight prefetch -- on SMs the previous
// layer has already left; the kernel then blocks in griddepcontrol.wait before touching activations.
bool pdl_enabled() {
    static int v = env_int("ADAPT_PDL", 1);
    return v != 0;
}

// A SIMT kernel of the stage program launched like the convolution kernels: with the programmatic-serialization
// attribute, so that its launch overlaps the tail of the kernel before it (the kernel waits with griddepcontrol).
template <typename... KArgs, typename... Args>
cudaError_t launch_simt_pdl(void (*kernel)(KArgs...), unsigned grid, unsigned block, cudaStream_t s, Args... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    unsigned n = 0;
    static const int simt_pdl = env_int("ADAPT_PDL_SIMT", 1);
    if (pdl_enabled() && simt_pdl) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

template <int BN, int MODE, int CL, bool PAIR>
cudaError_t launch_conv_ex(const ConvLaunch& L, cudaStream_t s) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)L.grid);
    cfg.blockDim = dim3((unsigned)adapt::conv_gemm_threads<MODE>());
    cfg.dynamicSmemBytes = (size_t)L.smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    unsigned n = 0;
    if (CL > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = CL;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, adapt::conv_gemm_kernel<BN, MODE, CL, PAIR>, L.tmap_a, L.tmap_b, L.tmap_out, L.tmap_res,
                              L.p);
}

template <int BN>
cudaError_t launch_conv_bn(const ConvLaunch& L, cudaStream_t s) {
    if (L.a_mode == adapt::A_GATHER) return launch_conv_ex<BN, adapt::A_GATHER, 1, false>(L, s);
    if (L.a_mode == adapt::A_S2DP) {
        if constexpr (BN == 256) return launch_conv_ex<256, adapt::A_S2DP, 1, false>(L, s);
        return cudaErrorInvalidConfiguration;
    }
    if (L.a_mode == adapt::A_ROW3) {
        if constexpr (BN == 128) {
            if (L.pair) return launch_conv_ex<128, adapt::A_ROW3, 2, true>(L, s);
        }
        if constexpr (BN == 64) return launch_conv_ex<64, adapt::A_ROW3, 1, false>(L, s);
        return cudaErrorInvalidConfiguration;
    }
    const bool tiled = L.a_mode == adapt::A_TILED;
    if constexpr (BN >= 128) {
        if (L.pair)
            return tiled ? launch_conv_ex<BN, adapt::A_TILED, 2, true>(L, s) : launch_conv_ex<BN, adapt::A_IM2COL, 2, true>(L, s);
        if (L.cluster == 2)
            return tiled ? launch_conv_ex<BN, adapt::A_TILED, 2, false>(L, s) : launch_conv_ex<BN, adapt::A_IM2COL, 2, false>(L, s);
        if (L.cluster == 4)
            return tiled ? launch_conv_ex<BN, adapt::A_TILED, 4, false>(L, s) : launch_conv_ex<BN, adapt::A_IM2COL, 4, false>(L, s);
    }
    return tiled ? launch_conv_ex<BN, adapt::A_TILED, 1, false>(L, s) : launch_conv_ex<BN, adapt::A_IM2COL, 1, false>(L, s);
}

int run_conv_launch(const ConvLaunch& L, cudaStream_t s) {
    static const int debug = env_int("ADAPT_DEBUG", 0);
    if (debug)
        fprintf(stderr, "[adapt] conv M=%d N=%d kblocks=%d a_mode=%d bn=%d pair=%d cluster=%d grid=%d stages=%d smem=%d b_res=%d epi=%d\n",
                L.p.M, L.p.N, L.p.num_k_blocks, L.a_mode, L.block_n, (int)L.pair, L.cluster, L.grid, L.p.stages, L.smem,
                L.p.b_resident, L.p.epi_tma);
    if (L.cluster > 1 && L.block_n < 128) return fail(ADAPT_EINVAL, "conv: clusters need block_n >= 128");
    if (L.block_n == 64)
        CUDA_TRY(launch_conv_bn<64>(L, s));
    else if (L.block_n == 128)
        CUDA_TRY(launch_conv_bn<128>(L, s));
    else
        CUDA_TRY(launch_conv_bn<256>(L, s));
    return ADAPT_OK;
}

// Chain eligibility: 3x3 / stride 1 / pad 1 with C in {64, 128} on a map the row-reuse walk pays for (W >= 28), followed by
// a 1x1 to 4C channels with a shortcut.
// Stem (space-to-depth form: 4x4 / stride 1 / VALID over 16-channel pixels stored in rows of 32 units, 64 output
// channels, +BN +ReLU) followed by the 3x3 / stride 2 / zero-pad 1 max-pool, in one kernel.
bool stem_pool_shape_ok(int H, int W, int Cin, int Ho, int Wo, int Cout) {
    return Cin == 16 && Cout == 64 && W == 128 && Ho % 4 == 0 && Wo % 4 == 0 && Wo / 4 <= 32 && Ho >= 4 && Ho <= H - 3 && Wo <= W - 3;
}

int build_stem_pool_launch(StemPoolLaunch* L, int batch, int H, int W, int Ho, int Wo, const void* in, const void* w_bf16,
                           const float* scale, const float* bias, void* out, int num_sms) {
    if (!stem_pool_shape_ok(H, W, 16, Ho, Wo, 64)) return fail(ADAPT_EINVAL, "stem+pool: unsupported shape H=%d W=%d Ho=%d Wo=%d", H, W, Ho, Wo);
    const long long units = (long long)batch * H * (W / 4);
    if (batch <= 0 || units > 0x7fffffffLL) return fail(ADAPT_EINVAL, "stem+pool: bad batch %d", batch);
    adapt::StemPoolParams& p = L->p;
    memset(&p, 0, sizeof(p));
    p.batch = batch;
    p.H = H;
    p.Ho = Ho;
    p.Wo = Wo;
    p.tiles_per_image = Ho / 4;
    p.total_tiles = batch * p.tiles_per_image;
    p.scale = scale;
    p.bias = bias;
    p.out = reinterpret_cast<__nv_bfloat16*>(out);
    int stages = env_int("ADAPT_STEM_POOL_STAGES", adapt::kStemPoolMaxStages);
    if (stages > adapt::kStemPoolMaxStages) stages = adapt::kStemPoolMaxStages;
    while (stages > 2 && adapt::stem_pool_smem_bytes(stages) > adapt::kMaxSmemBytes) --stages;
    if (stages < 2) stages = 2;
    p.stages = stages;
    L->smem = adapt::stem_pool_smem_bytes(stages);
    L->grid = p.total_tiles < num_sms ? p.total_tiles : num_sms;
    int rc = encode_tiled_2d_bf16(&L->tmap_b, w_bf16, 256, 64, adapt::kBlockK, 64);
    if (rc) return rc;
    return encode_tiled_2d_bf16(&L->tmap_a, in, adapt::kBlockK, (uint64_t)units, adapt::kBlockK, adapt::kRow3Rows);
}

int run_stem_pool_launch(const StemPoolLaunch& L, cudaStream_t s) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)L.grid);
    cfg.blockDim = dim3((unsigned)adapt::kStemPoolThreads);
    cfg.dynamicSmemBytes = (size_t)L.smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    unsigned n = 0;
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    CUDA_TRY(cudaLaunchKernelEx(&cfg, adapt::stem_pool_kernel, L.tmap_a, L.tmap_b, L.p));
    return ADAPT_OK;
}

bool chain_shape_ok(int C, int H, int W) { return (C == 64 || C == 128) && W >= 28 && H >= 1; }

int build_chain_launch(ChainLaunch* L, int batch, int H, int W, int C, const void* in, const void* w2, const float* scale1,
                       const float* bias1, const void* w3, const float* scale2, const float* bias2, const void* residual,
                       void* out, int num_sms) {
    if (!chain_shape_ok(C, H, W)) return fail(ADAPT_EINVAL, "conv chain: unsupported shape C=%d H=%d W=%d", C, H, W);
    if (!residual) return fail(ADAPT_EINVAL, "conv chain: the 1x1 needs its shortcut tensor");
    const long long M_walk = (long long)batch * H * (W + 2);
    if (M_walk <= 0 || M_walk > 0x7fffffffLL) return fail(ADAPT_EINVAL, "conv chain: bad M=%lld", M_walk);
    const bool pair = C == 128;
    adapt::ConvChainParams& p = L->p;
    memset(&p, 0, sizeof(p));
    p.M = (int)M_walk;
    p.m_tiles = (int)((M_walk + adapt::kBlockM - 1) / adapt::kBlockM);
    p.H = H;
    p.W = W;
    p.row_pitch = W + 2;
    p.N2 = 4 * C;
    p.scale1 = scale1;
    p.bias1 = bias1;
    p.scale2 = scale2;
    p.bias2 = bias2;
    p.out = reinterpret_cast<__nv_bfloat16*>(out);
    p.ring = env_int("ADAPT_CHAIN_RING", 2);
    if (p.ring < 2) p.ring = 2;
    if (p.ring > adapt::kChainMaxRing) p.ring = adapt::kChainMaxRing;
    p.l2_prefetch = env_int("ADAPT_CHAIN_L2PF", 0);
    int stages = env_int("ADAPT_CHAIN_STAGES", adapt::kChainMaxStages);
    if (stages > adapt::kChainMaxStages) stages = adapt::kChainMaxStages;
    auto bytes = [&](int st) {
        return pair ? adapt::conv_chain_smem_bytes<128, true>(st, p.ring) : adapt::conv_chain_smem_bytes<64, false>(st, p.ring);
    };
    while (stages > 2 && bytes(stages) > adapt::kMaxSmemBytes) --stages;
    if (bytes(stages) > adapt::kMaxSmemBytes) return fail(ADAPT_EINVAL, "conv chain: no shared-memory configuration fits");
    p.stages = stages;
    L->smem = bytes(stages);
    L->C = C;
    const int cl = pair ? 2 : 1;
    const long long units = (p.m_tiles + cl - 1) / cl;
    long long max_clusters = num_sms / cl;
    if (pair) {
        const int resident = max_resident_clusters(2);
        if (resident > 0 && resident < max_clusters) max_clusters = resident;
    }
    L->grid = (int)((units < max_clusters ? units : max_clusters) * cl);
    int rc = encode_row3_bf16(&L->tmap_a, in, C, W, H, batch);
    if (rc) return fail(ADAPT_ECUDA, "conv chain: row-reuse tensor map refused (C=%d W=%d H=%d N=%d)", C, W, H, batch);
    rc = encode_tiled_2d_bf16(&L->tmap_w2, w2, (uint64_t)9 * C, (uint64_t)C, adapt::kBlockK, (uint32_t)(pair ? C / 2 : C));
    if (rc) return rc;
    rc = encode_tiled_2d_bf16(&L->tmap_w3, w3, (uint64_t)C, (uint64_t)4 * C, adapt::kBlockK,
                              (uint32_t)(pair ? adapt::kChainBN2 / 2 : adapt::kChainBN2));
    if (rc) return rc;
    return encode_padded_flat_slab_bf16(&L->tmap_res, residual, 4 * C, W, H, batch);
}

template <int C, bool PAIR>
cudaError_t launch_chain_ex(const ChainLaunch& L, cudaStream_t s) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3((unsigned)L.grid);
    cfg.blockDim = dim3((unsigned)adapt::kChainThreads);
    cfg.dynamicSmemBytes = (size_t)L.smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[2];
    unsigned n = 0;
    if (PAIR) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = 2;
        attr[n].val.clusterDim.y = 1;
        attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr;
    cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, adapt::conv_chain_kernel<C, PAIR>, L.tmap_a, L.tmap_w2, L.tmap_w3, L.tmap_res, L.p);
}

int run_chain_launch(const ChainLaunch& L, cudaStream_t s) {
    static const int debug = env_int("ADAPT_DEBUG", 0);
    if (debug)
        fprintf(stderr, "[adapt] chain C=%d M=%d m_tiles=%d grid=%d stages=%d ring=%d smem=%d\n", L.C, L.p.M, L.p.m_tiles, L.grid,
                L.p.stages, L.p.ring, L.smem);
    if (L.C == 64)
        CUDA_TRY((launch_chain_ex<64, false>(L, s)));
    else
        CUDA_TRY((launch_chain_ex<128, true>(L, s)));
    return ADAPT_OK;
}

// ---------------------------------------------------------------- stage
struct WeightDev {
    int kh = 0, kw = 0, cin = 0, cin_store = 0, cout = 0;
    void* kernel = nullptr;  // bf16 [cout][kh][kw][cin_store]  or  fp32 HWIO in check mode
    float* scale = nullptr;
    float* bias = nullptr;
};

struct Step {
    int kind = 0;
    ConvLaunch conv;  // bf16 CONV
    ChainLaunch chain;  // bf16 CONV_CHAIN
    StemPoolLaunch stem_pool;  // bf16 STEM_POOL
    // generic pointers / geometry for the SIMT kernels
    const void* in = nullptr;
    const void* res = nullptr;
    void* out = nullptr;
    adapt_op_desc op;
    int batch = 0;
};

struct Plan {
    int batch = 0;
    std::vector<Step> steps[2];
    bool built[2] = {false, false};
    cudaGraphExec_t exec[2] = {nullptr, nullptr};
};

}  // namespace

struct adapt_stage {
    int device = 0;
    int mode = 0;
    int max_batch = 0;
    int num_sms = 0;
    bool use_graph = true;
    unsigned warmed_kinds = 0;  // op kinds whose kernels this stage has already launched outside a capture
    std::vector<WeightDev> weights;
    std::vector<adapt_tensor_desc> tdesc;
    std::vector<void*> tptr;  // internal tensors: sub-ranges of `arena`
    // One block for all internal activation tensors of the installed program.  It only ever grows, so a
    // re-slice to a program that fits (every re-slice, when the stage was created with its largest program)
    // allocates nothing: cudaMalloc / cudaFree synchronise the device and, with peer access enabled, re-map the
    // block into every peer, which is what a fail-over must not wait for.
    void* arena = nullptr;
    size_t arena_bytes = 0;
    std::vector<adapt_op_desc> ops;
    void* io_in[2] = {nullptr, nullptr};
    void* io_out[2] = {nullptr, nullptr};
    std::map<int, std::unique_ptr<Plan>> plans;
};

namespace {

size_t elem_size(const adapt_stage* st, int dtype) {
    if (dtype == ADAPT_DT_F32) return 4;
    if (dtype == ADAPT_DT_U8) return 1;
    return st->mode == ADAPT_MODE_BF16 ? 2 : 4;
}

void drop_plans(adapt_stage* st) {
    for (auto& kv : st->plans)
        for (int s = 0; s < 2; ++s)
            if (kv.second->exec[s]) cudaGraphExecDestroy(kv.second->exec[s]);
    st->plans.clear();
}

void free_tensors(adapt_stage* st, bool release_arena) {
    st->tptr.clear();
    st->tdesc.clear();
    if (release_arena && st->arena) {
        cudaFree(st->arena);
        st->arena = nullptr;
        st->arena_bytes = 0;
    }
}

int install_program(adapt_stage* st, const adapt_tensor_desc* tensors, int n_tensors, const adapt_op_desc* ops,
                    int n_ops) {
    if (n_tensors <= 0 || n_ops <= 0 || !tensors || !ops) return fail(ADAPT_EINVAL, "empty stage program");
    // validate before touching state
    int n_in = 0, n_out = 0;
    for (int i = 0; i < n_tensors; ++i) {
        if (tensors[i].elems_per_image <= 0) return fail(ADAPT_EINVAL, "tensor %d: bad size", i);
        if (tensors[i].role == ADAPT_T_INPUT) ++n_in;
        if (tensors[i].role == ADAPT_T_OUTPUT) ++n_out;
    }
    if (n_in != 1 || n_out != 1) return fail(ADAPT_EINVAL, "program needs exactly one INPUT and one OUTPUT tensor");
    for (int i = 0; i < n_ops; ++i) {
        const adapt_op_desc& o = ops[i];
        auto bad_t = [&](int t, bool optional) { return !(optional && t == -1) && (t < 0 || t >= n_tensors); };
        if (bad_t(o.in_tensor, false) || bad_t(o.out_tensor, false) || bad_t(o.res_tensor, true))
            return fail(ADAPT_EINVAL, "op %d: tensor index out of range", i);
        if (o.kind == ADAPT_OP_CONV) {
            if (o.weight < 0 || o.weight >= (int)st->weights.size())
                return fail(ADAPT_EINVAL, "op %d: weight index %d out of range", i, o.weight);
            const WeightDev& w = st->weights[o.weight];
            const int cin_expect = st->mode == ADAPT_MODE_BF16 ? w.cin_store : w.cin;
            if (w.kh != o.kh || w.kw != o.kw || w.cout != o.Cout || cin_expect != o.Cin)
                return fail(ADAPT_EINVAL, "op %d: geometry %dx%dx%d->%d does not match weight %d (%dx%dx%d->%d)", i,
                            o.kh, o.kw, o.Cin, o.Cout, o.weight, w.kh, w.kw, cin_expect, w.cout);
        } else if (o.kind == ADAPT_OP_AFFINE) {
            if (o.weight >= (int)st->weights.size()) return fail(ADAPT_EINVAL, "op %d: weight index out of range", i);
            if (o.weight >= 0 && st->weights[o.weight].cout != o.Cin)
                return fail(ADAPT_EINVAL, "op %d: affine channels %d != weight %d", i, o.Cin, st->weights[o.weight].cout);
            if (o.Cin % 8 != 0) return fail(ADAPT_EINVAL, "op %d: affine needs C %% 8 == 0", i);
        } else if (o.kind == ADAPT_OP_CONV_CHAIN) {
            if (st->mode != ADAPT_MODE_BF16) return fail(ADAPT_EINVAL, "op %d: CONV_CHAIN exists in bf16 mode only", i);
            if (o.weight < 0 || o.weight >= (int)st->weights.size() || o.weight2 < 0 || o.weight2 >= (int)st->weights.size())
                return fail(ADAPT_EINVAL, "op %d: chain weight index out of range", i);
            const WeightDev& w2 = st->weights[o.weight];
            const WeightDev& w3 = st->weights[o.weight2];
            if (w2.kh != 3 || w2.kw != 3 || w2.cin_store != o.Cin || w2.cout != o.Cin || w3.kh != 1 || w3.kw != 1 ||
                w3.cin_store != o.Cin || w3.cout != o.Cout || o.Cout != 4 * o.Cin || o.res_tensor < 0 ||
                !chain_shape_ok(o.Cin, o.H, o.W) || o.Ho != o.H || o.Wo != o.W)
                return fail(ADAPT_EINVAL, "op %d: not a chainable 3x3 -> 1x1 block (C=%d H=%d W=%d Cout=%d)", i, o.Cin, o.H, o.W,
                            o.Cout);
        } else if (o.kind == ADAPT_OP_STEM_POOL) {
            if (st->mode != ADAPT_MODE_BF16) return fail(ADAPT_EINVAL, "op %d: STEM_POOL exists in bf16 mode only", i);
            if (o.weight < 0 || o.weight >= (int)st->weights.size()) return fail(ADAPT_EINVAL, "op %d: weight index out of range", i);
            const WeightDev& w = st->weights[o.weight];
            if (w.kh != 4 || w.kw != 4 || w.cin_store != 16 || w.cout != 64 || o.kh != 4 || o.kw != 4 || o.stride != 1 || o.pad != 0 ||
                !o.relu || !stem_pool_shape_ok(o.H, o.W, o.Cin, o.Ho, o.Wo, o.Cout))
                return fail(ADAPT_EINVAL, "op %d: not the space-to-depth stem followed by its 3x3/2 max-pool (H=%d W=%d Cin=%d Ho=%d Wo=%d)", i,
                            o.H, o.W, o.Cin, o.Ho, o.Wo);
        } else if (o.kind < ADAPT_OP_INGEST || o.kind > ADAPT_OP_STEM_POOL) {
            return fail(ADAPT_EINVAL, "op %d: unknown kind %d", i, o.kind);
        }
    }
    drop_plans(st);
    free_tensors(st, false);
    constexpr size_t kAlign = 1024;  // (TMA needs 16 B; whole swizzle atoms keep the boxes of different tensors apart)
    size_t need = 0;
    std::vector<size_t> offs(n_tensors, 0);
    for (int i = 0; i < n_tensors; ++i) {
        if (tensors[i].role != ADAPT_T_INTERNAL) continue;
        const size_t bytes = (size_t)tensors[i].elems_per_image * st->max_batch * elem_size(st, tensors[i].dtype);
        offs[i] = need;
        need += (bytes + kAlign - 1) / kAlign * kAlign;
    }
    if (need > st->arena_bytes) {
        free_tensors(st, true);
        CUDA_TRY(cudaMalloc(&st->arena, need));
        st->arena_bytes = need;
    }
    st->tdesc.assign(tensors, tensors + n_tensors);
    st->tptr.assign(n_tensors, nullptr);
    st->ops.assign(ops, ops + n_ops);
    for (int i = 0; i < n_tensors; ++i)
        if (tensors[i].role == ADAPT_T_INTERNAL) st->tptr[i] = static_cast<uint8_t*>(st->arena) + offs[i];
    return ADAPT_OK;
}

void* resolve(adapt_stage* st, int t, int slot) {
    if (t < 0) return nullptr;
    const int role = st->tdesc[t].role;
    if (role == ADAPT_T_INPUT) return st->io_in[slot];
    if (role == ADAPT_T_OUTPUT) return st->io_out[slot];
    return st->tptr[t];
}

int build_steps(adapt_stage* st, Plan* plan, int slot) {
    if (!st->io_in[slot] || !st->io_out[slot]) return fail(ADAPT_ESTATE, "slot %d has no bound I/O buffers", slot);
    std::vector<Step>& steps = plan->steps[slot];
    steps.clear();
    for (size_t i = 0; i < st->ops.size(); ++i) {
        const adapt_op_desc& o = st->ops[i];
        Step s;
        s.kind = o.kind;
        s.op = o;
        s.batch = plan->batch;
        s.in = resolve(st, o.in_tensor, slot);
        s.res = resolve(st, o.res_tensor, slot);
        s.out = resolve(st, o.out_tensor, slot);
        if (o.kind == ADAPT_OP_CONV && st->mode == ADAPT_MODE_BF16) {
            const WeightDev& w = st->weights[o.weight];
            ConvGeom g{plan->batch, o.H, o.W, o.Cin, o.Ho, o.Wo, o.Cout, o.kh, o.kw, o.stride, o.pad};
            // (w.kernel is this stage's own cudaMalloc on st->device: upload_weights)
            int rc = build_conv_launch(&s.conv, g, s.in, w.kernel, w.scale, w.bias, s.res, s.out, o.out_f32 != 0,
                                       o.relu, -1, 0, st->num_sms, /*w_local=*/true);
            if (rc) return rc;
        } else if (o.kind == ADAPT_OP_CONV_CHAIN) {
            const WeightDev& w2 = st->weights[o.weight];
            const WeightDev& w3 = st->weights[o.weight2];
            int rc = build_chain_launch(&s.chain, plan->batch, o.H, o.W, o.Cin, s.in, w2.kernel, w2.scale, w2.bias, w3.kernel,
                                        w3.scale, w3.bias, s.res, s.out, st->num_sms);
            if (rc) return rc;
        } else if (o.kind == ADAPT_OP_STEM_POOL) {
            const WeightDev& w = st->weights[o.weight];
            int rc = build_stem_pool_launch(&s.stem_pool, plan->batch, o.H, o.W, o.Ho, o.Wo, s.in, w.kernel, w.scale, w.bias, s.out,
                                            st->num_sms);
            if (rc) return rc;
        }
        steps.push_back(s);
    }
    plan->built[slot] = true;
    return ADAPT_OK;
}

int run_steps(adapt_stage* st, const std::vector<Step>& steps, cudaStream_t stream) {
    for (const Step& s : steps) {
        const adapt_op_desc& o = s.op;
        const int B = s.batch;
        switch (s.kind) {
            case ADAPT_OP_INGEST: {
                const long long pixels = (long long)B * o.H * o.W;
                const bool u8 = o.in_format == ADAPT_IN_U8_RGB_CAFFE;  // raw uint8 RGB, preprocessing fused
                if (o.in_format != ADAPT_IN_F32 && !u8) return fail(ADAPT_EINVAL, "ingest: unknown input format %d", o.in_format);
                if (u8 && o.Cin != 3) return fail(ADAPT_EINVAL, "ingest: uint8 input needs 3 channels");
                const int threads = 256;
                if (st->mode == ADAPT_MODE_BF16 && o.Cout == 16) {
                    // space-to-depth ingest for the stem: (H, W, 3) -> (Ho, Wo, 16) with `pad` folded in
                    // (Wo may be wider than the image needs: the extra pixels of each stored row are written as zeros)
                    if (o.Cin != 3 || o.Ho * 2 != o.H + 2 * o.pad || o.Wo * 2 < o.W + 2 * o.pad)
                        return fail(ADAPT_EINVAL, "ingest(s2d): expects 3 channels, Ho = (H + 2*pad) / 2, Wo >= (W + 2*pad) / 2");
                    const long long outs = (long long)B * o.Ho * o.Wo;
                    const unsigned blocks = (unsigned)((outs + threads - 1) / threads);
                    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(s.out);
                    if (u8)
                        CUDA_TRY(launch_simt_pdl(adapt::ingest_s2d_kernel<adapt::SrcU8Caffe>, blocks, threads, stream,
                                                 reinterpret_cast<const uint8_t*>(s.in), dst, B, o.H, o.W, o.Ho, o.Wo, o.pad));
                    else
                        CUDA_TRY(launch_simt_pdl(adapt::ingest_s2d_kernel<adapt::SrcF32>, blocks, threads, stream,
                                                 reinterpret_cast<const float*>(s.in), dst, B, o.H, o.W, o.Ho, o.Wo, o.pad));
                } else if (st->mode == ADAPT_MODE_BF16) {
                    if (o.Cout != 8 || o.Cin > 8) return fail(ADAPT_EINVAL, "ingest: expects Cin<=8 -> 8 channels");
                    const unsigned blocks = (unsigned)((pixels + threads - 1) / threads);
                    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(s.out);
                    if (u8)
                        adapt::ingest_pad8_kernel<adapt::SrcU8Caffe><<<blocks, threads, 0, stream>>>(
                            reinterpret_cast<const uint8_t*>(s.in), dst, pixels, o.Cin);
                    else
                        adapt::ingest_pad8_kernel<adapt::SrcF32><<<blocks, threads, 0, stream>>>(
                            reinterpret_cast<const float*>(s.in), dst, pixels, o.Cin);
                } else if (u8) {
                    adapt::ingest_u8_to_f32_kernel<<<(unsigned)((pixels + threads - 1) / threads), threads, 0, stream>>>(
                        reinterpret_cast<const uint8_t*>(s.in), reinterpret_cast<float*>(s.out), pixels);
                } else {
                    CUDA_TRY(cudaMemcpyAsync(s.out, s.in, (size_t)pixels * o.Cin * 4, cudaMemcpyDeviceToDevice,
                                             stream));
                }
                break;
            }
            case ADAPT_OP_CONV: {
                if (st->mode == ADAPT_MODE_BF16) {
                    int rc = run_conv_launch(s.conv, stream);
                    if (rc) return rc;
                } else {
                    const WeightDev& w = st->weights[o.weight];
                    const int M = B * o.Ho * o.Wo;
                    constexpr int PIX = 8;
                    dim3 block(o.Cout >= 128 ? 128 : 64);
                    dim3 grid((o.Cout + block.x - 1) / block.x, (M + PIX - 1) / PIX);
                    adapt::conv_fp32_direct_kernel<PIX><<<grid, block, 0, stream>>>(
                        reinterpret_cast<const float*>(s.in), reinterpret_cast<const float*>(w.kernel), w.scale,
                        w.bias, reinterpret_cast<const float*>(s.res), reinterpret_cast<float*>(s.out), M, o.H, o.W,
                        o.Cin, o.Ho, o.Wo, o.Cout, o.kh, o.kw, o.stride, o.pad, o.relu);
                }
                break;
            }
            case ADAPT_OP_CONV_CHAIN: {
                int rc = run_chain_launch(s.chain, stream);
                if (rc) return rc;
                break;
            }
            case ADAPT_OP_STEM_POOL: {
                int rc = run_stem_pool_launch(s.stem_pool, stream);
                if (rc) return rc;
                break;
            }
            case ADAPT_OP_MAXPOOL: {
                const long long total = (long long)B * o.Ho * o.Wo * (o.Cin / 8);
                const int threads = 256;
                const unsigned blocks = (unsigned)((total + threads - 1) / threads);
                if (st->mode == ADAPT_MODE_BF16 && o.kh == 3 && o.stride == 2 && o.pad == 1) {
                    // the ResNet stem's pool: 2 x 2 outputs per thread from a shared 5 x 5 window
                    const long long blocks2 = ((long long)B * ((o.Ho + 1) / 2) * ((o.Wo + 1) / 2) * (o.Cin / 8) + threads - 1) / threads;
                    adapt::maxpool3x3s2p1_bf16_kernel<<<(unsigned)blocks2, threads, 0, stream>>>(
                        reinterpret_cast<const __nv_bfloat16*>(s.in), reinterpret_cast<__nv_bfloat16*>(s.out), B, o.H, o.W, o.Cin,
                        o.Ho, o.Wo);
                } else if (st->mode == ADAPT_MODE_BF16)
                    adapt::maxpool_zeropad_bf16_kernel<<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const __nv_bfloat16*>(s.in), reinterpret_cast<__nv_bfloat16*>(s.out), B, o.H,
                        o.W, o.Cin, o.Ho, o.Wo, o.kh, o.stride, o.pad);
                else
                    adapt::maxpool_zeropad_kernel<float><<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const float*>(s.in), reinterpret_cast<float*>(s.out), B, o.H, o.W, o.Cin,
                        o.Ho, o.Wo, o.kh, o.stride, o.pad);
                break;
            }
            case ADAPT_OP_GAP: {
                const int total = B * (o.Cin / 8);
                const int threads = 128;
                const unsigned blocks = (unsigned)((total + threads - 1) / threads);
                if (st->mode == ADAPT_MODE_BF16)
                    CUDA_TRY(launch_simt_pdl(adapt::global_avgpool_bf16_kernel, (unsigned)((total + 63) / 64), 64u, stream,
                                             reinterpret_cast<const __nv_bfloat16*>(s.in), reinterpret_cast<__nv_bfloat16*>(s.out), B,
                                             o.H * o.W, o.Cin));
                else
                    adapt::global_avgpool_kernel<float, float><<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const float*>(s.in), reinterpret_cast<float*>(s.out), B, o.H * o.W, o.Cin);
                break;
            }
            case ADAPT_OP_SOFTMAX: {
                const int warps_per_block = 4;
                const unsigned blocks = (unsigned)((B + warps_per_block - 1) / warps_per_block);
                CUDA_TRY(launch_simt_pdl(adapt::softmax_rows_kernel, blocks, (unsigned)(warps_per_block * 32), stream,
                                         reinterpret_cast<const float*>(s.in), reinterpret_cast<float*>(s.out), B, o.Cin));
                break;
            }
            case ADAPT_OP_AFFINE: {
                const long long vecs = (long long)B * o.H * o.W * o.Cin / 8;
                const float* sc = o.weight >= 0 ? st->weights[o.weight].scale : nullptr;
                const float* bi = o.weight >= 0 ? st->weights[o.weight].bias : nullptr;
                const int threads = 256;
                const unsigned blocks = (unsigned)((vecs + threads - 1) / threads);
                if (st->mode == ADAPT_MODE_BF16)
                    adapt::affine_relu_kernel<__nv_bfloat16><<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const __nv_bfloat16*>(s.in), reinterpret_cast<__nv_bfloat16*>(s.out), sc, bi,
                        vecs, o.Cin, o.relu);
                else
                    adapt::affine_relu_kernel<float><<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const float*>(s.in), reinterpret_cast<float*>(s.out), sc, bi, vecs, o.Cin,
                        o.relu);
                break;
            }
            case ADAPT_OP_ZEROPAD: {
                const long long total = (long long)B * (o.H + 2 * o.pad) * (o.W + 2 * o.pad) * o.Cin;
                const int threads = 256;
                const unsigned blocks = (unsigned)((total + threads - 1) / threads);
                if (st->mode == ADAPT_MODE_BF16)
                    adapt::zeropad_kernel<__nv_bfloat16><<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const __nv_bfloat16*>(s.in), reinterpret_cast<__nv_bfloat16*>(s.out), B, o.H,
                        o.W, o.Cin, o.pad);
                else
                    adapt::zeropad_kernel<float><<<blocks, threads, 0, stream>>>(
                        reinterpret_cast<const float*>(s.in), reinterpret_cast<float*>(s.out), B, o.H, o.W, o.Cin,
                        o.pad);
                break;
            }
            default:
                return fail(ADAPT_EINVAL, "unknown op kind %d", s.kind);
        }
        CUDA_TRY(cudaGetLastError());
    }
    return ADAPT_OK;
}

int upload_weights(adapt_stage* st, const adapt_weight_desc* weights, int n_weights) {
    st->weights.resize(n_weights);
    for (int i = 0; i < n_weights; ++i) {
        const adapt_weight_desc& h = weights[i];
        const bool affine_only = h.kernel == nullptr && h.kh == 0 && h.kw == 0;
        if (!h.scale || !h.bias || h.cout <= 0 || (!affine_only && (!h.kernel || h.kh <= 0 || h.kw <= 0 || h.cin <= 0)))
            return fail(ADAPT_EINVAL, "weight %d: bad descriptor", i);
        WeightDev& w = st->weights[i];
        w.kh = h.kh; w.kw = h.kw; w.cin = h.cin; w.cout = h.cout;
        w.cin_store = (h.cin + 7) / 8 * 8;
        const size_t cout_pad = (size_t)(h.cout + 7) / 8 * 8;
        CUDA_TRY(cudaMalloc(&w.scale, cout_pad * sizeof(float)));
        CUDA_TRY(cudaMalloc(&w.bias, cout_pad * sizeof(float)));
        CUDA_TRY(cudaMemset(w.scale, 0, cout_pad * sizeof(float)));
        CUDA_TRY(cudaMemset(w.bias, 0, cout_pad * sizeof(float)));
        CUDA_TRY(cudaMemcpy(w.scale, h.scale, (size_t)h.cout * sizeof(float), cudaMemcpyHostToDevice));
        CUDA_TRY(cudaMemcpy(w.bias, h.bias, (size_t)h.cout * sizeof(float), cudaMemcpyHostToDevice));
        if (affine_only) continue;
        if (st->mode == ADAPT_MODE_BF16) {
            // HWIO fp32 -> [cout][kh][kw][cin_store] bf16 (round-to-nearest-even), zero channel padding
            const size_t K = (size_t)h.kh * h.kw * w.cin_store;
            std::vector<__nv_bfloat16> t((size_t)h.cout * K, __float2bfloat16_rn(0.0f));
            for (int r = 0; r < h.kh; ++r)
                for (int s = 0; s < h.kw; ++s)
                    for (int c = 0; c < h.cin; ++c) {
                        const float* src = h.kernel + ((size_t)(r * h.kw + s) * h.cin + c) * h.cout;
                        for (int o = 0; o < h.cout; ++o)
                            t[(size_t)o * K + (size_t)(r * h.kw + s) * w.cin_store + c] = __float2bfloat16_rn(src[o]);
                    }
            CUDA_TRY(cudaMalloc(&w.kernel, t.size() * sizeof(__nv_bfloat16)));
            CUDA_TRY(cudaMemcpy(w.kernel, t.data(), t.size() * sizeof(__nv_bfloat16), cudaMemcpyHostToDevice));
        } else {
            const size_t n = (size_t)h.kh * h.kw * h.cin * h.cout;
            CUDA_TRY(cudaMalloc(&w.kernel, n * sizeof(float)));
            CUDA_TRY(cudaMemcpy(w.kernel, h.kernel, n * sizeof(float), cudaMemcpyHostToDevice));
        }
    }
    return ADAPT_OK;
}

}  // namespace

// =============================================================================== C ABI
extern "C" {

const char* adapt_version(void) { return "adapt_b200 0.1 (sm_100a)"; }

const char* adapt_last_error(void) { return g_last_error.c_str(); }

int adapt_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int adapt_stage_create(int device, int dtype_mode, int max_batch, const adapt_weight_desc* weights, int n_weights,
                       const adapt_tensor_desc* tensors, int n_tensors, const adapt_op_desc* ops, int n_ops,
                       adapt_stage_t** out) {
    CONFIG_LOCK();
    if (!out) return fail(ADAPT_EINVAL, "out is null");
    *out = nullptr;
    if (dtype_mode != ADAPT_MODE_BF16 && dtype_mode != ADAPT_MODE_FP32_CHECK)
        return fail(ADAPT_EINVAL, "unknown dtype_mode %d", dtype_mode);
    if (max_batch <= 0 || !weights || n_weights <= 0) return fail(ADAPT_EINVAL, "bad max_batch / weights");
    int rc = require_device(device);
    if (rc) return rc;
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10) return fail(ADAPT_ENODEV, "device %d is sm_%d%d; this library is sm_100a only", device,
                                      prop.major, prop.minor);
    rc = init_conv_kernels_for_device();
    if (rc) return rc;
    std::unique_ptr<adapt_stage> st(new adapt_stage());
    st->device = device;
    st->mode = dtype_mode;
    st->max_batch = max_batch;
    st->num_sms = prop.multiProcessorCount;
    st->use_graph = env_int("ADAPT_NO_GRAPH", 0) == 0;
    rc = upload_weights(st.get(), weights, n_weights);
    if (rc == ADAPT_OK) rc = install_program(st.get(), tensors, n_tensors, ops, n_ops);
    if (rc) {
        adapt_stage_destroy(st.release());
        return rc;
    }
    *out = st.release();
    return ADAPT_OK;
}

int adapt_stage_reslice(adapt_stage_t* st, const adapt_tensor_desc* tensors, int n_tensors, const adapt_op_desc* ops,
                        int n_ops) {
    CONFIG_LOCK();
    if (!st) return fail(ADAPT_EINVAL, "stage is null");
    int rc = require_device(st->device);
    if (rc) return rc;
    CUDA_TRY(cudaDeviceSynchronize());
    st->io_in[0] = st->io_in[1] = st->io_out[0] = st->io_out[1] = nullptr;
    return install_program(st, tensors, n_tensors, ops, n_ops);
}

int adapt_stage_bind_io(adapt_stage_t* st, int slot, void* in_ptr, void* out_ptr) {
    if (!st || slot < 0 || slot > 1 || !in_ptr || !out_ptr) return fail(ADAPT_EINVAL, "bind_io: bad argument");
    if (st->io_in[slot] == in_ptr && st->io_out[slot] == out_ptr) return ADAPT_OK;
    st->io_in[slot] = in_ptr;
    st->io_out[slot] = out_ptr;
    // cached launch records / graphs of this slot hold the old pointers
    for (auto& kv : st->plans) {
        Plan* p = kv.second.get();
        p->built[slot] = false;
        if (p->exec[slot]) {
            cudaGraphExecDestroy(p->exec[slot]);
            p->exec[slot] = nullptr;
        }
    }
    return ADAPT_OK;
}

namespace {
int get_plan(adapt_stage_t* st, int slot, int batch, Plan** out) {
    if (!st || slot < 0 || slot > 1) return fail(ADAPT_EINVAL, "stage call: bad argument");
    if (batch <= 0 || batch > st->max_batch) return fail(ADAPT_EINVAL, "batch %d not in [1, %d]", batch, st->max_batch);
    if (st->ops.empty()) return fail(ADAPT_ESTATE, "no program installed");
    int rc = require_device(st->device);
    if (rc) return rc;
    auto it = st->plans.find(batch);
    if (it == st->plans.end()) {
        std::unique_ptr<Plan> p(new Plan());
        p->batch = batch;
        it = st->plans.emplace(batch, std::move(p)).first;
    }
    Plan* plan = it->second.get();
    if (!plan->built[slot]) {
        rc = build_steps(st, plan, slot);  // host-only work: tensor-map encodes
        if (rc) return rc;
    }
    *out = plan;
    return ADAPT_OK;
}
}  // namespace

int adapt_stage_prepare(adapt_stage_t* st, int slot, int batch) {
    CONFIG_LOCK();
    Plan* plan = nullptr;
    int rc = get_plan(st, slot, batch, &plan);
    if (rc) return rc;
    if (!st->use_graph || plan->exec[slot]) return ADAPT_OK;
    cudaStream_t cap;
    CUDA_TRY(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
    // A kernel's very first launch loads its module (lazy loading), which is not something to do inside a
    // capture: run the program once as plain launches if it contains a kind this stage has not launched yet.
    unsigned kinds = 0;
    for (const Step& s : plan->steps[slot]) kinds |= 1u << s.kind;
    if (kinds & ~st->warmed_kinds) {
        rc = run_steps(st, plan->steps[slot], cap);
        cudaError_t we = cudaStreamSynchronize(cap);
        if (rc == ADAPT_OK && we != cudaSuccess) rc = fail(ADAPT_ECUDA, "prepare: warm-up run: %s", cudaGetErrorString(we));
        if (rc) {
            cudaStreamDestroy(cap);
            return rc;
        }
        st->warmed_kinds |= kinds;
    }
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal);
    if (e != cudaSuccess) {
        cudaStreamDestroy(cap);
        return fail(ADAPT_ECUDA, "cudaStreamBeginCapture: %s", cudaGetErrorString(e));
    }
    rc = run_steps(st, plan->steps[slot], cap);
    e = cudaStreamEndCapture(cap, &graph);
    cudaStreamDestroy(cap);
    if (rc) {
        if (graph) cudaGraphDestroy(graph);
        return rc;
    }
    if (e != cudaSuccess) return fail(ADAPT_ECUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
    e = cudaGraphInstantiate(&plan->exec[slot], graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return fail(ADAPT_ECUDA, "cudaGraphInstantiate: %s", cudaGetErrorString(e));
    return ADAPT_OK;
}

int adapt_stage_enqueue(adapt_stage_t* st, int slot, int batch, void* stream_v) {
    Plan* plan = nullptr;
    int rc = get_plan(st, slot, batch, &plan);
    if (rc) return rc;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
    // a batch size that was prepared runs as one CUDA graph; any other runs as plain launches
    // (graph capture / instantiation never happens here: it can synchronise the device, which
    // must not occur while peer stages are blocked on this stage's flags)
    if (plan->exec[slot]) {
        CUDA_TRY(cudaGraphLaunch(plan->exec[slot], stream));
        return ADAPT_OK;
    }
    return run_steps(st, plan->steps[slot], stream);
}

int adapt_stage_read_tensor(adapt_stage_t* st, int tensor, void* host_dst, size_t bytes, void* stream_v) {
    if (!st || tensor < 0 || tensor >= (int)st->tdesc.size() || !host_dst)
        return fail(ADAPT_EINVAL, "read_tensor: bad argument");
    if (st->tdesc[tensor].role != ADAPT_T_INTERNAL) return fail(ADAPT_EINVAL, "read_tensor: tensor is not internal");
    const size_t cap = (size_t)st->tdesc[tensor].elems_per_image * st->max_batch * elem_size(st, st->tdesc[tensor].dtype);
    if (bytes > cap) return fail(ADAPT_EINVAL, "read_tensor: %zu bytes > capacity %zu", bytes, cap);
    int rc = require_device(st->device);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream_v)));
    CUDA_TRY(cudaMemcpy(host_dst, st->tptr[tensor], bytes, cudaMemcpyDeviceToHost));
    return ADAPT_OK;
}

int adapt_stage_profile(adapt_stage_t* st, int slot, int batch, void* stream_v, int reps, float* ms_per_op, int n_ops) {
    Plan* plan = nullptr;
    int rc = get_plan(st, slot, batch, &plan);
    if (rc) return rc;
    const std::vector<Step>& steps = plan->steps[slot];
    if (!ms_per_op || n_ops != (int)steps.size() || reps <= 0)
        return fail(ADAPT_EINVAL, "profile: need reps > 0 and n_ops == %d", (int)steps.size());
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
    const size_t n = steps.size();
    std::vector<cudaEvent_t> ev((n + 1) * (size_t)reps);
    for (auto& e : ev) CUDA_TRY(cudaEventCreate(&e));
    for (int r = 0; r < reps; ++r) {
        CUDA_TRY(cudaEventRecord(ev[(size_t)r * (n + 1)], stream));
        for (size_t i = 0; i < n; ++i) {
            std::vector<Step> one(1, steps[i]);
            rc = run_steps(st, one, stream);
            if (rc) return rc;
            CUDA_TRY(cudaEventRecord(ev[(size_t)r * (n + 1) + i + 1], stream));
        }
    }
    CUDA_TRY(cudaStreamSynchronize(stream));
    for (size_t i = 0; i < n; ++i) {
        float acc = 0.f;
        for (int r = 0; r < reps; ++r) {
            float ms = 0.f;
            CUDA_TRY(cudaEventElapsedTime(&ms, ev[(size_t)r * (n + 1) + i], ev[(size_t)r * (n + 1) + i + 1]));
            acc += ms;
        }
        ms_per_op[i] = acc / (float)reps;
    }
    for (auto& e : ev) cudaEventDestroy(e);
    return ADAPT_OK;
}

int adapt_stage_time_kind(adapt_stage_t* st, int slot, int batch, void* stream_v, int kind_mask, int reps, float* ms_per_rep,
                          int* n_launches) {
    CONFIG_LOCK();
    Plan* plan = nullptr;
    int rc = get_plan(st, slot, batch, &plan);
    if (rc) return rc;
    if (!ms_per_rep || reps <= 0) return fail(ADAPT_EINVAL, "time_kind: need reps > 0 and an output pointer");
    std::vector<Step> only;
    for (const Step& s : plan->steps[slot])
        if ((kind_mask >> s.kind) & 1) only.push_back(s);
    if (n_launches) *n_launches = (int)only.size();
    if (only.empty()) {
        *ms_per_rep = 0.f;
        return ADAPT_OK;
    }
    // the same launches (attributes included: programmatic dependent launch between consecutive convolutions)
    // captured into their own graph, so that the class is timed the way it runs inside the stage's graph
    cudaStream_t cap;
    CUDA_TRY(cudaStreamCreateWithFlags(&cap, cudaStreamNonBlocking));
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    cudaError_t e = cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal);
    if (e == cudaSuccess) {
        rc = run_steps(st, only, cap);
        e = cudaStreamEndCapture(cap, &graph);
    }
    cudaStreamDestroy(cap);
    if (rc == ADAPT_OK && e != cudaSuccess) rc = fail(ADAPT_ECUDA, "time_kind: capture failed: %s", cudaGetErrorString(e));
    if (rc == ADAPT_OK && (e = cudaGraphInstantiate(&exec, graph, 0)) != cudaSuccess)
        rc = fail(ADAPT_ECUDA, "time_kind: cudaGraphInstantiate: %s", cudaGetErrorString(e));
    if (graph) cudaGraphDestroy(graph);
    if (rc) return rc;
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    for (int r = 0; r < 3; ++r) CUDA_TRY(cudaGraphLaunch(exec, stream));  // warm-up
    CUDA_TRY(cudaEventRecord(e0, stream));
    for (int r = 0; r < reps; ++r) CUDA_TRY(cudaGraphLaunch(exec, stream));
    CUDA_TRY(cudaEventRecord(e1, stream));
    CUDA_TRY(cudaStreamSynchronize(stream));
    float ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    *ms_per_rep = ms / (float)reps;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaGraphExecDestroy(exec);
    return ADAPT_OK;
}

int adapt_stage_num_launches(adapt_stage_t* st, int batch) {
    if (!st) return fail(ADAPT_EINVAL, "stage is null");
    (void)batch;
    return (int)st->ops.size();
}

void adapt_stage_destroy(adapt_stage_t* st) {
    CONFIG_LOCK();
    if (!st) return;
    if (cudaSetDevice(st->device) == cudaSuccess) {
        cudaDeviceSynchronize();
        drop_plans(st);
        free_tensors(st, true);
        for (WeightDev& w : st->weights) {
            if (w.kernel) cudaFree(w.kernel);
            if (w.scale) cudaFree(w.scale);
            if (w.bias) cudaFree(w.bias);
        }
    }
    cudaGetLastError();
    delete st;
}

int adapt_conv2d_bf16(int device, const void* in, const void* w, const float* scale, const float* bias,
                      const void* residual, void* out, int out_is_f32, int batch, int H, int W, int Cin, int Cout,
                      int kh, int kw, int stride, int pad, int relu, int a_mode, int block_n, void* stream) {
    if (!in || !w || !scale || !bias || !out) return fail(ADAPT_EINVAL, "conv2d: null pointer");
    if (batch <= 0 || H <= 0 || W <= 0 || kh <= 0 || kw <= 0 || stride <= 0 || pad < 0)
        return fail(ADAPT_EINVAL, "conv2d: bad geometry");
    int rc = require_device(device);
    if (rc) return rc;
    rc = init_conv_kernels_once(device);
    if (rc) return rc;
    int num_sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device));
    const int Ho = (H + 2 * pad - kh) / stride + 1;
    const int Wo = (W + 2 * pad - kw) / stride + 1;
    if (Ho <= 0 || Wo <= 0) return fail(ADAPT_EINVAL, "conv2d: empty output");
    ConvGeom g{batch, H, W, Cin, Ho, Wo, Cout, kh, kw, stride, pad};
    ConvLaunch L;
    rc = build_conv_launch(&L, g, in, w, scale, bias, residual, out, out_is_f32 != 0, relu, a_mode, block_n, num_sms,
                           pointer_is_local_device_memory(w, device));
    if (rc) return rc;
    // Developer timeline (tools/kernel_check.py --group trace): with ADAPT_TRACE_FILE set, every call of this
    // entry point runs synchronously and appends the per-CTA, per-tile role timestamps to that file.
    static const char* trace_file = ADAPT_ENABLE_TRACE ? getenv("ADAPT_TRACE_FILE") : nullptr;
    if (trace_file == nullptr || trace_file[0] == 0) return run_conv_launch(L, reinterpret_cast<cudaStream_t>(stream));
    const size_t n_stamps = (size_t)L.grid * adapt::kTraceTiles * adapt::kTraceEvents;
    long long* dtrace = nullptr;
    CUDA_TRY(cudaMalloc(&dtrace, n_stamps * sizeof(long long)));
    CUDA_TRY(cudaMemset(dtrace, 0, n_stamps * sizeof(long long)));
    L.p.trace = dtrace;
    rc = run_conv_launch(L, reinterpret_cast<cudaStream_t>(stream));
    if (rc == ADAPT_OK) {
        std::vector<long long> host(n_stamps);
        cudaError_t e = cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream));
        if (e == cudaSuccess) e = cudaMemcpy(host.data(), dtrace, n_stamps * sizeof(long long), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            rc = fail(ADAPT_ECUDA, "trace read-back: %s", cudaGetErrorString(e));
        } else if (FILE* f = fopen(trace_file, "a")) {
            fprintf(f, "{\"grid\": %d, \"block_n\": %d, \"a_mode\": %d, \"pair\": %d, \"m_tiles\": %d, \"n_tiles\": %d, "
                       "\"k_blocks\": %d, \"stages\": %d, \"tiles\": %d, \"events\": %d, \"stamps\": [",
                    L.grid, L.block_n, L.a_mode, L.pair ? 1 : 0, L.p.m_tiles, L.p.n_tiles, L.p.num_k_blocks, L.p.stages,
                    adapt::kTraceTiles, adapt::kTraceEvents);
            for (size_t i = 0; i < n_stamps; ++i) fprintf(f, i ? ",%lld" : "%lld", host[i]);
            fprintf(f, "]}\n");
            fclose(f);
        }
    }
    cudaFree(dtrace);
    return rc;
}

int adapt_stem_pool_bf16(int device, const void* in, const void* w, const float* scale, const float* bias, void* out, int batch,
                         int H, int W, int Ho, int Wo, void* stream) {
    if (!in || !w || !scale || !bias || !out) return fail(ADAPT_EINVAL, "stem_pool: null pointer");
    int rc = require_device(device);
    if (rc) return rc;
    rc = init_conv_kernels_once(device);
    if (rc) return rc;
    int num_sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device));
    StemPoolLaunch L;
    rc = build_stem_pool_launch(&L, batch, H, W, Ho, Wo, in, w, scale, bias, out, num_sms);
    if (rc) return rc;
    return run_stem_pool_launch(L, reinterpret_cast<cudaStream_t>(stream));
}

int adapt_conv_chain_bf16(int device, const void* in, const void* w2, const float* scale1, const float* bias1,
                          const void* w3, const float* scale2, const float* bias2, const void* residual, void* out,
                          int batch, int H, int W, int C, void* stream) {
    if (!in || !w2 || !scale1 || !bias1 || !w3 || !scale2 || !bias2 || !residual || !out)
        return fail(ADAPT_EINVAL, "conv_chain: null pointer");
    if (batch <= 0 || H <= 0 || W <= 0) return fail(ADAPT_EINVAL, "conv_chain: bad geometry");
    int rc = require_device(device);
    if (rc) return rc;
    rc = init_conv_kernels_once(device);
    if (rc) return rc;
    int num_sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device));
    ChainLaunch L;
    rc = build_chain_launch(&L, batch, H, W, C, in, w2, scale1, bias1, w3, scale2, bias2, residual, out, num_sms);
    if (rc) return rc;
    // Developer timeline (tools/chain_trace.py): with ADAPT_TRACE_FILE set (trace build only) the call runs
    // synchronously and appends the per-CTA, per-unit role timestamps to that file.
    static const char* trace_file = ADAPT_ENABLE_TRACE ? getenv("ADAPT_TRACE_FILE") : nullptr;
    if (trace_file == nullptr || trace_file[0] == 0) return run_chain_launch(L, reinterpret_cast<cudaStream_t>(stream));
    const size_t n_stamps = (size_t)L.grid * adapt::kTraceTiles * adapt::kChainTraceEvents;
    long long* dtrace = nullptr;
    CUDA_TRY(cudaMalloc(&dtrace, n_stamps * sizeof(long long)));
    CUDA_TRY(cudaMemset(dtrace, 0, n_stamps * sizeof(long long)));
    L.p.trace = dtrace;
    rc = run_chain_launch(L, reinterpret_cast<cudaStream_t>(stream));
    if (rc == ADAPT_OK) {
        std::vector<long long> host(n_stamps);
        cudaError_t e = cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream));
        if (e == cudaSuccess) e = cudaMemcpy(host.data(), dtrace, n_stamps * sizeof(long long), cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) {
            rc = fail(ADAPT_ECUDA, "trace read-back: %s", cudaGetErrorString(e));
        } else if (FILE* f = fopen(trace_file, "a")) {
            fprintf(f, "{\"kernel\": \"chain\", \"C\": %d, \"grid\": %d, \"m_tiles\": %d, \"stages\": %d, \"ring\": %d, "
                       "\"tiles\": %d, \"events\": %d, \"stamps\": [",
                    C, L.grid, L.p.m_tiles, L.p.stages, L.p.ring, adapt::kTraceTiles, adapt::kChainTraceEvents);
            for (size_t i = 0; i < n_stamps; ++i) fprintf(f, i ? ",%lld" : "%lld", host[i]);
            fprintf(f, "]}\n");
            fclose(f);
        }
    }
    cudaFree(dtrace);
    return rc;
}

// ---------------------------------------------------------------- classifier head without allocations
}  // extern "C"  (the scratch array below needs C++ linkage-free file scope)

// Fixed scratch for adapt_dense_softmax_noalloc: the fp32 logits that sit between the Dense
// GEMM and the row softmax.  ADAPT_NOALLOC_MAX_LOGITS floats = 1 MiB, reserved on every GPU
// that loads this library for as long as the process lives.
static __device__ __align__(256) float g_noalloc_logits[ADAPT_NOALLOC_MAX_LOGITS];

namespace {
// Device address of g_noalloc_logits per device.  In host code the array's name is only a
// stand-in symbol, never an address a kernel could use, so the real address is looked up once
// per device (with that device current) and cached.
float* g_noalloc_logits_addr[64] = {nullptr};
std::mutex g_noalloc_mu;  // first calls on one device from several threads

int noalloc_scratch(int device, float** out) {
    if (device < 0 || device >= 64) return fail(ADAPT_EINVAL, "noalloc: device %d unsupported", device);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_logits_addr[device] == nullptr) {
        float* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_logits);
        if (e != cudaSuccess || ptr == nullptr)
            return fail(ADAPT_ECUDA, "cudaGetSymbolAddress(g_noalloc_logits) on device %d: %s", device,
                        cudaGetErrorString(e));
        g_noalloc_logits_addr[device] = ptr;
    }
    *out = g_noalloc_logits_addr[device];
    return ADAPT_OK;
}
}  // namespace

extern "C" {

int adapt_dense_softmax_noalloc(int device, const void* in, const void* w, const float* scale, const float* bias,
                                float* probs, int batch, int cin, int cout, void* stream_v) {
    if (!in || !w || !scale || !bias || !probs) return fail(ADAPT_EINVAL, "dense_softmax_noalloc: null pointer");
    if (batch <= 0 || cin <= 0 || cout <= 0 || cin % 8 != 0 || cout % 8 != 0)
        return fail(ADAPT_EINVAL, "dense_softmax_noalloc: need batch > 0 and cin, cout positive multiples of 8");
    if ((long long)batch * cout > (long long)ADAPT_NOALLOC_MAX_LOGITS)
        return fail(ADAPT_EINVAL,
                    "dense_softmax_noalloc: batch %d x cout %d = %lld logits exceeds the fixed scratch of %d floats "
                    "(ADAPT_NOALLOC_MAX_LOGITS); split the batch or use the stage API",
                    batch, cout, (long long)batch * cout, (int)ADAPT_NOALLOC_MAX_LOGITS);
    // the kernels read these with 16-byte vector loads / TMA: a caller carving them out of its own arena must keep
    // that alignment (a misaligned address is a sticky fault of the whole context, so it is refused here)
    if (((uintptr_t)in | (uintptr_t)w | (uintptr_t)scale | (uintptr_t)bias | (uintptr_t)probs) & 15)
        return fail(ADAPT_EINVAL, "dense_softmax_noalloc: in, w, scale, bias and probs must be 16-byte aligned");
    int rc = require_device(device);
    if (rc) return rc;
    rc = init_conv_kernels_once(device);  // first call on this device: one-time set-up
    if (rc) return rc;
    float* logits = nullptr;
    rc = noalloc_scratch(device, &logits);
    if (rc) return rc;
    int num_sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device));
    cudaStream_t stream = reinterpret_cast<cudaStream_t>(stream_v);
    // Dense as a 1x1 convolution over M = batch rows, fp32 logits into the fixed scratch (every
    // element of logits[0 : batch*cout) is overwritten, so nothing has to be cleared first) ...
    ConvGeom g{batch, 1, 1, cin, 1, 1, cout, 1, 1, 1, 0};
    ConvLaunch L;
    rc = build_conv_launch(&L, g, in, w, scale, bias, nullptr, logits, /*out_f32=*/true, /*relu=*/0, -1, 0, num_sms,
                           /*w_local=*/false);  // (fp32 logits never take the cluster variant anyway)
    if (rc) return rc;
    rc = run_conv_launch(L, stream);
    if (rc) return rc;
    // ... then the row softmax from the scratch into the caller's output.
    const int warps_per_block = 4;
    const unsigned blocks = (unsigned)((batch + warps_per_block - 1) / warps_per_block);
    adapt::softmax_rows_kernel<<<blocks, warps_per_block * 32, 0, stream>>>(logits, probs, batch, cout);
    CUDA_TRY(cudaGetLastError());
    return ADAPT_OK;
}

// ---------------------------------------------------------------- hand-off plumbing
int adapt_device_alloc(int device, size_t bytes, void** out) {
    CONFIG_LOCK();
    if (!out || bytes == 0) return fail(ADAPT_EINVAL, "device_alloc: bad argument");
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaMalloc(out, bytes));
    CUDA_TRY(cudaMemset(*out, 0, bytes));
    return ADAPT_OK;
}

int adapt_device_free(int device, void* ptr) {
    CONFIG_LOCK();
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaFree(ptr));
    return ADAPT_OK;
}

int adapt_peer_enable(int a, int b) {
    CONFIG_LOCK();
    if (a == b) return ADAPT_OK;
    int rc = require_device(a);
    if (rc) return rc;
    for (int dir = 0; dir < 2; ++dir) {
        const int from = dir == 0 ? a : b, to = dir == 0 ? b : a;
        int can = 0;
        CUDA_TRY(cudaDeviceCanAccessPeer(&can, from, to));
        if (!can) return fail(ADAPT_ECUDA, "device %d cannot access peer %d", from, to);
        CUDA_TRY(cudaSetDevice(from));
        cudaError_t e = cudaDeviceEnablePeerAccess(to, 0);
        if (e == cudaErrorPeerAccessAlreadyEnabled) {
            cudaGetLastError();
        } else if (e != cudaSuccess) {
            return fail(ADAPT_ECUDA, "cudaDeviceEnablePeerAccess(%d->%d): %s", from, to, cudaGetErrorString(e));
        }
    }
    return ADAPT_OK;
}

int adapt_ipc_export(int device, void* ptr, void* handle64_out) {
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!ptr || !handle64_out) return fail(ADAPT_EINVAL, "ipc_export: null");
    int rc = require_device(device);
    if (rc) return rc;
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, ptr));
    memcpy(handle64_out, &h, 64);
    return ADAPT_OK;
}

int adapt_ipc_import(int device, const void* handle64, void** out) {
    CONFIG_LOCK();
    if (!handle64 || !out) return fail(ADAPT_EINVAL, "ipc_import: null");
    int rc = require_device(device);
    if (rc) return rc;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    CUDA_TRY(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
    return ADAPT_OK;
}

int adapt_ipc_close(int device, void* ptr) {
    CONFIG_LOCK();
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaIpcCloseMemHandle(ptr));
    return ADAPT_OK;
}

int adapt_stream_write_flag(int device, void* stream, void* flag_ptr, uint32_t value) {
    int rc = require_device(device);
    if (rc) return rc;
    DriverApi* d;
    rc = get_driver(&d);
    if (rc) return rc;
    CUresult r = d->writeValue32(reinterpret_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(flag_ptr), value,
                                 CU_STREAM_WRITE_VALUE_DEFAULT);
    if (r != CUDA_SUCCESS) return fail(ADAPT_ECUDA, "cuStreamWriteValue32 failed (%d)", (int)r);
    return ADAPT_OK;
}

int adapt_stream_wait_flag_geq(int device, void* stream, void* flag_ptr, uint32_t value) {
    int rc = require_device(device);
    if (rc) return rc;
    DriverApi* d;
    rc = get_driver(&d);
    if (rc) return rc;
    CUresult r = d->waitValue32(reinterpret_cast<CUstream>(stream), reinterpret_cast<CUdeviceptr>(flag_ptr), value,
                                CU_STREAM_WAIT_VALUE_GEQ);
    if (r != CUDA_SUCCESS) return fail(ADAPT_ECUDA, "cuStreamWaitValue32 failed (%d)", (int)r);
    return ADAPT_OK;
}

int adapt_memcpy_h2d(int device, void* dst, const void* src, size_t bytes, void* stream) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}
int adapt_memcpy_d2h(int device, void* dst, const void* src, size_t bytes, void* stream) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}
int adapt_memcpy_d2d(int device, void* dst, const void* src, size_t bytes, void* stream) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}
int adapt_memset(int device, void* dst, int value, size_t bytes, void* stream) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaMemsetAsync(dst, value, bytes, reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}
int adapt_stream_sync(int device, void* stream) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaStreamSynchronize(reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}

struct adapt_heartbeat {
    std::thread th;
    std::atomic<bool> stop{false};
};

int adapt_heartbeat_start(volatile int64_t* word, int interval_us, void** handle_out) {
    if (!word || interval_us <= 0 || !handle_out) return fail(ADAPT_EINVAL, "heartbeat_start: bad argument");
    adapt_heartbeat* hb = new adapt_heartbeat();
    hb->th = std::thread([hb, word, interval_us]() {
        while (!hb->stop.load(std::memory_order_relaxed)) {
            timespec ts;
            clock_gettime(CLOCK_MONOTONIC, &ts);
            __atomic_store_n(const_cast<int64_t*>(word), (int64_t)ts.tv_sec * 1000000000LL + ts.tv_nsec, __ATOMIC_RELEASE);
            std::this_thread::sleep_for(std::chrono::microseconds(interval_us));
        }
    });
    *handle_out = hb;
    return ADAPT_OK;
}

int adapt_heartbeat_stop(void* handle) {
    if (!handle) return ADAPT_OK;
    adapt_heartbeat* hb = reinterpret_cast<adapt_heartbeat*>(handle);
    hb->stop.store(true);
    if (hb->th.joinable()) hb->th.join();
    delete hb;
    return ADAPT_OK;
}

int adapt_stream_query(int device, void* stream, int* busy_out) {
    if (!busy_out) return fail(ADAPT_EINVAL, "stream_query: null");
    int rc = require_device(device);
    if (rc) return rc;
    cudaError_t e = cudaStreamQuery(reinterpret_cast<cudaStream_t>(stream));
    if (e == cudaSuccess) {
        *busy_out = 0;
    } else if (e == cudaErrorNotReady) {
        cudaGetLastError();
        *busy_out = 1;
    } else {
        return fail(ADAPT_ECUDA, "cudaStreamQuery: %s", cudaGetErrorString(e));
    }
    return ADAPT_OK;
}
int adapt_host_register(int device, void* host_ptr, size_t bytes, void** dev_ptr_out) {
    CONFIG_LOCK();
    if (!host_ptr || !bytes || !dev_ptr_out) return fail(ADAPT_EINVAL, "host_register: bad argument");
    int rc = require_device(device);
    if (rc) return rc;
    // several workers of one process share the result ring: the second registration is a no-op
    cudaError_t e = cudaHostRegister(host_ptr, bytes, cudaHostRegisterPortable | cudaHostRegisterMapped);
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        cudaGetLastError();
    } else if (e != cudaSuccess) {
        return fail(ADAPT_ECUDA, "cudaHostRegister failed: %s", cudaGetErrorString(e));
    }
    CUDA_TRY(cudaHostGetDevicePointer(dev_ptr_out, host_ptr, 0));
    return ADAPT_OK;
}
int adapt_host_unregister(int device, void* host_ptr) {
    CONFIG_LOCK();
    int rc = require_device(device);
    if (rc) return rc;
    cudaError_t e = cudaHostUnregister(host_ptr);
    if (e == cudaErrorHostMemoryNotRegistered) {
        cudaGetLastError();  // another worker of this process already released the shared range
    } else if (e != cudaSuccess) {
        return fail(ADAPT_ECUDA, "cudaHostUnregister failed: %s", cudaGetErrorString(e));
    }
    return ADAPT_OK;
}
int adapt_stream_create(int device, void** stream_out) {
    CONFIG_LOCK();
    if (!stream_out) return fail(ADAPT_EINVAL, "stream_create: null");
    int rc = require_device(device);
    if (rc) return rc;
    cudaStream_t s;
    CUDA_TRY(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    *stream_out = s;
    return ADAPT_OK;
}
int adapt_stream_destroy(int device, void* stream) {
    CONFIG_LOCK();
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaStreamDestroy(reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}
int adapt_event_create(int device, void** event_out) {
    if (!event_out) return fail(ADAPT_EINVAL, "event_create: null");
    int rc = require_device(device);
    if (rc) return rc;
    cudaEvent_t e;
    CUDA_TRY(cudaEventCreate(&e));
    *event_out = e;
    return ADAPT_OK;
}
int adapt_event_record(int device, void* event, void* stream) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaEventRecord(reinterpret_cast<cudaEvent_t>(event), reinterpret_cast<cudaStream_t>(stream)));
    return ADAPT_OK;
}
int adapt_event_sync(int device, void* event) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaEventSynchronize(reinterpret_cast<cudaEvent_t>(event)));
    return ADAPT_OK;
}
int adapt_event_elapsed_ms(int device, void* start, void* stop, float* ms_out) {
    if (!ms_out) return fail(ADAPT_EINVAL, "event_elapsed: null");
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaEventElapsedTime(ms_out, reinterpret_cast<cudaEvent_t>(start), reinterpret_cast<cudaEvent_t>(stop)));
    return ADAPT_OK;
}
int adapt_event_query(int device, void* event, int* done_out) {
    if (!done_out) return fail(ADAPT_EINVAL, "event_query: null");
    int rc = require_device(device);
    if (rc) return rc;
    cudaError_t e = cudaEventQuery(reinterpret_cast<cudaEvent_t>(event));
    if (e == cudaSuccess) {
        *done_out = 1;
    } else if (e == cudaErrorNotReady) {
        cudaGetLastError();
        *done_out = 0;
    } else {
        return fail(ADAPT_ECUDA, "cudaEventQuery: %s", cudaGetErrorString(e));
    }
    return ADAPT_OK;
}

int adapt_debug_spin(int device, void* stream, int microseconds) {
    if (microseconds <= 0) return ADAPT_OK;
    int rc = require_device(device);
    if (rc) return rc;
    adapt::spin_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>((long long)microseconds * 1000LL);
    CUDA_TRY(cudaGetLastError());
    return ADAPT_OK;
}

int adapt_event_destroy(int device, void* event) {
    int rc = require_device(device);
    if (rc) return rc;
    CUDA_TRY(cudaEventDestroy(reinterpret_cast<cudaEvent_t>(event)));
    return ADAPT_OK;
}

}  // extern "C"
