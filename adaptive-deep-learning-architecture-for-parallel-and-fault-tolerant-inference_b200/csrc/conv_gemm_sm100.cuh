// Implicit-GEMM convolution for one pipeline stage's Conv2D+bias+BN(+Add)(+ReLU) layers.
//
// Replaces the arithmetic the reference delegates to `model.predict` (reference
// src/node.py:177) for every Conv2D / BatchNormalization / Add / ReLU group of the Keras
// ResNet-50 graph, and for the final Dense layer (as a 1x1 "convolution" over M = batch).
//
//   out[m, n] = act( (sum_k A[m, k] * Wt[n, k]) * scale[n] + bias[n] (+ residual[m, n]) )
//
//   A  : NHWC bf16 activations, viewed as the im2col matrix [M = B*Ho*Wo, K = kh*kw*Cin]
//   Wt : bf16 weights [Cout, kh, kw, Cin]  (K-major "B" operand)
//   acc: fp32 in TMEM (tcgen05.mma kind::f16), epilogue in fp32, output bf16 (or fp32)
//
// Structure: persistent CTAs (one per SM), warp-specialised:
//   warp 0        TMA producer (A: tiled 2-D box, im2col box, or a padded-flat row segment that serves the three
//                 horizontal taps of a 3x3 filter row (A_ROW3); B: tiled 2-D box, optionally resident)
//   warp 1        MMA issuer   (one elected lane issues tcgen05.mma; owns the TMEM allocation; in CTA-pair mode
//                 only the leader CTA issues, for both)
//   warps 2..9    epilogue, two groups of four: group g drains TMEM accumulator stage g
//                 (tcgen05.ld -> scale/bias(/residual) -> cvt[.relu].bf16x2 -> 4 KiB smem slab -> TMA store;
//                 with a residual the slab is used in place, filled by a per-warp TMA prefetch that runs ahead
//                 across tile boundaries)
//   warps 10..13  (A_GATHER only) software im2col producers for thin-channel inputs (odd stem cuts)
// Three pipelines: smem full/empty ring (TMA <-> MMA), TMEM full/empty double buffer (MMA <-> epilogue), and a
// static round-robin tile schedule.  The shared-memory budget is split at launch time between operand stages and
// epilogue slabs (ConvGemmParams::stages/d_out/d_res/b_resident).  Launched with programmatic stream
// serialization: the prologue overlaps the previous layer's tail, griddepcontrol.wait precedes the first access
// to activations.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cuda.h>
#include <cstdint>
#include <type_traits>

#include "ptx_sm100.cuh"

namespace adapt {

// A_ROW3: 3x3 / stride 1 / pad 1 convolutions with the three horizontal taps of a filter row served from ONE
// shared-memory tile.  The output is walked in "padded-flat" order (rows of W + 2 positions, the last two of
// each row are discarded), so that tap s of output position m is input position m + s of the same walk: one
// im2col load of 128 + 2 (rounded to 136) positions per (filter row, channel block) feeds three k-blocks
// through descriptors whose start is shifted by s pixel rows.  The A bytes an SM ingests drop 3x.
//
// A_S2DP: the space-to-depth stem (4x4 / stride 1 over 16-channel pixels, 64 output channels) by column PHASE.
// The image rows are stored 4k pixels wide, so the image is a sequence of 128-byte UNITS of four pixels.  An
// output pixel q = 4j + phi needs pixels q .. q+3 of filter row r, i.e. the 32-byte chunks phi .. phi+3 counted
// from unit j: chunk (phi + s) lives in unit row (phi+s)/4 at byte ((phi+s) mod 4) * 32.  One plain 2-D load of
// 128 + 1 (rounded to 136) consecutive units per filter row therefore serves 128 output positions j of ALL FOUR
// phases -- 512 output pixels -- through descriptors shifted by whole rows and by K-chunks, into four 64-column
// accumulators (the "wide pixel" im2col path fetches the same bytes four times, at 32-byte offsets).  The walk
// runs over units of all H rows (output rows >= Ho and the unit column beyond Wo are discarded).
enum AMode : int { A_TILED = 0, A_IM2COL = 1, A_GATHER = 2, A_ROW3 = 3, A_S2DP = 4 };

struct ConvGemmParams {
    int M;             // output pixels in this launch: batch * Ho * Wo
    int N;             // output channels
    int num_k_blocks;  // ceil(K / 64)
    int m_tiles, n_tiles;
    // convolution geometry (A_IM2COL / A_GATHER)
    int H, W, Cin;  // stored input channels per pixel
    int Ho, Wo;
    int kh, kw, stride, pad;
    int c_blocks;  // Cin / 64 (A_IM2COL)
    int K_real;    // kh*kw*Cin (A_GATHER zero-fills beyond it)
    // epilogue
    const float* scale;             // [N] folded BN scale (gamma / sqrt(var + eps)), or 1
    const float* bias;              // [N] folded (conv_bias - mean) * scale + beta
    const __nv_bfloat16* residual;  // [M, N] or nullptr
    __nv_bfloat16* out_bf16;        // [M, N] or nullptr
    float* out_f32;                 // [M, N] or nullptr (Dense logits)
    int relu;
    int epi_tma;  // 1: staged epilogue (TMA residual ring + TMA store); 2: staged + coalesced st.global; 0: direct
    // shared-memory split (chosen by the host per layer)
    int stages;  // operand pipeline depth, 2..kMaxStages
    int d_out;   // output staging slabs per epilogue warp (2 or 4)
    int d_res;   // residual staging slabs per epilogue warp (2..kMaxRes; ignored without residual)
    // 1: the whole B operand (n_tiles == 1, num_k_blocks tiles) stays in shared memory; loaded with the CTA's
    // first tile, after which the operand ring carries A only
    int b_resident;
    // A_ROW3: positions per padded-flat row (W + 2), valid positions per row (W); M and m_tiles count padded-flat
    // positions.
    int row_pitch, row_valid;
    // A_GATHER source
    const __nv_bfloat16* in;
    // developer timeline (nullptr in production): [grid][kTraceTiles][kTraceEvents] SM clock stamps
    long long* trace;
};

#ifndef ADAPT_ENABLE_TRACE
#define ADAPT_ENABLE_TRACE 0
#endif

// Timeline slots written when ConvGemmParams::trace is set (tools/kernel_check.py --group trace).
constexpr int kTraceTiles = 24;
constexpr int kTraceEvents = 8;
enum : int {
    TR_CTA_START = 0,   // (tile 0 only) kernel entry
    TR_PROD_FIRST = 1,  // producer: first load of the tile issued
    TR_PROD_LAST = 2,   // producer: last load of the tile issued
    TR_MMA_START = 3,   // MMA warp: accumulator stage acquired
    TR_MMA_FIRST = 4,   // MMA warp: first operand stage arrived
    TR_MMA_DONE = 5,    // MMA warp: last commit issued
    TR_EPI_START = 6,   // epilogue warp 0 of the tile's group: accumulator ready
    TR_EPI_DONE = 7,    // epilogue warp 0 of the tile's group: last store issued
};

constexpr int kBlockM = 128;
constexpr int kBlockK = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int kUmmaK = 16;
constexpr int kABytes = kBlockM * kBlockK * 2;
constexpr int kRow3Rows = 136;  // 128 + 2 shifted rows, rounded up to whole 8-row swizzle groups
constexpr int kRow3ABytes = kRow3Rows * kBlockK * 2;  // 17 KiB
constexpr int kMaxStages = 8;
constexpr int kMaxRes = 8;

// Epilogue staging unit: one warp's 32-row x 64-column bf16 slab in the 128-byte-swizzled layout.
constexpr int kEpiCols = 64;
constexpr int kEpiSlabBytes = 32 * kEpiCols * 2;  // 4 KiB
constexpr int kBarrierBytes = 1024;               // full[8] empty[8] tmem_full[2] tmem_empty[2] res[8][8] + tmem slot
constexpr int kMaxSmemBytes = 227 * 1024;
constexpr int kEpiWarps = 8;  // two groups of four: group g drains accumulator stage g (tiles g, g+2, ...)

// (CTA-pair mode: each CTA of the pair holds only half of the B tile)
__host__ __device__ constexpr int conv_gemm_stage_bytes(int block_n, bool pair = false) {
    return kABytes + (pair ? block_n / 2 : block_n) * kBlockK * 2;
}
__host__ __device__ constexpr int conv_gemm_epi_bytes(int block_n, int d_out, int d_res) {
    return kEpiWarps * (d_out + d_res) * kEpiSlabBytes + kEpiWarps * 2 * block_n * 4;
}
// (b_resident_blocks > 0: B is held once, `b_resident_blocks` k-blocks deep, instead of once per stage)
// (row3: a stage is one 17 KiB A tile plus the three B tiles of a filter row)
__host__ __device__ constexpr int conv_gemm_smem_bytes(int block_n, int stages, int d_out, int d_res, bool pair = false,
                                                       int b_resident_blocks = 0, bool row3 = false) {
    const int a_bytes = row3 ? kRow3ABytes : kABytes;
    const int b_bytes = conv_gemm_stage_bytes(block_n, pair) - kABytes;
    return (b_resident_blocks > 0 ? stages * a_bytes + b_resident_blocks * b_bytes
                                  : stages * (a_bytes + (row3 ? 3 : 1) * b_bytes)) +
           conv_gemm_epi_bytes(block_n, d_out, d_res) + kBarrierBytes + 1024;
}

template <int A_MODE>
constexpr int conv_gemm_threads() {
    return A_MODE == A_GATHER ? 448 : 320;  // TMA + MMA + 8 epilogue warps (+ 4 gather warps)
}

__device__ __forceinline__ void epi_store_wait_read(int allowed_pending) {
    if (allowed_pending >= 3)
        ptx::tma_store_wait_read<3>();
    else if (allowed_pending >= 1)
        ptx::tma_store_wait_read<1>();
    else
        ptx::tma_store_wait_read<0>();
}

// CL = thread-block cluster size (1, 2 or 4).  With CL > 1 the CL CTAs of a cluster work on CL
// consecutive m-tiles of the same n-tile and share the B (weight) operand: each CTA loads 1/CL of
// the B tile and TMA-multicasts it into every CTA's shared memory, so the L2 -> SM traffic for B
// drops CL-fold (the 3x3 and conv5 layers are L2-bandwidth-bound with private B tiles).
//
// PAIR (needs CL == 2) = CTA-pair tensor-core mode: the two CTAs of a cluster form one 256 x BLOCK_N tile
// computed by `tcgen05.mma.cta_group::2`, issued by the leader (even) CTA.  Each CTA loads its own 128
// A rows and only HALF of the B tile, and drains its own 128 accumulator rows from its own TMEM.  This
// halves the shared-memory traffic for B (TMA fill + operand read), which is what bounds the 3x3 and
// conv5 layers with single-CTA tiles.
template <int BLOCK_N, int A_MODE, int CL = 1, bool PAIR = false>
__global__ void __launch_bounds__(conv_gemm_threads<A_MODE>(), 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_res,
                 const ConvGemmParams p) {
    static_assert(BLOCK_N % 64 == 0 && BLOCK_N >= 64 && BLOCK_N <= 256, "BLOCK_N");
    static_assert(CL == 1 || CL == 2 || CL == 4, "cluster size");
    static_assert(CL == 1 || A_MODE != A_GATHER, "the gather path does not multicast");
    static_assert(!PAIR || CL == 2, "a CTA pair is a cluster of two");
    static_assert(A_MODE != A_ROW3 || CL == 1 || PAIR, "row-reuse mode: single CTA or CTA pair");
    constexpr bool ROW3 = A_MODE == A_ROW3;
    constexpr bool S2DP = A_MODE == A_S2DP;
    static_assert(!S2DP || (BLOCK_N == 256 && CL == 1 && !PAIR), "stem-by-phase: four 64-column accumulators, single CTA");
    constexpr int kMmaN = S2DP ? 64 : BLOCK_N;  // N of one tcgen05.mma
    constexpr int kBBytes = (PAIR ? kMmaN / 2 : kMmaN) * kBlockK * 2;
    constexpr int kATile = (ROW3 || S2DP) ? kRow3ABytes : kABytes;  // A bytes per stage
    constexpr int kBPerStage = ROW3 ? 3 : 1;              // B tiles (k-blocks) per stage
    constexpr int kStageBytes = kATile + kBPerStage * kBBytes;
    constexpr int kTmemCols = 2 * BLOCK_N;  // double-buffered accumulator (power of 2 >= 32)
    const int kStages = p.stages;
    const int d_out = p.d_out;
    const int d_res = p.residual != nullptr ? p.d_res : 0;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + kStages * kATile;
    const int b_slots = p.b_resident ? p.num_k_blocks : kStages * kBPerStage;  // B tiles held: whole operand, or per stage
    uint8_t* smem_epi = smem + kStages * kATile + b_slots * kBBytes;  // [8 warps][ring] slabs, 1024-aligned
    const int k_iters = ROW3 ? 3 * p.c_blocks : p.num_k_blocks;  // pipeline steps per tile
    const int epi_warp_bytes = (p.d_out + p.d_res) * kEpiSlabBytes;
    float* smem_sb = reinterpret_cast<float*>(smem_epi + kEpiWarps * epi_warp_bytes);  // per warp [scale | bias]
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(smem_sb) + kEpiWarps * 2 * BLOCK_N * 4);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + kMaxStages;
    uint64_t* tmem_full_bar = bars + 2 * kMaxStages;
    uint64_t* tmem_empty_bar = bars + 2 * kMaxStages + 2;
    uint64_t* res_bar = bars + 2 * kMaxStages + 4;  // [kEpiWarps][kMaxRes]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kMaxStages + 4 + kEpiWarps * kMaxRes);

    const int warp_idx = __shfl_sync(0xffffffff, threadIdx.x / 32, 0);
    const int lane = threadIdx.x % 32;
    // Work distribution.  A "unit" is one tile (CL == 1) or CL consecutive m-tiles of one n-tile (CL > 1);
    // units are dealt round-robin to CTAs / clusters, n fastest so that concurrently running CTAs share
    // A through L2.  Every role of a CTA walks the same sequence with tile_at().
    const int cl_rank = CL > 1 ? static_cast<int>(ptx::cluster_ctarank()) : 0;
    const long long num_units = CL > 1 ? static_cast<long long>((p.m_tiles + CL - 1) / CL) * p.n_tiles
                                       : static_cast<long long>(p.m_tiles) * p.n_tiles;
    const int unit0 = blockIdx.x / CL;
    const int unit_step = gridDim.x / CL;
    constexpr uint16_t kClusterMask = static_cast<uint16_t>((1u << CL) - 1);
    auto tile_at = [&](long long i, int& m_blk, int& n_blk, bool& m_valid) -> bool {
        const long long u = unit0 + i * unit_step;
        if (u >= num_units) return false;
        const int g = static_cast<int>(u / p.n_tiles);
        n_blk = static_cast<int>(u - static_cast<long long>(g) * p.n_tiles);
        m_blk = g * CL + cl_rank;
        m_valid = m_blk < p.m_tiles;  // (a cluster's last unit may hang over the end of M)
        return true;
    };

    // (compiled in only for the developer build `ADAPT_BUILD_TRACE=1 python csrc/build.py`: even a never-taken
    // branch per k-block in the MMA / producer loops costs the 3x3 layers a few percent)
    auto stamp = [&](long long tile, int ev) {
#if ADAPT_ENABLE_TRACE
        if (p.trace != nullptr && tile < kTraceTiles)
            p.trace[(static_cast<long long>(blockIdx.x) * kTraceTiles + tile) * kTraceEvents + ev] = clock64();
#else
        (void)tile;
        (void)ev;
#endif
    };
    if (threadIdx.x == 0) stamp(0, TR_CTA_START);

    if (warp_idx == 0) {
        // 84 mbarriers, initialised by the 32 lanes in parallel
        constexpr uint32_t full_count = (A_MODE == A_GATHER) ? (1 + 128) : 1;
        constexpr int kNumBars = 2 * kMaxStages + 4 + kEpiWarps * kMaxRes;
        for (int i = lane; i < kNumBars; i += 32) {
            uint32_t count = 1;  // tmem_full, residual slots
            if (i < kMaxStages)
                count = full_count;
            else if (i < 2 * kMaxStages)
                count = PAIR ? 1 : CL;  // multicast: every CTA of the cluster must have drained the stage; pair: the leader's MMA commit
            else if (i >= 2 * kMaxStages + 2 && i < 2 * kMaxStages + 4)
                count = PAIR ? 8 : 4;  // tmem_empty: one arrive per epilogue warp (of both CTAs of a pair)
            ptx::mbar_init(&bars[i], count);
        }
        ptx::fence_mbar_init();
    } else if (warp_idx == 2 && lane == 0) {
        // descriptor fetches start at once, on a warp that has nothing else to do yet
        if (A_MODE != A_GATHER) ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_b);
        if (p.epi_tma) {
            ptx::prefetch_tmap(&tmap_out);
            if (p.residual != nullptr) ptx::prefetch_tmap(&tmap_res);
        }
    }
    if (warp_idx == 1) {
        if (PAIR) {
            ptx::tmem_alloc_pair(tmem_slot, kTmemCols);
            ptx::tmem_relinquish_pair();
        } else {
            ptx::tmem_alloc(tmem_slot, kTmemCols);
            ptx::tmem_relinquish();
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (CL > 1) ptx::cluster_sync();  // peers' barriers are initialised before anyone multicasts into them
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    // Programmatic dependent launch: everything above ran while the previous layer may still have been
    // finishing on other SMs.  Let the next layer start the same way, pull this CTA's first weight tiles
    // (which no kernel writes) towards L2, then wait for the previous layer's activations to be complete.
    ptx::griddep_launch_dependents();
    if (warp_idx == 0 && ptx::elect_one()) {
        int mb = 0, nb = 0;
        bool mv = true;
        if (tile_at(0, mb, nb, mv)) {
            const int rows = PAIR ? BLOCK_N / 2 : (CL == 1 ? kMmaN : BLOCK_N / CL);
            const int row0 = nb * BLOCK_N + (CL > 1 ? cl_rank * rows : 0);
            const int nkb = p.num_k_blocks < kStages ? p.num_k_blocks : kStages;
            for (int kb = 0; kb < nkb; ++kb) ptx::tma_prefetch_2d(&tmap_b, kb * kBlockK, row0);
        }
    }
    ptx::griddep_wait();

    if (warp_idx == 0) {
        // ===================== TMA producer =====================
        // The whole warp walks the loop (uniform control flow keeps the coordinates in uniform registers and
        // the per-k-block instruction count low -- measured: a single-lane loop with a tap division per
        // k-block, not the memory system, was pacing the 3x3 layers); one elected lane issues.
        int stage = 0;
        uint32_t phase = 0;
        int m_blk_raw = 0, n_blk = 0;
        bool m_valid = true;
        for (long long ti = 0; tile_at(ti, m_blk_raw, n_blk, m_valid); ++ti) {
            const int m_blk = m_valid ? m_blk_raw : p.m_tiles - 1;  // overhanging CTA: re-load a valid tile
            if constexpr (S2DP) {
                // one filter row per stage: 136 consecutive units starting at this tile's first unit, r rows down;
                // the whole weight matrix (4 k-blocks of 64 x 64) arrives with the CTA's first tile and stays
                for (int r4 = 0; r4 < k_iters; ++r4) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (ptx::elect_one()) {
                        if (r4 == 0) stamp(ti, TR_PROD_FIRST);
                        if (r4 == k_iters - 1) stamp(ti, TR_PROD_LAST);
                        const bool load_b = ti == 0;
                        ptx::mbar_expect_tx(&full_bar[stage], kATile + (load_b ? kBBytes : 0));
                        ptx::tma_load_2d(smem_a + stage * kATile, &tmap_a, &full_bar[stage], 0, m_blk * kBlockM + r4 * p.row_pitch);
                        if (load_b) ptx::tma_load_2d(smem_b + r4 * kBBytes, &tmap_b, &full_bar[stage], r4 * kBlockK, 0);
                    }
                    __syncwarp();
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                continue;
            }
            if constexpr (ROW3) {
                // start of this tile in the padded-flat walk: position q0 of row p0 of image img0
                const int m0 = m_blk * kBlockM;
                const int rest = m0 / p.row_pitch;
                const int q0 = m0 - rest * p.row_pitch;
                const int img0 = rest / p.H;
                const int p0 = rest - img0 * p.H;
                const int b_row3 = PAIR ? n_blk * BLOCK_N + cl_rank * (BLOCK_N / 2) : n_blk * BLOCK_N;
                int tap_r3 = 0, cb3 = 0;
                for (int it = 0; it < k_iters; ++it) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    if (ptx::elect_one()) {
                        if (it == 0) stamp(ti, TR_PROD_FIRST);
                        if (it == k_iters - 1) stamp(ti, TR_PROD_LAST);
                        const bool load_b = !p.b_resident || ti == 0;  // resident B arrives with the CTA's first tile
                        const uint32_t bytes = kATile + (load_b ? 3 * kBBytes : 0);
                        uint8_t* a_dst = smem_a + stage * kATile;
                        if (PAIR) {
                            if (cl_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2 * bytes);
                            ptx::tma_load_im2col_4d_pair(a_dst, &tmap_a, &full_bar[stage], cb3 * kBlockK, q0 - 1, p0 - 1, img0, 0,
                                                         static_cast<uint16_t>(tap_r3));
                        } else {
                            ptx::mbar_expect_tx(&full_bar[stage], bytes);
                            ptx::tma_load_im2col_4d(a_dst, &tmap_a, &full_bar[stage], cb3 * kBlockK, q0 - 1, p0 - 1, img0, 0,
                                                    static_cast<uint16_t>(tap_r3));
                        }
                        if (load_b) {
#pragma unroll
                            for (int s3 = 0; s3 < 3; ++s3) {
                                const int kbi = (tap_r3 * 3 + s3) * p.c_blocks + cb3;  // k-block of tap (r, s), channels cb
                                uint8_t* b_dst = smem_b + (p.b_resident ? kbi : stage * 3 + s3) * kBBytes;
                                if (PAIR)
                                    ptx::tma_load_2d_pair(b_dst, &tmap_b, &full_bar[stage], kbi * kBlockK, b_row3);
                                else
                                    ptx::tma_load_2d(b_dst, &tmap_b, &full_bar[stage], kbi * kBlockK, b_row3);
                            }
                        }
                    }
                    __syncwarp();
                    if (++cb3 == p.c_blocks) {
                        cb3 = 0;
                        ++tap_r3;
                    }
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                continue;
            }
            int img = 0, base_w = 0, base_h = 0;
            if (A_MODE == A_IM2COL) {
                const int m0 = m_blk * kBlockM;
                const int hw = p.Ho * p.Wo;
                img = m0 / hw;
                const int rem = m0 - img * hw;
                const int po = rem / p.Wo;
                const int qo = rem - po * p.Wo;
                base_w = qo * p.stride - p.pad;
                base_h = po * p.stride - p.pad;
            }
            const int a_row = m_blk * kBlockM;
            const int b_row = PAIR ? n_blk * BLOCK_N + cl_rank * (BLOCK_N / 2)
                                   : (CL == 1 ? n_blk * BLOCK_N : n_blk * BLOCK_N + cl_rank * (BLOCK_N / CL));
            int tap_r = 0, tap_s = 0, cb = 0;  // (filter tap, 64-channel block) of the current k-block
            int k_elem = 0;
            for (int kb = 0; kb < p.num_k_blocks; ++kb, k_elem += kBlockK) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                if (ptx::elect_one()) {
                    if (kb == 0) stamp(ti, TR_PROD_FIRST);
                    if (kb == p.num_k_blocks - 1) stamp(ti, TR_PROD_LAST);
                    uint8_t* a_dst = smem_a + stage * kABytes;
                    uint8_t* b_dst = smem_b + (p.b_resident ? kb : stage) * kBBytes;
                    const bool load_b = !p.b_resident || ti == 0;  // resident B arrives with the CTA's first tile
                    if (!PAIR && CL == 1 && A_MODE != A_GATHER && !load_b) {
                        ptx::mbar_expect_tx(&full_bar[stage], kABytes);
                    } else if (PAIR) {
                        // both CTAs' loads report to the LEADER's barrier, which the leader arms for all of them
                        if (cl_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2 * kStageBytes);
                    } else if (A_MODE == A_GATHER) {
                        ptx::mbar_expect_tx(&full_bar[stage], kBBytes);
                    } else {
                        ptx::mbar_expect_tx(&full_bar[stage], kStageBytes);
                    }
                    if (A_MODE == A_TILED) {
                        if (PAIR)
                            ptx::tma_load_2d_pair(a_dst, &tmap_a, &full_bar[stage], k_elem, a_row);
                        else
                            ptx::tma_load_2d(a_dst, &tmap_a, &full_bar[stage], k_elem, a_row);
                    } else if (A_MODE == A_IM2COL) {
                        if (PAIR)
                            ptx::tma_load_im2col_4d_pair(a_dst, &tmap_a, &full_bar[stage], cb * kBlockK, base_w, base_h, img,
                                                         static_cast<uint16_t>(tap_s), static_cast<uint16_t>(tap_r));
                        else
                            ptx::tma_load_im2col_4d(a_dst, &tmap_a, &full_bar[stage], cb * kBlockK, base_w, base_h, img,
                                                    static_cast<uint16_t>(tap_s), static_cast<uint16_t>(tap_r));
                    }
                    if (PAIR) {
                        // this CTA's half of the B tile (the pair's MMA reads both halves in place)
                        ptx::tma_load_2d_pair(b_dst, &tmap_b, &full_bar[stage], k_elem, b_row);
                    } else if (CL == 1) {
                        if (load_b) ptx::tma_load_2d(b_dst, &tmap_b, &full_bar[stage], k_elem, b_row);
                    } else {
                        // this CTA's 1/CL slice of the B tile, delivered to all CTAs of the cluster
                        constexpr int kSliceRows = BLOCK_N / CL;
                        ptx::tma_load_2d_multicast(b_dst + cl_rank * kSliceRows * kBlockK * 2, &tmap_b, &full_bar[stage],
                                                   k_elem, b_row, kClusterMask);
                    }
                }
                __syncwarp();
                if (A_MODE == A_IM2COL) {  // next (tap, channel block), without dividing
                    if (++cb == p.c_blocks) {
                        cb = 0;
                        if (++tap_s == p.kw) {
                            tap_s = 0;
                            ++tap_r;
                        }
                    }
                }
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp_idx == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(PAIR ? 2 * kBlockM : kBlockM, kMmaN);
        int stage = 0;
        uint32_t phase = 0;
        int m_blk_u = 0, n_blk_u = 0;
        bool m_valid_u = true;
        // (pair mode: only the leader CTA issues MMAs; the peer's warp 1 just owns its TMEM allocation)
        for (long long tile_iter = 0; !(PAIR && cl_rank != 0) && tile_at(tile_iter, m_blk_u, n_blk_u, m_valid_u);
             ++tile_iter) {
            const int acc_stage = static_cast<int>(tile_iter & 1);
            const uint32_t acc_phase = static_cast<uint32_t>(tile_iter >> 1) & 1;
            ptx::mbar_wait(&tmem_empty_bar[acc_stage], acc_phase ^ 1);
            ptx::tc_fence_after();
            if (lane == 0) stamp(tile_iter, TR_MMA_START);
            const uint32_t tmem_d = tmem_base + acc_stage * BLOCK_N;
            if constexpr (S2DP) {
                for (int r4 = 0; r4 < k_iters; ++r4) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    if (r4 == 0 && lane == 0) stamp(tile_iter, TR_MMA_FIRST);
                    if (ptx::elect_one()) {
                        const uint32_t a_base = ptx::smem_u32(smem_a + stage * kATile);
                        const uint64_t b_desc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(smem_b + r4 * kBBytes));
#pragma unroll
                        for (int phi = 0; phi < 4; ++phi) {      // column phase: accumulator phi
#pragma unroll
                            for (int s4 = 0; s4 < 4; ++s4) {     // filter column = 32-byte chunk (phi + s) from unit j
                                const int chunk = phi + s4;
                                const uint64_t a_desc =
                                    ptx::make_kmajor_sw128_desc(a_base + (chunk >> 2) * (kBlockK * 2) + (chunk & 3) * 32);
                                ptx::umma_bf16_ss(tmem_d + phi * kMmaN, a_desc, b_desc + 2 * s4, idesc, (r4 | s4) != 0 ? 1u : 0u);
                            }
                        }
                        ptx::umma_commit(&empty_bar[stage]);
                        if (r4 == k_iters - 1) {
                            ptx::umma_commit(&tmem_full_bar[acc_stage]);
                            stamp(tile_iter, TR_MMA_DONE);
                        }
                    }
                    __syncwarp();
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
            if constexpr (ROW3) {
                int tap_r3 = 0, cb3 = 0;
                for (int it = 0; it < k_iters; ++it) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    if (it == 0 && lane == 0) stamp(tile_iter, TR_MMA_FIRST);
                    if (ptx::elect_one()) {
                        const uint32_t a_base = ptx::smem_u32(smem_a + stage * kATile);
#pragma unroll
                        for (int s3 = 0; s3 < 3; ++s3) {
                            const int kbi = (tap_r3 * 3 + s3) * p.c_blocks + cb3;
                            // tap s reads the same tile one pixel row (128 B) further on.  The 128-byte swizzle is a
                            // function of the absolute shared-memory address (measured: a start shifted by whole
                            // rows is read correctly with the descriptor's base-offset field left at 0, and wrongly
                            // with the row phase written into it), so only the start address moves.
                            const uint64_t a_desc = ptx::make_kmajor_sw128_desc(a_base + s3 * (kBlockK * 2));
                            const uint64_t b_desc = ptx::make_kmajor_sw128_desc(
                                ptx::smem_u32(smem_b + (p.b_resident ? kbi : stage * 3 + s3) * kBBytes));
#pragma unroll
                            for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                                if (PAIR)
                                    ptx::umma_bf16_ss_pair(tmem_d, a_desc + 2 * k, b_desc + 2 * k, idesc, (it | s3 | k) != 0 ? 1u : 0u);
                                else
                                    ptx::umma_bf16_ss(tmem_d, a_desc + 2 * k, b_desc + 2 * k, idesc, (it | s3 | k) != 0 ? 1u : 0u);
                            }
                        }
                        if (PAIR)
                            ptx::umma_commit_pair(&empty_bar[stage], kClusterMask);
                        else
                            ptx::umma_commit(&empty_bar[stage]);
                        if (it == k_iters - 1) {
                            if (PAIR)
                                ptx::umma_commit_pair(&tmem_full_bar[acc_stage], kClusterMask);
                            else
                                ptx::umma_commit(&tmem_full_bar[acc_stage]);
                            stamp(tile_iter, TR_MMA_DONE);
                        }
                    }
                    __syncwarp();
                    if (++cb3 == p.c_blocks) {
                        cb3 = 0;
                        ++tap_r3;
                    }
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
            if constexpr (!ROW3 && !S2DP) for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                ptx::mbar_wait(&full_bar[stage], phase);
                ptx::tc_fence_after();
                if (kb == 0 && lane == 0) stamp(tile_iter, TR_MMA_FIRST);
                if (ptx::elect_one()) {
                    const uint64_t a_desc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(smem_a + stage * kABytes));
                    const uint64_t b_desc =
                        ptx::make_kmajor_sw128_desc(ptx::smem_u32(smem_b + (p.b_resident ? kb : stage) * kBBytes));
#pragma unroll
                    for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                        // advance 16 elements (32 B) along K inside the swizzle atom: +2 in 16-byte units
                        if (PAIR)
                            ptx::umma_bf16_ss_pair(tmem_d, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                        else
                            ptx::umma_bf16_ss(tmem_d, a_desc + 2 * k, b_desc + 2 * k, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    // free the smem slot once these MMAs retire (in every CTA whose loads refill it)
                    if (PAIR)
                        ptx::umma_commit_pair(&empty_bar[stage], kClusterMask);
                    else if (CL == 1)
                        ptx::umma_commit(&empty_bar[stage]);
                    else
                        ptx::umma_commit_multicast(&empty_bar[stage], kClusterMask);
                    if (kb == p.num_k_blocks - 1) {
                        if (PAIR)
                            ptx::umma_commit_pair(&tmem_full_bar[acc_stage], kClusterMask);  // both CTAs' epilogues
                        else
                            ptx::umma_commit(&tmem_full_bar[acc_stage]);
                        stamp(tile_iter, TR_MMA_DONE);
                    }
                }
                __syncwarp();
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp_idx < 2 + kEpiWarps) {
        // ===================== epilogue =====================
        // Two groups of four warps; group g owns accumulator stage g, i.e. this CTA's tiles g, g+2, g+4, ...
        // so that two tiles are being drained at any time (one warp per SM sub-partition per group).
        const int epi_warp = warp_idx - 2;   // 0..7: staging buffers, barriers
        const int group = epi_warp >> 2;     // accumulator stage this warp drains
        const int lane_group = warp_idx & 3;  // TMEM lanes this warp may touch: [32*lane_group, +32)
        constexpr int kChunks = BLOCK_N / kEpiCols;
        // Per-warp ring of 4 KiB slabs.  Without a residual a slab only stages the outgoing chunk.  With one,
        // the slab is used IN PLACE: TMA loads the residual chunk into it, the warp adds and overwrites, TMA
        // stores it; the slab stored one chunk ago is then refilled with the residual `ring - 1` chunks ahead.
        const bool has_res = d_res > 0;
        const uint32_t ring = static_cast<uint32_t>(has_res ? d_res : d_out);
        uint8_t* my_slabs = smem_epi + epi_warp * epi_warp_bytes;
        const uint32_t slabs_sa = ptx::smem_u32(my_slabs);
        float* my_scale = smem_sb + epi_warp * 2 * BLOCK_N;
        const uint32_t scale_sa = ptx::smem_u32(my_scale);  // shared-space addresses for explicit ld/st.shared
        const uint32_t bias_sa = scale_sa + BLOCK_N * 4;
        uint64_t* my_res_bar = res_bar + epi_warp * kMaxRes;
        uint32_t res_par = 0;  // bit s: parity the next wait on slab s expects
        uint32_t slot = 0;     // slab of the chunk being drained

        // Residual prefetch cursor: walks this warp's chunks in consumption order, one tile_at() per tile.
        long long ri_iter = group;
        int ri_c = 0, ri_row0 = 0, ri_col0 = 0;
        uint32_t ri_slot = 0;
        bool ri_live = false;
        auto ri_locate = [&]() {
            int mb = 0, nb = 0;
            bool mv = true;
            ri_live = false;
            if (tile_at(ri_iter, mb, nb, mv)) {
                ri_row0 = mb * kBlockM + lane_group * 32;
                ri_col0 = nb * BLOCK_N;
                ri_live = mv && ri_row0 < p.M;
            }
        };
        auto issue_residual = [&]() {
            if (ri_live && lane == 0) {
                ptx::mbar_expect_tx(&my_res_bar[ri_slot], kEpiSlabBytes);
                ptx::tma_load_2d(my_slabs + ri_slot * kEpiSlabBytes, &tmap_res, &my_res_bar[ri_slot],
                                 ri_col0 + ri_c * kEpiCols, ri_row0);
            }
            if (++ri_slot == ring) ri_slot = 0;
            if (++ri_c == kChunks) {
                ri_c = 0;
                ri_iter += 2;
                ri_locate();
            }
        };
        if (p.epi_tma && has_res) {
            ri_locate();
            for (uint32_t i = 0; i + 1 < ring; ++i) issue_residual();  // prime: ring - 1 chunks in flight
        }

        int m_blk = 0, n_blk = 0;
        bool m_valid = true;
        for (long long tile_iter = group; tile_at(tile_iter, m_blk, n_blk, m_valid); tile_iter += 2) {
            const int acc_stage = group;  // == tile_iter & 1
            const uint32_t acc_phase = static_cast<uint32_t>(tile_iter >> 1) & 1;
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(lane_group * 32) << 16) + acc_stage * BLOCK_N;
            if (p.epi_tma) {
                // ---------- staged epilogue: TMEM -> regs -> swizzled smem slab -> TMA store ----------
                const int m_row0 = m_blk * kBlockM + lane_group * 32;
                const int n_tile0 = n_blk * BLOCK_N;
                const bool slab_ok = m_valid && m_row0 < p.M;
                // this tile's scale / bias -> per-warp smem (read back as broadcasts)
                for (int i = lane; i < kMmaN / 4; i += 32) {  // (stem-by-phase: the four accumulators share 64 channels)
                    ptx::sts128f(scale_sa + i * 16, __ldg(reinterpret_cast<const float4*>(p.scale + n_tile0) + i));
                    ptx::sts128f(bias_sa + i * 16, __ldg(reinterpret_cast<const float4*>(p.bias + n_tile0) + i));
                }
                __syncwarp();
                ptx::mbar_wait(&tmem_full_bar[acc_stage], acc_phase);
                ptx::tc_fence_after();
                if ((epi_warp & 3) == 0 && lane == 0) stamp(tile_iter, TR_EPI_START);
#pragma unroll 1
                for (int c = 0; c < kChunks; ++c) {
                    // (a barrier probe takes ~200 cycles to answer even when the load has landed: issued ahead of the
                    // TMEM loads, looked at after them)
                    const bool res_landed = has_res && slab_ok && ptx::mbar_try_wait(&my_res_bar[slot], (res_par >> slot) & 1);
                    uint32_t v0[32], v1[32];
                    ptx::tmem_ld_32x32b_x32(taddr + c * kEpiCols, v0);
                    ptx::tmem_ld_32x32b_x32(taddr + c * kEpiCols + 32, v1);
                    ptx::tmem_ld_wait();
                    if (c == kChunks - 1) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if (PAIR)
                                ptx::mbar_arrive_remote(&tmem_empty_bar[acc_stage], 0);  // the leader's MMA warp waits for both CTAs
                            else
                                ptx::mbar_arrive(&tmem_empty_bar[acc_stage]);
                        }
                    }
                    const uint32_t slab = slabs_sa + slot * kEpiSlabBytes;
                    uint8_t* slab_ptr = my_slabs + slot * kEpiSlabBytes;
                    const uint32_t this_slot = slot;
                    if (++slot == ring) slot = 0;
                    if (!slab_ok) {
                        if (has_res) {
                            // (rare: rows past the end of M) nothing was stored from this chunk, so the newest
                            // store group may still be reading the slab the prefetch is about to refill
                            if (p.epi_tma == 1 && lane == 0) ptx::tma_store_wait_read<0>();
                            __syncwarp();
                            issue_residual();
                        }
                        continue;
                    }
                    if (has_res) {
                        if (!res_landed) ptx::mbar_wait(&my_res_bar[this_slot], (res_par >> this_slot) & 1);
                        res_par ^= (1u << this_slot);
                    } else {
                        // the TMA store issued from this slab `ring` chunks ago must have read it
                        if (p.epi_tma == 1 && lane == 0) epi_store_wait_read(static_cast<int>(ring) - 1);
                        __syncwarp();
                    }
                    const uint32_t sc = scale_sa + (S2DP ? 0 : c * kEpiCols * 4);
                    const uint32_t bi = bias_sa + (S2DP ? 0 : c * kEpiCols * 4);
                    auto drain = [&](auto relu_tag) {
                        constexpr bool kRelu = decltype(relu_tag)::value;
                        // (the shared-memory loads of piece j + 1 -- constants and shortcut -- are issued before the
                        // arithmetic of piece j: they are ordered asm statements, the compiler cannot hoist them)
                        float4 s0 = ptx::lds128f(sc), s1 = ptx::lds128f(sc + 16);
                        float4 b0 = ptx::lds128f(bi), b1 = ptx::lds128f(bi + 16);
                        uint4 rv = make_uint4(0u, 0u, 0u, 0u);
                        if (has_res) rv = ptx::lds128(slab + ((0 ^ (lane & 7)) << 4) + lane * 128);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {  // 8 columns = one 16-byte unit of this thread's row
                            const uint32_t* vv = (j < 4) ? &v0[j * 8] : &v1[(j - 4) * 8];
                            const float4 cs0 = s0, cs1 = s1, cb0 = b0, cb1 = b1;
                            const uint32_t r[4] = {rv.x, rv.y, rv.z, rv.w};
                            const uint32_t sw = ((j ^ (lane & 7)) << 4) + lane * 128;
                            if (j < 7) {
                                s0 = ptx::lds128f(sc + (j + 1) * 32), s1 = ptx::lds128f(sc + (j + 1) * 32 + 16);
                                b0 = ptx::lds128f(bi + (j + 1) * 32), b1 = ptx::lds128f(bi + (j + 1) * 32 + 16);
                                if (has_res) rv = ptx::lds128(slab + (((j + 1) ^ (lane & 7)) << 4) + lane * 128);
                            }
                            float y[8];
                            y[0] = fmaf(__uint_as_float(vv[0]), cs0.x, cb0.x);
                            y[1] = fmaf(__uint_as_float(vv[1]), cs0.y, cb0.y);
                            y[2] = fmaf(__uint_as_float(vv[2]), cs0.z, cb0.z);
                            y[3] = fmaf(__uint_as_float(vv[3]), cs0.w, cb0.w);
                            y[4] = fmaf(__uint_as_float(vv[4]), cs1.x, cb1.x);
                            y[5] = fmaf(__uint_as_float(vv[5]), cs1.y, cb1.y);
                            y[6] = fmaf(__uint_as_float(vv[6]), cs1.z, cb1.z);
                            y[7] = fmaf(__uint_as_float(vv[7]), cs1.w, cb1.w);
                            if (has_res) {
#pragma unroll
                                for (int e = 0; e < 4; ++e) {  // bf16 pair -> two fp32: shift / mask, no permute
                                    y[2 * e] += __uint_as_float(r[e] << 16);
                                    y[2 * e + 1] += __uint_as_float(r[e] & 0xffff0000u);
                                }
                            }
                            uint4 ov;  // (ReLU rides on the fp32 -> bf16x2 conversion)
                            ov.x = kRelu ? ptx::pack_bf16x2_relu(y[0], y[1]) : ptx::pack_bf16x2(y[0], y[1]);
                            ov.y = kRelu ? ptx::pack_bf16x2_relu(y[2], y[3]) : ptx::pack_bf16x2(y[2], y[3]);
                            ov.z = kRelu ? ptx::pack_bf16x2_relu(y[4], y[5]) : ptx::pack_bf16x2(y[4], y[5]);
                            ov.w = kRelu ? ptx::pack_bf16x2_relu(y[6], y[7]) : ptx::pack_bf16x2(y[6], y[7]);
                            ptx::sts128(slab + sw, ov);
                        }
                    };
                    if (p.relu)
                        drain(std::true_type{});
                    else
                        drain(std::false_type{});
                    if (p.epi_tma == 1) ptx::fence_proxy_async_smem();  // slab writes -> visible to the TMA store
                    __syncwarp();
                    if (p.epi_tma == 1) {
                        if (lane == 0) {
                            if constexpr (S2DP) {
                                // rows of exactly 32 units: this warp's slab is one image row of phase c; the unit
                                // columns past Wo and the walk's rows past Ho fall outside the tensor and are clipped
                                const int walk_row = m_row0 >> 5;
                                const int img = walk_row / p.H;
                                ptx::tma_store_5d(&tmap_out, slab_ptr, 0, c, 0, walk_row - img * p.H, img);
                            } else {
                                ptx::tma_store_2d(&tmap_out, slab_ptr, n_tile0 + c * kEpiCols, m_row0);
                            }
                            ptx::tma_store_commit();
                        }
                    } else {
                        // coalesced write-out: 8 lanes cover one 128-byte row segment, 4 rows per instruction
                        const int ch = lane & 7;
                        if constexpr (S2DP) {
                            // chunk c is column phase c: unit position m -> (image, row, unit column j) -> pixel 4j + c.
                            // Rows past Ho and the unit column past Wo exist only in the walk.
                            __nv_bfloat16* gcol = p.out_bf16 + ch * 8;
                            int m = m_row0 + (lane >> 3);
                            int rest = m / p.row_pitch;
                            int j = m - rest * p.row_pitch;
                            int img = rest / p.H;
                            int prow = rest - img * p.H;
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int row = i * 4 + (lane >> 3);
                                const uint4 val = ptx::lds128(slab + row * 128 + ((ch ^ (row & 7)) << 4));
                                const int q = 4 * j + c;
                                if (m < p.M && prow < p.Ho && q < p.Wo)
                                    *reinterpret_cast<uint4*>(
                                        gcol + ((static_cast<size_t>(img) * p.Ho + prow) * p.Wo + q) * p.N) = val;
                                m += 4;
                                j += 4;
                                while (j >= p.row_pitch) {  // (rows narrower than four units wrap more than once)
                                    j -= p.row_pitch;
                                    if (++prow == p.H) {
                                        prow = 0;
                                        ++img;
                                    }
                                }
                            }
                        } else if constexpr (ROW3) {
                            // padded-flat position -> output pixel; the two overhang positions of every row are dropped
                            __nv_bfloat16* gcol = p.out_bf16 + n_tile0 + c * kEpiCols + ch * 8;
                            int m = m_row0 + (lane >> 3);
                            int rest = m / p.row_pitch;
                            int q = m - rest * p.row_pitch;
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int row = i * 4 + (lane >> 3);
                                const uint4 val = ptx::lds128(slab + row * 128 + ((ch ^ (row & 7)) << 4));
                                if (m < p.M && q < p.row_valid)
                                    *reinterpret_cast<uint4*>(gcol + (static_cast<size_t>(rest) * p.row_valid + q) * p.N) = val;
                                m += 4;
                                q += 4;
                                if (q >= p.row_pitch) {
                                    q -= p.row_pitch;
                                    ++rest;
                                }
                            }
                        } else {
                            __nv_bfloat16* gout = p.out_bf16 + static_cast<size_t>(m_row0) * p.N + n_tile0 + c * kEpiCols;
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int row = i * 4 + (lane >> 3);
                                const uint4 val = ptx::lds128(slab + row * 128 + ((ch ^ (row & 7)) << 4));
                                if (m_row0 + row < p.M)
                                    *reinterpret_cast<uint4*>(gout + static_cast<size_t>(row) * p.N + ch * 8) = val;
                            }
                        }
                        __syncwarp();
                    }
                    if (has_res) {
                        // every store but the one just issued has read its slab: the previous chunk's slab is free
                        // for the residual `ring - 1` chunks ahead
                        if (p.epi_tma == 1 && lane == 0) ptx::tma_store_wait_read<1>();
                        __syncwarp();
                        issue_residual();
                    }
                }
                if ((epi_warp & 3) == 0 && lane == 0) stamp(tile_iter, TR_EPI_DONE);
                continue;
            }
            // ---------- direct epilogue (fp32 logits; debugging switch) ----------
            ptx::mbar_wait(&tmem_full_bar[acc_stage], acc_phase);
            ptx::tc_fence_after();
            const int m = m_blk * kBlockM + lane_group * 32 + lane;
            const bool row_ok = m < p.M;
#pragma unroll 1
            for (int chunk = 0; chunk < BLOCK_N / 32; ++chunk) {
                uint32_t v[32];
                ptx::tmem_ld_32x32b_x32(taddr + chunk * 32, v);
                ptx::tmem_ld_wait();
                if (chunk == BLOCK_N / 32 - 1) {
                    // accumulator fully drained into registers: hand the TMEM buffer back to the MMA warp
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive(&tmem_empty_bar[acc_stage]);
                }
                const int n0 = n_blk * BLOCK_N + chunk * 32;
                if (n0 >= p.N) continue;
                const size_t row_off = static_cast<size_t>(m) * p.N + n0;
#pragma unroll
                for (int g = 0; g < 4; ++g) {  // 8 columns per group
                    const int n = n0 + g * 8;
                    if (n >= p.N) break;  // N is a multiple of 8
                    const float4 s0 = __ldg(reinterpret_cast<const float4*>(p.scale + n));
                    const float4 s1 = __ldg(reinterpret_cast<const float4*>(p.scale + n + 4));
                    const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + n));
                    const float4 b1 = __ldg(reinterpret_cast<const float4*>(p.bias + n + 4));
                    float y[8];
                    y[0] = fmaf(__uint_as_float(v[g * 8 + 0]), s0.x, b0.x);
                    y[1] = fmaf(__uint_as_float(v[g * 8 + 1]), s0.y, b0.y);
                    y[2] = fmaf(__uint_as_float(v[g * 8 + 2]), s0.z, b0.z);
                    y[3] = fmaf(__uint_as_float(v[g * 8 + 3]), s0.w, b0.w);
                    y[4] = fmaf(__uint_as_float(v[g * 8 + 4]), s1.x, b1.x);
                    y[5] = fmaf(__uint_as_float(v[g * 8 + 5]), s1.y, b1.y);
                    y[6] = fmaf(__uint_as_float(v[g * 8 + 6]), s1.z, b1.z);
                    y[7] = fmaf(__uint_as_float(v[g * 8 + 7]), s1.w, b1.w);
                    if (!row_ok) continue;
                    if (p.residual != nullptr) {
                        const uint4 rv = *reinterpret_cast<const uint4*>(p.residual + row_off + g * 8);
                        const __nv_bfloat162* r2 = reinterpret_cast<const __nv_bfloat162*>(&rv);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            const float2 f = __bfloat1622float2(r2[j]);
                            y[2 * j] += f.x;
                            y[2 * j + 1] += f.y;
                        }
                    }
                    if (p.relu) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) y[j] = fmaxf(y[j], 0.0f);
                    }
                    if (p.out_f32 != nullptr) {
                        float4* o = reinterpret_cast<float4*>(p.out_f32 + row_off + g * 8);
                        o[0] = make_float4(y[0], y[1], y[2], y[3]);
                        o[1] = make_float4(y[4], y[5], y[6], y[7]);
                    }
                    if (p.out_bf16 != nullptr) {
                        uint4 ov;
                        __nv_bfloat162* o2 = reinterpret_cast<__nv_bfloat162*>(&ov);
#pragma unroll
                        for (int j = 0; j < 4; ++j) o2[j] = __floats2bfloat162_rn(y[2 * j], y[2 * j + 1]);
                        *reinterpret_cast<uint4*>(p.out_bf16 + row_off + g * 8) = ov;
                    }
                }
            }
        }
        // all of this warp's bulk stores must have landed before the CTA may exit
        if (p.epi_tma == 1 && lane == 0) ptx::tma_store_wait<0>();
    } else {
        // ===================== software im2col producers (A_GATHER) =====================
        if (A_MODE == A_GATHER) {
            const int row = threadIdx.x - 320;  // 0..127 : one A-tile row per thread
            const int hw = p.Ho * p.Wo;
            int stage = 0;
            uint32_t phase = 0;
            int m_blk = 0, n_blk_g = 0;
            bool m_valid_g = true;
            for (long long ti = 0; tile_at(ti, m_blk, n_blk_g, m_valid_g); ++ti) {
                const int m = m_blk * kBlockM + row;
                const bool row_ok = m < p.M;
                int img = 0, hi0 = 0, wi0 = 0;
                if (row_ok) {
                    img = m / hw;
                    const int rem = m - img * hw;
                    const int po = rem / p.Wo;
                    const int qo = rem - po * p.Wo;
                    hi0 = po * p.stride - p.pad;
                    wi0 = qo * p.stride - p.pad;
                }
                for (int kb = 0; kb < p.num_k_blocks; ++kb) {
                    uint4 vals[8];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const int k = kb * kBlockK + j * 8;
                        uint4 val = make_uint4(0u, 0u, 0u, 0u);
                        if (row_ok && k < p.K_real) {
                            const int tap = k / p.Cin;
                            const int c = k - tap * p.Cin;
                            const int r = tap / p.kw;
                            const int s = tap - r * p.kw;
                            const int hi = hi0 + r;
                            const int wi = wi0 + s;
                            if (hi >= 0 && hi < p.H && wi >= 0 && wi < p.W) {
                                val = __ldg(reinterpret_cast<const uint4*>(
                                    p.in + (static_cast<size_t>(img * p.H + hi) * p.W + wi) * p.Cin + c));
                            }
                        }
                        vals[j] = val;
                    }
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* dst = smem_a + stage * kABytes + row * 128;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        *reinterpret_cast<uint4*>(dst + ((j ^ (row & 7)) << 4)) = vals[j];
                    }
                    ptx::fence_proxy_async_smem();
                    ptx::mbar_arrive(&full_bar[stage]);
                    if (++stage == kStages) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    }

    // ===================== teardown =====================
    ptx::tc_fence_before();
    __syncthreads();
    if (CL > 1) ptx::cluster_sync();  // no CTA leaves while a peer may still multicast into it / arrive on its barriers
    if (warp_idx == 1) {
        ptx::tc_fence_after();
        if (PAIR)
            ptx::tmem_dealloc_pair(tmem_base, kTmemCols);
        else
            ptx::tmem_dealloc(tmem_base, kTmemCols);
    }
}

}  // namespace adapt
