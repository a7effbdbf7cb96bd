// Bottleneck-block chain: `_2_conv` (3x3 / stride 1 / pad 1, C -> C, +BN +ReLU) and `_3_conv` (1x1, C -> 4C, +BN
// +shortcut Add +ReLU) of one Keras residual block in ONE kernel.
//
// Replaces, inside a stage's `model.predict` (reference src/node.py:177), the two launches
//     y2  = relu(conv3x3(x) * s2 + b2)                 (written to HBM, read back)
//     out = relu(conv1x1(y2) * s3 + b3 + shortcut)
// by one: the 3x3's 128 x C accumulator tile, after scale / bias / ReLU and rounding to bf16 (the SAME rounding
// point as the two-launch form), is exactly a 128 x K A-operand tile of the following 1x1 (point-wise: no halo), so
// the epilogue warps write it into a swizzled shared-memory A tile and the MMA warp issues the second GEMM against
// `_3_conv`'s weights.  The `_2_relu` tensor never exists in HBM (conv2: 103 MB of traffic per block at batch 128,
// conv3: 51 MB) and a launch per block disappears.
//
// Phase 1 is the row-reuse form of conv_gemm_sm100.cuh (A_ROW3): the output is walked in "padded-flat" order (rows
// of W + 2 positions, the last two of each row discarded), one im2col load of 136 positions per (filter row, channel
// block) feeds the three horizontal taps through shifted descriptors.  Phase 2 therefore also runs on padded-flat
// positions; its residual tile is fetched with an im2col-mode TMA load that walks the same positions (zero-filled
// where a position is not a pixel) and its output leaves through a coalesced st.global write-out whose position ->
// pixel offsets are worked out once per unit and warp, not per chunk.  (TMA stores were tried for this and fault on
// B200 -- "illegal instruction" -- both as tiled stores with a negative start column, the way to express a slab that
// straddles image rows, and as im2col-mode stores over the padded-flat walk.)
//
//   warp 0      TMA producer: the phase-1 stages of a unit (A row segments; the three B tiles of a filter row unless
//               the 3x3's weights are resident) and one stage per phase-2 n-tile (`_3_conv` weights), all through
//               ONE operand ring in the order the MMA warp consumes them
//   warp 1      MMA issuer (phase 1 into acc1, phase 2 per 128-column n-tile into acc2[n & 1]); in the peer CTA of a
//               pair it forwards that CTA's "A2 half written" to the leader (one cluster-scope release instead of
//               one per epilogue warp, each of which would wait for that warp's outstanding global stores)
//   warps 2..9  epilogue, two groups of four.  Both groups turn acc1 into the A2 tile (32-column units alternate
//               between the groups); group g then drains acc2[g], i.e. n-tiles g, g+2, ...
//
// Order on the tensor pipe:  M1(0) | N[0,1](u)  M1(u+1)  N[2..](u) | ...   (M1 = 3x3 of a unit, N[k] = n-tile k of its
// 1x1): the next unit's 3x3 runs while the first two n-tiles are drained.  The epilogue warps build the A2 tile of
// unit u+1 BEFORE their last chunk of unit u, so that the first n-tiles of u+1 are computed while that chunk is
// written out and no epilogue warp idles on the tensor pipe between units.
//
//   C = 64  (conv2): single CTA, N1 = 64; the 3x3's weights (72 KiB) stay resident in shared memory
//   C = 128 (conv3): CTA pair (tcgen05 cta_group::2, 256 positions per unit), all weights streamed
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cuda.h>
#include <cstdint>
#include <type_traits>

#include "conv_gemm_sm100.cuh"
#include "ptx_sm100.cuh"

namespace adapt {

struct ConvChainParams {
    int M;          // padded-flat positions: batch * H * (W + 2)
    int m_tiles;    // ceil(M / 128)
    int H, W;       // spatial size of input, intermediate and output
    int row_pitch;  // W + 2
    int N2;         // output channels = 4 * C
    const float* scale1;  // [C]   folded BN of the 3x3
    const float* bias1;
    const float* scale2;  // [N2]  folded BN of the 1x1
    const float* bias2;
    __nv_bfloat16* out;   // [batch * H * W, N2]
    int stages;           // phase-1 operand ring depth
    int ring;             // residual slabs per epilogue warp (2..4)
    int l2_prefetch;      // 1: every epilogue warp pulls its residual rows of the NEXT unit towards L2 while it works
                          // on this one (HBM latency is then hidden by L2, not by the depth of the slab ring)
    long long* trace;     // developer timeline (nullptr in production): [grid][kTraceTiles][kChainTraceEvents] SM clocks
};

// Timeline slots of the chain kernel (developer build, tools/chain_trace.py)
constexpr int kChainTraceEvents = 21;
enum : int {
    CT_CTA_START = 0,    // (unit 0 only) kernel entry
    CT_PROD_FIRST = 1,   // producer: first phase-1 load of the unit issued
    CT_PROD_LAST = 2,    // producer: last phase-2 weight tile of the unit issued
    CT_MMA1_FIRST = 3,   // MMA warp: first operand stage of the unit arrived
    CT_MMA1_DONE = 4,    // MMA warp: acc1 commit issued
    CT_A2_READY = 5,     // MMA warp: A2 tile complete (every epilogue warp arrived)
    CT_MMA2_DONE = 6,    // MMA warp: last n-tile's commit issued
    CT_EPI1_START = 7,   // epilogue warp 0: acc1 ready
    CT_EPI1_DONE = 8,    // epilogue warp 0: its part of A2 written and announced
    CT_EPI2_FIRST = 9,   // epilogue warp 0: first acc2 of the unit ready
    CT_EPI2_DONE = 10,   // epilogue warp 0: last chunk of the unit written out
    CT_EPI2_G1_DONE = 11,  // epilogue warp 4 (group 1): last chunk of the unit written out
    // inside epilogue warp 0's first chunk of the unit
    CT_C0_START = 12,    // before the TMEM loads
    CT_C0_TMEM = 13,     // accumulator chunk in registers
    CT_C0_RES = 14,      // residual slab arrived
    CT_C0_MATH = 15,     // scale / bias / add / ReLU / pack done, slab rewritten
    CT_C0_OUT = 16,      // write-out instructions issued
    CT_C0_END = 17,      // next residual load issued
    CT_C1_END = 18,      // second chunk finished
    CT_C0_SYNC = 19,     // warp re-converged after the write-out
    CT_C0_EXPECT = 20,   // expect-tx of the next residual load posted
};

constexpr int kChainBN2 = 128;      // phase-2 n-tile width
constexpr int kChainThreads = 320;  // TMA + MMA + 8 epilogue warps
constexpr int kChainMaxStages = 6;
constexpr int kChainMaxRing = 4;

template <int C, bool PAIR>
__host__ __device__ constexpr int conv_chain_smem_bytes(int stages, int ring) {
    constexpr int b1 = (PAIR ? C / 2 : C) * 128;  // one k-block of the 3x3's weights held by this CTA
    constexpr bool resident = !PAIR;
    constexpr int c_blocks = C / 64;
    // (a phase-2 stage -- one n-tile of the 1x1's weights, 16 KiB per CTA -- lives in the A slot of a ring stage)
    const int ring1 = stages * kRow3ABytes + (resident ? 9 * c_blocks * b1 : stages * 3 * b1);
    const int a2 = c_blocks * kABytes;
    const int slabs = kEpiWarps * ring * kEpiSlabBytes;
    const int sb = 2 * (C + 4 * C) * 4;
    return ring1 + a2 + slabs + sb + 1024 /*barriers*/ + 1024 /*alignment*/;
}

template <int C, bool PAIR>
__global__ void __launch_bounds__(kChainThreads, 1)
conv_chain_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_w2,
                  const __grid_constant__ CUtensorMap tmap_w3, const __grid_constant__ CUtensorMap tmap_res,
                  const ConvChainParams p) {
    static_assert(C == 64 || C == 128, "chain kernel: C = 64 (single CTA) or 128 (CTA pair)");
    static_assert(PAIR == (C == 128), "C = 128 runs on CTA pairs, C = 64 on single CTAs");
    constexpr int CL = PAIR ? 2 : 1;
    constexpr bool RESIDENT = !PAIR;             // the 3x3's weights stay in shared memory
    constexpr int c_blocks = C / 64;             // k-blocks per tap (phase 1) = k-blocks of phase 2
    constexpr int k_iters = 3 * c_blocks;        // phase-1 pipeline steps per tile (one per filter row and channel block)
    constexpr int n2_tiles = 4 * C / kChainBN2;  // 2 or 4 (even: n-tile n uses accumulator n & 1)
    constexpr int kB1 = (PAIR ? C / 2 : C) * 128;
    constexpr int kB2 = (PAIR ? kChainBN2 / 2 : kChainBN2) * 128;
    constexpr int kW3Tile = c_blocks * kB2;      // one phase-2 n-tile of weights held by this CTA (16 KiB)
    static_assert(kW3Tile <= kRow3ABytes, "a phase-2 weight tile must fit the A slot of a ring stage");
    constexpr uint32_t kAcc2Col = 256;           // TMEM: acc1 at columns [0, C), acc2[b] at 256 + 128 b
    constexpr uint16_t kClusterMask = static_cast<uint16_t>((1u << CL) - 1);
    const int kStages = p.stages;
    const uint32_t ring = static_cast<uint32_t>(p.ring);

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b1 = smem_a + kStages * kRow3ABytes;
    uint8_t* smem_a2 = smem_b1 + (RESIDENT ? 9 * c_blocks * kB1 : kStages * 3 * kB1);
    uint8_t* smem_epi = smem_a2 + c_blocks * kABytes;
    float* smem_sb = reinterpret_cast<float*>(smem_epi + kEpiWarps * p.ring * kEpiSlabBytes);  // scale1 bias1 scale2 bias2
    uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(smem_sb) + 2 * (C + 4 * C) * 4);
    uint64_t* full_bar = bars;                               // [kChainMaxStages]
    uint64_t* empty_bar = bars + kChainMaxStages;            // [kChainMaxStages]
    uint64_t* acc1_full = bars + 2 * kChainMaxStages;        // [1]
    uint64_t* a2_full = acc1_full + 1;                       // [1]
    uint64_t* acc2_full = a2_full + 1;                       // [2]
    uint64_t* acc2_empty = acc2_full + 2;                    // [2]
    uint64_t* w_full = acc2_empty + 2;                       // [1] resident weights
    uint64_t* a2_free = w_full + 1;                          // [1] the unit's 1x1 has read the A2 tile (MMA commit)
    uint64_t* a2_local = a2_free + 1;                        // [1] pair, peer CTA: its epilogue warps' A2 arrivals
    uint64_t* res_bar = a2_local + 1;                        // [kEpiWarps][kChainMaxRing]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(res_bar + kEpiWarps * kChainMaxRing);
    constexpr int kNumBars = 2 * kChainMaxStages + 1 + 1 + 2 + 2 + 1 + 2 + kEpiWarps * kChainMaxRing;

    const int warp_idx = __shfl_sync(0xffffffff, threadIdx.x / 32, 0);
    const int lane = threadIdx.x % 32;
    const int cl_rank = PAIR ? static_cast<int>(ptx::cluster_ctarank()) : 0;
    const int num_units = (p.m_tiles + CL - 1) / CL;
    const int unit0 = blockIdx.x / CL;
    const int unit_step = gridDim.x / CL;
    auto tile_at = [&](int i, int& m_blk, bool& m_valid) -> bool {
        const int u = unit0 + i * unit_step;
        if (u >= num_units) return false;
        m_blk = u * CL + cl_rank;
        m_valid = m_blk < p.m_tiles;  // (a pair's last unit may hang over the end of M)
        return true;
    };

    auto stamp = [&](int unit, int ev) {
#if ADAPT_ENABLE_TRACE
        if (p.trace != nullptr && unit < kTraceTiles)
            p.trace[(static_cast<long long>(blockIdx.x) * kTraceTiles + unit) * kChainTraceEvents + ev] = clock64();
#else
        (void)unit;
        (void)ev;
#endif
    };
    if (threadIdx.x == 0) stamp(0, CT_CTA_START);

    if (warp_idx == 0) {
        for (int i = lane; i < kNumBars; i += 32) {
            uint32_t count = 1;
            if (&bars[i] == a2_full) count = kEpiWarps + (PAIR ? 1 : 0);         // own epilogue warps (+ the peer's forwarder)
            if (&bars[i] == a2_local) count = kEpiWarps;
            if (&bars[i] == acc2_empty || &bars[i] == acc2_empty + 1) count = 4 * CL;  // the draining group (of both CTAs)
            ptx::mbar_init(&bars[i], count);
        }
        ptx::fence_mbar_init();
    } else if (warp_idx == 2 && lane == 0) {
        ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_w2);
        ptx::prefetch_tmap(&tmap_w3);
        ptx::prefetch_tmap(&tmap_res);
    }
    if (warp_idx == 1) {
        if (PAIR) {
            ptx::tmem_alloc_pair(tmem_slot, 512);
            ptx::tmem_relinquish_pair();
        } else {
            ptx::tmem_alloc(tmem_slot, 512);
            ptx::tmem_relinquish();
        }
    }
    // folded BatchNorm constants of both layers -> shared memory, once per CTA (weights: nothing upstream writes them)
    for (int i = threadIdx.x; i < C; i += blockDim.x) {
        smem_sb[i] = __ldg(p.scale1 + i);
        smem_sb[C + i] = __ldg(p.bias1 + i);
    }
    for (int i = threadIdx.x; i < 4 * C; i += blockDim.x) {
        smem_sb[2 * C + i] = __ldg(p.scale2 + i);
        smem_sb[2 * C + 4 * C + i] = __ldg(p.bias2 + i);
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (CL > 1) ptx::cluster_sync();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    ptx::griddep_launch_dependents();
    if (RESIDENT && warp_idx == 0 && ptx::elect_one()) {
        // the 3x3's weights, once: no kernel writes them, so this runs ahead of the dependency wait
        ptx::mbar_expect_tx(w_full, 9 * c_blocks * kB1);
        for (int kb = 0; kb < 9 * c_blocks; ++kb) ptx::tma_load_2d(smem_b1 + kb * kB1, &tmap_w2, w_full, kb * kBlockK, 0);
    }
    ptx::griddep_wait();

    // Schedule (u = unit): the tensor pipe runs  M1(0) | N[0,1](u) M1(u+1) N[2..](u) | ...  -- the next unit's 3x3 sits
    // between the early and the late n-tiles of this unit's 1x1, so it overlaps the drain of acc2 -- and the epilogue
    // warps turn acc1(u+1) into the A2 tile BEFORE their last chunk of unit u ("splice"): by the time that chunk is
    // written out, the first n-tiles of unit u+1 are in acc2 and no epilogue warp waits for the tensor pipe.
    constexpr int n2_early = 2;  // n-tiles issued before the next unit's phase 1 (one per accumulator buffer)

    if (warp_idx == 0) {
        // ===================== producer: operand stages in the order the MMA warp consumes them =====================
        int stage = 0;
        uint32_t phase = 0;
        auto advance = [&]() {
            if (++stage == kStages) {
                stage = 0;
                phase ^= 1;
            }
        };
        auto load_phase1 = [&](int ti, int m_blk_raw, bool m_valid) {
            const int m_blk = m_valid ? m_blk_raw : p.m_tiles - 1;
            const int m0 = m_blk * kBlockM;
            const int rest = m0 / p.row_pitch;
            const int q0 = m0 - rest * p.row_pitch;
            const int img0 = rest / p.H;
            const int p0 = rest - img0 * p.H;
            int tap_r = 0, cb = 0;
            for (int it = 0; it < k_iters; ++it) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                if (ptx::elect_one()) {
                    if (it == 0) stamp(ti, CT_PROD_FIRST);
                    uint8_t* a_dst = smem_a + stage * kRow3ABytes;
                    if (PAIR) {
                        if (cl_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2 * (kRow3ABytes + 3 * kB1));
                        ptx::tma_load_im2col_4d_pair(a_dst, &tmap_a, &full_bar[stage], cb * kBlockK, q0 - 1, p0 - 1, img0, 0,
                                                     static_cast<uint16_t>(tap_r));
#pragma unroll
                        for (int s3 = 0; s3 < 3; ++s3) {
                            const int kbi = (tap_r * 3 + s3) * c_blocks + cb;
                            ptx::tma_load_2d_pair(smem_b1 + (stage * 3 + s3) * kB1, &tmap_w2, &full_bar[stage], kbi * kBlockK,
                                                  cl_rank * (C / 2));
                        }
                    } else {
                        ptx::mbar_expect_tx(&full_bar[stage], kRow3ABytes);
                        ptx::tma_load_im2col_4d(a_dst, &tmap_a, &full_bar[stage], cb * kBlockK, q0 - 1, p0 - 1, img0, 0,
                                                static_cast<uint16_t>(tap_r));
                    }
                }
                __syncwarp();
                if (++cb == c_blocks) {
                    cb = 0;
                    ++tap_r;
                }
                advance();
            }
        };
        // one n-tile of the 1x1's weights, in the A slot of a ring stage
        auto load_w3 = [&](int ti, int n2) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            if (ptx::elect_one()) {
                if (n2 == n2_tiles - 1) stamp(ti, CT_PROD_LAST);
                uint8_t* w_dst = smem_a + stage * kRow3ABytes;
                if (PAIR) {
                    if (cl_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], 2 * kW3Tile);
#pragma unroll
                    for (int kb = 0; kb < c_blocks; ++kb)
                        ptx::tma_load_2d_pair(w_dst + kb * kB2, &tmap_w3, &full_bar[stage], kb * kBlockK,
                                              n2 * kChainBN2 + cl_rank * (kChainBN2 / 2));
                } else {
                    ptx::mbar_expect_tx(&full_bar[stage], kW3Tile);
#pragma unroll
                    for (int kb = 0; kb < c_blocks; ++kb)
                        ptx::tma_load_2d(w_dst + kb * kB2, &tmap_w3, &full_bar[stage], kb * kBlockK, n2 * kChainBN2);
                }
            }
            __syncwarp();
            advance();
        };
        int m_blk = 0, m_next = 0;
        bool m_valid = true, v_next = true;
        bool have = tile_at(0, m_blk, m_valid);
        if (have) load_phase1(0, m_blk, m_valid);
        for (int ti = 0; have; ++ti) {
            const bool more = tile_at(ti + 1, m_next, v_next);
            for (int n2 = 0; n2 < n2_early; ++n2) load_w3(ti, n2);
            if (more) load_phase1(ti + 1, m_next, v_next);
            for (int n2 = n2_early; n2 < n2_tiles; ++n2) load_w3(ti, n2);
            have = more;
        }
    } else if (warp_idx == 1) {
        if (!(PAIR && cl_rank != 0)) {
            // ===================== MMA issuer =====================
            constexpr uint32_t idesc1 = ptx::make_idesc_bf16_f32(PAIR ? 2 * kBlockM : kBlockM, C);
            constexpr uint32_t idesc2 = ptx::make_idesc_bf16_f32(PAIR ? 2 * kBlockM : kBlockM, kChainBN2);
            int stage = 0;
            uint32_t phase = 0;
            auto advance = [&]() {
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            };
            // acc1 = im2col(x) . W2^T.  acc1 is free: the A2 tile of the previous unit was complete, i.e. acc1 fully
            // read, before that unit's phase 2 was issued.
            auto mma_phase1 = [&](int ti) {
                ptx::tc_fence_after();
                int tap_r = 0, cb = 0;
                for (int it = 0; it < k_iters; ++it) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    if (it == 0 && lane == 0) stamp(ti, CT_MMA1_FIRST);
                    if (ptx::elect_one()) {
                        const uint32_t a_base = ptx::smem_u32(smem_a + stage * kRow3ABytes);
#pragma unroll
                        for (int s3 = 0; s3 < 3; ++s3) {
                            const int kbi = (tap_r * 3 + s3) * c_blocks + cb;
                            const uint64_t a_desc = ptx::make_kmajor_sw128_desc(a_base + s3 * (kBlockK * 2));
                            const uint64_t b_desc =
                                ptx::make_kmajor_sw128_desc(ptx::smem_u32(smem_b1 + (RESIDENT ? kbi : stage * 3 + s3) * kB1));
#pragma unroll
                            for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                                if (PAIR)
                                    ptx::umma_bf16_ss_pair(tmem_base, a_desc + 2 * k, b_desc + 2 * k, idesc1, (it | s3 | k) != 0 ? 1u : 0u);
                                else
                                    ptx::umma_bf16_ss(tmem_base, a_desc + 2 * k, b_desc + 2 * k, idesc1, (it | s3 | k) != 0 ? 1u : 0u);
                            }
                        }
                        if (PAIR) {
                            ptx::umma_commit_pair(&empty_bar[stage], kClusterMask);
                            if (it == k_iters - 1) ptx::umma_commit_pair(acc1_full, kClusterMask);
                        } else {
                            ptx::umma_commit(&empty_bar[stage]);
                            if (it == k_iters - 1) ptx::umma_commit(acc1_full);
                        }
                        if (it == k_iters - 1) stamp(ti, CT_MMA1_DONE);
                    }
                    __syncwarp();
                    if (++cb == c_blocks) {
                        cb = 0;
                        ++tap_r;
                    }
                    advance();
                }
            };
            // acc2[j & 1] = A2 . W3[n2]^T (j counts n-tiles over the whole kernel)
            uint32_t j = 0;
            auto mma_ntile = [&](int ti, int n2) {
                const uint32_t buf = j & 1, use = j >> 1;
                ++j;
                ptx::mbar_wait(&acc2_empty[buf], (use & 1) ^ 1);
                ptx::mbar_wait(&full_bar[stage], phase);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t tmem_d = tmem_base + kAcc2Col + buf * kChainBN2;
                    const uint32_t w3_base = ptx::smem_u32(smem_a + stage * kRow3ABytes);
#pragma unroll
                    for (int kb = 0; kb < c_blocks; ++kb) {
                        const uint64_t a_desc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(smem_a2 + kb * kABytes));
                        const uint64_t b_desc = ptx::make_kmajor_sw128_desc(w3_base + kb * kB2);
#pragma unroll
                        for (int k = 0; k < kBlockK / kUmmaK; ++k) {
                            if (PAIR)
                                ptx::umma_bf16_ss_pair(tmem_d, a_desc + 2 * k, b_desc + 2 * k, idesc2, (kb | k) != 0 ? 1u : 0u);
                            else
                                ptx::umma_bf16_ss(tmem_d, a_desc + 2 * k, b_desc + 2 * k, idesc2, (kb | k) != 0 ? 1u : 0u);
                        }
                    }
                    if (PAIR) {
                        ptx::umma_commit_pair(&empty_bar[stage], kClusterMask);
                        ptx::umma_commit_pair(&acc2_full[buf], kClusterMask);
                        if (n2 == n2_tiles - 1) ptx::umma_commit_pair(a2_free, kClusterMask);  // A2 of this unit fully read
                    } else {
                        ptx::umma_commit(&empty_bar[stage]);
                        ptx::umma_commit(&acc2_full[buf]);
                        if (n2 == n2_tiles - 1) ptx::umma_commit(a2_free);
                    }
                    if (n2 == n2_tiles - 1) stamp(ti, CT_MMA2_DONE);
                }
                __syncwarp();
                advance();
            };
            int m_blk = 0;
            bool m_valid = true;
            if (RESIDENT) ptx::mbar_wait(w_full, 0);
            bool have = tile_at(0, m_blk, m_valid);
            if (have) mma_phase1(0);
            for (int ti = 0; have; ++ti) {
                const bool more = tile_at(ti + 1, m_blk, m_valid);
                if (PAIR)
                    ptx::mbar_wait_cluster(a2_full, static_cast<uint32_t>(ti) & 1);  // (the peer CTA's arrival is forwarded)
                else
                    ptx::mbar_wait(a2_full, static_cast<uint32_t>(ti) & 1);
                ptx::tc_fence_after();
                if (lane == 0) stamp(ti, CT_A2_READY);
                for (int n2 = 0; n2 < n2_early; ++n2) mma_ntile(ti, n2);
                if (more) mma_phase1(ti + 1);
                for (int n2 = n2_early; n2 < n2_tiles; ++n2) mma_ntile(ti, n2);
                have = more;
            }
        } else {
            // ===================== peer CTA of a pair: forward "my half of the A2 tile is written" =====================
            // The eight epilogue warps of this CTA arrive on a LOCAL barrier (release at CTA scope: cheap); this
            // otherwise idle warp acquires that and performs the ONE cluster-scope release towards the leader's
            // barrier.  (A release at cluster scope from every epilogue warp costs each of them a MEMBAR.GPU that
            // waits for the global stores of their previous chunk.)
            int m_blk = 0;
            bool m_valid = true;
            for (int ti = 0; tile_at(ti, m_blk, m_valid); ++ti) {
                ptx::mbar_wait(a2_local, static_cast<uint32_t>(ti) & 1);
                if (lane == 0) ptx::mbar_arrive_remote_release(a2_full, 0);
                __syncwarp();
            }
        }
    } else if (warp_idx < 2 + kEpiWarps) {
        // ===================== epilogue =====================
        const int epi_warp = warp_idx - 2;
        const int group = epi_warp >> 2;
        const int lane_group = warp_idx & 3;  // TMEM lanes this warp may touch: [32 * lane_group, +32)
        uint8_t* my_slabs = smem_epi + epi_warp * p.ring * kEpiSlabBytes;
        const uint32_t slabs_sa = ptx::smem_u32(my_slabs);
        const uint32_t sb_sa = ptx::smem_u32(smem_sb);
        const uint32_t a2_sa = ptx::smem_u32(smem_a2);
        uint64_t* my_res_bar = res_bar + epi_warp * kChainMaxRing;
        uint32_t res_par = 0, slot = 0;
        const uint32_t taddr_lane = tmem_base + (static_cast<uint32_t>(lane_group * 32) << 16);

        // residual prefetch cursor: this warp's chunks in consumption order (tile, n-tile = group, group + 2, .., chunk)
        int ri_ti = 0, ri_n2 = group, ri_c = 0, ri_q = 0, ri_p = 0, ri_img = 0;
        uint32_t ri_slot = 0;
        bool ri_live = false, ri_more = true;
        auto ri_locate = [&]() {
            int mb = 0;
            bool mv = true;
            ri_live = false;
            ri_more = tile_at(ri_ti, mb, mv);
            if (ri_more) {
                const int m0 = mb * kBlockM + lane_group * 32;
                const int rest = m0 / p.row_pitch;
                ri_q = m0 - rest * p.row_pitch;
                ri_img = rest / p.H;
                ri_p = rest - ri_img * p.H;
                ri_live = mv && m0 < p.M;
            }
        };
        auto issue_residual = [&](int st_unit) {
            if (ri_live && lane == 0) {
                ptx::mbar_expect_tx(&my_res_bar[ri_slot], kEpiSlabBytes);
                if (st_unit >= 0) stamp(st_unit, CT_C0_EXPECT);
                ptx::tma_load_im2col_4d(my_slabs + ri_slot * kEpiSlabBytes, &tmap_res, &my_res_bar[ri_slot],
                                        ri_n2 * kChainBN2 + ri_c * kEpiCols, ri_q, ri_p, ri_img, 0, 0);
            }
            if (++ri_slot == ring) ri_slot = 0;
            if (++ri_c == kChainBN2 / kEpiCols) {
                ri_c = 0;
                ri_n2 += 2;
                if (ri_n2 >= n2_tiles) {
                    ri_n2 = group;
                    ++ri_ti;
                    ri_locate();
                }
            }
        };
        ri_locate();
        // prime every slot: a chunk's write-out is synchronous (st.global from the slab), so at the end of chunk c its
        // slot is free and takes the load of chunk c + ring
        for (uint32_t i = 0; i < ring; ++i) issue_residual(-1);

        // ---- epilogue 1 of unit `ti`: acc1 -> relu(acc * s + b) -> bf16 -> the A2 operand tile (K-major, 128-byte swizzle)
        auto epilogue1 = [&](int ti) {
            if (ti > 0) {
                ptx::mbar_wait(a2_free, static_cast<uint32_t>(ti - 1) & 1);  // the previous unit's 1x1 has read the tile
            }
            ptx::mbar_wait(acc1_full, static_cast<uint32_t>(ti) & 1);
            ptx::tc_fence_after();
            if (epi_warp == 0 && lane == 0) stamp(ti, CT_EPI1_START);
            const int row = lane_group * 32 + lane;
#pragma unroll 1
            for (int u = group; u < C / 32; u += 2) {  // 32-column units alternate between the two groups
                uint32_t v[32];
                ptx::tmem_ld_32x32b_x32(taddr_lane + u * 32, v);
                ptx::tmem_ld_wait();
                const uint32_t sc = sb_sa + u * 32 * 4, bi = sb_sa + (C + u * 32) * 4;
                const uint32_t dst = a2_sa + (u >> 1) * kABytes + row * 128;
                float4 s0 = ptx::lds128f(sc), s1 = ptx::lds128f(sc + 16);
                float4 b0 = ptx::lds128f(bi), b1 = ptx::lds128f(bi + 16);
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) {  // 8 columns = one 16-byte unit of this thread's row
                    const float4 cs0 = s0, cs1 = s1, cb0 = b0, cb1 = b1;
                    if (jj < 3) {  // next piece's constants before this piece's arithmetic
                        s0 = ptx::lds128f(sc + (jj + 1) * 32), s1 = ptx::lds128f(sc + (jj + 1) * 32 + 16);
                        b0 = ptx::lds128f(bi + (jj + 1) * 32), b1 = ptx::lds128f(bi + (jj + 1) * 32 + 16);
                    }
                    const uint32_t* vv = &v[jj * 8];
                    uint4 ov;
                    ov.x = ptx::pack_bf16x2_relu(fmaf(__uint_as_float(vv[0]), cs0.x, cb0.x), fmaf(__uint_as_float(vv[1]), cs0.y, cb0.y));
                    ov.y = ptx::pack_bf16x2_relu(fmaf(__uint_as_float(vv[2]), cs0.z, cb0.z), fmaf(__uint_as_float(vv[3]), cs0.w, cb0.w));
                    ov.z = ptx::pack_bf16x2_relu(fmaf(__uint_as_float(vv[4]), cs1.x, cb1.x), fmaf(__uint_as_float(vv[5]), cs1.y, cb1.y));
                    ov.w = ptx::pack_bf16x2_relu(fmaf(__uint_as_float(vv[6]), cs1.z, cb1.z), fmaf(__uint_as_float(vv[7]), cs1.w, cb1.w));
                    const int unit16 = (u & 1) * 4 + jj;
                    ptx::sts128(dst + ((unit16 ^ (row & 7)) << 4), ov);
                }
            }
            ptx::fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's reads
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) {
                if (PAIR && cl_rank != 0)
                    ptx::mbar_arrive(a2_local);  // forwarded to the leader by this CTA's warp 1
                else
                    ptx::mbar_arrive(a2_full);
                if (epi_warp == 0) stamp(ti, CT_EPI1_DONE);
            }
        };

        int m_blk = 0, m_next = 0;
        bool m_valid = true, v_next = true;
        uint32_t uses = 0;  // n-tiles this group has drained so far (accumulator phase)
        bool have = tile_at(0, m_blk, m_valid);
        if (have) epilogue1(0);
        for (int ti = 0; have; ++ti) {
            const bool more = tile_at(ti + 1, m_next, v_next);
            const int m_row0 = m_blk * kBlockM + lane_group * 32;
            const bool slab_ok = m_valid && m_row0 < p.M;
            // write-out offsets, once per unit: lane (8 r + ch) moves 16-byte piece ch of slab rows 4 i + r (i = 0..7);
            // out_off[i] = element offset of that row's pixel in the output, or -1 for a position that is not a pixel
            int out_off[8];
            {
                int m = m_row0 + (lane >> 3);
                int rest = m / p.row_pitch;
                int q = m - rest * p.row_pitch;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    out_off[i] = (m < p.M && q < p.W) ? (rest * p.W + q) * p.N2 + (lane & 7) * 8 : -1;
                    m += 4;
                    q += 4;
                    if (q >= p.row_pitch) {
                        q -= p.row_pitch;
                        ++rest;
                    }
                }
            }
            if (p.l2_prefetch && lane == 0) {
                if (more && v_next && m_next * kBlockM + lane_group * 32 < p.M) {
                    const int m0 = m_next * kBlockM + lane_group * 32;
                    const int rest = m0 / p.row_pitch;
                    const int q = m0 - rest * p.row_pitch;
                    const int img = rest / p.H;
                    const int pr = rest - img * p.H;
                    for (int n2 = group; n2 < n2_tiles; n2 += 2)
                        for (int c = 0; c < kChainBN2 / kEpiCols; ++c)
                            ptx::tma_prefetch_im2col_4d(&tmap_res, n2 * kChainBN2 + c * kEpiCols, q, pr, img, 0, 0);
                }
            }
            // ---- epilogue 2: this group's n-tiles; the next unit's epilogue 1 is spliced in before the last chunk
#pragma unroll 1
            for (int n2 = group; n2 < n2_tiles; n2 += 2, ++uses) {
                ptx::mbar_wait(&acc2_full[group], uses & 1);
                ptx::tc_fence_after();
                if (epi_warp == 0 && lane == 0 && n2 == group) stamp(ti, CT_EPI2_FIRST);
                const uint32_t taddr = taddr_lane + kAcc2Col + group * kChainBN2;
#pragma unroll 1
                for (int c = 0; c < kChainBN2 / kEpiCols; ++c) {
                    const bool last_chunk = n2 + 2 >= n2_tiles && c == kChainBN2 / kEpiCols - 1;
                    if (last_chunk && more) epilogue1(ti + 1);
                    const bool st_c0 = epi_warp == 0 && lane == 0 && n2 == group && c == 0;
                    if (st_c0) stamp(ti, CT_C0_START);
                    // (the barrier probe takes ~200 cycles to answer even when the load has landed: it is issued
                    // ahead of the TMEM loads, its answer is looked at after them)
                    const bool res_landed = slab_ok && ptx::mbar_try_wait(&my_res_bar[slot], (res_par >> slot) & 1);
                    uint32_t v0[32], v1[32];
                    ptx::tmem_ld_32x32b_x32(taddr + c * kEpiCols, v0);
                    ptx::tmem_ld_32x32b_x32(taddr + c * kEpiCols + 32, v1);
                    ptx::tmem_ld_wait();
                    if (st_c0) stamp(ti, CT_C0_TMEM);
                    if (c == kChainBN2 / kEpiCols - 1) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) {
                            if (PAIR)
                                ptx::mbar_arrive_remote(&acc2_empty[group], 0);
                            else
                                ptx::mbar_arrive(&acc2_empty[group]);
                        }
                    }
                    const uint32_t slab = slabs_sa + slot * kEpiSlabBytes;
                    const uint32_t this_slot = slot;
                    if (++slot == ring) slot = 0;
                    if (!slab_ok) {
                        __syncwarp();
                        issue_residual(-1);
                        continue;
                    }
                    if (!res_landed) ptx::mbar_wait(&my_res_bar[this_slot], (res_par >> this_slot) & 1);
                    res_par ^= (1u << this_slot);
                    if (st_c0) stamp(ti, CT_C0_RES);
                    const int col0 = n2 * kChainBN2 + c * kEpiCols;
                    const uint32_t sc = sb_sa + (2 * C + col0) * 4, bi = sb_sa + (2 * C + 4 * C + col0) * 4;
                    // (the loads of 16-byte piece jj + 1 -- constants and shortcut -- are issued before the arithmetic of
                    // piece jj: the shared-memory accesses are ordered asm statements, the compiler cannot hoist them)
                    float4 s0 = ptx::lds128f(sc), s1 = ptx::lds128f(sc + 16);
                    float4 b0 = ptx::lds128f(bi), b1 = ptx::lds128f(bi + 16);
                    uint4 rv = ptx::lds128(slab + ((0 ^ (lane & 7)) << 4) + lane * 128);
#pragma unroll
                    for (int jj = 0; jj < 8; ++jj) {
                        const uint32_t* vv = (jj < 4) ? &v0[jj * 8] : &v1[(jj - 4) * 8];
                        const float4 cs0 = s0, cs1 = s1, cb0 = b0, cb1 = b1;
                        const uint32_t r[4] = {rv.x, rv.y, rv.z, rv.w};
                        const uint32_t sw = ((jj ^ (lane & 7)) << 4) + lane * 128;
                        if (jj < 7) {
                            s0 = ptx::lds128f(sc + (jj + 1) * 32), s1 = ptx::lds128f(sc + (jj + 1) * 32 + 16);
                            b0 = ptx::lds128f(bi + (jj + 1) * 32), b1 = ptx::lds128f(bi + (jj + 1) * 32 + 16);
                            rv = ptx::lds128(slab + (((jj + 1) ^ (lane & 7)) << 4) + lane * 128);
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
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            y[2 * e] += __uint_as_float(r[e] << 16);
                            y[2 * e + 1] += __uint_as_float(r[e] & 0xffff0000u);
                        }
                        uint4 ov;
                        ov.x = ptx::pack_bf16x2_relu(y[0], y[1]);
                        ov.y = ptx::pack_bf16x2_relu(y[2], y[3]);
                        ov.z = ptx::pack_bf16x2_relu(y[4], y[5]);
                        ov.w = ptx::pack_bf16x2_relu(y[6], y[7]);
                        ptx::sts128(slab + sw, ov);
                    }
                    __syncwarp();
                    if (st_c0) stamp(ti, CT_C0_MATH);
                    {
                        // coalesced write-out: 8 lanes cover one 128-byte row segment, 4 rows per instruction (all
                        // loads first: a store behind each load serialises on the shared-memory latency)
                        const int ch = lane & 7;
                        __nv_bfloat16* gcol = p.out + col0;
                        uint4 val[8];
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int row = i * 4 + (lane >> 3);
                            val[i] = ptx::lds128(slab + row * 128 + ((ch ^ (row & 7)) << 4));
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            if (out_off[i] >= 0) *reinterpret_cast<uint4*>(gcol + out_off[i]) = val[i];
                    }
                    if (st_c0) stamp(ti, CT_C0_OUT);
                    __syncwarp();
                    if (st_c0) stamp(ti, CT_C0_SYNC);
                    issue_residual(st_c0 ? ti : -1);
                    if (st_c0) stamp(ti, CT_C0_END);
                    if (epi_warp == 0 && lane == 0 && n2 == group && c == 1) stamp(ti, CT_C1_END);
                }
            }
            if ((epi_warp & 3) == 0 && lane == 0) stamp(ti, epi_warp == 0 ? CT_EPI2_DONE : CT_EPI2_G1_DONE);
            have = more;
            m_blk = m_next;
            m_valid = v_next;
        }
    }

    // ===================== teardown =====================
    ptx::tc_fence_before();
    __syncthreads();
    if (CL > 1) ptx::cluster_sync();
    if (warp_idx == 1) {
        ptx::tc_fence_after();
        if (PAIR)
            ptx::tmem_dealloc_pair(tmem_base, 512);
        else
            ptx::tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace adapt
