// Stem + max-pool in ONE kernel: `conv1_conv` (7x7 / stride 2 as the space-to-depth 4x4 / stride 1 form, +BN +ReLU) and
// `pool1_pad` + `pool1_pool` (3x3 / stride 2 / zero pad 1) of the Keras ResNet stem.
//
// Replaces, inside a stage's `model.predict` (reference src/node.py:177), the launches
//     y   = relu(conv(x) * s + b)          112 x 112 x 64 per image, 205 MB at batch 128: written to HBM, read back
//     out = maxpool3x3s2(zero_pad(y))       56 x 56 x 64
// by one whose only output is the pooled tensor (51 MB).  Values are identical to the two-launch form: the
// convolution accumulates in the same order, and max commutes with the two monotone steps that follow it there
// (ReLU, rounding to bf16), so "pool, then ReLU and round once" equals "ReLU and round, then pool"; zero padding is the
// neutral element because everything pooled is >= 0.
//
// Operand side = the stem-by-phase form of conv_gemm_sm100.cuh (A_S2DP): the image is a sequence of 128-byte units of
// four 16-channel pixels, one plain 2-D load of 136 consecutive units per filter row serves 128 output positions of all
// four column phases through shifted descriptors, into four 64-column accumulators.  A tile is FOUR WHOLE conv rows
// of one image (rows are 32 units wide), i.e. exactly the rows of two pooled rows minus one:
//     pooled row 2t   = max(conv rows 4t-1, 4t, 4t+1)        pooled row 2t+1 = max(conv rows 4t+1, 4t+2, 4t+3)
// so every CTA walks a CONTIGUOUS range of tiles and carries the horizontally pooled row 4t+3 to its next tile in
// shared memory (a CTA whose range starts inside an image first runs the tile before it for that one row).
//
//   warp 0      TMA producer (one stage per filter row; the 32 KiB weight matrix arrives with the first tile and stays)
//   warp 1      MMA issuer (16 tcgen05.mma per stage: 4 phases x 4 filter columns), accumulators double buffered
//   warps 2..9  epilogue, two groups of four alternating tiles; warp w of a group owns conv row 4t + w:
//               TMEM -> scale / bias -> max over the 3 columns of each pooled column (own unit's four phases + phase 3
//               of the lane to the left) -> ReLU + bf16 -> its row buffer; then, behind a group barrier, the 128
//               threads take the vertical max of three row buffers and write the two pooled rows, coalesced.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cuda.h>
#include <cstdint>

#include "conv_gemm_sm100.cuh"
#include "ptx_sm100.cuh"

namespace adapt {

struct StemPoolParams {
    int batch;
    int H;                // stored rows per image (s2d rows, >= Ho + 3)
    int Ho, Wo;           // conv output size (multiples of 4; Wo / 4 <= 32)
    int tiles_per_image;  // Ho / 4
    int total_tiles;      // batch * tiles_per_image
    const float* scale;   // [64] folded BN
    const float* bias;
    __nv_bfloat16* out;   // [batch, Ho / 2, Wo / 2, 64]
    int stages;           // operand ring depth (one filter row per stage)
};

constexpr int kStemPoolThreads = 320;
constexpr int kStemPoolMaxStages = 6;
constexpr int kStemPoolRowBytes = 64 * 128;  // a horizontally pooled row: up to 64 pixels x 64 channels bf16
constexpr int kStemPoolWBytes = 4 * 64 * kBlockK * 2;

__host__ __device__ constexpr int stem_pool_smem_bytes(int stages) {
    return stages * kRow3ABytes + kStemPoolWBytes + (2 * 4 + 2) * kStemPoolRowBytes + 2 * 64 * 4 + 1024 /*barriers*/ +
           1024 /*alignment*/;
}

__global__ void __launch_bounds__(kStemPoolThreads, 1)
stem_pool_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b, const StemPoolParams p) {
    constexpr int kMmaN = 64;
    constexpr int kBBytes = kMmaN * kBlockK * 2;  // one filter row of the weights: 64 x 64
    constexpr int kUnitsPerRow = 32;
    const int kStages = p.stages;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem_a + kStages * kRow3ABytes;
    uint8_t* smem_rows = smem_b + kStemPoolWBytes;                     // [2 groups][4 rows]
    uint8_t* smem_carry = smem_rows + 8 * kStemPoolRowBytes;           // [2]
    float* smem_sb = reinterpret_cast<float*>(smem_carry + 2 * kStemPoolRowBytes);  // scale[64] bias[64]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_sb + 128);
    uint64_t* full_bar = bars;                                  // [kStemPoolMaxStages]
    uint64_t* empty_bar = bars + kStemPoolMaxStages;            // [kStemPoolMaxStages]
    uint64_t* tmem_full_bar = bars + 2 * kStemPoolMaxStages;    // [2]
    uint64_t* tmem_empty_bar = tmem_full_bar + 2;               // [2]
    uint64_t* carry_full = tmem_empty_bar + 2;                  // [2] row 3 of a tile is in the carry slot
    uint64_t* carry_free = carry_full + 2;                      // [2] the next tile has consumed it
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(carry_free + 2);
    constexpr int kNumBars = 2 * kStemPoolMaxStages + 8;

    const int warp_idx = __shfl_sync(0xffffffff, threadIdx.x / 32, 0);
    const int lane = threadIdx.x % 32;

    // this CTA's contiguous range of tiles; `warm`: one tile before it, run only for the row it hands on
    const long long u0 = static_cast<long long>(blockIdx.x) * p.total_tiles / gridDim.x;
    const long long u1 = static_cast<long long>(blockIdx.x + 1) * p.total_tiles / gridDim.x;
    const int warm = (u1 > u0 && (u0 % p.tiles_per_image) != 0) ? 1 : 0;
    const int n_seq = static_cast<int>(u1 - u0) + warm;
    const int first_tile = static_cast<int>(u0) - warm;

    if (warp_idx == 0) {
        for (int i = lane; i < kNumBars; i += 32) {
            uint32_t count = 1;
            if (&bars[i] == tmem_empty_bar || &bars[i] == tmem_empty_bar + 1) count = 4;  // one arrive per warp of the group
            ptx::mbar_init(&bars[i], count);
        }
        ptx::fence_mbar_init();
    } else if (warp_idx == 2 && lane == 0) {
        ptx::prefetch_tmap(&tmap_a);
        ptx::prefetch_tmap(&tmap_b);
    }
    if (warp_idx == 1) {
        ptx::tmem_alloc(tmem_slot, 512);
        ptx::tmem_relinquish();
    }
    for (int i = threadIdx.x; i < 64; i += blockDim.x) {  // weights: nothing upstream writes them
        smem_sb[i] = __ldg(p.scale + i);
        smem_sb[64 + i] = __ldg(p.bias + i);
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    ptx::griddep_launch_dependents();
    ptx::griddep_wait();

    if (warp_idx == 0) {
        // ===================== TMA producer =====================
        int stage = 0;
        uint32_t phase = 0;
        for (int i = 0; i < n_seq; ++i) {
            const int u = first_tile + i;
            const int img = u / p.tiles_per_image;
            const int t = u - img * p.tiles_per_image;
            const int unit_row0 = (img * p.H + 4 * t) * kUnitsPerRow;
            for (int r4 = 0; r4 < 4; ++r4) {
                ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                if (ptx::elect_one()) {
                    const bool load_b = i == 0;
                    ptx::mbar_expect_tx(&full_bar[stage], kRow3ABytes + (load_b ? kBBytes : 0));
                    ptx::tma_load_2d(smem_a + stage * kRow3ABytes, &tmap_a, &full_bar[stage], 0, unit_row0 + r4 * kUnitsPerRow);
                    if (load_b) ptx::tma_load_2d(smem_b + r4 * kBBytes, &tmap_b, &full_bar[stage], r4 * kBlockK, 0);
                }
                __syncwarp();
                if (++stage == kStages) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
    } else if (warp_idx == 1) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(kBlockM, kMmaN);
        int stage = 0;
        uint32_t phase = 0;
        for (int i = 0; i < n_seq; ++i) {
            const int acc = i & 1;
            ptx::mbar_wait(&tmem_empty_bar[acc], ((static_cast<uint32_t>(i) >> 1) & 1) ^ 1);
            ptx::tc_fence_after();
            const uint32_t tmem_d = tmem_base + acc * 256;
            for (int r4 = 0; r4 < 4; ++r4) {
                ptx::mbar_wait(&full_bar[stage], phase);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t a_base = ptx::smem_u32(smem_a + stage * kRow3ABytes);
                    const uint64_t b_desc = ptx::make_kmajor_sw128_desc(ptx::smem_u32(smem_b + r4 * kBBytes));
#pragma unroll
                    for (int phi = 0; phi < 4; ++phi) {   // column phase: accumulator phi
#pragma unroll
                        for (int s4 = 0; s4 < 4; ++s4) {  // filter column = 32-byte chunk (phi + s) counted from unit j
                            const int chunk = phi + s4;
                            const uint64_t a_desc =
                                ptx::make_kmajor_sw128_desc(a_base + (chunk >> 2) * (kBlockK * 2) + (chunk & 3) * 32);
                            ptx::umma_bf16_ss(tmem_d + phi * kMmaN, a_desc, b_desc + 2 * s4, idesc, (r4 | s4) != 0 ? 1u : 0u);
                        }
                    }
                    ptx::umma_commit(&empty_bar[stage]);
                    if (r4 == 3) ptx::umma_commit(&tmem_full_bar[acc]);
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
        const int epi_warp = warp_idx - 2;
        const int group = epi_warp >> 2;
        const int w = warp_idx & 3;  // TMEM lanes [32 w, +32) = units of conv row 4t + w
        const int tg = w * 32 + lane;  // thread of the group
        const uint32_t rows_sa = ptx::smem_u32(smem_rows + group * 4 * kStemPoolRowBytes);
        const uint32_t carry_sa = ptx::smem_u32(smem_carry);
        const uint32_t sb_sa = ptx::smem_u32(smem_sb);
        const int Po = p.Ho / 2, Qo = p.Wo / 2;
        const bool lane_live = lane < p.Wo / 4;  // (unit columns past Wo exist only in the walk)

        for (int i = group; i < n_seq; i += 2) {
            const int u = first_tile + i;
            const int img = u / p.tiles_per_image;
            const int t = u - img * p.tiles_per_image;
            const bool is_warm = warm && i == 0;
            const bool hands_on = i + 1 < n_seq;  // row 3 goes to the next tile of this CTA
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(w * 32) << 16) + group * 256;
            ptx::mbar_wait(&tmem_full_bar[group], (static_cast<uint32_t>(i) >> 1) & 1);
            ptx::tc_fence_after();
            const bool do_row = !is_warm || w == 3;
            const bool to_carry = w == 3 && hands_on;
            if (to_carry) {
                // the carry slot's previous content (tile i - 2) has been consumed by tile i - 1
                ptx::mbar_wait(&carry_free[i & 1], ((static_cast<uint32_t>(i) >> 1) & 1) ^ 1);
            }
            if (do_row) {
                const uint32_t row_dst = rows_sa + w * kStemPoolRowBytes;
                const uint32_t carry_dst = carry_sa + (i & 1) * kStemPoolRowBytes;
#pragma unroll 1
                for (int cg = 0; cg < 4; ++cg) {  // 16 channels at a time, all four phases
                    uint32_t v[4][16];
#pragma unroll
                    for (int phi = 0; phi < 4; ++phi) ptx::tmem_ld_32x32b_x16(taddr + phi * kMmaN + cg * 16, v[phi]);
                    ptx::tmem_ld_wait();
                    float sc[16], bi[16];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 s4 = ptx::lds128f(sb_sa + (cg * 16 + q * 4) * 4);
                        const float4 b4 = ptx::lds128f(sb_sa + (64 + cg * 16 + q * 4) * 4);
                        sc[q * 4 + 0] = s4.x; sc[q * 4 + 1] = s4.y; sc[q * 4 + 2] = s4.z; sc[q * 4 + 3] = s4.w;
                        bi[q * 4 + 0] = b4.x; bi[q * 4 + 1] = b4.y; bi[q * 4 + 2] = b4.z; bi[q * 4 + 3] = b4.w;
                    }
                    uint32_t h0[8], h1[8];  // pooled columns 2j and 2j + 1, 16 channels each, bf16 pairs
#pragma unroll
                    for (int c = 0; c < 16; c += 2) {
                        float e[2], o[2];
#pragma unroll
                        for (int k = 0; k < 2; ++k) {
                            const float y0 = fmaf(__uint_as_float(v[0][c + k]), sc[c + k], bi[c + k]);
                            const float y1 = fmaf(__uint_as_float(v[1][c + k]), sc[c + k], bi[c + k]);
                            const float y2 = fmaf(__uint_as_float(v[2][c + k]), sc[c + k], bi[c + k]);
                            const float y3 = fmaf(__uint_as_float(v[3][c + k]), sc[c + k], bi[c + k]);
                            float left = __shfl_up_sync(0xffffffff, y3, 1);  // pixel 4j - 1 = phase 3 of the unit to the left
                            if (lane == 0) left = 0.0f;                      // (zero padding; everything pooled is >= 0 after ReLU)
                            e[k] = fmaxf(fmaxf(left, y0), y1);
                            o[k] = fmaxf(fmaxf(y1, y2), y3);
                        }
                        h0[c >> 1] = ptx::pack_bf16x2_relu(e[0], e[1]);
                        h1[c >> 1] = ptx::pack_bf16x2_relu(o[0], o[1]);
                    }
                    if (lane_live) {
                        // pixel px, 16-byte piece k of its 128-byte row at ((k ^ ((px >> 1) & 7)) << 4): conflict-free both
                        // for this write (lanes = every other pixel) and for the row-wise reads of the vertical step
                        const uint32_t sw = static_cast<uint32_t>(lane & 7);
                        const uint32_t px0 = static_cast<uint32_t>(2 * lane) * 128, px1 = px0 + 128;
                        const uint4 a0 = make_uint4(h0[0], h0[1], h0[2], h0[3]), a1 = make_uint4(h0[4], h0[5], h0[6], h0[7]);
                        const uint4 c0 = make_uint4(h1[0], h1[1], h1[2], h1[3]), c1 = make_uint4(h1[4], h1[5], h1[6], h1[7]);
                        const uint32_t k0 = ((2 * cg) ^ sw) << 4, k1 = ((2 * cg + 1) ^ sw) << 4;
                        if (!is_warm) {
                            ptx::sts128(row_dst + px0 + k0, a0);
                            ptx::sts128(row_dst + px0 + k1, a1);
                            ptx::sts128(row_dst + px1 + k0, c0);
                            ptx::sts128(row_dst + px1 + k1, c1);
                        }
                        if (to_carry) {
                            ptx::sts128(carry_dst + px0 + k0, a0);
                            ptx::sts128(carry_dst + px0 + k1, a1);
                            ptx::sts128(carry_dst + px1 + k0, c0);
                            ptx::sts128(carry_dst + px1 + k1, c1);
                        }
                    }
                }
            }
            // accumulator drained (or not needed): the MMA warp may start the tile after next
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tmem_empty_bar[group]);
            if (to_carry && lane == 0) ptx::mbar_arrive(&carry_full[i & 1]);  // (release: this lane's and, via syncwarp, the warp's writes)
            if (is_warm) continue;

            // ---- vertical step: pooled rows 2t (carry | row0 | row1) and 2t + 1 (row1 | row2 | row3)
            ptx::named_bar_sync(1 + group, 128);  // the four row buffers are written
            const bool have_prev = i >= 1;       // a tile of this CTA ran before this one: its row 3 is (or will be) in the carry
            const bool use_carry = t != 0;       // (t == 0: the row above is zero padding)
            if (have_prev) ptx::mbar_wait(&carry_full[(i - 1) & 1], (static_cast<uint32_t>(i - 1) >> 1) & 1);
            const uint32_t prev_sa = carry_sa + ((i - 1) & 1) * kStemPoolRowBytes;
            const int vecs_per_row = Qo * 8;
            __nv_bfloat16* out_img = p.out + (static_cast<size_t>(img) * Po + 2 * t) * Qo * 64;
            for (int vec = tg; vec < 2 * vecs_per_row; vec += 128) {
                const int R = vec >= vecs_per_row ? 1 : 0;
                const int rem = vec - R * vecs_per_row;
                const int px = rem >> 3, ch = rem & 7;
                const uint32_t off = static_cast<uint32_t>(px) * 128 + ((static_cast<uint32_t>(ch) ^ ((px >> 1) & 7)) << 4);
                uint4 a, b, c;
                if (R == 0) {
                    a = use_carry ? ptx::lds128(prev_sa + off) : make_uint4(0u, 0u, 0u, 0u);
                    b = ptx::lds128(rows_sa + 0 * kStemPoolRowBytes + off);
                    c = ptx::lds128(rows_sa + 1 * kStemPoolRowBytes + off);
                } else {
                    a = ptx::lds128(rows_sa + 1 * kStemPoolRowBytes + off);
                    b = ptx::lds128(rows_sa + 2 * kStemPoolRowBytes + off);
                    c = ptx::lds128(rows_sa + 3 * kStemPoolRowBytes + off);
                }
                uint4 m;
                m.x = ptx::max_bf16x2(ptx::max_bf16x2(a.x, b.x), c.x);
                m.y = ptx::max_bf16x2(ptx::max_bf16x2(a.y, b.y), c.y);
                m.z = ptx::max_bf16x2(ptx::max_bf16x2(a.z, b.z), c.z);
                m.w = ptx::max_bf16x2(ptx::max_bf16x2(a.w, b.w), c.w);
                *reinterpret_cast<uint4*>(out_img + static_cast<size_t>(R) * Qo * 64 + rem * 8) = m;
            }
            ptx::named_bar_sync(1 + group, 128);  // row buffers and the carry slot are free again
            if (have_prev && tg == 0) ptx::mbar_arrive(&carry_free[(i - 1) & 1]);
        }
    }

    // ===================== teardown =====================
    ptx::tc_fence_before();
    __syncthreads();
    if (warp_idx == 1) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace adapt
