// K1/K2: stride-1 "same" convolution (3x3 or 5x5) over NHWC fp32 as an implicit GEMM on the 5th-gen tensor
// cores (tcgen05, kind::tf32, fp32 accumulate in TMEM) with 3xTF32 error compensation, so fp32-level
// tolerances hold:   x = x_hi + x_lo,  w = w_hi + w_lo   (each rounded to tf32)
//     y = x_hi*w_hi                      -> "main"  accumulator, TMEM columns [0, N)
//       + x_hi*w_lo + x_lo*w_hi          -> "corr"  accumulator, TMEM columns [N, 2N)
// The weights are pre-split and packed [w_hi ; w_lo] along N per (tap, 32-channel) k-block, so x_hi needs ONE
// MMA of width 2N per k8 and x_lo one of width N.
//
// Replaces the ATen conv2d calls of the reference at helpers/utils.py:167-177 (3x3, via
// modules/DiffEqSolver.py:77) and modules/ConvGRUCell.py:41-43,47-49 (5x5); torch.cat((x, h)) of
// modules/ConvGRUCell.py:72,79 is folded in as two K ranges ("sources").
//
// One CTA = 128 output pixels (TH full rows of one image) x N output channels.  Warp roles:
//   warps 0-3  transform: read the tap-shifted pixel row from the halo tile in smem (TMA-loaded once per
//              64-channel group, zero-filled out of bounds = the conv padding), split hi/lo, stage the A
//              operand in TMEM (tcgen05.st; default) or in swizzled smem (SS mode); then the epilogue.
//   warp  4    TMA producer: halo tiles (4-D tensor map over NHWC) and weight k-blocks (2-D tensor map).
//   warp  5    TMEM allocator + single-thread tcgen05.mma issuer.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace odex {

struct ConvGeom {
    int B, H, W;       // images, rows, cols (W power of two, 8 <= W <= 128, (128 / W) divides H)
    int nsrc, csrc;    // 1 or 2 source tensors of csrc channels each (csrc % 64 == 0)
    int R;             // taps per side, 3 or 5
    int TH;            // rows per tile = 128 / W
    int log2W;
    int halo_stages;   // 1 or 2 (2 only useful with more than one 64-channel group)
    int w_rows;        // packed-weight rows per k-block = 2 * (output channels of the whole convolution)
    int w_lo;          // row offset of the lo half inside a k-block = output channels of the whole convolution
    int w_off;         // first output channel of this launch (N-blocks of 128 when the convolution has more than 128)
    int halo_bytes;    // per 32-channel chunk, rounded up to 1024
};

struct ConvEpilogue {
    const float* bias;   // [N] or nullptr
    float* out;          // NHWC, row (pixel) stride ldo floats
    int ldo;
    const float* res;    // optional: out = res + alpha * v
    int ldr;
    float alpha;
    const double* alpha_dev;   // optional: alpha *= (float)*alpha_dev (step size kept on the device)
    const float* dact;   // optional saved activation a: v *= act'(a)  (1 = ReLU: a > 0; 2 = tanh: 1 - a*a)
    int ldd;
    int dact_kind;
    int act;             // 0 none, 1 ReLU, 2 tanh -- applied to acc + bias
    const int* skip;     // optional device flag: non-zero -> the whole grid exits immediately
};

template <int N>
struct ConvCfg {
    static constexpr int kThreads = 192;
    static constexpr int kBStages = 3;
    static constexpr int kMinCtas = (N == 64) ? 2 : 1;                  // 2 CTAs/SM: one's epilogue hides under the other's MMAs
    static constexpr int kAStages = (N == 64) ? 2 : 4;                  // x 64 TMEM columns (or 32 KB smem)
    static constexpr int kAStagesSS = 2;
    static constexpr int kBStageBytes = 2 * N * 128;
    static constexpr int kTmemCols = (2 * N + kAStages * 64 <= 256) ? 256 : 512;
};

__host__ __device__ inline int conv_smem_bytes(int N, int halo_bytes, int halo_stages, int b_stages, int a_stages_ss) {
    return 1024 /*align slack*/ + halo_stages * 2 * halo_bytes + b_stages * 2 * N * 128 + a_stages_ss * 32768 + 256;
}

template <int N, bool kATmem>
__global__ void __launch_bounds__(192, ConvCfg<N>::kMinCtas)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
               const __grid_constant__ CUtensorMap tmB, const ConvGeom g, const ConvEpilogue ep) {
    using Cfg = ConvCfg<N>;
    constexpr int SB = Cfg::kBStages;
    constexpr int SA = kATmem ? Cfg::kAStages : Cfg::kAStagesSS;

    extern __shared__ uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int G = g.nsrc * g.csrc / 64;              // 64-channel groups
    const int HS = g.halo_stages;
    const int taps = g.R * g.R;
    const int pad = g.R >> 1;
    const int HWp = g.W + 2 * pad;                   // halo row pitch in pixels
    const int HP = (g.TH + 2 * pad) * HWp;           // halo pixels
    const uint32_t halo0 = smem_base;
    const uint32_t bst0 = halo0 + HS * 2 * g.halo_bytes;
    const uint32_t ast0 = bst0 + SB * Cfg::kBStageBytes;           // SS mode only
    const uint32_t bars = ast0 + (kATmem ? 0 : SA * 32768);
    // barrier map (8 bytes each)
    const uint32_t bar_halo_full = bars;             // [2]
    const uint32_t bar_halo_empty = bars + 16;       // [2]
    const uint32_t bar_b_full = bars + 32;           // [SB<=4]
    const uint32_t bar_b_empty = bars + 64;          // [SB<=4]
    const uint32_t bar_a_full = bars + 96;           // [SA<=4]
    const uint32_t bar_a_empty = bars + 128;         // [SA<=4]
    const uint32_t bar_acc = bars + 160;
    const uint32_t tmem_slot = bars + 168;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) { mbar_init(bar_halo_full + 8 * i, 1); mbar_init(bar_halo_empty + 8 * i, 128); }
        for (int i = 0; i < SB; ++i) { mbar_init(bar_b_full + 8 * i, 1); mbar_init(bar_b_empty + 8 * i, 1); }
        for (int i = 0; i < SA; ++i) { mbar_init(bar_a_full + 8 * i, 128); mbar_init(bar_a_empty + 8 * i, 1); }
        mbar_init(bar_acc, 1);
        fence_barrier_init();
    }
    if (warp == 4 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmB);
    }
    if (warp == 5) {
        tmem_alloc(tmem_slot, Cfg::kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    const int tiles_per_img = g.H / g.TH;
    const int img = blockIdx.x / tiles_per_img;
    const int y0 = (blockIdx.x % tiles_per_img) * g.TH;
    const int KB = G * taps * 2;

    // shared memory / TMEM set-up above may overlap the previous kernel's tail; global memory is first read below
    pdl_wait();
    pdl_trigger();
    const bool skip = ep.skip != nullptr && *ep.skip != 0;      // CTA-uniform: a skipped attempt only tears down

    if (skip) {
    } else if (warp == 4) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int sb = 0; uint32_t pb = 0;
            int kb = 0;
            int next_halo = 0;
            for (int grp = 0; grp < G; ++grp) {
                // halo tiles run up to HS groups ahead of the weight stream
                for (; next_halo < G && next_halo < grp + HS; ++next_halo) {
                    const int hg = next_halo;
                    const int hs = hg % HS;
                    const uint32_t use = static_cast<uint32_t>(hg / HS);
                    mbar_wait(bar_halo_empty + 8 * hs, (use & 1u) ^ 1u);
                    mbar_arrive_expect_tx(bar_halo_full + 8 * hs, 2u * HP * 128u);
                    const int ch = hg * 64;
                    const int src = ch / g.csrc;
                    const int c0 = ch % g.csrc;
                    const CUtensorMap* tm = src == 0 ? &tmA0 : &tmA1;
                    const uint32_t dst = halo0 + hs * 2 * g.halo_bytes;
                    tma_load_4d(dst, tm, bar_halo_full + 8 * hs, c0, -pad, y0 - pad, img);
                    tma_load_4d(dst + g.halo_bytes, tm, bar_halo_full + 8 * hs, c0 + 32, -pad, y0 - pad, img);
                }
                for (int i = 0; i < taps * 2; ++i, ++kb) {
                    mbar_wait(bar_b_empty + 8 * sb, pb ^ 1u);
                    mbar_arrive_expect_tx(bar_b_full + 8 * sb, Cfg::kBStageBytes);
                    if (N == 128) {   // box rows are limited to 256: two boxes of N rows
                        tma_load_2d(bst0 + sb * Cfg::kBStageBytes, &tmB, bar_b_full + 8 * sb, 0, kb * g.w_rows + g.w_off);
                        tma_load_2d(bst0 + sb * Cfg::kBStageBytes + N * 128, &tmB, bar_b_full + 8 * sb, 0, kb * g.w_rows + g.w_lo + g.w_off);
                    } else {
                        tma_load_2d(bst0 + sb * Cfg::kBStageBytes, &tmB, bar_b_full + 8 * sb, 0, kb * g.w_rows);
                    }
                    if (++sb == SB) { sb = 0; pb ^= 1u; }
                }
            }
        }
    } else if (warp == 5) {
        // ===================== MMA issuer =====================
        // whole warp runs the loop (warp-uniform operands stay in uniform registers); one elected lane issues
        {
            constexpr uint32_t idesc_wide = umma_idesc_tf32(128, 2 * N);
            constexpr uint32_t idesc_narrow = umma_idesc_tf32(128, N);
            int sb = 0, sa = 0; uint32_t pb = 0, pa = 0;
            for (int kb = 0; kb < KB; ++kb) {
                mbar_wait(bar_b_full + 8 * sb, pb);
                mbar_wait(bar_a_full + 8 * sa, pa);
                tc_fence_after_sync();
                const uint32_t bsm = bst0 + sb * Cfg::kBStageBytes;
                if (elect_one()) {
#pragma unroll
                for (int k8 = 0; k8 < 4; ++k8) {
                    const uint64_t bd = umma_desc_sw128(bsm + k8 * 32, 16, 1024);
                    if (kATmem) {
                        const uint32_t a_hi = tmem_base + 2 * N + sa * 64 + k8 * 8;
                        umma_tf32_ts(tmem_base, a_hi, bd, idesc_wide, (kb | k8) != 0);
                        umma_tf32_ts(tmem_base + N, a_hi + 32, bd, idesc_narrow, 1u);
                    } else {
                        const uint32_t asm_ = ast0 + sa * 32768;
                        const uint64_t ahi = umma_desc_sw128(asm_ + k8 * 32, 16, 1024);
                        const uint64_t alo = umma_desc_sw128(asm_ + 16384 + k8 * 32, 16, 1024);
                        umma_tf32_ss(tmem_base, ahi, bd, idesc_wide, (kb | k8) != 0);
                        umma_tf32_ss(tmem_base + N, alo, bd, idesc_narrow, 1u);
                    }
                }
                umma_commit(bar_a_empty + 8 * sa);
                umma_commit(bar_b_empty + 8 * sb);
                }
                __syncwarp();
                if (++sb == SB) { sb = 0; pb ^= 1u; }
                if (++sa == SA) { sa = 0; pa ^= 1u; }
            }
            if (elect_one()) umma_commit(bar_acc);
            __syncwarp();
        }
    } else {
        // ===================== transform (A operand) + epilogue: thread <-> output pixel =====================
        const int m = threadIdx.x;                       // 0..127 == TMEM lane
        const int ty = m >> g.log2W;
        const int tx = m & (g.W - 1);
        const uint32_t lane_addr = static_cast<uint32_t>(warp * 32) << 16;
        int sa = 0; uint32_t pa = 0;
        for (int grp = 0; grp < G; ++grp) {
            const int hs = grp % HS;
            mbar_wait(bar_halo_full + 8 * hs, static_cast<uint32_t>(grp / HS) & 1u);
            const uint32_t hbase = halo0 + hs * 2 * g.halo_bytes;
            for (int r = 0; r < g.R; ++r) {
                for (int s = 0; s < g.R; ++s) {
                    const int hp = (ty + r) * HWp + tx + s;
                    const uint32_t rowoff = static_cast<uint32_t>(hp) * 128u;
                    const uint32_t sw = static_cast<uint32_t>(hp & 7);
#pragma unroll 1
                    for (int cc = 0; cc < 2; ++cc) {
                        const uint32_t src = hbase + cc * g.halo_bytes + rowoff;
                        uint32_t hi[32], lo[32];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            uint32_t a, b, c, d;
                            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                                         : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
                                         : "r"(src + ((static_cast<uint32_t>(j) ^ sw) << 4)));
                            const uint32_t x[4] = {a, b, c, d};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                const uint32_t h = tf32_rna_bits(x[e]);
                                hi[4 * j + e] = h;
                                lo[4 * j + e] = tf32_rna_bits(__float_as_uint(__uint_as_float(x[e]) - __uint_as_float(h)));
                            }
                        }
                        mbar_wait(bar_a_empty + 8 * sa, pa ^ 1u);
                        if (kATmem) {
                            tc_fence_after_sync();
                            const uint32_t dst = tmem_base + lane_addr + 2 * N + sa * 64;
                            tmem_st32(dst, hi);
                            tmem_st32(dst + 32, lo);
                            tmem_st_wait();
                            tc_fence_before_sync();
                        } else {
                            const uint32_t dst = ast0 + sa * 32768 + static_cast<uint32_t>(m) * 128u;
                            const uint32_t swm = static_cast<uint32_t>(m & 7);
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const uint32_t o = (static_cast<uint32_t>(j) ^ swm) << 4;
                                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst + o), "r"(hi[4 * j]),
                                             "r"(hi[4 * j + 1]), "r"(hi[4 * j + 2]), "r"(hi[4 * j + 3]) : "memory");
                                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16384 + o),
                                             "r"(lo[4 * j]), "r"(lo[4 * j + 1]), "r"(lo[4 * j + 2]), "r"(lo[4 * j + 3])
                                             : "memory");
                            }
                            fence_proxy_async_smem();
                        }
                        mbar_arrive(bar_a_full + 8 * sa);
                        if (++sa == SA) { sa = 0; pa ^= 1u; }
                    }
                }
            }
            mbar_arrive(bar_halo_empty + 8 * hs);
        }

        // ---------- epilogue: (main + corr) + bias -> relu -> mask -> res + alpha*v -> NHWC store ----------
        mbar_wait(bar_acc, 0u);
        tc_fence_after_sync();
        const long pix = (static_cast<long>(img) * g.H + y0 + ty) * g.W + tx;
        float* orow = ep.out + pix * ep.ldo;
        const float* rrow = ep.res ? ep.res + pix * ep.ldr : nullptr;
        const float* mrow = ep.dact ? ep.dact + pix * ep.ldd : nullptr;
        float alpha = ep.alpha;
        if (ep.alpha_dev) alpha *= static_cast<float>(*ep.alpha_dev);
#pragma unroll 1
        for (int c0 = 0; c0 < N; c0 += 32) {
            uint32_t vm[32], vc[32];
            tmem_ld32(tmem_base + lane_addr + c0, vm);
            tmem_ld32(tmem_base + lane_addr + N + c0, vc);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float v[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float t = __uint_as_float(vm[4 * j + e]) + __uint_as_float(vc[4 * j + e]);
                    if (ep.bias) t += __ldg(ep.bias + c0 + 4 * j + e);
                    if (ep.act == 1) t = relu_nan(t);
                    else if (ep.act == 2) t = tanhf(t);
                    v[e] = t;
                }
                if (mrow) {
                    const float4 mk = *reinterpret_cast<const float4*>(mrow + c0 + 4 * j);
                    if (ep.dact_kind == 1) {
                        v[0] = mk.x > 0.f ? v[0] : 0.f; v[1] = mk.y > 0.f ? v[1] : 0.f;
                        v[2] = mk.z > 0.f ? v[2] : 0.f; v[3] = mk.w > 0.f ? v[3] : 0.f;
                    } else {
                        v[0] *= 1.f - mk.x * mk.x; v[1] *= 1.f - mk.y * mk.y;
                        v[2] *= 1.f - mk.z * mk.z; v[3] *= 1.f - mk.w * mk.w;
                    }
                }
                if (rrow) {
                    const float4 rr = *reinterpret_cast<const float4*>(rrow + c0 + 4 * j);
                    v[0] = fmaf(alpha, v[0], rr.x); v[1] = fmaf(alpha, v[1], rr.y);
                    v[2] = fmaf(alpha, v[2], rr.z); v[3] = fmaf(alpha, v[3], rr.w);
                }
                *reinterpret_cast<float4*>(orow + c0 + 4 * j) = make_float4(v[0], v[1], v[2], v[3]);
            }
        }
        tc_fence_before_sync();
    }
    __syncthreads();
    if (warp == 5) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, Cfg::kTmemCols);
    }
}

}  // namespace odex
