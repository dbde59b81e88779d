// K1 v2: stride-1 "same" convolution over NHWC fp32 as an implicit GEMM on tcgen05 (kind::tf32, 3xTF32, fp32
// accumulate in TMEM) where the tap-shifted A operand is read STRAIGHT FROM THE HALO TILE in shared memory.
//
// v1 (conv_tc.cuh) re-split every pixel row into tf32 hi/lo for each of the R*R taps and staged it through TMEM; ncu
// showed it latency-bound on that ping-pong (profiles/r01_conv_tc_v1_ncu.md: 24-29 us against a 3.5 us tensor floor).
// Here the halo of a 16x16-pixel super-tile is loaded once by TMA (SWIZZLE_128B, out-of-bounds zero fill = padding),
// split once into hi / lo planes in place, and every tap is just a different START ADDRESS of the K-major UMMA
// descriptor: pixel (ty, tx) of an 8-column M-tile is halo row (ty + r) * P + tx + s, i.e. 8-row groups at a uniform
// stride SBO = P * 128 bytes.  The hardware applies the 128-byte swizzle to absolute shared-memory addresses
// (tools/umma_shift_probe.cu, measured on B200), so a start address that is a multiple of 128 B but not of 1024 B
// reads exactly what TMA wrote.
//
// One CTA = one 16x16 super-tile = two M=128 tiles (x in [0,8) and [8,16)) sharing the halo and every weight k-block.
//   warp 0      TMA producer: all halo chunks up front, then the weight k-blocks [w_hi ; w_lo] through a ring
//   warp 1      TMEM allocator + single-thread MMA issuer
//   warps 2-5   hi/lo split per 32-channel chunk, then the epilogue (bias, act, act', res + alpha*v, NHWC store)
// Shapes: all input channels resident (hi + lo planes), i.e. (c0 + c1) / 32 * 2 * plane + weight ring <= 227 KB:
// the 3x3 and 1x1 convolutions with 64 input channels (the whole ODE function); everything else stays on v1.
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "conv_tc.cuh"

namespace odex {

struct Conv2Geom {
    int B, H, W, R;
    int nsrc, csrc;     // 1 or 2 sources of csrc channels (csrc % 32 == 0)
    int P;              // halo pitch in pixel slots = 16 + 2 * pad
    int plane;          // bytes of one 32-channel halo plane (hi or lo), multiple of 1024
    int b_stages;
    int aux_mode;       // 1: the epilogue's one extra operand tile (act' mask or residual) is TMA-prefetched into shared memory
    int stage_off;      // byte offset of the output staging tile inside the halo planes
    long long* dbg;     // diagnostic cycle counters of CTA 0 (ODEX_CONV_DEBUG=1), else nullptr
};

template <int N, int R>
__global__ void __launch_bounds__(192, 1)
conv_tc2_kernel(const __grid_constant__ CUtensorMap tmA0, const __grid_constant__ CUtensorMap tmA1,
                const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmAux,
                const Conv2Geom g,
                const ConvEpilogue ep) {
    constexpr int kBStageBytes = 2 * N * 128;
    constexpr uint32_t kTmemCols = 4 * N;          // 2 M-tiles x (main N + corr N)
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int NC = g.nsrc * g.csrc / 32;           // 32-channel chunks
    const int SB = g.b_stages;
    const uint32_t bst0 = base + static_cast<uint32_t>(NC) * 2u * g.plane;
    const uint32_t bars = bst0 + SB * kBStageBytes;
    const uint32_t bar_raw_full = bars;            // [NC <= 8]
    const uint32_t bar_split_full = bars + 64;     // [NC <= 8]
    const uint32_t bar_b_full = bars + 128;        // [SB <= 4]
    const uint32_t bar_b_empty = bars + 160;       // [SB <= 4]
    const uint32_t bar_acc = bars + 192;
    const uint32_t tmem_slot = bars + 200;
    const uint32_t bar_early = bars + 208;         // MMAs of every chunk but the last have completed
    const uint32_t bar_aux = bars + 216;           // the epilogue operand tile has landed
    const uint32_t bias_s = bars + 256;            // [N] floats, staged once so the epilogue never waits on global memory

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    constexpr int taps = R * R;
    constexpr int pad = R >> 1;
    constexpr int P = 16 + 2 * pad;            // == g.P

    if (threadIdx.x == 0) {
        for (int i = 0; i < NC; ++i) { mbar_init(bar_raw_full + 8 * i, 1); mbar_init(bar_split_full + 8 * i, 128); }
        for (int i = 0; i < SB; ++i) { mbar_init(bar_b_full + 8 * i, 1); mbar_init(bar_b_empty + 8 * i, 1); }
        mbar_init(bar_acc, 1);
        mbar_init(bar_early, 1);
        mbar_init(bar_aux, 1);
        fence_barrier_init();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA0);
        tma_prefetch_desc(&tmA1);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmOut);
        tma_prefetch_desc(&tmAux);
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    const int tiles_x = g.W >> 4, tiles_y = g.H >> 4;
    const int x0 = (blockIdx.x % tiles_x) << 4;
    const int y0 = ((blockIdx.x / tiles_x) % tiles_y) << 4;
    const int img = blockIdx.x / (tiles_x * tiles_y);
    const int hrows = (16 + 2 * pad) * g.P;

    // Everything above touched only shared memory / TMEM, so it may overlap the tail of the previous kernel on the
    // stream; from here on global memory is read.
    pdl_wait();
    pdl_trigger();
    const bool skip = ep.skip != nullptr && *ep.skip != 0;      // CTA-uniform: a skipped attempt only tears down

    if (skip) {
    } else if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            for (int c = 0; c < NC; ++c) {
                const int ch = c * 32;
                const int src = ch / g.csrc;
                const CUtensorMap* tm = src == 0 ? &tmA0 : &tmA1;
                mbar_arrive_expect_tx(bar_raw_full + 8 * c, static_cast<uint32_t>(hrows) * 128u);
                tma_load_4d(base + static_cast<uint32_t>(c) * 2u * g.plane, tm, bar_raw_full + 8 * c, ch - src * g.csrc,
                            x0 - pad, y0 - pad, img);
            }
            int sb = 0; uint32_t pb = 0;
            long long w_be = 0; const long long t0_ = clock64();
            for (int c = 0; c < NC; ++c) {
                for (int tap = 0; tap < taps; ++tap) {
                    const int kb = ((c >> 1) * taps + tap) * 2 + (c & 1);       // packing order of odex_pack_weights(tc)
                    const long long c0_ = clock64();
                    mbar_wait(bar_b_empty + 8 * sb, pb ^ 1u);
                    w_be += clock64() - c0_;
                    mbar_arrive_expect_tx(bar_b_full + 8 * sb, kBStageBytes);
                    const uint32_t dst = bst0 + sb * kBStageBytes;
                    if (N == 128) {   // box rows are limited to 256: two boxes of N rows
                        tma_load_2d(dst, &tmB, bar_b_full + 8 * sb, 0, kb * 2 * N);
                        tma_load_2d(dst + N * 128, &tmB, bar_b_full + 8 * sb, 0, kb * 2 * N + N);
                    } else {
                        tma_load_2d(dst, &tmB, bar_b_full + 8 * sb, 0, kb * 2 * N);
                    }
                    if (++sb == SB) { sb = 0; pb ^= 1u; }
                }
            }
            if (g.dbg && blockIdx.x == 0) { g.dbg[0] = w_be; g.dbg[1] = clock64() - t0_; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        // whole warp runs the loop (warp-uniform operands stay in uniform registers); one elected lane issues
        {
            constexpr uint32_t idesc_wide = umma_idesc_tf32(128, 2 * N);
            constexpr uint32_t idesc_narrow = umma_idesc_tf32(128, N);
            // descriptor words (K-major SWIZZLE_128B): hi = SBO | version | layout, lo = LBO(16 B) << 16 | start >> 4;
            // stepping through taps / k8 / M-tiles is one 32-bit add on the low word
            const uint32_t a_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, P * 128) >> 32);
            const uint32_t b_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024) >> 32);
            constexpr uint32_t lbo16 = 1u << 16;
            int sb = 0; uint32_t pb = 0;
            long long w_sp = 0, w_bf = 0; const long long t0_ = clock64();
            for (int c = 0; c < NC; ++c) {
                { const long long c0_ = clock64(); mbar_wait(bar_split_full + 8 * c, 0u); w_sp += clock64() - c0_; }
                tc_fence_after_sync();
                const uint32_t hi = base + static_cast<uint32_t>(c) * 2u * g.plane;
                const uint32_t lo_delta = static_cast<uint32_t>(g.plane) >> 4;
                int r = 0, s = 0;
                for (int tap = 0; tap < taps; ++tap) {
                    { const long long c0_ = clock64(); mbar_wait(bar_b_full + 8 * sb, pb); w_bf += clock64() - c0_; }
                    tc_fence_after_sync();
                    const uint32_t b_lo0 = lbo16 | ((bst0 + sb * kBStageBytes) >> 4);
                    const uint32_t a_lo0 = lbo16 | ((hi >> 4) + static_cast<uint32_t>(r * P + s) * 8u);
                    if (elect_one()) {
#pragma unroll
                        for (int k8 = 0; k8 < 4; ++k8) {
#pragma unroll
                            for (int mt = 0; mt < 2; ++mt) {
                                const uint32_t a_lo = a_lo0 + mt * 64u + k8 * 2u;
                                const uint32_t d = tmem_base + mt * 2 * N;
                                umma_tf32_ss_w(d, a_lo, a_hi32, b_lo0 + k8 * 2u, b_hi32, idesc_wide, (c | tap | k8) != 0);
                                umma_tf32_ss_w(d + N, a_lo + lo_delta, a_hi32, b_lo0 + k8 * 2u, b_hi32, idesc_narrow, 1u);
                            }
                        }
                        umma_commit(bar_b_empty + 8 * sb);
                    }
                    __syncwarp();
                    if (++sb == SB) { sb = 0; pb ^= 1u; }
                    if (++s == R) { s = 0; ++r; }
                }
                if (g.aux_mode && c == NC - 2) {        // chunks 0 .. NC-2 are dead once these MMAs retire
                    if (elect_one()) umma_commit(bar_early);
                    __syncwarp();
                }
            }
            if (elect_one()) umma_commit(bar_acc);
            __syncwarp();
            if (g.dbg && blockIdx.x == 0 && lane == 0) { g.dbg[2] = w_sp; g.dbg[3] = w_bf; g.dbg[4] = clock64() - t0_; }
        }
    } else {
        // ===================== hi/lo split, then epilogue =====================
        const int tid = threadIdx.x - 64;
        const int nf4 = hrows * 8;
        if (tid < N / 4) {
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ep.bias) bv = __ldg(reinterpret_cast<const float4*>(ep.bias) + tid);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(bias_s + 16u * tid), "f"(bv.x), "f"(bv.y), "f"(bv.z), "f"(bv.w) : "memory");
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        long long w_raw = 0; const long long t0_ = clock64();
        for (int c = 0; c < NC; ++c) {
            { const long long c0_ = clock64(); mbar_wait(bar_raw_full + 8 * c, 0u); w_raw += clock64() - c0_; }
            const uint32_t hi = base + static_cast<uint32_t>(c) * 2u * g.plane;
            // 4 independent 16-byte loads in flight per thread (the volatile asm keeps program order: a plain loop is one
            // load-use latency per 16 bytes)
            for (int i0 = tid; i0 < nf4; i0 += 4 * 128) {
                uint32_t x[4][4];
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int i = i0 + 128 * b;
                    x[b][0] = x[b][1] = x[b][2] = x[b][3] = 0u;
                    if (i < nf4)
                        asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                                     : "=r"(x[b][0]), "=r"(x[b][1]), "=r"(x[b][2]), "=r"(x[b][3])
                                     : "r"(hi + static_cast<uint32_t>(i) * 16u));
                }
#pragma unroll
                for (int b = 0; b < 4; ++b) {
                    const int i = i0 + 128 * b;
                    const uint32_t a = hi + static_cast<uint32_t>(i) * 16u;
                    uint32_t h[4], l[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        h[e] = tf32_rna_bits(x[b][e]);
                        l[e] = tf32_rna_bits(__float_as_uint(__uint_as_float(x[b][e]) - __uint_as_float(h[e])));
                    }
                    if (i < nf4) {
                        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(h[0]), "r"(h[1]), "r"(h[2]), "r"(h[3]) : "memory");
                        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(a + g.plane), "r"(l[0]), "r"(l[1]), "r"(l[2]), "r"(l[3]) : "memory");
                    }
                }
            }
            fence_proxy_async_smem();
            mbar_arrive(bar_split_full + 8 * c);
        }

        // ---------- epilogue: (main + corr) + bias -> act -> act'(saved) -> res + alpha*v -> NHWC store ----------
        const long long t_split = clock64() - t0_;
        const uint32_t stage = base + static_cast<uint32_t>(g.stage_off);
        if (g.aux_mode) {
            // the act' mask / residual tile goes into the planes of the chunks the tensor core is done with, while the
            // last chunk's MMAs are still running; the output is staged in the last chunk's planes afterwards
            if (tid == 0) {
                mbar_wait(bar_early, 0u);
                mbar_arrive_expect_tx(bar_aux, 256u * N * 4u);
#pragma unroll
                for (int pl = 0; pl < N / 32; ++pl) tma_load_4d(base + pl * 32768u, &tmAux, bar_aux, pl * 32, x0, y0, img);
            }
            __syncwarp();
        }
        mbar_wait(bar_acc, 0u);
        tc_fence_after_sync();
        if (g.aux_mode) mbar_wait(bar_aux, 0u);
        const long long t_acc = clock64() - t0_;
        // Phase 1: TMEM -> registers -> shared staging tile (the halo planes are idle now), laid out as the TMA store box:
        // N / 32 planes of [256 pixels][32 channels], 128-byte rows with the SWIZZLE_128B chunk XOR (conflict-free for
        // this row-per-thread write).  With no residual / act' operand the whole epilogue (bias, activation) happens
        // here in registers; otherwise the raw sums are staged and Phase 2 finishes them in place with coalesced loads.
        const int q = warp & 3;
        const int m = q * 32 + lane;                  // TMEM lane = pixel of the M-tile: (ty, tx) = (m / 8, m % 8)
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        constexpr int kChunks = N / 4;                // 16-byte chunks per pixel (16 or 32)
        const bool fast = g.aux_mode || (ep.dact == nullptr && ep.res == nullptr);     // whole epilogue in registers
        const int aux_kind = !g.aux_mode ? 0 : ep.res ? 3 : ep.dact_kind;                // 1 relu', 2 tanh', 3 residual
        float alpha = ep.alpha;
        if (ep.alpha_dev) alpha *= static_cast<float>(*ep.alpha_dev);
        const float4* bias4 = reinterpret_cast<const float4*>(smem_raw + (bias_s - smem_u32(smem_raw)));
        // 64 accumulator columns (main + correction) per trip: four TMEM loads in flight, one wait
#pragma unroll 1
        for (int it = 0; it < N / 32; ++it) {
            const int mt = it / (N / 64);
            const int c0 = (it % (N / 64)) * 64;
            const int p = (m >> 3) * 16 + mt * 8 + (m & 7);
            uint32_t vm[2][32], vc[2][32];
            tmem_ld32(tmem_base + lane_addr + mt * 2 * N + c0, vm[0]);
            tmem_ld32(tmem_base + lane_addr + mt * 2 * N + N + c0, vc[0]);
            tmem_ld32(tmem_base + lane_addr + mt * 2 * N + c0 + 32, vm[1]);
            tmem_ld32(tmem_base + lane_addr + mt * 2 * N + N + c0 + 32, vc[1]);
            tmem_ld_wait();
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float4 bv[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) bv[j] = fast ? bias4[(c0 + 32 * h) / 4 + j] : make_float4(0.f, 0.f, 0.f, 0.f);
                const uint32_t rel = static_cast<uint32_t>(c0 / 32 + h) * 32768u + static_cast<uint32_t>(p) * 128u;
                const uint32_t row = stage + rel;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float a[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) a[e] = __uint_as_float(vm[h][4 * j + e]) + __uint_as_float(vc[h][4 * j + e]);
                    a[0] += bv[j].x; a[1] += bv[j].y; a[2] += bv[j].z; a[3] += bv[j].w;
                    if (fast) {
                        if (ep.act == 1) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) a[e] = relu_nan(a[e]);
                        } else if (ep.act == 2) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) a[e] = tanhf(a[e]);
                        }
                    }
                    const uint32_t sw = (static_cast<uint32_t>(j) ^ (static_cast<uint32_t>(p) & 7u)) << 4;
                    if (aux_kind) {
                        float x[4];
                        asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(x[0]), "=f"(x[1]), "=f"(x[2]), "=f"(x[3]) : "r"(base + rel + sw));
                        if (aux_kind == 1) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) a[e] = x[e] > 0.f ? a[e] : 0.f;
                        } else if (aux_kind == 2) {
#pragma unroll
                            for (int e = 0; e < 4; ++e) a[e] *= 1.f - x[e] * x[e];
                        } else {
#pragma unroll
                            for (int e = 0; e < 4; ++e) a[e] = fmaf(alpha, a[e], x[e]);
                        }
                    }
                    const uint32_t dst = row + sw;
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(dst), "f"(a[0]), "f"(a[1]), "f"(a[2]), "f"(a[3]));
                }
            }
        }
        tc_fence_before_sync();
        if (fast) fence_proxy_async_smem();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const long long t_ph1 = clock64() - t0_;

        if (!fast) {
            // Phase 2: finish in place.  Consecutive threads take consecutive 16-byte chunks of a pixel row, so the residual /
            // saved-activation loads are full 128-byte lines.
            const int jch = tid & (kChunks - 1);      // this thread's chunk of every row it visits (128 % kChunks == 0)
            float4 bv;
            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(bv.x), "=f"(bv.y), "=f"(bv.z), "=f"(bv.w) : "r"(bias_s + 16u * jch));
#pragma unroll 4
            for (int i = tid; i < 256 * kChunks; i += 128) {
                const int p = i / kChunks;
                const uint32_t src = stage + static_cast<uint32_t>(jch >> 3) * 32768u + static_cast<uint32_t>(p) * 128u +
                                     ((static_cast<uint32_t>(jch & 7) ^ (static_cast<uint32_t>(p) & 7u)) << 4);
                float v[4];
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]) : "r"(src));
                const long pix = (static_cast<long>(img) * g.H + y0 + (p >> 4)) * g.W + x0 + (p & 15);
                v[0] += bv.x; v[1] += bv.y; v[2] += bv.z; v[3] += bv.w;
                if (ep.act == 1) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) v[e] = relu_nan(v[e]);
                } else if (ep.act == 2) {
#pragma unroll
                    for (int e = 0; e < 4; ++e) v[e] = tanhf(v[e]);
                }
                if (ep.dact) {
                    const float4 mk = *reinterpret_cast<const float4*>(ep.dact + pix * ep.ldd + 4 * jch);
                    if (ep.dact_kind == 1) {
                        v[0] = mk.x > 0.f ? v[0] : 0.f; v[1] = mk.y > 0.f ? v[1] : 0.f;
                        v[2] = mk.z > 0.f ? v[2] : 0.f; v[3] = mk.w > 0.f ? v[3] : 0.f;
                    } else {
                        v[0] *= 1.f - mk.x * mk.x; v[1] *= 1.f - mk.y * mk.y;
                        v[2] *= 1.f - mk.z * mk.z; v[3] *= 1.f - mk.w * mk.w;
                    }
                }
                if (ep.res) {
                    const float4 rr = *reinterpret_cast<const float4*>(ep.res + pix * ep.ldr + 4 * jch);
                    v[0] = fmaf(alpha, v[0], rr.x); v[1] = fmaf(alpha, v[1], rr.y);
                    v[2] = fmaf(alpha, v[2], rr.z); v[3] = fmaf(alpha, v[3], rr.w);
                }
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(src), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
            }
            fence_proxy_async_smem();
            asm volatile("bar.sync 1, 128;" ::: "memory");
        }
        // Phase 3: one thread hands the tile to the TMA store engine (full-line asynchronous writes)
        if (tid == 0) {
#pragma unroll
            for (int pl = 0; pl < N / 32; ++pl) tma_store_4d(&tmOut, stage + pl * 32768u, pl * 32, x0, y0, img);
            tma_store_commit();
            tma_store_wait_read();
        }
        tc_fence_before_sync();
        if (g.dbg && blockIdx.x == 0 && tid == 0) { g.dbg[5] = w_raw; g.dbg[6] = t_split; g.dbg[7] = t_acc; g.dbg[8] = clock64() - t0_; g.dbg[9] = t_ph1; }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn2 encode_fn2() {
    static EncodeTiledFn2 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn2>(p);
    });
    return fn;
}

static constexpr int kSmemMax2 = 232448;

static int conv2_plan(const odex_conv_desc* d, Conv2Geom* g, int* smem) {
    const int pad = d->R / 2;
    g->B = d->B; g->H = d->H; g->W = d->W; g->R = d->R;
    g->nsrc = d->x1 ? 2 : 1; g->csrc = d->c0;
    g->P = 16 + 2 * pad;
    const int hrows = (16 + 2 * pad) * g->P;
    g->plane = (hrows * 128 + 1023) / 1024 * 1024;
    const int NC = g->nsrc * g->csrc / 32;
    const int fixed = 1024 + NC * 2 * g->plane + 256 + 512;     // alignment slack, halo planes, barriers, bias
    const int bstage = 2 * d->N * 128;
    int sb = (kSmemMax2 - fixed) / bstage;
    if (sb > 4) sb = 4;
    g->b_stages = sb;
    *smem = fixed + sb * bstage;
    g->dbg = nullptr;
    g->aux_mode = 0;
    g->stage_off = 0;
    {
        const int tile = 256 * d->N * 4;
        const bool one = (d->dact != nullptr) != (d->res != nullptr);
        if (one && NC >= 2 && tile <= 2 * g->plane && tile <= (NC - 1) * 2 * g->plane) {
            g->aux_mode = 1;
            g->stage_off = (NC - 1) * 2 * g->plane;
        }
    }
    if (256 * d->N * 4 > NC * 2 * g->plane) return 1;      // the epilogue stages the output tile in the halo planes
    return sb >= 2 ? 0 : 1;
}

bool conv_tc2_supported(const odex_conv_desc* d) {
    if (!d->w_tc) return false;
    if (d->N != 64 && d->N != 128) return false;
    if (d->R != 1 && d->R != 3 && d->R != 5) return false;
    if (d->c0 % 64 != 0 || (d->x1 && d->c1 != d->c0)) return false;     // weights are packed per 64-channel group
    if (d->H % 16 != 0 || d->W % 16 != 0) return false;
    if ((d->x1 ? 2 : 1) * d->c0 / 32 > 8) return false;
    if (d->ld0 % 4 || (d->x1 && d->ld1 % 4) || d->ldo % 4) return false;
    Conv2Geom g; int smem;
    return conv2_plan(d, &g, &smem) == 0;
}

static int make_act_map2(CUtensorMap* tm, const float* x, int ld, int C, int B, int H, int W, int bw, int bh) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, static_cast<cuuint64_t>(W) * ld * 4,
                                   static_cast<cuuint64_t>(H) * W * ld * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(bw), static_cast<cuuint32_t>(bh), 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = encode_fn2()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(activation, v2) failed: %d", static_cast<int>(r));
    return 0;
}

template <int N, int R>
static int launch_tc2(const odex_conv_desc* d, cudaStream_t s) {
    if (!encode_fn2()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    Conv2Geom g; int smem;
    if (conv2_plan(d, &g, &smem)) return fail(ODEX_E_UNSUPPORTED, "conv(tc2): tile does not fit shared memory");
    const int pad = d->R / 2;
    CUtensorMap tmA0, tmA1, tmB, tmOut, tmAux;
    if (int rc = make_act_map2(&tmOut, d->out, d->ldo, d->N, d->B, d->H, d->W, 16, 16)) return rc;
    tmAux = tmOut;
    if (g.aux_mode) {
        const float* aux = d->res ? d->res : d->dact;
        if (int rc = make_act_map2(&tmAux, aux, d->res ? d->ldr : d->ldd, d->N, d->B, d->H, d->W, 16, 16)) return rc;
    }
    if (int rc = make_act_map2(&tmA0, d->x0, d->ld0, d->c0, d->B, d->H, d->W, g.P, 16 + 2 * pad)) return rc;
    if (d->x1) {
        if (int rc = make_act_map2(&tmA1, d->x1, d->ld1, d->c1, d->B, d->H, d->W, g.P, 16 + 2 * pad)) return rc;
    } else {
        tmA1 = tmA0;
    }
    {
        const long rows = static_cast<long>(g.nsrc * g.csrc / 32) * d->R * d->R * 2 * N;
        const cuuint64_t dims[2] = {32, static_cast<cuuint64_t>(rows)};
        const cuuint64_t strides[1] = {128};
        const cuuint32_t box[2] = {32, 128};
        const cuuint32_t es[2] = {1, 1};
        CUresult r = encode_fn2()(&tmB, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(d->w_tc), dims, strides, box,
                                  es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(weights, v2) failed: %d", static_cast<int>(r));
    }
    ConvEpilogue ep;
    ep.bias = d->bias; ep.out = d->out; ep.ldo = d->ldo;
    ep.res = d->res; ep.ldr = d->ldr; ep.alpha = d->alpha; ep.alpha_dev = d->alpha_dev;
    ep.dact = d->dact; ep.ldd = d->ldd; ep.dact_kind = d->dact_kind;
    ep.act = d->act; ep.skip = d->skip;

    auto kern = conv_tc2_kernel<N, R>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] { attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax2); });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "conv(tc2): cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    const int grid = d->B * (d->H / 16) * (d->W / 16);
    static const bool debug = getenv("ODEX_CONV_DEBUG") != nullptr;
    if (debug) {
        static long long* dbuf = nullptr;
        if (!dbuf) cudaMalloc(&dbuf, 16 * sizeof(long long));
        cudaMemsetAsync(dbuf, 0, 16 * sizeof(long long), s);
        g.dbg = dbuf;
    }
    launch_pdl(kern, grid, 192, smem, s, tmA0, tmA1, tmB, tmOut, tmAux, g, ep);
    if (debug) {
        long long h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, g.dbg, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[conv_tc2 R=%d C=%d N=%d B=%d grid=%d SB=%d smem=%d] producer: wait_b_empty=%lld total=%lld | mma: wait_split=%lld "
                "wait_b_full=%lld total=%lld | split: wait_raw=%lld split_done=%lld acc_done=%lld staged=%lld end=%lld (cycles)\n", R, g.nsrc * g.csrc,
                N, d->B, grid, g.b_stages, smem, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7], h[9], h[8]);
    }
    return launched("conv_tc2_kernel");
}

int launch_conv_tc2(const odex_conv_desc* d, cudaStream_t s) {
    if (d->N == 64) {
        if (d->R == 1) return launch_tc2<64, 1>(d, s);
        if (d->R == 3) return launch_tc2<64, 3>(d, s);
        return launch_tc2<64, 5>(d, s);
    }
    if (d->R == 1) return launch_tc2<128, 1>(d, s);
    if (d->R == 3) return launch_tc2<128, 3>(d, s);
    return launch_tc2<128, 5>(d, s);
}

}  // namespace odex
