// Weight gradient of the stride-1 "same" convolution on tcgen05 kind::f16 with the two-term fp16 operand split (C-ABI
// odex_conv2d_wgrad_f16; the kind::tf32 / 3xTF32 original is wgrad_tc.cu, whose structure this file keeps):
//     dw[tap][n][c] += sum_p dy[p][n] * X[p + tap][c]          (reduction over PIXELS)
// Replaces the autograd conv backward-weight of helpers/utils.py:167-177 (3x3) and modules/ConvGRUCell.py:41-43,47-49 (5x5).
//
// K runs over pixels, which are shared-memory ROWS in NHWC (one pixel slot = 64 channels = 128 bytes of fp16), so X is an
// MN-major B operand (SWIZZLE_128B) and the tap shift (r, s) is only the start address of its descriptor.  dy is the K-major A
// operand [dy_hi ; dy_lo] (M = 128) in TMEM (TS mode: the tensor core reads only the X atoms from shared memory), two K
// elements (pixels) packed per 32-bit column.
//
// Operand split: x * s = hi + lo with hi = fp16(x * s), lo = fp16(x * s - hi).  The accumulator of a tap lives in TMEM over ALL
// tiles a CTA visits, so s must be one number per tensor: odex_conv2d_wgrad_f16 first runs a small max-|.| kernel over X and dy
// (two floats in a caller-provided scratch) and both scales put the tensor's max in [2^12, 2^13).  At that position lo needs no
// extra 2^11 (it is resolved to 2^-24 absolute against a max of 2^13: 2^-37 relative), so all four partial products accumulate
// into the same 64 TMEM columns exactly as in the tf32 kernel -- at twice its MMA rate, K = 16 pixels per instruction.
//
// One CTA = (tap group, 64-channel group of X, 64-channel block of dy, a strided set of 64-pixel tiles).
//   warp 0      TMA producer: raw fp32 X halo + dy tile per stage
//   warp 1      TMEM allocator + MMA issue
//   warps 2-9   fp32 -> fp16 hi / lo (X in place: the two 32-channel fp32 planes of a stage become its hi and lo planes; dy
//               into TMEM), bias-gradient column sums; warps 2-5 then run the epilogue (red.global.add.v4.f32)
#include <cuda.h>
#include <cuda_fp16.h>

#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"
#include "f16_split.cuh"

namespace odex {

constexpr int kW16TH = 4;            // tile = 4 rows x 16 pixels
constexpr int kW16TW = 16;
constexpr int kW16DyPlane = kW16TH * kW16TW * 128;     // 8 KB per 32-channel fp32 plane of dy

struct Wgrad16Geom {
    int B, H, W, R;
    int Ctot, N;        // channels of X (all sources) and of dy
    int c0;             // channels of source 0 (groups beyond it come from source 1)
    int xplane;         // bytes per halo plane (hrows * 128), multiple of 1024
    int TG, CG, NB, PS; // grid factors: tap groups, 64-channel groups of X, 64-channel blocks of dy, pixel splits
    int taps_per_group;
    int ntiles;         // B * (H / 4) * (W / 16)
    float* dwp;         // [R*R][N][Ctot], accumulated
    float* db;          // [N] or nullptr, accumulated
    // operand maxima: one pair for the whole tensor (amax_stride 0: absmax2_kernel's, or the caller's) or one pair PER IMAGE
    // (amax_stride 1: amax_x[b] >= max |X[b]|, amax_dy[b] >= max |dy[b]|, as the producing kernels measured them)
    const float* amax_x; const float* amax_dy; int amax_stride;
    long long* dbg;     // diagnostic: per-role cycle counters of CTA 0 (ODEX_WGRAD_DEBUG=1), else nullptr
};

// max |.| of up to three strided [P, C] fp32 matrices: out[0] <- max over a and b (the X sources), out[1] <- c (dy).
// C / 4 is a power of two for every shape of the path (64 or 128 channels): shifts, four independent 16-byte loads in flight.
__global__ void __launch_bounds__(256) absmax2_kernel(const float* __restrict__ a, int lda, int ca, const float* __restrict__ b, int ldb,
                                                      int cb, const float* __restrict__ c, int ldc, int cc, long P, float* out) {
    pdl_wait();
    pdl_trigger();
    float m[2] = {0.f, 0.f};
    const float* ptr[3] = {a, b, c};
    const int ld[3] = {lda, ldb, ldc}, ch[3] = {ca, cb, cc};
#pragma unroll
    for (int t = 0; t < 3; ++t) {
        if (!ptr[t]) continue;
        const int c4 = ch[t] >> 2;
        const int sh = 31 - __clz(c4);
        const bool pow2 = (c4 & (c4 - 1)) == 0;
        const long n4 = P * c4;
        const long stride = static_cast<long>(gridDim.x) * blockDim.x;
        float mx = 0.f;
        long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x;
        for (; i + 3 * stride < n4; i += 4 * stride) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const long j = i + u * stride;
                const long r = pow2 ? (j >> sh) : (j / c4);
                const long col = pow2 ? (j & (c4 - 1)) : (j % c4);
                v[u] = ld4(ptr[t] + r * ld[t] + col * 4);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) mx = fmaxf(mx, fmaxf(fmaxf(fabsf(v[u].x), fabsf(v[u].y)), fmaxf(fabsf(v[u].z), fabsf(v[u].w))));
        }
        for (; i < n4; i += stride) {
            const long r = pow2 ? (i >> sh) : (i / c4);
            const long col = pow2 ? (i & (c4 - 1)) : (i % c4);
            const float4 v = ld4(ptr[t] + r * ld[t] + col * 4);
            mx = fmaxf(mx, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
        }
        m[t == 2 ? 1 : 0] = fmaxf(m[t == 2 ? 1 : 0], mx);
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        float v = m[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
        if ((threadIdx.x & 31) == 0 && v > 0.f) atomicMax(reinterpret_cast<unsigned int*>(out + k), __float_as_uint(v));
    }
}

__device__ __forceinline__ void w16_red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void w16_tile_coords(const Wgrad16Geom& g, int t, int& img, int& y0, int& x0) {
    const int tx = g.W / kW16TW, ty = g.H / kW16TH;
    x0 = (t % tx) * kW16TW;
    y0 = ((t / tx) % ty) * kW16TH;
    img = t / (tx * ty);
}

// kind::f16, A K-major from TMEM, B MN-major (bit 16) from shared memory
__host__ __device__ constexpr uint32_t w16_idesc(int M, int N) {
    return (1u << 4) | (1u << 16) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void umma_f16_ts_w(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                              uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

// Operand scales.  Every tap accumulator is shared by all the tiles (images) a CTA visits, so what has to be common is the scale
// of the PRODUCT, 2^t; how t is split between the two operands is free per image:
//     s_x(b) puts max |X[b]| in [2^12, 2^13),      s_dy(b) = 2^t / s_x(b),
//     t from  max_b (max |X[b]| * max |dy[b]|) * 2^t  in (2^24, 2^26]      =>  max |dy[b]| * s_dy(b) < 2^14 for every b.
// An image whose product of maxima is the largest keeps >= 11 + 11 bits in both operands; one whose product is 2^-k of that
// loses k bits in dy -- and contributes 2^-k of the result.  So maxima that differ by any factor between the images of a launch
// (a state that grows by orders of magnitude over a rollout while its adjoint shrinks) cost nothing, where one pair of maxima
// for the whole tensor would flush the small images' operands to zero.  Exponent arithmetic on the biased exponents:
// v in [2^(e-127), 2^(e-126)).
__device__ __forceinline__ int w16_bexp(float v) { return static_cast<int>((__float_as_uint(v) >> 23) & 0xFFu); }
__device__ __forceinline__ float w16_pow2(int biased) {
    biased = biased < 1 ? 1 : (biased > 254 ? 254 : biased);
    return __uint_as_float(static_cast<uint32_t>(biased) << 23);
}
// max over the launch's images of e_x + e_d (0 when every product is zero); whole warp, result in every lane
__device__ __forceinline__ int w16_pmax_exp(const Wgrad16Geom& g, int lane) {
    int pe = 0;
    const int n = g.amax_stride ? g.B : 1;
    for (int b = lane; b < n; b += 32) {
        const int ex = w16_bexp(__ldg(g.amax_x + b)), ed = w16_bexp(__ldg(g.amax_dy + b));
        if (ex != 0 && ed != 0) pe = max(pe, ex + ed);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) pe = max(pe, __shfl_xor_sync(0xffffffffu, pe, o));
    return pe;
}
// scales of image b: x * sx <= 2^13, dy * sdy < 2^14, sx * sdy = 2^(278 - pmax_e)
__device__ __forceinline__ void w16_scales(const Wgrad16Geom& g, int b, int pmax_e, float* sx, float* sdy) {
    const int ex = w16_bexp(__ldg(g.amax_x + b * g.amax_stride));
    int sxe = (ex == 0 || ex == 255) ? 127 : 266 - ex;
    sxe = sxe < 1 ? 1 : (sxe > 254 ? 254 : sxe);
    *sx = w16_pow2(sxe);
    // an all-zero (or denormal) X image contributes nothing: dy * 0 keeps an unbounded dy from overflowing into inf * 0
    *sdy = ex == 0 ? 0.f : w16_pow2(pmax_e == 0 ? 127 : (278 - pmax_e) - (sxe - 127) + 127);
}

template <int R>
__global__ void __launch_bounds__(320, 1)
wgrad_f16_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                 const __grid_constant__ CUtensorMap tmDY, const Wgrad16Geom g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t stage_bytes = 2u * g.xplane + 2u * kW16DyPlane;   // [X c0-31 -> X hi][X c32-63 -> X lo][dy n0-31][dy n32-63]
    constexpr int S = (R <= 3) ? 5 : 4;            // pipeline stages (shared-memory tiles and TMEM A slots): all that fit 227 KB
    const uint32_t bars = base + static_cast<uint32_t>(S) * stage_bytes;
    const uint32_t bar_raw_full = bars;            // [S <= 6]
    const uint32_t bar_split_full = bars + 48;     // [S]
    const uint32_t bar_stage_empty = bars + 96;    // [S]
    const uint32_t bar_acc = bars + 144;
    const uint32_t tmem_slot = bars + 152;

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    // ---- job decode
    int j = blockIdx.x;
    const int ps = j % g.PS; j /= g.PS;
    const int nb = j % g.NB; j /= g.NB;
    const int cg = j % g.CG; j /= g.CG;
    const int tg = j;
    constexpr int taps = R * R;
    const int tap0 = tg * g.taps_per_group;
    const int ntap = min(g.taps_per_group, taps - tap0);
    constexpr int pad = R >> 1;
    constexpr int P = kW16TW + 2 * pad;
    constexpr int hrows = (kW16TH + 2 * pad) * P;   // pixel slots of one halo plane
    const uint32_t a_col0 = static_cast<uint32_t>(ntap * 64);        // S slots x 32 columns (64 pixels, two per column) of A
    uint32_t tmem_cols = 32;
    while (tmem_cols < a_col0 + 32u * S) tmem_cols <<= 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < S; ++i) {
            mbar_init(bar_raw_full + 8 * i, 1);
            mbar_init(bar_split_full + 8 * i, 128);
            mbar_init(bar_stage_empty + 8 * i, 1);
        }
        mbar_init(bar_acc, 1);
        fence_barrier_init();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmX0);
        tma_prefetch_desc(&tmX1);
        tma_prefetch_desc(&tmDY);
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, tmem_cols);
        tmem_relinquish();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    uint32_t tmem_base;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    pdl_wait();
    pdl_trigger();

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int cabs = cg * 64;
            const bool second = cabs >= g.c0;
            const CUtensorMap* tmX = second ? &tmX1 : &tmX0;
            const int cx = second ? cabs - g.c0 : cabs;
            constexpr uint32_t tx_bytes = 2u * hrows * 128u + 2u * kW16DyPlane;
            int it = 0;
            long long w_empty = 0, t_begin = clock64();
            for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
                const int st = it % S;
                const uint32_t use = static_cast<uint32_t>(it / S);
                const long long c0_ = clock64();
                mbar_wait(bar_stage_empty + 8 * st, (use & 1u) ^ 1u);
                w_empty += clock64() - c0_;
                int img, y0, x0;
                w16_tile_coords(g, t, img, y0, x0);
                const uint32_t xs = base + st * stage_bytes;
                const uint32_t ds = xs + 2u * g.xplane;
                const uint32_t bar = bar_raw_full + 8 * st;
                mbar_arrive_expect_tx(bar, tx_bytes);
                tma_load_4d(xs, tmX, bar, cx, x0 - pad, y0 - pad, img);
                tma_load_4d(xs + g.xplane, tmX, bar, cx + 32, x0 - pad, y0 - pad, img);
                tma_load_4d(ds, &tmDY, bar, nb * 64, x0, y0, img);
                tma_load_4d(ds + kW16DyPlane, &tmDY, bar, nb * 64 + 32, x0, y0, img);
            }
            if (g.dbg && blockIdx.x == 0) { g.dbg[0] = w_empty; g.dbg[1] = clock64() - t_begin; g.dbg[2] = it; }
        }
    } else if (warp == 1) {
        // ===================== MMA issue (whole warp loops, one elected lane issues) =====================
        constexpr uint32_t idesc = w16_idesc(128, 64);
        const uint32_t d_hi = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024) >> 32);      // SBO 1024: 8-pixel groups along K
        const uint32_t b_lbo = ((static_cast<uint32_t>(g.xplane) >> 4) & 0x3FFFu) << 16;      // (one 64-channel atom along N: unused)
        const uint32_t lo_delta = static_cast<uint32_t>(g.xplane) >> 4;                       // X hi plane -> X lo plane
        int it = 0;
        long long w_split = 0, t_begin = clock64();
        for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
            const int st = it % S;
            const uint32_t use = static_cast<uint32_t>(it / S);
            const long long c0_ = clock64();
            mbar_wait(bar_split_full + 8 * st, use & 1u);
            w_split += clock64() - c0_;
            tc_fence_after_sync();
            const uint32_t xs = base + st * stage_bytes;
            const uint32_t a_tmem = tmem_base + a_col0 + static_cast<uint32_t>(st) * 32u;
            int r = tap0 / R, s = tap0 - r * R;
            for (int tl = 0; tl < ntap; ++tl) {
                const uint32_t d_tmem = tmem_base + tl * 64;
                const uint32_t b_lo0 = b_lbo | ((xs >> 4) + static_cast<uint32_t>(r * P + s) * 8u);
                if (elect_one()) {
#pragma unroll
                    for (int ty = 0; ty < kW16TH; ++ty) {
                        const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(ty * P) * 8u;        // 16 pixels of tile row ty, shifted by the tap
                        const uint32_t a_at = a_tmem + static_cast<uint32_t>(ty * 8);            // the same 16 pixels of dy: 8 columns
                        umma_f16_ts_w(d_tmem, a_at, b_lo, d_hi, idesc, (it | ty) != 0);
                        umma_f16_ts_w(d_tmem, a_at, b_lo + lo_delta, d_hi, idesc, 1u);
                    }
                }
                __syncwarp();
                if (++s == R) { s = 0; ++r; }
            }
            if (elect_one()) umma_commit(bar_stage_empty + 8 * st);
            __syncwarp();
        }
        if (elect_one()) umma_commit(bar_acc);
        __syncwarp();
        if (g.dbg && blockIdx.x == 0 && lane == 0) { g.dbg[3] = w_split; g.dbg[4] = clock64() - t_begin; }
    } else {
        // ===================== operand preparation (8 warps), then epilogue (the first 4 of them) =====================
        // thread <-> TMEM lane L = 32 * (warp % 4) + lane: dy channel n = L % 64, part = L / 64 (0: hi rows, 1: lo rows).  The
        // preparation of a tile is a latency chain that did not overlap the MMAs of the previous one (tools/wgrad_probe.py);
        // two groups of four warps take alternate tiles, so each chain spans two tile periods.
        const int grp = (warp - 2) >> 2;                  // two groups of four warps take alternate tiles (see wgrad_tc.cu)
        const int tid = (threadIdx.x - 64) & 127;         // 0..127 within the group
        const int q = warp & 3;
        const int L = q * 32 + lane;
        const int n = L & 63;
        const int part = L >> 6;
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        const bool do_db = g.db != nullptr && tg == 0 && cg == 0 && part == 0;
        float dbsum = 0.f;
        const int pmax_e = w16_pmax_exp(g, lane);
        const int tiles_per_img = (g.H / kW16TH) * (g.W / kW16TW);
        // dy element (pixel k, channel n) of the raw SWIZZLE_128B tile: plane n / 32, row k, 16-byte chunk ((n % 32) / 4) ^ (k % 8)
        const uint32_t dy_off = static_cast<uint32_t>(n >> 5) * kW16DyPlane + static_cast<uint32_t>(n & 3) * 4u;
        const uint32_t dy_chunk = static_cast<uint32_t>((n & 31) >> 2);
        const int xhalf = tid & 1;                        // X: a lane pair owns a pixel slot, one 32-channel raw plane each
        int it = 0;
        long long w_raw = 0, t_x = 0, t_dy = 0, t_begin = clock64();
        for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
            if ((it & 1) != grp) continue;
            float sx, sdy;
            w16_scales(g, t / tiles_per_img, pmax_e, &sx, &sdy);
            const int st = it % S;
            const uint32_t use = static_cast<uint32_t>(it / S);
            const long long c0_ = clock64();
            mbar_wait(bar_raw_full + 8 * st, use & 1u);
            const long long c1_ = clock64();
            w_raw += c1_ - c0_;
            const uint32_t xs = base + st * stage_bytes;
            const uint32_t ds = xs + 2u * g.xplane + dy_off;
            // ---- X: slot i holds 32 + 32 fp32 channels in the two raw planes; it becomes 64 fp16 hi (plane 0) + 64 fp16 lo
            // (plane 1), in place: both lanes of the pair read their raw plane, then (after __syncwarp) write their half rows
#pragma unroll 1
            for (int i0_ = 0; i0_ < hrows; i0_ += 64) {
                const int i = i0_ + (tid >> 1);
                const bool ok = i < hrows;
                const uint32_t row = xs + static_cast<uint32_t>(ok ? i : 0) * 128u;
                const uint32_t sw = static_cast<uint32_t>(i) & 7u;
                float4 xv[8];
                // the two lanes of a pair read the same slot of the two raw planes, which are a multiple of 1 KB apart: the same
                // banks.  Lane 1 therefore walks its eight 16-byte chunks rotated by four (xv[c] = chunk c ^ 4).
                const uint32_t rot = static_cast<uint32_t>(xhalf) << 2;
                if (ok) {
#pragma unroll
                    for (int c = 0; c < 8; ++c)
                        xv[c] = lds128f(row + static_cast<uint32_t>(xhalf) * g.xplane + (((static_cast<uint32_t>(c) ^ rot) ^ sw) << 4));
                }
                __syncwarp();
                if (ok) {
#pragma unroll
                    for (int c8 = 0; c8 < 4; ++c8) {
                        const float4 u = xv[2 * c8], w = xv[2 * c8 + 1];
                        const float o[8] = {u.x * sx, u.y * sx, u.z * sx, u.w * sx, w.x * sx, w.y * sx, w.z * sx, w.w * sx};
                        uint32_t h[4], l[4];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const __half2 hh = __floats2half2_rn(o[2 * e], o[2 * e + 1]);
                            const float2 hf = __half22float2(hh);
                            const __half2 ll = __floats2half2_rn(o[2 * e] - hf.x, o[2 * e + 1] - hf.y);
                            h[e] = *reinterpret_cast<const uint32_t*>(&hh);
                            l[e] = *reinterpret_cast<const uint32_t*>(&ll);
                        }
                        const uint32_t off = (static_cast<uint32_t>(4 * xhalf + (c8 ^ (2 * xhalf))) ^ sw) << 4;   // xv[2 c8], xv[2 c8 + 1] = 8 channels (c8 ^ 2 xhalf)
                        sts128(row + off, make_uint4(h[0], h[1], h[2], h[3]));
                        sts128(row + g.xplane + off, make_uint4(l[0], l[1], l[2], l[3]));
                    }
                }
            }
            // ---- dy: this lane's channel at 32 of the tile's pixels -> A operand row in TMEM (K-major, two pixels per column)
            const long long c2_ = clock64();
            t_x += c2_ - c1_;
            tc_fence_after_sync();
            uint32_t a[32];
#pragma unroll
            for (int k0 = 0; k0 < 64; k0 += 16) {
                uint32_t v[16];
#pragma unroll
                for (int b = 0; b < 16; ++b) {
                    const uint32_t k = static_cast<uint32_t>(k0 + b);
                    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[b]) : "r"(ds + k * 128u + ((dy_chunk ^ (k & 7u)) << 4)));
                }
#pragma unroll
                for (int b = 0; b < 16; b += 2) {
                    const float x0 = __uint_as_float(v[b]), x1 = __uint_as_float(v[b + 1]);
                    dbsum += x0 + x1;
                    const __half2 hh = __floats2half2_rn(x0 * sdy, x1 * sdy);
                    __half2 outv = hh;
                    if (part == 1) {
                        const float2 hf = __half22float2(hh);
                        outv = __floats2half2_rn(x0 * sdy - hf.x, x1 * sdy - hf.y);
                    }
                    a[(k0 + b) >> 1] = *reinterpret_cast<const uint32_t*>(&outv);        // low half = the even pixel
                }
            }
            tmem_st32(tmem_base + lane_addr + a_col0 + static_cast<uint32_t>(st) * 32u, a);
            tmem_st_wait();
            tc_fence_before_sync();
            fence_proxy_async_smem();
            mbar_arrive(bar_split_full + 8 * st);
            t_dy += clock64() - c2_;
        }
        if (g.dbg && blockIdx.x == 0 && tid == 0 && grp == 0) { g.dbg[5] = w_raw; g.dbg[6] = t_x; g.dbg[7] = t_dy; g.dbg[8] = clock64() - t_begin; }
        if (do_db) atomicAdd(g.db + nb * 64 + n, dbsum);
        if (grp == 0) {
        mbar_wait(bar_acc, 0u);             // every MMA has completed: the stage buffers are free as well
        tc_fence_after_sync();

        // ---- epilogue: dwp[tap][n][c] += (D[n][c] + D[64 + n][c]) / (s_x * s_dy); the lo-row warps hand their sums over
        // through shared memory, so a CTA issues 64 x 16 vector reductions per tap
        // 1 / (s_x * s_dy) = 2^-(278 - pmax_e), in two factors (the exponent can exceed what one float holds)
        const int tsh = pmax_e == 0 ? 0 : 278 - pmax_e;
        const float un1 = w16_pow2(127 - tsh / 2), un2 = w16_pow2(127 - (tsh - tsh / 2));
        for (int tl = 0; tl < ntap; ++tl) {
            float acc[64];
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tmem_base + lane_addr + tl * 64 + c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 32; ++k) acc[c0 + k] = __uint_as_float(v[k]);
            }
            const uint32_t sc = base + static_cast<uint32_t>(tl & 1) * 16384u + static_cast<uint32_t>(n) * 256u;
            if (part == 1) {
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sc + ((static_cast<uint32_t>(k) ^ (n & 15)) << 4)),
                                 "f"(acc[4 * k]), "f"(acc[4 * k + 1]), "f"(acc[4 * k + 2]), "f"(acc[4 * k + 3]) : "memory");
            }
            asm volatile("bar.sync 1, 128;" ::: "memory");
            if (part == 0) {
                float* dst = g.dwp + (static_cast<long>(tap0 + tl) * g.N + nb * 64 + n) * g.Ctot + cg * 64;
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    float4 o;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
                                 : "r"(sc + ((static_cast<uint32_t>(k) ^ (n & 15)) << 4)));
                    w16_red_add_v4(dst + 4 * k, (acc[4 * k] + o.x) * un1 * un2, (acc[4 * k + 1] + o.y) * un1 * un2,
                                   (acc[4 * k + 2] + o.z) * un1 * un2, (acc[4 * k + 3] + o.w) * un1 * un2);
                }
            }
        }
        tc_fence_before_sync();
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

typedef CUresult (*EncodeTiledFnW)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnW w16_encode_fn() {
    static EncodeTiledFnW fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnW>(p);
    });
    return fn;
}

// 4-D map over NHWC fp32 (C, W, H, B), SWIZZLE_128B; box = (32 channels, bw pixels, bh rows, 1); out-of-bounds = 0 (the padding)
static int w16_make_map(CUtensorMap* tm, const float* x, int ld, int C, int B, int H, int W, int bw, int bh) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, static_cast<cuuint64_t>(W) * ld * 4,
                                   static_cast<cuuint64_t>(H) * W * ld * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(bw), static_cast<cuuint32_t>(bh), 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = w16_encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(wgrad f16) failed: %d", static_cast<int>(r));
    return 0;
}

bool wgrad_f16_supported(int c0, int c1, int N, int H, int W, int R) {
    if (R != 1 && R != 3 && R != 5) return false;
    if (c0 % 64 != 0 || c1 % 64 != 0 || N % 64 != 0) return false;
    if (W % kW16TW != 0 || H % kW16TH != 0) return false;
    return true;
}

int launch_wgrad_f16(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy, int B,
                     int H, int W, int R, int N, float* dwp, float* db, float* scratch2, const float* amax_x, const float* amax_dy,
                     cudaStream_t s) {
    if (!w16_encode_fn()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    // ---- per-tensor maxima -> scratch2 (unless the caller has per-image upper bounds already)
    const long P = static_cast<long>(B) * H * W;
    if (!amax_x) {
        cudaMemsetAsync(scratch2, 0, 2 * sizeof(float), s);
        const long work = P * ((c0 + c1 + N) / 4);
        const int blocks = static_cast<int>(std::min<long>((work + 1023) / 1024, 8L * kSMs));
        launch_pdl(absmax2_kernel, blocks, 256, 0, s, x0, ld0, c0, x1, ld1, c1, dy, ldy, N, P, scratch2);
        if (int rc = launched("absmax2_kernel")) return rc;
    }
    Wgrad16Geom g;
    g.B = B; g.H = H; g.W = W; g.R = R;
    g.Ctot = c0 + c1; g.N = N; g.c0 = c0;
    const int pad = R / 2;
    const int Pp = kW16TW + 2 * pad;
    const int hrows = (kW16TH + 2 * pad) * Pp;
    g.xplane = (hrows * 128 + 1023) / 1024 * 1024;
    const int taps = R * R;
    // taps x 64 accumulator columns + stages x 32 operand columns <= 512: up to 6 taps per CTA.  3x3: 5 + 4 instead of 3 + 3 + 3
    // -- a tile's load and fp16 conversion then feed 40 / 32 MMAs instead of 24 and are repeated by 2 CTAs instead of 3
    static const int taps3 = getenv("ODEX_W16_TAPS3") ? atoi(getenv("ODEX_W16_TAPS3")) : 5;
    g.taps_per_group = R == 3 ? std::max(1, std::min(6, taps3)) : R;
    g.TG = (taps + g.taps_per_group - 1) / g.taps_per_group;
    g.CG = g.Ctot / 64;
    g.NB = N / 64;
    g.ntiles = B * (H / kW16TH) * (W / kW16TW);
    const int units = g.TG * g.CG * g.NB;
    g.PS = std::max(1, std::min(g.ntiles, kSMs / units));
    g.dwp = dwp; g.db = db;
    g.amax_x = amax_x ? amax_x : scratch2;
    g.amax_dy = amax_x ? amax_dy : scratch2 + 1;
    g.amax_stride = amax_x ? 1 : 0;
    const int stages = R <= 3 ? 5 : 4;
    const int smem = 1024 + std::max(stages * (2 * g.xplane + 2 * kW16DyPlane), 32768) + 256;
    if (smem > 232448) return fail(ODEX_E_UNSUPPORTED, "wgrad(f16): needs %d bytes of shared memory", smem);

    CUtensorMap tmX0, tmX1, tmDY;
    if (int rc = w16_make_map(&tmX0, x0, ld0, c0, B, H, W, Pp, kW16TH + 2 * pad)) return rc;
    if (x1) {
        if (int rc = w16_make_map(&tmX1, x1, ld1, c1, B, H, W, Pp, kW16TH + 2 * pad)) return rc;
    } else {
        tmX1 = tmX0;
    }
    if (int rc = w16_make_map(&tmDY, dy, ldy, N, B, H, W, kW16TW, kW16TH)) return rc;

    auto kern = R == 1 ? wgrad_f16_kernel<1> : (R == 3 ? wgrad_f16_kernel<3> : wgrad_f16_kernel<5>);
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] {
        attr_err = cudaFuncSetAttribute(wgrad_f16_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        if (attr_err == cudaSuccess) attr_err = cudaFuncSetAttribute(wgrad_f16_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        if (attr_err == cudaSuccess) attr_err = cudaFuncSetAttribute(wgrad_f16_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "wgrad(f16): cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    g.dbg = nullptr;
    static const bool debug = getenv("ODEX_WGRAD_DEBUG") != nullptr;
    if (debug) {
        static long long* dbuf = nullptr;
        if (!dbuf) cudaMalloc(&dbuf, 16 * sizeof(long long));
        cudaMemsetAsync(dbuf, 0, 16 * sizeof(long long), s);
        g.dbg = dbuf;
    }
    launch_pdl(kern, units * g.PS, 320, smem, s, tmX0, tmX1, tmDY, g);
    if (debug) {
        long long h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, g.dbg, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[wgrad_f16 R=%d C=%d N=%d B=%d grid=%d tiles/cta=%lld] producer: wait_empty=%lld total=%lld | mma: wait_split=%lld total=%lld | "
                "transform group 0 (every other tile): wait_raw=%lld x=%lld dy=%lld total=%lld (cycles)\n", R, g.Ctot, N, B, units * g.PS, h[2], h[0],
                h[1], h[3], h[4], h[5], h[6], h[7], h[8]);
    }
    return launched("wgrad_f16_kernel");
}

}  // namespace odex

using namespace odex;

extern "C" int odex_conv2d_wgrad_f16(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy,
                                     int B, int H, int W, int R, int N, float* dwp, float* db, float* scratch2, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!x0 || !dy || !dwp || !scratch2) return fail(ODEX_E_ARG, "wgrad_f16: null pointer");
    if (!aligned16(x0) || !aligned16(dy) || (x1 && !aligned16(x1)) || ld0 % 4 || ldy % 4 || (x1 && ld1 % 4))
        return fail(ODEX_E_ALIGN, "wgrad_f16: pointers / strides must be 16-byte aligned");
    if (B <= 0 || (x1 == nullptr) != (c1 == 0)) return fail(ODEX_E_ARG, "wgrad_f16: bad batch / second source");
    if (!wgrad_f16_supported(c0, c1, N, H, W, R))
        return fail(ODEX_E_UNSUPPORTED, "wgrad_f16: shape not supported (c0=%d c1=%d N=%d HxW=%dx%d R=%d)", c0, c1, N, H, W, R);
    return launch_wgrad_f16(x0, ld0, c0, x1, ld1, c1, dy, ldy, B, H, W, R, N, dwp, db, scratch2, nullptr, nullptr, s);
}

extern "C" int odex_conv2d_wgrad_f16_scaled(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy,
                                            int ldy, int B, int H, int W, int R, int N, float* dwp, float* db,
                                            const float* amax_x, const float* amax_dy, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!x0 || !dy || !dwp || !amax_x || !amax_dy) return fail(ODEX_E_ARG, "wgrad_f16_scaled: null pointer");
    if (!aligned16(x0) || !aligned16(dy) || (x1 && !aligned16(x1)) || ld0 % 4 || ldy % 4 || (x1 && ld1 % 4))
        return fail(ODEX_E_ALIGN, "wgrad_f16_scaled: pointers / strides must be 16-byte aligned");
    if (B <= 0 || (x1 == nullptr) != (c1 == 0)) return fail(ODEX_E_ARG, "wgrad_f16_scaled: bad batch / second source");
    if (!wgrad_f16_supported(c0, c1, N, H, W, R))
        return fail(ODEX_E_UNSUPPORTED, "wgrad_f16_scaled: shape not supported (c0=%d c1=%d N=%d HxW=%dx%d R=%d)", c0, c1, N, H, W, R);
    return launch_wgrad_f16(x0, ld0, c0, x1, ld1, c1, dy, ldy, B, H, W, R, N, dwp, db, nullptr, amax_x, amax_dy, s);
}
