// Image-resident convolution for 16x16 latents on tcgen05 kind::f16 (two-term fp16 operand split, f16_split.cuh) with the
// ConvGRU gate math fused into its epilogue -- north_star (c), SURVEY.md 2.4 "K2" (C-ABI odex_conv16, include/odex.h).
//
//   out[p, n] = epi( bias[n] + sum_{src, r, s, c} X_src[p + (r, s) - pad, c] * W[n, src*64 + c, r, s] ),   R = 3 or 5,
// X the channel concatenation of one or two 64-channel NHWC sources (torch.cat((x, h), 1) folded in as two K ranges).
// One CTA = one image x one block of 64 output channels.  At 16x16 the 256-pixel x 64-channel accumulator tile holds two whole
// GroupNorm groups (32 channels x all pixels), so the statistics never leave the CTA:
//   PLAIN      v = act(acc + bias) [* act'(aux) | aux + alpha * v]                              (odex_conv2d semantics)
//   GRU_GATES  g = acc + bias; z / r = sigmoid(GroupNorm(g)); block < C: z out; block >= C: r * h out   (modules/ConvGRUCell.py:72-77)
//   GRU_OUT    c = tanh(GroupNorm(acc + bias)); h' = (1 - z) * h + z * c                                 (modules/ConvGRUCell.py:79-82)
// which replaces conv -> HBM -> gru_gates_fwd_kernel -> conv -> HBM -> gru_out_fwd_kernel (4 launches, 2 activation round trips)
// by two launches whose only global traffic is the operands and the results.
//
// Layout / roles follow convnet_f16.cu: the image is converted once per source into fp16 hi / lo halo planes ((16+2p)^2 slots
// of 128 B, zero border, SWIZZLE_128B by absolute address), tap (r, s) = start address of the K-major A descriptor, weights
// [w_hi ; w_lo] of this CTA's 64 output channels stream per (source, tap) through a 3-stage TMA ring.  The two sources use the
// SAME pair of planes one after the other (both must carry the same power-of-two scale, taken from the larger of their maxima):
// source 1 is TMA-loaded into the fp32 staging tile S while source 0's MMAs run and converted once they have retired.
//   warp 0  TMA producer (source 0 tile, weight ring)      warp 1  TMEM allocator + MMA issue
//   warps 2-9  conversion (+ source-1 / aux TMA by its thread 0), then the epilogue: two groups of four warps, one per 32-channel
//              half of the block's 64 outputs (= one GroupNorm group), so that every scheduler holds two epilogue warps
// GroupNorm statistics: per-thread chunk mean / M2 -> Chan merges (warp shuffles, then 4 warps through shared memory): the
// centred sum of squares, not E[x^2] - mean^2; a second pass over TMEM applies them.
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"
#include "f16_split.cuh"

namespace odex {

constexpr int kC16SB = 3;
constexpr int kC16BStage = 2 * 64 * 128;         // [w_hi ; w_lo] rows of this CTA's 64 output channels, one (source, tap)

template <int R>
struct C16Cfg {
    static constexpr int pad = R / 2;
    static constexpr int P = 16 + 2 * pad;
    static constexpr int taps = R * R;
    static constexpr int plane = (P * P * 128 + 1023) / 1024 * 1024;
    static constexpr int smem = 1024 + 2 * plane + 2 * kF16SHalf + kC16SB * kC16BStage + 1024;
};

struct C16Maps {
    CUtensorMap x0, x1;             // fp32 [B][16][16][64], box [32][16][16][1]
    CUtensorMap w;                  // packed fp16 weights, rows of 128 B, box [64][64]
    CUtensorMap aux, out;           // PLAIN: mask / residual operand and result (channel coordinate = n_off + 32 h)
    CUtensorMap graw, omain, ocand; // GRU modes: the pre-norm convolution, z | r*h | h', the candidate
};

constexpr int kC16EpiThreads = 256;      // 8 epilogue warps: two groups of four, one per 32-channel half (= one GroupNorm group)
constexpr int kC16Threads = 64 + kC16EpiThreads;

struct C16Args {
    int nsrc, N, mode;
    const float* x1; int ld1;       // source 1 again as a plain pointer (its max is read directly)
    const float* bias;
    const float* wtail;
    int act, aux_kind;
    float alpha; const double* alpha_dev;
    const float* gamma; const float* beta; float eps;
    const float* h; int ldh; const float* z_in;
    float* z; float* rh; float* hnew; int ld_hnew; float* cand; float* g_raw; float* stats;
    const int* skip;
    float* amax_out;
    int out_split;                  // PLAIN: channels >= out_split are stored through C16Maps::ocand at column c - out_split (0: off)
    long long* dbg;
};

// sigmoid / tanh from ex2.approx (max rel. error 2^-22) and rcp.approx (1 ulp): ~3e-7 relative on the sigmoid, 6e-7 absolute on
// the tanh -- fp32-rounding level, far inside the 1e-5 bar (tests: test_conv16_fused_gru_step_*), at 6 instructions instead of
// the ~30 of 1 / (1 + expf(-x)) / tanhf: with 128 values per thread the gate math was the longest part of the epilogue.
__device__ __forceinline__ float sigmoid_fast(float x) {
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(-1.4426950408889634f * x));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(1.f + e));
    return r;                                   // x -> -inf: e = inf, r = 0;  x -> +inf: e = 0, r = 1;  NaN propagates
}
__device__ __forceinline__ float tanh_fast(float x) { return fmaf(2.f, sigmoid_fast(2.f * x), -1.f); }

// fp32 tile in S (TMA box layout) -> hi / lo planes, this thread's two pixels, channels [32 hh, +32)
template <int R>
__device__ __forceinline__ void c16_convert(uint32_t s_base, uint32_t base, int m, int hh, float scale) {
    using Cfg = C16Cfg<R>;
#pragma unroll 1
    for (int mt = 0; mt < 2; ++mt) {
        const int y = m >> 3, x = mt * 8 + (m & 7);
        const uint32_t p = static_cast<uint32_t>(y * 16 + x);
        const uint32_t srow = static_cast<uint32_t>((y + Cfg::pad) * Cfg::P + x + Cfg::pad);
        const uint32_t row = base + srow * 128u;
        float4 xv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j)
            xv[j] = lds128f(s_base + static_cast<uint32_t>(hh) * kF16SHalf + p * 128u + ((static_cast<uint32_t>(j) ^ (p & 7u)) << 4));
#pragma unroll
        for (int c8 = 0; c8 < 4; ++c8) {
            const float4 u = xv[2 * c8], w = xv[2 * c8 + 1];
            const float o[8] = {u.x * scale, u.y * scale, u.z * scale, u.w * scale, w.x * scale, w.y * scale, w.z * scale, w.w * scale};
            uint4 hi, lo;
            f16_split8(o, &hi, &lo);
            const uint32_t off = (static_cast<uint32_t>(4 * hh + c8) ^ (srow & 7u)) << 4;
            sts128(row + off, hi);
            sts128(row + Cfg::plane + off, lo);
        }
    }
}

__device__ __forceinline__ float absmax4(float m, float4 v) {
    return fmaxf(m, fmaxf(fmaxf(fabsf(v.x), fabsf(v.y)), fmaxf(fabsf(v.z), fabsf(v.w))));
}

// acc + bias of the 32 channels [32 hh, +32) of this thread's pixel in M-tile mt
__device__ __forceinline__ void c16_load32(uint32_t tmem_base, uint32_t lane_addr, int mt, int hh, const float* bias_sm, float unscale,
                                           float (&v)[32]) {
    uint32_t vm[32], vc[32];
    tmem_ld32(tmem_base + lane_addr + mt * 128 + hh * 32, vm);
    tmem_ld32(tmem_base + lane_addr + mt * 128 + 64 + hh * 32, vc);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j)
        v[j] = fmaf(fmaf(__uint_as_float(vc[j]), 4.8828125e-4f, __uint_as_float(vm[j])), unscale, bias_sm[hh * 32 + j]);
}

template <int R>
__global__ void __launch_bounds__(kC16Threads, 1)
conv16_f16_kernel(const __grid_constant__ C16Maps tm, const C16Args a) {
    using Cfg = C16Cfg<R>;
    constexpr int taps = Cfg::taps;
    constexpr uint32_t kTmemCols = 256;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // hi plane; lo plane at + Cfg::plane
    const uint32_t s_base = base + 2u * Cfg::plane;
    const uint32_t bst0 = s_base + 2u * kF16SHalf;
    const uint32_t bars = bst0 + kC16SB * kC16BStage;
    const uint32_t bar_in = bars;                  // [2] source tile landed in S
    const uint32_t bar_b_full = bars + 16;         // [3]
    const uint32_t bar_b_empty = bars + 40;        // [3]
    const uint32_t bar_next = bars + 64;           // operands of a source are in the planes (128 arrivals, one phase per source)
    const uint32_t bar_mma = bars + 72;            // the MMAs of a source have retired (one phase per source)
    const uint32_t bar_aux = bars + 80;
    const uint32_t tmem_slot = bars + 88;
    const uint32_t amax_s = bars + 96;
    const uint32_t gn_s = bars + 128;              // [4 warps][2 groups][mean, M2]
    const uint32_t bias_s = bars + 256;            // [64] bias, [64] gamma, [64] beta

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int NB = a.N >> 6;
    const int img = blockIdx.x / NB;
    const int n_off = (blockIdx.x % NB) * 64;
    const int nsrc = a.nsrc;

    if (threadIdx.x == 0) {
        mbar_init(bar_in, 1); mbar_init(bar_in + 8, 1);
        for (int i = 0; i < kC16SB; ++i) { mbar_init(bar_b_full + 8 * i, 1); mbar_init(bar_b_empty + 8 * i, 1); }
        mbar_init(bar_next, kC16EpiThreads);
        mbar_init(bar_mma, 1);
        mbar_init(bar_aux, 1);
        asm volatile("st.shared.u32 [%0], %1;" ::"r"(amax_s), "r"(0u) : "memory");
        fence_barrier_init();
    }
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm.x0);
        tma_prefetch_desc(&tm.w);
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

    pdl_wait();
    pdl_trigger();
    const bool skip = a.skip != nullptr && *a.skip != 0;
    const long long t_start = clock64();
#define C16_MARK(i) do { if (a.dbg && blockIdx.x == 0 && threadIdx.x == 64) a.dbg[i] = clock64() - t_start; } while (0)

    if (skip) {
    } else if (warp == 0) {
        // ===================== TMA producer: source 0, then the weight ring =====================
        if (lane == 0) {
            mbar_arrive_expect_tx(bar_in, 2u * kF16SHalf);
            tma_load_4d(s_base, &tm.x0, bar_in, 0, 0, 0, img);
            tma_load_4d(s_base + kF16SHalf, &tm.x0, bar_in, 32, 0, 0, img);
            int sb = 0; uint32_t pb = 0;
            for (int kb = 0; kb < nsrc * taps; ++kb) {
                mbar_wait(bar_b_empty + 8 * sb, pb ^ 1u);
                mbar_arrive_expect_tx(bar_b_full + 8 * sb, kC16BStage);
                const uint32_t dst = bst0 + sb * kC16BStage;
                tma_load_2d(dst, &tm.w, bar_b_full + 8 * sb, 0, kb * 2 * a.N + n_off);                  // w_hi rows of this block
                tma_load_2d(dst + kC16BStage / 2, &tm.w, bar_b_full + 8 * sb, 0, kb * 2 * a.N + a.N + n_off);   // w_lo rows
                if (++sb == kC16SB) { sb = 0; pb ^= 1u; }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issue =====================
        constexpr uint32_t idesc_wide = umma_idesc_f16(128, 128);
        constexpr uint32_t idesc_narrow = umma_idesc_f16(128, 64);
        const uint32_t a_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, Cfg::P * 128) >> 32);
        const uint32_t b_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024) >> 32);
        constexpr uint32_t lbo16 = 1u << 16;
        constexpr uint32_t lo_delta = static_cast<uint32_t>(Cfg::plane) >> 4;
        int sb = 0; uint32_t pb = 0;
        for (int src = 0; src < nsrc; ++src) {
            mbar_wait(bar_next, static_cast<uint32_t>(src) & 1u);
            tc_fence_after_sync();
            int r = 0, s = 0;
            for (int tap = 0; tap < taps; ++tap) {
                mbar_wait(bar_b_full + 8 * sb, pb);
                tc_fence_after_sync();
                const uint32_t b_lo0 = lbo16 | ((bst0 + sb * kC16BStage) >> 4);
                const uint32_t a_lo0 = lbo16 | ((base >> 4) + static_cast<uint32_t>(r * Cfg::P + s) * 8u);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
#pragma unroll
                        for (int mt = 0; mt < 2; ++mt) {
                            const uint32_t a_lo = a_lo0 + mt * 64u + k * 2u;
                            const uint32_t d = tmem_base + mt * 128;
                            umma_f16_ss_w(d, a_lo, a_hi32, b_lo0 + k * 2u, b_hi32, idesc_wide, (src | tap | k) != 0);
                            umma_f16_ss_w(d + 64, a_lo + lo_delta, a_hi32, b_lo0 + k * 2u, b_hi32, idesc_narrow, 1u);
                        }
                    }
                    umma_commit(bar_b_empty + 8 * sb);
                }
                __syncwarp();
                if (++sb == kC16SB) { sb = 0; pb ^= 1u; }
                if (++s == R) { s = 0; ++r; }
            }
            if (elect_one()) umma_commit(bar_mma);
            __syncwarp();
        }
    } else {
        // ===================== conversion, then the epilogue =====================
        const int tid = threadIdx.x - 64;
        const bool gn = a.mode != ODEX_C16_PLAIN;
        if (tid < 16) {
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a.bias) bv = __ldg(reinterpret_cast<const float4*>(a.bias + n_off) + tid);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(bias_s + 16u * tid), "f"(bv.x), "f"(bv.y), "f"(bv.z), "f"(bv.w) : "memory");
        } else if (gn && tid < 48) {
            const int j = tid & 15;
            const float* src = tid < 32 ? a.gamma : a.beta;
            const float4 v = __ldg(reinterpret_cast<const float4*>(src + n_off) + j);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(bias_s + 256u * (tid < 32 ? 1u : 2u) + 16u * j), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
        }
        for (uint32_t i = tid; i < 2u * Cfg::plane / 16u; i += kC16EpiThreads) sts128(base + i * 16u, make_uint4(0u, 0u, 0u, 0u));
        float alpha = a.alpha;
        if (a.alpha_dev) alpha *= static_cast<float>(*a.alpha_dev);

        const int q = warp & 3;
        const int hh_own = tid >> 7;                  // this warp group's 32-channel half of the 64 outputs (= one GroupNorm group)
        const int m = q * 32 + lane;                  // TMEM lane = pixel of the M-tile: (y, x) = (m / 8, mt * 8 + m % 8)
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        const uint32_t p0 = static_cast<uint32_t>((m >> 3) * 16 + (m & 7));          // pixel of M-tile 0; M-tile 1: + 8

        // ---- the scale shared by both sources: max |x| over this image (source 1 read directly, its tile comes later)
        float mx = 0.f;
        if (nsrc == 2) {
            // a max does not care which thread sees which element: fully coalesced 16-byte reads (pixel = i / 16, chunk = i % 16)
            const float* xb = a.x1 + static_cast<long>(img) * 256 * a.ld1;
#pragma unroll 1
            for (int i0_ = 0; i0_ < 4096; i0_ += 8 * kC16EpiThreads) {
                float4 t[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const int i = i0_ + k * kC16EpiThreads + tid;
                    t[k] = __ldg(reinterpret_cast<const float4*>(xb + static_cast<long>(i >> 4) * a.ld1) + (i & 15));
                }
#pragma unroll
                for (int k = 0; k < 8; ++k) mx = absmax4(mx, t[k]);
            }
        }
        mbar_wait(bar_in, 0u);
#pragma unroll 1
        for (int mt = 0; mt < 2; ++mt) {
            const uint32_t p = p0 + 8u * mt;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                mx = absmax4(mx, lds128f(s_base + static_cast<uint32_t>(hh_own) * kF16SHalf + p * 128u + ((static_cast<uint32_t>(j) ^ (p & 7u)) << 4)));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        asm volatile("bar.sync 1, 256;" ::: "memory");          // planes zeroed, bias in place, amax_s initialised
        if (lane == 0) asm volatile("atom.shared.max.u32 _, [%0], %1;" ::"r"(amax_s), "r"(__float_as_uint(mx)) : "memory");
        asm volatile("bar.sync 1, 256;" ::: "memory");
        uint32_t mb;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(mb) : "r"(amax_s));
        float a_inv;
        const float a_scale = f16_scale_for(__uint_as_float(mb), &a_inv);
        if (tid == 0 && a.amax_out) a.amax_out[img] = __uint_as_float(mb);      // (every output-channel block of the image writes the same value)
        C16_MARK(1);

        c16_convert<R>(s_base, base, m, hh_own, a_scale);
        fence_proxy_async_smem();
        mbar_arrive(bar_next);
        asm volatile("bar.sync 1, 256;" ::: "memory");          // S (source 0) has been consumed by every thread
        C16_MARK(2);
        if (nsrc == 2) {
            if (tid == 0) {
                mbar_arrive_expect_tx(bar_in + 8, 2u * kF16SHalf);
                tma_load_4d(s_base, &tm.x1, bar_in + 8, 0, 0, 0, img);
                tma_load_4d(s_base + kF16SHalf, &tm.x1, bar_in + 8, 32, 0, 0, img);
            }
            mbar_wait(bar_in + 8, 0u);
            C16_MARK(3);
            mbar_wait(bar_mma, 0u);                              // source 0's MMAs have read the planes
            C16_MARK(4);
            c16_convert<R>(s_base, base, m, hh_own, a_scale);
            fence_proxy_async_smem();
            mbar_arrive(bar_next);
            asm volatile("bar.sync 1, 256;" ::: "memory");      // S (source 1) consumed
            C16_MARK(5);
        }
        const int kind = a.mode == ODEX_C16_PLAIN ? a.aux_kind : 0;
        if (kind && tid == 0) {
            mbar_arrive_expect_tx(bar_aux, 2u * kF16SHalf);
            tma_load_4d(s_base, &tm.aux, bar_aux, n_off, 0, 0, img);
            tma_load_4d(s_base + kF16SHalf, &tm.aux, bar_aux, n_off + 32, 0, 0, img);
        }
        const float unscale = a_inv * __ldg(a.wtail);
        const float* bias_sm = reinterpret_cast<const float*>(smem_raw + (bias_s - smem_u32(smem_raw)));

        mbar_wait(bar_mma, static_cast<uint32_t>(nsrc - 1) & 1u);
        tc_fence_after_sync();
        C16_MARK(6);

        if (a.mode == ODEX_C16_PLAIN) {
            if (kind) mbar_wait(bar_aux, 0u);
            const float4* bl = reinterpret_cast<const float4*>(bias_sm);
#define ODEX_C16_EPI(A, K) f16_epilogue<A, K, true, true>(tmem_base, lane_addr, base, s_base, m, bl, alpha, unscale, 1.f, hh_own, hh_own + 1)
            switch (a.act * 4 + kind) {
                case ODEX_ACT_NONE * 4 + 0: ODEX_C16_EPI(ODEX_ACT_NONE, 0); break;
                case ODEX_ACT_NONE * 4 + 1: ODEX_C16_EPI(ODEX_ACT_NONE, 1); break;
                case ODEX_ACT_NONE * 4 + 2: ODEX_C16_EPI(ODEX_ACT_NONE, 2); break;
                case ODEX_ACT_NONE * 4 + 3: ODEX_C16_EPI(ODEX_ACT_NONE, 3); break;
                case ODEX_ACT_RELU * 4 + 0: ODEX_C16_EPI(ODEX_ACT_RELU, 0); break;
                case ODEX_ACT_RELU * 4 + 3: ODEX_C16_EPI(ODEX_ACT_RELU, 3); break;
                case ODEX_ACT_TANH * 4 + 0: ODEX_C16_EPI(ODEX_ACT_TANH, 0); break;
                case ODEX_ACT_TANH * 4 + 3: ODEX_C16_EPI(ODEX_ACT_TANH, 3); break;
                default: break;
            }
#undef ODEX_C16_EPI
            tc_fence_before_sync();
            fence_proxy_async_smem();
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (tid == 0) {
                const bool hi_part = a.out_split > 0 && n_off >= a.out_split;
                const CUtensorMap* mo = hi_part ? &tm.ocand : &tm.out;
                const int co = hi_part ? n_off - a.out_split : n_off;
                tma_store_4d(mo, s_base, co, 0, 0, img);
                tma_store_4d(mo, s_base + kF16SHalf, co + 32, 0, 0, img);
                tma_store_commit();
                tma_store_wait_read();
            }
        } else {
            // ---- GroupNorm statistics of the two 32-channel groups of this block: chunk mean / M2, merged pairwise (Chan)
            float gmean_o = 0.f, gM2_o = 0.f;                // this group's (mean, M2)
#pragma unroll 1
            for (int mt = 0; mt < 2; ++mt) {
#pragma unroll 1
                for (int hh = hh_own; hh <= hh_own; ++hh) {
                    float v[32];
                    c16_load32(tmem_base, lane_addr, mt, hh, bias_sm, unscale, v);
                    if (a.g_raw) {              // the pre-norm convolution goes out through S (TMA store below)
                        const uint32_t p = p0 + 8u * mt;
                        const uint32_t row = s_base + static_cast<uint32_t>(hh) * kF16SHalf + p * 128u;
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(row + ((static_cast<uint32_t>(j) ^ (p & 7u)) << 4)),
                                         "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
                    }
                    float sum = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; ++j) sum += v[j];
                    const float cm = sum * (1.f / 32.f);
                    float cq = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; ++j) { const float d = v[j] - cm; cq = fmaf(d, d, cq); }
                    if (mt == 0) { gmean_o = cm; gM2_o = cq; }
                    else { const float d = cm - gmean_o; gmean_o += 0.5f * d; gM2_o += cq + d * d * 16.f; }
                }
            }
            float n_half = 32.f;                        // n / 2 of each side of the next merge (n = 64 per thread)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
#pragma unroll 1
                for (int hh = hh_own; hh <= hh_own; ++hh) {
                    const float mo = __shfl_xor_sync(0xffffffffu, gmean_o, o), qo = __shfl_xor_sync(0xffffffffu, gM2_o, o);
                    const float d = mo - gmean_o;
                    gM2_o = (gM2_o + qo) + d * d * n_half;
                    gmean_o = 0.5f * (gmean_o + mo);
                }
                n_half *= 2.f;
            }
            if (lane == 0)
                asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(gn_s + static_cast<uint32_t>(q * 2 + hh_own) * 8u),
                             "f"(gmean_o), "f"(gM2_o) : "memory");
            if (a.g_raw) fence_proxy_async_smem();
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (a.g_raw && tid == 0) {
                tma_store_4d(&tm.graw, s_base, n_off, 0, 0, img);
                tma_store_4d(&tm.graw, s_base + kF16SHalf, n_off + 32, 0, 0, img);
                tma_store_commit();
            }
            float mean_o = 0.f, rstd_o = 0.f;
#pragma unroll 1
            for (int hh = hh_own; hh <= hh_own; ++hh) {
                float mu, q2;
                asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(mu) , "=f"(q2) : "r"(gn_s + static_cast<uint32_t>(hh) * 8u));
                float n = 2048.f;
#pragma unroll
                for (int w = 1; w < 4; ++w) {
                    float mo, qo;
                    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(mo), "=f"(qo) : "r"(gn_s + static_cast<uint32_t>(w * 2 + hh) * 8u));
                    const float d = mo - mu, nn = n + 2048.f;
                    q2 = (q2 + qo) + d * d * (n * 2048.f / nn);
                    mu += d * (2048.f / nn);
                    n = nn;
                }
                mean_o = mu;
                rstd_o = 1.f / sqrtf(q2 * (1.f / 8192.f) + a.eps);
            }
            const int G = a.N >> 5;
            if ((tid & 127) == 0) {
                float* st = a.stats + (static_cast<long>(img) * G + (n_off >> 5) + hh_own) * 2;
                st[0] = mean_o; st[1] = rstd_o;
            }
            // ---- apply: normalise, gate, blend.  Results are staged in shared memory in the TMA box layout and stored by TMA: the
            // main result (z | r*h | h') in the operand planes, which every MMA has finished reading; the candidate in S once
            // the store of the pre-norm convolution has read it.  h / z are read straight from global memory (L2).
            const float* gam = bias_sm + 64;
            const float* bet = bias_sm + 128;
            const int C = a.mode == ODEX_C16_GRU_GATES ? a.N >> 1 : a.N;
            const bool z_block = a.mode == ODEX_C16_GRU_GATES && n_off < C;
            const bool want_cand = a.mode == ODEX_C16_GRU_OUT && a.cand != nullptr;
            if (want_cand) {
                if (tid == 0) tma_store_wait_read();
                asm volatile("bar.sync 1, 256;" ::: "memory");
            }
#pragma unroll 1
            for (int mt = 0; mt < 2; ++mt) {
                const uint32_t p = p0 + 8u * mt;
                const long gp = static_cast<long>(img) * 256 + p;
#pragma unroll 1
                for (int hh = hh_own; hh <= hh_own; ++hh) {
                    const int cb = n_off + hh * 32;                   // first channel of this group among the N outputs
                    float4 hv[8], zv[8];
                    if (a.mode == ODEX_C16_GRU_OUT) {
                        const float4* hp = reinterpret_cast<const float4*>(a.h + gp * a.ldh + cb);
                        const float4* zp = reinterpret_cast<const float4*>(a.z_in + gp * C + cb);
#pragma unroll
                        for (int j = 0; j < 8; ++j) { hv[j] = __ldg(hp + j); zv[j] = __ldg(zp + j); }
                    } else if (!z_block) {
                        const float4* hp = reinterpret_cast<const float4*>(a.h + gp * a.ldh + (cb - C));
#pragma unroll
                        for (int j = 0; j < 8; ++j) hv[j] = __ldg(hp + j);
                    }
                    float v[32];
                    c16_load32(tmem_base, lane_addr, mt, hh, bias_sm, unscale, v);
                    const float mu = mean_o, rs = rstd_o;
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = fmaf((v[j] - mu) * rs, gam[hh * 32 + j], bet[hh * 32 + j]);
                    const uint32_t t_row = base + static_cast<uint32_t>(hh) * kF16SHalf + p * 128u;
                    const uint32_t s_row = s_base + static_cast<uint32_t>(hh) * kF16SHalf + p * 128u;
                    if (a.mode == ODEX_C16_GRU_GATES) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = sigmoid_fast(v[j]);
                        if (!z_block) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) { v[4 * j] *= hv[j].x; v[4 * j + 1] *= hv[j].y; v[4 * j + 2] *= hv[j].z; v[4 * j + 3] *= hv[j].w; }
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = tanh_fast(v[j]);
                        if (want_cand) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(s_row + ((static_cast<uint32_t>(j) ^ (p & 7u)) << 4)),
                                             "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            v[4 * j] = (1.f - zv[j].x) * hv[j].x + zv[j].x * v[4 * j];
                            v[4 * j + 1] = (1.f - zv[j].y) * hv[j].y + zv[j].y * v[4 * j + 1];
                            v[4 * j + 2] = (1.f - zv[j].z) * hv[j].z + zv[j].z * v[4 * j + 2];
                            v[4 * j + 3] = (1.f - zv[j].w) * hv[j].w + zv[j].w * v[4 * j + 3];
                        }
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(t_row + ((static_cast<uint32_t>(j) ^ (p & 7u)) << 4)),
                                     "f"(v[4 * j]), "f"(v[4 * j + 1]), "f"(v[4 * j + 2]), "f"(v[4 * j + 3]) : "memory");
                }
            }
            tc_fence_before_sync();
            fence_proxy_async_smem();
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (tid == 0) {
                const bool r_block = a.mode == ODEX_C16_GRU_GATES && !z_block;
                const int c_main = r_block ? n_off - C : n_off;
                const CUtensorMap* mm = r_block ? &tm.ocand : &tm.omain;      // gates: ocand is the map over r*h
                tma_store_4d(mm, base, c_main, 0, 0, img);
                tma_store_4d(mm, base + kF16SHalf, c_main + 32, 0, 0, img);
                if (want_cand) {
                    tma_store_4d(&tm.ocand, s_base, n_off, 0, 0, img);
                    tma_store_4d(&tm.ocand, s_base + kF16SHalf, n_off + 32, 0, 0, img);
                }
                tma_store_commit();
                tma_store_wait_read();
            }
            tc_fence_before_sync();
        }
    }
    C16_MARK(7);
#undef C16_MARK
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

typedef CUresult (*EncodeTiledFnC)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnC c16_encode_fn() {
    static EncodeTiledFnC fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnC>(p);
    });
    return fn;
}

// [B][16][16][C] NHWC fp32 view with pixel-row stride ld; box = 32 channels x 16 x 16 pixels (one half of S)
static int c16_act_map(CUtensorMap* tm, const float* x, int ld, int C, int B) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), 16, 16, static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, 16ull * ld * 4, 256ull * ld * 4};
    const cuuint32_t box[4] = {32, 16, 16, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = c16_encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(conv16 activation) failed: %d", static_cast<int>(r));
    return 0;
}

static const char* c16_check(const odex_conv16_desc* d) {
    if (!d || !d->x0 || !d->w_f16) return "null descriptor / x0 / w_f16";
    if (d->B <= 0 || (d->R != 3 && d->R != 5) || d->N <= 0 || d->N % 64 != 0) return "needs R in {3,5} and N a multiple of 64";
    if (d->ld0 % 4 || d->ld0 < 64 || !aligned16(d->x0) || !aligned16(d->w_f16)) return "x0 / w_f16 alignment or stride";
    if (d->x1 && (d->ld1 % 4 || d->ld1 < 64 || !aligned16(d->x1))) return "x1 alignment or stride";
    if (d->bias && !aligned16(d->bias)) return "bias alignment";
    if (d->mode == ODEX_C16_PLAIN) {
        if (!d->out || d->ldo % 4 || d->ldo < (d->out_split > 0 ? d->out_split : d->N) || !aligned16(d->out)) return "out";
        if (d->act < 0 || d->act > 2 || d->aux_kind < 0 || d->aux_kind > 3) return "act / aux_kind";
        if (d->act != ODEX_ACT_NONE && (d->aux_kind == 1 || d->aux_kind == 2)) return "activation together with a mask";
        if (d->aux_kind && (!d->aux || d->ld_aux % 4 || d->ld_aux < d->N || !aligned16(d->aux))) return "aux";
        return nullptr;
    }
    if (d->mode != ODEX_C16_GRU_GATES && d->mode != ODEX_C16_GRU_OUT) return "unknown mode";
    if (!d->gamma || !d->beta || !d->stats || !d->h || d->ldh % 4 || !aligned16(d->h) || !aligned16(d->gamma) || !aligned16(d->beta))
        return "GroupNorm parameters / h / stats";
    if (d->g_raw && !aligned16(d->g_raw)) return "g_raw alignment";
    if (d->mode == ODEX_C16_GRU_GATES) {
        if (d->N % 128 != 0 || !d->z || !d->rh || !aligned16(d->z) || !aligned16(d->rh) || d->ldh < d->N / 2) return "gates: N = 2C with C % 64 == 0, z, rh";
    } else {
        if (!d->z_in || !d->hnew || d->ld_hnew % 4 || d->ld_hnew < d->N || d->ldh < d->N || !aligned16(d->z_in) || !aligned16(d->hnew) ||
            (d->cand && !aligned16(d->cand)))
            return "out: z_in, hnew";
    }
    return nullptr;
}

template <int R>
static int c16_launch(const odex_conv16_desc* d, cudaStream_t s) {
    using Cfg = C16Cfg<R>;
    C16Maps tm;
    C16Args a;
    const int nsrc = d->x1 ? 2 : 1;
    if (int rc = c16_act_map(&tm.x0, d->x0, d->ld0, 64, d->B)) return rc;
    tm.x1 = tm.x0; tm.aux = tm.x0; tm.out = tm.x0;
    if (d->x1)
        if (int rc = c16_act_map(&tm.x1, d->x1, d->ld1, 64, d->B)) return rc;
    {
        const cuuint64_t rows = static_cast<cuuint64_t>(nsrc) * Cfg::taps * 2 * d->N;
        const cuuint64_t dims[2] = {64, rows};
        const cuuint64_t strides[1] = {128};
        const cuuint32_t box[2] = {64, 64};
        const cuuint32_t es[2] = {1, 1};
        CUresult r = c16_encode_fn()(&tm.w, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<float*>(d->w_f16), dims, strides, box, es,
                                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(conv16 weights) failed: %d", static_cast<int>(r));
    }
    tm.graw = tm.x0; tm.omain = tm.x0; tm.ocand = tm.x0;
    if (d->mode == ODEX_C16_PLAIN) {
        if (d->aux_kind)
            if (int rc = c16_act_map(&tm.aux, d->aux, d->ld_aux, d->N, d->B)) return rc;
        if (d->out_split > 0) {
            if (d->out_split % 64 || d->out_split >= d->N || !d->out_hi || d->ld_out_hi % 4 || d->ld_out_hi < d->N - d->out_split ||
                !aligned16(d->out_hi) || d->ldo < d->out_split)
                return fail(ODEX_E_ARG, "conv16: bad out_split / out_hi (split=%d N=%d)", d->out_split, d->N);
            if (int rc = c16_act_map(&tm.out, d->out, d->ldo, d->out_split, d->B)) return rc;
            if (int rc = c16_act_map(&tm.ocand, d->out_hi, d->ld_out_hi, d->N - d->out_split, d->B)) return rc;
        } else {
            if (int rc = c16_act_map(&tm.out, d->out, d->ldo, d->N, d->B)) return rc;
        }
    } else {
        const int C = d->mode == ODEX_C16_GRU_GATES ? d->N / 2 : d->N;
        if (d->g_raw)
            if (int rc = c16_act_map(&tm.graw, d->g_raw, d->N, d->N, d->B)) return rc;
        if (d->mode == ODEX_C16_GRU_GATES) {
            // one launch covers both kinds of block: the map of a z block is over z, that of an r block over r*h.  Two maps would
            // need a per-block choice in the kernel; z and r*h have the same shape, so the kernel picks the base by block instead
            if (int rc = c16_act_map(&tm.omain, d->z, C, C, d->B)) return rc;
            if (int rc = c16_act_map(&tm.ocand, d->rh, C, C, d->B)) return rc;
        } else {
            if (int rc = c16_act_map(&tm.omain, d->hnew, d->ld_hnew, C, d->B)) return rc;
            if (d->cand)
                if (int rc = c16_act_map(&tm.ocand, d->cand, C, C, d->B)) return rc;
        }
    }
    a.nsrc = nsrc; a.N = d->N; a.mode = d->mode;
    a.x1 = d->x1; a.ld1 = d->ld1;
    a.bias = d->bias;
    a.wtail = d->w_f16 + static_cast<long>(d->N) * 64 * nsrc * R * R;
    a.act = d->act; a.aux_kind = d->aux_kind; a.alpha = d->alpha; a.alpha_dev = d->alpha_dev;
    a.gamma = d->gamma; a.beta = d->beta; a.eps = d->eps;
    a.h = d->h; a.ldh = d->ldh; a.z_in = d->z_in;
    a.z = d->z; a.rh = d->rh; a.hnew = d->hnew; a.ld_hnew = d->ld_hnew; a.cand = d->cand; a.g_raw = d->g_raw; a.stats = d->stats;
    a.skip = d->skip;
    a.amax_out = d->amax_out;
    a.out_split = d->mode == ODEX_C16_PLAIN ? d->out_split : 0;
    a.dbg = nullptr;
    static const bool debug = getenv("ODEX_CONV_DEBUG") != nullptr;
    if (debug) {
        static long long* dbuf = nullptr;
        if (!dbuf) cudaMalloc(&dbuf, 16 * sizeof(long long));
        cudaMemsetAsync(dbuf, 0, 16 * sizeof(long long), s);
        a.dbg = dbuf;
    }
    auto kern = conv16_f16_kernel<R>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] { attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::smem); });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "conv16: cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    launch_pdl(kern, d->B * (d->N / 64), kC16Threads, Cfg::smem, s, tm, a);
    if (debug) {
        long long h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, a.dbg, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[conv16 R=%d nsrc=%d N=%d mode=%d B=%d] cta0 cycles: scale %lld | src0 converted %lld | src1 landed %lld | mma0 done %lld | src1 converted %lld | mma done %lld | end %lld\n",
                R, nsrc, d->N, d->mode, d->B, h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
    }
    return launched("conv16_f16_kernel");
}

}  // namespace odex

using namespace odex;

extern "C" int odex_conv16_supported(const odex_conv16_desc* d) { return c16_check(d) == nullptr ? 1 : 0; }

extern "C" int odex_conv16(const odex_conv16_desc* d, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (const char* why = c16_check(d)) return fail(ODEX_E_UNSUPPORTED, "conv16: %s", why);
    if (!c16_encode_fn()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    return d->R == 3 ? c16_launch<3>(d, s) : c16_launch<5>(d, s);
}
