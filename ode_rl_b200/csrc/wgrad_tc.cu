// Weight gradient of the stride-1 "same" convolution on the 5th-gen tensor cores (tcgen05, kind::tf32, 3xTF32):
//     dw[tap][c][n] += sum_p X[p + tap][c] * dy[p][n]          (reduction over PIXELS)
// Replaces the autograd conv backward-weight of helpers/utils.py:167-177 (3x3) and modules/ConvGRUCell.py:41-43,47-49
// (5x5) -- 52 % of the training step in the first profile (profiles/r01_launches_v1.md) as an FFMA kernel.
//
// Both GEMM operands are read as they lie in NHWC: a shared-memory row is one PIXEL (32 channels = 128 bytes), so the
// reduction dimension K runs over rows: "MN-major" operands.  For tf32 the only MN-major layout the tensor core takes
// is SWIZZLE_128B_BASE32B (32-byte atoms), which TMA writes with CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B.  The swizzle is a
// function of the absolute shared-memory address (tools/umma_shift_probe.cu, measured on B200), so the tap shift
// (r, s) is nothing but a different START ADDRESS of the X operand inside the halo tile: no im2col, no copies.
//
// 3xTF32: the transform warps split the raw tiles once per tile into tf32 hi / lo planes, in place.  With the planes
// laid out [Xhi c0-31][Xhi c32-63][Xlo c0-31][Xlo c32-63] at a uniform stride, ONE descriptor covers M = 128 rows
// = [Xhi ; Xlo], and per 8-pixel K atom two MMAs (B = dy_hi, then B = dy_lo, N = 64) accumulate
//     lanes 0-63 : Xhi*dyhi + Xhi*dylo        lanes 64-127 : Xlo*dyhi + Xlo*dylo
// into the same 64 TMEM columns of that tap; the epilogue adds the two lane halves (red.global.add.v4.f32).
//
// One CTA = (tap group <= 8 taps, 64-channel group of X, 64-channel block of dy, a strided set of 64-pixel tiles).
//   warp 0      TMA producer: X halo (TH+2p rows x 16+2p pixel slots) + dy tile per stage, 2 stages
//   warp 1      TMEM allocator + single-thread MMA issuer
//   warps 2-5   hi/lo split (in place), bias-gradient column sums, then the epilogue
#include <cuda.h>

#include <algorithm>
#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace odex {

constexpr int kWgTH = 4;            // tile = 4 rows x 16 pixels
constexpr int kWgTW = 16;
constexpr int kWgDyPlane = kWgTH * kWgTW * 128;     // 8 KB per 32-channel plane of dy

struct WgradTcGeom {
    int B, H, W, R;
    int Ctot, N;        // channels of X (all sources) and of dy
    int c0;             // channels of source 0 (groups beyond it come from source 1)
    int P;              // halo pitch in pixel slots = 16 + 2 * pad
    int xplane;         // bytes per 32-channel halo plane, multiple of 1024
    int TG, CG, NB, PS; // grid factors: tap groups, 64-channel groups of X, 64-channel blocks of dy, pixel splits
    int taps_per_group;
    int ntiles;         // B * (H / 4) * (W / 16)
    float* dwp;         // [R*R][Ctot][N], accumulated
    float* db;          // [N] or nullptr, accumulated
    int trunc_hi;       // 1: hi operand = raw word (hardware truncation), see split4
    long long* dbg;     // diagnostic: per-role cycle counters of CTA 0 (ODEX_WGRAD_DEBUG=1), else nullptr
};

// MN-major SWIZZLE_128B_BASE32B descriptor: LBO = stride between 32-element groups along M/N, SBO = stride between
// 4-row groups along K (512 B for dense rows).  layout type 1 in bits [61,64).
__device__ __forceinline__ uint64_t umma_desc_mn32(uint32_t saddr, uint32_t lbo_bytes) {
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr & 0x3FFFF) >> 4);
    d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= static_cast<uint64_t>((512u >> 4) & 0x3FFF) << 32;
    d |= 1ull << 46;
    d |= 1ull << 61;
    return d;
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// trunc != 0: leave the raw fp32 word as the "hi" operand -- the tensor core reads only the top 19 bits of a tf32
// operand, i.e. hi = trunc(x) -- and store lo = tf32(x - trunc(x)); saves one shared-memory store per element.
__device__ __forceinline__ void split4(uint32_t src, uint32_t dst_hi, uint32_t dst_lo, float4& raw, int trunc = 0) {
    uint32_t a, b, c, d;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(src));
    raw = make_float4(__uint_as_float(a), __uint_as_float(b), __uint_as_float(c), __uint_as_float(d));
    const uint32_t x[4] = {a, b, c, d};
    uint32_t hi[4], lo[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        hi[e] = trunc ? (x[e] & 0xFFFFE000u) : tf32_rna_bits(x[e]);
        lo[e] = tf32_rna_bits(__float_as_uint(__uint_as_float(x[e]) - __uint_as_float(hi[e])));
    }
    if (!trunc)
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst_hi), "r"(hi[0]), "r"(hi[1]), "r"(hi[2]), "r"(hi[3]) : "memory");
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst_lo), "r"(lo[0]), "r"(lo[1]), "r"(lo[2]), "r"(lo[3]) : "memory");
}

// Software-pipelined form: NB independent 16-byte loads are issued before the first one is consumed (the inline-asm
// loads / stores are volatile, so the compiler keeps them in program order: without this the loop is one
// load-use latency per 16 bytes and the split, not the tensor core, bounds the kernel).
template <int NB>
__device__ __forceinline__ void split_batch(const uint32_t (&src)[NB], const bool (&ok)[NB], uint32_t lo_delta, int trunc,
                                            float4& sum) {
    uint32_t x[NB][4];
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        x[b][0] = x[b][1] = x[b][2] = x[b][3] = 0u;
        if (ok[b])
            asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(x[b][0]), "=r"(x[b][1]), "=r"(x[b][2]), "=r"(x[b][3]) : "r"(src[b]));
    }
#pragma unroll
    for (int b = 0; b < NB; ++b) {
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            hi[e] = trunc ? (x[b][e] & 0xFFFFE000u) : tf32_rna_bits(x[b][e]);
            lo[e] = tf32_rna_bits(__float_as_uint(__uint_as_float(x[b][e]) - __uint_as_float(hi[e])));
        }
        sum.x += __uint_as_float(x[b][0]); sum.y += __uint_as_float(x[b][1]);
        sum.z += __uint_as_float(x[b][2]); sum.w += __uint_as_float(x[b][3]);
        if (ok[b]) {
            if (!trunc)
                asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(src[b]), "r"(hi[0]), "r"(hi[1]), "r"(hi[2]), "r"(hi[3]) : "memory");
            asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(src[b] + lo_delta), "r"(lo[0]), "r"(lo[1]), "r"(lo[2]), "r"(lo[3]) : "memory");
        }
    }
}

__device__ __forceinline__ void tile_coords(const WgradTcGeom& g, int t, int& img, int& y0, int& x0) {
    const int tx = g.W / kWgTW, ty = g.H / kWgTH;
    x0 = (t % tx) * kWgTW;
    y0 = ((t / tx) % ty) * kWgTH;
    img = t / (tx * ty);
}

template <int R>
__global__ void __launch_bounds__(192, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                const __grid_constant__ CUtensorMap tmDY, const WgradTcGeom g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t stage_bytes = 4u * g.xplane + 4u * kWgDyPlane;
    const uint32_t bars = base + 2u * stage_bytes;
    const uint32_t bar_raw_full = bars;            // [2]
    const uint32_t bar_split_full = bars + 16;     // [2]
    const uint32_t bar_stage_empty = bars + 32;    // [2]
    const uint32_t bar_acc = bars + 48;
    const uint32_t tmem_slot = bars + 56;

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
    const int hrows = (kWgTH + 2 * pad) * g.P;     // pixel slots of one halo plane
    // R <= 3: one N = 128 MMA per K atom with B = [dy_hi | dy_lo] (8 KB of operand reads instead of 2 x 6 KB -- the
    // kernel is bound by shared-memory bandwidth), 128 accumulator columns per tap; R = 5: two N = 64 MMAs into 64 columns
    constexpr bool kWide = R <= 3;
    constexpr int DC = kWide ? 128 : 64;
    uint32_t tmem_cols = 32;
    while (tmem_cols < static_cast<uint32_t>(ntap * DC)) tmem_cols <<= 1;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
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

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int cabs = cg * 64;
            const bool second = cabs >= g.c0;
            const CUtensorMap* tmX = second ? &tmX1 : &tmX0;
            const int cx = second ? cabs - g.c0 : cabs;
            const uint32_t tx_bytes = 2u * static_cast<uint32_t>(hrows) * 128u + 2u * kWgDyPlane;
            int it = 0;
            long long w_empty = 0, t_begin = clock64();
            for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
                const int st = it & 1;
                const uint32_t use = static_cast<uint32_t>(it >> 1);
                const long long c0_ = clock64();
                mbar_wait(bar_stage_empty + 8 * st, (use & 1u) ^ 1u);
                w_empty += clock64() - c0_;
                int img, y0, x0;
                tile_coords(g, t, img, y0, x0);
                const uint32_t xs = base + st * stage_bytes;
                const uint32_t ds = xs + 4u * g.xplane;
                const uint32_t bar = bar_raw_full + 8 * st;
                mbar_arrive_expect_tx(bar, tx_bytes);
                tma_load_4d(xs, tmX, bar, cx, x0 - pad, y0 - pad, img);
                tma_load_4d(xs + g.xplane, tmX, bar, cx + 32, x0 - pad, y0 - pad, img);
                tma_load_4d(ds, &tmDY, bar, nb * 64, x0, y0, img);
                tma_load_4d(ds + kWgDyPlane, &tmDY, bar, nb * 64 + 32, x0, y0, img);
            }
            if (g.dbg && blockIdx.x == 0) { g.dbg[0] = w_empty; g.dbg[1] = clock64() - t_begin; g.dbg[2] = it; }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_tf32(128, DC, 1, 1);
            constexpr int P = kWgTW + 2 * (R / 2);
            // descriptor words: hi = SBO 512 | version 1 | layout SWIZZLE_128B_BASE32B; lo = LBO << 16 | start >> 4
            const uint32_t d_hi = static_cast<uint32_t>(umma_desc_mn32(0, 0) >> 32);
            const uint32_t a_lbo = ((static_cast<uint32_t>(g.xplane) >> 4) & 0x3FFFu) << 16;
            constexpr uint32_t b_lbo = ((static_cast<uint32_t>(kWgDyPlane) >> 4) & 0x3FFFu) << 16;
            int it = 0;
            long long w_split = 0, t_begin = clock64();
            for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
                const int st = it & 1;
                const uint32_t use = static_cast<uint32_t>(it >> 1);
                const long long c0_ = clock64();
                mbar_wait(bar_split_full + 8 * st, use & 1u);
                w_split += clock64() - c0_;
                tc_fence_after_sync();
                const uint32_t xs = base + st * stage_bytes;
                const uint32_t ds = xs + 4u * g.xplane;
                const uint32_t b_lo0 = b_lbo | (ds >> 4);
                int r = tap0 / R, s = tap0 - r * R;
                for (int tl = 0; tl < ntap; ++tl) {
                    const uint32_t d_tmem = tmem_base + tl * DC;
                    const uint32_t a_lo0 = a_lbo | ((xs >> 4) + static_cast<uint32_t>(r * P + s) * 8u);
#pragma unroll
                    for (int ty = 0; ty < kWgTH; ++ty) {
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            const uint32_t a_lo = a_lo0 + static_cast<uint32_t>(ty * P + 8 * h) * 8u;
                            const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(ty * kWgTW + 8 * h) * 8u;
                            umma_tf32_ss_w(d_tmem, a_lo, d_hi, b_lo, d_hi, idesc, (it | ty | h) != 0);
                            if (!kWide) umma_tf32_ss_w(d_tmem, a_lo, d_hi, b_lo + (2u * kWgDyPlane >> 4), d_hi, idesc, 1u);
                        }
                    }
                    if (++s == R) { s = 0; ++r; }
                }
                umma_commit(bar_stage_empty + 8 * st);
            }
            umma_commit(bar_acc);
            if (g.dbg && blockIdx.x == 0) { g.dbg[3] = w_split; g.dbg[4] = clock64() - t_begin; }
        }
    } else {
        // ===================== hi/lo split (+ bias-gradient sums), then epilogue =====================
        const int tid = threadIdx.x - 64;                 // 0..127
        const bool do_db = g.db != nullptr && tg == 0 && cg == 0;
        float4 db0 = make_float4(0.f, 0.f, 0.f, 0.f), db1 = db0;
        const int nf4x = hrows * 8;
        int it = 0;
        long long w_raw = 0, t_begin = clock64();
        for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
            const int st = it & 1;
            const uint32_t use = static_cast<uint32_t>(it >> 1);
            const long long c0_ = clock64();
            mbar_wait(bar_raw_full + 8 * st, use & 1u);
            w_raw += clock64() - c0_;
            const uint32_t xs = base + st * stage_bytes;
            const uint32_t ds = xs + 4u * g.xplane;
            {
                // X halo planes: items (plane, i) with i = tid + 128 * k, in batches of 4 loads in flight
                constexpr int kPer = (((kWgTH + 2 * pad) * (kWgTW + 2 * pad) * 8) + 127) / 128;   // per plane, per thread
                float4 dummy = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int k0 = 0; k0 < 2 * kPer; k0 += 4) {
                    uint32_t src[4]; bool ok[4];
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        const int item = k0 + b;
                        const int i = tid + 128 * (item >> 1);
                        ok[b] = item < 2 * kPer && i < nf4x;
                        src[b] = xs + static_cast<uint32_t>(item & 1) * g.xplane + static_cast<uint32_t>(i) * 16u;
                    }
                    split_batch<4>(src, ok, 2u * g.xplane, g.trunc_hi, dummy);
                }
                // dy planes (sums of the raw values feed the bias gradient)
#pragma unroll
                for (int pl = 0; pl < 2; ++pl) {
                    uint32_t src[4]; bool ok[4];
#pragma unroll
                    for (int b = 0; b < 4; ++b) {
                        ok[b] = true;
                        src[b] = ds + static_cast<uint32_t>(pl) * kWgDyPlane + static_cast<uint32_t>(tid + 128 * b) * 16u;
                    }
                    split_batch<4>(src, ok, 2u * kWgDyPlane, g.trunc_hi, pl == 0 ? db0 : db1);
                }
            }
            fence_proxy_async_smem();
            mbar_arrive(bar_split_full + 8 * st);
        }

        // ---- epilogue: dwp[tap][c][n] += D[c][n] + D[64 + c][n]
        const long long t_loop = clock64() - t_begin;
        mbar_wait(bar_acc, 0u);             // every MMA has completed: the stage buffers are free as well
        tc_fence_after_sync();
        const long long t_acc = clock64() - t_begin;
        if (do_db) {
            // thread tid always touched physical 16-byte chunk (tid & 7) of rows with (row & 3) == ((tid >> 3) & 3):
            // logical 32-byte group = physical group ^ (row & 3)  (SWIZZLE_128B_BASE32B)
            float* sc = reinterpret_cast<float*>(smem_raw + (base - smem_u32(smem_raw)));   // stage 0 is idle now
            sc[tid * 8 + 0] = db0.x; sc[tid * 8 + 1] = db0.y; sc[tid * 8 + 2] = db0.z; sc[tid * 8 + 3] = db0.w;
            sc[tid * 8 + 4] = db1.x; sc[tid * 8 + 5] = db1.y; sc[tid * 8 + 6] = db1.z; sc[tid * 8 + 7] = db1.w;
            named_bar_sync(1, 128);
            if (tid < 64) {
                const int plane = tid >> 5, l32 = (tid & 31) >> 3, b4 = (tid & 7) >> 2, e = tid & 3;
                float sum = 0.f;
#pragma unroll
                for (int rr = 0; rr < 4; ++rr) {
                    const int c16 = ((l32 ^ rr) << 1) | b4;
#pragma unroll
                    for (int rb = 0; rb < 4; ++rb) sum += sc[((rr + 4 * rb) * 8 + c16) * 8 + plane * 4 + e];
                }
                atomicAdd(g.db + nb * 64 + tid, sum);
            }
        }

        // D rows 0-63 = Xhi * [dyhi | dylo], rows 64-127 = Xlo * [...]: dw[c][n] is the sum over both row halves (and, in
        // wide mode, both column halves).  The lo-row warps hand their sums over through shared memory (the stage
        // buffers are idle), so each CTA issues 64 x 16 vector reductions per tap.
        const int q = warp & 3;
        const int row = q * 32 + lane;
        const int cl = row & 63;
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        for (int tl = 0; tl < ntap; ++tl) {
            float acc[64];
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(tmem_base + lane_addr + tl * DC + c0, v);
                tmem_ld_wait();
#pragma unroll
                for (int k = 0; k < 32; ++k) acc[c0 + k] = __uint_as_float(v[k]);
                if (kWide) {
                    tmem_ld32(tmem_base + lane_addr + tl * DC + 64 + c0, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int k = 0; k < 32; ++k) acc[c0 + k] += __uint_as_float(v[k]);
                }
            }
            const uint32_t sc = base + 16384u + static_cast<uint32_t>(tl & 1) * 16384u + static_cast<uint32_t>(cl) * 256u;
            if (row >= 64) {
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sc + ((static_cast<uint32_t>(k) ^ (cl & 15)) << 4)),
                                 "f"(acc[4 * k]), "f"(acc[4 * k + 1]), "f"(acc[4 * k + 2]), "f"(acc[4 * k + 3]) : "memory");
            }
            named_bar_sync(1, 128);
            if (row < 64) {
                float* dst = g.dwp + (static_cast<long>(tap0 + tl) * g.Ctot + cg * 64 + cl) * g.N + nb * 64;
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    float4 o;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
                                 : "r"(sc + ((static_cast<uint32_t>(k) ^ (cl & 15)) << 4)));
                    red_add_v4(dst + 4 * k, acc[4 * k] + o.x, acc[4 * k + 1] + o.y, acc[4 * k + 2] + o.z, acc[4 * k + 3] + o.w);
                }
            }
        }
        tc_fence_before_sync();
        if (g.dbg && blockIdx.x == 0 && tid == 0) { g.dbg[5] = w_raw; g.dbg[6] = t_loop; g.dbg[7] = t_acc; g.dbg[8] = clock64() - t_begin; }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, tmem_cols);
    }
}

// dw[n][c][r][s] += dwp[tap][c][n]   (torch OIHW <- the kernels' accumulation layout)
__global__ void __launch_bounds__(256) wgrad_unpack_kernel(const float* __restrict__ dwp, int N, int C, int taps,
                                                           float* __restrict__ dw) {
    const long total = static_cast<long>(N) * C * taps;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int tap = static_cast<int>(i % taps);
        const int c = static_cast<int>((i / taps) % C);
        const int n = static_cast<int>(i / (static_cast<long>(taps) * C));
        dw[i] += dwp[(static_cast<long>(tap) * C + c) * N + n];
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn wg_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// 4-D map over NHWC (C, W, H, B) with the 32-byte-atom 128-byte swizzle; box = (32 channels, bw pixels, bh rows, 1)
static int make_map32(CUtensorMap* tm, const float* x, int ld, int C, int B, int H, int W, int bw, int bh) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, static_cast<cuuint64_t>(W) * ld * 4,
                                   static_cast<cuuint64_t>(H) * W * ld * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(bw), static_cast<cuuint32_t>(bh), 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = wg_encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                                CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B,
                                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(wgrad) failed: %d", static_cast<int>(r));
    return 0;
}

bool wgrad_tc_supported(int c0, int c1, int N, int H, int W, int R) {
    if (R != 1 && R != 3 && R != 5) return false;
    if (c0 % 64 != 0 || c1 % 64 != 0 || N % 64 != 0) return false;
    if (W % kWgTW != 0 || H % kWgTH != 0) return false;
    return true;
}

int launch_wgrad_tc(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy, int B,
                    int H, int W, int R, int N, float* dwp, float* db, cudaStream_t s) {
    if (!wg_encode_fn()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    WgradTcGeom g;
    g.B = B; g.H = H; g.W = W; g.R = R;
    g.Ctot = c0 + c1; g.N = N; g.c0 = c0;
    const int pad = R / 2;
    g.P = kWgTW + 2 * pad;
    const int hrows = (kWgTH + 2 * pad) * g.P;
    g.xplane = (hrows * 128 + 1023) / 1024 * 1024;
    const int taps = R * R;
    g.taps_per_group = R;                       // 1, 3 or 5 taps (<= 8 x 64 TMEM columns)
    g.TG = taps / g.taps_per_group;
    g.CG = g.Ctot / 64;
    g.NB = N / 64;
    g.ntiles = B * (H / kWgTH) * (W / kWgTW);
    const int units = g.TG * g.CG * g.NB;
    g.PS = std::max(1, std::min(g.ntiles, kSMs / units));
    g.dwp = dwp; g.db = db;
    static const int trunc_hi = getenv("ODEX_WGRAD_TRUNC") ? atoi(getenv("ODEX_WGRAD_TRUNC")) : 1;
    g.trunc_hi = trunc_hi;
    g.dbg = nullptr;
    static const bool debug = getenv("ODEX_WGRAD_DEBUG") != nullptr;
    if (debug) {
        static long long* dbuf = nullptr;
        if (!dbuf) cudaMalloc(&dbuf, 16 * sizeof(long long));
        cudaMemsetAsync(dbuf, 0, 16 * sizeof(long long), s);
        g.dbg = dbuf;
    }
    const int smem = 1024 + 2 * (4 * g.xplane + 4 * kWgDyPlane) + 64;
    if (smem > 232448) return fail(ODEX_E_UNSUPPORTED, "wgrad(tc): needs %d bytes of shared memory", smem);

    CUtensorMap tmX0, tmX1, tmDY;
    if (int rc = make_map32(&tmX0, x0, ld0, c0, B, H, W, g.P, kWgTH + 2 * pad)) return rc;
    if (x1) {
        if (int rc = make_map32(&tmX1, x1, ld1, c1, B, H, W, g.P, kWgTH + 2 * pad)) return rc;
    } else {
        tmX1 = tmX0;
    }
    if (int rc = make_map32(&tmDY, dy, ldy, N, B, H, W, kWgTW, kWgTH)) return rc;

    auto kern = R == 1 ? wgrad_tc_kernel<1> : (R == 3 ? wgrad_tc_kernel<3> : wgrad_tc_kernel<5>);
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] {
        attr_err = cudaFuncSetAttribute(wgrad_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        if (attr_err == cudaSuccess) attr_err = cudaFuncSetAttribute(wgrad_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
        if (attr_err == cudaSuccess) attr_err = cudaFuncSetAttribute(wgrad_tc_kernel<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448);
    });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "wgrad(tc): cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    kern<<<units * g.PS, 192, smem, s>>>(tmX0, tmX1, tmDY, g);
    if (debug) {
        long long h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, g.dbg, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[wgrad_tc R=%d C=%d N=%d B=%d grid=%d tiles/cta=%lld] producer: wait_empty=%lld total=%lld | mma: wait_split=%lld "
                "total=%lld | split: wait_raw=%lld loop=%lld until_acc=%lld end=%lld (cycles)\n", R, g.Ctot, N, B, units * g.PS, h[2],
                h[0], h[1], h[3], h[4], h[5], h[6], h[7], h[8]);
    }
    return launched("wgrad_tc_kernel");
}

}  // namespace odex

using namespace odex;

extern "C" int odex_wgrad_unpack(const float* dwp, int N, int C, int R, float* dw, odex_stream_t s) {
    if (!dwp || !dw || N <= 0 || C <= 0 || R <= 0) return fail(ODEX_E_ARG, "wgrad_unpack: bad argument");
    const long total = static_cast<long>(N) * C * R * R;
    const int blocks = static_cast<int>(std::min<long>((total + 255) / 256, 4L * kSMs));
    wgrad_unpack_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(s)>>>(dwp, N, C, R * R, dw);
    return launched("wgrad_unpack_kernel");
}
