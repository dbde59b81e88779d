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
    float* dwp;         // [R*R][N][Ctot], accumulated
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

// TS-mode MMA: A (dy, K-major) from TMEM, B (X atom, MN-major) from shared memory.
__device__ __forceinline__ void umma_tf32_ts_w(uint32_t d_tmem, uint32_t a_tmem, uint32_t b_lo, uint32_t b_hi, uint32_t idesc,
                                               uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 db;\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], db, %4, p;\n\t}"
        ::"r"(d_tmem), "r"(a_tmem), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

template <int R>
__global__ void __launch_bounds__(192, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmX0, const __grid_constant__ CUtensorMap tmX1,
                const __grid_constant__ CUtensorMap tmDY, const WgradTcGeom g) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t stage_bytes = 4u * g.xplane + 2u * kWgDyPlane;   // [Xhi c0-31][Xhi c32-63][Xlo ..][Xlo ..][dy n0-31][dy n32-63]
    constexpr int S = (R <= 3) ? 3 : 2;            // pipeline stages (shared-memory tiles and TMEM A slots)
    const uint32_t bars = base + static_cast<uint32_t>(S) * stage_bytes;
    const uint32_t bar_raw_full = bars;            // [S]
    const uint32_t bar_split_full = bars + 24;     // [S]
    const uint32_t bar_stage_empty = bars + 48;    // [S]
    const uint32_t bar_acc = bars + 72;
    const uint32_t tmem_slot = bars + 80;

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
    constexpr int P = kWgTW + 2 * pad;
    constexpr int hrows = (kWgTH + 2 * pad) * P;   // pixel slots of one halo plane
    // The A operand [dy_hi ; dy_lo] (M = 128 rows = 64 channels x {hi, lo}, K = the tile's 64 pixels) lives in TMEM, so the
    // tensor core reads only the X atoms from shared memory: per 8-pixel atom two N = 64 MMAs (B = Xhi atom, then Xlo atom)
    // accumulate all four partial products into the tap's 64 columns (rows 0-63: dy_hi x ., rows 64-127: dy_lo x .).
    constexpr bool kWide = false;
    constexpr int DC = kWide ? 128 : 64;
    const uint32_t a_col0 = static_cast<uint32_t>(ntap * DC);       // S slots x 64 columns of A behind the accumulators
    uint32_t tmem_cols = 32;
    while (tmem_cols < a_col0 + 64u * S) tmem_cols <<= 1;

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

    // shared memory / TMEM set-up above may overlap the previous kernel's tail; global memory is first touched below
    pdl_wait();
    pdl_trigger();

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            const int cabs = cg * 64;
            const bool second = cabs >= g.c0;
            const CUtensorMap* tmX = second ? &tmX1 : &tmX0;
            const int cx = second ? cabs - g.c0 : cabs;
            constexpr uint32_t tx_bytes = 2u * hrows * 128u + 2u * kWgDyPlane;
            int it = 0;
            long long w_empty = 0, t_begin = clock64();
            for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
                const int st = it % S;
                const uint32_t use = static_cast<uint32_t>(it / S);
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
        // The whole warp runs this loop so that every operand (descriptor words, TMEM addresses) is warp-uniform and stays
        // in uniform registers; one elected lane issues.  (Wrapped in `if (lane == 0)` the operands are per-thread
        // values and every UTCHMMA is preceded by an ELECT / R2UR.BROADCAST sequence: ~60 cycles per MMA, measured.)
        {
            constexpr uint32_t idesc = umma_idesc_tf32(128, DC, 0, 1);        // A K-major (TMEM), B MN-major (smem)
            // B descriptor words: hi = SBO 512 | version 1 | layout SWIZZLE_128B_BASE32B; lo = LBO << 16 | start >> 4
            const uint32_t d_hi = static_cast<uint32_t>(umma_desc_mn32(0, 0) >> 32);
            const uint32_t b_lbo = ((static_cast<uint32_t>(g.xplane) >> 4) & 0x3FFFu) << 16;
            const uint32_t lo_delta = (2u * static_cast<uint32_t>(g.xplane)) >> 4;   // Xhi planes -> Xlo planes
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
                const uint32_t a_tmem = tmem_base + a_col0 + static_cast<uint32_t>(st) * 64u;
                int r = tap0 / R, s = tap0 - r * R;
                for (int tl = 0; tl < ntap; ++tl) {
                    const uint32_t d_tmem = tmem_base + tl * DC;
                    const uint32_t b_lo0 = b_lbo | ((xs >> 4) + static_cast<uint32_t>(r * P + s) * 8u);
                    if (elect_one()) {
#pragma unroll
                        for (int ty = 0; ty < kWgTH; ++ty) {
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                const uint32_t b_lo = b_lo0 + static_cast<uint32_t>(ty * P + 8 * h) * 8u;
                                const uint32_t a_at = a_tmem + static_cast<uint32_t>(ty * kWgTW + 8 * h);
                                umma_tf32_ts_w(d_tmem, a_at, b_lo, d_hi, idesc, (it | ty | h) != 0);
                                if (!kWide) umma_tf32_ts_w(d_tmem, a_at, b_lo + lo_delta, d_hi, idesc, 1u);
                            }
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
        }
    } else {
        // ===================== operand preparation, then epilogue =====================
        // thread <-> TMEM lane L = 32 * (warp % 4) + lane: dy channel n = L % 64, part = L / 64 (0: hi rows, 1: lo rows)
        const int tid = threadIdx.x - 64;                 // 0..127 (X split work distribution)
        const int q = warp & 3;
        const int L = q * 32 + lane;
        const int n = L & 63;
        const int part = L >> 6;
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        const bool do_db = g.db != nullptr && tg == 0 && cg == 0 && part == 0;
        float dbsum = 0.f;
        constexpr int nf4x = hrows * 8;
        // dy element (pixel k, channel n) inside the SWIZZLE_128B_ATOM_32B tile: 32-byte group ((n % 32) / 8) ^ (k % 4)
        const uint32_t dy_off = static_cast<uint32_t>(n >> 5) * kWgDyPlane + static_cast<uint32_t>(n & 7) * 4u;
        const uint32_t dy_grp = static_cast<uint32_t>((n & 31) >> 3);
        int it = 0;
        long long w_raw = 0, t_begin = clock64();
        for (int t = ps; t < g.ntiles; t += g.PS, ++it) {
            const int st = it % S;
            const uint32_t use = static_cast<uint32_t>(it / S);
            const long long c0_ = clock64();
            mbar_wait(bar_raw_full + 8 * st, use & 1u);
            w_raw += clock64() - c0_;
            const uint32_t xs = base + st * stage_bytes;
            const uint32_t ds = xs + 4u * g.xplane + dy_off;
            {
                // X halo planes: hi = the raw word (the tensor core reads the top 19 bits), lo = tf32(x - trunc(x)) into
                // the lo planes; items (plane, i) with i = tid + 128 * k, 4 loads in flight
                constexpr int kPer = (nf4x + 127) / 128;   // per plane, per thread
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
                    split_batch<4>(src, ok, 2u * g.xplane, 1, dummy);
                }
            }
            // dy: this lane's channel at the tile's 64 pixels -> A operand row in TMEM (K-major: column = pixel)
            tc_fence_after_sync();
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
                uint32_t a[32];
#pragma unroll
                for (int k0 = 0; k0 < 32; k0 += 16) {
                    uint32_t v[16];
#pragma unroll
                    for (int b = 0; b < 16; ++b) {
                        const uint32_t k = static_cast<uint32_t>(c0 + k0 + b);
                        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v[b]) : "r"(ds + k * 128u + ((dy_grp ^ (k & 3u)) << 5)));
                    }
#pragma unroll
                    for (int b = 0; b < 16; ++b) {
                        const float x = __uint_as_float(v[b]);
                        dbsum += x;
                        const float hi = __uint_as_float(v[b] & 0xFFFFE000u);
                        a[k0 + b] = part == 0 ? v[b] : tf32_rna_bits(__float_as_uint(x - hi));
                    }
                }
                tmem_st32(tmem_base + lane_addr + a_col0 + static_cast<uint32_t>(st) * 64u + c0, a);
            }
            tmem_st_wait();
            tc_fence_before_sync();
            fence_proxy_async_smem();
            mbar_arrive(bar_split_full + 8 * st);
        }
        const long long t_loop = clock64() - t_begin;
        mbar_wait(bar_acc, 0u);             // every MMA has completed: the stage buffers are free as well
        tc_fence_after_sync();
        const long long t_acc = clock64() - t_begin;
        if (do_db) atomicAdd(g.db + nb * 64 + n, dbsum);

        // ---- epilogue: dwp[tap][n][c] += D[n][c] (+ D[n][64 + c]) + the same from the lo rows D[64 + n][.]; the lo-row
        // warps hand their sums over through shared memory, so a CTA issues 64 x 16 vector reductions per tap
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
            const uint32_t sc = base + 16384u + static_cast<uint32_t>(tl & 1) * 16384u + static_cast<uint32_t>(n) * 256u;
            if (part == 1) {
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(sc + ((static_cast<uint32_t>(k) ^ (n & 15)) << 4)),
                                 "f"(acc[4 * k]), "f"(acc[4 * k + 1]), "f"(acc[4 * k + 2]), "f"(acc[4 * k + 3]) : "memory");
            }
            named_bar_sync(1, 128);
            if (part == 0) {
                float* dst = g.dwp + (static_cast<long>(tap0 + tl) * g.N + nb * 64 + n) * g.Ctot + cg * 64;
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    float4 o;
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w)
                                 : "r"(sc + ((static_cast<uint32_t>(k) ^ (n & 15)) << 4)));
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

// dw[n][c][r][s] += dwp[tap][n][c]   (torch OIHW <- the kernels' accumulation layout)
__global__ void __launch_bounds__(256) wgrad_unpack_kernel(const float* __restrict__ dwp, int N, int C, int taps,
                                                           float* __restrict__ dw) {
    const long total = static_cast<long>(N) * C * taps;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int tap = static_cast<int>(i % taps);
        const int c = static_cast<int>((i / taps) % C);
        const int n = static_cast<int>(i / (static_cast<long>(taps) * C));
        dw[i] += dwp[(static_cast<long>(tap) * N + n) * C + c];
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
    const int stages = R <= 3 ? 3 : 2;
    const int smem = 1024 + stages * (4 * g.xplane + 2 * kWgDyPlane) + 128;
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
    launch_pdl(kern, units * g.PS, 192, smem, s, tmX0, tmX1, tmDY, g);
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
