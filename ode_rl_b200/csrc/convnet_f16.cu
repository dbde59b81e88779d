// Fused chain of 3x3 64->64 convolutions over 16x16 NHWC fp32 images on tcgen05 kind::f16 with a two-term fp16 split of both
// operands (C-ABI odex_convnet with split = 1; the kind::tf32 / 3xTF32 version of the same chain is convnet_tc.cu).
//
// Why: fp32-level results (<= 1e-5 after 36 sequential f-evals) need ~22 significant bits per operand.  3xTF32 gets them
// from tf32 hi + tf32 lo at the tf32 MMA rate; x = (hi + lo * 2^-11) / s with hi = fp16(x*s), lo = fp16((x*s - hi) * 2^11)
// gets the same 22 bits from fp16 operands, whose MMAs run at TWICE the tf32 rate and read half the shared-memory bytes.
// Same three partial products: main = hi.w_hi, corr = hi.w_lo + lo.w_hi, result = (main + corr * 2^-11) / (s_a * s_w).
// fp16 has a 5-bit exponent, so every operand tensor carries a power-of-two scale s:
//   weights      s_w per tensor, chosen by odex_pack_weights(tc = 2) (max |w| * s_w in [2^12, 2^13)), stored behind the rows;
//   activations  s_a per CTA (= per image) and per layer.  The input's is taken from its measured max |x|.  A layer's OUTPUT
//                must be converted in the same epilogue pass that produces it, before its max is known, so its scale comes
//                from the rigorous bound  max |v| <= G * max |a_in| + max |bias|  (G = max_n sum_k |w[n,k]|, from the packer;
//                act / act' masks only shrink |v|): no overflow for any input, and since max |a_in| is the MEASURED max of the
//                previous layer's output the slack (a factor 10-50 for conv nets) does not compound.  Elements far below the
//                max become fp16 subnormals of hi, which lo then resolves: absolute error <= 2^-35 * 2^13 / s, i.e. < 2^-30 of
//                the tensor's max -- below the fp32 rounding of the accumulation.
//
// Layout: the image lives in shared memory as two 18x18-slot halo planes [hi | lo], one pixel slot = 128 B = 64 fp16 channels
// (SWIZZLE_128B by absolute address, zero border = padding); tap (r, s) = start address of the K-major A descriptor
// (conv_tc2.cu's trick).  A dense fp32 staging tile S (2 x 256 px x 32 ch, the TMA box layout) receives the input and each
// layer's mask / residual operand by TMA and is the source of the TMA stores of the layer outputs that go to global memory.
//   warp 0   TMA producer: input tile, weight ring [w_hi ; w_lo] per tap (runs ahead across layers)
//   warp 1   TMEM allocator + MMA issue (whole warp loops, one elected lane issues)
//   warps 2-9 input conversion, then every layer's epilogue: TMEM -> registers -> unscale / bias / act / act' / residual ->
//            fp32 into S (if stored) and fp16 hi / lo into the planes (next layer's operands); TMA aux loads and stores by tid 0.
//            Two groups of four warps (a warp reaches the TMEM lanes of its quarter only), one per 32-channel half of every pixel:
//            the epilogue is a latency chain (tcgen05.ld -> FMA -> cvt -> st.shared) and runs between two layers' MMAs, not under
//            them, so two warps per scheduler instead of one shorten every layer
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"
#include "f16_split.cuh"

namespace odex {

constexpr int kF16MaxL = ODEX_NET_MAX_LAYERS;
constexpr int kF16BStage = 2 * kF16C * 128;    // [w_hi ; w_lo] of one tap: 16 KB
constexpr int kF16SB = 4;
constexpr int kF16EpiThreads = 256;             // 8 epilogue warps
constexpr int kF16Threads = 64 + kF16EpiThreads;
constexpr int kF16Smem = 1024 + 2 * kF16Plane + 2 * kF16SHalf + kF16SB * kF16BStage + 256 + kF16MaxL * 256 + 64;

struct F16Maps {
    CUtensorMap x;                  // fp32 tile box [32][16][16][1]
    CUtensorMap w[kF16MaxL];        // packed fp16 weights, rows of 128 B, box [64][128]
    CUtensorMap aux[kF16MaxL];
    CUtensorMap out[kF16MaxL];
};

struct F16Args {
    int L;
    int act[kF16MaxL];
    int aux_kind[kF16MaxL];
    int store[kF16MaxL];
    const float* bias[kF16MaxL];
    const float* wtail[kF16MaxL];   // {1 / s_w, G, max |w|, -} behind the packed rows
    float alpha;
    const double* alpha_dev;
    const int* skip;
    float* amax_out[kF16MaxL];      // nullable: amax_out[l][img] <- max |input of layer l| of image img
    long long* dbg;
};

template <bool kTanh>
__global__ void __launch_bounds__(kF16Threads, 1)
convnet_f16_kernel(const __grid_constant__ F16Maps tm, const F16Args a) {
    constexpr int N = kF16C;
    constexpr int taps = 9;
    constexpr uint32_t kTmemCols = 4 * N;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;       // hi plane; lo plane at + kF16Plane
    const uint32_t s_base = base + 2u * kF16Plane;                     // S: [c0-31 | c32-63] x 256 px, fp32
    const uint32_t bst0 = s_base + 2u * kF16SHalf;
    const uint32_t bars = bst0 + kF16SB * kF16BStage;
    const uint32_t bar_in = bars;                  // input tile landed in S
    const uint32_t bar_b_full = bars + 8;          // [kF16SB <= 4]
    const uint32_t bar_b_empty = bars + 40;        // [kF16SB <= 4]
    const uint32_t bar_acc = bars + 72;            // a layer's MMAs have retired            (one phase per layer)
    const uint32_t bar_next = bars + 80;           // a layer's operands are in the planes   (128 arrivals per layer, layer 0 included)
    const uint32_t bar_aux = bars + 88;            // a mask / residual tile landed in S     (one phase per use)
    const uint32_t tmem_slot = bars + 96;
    const uint32_t amax_s = bars + 112;            // [L + 1] measured max |a_l| (float bits, >= 0)
    const uint32_t bias_s = bars + 256;            // [L][64] floats

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int img = blockIdx.x;
    const int L = a.L;

    if (threadIdx.x == 0) {
        mbar_init(bar_in, 1);
        for (int i = 0; i < kF16SB; ++i) { mbar_init(bar_b_full + 8 * i, 1); mbar_init(bar_b_empty + 8 * i, 1); }
        mbar_init(bar_acc, 1);
        mbar_init(bar_next, kF16EpiThreads);
        mbar_init(bar_aux, 1);
        fence_barrier_init();
    }
    if (threadIdx.x < 8) asm volatile("st.shared.u32 [%0], %1;" ::"r"(amax_s + 4u * threadIdx.x), "r"(0u) : "memory");
    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tm.x);
        for (int l = 0; l < L; ++l) tma_prefetch_desc(&tm.w[l]);
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

    if (skip) {
    } else if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            mbar_arrive_expect_tx(bar_in, 2u * kF16SHalf);
            tma_load_4d(s_base, &tm.x, bar_in, 0, 0, 0, img);
            tma_load_4d(s_base + kF16SHalf, &tm.x, bar_in, 32, 0, 0, img);
            int sb = 0; uint32_t pb = 0;
            for (int l = 0; l < L; ++l) {
                for (int tap = 0; tap < taps; ++tap) {
                    mbar_wait(bar_b_empty + 8 * sb, pb ^ 1u);
                    mbar_arrive_expect_tx(bar_b_full + 8 * sb, kF16BStage);
                    tma_load_2d(bst0 + sb * kF16BStage, &tm.w[l], bar_b_full + 8 * sb, 0, tap * 2 * N);
                    if (++sb == kF16SB) { sb = 0; pb ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issue =====================
        constexpr uint32_t idesc_wide = umma_idesc_f16(128, 2 * N);
        constexpr uint32_t idesc_narrow = umma_idesc_f16(128, N);
        const uint32_t a_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, kF16P * 128) >> 32);
        const uint32_t b_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024) >> 32);
        constexpr uint32_t lbo16 = 1u << 16;
        constexpr uint32_t lo_delta = static_cast<uint32_t>(kF16Plane) >> 4;
        int sb = 0; uint32_t pb = 0;
        for (int l = 0; l < L; ++l) {
            mbar_wait(bar_next, static_cast<uint32_t>(l) & 1u);
            tc_fence_after_sync();
            int r = 0, s = 0;
            for (int tap = 0; tap < taps; ++tap) {
                mbar_wait(bar_b_full + 8 * sb, pb);
                tc_fence_after_sync();
                const uint32_t b_lo0 = lbo16 | ((bst0 + sb * kF16BStage) >> 4);
                const uint32_t a_lo0 = lbo16 | ((base >> 4) + static_cast<uint32_t>(r * kF16P + s) * 8u);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
#pragma unroll
                        for (int mt = 0; mt < 2; ++mt) {
                            const uint32_t a_lo = a_lo0 + mt * 64u + k * 2u;
                            const uint32_t d = tmem_base + mt * 2 * N;
                            umma_f16_ss_w(d, a_lo, a_hi32, b_lo0 + k * 2u, b_hi32, idesc_wide, (tap | k) != 0);
                            umma_f16_ss_w(d + N, a_lo + lo_delta, a_hi32, b_lo0 + k * 2u, b_hi32, idesc_narrow, 1u);
                        }
                    }
                    umma_commit(bar_b_empty + 8 * sb);
                }
                __syncwarp();
                if (++sb == kF16SB) { sb = 0; pb ^= 1u; }
                if (++s == 3) { s = 0; ++r; }
            }
            if (elect_one()) umma_commit(bar_acc);
            __syncwarp();
        }
    } else {
        // ===================== input conversion, then every layer's epilogue =====================
        const int tid = threadIdx.x - 64;
        for (int i = tid; i < L * (N / 4); i += kF16EpiThreads) {
            const int l = i / (N / 4), j = i % (N / 4);
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a.bias[l]) bv = __ldg(reinterpret_cast<const float4*>(a.bias[l]) + j);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(bias_s + 16u * i), "f"(bv.x), "f"(bv.y), "f"(bv.z), "f"(bv.w) : "memory");
        }
        // zero both planes (the 1-pixel border is the convolution's padding and is never written again)
        for (uint32_t i = tid; i < 2u * kF16Plane / 16u; i += kF16EpiThreads) sts128(base + i * 16u, make_uint4(0u, 0u, 0u, 0u));
        float alpha = a.alpha;
        if (a.alpha_dev) alpha *= static_cast<float>(*a.alpha_dev);
        asm volatile("bar.sync 1, 256;" ::: "memory");

        const int q = warp & 3;
        const int hh = tid >> 7;                      // this warp group's 32-channel half
        const int m = q * 32 + lane;                  // TMEM lane = pixel of the M-tile: (y, x) = (m / 8, mt * 8 + m % 8)
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        const float4* bias4 = reinterpret_cast<const float4*>(smem_raw + (bias_s - smem_u32(smem_raw)));

        // ---- layer 0 operands: max |x| of the image, then hi / lo of x * s0
        mbar_wait(bar_in, 0u);
        float cur_inv;                                // 1 / (scale of the operands now in the planes)
        {
            float mx = 0.f;
            float4 xv[2][8];
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                const uint32_t p = static_cast<uint32_t>((m >> 3) * 16 + mt * 8 + (m & 7));
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    xv[mt][j] = lds128f(s_base + static_cast<uint32_t>(hh) * kF16SHalf + p * 128u + ((static_cast<uint32_t>(j) ^ (p & 7u)) << 4));
                    mx = fmaxf(mx, fmaxf(fmaxf(fabsf(xv[mt][j].x), fabsf(xv[mt][j].y)), fmaxf(fabsf(xv[mt][j].z), fabsf(xv[mt][j].w))));
                }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            if (lane == 0) asm volatile("atom.shared.max.u32 _, [%0], %1;" ::"r"(amax_s), "r"(__float_as_uint(mx)) : "memory");   // NaN compares high: it propagates
            asm volatile("bar.sync 1, 256;" ::: "memory");
            uint32_t mb;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(mb) : "r"(amax_s));
            const float s0 = f16_scale_for(__uint_as_float(mb), &cur_inv);
#pragma unroll
            for (int mt = 0; mt < 2; ++mt) {
                const uint32_t srow = f16_slot(m >> 3, mt * 8 + (m & 7));
                const uint32_t row = base + srow * 128u;
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    const float4 u = xv[mt][2 * c8], w = xv[mt][2 * c8 + 1];
                    const float o[8] = {u.x * s0, u.y * s0, u.z * s0, u.w * s0, w.x * s0, w.y * s0, w.z * s0, w.w * s0};
                    uint4 hi, lo;
                    f16_split8(o, &hi, &lo);
                    const uint32_t off = (static_cast<uint32_t>(4 * hh + c8) ^ (srow & 7u)) << 4;
                    sts128(row + off, hi);
                    sts128(row + kF16Plane + off, lo);
                }
            }
        }
        fence_proxy_async_smem();
        mbar_arrive(bar_next);
        asm volatile("bar.sync 1, 256;" ::: "memory");          // S (the input) has been consumed by every thread

        uint32_t aux_phase = 0;
        bool stores_pending = false;
        for (int l = 0; l < L; ++l) {
            const bool last = l == L - 1;
            const int kind = a.aux_kind[l];
            const bool touch_s = kind != 0 || a.store[l] != 0;      // this layer reads an operand from S and / or writes its result there
            if (touch_s) {
                if (tid == 0) {
                    if (stores_pending) tma_store_wait_read();    // the last stores issued have read S
                    if (kind) {
                        mbar_arrive_expect_tx(bar_aux, 2u * kF16SHalf);
                        tma_load_4d(s_base, &tm.aux[l], bar_aux, 0, 0, 0, img);
                        tma_load_4d(s_base + kF16SHalf, &tm.aux[l], bar_aux, 32, 0, 0, img);
                    }
                }
                stores_pending = false;
            }
            // scales: this layer's accumulators carry s_a * s_w; its output is converted with a scale sized from the bound
            const float w_inv = __ldg(a.wtail[l]), G = __ldg(a.wtail[l] + 1);
            uint32_t inb;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(inb) : "r"(amax_s + 4u * l));
            float bmax = 0.f;
            {
                const float4* bl = bias4 + l * (N / 4);
#pragma unroll 4
                for (int j = 0; j < N / 4; ++j) bmax = fmaxf(bmax, fmaxf(fmaxf(fabsf(bl[j].x), fabsf(bl[j].y)), fmaxf(fabsf(bl[j].z), fabsf(bl[j].w))));
            }
            const float unscale = cur_inv * w_inv;
            float next_inv = 1.f;
            const float next_scale = last ? 1.f : f16_scale_for(fmaf(G, __uint_as_float(inb), bmax), &next_inv);

            mbar_wait(bar_acc, static_cast<uint32_t>(l) & 1u);
            tc_fence_after_sync();
            if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[1 + 2 * l] = clock64() - t_start;
            if (touch_s) asm volatile("bar.sync 1, 256;" ::: "memory");      // tid 0's wait_read / aux issue above is done
            if (kind) { mbar_wait(bar_aux, aux_phase); aux_phase ^= 1u; }
            const int act = a.act[l];
            const float4* bl = bias4 + l * (N / 4);
            const bool st = a.store[l] != 0;
            float amax = 0.f;
#define ODEX_F16_EPI(A, K) \
    amax = last ? f16_epilogue<A, K, true, true>(tmem_base, lane_addr, base, s_base, m, bl, alpha, unscale, next_scale, hh, hh + 1) \
                : (st ? f16_epilogue<A, K, true, false>(tmem_base, lane_addr, base, s_base, m, bl, alpha, unscale, next_scale, hh, hh + 1) \
                      : f16_epilogue<A, K, false, false>(tmem_base, lane_addr, base, s_base, m, bl, alpha, unscale, next_scale, hh, hh + 1))
            switch (act * 4 + kind) {
                case ODEX_ACT_NONE * 4 + 0: ODEX_F16_EPI(ODEX_ACT_NONE, 0); break;
                case ODEX_ACT_NONE * 4 + 1: ODEX_F16_EPI(ODEX_ACT_NONE, 1); break;
                case ODEX_ACT_NONE * 4 + 2: if (kTanh) ODEX_F16_EPI(ODEX_ACT_NONE, 2); break;
                case ODEX_ACT_NONE * 4 + 3: ODEX_F16_EPI(ODEX_ACT_NONE, 3); break;
                case ODEX_ACT_RELU * 4 + 0: ODEX_F16_EPI(ODEX_ACT_RELU, 0); break;
                case ODEX_ACT_RELU * 4 + 3: ODEX_F16_EPI(ODEX_ACT_RELU, 3); break;
                case ODEX_ACT_TANH * 4 + 0: if (kTanh) ODEX_F16_EPI(ODEX_ACT_TANH, 0); break;
                case ODEX_ACT_TANH * 4 + 3: if (kTanh) ODEX_F16_EPI(ODEX_ACT_TANH, 3); break;
                default: break;
            }
#undef ODEX_F16_EPI
            if (!last) {
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
                if (lane == 0) asm volatile("atom.shared.max.u32 _, [%0], %1;" ::"r"(amax_s + 4u * (l + 1)), "r"(__float_as_uint(amax)) : "memory");
            }
            cur_inv = next_inv;
            tc_fence_before_sync();
            fence_proxy_async_smem();
            if (!last) mbar_arrive(bar_next);
            if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[2 + 2 * l] = clock64() - t_start;
            asm volatile("bar.sync 1, 256;" ::: "memory");       // every thread has left S (aux read / result written) and amax_s[l+1] is final
            if (a.store[l]) {
                if (tid == 0) {
                    tma_store_4d(&tm.out[l], s_base, 0, 0, 0, img);
                    tma_store_4d(&tm.out[l], s_base + kF16SHalf, 32, 0, 0, img);
                    tma_store_commit();
                }
                stores_pending = true;
            }
        }
        if (tid == 0) tma_store_wait_read();
        // this image's maxima of the chain input and of the outputs of layers 0 .. L-2 (all final since the last layer's closing
        // bar.sync): the operands of the weight gradients, whose kind::f16 kernel then needs no max pass of its own
        if (tid < L && a.amax_out[tid]) {
            uint32_t mb;
            asm volatile("ld.shared.u32 %0, [%1];" : "=r"(mb) : "r"(amax_s + 4u * tid));
            a.amax_out[tid][img] = __uint_as_float(mb);
        }
        tc_fence_before_sync();
        if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[0] = clock64() - t_start;
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

typedef CUresult (*EncodeTiledFnF)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnF f16_encode_fn() {
    static EncodeTiledFnF fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnF>(p);
    });
    return fn;
}

// [B][16][16][64] NHWC fp32 view with pixel-row stride ld; box = 32 channels x 16 x 16 pixels (one half of S)
static int f16_act_map(CUtensorMap* tm, const float* x, int ld, int B) {
    const cuuint64_t dims[4] = {kF16C, 16, 16, static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, 16ull * ld * 4, 256ull * ld * 4};
    const cuuint32_t box[4] = {32, 16, 16, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = f16_encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(convnet f16 activation) failed: %d", static_cast<int>(r));
    return 0;
}

bool convnet_f16_supported(const odex_convnet_desc* d) {
    if (!d || !d->x || d->H != 16 || d->W != 16 || d->C != kF16C || d->B <= 0) return false;
    if (d->L < 1 || d->L > kF16MaxL || !d->out[d->L - 1]) return false;
    if (d->ldx % 4 || !aligned16(d->x)) return false;
    for (int l = 0; l < d->L; ++l) {
        if (!d->w_tc[l] || !aligned16(d->w_tc[l]) || d->act[l] < 0 || d->act[l] > 2 || d->aux_kind[l] < 0 || d->aux_kind[l] > 3) return false;
        if (d->act[l] != ODEX_ACT_NONE && (d->aux_kind[l] == 1 || d->aux_kind[l] == 2)) return false;
        if (d->aux_kind[l] == 3 && l != d->L - 1) return false;     // a residual in the middle of the chain would need its max in the bound
        if (d->aux_kind[l] && (!d->aux[l] || d->ld_aux[l] % 4 || !aligned16(d->aux[l]))) return false;
        if (d->out[l] && (d->ld_out[l] % 4 || !aligned16(d->out[l]))) return false;
        if (d->bias[l] && !aligned16(d->bias[l])) return false;
    }
    return true;
}

int launch_convnet_f16(const odex_convnet_desc* d, cudaStream_t s) {
    if (!convnet_f16_supported(d))
        return fail(ODEX_E_UNSUPPORTED, "convnet(f16): shape not supported (B=%d HxW=%dx%d C=%d L=%d; needs 16x16, C=64, L<=%d, f16-packed weights)",
                    d->B, d->H, d->W, d->C, d->L, kF16MaxL);
    if (!f16_encode_fn()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    F16Maps tm;
    F16Args a;
    if (int rc = f16_act_map(&tm.x, d->x, d->ldx, d->B)) return rc;
    a.L = d->L; a.alpha = d->alpha; a.alpha_dev = d->alpha_dev; a.skip = d->skip; a.dbg = nullptr;
    constexpr long kRowsFloats = 9L * 2 * kF16C * 64 / 2;            // halves of all k-blocks, counted in floats
    for (int l = 0; l < kF16MaxL; ++l) {
        const bool on = l < d->L;
        a.act[l] = on ? d->act[l] : 0;
        a.aux_kind[l] = on ? d->aux_kind[l] : 0;
        a.store[l] = on && d->out[l] ? 1 : 0;
        a.bias[l] = on ? d->bias[l] : nullptr;
        a.wtail[l] = on ? d->w_tc[l] + kRowsFloats : nullptr;
        a.amax_out[l] = on ? d->amax_out[l] : nullptr;
        tm.w[l] = tm.x; tm.aux[l] = tm.x; tm.out[l] = tm.x;
        if (!on) continue;
        {
            const cuuint64_t dims[2] = {64, static_cast<cuuint64_t>(9 * 2 * kF16C)};
            const cuuint64_t strides[1] = {128};
            const cuuint32_t box[2] = {64, 128};
            const cuuint32_t es[2] = {1, 1};
            CUresult r = f16_encode_fn()(&tm.w[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<float*>(d->w_tc[l]), dims, strides,
                                         box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(convnet f16 weights) failed: %d", static_cast<int>(r));
        }
        if (a.aux_kind[l])
            if (int rc = f16_act_map(&tm.aux[l], d->aux[l], d->ld_aux[l], d->B)) return rc;
        if (a.store[l])
            if (int rc = f16_act_map(&tm.out[l], d->out[l], d->ld_out[l], d->B)) return rc;
    }
    bool has_tanh = false;
    for (int l = 0; l < d->L; ++l) has_tanh = has_tanh || d->act[l] == ODEX_ACT_TANH || d->aux_kind[l] == 2;
    auto kern = has_tanh ? convnet_f16_kernel<true> : convnet_f16_kernel<false>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] {
        attr_err = cudaFuncSetAttribute(convnet_f16_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kF16Smem);
        if (attr_err == cudaSuccess)
            attr_err = cudaFuncSetAttribute(convnet_f16_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kF16Smem);
    });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "convnet(f16): cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    static const bool debug = getenv("ODEX_CONV_DEBUG") != nullptr;
    if (debug) {
        static long long* dbuf = nullptr;
        if (!dbuf) cudaMalloc(&dbuf, 16 * sizeof(long long));
        cudaMemsetAsync(dbuf, 0, 16 * sizeof(long long), s);
        a.dbg = dbuf;
    }
    launch_pdl(kern, d->B, kF16Threads, kF16Smem, s, tm, a);
    if (debug) {
        long long h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, a.dbg, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[convnet_f16 L=%d B=%d smem=%d] cta0 cycles=%lld (%.0f per layer); (acc_done, epilogue_done) per layer:", d->L, d->B,
                kF16Smem, h[0], static_cast<double>(h[0]) / d->L);
        for (int l = 0; l < d->L; ++l) fprintf(stderr, " (%lld, %lld)", h[1 + 2 * l], h[2 + 2 * l]);
        fprintf(stderr, "\n");
    }
    return launched("convnet_f16_kernel");
}

}  // namespace odex
