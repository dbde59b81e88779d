// Fused chain of 3x3 64->64 convolutions over 16x16 NHWC fp32 images: the whole ODE function f (forward) or its whole
// data-gradient chain (backward) in ONE launch (C-ABI odex_convnet, include/odex.h).
//
// At config A the conv tile of conv_tc2.cu (16x16 pixels) IS one latent image, so a layer needs no halo from another
// CTA: one CTA keeps its image in shared memory as four 18x18-slot halo planes [hi c0-31][lo c0-31][hi c32-63]
// [lo c32-63] (rows of 128 B, SWIZZLE_128B by absolute address, zero border = padding) and runs every layer on it:
//   MMA phase     exactly conv_tc2's: tap = start address of the K-major descriptor, weights [w_hi ; w_lo] of the layer
//                 streamed through a TMA ring (the ring runs ahead into the next layer during the epilogue);
//                 hi plane = the raw fp32 value (the tensor core truncates tf32 operands), lo = tf32(x - trunc x).
//   epilogue      TMEM -> registers -> bias / act / act'(mask) / res + alpha*v -> written back INTO the halo planes as
//                 the next layer's hi / lo operands.  No TMA reload, no separate split pass, no launch gap.
//   mask/residual the layer's one extra operand tile is TMA-loaded row by row into chunk 0's planes once that chunk's
//                 MMAs have retired (each thread later reads and overwrites only its own pixel slot).
//   stores        a layer output that is needed in global memory (saved activation, gradient for the weight-gradient
//                 kernel, final result) is stored by TMA straight from the hi planes, one 16-pixel row per request
//                 (tools/tma_store_probe.cu: TMA swizzle follows the absolute shared address), overlapping the next
//                 layer's MMAs.
//   warp 0  TMA producer (input halo, weight ring)      warp 1  TMEM allocator + MMA issue
//   warps 2-5  lo-split of the input, then every layer's epilogue
#include <cuda.h>

#include <cstdlib>
#include <mutex>

#include "common.cuh"
#include "ptx_sm100.cuh"

namespace odex {

constexpr int kNetMaxL = ODEX_NET_MAX_LAYERS;
constexpr int kNetC = 64;                       // channels of every layer
constexpr int kNetP = 18;                       // halo pitch in pixel slots
constexpr int kNetPlane = 41984;                // 18*18*128 = 41472 rounded up to 1 KB
constexpr int kNetBStage = 2 * kNetC * 128;     // [w_hi ; w_lo] of one (chunk, tap): 16 KB
constexpr int kNetSB = 3;
constexpr int kNetSmem = 1024 + 4 * kNetPlane + kNetSB * kNetBStage + 256 + kNetMaxL * 256;

struct NetMaps {
    CUtensorMap x;                  // input halo box [32][18][18][1]
    CUtensorMap w[kNetMaxL];        // packed weights, rows of 128 B
    CUtensorMap aux[kNetMaxL];      // row box [32][16][1][1] of the layer's mask / residual operand
    CUtensorMap out[kNetMaxL];      // row box [32][16][1][1] of the layer's stored result
};

struct NetArgs {
    int L;
    int act[kNetMaxL];
    int aux_kind[kNetMaxL];         // 0 none, 1 relu' mask, 2 tanh' mask, 3 residual
    int store[kNetMaxL];
    const float* bias[kNetMaxL];
    float alpha;
    const double* alpha_dev;
    const int* skip;
    long long* dbg;
};

__device__ __forceinline__ uint32_t net_slot(int y, int x) { return static_cast<uint32_t>((y + 1) * kNetP + x + 1); }

// One layer's epilogue for the two pixels (M-tiles) of this thread: TMEM accumulators -> bias -> ACT -> KIND (1 relu' mask,
// 2 tanh' mask, 3 residual; operand read from chunk 0's planes, own slot) -> hi plane (raw fp32) and, unless LAST, lo plane.
template <int ACT, int KIND, bool LAST>
__device__ __forceinline__ void net_epilogue(uint32_t tmem_base, uint32_t lane_addr, uint32_t base, int m, const float4* bl,
                                             float alpha) {
    constexpr int N = kNetC;
#pragma unroll 1
    for (int mt = 0; mt < 2; ++mt) {
        const uint32_t srow = net_slot(m >> 3, mt * 8 + (m & 7));
        const uint32_t sw7 = srow & 7u;
        uint32_t vm[2][32], vc[2][32];
        tmem_ld32(tmem_base + lane_addr + mt * 2 * N, vm[0]);
        tmem_ld32(tmem_base + lane_addr + mt * 2 * N + N, vc[0]);
        tmem_ld32(tmem_base + lane_addr + mt * 2 * N + 32, vm[1]);
        tmem_ld32(tmem_base + lane_addr + mt * 2 * N + N + 32, vc[1]);
        tmem_ld_wait();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
            const int h = 1 - hh;              // channels 32-63 first: their operand lives in hi0, which h = 0 overwrites
            const uint32_t hi_row = base + static_cast<uint32_t>(h) * 2u * kNetPlane + srow * 128u;
            const uint32_t aux_row = base + (h ? 0u : static_cast<uint32_t>(kNetPlane)) + srow * 128u;
            float4 bv[8];
            float xa[8][4];
#pragma unroll
            for (int j = 0; j < 8; ++j) bv[j] = bl[h * 8 + j];
            if (KIND != 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];"
                                 : "=f"(xa[j][0]), "=f"(xa[j][1]), "=f"(xa[j][2]), "=f"(xa[j][3])
                                 : "r"(aux_row + ((static_cast<uint32_t>(j) ^ sw7) << 4)));
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float v[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) v[e] = __uint_as_float(vm[h][4 * j + e]) + __uint_as_float(vc[h][4 * j + e]);
                v[0] += bv[j].x; v[1] += bv[j].y; v[2] += bv[j].z; v[3] += bv[j].w;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (ACT == ODEX_ACT_RELU) v[e] = relu_nan(v[e]);
                    if (ACT == ODEX_ACT_TANH) v[e] = tanhf(v[e]);
                    if (KIND == 1) v[e] = xa[j][e] > 0.f ? v[e] : 0.f;
                    if (KIND == 2) v[e] *= 1.f - xa[j][e] * xa[j][e];
                    if (KIND == 3) v[e] = fmaf(alpha, v[e], xa[j][e]);
                }
                const uint32_t sw = (static_cast<uint32_t>(j) ^ sw7) << 4;
                asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(hi_row + sw), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]));
                if (!LAST) {
                    uint32_t lw[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const uint32_t b = __float_as_uint(v[e]);
                        lw[e] = tf32_rna_bits(__float_as_uint(v[e] - __uint_as_float(b & 0xFFFFE000u)));
                    }
                    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(hi_row + kNetPlane + sw), "r"(lw[0]), "r"(lw[1]), "r"(lw[2]), "r"(lw[3]));
                }
            }
        }
    }
}

// kTanh: some layer uses Tanh (the inlined tanhf is ~60 instructions per 4 values; leaving it out of the ReLU-only
// instantiation keeps the unrolled epilogue small)
template <bool kTanh>
__global__ void __launch_bounds__(192, 1)
convnet_tc_kernel(const __grid_constant__ NetMaps tm, const NetArgs a) {
    constexpr int N = kNetC;
    constexpr int taps = 9;
    constexpr uint32_t kTmemCols = 4 * N;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bst0 = base + 4u * kNetPlane;
    const uint32_t bars = bst0 + kNetSB * kNetBStage;
    const uint32_t bar_raw_full = bars;            // [2]  input halo chunk landed
    const uint32_t bar_split_full = bars + 16;     // [2]  lo plane of the input chunk written (128 arrivals)
    const uint32_t bar_b_full = bars + 32;         // [3]
    const uint32_t bar_b_empty = bars + 64;        // [3]
    const uint32_t bar_acc = bars + 96;            // a layer's MMAs have retired          (one phase per layer)
    const uint32_t bar_early = bars + 104;         // a layer's chunk-0 MMAs have retired  (one phase per layer)
    const uint32_t bar_next = bars + 112;          // the next layer's operands are in the planes (128 arrivals per layer)
    const uint32_t bar_aux = bars + 120;           // a mask / residual tile landed        (one phase per use)
    const uint32_t tmem_slot = bars + 128;
    const uint32_t bias_s = bars + 256;            // [L][64] floats

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int img = blockIdx.x;
    const int L = a.L;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) { mbar_init(bar_raw_full + 8 * i, 1); mbar_init(bar_split_full + 8 * i, 128); }
        for (int i = 0; i < kNetSB; ++i) { mbar_init(bar_b_full + 8 * i, 1); mbar_init(bar_b_empty + 8 * i, 1); }
        mbar_init(bar_acc, 1);
        mbar_init(bar_early, 1);
        mbar_init(bar_next, 128);
        mbar_init(bar_aux, 1);
        fence_barrier_init();
    }
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

    // shared memory / TMEM set-up above may overlap the previous kernel's tail; global memory is first read below
    pdl_wait();
    pdl_trigger();
    const bool skip = a.skip != nullptr && *a.skip != 0;
    const long long t_start = clock64();

    if (skip) {
    } else if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            for (int c = 0; c < 2; ++c) {
                mbar_arrive_expect_tx(bar_raw_full + 8 * c, kNetP * kNetP * 128u);
                tma_load_4d(base + static_cast<uint32_t>(c) * 2u * kNetPlane, &tm.x, bar_raw_full + 8 * c, c * 32, -1, -1, img);
            }
            int sb = 0; uint32_t pb = 0;
            for (int l = 0; l < L; ++l) {
                for (int c = 0; c < 2; ++c) {
                    for (int tap = 0; tap < taps; ++tap) {
                        const int kb = tap * 2 + c;                  // packing order of odex_pack_weights(tc), one 64-ch group
                        mbar_wait(bar_b_empty + 8 * sb, pb ^ 1u);
                        mbar_arrive_expect_tx(bar_b_full + 8 * sb, kNetBStage);
                        tma_load_2d(bst0 + sb * kNetBStage, &tm.w[l], bar_b_full + 8 * sb, 0, kb * 2 * N);
                        if (++sb == kNetSB) { sb = 0; pb ^= 1u; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issue (whole warp runs the loop, one elected lane issues) =====================
        constexpr uint32_t idesc_wide = umma_idesc_tf32(128, 2 * N);
        constexpr uint32_t idesc_narrow = umma_idesc_tf32(128, N);
        const uint32_t a_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, kNetP * 128) >> 32);
        const uint32_t b_hi32 = static_cast<uint32_t>(umma_desc_sw128(0, 16, 1024) >> 32);
        constexpr uint32_t lbo16 = 1u << 16;
        constexpr uint32_t lo_delta = static_cast<uint32_t>(kNetPlane) >> 4;
        int sb = 0; uint32_t pb = 0;
        for (int l = 0; l < L; ++l) {
            if (l > 0) mbar_wait(bar_next, static_cast<uint32_t>(l - 1) & 1u);
            for (int c = 0; c < 2; ++c) {
                if (l == 0) mbar_wait(bar_split_full + 8 * c, 0u);
                tc_fence_after_sync();
                const uint32_t hi = base + static_cast<uint32_t>(c) * 2u * kNetPlane;
                int r = 0, s = 0;
                for (int tap = 0; tap < taps; ++tap) {
                    mbar_wait(bar_b_full + 8 * sb, pb);
                    tc_fence_after_sync();
                    const uint32_t b_lo0 = lbo16 | ((bst0 + sb * kNetBStage) >> 4);
                    const uint32_t a_lo0 = lbo16 | ((hi >> 4) + static_cast<uint32_t>(r * kNetP + s) * 8u);
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
                    if (++sb == kNetSB) { sb = 0; pb ^= 1u; }
                    if (++s == 3) { s = 0; ++r; }
                }
                if (elect_one()) umma_commit(c == 0 ? bar_early : bar_acc);
                __syncwarp();
            }
        }
    } else {
        // ===================== input lo-split, then every layer's epilogue =====================
        const int tid = threadIdx.x - 64;
        for (int i = tid; i < L * (N / 4); i += 128) {
            const int l = i / (N / 4), j = i % (N / 4);
            float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
            if (a.bias[l]) bv = __ldg(reinterpret_cast<const float4*>(a.bias[l]) + j);
            asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(bias_s + 16u * i), "f"(bv.x), "f"(bv.y), "f"(bv.z), "f"(bv.w) : "memory");
        }
        float alpha = a.alpha;
        if (a.alpha_dev) alpha *= static_cast<float>(*a.alpha_dev);
        asm volatile("bar.sync 1, 128;" ::: "memory");
        constexpr int nf4 = kNetP * kNetP * 8;
        for (int c = 0; c < 2; ++c) {
            mbar_wait(bar_raw_full + 8 * c, 0u);
            const uint32_t hi = base + static_cast<uint32_t>(c) * 2u * kNetPlane;
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
                    uint32_t lw[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e)
                        lw[e] = tf32_rna_bits(__float_as_uint(__uint_as_float(x[b][e]) - __uint_as_float(x[b][e] & 0xFFFFE000u)));
                    if (i < nf4)
                        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(hi + kNetPlane + static_cast<uint32_t>(i) * 16u),
                                     "r"(lw[0]), "r"(lw[1]), "r"(lw[2]), "r"(lw[3]) : "memory");
                }
            }
            fence_proxy_async_smem();
            mbar_arrive(bar_split_full + 8 * c);
        }

        const int q = warp & 3;
        const int m = q * 32 + lane;                  // TMEM lane = pixel of the M-tile: (y, x) = (m / 8, mt * 8 + m % 8)
        const uint32_t lane_addr = static_cast<uint32_t>(q * 32) << 16;
        const float4* bias4 = reinterpret_cast<const float4*>(smem_raw + (bias_s - smem_u32(smem_raw)));
        uint32_t aux_phase = 0;
        bool stores_pending = false;
        for (int l = 0; l < L; ++l) {
            const bool last = l == L - 1;
            const int kind = a.aux_kind[l];
            if (kind && tid == 0) {
                // chunk 0's planes are dead for this layer once its MMAs retire: aux channels 32-63 -> hi0, 0-31 -> lo0
                mbar_wait(bar_early, static_cast<uint32_t>(l) & 1u);
                if (stores_pending) tma_store_wait_read();
                mbar_arrive_expect_tx(bar_aux, 256u * N * 4u);
#pragma unroll 1
                for (int y = 0; y < 16; ++y) {
                    const uint32_t off = net_slot(y, 0) * 128u;
                    tma_load_4d(base + off, &tm.aux[l], bar_aux, 32, 0, y, img);
                    tma_load_4d(base + kNetPlane + off, &tm.aux[l], bar_aux, 0, 0, y, img);
                }
            }
            __syncwarp();
            mbar_wait(bar_acc, static_cast<uint32_t>(l) & 1u);
            tc_fence_after_sync();
            if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[1 + 2 * l] = clock64() - t_start;
            if (stores_pending) {
                if (tid == 0) tma_store_wait_read();          // the previous layer's stores have read the hi planes
                asm volatile("bar.sync 1, 128;" ::: "memory");
            }
            if (kind) { mbar_wait(bar_aux, aux_phase); aux_phase ^= 1u; }
            const int act = a.act[l];
            const float4* bl = bias4 + l * (N / 4);
            // one straight-line instantiation per (activation, operand kind, last layer): the unrolled body is fetched once per
            // tile, so every variant it does not contain is instruction-cache footprint saved
#define ODEX_NET_EPI(A, K) \
    (last ? net_epilogue<A, K, true>(tmem_base, lane_addr, base, m, bl, alpha) : net_epilogue<A, K, false>(tmem_base, lane_addr, base, m, bl, alpha))
            switch (act * 4 + kind) {
                case ODEX_ACT_NONE * 4 + 0: ODEX_NET_EPI(ODEX_ACT_NONE, 0); break;
                case ODEX_ACT_NONE * 4 + 1: ODEX_NET_EPI(ODEX_ACT_NONE, 1); break;
                case ODEX_ACT_NONE * 4 + 2: if (kTanh) ODEX_NET_EPI(ODEX_ACT_NONE, 2); break;
                case ODEX_ACT_NONE * 4 + 3: ODEX_NET_EPI(ODEX_ACT_NONE, 3); break;
                case ODEX_ACT_RELU * 4 + 0: ODEX_NET_EPI(ODEX_ACT_RELU, 0); break;
                case ODEX_ACT_RELU * 4 + 3: ODEX_NET_EPI(ODEX_ACT_RELU, 3); break;
                case ODEX_ACT_TANH * 4 + 0: if (kTanh) ODEX_NET_EPI(ODEX_ACT_TANH, 0); break;
                case ODEX_ACT_TANH * 4 + 3: if (kTanh) ODEX_NET_EPI(ODEX_ACT_TANH, 3); break;
                default: break;      // an activation together with a mask is not a chain this kernel is launched for (host check)
            }
#undef ODEX_NET_EPI
            tc_fence_before_sync();
            fence_proxy_async_smem();
            if (!last) mbar_arrive(bar_next);
            if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[2 + 2 * l] = clock64() - t_start;
            stores_pending = false;
            if (a.store[l]) {
                asm volatile("bar.sync 1, 128;" ::: "memory");
                if (tid == 0) {
#pragma unroll 1
                    for (int y = 0; y < 16; ++y) {
                        const uint32_t off = net_slot(y, 0) * 128u;
                        tma_store_4d(&tm.out[l], base + off, 0, 0, y, img);
                        tma_store_4d(&tm.out[l], base + 2u * kNetPlane + off, 32, 0, y, img);
                    }
                    tma_store_commit();
                }
                stores_pending = true;
            }
        }
        if (tid == 0) tma_store_wait_read();
        tc_fence_before_sync();
        if (a.dbg && blockIdx.x == 0 && tid == 0) a.dbg[0] = clock64() - t_start;
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

typedef CUresult (*EncodeTiledFnN)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFnN net_encode_fn() {
    static EncodeTiledFnN fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnN>(p);
    });
    return fn;
}

// [B][16][16][64] NHWC view with pixel-row stride ld; box = 32 channels x bw x bh pixels
static int net_act_map(CUtensorMap* tm, const float* x, int ld, int B, int bw, int bh) {
    const cuuint64_t dims[4] = {kNetC, 16, 16, static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, 16ull * ld * 4, 256ull * ld * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(bw), static_cast<cuuint32_t>(bh), 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = net_encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                                 CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                 CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(convnet activation) failed: %d", static_cast<int>(r));
    return 0;
}

static bool net_supported(const odex_convnet_desc* d) {
    if (!d || !d->x || d->H != 16 || d->W != 16 || d->C != kNetC || d->B <= 0) return false;
    if (d->L < 1 || d->L > kNetMaxL || !d->out[d->L - 1]) return false;
    if (d->ldx % 4 || !aligned16(d->x)) return false;
    for (int l = 0; l < d->L; ++l) {
        if (!d->w_tc[l] || d->act[l] < 0 || d->act[l] > 2 || d->aux_kind[l] < 0 || d->aux_kind[l] > 3) return false;
        if (d->act[l] != ODEX_ACT_NONE && (d->aux_kind[l] == 1 || d->aux_kind[l] == 2)) return false;   // not a chain we build
        if (d->aux_kind[l] && (!d->aux[l] || d->ld_aux[l] % 4 || !aligned16(d->aux[l]))) return false;
        if (d->out[l] && (d->ld_out[l] % 4 || !aligned16(d->out[l]))) return false;
        if (d->bias[l] && !aligned16(d->bias[l])) return false;
    }
    return true;
}

}  // namespace odex

namespace odex {
bool convnet_f16_supported(const odex_convnet_desc* d);
int launch_convnet_f16(const odex_convnet_desc* d, cudaStream_t s);
}  // namespace odex

using namespace odex;

extern "C" int odex_convnet_supported(const odex_convnet_desc* d) {
    if (d && d->split == 1) return convnet_f16_supported(d) ? 1 : 0;
    return net_supported(d) ? 1 : 0;
}

extern "C" int odex_convnet(const odex_convnet_desc* d, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!d) return fail(ODEX_E_ARG, "convnet: null descriptor");
    if (d->split == 1) return launch_convnet_f16(d, s);
    if (d->split != 0) return fail(ODEX_E_ARG, "convnet: unknown operand split %d", d->split);
    for (int l = 0; l < ODEX_NET_MAX_LAYERS; ++l)
        if (d->amax_out[l]) return fail(ODEX_E_ARG, "convnet: amax_out needs split = 1 (the kind::f16 chain measures the maxima anyway)");
    if (!net_supported(d))
        return fail(ODEX_E_UNSUPPORTED, "convnet: shape not supported (B=%d HxW=%dx%d C=%d L=%d; needs 16x16, C=64, L<=%d, packed tc weights)",
                    d->B, d->H, d->W, d->C, d->L, kNetMaxL);
    if (!net_encode_fn()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    NetMaps tm;
    NetArgs a;
    if (int rc = net_act_map(&tm.x, d->x, d->ldx, d->B, kNetP, kNetP)) return rc;
    a.L = d->L; a.alpha = d->alpha; a.alpha_dev = d->alpha_dev; a.skip = d->skip; a.dbg = nullptr;
    for (int l = 0; l < kNetMaxL; ++l) {
        const bool on = l < d->L;
        a.act[l] = on ? d->act[l] : 0;
        a.aux_kind[l] = on ? d->aux_kind[l] : 0;
        a.store[l] = on && d->out[l] ? 1 : 0;
        a.bias[l] = on ? d->bias[l] : nullptr;
        tm.w[l] = tm.x; tm.aux[l] = tm.x; tm.out[l] = tm.x;
        if (!on) continue;
        {
            const cuuint64_t dims[2] = {32, static_cast<cuuint64_t>(2 * 9 * 2 * kNetC)};
            const cuuint64_t strides[1] = {128};
            const cuuint32_t box[2] = {32, 128};
            const cuuint32_t es[2] = {1, 1};
            CUresult r = net_encode_fn()(&tm.w[l], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(d->w_tc[l]), dims, strides,
                                         box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
            if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(convnet weights) failed: %d", static_cast<int>(r));
        }
        if (a.aux_kind[l])
            if (int rc = net_act_map(&tm.aux[l], d->aux[l], d->ld_aux[l], d->B, 16, 1)) return rc;
        if (a.store[l])
            if (int rc = net_act_map(&tm.out[l], d->out[l], d->ld_out[l], d->B, 16, 1)) return rc;
    }
    bool has_tanh = false;
    for (int l = 0; l < d->L; ++l) has_tanh = has_tanh || d->act[l] == ODEX_ACT_TANH || d->aux_kind[l] == 2;
    auto kern = has_tanh ? convnet_tc_kernel<true> : convnet_tc_kernel<false>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] {
        attr_err = cudaFuncSetAttribute(convnet_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kNetSmem);
        if (attr_err == cudaSuccess)
            attr_err = cudaFuncSetAttribute(convnet_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kNetSmem);
    });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "convnet: cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    static const bool debug = getenv("ODEX_CONV_DEBUG") != nullptr;
    if (debug) {
        static long long* dbuf = nullptr;
        if (!dbuf) cudaMalloc(&dbuf, 16 * sizeof(long long));
        cudaMemsetAsync(dbuf, 0, 16 * sizeof(long long), s);
        a.dbg = dbuf;
    }
    launch_pdl(kern, d->B, 192, kNetSmem, s, tm, a);
    if (debug) {
        long long h[16];
        cudaStreamSynchronize(s);
        cudaMemcpy(h, a.dbg, sizeof(h), cudaMemcpyDeviceToHost);
        fprintf(stderr, "[convnet_tc L=%d B=%d smem=%d] cta0 cycles=%lld (%.0f per layer); (acc_done, epilogue_done) per layer:", d->L, d->B,
                kNetSmem, h[0], static_cast<double>(h[0]) / d->L);
        for (int l = 0; l < d->L; ++l) fprintf(stderr, " (%lld, %lld)", h[1 + 2 * l], h[2 + 2 * l]);
        fprintf(stderr, "\n");
    }
    return launched("convnet_tc_kernel");
}
