// Host side of the tcgen05 implicit-GEMM convolution (kernel in conv_tc.cuh): geometry, TMA tensor maps, launch.
#include "conv_tc.cuh"

#include <cstdlib>
#include <mutex>

#include "common.cuh"

namespace odex {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
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

// 4-D map over an NHWC activation tensor with pixel-row stride ld: dims (C, W, H, B); box (32, W+2p, TH+2p, 1)
static int make_act_map(CUtensorMap* tm, const float* x, int ld, int C, int B, int H, int W, int TH, int pad) {
    const cuuint64_t dims[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                                static_cast<cuuint64_t>(B)};
    const cuuint64_t strides[3] = {static_cast<cuuint64_t>(ld) * 4, static_cast<cuuint64_t>(W) * ld * 4,
                                   static_cast<cuuint64_t>(H) * W * ld * 4};
    const cuuint32_t box[4] = {32, static_cast<cuuint32_t>(W + 2 * pad), static_cast<cuuint32_t>(TH + 2 * pad), 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    CUresult r = encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(x), dims, strides, box, es,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(activation) failed: %d", static_cast<int>(r));
    return 0;
}

// 2-D map over the packed weights: rows of 32 floats; box = 128 rows
static int make_w_map(CUtensorMap* tm, const float* w, long rows) {
    const cuuint64_t dims[2] = {32, static_cast<cuuint64_t>(rows)};
    const cuuint64_t strides[1] = {128};
    const cuuint32_t box[2] = {32, 128};
    const cuuint32_t es[2] = {1, 1};
    CUresult r = encode_fn()(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(w), dims, strides, box, es,
                             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                             CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled(weights) failed: %d", static_cast<int>(r));
    return 0;
}

bool conv_tc_supported(const odex_conv_desc* d) {
    if (!d->w_tc) return false;
    // more than 128 output channels: N-blocks of 128, one launch each (the VidODE ConvGRU gates: 256)
    if (d->N != 64 && (d->N % 128 != 0 || d->N > 1024)) return false;
    if (d->R != 1 && d->R != 3 && d->R != 5) return false;
    if (d->c0 % 64 != 0) return false;
    if (d->x1 && d->c1 != d->c0) return false;
    if (d->W < 8 || d->W > 128 || (d->W & (d->W - 1)) != 0) return false;
    if (d->H % (128 / d->W) != 0) return false;
    if (d->ld0 % 4 || (d->x1 && d->ld1 % 4) || d->ldo % 4) return false;
    return true;
}

// n_total: output channels of the whole convolution (the packed weights hold [hi ; lo] rows of ALL of them per k-block);
// n_off: first output channel of this launch.  d describes the N-block (N, out, bias, res, dact already offset).
template <int N>
static int launch_tc(const odex_conv_desc* d, cudaStream_t s, int n_total, int n_off) {
    constexpr int kSmemMax = 232448;
    ConvGeom g;
    g.B = d->B; g.H = d->H; g.W = d->W;
    g.nsrc = d->x1 ? 2 : 1; g.csrc = d->c0;
    g.R = d->R; g.TH = 128 / d->W;
    g.log2W = 0;
    while ((1 << g.log2W) < d->W) ++g.log2W;
    const int pad = d->R / 2;
    const int hp = (g.TH + 2 * pad) * (d->W + 2 * pad);
    g.halo_bytes = (hp * 128 + 1023) / 1024 * 1024;
    const int G = g.nsrc * g.csrc / 64;
    g.halo_stages = 1;
    if (G > 1) {
        const int two = conv_smem_bytes(N, g.halo_bytes, 2, ConvCfg<N>::kBStages, 0);
        const int one = conv_smem_bytes(N, g.halo_bytes, 1, ConvCfg<N>::kBStages, 0);
        const bool one_fits_pair = ConvCfg<N>::kMinCtas == 2 && 2 * (one + 1024) <= 233472;
        if (two <= kSmemMax && !one_fits_pair) g.halo_stages = 2;
    }
    const int smem = conv_smem_bytes(N, g.halo_bytes, g.halo_stages, ConvCfg<N>::kBStages, 0);
    if (smem > kSmemMax) return fail(ODEX_E_UNSUPPORTED, "conv(tc): tile needs %d bytes of shared memory", smem);

    if (!encode_fn()) return fail(ODEX_E_DRIVER, "cuTensorMapEncodeTiled not available from the driver");
    CUtensorMap tmA0, tmA1, tmB;
    if (int rc = make_act_map(&tmA0, d->x0, d->ld0, d->c0, d->B, d->H, d->W, g.TH, pad)) return rc;
    if (d->x1) {
        if (int rc = make_act_map(&tmA1, d->x1, d->ld1, d->c1, d->B, d->H, d->W, g.TH, pad)) return rc;
    } else {
        tmA1 = tmA0;
    }
    const long rows = static_cast<long>(G) * d->R * d->R * 2 * 2 * n_total;
    if (int rc = make_w_map(&tmB, d->w_tc, rows)) return rc;
    g.w_rows = 2 * n_total; g.w_lo = n_total; g.w_off = n_off;

    ConvEpilogue ep;
    ep.bias = d->bias; ep.out = d->out; ep.ldo = d->ldo;
    ep.res = d->res; ep.ldr = d->ldr; ep.alpha = d->alpha; ep.alpha_dev = d->alpha_dev;
    ep.dact = d->dact; ep.ldd = d->ldd; ep.dact_kind = d->dact_kind;
    ep.act = d->act; ep.skip = d->skip;

    auto kern = conv_tc_kernel<N, true>;
    static std::once_flag once;
    static cudaError_t attr_err = cudaSuccess;
    std::call_once(once, [&] { attr_err = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemMax); });
    if (attr_err != cudaSuccess) return fail(-static_cast<int>(attr_err), "conv(tc): cudaFuncSetAttribute: %s", cudaGetErrorString(attr_err));
    const int grid = d->B * (d->H / g.TH);
    launch_pdl(kern, grid, ConvCfg<N>::kThreads, smem, s, tmA0, tmA1, tmB, g, ep);
    return launched("conv_tc_kernel");
}

bool conv_tc2_supported(const odex_conv_desc* d);
int launch_conv_tc2(const odex_conv_desc* d, cudaStream_t s);

int launch_conv_tc(const odex_conv_desc* d, cudaStream_t s) {
    // v2 (tap shift = descriptor start address, halo split once) wherever all input channels fit in shared memory
    static const bool no_v2 = getenv("ODEX_CONV_TC_V1") != nullptr;
    if (!no_v2 && conv_tc2_supported(d)) return launch_conv_tc2(d, s);
    if (!conv_tc_supported(d)) return fail(ODEX_E_UNSUPPORTED, "conv(tc): shape not supported (N=%d R=%d c0=%d c1=%d HxW=%dx%d)", d->N, d->R, d->c0, d->c1, d->H, d->W);
    if (d->N == 64) return launch_tc<64>(d, s, 64, 0);
    for (int n_off = 0; n_off < d->N; n_off += 128) {
        odex_conv_desc b = *d;
        b.N = 128;
        b.out = d->out + n_off;
        if (d->bias) b.bias = d->bias + n_off;
        if (d->res) b.res = d->res + n_off;
        if (d->dact) b.dact = d->dact + n_off;
        if (int rc = launch_tc<128>(&b, s, d->N, n_off)) return rc;
    }
    return 0;
}

}  // namespace odex
