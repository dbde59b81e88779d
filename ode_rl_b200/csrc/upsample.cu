// nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False) of the VidODE per-step decoder (models/VidODE.py:28-44)
// on NHWC tensors, forward and adjoint.  HBM-bound: the forward writes 4x what it reads (16 B/elem of input: 1R + 4W... per
// OUTPUT element 4 B written + 1 B-equivalent read); torch's channels-last kernel runs at ~0.3-0.6 TB/s on these shapes
// (profiles/r01_vidode_forward_cupti.log: 415 us per call), which made it 13 % of the VidODE forward.
// Source index (torch's area_pixel_compute_source_index, align_corners=False, scale 1/2): src = max(0.5*(dst+0.5)-0.5, 0),
// i0 = floor(src), i1 = min(i0+1, H-1), l1 = src - i0, l0 = 1 - l1; out = l0h*(l0w*v00 + l1w*v01) + l1h*(l0w*v10 + l1w*v11).
#include <cuda_runtime.h>

#include "common.cuh"

namespace odex {

__device__ __forceinline__ void up_src(int dst, int n_in, int& i0, int& i1, float& l0, float& l1) {
    float src = 0.5f * (static_cast<float>(dst) + 0.5f) - 0.5f;
    src = src < 0.f ? 0.f : src;
    i0 = static_cast<int>(src);
    i1 = i0 + (i0 < n_in - 1 ? 1 : 0);
    l1 = src - static_cast<float>(i0);
    l0 = 1.f - l1;
}


// one thread = one output pixel x 4 channels; consecutive threads = consecutive channel quads (coalesced both ways)
__global__ void __launch_bounds__(256)
upsample2x_fwd_kernel(const float* __restrict__ x, float* __restrict__ out, int H, int W, int c4, long total) {
    pdl_wait();
    pdl_trigger();
    for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
        const int c = static_cast<int>(i % c4);
        long r = i / c4;
        const int ox = static_cast<int>(r % (2 * W));
        r /= 2 * W;
        const int oy = static_cast<int>(r % (2 * H));
        const long b = r / (2 * H);
        int y0, y1, x0, x1;
        float h0, h1, w0, w1;
        up_src(oy, H, y0, y1, h0, h1);
        up_src(ox, W, x0, x1, w0, w1);
        const float* base = x + (b * H * W) * 4L * c4 + 4 * c;
        const float4 a = ld4(base + (static_cast<long>(y0) * W + x0) * 4L * c4);
        const float4 bb = ld4(base + (static_cast<long>(y0) * W + x1) * 4L * c4);
        const float4 cc = ld4(base + (static_cast<long>(y1) * W + x0) * 4L * c4);
        const float4 d = ld4(base + (static_cast<long>(y1) * W + x1) * 4L * c4);
        float4 o;
        o.x = h0 * (w0 * a.x + w1 * bb.x) + h1 * (w0 * cc.x + w1 * d.x);
        o.y = h0 * (w0 * a.y + w1 * bb.y) + h1 * (w0 * cc.y + w1 * d.y);
        o.z = h0 * (w0 * a.z + w1 * bb.z) + h1 * (w0 * cc.z + w1 * d.z);
        o.w = h0 * (w0 * a.w + w1 * bb.w) + h1 * (w0 * cc.w + w1 * d.w);
        __stcs(reinterpret_cast<float4*>(out) + i, o);
    }
}

// adjoint in gather form: input row iy is read by output rows 2iy-1 .. 2iy+2 with weights (1/4, 3/4, 3/4, 1/4); at the
// borders the clamped neighbour folds its weight onto the edge row (row 0: (0, 1, 3/4, 1/4); row H-1: (1/4, 3/4, 1, 0)).
__device__ __forceinline__ float up_adj_w(int i, int j, int n_in) {
    if (j == 0) return i >= 1 ? 0.25f : 0.f;
    if (j == 1) return i >= 1 ? 0.75f : 1.f;
    if (j == 2) return i < n_in - 1 ? 0.75f : 1.f;
    return i < n_in - 1 ? 0.25f : 0.f;
}

__global__ void __launch_bounds__(256)
upsample2x_bwd_kernel(const float* __restrict__ dout, float* __restrict__ dx, int H, int W, int c4, long total) {
    pdl_wait();
    pdl_trigger();
    for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
        const int c = static_cast<int>(i % c4);
        long r = i / c4;
        const int ix = static_cast<int>(r % W);
        r /= W;
        const int iy = static_cast<int>(r % H);
        const long b = r / H;
        const float* base = dout + (b * 4L * H * W) * 4L * c4 + 4 * c;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int jy = 0; jy < 4; ++jy) {
            const float wy = up_adj_w(iy, jy, H);
            const int oy = 2 * iy - 1 + jy;
            if (wy == 0.f) continue;
#pragma unroll
            for (int jx = 0; jx < 4; ++jx) {
                const float wx = up_adj_w(ix, jx, W);
                const int ox = 2 * ix - 1 + jx;
                if (wx == 0.f) continue;
                const float4 v = ld4(base + (static_cast<long>(oy) * 2 * W + ox) * 4L * c4);
                const float wgt = wy * wx;
                acc.x = fmaf(wgt, v.x, acc.x);
                acc.y = fmaf(wgt, v.y, acc.y);
                acc.z = fmaf(wgt, v.z, acc.z);
                acc.w = fmaf(wgt, v.w, acc.w);
            }
        }
        reinterpret_cast<float4*>(dx)[i] = acc;
    }
}

static int up_check(const void* a, const void* b, int n, int H, int W, int C) {
    if (!a || !b || n <= 0 || H <= 0 || W <= 0) return fail(ODEX_E_ARG, "upsample2x: null pointer or empty tensor");
    if (C <= 0 || (C & 3)) return fail(ODEX_E_UNSUPPORTED, "upsample2x: channels must be a multiple of 4 (got %d)", C);
    if (!aligned16(a) || !aligned16(b)) return fail(ODEX_E_ALIGN, "upsample2x: pointers must be 16-byte aligned");
    return 0;
}

static int up_grid(long total) {
    const long blocks = (total + 255) / 256;
    return static_cast<int>(blocks < 148L * 32 ? blocks : 148L * 32);     // grid-stride beyond 32 CTAs per SM
}

}  // namespace odex

using namespace odex;

extern "C" int odex_upsample2x_fwd(const float* x, int n, int H, int W, int C, float* out, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = up_check(x, out, n, H, W, C)) return rc;
    const long total = static_cast<long>(n) * 4 * H * W * (C / 4);
    launch_pdl(upsample2x_fwd_kernel, up_grid(total), 256, 0, s, x, out, H, W, C / 4, total);
    return launched("upsample2x_fwd_kernel");
}

extern "C" int odex_upsample2x_bwd(const float* dout, int n, int H, int W, int C, float* dx, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = up_check(dout, dx, n, H, W, C)) return rc;
    const long total = static_cast<long>(n) * H * W * (C / 4);
    launch_pdl(upsample2x_bwd_kernel, up_grid(total), 256, 0, s, dout, dx, H, W, C / 4, total);
    return launched("upsample2x_bwd_kernel");
}
