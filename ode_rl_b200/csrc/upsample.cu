// nn.Upsample(scale_factor=2, mode='bilinear', align_corners=False) of the VidODE per-step decoder (models/VidODE.py:28-44)
// on NHWC tensors, forward and adjoint.  HBM-bound: per input element the forward reads 4 B and writes 16 B, the adjoint the
// reverse; torch's channels-last kernel runs at ~0.3-0.6 TB/s on these shapes (profiles/r01_vidode_forward_cupti.log: 415 us
// per call), which made it 13 % of the VidODE forward.
// Source index (torch's area_pixel_compute_source_index, align_corners=False, scale 1/2): src = max(0.5*(dst+0.5)-0.5, 0),
// i0 = floor(src), i1 = min(i0+1, H-1), l1 = src - i0, l0 = 1 - l1; out = l0h*(l0w*v00 + l1w*v01) + l1h*(l0w*v10 + l1w*v11).
// For scale 2 the weights are the constants 1/4 and 3/4 (and 0 / 1 on the first row / column, where src is clamped).
#include <cuda_runtime.h>

#include "common.cuh"

namespace odex {

__device__ __forceinline__ float4 lerp4(float w0, float4 a, float w1, float4 b) {
    return make_float4(w0 * a.x + w1 * b.x, w0 * a.y + w1 * b.y, w0 * a.z + w1 * b.z, w0 * a.w + w1 * b.w);
}

// One thread = one INPUT pixel x 4 channels -> its 2x2 block of output pixels: 9 loads (3x3 neighbourhood, clamped at the
// border; neighbours are L1 / L2 hits) and 4 streaming stores per thread.  (The first version, one thread per output pixel
// with 64-bit index arithmetic, was issue-bound: ncu 76 % issue slots busy at 21-27 % of DRAM throughput.)
// Even output row 2y reads input rows (y-1, y) with weights (1/4, 3/4) -- (0, 1) on row 0, where torch clamps the source
// index to 0 -- odd row 2y+1 reads (y, min(y+1, H-1)) with (3/4, 1/4); same for columns.  Grouping as in torch:
// out = h0 * (w0 * v00 + w1 * v01) + h1 * (w0 * v10 + w1 * v11).
// grid.x = n * H input rows, grid.y * 256 threads cover W * c4 (pixel, channel-quad) pairs of the row.
__global__ void __launch_bounds__(256)
upsample2x_fwd_kernel(const float* __restrict__ x, float* __restrict__ out, int H, int W, int c4) {
    pdl_wait();
    pdl_trigger();
    const int col = blockIdx.y * 256 + threadIdx.x;             // ix * c4 + c
    if (col >= W * c4) return;
    const int ix = col / c4, c = col - ix * c4;
    const int row = blockIdx.x;                                 // b * H + iy
    const int iy = row % H;
    const int ym = iy > 0 ? -1 : 0, yp = iy < H - 1 ? 1 : 0, xm = ix > 0 ? -1 : 0, xp = ix < W - 1 ? 1 : 0;
    const float he0 = iy > 0 ? 0.25f : 0.f, he1 = iy > 0 ? 0.75f : 1.f;      // even output row: rows (y-1, y)
    const float we0 = ix > 0 ? 0.25f : 0.f, we1 = ix > 0 ? 0.75f : 1.f;      // even output column: columns (x-1, x)
    const long rs = static_cast<long>(W) * c4 * 4;                            // floats per input row
    const float* p = x + static_cast<long>(row) * rs + (static_cast<long>(ix) * c4 + c) * 4;
    float4 ev[3], od[3];                                        // horizontally interpolated rows y-1, y, y+1
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float* q = p + (r == 0 ? ym : r == 2 ? yp : 0) * rs;
        const float4 vm = ld4(q + xm * c4 * 4), vc = ld4(q), vp = ld4(q + xp * c4 * 4);
        ev[r] = lerp4(we0, vm, we1, vc);
        od[r] = lerp4(0.75f, vc, 0.25f, vp);
    }
    const long ors = 2 * rs;                                    // floats per output row
    float* o = out + static_cast<long>(row / H) * (2L * H) * ors + (2L * iy) * ors + (2L * ix * c4 + c) * 4;
    __stcs(reinterpret_cast<float4*>(o), lerp4(he0, ev[0], he1, ev[1]));
    __stcs(reinterpret_cast<float4*>(o + c4 * 4), lerp4(he0, od[0], he1, od[1]));
    __stcs(reinterpret_cast<float4*>(o + ors), lerp4(0.75f, ev[1], 0.25f, ev[2]));
    __stcs(reinterpret_cast<float4*>(o + ors + c4 * 4), lerp4(0.75f, od[1], 0.25f, od[2]));
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
upsample2x_bwd_kernel(const float* __restrict__ dout, float* __restrict__ dx, int H, int W, int c4) {
    pdl_wait();
    pdl_trigger();
    const int col = blockIdx.y * 256 + threadIdx.x;             // ix * c4 + c
    if (col >= W * c4) return;
    const int ix = col / c4, c = col - ix * c4;
    const int row = blockIdx.x;                                 // b * H + iy
    const int iy = row % H;
    const long ors = 2L * W * c4 * 4;                           // floats per output row
    const float* base = dout + static_cast<long>(row / H) * (2L * H) * ors + c * 4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int jy = 0; jy < 4; ++jy) {
        const float wy = up_adj_w(iy, jy, H);                   // 0 exactly where 2*iy-1+jy falls outside the tensor
        const int oy = min(max(2 * iy - 1 + jy, 0), 2 * H - 1);
#pragma unroll
        for (int jx = 0; jx < 4; ++jx) {
            const float wgt = wy * up_adj_w(ix, jx, W);
            const int ox = min(max(2 * ix - 1 + jx, 0), 2 * W - 1);
            const float4 v = ld4(base + oy * ors + static_cast<long>(ox) * c4 * 4);
            acc.x = fmaf(wgt, v.x, acc.x);
            acc.y = fmaf(wgt, v.y, acc.y);
            acc.z = fmaf(wgt, v.z, acc.z);
            acc.w = fmaf(wgt, v.w, acc.w);
        }
    }
    *reinterpret_cast<float4*>(dx + (static_cast<long>(row) * W * c4 + col) * 4) = acc;
}

static int up_check(const void* a, const void* b, int n, int H, int W, int C) {
    if (!a || !b || n <= 0 || H <= 0 || W <= 0) return fail(ODEX_E_ARG, "upsample2x: null pointer or empty tensor");
    if (C <= 0 || (C & 3)) return fail(ODEX_E_UNSUPPORTED, "upsample2x: channels must be a multiple of 4 (got %d)", C);
    if (!aligned16(a) || !aligned16(b)) return fail(ODEX_E_ALIGN, "upsample2x: pointers must be 16-byte aligned");
    return 0;
}

static int up_launch_dims(int n, int H, int W, int C, dim3& grid) {
    const long cols = static_cast<long>(W) * (C / 4);
    if (static_cast<long>(n) * H > 0x7fffffffL || (cols + 255) / 256 > 65535)
        return fail(ODEX_E_UNSUPPORTED, "upsample2x: tensor too large for the launch grid");
    grid = dim3(n * H, static_cast<unsigned>((cols + 255) / 256));       // one CTA row per input row
    return 0;
}

}  // namespace odex

using namespace odex;

extern "C" int odex_upsample2x_fwd(const float* x, int n, int H, int W, int C, float* out, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = up_check(x, out, n, H, W, C)) return rc;
    dim3 grid;
    if (int rc = up_launch_dims(n, H, W, C, grid)) return rc;
    launch_pdl(upsample2x_fwd_kernel, grid, 256, 0, s, x, out, H, W, C / 4);
    return launched("upsample2x_fwd_kernel");
}

extern "C" int odex_upsample2x_bwd(const float* dout, int n, int H, int W, int C, float* dx, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = up_check(dout, dx, n, H, W, C)) return rc;
    dim3 grid;
    if (int rc = up_launch_dims(n, H, W, C, grid)) return rc;
    launch_pdl(upsample2x_bwd_kernel, grid, 256, 0, s, dout, dx, H, W, C / 4);
    return launched("upsample2x_bwd_kernel");
}
