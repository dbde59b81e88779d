// The stem of ConvEncode (models/ODEConvGRU.py:101-119): Conv2d(1, 16, 3, 2, 1) + LeakyReLU on the 64x64 frames, written
// straight into the SPACE-TO-DEPTH layout the tensor-core strided convolution reads (modules.conv3x3s2):
//   cells[n][Y][X][(a*2+b)*16 + c] = leaky(conv(x))[n][c][2Y+a][2X+b],   Y, X < 16.
// One input channel: 9 MACs per output against 4 B written -- HBM-bound work (21 MB of frames in, 84 MB out at B*T = 1280)
// that cuDNN runs as an SGEMM plus separate activation, layout-shuffle and (backward) wgrad / bias-reduction kernels.
//   forward : one CTA per image; the frame sits in shared memory with a zero border; a work item = (cell, phase, channel quad),
//             consecutive lanes = consecutive 16-byte pieces of the output (fully coalesced stores).
//   backward: dpre = dcells * leaky'(cells) (the sign of the saved output is the sign of the pre-activation, slope > 0);
//             dw[c][kh][kw] += sum dpre * x[2y-1+kh][2x-1+kw], db[c] += sum dpre.  Persistent CTAs keep the 36 + 4 partial sums
//             of their channel quad in registers over all their images; one shuffle / shared-memory reduction and 160 atomics
//             per CTA at the end.  The frames need no gradient (the caller falls back to torch when they do).
#include <cuda_runtime.h>

#include "common.cuh"

namespace odex {

constexpr int kStHW = 64;                    // frame side
constexpr int kStC = 16;                     // output channels
constexpr int kStPad = kStHW + 2;            // padded row length in shared memory (index +1)
constexpr int kStCells = 16;                 // cells per side (output 32x32 = 16x16 cells of 2x2)

__device__ __forceinline__ void stem_load_frame(const float* __restrict__ src, float* tile, int tid) {
    for (int i = tid; i < kStPad * kStPad; i += 256) {
        const int r = i / kStPad, c = i % kStPad;
        if (r == 0 || c == 0 || r == kStPad - 1 || c == kStPad - 1) tile[i] = 0.f;
    }
    for (int i = tid; i < kStHW * kStHW / 4; i += 256) {
        const int r = i / (kStHW / 4), c4 = i % (kStHW / 4);
        const float4 v = ld4(src + r * kStHW + 4 * c4);
        float* d = tile + (r + 1) * kStPad + 4 * c4 + 1;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
}

__global__ void __launch_bounds__(256)
encode_stem_fwd_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b, float slope,
                       float* __restrict__ cells) {
    __shared__ float tile[kStPad * kStPad];
    __shared__ __align__(16) float ws[9 * kStC];          // [tap][c]
    __shared__ __align__(16) float bs[kStC];
    pdl_wait();
    pdl_trigger();
    const int tid = threadIdx.x;
    if (tid < 9 * kStC) ws[(tid % 9) * kStC + tid / 9] = w[tid];            // w[c][tap] -> ws[tap][c]
    if (tid < kStC) bs[tid] = b ? b[tid] : 0.f;
    stem_load_frame(x + static_cast<long>(blockIdx.x) * kStHW * kStHW, tile, tid);
    __syncthreads();
    float* dst = cells + static_cast<long>(blockIdx.x) * kStCells * kStCells * 4 * kStC;
    const int q = tid & 15, cq = q & 3, a = q >> 3, bb = (q >> 2) & 1;
    const float4 bias = *reinterpret_cast<const float4*>(bs + 4 * cq);
#pragma unroll 4
    for (int it = 0; it < 16; ++it) {
        const int cell = it * 16 + (tid >> 4);
        const int oy = 2 * (cell >> 4) + a, ox = 2 * (cell & 15) + bb;
        const float* p = tile + (2 * oy) * kStPad + 2 * ox;             // input (2oy-1, 2ox-1) in padded coordinates
        float4 acc = bias;
#pragma unroll
        for (int kh = 0; kh < 3; ++kh)
#pragma unroll
            for (int kw = 0; kw < 3; ++kw) {
                const float v = p[kh * kStPad + kw];
                const float4 wv = *reinterpret_cast<const float4*>(ws + (kh * 3 + kw) * kStC + 4 * cq);
                acc.x = fmaf(v, wv.x, acc.x); acc.y = fmaf(v, wv.y, acc.y);
                acc.z = fmaf(v, wv.z, acc.z); acc.w = fmaf(v, wv.w, acc.w);
            }
        acc.x = acc.x > 0.f ? acc.x : acc.x * slope; acc.y = acc.y > 0.f ? acc.y : acc.y * slope;
        acc.z = acc.z > 0.f ? acc.z : acc.z * slope; acc.w = acc.w > 0.f ? acc.w : acc.w * slope;
        st4(dst + cell * 4 * kStC + 4 * q, acc);
    }
}

__global__ void __launch_bounds__(256)
encode_stem_bwd_kernel(const float* __restrict__ x, const float* __restrict__ cells, const float* __restrict__ dcells, int n,
                       float slope, float* __restrict__ dw, float* __restrict__ db) {
    __shared__ float tile[kStPad * kStPad];
    __shared__ float red[8][4][40];
    pdl_wait();
    pdl_trigger();
    const int tid = threadIdx.x;
    const int q = tid & 15, a = q >> 3, bb = (q >> 2) & 1;
    float acc[10][4];                                           // 9 taps + the bias sum, for this thread's channel quad
#pragma unroll
    for (int t = 0; t < 10; ++t)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[t][j] = 0.f;
    for (int img = blockIdx.x; img < n; img += gridDim.x) {
        __syncthreads();                                        // the previous image's tile has been consumed
        stem_load_frame(x + static_cast<long>(img) * kStHW * kStHW, tile, tid);
        __syncthreads();
        const long base = static_cast<long>(img) * kStCells * kStCells * 4 * kStC;
#pragma unroll 2
        for (int it = 0; it < 16; ++it) {
            const int cell = it * 16 + (tid >> 4);
            const float4 y = ld4(cells + base + cell * 4 * kStC + 4 * q);
            float4 d = ld4(dcells + base + cell * 4 * kStC + 4 * q);
            d.x *= y.x > 0.f ? 1.f : slope; d.y *= y.y > 0.f ? 1.f : slope;
            d.z *= y.z > 0.f ? 1.f : slope; d.w *= y.w > 0.f ? 1.f : slope;
            const int oy = 2 * (cell >> 4) + a, ox = 2 * (cell & 15) + bb;
            const float* p = tile + (2 * oy) * kStPad + 2 * ox;
#pragma unroll
            for (int kh = 0; kh < 3; ++kh)
#pragma unroll
                for (int kw = 0; kw < 3; ++kw) {
                    const float v = p[kh * kStPad + kw];
                    float* ac = acc[kh * 3 + kw];
                    ac[0] = fmaf(v, d.x, ac[0]); ac[1] = fmaf(v, d.y, ac[1]);
                    ac[2] = fmaf(v, d.z, ac[2]); ac[3] = fmaf(v, d.w, ac[3]);
                }
            acc[9][0] += d.x; acc[9][1] += d.y; acc[9][2] += d.z; acc[9][3] += d.w;
        }
    }
    // lanes with equal (lane & 3) hold the same channel quad: reduce over lane bits 2..4, then over the 8 warps
#pragma unroll
    for (int t = 0; t < 10; ++t)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float v = acc[t][j];
            v += __shfl_xor_sync(0xffffffffu, v, 4);
            v += __shfl_xor_sync(0xffffffffu, v, 8);
            v += __shfl_xor_sync(0xffffffffu, v, 16);
            acc[t][j] = v;
        }
    const int lane = tid & 31, warp = tid >> 5;
    if (lane < 4) {
#pragma unroll
        for (int t = 0; t < 10; ++t)
#pragma unroll
            for (int j = 0; j < 4; ++j) red[warp][lane][t * 4 + j] = acc[t][j];
    }
    __syncthreads();
    if (tid < 160) {
        const int cq = tid / 40, k = tid % 40;
        float s = 0.f;
#pragma unroll
        for (int wp = 0; wp < 8; ++wp) s += red[wp][cq][k];
        const int c = cq * 4 + (k & 3), tap = k >> 2;
        if (tap < 9) {
            if (dw) atomicAdd(dw + c * 9 + tap, s);
        } else if (db) {
            atomicAdd(db + c, s);
        }
    }
}

static int stem_check(const void* x, const void* cells, int n, int H, int W, int C) {
    if (!x || !cells || n <= 0) return fail(ODEX_E_ARG, "encode_stem: null pointer or empty batch");
    if (H != kStHW || W != kStHW || C != kStC)
        return fail(ODEX_E_UNSUPPORTED, "encode_stem: needs 64x64 frames and 16 output channels (got %dx%d, %d)", H, W, C);
    if (!aligned16(x) || !aligned16(cells)) return fail(ODEX_E_ALIGN, "encode_stem: pointers must be 16-byte aligned");
    return 0;
}

}  // namespace odex

using namespace odex;

extern "C" int odex_encode_stem_fwd(const float* x, int n, int H, int W, int C, const float* w, const float* b, float slope,
                                    float* cells, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = stem_check(x, cells, n, H, W, C)) return rc;
    if (!w) return fail(ODEX_E_ARG, "encode_stem_fwd: null weight");
    launch_pdl(encode_stem_fwd_kernel, n, 256, 0, s, x, w, b, slope, cells);
    return launched("encode_stem_fwd_kernel");
}

extern "C" int odex_encode_stem_bwd(const float* x, const float* cells, const float* dcells, int n, int H, int W, int C,
                                    float slope, float* dw, float* db, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = stem_check(x, cells, n, H, W, C)) return rc;
    if (!dcells || !aligned16(dcells)) return fail(ODEX_E_ARG, "encode_stem_bwd: null / misaligned gradient");
    const int grid = n < 148 * 2 ? n : 148 * 2;               // persistent: two CTAs per SM
    launch_pdl(encode_stem_bwd_kernel, grid, 256, 0, s, x, cells, dcells, n, slope, dw, db);
    return launched("encode_stem_bwd_kernel");
}
