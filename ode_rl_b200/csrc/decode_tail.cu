// The tail of ConvDecode (models/ODEConvGRU.py:122-140): LeakyReLU -> ConvTranspose2d(32, 1, 4, 2, 1) on the phase-blocked
// output of the first transposed convolution (modules.conv_transpose4x4s2: P[n][Y][X][(a*2+b)*32 + ci] is the value of
// channel ci at fine pixel (2Y+a, 2X+b), BEFORE the activation).  One channel out: 128 MACs per output pixel against
// 128 bytes read per input pixel -- HBM-bound work that cuDNN runs as FFT / SIMT GEMMs plus layout shuffles.  Here one CTA
// owns one image: the tile is read once (coalesced, LeakyReLU on the fly) and everything else happens on chip.
//   out[oy][ox] = b + sum_ci sum_{kh,kw} h[ci][iy][ix] * w[ci][kh][kw],  oy = 2*iy - 1 + kh, ox = 2*ix - 1 + kw
// Backward: dP = (sum_{kh,kw} dout[2*iy-1+kh][2*ix-1+kw] * w[ci][kh][kw]) * leaky'(P),  dw[ci][kh][kw] += sum h * dout,
// db += sum dout.  Shapes: 16 x 16 cells (64 x 64 output), 32 channels.
#include <cuda_runtime.h>

#include <mutex>

#include "common.cuh"

namespace odex {

constexpr int kDtCells = 16;                 // cells per side
constexpr int kDtC = 32;                     // channels of the hidden layer
constexpr int kDtCh = 4 * kDtC;              // phase-blocked channels per cell
constexpr int kDtPad = kDtCells + 2;
constexpr int kDtFwdSmem = (kDtPad * kDtPad * kDtCh + 16 * kDtC) * 4;
constexpr int kDtBwdSmem = (kDtCells * kDtCells * kDtCh + 66 * 66 + 16 * kDtC + 8) * 4;

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

// fine row 2Y - 1 + u  ->  cell row offset / phase
__device__ __forceinline__ int dt_dr(int u) { return (u + 1) / 2 - 1; }      // -1, 0, 0, 1
__device__ __forceinline__ int dt_ph(int u) { return (u + 1) & 1; }          //  1, 0, 1, 0

__global__ void __launch_bounds__(256, 1)
decode_tail_fwd_kernel(const float* __restrict__ p, const float* __restrict__ w, const float* __restrict__ b, float slope,
                       float* __restrict__ out) {
    extern __shared__ float sm[];
    float* tile = sm;                                   // [18][18][128], zero border
    float* ws = sm + kDtPad * kDtPad * kDtCh;           // [16 taps][32 ci]
    pdl_wait();
    pdl_trigger();
    const int tid = threadIdx.x;
    const float* src = p + static_cast<long>(blockIdx.x) * kDtCells * kDtCells * kDtCh;
    for (int i = tid; i < 16 * kDtC; i += 256) ws[(i % 16) * kDtC + i / 16] = w[i];        // w[ci][tap] -> ws[tap][ci]
    for (int i = tid; i < kDtPad * kDtPad * (kDtCh / 4); i += 256) {
        const int cell = i / (kDtCh / 4), q = i % (kDtCh / 4);
        const int yy = cell / kDtPad, xx = cell % kDtPad;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (yy >= 1 && yy <= kDtCells && xx >= 1 && xx <= kDtCells) {
            v = ld4(src + ((yy - 1) * kDtCells + (xx - 1)) * kDtCh + 4 * q);
            v.x = leaky(v.x, slope); v.y = leaky(v.y, slope); v.z = leaky(v.z, slope); v.w = leaky(v.w, slope);
        }
        st4(tile + cell * kDtCh + 4 * q, v);
    }
    __syncthreads();
    const int Y = tid / kDtCells, X = tid % kDtCells;
    float acc[4][4];
    const float b0 = b ? __ldg(b) : 0.f;
#pragma unroll
    for (int v = 0; v < 4; ++v)
#pragma unroll
        for (int x = 0; x < 4; ++x) acc[v][x] = b0;
#pragma unroll 1
    for (int q = 0; q < kDtC / 4; ++q) {
        float4 h[4][4], wt[16];
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int c = 0; c < 4; ++c)
                h[u][c] = ld4(tile + ((Y + 1 + dt_dr(u)) * kDtPad + (X + 1 + dt_dr(c))) * kDtCh +
                              (dt_ph(u) * 2 + dt_ph(c)) * kDtC + 4 * q);
#pragma unroll
        for (int t = 0; t < 16; ++t) wt[t] = ld4(ws + t * kDtC + 4 * q);
#pragma unroll
        for (int u = 0; u < 4; ++u)
#pragma unroll
            for (int v = 0; v < 4; ++v) {
                const int kh = v + 3 - 2 * u;
                if (kh < 0 || kh > 3) continue;
#pragma unroll
                for (int c = 0; c < 4; ++c)
#pragma unroll
                    for (int x = 0; x < 4; ++x) {
                        const int kw = x + 3 - 2 * c;
                        if (kw < 0 || kw > 3) continue;
                        const float4 a = h[u][c], ww = wt[kh * 4 + kw];
                        acc[v][x] = fmaf(a.x, ww.x, fmaf(a.y, ww.y, fmaf(a.z, ww.z, fmaf(a.w, ww.w, acc[v][x]))));
                    }
            }
    }
    float* dst = out + static_cast<long>(blockIdx.x) * 64 * 64 + (4 * Y) * 64 + 4 * X;
#pragma unroll
    for (int v = 0; v < 4; ++v) st4(dst + v * 64, make_float4(acc[v][0], acc[v][1], acc[v][2], acc[v][3]));
}

__global__ void __launch_bounds__(256, 1)
decode_tail_bwd_kernel(const float* __restrict__ p, const float* __restrict__ dout, const float* __restrict__ w, float slope,
                       float* __restrict__ dp, float* __restrict__ dw, float* __restrict__ db) {
    extern __shared__ float sm[];
    float* ht = sm;                                            // [256 cells][128]: h = leaky(P), for the weight gradient
    float* dt = ht + kDtCells * kDtCells * kDtCh;              // [66][66]: dout with a zero border, index (oy + 1, ox + 1)
    float* ws = dt + 66 * 66;                                  // [16 taps][32 ci]
    float* red = ws + 16 * kDtC;                               // [8]
    pdl_wait();
    pdl_trigger();
    const int tid = threadIdx.x;
    const long img = blockIdx.x;
    const float* src = p + img * kDtCells * kDtCells * kDtCh;
    const float* dsrc = dout + img * 64 * 64;
    for (int i = tid; i < 16 * kDtC; i += 256) ws[(i % 16) * kDtC + i / 16] = w[i];
    float bsum = 0.f;
    for (int i = tid; i < 66 * 66; i += 256) {
        const int yy = i / 66, xx = i % 66;
        float v = 0.f;
        if (yy >= 1 && yy <= 64 && xx >= 1 && xx <= 64) v = dsrc[(yy - 1) * 64 + xx - 1];
        dt[i] = v;
        bsum += v;
    }
    __syncthreads();
    // ---- phase A: thread = cell; dP for its 4 fine pixels x 32 channels, h into shared memory
    {
        const int Y = tid / kDtCells, X = tid % kDtCells;
#pragma unroll 1
        for (int ph = 0; ph < 4; ++ph) {
            const int fy = 2 * Y + (ph >> 1), fx = 2 * X + (ph & 1);
            float dv[16];
#pragma unroll
            for (int kh = 0; kh < 4; ++kh)
#pragma unroll
                for (int kw = 0; kw < 4; ++kw) dv[kh * 4 + kw] = dt[(2 * fy + kh) * 66 + 2 * fx + kw];   // dout[2fy-1+kh][2fx-1+kw]
#pragma unroll 2
            for (int q = 0; q < kDtC / 4; ++q) {
                const int off = tid * kDtCh + ph * kDtC + 4 * q;
                const float4 pv = ld4(src + off);
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int t = 0; t < 16; ++t) {
                    const float4 ww = ld4(ws + t * kDtC + 4 * q);
                    g.x = fmaf(dv[t], ww.x, g.x); g.y = fmaf(dv[t], ww.y, g.y);
                    g.z = fmaf(dv[t], ww.z, g.z); g.w = fmaf(dv[t], ww.w, g.w);
                }
                g.x *= pv.x > 0.f ? 1.f : slope; g.y *= pv.y > 0.f ? 1.f : slope;
                g.z *= pv.z > 0.f ? 1.f : slope; g.w *= pv.w > 0.f ? 1.f : slope;
                st4(dp + img * kDtCells * kDtCells * kDtCh + off, g);
                st4(ht + off, make_float4(leaky(pv.x, slope), leaky(pv.y, slope), leaky(pv.z, slope), leaky(pv.w, slope)));
            }
        }
    }
    __syncthreads();
    // ---- phase B: thread = (channel, pair of taps); dw over the image's 1024 fine pixels
    if (dw) {
        const int ci = tid & 31, t0 = (tid >> 5) * 2;
        const int kh0 = t0 >> 2, kw0 = t0 & 3;                 // taps t0, t0 + 1 share kh (t0 is even)
        float a0 = 0.f, a1 = 0.f;
#pragma unroll 4
        for (int px = 0; px < 4 * kDtCells * kDtCells; ++px) {
            const int cell = px >> 2, ph = px & 3;
            const int fy = 2 * (cell / kDtCells) + (ph >> 1), fx = 2 * (cell % kDtCells) + (ph & 1);
            const float h = ht[cell * kDtCh + ph * kDtC + ci];
            const float* d = dt + (2 * fy + kh0) * 66 + 2 * fx + kw0;
            a0 = fmaf(h, d[0], a0);
            a1 = fmaf(h, d[1], a1);
        }
        atomicAdd(dw + ci * 16 + t0, a0);
        atomicAdd(dw + ci * 16 + t0 + 1, a1);
    }
    if (db) {
        bsum = warp_sum(bsum);
        if ((tid & 31) == 0) red[tid >> 5] = bsum;
        __syncthreads();
        if (tid == 0) {
            float s = 0.f;
            for (int i = 0; i < 8; ++i) s += red[i];
            atomicAdd(db, s);
        }
    }
}

static int dt_check(const void* p, int n, int cells_h, int cells_w, int C) {
    if (!p || n <= 0) return fail(ODEX_E_ARG, "decode_tail: null pointer or empty batch");
    if (cells_h != kDtCells || cells_w != kDtCells || C != kDtC)
        return fail(ODEX_E_UNSUPPORTED, "decode_tail: needs 16x16 cells and 32 channels (got %dx%d, %d)", cells_h, cells_w, C);
    if (!aligned16(p)) return fail(ODEX_E_ALIGN, "decode_tail: pointers must be 16-byte aligned");
    return 0;
}

}  // namespace odex

using namespace odex;

extern "C" int odex_decode_tail_fwd(const float* p, int n, int cells_h, int cells_w, int C, const float* w, const float* b,
                                    float slope, float* out, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = dt_check(p, n, cells_h, cells_w, C)) return rc;
    if (!w || !out || !aligned16(out)) return fail(ODEX_E_ARG, "decode_tail_fwd: null / misaligned pointer");
    static std::once_flag once;
    static cudaError_t err = cudaSuccess;
    std::call_once(once, [] { err = cudaFuncSetAttribute(decode_tail_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDtFwdSmem); });
    if (err != cudaSuccess) return fail(-static_cast<int>(err), "decode_tail_fwd: cudaFuncSetAttribute: %s", cudaGetErrorString(err));
    launch_pdl(decode_tail_fwd_kernel, n, 256, kDtFwdSmem, s, p, w, b, slope, out);
    return launched("decode_tail_fwd_kernel");
}

extern "C" int odex_decode_tail_bwd(const float* p, const float* dout, int n, int cells_h, int cells_w, int C, const float* w,
                                    float slope, float* dp, float* dw, float* db, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = dt_check(p, n, cells_h, cells_w, C)) return rc;
    if (!dout || !w || !dp || !aligned16(dp)) return fail(ODEX_E_ARG, "decode_tail_bwd: null / misaligned pointer");
    static std::once_flag once;
    static cudaError_t err = cudaSuccess;
    std::call_once(once, [] { err = cudaFuncSetAttribute(decode_tail_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kDtBwdSmem); });
    if (err != cudaSuccess) return fail(-static_cast<int>(err), "decode_tail_bwd: cudaFuncSetAttribute: %s", cudaGetErrorString(err));
    launch_pdl(decode_tail_bwd_kernel, n, 256, kDtBwdSmem, s, p, dout, w, slope, dp, dw, db);
    return launched("decode_tail_bwd_kernel");
}
