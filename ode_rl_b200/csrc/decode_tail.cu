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
constexpr int kDtChP = kDtCh + 4;            // cell pitch in shared memory: 132 floats, so the 8 lanes of a 128-bit access that sit
                                             // on neighbouring cells fall on different banks (a 128-float pitch is an 8-way conflict)
constexpr int kDtFwdSmem = (kDtPad * kDtPad * kDtChP + 16 * kDtC) * 4;
constexpr int kDtBwdSmem = (kDtCells * kDtCells * kDtChP + 66 * 66 + 16 * kDtC + 8) * 4;

__device__ __forceinline__ float leaky(float v, float slope) { return v > 0.f ? v : v * slope; }

// fine row 2Y - 1 + u  ->  cell row offset / phase
__device__ __forceinline__ int dt_dr(int u) { return (u + 1) / 2 - 1; }      // -1, 0, 0, 1
__device__ __forceinline__ int dt_ph(int u) { return (u + 1) & 1; }          //  1, 0, 1, 0

__global__ void __launch_bounds__(256, 1)
decode_tail_fwd_kernel(const float* __restrict__ p, const float* __restrict__ w, const float* __restrict__ b, float slope,
                       float* __restrict__ out) {
    extern __shared__ float sm[];
    float* tile = sm;                                   // [18][18][132], zero border
    float* ws = sm + kDtPad * kDtPad * kDtChP;          // [16 taps][32 ci]
    pdl_wait();
    pdl_trigger();
    const int tid = threadIdx.x;
    const float* src = p + static_cast<long>(blockIdx.x) * kDtCells * kDtCells * kDtCh;
    for (int i = tid; i < 16 * kDtC; i += 256) ws[(i % 16) * kDtC + i / 16] = w[i];        // w[ci][tap] -> ws[tap][ci]
    // zero border: 68 cells
    for (int i = tid; i < 68 * (kDtCh / 4); i += 256) {
        const int c = i / (kDtCh / 4), q = i % (kDtCh / 4);
        const int yy = c < 18 ? 0 : (c < 36 ? 17 : 1 + (c - 36) / 2), xx = c < 18 ? c : (c < 36 ? c - 18 : ((c - 36) & 1) * 17);
        st4(tile + (yy * kDtPad + xx) * kDtChP + 4 * q, make_float4(0.f, 0.f, 0.f, 0.f));
    }
    // interior: the image's 128 KB read once, fully coalesced, 8 independent 16-byte loads in flight per thread
#pragma unroll 1
    for (int i0 = 0; i0 < kDtCells * kDtCells * (kDtCh / 4); i0 += 8 * 256) {
        float4 v[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = ld4(src + 4 * (i0 + k * 256 + tid));
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int i = i0 + k * 256 + tid, cell = i >> 5, q = i & 31;
            v[k].x = leaky(v[k].x, slope); v[k].y = leaky(v[k].y, slope); v[k].z = leaky(v[k].z, slope); v[k].w = leaky(v[k].w, slope);
            st4(tile + ((1 + (cell >> 4)) * kDtPad + 1 + (cell & 15)) * kDtChP + 4 * q, v[k]);
        }
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
                h[u][c] = ld4(tile + ((Y + 1 + dt_dr(u)) * kDtPad + (X + 1 + dt_dr(c))) * kDtChP +
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
    float* ht = sm;                                            // [256 cells][132]: h = leaky(P), for the weight gradient
    float* dt = ht + kDtCells * kDtCells * kDtChP;             // [66][66]: dout with a zero border, index (oy + 1, ox + 1)
    float* ws = dt + 66 * 66;                                  // [16 taps][32 ci]
    float* red = ws + 16 * kDtC;                               // [8]
    pdl_wait();
    pdl_trigger();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const long img = blockIdx.x;
    const float* src = p + img * kDtCells * kDtCells * kDtCh;
    const float* dsrc = dout + img * 64 * 64;
    for (int i = tid; i < 16 * kDtC; i += 256) ws[(i % 16) * kDtC + i / 16] = w[i];
    // dout tile: zero border (260 entries), interior by coalesced 16-byte loads
    for (int i = tid; i < 260; i += 256) {
        const int yy = i < 66 ? 0 : (i < 132 ? 65 : 1 + (i - 132) / 2), xx = i < 66 ? i : (i < 132 ? i - 66 : ((i - 132) & 1) * 65);
        dt[yy * 66 + xx] = 0.f;
    }
    float bsum = 0.f;
    {
        float4 v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = ld4(dsrc + 4 * (k * 256 + tid));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int i = k * 256 + tid, yy = 1 + (i >> 4), xx = 1 + 4 * (i & 15);
            float* d = dt + yy * 66 + xx;
            d[0] = v[k].x; d[1] = v[k].y; d[2] = v[k].z; d[3] = v[k].w;
            bsum += (v[k].x + v[k].y) + (v[k].z + v[k].w);
        }
    }
    __syncthreads();
    // ---- phase A: lane = (cell of a group of 4, 16-byte chunk q of the 32 channels): P read and dP written in 128-byte runs;
    // dP[cell][ph][ci] = leaky'(P) * sum_t dout[2fy-1+kh][2fx-1+kw] * w[ci][t]; h = leaky(P) into shared memory
    {
        const int q = lane & 7, sub = lane >> 3;
        float4 wt[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) wt[t] = ld4(ws + t * kDtC + 4 * q);
#pragma unroll 2
        for (int it = 0; it < 8; ++it) {
            const int cell = warp * 32 + it * 4 + sub;
            const int Y = cell >> 4, X = cell & 15;
            float4 pv[4];
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) pv[ph] = ld4(src + cell * kDtCh + ph * kDtC + 4 * q);
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) {
                const int fy = 2 * Y + (ph >> 1), fx = 2 * X + (ph & 1);
                float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
                for (int kh = 0; kh < 4; ++kh) {
                    const float2 d01 = *reinterpret_cast<const float2*>(dt + (2 * fy + kh) * 66 + 2 * fx);
                    const float2 d23 = *reinterpret_cast<const float2*>(dt + (2 * fy + kh) * 66 + 2 * fx + 2);
                    const float dv[4] = {d01.x, d01.y, d23.x, d23.y};
#pragma unroll
                    for (int kw = 0; kw < 4; ++kw) {
                        const float4 ww = wt[kh * 4 + kw];
                        g.x = fmaf(dv[kw], ww.x, g.x); g.y = fmaf(dv[kw], ww.y, g.y);
                        g.z = fmaf(dv[kw], ww.z, g.z); g.w = fmaf(dv[kw], ww.w, g.w);
                    }
                }
                const float4 pp = pv[ph];
                g.x *= pp.x > 0.f ? 1.f : slope; g.y *= pp.y > 0.f ? 1.f : slope;
                g.z *= pp.z > 0.f ? 1.f : slope; g.w *= pp.w > 0.f ? 1.f : slope;
                st4(dp + img * kDtCells * kDtCells * kDtCh + cell * kDtCh + ph * kDtC + 4 * q, g);
                st4(ht + cell * kDtChP + ph * kDtC + 4 * q,
                    make_float4(leaky(pp.x, slope), leaky(pp.y, slope), leaky(pp.z, slope), leaky(pp.w, slope)));
            }
        }
    }
    __syncthreads();
    // ---- phase B: lane = channel, warp = 32 cells; all 16 taps per h value (dout window by broadcast 8-byte loads), then the
    // 8 warps' partial sums meet in shared memory: 512 atomics per image
    if (dw) {
        float acc[16];
#pragma unroll
        for (int t = 0; t < 16; ++t) acc[t] = 0.f;
#pragma unroll 1
        for (int c = 0; c < 32; ++c) {
            const int cell = warp * 32 + c;
            const int Y = cell >> 4, X = cell & 15;
#pragma unroll
            for (int ph = 0; ph < 4; ++ph) {
                const int fy = 2 * Y + (ph >> 1), fx = 2 * X + (ph & 1);
                const float h = ht[cell * kDtChP + ph * kDtC + lane];
#pragma unroll
                for (int kh = 0; kh < 4; ++kh) {
                    const float2 d01 = *reinterpret_cast<const float2*>(dt + (2 * fy + kh) * 66 + 2 * fx);
                    const float2 d23 = *reinterpret_cast<const float2*>(dt + (2 * fy + kh) * 66 + 2 * fx + 2);
                    acc[kh * 4 + 0] = fmaf(h, d01.x, acc[kh * 4 + 0]); acc[kh * 4 + 1] = fmaf(h, d01.y, acc[kh * 4 + 1]);
                    acc[kh * 4 + 2] = fmaf(h, d23.x, acc[kh * 4 + 2]); acc[kh * 4 + 3] = fmaf(h, d23.y, acc[kh * 4 + 3]);
                }
            }
        }
        __syncthreads();                                       // every warp has finished reading ht: reuse it for the partials
#pragma unroll
        for (int t = 0; t < 16; ++t) ht[(warp * 16 + t) * 32 + lane] = acc[t];
        __syncthreads();
        for (int o = tid; o < 16 * kDtC; o += 256) {           // o = t * 32 + ci
            float sacc = 0.f;
#pragma unroll
            for (int ww = 0; ww < 8; ++ww) sacc += ht[ww * 512 + o];
            atomicAdd(dw + (o & 31) * 16 + (o >> 5), sacc);
        }
    }
    if (db) {
        bsum = warp_sum(bsum);
        if (lane == 0) red[warp] = bsum;
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
