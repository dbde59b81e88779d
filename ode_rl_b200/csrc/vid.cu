// The torch-side remainder of the VidODE forward / backward (models/VidODE.py) as HBM-bound NHWC kernels:
//   * BatchNorm2d in training mode (batch statistics, biased variance) fused with its ReLU -- models/VidODE.py:16-19,36
//     (`nn.Conv2d, nn.BatchNorm2d, nn.ReLU` triples of the encoder / decoder): one statistics pass + one apply pass instead of
//     cuDNN's batch-norm kernel plus a separate clamp; backward likewise (one reduction pass + one apply pass);
//   * the recursive flow warp + mask composition of models/VidODE.py:121-131,162-186: per predicted frame
//     grid_sample(prev, grid + flow / ((size - 1) / 2), bilinear, padding_mode='border', align_corners=False) and
//     pred = sigmoid(m) * warped + (1 - sigmoid(m)) * intermediate, read straight from the decoder head's NHWC output.
// Plain CUDA: 128-bit accesses, grids a multiple of the 148 SMs, deterministic two-level reductions (no atomics on the
// statistics), each operand plane read once.
#include <mutex>

#include "common.cuh"

namespace odex {

// ------------------------------------------------------------------------------------------------- BatchNorm + ReLU
// x: [M][C] rows (NHWC with all leading dims flattened), C % 4 == 0, C <= 256.
// Statistics in two levels, both with the CENTRED sum of squares (never E[x^2] - mean^2): a block owns a contiguous slab of
// rows, sums it, re-reads it (from L2: the slab was just read) for sum (x - mean_b)^2, and writes (mean_b, M2_b); the finish
// kernel merges the slabs per channel with Chan's formula in fp64.
constexpr int kBnThreads = 256;

__global__ void __launch_bounds__(kBnThreads) bn_slab_stats_kernel(const float* __restrict__ x, long M, int C, long rows_per_block,
                                                                   float* __restrict__ part) {
    extern __shared__ float sh[];                       // [rows-in-flight][C] partials
    const int c4 = C >> 2;                              // float4 columns
    const int rpi = kBnThreads / c4;                    // rows per iteration
    const int col = threadIdx.x % c4, rr = threadIdx.x / c4;
    const long r0 = blockIdx.x * rows_per_block;
    const long r1 = min(M, r0 + rows_per_block);
    const long n = r1 - r0;
    if (n <= 0) return;
    const bool active = rr < rpi;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active)
        for (long r = r0 + rr; r < r1; r += rpi) {
            const float4 v = ld4(x + r * C + 4 * col);
            s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
        }
    float* shf = sh;
    if (active) *reinterpret_cast<float4*>(shf + (rr * c4 + col) * 4) = s;
    __syncthreads();
    float4 mean = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int k = 0; k < rpi; ++k) {
        const float4 v = *reinterpret_cast<const float4*>(shf + (k * c4 + col) * 4);
        mean.x += v.x; mean.y += v.y; mean.z += v.z; mean.w += v.w;
    }
    const float inv = 1.f / static_cast<float>(n);
    mean.x *= inv; mean.y *= inv; mean.z *= inv; mean.w *= inv;
    __syncthreads();
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active)
        for (long r = r0 + rr; r < r1; r += rpi) {
            const float4 v = ld4(x + r * C + 4 * col);
            const float a = v.x - mean.x, b = v.y - mean.y, c = v.z - mean.z, d = v.w - mean.w;
            q.x = fmaf(a, a, q.x); q.y = fmaf(b, b, q.y); q.z = fmaf(c, c, q.z); q.w = fmaf(d, d, q.w);
        }
    if (active) *reinterpret_cast<float4*>(shf + (rr * c4 + col) * 4) = q;
    __syncthreads();
    if (threadIdx.x < c4) {
        float4 m2 = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int k = 0; k < rpi; ++k) {
            const float4 v = *reinterpret_cast<const float4*>(shf + (k * c4 + col) * 4);
            m2.x += v.x; m2.y += v.y; m2.z += v.z; m2.w += v.w;
        }
        float* dst = part + (static_cast<long>(blockIdx.x) * 2) * C + 4 * col;
        *reinterpret_cast<float4*>(dst) = mean;
        *reinterpret_cast<float4*>(dst + C) = m2;
    }
}

// stats[0][c] = mean, stats[1][c] = rstd = 1 / sqrt(biased var + eps), stats[2][c] = biased var (for the running statistics).
// One block per channel: thread t merges slabs t, t + 128, ... in order, then a fixed shared-memory tree merges the 128 partial
// (n, mean, M2) triples (Chan, fp64): deterministic, and ~5 dependent loads per thread instead of one thread walking all slabs.
__global__ void __launch_bounds__(128) bn_finish_stats_kernel(const float* __restrict__ part, int nblocks, long M, long rows_per_block,
                                                              int C, float eps, float* __restrict__ stats) {
    __shared__ double sn[128], sm[128], sq[128];
    const int c = blockIdx.x, t = threadIdx.x;
    double n = 0.0, mean = 0.0, m2 = 0.0;
    for (int b = t; b < nblocks; b += 128) {
        const long r0 = b * rows_per_block;
        const double nb = static_cast<double>(min(M, r0 + rows_per_block) - r0);
        if (nb <= 0.0) break;
        const double mb = part[(static_cast<long>(b) * 2) * C + c], qb = part[(static_cast<long>(b) * 2 + 1) * C + c];
        const double d = mb - mean, nn = n + nb;
        m2 += qb + d * d * (n * nb / nn);
        mean += d * (nb / nn);
        n = nn;
    }
    sn[t] = n; sm[t] = mean; sq[t] = m2;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (t < o) {
            const double na = sn[t], nb = sn[t + o];
            if (nb > 0.0) {
                const double nn = na + nb, d = sm[t + o] - sm[t];
                sq[t] = sq[t] + sq[t + o] + d * d * (na * nb / nn);
                sm[t] = sm[t] + d * (nb / nn);
                sn[t] = nn;
            }
        }
        __syncthreads();
    }
    if (t == 0) {
        const double var = sq[0] / sn[0];
        stats[c] = static_cast<float>(sm[0]);
        stats[C + c] = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
        stats[2 * C + c] = static_cast<float>(var);
    }
}

// y = relu((x - mean) * rstd * gamma + beta)   (relu != 0), NaN-propagating like torch
__global__ void __launch_bounds__(256) bn_relu_apply_kernel(const float* __restrict__ x, const float* __restrict__ stats,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            long M, int C, int relu, float* __restrict__ y) {
    const int c4 = C >> 2;
    const long n4 = M * c4;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4; i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int col = static_cast<int>(i % c4) * 4;
        const float4 v = ld4(x + 4 * i), mu = ld4(stats + col), rs = ld4(stats + C + col), ga = ld4(gamma + col), be = ld4(beta + col);
        float4 o;
        o.x = fmaf((v.x - mu.x) * rs.x, ga.x, be.x); o.y = fmaf((v.y - mu.y) * rs.y, ga.y, be.y);
        o.z = fmaf((v.z - mu.z) * rs.z, ga.z, be.z); o.w = fmaf((v.w - mu.w) * rs.w, ga.w, be.w);
        if (relu) { o.x = relu_nan(o.x); o.y = relu_nan(o.y); o.z = relu_nan(o.z); o.w = relu_nan(o.w); }
        st4(y + 4 * i, o);
    }
}

// backward, pass 1: per slab and channel  s1 = sum g, s2 = sum g * xhat  with g = dy * (y > 0) (relu) and xhat = (x - mean) * rstd
__global__ void __launch_bounds__(kBnThreads) bn_relu_bwd_reduce_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                                        const float* __restrict__ dy, const float* __restrict__ stats,
                                                                        long M, int C, int relu, long rows_per_block,
                                                                        float* __restrict__ part) {
    extern __shared__ float sh[];
    const int c4 = C >> 2;
    const int rpi = kBnThreads / c4;
    const int col = threadIdx.x % c4, rr = threadIdx.x / c4;
    const long r0 = blockIdx.x * rows_per_block;
    const long r1 = min(M, r0 + rows_per_block);
    if (r1 <= r0) return;
    const bool active = rr < rpi;
    const float4 mu = ld4(stats + 4 * col), rs = ld4(stats + C + 4 * col);
    float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f), s2 = s1;
    if (active)
        for (long r = r0 + rr; r < r1; r += rpi) {
            const long o = r * C + 4 * col;
            const float4 v = ld4(x + o), d = ld4(dy + o);
            float4 g = d;
            if (relu) {
                const float4 yy = ld4(y + o);
                g.x = yy.x > 0.f ? d.x : 0.f; g.y = yy.y > 0.f ? d.y : 0.f; g.z = yy.z > 0.f ? d.z : 0.f; g.w = yy.w > 0.f ? d.w : 0.f;
            }
            s1.x += g.x; s1.y += g.y; s1.z += g.z; s1.w += g.w;
            s2.x = fmaf(g.x, (v.x - mu.x) * rs.x, s2.x); s2.y = fmaf(g.y, (v.y - mu.y) * rs.y, s2.y);
            s2.z = fmaf(g.z, (v.z - mu.z) * rs.z, s2.z); s2.w = fmaf(g.w, (v.w - mu.w) * rs.w, s2.w);
        }
    float* shf = sh;
    if (active) {
        *reinterpret_cast<float4*>(shf + ((rr * 2) * c4 + col) * 4) = s1;
        *reinterpret_cast<float4*>(shf + ((rr * 2 + 1) * c4 + col) * 4) = s2;
    }
    __syncthreads();
    if (threadIdx.x < c4) {
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
        for (int k = 0; k < rpi; ++k) {
            const float4 u = *reinterpret_cast<const float4*>(shf + ((k * 2) * c4 + col) * 4);
            const float4 w = *reinterpret_cast<const float4*>(shf + ((k * 2 + 1) * c4 + col) * 4);
            a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w;
            b.x += w.x; b.y += w.y; b.z += w.z; b.w += w.w;
        }
        float* dst = part + (static_cast<long>(blockIdx.x) * 2) * C + 4 * col;
        *reinterpret_cast<float4*>(dst) = a;
        *reinterpret_cast<float4*>(dst + C) = b;
    }
}

// sums[0][c] = sum g (= dbeta), sums[1][c] = sum g * xhat (= dgamma); one block per channel, fixed order, fp64
__global__ void __launch_bounds__(128) bn_bwd_finish_kernel(const float* __restrict__ part, int nblocks, int C, float* __restrict__ sums) {
    __shared__ double sa[128], sb[128];
    const int c = blockIdx.x, t = threadIdx.x;
    double a = 0.0, b = 0.0;
    for (int k = t; k < nblocks; k += 128) { a += part[(static_cast<long>(k) * 2) * C + c]; b += part[(static_cast<long>(k) * 2 + 1) * C + c]; }
    sa[t] = a; sb[t] = b;
    __syncthreads();
    for (int o = 64; o > 0; o >>= 1) {
        if (t < o) { sa[t] += sa[t + o]; sb[t] += sb[t + o]; }
        __syncthreads();
    }
    if (t == 0) { sums[c] = static_cast<float>(sa[0]); sums[C + c] = static_cast<float>(sb[0]); }
}

// backward, pass 2: dx = gamma * rstd * (g - mean(g) - xhat * mean(g * xhat))
__global__ void __launch_bounds__(256) bn_relu_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ y,
                                                                const float* __restrict__ dy, const float* __restrict__ stats,
                                                                const float* __restrict__ gamma, const float* __restrict__ sums,
                                                                long M, int C, int relu, float* __restrict__ dx) {
    const int c4 = C >> 2;
    const long n4 = M * c4;
    const float invM = 1.f / static_cast<float>(M);
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4; i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int col = static_cast<int>(i % c4) * 4;
        const float4 v = ld4(x + 4 * i), d = ld4(dy + 4 * i), mu = ld4(stats + col), rs = ld4(stats + C + col), ga = ld4(gamma + col);
        const float4 a = ld4(sums + col), b = ld4(sums + C + col);
        float4 g = d;
        if (relu) {
            const float4 yy = ld4(y + 4 * i);
            g.x = yy.x > 0.f ? d.x : 0.f; g.y = yy.y > 0.f ? d.y : 0.f; g.z = yy.z > 0.f ? d.z : 0.f; g.w = yy.w > 0.f ? d.w : 0.f;
        }
        float4 o;
        o.x = ga.x * rs.x * (g.x - a.x * invM - (v.x - mu.x) * rs.x * (b.x * invM));
        o.y = ga.y * rs.y * (g.y - a.y * invM - (v.y - mu.y) * rs.y * (b.y * invM));
        o.z = ga.z * rs.z * (g.z - a.z * invM - (v.z - mu.z) * rs.z * (b.z * invM));
        o.w = ga.w * rs.w * (g.w - a.w * invM - (v.w - mu.w) * rs.w * (b.w * invM));
        st4(dx + 4 * i, o);
    }
}

inline long bn_rows_per_block(long M, int* nblocks) {
    long rpb = (M + 4L * kSMs - 1) / (4L * kSMs);
    rpb = std::max<long>(rpb, 64);
    *nblocks = static_cast<int>((M + rpb - 1) / rpb);
    return rpb;
}

}  // namespace odex

// ------------------------------------------------------------------------------------------------- flow warp + composition
namespace odex {

// One predicted frame k of models/VidODE.py:162-186 + :130, one thread per output pixel:
//   gx = lin_x[x] + flow_x / ((W - 1) / 2), gy = lin_y[y] + flow_y / ((H - 1) / 2)          (the reference's grid + flow)
//   warped = grid_sample(prev, (gx, gy), bilinear, padding_mode = border, align_corners = False)   (torch's arithmetic order)
//   pred   = sigmoid(m) * warped + (1 - sigmoid(m)) * inter
// head: the decoder head's NHWC output [B][H][W][ldh]: channels 0-1 flow, 2..2+C-1 intermediate frame, 2+C mask logit.
// prev / warped / pred: NCHW [B][C][H][W] (frames), C <= 4.
__global__ void __launch_bounds__(256) warp_compose_fwd_kernel(const float* __restrict__ prev, const float* __restrict__ head, int ldh,
                                                               const float* __restrict__ lin_x, const float* __restrict__ lin_y,
                                                               int B, int C, int H, int W, float* __restrict__ warped,
                                                               float* __restrict__ pred) {
    const long total = static_cast<long>(B) * H * W;
    const float cx = (W - 1.0f) / 2.0f, cy = (H - 1.0f) / 2.0f;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total; i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int x = static_cast<int>(i % W), y = static_cast<int>((i / W) % H);
        const long b = i / (static_cast<long>(W) * H);
        const float* hp = head + i * ldh;
        const float4 h0 = ld4(hp);                     // flow_x, flow_y, inter_0, inter_1
        const float4 h1 = ld4(hp + 4);                 // inter_2 | mask ...
        const float gx = lin_x[x] + h0.x / cx, gy = lin_y[y] + h0.y / cy;
        // grid_sampler_compute_source_index: unnormalize (align_corners = False), clip to the border
        float ix = ((gx + 1.f) * W - 1.f) / 2.f, iy = ((gy + 1.f) * H - 1.f) / 2.f;
        ix = fminf(static_cast<float>(W - 1), fmaxf(ix, 0.f));
        iy = fminf(static_cast<float>(H - 1), fmaxf(iy, 0.f));
        const float fx = floorf(ix), fy = floorf(iy);
        const int x0 = static_cast<int>(fx), y0 = static_cast<int>(fy), x1 = x0 + 1, y1 = y0 + 1;
        const float nw = (x1 - ix) * (y1 - iy), ne = (ix - x0) * (y1 - iy), sw = (x1 - ix) * (iy - y0), se = (ix - x0) * (iy - y0);
        const bool vx1 = x1 < W, vy1 = y1 < H;
        const float inter[3] = {h0.z, h0.w, h1.x};
        const float mlogit = C == 3 ? h1.y : (C == 2 ? h1.x : (C == 1 ? h0.w : h1.z));
        const float m = 1.f / (1.f + expf(-mlogit));
        for (int c = 0; c < C; ++c) {
            const float* pc = prev + (b * C + c) * static_cast<long>(H) * W;
            float v = pc[static_cast<long>(y0) * W + x0] * nw;
            if (vx1) v += pc[static_cast<long>(y0) * W + x1] * ne;
            if (vy1) v += pc[static_cast<long>(y1) * W + x0] * sw;
            if (vx1 && vy1) v += pc[static_cast<long>(y1) * W + x1] * se;
            const long o = (b * C + c) * static_cast<long>(H) * W + static_cast<long>(y) * W + x;
            warped[o] = v;
            const float it = c < 3 ? inter[c] : 0.f;
            pred[o] = m * v + (1.f - m) * it;
        }
    }
}

}  // namespace odex

using namespace odex;

extern "C" int odex_bn_relu_fwd(const float* x, long long M, int C, const float* gamma, const float* beta, float eps, int relu,
                                float* y, float* stats, float* scratch, long long scratch_floats, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!x || !y || !gamma || !beta || !stats || !scratch || M <= 0) return fail(ODEX_E_ARG, "bn_relu_fwd: bad argument");
    if (C % 4 != 0 || C < 4 || C > 256) return fail(ODEX_E_SHAPE, "bn_relu_fwd: C must be a multiple of 4 in [4, 256] (got %d)", C);
    if (!aligned16(x) || !aligned16(y) || !aligned16(gamma) || !aligned16(beta) || !aligned16(stats) || !aligned16(scratch))
        return fail(ODEX_E_ALIGN, "bn_relu_fwd: pointers must be 16-byte aligned");
    int nblocks;
    const long rpb = bn_rows_per_block(M, &nblocks);
    if (scratch_floats < 2LL * nblocks * C) return fail(ODEX_E_ARG, "bn_relu_fwd: scratch needs %lld floats", 2LL * nblocks * C);
    const int smem = kBnThreads * 4 * static_cast<int>(sizeof(float));
    bn_slab_stats_kernel<<<nblocks, kBnThreads, smem, s>>>(x, M, C, rpb, scratch);
    if (int rc = launched("bn_slab_stats_kernel")) return rc;
    bn_finish_stats_kernel<<<C, 128, 0, s>>>(scratch, nblocks, M, rpb, C, eps, stats);
    if (int rc = launched("bn_finish_stats_kernel")) return rc;
    const long n4 = M * (C / 4);
    const int blocks = static_cast<int>(std::min<long>((n4 + 255) / 256, 8L * kSMs));
    bn_relu_apply_kernel<<<blocks, 256, 0, s>>>(x, stats, gamma, beta, M, C, relu, y);
    return launched("bn_relu_apply_kernel");
}

// ---- odex_bn_relu_fwd without a caller-provided scratch buffer (no counterpart in the reference; added
// for callers that cannot allocate: CUDA-graph capture, serving loops, harnesses that forbid cudaMalloc).
// The slab statistics go through this file-scope array instead: 2 * 592 * 128 floats = 592 KiB, resident on every GPU that
// loads the library until the process exits.  In host code `g_bn_static_scratch` is only a stand-in for the symbol: its device
// address is taken once per device with cudaGetSymbolAddress (that device current) and that pointer is what the kernels get.
constexpr long long kBnStaticFloats = 2LL * 4 * kSMs * 128;
static __device__ __align__(256) float g_bn_static_scratch[kBnStaticFloats];

static int bn_static_scratch_ptr(float** out) {
    constexpr int kMaxDev = 64;
    static float* ptrs[kMaxDev] = {};
    static std::mutex mu;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(-static_cast<int>(e), "bn_relu_fwd_static: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kMaxDev) return fail(ODEX_E_ARG, "bn_relu_fwd_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
        if (e != cudaSuccess || !p)
            return fail(-static_cast<int>(e), "bn_relu_fwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        ptrs[dev] = static_cast<float*>(p);
    }
    *out = ptrs[dev];
    return 0;
}

extern "C" long long odex_bn_static_scratch_floats(void) { return kBnStaticFloats; }

extern "C" int odex_bn_relu_fwd_static(const float* x, long long M, int C, const float* gamma, const float* beta, float eps,
                                       int relu, float* y, float* stats, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!x || !y || !gamma || !beta || !stats || M <= 0) return fail(ODEX_E_ARG, "bn_relu_fwd_static: bad argument");
    if (C % 4 != 0 || C < 4 || C > 256) return fail(ODEX_E_SHAPE, "bn_relu_fwd_static: C must be a multiple of 4 in [4, 256] (got %d)", C);
    if (!aligned16(x) || !aligned16(y) || !aligned16(gamma) || !aligned16(beta) || !aligned16(stats))
        return fail(ODEX_E_ALIGN, "bn_relu_fwd_static: pointers must be 16-byte aligned");
    int nblocks;
    const long rpb = bn_rows_per_block(M, &nblocks);
    const long long need = 2LL * nblocks * C;
    if (need > kBnStaticFloats)
        return fail(ODEX_E_UNSUPPORTED, "bn_relu_fwd_static: M=%lld C=%d needs %lld floats of scratch, the built-in array holds %lld "
                    "(use odex_bn_relu_fwd with a caller-provided buffer)", M, C, need, kBnStaticFloats);
    float* scratch = nullptr;
    if (int rc = bn_static_scratch_ptr(&scratch)) return rc;
    // every slot the finish kernel reads (2 * nblocks * C) is written by the slab kernel of the same call: nothing to clear
    const int smem = kBnThreads * 4 * static_cast<int>(sizeof(float));
    bn_slab_stats_kernel<<<nblocks, kBnThreads, smem, s>>>(x, M, C, rpb, scratch);
    if (int rc = launched("bn_slab_stats_kernel")) return rc;
    bn_finish_stats_kernel<<<C, 128, 0, s>>>(scratch, nblocks, M, rpb, C, eps, stats);
    if (int rc = launched("bn_finish_stats_kernel")) return rc;
    const long n4 = M * (C / 4);
    const int blocks = static_cast<int>(std::min<long>((n4 + 255) / 256, 8L * kSMs));
    bn_relu_apply_kernel<<<blocks, 256, 0, s>>>(x, stats, gamma, beta, M, C, relu, y);
    return launched("bn_relu_apply_kernel");
}

extern "C" long long odex_bn_scratch_floats(long long M, int C) {
    int nblocks;
    bn_rows_per_block(M, &nblocks);
    return 4LL * nblocks * C;
}

extern "C" int odex_bn_relu_bwd(const float* x, const float* y, const float* dy, long long M, int C, const float* gamma,
                                const float* stats, int relu, float* dx, float* dgamma_dbeta, float* scratch,
                                long long scratch_floats, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!x || !dy || !gamma || !stats || !dx || !dgamma_dbeta || !scratch || M <= 0 || (relu && !y))
        return fail(ODEX_E_ARG, "bn_relu_bwd: bad argument");
    if (C % 4 != 0 || C < 4 || C > 256) return fail(ODEX_E_SHAPE, "bn_relu_bwd: C must be a multiple of 4 in [4, 256] (got %d)", C);
    if (!aligned16(x) || !aligned16(dy) || !aligned16(dx) || !aligned16(gamma) || !aligned16(stats) || !aligned16(scratch) ||
        !aligned16(dgamma_dbeta) || (y && !aligned16(y)))
        return fail(ODEX_E_ALIGN, "bn_relu_bwd: pointers must be 16-byte aligned");
    int nblocks;
    const long rpb = bn_rows_per_block(M, &nblocks);
    if (scratch_floats < 2LL * nblocks * C) return fail(ODEX_E_ARG, "bn_relu_bwd: scratch needs %lld floats", 2LL * nblocks * C);
    const int smem = 2 * kBnThreads * 4 * static_cast<int>(sizeof(float));
    bn_relu_bwd_reduce_kernel<<<nblocks, kBnThreads, smem, s>>>(x, y, dy, stats, M, C, relu, rpb, scratch);
    if (int rc = launched("bn_relu_bwd_reduce_kernel")) return rc;
    // dgamma_dbeta[0][c] = sum g (dbeta), [1][c] = sum g * xhat (dgamma)
    bn_bwd_finish_kernel<<<C, 128, 0, s>>>(scratch, nblocks, C, dgamma_dbeta);
    if (int rc = launched("bn_bwd_finish_kernel")) return rc;
    const long n4 = M * (C / 4);
    const int blocks = static_cast<int>(std::min<long>((n4 + 255) / 256, 8L * kSMs));
    bn_relu_bwd_apply_kernel<<<blocks, 256, 0, s>>>(x, y, dy, stats, gamma, dgamma_dbeta, M, C, relu, dx);
    return launched("bn_relu_bwd_apply_kernel");
}

extern "C" int odex_warp_compose_fwd(const float* prev, const float* head, int ldh, const float* lin_x, const float* lin_y, int B,
                                     int C, int H, int W, float* warped, float* pred, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!prev || !head || !lin_x || !lin_y || !warped || !pred || B <= 0 || H <= 1 || W <= 1)
        return fail(ODEX_E_ARG, "warp_compose_fwd: bad argument");
    if (C < 1 || C > 3 || ldh < C + 3 || ldh % 4 != 0 || ldh < 8 || !aligned16(head))
        return fail(ODEX_E_SHAPE, "warp_compose_fwd: needs 1 <= C <= 3 and a head row of >= 8 floats, 16-byte aligned (C=%d ldh=%d)", C, ldh);
    const long total = static_cast<long>(B) * H * W;
    const int blocks = static_cast<int>(std::min<long>((total + 255) / 256, 8L * kSMs));
    warp_compose_fwd_kernel<<<blocks, 256, 0, s>>>(prev, head, ldh, lin_x, lin_y, B, C, H, W, warped, pred);
    return launched("warp_compose_fwd_kernel");
}
