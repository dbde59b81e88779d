// Exact-fp32 (FFMA) implicit-GEMM convolution, its weight gradient and the weight packers.
//
// conv_simt_kernel is the always-available CUDA path of odex_conv2d (shapes the tcgen05 kernel does not take,
// e.g. the 1x1 transform_z0 convs, and the cross-check for the tensor-core kernel in tests/).  The weight
// gradient (conv_wgrad_kernel) is the reduction-over-pixels GEMM of the autograd backward.
// Reference call sites replaced: helpers/utils.py:167-177 (3x3), modules/ConvGRUCell.py:41-43,47-49 (5x5),
// modules/ODEConvGRUCell.py:27-30 (1x1) and their autograd backward.
#include <cuda_fp16.h>

#include "common.cuh"

namespace odex {

// ------------------------------------------------------------------------------------------------- packers
// w: OIHW [N][C][R][R].  dgrad: operator of the input gradient = channels swapped + taps flipped:
//   Wd[n'=c][c'=n][r][s] = W[n][c][R-1-r][R-1-s].   Below (n_out, c_in) are the dims of the packed operator.
__device__ __forceinline__ float packed_src(const float* w, int N, int C, int R, int dgrad, int n_out, int c_in,
                                            int r, int s) {
    if (!dgrad) return w[((static_cast<long>(n_out) * C + c_in) * R + r) * R + s];
    return w[((static_cast<long>(c_in) * C + n_out) * R + (R - 1 - r)) * R + (R - 1 - s)];
}

__global__ void pack_simt_kernel(const float* __restrict__ w, int N, int C, int R, int dgrad, float* __restrict__ out) {
    const int No = dgrad ? C : N, Ci = dgrad ? N : C;
    const long total = static_cast<long>(R) * R * Ci * No;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int n = static_cast<int>(i % No);
        const int c = static_cast<int>((i / No) % Ci);
        const int tap = static_cast<int>(i / (static_cast<long>(No) * Ci));
        out[i] = packed_src(w, N, C, R, dgrad, n, c, tap / R, tap % R);
    }
}

__host__ __device__ __forceinline__ uint32_t tf32_round_bits(uint32_t b) { return (b + 0x1000u) & 0xFFFFE000u; }

// TC layout: k-block kb = (grp * taps + tap) * 2 + half covers channels [grp*64 + half*32, +32) of tap.
// Rows [0, No): tf32(hi) ; rows [No, 2No): tf32(w - hi).  Row = 32 floats.
__global__ void pack_tc_kernel(const float* __restrict__ w, int N, int C, int R, int dgrad, float* __restrict__ out) {
    const int No = dgrad ? C : N, Ci = dgrad ? N : C;
    const int taps = R * R;
    const long total = static_cast<long>(Ci / 32) * taps * No * 32;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int kk = static_cast<int>(i & 31);
        const int n = static_cast<int>((i >> 5) % No);
        const long kb = (i >> 5) / No;
        const int half = static_cast<int>(kb & 1);
        const int tap = static_cast<int>((kb >> 1) % taps);
        const int grp = static_cast<int>((kb >> 1) / taps);
        const int c = grp * 64 + half * 32 + kk;
        const float v = packed_src(w, N, C, R, dgrad, n, c, tap / R, tap % R);
        const float hi = __uint_as_float(tf32_round_bits(__float_as_uint(v)));
        const float lo = __uint_as_float(tf32_round_bits(__float_as_uint(v - hi)));
        const long base = kb * (2L * No * 32);
        out[base + static_cast<long>(n) * 32 + kk] = hi;
        out[base + static_cast<long>(No + n) * 32 + kk] = lo;
    }
}

// F16-split layout (tc = 2): fp32 w = (hi + lo * 2^-11) / s with hi = fp16(w * s), lo = fp16((w * s - hi) * 2^11), s a power of
// two chosen so that max |w| * s lies in [2^12, 2^13): 22 significant bits, the accuracy of the 3xTF32 split, on
// tcgen05 kind::f16 at twice the tf32 rate.  k-block kb = grp * taps + tap covers the 64 channels [grp*64, +64) of tap:
// 2*No rows of 64 halves (128 B): rows [0, No) = hi, [No, 2No) = lo.  A 4-float tail follows the rows:
//   tail[0] = 1 / s,  tail[1] = G = max over output channels n of sum_{c,tap} |w[n, c, tap]|  (so |conv(x)[., n]| <= G * max|x|: the
//   kernels size the fp16 scale of a layer's OUTPUT from it before that output exists),  tail[2] = max |w|.
__global__ void pack_f16_stats_kernel(const float* __restrict__ w, int N, int C, int R, int dgrad, float* __restrict__ tail) {
    const int No = dgrad ? C : N, Ci = dgrad ? N : C;
    const int n = blockIdx.x;
    if (n >= No) return;
    __shared__ float sh[33];
    float sum = 0.f, mx = 0.f;
    const int per = Ci * R * R;
    for (int i = threadIdx.x; i < per; i += blockDim.x) {
        const int c = i / (R * R), t = i % (R * R);
        const float v = fabsf(packed_src(w, N, C, R, dgrad, n, c, t / R, t % R));
        sum += v;
        mx = fmaxf(mx, v);
    }
    sum = block_sum(sum, sh);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned int*>(tail + 2), __float_as_uint(mx));
    if (threadIdx.x == 0) atomicMax(reinterpret_cast<unsigned int*>(tail + 1), __float_as_uint(sum));
}

__device__ __forceinline__ float f16_pack_scale(float maxabs) {
    // power of two s with maxabs * s in [2^12, 2^13); 1 for an all-zero (or non-finite) tensor
    if (!(maxabs > 0.f) || !isfinite(maxabs)) return 1.f;
    int e;
    frexpf(maxabs, &e);                       // maxabs = m * 2^e, m in [0.5, 1)
    return ldexpf(1.f, 13 - e);
}

__global__ void pack_f16_kernel(const float* __restrict__ w, int N, int C, int R, int dgrad, __half* __restrict__ out,
                                float* __restrict__ tail) {
    const int No = dgrad ? C : N, Ci = dgrad ? N : C;
    const int taps = R * R;
    const float s = f16_pack_scale(tail[2]);
    const long total = static_cast<long>(Ci / 64) * taps * No * 64;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < total;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const int kk = static_cast<int>(i & 63);
        const int n = static_cast<int>((i >> 6) % No);
        const long kb = (i >> 6) / No;
        const int tap = static_cast<int>(kb % taps);
        const int grp = static_cast<int>(kb / taps);
        const float v = packed_src(w, N, C, R, dgrad, n, grp * 64 + kk, tap / R, tap % R) * s;
        const __half hi = __float2half_rn(v);
        const __half lo = __float2half_rn((v - __half2float(hi)) * 2048.f);
        const long base = kb * (2L * No * 64);
        out[base + static_cast<long>(n) * 64 + kk] = hi;
        out[base + static_cast<long>(No + n) * 64 + kk] = lo;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) tail[0] = 1.f / s;
}

// Both orientations of one weight tensor in one pair of launches (odex_pack_weights_f16_pair): blocks [0, N) are the forward
// operator's rows, blocks [N, N + C) the data-gradient operator's; same arithmetic as the single-orientation kernels.
__global__ void pack_f16_stats2_kernel(const float* __restrict__ w, int N, int C, int R, float* __restrict__ tail_f,
                                       float* __restrict__ tail_d) {
    const int dgrad = static_cast<int>(blockIdx.x) >= N;
    const int No = dgrad ? C : N, Ci = dgrad ? N : C;
    const int n = dgrad ? static_cast<int>(blockIdx.x) - N : static_cast<int>(blockIdx.x);
    if (n >= No) return;
    float* tail = dgrad ? tail_d : tail_f;
    __shared__ float sh[33];
    float sum = 0.f, mx = 0.f;
    const int per = Ci * R * R;
    for (int i = threadIdx.x; i < per; i += blockDim.x) {
        const int c = i / (R * R), t = i % (R * R);
        const float v = fabsf(packed_src(w, N, C, R, dgrad, n, c, t / R, t % R));
        sum += v;
        mx = fmaxf(mx, v);
    }
    sum = block_sum(sum, sh);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    if ((threadIdx.x & 31) == 0) atomicMax(reinterpret_cast<unsigned int*>(tail + 2), __float_as_uint(mx));
    if (threadIdx.x == 0) atomicMax(reinterpret_cast<unsigned int*>(tail + 1), __float_as_uint(sum));
}

__global__ void pack_f16_2_kernel(const float* __restrict__ w, int N, int C, int R, __half* __restrict__ out_f, float* __restrict__ tail_f,
                                  __half* __restrict__ out_d, float* __restrict__ tail_d) {
    const int taps = R * R;
    const long total = static_cast<long>(N) * C * taps;            // elements per orientation
    const float s_f = f16_pack_scale(tail_f[2]), s_d = f16_pack_scale(tail_d[2]);
    for (long j = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; j < 2 * total;
         j += static_cast<long>(gridDim.x) * blockDim.x) {
        const int dgrad = j >= total;
        const long i = dgrad ? j - total : j;
        const int No = dgrad ? C : N;
        const int kk = static_cast<int>(i & 63);
        const int n = static_cast<int>((i >> 6) % No);
        const long kb = (i >> 6) / No;
        const int tap = static_cast<int>(kb % taps);
        const int grp = static_cast<int>(kb / taps);
        const float v = packed_src(w, N, C, R, dgrad, n, grp * 64 + kk, tap / R, tap % R) * (dgrad ? s_d : s_f);
        const __half hi = __float2half_rn(v);
        const __half lo = __float2half_rn((v - __half2float(hi)) * 2048.f);
        const long base = kb * (2L * No * 64);
        __half* out = dgrad ? out_d : out_f;
        out[base + static_cast<long>(n) * 64 + kk] = hi;
        out[base + static_cast<long>(No + n) * 64 + kk] = lo;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) { tail_f[0] = 1.f / s_f; tail_d[0] = 1.f / s_d; }
}

// ------------------------------------------------------------------------------------------------- forward
struct SimtConvArgs {
    const float* x0; int ld0; int c0;
    const float* x1; int ld1; int c1;
    int B, H, W, N;
    const float* w;      // [tap][Ctot][N]
    const float* bias;
    int act;
    const float* dact; int ldd; int dact_kind;
    const float* res; int ldr;
    float alpha; const double* alpha_dev;
    float* out; int ldo;
    const int* skip;
};

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == ODEX_ACT_RELU) return relu_nan(v);
    if (act == ODEX_ACT_TANH) return tanhf(v);
    return v;
}
__device__ __forceinline__ float act_deriv(float a, int kind) {
    if (kind == ODEX_ACT_RELU) return a > 0.f ? 1.f : 0.f;
    if (kind == ODEX_ACT_TANH) return 1.f - a * a;
    return 1.f;
}

// CTA tile: 8 rows x 16 cols of one image x 64 output channels; thread tile 8 pixels (one row segment) x 4 ch.
template <int R>
__global__ void __launch_bounds__(256) conv_simt_kernel(const SimtConvArgs a) {
    constexpr int KC = 16, TH = 8, TW = 16, PAD = R / 2;
    constexpr int HWp = TW + R - 1, HP = (TH + R - 1) * HWp;
    constexpr int HPP = ((HP + 5) / 8) * 8 + 2;          // == 2 (mod 8): conflict-free transposed stores
    __shared__ float halo[KC][HPP];
    __shared__ __align__(16) float wrow[R][KC][64];
    if (a.skip != nullptr && *a.skip != 0) return;

    const int tid = threadIdx.x;
    const int tiles_x = a.W / TW, tiles_y = a.H / TH;
    int t = blockIdx.x;
    const int tx0 = (t % tiles_x) * TW; t /= tiles_x;
    const int ty0 = (t % tiles_y) * TH; t /= tiles_y;
    const int b = t;
    const int n0 = blockIdx.y * 64;
    const int cg = tid & 15, pg = tid >> 4;
    const int ty = pg >> 1, x0 = (pg & 1) * 8;
    const int Ctot = a.c0 + a.c1;

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int cb = 0; cb < Ctot; cb += KC) {
        const float* src; int ld, coff;
        if (cb < a.c0) { src = a.x0; ld = a.ld0; coff = cb; } else { src = a.x1; ld = a.ld1; coff = cb - a.c0; }
        __syncthreads();
        for (int idx = tid; idx < HP * (KC / 4); idx += 256) {
            const int q = idx & 3, pix = idx >> 2;
            const int hy = pix / HWp, hx = pix - hy * HWp;
            const int gy = ty0 + hy - PAD, gx = tx0 + hx - PAD;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W)
                v = ld4(src + (static_cast<long>(b * a.H + gy) * a.W + gx) * ld + coff + q * 4);
            halo[q * 4 + 0][pix] = v.x; halo[q * 4 + 1][pix] = v.y;
            halo[q * 4 + 2][pix] = v.z; halo[q * 4 + 3][pix] = v.w;
        }
        for (int r = 0; r < R; ++r) {
            __syncthreads();
            for (int idx = tid; idx < R * KC * 16; idx += 256) {
                const int n4 = idx & 15, k = (idx >> 4) % KC, s = idx / (16 * KC);
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                const int n = n0 + n4 * 4;
                if (n < a.N) v = ld4(a.w + (static_cast<long>(r * R + s) * Ctot + cb + k) * a.N + n);
                st4(&wrow[s][k][n4 * 4], v);
            }
            __syncthreads();
#pragma unroll 4
            for (int k = 0; k < KC; ++k) {
                float av[8 + R - 1];
                const float* hp = &halo[k][(ty + r) * HWp + x0];
#pragma unroll
                for (int i = 0; i < 8 + R - 1; ++i) av[i] = hp[i];
#pragma unroll
                for (int s = 0; s < R; ++s) {
                    const float4 bv = ld4(&wrow[s][k][cg * 4]);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        acc[i][0] = fmaf(av[i + s], bv.x, acc[i][0]);
                        acc[i][1] = fmaf(av[i + s], bv.y, acc[i][1]);
                        acc[i][2] = fmaf(av[i + s], bv.z, acc[i][2]);
                        acc[i][3] = fmaf(av[i + s], bv.w, acc[i][3]);
                    }
                }
            }
        }
    }

    const int n = n0 + cg * 4;
    if (n >= a.N) return;
    float bias[4] = {0.f, 0.f, 0.f, 0.f};
    if (a.bias) { const float4 bb = ld4(a.bias + n); bias[0] = bb.x; bias[1] = bb.y; bias[2] = bb.z; bias[3] = bb.w; }
    float alpha = a.alpha;
    if (a.alpha_dev) alpha *= static_cast<float>(*a.alpha_dev);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const long pix = static_cast<long>(b * a.H + ty0 + ty) * a.W + tx0 + x0 + i;
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = apply_act(acc[i][j] + bias[j], a.act);
        if (a.dact) {
            const float4 d = ld4(a.dact + pix * a.ldd + n);
            v[0] *= act_deriv(d.x, a.dact_kind); v[1] *= act_deriv(d.y, a.dact_kind);
            v[2] *= act_deriv(d.z, a.dact_kind); v[3] *= act_deriv(d.w, a.dact_kind);
        }
        if (a.res) {
            const float4 rr = ld4(a.res + pix * a.ldr + n);
            v[0] = fmaf(alpha, v[0], rr.x); v[1] = fmaf(alpha, v[1], rr.y);
            v[2] = fmaf(alpha, v[2], rr.z); v[3] = fmaf(alpha, v[3], rr.w);
        }
        st4(a.out + pix * a.ldo + n, make_float4(v[0], v[1], v[2], v[3]));
    }
}

int launch_conv_simt(const odex_conv_desc* d, cudaStream_t s) {
    if (d->W % 16 != 0 || d->H % 8 != 0) return fail(ODEX_E_SHAPE, "conv(simt): need W %% 16 == 0 and H %% 8 == 0, got %dx%d", d->H, d->W);
    if (d->c0 % 16 != 0 || (d->x1 && d->c1 % 16 != 0) || d->N % 4 != 0)
        return fail(ODEX_E_SHAPE, "conv(simt): need c0, c1 %% 16 == 0 and N %% 4 == 0 (c0=%d c1=%d N=%d)", d->c0, d->c1, d->N);
    if (!d->w_simt) return fail(ODEX_E_ARG, "conv(simt): w_simt is NULL");
    SimtConvArgs a;
    a.x0 = d->x0; a.ld0 = d->ld0; a.c0 = d->c0;
    a.x1 = d->x1; a.ld1 = d->ld1; a.c1 = d->x1 ? d->c1 : 0;
    a.B = d->B; a.H = d->H; a.W = d->W; a.N = d->N;
    a.w = d->w_simt; a.bias = d->bias; a.act = d->act;
    a.dact = d->dact; a.ldd = d->ldd; a.dact_kind = d->dact_kind;
    a.res = d->res; a.ldr = d->ldr; a.alpha = d->alpha; a.alpha_dev = d->alpha_dev;
    a.out = d->out; a.ldo = d->ldo; a.skip = d->skip;
    dim3 grid(d->B * (d->H / 8) * (d->W / 16), (d->N + 63) / 64);
    switch (d->R) {
        case 1: conv_simt_kernel<1><<<grid, 256, 0, s>>>(a); break;
        case 3: conv_simt_kernel<3><<<grid, 256, 0, s>>>(a); break;
        case 5: conv_simt_kernel<5><<<grid, 256, 0, s>>>(a); break;
        default: return fail(ODEX_E_SHAPE, "conv: R must be 1, 3 or 5 (got %d)", d->R);
    }
    return launched("conv_simt_kernel");
}

// ------------------------------------------------------------------------------------------------- wgrad
// dw[n][c][r][s] += sum_p dy[p][n] * X[p + (r,s) - pad][c].  One CTA = one tap x 64 c x 64 n x a pixel range;
// 256 threads x (4 c x 4 n); pixels streamed through smem 32 at a time.  Partial tiles are added with atomics.
struct WgradArgs {
    const float* x0; int ld0; int c0;
    const float* x1; int ld1; int c1;
    const float* dy; int ldy;
    int B, H, W, R, N;
    float* dw;
    int nsplit;
    int packed;      // 0: dw is torch OIHW; 1: dw is [tap][N][Ctot] (the accumulation layout of the tensor-core kernel)
};

__global__ void __launch_bounds__(256) conv_wgrad_kernel(const WgradArgs a) {
    constexpr int KP = 32;
    __shared__ __align__(16) float As[KP][64];
    __shared__ __align__(16) float Bs[KP][64];
    const int tid = threadIdx.x;
    const int Ctot = a.c0 + a.c1;
    const int cblks = Ctot / 64, nblks = (a.N + 63) / 64;
    int t = blockIdx.x;
    const int split = t % a.nsplit; t /= a.nsplit;
    const int nb = t % nblks; t /= nblks;
    const int cb = t % cblks; t /= cblks;
    const int tap = t;
    const int r = tap / a.R, s = tap % a.R, pad = a.R / 2;
    const int cbase = cb * 64;
    const float* src; int ld, coff;
    if (cbase < a.c0) { src = a.x0; ld = a.ld0; coff = cbase; } else { src = a.x1; ld = a.ld1; coff = cbase - a.c0; }
    const long P = static_cast<long>(a.B) * a.H * a.W;
    const long per = ((P + a.nsplit - 1) / a.nsplit + KP - 1) / KP * KP;
    const long p_begin = split * per, p_end = min(P, p_begin + per);
    const int tc = tid >> 4, tn = tid & 15;

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (long p0 = p_begin; p0 < p_end; p0 += KP) {
        __syncthreads();
#pragma unroll
        for (int it = 0; it < 2; ++it) {
            const int idx = tid + it * 256;
            const int pp = idx >> 4, q = idx & 15;
            const long p = p0 + pp;
            float4 xa = make_float4(0.f, 0.f, 0.f, 0.f), yb = xa;
            if (p < p_end) {
                const int x = static_cast<int>(p % a.W);
                const int y = static_cast<int>((p / a.W) % a.H);
                const int gy = y + r - pad, gx = x + s - pad;
                if (gy >= 0 && gy < a.H && gx >= 0 && gx < a.W)
                    xa = ld4(src + (p + static_cast<long>(r - pad) * a.W + (s - pad)) * ld + coff + q * 4);
                const int n = nb * 64 + q * 4;
                if (n < a.N) yb = ld4(a.dy + p * a.ldy + n);
            }
            st4(&As[pp][q * 4], xa);
            st4(&Bs[pp][q * 4], yb);
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < KP; ++k) {
            const float4 av = ld4(&As[k][tc * 4]);
            const float4 bv = ld4(&Bs[k][tn * 4]);
            const float aa[4] = {av.x, av.y, av.z, av.w};
            const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(aa[i], bb[j], acc[i][j]);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int c = cbase + tc * 4 + i;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = nb * 64 + tn * 4 + j;
            if (n < a.N) {
                float* dst = a.packed ? a.dw + (static_cast<long>(tap) * a.N + n) * Ctot + c
                                      : a.dw + ((static_cast<long>(n) * Ctot + c) * a.R + r) * a.R + s;
                atomicAdd(dst, acc[i][j]);
            }
        }
    }
}

// db[n] += sum_p dy[p][n].  Thread -> (row lane, 4-channel column); the row lanes of a block are reduced through
// shared memory so that each block issues N atomics (not 256 x 4: contended atomics on N addresses were 24 % of the
// training step in the first profile).
__global__ void __launch_bounds__(256) colsum_kernel(const float* __restrict__ dy, int ldy, long P, int N, float* db) {
    __shared__ float4 part[256];
    const int ncol4 = N / 4;
    const int col = threadIdx.x % ncol4;
    const int rl = threadIdx.x / ncol4;
    const int rlanes = blockDim.x / ncol4;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (rl < rlanes) {
        const long stride = static_cast<long>(gridDim.x) * rlanes;
        long p = static_cast<long>(blockIdx.x) * rlanes + rl;
        // two independent loads in flight per thread
        for (; p + stride < P; p += 2 * stride) {
            const float4 v = ld4(dy + p * ldy + col * 4);
            const float4 w = ld4(dy + (p + stride) * ldy + col * 4);
            acc.x += v.x + w.x; acc.y += v.y + w.y; acc.z += v.z + w.z; acc.w += v.w + w.w;
        }
        if (p < P) {
            const float4 v = ld4(dy + p * ldy + col * 4);
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    if (threadIdx.x < ncol4) {
        float4 t = part[threadIdx.x];
        for (int r = 1; r < rlanes; ++r) {
            const float4 v = part[r * ncol4 + threadIdx.x];
            t.x += v.x; t.y += v.y; t.z += v.z; t.w += v.w;
        }
        atomicAdd(db + col * 4 + 0, t.x); atomicAdd(db + col * 4 + 1, t.y);
        atomicAdd(db + col * 4 + 2, t.z); atomicAdd(db + col * 4 + 3, t.w);
    }
}

}  // namespace odex

using namespace odex;

extern "C" size_t odex_packed_weight_floats(int n_out, int c_in, int R, int tc) {
    const size_t base = static_cast<size_t>(n_out) * c_in * R * R;
    if (tc == 2) return base + 4;            // [hi ; lo] halves = base floats, + the 4-float tail
    return tc ? 2 * base : base;
}

extern "C" int odex_pack_weights(const float* w, int N, int C, int R, int dgrad, int tc, float* out, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!w || !out || N <= 0 || C <= 0 || R <= 0) return fail(ODEX_E_ARG, "pack_weights: bad argument");
    const long total = static_cast<long>(N) * C * R * R;
    const int blocks = static_cast<int>(std::min<long>((total + 255) / 256, 8L * kSMs));
    if (tc == 2) {
        const int Ci = dgrad ? N : C, No = dgrad ? C : N;
        if (Ci % 64 != 0) return fail(ODEX_E_SHAPE, "pack_weights(f16): input channels %d not a multiple of 64", Ci);
        if (!aligned16(out)) return fail(ODEX_E_ALIGN, "pack_weights(f16): out not 16-byte aligned");
        float* tail = out + total;
        cudaMemsetAsync(tail, 0, 4 * sizeof(float), s);
        pack_f16_stats_kernel<<<No, 128, 0, s>>>(w, N, C, R, dgrad, tail);
        int rc = launched("pack_f16_stats_kernel");
        if (rc) return rc;
        pack_f16_kernel<<<blocks, 256, 0, s>>>(w, N, C, R, dgrad, reinterpret_cast<__half*>(out), tail);
        return launched("pack_f16_kernel");
    }
    if (tc) {
        const int Ci = dgrad ? N : C;
        if (Ci % 64 != 0) return fail(ODEX_E_SHAPE, "pack_weights(tc): input channels %d not a multiple of 64", Ci);
        pack_tc_kernel<<<blocks, 256, 0, s>>>(w, N, C, R, dgrad, out);
        return launched("pack_tc_kernel");
    }
    pack_simt_kernel<<<blocks, 256, 0, s>>>(w, N, C, R, dgrad, out);
    return launched("pack_simt_kernel");
}

extern "C" int odex_pack_weights_f16_pair(const float* w, int N, int C, int R, float* out_fwd, float* out_dgrad, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!w || !out_fwd || !out_dgrad || N <= 0 || C <= 0 || R <= 0) return fail(ODEX_E_ARG, "pack_weights_f16_pair: bad argument");
    if (N % 64 != 0 || C % 64 != 0) return fail(ODEX_E_SHAPE, "pack_weights_f16_pair: N and C must be multiples of 64 (got %d, %d)", N, C);
    if (!aligned16(out_fwd) || !aligned16(out_dgrad)) return fail(ODEX_E_ALIGN, "pack_weights_f16_pair: outputs not 16-byte aligned");
    const long total = static_cast<long>(N) * C * R * R;
    float* tail_f = out_fwd + total;
    float* tail_d = out_dgrad + total;
    cudaMemsetAsync(tail_f, 0, 4 * sizeof(float), s);
    cudaMemsetAsync(tail_d, 0, 4 * sizeof(float), s);
    pack_f16_stats2_kernel<<<N + C, 128, 0, s>>>(w, N, C, R, tail_f, tail_d);
    if (int rc = launched("pack_f16_stats2_kernel")) return rc;
    const int blocks = static_cast<int>(std::min<long>((2 * total + 255) / 256, 8L * kSMs));
    pack_f16_2_kernel<<<blocks, 256, 0, s>>>(w, N, C, R, reinterpret_cast<__half*>(out_fwd), tail_f, reinterpret_cast<__half*>(out_dgrad), tail_d);
    return launched("pack_f16_2_kernel");
}

namespace odex {
bool wgrad_tc_supported(int c0, int c1, int N, int H, int W, int R);
int launch_wgrad_tc(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy, int B,
                    int H, int W, int R, int N, float* dwp, float* db, cudaStream_t s);

static int wgrad_simt(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy, int ldy,
                      int B, int H, int W, int R, int N, float* dw, float* db, int packed, cudaStream_t s) {
    if (c0 % 64 != 0 || c1 % 64 != 0 || N % 4 != 0)
        return fail(ODEX_E_SHAPE, "wgrad: need c0, c1 %% 64 == 0 and N %% 4 == 0 (c0=%d c1=%d N=%d)", c0, c1, N);
    WgradArgs a;
    a.x0 = x0; a.ld0 = ld0; a.c0 = c0; a.x1 = x1; a.ld1 = ld1; a.c1 = c1;
    a.dy = dy; a.ldy = ldy; a.B = B; a.H = H; a.W = W; a.R = R; a.N = N; a.dw = dw; a.packed = packed;
    const int base = R * R * ((c0 + c1) / 64) * ((N + 63) / 64);
    const long P = static_cast<long>(B) * H * W;
    int nsplit = std::max(1, (4 * kSMs + base - 1) / base);
    nsplit = static_cast<int>(std::min<long>(nsplit, std::max<long>(1, P / 256)));
    a.nsplit = nsplit;
    conv_wgrad_kernel<<<base * nsplit, 256, 0, s>>>(a);
    int rc = launched("conv_wgrad_kernel");
    if (rc) return rc;
    if (db) {
        if (N > 1024 || 256 % (N / 4) != 0) return fail(ODEX_E_SHAPE, "wgrad: bias gradient needs N/4 to divide 256 (N=%d)", N);
        const int rlanes = 256 / (N / 4);
        const int blocks = static_cast<int>(std::min<long>((P + rlanes * 8 - 1) / (rlanes * 8), 4L * kSMs));
        colsum_kernel<<<blocks, 256, 0, s>>>(dy, ldy, P, N, db);
        rc = launched("colsum_kernel");
    }
    return rc;
}

static int wgrad_check(const float* x0, int ld0, const float* x1, int ld1, const float* dy, int ldy, const float* dw) {
    if (!x0 || !dy || !dw) return fail(ODEX_E_ARG, "wgrad: null pointer");
    if (!aligned16(x0) || !aligned16(dy) || (x1 && !aligned16(x1)) || ld0 % 4 || ldy % 4 || (x1 && ld1 % 4))
        return fail(ODEX_E_ALIGN, "wgrad: pointers / strides must be 16-byte aligned");
    return 0;
}
}  // namespace odex

extern "C" int odex_conv2d_wgrad(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1, const float* dy,
                                 int ldy, int B, int H, int W, int R, int N, float* dw, float* db, odex_stream_t s_) {
    if (int rc = wgrad_check(x0, ld0, x1, ld1, dy, ldy, dw)) return rc;
    if (!x1) c1 = 0;
    return wgrad_simt(x0, ld0, c0, x1, ld1, c1, dy, ldy, B, H, W, R, N, dw, db, 0, static_cast<cudaStream_t>(s_));
}

extern "C" int odex_conv2d_wgrad_packed(const float* x0, int ld0, int c0, const float* x1, int ld1, int c1,
                                        const float* dy, int ldy, int B, int H, int W, int R, int N, float* dwp,
                                        float* db, int impl, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (int rc = wgrad_check(x0, ld0, x1, ld1, dy, ldy, dwp)) return rc;
    if (!x1) c1 = 0;
    const bool tc_ok = wgrad_tc_supported(c0, c1, N, H, W, R);
    if (impl == ODEX_CONV_TC && !tc_ok)
        return fail(ODEX_E_UNSUPPORTED, "wgrad(tc): shape not supported (c0=%d c1=%d N=%d HxW=%dx%d R=%d)", c0, c1, N, H, W, R);
    if (impl != ODEX_CONV_SIMT && tc_ok) return launch_wgrad_tc(x0, ld0, c0, x1, ld1, c1, dy, ldy, B, H, W, R, N, dwp, db, s);
    return wgrad_simt(x0, ld0, c0, x1, ld1, c1, dy, ldy, B, H, W, R, N, dwp, db, 1, s);
}
