// ConvGRU gate math fused with its GroupNorm (forward and backward).  One CTA per (sample, 32-channel group):
// the group's HW x 32 slab is read with 128-bit loads (each pixel's 32 channels are one 128-byte line), the
// statistics are reduced in the CTA, and the non-linearity / blend is applied in the same kernel.
// Replaces modules/ConvGRUCell.py:44,50 (GroupNorm) and :75-82 (split, sigmoid, r*h, tanh, blend) plus their
// autograd backward.  HBM-bound: algorithmic bytes per element are listed in DESIGN.md.
#include "common.cuh"

namespace odex {

// full-precision expf: these kernels are HBM-bound, the ~15 extra instructions are free
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }

// mean / rstd of the slab x[p*ld + c], p < HW, c in [c0, c0+32): two-pass (exact centred variance).
__device__ __forceinline__ void group_stats(const float* __restrict__ x, long ld, int HW, float eps, float* sh,
                                            float& mean, float& rstd) {
    const int units = HW * 8;
    float s = 0.f;
    for (int u = threadIdx.x; u < units; u += blockDim.x) {
        const float4 v = ld4(x + static_cast<long>(u >> 3) * ld + (u & 7) * 4);
        s += (v.x + v.y) + (v.z + v.w);
    }
    const float inv = 1.f / static_cast<float>(HW * 32);
    mean = block_sum(s, sh) * inv;
    float q = 0.f;
    for (int u = threadIdx.x; u < units; u += blockDim.x) {
        const float4 v = ld4(x + static_cast<long>(u >> 3) * ld + (u & 7) * 4);
        const float a = v.x - mean, b = v.y - mean, c = v.z - mean, d = v.w - mean;
        q += (a * a + b * b) + (c * c + d * d);
    }
    const float var = block_sum(q, sh) * inv;
    rstd = 1.f / sqrtf(var + eps);
}

__global__ void __launch_bounds__(256) gru_gates_fwd_kernel(const float* __restrict__ g, const float* __restrict__ h,
                                                            int ldh, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps, int HW, int C,
                                                            float* __restrict__ z, float* __restrict__ rh,
                                                            float* __restrict__ stats) {
    __shared__ float sh[33];
    const int G = 2 * C / 32;
    const int b = blockIdx.x / G, grp = blockIdx.x % G;
    const int c0 = grp * 32;
    const float* x = g + static_cast<long>(b) * HW * 2 * C + c0;
    float mean, rstd;
    group_stats(x, 2 * C, HW, eps, sh, mean, rstd);
    if (threadIdx.x == 0) { stats[blockIdx.x * 2] = mean; stats[blockIdx.x * 2 + 1] = rstd; }
    const int c4 = (threadIdx.x & 7) * 4;
    const float4 ga = ld4(gamma + c0 + c4), be = ld4(beta + c0 + c4);
    const bool is_z = c0 < C;
    for (int u = threadIdx.x; u < HW * 8; u += blockDim.x) {
        const long p = static_cast<long>(b) * HW + (u >> 3);
        const float4 v = ld4(x + static_cast<long>(u >> 3) * 2 * C + c4);
        float4 a;
        a.x = sigmoidf_((v.x - mean) * rstd * ga.x + be.x);
        a.y = sigmoidf_((v.y - mean) * rstd * ga.y + be.y);
        a.z = sigmoidf_((v.z - mean) * rstd * ga.z + be.z);
        a.w = sigmoidf_((v.w - mean) * rstd * ga.w + be.w);
        if (is_z) {
            st4(z + p * C + c0 + c4, a);
        } else {
            const float4 hv = ld4(h + p * ldh + (c0 - C) + c4);
            st4(rh + p * C + (c0 - C) + c4, make_float4(a.x * hv.x, a.y * hv.y, a.z * hv.z, a.w * hv.w));
        }
    }
}

__global__ void __launch_bounds__(256) gru_out_fwd_kernel(const float* __restrict__ cpre, const float* __restrict__ z,
                                                          const float* __restrict__ h, int ldh,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          float eps, int HW, int C, float* __restrict__ hnew,
                                                          float* __restrict__ cand, float* __restrict__ stats) {
    __shared__ float sh[33];
    const int G = C / 32;
    const int b = blockIdx.x / G, grp = blockIdx.x % G;
    const int c0 = grp * 32;
    const float* x = cpre + static_cast<long>(b) * HW * C + c0;
    float mean, rstd;
    group_stats(x, C, HW, eps, sh, mean, rstd);
    if (threadIdx.x == 0) { stats[blockIdx.x * 2] = mean; stats[blockIdx.x * 2 + 1] = rstd; }
    const int c4 = (threadIdx.x & 7) * 4;
    const float4 ga = ld4(gamma + c0 + c4), be = ld4(beta + c0 + c4);
    for (int u = threadIdx.x; u < HW * 8; u += blockDim.x) {
        const long p = static_cast<long>(b) * HW + (u >> 3);
        const float4 v = ld4(x + static_cast<long>(u >> 3) * C + c4);
        const float4 zz = ld4(z + p * C + c0 + c4);
        const float4 hv = ld4(h + p * ldh + c0 + c4);
        float4 c;
        c.x = tanhf((v.x - mean) * rstd * ga.x + be.x);
        c.y = tanhf((v.y - mean) * rstd * ga.y + be.y);
        c.z = tanhf((v.z - mean) * rstd * ga.z + be.z);
        c.w = tanhf((v.w - mean) * rstd * ga.w + be.w);
        if (cand) st4(cand + p * C + c0 + c4, c);
        float4 o;
        o.x = (1.f - zz.x) * hv.x + zz.x * c.x;
        o.y = (1.f - zz.y) * hv.y + zz.y * c.y;
        o.z = (1.f - zz.z) * hv.z + zz.z * c.z;
        o.w = (1.f - zz.w) * hv.w + zz.w * c.w;
        st4(hnew + p * C + c0 + c4, o);
    }
}

// shared tail of both backward kernels: per-channel dgamma/dbeta partials -> smem -> global atomics
__device__ __forceinline__ void flush_affine_grads(const float (&dgm)[4], const float (&dbt)[4], float* shg, float* shb,
                                                   float* dgamma, float* dbeta, int c0) {
    __syncthreads();
    if (threadIdx.x < 32) { shg[threadIdx.x] = 0.f; shb[threadIdx.x] = 0.f; }
    __syncthreads();
    const int c4 = (threadIdx.x & 7) * 4;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        // reduce over the 4 lanes of a warp that share c4 first (lanes l, l+8, l+16, l+24)
        float a = dgm[e], b = dbt[e];
        a += __shfl_xor_sync(0xffffffffu, a, 8); a += __shfl_xor_sync(0xffffffffu, a, 16);
        b += __shfl_xor_sync(0xffffffffu, b, 8); b += __shfl_xor_sync(0xffffffffu, b, 16);
        if ((threadIdx.x & 31) < 8) { atomicAdd(&shg[c4 + e], a); atomicAdd(&shb[c4 + e], b); }
    }
    __syncthreads();
    if (threadIdx.x < 32) {
        atomicAdd(dgamma + c0 + threadIdx.x, shg[threadIdx.x]);
        atomicAdd(dbeta + c0 + threadIdx.x, shb[threadIdx.x]);
    }
}

__global__ void __launch_bounds__(256) gru_out_bwd_kernel(const float* __restrict__ dhnew, const float* __restrict__ cpre,
                                                          const float* __restrict__ cand, const float* __restrict__ z,
                                                          const float* __restrict__ h, int ldh,
                                                          const float* __restrict__ gamma, const float* __restrict__ stats,
                                                          int HW, int C, float* __restrict__ dcpre, float* __restrict__ dz,
                                                          float* __restrict__ dh, float* dgamma, float* dbeta) {
    __shared__ float sh[33];
    __shared__ float shg[32], shb[32];
    const int G = C / 32;
    const int b = blockIdx.x / G, grp = blockIdx.x % G;
    const int c0 = grp * 32;
    const float mean = stats[blockIdx.x * 2], rstd = stats[blockIdx.x * 2 + 1];
    const int c4 = (threadIdx.x & 7) * 4;
    const float4 ga4 = ld4(gamma + c0 + c4);
    const float ga[4] = {ga4.x, ga4.y, ga4.z, ga4.w};
    float s1 = 0.f, s2 = 0.f;
    float dgm[4] = {0.f, 0.f, 0.f, 0.f}, dbt[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int u = threadIdx.x; u < HW * 8; u += blockDim.x) {
        const long o = (static_cast<long>(b) * HW + (u >> 3)) * C + c0 + c4;
        const float4 d4 = ld4(dhnew + o), x4 = ld4(cpre + o), cc4 = ld4(cand + o), z4 = ld4(z + o);
        const float d[4] = {d4.x, d4.y, d4.z, d4.w}, x[4] = {x4.x, x4.y, x4.z, x4.w};
        const float cc[4] = {cc4.x, cc4.y, cc4.z, cc4.w}, zz[4] = {z4.x, z4.y, z4.z, z4.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float dy = d[e] * zz[e] * (1.f - cc[e] * cc[e]);
            const float xh = (x[e] - mean) * rstd;
            dgm[e] += dy * xh; dbt[e] += dy;
            const float dxh = dy * ga[e];
            s1 += dxh; s2 += dxh * xh;
        }
    }
    const float invM = 1.f / static_cast<float>(HW * 32);
    const float m1 = block_sum(s1, sh) * invM;
    const float m2 = block_sum(s2, sh) * invM;
#pragma unroll 4
    for (int u = threadIdx.x; u < HW * 8; u += blockDim.x) {
        const long p = static_cast<long>(b) * HW + (u >> 3);
        const long o = p * C + c0 + c4;
        const float4 d4 = ld4(dhnew + o), x4 = ld4(cpre + o), cc4 = ld4(cand + o), z4 = ld4(z + o);
        const float4 h4 = ld4(h + p * ldh + c0 + c4);
        const float d[4] = {d4.x, d4.y, d4.z, d4.w}, x[4] = {x4.x, x4.y, x4.z, x4.w};
        const float cc[4] = {cc4.x, cc4.y, cc4.z, cc4.w}, zz[4] = {z4.x, z4.y, z4.z, z4.w};
        const float hh[4] = {h4.x, h4.y, h4.z, h4.w};
        float ox[4], oz[4], oh[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float dy = d[e] * zz[e] * (1.f - cc[e] * cc[e]);
            const float xh = (x[e] - mean) * rstd;
            ox[e] = rstd * (dy * ga[e] - m1 - xh * m2);
            oz[e] = d[e] * (cc[e] - hh[e]);
            oh[e] = d[e] * (1.f - zz[e]);
        }
        st4(dcpre + o, make_float4(ox[0], ox[1], ox[2], ox[3]));
        st4(dz + o, make_float4(oz[0], oz[1], oz[2], oz[3]));
        st4(dh + o, make_float4(oh[0], oh[1], oh[2], oh[3]));
    }
    flush_affine_grads(dgm, dbt, shg, shb, dgamma, dbeta, c0);
}

__global__ void __launch_bounds__(256) gru_gates_bwd_kernel(const float* __restrict__ dz, const float* drh,
                                                            int lddrh, const float* __restrict__ dh_in,
                                                            const float* __restrict__ g,
                                                            const float* __restrict__ h, int ldh,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            const float* __restrict__ stats, int HW, int C,
                                                            float* __restrict__ dg, float* dh_out, int lddh,
                                                            float* dgamma, float* dbeta) {
    __shared__ float sh[33];
    __shared__ float shg[32], shb[32];
    const int G = 2 * C / 32;
    const int b = blockIdx.x / G, grp = blockIdx.x % G;
    const int c0 = grp * 32;
    const bool is_z = c0 < C;
    const int cg0 = is_z ? c0 : c0 - C;   // channel base inside the C-wide gate tensors
    const float mean = stats[blockIdx.x * 2], rstd = stats[blockIdx.x * 2 + 1];
    const int c4 = (threadIdx.x & 7) * 4;
    const float4 ga4 = ld4(gamma + c0 + c4), be4 = ld4(beta + c0 + c4);
    const float ga[4] = {ga4.x, ga4.y, ga4.z, ga4.w}, be[4] = {be4.x, be4.y, be4.z, be4.w};
    float s1 = 0.f, s2 = 0.f;
    float dgm[4] = {0.f, 0.f, 0.f, 0.f}, dbt[4] = {0.f, 0.f, 0.f, 0.f};
    for (int pass = 0; pass < 2; ++pass) {
        float m1 = 0.f, m2 = 0.f;
        if (pass == 1) {
            const float invM = 1.f / static_cast<float>(HW * 32);
            m1 = block_sum(s1, sh) * invM;
            m2 = block_sum(s2, sh) * invM;
        }
#pragma unroll 4
        for (int u = threadIdx.x; u < HW * 8; u += blockDim.x) {
            const long p = static_cast<long>(b) * HW + (u >> 3);
            const float4 x4 = ld4(g + p * 2 * C + c0 + c4);
            const float x[4] = {x4.x, x4.y, x4.z, x4.w};
            float da[4], hh[4] = {0.f, 0.f, 0.f, 0.f}, dr[4] = {0.f, 0.f, 0.f, 0.f};
            if (is_z) {
                const float4 d4 = ld4(dz + p * C + cg0 + c4);
                da[0] = d4.x; da[1] = d4.y; da[2] = d4.z; da[3] = d4.w;
            } else {
                const float4 d4 = ld4(drh + p * lddrh + cg0 + c4);
                const float4 h4 = ld4(h + p * ldh + cg0 + c4);
                dr[0] = d4.x; dr[1] = d4.y; dr[2] = d4.z; dr[3] = d4.w;
                hh[0] = h4.x; hh[1] = h4.y; hh[2] = h4.z; hh[3] = h4.w;
#pragma unroll
                for (int e = 0; e < 4; ++e) da[e] = dr[e] * hh[e];
            }
            float ox[4], racc[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float xh = (x[e] - mean) * rstd;
                const float a = sigmoidf_(xh * ga[e] + be[e]);
                const float dy = da[e] * a * (1.f - a);
                if (pass == 0) {
                    dgm[e] += dy * xh; dbt[e] += dy;
                    const float dxh = dy * ga[e];
                    s1 += dxh; s2 += dxh * xh;
                } else {
                    ox[e] = rstd * (dy * ga[e] - m1 - xh * m2);
                    racc[e] = dr[e] * a;
                }
            }
            if (pass == 1) {
                st4(dg + p * 2 * C + c0 + c4, make_float4(ox[0], ox[1], ox[2], ox[3]));
                if (!is_z) {
                    // dh_out may alias drh (same element, read above by this thread)
                    const float4 cur = ld4(dh_in + p * C + cg0 + c4);
                    st4(dh_out + p * lddh + cg0 + c4,
                        make_float4(cur.x + racc[0], cur.y + racc[1], cur.z + racc[2], cur.w + racc[3]));
                }
            }
        }
    }
    flush_affine_grads(dgm, dbt, shg, shb, dgamma, dbeta, c0);
}

// ------------------------------------------------------------------------------------------------- layout
__global__ void transpose_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
    // per batch: in [rows][cols] -> out [cols][rows]
    __shared__ float tile[32][33];
    const long boff = static_cast<long>(blockIdx.z) * rows * cols;
    const int c = blockIdx.x * 32 + threadIdx.x;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int r = blockIdx.y * 32 + j;
        if (r < rows && c < cols) tile[j][threadIdx.x] = in[boff + static_cast<long>(r) * cols + c];
    }
    __syncthreads();
    const int r2 = blockIdx.y * 32 + threadIdx.x;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int c2 = blockIdx.x * 32 + j;
        if (r2 < rows && c2 < cols) out[boff + static_cast<long>(c2) * rows + r2] = tile[threadIdx.x][j];
    }
}

}  // namespace odex

using namespace odex;

static int check_gru(int B, int HW, int C) {
    if (B <= 0 || HW <= 0 || C <= 0 || C % 32 != 0) return fail(ODEX_E_SHAPE, "gru: C must be a positive multiple of 32 (B=%d HW=%d C=%d)", B, HW, C);
    return 0;
}

extern "C" int odex_gru_gates_fwd(const float* g, const float* h, int ldh, const float* gamma, const float* beta,
                                  float eps, int B, int HW, int C, float* z, float* rh, float* stats, odex_stream_t s) {
    if (!g || !h || !gamma || !beta || !z || !rh || !stats) return fail(ODEX_E_ARG, "gru_gates_fwd: null pointer");
    if (int rc = check_gru(B, HW, C)) return rc;
    gru_gates_fwd_kernel<<<B * (2 * C / 32), 256, 0, static_cast<cudaStream_t>(s)>>>(g, h, ldh, gamma, beta, eps, HW, C, z, rh, stats);
    return launched("gru_gates_fwd_kernel");
}

extern "C" int odex_gru_out_fwd(const float* cpre, const float* z, const float* h, int ldh, const float* gamma,
                                const float* beta, float eps, int B, int HW, int C, float* hnew, float* cand,
                                float* stats, odex_stream_t s) {
    if (!cpre || !z || !h || !gamma || !beta || !hnew || !stats) return fail(ODEX_E_ARG, "gru_out_fwd: null pointer");
    if (int rc = check_gru(B, HW, C)) return rc;
    gru_out_fwd_kernel<<<B * (C / 32), 256, 0, static_cast<cudaStream_t>(s)>>>(cpre, z, h, ldh, gamma, beta, eps, HW, C, hnew, cand, stats);
    return launched("gru_out_fwd_kernel");
}

extern "C" int odex_gru_out_bwd(const float* dhnew, const float* cpre, const float* cand, const float* z,
                                const float* h, int ldh, const float* gamma, const float* stats, int B, int HW, int C,
                                float* dcpre, float* dz, float* dh, float* dgamma, float* dbeta, odex_stream_t s) {
    if (!dhnew || !cpre || !cand || !z || !h || !gamma || !stats || !dcpre || !dz || !dh || !dgamma || !dbeta)
        return fail(ODEX_E_ARG, "gru_out_bwd: null pointer");
    if (int rc = check_gru(B, HW, C)) return rc;
    gru_out_bwd_kernel<<<B * (C / 32), 256, 0, static_cast<cudaStream_t>(s)>>>(dhnew, cpre, cand, z, h, ldh, gamma, stats, HW, C, dcpre, dz, dh, dgamma, dbeta);
    return launched("gru_out_bwd_kernel");
}

extern "C" int odex_gru_gates_bwd(const float* dz, const float* drh, int lddrh, const float* dh_in, const float* g,
                                  const float* h, int ldh, const float* gamma, const float* beta, const float* stats,
                                  int B, int HW, int C, float* dg, float* dh_out, int lddh, float* dgamma,
                                  float* dbeta, odex_stream_t s) {
    if (!dz || !drh || !dh_in || !g || !h || !gamma || !beta || !stats || !dg || !dh_out || !dgamma || !dbeta)
        return fail(ODEX_E_ARG, "gru_gates_bwd: null pointer");
    if (int rc = check_gru(B, HW, C)) return rc;
    gru_gates_bwd_kernel<<<B * (2 * C / 32), 256, 0, static_cast<cudaStream_t>(s)>>>(dz, drh, lddrh, dh_in, g, h, ldh, gamma, beta, stats, HW, C, dg, dh_out, lddh, dgamma, dbeta);
    return launched("gru_gates_bwd_kernel");
}

static int transpose(const float* in, float* out, int B, int rows, int cols, odex_stream_t s) {
    if (!in || !out || B <= 0 || rows <= 0 || cols <= 0) return fail(ODEX_E_ARG, "transpose: bad argument");
    dim3 grid((cols + 31) / 32, (rows + 31) / 32, B), block(32, 8);
    transpose_kernel<<<grid, block, 0, static_cast<cudaStream_t>(s)>>>(in, out, rows, cols);
    return launched("transpose_kernel");
}
extern "C" int odex_nchw_to_nhwc(const float* in, float* out, int B, int C, int HW, odex_stream_t s) {
    return transpose(in, out, B, C, HW, s);
}
extern "C" int odex_nhwc_to_nchw(const float* in, float* out, int B, int C, int HW, odex_stream_t s) {
    return transpose(in, out, B, HW, C, s);
}
