// Device helpers shared by the kind::f16 kernels (convnet_f16.cu, conv16_f16.cu): the two-term fp16 operand split
//   x = (hi + lo * 2^-11) / s,  hi = fp16(x * s),  lo = fp16((x * s - hi) * 2^11),  s a power of two,
// the scale choice, and the kind::f16 instruction descriptor / MMA issue.  See convnet_f16.cu for the accuracy argument.
#pragma once
#include <cuda_fp16.h>

#include "ptx_sm100.cuh"

namespace odex {

// power of two s with bound * s in [2^12, 2^13) (1 when bound is 0 or not finite); *inv = 1 / s
__device__ __forceinline__ float f16_scale_for(float bound, float* inv) {
    const uint32_t b = __float_as_uint(bound);
    const int e = static_cast<int>((b >> 23) & 0xFF);              // biased exponent: bound in [2^(e-127), 2^(e-126))
    if (e == 0 || e == 255) { *inv = 1.f; return 1.f; }
    int se = 127 + 12 - (e - 127);                                  // biased exponent of s
    se = se < 1 ? 1 : (se > 254 ? 254 : se);
    *inv = __uint_as_float(static_cast<uint32_t>(254 - se) << 23);
    return __uint_as_float(static_cast<uint32_t>(se) << 23);
}

// 8 consecutive channels (fp32, already multiplied by the scale) -> 16 B of hi and 16 B of lo halves
__device__ __forceinline__ void f16_split8(const float (&v)[8], uint4* hi, uint4* lo) {
    uint32_t h[4], l[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __half2 hh = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
        const float2 hf = __half22float2(hh);
        const __half2 ll = __floats2half2_rn((v[2 * i] - hf.x) * 2048.f, (v[2 * i + 1] - hf.y) * 2048.f);
        h[i] = *reinterpret_cast<const uint32_t*>(&hh);
        l[i] = *reinterpret_cast<const uint32_t*>(&ll);
    }
    *hi = make_uint4(h[0], h[1], h[2], h[3]);
    *lo = make_uint4(l[0], l[1], l[2], l[3]);
}

__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float4 lds128f(uint32_t addr) {
    float4 r;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "r"(addr));
    return r;
}

// kind::f16 (fp16 A and B, fp32 accumulate): a/b format 0 = F16, c format 1 = F32
__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N) {
    return (1u << 4) | (static_cast<uint32_t>(N >> 3) << 17) | (static_cast<uint32_t>(M >> 4) << 24);
}
__device__ __forceinline__ void umma_f16_ss_w(uint32_t d_tmem, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                              uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %2};\n\t"
        "mov.b64 db, {%3, %4};\n\t"
        "setp.ne.b32 p, %6, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
        ::"r"(d_tmem), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate)
        : "memory");
}

// geometry shared by the 16x16-image kernels: 64 channels, fp32 staging tile S = 2 halves of 256 px x 32 ch
constexpr int kF16C = 64;
constexpr int kF16P = 18;                      // halo pitch of the 3x3 chain
constexpr int kF16Plane = 41984;               // 18*18*128 = 41472 rounded up to 1 KB
constexpr int kF16SHalf = 256 * 128;           // 32 channels x 256 pixels, fp32

__device__ __forceinline__ uint32_t f16_slot(int y, int x) { return static_cast<uint32_t>((y + 1) * kF16P + x + 1); }

// One layer's epilogue for the two pixels (M-tiles) of this thread and the 32-channel halves [h0, h1) (both by default; the fused
// chain runs two warp groups, one per half).  ACT / KIND as in convnet_tc.cu; STORE: fp32 result into S; LAST: no operands for a
// next layer.  Returns max |v| of this thread's outputs.
template <int ACT, int KIND, bool STORE, bool LAST>
__device__ __forceinline__ float f16_epilogue(uint32_t tmem_base, uint32_t lane_addr, uint32_t base, uint32_t s_base, int m,
                                              const float4* bl, float alpha, float unscale, float next_scale, int h0 = 0, int h1 = 2) {
    constexpr int N = kF16C;
    float amax = 0.f;
#pragma unroll 1
    for (int mt = 0; mt < 2; ++mt) {
        const int y = m >> 3, x = mt * 8 + (m & 7);
        const uint32_t srow = f16_slot(y, x);
        const uint32_t p = static_cast<uint32_t>(y * 16 + x);
        const uint32_t plane_row = base + srow * 128u;
        const uint32_t psw = srow & 7u, ssw = p & 7u;
#pragma unroll 1
        for (int h = h0; h < h1; ++h) {
            uint32_t vm[32], vc[32];
            tmem_ld32(tmem_base + lane_addr + mt * 2 * N + h * 32, vm);
            tmem_ld32(tmem_base + lane_addr + mt * 2 * N + N + h * 32, vc);
            const uint32_t s_row = s_base + static_cast<uint32_t>(h) * kF16SHalf + p * 128u;
            float4 xa[8];
            if (KIND != 0) {
#pragma unroll
                for (int j = 0; j < 8; ++j) xa[j] = lds128f(s_row + ((static_cast<uint32_t>(j) ^ ssw) << 4));
            }
            tmem_ld_wait();
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) {
                float o[8];
#pragma unroll
                for (int q = 0; q < 2; ++q) {
                    const int j = 2 * jj + q;
                    const float4 bv = bl[h * 8 + j];
                    const float bb[4] = {bv.x, bv.y, bv.z, bv.w};
                    const float aa[4] = {KIND ? xa[j].x : 0.f, KIND ? xa[j].y : 0.f, KIND ? xa[j].z : 0.f, KIND ? xa[j].w : 0.f};
                    float v[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float acc = fmaf(__uint_as_float(vc[4 * j + e]), 4.8828125e-4f, __uint_as_float(vm[4 * j + e]));
                        v[e] = fmaf(acc, unscale, bb[e]);
                        if (ACT == ODEX_ACT_RELU) v[e] = relu_nan(v[e]);
                        if (ACT == ODEX_ACT_TANH) v[e] = tanhf(v[e]);
                        if (KIND == 1) v[e] = aa[e] > 0.f ? v[e] : 0.f;
                        if (KIND == 2) v[e] *= 1.f - aa[e] * aa[e];
                        if (KIND == 3) v[e] = fmaf(alpha, v[e], aa[e]);
                        o[4 * q + e] = v[e];
                    }
                    if (STORE)
                        asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(s_row + ((static_cast<uint32_t>(j) ^ ssw) << 4)),
                                     "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]) : "memory");
                }
                if (!LAST) {
#pragma unroll
                    for (int e = 0; e < 8; ++e) { amax = fmaxf(amax, fabsf(o[e])); o[e] *= next_scale; }
                    uint4 hi, lo;
                    f16_split8(o, &hi, &lo);
                    const uint32_t off = ((static_cast<uint32_t>(h * 4 + jj)) ^ psw) << 4;
                    sts128(plane_row + off, hi);
                    sts128(plane_row + kF16Plane + off, lo);
                }
            }
        }
    }
    return amax;
}

}  // namespace odex
