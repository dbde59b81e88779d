// Probe: does a tcgen05 K-major SWIZZLE_128B shared-memory descriptor accept a start address that is a multiple of
// 128 B but not of 1024 B (a row-shifted window into a TMA-written tile), and a stride between 8-row groups (SBO)
// that is not a multiple of 1024 B?  This decides whether the convolution can take its tap-shifted A operand
// straight from the halo tile.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -I ode_rl_b200/csrc
//   tools/umma_shift_probe.cu -o gpurun_out/umma_shift_probe ; run on the B200.
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <vector>

#include "ptx_sm100.cuh"

using namespace odex;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int kRows = 256;   // rows of the A source tile in smem (32 floats each)
constexpr int kN = 64;

__device__ __forceinline__ uint64_t desc_bo(uint32_t saddr, uint32_t sbo_bytes, uint32_t base_off) {
    uint64_t d = umma_desc_sw128(saddr, 16, sbo_bytes);
    d |= static_cast<uint64_t>(base_off & 7u) << 49;
    return d;
}

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                             int shift, int pitch, int bo_mode, float* out) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t a_sm = base;                       // kRows * 128 B
    const uint32_t b_sm = base + kRows * 128;         // kN * 128 B
    const uint32_t bar_ld = b_sm + kN * 128;
    const uint32_t bar_mma = bar_ld + 8;
    const uint32_t slot = bar_ld + 16;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 1);
        mbar_init(bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) { tmem_alloc(slot, 64); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar_ld, (kRows + kN) * 128);
        tma_load_2d(a_sm, &tmA, bar_ld, 0, 0);
        tma_load_2d(b_sm, &tmB, bar_ld, 0, 0);
        mbar_wait(bar_ld, 0);
        tc_fence_after_sync();
        const uint32_t start = a_sm + shift * 128;
        const uint32_t bo = bo_mode == 0 ? 0u : ((start >> 7) & 7u);
        constexpr uint32_t idesc = umma_idesc_tf32(128, kN);
        for (int k8 = 0; k8 < 4; ++k8) {
            const uint64_t ad = desc_bo(start + k8 * 32, pitch * 128, bo);
            const uint64_t bd = umma_desc_sw128(b_sm + k8 * 32, 16, 1024);
            umma_tf32_ss(tmem, ad, bd, idesc, k8 != 0);
        }
        umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, 0);
    tc_fence_after_sync();
    uint32_t v[32];
    for (int c0 = 0; c0 < kN; c0 += 32) {
        tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[threadIdx.x * kN + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem, 64); }
}

// ---- MN-major variant (the weight-gradient GEMM: K = pixel rows, M/N = channels contiguous in a 128-byte row)
// A: 4 planes (32 M-values each) x kRowsT rows x 128 B; B: 2 planes.  D[m][n] = sum_k A[k + shift][m] * B[k][n].
constexpr int kRowsT = 64;
__global__ void __launch_bounds__(128) probe_mn(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                int shift, int bo_mode, int layout_type, int sbo, float* out) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t plane = kRowsT * 128;
    const uint32_t a_sm = base;                       // 4 planes
    const uint32_t b_sm = base + 4 * plane;           // 2 planes
    const uint32_t bar_ld = b_sm + 2 * plane;
    const uint32_t bar_mma = bar_ld + 8;
    const uint32_t slot = bar_ld + 16;
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) { mbar_init(bar_ld, 1); mbar_init(bar_mma, 1); fence_barrier_init(); }
    if (warp == 0) { tmem_alloc(slot, 64); tmem_relinquish(); }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(slot));
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar_ld, 6 * plane);
        for (int pl = 0; pl < 4; ++pl) tma_load_2d(a_sm + pl * plane, &tmA, bar_ld, 0, pl * kRowsT);
        for (int pl = 0; pl < 2; ++pl) tma_load_2d(b_sm + pl * plane, &tmB, bar_ld, 0, pl * kRowsT);
        mbar_wait(bar_ld, 0);
        tc_fence_after_sync();
        constexpr uint32_t idesc = umma_idesc_tf32(128, kN, 1, 1);
        for (int k8 = 0; k8 < 4; ++k8) {
            const uint32_t sa = a_sm + (shift + 8 * k8) * 128;
            const uint32_t sb = b_sm + (8 * k8) * 128;
            const uint32_t bo = bo_mode == 0 ? 0u : ((sa >> 7) & 7u);
            uint64_t ad = umma_desc_sw128(sa, plane, sbo);
            ad |= static_cast<uint64_t>(bo) << 49;
            uint64_t bd = umma_desc_sw128(sb, plane, sbo);
            // layout type lives in bits [61,64): 2 = SWIZZLE_128B (16-byte atoms), 1 = SWIZZLE_128B_BASE32B
            ad = (ad & ~(7ull << 61)) | (static_cast<uint64_t>(layout_type) << 61);
            bd = (bd & ~(7ull << 61)) | (static_cast<uint64_t>(layout_type) << 61);
            umma_tf32_ss(tmem, ad, bd, idesc, k8 != 0);
        }
        umma_commit(bar_mma);
    }
    mbar_wait(bar_mma, 0);
    tc_fence_after_sync();
    uint32_t v[32];
    for (int c0 = 0; c0 < kN; c0 += 32) {
        tmem_ld32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + c0, v);
        tmem_ld_wait();
        for (int j = 0; j < 32; ++j) out[threadIdx.x * kN + c0 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem, 64); }
}

int main() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) { printf("no encode fn\n"); return 2; }
    EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(p);
    std::vector<float> hA(kRows * 32), hB(kN * 32);
    for (int r = 0; r < kRows; ++r) for (int k = 0; k < 32; ++k) hA[r * 32 + k] = static_cast<float>((r * 7 + k * 3) % 13 - 6);
    for (int n = 0; n < kN; ++n) for (int k = 0; k < 32; ++k) hB[n * 32 + k] = static_cast<float>((n * 5 + k) % 7 - 3);
    float *dA, *dB, *dO;
    cudaMalloc(&dA, hA.size() * 4); cudaMalloc(&dB, hB.size() * 4); cudaMalloc(&dO, 128 * kN * 4);
    cudaMemcpy(dA, hA.data(), hA.size() * 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dB, hB.data(), hB.size() * 4, cudaMemcpyHostToDevice);
    CUtensorMap tmA, tmB;
    auto mk = [&](CUtensorMap* tm, float* ptr, int rows) {
        const cuuint64_t dims[2] = {32, static_cast<cuuint64_t>(rows)};
        const cuuint64_t strides[1] = {128};
        const cuuint32_t box[2] = {32, static_cast<cuuint32_t>(rows)};
        const cuuint32_t es[2] = {1, 1};
        return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    };
    if (mk(&tmA, dA, kRows) != CUDA_SUCCESS || mk(&tmB, dB, kN) != CUDA_SUCCESS) { printf("encode failed\n"); return 2; }
    const int smem = 1024 + (kRows + kN) * 128 + 64;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    std::vector<float> hO(128 * kN);
    int bad_total = 0;
    const int pitches[] = {8, 10, 12, 16};
    for (int pitch : pitches)
        for (int shift = 0; shift < 12; ++shift)
            for (int bo = 0; bo < 2; ++bo) {
                if (shift + 15 * pitch + 8 > kRows) continue;
                cudaMemset(dO, 0, 128 * kN * 4);
                probe<<<1, 128, smem>>>(tmA, tmB, shift, pitch, bo, dO);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("pitch %d shift %d bo %d: CUDA error %s\n", pitch, shift, bo, cudaGetErrorString(e)); return 3; }
                cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost);
                int bad = 0;
                for (int m = 0; m < 128; ++m) {
                    const int row = shift + (m / 8) * pitch + (m % 8);
                    for (int n = 0; n < kN; ++n) {
                        float ref = 0.f;
                        for (int k = 0; k < 32; ++k) ref += hA[row * 32 + k] * hB[n * 32 + k];
                        if (ref != hO[m * kN + n]) ++bad;
                    }
                }
                printf("pitch %2d shift %2d base_offset_mode %d : %s (%d mismatches)\n", pitch, shift, bo, bad ? "WRONG" : "ok", bad);
                if (bad && bo == 0) ++bad_total;
            }
    printf("summary: %d failing (pitch, shift) combinations with base_offset = 0\n", bad_total);
    // ---------------- MN-major
    {
        std::vector<float> tA(4 * kRowsT * 32), tB(2 * kRowsT * 32);
        for (int pl = 0; pl < 4; ++pl) for (int r = 0; r < kRowsT; ++r) for (int j = 0; j < 32; ++j)
            tA[(pl * kRowsT + r) * 32 + j] = static_cast<float>((r * 5 + (pl * 32 + j) * 3) % 11 - 5);
        for (int pl = 0; pl < 2; ++pl) for (int r = 0; r < kRowsT; ++r) for (int j = 0; j < 32; ++j)
            tB[(pl * kRowsT + r) * 32 + j] = static_cast<float>((r * 3 + (pl * 32 + j)) % 7 - 3);
        float *dtA, *dtB;
        cudaMalloc(&dtA, tA.size() * 4); cudaMalloc(&dtB, tB.size() * 4);
        cudaMemcpy(dtA, tA.data(), tA.size() * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(dtB, tB.data(), tB.size() * 4, cudaMemcpyHostToDevice);
        for (int lt = 1; lt <= 2; ++lt)
        for (int sbo = 512; sbo <= 1024; sbo *= 2) {
        CUtensorMap mA, mB;
        auto mk2 = [&](CUtensorMap* tm, float* ptr, int rows) {
            const cuuint64_t dims[2] = {32, static_cast<cuuint64_t>(rows)};
            const cuuint64_t strides[1] = {128};
            const cuuint32_t box[2] = {32, static_cast<cuuint32_t>(kRowsT)};
            const cuuint32_t es[2] = {1, 1};
            return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                       lt == 1 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        };
        if (mk2(&mA, dtA, 4 * kRowsT) != CUDA_SUCCESS || mk2(&mB, dtB, 2 * kRowsT) != CUDA_SUCCESS) { printf("encode failed\n"); return 2; }
        const int smem2 = 1024 + 6 * kRowsT * 128 + 64;
        cudaFuncSetAttribute(probe_mn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2);
        for (int shift = 0; shift < 12; ++shift)
            for (int bo = 0; bo < 2; ++bo) {
                cudaMemset(dO, 0, 128 * kN * 4);
                probe_mn<<<1, 128, smem2>>>(mA, mB, shift, bo, lt, sbo, dO);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("MN shift %d bo %d: CUDA error %s\n", shift, bo, cudaGetErrorString(e)); return 3; }
                cudaMemcpy(hO.data(), dO, hO.size() * 4, cudaMemcpyDeviceToHost);
                int bad = 0;
                for (int m = 0; m < 128; ++m)
                    for (int n = 0; n < kN; ++n) {
                        float ref = 0.f;
                        for (int k = 0; k < 32; ++k)
                            ref += tA[((m / 32) * kRowsT + k + shift) * 32 + (m % 32)] * tB[((n / 32) * kRowsT + k) * 32 + (n % 32)];
                        if (ref != hO[m * kN + n]) ++bad;
                    }
                printf("MN-major layout_type %d sbo %4d shift %2d base_offset_mode %d : %s (%d mismatches)\n", lt, sbo, shift, bo, bad ? "WRONG" : "ok", bad);
            }
        }
    }
    return 0;
}
