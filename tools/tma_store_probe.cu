// Probe: can a SWIZZLE_128B TMA store (and load) use a shared-memory address that is 128-byte but not 1024-byte aligned,
// i.e. is the swizzle a function of the ABSOLUTE shared address (as it is for UMMA descriptors, umma_shift_probe.cu)?
// Decides whether a kernel that keeps an image as an 18x18-slot halo plane can store its 16-pixel interior rows
// straight from the plane.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -I ode_rl_b200/csrc
//   tools/tma_store_probe.cu -o tools/build/tma_store_probe ; run on the B200.
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <vector>

#include "ptx_sm100.cuh"

using namespace odex;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

constexpr int kSlots = 64;   // plane of 64 pixel slots x 32 channels

// mode 0: generic stores write the plane with the absolute-address swizzle, TMA stores rows of 16 slots from slot s0
// mode 1: TMA loads 16 slots to slot s0, generic loads read the plane back assuming the absolute-address swizzle
__global__ void probe_kernel(const __grid_constant__ CUtensorMap tm, int s0, int mode, float* dump) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t bar = base + kSlots * 128;
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    __syncthreads();
    if (mode == 0) {
        for (int i = threadIdx.x; i < kSlots * 32; i += blockDim.x) {
            const int s = i / 32, c = i % 32;
            const uint32_t a = base + s * 128 + (((c >> 2) ^ (s & 7)) << 4) + (c & 3) * 4;
            const float v = s * 100.f + c;
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory");
        }
        fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) {
            tma_store_4d(&tm, base + s0 * 128, 0, 0, 0, 0);
            tma_store_commit();
            tma_store_wait_read();
        }
    } else {
        for (int i = threadIdx.x; i < kSlots * 32; i += blockDim.x)
            asm volatile("st.shared.f32 [%0], %1;" ::"r"(base + i * 4), "f"(-1.f) : "memory");
        fence_proxy_async_smem();
        __syncthreads();
        if (threadIdx.x == 0) {
            mbar_arrive_expect_tx(bar, 16 * 128);
            tma_load_4d(base + s0 * 128, &tm, bar, 0, 0, 0, 0);
        }
        mbar_wait(bar, 0u);
        for (int i = threadIdx.x; i < kSlots * 32; i += blockDim.x) {
            const int s = i / 32, c = i % 32;
            const uint32_t a = base + s * 128 + (((c >> 2) ^ (s & 7)) << 4) + (c & 3) * 4;
            float v;
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
            dump[i] = v;
        }
    }
}

int main() {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) { printf("no encode fn\n"); return 2; }
    EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(p);
    float *g, *dump;
    cudaMalloc(&g, 16 * 32 * 4);
    cudaMalloc(&dump, kSlots * 32 * 4);
    CUtensorMap tm;
    const cuuint64_t dims[4] = {32, 16, 1, 1};
    const cuuint64_t strides[3] = {128, 16 * 128, 16 * 128};
    const cuuint32_t box[4] = {32, 16, 1, 1};
    const cuuint32_t es[4] = {1, 1, 1, 1};
    if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, g, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
        printf("encode failed\n");
        return 2;
    }
    cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
    int bad_total = 0;
    const int shifts[] = {0, 1, 3, 8, 19, 21, 37};
    for (int s0 : shifts) {
        // ---- store
        cudaMemset(g, 0, 16 * 32 * 4);
        probe_kernel<<<1, 128, 32768>>>(tm, s0, 0, dump);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("store s0=%d: CUDA error %s\n", s0, cudaGetErrorString(e)); return 3; }
        std::vector<float> h(16 * 32);
        cudaMemcpy(h.data(), g, h.size() * 4, cudaMemcpyDeviceToHost);
        int bad = 0;
        for (int x = 0; x < 16; ++x)
            for (int c = 0; c < 32; ++c)
                if (h[x * 32 + c] != (s0 + x) * 100.f + c) ++bad;
        printf("TMA store from slot %2d (addr %% 1024 = %4d): %s (%d mismatches; first row %g %g %g %g)\n", s0, (s0 * 128) % 1024,
               bad ? "WRONG" : "OK", bad, h[0], h[1], h[4], h[8]);
        bad_total += bad;
        // ---- load
        std::vector<float> src(16 * 32);
        for (int x = 0; x < 16; ++x)
            for (int c = 0; c < 32; ++c) src[x * 32 + c] = 7000.f + x * 100.f + c;
        cudaMemcpy(g, src.data(), src.size() * 4, cudaMemcpyHostToDevice);
        probe_kernel<<<1, 128, 32768>>>(tm, s0, 1, dump);
        e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("load s0=%d: CUDA error %s\n", s0, cudaGetErrorString(e)); return 3; }
        std::vector<float> d(kSlots * 32);
        cudaMemcpy(d.data(), dump, d.size() * 4, cudaMemcpyDeviceToHost);
        bad = 0;
        for (int s = 0; s < kSlots; ++s)
            for (int c = 0; c < 32; ++c) {
                const float want = (s >= s0 && s < s0 + 16) ? 7000.f + (s - s0) * 100.f + c : -1.f;
                if (d[s * 32 + c] != want) ++bad;
            }
        printf("TMA load  to   slot %2d: %s (%d mismatches)\n", s0, bad ? "WRONG" : "OK", bad);
        bad_total += bad;
    }
    printf(bad_total ? "RESULT: swizzle is NOT a pure function of the absolute address for TMA\n"
                     : "RESULT: TMA SWIZZLE_128B follows the absolute shared address (any 128-B aligned start works)\n");
    return 0;
}
