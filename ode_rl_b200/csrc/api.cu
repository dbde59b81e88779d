// C-ABI entry points that are not tied to one kernel family: version / error / accounting and the convolution
// dispatcher (tcgen05 path where the shape allows it, exact-fp32 FFMA path otherwise).
#include "common.cuh"

namespace odex {
thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

int launch_conv_simt(const odex_conv_desc* d, cudaStream_t s);
int launch_conv_tc(const odex_conv_desc* d, cudaStream_t s);
bool conv_tc_supported(const odex_conv_desc* d);
}  // namespace odex

using namespace odex;

extern "C" int odex_version(void) { return ODEX_VERSION; }
extern "C" const char* odex_last_error(void) { return g_err; }
extern "C" long long odex_launch_count(void) { return g_launches.load(); }

extern "C" int odex_conv2d(const odex_conv_desc* d, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!d || !d->x0 || !d->out) return fail(ODEX_E_ARG, "conv2d: null descriptor / x0 / out");
    if (d->B <= 0 || d->H <= 0 || d->W <= 0 || d->N <= 0 || d->c0 <= 0) return fail(ODEX_E_ARG, "conv2d: non-positive dimension");
    if (d->x1 && d->c1 <= 0) return fail(ODEX_E_ARG, "conv2d: x1 given with c1 <= 0");
    if (!aligned16(d->x0) || !aligned16(d->out) || (d->x1 && !aligned16(d->x1)) || (d->res && !aligned16(d->res)) ||
        (d->dact && !aligned16(d->dact)) || (d->bias && !aligned16(d->bias)))
        return fail(ODEX_E_ALIGN, "conv2d: pointers must be 16-byte aligned");
    if (d->ld0 % 4 || d->ldo % 4 || (d->x1 && d->ld1 % 4) || (d->res && d->ldr % 4) || (d->dact && d->ldd % 4))
        return fail(ODEX_E_ALIGN, "conv2d: row strides must be multiples of 4 floats");
    if (d->ld0 < d->c0 || (d->x1 && d->ld1 < d->c1) || d->ldo < d->N) return fail(ODEX_E_ARG, "conv2d: row stride smaller than channel count");
    switch (d->impl) {
        case ODEX_CONV_SIMT: return launch_conv_simt(d, s);
        case ODEX_CONV_TC: return launch_conv_tc(d, s);
        case ODEX_CONV_AUTO:
            if (conv_tc_supported(d)) return launch_conv_tc(d, s);
            return launch_conv_simt(d, s);
        default: return fail(ODEX_E_ARG, "conv2d: unknown impl %d", d->impl);
    }
}

// Measurement helper: enqueue the same convolution `reps` times from C (a Python loop over ctypes is slower than the
// kernel, so events around it would time the host).  bench.py brackets this with CUDA events on the same stream.
extern "C" int odex_conv2d_repeat(const odex_conv_desc* d, int reps, odex_stream_t s) {
    for (int i = 0; i < reps; ++i)
        if (int rc = odex_conv2d(d, s)) return rc;
    return 0;
}
