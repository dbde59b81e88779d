// Runge-Kutta stage arithmetic and the device-resident dopri5 step controller.
//
// All of these are HBM-bandwidth kernels (<= 0.5 FLOP/B): 128-bit loads, every operand plane read once, grids a
// multiple of the 148 SMs, warp-shuffle -> block -> last-block reductions (deterministic order), and the step
// decision (t, dt, accept/reject, which outputs to interpolate) stays in device memory -- no host sync per step.
// They restate torchdiffeq 0.2.1 (rk_common.py, misc.py, interp.py, dopri5.py; third-party, pinned at
// requirements.txt:102, called from modules/DiffEqSolver.py:45-46) -- see SURVEY.md Appendix A.
#include "common.cuh"

namespace odex {

// Dormand-Prince 5(4) coefficients used on the device
__constant__ double kCError[7] = {35.0 / 384.0 - 1951.0 / 21600.0, 0.0, 500.0 / 1113.0 - 22642.0 / 50085.0,
                                  125.0 / 192.0 - 451.0 / 720.0, -2187.0 / 6784.0 - -12231.0 / 42400.0,
                                  11.0 / 84.0 - 649.0 / 6300.0, -1.0 / 60.0};
__constant__ double kCMid[7] = {6025192743.0 / 30085553152.0 / 2.0, 0.0, 51252292925.0 / 65400821598.0 / 2.0,
                                -2691868925.0 / 45128329728.0 / 2.0, 187940372067.0 / 1594534317056.0 / 2.0,
                                -1776094331.0 / 19743644256.0 / 2.0, 11237099.0 / 235043384.0 / 2.0};

constexpr int kRedBlocks = 4 * kSMs;   // must match ODEX_RED_BLOCKS

struct LinArgs {
    const float* in[8];
    float coef[8];
    uint32_t dtmask;
    int nin;
};

template <int NIN>
__global__ void __launch_bounds__(256) lincomb_kernel(float* __restrict__ out, const LinArgs a, const double* dt_dev,
                                                      long n4, const int* skip) {
    pdl_wait();
    pdl_trigger();
    if (skip != nullptr && *skip != 0) return;
    float c[NIN];
    const float dt = dt_dev ? static_cast<float>(*dt_dev) : 1.f;
#pragma unroll
    for (int j = 0; j < NIN; ++j) c[j] = (a.dtmask >> j & 1u) ? a.coef[j] * dt : a.coef[j];
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        float4 v[NIN];
#pragma unroll
        for (int j = 0; j < NIN; ++j) v[j] = ld4(a.in[j] + 4 * i);
        float4 o = make_float4(v[0].x * c[0], v[0].y * c[0], v[0].z * c[0], v[0].w * c[0]);
#pragma unroll
        for (int j = 1; j < NIN; ++j) {
            o.x = fmaf(v[j].x, c[j], o.x); o.y = fmaf(v[j].y, c[j], o.y);
            o.z = fmaf(v[j].z, c[j], o.z); o.w = fmaf(v[j].w, c[j], o.w);
        }
        st4(out + 4 * i, o);
    }
}


// out = dy * act'(a): gradient through a trailing activation of a conv stack (a = saved activation output)
__global__ void __launch_bounds__(256) act_bwd_kernel(float* __restrict__ out, const float* __restrict__ dy,
                                                      const float* __restrict__ a, int kind, long n4) {
    pdl_wait();
    pdl_trigger();
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const float4 d = ld4(dy + 4 * i), v = ld4(a + 4 * i);
        float4 o;
        if (kind == ODEX_ACT_RELU) {
            o = make_float4(v.x > 0.f ? d.x : 0.f, v.y > 0.f ? d.y : 0.f, v.z > 0.f ? d.z : 0.f, v.w > 0.f ? d.w : 0.f);
        } else if (kind == ODEX_ACT_TANH) {
            o = make_float4(d.x * (1.f - v.x * v.x), d.y * (1.f - v.y * v.y), d.z * (1.f - v.z * v.z), d.w * (1.f - v.w * v.w));
        } else {
            o = d;
        }
        st4(out + 4 * i, o);
    }
}

// ------------------------------------------------------------------------------------------------- reductions
// Each block writes up to 3 double partials; the last block to finish sums them in block order (deterministic)
// into st->red[0..2] and resets the ticket.
struct DopriDev {          // device view of odex_dopri_state (same layout) + reduction scratch behind it
    odex_dopri_state s;
};

// Returns true in thread 0 of the LAST block, after st->red[] holds the full sums: the fused kernels run the 1-thread
// controller right there instead of in a launch of their own.
__device__ __forceinline__ bool finish_reduce(odex_dopri_state* st, double p0, double p1, double p2, double count,
                                              double* sh) {
    double* partial = reinterpret_cast<double*>(reinterpret_cast<char*>(st) + ODEX_STATE_HEAD_BYTES);
    unsigned int* ticket = reinterpret_cast<unsigned int*>(partial + 3 * kRedBlocks);
    p0 = block_sum(p0, sh);
    p1 = block_sum(p1, sh);
    p2 = block_sum(p2, sh);
    __shared__ bool last;
    if (threadIdx.x == 0) {
        partial[3 * blockIdx.x + 0] = p0;
        partial[3 * blockIdx.x + 1] = p1;
        partial[3 * blockIdx.x + 2] = p2;
        __threadfence();
        const unsigned int tk = atomicAdd(ticket, 1u);
        last = (tk == gridDim.x - 1);
    }
    __syncthreads();
    if (last) {
        __threadfence();
        double a0 = 0.0, a1 = 0.0, a2 = 0.0;
        // fixed assignment of partials to threads + fixed tree -> run-to-run deterministic
        for (int i = threadIdx.x; i < static_cast<int>(gridDim.x); i += blockDim.x) {
            a0 += partial[3 * i + 0]; a1 += partial[3 * i + 1]; a2 += partial[3 * i + 2];
        }
        a0 = block_sum(a0, sh); a1 = block_sum(a1, sh); a2 = block_sum(a2, sh);
        if (threadIdx.x == 0) {
            st->red[0] = a0; st->red[1] = a1; st->red[2] = a2; st->red[3] = count;
            *ticket = 0u;
            return true;
        }
    }
    return false;
}

// rk_common._adaptive_step's asserts for the NEXT attempt (`t + dt > t`, max_num_steps), evaluated as soon as its dt is known
__device__ __forceinline__ void begin_next_attempt(odex_dopri_state* st) {
    st->dt_try = st->dt;
    if (!(st->t1 + st->dt > st->t1)) { st->status |= 1; st->done = 1; }
    if (st->n_attempt >= st->max_attempts) { st->status |= 4; st->done = 1; }
}

// torchdiffeq misc._select_initial_step, computed in fp32 like the reference (y0.dtype), order = 4
__device__ __forceinline__ void initctrl_body(odex_dopri_state* st, int phase) {
    if (phase == 0) {
        const float d0 = static_cast<float>(sqrt(st->red[0] / st->red[3]));
        const float d1 = static_cast<float>(sqrt(st->red[1] / st->red[3]));
        st->d0 = d0; st->d1 = d1;
        const float h0 = (d0 < 1e-5f || d1 < 1e-5f) ? 1e-6f : 0.01f * d0 / d1;
        st->h0 = h0;
        st->n_feval += 2;    // f0 and the probe evaluation f(y0 + h0*f0)
    } else {
        const float h0 = static_cast<float>(st->h0), d1 = static_cast<float>(st->d1);
        const float d2 = static_cast<float>(sqrt(st->red[0] / st->red[3])) / h0;
        st->d2 = d2;
        float h1;
        if (d1 <= 1e-15f && d2 <= 1e-15f) h1 = fmaxf(1e-6f, h0 * 1e-3f);
        else h1 = powf(0.01f / fmaxf(d1, d2), 1.0f / 5.0f);
        st->dt = static_cast<double>(fminf(100.f * h0, h1));
        begin_next_attempt(st);
    }
}

// accept/reject + misc._optimal_step_size (safety 0.9, ifactor 10, dfactor 0.2, order 5), all fp64; one thread.
// `seq` is the host's running number of this attempt: an accepted attempt publishes it in st->adv_seq, which is what the
// dense-output / advance kernel of the SAME attempt checks (st->done may already be set by then: the last step).
__device__ __forceinline__ void ctrl_body(odex_dopri_state* st, const double* t, double* log, int log_cap, int seq) {
    const double dt = st->dt_try;
    if (st->red[1] > 0.0) { st->status |= 2; st->done = 1; return; }
    // the reference's ratio is an fp32 tensor (mean of fp32 squares, sqrt) cast to fp64 for the controller
    const double ratio = static_cast<double>(static_cast<float>(sqrt(st->red[0] / st->red[3])));
    st->ratio = ratio;
    const int accept = ratio <= 1.0 ? 1 : 0;
    st->accept = accept;
    st->n_feval += 6;
    const int a = st->n_attempt;
    if (log != nullptr && a < log_cap) {
        log[4 * a + 0] = st->t1; log[4 * a + 1] = dt; log[4 * a + 2] = ratio; log[4 * a + 3] = accept;
    }
    st->n_attempt = a + 1;
    bool finished = false;
    if (accept) {
        if (st->save_cap > 0 && st->n_accept >= st->save_cap) {
            // the backward checkpoints are full: stop HERE (sticky bit 3) instead of silently dropping this step's checkpoint
            st->status |= 8; st->done = 1;
            return;
        }
        st->t0 = st->t1;
        st->t1 = st->t1 + dt;
        st->n_accept += 1;
        st->dt_acc = dt;
        st->adv_seq = seq;
        int j = st->jnext;
        st->jlo = j;
        while (j < st->n_out && !(t[j] > st->t1)) ++j;
        st->jhi = j;
        st->jnext = j;
        finished = j >= st->n_out;
    } else {
        st->n_reject += 1;
        st->jlo = st->jhi = st->jnext;
    }
    double factor;
    if (ratio == 0.0) {
        factor = 10.0;
    } else {
        const double dfactor = ratio < 1.0 ? 1.0 : 0.2;
        factor = fmin(10.0, fmax(0.9 / pow(ratio, 0.2), dfactor));
    }
    st->dt = dt * factor;
    if (finished) { st->done = 1; return; }
    begin_next_attempt(st);
}

__global__ void __launch_bounds__(256) dopri_initnorm_kernel(odex_dopri_state* st, int phase, const float* __restrict__ y0,
                                                             const float* __restrict__ f0, const float* __restrict__ f1,
                                                             long n4, int fuse_ctrl) {
    __shared__ double sh[33];
    if (st->done) return;
    const float rtol = static_cast<float>(st->rtol), atol = static_cast<float>(st->atol);
    double a = 0.0, b = 0.0;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const float4 y = ld4(y0 + 4 * i), f = ld4(f0 + 4 * i);
        const float yy[4] = {y.x, y.y, y.z, y.w}, ff[4] = {f.x, f.y, f.z, f.w};
        float gg[4] = {0.f, 0.f, 0.f, 0.f};
        if (phase == 1) { const float4 g = ld4(f1 + 4 * i); gg[0] = g.x; gg[1] = g.y; gg[2] = g.z; gg[3] = g.w; }
        float sa = 0.f, sb = 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const float scale = atol + fabsf(yy[e]) * rtol;
            if (phase == 0) {
                const float u = yy[e] / scale, v = ff[e] / scale;
                sa += u * u; sb += v * v;
            } else {
                const float u = (gg[e] - ff[e]) / scale;
                sa += u * u;
            }
        }
        a += sa; b += sb;
    }
    if (finish_reduce(st, a, b, 0.0, static_cast<double>(n4) * 4.0, sh) && fuse_ctrl) initctrl_body(st, phase);
}

__global__ void dopri_init_kernel(odex_dopri_state* st, const double* t, int n_out, double rtol, double atol,
                                  int max_attempts, int save_cap) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    char* raw = reinterpret_cast<char*>(st);
    for (int i = 0; i < ODEX_STATE_HEAD_BYTES; ++i) raw[i] = 0;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(raw + ODEX_STATE_HEAD_BYTES + 3 * kRedBlocks * sizeof(double));
    *ticket = 0u;
    st->t0 = t[0]; st->t1 = t[0];
    st->rtol = rtol; st->atol = atol;
    st->n_out = n_out; st->jnext = 1;
    st->max_attempts = max_attempts;
    st->save_cap = save_cap;
    st->adv_seq = -1;
    st->done = n_out <= 1 ? 1 : 0;
}

__global__ void dopri_initctrl_kernel(odex_dopri_state* st, int phase) {
    if (threadIdx.x != 0 || blockIdx.x != 0 || st->done) return;
    initctrl_body(st, phase);
}

struct K7 { const float* k[7]; };
struct K7w { float* k[7]; };

__global__ void __launch_bounds__(256) dopri_errnorm_kernel(odex_dopri_state* st, const float* __restrict__ y0,
                                                            const float* __restrict__ y1, const K7 k, long n4,
                                                            const double* t, double* log, int log_cap, int seq, int fuse_ctrl) {
    __shared__ double sh[33];
    if (st->done) return;
    const float rtol = static_cast<float>(st->rtol), atol = static_cast<float>(st->atol);
    const float dt = static_cast<float>(st->dt_try);
    float ce[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) ce[j] = dt * static_cast<float>(kCError[j]);
    double acc = 0.0, bad = 0.0;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const float4 a = ld4(y0 + 4 * i), b = ld4(y1 + 4 * i);
        float4 e = ld4(k.k[0] + 4 * i);
        e.x *= ce[0]; e.y *= ce[0]; e.z *= ce[0]; e.w *= ce[0];
#pragma unroll
        for (int j = 2; j < 7; ++j) {
            const float4 v = ld4(k.k[j] + 4 * i);
            e.x = fmaf(v.x, ce[j], e.x); e.y = fmaf(v.y, ce[j], e.y);
            e.z = fmaf(v.z, ce[j], e.z); e.w = fmaf(v.w, ce[j], e.w);
        }
        const float t0 = atol + rtol * fmaxf(fabsf(a.x), fabsf(b.x));
        const float t1 = atol + rtol * fmaxf(fabsf(a.y), fabsf(b.y));
        const float t2 = atol + rtol * fmaxf(fabsf(a.z), fabsf(b.z));
        const float t3 = atol + rtol * fmaxf(fabsf(a.w), fabsf(b.w));
        const float r0 = e.x / t0, r1 = e.y / t1, r2 = e.z / t2, r3 = e.w / t3;
        acc += static_cast<double>((r0 * r0 + r1 * r1) + (r2 * r2 + r3 * r3));
        if (!(isfinite(a.x) && isfinite(a.y) && isfinite(a.z) && isfinite(a.w))) bad += 1.0;
    }
    if (finish_reduce(st, acc, bad, 0.0, static_cast<double>(n4) * 4.0, sh) && fuse_ctrl) ctrl_body(st, t, log, log_cap, seq);
}

__global__ void dopri_ctrl_kernel(odex_dopri_state* st, const double* t, double* log, int log_cap, int seq) {
    if (threadIdx.x != 0 || blockIdx.x != 0 || st->done) return;
    ctrl_body(st, t, log, log_cap, seq);
}

// dense output + state advance of the attempt numbered `seq`, iff the controller accepted it (st->adv_seq == seq; a
// speculative launch past the end of the integration, or after a rejection, finds another number there and returns).
__global__ void __launch_bounds__(256) dopri_advance_kernel(odex_dopri_state* st, const double* __restrict__ t,
                                                            float* __restrict__ y0, const float* __restrict__ y1,
                                                            const K7w k, float* __restrict__ sol, long n4,
                                                            float* __restrict__ ysave, float* __restrict__ fsave,
                                                            int save_cap, int seq) {
    if (st->adv_seq != seq) return;
    const int jlo = st->jlo, jhi = st->jhi;
    const float dt = static_cast<float>(st->dt_acc);
    const double T0 = st->t0, T1 = st->t1;
    const int step = st->n_accept - 1;
    float cm[7];
#pragma unroll
    for (int j = 0; j < 7; ++j) cm[j] = dt * static_cast<float>(kCMid[j]);
    const long n = n4 * 4;
    const bool save = ysave != nullptr && step < save_cap;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const float4 a = ld4(y0 + 4 * i), b = ld4(y1 + 4 * i);
        const float4 f0 = ld4(k.k[0] + 4 * i), f1 = ld4(k.k[6] + 4 * i);
        if (save) { st4(ysave + step * n + 4 * i, a); st4(fsave + step * n + 4 * i, f0); }
        if (jhi > jlo) {
            float4 m = make_float4(f0.x * cm[0], f0.y * cm[0], f0.z * cm[0], f0.w * cm[0]);
#pragma unroll
            for (int j = 2; j < 6; ++j) {
                const float4 v = ld4(k.k[j] + 4 * i);
                m.x = fmaf(v.x, cm[j], m.x); m.y = fmaf(v.y, cm[j], m.y);
                m.z = fmaf(v.z, cm[j], m.z); m.w = fmaf(v.w, cm[j], m.w);
            }
            m.x = fmaf(f1.x, cm[6], m.x); m.y = fmaf(f1.y, cm[6], m.y);
            m.z = fmaf(f1.z, cm[6], m.z); m.w = fmaf(f1.w, cm[6], m.w);
            const float ya[4] = {a.x, a.y, a.z, a.w}, yb[4] = {b.x, b.y, b.z, b.w};
            const float fa[4] = {f0.x, f0.y, f0.z, f0.w}, fb[4] = {f1.x, f1.y, f1.z, f1.w};
            const float mm[4] = {m.x, m.y, m.z, m.w};
            float ca[4], cb[4], cc[4], cd[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const float ymid = ya[e] + mm[e];
                ca[e] = 2.f * dt * (fb[e] - fa[e]) - 8.f * (yb[e] + ya[e]) + 16.f * ymid;
                cb[e] = dt * (5.f * fa[e] - 3.f * fb[e]) + 18.f * ya[e] + 14.f * yb[e] - 32.f * ymid;
                cc[e] = dt * (fb[e] - 4.f * fa[e]) - 11.f * ya[e] - 5.f * yb[e] + 16.f * ymid;
                cd[e] = dt * fa[e];
            }
            for (int j = jlo; j < jhi; ++j) {
                const float x = static_cast<float>((t[j] - T0) / (T1 - T0));
                const float x2 = x * x, x3 = x2 * x, x4 = x3 * x;
                float o[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) o[e] = (((ya[e] + x * cd[e]) + x2 * cc[e]) + x3 * cb[e]) + x4 * ca[e];
                st4(sol + j * n + 4 * i, make_float4(o[0], o[1], o[2], o[3]));
            }
        }
        st4(y0 + 4 * i, b);
        st4(k.k[0] + 4 * i, f1);
    }
}

inline int ew_blocks(long n4) { return static_cast<int>(std::min<long>((n4 + 255) / 256, static_cast<long>(kRedBlocks))); }

}  // namespace odex

using namespace odex;

static_assert(sizeof(odex_dopri_state) <= ODEX_STATE_HEAD_BYTES, "state header too large");
static_assert(ODEX_RED_BLOCKS == kRedBlocks, "ODEX_RED_BLOCKS must equal 4 * 148");

extern "C" int odex_lincomb(float* out, int nin, const float* const* in, const double* coef, uint32_t dtmask,
                            const double* dt_dev, long long n, const int* skip, odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!out || !in || !coef || nin < 1 || nin > 8 || n <= 0) return fail(ODEX_E_ARG, "lincomb: bad argument (nin=%d n=%lld)", nin, n);
    if (n % 4 != 0) return fail(ODEX_E_SHAPE, "lincomb: n must be a multiple of 4 (got %lld)", n);
    if (dtmask != 0 && !dt_dev) return fail(ODEX_E_ARG, "lincomb: dtmask set but dt_dev is NULL");
    LinArgs a;
    a.nin = nin; a.dtmask = dtmask;
    for (int j = 0; j < 8; ++j) { a.in[j] = j < nin ? in[j] : nullptr; a.coef[j] = j < nin ? static_cast<float>(coef[j]) : 0.f; }
    if (!aligned16(out)) return fail(ODEX_E_ALIGN, "lincomb: out not 16-byte aligned");
    for (int j = 0; j < nin; ++j)
        if (!a.in[j] || !aligned16(a.in[j])) return fail(ODEX_E_ALIGN, "lincomb: in[%d] null or not 16-byte aligned", j);
    const long n4 = n / 4;
    const int blocks = ew_blocks(n4);
    switch (nin) {
        case 1: launch_pdl(lincomb_kernel<1>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        case 2: launch_pdl(lincomb_kernel<2>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        case 3: launch_pdl(lincomb_kernel<3>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        case 4: launch_pdl(lincomb_kernel<4>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        case 5: launch_pdl(lincomb_kernel<5>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        case 6: launch_pdl(lincomb_kernel<6>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        case 7: launch_pdl(lincomb_kernel<7>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
        default: launch_pdl(lincomb_kernel<8>, blocks, 256, 0, s, out, a, dt_dev, n4, skip); break;
    }
    return launched("lincomb_kernel");
}


extern "C" int odex_act_bwd(float* out, const float* dy, const float* a, int kind, long long n, odex_stream_t s) {
    if (!out || !dy || !a || n <= 0 || n % 4) return fail(ODEX_E_ARG, "act_bwd: bad argument");
    if (!aligned16(out) || !aligned16(dy) || !aligned16(a)) return fail(ODEX_E_ALIGN, "act_bwd: pointers must be 16-byte aligned");
    launch_pdl(act_bwd_kernel, ew_blocks(n / 4), 256, 0, static_cast<cudaStream_t>(s), out, dy, a, kind, n / 4);
    return launched("act_bwd_kernel");
}

extern "C" size_t odex_dopri_state_bytes(void) {
    return ODEX_STATE_HEAD_BYTES + 3 * ODEX_RED_BLOCKS * sizeof(double) + 64;
}

extern "C" int odex_dopri_init(odex_dopri_state* st, const double* t, int n_out, double rtol, double atol,
                               int max_attempts, int save_cap, odex_stream_t s) {
    if (!st || !t || n_out < 1 || save_cap < 0) return fail(ODEX_E_ARG, "dopri_init: bad argument");
    dopri_init_kernel<<<1, 32, 0, static_cast<cudaStream_t>(s)>>>(st, t, n_out, rtol, atol, max_attempts, save_cap);
    return launched("dopri_init_kernel");
}

extern "C" int odex_dopri_initnorm(odex_dopri_state* st, int phase, const float* y0, const float* f0, const float* f1,
                                   long long n, int fuse_ctrl, odex_stream_t s) {
    if (!st || !y0 || !f0 || (phase == 1 && !f1) || n <= 0 || n % 4) return fail(ODEX_E_ARG, "dopri_initnorm: bad argument");
    dopri_initnorm_kernel<<<ew_blocks(n / 4), 256, 0, static_cast<cudaStream_t>(s)>>>(st, phase, y0, f0, f1, n / 4, fuse_ctrl);
    return launched("dopri_initnorm_kernel");
}

extern "C" int odex_dopri_initctrl(odex_dopri_state* st, int phase, odex_stream_t s) {
    if (!st) return fail(ODEX_E_ARG, "dopri_initctrl: null state");
    dopri_initctrl_kernel<<<1, 32, 0, static_cast<cudaStream_t>(s)>>>(st, phase);
    return launched("dopri_initctrl_kernel");
}

extern "C" int odex_dopri_errnorm(odex_dopri_state* st, const float* y0, const float* y1, const float* const* k7,
                                  long long n, odex_stream_t s) {
    if (!st || !y0 || !y1 || !k7 || n <= 0 || n % 4) return fail(ODEX_E_ARG, "dopri_errnorm: bad argument");
    K7 k;
    for (int j = 0; j < 7; ++j) k.k[j] = k7[j];
    dopri_errnorm_kernel<<<ew_blocks(n / 4), 256, 0, static_cast<cudaStream_t>(s)>>>(st, y0, y1, k, n / 4, nullptr, nullptr, 0, 0, 0);
    return launched("dopri_errnorm_kernel");
}

extern "C" int odex_dopri_errctrl(odex_dopri_state* st, const double* t, const float* y0, const float* y1,
                                  const float* const* k7, long long n, double* log, int log_cap, int seq, odex_stream_t s) {
    if (!st || !t || !y0 || !y1 || !k7 || n <= 0 || n % 4 || seq < 0) return fail(ODEX_E_ARG, "dopri_errctrl: bad argument");
    K7 k;
    for (int j = 0; j < 7; ++j) k.k[j] = k7[j];
    dopri_errnorm_kernel<<<ew_blocks(n / 4), 256, 0, static_cast<cudaStream_t>(s)>>>(st, y0, y1, k, n / 4, t, log, log_cap, seq, 1);
    return launched("dopri_errnorm_kernel");
}

extern "C" int odex_dopri_ctrl(odex_dopri_state* st, const double* t, double* log, int log_cap, int seq, odex_stream_t s) {
    if (!st || !t || seq < 0) return fail(ODEX_E_ARG, "dopri_ctrl: bad argument");
    dopri_ctrl_kernel<<<1, 32, 0, static_cast<cudaStream_t>(s)>>>(st, t, log, log_cap, seq);
    return launched("dopri_ctrl_kernel");
}

extern "C" int odex_dopri_advance(odex_dopri_state* st, const double* t, float* y0, const float* y1, float* const* k7,
                                  float* sol, long long n, float* ysave, float* fsave, int save_cap, int seq,
                                  odex_stream_t s_) {
    cudaStream_t s = static_cast<cudaStream_t>(s_);
    if (!st || !t || !y0 || !y1 || !k7 || !sol || n <= 0 || n % 4 || seq < 0) return fail(ODEX_E_ARG, "dopri_advance: bad argument");
    if ((ysave == nullptr) != (fsave == nullptr)) return fail(ODEX_E_ARG, "dopri_advance: ysave and fsave go together");
    K7w k;
    for (int j = 0; j < 7; ++j) k.k[j] = k7[j];
    dopri_advance_kernel<<<ew_blocks(n / 4), 256, 0, s>>>(st, t, y0, y1, k, sol, n / 4, ysave, fsave, save_cap, seq);
    return launched("dopri_advance_kernel");
}

// *flag |= bit when x holds a NaN: the encoder recurrence's `assert not torch.isnan(...)` (modules/ODEConvGRUCell.py:56,59)
// as a sticky device word the host reads once per forward.  A NaN in any iteration reaches the last hidden state (every
// later update reads it through the convolutions), so one pass over the final state replaces the reference's 2 syncs per frame.
namespace odex {
__global__ void __launch_bounds__(256) nan_flag_kernel(const float* __restrict__ x, long n4, int* flag, int bit) {
    bool bad = false;
    for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n4;
         i += static_cast<long>(gridDim.x) * blockDim.x) {
        const float4 v = ld4(x + 4 * i);
        bad = bad || isnan(v.x) || isnan(v.y) || isnan(v.z) || isnan(v.w);
    }
    if (__syncthreads_or(bad ? 1 : 0) && threadIdx.x == 0) atomicOr(flag, bit);
}
}  // namespace odex

extern "C" int odex_nan_flag(const float* x, long long n, int* flag, int bit, odex_stream_t s) {
    if (!x || !flag || n <= 0 || n % 4) return fail(ODEX_E_ARG, "nan_flag: bad argument");
    if (!aligned16(x)) return fail(ODEX_E_ALIGN, "nan_flag: x not 16-byte aligned");
    nan_flag_kernel<<<ew_blocks(n / 4), 256, 0, static_cast<cudaStream_t>(s)>>>(x, n / 4, flag, bit);
    return launched("nan_flag_kernel");
}
