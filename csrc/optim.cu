// Multi-tensor fused optimizer, AMP and gradient-norm kernels (HBM-bound; W10, K13, K8).
//
// One CTA handles one B200DP_CHUNK-element chunk of one tensor.  Every element is read
// once and written once; 16-byte vector accesses are used whenever the chunk's pointers
// are 16-byte aligned (they are for flat bucket layouts), scalar accesses otherwise.
#include <stdarg.h>
#include <string.h>

#include "common.cuh"

namespace b200dp {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

constexpr int kThreads = 256;

template <typename GT, typename LT>
__global__ void __launch_bounds__(kThreads) multi_adam_kernel(
    const b200dp_tensor_desc* __restrict__ descs, const b200dp_chunk* __restrict__ chunks,
    b200dp_adam_hparams hp, const float* __restrict__ step_p, const float* __restrict__ grad_mult_p,
    const float* __restrict__ found_inf_p) {
    if (found_inf_p != nullptr && *found_inf_p != 0.f) return;  // device-side "skip step"
    const b200dp_chunk ck = chunks[blockIdx.x];
    const b200dp_tensor_desc d = descs[ck.x];
    const int64_t start = (int64_t)ck.y * B200DP_CHUNK;
    const int64_t len = min((int64_t)B200DP_CHUNK, d.n - start);
    float* p = (float*)d.p + start;
    const GT* g = (const GT*)d.g + start;
    float* m = (float*)d.m + start;
    float* v = (float*)d.v + start;
    LT* lp = d.p_lp ? (LT*)d.p_lp + start : nullptr;

    const float step = *step_p;
    const float gm = grad_mult_p ? *grad_mult_p : 1.f;
    const float lr = hp.lr_dev ? *hp.lr_dev : hp.lr;
    const float bc1 = 1.f - powf(hp.beta1, step);
    const float bc2 = 1.f - powf(hp.beta2, step);
    const float step_size = lr / bc1;
    const float inv_sqrt_bc2 = rsqrtf(bc2);
    const float decay = hp.adamw ? 1.f - lr * hp.weight_decay : 1.f;
    const float l2 = hp.adamw ? 0.f : hp.weight_decay;
    const float sign = hp.maximize ? -1.f : 1.f;
    const float b1 = hp.beta1, b2 = hp.beta2, eps = hp.eps;

    auto upd = [&](float pv, float gv, float& mv, float& vv) -> float {
        gv = gv * gm * sign + l2 * pv;
        pv *= decay;
        mv = b1 * mv + (1.f - b1) * gv;
        vv = b2 * vv + (1.f - b2) * gv * gv;
        const float denom = sqrtf(vv) * inv_sqrt_bc2 + eps;
        return pv - step_size * (mv / denom);
    };

    const bool vec = is_aligned16(p) && is_aligned16(g) && is_aligned16(m) && is_aligned16(v) &&
                     (lp == nullptr || is_aligned16(lp)) && (sizeof(GT) == 4 || is_aligned16(g));
    int64_t done = 0;
    if (vec) {
        const int64_t nvec = len / 8;
        for (int64_t i = threadIdx.x; i < nvec; i += kThreads) {
            float pv[8], gv[8], mv[8], vv[8];
            load8<float>(p + i * 8, pv);
            load8<GT>(g + i * 8, gv);
            load8<float>(m + i * 8, mv);
            load8<float>(v + i * 8, vv);
#pragma unroll
            for (int k = 0; k < 8; ++k) pv[k] = upd(pv[k], gv[k], mv[k], vv[k]);
            store8<float>(p + i * 8, pv);
            store8<float>(m + i * 8, mv);
            store8<float>(v + i * 8, vv);
            if (lp) store8<LT>(lp + i * 8, pv);
        }
        done = nvec * 8;
    }
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) {
        float mv = m[i], vv = v[i];
        const float pv = upd(p[i], Cvt<GT>::to_f(g[i]), mv, vv);
        p[i] = pv; m[i] = mv; v[i] = vv;
        if (lp) lp[i] = Cvt<LT>::from_f(pv);
    }
}

template <typename GT, typename LT>
__global__ void __launch_bounds__(kThreads) multi_sgd_kernel(
    const b200dp_tensor_desc* __restrict__ descs, const b200dp_chunk* __restrict__ chunks,
    b200dp_sgd_hparams hp, int first_step, const float* __restrict__ grad_mult_p,
    const float* __restrict__ found_inf_p) {
    if (found_inf_p != nullptr && *found_inf_p != 0.f) return;
    const b200dp_chunk ck = chunks[blockIdx.x];
    const b200dp_tensor_desc d = descs[ck.x];
    const int64_t start = (int64_t)ck.y * B200DP_CHUNK;
    const int64_t len = min((int64_t)B200DP_CHUNK, d.n - start);
    float* p = (float*)d.p + start;
    const GT* g = (const GT*)d.g + start;
    float* m = d.m ? (float*)d.m + start : nullptr;
    LT* lp = d.p_lp ? (LT*)d.p_lp + start : nullptr;
    const float gm = grad_mult_p ? *grad_mult_p : 1.f;
    const float sign = hp.maximize ? -1.f : 1.f;
    const float lr = hp.lr_dev ? *hp.lr_dev : hp.lr;
    for (int64_t i = threadIdx.x; i < len; i += kThreads) {
        float pv = p[i];
        float gv = Cvt<GT>::to_f(g[i]) * gm * sign + hp.weight_decay * pv;
        if (hp.momentum != 0.f && m != nullptr) {
            float mv = first_step ? gv : hp.momentum * m[i] + (1.f - hp.dampening) * gv;
            m[i] = mv;
            gv = hp.nesterov ? gv + hp.momentum * mv : mv;
        }
        pv -= lr * gv;
        p[i] = pv;
        if (lp) lp[i] = Cvt<LT>::from_f(pv);
    }
}

__global__ void step_advance_kernel(float* step, const float* found_inf) {
    if (found_inf == nullptr || *found_inf == 0.f) *step += 1.f;
}

template <typename GT>
__global__ void __launch_bounds__(kThreads) multi_unscale_kernel(
    const b200dp_tensor_desc* __restrict__ descs, const b200dp_chunk* __restrict__ chunks,
    const float* __restrict__ inv_scale_p, float* found_inf) {
    const b200dp_chunk ck = chunks[blockIdx.x];
    const b200dp_tensor_desc d = descs[ck.x];
    const int64_t start = (int64_t)ck.y * B200DP_CHUNK;
    const int64_t len = min((int64_t)B200DP_CHUNK, d.n - start);
    GT* g = (GT*)d.g + start;
    const float s = *inv_scale_p;
    bool bad = false;
    int64_t done = 0;
    if (is_aligned16(g)) {
        const int64_t nvec = len / 8;
        for (int64_t i = threadIdx.x; i < nvec; i += kThreads) {
            float gv[8];
            load8<GT>(g + i * 8, gv);
#pragma unroll
            for (int k = 0; k < 8; ++k) { bad |= !isfinite(gv[k]); gv[k] *= s; }
            store8<GT>(g + i * 8, gv);
        }
        done = nvec * 8;
    }
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) {
        float gv = Cvt<GT>::to_f(g[i]);
        bad |= !isfinite(gv);
        g[i] = Cvt<GT>::from_f(gv * s);
    }
    if (__syncthreads_or(bad) && threadIdx.x == 0) *found_inf = 1.f;
}

// Sum of squares + non-finite check; per-CTA partial -> scratch, last CTA (ticket) sums
// the partials in index order so the result does not depend on scheduling.
template <typename GT>
__global__ void __launch_bounds__(kThreads) multi_sumsq_kernel(
    const b200dp_tensor_desc* __restrict__ descs, const b200dp_chunk* __restrict__ chunks, int n_chunks,
    float* __restrict__ scratch, unsigned int* __restrict__ counter, float* __restrict__ out) {
    __shared__ float red[32];
    __shared__ bool is_last;
    const b200dp_chunk ck = chunks[blockIdx.x];
    const b200dp_tensor_desc d = descs[ck.x];
    const int64_t start = (int64_t)ck.y * B200DP_CHUNK;
    const int64_t len = min((int64_t)B200DP_CHUNK, d.n - start);
    const GT* g = (const GT*)d.g + start;
    float acc = 0.f;
    bool bad = false;
    int64_t done = 0;
    if (is_aligned16(g)) {
        const int64_t nvec = len / 8;
        for (int64_t i = threadIdx.x; i < nvec; i += kThreads) {
            float gv[8];
            load8<GT>(g + i * 8, gv);
#pragma unroll
            for (int k = 0; k < 8; ++k) { bad |= !isfinite(gv[k]); acc += gv[k] * gv[k]; }
        }
        done = nvec * 8;
    }
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) {
        const float gv = Cvt<GT>::to_f(g[i]);
        bad |= !isfinite(gv);
        acc += gv * gv;
    }
    const float tot = block_sum(acc, red);
    const bool any_bad = __syncthreads_or(bad);
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = tot;
        if (any_bad) out[1] = 1.f;
        __threadfence();
        const unsigned int t = atomicAdd(counter, 1u);
        is_last = (t == (unsigned int)n_chunks - 1u);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float s = 0.f;
        for (int i = threadIdx.x; i < n_chunks; i += kThreads) s += __ldcg(scratch + i);
        s = block_sum(s, red);
        if (threadIdx.x == 0) { out[0] = s; *counter = 0u; }
    }
}

__global__ void clip_coef_kernel(const float* sumsq, const float* inv_scale, float max_norm, float* grad_mult,
                                 float* total_norm) {
    const float is = inv_scale ? *inv_scale : 1.f;
    const float norm = sqrtf(*sumsq) * is;
    if (total_norm) *total_norm = norm;
    *grad_mult = is * fminf(1.f, max_norm / (norm + 1e-6f));
}

__global__ void scaler_update_kernel(float* scale, int* tracker, float* found_inf, float growth, float backoff,
                                     int interval) {
    if (*found_inf != 0.f) {
        *scale *= backoff;
        *tracker = 0;
    } else {
        const int t = *tracker + 1;
        if (t == interval) {
            const float ns = *scale * growth;
            if (isfinite(ns)) *scale = ns;
            *tracker = 0;
        } else {
            *tracker = t;
        }
    }
    *found_inf = 0.f;
}

constexpr int kFlatPerBlock = 16384;

__global__ void __launch_bounds__(kThreads) sumsq_f32_kernel(const float* __restrict__ x, int64_t n,
                                                             float* __restrict__ scratch,
                                                             unsigned int* __restrict__ counter,
                                                             float* __restrict__ out) {
    __shared__ float red[32];
    __shared__ bool is_last;
    const int64_t start = (int64_t)blockIdx.x * kFlatPerBlock;
    const int64_t len = min((int64_t)kFlatPerBlock, n - start);
    const float* xp = x + start;
    float acc = 0.f;
    int64_t done = 0;
    if (is_aligned16(xp)) {
        const int64_t nvec = len / 4;
        for (int64_t i = threadIdx.x; i < nvec; i += kThreads) {
            const float4 v = reinterpret_cast<const float4*>(xp)[i];
            acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
        done = nvec * 4;
    }
    for (int64_t i = done + threadIdx.x; i < len; i += kThreads) acc += xp[i] * xp[i];
    const float tot = block_sum(acc, red);
    if (threadIdx.x == 0) {
        scratch[blockIdx.x] = tot;
        __threadfence();
        is_last = (atomicAdd(counter, 1u) == gridDim.x - 1u);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float s = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += kThreads) s += __ldcg(scratch + i);
        s = block_sum(s, red);
        if (threadIdx.x == 0) { *out = s; *counter = 0u; }
    }
}

}  // namespace b200dp

using namespace b200dp;

extern "C" {

const char* b200dp_last_error(void) { return b200dp::g_err; }
int b200dp_version(void) { return 100; }

int b200dp_multi_adam(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                      b200dp_dtype grad_dtype, b200dp_dtype lp_dtype, b200dp_adam_hparams hp, const float* step,
                      const float* grad_mult, const float* found_inf, void* stream) {
    if (n_chunks <= 0) return 0;
    B200DP_REQUIRE(descs && chunks && step, "multi_adam: null table or step pointer");
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_DTYPE(grad_dtype, GT, {
        if (lp_dtype == B200DP_BF16)
            multi_adam_kernel<GT, __nv_bfloat16><<<n_chunks, kThreads, 0, s>>>(descs, chunks, hp, step, grad_mult, found_inf);
        else
            multi_adam_kernel<GT, __half><<<n_chunks, kThreads, 0, s>>>(descs, chunks, hp, step, grad_mult, found_inf);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_multi_sgd(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                     b200dp_dtype grad_dtype, b200dp_dtype lp_dtype, b200dp_sgd_hparams hp, int first_step,
                     const float* grad_mult, const float* found_inf, void* stream) {
    if (n_chunks <= 0) return 0;
    B200DP_REQUIRE(descs && chunks, "multi_sgd: null table");
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_DTYPE(grad_dtype, GT, {
        if (lp_dtype == B200DP_BF16)
            multi_sgd_kernel<GT, __nv_bfloat16><<<n_chunks, kThreads, 0, s>>>(descs, chunks, hp, first_step, grad_mult, found_inf);
        else
            multi_sgd_kernel<GT, __half><<<n_chunks, kThreads, 0, s>>>(descs, chunks, hp, first_step, grad_mult, found_inf);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_step_advance(float* step, const float* found_inf, void* stream) {
    B200DP_REQUIRE(step, "step_advance: null step");
    step_advance_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step, found_inf);
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_multi_unscale_check(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                               b200dp_dtype grad_dtype, const float* inv_scale, float* found_inf, void* stream) {
    if (n_chunks <= 0) return 0;
    B200DP_REQUIRE(descs && chunks && inv_scale && found_inf, "multi_unscale_check: null argument");
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_DTYPE(grad_dtype, GT, {
        multi_unscale_kernel<GT><<<n_chunks, kThreads, 0, s>>>(descs, chunks, inv_scale, found_inf);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_multi_sumsq_check(const b200dp_tensor_desc* descs, const b200dp_chunk* chunks, int n_chunks,
                             b200dp_dtype grad_dtype, float* scratch, unsigned int* counter, float* out,
                             void* stream) {
    B200DP_REQUIRE(out, "multi_sumsq_check: null out");
    cudaStream_t s = (cudaStream_t)stream;
    if (n_chunks <= 0) {
        B200DP_CUDA(cudaMemsetAsync(out, 0, sizeof(float), s));
        return 0;
    }
    B200DP_REQUIRE(descs && chunks && scratch && counter, "multi_sumsq_check: null argument");
    B200DP_DISPATCH_DTYPE(grad_dtype, GT, {
        multi_sumsq_kernel<GT><<<n_chunks, kThreads, 0, s>>>(descs, chunks, n_chunks, scratch, counter, out);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_clip_coef(const float* sumsq, const float* inv_scale, float max_norm, float* grad_mult,
                     float* total_norm, void* stream) {
    B200DP_REQUIRE(sumsq && grad_mult, "clip_coef: null argument");
    clip_coef_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sumsq, inv_scale, max_norm, grad_mult, total_norm);
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_scaler_update(float* scale, int* growth_tracker, float* found_inf, float growth_factor,
                         float backoff_factor, int growth_interval, void* stream) {
    B200DP_REQUIRE(scale && growth_tracker && found_inf, "scaler_update: null argument");
    scaler_update_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(scale, growth_tracker, found_inf, growth_factor,
                                                            backoff_factor, growth_interval);
    B200DP_CHECK_LAUNCH();
    return 0;
}

int64_t b200dp_sumsq_scratch_floats(int64_t n) { return (n + kFlatPerBlock - 1) / kFlatPerBlock + 1; }

int b200dp_sumsq_f32(const float* x, int64_t n, float* scratch, unsigned int* counter, float* out, void* stream) {
    B200DP_REQUIRE(out, "sumsq_f32: null out");
    cudaStream_t s = (cudaStream_t)stream;
    if (n <= 0) {
        B200DP_CUDA(cudaMemsetAsync(out, 0, sizeof(float), s));
        return 0;
    }
    B200DP_REQUIRE(x && scratch && counter, "sumsq_f32: null argument");
    const int64_t blocks = (n + kFlatPerBlock - 1) / kFlatPerBlock;
    B200DP_REQUIRE(blocks < (1ll << 31), "sumsq_f32: n too large");
    sumsq_f32_kernel<<<(unsigned)blocks, kThreads, 0, s>>>(x, n, scratch, counter, out);
    B200DP_CHECK_LAUNCH();
    return 0;
}

}  // extern "C"
