// Loss kernels (W9): softmax cross-entropy forward+backward in one pass, MSE.
#include "common.cuh"

namespace b200dp {

// One CTA per row.  The row is read once into registers (classes <= kThreads*kPer),
// max / sum-exp are block reductions, and the gradient (softmax - onehot)/rows is
// written from the same registers.  The last CTA to finish averages the per-row losses
// in index order (deterministic).
constexpr int kCEThreads = 256;
constexpr int kCEPer = 16;  // up to 4096 classes in registers; larger rows take the looped path

template <typename T>
__global__ void __launch_bounds__(kCEThreads) cross_entropy_kernel(
    const T* __restrict__ logits, const int64_t* __restrict__ labels, int rows, int classes,
    float* __restrict__ loss_rows, T* __restrict__ dlogits, float* __restrict__ loss_mean,
    unsigned int* __restrict__ counter_in_loss_rows_tail, float smoothing) {
    __shared__ float red[32];
    __shared__ bool is_last;
    const int row = blockIdx.x;
    const T* x = logits + (int64_t)row * classes;
    const int label = (int)labels[row];
    const bool in_regs = classes <= kCEThreads * kCEPer;
    float vals[kCEPer];
    float mx = -INFINITY;
    if (in_regs) {
#pragma unroll
        for (int k = 0; k < kCEPer; ++k) {
            const int c = threadIdx.x + k * kCEThreads;
            vals[k] = c < classes ? Cvt<T>::to_f(x[c]) : -INFINITY;
            mx = fmaxf(mx, vals[k]);
        }
    } else {
        for (int c = threadIdx.x; c < classes; c += kCEThreads) mx = fmaxf(mx, Cvt<T>::to_f(x[c]));
    }
    mx = block_max(mx, red);
    float se = 0.f, sx = 0.f;
    if (in_regs) {
#pragma unroll
        for (int k = 0; k < kCEPer; ++k) {
            const int c = threadIdx.x + k * kCEThreads;
            if (c < classes) { se += __expf(vals[k] - mx); sx += vals[k]; }
        }
    } else {
        for (int c = threadIdx.x; c < classes; c += kCEThreads) {
            const float v = Cvt<T>::to_f(x[c]);
            se += __expf(v - mx); sx += v;
        }
    }
    se = block_sum(se, red);
    sx = block_sum(sx, red);
    const float lse = mx + __logf(se);
    const float x_label = Cvt<T>::to_f(x[label]);
    // loss = (1-s) * (lse - x_label) + s * (lse - mean(x))
    const float loss = (1.f - smoothing) * (lse - x_label) + smoothing * (lse - sx / classes);
    if (dlogits != nullptr) {
        T* dx = dlogits + (int64_t)row * classes;
        const float inv_rows = 1.f / rows;
        const float inv_se = 1.f / se;
        const float off = smoothing / classes;
        if (in_regs) {
#pragma unroll
            for (int k = 0; k < kCEPer; ++k) {
                const int c = threadIdx.x + k * kCEThreads;
                if (c < classes) {
                    const float p = __expf(vals[k] - mx) * inv_se;
                    const float t = (c == label ? 1.f - smoothing : 0.f) + off;
                    dx[c] = Cvt<T>::from_f((p - t) * inv_rows);
                }
            }
        } else {
            for (int c = threadIdx.x; c < classes; c += kCEThreads) {
                const float p = __expf(Cvt<T>::to_f(x[c]) - mx) * inv_se;
                const float t = (c == label ? 1.f - smoothing : 0.f) + off;
                dx[c] = Cvt<T>::from_f((p - t) * inv_rows);
            }
        }
    }
    if (threadIdx.x == 0) {
        loss_rows[row] = loss;
        __threadfence();
        is_last = (atomicAdd(counter_in_loss_rows_tail, 1u) == (unsigned)rows - 1u);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float s = 0.f;
        for (int i = threadIdx.x; i < rows; i += kCEThreads) s += __ldcg(loss_rows + i);
        s = block_sum(s, red);
        if (threadIdx.x == 0) { *loss_mean = s / rows; *counter_in_loss_rows_tail = 0u; }
    }
}

template <typename T>
__global__ void scale_inplace_kernel(T* x, int64_t n, const float* scale) {
    const float s = *scale;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        x[i] = Cvt<T>::from_f(Cvt<T>::to_f(x[i]) * s);
}

constexpr int kMSEPerBlock = 16384;
template <typename T>
__global__ void __launch_bounds__(256) mse_kernel(const T* __restrict__ a, const T* __restrict__ b, int64_t n,
                                                  T* __restrict__ grad, float* __restrict__ scratch,
                                                  unsigned int* __restrict__ counter, float* __restrict__ out) {
    __shared__ float red[32];
    __shared__ bool is_last;
    const int64_t start = (int64_t)blockIdx.x * kMSEPerBlock;
    const int64_t len = min((int64_t)kMSEPerBlock, n - start);
    const float gs = 2.f / (float)n;
    float acc = 0.f;
    for (int64_t i = threadIdx.x; i < len; i += 256) {
        const float d = Cvt<T>::to_f(a[start + i]) - Cvt<T>::to_f(b[start + i]);
        acc += d * d;
        if (grad) grad[start + i] = Cvt<T>::from_f(d * gs);
    }
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
        for (int i = threadIdx.x; i < (int)gridDim.x; i += 256) s += __ldcg(scratch + i);
        s = block_sum(s, red);
        if (threadIdx.x == 0) { *out = s / (float)n; *counter = 0u; }
    }
}

}  // namespace b200dp

using namespace b200dp;

extern "C" {

int b200dp_cross_entropy(const void* logits, const int64_t* labels, int rows, int classes, b200dp_dtype dtype,
                         float* loss_rows, void* dlogits, float* loss_mean, float label_smoothing, void* stream) {
    B200DP_REQUIRE(logits && labels && loss_rows && loss_mean, "cross_entropy: null argument");
    B200DP_REQUIRE(rows > 0 && classes > 0, "cross_entropy: empty input");
    // loss_rows must have rows+1 floats: the extra word is the (zeroed) completion ticket
    unsigned int* counter = reinterpret_cast<unsigned int*>(loss_rows + rows);
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_DTYPE(dtype, T, {
        cross_entropy_kernel<T><<<rows, kCEThreads, 0, s>>>((const T*)logits, labels, rows, classes, loss_rows,
                                                            (T*)dlogits, loss_mean, counter, label_smoothing);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_scale_inplace(void* x, int64_t n, b200dp_dtype dtype, const float* scale, void* stream) {
    if (n <= 0) return 0;
    B200DP_REQUIRE(x && scale, "scale_inplace: null argument");
    const int blocks = (int)min((int64_t)B200DP_SMS * 8, (n + 255) / 256);
    B200DP_DISPATCH_DTYPE(dtype, T, {
        scale_inplace_kernel<T><<<blocks, 256, 0, (cudaStream_t)stream>>>((T*)x, n, scale);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_mse(const void* a, const void* b, int64_t n, b200dp_dtype dtype, void* grad, float* scratch,
               unsigned int* counter, float* out, void* stream) {
    B200DP_REQUIRE(a && b && scratch && counter && out && n > 0, "mse: bad argument");
    const int64_t blocks = (n + kMSEPerBlock - 1) / kMSEPerBlock;
    B200DP_DISPATCH_DTYPE(dtype, T, {
        mse_kernel<T><<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const T*)a, (const T*)b, n, (T*)grad,
                                                                         scratch, counter, out);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

}  // extern "C"
