// NHWC max-pooling forward / backward (HBM-bound).  The forward stores, next to the output, the
// position (0 .. K*K-1, first maximum in row-major window order, like ATen) of the winning tap in
// ONE byte per output element; the backward is then a pure gather: every input pixel looks at the
// <= ceil(K/stride)^2 windows that cover it and takes dy where the stored position points back at
// it.  No 8-byte index tensor, no zero-fill pass, no atomics: forward moves x + y + y/2 bytes,
// backward moves dx + (dy + idx, re-read through L1/L2).
#include "common.cuh"

namespace b200dp {

// KT / ST > 0: compile-time window and stride (all taps are issued as independent predicated loads,
// so 9 x 16 B are in flight per thread); KT == 0: generic run-time loop.
template <typename T, int KT, int ST>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y,
                                                          uint8_t* __restrict__ idx, int N, int H, int W, int C, int P, int Q,
                                                          int Krt, int stride_rt, int pad) {
    const int K = KT > 0 ? KT : Krt, stride = KT > 0 ? ST : stride_rt;
    const int cgs = C / 8;
    const int64_t total = (int64_t)N * P * Q * cgs;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int cg = (int)(t % cgs);
        int64_t pix = t / cgs;
        const int q = (int)(pix % Q); pix /= Q;
        const int p = (int)(pix % P);
        const int n = (int)(pix / P);
        float best[8];
        uint8_t bi[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { best[k] = -INFINITY; bi[k] = 0; }
        const T* xn = x + (int64_t)n * H * W * C + cg * 8;
        if (KT > 0) {
            constexpr int kTaps = KT > 0 ? KT * KT : 1;
            Vec8<T> v[kTaps];
            bool ok[kTaps];
#pragma unroll
            for (int r = 0; r < KT; ++r)
#pragma unroll
                for (int s = 0; s < KT; ++s) {
                    const int h = p * stride - pad + r, w = q * stride - pad + s;
                    ok[r * KT + s] = h >= 0 && h < H && w >= 0 && w < W;
                    const int hc = min(max(h, 0), H - 1), wc = min(max(w, 0), W - 1);
                    v[r * KT + s] = *reinterpret_cast<const Vec8<T>*>(xn + ((int64_t)hc * W + wc) * C);
                }
#pragma unroll
            for (int tap = 0; tap < KT * KT; ++tap)
                if (ok[tap]) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const float f = Cvt<T>::to_f(v[tap].v[k]);
                        if (f > best[k] || (f != f && best[k] == best[k])) { best[k] = f; bi[k] = (uint8_t)tap; }
                    }
                }
        } else {
            for (int r = 0; r < K; ++r) {
                const int h = p * stride - pad + r;
                if (h < 0 || h >= H) continue;
                for (int s = 0; s < K; ++s) {
                    const int w = q * stride - pad + s;
                    if (w < 0 || w >= W) continue;
                    float f[8];
                    load8<T>(xn + ((int64_t)h * W + w) * C, f);
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        if (f[k] > best[k] || (f[k] != f[k] && best[k] == best[k])) { best[k] = f[k]; bi[k] = (uint8_t)(r * K + s); }
                }
            }
        }
        const int64_t o = (((int64_t)n * P + p) * Q + q) * C + cg * 8;
        store8<T>(y + o, best);
        *reinterpret_cast<uint2*>(idx + o) = *reinterpret_cast<const uint2*>(bi);
    }
}

template <typename T, int KT, int ST>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ idx,
                                                          T* __restrict__ dx, int N, int H, int W, int C, int P, int Q,
                                                          int Krt, int stride_rt, int pad) {
    const int K = KT > 0 ? KT : Krt, stride = KT > 0 ? ST : stride_rt;
    constexpr int kMaxWin = KT > 0 ? (KT + ST - 1) / ST : 1;       // windows covering a pixel, per dimension
    const int cgs = C / 8;
    const int64_t total = (int64_t)N * H * W * cgs;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int cg = (int)(t % cgs);
        int64_t pix = t / cgs;
        const int w = (int)(pix % W); pix /= W;
        const int h = (int)(pix % H);
        const int n = (int)(pix / H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // windows p with p*stride - pad <= h <= p*stride - pad + K - 1
        const int p_lo = max(0, (h + pad - (K - 1) + stride - 1) / stride), p_hi = min(P - 1, (h + pad) / stride);
        const int q_lo = max(0, (w + pad - (K - 1) + stride - 1) / stride), q_hi = min(Q - 1, (w + pad) / stride);
        if (KT > 0) {
            uint2 packed[kMaxWin * kMaxWin];
            Vec8<T> g[kMaxWin * kMaxWin];
            bool ok[kMaxWin * kMaxWin];
#pragma unroll
            for (int a = 0; a < kMaxWin; ++a)
#pragma unroll
                for (int b = 0; b < kMaxWin; ++b) {
                    const int p = p_lo + a, q = q_lo + b;
                    ok[a * kMaxWin + b] = p <= p_hi && q <= q_hi;
                    const int64_t o = (((int64_t)n * P + min(p, P - 1)) * Q + min(q, Q - 1)) * C + cg * 8;
                    packed[a * kMaxWin + b] = *reinterpret_cast<const uint2*>(idx + o);
                    g[a * kMaxWin + b] = *reinterpret_cast<const Vec8<T>*>(dy + o);
                }
#pragma unroll
            for (int a = 0; a < kMaxWin; ++a)
#pragma unroll
                for (int b = 0; b < kMaxWin; ++b)
                    if (ok[a * kMaxWin + b]) {
                        const int r = h + pad - (p_lo + a) * stride, s = w + pad - (q_lo + b) * stride;
                        const uint8_t me = (uint8_t)(r * K + s);
                        const uint8_t* bi = reinterpret_cast<const uint8_t*>(&packed[a * kMaxWin + b]);
#pragma unroll
                        for (int k = 0; k < 8; ++k) acc[k] += (bi[k] == me) ? Cvt<T>::to_f(g[a * kMaxWin + b].v[k]) : 0.f;
                    }
        } else {
            for (int p = p_lo; p <= p_hi; ++p) {
                const int r = h + pad - p * stride;
                for (int q = q_lo; q <= q_hi; ++q) {
                    const int s = w + pad - q * stride;
                    const uint8_t me = (uint8_t)(r * K + s);
                    const int64_t o = (((int64_t)n * P + p) * Q + q) * C + cg * 8;
                    const uint2 packed = *reinterpret_cast<const uint2*>(idx + o);
                    const uint8_t* bi = reinterpret_cast<const uint8_t*>(&packed);
                    float g[8];
                    load8<T>(dy + o, g);
#pragma unroll
                    for (int k = 0; k < 8; ++k) acc[k] += (bi[k] == me) ? g[k] : 0.f;
                }
            }
        }
        store8<T>(dx + (((int64_t)n * H + h) * W + w) * C + cg * 8, acc);
    }
}

}  // namespace b200dp

using namespace b200dp;

extern "C" {

int b200dp_maxpool_fwd(const void* x, void* y, unsigned char* idx, int N, int H, int W, int C, int K, int stride, int pad,
                       b200dp_dtype dtype, void* stream) {
    B200DP_REQUIRE(x && y && idx && C % 8 == 0 && K >= 1 && K <= 15 && stride >= 1, "maxpool_fwd: bad arguments");
    const int P = (H + 2 * pad - K) / stride + 1, Q = (W + 2 * pad - K) / stride + 1;
    const int64_t total = (int64_t)N * P * Q * (C / 8);
    const int blocks = (int)min((int64_t)B200DP_SMS * 16, (total + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_HALF(dtype, T, {
        if (K == 3 && stride == 2) maxpool_fwd_kernel<T, 3, 2><<<blocks, 256, 0, s>>>((const T*)x, (T*)y, idx, N, H, W, C, P, Q, K, stride, pad);
        else if (K == 2 && stride == 2) maxpool_fwd_kernel<T, 2, 2><<<blocks, 256, 0, s>>>((const T*)x, (T*)y, idx, N, H, W, C, P, Q, K, stride, pad);
        else maxpool_fwd_kernel<T, 0, 0><<<blocks, 256, 0, s>>>((const T*)x, (T*)y, idx, N, H, W, C, P, Q, K, stride, pad);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_maxpool_bwd(const void* dy, const unsigned char* idx, void* dx, int N, int H, int W, int C, int K, int stride,
                       int pad, b200dp_dtype dtype, void* stream) {
    B200DP_REQUIRE(dy && dx && idx && C % 8 == 0 && K >= 1 && K <= 15 && stride >= 1, "maxpool_bwd: bad arguments");
    const int P = (H + 2 * pad - K) / stride + 1, Q = (W + 2 * pad - K) / stride + 1;
    const int64_t total = (int64_t)N * H * W * (C / 8);
    const int blocks = (int)min((int64_t)B200DP_SMS * 16, (total + 255) / 256);
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_HALF(dtype, T, {
        if (K == 3 && stride == 2) maxpool_bwd_kernel<T, 3, 2><<<blocks, 256, 0, s>>>((const T*)dy, idx, (T*)dx, N, H, W, C, P, Q, K, stride, pad);
        else if (K == 2 && stride == 2) maxpool_bwd_kernel<T, 2, 2><<<blocks, 256, 0, s>>>((const T*)dy, idx, (T*)dx, N, H, W, C, P, Q, K, stride, pad);
        else maxpool_bwd_kernel<T, 0, 0><<<blocks, 256, 0, s>>>((const T*)dy, idx, (T*)dx, N, H, W, C, P, Q, K, stride, pad);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

}  // extern "C"
