// Normalisation kernels (W8, W6 local part, LayerNorm for ViT / SwinIR).
//
// BatchNorm works on channels-last activations viewed as x[rows, C].  A thread owns one
// group of 8 consecutive channels (one 16-byte access) for the whole kernel, so the
// per-channel constants live in registers and the row loop is a pure stream.  The two
// reductions write one partial per CTA to caller-provided scratch and the last CTA of a
// channel column adds them in a fixed order (deterministic, no float atomics).
#include "common.cuh"

namespace b200dp {

constexpr int kBNThreads = 256;
constexpr int kBNMaxCols = 64;  // counters available per call (channel columns of 256 channels)

struct BNGeom {
    int tpr;   // 8-channel groups per row
    int cgb;   // channel groups per CTA (<= 32)
    int rpb;   // rows per CTA iteration
    int gx;    // channel columns
    int gy;    // row slices
};

static inline BNGeom bn_geom(int64_t rows, int C, int per_sm = 4) {
    BNGeom g;
    g.tpr = C / 8;
    g.cgb = g.tpr < 32 ? g.tpr : 32;
    g.rpb = kBNThreads / g.cgb;
    g.gx = (g.tpr + g.cgb - 1) / g.cgb;
    int64_t want = (B200DP_SMS * per_sm + g.gx - 1) / g.gx;   // ~per_sm CTAs per SM in total
    int64_t maxy = (rows + (int64_t)g.rpb * 4 - 1) / ((int64_t)g.rpb * 4);  // >= 4 rows per thread
    if (maxy < 1) maxy = 1;
    g.gy = (int)(want < maxy ? want : maxy);
    if (g.gy < 1) g.gy = 1;
    return g;
}

// Shared tail of the two reductions: in-CTA reduce across the rpb row lanes, partial to
// scratch, last CTA of the column sums the partials.
__device__ __forceinline__ void bn_reduce_tail(float (&a)[8], float (&b)[8], int C, int cgb, int rpb, int rl,
                                               int cgl, int cg, bool active, float* __restrict__ scratch,
                                               unsigned int* __restrict__ counters, float* __restrict__ out) {
    __shared__ float sm[2][kBNThreads * 8];
    __shared__ bool is_last;
    const int nch = cgb * 8;  // channels handled by this CTA
    if (active) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            sm[0][rl * nch + cgl * 8 + k] = a[k];
            sm[1][rl * nch + cgl * 8 + k] = b[k];
        }
    }
    __syncthreads();
    const int c0 = blockIdx.x * nch;
    if ((int)threadIdx.x < nch && c0 + (int)threadIdx.x < C) {
        float s0 = 0.f, s1 = 0.f;
        for (int r = 0; r < rpb; ++r) {
            s0 += sm[0][r * nch + threadIdx.x];
            s1 += sm[1][r * nch + threadIdx.x];
        }
        float* part = scratch + (int64_t)blockIdx.y * 2 * C;
        part[c0 + threadIdx.x] = s0;
        part[C + c0 + threadIdx.x] = s1;
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(counters + blockIdx.x, 1u) == gridDim.y - 1u);
    __syncthreads();
    if (is_last) {
        __threadfence();
        if ((int)threadIdx.x < nch && c0 + (int)threadIdx.x < C) {
            float s0 = 0.f, s1 = 0.f;
            for (int y = 0; y < (int)gridDim.y; ++y) {
                const float* part = scratch + (int64_t)y * 2 * C;
                s0 += __ldcg(part + c0 + threadIdx.x);
                s1 += __ldcg(part + C + c0 + threadIdx.x);
            }
            out[c0 + threadIdx.x] = s0;
            out[C + c0 + threadIdx.x] = s1;
        }
        if (threadIdx.x == 0) counters[blockIdx.x] = 0u;
    }
}

template <typename T>
__global__ void __launch_bounds__(kBNThreads, 4) bn_reduce_kernel(const T* __restrict__ x, int64_t rows, int C,
                                                               int tpr, int cgb, int rpb,
                                                               float* __restrict__ scratch,
                                                               unsigned int* __restrict__ counters,
                                                               float* __restrict__ sums) {
    const int rl = threadIdx.x / cgb, cgl = threadIdx.x % cgb;
    const int cg = blockIdx.x * cgb + cgl;
    const bool active = rl < rpb && cg < tpr;
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (active) {
        const int64_t rstep = (int64_t)gridDim.y * rpb;
        const T* xp = x + (int64_t)cg * 8;
        int64_t r = (int64_t)blockIdx.y * rpb + rl;
        for (; r + 3 * rstep < rows; r += 4 * rstep) {
            float v0[8], v1[8], v2[8], v3[8];
            load8<T>(xp + r * C, v0);
            load8<T>(xp + (r + rstep) * C, v1);
            load8<T>(xp + (r + 2 * rstep) * C, v2);
            load8<T>(xp + (r + 3 * rstep) * C, v3);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                s[k] += (v0[k] + v1[k]) + (v2[k] + v3[k]);
                q[k] += (v0[k] * v0[k] + v1[k] * v1[k]) + (v2[k] * v2[k] + v3[k] * v3[k]);
            }
        }
        for (; r < rows; r += rstep) {
            float v[8];
            load8<T>(xp + r * C, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] += v[k] * v[k]; }
        }
    }
    bn_reduce_tail(s, q, C, cgb, rpb, rl, cgl, cg, active, scratch, counters, sums);
}

// Column sums only (bias gradients of linear layers): per-CTA partials [gridDim.y][C]; ln_partials_reduce_kernel sums
// them.  A transformer's bias gradients are many SMALL reductions (tens of MB): the single-CTA tail of the moment
// kernel above (one CTA walking ~600 partials) costs more than streaming the data, so this pair keeps the partial count
// low (>= 16 rows per thread) and spreads the second stage over several CTAs.
template <typename T>
__global__ void __launch_bounds__(kBNThreads, 4) colsum_kernel(const T* __restrict__ x, int64_t rows, int C, int tpr,
                                                            int cgb, int rpb, float* __restrict__ scratch) {
    __shared__ float sm[kBNThreads * 8];
    const int rl = threadIdx.x / cgb, cgl = threadIdx.x % cgb;
    const int cg = blockIdx.x * cgb + cgl;
    const bool active = rl < rpb && cg < tpr;
    const int nch = cgb * 8;
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (active) {
        const int64_t rstep = (int64_t)gridDim.y * rpb;
        const T* xp = x + (int64_t)cg * 8;
        int64_t r = (int64_t)blockIdx.y * rpb + rl;
        for (; r + 3 * rstep < rows; r += 4 * rstep) {
            float v0[8], v1[8], v2[8], v3[8];
            load8<T>(xp + r * C, v0);
            load8<T>(xp + (r + rstep) * C, v1);
            load8<T>(xp + (r + 2 * rstep) * C, v2);
            load8<T>(xp + (r + 3 * rstep) * C, v3);
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] += (v0[k] + v1[k]) + (v2[k] + v3[k]);
        }
        for (; r < rows; r += rstep) {
            float v[8];
            load8<T>(xp + r * C, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) s[k] += v[k];
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) sm[rl * nch + cgl * 8 + k] = s[k];
    }
    __syncthreads();
    const int c0 = blockIdx.x * nch;
    if ((int)threadIdx.x < nch && c0 + (int)threadIdx.x < C) {
        float t = 0.f;
        for (int r = 0; r < rpb; ++r) t += sm[r * nch + threadIdx.x];
        scratch[(int64_t)blockIdx.y * C + c0 + threadIdx.x] = t;
    }
}

template <typename T, bool kTrain>
__global__ void __launch_bounds__(kBNThreads, 3) bn_apply_kernel(
    const T* __restrict__ x, const T* __restrict__ res, T* __restrict__ y, int64_t rows, int C, int tpr, int cgb,
    int rpb, const float* __restrict__ sums, float count, const float* __restrict__ gamma,
    const float* __restrict__ beta, float eps, float momentum, float* __restrict__ running_mean,
    float* __restrict__ running_var, float* __restrict__ mean_out, float* __restrict__ invstd_out, int relu,
    uint8_t* __restrict__ mask_out, long long* __restrict__ num_batches_tracked = nullptr) {
    const int rl = threadIdx.x / cgb, cgl = threadIdx.x % cgb;
    const int cg = blockIdx.x * cgb + cgl;
    // nn.BatchNorm's step counter rides along (one thread of the grid) instead of costing a launch per layer and step
    if (kTrain && num_batches_tracked != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) *num_batches_tracked += 1;
    if (!(rl < rpb && cg < tpr)) return;
    float sc[8], sh[8];
    const int c0 = cg * 8;
    // count <= 0: the element count rides in sums[2C] (cross-rank statistics: the all-reduced TRUE count, so ranks with
    // ragged last batches still normalise by what was really summed)
    if (kTrain && !(count > 0.f)) count = sums[2 * C];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float mean, invstd;
        if (kTrain) {
            mean = sums[c0 + k] / count;
            const float var = fmaxf(sums[C + c0 + k] / count - mean * mean, 0.f);
            invstd = rsqrtf(var + eps);
            if (blockIdx.y == 0 && rl == 0) {
                mean_out[c0 + k] = mean;
                invstd_out[c0 + k] = invstd;
                if (running_mean != nullptr) {
                    const float unbiased = var * (count / fmaxf(count - 1.f, 1.f));
                    running_mean[c0 + k] = (1.f - momentum) * running_mean[c0 + k] + momentum * mean;
                    running_var[c0 + k] = (1.f - momentum) * running_var[c0 + k] + momentum * unbiased;
                }
            }
        } else {
            mean = running_mean[c0 + k];
            invstd = rsqrtf(running_var[c0 + k] + eps);
        }
        sc[k] = gamma[c0 + k] * invstd;
        sh[k] = beta[c0 + k] - mean * sc[k];
    }
    const int64_t rstep = (int64_t)gridDim.y * rpb;
    auto finish = [&](float (&v)[8], int64_t r) {
        if (relu) {
            if (mask_out != nullptr) {      // 1 bit per element: the backward never has to read y
                unsigned m = 0;
#pragma unroll
                for (int k = 0; k < 8; ++k) m |= (v[k] > 0.f ? 1u : 0u) << k;
                mask_out[r * tpr + cg] = (uint8_t)m;
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], 0.f);
        }
        store8<T>(y + r * C + c0, v);
    };
    // The tensor is walked from its END to its start: x was just written front-to-back by the producing
    // convolution, so its tail is what is still resident in the 126 MB L2 when this kernel starts.
    const int64_t last = rows - 1;
    int64_t r = (int64_t)blockIdx.y * rpb + rl;
    // four rows per iteration while they last: up to eight independent 16-byte loads in flight per thread
    for (; r + 3 * rstep < rows; r += 4 * rstep) {
        Vec8<T> xr[4], rr[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t o = (last - (r + u * rstep)) * C + c0;
            xr[u] = *reinterpret_cast<const Vec8<T>*>(x + o);
            if (res != nullptr) rr[u] = *reinterpret_cast<const Vec8<T>*>(res + o);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = Cvt<T>::to_f(xr[u].v[k]) * sc[k] + sh[k];
            if (res != nullptr) {
#pragma unroll
                for (int k = 0; k < 8; ++k) v[k] += Cvt<T>::to_f(rr[u].v[k]);
            }
            finish(v, last - (r + u * rstep));
        }
    }
    for (; r + rstep < rows; r += 2 * rstep) {
        const int64_t p0 = last - r, p1 = last - (r + rstep);
        float v0[8], v1[8];
        load8<T>(x + p0 * C + c0, v0);
        load8<T>(x + p1 * C + c0, v1);
        if (res != nullptr) {
            float r0[8], r1[8];
            load8<T>(res + p0 * C + c0, r0);
            load8<T>(res + p1 * C + c0, r1);
#pragma unroll
            for (int k = 0; k < 8; ++k) { v0[k] = v0[k] * sc[k] + sh[k] + r0[k]; v1[k] = v1[k] * sc[k] + sh[k] + r1[k]; }
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) { v0[k] = v0[k] * sc[k] + sh[k]; v1[k] = v1[k] * sc[k] + sh[k]; }
        }
        finish(v0, p0);
        finish(v1, p1);
    }
    for (; r < rows; r += rstep) {
        const int64_t p0 = last - r;
        float v[8];
        load8<T>(x + p0 * C + c0, v);
        if (res != nullptr) {
            float rv[8];
            load8<T>(res + p0 * C + c0, rv);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = v[k] * sc[k] + sh[k] + rv[k];
        } else {
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = v[k] * sc[k] + sh[k];
        }
        finish(v, p0);
    }
}

template <typename T>
__global__ void __launch_bounds__(kBNThreads, 3) bn_bwd_reduce_kernel(
    const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, int64_t rows, int C, int tpr,
    int cgb, int rpb, const float* __restrict__ mean, const float* __restrict__ invstd, int relu,
    float* __restrict__ scratch, unsigned int* __restrict__ counters, float* __restrict__ dsums,
    const uint8_t* __restrict__ mask) {
    const int rl = threadIdx.x / cgb, cgl = threadIdx.x % cgb;
    const int cg = blockIdx.x * cgb + cgl;
    const bool active = rl < rpb && cg < tpr;
    float a[8] = {0, 0, 0, 0, 0, 0, 0, 0}, b[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (active) {
        const int c0 = cg * 8;
        const int64_t rstep = (int64_t)gridDim.y * rpb;
        auto masked = [&](float (&d)[8], int64_t off) {      // off = row * C + c0
            if (!relu) return;
            if (mask != nullptr) {
                const unsigned m = mask[off >> 3];
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] = ((m >> k) & 1u) ? d[k] : 0.f;
            } else {
                float yv[8];
                load8<T>(y + off, yv);
#pragma unroll
                for (int k = 0; k < 8; ++k) d[k] = yv[k] > 0.f ? d[k] : 0.f;
            }
        };
        // The loop keeps only sum(dz) and sum(dz * x) (16 accumulators): sum(dz * xhat) is recovered as
        // (sum(dz*x) - mean * sum(dz)) * invstd afterwards.  Few live registers => two rows (four 16-byte
        // loads) in flight per thread at 4 CTAs/SM, which is what the HBM latency needs.
        // Walked from the END of the tensor to its start: the tail of dy (just written front-to-back by the
        // kernel before this one) is still in L2, and bn_bwd_apply, which runs next and walks front-to-back,
        // then starts on the rows this kernel touched last.
        // element offset of (row, c0); the mask byte of that group is at offset / 8
        const int64_t rs = rstep * C;
        int64_t left = rows - ((int64_t)blockIdx.y * rpb + rl);          // rows still ahead of this thread (incl. its current one)
        int64_t off = (rows - 1 - ((int64_t)blockIdx.y * rpb + rl)) * C + c0;
        // four rows per iteration while they last: eight independent 16-byte loads (plus the mask bytes) in flight per
        // thread, held as raw vectors until used — a read-only kernel needs more bytes in flight than a copy does
        for (; left > 3 * rstep; left -= 4 * rstep, off -= 4 * rs) {
            Vec8<T> dr[4], xr[4];
            unsigned mk[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                dr[u] = *reinterpret_cast<const Vec8<T>*>(dy + off - u * rs);
                xr[u] = *reinterpret_cast<const Vec8<T>*>(x + off - u * rs);
                mk[u] = (relu && mask != nullptr) ? mask[(off - u * rs) >> 3] : 0xffu;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float d = Cvt<T>::to_f(dr[u].v[k]);
                    const float xv = Cvt<T>::to_f(xr[u].v[k]);
                    if (relu && mask == nullptr) d = Cvt<T>::to_f(y[off - u * rs + k]) > 0.f ? d : 0.f;
                    d = ((mk[u] >> k) & 1u) ? d : 0.f;
                    a[k] += d;
                    b[k] = fmaf(d, xv, b[k]);
                }
            }
        }
        for (; left > rstep; left -= 2 * rstep, off -= 2 * rs) {
            float d0[8], x0[8], d1[8], x1[8];
            load8<T>(dy + off, d0);
            load8<T>(x + off, x0);
            load8<T>(dy + off - rs, d1);
            load8<T>(x + off - rs, x1);
            masked(d0, off);
            masked(d1, off - rs);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                a[k] += d0[k] + d1[k];
                b[k] += d0[k] * x0[k] + d1[k] * x1[k];
            }
        }
        if (left > 0) {
            float d[8], xv[8];
            load8<T>(dy + off, d);
            load8<T>(x + off, xv);
            masked(d, off);
#pragma unroll
            for (int k = 0; k < 8; ++k) { a[k] += d[k]; b[k] += d[k] * xv[k]; }
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) b[k] = (b[k] - mean[c0 + k] * a[k]) * invstd[c0 + k];
    }
    bn_reduce_tail(a, b, C, cgb, rpb, rl, cgl, cg, active, scratch, counters, dsums);
}

template <typename T>
__global__ void __launch_bounds__(kBNThreads, 4) bn_bwd_apply_kernel(
    const T* __restrict__ dy, const T* __restrict__ y, const T* __restrict__ x, T* __restrict__ dx,
    T* __restrict__ dres, int64_t rows, int C, int tpr, int cgb, int rpb, const float* __restrict__ mean,
    const float* __restrict__ invstd, const float* __restrict__ gamma, const float* __restrict__ dsums,
    float count, int relu, const uint8_t* __restrict__ mask) {
    const int rl = threadIdx.x / cgb, cgl = threadIdx.x % cgb;
    const int cg = blockIdx.x * cgb + cgl;
    if (!(rl < rpb && cg < tpr)) return;
    const int c0 = cg * 8;
    // dx = ga*(dz - m1 - xhat*m2) with xhat = (x - mu)*is  ==  A*dz + B*x + K  (three constants per channel)
    float A[8], B[8], K[8];
    if (!(count > 0.f)) count = dsums[2 * C];          // same convention: the true cross-rank count in slot 2C
    const float inv_n = 1.f / count;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float mu = mean[c0 + k], is = invstd[c0 + k];
        const float ga = gamma[c0 + k] * is;
        const float m1 = dsums[c0 + k] * inv_n, m2 = dsums[C + c0 + k] * inv_n;
        A[k] = ga;
        B[k] = -ga * m2 * is;
        K[k] = -ga * m1 - B[k] * mu;
    }
    const int64_t rstep = (int64_t)gridDim.y * rpb;
    auto masked = [&](float (&d)[8], int64_t r) {
        if (!relu) return;
        if (mask != nullptr) {
            const unsigned m = mask[r * tpr + cg];
#pragma unroll
            for (int k = 0; k < 8; ++k) d[k] = ((m >> k) & 1u) ? d[k] : 0.f;
        } else {
            float yv[8];
            load8<T>(y + r * C + c0, yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) d[k] = yv[k] > 0.f ? d[k] : 0.f;
        }
    };
    int64_t r = (int64_t)blockIdx.y * rpb + rl;
    for (; r + rstep < rows; r += 2 * rstep) {
        float d0[8], x0[8], d1[8], x1[8];
        load8<T>(dy + r * C + c0, d0);
        load8<T>(x + r * C + c0, x0);
        load8<T>(dy + (r + rstep) * C + c0, d1);
        load8<T>(x + (r + rstep) * C + c0, x1);
        masked(d0, r);
        masked(d1, r + rstep);
        if (dres != nullptr) { store8<T>(dres + r * C + c0, d0); store8<T>(dres + (r + rstep) * C + c0, d1); }
#pragma unroll
        for (int k = 0; k < 8; ++k) { d0[k] = A[k] * d0[k] + B[k] * x0[k] + K[k]; d1[k] = A[k] * d1[k] + B[k] * x1[k] + K[k]; }
        store8<T>(dx + r * C + c0, d0);
        store8<T>(dx + (r + rstep) * C + c0, d1);
    }
    for (; r < rows; r += rstep) {
        float d[8], xv[8];
        load8<T>(dy + r * C + c0, d);
        load8<T>(x + r * C + c0, xv);
        masked(d, r);
        if (dres != nullptr) store8<T>(dres + r * C + c0, d);
#pragma unroll
        for (int k = 0; k < 8; ++k) d[k] = A[k] * d[k] + B[k] * xv[k] + K[k];
        store8<T>(dx + r * C + c0, d);
    }
}

// ---------------------------------------------------------------- LayerNorm
constexpr int kLNWarps = 4;
constexpr int kLNMaxChunks = 8;  // D <= 2048

template <typename T>
__global__ void __launch_bounds__(kLNWarps * 32) layernorm_fwd_kernel(
    const T* __restrict__ x, T* __restrict__ y, int64_t rows, int D, const float* __restrict__ gamma,
    const float* __restrict__ beta, float eps, float* __restrict__ mean, float* __restrict__ rstd) {
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kLNWarps + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nchunk = D / 8;
    float v[kLNMaxChunks][8];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < kLNMaxChunks; ++j) {
        const int ch = lane + j * 32;
        if (ch < nchunk) {
            load8<T>(x + row * D + ch * 8, v[j]);
#pragma unroll
            for (int k = 0; k < 8; ++k) s += v[j][k];
        }
    }
    const float mu = warp_sum(s) / D;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < kLNMaxChunks; ++j) {
        const int ch = lane + j * 32;
        if (ch < nchunk) {
#pragma unroll
            for (int k = 0; k < 8; ++k) { const float d = v[j][k] - mu; q += d * d; }
        }
    }
    const float rs = rsqrtf(warp_sum(q) / D + eps);
    if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
#pragma unroll
    for (int j = 0; j < kLNMaxChunks; ++j) {
        const int ch = lane + j * 32;
        if (ch < nchunk) {
            float g[8], b[8], o[8];
            load8<float>(gamma + ch * 8, g);
            load8<float>(beta + ch * 8, b);
#pragma unroll
            for (int k = 0; k < 8; ++k) o[k] = (v[j][k] - mu) * rs * g[k] + b[k];
            store8<T>(y + row * D + ch * 8, o);
        }
    }
}

// Fused LayerNorm backward: ONE pass over dy and x produces dx and the parameter gradients.  Persistent
// warps walk the rows (warp per row, lane owns chunks lane, lane+32, ... of 8 columns, as in the
// forward), so each lane's share of dbeta = sum dy and dgamma = sum dy * xhat stays in registers across
// all its rows; the CTA's warps combine in shared memory and write one partial per CTA, which a small
// second kernel adds in a fixed order (deterministic, no float atomics).
template <typename T, int NCH>
__global__ void __launch_bounds__(kLNWarps * 32, 3) layernorm_bwd_fused_kernel(
    const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t rows, int D,
    const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
    float* __restrict__ scratch /* one [dbeta(D) | dgamma(D)] partial per CTA */,
    const T* __restrict__ add_in /* optional: gradient of a branch that bypassed the layer, added to dx */) {
    __shared__ float red[2 * NCH * 256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nchunk = D / 8;
    float pa[NCH][8], pb[NCH][8];
#pragma unroll
    for (int j = 0; j < NCH; ++j)
#pragma unroll
        for (int k = 0; k < 8; ++k) { pa[j][k] = 0.f; pb[j][k] = 0.f; }

    for (int64_t row = (int64_t)blockIdx.x * kLNWarps + warp; row < rows; row += (int64_t)gridDim.x * kLNWarps) {
        const float mu = mean[row], rs = rstd[row];
        Vec8<T> dv[NCH], xv[NCH], av[NCH];
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int ch = lane + j * 32;
            if (ch < nchunk) {
                dv[j] = *reinterpret_cast<const Vec8<T>*>(dy + row * D + ch * 8);
                xv[j] = *reinterpret_cast<const Vec8<T>*>(x + row * D + ch * 8);
                if (add_in != nullptr) av[j] = *reinterpret_cast<const Vec8<T>*>(add_in + row * D + ch * 8);   // in flight with the rest
            }
        }
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int ch = lane + j * 32;
            if (ch < nchunk) {
                float ga[8];
                load8<float>(gamma + ch * 8, ga);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float d = Cvt<T>::to_f(dv[j].v[k]);
                    const float xh = (Cvt<T>::to_f(xv[j].v[k]) - mu) * rs;
                    const float g = d * ga[k];
                    c1 += g * xh;
                    c2 += g;
                    pa[j][k] += d;
                    pb[j][k] += d * xh;
                }
            }
        }
        c1 = warp_sum(c1) / D;
        c2 = warp_sum(c2) / D;
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int ch = lane + j * 32;
            if (ch < nchunk) {
                float ga[8], o[8];
                load8<float>(gamma + ch * 8, ga);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const float xh = (Cvt<T>::to_f(xv[j].v[k]) - mu) * rs;
                    o[k] = rs * (Cvt<T>::to_f(dv[j].v[k]) * ga[k] - c2 - xh * c1);
                }
                if (add_in != nullptr) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) o[k] += Cvt<T>::to_f(av[j].v[k]);
                }
                store8<T>(dx + row * D + ch * 8, o);
            }
        }
    }
    // combine the CTA's warps (fixed order), publish the CTA partial
    for (int w = 0; w < kLNWarps; ++w) {
        if (warp == w) {
#pragma unroll
            for (int j = 0; j < NCH; ++j) {
                const int ch = lane + j * 32;
                if (ch < nchunk) {
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int c = ch * 8 + k;
                        red[c] = (w == 0 ? 0.f : red[c]) + pa[j][k];
                        red[D + c] = (w == 0 ? 0.f : red[D + c]) + pb[j][k];
                    }
                }
            }
        }
        __syncthreads();
    }
    float* mine = scratch + (size_t)blockIdx.x * 2 * D;
    for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) mine[i] = red[i];
}

// ---- narrow rows (D % 4 == 0, D <= 256, e.g. SwinIR-S: 60): the same two kernels on 4-element vectors (8-byte accesses
//      for 16-bit data — rows of such widths are not 16-byte aligned)
template <typename T> struct alignas(sizeof(T) * 4) Vec4 { T v[4]; };

template <typename T, int NCH>
__global__ void __launch_bounds__(kLNWarps * 32) layernorm_fwd4_kernel(
    const T* __restrict__ x, T* __restrict__ y, int64_t rows, int D, const float* __restrict__ gamma,
    const float* __restrict__ beta, float eps, float* __restrict__ mean, float* __restrict__ rstd, int pitch) {
    // pitch: elements between rows (>= D, multiple of 4); columns D .. pitch of y are written as zeros, so a row padded
    // to a tensor-core friendly width (60 -> 64) stays padded with zeros through the layer
    const int lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * kLNWarps + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int nchunk = D / 4, pchunk = pitch / 4;
    x += row * (pitch - D); y += row * (pitch - D);          // the row * D indexing below then lands on row * pitch
    float v[NCH][4];
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
        const int ch = lane + j * 32;
        if (ch < nchunk) {
            const Vec4<T> r = *reinterpret_cast<const Vec4<T>*>(x + row * D + ch * 4);
#pragma unroll
            for (int k = 0; k < 4; ++k) { v[j][k] = Cvt<T>::to_f(r.v[k]); s += v[j][k]; }
        }
    }
    const float mu = warp_sum(s) / D;
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < NCH; ++j)
        if (lane + j * 32 < nchunk) {
#pragma unroll
            for (int k = 0; k < 4; ++k) { const float d = v[j][k] - mu; q += d * d; }
        }
    const float rs = rsqrtf(warp_sum(q) / D + eps);
    if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
        const int ch = lane + j * 32;
        if (ch < nchunk) {
            const float4 g = *reinterpret_cast<const float4*>(gamma + ch * 4), b = *reinterpret_cast<const float4*>(beta + ch * 4);
            const float ga[4] = {g.x, g.y, g.z, g.w}, be[4] = {b.x, b.y, b.z, b.w};
            Vec4<T> o;
#pragma unroll
            for (int k = 0; k < 4; ++k) o.v[k] = Cvt<T>::from_f((v[j][k] - mu) * rs * ga[k] + be[k]);
            *reinterpret_cast<Vec4<T>*>(y + row * D + ch * 4) = o;
        } else if (ch < pchunk) {
            Vec4<T> o;
#pragma unroll
            for (int k = 0; k < 4; ++k) o.v[k] = Cvt<T>::from_f(0.f);
            *reinterpret_cast<Vec4<T>*>(y + row * D + ch * 4) = o;
        }
    }
}

template <typename T, int NCH>
__global__ void __launch_bounds__(kLNWarps * 32, 3) layernorm_bwd4_kernel(
    const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t rows, int D,
    const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
    float* __restrict__ scratch, const T* __restrict__ add_in, int pitch) {
    __shared__ float red[2 * NCH * 128];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nchunk = D / 4, pchunk = pitch / 4;
    const int64_t P = pitch;             // rows of dy / x / dx / add_in are `pitch` elements apart; padding of dx is zeroed
    float pa[NCH][4], pb[NCH][4], ga[NCH][4];
#pragma unroll
    for (int j = 0; j < NCH; ++j) {
        const int ch = lane + j * 32;
        const float4 g = ch < nchunk ? *reinterpret_cast<const float4*>(gamma + ch * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        ga[j][0] = g.x; ga[j][1] = g.y; ga[j][2] = g.z; ga[j][3] = g.w;
#pragma unroll
        for (int k = 0; k < 4; ++k) { pa[j][k] = 0.f; pb[j][k] = 0.f; }
    }
    for (int64_t row = (int64_t)blockIdx.x * kLNWarps + warp; row < rows; row += (int64_t)gridDim.x * kLNWarps) {
        const float mu = mean[row], rs = rstd[row];
        Vec4<T> dv[NCH], xv[NCH], av[NCH];
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int ch = lane + j * 32;
            if (ch < nchunk) {
                dv[j] = *reinterpret_cast<const Vec4<T>*>(dy + row * P + ch * 4);
                xv[j] = *reinterpret_cast<const Vec4<T>*>(x + row * P + ch * 4);
                if (add_in != nullptr) av[j] = *reinterpret_cast<const Vec4<T>*>(add_in + row * P + ch * 4);
            }
        }
        float c1 = 0.f, c2 = 0.f;
#pragma unroll
        for (int j = 0; j < NCH; ++j)
            if (lane + j * 32 < nchunk) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float d = Cvt<T>::to_f(dv[j].v[k]);
                    const float xh = (Cvt<T>::to_f(xv[j].v[k]) - mu) * rs;
                    const float g = d * ga[j][k];
                    c1 += g * xh; c2 += g;
                    pa[j][k] += d; pb[j][k] += d * xh;
                }
            }
        c1 = warp_sum(c1) / D;
        c2 = warp_sum(c2) / D;
#pragma unroll
        for (int j = 0; j < NCH; ++j) {
            const int ch = lane + j * 32;
            if (ch < nchunk) {
                Vec4<T> o;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    const float xh = (Cvt<T>::to_f(xv[j].v[k]) - mu) * rs;
                    float r = rs * (Cvt<T>::to_f(dv[j].v[k]) * ga[j][k] - c2 - xh * c1);
                    if (add_in != nullptr) r += Cvt<T>::to_f(av[j].v[k]);
                    o.v[k] = Cvt<T>::from_f(r);
                }
                *reinterpret_cast<Vec4<T>*>(dx + row * P + ch * 4) = o;
            } else if (ch < pchunk) {
                Vec4<T> o;
#pragma unroll
                for (int k = 0; k < 4; ++k) o.v[k] = Cvt<T>::from_f(0.f);
                *reinterpret_cast<Vec4<T>*>(dx + row * P + ch * 4) = o;
            }
        }
    }
    for (int w = 0; w < kLNWarps; ++w) {      // combine the CTA's warps in a fixed order
        if (warp == w) {
#pragma unroll
            for (int j = 0; j < NCH; ++j) {
                const int ch = lane + j * 32;
                if (ch < nchunk) {
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int c = ch * 4 + k;
                        red[c] = (w == 0 ? 0.f : red[c]) + pa[j][k];
                        red[D + c] = (w == 0 ? 0.f : red[D + c]) + pb[j][k];
                    }
                }
            }
        }
        __syncthreads();
    }
    float* mine = scratch + (size_t)blockIdx.x * 2 * D;
    for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) mine[i] = red[i];
}

// ---- short rows (pitch % 8 == 0, pitch <= 256; D <= pitch, any D): LPR = 8 / 16 / 32 lanes per row, 32 / LPR rows per
//      warp, every lane owns ONE 8-element chunk (a 16-byte access for 16-bit data).  A 60- or 64-wide row keeps all
//      32 lanes of a warp busy on four rows instead of 8 - 15 lanes on one.  Columns D .. pitch are written as zeros.
template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
    for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <typename T, int LPR>
__global__ void __launch_bounds__(kLNWarps * 32) layernorm_fwdR_kernel(
    const T* __restrict__ x, T* __restrict__ y, int64_t rows, int D, int pitch, const float* __restrict__ gamma,
    const float* __restrict__ beta, float eps, float* __restrict__ mean, float* __restrict__ rstd) {
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31, sub = lane / LPR, li = lane % LPR;
    const int64_t row = ((int64_t)blockIdx.x * kLNWarps + (threadIdx.x >> 5)) * RPW + sub;
    const int c0 = li * 8;
    const bool live = row < rows && c0 < pitch;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (live) load8<T>(x + row * pitch + c0, v);
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { if (c0 + k >= D) v[k] = 0.f; s += v[k]; }
    const float mu = group_sum<LPR>(s) / D;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) { const float d = c0 + k < D ? v[k] - mu : 0.f; q += d * d; }
    const float rs = rsqrtf(group_sum<LPR>(q) / D + eps);
    if (li == 0 && row < rows) { mean[row] = mu; rstd[row] = rs; }
    if (live) {
        float o[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) o[k] = c0 + k < D ? (v[k] - mu) * rs * __ldg(gamma + c0 + k) + __ldg(beta + c0 + k) : 0.f;
        store8<T>(y + row * pitch + c0, o);
    }
}

template <typename T, int LPR>
__global__ void __launch_bounds__(kLNWarps * 32, 3) layernorm_bwdR_kernel(
    const T* __restrict__ dy, const T* __restrict__ x, T* __restrict__ dx, int64_t rows, int D, int pitch,
    const float* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd,
    float* __restrict__ scratch, const T* __restrict__ add_in) {
    constexpr int RPW = 32 / LPR;
    __shared__ float red[2 * 256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane / LPR, li = lane % LPR;
    const int c0 = li * 8;
    const bool col_live = c0 < pitch;
    float ga[8], pa[8], pb[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { ga[k] = c0 + k < D ? __ldg(gamma + c0 + k) : 0.f; pa[k] = 0.f; pb[k] = 0.f; }
    const int64_t groups = (rows + RPW - 1) / RPW;
    for (int64_t grp = (int64_t)blockIdx.x * kLNWarps + warp; grp < groups; grp += (int64_t)gridDim.x * kLNWarps) {
        const int64_t row = grp * RPW + sub;
        const bool live = row < rows && col_live;
        float d[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, xv[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, av[8];
        float mu = 0.f, rs = 0.f;
        if (live) {
            load8<T>(dy + row * pitch + c0, d);
            load8<T>(x + row * pitch + c0, xv);
            if (add_in != nullptr) load8<T>(add_in + row * pitch + c0, av);
        }
        if (row < rows) { mu = mean[row]; rs = rstd[row]; }
        float c1 = 0.f, c2 = 0.f, xh[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const bool in = live && c0 + k < D;
            if (!in) d[k] = 0.f;
            xh[k] = in ? (xv[k] - mu) * rs : 0.f;
            const float g = d[k] * ga[k];
            c1 += g * xh[k]; c2 += g;
            pa[k] += d[k]; pb[k] += d[k] * xh[k];
        }
        c1 = group_sum<LPR>(c1) / D;
        c2 = group_sum<LPR>(c2) / D;
        if (live) {
            float o[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float r = c0 + k < D ? rs * (d[k] * ga[k] - c2 - xh[k] * c1) : 0.f;
                if (add_in != nullptr && c0 + k < D) r += av[k];
                o[k] = r;
            }
            store8<T>(dx + row * pitch + c0, o);
        }
    }
    // lanes that own the same columns (the warp's RPW rows), then the CTA's warps in a fixed order
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            pa[k] += __shfl_xor_sync(0xffffffffu, pa[k], o);
            pb[k] += __shfl_xor_sync(0xffffffffu, pb[k], o);
        }
    }
    for (int w = 0; w < kLNWarps; ++w) {
        if (warp == w && sub == 0) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                const int c = c0 + k;
                if (c < D) {
                    red[c] = (w == 0 ? 0.f : red[c]) + pa[k];
                    red[D + c] = (w == 0 ? 0.f : red[D + c]) + pb[k];
                }
            }
        }
        __syncthreads();
    }
    float* mine = scratch + (size_t)blockIdx.x * 2 * D;
    for (int i = threadIdx.x; i < 2 * D; i += blockDim.x) mine[i] = red[i];
}

// out[c] = sum over the CTA partials of column c, in a fixed order: 32 columns x 8 partial slices per CTA
__global__ void __launch_bounds__(256) ln_partials_reduce_kernel(const float* __restrict__ scratch, int parts, int cols,
                                                                 float* __restrict__ out) {
    __shared__ float sm[8][33];
    const int cl = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + cl;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    if (c < cols) {
        int p = sl;
        for (; p + 24 < parts; p += 32) {
            a0 += scratch[(size_t)p * cols + c];
            a1 += scratch[(size_t)(p + 8) * cols + c];
            a2 += scratch[(size_t)(p + 16) * cols + c];
            a3 += scratch[(size_t)(p + 24) * cols + c];
        }
        for (; p < parts; p += 8) a0 += scratch[(size_t)p * cols + c];
    }
    sm[sl][cl] = (a0 + a1) + (a2 + a3);
    __syncthreads();
    if (sl == 0 && c < cols) {
        float t = 0.f;
#pragma unroll
        for (int i = 0; i < 8; ++i) t += sm[i][cl];
        out[c] = t;
    }
}

}  // namespace b200dp

using namespace b200dp;

#define BN_COMMON_CHECKS(name)                                                                       \
    B200DP_REQUIRE(rows > 0 && C > 0 && C % 8 == 0, name ": C must be a positive multiple of 8");    \
    const BNGeom g = bn_geom(rows, C);                                                               \
    B200DP_REQUIRE(g.gx <= kBNMaxCols, name ": C too large (max %d)", kBNMaxCols * 256);             \
    const dim3 grid(g.gx, g.gy);                                                                     \
    cudaStream_t s = (cudaStream_t)stream;

static inline BNGeom colsum_geom(int64_t rows, int C) {
    BNGeom g = bn_geom(rows, C, 2);
    int64_t maxy = rows / ((int64_t)g.rpb * 16);          // >= 16 rows per thread: few partials
    if (maxy < 1) maxy = 1;
    if (g.gy > maxy) g.gy = (int)maxy;
    return g;
}

extern "C" {

int64_t b200dp_column_sums_scratch_floats(int64_t rows, int C) {
    if (rows <= 0 || C <= 0 || C % 8) return 0;
    return (int64_t)colsum_geom(rows, C).gy * C;
}

// out[C] (fp32) = column sums of the 16-bit / fp32 matrix x[rows][C] (C % 8 == 0); scratch: b200dp_column_sums_scratch_floats.
int b200dp_column_sums(const void* x, int64_t rows, int C, b200dp_dtype dtype, float* scratch, float* out, void* stream) {
    B200DP_REQUIRE(x && scratch && out, "column_sums: null argument");
    B200DP_REQUIRE(rows > 0 && C > 0 && C % 8 == 0, "column_sums: C must be a positive multiple of 8");
    const BNGeom g = colsum_geom(rows, C);
    cudaStream_t s = (cudaStream_t)stream;
    B200DP_DISPATCH_DTYPE(dtype, T, {
        colsum_kernel<T><<<dim3(g.gx, g.gy), kBNThreads, 0, s>>>((const T*)x, rows, C, g.tpr, g.cgb, g.rpb, scratch);
    });
    ln_partials_reduce_kernel<<<(C + 31) / 32, 256, 0, s>>>(scratch, g.gy, C, out);
    B200DP_CHECK_LAUNCH();
    return 0;
}

int64_t b200dp_bn_scratch_floats(int64_t rows, int C) {
    if (rows <= 0 || C <= 0 || C % 8) return 0;
    const BNGeom g = bn_geom(rows, C);
    return (int64_t)g.gy * 2 * C;
}

int b200dp_bn_reduce(const void* x, int64_t rows, int C, b200dp_dtype dtype, float* scratch,
                     unsigned int* counter, float* sums, void* stream) {
    B200DP_REQUIRE(x && scratch && counter && sums, "bn_reduce: null argument");
    BN_COMMON_CHECKS("bn_reduce")
    B200DP_DISPATCH_HALF(dtype, T, {
        bn_reduce_kernel<T><<<grid, kBNThreads, 0, s>>>((const T*)x, rows, C, g.tpr, g.cgb, g.rpb, scratch,
                                                         counter, sums);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_bn_apply_tracked(const void* x, const void* residual, void* y, int64_t rows, int C, b200dp_dtype dtype,
                            const float* sums, float count, const float* gamma, const float* beta, float eps,
                            float momentum, float* running_mean, float* running_var, float* mean_out,
                            float* invstd_out, int relu, unsigned char* relu_mask, long long* num_batches_tracked,
                            void* stream) {
    B200DP_REQUIRE(x && y && sums && gamma && beta && mean_out && invstd_out, "bn_apply: null argument");
    BN_COMMON_CHECKS("bn_apply")
    B200DP_DISPATCH_HALF(dtype, T, {
        bn_apply_kernel<T, true><<<dim3(bn_geom(rows, C, 3).gx, bn_geom(rows, C, 3).gy), kBNThreads, 0, s>>>((const T*)x, (const T*)residual, (T*)y, rows, C, g.tpr,
                                                             g.cgb, g.rpb, sums, count, gamma, beta, eps, momentum,
                                                             running_mean, running_var, mean_out, invstd_out, relu, relu_mask,
                                                             num_batches_tracked);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_bn_apply(const void* x, const void* residual, void* y, int64_t rows, int C, b200dp_dtype dtype,
                    const float* sums, float count, const float* gamma, const float* beta, float eps,
                    float momentum, float* running_mean, float* running_var, float* mean_out,
                    float* invstd_out, int relu, unsigned char* relu_mask, void* stream) {
    return b200dp_bn_apply_tracked(x, residual, y, rows, C, dtype, sums, count, gamma, beta, eps, momentum, running_mean,
                                   running_var, mean_out, invstd_out, relu, relu_mask, nullptr, stream);
}

int b200dp_bn_infer(const void* x, const void* residual, void* y, int64_t rows, int C, b200dp_dtype dtype,
                    const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                    float eps, int relu, void* stream) {
    B200DP_REQUIRE(x && y && gamma && beta && running_mean && running_var, "bn_infer: null argument");
    BN_COMMON_CHECKS("bn_infer")
    B200DP_DISPATCH_HALF(dtype, T, {
        bn_apply_kernel<T, false><<<grid, kBNThreads, 0, s>>>(
            (const T*)x, (const T*)residual, (T*)y, rows, C, g.tpr, g.cgb, g.rpb, nullptr, 1.f, gamma, beta, eps,
            0.f, const_cast<float*>(running_mean), const_cast<float*>(running_var), nullptr, nullptr, relu, nullptr);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_bn_bwd_reduce(const void* dy, const void* y, const void* x, int64_t rows, int C, b200dp_dtype dtype,
                         const float* mean, const float* invstd, int relu, float* scratch,
                         unsigned int* counter, float* dsums, const unsigned char* relu_mask, void* stream) {
    B200DP_REQUIRE(dy && x && mean && invstd && scratch && counter && dsums, "bn_bwd_reduce: null argument");
    B200DP_REQUIRE(!relu || y || relu_mask, "bn_bwd_reduce: relu needs the forward output y or the relu mask");
    BN_COMMON_CHECKS("bn_bwd_reduce")
    // three CTAs per SM with eight 16-byte loads in flight per thread (the kernel is register-heavier than its siblings);
    // fewer row slices than b200dp_bn_scratch_floats provides for, never more
    const BNGeom g3 = bn_geom(rows, C, 3);
    const dim3 grid3(g3.gx, g3.gy);
    B200DP_DISPATCH_HALF(dtype, T, {
        bn_bwd_reduce_kernel<T><<<grid3, kBNThreads, 0, s>>>((const T*)dy, (const T*)y, (const T*)x, rows, C, g3.tpr,
                                                             g3.cgb, g3.rpb, mean, invstd, relu, scratch, counter,
                                                             dsums, relu_mask);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_bn_bwd_apply(const void* dy, const void* y, const void* x, void* dx, void* dres, int64_t rows, int C,
                        b200dp_dtype dtype, const float* mean, const float* invstd, const float* gamma,
                        const float* dsums, float count, int relu, const unsigned char* relu_mask, void* stream) {
    B200DP_REQUIRE(dy && x && dx && mean && invstd && gamma && dsums, "bn_bwd_apply: null argument");
    B200DP_REQUIRE(!relu || y || relu_mask, "bn_bwd_apply: relu needs the forward output y or the relu mask");
    BN_COMMON_CHECKS("bn_bwd_apply")
    B200DP_DISPATCH_HALF(dtype, T, {
        bn_bwd_apply_kernel<T><<<grid, kBNThreads, 0, s>>>((const T*)dy, (const T*)y, (const T*)x, (T*)dx, (T*)dres,
                                                           rows, C, g.tpr, g.cgb, g.rpb, mean, invstd, gamma, dsums,
                                                           count, relu, relu_mask);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

// persistent grid: never more CTAs than are resident at once (a second, partial wave would double the time);
// 3 per SM is what the heaviest instantiation (D = 768 with the bypass gradient) fits
constexpr int kLNBwdCtasPerSM = 3;
static unsigned ln_bwd_grid(int64_t rows) {
    const int64_t want = (rows + kLNWarps - 1) / kLNWarps;
    const int64_t cap = (int64_t)B200DP_SMS * kLNBwdCtasPerSM;
    return (unsigned)(want < cap ? want : cap);
}

// one [dbeta | dgamma] partial per CTA of the fused backward
int64_t b200dp_ln_scratch_floats(int64_t rows, int D) { return rows <= 0 || D <= 0 ? 0 : (int64_t)ln_bwd_grid(rows) * 2 * D; }

static int layernorm_fwd4(const void* x, void* y, int64_t rows, int D, int pitch, b200dp_dtype dtype, const float* gamma,
                          const float* beta, float eps, float* mean, float* rstd, void* stream) {
    if (pitch % 8 == 0) {                // several rows per warp, 16-byte accesses
#define LNR(LPR_)                                                                                                      \
        {                                                                                                              \
            const unsigned nb = (unsigned)((rows + kLNWarps * (32 / LPR_) - 1) / (kLNWarps * (32 / LPR_)));            \
            B200DP_DISPATCH_DTYPE(dtype, T, { layernorm_fwdR_kernel<T, LPR_><<<nb, kLNWarps * 32, 0, (cudaStream_t)stream>>>( \
                (const T*)x, (T*)y, rows, D, pitch, gamma, beta, eps, mean, rstd); });                                 \
        }
        if (pitch <= 64) LNR(8) else if (pitch <= 128) LNR(16) else LNR(32)
#undef LNR
        B200DP_CHECK_LAUNCH();
        return 0;
    }
    const unsigned blocks = (unsigned)((rows + kLNWarps - 1) / kLNWarps);
    if (pitch <= 128) {
        B200DP_DISPATCH_DTYPE(dtype, T, { layernorm_fwd4_kernel<T, 1><<<blocks, kLNWarps * 32, 0, (cudaStream_t)stream>>>(
            (const T*)x, (T*)y, rows, D, gamma, beta, eps, mean, rstd, pitch); });
    } else {
        B200DP_DISPATCH_DTYPE(dtype, T, { layernorm_fwd4_kernel<T, 2><<<blocks, kLNWarps * 32, 0, (cudaStream_t)stream>>>(
            (const T*)x, (T*)y, rows, D, gamma, beta, eps, mean, rstd, pitch); });
    }
    B200DP_CHECK_LAUNCH();
    return 0;
}

// LayerNorm over the first D elements of rows that are `pitch` elements apart (D % 4 == 0, D <= pitch <= 256,
// pitch % 4 == 0); y has the same pitch and its columns D .. pitch come out as zeros.
int b200dp_layernorm_fwd_pitched(const void* x, void* y, int64_t rows, int D, int pitch, b200dp_dtype dtype,
                                 const float* gamma, const float* beta, float eps, float* mean, float* rstd, void* stream) {
    B200DP_REQUIRE(x && y && gamma && beta && mean && rstd, "layernorm_fwd_pitched: null argument");
    B200DP_REQUIRE(rows > 0 && D > 0 && D % 4 == 0 && pitch % 4 == 0 && pitch >= D && pitch <= 256,
                   "layernorm_fwd_pitched: D=%d pitch=%d (multiples of 4, D <= pitch <= 256)", D, pitch);
    return layernorm_fwd4(x, y, rows, D, pitch, dtype, gamma, beta, eps, mean, rstd, stream);
}

int b200dp_layernorm_fwd(const void* x, void* y, int64_t rows, int D, b200dp_dtype dtype, const float* gamma,
                         const float* beta, float eps, float* mean, float* rstd, void* stream) {
    B200DP_REQUIRE(x && y && gamma && beta && mean && rstd, "layernorm_fwd: null argument");
    const unsigned blocks = (unsigned)((rows + kLNWarps - 1) / kLNWarps);
    if (rows > 0 && D % 4 == 0 && D <= 256)                      // short rows: several per warp / 4-element vectors
        return layernorm_fwd4(x, y, rows, D, D, dtype, gamma, beta, eps, mean, rstd, stream);
    B200DP_REQUIRE(rows > 0 && D % 8 == 0 && D <= kLNMaxChunks * 256, "layernorm_fwd: D must be a multiple of 8 (or of 4 up to 256), <= %d",
                   kLNMaxChunks * 256);
    B200DP_DISPATCH_DTYPE(dtype, T, {
        layernorm_fwd_kernel<T><<<blocks, kLNWarps * 32, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, rows, D, gamma,
                                                                                    beta, eps, mean, rstd);
    });
    B200DP_CHECK_LAUNCH();
    return 0;
}

static int layernorm_bwd_impl(const void* dy, const void* x, void* dx, int64_t rows, int D, b200dp_dtype dtype,
                              const float* gamma, const float* mean, const float* rstd, float* scratch,
                              unsigned int* counter, float* dgamma, float* dbeta, const void* add_in, void* stream,
                              int pitch = 0) {
    B200DP_REQUIRE(dy && x && dx && gamma && mean && rstd && scratch && counter && dgamma && dbeta,
                   "layernorm_bwd: null argument");
    B200DP_REQUIRE(dgamma == dbeta + D, "layernorm_bwd: dbeta and dgamma must be one buffer [dbeta(D) | dgamma(D)]");
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned grid = ln_bwd_grid(rows);
    if (rows > 0 && (pitch > 0 || (D % 4 == 0 && D <= 256))) {      // short rows: several per warp / 4-element vectors
        if (pitch <= 0) pitch = D;
        if (pitch % 8 == 0) {
#define LNR(LPR_)                                                                                                      \
            {                                                                                                          \
                const int64_t want = ((rows + (32 / LPR_) - 1) / (32 / LPR_) + kLNWarps - 1) / kLNWarps;               \
                const unsigned gr = (unsigned)(want < (int64_t)grid ? want : (int64_t)grid);                           \
                B200DP_DISPATCH_DTYPE(dtype, T, { layernorm_bwdR_kernel<T, LPR_><<<gr, kLNWarps * 32, 0, s>>>(         \
                    (const T*)dy, (const T*)x, (T*)dx, rows, D, pitch, gamma, mean, rstd, scratch, (const T*)add_in); }); \
                ln_partials_reduce_kernel<<<(2 * D + 31) / 32, 256, 0, s>>>(scratch, (int)gr, 2 * D, dbeta);           \
            }
            if (pitch <= 64) LNR(8) else if (pitch <= 128) LNR(16) else LNR(32)
#undef LNR
            B200DP_CHECK_LAUNCH();
            return 0;
        }
        if (pitch <= 128) {
            B200DP_DISPATCH_DTYPE(dtype, T, { layernorm_bwd4_kernel<T, 1><<<grid, kLNWarps * 32, 0, s>>>(
                (const T*)dy, (const T*)x, (T*)dx, rows, D, gamma, mean, rstd, scratch, (const T*)add_in, pitch); });
        } else {
            B200DP_DISPATCH_DTYPE(dtype, T, { layernorm_bwd4_kernel<T, 2><<<grid, kLNWarps * 32, 0, s>>>(
                (const T*)dy, (const T*)x, (T*)dx, rows, D, gamma, mean, rstd, scratch, (const T*)add_in, pitch); });
        }
        ln_partials_reduce_kernel<<<(2 * D + 31) / 32, 256, 0, s>>>(scratch, (int)grid, 2 * D, dbeta);
        B200DP_CHECK_LAUNCH();
        return 0;
    }
    B200DP_REQUIRE(rows > 0 && D % 8 == 0 && D <= kLNMaxChunks * 256, "layernorm_bwd: D must be a multiple of 8 (or of 4 up to 256), <= %d",
                   kLNMaxChunks * 256);
    const int nch = (D + 255) / 256;
#define LNB(N_)                                                                                                     \
    B200DP_DISPATCH_DTYPE(dtype, T, {                                                                               \
        layernorm_bwd_fused_kernel<T, N_><<<grid, kLNWarps * 32, 0, s>>>((const T*)dy, (const T*)x, (T*)dx, rows, D, \
                                                                         gamma, mean, rstd, scratch, (const T*)add_in); \
    })
    if (nch <= 1) LNB(1) else if (nch <= 2) LNB(2) else if (nch <= 3) LNB(3) else if (nch <= 4) LNB(4)
    else if (nch <= 6) LNB(6) else LNB(8)
#undef LNB
    ln_partials_reduce_kernel<<<(2 * D + 31) / 32, 256, 0, s>>>(scratch, (int)grid, 2 * D, dbeta);
    B200DP_CHECK_LAUNCH();
    return 0;
}

int b200dp_layernorm_bwd(const void* dy, const void* x, void* dx, int64_t rows, int D, b200dp_dtype dtype,
                         const float* gamma, const float* mean, const float* rstd, float* scratch,
                         unsigned int* counter, float* dgamma, float* dbeta, void* stream) {
    return layernorm_bwd_impl(dy, x, dx, rows, D, dtype, gamma, mean, rstd, scratch, counter, dgamma, dbeta, nullptr, stream);
}

// Same, plus dx += add_in: the gradient of a branch that bypassed the LayerNorm (the residual stream of a
// transformer block) is folded into the input gradient instead of a separate add kernel.
int b200dp_layernorm_bwd_add(const void* dy, const void* x, void* dx, int64_t rows, int D, b200dp_dtype dtype,
                             const float* gamma, const float* mean, const float* rstd, float* scratch,
                             unsigned int* counter, float* dgamma, float* dbeta, const void* add_in, void* stream) {
    return layernorm_bwd_impl(dy, x, dx, rows, D, dtype, gamma, mean, rstd, scratch, counter, dgamma, dbeta, add_in, stream);
}

// Backward of b200dp_layernorm_fwd_pitched: dy / x / dx (/ add_in, may be NULL) all have rows `pitch` elements apart;
// columns D .. pitch of dx are written as zeros.
int b200dp_layernorm_bwd_pitched(const void* dy, const void* x, void* dx, int64_t rows, int D, int pitch, b200dp_dtype dtype,
                                 const float* gamma, const float* mean, const float* rstd, float* scratch,
                                 unsigned int* counter, float* dgamma, float* dbeta, const void* add_in, void* stream) {
    B200DP_REQUIRE(rows > 0 && D > 0 && D % 4 == 0 && pitch % 4 == 0 && pitch >= D && pitch <= 256,
                   "layernorm_bwd_pitched: D=%d pitch=%d (multiples of 4, D <= pitch <= 256)", D, pitch);
    return layernorm_bwd_impl(dy, x, dx, rows, D, dtype, gamma, mean, rstd, scratch, counter, dgamma, dbeta, add_in, stream, pitch);
}

}  // extern "C"
