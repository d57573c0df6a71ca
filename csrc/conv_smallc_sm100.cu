// First-layer convolutions (RGB stems: Cin <= 4) on the 5th-generation tensor cores (W7 / K15) — with NO im2col.
//
// The implicit-GEMM kernels of conv_sm100.cu let TMA fetch 128-byte channel rows, which needs Cin % 64 == 0; a
// 3-channel image has 6-byte pixels.  Here the input rows a tile needs are staged in shared memory with the
// channels padded to Cp = 8 / stride (4 for the stride-2 ResNet stem, 8 for stride-1 layers), so that consecutive
// OUTPUT pixels start exactly 16 bytes apart.  The window of one output pixel in one filter row is then a run of
// S * Cp contiguous elements, and the windows of consecutive pixels are the SAME buffer shifted by 16 bytes: the
// un-swizzled UMMA shared-memory layout (8 x 16-byte core matrices, rows 16 bytes apart) describes that overlapping
// "Toeplitz" matrix directly with LBO = 16 and SBO = 128 — the tensor core reads the staged image rows as the
// im2col matrix without it ever being materialised.
//
//   forward   per tile of 128 output pixels of one output row:  Y[pix][co] = sum_r  X_r[pix][k] W_r[co][k]^T
//             (R chains of K = S*Cp, one per filter row, all accumulating into the same 64 TMEM columns)
//   wgrad     per tile:  D[co][r][k] += sum_pix dY[pix][co] X_r[pix][k]   (dY MN-major as loaded, the Toeplitz rows
//             as the MN-major B operand, N = K of one filter row); D stays in TMEM over all tiles of the CTA and is
//             added to the fp32 gradient once at the end
// The input gradient of a first layer is rarely needed; when it is (perceptual loss on an RGB output) the front end
// pads Cin to 8 and uses the regular dgrad kernel.
//
// Reference: the first convolution of the CNN models — ResNet stem 7x7/2 (BASELINE configs 1-3), the SR net's 5x5
// (Fairscale-DDP.py:74), SwinIR's conv_first and the VGG layer of feat_loss (Stoke-DDP.py:206-224); cuDNN's
// first_layer kernels there.
#include "tc_common.cuh"

namespace b200dp {
namespace convsc {

using namespace b200dp::tc;

struct Geom {
    int N, H, W, Cin, Cout, R, S, stride, pad, P, Q;
    int Cp;         // channels as staged: 8 / stride
    int Kr;         // K of one filter row, S * Cp rounded up to 16
    int rowb;       // bytes of one staged input row buffer (multiple of 128)
    int npix;       // input pixels staged per row
    int qblocks;    // 128-pixel tiles per output row
    long long tiles;
};

// un-swizzled UMMA descriptor (layout type 0): core matrices of 8 rows x 16 bytes.  K-major operand: SBO = distance
// between 8-row (M/N) groups, LBO = distance between 16-byte K chunks.  MN-major operand: the roles are the other way
// round from the swizzled modes — SBO = distance between 8-element MN chunks, LBO = distance between 8-row K groups.
__device__ __forceinline__ uint64_t make_desc_ns(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}

__device__ __forceinline__ void zero_smem(uint8_t* p, int bytes) {
    for (int i = threadIdx.x * 16; i < bytes; i += blockDim.x * 16) *reinterpret_cast<uint4*>(p + i) = make_uint4(0, 0, 0, 0);
}

struct TileId { int n, p, q0; };
__device__ __forceinline__ TileId tile_id(const Geom& g, long long t) {
    TileId o;
    o.q0 = (int)(t % g.qblocks) * 128;
    const long long np = t / g.qblocks;
    o.p = (int)(np % g.P);
    o.n = (int)(np / g.P);
    return o;
}

// Input rows ih = p*stride - pad + r (r < R), pixels iw = q0*stride - pad + j (j < npix), channels padded to Cp, zeros
// outside the image: row r at sIn + r * rowb, pixel j at + j * Cp * 2.
template <typename T>
__device__ __forceinline__ void stage_input(uint8_t* sIn, const T* __restrict__ x, const Geom& g, const TileId& t) {
    const unsigned short* xs = reinterpret_cast<const unsigned short*>(x);
    const uint32_t s0 = smem_u32(sIn);
    const int total = g.R * g.npix;
    const int iw0 = t.q0 * g.stride - g.pad;
    // four items per iteration: every load of the group is issued before the first store (the loop is latency-bound)
    for (int i0 = threadIdx.x; i0 < total; i0 += 4 * blockDim.x) {
        unsigned short v[4][4];
        uint32_t addr[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * blockDim.x;
            const int r = i / g.npix, j = i - r * g.npix;
            const int ih = t.p * g.stride - g.pad + r, iw = iw0 + j;
            const bool ok = i < total && ih >= 0 && ih < g.H && iw >= 0 && iw < g.W;
            const unsigned short* src = xs + (((long long)t.n * g.H + (ok ? ih : 0)) * g.W + (ok ? iw : 0)) * g.Cin;
#pragma unroll
            for (int c = 0; c < 4; ++c) v[u][c] = (ok && c < g.Cin) ? __ldg(src + c) : (unsigned short)0;
            addr[u] = i < total ? s0 + r * g.rowb + j * g.Cp * 2 : 0xffffffffu;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (addr[u] == 0xffffffffu) continue;
            const uint32_t lo = (uint32_t)v[u][0] | ((uint32_t)v[u][1] << 16), hi = (uint32_t)v[u][2] | ((uint32_t)v[u][3] << 16);
            if (g.Cp == 4) asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(addr[u]), "r"(lo), "r"(hi) : "memory");
            else asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr[u]), "r"(lo), "r"(hi), "r"(0u), "r"(0u) : "memory");
        }
    }
}

constexpr int kMaxR = 7;
constexpr int kRowBufBytes = kMaxR * 2560;                 // 7 staged rows of up to 2.5 KB
constexpr int kFwdSmem = kRowBufBytes + kMaxR * 8192 /*W_r: [64 co][64 k] each*/ + 1024 + 64;

template <typename T>
__global__ void __launch_bounds__(256, 2) convsc_fwd_kernel(const T* __restrict__ x, const T* __restrict__ w,
                                                            const float* __restrict__ bias, T* __restrict__ y,
                                                            const Geom g, int is_bf16) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sW = smem;                              // R x [64 co][64 k], K-major, 128B swizzle
    uint8_t* sIn = smem + kMaxR * 8192;              // R staged input rows
    uint64_t* bar = reinterpret_cast<uint64_t*>(sIn + kRowBufBytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    zero_smem(sW, g.R * 8192);
    zero_smem(sIn, kRowBufBytes);
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (warp == 2) tmem_alloc(tmem_slot, 64);
    __syncthreads();
    {   // filter [Cout][R][S][Cin] -> per filter row a K-major tile with k = s * Cp + c (zeros elsewhere)
        const unsigned short* ws = reinterpret_cast<const unsigned short*>(w);
        const uint32_t sw = smem_u32(sW);
        const int per_co = g.R * g.S * g.Cin;
        for (int idx = threadIdx.x; idx < g.Cout * per_co; idx += blockDim.x) {
            const int co = idx / per_co, rem = idx - co * per_co;
            const int r = rem / (g.S * g.Cin), sc = rem - r * (g.S * g.Cin);
            const int s = sc / g.Cin, c = sc - s * g.Cin;
            const int k = s * g.Cp + c;
            const unsigned short v = __ldg(ws + idx);
            const uint32_t a = sw + r * 8192 + co * 128 + (((k >> 3) ^ (co & 7)) << 4) + (k & 7) * 2;
            asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"(v) : "memory");
        }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int ksteps = g.Kr >> 4;
    const int row = (warp & 3) * 32 + lane, half = warp >> 2;
    uint32_t phase = 0;
    for (long long tile = blockIdx.x; tile < g.tiles; tile += gridDim.x) {
        const TileId t = tile_id(g, tile);
        stage_input<T>(sIn, x, g, t);
        fence_async_smem();
        __syncthreads();
        if (warp == 1) {
            if (lane == 0) {
                tc_fence_after();
                const uint32_t idesc = make_idesc(is_bf16, 0, 0, 128, 64);
                const uint32_t ia = smem_u32(sIn), wa = smem_u32(sW);
                bool acc = false;
                for (int r = 0; r < g.R; ++r)
                    for (int k = 0; k < ksteps; ++k) {
                        umma_f16(tmem_base, make_desc_ns(ia + r * g.rowb + k * 32, 16, 128),
                                 make_desc(wa + r * 8192 + k * 32, 16, 1024), idesc, acc ? 1u : 0u);
                        acc = true;
                    }
                umma_commit(bar);
            }
            __syncwarp();
        }
        mbar_wait(bar, phase);
        phase ^= 1;
        tc_fence_after();
        {
            uint32_t v[32];
            tmem_ld_32x32b_x32(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + half * 32, v);
            tmem_ld_wait();
            const int q = t.q0 + row;
            if (q < g.Q) {
                T* dst = y + (((long long)t.n * g.P + t.p) * g.Q + q) * g.Cout + half * 32;
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    const int c = half * 32 + c8 * 8;
                    if (c < g.Cout) {
                        float o[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) o[e] = __uint_as_float(v[c8 * 8 + e]) + (bias ? __ldg(bias + c + e) : 0.f);
                        store8<T>(dst + c8 * 8, o);
                    }
                }
            }
        }
        tc_fence_before();
        __syncthreads();           // the staged rows and the accumulator are free again
    }
    if (warp == 2) tmem_dealloc(tmem_base, 64);
}

constexpr int kWgSmem = kRowBufBytes + 2 * 16384 /*dY tile + zero block*/ + 1024 + 64;

template <typename T>
__global__ void __launch_bounds__(256, 2) convsc_wgrad_kernel(const T* __restrict__ x, const T* __restrict__ dy,
                                                              float* __restrict__ dw, const Geom g, int is_bf16) {
    const uint32_t tmem_cols = g.R * g.Kr <= 128 ? 128u : (g.R * g.Kr <= 256 ? 256u : 512u);
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sD = smem;                              // [128 pixels][64 co], then 16 KB of zeros (the second 64-wide M block)
    uint8_t* sIn = smem + 2 * 16384;
    uint64_t* bar = reinterpret_cast<uint64_t*>(sIn + kRowBufBytes);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar + 1);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    zero_smem(sD, 2 * 16384);
    zero_smem(sIn, kRowBufBytes);
    if (threadIdx.x == 0) { mbar_init(bar, 1); fence_barrier_init(); }
    if (warp == 2) tmem_alloc(tmem_slot, tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    uint32_t phase = 0;
    bool first = true;
    const uint32_t sd = smem_u32(sD);
    for (long long tile = blockIdx.x; tile < g.tiles; tile += gridDim.x) {
        const TileId t = tile_id(g, tile);
        stage_input<T>(sIn, x, g, t);
        // dY rows of the tile: 128 x Cout, 16-byte pieces, coalesced; pixels past the row end are zeros
        for (int idx = threadIdx.x; idx < 128 * 8; idx += blockDim.x) {
            const int rw = idx >> 3, pc = idx & 7;
            const int q = t.q0 + rw;
            uint4 v = make_uint4(0, 0, 0, 0);
            if (q < g.Q && pc * 8 < g.Cout)
                v = __ldg(reinterpret_cast<const uint4*>(dy + (((long long)t.n * g.P + t.p) * g.Q + q) * g.Cout + pc * 8));
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sd + rw * 128 + ((pc ^ (rw & 7)) << 4)), "r"(v.x), "r"(v.y),
                         "r"(v.z), "r"(v.w) : "memory");
        }
        fence_async_smem();
        __syncthreads();
        if (warp == 1) {
            if (lane == 0) {
                tc_fence_after();
                const uint32_t idesc = make_idesc(is_bf16, 1, 1, 128, g.Kr);
                const uint32_t ia = smem_u32(sIn);
                for (int r = 0; r < g.R; ++r)
                    for (int k = 0; k < 8; ++k)        // 128 pixels = 8 x 16
                        umma_f16(tmem_base + r * g.Kr, make_desc(sd + k * 2048, 16384, 1024),
                                 make_desc_ns(ia + r * g.rowb + k * 256, 128, 16), idesc, (first && k == 0) ? 0u : 1u);
                umma_commit(bar);
            }
            __syncwarp();
        }
        first = false;
        mbar_wait(bar, phase);          // the MMAs have read the tiles: they may be restaged
        phase ^= 1;
        __syncthreads();
    }
    if (!first && (warp & 3) < 2) {
        // lanes 0..63 hold the output channels; columns r * Kr + s * Cp + c hold the filter taps
        tc_fence_after();
        const int co = (warp & 3) * 32 + lane, half = warp >> 2;
        const int cols = g.R * g.Kr;
        for (int c0 = half * 32; c0 < cols; c0 += 64) {
            uint32_t v[32];
            tmem_ld_32x32b_x32(tmem_base + ((uint32_t)((warp & 3) * 32) << 16) + c0, v);
            tmem_ld_wait();
            if (co < g.Cout) {
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int col = c0 + j;
                    const int r = col / g.Kr, k = col - r * g.Kr;
                    const int s = k / g.Cp, c = k - s * g.Cp;
                    if (col < cols && s < g.S && c < g.Cin)
                        atomicAdd(dw + (((long long)co * g.R + r) * g.S + s) * g.Cin + c, __uint_as_float(v[j]));
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, tmem_cols);
}

static int make_geom(Geom* g, int N, int H, int W, int Cin, int Cout, int R, int S, int stride, int pad) {
    B200DP_REQUIRE(N > 0 && H > 0 && W > 0 && Cin >= 1 && Cin <= 4 && R >= 1 && R <= kMaxR && S >= 1,
                   "conv2d_smallc: needs Cin <= 4 and R <= %d (got Cin=%d R=%d S=%d)", kMaxR, Cin, R, S);
    B200DP_REQUIRE(stride == 1 || stride == 2, "conv2d_smallc: stride must be 1 or 2 (got %d)", stride);
    B200DP_REQUIRE(Cout >= 8 && Cout <= 64 && Cout % 8 == 0, "conv2d_smallc: Cout must be a multiple of 8 in [8, 64] (got %d)", Cout);
    B200DP_REQUIRE(pad >= 0 && H + 2 * pad >= R && W + 2 * pad >= S, "conv2d_smallc: bad geometry");
    g->N = N; g->H = H; g->W = W; g->Cin = Cin; g->Cout = Cout; g->R = R; g->S = S; g->stride = stride; g->pad = pad;
    g->P = (H + 2 * pad - R) / stride + 1;
    g->Q = (W + 2 * pad - S) / stride + 1;
    g->Cp = 8 / stride;
    g->Kr = (S * g->Cp + 15) / 16 * 16;
    B200DP_REQUIRE(g->Kr <= 64 && R * g->Kr <= 512, "conv2d_smallc: filter too wide (S=%d gives K=%d per filter row)", S, g->Kr);
    // pixel j of the staged row is read by output pixel m at element offset (m * stride + s): 127 * stride + Kr / Cp pixels
    g->npix = 127 * stride + g->Kr / g->Cp + 1;
    g->rowb = (g->npix * g->Cp * 2 + 127) / 128 * 128;
    B200DP_REQUIRE(g->rowb <= 2560, "conv2d_smallc: staged row does not fit (%d bytes)", g->rowb);
    g->qblocks = (g->Q + 127) / 128;
    g->tiles = (long long)N * g->P * g->qblocks;
    return 0;
}

}  // namespace convsc
}  // namespace b200dp

using namespace b200dp;
using namespace b200dp::convsc;

extern "C" int b200dp_conv2d_smallc_supported(int Cin, int Cout, int R, int S, int stride) {
    if (!(Cin >= 1 && Cin <= 4 && R >= 1 && R <= kMaxR && S >= 1 && (stride == 1 || stride == 2))) return 0;
    const int kr = (S * (8 / stride) + 15) / 16 * 16;
    return kr <= 64 && R * kr <= 512 && Cout >= 8 && Cout <= 64 && Cout % 8 == 0;
}

// y[N][P][Q][Cout] = conv(x[N][H][W][Cin], w[Cout][R][S][Cin]) (+ bias), 16-bit NHWC tensors.
extern "C" int b200dp_conv2d_smallc_fwd(const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin,
                                        int Cout, int R, int S, int stride, int pad, b200dp_dtype dtype, void* stream) {
    B200DP_REQUIRE(x && w && y, "conv2d_smallc_fwd: null pointer");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "conv2d_smallc_fwd: fp16 / bf16 only");
    B200DP_REQUIRE(((uintptr_t)y & 15) == 0, "conv2d_smallc_fwd: y must be 16-byte aligned");
    Geom g;
    int rc = make_geom(&g, N, H, W, Cin, Cout, R, S, stride, pad);
    if (rc) return rc;
    const int grid = (int)(g.tiles < 2 * B200DP_SMS ? g.tiles : 2 * B200DP_SMS);      // two CTAs per SM
    cudaStream_t s = (cudaStream_t)stream;
    const int is_bf16 = dtype == B200DP_BF16;
#define LAUNCH(TT)                                                                                              \
    {                                                                                                           \
        auto kern = convsc_fwd_kernel<TT>;                                                                      \
        static bool configured = false;                                                                         \
        if (!configured) {                                                                                      \
            B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));     \
            configured = true;                                                                                  \
        }                                                                                                       \
        kern<<<grid, 256, kFwdSmem, s>>>((const TT*)x, (const TT*)w, bias, (TT*)y, g, is_bf16);                 \
    }
    if (is_bf16) LAUNCH(__nv_bfloat16) else LAUNCH(__half)
#undef LAUNCH
    B200DP_CHECK_LAUNCH();
    return 0;
}

// dw[Cout][R][S][Cin] (fp32) += sum over pixels of dy x window(x).  The caller zeroes dw (atomic accumulation).
extern "C" int b200dp_conv2d_smallc_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R,
                                          int S, int stride, int pad, b200dp_dtype dtype, void* stream) {
    B200DP_REQUIRE(x && dy && dw, "conv2d_smallc_wgrad: null pointer");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "conv2d_smallc_wgrad: fp16 / bf16 only");
    B200DP_REQUIRE(((uintptr_t)dy & 15) == 0, "conv2d_smallc_wgrad: dy must be 16-byte aligned");
    Geom g;
    int rc = make_geom(&g, N, H, W, Cin, Cout, R, S, stride, pad);
    if (rc) return rc;
    const int per_sm = g.R * g.Kr <= 256 ? 2 : 1;          // TMEM columns decide how many CTAs share an SM
    const int grid = (int)(g.tiles < per_sm * B200DP_SMS ? g.tiles : per_sm * B200DP_SMS);
    cudaStream_t s = (cudaStream_t)stream;
    const int is_bf16 = dtype == B200DP_BF16;
#define LAUNCH(TT)                                                                                              \
    {                                                                                                           \
        auto kern = convsc_wgrad_kernel<TT>;                                                                    \
        static bool configured = false;                                                                         \
        if (!configured) {                                                                                      \
            B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kWgSmem));      \
            configured = true;                                                                                  \
        }                                                                                                       \
        kern<<<grid, 256, kWgSmem, s>>>((const TT*)x, (const TT*)dy, dw, g, is_bf16);                           \
    }
    if (is_bf16) LAUNCH(__nv_bfloat16) else LAUNCH(__half)
#undef LAUNCH
    B200DP_CHECK_LAUNCH();
    return 0;
}
