// Attention core on the 5th-generation tensor cores (K17): softmax(Q K^T * scale + bias) V, forward and backward,
// for the two shapes the transformer models use — ViT (197 tokens, head dim 64) and Swin / SwinIR windows
// (<= 64 tokens, small head dims, additive relative-position bias + shift mask).
//
// Everything is read from / written to the PACKED token-major projection tensors, so no per-head copies exist:
//     qkv  [B][S][3][H][d]     (output of the qkv linear)         out [B][S][H][d]   (input of the out projection)
//     dqkv [B][S][3][H][d]     (gradient of the qkv linear)       dout[B][S][H][d]
//
// One 128-row tile of queries per step (TMEM has 128 lanes).  Sequences of at most 64 tokens are stacked two
// problems to a tile (same head, consecutive batch / window index): the score matrix is then block diagonal
// and the off-diagonal blocks are written as exact zeros, so the second product needs no special casing.
//
// Forward (one CTA per query tile, two CTAs per SM):
//     S = Q K^T                 tcgen05.mma, M = 128, N = keys (<= 256), fp32 in TMEM
//     P = softmax rows          4 warps, one TMEM lane (row) per thread: tcgen05.ld, two passes (max, then exp / sum),
//                               P written as 16-bit into 128B-swizzled shared memory over the dead Q / K tiles
//     O = P V                   tcgen05.mma, A = P (K-major), B = V as loaded ([key][d], i.e. MN-major), into the
//                               TMEM columns S no longer needs; rows scaled by 1 / sum on the way out
// Backward (one CTA per (batch, head), 128 x 128 blocks of the score matrix, everything accumulated in TMEM):
//     S = Q K^T, dP = dO V^T                        recomputed per block
//     P = exp(S * scale + bias - lse), dS = P o (dP - delta) * scale       delta = rowsum(dO o O)
//     dV += P^T dO,  dK += dS^T Q,  dQ += dS K      A operands P / dS are the SAME shared-memory tiles read
//                                                   K-major (dQ) or MN-major (dV, dK): no transposes anywhere
//   TMEM columns: S 0-127 | dP 128-255 | dQ tile0 256-319 | dQ tile1 320-383 | dK 384-447 | dV 448-511
// Operand staging: TMA (3-D tensor maps over the packed tensors, out-of-range tokens zero-filled) when the head
// dimension is 64; otherwise the threads copy the rows into the same swizzled layout, zero-padded to 64.
//
// Reference: the attention inside nn.MultiheadAttention / WindowAttention of the transformer models
// (SwinIR built at Stoke-DDP.py:206-208; ViT-B/16 is BASELINE config 4); the library path there is
// F.scaled_dot_product_attention (flash / efficient kernels).
#include <string.h>

#include "tc_common.cuh"

namespace b200dp {
namespace attn {

using namespace b200dp::tc;

constexpr int kTile = 128;
constexpr int kTileBytes = kTile * 128;      // 128 rows x 64 x 2 B
constexpr float kLog2e = 1.4426950408889634f;

struct Geom {
    int B, H, S, d;    // B x H problems of S tokens, head dimension d (even, <= 64)
    int G;             // problems stacked per tile: 2 when S <= 64, else 1
    int spad;          // tile rows per stacked problem (128 / G)
    int nq;            // 128-row tiles per problem (1 or 2)
    int nk;            // forward: keys staged per tile (multiple of 16, <= 256)
    int manual;        // 1: threads stage the operands (head dim != 64 or stacked problems); 0: TMA
    int nW;            // number of mask windows (mask index = batch % nW); 0: no mask
    float scale;
    int qp, op, dop;   // elements between consecutive tokens of qkv / out / dout (>= 3 H d, H d, H d)
    // window mode (win != 0): problem b is window (b % wpi) of image b / wpi, an ws x ws patch of an Hi x Wi token grid
    // cyclically shifted by `shift`; token rows are then addressed in IMAGE order, so no partition / roll copy exists
    int win, ws, nWx, wpi, Hi, Wi, shift;
    // shifts for the index arithmetic (the kernels are issue-bound on these shapes: no runtime divisions on the hot path)
    int sp_sh;         // log2(spad)
    int p2;            // window geometry is all powers of two: ws, nWx, wpi (then *_sh are their logs)
    int ws_sh, nwx_sh, wpi_sh;
    int nw_mask;       // nW - 1 when nW is a power of two, else -1
};

__device__ __forceinline__ int mask_index(const Geom& g, int b) { return g.nw_mask >= 0 ? (b & g.nw_mask) : (b % g.nW); }

// Row index (in units of the token pitch) of token s of problem b.
__device__ __forceinline__ int64_t tok(const Geom& g, int b, int s) {
    if (g.win == 0) return (int64_t)b * g.S + s;
    int img, wy, wx, ty, tx;
    if (g.p2) {
        img = b >> g.wpi_sh;
        const int w = b & (g.wpi - 1);
        wy = w >> g.nwx_sh; wx = w & (g.nWx - 1);
        ty = s >> g.ws_sh; tx = s & (g.ws - 1);
    } else {
        img = b / g.wpi;
        const int w = b - img * g.wpi;
        wy = w / g.nWx; wx = w - wy * g.nWx;
        ty = s / g.ws; tx = s - ty * g.ws;
    }
    int y = wy * g.ws + ty + g.shift, x = wx * g.ws + tx + g.shift;
    if (y >= g.Hi) y -= g.Hi;
    if (x >= g.Wi) x -= g.Wi;
    return ((int64_t)img * g.Hi + y) * g.Wi + x;
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

// Copy `rows` token rows of one projection (q, k or v of head h; or dout) into a 128B-swizzled K-major tile,
// zero-padding the head dimension to 64 and rows that do not exist.  src points at element [0][0][which][h][0].
template <typename T>
__device__ __forceinline__ void stage_rows(uint8_t* dst, const T* __restrict__ src, int64_t pitch /* elements between tokens */,
                                           const Geom& g, int b0, int s0, int rows) {
    const int sp_sh = g.G == 1 ? 20 : g.sp_sh;              // one problem: every row is a token of it
    const uint32_t sdst = smem_u32(dst);
    for (int row = threadIdx.x; row < rows; row += blockDim.x) {
        const int gi = row >> sp_sh, lr = row - (gi << sp_sh);
        const int b = b0 + gi, s = s0 + lr;
        const bool ok = b < g.B && s < g.S && gi < g.G;
        const uint32_t* sp = reinterpret_cast<const uint32_t*>(src + tok(g, ok ? b : 0, ok ? s : 0) * pitch);
        const int nw = ok ? (g.d >> 1) : 0;                  // 32-bit words of this row that exist
        // every load is issued before the first store; words past the head dimension stay zero
        if (g.d <= 16) {                                     // small heads: 8 words instead of 32 predicated loads
            uint32_t w[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) w[e] = e < nw ? __ldg(sp + e) : 0u;
#pragma unroll
            for (int j = 0; j < 8; ++j)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sdst + row * 128 + ((j ^ (row & 7)) << 4)),
                             "r"(j < 2 ? w[4 * (j & 1)] : 0u), "r"(j < 2 ? w[4 * (j & 1) + 1] : 0u),
                             "r"(j < 2 ? w[4 * (j & 1) + 2] : 0u), "r"(j < 2 ? w[4 * (j & 1) + 3] : 0u) : "memory");
            continue;
        }
        uint32_t w[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) w[e] = e < nw ? __ldg(sp + e) : 0u;
#pragma unroll
        for (int j = 0; j < 8; ++j)
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sdst + row * 128 + ((j ^ (row & 7)) << 4)),
                         "r"(w[4 * j]), "r"(w[4 * j + 1]), "r"(w[4 * j + 2]), "r"(w[4 * j + 3]) : "memory");
    }
}

// Up to four tiles staged in ONE round: every thread first issues the loads of all its rows (of all tiles), then stores.
// The staging of a tile is a pure latency chain (global load -> shared store); doing the tiles one after the other pays
// that latency once per tile.
template <typename T>
struct StageJob { uint8_t* dst; const T* src; int64_t pitch; int s0; int rows; };

template <typename T, int NJ>
__device__ __forceinline__ void stage_many(const StageJob<T> (&jobs)[NJ], const Geom& g, int b0) {
    const int sp_sh = g.G == 1 ? 20 : g.sp_sh;
    // 128-row tiles, 256 threads: thread t serves row (t & 127) of jobs {t >> 7, (t >> 7) + 2}
    const int row = threadIdx.x & 127, jsel = threadIdx.x >> 7;
    constexpr int kPer = (NJ + 1) / 2;
    const int gi = row >> sp_sh, lr = row - (gi << sp_sh);
    const int b = b0 + gi;
    const bool small = g.d <= 16;
    uint32_t w[kPer][32];
    bool live[kPer];
    int64_t t_prev = 0;
    int s_prev = -1;
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        const int j = jsel + 2 * k;
        live[k] = j < NJ && row < jobs[j < NJ ? j : 0].rows;
        const StageJob<T>& jb = jobs[j < NJ ? j : 0];
        const int sidx = jb.s0 + lr;
        const bool okp = b < g.B && sidx < g.S && gi < g.G;
        const bool ok = live[k] && okp;
        // the jobs of one call usually start at the same token: one index computation serves them all
        if (sidx != s_prev) { t_prev = tok(g, okp ? b : 0, okp ? sidx : 0); s_prev = sidx; }
        const uint32_t* sp = reinterpret_cast<const uint32_t*>(jb.src + t_prev * jb.pitch);
        const int nw = ok ? (g.d >> 1) : 0;
        if (small) {
#pragma unroll
            for (int e = 0; e < 8; ++e) w[k][e] = e < nw ? __ldg(sp + e) : 0u;
        } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) w[k][e] = e < nw ? __ldg(sp + e) : 0u;
        }
    }
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
        if (!live[k]) continue;
        const int j = jsel + 2 * k;
        const uint32_t sdst = smem_u32(jobs[j].dst);
        if (small) {
#pragma unroll
            for (int p = 0; p < 8; ++p)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sdst + row * 128 + ((p ^ (row & 7)) << 4)),
                             "r"(p < 2 ? w[k][4 * (p & 1)] : 0u), "r"(p < 2 ? w[k][4 * (p & 1) + 1] : 0u),
                             "r"(p < 2 ? w[k][4 * (p & 1) + 2] : 0u), "r"(p < 2 ? w[k][4 * (p & 1) + 3] : 0u) : "memory");
        } else {
#pragma unroll
            for (int p = 0; p < 8; ++p)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(sdst + row * 128 + ((p ^ (row & 7)) << 4)),
                             "r"(w[k][4 * p]), "r"(w[k][4 * p + 1]), "r"(w[k][4 * p + 2]), "r"(w[k][4 * p + 3]) : "memory");
        }
    }
}

// Which (batch, token) a tile row / block column stands for.
struct RowId { int b, s, gi; bool ok; };
__device__ __forceinline__ RowId row_id(const Geom& g, int b0, int tile, int r) {
    RowId o;
    o.gi = r >> g.sp_sh;
    o.b = b0 + o.gi;
    o.s = tile * kTile + (r - (o.gi << g.sp_sh));
    o.ok = o.b < g.B && o.s < g.S && o.gi < g.G;
    return o;
}

// First d of N fp32 values -> 16-bit row in global memory (fully unrolled with predicates: v stays in registers).
template <typename T, int N>
__device__ __forceinline__ void store_row(T* dst, const float (&v)[N], int d) {
    if ((d & 7) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
#pragma unroll
        for (int e = 0; e < N; e += 8) {
            if (e < d) {
                const float t[8] = {v[e], v[e + 1], v[e + 2], v[e + 3], v[e + 4], v[e + 5], v[e + 6], v[e + 7]};
                store8<T>(dst + e, t);
            }
        }
    } else {
#pragma unroll
        for (int e = 0; e < N; e += 2) {
            if (e < d) {
                T pr[2] = {Cvt<T>::from_f(v[e]), Cvt<T>::from_f(v[e + 1])};
                *reinterpret_cast<uint32_t*>(dst + e) = *reinterpret_cast<const uint32_t*>(pr);
            }
        }
    }
}

// 32 consecutive additive terms (bias + mask) of one score row, starting at key k0: every load is issued before the
// first use (no branches around them), out-of-range keys read a clamped address and are masked by the caller.
__device__ __forceinline__ void load_add32(float (&add)[32], const float* __restrict__ brow, const float* __restrict__ mrow,
                                           int k0, int S) {
#pragma unroll
    for (int j = 0; j < 32; ++j) add[j] = 0.f;
    const bool vec = (S & 3) == 0 && k0 >= 0 && k0 + 32 <= S;
#pragma unroll
    for (int w = 0; w < 2; ++w) {
        const float* row = w == 0 ? brow : mrow;
        if (row == nullptr) continue;
        if (vec) {
            float4 t[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) t[j] = __ldg(reinterpret_cast<const float4*>(row + k0) + j);
#pragma unroll
            for (int j = 0; j < 8; ++j) { add[4 * j] += t[j].x; add[4 * j + 1] += t[j].y; add[4 * j + 2] += t[j].z; add[4 * j + 3] += t[j].w; }
        } else {
            float t[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) t[j] = __ldg(row + min(max(k0 + j, 0), S - 1));
#pragma unroll
            for (int j = 0; j < 32; ++j) add[j] += t[j];
        }
    }
}

// Window attention (two stacked problems of <= 64 tokens): bias[h] + mask[window] for both problems of the tile is
// staged ONCE in shared memory with coalesced reads, as sAdd[gi][row][k ^ (row & 31)] — the XOR makes the row-wise
// reads of the softmax threads (lane = row) hit 32 different banks.  Reading the [S][S] rows straight from global
// memory costs one 128-byte line per lane per load.
__device__ __forceinline__ void stage_add(float* sAdd, const float* __restrict__ bias, const float* __restrict__ mask,
                                          const Geom& g, int h, int b0) {
    const int S = g.S;
    if (S == 64 && blockDim.x == 256) {
        // the common window (8 x 8): 2 x 64 x 64 floats = 2048 float4, 8 per thread, every load issued before the first store
        const bool has_m = mask != nullptr && g.nW > 0;
        float4 bv[8], mv[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx4 = threadIdx.x + it * 256, gi = idx4 >> 10, row = (idx4 >> 4) & 63, k4 = (idx4 & 15) * 4;
            const bool live = b0 + gi < g.B;
            bv[it] = (bias && live) ? __ldg(reinterpret_cast<const float4*>(bias + ((int64_t)h * 64 + row) * 64 + k4)) : make_float4(0.f, 0.f, 0.f, 0.f);
            mv[it] = (has_m && live) ? __ldg(reinterpret_cast<const float4*>(mask + ((int64_t)mask_index(g, b0 + gi) * 64 + row) * 64 + k4))
                                     : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx4 = threadIdx.x + it * 256, gi = idx4 >> 10, row = (idx4 >> 4) & 63, k4 = (idx4 & 15) * 4;
            float* dst = sAdd + gi * 4096 + row * 64;
            const int m = row & 31;
            dst[(k4 + 0) ^ m] = bv[it].x + mv[it].x;
            dst[(k4 + 1) ^ m] = bv[it].y + mv[it].y;
            dst[(k4 + 2) ^ m] = bv[it].z + mv[it].z;
            dst[(k4 + 3) ^ m] = bv[it].w + mv[it].w;
        }
        return;
    }
    for (int idx = threadIdx.x; idx < 2 * S * S; idx += blockDim.x) {
        const int gi = idx / (S * S), rem = idx - gi * S * S;
        const int row = rem / S, k = rem - row * S;
        float v = 0.f;
        if (b0 + gi < g.B) {
            if (bias) v += __ldg(bias + ((int64_t)h * S + row) * S + k);
            if (mask && g.nW > 0) v += __ldg(mask + ((int64_t)mask_index(g, b0 + gi) * S + row) * S + k);
        }
        sAdd[gi * 4096 + row * 64 + (k ^ (row & 31))] = v;
    }
}
__device__ __forceinline__ void smem_add32(float (&add)[32], const float* sAddRow /* sAdd + gi*4096 + row*64 */, int row, int k0) {
#pragma unroll
    for (int j = 0; j < 32; ++j) add[j] = sAddRow[(k0 + j) ^ (row & 31)];
}

// Row maximum over one 32-column chunk.  kAdd: scores carry an additive term (then the scaled score is compared);
// otherwise the RAW accumulator is compared and the caller scales the result (scale > 0).  kFull: every column
// of the chunk is a real key, so no per-element range test is needed.
template <bool kAdd, bool kFull>
__device__ __forceinline__ float chunk_max(const uint32_t (&v)[32], const float (&add)[32], float scale, int k0, int S, float m) {
    float m4[4] = {m, -INFINITY, -INFINITY, -INFINITY};        // four chains instead of one 32-long dependency
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        float x = kAdd ? fmaf(__uint_as_float(v[j]), scale, add[j]) : __uint_as_float(v[j]);
        if (!kFull) x = (k0 + j >= 0 && k0 + j < S) ? x : -INFINITY;
        m4[j & 3] = fmaxf(m4[j & 3], x);
    }
    return fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
}
// p[j] = 2^(v[j] * sl + add[j] * log2e - off)   (exp of the shifted score; off = log2e * (row max | row lse))
template <bool kAdd, bool kFull>
__device__ __forceinline__ void chunk_exp(const uint32_t (&v)[32], const float (&add)[32], float sl, float off, int k0, int S,
                                          float (&p)[32]) {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        float x = exp2f(kAdd ? fmaf(__uint_as_float(v[j]), sl, fmaf(add[j], kLog2e, -off)) : fmaf(__uint_as_float(v[j]), sl, -off));
        if (!kFull) x = (k0 + j >= 0 && k0 + j < S) ? x : 0.f;
        p[j] = x;
    }
}
#define ATTN_DISPATCH(has_add, full, CALL)                  \
    do {                                                    \
        if (has_add) { if (full) { CALL(true, true); } else { CALL(true, false); } }   \
        else { if (full) { CALL(false, true); } else { CALL(false, false); } }         \
    } while (0)

// 8 values -> 16 bytes of shared memory (explicit shared-space store: the tile pointers are generic)
template <typename T>
__device__ __forceinline__ void sts8(uint32_t saddr, const float (&v)[8]) {
    Vec8<T> r;
#pragma unroll
    for (int i = 0; i < 8; ++i) r.v[i] = Cvt<T>::from_f(v[i]);
    const uint4 u = *reinterpret_cast<const uint4*>(&r);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(u.x), "r"(u.y), "r"(u.z), "r"(u.w) : "memory");
}

// ---------------------------------------------------------------- forward
constexpr int kFwdSmem = 65536 + 32768 + 1024 /*align*/ + 128 /*barriers*/ + 2048 /*row partials*/;

template <typename T>
__global__ void __launch_bounds__(256, 2) attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQ,
                                                          const __grid_constant__ CUtensorMap tmKV,
                                                          const T* __restrict__ qkv, T* __restrict__ out,
                                                          float* __restrict__ lse, const float* __restrict__ bias,
                                                          const float* __restrict__ mask, const Geom g, int is_bf16) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;                     // 16 KB
    uint8_t* sK = smem + kTileBytes;        // up to 256 keys: 32 KB
    uint8_t* sP = smem;                     // 4 chunks of 64 keys x 128 rows: 64 KB, overlays Q and K once S exists
    uint8_t* sV = smem + 65536;             // up to 256 keys: 32 KB
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 65536 + 32768);
    uint64_t* bar_ld = bars;
    uint64_t* bar_s = bars + 1;
    uint64_t* bar_o = bars + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);
    float* s_max = reinterpret_cast<float*>(bars + 16);      // [2][128] partial row maxima of the two warp halves
    float* s_sum = s_max + 2 * kTile;                        // [2][128] partial row sums

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x, h = blockIdx.y, b0 = blockIdx.z * g.G;
    const int64_t pitch = g.qp;

    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 1); mbar_init(bar_s, 1); mbar_init(bar_o, 1);
        fence_barrier_init();
        if (!g.manual) { tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmKV); }
    }
    if (warp == 2) tmem_alloc(tmem_slot, 256);
    // window attention: the additive tile lives in the upper half of the P region (nk = 128 uses only the lower half)
    float* sAdd = reinterpret_cast<float*>(smem + 32768);
    const bool use_sadd = g.G == 2 && (bias != nullptr || (mask != nullptr && g.nW > 0));
    if (use_sadd) stage_add(sAdd, bias, mask, g, h, b0);
    if (g.manual) {
        stage_rows<T>(sQ, qkv + (int64_t)h * g.d, pitch, g, b0, qt * kTile, kTile);
        stage_rows<T>(sK, qkv + (int64_t)(g.H + h) * g.d, pitch, g, b0, 0, g.nk);
        stage_rows<T>(sV, qkv + (int64_t)(2 * g.H + h) * g.d, pitch, g, b0, 0, g.nk);
        fence_async_smem();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const int ksteps = (g.d + 15) >> 4;

    if (warp == 1) {
        // lanes 1..31 park at the __syncwarp below instead of sleeping on an mbarrier: a suspended
        // try_wait of the sibling lanes must never delay the one lane that issues the TMA / MMA work
        if (lane == 0) {
            if (!g.manual) {
                mbar_expect_tx(bar_ld, kTileBytes + 2 * g.nk * 128);
                tma_load_3d(sQ, &tmQ, bar_ld, h * 64, qt * kTile, b0);
                tma_load_3d(sK, &tmKV, bar_ld, (g.H + h) * 64, 0, b0);
                tma_load_3d(sV, &tmKV, bar_ld, (2 * g.H + h) * 64, 0, b0);
                mbar_wait(bar_ld, 0);
                tc_fence_after();
            }
            const uint32_t idesc = make_idesc(is_bf16, 0, 0, kTile, g.nk);
            const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK);
            for (int k = 0; k < ksteps; ++k)
                umma_f16(tmem_base, make_desc(qa + k * 32, 16, 1024), make_desc(ka + k * 32, 16, 1024), idesc, k ? 1u : 0u);
            umma_commit(bar_s);
        }
        __syncwarp();
    }

    // ---- softmax: ALL eight warps.  Warps w and w + 4 reach the same TMEM lane quarter; they split a row's
    //      32-column chunks by parity and combine their row maxima / sums through shared memory.
    const int r = (warp & 3) * 32 + lane, half = warp >> 2;       // tile row == TMEM lane
    const RowId me = row_id(g, b0, qt, r);
    const int klo = g.G == 1 ? 0 : me.gi * g.spad;                // warp-uniform: a warp's 32 rows belong to one problem
    const float* brow = bias ? bias + ((int64_t)h * g.S + (me.ok ? me.s : 0)) * g.S : nullptr;
    const float* mrow = (mask && g.nW > 0) ? mask + ((int64_t)mask_index(g, me.ok ? me.b : 0) * g.S + (me.ok ? me.s : 0)) * g.S : nullptr;
    const uint32_t taddr = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
    const int nch = (g.nk + 31) >> 5;
    const int ch_lo = klo >> 5, ch_hi = min(nch, (klo + g.S + 31) >> 5);          // chunks holding this problem's keys
    const uint32_t p_row = smem_u32(sP) + r * 128;
    const bool has_add = brow != nullptr || mrow != nullptr || !(g.scale > 0.f);
    const float* sadd_row = sAdd + me.gi * 4096 + (me.ok ? me.s : 0) * 64;
    mbar_wait(bar_s, 0);
    tc_fence_after();
    float mx = -INFINITY;
    const bool plain = !has_add && g.G == 1;          // ViT: no additive term, one problem per tile
    if (plain) {
        // two register buffers, the TMEM load of the next chunk in flight while this one is reduced
        uint32_t va[32], vb[32];
        float dummy[32];
        int ch = ch_lo + ((ch_lo ^ half) & 1);
        if (ch < ch_hi) tmem_ld_32x32b_x32(taddr + ch * 32, va);
        while (ch < ch_hi) {
            tmem_ld_wait();
            if (ch + 2 < ch_hi) tmem_ld_32x32b_x32(taddr + (ch + 2) * 32, vb);
            if (ch * 32 + 32 <= g.S) mx = chunk_max<false, true>(va, dummy, 1.f, ch * 32, g.S, mx);
            else mx = chunk_max<false, false>(va, dummy, 1.f, ch * 32, g.S, mx);
            ch += 2;
            if (ch >= ch_hi) break;
            tmem_ld_wait();
            if (ch + 2 < ch_hi) tmem_ld_32x32b_x32(taddr + (ch + 2) * 32, va);
            if (ch * 32 + 32 <= g.S) mx = chunk_max<false, true>(vb, dummy, 1.f, ch * 32, g.S, mx);
            else mx = chunk_max<false, false>(vb, dummy, 1.f, ch * 32, g.S, mx);
            ch += 2;
        }
    }
    for (int ch = ch_lo + ((ch_lo ^ half) & 1); !plain && ch < ch_hi; ch += 2) {
        uint32_t v[32];
        float add[32];
        const int k0 = ch * 32 - klo;
        tmem_ld_32x32b_x32(taddr + ch * 32, v);
        if (use_sadd) smem_add32(add, sadd_row, me.ok ? me.s : 0, k0);
        else if (has_add) load_add32(add, brow, mrow, k0, g.S);
        tmem_ld_wait();
        const bool full = k0 >= 0 && k0 + 32 <= g.S;
#define CALL(A, F) mx = chunk_max<A, F>(v, add, g.scale, k0, g.S, mx)
        ATTN_DISPATCH(has_add, full, CALL);
#undef CALL
    }
    if (!has_add) mx *= g.scale;
    s_max[half * kTile + r] = mx;
    __syncthreads();
    mx = fmaxf(s_max[r], s_max[kTile + r]);
    if (!me.ok || mx == -INFINITY) mx = 0.f;
    const float mxl = mx * kLog2e, sl = g.scale * kLog2e;
    float sum = 0.f;
    auto put_p = [&](int ch, const float (&p)[32]) {
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 32; ++j) s4[j & 3] += p[j];
        sum += (s4[0] + s4[1]) + (s4[2] + s4[3]);
#pragma unroll
        for (int q8 = 0; q8 < 4; ++q8) {
            const float t[8] = {p[q8 * 8], p[q8 * 8 + 1], p[q8 * 8 + 2], p[q8 * 8 + 3], p[q8 * 8 + 4], p[q8 * 8 + 5], p[q8 * 8 + 6], p[q8 * 8 + 7]};
            const int c0 = ch * 32 + q8 * 8;
            sts8<T>(p_row + (c0 >> 6) * kTileBytes + ((((c0 >> 3) & 7) ^ (r & 7)) << 4), t);
        }
    };
    if (plain) {
        // every chunk of [0, nch) is live
        uint32_t va[32], vb[32];
        float dummy[32], p[32];
        int ch = half;
        const int end = nch;
        if (ch < end) tmem_ld_32x32b_x32(taddr + ch * 32, va);
        while (ch < end) {
            tmem_ld_wait();
            if (ch + 2 < end) tmem_ld_32x32b_x32(taddr + (ch + 2) * 32, vb);
            if (ch * 32 + 32 <= g.S) chunk_exp<false, true>(va, dummy, sl, mxl, ch * 32, g.S, p);
            else chunk_exp<false, false>(va, dummy, sl, mxl, ch * 32, g.S, p);
            put_p(ch, p);
            ch += 2;
            if (ch >= end) break;
            tmem_ld_wait();
            if (ch + 2 < end) tmem_ld_32x32b_x32(taddr + (ch + 2) * 32, va);
            if (ch * 32 + 32 <= g.S) chunk_exp<false, true>(vb, dummy, sl, mxl, ch * 32, g.S, p);
            else chunk_exp<false, false>(vb, dummy, sl, mxl, ch * 32, g.S, p);
            put_p(ch, p);
            ch += 2;
        }
    }
    for (int ch = half; !plain && ch < nch; ch += 2) {
        float p[32];
        if (ch >= ch_lo && ch < ch_hi) {                                          // warp-uniform
            uint32_t v[32];
            float add[32];
            const int k0 = ch * 32 - klo;
            tmem_ld_32x32b_x32(taddr + ch * 32, v);
            if (use_sadd) smem_add32(add, sadd_row, me.ok ? me.s : 0, k0);
            else if (has_add) load_add32(add, brow, mrow, k0, g.S);
            tmem_ld_wait();
            const bool full = k0 >= 0 && k0 + 32 <= g.S;
#define CALL(A, F) chunk_exp<A, F>(v, add, sl, mxl, k0, g.S, p)
            ATTN_DISPATCH(has_add, full, CALL);
#undef CALL
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) p[j] = 0.f;
        }
        put_p(ch, p);
    }
    s_sum[half * kTile + r] = sum;
    fence_async_smem();
    tc_fence_before();
    __syncthreads();                                      // P is complete; every thread is done with the S columns
    if (warp == 1) {
        // lanes 1..31 park at the __syncwarp below instead of sleeping on an mbarrier: a suspended
        // try_wait of the sibling lanes must never delay the one lane that issues the TMA / MMA work
        if (lane == 0) {
            tc_fence_after();
            const uint32_t idesc = make_idesc(is_bf16, 0, 1, kTile, 64);
            const uint32_t pa = smem_u32(sP), va = smem_u32(sV);
            const int steps = g.nk >> 4;
            for (int k = 0; k < steps; ++k)
                umma_f16(tmem_base, make_desc(pa + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                         make_desc(va + k * 2048, 8192, 1024), idesc, k ? 1u : 0u);
            umma_commit(bar_o);
        }
        __syncwarp();
    }
    sum = s_sum[r] + s_sum[kTile + r];
    const float inv_sum = sum > 0.f ? 1.f / sum : 0.f;
    mbar_wait(bar_o, 0);
    tc_fence_after();
    {
        // each half of the warps drains 32 of the 64 output columns of its rows
        uint32_t v[32];
        tmem_ld_32x32b_x32(taddr + half * 32, v);
        tmem_ld_wait();
        float o[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) o[j] = __uint_as_float(v[j]) * inv_sum;
        if (me.ok && g.d > half * 32)
            store_row<T, 32>(out + tok(g, me.b, me.s) * g.op + h * g.d + half * 32, o, min(32, g.d - half * 32));
        if (me.ok && half == 0) lse[((int64_t)me.b * g.H + h) * g.S + me.s] = mx + logf(fmaxf(sum, 1e-30f));
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 256);
}

// ---------------------------------------------------------------- backward
constexpr int kDbStride = 65;    // floats per row of the per-CTA bias-gradient accumulator (bank spread)
constexpr int kBwdSmemBase = 10 * kTileBytes;                      // Q0 Q1 dO0 dO1 K V P(2) dS(2)
constexpr int kBwdSmem = kBwdSmemBase + 2 * 64 * kDbStride * 4 + 1024 + 128;

template <typename T>
__device__ __forceinline__ float row_dot(const T* __restrict__ a, const T* __restrict__ b, int d) {
    float s = 0.f;
    if ((d & 7) == 0 && ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) & 15) == 0) {
        for (int e = 0; e < d; e += 8) {
            float x[8], y[8];
            load8<T>(a + e, x);
            load8<T>(b + e, y);
#pragma unroll
            for (int k = 0; k < 8; ++k) s = fmaf(x[k], y[k], s);
        }
    } else {
        for (int e = 0; e < d; e += 2) {
            const float2 x = unpack2<T>(*reinterpret_cast<const uint32_t*>(a + e));
            const float2 y = unpack2<T>(*reinterpret_cast<const uint32_t*>(b + e));
            s = fmaf(x.x, y.x, fmaf(x.y, y.y, s));
        }
    }
    return s;
}

template <typename T>
__global__ void __launch_bounds__(256, 1) attn_bwd_kernel(const __grid_constant__ CUtensorMap tmQKV,
                                                          const __grid_constant__ CUtensorMap tmDO,
                                                          const T* __restrict__ qkv, const T* __restrict__ out,
                                                          const T* __restrict__ dout, const float* __restrict__ lse,
                                                          T* __restrict__ dqkv, const float* __restrict__ bias,
                                                          const float* __restrict__ mask, float* __restrict__ dbias,
                                                          const Geom g, int is_bf16) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* sQ = smem;                          // 2 tiles
    // stacked windows use one query tile: dO moves next to Q and the freed 32 KB hold the additive (bias + mask) tile
    uint8_t* sDO = g.G == 2 ? smem + kTileBytes : smem + 2 * kTileBytes;        // 2 tiles (1 when G == 2)
    float* sAdd = reinterpret_cast<float*>(smem + 2 * kTileBytes);
    uint8_t* sK = smem + 4 * kTileBytes;
    uint8_t* sV = smem + 5 * kTileBytes;
    uint8_t* sP = smem + 6 * kTileBytes;         // 2 chunks of 64 keys
    uint8_t* sDS = smem + 8 * kTileBytes;        // 2 chunks of 64 keys
    float* sDB = reinterpret_cast<float*>(smem + kBwdSmemBase);          // [2][64][kDbStride]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kBwdSmemBase + 2 * 64 * kDbStride * 4);
    uint64_t* bar_ld = bars;
    uint64_t* bar_s = bars + 1;
    uint64_t* bar_m = bars + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 3);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int h = blockIdx.y;
    const int64_t pitch = g.qp, opitch = g.dop;
    const int nunits = (g.B + g.G - 1) / g.G;
    const int nkb = g.nq;                         // 128-key blocks per problem (== query tiles)
    const int ksteps = (g.d + 15) >> 4;

    if (threadIdx.x == 0) {
        mbar_init(bar_ld, 1); mbar_init(bar_s, 1); mbar_init(bar_m, 1);
        fence_barrier_init();
        if (!g.manual) { tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); }
    }
    if (warp == 2) tmem_alloc(tmem_slot, 512);
    if (dbias != nullptr)
        for (int i = threadIdx.x; i < 2 * 64 * kDbStride; i += blockDim.x) sDB[i] = 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tS = tmem_base, tDP = tmem_base + 128, tDQ = tmem_base + 256, tDK = tmem_base + 384, tDV = tmem_base + 448;

    uint32_t n_ld = 0, n_s = 0, n_m = 0;          // commits / transactions issued so far on each barrier (all threads count)
    // elementwise work and the TMEM drains use ALL eight warps: warps w and w + 4 share a lane quarter and split
    // the 32-column chunks (resp. the 64 output columns) between them
    const int r = (warp & 3) * 32 + lane, half = warp >> 2;
    const uint32_t lane_sel = (uint32_t)((warp & 3) * 32) << 16;

    const bool use_sadd = g.G == 2 && (bias != nullptr || (mask != nullptr && g.nW > 0));
    bool sadd_valid = false;
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
        const int b0 = unit * g.G;
        if (use_sadd && (!sadd_valid || (mask != nullptr && g.nW > 0))) {
            // every thread is past its last read of the previous tile (the barrier before the second group of MMAs)
            stage_add(sAdd, bias, mask, g, h, b0);
            sadd_valid = true;
        }
        float delta[2] = {0.f, 0.f}, row_lse[2] = {0.f, 0.f};
        for (int j = 0; j < nkb; ++j) {
            for (int i = 0; i < g.nq; ++i) {
                const bool need_q = j == 0, need_kv = i == 0;
                // ---- operands.  Tiles may be overwritten only once every MMA issued so far has retired.
                if (g.manual && (need_q || need_kv)) {
                    if (n_m > 0) mbar_wait(bar_m, (n_m - 1) & 1);
                    if (need_q && need_kv) {
                        const StageJob<T> jobs[4] = {
                            {sQ + i * kTileBytes, qkv + (int64_t)h * g.d, pitch, i * kTile, kTile},
                            {sDO + i * kTileBytes, dout + (int64_t)h * g.d, opitch, i * kTile, kTile},
                            {sK, qkv + (int64_t)(g.H + h) * g.d, pitch, j * kTile, kTile},
                            {sV, qkv + (int64_t)(2 * g.H + h) * g.d, pitch, j * kTile, kTile}};
                        stage_many<T, 4>(jobs, g, b0);
                    } else if (need_q) {
                        const StageJob<T> jobs[2] = {{sQ + i * kTileBytes, qkv + (int64_t)h * g.d, pitch, i * kTile, kTile},
                                                     {sDO + i * kTileBytes, dout + (int64_t)h * g.d, opitch, i * kTile, kTile}};
                        stage_many<T, 2>(jobs, g, b0);
                    } else {
                        const StageJob<T> jobs[2] = {{sK, qkv + (int64_t)(g.H + h) * g.d, pitch, j * kTile, kTile},
                                                     {sV, qkv + (int64_t)(2 * g.H + h) * g.d, pitch, j * kTile, kTile}};
                        stage_many<T, 2>(jobs, g, b0);
                    }
                    fence_async_smem();
                }
                __syncthreads();
                if (warp == 1) {
                    // lanes 1..31 park at the __syncwarp below instead of sleeping on an mbarrier: a suspended
                    // try_wait of the sibling lanes must never delay the one lane that issues the TMA / MMA work
                    if (lane == 0) {
                        if (!g.manual && (need_q || need_kv)) {
                            if (n_m > 0) mbar_wait(bar_m, (n_m - 1) & 1);
                            mbar_expect_tx(bar_ld, (need_q ? 2 : 0) * kTileBytes + (need_kv ? 2 : 0) * kTileBytes);
                            if (need_q) {
                                tma_load_3d(sQ + i * kTileBytes, &tmQKV, bar_ld, h * 64, i * kTile, b0);
                                tma_load_3d(sDO + i * kTileBytes, &tmDO, bar_ld, h * 64, i * kTile, b0);
                            }
                            if (need_kv) {
                                tma_load_3d(sK, &tmQKV, bar_ld, (g.H + h) * 64, j * kTile, b0);
                                tma_load_3d(sV, &tmQKV, bar_ld, (2 * g.H + h) * 64, j * kTile, b0);
                            }
                            mbar_wait(bar_ld, n_ld & 1);
                        }
                        tc_fence_after();
                        const uint32_t idesc = make_idesc(is_bf16, 0, 0, kTile, kTile);
                        const uint32_t qa = smem_u32(sQ + i * kTileBytes), doa = smem_u32(sDO + i * kTileBytes);
                        const uint32_t ka = smem_u32(sK), va = smem_u32(sV);
                        for (int k = 0; k < ksteps; ++k)
                            umma_f16(tS, make_desc(qa + k * 32, 16, 1024), make_desc(ka + k * 32, 16, 1024), idesc, k ? 1u : 0u);
                        for (int k = 0; k < ksteps; ++k)
                            umma_f16(tDP, make_desc(doa + k * 32, 16, 1024), make_desc(va + k * 32, 16, 1024), idesc, k ? 1u : 0u);
                        umma_commit(bar_s);
                    }
                    __syncwarp();
                }
                if (!g.manual && (need_q || need_kv)) ++n_ld;
                ++n_s;
                {
                    const RowId me = row_id(g, b0, i, r);
                    if (j == 0) {
                        // per-row constants of this query tile, fetched while the tensor core works
                        if (me.ok) {
                            const int64_t t = tok(g, me.b, me.s);
                            delta[i] = row_dot<T>(out + t * g.op + h * g.d, dout + t * g.dop + h * g.d, g.d);
                            row_lse[i] = lse[((int64_t)me.b * g.H + h) * g.S + me.s];
                        } else {
                            delta[i] = 0.f; row_lse[i] = 0.f;
                        }
                    }
                    // key index of tile column c is c - koff (warp-uniform: a warp's rows belong to one problem)
                    const int koff = g.G == 1 ? -j * kTile : me.gi * g.spad;
                    const int ch_lo = g.G == 1 ? 0 : (koff >> 5), ch_hi = g.G == 1 ? 4 : min(4, (koff + g.S + 31) >> 5);
                    const float* brow = bias ? bias + ((int64_t)h * g.S + (me.ok ? me.s : 0)) * g.S : nullptr;
                    const float* mrow = (mask && g.nW > 0) ? mask + ((int64_t)mask_index(g, me.ok ? me.b : 0) * g.S + (me.ok ? me.s : 0)) * g.S : nullptr;
                    float* dbrow = (dbias && me.ok) ? sDB + (me.gi * 64 + me.s) * kDbStride : nullptr;     // rows are thread-private
                    // rows that stand for no token get lse = +inf: their probabilities come out as exact zeros
                    const float lsel = me.ok ? row_lse[i] * kLog2e : INFINITY, sl = g.scale * kLog2e, dl = delta[i];
                    const bool has_add = brow != nullptr || mrow != nullptr;
                    const uint32_t p_row = smem_u32(sP) + r * 128, ds_row = smem_u32(sDS) + r * 128;
                    mbar_wait(bar_s, (n_s - 1) & 1);
                    tc_fence_after();
#pragma unroll 1
                    for (int ch = half; ch < 4; ch += 2) {
                        float p[32], ds[32];
                        if (ch >= ch_lo && ch < ch_hi) {                          // warp-uniform
                            uint32_t sv[32], dv[32];
                            float add[32];
                            const int k0 = ch * 32 - koff;
                            tmem_ld_32x32b_x32(tS + lane_sel + ch * 32, sv);
                            tmem_ld_32x32b_x32(tDP + lane_sel + ch * 32, dv);
                            if (use_sadd) smem_add32(add, sAdd + me.gi * 4096 + (me.ok ? me.s : 0) * 64, me.ok ? me.s : 0, k0);
                            else if (has_add) load_add32(add, brow, mrow, k0, g.S);
                            tmem_ld_wait();
                            const bool full = k0 >= 0 && k0 + 32 <= g.S;
#define CALL(A, F) chunk_exp<A, F>(sv, add, sl, lsel, k0, g.S, p)
                            ATTN_DISPATCH(has_add, full, CALL);
#undef CALL
#pragma unroll
                            for (int jj = 0; jj < 32; ++jj) ds[jj] = p[jj] * (__uint_as_float(dv[jj]) - dl);
                            if (dbrow) {       // the two halves of a row own different chunks: no two threads share an entry
#pragma unroll
                                for (int jj = 0; jj < 32; ++jj)
                                    if (full || (k0 + jj >= 0 && k0 + jj < g.S)) dbrow[k0 + jj] += ds[jj];
                            }
                        } else {
#pragma unroll
                            for (int jj = 0; jj < 32; ++jj) { p[jj] = 0.f; ds[jj] = 0.f; }
                        }
#pragma unroll
                        for (int q8 = 0; q8 < 4; ++q8) {
                            float tp[8], td[8];
#pragma unroll
                            for (int e = 0; e < 8; ++e) { tp[e] = p[q8 * 8 + e]; td[e] = ds[q8 * 8 + e] * g.scale; }
                            const int c0 = ch * 32 + q8 * 8;
                            const uint32_t off = (c0 >> 6) * kTileBytes + ((((c0 >> 3) & 7) ^ (r & 7)) << 4);
                            sts8<T>(p_row + off, tp);
                            sts8<T>(ds_row + off, td);
                        }
                    }
                    fence_async_smem();
                    tc_fence_before();
                }
                __syncthreads();
                if (warp == 1) {
                    // lanes 1..31 park at the __syncwarp below instead of sleeping on an mbarrier: a suspended
                    // try_wait of the sibling lanes must never delay the one lane that issues the TMA / MMA work
                    if (lane == 0) {
                        tc_fence_after();
                        const uint32_t id_t = make_idesc(is_bf16, 1, 1, kTile, 64);     // A^T forms: dV, dK
                        const uint32_t id_n = make_idesc(is_bf16, 0, 1, kTile, 64);     // dQ
                        const uint32_t pa = smem_u32(sP), dsa = smem_u32(sDS), ka = smem_u32(sK);
                        const uint32_t qa = smem_u32(sQ + i * kTileBytes), doa = smem_u32(sDO + i * kTileBytes);
                        for (int k = 0; k < 8; ++k)         // dV[key][d] += sum_q P[q][key] dO[q][d]
                            umma_f16(tDV, make_desc(pa + k * 2048, kTileBytes, 1024), make_desc(doa + k * 2048, 8192, 1024), id_t,
                                     (i > 0 || k > 0) ? 1u : 0u);
                        for (int k = 0; k < 8; ++k)         // dK[key][d] += sum_q dS[q][key] Q[q][d]
                            umma_f16(tDK, make_desc(dsa + k * 2048, kTileBytes, 1024), make_desc(qa + k * 2048, 8192, 1024), id_t,
                                     (i > 0 || k > 0) ? 1u : 0u);
                        for (int k = 0; k < 8; ++k)         // dQ[q][d] += sum_key dS[q][key] K[key][d]
                            umma_f16(tDQ + i * 64, make_desc(dsa + (k >> 2) * kTileBytes + (k & 3) * 32, 16, 1024),
                                     make_desc(ka + k * 2048, 8192, 1024), id_n, (j > 0 || k > 0) ? 1u : 0u);
                        umma_commit(bar_m);
                    }
                    __syncwarp();
                }
                ++n_m;
            }
            // ---- dK, dV of key block j are complete: each warp half drains 32 of the 64 columns
            {
                mbar_wait(bar_m, (n_m - 1) & 1);
                tc_fence_after();
                const RowId kr = row_id(g, b0, j, r);
#pragma unroll 1
                for (int w = 0; w < 2; ++w) {
                    uint32_t v[32];
                    tmem_ld_32x32b_x32((w == 0 ? tDK : tDV) + lane_sel + half * 32, v);
                    tmem_ld_wait();
                    float o[32];
#pragma unroll
                    for (int q = 0; q < 32; ++q) o[q] = __uint_as_float(v[q]);
                    if (kr.ok && g.d > half * 32)
                        store_row<T, 32>(dqkv + tok(g, kr.b, kr.s) * g.qp + ((w + 1) * g.H + h) * g.d + half * 32, o,
                                         min(32, g.d - half * 32));
                }
                tc_fence_before();
            }
        }
        // ---- dQ of every query tile is complete (the last wait above covered the last commit)
        {
            tc_fence_after();
#pragma unroll 1
            for (int i = 0; i < g.nq; ++i) {
                const RowId qr = row_id(g, b0, i, r);
                uint32_t v[32];
                tmem_ld_32x32b_x32(tDQ + i * 64 + lane_sel + half * 32, v);
                tmem_ld_wait();
                float o[32];
#pragma unroll
                for (int q = 0; q < 32; ++q) o[q] = __uint_as_float(v[q]);
                if (qr.ok && g.d > half * 32)
                    store_row<T, 32>(dqkv + tok(g, qr.b, qr.s) * g.qp + h * g.d + half * 32, o,
                                     min(32, g.d - half * 32));
            }
            tc_fence_before();
        }
    }
    __syncthreads();
    if (dbias != nullptr) {
        // rows of the two stacked problems add into the same [H][S][S] gradient
        for (int idx = threadIdx.x; idx < g.S * g.S; idx += blockDim.x) {
            const int q = idx / g.S, k = idx - q * g.S;
            const float v = sDB[q * kDbStride + k] + sDB[(64 + q) * kDbStride + k];
            if (v != 0.f) atomicAdd(dbias + ((int64_t)h * g.S + q) * g.S + k, v);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// ---------------------------------------------------------------- host
static int make_tmap3(CUtensorMap* tm, const void* ptr, int is_bf16, int64_t cols, int64_t S, int64_t B, int box_rows) {
    static PFN_cuTensorMapEncodeTiled_v12000 enc =
        reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(driver_fn("cuTensorMapEncodeTiled"));
    B200DP_REQUIRE(enc != nullptr, "attention: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[3] = {(cuuint64_t)cols, (cuuint64_t)S, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)cols * 2, (cuuint64_t)S * cols * 2};
    cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    const CUtensorMapDataType dt = is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    CUresult r = enc(tm, dt, 3, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200DP_REQUIRE(r == CUDA_SUCCESS, "attention: cuTensorMapEncodeTiled failed (%d) cols=%lld S=%lld B=%lld", (int)r,
                   (long long)cols, (long long)S, (long long)B);
    return 0;
}

static int make_geom(Geom* g, int B, int H, int S, int d, float scale, int nW, const void* p0, const void* p1) {
    B200DP_REQUIRE(B > 0 && H > 0 && S > 0 && S <= 256 && d >= 2 && d <= 64 && (d & 1) == 0,
                   "attention: unsupported shape B=%d H=%d S=%d d=%d (S <= 256, even d <= 64)", B, H, S, d);
    g->B = B; g->H = H; g->S = S; g->d = d; g->scale = scale; g->nW = nW;
    g->G = S <= 64 ? 2 : 1;
    g->spad = kTile / g->G;
    g->nq = (S + kTile - 1) / kTile;
    g->nk = g->G == 2 ? kTile : ((S + 15) / 16) * 16;
    g->manual = !(d == 64 && g->G == 1 && ((uintptr_t)p0 & 15) == 0 && ((uintptr_t)p1 & 15) == 0);
    g->qp = 3 * H * d; g->op = H * d; g->dop = H * d;
    g->win = 0; g->ws = g->nWx = g->wpi = g->Hi = g->Wi = g->shift = 0;
    g->sp_sh = g->G == 2 ? 6 : 7;
    g->p2 = 0; g->ws_sh = g->nwx_sh = g->wpi_sh = 0;
    g->nw_mask = (nW > 0 && (nW & (nW - 1)) == 0) ? nW - 1 : -1;
    return 0;
}
static inline int ilog2(int v) { int r = 0; while ((1 << r) < v) ++r; return r; }

// Window mode on top of make_geom: images of Hi x Wi tokens cut into ws x ws windows after a cyclic shift; pitches in elements.
static int set_windows(Geom* g, int Hi, int Wi, int ws, int shift, int qp, int op, int dop) {
    B200DP_REQUIRE(ws > 0 && ws * ws == g->S && Hi % ws == 0 && Wi % ws == 0 && shift >= 0 && shift < ws,
                   "window attention: %d x %d tokens do not tile into %d x %d windows (S=%d, shift %d)", Hi, Wi, ws, ws, g->S, shift);
    const int wpi = (Hi / ws) * (Wi / ws);
    B200DP_REQUIRE(g->B % wpi == 0, "window attention: %d windows are not whole images of %d", g->B, wpi);
    B200DP_REQUIRE(qp >= 3 * g->H * g->d && op >= g->H * g->d && dop >= g->H * g->d && ((qp | op | dop) & 1) == 0,
                   "window attention: pitches %d / %d / %d too small or odd", qp, op, dop);
    g->win = 1; g->ws = ws; g->nWx = Wi / ws; g->wpi = wpi; g->Hi = Hi; g->Wi = Wi; g->shift = shift;
    g->qp = qp; g->op = op; g->dop = dop;
    g->manual = 1;
    const int nwx = Wi / ws;
    g->p2 = (ws & (ws - 1)) == 0 && (nwx & (nwx - 1)) == 0 && (wpi & (wpi - 1)) == 0;
    g->ws_sh = ilog2(ws); g->nwx_sh = ilog2(nwx); g->wpi_sh = ilog2(wpi);
    return 0;
}

}  // namespace attn
}  // namespace b200dp

using namespace b200dp;
using namespace b200dp::attn;

extern "C" int b200dp_attention_supported(int S, int d) { return S > 0 && S <= 256 && d >= 2 && d <= 64 && (d & 1) == 0; }

// out[B][S][H][d] = softmax(q k^T * scale + bias[H][S][S] + mask[b % nW][S][S]) v over the packed qkv[B][S][3][H][d];
// lse[B][H][S] (fp32) keeps the row log-sum-exp for the backward.  bias / mask may be NULL.
struct WinArgs { int Hi, Wi, ws, shift, qp, op, dop; };

static int attention_fwd_impl(const void* qkv, void* out, float* lse, const float* bias, const float* mask, int nW,
                              int B, int H, int S, int d, float scale, b200dp_dtype dtype, void* stream, const WinArgs* win) {
    B200DP_REQUIRE(qkv && out && lse, "attention_fwd: null pointer");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "attention_fwd: fp16 / bf16 only");
    Geom g;
    int rc = make_geom(&g, B, H, S, d, scale, mask ? nW : 0, qkv, out);
    if (rc) return rc;
    if (win) { rc = set_windows(&g, win->Hi, win->Wi, win->ws, win->shift, win->qp, win->op, win->op); if (rc) return rc; }
    const int is_bf16 = dtype == B200DP_BF16;
    CUtensorMap tq, tkv;
    memset(&tq, 0, sizeof(tq)); memset(&tkv, 0, sizeof(tkv));
    if (!g.manual) {
        bind_context_once();
        rc = make_tmap3(&tq, qkv, is_bf16, (int64_t)3 * H * d, S, B, kTile);
        if (rc) return rc;
        rc = make_tmap3(&tkv, qkv, is_bf16, (int64_t)3 * H * d, S, B, g.nk);
        if (rc) return rc;
    }
    const dim3 grid(g.nq, H, (B + g.G - 1) / g.G);
    cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH(TT)                                                                                                      \
    {                                                                                                                   \
        auto kern = attn_fwd_kernel<TT>;                                                                                \
        static bool configured = false;                                                                                 \
        if (!configured) {                                                                                              \
            B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFwdSmem));             \
            configured = true;                                                                                          \
        }                                                                                                               \
        kern<<<grid, 256, kFwdSmem, s>>>(tq, tkv, (const TT*)qkv, (TT*)out, lse, bias, mask, g, is_bf16);               \
    }
    if (is_bf16) LAUNCH(__nv_bfloat16) else LAUNCH(__half)
#undef LAUNCH
    B200DP_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200dp_attention_fwd(const void* qkv, void* out, float* lse, const float* bias, const float* mask, int nW,
                                    int B, int H, int S, int d, float scale, b200dp_dtype dtype, void* stream) {
    return attention_fwd_impl(qkv, out, lse, bias, mask, nW, B, H, S, d, scale, dtype, stream, nullptr);
}

// Window attention straight on the image-ordered token grid: qkv is [images][Hi][Wi][qkv_pitch] (first 3*H*d elements of
// a row are q|k|v), out [images][Hi][Wi][out_pitch]; the cyclic shift and the cut into ws x ws windows happen in the
// kernel's addressing.  lse / bias / mask / dbias keep the window-major layouts of the plain entry points
// (B = images * windows-per-image problems of S = ws * ws tokens; mask index = window index inside its image).
extern "C" int b200dp_window_attention_fwd(const void* qkv, void* out, float* lse, const float* bias, const float* mask,
                                           int images, int Hi, int Wi, int ws, int shift, int H, int d, int qkv_pitch,
                                           int out_pitch, float scale, b200dp_dtype dtype, void* stream) {
    B200DP_REQUIRE(ws > 0 && Hi > 0 && Wi > 0 && Hi % ws == 0 && Wi % ws == 0, "window_attention_fwd: bad window geometry");
    const int wpi = (Hi / ws) * (Wi / ws);
    const WinArgs w{Hi, Wi, ws, shift, qkv_pitch, out_pitch, out_pitch};
    return attention_fwd_impl(qkv, out, lse, bias, mask, wpi, images * wpi, H, ws * ws, d, scale, dtype, stream, &w);
}

// dqkv[B][S][3][H][d] from dout / out [B][S][H][d], the packed qkv and lse.  dbias ([H][S][S] fp32, S <= 64 only)
// is ACCUMULATED into (atomic adds; the caller zeroes it); pass NULL when the bias needs no gradient.
static int attention_bwd_impl(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                              const float* bias, const float* mask, int nW, float* dbias, int B, int H, int S, int d,
                              float scale, b200dp_dtype dtype, void* stream, const WinArgs* win) {
    B200DP_REQUIRE(qkv && out && dout && lse && dqkv, "attention_bwd: null pointer");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "attention_bwd: fp16 / bf16 only");
    B200DP_REQUIRE(dbias == nullptr || S <= 64, "attention_bwd: the bias gradient is implemented for S <= 64 (windows)");
    Geom g;
    int rc = make_geom(&g, B, H, S, d, scale, mask ? nW : 0, qkv, dout);
    if (rc) return rc;
    if (((uintptr_t)dqkv & 15) != 0 || ((uintptr_t)out & 15) != 0) g.manual = 1;
    if (win) { rc = set_windows(&g, win->Hi, win->Wi, win->ws, win->shift, win->qp, win->op, win->dop); if (rc) return rc; }
    const int is_bf16 = dtype == B200DP_BF16;
    CUtensorMap tq, tdo;
    memset(&tq, 0, sizeof(tq)); memset(&tdo, 0, sizeof(tdo));
    if (!g.manual) {
        bind_context_once();
        rc = make_tmap3(&tq, qkv, is_bf16, (int64_t)3 * H * d, S, B, kTile);
        if (rc) return rc;
        rc = make_tmap3(&tdo, dout, is_bf16, (int64_t)H * d, S, B, kTile);
        if (rc) return rc;
    }
    const int nunits = (B + g.G - 1) / g.G;
    int gx = nunits;
    if (dbias != nullptr) {
        // persistent CTAs, exactly one wave (one CTA per SM: the kernel owns all 512 TMEM columns): each walks its share
        // of the windows of one head and flushes its bias-gradient partial once
        gx = B200DP_SMS / H;
        if (gx < 1) gx = 1;
        if (gx > nunits) gx = nunits;
    }
    const dim3 grid(gx, H, 1);
    cudaStream_t s = (cudaStream_t)stream;
#define LAUNCH(TT)                                                                                                      \
    {                                                                                                                   \
        auto kern = attn_bwd_kernel<TT>;                                                                                \
        static bool configured = false;                                                                                 \
        if (!configured) {                                                                                              \
            B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kBwdSmem));             \
            configured = true;                                                                                          \
        }                                                                                                               \
        kern<<<grid, 256, kBwdSmem, s>>>(tq, tdo, (const TT*)qkv, (const TT*)out, (const TT*)dout, lse, (TT*)dqkv, bias, \
                                         mask, dbias, g, is_bf16);                                                      \
    }
    if (is_bf16) LAUNCH(__nv_bfloat16) else LAUNCH(__half)
#undef LAUNCH
    B200DP_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200dp_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                    const float* bias, const float* mask, int nW, float* dbias, int B, int H, int S, int d,
                                    float scale, b200dp_dtype dtype, void* stream) {
    return attention_bwd_impl(qkv, out, dout, lse, dqkv, bias, mask, nW, dbias, B, H, S, d, scale, dtype, stream, nullptr);
}

// Backward of b200dp_window_attention_fwd.  dqkv has the layout (and pitch) of qkv; only its first 3*H*d elements per
// token are written, so the caller zeroes any padding columns.  dout may have its own pitch.
extern "C" int b200dp_window_attention_bwd(const void* qkv, const void* out, const void* dout, const float* lse, void* dqkv,
                                           const float* bias, const float* mask, float* dbias, int images, int Hi, int Wi,
                                           int ws, int shift, int H, int d, int qkv_pitch, int out_pitch, int dout_pitch,
                                           float scale, b200dp_dtype dtype, void* stream) {
    B200DP_REQUIRE(ws > 0 && Hi > 0 && Wi > 0 && Hi % ws == 0 && Wi % ws == 0, "window_attention_bwd: bad window geometry");
    const int wpi = (Hi / ws) * (Wi / ws);
    const WinArgs w{Hi, Wi, ws, shift, qkv_pitch, out_pitch, dout_pitch};
    return attention_bwd_impl(qkv, out, dout, lse, dqkv, bias, mask, wpi, dbias, images * wpi, H, ws * ws, d, scale, dtype,
                              stream, &w);
}


