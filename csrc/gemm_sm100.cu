// bf16 / fp16 GEMM on the 5th-generation tensor cores (W7 for 1x1 convolutions, K17 for
// linears, and the compute half of W4).
//
//   C[M, N] (+bias[N]) = A(M x K) * B(N x K)^T,   fp32 accumulation in TMEM
//
// Each operand may be "K-major" (the reduction index is the contiguous one in global memory)
// or "MN-major" (the output index is contiguous).  That covers forward, dgrad and wgrad of a
// linear layer with no transposes:
//     forward  Y[M,N]  = X[M,K]  W[N,K]^T     A = X  (K-major),   B = W  (K-major)
//     dgrad    dX[M,K] = dY[M,N] W[N,K]       A = dY (K-major),   B = W  (MN-major)
//     wgrad    dW[N,K] = dY[M,N]^T X[M,K]     A = dY (MN-major),  B = X  (MN-major)
//
// Structure (one CTA per SM, persistent over 128 x BN output tiles):
//   warp 0      TMA producer: cp.async.bulk.tensor into a kStages-deep ring of 128B-swizzled
//               smem tiles, completion signalled on "full" mbarriers (complete_tx bytes)
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma (M=128, N=BN, K=16) reading smem
//               through UMMA descriptors, accumulating into one of two TMEM buffers;
//               tcgen05.commit releases the smem stage ("empty") and finally hands the
//               accumulator to the epilogue ("tmem_full")
//   warp 2      allocates / frees the 2*BN TMEM columns
//   warps 4-7   epilogue: tcgen05.ld (32 lanes x 32 columns per warp per step) -> bias -> pack ->
//               128B-swizzled smem staging (double-buffered per warp) -> TMA store (or TMA
//               reduce-add when accumulating into C); releases the accumulator ("tmem_empty")
//               as soon as the tile has left TMEM
// The two accumulator buffers let the epilogue of tile i overlap the MMAs of tile i+1.
// Ragged M / N / K edges: TMA zero-fills out-of-bounds reads, the epilogue masks the stores.
#include "tc_common.cuh"

namespace b200dp {
namespace gemm {

using namespace b200dp::tc;

template <int BN> struct Cfg {
    static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 5 : 7);
    static constexpr int kABytes = kBM * kBK * 2;   // 16 KB
    static constexpr int kBBytes = BN * kBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kEpiBufBytes = 32 * 128;   // 32 rows x 128 B per staging buffer
    static constexpr int kEpiBytes = 4 * 2 * kEpiBufBytes;   // 4 warps x double buffer
    static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 /*align*/ + 256 /*barriers*/;
    static constexpr int kTmemCols = 2 * BN;        // two accumulators (power of two: 128 / 256 / 512)
};

// kStats: the epilogue also accumulates, per output column, the sum and the sum of squares of the
// (rounded) values it stores — the batch statistics the BatchNorm that follows a 1x1 convolution
// needs — so that layer does not have to read the activation once more (colstats[0:N] += sum,
// colstats[N:2N] += sum of squares; per-lane register partials, one red.add per column per change
// of column tile).
template <int BN, bool kAMN, bool kBMN, typename TOut, bool kStats = false>
__global__ void __launch_bounds__(kThreads, 1) gemm_kernel(const __grid_constant__ CUtensorMap tmA,
                                                           const __grid_constant__ CUtensorMap tmB,
                                                           const __grid_constant__ CUtensorMap tmC,
                                                           const float* __restrict__ bias, int M, int N, int K,
                                                           int is_bf16, int accumulate_c, int splits,
                                                           float* __restrict__ colstats) {
    using cfg = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smemA = smem;
    uint8_t* smemB = smem + cfg::kStages * cfg::kABytes;
    uint8_t* smemC = smem + cfg::kStages * cfg::kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemC + cfg::kEpiBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + cfg::kStages;
    uint64_t* tmem_full = bars + 2 * cfg::kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_m = (M + kBM - 1) / kBM, num_n = (N + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int num_kb = (K + kBK - 1) / kBK;
    // split-K: work item = (tile, split); split s reduces k-blocks [s*kb_per, min(num_kb, (s+1)*kb_per))
    // and is added into C by TMA reduce-add (C pre-zeroed by the host unless accumulating)
    const int kb_per = (num_kb + splits - 1) / splits;
    const int num_items = num_tiles * splits;

    if (warp == 0 && lane == 0) {
        tma_prefetch_desc(&tmA);
        tma_prefetch_desc(&tmB);
        tma_prefetch_desc(&tmC);
    }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < cfg::kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 4); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, cfg::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                const int tile = item % num_tiles, split = item / num_tiles;
                const int m0 = (tile % num_m) * kBM, n0 = (tile / num_m) * BN;
                const int kb_begin = split * kb_per, kb_end = min(num_kb, kb_begin + kb_per);
                for (int kb = kb_begin; kb < kb_end; ++kb) {
                    mbar_wait(empty + stage, phase ^ 1);
                    mbar_expect_tx(full + stage, cfg::kStageBytes);
                    uint8_t* a = smemA + stage * cfg::kABytes;
                    uint8_t* b = smemB + stage * cfg::kBBytes;
                    const int k0 = kb * kBK;
                    if (kAMN) {
#pragma unroll
                        for (int j = 0; j < kBM / 64; ++j) tma_load_2d(a + j * (kBK * 128), &tmA, full + stage, m0 + j * 64, k0);
                    } else {
                        tma_load_2d(a, &tmA, full + stage, k0, m0);
                    }
                    if (kBMN) {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j) tma_load_2d(b + j * (kBK * 128), &tmB, full + stage, n0 + j * 64, k0);
                    } else {
                        tma_load_2d(b, &tmB, full + stage, k0, n0);
                    }
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, kAMN ? 1 : 0, kBMN ? 1 : 0, kBM, BN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                const int split = item / num_tiles;
                const int kb_begin = split * kb_per, kb_end = min(num_kb, kb_begin + kb_per);
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = kb_begin; kb < kb_end; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smemA + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smemB + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k) {
                        const uint64_t da = kAMN ? make_desc(a_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                 : make_desc(a_addr + k * (kUmmaK * 2), 16, 1024);
                        const uint64_t db = kBMN ? make_desc(b_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                 : make_desc(b_addr + k * (kUmmaK * 2), 16, 1024);
                        umma_f16(d_tmem, da, db, idesc, (kb != kb_begin || k != 0) ? 1u : 0u);
                    }
                    umma_commit(empty + stage);            // smem stage reusable once these MMAs retire
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full + acc);              // accumulator complete
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===== epilogue =====
        const int quarter = warp - kEpiWarp0;              // TMEM lanes [32q, 32q+32) belong to warp q (mod 4)
        constexpr int kChunkCols = 128 / (int)sizeof(TOut);     // columns per 128-byte staging row
        constexpr int kLoadsPerChunk = kChunkCols / 32;         // tcgen05.ld.x32 per chunk (2 for 16-bit, 1 for fp32)
        uint8_t* my_stage = smemC + quarter * 2 * cfg::kEpiBufBytes;
        int acc = 0; uint32_t acc_phase = 0;
        uint32_t buf_idx = 0;
        constexpr int kChunks = BN / kChunkCols;
        // column statistics: lane l owns columns 2l, 2l+1 of every 64-column chunk of the current column tile
        float st[kStats ? kChunks * 4 : 1];
        int st_n0 = -1;
#pragma unroll
        for (int i = 0; i < (kStats ? kChunks * 4 : 1); ++i) st[i] = 0.f;
        auto flush_stats = [&]() {
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) {
                    const int col = st_n0 + ch * kChunkCols + 2 * lane;
                    if (col < N) {
                        atomicAdd(colstats + col, st[ch * 4 + 0]);
                        atomicAdd(colstats + N + col, st[ch * 4 + 2]);
                    }
                    if (col + 1 < N) {
                        atomicAdd(colstats + col + 1, st[ch * 4 + 1]);
                        atomicAdd(colstats + N + col + 1, st[ch * 4 + 3]);
                    }
                    st[ch * 4 + 0] = st[ch * 4 + 1] = st[ch * 4 + 2] = st[ch * 4 + 3] = 0.f;
                }
            }
        };
        for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
            const int tile = item % num_tiles, split = item / num_tiles;
            const int m0 = (tile % num_m) * kBM, n0 = (tile / num_m) * BN;
            const bool add_bias = bias != nullptr && split == 0;
            if (kStats && n0 != st_n0) {
                if (st_n0 >= 0) flush_stats();
                st_n0 = n0;
            }
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            const int row0 = m0 + quarter * 32;
            auto do_chunk = [&](const int ch) {
                uint8_t* buf = my_stage + (buf_idx & 1) * cfg::kEpiBufBytes;
                // the TMA store that last read this buffer (two chunks ago) must have drained it
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
#pragma unroll
                for (int l = 0; l < kLoadsPerChunk; ++l) {
                    uint32_t r[32];
                    const int colbase = ch * kChunkCols + l * 32;
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN + colbase, r);
                    tmem_ld_wait();
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (add_bias) {
                        const int gcol = n0 + colbase;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (gcol + j < N) ? __ldg(bias + gcol + j) : 0.f;
                    }
                    // 128-byte swizzle: 16-byte chunk j of row r lives at chunk (j ^ (r & 7))
                    uint8_t* rowp = buf + lane * 128;
                    if (sizeof(TOut) == 2) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float t[8] = {v[q * 8], v[q * 8 + 1], v[q * 8 + 2], v[q * 8 + 3],
                                          v[q * 8 + 4], v[q * 8 + 5], v[q * 8 + 6], v[q * 8 + 7]};
                            const int j = l * 4 + q;
                            store8<TOut>(reinterpret_cast<TOut*>(rowp + ((j ^ (lane & 7)) << 4)), t);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(rowp + ((j ^ (lane & 7)) << 4)) =
                                make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
                    }
                }
                fence_async_smem();        // generic-proxy smem writes -> visible to the TMA engine
                __syncwarp();
                if (lane == 0 && n0 + ch * kChunkCols < N && row0 < M) {
                    if (accumulate_c || splits > 1) tma_reduce_add_2d(&tmC, buf, n0 + ch * kChunkCols, row0);
                    else tma_store_2d(&tmC, buf, n0 + ch * kChunkCols, row0);
                }
                if (lane == 0) tma_store_commit();
                if constexpr (kStats && sizeof(TOut) == 2) {
                    // column sums over this warp's 32 staged rows (rows past M hold zeros): lane l reads
                    // the 4-byte word of columns 2l, 2l+1 in every row — one full 128-byte row per
                    // request, no bank conflicts under the swizzle
                    const int j = lane >> 2, w = lane & 3;
                    float s0 = 0.f, s1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll 8
                    for (int rr = 0; rr < 32; ++rr) {
                        const uint32_t u = *reinterpret_cast<const uint32_t*>(buf + rr * 128 + ((j ^ (rr & 7)) << 4) + w * 4);
                        const float2 f = unpack2<TOut>(u);
                        s0 += f.x; s1 += f.y;
                        q0 = fmaf(f.x, f.x, q0); q1 = fmaf(f.y, f.y, q1);
                    }
                    st[ch * 4 + 0] += s0; st[ch * 4 + 1] += s1; st[ch * 4 + 2] += q0; st[ch * 4 + 3] += q1;
                }
                ++buf_idx;
            };
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            } else {
#pragma unroll 1
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            }
            // the accumulator has left TMEM: hand it back to the MMA warp before the stores finish
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + acc);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (kStats && st_n0 >= 0) flush_stats();
        if (lane == 0) tma_store_wait_read<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, cfg::kTmemCols);
}

// ---------------------------------------------------------------- CTA-pair kernel (cta_group::2)
// C[M,N] = A[M,K] B[N,K]^T (+bias), K-major operands, 16-bit output.  Two CTAs of a cluster (one TPC)
// work on one 256 x 256 output tile: each loads ITS 128 rows of A and ITS 128 rows of B per k-block
// (32 KB instead of 48 KB per 128 x 256 x 64 of math — operand traffic from L2 and shared-memory
// reads per FLOP both drop by a third), the leader's MMA thread issues one M = 256 tcgen05.mma for
// the pair, and each CTA's tensor core accumulates its own 128 rows in its own TMEM.
//   full[stage]      leader only; armed by the leader's producer with the pair's 64 KB, credited by both CTAs' TMA
//   empty[stage]     one per CTA; the pair MMA's commit arrives on both
//   tmem_full[acc]   one per CTA; commit multicast
//   tmem_empty[acc]  leader only; the 16 epilogue warps of the pair arrive on it
namespace pair {
constexpr int kBN = 256;
constexpr int kHalfN = 128;
constexpr int kStages = 5;
constexpr int kABytes = kBM * kBK * 2;
constexpr int kBBytes = kHalfN * kBK * 2;
constexpr int kStageBytes = kABytes + kBBytes;
constexpr int kEpiBufBytes = 32 * 128;
// EIGHT epilogue warps per CTA: two per TMEM lane quarter, splitting the column chunks between them.  With
// short reductions (K = 768: ~6k MMA cycles per tile) four warps cannot drain a 128 x 256 accumulator in
// the time the tensor core needs for the next one; eight can, also with a GELU in the epilogue.
constexpr int kEpiWarps = 8;
constexpr int kPairThreads = (kEpiWarp0 + kEpiWarps) * 32;
constexpr int kEpiBytes = kEpiWarps * 2 * kEpiBufBytes;
constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
constexpr int kTmemCols = 512;
}  // namespace pair

// Epilogue variants (16-bit outputs), all applied to the fp32 accumulator before it is rounded:
//   kEpiGeluDual  C = acc (+bias) AND C2 = gelu(C): the pre-activation the backward needs and the activation
//                 the next layer reads leave the same epilogue (exact erf GELU)
//   kEpiDGelu     C = acc * gelu'(aux): the input gradient of the layer after a GELU becomes the gradient of
//                 the pre-activation on the way out (aux = saved pre-activation, [M][ld_aux])
//   kEpiAddAux    C = acc (+bias) + aux: residual add
enum : int { kEpiNone = 0, kEpiGeluDual = 1, kEpiDGelu = 2, kEpiAddAux = 3 };

// GELU pieces by Abramowitz-Stegun 7.1.26 (|erf error| < 1.5e-7, far below 16-bit output rounding) with the
// hardware approximations for the reciprocal and the exponential: ~17 instructions per element, no branches
// (erff is ~40 with a slow path).  With x = |z| / sqrt(2):  w = poly(t) * exp(-x^2), t = 1 / (1 + p x), and
//   Phi(z) = 1 - w/2 (z >= 0),  w/2 (z < 0);   gelu(z) = z Phi(z);   gelu'(z) = Phi(z) + z * pdf(z),
// where exp(-x^2) = exp(-z^2 / 2) is also the Gaussian factor of pdf(z).
__device__ __forceinline__ float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ float fast_ex2(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ void gelu_parts(float z, float& half_w, float& gauss) {
    const float t = fast_rcp(fmaf(0.3275911f * 0.70710678118654752f, fabsf(z), 1.f));
    gauss = fast_ex2(z * z * (-0.5f * 1.4426950408889634f));
    float p = fmaf(t, 1.061405429f, -1.453152027f);
    p = fmaf(p, t, 1.421413741f);
    p = fmaf(p, t, -0.284496736f);
    p = fmaf(p, t, 0.254829592f);
    half_w = 0.5f * p * t * gauss;
}
__device__ __forceinline__ float gelu_f(float z) {
    float hw, g;
    gelu_parts(z, hw, g);
    return z >= 0.f ? fmaf(-z, hw, z) : z * hw;
}
__device__ __forceinline__ float dgelu_f(float z) {
    float hw, g;
    gelu_parts(z, hw, g);
    const float phi = z >= 0.f ? 1.f - hw : hw;
    return fmaf(z * 0.39894228040143268f, g, phi);
}

template <bool kAMN, bool kBMN, typename TOut, int kEpi = kEpiNone>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(pair::kPairThreads, 1)
gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmC, const float* __restrict__ bias, int M, int N, int K, int is_bf16,
                 int accumulate_c, int splits, const __grid_constant__ CUtensorMap tmC2, const TOut* __restrict__ aux,
                 int ld_aux) {
    using namespace pair;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smemA = smem;
    uint8_t* smemB = smem + kStages * kABytes;
    uint8_t* smemC = smem + kStages * kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemC + kEpiBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + kStages;
    uint64_t* tmem_full = bars + 2 * kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta_rank = cluster_ctarank();
    const bool leader = cta_rank == 0;
    const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    const int num_m = (M + 2 * kBM - 1) / (2 * kBM), num_n = (N + kBN - 1) / kBN;
    const int num_tiles = num_m * num_n;
    const int num_kb = (K + kBK - 1) / kBK;
    // split-K as in gemm_kernel: item = (tile, split), partial products combined by TMA reduce-add
    const int kb_per = (num_kb + splits - 1) / splits;
    const int num_items = num_tiles * splits;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); tma_prefetch_desc(&tmC); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 2 * kEpiWarps); }
        fence_barrier_init();
    }
    cluster_sync_all();                       // both CTAs' barriers exist before anything remote touches them
    if (warp == 2) tmem_alloc_pair(tmem_slot, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer (both CTAs) =====
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int item = pair_id; item < num_items; item += num_pairs) {
                const int tile = item % num_tiles, split = item / num_tiles;
                const int m0 = (tile % num_m) * (2 * kBM) + (int)cta_rank * kBM;
                const int n0 = (tile / num_m) * kBN + (int)cta_rank * kHalfN;
                const int kb_begin = split * kb_per, kb_end = min(num_kb, kb_begin + kb_per);
                for (int kb = kb_begin; kb < kb_end; ++kb) {
                    mbar_wait(empty + stage, phase ^ 1);
                    if (leader) mbar_expect_tx(full + stage, 2 * kStageBytes);
                    uint8_t* a = smemA + stage * kABytes;
                    uint8_t* b = smemB + stage * kBBytes;
                    const int k0 = kb * kBK;
                    if (kAMN) {
#pragma unroll
                        for (int j = 0; j < kBM / 64; ++j) tma_load_2d_pair(a + j * (kBK * 128), &tmA, full + stage, m0 + j * 64, k0);
                    } else {
                        tma_load_2d_pair(a, &tmA, full + stage, k0, m0);
                    }
                    if (kBMN) {
#pragma unroll
                        for (int j = 0; j < kHalfN / 64; ++j) tma_load_2d_pair(b + j * (kBK * 128), &tmB, full + stage, n0 + j * 64, k0);
                    } else {
                        tma_load_2d_pair(b, &tmB, full + stage, k0, n0);
                    }
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (leader CTA only) =====
        if (leader && lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, kAMN ? 1 : 0, kBMN ? 1 : 0, 2 * kBM, kBN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int item = pair_id; item < num_items; item += num_pairs) {
                const int split = item / num_tiles;
                const int kb_begin = split * kb_per, kb_end = min(num_kb, kb_begin + kb_per);
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * kBN;
                for (int kb = kb_begin; kb < kb_end; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smemA + stage * kABytes);
                    const uint32_t b_addr = smem_u32(smemB + stage * kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k) {
                        const uint64_t da = kAMN ? make_desc(a_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                 : make_desc(a_addr + k * (kUmmaK * 2), 16, 1024);
                        const uint64_t db = kBMN ? make_desc(b_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                 : make_desc(b_addr + k * (kUmmaK * 2), 16, 1024);
                        umma_f16_pair(d_tmem, da, db, idesc, (kb != kb_begin || k != 0) ? 1u : 0u);
                    }
                    umma_commit_pair(empty + stage, 3);          // both CTAs' producers may refill this stage
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit_pair(tmem_full + acc, 3);            // both CTAs' epilogues may read their half
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // ===== epilogue (both CTAs; each drains its own 128 rows) =====
        const int ew = warp - kEpiWarp0;                   // 0..7
        const int quarter = ew & 3;                        // TMEM lanes [32q, 32q+32) are reachable from warps with warp % 4 == q
        const int half = ew >> 2;                          // which of the two warps of this lane quarter: even / odd chunks
        constexpr int kChunkCols = 128 / (int)sizeof(TOut);
        constexpr int kLoadsPerChunk = kChunkCols / 32;
        uint8_t* my_stage = smemC + ew * 2 * kEpiBufBytes;
        int acc = 0; uint32_t acc_phase = 0;
        uint32_t buf_idx = 0;
        for (int item = pair_id; item < num_items; item += num_pairs) {
            const int tile = item % num_tiles, split = item / num_tiles;
            const int m0 = (tile % num_m) * (2 * kBM) + (int)cta_rank * kBM;
            const int n0 = (tile / num_m) * kBN;
            const bool add_bias = bias != nullptr && split == 0;
            const bool has_k = split * kb_per < num_kb;
            const int row0 = m0 + quarter * 32;
            if constexpr (kEpi == kEpiDGelu || kEpi == kEpiAddAux) {
                // pull this lane's aux rows for the tile towards L2 while the tensor core is still busy
                if (row0 + lane < M) {
                    for (int ch = half; ch < kBN / kChunkCols; ch += 2)
                        if (n0 + ch * kChunkCols < N)
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(aux + (size_t)(row0 + lane) * ld_aux + n0 + ch * kChunkCols));
                }
            }
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
#pragma unroll 1
            for (int ch = half; ch < kBN / kChunkCols; ch += 2) {
                uint8_t* buf = my_stage + (buf_idx & 1) * kEpiBufBytes;
                uint8_t* buf2 = my_stage + ((buf_idx & 1) ^ 1) * kEpiBufBytes;      // kEpiGeluDual: second output
                if (lane == 0) { if (kEpi == kEpiGeluDual) tma_store_wait_read<0>(); else tma_store_wait_read<1>(); }
                __syncwarp();
                if constexpr (kEpi == kEpiDGelu || kEpi == kEpiAddAux) {
                    // the aux chunk (32 rows x 128 B) goes straight into the staging buffer with 16-byte async
                    // copies — 8 lanes per row, fully coalesced, same swizzle as the output — while the
                    // accumulator is still on its way out of TMEM
#pragma unroll
                    for (int rr = 0; rr < 32; rr += 4) {
                        const int row = rr + (lane >> 3), c = lane & 7;
                        const int grow = row0 + row, gcol = n0 + ch * kChunkCols + c * 8;
                        if (grow < M && gcol < N)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(buf + row * 128 + ((c ^ (row & 7)) << 4))),
                                         "l"(aux + (size_t)grow * ld_aux + gcol) : "memory");
                    }
                    asm volatile("cp.async.commit_group;" ::: "memory");
                }
#pragma unroll
                for (int l = 0; l < kLoadsPerChunk; ++l) {
                    uint32_t r[32];
                    const int colbase = ch * kChunkCols + l * 32;
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * kBN + colbase, r);
                    tmem_ld_wait();
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (add_bias) {
                        const int gcol = n0 + colbase;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (gcol + j < N) ? __ldg(bias + gcol + j) : 0.f;
                    }
                    if constexpr (kEpi == kEpiDGelu || kEpi == kEpiAddAux) {
                        if (l == 0) {
                            asm volatile("cp.async.wait_group 0;" ::: "memory");
                            __syncwarp();
                        }
                        // this lane's row of the aux chunk, read back from the staging buffer (rows past M / columns
                        // past N hold stale bytes: their results are clipped by the TMA store)
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float a8[8];
                            load8<TOut>(reinterpret_cast<const TOut*>(buf + lane * 128 + (((l * 4 + q) ^ (lane & 7)) << 4)), a8);
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                if (kEpi == kEpiDGelu) v[q * 8 + e] *= dgelu_f(a8[e]);
                                else v[q * 8 + e] += a8[e];
                            }
                        }
                    }
                    uint8_t* rowp = buf + lane * 128;
                    if (sizeof(TOut) == 2) {
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            float t[8] = {v[q * 8], v[q * 8 + 1], v[q * 8 + 2], v[q * 8 + 3],
                                          v[q * 8 + 4], v[q * 8 + 5], v[q * 8 + 6], v[q * 8 + 7]};
                            store8<TOut>(reinterpret_cast<TOut*>(rowp + (((l * 4 + q) ^ (lane & 7)) << 4)), t);
                            if constexpr (kEpi == kEpiGeluDual) {
#pragma unroll
                                for (int e = 0; e < 8; ++e) t[e] = gelu_f(t[e]);
                                store8<TOut>(reinterpret_cast<TOut*>(buf2 + lane * 128 + (((l * 4 + q) ^ (lane & 7)) << 4)), t);
                            }
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            *reinterpret_cast<float4*>(rowp + ((j ^ (lane & 7)) << 4)) =
                                make_float4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
                    }
                }
                fence_async_smem();
                __syncwarp();
                if (lane == 0 && n0 + ch * kChunkCols < N && row0 < M && has_k) {
                    if (accumulate_c || splits > 1) tma_reduce_add_2d(&tmC, buf, n0 + ch * kChunkCols, row0);
                    else tma_store_2d(&tmC, buf, n0 + ch * kChunkCols, row0);
                    if (kEpi == kEpiGeluDual) tma_store_2d(&tmC2, buf2, n0 + ch * kChunkCols, row0);
                }
                if (lane == 0) tma_store_commit();
                ++buf_idx;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(tmem_empty + acc);     // the leader's MMA thread counts all 16 warps
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (lane == 0) tma_store_wait_read<0>();
    }
    tc_fence_before();
    cluster_sync_all();                       // neither CTA frees TMEM / exits while its peer may still use it
    if (warp == 2) tmem_dealloc_pair(tmem_base, kTmemCols);
}

// ---------------------------------------------------------------- host: tensor maps
static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn =
        reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(driver_fn("cuTensorMapEncodeTiled"));
    return fn;
}

// 2-D row-major tensor [rows][cols] of 16-bit elements, row pitch ld elements; box = {box_cols, box_rows}
static int make_tmap(CUtensorMap* tm, const void* ptr, int is_bf16, int64_t rows, int64_t cols, int64_t ld, int box_cols,
                     int box_rows, int elem_bytes = 2) {
    auto enc = get_encode();
    B200DP_REQUIRE(enc != nullptr, "gemm: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * elem_bytes};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    const CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                   : (is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
    CUresult r = enc(tm, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r == CUDA_ERROR_INVALID_CONTEXT || r == CUDA_ERROR_NOT_INITIALIZED) {
        // called from a thread that has only used the runtime API lazily (e.g. an autograd worker):
        // bind the primary context of the current device to this thread, then retry
        cudaFree(nullptr);
        r = enc(tm, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    B200DP_REQUIRE(r == CUDA_SUCCESS, "gemm: cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld", (int)r,
                   (long long)rows, (long long)cols, (long long)ld);
    return 0;
}

template <int BN, bool kAMN, bool kBMN, typename TOut, bool kStats = false>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const float* bias, int M, int N, int K,
                  int is_bf16, int accumulate_c, int splits, int sms, cudaStream_t s, float* colstats = nullptr) {
    auto kern = gemm_kernel<BN, kAMN, kBMN, TOut, kStats>;
    static bool configured = false;
    if (!configured) {
        B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmemBytes));
        configured = true;
    }
    const int items = ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN) * splits;
    const int grid = items < sms ? items : sms;
    kern<<<grid, kThreads, Cfg<BN>::kSmemBytes, s>>>(ta, tb, tc, bias, M, N, K, is_bf16, accumulate_c, splits, colstats);
    B200DP_CHECK_LAUNCH();
    return 0;
}

template <bool kAMN, bool kBMN, typename TOut, int kEpi = kEpiNone>
static int launch_pair(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const float* bias, int M, int N,
                       int K, int is_bf16, int accumulate_c, int splits, int sms, cudaStream_t s,
                       const CUtensorMap* tc2 = nullptr, const void* aux = nullptr, int ld_aux = 0) {
    auto kern = gemm_pair_kernel<kAMN, kBMN, TOut, kEpi>;
    static bool configured = false;
    if (!configured) {
        B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pair::kSmemBytes));
        configured = true;
    }
    const int items = ((M + 2 * kBM - 1) / (2 * kBM)) * ((N + pair::kBN - 1) / pair::kBN) * splits;
    int pairs = sms / 2;
    if (items < pairs) pairs = items;
    kern<<<2 * pairs, pair::kPairThreads, pair::kSmemBytes, s>>>(ta, tb, tc, bias, M, N, K, is_bf16, accumulate_c, splits,
                                                       tc2 ? *tc2 : tc, (const TOut*)aux, ld_aux);
    B200DP_CHECK_LAUNCH();
    return 0;
}

// B200DP_GEMM_PAIR=0 turns the cta_group::2 kernel off (A/B measurements)
static bool pair_enabled() {
    static const bool on = [] { const char* e = getenv("B200DP_GEMM_PAIR"); return !(e && e[0] == '0'); }();
    return on;
}

}  // namespace gemm
}  // namespace b200dp

using namespace b200dp;
using namespace b200dp::gemm;

// C[M,N] (+)= A * B^T (+ bias).  a_mn / b_mn: 0 = operand is K-major ([rows][K], K contiguous),
// 1 = MN-major ([K][rows], rows contiguous).  lda / ldb are the row pitches (elements) of the
// arrays as stored.  out_dtype: B200DP_F16 / B200DP_BF16 (same as the inputs) or B200DP_F32.
static int gemm_impl(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb, int ldc,
                     int a_mn, int b_mn, b200dp_dtype in_dtype, b200dp_dtype out_dtype, int accumulate_c, float* colstats,
                     void* stream) {
    B200DP_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "gemm: bad arguments");
    B200DP_REQUIRE(in_dtype == B200DP_F16 || in_dtype == B200DP_BF16, "gemm: inputs must be fp16 or bf16");
    B200DP_REQUIRE(out_dtype == in_dtype || out_dtype == B200DP_F32, "gemm: output must match the inputs or be fp32");
    B200DP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0, "gemm: lda / ldb must be multiples of 8 elements (16-byte TMA strides)");
    B200DP_REQUIRE(((uintptr_t)A & 15) == 0 && ((uintptr_t)B & 15) == 0, "gemm: A and B must be 16-byte aligned");
    const int is_bf16 = in_dtype == B200DP_BF16;
    bind_context_once();
    B200DP_REQUIRE(((uintptr_t)C & 15) == 0 && (ldc * (out_dtype == B200DP_F32 ? 4 : 2)) % 16 == 0,
                   "gemm: C must be 16-byte aligned with a 16-byte row pitch (TMA store)");
    const int BN = N <= 64 ? 64 : ((N >= 256 && (int64_t)M * N >= (int64_t)148 * 128 * 256) ? 256 : 128);
    CUtensorMap ta, tb, tc;
    int rc;
    // large products: the CTA-pair kernel (256 x 256 tiles).  16-bit results need >= one tile per pair;
    // fp32 results (weight gradients: few tiles, long K) are split along K to give every pair work.
    if (colstats == nullptr && pair_enabled() && M >= 256 && N >= 256) {
        const int64_t tiles2 = (int64_t)((M + 255) / 256) * ((N + 255) / 256);
        const int num_kb2 = (K + kBK - 1) / kBK;
        int splits2 = 1;
        bool use = false;
        if (out_dtype == in_dtype) {
            use = tiles2 >= 74;              // accumulate_c: the epilogue's TMA store becomes a reduce-add
        } else {
            if (tiles2 < 74) {
                splits2 = (int)(74 / tiles2);
                if (splits2 > num_kb2 / 8) splits2 = num_kb2 / 8;
                if (splits2 < 1) splits2 = 1;
                const int kbp = (num_kb2 + splits2 - 1) / splits2;
                splits2 = (num_kb2 + kbp - 1) / kbp;
            }
            use = tiles2 * splits2 >= 40;
        }
        if (use) {
            const int cb = out_dtype == B200DP_F32 ? 4 : 2;
            rc = make_tmap(&tc, C, is_bf16, M, N, ldc, 128 / cb, 32, cb);
            if (rc) return rc;
            rc = a_mn ? make_tmap(&ta, A, is_bf16, K, M, lda, 64, kBK) : make_tmap(&ta, A, is_bf16, M, K, lda, kBK, kBM);
            if (rc) return rc;
            rc = b_mn ? make_tmap(&tb, B, is_bf16, K, N, ldb, 64, kBK) : make_tmap(&tb, B, is_bf16, N, K, ldb, kBK, pair::kHalfN);
            if (rc) return rc;
            cudaStream_t ps = (cudaStream_t)stream;
            if (splits2 > 1 && !accumulate_c) B200DP_CUDA(cudaMemsetAsync(C, 0, (size_t)M * ldc * 4, ps));
#define GP(AM, BM_, T) return launch_pair<AM, BM_, T>(ta, tb, tc, bias, M, N, K, is_bf16, accumulate_c, splits2, B200DP_SMS, ps)
#define GP_OUT(AM, BM_)                                          \
    if (out_dtype == B200DP_F32) { GP(AM, BM_, float); }          \
    else if (is_bf16) { GP(AM, BM_, __nv_bfloat16); }             \
    else { GP(AM, BM_, __half); }
            if (!a_mn && !b_mn) { GP_OUT(false, false) }
            else if (!a_mn && b_mn) { GP_OUT(false, true) }
            else if (a_mn && b_mn) { GP_OUT(true, true) }
            else { GP_OUT(true, false) }
#undef GP_OUT
#undef GP
        }
    }
    const int c_bytes = out_dtype == B200DP_F32 ? 4 : 2;
    rc = make_tmap(&tc, C, is_bf16, M, N, ldc, 128 / c_bytes, 32, c_bytes);   // staging rows are 128 B
    if (rc) return rc;
    // K-major operand [rows][K]: box = 64 (K) x tile rows.  MN-major [K][rows]: box = 64 (rows) x 64 (K).
    rc = a_mn ? make_tmap(&ta, A, is_bf16, K, M, lda, 64, kBK) : make_tmap(&ta, A, is_bf16, M, K, lda, kBK, kBM);
    if (rc) return rc;
    rc = b_mn ? make_tmap(&tb, B, is_bf16, K, N, ldb, 64, kBK) : make_tmap(&tb, B, is_bf16, N, K, ldb, kBK, BN);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int sms = B200DP_SMS;
    // split-K when the output has too few tiles to fill the GPU and the reduction is long (wgrad):
    // fp32 outputs only (partial sums are combined by TMA reduce-add, whose order is not fixed)
    int splits = 1;
    {
        const int tiles = ((M + kBM - 1) / kBM) * ((N + BN - 1) / BN);
        const int num_kb = (K + kBK - 1) / kBK;
        if (out_dtype == B200DP_F32 && tiles * 2 <= sms && num_kb >= 16) {
            splits = sms / tiles;
            if (splits > num_kb / 8) splits = num_kb / 8;     // at least 8 k-blocks per split
            if (splits < 1) splits = 1;
            const int kb_per = (num_kb + splits - 1) / splits;
            splits = (num_kb + kb_per - 1) / kb_per;          // no empty split (an empty one would add garbage)
        }
        if (splits > 1 && !accumulate_c)
            B200DP_CUDA(cudaMemsetAsync(C, 0, (size_t)M * ldc * 4, s));
    }
    if (colstats != nullptr) {
        // statistics variant: K-major operands, 16-bit output, one pass over K (checked by the caller)
#define GS(BNv)                                                                                                        \
    {                                                                                                                  \
        if (is_bf16) return launch<BNv, false, false, __nv_bfloat16, true>(ta, tb, tc, nullptr, M, N, K, 1, 0, 1, sms, s, colstats); \
        return launch<BNv, false, false, __half, true>(ta, tb, tc, nullptr, M, N, K, 0, 0, 1, sms, s, colstats);        \
    }
        if (BN == 256) GS(256) else if (BN == 128) GS(128) else GS(64)
#undef GS
    }
#define GO(BNv, AM, BM_, T) return launch<BNv, AM, BM_, T>(ta, tb, tc, bias, M, N, K, is_bf16, accumulate_c, splits, sms, s)
#define BY_OUT(BNv, AM, BM_)                                                      \
    if (out_dtype == B200DP_F32) { GO(BNv, AM, BM_, float); }                     \
    else if (is_bf16) { GO(BNv, AM, BM_, __nv_bfloat16); }                        \
    else { GO(BNv, AM, BM_, __half); }
#define BY_MAJOR(BNv)                                                             \
    if (!a_mn && !b_mn) { BY_OUT(BNv, false, false) }                             \
    else if (!a_mn && b_mn) { BY_OUT(BNv, false, true) }                          \
    else if (a_mn && b_mn) { BY_OUT(BNv, true, true) }                            \
    else { BY_OUT(BNv, true, false) }
    if (BN == 256) { BY_MAJOR(256) } else if (BN == 128) { BY_MAJOR(128) } else { BY_MAJOR(64) }
#undef BY_MAJOR
#undef BY_OUT
#undef GO
    return 0;
}

extern "C" int b200dp_gemm(const void* A, const void* B, void* C, const float* bias, int M, int N, int K, int lda, int ldb,
                           int ldc, int a_mn, int b_mn, b200dp_dtype in_dtype, b200dp_dtype out_dtype, int accumulate_c,
                           void* stream) {
    return gemm_impl(A, B, C, bias, M, N, K, lda, ldb, ldc, a_mn, b_mn, in_dtype, out_dtype, accumulate_c, nullptr, stream);
}

// C[M,N] = A[M,K] B[N,K]^T (16-bit in / out) and, from the same epilogue, colstats[0:N] += column sums
// of C, colstats[N:2N] += column sums of squares (fp32 atomics; the caller zeroes colstats).
extern "C" int b200dp_gemm_colstats(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                                    b200dp_dtype dtype, float* colstats, void* stream) {
    B200DP_REQUIRE(colstats != nullptr, "gemm_colstats: null statistics buffer");
    return gemm_impl(A, B, C, nullptr, M, N, K, lda, ldb, ldc, 0, 0, dtype, dtype, 0, colstats, stream);
}

// Large-product GEMM with a fused epilogue (CTA-pair kernel only; see the kEpi* modes above):
//   mode 1  C = A B^T (+bias), C2 = gelu(C)            (b_mn must be 0)
//   mode 2  C = (A B^T) * gelu'(aux)                    (dgrad form: b_mn = 1)
//   mode 3  C = A B^T (+bias) + aux                     (b_mn = 0)
// 16-bit operands and outputs, A K-major.  Returns B200DP_ERR_INVALID when the shape is below the pair
// kernel's envelope (M, N >= 256 and at least 74 output tiles of 256 x 256) — b200dp_gemm_fused_supported tells.
extern "C" int b200dp_gemm_fused_supported(int M, int N) {
    return pair_enabled() && M >= 256 && N >= 256 && (int64_t)((M + 255) / 256) * ((N + 255) / 256) >= 74;
}

extern "C" int b200dp_gemm_fused(const void* A, const void* B, void* C, void* C2, const float* bias, const void* aux, int M,
                                 int N, int K, int lda, int ldb, int ldc, int ldc2, int ld_aux, int b_mn, b200dp_dtype dtype,
                                 int mode, void* stream) {
    B200DP_REQUIRE(A && B && C && M > 0 && N > 0 && K > 0, "gemm_fused: bad arguments");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "gemm_fused: fp16 / bf16 only");
    B200DP_REQUIRE(mode >= 1 && mode <= 3, "gemm_fused: mode must be 1 (gelu dual), 2 (dgelu) or 3 (add)");
    B200DP_REQUIRE(b200dp_gemm_fused_supported(M, N), "gemm_fused: shape %d x %d is below the CTA-pair kernel's envelope", M, N);
    B200DP_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % 8 == 0 && N % 8 == 0, "gemm_fused: pitches and N must be multiples of 8");
    B200DP_REQUIRE((mode == 1) == (C2 != nullptr) && (mode == 1 || aux != nullptr), "gemm_fused: C2 / aux do not match the mode");
    B200DP_REQUIRE((mode == 2) == (b_mn != 0), "gemm_fused: mode 2 is the dgrad form (b_mn = 1), modes 1 / 3 the forward form");
    B200DP_REQUIRE(mode == 1 || (ld_aux % 8 == 0 && ((uintptr_t)aux & 15) == 0), "gemm_fused: aux must be 16-byte aligned rows");
    bind_context_once();
    const int is_bf16 = dtype == B200DP_BF16;
    CUtensorMap ta, tb, tc, tc2;
    int rc = make_tmap(&tc, C, is_bf16, M, N, ldc, 64, 32, 2);
    if (rc) return rc;
    if (mode == 1) { rc = make_tmap(&tc2, C2, is_bf16, M, N, ldc2, 64, 32, 2); if (rc) return rc; }
    rc = make_tmap(&ta, A, is_bf16, M, K, lda, kBK, kBM);
    if (rc) return rc;
    rc = b_mn ? make_tmap(&tb, B, is_bf16, K, N, ldb, 64, kBK) : make_tmap(&tb, B, is_bf16, N, K, ldb, kBK, pair::kHalfN);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
#define GF(BM_, T, E) return launch_pair<false, BM_, T, E>(ta, tb, tc, bias, M, N, K, is_bf16, 0, 1, B200DP_SMS, s, &tc2, aux, ld_aux)
    if (mode == 1) { if (is_bf16) GF(false, __nv_bfloat16, kEpiGeluDual); else GF(false, __half, kEpiGeluDual); }
    if (mode == 2) { if (is_bf16) GF(true, __nv_bfloat16, kEpiDGelu); else GF(true, __half, kEpiDGelu); }
    if (is_bf16) GF(false, __nv_bfloat16, kEpiAddAux); else GF(false, __half, kEpiAddAux);
#undef GF
    return 0;
}
