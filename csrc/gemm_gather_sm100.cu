// All-gather fused into the consuming GEMM (W4): Y[M,N] = X[M,K] W[N,K]^T where W is row-sharded
// across the ranks of an NVLink domain (rank p holds rows [p*rpr, (p+1)*rpr) in peer-mapped memory).
//
// ONE kernel per rank does both jobs.  Warp 3 of every CTA (otherwise idle in the GEMM) is a gather
// engine: W is cut into chunks of 1-8 rows, chunk j is pulled by the gather warp of CTA j mod grid
// from the owning rank's shard over NVLink (16-byte peer loads, 16 in flight per lane) into a local
// full copy of W, and each finished chunk bumps the arrival counter of its 128-row weight tile.  The
// TMA producer of any CTA that is about to consume weight tile t first waits until that counter is
// complete, then streams the tile from the local copy like a normal GEMM.  Output tiles
// are visited starting with the weight tiles this rank owns (local copy, ready almost at once), so
// the remote transfers overlap the math of the early tiles; every remote byte crosses NVLink exactly
// once per rank.  Cross-rank ordering: the gather warps of CTA b on all ranks meet at a flag barrier
// before the first remote read (all shards are final) and after their last one (owners may
// overwrite their shards again); the MMA / epilogue warps never wait on another rank.
#include "tc_common.cuh"

namespace b200dp {
namespace gg {

using namespace b200dp::tc;

constexpr int kABytes = kBM * kBK * 2;
constexpr int kEpiBufBytes = 32 * 128;
constexpr int kEpiBytes = 4 * 2 * kEpiBufBytes;
// 128 x 256 tiles (the shape that keeps the tensor pipe fed from shared memory) unless N is small
template <int BN> struct GCfg {
    static constexpr int kStages = BN == 256 ? 4 : 5;
    static constexpr int kBBytes = BN * kBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
    static constexpr int kTmemCols = 2 * BN;
};

struct Shards { const void* p[B200DP_MAX_RANKS]; };

__device__ __forceinline__ uint4 ld_peer16(const void* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

// Warp-level cross-rank barrier on the context's flag row of this CTA (same protocol as
// csrc/collectives.cu, executed by one warp instead of the whole CTA).
__device__ __forceinline__ void warp_barrier_all_ranks(const b200dp_peer_ctx& c, int lane) {
    __syncwarp();
    uint32_t e = 0;
    if (lane == 0) {
        uint32_t* ctr = c.epochs[c.rank] + blockIdx.x;
        e = *ctr + 1u;
        *ctr = e;
        __threadfence_system();
    }
    e = __shfl_sync(0xffffffffu, e, 0);
    if (lane < c.world) {
        uint32_t* theirs = c.flags[lane] + (size_t)blockIdx.x * B200DP_MAX_RANKS + c.rank;
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(theirs), "r"(e) : "memory");
        const uint32_t* mine = c.flags[c.rank] + (size_t)blockIdx.x * B200DP_MAX_RANKS + lane;
        uint64_t t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            uint32_t v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
            if ((int32_t)(v - e) >= 0) break;
            uint64_t t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (c.timeout_ns != 0 && t1 - t0 > c.timeout_ns) { *c.status[c.rank] = (uint32_t)(-B200DP_ERR_TIMEOUT); break; }
            __nanosleep(20);
        }
    }
    __syncwarp();
}

template <int BN, typename TOut>
__global__ void __launch_bounds__(kThreads, 1) gemm_gather_kernel(
    const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmC,
    const float* __restrict__ bias, int M, int N, int K, int is_bf16, Shards shards, TOut* __restrict__ w_full, int rank,
    int rows_per_rank, int chunk_rows, unsigned int* __restrict__ tile_flags, b200dp_peer_ctx ctx, int use_ctx) {
    constexpr int kStages = GCfg<BN>::kStages, kBBytes = GCfg<BN>::kBBytes, kStageBytes = GCfg<BN>::kStageBytes;
    constexpr int kTmemCols = GCfg<BN>::kTmemCols;
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
    const int num_m = (M + kBM - 1) / kBM, num_n = (N + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int num_kb = (K + kBK - 1) / kBK;
    // weight tiles are visited starting with the ones this rank owns
    // (the first tile that lies entirely inside the local shard, when there is one)
    const int first_n = (int)(((long long)rank * rows_per_rank + BN - 1) / BN) % num_n;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); tma_prefetch_desc(&tmC); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 4); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 3) {
        // ===== gather engine =====
        // W is fetched in chunks of chunk_rows rows; chunk j (in the order the weight tiles will be
        // consumed, locally owned tiles first) is copied by the gather warp of CTA j mod grid, so all
        // ~148 gather warps pull in parallel (16 x 16 B in flight per lane) and every tile is complete
        // after a few microseconds; each finished chunk bumps its tile's counter.
        const int vec_per_row = K / 8;                            // 16-byte vectors per weight row
        const int cpt = BN / chunk_rows;                          // chunks per weight tile
        const int total_chunks = num_n * cpt;
        bool met = !use_ctx;                                      // opening cross-rank barrier done?
        for (int j = blockIdx.x; j < total_chunks; j += gridDim.x) {
            const int t = (first_n + j / cpt) % num_n;
            const int row0 = t * BN + (j % cpt) * chunk_rows;
            const int rows = max(0, min(chunk_rows, N - row0));
            // chunks of the local shard need no handshake; the first remote one waits until every
            // rank has entered the kernel (its shard is final by stream order)
            if (!met && rows > 0 && (row0 / rows_per_rank != rank || (row0 + rows - 1) / rows_per_rank != rank)) {
                warp_barrier_all_ranks(ctx, lane);
                met = true;
            }
            const int total = rows * vec_per_row;
            for (int i0 = 0; i0 < total; i0 += 32 * 16) {
                uint4 v[16];
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    const int i = i0 + u * 32 + lane;
                    if (i < total) {
                        const int r = row0 + i / vec_per_row, cv = i % vec_per_row;
                        const int owner = r / rows_per_rank;
                        const TOut* src = (const TOut*)shards.p[owner] + (size_t)(r - owner * rows_per_rank) * K + cv * 8;
                        v[u] = ld_peer16(src);
                    }
                }
#pragma unroll
                for (int u = 0; u < 16; ++u) {
                    const int i = i0 + u * 32 + lane;
                    if (i < total) {
                        const int r = row0 + i / vec_per_row, cv = i % vec_per_row;
                        *reinterpret_cast<uint4*>(w_full + (size_t)r * K + cv * 8) = v[u];
                    }
                }
            }
            __threadfence();
            __syncwarp();
            if (lane == 0) asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(tile_flags + t), "r"(1u) : "memory");
        }
        if (use_ctx) {
            if (!met) warp_barrier_all_ranks(ctx, lane);          // keep the barrier count equal on every rank
            warp_barrier_all_ranks(ctx, lane);                    // nobody reads a peer's shard after this point
        }
    } else if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const int nt = (first_n + tile / num_m) % num_n;
                const int m0 = (tile % num_m) * kBM, n0 = nt * BN;
                // wait until weight tile nt has landed in the local copy
                for (;;) {
                    uint32_t f;
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(f) : "l"(tile_flags + nt) : "memory");
                    if (f >= (uint32_t)(BN / chunk_rows)) break;
                    __nanosleep(40);
                }
                asm volatile("fence.proxy.async.global;" ::: "memory");    // generic-proxy writes -> TMA reads
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(empty + stage, phase ^ 1);
                    mbar_expect_tx(full + stage, kStageBytes);
                    tma_load_2d(smemA + stage * kABytes, &tmA, full + stage, kb * kBK, m0);
                    tma_load_2d(smemB + stage * kBBytes, &tmB, full + stage, kb * kBK, n0);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, 0, 0, kBM, BN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smemA + stage * kABytes);
                    const uint32_t b_addr = smem_u32(smemB + stage * kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k)
                        umma_f16(d_tmem, make_desc(a_addr + k * (kUmmaK * 2), 16, 1024), make_desc(b_addr + k * (kUmmaK * 2), 16, 1024),
                                 idesc, (kb | k) != 0 ? 1u : 0u);
                    umma_commit(empty + stage);
                    if (++stage == kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full + acc);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        const int quarter = warp - kEpiWarp0;
        uint8_t* my_stage = smemC + quarter * 2 * kEpiBufBytes;
        int acc = 0; uint32_t acc_phase = 0;
        uint32_t buf_idx = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const int nt = (first_n + tile / num_m) % num_n;
            const int m0 = (tile % num_m) * kBM, n0 = nt * BN;
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            const int row0 = m0 + quarter * 32;
#pragma unroll 1
            for (int ch = 0; ch < BN / 64; ++ch) {
                uint8_t* buf = my_stage + (buf_idx & 1) * kEpiBufBytes;
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
#pragma unroll
                for (int l = 0; l < 2; ++l) {
                    uint32_t r[32];
                    const int colbase = ch * 64 + l * 32;
                    tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN + colbase, r);
                    tmem_ld_wait();
                    float v[32];
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
                    if (bias != nullptr) {
                        const int gcol = n0 + colbase;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (gcol + j < N) ? __ldg(bias + gcol + j) : 0.f;
                    }
                    uint8_t* rowp = buf + lane * 128;
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        float t[8] = {v[q * 8], v[q * 8 + 1], v[q * 8 + 2], v[q * 8 + 3], v[q * 8 + 4], v[q * 8 + 5], v[q * 8 + 6], v[q * 8 + 7]};
                        store8<TOut>(reinterpret_cast<TOut*>(rowp + (((l * 4 + q) ^ (lane & 7)) << 4)), t);
                    }
                }
                fence_async_smem();
                __syncwarp();
                if (lane == 0 && n0 + ch * 64 < N && row0 < M) tma_store_2d(&tmC, buf, n0 + ch * 64, row0);
                if (lane == 0) tma_store_commit();
                ++buf_idx;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + acc);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (lane == 0) tma_store_wait_read<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
    // the last CTA to finish clears the arrival counters, so the next call needs no memset
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(tile_flags + num_n, 1u) == gridDim.x - 1) {
            for (int i = 0; i <= num_n; ++i) tile_flags[i] = 0u;
        }
    }
}

static int make_map(CUtensorMap* tm, const void* ptr, int is_bf16, int64_t rows, int64_t cols, int64_t ld, int box_cols, int box_rows) {
    static PFN_cuTensorMapEncodeTiled_v12000 enc =
        reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(driver_fn("cuTensorMapEncodeTiled"));
    B200DP_REQUIRE(enc != nullptr, "gemm_gather: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(ptr),
                     dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200DP_REQUIRE(r == CUDA_SUCCESS, "gemm_gather: cuTensorMapEncodeTiled failed (%d)", (int)r);
    return 0;
}

}  // namespace gg
}  // namespace b200dp

using namespace b200dp;
using namespace b200dp::gg;

// Y[M,N] = X[M,K] W^T (+bias) with W[N,K] row-sharded over `world` ranks: w_shards[p] is rank p's
// [rows_per_rank][K] shard as mapped into this process; w_full is a local [N][K] buffer the kernel
// fills while it computes; tile_flags: ceil(N/128)+1 words, zero before the first call (the kernel leaves them zero again).
// ctx == NULL: all shards are local (single-GPU emulation / tests), no cross-rank barrier.
extern "C" int b200dp_gemm_gather(const void* X, const void* const* w_shards, void* w_full, void* Y, const float* bias, int M,
                                  int N, int K, int ldx, int ldy, int world, int rank, int rows_per_rank, b200dp_dtype dtype,
                                  unsigned int* tile_flags, const b200dp_peer_ctx* ctx, void* stream) {
    B200DP_REQUIRE(X && w_shards && w_full && Y && tile_flags, "gemm_gather: null argument");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "gemm_gather: fp16 / bf16 only");
    B200DP_REQUIRE(world >= 1 && world <= B200DP_MAX_RANKS && rank >= 0 && rank < world, "gemm_gather: bad world / rank");
    B200DP_REQUIRE(K % 8 == 0 && ldx % 8 == 0 && ldy % 8 == 0, "gemm_gather: K, ldx, ldy must be multiples of 8");
    B200DP_REQUIRE((long long)rows_per_rank * world >= N && rows_per_rank > 0, "gemm_gather: shards do not cover N");
    bind_context_once();
    const int is_bf16 = dtype == B200DP_BF16;
    CUtensorMap ta, tb, tc;
    int rc = make_map(&ta, X, is_bf16, M, K, ldx, kBK, kBM);
    if (rc) return rc;
    const int BN = (N >= 256 && (int64_t)M * N >= (int64_t)B200DP_SMS * 128 * 256) ? 256 : 128;
    rc = make_map(&tb, w_full, is_bf16, N, K, K, kBK, BN);
    if (rc) return rc;
    rc = make_map(&tc, Y, is_bf16, M, N, ldy, 64, 32);
    if (rc) return rc;
    Shards sh;
    for (int i = 0; i < B200DP_MAX_RANKS; ++i) sh.p[i] = i < world ? w_shards[i] : nullptr;
    cudaStream_t s = (cudaStream_t)stream;
    const int num_n = (N + BN - 1) / BN;
    const int tiles = ((M + kBM - 1) / kBM) * num_n;
    (void)tiles;
    const int grid = B200DP_SMS;                                   // persistent; every SM contributes a gather warp
    int chunk_rows = 8;                                            // at least two chunks per gather warp when W is small
    while (chunk_rows > 1 && num_n * (BN / chunk_rows) < 2 * grid) chunk_rows >>= 1;
    b200dp_peer_ctx c{};
    if (ctx) {
        c = *ctx;
        B200DP_REQUIRE(grid <= c.max_blocks, "gemm_gather: %d CTAs but the peer context has %d flag rows", grid, c.max_blocks);
    }
#define GG(BNv, T)                                                                                                       \
    {                                                                                                                    \
        auto kern = gemm_gather_kernel<BNv, T>;                                                                          \
        static bool configured = false;                                                                                  \
        if (!configured) { B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, GCfg<BNv>::kSmemBytes)); configured = true; } \
        kern<<<grid, kThreads, GCfg<BNv>::kSmemBytes, s>>>(ta, tb, tc, bias, M, N, K, is_bf16, sh, (T*)w_full, rank, rows_per_rank,  \
                                                chunk_rows, tile_flags, c, ctx ? 1 : 0);                                             \
    }
    if (BN == 256) { if (is_bf16) GG(256, __nv_bfloat16) else GG(256, __half) }
    else { if (is_bf16) GG(128, __nv_bfloat16) else GG(128, __half) }
#undef GG
    B200DP_CHECK_LAUNCH();
    return 0;
}
