// Implicit-GEMM convolution on tcgen05 / TMEM, operands fetched by TMA straight from the NHWC
// activation tensor — no im2col buffer, no cuDNN (W7).
//
// The GEMM view:  M = output pixels (tiles of bn x bp x bq pixels, <= 128 rows),
//                 N = output channels, K = (filter tap r,s) x (64-channel chunk).
// For one k-block the A tile is "the input pixels seen by this tile of output pixels through
// tap (r,s)": a box of the 4-D tensor [N][H][W][C] starting at (q0*stride + s - pad,
// p0*stride + r - pad) and traversed with the convolution stride.  That is exactly what a tiled
// TMA load with elementStrides = {1, stride, stride, 1} fetches; out-of-image pixels (padding,
// ragged tiles) are zero-filled by the TMA unit.  Rows land in smem in (n, p, q) order with 128 B
// per row (64 channels), i.e. the K-major SWIZZLE_128B layout the UMMA descriptor expects.
//
//   fprop  Y[n,p,q,co]  = sum_{r,s,ci} X[n, p*st+r-pad, q*st+s-pad, ci] * W[co,r,s,ci]
//          A = X via 4-D TMA, B = W viewed as [Cout][R*S*Cin] (K-major)
//   dgrad  dX[n,h,w,ci] = sum_{r,s,co} dY[n, h+pad-r, w+pad-s, co] * W[co,r,s,ci]     (stride 1)
//          A = dY via 4-D TMA (tap offsets negated), B = W viewed as [Cout][R*S*Cin] read MN-major
//   wgrad  dW[co,r,s,ci] = sum_{n,p,q} dY[n,p,q,co] * X[n, p*st+r-pad, q*st+s-pad, ci]
//          both operands MN-major via 4-D TMA, K = pixels (exactly 64 per k-block), split over
//          pixel tiles and combined with TMA reduce-add into the fp32 gradient
//
//          stride 2: four stride-1 problems, one per parity class of the input pixel, each with an
//          explicit tap list (Geom::ntaps / tap_id / tap_dh / tap_dw) and a strided output mapping
//
// Kernels in this file:
//   conv_fd_kernel          fprop / dgrad, one tap and one 64-channel chunk per k-block (any filter, stride 1/2)
//   conv_rows_kernel        3x3 stride-1 fprop / dgrad for < 256 output channels: one haloed activation patch
//                           feeds the three filter rows (patch ring + weight ring, or the filter resident)
//   conv_fd_pair_kernel     fprop / dgrad for 256-channel tiles on a CTA pair (cta_group::2)
//   conv_wgrad_kernel       weight gradient, one tap per work item
//   conv_wgrad_rows_kernel  3x3 stride-1 weight gradient, three taps per work item from one dY box + one X patch
// fprop kernels can also emit the per-channel sum / sum of squares of their output (kStats).
//
// Warp roles, pipeline and TMEM double buffering are those of gemm_sm100.cu.  The epilogue
// stages 32 rows x 128 B per warp in swizzled smem and writes them out 4 rows per instruction
// (8 lanes x 16 B per row) to the pixel's address, masking rows that fall outside the image.
#include "tc_common.cuh"

namespace b200dp {
namespace conv {

using namespace b200dp::tc;

struct Geom {
    int n_img, out_h, out_w;      // pixel space this pass writes (fprop: P,Q; dgrad: H,W)
    int c_out, c_red;             // GEMM N and the per-tap reduction length
    int R, S, stride;
    int off_h, off_w, tap_sign;   // input coordinate = out*stride + off + tap_sign*tap
    int bq, bp, bn;               // tile box (pixels): bq*bp*bn <= 128
    int tiles_q, tiles_p, tiles_n;
    int w_taps_ld;                // R*S*Cin: row pitch of the weight matrix view (elements)
    // Strided input gradient, one launch per output parity class (conv_fd_kernel only): ntaps > 0 selects an
    // explicit tap list — weight tap index and the (unit-stride) dY offset it reads — and the pass writes
    // pixel (p * out_mul + out_off_h, q * out_mul + out_off_w) of a full_h x full_w image.
    int ntaps;
    int out_mul, out_off_h, out_off_w, full_h, full_w;
    signed char tap_id[16], tap_dh[16], tap_dw[16];
};

__device__ __forceinline__ void tma_load_4d(void* smem_dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1, int c2,
                                            int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

template <int BN> struct Cfg {
    static constexpr int kStages = BN == 256 ? 4 : (BN == 128 ? 5 : 7);
    static constexpr int kABytes = kBM * kBK * 2;
    static constexpr int kBBytes = BN * kBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kEpiBufBytes = 32 * 128;
    static constexpr int kEpiBytes = 4 * kEpiBufBytes;      // one staging buffer per epilogue warp
    static constexpr int kSmemBytes = kStages * kStageBytes + 2 * kEpiBytes + 1024 + 256;   // wgrad double-buffers its staging
    static constexpr int kTmemCols = 2 * BN;
};

// fprop (kBMN = false) and stride-1 dgrad (kBMN = true).  kStats (fprop): the epilogue also adds the
// per-channel sum / sum of squares of the stored values into colstats[0:Cout] / [Cout:2Cout] — the
// batch statistics of the BatchNorm that follows — so that layer skips its own pass over the tensor.
template <int BN, bool kBMN, typename TOut, bool kStats = false>
__global__ void __launch_bounds__(kThreads, 1) conv_fd_kernel(const __grid_constant__ CUtensorMap tmA,
                                                              const __grid_constant__ CUtensorMap tmB,
                                                              TOut* __restrict__ out, const float* __restrict__ bias,
                                                              Geom g, int is_bf16, float* __restrict__ colstats) {
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
    const int num_m = g.tiles_q * g.tiles_p * g.tiles_n;
    const int num_n = (g.c_out + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int chunks = g.c_red / kBK;
    const int num_kb = (g.ntaps > 0 ? g.ntaps : g.R * g.S) * chunks;
    const int rows_valid = g.bq * g.bp * g.bn;
    const uint32_t a_bytes = (uint32_t)rows_valid * 128u;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < cfg::kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 4); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, cfg::kTmemCols);
    // rows of the A tile beyond rows_valid are never written by TMA: clear them once so the MMA
    // reads zeros (their results are discarded, but NaN bit patterns must not linger in smem)
    for (int i = threadIdx.x; i < cfg::kStages * cfg::kABytes / 16; i += kThreads)
        reinterpret_cast<uint4*>(smemA)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    auto decode_m = [&](int mt, int& n0, int& p0, int& q0) {
        q0 = (mt % g.tiles_q) * g.bq;
        p0 = ((mt / g.tiles_q) % g.tiles_p) * g.bp;
        n0 = (mt / (g.tiles_q * g.tiles_p)) * g.bn;
    };

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int n0, p0, q0;
                decode_m(tile % num_m, n0, p0, q0);
                const int nt0 = (tile / num_m) * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int t = kb / chunks, c0 = (kb % chunks) * kBK;
                    int tap, cw, chh;
                    if (g.ntaps > 0) {                  // parity class of a strided dgrad: explicit tap list
                        tap = g.tap_id[t];
                        cw = q0 + g.tap_dw[t];
                        chh = p0 + g.tap_dh[t];
                    } else {
                        tap = t;
                        const int r = tap / g.S, s = tap % g.S;
                        cw = q0 * g.stride + g.off_w + g.tap_sign * s;
                        chh = p0 * g.stride + g.off_h + g.tap_sign * r;
                    }
                    mbar_wait(empty + stage, phase ^ 1);
                    mbar_expect_tx(full + stage, a_bytes + cfg::kBBytes);
                    tma_load_4d(smemA + stage * cfg::kABytes, &tmA, full + stage, c0, cw, chh, n0);
                    uint8_t* b = smemB + stage * cfg::kBBytes;
                    if (kBMN) {
#pragma unroll
                        for (int j = 0; j < BN / 64; ++j)
                            tma_load_2d(b + j * (kBK * 128), &tmB, full + stage, tap * g.c_out + nt0 + j * 64, c0);
                    } else {
                        tma_load_2d(b, &tmB, full + stage, tap * g.c_red + c0, nt0);
                    }
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, 0, kBMN ? 1 : 0, kBM, BN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smemA + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smemB + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k) {
                        const uint64_t da = make_desc(a_addr + k * (kUmmaK * 2), 16, 1024);
                        const uint64_t db = kBMN ? make_desc(b_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                 : make_desc(b_addr + k * (kUmmaK * 2), 16, 1024);
                        umma_f16(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit(empty + stage);
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full + acc);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        const int quarter = warp - kEpiWarp0;
        uint8_t* buf = smemC + quarter * cfg::kEpiBufBytes;
        int acc = 0; uint32_t acc_phase = 0;
        constexpr int kChunks = BN / 64;
        // channel statistics: lane l owns channels 2l, 2l+1 of every 64-channel chunk of the current channel tile
        float st[kStats ? kChunks * 4 : 1];
        int st_n0 = -1;
#pragma unroll
        for (int i = 0; i < (kStats ? kChunks * 4 : 1); ++i) st[i] = 0.f;
        auto flush_stats = [&]() {
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) {
                    const int col = st_n0 + ch * 64 + 2 * lane;
                    if (col < g.c_out) {          // c_out is even (multiple of 8): col + 1 is valid too
                        atomicAdd(colstats + col, st[ch * 4 + 0]);
                        atomicAdd(colstats + col + 1, st[ch * 4 + 1]);
                        atomicAdd(colstats + g.c_out + col, st[ch * 4 + 2]);
                        atomicAdd(colstats + g.c_out + col + 1, st[ch * 4 + 3]);
                    }
                    st[ch * 4 + 0] = st[ch * 4 + 1] = st[ch * 4 + 2] = st[ch * 4 + 3] = 0.f;
                }
            }
        };
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            int n0, p0, q0;
            decode_m(tile % num_m, n0, p0, q0);
            const int nt0 = (tile / num_m) * BN;
            if (kStats && nt0 != st_n0) {
                if (st_n0 >= 0) flush_stats();
                st_n0 = nt0;
            }
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            // this lane's row of the tile -> output pixel
            const int t = quarter * 32 + lane;
            const int bq_i = t % g.bq, bp_i = (t / g.bq) % g.bp, bn_i = t / (g.bq * g.bp);
            const int n = n0 + bn_i, p = p0 + bp_i, q = q0 + bq_i;
            const bool valid = t < rows_valid && n < g.n_img && p < g.out_h && q < g.out_w;
            const long long pix = valid ? (((long long)n * g.full_h + p * g.out_mul + g.out_off_h) * g.full_w + q * g.out_mul + g.out_off_w) : -1;
            auto do_chunk = [&](const int ch) {
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
                        const int gcol = nt0 + colbase;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (gcol + j < g.c_out) ? __ldg(bias + gcol + j) : 0.f;
                    }
                    if (kStats && !valid) {       // pixels outside the image must not count
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = 0.f;
                    }
                    uint8_t* rowp = buf + lane * 128;
#pragma unroll
                    for (int qd = 0; qd < 4; ++qd) {
                        float tt[8] = {v[qd * 8], v[qd * 8 + 1], v[qd * 8 + 2], v[qd * 8 + 3],
                                       v[qd * 8 + 4], v[qd * 8 + 5], v[qd * 8 + 6], v[qd * 8 + 7]};
                        const int j = l * 4 + qd;
                        store8<TOut>(reinterpret_cast<TOut*>(rowp + ((j ^ (lane & 7)) << 4)), tt);
                    }
                }
                __syncwarp();
                // write-out: 8 lanes per row, 4 rows per instruction, 128 contiguous bytes per row
                const int col0 = nt0 + ch * 64;
#pragma unroll
                for (int rr = 0; rr < 32; rr += 4) {
                    const int row = rr + (lane >> 3), chunk = lane & 7;
                    const long long rpix = __shfl_sync(0xffffffffu, pix, row);
                    const uint4 val = *reinterpret_cast<const uint4*>(buf + row * 128 + ((chunk ^ (row & 7)) << 4));
                    if (rpix >= 0 && col0 + chunk * 8 < g.c_out)
                        *reinterpret_cast<uint4*>(out + rpix * g.c_out + col0 + chunk * 8) = val;
                }
                if constexpr (kStats) {
                    // lane l sums the 4-byte word of channels 2l, 2l+1 over the 32 staged rows
                    const int j = lane >> 2, w = lane & 3;
                    float s0 = 0.f, s1 = 0.f, q0s = 0.f, q1s = 0.f;
#pragma unroll 8
                    for (int rr = 0; rr < 32; ++rr) {
                        const uint32_t u = *reinterpret_cast<const uint32_t*>(buf + rr * 128 + ((j ^ (rr & 7)) << 4) + w * 4);
                        const float2 f = unpack2<TOut>(u);
                        s0 += f.x; s1 += f.y;
                        q0s = fmaf(f.x, f.x, q0s); q1s = fmaf(f.y, f.y, q1s);
                    }
                    st[ch * 4 + 0] += s0; st[ch * 4 + 1] += s1; st[ch * 4 + 2] += q0s; st[ch * 4 + 3] += q1s;
                }
            };
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            } else {
#pragma unroll 1
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + acc);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (kStats && st_n0 >= 0) flush_stats();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, cfg::kTmemCols);
}


// ---------------------------------------------------------------- fprop / dgrad, 3 x 3 stride 1, row-grouped
// The tap-at-a-time kernel above is bound by operand traffic on the shallow layers (64 / 128 channels: few
// MMA cycles per byte loaded; ~64 B/clk/SM is what TMA + L2 deliver).  Here one activation PATCH of
// (bp + 2) x bq pixels (bq a multiple of 8) is loaded once per filter column s and 64-channel chunk, and the
// A operand of filter row r is that same patch read from byte offset r * bq * 128 — a multiple of the
// 1024-byte swizzle atom, so the UMMA descriptor just starts there.  Activation traffic drops ~2.5x.
// Two rings: activation patches (kRowsAStages) and per-tap weight tiles (kRowsBStages).
constexpr int kRowsMaxAStages = 8;     // activation patches in flight: as many as fit (host decides), at most 8
constexpr int kRowsBStages = 8;
constexpr int kRowsEpiBufBytes = 32 * 128;
constexpr int kRowsEpiWarps = 8;
constexpr int kRowsThreads = (kEpiWarp0 + kRowsEpiWarps) * 32;
constexpr int kRowsEpiBytes = kRowsEpiWarps * kRowsEpiBufBytes;
template <int BN> struct RCfg {
    static constexpr int kBBytes = BN * kBK * 2;                      // one tap's weight tile
    static constexpr int kBStages = BN == 128 ? 6 : kRowsBStages;
    // patch: up to (2 + 2) * 64 = 256 rows of 128 B; sized by the host per geometry (a_patch_bytes)
    static int smem_bytes(int a_stages, int a_patch_bytes, int b_tiles) { return a_stages * a_patch_bytes + b_tiles * kBBytes + kRowsEpiBytes + 1024 + 512; }
};

template <int BN, bool kBMN, typename TOut, bool kStats = false>
__global__ void __launch_bounds__(kRowsThreads, 1) conv_rows_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                TOut* __restrict__ out, const float* __restrict__ bias,
                                                                Geom g, int is_bf16, float* __restrict__ colstats,
                                                                int a_patch_bytes, int w_resident, int a_stages) {
    using cfg = RCfg<BN>;
    // (Splitting a tile's k-steps over several TMEM accumulators — in case back-to-back MMAs on one accumulator
    // serialised — was tried and changed nothing; ncu shows the tensor pipe waiting on operands instead.)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smemA = smem;
    uint8_t* smemB = smem + a_stages * a_patch_bytes;
    uint8_t* smemC = smemB + (w_resident ? 9 * (g.c_red / kBK) : cfg::kBStages) * cfg::kBBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemC + kRowsEpiBytes);
    uint64_t* a_full = bars;
    uint64_t* a_empty = a_full + kRowsMaxAStages;
    uint64_t* b_full = a_empty + kRowsMaxAStages;
    uint64_t* b_empty = b_full + cfg::kBStages;
    uint64_t* tmem_full = b_empty + cfg::kBStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_m = g.tiles_q * g.tiles_p * g.tiles_n;
    const int num_n = (g.c_out + BN - 1) / BN;
    const int num_tiles = num_m * num_n;
    const int chunks = g.c_red / kBK;
    const int rows_valid = g.bq * g.bp * g.bn;           // == 128 here
    const uint32_t a_bytes = (uint32_t)(g.bp + 2) * g.bq * 128u;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < a_stages; ++i) { mbar_init(a_full + i, 1); mbar_init(a_empty + i, 1); }
        for (int i = 0; i < cfg::kBStages; ++i) { mbar_init(b_full + i, 1); mbar_init(b_empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 4); }
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, 2 * BN);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    auto decode_m = [&](int mt, int& n0, int& p0, int& q0) {
        q0 = (mt % g.tiles_q) * g.bq;
        p0 = ((mt / g.tiles_q) % g.tiles_p) * g.bp;
        n0 = (mt / (g.tiles_q * g.tiles_p)) * g.bn;
    };
    // input row seen by output row p through filter row r: p + off_h + tap_sign * r.  The patch starts at the
    // lowest of the three; filter row r then sits r (fprop) or 2 - r (dgrad, tap_sign = -1) patch rows down.
    const int patch_dh = g.tap_sign > 0 ? 0 : -2;

    if (warp == 0) {
        if (lane == 0) {
            int sa = 0; uint32_t pa = 0;
            int sb = 0; uint32_t pb = 0;
            if (w_resident) {
                // the whole filter (9 taps x chunks weight tiles, one channel tile) fits next to the patches:
                // fetch it once per CTA, tile (tap, chunk) at smemB + (tap * chunks + chunk) * kBBytes
                mbar_expect_tx(b_full, (uint32_t)(9 * chunks) * cfg::kBBytes);
                for (int tap = 0; tap < 9; ++tap)
                    for (int ck = 0; ck < chunks; ++ck) {
                        uint8_t* b = smemB + (tap * chunks + ck) * cfg::kBBytes;
                        if (kBMN) {
#pragma unroll
                            for (int j = 0; j < BN / 64; ++j)
                                tma_load_2d(b + j * (kBK * 128), &tmB, b_full, tap * g.c_out + j * 64, ck * kBK);
                        } else {
                            tma_load_2d(b, &tmB, b_full, tap * g.c_red + ck * kBK, 0);
                        }
                    }
            }
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                int n0, p0, q0;
                decode_m(tile % num_m, n0, p0, q0);
                const int nt0 = (tile / num_m) * BN;
                for (int s = 0; s < 3; ++s) {
                    for (int ck = 0; ck < chunks; ++ck) {
                        const int c0 = ck * kBK;
                        mbar_wait(a_empty + sa, pa ^ 1);
                        mbar_expect_tx(a_full + sa, a_bytes);
                        tma_load_4d(smemA + sa * a_patch_bytes, &tmA, a_full + sa, c0, q0 + g.off_w + g.tap_sign * s,
                                    p0 + g.off_h + patch_dh, n0);
                        if (++sa == a_stages) { sa = 0; pa ^= 1; }
                        for (int r = 0; r < 3 && !w_resident; ++r) {
                            const int tap = r * 3 + s;
                            mbar_wait(b_empty + sb, pb ^ 1);
                            mbar_expect_tx(b_full + sb, cfg::kBBytes);
                            uint8_t* b = smemB + sb * cfg::kBBytes;
                            if (kBMN) {
#pragma unroll
                                for (int j = 0; j < BN / 64; ++j)
                                    tma_load_2d(b + j * (kBK * 128), &tmB, b_full + sb, tap * g.c_out + nt0 + j * 64, c0);
                            } else {
                                tma_load_2d(b, &tmB, b_full + sb, tap * g.c_red + c0, nt0);
                            }
                            if (++sb == cfg::kBStages) { sb = 0; pb ^= 1; }
                        }
                    }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, 0, kBMN ? 1 : 0, kBM, BN);
            int sa = 0; uint32_t pa = 0;
            int sb = 0; uint32_t pb = 0;
            int acc = 0; uint32_t acc_phase = 0;
            if (w_resident) { mbar_wait(b_full, 0); tc_fence_after(); }
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                bool first = true;
                for (int grp = 0; grp < 3 * chunks; ++grp) {
                    mbar_wait(a_full + sa, pa);
                    tc_fence_after();
                    const uint32_t a_base = smem_u32(smemA + sa * a_patch_bytes);
                    for (int r = 0; r < 3; ++r) {
                        if (!w_resident) {
                            mbar_wait(b_full + sb, pb);
                            tc_fence_after();
                        }
                        const uint32_t a_addr = a_base + (uint32_t)((g.tap_sign > 0 ? r : 2 - r) * g.bq) * 128u;
                        // producer order: s outermost, then chunk: grp = s * chunks + chunk
                        const uint32_t b_addr = w_resident
                            ? smem_u32(smemB + ((r * 3 + grp / chunks) * chunks + grp % chunks) * cfg::kBBytes)
                            : smem_u32(smemB + sb * cfg::kBBytes);
#pragma unroll
                        for (int k = 0; k < kBK / kUmmaK; ++k) {
                            const uint64_t da = make_desc(a_addr + k * (kUmmaK * 2), 16, 1024);
                            const uint64_t db = kBMN ? make_desc(b_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                     : make_desc(b_addr + k * (kUmmaK * 2), 16, 1024);
                            umma_f16(d_tmem, da, db, idesc, (first && k == 0) ? 0u : 1u);
                        }
                        first = false;
                        if (!w_resident) {
                            umma_commit(b_empty + sb);
                            if (++sb == cfg::kBStages) { sb = 0; pb ^= 1; }
                        }
                    }
                    umma_commit(a_empty + sa);
                    if (++sa == a_stages) { sa = 0; pa ^= 1; }
                }
                umma_commit(tmem_full + acc);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        // EIGHT epilogue warps: warps 4-7 drain the tiles that land in accumulator 0 (this CTA's even tiles),
        // warps 8-11 those in accumulator 1 — with K as short as 9 x 64 a tile's MMAs take ~1.2k cycles, far
        // less than one warp set needs to convert, store and sum a 128 x BN tile
        const int ew = warp - kEpiWarp0;
        const int quarter = ew & 3, eset = ew >> 2;
        uint8_t* buf = smemC + ew * kRowsEpiBufBytes;
        const int acc = eset; uint32_t acc_phase = 0;
        constexpr int kChunks = BN / 64;
        // channel statistics: lane l owns channels 2l, 2l+1 of every 64-channel chunk of the current channel tile
        float st[kStats ? kChunks * 4 : 1];
        int st_n0 = -1;
#pragma unroll
        for (int i = 0; i < (kStats ? kChunks * 4 : 1); ++i) st[i] = 0.f;
        auto flush_stats = [&]() {
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) {
                    const int col = st_n0 + ch * 64 + 2 * lane;
                    if (col < g.c_out) {          // c_out is even (multiple of 8): col + 1 is valid too
                        atomicAdd(colstats + col, st[ch * 4 + 0]);
                        atomicAdd(colstats + col + 1, st[ch * 4 + 1]);
                        atomicAdd(colstats + g.c_out + col, st[ch * 4 + 2]);
                        atomicAdd(colstats + g.c_out + col + 1, st[ch * 4 + 3]);
                    }
                    st[ch * 4 + 0] = st[ch * 4 + 1] = st[ch * 4 + 2] = st[ch * 4 + 3] = 0.f;
                }
            }
        };
        for (int tile = blockIdx.x + eset * gridDim.x; tile < num_tiles; tile += 2 * gridDim.x) {
            int n0, p0, q0;
            decode_m(tile % num_m, n0, p0, q0);
            const int nt0 = (tile / num_m) * BN;
            if (kStats && nt0 != st_n0) {
                if (st_n0 >= 0) flush_stats();
                st_n0 = nt0;
            }
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            // this lane's row of the tile -> output pixel
            const int t = quarter * 32 + lane;
            const int bq_i = t % g.bq, bp_i = (t / g.bq) % g.bp, bn_i = t / (g.bq * g.bp);
            const int n = n0 + bn_i, p = p0 + bp_i, q = q0 + bq_i;
            const bool valid = t < rows_valid && n < g.n_img && p < g.out_h && q < g.out_w;
            const long long pix = valid ? (((long long)n * g.out_h + p) * g.out_w + q) : -1;
            auto do_chunk = [&](const int ch) {
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
                        const int gcol = nt0 + colbase;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (gcol + j < g.c_out) ? __ldg(bias + gcol + j) : 0.f;
                    }
                    if (kStats && !valid) {       // pixels outside the image must not count
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = 0.f;
                    }
                    uint8_t* rowp = buf + lane * 128;
#pragma unroll
                    for (int qd = 0; qd < 4; ++qd) {
                        float tt[8] = {v[qd * 8], v[qd * 8 + 1], v[qd * 8 + 2], v[qd * 8 + 3],
                                       v[qd * 8 + 4], v[qd * 8 + 5], v[qd * 8 + 6], v[qd * 8 + 7]};
                        const int j = l * 4 + qd;
                        store8<TOut>(reinterpret_cast<TOut*>(rowp + ((j ^ (lane & 7)) << 4)), tt);
                    }
                }
                __syncwarp();
                // write-out: 8 lanes per row, 4 rows per instruction, 128 contiguous bytes per row
                const int col0 = nt0 + ch * 64;
#pragma unroll
                for (int rr = 0; rr < 32; rr += 4) {
                    const int row = rr + (lane >> 3), chunk = lane & 7;
                    const long long rpix = __shfl_sync(0xffffffffu, pix, row);
                    const uint4 val = *reinterpret_cast<const uint4*>(buf + row * 128 + ((chunk ^ (row & 7)) << 4));
                    if (rpix >= 0 && col0 + chunk * 8 < g.c_out)
                        *reinterpret_cast<uint4*>(out + rpix * g.c_out + col0 + chunk * 8) = val;
                }
                if constexpr (kStats) {
                    // lane l sums the 4-byte word of channels 2l, 2l+1 over the 32 staged rows
                    const int j = lane >> 2, w = lane & 3;
                    float s0 = 0.f, s1 = 0.f, q0s = 0.f, q1s = 0.f;
#pragma unroll 8
                    for (int rr = 0; rr < 32; ++rr) {
                        const uint32_t u = *reinterpret_cast<const uint32_t*>(buf + rr * 128 + ((j ^ (rr & 7)) << 4) + w * 4);
                        const float2 f = unpack2<TOut>(u);
                        s0 += f.x; s1 += f.y;
                        q0s = fmaf(f.x, f.x, q0s); q1s = fmaf(f.y, f.y, q1s);
                    }
                    st[ch * 4 + 0] += s0; st[ch * 4 + 1] += s1; st[ch * 4 + 2] += q0s; st[ch * 4 + 3] += q1s;
                }
            };
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            } else {
#pragma unroll 1
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty + acc);
            acc_phase ^= 1;
        }
        if (kStats && st_n0 >= 0) flush_stats();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, 2 * BN);
}



// ---------------------------------------------------------------- fprop / dgrad, CTA pair (cta_group::2)
// Same implicit GEMM with two CTAs of a cluster on one 256-pixel x BN tile: each CTA fetches ITS 128 pixels
// and only HALF of the weight tile per k-block, the leader issues one M = 256 tcgen05.mma for the pair.
// The weight tile is the larger operand for BN >= 128, so L2 -> SM operand traffic drops 25-33 %, which is
// what bounds these kernels (protocol as in gemm_sm100.cu: leader-only full / tmem_empty barriers, multicast
// commits).
template <int BN> struct PCfg {
    static constexpr int kABytes = kBM * kBK * 2;
    static constexpr int kBBytes = (BN / 2) * kBK * 2;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = BN == 256 ? 6 : 8;
    static constexpr int kEpiBufBytes = 32 * 128;
    static constexpr int kEpiBytes = 4 * kEpiBufBytes;
    static constexpr int kSmemBytes = kStages * kStageBytes + 2 * kEpiBytes + 1024 + 256;
    static constexpr int kTmemCols = 2 * BN;
};

__device__ __forceinline__ void tma_load_4d_pair(void* smem_dst, const CUtensorMap* tm, uint64_t* leader_bar, int c0, int c1,
                                                 int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(smem_dst)), "l"(tm), "r"(smem_u32(leader_bar) & kPeerBitMask), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}

template <int BN, bool kBMN, typename TOut, bool kStats = false>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_fd_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, TOut* __restrict__ out,
                    const float* __restrict__ bias, Geom g, int is_bf16, float* __restrict__ colstats) {
    using cfg = PCfg<BN>;
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
    const int num_m = g.tiles_q * g.tiles_p * g.tiles_n;
    const int num_n = (g.c_out + BN - 1) / BN;
    const uint32_t cta_rank = cluster_ctarank();
    const bool leader = cta_rank == 0;
    const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
    const int num_mp = (num_m + 1) / 2;                  // pixel tiles are handed out two at a time
    const int num_tiles = num_mp * num_n;
    const int chunks = g.c_red / kBK;
    const int num_kb = g.R * g.S * chunks;
    const int rows_valid = g.bq * g.bp * g.bn;
    const uint32_t a_bytes = (uint32_t)rows_valid * 128u;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < cfg::kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        for (int i = 0; i < 2; ++i) { mbar_init(tmem_full + i, 1); mbar_init(tmem_empty + i, 8); }
        fence_barrier_init();
    }
    // rows of the A tile beyond rows_valid are never written by TMA: clear them once so the MMA
    // reads zeros (their results are discarded, but NaN bit patterns must not linger in smem)
    for (int i = threadIdx.x; i < cfg::kStages * cfg::kABytes / 16; i += kThreads)
        reinterpret_cast<uint4*>(smemA)[i] = make_uint4(0, 0, 0, 0);
    fence_async_smem();
    cluster_sync_all();                       // both CTAs' barriers exist before anything remote touches them
    if (warp == 2) tmem_alloc_pair(tmem_slot, cfg::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    auto decode_m = [&](int mt, int& n0, int& p0, int& q0) {
        q0 = (mt % g.tiles_q) * g.bq;
        p0 = ((mt / g.tiles_q) % g.tiles_p) * g.bp;
        n0 = (mt / (g.tiles_q * g.tiles_p)) * g.bn;
    };

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int tile = pair_id; tile < num_tiles; tile += num_pairs) {
                int n0, p0, q0;
                decode_m(2 * (tile % num_mp) + (int)cta_rank, n0, p0, q0);      // an index past the last tile decodes to n0 >= n_img: all rows out of range
                const int nt0 = (tile / num_mp) * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    const int tap = kb / chunks, c0 = (kb % chunks) * kBK;
                    const int r = tap / g.S, s = tap % g.S;
                    mbar_wait(empty + stage, phase ^ 1);
                    if (leader) mbar_expect_tx(full + stage, 2 * (a_bytes + cfg::kBBytes));
                    tma_load_4d_pair(smemA + stage * cfg::kABytes, &tmA, full + stage, c0,
                                q0 * g.stride + g.off_w + g.tap_sign * s, p0 * g.stride + g.off_h + g.tap_sign * r, n0);
                    uint8_t* b = smemB + stage * cfg::kBBytes;
                    if (kBMN) {
#pragma unroll
                        for (int j = 0; j < BN / 128; ++j)
                            tma_load_2d_pair(b + j * (kBK * 128), &tmB, full + stage,
                                             tap * g.c_out + nt0 + (int)cta_rank * (BN / 2) + j * 64, c0);
                    } else {
                        tma_load_2d_pair(b, &tmB, full + stage, tap * g.c_red + c0, nt0 + (int)cta_rank * (BN / 2));
                    }
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (leader && lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, 0, kBMN ? 1 : 0, 2 * kBM, BN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int tile = pair_id; tile < num_tiles; tile += num_pairs) {
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = 0; kb < num_kb; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smemA + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smemB + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k) {
                        const uint64_t da = make_desc(a_addr + k * (kUmmaK * 2), 16, 1024);
                        const uint64_t db = kBMN ? make_desc(b_addr + k * (kUmmaK * 128), kBK * 128, 1024)
                                                 : make_desc(b_addr + k * (kUmmaK * 2), 16, 1024);
                        umma_f16_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
                    }
                    umma_commit_pair(empty + stage, 3);
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit_pair(tmem_full + acc, 3);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        const int quarter = warp - kEpiWarp0;
        uint8_t* buf = smemC + quarter * cfg::kEpiBufBytes;
        int acc = 0; uint32_t acc_phase = 0;
        constexpr int kChunks = BN / 64;
        // channel statistics: lane l owns channels 2l, 2l+1 of every 64-channel chunk of the current channel tile
        float st[kStats ? kChunks * 4 : 1];
        int st_n0 = -1;
#pragma unroll
        for (int i = 0; i < (kStats ? kChunks * 4 : 1); ++i) st[i] = 0.f;
        auto flush_stats = [&]() {
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) {
                    const int col = st_n0 + ch * 64 + 2 * lane;
                    if (col < g.c_out) {          // c_out is even (multiple of 8): col + 1 is valid too
                        atomicAdd(colstats + col, st[ch * 4 + 0]);
                        atomicAdd(colstats + col + 1, st[ch * 4 + 1]);
                        atomicAdd(colstats + g.c_out + col, st[ch * 4 + 2]);
                        atomicAdd(colstats + g.c_out + col + 1, st[ch * 4 + 3]);
                    }
                    st[ch * 4 + 0] = st[ch * 4 + 1] = st[ch * 4 + 2] = st[ch * 4 + 3] = 0.f;
                }
            }
        };
        for (int tile = pair_id; tile < num_tiles; tile += num_pairs) {
            int n0, p0, q0;
            decode_m(2 * (tile % num_mp) + (int)cta_rank, n0, p0, q0);      // an index past the last tile decodes to n0 >= n_img: all rows out of range
            const int nt0 = (tile / num_mp) * BN;
            if (kStats && nt0 != st_n0) {
                if (st_n0 >= 0) flush_stats();
                st_n0 = nt0;
            }
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            // this lane's row of the tile -> output pixel
            const int t = quarter * 32 + lane;
            const int bq_i = t % g.bq, bp_i = (t / g.bq) % g.bp, bn_i = t / (g.bq * g.bp);
            const int n = n0 + bn_i, p = p0 + bp_i, q = q0 + bq_i;
            const bool valid = t < rows_valid && n < g.n_img && p < g.out_h && q < g.out_w;
            const long long pix = valid ? (((long long)n * g.out_h + p) * g.out_w + q) : -1;
            auto do_chunk = [&](const int ch) {
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
                        const int gcol = nt0 + colbase;
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] += (gcol + j < g.c_out) ? __ldg(bias + gcol + j) : 0.f;
                    }
                    if (kStats && !valid) {       // pixels outside the image must not count
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = 0.f;
                    }
                    uint8_t* rowp = buf + lane * 128;
#pragma unroll
                    for (int qd = 0; qd < 4; ++qd) {
                        float tt[8] = {v[qd * 8], v[qd * 8 + 1], v[qd * 8 + 2], v[qd * 8 + 3],
                                       v[qd * 8 + 4], v[qd * 8 + 5], v[qd * 8 + 6], v[qd * 8 + 7]};
                        const int j = l * 4 + qd;
                        store8<TOut>(reinterpret_cast<TOut*>(rowp + ((j ^ (lane & 7)) << 4)), tt);
                    }
                }
                __syncwarp();
                // write-out: 8 lanes per row, 4 rows per instruction, 128 contiguous bytes per row
                const int col0 = nt0 + ch * 64;
#pragma unroll
                for (int rr = 0; rr < 32; rr += 4) {
                    const int row = rr + (lane >> 3), chunk = lane & 7;
                    const long long rpix = __shfl_sync(0xffffffffu, pix, row);
                    const uint4 val = *reinterpret_cast<const uint4*>(buf + row * 128 + ((chunk ^ (row & 7)) << 4));
                    if (rpix >= 0 && col0 + chunk * 8 < g.c_out)
                        *reinterpret_cast<uint4*>(out + rpix * g.c_out + col0 + chunk * 8) = val;
                }
                if constexpr (kStats) {
                    // lane l sums the 4-byte word of channels 2l, 2l+1 over the 32 staged rows
                    const int j = lane >> 2, w = lane & 3;
                    float s0 = 0.f, s1 = 0.f, q0s = 0.f, q1s = 0.f;
#pragma unroll 8
                    for (int rr = 0; rr < 32; ++rr) {
                        const uint32_t u = *reinterpret_cast<const uint32_t*>(buf + rr * 128 + ((j ^ (rr & 7)) << 4) + w * 4);
                        const float2 f = unpack2<TOut>(u);
                        s0 += f.x; s1 += f.y;
                        q0s = fmaf(f.x, f.x, q0s); q1s = fmaf(f.y, f.y, q1s);
                    }
                    st[ch * 4 + 0] += s0; st[ch * 4 + 1] += s1; st[ch * 4 + 2] += q0s; st[ch * 4 + 3] += q1s;
                }
            };
            if constexpr (kStats) {
#pragma unroll
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            } else {
#pragma unroll 1
                for (int ch = 0; ch < kChunks; ++ch) do_chunk(ch);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_leader(tmem_empty + acc);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (kStats && st_n0 >= 0) flush_stats();
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 2) tmem_dealloc_pair(tmem_base, cfg::kTmemCols);
}



// ---------------------------------------------------------------- wgrad
// dW[co][tap][ci] += sum over a 64-pixel box of dY[pix][co] * X[pix shifted by tap][ci].
// Work item = (tap, co tile, ci tile, split); a split walks its share of the pixel boxes.  Both
// operands are MN-major (channels contiguous, pixels = K rows), fetched by 4-D TMA; out-of-image
// pixels are zero-filled for both, so ragged edges and padding contribute nothing.
struct WGeom {
    int n_img, out_h, out_w;      // dY pixel space (P, Q)
    int c_out, c_in, R, S, stride, pad;
    int tiles_q, tiles_p;         // 8 x 8 pixel boxes per image
    int splits, kb_per;           // pixel boxes per split
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 1) conv_wgrad_kernel(const __grid_constant__ CUtensorMap tmDY,
                                                                 const __grid_constant__ CUtensorMap tmX,
                                                                 const __grid_constant__ CUtensorMap tmDW, WGeom g,
                                                                 int is_bf16) {
    using cfg = Cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smemA = smem;
    uint8_t* smemB = smem + cfg::kStages * cfg::kABytes;
    uint8_t* smemC = smem + cfg::kStages * cfg::kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemC + 2 * cfg::kEpiBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + cfg::kStages;
    uint64_t* tmem_full = bars + 2 * cfg::kStages;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_m = (g.c_out + kBM - 1) / kBM, num_n = (g.c_in + BN - 1) / BN;
    const int taps = g.R * g.S;
    const int boxes_per_img = g.tiles_p * g.tiles_q;
    const int num_kb = g.n_img * boxes_per_img;
    const int num_items = taps * num_m * num_n * g.splits;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmDY); tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmDW); }
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

    auto decode = [&](int item, int& tap, int& m0, int& n0, int& kb0, int& kb1) {
        const int split = item % g.splits;
        int rest = item / g.splits;
        n0 = (rest % num_n) * BN; rest /= num_n;
        m0 = (rest % num_m) * kBM; rest /= num_m;
        tap = rest;
        kb0 = split * g.kb_per;
        kb1 = min(num_kb, kb0 + g.kb_per);
    };

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                int tap, m0, n0, kb0, kb1;
                decode(item, tap, m0, n0, kb0, kb1);
                const int r = tap / g.S, s = tap % g.S;
                for (int kb = kb0; kb < kb1; ++kb) {
                    const int img = kb / boxes_per_img, bx = kb % boxes_per_img;
                    const int p0 = (bx / g.tiles_q) * 8, q0 = (bx % g.tiles_q) * 8;
                    mbar_wait(empty + stage, phase ^ 1);
                    mbar_expect_tx(full + stage, cfg::kStageBytes);
                    uint8_t* a = smemA + stage * cfg::kABytes;
                    uint8_t* b = smemB + stage * cfg::kBBytes;
#pragma unroll
                    for (int j = 0; j < kBM / 64; ++j) tma_load_4d(a + j * (kBK * 128), &tmDY, full + stage, m0 + j * 64, q0, p0, img);
#pragma unroll
                    for (int j = 0; j < BN / 64; ++j)
                        tma_load_4d(b + j * (kBK * 128), &tmX, full + stage, n0 + j * 64, q0 * g.stride + s - g.pad,
                                    p0 * g.stride + r - g.pad, img);
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, 1, 1, kBM, BN);
            int stage = 0; uint32_t phase = 0;
            int acc = 0; uint32_t acc_phase = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                int tap, m0, n0, kb0, kb1;
                decode(item, tap, m0, n0, kb0, kb1);
                mbar_wait(tmem_empty + acc, acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * BN;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smemA + stage * cfg::kABytes);
                    const uint32_t b_addr = smem_u32(smemB + stage * cfg::kBBytes);
#pragma unroll
                    for (int k = 0; k < kBK / kUmmaK; ++k) {
                        const uint64_t da = make_desc(a_addr + k * (kUmmaK * 128), kBK * 128, 1024);
                        const uint64_t db = make_desc(b_addr + k * (kUmmaK * 128), kBK * 128, 1024);
                        umma_f16(d_tmem, da, db, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
                    }
                    umma_commit(empty + stage);
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full + acc);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= kEpiWarp0) {
        const int quarter = warp - kEpiWarp0;
        uint8_t* my_stage = smemC + quarter * 2 * cfg::kEpiBufBytes;
        int acc = 0; uint32_t acc_phase = 0;
        uint32_t buf_idx = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
            int tap, m0, n0, kb0, kb1;
            decode(item, tap, m0, n0, kb0, kb1);
            mbar_wait(tmem_full + acc, acc_phase);
            tc_fence_after();
            const int row0 = m0 + quarter * 32;
#pragma unroll 1
            for (int ch = 0; ch < BN / 32; ++ch) {
                uint8_t* buf = my_stage + (buf_idx & 1) * cfg::kEpiBufBytes;
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
                uint32_t r[32];
                tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + acc * BN + ch * 32, r);
                tmem_ld_wait();
                uint8_t* rowp = buf + lane * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4)) = make_uint4(r[j * 4], r[j * 4 + 1], r[j * 4 + 2], r[j * 4 + 3]);
                fence_async_smem();
                __syncwarp();
                const int col = n0 + ch * 32;
                if (lane == 0 && col < g.c_in && row0 < g.c_out && kb1 > kb0)
                    tma_reduce_add_2d(&tmDW, buf, tap * g.c_in + col, row0);
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
    if (warp == 2) tmem_dealloc(tmem_base, cfg::kTmemCols);
}

// ---------------------------------------------------------------- wgrad, 3 x S filters, stride 1
// Row-grouped variant: one work item produces the THREE taps (r = 0,1,2) of one filter column s.
// Per 8 x 8 pixel box it loads dY once and an X patch of 10 x 8 pixels (two halo rows); because a
// patch row is 8 pixels x 128 B = 1024 B — one swizzle atom — the operand for tap r is the same
// patch read from byte offset r * 1024, so the three MMAs of a k-step share both loads.  L2 -> SM
// traffic per FLOP drops ~2.7x against the tap-at-a-time kernel above, which is what bounds it.
// The 3 x BN fp32 accumulators fill TMEM (single-buffered); items are long (>= 16 k-blocks).
template <int BN> struct WCfg {
    static constexpr int kPatchBytes = 10 * 8 * 128;                    // per 64-channel chunk of X
    static constexpr int kABytes = kBM * kBK * 2;
    static constexpr int kBBytes = (BN / 64) * kPatchBytes;
    static constexpr int kStageBytes = kABytes + kBBytes;
    static constexpr int kStages = BN == 128 ? 5 : 7;
    static constexpr int kEpiBufBytes = 32 * 128;
    static constexpr int kEpiBytes = 4 * 2 * kEpiBufBytes;
    static constexpr int kSmemBytes = kStages * kStageBytes + kEpiBytes + 1024 + 256;
    static constexpr int kTmemCols = BN == 128 ? 512 : 256;             // 3 * BN rounded up to a power of two
};

template <int BN>
__global__ void __launch_bounds__(kThreads, 1) conv_wgrad_rows_kernel(const __grid_constant__ CUtensorMap tmDY,
                                                                      const __grid_constant__ CUtensorMap tmX,
                                                                      const __grid_constant__ CUtensorMap tmDW, WGeom g,
                                                                      int is_bf16) {
    using cfg = WCfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* smemC = smem + cfg::kStages * cfg::kStageBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smemC + cfg::kEpiBytes);
    uint64_t* full = bars;
    uint64_t* empty = bars + cfg::kStages;
    uint64_t* tmem_full = bars + 2 * cfg::kStages;
    uint64_t* tmem_empty = tmem_full + 1;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_empty + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_m = (g.c_out + kBM - 1) / kBM, num_n = (g.c_in + BN - 1) / BN;
    const int boxes_per_img = g.tiles_p * g.tiles_q;
    const int num_kb = g.n_img * boxes_per_img;
    const int num_items = g.S * num_m * num_n * g.splits;

    if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmDY); tma_prefetch_desc(&tmX); tma_prefetch_desc(&tmDW); }
    if (warp == 1 && lane == 0) {
        for (int i = 0; i < cfg::kStages; ++i) { mbar_init(full + i, 1); mbar_init(empty + i, 1); }
        mbar_init(tmem_full, 1);
        mbar_init(tmem_empty, 4);
        fence_barrier_init();
    }
    if (warp == 2) tmem_alloc(tmem_slot, cfg::kTmemCols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    auto decode = [&](int item, int& s, int& m0, int& n0, int& kb0, int& kb1) {
        const int split = item % g.splits;
        int rest = item / g.splits;
        n0 = (rest % num_n) * BN; rest /= num_n;
        m0 = (rest % num_m) * kBM; rest /= num_m;
        s = rest;
        kb0 = split * g.kb_per;
        kb1 = min(num_kb, kb0 + g.kb_per);
    };

    if (warp == 0) {
        if (lane == 0) {
            int stage = 0; uint32_t phase = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                int s, m0, n0, kb0, kb1;
                decode(item, s, m0, n0, kb0, kb1);
                for (int kb = kb0; kb < kb1; ++kb) {
                    const int img = kb / boxes_per_img, bx = kb % boxes_per_img;
                    const int p0 = (bx / g.tiles_q) * 8, q0 = (bx % g.tiles_q) * 8;
                    mbar_wait(empty + stage, phase ^ 1);
                    mbar_expect_tx(full + stage, cfg::kStageBytes);
                    uint8_t* a = smem + stage * cfg::kStageBytes;
                    uint8_t* b = a + cfg::kABytes;
#pragma unroll
                    for (int j = 0; j < kBM / 64; ++j) tma_load_4d(a + j * (kBK * 128), &tmDY, full + stage, m0 + j * 64, q0, p0, img);
#pragma unroll
                    for (int j = 0; j < BN / 64; ++j)
                        tma_load_4d(b + j * cfg::kPatchBytes, &tmX, full + stage, n0 + j * 64, q0 + s - g.pad, p0 - g.pad, img);
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc(is_bf16, 1, 1, kBM, BN);
            int stage = 0; uint32_t phase = 0;
            uint32_t acc_phase = 0;
            for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
                int s, m0, n0, kb0, kb1;
                decode(item, s, m0, n0, kb0, kb1);
                mbar_wait(tmem_empty, acc_phase ^ 1);
                tc_fence_after();
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(full + stage, phase);
                    tc_fence_after();
                    const uint32_t a_addr = smem_u32(smem + stage * cfg::kStageBytes);
                    const uint32_t b_addr = a_addr + cfg::kABytes;
#pragma unroll
                    for (int r = 0; r < 3; ++r) {
#pragma unroll
                        for (int k = 0; k < kBK / kUmmaK; ++k) {
                            const uint64_t da = make_desc(a_addr + k * (kUmmaK * 128), kBK * 128, 1024);
                            const uint64_t db = make_desc(b_addr + r * 1024 + k * (kUmmaK * 128), cfg::kPatchBytes, 1024);
                            umma_f16(tmem_base + r * BN, da, db, idesc, (kb != kb0 || k != 0) ? 1u : 0u);
                        }
                    }
                    umma_commit(empty + stage);
                    if (++stage == cfg::kStages) { stage = 0; phase ^= 1; }
                }
                umma_commit(tmem_full);
                acc_phase ^= 1;
            }
        }
    } else if (warp >= kEpiWarp0) {
        const int quarter = warp - kEpiWarp0;
        uint8_t* my_stage = smemC + quarter * 2 * cfg::kEpiBufBytes;
        uint32_t acc_phase = 0;
        uint32_t buf_idx = 0;
        for (int item = blockIdx.x; item < num_items; item += gridDim.x) {
            int s, m0, n0, kb0, kb1;
            decode(item, s, m0, n0, kb0, kb1);
            mbar_wait(tmem_full, acc_phase);
            tc_fence_after();
            const int row0 = m0 + quarter * 32;
#pragma unroll 1
            for (int ch = 0; ch < 3 * (BN / 32); ++ch) {
                const int r = ch / (BN / 32), cc = ch % (BN / 32);
                uint8_t* buf = my_stage + (buf_idx & 1) * cfg::kEpiBufBytes;
                if (lane == 0) tma_store_wait_read<1>();
                __syncwarp();
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_base + ((uint32_t)(quarter * 32) << 16) + r * BN + cc * 32, v);
                tmem_ld_wait();
                uint8_t* rowp = buf + lane * 128;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4*>(rowp + ((j ^ (lane & 7)) << 4)) = make_uint4(v[j * 4], v[j * 4 + 1], v[j * 4 + 2], v[j * 4 + 3]);
                fence_async_smem();
                __syncwarp();
                const int col = n0 + cc * 32;
                if (lane == 0 && col < g.c_in && row0 < g.c_out && kb1 > kb0)
                    tma_reduce_add_2d(&tmDW, buf, (r * g.S + s) * g.c_in + col, row0);
                if (lane == 0) tma_store_commit();
                ++buf_idx;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(tmem_empty);
            acc_phase ^= 1;
        }
        if (lane == 0) tma_store_wait_read<0>();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) tmem_dealloc(tmem_base, cfg::kTmemCols);
}

// ---------------------------------------------------------------- host
static PFN_cuTensorMapEncodeTiled_v12000 get_encode() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn =
        reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(driver_fn("cuTensorMapEncodeTiled"));
    return fn;
}

// NHWC activation [N][H][W][C] as a 4-D map {C, W, H, N}; box = {64 ch, bq, bp, bn} pixels traversed with `stride`
static int make_act_map(CUtensorMap* tm, const void* ptr, int is_bf16, int N, int H, int W, int C, int bq, int bp, int bn,
                        int stride, int box_c = 64) {
    auto enc = get_encode();
    B200DP_REQUIRE(enc != nullptr, "conv: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    cuuint32_t box[4] = {(cuuint32_t)box_c, (cuuint32_t)((bq - 1) * stride + 1), (cuuint32_t)((bp - 1) * stride + 1),
                         (cuuint32_t)bn};
    cuuint32_t estr[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
    B200DP_REQUIRE(box[1] <= 256 && box[2] <= 256 && box[3] <= 256, "conv: TMA box too large (%u,%u,%u)", box[1], box[2], box[3]);
    CUresult r = enc(tm, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4,
                     const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200DP_REQUIRE(r == CUDA_SUCCESS, "conv: cuTensorMapEncodeTiled(4d) failed (%d) N=%d H=%d W=%d C=%d box=(%u,%u,%u,%u) stride=%d",
                   (int)r, N, H, W, C, box[0], box[1], box[2], box[3], stride);
    return 0;
}

static int make_mat_map(CUtensorMap* tm, const void* ptr, int is_bf16, int64_t rows, int64_t cols, int64_t ld, int box_cols,
                        int box_rows) {
    auto enc = get_encode();
    B200DP_REQUIRE(enc != nullptr, "conv: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
    cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2,
                     const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200DP_REQUIRE(r == CUDA_SUCCESS, "conv: cuTensorMapEncodeTiled(2d) failed (%d)", (int)r);
    return 0;
}

// fp32 dW as a matrix [Cout][R*S*Cin]; staging rows are 128 B = 32 floats
static int make_dw_map(CUtensorMap* tm, float* dw, int c_out, int64_t row_len) {
    auto enc = get_encode();
    B200DP_REQUIRE(enc != nullptr, "conv: cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[2] = {(cuuint64_t)row_len, (cuuint64_t)c_out};
    cuuint64_t strides[1] = {(cuuint64_t)row_len * 4};
    cuuint32_t box[2] = {32, 32};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, dw, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200DP_REQUIRE(r == CUDA_SUCCESS, "conv2d_wgrad: cuTensorMapEncodeTiled(dW) failed (%d)", (int)r);
    return 0;
}

// pick the pixel box: full rows when they fit, several rows / images per tile up to 128 pixels
static void pick_box(int n_img, int out_h, int out_w, int stride, int& bq, int& bp, int& bn) {
    const int max_span = stride == 1 ? 256 : 128;       // (b-1)*stride+1 <= 256
    bq = out_w < 128 ? out_w : 128;
    if (bq > max_span) bq = max_span;
    bp = 128 / bq;
    if (bp > out_h) bp = out_h;
    if (bp < 1) bp = 1;
    bn = 1;
    if (bq == out_w && bp == out_h) {
        bn = 128 / (bq * bp);
        if (bn > n_img) bn = n_img;
        if (bn < 1) bn = 1;
    }
}

template <int BN, bool kBMN, typename TOut, bool kStats = false>
static int launch_fd(const CUtensorMap& ta, const CUtensorMap& tb, void* out, const float* bias, const Geom& g, int is_bf16,
                     cudaStream_t s, float* colstats = nullptr) {
    auto kern = conv_fd_kernel<BN, kBMN, TOut, kStats>;
    static bool configured = false;
    if (!configured) {
        B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BN>::kSmemBytes));
        configured = true;
    }
    const int tiles = g.tiles_q * g.tiles_p * g.tiles_n * ((g.c_out + BN - 1) / BN);
    const int grid = tiles < B200DP_SMS ? tiles : B200DP_SMS;
    kern<<<grid, kThreads, Cfg<BN>::kSmemBytes, s>>>(ta, tb, (TOut*)out, bias, g, is_bf16, colstats);
    B200DP_CHECK_LAUNCH();
    return 0;
}

template <int BN, bool kBMN, typename TOut, bool kStats = false>
static int launch_fd_pair(const CUtensorMap& ta, const CUtensorMap& tb, void* out, const float* bias, const Geom& g, int is_bf16,
                          cudaStream_t s, float* colstats = nullptr) {
    auto kern = conv_fd_pair_kernel<BN, kBMN, TOut, kStats>;
    static bool configured = false;
    if (!configured) {
        B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PCfg<BN>::kSmemBytes));
        configured = true;
    }
    const int num_m = g.tiles_q * g.tiles_p * g.tiles_n;
    const int tiles = ((num_m + 1) / 2) * ((g.c_out + BN - 1) / BN);
    int pairs = B200DP_SMS / 2;
    if (tiles < pairs) pairs = tiles;
    kern<<<2 * pairs, kThreads, PCfg<BN>::kSmemBytes, s>>>(ta, tb, (TOut*)out, bias, g, is_bf16, colstats);
    B200DP_CHECK_LAUNCH();
    return 0;
}

template <int BN, bool kBMN, typename TOut, bool kStats = false>
static int launch_rows(const CUtensorMap& ta, const CUtensorMap& tb, void* out, const float* bias, const Geom& g, int is_bf16,
                       cudaStream_t s, float* colstats, int a_patch_bytes) {
    auto kern = conv_rows_kernel<BN, kBMN, TOut, kStats>;
    static bool configured = false;
    if (!configured) {
        B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        configured = true;
    }
    const int tiles = g.tiles_q * g.tiles_p * g.tiles_n * ((g.c_out + BN - 1) / BN);
    const int grid = tiles < B200DP_SMS ? tiles : B200DP_SMS;
    // weights stay resident in shared memory when the whole filter of the (single) channel tile fits next to at
    // least three activation patches; whatever shared memory is left goes to more patches in flight (the
    // shallow layers are bound by the TMA round trip, not by bandwidth)
    const int chunks = g.c_red / kBK, all_tiles = 9 * chunks;
    const int w_resident = (g.c_out <= BN && RCfg<BN>::smem_bytes(3, a_patch_bytes, all_tiles) <= 232448) ? 1 : 0;
    const int b_tiles = w_resident ? all_tiles : RCfg<BN>::kBStages;
    int a_stages = (232448 - RCfg<BN>::smem_bytes(0, a_patch_bytes, b_tiles)) / a_patch_bytes;
    if (a_stages > kRowsMaxAStages) a_stages = kRowsMaxAStages;
    B200DP_REQUIRE(a_stages >= 2, "conv2d: activation patch of %d bytes does not fit twice in shared memory", a_patch_bytes);
    const int smem = RCfg<BN>::smem_bytes(a_stages, a_patch_bytes, b_tiles);
    kern<<<grid, kRowsThreads, smem, s>>>(ta, tb, (TOut*)out, bias, g, is_bf16, colstats, a_patch_bytes, w_resident, a_stages);
    B200DP_CHECK_LAUNCH();
    return 0;
}

// pixel tile for the row-grouped kernel: bq x bp = 128 with bq a multiple of 8; the shape that wastes the
// fewest rows on ragged edges wins (ties: the narrower one, its halo patch is smaller).  Returns the fraction
// of useful rows.
static double pick_rows_box(int out_h, int out_w, int& bq, int& bp) {
    double best = 0.0;
    bq = 8; bp = 16;
    for (int q = 8; q <= 64; q *= 2) {
        const int p = 128 / q;
        const double u = ((double)out_w / (((out_w + q - 1) / q) * q)) * ((double)out_h / (((out_h + p - 1) / p) * p));
        if (u > best + 1e-9) { best = u; bq = q; bp = p; }
    }
    return best;
}

// B200DP_CONV_ROWS=0 turns the row-grouped fprop / dgrad kernel off, B200DP_CONV_PAIR=0 the cta_group::2 one
static bool conv_rows_enabled() {
    static const bool on = [] { const char* e = getenv("B200DP_CONV_ROWS"); return !(e && e[0] == '0'); }();
    return on;
}

// B200DP_CONV_PAIR=0 turns the cta_group::2 convolution off (A/B measurements)
static bool conv_pair_enabled() {
    static const bool on = [] { const char* e = getenv("B200DP_CONV_PAIR"); return !(e && e[0] == '0'); }();
    return on;
}

}  // namespace conv
}  // namespace b200dp

using namespace b200dp;
using namespace b200dp::conv;

// NHWC convolution forward (mode 0) or stride-1 input gradient (mode 1).
//   mode 0: x = input  [N][H][W][Cin],  y = output [N][P][Q][Cout]
//   mode 1: x = dY     [N][P][Q][Cout], y = dX     [N][H][W][Cin]   (stride must be 1)
//   w = [Cout][R][S][Cin] (channels-last weight), 16-bit like x.  Cin and Cout multiples of 64
//   on the reduction side (Cin for fprop, Cout for dgrad) and of 8 on the output side.
static int conv2d_impl(int mode, const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin,
                       int Cout, int R, int S, int stride, int pad, b200dp_dtype dtype, float* colstats, void* stream) {
    B200DP_REQUIRE(x && w && y, "conv2d: null argument");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "conv2d: fp16 / bf16 only");
    B200DP_REQUIRE(mode == 0 || mode == 1, "conv2d: mode 0 (forward) or 1 (input gradient)");
    B200DP_REQUIRE(mode == 0 || stride == 1 || (R <= 7 && S <= 7), "conv2d: strided dgrad supports filters up to 7 x 7");
    B200DP_REQUIRE(stride >= 1 && stride <= 2 && R >= 1 && S >= 1, "conv2d: unsupported stride / filter");
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    const int c_red = mode == 0 ? Cin : Cout, c_out = mode == 0 ? Cout : Cin;
    B200DP_REQUIRE(c_red % 64 == 0, "conv2d: the reduction channel count (%d) must be a multiple of 64", c_red);
    B200DP_REQUIRE(c_out % 8 == 0, "conv2d: the output channel count (%d) must be a multiple of 8", c_out);
    bind_context_once();
    const int is_bf16 = dtype == B200DP_BF16;
    Geom g{};
    g.n_img = N; g.R = R; g.S = S; g.c_out = c_out; g.c_red = c_red; g.w_taps_ld = R * S * Cin;
    g.ntaps = 0; g.out_mul = 1; g.out_off_h = 0; g.out_off_w = 0;
    int src_h, src_w;
    if (mode == 0) { g.out_h = P; g.out_w = Q; g.stride = stride; g.off_h = -pad; g.off_w = -pad; g.tap_sign = 1; src_h = H; src_w = W; }
    else           { g.out_h = H; g.out_w = W; g.stride = 1; g.off_h = pad; g.off_w = pad; g.tap_sign = -1; src_h = P; src_w = Q; }
    g.full_h = g.out_h; g.full_w = g.out_w;
    if (mode == 1 && stride == 2) {
        // Strided input gradient = four stride-1 problems, one per parity class (ph, pw) of the input pixel:
        //   dX[n, 2a+ph, 2b+pw] = sum over taps (r, s) with (ph+pad-r), (pw+pad-s) even of
        //                         dY[n, a + (ph+pad-r)/2, b + (pw+pad-s)/2] * W[:, r, s, :]
        // Each class runs the dgrad kernel over dY with its own tap list and writes every second pixel.
        CUtensorMap tb;
        int rc = make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, 64, kBK);
        if (rc) return rc;
        cudaStream_t cs = (cudaStream_t)stream;
        const int BNc = c_out <= 64 ? 64 : (c_out >= 256 ? 256 : 128);
        const int esz = 2;
        int nonempty = 0;
        for (int ph = 0; ph < 2; ++ph)
            for (int pw = 0; pw < 2; ++pw) {
                int cnt = 0;
                for (int r = 0; r < R; ++r) for (int q = 0; q < S; ++q)
                    if (((ph + pad - r) & 1) == 0 && ((pw + pad - q) & 1) == 0) ++cnt;
                if (cnt > 0 && (H - ph + 1) / 2 > 0 && (W - pw + 1) / 2 > 0) ++nonempty;
            }
        if (nonempty < 4) B200DP_CUDA(cudaMemsetAsync(y, 0, (size_t)N * H * W * Cin * esz, cs));
        for (int ph = 0; ph < 2; ++ph)
            for (int pw = 0; pw < 2; ++pw) {
                Geom gc = g;
                gc.stride = 1; gc.out_mul = 2; gc.out_off_h = ph; gc.out_off_w = pw; gc.full_h = H; gc.full_w = W;
                gc.out_h = (H - ph + 1) / 2; gc.out_w = (W - pw + 1) / 2;
                gc.ntaps = 0;
                for (int r = 0; r < R; ++r)
                    for (int q = 0; q < S; ++q) {
                        const int eh = ph + pad - r, ew = pw + pad - q;
                        if ((eh & 1) || (ew & 1)) continue;
                        gc.tap_id[gc.ntaps] = (signed char)(r * S + q);
                        gc.tap_dh[gc.ntaps] = (signed char)(eh / 2);
                        gc.tap_dw[gc.ntaps] = (signed char)(ew / 2);
                        ++gc.ntaps;
                    }
                if (gc.ntaps == 0 || gc.out_h <= 0 || gc.out_w <= 0) continue;
                pick_box(N, gc.out_h, gc.out_w, 1, gc.bq, gc.bp, gc.bn);
                gc.tiles_q = (gc.out_w + gc.bq - 1) / gc.bq;
                gc.tiles_p = (gc.out_h + gc.bp - 1) / gc.bp;
                gc.tiles_n = (N + gc.bn - 1) / gc.bn;
                CUtensorMap ta;
                rc = make_act_map(&ta, x, is_bf16, N, P, Q, c_red, gc.bq, gc.bp, gc.bn, 1);
                if (rc) return rc;
#define GC(BNv)                                                                                          \
    { rc = is_bf16 ? launch_fd<BNv, true, __nv_bfloat16>(ta, tb, y, nullptr, gc, 1, cs)                    \
                   : launch_fd<BNv, true, __half>(ta, tb, y, nullptr, gc, 0, cs); }
                if (BNc == 256) GC(256) else if (BNc == 128) GC(128) else GC(64)
#undef GC
                if (rc) return rc;
            }
        return 0;
    }
    if (R == 3 && S == 3 && stride == 1 && c_out < 256 && conv_rows_enabled()) {
        // shallow 3x3 layers: row-grouped kernel (one haloed activation patch feeds the three filter rows)
        int rq, rp;
        if (pick_rows_box(g.out_h, g.out_w, rq, rp) >= 0.75) {
            g.bq = rq; g.bp = rp; g.bn = 1;
            g.tiles_q = (g.out_w + g.bq - 1) / g.bq;
            g.tiles_p = (g.out_h + g.bp - 1) / g.bp;
            g.tiles_n = N;
            const int BNr = c_out <= 64 ? 64 : 128;
            const int a_patch_bytes = (g.bp + 2) * g.bq * 128;
            CUtensorMap ta, tb;
            int rc = make_act_map(&ta, x, is_bf16, N, src_h, src_w, c_red, g.bq, g.bp + 2, 1, 1);
            if (rc) return rc;
            rc = mode == 0 ? make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, kBK, BNr)
                           : make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, 64, kBK);
            if (rc) return rc;
            cudaStream_t rs = (cudaStream_t)stream;
#define GR(BNv, MN, T, ST) return launch_rows<BNv, MN, T, ST>(ta, tb, y, bias, g, is_bf16, rs, colstats, a_patch_bytes)
#define GR_T(BNv, MN, ST) if (is_bf16) { GR(BNv, MN, __nv_bfloat16, ST); } else { GR(BNv, MN, __half, ST); }
#define GR_MODE(BNv)                                                 \
    if (mode == 0 && colstats != nullptr) { GR_T(BNv, false, true) }  \
    else if (mode == 0) { GR_T(BNv, false, false) }                   \
    else { GR_T(BNv, true, false) }
            if (BNr == 128) { GR_MODE(128) } else { GR_MODE(64) }
#undef GR_MODE
#undef GR_T
#undef GR
        }
    }
    pick_box(N, g.out_h, g.out_w, g.stride, g.bq, g.bp, g.bn);
    g.tiles_q = (g.out_w + g.bq - 1) / g.bq;
    g.tiles_p = (g.out_h + g.bp - 1) / g.bp;
    g.tiles_n = (N + g.bn - 1) / g.bn;
    const int BN = c_out <= 64 ? 64 : (c_out >= 256 ? 256 : 128);
    CUtensorMap ta, tb;
    int rc = make_act_map(&ta, x, is_bf16, N, src_h, src_w, c_red, g.bq, g.bp, g.bn, g.stride);
    if (rc) return rc;
    {
        const int num_m_tiles = g.tiles_q * g.tiles_p * g.tiles_n;
        // measured (scripts/bench_conv.py, batch 256): 5-15 % faster than the single-CTA kernel for 256-channel
        // tiles, no gain at 128 (there the activation operand, not the weights, is the larger load)
        if (BN == 256 && conv_pair_enabled() && ((num_m_tiles + 1) / 2) * ((c_out + BN - 1) / BN) >= 74) {
            // CTA-pair kernel: each CTA of a pair fetches half of the weight tile (box of BN/2 rows / 64-column boxes)
            rc = mode == 0 ? make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, kBK, BN / 2)
                           : make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, 64, kBK);
            if (rc) return rc;
            cudaStream_t ps = (cudaStream_t)stream;
#define GP(BNv, MN, T, ST) return launch_fd_pair<BNv, MN, T, ST>(ta, tb, y, bias, g, is_bf16, ps, colstats)
#define GP_T(BNv, MN, ST) if (is_bf16) { GP(BNv, MN, __nv_bfloat16, ST); } else { GP(BNv, MN, __half, ST); }
#define GP_MODE(BNv)                                           \
    if (mode == 0 && colstats != nullptr) { GP_T(BNv, false, true) } \
    else if (mode == 0) { GP_T(BNv, false, false) }             \
    else { GP_T(BNv, true, false) }
            if (BN == 256) { GP_MODE(256) } else { GP_MODE(128) }
#undef GP_MODE
#undef GP_T
#undef GP
        }
    }
    // weights as a matrix [Cout][R*S*Cin]: fprop reads it K-major (box 64 x BN rows), dgrad MN-major (box 64 x 64)
    rc = mode == 0 ? make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, kBK, BN)
                   : make_mat_map(&tb, w, is_bf16, Cout, (int64_t)R * S * Cin, (int64_t)R * S * Cin, 64, kBK);
    if (rc) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (colstats != nullptr) {
#define GS(BNv)                                                                                               \
    {                                                                                                         \
        if (is_bf16) return launch_fd<BNv, false, __nv_bfloat16, true>(ta, tb, y, bias, g, 1, s, colstats);    \
        return launch_fd<BNv, false, __half, true>(ta, tb, y, bias, g, 0, s, colstats);                        \
    }
        if (BN == 256) GS(256) else if (BN == 128) GS(128) else GS(64)
#undef GS
    }
#define GO(BNv, MN, T) return launch_fd<BNv, MN, T>(ta, tb, y, bias, g, is_bf16, s)
#define BY_T(BNv, MN) if (is_bf16) { GO(BNv, MN, __nv_bfloat16); } else { GO(BNv, MN, __half); }
#define BY_MODE(BNv) if (mode == 0) { BY_T(BNv, false) } else { BY_T(BNv, true) }
    if (BN == 256) { BY_MODE(256) } else if (BN == 128) { BY_MODE(128) } else { BY_MODE(64) }
#undef BY_MODE
#undef BY_T
#undef GO
    return 0;
}

extern "C" int b200dp_conv2d(int mode, const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin,
                             int Cout, int R, int S, int stride, int pad, b200dp_dtype dtype, void* stream) {
    return conv2d_impl(mode, x, w, y, bias, N, H, W, Cin, Cout, R, S, stride, pad, dtype, nullptr, stream);
}

// Forward convolution that also accumulates the per-channel sum / sum of squares of its output into
// colstats[0:Cout] / colstats[Cout:2Cout] (fp32 atomics; the caller zeroes colstats).
extern "C" int b200dp_conv2d_colstats(const void* x, const void* w, void* y, const float* bias, int N, int H, int W, int Cin,
                                      int Cout, int R, int S, int stride, int pad, b200dp_dtype dtype, float* colstats,
                                      void* stream) {
    B200DP_REQUIRE(colstats != nullptr, "conv2d_colstats: null statistics buffer");
    return conv2d_impl(0, x, w, y, bias, N, H, W, Cin, Cout, R, S, stride, pad, dtype, colstats, stream);
}

// Weight gradient: dw[Cout][R][S][Cin] (fp32, channels-last weight layout) = sum over pixels of
// dy[N][P][Q][Cout] (x) x[N][H][W][Cin] shifted by the tap.  Cin and Cout multiples of 8; stride 1 or 2.
// The result is accumulated with TMA reduce-add: dw is zeroed here first unless accumulate != 0.
extern "C" int b200dp_conv2d_wgrad(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin, int Cout, int R,
                                   int S, int stride, int pad, b200dp_dtype dtype, int accumulate, void* stream) {
    B200DP_REQUIRE(x && dy && dw, "conv2d_wgrad: null argument");
    B200DP_REQUIRE(dtype == B200DP_F16 || dtype == B200DP_BF16, "conv2d_wgrad: fp16 / bf16 only");
    B200DP_REQUIRE(Cin % 8 == 0 && Cout % 8 == 0, "conv2d_wgrad: channel counts must be multiples of 8");
    B200DP_REQUIRE(stride >= 1 && stride <= 2, "conv2d_wgrad: stride 1 or 2");
    B200DP_REQUIRE(((int64_t)R * S * Cin * 4) % 16 == 0, "conv2d_wgrad: weight row pitch must be a multiple of 16 bytes");
    bind_context_once();
    const int is_bf16 = dtype == B200DP_BF16;
    const int P = (H + 2 * pad - R) / stride + 1, Q = (W + 2 * pad - S) / stride + 1;
    WGeom g;
    g.n_img = N; g.out_h = P; g.out_w = Q; g.c_out = Cout; g.c_in = Cin; g.R = R; g.S = S; g.stride = stride; g.pad = pad;
    g.tiles_q = (Q + 7) / 8; g.tiles_p = (P + 7) / 8;
    const int num_kb = N * g.tiles_p * g.tiles_q;
    cudaStream_t s = (cudaStream_t)stream;
    if (R == 3 && stride == 1 && Cin % 64 == 0 && num_kb >= 16) {
        // row-grouped kernel: work item = (s, co tile, ci tile, split), three taps per item
        const int BNr = Cin <= 64 ? 64 : 128;
        const int base = S * ((Cout + kBM - 1) / kBM) * ((Cin + BNr - 1) / BNr);
        // items are long on purpose (single-buffered accumulators): pick the split count that fills whole
        // waves of 148 CTAs best while keeping >= 16 pixel boxes per item
        int best_splits = 1; double best_eff = 0.0;
        for (int w = 1; w <= 3; ++w) {
            int sp = (B200DP_SMS * w) / base;
            if (sp < 1) sp = 1;
            if (sp > num_kb / 16) sp = num_kb / 16;
            if (sp < 1) sp = 1;
            const int kbp = (num_kb + sp - 1) / sp;
            sp = (num_kb + kbp - 1) / kbp;
            const int items = base * sp;
            const int waves = (items + B200DP_SMS - 1) / B200DP_SMS;
            const double eff = (double)items / ((double)waves * B200DP_SMS);
            if (eff > best_eff + 0.02) { best_eff = eff; best_splits = sp; }
        }
        g.kb_per = (num_kb + best_splits - 1) / best_splits;
        g.splits = (num_kb + g.kb_per - 1) / g.kb_per;
        CUtensorMap tdy, tx, tdw;
        int rc = make_act_map(&tdy, dy, is_bf16, N, P, Q, Cout, 8, 8, 1, 1);
        if (rc) return rc;
        rc = make_act_map(&tx, x, is_bf16, N, H, W, Cin, 8, 10, 1, 1);          // 8 x 10 pixel patch (two halo rows)
        if (rc) return rc;
        rc = make_dw_map(&tdw, dw, Cout, R * S * Cin);
        if (rc) return rc;
        if (!accumulate) B200DP_CUDA(cudaMemsetAsync(dw, 0, (size_t)Cout * R * S * Cin * 4, s));
        const int items = base * g.splits;
        const int grid = items < B200DP_SMS ? items : B200DP_SMS;
#define WR(BNv)                                                                                                  \
    {                                                                                                            \
        auto kern = conv_wgrad_rows_kernel<BNv>;                                                                 \
        static bool configured = false;                                                                          \
        if (!configured) {                                                                                       \
            B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, WCfg<BNv>::kSmemBytes)); \
            configured = true;                                                                                   \
        }                                                                                                        \
        kern<<<grid, kThreads, WCfg<BNv>::kSmemBytes, s>>>(tdy, tx, tdw, g, is_bf16);                            \
    }
        if (BNr == 128) WR(128) else WR(64)
#undef WR
        B200DP_CHECK_LAUNCH();
        return 0;
    }
    const int BN = Cin <= 64 ? 64 : (Cin >= 256 ? 256 : 128);
    const int base_items = R * S * ((Cout + kBM - 1) / kBM) * ((Cin + BN - 1) / BN);
    int splits = (2 * B200DP_SMS + base_items - 1) / base_items;         // ~2 waves of work items
    if (splits > num_kb / 4) splits = num_kb / 4;
    if (splits < 1) splits = 1;
    g.kb_per = (num_kb + splits - 1) / splits;
    g.splits = (num_kb + g.kb_per - 1) / g.kb_per;
    CUtensorMap tdy, tx, tdw;
    int rc = make_act_map(&tdy, dy, is_bf16, N, P, Q, Cout, 8, 8, 1, 1);
    if (rc) return rc;
    rc = make_act_map(&tx, x, is_bf16, N, H, W, Cin, 8, 8, 1, stride);
    if (rc) return rc;
    rc = make_dw_map(&tdw, dw, Cout, R * S * Cin);
    if (rc) return rc;
    if (!accumulate) B200DP_CUDA(cudaMemsetAsync(dw, 0, (size_t)Cout * R * S * Cin * 4, s));
    const int items = base_items * g.splits;
    const int grid = items < B200DP_SMS ? items : B200DP_SMS;
#define WG(BNv)                                                                                                  \
    {                                                                                                            \
        auto kern = conv_wgrad_kernel<BNv>;                                                                      \
        static bool configured = false;                                                                          \
        if (!configured) {                                                                                       \
            B200DP_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg<BNv>::kSmemBytes)); \
            configured = true;                                                                                   \
        }                                                                                                        \
        kern<<<grid, kThreads, Cfg<BNv>::kSmemBytes, s>>>(tdy, tx, tdw, g, is_bf16);                             \
    }
    if (BN == 256) WG(256) else if (BN == 128) WG(128) else WG(64)
#undef WG
    B200DP_CHECK_LAUNCH();
    return 0;
}
