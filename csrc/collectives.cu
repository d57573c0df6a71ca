// Peer-memory collectives fused with their neighbouring compute (W1, W2, W3, W5, W6/W11).
//
// Every rank maps every other rank's buffers (CUDA IPC / fabric handles via torch
// symmetric memory) and, when the NVSwitch offers it, one multicast address that fans a
// store out to all ranks and reduces a load across them (NVLS: multimem.st /
// multimem.ld_reduce).  A collective is ONE kernel per rank; ranks meet at flag barriers
// kept in peer-mapped memory.
//
// Barrier protocol.  flags[p] is rank p's array of [max_blocks][B200DP_MAX_RANKS] words.
// CTA b of rank r announces barrier number e by storing e (release, system scope) into
// flags[p][b][r] of every peer p, then spins (acquire, system scope) until its own
// flags[r][b][p] >= e for all p.  e comes from a per-CTA counter in local memory that the
// kernel itself advances, so the protocol is re-entrant, needs no reset, and survives CUDA
// graph replay.  Work is laid out so CTA b only ever consumes what CTA b of a peer
// produced: barriers are per CTA, no CTA waits on another CTA of its own grid, and the
// grid need not be co-resident.  A spin that exceeds timeout_ns gives up, records
// B200DP_ERR_TIMEOUT in the rank's status word and lets the kernel drain (a stuck peer
// must not hang the GPU).
//
// Emulation.  With ctx.emulate the same kernels run all ranks on ONE GPU: gridDim.y is the
// world size, blockIdx.y the emulated rank, every pointer table holds local addresses,
// and the launch is cooperative so the CTAs that wait on each other are co-resident.
// That is how the multi-rank logic is tested on a single GPU (no multicast there).
#include <cooperative_groups.h>

#include "common.cuh"

namespace b200dp {

constexpr int kCommThreads = 512;

// ---------------------------------------------------------------- sys-scope primitives
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint4 ld_peer16(const void* p) {  // coherent 16-byte load of peer memory (bypasses L1)
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_peer16(void* p, uint4 v) {
    asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// NVLS: reduce-on-load across all ranks, fp32 accumulation for the 16-bit types.
template <typename T> __device__ __forceinline__ uint4 multimem_ld_reduce(const void* mc);
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<float>(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<__half>(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<__nv_bfloat16>(const void* mc) {
    uint4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(void* mc, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(v.x), "r"(v.y),
                 "r"(v.z), "r"(v.w) : "memory");
}

// 16 bytes <-> floats
template <typename T> struct Pack;
template <> struct Pack<float> {
    static constexpr int N = 4;
    __device__ __forceinline__ static void unpack(uint4 v, float* f) {
        f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y); f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
    }
    __device__ __forceinline__ static uint4 pack(const float* f) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
    }
};
template <> struct Pack<__half> {
    static constexpr int N = 8;
    __device__ __forceinline__ static void unpack(uint4 v, float* f) {
        const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 t = __half22float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
    }
    __device__ __forceinline__ static uint4 pack(const float* f) {
        uint4 v; __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
        return v;
    }
};
template <> struct Pack<__nv_bfloat16> {
    static constexpr int N = 8;
    __device__ __forceinline__ static void unpack(uint4 v, float* f) {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
    }
    __device__ __forceinline__ static uint4 pack(const float* f) {
        uint4 v; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
        return v;
    }
};

// ---------------------------------------------------------------- barrier
struct Rank {
    int rank, world;
    __device__ __forceinline__ Rank(const b200dp_peer_ctx& c) : rank(c.emulate ? (int)blockIdx.y : c.rank), world(c.world) {}
};

// All threads of the CTA call this.  Orders everything the CTA wrote before the barrier
// ahead of everything peers read after it, and vice versa.
__device__ __forceinline__ void cta_barrier_all_ranks(const b200dp_peer_ctx& c, const Rank& me) {
    if (c.skew_ns != 0 && threadIdx.x == 0) {
        // stress mode (tests): shuffle the arrival order of ranks and CTAs at every barrier
        uint32_t h = (uint32_t)me.rank * 2654435761u ^ blockIdx.x * 40503u ^ (c.epochs[me.rank][blockIdx.x] + 1u) * 2246822519u;
        h ^= h >> 15; h *= 2654435761u; h ^= h >> 13;
        const uint64_t wait = h % c.skew_ns, t0 = globaltimer_ns();
        while (globaltimer_ns() - t0 < wait) __nanosleep(64);
    }
    __syncthreads();
    __shared__ uint32_t s_epoch;
    if (threadIdx.x == 0) {
        uint32_t* ctr = c.epochs[me.rank] + blockIdx.x;
        s_epoch = *ctr + 1u;
        *ctr = s_epoch;
        __threadfence_system();
    }
    __syncthreads();
    const uint32_t e = s_epoch;
    if ((int)threadIdx.x < me.world) {
        const int p = threadIdx.x;
        st_release_sys(c.flags[p] + (size_t)blockIdx.x * B200DP_MAX_RANKS + me.rank, e);
        const uint32_t* mine = c.flags[me.rank] + (size_t)blockIdx.x * B200DP_MAX_RANKS + p;
        const uint64_t t0 = globaltimer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - e) < 0) {
            if (c.timeout_ns != 0 && globaltimer_ns() - t0 > c.timeout_ns) {
                *c.status[me.rank] = (uint32_t)(-B200DP_ERR_TIMEOUT);
                break;
            }
            __nanosleep(20);
        }
    }
    __syncthreads();
}

// Sum over all ranks of U x 16 bytes (element offsets off[u], valid where ok[u]) as floats.
// All loads are issued before the first add so that U x world (P2P) or U (NVLS) 16-byte
// requests per thread are in flight: NVLink latency is ~2 us, and ~2 MB must be outstanding
// per GPU to fill the link.  WMAX bounds the peer loop at compile time (registers).
template <typename T, bool kMC, int WMAX, int U>
__device__ __forceinline__ void reduce16_batch(const Rank& me, void* const* bufs, const void* mc,
                                               const int64_t (&off)[U], const bool (&ok)[U],
                                               float (&acc)[U][Pack<T>::N]) {
    if (kMC) {
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (ok[u]) v[u] = multimem_ld_reduce<T>((const T*)mc + off[u]);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (ok[u]) Pack<T>::unpack(v[u], acc[u]);
    } else {
        uint4 v[U][WMAX];
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int p = 0; p < WMAX; ++p)
                if (ok[u] && p < me.world) v[u][p] = ld_peer16((const T*)bufs[p] + off[u]);
        // peers are added in rank order on every rank => bitwise identical results everywhere
#pragma unroll
        for (int u = 0; u < U; ++u) {
#pragma unroll
            for (int k = 0; k < Pack<T>::N; ++k) acc[u][k] = 0.f;
#pragma unroll
            for (int p = 0; p < WMAX; ++p)
                if (ok[u] && p < me.world) {
                    float f[Pack<T>::N];
                    Pack<T>::unpack(v[u][p], f);
#pragma unroll
                    for (int k = 0; k < Pack<T>::N; ++k) acc[u][k] += f[k];
                }
        }
    }
}

template <bool kMC, int WMAX> struct Unroll { static constexpr int U = kMC ? 8 : (8 / WMAX); };

struct PtrTable { void* p[B200DP_MAX_RANKS]; };
struct CPtrTable { const void* p[B200DP_MAX_RANKS]; };

__device__ __forceinline__ void slice_range(int64_t n_vec, int parts, int idx, int64_t& b, int64_t& e) {
    const int64_t per = (n_vec + parts - 1) / parts;
    b = min(n_vec, per * idx);
    e = min(n_vec, b + per);
}

// ---------------------------------------------------------------- barrier-only kernel (tests the protocol)
__global__ void __launch_bounds__(kCommThreads) peer_barrier_kernel(b200dp_peer_ctx c, int rounds) {
    const Rank me(c);
    for (int i = 0; i < rounds; ++i) cta_barrier_all_ranks(c, me);
}

// ---------------------------------------------------------------- all-reduce (+unscale, +inf check), in place
// algo 0: one-shot   every rank reads everything and writes its own copy (small buckets)
// algo 1: two-shot   rank r reduces slice r, then every rank gathers all slices
// kMC: NVLS — slice r is reduced in the switch and multicast-stored to every rank
template <typename T, bool kMC, int WMAX>
__global__ void __launch_bounds__(kCommThreads) peer_allreduce_kernel(
    b200dp_peer_ctx c, PtrTable bufs, void* mc, int64_t n, float pre_mult, CPtrTable mult_dev, PtrTable found_inf,
    int algo) {
    const Rank me(c);
    constexpr int V = Pack<T>::N;
    constexpr int U = Unroll<kMC, WMAX>::U;
    const int64_t n_vec = n / V;  // n is a multiple of V (buckets are padded)
    const float* md = (const float*)mult_dev.p[me.rank];
    const float mult = pre_mult * (md ? *md : 1.f);
    bool bad = false;
    T* mine = (T*)bufs.p[me.rank];
    cta_barrier_all_ranks(c, me);  // every rank's bucket is complete and visible
    if (algo == 0 && !kMC) {
        int64_t b, e;
        slice_range(n_vec, gridDim.x, blockIdx.x, b, e);
        // results go to registers first: nobody may overwrite a bucket while peers still read it
        for (int64_t i0 = b; i0 < e; i0 += (int64_t)kCommThreads * U) {
            float acc[U][V];
            int64_t off[U];
            bool ok[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
                ok[u] = i < e;
                off[u] = i * V;
            }
            reduce16_batch<T, false, WMAX, U>(me, bufs.p, nullptr, off, ok, acc);
            cta_barrier_all_ranks(c, me);  // all peers have read this chunk group
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (ok[u]) {
#pragma unroll
                    for (int k = 0; k < V; ++k) { acc[u][k] *= mult; bad |= !isfinite(acc[u][k]); }
                    *reinterpret_cast<uint4*>(mine + off[u]) = Pack<T>::pack(acc[u]);
                }
        }
    } else {
        // phase 1: reduce my slice; CTA b owns sub-range b of every slice on every rank
        int64_t sb, se, b, e;
        slice_range(n_vec, me.world, me.rank, sb, se);
        slice_range(se - sb, gridDim.x, blockIdx.x, b, e);
        b += sb; e += sb;
        for (int64_t i0 = b; i0 < e; i0 += (int64_t)kCommThreads * U) {
            float acc[U][V];
            int64_t off[U];
            bool ok[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
                ok[u] = i < e;
                off[u] = i * V;
            }
            reduce16_batch<T, kMC, WMAX, U>(me, bufs.p, mc, off, ok, acc);
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (ok[u]) {
#pragma unroll
                    for (int k = 0; k < V; ++k) { acc[u][k] *= mult; bad |= !isfinite(acc[u][k]); }
                    const uint4 r = Pack<T>::pack(acc[u]);
                    if (kMC) multimem_st((T*)mc + off[u], r);               // lands in every rank's bucket
                    else *reinterpret_cast<uint4*>(mine + off[u]) = r;      // slice r of my own bucket
                }
        }
        // NVLS: only the owner of a slice sees its reduced values, but the non-finite flag has to come up on
        // EVERY rank (all ranks must skip the same optimizer step).  The owner posts the epoch of the barrier
        // below into its slot of the inf row of this CTA on every rank; the barrier's release / acquire pair
        // publishes it, and each rank then compares the slots with that epoch (no reset ever needed).
        uint32_t inf_epoch = 0;
        if (kMC && found_inf.p[me.rank] != nullptr) {
            inf_epoch = c.epochs[me.rank][blockIdx.x] + 1u;       // written by thread 0 inside the barrier only
            const bool cta_bad = __syncthreads_or(bad);
            if (cta_bad && (int)threadIdx.x < me.world)
                st_release_sys(c.flags[threadIdx.x] + ((size_t)c.max_blocks + blockIdx.x) * B200DP_MAX_RANKS + me.rank, inf_epoch);
        }
        cta_barrier_all_ranks(c, me);  // every slice is reduced (and, with NVLS, already delivered)
        if (kMC && found_inf.p[me.rank] != nullptr && (int)threadIdx.x < me.world)
            bad |= ld_acquire_sys(c.flags[me.rank] + ((size_t)c.max_blocks + blockIdx.x) * B200DP_MAX_RANKS + threadIdx.x) == inf_epoch;
        if (!kMC) {
            // phase 2: gather the other slices from their owners, 4 loads in flight per thread
            for (int p = 1; p < me.world; ++p) {
                const int src = (me.rank + p) % me.world;
                slice_range(n_vec, me.world, src, sb, se);
                slice_range(se - sb, gridDim.x, blockIdx.x, b, e);
                b += sb; e += sb;
                const T* from = (const T*)bufs.p[src];
                for (int64_t i0 = b; i0 < e; i0 += (int64_t)kCommThreads * 4) {
                    uint4 r[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
                        if (i < e) r[u] = ld_peer16(from + i * V);
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
                        if (i < e) {
                            float f[V];
                            Pack<T>::unpack(r[u], f);
#pragma unroll
                            for (int k = 0; k < V; ++k) bad |= !isfinite(f[k]);
                            *reinterpret_cast<uint4*>(mine + i * V) = r[u];
                        }
                    }
                }
            }
            cta_barrier_all_ranks(c, me);  // nobody still reads a slice its owner may now overwrite
        }
    }
    float* fi = (float*)found_inf.p[me.rank];
    if (fi != nullptr && __syncthreads_or(bad) && threadIdx.x == 0) *fi = 1.f;
}

// ---------------------------------------------------------------- reduce-scatter (+scale, fp32 shard out)
// out[rank][i] (=|+=) mult * sum_p in[p][rank*n_per + i]
template <typename T, bool kMC, int WMAX>
__global__ void __launch_bounds__(kCommThreads) peer_reduce_scatter_kernel(
    b200dp_peer_ctx c, PtrTable in_bufs, void* mc, int64_t n_per, PtrTable outs, float pre_mult, CPtrTable mult_dev,
    PtrTable found_inf, int accumulate) {
    const Rank me(c);
    constexpr int V = Pack<T>::N;
    constexpr int U = Unroll<kMC, WMAX>::U;
    const int64_t n_vec = n_per / V;
    const float* md = (const float*)mult_dev.p[me.rank];
    const float mult = pre_mult * (md ? *md : 1.f);
    float* out = (float*)outs.p[me.rank];
    bool bad = false;
    cta_barrier_all_ranks(c, me);
    int64_t b, e;
    slice_range(n_vec, gridDim.x, blockIdx.x, b, e);
    const int64_t base = (int64_t)me.rank * n_per;
    for (int64_t i0 = b; i0 < e; i0 += (int64_t)kCommThreads * U) {
        float acc[U][V];
        int64_t off[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
            ok[u] = i < e;
            off[u] = base + i * V;
        }
        reduce16_batch<T, kMC, WMAX, U>(me, in_bufs.p, mc, off, ok, acc);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (ok[u]) {
#pragma unroll
                for (int k = 0; k < V; ++k) { acc[u][k] *= mult; bad |= !isfinite(acc[u][k]); }
                float* o = out + (off[u] - base);
#pragma unroll
                for (int k = 0; k < V; k += 4) {
                    float4 t = make_float4(acc[u][k], acc[u][k + 1], acc[u][k + 2], acc[u][k + 3]);
                    if (accumulate) {
                        const float4 old = *reinterpret_cast<float4*>(o + k);
                        t.x += old.x; t.y += old.y; t.z += old.z; t.w += old.w;
                    }
                    *reinterpret_cast<float4*>(o + k) = t;
                }
            }
    }
    cta_barrier_all_ranks(c, me);  // peers are done reading my input: it may be reused
    float* fi = (float*)found_inf.p[me.rank];
    if (fi != nullptr && __syncthreads_or(bad) && threadIdx.x == 0) *fi = 1.f;
}

// ---------------------------------------------------------------- all-gather (+cast), push based
// every rank writes cast<TO>(shard[rank][i]) into out[p][rank*n_per + i] of every rank p
template <typename TI, typename TO, bool kMC>
__global__ void __launch_bounds__(kCommThreads) peer_all_gather_kernel(b200dp_peer_ctx c, PtrTable outs, void* mc,
                                                                      CPtrTable shards, int64_t n_per) {
    const Rank me(c);
    constexpr int V = Pack<TO>::N;   // elements per 16-byte store
    const int64_t n_vec = n_per / V;
    const TI* src = (const TI*)shards.p[me.rank];
    cta_barrier_all_ranks(c, me);    // every rank is past its last use of the previous contents
    int64_t b, e;
    slice_range(n_vec, gridDim.x, blockIdx.x, b, e);
    const int64_t base = (int64_t)me.rank * n_per;
    for (int64_t i = b + threadIdx.x; i < e; i += kCommThreads) {
        float f[V];
        if (sizeof(TI) == 4) {
#pragma unroll
            for (int k = 0; k < V; k += 4) {
                const float4 t = *reinterpret_cast<const float4*>((const float*)src + i * V + k);
                f[k] = t.x; f[k + 1] = t.y; f[k + 2] = t.z; f[k + 3] = t.w;
            }
        } else {
            Pack<TO>::unpack(*reinterpret_cast<const uint4*>((const TO*)src + i * V), f);
        }
        const uint4 r = Pack<TO>::pack(f);
        if (kMC) {
            multimem_st((TO*)mc + base + i * V, r);
        } else {
#pragma unroll
            for (int p = 0; p < B200DP_MAX_RANKS; ++p)
                if (p < me.world) {
                    const int dst = (me.rank + p) % me.world;
                    st_peer16((TO*)outs.p[dst] + base + i * V, r);
                }
        }
    }
    cta_barrier_all_ranks(c, me);    // every shard has landed everywhere
}

// ---------------------------------------------------------------- ZeRO step: reduce-to-owner + AdamW + publish
// The flat gradient / parameter buffers are ordered by owner: rank r owns elements
// [seg[r], seg[r+1]).  One kernel on the owner: reduce the segment across ranks (or take it
// pre-reduced), apply AdamW to the fp32 master / moments it alone holds, and publish the new
// values into every rank's parameter buffer (multicast store, or one peer store per rank).
struct SegTable { int64_t s[B200DP_MAX_RANKS + 1]; };

template <typename TW, bool kMC, int WMAX>   // TW: wire / parameter-buffer dtype
__global__ void __launch_bounds__(kCommThreads) peer_zero_adam_kernel(
    b200dp_peer_ctx c, PtrTable grad_bufs, void* grad_mc, PtrTable param_bufs, void* param_mc, SegTable seg,
    PtrTable master, PtrTable m_tab, PtrTable v_tab, b200dp_adam_hparams hp, CPtrTable step_tab, float pre_mult,
    CPtrTable mult_dev, CPtrTable found_inf, int prereduced, CPtrTable local_grad) {
    const Rank me(c);
    constexpr int V = 4;  // fp32 gradients: 4 per 16 bytes
    constexpr int U = kMC ? 4 : (WMAX >= 8 ? 1 : 4 / WMAX);   // fewer in flight than the pure reduce: state traffic shares the registers
    const int64_t s0 = seg.s[me.rank], s1 = seg.s[me.rank + 1];
    const int64_t n_vec = (s1 - s0) / V;   // segments are padded to multiples of 8
    const float* md = (const float*)mult_dev.p[me.rank];
    const float mult = pre_mult * (md ? *md : 1.f);
    const float* fi = (const float*)found_inf.p[me.rank];
    const bool skip = fi != nullptr && *fi != 0.f;       // identical on all ranks (flag is all-reduced)
    float* pm = (float*)master.p[me.rank];
    float* mm = (float*)m_tab.p[me.rank];
    float* vm = (float*)v_tab.p[me.rank];
    const float step = *(const float*)step_tab.p[me.rank];
    const float bc1 = 1.f - powf(hp.beta1, step), bc2 = 1.f - powf(hp.beta2, step);
    const float lr = hp.lr_dev ? *hp.lr_dev : hp.lr;
    const float step_size = lr / bc1, inv_sqrt_bc2 = rsqrtf(bc2);
    const bool sgd = hp.adamw == 2;                      // the same fused step for SGD (momentum / Nesterov / L2): see below
    const float decay = hp.adamw == 1 ? 1.f - lr * hp.weight_decay : 1.f, l2 = hp.adamw == 1 ? 0.f : hp.weight_decay;
    if (!prereduced) cta_barrier_all_ranks(c, me);   // all ranks' gradients are complete
    int64_t b, e;
    slice_range(n_vec, gridDim.x, blockIdx.x, b, e);
    if (!skip) {
        for (int64_t i0 = b; i0 < e; i0 += (int64_t)kCommThreads * U) {
            float g[U][V];
            int64_t off[U];
            bool ok[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
                ok[u] = i < e;
                off[u] = s0 + i * V;
            }
            if (prereduced) {
#pragma unroll
                for (int u = 0; u < U; ++u)
                    if (ok[u]) {
                        const float4 t = *reinterpret_cast<const float4*>((const float*)local_grad.p[me.rank] + (off[u] - s0));
                        g[u][0] = t.x; g[u][1] = t.y; g[u][2] = t.z; g[u][3] = t.w;
                    }
            } else {
                reduce16_batch<float, kMC, WMAX, U>(me, grad_bufs.p, grad_mc, off, ok, g);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!ok[u]) continue;
                const int64_t li = off[u] - s0;     // element index inside the owned segment
                float4 p4 = *reinterpret_cast<float4*>(pm + li), m4 = *reinterpret_cast<float4*>(mm + li),
                       v4 = sgd ? make_float4(0.f, 0.f, 0.f, 0.f) : *reinterpret_cast<float4*>(vm + li);
                float pv[V] = {p4.x, p4.y, p4.z, p4.w}, mv[V] = {m4.x, m4.y, m4.z, m4.w}, vv[V] = {v4.x, v4.y, v4.z, v4.w};
                if (sgd) {
                    // hp: beta1 = momentum, beta2 = dampening, eps != 0 selects Nesterov, weight_decay is L2; the first
                    // step seeds the momentum buffer with the gradient (torch.optim.SGD)
#pragma unroll
                    for (int k = 0; k < V; ++k) {
                        float gv = g[u][k] * mult + hp.weight_decay * pv[k];
                        if (hp.beta1 != 0.f) {
                            mv[k] = step <= 1.f ? gv : hp.beta1 * mv[k] + (1.f - hp.beta2) * gv;
                            gv = hp.eps != 0.f ? gv + hp.beta1 * mv[k] : mv[k];
                        }
                        pv[k] -= lr * gv;
                    }
                } else {
#pragma unroll
                    for (int k = 0; k < V; ++k) {
                        const float gv = g[u][k] * mult + l2 * pv[k];
                        pv[k] *= decay;
                        mv[k] = hp.beta1 * mv[k] + (1.f - hp.beta1) * gv;
                        vv[k] = hp.beta2 * vv[k] + (1.f - hp.beta2) * gv * gv;
                        pv[k] -= step_size * (mv[k] / (sqrtf(vv[k]) * inv_sqrt_bc2 + hp.eps));
                    }
                    *reinterpret_cast<float4*>(vm + li) = make_float4(vv[0], vv[1], vv[2], vv[3]);
                }
                *reinterpret_cast<float4*>(pm + li) = make_float4(pv[0], pv[1], pv[2], pv[3]);
                *reinterpret_cast<float4*>(mm + li) = make_float4(mv[0], mv[1], mv[2], mv[3]);
                // publish (W3): fp32 wire = one 16-byte store per 4 values; 16-bit wire = 8 bytes
                if (sizeof(TW) == 4) {
                    const uint4 r = Pack<float>::pack(pv);
                    if (kMC) multimem_st((float*)param_mc + off[u], r);
                    else {
#pragma unroll
                        for (int p = 0; p < WMAX; ++p)
                            if (p < me.world) st_peer16((float*)param_bufs.p[(me.rank + p) % me.world] + off[u], r);
                    }
                } else {
                    TW w[V];
#pragma unroll
                    for (int k = 0; k < V; ++k) w[k] = Cvt<TW>::from_f(pv[k]);
                    const uint2 r = *reinterpret_cast<const uint2*>(w);
#pragma unroll
                    for (int p = 0; p < WMAX; ++p)
                        if (p < me.world) {
                            TW* dst = (TW*)param_bufs.p[(me.rank + p) % me.world] + off[u];
                            asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1,%2};" ::"l"(dst), "r"(r.x), "r"(r.y) : "memory");
                        }
                }
            }
        }
    }
    cta_barrier_all_ranks(c, me);   // every rank's parameters are complete; gradients are free again
}

// reduce-to-owner only (used when clipping / AMP need the global norm and inf flag before the
// update): local_out[rank][i] = mult * sum_p grad[p][seg[rank]+i]; sum of squares -> norm_out
template <int WMAX>
__global__ void __launch_bounds__(kCommThreads) peer_reduce_to_owner_kernel(
    b200dp_peer_ctx c, PtrTable grad_bufs, SegTable seg, PtrTable local_out, float pre_mult, CPtrTable mult_dev,
    PtrTable norm_out, PtrTable partials, PtrTable counters) {
    __shared__ float red[32];
    __shared__ bool is_last;
    const Rank me(c);
    constexpr int V = 4;
    constexpr int U = Unroll<false, WMAX>::U;
    const int64_t s0 = seg.s[me.rank], s1 = seg.s[me.rank + 1];
    const int64_t n_vec = (s1 - s0) / V;
    const float* md = (const float*)mult_dev.p[me.rank];
    const float mult = pre_mult * (md ? *md : 1.f);
    float* out = (float*)local_out.p[me.rank];
    cta_barrier_all_ranks(c, me);
    int64_t b, e;
    slice_range(n_vec, gridDim.x, blockIdx.x, b, e);
    float acc2 = 0.f;
    bool bad = false;
    for (int64_t i0 = b; i0 < e; i0 += (int64_t)kCommThreads * U) {
        float g[U][V];
        int64_t off[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int64_t i = i0 + (int64_t)u * kCommThreads + threadIdx.x;
            ok[u] = i < e;
            off[u] = s0 + i * V;
        }
        reduce16_batch<float, false, WMAX, U>(me, grad_bufs.p, nullptr, off, ok, g);
#pragma unroll
        for (int u = 0; u < U; ++u)
            if (ok[u]) {
#pragma unroll
                for (int k = 0; k < V; ++k) { g[u][k] *= mult; bad |= !isfinite(g[u][k]); acc2 += g[u][k] * g[u][k]; }
                *reinterpret_cast<float4*>(out + (off[u] - s0)) = make_float4(g[u][0], g[u][1], g[u][2], g[u][3]);
            }
    }
    cta_barrier_all_ranks(c, me);
    const float tot = block_sum(acc2, red);
    const bool any_bad = __syncthreads_or(bad);
    float* part = (float*)partials.p[me.rank];
    unsigned int* ctr = (unsigned int*)counters.p[me.rank];
    float* no = (float*)norm_out.p[me.rank];
    if (threadIdx.x == 0) {
        part[blockIdx.x] = tot;
        if (any_bad) no[1] = 1.f;
        __threadfence();
        is_last = (atomicAdd(ctr, 1u) == gridDim.x - 1u);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float s = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += kCommThreads) s += __ldcg(part + i);
        s = block_sum(s, red);
        if (threadIdx.x == 0) { no[0] = s; *ctr = 0u; }
    }
}

// ---------------------------------------------------------------- broadcast (root -> all), raw bytes
// Module buffers (BatchNorm running statistics) replicated from the root every forward
// (K11 hot half): one multicast store per 16 bytes with NVLS, else one peer store per rank.
template <bool kMC>
__global__ void __launch_bounds__(kCommThreads) peer_broadcast_kernel(b200dp_peer_ctx c, PtrTable bufs, void* mc,
                                                                     int64_t n_vec16, int root) {
    const Rank me(c);
    cta_barrier_all_ranks(c, me);            // everyone is done reading the previous contents
    if (me.rank == root) {
        int64_t b, e;
        slice_range(n_vec16, gridDim.x, blockIdx.x, b, e);
        const uint4* src = (const uint4*)bufs.p[root];
        for (int64_t i = b + threadIdx.x; i < e; i += kCommThreads) {
            const uint4 v = src[i];
            if (kMC) {
                multimem_st((uint4*)mc + i, v);
            } else {
#pragma unroll
                for (int p = 0; p < B200DP_MAX_RANKS; ++p)
                    if (p < me.world && p != root) st_peer16((uint4*)bufs.p[p] + i, v);
            }
        }
    }
    cta_barrier_all_ranks(c, me);            // the root's stores have landed everywhere
}

// ---------------------------------------------------------------- small one-shot all-reduce (<= 64 KB fp32)
// Double-buffered staging slots in peer memory: call k uses slot k&1, so a single barrier
// per call is enough (a peer can only be two calls behind after passing that barrier).
__global__ void __launch_bounds__(kCommThreads) peer_small_allreduce_kernel(b200dp_peer_ctx c, PtrTable staging,
                                                                           PtrTable io, int n, int slot_floats,
                                                                           PtrTable call_counter) {
    const Rank me(c);
    unsigned int* cc = (unsigned int*)call_counter.p[me.rank];
    const unsigned int k = *cc;
    float* data = (float*)io.p[me.rank];
    float* my_slot = (float*)staging.p[me.rank] + (size_t)(k & 1u) * slot_floats;
    for (int i = threadIdx.x; i < n; i += kCommThreads) my_slot[i] = data[i];
    __threadfence_system();
    cta_barrier_all_ranks(c, me);
    for (int i = threadIdx.x; i < n; i += kCommThreads) {
        float s = 0.f;
        for (int p = 0; p < me.world; ++p) {
            const float* slot = (const float*)staging.p[p] + (size_t)(k & 1u) * slot_floats;
            float v;
            asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(slot + i) : "memory");
            s += v;
        }
        data[i] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) *cc = k + 1u;
}

// ---------------------------------------------------------------- launch helper
template <typename... Args>
static int launch_comm(void (*kernel)(Args...), const b200dp_peer_ctx& ctx, int blocks, cudaStream_t s, Args... args) {
    dim3 grid(blocks, ctx.emulate ? ctx.world : 1);
    if (ctx.emulate) {
        void* params[] = {(void*)&args...};
        B200DP_CUDA(cudaLaunchCooperativeKernel((const void*)kernel, grid, dim3(kCommThreads), params, 0, s));
    } else {
        kernel<<<grid, kCommThreads, 0, s>>>(args...);
        B200DP_CHECK_LAUNCH();
    }
    return 0;
}

static int check_ctx(const b200dp_peer_ctx* c, int blocks) {
    B200DP_REQUIRE(c != nullptr, "peer: null context");
    B200DP_REQUIRE(c->world >= 1 && c->world <= B200DP_MAX_RANKS, "peer: world %d out of range", c->world);
    B200DP_REQUIRE(c->emulate || (c->rank >= 0 && c->rank < c->world), "peer: bad rank %d", c->rank);
    B200DP_REQUIRE(blocks >= 1 && blocks <= c->max_blocks, "peer: %d CTAs requested, context has flags for %d", blocks,
                   c->max_blocks);
    return 0;
}

static void fill(PtrTable& t, void* const* src) { for (int i = 0; i < B200DP_MAX_RANKS; ++i) t.p[i] = src ? src[i] : nullptr; }
static void fill(CPtrTable& t, const void* const* src) { for (int i = 0; i < B200DP_MAX_RANKS; ++i) t.p[i] = src ? src[i] : nullptr; }

}  // namespace b200dp

using namespace b200dp;

extern "C" {

int b200dp_peer_barrier(const b200dp_peer_ctx* ctx, int blocks, int rounds, void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    return launch_comm(peer_barrier_kernel, *ctx, blocks, (cudaStream_t)stream, *ctx, rounds);
}

int b200dp_peer_allreduce(const b200dp_peer_ctx* ctx, void* const* bufs, void* mc, int64_t n, b200dp_dtype dtype,
                          float pre_mult, const void* const* mult_dev, void* const* found_inf, int algo, int blocks,
                          void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    B200DP_REQUIRE(bufs != nullptr && n > 0, "peer_allreduce: bad arguments");
    B200DP_REQUIRE(n % (dtype == B200DP_F32 ? 4 : 8) == 0, "peer_allreduce: n must be a multiple of 16 bytes");
    B200DP_REQUIRE(mc == nullptr || !ctx->emulate, "peer_allreduce: multicast cannot be emulated on one GPU");
    PtrTable tb, tf; CPtrTable tm;
    fill(tb, bufs); fill(tf, found_inf); fill(tm, mult_dev);
    cudaStream_t s = (cudaStream_t)stream;
#define AR(W)                                                                                                        \
    B200DP_DISPATCH_DTYPE(dtype, T, {                                                                               \
        if (mc != nullptr) rc = launch_comm(peer_allreduce_kernel<T, true, W>, *ctx, blocks, s, *ctx, tb, mc, n, pre_mult, tm, tf, 1); \
        else rc = launch_comm(peer_allreduce_kernel<T, false, W>, *ctx, blocks, s, *ctx, tb, mc, n, pre_mult, tm, tf, algo); \
    })
    if (ctx->world <= 2) { AR(2); } else if (ctx->world <= 4) { AR(4); } else { AR(8); }
#undef AR
    return rc;
}

int b200dp_peer_reduce_scatter(const b200dp_peer_ctx* ctx, void* const* in_bufs, void* mc, int64_t n_per_rank,
                               b200dp_dtype dtype, void* const* outs, float pre_mult, const void* const* mult_dev,
                               void* const* found_inf, int accumulate, int blocks, void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    B200DP_REQUIRE(in_bufs && outs && n_per_rank > 0 && n_per_rank % 8 == 0,
                   "peer_reduce_scatter: n_per_rank must be a positive multiple of 8");
    B200DP_REQUIRE(mc == nullptr || !ctx->emulate, "peer_reduce_scatter: multicast cannot be emulated on one GPU");
    PtrTable ti, to, tf; CPtrTable tm;
    fill(ti, in_bufs); fill(to, outs); fill(tf, found_inf); fill(tm, mult_dev);
    cudaStream_t s = (cudaStream_t)stream;
#define RS(W)                                                                                                        \
    B200DP_DISPATCH_DTYPE(dtype, T, {                                                                               \
        if (mc != nullptr) rc = launch_comm(peer_reduce_scatter_kernel<T, true, W>, *ctx, blocks, s, *ctx, ti, mc, n_per_rank, to, pre_mult, tm, tf, accumulate); \
        else rc = launch_comm(peer_reduce_scatter_kernel<T, false, W>, *ctx, blocks, s, *ctx, ti, mc, n_per_rank, to, pre_mult, tm, tf, accumulate); \
    })
    if (ctx->world <= 2) { RS(2); } else if (ctx->world <= 4) { RS(4); } else { RS(8); }
#undef RS
    return rc;
}

int b200dp_peer_all_gather(const b200dp_peer_ctx* ctx, void* const* outs, void* mc, const void* const* shards,
                           int64_t n_per_rank, b200dp_dtype shard_dtype, b200dp_dtype out_dtype, int blocks,
                           void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    B200DP_REQUIRE(outs && shards && n_per_rank > 0 && n_per_rank % 8 == 0,
                   "peer_all_gather: n_per_rank must be a positive multiple of 8");
    B200DP_REQUIRE(mc == nullptr || !ctx->emulate, "peer_all_gather: multicast cannot be emulated on one GPU");
    B200DP_REQUIRE(shard_dtype == B200DP_F32 || shard_dtype == out_dtype,
                   "peer_all_gather: shard must be fp32 (cast on the fly) or already in the output dtype");
    PtrTable to; CPtrTable ts;
    fill(to, outs); fill(ts, shards);
    cudaStream_t s = (cudaStream_t)stream;
#define AG(TI, TO)                                                                                              \
    rc = (mc != nullptr) ? launch_comm(peer_all_gather_kernel<TI, TO, true>, *ctx, blocks, s, *ctx, to, mc, ts, n_per_rank) \
                         : launch_comm(peer_all_gather_kernel<TI, TO, false>, *ctx, blocks, s, *ctx, to, mc, ts, n_per_rank)
    if (out_dtype == B200DP_F32) { AG(float, float); }
    else if (out_dtype == B200DP_F16) { if (shard_dtype == B200DP_F32) { AG(float, __half); } else { AG(__half, __half); } }
    else { if (shard_dtype == B200DP_F32) { AG(float, __nv_bfloat16); } else { AG(__nv_bfloat16, __nv_bfloat16); } }
#undef AG
    return rc;
}

int b200dp_peer_zero_adam(const b200dp_peer_ctx* ctx, void* const* grad_bufs, void* grad_mc, void* const* param_bufs,
                          void* param_mc, const int64_t* seg, b200dp_dtype wire_dtype, void* const* master,
                          void* const* m, void* const* v, b200dp_adam_hparams hp, const void* const* step,
                          float pre_mult, const void* const* mult_dev, const void* const* found_inf, int prereduced,
                          const void* const* local_grad, int blocks, void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    B200DP_REQUIRE(param_bufs && seg && master && m && v && step, "peer_zero_adam: null argument");
    B200DP_REQUIRE(prereduced ? local_grad != nullptr : grad_bufs != nullptr, "peer_zero_adam: no gradient source");
    B200DP_REQUIRE((grad_mc == nullptr && param_mc == nullptr) || !ctx->emulate,
                   "peer_zero_adam: multicast cannot be emulated on one GPU");
    SegTable st;
    for (int i = 0; i <= B200DP_MAX_RANKS; ++i) st.s[i] = i <= ctx->world ? seg[i] : seg[ctx->world];
    for (int i = 0; i <= ctx->world; ++i)
        B200DP_REQUIRE(st.s[i] % 8 == 0, "peer_zero_adam: segment boundaries must be multiples of 8 elements");
    PtrTable tg, tp, tma, tm_, tv; CPtrTable ts, tmu, tfi, tlg;
    fill(tg, grad_bufs); fill(tp, param_bufs); fill(tma, master); fill(tm_, m); fill(tv, v);
    fill(ts, step); fill(tmu, mult_dev); fill(tfi, found_inf); fill(tlg, local_grad);
    cudaStream_t s = (cudaStream_t)stream;
    const bool mc = (grad_mc != nullptr || prereduced) && param_mc != nullptr && wire_dtype == B200DP_F32 && !ctx->emulate &&
                    (prereduced || grad_mc != nullptr);
#define ZA(TW, MC, W) rc = launch_comm(peer_zero_adam_kernel<TW, MC, W>, *ctx, blocks, s, *ctx, tg, grad_mc, tp, param_mc, st, tma, tm_, tv, hp, ts, pre_mult, tmu, tfi, prereduced, tlg)
#define ZAW(W)                                                                              \
    if (wire_dtype == B200DP_F32) { if (mc) { ZA(float, true, W); } else { ZA(float, false, W); } } \
    else if (wire_dtype == B200DP_F16) { ZA(__half, false, W); }                            \
    else { ZA(__nv_bfloat16, false, W); }
    if (ctx->world <= 2) { ZAW(2) } else if (ctx->world <= 4) { ZAW(4) } else { ZAW(8) }
#undef ZAW
#undef ZA
    return rc;
}

int b200dp_peer_reduce_to_owner(const b200dp_peer_ctx* ctx, void* const* grad_bufs, const int64_t* seg,
                                void* const* local_out, float pre_mult, const void* const* mult_dev,
                                void* const* norm_out, void* const* partials, void* const* counters, int blocks,
                                void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    B200DP_REQUIRE(grad_bufs && seg && local_out && norm_out && partials && counters, "peer_reduce_to_owner: null argument");
    SegTable st;
    for (int i = 0; i <= B200DP_MAX_RANKS; ++i) st.s[i] = i <= ctx->world ? seg[i] : seg[ctx->world];
    PtrTable tg, tl, tn, tpa, tc; CPtrTable tmu;
    fill(tg, grad_bufs); fill(tl, local_out); fill(tn, norm_out); fill(tpa, partials); fill(tc, counters); fill(tmu, mult_dev);
    cudaStream_t s = (cudaStream_t)stream;
    if (ctx->world <= 2) return launch_comm(peer_reduce_to_owner_kernel<2>, *ctx, blocks, s, *ctx, tg, st, tl, pre_mult, tmu, tn, tpa, tc);
    if (ctx->world <= 4) return launch_comm(peer_reduce_to_owner_kernel<4>, *ctx, blocks, s, *ctx, tg, st, tl, pre_mult, tmu, tn, tpa, tc);
    return launch_comm(peer_reduce_to_owner_kernel<8>, *ctx, blocks, s, *ctx, tg, st, tl, pre_mult, tmu, tn, tpa, tc);
}

int b200dp_peer_broadcast(const b200dp_peer_ctx* ctx, void* const* bufs, void* mc, int64_t nbytes, int root, int blocks,
                          void* stream) {
    int rc = check_ctx(ctx, blocks);
    if (rc) return rc;
    B200DP_REQUIRE(bufs && nbytes > 0 && nbytes % 16 == 0, "peer_broadcast: size must be a positive multiple of 16 bytes");
    B200DP_REQUIRE(root >= 0 && root < ctx->world, "peer_broadcast: bad root %d", root);
    B200DP_REQUIRE(mc == nullptr || !ctx->emulate, "peer_broadcast: multicast cannot be emulated on one GPU");
    PtrTable tb;
    fill(tb, bufs);
    cudaStream_t s = (cudaStream_t)stream;
    if (mc != nullptr) return launch_comm(peer_broadcast_kernel<true>, *ctx, blocks, s, *ctx, tb, mc, nbytes / 16, root);
    return launch_comm(peer_broadcast_kernel<false>, *ctx, blocks, s, *ctx, tb, mc, nbytes / 16, root);
}

int b200dp_peer_small_allreduce(const b200dp_peer_ctx* ctx, void* const* staging, void* const* io, int n,
                                int slot_floats, void* const* call_counter, void* stream) {
    int rc = check_ctx(ctx, 1);
    if (rc) return rc;
    B200DP_REQUIRE(staging && io && call_counter && n > 0 && n <= slot_floats, "peer_small_allreduce: bad arguments");
    PtrTable ts, ti, tc;
    fill(ts, staging); fill(ti, io); fill(tc, call_counter);
    return launch_comm(peer_small_allreduce_kernel, *ctx, 1, (cudaStream_t)stream, *ctx, ts, ti, n, slot_floats, tc);
}

}  // extern "C"
