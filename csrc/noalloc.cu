// Sum of squares for callers that must not allocate device memory.
//
// Who this is for: code that runs where cudaMalloc is off limits — inside a host
// callback or a latency-critical serving loop where an allocation would synchronise the
// device, inside a process whose memory is fully pre-partitioned by an outer allocator
// (no head-room left to ask the driver for more), or a plain C/Fortran host that links
// this library and has no caching allocator to borrow a workspace from.  The regular
// entry point, b200dp_sumsq_f32, makes the caller bring a scratch buffer and a zeroed
// ticket word; those callers have nowhere to get them.  Here the scratch lives in the
// library image instead: one fixed, file-scope device array that every GPU which loads
// the library carries until the process exits (65,792 bytes per GPU).
//
// Consequences the caller has to live with:
//   * the input is limited to B200DP_NOALLOC_MAX_ELEMS floats (1 GiB of fp32); a larger n
//     is refused with B200DP_ERR_INVALID and nothing is launched;
//   * there is ONE array per GPU, shared by every call: it still holds the previous
//     call's partial sums, so each call clears the word that must start at zero (the
//     completion ticket) on the caller's stream before launching; two calls that can
//     overlap on the same GPU (different streams, or different host threads on different
//     streams) race on the array and give garbage — serialise them on one stream;
//   * the very first call on each GPU looks up the array's device address
//     (cudaGetSymbolAddress, with that GPU current) and caches it; make that first call
//     before starting any stream capture on that GPU.
//
// No cudaMalloc / cudaFree / allocating library routine is called anywhere in this file.
#include <mutex>

#include "common.cuh"

namespace b200dp {
namespace {

constexpr int kNAThreads = 256;
constexpr int kNAPerBlock = 16384;                                   // floats per CTA
constexpr int64_t kNAMaxElems = B200DP_NOALLOC_MAX_ELEMS;            // 2^28
constexpr int kNAMaxBlocks = (int)(kNAMaxElems / kNAPerBlock);       // 16384 partial sums
constexpr int kNATicketWords = 64;                                   // 256 B: ticket word + padding

// [0, 64)  : completion ticket (word 0) padded to 256 B so the partials start aligned
// [64, ...) : one partial sum per CTA
static __device__ __align__(256) float g_noalloc_scratch[kNATicketWords + kNAMaxBlocks];

constexpr int kMaxDevices = 64;
float* g_scratch_addr[kMaxDevices] = {nullptr};   // device address of the array, per GPU
std::mutex g_setup_mutex;

__global__ void __launch_bounds__(kNAThreads) sumsq_noalloc_kernel(const float* __restrict__ x, int64_t n,
                                                                  float* __restrict__ partials,
                                                                  unsigned int* __restrict__ ticket,
                                                                  float* __restrict__ out) {
    __shared__ float red[32];
    __shared__ bool is_last;
    const int64_t start = (int64_t)blockIdx.x * kNAPerBlock;
    const int64_t len = min((int64_t)kNAPerBlock, n - start);
    const float* xp = x + start;
    float acc = 0.f;
    int64_t done = 0;
    if (is_aligned16(xp)) {
        const int64_t nvec = len / 4;
        for (int64_t i = threadIdx.x; i < nvec; i += kNAThreads) {
            const float4 v = reinterpret_cast<const float4*>(xp)[i];
            acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
        }
        done = nvec * 4;
    }
    for (int64_t i = done + threadIdx.x; i < len; i += kNAThreads) acc += xp[i] * xp[i];
    const float tot = block_sum(acc, red);
    if (threadIdx.x == 0) {
        partials[blockIdx.x] = tot;
        __threadfence();
        is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1u);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        float s = 0.f;
        for (int i = threadIdx.x; i < (int)gridDim.x; i += kNAThreads) s += __ldcg(partials + i);
        s = block_sum(s, red);
        if (threadIdx.x == 0) *out = s;
    }
}

// Device address of the static array on the CURRENT device, looked up once per device.
// In host code the name g_noalloc_scratch is only a stand-in for the symbol; it is not
// something a kernel may dereference, hence the explicit address query.
int scratch_for_current_device(float** out) {
    int dev = -1;
    B200DP_CUDA(cudaGetDevice(&dev));
    B200DP_REQUIRE(dev >= 0 && dev < kMaxDevices, "sumsq_f32_noalloc: device index %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_setup_mutex);
    if (g_scratch_addr[dev] == nullptr) {
        float* p = nullptr;
        B200DP_CUDA(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        B200DP_REQUIRE(p != nullptr, "sumsq_f32_noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}

}  // namespace
}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_sumsq_f32_noalloc(const float* x, int64_t n, float* out, void* stream) {
    B200DP_REQUIRE(out != nullptr, "sumsq_f32_noalloc: null out");
    B200DP_REQUIRE(n >= 0, "sumsq_f32_noalloc: negative n");
    B200DP_REQUIRE(n <= kNAMaxElems,
                   "sumsq_f32_noalloc: n = %lld exceeds the built-in scratch (max %lld floats); use "
                   "b200dp_sumsq_f32 with caller-provided scratch for larger inputs",
                   (long long)n, (long long)kNAMaxElems);
    cudaStream_t s = (cudaStream_t)stream;
    if (n == 0) {
        B200DP_CUDA(cudaMemsetAsync(out, 0, sizeof(float), s));
        return 0;
    }
    B200DP_REQUIRE(x != nullptr, "sumsq_f32_noalloc: null input");
    float* base = nullptr;
    const int rc = scratch_for_current_device(&base);
    if (rc != 0) return rc;
    unsigned int* ticket = reinterpret_cast<unsigned int*>(base);
    float* partials = base + kNATicketWords;
    // The array is shared and keeps whatever the last call (or an aborted one) left in it:
    // the ticket is the only word that has to start at zero, so clear it in stream order.
    B200DP_CUDA(cudaMemsetAsync(ticket, 0, sizeof(unsigned int), s));
    const unsigned blocks = (unsigned)((n + kNAPerBlock - 1) / kNAPerBlock);
    sumsq_noalloc_kernel<<<blocks, kNAThreads, 0, s>>>(x, n, partials, ticket, out);
    B200DP_CHECK_LAUNCH();
    return 0;
}
