// Layout moves around the attention core of the transformer models (ViT-B/16, SwinIR): the library
// attention kernel hands back per-head tensors [B][H][S][d]; the projections before / after it work
// on token-major matrices [B*S][W*H*d].  One streaming pass each way (16-byte vectors, every 2*d-byte
// head row read and written as one contiguous segment) instead of ATen's stack / strided-copy chains.
#include "common.cuh"

namespace b200dp {

struct HeadSrc { const void* p[3]; };
struct HeadDst { void* p[3]; };

// out[b][s][w][h][:] = src_w[b][h][s][:]      (W sources, each logically [B][H][S][d] with element strides
// (sb, sh, ss) and a contiguous last dimension — contiguous per-head tensors and token-major views alike)
__global__ void __launch_bounds__(256) heads_to_tokens_kernel(HeadSrc src, uint4* __restrict__ out, int B, int H, int S,
                                                              int W, int vpr /* 16-byte vectors per head row */,
                                                              long long sb, long long sh, long long ss /* in vectors */) {
    const long long total = (long long)B * S * W * H * vpr;
    for (long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += (long long)gridDim.x * blockDim.x) {
        const int v = (int)(o % vpr);
        long long t = o / vpr;
        const int h = (int)(t % H); t /= H;
        const int w = (int)(t % W); t /= W;
        const int s = (int)(t % S);
        const int b = (int)(t / S);
        const uint4* sp = reinterpret_cast<const uint4*>(src.p[w]);
        out[o] = __ldg(sp + b * sb + h * sh + s * ss + v);
    }
}

// dst_w[b][h][s][:] = in[b][s][w][h][:]
__global__ void __launch_bounds__(256) tokens_to_heads_kernel(const uint4* __restrict__ in, HeadDst dst, int B, int H, int S,
                                                              int W, int vpr) {
    const long long per = (long long)B * H * S * vpr;
    const long long total = per * W;
    for (long long o = (long long)blockIdx.x * blockDim.x + threadIdx.x; o < total; o += (long long)gridDim.x * blockDim.x) {
        const int w = (int)(o / per);
        long long t = o % per;
        const int v = (int)(t % vpr); t /= vpr;
        const int s = (int)(t % S); t /= S;
        const int h = (int)(t % H);
        const int b = (int)(t / H);
        uint4* dp = reinterpret_cast<uint4*>(dst.p[w]);
        dp[o % per] = __ldg(in + ((((long long)b * S + s) * W + w) * H + h) * vpr + v);
    }
}

}  // namespace b200dp

using namespace b200dp;

extern "C" int b200dp_heads_to_tokens(const void* const* srcs, void* out, int B, int H, int S, int d, int W, int elem_bytes,
                                      long long stride_b, long long stride_h, long long stride_s, void* stream) {
    B200DP_REQUIRE(srcs && out && W >= 1 && W <= 3, "heads_to_tokens: bad arguments");
    B200DP_REQUIRE((d * elem_bytes) % 16 == 0, "heads_to_tokens: head rows must be multiples of 16 bytes");
    HeadSrc s;
    for (int i = 0; i < 3; ++i) s.p[i] = i < W ? srcs[i] : nullptr;
    const int vpr = d * elem_bytes / 16;
    const long long total = (long long)B * S * W * H * vpr;
    if (total == 0) return 0;
    long long blocks = (total + 255) / 256;
    if (blocks > B200DP_SMS * 16) blocks = B200DP_SMS * 16;
    const long long epv = 16 / elem_bytes;      // elements per vector
    B200DP_REQUIRE(stride_b % epv == 0 && stride_h % epv == 0 && stride_s % epv == 0,
                   "heads_to_tokens: source strides must be multiples of 16 bytes");
    heads_to_tokens_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(s, (uint4*)out, B, H, S, W, vpr, stride_b / epv,
                                                                               stride_h / epv, stride_s / epv);
    B200DP_CHECK_LAUNCH();
    return 0;
}

extern "C" int b200dp_tokens_to_heads(const void* in, void* const* dsts, int B, int H, int S, int d, int W, int elem_bytes,
                                      void* stream) {
    B200DP_REQUIRE(in && dsts && W >= 1 && W <= 3, "tokens_to_heads: bad arguments");
    B200DP_REQUIRE((d * elem_bytes) % 16 == 0, "tokens_to_heads: head rows must be multiples of 16 bytes");
    HeadDst t;
    for (int i = 0; i < 3; ++i) t.p[i] = i < W ? dsts[i] : nullptr;
    const int vpr = d * elem_bytes / 16;
    const long long total = (long long)B * S * W * H * vpr;
    if (total == 0) return 0;
    long long blocks = (total + 255) / 256;
    if (blocks > B200DP_SMS * 16) blocks = B200DP_SMS * 16;
    tokens_to_heads_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>((const uint4*)in, t, B, H, S, W, vpr);
    B200DP_CHECK_LAUNCH();
    return 0;
}
