/* Minimal C host for the kernel library: links libb200dp.so through its C ABI (include/b200dp.h), queries the
 * version and the scratch sizes a caller has to provide, and — when a GPU is present — runs one fused
 * sum-of-squares reduction.  Build:
 *   gcc -I include examples/c_host/main.c -L pytorch_distributedtraining_b200/ops -lb200dp \
 *       -L /usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/pytorch_distributedtraining_b200/ops -o /tmp/b200dp_host
 * (tests/test_misc_cpu.py::test_c_host_links_against_the_kernel_library does exactly this). */
#include <stdio.h>
#include <stdint.h>
#include "b200dp.h"

int main(void) {
    printf("b200dp version %d\n", b200dp_version());
    /* host-side planning calls: no device needed */
    printf("bn scratch floats (rows=802816, C=256): %lld\n", (long long)b200dp_bn_scratch_floats(802816, 256));
    printf("ln scratch floats (rows=25216, D=768): %lld\n", (long long)b200dp_ln_scratch_floats(25216, 768));
    printf("fused epilogue envelope 25216 x 3072: %d, 128 x 128: %d\n", b200dp_gemm_fused_supported(25216, 3072),
           b200dp_gemm_fused_supported(128, 128));
    /* an invalid call must fail with a message, not crash */
    int rc = b200dp_gemm(0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, B200DP_BF16, B200DP_BF16, 0, 0);
    printf("invalid gemm call -> rc %d (%s)\n", rc, b200dp_last_error());
    return rc == 0 ? 1 : 0;
}
