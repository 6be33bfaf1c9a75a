// microbench.cu -- integer issue-rate probe used as the roofline denominator of the DP kernel
// ("int_lanes_per_SM must be measured on the box").
// Each thread runs `iters` rounds of 8 independent dependency chains of one instruction kind, so the
// measured rate is the pipe's issue limit, not its latency.
#include "lara_gpu.cuh"

namespace lara
{

namespace
{

template <int KIND>
__global__ void int_issue_kernel(int iters, int32_t * out)
{
    int32_t a[8];
    int32_t const s = (int32_t)threadIdx.x, b = (int32_t)blockIdx.x + 3;
    #pragma unroll
    for (int i = 0; i < 8; ++i) a[i] = s * (i + 1) + b;
    uint32_t tw = 0;
    for (int it = 0; it < iters; ++it)
    {
        #pragma unroll
        for (int u = 0; u < 4; ++u)
        {
            #pragma unroll
            for (int i = 0; i < 8; ++i)
            {
                if (KIND == 0) a[i] = a[i] + a[(i + 3) & 7];               // IADD3 (two live operands: cannot be folded)
                else if (KIND == 1) a[i] = max(a[i] ^ it, b);               // LOP3 + VIMNMX
                else
                {   // the DP's decision step: difference, max, funnel shift
                    int32_t const x = a[i] + b, y = a[(i + 1) & 7] + it;
                    tw = __funnelshift_l((uint32_t)(x - y), tw, 1);
                    a[i] = max(x, y);
                }
            }
        }
    }
    int32_t r = (int32_t)tw;
    #pragma unroll
    for (int i = 0; i < 8; ++i) r ^= a[i];
    if (r == 0x7fffffff) out[0] = r;                                        // keep the chains alive
}

} // namespace

cudaError_t launch_int_issue(int kind, int blocks, int threads, int iters, int32_t * out, cudaStream_t stream)
{
    switch (kind)
    {
        case 0: int_issue_kernel<0><<<blocks, threads, 0, stream>>>(iters, out); break;
        case 1: int_issue_kernel<1><<<blocks, threads, 0, stream>>>(iters, out); break;
        default: int_issue_kernel<2><<<blocks, threads, 0, stream>>>(iters, out); break;
    }
    return cudaGetLastError();
}

} // namespace lara
