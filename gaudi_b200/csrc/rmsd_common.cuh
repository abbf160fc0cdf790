// Pieces shared by the similarity-bitmask kernels (rmsd.cu: FP32 and mma.sync paths; rmsd_tc.cu:
// the tcgen05 / TMEM / TMA path).
#pragma once
#include <cuda_fp16.h>

#include "common.cuh"

namespace gmm {

constexpr int kRT = 128;         // tile edge (pairs)
constexpr int kMQ = 4096;        // queue of pairs to re-decide, per tile (mma.sync path)

struct RmsdMmaArgs {
    int N, K, Kp, row0, row1, words;
    const float* xyz;            // original coordinates [N][K]
    const __half* xc;            // re-centred, rounded to half, zero-padded [N][Kp] (tcgen05 path: row role)
    const __half* xcb;           // tcgen05 path: the column-role copy (rmsd_center_kernel)
    const int* overflow;         // != 0: some re-centred coordinate does not fit half -> nothing is pre-decided
    const float* norm;           // [N] |xc|^2
    uint32_t* bits;
    float thr2, inv_a, cut;      // cut = thr2 * A: the threshold on the un-normalised sum
    int sym, R0, R1;
};

// sum_k (x_k - y_k)^2 accumulated in k order with one FFMA per term: the very sequence of roundings
// the FP32 kernel's accumulators go through, so the decision taken from it is that kernel's.
__device__ __forceinline__ float exact_sum(const float* __restrict__ x, const float* __restrict__ y, int K, bool vec) {
    float s = 0.f;
    if (vec) {      // K a multiple of 4 and both rows 16-byte aligned: same order, a quarter of the loads
        const float4* x4 = reinterpret_cast<const float4*>(x);
        const float4* y4 = reinterpret_cast<const float4*>(y);
        // The call is bound by the latency of its loads (one thread, two rows in L2), so they are
        // issued fifteen float4 pairs at a time before any is consumed; the additions still run in
        // k order.
        constexpr int kBatch = 15;
        int k = 0;
        for (; k + kBatch <= K / 4; k += kBatch) {
            float4 p[kBatch], q[kBatch];
#pragma unroll
            for (int u = 0; u < kBatch; ++u) { p[u] = x4[k + u]; q[u] = y4[k + u]; }
#pragma unroll
            for (int u = 0; u < kBatch; ++u) {
                float d = p[u].x - q[u].x; s = fmaf(d, d, s);
                d = p[u].y - q[u].y; s = fmaf(d, d, s);
                d = p[u].z - q[u].z; s = fmaf(d, d, s);
                d = p[u].w - q[u].w; s = fmaf(d, d, s);
            }
        }
        for (; k < K / 4; ++k) {
            const float4 p = x4[k], q = y4[k];
            float d = p.x - q.x; s = fmaf(d, d, s);
            d = p.y - q.y; s = fmaf(d, d, s);
            d = p.z - q.z; s = fmaf(d, d, s);
            d = p.w - q.w; s = fmaf(d, d, s);
        }
    } else {
        for (int k = 0; k < K; ++k) { const float d = x[k] - y[k]; s = fmaf(d, d, s); }
    }
    return s;
}

// rmsd_tc.cu
int launch_rmsd_tc(gmm_ctx* c, const RmsdMmaArgs& a, cudaStream_t s, int* handled);

}  // namespace gmm
