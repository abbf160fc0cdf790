// Non-dominated sorting of a population's fitness rows on the device ("next" row f.1: the
// selection step of gaudi/algorithms.py:160, DEAP's selNSGA2 -> sortNondominated).
//
// DEAP peels fronts one at a time with O(M N^2) Python-level comparisons per call; the array GA's
// numpy restatement of that takes 11 s at N = 4,000 and 277 s at N = 16,000 with three objectives.
// Here the rows arrive sorted lexicographically, best first, so every row that dominates row i
// stands before it, and
//     front(i) = 0 if nothing dominates i, else 1 + max { front(j) : j dominates i }
// (the length of the longest chain of dominators, which is what peeling produces).  Rows are
// taken in blocks of kNdsBlock: one kernel finishes the block against itself (a wavefront over
// the block, one row per step), a second pushes the finished block's fronts to every later row
// (all pairs in parallel, the N^2 / 2 part).  Equal rows dominate neither each other nor anything
// differently, so they come out on the same front, as DEAP's grouping by fitness value does.
#include "common.cuh"

namespace gmm {

constexpr int kNdsBlock = 512;       // rows finished per step (= threads of the block kernel)
constexpr int kNdsMaxObj = 8;

// j dominates i (maximisation): no component worse, at least one better
template <int M>
__device__ __forceinline__ bool dominates(const double* __restrict__ wj, const double* __restrict__ wi) {
    bool ge = true, gt = false;
#pragma unroll
    for (int c = 0; c < M; ++c) {
        ge = ge && (wj[c] >= wi[c]);
        gt = gt || (wj[c] > wi[c]);
    }
    return ge && gt;
}

// Finishes rows [b0, b0 + nb): `front` already holds, for each of them, 1 + the best front of its
// dominators among the rows before b0 (0 if none).  One thread per row.  Thread t first marks the
// later rows of the block it dominates (a bit each, in registers); then the rows are finalised in
// order, one per step: every thread whose bit for row s is set offers its own, already final,
// front + 1, the offers are max-reduced per warp and merged with one shared-memory atomic, one
// barrier per step.  (Having row s walk the list of its dominators itself put a serial chain of a
// hundred dependent shared-memory reads into every step: 1.6 ms per block instead of 0.05.)
template <int M>
__global__ void __launch_bounds__(kNdsBlock)
nds_block_kernel(int b0, int nb, const double* __restrict__ w, int32_t* __restrict__ front) {
    __shared__ double sw[kNdsBlock * M];
    __shared__ int32_t sf[kNdsBlock];
    const int t = threadIdx.x, lane = t & 31;
    if (t < nb) {
#pragma unroll
        for (int c = 0; c < M; ++c) sw[t * M + c] = w[(size_t)(b0 + t) * M + c];
        sf[t] = front[b0 + t];
    }
    __syncthreads();
    double mine[M];
#pragma unroll
    for (int c = 0; c < M; ++c) mine[c] = t < nb ? sw[t * M + c] : 0.0;
    uint32_t mask[kNdsBlock / 32];
#pragma unroll
    for (int q = 0; q < kNdsBlock / 32; ++q) {
        uint32_t word = 0u;
        if (t < nb && (q << 5) + 31 > t) {
#pragma unroll 1
            for (int bit = 0; bit < 32; ++bit) {
                const int i = (q << 5) + bit;
                if (i > t && i < nb && dominates<M>(mine, &sw[i * M])) word |= 1u << bit;
            }
        }
        mask[q] = word;
    }
    int my_front = 0;
#pragma unroll
    for (int q = 0; q < kNdsBlock / 32; ++q) {
        if ((q << 5) >= nb) break;
#pragma unroll 1
        for (int bit = 0; bit < 32; ++bit) {
            const int s = (q << 5) + bit;
            if (s >= nb) break;
            int offer = ((mask[q] >> bit) & 1u) ? my_front + 1 : 0;
            offer = __reduce_max_sync(0xffffffffu, offer);
            if (lane == 0 && offer > 0) atomicMax(&sf[s], offer);
            __syncthreads();
            if (t == s) my_front = sf[s];
        }
    }
    if (t < nb) front[b0 + t] = my_front;
}

// Pushes the finished rows [b0, b0 + nb) to every later row: front(i) = max(front(i), 1 + front(j))
// for the j of the block that dominate i.  One thread per later row, the block's rows in shared memory.
template <int M>
__global__ void __launch_bounds__(256)
nds_push_kernel(int b0, int nb, int n, const double* __restrict__ w, int32_t* __restrict__ front) {
    __shared__ double sw[kNdsBlock * M];
    __shared__ int32_t sf[kNdsBlock];
    for (int e = threadIdx.x; e < nb; e += blockDim.x) {
#pragma unroll
        for (int c = 0; c < M; ++c) sw[e * M + c] = w[(size_t)(b0 + e) * M + c];
        sf[e] = front[b0 + e];
    }
    __syncthreads();
    const int i = b0 + nb + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double mine[M];
#pragma unroll
    for (int c = 0; c < M; ++c) mine[c] = w[(size_t)i * M + c];
    int best = front[i];
    for (int j = 0; j < nb; ++j)
        if (dominates<M>(&sw[j * M], mine)) best = max(best, sf[j] + 1);
    front[i] = best;
}

// 1 in *bad when two consecutive rows are out of (descending lexicographic) order
__global__ void nds_check_sorted_kernel(int n, int m, const double* __restrict__ w, int* __restrict__ bad) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i + 1 >= n) return;
    for (int c = 0; c < m; ++c) {
        const double a = w[(size_t)i * m + c], b = w[(size_t)(i + 1) * m + c];
        if (a > b) return;                 // strictly better here: in order
        if (a < b || a != a || b != b) { atomicExch(bad, 1); return; }
    }
}

template <int M>
static int nds_run(int n, const double* w, int32_t* front, cudaStream_t s) {
    for (int b0 = 0; b0 < n; b0 += kNdsBlock) {
        const int nb = n - b0 < kNdsBlock ? n - b0 : kNdsBlock;
        nds_block_kernel<M><<<1, kNdsBlock, 0, s>>>(b0, nb, w, front);
        GMM_LAUNCH_CHECK("nds_block_kernel");
        const int later = n - (b0 + nb);
        if (later > 0) {
            nds_push_kernel<M><<<(later + 255) / 256, 256, 0, s>>>(b0, nb, n, w, front);
            GMM_LAUNCH_CHECK("nds_push_kernel");
        }
    }
    return GMM_OK;
}

}  // namespace gmm

extern "C" int gmm_nondominated_fronts(gmm_ctx* c, int n, int m, const double* w_sorted_d, int32_t* front_d,
                                       void* stream) {
    if (c == nullptr) { gmm::set_error("gmm_nondominated_fronts: ctx is NULL"); return GMM_EINVAL; }
    if (n < 0 || m < 1 || m > gmm::kNdsMaxObj) {
        gmm::set_error("gmm_nondominated_fronts: n=%d, m=%d outside n >= 0, 1 <= m <= %d", n, m, gmm::kNdsMaxObj);
        return GMM_EINVAL;
    }
    if (n == 0) return GMM_OK;
    if (w_sorted_d == nullptr || front_d == nullptr) { gmm::set_error("gmm_nondominated_fronts: NULL buffer"); return GMM_EINVAL; }
    gmm::DeviceGuard g(c->device);
    cudaStream_t s = (cudaStream_t)stream;
    // the recurrence is only valid on rows sorted best first: check it (one flag comes back)
    int* bad = reinterpret_cast<int*>(c->d_counters + 2);
    GMM_CUDA(cudaMemsetAsync(bad, 0, sizeof(int), s));
    gmm::nds_check_sorted_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, m, w_sorted_d, bad);
    GMM_LAUNCH_CHECK("nds_check_sorted_kernel");
    int bad_h = 0;
    GMM_CUDA(cudaMemcpyAsync(&bad_h, bad, sizeof(int), cudaMemcpyDeviceToHost, s));
    GMM_CUDA(cudaStreamSynchronize(s));
    if (bad_h) {
        gmm::set_error("gmm_nondominated_fronts: rows are not sorted lexicographically, best first (or hold NaN)");
        return GMM_EINVAL;
    }
    GMM_CUDA(cudaMemsetAsync(front_d, 0, sizeof(int32_t) * (size_t)n, s));
    switch (m) {
        case 1: return gmm::nds_run<1>(n, w_sorted_d, front_d, s);
        case 2: return gmm::nds_run<2>(n, w_sorted_d, front_d, s);
        case 3: return gmm::nds_run<3>(n, w_sorted_d, front_d, s);
        case 4: return gmm::nds_run<4>(n, w_sorted_d, front_d, s);
        case 5: return gmm::nds_run<5>(n, w_sorted_d, front_d, s);
        case 6: return gmm::nds_run<6>(n, w_sorted_d, front_d, s);
        case 7: return gmm::nds_run<7>(n, w_sorted_d, front_d, s);
        default: return gmm::nds_run<8>(n, w_sorted_d, front_d, s);
    }
}
