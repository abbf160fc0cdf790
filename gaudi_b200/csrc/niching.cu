// Similarity bookkeeping of the GA loop ("next" row f.1: GPU-assisted niching).
//
// gaudi/algorithms.py:139-155 walks the candidates in order and asks ind.similar(other) for
// every kept individual: O(N^2) Python calls.  Here gmm_rmsd_matrix has already produced the
// N x N similarity bitmask on the device (2 GiB at N = 131,072, far too much to ship to the host
// every generation).  Similar pairs are rare, so the bitmask is compacted on the GPU into one
// partner list per candidate (CSR, self excluded, partners ascending) and only those lists go to
// the host, where the inherently sequential greedy pass runs over them in a few milliseconds.
#include "common.cuh"

namespace gmm {

// one warp per row: number of set bits, not counting the row's own column
__global__ void sim_count_kernel(int rows, int row0, int words, const uint32_t* __restrict__ bits,
                                 int32_t* __restrict__ counts) {
    const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= rows) return;
    const uint32_t* row = bits + (size_t)r * words;
    const int self = row0 + r;
    int n = 0;
    for (int w = lane; w < words; w += 32) {
        uint32_t v = row[w];
        if (w == (self >> 5)) v &= ~(1u << (self & 31));
        n += __popc(v);
    }
    for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
    if (lane == 0) counts[r] = n;
}

// one warp per row: partner indices in ascending order at idx[ptr[r] ...]
__global__ void sim_fill_kernel(int rows, int row0, int words, const uint32_t* __restrict__ bits,
                                const int64_t* __restrict__ ptr, int32_t* __restrict__ idx, int64_t cap) {
    const int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (r >= rows) return;
    const uint32_t* row = bits + (size_t)r * words;
    const int self = row0 + r;
    int64_t base = ptr[r];
    for (int w0 = 0; w0 < words; w0 += 32) {
        const int w = w0 + lane;
        uint32_t v = w < words ? row[w] : 0u;
        if (w == (self >> 5)) v &= ~(1u << (self & 31));
        const int c = __popc(v);
        int incl = c;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= o) incl += t;
        }
        int64_t at = base + incl - c;
        while (v) {
            if (at < cap) idx[at] = (w << 5) + (__ffs(v) - 1);     // entries past the caller's capacity are dropped
            ++at;
            v &= v - 1;
        }
        base += __shfl_sync(0xffffffffu, incl, 31);
    }
}

}  // namespace gmm

extern "C" int gmm_similar_lists(gmm_ctx* c, int N, int row0, int row1, const uint32_t* bits_d, int32_t* counts_d,
                                 const int64_t* ptr_d, int32_t* idx_d, long long idx_capacity, void* stream) {
    if (c == nullptr) { gmm::set_error("gmm_similar_lists: ctx is NULL"); return GMM_EINVAL; }
    if (N < 0 || row0 < 0 || row0 > row1 || row1 > N) {
        gmm::set_error("gmm_similar_lists: rows [%d,%d) outside [0,%d]", row0, row1, N);
        return GMM_EINVAL;
    }
    const int rows = row1 - row0;
    if (rows == 0) return GMM_OK;
    if (bits_d == nullptr || (counts_d == nullptr && (ptr_d == nullptr || idx_d == nullptr))) {
        gmm::set_error("gmm_similar_lists: need bits_d and either counts_d (count pass) or ptr_d + idx_d (fill pass)");
        return GMM_EINVAL;
    }
    gmm::DeviceGuard g(c->device);
    const int words = (N + 31) / 32, threads = 256;
    const int blocks = (int)(((size_t)rows * 32 + threads - 1) / threads);
    cudaStream_t s = (cudaStream_t)stream;
    if (counts_d) {
        gmm::sim_count_kernel<<<blocks, threads, 0, s>>>(rows, row0, words, bits_d, counts_d);
        GMM_LAUNCH_CHECK("sim_count_kernel");
    }
    if (ptr_d && idx_d) {
        gmm::sim_fill_kernel<<<blocks, threads, 0, s>>>(rows, row0, words, bits_d, ptr_d, idx_d,
                                                        idx_capacity > 0 ? (int64_t)idx_capacity : INT64_MAX);
        GMM_LAUNCH_CHECK("sim_fill_kernel");
    }
    return GMM_OK;
}

// Host: the greedy pass of gaudi/algorithms.py:139-155 over the partner lists.  A candidate is
// dropped when a kept partner is at least as fit (rank >=); otherwise every kept partner is
// evicted and the candidate is kept.
extern "C" int gmm_niching_host(int N, const int64_t* ptr_h, const int32_t* idx_h, const int64_t* rank_h,
                                uint8_t* kept_h) {
    if (N < 0 || (N > 0 && (!ptr_h || !rank_h || !kept_h))) {
        gmm::set_error("gmm_niching_host: bad argument");
        return GMM_EINVAL;
    }
    for (int i = 0; i < N; ++i) kept_h[i] = 0;
    for (int i = 0; i < N; ++i) {
        bool drop = false;
        for (int64_t k = ptr_h[i]; k < ptr_h[i + 1]; ++k) {
            const int j = idx_h[k];
            if (j < 0 || j >= N) { gmm::set_error("gmm_niching_host: partner %d outside [0,%d)", j, N); return GMM_EINVAL; }
            if (kept_h[j] && rank_h[j] >= rank_h[i]) { drop = true; break; }
        }
        if (drop) continue;
        for (int64_t k = ptr_h[i]; k < ptr_h[i + 1]; ++k) kept_h[idx_h[k]] = 0;
        kept_h[i] = 1;
    }
    return GMM_OK;
}

// Host: fronts of TWO-objective rows sorted best first (column 0 descending, ties column 1
// descending): every dominator of a row stands before it, front f remembers the largest second
// objective it holds (non-increasing in f), and a row joins the first front whose best is strictly
// below its own second objective (binary search): O(N log N), the host-side counterpart of
// gmm_nondominated_fronts for the common two-objective job.  Equal rows share a front.
extern "C" int gmm_fronts2_host(int N, const double* w_sorted_h, int32_t* front_h) {
    if (N < 0 || (N > 0 && (!w_sorted_h || !front_h))) { gmm::set_error("gmm_fronts2_host: bad argument"); return GMM_EINVAL; }
    std::vector<double> best;            // best[f], non-increasing
    for (int i = 0; i < N; ++i) {
        const double x = w_sorted_h[2 * (size_t)i], y = w_sorted_h[2 * (size_t)i + 1];
        if (i > 0) {
            const double px = w_sorted_h[2 * (size_t)i - 2], py = w_sorted_h[2 * (size_t)i - 1];
            if (px < x || (px == x && py < y) || x != x || y != y) {
                gmm::set_error("gmm_fronts2_host: rows are not sorted best first (or hold NaN)");
                return GMM_EINVAL;
            }
            if (px == x && py == y) { front_h[i] = front_h[i - 1]; continue; }
        }
        // first f with best[f] < y
        size_t lo = 0, hi = best.size();
        while (lo < hi) {
            const size_t mid = (lo + hi) / 2;
            if (best[mid] < y) hi = mid; else lo = mid + 1;
        }
        if (lo == best.size()) best.push_back(y); else best[lo] = y;
        front_h[i] = (int32_t)lo;
    }
    return GMM_OK;
}
