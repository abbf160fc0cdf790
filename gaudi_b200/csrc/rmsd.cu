// All-pairs RMSD without superposition: gaudi/similarity.py:36-75 (`rmsd`, `_rmsd_squared`),
// evaluated for a block of rows of the N x N matrix in one launch instead of once per
// Python-level pair as gaudi/algorithms.py:141-155 does.
//
// The reference's formula has no alignment step, so there is no cross-covariance matrix and
// nothing GEMM-shaped: rmsd^2 = sum_k (a_k - b_k)^2 / A.  Expanding the square into
// |a|^2 + |b|^2 - 2 a.b would make it a GEMM but cannot hold 1e-5 A in fp32 (near-identical
// conformers come out as sqrt(rounding noise)), let alone in TF32/BF16, so this is a
// register-tiled FP32 SIMT kernel on direct differences: a 128 x 128 tile of pairs per CTA,
// 8 x 8 pairs per thread, operands staged k-major in shared memory so that every inner step is
// 4 LDS.128 feeding 64 FADD + 64 FFMA.  Staging is double-buffered with cp.async (LDGSTS):
// the transposing scatter of chunk c+1 runs while chunk c is being consumed, without passing
// through registers.
#include <cuda_fp16.h>
#include <mma.h>

#include "rmsd_common.cuh"

namespace gmm {

constexpr int kRS = kRT + 4;     // padded row stride (floats), keeps 16-byte alignment

struct RmsdArgs {
    int N, K, A, row0, row1;
    const float* xyz;
    float* out; uint32_t* bits; int words;
    float thr2, inv_a;
    int sym, R0, R1;      // bitmask-only calls: tiles below the diagonal of the square part are
                          // skipped and filled in by their mirror tile (the matrix is symmetric)
};

__device__ __forceinline__ void cp_async_4(float* smem_dst, const float* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N)); }

// RK = coordinates per shared-memory chunk (36 when 3A is a multiple of 36, e.g. A = 60: no
// padding work; 32 otherwise).
template <int RK>
__global__ void __launch_bounds__(256, 2)
rmsd_tile_kernel(const RmsdArgs a) {
    extern __shared__ __align__(16) float smem_f[];
    float (*As)[RK][kRS] = reinterpret_cast<float (*)[RK][kRS]>(smem_f);                    // [2][RK][kRS]
    float (*Bs)[RK][kRS] = reinterpret_cast<float (*)[RK][kRS]>(smem_f + 2 * RK * kRS);     // [2][RK][kRS]
    const int tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;
    const int r_base = a.row0 + blockIdx.y * kRT;   // first row (conformer index) of the tile
    const int c_base = blockIdx.x * kRT;            // first column
    const int n_chunks = (a.K + RK - 1) / RK;
    const int rb = a.R0 + blockIdx.y, cbk = blockIdx.x;      // tile coordinates in units of kRT
    const bool in_square = a.sym && cbk >= a.R0 && cbk < a.R1;
    if (in_square && cbk < rb) return;                        // written by tile (cbk, rb)
    const bool mirror = in_square && cbk > rb;

    // stage chunk `c` into buffer `buf`, transposed (k-major).  A warp moves 8 conformers x 4 consecutive
    // coordinates per step: lane -> (k = 4 kg + (lane & 3), r = base + (lane >> 2 & 3) + 16 (lane >> 4)).
    // With the row stride of 132 floats the shared-memory bank of element (k, r) is (4 k + r) mod 32,
    // which this mapping makes distinct for the 32 lanes (the flat mapping e -> (e / RK, e % RK) put
    // consecutive lanes 4 banks apart: 4-way conflicts on every scatter); each conformer still
    // contributes a contiguous 16-byte piece to the global read.  Out-of-range elements are zero-filled
    // with plain stores.
    static_assert(RK % 4 == 0, "chunks are moved in groups of 4 coordinates");
    auto stage = [&](int c, int buf) {
        const int k0 = c * RK;
        const int warp_id = tid >> 5, ln = tid & 31;
        const int k_l = ln & 3, r_l = ((ln >> 2) & 3) + 16 * (ln >> 4);
        for (int u = warp_id; u < 16 * (RK / 4); u += 8) {
            const int rg = u & 15, kg = u >> 4;
            const int r = 32 * (rg >> 2) + 4 * (rg & 3) + r_l, k = 4 * kg + k_l;
            const int gk = k0 + k, ra = r_base + r, cb = c_base + r;
            if (gk < a.K && ra < a.row1) cp_async_4(&As[buf][k][r], a.xyz + (size_t)ra * a.K + gk);
            else As[buf][k][r] = 0.f;
            if (gk < a.K && cb < a.N) cp_async_4(&Bs[buf][k][r], a.xyz + (size_t)cb * a.K + gk);
            else Bs[buf][k][r] = 0.f;
        }
        cp_async_commit();
    };

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    stage(0, 0);
    for (int c = 0; c < n_chunks; ++c) {
        const int buf = c & 1;
        if (c + 1 < n_chunks) { stage(c + 1, buf ^ 1); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
#pragma unroll 4
        for (int kk = 0; kk < RK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
            const float av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float d = av[i] - bv[j];
                    acc[i][j] = fmaf(d, d, acc[i][j]);
                }
        }
        __syncthreads();
    }

    // epilogue: values straight to global; similarity bits through a 128 x 128-bit tile in shared
    // memory (the operand buffers are free now), written out coalesced and, for tiles above the
    // diagonal of a bitmask-only call, once more transposed
    uint32_t* Wn = reinterpret_cast<uint32_t*>(smem_f);       // [128 rows][4 words]
    const bool want_values = a.out != nullptr;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = r_base + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        const bool row_ok = row < a.row1;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int col0 = c_base + h * 64 + tx * 4;
            unsigned nib = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float ms = acc[i][h * 4 + j] * a.inv_a;   // similarity.py:75 (sum / N)
                // similarity.py:67-69: "if rmsd_squared > threshold*threshold: return False"
                if (row_ok && col0 + j < a.N && !(ms > a.thr2)) nib |= 1u << j;
                if (want_values && row_ok && col0 + j < a.N)
                    a.out[(size_t)(row - a.row0) * a.N + col0 + j] = sqrtf(ms);
            }
            if (a.bits != nullptr) {
                unsigned word = nib << ((tx & 7) * 4);
                word |= __shfl_xor_sync(0xffffffffu, word, 1);
                word |= __shfl_xor_sync(0xffffffffu, word, 2);
                word |= __shfl_xor_sync(0xffffffffu, word, 4);
                if ((tx & 7) == 0) Wn[(i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4)) * 4 + h * 2 + (tx >> 3)] = word;
            }
        }
    }
    if (a.bits != nullptr) {
        __syncthreads();
        for (int w = tid; w < kRT * 4; w += 256) {
            const int r = w >> 2, q = w & 3, row = r_base + r, wc = c_base / 32 + q;
            if (row < a.row1 && wc < a.words) a.bits[(size_t)(row - a.row0) * a.words + wc] = Wn[w];
        }
        if (mirror) {
            for (int w = tid; w < kRT * 4; w += 256) {
                const int cidx = w & 127, rw = w >> 7;
                unsigned t = 0;
#pragma unroll 8
                for (int r = 0; r < 32; ++r) t |= ((Wn[(32 * rw + r) * 4 + (cidx >> 5)] >> (cidx & 31)) & 1u) << r;
                const int mrow = c_base + cidx, wc = r_base / 32 + rw;
                if (mrow < a.row1 && mrow < a.N && wc < a.words) a.bits[(size_t)(mrow - a.row0) * a.words + wc] = t;
            }
        }
    }
}


// ---- similarity bitmask through the tensor cores ------------------------------------------------
// A bitmask-only call does not need the RMSD values, only on which side of the threshold each
// pair falls.  With the conformers re-centred on their column means, sum_k (a_k - b_k)^2 =
// |a|^2 + |b|^2 - 2 a.b, and a.b over all pairs of a tile is a GEMM: it runs on the tensor
// cores in FP16 with fp32 accumulation (mma.sync through the WMMA interface; the re-centred
// coordinates are a few Angstrom, well inside half range, and half carries the same 11
// significant bits as TF32 at half the shared-memory traffic and a quarter of the MMA
// instructions).  That product is only a PRE-FILTER: its error is bounded by
// 2^-10 (|a|^2 + |b|^2) (Cauchy-Schwarz on the rounded inputs, plus slack for the
// accumulation), so pairs whose approximate sum is further than that from threshold^2 * A are
// decided, and the others (about one in a hundred) are queued and re-decided by the same
// sequential fp32 direct-difference sum the FP32 kernel above computes, on the original
// coordinates.  The bits are therefore identical to that kernel's (tests compare them).
constexpr int kMK = 64;          // coordinates per staged chunk (4 k-steps of the m16n16k16 MMA)
constexpr int kMS = kMK + 8;     // padded row stride of a staged chunk (halves)
constexpr int kMD = kRT + 4;     // padded row stride of the dot-product tile (floats)

__global__ void rmsd_colsum_kernel(int N, int K, const float* __restrict__ xyz, float* __restrict__ colsum) {
    // grid.x blocks of 32 conformers each add their column sums into colsum[K] (pre-zeroed)
    const int n0 = blockIdx.x * 32;
    for (int k = threadIdx.x; k < K; k += blockDim.x) {
        float s = 0.f;
        const int n1 = min(N, n0 + 32);
#pragma unroll 8
        for (int n = n0; n < n1; ++n) s += xyz[(size_t)n * K + k];
        atomicAdd(&colsum[k], s);
    }
}

// Re-centred fp16 copies of the conformers and their squared norms.  ext = 0: one array, zero
// padded to Kp columns (mma.sync kernel).  ext = 1 (tcgen05 kernel): two arrays whose first K
// columns are the same and whose next six fold the norms and the threshold into the product itself:
//     row role     [x, n_hi, n_lo, n_ll, 1, 1, 1]      n = -|x|^2 / 2 split over three halves
//     column role  [x, 1, 1, 1, m_hi, m_lo, m_ll]      m = -|x|^2 / 2 + cut / 2
// so that  a' . b' = a . b - (|a|^2 + |b|^2 - cut) / 2 = -(approximate sum - cut) / 2  comes straight
// out of the tensor cores and the epilogue compares it with the error bound alone.
__global__ void rmsd_center_kernel(int N, int K, int Kp, const float* __restrict__ xyz, const float* __restrict__ colsum,
                                   __half* __restrict__ xc, __half* __restrict__ xcb, float half_cut, int ext,
                                   float* __restrict__ norm, int* __restrict__ overflow) {
    const int n = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (n >= N) return;
    const float inv = 1.0f / (float)N;
    float s = 0.f;
    bool big = false;
    for (int k = lane; k < Kp; k += 32) {
        float v = 0.f;
        if (k < K) v = xyz[(size_t)n * K + k] - colsum[k] * inv;
        big = big || !(fabsf(v) < 6.0e4f);
        const __half h = __float2half_rn(v);
        xc[(size_t)n * Kp + k] = h;
        if (ext) xcb[(size_t)n * Kp + k] = h;
        s = fmaf(v, v, s);                       // the norm is of the unrounded value
    }
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) norm[n] = s;
    if (ext) {
        __syncwarp();
        big = big || !(s < 1.0e5f);              // -|x|^2 / 2 has to fit a half
        if (lane < 6) {
            const float val = (lane < 3) ? -0.5f * s : -0.5f * s + half_cut;     // lanes 0-2: n, lanes 3-5: m
            const __half hi = __float2half_rn(val);
            const __half lo = __float2half_rn(val - __half2float(hi));
            const __half ll = __float2half_rn(val - __half2float(hi) - __half2float(lo));
            const int piece = lane % 3;
            const __half pc = piece == 0 ? hi : piece == 1 ? lo : ll;
            const __half one = __float2half_rn(1.0f);
            // row role: pieces of n at K .. K+2, ones at K+3 .. K+5; column role: the other way round
            xc[(size_t)n * Kp + K + lane] = lane < 3 ? pc : one;
            xcb[(size_t)n * Kp + K + lane] = lane < 3 ? one : pc;
        }
    }
    if (big) atomicExch(overflow, 1);
}

__device__ __forceinline__ void cp_async_16(void* smem_dst, const void* gmem_src) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}

__global__ void __launch_bounds__(256, 2)
rmsd_bits_mma_kernel(const RmsdMmaArgs a) {
    using namespace nvcuda;
    extern __shared__ __align__(128) float smem_m[];
    // staging: As[2][128][kMS], Bs[2][128][kMS] halves; afterwards the same memory holds the
    // dot-product tile D[128][kMD] floats; behind it: norms, bit tile, queue
    constexpr size_t kStageBytes = (size_t)4 * kRT * kMS * sizeof(__half), kDBytes = (size_t)kRT * kMD * sizeof(float);
    constexpr size_t kUnion = kStageBytes > kDBytes ? kStageBytes : kDBytes;
    __half* As = reinterpret_cast<__half*>(smem_m);
    __half* Bs = As + 2 * kRT * kMS;
    float* D = smem_m;
    float* na = smem_m + kUnion / sizeof(float);   // [128]
    float* nb = na + kRT;                          // [128]
    uint32_t* Wn = reinterpret_cast<uint32_t*>(nb + kRT);   // [128][4]
    uint32_t* queue = Wn + kRT * 4;                // [kMQ] row << 16 | col
    int* nq = reinterpret_cast<int*>(queue + kMQ);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int r_base = a.row0 + blockIdx.y * kRT, c_base = blockIdx.x * kRT;
    const int rb = a.R0 + blockIdx.y, cbk = blockIdx.x;
    const bool in_square = a.sym && cbk >= a.R0 && cbk < a.R1;
    if (in_square && cbk < rb) return;                        // written by tile (cbk, rb)
    const bool mirror = in_square && cbk > rb;
    if (tid == 0) *nq = 0;
    if (tid < kRT) {
        na[tid] = (r_base + tid < a.row1) ? a.norm[r_base + tid] : 0.f;
        nb[tid] = (c_base + tid < a.N) ? a.norm[c_base + tid] : 0.f;
    }
    for (int w = tid; w < kRT * 4; w += 256) Wn[w] = 0u;

    const int n_chunks = (a.Kp + kMK - 1) / kMK;
    // 128 rows x 8 sixteen-byte pieces per operand and chunk: thread t always moves piece t % 8 of
    // rows t / 8 + 32 i, so its source rows and validity are fixed for the whole tile
    const int sq = tid & 7, sr = tid >> 3;
    const __half* srcA[4]; const __half* srcB[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int ra = r_base + sr + 32 * i, cb = c_base + sr + 32 * i;
        srcA[i] = ra < a.row1 ? a.xc + (size_t)ra * a.Kp + 8 * sq : nullptr;
        srcB[i] = cb < a.N ? a.xc + (size_t)cb * a.Kp + 8 * sq : nullptr;
    }
    auto stage = [&](int c, int buf) {
        const int k0 = c * kMK;
        const bool k_ok = k0 + 8 * sq < a.Kp;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half* da = As + ((size_t)buf * kRT + sr + 32 * i) * kMS + 8 * sq;
            __half* db = Bs + ((size_t)buf * kRT + sr + 32 * i) * kMS + 8 * sq;
            if (k_ok && srcA[i]) cp_async_16(da, srcA[i] + k0);
            else *reinterpret_cast<uint4*>(da) = make_uint4(0u, 0u, 0u, 0u);
            if (k_ok && srcB[i]) cp_async_16(db, srcB[i] + k0);
            else *reinterpret_cast<uint4*>(db) = make_uint4(0u, 0u, 0u, 0u);
        }
        cp_async_commit();
    };

    // warp (wy, wx) of a 2 x 4 grid owns rows [64 wy, 64 wy + 64) x columns [32 wx, 32 wx + 32)
    const int wy = warp >> 2, wx = warp & 3;
    wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) wmma::fill_fragment(acc[i][j], 0.0f);

    stage(0, 0);
    for (int c = 0; c < n_chunks; ++c) {
        const int buf = c & 1;
        if (c + 1 < n_chunks) { stage(c + 1, buf ^ 1); cp_async_wait<1>(); }
        else cp_async_wait<0>();
        __syncthreads();
        const __half* Ab = As + (size_t)buf * kRT * kMS;
        const __half* Bb = Bs + (size_t)buf * kRT * kMS;
#pragma unroll
        for (int k16 = 0; k16 < kMK; k16 += 16) {
            wmma::fragment<wmma::matrix_a, 16, 16, 16, __half, wmma::row_major> fa[4];
            wmma::fragment<wmma::matrix_b, 16, 16, 16, __half, wmma::col_major> fb[2];
#pragma unroll
            for (int i = 0; i < 4; ++i) wmma::load_matrix_sync(fa[i], Ab + (size_t)(wy * 64 + i * 16) * kMS + k16, kMS);
#pragma unroll
            for (int j = 0; j < 2; ++j) wmma::load_matrix_sync(fb[j], Bb + (size_t)(wx * 32 + j * 16) * kMS + k16, kMS);
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 2; ++j) wmma::mma_sync(acc[i][j], fa[i], fb[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j)
            wmma::store_matrix_sync(D + (size_t)(wy * 64 + i * 16) * kMD + wx * 32 + j * 16, acc[i][j], kMD, wmma::mem_row_major);
    __syncthreads();
    const bool nothing_decided = *a.overflow != 0;
    const bool vec = (a.K % 4 == 0) && (reinterpret_cast<uintptr_t>(a.xyz) % 16 == 0);

    // pass 1: decided pairs -> bits, undecided pairs -> queue.  Warp w takes rows 16 w .. 16 w + 15,
    // lanes take 32 consecutive columns: one ballot is one word of the bit tile.  With
    // e = 1.2e-3 (|a|^2 + |b|^2) + c0 the bound on the error of the approximate sum
    // |a|^2 + |b|^2 - 2 D, a pair is certainly similar when D > ((1 + 1.2e-3)(|a|^2 + |b|^2) - cut + c0) / 2
    // and certainly not when D < ((1 - 1.2e-3)(|a|^2 + |b|^2) - cut - c0) / 2.
    const float c0 = 1e-4f * a.cut + 1e-4f;
    float hi_c[4], lo_c[4];
    bool col_ok[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const float nc = nb[q * 32 + lane];
        hi_c[q] = 0.5f * (1.0012f * nc - a.cut + c0);
        lo_c[q] = 0.5f * (0.9988f * nc - a.cut - c0);
        col_ok[q] = c_base + q * 32 + lane < a.N;
    }
    for (int rr = 0; rr < 16; ++rr) {
        const int r = warp * 16 + rr;
        if (r_base + r >= a.row1) break;                     // warp-uniform; the bit tile is zeroed
        const float nr = na[r];
        const float hi_r = 0.5f * 1.0012f * nr, lo_r = 0.5f * 0.9988f * nr;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float d = D[(size_t)r * kMD + q * 32 + lane];
            const bool yes = col_ok[q] && !nothing_decided && d > hi_r + hi_c[q];
            const bool maybe = col_ok[q] && !yes && (nothing_decided || !(d < lo_r + lo_c[q]));
            const unsigned word = __ballot_sync(0xffffffffu, yes);
            if (lane == 0) Wn[r * 4 + q] = word;
            if (__any_sync(0xffffffffu, maybe)) {
                __syncwarp();
                if (maybe) {
                    const int col = q * 32 + lane;
                    const int at = atomicAdd(nq, 1);
                    if (at < kMQ) queue[at] = ((uint32_t)r << 16) | (uint32_t)col;
                    else {
                        // queue full (threshold in the middle of a dense cluster): decide here and now
                        const float* x = a.xyz + (size_t)(r_base + r) * a.K;
                        const float* y = a.xyz + (size_t)(c_base + col) * a.K;
                        const float s = exact_sum(x, y, a.K, vec);
                        if (!(s * a.inv_a > a.thr2)) atomicOr(&Wn[r * 4 + q], 1u << lane);
                    }
                }
            }
        }
    }
    __syncthreads();
    // pass 2: the queue, every thread busy; the sum runs over k in order, exactly as the FP32 kernel's
    const int n_queued = min(*nq, kMQ);
    for (int e = tid; e < n_queued; e += 256) {
        const int r = queue[e] >> 16, col = queue[e] & 0xffff;
        const float* x = a.xyz + (size_t)(r_base + r) * a.K;
        const float* y = a.xyz + (size_t)(c_base + col) * a.K;
        const float s = exact_sum(x, y, a.K, vec);
        // similarity.py:67-69: "if rmsd_squared > threshold*threshold: return False"
        if (!(s * a.inv_a > a.thr2)) atomicOr(&Wn[r * 4 + (col >> 5)], 1u << (col & 31));
    }
    __syncthreads();
    for (int w = tid; w < kRT * 4; w += 256) {
        const int r = w >> 2, q = w & 3, row = r_base + r, wc = c_base / 32 + q;
        if (row < a.row1 && wc < a.words) a.bits[(size_t)(row - a.row0) * a.words + wc] = Wn[w];
    }
    if (mirror) {
        // the tile transposed: sixteen 32 x 32 bit blocks, each turned by one warp with a five-step
        // butterfly of shuffles (lane = row of the block in, column of the block out)
        for (int blk = warp; blk < 16; blk += 8) {
            const int rw = blk >> 2, q = blk & 3;
            unsigned x = Wn[(32 * rw + lane) * 4 + q];
#pragma unroll
            for (int step = 0; step < 5; ++step) {
                const int j = 16 >> step;
                const unsigned m = step == 0 ? 0x0000ffffu : step == 1 ? 0x00ff00ffu : step == 2 ? 0x0f0f0f0fu
                                 : step == 3 ? 0x33333333u : 0x55555555u;
                const unsigned y = __shfl_xor_sync(0xffffffffu, x, j);
                x = (lane & j) ? ((x & ~m) | ((y >> j) & m)) : ((x & m) | ((y & m) << j));
            }
            const int mrow = c_base + 32 * q + lane, wc = r_base / 32 + rw;
            if (mrow < a.row1 && mrow < a.N && wc < a.words) a.bits[(size_t)(mrow - a.row0) * a.words + wc] = x;
        }
    }
}

// Workspace of the bitmask path: re-centred fp16 conformers, their squared norms, column sums, flag.
// Grown on demand (device-wide synchronisation + cudaMalloc); gmm_reserve_rmsd() pre-sizes it.
// Kt: columns of the staged copies = 3 A coordinates + the six folded columns of the tcgen05 path,
// rounded up to a multiple of 16
static int staged_columns(int A) { return (3 * A + 6 + 15) & ~15; }

int reserve_rmsd_workspace(gmm_ctx* c, int N, int A) {
    const int Kt = staged_columns(A);
    const size_t xc_bytes = ((size_t)N * Kt * sizeof(__half) + 255) / 256 * 256;
    const size_t need = 2 * xc_bytes + ((size_t)N + Kt + 64) * sizeof(float);
    if (need > c->rmsd_ws_bytes) {
        if (c->d_rmsd_ws) { GMM_CUDA(cudaDeviceSynchronize()); cudaFree(c->d_rmsd_ws); c->d_rmsd_ws = nullptr; c->rmsd_ws_bytes = 0; }
        GMM_CUDA(cudaMalloc(&c->d_rmsd_ws, need));
        c->rmsd_ws_bytes = need;
    }
    return GMM_OK;
}

static int launch_rmsd_mma(gmm_ctx* c, int N, int A, const float* xyz, int row0, int row1, uint32_t* out_bits,
                           double thr, cudaStream_t s) {
    const bool tc = !(c->prm.options & GMM_OPT_RMSD_HMMA);
    const int K = 3 * A, Kt = staged_columns(A), Kp = tc ? Kt : ((K + 15) & ~15);
    const size_t xc_bytes = ((size_t)N * Kt * sizeof(__half) + 255) / 256 * 256;
    { const int rc = reserve_rmsd_workspace(c, N, A); if (rc) return rc; }
    __half* xc = static_cast<__half*>(c->d_rmsd_ws);
    __half* xcb = reinterpret_cast<__half*>(static_cast<unsigned char*>(c->d_rmsd_ws) + xc_bytes);
    float* norm = reinterpret_cast<float*>(static_cast<unsigned char*>(c->d_rmsd_ws) + 2 * xc_bytes);
    float* colsum = norm + N;
    int* overflow = reinterpret_cast<int*>(colsum + Kt);
    const float cut_f = (float)(thr * thr) * (float)A;
    GMM_CUDA(cudaMemsetAsync(colsum, 0, sizeof(float) * (Kt + 1), s));       // column sums and the flag
    rmsd_colsum_kernel<<<(N + 31) / 32, 192, 0, s>>>(N, K, xyz, colsum);
    GMM_LAUNCH_CHECK("rmsd_colsum_kernel");
    rmsd_center_kernel<<<(int)(((size_t)N * 32 + 255) / 256), 256, 0, s>>>(N, K, Kp, xyz, colsum, xc, xcb, 0.5f * cut_f,
                                                                           tc ? 1 : 0, norm, overflow);
    GMM_LAUNCH_CHECK("rmsd_center_kernel");
    RmsdMmaArgs a;
    a.N = N; a.K = K; a.Kp = Kp; a.row0 = row0; a.row1 = row1; a.words = (N + 31) / 32;
    a.xyz = xyz; a.xc = xc; a.xcb = xcb; a.norm = norm; a.overflow = overflow; a.bits = out_bits;
    a.thr2 = (float)(thr * thr); a.inv_a = 1.0f / (float)A; a.cut = a.thr2 * (float)A;
    a.sym = (row0 % kRT == 0) ? 1 : 0;
    a.R0 = row0 / kRT; a.R1 = (row1 + kRT - 1) / kRT;
    if (tc) {                                          // the tcgen05 / TMEM / TMA kernel (rmsd_tc.cu)
        int handled = 0;
        const int rc = launch_rmsd_tc(c, a, s, &handled);
        if (rc != GMM_OK || handled) return rc;
        // no tensor map on this driver: stage the plain copy and run the mma.sync kernel
        a.Kp = (K + 15) & ~15;
        rmsd_center_kernel<<<(int)(((size_t)N * 32 + 255) / 256), 256, 0, s>>>(N, K, a.Kp, xyz, colsum, xc, xcb, 0.f, 0, norm,
                                                                               overflow);
        GMM_LAUNCH_CHECK("rmsd_center_kernel");
    }
    dim3 grid((N + kRT - 1) / kRT, (row1 - row0 + kRT - 1) / kRT);
    const size_t staging = (size_t)4 * kRT * kMS * sizeof(__half), dtile = (size_t)kRT * kMD * sizeof(float);
    const size_t smem = (staging > dtile ? staging : dtile) + (2 * kRT + kRT * 4 + kMQ + 4) * sizeof(float);
    { const int rc = allow_dynamic_smem((const void*)rmsd_bits_mma_kernel, c->device, smem); if (rc) return rc; }
    rmsd_bits_mma_kernel<<<grid, 256, smem, s>>>(a);
    GMM_LAUNCH_CHECK("rmsd_bits_mma_kernel");
    return GMM_OK;
}

template <int RK>
static int launch_rmsd_rk(const RmsdArgs& a, dim3 grid, cudaStream_t s, int device) {
    const size_t smem = (size_t)4 * RK * kRS * sizeof(float);
    { const int rc = allow_dynamic_smem((const void*)rmsd_tile_kernel<RK>, device, smem); if (rc) return rc; }
    rmsd_tile_kernel<RK><<<grid, 256, smem, s>>>(a);
    GMM_LAUNCH_CHECK("rmsd_tile_kernel");
    return GMM_OK;
}

int launch_rmsd(gmm_ctx* c, int N, int A, const float* xyz, int row0, int row1, float* out_rmsd,
                uint32_t* out_bits, double thr, cudaStream_t s) {
    if (row1 <= row0 || N == 0) return GMM_OK;
    // bitmask only: tensor-core pre-filter + exact re-decision (identical bits, several times faster)
    if (out_rmsd == nullptr && out_bits != nullptr && !(c->prm.options & GMM_OPT_RMSD_SIMT) && A >= 8)
        return launch_rmsd_mma(c, N, A, xyz, row0, row1, out_bits, thr, s);
    RmsdArgs a;
    a.N = N; a.A = A; a.K = 3 * A; a.row0 = row0; a.row1 = row1;
    a.xyz = xyz; a.out = out_rmsd; a.bits = out_bits; a.words = (N + 31) / 32;
    a.thr2 = (float)(thr * thr); a.inv_a = 1.0f / (float)A;
    a.sym = (out_rmsd == nullptr && out_bits != nullptr && row0 % kRT == 0) ? 1 : 0;
    a.R0 = row0 / kRT; a.R1 = (row1 + kRT - 1) / kRT;
    dim3 grid((N + kRT - 1) / kRT, (row1 - row0 + kRT - 1) / kRT);
    if (a.K % 36 == 0) return launch_rmsd_rk<36>(a, grid, s, c->device);
    return launch_rmsd_rk<32>(a, grid, s, c->device);
}

}  // namespace gmm
