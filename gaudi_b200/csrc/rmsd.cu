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
#include "common.cuh"

namespace gmm {

constexpr int kRT = 128;         // tile edge (pairs)
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

    // stage chunk `c` into buffer `buf`: flat element e -> (conformer r = e / RK, coordinate k = e % RK);
    // consecutive threads read consecutive coordinates of a conformer (coalesced), and scatter
    // them transposed; out-of-range elements are zero-filled with plain stores.
    auto stage = [&](int c, int buf) {
        const int k0 = c * RK;
        for (int e = tid; e < RK * kRT; e += 256) {
            const int r = e / RK, k = e - r * RK;
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
