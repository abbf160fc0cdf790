// All-pairs RMSD without superposition: gaudi/similarity.py:36-75 (`rmsd`, `_rmsd_squared`),
// evaluated for a block of rows of the N x N matrix in one launch instead of once per
// Python-level pair as gaudi/algorithms.py:141-155 does.
//
// The reference's formula has no alignment step, so there is no cross-covariance matrix and
// nothing GEMM-shaped: rmsd^2 = sum_k (a_k - b_k)^2 / A.  Expanding the square into
// |a|^2 + |b|^2 - 2 a.b would make it a GEMM but cannot hold 1e-5 A in fp32 (let alone
// TF32/BF16), so this is a register-tiled FP32 SIMT kernel on direct differences:
// a 128 x 128 tile of pairs per CTA, 8 x 8 pairs per thread, operands staged k-major in
// shared memory so that every inner step is 4 LDS.128 feeding 64 FADD + 64 FFMA.
#include "common.cuh"

namespace gmm {

constexpr int kRT = 128;         // tile edge (pairs)
constexpr int kRK = 32;          // coordinates per shared-memory chunk
constexpr int kRS = kRT + 4;     // padded row stride (floats), keeps 16-byte alignment

struct RmsdArgs {
    int N, K, A, row0, row1;
    const float* xyz;
    float* out; uint32_t* bits; int words;
    float thr2, fa;
};

__global__ void __launch_bounds__(256, 2)
rmsd_tile_kernel(const RmsdArgs a) {
    __shared__ __align__(16) float As[kRK][kRS];
    __shared__ __align__(16) float Bs[kRK][kRS];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ty = tid >> 4, tx = tid & 15;
    const int r_base = a.row0 + blockIdx.y * kRT;   // first row (conformer index) of the tile
    const int c_base = blockIdx.x * kRT;            // first column

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < a.K; k0 += kRK) {
        // stage: lane -> coordinate (coalesced), warp -> conformers warp, warp+8, ...
        const int k = k0 + lane;
#pragma unroll 4
        for (int r = warp; r < kRT; r += 8) {
            const int ra = r_base + r, cb = c_base + r;
            float va = 0.f, vb = 0.f;
            if (k < a.K) {
                if (ra < a.row1) va = a.xyz[(size_t)ra * a.K + k];
                if (cb < a.N) vb = a.xyz[(size_t)cb * a.K + k];
            }
            As[lane][r] = va;
            Bs[lane][r] = vb;
        }
        __syncthreads();
#pragma unroll 4
        for (int kk = 0; kk < kRK; ++kk) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[kk][64 + tx * 4]);
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

    // epilogue
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int row = r_base + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        const bool row_ok = row < a.row1;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int col0 = c_base + h * 64 + tx * 4;
            unsigned nib = 0;
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float ms = acc[i][h * 4 + j] / a.fa;   // similarity.py:75
                v[j] = sqrtf(ms);
                // similarity.py:67-69: "if rmsd_squared > threshold*threshold: return False"
                if (row_ok && col0 + j < a.N && !(ms > a.thr2)) nib |= 1u << j;
            }
            if (a.out != nullptr && row_ok) {
                float* o = a.out + (size_t)(row - a.row0) * a.N + col0;
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (col0 + j < a.N) o[j] = v[j];
            }
            if (a.bits != nullptr) {
                unsigned word = nib << ((tx & 7) * 4);
                word |= __shfl_xor_sync(0xffffffffu, word, 1);
                word |= __shfl_xor_sync(0xffffffffu, word, 2);
                word |= __shfl_xor_sync(0xffffffffu, word, 4);
                const int w = (c_base + h * 64) / 32 + (tx >> 3);
                if ((tx & 7) == 0 && row_ok && w < a.words)
                    a.bits[(size_t)(row - a.row0) * a.words + w] = word;
            }
        }
    }
}

int launch_rmsd(gmm_ctx* c, int N, int A, const float* xyz, int row0, int row1, float* out_rmsd,
                uint32_t* out_bits, double thr, cudaStream_t s) {
    (void)c;
    if (row1 <= row0 || N == 0) return GMM_OK;
    RmsdArgs a;
    a.N = N; a.A = A; a.K = 3 * A; a.row0 = row0; a.row1 = row1;
    a.xyz = xyz; a.out = out_rmsd; a.bits = out_bits; a.words = (N + 31) / 32;
    a.thr2 = (float)(thr * thr); a.fa = (float)A;
    dim3 grid((N + kRT - 1) / kRT, (row1 - row0 + kRT - 1) / kRT);
    rmsd_tile_kernel<<<grid, 256, 0, s>>>(a);
    GMM_LAUNCH_CHECK("rmsd_tile_kernel");
    return GMM_OK;
}

}  // namespace gmm
