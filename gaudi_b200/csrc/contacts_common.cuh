// Pieces shared by the Contacts kernels (contacts.cu: tiled sweep + finalize and the brute-force
// pair dump; contacts_fused.cu: the one-kernel path).
#pragma once
#include "common.cuh"

namespace gmm {

struct SweepArgs {
    int P, M, Mp, MW, n_seed, n_hseed, n_tiles, tiles_per_chunk, n_chunks, ZW, n_sseed, cull, all_seed;
    const float* xyz;
    const int32_t* slot2atom; const float* mrad; const uint8_t* mflags;
    const float4* rec4; const uint8_t* rflags; const float* tile_box; const uint32_t* always;
    const float4* sseed;
    const int32_t* ms_ptr; const int32_t* ms_idx;
    // flexible side chains (fused kernel): runs of non-seed slots that move together ("bodies", lo | hi << 16,
    // at most 32 slots each) and per receptor atom the bodies whose range of motion it is within reach of
    const uint32_t* flex_mask; const uint32_t* body_range; int n_bodies;
    uint32_t* zone; long long* partial; unsigned long long* counters;
    Params prm;
    float rad2, thr_min, thr_tight, rmax_mobile, reach_extra;
};

__device__ __forceinline__ float dist2f(float ax, float ay, float az, float bx, float by, float bz) {
    const float dx = ax - bx, dy = ay - by, dz = az - bz;
    return fmaf(dz, dz, fmaf(dy, dy, dx * dx));
}

// fp32 pre-filter cutoff, slightly widened so that the fp64 decision in score_pair is the
// only one that matters.
__device__ __forceinline__ float loose_c2(float ri, float rj, float thr_min) {
    const float c = ri + rj - thr_min;
    return c * c * 1.000004f + 1e-6f;
}

__device__ __forceinline__ bool ms_excluded(const int32_t* ms_ptr, const int32_t* ms_idx, int slot, int j) {
    if (ms_ptr == nullptr) return false;
    for (int e = ms_ptr[slot]; e < ms_ptr[slot + 1]; ++e)
        if (ms_idx[e] == j) return true;
    return false;
}


// Load G individuals' mobile atoms into slot order and decide which mobile atoms are in the zone.
template <int THREADS>
__device__ __forceinline__ void load_ligands(const SweepArgs& a, int p0, int Gn, float4* lig, uint32_t* zm) {
    for (int idx = threadIdx.x; idx < Gn * a.Mp; idx += THREADS) {
        const int g = idx / a.Mp, s = idx - g * a.Mp;
        const int atom = a.slot2atom[s];
        float4 v = make_float4(kFar, kFar, kFar, 0.f);
        if (atom >= 0) {
            const float* src = a.xyz + ((size_t)(p0 + g) * a.M + atom) * 3;
            v = make_float4(src[0], src[1], src[2], a.mrad[s]);
        }
        lig[idx] = v;
    }
    __syncthreads();
    // zone membership of mobile atoms: seeds always; others iff within radius of a seed
    for (int idx = threadIdx.x; idx < Gn * a.MW; idx += THREADS) {
        const int g = idx / a.MW, w = idx - g * a.MW;
        uint32_t bits = 0;
        for (int b = 0; b < 32; ++b) {
            const int s = w * 32 + b;
            if (s >= a.M) break;
            bool in = s < a.n_seed;
            if (!in && !(a.prm.options & GMM_OPT_ONLY_INTERNAL)) {
                const float4 q = lig[g * a.Mp + s];
                for (int k = 0; k < a.n_seed && !in; ++k) {
                    const float4 t = lig[g * a.Mp + k];
                    in = dist2f(t.x, t.y, t.z, q.x, q.y, q.z) <= a.rad2;
                }
                for (int k = 0; k < a.n_sseed && !in; ++k) {
                    const float4 t = a.sseed[k];
                    in = dist2f(t.x, t.y, t.z, q.x, q.y, q.z) <= a.rad2;
                }
            }
            if (in) bits |= 1u << b;
        }
        zm[idx] = bits;
    }
    __syncthreads();
}

}  // namespace gmm
