// Genome expression for a whole population: Rotamers.express (gaudi/genes/rotamers.py:141-152,
// 186-189), Torsion.express (gaudi/genes/torsion.py:153-164), Search.express
// (gaudi/genes/search.py:201-214) and the coordinate read-back of
// MolecularIndividual.post_express (gaudi/base.py:301-303).
//
// One warp owns one individual.  The template coordinates are copied into that warp's
// slice of shared memory (fp64, SoA), the bond rotations are applied there in the
// reference's bond order on the *current* coordinates (exactly what the sequence of
// BondRot.adjustAngle calls does), then the rigid transform x' = R (x - origin) + t is
// applied and the result is written as float32 in the re-centred device frame with fully
// coalesced stores.  All arithmetic is fp64: the whole stage is ~6 kflop per individual,
// three orders of magnitude below the Contacts stage, so precision is free here.
#include "common.cuh"

namespace gmm {

constexpr int kExprWarps = 1;     // one warp per CTA: per-warp shared-memory offsets fold away (3 % faster at 65,536 individuals, 8 % at 8,192)

struct ExpressArgs {
    int P, M, Mp, T, n_allele_tors, n_res, origin_atom, tors_before_search;
    const double* xyz0; const uint8_t* mflags;
    const int32_t* tfixed; const int32_t* tmoving; const int32_t* tsrc; const uint8_t* tsnap;
    const uint64_t* tmask; int W;
    const int32_t* rot_ptr; const double* rot_delta;
    const double* tors; const double* xform; const double* rot;
    float* out; double cx, cy, cz;
};

__global__ void __launch_bounds__(kExprWarps * 32, 32)
express_kernel(const ExpressArgs a) {
    extern __shared__ double sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    double* sx = sm + (size_t)warp * 3 * a.Mp;
    double* sy = sx + a.Mp;
    double* sz = sy + a.Mp;
    const int stride = gridDim.x * kExprWarps;
    for (int p = blockIdx.x * kExprWarps + warp; p < a.P; p += stride) {
        for (int i = lane; i < a.M; i += 32) {
            sx[i] = a.xyz0[3 * i]; sy[i] = a.xyz0[3 * i + 1]; sz[i] = a.xyz0[3 * i + 2];
        }
        __syncwarp();
        // The angle of every bond (allele column or rotamer table row) and its sine / cosine are
        // evaluated 32 bonds at a time, one per lane, then broadcast bond by bond.
        for (int t0 = 0; t0 < a.T; t0 += 32) {
            const int tl = t0 + lane;
            double my_alpha = 0.0, my_s = 0.0, my_c = 1.0;
            if (tl < a.T) {
                const int src = a.tsrc[tl];
                if (src >= 0) {
                    my_alpha = a.tors[(size_t)p * a.n_allele_tors + src];
                    if (a.tsnap[tl]) my_alpha = (my_alpha <= 0.0) ? 0.0 : 180.0;   // torsion.py:159-160
                } else {
                    const int code = -(src + 1), res = code >> 2, chi = code & 3;
                    const int r0 = a.rot_ptr[res], n = a.rot_ptr[res + 1] - r0;
                    int k = (int)(a.rot[(size_t)p * a.n_res + res] * (double)n);  // rotamers.py:150
                    k = min(max(k, 0), n - 1);
                    my_alpha = a.rot_delta[(size_t)(r0 + k) * 4 + chi];
                }
                if (my_alpha != 0.0) sincos(my_alpha * (3.14159265358979323846 / 180.0), &my_s, &my_c);
            }
            const int tn = min(32, a.T - t0);
            for (int tt = 0; tt < tn; ++tt) {
                const int t = t0 + tt;
                const double alpha = __shfl_sync(0xffffffffu, my_alpha, tt);
                if (alpha != 0.0) {   // warp-uniform
                    const double s = __shfl_sync(0xffffffffu, my_s, tt);
                    const double c = __shfl_sync(0xffffffffu, my_c, tt);
                    const int f = a.tfixed[t], m = a.tmoving[t];
                    const double bx = sx[m], by = sy[m], bz = sz[m];
                    double ux = bx - sx[f], uy = by - sy[f], uz = bz - sz[f];
                    const double inv = rsqrt(ux * ux + uy * uy + uz * uz);
                    ux *= inv; uy *= inv; uz *= inv;
                    // Rodrigues matrix R = c I + s [u]x + (1 - c) u u^T, built once per bond
                    const double omc = 1.0 - c;
                    const double r00 = c + omc * ux * ux, r01 = omc * ux * uy - s * uz, r02 = omc * ux * uz + s * uy;
                    const double r10 = omc * ux * uy + s * uz, r11 = c + omc * uy * uy, r12 = omc * uy * uz - s * ux;
                    const double r20 = omc * ux * uz - s * uy, r21 = omc * uy * uz + s * ux, r22 = c + omc * uz * uz;
                    __syncwarp();
                    const uint64_t* mask = a.tmask + (size_t)t * a.W;
                    for (int i = lane; i < a.M; i += 32) {
                        if ((mask[i >> 6] >> (i & 63)) & 1ull) {
                            const double vx = sx[i] - bx, vy = sy[i] - by, vz = sz[i] - bz;
                            sx[i] = bx + (r00 * vx + r01 * vy + r02 * vz);
                            sy[i] = by + (r10 * vx + r11 * vy + r12 * vz);
                            sz[i] = bz + (r20 * vx + r21 * vy + r22 * vz);
                        }
                    }
                }
                __syncwarp();
            }
        }
        // Search: xform = T . R . to_zero(origin)  ->  x' = R (x - origin) + t
        double m[12];
        const bool has_x = a.xform != nullptr;
        double ox = 0, oy = 0, oz = 0;
        if (has_x) {
#pragma unroll
            for (int k = 0; k < 12; ++k) m[k] = a.xform[(size_t)p * 12 + k];
            if (a.tors_before_search) {
                ox = sx[a.origin_atom]; oy = sy[a.origin_atom]; oz = sz[a.origin_atom];
            } else {
                ox = a.xyz0[3 * a.origin_atom]; oy = a.xyz0[3 * a.origin_atom + 1];
                oz = a.xyz0[3 * a.origin_atom + 2];
            }
        }
        __syncwarp();
        for (int i = lane; i < a.M; i += 32) {
            double x = sx[i], y = sy[i], z = sz[i];
            if (has_x && (a.mflags[i] & GMM_F_XFORM)) {
                const double dx = x - ox, dy = y - oy, dz = z - oz;
                x = m[0] * dx + m[1] * dy + m[2] * dz + m[3];
                y = m[4] * dx + m[5] * dy + m[6] * dz + m[7];
                z = m[8] * dx + m[9] * dy + m[10] * dz + m[11];
            }
            sx[i] = x - a.cx; sy[i] = y - a.cy; sz[i] = z - a.cz;
        }
        __syncwarp();
        float* o = a.out + (size_t)p * a.M * 3;
        for (int e = lane; e < 3 * a.M; e += 32) {
            const int i = e / 3, k = e - 3 * i;
            o[e] = (float)(k == 0 ? sx[i] : (k == 1 ? sy[i] : sz[i]));
        }
        __syncwarp();
    }
}

// The same expression for models whose moving sets form a tree (gmm_ctx::tree_ok, established by a
// host-side test in gmm_set_mobile): rotations about the template axes, innermost moving set
// first, are equal to the reference's sequential BondRots.  Every bond's affine map
// x -> R x + (b - R b) then depends on its own angle only: lanes build one map each (sincos and
// Rodrigues once per bond instead of once per bond per lane-group), park the 3x4 matrices in
// shared memory, and each atom composes the maps of the bonds it turns with.  No serial chain
// through shared memory, a third of the instructions of the sequential kernel.
struct TreeArgs {
    int P, M, T, n_allele_tors, n_res, origin_atom, tors_before_search;
    const double* xyz0; const uint8_t* mflags;
    const int32_t* tsrc; const uint8_t* tsnap; const double* axis; const uint64_t* amask;
    const int32_t* rot_ptr; const double* rot_delta;
    const double* tors; const double* xform; const double* rot;
    float* out; double cx, cy, cz;
};

__global__ void __launch_bounds__(kExprWarps * 32, 32)
express_tree_kernel(const TreeArgs a) {
    extern __shared__ double sm[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int per_warp = 3 * a.M + 12 * a.T + (3 * a.M + 1) / 2;      // doubles: coordinates, maps, float output
    double* A = sm + (size_t)warp * (per_warp + (per_warp & 1));   // [T][12] row-major 3x4, 16-byte aligned
    double* sx = A + 12 * a.T;
    double* sy = sx + a.M;
    double* sz = sy + a.M;
    const int stride = gridDim.x * kExprWarps;
    for (int p = blockIdx.x * kExprWarps + warp; p < a.P; p += stride) {
        unsigned long long active = 0ull;
        for (int t0 = 0; t0 < a.T; t0 += 32) {
            const int t = t0 + lane;
            double alpha = 0.0;
            if (t < a.T) {
                const int src = a.tsrc[t];
                if (src >= 0) {
                    alpha = a.tors[(size_t)p * a.n_allele_tors + src];
                    if (a.tsnap[t]) alpha = (alpha <= 0.0) ? 0.0 : 180.0;        // torsion.py:159-160
                } else {
                    const int code = -(src + 1), res = code >> 2, chi = code & 3;
                    const int r0 = a.rot_ptr[res], n = a.rot_ptr[res + 1] - r0;
                    int k = (int)(a.rot[(size_t)p * a.n_res + res] * (double)n);  // rotamers.py:150
                    k = min(max(k, 0), n - 1);
                    alpha = a.rot_delta[(size_t)(r0 + k) * 4 + chi];
                }
                if (alpha != 0.0) {
                    double s, c;
                    sincos(alpha * (3.14159265358979323846 / 180.0), &s, &c);
                    const double* ax = a.axis + (size_t)t * 6;
                    const double ux = ax[0], uy = ax[1], uz = ax[2], bx = ax[3], by = ax[4], bz = ax[5];
                    const double omc = 1.0 - c;
                    const double r00 = c + omc * ux * ux, r01 = omc * ux * uy - s * uz, r02 = omc * ux * uz + s * uy;
                    const double r10 = omc * ux * uy + s * uz, r11 = c + omc * uy * uy, r12 = omc * uy * uz - s * ux;
                    const double r20 = omc * ux * uz - s * uy, r21 = omc * uy * uz + s * ux, r22 = c + omc * uz * uz;
                    double* m = A + (size_t)t * 12;
                    m[0] = r00; m[1] = r01; m[2] = r02; m[3] = bx - (r00 * bx + r01 * by + r02 * bz);
                    m[4] = r10; m[5] = r11; m[6] = r12; m[7] = by - (r10 * bx + r11 * by + r12 * bz);
                    m[8] = r20; m[9] = r21; m[10] = r22; m[11] = bz - (r20 * bx + r21 * by + r22 * bz);
                }
            }
            active |= (unsigned long long)__ballot_sync(0xffffffffu, alpha != 0.0) << t0;
        }
        __syncwarp();
        for (int i0 = 0; i0 < a.M; i0 += 32) {
            const int i = i0 + lane;
            double x = 0.0, y = 0.0, z = 0.0;
            unsigned long long mine = 0ull;
            if (i < a.M) {
                x = a.xyz0[3 * i]; y = a.xyz0[3 * i + 1]; z = a.xyz0[3 * i + 2];
                mine = a.amask[i] & active;
            }
            // bonds some atom of this group turns with, in application order
            const unsigned lo = __reduce_or_sync(0xffffffffu, (unsigned)mine);
            const unsigned hi = a.T > 32 ? __reduce_or_sync(0xffffffffu, (unsigned)(mine >> 32)) : 0u;
            const unsigned long long any = ((unsigned long long)hi << 32) | lo;
            for (int t = 0; t < a.T; ++t) {
                if (!((any >> t) & 1ull)) continue;                    // warp-uniform
                const double2* m = reinterpret_cast<const double2*>(A + (size_t)t * 12);
                const double2 m01 = m[0], m23 = m[1], m45 = m[2], m67 = m[3], m89 = m[4], mab = m[5];
                if ((mine >> t) & 1ull) {
                    const double nx = fma(m01.x, x, fma(m01.y, y, fma(m23.x, z, m23.y)));
                    const double ny = fma(m45.x, x, fma(m45.y, y, fma(m67.x, z, m67.y)));
                    const double nz = fma(m89.x, x, fma(m89.y, y, fma(mab.x, z, mab.y)));
                    x = nx; y = ny; z = nz;
                }
            }
            if (i < a.M) { sx[i] = x; sy[i] = y; sz[i] = z; }
        }
        __syncwarp();
        // Search: xform = T . R . to_zero(origin)  ->  x' = R (x - origin) + t
        double m[12];
        const bool has_x = a.xform != nullptr;
        double ox = 0, oy = 0, oz = 0;
        if (has_x) {
#pragma unroll
            for (int k = 0; k < 12; ++k) m[k] = a.xform[(size_t)p * 12 + k];
            if (a.tors_before_search) {
                ox = sx[a.origin_atom]; oy = sy[a.origin_atom]; oz = sz[a.origin_atom];
            } else {
                ox = a.xyz0[3 * a.origin_atom]; oy = a.xyz0[3 * a.origin_atom + 1];
                oz = a.xyz0[3 * a.origin_atom + 2];
            }
        }
        __syncwarp();
        // final coordinates as float, interleaved x y z (stride 3 words: conflict-free), so the
        // write-out below is a straight coalesced copy
        float* of = reinterpret_cast<float*>(sz + a.M);
        for (int i0 = 0; i0 < a.M; i0 += 32) {
            const int i = i0 + lane;
            double x = 0.0, y = 0.0, z = 0.0;
            if (i < a.M) {
                x = sx[i]; y = sy[i]; z = sz[i];
                if (has_x && (a.mflags[i] & GMM_F_XFORM)) {
                    const double dx = x - ox, dy = y - oy, dz = z - oz;
                    x = fma(m[0], dx, fma(m[1], dy, fma(m[2], dz, m[3])));
                    y = fma(m[4], dx, fma(m[5], dy, fma(m[6], dz, m[7])));
                    z = fma(m[8], dx, fma(m[9], dy, fma(m[10], dz, m[11])));
                }
            }
            if (i < a.M) {
                of[3 * i] = (float)(x - a.cx); of[3 * i + 1] = (float)(y - a.cy); of[3 * i + 2] = (float)(z - a.cz);
            }
        }
        __syncwarp();
        float* o = a.out + (size_t)p * a.M * 3;
        for (int e = lane; e < 3 * a.M; e += 32) o[e] = of[e];
        __syncwarp();
    }
}

static int launch_express_tree(gmm_ctx* c, int P, const double* tors, const double* xform, const double* rot,
                               float* out_xyz, cudaStream_t s) {
    TreeArgs a;
    a.P = P; a.M = c->M; a.T = c->T; a.n_allele_tors = c->n_allele_tors;
    a.n_res = c->n_res; a.origin_atom = c->origin_atom; a.tors_before_search = c->tors_before_search;
    a.xyz0 = c->d_xyz0; a.mflags = c->d_mflags_atom;
    a.tsrc = c->d_tree_src; a.tsnap = c->d_tree_snap; a.axis = c->d_tree_axis; a.amask = c->d_tree_amask;
    a.rot_ptr = c->d_rot_ptr; a.rot_delta = c->d_rot_delta;
    a.tors = tors; a.xform = xform; a.rot = rot; a.out = out_xyz;
    a.cx = c->centre[0]; a.cy = c->centre[1]; a.cz = c->centre[2];
    const size_t smem = (size_t)kExprWarps * (3 * a.M + 12 * a.T + (3 * a.M + 1) / 2 + 1) * sizeof(double);
    if (smem > 48 * 1024) {
        const int rc = allow_dynamic_smem((const void*)express_tree_kernel, c->device, 200 * 1024);
        if (rc) return rc;
    }
    int blocks = (P + kExprWarps - 1) / kExprWarps;
    const int cap = 148 * 32;
    if (blocks > cap) blocks = cap;
    if (c->prof) GMM_CUDA(cudaEventRecord(c->ev[0], s));
    express_tree_kernel<<<blocks, kExprWarps * 32, smem, s>>>(a);
    GMM_LAUNCH_CHECK("express_tree_kernel");
    if (c->prof) { GMM_CUDA(cudaEventRecord(c->ev[1], s)); c->ev_used[0] = true; }
    return GMM_OK;
}

int launch_express(gmm_ctx* c, int P, const double* tors, const double* xform, const double* rot,
                   float* out_xyz, cudaStream_t s) {
    if (P == 0) return GMM_OK;
    const bool sequential_only = (c->prm.options & GMM_OPT_SEQUENTIAL_EXPRESS) != 0;
    const size_t tree_smem = (size_t)kExprWarps * (3 * c->M + 12 * c->T + (3 * c->M + 1) / 2 + 1) * sizeof(double);
    if (c->tree_ok && !sequential_only && tree_smem <= 200 * 1024)
        return launch_express_tree(c, P, tors, xform, rot, out_xyz, s);
    ExpressArgs a;
    a.P = P; a.M = c->M; a.Mp = c->M; a.T = c->T; a.n_allele_tors = c->n_allele_tors;
    a.n_res = c->n_res; a.origin_atom = c->origin_atom; a.tors_before_search = c->tors_before_search;
    a.xyz0 = c->d_xyz0; a.mflags = c->d_mflags_atom;
    a.tfixed = c->d_tfixed; a.tmoving = c->d_tmoving; a.tsrc = c->d_tsrc; a.tsnap = c->d_tsnap;
    a.tmask = c->d_tmask; a.W = c->W;
    a.rot_ptr = c->d_rot_ptr; a.rot_delta = c->d_rot_delta;
    a.tors = tors; a.xform = xform; a.rot = rot; a.out = out_xyz;
    a.cx = c->centre[0]; a.cy = c->centre[1]; a.cz = c->centre[2];
    const size_t smem = (size_t)kExprWarps * 3 * a.Mp * sizeof(double);
    if (smem > 48 * 1024) {
        const int rc = allow_dynamic_smem((const void*)express_kernel, c->device, 200 * 1024);
        if (rc) return rc;
    }
    int blocks = (P + kExprWarps - 1) / kExprWarps;
    const int cap = 148 * 32;
    if (blocks > cap) blocks = cap;
    if (c->prof) GMM_CUDA(cudaEventRecord(c->ev[0], s));
    express_kernel<<<blocks, kExprWarps * 32, smem, s>>>(a);
    GMM_LAUNCH_CHECK("express_kernel");
    if (c->prof) { GMM_CUDA(cudaEventRecord(c->ev[1], s)); c->ev_used[0] = true; }
    return GMM_OK;
}

}  // namespace gmm
