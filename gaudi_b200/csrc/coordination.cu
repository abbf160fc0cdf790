// Coordination objective on the expressed population ("next" row 4;
// gaudi/objectives/coordination.py:180-283 with gaudi/_cpdrift.py:45-93,127-199 and the
// directionality term coordination.py:286-372).  Per individual and per metal site:
//
//   coord_sphere_kernel   one warp per individual: atoms within `radius` of the metal
//                         (specifier.zone, d <= radius) are collected with a ballot-compacted scan,
//                         ranked by |distance - d| (rank = number of smaller keys, no sort network:
//                         the zone holds a few dozen atoms), the walk of coordination.py:256-262 is
//                         a prefix over the ranked list, the ligand atoms' directionality terms run
//                         one per lane, and the point set (ligands + metal) goes to a scratch row;
//   coord_cpd_kernel      one thread per (individual, start rotation): rigid coherent point drift
//                         (2^3 starts x <= 10 EM steps, as the reference calls it) in fp64; the 8
//                         lanes of an individual then take the first smallest RMSD by shuffles.
//
// The rotation of an EM step is the proper rotation maximising trace(A^T R) — the reference's
// U C V^T of the SVD of A — found as the top eigenvector of Horn's 4 x 4 quaternion matrix by
// Jacobi rotations: no special cases for rank-deficient A (planar and linear point sets are the rule here).
// Everything is fp64 (the reference is numpy fp64); coordinates are the expressed float32 poses
// plus the frame centre, and the caller's doubles for static atoms.
#include <cmath>
#include <vector>

#include "common.cuh"

namespace gmm {

constexpr int kCoordMaxVertices = 8;     // cube / hexagonal bipyramid: 8 vertices + the metal
constexpr int kCoordMaxPoints = kCoordMaxVertices + 1;
constexpr int kCoordZoneCap = 192;       // atoms within the radius of a metal, per individual
constexpr int kCoordStarts = 8;          // guess_steps = 2 -> 2^3 start rotations (coordination.py:204)
constexpr int kCoordIters = 10;          // max_iterations = 10
constexpr int kCoordMaxNb = 8;           // substituents of a ligand atom that are looked at
constexpr double kRad = 3.14159265358979323846 / 180.0;

struct CoordSite {
    int n_atoms, n_scan;
    const int32_t* scan;     // [n_scan] atoms that can be within the radius of the metal
    const int32_t* row;      // [n_atoms] row of the expressed pose, -1: static
    const double* sxyz;      // [n_atoms, 3] world frame, used for static atoms
    const uint8_t* cand;     // [n_atoms] valid ligand atom of a requested residue
    const uint8_t* geom;     // [n_atoms] 4 tetrahedral, 3 planar, 2 linear, 1 single
    const double* brad;      // [n_atoms] covalent bond radius
    const int32_t* nb_ptr; const int32_t* nb_idx;   // bonded atoms, in bond order
    int metal, n_vertices, min_atoms, flags;
    double radius, distance, weight;
    double tmpl[kCoordMaxPoints * 3];
    double start[kCoordStarts * 9];
};

struct CoordWork {           // scratch row per (individual, site)
    double X[kCoordMaxPoints * 3];
    double partial;          // directionality + corrections, or the final penalty
    int n_pts;               // ligands + 1
    int status;              // 0: run the CPD; 1: `partial` is final; 2: zone overflow
};

struct CoordState {          // host side, owned by the context
    std::vector<void*> owned;
    CoordSite* d_sites = nullptr;
    int n_sites = 0;
    CoordWork* d_work = nullptr; size_t work_rows = 0;
};

__device__ __forceinline__ void coord_xyz(const CoordSite& s, const float* pose, const double* ctr, int g, double* o) {
    const int r = s.row[g];
    if (r >= 0) {
        o[0] = (double)pose[3 * r] + ctr[0]; o[1] = (double)pose[3 * r + 1] + ctr[1]; o[2] = (double)pose[3 * r + 2] + ctr[2];
    } else {
        o[0] = s.sxyz[3 * (size_t)g]; o[1] = s.sxyz[3 * (size_t)g + 1]; o[2] = s.sxyz[3 * (size_t)g + 2];
    }
}

__device__ __forceinline__ double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
__device__ __forceinline__ void sub3(const double* a, const double* b, double* o) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
__device__ __forceinline__ void cross3(const double* a, const double* b, double* o) {
    o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
__device__ __forceinline__ void unit3(double* v) { const double n = sqrt(dot3(v, v)); v[0] /= n; v[1] /= n; v[2] /= n; }

// chimera.angle(a, b, c): the angle at b, degrees
__device__ double coord_angle(const double* a, const double* b, const double* c) {
    double u[3], v[3];
    sub3(a, b, u); sub3(c, b, v);
    const double cs = dot3(u, v) / sqrt(dot3(u, u) * dot3(v, v));
    return acos(fmin(1.0, fmax(-1.0, cs))) / kRad;
}

// chimera.dihedral(p0, p1, p2, p3), IUPAC sign, degrees (same form as geometry.cu)
__device__ double coord_dihedral(const double* p0, const double* p1, const double* p2, const double* p3) {
    double b0[3], b1[3], b2[3], v[3], w[3], cr[3];
    sub3(p0, p1, b0); sub3(p2, p1, b1); sub3(p3, p2, b2);
    unit3(b1);
    const double d0 = dot3(b0, b1), d2 = dot3(b2, b1);
    for (int k = 0; k < 3; ++k) { v[k] = b0[k] - d0 * b1[k]; w[k] = b2[k] - d2 * b1[k]; }
    cross3(b1, v, cr);
    return atan2(dot3(cr, w), dot3(v, w)) / kRad;
}

// The direction perpendicular to x in the plane of x and the y axis (the x axis when x is along
// y): Chimera's choice for a planar atom with one substituent, fixed by the reference's golden
// score 2.4696 (tests/test_objectives_coordination.py:33-47).
__device__ void coord_perpendicular(const double* x, double* y) {
    double e[3] = {0.0, 1.0, 0.0};
    double d = dot3(e, x);
    for (int k = 0; k < 3; ++k) y[k] = e[k] - d * x[k];
    if (dot3(y, y) <= 1e-12) {
        e[0] = 1.0; e[1] = 0.0;
        d = dot3(e, x);
        for (int k = 0; k < 3; ++k) y[k] = e[k] - d * x[k];
    }
    unit3(y);
}

// First free valence position of an atom at L with substituents nb[0..n) (coordination.py:363-372
// -> chimera.bondGeom.bondPositions); false when every position is taken.
__device__ bool coord_free_position(const double* L, int geom, double len, const double (*nb)[3], int n, double* pos) {
    if (geom < 1 || n >= geom) return false;
    if (n == 0) { pos[0] = L[0] + len; pos[1] = L[1]; pos[2] = L[2]; return true; }
    double u[3][3];
    for (int k = 0; k < n && k < 3; ++k) { sub3(nb[k], L, u[k]); unit3(u[k]); }
    double x[3] = {-u[0][0], -u[0][1], -u[0][2]}, dir[3];
    if (geom == 2) {
        for (int k = 0; k < 3; ++k) dir[k] = x[k];
    } else if (geom == 3 && n == 2) {
        for (int k = 0; k < 3; ++k) dir[k] = -(u[0][k] + u[1][k]);
        unit3(dir);
    } else if (geom == 3) {
        double y[3];
        coord_perpendicular(x, y);
        const double c = cos(60.0 * kRad), s = sin(60.0 * kRad);
        for (int k = 0; k < 3; ++k) dir[k] = c * x[k] + s * y[k];
    } else if (n == 3) {
        for (int k = 0; k < 3; ++k) dir[k] = -(u[0][k] + u[1][k] + u[2][k]);
        unit3(dir);
    } else if (n == 2) {
        double v[3], w[3];
        for (int k = 0; k < 3; ++k) v[k] = -(u[0][k] + u[1][k]);
        unit3(v);
        cross3(u[0], u[1], w);
        unit3(w);
        const double half = 0.5 * 109.4712206 * kRad;
        for (int k = 0; k < 3; ++k) dir[k] = cos(half) * v[k] + sin(half) * w[k];
    } else {
        double y[3];
        coord_perpendicular(x, y);
        const double a = (180.0 - 109.4712206) * kRad;
        for (int k = 0; k < 3; ++k) dir[k] = cos(a) * x[k] + sin(a) * y[k];
    }
    for (int k = 0; k < 3; ++k) pos[k] = L[k] + len * dir[k];
    return true;
}

// coordination.py:286-360
__device__ double coord_deviation(const CoordSite& s, const float* pose, const double* ctr, const double* Mx,
                                  int ligand, const int* ligands, int n_lig) {
    const int e0 = s.nb_ptr[ligand], nn = s.nb_ptr[ligand + 1] - e0;
    const int geom = s.geom[ligand];
    double L[3], nbx[kCoordMaxNb][3];
    coord_xyz(s, pose, ctr, ligand, L);
    const int nload = min(nn, kCoordMaxNb);
    for (int k = 0; k < nload; ++k) coord_xyz(s, pose, ctr, s.nb_idx[e0 + k], nbx[k]);
    double pos[3];
    if (nn > kCoordMaxNb || !coord_free_position(L, geom, s.brad[ligand] + s.brad[s.metal], nbx, nn, pos)) return 1.0;
    bool rotates = geom == 4 && nn == 1;
    // bidentate: another ligand atom bonded to one of this one's substituents (the first such ligand;
    // of several shared substituents the lowest index)
    int mate = -1, shared = -1;
    for (int k = 0; k < n_lig && mate < 0; ++k) {
        const int o = ligands[k];
        if (o == ligand) continue;
        for (int a = 0; a < nn; ++a) {
            const int x = s.nb_idx[e0 + a];
            bool common = false;
            for (int b = s.nb_ptr[o]; b < s.nb_ptr[o + 1]; ++b) common = common || s.nb_idx[b] == x;
            if (common && (shared < 0 || x < shared)) { shared = x; mate = o; }
        }
    }
    if (nn == 0) return 0.0;
    const int nb = s.nb_idx[e0];
    int nnb = -1;
    for (int b = s.nb_ptr[nb]; b < s.nb_ptr[nb + 1] && nnb < 0; ++b)
        if (s.nb_idx[b] != ligand) nnb = s.nb_idx[b];
    if (nnb < 0) rotates = true;
    const double* Nb = nbx[0];
    const double actual_angle = coord_angle(Nb, L, Mx);
    double angle_diff = coord_angle(Nb, L, pos) - actual_angle;
    double NN[3] = {0, 0, 0}, bi[3];
    const bool bidentate = mate >= 0 && !rotates;
    if (nnb >= 0) coord_xyz(s, pose, ctr, nnb, NN);
    if (bidentate) {
        double c[3], b[3];
        coord_xyz(s, pose, ctr, shared, c);
        coord_xyz(s, pose, ctr, mate, b);
        for (int k = 0; k < 3; ++k) bi[k] = c[k] + 1.5 * ((L[k] - c[k]) + (b[k] - c[k]));
        angle_diff = fmin(angle_diff, coord_angle(Nb, L, bi) - actual_angle);
    }
    const double out = fabs(sin(angle_diff * kRad));
    if (rotates) return out;
    const double actual_dihedral = coord_dihedral(NN, Nb, L, Mx);
    double dihedral_diff = coord_dihedral(NN, Nb, L, pos) - actual_dihedral;
    if (bidentate) dihedral_diff = fmin(dihedral_diff, coord_dihedral(NN, Nb, L, bi) - actual_dihedral);
    return out + fabs(sin(dihedral_diff * kRad));
}

__global__ void __launch_bounds__(128)
coord_sphere_kernel(int P, int M, const float* xyz, const CoordSite* sites, int n_sites, double cx, double cy, double cz,
                    CoordWork* work, unsigned long long* counters) {
    __shared__ double s_key[4][kCoordZoneCap], s_skey[4][kCoordZoneCap];
    __shared__ int s_id[4][kCoordZoneCap], s_sid[4][kCoordZoneCap];
    __shared__ int s_lig[4][kCoordMaxVertices];
    __shared__ double s_dev[4][kCoordMaxVertices];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int p = blockIdx.x * 4 + warp;
    if (p >= P) return;
    const CoordSite& s = sites[blockIdx.y];
    CoordWork& w = work[(size_t)p * n_sites + blockIdx.y];
    const float* pose = xyz + (size_t)p * M * 3;
    const double ctr[3] = {cx, cy, cz};
    double Mx[3];
    coord_xyz(s, pose, ctr, s.metal, Mx);
    double* key = s_key[warp]; int* id = s_id[warp];
    double* skey = s_skey[warp]; int* sid = s_sid[warp];

    // ---- atoms within the radius ---------------------------------------------------------
    int n = 0;
    bool overflow = false;
    for (int base = 0; base < s.n_scan; base += 32) {
        const int i = base + lane;
        bool in = false;
        double k = 0.0;
        int g = -1;
        if (i < s.n_scan) {
            g = s.scan[i];
            double q[3];
            coord_xyz(s, pose, ctr, g, q);
            const double dx = q[0] - Mx[0], dy = q[1] - Mx[1], dz = q[2] - Mx[2];
            // (x^2 + y^2) + z^2 with every product rounded, as numpy evaluates it: the keys order
            // the walk, so they are kept free of contraction into FMAs
            const double d = sqrt(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)));
            in = g != s.metal && d <= s.radius;
            k = fabs(s.distance - d);
        }
        const unsigned m = __ballot_sync(0xffffffffu, in);
        const int at = n + __popc(m & ((1u << lane) - 1u));
        if (in) {
            if (at < kCoordZoneCap) { key[at] = k; id[at] = g; } else overflow = true;
        }
        n += __popc(m);
    }
    overflow = __any_sync(0xffffffffu, overflow);
    __syncwarp();
    if (overflow) {
        if (lane == 0) {
            w.status = 2; w.n_pts = 0; w.partial = __longlong_as_double(0x7ff8000000000000ll);
            atomicAdd(counters + 2, 1ull);
        }
        return;
    }
    // ---- rank by (key, atom index) -------------------------------------------------------
    for (int e = lane; e < n; e += 32) {
        const double ke = key[e]; const int ie = id[e];
        int rank = 0;
        for (int f = 0; f < n; ++f) rank += (key[f] < ke || (key[f] == ke && id[f] < ie)) ? 1 : 0;
        skey[rank] = ke; sid[rank] = ie;
    }
    __syncwarp();
    // ---- the walk (coordination.py:256-262) ----------------------------------------------
    const bool prevent = s.flags & 1;
    int n_found = 0, n_lig = 0;
    bool stopped = false;
    for (int base = 0; base < n && !stopped; base += 32) {
        const int i = base + lane;
        bool take = false;
        if (i < n) take = s.cand[sid[i]] && skey[i] > 1.0;
        unsigned m = __ballot_sync(0xffffffffu, take);
        const unsigned live = (n - base >= 32) ? 0xffffffffu : ((1u << (n - base)) - 1u);
        if (prevent) {
            const unsigned bad = ~m & live;
            if (bad) { m &= (bad & (0u - bad)) - 1u; stopped = true; }     // keep what precedes the first intruder
        }
        if (take && ((m >> lane) & 1u)) {
            const int at = n_found + __popc(m & ((1u << lane) - 1u));
            if (at < s.n_vertices) s_lig[warp][at] = sid[i];
        }
        n_found += __popc(m);
    }
    n_lig = min(n_found, s.n_vertices);
    __syncwarp();
    if ((s.flags & 2) || s.min_atoms - n_found > 0) {
        if (lane == 0) {
            // enforce_all_residues compares a set with a list (coordination.py:264): never satisfied
            w.partial = (s.flags & 2) ? -1000.0 * s.weight : -100.0 * (s.min_atoms - n_found) * s.weight;
            w.status = 1; w.n_pts = 0;
        }
        return;
    }
    // ---- point set, directionality, corrections ---------------------------------------------
    if (lane < n_lig) {
        double q[3];
        coord_xyz(s, pose, ctr, s_lig[warp][lane], q);
        w.X[3 * lane] = q[0]; w.X[3 * lane + 1] = q[1]; w.X[3 * lane + 2] = q[2];
        double v = coord_deviation(s, pose, ctr, Mx, s_lig[warp][lane], s_lig[warp], n_lig);
        if (s.flags & 8) {        // distance_correction: |bond length - ideal|
            const double dx = q[0] - Mx[0], dy = q[1] - Mx[1], dz = q[2] - Mx[2];
            v += fabs(sqrt(dx * dx + dy * dy + dz * dz) - (s.brad[s_lig[warp][lane]] + s.brad[s.metal]));
        }
        s_dev[warp][lane] = v;
    }
    __syncwarp();
    if (lane == 0) {
        double total = 0.0, com[3] = {0, 0, 0};
        for (int k = 0; k < n_lig; ++k) {
            total += s_dev[warp][k];
            com[0] += w.X[3 * k]; com[1] += w.X[3 * k + 1]; com[2] += w.X[3 * k + 2];
        }
        if ((s.flags & 4) && n_lig > 0) {     // center_of_mass_correction
            const double dx = Mx[0] - com[0] / n_lig, dy = Mx[1] - com[1] / n_lig, dz = Mx[2] - com[2] / n_lig;
            total += sqrt(dx * dx + dy * dy + dz * dz);
        }
        w.X[3 * n_lig] = Mx[0]; w.X[3 * n_lig + 1] = Mx[1]; w.X[3 * n_lig + 2] = Mx[2];
        w.partial = total; w.n_pts = n_lig + 1; w.status = 0;
    }
}

// Cyclic Jacobi on a symmetric 4 x 4 matrix held in registers (every index below is a compile-time
// constant after unrolling).  `v` comes in as an orthogonal basis in which `a` is expressed and
// leaves as the eigenvectors: an EM step starts from the eigenvectors of the step before, in which
// the new matrix is nearly diagonal, so two or three sweeps do where a cold start needs six.
__device__ __forceinline__ void coord_jacobi(double (&a)[4][4], double (&v)[4][4]) {
#pragma unroll 1
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0, diag = 0.0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            diag += a[i][i] * a[i][i];
#pragma unroll
            for (int j = i + 1; j < 4; ++j) off += a[i][j] * a[i][j];
        }
        if (off <= 1e-32 * diag || off == 0.0) break;
#pragma unroll
        for (int p = 0; p < 3; ++p)
#pragma unroll
            for (int q = p + 1; q < 4; ++q) {
                const double apq = a[p][q];
                if (apq != 0.0) {
                    // t = sgn(theta) / (|theta| + sqrt(theta^2 + 1)), theta = (a_qq - a_pp) / (2 a_pq),
                    // with numerator and denominator scaled by 2 |a_pq|: one square root, one division
                    const double d = a[q][q] - a[p][p];
                    const double t = (d >= 0.0 ? 2.0 * apq : -2.0 * apq) / (fabs(d) + sqrt(d * d + 4.0 * apq * apq));
                    const double c = rsqrt(t * t + 1.0), sn = t * c;
                    a[p][p] -= t * apq; a[q][q] += t * apq;
                    a[p][q] = 0.0; a[q][p] = 0.0;
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        if (r != p && r != q) {
                            const double arp = a[r][p], arq = a[r][q];
                            a[r][p] = c * arp - sn * arq; a[p][r] = a[r][p];
                            a[r][q] = sn * arp + c * arq; a[q][r] = a[r][q];
                        }
                    }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const double vkp = v[k][p], vkq = v[k][q];
                        v[k][p] = c * vkp - sn * vkq; v[k][q] = sn * vkp + c * vkq;
                    }
                }
            }
    }
}

__device__ __forceinline__ void rot_apply(const double* R, const double* y, double* o) {
    o[0] = R[0] * y[0] + R[1] * y[1] + R[2] * y[2];
    o[1] = R[3] * y[0] + R[4] * y[1] + R[5] * y[2];
    o[2] = R[6] * y[0] + R[7] * y[1] + R[8] * y[2];
}

// Rigid coherent point drift, one thread per (individual, start rotation) (_cpdrift.py:45-93,127-149
// with w = 0).  Written in coordinates centred on the mean of X.  Every column of the reference's P
// sums to one, so its row-sum vector over X, N_p and mu_x are N ones, N and the mean of X up to
// rounding; they are still accumulated as sums here, but that is why A needs no stored P: with
// q_m = sum_n P_mn (x_n - mean) the reference's X^T P^T Y (hats omitted) is sum_m q_m (y_m - mu_y)^T.
__global__ void __launch_bounds__(128)
coord_cpd_kernel(int P, const CoordSite* sites, int n_sites, const CoordWork* work, double* out) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    const int item = idx / kCoordStarts, start = idx % kCoordStarts;
    const bool active = item < P * n_sites;
    const int site = active ? item % n_sites : 0;
    const CoordSite& s = sites[site];
    const CoordWork& w = work[active ? item : 0];
    double rmsd = 1000.0;
    bool ok = true;
    if (active && w.status == 0) {
        const int N = w.n_pts, Mv = s.n_vertices + 1;
        double X[kCoordMaxPoints][3], TY[kCoordMaxPoints][3], q[kCoordMaxPoints][3], p1[kCoordMaxPoints], o[kCoordMaxPoints];
        double mean[3] = {0, 0, 0}, xvar = 0.0;
        for (int i = 0; i < N; ++i) for (int k = 0; k < 3; ++k) { X[i][k] = w.X[3 * i + k]; mean[k] += X[i][k]; }
        for (int k = 0; k < 3; ++k) mean[k] /= N;
        for (int i = 0; i < N; ++i) for (int k = 0; k < 3; ++k) { X[i][k] -= mean[k]; xvar += X[i][k] * X[i][k]; }
        const double* Y = s.tmpl;
        const double sx = sqrt(xvar / (3.0 * N));           // _cpdrift.py:156-157 of X
        double R[9];
        for (int k = 0; k < 9; ++k) R[k] = s.start[9 * start + k];
        // s = std(X) / std(R Y); t = mean(X) - mean(s R Y); sigma^2 = mean squared distance / 3
        double ym[3] = {0, 0, 0}, yvar = 0.0;
        for (int m = 0; m < Mv; ++m) { rot_apply(R, Y + 3 * m, TY[m]); for (int k = 0; k < 3; ++k) ym[k] += TY[m][k]; }
        for (int k = 0; k < 3; ++k) ym[k] /= Mv;
        for (int m = 0; m < Mv; ++m) for (int k = 0; k < 3; ++k) yvar += (TY[m][k] - ym[k]) * (TY[m][k] - ym[k]);
        double sc = sx / sqrt(yvar / (3.0 * Mv));
        double t[3] = {-sc * ym[0], -sc * ym[1], -sc * ym[2]};
        double sigma2 = 0.0;
        for (int m = 0; m < Mv; ++m)
            for (int n = 0; n < N; ++n)
                for (int k = 0; k < 3; ++k) { const double d = sc * TY[m][k] + t[k] - X[n][k]; sigma2 += d * d; }
        sigma2 /= 3.0 * Mv * N;
        double V[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
        int done = 0;
#pragma unroll 1
        for (int it = 0; it < kCoordIters; ++it) {
            if (!(sigma2 > 0.0)) break;          // the step's arithmetic would be invalid (_cpdrift.py:79-86)
            for (int m = 0; m < Mv; ++m) {
                rot_apply(R, Y + 3 * m, TY[m]);
                for (int k = 0; k < 3; ++k) { TY[m][k] = sc * TY[m][k] + t[k]; q[m][k] = 0.0; }
                p1[m] = 0.0;
            }
            const double inv2s = -1.0 / (2.0 * sigma2);
            bool valid = true;
            double e[3] = {0, 0, 0}, xv = 0.0;
#pragma unroll 1
            for (int n = 0; n < N; ++n) {
                double col = 0.0;
                for (int m = 0; m < Mv; ++m) {
                    const double dx = TY[m][0] - X[n][0], dy = TY[m][1] - X[n][1], dz = TY[m][2] - X[n][2];
                    o[m] = exp((dx * dx + dy * dy + dz * dz) * inv2s);
                    col += o[m];
                }
                if (col == 0.0 || !isfinite(col)) { valid = false; break; }
                const double inv = 1.0 / col;
                double pt = 0.0;
                for (int m = 0; m < Mv; ++m) {
                    const double pm = o[m] * inv;
                    pt += pm; p1[m] += pm;
                    q[m][0] += pm * X[n][0]; q[m][1] += pm * X[n][1]; q[m][2] += pm * X[n][2];
                }
                e[0] += pt * X[n][0]; e[1] += pt * X[n][1]; e[2] += pt * X[n][2];
                xv += pt * (X[n][0] * X[n][0] + X[n][1] * X[n][1] + X[n][2] * X[n][2]);
            }
            if (!valid) break;
            double Np = 0.0, muy[3] = {0, 0, 0};
            for (int m = 0; m < Mv; ++m) {
                Np += p1[m];
                for (int k = 0; k < 3; ++k) muy[k] += Y[3 * m + k] * p1[m];
            }
            for (int k = 0; k < 3; ++k) { muy[k] /= Np; e[k] /= Np; }          // e: mu_x in the centred frame (~1e-17)
            xv -= Np * (e[0] * e[0] + e[1] * e[1] + e[2] * e[2]);
            double A[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}}, den = 0.0;
            for (int m = 0; m < Mv; ++m) {
                const double y0 = Y[3 * m] - muy[0], y1 = Y[3 * m + 1] - muy[1], y2 = Y[3 * m + 2] - muy[2];
                den += p1[m] * (y0 * y0 + y1 * y1 + y2 * y2);
#pragma unroll
                for (int a = 0; a < 3; ++a) {
                    const double qa = q[m][a] - p1[m] * e[a];
                    A[a][0] += qa * y0; A[a][1] += qa * y1; A[a][2] += qa * y2;
                }
            }
            // Horn's quaternion matrix of S = A^T (S_ab = sum w y_a x_b), expressed in the basis V
            const double Sxx = A[0][0], Sxy = A[1][0], Sxz = A[2][0], Syx = A[0][1], Syy = A[1][1], Syz = A[2][1],
                         Szx = A[0][2], Szy = A[1][2], Szz = A[2][2];
            const double Nq[4][4] = {{Sxx + Syy + Szz, Syz - Szy, Szx - Sxz, Sxy - Syx},
                                     {Syz - Szy, Sxx - Syy - Szz, Sxy + Syx, Szx + Sxz},
                                     {Szx - Sxz, Sxy + Syx, -Sxx + Syy - Szz, Syz + Szy},
                                     {Sxy - Syx, Szx + Sxz, Syz + Szy, -Sxx - Syy + Szz}};
            double B[4][4], Wm[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    Wm[i][j] = Nq[i][0] * V[0][j] + Nq[i][1] * V[1][j] + Nq[i][2] * V[2][j] + Nq[i][3] * V[3][j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = i; j < 4; ++j) {
                    B[i][j] = V[0][i] * Wm[0][j] + V[1][i] * Wm[1][j] + V[2][i] * Wm[2][j] + V[3][i] * Wm[3][j];
                    B[j][i] = B[i][j];
                }
            coord_jacobi(B, V);
            int top = 0;
#pragma unroll
            for (int i = 1; i < 4; ++i) if (B[i][i] > B[top][top]) top = i;
            double qw = 0, qx = 0, qy = 0, qz = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) if (i == top) { qw = V[0][i]; qx = V[1][i]; qy = V[2][i]; qz = V[3][i]; }
            const double qn = rsqrt(qw * qw + qx * qx + qy * qy + qz * qz);
            qw *= qn; qx *= qn; qy *= qn; qz *= qn;
            R[0] = 1 - 2 * (qy * qy + qz * qz); R[1] = 2 * (qx * qy - qz * qw); R[2] = 2 * (qx * qz + qy * qw);
            R[3] = 2 * (qx * qy + qz * qw); R[4] = 1 - 2 * (qx * qx + qz * qz); R[5] = 2 * (qy * qz - qx * qw);
            R[6] = 2 * (qx * qz - qy * qw); R[7] = 2 * (qy * qz + qx * qw); R[8] = 1 - 2 * (qx * qx + qy * qy);
            double tr = 0.0;
#pragma unroll
            for (int a = 0; a < 3; ++a) tr += A[a][0] * R[3 * a] + A[a][1] * R[3 * a + 1] + A[a][2] * R[3 * a + 2];
            sc = tr / den;
            double rm[3];
            rot_apply(R, muy, rm);
            for (int k = 0; k < 3; ++k) t[k] = e[k] - sc * rm[k];
            sigma2 = (xv - sc * tr) / (Np * 3.0);
            if (!isfinite(sc) || !isfinite(t[0]) || !isfinite(t[1]) || !isfinite(t[2])) { done = 0; break; }
            ++done;
        }
        if (done == 0) {
            ok = false;            // the reference fails in last() of an empty sequence: evaluate's penalty
        } else {
            double acc = 0.0;
            for (int m = 0; m < Mv; ++m) { rot_apply(R, Y + 3 * m, TY[m]); for (int k = 0; k < 3; ++k) TY[m][k] = sc * TY[m][k] + t[k]; }
            for (int n = 0; n < N; ++n) {
                double best = 1.79e308;
                for (int m = 0; m < Mv; ++m) {
                    const double dx = TY[m][0] - X[n][0], dy = TY[m][1] - X[n][1], dz = TY[m][2] - X[n][2];
                    best = fmin(best, dx * dx + dy * dy + dz * dz);
                }
                acc += best;
            }
            rmsd = sqrt(acc / N) / sx;      // _cpdrift.py:193-198
        }
    }
    // the individual's 8 starts sit in 8 consecutive lanes: first smallest RMSD, any failed start fails all
    const int lane = threadIdx.x & 31, base = lane & ~(kCoordStarts - 1);
    double best = 1000.0;
    bool all_ok = true;
    for (int r = 0; r < kCoordStarts; ++r) {
        const double v = __shfl_sync(0xffffffffu, rmsd, base + r);
        const int o2 = __shfl_sync(0xffffffffu, (int)ok, base + r);
        all_ok = all_ok && o2;
        if (v < best) best = v;
    }
    if (active && start == 0) {
        double v = w.partial;
        if (w.status == 0) v = all_ok ? w.partial + best : -1000.0 * s.weight;
        out[item] = v;
    }
}

static void coord_release(gmm_ctx* c) {
    CoordState* st = static_cast<CoordState*>(c->coordination);
    if (!st) return;
    for (void* p : st->owned) cudaFree(p);
    if (st->d_sites) cudaFree(st->d_sites);
    if (st->d_work) cudaFree(st->d_work);
    delete st;
    c->coordination = nullptr;
}

void coordination_destroy(gmm_ctx* c) { coord_release(c); }

// _cpdrift.py:168-185 (Shoemake's uniform quaternions on a regular grid -> angle / axis -> matrix)
static void coord_start_rotations(double* out) {
    const int steps = 2;
    const double two_pi = 2.0 * 3.14159265358979323846;
    int n = 0;
    for (int ix = 0; ix < steps; ++ix)
        for (int i0 = 0; i0 < steps; ++i0)
            for (int i1 = 0; i1 < steps; ++i1, ++n) {
                const double x = (double)ix / steps, t0 = two_pi * i0 / steps, t1 = two_pi * i1 / steps;
                const double r0 = std::sqrt(1.0 - x), r1 = std::sqrt(x);
                const double qs = std::sin(t0) * r0, qi = std::cos(t0) * r0, qj = std::sin(t1) * r1, qk = std::cos(t1) * r1;
                const double theta = 2.0 * std::acos(qs), sh = std::sin(theta / 2.0);
                double ux = 1.0, uy = 0.0, uz = 0.0;
                if (sh != 0.0) { ux = qi / sh; uy = qj / sh; uz = qk / sh; }
                const double c = std::cos(theta), s = std::sin(theta);
                double* R = out + 9 * n;
                R[0] = c + ux * ux * (1 - c); R[1] = ux * uy * (1 - c) - uz * s; R[2] = (1 - c) * ux * uz + uy * s;
                R[3] = uy * ux * (1 - c) + uz * s; R[4] = c + uy * uy * (1 - c); R[5] = uy * uz * (1 - c) - ux * s;
                R[6] = uz * ux * (1 - c) - uy * s; R[7] = uz * uy * (1 - c) + ux * s; R[8] = c + uz * uz * (1 - c);
            }
}

template <typename T>
static int coord_upload(CoordState* st, const T** dst, const T* src, size_t n) {
    void* p = nullptr;
    GMM_CUDA(cudaMalloc(&p, (n ? n : 1) * sizeof(T)));
    st->owned.push_back(p);
    if (n) GMM_CUDA(cudaMemcpy(p, src, n * sizeof(T), cudaMemcpyHostToDevice));
    *dst = static_cast<const T*>(p);
    return GMM_OK;
}

int coordination_set(gmm_ctx* c, int n_sites, const gmm_coord_site* sites) {
    coord_release(c);
    if (n_sites == 0) return GMM_OK;
    CoordState* st = new CoordState();
    c->coordination = st;
    std::vector<CoordSite> dev(n_sites);
    for (int k = 0; k < n_sites; ++k) {
        const gmm_coord_site& h = sites[k];
        CoordSite& d = dev[k];
        d.n_atoms = h.n_atoms; d.metal = h.metal; d.n_vertices = h.n_vertices; d.min_atoms = h.min_atoms; d.flags = h.flags;
        d.radius = h.radius; d.distance = h.distance; d.weight = h.weight;
        for (int i = 0; i < (h.n_vertices + 1) * 3; ++i) d.tmpl[i] = h.template_xyz[i];
        coord_start_rotations(d.start);
        // atoms that can come within the radius of the metal: every mobile atom, and the static ones
        // near enough to a static metal (all of them when the metal itself moves)
        std::vector<int32_t> scan;
        const bool metal_static = h.mobile_row[h.metal] < 0;
        const double* m = h.xyz + 3 * (size_t)h.metal;
        for (int i = 0; i < h.n_atoms; ++i) {
            if (i == h.metal) continue;
            if (h.mobile_row[i] < 0 && metal_static) {
                const double* q = h.xyz + 3 * (size_t)i;
                const double d2 = (q[0] - m[0]) * (q[0] - m[0]) + (q[1] - m[1]) * (q[1] - m[1]) + (q[2] - m[2]) * (q[2] - m[2]);
                if (d2 > (h.radius + 1e-6) * (h.radius + 1e-6)) continue;
            }
            scan.push_back(i);
        }
        d.n_scan = (int)scan.size();
        int rc;
        if ((rc = coord_upload(st, &d.scan, scan.data(), scan.size()))) return rc;
        if ((rc = coord_upload(st, &d.row, h.mobile_row, (size_t)h.n_atoms))) return rc;
        if ((rc = coord_upload(st, &d.sxyz, h.xyz, (size_t)h.n_atoms * 3))) return rc;
        if ((rc = coord_upload(st, &d.cand, h.candidate, (size_t)h.n_atoms))) return rc;
        if ((rc = coord_upload(st, &d.geom, h.geometry, (size_t)h.n_atoms))) return rc;
        if ((rc = coord_upload(st, &d.brad, h.bond_radius, (size_t)h.n_atoms))) return rc;
        if ((rc = coord_upload(st, &d.nb_ptr, h.nb_ptr, (size_t)h.n_atoms + 1))) return rc;
        if ((rc = coord_upload(st, &d.nb_idx, h.nb_idx, (size_t)h.nb_ptr[h.n_atoms]))) return rc;
    }
    GMM_CUDA(cudaMalloc((void**)&st->d_sites, sizeof(CoordSite) * n_sites));
    GMM_CUDA(cudaMemcpy(st->d_sites, dev.data(), sizeof(CoordSite) * n_sites, cudaMemcpyHostToDevice));
    st->n_sites = n_sites;
    return GMM_OK;
}

int coordination_sites(const gmm_ctx* c) {
    const CoordState* st = static_cast<const CoordState*>(c->coordination);
    return st ? st->n_sites : 0;
}

int launch_coordination(gmm_ctx* c, int P, const float* xyz, double* out, cudaStream_t s) {
    CoordState* st = static_cast<CoordState*>(c->coordination);
    if (P == 0 || !st || st->n_sites == 0) return GMM_OK;
    const size_t rows = (size_t)P * st->n_sites;
    if (rows > st->work_rows) {          // grown on first use / when the population grows (synchronous)
        if (st->d_work) { GMM_CUDA(cudaDeviceSynchronize()); cudaFree(st->d_work); st->d_work = nullptr; st->work_rows = 0; }
        GMM_CUDA(cudaMalloc((void**)&st->d_work, rows * sizeof(CoordWork)));
        st->work_rows = rows;
    }
    coord_sphere_kernel<<<dim3((P + 3) / 4, st->n_sites), 128, 0, s>>>(P, c->M, xyz, st->d_sites, st->n_sites, c->centre[0],
                                                                    c->centre[1], c->centre[2], st->d_work, c->d_counters);
    GMM_LAUNCH_CHECK("coord_sphere_kernel");
    const size_t threads = rows * kCoordStarts;
    coord_cpd_kernel<<<(unsigned)((threads + 127) / 128), 128, 0, s>>>(P, st->d_sites, st->n_sites, st->d_work, out);
    GMM_LAUNCH_CHECK("coord_cpd_kernel");
    return GMM_OK;
}

}  // namespace gmm
