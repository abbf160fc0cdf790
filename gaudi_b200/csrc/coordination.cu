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

// Top eigenvector of a symmetric 4 x 4 matrix (cyclic Jacobi): q = (w, x, y, z)
__device__ void coord_top_eigenvector(double a[4][4], double* q) {
    double v[4][4] = {{1, 0, 0, 0}, {0, 1, 0, 0}, {0, 0, 1, 0}, {0, 0, 0, 1}};
    for (int sweep = 0; sweep < 30; ++sweep) {
        double off = 0.0, diag = 0.0;
        for (int i = 0; i < 4; ++i) { diag += a[i][i] * a[i][i]; for (int j = i + 1; j < 4; ++j) off += a[i][j] * a[i][j]; }
        if (off <= 1e-32 * diag || off == 0.0) break;
        for (int pi = 0; pi < 3; ++pi)
            for (int qi = pi + 1; qi < 4; ++qi) {
                const double apq = a[pi][qi];
                if (apq == 0.0) continue;
                const double theta = (a[qi][qi] - a[pi][pi]) / (2.0 * apq);
                const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                const double c = 1.0 / sqrt(t * t + 1.0), sn = t * c;
                for (int k = 0; k < 4; ++k) {
                    const double akp = a[k][pi], akq = a[k][qi];
                    a[k][pi] = c * akp - sn * akq; a[k][qi] = sn * akp + c * akq;
                }
                for (int k = 0; k < 4; ++k) {
                    const double apk = a[pi][k], aqk = a[qi][k];
                    a[pi][k] = c * apk - sn * aqk; a[qi][k] = sn * apk + c * aqk;
                }
                for (int k = 0; k < 4; ++k) {
                    const double vkp = v[k][pi], vkq = v[k][qi];
                    v[k][pi] = c * vkp - sn * vkq; v[k][qi] = sn * vkp + c * vkq;
                }
            }
    }
    int best = 0;
    for (int i = 1; i < 4; ++i) if (a[i][i] > a[best][best]) best = i;
    for (int k = 0; k < 4; ++k) q[k] = v[k][best];
}

__device__ __forceinline__ void rot_apply(const double* R, const double* y, double* o) {
    o[0] = R[0] * y[0] + R[1] * y[1] + R[2] * y[2];
    o[1] = R[3] * y[0] + R[4] * y[1] + R[5] * y[2];
    o[2] = R[6] * y[0] + R[7] * y[1] + R[8] * y[2];
}

// _cpdrift.py:156-157: sqrt of the mean over the axes of the variance over the points
__device__ double coord_spread(const double (*pts)[3], int n) {
    double tot = 0.0;
    for (int k = 0; k < 3; ++k) {
        double mean = 0.0, var = 0.0;
        for (int i = 0; i < n; ++i) mean += pts[i][k];
        mean /= n;
        for (int i = 0; i < n; ++i) var += (pts[i][k] - mean) * (pts[i][k] - mean);
        tot += var / n;
    }
    return sqrt(tot / 3.0);
}

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
        double X[kCoordMaxPoints][3], Y[kCoordMaxPoints][3], TY[kCoordMaxPoints][3], Pm[kCoordMaxPoints][kCoordMaxPoints];
        for (int i = 0; i < N; ++i) for (int k = 0; k < 3; ++k) X[i][k] = w.X[3 * i + k];
        for (int i = 0; i < Mv; ++i) for (int k = 0; k < 3; ++k) Y[i][k] = s.tmpl[3 * i + k];
        double R[9];
        for (int k = 0; k < 9; ++k) R[k] = s.start[9 * start + k];
        const double sx = coord_spread(X, N);
        for (int i = 0; i < Mv; ++i) rot_apply(R, Y[i], TY[i]);
        double sc = sx / coord_spread(TY, Mv);
        double t[3];
        for (int k = 0; k < 3; ++k) {
            double mx = 0.0, my = 0.0;
            for (int i = 0; i < N; ++i) mx += X[i][k];
            for (int i = 0; i < Mv; ++i) my += sc * TY[i][k];
            t[k] = mx / N - my / Mv;
        }
        double sigma2 = 0.0;
        for (int m = 0; m < Mv; ++m)
            for (int n = 0; n < N; ++n)
                for (int k = 0; k < 3; ++k) { const double d = sc * TY[m][k] + t[k] - X[n][k]; sigma2 += d * d; }
        sigma2 /= 3.0 * Mv * N;
        int done = 0;
        for (int it = 0; it < kCoordIters; ++it) {
            if (!(sigma2 > 0.0)) break;          // the step's arithmetic would be invalid (_cpdrift.py:79-86)
            for (int m = 0; m < Mv; ++m) { rot_apply(R, Y[m], TY[m]); for (int k = 0; k < 3; ++k) TY[m][k] = sc * TY[m][k] + t[k]; }
            bool valid = true;
            for (int n = 0; n < N; ++n) {
                double col = 0.0;
                for (int m = 0; m < Mv; ++m) {
                    const double dx = TY[m][0] - X[n][0], dy = TY[m][1] - X[n][1], dz = TY[m][2] - X[n][2];
                    Pm[m][n] = exp(-(dx * dx + dy * dy + dz * dz) / (2.0 * sigma2));
                    col += Pm[m][n];
                }
                if (col == 0.0 || !isfinite(col)) { valid = false; break; }
                for (int m = 0; m < Mv; ++m) Pm[m][n] /= col;
            }
            if (!valid) break;
            double Np = 0.0, p1[kCoordMaxPoints], pt1[kCoordMaxPoints], mux[3] = {0, 0, 0}, muy[3] = {0, 0, 0};
            for (int m = 0; m < Mv; ++m) { p1[m] = 0.0; for (int n = 0; n < N; ++n) p1[m] += Pm[m][n]; Np += p1[m]; }
            for (int n = 0; n < N; ++n) { pt1[n] = 0.0; for (int m = 0; m < Mv; ++m) pt1[n] += Pm[m][n]; }
            for (int k = 0; k < 3; ++k) {
                for (int n = 0; n < N; ++n) mux[k] += X[n][k] * pt1[n];
                for (int m = 0; m < Mv; ++m) muy[k] += Y[m][k] * p1[m];
                mux[k] /= Np; muy[k] /= Np;
            }
            double A[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
            for (int m = 0; m < Mv; ++m)
                for (int n = 0; n < N; ++n)
                    for (int a = 0; a < 3; ++a)
                        for (int b = 0; b < 3; ++b) A[a][b] += Pm[m][n] * (X[n][a] - mux[a]) * (Y[m][b] - muy[b]);
            // Horn: S = A^T (S_ab = sum w y_a x_b)
            const double Sxx = A[0][0], Sxy = A[1][0], Sxz = A[2][0], Syx = A[0][1], Syy = A[1][1], Syz = A[2][1],
                         Szx = A[0][2], Szy = A[1][2], Szz = A[2][2];
            double Nq[4][4] = {{Sxx + Syy + Szz, Syz - Szy, Szx - Sxz, Sxy - Syx},
                               {Syz - Szy, Sxx - Syy - Szz, Sxy + Syx, Szx + Sxz},
                               {Szx - Sxz, Sxy + Syx, -Sxx + Syy - Szz, Syz + Szy},
                               {Sxy - Syx, Szx + Sxz, Syz + Szy, -Sxx - Syy + Szz}};
            double q[4];
            coord_top_eigenvector(Nq, q);
            const double qn = sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
            const double qw = q[0] / qn, qx = q[1] / qn, qy = q[2] / qn, qz = q[3] / qn;
            R[0] = 1 - 2 * (qy * qy + qz * qz); R[1] = 2 * (qx * qy - qz * qw); R[2] = 2 * (qx * qz + qy * qw);
            R[3] = 2 * (qx * qy + qz * qw); R[4] = 1 - 2 * (qx * qx + qz * qz); R[5] = 2 * (qy * qz - qx * qw);
            R[6] = 2 * (qx * qz - qy * qw); R[7] = 2 * (qy * qz + qx * qw); R[8] = 1 - 2 * (qx * qx + qy * qy);
            double tr = 0.0, den = 0.0, xvar = 0.0;
            for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) tr += A[a][b] * R[3 * a + b];
            for (int m = 0; m < Mv; ++m)
                for (int k = 0; k < 3; ++k) den += p1[m] * (Y[m][k] - muy[k]) * (Y[m][k] - muy[k]);
            for (int n = 0; n < N; ++n)
                for (int k = 0; k < 3; ++k) xvar += pt1[n] * (X[n][k] - mux[k]) * (X[n][k] - mux[k]);
            sc = tr / den;
            double rm[3];
            rot_apply(R, muy, rm);
            for (int k = 0; k < 3; ++k) t[k] = mux[k] - sc * rm[k];
            sigma2 = (xvar - sc * tr) / (Np * 3.0);
            if (!isfinite(sc) || !isfinite(t[0]) || !isfinite(t[1]) || !isfinite(t[2])) { done = 0; break; }
            ++done;
        }
        if (done == 0) {
            ok = false;            // the reference fails in last() of an empty sequence: evaluate's penalty
        } else {
            double acc = 0.0;
            for (int m = 0; m < Mv; ++m) { rot_apply(R, Y[m], TY[m]); for (int k = 0; k < 3; ++k) TY[m][k] = sc * TY[m][k] + t[k]; }
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
        const int o = __shfl_sync(0xffffffffu, (int)ok, base + r);
        all_ok = all_ok && o;
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
