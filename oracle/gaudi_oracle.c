/*
 * CPU oracle in plain C (TEST INFRASTRUCTURE — not product code).
 *
 * A second, independent restatement of the reference's CPU path for the GaudiMM hot path
 * (genome expression -> Contacts objective -> similarity.rmsd), fp64 throughout, written so
 * that whole populations can be checked and timed on the host cores (OpenMP over individuals).
 * oracle/gaudi_oracle.py is the primary restatement and the one pinned to the reference's golden
 * values; tests/test_c_oracle.py checks this file against it (and against the same goldens).
 * Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may load it.
 *
 * Reference citations are paths relative to the reference repository root.
 *
 * Build (done by __graft_entry__.build() through oracle/Makefile):
 *     gcc -O2 [-fopenmp] -shared -fPIC -o oracle/libgaudi_oracle.so oracle/gaudi_oracle.c -lm
 * Without OpenMP, oracle/c_oracle.py runs slices of the population on Python threads (ctypes
 * releases the GIL).
 */
#include <math.h>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <stdlib.h>
#include <string.h>

#define F_HYDROPHOBIC 1
#define F_DONOR 2
#define F_ACCEPTOR 4
#define F_SEED 8
#define F_XFORM 16

/* ------------------------------------------------------------------------------------------
 * Genome expression
 * ---------------------------------------------------------------------------------------- */

/* chimera.BondRot.adjustAngle as used at gaudi/genes/torsion.py:161 and
 * gaudi/genes/rotamers.py:189: right-handed rotation of the moving side about the axis
 * fixed -> moving (sign pinned by tests/test_genes_torsion.py:54-63). */
static void rotate_about_bond(double* x, int fixed, int moving, const int* idx, int n_idx, double deg) {
    if (deg == 0.0 || n_idx == 0) return;
    const double* a = x + 3 * fixed;
    double b[3] = {x[3 * moving], x[3 * moving + 1], x[3 * moving + 2]};
    double u[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]};
    const double inv = 1.0 / sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
    u[0] *= inv; u[1] *= inv; u[2] *= inv;
    const double th = deg * (M_PI / 180.0), c = cos(th), s = sin(th);
    for (int k = 0; k < n_idx; ++k) {
        double* p = x + 3 * idx[k];
        const double v[3] = {p[0] - b[0], p[1] - b[1], p[2] - b[2]};
        const double cr[3] = {u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0]};
        const double dot = u[0] * v[0] + u[1] * v[1] + u[2] * v[2];
        for (int d = 0; d < 3; ++d) p[d] = b[d] + v[d] * c + cr[d] * s + u[d] * dot * (1.0 - c);
    }
}

/* One individual.  Bonds are applied in the given order on current coordinates
 * (torsion.py:153-164); tors_src >= 0 reads the torsion allele (with the C2/N2 snap of
 * torsion.py:159-160), tors_src < 0 encodes -(1 + 4 * residue + chi) and reads the rotamer table
 * row int(allele * n_rows) (rotamers.py:141-152) holding chi - chi0 (rotamers.py:186-189).
 * Search: x' = R (x - origin) + t with origin = current coordinate of the target atom when
 * Torsion precedes Search (search.py:186-214, base.py:157-168). */
void go_express(int M, const double* xyz0, int T, const int* fixed, const int* moving, const int* mask_ptr,
                const int* mask_idx, const unsigned char* snap, const int* tors_src, const double* tors_allele,
                const double* xform, int origin_atom, int torsion_before_search, const unsigned char* flags,
                const int* rot_ptr, const double* rot_delta, const double* rot_allele, double* out) {
    memcpy(out, xyz0, sizeof(double) * 3 * M);
    double origin_pre[3] = {xyz0[3 * origin_atom], xyz0[3 * origin_atom + 1], xyz0[3 * origin_atom + 2]};
    for (int t = 0; t < T; ++t) {
        double alpha;
        if (tors_src[t] >= 0) {
            alpha = tors_allele[tors_src[t]];
            if (snap[t]) alpha = (alpha <= 0.0) ? 0.0 : 180.0;
        } else {
            const int code = -(tors_src[t] + 1), res = code >> 2, chi = code & 3;
            const int r0 = rot_ptr[res], n = rot_ptr[res + 1] - r0;
            int k = (int)(rot_allele[res] * (double)n);
            if (k > n - 1) k = n - 1;
            if (k < 0) k = 0;
            alpha = rot_delta[4 * (r0 + k) + chi];
        }
        rotate_about_bond(out, fixed[t], moving[t], mask_idx + mask_ptr[t], mask_ptr[t + 1] - mask_ptr[t], alpha);
    }
    if (xform) {
        double o[3];
        for (int d = 0; d < 3; ++d) o[d] = torsion_before_search ? out[3 * origin_atom + d] : origin_pre[d];
        for (int i = 0; i < M; ++i) {
            if (!(flags[i] & F_XFORM)) continue;
            const double v[3] = {out[3 * i] - o[0], out[3 * i + 1] - o[1], out[3 * i + 2] - o[2]};
            for (int r = 0; r < 3; ++r)
                out[3 * i + r] = xform[4 * r] * v[0] + xform[4 * r + 1] * v[1] + xform[4 * r + 2] * v[2] + xform[4 * r + 3];
        }
    }
}

/* ------------------------------------------------------------------------------------------
 * Contacts objective
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    double zone_radius, clash_thr, hydro_thr, hb_allow, amv, cutoff, weight;
    int only_internal;
} go_params;

/* uniform grid over the static atoms (built once per receptor) */
typedef struct {
    int n, nx, ny, nz;
    double lo[3], cell;
    int* start;   /* [ncells + 1] */
    int* items;   /* [n] static atom indices sorted by cell */
} go_grid;

go_grid* go_grid_build(int n, const double* xyz, double cell) {
    go_grid* g = (go_grid*)calloc(1, sizeof(go_grid));
    g->n = n; g->cell = cell;
    double hi[3] = {-1e300, -1e300, -1e300};
    g->lo[0] = g->lo[1] = g->lo[2] = 1e300;
    for (int i = 0; i < n; ++i)
        for (int d = 0; d < 3; ++d) {
            if (xyz[3 * i + d] < g->lo[d]) g->lo[d] = xyz[3 * i + d];
            if (xyz[3 * i + d] > hi[d]) hi[d] = xyz[3 * i + d];
        }
    if (n == 0) { g->lo[0] = g->lo[1] = g->lo[2] = 0; hi[0] = hi[1] = hi[2] = 0; }
    g->nx = (int)((hi[0] - g->lo[0]) / cell) + 1;
    g->ny = (int)((hi[1] - g->lo[1]) / cell) + 1;
    g->nz = (int)((hi[2] - g->lo[2]) / cell) + 1;
    const int nc = g->nx * g->ny * g->nz;
    g->start = (int*)calloc(nc + 1, sizeof(int));
    g->items = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    int* cid = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    for (int i = 0; i < n; ++i) {
        const int cx = (int)((xyz[3 * i] - g->lo[0]) / cell), cy = (int)((xyz[3 * i + 1] - g->lo[1]) / cell),
                  cz = (int)((xyz[3 * i + 2] - g->lo[2]) / cell);
        cid[i] = (cx * g->ny + cy) * g->nz + cz;
        g->start[cid[i] + 1]++;
    }
    for (int c = 0; c < nc; ++c) g->start[c + 1] += g->start[c];
    int* fill = (int*)calloc(nc, sizeof(int));
    for (int i = 0; i < n; ++i) g->items[g->start[cid[i]] + fill[cid[i]]++] = i;
    free(fill); free(cid);
    return g;
}

void go_grid_free(go_grid* g) {
    if (!g) return;
    free(g->start); free(g->items); free(g);
}

static int excluded(const int* excl_ptr, const int* excl_idx, int a, int b) {
    int lo = excl_ptr[a], hi = excl_ptr[a + 1] - 1;      /* sorted: binary search */
    while (lo <= hi) {
        const int mid = (lo + hi) >> 1;
        if (excl_idx[mid] == b) return 1;
        if (excl_idx[mid] < b) lo = mid + 1; else hi = mid - 1;
    }
    return 0;
}

/* gaudi/objectives/contacts.py:215-241 */
static double lennard_jones(double r1, double r2, double overlap) {
    const double r0 = r1 + r2, distance = r0 - overlap;
    const double x = pow(r0 / distance, 6.0);
    return x * x - 2 * x;
}

/* gaudi/objectives/contacts.py:243-268 */
static double vdw_vol_overlap(double r1, double r2, double overlap) {
    const double PI = 3.14159265359;
    const double d = r1 + r2 - overlap;
    if (d == 0) return 1000;
    double ha = 0, hb = 0;
    if (d < r1 + r2) {
        ha = (r2 * r2 - (d - r1) * (d - r1)) / (2 * d);
        hb = (r1 * r1 - (d - r2) * (d - r2)) / (2 * d);
    }
    return (PI / 3) * (ha * ha * (3 * r1 - ha) + hb * hb * (3 * r2 - hb));
}

/* One unordered pair through DetectClash's rule (contacts.py:142-145) for both
 * `which` modes and through _analyze_interactions (contacts.py:148-188). */
static void pair(const double* xa, const double* xb, double ra, double rb, unsigned fa, unsigned fb,
                 const go_params* p, double* clash, double* hydro, int* nc, int* nh) {
    const double dx = xa[0] - xb[0], dy = xa[1] - xb[1], dz = xa[2] - xb[2];
    const double d = sqrt(dx * dx + dy * dy + dz * dz);
    const int hb = ((fa & F_DONOR) && (fb & F_ACCEPTOR)) || ((fb & F_DONOR) && (fa & F_ACCEPTOR));
    const double ov = ra + rb - d - ((hb && p->hb_allow != 0.0) ? p->hb_allow : 0.0);
    const double maxr = ra > rb ? ra : rb;
    if (d <= maxr + p->amv - p->hydro_thr && ov >= p->hydro_thr && ov <= p->clash_thr &&
        (fa & F_HYDROPHOBIC) && (fb & F_HYDROPHOBIC)) {
        *hydro += lennard_jones(ra, rb, ov);
        ++*nh;
    }
    if (d <= maxr + p->amv - p->clash_thr && ov >= p->clash_thr && ov > p->clash_thr) {
        *clash += fabs(vdw_vol_overlap(ra, rb, ov));
        ++*nc;
    }
}

/* Contacts.evaluate_clashes / evaluate_hydrophobic (contacts.py:129-138) of one individual.
 * Atoms: n_mobile mobile ones (coordinates `mob`) then n_static static ones (`sta`, with grid).
 * excl: CSR over all atoms of the atoms within <= bond_separation bonds (sorted, global ids). */
void go_contacts(int n_mobile, const double* mob, int n_static, const double* sta, const go_grid* grid,
                 const double* radius, const unsigned char* flags, const int* excl_ptr, const int* excl_idx,
                 const go_params* p, double* scores, int* counts) {
    const int n = n_mobile + n_static;
    unsigned char* inz = (unsigned char*)calloc(n > 0 ? n : 1, 1);
    const double r2 = p->zone_radius * p->zone_radius;
    /* zone (contacts.py:190-213): seeds, plus everything within `radius` of a seed */
    for (int i = 0; i < n; ++i)
        if (flags[i] & F_SEED) inz[i] = 1;
    if (!p->only_internal) {
        for (int s = 0; s < n; ++s) {
            if (!(flags[s] & F_SEED)) continue;
            const double* xs = s < n_mobile ? mob + 3 * s : sta + 3 * (s - n_mobile);
            for (int i = 0; i < n_mobile; ++i) {
                const double dx = mob[3 * i] - xs[0], dy = mob[3 * i + 1] - xs[1], dz = mob[3 * i + 2] - xs[2];
                if (dx * dx + dy * dy + dz * dz <= r2) inz[i] = 1;
            }
            if (n_static == 0) continue;
            const int span = (int)(p->zone_radius / grid->cell) + 1;
            const int cx = (int)floor((xs[0] - grid->lo[0]) / grid->cell), cy = (int)floor((xs[1] - grid->lo[1]) / grid->cell),
                      cz = (int)floor((xs[2] - grid->lo[2]) / grid->cell);
            for (int ix = cx - span; ix <= cx + span; ++ix) {
                if (ix < 0 || ix >= grid->nx) continue;
                for (int iy = cy - span; iy <= cy + span; ++iy) {
                    if (iy < 0 || iy >= grid->ny) continue;
                    for (int iz = cz - span; iz <= cz + span; ++iz) {
                        if (iz < 0 || iz >= grid->nz) continue;
                        const int c = (ix * grid->ny + iy) * grid->nz + iz;
                        for (int e = grid->start[c]; e < grid->start[c + 1]; ++e) {
                            const int j = grid->items[e];
                            const double dx = sta[3 * j] - xs[0], dy = sta[3 * j + 1] - xs[1], dz = sta[3 * j + 2] - xs[2];
                            if (dx * dx + dy * dy + dz * dz <= r2) inz[n_mobile + j] = 1;
                        }
                    }
                }
            }
        }
    }
    double clash = 0, hydro = 0;
    int nc = 0, nh = 0;
    double rmax = 0;
    for (int i = 0; i < n; ++i) if (radius[i] > rmax) rmax = radius[i];
    const double thr_min = p->hydro_thr < p->clash_thr ? p->hydro_thr : p->clash_thr;
    const double reach = 2 * rmax - thr_min + 1e-9;       /* no pair beyond this can pass */
    /* mobile - mobile */
    for (int a = 0; a < n_mobile; ++a) {
        if (!inz[a]) continue;
        for (int b = a + 1; b < n_mobile; ++b) {
            if (!inz[b] || excluded(excl_ptr, excl_idx, a, b)) continue;
            pair(mob + 3 * a, mob + 3 * b, radius[a], radius[b], flags[a], flags[b], p, &clash, &hydro, &nc, &nh);
        }
    }
    if (n_static > 0) {
        const int span = (int)(reach / grid->cell) + 1;
        /* every zone atom a against static zone atoms b (b > a in global numbering) */
        for (int a = 0; a < n; ++a) {
            if (!inz[a]) continue;
            const double* xa = a < n_mobile ? mob + 3 * a : sta + 3 * (a - n_mobile);
            const int cx = (int)floor((xa[0] - grid->lo[0]) / grid->cell), cy = (int)floor((xa[1] - grid->lo[1]) / grid->cell),
                      cz = (int)floor((xa[2] - grid->lo[2]) / grid->cell);
            for (int ix = cx - span; ix <= cx + span; ++ix) {
                if (ix < 0 || ix >= grid->nx) continue;
                for (int iy = cy - span; iy <= cy + span; ++iy) {
                    if (iy < 0 || iy >= grid->ny) continue;
                    for (int iz = cz - span; iz <= cz + span; ++iz) {
                        if (iz < 0 || iz >= grid->nz) continue;
                        const int c = (ix * grid->ny + iy) * grid->nz + iz;
                        for (int e = grid->start[c]; e < grid->start[c + 1]; ++e) {
                            const int b = n_mobile + grid->items[e];
                            if (b <= a || !inz[b]) continue;
                            const double* xb = sta + 3 * (b - n_mobile);
                            const double dx = xa[0] - xb[0], dy = xa[1] - xb[1], dz = xa[2] - xb[2];
                            if (dx * dx + dy * dy + dz * dz > reach * reach) continue;
                            if (excluded(excl_ptr, excl_idx, a, b)) continue;
                            pair(xa, xb, radius[a], radius[b], flags[a], flags[b], p, &clash, &hydro, &nc, &nh);
                        }
                    }
                }
            }
        }
    }
    free(inz);
    clash *= 2.0;            /* the reference sees every pair from both atoms (contacts.py:176-177) */
    hydro *= 2.0;
    if (p->cutoff != 0.0 && clash > p->cutoff) clash = -1000 * p->weight;   /* contacts.py:132-133 */
    scores[0] = clash; scores[1] = hydro;
    if (counts) { counts[0] = nc; counts[1] = nh; }
}

/* Environment.evaluate (base.py:353-364) for a population: express + Contacts per individual,
 * OpenMP over individuals.  out_xyz may be NULL. */
void go_evaluate_population(int P, int M, const double* xyz0, int T, const int* fixed, const int* moving,
                            const int* mask_ptr, const int* mask_idx, const unsigned char* snap, const int* tors_src,
                            int n_allele, const double* tors, const double* xforms, int origin_atom,
                            int torsion_before_search, int n_res, const int* rot_ptr, const double* rot_delta,
                            const double* rot, int n_static, const double* sta, const go_grid* grid,
                            const double* radius, const unsigned char* flags, const int* excl_ptr, const int* excl_idx,
                            const go_params* prm, double* scores, int* counts, double* out_xyz) {
#pragma omp parallel
    {
        double* x = (double*)malloc(sizeof(double) * 3 * (M > 0 ? M : 1));
#pragma omp for schedule(dynamic, 8)
        for (int p = 0; p < P; ++p) {
            go_express(M, xyz0, T, fixed, moving, mask_ptr, mask_idx, snap, tors_src, tors + (size_t)p * n_allele,
                       xforms ? xforms + (size_t)p * 12 : NULL, origin_atom, torsion_before_search, flags, rot_ptr,
                       rot_delta, rot ? rot + (size_t)p * n_res : NULL, x);
            if (out_xyz) memcpy(out_xyz + (size_t)p * 3 * M, x, sizeof(double) * 3 * M);
            go_contacts(M, x, n_static, sta, grid, radius, flags, excl_ptr, excl_idx, prm, scores + 2 * (size_t)p,
                        counts ? counts + 2 * (size_t)p : NULL);
        }
        free(x);
    }
}

/* gaudi/similarity.py:73-75 for rows [row0, row1) of the all-pairs matrix (no superposition) */
void go_rmsd_matrix(int N, int A, const double* xyz, int row0, int row1, double* out) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int i = row0; i < row1; ++i) {
        const double* a = xyz + (size_t)i * 3 * A;
        for (int j = 0; j < N; ++j) {
            const double* b = xyz + (size_t)j * 3 * A;
            double s = 0;
            for (int k = 0; k < 3 * A; ++k) { const double d = a[k] - b[k]; s += d * d; }
            out[(size_t)(i - row0) * N + j] = sqrt(s / A);
        }
    }
}

int go_num_threads(void) {
    int n = 1;
#ifdef _OPENMP
#pragma omp parallel
    {
#pragma omp single
        n = omp_get_num_threads();
    }
#endif
    return n;
}
