/*
 * gaudimm_b200.h — C ABI of libgaudimm_b200.so
 *
 * Batched (population-wide) replacement for the data-parallel hot path of GaudiMM
 * (insilichem/gaudi): genome expression (Torsion, Rotamers, Search genes) -> Contacts
 * objective -> pairwise similarity.rmsd.  The reference evaluates one individual at a
 * time through UCSF Chimera; this library evaluates a whole population tensor per call on
 * one B200.  Reference citations are paths relative to the reference repository root.
 *
 * Conventions
 *   - every function returns 0 on success and a negative GMM_E* code on failure; the
 *     message is available from gmm_last_error() (thread local).  Nothing throws.
 *   - pointers suffixed _h are HOST pointers, read during the call and not retained;
 *     pointers suffixed _d are DEVICE pointers on the context's device (for a torch
 *     tensor: tensor.data_ptr() of a contiguous CUDA tensor).  The library never frees or
 *     reallocates caller memory.
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Work is
 *     enqueued on it; the batched calls do not synchronise, except that the first call
 *     with a larger population than any before it grows the context's scratch buffer
 *     (one cudaMalloc); call gmm_reserve() up front to avoid that.
 *   - device coordinates are float32 in a frame re-centred at `centre` given to
 *     gmm_set_receptor(): x_dev = (float)(x_world - centre).
 *   - one context serves one (mobile model, receptor, parameter set); contexts are
 *     independent.  Calls on one context must not overlap in time unless they are on the
 *     same stream (they share the scratch buffer).
 */
#ifndef GAUDIMM_B200_H
#define GAUDIMM_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GMM_OK            0
#define GMM_EINVAL       (-1)   /* bad argument */
#define GMM_ECUDA        (-2)   /* CUDA runtime error (see gmm_last_error) */
#define GMM_ESTATE       (-3)   /* call order: model/receptor/params not set */
#define GMM_ELIMIT       (-4)   /* size beyond a documented limit */

/* per-atom flag bits (uint8) */
#define GMM_F_HYDROPHOBIC  1    /* element in Contacts.hydrophobic_elements (contacts.py:103) */
#define GMM_F_DONOR        2    /* H-bond donor   (DetectClash hbondAllowance, contacts.py:144) */
#define GMM_F_ACCEPTOR     4    /* H-bond acceptor */
#define GMM_F_SEED         8    /* zone seed: probe atom / CB of a Rotamers residue (contacts.py:195-207) */
#define GMM_F_XFORM       16    /* mobile atom moved by the Search xform (search.py:201-214) */

/* gmm_set_params option bits */
#define GMM_OPT_ONLY_INTERNAL  1  /* Contacts(only_internal=True), contacts.py:199 */
#define GMM_OPT_NO_CULL        2  /* score every receptor tile (dense), no bounding-box culling */
#define GMM_OPT_TILED          4  /* use the two-kernel tiled Contacts path instead of the fused one */
#define GMM_OPT_RMSD_SIMT      16  /* bitmask-only gmm_rmsd_matrix calls: use the FP32 direct-difference kernel
                                    * instead of the tensor-core pre-filter + exact re-decision (same bits) */
#define GMM_OPT_RMSD_HMMA      32  /* bitmask-only gmm_rmsd_matrix calls: run the product of the pre-filter with
                                    * mma.sync (the first tensor-core version) instead of tcgen05 + TMEM + TMA;
                                    * same bits, kept as a cross-check */
#define GMM_OPT_SEQUENTIAL_EXPRESS 8 /* apply the bond rotations one after another on the current coordinates
                                     * (the reference's own order of operations) even when the model allows the
                                     * faster kinematic-tree form; results agree to fp64 rounding */

#define GMM_MAX_PEERS          16    /* ranks (GPUs of one node) in a fitness exchange */
#define GMM_IPC_HANDLE_BYTES   64    /* sizeof(cudaIpcMemHandle_t) */
#define GMM_MAX_MOBILE_ATOMS   1024
#define GMM_MAX_TORSIONS       1024

typedef struct gmm_ctx gmm_ctx;

/* Library / context -------------------------------------------------------------- */
int         gmm_create(gmm_ctx** out, int device);
int         gmm_destroy(gmm_ctx* ctx);
const char* gmm_last_error(void);
int         gmm_version(void);
/* Pre-size the context's scratch for populations up to max_pop (optional). */
int         gmm_reserve(gmm_ctx* ctx, int max_pop);

/* Pre-size the workspace of bitmask-only gmm_rmsd_matrix calls (re-centred fp16 copies of the
 * conformers and their norms) for up to max_n conformers of `atoms` atoms (optional).  Without it
 * the first such call at a larger size synchronises the device and allocates, like the scratch. */
int         gmm_reserve_rmsd(gmm_ctx* ctx, int max_n, int atoms);

/* Static model (host pointers, uploaded once) ------------------------------------ */

/* The atoms whose coordinates depend on the genome.  Replaces the live chimera.Molecule
 * plus the BondRot list of gaudi/genes/torsion.py:215-307 and the per-residue chi BondRots
 * of gaudi/genes/rotamers.py:186-210.
 *   xyz0_h      [M,3]  template coordinates (input-file frame), fp64
 *   tors_*      T rotatable bonds in the reference's bond order (torsion.py:246-248);
 *               fixed = atom on the anchor side (torsion.py:305), moving = the other atom
 *   tors_mask_h [T, ceil(M/64)]  atoms on the moving side (bit i of word i/64)
 *   tors_snap_h [T]    1 = both atoms C2/N2: allele <= 0 -> 0 deg else 180 (torsion.py:159-160)
 *   tors_src_h  [T]    >= 0: column of torsion_deg_d driving this bond;
 *                      <  0: rotamer driven, residue r and chi k with -(1 + 4 r + k)
 *   excl_mask_h [M, ceil(M/64)]  atoms within <= bond_separation bonds (DetectClash exclusion)
 *   origin_atom        Search target atom (search.py:186-188)
 *   torsion_before_search  1 if Torsion is listed before Search in the YAML (base.py:157-168)
 */
int gmm_set_mobile(gmm_ctx* ctx, int M,
                   const double* xyz0_h, const float* radius_h, const uint8_t* flags_h,
                   int T, const int32_t* tors_fixed_h, const int32_t* tors_moving_h,
                   const uint64_t* tors_mask_h, const uint8_t* tors_snap_h,
                   const int32_t* tors_src_h, const uint64_t* excl_mask_h,
                   int origin_atom, int torsion_before_search);

/* Rotamer chi tables (gaudi/genes/rotamers.py:141-152,165-189): residue r owns rows
 * rot_ptr_h[r] .. rot_ptr_h[r+1] of rot_delta_h [rows,4] = (chi - chi0) in degrees, sorted
 * as the reference sorts its library.  Allele a in [0,1) selects row int(a * n_rows). */
int gmm_set_rotamers(gmm_ctx* ctx, int n_res, const int32_t* rot_ptr_h, const double* rot_delta_h);

/* Every atom with fixed coordinates (all other atoms of all Molecule genes).  Replaces the
 * chimera.specifier.zone / DetectClash spatial searches of contacts.py:140-146,190-213.
 *   rr_ptr_h/rr_idx_h  CSR over static atoms a of partners b > a that are more than
 *                      bond_separation bonds apart and close enough to ever be reported;
 *                      their score contributions are computed on the device.
 *   ms_ptr_h/ms_idx_h  CSR over MOBILE atoms of static atoms within <= bond_separation
 *                      bonds (only with flexible side chains; may be NULL).
 *   centre_h[3]        origin of the fp32 device frame (e.g. the Search centre).
 * Requires gmm_set_params() first (the static pair values depend on the thresholds). */
int gmm_set_receptor(gmm_ctx* ctx, int R,
                     const double* xyz_h, const float* radius_h, const uint8_t* flags_h,
                     const int32_t* rr_ptr_h, const int32_t* rr_idx_h,
                     const int32_t* ms_ptr_h, const int32_t* ms_idx_h,
                     const double* centre_h);

/* Contacts objective settings (contacts.py:101-121) and DetectClash's fixed options
 * (contacts.py:142-145).  `cutoff`/`weight`: evaluate_clashes' penalty (contacts.py:132-133). */
int gmm_set_params(gmm_ctx* ctx, double zone_radius, double clash_threshold,
                   double hydrophobic_threshold, double hbond_allowance,
                   double assumed_max_vdw, double cutoff, double weight, int options);

/* Batched hot path (device pointers) --------------------------------------------- */

/* Individual.express() for P individuals at once: Rotamers.express (rotamers.py:141-152),
 * Torsion.express (torsion.py:153-164), Search.express (search.py:201-214) and the
 * coordinate read-back of MolecularIndividual.post_express (base.py:301-303).
 *   torsion_deg_d [P, n_allele_torsions] fp64 degrees (may be NULL if none)
 *   xform_d       [P, 12] fp64 row-major 3x4 product T.R of the Search allele (NULL = identity)
 *   rot_allele_d  [P, n_res] fp64 in [0,1) (NULL if no rotamers)
 *   out_xyz_d     [P, M, 3] fp32, device frame */
int gmm_express(gmm_ctx* ctx, int P, const double* torsion_deg_d, const double* xform_d,
                const double* rot_allele_d, float* out_xyz_d, void* stream);

/* Contacts.evaluate_clashes and Contacts.evaluate_hydrophobic (contacts.py:129-138) for P
 * expressed individuals.
 *   xyz_d        [P, M, 3] fp32 device-frame coordinates of the mobile atoms
 *   out_scores_d [P, 2] fp64 = (clashes, hydrophobic), before the caller's round()
 *   out_counts_d [P, 2] int32 = unique (clash, hydrophobic) pairs counted, or NULL
 * A pose whose staging copy has not landed after two seconds (never seen; the wait is bounded so
 * that a fault cannot hang the GPU) gets NaN scores and counts of -1 instead of stopping the context. */
int gmm_contacts(gmm_ctx* ctx, int P, const float* xyz_d, double* out_scores_d,
                 int32_t* out_counts_d, void* stream);

/* Environment.evaluate (base.py:353-364) for the whole population: express + Contacts.
 * out_xyz_d may be NULL (coordinates then live only in the context's scratch). */
int gmm_evaluate(gmm_ctx* ctx, int P, const double* torsion_deg_d, const double* xform_d,
                 const double* rot_allele_d, float* out_xyz_d, double* out_scores_d,
                 int32_t* out_counts_d, void* stream);

/* Same as gmm_evaluate with HOST buffers: copies alleles in, runs, copies scores (and
 * coordinates if out_xyz_h != NULL) back, and synchronises the stream before returning.
 * This is what toolbox.map(toolbox.evaluate, population) (algorithms.py:99-102) binds to. */
int gmm_evaluate_host(gmm_ctx* ctx, int P, const double* torsion_deg_h, const double* xform_h,
                      const double* rot_allele_h, float* out_xyz_h, double* out_scores_h,
                      int32_t* out_counts_h, void* stream);

/* gmm_evaluate for callers that may not allocate device memory (no counterpart in the
 * reference, which runs on the CPU and has no device memory to manage).  Same inputs and
 * outputs as gmm_evaluate without out_xyz_d, but the scratch the evaluation needs (the
 * expressed coordinates, 12 bytes per mobile atom per individual; with GMM_OPT_TILED also a
 * zone bitmask and partial sums) is a 16 MiB array built into the library instead of the
 * context's cudaMalloc'ed buffer: the call never calls cudaMalloc/cudaFree.
 *   - populations whose scratch need exceeds 16 MiB are refused with GMM_ELIMIT (the message
 *     states both sizes); split them into smaller batches;
 *   - the first call on each device looks the array's device address up (one-time set-up, guarded
 *     by a lock so that host threads may race to it); with CUDA's lazy module loading that is also
 *     the moment the driver places the 16 MiB array on that GPU, so free device memory drops by
 *     about that much once, there, and never afterwards; make that first call before starting
 *     any stream capture;
 *   - all calls, from every context on that device, share the one array: calls that overlap on
 *     the GPU (different streams) race with each other; on one stream they are safe.  Nothing
 *     in it has to be zero on entry (every word is written before it is read). */
int gmm_evaluate_static(gmm_ctx* ctx, int P, const double* torsion_deg_d, const double* xform_d,
                        const double* rot_allele_d, double* out_scores_d, int32_t* out_counts_d,
                        void* stream);

/* gaudi.similarity.rmsd / _rmsd_squared (similarity.py:36-75) for rows [row0,row1) of the
 * all-pairs matrix of N conformers x A atoms (no superposition).
 *   xyz_d            [N, A, 3] fp32 (same frame for all conformers)
 *   out_rmsd_d       [row1-row0, N] fp32 RMSD values, or NULL
 *   out_similar_d    [row1-row0, ceil(N/32)] uint32 bitmask of NOT(rmsd^2 > thr^2), or NULL */
int gmm_rmsd_matrix(gmm_ctx* ctx, int N, int A, const float* xyz_d, int row0, int row1,
                    float* out_rmsd_d, uint32_t* out_similar_d, double threshold, void* stream);

/* Multi-GPU: fitness exchange between the GPUs of one node, one process per GPU ------------------
 * The reference's seam is toolbox.map(toolbox.evaluate, invalid_ind) (gaudi/algorithms.py:99-102,
 * 133-137; gaudi/cli/gaudi_run.py:97-99, where a multiprocessing pool was once meant to go): the
 * population is split by individual, every rank scores its block with gmm_evaluate, and the rows
 * are gathered so that the rank running selection (gaudi/algorithms.py:160) holds them all.
 *   gmm_exchange_create   allocates this rank's block (room for max_rows = the whole population,
 *                         twice: two buffers alternate) and returns its CUDA IPC handle;
 *   gmm_exchange_connect  maps the other ranks' blocks: handles_h = the world handles in rank
 *                         order, exchanged by the caller (any transport: torch.distributed, MPI,
 *                         a file); ranks may share one GPU;
 *   gmm_exchange_publish  enqueues ONE kernel on `stream` that stores this rank's rows
 *                         [row0, row0 + n_rows) into every rank's block with plain stores through the
 *                         peer mappings, publishes the epoch and waits (on the GPU, at most 10 s)
 *                         for every other rank's publish of the same epoch.  No host
 *                         synchronisation, no allocation: capturable in a CUDA graph.  Every rank
 *                         must publish the same number of times.  *gathered_*_out: where the rows
 *                         of THIS call will be once the kernel has run ([max_rows, 2] fp64 scores,
 *                         [max_rows, 2] int32 counts); valid until this rank's publish after next;
 *   gmm_exchange_state    synchronises the device; returns the epochs completed, whether a wait
 *                         ever timed out, and the buffers holding the LAST completed epoch (use it
 *                         after replaying publishes from a CUDA graph, which the host-side epoch
 *                         count behind gmm_exchange_publish's out pointers does not see). */
int gmm_exchange_create(gmm_ctx* ctx, int world, int rank, int max_rows, unsigned char* handle_out /*[64]*/);
int gmm_exchange_connect(gmm_ctx* ctx, const unsigned char* handles_h /*[world, 64]*/);
int gmm_exchange_publish(gmm_ctx* ctx, const double* scores_d, const int32_t* counts_d, int row0, int n_rows,
                         double** gathered_scores_out, int32_t** gathered_counts_out, void* stream);
/* gmm_evaluate followed by gmm_exchange_publish of its rows, as one call: with the one-kernel
 * Contacts path every individual's row is stored into every rank's table by the Contacts kernel
 * itself the moment it is finished (the transfer hides behind the scoring of the others), and only
 * the epoch flags and the wait remain behind it.  Same outputs as the two calls. */
int gmm_evaluate_publish(gmm_ctx* ctx, int P, const double* torsion_deg_d, const double* xform_d,
                         const double* rot_allele_d, float* out_xyz_d, double* out_scores_d, int32_t* out_counts_d,
                         int row0, double** gathered_scores_out, int32_t** gathered_counts_out, void* stream);
int gmm_exchange_state(gmm_ctx* ctx, unsigned* epoch_out, int* error_out, double** gathered_scores_out,
                       int32_t** gathered_counts_out);
int gmm_exchange_destroy(gmm_ctx* ctx);

/* 1 when gmm_express / gmm_evaluate use the kinematic-tree form for this model (moving sets nested
 * or disjoint, verified numerically in gmm_set_mobile), 0 when they apply the bond rotations one
 * after another, negative on error.  Informational: both forms give the reference's coordinates. */
int gmm_express_is_tree(gmm_ctx* ctx);

/* Partner lists of the similarity bitmask gmm_rmsd_matrix wrote, for the niching pass of
 * gaudi/algorithms.py:139-155 (which asks ind.similar(other) pair by pair).  bits_d holds rows
 * [row0,row1) of the N x N matrix, ceil(N/32) uint32 words per row.  Two passes:
 *   count pass  counts_d != NULL: counts_d[r] = set bits of row r, its own column excluded;
 *   fill pass   ptr_d + idx_d   : idx_d[ptr_d[r] ...] = partner columns of row r, ascending;
 * ptr_d is the exclusive prefix sum of the counts (int64, rows + 1 entries), made by the caller.
 * idx_capacity > 0: idx_d holds that many entries and entries beyond it are dropped (a caller that
 * sized idx_d before knowing the counts compares them with its capacity afterwards); <= 0: no bound. */
int gmm_similar_lists(gmm_ctx* ctx, int N, int row0, int row1, const uint32_t* bits_d, int32_t* counts_d,
                      const int64_t* ptr_d, int32_t* idx_d, long long idx_capacity, void* stream);

/* Non-dominated sorting for the selection step (gaudi/algorithms.py:160 -> DEAP selNSGA2 ->
 * sortNondominated).  w_sorted_d: [n, m] fp64 weighted fitness values (larger is better in every
 * column, DEAP's wvalues), rows sorted lexicographically in DESCENDING order (checked; GMM_EINVAL
 * otherwise), m <= 8.  front_d[i] = index of the front row i belongs to (0 = non-dominated);
 * equal rows share a front.  Synchronises the stream once (the order check). */
int gmm_nondominated_fronts(gmm_ctx* ctx, int n, int m, const double* w_sorted_d, int32_t* front_d, void* stream);

/* HOST function: gmm_nondominated_fronts for the common two-objective job, O(N log N) on the host.
 * w_sorted_h [N, 2] fp64 weighted values, rows sorted by column 0 descending, ties by column 1
 * descending (checked); front_h [N] int32 out.  Equal rows share a front. */
int gmm_fronts2_host(int N, const double* w_sorted_h, int32_t* front_h);

/* HOST function: the greedy pass of gaudi/algorithms.py:139-155 over HOST partner lists
 * (ptr_h int64 [N+1], idx_h int32) and the candidates' lexicographic fitness ranks (rank_h,
 * larger = fitter, as ``i.fitness <= n_i.fitness`` compares wvalues): visit candidates in order;
 * drop one when a kept partner has rank >= its own, else evict its kept partners and keep it.
 * kept_h [N] uint8 out. */
int gmm_niching_host(int N, const int64_t* ptr_h, const int32_t* idx_h, const int64_t* rank_h, uint8_t* kept_h);

/* Geometric restraint objectives on expressed coordinates ("next" row of the scope: Distance,
 * gaudi/objectives/distance.py:102-129, and Angle, gaudi/objectives/angle.py:86-96).  A term is a
 * distance (kind 0, points 0-1), an angle at point 1 (kind 1, points 0-2) or an IUPAC dihedral
 * (kind 2, points 0-3); a point is a mobile atom (atoms_h >= 0) or fixed world coordinates
 * (atoms_h == -1, points_h).  Terms with the same obj_h are averaged into one output column:
 * distance terms contribute |d - threshold| (or |-1000 * weight| when d - threshold < tolerance;
 * tolerance NaN = none), angle terms |threshold - angle| in degrees, or |sin(angle)| if planar.
 * Kind 3 (Distance with center_of_mass, distance.py:131-148): point 0 is a probe atom, threshold_h
 * its mass, point 1 the target; a column made of kind-3 terms is the distance between the
 * mass-weighted centre of its probes and the target. */
int gmm_set_geometry(gmm_ctx* ctx, int n_terms, int n_obj, const int32_t* kind_h, const int32_t* obj_h,
                     const int32_t* atoms_h /*[n,4]*/, const double* points_h /*[n,4,3]*/,
                     const double* threshold_h, const double* tolerance_h, const double* weight_h,
                     const int32_t* planar_h);
/*   xyz_d [P, M, 3] fp32 device-frame coordinates; out_d [P, n_obj] fp64 */
int gmm_geometry(gmm_ctx* ctx, int P, const float* xyz_d, double* out_d, void* stream);

/* Coordination objective on expressed coordinates ("next" row of the scope:
 * gaudi/objectives/coordination.py:180-283 — sphere search around a metal, rigid coherent point
 * drift of the ligand atoms onto a geometry template (gaudi/_cpdrift.py:45-93), directionality of
 * the ligand atoms' free valences (:286-372) and the two optional corrections).  A site is one
 * Coordination objective; its tables run over all atoms of the job (every Molecule gene, caller's
 * numbering).  What Chimera supplies to the reference — bonded atoms in bond order, the geometry of
 * each atom's IDATM type, covalent bond radii — comes in as plain arrays. */
typedef struct gmm_coord_site {
    int n_atoms;
    const int32_t* mobile_row;     /* [n_atoms] row of the atom in the expressed poses, -1: static */
    const double* xyz;             /* [n_atoms, 3] world coordinates (read for static atoms) */
    const uint8_t* candidate;      /* [n_atoms] 1: passes atom_types / atom_elements / atom_names and
                                      belongs to one of `residues` (coordination.py:244-252,258) */
    const uint8_t* geometry;       /* [n_atoms] chimera.idatm geometry: 4 tetrahedral, 3 planar, 2 linear, 1 single */
    const double* bond_radius;     /* [n_atoms] chimera.Element.bondLength(a, b) = radius[a] + radius[b] */
    const int32_t* nb_ptr;         /* [n_atoms + 1] CSR of bonded atoms, in bond order (atom.neighbors) */
    const int32_t* nb_idx;
    int metal;                     /* `probe` */
    int n_vertices;                /* <= 8 */
    const double* template_xyz;    /* [(n_vertices + 1), 3] `geometry`, the metal's place last */
    double radius, distance, weight;
    int min_atoms;
    int flags;                     /* GMM_COORD_* */
} gmm_coord_site;
#define GMM_COORD_PREVENT_INTRUDERS 1
#define GMM_COORD_ENFORCE_ALL_RESIDUES 2   /* as in the reference, never satisfied: -1000 * weight */
#define GMM_COORD_COM_CORRECTION 4
#define GMM_COORD_DISTANCE_CORRECTION 8
/* Copies everything it needs; n_sites = 0 clears.  gmm_set_mobile / gmm_set_params first. */
int gmm_set_coordination(gmm_ctx* ctx, int n_sites, const gmm_coord_site* sites_h);
/*   xyz_d [P, M, 3] fp32 device-frame coordinates; out_d [P, n_sites] fp64.  An individual with
 *   more than 192 atoms within the radius of a metal gets NaN.  The first call for a larger P
 *   allocates its scratch rows (synchronous); later calls only enqueue two kernels. */
int gmm_coordination(gmm_ctx* ctx, int P, const float* xyz_d, double* out_d, void* stream);

/* Parity / debugging: the pair lists the Contacts score of ONE individual is made of
 * (what Contacts.find_interactions returns, contacts.py:140-146).  Atom ids: mobile atoms
 * 0..M-1, static atoms M..M+R-1 (caller's receptor order).  kind: 0 = clash (volume),
 * 1 = hydrophobic (LJ).  Synchronises.  Returns the number of pairs via n_out_h (may exceed
 * max_pairs; only max_pairs are written). */
int gmm_contacts_pairs(gmm_ctx* ctx, const float* xyz_one_d, int max_pairs,
                       int32_t* out_pairs_h /*[max,2]*/, int32_t* out_kind_h /*[max]*/,
                       double* out_overlap_h /*[max]*/, double* out_value_h /*[max]*/,
                       int* n_out_h, void* stream);

/* Measurement helpers ------------------------------------------------------------ */
/* Dependent-free FFMA loop; returns achieved FP32 TFLOP/s (the roofline denominator that
 * MEASURED_PEAKS.json does not carry).  Synchronises. */
int gmm_measure_fp32_peak(int device, double* tflops_out);
/* Pair tests executed by the last gmm_contacts/gmm_evaluate on this context (receptor
 * tiles that survived culling x mobile atoms + intra-mobile pairs).  The one-kernel Contacts path
 * keeps this tally only while profiling is on (gmm_set_profiling(ctx, 1)): it is instrumentation,
 * one global atomic per individual, and costs 1 % of the kernel.  Synchronises. */
int gmm_last_pair_tests(gmm_ctx* ctx, unsigned long long* n_out);
/* Per-kernel timing for roofline reports: when on, CUDA events are recorded around the
 * express, Contacts-sweep and Contacts-finalize kernels on the caller's stream;
 * gmm_kernel_times() synchronises and returns the durations (ms) of the LAST call of each. */
int gmm_set_profiling(gmm_ctx* ctx, int on);
int gmm_kernel_times(gmm_ctx* ctx, float* ms_out /*[3] express, sweep, finalize*/);
/* Kernel launches issued by this library since load (for bench.py's gpu_launches). */
unsigned long long gmm_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* GAUDIMM_B200_H */
