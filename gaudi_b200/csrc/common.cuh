// Shared declarations of libgaudimm_b200 (context, parameters, pair scoring).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <string>
#include <vector>

#include "../../include/gaudimm_b200.h"

namespace gmm {

// ---- error handling ---------------------------------------------------------------
void set_error(const char* fmt, ...);
extern std::atomic<unsigned long long> g_launches;   // kernels launched by this library (any thread)

#define GMM_CUDA(call)                                                              \
    do {                                                                            \
        cudaError_t _e = (call);                                                    \
        if (_e != cudaSuccess) {                                                    \
            gmm::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e),  \
                           __FILE__, __LINE__);                                     \
            return GMM_ECUDA;                                                       \
        }                                                                           \
    } while (0)

#define GMM_LAUNCH_CHECK(name)                                                      \
    do {                                                                            \
        ++gmm::g_launches;                                                          \
        cudaError_t _e = cudaGetLastError();                                        \
        if (_e != cudaSuccess) {                                                    \
            gmm::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));\
            return GMM_ECUDA;                                                       \
        }                                                                           \
    } while (0)

// selects the context's device for the duration of a call and restores the caller's
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// ---- tunables ---------------------------------------------------------------------
constexpr int kTileAtoms = 128;      // receptor atoms per tile (4 per lane)
constexpr int kSubTile = 32;         // receptor atoms per sub-tile (k-d leaf), one per lane
constexpr int kSweepWarps = 4;       // warps per sweep CTA
constexpr int kGroup = 8;            // individuals per sweep CTA
constexpr int kCandQ = 256;          // per-warp candidate queue entries
constexpr int kPairQ = 128;          // per-warp pair queue entries
constexpr double kFxScale = 4294967296.0;  // 2^32 fixed-point scale of score sums
constexpr float kFar = 1.0e15f;      // coordinate of padding atoms
constexpr unsigned kRflagNearFlex = 128;   // internal bit of d_rflags: the atom is within contact reach of where a non-seed mobile atom can be

// Contacts settings (gaudi/objectives/contacts.py:101-121,142-145), device copy.
struct Params {
    double zone_radius;
    double clash_thr;
    double hydro_thr;
    double hb_allow;
    double amv;          // assumedMaxVdw
    double cutoff;
    double weight;
    int options;
    float f_clash_thr, f_hydro_thr, f_allow, f_amv;   // fp32 copies for score_pair
};

// One term of a geometric restraint objective (geometry.cu)
struct GeoTerm {
    int kind;            // 0 distance, 1 angle (3 points), 2 dihedral (4 points), 3 mass-weighted centre (threshold = mass)
    int obj;             // which output column this term averages into
    int atom[4];         // mobile atom index, or -1 = use point[k]
    double point[4][3];  // world-frame coordinates of static atoms / literal points
    double threshold, tolerance, weight;
    int has_tolerance, planar;
};

// Where a Contacts kernel additionally stores each finished individual's scores when a fitness
// exchange is attached (exchange.cu): every rank's gather block, reached through peer mappings with
// plain stores.  Lives in device memory; the kernel reads the owner's epoch to pick the buffer.
struct MirrorTable {
    int n, rank, max_rows, pad;
    unsigned char* base[GMM_MAX_PEERS];
};

struct PairScore {
    long long clash_fx;  // |volume overlap| * 2^32 (0 if not a clash)
    long long hydro_fx;  // LJ-like energy * 2^32   (0 if not a hydrophobic contact)
    int kind;            // bit0 clash counted, bit1 hydrophobic counted
    double overlap;
    double clash_val, hydro_val;
};

// One unordered atom pair through DetectClash's acceptance rule and Contacts' scoring.
//   d, overlap, allowance:          SURVEY.md A.3 / contacts.py:142-145
//   positive / negative split:      contacts.py:176-187
//   LJ-like energy:                 contacts.py:215-241
//   spherical-cap overlap volume:   contacts.py:243-268 (pi literal 3.14159265359)
// Arithmetic is IEEE fp32 (correctly rounded sqrt and divisions, no fast-math): per pair the
// values carry ~1e-6 relative error, two orders below the 1e-4 the sums are held to, and the
// accept/reject decisions move by < 1e-6 A, inside the 1e-4 A band north_star exempts.  An fp64
// version of this function cost 16 % of the whole kernel (DSQRT + 3 DDIV per pair) for digits
// nobody reads.  Every kernel (fused, tiled, static-pair set-up, pair dump) calls this one
// function, so their results stay bit-identical to each other.
__device__ __forceinline__ PairScore score_pair(float ax, float ay, float az, float ra, unsigned fa,
                                                float bx, float by, float bz, float rb, unsigned fb,
                                                const Params& p) {
    PairScore o;
    o.clash_fx = 0; o.hydro_fx = 0; o.kind = 0; o.clash_val = 0.0; o.hydro_val = 0.0;
    const float dx = ax - bx, dy = ay - by, dz = az - bz;
    const float d = sqrtf(fmaf(dz, dz, fmaf(dy, dy, dx * dx)));
    const bool hb = ((fa & GMM_F_DONOR) && (fb & GMM_F_ACCEPTOR)) ||
                    ((fb & GMM_F_DONOR) && (fa & GMM_F_ACCEPTOR));
    const float r0 = ra + rb;
    const float ov = r0 - d - ((hb && p.f_allow != 0.f) ? p.f_allow : 0.f);
    o.overlap = (double)ov;
    const float maxr = fmaxf(ra, rb);
    // which == 'hydrophobic': DetectClash threshold = hydrophobic_threshold
    if (d <= maxr + p.f_amv - p.f_hydro_thr && ov >= p.f_hydro_thr && ov <= p.f_clash_thr &&
        (fa & GMM_F_HYDROPHOBIC) && (fb & GMM_F_HYDROPHOBIC)) {
        const float dist = r0 - ov;
        float x = r0 / dist;
        x = x * x; x = x * x * x;                 // (r0 / distance) ** 6
        float e = fmaf(x, x, -2.f * x);
        e = fminf(fmaxf(e, -1.0e9f), 1.0e9f);
        o.hydro_val = (double)e;
        o.hydro_fx = __double2ll_rn((double)e * kFxScale);
        o.kind |= 2;
    }
    // which == 'clashes': DetectClash threshold = clash_threshold, then overlap > clash_threshold
    if (d <= maxr + p.f_amv - p.f_clash_thr && ov > p.f_clash_thr) {
        const float dd = r0 - ov;
        float v;
        if (dd == 0.f) {
            v = 1000.f;
        } else {
            float ha = 0.f, hbb = 0.f;
            if (dd < r0) {
                const float inv = 1.f / (2.f * dd);
                ha = (rb * rb - (dd - ra) * (dd - ra)) * inv;
                hbb = (ra * ra - (dd - rb) * (dd - rb)) * inv;
            }
            v = (3.14159265359f / 3.f) * (ha * ha * (3.f * ra - ha) + hbb * hbb * (3.f * rb - hbb));
        }
        v = fminf(fabsf(v), 1.0e9f);
        o.clash_val = (double)v;
        o.clash_fx = __double2ll_rn((double)v * kFxScale);
        o.kind |= 1;
    }
    return o;
}

}  // namespace gmm

// ---- the context ------------------------------------------------------------------
struct gmm_ctx {
    int device = 0;
    bool has_params = false, has_mobile = false, has_receptor = false;
    gmm::Params prm{};

    // mobile model (slot order = seeds first; d_slot2atom maps back to the caller's order)
    int M = 0, Mp = 0, T = 0, W = 0, n_allele_tors = 0, n_seed = 0;
    int n_hseed = 0;                 // zone seeds that are hydrophobic: slots [0, n_hseed)
    int origin_atom = 0, tors_before_search = 1;
    double* d_xyz0 = nullptr;        // [M,3]   caller order
    uint8_t* d_mflags_atom = nullptr;// [M]     caller order (express needs F_XFORM)
    int32_t* d_tfixed = nullptr; int32_t* d_tmoving = nullptr; int32_t* d_tsrc = nullptr;
    uint8_t* d_tsnap = nullptr; uint64_t* d_tmask = nullptr;
    // kinematic-tree form of the same torsions (express.cu), valid when tree_ok: bonds in
    // application order (innermost moving set first), template axes, per-atom bond bitmask
    bool tree_ok = false;
    int32_t* d_tree_src = nullptr; uint8_t* d_tree_snap = nullptr;
    double* d_tree_axis = nullptr;   // [T,6] unit axis (fixed -> moving) and the moving atom's template position
    uint64_t* d_tree_amask = nullptr;  // [M] bit t: atom turns with bond t (T <= 64)
    int32_t* d_slot2atom = nullptr;  // [Mp]    slot -> caller atom index (-1 = padding)
    float* d_mrad_slot = nullptr;    // [Mp]
    uint8_t* d_mflags_slot = nullptr;// [Mp]
    uint32_t* d_mm_pairs = nullptr;  // [n_mm]  slotA | slotB << 16, not excluded
    int n_mm = 0;
    float* d_mm_c2 = nullptr;        // [n_mm]  squared pre-filter cutoff of each pair (set with the receptor)
    std::vector<uint32_t> h_mm_pairs; std::vector<uint8_t> h_mflags_slot;
    float rmax_mobile = 0.f;
    bool all_seed = true;
    std::vector<int32_t> h_slot2atom, h_atom2slot;
    std::vector<float> h_mrad;
    std::vector<float4> h_ball;       // [Mp] slot order: where the atom can ever be (template frame), w = radius, 3e38 = anywhere
    // rotamers
    int n_res = 0; int32_t* d_rot_ptr = nullptr; double* d_rot_delta = nullptr;

    // receptor (Morton-sorted, padded to whole tiles)
    int R = 0, Rpad = 0, n_tiles = 0;
    float4* d_rec4 = nullptr;        // [Rpad] x,y,z (device frame), radius
    uint8_t* d_rflags = nullptr;     // [Rpad]
    float* d_tile_box = nullptr;     // [n_tiles, 8] lo.xyz, hi.xyz, max radius, pad
    uint32_t* d_always = nullptr;    // [n_tiles*4] zone words of always-in-zone atoms
    int32_t* d_rr_ptr = nullptr; int32_t* d_rr_idx = nullptr;   // sorted index space
    long long* d_rr_val = nullptr;   // [nnz,2] clash_fx, hydro_fx
    uint8_t* d_rr_kind = nullptr;    // [nnz]
    uint4* d_edge = nullptr;         // [nnz]   x: b | a_local << 24 | kind << 29, (z, w): the pair's value
                                     //         (clash or hydrophobic, never both) as int64 fixed point (fused kernel)
    float* d_sub_box = nullptr;      // [Rpad/32 + 1, 8] sub-tile boxes (last = dummy leaf)
    int rr_nnz = 0;
    uint32_t* d_flex_mask = nullptr; uint32_t* d_body_range = nullptr; int n_bodies = 0;   // see SweepArgs
    float4* d_sseed = nullptr; int n_sseed = 0;                  // static zone seeds (CB atoms)
    int32_t* d_ms_ptr = nullptr; int32_t* d_ms_idx = nullptr;   // CSR over slots -> sorted static idx
    int ms_nnz = 0;
    std::vector<int32_t> h_sorted2orig;
    float rmax_static = 0.f;
    double centre[3] = {0, 0, 0};

    // geometric restraint objectives (gmm_set_geometry)
    gmm::GeoTerm* d_geo_terms = nullptr; int n_geo_terms = 0, n_geo_obj = 0;

    void* coordination = nullptr;    // gmm::CoordState (coordination.cu)

    // scratch (grown on demand)
    void* d_scratch = nullptr; size_t scratch_bytes = 0;
    unsigned long long* d_counters = nullptr;   // [4] pair tests, ...
    // optional per-kernel timing (gmm_set_profiling): events around express / sweep / finalize
    bool prof = false;
    cudaEvent_t ev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    bool ev_used[3] = {false, false, false};
    // second stream + events for the chunked upload/compute overlap of gmm_evaluate_host
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t ev_copy[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    // pinned staging for gmm_evaluate_host
    void* h_pin = nullptr; size_t pin_bytes = 0;
    void* d_stage = nullptr; size_t stage_bytes = 0;
    void* d_rmsd_ws = nullptr; size_t rmsd_ws_bytes = 0;   // centred conformers + norms (rmsd.cu, bitmask path)
    void* exchange = nullptr;        // gmm::Exchange (exchange.cu): fitness exchange with the other ranks' GPUs
    const gmm::MirrorTable* d_mirror = nullptr;   // device copy of the peers' blocks (set by gmm_exchange_connect)
    int pub_row0 = -1;               // >= 0 while gmm_evaluate_publish runs: row of individual 0 in the gathered table
    bool pub_done = false;           // the Contacts kernel stored the rows itself
};

namespace gmm {
// kernels' host launchers (each returns a GMM_* code)
int launch_express(gmm_ctx* c, int P, const double* tors, const double* xform, const double* rot,
                   float* out_xyz, cudaStream_t s);
int launch_contacts(gmm_ctx* c, int P, const float* xyz, double* scores, int32_t* counts,
                    cudaStream_t s, void* workspace = nullptr);
int launch_contacts_fused(gmm_ctx* c, int P, const float* xyz, double* scores, int32_t* counts, cudaStream_t s,
                          int* handled);
void prefilter_cutoffs(const gmm_ctx* c, float* thr_min, float* thr_tight);
bool fused_applicable(const gmm_ctx* c);
int launch_geometry(gmm_ctx* c, int P, const float* xyz, double* out, cudaStream_t s);
int launch_rr_values(gmm_ctx* c, cudaStream_t s);
int coordination_set(gmm_ctx* c, int n_sites, const gmm_coord_site* sites);
int coordination_sites(const gmm_ctx* c);
void coordination_destroy(gmm_ctx* c);
int launch_coordination(gmm_ctx* c, int P, const float* xyz, double* out, cudaStream_t s);
int launch_rmsd(gmm_ctx* c, int N, int A, const float* xyz, int row0, int row1, float* out_rmsd,
                uint32_t* out_bits, double thr, cudaStream_t s);
int reserve_rmsd_workspace(gmm_ctx* c, int N, int A);
int launch_pairs_dump(gmm_ctx* c, const float* xyz_one, int max_pairs, int32_t* pairs_h,
                      int32_t* kind_h, double* ov_h, double* val_h, int* n_out, cudaStream_t s);
int ensure_scratch(gmm_ctx* c, size_t bytes);
// cudaFuncAttributeMaxDynamicSharedMemorySize, applied once per (kernel, device)
int allow_dynamic_smem(const void* kernel, int device, size_t bytes);
size_t contacts_scratch_bytes(const gmm_ctx* c, int P);
}  // namespace gmm
