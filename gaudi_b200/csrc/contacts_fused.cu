// Contacts objective, one-kernel path (gaudi/objectives/contacts.py:129-268; Chimera's
// specifier.zone at :210-212 and DetectClash.detectClash at :145): a team of S warps (S = 1 for
// large populations, S = 4 = one CTA for small ones) owns one individual from its expressed
// coordinates to its two scores, so nothing but the scores leaves the SM.
//
//   0  stage    the pose's mobile atoms go to shared memory twice: exact (x, y, z, radius) and
//               re-centred on the pose's own centre c (x-c, y-c, z-c, |.|^2);
//   1  cull     lanes test the boxes of 32-atom receptor leaves against the pose in parallel
//               and ballot-compact the survivors into a list (all leaves when culling is off);
//               a second pass keeps a leaf only if some mobile atom is within reach of its box;
//   2  sweep    four leaves at a time, one atom of each per lane.  With a = -2 (r - c) and
//               n = |r - c|^2 kept in registers, d^2 = n + min_i (a . m_i + |m_i|^2): THREE FFMA
//               per pair test and one FMNMX3 per two, branch-free.  Working relative to the pose centre
//               keeps every magnitude small, so this expanded form is accurate to ~1e-4 A^2;
//               receptor atoms whose min d^2 lands within a margin of radius^2 get their zone
//               membership re-decided with exact differences (rare), so zone bits are identical
//               to the direct-difference definition.  Candidates (atoms that can have a contact
//               at all: in the zone and within contact reach of a seed, or flagged at set-up as within
//               reach of a flexible side chain's range of motion) are expanded against the mobile atoms
//               and the surviving pairs scored with all lanes busy (score_pair), as in contacts.cu;
//   3  static   static-static pairs are stored as per-leaf edge lists of 16-byte records (partner,
//               kind, precomputed value); leaves with a non-empty zone word walk their edges with
//               coalesced 128-bit loads and test both zone bits in shared memory;
//   4  intra    pairs inside the mobile set (exclusions pre-removed);
//   5  finish   fixed-point sums -> doubles, x2 (contacts.py:176-177), cutoff penalty
//               (contacts.py:132-133).
#include "contacts_common.cuh"

namespace gmm {

struct FusedArgs {
    SweepArgs s;
    int NST, R;
    const float* sub_box;
    const int32_t* rr_ptr; const uint4* edge;
    const uint32_t* mm_pairs; const float* mm_c2; int n_mm;
    double* scores; int32_t* counts;
    int ind_bytes;
    int count_tests;  // 1: tally the executed pair tests (gmm_last_pair_tests); on while profiling is enabled
    int tma_pose;     // 1: stage each pose with one TMA bulk copy (size and alignment permit)
    const MirrorTable* mirror;   // != NULL: also store the rows into every rank's gather block (exchange.cu)
    int mirror_row0;
};

// Per-warp candidate queue of this kernel (contacts.cu keeps kCandQ = 256).  192 entries instead of 256:
// the configs[4] job (222 mobile slots, 145 leaves) then fits six CTAs per SM instead of five (3 % faster);
// configs[2] is unchanged.  A sweep step adds at most 128 candidates, so the queue is expanded above 64.
constexpr int kFCandQ = 192;

struct Acc { long long cfx, hfx; int nc, nh; };

// Warps per CTA.  A team of four needs its four.  The warp-per-individual PLAIN variants run ONE warp
// per CTA: with `warp`, `team` and every per-warp shared-memory offset compile-time zeros the kernel
// drops to 63 registers without a spill and the 65,536-individual step is 4 % faster (8,192: 7 %);
// the gain is the code, not the scheduling granularity — the same kernel compiled for four warps and
// launched with one is slower than either.  The variants with flexible side
// chains stay at four warps: one warp per CTA costs them occupancy (22 resident warps instead of 24 on
// configs[4]: 5.92 ms against 5.75).  Built for one warp they also hit an illegal address in the
// candidate expansion on configs[4] for as long as the OR over the lanes' body masks was a
// __reduce_or_sync (REDUX); with a shuffle butterfly in its place the build is clean — cause not found,
// the shuffles stay.
__host__ __device__ constexpr int fused_warps(int S, bool plain) { return (S == 1 && plain) ? 1 : 4; }

// Warp reductions on the REDUX unit (one instruction per 32-bit word) instead of five rounds of
// shuffles: floats through the usual order-preserving map to unsigned integers, 64-bit sums in
// three pieces of at most 24 bits so that 32 lanes cannot overflow a word.
__device__ __forceinline__ unsigned float_key(float v) {
    const unsigned b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_float(unsigned k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
__device__ __forceinline__ float warp_min(float v) { return key_float(__reduce_min_sync(0xffffffffu, float_key(v))); }
__device__ __forceinline__ float warp_max(float v) { return key_float(__reduce_max_sync(0xffffffffu, float_key(v))); }
__device__ __forceinline__ long long warp_sum(long long v) {
    const unsigned long long u = (unsigned long long)v;
    const unsigned long long a = __reduce_add_sync(0xffffffffu, (unsigned)(u & 0xffffffu));
    const unsigned long long b = __reduce_add_sync(0xffffffffu, (unsigned)((u >> 24) & 0xffffffu));
    const unsigned long long c = __reduce_add_sync(0xffffffffu, (unsigned)(u >> 48));
    return (long long)(a + (b << 24) + (c << 48));          // modulo 2^64, as the lane-wise sum would be
}

template <int S>
__device__ __forceinline__ void team_sync() {
    if (S == 1) __syncwarp(); else __syncthreads();
}

template <bool PLAIN>
__device__ __forceinline__ void fused_drain_pairs(const SweepArgs& a, uint32_t* pq, int& npq, const float4* ligx,
                                                  Acc& acc, int lane) {
    for (int base = 0; base < npq; base += 32) {
        const int e = base + lane;
        if (e < npq) {
            const uint32_t pr = pq[e];
            const int s = pr >> 22, j = pr & 0x3fffff;
            if (PLAIN || !ms_excluded(a.ms_ptr, a.ms_idx, s, j)) {
                const float4 m = ligx[s];
                const float4 r = a.rec4[j];
                const PairScore ps = score_pair(m.x, m.y, m.z, m.w, a.mflags[s], r.x, r.y, r.z, r.w, a.rflags[j], a.prm);
                if (ps.kind & 1) { acc.cfx += ps.clash_fx; ++acc.nc; }
                if (ps.kind & 2) { acc.hfx += ps.hydro_fx; ++acc.nh; }
            }
        }
    }
    __syncwarp();
    npq = 0;
}

// Expand candidates (receptor atoms that can have a contact at all) against the mobile atoms.
// One candidate per lane.  The pre-filter uses the sweep's expanded form relative to the pose
// centre: d^2 ~ n + (a . m + |m|^2) with a = -2 (r - c), three FFMA per pair, accepted when
// d^2 <= (r_m + r_r - thr_min)^2 widened by the same error bound the sweep uses; it only has to
// be a superset, score_pair decides every pair with exact differences.  Hits are collected as a
// bit per mobile atom and the set bits of all lanes are then compacted into the pair queue one
// ballot per round (rounds = the largest hit count of any lane, a handful) instead of one ballot
// per mobile atom.
template <bool PLAIN>
__device__ __forceinline__ void fused_drain_candidates(const SweepArgs& a, uint32_t* cq, int& ncq, uint32_t* pq,
                                                       int& npq, const float4* ligx, const float4* ligc,
                                                       const float4 sph, const uint32_t* zm, Acc& acc, int lane) {
    const unsigned lt = (1u << lane) - 1u;
    for (int base = 0; base < ncq; base += 32) {
        const int e = base + lane;
        int j = 0;
        unsigned near_seed = 0, bodies = 0;
        float ax = 0.f, ay = 0.f, az = 0.f, cr_h = 0.f, cr_c = 0.f, slack = -3e38f;   // idle lane: nothing passes
        if (e < ncq) {
            j = (int)cq[e];
            if (!PLAIN) {
                near_seed = (unsigned)j >> 31;          // set by the sweep: within contact reach of a zone seed
                j &= 0x3fffff;
                bodies = a.n_bodies ? a.flex_mask[j] : 0u;      // no non-seed atoms: no table
            }
            const float4 r = a.rec4[j];
            const bool rh = (a.rflags[j] & GMM_F_HYDROPHOBIC) != 0;
            const float x = r.x - sph.x, y = r.y - sph.y, z = r.z - sph.z;
            const float nr = fmaf(z, z, fmaf(y, y, x * x));
            ax = -2.f * x; ay = -2.f * y; az = -2.f * z;
            cr_c = r.w - a.thr_tight;                    // partner that cannot make a hydrophobic contact
            cr_h = rh ? r.w - a.thr_min : cr_c;          // partner that can (if this atom can)
            slack = 2e-4f + 2e-6f * (nr + sph.w) - nr;   // error bound of the form, minus n
        }
        // PLAIN: every slot, in chunks of 32.  Otherwise: the seed slots in chunks of 32 if some
        // candidate of this batch is near a seed, then only the bodies (side chains) some candidate
        // is within reach of — a batch comes from one or two leaves, i.e. from one corner of the zone.
        const unsigned any_seed = PLAIN ? 1u : __ballot_sync(0xffffffffu, near_seed != 0u);
        // (a butterfly of shuffles, not __reduce_or_sync: built for one warp per CTA, the variant with the
        // REDUX here faulted on configs[4] and the one with the shuffles does not — see fused_warps)
        unsigned any_body = 0u;
        if (!PLAIN) {
            any_body = bodies;
#pragma unroll
            for (int sh = 16; sh > 0; sh >>= 1) any_body |= __shfl_xor_sync(0xffffffffu, any_body, sh);
        }
        const int n_schunks = ((PLAIN ? a.M : a.n_seed) + 31) >> 5;
        for (int c = 0; c < n_schunks + (PLAIN ? 0 : a.n_bodies); ++c) {
            int s0, n;
            unsigned zbits = 0xffffffffu;
            if (PLAIN || c < n_schunks) {
                if (!PLAIN && !any_seed) continue;
                s0 = c << 5;
                n = min(32, (PLAIN ? a.M : a.n_seed) - s0);
                if (!PLAIN) zbits = zm[c];
            } else {
                if (!((any_body >> (c - n_schunks)) & 1u)) continue;
                const uint32_t br = a.body_range[c - n_schunks];
                s0 = (int)(br & 0xffffu);
                n = (int)(br >> 16) - s0;
                zbits = __funnelshift_r(zm[s0 >> 5], zm[(s0 >> 5) + 1], s0 & 31);   // zm has one spare word
            }
            unsigned hits = 0;
            // slots [0, n_hseed) are hydrophobic seeds, [n_hseed, n_seed) the other seeds; mobile
            // atoms that are not seeds (flexible side chains) are not sorted and keep the wide cutoff
#pragma unroll 4
            for (int k = 0; k < n; ++k) {
                const int s = s0 + k;
                const float4 m = ligc[s];
                const float t = ligx[s].w + ((s < a.n_hseed || (!PLAIN && s >= a.n_seed)) ? cr_h : cr_c);
                const float dot = fmaf(ax, m.x, fmaf(ay, m.y, fmaf(az, m.z, m.w)));
                if (dot <= fmaf(t * 1.000004f, t, slack)) hits |= 1u << k;
            }
            if (!PLAIN) hits &= zbits;
            unsigned live = __ballot_sync(0xffffffffu, hits != 0u);
            while (live) {
                if (hits) {
                    const int k = __ffs(hits) - 1;
                    hits &= hits - 1u;
                    pq[npq + __popc(live & lt)] = ((uint32_t)(s0 + k) << 22) | (uint32_t)j;
                }
                npq += __popc(live);
                __syncwarp();
                if (npq > kPairQ - 32) fused_drain_pairs<PLAIN>(a, pq, npq, ligx, acc, lane);
                live = __ballot_sync(0xffffffffu, hits != 0u);
            }
        }
    }
    __syncwarp();
    ncq = 0;                 // pairs still queued stay in pq; the caller flushes after the sweep
}

// S: warps per individual.  CULL: leaves out of reach of the pose are skipped (the normal mode);
// without it every leaf is swept (the dense mode the FP32 roofline is quoted for).  A template
// parameter rather than a run-time flag: the kernel is sensitive to its instruction footprint
// and each variant carries only its own code.  PLAIN: the common docking job — every mobile atom is a
// zone seed (a ligand, no flexible side chains), the receptor holds no seeds and no atom bonded to
// the mobile set: the mobile-zone bits, the always-in-zone words, the side-chain ("body") masks of
// the candidate expansion and the exclusion lookups are compiled out.
template <int S, bool CULL, bool PLAIN>
// (one-warp CTAs: 28 resident instead of 32, i.e. 72 registers instead of 64 — 2 % faster; shared
// memory allows 31 on configs[2] anyway; 80 registers / 25 CTAs is no better than 64 / 32.  The
// four-warp variants with flexible side chains are bound by shared memory to 6-7 CTAs per SM, so they
// get 72 registers as well (7 CTAs): no spills, configs[4] 4 % faster; 80 registers / 6 CTAs gives 1 % more
// on configs[4] but would cost the mid-sized jobs a CTA)
__global__ void __launch_bounds__(fused_warps(S, PLAIN) * 32, fused_warps(S, PLAIN) == 1 ? 28 : ((S == 1 && !PLAIN) ? 7 : 8))
contacts_fused_kernel(const FusedArgs f) {
    const SweepArgs& a = f.s;
    constexpr int IPC = fused_warps(S, PLAIN) / S;            // individuals per CTA
    constexpr int TL = 32 * S;            // lanes per team
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int team = warp / S, wt = warp % S, tlane = wt * 32 + lane;
    const int p = blockIdx.x * IPC + team;
    if (S == 1 && p >= a.P) return;       // whole warp leaves; S == 4 has one individual per CTA

    unsigned char* ib = smem_raw + (size_t)team * f.ind_bytes;
    float4* ligx = reinterpret_cast<float4*>(ib);                 // [Mp] exact x, y, z, radius
    float4* ligc = ligx + a.Mp;                                   // [Mp] centred x, y, z, |.|^2
    float4* misc = ligc + a.Mp;                                   // [0] sphere, [1] box lo, [2] box hi
    long long* accs = reinterpret_cast<long long*>(misc + 3);     // [4]
    int* nlist_p = reinterpret_cast<int*>(accs + 4);              // [4] (16 bytes)
    uint16_t* list = reinterpret_cast<uint16_t*>(nlist_p + 4);    // [NST + 4] leaf ids, 16-byte aligned
    uint32_t* zone = reinterpret_cast<uint32_t*>(list + (f.NST + 4));   // [NST + 1] (NST is a multiple of 4)
    uint32_t* zm = zone + (f.NST + 1);                            // [MW]
    uint32_t* cq = reinterpret_cast<uint32_t*>(smem_raw + (size_t)IPC * f.ind_bytes) + warp * (kFCandQ + kPairQ);
    uint32_t* pq = cq + kFCandQ;

    // ---- 0. stage the pose --------------------------------------------------------------
    // The pose's M x 3 floats are one contiguous block in HBM: a single TMA bulk copy (cp.async.bulk,
    // SASS UBLKCP), completion signalled on an mbarrier; lanes then permute it into slot order.
    // General form (!PLAIN): the 16-byte aligned span around the pose lands in the area of the centred
    // copy (written only after the bounding sphere is known); a pose whose span would end past the
    // array (the last one, when P x M x 12 is not a multiple of 16) is read with plain loads.
    // PLAIN keeps the narrower form — poses of whole 16-byte multiples landing in this team's queue
    // area: with the general one compiled in, the same sweep loop got a different register
    // assignment and configs[2] ran 1.3 % slower.  f.tma_pose == 0 (decided per
    // launch): plain loads for everything.
    const unsigned shift = PLAIN ? 0u : (((unsigned)p * (unsigned)(a.M * 3)) & 3u);
    if (f.tma_pose && (PLAIN || p + 1 < a.P || (((unsigned)a.P * (unsigned)(a.M * 3)) & 3u) == 0u)) {
        const float* raw = PLAIN ? reinterpret_cast<const float*>(smem_raw + (size_t)IPC * f.ind_bytes) +
                                       (size_t)(warp - wt) * (kFCandQ + kPairQ)
                                 : reinterpret_cast<const float*>(ligc);
        unsigned long long* mbar = reinterpret_cast<unsigned long long*>(nlist_p + 2);
        const unsigned mbar_s = (unsigned)__cvta_generic_to_shared(mbar);
        const unsigned bytes = PLAIN ? (unsigned)(a.M * 3 * sizeof(float))
                                     : ((shift + (unsigned)(a.M * 3) + 3u) & ~3u) * (unsigned)sizeof(float);
        if (tlane == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar_s));
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        team_sync<S>();
        if (tlane == 0) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar_s), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"((unsigned)__cvta_generic_to_shared(raw)), "l"(a.xyz + ((size_t)p * a.M * 3 - shift)),
                           "r"(bytes), "r"(mbar_s) : "memory");
        }
        // The wait is bounded by TIME (two seconds of the global timer, looked at every 4,096 polls),
        // not by a poll count: under a debugger, a profiler's replay or time-slicing a healthy copy can
        // take arbitrarily many polls.  A copy that has still not landed then is given up on without
        // taking the context down: the individual's scores become NaN and d_counters[1] counts it.
        unsigned done = 0, expired = 0;
        unsigned long long t_start = 0;
        for (int spin = 0; !done && !expired; ++spin) {
            asm volatile("{\n\t.reg .pred q;\n\t"
                         "mbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0;\n\t"
                         "selp.u32 %0, 1, 0, q;\n\t}" : "=r"(done) : "r"(mbar_s) : "memory");
            if (!done && (spin & 4095) == 4095) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t_start == 0) t_start = now;
                else if (now - t_start > 2000000000ull) expired = 1;
            }
        }
        if (S == 1 ? __any_sync(0xffffffffu, expired) : __syncthreads_or((int)expired)) {
            if (tlane == 0) {
                atomicAdd(a.counters + 1, 1ull);
                f.scores[2 * (size_t)p] = __longlong_as_double(0x7ff8000000000000ll);
                f.scores[2 * (size_t)p + 1] = __longlong_as_double(0x7ff8000000000000ll);
                if (f.counts) { f.counts[2 * (size_t)p] = -1; f.counts[2 * (size_t)p + 1] = -1; }
            }
            return;
        }
        for (int s = tlane; s < a.Mp; s += TL) {
            const int atom = a.slot2atom[s];
            float4 v = make_float4(kFar, kFar, kFar, 0.f);
            if (atom >= 0) v = make_float4(raw[shift + 3 * atom], raw[shift + 3 * atom + 1], raw[shift + 3 * atom + 2], a.mrad[s]);
            ligx[s] = v;
        }
    } else {
        for (int s = tlane; s < a.Mp; s += TL) {
            const int atom = a.slot2atom[s];
            float4 v = make_float4(kFar, kFar, kFar, 0.f);
            if (atom >= 0) {
                const float* src = a.xyz + ((size_t)p * a.M + atom) * 3;
                v = make_float4(src[0], src[1], src[2], a.mrad[s]);
            }
            ligx[s] = v;
        }
    }
    if (tlane < 4) accs[tlane] = 0;
    if (tlane == 0) { nlist_p[0] = 0; if (!PLAIN) zm[a.MW] = 0u; }
    team_sync<S>();
    const bool internal = a.prm.options & GMM_OPT_ONLY_INTERNAL;
    // mobile atoms in the zone: seeds always, the others (flexible side chains) iff within radius
    // of a seed; one atom per lane, 32 atoms = one word per ballot
    if (!PLAIN)
    for (int base = wt * 32; base < a.MW * 32; base += TL) {
        const int s = base + lane;
        bool in = s < a.n_seed;
        if (!in && s < a.M && !internal) {
            const float4 q = ligx[s];
            // static seeds first: a flexible side chain's atoms mostly sit within the radius of their own CB
#pragma unroll 1
            for (int k = 0; k < a.n_sseed && !in; ++k)
                in = dist2f(a.sseed[k].x, a.sseed[k].y, a.sseed[k].z, q.x, q.y, q.z) <= a.rad2;
#pragma unroll 1
            for (int k = 0; k < a.n_seed && !in; ++k)
                in = dist2f(ligx[k].x, ligx[k].y, ligx[k].z, q.x, q.y, q.z) <= a.rad2;
        }
        const unsigned bits = __ballot_sync(0xffffffffu, in && s < a.M);
        if (lane == 0) zm[base >> 5] = bits;
    }
    if (wt == 0) {                                     // bounding box and sphere of the pose
        float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
#pragma unroll 1
        for (int s = lane; s < a.M; s += 32) {
            const float4 m = ligx[s];
            lo[0] = fminf(lo[0], m.x); hi[0] = fmaxf(hi[0], m.x);
            lo[1] = fminf(lo[1], m.y); hi[1] = fmaxf(hi[1], m.y);
            lo[2] = fminf(lo[2], m.z); hi[2] = fmaxf(hi[2], m.z);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { lo[k] = warp_min(lo[k]); hi[k] = warp_max(hi[k]); }
        const float cx = 0.5f * (lo[0] + hi[0]), cy = 0.5f * (lo[1] + hi[1]), cz = 0.5f * (lo[2] + hi[2]);
        float r2 = 0.f;
#pragma unroll 1
        for (int s = lane; s < a.M; s += 32) r2 = fmaxf(r2, dist2f(ligx[s].x, ligx[s].y, ligx[s].z, cx, cy, cz));
        r2 = warp_max(r2);
        if (lane == 0) {
            misc[0] = make_float4(cx, cy, cz, r2);
            misc[1] = make_float4(lo[0], lo[1], lo[2], 0.f);
            misc[2] = make_float4(hi[0], hi[1], hi[2], 0.f);
        }
    }
    team_sync<S>();
    const float4 sph = misc[0];
    const float wmax = sph.w;
#pragma unroll 1
    for (int s = tlane; s < a.Mp; s += TL) {
        const float4 m = ligx[s];
        const float x = m.x - sph.x, y = m.y - sph.y, z = m.z - sph.z;
        ligc[s] = make_float4(x, y, z, fmaf(z, z, fmaf(y, y, x * x)));
    }
    // always-in-zone receptor atoms exist only when the receptor holds zone seeds (n_sseed > 0)
    const bool has_always = !PLAIN && a.n_sseed > 0;
#pragma unroll 2
    for (int st = tlane; st <= f.NST; st += TL) zone[st] = has_always ? a.always[st] : 0u;

    Acc acc; acc.cfx = 0; acc.hfx = 0; acc.nc = 0; acc.nh = 0;
    int swept_leaves = 0;
    const bool have_sweep = !internal && f.R > 0;

    if (have_sweep) {
        // ---- 1. surviving leaves ----------------------------------------------------------
        const float4 blo = misc[1], bhi = misc[2];
        const float prad = sqrtf(sph.w) * 1.00001f + 1e-4f;
        int nl = 0;
        if (S == 1 && CULL) {
            // coarse level: 128-atom tiles (4 consecutive leaves); the refine pass below does the
            // leaf-level work, so every leaf of a surviving tile simply goes on the list
            for (int base = 0; base < a.n_tiles; base += 32) {
                const int t = base + lane;
                bool keep = false;
                if (t < a.n_tiles) {
                    const float4 b0 = *reinterpret_cast<const float4*>(a.tile_box + (size_t)t * 8);
                    const float4 b1 = *reinterpret_cast<const float4*>(a.tile_box + (size_t)t * 8 + 4);
                    const float reach = fmaxf(sqrtf(a.rad2), b1.z + a.rmax_mobile - a.thr_min) + a.reach_extra;
                    const float ex = fmaxf(fmaxf(b0.x - sph.x, sph.x - b0.w), 0.f);
                    const float ey = fmaxf(fmaxf(b0.y - sph.y, sph.y - b1.x), 0.f);
                    const float ez = fmaxf(fmaxf(b0.z - sph.z, sph.z - b1.y), 0.f);
                    const float lim = prad + reach;
                    keep = ex * ex + ey * ey + ez * ez <= lim * lim &&
                           b0.x <= bhi.x + reach && b0.w >= blo.x - reach &&
                           b0.y <= bhi.y + reach && b1.x >= blo.y - reach &&
                           b0.z <= bhi.z + reach && b1.y >= blo.z - reach;
                    if (has_always)
                        keep = keep || (a.always[4 * t] | a.always[4 * t + 1] | a.always[4 * t + 2] | a.always[4 * t + 3]) != 0u;
                }
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                if (keep) {
                    const int pos = nl + 4 * __popc(mask & ((1u << lane) - 1u));
                    const unsigned l0 = 4u * (unsigned)t;
                    *reinterpret_cast<uint2*>(list + pos) = make_uint2(l0 | ((l0 + 1u) << 16), (l0 + 2u) | ((l0 + 3u) << 16));
                }
                nl += 4 * __popc(mask);
            }
        } else
        for (int base = 0; base < f.NST; base += TL) {
            const int st = base + tlane;
            bool keep = false;
            if (st < f.NST && st * kSubTile < f.R) {
                if (!CULL || (has_always && a.always[st] != 0u)) {
                    keep = true;
                } else {
                    const float* bx = f.sub_box + (size_t)st * 8;
                    const float reach = fmaxf(sqrtf(a.rad2), bx[6] + a.rmax_mobile - a.thr_min) + a.reach_extra;
                    const float ex = fmaxf(fmaxf(bx[0] - sph.x, sph.x - bx[3]), 0.f);
                    const float ey = fmaxf(fmaxf(bx[1] - sph.y, sph.y - bx[4]), 0.f);
                    const float ez = fmaxf(fmaxf(bx[2] - sph.z, sph.z - bx[5]), 0.f);
                    const float lim = prad + reach;
                    keep = ex * ex + ey * ey + ez * ez <= lim * lim &&
                           bx[0] <= bhi.x + reach && bx[3] >= blo.x - reach &&
                           bx[1] <= bhi.y + reach && bx[4] >= blo.y - reach &&
                           bx[2] <= bhi.z + reach && bx[5] >= blo.z - reach;
                }
            }
            const unsigned mask = __ballot_sync(0xffffffffu, keep);
            int pos;
            if (S == 1) {
                pos = nl;
                nl += __popc(mask);
            } else {
                int b0 = 0;
                if (lane == 0 && mask) b0 = atomicAdd(nlist_p, __popc(mask));
                pos = __shfl_sync(0xffffffffu, b0, 0);
            }
            if (keep) list[pos + __popc(mask & ((1u << lane) - 1u))] = (uint16_t)st;
        }
        if (S == 1 && CULL) {
            // Leaf level, cheap test first: the sphere/box test again, now on the 32-atom boxes
            // (about half of the leaves of the surviving tiles go).  Compacted in place.
            __syncwarp();
            int nmid = 0;
            for (int base = 0; base < nl; base += 32) {
                const int i = base + lane;
                uint32_t st = 0;
                bool keep = false;
                if (i < nl) {
                    st = list[i];
                    keep = has_always && a.always[st] != 0u;
                    if (!keep && st * kSubTile < (uint32_t)f.R) {
                        const float4 b0 = *reinterpret_cast<const float4*>(f.sub_box + (size_t)st * 8);
                        const float4 b1 = *reinterpret_cast<const float4*>(f.sub_box + (size_t)st * 8 + 4);
                        const float reach = fmaxf(sqrtf(a.rad2), b1.z + a.rmax_mobile - a.thr_min) + a.reach_extra;
                        const float ex = fmaxf(fmaxf(b0.x - sph.x, sph.x - b0.w), 0.f);
                        const float ey = fmaxf(fmaxf(b0.y - sph.y, sph.y - b1.x), 0.f);
                        const float ez = fmaxf(fmaxf(b0.z - sph.z, sph.z - b1.y), 0.f);
                        const float lim = prad + reach;
                        keep = ex * ex + ey * ey + ez * ez <= lim * lim &&
                               b0.x <= bhi.x + reach && b0.w >= blo.x - reach &&
                               b0.y <= bhi.y + reach && b1.x >= blo.y - reach &&
                               b0.z <= bhi.z + reach && b1.y >= blo.z - reach;
                    }
                }
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                __syncwarp();
                if (keep) list[nmid + __popc(mask & ((1u << lane) - 1u))] = (uint16_t)st;
                nmid += __popc(mask);
                __syncwarp();
            }
            nl = nmid;
            // Refine: a leaf stays only if some mobile atom is within reach of its box (the
            // sphere/box test above keeps ~40 leaves, this keeps ~27; the sweep of a leaf costs
            // more than this test).  One leaf per lane, compacted in place.
            int nout = 0;
            for (int base = 0; base < nl; base += 32) {
                const int i = base + lane;
                uint32_t st = 0;
                bool keep = false;
                if (i < nl) {
                    st = list[i];
                    keep = has_always && a.always[st] != 0u;
                    if (!keep) {
                        const float* bx = f.sub_box + (size_t)st * 8;
                        const float b0 = bx[0], b1 = bx[1], b2 = bx[2], b3 = bx[3], b4 = bx[4], b5 = bx[5];
                        const float reach = fmaxf(sqrtf(a.rad2), bx[6] + a.rmax_mobile - a.thr_min) + a.reach_extra;
                        float mn = 3e38f;
#pragma unroll 4
                        for (int s = 0; s < a.M; ++s) {
                            const float4 m = ligx[s];
                            const float ex = fmaxf(fmaxf(b0 - m.x, m.x - b3), 0.f);
                            const float ey = fmaxf(fmaxf(b1 - m.y, m.y - b4), 0.f);
                            const float ez = fmaxf(fmaxf(b2 - m.z, m.z - b5), 0.f);
                            mn = fminf(mn, fmaf(ez, ez, fmaf(ey, ey, ex * ex)));
                        }
                        keep = mn <= reach * reach;
                    }
                }
                const unsigned mask = __ballot_sync(0xffffffffu, keep);
                __syncwarp();
                if (keep) list[nout + __popc(mask & ((1u << lane) - 1u))] = (uint16_t)st;
                nout += __popc(mask);
                __syncwarp();
            }
            nl = nout;
        }
        if (S == 1) { if (lane == 0) nlist_p[0] = nl; }
        team_sync<S>();
        const int nlist = nlist_p[0];
        const int ngroups = (nlist + 3) >> 2;
        if (tlane < 4 && nlist + tlane < ngroups * 4) list[nlist + tlane] = (uint16_t)f.NST;   // dummy far leaf
        team_sync<S>();

        // ---- 2. sweep -----------------------------------------------------------------------
        int ncq = 0, npq = 0;
        const float near_c = a.rad2 + 2e-4f + 1e-6f * wmax, near_m = 1e-6f - 1.f;
        for (int gi = wt;; gi += S) {
            // one call site for the expansion keeps the kernel's code (and its instruction-cache
            // footprint) small: queue nearly full, or nothing left to sweep
            if (gi >= ngroups || ncq > kFCandQ - 128) {
                fused_drain_candidates<PLAIN>(a, cq, ncq, pq, npq, ligx, ligc, sph, zm, acc, lane);
                if (gi >= ngroups) break;
            }
            const uint2 st4 = *reinterpret_cast<const uint2*>(list + 4 * gi);
            const uint32_t stv[4] = {st4.x & 0xffffu, st4.x >> 16, st4.y & 0xffffu, st4.y >> 16};
            float ax[4], ay[4], az[4], nr[4], rw[4], ms[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float4 r = a.rec4[stv[k] * kSubTile + lane];
                const float x = r.x - sph.x, y = r.y - sph.y, z = r.z - sph.z;
                ax[k] = -2.f * x; ay[k] = -2.f * y; az[k] = -2.f * z;
                nr[k] = fmaf(z, z, fmaf(y, y, x * x));
                // candidate reach of this atom: its radius minus the overlap cutoff it can reach at
                // all (an atom that is not hydrophobic only takes part in clashes)
                const unsigned rf = a.rflags[stv[k] * kSubTile + lane];
                rw[k] = r.w - ((rf & GMM_F_HYDROPHOBIC) ? a.thr_min : a.thr_tight);
                // within reach of where a mobile atom that is not a zone seed (a flexible side chain) can
                // ever be, decided once per receptor (kRflagNearFlex): rides on the sign of rw
                if (!PLAIN && (rf & kRflagNearFlex)) rw[k] = -rw[k];
                ms[k] = 3e38f;
                swept_leaves += (stv[k] < (uint32_t)f.NST) ? 1 : 0;
            }
            // two mobile atoms per step: 6 FFMA + one 3-input FMNMX3 per receptor atom
            int s = 0;
            if (CULL) {
#pragma unroll 2
                for (; s + 1 < a.n_seed; s += 2) {
                    const float4 m = ligc[s], q = ligc[s + 1];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        ms[k] = fminf(ms[k], fminf(fmaf(ax[k], m.x, fmaf(ay[k], m.y, fmaf(az[k], m.z, m.w))),
                                                   fmaf(ax[k], q.x, fmaf(ay[k], q.y, fmaf(az[k], q.z, q.w)))));
                }
            } else {
#pragma unroll 4
                for (; s + 1 < a.n_seed; s += 2) {
                    const float4 m = ligc[s], q = ligc[s + 1];
#pragma unroll
                    for (int k = 0; k < 4; ++k)
                        ms[k] = fminf(ms[k], fminf(fmaf(ax[k], m.x, fmaf(ay[k], m.y, fmaf(az[k], m.z, m.w))),
                                                   fmaf(ax[k], q.x, fmaf(ay[k], q.y, fmaf(az[k], q.z, q.w)))));
                }
            }
            if (s < a.n_seed) {
                const float4 m = ligc[s];
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    ms[k] = fminf(ms[k], fmaf(ax[k], m.x, fmaf(ay[k], m.y, fmaf(az[k], m.z, m.w))));
            }
            // Far group (the common case without culling): no atom of these four leaves is within
            // radius + margin of a seed and none is always-in-zone -> their zone words stay as
            // initialised and there is nothing to expand.
            {
                // ms + nr <= rad2 + margin(nr)  <=>  ms <= near_c - near_m * nr
                bool near = false;
#pragma unroll
                for (int k = 0; k < 4; ++k) near = near || (ms[k] <= fmaf(near_m, nr[k], near_c));
                if (!PLAIN && a.n_sseed > 0) near = near || ((zone[stv[0]] | zone[stv[1]] | zone[stv[2]] | zone[stv[3]]) != 0u);
                if (!__any_sync(0xffffffffu, near)) continue;
            }
            bool in[4], band[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float margin = 2e-4f + 1e-6f * (nr[k] + wmax);
                const float dz = ms[k] + nr[k];
                in[k] = (!PLAIN && ((zone[stv[k]] >> lane) & 1u)) || dz <= a.rad2 - margin;
                band[k] = !in[k] && dz <= a.rad2 + margin;
            }
            if (__any_sync(0xffffffffu, band[0] || band[1] || band[2] || band[3])) {
                // rare (a handful of atoms per thousand): re-decide with exact differences, the same
                // arithmetic as the dump kernel.  One rolled copy of the loop for the four leaves.
#pragma unroll 1
                for (int k = 0; k < 4; ++k) {
                    const bool bk = k == 0 ? band[0] : k == 1 ? band[1] : k == 2 ? band[2] : band[3];
                    if (!bk) continue;
                    const uint32_t st = k == 0 ? stv[0] : k == 1 ? stv[1] : k == 2 ? stv[2] : stv[3];
                    const float4 r = a.rec4[st * kSubTile + lane];
                    float e = 3e38f;
#pragma unroll 2
                    for (int s = 0; s < a.n_seed; ++s)
                        e = fminf(e, dist2f(r.x, r.y, r.z, ligx[s].x, ligx[s].y, ligx[s].z));
                    const bool res = e <= a.rad2;
                    in[0] = k == 0 ? res : in[0]; in[1] = k == 1 ? res : in[1];
                    in[2] = k == 2 ? res : in[2]; in[3] = k == 3 ? res : in[3];
                }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float margin = 2e-4f + 1e-6f * (nr[k] + wmax);
                // candidate: in the zone and within contact reach of a seed, or flagged as within reach of
                // a side chain's range of motion (no sweep over the non-seed atoms: the expansion's
                // pre-filter sorts those out)
                const float c = (PLAIN ? rw[k] : fabsf(rw[k])) + a.rmax_mobile;
                const bool near_seed = ms[k] + nr[k] <= c * c * 1.000004f + 1e-6f + margin;
                const bool cand = in[k] && (near_seed || (!PLAIN && rw[k] < 0.f));
                const unsigned zb = __ballot_sync(0xffffffffu, in[k]);
                const unsigned cb = __ballot_sync(0xffffffffu, cand);
                __syncwarp();
                if (lane == 0 && stv[k] < (uint32_t)f.NST) zone[stv[k]] = zb;
                if (cb) {
                    if (cand) cq[ncq + __popc(cb & ((1u << lane) - 1u))] =
                        (stv[k] * kSubTile + lane) | ((!PLAIN && near_seed) ? 0x80000000u : 0u);
                    ncq += __popc(cb);
                }
            }
            __syncwarp();
        }
        fused_drain_pairs<PLAIN>(a, pq, npq, ligx, acc, lane);
        team_sync<S>();

        // ---- 3. static-static pairs with both atoms in the zone ------------------------------
        // Only swept leaves can have a non-empty zone word (leaves holding always-in-zone atoms are
        // never culled).  Lanes fetch the edge ranges of 32 leaves at once, then the warp
        // walks each non-empty leaf's edges with coalesced loads.
        if (f.rr_ptr != nullptr) {
            for (int base = wt * 32; base < nlist; base += TL) {
                const int i = base + lane;
                uint32_t my_z = 0;
                int my_e0 = 0, my_e1 = 0;
                if (i < nlist) {
                    const uint32_t my_st = list[i];
                    my_z = zone[my_st];
                    if (my_z) { my_e0 = f.rr_ptr[my_st * kSubTile]; my_e1 = f.rr_ptr[my_st * kSubTile + kSubTile]; }
                }
                unsigned todo = __ballot_sync(0xffffffffu, my_z != 0u && my_e1 > my_e0);
                while (todo) {
                    const int src = __ffs(todo) - 1;
                    todo &= todo - 1;
                    const uint32_t z = __shfl_sync(0xffffffffu, my_z, src);
                    const int e0 = __shfl_sync(0xffffffffu, my_e0, src), e1 = __shfl_sync(0xffffffffu, my_e1, src);
                    for (int e = e0 + lane; e < e1; e += 64) {      // two independent edges in flight
                        const int e2 = e + 32;
                        const uint4 en = f.edge[e];                 // the value rides with the edge: no
                        const uint4 en2 = e2 < e1 ? f.edge[e2] : make_uint4(0u, 0u, 0u, 0u);   // dependent gather
                        const uint32_t kind = en.x >> 29, al = (en.x >> 24) & 31u, b = en.x & 0x3fffffu;
                        const uint32_t kind2 = en2.x >> 29, al2 = (en2.x >> 24) & 31u, b2 = en2.x & 0x3fffffu;
                        if (kind && ((z >> al) & 1u) && ((zone[b >> 5] >> (b & 31u)) & 1u)) {
                            const long long v = (long long)(((unsigned long long)en.w << 32) | en.z);
                            if (kind & 1) { acc.cfx += v; ++acc.nc; } else { acc.hfx += v; ++acc.nh; }
                        }
                        if (kind2 && ((z >> al2) & 1u) && ((zone[b2 >> 5] >> (b2 & 31u)) & 1u)) {
                            const long long v = (long long)(((unsigned long long)en2.w << 32) | en2.z);
                            if (kind2 & 1) { acc.cfx += v; ++acc.nc; } else { acc.hfx += v; ++acc.nh; }
                        }
                    }
                }
            }
        }
    }

    // ---- 4. pairs inside the mobile set ------------------------------------------------------
    // (the next pair's record is fetched before this one is worked on: at the occupancy a large mobile
    // set leaves, nothing else hides the load)
    uint32_t pr_next = 0; float c2_next = 0.f;
    if (tlane < f.n_mm) { pr_next = f.mm_pairs[tlane]; c2_next = f.mm_c2[tlane]; }
    for (int e = tlane; e < f.n_mm; e += TL) {
        const uint32_t pr = pr_next;
        const float c2 = c2_next;                      // (r_a + r_b - cutoff)^2, fixed per pair
        if (e + TL < f.n_mm) { pr_next = f.mm_pairs[e + TL]; c2_next = f.mm_c2[e + TL]; }
        const int sa = pr & 0xffff, sb = pr >> 16;
        if (!PLAIN && !a.all_seed && (!((zm[sa >> 5] >> (sa & 31)) & 1u) || !((zm[sb >> 5] >> (sb & 31)) & 1u))) continue;
        const float4 A = ligx[sa], B = ligx[sb];
        if (dist2f(A.x, A.y, A.z, B.x, B.y, B.z) > c2) continue;
        const PairScore ps = score_pair(A.x, A.y, A.z, A.w, a.mflags[sa], B.x, B.y, B.z, B.w, a.mflags[sb], a.prm);
        if (ps.kind & 1) { acc.cfx += ps.clash_fx; ++acc.nc; }
        if (ps.kind & 2) { acc.hfx += ps.hydro_fx; ++acc.nh; }
    }

    // ---- 5. finish ---------------------------------------------------------------------------
    long long cfx = warp_sum(acc.cfx), hfx = warp_sum(acc.hfx);
    long long nc = __reduce_add_sync(0xffffffffu, acc.nc), nh = __reduce_add_sync(0xffffffffu, acc.nh);
    if (S > 1) {
        if (lane == 0) {
            atomicAdd((unsigned long long*)&accs[0], (unsigned long long)cfx);
            atomicAdd((unsigned long long*)&accs[1], (unsigned long long)hfx);
            atomicAdd((unsigned long long*)&accs[2], (unsigned long long)nc);
            atomicAdd((unsigned long long*)&accs[3], (unsigned long long)nh);
        }
        team_sync<S>();
        cfx = accs[0]; hfx = accs[1]; nc = accs[2]; nh = accs[3];
    }
    if (f.count_tests && lane == 0 && swept_leaves) atomicAdd(a.counters, (unsigned long long)swept_leaves * kSubTile * (PLAIN ? a.M : a.n_seed));
    if (tlane == 0) {
        if (f.count_tests && f.n_mm) atomicAdd(a.counters, (unsigned long long)f.n_mm);
        double clash = 2.0 * ((double)cfx / kFxScale);          // contacts.py:176-177
        const double hydro = 2.0 * ((double)hfx / kFxScale);
        if (a.prm.cutoff != 0.0 && clash > a.prm.cutoff) clash = -1000.0 * a.prm.weight;   // contacts.py:132-133
        f.scores[2 * (size_t)p] = clash;
        f.scores[2 * (size_t)p + 1] = hydro;
        if (f.counts) { f.counts[2 * (size_t)p] = (int32_t)nc; f.counts[2 * (size_t)p + 1] = (int32_t)nh; }
        if (f.mirror != nullptr) {
            // fitness exchange: the row goes to every rank's table now, while the other individuals
            // are still being scored; plain stores through the peer mappings.  The buffer is the one
            // of the epoch the signal kernel behind this one will publish (exchange.cu).
            const MirrorTable& mt = *f.mirror;
            const unsigned parity = (*reinterpret_cast<const unsigned*>(mt.base[mt.rank] + 4 * GMM_MAX_PEERS) + 1u) & 1u;
            const size_t buf = ((size_t)mt.max_rows * 24 + 255) / 256 * 256;
            const size_t row = (size_t)f.mirror_row0 + p;
#pragma unroll 1
            for (int r = 0; r < mt.n; ++r) {
                unsigned char* b = mt.base[r] + 256 + parity * buf;
                reinterpret_cast<double2*>(b)[row] = make_double2(clash, hydro);
                reinterpret_cast<int2*>(b + (size_t)mt.max_rows * 16)[row] = make_int2((int)nc, (int)nh);
            }
        }
    }
}

void fill_sweep_args(const gmm_ctx* c, int P, const float* xyz, SweepArgs& a);

static size_t fused_ind_bytes(const gmm_ctx* c) {
    const int NST = c->Rpad / kSubTile, MW = (c->Mp + 31) / 32;
    size_t ind = (size_t)2 * c->Mp * sizeof(float4) + 3 * sizeof(float4) + 4 * sizeof(long long) + 4 * sizeof(int) +
                 (size_t)(NST + 4) * 2 + (size_t)(NST + 1) * 4 + (size_t)(MW + 1) * 4;     // zm: one spare word
    return (ind + 15) / 16 * 16;
}
static const size_t kFusedQueues = (size_t)4 * (kFCandQ + kPairQ) * sizeof(uint32_t);
static const size_t kFusedSmemLimit = 200 * 1024;

// True when gmm_contacts / gmm_evaluate will run the one-kernel path for this context.
bool fused_applicable(const gmm_ctx* c) {
    if (c->prm.options & GMM_OPT_TILED) return false;
    if (c->Rpad / kSubTile >= 65535) return false;      // leaf ids are 16-bit in shared memory
    return fused_ind_bytes(c) + kFusedQueues <= kFusedSmemLimit;
}

// Returns GMM_OK and sets *handled = 1 when the fused kernel ran; *handled = 0 when the model
// does not fit its shared-memory plan (the caller then uses the tiled kernels of contacts.cu).
int launch_contacts_fused(gmm_ctx* c, int P, const float* xyz, double* scores, int32_t* counts, cudaStream_t s,
                          int* handled) {
    *handled = 0;
    if (!fused_applicable(c)) return GMM_OK;
    FusedArgs f;
    fill_sweep_args(c, P, xyz, f.s);
    f.NST = c->Rpad / kSubTile; f.R = c->R;
    f.sub_box = c->d_sub_box;
    f.rr_ptr = c->rr_nnz ? c->d_rr_ptr : nullptr; f.edge = c->d_edge;
    f.mm_pairs = c->d_mm_pairs; f.mm_c2 = c->d_mm_c2; f.n_mm = c->n_mm;
    f.scores = scores; f.counts = counts;
    f.count_tests = c->prof ? 1 : 0;
    f.mirror = (c->pub_row0 >= 0) ? c->d_mirror : nullptr;
    f.mirror_row0 = c->pub_row0;
    const size_t ind = fused_ind_bytes(c);
    f.ind_bytes = (int)ind;
    const bool plain = c->all_seed && c->n_sseed == 0 && c->ms_nnz == 0 && !(c->prm.options & GMM_OPT_ONLY_INTERNAL);
    const size_t pose_bytes = (size_t)c->M * 3 * sizeof(float);
    // PLAIN: whole 16-byte multiples that fit the queue area; otherwise the aligned span (at most
    // 12 M + 24 bytes) has to fit the centred copy's 16 Mp bytes
    f.tma_pose = ((reinterpret_cast<uintptr_t>(xyz) % 16) == 0 &&
                  (plain ? (pose_bytes % 16 == 0 && pose_bytes <= (kFCandQ + kPairQ) * sizeof(uint32_t)) : c->M >= 6)) ? 1 : 0;
    const size_t queues = kFusedQueues;
    const size_t limit = kFusedSmemLimit;
    // warp-per-individual from this population on; below it a CTA per individual fills the GPU better
    constexpr int kTeamSwitch = 148 * 7;        // measured (tools/latency_small_pop.py): 900 individuals 0.062 ms as teams, 0.067 as warps; 1,200: 0.075 / 0.071
    int S = (P >= kTeamSwitch) ? 1 : 4;
    if (S == 1 && 4 * ind + queues > limit) S = 4;
    if (ind + queues > limit) return GMM_OK;    // not handled
    const int warps = fused_warps(S, plain);
    const size_t smem = (size_t)(warps / S) * ind + queues * warps / 4;
    if (c->prof) GMM_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long), s));   // the pair-test tally
    const bool cull = f.s.cull != 0;
    void (*const variants[8])(const FusedArgs) = {
        contacts_fused_kernel<1, false, false>, contacts_fused_kernel<1, false, true>,
        contacts_fused_kernel<1, true, false>, contacts_fused_kernel<1, true, true>,
        contacts_fused_kernel<4, false, false>, contacts_fused_kernel<4, false, true>,
        contacts_fused_kernel<4, true, false>, contacts_fused_kernel<4, true, true>};
    void (*kernel)(const FusedArgs) = variants[(S == 1 ? 0 : 4) + (cull ? 2 : 0) + (plain ? 1 : 0)];
    { const int rc = allow_dynamic_smem((const void*)kernel, c->device, limit); if (rc) return rc; }
    if (c->prof) GMM_CUDA(cudaEventRecord(c->ev[2], s));
    kernel<<<(P + warps / S - 1) / (warps / S), warps * 32, smem, s>>>(f);
    GMM_LAUNCH_CHECK("contacts_fused_kernel");
    if (c->prof) { GMM_CUDA(cudaEventRecord(c->ev[3], s)); c->ev_used[1] = true; c->ev_used[2] = false; }
    c->pub_done = f.mirror != nullptr;
    *handled = 1;
    return GMM_OK;
}

}  // namespace gmm
