// Contacts objective for a whole population (gaudi/objectives/contacts.py:129-268 and the
// Chimera calls it makes: specifier.zone at :210-212, DetectClash.detectClash at :145).
//
// Per individual the reference builds the zone (probe atoms + CB seeds + everything within
// `radius` of them), runs an all-vs-all clash detection inside the zone, and scores the pairs.
// Here the same set of unordered pairs is found by three cooperating stages that live in
// ONE kernel (contacts_sweep_kernel) plus a small per-individual epilogue:
//
//   sweep      each lane keeps 4 receptor atoms of a 128-atom tile in registers and streams
//              the mobile atoms of G individuals past them from shared memory, tracking only
//              min d^2 per (individual, receptor atom): 3 FADD + FMUL + 2 FFMA + FMNMX per
//              pair test, branch-free.  The minimum over the seed atoms decides zone
//              membership exactly (d <= radius); the minimum over all mobile atoms flags the
//              few receptor atoms that can have a contact at all ("candidates").
//   expand     candidates are compacted into a per-warp queue; when it fills, every lane
//              takes one candidate and re-tests it against the individual's mobile atoms with
//              the exact per-pair cutoff r_i + r_j - threshold; hits go to a pair queue.
//   score      pairs are scored in fp64 with all 32 lanes busy (score_pair, common.cuh) and
//              added to per-individual 64-bit fixed-point sums, so the result is independent
//              of tiling, culling, grouping and GPU count (integer addition commutes).
//
//   finalize   one warp per individual: intra-mobile pairs, the static-static pairs whose
//              two atoms are both in this individual's zone (precomputed values gated by the
//              zone bitmask the sweep wrote), reduction of the per-chunk partial sums,
//              doubling (the reference counts every pair in both directions, contacts.py:
//              176-177), cutoff penalty (contacts.py:132-133).
//
// Receptor tiles are Morton-sorted; a tile is skipped for an individual when its bounding
// box is out of reach of the pose's bounding sphere (GMM_OPT_NO_CULL disables this).
#include "contacts_common.cuh"

namespace gmm {

// ---------------------------------------------------------------------------------------
// sweep + expand + score
// ---------------------------------------------------------------------------------------
struct WarpQueues {
    uint32_t* cq; int ncq;     // candidates: g << 24 | (sorted static index)
    uint2* pq; int npq;        // pairs: x = g << 16 | slot, y = static index
};

__device__ __forceinline__ void drain_pairs(const SweepArgs& a, WarpQueues& q, const float4* lig,
                                            long long* acc, int lane) {
    for (int base = 0; base < q.npq; base += 32) {
        const int e = base + lane;
        if (e < q.npq) {
            const uint2 pr = q.pq[e];
            const int g = pr.x >> 16, s = pr.x & 0xffff, j = (int)pr.y;
            if (!ms_excluded(a.ms_ptr, a.ms_idx, s, j)) {
                const float4 m = lig[g * a.Mp + s];
                const float4 r = a.rec4[j];
                const PairScore ps = score_pair(m.x, m.y, m.z, m.w, a.mflags[s],
                                                r.x, r.y, r.z, r.w, a.rflags[j], a.prm);
                if (ps.kind & 1) {
                    atomicAdd((unsigned long long*)&acc[g * 4 + 0], (unsigned long long)ps.clash_fx);
                    atomicAdd((unsigned long long*)&acc[g * 4 + 2], 1ull);
                }
                if (ps.kind & 2) {
                    atomicAdd((unsigned long long*)&acc[g * 4 + 1], (unsigned long long)ps.hydro_fx);
                    atomicAdd((unsigned long long*)&acc[g * 4 + 3], 1ull);
                }
            }
        }
    }
    __syncwarp();
    q.npq = 0;
}

__device__ __forceinline__ void drain_candidates(const SweepArgs& a, WarpQueues& q, const float4* lig,
                                                 const uint32_t* zm, long long* acc, int lane) {
    for (int base = 0; base < q.ncq; base += 32) {
        const int e = base + lane;
        const bool valid = e < q.ncq;
        int g = 0, j = 0;
        float4 r = make_float4(kFar, kFar, kFar, 0.f);
        if (valid) {
            const uint32_t c = q.cq[e];
            g = c >> 24; j = c & 0xffffff;
            r = a.rec4[j];
        }
        const float4* L = lig + g * a.Mp;
        const uint32_t* Z = zm + g * a.MW;
        for (int s = 0; s < a.M; ++s) {
            const float4 m = L[s];
            const float d2 = dist2f(r.x, r.y, r.z, m.x, m.y, m.z);
            const bool hit = valid && d2 <= loose_c2(m.w, r.w, a.thr_min) && ((Z[s >> 5] >> (s & 31)) & 1u);
            const unsigned mask = __ballot_sync(0xffffffffu, hit);
            if (mask) {
                if (hit) q.pq[q.npq + __popc(mask & ((1u << lane) - 1u))] = make_uint2((g << 16) | s, j);
                q.npq += __popc(mask);
                __syncwarp();
                if (q.npq > kPairQ - 32) drain_pairs(a, q, lig, acc, lane);
            }
        }
    }
    __syncwarp();
    q.ncq = 0;
    drain_pairs(a, q, lig, acc, lane);
}

__global__ void __launch_bounds__(kSweepWarps * 32)
contacts_sweep_kernel(const SweepArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* lig = reinterpret_cast<float4*>(smem_raw);                       // [G][Mp]
    float4* sphere = lig + kGroup * a.Mp;                                    // [G]
    long long* acc = reinterpret_cast<long long*>(sphere + kGroup);          // [G][4]
    uint2* pq_all = reinterpret_cast<uint2*>(acc + kGroup * 4);              // [W][kPairQ]
    uint32_t* cq_all = reinterpret_cast<uint32_t*>(pq_all + kSweepWarps * kPairQ);  // [W][kCandQ]
    uint32_t* zm = cq_all + kSweepWarps * kCandQ;                            // [G][MW]

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = blockIdx.x, p0 = blockIdx.y * kGroup;
    const int Gn = min(kGroup, a.P - p0);

    if (threadIdx.x < kGroup * 4) acc[threadIdx.x] = 0;
    load_ligands<kSweepWarps * 32>(a, p0, Gn, lig, zm);

    // bounding sphere of every pose (all mobile atoms), one warp per individual
    for (int g = warp; g < Gn; g += kSweepWarps) {
        float lo[3] = {3e38f, 3e38f, 3e38f}, hi[3] = {-3e38f, -3e38f, -3e38f};
        for (int s = lane; s < a.M; s += 32) {
            const float4 m = lig[g * a.Mp + s];
            lo[0] = fminf(lo[0], m.x); hi[0] = fmaxf(hi[0], m.x);
            lo[1] = fminf(lo[1], m.y); hi[1] = fmaxf(hi[1], m.y);
            lo[2] = fminf(lo[2], m.z); hi[2] = fmaxf(hi[2], m.z);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
                hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
            }
        const float cx = 0.5f * (lo[0] + hi[0]), cy = 0.5f * (lo[1] + hi[1]), cz = 0.5f * (lo[2] + hi[2]);
        float r2 = 0.f;
        for (int s = lane; s < a.M; s += 32) {
            const float4 m = lig[g * a.Mp + s];
            r2 = fmaxf(r2, dist2f(m.x, m.y, m.z, cx, cy, cz));
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) r2 = fmaxf(r2, __shfl_xor_sync(0xffffffffu, r2, o));
        if (lane == 0) sphere[g] = make_float4(cx, cy, cz, sqrtf(r2) * 1.00001f + 1e-4f);
    }
    __syncthreads();

    WarpQueues q;
    q.cq = cq_all + warp * kCandQ; q.ncq = 0;
    q.pq = pq_all + warp * kPairQ; q.npq = 0;
    unsigned long long tests = 0;

    const int t_begin = chunk * a.tiles_per_chunk;
    const int t_end = min(t_begin + a.tiles_per_chunk, a.n_tiles);
    for (int tile = t_begin + warp; tile < t_end; tile += kSweepWarps) {
        float rx[4], ry[4], rz[4], cc2[4];
        bool alw[4];
        const uint32_t* aw = a.always + tile * 4;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const float4 r = a.rec4[tile * kTileAtoms + k * 32 + lane];
            rx[k] = r.x; ry[k] = r.y; rz[k] = r.z;
            const float c = r.w + a.rmax_mobile - a.thr_min;
            cc2[k] = c * c * 1.000004f + 1e-6f;
            alw[k] = (aw[k] >> lane) & 1u;
        }
        const float* box = a.tile_box + tile * 8;
        const float blx = box[0], bly = box[1], blz = box[2], bhx = box[3], bhy = box[4], bhz = box[5];
        const float reach = fmaxf(sqrtf(a.rad2), box[6] + a.rmax_mobile - a.thr_min) + a.reach_extra;

        for (int g = 0; g < Gn; ++g) {
            uint32_t* zrow = a.zone + (size_t)(p0 + g) * a.ZW + tile * 4;
            if (a.cull) {
                const float4 sp = sphere[g];
                const float ex = fmaxf(fmaxf(blx - sp.x, sp.x - bhx), 0.f);
                const float ey = fmaxf(fmaxf(bly - sp.y, sp.y - bhy), 0.f);
                const float ez = fmaxf(fmaxf(blz - sp.z, sp.z - bhz), 0.f);
                const float lim = sp.w + reach;
                if (ex * ex + ey * ey + ez * ez > lim * lim) {
                    if (lane < 4) zrow[lane] = aw[lane];
                    continue;
                }
            }
            tests += (unsigned long long)kTileAtoms * a.M;
            const float4* L = lig + g * a.Mp;
            float m0 = 3e38f, m1 = 3e38f, m2 = 3e38f, m3 = 3e38f;
#pragma unroll 4
            for (int s = 0; s < a.n_seed; ++s) {
                const float4 v = L[s];
                m0 = fminf(m0, dist2f(rx[0], ry[0], rz[0], v.x, v.y, v.z));
                m1 = fminf(m1, dist2f(rx[1], ry[1], rz[1], v.x, v.y, v.z));
                m2 = fminf(m2, dist2f(rx[2], ry[2], rz[2], v.x, v.y, v.z));
                m3 = fminf(m3, dist2f(rx[3], ry[3], rz[3], v.x, v.y, v.z));
            }
            const bool z0 = alw[0] || m0 <= a.rad2, z1 = alw[1] || m1 <= a.rad2;
            const bool z2 = alw[2] || m2 <= a.rad2, z3 = alw[3] || m3 <= a.rad2;
            for (int s = a.n_seed; s < a.M; ++s) {
                const float4 v = L[s];
                m0 = fminf(m0, dist2f(rx[0], ry[0], rz[0], v.x, v.y, v.z));
                m1 = fminf(m1, dist2f(rx[1], ry[1], rz[1], v.x, v.y, v.z));
                m2 = fminf(m2, dist2f(rx[2], ry[2], rz[2], v.x, v.y, v.z));
                m3 = fminf(m3, dist2f(rx[3], ry[3], rz[3], v.x, v.y, v.z));
            }
            const unsigned zb0 = __ballot_sync(0xffffffffu, z0), zb1 = __ballot_sync(0xffffffffu, z1);
            const unsigned zb2 = __ballot_sync(0xffffffffu, z2), zb3 = __ballot_sync(0xffffffffu, z3);
            if (lane == 0) { zrow[0] = zb0; zrow[1] = zb1; zrow[2] = zb2; zrow[3] = zb3; }
            const bool c[4] = {z0 && m0 <= cc2[0], z1 && m1 <= cc2[1], z2 && m2 <= cc2[2], z3 && m3 <= cc2[3]};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const unsigned mask = __ballot_sync(0xffffffffu, c[k]);
                if (mask) {
                    if (c[k]) q.cq[q.ncq + __popc(mask & ((1u << lane) - 1u))] =
                        ((uint32_t)g << 24) | (uint32_t)(tile * kTileAtoms + k * 32 + lane);
                    q.ncq += __popc(mask);
                }
            }
            __syncwarp();
            if (q.ncq > kCandQ - kTileAtoms) drain_candidates(a, q, lig, zm, acc, lane);
        }
    }
    drain_candidates(a, q, lig, zm, acc, lane);
    if (lane == 0 && tests) atomicAdd(a.counters, tests);
    __syncthreads();
    if (threadIdx.x < Gn * 4) {
        const int g = threadIdx.x >> 2, k = threadIdx.x & 3;
        a.partial[((size_t)(p0 + g) * a.n_chunks + chunk) * 4 + k] = acc[g * 4 + k];
    }
}

// ---------------------------------------------------------------------------------------
// finalize: intra-mobile pairs + static-static pairs in zone + reduction + penalty
// ---------------------------------------------------------------------------------------
constexpr int kFinWarps = 4;

struct FinArgs {
    SweepArgs s;
    const uint32_t* mm_pairs; int n_mm;
    const int32_t* rr_ptr; const int32_t* rr_idx; const long long* rr_val; const uint8_t* rr_kind;
    double* scores; int32_t* counts;
    int have_sweep;
};

__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__global__ void __launch_bounds__(kFinWarps * 32)
contacts_finalize_kernel(const FinArgs f) {
    const SweepArgs& a = f.s;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* lig = reinterpret_cast<float4*>(smem_raw);                    // [kFinWarps][Mp]
    uint32_t* zm = reinterpret_cast<uint32_t*>(lig + kFinWarps * a.Mp);   // [kFinWarps][MW]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int p0 = blockIdx.x * kFinWarps;
    const int Gn = min(kFinWarps, a.P - p0);
    load_ligands<kFinWarps * 32>(a, p0, Gn, lig, zm);
    if (warp >= Gn) return;
    const int p = p0 + warp;
    const float4* L = lig + warp * a.Mp;
    const uint32_t* Z = zm + warp * a.MW;
    long long cfx = 0, hfx = 0, nc = 0, nh = 0;

    // intra-mobile pairs (exclusions already removed from the list)
    for (int e = lane; e < f.n_mm; e += 32) {
        const uint32_t pr = f.mm_pairs[e];
        const int sa = pr & 0xffff, sb = pr >> 16;
        if (!((Z[sa >> 5] >> (sa & 31)) & 1u) || !((Z[sb >> 5] >> (sb & 31)) & 1u)) continue;
        const float4 A = L[sa], B = L[sb];
        if (dist2f(A.x, A.y, A.z, B.x, B.y, B.z) > loose_c2(A.w, B.w, a.thr_min)) continue;
        const PairScore ps = score_pair(A.x, A.y, A.z, A.w, a.mflags[sa], B.x, B.y, B.z, B.w, a.mflags[sb], a.prm);
        if (ps.kind & 1) { cfx += ps.clash_fx; ++nc; }
        if (ps.kind & 2) { hfx += ps.hydro_fx; ++nh; }
    }
    if (f.have_sweep) {
        // static-static pairs with both atoms in this individual's zone
        const uint32_t* zone = a.zone + (size_t)p * a.ZW;
        if (f.rr_ptr != nullptr) {
            for (int w = lane; w < a.ZW; w += 32) {
                uint32_t bits = zone[w];
                while (bits) {
                    const int b = __ffs(bits) - 1;
                    bits &= bits - 1;
                    const int ia = w * 32 + b;
                    for (int e = f.rr_ptr[ia]; e < f.rr_ptr[ia + 1]; ++e) {
                        const int ib = f.rr_idx[e];
                        if ((zone[ib >> 5] >> (ib & 31)) & 1u) {
                            const int kind = f.rr_kind[e];
                            if (kind & 1) { cfx += f.rr_val[2 * (size_t)e]; ++nc; }
                            if (kind & 2) { hfx += f.rr_val[2 * (size_t)e + 1]; ++nh; }
                        }
                    }
                }
            }
        }
        // partial sums of the sweep chunks
        const long long* part = a.partial + (size_t)p * a.n_chunks * 4;
        for (int c = lane; c < a.n_chunks; c += 32) {
            cfx += part[c * 4 + 0]; hfx += part[c * 4 + 1]; nc += part[c * 4 + 2]; nh += part[c * 4 + 3];
        }
    }
    cfx = warp_sum_ll(cfx); hfx = warp_sum_ll(hfx); nc = warp_sum_ll(nc); nh = warp_sum_ll(nh);
    if (lane == 0) {
        // every pair is seen from both of its atoms (contacts.py:176-177)
        double clash = 2.0 * ((double)cfx / kFxScale);
        const double hydro = 2.0 * ((double)hfx / kFxScale);
        if (a.prm.cutoff != 0.0 && clash > a.prm.cutoff) clash = -1000.0 * a.prm.weight;  // contacts.py:132-133
        f.scores[2 * (size_t)p] = clash;
        f.scores[2 * (size_t)p + 1] = hydro;
        if (f.counts) { f.counts[2 * (size_t)p] = (int32_t)nc; f.counts[2 * (size_t)p + 1] = (int32_t)nh; }
    }
    if (p == 0 && lane == 0 && f.n_mm) atomicAdd(a.counters, (unsigned long long)f.n_mm * a.P);
}

// ---------------------------------------------------------------------------------------
// static-static pair values and always-in-zone atoms (run once per gmm_set_receptor)
// ---------------------------------------------------------------------------------------
__global__ void rr_values_kernel(int R, const int32_t* rr_ptr, const int32_t* rr_idx, const float4* rec4,
                                 const uint8_t* rflags, Params prm, long long* val, uint8_t* kind, uint4* edge) {
    const int ia = blockIdx.x * blockDim.x + threadIdx.x;
    if (ia >= R) return;
    const float4 A = rec4[ia];
    const unsigned fa = rflags[ia];
    for (int e = rr_ptr[ia]; e < rr_ptr[ia + 1]; ++e) {
        const int ib = rr_idx[e];
        const float4 B = rec4[ib];
        const PairScore ps = score_pair(A.x, A.y, A.z, A.w, fa, B.x, B.y, B.z, B.w, rflags[ib], prm);
        val[2 * (size_t)e] = ps.clash_fx;
        val[2 * (size_t)e + 1] = ps.hydro_fx;
        kind[e] = (uint8_t)ps.kind;
        // the two scores exclude each other (overlap above / not above the clash threshold)
        const unsigned long long v = (unsigned long long)((ps.kind & 1) ? ps.clash_fx : ps.hydro_fx);
        edge[e] = make_uint4((uint32_t)ib | ((uint32_t)(ia & 31) << 24) | ((uint32_t)ps.kind << 29), 0u,
                             (uint32_t)v, (uint32_t)(v >> 32));
    }
}

__global__ void always_zone_kernel(int Rpad, int R, const float4* rec4, const uint8_t* rflags,
                                   const float4* sseed, int n_sseed, float rad2, int only_internal,
                                   uint32_t* always) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;   // blockDim multiple of 32
    bool in = false;
    if (j < R) {
        in = rflags[j] & GMM_F_SEED;
        if (!in && !only_internal) {
            const float4 q = rec4[j];
            for (int k = 0; k < n_sseed && !in; ++k) {
                const float4 t = sseed[k];
                in = dist2f(t.x, t.y, t.z, q.x, q.y, q.z) <= rad2;
            }
        }
    }
    const unsigned bits = __ballot_sync(0xffffffffu, in);
    if ((threadIdx.x & 31) == 0 && j < Rpad) always[j >> 5] = bits;
}

// ---------------------------------------------------------------------------------------
// pair dump of one individual: an independent brute-force walk over the same definitions
// ---------------------------------------------------------------------------------------
struct DumpArgs {
    SweepArgs s;
    const uint32_t* mm_pairs; int n_mm;
    const int32_t* rr_ptr; const int32_t* rr_idx;
    const int32_t* sorted2orig;
    uint32_t* zone_tmp;           // [ZW]
    int R, max_pairs;
    int32_t* out_pairs; int32_t* out_kind; double* out_ov; double* out_val; int* out_n;
};

__device__ __forceinline__ void emit_pair(const DumpArgs& d, int ida, int idb, const PairScore& ps) {
    if (!ps.kind) return;
    const int idx = atomicAdd(d.out_n, 1);
    if (idx < d.max_pairs) {
        d.out_pairs[2 * idx] = min(ida, idb);
        d.out_pairs[2 * idx + 1] = max(ida, idb);
        d.out_kind[idx] = (ps.kind & 1) ? 0 : 1;
        d.out_ov[idx] = ps.overlap;
        d.out_val[idx] = (ps.kind & 1) ? ps.clash_val : ps.hydro_val;
    }
}

__global__ void __launch_bounds__(256)
contacts_dump_kernel(const DumpArgs d) {
    const SweepArgs& a = d.s;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float4* lig = reinterpret_cast<float4*>(smem_raw);
    uint32_t* zm = reinterpret_cast<uint32_t*>(lig + a.Mp);
    load_ligands<256>(a, 0, 1, lig, zm);
    const bool internal = a.prm.options & GMM_OPT_ONLY_INTERNAL;
    // zone of static atoms
    for (int w = threadIdx.x; w < a.ZW; w += 256) {
        uint32_t bits = a.always[w];
        if (!internal) {
            for (int b = 0; b < 32; ++b) {
                const int j = w * 32 + b;
                if (j >= d.R) break;
                const float4 r = a.rec4[j];
                float m = 3e38f;
                for (int s = 0; s < a.n_seed; ++s)
                    m = fminf(m, dist2f(r.x, r.y, r.z, lig[s].x, lig[s].y, lig[s].z));
                if (m <= a.rad2) bits |= 1u << b;
            }
        }
        d.zone_tmp[w] = bits;
    }
    __syncthreads();
    // mobile - static
    for (int j = threadIdx.x; j < d.R; j += 256) {
        if (!((d.zone_tmp[j >> 5] >> (j & 31)) & 1u)) continue;
        const float4 r = a.rec4[j];
        for (int s = 0; s < a.M; ++s) {
            if (!((zm[s >> 5] >> (s & 31)) & 1u)) continue;
            if (ms_excluded(a.ms_ptr, a.ms_idx, s, j)) continue;
            const float4 m = lig[s];
            emit_pair(d, a.slot2atom[s], a.M + d.sorted2orig[j],
                      score_pair(m.x, m.y, m.z, m.w, a.mflags[s], r.x, r.y, r.z, r.w, a.rflags[j], a.prm));
        }
    }
    // mobile - mobile
    for (int e = threadIdx.x; e < d.n_mm; e += 256) {
        const uint32_t pr = d.mm_pairs[e];
        const int sa = pr & 0xffff, sb = pr >> 16;
        if (!((zm[sa >> 5] >> (sa & 31)) & 1u) || !((zm[sb >> 5] >> (sb & 31)) & 1u)) continue;
        const float4 A = lig[sa], B = lig[sb];
        emit_pair(d, a.slot2atom[sa], a.slot2atom[sb],
                  score_pair(A.x, A.y, A.z, A.w, a.mflags[sa], B.x, B.y, B.z, B.w, a.mflags[sb], a.prm));
    }
    // static - static
    if (d.rr_ptr != nullptr) {
        for (int ia = threadIdx.x; ia < d.R; ia += 256) {
            if (!((d.zone_tmp[ia >> 5] >> (ia & 31)) & 1u)) continue;
            const float4 A = a.rec4[ia];
            for (int e = d.rr_ptr[ia]; e < d.rr_ptr[ia + 1]; ++e) {
                const int ib = d.rr_idx[e];
                if (!((d.zone_tmp[ib >> 5] >> (ib & 31)) & 1u)) continue;
                const float4 B = a.rec4[ib];
                emit_pair(d, a.M + d.sorted2orig[ia], a.M + d.sorted2orig[ib],
                          score_pair(A.x, A.y, A.z, A.w, a.rflags[ia], B.x, B.y, B.z, B.w, a.rflags[ib], a.prm));
            }
        }
    }
}

// ---------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------
static void plan_chunks(const gmm_ctx* c, int P, int* tiles_per_chunk, int* n_chunks) {
    const int groups = (P + kGroup - 1) / kGroup;
    const int want_ctas = 148 * 12;
    int nch = (want_ctas + groups - 1) / groups;
    const int max_ch = (c->n_tiles + kSweepWarps - 1) / kSweepWarps;
    if (nch > max_ch) nch = max_ch;
    if (nch < 1) nch = 1;
    int tpc = (c->n_tiles + nch - 1) / nch;
    tpc = ((tpc + kSweepWarps - 1) / kSweepWarps) * kSweepWarps;
    if (tpc < kSweepWarps) tpc = kSweepWarps;
    nch = (c->n_tiles + tpc - 1) / tpc;
    if (nch < 1) nch = 1;
    *tiles_per_chunk = tpc; *n_chunks = nch;
}

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// scratch layout: [coords P*M*3 f32] and, for the tiled path only, [zone P*ZW u32][partial P*n_chunks*4 i64]
size_t contacts_scratch_bytes(const gmm_ctx* c, int P) {
    const size_t ZW = (size_t)c->n_tiles * 4;
    size_t b = align_up((size_t)P * c->M * 3 * sizeof(float), 256);
    if (fused_applicable(c)) return b + 256;      // the one-kernel path only parks coordinates here
    b += align_up((size_t)P * ZW * sizeof(uint32_t), 256);
    // P * n_chunks(P) <= kGroup * want_ctas + P for every P (see plan_chunks): monotone bound
    b += align_up(((size_t)kGroup * 148 * 12 + kGroup + (size_t)P) * 4 * sizeof(long long), 256);
    return b + 256;
}

// Pre-filter cutoffs on the overlap r_a + r_b - d: a hydrophobic contact needs both atoms
// hydrophobic and overlap >= hydro_thr, a clash needs overlap > clash_thr (after an H-bond
// allowance that only ever lowers the overlap when it is positive).  thr_min admits either,
// thr_tight is enough for a pair that cannot be a hydrophobic contact.
void prefilter_cutoffs(const gmm_ctx* c, float* thr_min, float* thr_tight) {
    *thr_tight = (float)(c->prm.clash_thr + fmin(0.0, c->prm.hb_allow));
    *thr_min = fminf((float)c->prm.hydro_thr, *thr_tight);
}

void fill_sweep_args(const gmm_ctx* c, int P, const float* xyz, SweepArgs& a) {
    a.P = P; a.M = c->M; a.Mp = c->Mp; a.MW = (c->Mp + 31) / 32; a.n_seed = c->n_seed; a.n_hseed = c->n_hseed;
    a.n_tiles = c->n_tiles; a.ZW = c->n_tiles * 4;
    a.n_sseed = c->n_sseed; a.sseed = c->d_sseed;
    a.cull = (c->prm.options & GMM_OPT_NO_CULL) ? 0 : 1;
    a.all_seed = c->all_seed;
    a.xyz = xyz; a.slot2atom = c->d_slot2atom; a.mrad = c->d_mrad_slot; a.mflags = c->d_mflags_slot;
    a.rec4 = c->d_rec4; a.rflags = c->d_rflags; a.tile_box = c->d_tile_box; a.always = c->d_always;
    a.ms_ptr = c->ms_nnz ? c->d_ms_ptr : nullptr; a.ms_idx = c->d_ms_idx;
    a.flex_mask = c->d_flex_mask; a.body_range = c->d_body_range; a.n_bodies = c->n_bodies;
    a.counters = c->d_counters;
    a.prm = c->prm;
    a.rad2 = (float)(c->prm.zone_radius * c->prm.zone_radius);
    prefilter_cutoffs(c, &a.thr_min, &a.thr_tight);
    a.rmax_mobile = c->rmax_mobile;
    a.reach_extra = 2e-3f;
    plan_chunks(c, P, &a.tiles_per_chunk, &a.n_chunks);
}

int launch_contacts(gmm_ctx* c, int P, const float* xyz, double* scores, int32_t* counts, cudaStream_t s,
                    void* workspace) {
    if (P == 0) return GMM_OK;
    if (!(c->prm.options & GMM_OPT_TILED)) {
        int handled = 0;
        const int rc = launch_contacts_fused(c, P, xyz, scores, counts, s, &handled);
        if (rc != GMM_OK || handled) return rc;
    }
    SweepArgs a;
    fill_sweep_args(c, P, xyz, a);
    unsigned char* base = static_cast<unsigned char*>(workspace ? workspace : c->d_scratch);
    size_t off = align_up((size_t)P * c->M * 3 * sizeof(float), 256);
    a.zone = reinterpret_cast<uint32_t*>(base + off);
    off += align_up((size_t)P * a.ZW * sizeof(uint32_t), 256);
    a.partial = reinterpret_cast<long long*>(base + off);

    GMM_CUDA(cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long), s));
    const bool internal = c->prm.options & GMM_OPT_ONLY_INTERNAL;
    const bool have_sweep = !internal && c->R > 0;
    if (have_sweep) {
        const size_t smem = (size_t)kGroup * a.Mp * sizeof(float4) + kGroup * sizeof(float4) +
                            kGroup * 4 * sizeof(long long) + kSweepWarps * kPairQ * sizeof(uint2) +
                            kSweepWarps * kCandQ * sizeof(uint32_t) + (size_t)kGroup * a.MW * sizeof(uint32_t);
        if (smem > 48 * 1024) {
            const int rc = allow_dynamic_smem((const void*)contacts_sweep_kernel, c->device, 200 * 1024);
            if (rc) return rc;
        }
        dim3 grid(a.n_chunks, (P + kGroup - 1) / kGroup);
        if (c->prof) GMM_CUDA(cudaEventRecord(c->ev[2], s));
        contacts_sweep_kernel<<<grid, kSweepWarps * 32, smem, s>>>(a);
        GMM_LAUNCH_CHECK("contacts_sweep_kernel");
        if (c->prof) { GMM_CUDA(cudaEventRecord(c->ev[3], s)); c->ev_used[1] = true; }
    }
    FinArgs f;
    f.s = a; f.mm_pairs = c->d_mm_pairs; f.n_mm = c->n_mm;
    f.rr_ptr = c->rr_nnz ? c->d_rr_ptr : nullptr; f.rr_idx = c->d_rr_idx;
    f.rr_val = c->d_rr_val; f.rr_kind = c->d_rr_kind;
    f.scores = scores; f.counts = counts; f.have_sweep = have_sweep;
    const size_t fsmem = (size_t)kFinWarps * a.Mp * sizeof(float4) + (size_t)kFinWarps * a.MW * sizeof(uint32_t);
    if (fsmem > 48 * 1024) {
        const int rc = allow_dynamic_smem((const void*)contacts_finalize_kernel, c->device, 200 * 1024);
        if (rc) return rc;
    }
    if (c->prof) GMM_CUDA(cudaEventRecord(c->ev[4], s));
    contacts_finalize_kernel<<<(P + kFinWarps - 1) / kFinWarps, kFinWarps * 32, fsmem, s>>>(f);
    GMM_LAUNCH_CHECK("contacts_finalize_kernel");
    if (c->prof) { GMM_CUDA(cudaEventRecord(c->ev[5], s)); c->ev_used[2] = true; }
    return GMM_OK;
}

int launch_rr_values(gmm_ctx* c, cudaStream_t s) {
    const float4* sseed = c->d_sseed;
    const int n_sseed = c->n_sseed;
    if (c->Rpad > 0) {
        always_zone_kernel<<<c->Rpad / 128, 128, 0, s>>>(
            c->Rpad, c->R, c->d_rec4, c->d_rflags, sseed, n_sseed,
            (float)(c->prm.zone_radius * c->prm.zone_radius),
            (c->prm.options & GMM_OPT_ONLY_INTERNAL) ? 1 : 0, c->d_always);
        GMM_LAUNCH_CHECK("always_zone_kernel");
    }
    if (c->rr_nnz > 0) {
        rr_values_kernel<<<(c->R + 127) / 128, 128, 0, s>>>(c->R, c->d_rr_ptr, c->d_rr_idx, c->d_rec4,
                                                           c->d_rflags, c->prm, c->d_rr_val, c->d_rr_kind, c->d_edge);
        GMM_LAUNCH_CHECK("rr_values_kernel");
    }
    return GMM_OK;
}

int launch_pairs_dump(gmm_ctx* c, const float* xyz_one, int max_pairs, int32_t* pairs_h, int32_t* kind_h,
                      double* ov_h, double* val_h, int* n_out, cudaStream_t s) {
    DumpArgs d;
    fill_sweep_args(c, 1, xyz_one, d.s);
    d.mm_pairs = c->d_mm_pairs; d.n_mm = c->n_mm;
    d.rr_ptr = c->rr_nnz ? c->d_rr_ptr : nullptr; d.rr_idx = c->d_rr_idx;
    d.R = c->R; d.max_pairs = max_pairs;
    // temporaries of this diagnostic call; freed on every path out of the function
    struct Scratch {
        void* p[7] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
        ~Scratch() { for (void* q : p) if (q) cudaFree(q); }
    } tmp;
    const size_t zw = (size_t)(c->n_tiles * 4 + 1);
    GMM_CUDA(cudaMalloc(&tmp.p[0], sizeof(int32_t) * (c->R + 1)));
    GMM_CUDA(cudaMalloc(&tmp.p[1], sizeof(uint32_t) * zw));
    GMM_CUDA(cudaMalloc(&tmp.p[2], sizeof(int32_t) * 2 * (max_pairs + 1)));
    GMM_CUDA(cudaMalloc(&tmp.p[3], sizeof(int32_t) * (max_pairs + 1)));
    GMM_CUDA(cudaMalloc(&tmp.p[4], sizeof(double) * (max_pairs + 1)));
    GMM_CUDA(cudaMalloc(&tmp.p[5], sizeof(double) * (max_pairs + 1)));
    GMM_CUDA(cudaMalloc(&tmp.p[6], sizeof(int)));
    int32_t* d_s2o = static_cast<int32_t*>(tmp.p[0]); uint32_t* d_zone = static_cast<uint32_t*>(tmp.p[1]);
    int32_t* d_pairs = static_cast<int32_t*>(tmp.p[2]); int32_t* d_kind = static_cast<int32_t*>(tmp.p[3]);
    double* d_ov = static_cast<double*>(tmp.p[4]); double* d_val = static_cast<double*>(tmp.p[5]);
    int* d_n = static_cast<int*>(tmp.p[6]);
    GMM_CUDA(cudaMemsetAsync(d_n, 0, sizeof(int), s));
    if (c->R) GMM_CUDA(cudaMemcpyAsync(d_s2o, c->h_sorted2orig.data(), sizeof(int32_t) * c->R, cudaMemcpyHostToDevice, s));
    d.sorted2orig = d_s2o; d.zone_tmp = d_zone;
    d.out_pairs = d_pairs; d.out_kind = d_kind; d.out_ov = d_ov; d.out_val = d_val; d.out_n = d_n;
    const size_t smem = (size_t)d.s.Mp * sizeof(float4) + (size_t)d.s.MW * sizeof(uint32_t);
    contacts_dump_kernel<<<1, 256, smem, s>>>(d);
    GMM_LAUNCH_CHECK("contacts_dump_kernel");
    int n = 0;
    GMM_CUDA(cudaMemcpyAsync(&n, d_n, sizeof(int), cudaMemcpyDeviceToHost, s));
    GMM_CUDA(cudaStreamSynchronize(s));
    const int w = n < max_pairs ? n : max_pairs;
    if (w > 0) {
        GMM_CUDA(cudaMemcpy(pairs_h, d_pairs, sizeof(int32_t) * 2 * w, cudaMemcpyDeviceToHost));
        GMM_CUDA(cudaMemcpy(kind_h, d_kind, sizeof(int32_t) * w, cudaMemcpyDeviceToHost));
        GMM_CUDA(cudaMemcpy(ov_h, d_ov, sizeof(double) * w, cudaMemcpyDeviceToHost));
        GMM_CUDA(cudaMemcpy(val_h, d_val, sizeof(double) * w, cudaMemcpyDeviceToHost));
    }
    *n_out = n;
    return GMM_OK;
}

}  // namespace gmm
