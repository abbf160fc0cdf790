// Similarity bitmask on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
// Same idea as the mma.sync kernel in rmsd.cu (gaudi/similarity.py:62-75 decided for whole tiles of
// pairs: a . b over the re-centred fp16 conformers is a pre-filter with a proven error bound, the
// pairs inside the bound are re-decided by the exact sequential fp32 sum, so the bits equal the FP32
// kernel's), built the Blackwell way:
//
//   operands   one elected thread streams 128-row x 64-half blocks of both operand tiles into
//              shared memory with 2-D TMA loads (cp.async.bulk.tensor, 128-byte swizzle, tensor map
//              over this GPU's workspace only), completion on mbarriers, a ring of kStages blocks;
//   product    one elected thread issues tcgen05.mma.kind::f16 (M = N = 128, K = 16 per instruction,
//              both operands K-major from shared memory through UMMA descriptors); the 128 x 128
//              fp32 accumulator lives in 128 columns of tensor memory; tcgen05.commit hands shared
//              memory back to the loader and, at the end, signals the epilogue;
//   epilogue   thread t owns accumulator row t: four tcgen05.ld.32x32b.x32 bring 32 columns each
//              into registers, every value is compared with hi_r + hi_c[col] / lo_r + lo_c[col] and
//              the 32 outcomes ARE one word of the bitmask — no accumulator round trip through
//              shared memory, no ballots; undecided pairs go to a queue in the (now free) operand
//              buffers and are re-decided by all threads with the exact sum.
//
// One 128 x 128 tile per CTA, 128 threads, 36 KB of shared memory: four CTAs per SM overlap each
// other's phases.
#include <cuda.h>

#include "rmsd_common.cuh"

namespace gmm {

constexpr int kTcBlockK = 64;                         // halves per staged block: 128 bytes, one swizzle atom
constexpr int kTcStages = 1;                          // one 16 KB + 16 KB block in flight per CTA: what bounds the kernel is
                                                      // latency, not bandwidth, and 36 KB of shared memory per CTA lets
                                                      // four CTAs (4 x 128 TMEM columns = all 512) share an SM and
                                                      // overlap each other's load / MMA / epilogue / re-decision phases
                                                      // (measured: 3 stages, 2 CTAs per SM 0.267 ms; 2: 0.214; 1: 0.173)
constexpr int kTcStageBytes = 2 * kRT * kTcBlockK * 2;
constexpr int kTcQueue = 4096;                        // undecided pairs per tile (aliases the operand buffers)
constexpr int kTcTmemCols = 128;

struct TcSmem {                                       // behind the 1024-byte aligned operand ring
    float ecw[4];                                     // per 32-column word: the largest column part of the error bound
    uint32_t words[kRT * 4];                          // the tile's bits, [row][4 words]
    unsigned long long full[kTcStages], empty[kTcStages], done;
    uint32_t tmem_base;
    int n_queued;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(unsigned long long* b, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
// Bounded by time (two seconds), not by polls; gives up quietly so that a fault cannot hang the GPU
// (the tile's bits are then wrong, which every test of this kernel checks for).
__device__ __forceinline__ void mbar_wait(unsigned long long* b, unsigned parity) {
    unsigned done = 0;
    unsigned long long t0 = 0;
    for (unsigned spin = 0; !done; ++spin) {
        asm volatile("{\n\t.reg .pred q;\n\t"
                     "mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n\t"
                     "selp.u32 %0, 1, 0, q;\n\t}" : "=r"(done) : "r"(smem_u32(b)), "r"(parity) : "memory");
        if (!done && (spin & 4095u) == 4095u) {
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 2000000000ull) return;
        }
    }
}

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1,
                                            unsigned long long* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
                 : "memory");
}

// UMMA shared-memory descriptor of a K-major operand block stored as [rows][64 halves] with the
// 128-byte swizzle TMA applied: 8-row groups are 1024 bytes apart (SBO), version 1 (sm_100),
// layout type 2 (SWIZZLE_128B); the leading-dimension offset is not used by swizzled K-major layouts.
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) |
           (2ull << 61);
}
// kind::f16 instruction descriptor: fp32 accumulator (bits 4-5 = 1), A and B fp16 (0), both K-major,
// N >> 3 at bits 17-22, M >> 4 at bits 24-28.
constexpr uint32_t kTcInstrDesc = (1u << 4) | ((uint32_t)(kRT >> 3) << 17) | ((uint32_t)(kRT >> 4) << 24);

__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(kTcInstrDesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(unsigned long long* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void tmem_load_32(uint32_t taddr, float (&v)[32]) {
    uint32_t r[32];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}

__global__ void __launch_bounds__(128)
rmsd_bits_tc_kernel(const __grid_constant__ CUtensorMap tmap, const __grid_constant__ CUtensorMap tmap_b,
                    const RmsdMmaArgs a) {
    extern __shared__ unsigned char tc_smem_raw[];
    // the operand ring needs 1024-byte alignment (swizzle atom = 8 rows x 128 bytes)
    unsigned char* ring = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~(uintptr_t)1023);
    TcSmem* sm = reinterpret_cast<TcSmem*>(ring + (size_t)kTcStages * kTcStageBytes);
    uint32_t* queue = reinterpret_cast<uint32_t*>(ring);                 // reused once the MMAs are done
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int r_base = a.row0 + blockIdx.y * kRT, c_base = blockIdx.x * kRT;
    const int rb = a.R0 + blockIdx.y, cbk = blockIdx.x;
    const bool in_square = a.sym && cbk >= a.R0 && cbk < a.R1;
    if (in_square && cbk < rb) return;                                   // written by tile (cbk, rb)
    const bool mirror = in_square && cbk > rb;
    const int n_kb = (a.Kp + kTcBlockK - 1) / kTcBlockK;

    if (tid == 0) {
        for (int s = 0; s < kTcStages; ++s) { mbar_init(&sm->full[s], 1); mbar_init(&sm->empty[s], 1); }
        mbar_init(&sm->done, 1);
        sm->n_queued = 0;
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {                                                     // TMEM: 128 columns x 128 lanes of fp32
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(&sm->tmem_base)), "n"(kTcTmemCols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // The accumulator holds u = a . b - (|a|^2 + |b|^2 - cut) / 2 = -(approximate sum - cut) / 2
    // (norms and threshold ride in six extra operand columns, rmsd_center_kernel).  The error of the
    // approximate sum is bounded by 1.2e-3 (|a|^2 + |b|^2) + c0 (fp16 rounding of the inputs,
    // Cauchy-Schwarz, slack for the accumulation; rmsd.cu), so with e = half of that a pair is
    // certainly similar when u > e, certainly not when u < -e.  e = (row part) + (column part); the
    // column part is taken as the largest of the word's 32 columns, which only widens the band.
    const float c0 = 1e-4f * a.cut + 1e-4f;
    {
        const int col = c_base + tid;
        const float nc = col < a.N ? a.norm[col] : 0.f;
        const unsigned mx = __reduce_max_sync(0xffffffffu, __float_as_uint(nc));      // norms are >= 0
        if (lane == 0) sm->ecw[warp] = 0.0006f * __uint_as_float(mx);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = sm->tmem_base;

    if (tid == 0) {
        // ---- loader: both operand blocks of every K block through the ring ----
        for (int kb = 0; kb < n_kb; ++kb) {
            const int s = kb % kTcStages;
            if (kb >= kTcStages) mbar_wait(&sm->empty[s], ((kb / kTcStages) - 1) & 1);
            mbar_expect_tx(&sm->full[s], kTcStageBytes);
            unsigned char* dst = ring + (size_t)s * kTcStageBytes;
            tma_load_2d(dst, &tmap, kb * kTcBlockK, r_base, &sm->full[s]);
            tma_load_2d(dst + kTcStageBytes / 2, &tmap_b, kb * kTcBlockK, c_base, &sm->full[s]);
        }
    } else if (tid == 32) {
        // ---- MMA issuer ----
        for (int kb = 0; kb < n_kb; ++kb) {
            const int s = kb % kTcStages;
            mbar_wait(&sm->full[s], (kb / kTcStages) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a_addr = smem_u32(ring + (size_t)s * kTcStageBytes), b_addr = a_addr + kTcStageBytes / 2;
#pragma unroll
            for (int k = 0; k < kTcBlockK / 16; ++k)                     // 32 bytes along K per instruction
                umma_f16(tmem, umma_desc_sw128(a_addr + 32 * k), umma_desc_sw128(b_addr + 32 * k), (kb | k) != 0);
            umma_commit(&sm->empty[s]);                                  // frees the stage once these MMAs have read it
        }
        umma_commit(&sm->done);
    }
    __syncwarp();

    // ---- epilogue: thread t = accumulator row t ----
    const int row = r_base + tid;
    const bool row_ok = row < a.row1;
    const float er = 0.0006f * (row_ok ? a.norm[row] : 0.f) + 0.5f * c0;
    const bool nothing_decided = *a.overflow != 0;
    const bool vec = (a.K % 4 == 0) && (reinterpret_cast<uintptr_t>(a.xyz) % 16 == 0);
    mbar_wait(&sm->done, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t maybe_w[4];
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
        float v[32];
        tmem_load_32(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)(q * 32), v);
        const float e = er + sm->ecw[q];
        // v > e and |v| <= e as SIGN BITS of e - v and e - |v| (never -0: x - x is +0), each shifted into
        // its word by one funnel shift: 2 FADD on the FMA pipe + 2 SHF per value instead of 2 FSETP +
        // 2 SEL + an add on the ALU pipe, which this epilogue keeps half busy.  Descending j puts
        // value j at bit j.
        uint32_t yes = 0, outside = 0;
#pragma unroll
        for (int j = 31; j >= 0; --j) {
            yes = __funnelshift_l(__float_as_uint(e - v[j]), yes, 1);
            outside = __funnelshift_l(__float_as_uint(e - fabsf(v[j])), outside, 1);
        }
        uint32_t maybe = ~outside;
        if (nothing_decided) { yes = 0; maybe = 0xffffffffu; }
        // columns past N were zero-filled by TMA (u = 0, inside the band): mask them
        const int left = a.N - (c_base + q * 32);
        const uint32_t valid = left >= 32 ? 0xffffffffu : (left <= 0 ? 0u : ((1u << left) - 1u));
        if (!row_ok) { yes = 0; maybe = 0; }
        sm->words[tid * 4 + q] = yes & valid;
        maybe_w[q] = maybe & valid;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();                                                     // all accumulator reads done; operands dead
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTcTmemCols) : "memory");
    // undecided pairs -> queue (in the operand ring); beyond its capacity they are decided on the spot
#pragma unroll 1
    for (int q = 0; q < 4; ++q) {
        uint32_t m = maybe_w[q];
        while (m) {
            const int j = __ffs(m) - 1;
            m &= m - 1;
            const int col = q * 32 + j;
            const int at = atomicAdd(&sm->n_queued, 1);
            if (at < kTcQueue) queue[at] = ((uint32_t)tid << 16) | (uint32_t)col;
            else {
                const float s = exact_sum(a.xyz + (size_t)row * a.K, a.xyz + (size_t)(c_base + col) * a.K, a.K, vec);
                if (!(s * a.inv_a > a.thr2)) atomicOr(&sm->words[tid * 4 + q], 1u << j);
            }
        }
    }
    __syncthreads();
    // the queue, every thread busy; the sum runs over k in order, exactly as the FP32 kernel's
    const int n_queued = min(sm->n_queued, kTcQueue);
    for (int e = tid; e < n_queued; e += 128) {
        const int r = queue[e] >> 16, col = queue[e] & 0xffff;
        const float s = exact_sum(a.xyz + (size_t)(r_base + r) * a.K, a.xyz + (size_t)(c_base + col) * a.K, a.K, vec);
        // similarity.py:67-69: "if rmsd_squared > threshold*threshold: return False"
        if (!(s * a.inv_a > a.thr2)) atomicOr(&sm->words[r * 4 + (col >> 5)], 1u << (col & 31));
    }
    __syncthreads();
    for (int w = tid; w < kRT * 4; w += 128) {
        const int r = w >> 2, q = w & 3, orow = r_base + r, wc = c_base / 32 + q;
        if (orow < a.row1 && wc < a.words) a.bits[(size_t)(orow - a.row0) * a.words + wc] = sm->words[w];
    }
    if (mirror) {
        // the tile transposed: sixteen 32 x 32 bit blocks, each turned by one warp with a five-step
        // butterfly of shuffles (lane = row of the block in, column of the block out)
        for (int blk = warp; blk < 16; blk += 4) {
            const int rw = blk >> 2, q = blk & 3;
            unsigned x = sm->words[(32 * rw + lane) * 4 + q];
#pragma unroll
            for (int step = 0; step < 5; ++step) {
                const int j = 16 >> step;
                const unsigned m = step == 0 ? 0x0000ffffu : step == 1 ? 0x00ff00ffu : step == 2 ? 0x0f0f0f0fu
                                 : step == 3 ? 0x33333333u : 0x55555555u;
                const unsigned y = __shfl_xor_sync(0xffffffffu, x, j);
                x = (lane & j) ? ((x & ~m) | ((y >> j) & m)) : ((x & m) | ((y & m) << j));
            }
            const int mrow = c_base + 32 * q + lane, wc = r_base / 32 + rw;
            if (mrow < a.row1 && mrow < a.N && wc < a.words) a.bits[(size_t)(mrow - a.row0) * a.words + wc] = x;
        }
    }
}

}  // namespace gmm

// ---- host side ------------------------------------------------------------------------------------
namespace gmm {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point query (no link against libcuda)
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
        else
            cudaGetLastError();
    }
    return fn;
}

// Sets *handled = 1 when the tcgen05 kernel ran; 0 when the tensor map could not be made (the caller
// then runs the mma.sync kernel, same bits).
int launch_rmsd_tc(gmm_ctx* c, const RmsdMmaArgs& a, cudaStream_t s, int* handled) {
    *handled = 0;
    EncodeTiledFn encode = encode_tiled_fn();
    if (encode == nullptr) return GMM_OK;
    // [N rows][Kp halves] row-major: dimension 0 = K (contiguous).  The map covers the workspace of
    // THIS context's device and nothing else.
    CUtensorMap tmap, tmap_b;
    const cuuint64_t dims[2] = {(cuuint64_t)a.Kp, (cuuint64_t)a.N};
    const cuuint64_t strides[1] = {(cuuint64_t)a.Kp * sizeof(__half)};
    const cuuint32_t box[2] = {(cuuint32_t)kTcBlockK, (cuuint32_t)kRT};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(a.xc), dims, strides, box,
                              estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return GMM_OK;
    const CUresult rb = encode(&tmap_b, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<__half*>(a.xcb), dims, strides, box,
                               estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rb != CUDA_SUCCESS) return GMM_OK;
    const size_t smem = 1024 + (size_t)kTcStages * kTcStageBytes + sizeof(TcSmem);
    { const int rc = allow_dynamic_smem((const void*)rmsd_bits_tc_kernel, c->device, smem); if (rc) return rc; }
    dim3 grid((a.N + kRT - 1) / kRT, (a.row1 - a.row0 + kRT - 1) / kRT);
    rmsd_bits_tc_kernel<<<grid, 128, smem, s>>>(tmap, tmap_b, a);
    GMM_LAUNCH_CHECK("rmsd_bits_tc_kernel");
    *handled = 1;
    return GMM_OK;
}

}  // namespace gmm
