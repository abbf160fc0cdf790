// Fitness exchange between the GPUs of one node (one process per GPU): what the all-gather of
// SURVEY.md section 8(e) does, without a collective library in the step.
//
// The population is sharded by individual (gaudi/algorithms.py:99-102 is a pure map), every rank
// scores its block, and rank 0 — where NSGA-II selection runs (gaudi/algorithms.py:160) — needs
// the whole fitness vector.  A 1 MiB all-gather is pure latency: an NCCL call costs a host-side
// enqueue per step and cannot be replayed from the step's CUDA graph on this stack.  Here each
// rank owns one cudaMalloc'ed block that the other processes map through CUDA IPC; one small
// kernel per step
//     1. copies this rank's [n_rows, 2] scores (and pair counts) into EVERY rank's block with
//        plain 16-byte stores over NVLink (peer pointers; no TMA, no multicast),
//     2. publishes "epoch e is complete" to every rank with a release store at system scope,
//     3. waits (bounded by the global timer) until every rank has published epoch e.
// The kernel has no host-side state, so the whole step (express, Contacts, exchange) is one CUDA
// graph at any number of GPUs.  Two gather buffers alternate with the parity of the epoch: a rank
// that runs ahead writes epoch e + 1 into the other buffer while a slower rank may still be
// reading epoch e; it cannot reach e + 2 before that rank has published e + 1, i.e. has finished
// everything it enqueued before.
#include <cstring>

#include "common.cuh"

namespace gmm {

struct ExchangeHeader {              // first 256 bytes of every rank's block
    unsigned flags[GMM_MAX_PEERS];   // flags[r]: last epoch rank r has published here
    unsigned epoch;                  // epochs completed by the owner
    unsigned done_blocks;            // ticket counter of the running publish kernel
    unsigned error;                  // 1: a wait timed out
    unsigned pad[64 - GMM_MAX_PEERS - 3];
};
static_assert(sizeof(ExchangeHeader) == 256, "header layout");

struct ExchangeArgs {
    unsigned char* base[GMM_MAX_PEERS];   // base[r]: rank r's block (own: local pointer)
    int world, rank, max_rows;
    const double* scores; const int32_t* counts;
    int row0, n_rows;
};

__host__ __device__ inline size_t exchange_buffer_bytes(int max_rows) {
    return ((size_t)max_rows * (2 * sizeof(double) + 2 * sizeof(int32_t)) + 255) / 256 * 256;
}
__host__ __device__ inline double* exchange_scores(unsigned char* base, int max_rows, unsigned parity) {
    return reinterpret_cast<double*>(base + sizeof(ExchangeHeader) + parity * exchange_buffer_bytes(max_rows));
}
__host__ __device__ inline int32_t* exchange_counts(unsigned char* base, int max_rows, unsigned parity) {
    return reinterpret_cast<int32_t*>(base + sizeof(ExchangeHeader) + parity * exchange_buffer_bytes(max_rows) +
                                      (size_t)max_rows * 2 * sizeof(double));
}

// grid: world x kSlices blocks; block (r, s) copies slice s of the rows into rank r's block
constexpr int kExSlices = 4;

__global__ void __launch_bounds__(256)
exchange_publish_kernel(const ExchangeArgs a) {
    ExchangeHeader* self = reinterpret_cast<ExchangeHeader*>(a.base[a.rank]);
    const unsigned e = self->epoch + 1u;            // the same in every block: only the last block to
    const unsigned parity = e & 1u;                 // finish advances it, after all of them have read it
    const int dst = blockIdx.x % a.world, slice = blockIdx.x / a.world;
    const int per = (a.n_rows + kExSlices - 1) / kExSlices;
    const int lo = slice * per, hi = min(a.n_rows, lo + per);
    double2* out_s = reinterpret_cast<double2*>(exchange_scores(a.base[dst], a.max_rows, parity)) + a.row0;
    const double2* in_s = reinterpret_cast<const double2*>(a.scores);
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) out_s[i] = in_s[i];
    if (a.counts != nullptr) {
        int2* out_c = reinterpret_cast<int2*>(exchange_counts(a.base[dst], a.max_rows, parity)) + a.row0;
        const int2* in_c = reinterpret_cast<const int2*>(a.counts);
        for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) out_c[i] = in_c[i];
    }
    __threadfence_system();
    __syncthreads();
    __shared__ bool last;
    if (threadIdx.x == 0) last = atomicAdd(&self->done_blocks, 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    // every block of this rank has stored and fenced: publish, then wait for the others
    if ((int)threadIdx.x < a.world) {
        // (the release store orders everything this block has observed, i.e. every block's fenced
        // stores, before the flag: no separate fence)
        unsigned* flag = &reinterpret_cast<ExchangeHeader*>(a.base[threadIdx.x])->flags[a.rank];
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(e) : "memory");
        const unsigned* mine = &self->flags[threadIdx.x];
        unsigned long long t0 = 0;
        for (unsigned spin = 0;; ++spin) {
            unsigned seen;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
            if ((int)(seen - e) >= 0) break;
            if ((spin & 1023u) == 1023u) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > 10000000000ull) { self->error = 1u; break; }   // 10 s: a rank is gone
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) { self->done_blocks = 0u; self->epoch = e; }
}

// Publish + wait only: used when the Contacts kernel has already stored the rows itself (they are
// complete and visible system-wide once that kernel has finished, which stream order guarantees).
__global__ void __launch_bounds__(32)
exchange_signal_kernel(const ExchangeArgs a) {
    ExchangeHeader* self = reinterpret_cast<ExchangeHeader*>(a.base[a.rank]);
    const unsigned e = self->epoch + 1u;
    if ((int)threadIdx.x < a.world) {
        unsigned* flag = &reinterpret_cast<ExchangeHeader*>(a.base[threadIdx.x])->flags[a.rank];
        asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(e) : "memory");
        const unsigned* mine = &self->flags[threadIdx.x];
        unsigned long long t0 = 0;
        for (unsigned spin = 0;; ++spin) {
            unsigned seen;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(seen) : "l"(mine) : "memory");
            if ((int)(seen - e) >= 0) break;
            if ((spin & 1023u) == 1023u) {
                unsigned long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                if (t0 == 0) t0 = now;
                else if (now - t0 > 10000000000ull) { self->error = 1u; break; }
            }
        }
    }
    __syncwarp();
    if (threadIdx.x == 0) self->epoch = e;
}

struct Exchange {
    int world = 0, rank = 0, max_rows = 0;
    MirrorTable* d_mirror = nullptr;
    unsigned char* local = nullptr; size_t bytes = 0;
    unsigned char* peer[GMM_MAX_PEERS] = {nullptr};
    bool connected = false;
    unsigned host_epoch = 0;         // publishes ISSUED through the host API (graph replays are not seen)
};

}  // namespace gmm

using namespace gmm;

static Exchange* ex_of(gmm_ctx* c) { return static_cast<Exchange*>(c->exchange); }

static int upload_mirror(gmm_ctx* c, Exchange* x) {
    MirrorTable mt;
    memset(&mt, 0, sizeof(mt));
    mt.n = x->world; mt.rank = x->rank; mt.max_rows = x->max_rows;
    for (int r = 0; r < x->world; ++r) mt.base[r] = x->peer[r];
    if (!x->d_mirror) GMM_CUDA(cudaMalloc((void**)&x->d_mirror, sizeof(MirrorTable)));
    GMM_CUDA(cudaMemcpy(x->d_mirror, &mt, sizeof(mt), cudaMemcpyHostToDevice));
    c->d_mirror = x->d_mirror;
    return GMM_OK;
}

extern "C" {

int gmm_exchange_destroy(gmm_ctx* c) {
    if (!c || !c->exchange) return GMM_OK;
    DeviceGuard g(c->device);
    Exchange* x = ex_of(c);
    cudaDeviceSynchronize();
    for (int r = 0; r < x->world; ++r)
        if (r != x->rank && x->peer[r]) cudaIpcCloseMemHandle(x->peer[r]);
    if (x->local) cudaFree(x->local);
    if (x->d_mirror) cudaFree(x->d_mirror);
    delete x;
    c->exchange = nullptr;
    c->d_mirror = nullptr;
    return GMM_OK;
}

int gmm_exchange_create(gmm_ctx* c, int world, int rank, int max_rows, unsigned char* handle_out) {
    if (!c || !handle_out) { set_error("gmm_exchange_create: NULL argument"); return GMM_EINVAL; }
    if (world < 1 || world > GMM_MAX_PEERS || rank < 0 || rank >= world || max_rows < 1) {
        set_error("gmm_exchange_create: world=%d (1..%d) rank=%d max_rows=%d invalid", world, GMM_MAX_PEERS, rank, max_rows);
        return GMM_EINVAL;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == GMM_IPC_HANDLE_BYTES, "IPC handle size");
    gmm_exchange_destroy(c);
    DeviceGuard g(c->device);
    Exchange* x = new Exchange();
    x->world = world; x->rank = rank; x->max_rows = max_rows;
    x->bytes = sizeof(ExchangeHeader) + 2 * exchange_buffer_bytes(max_rows);
    if (cudaMalloc((void**)&x->local, x->bytes) != cudaSuccess) {
        set_error("gmm_exchange_create: cudaMalloc of %zu bytes failed: %s", x->bytes, cudaGetErrorString(cudaGetLastError()));
        delete x;
        return GMM_ECUDA;
    }
    cudaMemset(x->local, 0, x->bytes);
    cudaDeviceSynchronize();
    x->peer[rank] = x->local;
    cudaIpcMemHandle_t h;
    memset(&h, 0, sizeof(h));
    if (world > 1) {
        const cudaError_t e = cudaIpcGetMemHandle(&h, x->local);
        if (e != cudaSuccess) {
            set_error("gmm_exchange_create: cudaIpcGetMemHandle failed: %s", cudaGetErrorString(e));
            cudaFree(x->local);
            delete x;
            return GMM_ECUDA;
        }
    }
    memcpy(handle_out, &h, sizeof(h));
    x->connected = (world == 1);
    c->exchange = x;
    if (world == 1) return upload_mirror(c, x);
    return GMM_OK;
}

int gmm_exchange_connect(gmm_ctx* c, const unsigned char* handles_h) {
    if (!c || !c->exchange || !handles_h) { set_error("gmm_exchange_connect: no exchange on this context"); return GMM_ESTATE; }
    DeviceGuard g(c->device);
    Exchange* x = ex_of(c);
    for (int r = 0; r < x->world; ++r) {
        if (r == x->rank || x->peer[r]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles_h + (size_t)r * GMM_IPC_HANDLE_BYTES, sizeof(h));
        void* p = nullptr;
        const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            set_error("gmm_exchange_connect: cudaIpcOpenMemHandle for rank %d failed: %s", r, cudaGetErrorString(e));
            return GMM_ECUDA;
        }
        x->peer[r] = static_cast<unsigned char*>(p);
    }
    x->connected = true;
    return upload_mirror(c, x);
}

int gmm_exchange_publish(gmm_ctx* c, const double* scores_d, const int32_t* counts_d, int row0, int n_rows,
                         double** gathered_scores_out, int32_t** gathered_counts_out, void* stream) {
    if (!c || !c->exchange) { set_error("gmm_exchange_publish: no exchange on this context"); return GMM_ESTATE; }
    Exchange* x = ex_of(c);
    if (!x->connected) { set_error("gmm_exchange_publish: gmm_exchange_connect has not been called"); return GMM_ESTATE; }
    if (row0 < 0 || n_rows < 0 || row0 + n_rows > x->max_rows || (n_rows > 0 && !scores_d)) {
        set_error("gmm_exchange_publish: rows [%d,%d) outside [0,%d] or NULL scores", row0, row0 + n_rows, x->max_rows);
        return GMM_EINVAL;
    }
    DeviceGuard g(c->device);
    ExchangeArgs a;
    for (int r = 0; r < GMM_MAX_PEERS; ++r) a.base[r] = r < x->world ? x->peer[r] : nullptr;
    a.world = x->world; a.rank = x->rank; a.max_rows = x->max_rows;
    a.scores = scores_d; a.counts = counts_d; a.row0 = row0; a.n_rows = n_rows;
    exchange_publish_kernel<<<x->world * kExSlices, 256, 0, (cudaStream_t)stream>>>(a);
    GMM_LAUNCH_CHECK("exchange_publish_kernel");
    x->host_epoch += 1;
    const unsigned parity = x->host_epoch & 1u;
    if (gathered_scores_out) *gathered_scores_out = exchange_scores(x->local, x->max_rows, parity);
    if (gathered_counts_out) *gathered_counts_out = exchange_counts(x->local, x->max_rows, parity);
    return GMM_OK;
}

int gmm_evaluate_publish(gmm_ctx* c, int P, const double* torsion_deg_d, const double* xform_d,
                         const double* rot_allele_d, float* out_xyz_d, double* out_scores_d, int32_t* out_counts_d,
                         int row0, double** gathered_scores_out, int32_t** gathered_counts_out, void* stream) {
    if (!c || !c->exchange) { set_error("gmm_evaluate_publish: no exchange on this context"); return GMM_ESTATE; }
    Exchange* x = ex_of(c);
    if (!x->connected) { set_error("gmm_evaluate_publish: gmm_exchange_connect has not been called"); return GMM_ESTATE; }
    if (row0 < 0 || P < 0 || row0 + P > x->max_rows) {
        set_error("gmm_evaluate_publish: rows [%d,%d) outside [0,%d]", row0, row0 + P, x->max_rows);
        return GMM_EINVAL;
    }
    if (!out_scores_d || !out_counts_d) { set_error("gmm_evaluate_publish: NULL output"); return GMM_EINVAL; }
    // the one-kernel Contacts path stores every finished row into every rank's table itself
    c->pub_row0 = row0;
    c->pub_done = false;
    int rc = gmm_evaluate(c, P, torsion_deg_d, xform_d, rot_allele_d, out_xyz_d, out_scores_d, out_counts_d, stream);
    c->pub_row0 = -1;
    if (rc) return rc;
    if (!c->pub_done || P == 0)      // tiled Contacts path (or nothing to score): copy the rows, then signal
        return gmm_exchange_publish(c, out_scores_d, out_counts_d, row0, P, gathered_scores_out, gathered_counts_out, stream);
    DeviceGuard g(c->device);
    ExchangeArgs a;
    for (int r = 0; r < GMM_MAX_PEERS; ++r) a.base[r] = r < x->world ? x->peer[r] : nullptr;
    a.world = x->world; a.rank = x->rank; a.max_rows = x->max_rows;
    a.scores = nullptr; a.counts = nullptr; a.row0 = row0; a.n_rows = P;
    exchange_signal_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(a);
    GMM_LAUNCH_CHECK("exchange_signal_kernel");
    x->host_epoch += 1;
    const unsigned parity = x->host_epoch & 1u;
    if (gathered_scores_out) *gathered_scores_out = exchange_scores(x->local, x->max_rows, parity);
    if (gathered_counts_out) *gathered_counts_out = exchange_counts(x->local, x->max_rows, parity);
    return GMM_OK;
}

int gmm_exchange_state(gmm_ctx* c, unsigned* epoch_out, int* error_out, double** gathered_scores_out,
                       int32_t** gathered_counts_out) {
    if (!c || !c->exchange) { set_error("gmm_exchange_state: no exchange on this context"); return GMM_ESTATE; }
    DeviceGuard g(c->device);
    Exchange* x = ex_of(c);
    ExchangeHeader h;
    GMM_CUDA(cudaDeviceSynchronize());
    GMM_CUDA(cudaMemcpy(&h, x->local, sizeof(h), cudaMemcpyDeviceToHost));
    x->host_epoch = h.epoch;                       // graph replays advance the device epoch only
    if (epoch_out) *epoch_out = h.epoch;
    if (error_out) *error_out = (int)h.error;
    const unsigned parity = h.epoch & 1u;
    if (gathered_scores_out) *gathered_scores_out = exchange_scores(x->local, x->max_rows, parity);
    if (gathered_counts_out) *gathered_counts_out = exchange_counts(x->local, x->max_rows, parity);
    return GMM_OK;
}

}  // extern "C"
