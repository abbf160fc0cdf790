// gmm_evaluate_static: population evaluate for callers that must not allocate device memory.
//
// Who needs this.  gmm_evaluate() parks the expressed coordinates (when the caller does not ask
// for them; the tiled Contacts path also a per-individual zone bitmask and per-chunk partial
// sums) in a scratch buffer owned by the context, and grows that buffer with cudaMalloc the first
// time it sees a larger population.  Some hosts of this library cannot live with that:
//   * a GA driver that shares the GPU with a framework which has already claimed the device
//     memory for its own pool (a torch process capped with a memory fraction, a MIG / MPS
//     slice with a fixed budget): a raw cudaMalloc from inside a plugin call can fail there,
//     or succeed and push the framework itself into an out-of-memory error later;
//   * a service that evaluates populations of varying size in a latency-bound loop:
//     cudaMalloc and cudaFree synchronise the whole device, so one growth step stalls every
//     other stream of the process;
//   * code that is only allowed to enqueue work (for example inside a callback) and has no
//     way to report an allocation failure upwards.
// For them the scratch lives in a fixed array inside the library image itself.  Every GPU
// that loads libgaudimm_b200.so carries these kStaticScratchBytes until the process exits,
// which is why the size is modest and populations that need more are refused.
#include <mutex>

#include "common.cuh"

namespace {

// 16 MiB.  Enough for about 34,000 individuals of a 40-atom ligand on the default (fused)
// Contacts path; with GMM_OPT_TILED, e.g. 4,096 individuals against a 5,000-atom receptor.
constexpr size_t kStaticScratchBytes = 16u << 20;

// The scratch itself.  In HOST code this name is only a handle for the symbol, not an
// address the GPU can use; the device address is looked up once per device below.
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
void* g_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_scratch_mu;      // host threads making their first call on a device at the same time

// Device address of g_static_scratch on `device` (which must be the current device).
int static_scratch_address(int device, void** out) {
    if (device < 0 || device >= kMaxDevices) {
        gmm::set_error("gmm_evaluate_static: device %d outside [0,%d)", device, kMaxDevices);
        return GMM_ELIMIT;
    }
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (g_scratch_addr[device] == nullptr) {
        void* ptr = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
        if (e != cudaSuccess || ptr == nullptr) {
            gmm::set_error("gmm_evaluate_static: cudaGetSymbolAddress failed on device %d: %s", device,
                           cudaGetErrorString(e));
            return GMM_ECUDA;
        }
        g_scratch_addr[device] = ptr;
    }
    *out = g_scratch_addr[device];
    return GMM_OK;
}

}  // namespace

extern "C" int gmm_evaluate_static(gmm_ctx* c, int P, const double* torsion_deg_d, const double* xform_d,
                                   const double* rot_allele_d, double* out_scores_d, int32_t* out_counts_d,
                                   void* stream) {
    if (c == nullptr) { gmm::set_error("gmm_evaluate_static: ctx is NULL"); return GMM_EINVAL; }
    if (!c->has_mobile || !c->has_receptor) {
        gmm::set_error("gmm_evaluate_static: the static model is not complete (gmm_set_mobile / gmm_set_receptor)");
        return GMM_ESTATE;
    }
    if (P < 0) { gmm::set_error("gmm_evaluate_static: P < 0"); return GMM_EINVAL; }
    if (P == 0) return GMM_OK;
    if (out_scores_d == nullptr) { gmm::set_error("gmm_evaluate_static: out_scores_d is NULL"); return GMM_EINVAL; }

    const size_t need = gmm::contacts_scratch_bytes(c, P);
    if (need > kStaticScratchBytes) {
        gmm::set_error("gmm_evaluate_static: a population of %d individuals (%d mobile atoms, %d receptor atoms) "
                       "needs %zu bytes of scratch but the built-in buffer holds %zu; evaluate in smaller "
                       "batches or use gmm_evaluate", P, c->M, c->R, need, kStaticScratchBytes);
        return GMM_ELIMIT;
    }
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != c->device && cudaSetDevice(c->device) != cudaSuccess) {
        gmm::set_error("gmm_evaluate_static: cannot select device %d", c->device);
        return GMM_ECUDA;
    }
    void* ws = nullptr;
    int rc = static_scratch_address(c->device, &ws);
    if (rc == GMM_OK) {
        // Layout inside the buffer is the one launch_contacts uses for the context scratch:
        // coordinates first, then (tiled path only) zone bitmask and partial sums.  Every word
        // that is read is written earlier in the same call (express writes all coordinates, the
        // tiled sweep writes all zone words and partial sums of the P individuals), so nothing has
        // to be cleared even though the buffer still holds the previous call's data.
        float* xyz = static_cast<float*>(ws);
        rc = gmm_express(c, P, torsion_deg_d, xform_d, rot_allele_d, xyz, stream);
        if (rc == GMM_OK)
            rc = gmm::launch_contacts(c, P, xyz, out_scores_d, out_counts_d, (cudaStream_t)stream, ws);
    }
    if (prev >= 0 && prev != c->device) cudaSetDevice(prev);
    return rc;
}
