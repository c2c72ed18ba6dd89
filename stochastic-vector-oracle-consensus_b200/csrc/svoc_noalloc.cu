// svoc_noalloc.cu — svoc_consensus_float_host_noalloc: the host-buffer consensus call for callers that must
// not allocate device memory.
//
// Who that is: a plugin living inside a process whose device memory is owned and budgeted by something
// else (an inference server, a simulator with a fixed arena), code that runs after the application has
// frozen its memory map, and latency-bound loops, where cudaMalloc / cudaFree are unwelcome because they
// take driver-wide locks and can synchronise the whole device.  svoc_consensus_float_host keeps a staging
// cache that it grows with cudaMalloc on demand; this variant stages through a buffer that is part of the
// library image instead, so it never allocates — and in exchange takes only batches that fit that buffer.
//
// The staging buffer is the file-scope array below: SVOC_NOALLOC_SCRATCH_BYTES bytes of device memory on
// every GPU that loads the library, held until the process exits.  Inside host code the array's name is
// only a handle for the runtime, not a device address; the address is asked for once per device with
// cudaGetSymbolAddress (with that device current) and that pointer is what the copies and the kernel get.
//
// Nothing in the buffer has to start at zero: a call overwrites the prediction area by its own copy and the
// consensus kernels write every output element before anything reads it, so the previous call's leftovers
// are never observed.  The buffer is shared by every call a process makes on a device and keeps what the
// last call left in it: two calls that overlap on the GPU — two host threads calling at once — overwrite each
// other's staging and race.  Call this entry point from one thread at a time per device.  (Another process
// has its own copy of the array.)
//
// The first call on a device does the set-up (cudaGetSymbolAddress, kernel attribute set-up in the launch
// path); make that first call before any stream capture is begun in the process.
#include <mutex>
#include <string>

#include "consensus_kernel.cuh"

namespace svoc {

int consensus_device(const svoc_config *cfg, const void *pred, int64_t T, const svoc_outputs *out, void *stream,
                     bool fixed);

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[SVOC_NOALLOC_SCRATCH_BYTES];

std::mutex g_setup_mu;
unsigned char *g_scratch_ptr[64] = {};       // device address of the array, per device ordinal

#define NA_CK(call)                                                                                    \
    do {                                                                                               \
        cudaError_t e_ = (call);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            return fail(std::string("svoc_consensus_float_host_noalloc: " #call ": ") + cudaGetErrorString(e_)); \
    } while (0)

inline size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

// set-up, once per device, with that device current
int scratch_for_current_device(unsigned char **ptr) {
    int dev = 0;
    NA_CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return fail("svoc_consensus_float_host_noalloc: device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_setup_mu);
    if (!g_scratch_ptr[dev]) {
        void *p = nullptr;
        NA_CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        if (!p) return fail("svoc_consensus_float_host_noalloc: cudaGetSymbolAddress returned a null address");
        g_scratch_ptr[dev] = static_cast<unsigned char *>(p);
    }
    *ptr = g_scratch_ptr[dev];
    return 0;
}

int float_host_noalloc(const svoc_config *cfg, const void *pred, int64_t T, const svoc_outputs *out) {
    if (!cfg || !out) return fail("svoc_consensus_float_host_noalloc: null config or outputs");
    if (T < 0) return fail("svoc_consensus_float_host_noalloc: negative round count");
    if (T == 0) return 0;
    if (!pred) return fail("svoc_consensus_float_host_noalloc: null predictions");
    if (cfg->dtype != SVOC_F32 && cfg->dtype != SVOC_F64)
        return fail("svoc_consensus_float_host_noalloc: dtype must be SVOC_F32 or SVOC_F64");
    if (cfg->n_oracles < 1 || cfg->dimension < 1)
        return fail("svoc_consensus_float_host_noalloc: n_oracles and dimension must be >= 1");
    if (!out->consensus || !out->rel1 || !out->rel2 || !out->status)
        return fail("svoc_consensus_float_host_noalloc: consensus, rel1, rel2 and status outputs are required");

    const size_t es = cfg->dtype == SVOC_F32 ? 4 : 8;     // element size; results have the same type
    // skewness / kurtosis are produced only when cfg->want_moments is set: without it the two outputs are neither
    // staged nor copied back (the caller's buffers stay untouched), so nothing the kernels did not write in THIS call
    // can ever reach the caller
    void *const h_skew = cfg->want_moments ? out->skew : nullptr, *const h_kurt = cfg->want_moments ? out->kurt : nullptr;
    const size_t N = cfg->n_oracles, D = cfg->dimension, W = (N + 31) / 32, n = (size_t)T;

    // carve the staging buffer: predictions first, then one area per requested output
    const size_t cap = (size_t)SVOC_NOALLOC_SCRATCH_BYTES;
    const size_t per_round = N * D * es + (out->essence1 ? D * es : 0) + (out->qr ? N * es : 0) + es + (out->reliable ? W * 4 : 0) +
                             D * es + es + (h_skew ? D * es : 0) + (h_kurt ? D * es : 0) + 4 + (out->rank ? N * 4 : 0);
    const size_t fit = cap > 11 * 256 ? (cap - 11 * 256) / per_round : 0;      // every area is padded to 256 bytes
    size_t off = 0;
    auto take = [&](bool wanted, size_t bytes) { const size_t at = off; if (wanted) off += up256(bytes); return at; };
    size_t a_in = 0, a_e1 = 0, a_qr = 0, a_r1 = 0, a_rl = 0, a_co = 0, a_r2 = 0, a_sk = 0, a_ku = 0, a_st = 0, a_rk = 0;
    const bool too_big = n > cap / per_round;                                  // also keeps the products below small
    if (!too_big) {
        a_in = take(true, n * N * D * es);
        a_e1 = take(out->essence1 != nullptr, n * D * es);
        a_qr = take(out->qr != nullptr, n * N * es);
        a_r1 = take(true, n * es);
        a_rl = take(out->reliable != nullptr, n * W * 4);
        a_co = take(true, n * D * es);
        a_r2 = take(true, n * es);
        a_sk = take(h_skew != nullptr, n * D * es);
        a_ku = take(h_kurt != nullptr, n * D * es);
        a_st = take(true, n * 4);
        a_rk = take(out->rank != nullptr, n * N * 4);
    }
    if (too_big || off > cap)
        return fail("svoc_consensus_float_host_noalloc: " + std::to_string(T) + " rounds of " + std::to_string(N) + " x " +
                    std::to_string(D) + " need more device staging than the built-in " + std::to_string(cap) +
                    " bytes (SVOC_NOALLOC_SCRATCH_BYTES; " + std::to_string(fit) +
                    " rounds of this shape fit): split the batch, or use svoc_consensus_float_host");

    unsigned char *b = nullptr;
    if (int rc = scratch_for_current_device(&b)) return rc;

    // everything on the legacy default stream: no stream is created either
    cudaStream_t st = nullptr;
    // once the first copy is queued an error must not return while work is still in flight: the caller may free
    // its (pinned) buffers on failure, and the next call would overlap this one in the shared array
    auto queued = [&]() -> int {
    NA_CK(cudaMemcpyAsync(b + a_in, pred, n * N * D * es, cudaMemcpyHostToDevice, st));
    svoc_outputs d{};
    d.essence1 = out->essence1 ? b + a_e1 : nullptr;
    d.qr = out->qr ? b + a_qr : nullptr;
    d.rel1 = b + a_r1;
    d.reliable = out->reliable ? reinterpret_cast<uint32_t *>(b + a_rl) : nullptr;
    d.consensus = b + a_co;
    d.rel2 = b + a_r2;
    d.skew = h_skew ? b + a_sk : nullptr;
    d.kurt = h_kurt ? b + a_ku : nullptr;
    d.status = reinterpret_cast<int32_t *>(b + a_st);
    d.rank = out->rank ? reinterpret_cast<int32_t *>(b + a_rk) : nullptr;
    if (int rc = consensus_device(cfg, b + a_in, T, &d, st, false)) return rc;
    auto back = [&](void *host, size_t at, size_t bytes) -> cudaError_t {
        return host ? cudaMemcpyAsync(host, b + at, bytes, cudaMemcpyDeviceToHost, st) : cudaSuccess;
    };
    NA_CK(back(out->essence1, a_e1, n * D * es));
    NA_CK(back(out->qr, a_qr, n * N * es));
    NA_CK(back(out->rel1, a_r1, n * es));
    NA_CK(back(out->reliable, a_rl, n * W * 4));
    NA_CK(back(out->consensus, a_co, n * D * es));
    NA_CK(back(out->rel2, a_r2, n * es));
    NA_CK(back(h_skew, a_sk, n * D * es));
    NA_CK(back(h_kurt, a_ku, n * D * es));
    NA_CK(back(out->status, a_st, n * 4));
    NA_CK(back(out->rank, a_rk, n * N * 4));
    NA_CK(cudaStreamSynchronize(st));
    return 0;
    };
    const int rc = queued();
    if (rc != 0) cudaStreamSynchronize(st);      // drain; the first error's text is kept
    return rc;
}

}  // namespace
}  // namespace svoc

extern "C" int svoc_consensus_float_host_noalloc(const svoc_config *cfg, const void *pred, int64_t n_rounds,
                                                 const svoc_outputs *out) {
    return svoc::float_host_noalloc(cfg, pred, n_rounds, out);
}
