// svoc_host.cu — host-buffer entry points: stage predictions to the device in chunks on two
// streams (H2D copy, consensus kernel, D2H copies overlap across chunks) and return when every
// output has landed.  This is the end-to-end path bench.py reports as `e2e`.
//
// The device staging buffers are a per-device cache that grows on demand and is reused across
// calls (one cudaMalloc per high-water mark, not per call).
#include <algorithm>
#include <mutex>
#include <string>

#include "consensus_kernel.cuh"

namespace svoc {

int consensus_device(const svoc_config *cfg, const void *pred, int64_t T, const svoc_outputs *out, void *stream,
                     bool fixed);

namespace {

struct HostWorkspace {
    int device = -1;
    cudaStream_t st[2] = {nullptr, nullptr};
    unsigned char *buf[2] = {nullptr, nullptr};
    size_t cap = 0;
};

// one lock per device: calls on the same device share its staging buffers and streams and are serialised, calls
// that drive different GPUs run side by side
std::mutex g_ws_mu[16];
HostWorkspace g_ws[16];

#define CK(call)                                                                       \
    do {                                                                               \
        cudaError_t e_ = (call);                                                       \
        if (e_ != cudaSuccess) return fail(std::string(#call ": ") + cudaGetErrorString(e_)); \
    } while (0)

size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

struct ChunkLayout {
    size_t in, essence1, qr, rel1, reliable, consensus, rel2, skew, kurt, status, rank, total;
};

ChunkLayout layout(const svoc_config &c, const svoc_outputs &o, int64_t rounds, size_t es, size_t vs) {
    ChunkLayout L{};
    const size_t N = c.n_oracles, D = c.dimension, W = (c.n_oracles + 31) / 32, T = (size_t)rounds;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t at = off; off += al256(bytes); return at; };
    L.in = take(T * N * D * es);
    L.essence1 = take(o.essence1 ? T * D * vs : 0);
    L.qr = take(o.qr ? T * N * vs : 0);
    L.rel1 = take(T * vs);
    L.reliable = take(o.reliable ? T * W * 4 : 0);
    L.consensus = take(T * D * vs);
    L.rel2 = take(T * vs);
    // the kernels write skew / kurt only when moments are requested: no staging (and no copy back) otherwise
    L.skew = take(o.skew && c.want_moments ? T * D * vs : 0);
    L.kurt = take(o.kurt && c.want_moments ? T * D * vs : 0);
    L.status = take(T * 4);
    L.rank = take(o.rank ? T * N * 4 : 0);
    L.total = off;
    return L;
}

int consensus_host(const svoc_config *cfg, const void *pred, int64_t T, const svoc_outputs *out, bool fixed) {
    if (!cfg || !out) return fail("null config or outputs");
    if (T < 0) return fail("negative round count");
    if (T == 0) return 0;
    if (!pred) return fail("null predictions");
    if (cfg->n_oracles < 1 || cfg->dimension < 1) return fail("n_oracles and dimension must be >= 1");
    if (!out->consensus || !out->rel1 || !out->rel2 || !out->status)
        return fail("consensus, rel1, rel2 and status outputs are required");
    const size_t es = (cfg->dtype == SVOC_I32 || cfg->dtype == SVOC_F32) ? 4 : 8;
    const size_t vs = (cfg->dtype == SVOC_F32) ? 4 : 8;
    const size_t N = cfg->n_oracles, D = cfg->dimension, W = (cfg->n_oracles + 31) / 32;
    const size_t round_in = N * D * es;

    // ~32 MiB of predictions per chunk keeps both copy engines and the SMs busy
    int64_t chunk = std::max<int64_t>(1, (int64_t)((32u << 20) / round_in));
    chunk = std::min<int64_t>(chunk, T);
    if (T > chunk && T < 2 * chunk) chunk = (T + 1) / 2;
    const ChunkLayout L = layout(*cfg, *out, chunk, es, vs);

    int dev = 0;
    CK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 16) return fail("device ordinal out of range");
    std::lock_guard<std::mutex> lk(g_ws_mu[dev]);
    HostWorkspace &ws = g_ws[dev];
    if (ws.device != dev) {
        for (int k = 0; k < 2; ++k) CK(cudaStreamCreateWithFlags(&ws.st[k], cudaStreamNonBlocking));
        ws.device = dev;
    }
    if (ws.cap < L.total) {
        for (int k = 0; k < 2; ++k) {
            if (ws.buf[k]) CK(cudaFree(ws.buf[k]));
            ws.buf[k] = nullptr;
        }
        ws.cap = 0;
        for (int k = 0; k < 2; ++k) CK(cudaMalloc(&ws.buf[k], L.total));
        ws.cap = L.total;
    }

    const bool moments = cfg->want_moments != 0;
    void *const h_skew = moments ? out->skew : nullptr, *const h_kurt = moments ? out->kurt : nullptr;
    const unsigned char *hin = static_cast<const unsigned char *>(pred);
    // the pipeline proper; on any failure after the first enqueue the caller below drains both streams, so that no
    // copy into or out of the caller's buffers is still in flight when the error is returned
    auto pipeline = [&]() -> int {
    int k = 0;
    for (int64_t t0 = 0; t0 < T; t0 += chunk, k ^= 1) {
        const int64_t n = std::min<int64_t>(chunk, T - t0);
        unsigned char *b = ws.buf[k];
        cudaStream_t st = ws.st[k];
        CK(cudaMemcpyAsync(b + L.in, hin + (size_t)t0 * round_in, (size_t)n * round_in, cudaMemcpyHostToDevice, st));
        svoc_outputs d{};
        d.essence1 = out->essence1 ? b + L.essence1 : nullptr;
        d.qr = out->qr ? b + L.qr : nullptr;
        d.rel1 = b + L.rel1;
        d.reliable = out->reliable ? reinterpret_cast<uint32_t *>(b + L.reliable) : nullptr;
        d.consensus = b + L.consensus;
        d.rel2 = b + L.rel2;
        d.skew = h_skew ? b + L.skew : nullptr;
        d.kurt = h_kurt ? b + L.kurt : nullptr;
        d.status = reinterpret_cast<int32_t *>(b + L.status);
        d.rank = out->rank ? reinterpret_cast<int32_t *>(b + L.rank) : nullptr;
        if (int rc = consensus_device(cfg, b + L.in, n, &d, st, fixed)) return rc;
        auto back = [&](void *host, size_t at, size_t per_round) -> cudaError_t {
            if (!host) return cudaSuccess;
            return cudaMemcpyAsync(static_cast<unsigned char *>(host) + (size_t)t0 * per_round, b + at,
                                   (size_t)n * per_round, cudaMemcpyDeviceToHost, st);
        };
        CK(back(out->essence1, L.essence1, D * vs));
        CK(back(out->qr, L.qr, N * vs));
        CK(back(out->rel1, L.rel1, vs));
        CK(back(out->reliable, L.reliable, W * 4));
        CK(back(out->consensus, L.consensus, D * vs));
        CK(back(out->rel2, L.rel2, vs));
        CK(back(h_skew, L.skew, D * vs));
        CK(back(h_kurt, L.kurt, D * vs));
        CK(back(out->status, L.status, 4));
        CK(back(out->rank, L.rank, N * 4));
    }
    CK(cudaStreamSynchronize(ws.st[0]));
    CK(cudaStreamSynchronize(ws.st[1]));
    return 0;
    };
    const int rc = pipeline();
    if (rc != 0) {                       // keep the first error's text; the drain's own result is of no interest
        cudaStreamSynchronize(ws.st[0]);
        cudaStreamSynchronize(ws.st[1]);
    }
    return rc;
}

}  // namespace
}  // namespace svoc

extern "C" {
int svoc_consensus_fixed_host(const svoc_config *cfg, const void *pred, int64_t n_rounds, const svoc_outputs *out) {
    return svoc::consensus_host(cfg, pred, n_rounds, out, true);
}
int svoc_consensus_float_host(const svoc_config *cfg, const void *pred, int64_t n_rounds, const svoc_outputs *out) {
    return svoc::consensus_host(cfg, pred, n_rounds, out, false);
}
}
