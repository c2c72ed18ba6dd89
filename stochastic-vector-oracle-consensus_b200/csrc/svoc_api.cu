// svoc_api.cu — host side of libsvoc: argument checks, launch planning and the C ABI
// declared in include/svoc.h.
#include <cstdio>
#include <cstring>
#include <mutex>
#include <string>

#include <algorithm>
#include <cstdlib>

#include "consensus_mid.cuh"
#include "montecarlo.cuh"

namespace svoc {

// ------------------------------------------------------------------------------ host side
static thread_local std::string g_err;
static int64_t g_launches = 0;
static std::mutex g_mu;

int fail(const std::string &m) {
    g_err = m;
    return -1;
}

static int npad_for(int N) { return N <= 8 ? 8 : N <= 16 ? 16 : N <= 32 ? 32 : N <= 64 ? 64 : 0; }

static size_t elem_size(int dtype) { return (dtype == SVOC_I32 || dtype == SVOC_F32) ? 4 : 8; }

static size_t smem_bytes(const svoc_config &c, int rb, int tile_smem) {
    const size_t es = elem_size(c.dtype);
    const size_t vs = (c.dtype == SVOC_F32) ? 4 : 8;  // val_t / acc_t size
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
    size_t s = 0;
    if (tile_smem) s += al((size_t)rb * c.n_oracles * c.dimension * es);
    s += al((size_t)rb * c.dimension * vs);
    s += al((size_t)rb * c.n_oracles * vs);
    s += al((size_t)rb * sizeof(int));
    s += 16;
    s += al((size_t)rb * c.n_oracles);
    return s;
}

// SM count of the CURRENT device (cached per ordinal: a process may drive different GPUs)
static int sm_count() {
    static int cache[64] = {};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    int n = cache[dev];
    if (n == 0) {
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return n;
}

// wide variant: 64 < N <= 512, 32-bit elements, one (column, row-part) unit per thread
static bool plan_wide(const svoc_config &c, const void *pred, Plan &pl) {
    const int N = c.n_oracles, D = c.dimension, R = c.n_oracles - c.n_failing;
    if (!(c.dtype == SVOC_F32 || c.dtype == SVOC_I32)) return false;
    if (N <= 64 || N > 512 || D > 512 || R < 2 || c.n_failing < 0) return false;
    const int NT = 512;
    int P = 1;
    while (P * 2 <= 8 && P * 2 * D <= NT && P * 2 * 32 <= N) P *= 2;
    const int RP = (((N + P - 1) / P) + 31) & ~31;
    if (P * RP > 512) return false;
    int CS = P * RP;
    while ((CS / 4) % 32 != 4) CS += 16;
    int H = 1;
    while (H * 2 * N <= NT && H * 2 <= D) H *= 2;
    const bool use_g = false;
    const size_t smem = wide_smem_bytes(N, D, P, RP, CS, H, 4, c.dtype == SVOC_F32 ? 4 : 8, use_g);
    if (smem > 227 * 1024) return false;
    pl.variant = 1;
    pl.wP = P; pl.wRP = RP; pl.wCS = CS; pl.wH = H;
    pl.smem = smem;
    pl.threads = NT;
    pl.rb = 1;
    pl.npad = 0;
    pl.tile_smem = use_g ? 0 : 1;
    return true;
}

static const std::string &forced_variant() {
    static const std::string force = [] { const char *e = getenv("SVOC_VARIANT"); return std::string(e ? e : ""); }();
    return force;
}

// wide2 variant: parts of a column inside one warp, the round read from global memory / L2 (two CTAs per SM)
static bool plan_wide2(const svoc_config &c, const void *pred, Plan &pl) {
    const int N = c.n_oracles, D = c.dimension, R = c.n_oracles - c.n_failing;
    if (N <= 64 || N > 512 || D > 512 || R < 2 || c.n_failing < 0) return false;
    const std::string &force = forced_variant();
    if (!force.empty() && force != "wide2") return false;
    const size_t es = elem_size(c.dtype);            // 64-bit elements too: keys and counting are type-independent
    const size_t round_bytes = (size_t)N * D * es;
    if ((reinterpret_cast<uintptr_t>(pred) & 15) != 0 || (round_bytes & 15) != 0) return false;
    int P = 1;
    while (P * 2 <= 8 && P * 2 * D <= W2_THREADS && P * 2 * 32 <= N) P *= 2;
    const int RP = (((N + P - 1) / P) + 31) & ~31;
    if (P * RP > 512) return false;
    // every warp must hold whole columns: D * P threads in warps of 32 / P columns
    if ((D * P + 31) / 32 > W2_THREADS / 32) return false;
    int CS = P * RP;
    while ((CS / 4) % 32 != 4) CS += 16;
    int HR = 1;
    while (HR * 2 <= 4 && HR * 2 * N <= W2_THREADS && D / (HR * 2) >= 8) HR *= 2;
    if ((W2_THREADS / 32) * (32 / HR) < 1) return false;
    const size_t smem = wide2_smem_bytes(N, D, P, RP, CS, es, c.dtype == SVOC_F32 ? 4 : 8);
    if (smem > 110 * 1024) return false;
    pl.variant = 3;
    pl.wP = P; pl.wRP = RP; pl.wCS = CS; pl.wH = HR;
    pl.smem = smem;
    pl.threads = W2_THREADS;
    pl.rb = 1;
    pl.npad = 0;
    pl.tile_smem = 0;
    return true;
}

// wide3 variant: the round in registers, split by columns over a 2-CTA cluster, fetched by TMA tensor-map copies.
// Covers the hot shape (D = 128, N <= 256, 32-bit elements, median centres, no moments / rank output); everything
// else goes to wide2.
static bool plan_wide3(const svoc_config &c, const void *pred, const svoc_outputs *out, int64_t T, Plan &pl) {
    const int N = c.n_oracles, D = c.dimension, R = c.n_oracles - c.n_failing;
    if (!(c.dtype == SVOC_F32 || c.dtype == SVOC_I32)) return false;
    if (N <= 64 || N > W3_ROWS || D != W3_D || R < 2 || c.n_failing < 0) return false;
    const std::string &force = forced_variant();
    if (!force.empty() && force != "wide3") return false;
    if (c.want_moments || (out && out->rank)) return false;
    const int centre2 = c.centre2 == SVOC_CENTRE_DEFAULT ? (c.constrained ? c.centre : SVOC_CENTRE_MEAN) : c.centre2;
    if (c.centre == SVOC_CENTRE_MEAN || centre2 == SVOC_CENTRE_MEAN) return false;
    if ((reinterpret_cast<uintptr_t>(pred) & 15) != 0) return false;
    if ((long long)T * N > 0x7fffffffLL) return false;                  // TMA row coordinates are 32-bit
    pl.variant = 4;
    pl.smem = w3_smem_layout(c.dtype == SVOC_F32 ? 4 : 8).total;
    // lanes per column: 8 (512 threads, 64 registers, two CTAs per SM) or 4 (256 threads; three CTAs per SM for fp32)
    static const int lanes_env = [] { const char *e = getenv("SVOC_W3_LANES"); return e ? atoi(e) : 0; }();
    const int P = lanes_env == 4 || lanes_env == 8 ? lanes_env : W3_DEFAULT_LANES;
    pl.threads = W3_COLS * P;
    pl.rb = 1;
    pl.npad = 0;
    pl.tile_smem = 1;
    pl.wP = P; pl.wRP = W3_ROWS / P; pl.wCS = 0; pl.wH = 1;
    // clusters of two CTAs; as many clusters as the SMs hold at once
    pl.grid = 2 * std::min<long long>(T, (long long)sm_count() * w3_ctas_per_sm(P, c.dtype == SVOC_F32 ? 4 : 8) / 2);
    return true;
}

// mid variant: one warp per round for 32 < N <= 64, D <= 32, 32-bit elements, median centres, no moments / rank output
static bool plan_mid(const svoc_config &c, const svoc_outputs *out, int64_t T, Plan &pl) {
    const int N = c.n_oracles, D = c.dimension, R = c.n_oracles - c.n_failing;
    if (!(c.dtype == SVOC_F32 || c.dtype == SVOC_I32)) return false;
    if (N <= 32 || N > MID_ROWS || D < 1 || D > 32 || R < 1 || c.n_failing < 0) return false;
    const std::string &force = forced_variant();
    if (!force.empty() && force != "mid") return false;
    if (c.want_moments || (out && out->rank)) return false;
    const int centre2 = c.centre2 == SVOC_CENTRE_DEFAULT ? (c.constrained ? c.centre : SVOC_CENTRE_MEAN) : c.centre2;
    if (c.centre == SVOC_CENTRE_MEAN || centre2 == SVOC_CENTRE_MEAN) return false;
    const int P = D > 16 ? 1 : (D > 8 ? 2 : 4);
    const size_t smem = mid_smem_layout(N, D, c.dtype == SVOC_F32 ? 4 : 8).total;
    if (smem > 200 * 1024) return false;
    pl.variant = 5;
    pl.smem = smem;
    pl.threads = MID_WARPS * 32;
    pl.rb = MID_ROUNDS;
    pl.npad = 0;
    pl.tile_smem = 1;
    pl.wP = P; pl.wRP = MID_ROWS / P; pl.wCS = 0; pl.wH = 1;
    pl.grid = (T + MID_ROUNDS - 1) / MID_ROUNDS;
    return pl.grid <= 0x7fffffffLL;
}

W3EncodeFn w3_encode_fn() {
    static W3EncodeFn fn = [] {
        void *f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            f = nullptr;
        return reinterpret_cast<W3EncodeFn>(f);
    }();
    return fn;
}

static Plan make_plan(const svoc_config &c, const void *pred, const svoc_outputs *out, int64_t T) {
    Plan pl{};
    if (plan_wide3(c, pred, out, T, pl)) return pl;
    if (plan_mid(c, out, T, pl)) return pl;
    if (plan_wide2(c, pred, pl)) {
        pl.grid = std::min<long long>(T, (long long)sm_count() * (elem_size(c.dtype) == 8 ? 1 : W2_CTAS_PER_SM));
        return pl;
    }
    if (forced_variant() != "generic" && plan_wide(c, pred, pl)) {
        pl.grid = std::min<long long>(T, (long long)sm_count() * (pl.tile_smem ? 1 : 2));
        return pl;
    }
    pl.variant = 0;
    pl.npad = npad_for(c.n_oracles);
    pl.threads = pl.npad ? 128 : 256;
    const size_t budget = pl.npad ? 40 * 1024 : 200 * 1024;
    int rb = pl.threads / (c.dimension > 0 ? c.dimension : 1);
    if (rb < 1) rb = 1;
    while (rb > 1 && smem_bytes(c, rb, 1) > budget) rb >>= 1;
    pl.tile_smem = smem_bytes(c, rb, 1) <= 200 * 1024 ? 1 : 0;
    pl.rb = rb;
    pl.smem = smem_bytes(c, rb, pl.tile_smem);
    pl.grid = (T + rb - 1) / rb;
    return pl;
}

static int validate(const svoc_config *cfg, const void *pred, int64_t T, const svoc_outputs *out, bool fixed) {
    if (!cfg || !out) return fail("null config or outputs");
    if (T < 0) return fail("negative round count");
    if (T > 0 && !pred) return fail("null predictions");
    if (cfg->n_oracles < 1 || cfg->dimension < 1) return fail("n_oracles and dimension must be >= 1");
    if (cfg->n_oracles > 65535) return fail("n_oracles > 65535 is not supported");
    if (cfg->n_failing < 0) return fail("n_failing must be >= 0");
    if (fixed && !(cfg->dtype == SVOC_I32 || cfg->dtype == SVOC_I64)) return fail("fixed form needs dtype I32 or I64");
    if (!fixed && !(cfg->dtype == SVOC_F32 || cfg->dtype == SVOC_F64)) return fail("float form needs dtype F32 or F64");
    if (!fixed && (cfg->n_oracles < 2 || cfg->n_failing >= cfg->n_oracles))
        return fail("float form needs n_oracles >= 2 and n_failing < n_oracles");
    if (cfg->centre < 0 || cfg->centre > 2) return fail("bad centre");
    if (cfg->centre2 < -1 || cfg->centre2 > 2) return fail("bad centre2");
    if (cfg->tie_rule < 0 || cfg->tie_rule > 1) return fail("bad tie_rule");
    if (!out->consensus || !out->rel1 || !out->rel2 || !out->status)
        return fail("consensus, rel1, rel2 and status outputs are required");
    return 0;
}

static Params make_params(const svoc_config &c, const void *pred, int64_t T, const svoc_outputs &o, const Plan &pl) {
    Params p;
    p.pred = pred;
    p.T = T;
    p.N = c.n_oracles;
    p.D = c.dimension;
    p.F = c.n_failing;
    p.R = c.n_oracles - c.n_failing;
    p.constrained = c.constrained ? 1 : 0;
    p.ms_w = c.max_spread_wsad;
    p.ms_f = c.max_spread;
    p.centre = c.centre;
    p.centre2 = c.centre2 == SVOC_CENTRE_DEFAULT ? (c.constrained ? c.centre : SVOC_CENTRE_MEAN) : c.centre2;
    p.tie = c.tie_rule;
    p.want_moments = c.want_moments ? 1 : 0;
    p.flags = c.flags;
    p.essence1 = o.essence1;
    p.qr = o.qr;
    p.rel1 = o.rel1;
    p.reliable = o.reliable;
    p.consensus = o.consensus;
    p.rel2 = o.rel2;
    p.skew = o.skew;
    p.kurt = o.kurt;
    p.status = o.status;
    p.rank = o.rank;
    p.rb = pl.rb;
    p.words = (c.n_oracles + 31) / 32;
    p.tile_smem = pl.tile_smem;
    p.wP = pl.wP; p.wRP = pl.wRP; p.wCS = pl.wCS; p.wH = pl.wH;
    const size_t round_bytes = (size_t)c.n_oracles * c.dimension * elem_size(c.dtype);
    p.wBulk = ((reinterpret_cast<uintptr_t>(pred) & 15) == 0 && (round_bytes & 15) == 0 && round_bytes < (1u << 20)) ? 1 : 0;
    return p;
}

int consensus_device(const svoc_config *cfg, const void *pred, int64_t T, const svoc_outputs *out, void *stream,
                     bool fixed) {
    if (int rc = validate(cfg, pred, T, out, fixed)) return rc;
    if (T == 0) return 0;
    const Plan pl = make_plan(*cfg, pred, out, T);
    const Params p = make_params(*cfg, pred, T, *out, pl);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (cfg->dtype) {
        case SVOC_I32: return consensus_i32(p, pl, st);
        case SVOC_I64: return consensus_i64(p, pl, st);
        case SVOC_F32: return consensus_f32(p, pl, st);
        default: return consensus_f64(p, pl, st);
    }
}

int montecarlo_device(const svoc_config *cfg, const svoc_gen_config *gen, int64_t T, int64_t *counts, double *sums,
                      void *stream) {
    if (!cfg || !gen || !counts || !sums) return fail("null argument");
    if (T < 0) return fail("negative round count");
    if (T == 0) return 0;
    if (!(cfg->dtype == SVOC_I32 || cfg->dtype == SVOC_F32 || cfg->dtype == SVOC_F64))
        return fail("montecarlo supports dtype I32, F32 and F64");
    if (cfg->n_oracles < 2 || cfg->dimension < 1 || cfg->n_failing < 0 || cfg->n_failing >= cfg->n_oracles)
        return fail("bad shape");
    if (gen->n_oracles != cfg->n_oracles || gen->dimension != cfg->dimension || gen->n_failing != cfg->n_failing)
        return fail("consensus and generator configurations disagree on the shape");
    if (gen->n_params < 1 || gen->n_params > SVOC_MAX_LAW_PARAMS) return fail("n_params out of range");
    if (gen->failing_prob != 0.0) return fail("the fused Monte-Carlo kernel draws exactly n_failing failing oracles; use svoc_generate for failing_prob");
    if (cfg->centre < 0 || cfg->centre > 2 || cfg->centre2 < -1 || cfg->centre2 > 2 || cfg->tie_rule < 0 || cfg->tie_rule > 1)
        return fail("bad centre / tie rule");
    if ((i64)cfg->n_oracles * cfg->dimension >= 0xFFFF0000LL) return fail("n_oracles * dimension too large");
    svoc_config c = *cfg;
    c.want_moments = 0;
    const size_t es = elem_size(c.dtype), vs = (c.dtype == SVOC_F32) ? 4 : 8;
    const int npad = npad_for(c.n_oracles);
    const int threads = npad ? 128 : 256;
    int rb = threads / c.dimension;
    if (rb < 1) rb = 1;
    const size_t budget = npad ? 44 * 1024 : 200 * 1024;
    while (rb > 1 && mc_smem_bytes(c.n_oracles, c.dimension, rb, es, vs) > budget) rb >>= 1;
    const size_t smem = mc_smem_bytes(c.n_oracles, c.dimension, rb, es, vs);
    if (smem > 220 * 1024) return fail("a round does not fit in shared memory for the fused Monte-Carlo kernel");
    Plan pl{};
    pl.npad = npad; pl.rb = rb; pl.threads = threads; pl.tile_smem = 1; pl.smem = smem;
    svoc_outputs none{};
    Params p = make_params(c, nullptr, T, none, pl);
    McParams m;
    m.gen = *gen;
    m.gen.dtype = c.dtype;
    m.rk = philox_round_keys(gen->seed);
    m.counts = reinterpret_cast<u64 *>(counts);
    m.sums = sums;
    m.n_blocks = (T + rb - 1) / rb;
    const int per_sm = (int)std::max<size_t>(1, std::min<size_t>(16, (200 * 1024) / std::max<size_t>(smem, 1)));
    const int grid = (int)std::min<i64>(m.n_blocks, (i64)sm_count() * per_sm);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (c.dtype) {
        case SVOC_I32: return montecarlo_i32(p, m, npad, threads, smem, grid, st);
        case SVOC_F32: return montecarlo_f32(p, m, npad, threads, smem, grid, st);
        default: return montecarlo_f64(p, m, npad, threads, smem, grid, st);
    }
}

void set_error(const std::string &m) { g_err = m; }
void count_launch() {
    std::lock_guard<std::mutex> lk(g_mu);
    ++g_launches;
}

}  // namespace svoc

// ------------------------------------------------------------------------------ C ABI
extern "C" {

int svoc_consensus_fixed(const svoc_config *cfg, const void *pred, int64_t n_rounds, const svoc_outputs *out,
                         void *stream) {
    return svoc::consensus_device(cfg, pred, n_rounds, out, stream, true);
}

int svoc_consensus_float(const svoc_config *cfg, const void *pred, int64_t n_rounds, const svoc_outputs *out,
                         void *stream) {
    return svoc::consensus_device(cfg, pred, n_rounds, out, stream, false);
}

int svoc_montecarlo(const svoc_config *cfg, const svoc_gen_config *gen, int64_t n_rounds, int64_t *counts,
                    double *sums, void *stream) {
    return svoc::montecarlo_device(cfg, gen, n_rounds, counts, sums, stream);
}

const char *svoc_last_error(void) { return svoc::g_err.c_str(); }
int svoc_version(void) { return SVOC_VERSION; }
int64_t svoc_launch_count(void) {
    std::lock_guard<std::mutex> lk(svoc::g_mu);
    return svoc::g_launches;
}
const char *svoc_kernel_variant(const svoc_config *cfg) {
    static thread_local char buf[32];
    if (!cfg) return "invalid";
    const int np = svoc::npad_for(cfg->n_oracles);
    svoc::Plan pl{};
    if (svoc::plan_wide3(*cfg, nullptr, nullptr, 1, pl)) snprintf(buf, sizeof buf, "wide3");
    else if (svoc::plan_mid(*cfg, nullptr, 1, pl)) snprintf(buf, sizeof buf, "mid");
    else if (svoc::plan_wide2(*cfg, nullptr, pl)) snprintf(buf, sizeof buf, "wide2");
    else if (svoc::plan_wide(*cfg, nullptr, pl)) snprintf(buf, sizeof buf, "wide");
    else if (np) snprintf(buf, sizeof buf, "small%d", np);
    else snprintf(buf, sizeof buf, "generic");
    return buf;
}

}  // extern "C"
