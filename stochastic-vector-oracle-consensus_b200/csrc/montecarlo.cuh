// montecarlo.cuh — fused generate -> consensus -> accumulate (the reference's `benchmark` loop,
// ipynb beta :385-402, as one persistent kernel).  Predictions are drawn straight into shared memory
// and never touch HBM; the only global traffic is one flush of the CTA's accumulators at the end.
// For N <= 64 (the small<NPAD> variant) and the generic variant.
//
// Accumulators (include/svoc.h): the `svoc_accumulate_stats` layout plus, at the end,
//   counts[528 + 2N]     rounds whose reliable set equals the honest set ("identification success")
//   counts[528 + 2N + 1] rounds that entered the distance statistics
//   sums[4 + N]          sum of ||consensus - centre of the honest rows||  (the benchmark's `distance`)
//   sums[4 + N + 1]      sum of its square
#pragma once
#include "consensus_kernel.cuh"
#include "generate_device.cuh"

namespace svoc {

struct McParams {
    svoc_gen_config gen;
    u64 *counts;
    double *sums;
    i64 n_blocks;
    PhiloxKeys rk;
};

template <class Tr, int NPAD>
__global__ void __launch_bounds__(NPAD == 0 ? 256 : 128, (NPAD > 0 && NPAD <= 32 && !Tr::fixed && sizeof(typename Tr::acc_t) == 4) ? 8 : 1)
montecarlo_kernel(const Params p, const McParams m) {
    using in_t = typename Tr::in_t;
    using val_t = typename Tr::val_t;
    using acc_t = typename Tr::acc_t;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int N = p.N, D = p.D, R = p.R, F = p.F, rb = p.rb;
    const int tid = threadIdx.x, nt = blockDim.x;
    const size_t round_elems = (size_t)N * D;
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };

    size_t off = 0;
    in_t *tile = reinterpret_cast<in_t *>(smem_raw);            off += al((size_t)rb * round_elems * sizeof(in_t));
    val_t *c1 = reinterpret_cast<val_t *>(smem_raw + off);      off += al((size_t)rb * D * sizeof(val_t));
    acc_t *qr_s = reinterpret_cast<acc_t *>(smem_raw + off);    off += al((size_t)rb * N * sizeof(acc_t));
    val_t *cons_s = reinterpret_cast<val_t *>(smem_raw + off);  off += al((size_t)rb * D * sizeof(val_t));
    val_t *rel_s = reinterpret_cast<val_t *>(smem_raw + off);   off += al((size_t)2 * rb * sizeof(val_t));
    int *stw_s = reinterpret_cast<int *>(smem_raw + off);       off += al((size_t)rb * sizeof(int));
    int32_t *status_s = reinterpret_cast<int32_t *>(smem_raw + off); off += al((size_t)rb * sizeof(int));
    uint32_t *keys = reinterpret_cast<uint32_t *>(smem_raw + off);   off += al((size_t)rb * N * 4);
    unsigned *acc = reinterpret_cast<unsigned *>(smem_raw + off);    off += al((size_t)(16 + 512 + N + 2) * 4);
    uint8_t *keep_s = smem_raw + off;                           off += al((size_t)rb * N);
    uint8_t *honest_s = smem_raw + off;

    const int nacc = 16 + 512 + N + 2;
    for (int k = tid; k < nacc; k += nt) acc[k] = 0;
    double s_rel1 = 0, s_rel2 = 0, q_rel1 = 0, q_rel2 = 0, s_dist = 0, q_dist = 0;

    Params q = p;              // outputs of a block go to shared memory
    q.essence1 = nullptr; q.qr = nullptr; q.reliable = nullptr; q.skew = nullptr; q.kurt = nullptr; q.rank = nullptr;
    q.want_moments = 0;
    q.rel1 = rel_s; q.rel2 = rel_s + rb; q.consensus = cons_s; q.status = status_s;

    for (i64 blk = blockIdx.x; blk < m.n_blocks; blk += gridDim.x) {
        const i64 r0 = blk * rb;
        const int nr = (int)min((i64)rb, p.T - r0);
        __syncthreads();   // the previous block's epilogue is done with the shared arrays
        // ---- generate: failing positions, then every component
        for (int r = tid; r < rb; r += nt) stw_s[r] = STW_OK;
        for (int w = tid; w < nr * N; w += nt) {
            const int r = w / N, i = w - r * N;
            const u64 t = (u64)(m.gen.trial_offset + r0 + r);
            keys[w] = position_key((uint32_t)t, (uint32_t)(t >> 32), i, m.rk);
        }
        __syncthreads();
        for (int w = tid; w < nr * N; w += nt) {
            const int r = w / N, i = w - r * N;
            const int rank = rank_of<uint32_t, 1>(keys + (size_t)r * N, N, i, false);   // ties: lower index first
            honest_s[w] = rank < F ? 0 : 1;
        }
        __syncthreads();
        fill_components<in_t, true>(m.gen, honest_s, N, D, nr, m.gen.trial_offset + r0, tile, tid, nt, m.rk);
        __syncthreads();
        // ---- the consensus round(s)
        consensus_block<Tr, NPAD>(q, tile, r0, nr, r0, c1, qr_s, stw_s, keep_s);
        __syncthreads();
        // ---- centre of the honest rows (restricted_median(oracles, true_oracles), ipynb beta :395), into c1.
        //      Where the reliable set IS the honest set (most rounds) that centre is the consensus just computed —
        //      the same function of the same rows — and the third sort of every column is skipped.
        for (int r = tid; r < nr; r += nt) {
            bool same = true;
            for (int i = 0; i < N; ++i) same &= keep_s[r * N + i] == honest_s[r * N + i];
            stw_s[r] = same ? 1 : 0;           // consensus_block is done with its status words
        }
        __syncthreads();
        for (int c = tid; c < nr * D; c += nt) {
            const int r = c / D, d = c - r * D;
            if (stw_s[r]) { c1[c] = cons_s[c]; continue; }
            const in_t *col = tile + (size_t)r * round_elems + d;
            val_t sk, ku;
            int stw = STW_OK;
            if constexpr (NPAD > 0)
                c1[c] = column_centre_net<Tr, NPAD>(col, D, N, honest_s + (size_t)r * N, N - F, p.centre2, false, sk, ku, stw);
            else
                c1[c] = column_centre_generic<Tr>(col, D, N, honest_s + (size_t)r * N, N - F, p.centre2, false, sk, ku, stw);
        }
        for (int w = tid; w < nr * N; w += nt)
            if (status_s[w / N] == 0 && !keep_s[w]) atomicAdd(&acc[16 + 512 + (w % N)], 1u);
        __syncthreads();
        // ---- per-round statistics
        for (int r = tid; r < nr; r += nt) {
            const int st = status_s[r];
            atomicAdd(&acc[(st >= 0 && st < 14) ? 2 + st : 15], 1u);
            if (st != 0) continue;
            double a, b;
            if constexpr (Tr::fixed) { a = (double)rel_s[r] * 1e-6; b = (double)rel_s[rb + r] * 1e-6; }
            else { a = (double)rel_s[r]; b = (double)rel_s[rb + r]; }
            int ba = (int)(a * 256), bb = (int)(b * 256);
            ba = ba < 0 ? 0 : (ba > 255 ? 255 : ba);
            bb = bb < 0 ? 0 : (bb > 255 ? 255 : bb);
            atomicAdd(&acc[16 + ba], 1u);
            atomicAdd(&acc[16 + 256 + bb], 1u);
            s_rel1 += a; s_rel2 += b; q_rel1 += a * a; q_rel2 += b * b;
            const bool same = stw_s[r] != 0;
            double d2 = 0;
            for (int d = 0; d < D; ++d) {
                double df = (double)cons_s[r * D + d] - (double)c1[r * D + d];
                if constexpr (Tr::fixed) df *= 1e-6;
                d2 += df * df;
            }
            const double dist = sqrt(d2);
            s_dist += dist; q_dist += dist * dist;
            if (same) atomicAdd(&acc[16 + 512 + N], 1u);
            atomicAdd(&acc[16 + 512 + N + 1], 1u);
        }
    }
    __syncthreads();
    // ---- flush: block-reduce the float sums, one global atomic per touched counter
    __shared__ double red[6][8];
    double v[6] = {s_rel1, s_rel2, q_rel1, q_rel2, s_dist, q_dist};
    for (int k = 0; k < 6; ++k)
        for (int o = 16; o; o >>= 1) v[k] += __shfl_xor_sync(0xffffffffu, v[k], o);
    if ((tid & 31) == 0)
        for (int k = 0; k < 6; ++k) red[k][tid >> 5] = v[k];
    __syncthreads();
    if (tid < 6) {
        double s = 0;
        for (int w = 0; w < (nt >> 5); ++w) s += red[tid][w];
        if (s != 0) atomicAdd(&m.sums[tid < 4 ? tid : SVOC_STATS_SUMS_BASE + N + (tid - 4)], s);
    }
    for (int k = tid; k < nacc; k += nt) {
        const unsigned c = acc[k];
        if (!c) continue;
        int slot;
        if (k < 16 + 512) slot = k;
        else if (k < 16 + 512 + N) slot = SVOC_STATS_COUNTS_BASE + (k - 16 - 512);
        else slot = SVOC_STATS_COUNTS_BASE + 2 * N + (k - 16 - 512 - N);
        atomicAdd(&m.counts[slot], (u64)c);
        if (k >= 2 && k < 16) atomicAdd(&m.counts[0], (u64)c);
        if (k == 2) atomicAdd(&m.counts[1], (u64)c);
    }
}

inline size_t mc_smem_bytes(int N, int D, int rb, size_t es, size_t vs) {
    auto al = [](size_t x) { return (x + 15) & ~(size_t)15; };
    size_t s = 0;
    s += al((size_t)rb * N * D * es);
    s += al((size_t)rb * D * vs);
    s += al((size_t)rb * N * vs);
    s += al((size_t)rb * D * vs);
    s += al((size_t)2 * rb * vs);
    s += 2 * al((size_t)rb * 4);
    s += al((size_t)rb * N * 4);
    s += al((size_t)(16 + 512 + N + 2) * 4);
    s += 2 * al((size_t)rb * N);
    return s;
}

template <class Tr, int NPAD>
static int launch_mc(const Params &p, const McParams &m, int threads, size_t smem, int grid, cudaStream_t st) {
    auto kern = montecarlo_kernel<Tr, NPAD>;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail(std::string("cudaFuncSetAttribute: ") + cudaGetErrorString(e));
    }
    kern<<<grid, threads, smem, st>>>(p, m);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(std::string("kernel launch: ") + cudaGetErrorString(e));
    count_launch();
    return 0;
}

template <class Tr>
int dispatch_mc(const Params &p, const McParams &m, int npad, int threads, size_t smem, int grid, cudaStream_t st) {
    switch (npad) {
        case 8: return launch_mc<Tr, 8>(p, m, threads, smem, grid, st);
        case 16: return launch_mc<Tr, 16>(p, m, threads, smem, grid, st);
        case 32: return launch_mc<Tr, 32>(p, m, threads, smem, grid, st);
        case 64: return launch_mc<Tr, 64>(p, m, threads, smem, grid, st);
        default: return launch_mc<Tr, 0>(p, m, threads, smem, grid, st);
    }
}

int montecarlo_i32(const Params &p, const McParams &m, int npad, int threads, size_t smem, int grid, cudaStream_t st);
int montecarlo_f32(const Params &p, const McParams &m, int npad, int threads, size_t smem, int grid, cudaStream_t st);
int montecarlo_f64(const Params &p, const McParams &m, int npad, int threads, size_t smem, int grid, cudaStream_t st);

}  // namespace svoc
