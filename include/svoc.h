/* svoc.h — C ABI of libsvoc: batched two-pass robust oracle consensus on B200 (sm_100a).
 *
 * The reference (Ophiase/Stochastic-Vector-Oracle-Consensus) has no FFI layer: the
 * consensus update is reached through a Cairo contract ABI, a few notebook functions
 * and a few client helpers.  Each entry point below names the reference interface it
 * replaces (paths relative to the reference checkout).  A "round" is one full update
 * with all N oracle vectors present, i.e. what contract/src/contract.cairo:442-503
 * (constrained) / :370-434 (unconstrained) compute once every oracle has committed.
 *
 * Conventions
 *   - device entry points take DEVICE pointers and a cudaStream_t passed as void*;
 *     `_host` entry points take HOST pointers and do their own staging;
 *   - predictions are row-major [T][N][D]  (trial, oracle, dimension);
 *   - optional outputs may be NULL;
 *   - return value 0 = ok, negative = bad argument / CUDA error (text via
 *     svoc_last_error()); per-round data errors (the contract's panics) go to
 *     status[T], never to the return code;
 *   - no CPU fallback exists: without a CUDA device every compute call fails;
 *   - float form: inputs must be finite (NaN / inf are not detected; the results of such a round are unspecified).
 */
#ifndef SVOC_H
#define SVOC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVOC_VERSION 101

/* element types of the prediction tensor */
#define SVOC_I32 0
#define SVOC_I64 1
#define SVOC_F32 2
#define SVOC_F64 3

/* centre estimators */
#define SVOC_CENTRE_SMOOTH_MEDIAN_CAIRO 0 /* contract/src/math.cairo:113-126 (3-value branch is dead code) */
#define SVOC_CENTRE_MEDIAN 1              /* np.median, ipynb beta :355, client/oracle_scheduler.py:108 */
#define SVOC_CENTRE_MEAN 2                /* math.cairo:256-269; gaussian_algorithm_demo.ipynb:191-204 */
#define SVOC_CENTRE_DEFAULT (-1)          /* second pass: same as first when constrained, MEAN otherwise */

/* which of two oracles with EQUAL divergence stays reliable at the cut */
#define SVOC_TIE_HIGH_INDEX_KEPT 0 /* contract/src/sort.cairo:96-102: order (value asc, index desc) */
#define SVOC_TIE_LOW_INDEX_KEPT 1  /* stable argsort, ipynb beta :342-367 */

/* per-round status = first panic in the contract's evaluation order */
#define SVOC_ST_OK 0
#define SVOC_ST_INTERVAL_INPUT 1    /* math.cairo:298-310 via contract.cairo:591-593 */
#define SVOC_ST_SQRT_REL1 2         /* math.cairo:271-292: sqrt(1) divides by zero */
#define SVOC_ST_MAX_SPREAD_ZERO 3   /* contract.cairo:365-368 */
#define SVOC_ST_REL1_RANGE 4        /* math.cairo:294-296 after contract.cairo:467 / :397 */
#define SVOC_ST_SECOND_PASS_SHAPE 5 /* too few reliable rows for the second-pass centre: < 2 for the smooth median
                                       (math.cairo:117 reads sorted[mid-1]), < 1 for np.median / the mean */
#define SVOC_ST_SQRT_REL2 6
#define SVOC_ST_REL2_RANGE 7        /* contract.cairo:488 / :419 */
#define SVOC_ST_SKEW_PANIC 8        /* math.cairo:320-338: variance in {0,1} or R <= 2 */
#define SVOC_ST_KURT_PANIC 9        /* math.cairo:340-363: R <= 3 */
#define SVOC_ST_OVERFLOW 10         /* i128 overflow in the contract */
#define SVOC_ST_FIRST_PASS_SHAPE 11 /* N < 2 */
#define SVOC_ST_ENGINE_RANGE 12     /* |x| > 2^30 (unconstrained): beyond this engine's 64-bit fast path */

typedef struct svoc_config {
    int32_t n_oracles;        /* N  — constructor `oracles.len()`, contract.cairo:236-246 */
    int32_t dimension;        /* D  — constructor `dimension` */
    int32_t n_failing;        /* F  — constructor `n_failing_oracles`; R = N - F rows stay reliable */
    int32_t constrained;      /* constructor `constrained` */
    int64_t max_spread_wsad;  /* constructor `unconstrained_max_spread` (fixed form, 1e6-scaled) */
    double max_spread;        /* same, float form */
    int32_t centre;           /* SVOC_CENTRE_* for the first pass */
    int32_t centre2;          /* SVOC_CENTRE_* for the second pass or SVOC_CENTRE_DEFAULT */
    int32_t tie_rule;         /* SVOC_TIE_* */
    int32_t dtype;            /* SVOC_I32/I64 (fixed form) or SVOC_F32/F64 (float form) */
    int32_t want_moments;     /* also compute skewness / kurtosis (contract.cairo:491-500) */
    int32_t flags;            /* SVOC_FLAG_* */
} svoc_config;

/* constrained reliability of the OBSOLETE contract, 1 - 2*sqrt(avg qr) without the division by the dimension
 * (contract/obsolete/src/contract_nd.cairo:418, :437); it is the formula behind the "0.798" comment of
 * contract/tests/test_contract.cairo:188 */
#define SVOC_FLAG_OBSOLETE_RELIABILITY 1

/* Output block shared by the device entry points.  Element type of the value outputs:
 * int64 for the fixed form; the input type (float / double) for the float form.
 * All pointers except consensus, rel1, rel2 and status may be NULL. */
typedef struct svoc_outputs {
    void *essence1;     /* [T][D] first-pass essence        — essence_first_pass, contract.cairo:459 */
    void *qr;           /* [T][N] per-oracle quadratic risk — quadratic_risk_values, contract.cairo:463 */
    void *rel1;         /* [T]    get_first_pass_consensus_reliability(), contract.cairo:633-636 */
    uint32_t *reliable; /* [T][ceil(N/32)] bit i of word i/32 = oracle i reliable — OracleInfo.reliable */
    void *consensus;    /* [T][D] get_consensus_value(), contract.cairo:619-631 */
    void *rel2;         /* [T]    get_second_pass_consensus_reliability(), contract.cairo:638-641 */
    void *skew;         /* [T][D] get_skewness(), contract.cairo:643-655 */
    void *kurt;         /* [T][D] get_kurtosis() */
    int32_t *status;    /* [T]    SVOC_ST_* */
    int32_t *rank;      /* [T][N] position of every oracle in the selection order (0 = smallest qr, ties by
                           tie_rule) — `ordered_oracles`, contract.cairo:469; N-1-rank is `rank_array`'s
                           second return value (ipynb beta :342-352) */
} svoc_outputs;

/* Fixed-point ("wsad", value * 1e6 as signed integer) form — bit-exact restatement of
 * update_constrained_consensus / update_unconstrained_consensus
 * (contract/src/contract.cairo:442-503 / :370-434) with math.cairo, sort.cairo and
 * signed_decimal.cairo underneath.  cfg->dtype is SVOC_I32 or SVOC_I64. */
int svoc_consensus_fixed(const svoc_config *cfg, const void *pred, int64_t n_rounds,
                         const svoc_outputs *out, void *stream);

/* Floating-point form: the notebook estimator (identify_failing_oracles +
 * restricted_median, ipynb beta :354-371; predictions_to_eel_values,
 * client/oracle_scheduler.py:106-111) and the contract evaluated in real arithmetic.
 * cfg->dtype is SVOC_F32 or SVOC_F64. */
int svoc_consensus_float(const svoc_config *cfg, const void *pred, int64_t n_rounds,
                         const svoc_outputs *out, void *stream);

/* Same two calls with HOST buffers: predictions are staged to the device in chunks on
 * internal streams and the outputs copied back; the call returns when the outputs are
 * complete.  This is the call a client holding numpy arrays makes
 * (client/contract.py:200-208 `update_all_the_predictions` is its on-chain analogue).
 * out->skew / out->kurt are written only when cfg->want_moments is set (left untouched otherwise).  When an
 * error is returned no copy from or into the caller's buffers is still in flight. */
int svoc_consensus_fixed_host(const svoc_config *cfg, const void *pred, int64_t n_rounds,
                              const svoc_outputs *out);
int svoc_consensus_float_host(const svoc_config *cfg, const void *pred, int64_t n_rounds,
                              const svoc_outputs *out);

/* svoc_consensus_float_host for callers that must not allocate device memory (no counterpart in
 * the reference; it exists for hosts of this library, not for parity).  Same arguments, same
 * results.  The two calls above grow a device staging cache with cudaMalloc; a plugin inside a
 * process whose device memory is owned by someone else, code that runs after the memory map
 * is frozen, or a latency-bound loop (cudaMalloc / cudaFree take driver-wide locks and may
 * synchronise the device) cannot afford that.  This variant stages through a buffer built into
 * the library image — SVOC_NOALLOC_SCRATCH_BYTES of device memory on every GPU that loads the
 * library, held until the process exits — runs on the legacy default stream, and calls no
 * allocator.  A batch whose predictions plus requested outputs (each area padded to 256 bytes)
 * exceed that size is rejected with an error that says how many rounds fit.
 * The buffer is shared by all calls on a device and keeps the last call's contents; nothing in it
 * needs to be zero (every byte that is read was written earlier in the same call).  Calls that
 * overlap on the GPU race: one thread at a time per device.  The first call on a device looks
 * up the buffer's device address (cudaGetSymbolAddress); make it before any stream capture. */
#define SVOC_NOALLOC_SCRATCH_BYTES (16u << 20)
int svoc_consensus_float_host_noalloc(const svoc_config *cfg, const void *pred, int64_t n_rounds,
                                      const svoc_outputs *out);

/* 1 - 2*mean(|A - median(A)|) over each flattened [N*D] array — `expected_reliability`,
 * ipynb beta :260-267.  pred is [T][n_values] float64 on the device; out is [T] float64. */
int svoc_expected_reliability(const double *pred, int64_t n_arrays, int32_t n_values, double *out,
                              void *stream);

/* `rank_array` (ipynb beta :342-352 == client/oracle_scheduler.py:94-104) for T independent lists of n
 * float64 values on the device: ranks[t][i] = n - 1 - (position of value i in a stable ascending sort),
 * i.e. the largest value gets rank 0. */
int svoc_rank_array(const double *values, int64_t n_lists, int32_t n, int32_t *ranks, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Synthetic workloads — the notebook generators on a counter-based Philox4x32-10 stream:
 *   generate_beta_oracles / generate_2d_beta_oracles   ipynb beta :37-68
 *   generate_2d_gaussian_oracles                        gaussian_distribution_for_tests.ipynb:37-52
 * Honest oracles draw law(a[d], b[d]) per component, the n_failing failing ones are uniform (or
 * collude on one shared vector), and failing POSITIONS are a uniformly random subset, which is what
 * the reference's `np.random.shuffle` of [failing..., honest...] yields.  Every draw is a pure
 * function of (seed, global trial index, oracle, dimension): results do not depend on the launch
 * geometry or on how trials are sharded over GPUs.  numpy's MT19937 stream is not reproduced;
 * parity with the reference generators is distributional. */
#define SVOC_LAW_BETA 0      /* honest ~ beta(a, b); failing ~ uniform(0, 1) */
#define SVOC_LAW_GAUSSIAN 1  /* honest ~ normal(a = mu, b = sigma); failing ~ uniform(mu, 5 sigma) */
#define SVOC_LAW_KUMARASWAMY 2 /* honest ~ Kumaraswamy(a, b) by inverse CDF (only its mode formula is in the reference,
                                  ipynb beta :77-78); failing ~ uniform(0, 1) */
#define SVOC_LAW_ARCTAN_GAUSSIAN 3 /* honest = arctan(normal(tan(pi (a - 1/2)), b)) / pi + 1/2 with a = essence in (0,1),
                                  b = sigma (generate_oracles, gaussian_algorithm_demo.ipynb:46-64); failing ~ uniform(0, 1) */
#define SVOC_FAIL_UNIFORM 0
#define SVOC_FAIL_COLLUDING 1 /* every failing oracle submits clip(mode + shift) (beta) / mu + shift*sigma (gaussian) */
#define SVOC_MAX_LAW_PARAMS 8

typedef struct svoc_gen_config {
    int32_t n_oracles, dimension, n_failing;
    int32_t law;                  /* SVOC_LAW_* */
    int32_t failing;              /* SVOC_FAIL_* */
    int32_t dtype;                /* output type: SVOC_F32/F64, or SVOC_I32/I64 = int(x * 1e6) (to_wsad, ipynb beta :250) */
    int32_t n_params;             /* 1..SVOC_MAX_LAW_PARAMS; dimension d uses a[d % n_params], b[d % n_params] */
    int32_t reserved;
    double a[SVOC_MAX_LAW_PARAMS];
    double b[SVOC_MAX_LAW_PARAMS];
    double shift;                 /* colluders' offset */
    uint64_t seed;
    int64_t trial_offset;         /* global index of the first trial of this call (multi-GPU sharding) */
    double failing_prob;          /* 0: exactly n_failing oracles fail, at random positions (the notebook generators).
                                     > 0: every oracle fails independently with this probability (the failing-fraction
                                     sweep of BASELINE config 2; documentation/README.md:105-114 "an oracle is failing
                                     with probability p"); n_failing is ignored by the generator then */
} svoc_gen_config;

/* pred: [T][N][D] of cfg->dtype on the device; honest (may be NULL): [T][ceil(N/32)] bit mask, bit = 1
 * for honest oracles (the `validity` / `true_oracles` array the notebook generators return). */
int svoc_generate(const svoc_gen_config *cfg, int64_t n_rounds, void *pred, uint32_t *honest, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Per-oracle divergence accumulators and reliability histograms over many rounds — the inputs
 * of a replacement decision (contract.cairo:345-363 keeps only the last `reliable` flag; the
 * admin vote contract.cairo:547-580, 661-738 is left to humans).  All counters live in two
 * caller-owned device buffers so that several GPUs can combine them with one SUM all-reduce each:
 *   counts  int64  [SVOC_STATS_COUNTS_BASE + 2*N]
 *           [0] rounds  [1] rounds with status 0  [2..15] rounds per status code
 *           [16..271] histogram of rel1, 256 bins over [0,1]   [272..527] histogram of rel2
 *           [528..528+N) per-oracle count of rounds in which the oracle was NOT reliable
 *           [528+N..528+2N) per-oracle sum of qr (fixed form: exact wsad integers; float form: unused)
 *   sums    float64 [SVOC_STATS_SUMS_BASE + N]
 *           [0] sum rel1  [1] sum rel2  [2] sum rel1^2  [3] sum rel2^2  [4..4+N) per-oracle sum of qr (float form)
 * Only rounds with status 0 enter the histograms, sums and per-oracle accumulators. */
#define SVOC_STATS_COUNTS_BASE 528
#define SVOC_STATS_SUMS_BASE 4
int svoc_accumulate_stats(const svoc_config *cfg, int64_t n_rounds, const svoc_outputs *out, int64_t *counts,
                          double *sums, void *stream);

/* Fused Monte-Carlo run — the reference's `benchmark` loop (ipynb beta :385-402) as one persistent kernel:
 * every round is drawn (as svoc_generate would, same Philox stream), reduced by the consensus round and
 * folded into the accumulators without its predictions ever reaching HBM.  `cfg` describes the consensus
 * (dtype SVOC_I32, SVOC_F32 or SVOC_F64 = the element type the predictions are drawn in), `gen` the
 * workload; both must agree on n_oracles / dimension / n_failing.  counts / sums use the
 * svoc_accumulate_stats layout with two extra slots each:
 *   counts[528 + 2N]     rounds whose reliable set equals the honest set (identification success)
 *   counts[528 + 2N + 1] rounds that entered the distance statistics (status 0)
 *   sums[4 + N]          sum over those rounds of ||consensus - centre of the honest rows||
 *   sums[4 + N + 1]      sum of its square
 * so that 100 * (1 - 2 * sums[4+N] / counts[528+2N+1]) is the benchmark's "true consensus reliability". */
#define SVOC_MC_EXTRA 2
int svoc_montecarlo(const svoc_config *cfg, const svoc_gen_config *gen, int64_t n_rounds, int64_t *counts,
                    double *sums, void *stream);

const char *svoc_last_error(void);
int svoc_version(void);
/* number of kernels this library has launched in the calling process (for bench.py's gpu_launches) */
int64_t svoc_launch_count(void);
/* name of the kernel variant the dispatcher picks for (cfg), for diagnostics and bench labels: "small8" .. "small64"
 * (N <= 64), "wide3" / "wide2" / "wide" (64 < N <= 512, 32-bit elements) or "generic" */
const char *svoc_kernel_variant(const svoc_config *cfg);

#ifdef __cplusplus
}
#endif
#endif /* SVOC_H */
