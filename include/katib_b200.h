/*
 * katib_b200.h -- C ABI of libkatib_b200.so: the B200 (sm_100a) arithmetic behind Katib's
 * goptuna suggestion service (TPE, CMA-ES, Sobol).
 *
 * The reference has no C FFI on this path: pkg/suggestion/v1beta1/goptuna is pure Go built with
 * CGO_ENABLED=0 (cmd/suggestion/goptuna/v1beta1/Dockerfile:18) and calls the Go module
 * github.com/c-bata/goptuna v0.8.0 (go.mod:8).  Each entry point below names the goptuna call the
 * reference makes at that spot (reference file:line = the call site a cgo shim would replace);
 * INTEGRATION.md shows the cgo stub.
 *
 * Conventions
 *   - every function returns 0 (KB_OK) or a negative kb_status; kb_last_error(ctx) gives the text,
 *     valid until the next call on that ctx (ctx == NULL: the creating thread's last kb_create error)
 *   - one kb_ctx per GPU per process (one process per GPU when sharding); calls on one ctx are
 *     serialised by an internal mutex; the library never keeps a caller pointer after returning
 *   - "internal representation" = goptuna's: one float64 per parameter; categorical = index into
 *     Choices, int = float64(int)   (converter.go:277-343)
 *   - host pointers unless the name ends in _device; *_device entry points take device pointers in
 *     the ctx's GPU memory and a cudaStream_t passed as void* (NULL = the ctx's own stream)
 *   - there is no CPU fallback: without a usable sm_100 device kb_create fails
 */
#ifndef KATIB_B200_H
#define KATIB_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct kb_ctx kb_ctx;

typedef enum {
    KB_OK = 0,
    KB_ERR_INVALID = -1,     /* bad argument */
    KB_ERR_CUDA = -2,        /* CUDA runtime / launch failure */
    KB_ERR_STATE = -3,       /* call order (e.g. scoring before a history was set) */
    KB_ERR_COMM = -4,        /* NCCL missing or failed */
    KB_ERR_UNSUPPORTED = -5, /* size beyond a stated limit */
    KB_ERR_OVERFLOW = -6     /* internal near-tie table overflowed (see kb_tpe_score_select) */
} kb_status;

/* Parameter kinds = the goptuna distribution types toGoptunaSearchSpace emits (converter.go:110-184) */
typedef enum {
    KB_UNIFORM = 0,     /* goptuna.UniformDistribution{Low,High}             DOUBLE, no step */
    KB_DISCRETE = 1,    /* goptuna.DiscreteUniformDistribution{Low,High,Q}   DOUBLE with step */
    KB_INT = 2,         /* goptuna.IntUniformDistribution{Low,High}          INT, no step */
    KB_STEP_INT = 3,    /* goptuna.StepIntUniformDistribution{Low,High,Step} INT with step */
    KB_CATEGORICAL = 4  /* goptuna.CategoricalDistribution{Choices}          CATEGORICAL and DISCRETE */
} kb_kind;

#define KB_MAX_PARAMS 1024
#define KB_MAX_CHOICES 4096
#define KB_SOBOL_MAX_DIM 1024

int kb_version(void);
const char *kb_last_error(const kb_ctx *ctx);

/* One context on CUDA device `device`.  Replaces goptuna.CreateStudy (converter.go:369): the ctx is
 * the per-experiment state (search space, history, optimiser state) held in HBM. */
int kb_create(int device, kb_ctx **out);
int kb_destroy(kb_ctx *ctx);
/* out[0]=SM count, out[1]=cc major, out[2]=cc minor, out[3]=SM clock kHz, out[4]=L2 bytes */
int kb_device_info(const kb_ctx *ctx, int64_t out[5]);
/* the ctx's own stream as a cudaStream_t (for event timing on the launching stream) */
void *kb_stream(const kb_ctx *ctx);

/* ---- multi-GPU: one process per GPU, NCCL over NVLink --------------------------------------- */
/* 128-byte NCCL unique id made on rank 0 and handed to every rank by the host plumbing. */
int kb_comm_unique_id(void *out128);
int kb_comm_init(kb_ctx *ctx, int world, int rank, const void *unique_id128);
int kb_comm_world(const kb_ctx *ctx, int *world, int *rank);

/* ---- search space ---------------------------------------------------------------------------
 * Replaces goptuna.StudyOptionDefineSearchSpace(searchSpace) (converter.go:360); the arrays are
 * toGoptunaSearchSpace's output (converter.go:110-184) in a fixed parameter order chosen by the
 * caller.  step[] is Q / Step (ignored for UNIFORM, INT, CATEGORICAL); n_choices[] only for
 * CATEGORICAL.  low/high for INT kinds are the integer bounds as doubles. */
int kb_space_set(kb_ctx *ctx, int n_params, const int32_t *kind, const double *low,
                 const double *high, const double *step, const int32_t *n_choices);

/* ---- TPE ------------------------------------------------------------------------------------
 * kb_tpe_set_history replaces the per-parameter history walk of tpe.Sampler.Sample
 * (getObservationPairs over study.GetTrials(); reached from sample.go:48-90): Complete trials only
 * (service.go:134-137), values[n_trials x n_params] row-major internal representation,
 * loss[n_trials] = objective already sign-flipped so that lower is better
 * (toGoptunaDirection, converter.go:31-36).  It uploads the history; the below/above split and both
 * Parzen estimators of every parameter are built on the GPU when an entry point first reads them
 * (a small kb_tpe_suggest on a history of up to 512 trials builds what it needs inside its own
 * kernel and never asks for the tables). */
int kb_tpe_set_history(kb_ctx *ctx, int64_t n_trials, const double *values, const double *loss);
/* Appends n_new trials to the history already on the device and rebuilds the estimators: the same result as
 * kb_tpe_set_history with the concatenated arrays, but only the new rows cross the boundary.  (The reference
 * re-reads every trial on every call, service.go:62-69 + sample.go:98; trials never leave a goptuna study, so a
 * history only grows.)  Needs a preceding kb_tpe_set_history (n_trials = 0 is fine). */
int kb_tpe_append_history(kb_ctx *ctx, int64_t n_new, const double *values, const double *loss);

/* Options of the TPE path.  KB_OPT_MERGE_DUPLICATES (default 0): components of an estimator that have identical
 * (mu, sigma) are folded into one with the summed weight before scoring.  The mixture is the same function (scores
 * agree to rounding) but has far fewer terms for integer / stepped parameters, whose histories are mostly repeats.
 * goptuna evaluates every observation separately; the default reproduces that count.  Tables of a history already
 * set are rebuilt at once.  kb_service_* switches it on. */
#define KB_OPT_MERGE_DUPLICATES 1
/* KB_OPT_PRUNE_FAR (default 0): continuous parameters only.  A component whose sigma sits on the magic-clip floor and
 * whose mu is more than 7.75 sigma from a candidate contributes less than e^-30 of its own peak to that candidate's
 * mixture value; with the option on such (candidate, component) pairs are not evaluated (candidates are bucketed by
 * value first, so that a thread block meets one window of the mu-sorted table).  Scores change by a few 1e-6
 * relative (the omitted terms are far smaller; the FP32 summation order differs), winners -- decided in FP64 on the unpruned mixture -- do not.  goptuna evaluates every pair; so does the
 * default. */
#define KB_OPT_PRUNE_FAR 2
/* KB_OPT_LOCAL_ORIGIN (default 1): continuous parameters, how x - mu is formed in the FP32 scoring kernel; every
 * (candidate, component) pair is evaluated either way and the winners are decided in FP64 as always.
 *   0  candidates stay in their order; x and mu are (hi, lo) FP32 pairs on the parameter's whole range.
 *   1  with a history of at least 2048 trials (below that the bucketing costs more than it saves), chunks of at
 *      least 32 768 candidates of a selection pass are first bucketed by value (as under KB_OPT_PRUNE_FAR), so that
 *      the 2048 candidates of a thread block lie close together; the block then measures candidates and components
 *      from its own origin in units of the narrow components' sigma, where one FP32 number is enough, and needs 3
 *      instead of 5 packed instructions per (candidate pair, component) for nearly every component.
 *      Blocks whose candidates span more than 1/128 of the range keep the (hi, lo) form.  Everything else as under 0,
 *      kb_tpe_scores included: bucketing uses atomics, so the last bits of an FP32 score would change from run to run.
 *   2  as 1 for every chunk size, history length and entry point (tests). */
#define KB_OPT_LOCAL_ORIGIN 3
int kb_tpe_set_option(kb_ctx *ctx, int option, int value);

/* Estimator of parameter `param`, side 0 = below (l), 1 = above (g): K entries sorted by mu.
 * Buffers hold `cap` doubles; *K is set even when cap is too small (then KB_ERR_INVALID).
 * For CATEGORICAL parameters mu/sigma are NULL-able and w receives the n_choices probabilities. */
int kb_tpe_get_parzen(kb_ctx *ctx, int param, int side, int *K, double *w, double *mu,
                      double *sigma, int cap, double *p_accept);

/* Score every candidate: scores[i*n_params + d] = log l(x) - log g(x) (FP32 arithmetic, FP64
 * accumulation) -- tpe.Sampler.gmmLogPDF / categoricalLogPDF on a caller-supplied
 * candidate set.  cand[n_cand x n_params] row-major, in the sampler's working domain: UNIFORM /
 * INT as is, DISCRETE and STEP_INT shifted by -low (sampleDiscreteUniform shifts to [0, high-low]
 * so that rounding to the step lands on the grid {low + k step}), CATEGORICAL the choice index. */
int kb_tpe_scores(kb_ctx *ctx, int64_t n_cand, const double *cand, double *scores);

/* tpe.Sampler.compare on a caller-supplied candidate set, per parameter: index of the winning
 * candidate (first maximum), the value Sampler.Sample would return for it (DISCRETE: shifted back
 * and clamped), and its FP64 score.  index_offset is added to reported indices (this rank's first
 * global candidate index when the set is sharded).  With a communicator set (kb_comm_init) every
 * rank passes its own slice and all ranks receive the global winners (one NCCL collective).
 * Winners are decided in FP64: the FP32 pass keeps every distinct value within a margin of the
 * running maximum and those are re-scored in FP64.  KB_ERR_OVERFLOW if that table fills. */
int kb_tpe_score_select(kb_ctx *ctx, int64_t n_cand, const double *cand, int64_t index_offset,
                        int64_t *out_index, double *out_value, double *out_score);
/* Same with candidates already in HBM (device pointer, same layout). */
int kb_tpe_score_select_device(kb_ctx *ctx, int64_t n_cand, const double *cand_device,
                               int64_t index_offset, int64_t *out_index, double *out_value,
                               double *out_score, void *stream);

/* No counterpart in the reference.  kb_tpe_scores for ONE continuous (UNIFORM) parameter, for callers that cannot
 * allocate GPU memory when they need scores (a plugin inside an application that owns all device memory through its
 * own pool, a process whose memory was handed out at start-up, a loop that must not hit cudaMalloc's device
 * synchronisation).  It needs no kb_ctx and calls no cudaMalloc / cudaFree / allocating routine: the caller passes the
 * candidates, both Parzen estimators (as kb_tpe_get_parzen returns them: w, mu, sigma, K entries per side) and the
 * output as device pointers on `device`; scores[i] = log l(x_i) - log g(x_i).  Scratch is a static __device__ array
 * inside the library (256 KiB + 256 B, resident on every GPU that loads the library until the process exits), which
 * limits each side to 4096 components; larger K is refused with KB_ERR_UNSUPPORTED.  The first call on a device
 * looks up the array's address (cudaGetSymbolAddress) -- make it before any stream capture starts.  All calls on a
 * device share the array: it keeps the previous call's contents (every call rewrites what it reads, nothing has to be
 * zero), and two calls that overlap on the GPU race, so serialise them.  Errors: kb_last_error(NULL). */
int kb_tpe_scores_noalloc(int device, int64_t n_cand, const double *cand_device, int K_below,
                          const double *w_below_device, const double *mu_below_device,
                          const double *sigma_below_device, int K_above, const double *w_above_device,
                          const double *mu_above_device, const double *sigma_above_device, double low,
                          double high, double *scores_device, void *stream);

/* Full tpe.Sampler.Sample for every parameter (sample.go:48-90): draw n_ei_candidates from l(x)
 * on the device (Philox4x32-10 keyed by seed; goptuna's math/rand stream is not reproducible),
 * score, select.  out_value[n_params] in internal representation.  The candidates are the ones
 * kb_tpe_sample_candidates returns for the same seed, but they are never stored: each chunk is drawn
 * straight into the scoring kernels' staging columns, so the scratch does not grow with
 * n_ei_candidates.  With a communicator (kb_comm_init) and n_ei_candidates >= 65536 every rank must
 * call with the same arguments: rank r draws and scores only its contiguous index range of the one
 * stream, one collective merges the winners and every rank returns the unsharded call's values.
 * Below 65536 candidates there is no collective and a rank may call alone. */
int kb_tpe_suggest(kb_ctx *ctx, int64_t n_ei_candidates, uint64_t seed, double *out_value);
/* Draw only (sampleFromGMM / sampleFromCategoricalDist): cand[n x n_params] row-major, working
 * domain.  cand_device may be NULL when cand_host is given and vice versa. */
int kb_tpe_sample_candidates(kb_ctx *ctx, int64_t n, uint64_t seed, double *cand_host,
                             double *cand_device, void *stream);

/* ---- Sobol ----------------------------------------------------------------------------------
 * sobol.Engine.Draw() n_points times starting at sequence index first_index (sobol.NewSampler,
 * converter.go:94; reached through trial.CallRelativeSampler, sample.go:39).  Joe-Kuo direction
 * numbers, 32 bits, gray-code order, point 0 = origin.  out[p*dim + d] = u in [0,1); with low/high
 * non-NULL, low[d] + u*(high[d]-low[d]) with separately rounded multiply and add (Go does not fuse). */
int kb_sobol_fill(kb_ctx *ctx, int dim, uint64_t first_index, int64_t n_points, const double *low,
                  const double *high, double *out);
int kb_sobol_fill_device(kb_ctx *ctx, int dim, uint64_t first_index, int64_t n_points,
                         const double *low, const double *high, double *out_device, void *stream);

/* ---- CMA-ES ---------------------------------------------------------------------------------
 * cmaes.NewOptimizer(mean, sigma, OptimizerOptionBounds, [popsize]) -- created by
 * cmaes.Sampler.SampleRelative (converter.go:41-68, sample.go:39).  popsize 0 = 4+floor(3 ln n).
 * lo/hi NULL = unbounded. One optimiser per ctx. */
int kb_cmaes_create(kb_ctx *ctx, int n, const double *mean, double sigma, int popsize,
                    const double *lo, const double *hi);
/* Optimizer.Ask() lam times in one batch: x[k*n+i] = m_i + sigma*(B diag(D) z_k)_i.  z[lam x n]
 * row-major supplied by the caller (NULL: drawn on the device, Philox keyed by seed).  With bounds:
 * caller-supplied z is never redrawn, out-of-bounds coordinates are clipped (repairInfeasibleParams);
 * device-drawn z is redrawn for an infeasible row up to 100 times and only then clipped, as the
 * cmaes library's ask() does. */
int kb_cmaes_ask(kb_ctx *ctx, int lam, const double *z, uint64_t seed, double *x);
/* Optimizer.Tell(solutions): x[popsize x n] row-major, f[popsize] lower is better.
 * With a communicator of W ranks (kb_comm_init) the population is sharded by rows: every rank creates the optimiser
 * with the same (global) popsize, asks for its popsize / W rows (rank r owns rows [r, r+1) * popsize / W; device-drawn
 * z continues one global Philox stream) and tells them with their popsize / W objective values.  The ranking is
 * global (one all-gather of f), y_w and the rank-mu sum are all-reduced, everything else -- mean, sigma, paths, C and
 * its eigensystem -- is replicated, so afterwards all ranks hold the same state (equal to the single-GPU state to
 * rounding).  popsize must be a multiple of W. */
int kb_cmaes_tell(kb_ctx *ctx, const double *x, const double *f);
/* Any pointer may be NULL.  B[i*n+j] = component i of eigenvector j, D = sqrt(eigenvalues), both
 * canonical (ascending, largest-|component| positive); they refer to the current C. */
int kb_cmaes_get_state(kb_ctx *ctx, double *mean, double *sigma, double *C, double *p_sigma,
                       double *p_c, double *B, double *D, int *generation);
/* Override the cached factors (parity hook: feed another solver's B, D). */
int kb_cmaes_set_eig(kb_ctx *ctx, const double *B, const double *D);
/* Symmetric eigendecomposition (parallel cyclic Jacobi, FP64) standing in for gonum mat.EigenSym:
 * A[n x n] row-major -> w ascending, V[i*n+j] = component i of eigenvector j. */
int kb_eigh(kb_ctx *ctx, int n, const double *A, double *w, double *V);

/* Jacobi sweeps used by the most recent factorisation on this ctx (kb_eigh or the one inside ask/tell). */
int kb_eigh_stats(const kb_ctx *ctx, int *sweeps);

/* ---- random -----------------------------------------------------------------------------------
 * goptuna.RandomSampler (converter.go:96-106): out[i*n_params + d] uniform over each parameter's
 * distribution, internal representation, Philox keyed by (seed, first_index + i). */
int kb_random_fill(kb_ctx *ctx, uint64_t seed, uint64_t first_index, int64_t n, double *out);

/* ---- the suggestion service above the kernels ------------------------------------------------
 * Host-side mirror of the reference's Go service (pkg/suggestion/v1beta1/goptuna/service.go:47-215 with
 * converter.go and sample.go): search-space and settings parsing, trial conversion and state mapping, trial-name
 * matching, the sampling loop and value formatting.  Requests and replies are serialized protobuf messages of
 * pkg/apis/manager/v1beta1/api.proto (GetSuggestionsRequest -> GetSuggestionsReply,
 * ValidateAlgorithmSettingsRequest -> empty reply), i.e. exactly the bytes the katib-controller puts on the wire, so
 * a gRPC front end only forwards them.  Both calls return the gRPC status code the reference returns at the same spot
 * (0 OK, 3 InvalidArgument, 13 Internal); kb_service_last_message gives the status text.  One kb_service per
 * experiment, like one reference service process per Suggestion CR (composer.go:72-146). */
typedef struct kb_service kb_service;
int kb_service_create(int device, kb_service **out);
/* The same service over several GPUs of this machine (devices[0 .. n_devices), distinct).  The controller talks to ONE
 * service process per experiment (composer.go:72-146; main.go:43-57), so the GPUs are driven from inside it: one
 * kb_ctx and one host thread per GPU.  Requests are answered by devices[0] alone, except TPE requests whose
 * n_ei_candidates setting is at least 65536: then every GPU draws and scores its contiguous index range of the
 * candidate stream and the per-GPU winners (D x 24 bytes each) are merged on the host by the argmax rule (best score,
 * lowest candidate index) -- the reply is the one a single-GPU service gives for the same request.  Inside one
 * process no NCCL communicator is used (the winners already share an address space); one process per GPU
 * (kb_comm_init) is the deployment that merges over NCCL.  Contexts are made at the first GetSuggestions, like the
 * reference's study (service.go:153-171). */
int kb_service_create_multi(int n_devices, const int *devices, kb_service **out);
int kb_service_destroy(kb_service *svc);
/* SuggestionService.GetSuggestions (service.go:54-93).  *reply points into memory owned by svc, valid until the
 * next call on it. */
int kb_service_get_suggestions(kb_service *svc, const void *request, size_t request_len, const void **reply,
                               size_t *reply_len);
/* SuggestionService.ValidateAlgorithmSettings (service.go:173-211); request == NULL is the nil request.
 * Needs no GPU. */
int kb_service_validate_algorithm_settings(kb_service *svc, const void *request, size_t request_len);
const char *kb_service_last_message(const kb_service *svc);
/* Introspection for tests: 0 CMA-ES population size in use, 1 restarts so far (IPOP / BIPOP), 2 internal trials,
 * 3 generation tag, 4 Complete observations, 5 history rows sent to the GPU so far (kb_tpe_set_history +
 * kb_tpe_append_history), 6 trials adopted under KB_SVC_ADOPT_UNKNOWN_TRIALS, 7 resent Trial messages recognised by
 * their bytes as already synced into a finished state and therefore not parsed again. */
long long kb_service_stat(kb_service *svc, int which);
/* KB_SVC_ADOPT_UNKNOWN_TRIALS (default 0): a trial whose parameters match no trial this service suggested fails
 * GetSuggestions with "Same parameter is not found for Trial: <name>" in the reference (sample.go:126) -- which is
 * what happens to every finished trial after the suggestion pod restarts, since the study lives in memory
 * (service.go:39-52).  With the option on, such a trial joins the study and, once finished, the history.  A superset
 * of the reference's behaviour, so it is off unless asked for. */
#define KB_SVC_ADOPT_UNKNOWN_TRIALS 1
int kb_service_set_option(kb_service *svc, int option, int value);

/* ---- measurement helpers (used by bench.py and the profiles) ----------------------------------
 * which: 0 = MUFU.EX2 Gop/s, 1 = FP32 FFMA GFMA/s, 2 = FP64 DFMA GFMA/s, 3 = HBM write GB/s,
 * 4 = HBM copy GB/s, 5 = FP32 GFMA/s through packed FFMA2, 6 = FFMA2 and FFMA interleaved (one pipe or two?),
 * 7 = FFMA2 whose accumulator and multiplicand both change every instruction (the operand pattern of the quantised
 * grid kernel).  Runs a saturating micro-kernel on the ctx's device and returns the rate. */
int kb_peak_measure(kb_ctx *ctx, int which, double *out);
/* Per-kernel-class device timing with CUDA events on the launching stream.  kb_profile(ctx, 1) starts
 * (and clears) the counters, kb_profile(ctx, 0) stops; kb_profile_read returns the summed duration and
 * the launch count of one class:  0 tpe_stage  1 tpe_score_cont  2 tpe_score_quant  3 tpe_rescore+pick
 * 4 sobol_fill  5 tpe_build  6 tpe_sample  7 cmaes_ask  8 cmaes_tell  9 eigh  10 collective  11 dgemm (DMMA) */
int kb_profile(kb_ctx *ctx, int enable);
int kb_profile_read(kb_ctx *ctx, int kernel_class, double *ms_total, int64_t *launches);
/* Worst error of the FP32 scoring pipeline over a candidate set in HBM (same layout as kb_tpe_score_select_device):
 * out_max_err[d] = max_i |score32(i,d) - score64(i,d)| / max(1, |score64(i,d)|) for numeric parameters (0 for
 * categorical ones), where score32 is what selection sees (the pipeline's score rounded to float) and score64 is the
 * FP64 formula the winners are decided with, evaluated on the device for EVERY candidate.  Costs ~100 FP64
 * operations per (candidate, component): minutes at the headline size.  The admission margins of the near-tie
 * table must stay above these numbers. */
int kb_tpe_score_error_device(kb_ctx *ctx, int64_t n_cand, const double *cand_device, double *out_max_err);
/* KB_OPT_LOCAL_ORIGIN: thread blocks of tpe_score_cont_local_kernel run on this context so far that used their own
 * origin (out_blocks[0]) and that kept the (hi, lo) form because their candidates were spread too widely
 * (out_blocks[1]). */
int kb_tpe_local_stats(kb_ctx *ctx, uint64_t *out_blocks);
/* number of kernels this library has launched on this ctx since creation */
int64_t kb_launch_count(const kb_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif
