/*
 * rnafold_b200.h -- C ABI of the B200-native replacement for the reference's reward hot path:
 * batched Zuker MFE folding (Turner nearest-neighbour model, ViennaRNA defaults dangles=2,
 * MAXLOOP=30, integer dcal/mol) + dot-bracket Hamming distance to a target structure.
 *
 * Plain pointers and sizes only; no torch/CUDA types.  Every entry point returns 0 on success or a
 * negative RNAFOLD_E* code and records a message retrievable with rnafold_last_error().
 *
 * Reference interfaces each entry point replaces (paths relative to the reference checkout):
 *   rnafold_fold_batch / rnafold_fold_hamming_batch
 *        <- `fold_fn(string) -> (structure, mfe)` = RNA.fold, bound at rlif/settings.py:62-71 and called
 *           at rlif/rna/solution.py:134,140,204; rlif/rna/utils.py:23; rlif/CLI.py:216;
 *           fused with hamming_distance (rlif/rna/utils.py:67-75) and the reward of
 *           Solution.evaluate (rlif/rna/solution.py:145-149).
 *   rnafold_read_parameter_file
 *        <- RNA.read_parameter_file(path): rlif/rna/vienna.py:18, rlif/CLI.py:21,
 *           rlif/interface/widgets.py:31, rlif/interface/parameters.py:10.
 *   rnafold_set_md
 *        <- setattr(RNA.cvar, name, value): rlif/rna/vienna.py:21-26, rlif/interface/widgets.py:25-26;
 *           names temperature, dangles, noGU, no_closingGU, uniq_ML (rlif/settings.py:35-39).
 *
 * Buffers: every input/output pointer may be a host pointer or a device pointer on the context's
 * device (detected with cudaPointerGetAttributes).  Host buffers are staged through pinned memory
 * inside the call; device buffers are used in place and the call is asynchronous on `stream`
 * (a cudaStream_t passed as void*, NULL = the context's own stream) unless any buffer is a host one.
 * One context per process: fold calls share its scratch, so calls issued on different streams are
 * ordered one behind the other on the device (each waits for the previous call's event); they never
 * overlap.  A kernel-side failure of an asynchronous call (backtrack found no decomposition) is sticky:
 * it is reported by the next call that returns host data, by the environment step's end-of-step
 * synchronisation, or by rnafold_last_status().
 *
 * Batches are CSR: `seq` holds the ASCII nucleotides of all B sequences back to back and `off[B+1]`
 * their start offsets (off[0] = 0).  Structures are written with the same offsets.
 */
#ifndef RNAFOLD_B200_H
#define RNAFOLD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RNAFOLD_OK 0
#define RNAFOLD_EINVAL -1   /* bad argument */
#define RNAFOLD_ECUDA -2    /* CUDA runtime error (sticky) */
#define RNAFOLD_EIO -3      /* parameter file unreadable / malformed */
#define RNAFOLD_ENOTSUP -4  /* model detail other than the reference's defaults */
#define RNAFOLD_ENOINIT -5  /* rnafold_init not called */

#define RNAFOLD_MAX_LEN 4000 /* longest sequence accepted */

/* Create the per-process context on CUDA device `device`, load the built-in Turner 2004
 * parameters (ViennaRNA's compiled-in default) and upload them.  Idempotent per device. */
int rnafold_init(int device);
void rnafold_shutdown(void);
const char *rnafold_last_error(void);
const char *rnafold_version(void);

/* Parse an "RNAfold parameter file v2.0", rescale for 37 C / dangles=2 and upload.  Takes effect
 * on the next fold, like the process-global state of RNA.read_parameter_file. */
int rnafold_read_parameter_file(const char *path);
/* Flat int32 dump of the live parameter tables in the oracle's export order (tests only);
 * returns the number of ints written (<= cap) or a negative code. */
int rnafold_export_params(int32_t *out, int cap);

/* Host-only variant: parse `path` (NULL = built-in Turner 2004) without touching a device and dump
 * the tables the same way.  Used by the CPU test-suite to check the loader. */
int rnafold_parse_params(const char *path, int32_t *out, int cap);

/* Model details (RNA.cvar).  Only the reference's defaults are implemented: temperature=37,
 * dangles=2, noGU=0, no_closingGU=0, uniq_ML=0.  Any other value -> RNAFOLD_ENOTSUP. */
int rnafold_set_md(const char *name, double value);
int rnafold_get_md(const char *name, double *value);

/* Bytes of device scratch a batch of B sequences with maximum length nmax will use. */
int rnafold_workspace_bytes(int32_t B, int32_t nmax, size_t *bytes);

/* Fold B sequences.  db_out: total-length bytes ('(', ')', '.'); mfe_dcal_out: int32[B] in
 * dcal/mol (RNA.fold's float is (float)(e/100.)). */
int rnafold_fold_batch(const uint8_t *seq, const int32_t *off, int32_t B, uint8_t *db_out,
                       int32_t *mfe_dcal_out, void *stream);

/* Fold + Hamming distance to `target` (same CSR offsets as seq) + reward
 * r = (1 - hd/len)^reward_exp (+0.25 if hd == 0), as rlif/rna/solution.py:145-148.
 * hd_out int32[B]; reward_out float64[B] (may be NULL); mismatch_out: one byte per position
 * (1 where target != structure; may be NULL); db_out / mfe_dcal_out may be NULL. */
int rnafold_fold_hamming_batch(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target,
                               uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out,
                               double *reward_out, uint8_t *mismatch_out, double reward_exp,
                               void *stream);

/* Allocation-free fold + Hamming + reward.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same results as rnafold_fold_hamming_batch, but the call allocates nothing -- no cudaMalloc,
 * cudaFree or growing buffer: the kernels' scratch (queue counters, device-side batch count, one range-check flag
 * per sequence) is a file-scope static __device__ array in the library image, (8 + RNAFOLD_NOALLOC_MAX_BATCH) * 4 =
 * 262,176 bytes, held by every GPU that loads the library until exit.
 * Limits: B <= RNAFOLD_NOALLOC_MAX_BATCH; every sequence <= nmax <= RNAFOLD_NOALLOC_MAX_LEN (the shared-memory
 * kernels; nmax sizes the launch, a longer sequence makes the call fold nothing and sets status code 100, see
 * rnafold_last_status); larger values are rejected with RNAFOLD_EINVAL.  EVERY buffer (seq, off, target, outputs)
 * must be a device pointer; outputs may be NULL as in rnafold_fold_hamming_batch (no mismatch flags).  Sequences
 * are folded in batch order (sort by length yourself for mixed batches).  Asynchronous on `stream`.
 * All calls share the one array, which keeps the last call's contents (the call clears what it needs): calls that
 * overlap on the GPU -- different streams without an ordering between them -- race.  Not ordered against the other
 * fold entry points either, except through the stream you pass. */
#define RNAFOLD_NOALLOC_MAX_BATCH 65536
#define RNAFOLD_NOALLOC_MAX_LEN 200
int rnafold_fold_hamming_batch_noalloc(const uint8_t *seq, const int32_t *off, int32_t B, int32_t nmax,
                                       const uint8_t *target, uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out,
                                       double *reward_out, double reward_exp, void *stream);

/* Wait for the last fold call and report the sticky device status word: 0, or the failure code of a kernel whose
 * backtrack found no decomposition (then the return value is RNAFOLD_ECUDA).  clear != 0 resets it. */
int rnafold_last_status(int32_t *status, int32_t clear);

/* Counters of the last fold call: kernel launches issued, and sequences routed to each kernel
 * family (0: shared-memory CTA kernel, 1: cluster/DSMEM kernel, 2: global-memory kernel). */
int rnafold_last_launch_info(int32_t *launches, int32_t routed[3]);

/* Enable (1) / disable (0) CUDA-event timing of the kernels of each fold call. */
int rnafold_set_timing(int enabled);

/* Time of the last fold call's kernels in milliseconds, measured with CUDA events on the stream
 * the kernels were launched on (synchronises the stream). */
int rnafold_last_kernel_ms(float *ms);

/* Device the context is bound to (-1 before rnafold_init). */
int rnafold_current_device(void);

/* ---- Batched RnaDesign environment on the device -------------------------------------------------
 * Replaces N reference environments (rlif/environments/RLIFenv.py:6-114, one per forked worker of
 * SubprocVecEnv, rlif/learning/trainer.py:523-565) by one lock-step batch.  Per-environment logic is
 * Solution.insert_nucleotide / find_next_unfilled / get_state / evaluate (rlif/rna/solution.py:84-155)
 * over targets parsed like DotBracket (rlif/rna/dotbracket.py:94-111, 171-211, encoding_type 2).
 * Observations are uint8 [N][10][2*kernel_size] (6 motif rows O C I H M E + 4 nucleotide rows A C G U),
 * the window encoding[:, i:i+2k] of the reference.  Buffers may be host or device pointers. */
int rnafold_env_create(int32_t n_envs, int32_t kernel_size, int32_t *handle);
int rnafold_env_destroy(int32_t handle);
/* auto_reset: re-arm finished environments inside step (VecEnv behaviour); randomize: hashed random target
 * choice instead of round-robin (RLIFenv.py:87-92; Python's RNG stream is not reproduced). */
int rnafold_env_config(int32_t handle, int32_t auto_reset, int32_t randomize, uint64_t seed);
/* env.set_attr('current_sequence', v) of the reference's evaluation loop (rlif/learning/tester.py:63): sets every
 * environment's round-robin position; the next reset takes target (v + 1) % n_targets (RLIFenv.py:89-90). */
int rnafold_env_set_current(int32_t handle, int32_t value);
/* Reward shaping of Solution.evaluate (rlif/rna/solution.py:150-151, config keys gc_penalty / gc):
 * r = r * 0.85 + clip(fraction of U, 0, 0.15) on the terminal step. */
int rnafold_env_config_reward(int32_t handle, int32_t gc_penalty);
/* RnaDesign.set_data: CSR list of dot-bracket targets (host pointers). */
int rnafold_env_set_targets(int32_t handle, const uint8_t *bytes, const int32_t *off, int32_t n_targets);
int rnafold_env_reset(int32_t handle, uint8_t *obs_out);
/* actions int32[N] in 0..3; reward_out float64[N]; done_out uint8[N]; any output may be NULL.  The step is
 * enqueued without any host round trip (the terminal batch and its size stay on the device); it synchronises once,
 * at its end, iff an output is a host buffer -- with device (or NULL) outputs it returns as soon as it is enqueued. */
int rnafold_env_step(int32_t handle, const int32_t *actions, double reward_exp, uint8_t *obs_out,
                     double *reward_out, uint8_t *done_out);
/* istat = {episodes, solved, sum HD, env steps}; fstat = {sum reward, sum MFE kcal/mol}. */
/* Deferred rewards for rollouts that keep ONE REWARD BUFFER PER STEP (what a PPO rollout stores anyway; the reference's
 * trainer consumes rewards after n_steps, rlif/learning/trainer.py:307-309): the terminal sequences of up to `steps` (1..64)
 * consecutive steps are folded by one call -- a step is bound by the latency of a fold, not by its throughput -- and every
 * reward lands in the reward_out buffer that was passed to the step it belongs to, at the latest when the window closes:
 * after `steps` steps, on rnafold_env_flush, or before any call that reads or changes the environments' state (stats,
 * snapshot, reset, set_targets, config_*).  Values are identical to the undeferred step's; observations, dones and
 * auto-resets are never deferred.  The window is capped by eight times the shortest episode of the data set (an environment
 * may finish up to eight times within it; every finish gets its own batch entry); steps with a host output buffer, with
 * the local search or without auto-reset run undeferred.  steps = 1 (default): every step folds its own terminals.  No counterpart in the reference (its workers fold synchronously inside step()). */
/* Enqueue the environments' work on the CALLER's stream from now on (a cudaStream_t, e.g. torch's current stream, so that a
 * policy network and the step are ordered without host synchronisation; NULL = the legacy default stream); use_own != 0 returns to the environments' own
 * non-blocking stream, on which device outputs are only ordered against the library's own calls. */
int rnafold_env_set_stream(int32_t handle, void *stream, int32_t use_own);
int rnafold_env_config_delay(int32_t handle, int32_t steps);
int rnafold_env_flush(int32_t handle);
int rnafold_env_stats(int32_t handle, uint64_t istat[4], double fstat[2], int32_t clear);
int rnafold_env_snapshot(int32_t handle, uint8_t *seq, int32_t *idx, int32_t *cur, int32_t *ep, uint8_t *last_seq,
                         uint8_t *last_db, int32_t *last_hd, int32_t *last_mfe, int32_t *last_t, int32_t *lmax_out);
/* Host-only: base-pair partners (opener -> closer, -1 elsewhere) and motif codes 0..5 of one target. */
int rnafold_parse_target(const uint8_t *dotbracket, int32_t n, int16_t *pair_out, uint8_t *motif_out);

/* ---- Batched local search: Solution.permute (rlif/rna/solution.py:157-219) as reached from
 * Solution.evaluate(permute=True) (rlif/rna/solution.py:133-141).  Every sequence is folded; those with
 * 0 < HD <= threshold run up to `budget` rounds of mutate-around-the-mismatches / refold / keep-if-HD-improves
 * (config keys permutation_budget, permutation_radius, permutation_threshold, mutation_probability,
 * allow_gu_permutations of rlif/config/RnaDesign.yml:34-38, Testing.yml:4-8); all candidates of a round are
 * folded by one batched fold.  `seq` is updated in place with the best sequence found.  Outputs (any may be NULL):
 * structure / MFE / HD / reward of the best sequence, rounds used per sequence, total folds performed.
 * seq/target/outputs may be host or device pointers; `off` must be a host array.  Random draws are the
 * counter-based function rnafold_permute_draw(seed, sequence number, round, position, kind). */
int rnafold_permute_batch(uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target, int32_t budget,
                          int32_t radius, int32_t threshold, double mutation_probability, int32_t allow_gu,
                          uint64_t seed, double reward_exp, uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out,
                          double *reward_out, int32_t *rounds_out, int32_t *folds_out);
/* Order in which a round visits its candidate positions -- the reference iterates a Python set (solution.py:183,189),
 * and the order decides which write wins when two mutations touch the same nucleotide.  0 (default): CPython >= 3.7 set
 * order (what recorded tests/golden/permute_cases.json); 1: CPython 3.6 (the interpreter rlif.yml:5 pins: the table
 * grows at 2/3 full instead of 3/5); 2: ascending positions. */
int rnafold_permute_set_order(int32_t mode);
/* Host-only: that order for one sequence's mismatch flags (n bytes) and radius -> positions in out[<= n]; returns their number. */
int rnafold_permute_candidate_order(const uint8_t *mismatch, int32_t n, int32_t radius, int32_t mode, int32_t *out);
/* Host-only: the draw behind random.random() (kind 0: (v >> 11) * 2^-53), random.randint(0,3) (kind 1: (v >> 33) % 4)
 * and random.randint(0,5) (kind 2: (v >> 33) % 6) of round `step` at position `pos` of sequence `index`. */
uint64_t rnafold_permute_draw(uint64_t seed, int32_t index, int32_t step, int32_t pos, int32_t kind);
/* RnaDesign.permute for the batched environment (RLIFenv.py:13,69-73; set through env.set_attr('permute', ...),
 * rlif/learning/tester.py:65): terminal evaluations run rnafold_permute_batch instead of a plain fold. */
int rnafold_env_config_permute(int32_t handle, int32_t enabled, int32_t budget, int32_t radius, int32_t threshold,
                               double mutation_probability, int32_t allow_gu);

/* ---- Ensemble statistics: Solution.compute_statistics (rlif/rna/solution.py:221-238; printed by rlif/CLI.py:42-49, logged
 * as the `md` / `ed` columns by rlif/utils/comparison.py:225-233) for a batch: McCaskill partition function and base-pair
 * probabilities of the same loop model (dangles=2), fp64 on the GPU.  `structure` / `e_dcal` are the fold and energy of every
 * sequence (what rnafold_fold_batch returns: fc.mfe()); `target` (may be NULL) the dot-bracket targets for the mountain
 * distance.  stats_out[B][RNAFOLD_ENSEMBLE_STATS] = { ensemble free energy kcal/mol (fc.pf()[1]), ensemble defect of
 * `structure` (fc.ensemble_defect(fold)), its probability (fc.pr_structure(fold)), mean base-pair distance
 * (fc.mean_bp_distance()), distance of the centroid to the ensemble (fc.centroid()[1]), mountain distance
 * RNA.dist_mountain(target, fold), expected accuracy of the MEA structure (fc.MEA()[1]), energy of the centroid
 * (fc.eval_structure(centroid), kcal/mol), energy of the MEA structure (solution.py:232-234) }.
 * Optional: centroid_out (CSR bytes like seq: pairs with p > 0.5, fc.centroid()[0]); pos_out[total][3] = P(paired with a
 * base downstream), P(paired upstream), positional entropy in bits (fc.positional_entropy(); the first two give the pair
 * propensity string of fc.pf()[0]); bpp_out (single-sequence calls only: n x n row-major, upper triangle); mea_out (CSR
 * bytes: the maximum-expected-accuracy structure fc.MEA()[0] for pair weight mea_gamma > 0; the reference uses the default 1).
 * off must be a host array; the other buffers may be host or device pointers.  Synchronous. */
#define RNAFOLD_ENSEMBLE_MAX_LEN 1000
#define RNAFOLD_ENSEMBLE_STATS 9
int rnafold_ensemble_batch(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *structure,
                           const int32_t *e_dcal, const uint8_t *target, double *stats_out, uint8_t *centroid_out,
                           double *pos_out, double *bpp_out, double mea_gamma, uint8_t *mea_out);

/* Free energy of GIVEN structures (fc.eval_structure(structure) of ViennaRNA 2.4.14 on the reference's model, dangles=2;
 * rlif/rna/solution.py:232,234) for a batch: one CTA per sequence, the loops of a structure spread over its threads.
 * structure: dot-bracket bytes, CSR like seq.  e_dcal_out[B]: energies in dcal/mol (x / 100 = kcal/mol), loops of any size
 * (interior loops above 30 unpaired bases take the lxc * log(u / 30) extrapolation).  err_out[B] (may be NULL): 0, 1 = a
 * bracket pair that cannot pair or a hairpin of fewer than 3 bases, 3 = unbalanced brackets (energy 0 then).
 * off must be a host array; the other buffers may be host or device pointers.  Synchronous. */
int rnafold_eval_structure_batch(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *structure,
                                 int32_t *e_dcal_out, int32_t *err_out);

/* Measurement helper (bench.py): on-box peaks of the two on-chip resources that bound the fold DP,
 * whole device: conflict-free 4-byte shared-memory loads per second and integer add+min
 * (min-plus candidate) updates per second. */
int rnafold_measure_onchip_peaks(double *lds_per_s, double *minplus_per_s);

/* Host-only planning helper: dynamic shared memory (bytes) the wavefront kernel needs for sequences
 * up to nmax with int16 (elem_bytes 2) or int32 (4) tables. */
int rnafold_wave_smem_bytes(int32_t nmax, int32_t elem_bytes);

/* SM count and opt-in shared memory per CTA of the bound device. */
int rnafold_device_info(int32_t *num_sms, int32_t *smem_optin_bytes);

#ifdef __cplusplus
}
#endif
#endif
