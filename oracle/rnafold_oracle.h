/*
 * rnafold_oracle.h -- CPU restatement of the reference's fold call.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (rnagent_b200/) never links, imports or calls it.
 *
 * What it restates: `RNA.fold(seq) -> (dot-bracket, mfe)` of ViennaRNA 2.4.14, the third-party
 * C library the reference binds at rlif/settings.py:63-64 (pinned at rlif.yml:6, README.md:32) and
 * calls from rlif/rna/solution.py:134,140,204.  ViennaRNA's source is NOT under /root/reference
 * and is not installed in this image, so the algorithm is restated from the published
 * ViennaRNA 2.4.x design (Zuker/Stiegler recursions, Turner nearest-neighbour loop energies,
 * dangles=2, MAXLOOP=30, TURN=3, integer dcal/mol; see SURVEY.md section 8a').
 *
 * PARITY PIN: the 1,726 logged (sequence, HD-to-target, MFE) rows under rlif/logs/ (extracted
 * by tools/make_golden.py into tests/golden/vienna_logged_rows.jsonl).  Beyond those rows parity
 * with ViennaRNA is unpinned (no live RNA.fold on this box).
 */
#ifndef RNAFOLD_ORACLE_H
#define RNAFOLD_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_INF 10000000
#define ORC_TURN 3
#define ORC_MAXLOOP 30
#define ORC_NBPAIRS 7
#define ORC_MAXSPECIAL 200

typedef struct orc_params {
  int stack[8][8];
  int hairpin[31], bulge[31], interior[31];
  int mismatchH[8][5][5], mismatchI[8][5][5], mismatch1nI[8][5][5], mismatch23I[8][5][5];
  int mismatchM[8][5][5], mismatchExt[8][5][5];
  int dangle5[8][5], dangle3[8][5];
  int int11[8][8][5][5];
  int int21[8][8][5][5][5];
  int int22[8][8][5][5][5][5];
  int MLbase, MLclosing, MLintern;
  int ninio, max_ninio;
  int DuplexInit, TerminalAU;
  double lxc;
  int n_tetra, n_tri, n_hexa;
  char tetra[ORC_MAXSPECIAL][8];
  char tri[ORC_MAXSPECIAL][8];
  char hexa[ORC_MAXSPECIAL][12];
  int tetra_E[ORC_MAXSPECIAL], tri_E[ORC_MAXSPECIAL], hexa_E[ORC_MAXSPECIAL];
  /* as read from the file, before the <= 0 clamp of the MFE model (the partition function smooths them instead) */
  int raw_mismatchM[8][5][5], raw_mismatchExt[8][5][5], raw_dangle5[8][5], raw_dangle3[8][5];
} orc_params;

/* Allocate / free a parameter set (zero-initialised; must be loaded before use). */
orc_params *orc_params_new(void);
void orc_params_free(orc_params *P);

/* Parse an "RNAfold parameter file v2.0" (the format of rlif/utils/parameters/NAME.par, selected by
 * rlif/rna/vienna.py:7-19 -> RNA.read_parameter_file) and apply ViennaRNA's 37 C scaling rules
 * (dangles/mismatchM/mismatchExt clamped to <= 0 because dangles != 0).  Returns 0 on success. */
int orc_params_load(orc_params *P, const char *path);

/* Number of ints orc_params_export writes. */
int orc_params_export_size(void);
/* Flat int32 view of every table in a fixed order (for cross-checking the product's loader). */
int orc_params_export(const orc_params *P, int *out);

/* RNA.fold: fills db[0..n-1] (+NUL) and the MFE in dcal/mol.  Returns 0 on success. */
int orc_fold(const orc_params *P, const char *seq, int n, char *db, int *mfe_dcal);

/* Work counter: number of min-plus candidates actually evaluated (pairable (i,j), pairable
 * (p,q)) by the last orc_fold call on this thread -- W_eff of SURVEY.md section 8d. */
long long orc_last_fold_candidates(void);

/* Independent loop-decomposition energy of a structure (dangles=2 semantics). */
int orc_eval_structure(const orc_params *P, const char *seq, int n, const char *db, int *e_dcal);

/* Exhaustive minimum over all non-crossing structures (n <= 20), via orc_eval_structure. */
int orc_bruteforce_mfe(const orc_params *P, const char *seq, int n, int *mfe_dcal);

/* hamming_distance of rlif/rna/utils.py:67-75 (zip-truncating); mismatch flags optional. */
int orc_hamming(const char *a, int na, const char *b, int nb, unsigned char *mismatch);

/* Batched fold + Hamming on `nthreads` host threads (OpenMP).  seqs/targets are concatenated
 * bytes with CSR offsets off[B+1]; targets may be NULL.  db_out has the same layout as seqs. */
int orc_fold_batch(const orc_params *P, const char *seqs, const int *off, int B, const char *targets,
                   char *db_out, int *mfe_out, int *hd_out, int nthreads);

/* ---- ensemble statistics (rnafold_pf.c): Solution.compute_statistics, rlif/rna/solution.py:221-238 --------------
 * McCaskill partition function of the same loop model (dangles=2) + base-pair probabilities.
 * PARITY PIN: the `ed` (ensemble defect of the MFE structure) and `md` (mountain distance target <-> MFE structure)
 * columns of the same 1,726 logged rows (written by rlif/utils/comparison.py:225-233 with '{:.3f}').
 * stats[0] ensemble free energy (kcal/mol)         fc.pf()[1]
 * stats[1] ensemble defect of `structure`, / n      fc.ensemble_defect(fold)   (2.4.14 leaves position n out of the sum)
 * stats[2] probability of `structure`               fc.pr_structure(fold)   (needs its energy: e_dcal)
 * stats[3] mean base-pair distance                  fc.mean_bp_distance()
 * stats[4] distance of the centroid to the ensemble fc.centroid()[1]
 * centroid (n + 1 bytes, may be NULL): pairs with p > 0.5; bpp (may be NULL): n * n doubles, row-major, 0-based, upper triangle.
 * smooth != 0: ViennaRNA's SMOOTH() of the dangle / multiloop- and exterior-mismatch terms in the Boltzmann factors. */
#define ORC_PF_NSTATS 5
int orc_pf(const orc_params *P, const char *seq, int n, const char *structure, int e_dcal, int smooth, double *stats,
           char *centroid, double *bpp);
/* RNA.dist_mountain(a, b) as ViennaRNA 2.4.14 computes it (see rnafold_pf.c); weighted must be non-zero for parity. */
double orc_dist_mountain(const char *a, const char *b, int n, int weighted);
/* fc.MEA() (gamma = 1 in the reference): maximum expected accuracy structure from orc_pf's base-pair probability matrix
 * (n x n row-major upper triangle); structure holds n + 1 chars.  Returns 0, or -1 if backtracking failed. */
int orc_mea(const double *bpp, int n, double gamma, char *structure, double *mea_out);
/* Batched (OpenMP): stats_out[B][ORC_PF_NSTATS]; structures / e_dcal as returned by orc_fold_batch. */
int orc_pf_batch(const orc_params *P, const char *seqs, const int *off, int B, const char *structures, const int *e_dcal,
                 int smooth, double *stats_out, int nthreads);

#ifdef __cplusplus
}
#endif
#endif
