/*
 * rnafold_pf.c -- CPU restatement of the ensemble statistics behind Solution.compute_statistics
 * (rlif/rna/solution.py:221-238): fc.pf(), fc.ensemble_defect(fold), fc.pr_structure(fold), fc.mean_bp_distance(),
 * fc.centroid() and RNA.dist_mountain(target, fold) of ViennaRNA 2.4.14.  TEST INFRASTRUCTURE ONLY (see rnafold_oracle.h).
 *
 * The arithmetic is third-party (ViennaRNA part_func.c / equilibrium_probs.c / params/basic.c get_scaled_exp_params,
 * loops/ exp_E_xxx; absent from /root/reference).  Restated here from the published McCaskill algorithm on
 * the loop model the MFE oracle uses (dangles=2: every stem in a multiloop / exterior loop sees both neighbours; MAXLOOP=30;
 * TURN=3), with Boltzmann factors exp(-E * 10 / kT) of the integer dcal/mol parameters, kT = 1.98717 * 310.15 cal/mol.
 * Where ViennaRNA differs from the MFE model it is followed: long hairpins use the un-truncated lxc * log(u / 30) term, and the
 * dangle / mismatchM / mismatchExt terms go through its SMOOTH() function instead of the min(0, .) clamp (`smooth`).
 * The OUTSIDE recursion is written in a different (wavefront-friendly) but algebraically identical form: multiloop
 * contributions are collected through two helper tables X(i,l), Y(i,l) instead of ViennaRNA's rolling prm_l arrays.
 *
 * PARITY PIN: `ed` / `md` columns of the 1,726 logged rows (tests/test_oracle_golden.py); brute-force enumeration for n <= 14.
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <ctype.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "rnafold_oracle.h"

#define TURN ORC_TURN
#define MAXLOOP ORC_MAXLOOP
#define MIN2(a, b) ((a) < (b) ? (a) : (b))
#define MAX2(a, b) ((a) > (b) ? (a) : (b))

static const int RT[8] = {0, 2, 1, 4, 3, 6, 5, 7};
static const int PR[5][5] = {{0, 0, 0, 0, 0}, {0, 0, 0, 0, 5}, {0, 0, 0, 1, 0}, {0, 0, 2, 0, 3}, {0, 6, 0, 4, 0}};

static int enc(char c) {
  switch (toupper((unsigned char)c)) {
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 3;
    case 'U': case 'T': return 4;
    default: return 0;
  }
}

typedef struct {
  double kT;
  double stack[8][8], hairpin[31], bulge[31], interior[31];
  double mmH[8][5][5], mmI[8][5][5], mm1nI[8][5][5], mm23I[8][5][5], mmM[8][5][5], mmExt[8][5][5];
  double d5[8][5], d3[8][5];
  double int11[8][8][5][5], int21[8][8][5][5][5], int22[8][8][5][5][5][5];
  double MLbase, MLclosing, MLintern, TermAU, ninio[MAXLOOP + 1], lxc;
  double tetra[ORC_MAXSPECIAL], tri[ORC_MAXSPECIAL], hexa[ORC_MAXSPECIAL];
} expar;

/* ViennaRNA params/basic.c: SMOOTH(X), SCALE = 10 -- a differentiable stand-in for max(0, X) on bonus terms */
static double smooth_fn(double x) {
  const double s = x / 10.;
  if (s < -1.2283697) return 0.;
  if (s > 0.8660254) return x;
  const double t = sin(s - 0.34242663) + 1.;
  return 10. * 0.38490018 * t * t;
}

static void make_expar(const orc_params *P, int smooth, expar *E) {
  const double kT = (37. + 273.15) * 1.98717;
  E->kT = kT;
#define BF(x) exp(-(double)(x) * 10. / kT)
  for (int a = 0; a < 8; a++)
    for (int b = 0; b < 8; b++) E->stack[a][b] = BF(P->stack[a][b]);
  for (int k = 0; k <= 30; k++) {
    E->hairpin[k] = BF(P->hairpin[k]);
    E->bulge[k] = BF(P->bulge[k]);
    E->interior[k] = BF(P->interior[k]);
  }
  for (int t = 0; t < 8; t++)
    for (int a = 0; a < 5; a++) {
      for (int b = 0; b < 5; b++) {
        E->mmH[t][a][b] = BF(P->mismatchH[t][a][b]);
        E->mmI[t][a][b] = BF(P->mismatchI[t][a][b]);
        E->mm1nI[t][a][b] = BF(P->mismatch1nI[t][a][b]);
        E->mm23I[t][a][b] = BF(P->mismatch23I[t][a][b]);
        if (smooth) {
          E->mmM[t][a][b] = exp(smooth_fn(-(double)P->raw_mismatchM[t][a][b]) * 10. / kT);
          E->mmExt[t][a][b] = exp(smooth_fn(-(double)P->raw_mismatchExt[t][a][b]) * 10. / kT);
        } else {
          E->mmM[t][a][b] = BF(P->mismatchM[t][a][b]);
          E->mmExt[t][a][b] = BF(P->mismatchExt[t][a][b]);
        }
      }
      if (smooth) {
        E->d5[t][a] = exp(smooth_fn(-(double)P->raw_dangle5[t][a]) * 10. / kT);
        E->d3[t][a] = exp(smooth_fn(-(double)P->raw_dangle3[t][a]) * 10. / kT);
      } else {
        E->d5[t][a] = BF(P->dangle5[t][a]);
        E->d3[t][a] = BF(P->dangle3[t][a]);
      }
    }
  for (int a = 0; a < 8; a++)
    for (int b = 0; b < 8; b++)
      for (int c = 0; c < 5; c++)
        for (int d = 0; d < 5; d++) {
          E->int11[a][b][c][d] = BF(P->int11[a][b][c][d]);
          for (int e = 0; e < 5; e++) {
            E->int21[a][b][c][d][e] = BF(P->int21[a][b][c][d][e]);
            for (int f = 0; f < 5; f++) E->int22[a][b][c][d][e][f] = BF(P->int22[a][b][c][d][e][f]);
          }
        }
  E->MLbase = BF(P->MLbase);
  E->MLclosing = BF(P->MLclosing);
  E->MLintern = BF(P->MLintern);
  E->TermAU = BF(P->TerminalAU);
  for (int k = 0; k <= MAXLOOP; k++) E->ninio[k] = BF(MIN2(P->max_ninio, k * P->ninio));
  E->lxc = P->lxc;
  for (int k = 0; k < P->n_tetra; k++) E->tetra[k] = BF(P->tetra_E[k]);
  for (int k = 0; k < P->n_tri; k++) E->tri[k] = BF(P->tri_E[k]);
  for (int k = 0; k < P->n_hexa; k++) E->hexa[k] = BF(P->hexa_E[k]);
#undef BF
}

/* exp_E_Hairpin (loops/hairpin.h): u unpaired bases, str = sequence from the closing pair's 5' base */
static double x_hairpin(const orc_params *P, const expar *E, int u, int type, int si1, int sj1, const char *str) {
  double q = (u <= 30) ? E->hairpin[u] : E->hairpin[30] * exp(-(E->lxc * log(u / 30.)) * 10. / E->kT);
  if (u < 3) return q;
  if (u == 4) {
    for (int k = 0; k < P->n_tetra; k++)
      if (!strncmp(P->tetra[k], str, 6)) return E->tetra[k];
  } else if (u == 6) {
    for (int k = 0; k < P->n_hexa; k++)
      if (!strncmp(P->hexa[k], str, 8)) return E->hexa[k];
  } else if (u == 3) {
    for (int k = 0; k < P->n_tri; k++)
      if (!strncmp(P->tri[k], str, 5)) return E->tri[k];
    return (type > 2) ? q * E->TermAU : q;
  }
  return q * E->mmH[type][si1][sj1];
}

/* exp_E_IntLoop (loops/internal.h) */
static inline double x_intloop(const expar *E, int u1, int u2, int type, int type2, int si1, int sj1, int sp1, int sq1) {
  const int ul = u1 > u2 ? u1 : u2, us = u1 > u2 ? u2 : u1;
  if (ul == 0) return E->stack[type][type2];
  if (us == 0) {
    double z = E->bulge[ul];
    if (ul == 1) return z * E->stack[type][type2];
    if (type > 2) z *= E->TermAU;
    if (type2 > 2) z *= E->TermAU;
    return z;
  }
  if (us == 1) {
    if (ul == 1) return E->int11[type][type2][si1][sj1];
    if (ul == 2) return (u1 == 1) ? E->int21[type][type2][si1][sq1][sj1] : E->int21[type2][type][sq1][si1][sp1];
    return E->interior[ul + 1] * E->mm1nI[type][si1][sj1] * E->mm1nI[type2][sq1][sp1] * E->ninio[ul - us];
  }
  if (us == 2) {
    if (ul == 2) return E->int22[type][type2][si1][sp1][sq1][sj1];
    if (ul == 3) return E->interior[5] * E->mm23I[type][si1][sj1] * E->mm23I[type2][sq1][sp1] * E->ninio[1];
  }
  return E->interior[ul + us] * E->mmI[type][si1][sj1] * E->mmI[type2][sq1][sp1] * E->ninio[ul - us];
}

static inline double x_mlstem(const expar *E, int type, int s5, int s3) {
  double e = 1.;
  if (s5 >= 0 && s3 >= 0) e = E->mmM[type][s5][s3];
  else if (s5 >= 0) e = E->d5[type][s5];
  else if (s3 >= 0) e = E->d3[type][s3];
  if (type > 2) e *= E->TermAU;
  return e * E->MLintern;
}

static inline double x_extloop(const expar *E, int type, int s5, int s3) {
  double e = 1.;
  if (s5 >= 0 && s3 >= 0) e = E->mmExt[type][s5][s3];
  else if (s5 >= 0) e = E->d5[type][s5];
  else if (s3 >= 0) e = E->d3[type][s3];
  if (type > 2) e *= E->TermAU;
  return e;
}

/* ------------------------------------------------------------------ inside / outside */

int orc_pf(const orc_params *P, const char *seq_in, int n, const char *structure, int e_dcal, int smooth, double *stats,
           char *centroid, double *bpp_out) {
  for (int k = 0; k < ORC_PF_NSTATS; k++) stats[k] = 0.;
  if (centroid) {
    for (int k = 0; k < n; k++) centroid[k] = '.';
    centroid[n > 0 ? n : 0] = 0;
  }
  if (bpp_out) memset(bpp_out, 0, sizeof(double) * (size_t)n * (size_t)n);
  if (n <= 0) return 0;
  expar *E = (expar *)malloc(sizeof(expar));
  make_expar(P, smooth, E);
  const double kT = E->kT;
  const int N = n + 2;
  char *seq = (char *)malloc((size_t)n + 2);
  short *S = (short *)malloc(sizeof(short) * (size_t)(n + 2));
  for (int k = 0; k < n; k++) {
    seq[k] = (char)toupper((unsigned char)seq_in[k]);
    if (seq[k] == 'T') seq[k] = 'U';
    S[k + 1] = (short)enc(seq_in[k]);
  }
  seq[n] = 0;
  S[0] = S[n];
  S[n + 1] = S[1];
  /* vrna_exp_params_rescale(&mfe): pf_scale = exp(-(sfact * mfe_per_nt) / kT), sfact = 1.07, never below 1 */
  double pf_scale = exp(-(1.07 * ((double)e_dcal * 10. / n)) / kT);
  if (pf_scale < 1.) pf_scale = 1.;
  double *scale = (double *)malloc(sizeof(double) * (size_t)(n + 3));
  double *bpow = (double *)malloc(sizeof(double) * (size_t)(n + 3));   /* (expMLbase * scale[1])^k */
  scale[0] = bpow[0] = 1.;
  for (int k = 1; k <= n + 2; k++) {
    scale[k] = scale[k - 1] / pf_scale;
    bpow[k] = bpow[k - 1] * E->MLbase / pf_scale;
  }
#define IX(i, j) ((size_t)(i) * (size_t)N + (size_t)(j))
  const size_t cells = (size_t)N * (size_t)N;
  double *qb = (double *)calloc(cells, sizeof(double)), *qm = (double *)calloc(cells, sizeof(double));
  double *qm1 = (double *)calloc(cells, sizeof(double)), *q = (double *)calloc(cells, sizeof(double));
  double *q1 = (double *)calloc(cells, sizeof(double));
  for (int i = 1; i <= n + 1; i++) q[IX(i, i - 1)] = 1.;
  for (int i = 1; i <= n; i++)
    for (int j = i; j <= MIN2(n, i + TURN); j++) q[IX(i, j)] = scale[j - i + 1];
  const double clos2 = E->MLclosing * scale[2];
  for (int d = TURN + 1; d < n; d++)
    for (int i = 1; i + d <= n; i++) {
      const int j = i + d;
      const int type = PR[S[i]][S[j]];
      double b = 0.;
      if (type) {
        b = x_hairpin(P, E, d - 1, type, S[i + 1], S[j - 1], seq + i - 1) * scale[d + 1];
        for (int k = i + 1; k <= MIN2(i + MAXLOOP + 1, j - TURN - 2); k++) {
          const int u1 = k - i - 1;
          for (int l = MAX2(k + TURN + 1, j - 1 - MAXLOOP + u1); l < j; l++) {
            const int t2 = PR[S[k]][S[l]];
            if (!t2) continue;
            b += qb[IX(k, l)] * x_intloop(E, u1, j - l - 1, type, RT[t2], S[i + 1], S[j - 1], S[k - 1], S[l + 1]) * scale[u1 + j - l + 1];
          }
        }
        double t = 0.;
        for (int k = i + 2; k <= j - 1; k++) t += qm[IX(i + 1, k - 1)] * qm1[IX(k, j - 1)];
        b += t * clos2 * x_mlstem(E, RT[type], S[j - 1], S[i + 1]);
      }
      qb[IX(i, j)] = b;
      const int s5 = (i > 1) ? S[i - 1] : -1, s3 = (j < n) ? S[j + 1] : -1;
      qm1[IX(i, j)] = qm1[IX(i, j - 1)] * bpow[1] + (type ? b * x_mlstem(E, type, s5, s3) : 0.);
      double m = 0.;
      for (int k = i; k <= j; k++) m += (qm[IX(i, k - 1)] + bpow[k - i]) * qm1[IX(k, j)];
      qm[IX(i, j)] = m;
      q1[IX(i, j)] = q1[IX(i, j - 1)] * scale[1] + (type ? b * x_extloop(E, type, s5, s3) : 0.);
      double t = scale[d + 1] + q1[IX(i, j)];
      for (int k = i; k < j; k++) t += q[IX(i, k)] * q1[IX(k + 1, j)];
      q[IX(i, j)] = t;
    }
  const double Z = q[IX(1, n)];
  stats[0] = (-log(Z) - n * log(pf_scale)) * kT / 1000.;

  /* outside: po(k,l) = P(k,l) / qb(k,l); W(i,j) = po(i,j) * closing weight of a multiloop closed by (i,j) */
  double *po = (double *)calloc(cells, sizeof(double)), *W = (double *)calloc(cells, sizeof(double));
  double *X = (double *)calloc(cells, sizeof(double)), *Y = (double *)calloc(cells, sizeof(double));
  for (int d = n - 1; d > TURN; d--) {
    for (int k = 1; k + d <= n; k++) {
      const int l = k + d;
      const int t2 = PR[S[k]][S[l]];
      if (!t2 || qb[IX(k, l)] == 0.) continue;
      double p = q[IX(1, k - 1)] * q[IX(l + 1, n)] / Z * x_extloop(E, t2, k > 1 ? S[k - 1] : -1, l < n ? S[l + 1] : -1);
      for (int i = MAX2(1, k - MAXLOOP - 1); i < k; i++) {
        const int u1 = k - i - 1;
        for (int j = l + 1; j <= MIN2(n, l + 1 + MAXLOOP - u1); j++) {
          const int type = PR[S[i]][S[j]];
          if (!type || po[IX(i, j)] == 0.) continue;
          p += po[IX(i, j)] * x_intloop(E, u1, j - l - 1, type, RT[t2], S[i + 1], S[j - 1], S[k - 1], S[l + 1]) * scale[u1 + j - l + 1];
        }
      }
      double m = 0.;
      for (int i = 1; i < k; i++)
        m += (bpow[k - i - 1] + qm[IX(i + 1, k - 1)]) * X[IX(i, l)] - bpow[k - i - 1] * Y[IX(i, l)];
      if (l < n) p += m * x_mlstem(E, t2, S[k - 1], S[l + 1]);
      po[IX(k, l)] = p;
      W[IX(k, l)] = p * clos2 * x_mlstem(E, RT[t2], S[l - 1], S[k + 1]);
    }
    /* X(i,l), Y(i,l) for l - i = d: sums over the closing pairs (i,j), j > l -- all of span > d, final by now */
    for (int i = 1; i + d <= n; i++) {
      const int l = i + d;
      double x = 0., y = 0.;
      for (int j = l + 1; j <= n; j++) {
        const double w = W[IX(i, j)];
        if (w == 0.) continue;
        x += w * (bpow[j - l - 1] + qm[IX(l + 1, j - 1)]);
        y += w * bpow[j - l - 1];
      }
      X[IX(i, l)] = x;
      Y[IX(i, l)] = y;
    }
  }
  /* base-pair probabilities and the statistics built on them */
  double *pp = (double *)calloc((size_t)n + 2, sizeof(double));   /* probability of being paired */
  int *pt = (int *)calloc((size_t)n + 2, sizeof(int));
  {
    int *stk = (int *)malloc(sizeof(int) * (size_t)(n + 1)), top = 0;
    for (int k = 1; k <= n; k++) {
      const char c = structure ? structure[k - 1] : '.';
      if (c == '(') stk[top++] = k;
      else if (c == ')' && top > 0) {
        const int a = stk[--top];
        pt[a] = k;
        pt[k] = a;
      }
    }
    free(stk);
  }
  double mbd = 0., cdist = 0., ed = 0.;
  for (int i = 1; i <= n; i++)
    for (int j = i + TURN + 1; j <= n; j++) {
      const double p = po[IX(i, j)] * qb[IX(i, j)];
      if (p == 0.) continue;
      pp[i] += p;
      pp[j] += p;
      mbd += 2. * p * (1. - p);
      if (p > 0.5) {
        if (centroid) {
          centroid[i - 1] = '(';
          centroid[j - 1] = ')';
        }
        cdist += 1. - p;
      } else cdist += p;
      if (bpp_out) bpp_out[(size_t)(i - 1) * n + (j - 1)] = p;
    }
  /* vrna_ensemble_defect of ViennaRNA 2.4.14 leaves the LAST position out of the sum and still divides by n: pinned by the
   * logged `ed` column (every one of the 1,697 rows whose HD reproduces matches to 5e-4 with i < n, 70 % with i <= n) */
  for (int i = 1; i < n; i++) ed += pt[i] ? 1. - po[IX(MIN2(i, pt[i]), MAX2(i, pt[i]))] * qb[IX(MIN2(i, pt[i]), MAX2(i, pt[i]))] : pp[i];
  stats[1] = ed / n;
  stats[2] = exp((stats[0] * 1000. - (double)e_dcal * 10.) / kT);
  stats[3] = mbd;
  stats[4] = cdist;
#undef IX
  free(pp); free(pt); free(po); free(W); free(X); free(Y);
  free(qb); free(qm); free(qm1); free(q); free(q1); free(scale); free(bpow); free(S); free(seq); free(E);
  return 0;
}

/* RNA.dist_mountain(a, b) of ViennaRNA 2.4.14, pinned by the logged `md` column (all 597 rows with HD > 0 whose HD reproduces
 * match to 5e-4): a pair (i, j) raises the running height by 1 / (j - i) at i and lowers it at j; the height is RECORDED at
 * paired positions only (unpaired positions hold 0), and the distance is the L1 norm of the difference of the two records.
 * weighted == 0: every pair weighs 1 (not what the reference's library does; kept for comparison). */
double orc_dist_mountain(const char *a, const char *b, int n, int weighted) {
  double *f[2];
  const char *s[2] = {a, b};
  int *stk = (int *)malloc(sizeof(int) * (size_t)(n + 1));
  for (int w = 0; w < 2; w++) {
    int *pt = (int *)calloc((size_t)n + 2, sizeof(int)), top = 0;
    for (int k = 1; k <= n; k++) {
      if (s[w][k - 1] == '(') stk[top++] = k;
      else if (s[w][k - 1] == ')' && top > 0) {
        const int o = stk[--top];
        pt[o] = k;
        pt[k] = o;
      }
    }
    f[w] = (double *)calloc((size_t)n + 2, sizeof(double));
    double h = 0.;
    for (int k = 1; k <= n; k++) {
      if (!pt[k]) continue;
      h += (pt[k] > k ? 1. : -1.) * (weighted ? 1. / (double)abs(pt[k] - k) : 1.);
      f[w][k] = h;
    }
    free(pt);
  }
  double dist = 0.;
  for (int k = 1; k <= n; k++) dist += fabs(f[0][k] - f[1][k]);
  free(f[0]); free(f[1]); free(stk);
  return dist;
}

int orc_pf_batch(const orc_params *P, const char *seqs, const int *off, int B, const char *structures, const int *e_dcal,
                 int smooth, double *stats_out, int nthreads) {
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
  int rc = 0;
#pragma omp parallel for schedule(dynamic, 1)
  for (int b = 0; b < B; b++) {
    const int n = off[b + 1] - off[b];
    char *st = (char *)malloc((size_t)n + 1);
    memcpy(st, structures + off[b], (size_t)n);
    st[n] = 0;
    if (orc_pf(P, seqs + off[b], n, st, e_dcal[b], smooth, stats_out + (size_t)b * ORC_PF_NSTATS, NULL, NULL)) {
#pragma omp critical
      rc = -1;
    }
    free(st);
  }
  return rc;
}

/* ------------------------------------------------------------------ maximum expected accuracy structure
 * fc.MEA() of ViennaRNA 2.4.14 (rlif/rna/solution.py:233; gamma = 1): restated from the published algorithm of MEA.c --
 *   pu[i] = 1 - sum of the listed pair probabilities of i, the list holding pairs with p >= 1e-4 / (1 + gamma);
 *   a pair is a candidate when 2 gamma p > pu[i] + pu[j];
 *   M(i,j) = max( M(i,j-1) + pu[j],  max_k M(i,k-1) + A(k,j) ),  A(k,j) = 2 gamma p(k,j) + M(k+1,j-1),
 *   rows i = n .. 1 kept two at a time, A(k,j) remembered in a per-j list only where the pair strictly improved M(k,j);
 *   backtracking re-derives a row per enclosed interval and accepts equality within DBL_EPSILON * M.
 * PARITY UNPINNED: the reference logs no MEA column and the library is absent, so this is checked against the definition
 * only (enumeration over all structures, tests/test_oracle_golden.py); the value does not depend on tie-breaking, the
 * structure could where two optima tie exactly.
 * bpp: n x n row-major upper triangle (orc_pf's bpp_out).  structure: n + 1 chars. */
typedef struct { int i; double A; } mea_item;
typedef struct { mea_item *v; int n, cap; } mea_list;

static void mea_push(mea_list *L, int i, double A) {
  if (L->n == L->cap) {
    L->cap = L->cap ? 2 * L->cap : 4;
    L->v = (mea_item *)realloc(L->v, sizeof(mea_item) * (size_t)L->cap);
  }
  L->v[L->n].i = i;
  L->v[L->n].A = A;
  L->n++;
}

static void mea_row(const mea_list *C, const double *pu, double *Mi, int i, int j) {
  /* Mi[i-1 .. j] of the interval that starts at i, from the remembered candidates */
  Mi[i - 1] = 0.;
  Mi[i] = pu[i];
  for (int k = i + 1; k <= j; k++) {
    Mi[k] = Mi[k - 1] + pu[k];
    for (int t = 0; t < C[k].n && C[k].v[t].i >= i; t++) {
      const double EA = C[k].v[t].A + Mi[C[k].v[t].i - 1];
      if (EA > Mi[k]) Mi[k] = EA;
    }
  }
}

static int mea_backtrack(const mea_list *C, const double *pu, double *Mi, char *structure, int i, int j, int pair) {
  if (pair) {
    structure[i - 1] = '(';
    structure[j - 1] = ')';
    i++;
    j--;
    if (j < i) return 0;
    mea_row(C, pu, Mi, i, j);
  }
  const double prec = 2.220446049250313e-16 * Mi[j];
  while (j > i && Mi[j] <= Mi[j - 1] + pu[j] + prec) j--;
  if (j <= i) return 0;
  for (int t = 0; t < C[j].n && C[j].v[t].i >= i; t++) {
    const int k = C[j].v[t].i;
    if (Mi[j] <= C[j].v[t].A + Mi[k - 1] + prec) {
      if (k > i + 3) {
        /* the left part keeps the row of i: the recursion below only reads entries <= k-1 */
        if (mea_backtrack(C, pu, Mi, structure, i, k - 1, 0)) return -1;
      }
      return mea_backtrack(C, pu, Mi, structure, k, j, 1);
    }
  }
  return -1;
}

int orc_mea(const double *bpp, int n, double gamma, char *structure, double *mea_out) {
  for (int k = 0; k < n; k++) structure[k] = '.';
  structure[n > 0 ? n : 0] = 0;
  *mea_out = 0.;
  if (n <= 0) return 0;
  const double cut = 1e-4 / (1. + gamma);
  double *pu = (double *)malloc(sizeof(double) * (size_t)(n + 2));
  double *Mi = (double *)calloc((size_t)n + 2, sizeof(double)), *Mi1 = (double *)calloc((size_t)n + 2, sizeof(double));
  mea_list *C = (mea_list *)calloc((size_t)n + 2, sizeof(mea_list));
  for (int i = 0; i <= n + 1; i++) pu[i] = 1.;
  for (int i = 1; i <= n; i++)
    for (int j = i + 1; j <= n; j++) {
      const double p = bpp[(size_t)(i - 1) * n + (j - 1)];
      if (p >= cut) { pu[i] -= p; pu[j] -= p; }
    }
  for (int i = n; i > 0; i--) {
    Mi[i] = pu[i];
    for (int j = i + 1; j <= n; j++) {
      Mi[j] = Mi[j - 1] + pu[j];
      for (int t = 0; t < C[j].n; t++) {
        const double EA = C[j].v[t].A + Mi[C[j].v[t].i - 1];
        if (EA > Mi[j]) Mi[j] = EA;
      }
      const double p = bpp[(size_t)(i - 1) * n + (j - 1)];
      if (p >= cut && p * 2. * gamma > pu[i] + pu[j]) {
        const double EA = 2. * gamma * p + Mi1[j - 1];
        if (Mi[j] < EA) {
          Mi[j] = EA;
          mea_push(&C[j], i, EA);
        }
      }
    }
    double *t = Mi1; Mi1 = Mi; Mi = t;
  }
  *mea_out = Mi1[n];
  /* the row of i = 1 is in Mi1 */
  int rc = mea_backtrack(C, pu, Mi1, structure, 1, n, 0);
  for (int k = 0; k <= n + 1; k++) free(C[k].v);
  free(C); free(Mi); free(Mi1); free(pu);
  return rc;
}
