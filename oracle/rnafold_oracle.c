/*
 * rnafold_oracle.c -- CPU restatement of RNA.fold (ViennaRNA 2.4.14) for the reference's reward
 * hot path.  TEST INFRASTRUCTURE ONLY (see rnafold_oracle.h for who may load it and what pins it).
 *
 * Reference call sites restated here:
 *   rlif/settings.py:63-64        fold_fn = RNA.fold            -> orc_fold
 *   rlif/rna/vienna.py:7-19       RNA.read_parameter_file(path) -> orc_params_load
 *   rlif/rna/utils.py:67-75       hamming_distance              -> orc_hamming
 * The arithmetic itself lives in ViennaRNA 2.4.14 (third party, absent from /root/reference):
 *   params/io.c + params/basic.c  parameter file + 37 C scaling -> orc_params_load
 *   loops/hairpin.h  E_Hairpin, loops/internal.h E_IntLoop, loops/multibranch.h E_MLstem,
 *   loops/external.h E_ExtLoop                                 -> e_hairpin / e_intloop / ...
 *   mfe.c fill_arrays + backtrack, loops vrna_BT_xxx          -> orc_fold
 * Model details fixed to the reference's defaults (rlif/settings.py:35-39): 37 C, dangles=2,
 * noGU=0, no_closingGU=0, uniq_ML=0, noLP=0, special hairpins on, no constraints.
 */
#include "rnafold_oracle.h"

#include <ctype.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define INF ORC_INF
#define TURN ORC_TURN
#define MAXLOOP ORC_MAXLOOP
#define MIN2(a, b) ((a) < (b) ? (a) : (b))
#define MAX2(a, b) ((a) > (b) ? (a) : (b))

static const int RTYPE[8] = {0, 2, 1, 4, 3, 6, 5, 7};
/* pair type matrix, bases N,A,C,G,U = 0..4; CG=1 GC=2 GU=3 UG=4 AU=5 UA=6 */
static const int PAIR[5][5] = {
    {0, 0, 0, 0, 0}, {0, 0, 0, 0, 5}, {0, 0, 0, 1, 0}, {0, 0, 2, 0, 3}, {0, 6, 0, 4, 0}};

static int encode_base(char c) {
  switch (toupper((unsigned char)c)) {
    case 'A': return 1;
    case 'C': return 2;
    case 'G': return 3;
    case 'U':
    case 'T': return 4;
    default: return 0;
  }
}

/* ------------------------------------------------------------------ parameter file */

orc_params *orc_params_new(void) { return (orc_params *)calloc(1, sizeof(orc_params)); }
void orc_params_free(orc_params *P) { free(P); }

typedef struct {
  char **tok;
  int n, cap;
} toklist;

static void tl_push(toklist *t, char *s) {
  if (t->n == t->cap) {
    t->cap = t->cap ? 2 * t->cap : 256;
    t->tok = (char **)realloc(t->tok, sizeof(char *) * (size_t)t->cap);
  }
  t->tok[t->n++] = s;
}

static int tok_int(const char *s) {
  if (!strcmp(s, "INF")) return INF;
  if (!strcmp(s, "DEF")) return -50;
  if (!strcmp(s, "NST")) return 0;
  return atoi(s);
}

/* copy `count` ints of a section into dst via a strided writer */
static int need(const toklist *t, int count, const char *name) {
  if (t->n < count) {
    fprintf(stderr, "oracle: section %s has %d values, expected %d\n", name, t->n, count);
    return -1;
  }
  return 0;
}

static int load_mismatch(const toklist *t, int dst[8][5][5], const char *name) {
  if (need(t, 7 * 25, name)) return -1;
  int k = 0;
  for (int p = 1; p <= 7; p++)
    for (int a = 0; a < 5; a++)
      for (int b = 0; b < 5; b++) dst[p][a][b] = tok_int(t->tok[k++]);
  return 0;
}

static void fill_int22_nonstandard(int a22[8][8][5][5][5][5]) {
  /* entries touching base N (index 0): maximum over every A/C/G/U completion */
  for (int i = 1; i <= 6; i++)
    for (int j = 1; j <= 6; j++)
      for (int k = 0; k < 5; k++)
        for (int l = 0; l < 5; l++)
          for (int m = 0; m < 5; m++)
            for (int n = 0; n < 5; n++) {
              if (k && l && m && n) continue;
              int best = -INF;
              for (int kk = (k ? k : 1); kk <= (k ? k : 4); kk++)
                for (int ll = (l ? l : 1); ll <= (l ? l : 4); ll++)
                  for (int mm = (m ? m : 1); mm <= (m ? m : 4); mm++)
                    for (int nn = (n ? n : 1); nn <= (n ? n : 4); nn++)
                      best = MAX2(best, a22[i][j][kk][ll][mm][nn]);
              a22[i][j][k][l][m][n] = best;
            }
}

int orc_params_load(orc_params *P, const char *path) {
  FILE *f = fopen(path, "rb");
  if (!f) return -1;
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  char *buf = (char *)malloc((size_t)sz + 2);
  if (fread(buf, 1, (size_t)sz, f) != (size_t)sz) {
    fclose(f);
    free(buf);
    return -1;
  }
  fclose(f);
  buf[sz] = '\n';
  buf[sz + 1] = 0;
  /* blank out C comments */
  for (char *p = buf; (p = strstr(p, "/*"));) {
    char *e = strstr(p + 2, "*/");
    if (!e) e = buf + sz - 2;
    for (char *q = p; q < e + 2; q++)
      if (*q != '\n') *q = ' ';
    p = e + 2;
  }
  if (strncmp(buf, "## RNAfold parameter file v2.0", 30)) {
    free(buf);
    return -2;
  }
  memset(P, 0, sizeof(*P));
  P->lxc = 107.856;
  P->max_ninio = 300;

  int rc = 0;
  char *line = buf;
  char secname[64] = "";
  toklist t = {0, 0, 0};
  int done = 0;
  while (!done) {
    char *nl = line ? strchr(line, '\n') : NULL;
    int at_end = (nl == NULL);
    if (nl) *nl = 0;
    int is_header = (!at_end && line[0] == '#');
    if (is_header || at_end) {
      /* flush previous section */
      if (secname[0]) {
        const char *s = secname;
        int k = 0;
        if (!strcmp(s, "stack")) {
          if (need(&t, 49, s)) rc = -3;
          else
            for (int a = 1; a <= 7; a++)
              for (int b = 1; b <= 7; b++) P->stack[a][b] = tok_int(t.tok[k++]);
        } else if (!strcmp(s, "mismatch_hairpin")) {
          if (load_mismatch(&t, P->mismatchH, s)) rc = -3;
        } else if (!strcmp(s, "mismatch_interior")) {
          if (load_mismatch(&t, P->mismatchI, s)) rc = -3;
        } else if (!strcmp(s, "mismatch_interior_1n")) {
          if (load_mismatch(&t, P->mismatch1nI, s)) rc = -3;
        } else if (!strcmp(s, "mismatch_interior_23")) {
          if (load_mismatch(&t, P->mismatch23I, s)) rc = -3;
        } else if (!strcmp(s, "mismatch_multi")) {
          if (load_mismatch(&t, P->mismatchM, s)) rc = -3;
        } else if (!strcmp(s, "mismatch_exterior")) {
          if (load_mismatch(&t, P->mismatchExt, s)) rc = -3;
        } else if (!strcmp(s, "dangle5") || !strcmp(s, "dangle3")) {
          int(*d)[5] = !strcmp(s, "dangle5") ? P->dangle5 : P->dangle3;
          if (need(&t, 35, s)) rc = -3;
          else
            for (int a = 1; a <= 7; a++)
              for (int b = 0; b < 5; b++) d[a][b] = tok_int(t.tok[k++]);
        } else if (!strcmp(s, "int11")) {
          if (need(&t, 7 * 7 * 25, s)) rc = -3;
          else
            for (int a = 1; a <= 7; a++)
              for (int b = 1; b <= 7; b++)
                for (int x = 0; x < 5; x++)
                  for (int y = 0; y < 5; y++) P->int11[a][b][x][y] = tok_int(t.tok[k++]);
        } else if (!strcmp(s, "int21")) {
          if (need(&t, 7 * 7 * 125, s)) rc = -3;
          else
            for (int a = 1; a <= 7; a++)
              for (int b = 1; b <= 7; b++)
                for (int x = 0; x < 5; x++)
                  for (int y = 0; y < 5; y++)
                    for (int z = 0; z < 5; z++) P->int21[a][b][x][y][z] = tok_int(t.tok[k++]);
        } else if (!strcmp(s, "int22")) {
          if (need(&t, 6 * 6 * 256, s)) rc = -3;
          else {
            for (int a = 1; a <= 6; a++)
              for (int b = 1; b <= 6; b++)
                for (int x = 1; x < 5; x++)
                  for (int y = 1; y < 5; y++)
                    for (int z = 1; z < 5; z++)
                      for (int w = 1; w < 5; w++) P->int22[a][b][x][y][z][w] = tok_int(t.tok[k++]);
            fill_int22_nonstandard(P->int22);
          }
        } else if (!strcmp(s, "hairpin") || !strcmp(s, "bulge") || !strcmp(s, "interior")) {
          int *d = !strcmp(s, "hairpin") ? P->hairpin : !strcmp(s, "bulge") ? P->bulge : P->interior;
          if (need(&t, 31, s)) rc = -3;
          else
            for (int a = 0; a < 31; a++) d[a] = tok_int(t.tok[k++]);
        } else if (!strcmp(s, "ML_params")) {
          if (need(&t, 6, s)) rc = -3;
          else {
            P->MLbase = tok_int(t.tok[0]);
            P->MLclosing = tok_int(t.tok[2]);
            P->MLintern = tok_int(t.tok[4]);
          }
        } else if (!strcmp(s, "NINIO")) {
          if (need(&t, 3, s)) rc = -3;
          else {
            P->ninio = tok_int(t.tok[0]);
            P->max_ninio = tok_int(t.tok[2]);
          }
        } else if (!strcmp(s, "Misc")) {
          if (need(&t, 4, s)) rc = -3;
          else {
            P->DuplexInit = tok_int(t.tok[0]);
            P->TerminalAU = tok_int(t.tok[2]);
            if (t.n >= 5) P->lxc = atof(t.tok[4]);
          }
        } else if (!strcmp(s, "Tetraloops") || !strcmp(s, "Triloops") || !strcmp(s, "Hexaloops")) {
          int cnt = t.n / 3;
          if (cnt > ORC_MAXSPECIAL) cnt = ORC_MAXSPECIAL;
          for (int a = 0; a < cnt; a++) {
            const char *sq = t.tok[3 * a];
            int e = tok_int(t.tok[3 * a + 1]);
            if (s[0] == 'T' && s[1] == 'e') { /* Tetraloops */
              strncpy(P->tetra[a], sq, 6);
              P->tetra_E[a] = e;
            } else if (s[0] == 'T' && s[1] == 'r') {
              strncpy(P->tri[a], sq, 5);
              P->tri_E[a] = e;
            } else {
              strncpy(P->hexa[a], sq, 8);
              P->hexa_E[a] = e;
            }
          }
          if (s[0] == 'T' && s[1] == 'e') P->n_tetra = cnt;
          else if (s[0] == 'T' && s[1] == 'r') P->n_tri = cnt;
          else P->n_hexa = cnt;
        }
        /* *_enthalpies and unknown sections: parsed and ignored (unused at 37 C) */
      }
      t.n = 0;
      secname[0] = 0;
      if (at_end) break;
      /* new header: "# name", "#END", "# END", "## ..." */
      char *h = line + 1;
      while (*h == '#' || *h == ' ' || *h == '\t') h++;
      char nm[64];
      int k = 0;
      while (*h && !isspace((unsigned char)*h) && k < 63) nm[k++] = *h++;
      nm[k] = 0;
      if (!strcmp(nm, "END")) done = 1;
      else if (line[1] != '#') strcpy(secname, nm);
    } else if (!at_end) {
      for (char *p = strtok(line, " \t\r"); p; p = strtok(NULL, " \t\r")) tl_push(&t, p);
    }
    line = nl ? nl + 1 : NULL;
  }

  /* ViennaRNA get_scaled_params at 37 C with dangles != 0: exterior/multiloop mismatches and
   * dangles are bonuses only (clamped to <= 0). */
  memcpy(P->raw_mismatchM, P->mismatchM, sizeof(P->mismatchM));
  memcpy(P->raw_mismatchExt, P->mismatchExt, sizeof(P->mismatchExt));
  memcpy(P->raw_dangle5, P->dangle5, sizeof(P->dangle5));
  memcpy(P->raw_dangle3, P->dangle3, sizeof(P->dangle3));
  for (int p = 0; p < 8; p++)
    for (int a = 0; a < 5; a++) {
      if (P->dangle5[p][a] > 0) P->dangle5[p][a] = 0;
      if (P->dangle3[p][a] > 0) P->dangle3[p][a] = 0;
      for (int b = 0; b < 5; b++) {
        if (P->mismatchM[p][a][b] > 0) P->mismatchM[p][a][b] = 0;
        if (P->mismatchExt[p][a][b] > 0) P->mismatchExt[p][a][b] = 0;
      }
    }
  free(t.tok);
  free(buf);
  return rc;
}

int orc_params_export_size(void) {
  return 64 + 31 * 3 + 6 * 200 + 2 * 40 + 8 * 8 * 25 + 8 * 8 * 125 + 8 * 8 * 625 + 8 + 3 * (1 + ORC_MAXSPECIAL * 9);
}

int orc_params_export(const orc_params *P, int *out) {
  int k = 0;
  const int *src;
#define DUMP(arr, cnt)                       \
  src = (const int *)(arr);                  \
  for (int i_ = 0; i_ < (cnt); i_++) out[k++] = src[i_];
  DUMP(P->stack, 64)
  DUMP(P->hairpin, 31) DUMP(P->bulge, 31) DUMP(P->interior, 31)
  DUMP(P->mismatchH, 200) DUMP(P->mismatchI, 200) DUMP(P->mismatch1nI, 200)
  DUMP(P->mismatch23I, 200) DUMP(P->mismatchM, 200) DUMP(P->mismatchExt, 200)
  DUMP(P->dangle5, 40) DUMP(P->dangle3, 40)
  DUMP(P->int11, 8 * 8 * 25) DUMP(P->int21, 8 * 8 * 125) DUMP(P->int22, 8 * 8 * 625)
#undef DUMP
  out[k++] = P->MLbase;
  out[k++] = P->MLclosing;
  out[k++] = P->MLintern;
  out[k++] = P->ninio;
  out[k++] = P->max_ninio;
  out[k++] = P->DuplexInit;
  out[k++] = P->TerminalAU;
  out[k++] = (int)lround(P->lxc * 1000.0);
  for (int s = 0; s < 3; s++) {
    int n = s == 0 ? P->n_tetra : s == 1 ? P->n_tri : P->n_hexa;
    int w = s == 0 ? 6 : s == 1 ? 5 : 8;
    out[k++] = n;
    for (int a = 0; a < ORC_MAXSPECIAL; a++) {
      const char *sq = s == 0 ? P->tetra[a] : s == 1 ? P->tri[a] : P->hexa[a];
      for (int c = 0; c < 8; c++) out[k++] = (a < n && c < w) ? sq[c] : 0;
      out[k++] = a < n ? (s == 0 ? P->tetra_E[a] : s == 1 ? P->tri_E[a] : P->hexa_E[a]) : 0;
    }
  }
  return k;
}

/* ------------------------------------------------------------------ loop energies */

static inline int e_hairpin(const orc_params *P, int size, int type, int si1, int sj1, const char *str) {
  int e = (size <= 30) ? P->hairpin[size] : P->hairpin[30] + (int)(P->lxc * log(size / 30.));
  if (size < 3) return e;
  if (size == 4) {
    for (int k = 0; k < P->n_tetra; k++)
      if (!strncmp(P->tetra[k], str, 6)) return P->tetra_E[k];
  } else if (size == 6) {
    for (int k = 0; k < P->n_hexa; k++)
      if (!strncmp(P->hexa[k], str, 8)) return P->hexa_E[k];
  } else if (size == 3) {
    for (int k = 0; k < P->n_tri; k++)
      if (!strncmp(P->tri[k], str, 5)) return P->tri_E[k];
    return e + (type > 2 ? P->TerminalAU : 0);
  }
  return e + P->mismatchH[type][si1][sj1];
}

static inline int e_intloop(const orc_params *P, int n1, int n2, int type, int type_2, int si1, int sj1,
                            int sp1, int sq1) {
  int nl, ns, u, e;
  if (n1 > n2) { nl = n1; ns = n2; } else { nl = n2; ns = n1; }
  if (nl == 0) return P->stack[type][type_2];
  if (ns == 0) {
    e = (nl <= MAXLOOP) ? P->bulge[nl] : P->bulge[30] + (int)(P->lxc * log(nl / 30.));
    if (nl == 1) e += P->stack[type][type_2];
    else {
      if (type > 2) e += P->TerminalAU;
      if (type_2 > 2) e += P->TerminalAU;
    }
    return e;
  }
  if (ns == 1) {
    if (nl == 1) return P->int11[type][type_2][si1][sj1];
    if (nl == 2) {
      if (n1 == 1) return P->int21[type][type_2][si1][sq1][sj1];
      return P->int21[type_2][type][sq1][si1][sp1];
    }
    e = (nl + 1 <= MAXLOOP) ? P->interior[nl + 1] : P->interior[30] + (int)(P->lxc * log((nl + 1) / 30.));
    e += MIN2(P->max_ninio, (nl - ns) * P->ninio);
    e += P->mismatch1nI[type][si1][sj1] + P->mismatch1nI[type_2][sq1][sp1];
    return e;
  }
  if (ns == 2) {
    if (nl == 2) return P->int22[type][type_2][si1][sp1][sq1][sj1];
    if (nl == 3) {
      e = P->interior[5] + P->ninio;
      e += P->mismatch23I[type][si1][sj1] + P->mismatch23I[type_2][sq1][sp1];
      return e;
    }
  }
  u = nl + ns;
  e = (u <= MAXLOOP) ? P->interior[u] : P->interior[30] + (int)(P->lxc * log(u / 30.));
  e += MIN2(P->max_ninio, (nl - ns) * P->ninio);
  e += P->mismatchI[type][si1][sj1] + P->mismatchI[type_2][sq1][sp1];
  return e;
}

static inline int e_mlstem(const orc_params *P, int type, int si1, int sj1) {
  int e = 0;
  if (si1 >= 0 && sj1 >= 0) e += P->mismatchM[type][si1][sj1];
  else if (si1 >= 0) e += P->dangle5[type][si1];
  else if (sj1 >= 0) e += P->dangle3[type][sj1];
  if (type > 2) e += P->TerminalAU;
  return e + P->MLintern;
}

static inline int e_extloop(const orc_params *P, int type, int si1, int sj1) {
  int e = 0;
  if (si1 >= 0 && sj1 >= 0) e += P->mismatchExt[type][si1][sj1];
  else if (si1 >= 0) e += P->dangle5[type][si1];
  else if (sj1 >= 0) e += P->dangle3[type][sj1];
  if (type > 2) e += P->TerminalAU;
  return e;
}

/* ------------------------------------------------------------------ fold */

static __thread long long g_candidates;
long long orc_last_fold_candidates(void) { return g_candidates; }

typedef struct { int i, j, ml; } sector;

int orc_fold(const orc_params *P, const char *seq_in, int n, char *db, int *mfe_dcal) {
  g_candidates = 0;
  if (n < 0) return -1;
  for (int k = 0; k < n; k++) db[k] = '.';
  db[n] = 0;
  *mfe_dcal = 0;
  if (n <= TURN + 1) return 0;

  char *seq = (char *)malloc((size_t)n + 2);
  short *S = (short *)malloc(sizeof(short) * (size_t)(n + 2));
  for (int k = 0; k < n; k++) {
    seq[k] = (char)toupper((unsigned char)seq_in[k]);
    if (seq[k] == 'T') seq[k] = 'U'; /* DNA letters fold as RNA (as RNAfold's input filter does); unpinned */
    S[k + 1] = (short)encode_base(seq_in[k]);
  }
  seq[n] = 0;
  S[0] = S[n];
  S[n + 1] = S[1];

  size_t ncell = (size_t)(n + 1) * (size_t)(n + 2) / 2 + 2;
  int *indx = (int *)malloc(sizeof(int) * (size_t)(n + 2));
  for (int j = 0; j <= n + 1; j++) indx[j] = (j * (j - 1)) / 2;
  int *c = (int *)malloc(sizeof(int) * ncell);
  int *fML = (int *)malloc(sizeof(int) * ncell);
  int *DML = (int *)malloc(sizeof(int) * ncell);
  unsigned char *ptype = (unsigned char *)calloc(ncell, 1);
  int *f5 = (int *)malloc(sizeof(int) * (size_t)(n + 2));
  for (size_t k = 0; k < ncell; k++) c[k] = fML[k] = DML[k] = INF;
  for (int i = 1; i <= n; i++)
    for (int j = i + TURN + 1; j <= n; j++) ptype[indx[j] + i] = (unsigned char)PAIR[S[i]][S[j]];

  /* fill: mfe.c fill_arrays */
  for (int i = n - TURN - 1; i >= 1; i--) {
    for (int j = i + TURN + 1; j <= n; j++) {
      int ij = indx[j] + i;
      int type = ptype[ij];
      if (type) {
        int e = e_hairpin(P, j - i - 1, type, S[i + 1], S[j - 1], seq + i - 1);
        g_candidates++;
        int maxp = MIN2(j - 2 - TURN, i + MAXLOOP + 1);
        for (int p = i + 1; p <= maxp; p++) {
          int minq = j - i + p - MAXLOOP - 2;
          if (minq < p + 1 + TURN) minq = p + 1 + TURN;
          for (int q = j - 1; q >= minq; q--) {
            int t2 = ptype[indx[q] + p];
            if (!t2) continue;
            g_candidates++;
            int en = c[indx[q] + p] +
                     e_intloop(P, p - i - 1, j - q - 1, type, RTYPE[t2], S[i + 1], S[j - 1], S[p - 1], S[q + 1]);
            e = MIN2(e, en);
          }
        }
        /* multibranch closing: DML of row i+1 at column j-1 */
        {
          int d = DML[indx[j - 1] + i + 1];
          g_candidates++;
          int en = d + e_mlstem(P, RTYPE[type], S[j - 1], S[i + 1]) + P->MLclosing;
          e = MIN2(e, en);
        }
        c[ij] = e;
      }
      /* fML: multibranch.c vrna_E_ml_stems_fast, dangles=2 */
      {
        int e = fML[indx[j] + i + 1] + P->MLbase;
        e = MIN2(e, fML[indx[j - 1] + i] + P->MLbase);
        if (type) e = MIN2(e, c[ij] + e_mlstem(P, type, S[i - 1], S[j + 1]));
        int d = INF;
        for (int u = i + 1 + TURN; u <= j - 2 - TURN; u++) {
          int en = fML[indx[u] + i] + fML[indx[j] + u + 1];
          d = MIN2(d, en);
        }
        g_candidates += MAX2(0, j - i - 2 * TURN - 2) + 2;
        DML[ij] = d;
        fML[ij] = MIN2(e, d);
      }
    }
  }
  /* f5: external.c vrna_E_ext_loop_5, dangles=2 */
  for (int j = 0; j <= TURN + 1 && j <= n; j++) f5[j] = 0;
  for (int j = TURN + 2; j <= n; j++) {
    int e = f5[j - 1];
    int s3 = (j < n) ? S[j + 1] : -1;
    for (int i = j - TURN - 1; i >= 1; i--) {
      int type = ptype[indx[j] + i];
      if (!type) continue;
      g_candidates++;
      int en = f5[i - 1] + c[indx[j] + i] + e_extloop(P, type, (i > 1) ? S[i - 1] : -1, s3);
      e = MIN2(e, en);
    }
    f5[j] = e;
  }
  *mfe_dcal = f5[n];

  /* backtrack: mfe.c backtrack + vrna_BT_ext_loop_f5 / _hp_loop / _int_loop / _mb_loop / _mb_loop_split */
  sector *st = (sector *)malloc(sizeof(sector) * (size_t)(2 * n + 8));
  int s = 0, rc = 0;
  st[++s] = (sector){1, n, 0};
  while (s > 0) {
    int i = st[s].i, j = st[s].j, ml = st[s].ml;
    s--;
    if (ml == 0) {
      /* f5[j]: strip unpaired 3' bases, then first stem found scanning u downwards */
      int jj = j, fij, fi;
      do {
        fij = f5[jj];
        fi = f5[jj - 1];
        if (--jj == 0) break;
      } while (fij == fi);
      jj++;
      if (jj < TURN + 2) continue;
      int s3 = (jj < n) ? S[jj + 1] : -1, found = 0;
      for (int u = jj - TURN - 1; u >= 1; u--) {
        int type = ptype[indx[jj] + u];
        if (!type) continue;
        int en = c[indx[jj] + u] + e_extloop(P, type, (u > 1) ? S[u - 1] : -1, s3);
        if (fij == en + f5[u - 1]) {
          if (u - 1 > 0) st[++s] = (sector){1, u - 1, 0};
          i = u;
          j = jj;
          found = 1;
          break;
        }
      }
      if (!found) { rc = -10; break; }
    } else if (ml == 1) {
      int ii = i, jj = j, fij, fi;
      do { /* strip unpaired 3' bases */
        fij = fML[indx[jj] + ii];
        fi = fML[indx[jj - 1] + ii] + P->MLbase;
        if (--jj == 0) break;
      } while (fij == fi);
      jj++;
      do { /* strip unpaired 5' bases */
        fij = fML[indx[jj] + ii];
        fi = fML[indx[jj] + ii + 1] + P->MLbase;
        if (++ii == jj) break;
      } while (fij == fi);
      ii--;
      if (jj < ii + TURN + 1) { rc = -11; break; }
      int ij = indx[jj] + ii, type = ptype[ij];
      if (type && fij == c[ij] + e_mlstem(P, type, S[ii - 1], S[jj + 1])) {
        i = ii;
        j = jj; /* continue with pair (ii,jj) below */
      } else {
        int found = 0;
        for (int u = ii + 1 + TURN; u <= jj - 2 - TURN; u++) {
          if (fij == fML[indx[u] + ii] + fML[indx[jj] + u + 1]) {
            st[++s] = (sector){ii, u, 1};
            st[++s] = (sector){u + 1, jj, 1};
            found = 1;
            break;
          }
        }
        if (!found) { rc = -12; break; }
        continue;
      }
    }
    /* repeat1: (i,j) is a pair closing c[i,j] */
    for (;;) {
      db[i - 1] = '(';
      db[j - 1] = ')';
      int ij = indx[j] + i, cij = c[ij], type = ptype[ij];
      if (cij == e_hairpin(P, j - i - 1, type, S[i + 1], S[j - 1], seq + i - 1)) break;
      int found = 0;
      int maxp = MIN2(j - 2 - TURN, i + MAXLOOP + 1);
      for (int p = i + 1; p <= maxp && !found; p++) {
        int minq = j - i + p - MAXLOOP - 2;
        if (minq < p + 1 + TURN) minq = p + 1 + TURN;
        for (int q = j - 1; q >= minq; q--) {
          int t2 = ptype[indx[q] + p];
          if (!t2) continue;
          int en = c[indx[q] + p] +
                   e_intloop(P, p - i - 1, j - q - 1, type, RTYPE[t2], S[i + 1], S[j - 1], S[p - 1], S[q + 1]);
          if (en == cij) {
            i = p;
            j = q;
            found = 1;
            break;
          }
        }
      }
      if (found) continue;
      /* multibranch loop closed by (i,j) */
      int e = cij - e_mlstem(P, RTYPE[type], S[j - 1], S[i + 1]) - P->MLclosing;
      int p = i + 1, q = j - 1, r = q - TURN - 1;
      for (int u = p + TURN + 1; u < r; u++) {
        if (e == fML[indx[u] + p] + fML[indx[q] + u + 1]) {
          st[++s] = (sector){p, u, 1};
          st[++s] = (sector){u + 1, q, 1};
          found = 1;
          break;
        }
      }
      if (!found) rc = -13;
      break;
    }
    if (rc) break;
  }

  free(st);
  free(f5);
  free(ptype);
  free(DML);
  free(fML);
  free(c);
  free(indx);
  free(S);
  free(seq);
  return rc;
}

/* ------------------------------------------------------------------ independent evaluator */

static int eval_loop(const orc_params *P, const char *seq, const short *S, const int *pt, int n, int i, int j,
                     int *err) {
  /* energy of the loop closed by (i,j) plus everything inside */
  int type = PAIR[S[i]][S[j]];
  if (!type || j - i - 1 < TURN) { *err = 1; return 0; }
  /* collect branches */
  int nb = 0, p1 = 0, q1 = 0, unp = 0, sum = 0;
  for (int k = i + 1; k < j;) {
    if (pt[k] > k) {
      if (nb == 0) { p1 = k; q1 = pt[k]; }
      nb++;
      k = pt[k] + 1;
    } else { unp++; k++; }
  }
  if (nb == 0) return e_hairpin(P, j - i - 1, type, S[i + 1], S[j - 1], seq + i - 1);
  if (nb == 1) {
    int t2 = PAIR[S[p1]][S[q1]];
    if (!t2) { *err = 1; return 0; }
    /* any loop size: fc.eval_structure extrapolates above MAXLOOP (the folding recursions never build such a loop) */
    return e_intloop(P, p1 - i - 1, j - q1 - 1, type, RTYPE[t2], S[i + 1], S[j - 1], S[p1 - 1], S[q1 + 1]) +
           eval_loop(P, seq, S, pt, n, p1, q1, err);
  }
  sum = P->MLclosing + e_mlstem(P, RTYPE[type], S[j - 1], S[i + 1]) + unp * P->MLbase;
  for (int k = i + 1; k < j;) {
    if (pt[k] > k) {
      int t2 = PAIR[S[k]][S[pt[k]]];
      if (!t2) { *err = 1; return 0; }
      sum += e_mlstem(P, t2, S[k - 1], S[pt[k] + 1]) + eval_loop(P, seq, S, pt, n, k, pt[k], err);
      k = pt[k] + 1;
    } else k++;
  }
  return sum;
}

static int eval_pt(const orc_params *P, const char *seq, const short *S, const int *pt, int n, int *err) {
  int e = 0;
  for (int k = 1; k <= n;) {
    if (pt[k] > k) {
      int i = k, j = pt[k], type = PAIR[S[i]][S[j]];
      if (!type) { *err = 1; return 0; }
      e += e_extloop(P, type, i > 1 ? S[i - 1] : -1, j < n ? S[j + 1] : -1) + eval_loop(P, seq, S, pt, n, i, j, err);
      k = j + 1;
    } else k++;
  }
  return e;
}

int orc_eval_structure(const orc_params *P, const char *seq_in, int n, const char *db, int *e_dcal) {
  char *seq = (char *)malloc((size_t)n + 2);
  short *S = (short *)malloc(sizeof(short) * (size_t)(n + 2));
  int *pt = (int *)calloc((size_t)n + 2, sizeof(int));
  int *stk = (int *)malloc(sizeof(int) * (size_t)(n + 1));
  int sp = 0, err = 0;
  for (int k = 0; k < n; k++) {
    seq[k] = (char)toupper((unsigned char)seq_in[k]);
    if (seq[k] == 'T') seq[k] = 'U'; /* DNA letters fold as RNA (as RNAfold's input filter does); unpinned */
    S[k + 1] = (short)encode_base(seq_in[k]);
  }
  seq[n] = 0;
  if (n > 0) { S[0] = S[n]; S[n + 1] = S[1]; }
  for (int k = 1; k <= n && !err; k++) {
    if (db[k - 1] == '(') stk[sp++] = k;
    else if (db[k - 1] == ')') {
      if (!sp) err = 3;
      else { int o = stk[--sp]; pt[o] = k; pt[k] = o; }
    }
  }
  if (sp) err = 3;
  int e = 0;
  if (!err) e = eval_pt(P, seq, S, pt, n, &err);
  *e_dcal = e;
  free(stk); free(pt); free(S); free(seq);
  return err ? -err : 0;
}

typedef struct {
  const orc_params *P;
  const char *seq;
  const short *S;
  int n, best;
  int *pt;
} bf_ctx;

static void bf_rec(bf_ctx *x, int k, int *open, int nopen) {
  /* position k: unpaired, open, or close the innermost open bracket */
  if (k > x->n) {
    if (nopen) return;
    int err = 0, e = eval_pt(x->P, x->seq, x->S, x->pt, x->n, &err);
    if (!err && e < x->best) x->best = e;
    return;
  }
  bf_rec(x, k + 1, open, nopen);
  if (x->n - k >= TURN + 1 + nopen) {
    open[nopen] = k;
    bf_rec(x, k + 1, open, nopen + 1);
  }
  if (nopen) {
    int o = open[nopen - 1];
    if (k - o > TURN && PAIR[x->S[o]][x->S[k]]) {
      x->pt[o] = k; x->pt[k] = o;
      bf_rec(x, k + 1, open, nopen - 1);
      x->pt[o] = 0; x->pt[k] = 0;
      open[nopen - 1] = o;
    }
  }
}

int orc_bruteforce_mfe(const orc_params *P, const char *seq_in, int n, int *mfe_dcal) {
  if (n > 20) return -1;
  char seq[32];
  short S[34] = {0};
  int pt[34] = {0}, open[34];
  for (int k = 0; k < n; k++) {
    seq[k] = (char)toupper((unsigned char)seq_in[k]);
    if (seq[k] == 'T') seq[k] = 'U'; /* DNA letters fold as RNA (as RNAfold's input filter does); unpinned */
    S[k + 1] = (short)encode_base(seq_in[k]);
  }
  seq[n] = 0;
  if (n > 0) { S[0] = S[n]; S[n + 1] = S[1]; }
  bf_ctx x = {P, seq, S, n, INF, pt};
  bf_rec(&x, 1, open, 0);
  *mfe_dcal = x.best;
  return 0;
}

/* ------------------------------------------------------------------ hamming + batch */

int orc_hamming(const char *a, int na, const char *b, int nb, unsigned char *mismatch) {
  int n = na < nb ? na : nb, hd = 0;
  for (int k = 0; k < n; k++) {
    int m = a[k] != b[k];
    hd += m;
    if (mismatch) mismatch[k] = (unsigned char)m;
  }
  return hd;
}

int orc_fold_batch(const orc_params *P, const char *seqs, const int *off, int B, const char *targets,
                   char *db_out, int *mfe_out, int *hd_out, int nthreads) {
  int rc = 0;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#pragma omp parallel for schedule(dynamic, 4)
#endif
  for (int b = 0; b < B; b++) {
    int n = off[b + 1] - off[b];
    char *tmp = (char *)malloc((size_t)n + 1);
    int e = 0;
    int r = orc_fold(P, seqs + off[b], n, tmp, &e);
    if (r) {
#ifdef _OPENMP
#pragma omp critical
#endif
      rc = r;
    }
    memcpy(db_out + off[b], tmp, (size_t)n);
    mfe_out[b] = e;
    if (targets && hd_out) hd_out[b] = orc_hamming(targets + off[b], n, tmp, n, NULL);
    free(tmp);
  }
  return rc;
}
