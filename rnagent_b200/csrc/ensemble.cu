// ensemble.cu -- ensemble statistics of Solution.compute_statistics (reference: rlif/rna/solution.py:221-238:
// fc.pf(), fc.positional_entropy(), fc.centroid(), fc.pr_structure(fold), fc.mean_bp_distance(), fc.ensemble_defect(fold),
// RNA.dist_mountain(target, fold) of ViennaRNA 2.4.14) for a batch of sequences: McCaskill's partition function and
// base-pair probabilities on the loop model of the fold kernels (dangles=2, MAXLOOP=30, TURN=3), fp64.
//
// One CTA per sequence (persistent, queue), one WARP per cell: a cell's sums (interior loops <= 496 terms, multiloop /
// exterior splits O(n) terms) are spread over the lanes and reduced with shuffles; cells of an anti-diagonal are independent,
// one __syncthreads per diagonal -- the same wavefront as the MFE kernels, upwards for the inside pass and downwards for
// the outside pass.  The outside multiloop term is collected through two helper tables
//     X(i,l) = sum_{j>l} W(i,j) (b^(j-l-1) + qm(l+1,j-1)),   Y(i,l) = sum_{j>l} W(i,j) b^(j-l-1),
//     W(i,j) = P_out(i,j) * expMLclosing * expMLstem(closing pair) * scale^2,
// so that a pair (k,l) inside a multiloop costs one O(n) sum like every other term (O(n^3) in all, every sum a contiguous
// row: the tables read along their second index are kept transposed).
// Tables are fp64 in a global workspace (10 (n+2)^2 doubles per resident CTA; L2-resident at n ~ 100): this path serves
// the report columns of the reference's CLI / comparison logs, not the RL inner loop.
//
// Boltzmann factors exp(-E * 10 / kT) of the integer dcal/mol parameters are tabulated on the HOST in double (libm);
// as in ViennaRNA's get_scaled_exp_params the dangle / mismatchM / mismatchExt terms go through SMOOTH() instead of the
// min(0, .) clamp of the MFE model, and long hairpins use the un-truncated lxc * log(u / 30) term.
// Two quirks of the reference's library are kept because its logged numbers pin them (tests/test_oracle_golden.py):
// ensemble_defect leaves the last position out of its sum, and dist_mountain records the mountain height at paired
// positions only.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/rnafold_b200.h"
#include "energy.cuh"
#include "evaluate.cuh"
#include "fold_kernels.h"

namespace rnafold {

struct ExpParams {
  double kT;
  double stack[8][8], bulge[32], interior[32];
  double mmH[8][5][5], mmI[8][5][5], mm1nI[8][5][5], mm23I[8][5][5], mmM[8][5][5], mmExt[8][5][5];
  double d5[8][5], d3[8][5];
  double MLbase, MLclosing, MLintern, TermAU, ninio[kMaxLoop + 2];
  double tetra[kMaxSpecial], tri[kMaxSpecial], hexa[kMaxSpecial];
  double int11[8][8][5][5], int21[8][8][5][5][5], int22[8][8][5][5][5][5];
  double hairpin_len[kMaxLen + 1];  // hairpin Boltzmann factor by loop size (the lxc * log term above 30, un-truncated)
};

namespace {

double smooth_fn(double x) {  // ViennaRNA params/basic.c SMOOTH(X), SCALE = 10
  const double s = x / 10.;
  if (s < -1.2283697) return 0.;
  if (s > 0.8660254) return x;
  const double t = std::sin(s - 0.34242663) + 1.;
  return 10. * 0.38490018 * t * t;
}

void make_exp_params(const HostParams &hp, ExpParams *E) {
  const SmallParams &s = hp.dev.s;
  const double kT = (37. + 273.15) * 1.98717;
  E->kT = kT;
  auto BF = [&](double e) { return std::exp(-e * 10. / kT); };
  auto SM = [&](int raw) { return std::exp(smooth_fn(-(double)raw) * 10. / kT); };
  for (int a = 0; a < 8; ++a)
    for (int b = 0; b < 8; ++b) E->stack[a][b] = BF(s.stack[a][b]);
  for (int k = 0; k <= 30; ++k) {
    E->bulge[k] = BF(s.bulge[k]);
    E->interior[k] = BF(s.interior[k]);
  }
  for (int t = 0; t < 8; ++t)
    for (int a = 0; a < 5; ++a) {
      for (int b = 0; b < 5; ++b) {
        E->mmH[t][a][b] = BF(s.mmH[t][a][b]);
        E->mmI[t][a][b] = BF(s.mmI[t][a][b]);
        E->mm1nI[t][a][b] = BF(s.mm1nI[t][a][b]);
        E->mm23I[t][a][b] = BF(s.mm23I[t][a][b]);
        E->mmM[t][a][b] = SM(hp.raw_mmM[t][a][b]);
        E->mmExt[t][a][b] = SM(hp.raw_mmExt[t][a][b]);
      }
      E->d5[t][a] = SM(hp.raw_dangle5[t][a]);
      E->d3[t][a] = SM(hp.raw_dangle3[t][a]);
    }
  for (int a = 0; a < 8; ++a)
    for (int b = 0; b < 8; ++b)
      for (int c = 0; c < 5; ++c)
        for (int d = 0; d < 5; ++d) {
          E->int11[a][b][c][d] = BF(hp.dev.int11[a][b][c][d]);
          for (int e = 0; e < 5; ++e) {
            E->int21[a][b][c][d][e] = BF(hp.dev.int21[a][b][c][d][e]);
            for (int f = 0; f < 5; ++f) E->int22[a][b][c][d][e][f] = BF(hp.dev.int22[a][b][c][d][e][f]);
          }
        }
  E->MLbase = BF(s.MLbase);
  E->MLclosing = BF(s.MLclosing);
  E->MLintern = BF(s.MLintern);
  E->TermAU = BF(s.TerminalAU);
  for (int k = 0; k <= kMaxLoop + 1; ++k) E->ninio[k] = BF(std::min(s.max_ninio, k * s.ninio));
  for (int k = 0; k < kMaxSpecial; ++k) {
    E->tetra[k] = BF(hp.dev.tetra_E[k]);
    E->tri[k] = BF(hp.dev.tri_E[k]);
    E->hexa[k] = BF(hp.dev.hexa_E[k]);
  }
  const double h30 = BF(s.hairpin[30]);
  for (int u = 0; u <= kMaxLen; ++u)
    E->hairpin_len[u] = (u <= 30) ? BF(s.hairpin[u]) : h30 * std::exp(-(hp.lxc * std::log(u / 30.)) * 10. / kT);
}

// ---- device loop weights (exp_E_Hairpin / exp_E_IntLoop / exp_E_MLstem / exp_E_ExtLoop)
__device__ __forceinline__ double x_hairpin(const ExpParams *__restrict__ E, const FoldParams *__restrict__ gp,
                                            const unsigned char *S, int i, int j, int type) {
  const int u = j - i - 1;
  const double q = E->hairpin_len[u];
  if (u == 4) {
    const unsigned key = special_key(S, i, 6);
    for (int k = 0; k < gp->s.n_tetra; ++k)
      if (gp->tetra_key[k] == key) return E->tetra[k];
  } else if (u == 6) {
    const unsigned key = special_key(S, i, 8);
    for (int k = 0; k < gp->s.n_hexa; ++k)
      if (gp->hexa_key[k] == key) return E->hexa[k];
  } else if (u == 3) {
    const unsigned key = special_key(S, i, 5);
    for (int k = 0; k < gp->s.n_tri; ++k)
      if (gp->tri_key[k] == key) return E->tri[k];
    return type > 2 ? q * E->TermAU : q;
  }
  return q * E->mmH[type][S[i + 1]][S[j - 1]];
}

__device__ __forceinline__ double x_intloop(const ExpParams *__restrict__ E, int u1, int u2, int type, int type2, int si1,
                                            int sj1, int sp1, int sq1) {
  const int ul = max(u1, u2), us = min(u1, u2);
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

__device__ __forceinline__ double x_mlstem(const ExpParams *__restrict__ E, int type, int s5, int s3) {
  double e = 1.;
  if (s5 >= 0 && s3 >= 0) e = E->mmM[type][s5][s3];
  else if (s5 >= 0) e = E->d5[type][s5];
  else if (s3 >= 0) e = E->d3[type][s3];
  if (type > 2) e *= E->TermAU;
  return e * E->MLintern;
}

__device__ __forceinline__ double x_extloop(const ExpParams *__restrict__ E, int type, int s5, int s3) {
  double e = 1.;
  if (s5 >= 0 && s3 >= 0) e = E->mmExt[type][s5][s3];
  else if (s5 >= 0) e = E->d5[type][s5];
  else if (s3 >= 0) e = E->d3[type][s3];
  if (type > 2) e *= E->TermAU;
  return e;
}

__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kEnsThreads = 256;
constexpr int kEnsWarps = kEnsThreads / 32;
constexpr int kEnsTables = 10;
// F, ensemble defect, p(structure), mean bp distance, centroid distance, mountain distance, MEA, centroid energy, MEA energy
constexpr int kEnsStats = RNAFOLD_ENSEMBLE_STATS;

struct EnsArgs {
  const uint8_t *seq;
  const int32_t *off;
  int32_t B;
  const uint8_t *db;       // structure whose defect / probability is reported (the MFE fold), CSR like seq
  const int32_t *e_dcal;   // its energy, dcal/mol
  const uint8_t *target;   // dot-bracket targets for the mountain distance, or nullptr
  double *stats;           // [B][kEnsStats]
  uint8_t *centroid;       // CSR like seq, or nullptr
  double *pos;             // [total][3]: P(paired with a downstream base), P(paired upstream), positional entropy (bits); or nullptr
  double *bpp;             // single-sequence calls only: n * n row-major upper triangle, or nullptr
  uint8_t *mea;            // maximum-expected-accuracy structure, CSR like seq, or nullptr
  double gamma;            // its pair weight (fc.MEA() default 1)
};

__global__ void __launch_bounds__(kEnsThreads)
ensemble_kernel(const ExpParams *__restrict__ E, const FoldParams *__restrict__ gp, EnsArgs a, int nmax, double *ws,
                long long ws_stride, int *queue) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int np = (nmax + 4 + 3) & ~3;
  double *scale = reinterpret_cast<double *>(smem);
  double *bpow = scale + np;
  double *red = bpow + np;                       // [8] block-level accumulators
  int *pt = reinterpret_cast<int *>(red + 8);    // pair table of the reported structure
  int *pt2 = pt + np;                            // pair table of the target / stack scratch
  int *ptc = pt2 + np;                           // pair table of the centroid
  int *ptm = ptc + np;                           // pair table of the MEA structure
  int *misc = ptm + np;
  unsigned char *S = reinterpret_cast<unsigned char *>(misc + 8);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double *base = ws + (long long)blockIdx.x * ws_stride;

  for (;;) {
    __syncthreads();
    if (tid == 0) misc[0] = atomicAdd(queue, 1);
    __syncthreads();
    const int b = misc[0];
    if (b >= a.B) break;
    const int beg = a.off[b], n = a.off[b + 1] - beg;
    double *st = a.stats + (size_t)b * kEnsStats;
    if (n <= 0) {
      if (tid < kEnsStats) st[tid] = 0.;
      continue;
    }
    const int N = n + 2;
    const size_t NN = (size_t)N * N;
    double *qb = base, *qm = qb + NN, *qmT = qm + NN, *qm1T = qmT + NN, *q = qm1T + NN, *q1T = q + NN, *po = q1T + NN,
           *W = po + NN, *XT = W + NN, *YT = XT + NN;
#define IX(i, j) ((size_t)(i) * N + (j))
    const double kT = E->kT;
    // vrna_exp_params_rescale(&mfe): pf_scale = exp(-(1.07 * mfe per nucleotide) / kT), never below 1
    double pf_scale = exp(-(1.07 * ((double)a.e_dcal[b] * 10. / n)) / kT);
    if (pf_scale < 1.) pf_scale = 1.;
    for (int k = tid; k < n; k += kEnsThreads) {
      int c = 0;
      switch (a.seq[beg + k]) {
        case 'A': case 'a': c = 1; break;
        case 'C': case 'c': c = 2; break;
        case 'G': case 'g': c = 3; break;
        case 'U': case 'u': case 'T': case 't': c = 4; break;
      }
      S[k + 1] = (unsigned char)c;
    }
    if (tid == 0) {
      scale[0] = bpow[0] = 1.;
      for (int k = 1; k <= n + 2; ++k) {
        scale[k] = scale[k - 1] / pf_scale;
        bpow[k] = bpow[k - 1] * E->MLbase / pf_scale;
      }
    }
    for (size_t k = tid; k < (size_t)kEnsTables * NN; k += kEnsThreads) base[k] = 0.;
    __syncthreads();
    if (tid == 0) {
      S[0] = S[n];
      S[n + 1] = S[1];
    }
    for (int i = 1 + tid; i <= n + 1; i += kEnsThreads) {
      q[IX(i, i - 1)] = 1.;
      for (int j = i; j <= min(n, i + kTurn); ++j) q[IX(i, j)] = scale[j - i + 1];
    }
    __syncthreads();
    const double clos2 = E->MLclosing * scale[2];

    // ---------------- inside: diagonals upwards
    for (int d = kTurn + 1; d < n; ++d) {
      for (int i = 1 + warp; i + d <= n; i += kEnsWarps) {
        const int j = i + d;
        const int type = pair_type(S[i], S[j]);
        double acc = 0.;
        if (type) {
          if (lane == 0) acc = x_hairpin(E, gp, S, i, j, type) * scale[d + 1];
          const int si1 = S[i + 1], sj1 = S[j - 1];
          for (int u1 = 0; u1 <= kMaxLoop; ++u1) {
            const int k = i + 1 + u1;
            if (k > j - kTurn - 2) break;
            const int lmin = max(k + kTurn + 1, j - 1 - kMaxLoop + u1);
            for (int l = lmin + lane; l < j; l += 32) {
              const int t2 = pair_type(S[k], S[l]);
              if (t2) acc += qb[IX(k, l)] * x_intloop(E, u1, j - l - 1, type, rtype_of(t2), si1, sj1, S[k - 1], S[l + 1]) * scale[u1 + j - l + 1];
            }
          }
          double t = 0.;
          for (int k = i + 2 + lane; k <= j - 1; k += 32) t += qm[IX(i + 1, k - 1)] * qm1T[IX(j - 1, k)];
          acc += t * clos2 * x_mlstem(E, rtype_of(type), sj1, si1);
        }
        const double bq = warp_sum(acc);
        const int s5 = (i > 1) ? (int)S[i - 1] : -1, s3 = (j < n) ? (int)S[j + 1] : -1;
        const double m1 = qm1T[IX(j - 1, i)] * bpow[1] + (type ? bq * x_mlstem(E, type, s5, s3) : 0.);
        const double qq = q1T[IX(j - 1, i)] * scale[1] + (type ? bq * x_extloop(E, type, s5, s3) : 0.);
        double mm = 0., tq = 0.;
        for (int k = i + 1 + lane; k <= j; k += 32) mm += (qm[IX(i, k - 1)] + bpow[k - i]) * qm1T[IX(j, k)];
        for (int k = i + lane; k < j; k += 32) tq += q[IX(i, k)] * q1T[IX(j, k + 1)];
        mm = warp_sum(mm);
        tq = warp_sum(tq);
        if (lane == 0) {
          qb[IX(i, j)] = bq;
          qm1T[IX(j, i)] = m1;
          qm[IX(i, j)] = mm + m1;
          qmT[IX(j, i)] = mm + m1;
          q1T[IX(j, i)] = qq;
          q[IX(i, j)] = scale[d + 1] + qq + tq;
        }
      }
      __syncthreads();
    }
    const double Z = q[IX(1, n)];

    // ---------------- outside: diagonals downwards
    for (int d = n - 1; d > kTurn; --d) {
      for (int k = 1 + warp; k + d <= n; k += kEnsWarps) {
        const int l = k + d;
        const int t2 = pair_type(S[k], S[l]);
        const double qbkl = qb[IX(k, l)];
        double p = 0., w = 0.;
        if (t2 && qbkl != 0.) {   // warp-uniform
          double acc = 0.;
          if (lane == 0) acc = q[IX(1, k - 1)] * q[IX(l + 1, n)] / Z * x_extloop(E, t2, k > 1 ? (int)S[k - 1] : -1, l < n ? (int)S[l + 1] : -1);
          const int rt2 = rtype_of(t2), sp1 = S[k - 1], sq1 = S[l + 1];
          for (int u1 = 0; u1 <= kMaxLoop; ++u1) {
            const int i = k - 1 - u1;
            if (i < 1) break;
            for (int u2 = lane; u2 <= kMaxLoop - u1; u2 += 32) {
              const int j = l + 1 + u2;
              if (j > n) break;
              const int type = pair_type(S[i], S[j]);
              if (!type) continue;
              const double pij = po[IX(i, j)];
              if (pij != 0.) acc += pij * x_intloop(E, u1, u2, type, rt2, S[i + 1], S[j - 1], sp1, sq1) * scale[u1 + u2 + 2];
            }
          }
          if (l < n) {
            double m = 0.;
            for (int i = 1 + lane; i < k; i += 32)
              m += (bpow[k - i - 1] + qmT[IX(k - 1, i + 1)]) * XT[IX(l, i)] - bpow[k - i - 1] * YT[IX(l, i)];
            acc += m * x_mlstem(E, t2, sp1, sq1);
          }
          p = warp_sum(acc);
          w = p * clos2 * x_mlstem(E, rt2, S[l - 1], S[k + 1]);
        }
        if (lane == 0) {
          po[IX(k, l)] = p;
          W[IX(k, l)] = w;
        }
      }
      __syncthreads();
      for (int i = 1 + warp; i + d <= n; i += kEnsWarps) {
        const int l = i + d;
        double x = 0., y = 0.;
        for (int j = l + 1 + lane; j <= n; j += 32) {
          const double w = W[IX(i, j)];
          if (w != 0.) {
            x += w * (bpow[j - l - 1] + qm[IX(l + 1, j - 1)]);
            y += w * bpow[j - l - 1];
          }
        }
        x = warp_sum(x);
        y = warp_sum(y);
        if (lane == 0) {
          XT[IX(l, i)] = x;
          YT[IX(l, i)] = y;
        }
      }
      __syncthreads();
    }

    // ---------------- statistics on the base-pair probabilities P(i,j) = po(i,j) qb(i,j)
    if (tid < 8) red[tid] = 0.;
    if (tid == 0) {   // pair tables of the reported structure and of the target (unbalanced brackets are skipped)
      for (int w = 0; w < 2; ++w) {
        int *tab = w ? pt2 : pt;
        const uint8_t *src = w ? a.target : a.db;
        for (int k = 0; k <= n + 1; ++k) tab[k] = 0;
        if (!src) continue;
        int *stk = reinterpret_cast<int *>(XT);   // the helper tables are free now
        int top = 0;
        for (int k = 1; k <= n; ++k) {
          const uint8_t c = src[beg + k - 1];
          if (c == '(') stk[top++] = k;
          else if (c == ')' && top > 0) {
            const int o = stk[--top];
            tab[o] = k;
            tab[k] = o;
          }
        }
      }
    }
    __syncthreads();
    double mbd = 0., cdist = 0., ed = 0.;
    const double mea_cut = 1e-4 / (1. + a.gamma);   // fc.MEA() lists the pairs with p >= 1e-4 / (1 + gamma)
    double *pu_mea = scale;                         // the scaling powers are not needed any more
    for (int i = 1 + tid; i <= n; i += kEnsThreads) {
      double up = 0., down = 0., ent = 0.;   // paired with a downstream base (j > i) / an upstream one
      double listed = 0.;
      char cen = '.';
      int cpart = 0;
      for (int j = 1; j <= n; ++j) {
        if (j == i) continue;
        const int lo = min(i, j), hi = max(i, j);
        if (hi - lo <= kTurn) continue;
        const double p = po[IX(lo, hi)] * qb[IX(lo, hi)];
        if (p == 0.) continue;
        if (j > i) {
          down += p;
          mbd += 2. * p * (1. - p);
          cdist += (p > 0.5) ? 1. - p : p;
          if (a.bpp) a.bpp[(size_t)(i - 1) * n + (j - 1)] = p;
        } else up += p;
        if (p > 0.5) {
          cen = (j > i) ? '(' : ')';
          cpart = j;
        }
        if (p >= mea_cut) listed += p;
        ent -= p * log(p);
      }
      ptc[i] = cpart;
      ptm[i] = 0;
      pu_mea[i] = 1. - listed;
      const double pu = 1. - up - down;
      if (pu > 0.) ent -= pu * log(pu);
      if (a.pos) {
        double *o = a.pos + (size_t)(beg + i - 1) * 3;
        o[0] = down;
        o[1] = up;
        o[2] = ent / log(2.);
      }
      if (a.centroid) a.centroid[beg + i - 1] = (uint8_t)cen;
      // vrna_ensemble_defect of ViennaRNA 2.4.14: positions 1 .. n-1 only, divided by n
      if (i < n) {
        const int pr = pt[i];
        ed += pr ? 1. - po[IX(min(i, pr), max(i, pr))] * qb[IX(min(i, pr), max(i, pr))] : up + down;
      }
    }
    mbd = warp_sum(mbd);
    cdist = warp_sum(cdist);
    ed = warp_sum(ed);
    if (lane == 0) {
      atomicAdd(&red[0], mbd);
      atomicAdd(&red[1], cdist);
      atomicAdd(&red[2], ed);
    }
    __syncthreads();
    if (tid == 0) {
      const double F = (-log(Z) - n * log(pf_scale)) * kT / 1000.;
      st[0] = F;
      st[1] = red[2] / n;
      st[2] = exp((F * 1000. - (double)a.e_dcal[b] * 10.) / kT);
      st[3] = red[0];
      st[4] = red[1];
      // RNA.dist_mountain(target, structure) of ViennaRNA 2.4.14: height recorded at paired positions only, L1
      double md = 0.;
      if (a.target) {
        double h1 = 0., h2 = 0.;
        for (int k = 1; k <= n; ++k) {
          double f1 = 0., f2 = 0.;
          if (pt2[k]) {
            h1 += (pt2[k] > k ? 1. : -1.) / (double)abs(pt2[k] - k);
            f1 = h1;
          }
          if (pt[k]) {
            h2 += (pt[k] > k ? 1. : -1.) / (double)abs(pt[k] - k);
            f2 = h2;
          }
          md += fabs(f1 - f2);
        }
      }
      st[5] = md;
    }
    // ---------------- fc.MEA(gamma): maximum expected accuracy structure (rlif/rna/solution.py:233)
    //   M(i,j) = max( M(i,j-1) + pu[j],  max_{i<k} M(i,k-1) + A(k,j),  A(i,j) ),   A(i,j) = 2 gamma P(i,j) + M(i+1,j-1)
    // over the listed pairs with 2 gamma P > pu[i] + pu[j]; as the library does, a pair's A is remembered only where it
    // strictly improved M(i,j).  The library rolls two rows and re-derives rows while backtracking; here the whole table is
    // filled by anti-diagonals (a warp per cell, lanes across k), M in the W table, A transposed in XT (0 = no candidate).
    {
      double *M = W, *AT = XT;
      const double g2 = 2. * a.gamma;
      __syncthreads();
      for (size_t k = tid; k < NN; k += kEnsThreads) AT[k] = 0.;
      for (int i = 1 + tid; i <= n + 1; i += kEnsThreads) {
        M[IX(i, i - 1)] = 0.;
        if (i <= n) M[IX(i, i)] = pu_mea[i];
      }
      __syncthreads();
      for (int d = 1; d < n; ++d) {
        for (int i = 1 + warp; i + d <= n; i += kEnsWarps) {
          const int j = i + d;
          double best = M[IX(i, j - 1)] + pu_mea[j];
          for (int k = i + 1 + lane; k <= j - kTurn - 1; k += 32) {
            const double A = AT[IX(j, k)];
            if (A > 0.) best = fmax(best, A + M[IX(i, k - 1)]);
          }
          for (int o = 16; o; o >>= 1) best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
          if (lane == 0) {
            if (d > kTurn) {
              const double p = po[IX(i, j)] * qb[IX(i, j)];
              if (p >= mea_cut && p * g2 > pu_mea[i] + pu_mea[j]) {
                const double EA = g2 * p + M[IX(i + 1, j - 1)];
                if (best < EA) {
                  best = EA;
                  AT[IX(j, i)] = EA;
                }
              }
            }
            M[IX(i, j)] = best;
          }
        }
        __syncthreads();
      }
      if (tid == 0) {   // backtracking: intervals on a stack (the YT table is free), accepting equality within DBL_EPSILON * M
        int *stk = reinterpret_cast<int *>(YT);
        int top = 0;
        stk[top++] = 1;
        stk[top++] = n;
        while (top > 0) {
          int j = stk[--top];
          const int i = stk[--top];
          if (j <= i) continue;
          const double prec = 2.220446049250313e-16 * M[IX(i, j)];
          while (j > i && M[IX(i, j)] <= M[IX(i, j - 1)] + pu_mea[j] + prec) --j;
          if (j <= i) continue;
          for (int k = j - kTurn - 1; k >= i; --k) {
            const double A = AT[IX(j, k)];
            if (A > 0. && M[IX(i, j)] <= A + M[IX(i, k - 1)] + prec) {
              ptm[k] = j;
              ptm[j] = k;
              if (k > i + 3) {
                stk[top++] = i;
                stk[top++] = k - 1;
              }
              stk[top++] = k + 1;
              stk[top++] = j - 1;
              break;
            }
          }
        }
        st[6] = M[IX(1, n)];
      }
      __syncthreads();
      if (a.mea)
        for (int i = 1 + tid; i <= n; i += kEnsThreads) a.mea[beg + i - 1] = (uint8_t)(ptm[i] == 0 ? '.' : ptm[i] > i ? '(' : ')');
      // fc.eval_structure of the centroid and of the MEA structure (solution.py:232,234), kcal/mol
      int err = 0;
      const int e_cen = eval_structure_cta(gp->s, gp, S, ptc, n, misc + 2, &err);
      const int e_mea = eval_structure_cta(gp->s, gp, S, ptm, n, misc + 2, &err);
      if (tid == 0) {
        st[7] = e_cen / 100.;
        st[8] = e_mea / 100.;
      }
    }
#undef IX
  }
}

size_t ens_smem_bytes(int nmax) {
  const int np = (nmax + 4 + 3) & ~3;
  return (size_t)np * 16 + 64 + (size_t)np * 16 + 32 + (size_t)np + 16;
}

struct EnsContext {
  ExpParams *d_E = nullptr;
  uint64_t epoch = ~0ull;
  double *d_ws = nullptr;
  size_t ws_cap = 0;
  int *d_queue = nullptr;
  std::vector<void *> staged;
};
EnsContext g_ens;
std::mutex g_ens_mu;

}  // namespace
}  // namespace rnafold

using namespace rnafold;

#define ENS_TRY(expr)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      char buf_[256];                                                                   \
      std::snprintf(buf_, sizeof(buf_), "%s: %s", #expr, cudaGetErrorString(e_));       \
      for (void *p_ : tmp) cudaFree(p_);                                                \
      return set_error(RNAFOLD_ECUDA, buf_);                                            \
    }                                                                                   \
  } while (0)

extern "C" int rnafold_ensemble_batch(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *structure,
                                      const int32_t *e_dcal, const uint8_t *target, double *stats_out,
                                      uint8_t *centroid_out, double *pos_out, double *bpp_out, double mea_gamma,
                                      uint8_t *mea_out) {
  std::lock_guard<std::mutex> lk(g_ens_mu);
  std::vector<void *> tmp;
  CtxView cv;
  if (ctx_view(&cv)) return set_error(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (B < 0 || !off || (B > 0 && (!seq || !structure || !e_dcal || !stats_out)))
    return set_error(RNAFOLD_EINVAL, "ensemble: null seq / off / structure / energies / stats or negative batch size");
  if (B == 0) return RNAFOLD_OK;
  if (!(mea_gamma > 0.)) return set_error(RNAFOLD_EINVAL, "ensemble: the MEA pair weight gamma must be positive");
  if (bpp_out && B != 1) return set_error(RNAFOLD_EINVAL, "ensemble: the base-pair probability matrix is returned for single-sequence calls only");
  if (off[0] != 0) return set_error(RNAFOLD_EINVAL, "off[0] must be 0");
  int nmax = 0;
  for (int b = 0; b < B; ++b) {
    const int n = off[b + 1] - off[b];
    if (n < 0 || n > RNAFOLD_ENSEMBLE_MAX_LEN) {
      char buf[128];
      std::snprintf(buf, sizeof(buf), "ensemble: sequence %d has length %d (limit RNAFOLD_ENSEMBLE_MAX_LEN = %d)", b, n, RNAFOLD_ENSEMBLE_MAX_LEN);
      return set_error(RNAFOLD_EINVAL, buf);
    }
    nmax = std::max(nmax, n);
  }
  ENS_TRY(cudaSetDevice(cv.device));
  cudaStream_t st = cv.stream;
  EnsContext &G = g_ens;
  if (!G.d_E) ENS_TRY(cudaMalloc((void **)&G.d_E, sizeof(ExpParams)));
  if (!G.d_queue) ENS_TRY(cudaMalloc((void **)&G.d_queue, sizeof(int)));
  if (G.epoch != cv.params_epoch) {   // Boltzmann factors of the live parameter set
    ExpParams *h = new ExpParams();
    make_exp_params(*cv.hp, h);
    cudaError_t e = cudaMemcpy(G.d_E, h, sizeof(ExpParams), cudaMemcpyHostToDevice);
    delete h;
    ENS_TRY(e);
    G.epoch = cv.params_epoch;
  }
  const size_t total = (size_t)off[B];
  auto stage = [&](const void *src, size_t bytes, void **dev) -> cudaError_t {
    cudaError_t e = cudaMalloc(dev, bytes ? bytes : 1);
    if (e != cudaSuccess) return e;
    tmp.push_back(*dev);
    return src ? cudaMemcpyAsync(*dev, src, bytes, cudaMemcpyDefault, st) : cudaSuccess;
  };
  EnsArgs a{};
  void *p = nullptr;
  ENS_TRY(stage(seq, total, &p)); a.seq = (const uint8_t *)p;
  ENS_TRY(stage(off, sizeof(int32_t) * ((size_t)B + 1), &p)); a.off = (const int32_t *)p;
  ENS_TRY(stage(structure, total, &p)); a.db = (const uint8_t *)p;
  ENS_TRY(stage(e_dcal, sizeof(int32_t) * (size_t)B, &p)); a.e_dcal = (const int32_t *)p;
  if (target) { ENS_TRY(stage(target, total, &p)); a.target = (const uint8_t *)p; }
  ENS_TRY(stage(nullptr, sizeof(double) * kEnsStats * (size_t)B, &p)); a.stats = (double *)p;
  if (centroid_out) { ENS_TRY(stage(nullptr, total, &p)); a.centroid = (uint8_t *)p; }
  if (pos_out) { ENS_TRY(stage(nullptr, sizeof(double) * 3 * total, &p)); a.pos = (double *)p; }
  if (bpp_out) {
    ENS_TRY(stage(nullptr, sizeof(double) * (size_t)nmax * nmax, &p));
    a.bpp = (double *)p;
    ENS_TRY(cudaMemsetAsync(p, 0, sizeof(double) * (size_t)nmax * nmax, st));
  }
  if (mea_out) { ENS_TRY(stage(nullptr, total, &p)); a.mea = (uint8_t *)p; }
  a.gamma = mea_gamma;
  a.B = B;
  const long long stride = (long long)kEnsTables * (nmax + 2) * (nmax + 2);
  int grid = std::min(B, cv.num_sms);
  const size_t budget = (size_t)6 << 30;
  while (grid > 1 && (size_t)grid * (size_t)stride * sizeof(double) > budget) --grid;
  const size_t need = (size_t)grid * (size_t)stride * sizeof(double);
  if (need > G.ws_cap) {
    if (G.d_ws) cudaFree(G.d_ws);
    G.d_ws = nullptr;
    G.ws_cap = 0;
    ENS_TRY(cudaMalloc((void **)&G.d_ws, need));
    G.ws_cap = need;
  }
  ENS_TRY(cudaMemsetAsync(G.d_queue, 0, sizeof(int), st));
  const size_t smem = ens_smem_bytes(nmax);
  ENS_TRY(cudaFuncSetAttribute(ensemble_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  ensemble_kernel<<<grid, kEnsThreads, smem, st>>>(G.d_E, cv.d_params, a, nmax, G.d_ws, stride, G.d_queue);
  ENS_TRY(cudaGetLastError());
  ENS_TRY(cudaMemcpyAsync(stats_out, a.stats, sizeof(double) * kEnsStats * (size_t)B, cudaMemcpyDefault, st));
  if (centroid_out) ENS_TRY(cudaMemcpyAsync(centroid_out, a.centroid, total, cudaMemcpyDefault, st));
  if (pos_out) ENS_TRY(cudaMemcpyAsync(pos_out, a.pos, sizeof(double) * 3 * total, cudaMemcpyDefault, st));
  if (bpp_out) ENS_TRY(cudaMemcpyAsync(bpp_out, a.bpp, sizeof(double) * (size_t)nmax * nmax, cudaMemcpyDefault, st));
  if (mea_out) ENS_TRY(cudaMemcpyAsync(mea_out, a.mea, total, cudaMemcpyDefault, st));
  ENS_TRY(cudaStreamSynchronize(st));
  for (void *q : tmp) cudaFree(q);
  return RNAFOLD_OK;
}
