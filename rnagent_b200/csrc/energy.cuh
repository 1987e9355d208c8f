// energy.cuh -- Turner nearest-neighbour loop energies on the device (dangles=2, 37 C).
//
// Restates, for the GPU, the loop-energy evaluators that ViennaRNA 2.4.14 runs under the reference's
// `RNA.fold` (rlif/settings.py:63-64): E_Hairpin, E_IntLoop, E_MLstem, E_ExtLoop
// (SURVEY.md section 8a' rows F3-F5).  Small tables come from shared memory (`sp`), the int11 /
// int21 / int22 and special-hairpin tables from global memory through the read-only path (`gp`).
#pragma once
#include <cuda_runtime.h>

#include "fold_params.h"

namespace rnafold {

__device__ __forceinline__ int pair_type(int a, int b) {
  // bases N,A,C,G,U = 0..4 -> CG=1 GC=2 GU=3 UG=4 AU=5 UA=6, 0 = cannot pair.
  // Packed 5x5 table, 3 bits per entry, row-major (a*5+b).
  // rows: N:00000 A:0,0,0,0,5 C:0,0,0,1,0 G:0,0,2,0,3 U:0,6,0,4,0
  // rows N,A,C,G live in `lo` (20 entries), row U in `hi`.
  const unsigned long long lo =
      (5ull << (3 * 9)) | (1ull << (3 * 13)) | (2ull << (3 * 17)) | (3ull << (3 * 19));
  const unsigned hi = (6u << (3 * 1)) | (4u << (3 * 3));
  return (a < 4) ? (int)((lo >> (3 * (a * 5 + b))) & 7ull) : (int)((hi >> (3 * b)) & 7u);
}

__device__ __forceinline__ int rtype_of(int t) {
  // {0,2,1,4,3,6,5,7}
  return (int)((0x75634120u >> (4 * t)) & 0xf);
}

// window key for special hairpins: 3 bits per base, first base in the low bits; non-ACGU -> 7
__device__ __forceinline__ unsigned special_key(const unsigned char *S, int i, int len) {
  unsigned key = 0;
  for (int k = 0; k < len; ++k) {
    unsigned c = S[i + k];
    key |= (c ? c : 7u) << (3 * k);
  }
  return key;
}

// E_Hairpin for the loop closed by (i,j); S is the 1-based code array.
__device__ __forceinline__ int e_hairpin(const SmallParams &sp, const FoldParams *__restrict__ gp,
                                         const unsigned char *S, int i, int j, int type) {
  const int size = j - i - 1;
  int e = (size <= 30) ? sp.hairpin[size] : __ldg(&gp->hairpin_len[size]);
  if (size == 4) {
    const unsigned key = special_key(S, i, 6);
    for (int k = 0; k < sp.n_tetra; ++k)
      if (__ldg(&gp->tetra_key[k]) == key) return __ldg(&gp->tetra_E[k]);
  } else if (size == 6) {
    const unsigned key = special_key(S, i, 8);
    for (int k = 0; k < sp.n_hexa; ++k)
      if (__ldg(&gp->hexa_key[k]) == key) return __ldg(&gp->hexa_E[k]);
  } else if (size == 3) {
    const unsigned key = special_key(S, i, 5);
    for (int k = 0; k < sp.n_tri; ++k)
      if (__ldg(&gp->tri_key[k]) == key) return __ldg(&gp->tri_E[k]);
    return e + (type > 2 ? sp.TerminalAU : 0);
  }
  return e + sp.mmH[type][S[i + 1]][S[j - 1]];
}

// E_IntLoop: n1 = p-i-1 unpaired 5' of the inner pair, n2 = j-q-1 on the 3' side;
// type = type(i,j), type2 = rtype[type(p,q)].
__device__ __forceinline__ int e_intloop(const SmallParams &sp, const FoldParams *__restrict__ gp, int n1,
                                         int n2, int type, int type2, int si1, int sj1, int sp1,
                                         int sq1) {
  const int nl = max(n1, n2), ns = min(n1, n2);
  if (nl == 0) return sp.stack[type][type2];
  if (ns == 0) {
    int e = sp.bulge[nl];
    if (nl == 1) e += sp.stack[type][type2];
    else {
      if (type > 2) e += sp.TerminalAU;
      if (type2 > 2) e += sp.TerminalAU;
    }
    return e;
  }
  if (ns == 1) {
    if (nl == 1) return __ldg(&gp->int11[type][type2][si1][sj1]);
    if (nl == 2) {
      if (n1 == 1) return __ldg(&gp->int21[type][type2][si1][sq1][sj1]);
      return __ldg(&gp->int21[type2][type][sq1][si1][sp1]);
    }
    return sp.interior[nl + 1] + min(sp.max_ninio, (nl - ns) * sp.ninio) + sp.mm1nI[type][si1][sj1] +
           sp.mm1nI[type2][sq1][sp1];
  }
  if (ns == 2) {
    if (nl == 2) return __ldg(&gp->int22[type][type2][si1][sp1][sq1][sj1]);
    if (nl == 3)
      return sp.interior[5] + sp.ninio + sp.mm23I[type][si1][sj1] + sp.mm23I[type2][sq1][sp1];
  }
  return sp.interior[nl + ns] + min(sp.max_ninio, (nl - ns) * sp.ninio) + sp.mmI[type][si1][sj1] +
         sp.mmI[type2][sq1][sp1];
}

// E_MLstem with both neighbours present (dangles=2 inside the sequence).
__device__ __forceinline__ int e_mlstem(const SmallParams &sp, int type, int s5, int s3) {
  return sp.mmM[type][s5][s3] + (type > 2 ? sp.TerminalAU : 0) + sp.MLintern;
}

// E_ExtLoop; s5 / s3 < 0 mean "beyond the sequence end".
__device__ __forceinline__ int e_extloop(const SmallParams &sp, int type, int s5, int s3) {
  int e = (type > 2) ? sp.TerminalAU : 0;
  if (s5 >= 0 && s3 >= 0) e += sp.mmExt[type][s5][s3];
  else if (s5 >= 0) e += sp.dangle5[type][s5];
  else if (s3 >= 0) e += sp.dangle3[type][s3];
  return e;
}

// E_ExtLoop without its TerminalAU term (for callers that read c + TerminalAU(pair) tables)
__device__ __forceinline__ int e_extloop_mm(const SmallParams &sp, int type, int s5, int s3) {
  if (s5 >= 0 && s3 >= 0) return sp.mmExt[type][s5][s3];
  if (s5 >= 0) return sp.dangle5[type][s5];
  if (s3 >= 0) return sp.dangle3[type][s3];
  return 0;
}

}  // namespace rnafold
