// evaluate.cuh -- free energy of a GIVEN secondary structure on the device (fc.eval_structure of ViennaRNA 2.4.14 with
// the reference's model, dangles=2: rlif/rna/solution.py:232,234 evaluate the centroid and the MEA structure).
//
// A structure's energy is the sum of its loops' energies: every pair (i,j) closes one loop (hairpin, interior loop or
// multiloop, decided by the pairs directly inside it), plus the exterior loop.  The loops are independent, so a CTA
// spreads the opening positions over its threads and adds the terms up in shared memory.  Unlike the folding recursions
// the evaluator accepts interior loops and bulges of any size (the lxc * log(u / 30) extrapolation of E_IntLoop, taken from
// the host-evaluated hairpin_len table: hairpin_len[u] - hairpin[30] is that term, truncated by the host libm).
#pragma once
#include "energy.cuh"

namespace rnafold {

// loop-size extrapolation above 30 unpaired bases: (int)(lxc * log(u / 30.))
__device__ __forceinline__ int size_extra(const FoldParams *__restrict__ gp, int u) {
  return __ldg(&gp->hairpin_len[min(u, kMaxLen)]) - __ldg(&gp->s.hairpin[30]);
}

// E_IntLoop for any loop size (sizes <= 30 go through e_intloop)
__device__ __forceinline__ int e_intloop_any(const SmallParams &sp, const FoldParams *__restrict__ gp, int n1, int n2,
                                             int type, int type2, int si1, int sj1, int sp1, int sq1) {
  const int nl = max(n1, n2), ns = min(n1, n2), u = n1 + n2;
  if (u <= kMaxLoop) return e_intloop(sp, gp, n1, n2, type, type2, si1, sj1, sp1, sq1);
  if (ns == 0) return sp.bulge[30] + size_extra(gp, nl) + (type > 2 ? sp.TerminalAU : 0) + (type2 > 2 ? sp.TerminalAU : 0);
  if (ns == 1)
    return sp.interior[30] + size_extra(gp, nl + 1) + min(sp.max_ninio, (nl - ns) * sp.ninio) + sp.mm1nI[type][si1][sj1] +
           sp.mm1nI[type2][sq1][sp1];
  return sp.interior[30] + size_extra(gp, u) + min(sp.max_ninio, (nl - ns) * sp.ninio) + sp.mmI[type][si1][sj1] +
         sp.mmI[type2][sq1][sp1];
}

// Energy (dcal/mol) of the structure whose pair table is pt[1..n] (pt[k] = partner or 0) on the coded sequence S[1..n]
// (S[0] = S[n], S[n+1] = S[1] as everywhere).  Called by ALL threads of the CTA; `acc` points at two shared ints
// ({sum, error}) that the caller need not initialise.  Returns the energy to every thread; *err_out gets 0, 1 (a bracket
// pair that cannot pair, or a hairpin shorter than 3) as orc_eval_structure does.
__device__ inline int eval_structure_cta(const SmallParams &sp, const FoldParams *__restrict__ gp, const unsigned char *S,
                                         const int *pt, int n, int *acc, int *err_out) {
  const int tid = threadIdx.x, nt = blockDim.x;
  __syncthreads();
  if (tid == 0) acc[0] = acc[1] = 0;
  __syncthreads();
  int sum = 0, err = 0;
  for (int i = 1 + tid; i <= n; i += nt) {
    const int j = pt[i];
    if (j <= i) continue;
    const int type = pair_type(S[i], S[j]);
    if (!type || j - i - 1 < kTurn) {
      err = 1;
      continue;
    }
    int nb = 0, p1 = 0, q1 = 0, unp = 0;
    for (int k = i + 1; k < j;) {
      if (pt[k] > k) {
        if (nb == 0) {
          p1 = k;
          q1 = pt[k];
        }
        ++nb;
        k = pt[k] + 1;
      } else {
        ++unp;
        ++k;
      }
    }
    if (nb == 0) {
      sum += e_hairpin(sp, gp, S, i, j, type);
    } else if (nb == 1) {
      const int t2 = pair_type(S[p1], S[q1]);
      if (t2)
        sum += e_intloop_any(sp, gp, p1 - i - 1, j - q1 - 1, type, rtype_of(t2), S[i + 1], S[j - 1], S[p1 - 1], S[q1 + 1]);
    } else {
      int e = sp.MLclosing + e_mlstem(sp, rtype_of(type), S[j - 1], S[i + 1]) + unp * sp.MLbase;
      for (int k = i + 1; k < j;) {
        if (pt[k] > k) {
          const int t2 = pair_type(S[k], S[pt[k]]);
          if (t2) e += e_mlstem(sp, t2, S[k - 1], S[pt[k] + 1]);
          k = pt[k] + 1;
        } else {
          ++k;
        }
      }
      sum += e;
    }
  }
  if (tid == 0) {   // exterior loop
    for (int k = 1; k <= n;) {
      if (pt[k] > k) {
        const int j = pt[k], type = pair_type(S[k], S[j]);
        if (type) sum += e_extloop(sp, type, k > 1 ? (int)S[k - 1] : -1, j < n ? (int)S[j + 1] : -1);
        k = j + 1;
      } else {
        ++k;
      }
    }
  }
  for (int o = 16; o; o >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    err |= __shfl_xor_sync(0xffffffffu, err, o);
  }
  if ((tid & 31) == 0) {
    if (sum) atomicAdd(&acc[0], sum);
    if (err) atomicOr(&acc[1], err);
  }
  __syncthreads();
  const int total = acc[0];
  *err_out = acc[1];
  __syncthreads();
  return total;
}

}  // namespace rnafold
