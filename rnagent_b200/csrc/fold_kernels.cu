// fold_kernels.cu -- batched Zuker MFE fold + backtrack + Hamming/reward, sm_100a.
//
// One CTA folds one sequence at a time (persistent CTAs pull sequences from a length-sorted
// queue).  The c / fML tables are stored DIAGONAL-MAJOR: cell (i,j), d = j-i >= 4, lives at
// doff(d) + (i-1).  Every anti-diagonal only depends on shorter ones, so the fill is a wavefront
// with one __syncthreads per diagonal, and threads that own consecutive i on one diagonal touch
// consecutive words of every operand (bank-conflict free in shared memory, coalesced in global).
//
// Replaces ViennaRNA 2.4.14's fill_arrays + backtrack behind `RNA.fold` (reference binding
// rlif/settings.py:63-64; recursions restated in SURVEY.md section 8a' rows F6-F7) and fuses
// hamming_distance (rlif/rna/utils.py:67-75) and the reward of Solution.evaluate
// (rlif/rna/solution.py:145-148) into the tail of the same kernel.
#include <cuda_runtime.h>
#include <stdint.h>

#include "energy.cuh"
#include "fold_kernels.h"

namespace rnafold {

namespace {

__device__ __forceinline__ int doff(int n, int d) {
  // first cell of diagonal d (d >= 4) when diagonals 4..n-1 are stored back to back
  const int k = d - 4;
  return k * (n - 4) - (k * (k - 1)) / 2;
}

__device__ __forceinline__ int encode_base(unsigned char ch) {
  switch (ch) {
    case 'A': case 'a': return 1;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 3;
    case 'U': case 'u': case 'T': case 't': return 4;
    default: return 0;
  }
}

// Shared-memory carve-up shared by host (size) and device (pointers).
struct SmemLayout {
  int sp_off, f5_off, dml_off, bt_off, misc_off, S_off, db_off, tab_off, total;
  __host__ __device__ static SmemLayout make(int nmax, bool tables_in_smem) {
    SmemLayout L;
    int o = 0;
    L.sp_off = o;   o += (int)((sizeof(SmallParams) + 15) / 16 * 16);
    L.f5_off = o;   o += (nmax + 4) * 4;
    L.dml_off = o;  o += 3 * (nmax + 4) * 4;
    L.bt_off = o;   o += (nmax + 8) * 4;
    L.misc_off = o; o += 64;
    L.S_off = o;    o += (nmax + 8 + 15) / 16 * 16;
    L.db_off = o;   o += (nmax + 8 + 15) / 16 * 16;
    L.tab_off = o;
    if (tables_in_smem) o += 2 * table_cells(nmax) * 4;
    L.total = o;
    return L;
  }
};

}  // namespace

// ---------------------------------------------------------------------------------------------
// Generic CTA-per-sequence kernel.  kSmemTables: c/fML in shared memory (n <= ~230) or in a
// per-CTA slice of a global workspace (any n <= kMaxLen).
// ---------------------------------------------------------------------------------------------
template <bool kSmemTables>
__global__ void __launch_bounds__(kCtaThreads)
fold_cta_kernel(const FoldParams *__restrict__ gp, BatchArgs a, int nmax, int *__restrict__ gws,
                long long ws_stride, int *__restrict__ queue) {
  extern __shared__ __align__(16) unsigned char smem[];
  if (a.dyn && a.dyn[0] == 0) return;  // device-counted launch with nothing routed to it
  const SmemLayout L = SmemLayout::make(nmax, kSmemTables);
  SmallParams &sp = *reinterpret_cast<SmallParams *>(smem + L.sp_off);
  int *f5 = reinterpret_cast<int *>(smem + L.f5_off);
  int *dml = reinterpret_cast<int *>(smem + L.dml_off);
  int *bt = reinterpret_cast<int *>(smem + L.bt_off);
  int *misc = reinterpret_cast<int *>(smem + L.misc_off);
  unsigned char *S = smem + L.S_off;
  unsigned char *db = smem + L.db_off;
  int *c, *fml;
  if (kSmemTables) {
    c = reinterpret_cast<int *>(smem + L.tab_off);
    fml = c + table_cells(nmax);
  } else {
    c = gws + (long long)blockIdx.x * ws_stride;
    fml = c + ws_stride / 2;
  }
  const int tid = threadIdx.x;
  const int dml_stride = nmax + 4;

  // stage the small parameter tables once per CTA
  {
    const int *src = reinterpret_cast<const int *>(&gp->s);
    int *dst = reinterpret_cast<int *>(&sp);
    for (int k = tid; k < (int)(sizeof(SmallParams) / 4); k += kCtaThreads) dst[k] = __ldg(src + k);
  }

  for (;;) {
    __syncthreads();
    if (tid == 0) {
      misc[0] = atomicAdd(queue, 1);
      misc[1] = 0;
    }
    __syncthreads();
    const int slot = misc[0];
    if (slot >= batch_count(a)) break;
    const int b = batch_entry(a, slot);
    if (a.redo_filter && !a.redo[b]) continue;
    const int beg = a.off[b];
    const int n = a.off[b + 1] - beg;

    // ---- load + encode; S is 1-based with circular sentinels S[0]=S[n], S[n+1]=S[1]
    for (int k = tid; k < n; k += kCtaThreads) {
      S[k + 1] = (unsigned char)encode_base(a.seq[beg + k]);
      db[k] = '.';
    }
    for (int k = tid; k <= n; k += kCtaThreads) f5[k] = 0;
    __syncthreads();
    if (tid == 0 && n > 0) {
      S[0] = S[n];
      S[n + 1] = S[1];
    }
    __syncthreads();

    // ---- wavefront fill over anti-diagonals
    for (int d = kTurn + 1; d < n; ++d) {
      const int cells = n - d;
      const int base = doff(n, d);
      int *dml_d = dml + (d % 3) * dml_stride;
      const int *dml_d2 = dml + ((d + 1) % 3) * dml_stride;  // diagonal d-2
      for (int i = tid + 1; i <= cells; i += kCtaThreads) {
        const int j = i + d;
        const int type = pair_type(S[i], S[j]);
        int cij = kInf;
        if (type) {
          cij = e_hairpin(sp, gp, S, i, j, type);
          const int si1 = S[i + 1], sj1 = S[j - 1];
          // interior loops: p = i+1+n1, q = j-1-n2, n1+n2 <= 30, q-p >= 4
          const int amax = min(kMaxLoop, d - 6);
          for (int n1 = 0; n1 <= amax; ++n1) {
            const int p = i + 1 + n1;
            const int bmax = min(kMaxLoop - n1, d - 6 - n1);
            const int sp1 = S[p - 1], Sp = S[p];
            for (int n2 = 0; n2 <= bmax; ++n2) {
              const int q = j - 1 - n2;
              const int t2 = pair_type(Sp, S[q]);
              if (!t2) continue;
              const int cpq = c[doff(n, q - p) + p - 1];
              const int e = cpq + e_intloop(sp, gp, n1, n2, type, rtype_of(t2), si1, sj1, sp1, S[q + 1]);
              cij = min(cij, e);
            }
          }
          // multibranch loop closed by (i,j): DML of (i+1, j-1) from two diagonals back
          if (d >= 2 * kTurn + 4) {
            const int e = dml_d2[i + 1] + e_mlstem(sp, rtype_of(type), sj1, si1) + sp.MLclosing;
            cij = min(cij, e);
          }
        }
        c[base + i - 1] = cij;
        // fML(i,j)
        int e = kInf;
        if (d > kTurn + 1) {
          const int pb = doff(n, d - 1);
          e = min(fml[pb + i], fml[pb + i - 1]) + sp.MLbase;  // fML(i+1,j), fML(i,j-1)
        }
        if (type) e = min(e, cij + e_mlstem(sp, type, S[i - 1], S[j + 1]));
        int dm = kInf;
        for (int t = kTurn + 1; t <= d - kTurn - 2; ++t) {
          // u = i+t: fML(i,u) on diagonal t, fML(u+1,j) on diagonal d-t-1
          const int v = fml[doff(n, t) + i - 1] + fml[doff(n, d - t - 1) + i + t];
          dm = min(dm, v);
        }
        dml_d[i] = dm;
        fml[base + i - 1] = min(e, dm);
      }
      __syncthreads();
    }

    // ---- f5 (one warp; sequential in j, lanes over i)
    if (tid < 32) {
      for (int j = kTurn + 2; j <= n; ++j) {
        int best = kInf;
        const int s3 = (j < n) ? (int)S[j + 1] : -1;
        for (int i = j - kTurn - 1 - tid; i >= 1; i -= 32) {
          const int type = pair_type(S[i], S[j]);
          if (!type) continue;
          const int e = f5[i - 1] + c[doff(n, j - i) + i - 1] + e_extloop(sp, type, i > 1 ? (int)S[i - 1] : -1, s3);
          best = min(best, e);
        }
        for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
        if (tid == 0) f5[j] = min(f5[j - 1], best);
        __syncwarp();
      }
    }
    __syncthreads();

    // ---- backtrack (ViennaRNA's deterministic order), one thread
    if (tid == 0) {
      int sp_top = 0;
      if (n > 0) bt[sp_top++] = (1 << 14) | (n << 2) | 0;
      while (sp_top > 0) {
        const int w = bt[--sp_top];
        int i = w >> 14, j = (w >> 2) & 0xfff, ml = w & 3;
        if (ml == 0) {
          int jj = j, fij, fi;
          do {
            fij = f5[jj];
            fi = f5[jj - 1];
            if (--jj == 0) break;
          } while (fij == fi);
          jj++;
          if (jj < kTurn + 2) continue;
          const int s3 = (jj < n) ? (int)S[jj + 1] : -1;
          int u;
          for (u = jj - kTurn - 1; u >= 1; --u) {
            const int type = pair_type(S[u], S[jj]);
            if (!type) continue;
            const int e = c[doff(n, jj - u) + u - 1] + e_extloop(sp, type, u > 1 ? (int)S[u - 1] : -1, s3);
            if (fij == e + f5[u - 1]) break;
          }
          if (u < 1) { misc[1] = 1; break; }
          if (u - 1 > 0) bt[sp_top++] = (1 << 14) | ((u - 1) << 2) | 0;
          i = u;
          j = jj;
        } else {
          int ii = i, jj = j, fij, fi;
          do {
            fij = fml[doff(n, jj - ii) + ii - 1];
            fi = (jj - 1 - ii > kTurn) ? fml[doff(n, jj - 1 - ii) + ii - 1] + sp.MLbase : kInf;
            if (--jj == 0) break;
          } while (fij == fi);
          jj++;
          do {
            fij = fml[doff(n, jj - ii) + ii - 1];
            fi = (jj - ii - 1 > kTurn) ? fml[doff(n, jj - ii - 1) + ii] + sp.MLbase : kInf;
            if (++ii == jj) break;
          } while (fij == fi);
          ii--;
          if (jj < ii + kTurn + 1) { misc[1] = 2; break; }
          const int type = pair_type(S[ii], S[jj]);
          if (type && fij == c[doff(n, jj - ii) + ii - 1] + e_mlstem(sp, type, S[ii - 1], S[jj + 1])) {
            i = ii;
            j = jj;
          } else {
            int u;
            for (u = ii + 1 + kTurn; u <= jj - 2 - kTurn; ++u)
              if (fij == fml[doff(n, u - ii) + ii - 1] + fml[doff(n, jj - u - 1) + u]) break;
            if (u > jj - 2 - kTurn) { misc[1] = 3; break; }
            bt[sp_top++] = (ii << 14) | (u << 2) | 1;
            bt[sp_top++] = ((u + 1) << 14) | (jj << 2) | 1;
            continue;
          }
        }
        // (i,j) pairs; walk down the helix / interior loops
        bool fail = false;
        for (;;) {
          db[i - 1] = '(';
          db[j - 1] = ')';
          const int d = j - i;
          const int type = pair_type(S[i], S[j]);
          const int cij = c[doff(n, d) + i - 1];
          if (cij == e_hairpin(sp, gp, S, i, j, type)) break;
          const int si1 = S[i + 1], sj1 = S[j - 1];
          bool found = false;
          const int amax = min(kMaxLoop, d - 6);
          for (int n1 = 0; n1 <= amax && !found; ++n1) {
            const int p = i + 1 + n1;
            const int bmax = min(kMaxLoop - n1, d - 6 - n1);
            for (int n2 = 0; n2 <= bmax; ++n2) {
              const int q = j - 1 - n2;
              const int t2 = pair_type(S[p], S[q]);
              if (!t2) continue;
              const int e = c[doff(n, q - p) + p - 1] +
                            e_intloop(sp, gp, n1, n2, type, rtype_of(t2), si1, sj1, S[p - 1], S[q + 1]);
              if (e == cij) {
                i = p;
                j = q;
                found = true;
                break;
              }
            }
          }
          if (found) continue;
          const int e = cij - e_mlstem(sp, rtype_of(type), sj1, si1) - sp.MLclosing;
          const int p = i + 1, q = j - 1;
          int u;
          for (u = p + kTurn + 1; u < q - kTurn - 1; ++u)
            if (e == fml[doff(n, u - p) + p - 1] + fml[doff(n, q - u - 1) + u]) break;
          if (u >= q - kTurn - 1) { fail = true; break; }
          bt[sp_top++] = (p << 14) | (u << 2) | 1;
          bt[sp_top++] = ((u + 1) << 14) | (q << 2) | 1;
          break;
        }
        if (fail) { misc[1] = 4; break; }
      }
      misc[2] = 0;  // hamming accumulator
    }
    __syncthreads();

    // ---- emit structure, energy, Hamming distance, reward
    int local_hd = 0;
    for (int k = tid; k < n; k += kCtaThreads) {
      const unsigned char ch = db[k];
      if (a.db) a.db[beg + k] = ch;
      if (a.target) {
        const int mm = a.target[beg + k] != ch;
        local_hd += mm;
        if (a.mismatch) a.mismatch[beg + k] = (unsigned char)mm;
      }
    }
    if (a.target) {
      for (int o = 16; o; o >>= 1) local_hd += __shfl_xor_sync(0xffffffffu, local_hd, o);
      if ((tid & 31) == 0 && local_hd) atomicAdd(&misc[2], local_hd);
    }
    __syncthreads();
    if (tid == 0) {
      if (a.mfe) a.mfe[b] = (n > 0) ? f5[n] : 0;
      if (a.target) {
        const int hd = misc[2];
        if (a.hd) a.hd[b] = hd;
        if (a.reward) {
          double r = (n > 0) ? pow(1.0 - (double)hd / (double)n, a.reward_exp) : 0.0;
          if (hd == 0) r += 0.25;
          a.reward[b] = r;
        }
      }
      if (misc[1]) {
        atomicExch(a.status, misc[1]);
        misc[1] = 0;
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------
// host-side launchers
// ---------------------------------------------------------------------------------------------
size_t cta_kernel_smem_bytes(int nmax, bool tables_in_smem) {
  return (size_t)SmemLayout::make(nmax, tables_in_smem).total;
}

cudaError_t launch_fold_cta(const FoldParams *gp, const BatchArgs &a, int nmax, bool tables_in_smem,
                            int grid, int *gws, long long ws_stride, int *queue, cudaStream_t st) {
  const size_t smem = cta_kernel_smem_bytes(nmax, tables_in_smem);
  cudaError_t err;
  if (tables_in_smem) {
    err = cudaFuncSetAttribute(fold_cta_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    fold_cta_kernel<true><<<grid, kCtaThreads, smem, st>>>(gp, a, nmax, gws, ws_stride, queue);
  } else {
    err = cudaFuncSetAttribute(fold_cta_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    fold_cta_kernel<false><<<grid, kCtaThreads, smem, st>>>(gp, a, nmax, gws, ws_stride, queue);
  }
  return cudaGetLastError();
}

int cta_kernel_occupancy(int nmax, bool tables_in_smem) {
  int nb = 0;
  const size_t smem = cta_kernel_smem_bytes(nmax, tables_in_smem);
  if (tables_in_smem) {
    cudaFuncSetAttribute(fold_cta_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fold_cta_kernel<true>, kCtaThreads, smem);
  } else {
    cudaFuncSetAttribute(fold_cta_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fold_cta_kernel<false>, kCtaThreads, smem);
  }
  return nb;
}

}  // namespace rnafold
