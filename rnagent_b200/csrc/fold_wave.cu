// fold_wave.cu -- the fast fold kernel: anti-diagonal wavefront with warp-level work items.
//
// One CTA folds one sequence at a time (persistent CTAs, length-sorted queue).  Per anti-diagonal
// d ("phase") the CTA's warps share a list of independent work items, one __syncthreads per phase:
//
//   G(c)    the 375 generic interior-loop shapes of ONE pairable cell, lanes across shapes: per candidate one
//           LDS [lane slot + uniform cell offset] and one DPX VIADDMNMX.
//   E(b,r)  bulges and 1xn loops of loop-size range r for a block of 32 pairable cells (lanes across cells), the
//           loop-size terms warp-uniform from the constant bank.
//   M(b,m)  multibranch split partial minima  min_t fML(i,i+t) + fML(i+t+1,j)  of diagonal d.
//   T(b)    the 9 tabulated small loop shapes (stack, 1-bulge, 1x1, 1x2, 2x1, 2x2, 2x3, 3x2) of diagonal d.
//   F(b)    finalisation of diagonal d-1: hairpin, multiloop closing, combination of the partials,
//           fML, and the pre-added table variants read by G() and E().
//   X       one step of the exterior recursion f5[d-1].
//
// Storage (diagonal-major): full triangles of cB = c + TerminalAU(pair) and fML, plus 32-diagonal
// rolling windows of cI = c + mismatchI(inner side) and c1 = c + mismatch1nI(inner side), so that a
// generic / 1xn / bulge candidate is a single add.  Tables are int16 (fast path, range
// checked on every store; a violation flags the sequence for the int32 kernels) or int32.
//
// Replaces ViennaRNA 2.4.14 fill_arrays + backtrack behind `RNA.fold` (rlif/settings.py:63-64;
// recursions SURVEY.md 8a' F3-F7), fused with hamming_distance (rlif/rna/utils.py:67-75) and the
// reward of Solution.evaluate (rlif/rna/solution.py:145-148).
#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstring>

#include "energy.cuh"
#include "fold_kernels.h"

namespace rnafold {

// warp-uniform constants of the live parameter set (constant bank -> usable as instruction operands)
struct WaveConst {
  int GI[31][32];  // generic shapes (n1, L-n1): interior[L] + asymmetry penalty; kBig where the shape is not generic
  int GB[32];      // bulges (0,L) and (L,0): bulge[L]
  int G1[32];      // 1xn loops (1,L-1) and (L-1,1): interior[L] + asymmetry penalty
  int MLbase, MLclosing, MLintern, TerminalAU;
  // mismatch tables of the fill's finalisation items: constant-cache copies (lane-divergent index, but only a
  // handful of distinct entries per warp) instead of global loads that miss the small L1 left beside 225 KB of smem
  int mmH[8][5][5], mmI[8][5][5], mm1nI[8][5][5], mmM[8][5][5], mmExt[8][5][5];
};
__constant__ WaveConst cW;
// The small tables once more in the constant bank, for the backtrack of the latency-bound instances (one CTA per
// SM: the wide-CTA kernel of small batches / long sequences and the cluster path's f5_backtrack_kernel): the
// backtrack's lookups are warp-uniform (one pair at a time), so they become constant-cache broadcasts instead of
// dependent global loads.  The throughput instance (128 threads, 5 CTAs/SM) keeps the global path: its register
// allocation is at the edge and must not change.
__constant__ SmallParams cSP;
template <bool CP>
__device__ __forceinline__ const SmallParams &small_params(const FoldParams *__restrict__ gp) {
  if constexpr (CP) return cSP;
  else return gp->s;
}

#ifdef WAVE_TRACE   // debug builds only (tools/wave_trace.py): per-warp time stamps of every phase of one fold (CTA 0, first fold)
__device__ unsigned g_wave_trace[512 * 16 * 8];
#define WTRACE(slot_) do { if (wtracing && lane == 0 && d < 512) g_wave_trace[(d * 16 + (warp & 15)) * 8 + (slot_)] = (unsigned)(clock64() - wtrace_t0); } while (0)
#define WTRACE_VAL(slot_, v_) do { if (wtracing && lane == 0 && d < 512) g_wave_trace[(d * 16 + (warp & 15)) * 8 + (slot_)] = (unsigned)(v_); } while (0)
#else
#define WTRACE(slot_) do { } while (0)
#define WTRACE_VAL(slot_, v_) do { } while (0)
#endif
#ifdef WAVE_PHASE_CLK  // debug builds only: cycles per phase, summed over all folds (tools/phase_clocks.py)
__device__ unsigned long long g_phase_clk[640];
#ifndef WAVE_SKIP
#define WAVE_SKIP 0  // debug: bit mask of item kinds to skip (timing experiments only, results are wrong): 1 G, 2 E, 4 T, 8 M, 16 F, 32 X
#endif
#define PHASE_CLK_BEGIN() long long clk_t0 = (tid == 0) ? clock64() : 0
#define PHASE_CLK_END(slot) do { if (tid == 0) { const long long t1 = clock64(); atomicAdd(&g_phase_clk[slot], (unsigned long long)(t1 - clk_t0)); clk_t0 = t1; } } while (0)
#define ITEM_CLK_BEGIN() const long long item_t0 = clock64()
#define ITEM_CLK_END(kind, late) do { if (lane == 0) { atomicAdd(&g_phase_clk[610 + 8 * (late) + (kind)], (unsigned long long)(clock64() - item_t0)); atomicAdd(&g_phase_clk[626 + 8 * (late) + (kind)], 1ull); } } while (0)
#else
#define WAVE_SKIP 0
#define ITEM_CLK_BEGIN() do { } while (0)
#define ITEM_CLK_END(kind, late) do { } while (0)
#define PHASE_CLK_BEGIN() do { } while (0)
#define PHASE_CLK_END(slot) do { } while (0)
#endif

namespace {

constexpr int kWin = 32;    // rolling window depth (diagonals d-2 .. d-32 are read by phase d)
#ifndef WAVE_CMM
#define WAVE_CMM 0  // 1: the fill reads the mismatch tables from constant memory
#endif
#ifndef WAVE_MLWANT
#define WAVE_MLWANT 1  // cluster kernel: multiloop work items per warp the split chunks aim for
#endif
#ifndef WAVE_EPARTS
#define WAVE_EPARTS 2
#endif
#ifndef WAVE_EMID
#define WAVE_EMID 16  // first loop size of the second bulge / 1xn part
#endif
constexpr int kEParts = WAVE_EPARTS;  // loop-size ranges of the bulge / 1xn work items of the wave kernel
constexpr int kWSlots = 2 + kEParts;  // wave kernel scratch rows: generic shapes, tabulated shapes, one per E part
__host__ __device__ constexpr int epart_beg(int r) {
#if WAVE_EPARTS == 2
  return r == 0 ? 2 : r == 1 ? WAVE_EMID : 31;
#elif WAVE_EPARTS == 1
  return r == 0 ? 2 : 31;
#else
  return r == 0 ? 2 : r == 1 ? 12 : r == 2 ? 22 : 31;
#endif
}
constexpr int kMaxMlParts = 4;
constexpr int kBig = 10000000;  // additive term that can never win a minimum
constexpr int kFullWinMax = 128;  // up to this length the cI / c1 tables are full triangles (cheaper row offsets)
#ifndef WAVE_MLCHUNK
#define WAVE_MLCHUNK 128
#endif
constexpr int kMlChunk = WAVE_MLCHUNK;  // split points per multiloop work item
constexpr int kGenSlots = 12 * 32;  // lane slots of the shape-parallel generic interior item (375 shapes + padding)

template <typename ST> struct Store;
template <> struct Store<int16_t> {
  static constexpr int kInfS = 30000;   // stored "infinity"
  static constexpr int kThr = 15000;    // anything >= kThr is infinity
  static constexpr int kLo = -25000;    // real values must stay inside (kLo, kHi) or the sequence is
  static constexpr int kHi = 3000;      // re-run by the int32 kernel
  static constexpr bool kChecked = true;
};
template <> struct Store<int32_t> {
  static constexpr int kInfS = kInf;
  static constexpr int kThr = 5000000;
  static constexpr int kLo = -2000000000;
  static constexpr int kHi = 5000000;
  static constexpr bool kChecked = false;
};

__device__ __forceinline__ int tri(int n, int d) {
  const int k = d - 4;
  return k * (n - 4) - (k * (k - 1)) / 2;
}

// fML lives in a FOLDED diagonal-major layout with a fixed row pitch rp (a function of the launch's
// length class only, hence warp-uniform and known at kernel entry): row r holds diagonal r followed
// by diagonal rp+4-r, so the triangle still packs into ~rp*rp/2 cells, but stepping from one diagonal
// to the next is a CONSTANT stride (+rp, or 1-rp in the folded half).  The multibranch split scan
// min_t fML(i,i+t) + fML(i+t+1,j) walks both operands diagonal by diagonal, so with this layout its
// addresses are base + k*stride and the loop body is 2 LDS + 1 VIADDMNMX per candidate.
__host__ __device__ __forceinline__ int fold_rows(int rp) { return rp / 2 - 1; }
__device__ __forceinline__ int fold_idx(int rp, int t, int x) {
  return (t <= rp / 2 + 2) ? (t - 4) * rp + x : (rp - t) * rp + (t - 4) + x;
}

template <typename ST, bool UPA, bool UPB>
__device__ __forceinline__ void split_run(const ST *__restrict__ pa, const ST *__restrict__ pb, int cnt, int rp,
                                          int &acc0, int &acc1) {
  const int sA = UPA ? rp : 1 - rp, sB = UPB ? rp : 1 - rp;
  int k = 0;
#pragma unroll 1
  for (; k + 8 <= cnt; k += 8) {
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (u & 1) acc1 = __viaddmin_s32((int)pa[u * sA], (int)pb[u * sB], acc1);
      else acc0 = __viaddmin_s32((int)pa[u * sA], (int)pb[u * sB], acc0);
    }
    pa += 8 * sA;
    pb += 8 * sB;
  }
#pragma unroll 1
  for (; k < cnt; ++k) {
    acc0 = __viaddmin_s32((int)*pa, (int)*pb, acc0);
    pa += sA;
    pb += sB;
  }
}

// min over split points t in [t0, t1) of fML(diag t, x) + fML(diag d-1-t, x+t+1) for the lane's cell x
template <typename ST>
__device__ __forceinline__ int split_scan(const ST *__restrict__ F, int rp, int d, int x, int t0, int t1) {
  const int T0 = rp / 2 + 2;
  int acc0 = kInf, acc1 = kInf;
  // operand B = diagonal d-1-t sits in the folded half while t < d-1-T0; operand A (diagonal t) while t > T0
  const int tb = min(max(d - 1 - T0, t0), t1);  // [t0, tb): A plain, B folded
  const int ta = min(max(T0 + 1, tb), t1);      // [tb, ta): both plain; [ta, t1): A folded, B plain
  if (tb > t0) split_run<ST, true, true>(F + fold_idx(rp, t0, x), F + fold_idx(rp, d - 1 - t0, x + t0 + 1), tb - t0, rp, acc0, acc1);
  if (ta > tb) split_run<ST, true, false>(F + fold_idx(rp, tb, x), F + fold_idx(rp, d - 1 - tb, x + tb + 1), ta - tb, rp, acc0, acc1);
  if (t1 > ta) split_run<ST, false, false>(F + fold_idx(rp, ta, x), F + fold_idx(rp, d - 1 - ta, x + ta + 1), t1 - ta, rp, acc0, acc1);
  return min(acc0, acc1);
}

__device__ __forceinline__ int enc_base(unsigned char ch) {
  switch (ch) {
    case 'A': case 'a': return 1;
    case 'C': case 'c': return 2;
    case 'G': case 'g': return 3;
    case 'U': case 'u': case 'T': case 't': return 4;
    default: return 0;
  }
}

// loop-shape classes of the unrolled interior scan
enum { kSkip = 0, kGen = 1, k1n = 2, kBul = 3 };
__host__ __device__ constexpr int shape_class(int n1, int n2) {
  const int nl = n1 > n2 ? n1 : n2, ns = n1 > n2 ? n2 : n1;
  if (ns == 0) return nl >= 2 ? kBul : kSkip;          // stack, 1-bulge are tabulated specials
  if (ns == 1) return nl >= 3 ? k1n : kSkip;           // 1x1, 1x2 specials
  if (ns == 2) return nl >= 4 ? kGen : kSkip;          // 2x2, 2x3 specials
  return kGen;
}

__host__ __device__ constexpr int wave_np(int nmax) { return (nmax + 9) & ~1; }  // even row pitch of the scratch rows

struct WaveSmem {
  int f5_off, dml_off, bt_off, misc_off, ptab_off, S_off, db_off, part_off, mlp_off, C_off, F_off, wI_off, w1_off, pl_off, plo_off, gd_off, mlf_off, total;
  int rs;        // window row stride (elements)
  int mlparts;   // max multiloop parts per cell
  __host__ __device__ static WaveSmem make(int nmax, int esz) {
    WaveSmem L;
    int o = 0;
    auto take = [&](int bytes) { int r = o; o += (bytes + 15) / 16 * 16; return r; };
    const int np = wave_np(nmax);
    L.rs = (nmax + 7) / 8 * 8;
    L.mlparts = (nmax + kMlChunk - 1) / kMlChunk;
    L.f5_off = take(np * 4);
    L.dml_off = take(3 * np * 4);
    L.misc_off = take(64);
    L.ptab_off = take(64);
    L.S_off = take(np);
    L.db_off = take(np);
    L.part_off = take(2 * kWSlots * np * esz);      // [parity][slot][cell]; slot 0 generic, 1 tabulated, 2.. bulge/1xn parts
    L.bt_off = L.part_off;                          // the backtrack stack reuses the scratch rows after the fill
    L.mlp_off = take(2 * L.mlparts * np * esz);     // [parity][mlpart][cell]
    const int cells = table_cells(nmax) + 64;
    L.C_off = take(cells * esz);
    L.F_off = take((fold_rows(L.rs) * L.rs + 64) * esz);  // folded layout, pitch rs
    // short classes may also run in a windowed (wide-CTA) instance: room for whichever layout is larger
    const int wcells = nmax <= kFullWinMax ? (cells > kWin * L.rs + 64 ? cells : kWin * L.rs + 64) : kWin * L.rs + 64;
    L.wI_off = take(wcells * esz);
    L.w1_off = take((kWin * L.rs + 64) * esz);  // c1 is only read at the two ends of a row: always a window
    L.pl_off = take(nmax * nmax / 4 + 64);          // cells whose ends can pair, diagonal by diagonal (uint8 index)
    L.plo_off = take(np * 2);                       // first list entry of each diagonal (+ total at [n])
    L.gd_off = take(kGenSlots * 2);                 // (L, n1) of the generic shape in each lane slot
    L.mlf_off = take((table_cells(nmax) + 31) / 32 * 4);  // one bit per cell: "closed by a multiloop and by nothing else"
    L.total = o;
    return L;
  }
};


// ---------------------------------------------------------------------------------------------
// G(cell): the generic interior shapes (n1, n2 >= 2, not 2x2 / 2x3 / 3x2) of ONE pairable cell, lanes
// across SHAPES.  The 375 shapes with loop size 6..30 are numbered by (L, n1) ascending; lane l owns
// shapes l, l+32, ... (12 register slots: element offset into the cI table, its per-diagonal step and
// the shape's size + asymmetry term).  Offsets follow the quadratic recurrence of the diagonal-major
// layout (off += step, step -= 1 per diagonal), so a candidate costs one address add, one LDS and
// one VIADDMNMX with every lane busy, however few pairable cells a diagonal has.  Shapes with
// L > d-6 are not valid yet: because the numbering is L-major they are exactly the slots >= nvalid(d).
// ---------------------------------------------------------------------------------------------
constexpr int kGenShapes = 375;
constexpr int kGenIters = 12;
static_assert(kGenIters * 32 == kGenSlots, "slot bookkeeping");
__host__ __device__ constexpr int gen_before(int L) { return L <= 6 ? 0 : (L - 4) * (L - 3) / 2 - 3; }  // shapes with size < L
__device__ __forceinline__ int gen_valid(int d) {
  const int lmax = min(kMaxLoop, d - 6);
  return lmax < 6 ? 0 : (lmax == kMaxLoop ? kGenIters * 32 : gen_before(lmax + 1));
}

// shared-memory load at a 32-bit shared-space address (lane part + warp-uniform part: ptxas folds a uniform
// register into the LDS address, which the generic-pointer form does not get)
template <typename ST>
__device__ __forceinline__ int lds_sa(unsigned addr) {
  int v;
  if (sizeof(ST) == 2) asm volatile("ld.shared.s16 %0, [%1];" : "=r"(v) : "r"(addr));
  else asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

// E item body (bulges and 1xn loops of loop sizes l0..l1 for the lane's cell) on shared-space addresses: p1 / pB are
// the lane's addresses of element (row 0, its cell + 1) in the c1 window / the cB table; every row and column offset
// is warp-uniform, so each candidate is one LDS [lane + uniform] and one VIADDMNMX with a constant-bank term.
template <typename ST, bool BWIN>
__device__ __forceinline__ void e_item_sa(unsigned p1, unsigned pB, int d, int n, int rs, int l0, int l1, int &a1, int &aB) {
  constexpr int esz = (int)sizeof(ST);
  int off = tri(n, d - 2 - l0);      // first cell of diagonal d-2-L in the full triangle
  int step = n - d + 3 + l0;         // off(L+1) = off(L) - step(L)
  for (int L = l0; L <= l1; ++L) {
    const int offw = ((d - 2 - L) & (kWin - 1)) * rs;
    const int gb = cW.GB[L], g1 = cW.G1[L];
    const int offb = BWIN ? offw : off;
    aB = __viaddmin_s32(lds_sa<ST>(pB + (unsigned)(offb * esz)), gb, aB);
    aB = __viaddmin_s32(lds_sa<ST>(pB + (unsigned)((offb + L) * esz)), gb, aB);
    a1 = __viaddmin_s32(lds_sa<ST>(p1 + (unsigned)((offw + 1) * esz)), g1, a1);
    a1 = __viaddmin_s32(lds_sa<ST>(p1 + (unsigned)((offw + L - 1) * esz)), g1, a1);
    off -= step;
    ++step;
  }
}

template <typename ST, bool ALL>
__device__ __forceinline__ int gen_cell_sa(unsigned xb, const int (&goff)[kGenIters], const int (&g)[kGenIters],
                                           int nvalid, int lane) {
  // goff[] are shared-space byte ADDRESSES of the lane's slots for cell 0, xb the warp-uniform byte offset of the cell
  int a0 = Store<ST>::kInfS, a1 = Store<ST>::kInfS, a2 = Store<ST>::kInfS, a3 = Store<ST>::kInfS;
  if (ALL) {
#pragma unroll
    for (int k = 0; k < kGenIters; ++k) {
      const int v = lds_sa<ST>((unsigned)goff[k] + xb);
      if ((k & 3) == 0) a0 = __viaddmin_s32(v, g[k], a0);
      else if ((k & 3) == 1) a1 = __viaddmin_s32(v, g[k], a1);
      else if ((k & 3) == 2) a2 = __viaddmin_s32(v, g[k], a2);
      else a3 = __viaddmin_s32(v, g[k], a3);
    }
  } else {
    const int nv = nvalid - lane;
#pragma unroll
    for (int k = 0; k < kGenIters; ++k) {
      if (32 * k < nvalid) {           // warp-uniform early out
        if (32 * k < nv) {             // this lane's slot is a valid shape on this diagonal
          const int v = lds_sa<ST>((unsigned)goff[k] + xb);
          if (k & 1) a1 = __viaddmin_s32(v, g[k], a1);
          else a0 = __viaddmin_s32(v, g[k], a0);
        }
      }
    }
  }
  return __reduce_min_sync(0xffffffffu, min(min(a0, a1), min(a2, a3)));
}

template <typename ST>
__device__ __forceinline__ void gen_cell2_sa(unsigned xb0, unsigned xb1, const int (&goff)[kGenIters], const int (&g)[kGenIters],
                                             int &r0, int &r1) {
  int a0 = Store<ST>::kInfS, a1 = Store<ST>::kInfS, c0 = Store<ST>::kInfS, c1 = Store<ST>::kInfS;
#pragma unroll
  for (int k = 0; k < kGenIters; ++k) {
    const int v = lds_sa<ST>((unsigned)goff[k] + xb0);
    const int w = lds_sa<ST>((unsigned)goff[k] + xb1);
    if (k & 1) {
      a1 = __viaddmin_s32(v, g[k], a1);
      c1 = __viaddmin_s32(w, g[k], c1);
    } else {
      a0 = __viaddmin_s32(v, g[k], a0);
      c0 = __viaddmin_s32(w, g[k], c0);
    }
  }
  r0 = __reduce_min_sync(0xffffffffu, min(a0, a1));
  r1 = __reduce_min_sync(0xffffffffu, min(c0, c1));
}

template <typename ST>
__device__ __forceinline__ ST put(int v, int &flag) {
  if (v >= Store<ST>::kThr) return (ST)Store<ST>::kInfS;
  if (Store<ST>::kChecked && (v <= Store<ST>::kLo || v >= Store<ST>::kHi)) flag = 1;
  return (ST)v;
}

__device__ __forceinline__ int mlparts_of(int d, int n) {
  const int nsplit = d - 2 * kTurn - 2;  // t = 4 .. d-5
  return (d < n && nsplit > 0) ? (nsplit + kMlChunk - 1) / kMlChunk : 0;
}

// k / nb for small quotients (warp-uniform, no integer division)
__device__ __forceinline__ int div_small(int k, int nb, int qmax) {
  int q = 0;
  for (int t = 1; t < qmax; ++t) q += (k >= t * nb) ? 1 : 0;
  return q;
}

// ---------------------------------------------------------------------------------------------
// Backtrack in ViennaRNA's deterministic order (SURVEY.md 8a' F7), run by ONE warp: the sector
// stack is warp-uniform and the lanes probe the candidates of each decision in ViennaRNA's scan
// order, the first hit (ballot + ffs) wins.  Tables may live in shared or global memory.
// Returns 0 or a failure code (no decomposition found -- would mean a fill error).
// ---------------------------------------------------------------------------------------------
template <typename ST, bool FOLDED, bool CP = false>
__device__ int backtrack_warp(int n, const unsigned char *S, const int *f5, const ST *CB, const ST *F, int rp, int *bt,
                              unsigned char *db, const FoldParams *__restrict__ gp, int lane,
                              const unsigned *mlf = nullptr) {
  const SmallParams &sp = small_params<CP>(gp);
  auto fx = [&](int t, int x) -> int { return FOLDED ? fold_idx(rp, t, x) : tri(n, t) + x; };
  const int tau = cW.TerminalAU;
  auto craw = [&](int cb, int type) -> int { return cb - (type > 2 ? tau : 0); };
  const unsigned full = 0xffffffffu;
  int top = 0, failc = 0;
  if (n > 0) {
    if (lane == 0) bt[0] = (1 << 14) | (n << 2) | 0;
    top = 1;
  }
  __syncwarp();
  while (top > 0 && !failc) {
    const int w = bt[--top];
    __syncwarp();
    int i = w >> 14, j = (w >> 2) & 0xfff;
    const int ml = w & 3;
    if (ml == 0) {
      // largest jj <= j with f5[jj] != f5[jj-1]
      int jj = 0;
      for (int base = j; base >= 1 && !jj; base -= 32) {
        const int k = base - lane;
        const unsigned m = __ballot_sync(full, k >= 1 && f5[k] != f5[k - 1]);
        if (m) jj = base - (__ffs(m) - 1);
      }
      if (jj < kTurn + 2) continue;
      const int fij = f5[jj];
      const int s3 = (jj < n) ? (int)S[jj + 1] : -1;
      int u = 0;
      for (int base = jj - kTurn - 1; base >= 1 && !u; base -= 32) {
        const int k = base - lane;
        bool hit = false;
        if (k >= 1) {
          const int type = pair_type(S[k], S[jj]);
          if (type) {
            const int e = craw((int)CB[tri(n, jj - k) + k - 1], type) + e_extloop(sp, type, k > 1 ? (int)S[k - 1] : -1, s3);
            hit = (fij == e + f5[k - 1]);
          }
        }
        const unsigned m = __ballot_sync(full, hit);
        if (m) u = base - (__ffs(m) - 1);
      }
      if (!u) { failc = 1; break; }
      if (u - 1 > 0) {
        if (lane == 0) bt[top] = (1 << 14) | ((u - 1) << 2) | 0;
        top++;
      }
      i = u;
      j = jj;
    } else {
      // fML sector: strip 3' bases, then 5' bases, then stem or split
      int ii = i, jj = 0;
      for (int base = j; base >= ii + kTurn + 1 && !jj; base -= 32) {
        const int k = base - lane;
        bool stop = false;
        if (k >= ii + kTurn + 1) {
          const int fij = (int)F[fx(k - ii, ii - 1)];
          const int fi = (k - 1 - ii > kTurn) ? (int)F[fx(k - 1 - ii, ii - 1)] + cW.MLbase : kInf;
          stop = (fij != fi);
        }
        const unsigned m = __ballot_sync(full, stop);
        if (m) jj = base - (__ffs(m) - 1);
      }
      if (!jj) { failc = 2; break; }
      int i2 = 0;
      for (int base = ii; base <= jj - kTurn - 1 && !i2; base += 32) {
        const int k = base + lane;
        bool stop = false;
        if (k <= jj - kTurn - 1) {
          const int fij = (int)F[fx(jj - k, k - 1)];
          const int fi = (jj - k - 1 > kTurn) ? (int)F[fx(jj - k - 1, k)] + cW.MLbase : kInf;
          stop = (fij != fi);
        }
        const unsigned m = __ballot_sync(full, stop);
        if (m) i2 = base + (__ffs(m) - 1);
      }
      if (!i2) { failc = 2; break; }
      ii = i2;
      const int fij = (int)F[fx(jj - ii, ii - 1)];
      const int type = pair_type(S[ii], S[jj]);
      if (type && fij == craw((int)CB[tri(n, jj - ii) + ii - 1], type) + e_mlstem(sp, type, S[ii - 1], S[jj + 1])) {
        i = ii;
        j = jj;
      } else {
        int u = 0;
        for (int base = ii + 1 + kTurn; base <= jj - 2 - kTurn && !u; base += 32) {
          const int k = base + lane;
          bool hit = false;
          if (k <= jj - 2 - kTurn)
            hit = (fij == (int)F[fx(k - ii, ii - 1)] + (int)F[fx(jj - k - 1, k)]);
          const unsigned m = __ballot_sync(full, hit);
          if (m) u = base + (__ffs(m) - 1);
        }
        if (!u) { failc = 3; break; }
        if (lane == 0) {
          bt[top] = (ii << 14) | (u << 2) | 1;
          bt[top + 1] = ((u + 1) << 14) | (jj << 2) | 1;
        }
        top += 2;
        __syncwarp();
        continue;
      }
    }
    // (i,j) pairs: walk down the helix / interior loops until a hairpin or a multiloop
    for (;;) {
      if (lane == 0) {
        db[i - 1] = '(';
        db[j - 1] = ')';
      }
      const int d = j - i;
      const int type = pair_type(S[i], S[j]);
      const int cij = craw((int)CB[tri(n, d) + i - 1], type);
      // hairpin first (ViennaRNA's order), then the stacked pair -- by far the most frequent step -- as a
      // lane-uniform check whose lookups are issued together with the hairpin's
      const int si1 = S[i + 1], sj1 = S[j - 1];
      const int hp = e_hairpin(sp, gp, S, i, j, type);
      const int t2s = (d >= 6) ? pair_type(si1, sj1) : 0;
      const int st = craw((int)CB[tri(n, max(d - 2, kTurn + 1)) + i], t2s) + sp.stack[type][rtype_of(t2s)];
      if (cij == hp) break;
      if (t2s && cij == st) {
        ++i;
        --j;
        continue;
      }
      bool found = false;
      const int cidx = tri(n, d) + i - 1;
      const bool ml_only = mlf != nullptr && ((mlf[cidx >> 5] >> (cidx & 31)) & 1u);  // recorded by the fill
      const int amax = ml_only ? -1 : min(kMaxLoop, d - 6);
      for (int n1 = 0; n1 <= amax && !found; ++n1) {
        const int p = i + 1 + n1;
        const int bmax = min(kMaxLoop - n1, d - 6 - n1);
        const int n2 = lane, q = j - 1 - n2;
        bool hit = false;
        if (n2 <= bmax) {
          const int t2 = pair_type(S[p], S[q]);
          if (t2) {
            const int e = craw((int)CB[tri(n, q - p) + p - 1], t2) +
                          e_intloop(sp, gp, n1, n2, type, rtype_of(t2), si1, sj1, S[p - 1], S[q + 1]);
            hit = (e == cij);
          }
        }
        const unsigned m = __ballot_sync(full, hit);
        if (m) {
          i = p;
          j = j - 1 - (__ffs(m) - 1);
          found = true;
        }
      }
      if (found) continue;
      const int e = cij - e_mlstem(sp, rtype_of(type), sj1, si1) - cW.MLclosing;
      const int p = i + 1, q = j - 1;
      int u = 0;
      for (int base = p + kTurn + 1; base < q - kTurn - 1 && !u; base += 32) {
        const int k = base + lane;
        bool hit = false;
        if (k < q - kTurn - 1) hit = (e == (int)F[fx(k - p, p - 1)] + (int)F[fx(q - k - 1, k)]);
        const unsigned m = __ballot_sync(full, hit);
        if (m) u = base + (__ffs(m) - 1);
      }
      if (!u) { failc = 4; break; }
      if (lane == 0) {
        bt[top] = (p << 14) | (u << 2) | 1;
        bt[top + 1] = ((u + 1) << 14) | (q << 2) | 1;
      }
      top += 2;
      break;
    }
    __syncwarp();
  }
  return failc;
}

// ---------------------------------------------------------------------------------------------
// the kernel
// ---------------------------------------------------------------------------------------------
// Register budgets: 96 registers keep the 5 CTAs/SM that shared memory allows at n = 100 (128-thread
// instance); 128 keep 2 CTAs of 256 threads (n <= 160) and the single 512-thread CTA.  __maxnreg__ is used
// instead of __launch_bounds__' minBlocks because the latter changes ptxas' unrolling heuristics as well
// (measured: +20 % instructions, instruction-cache misses).
#ifndef WAVE_MAXREG128
#define WAVE_MAXREG128 96
#endif
#ifndef WAVE_T128
#define WAVE_T128 128  // threads per CTA of the short-sequence instance (n <= 128)
#endif
constexpr int kSmallThreads = WAVE_T128;
// RS > 0: the row pitch of the folded fML table and of the windows is this compile-time constant (the launch's
// length class must round up to it): the strides of the split scan become LDS immediates and the row offsets
// shifts / small multiplies.  Instantiated for the pitches of the most common length classes; RS = 0 is generic.
template <typename ST, int MAXT, int RS = 0>
__global__ void __launch_bounds__(MAXT) __maxnreg__(MAXT == kSmallThreads && sizeof(ST) == 2 ? WAVE_MAXREG128 : 128) fold_wave_kernel(const FoldParams *__restrict__ gp, BatchArgs a, int nmax,
                                                        int *__restrict__ queue) {
  extern __shared__ __align__(16) unsigned char smem[];
  if (a.dyn && a.dyn[0] == 0) return;  // device-counted launch with nothing routed to it
  const WaveSmem Ls = WaveSmem::make(nmax, (int)sizeof(ST));
  int *f5 = reinterpret_cast<int *>(smem + Ls.f5_off);
  int *dml = reinterpret_cast<int *>(smem + Ls.dml_off);
  int *bt = reinterpret_cast<int *>(smem + Ls.bt_off);
  int *misc = reinterpret_cast<int *>(smem + Ls.misc_off);
  unsigned char *S = smem + Ls.S_off;
  unsigned char *db = smem + Ls.db_off;
  ST *part = reinterpret_cast<ST *>(smem + Ls.part_off);
  ST *mlp = reinterpret_cast<ST *>(smem + Ls.mlp_off);
  ST *CB = reinterpret_cast<ST *>(smem + Ls.C_off);
  ST *F = reinterpret_cast<ST *>(smem + Ls.F_off);
  ST *wI = reinterpret_cast<ST *>(smem + Ls.wI_off);
  ST *w1 = reinterpret_cast<ST *>(smem + Ls.w1_off);
  unsigned char *ptab = smem + Ls.ptab_off;
  unsigned char *pl = smem + Ls.pl_off;
  unsigned short *plo = reinterpret_cast<unsigned short *>(smem + Ls.plo_off);
  unsigned *mlf = reinterpret_cast<unsigned *>(smem + Ls.mlf_off);
  const int np = wave_np(nmax), rs = RS > 0 ? RS : Ls.rs, mlparts = Ls.mlparts;
  const int tid = threadIdx.x, lane = tid & 31;
  constexpr int NW = MAXT / 32, NT = MAXT;  // the kernel is always launched with MAXT threads
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // tells the compiler it is warp-uniform
  const SmallParams &sp = gp->s;
  constexpr int kInfS = Store<ST>::kInfS;
  constexpr bool kFullW = (MAXT == kSmallThreads);  // wave_threads(): the short-sequence instance <=> nmax <= kFullWinMax
  const int tau = cW.TerminalAU;

  if (tid < 64) ptab[tid] = (unsigned char)(((tid >> 3) < 5 && (tid & 7) < 5) ? pair_type(tid >> 3, tid & 7) : 0);
  // generic interior shapes, numbered by (L, n1): slot s -> L | n1 << 8 (padding slots repeat the last shape)
  unsigned short *gdesc = reinterpret_cast<unsigned short *>(smem + Ls.gd_off);
  for (int sidx = tid; sidx < kGenSlots; sidx += NT) {
    const int sc = min(sidx, kGenShapes - 1);
    int L = 6;
    while (gen_before(L + 1) <= sc) ++L;
    gdesc[sidx] = (unsigned short)(L | ((2 + sc - gen_before(L)) << 8));
  }
  __syncthreads();
  auto PT = [&](int x, int y) -> int { return ptab[x * 8 + y]; };
  auto craw = [&](int cb, int type) -> int { return cb - (type > 2 ? tau : 0); };
#if WAVE_CMM
  auto MMI = [&](int t, int a, int b) -> int { return cW.mmI[t][a][b]; };
  auto MM1 = [&](int t, int a, int b) -> int { return cW.mm1nI[t][a][b]; };
  auto MLSTEM = [&](int t, int s5, int s3) -> int { return cW.mmM[t][s5][s3] + (t > 2 ? tau : 0) + cW.MLintern; };
#else
  auto MMI = [&](int t, int a, int b) -> int { return __ldg(&sp.mmI[t][a][b]); };
  auto MM1 = [&](int t, int a, int b) -> int { return __ldg(&sp.mm1nI[t][a][b]); };
  auto MLSTEM = [&](int t, int s5, int s3) -> int { return e_mlstem(sp, t, s5, s3); };
#endif

  for (;;) {
    __syncthreads();
    if (tid == 0) {
      misc[0] = atomicAdd(queue, 1);
      misc[1] = 0;  // backtrack failure code
      misc[2] = 0;  // Hamming accumulator
      misc[3] = 0;  // range-check flag
    }
    __syncthreads();
    const int slot = misc[0];
    if (slot >= batch_count(a)) break;
    PHASE_CLK_BEGIN();
    const int b = batch_entry(a, slot);
    const int beg = a.off[b];
    const int n = a.off[b + 1] - beg;
    int flag = 0;

    for (int k = tid; k < n; k += NT) {
      S[k + 1] = (unsigned char)enc_base(a.seq[beg + k]);
      db[k] = '.';
    }
    for (int k = tid; k <= n; k += NT) f5[k] = 0;
    for (int k = tid; k < (tri(n, n) + 31) / 32; k += NT) mlf[k] = 0;
    __syncthreads();
    if (tid == 0 && n > 0) {
      S[0] = S[n];
      S[n + 1] = S[1];
    }
    __syncthreads();

    // ---- lists of the cells whose ends can pair, one list per diagonal (at most n*n/4 entries in all):
    // the interior-loop and tabulated-shape work items run over these lists, not over all cells.
    // Bit masks of the positions holding each base (one ballot per 32 positions), then one THREAD per
    // diagonal: pairable(i, i+d) for 32 values of i is six AND/ORs of mask words against funnel-shifted
    // mask words, so counting and listing a diagonal is a few dozen instructions, not a pass over its cells.
    const int mstride = (n >> 5) + 3;
    unsigned *bm = reinterpret_cast<unsigned *>(dml);  // [4][mstride], dml is idle until the first phases
    for (int w = warp; w < mstride; w += NW) {
      const int pp = w * 32 + lane;
      const int code = (pp >= 1 && pp <= n) ? (int)S[pp] : 0;
#pragma unroll
      for (int c = 1; c <= 4; ++c) {
        const unsigned m = __ballot_sync(0xffffffffu, code == c);
        if (lane == 0) bm[(c - 1) * mstride + w] = m;
      }
    }
    __syncthreads();
    auto dword = [&](int d, int w) -> unsigned {  // bit b: positions i = 32w+b and i+d can pair, 1 <= i <= n-d
      const int sh = d & 31, wo = w + (d >> 5);
      const unsigned *mA = bm, *mC = bm + mstride, *mG = bm + 2 * mstride, *mU = bm + 3 * mstride;
      const unsigned sA = __funnelshift_r(mA[wo], mA[wo + 1], sh), sC = __funnelshift_r(mC[wo], mC[wo + 1], sh);
      const unsigned sG = __funnelshift_r(mG[wo], mG[wo + 1], sh), sU = __funnelshift_r(mU[wo], mU[wo + 1], sh);
      unsigned m = (mA[w] & sU) | (mC[w] & sG) | (mG[w] & (sC | sU)) | (mU[w] & (sA | sG));
      const int hi = n - d - 32 * w;  // highest valid bit
      if (hi < 31) m &= (hi < 0) ? 0u : ((2u << hi) - 1u);
      if (w == 0) m &= ~1u;
      return m;
    };
    for (int d = kTurn + 1 + tid; d < n; d += NT) {
      int cnt = 0;
      for (int w = 0; w <= (n - d) >> 5; ++w) cnt += __popc(dword(d, w));
      plo[d] = (unsigned short)cnt;
    }
    __syncthreads();
    if (warp == 0) {
      int carry = 0;
      for (int base = kTurn + 1; base < n; base += 32) {
        const int d = base + lane;
        const int v = d < n ? (int)plo[d] : 0;
        int inc = v;
        for (int o = 1; o < 32; o <<= 1) {
          const int t = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += t;
        }
        if (d < n) plo[d] = (unsigned short)(carry + inc - v);
        carry += __shfl_sync(0xffffffffu, inc, 31);
      }
      if (lane == 0 && n > kTurn + 1) plo[n] = (unsigned short)carry;
    }
    __syncthreads();
    for (int d = kTurn + 1 + tid; d < n; d += NT) {
      int pos = plo[d];
      for (int w = 0; w <= (n - d) >> 5; ++w) {
        unsigned m = dword(d, w);
        while (m) {
          pl[pos++] = (unsigned char)(32 * w + __ffs(m) - 2);  // cell index x = i - 1
          m &= m - 1;
        }
      }
    }
    __syncthreads();

    PHASE_CLK_END(600);  // prologue
    // generic-shape slots at d = 4: element offset of (row d-2-L, column 1+n1), its per-diagonal step and
    // the shape's size + asymmetry term (nothing of this is live across the backtrack)
    int gterm[kGenIters], goff[kGenIters];
    const unsigned wI_sa = (unsigned)__cvta_generic_to_shared(wI);
    const unsigned w1_sa = (unsigned)__cvta_generic_to_shared(w1), CB_sa = (unsigned)__cvta_generic_to_shared(CB);
    unsigned gpack[kGenIters / 4] = {};  // one byte per slot: L * esz (full tables) or (1 + n1) * esz (windows)
#pragma unroll
    for (int k = 0; k < kGenIters; ++k) {
      const int ds = gdesc[32 * k + lane];
      const int L = ds & 255, n1 = ds >> 8;
      gterm[k] = (32 * k + lane < kGenShapes) ? cW.GI[L][n1] : kBig;
      if (kFullW) {
        goff[k] = (int)wI_sa + (tri(n, kTurn + 1 - 2 - L) + 1 + n1) * (int)sizeof(ST);  // shared-space addresses
        gpack[k >> 2] |= (unsigned)(L * (int)sizeof(ST)) << (8 * (k & 3));
      } else {
        goff[k] = (int)wI_sa + (((kTurn + 1 - 2 - L) & (kWin - 1)) * rs + 1 + n1) * (int)sizeof(ST);
        gpack[k >> 2] |= (unsigned)((1 + n1) * (int)sizeof(ST)) << (8 * (k & 3));
      }
    }

    // carried across phases instead of being re-derived from d by every warp in every phase: the first cells of
    // the six rows the tabulated shapes read (diagonals max(d-2-k, 4), k = 0..5) and d mod 3 for the dml rows
    int trow[6] = {0, 0, 0, 0, 0, 0};
    int dm3 = (kTurn + 1) % 3;
    // ... and what the finalisation of diagonal d1 = d - 1 indexes: first cell of d1 in the triangle (tri1) and in
    // the folded fML layout (fi1), and the latter for d1 - 1 (fi0); the folded layout turns around after row T0
    int tri1 = 0, fi1 = 0, fi0 = 0;
    const int foldT0 = rs / 2 + 2, fold_jump = fold_idx(rs, foldT0 + 1, 0) - fold_idx(rs, foldT0, 0);
#ifdef WAVE_TRACE
    __shared__ long long wtrace_base;
    if (tid == 0) wtrace_base = clock64();
    __syncthreads();
    const long long wtrace_t0 = wtrace_base;
    const bool wtracing = blockIdx.x == 0 && slot < (int)gridDim.x;
#endif
    for (int d = kTurn + 1; d <= n + 1; ++d) {
      WTRACE(0);
      const int par = d & 1, par1 = par ^ 1;
      const int cells_d = (d < n) ? n - d : 0;
      const int nb_d = (cells_d + 31) >> 5;
      const int lmax = min(kMaxLoop, d - 6);
      const int mpd = mlparts_of(d, n);
      const int d1 = d - 1;
      const int cells_1 = (d1 >= kTurn + 1 && d1 < n) ? n - d1 : 0;
      const int nb_1 = (cells_1 + 31) >> 5;
      const int has_x = (d1 >= kTurn + 2 && d1 <= n) ? 1 : 0;
      const int pl0 = (d < n) ? (int)plo[d] : 0;
      const int npair = (d < n) ? (int)plo[d + 1] - pl0 : 0;
      const int nbI = (npair + 31) >> 5;
      // ---------------- G(cell): generic interior shapes, lanes across shapes, one pairable cell at a time
      const int nvalid = (d < n && !(WAVE_SKIP & 1)) ? gen_valid(d) : 0;
      ITEM_CLK_BEGIN();
      if (nvalid == kGenSlots) {
#pragma unroll 1
        for (int c = warp; c < npair; c += NW) {
          const int xv = pl[pl0 + c];
          // the cell is the same for every lane: through REDUX its byte offset lands in a uniform register and
          // the twelve table loads address [lane offset + uniform base] without an add each
          const int xb = (int)__reduce_max_sync(0xffffffffu, (unsigned)xv * (unsigned)sizeof(ST));
          const int r = gen_cell_sa<ST, true>((unsigned)xb, goff, gterm, nvalid, lane);
          if (lane == 0) part[(par * kWSlots + 0) * np + xv] = (ST)min(r, kInfS);
        }
      } else if (nvalid > 0) {
#pragma unroll 1
        for (int c = warp; c < npair; c += NW) {
          const int xv = pl[pl0 + c];
          const int xb = (int)__reduce_max_sync(0xffffffffu, (unsigned)xv * (unsigned)sizeof(ST));
          const int r = gen_cell_sa<ST, false>((unsigned)xb, goff, gterm, nvalid, lane);
          if (lane == 0) part[(par * kWSlots + 0) * np + xv] = (ST)min(r, kInfS);
        }
      }
      ITEM_CLK_END(5, d >= 68 ? 1 : 0);
      WTRACE(1);
      WTRACE_VAL(4, 0);
      // the block items start with the warp after the last one that took a cell, to even out the phase
      const int nG = nvalid > 0 ? npair : 0;
      int neparts = 0;
#pragma unroll
      for (int r = 0; r < kEParts; ++r) neparts += (epart_beg(r) <= lmax) ? 1 : 0;
      const int nI = (WAVE_SKIP & 2) ? 0 : nbI * neparts, nT = (lmax >= 0 && !(WAVE_SKIP & 4)) ? nbI : 0, nM = (WAVE_SKIP & 8) ? 0 : nb_d * mpd, nF = (WAVE_SKIP & 16) ? 0 : nb_1;
      const int total = nI + nT + nM + nF + ((WAVE_SKIP & 32) ? 0 : has_x);
      int k0 = warp - (nG % NW);
      if (k0 < 0) k0 += NW;
      for (int k = k0; k < total; k += NW) {
        ITEM_CLK_BEGIN();
#ifdef WAVE_PHASE_CLK
        const int item_kind = k < nI ? 0 : k < nI + nT ? 1 : k < nI + nT + nM ? 2 : k < nI + nT + nM + nF ? 3 : 4;
#endif
        if (k < nI) {
          // ---------------- E(b,r): bulges and 1xn loops of the loop sizes in part r, lanes across pairable cells
          const int r = div_small(k, nbI, kEParts), blk = k - r * nbI;
          const int l0 = r == 0 ? epart_beg(0) : r == 1 ? epart_beg(1) : epart_beg(2);
          const int l1 = min((r == 0 ? epart_beg(1) : r == 1 ? epart_beg(2) : epart_beg(3)) - 1, lmax);
          const int idx = blk * 32 + lane;
          const int x = pl[pl0 + min(idx, npair - 1)];
          // the outer pair's term is fetched before the scan, not after it (its latency hides behind the scan)
          const int oi = x + 1, oj = oi + d;
          const int type = PT(S[oi], S[oj]);
          const int v_mm1 = MM1(type, S[oi + 1], S[oj - 1]);
          int a1 = kInfS, aB = kInfS;
          e_item_sa<ST, false>(w1_sa + (unsigned)((x + 1) * (int)sizeof(ST)), CB_sa + (unsigned)((x + 1) * (int)sizeof(ST)), d, n,
                               rs, l0, l1, a1, aB);
          if (idx < npair) {
            const int e = min(a1 + v_mm1, aB + (type > 2 ? tau : 0));
            part[(par * kWSlots + 2 + r) * np + x] = (ST)min(e, kInfS);
          }
        } else if (k < nI + nT) {
          // ---------------- T(b): the 9 tabulated shapes of diagonal d
          const int idx = (k - nI) * 32 + lane;
          if (idx < npair) {
            const int x = pl[pl0 + idx];
            const int i = x + 1, j = i + d;
            const int type = PT(S[i], S[j]);
            int best = kInfS;
            {
              int si[5], sj[5];
#pragma unroll
              for (int u = 0; u < 5; ++u) {
                si[u] = S[i + u];
                sj[u] = S[j - u];
              }
              // branch-free: every shape's table lookups are issued unconditionally (safe indices), invalid
              // shapes are masked at the end -- the nine dependent load chains overlap instead of serialising
#pragma unroll
              for (int s = 0; s < 9; ++s) {
                constexpr int kN1[9] = {0, 0, 1, 1, 1, 2, 2, 2, 3};
                constexpr int kN2[9] = {0, 1, 0, 1, 2, 1, 2, 3, 2};
                const int n1 = kN1[s], n2 = kN2[s];
                const int t2 = PT(si[n1 + 1], sj[n2 + 1]);
                const int cpq = craw((int)CB[trow[n1 + n2] + x + 1 + n1], t2);
                const int e = cpq + e_intloop(sp, gp, n1, n2, type, rtype_of(t2), si[1], sj[1], si[n1], sj[n2]);
                best = (n1 + n2 <= d - 6 && t2 != 0) ? min(best, e) : best;
              }
            }
            part[(par * kWSlots + 1) * np + x] = (ST)min(best, kInfS);
          }
        } else if (k < nI + nT + nM) {
          // ---------------- M(b,m): multibranch split points t0..t1-1 of diagonal d
          const int kk = k - nI - nT;
          const int m = (mpd == 1) ? 0 : div_small(kk, nb_d, kMaxMlParts), blk = kk - m * nb_d;
          const int x = blk * 32 + lane, xc = min(x, cells_d - 1);
          const int t0 = kTurn + 1 + m * kMlChunk, t1 = min(d - kTurn - 1, t0 + kMlChunk);
          const int best = split_scan<ST>(F, rs, d, xc, t0, t1);
          if (x < cells_d) mlp[(par * mlparts + m) * np + x] = put<ST>(best, flag);
        } else if (k < nI + nT + nM + nF) {
          // ---------------- F(b): finalise diagonal d1
          // Every table lookup is issued up front and unconditionally (safe indices; unpairable cells and
          // lanes beyond the diagonal are masked afterwards): the parameter loads of a cell are independent of
          // each other, but behind branches they would be fetched one L2 round trip after the other.
          const int blk = k - nI - nT - nM;
          const int x = blk * 32 + lane;
          const bool live = x < cells_1;
          const int xs = live ? x : 0;
          const int i = xs + 1, j = i + d1;
          const int type = PT(S[i], S[j]);
          const int si1 = S[i + 1], sj1 = S[j - 1], sim1 = S[i - 1], sjp1 = S[j + 1];
          const int rt = rtype_of(type);
          const int v_mmh = __ldg(&sp.mmH[type][si1][sj1]);
          const int v_mmio = MMI(type, si1, sj1);
          const int v_mlc = MLSTEM(rt, sj1, si1);
          const int v_mmii = MMI(rt, sjp1, sim1);
          const int v_mm1i = MM1(rt, sjp1, sim1);
          const int v_mls = MLSTEM(type, sim1, sjp1);
          const int hbase = (d1 - 1 <= 30) ? sp.hairpin[d1 - 1] : __ldg(&gp->hairpin_len[d1 - 1]);  // warp-uniform
          const ST *src = part + (par1 * kWSlots) * np + xs;
          const int s_g = (int)src[0], s_t = (int)src[np];
          int s_e = kInfS;
#pragma unroll
          for (int r = 0; r < kEParts; ++r) s_e = (d1 - 6 >= epart_beg(r)) ? min(s_e, (int)src[(2 + r) * np]) : s_e;
          const int v_dml = dml[dm3 * np + xs + 1];  // row (d1 - 2) mod 3 = d mod 3
          const int pb = fi0 + xs;
          const int f_a = (int)F[pb + 1], f_b = (int)F[pb];
          int dm = kInf;
          const int mp1 = mlparts_of(d1, n);
          if constexpr (kFullW) {  // n <= 128: one multiloop part per diagonal
            static_assert(kMlChunk >= kFullWinMax, "the short instance assumes a single multiloop part");
            dm = mp1 > 0 ? (int)mlp[par1 * mlparts * np + xs] : kInf;
          } else {
            for (int m = 0; m < mp1; ++m) dm = min(dm, (int)mlp[(par1 * mlparts + m) * np + xs]);
          }
          if (dm >= Store<ST>::kThr) dm = kInf;

          int cij = hbase + v_mmh;
          if (d1 == 4 || d1 == 5 || d1 == 7) cij = type ? e_hairpin(sp, gp, S, i, j, type) : cij;  // sizes with special loops
          cij = (d1 >= 12) ? min(cij, s_g + v_mmio) : cij;  // generic shapes (L >= 6)
          cij = (d1 >= 6) ? min(cij, s_t) : cij;            // tabulated shapes
          cij = (d1 >= 8) ? min(cij, s_e) : cij;            // bulges, 1xn
          const int eml = v_dml + v_mlc + cW.MLclosing;
          // strictly better than the hairpin and every interior loop: the backtrack may skip its search for an
          // interior loop at this cell (the long search that would end empty-handed)
          const bool ml_only = d1 >= 2 * kTurn + 4 && eml < cij;
          cij = ml_only ? eml : cij;
          int e = (d1 > kTurn + 1) ? min(f_a, f_b) + cW.MLbase : kInf;
          e = type ? min(e, cij + v_mls) : e;
          if (live) {
            const int wrow1 = (d1 & (kWin - 1)) * rs + x;
            const int wrow = kFullW ? tri1 + x : wrow1;
            const int cidx = tri1 + x;
            if (type) {
              if (ml_only) atomicOr(&mlf[cidx >> 5], 1u << (cidx & 31));
              CB[cidx] = put<ST>(cij + (type > 2 ? tau : 0), flag);
              wI[wrow] = put<ST>(cij + v_mmii, flag);
              w1[wrow1] = put<ST>(cij + v_mm1i, flag);
            } else {
              CB[cidx] = (ST)kInfS;
              wI[wrow] = (ST)kInfS;
              w1[wrow1] = (ST)kInfS;
            }
            dml[(dm3 == 0 ? 2 : dm3 - 1) * np + x] = dm;  // row d1 mod 3
            F[fi1 + x] = put<ST>(min(e, dm), flag);
          }
        } else {
          // ---------------- X: f5[d1]
          const int j = d1;
          const int s3 = (j < n) ? (int)S[j + 1] : -1;
          int best = kInf;
          for (int i = j - kTurn - 1 - lane; i >= 1; i -= 32) {
            const int type = PT(S[i], S[j]);
            if (!type) continue;
            // cB already carries TerminalAU(pair), which is also E_ExtLoop's first term: the stored value plus the
            // mismatch / dangle part is c + E_ExtLoop
            best = min(best, f5[i - 1] + (int)CB[tri(n, j - i) + i - 1] + e_extloop_mm(sp, type, i > 1 ? (int)S[i - 1] : -1, s3));
          }
          for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
          if (lane == 0) f5[j] = min(f5[j - 1], best);
        }
        ITEM_CLK_END(item_kind, d >= 68 ? 1 : 0);
        WTRACE_VAL(4, 1 + (k < nI ? 0 : k < nI + nT ? 1 : k < nI + nT + nM ? 2 : k < nI + nT + nM + nF ? 3 : 4));   // kind of the last item
      }
      WTRACE(2);
      // next diagonal: the tabulated shapes' rows shift by one (row r+1 starts n-r cells after row r)
#pragma unroll
      for (int k = 5; k >= 1; --k) trow[k] = trow[k - 1];
      if (d - 2 >= kTurn + 1) trow[0] += n - (d - 2);
      dm3 = (dm3 == 2) ? 0 : dm3 + 1;
      if (d1 >= kTurn + 1) {
        tri1 += n - d1;
        fi0 = fi1;
        fi1 += (d1 < foldT0) ? rs : (d1 == foldT0 ? fold_jump : 1 - rs);
      }
      // ... and every generic-shape slot moves one row down its table (row d-2-L has n-d+2+L cells)
      {
        const int ustep = (n - d + 2) * (int)sizeof(ST);
#pragma unroll
        for (int k = 0; k < kGenIters; ++k) {
          const int b = (int)((gpack[k >> 2] >> (8 * (k & 3))) & 0xffu);
          if (kFullW) {
            goff[k] += ustep + b;
          } else {
            const int t = goff[k] + rs * (int)sizeof(ST), u = t - kWin * rs * (int)sizeof(ST);
            goff[k] = (u >= b + (int)wI_sa) ? u : t;
          }
        }
      }
      __syncthreads();
      WTRACE(3);
      PHASE_CLK_END(d);
    }

    if (flag) misc[3] = 1;
    __syncthreads();

    // ---- backtrack (ViennaRNA's deterministic order); warp 0, lanes scan candidates in that order
    if (warp == 0 && !misc[3]) {
      const int failc = backtrack_warp<ST, true, (MAXT == 512)>(n, S, f5, CB, F, rs, bt, db, gp, lane, mlf);
      if (lane == 0 && failc) misc[1] = failc;
    }
    __syncthreads();

    PHASE_CLK_END(601);  // backtrack
    // ---- emit structure, energy, Hamming distance, reward
    int local_hd = 0;
    for (int k = tid; k < n; k += NT) {
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
      if (lane == 0 && local_hd) atomicAdd(&misc[2], local_hd);
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
      if (Store<ST>::kChecked && a.redo) a.redo[b] = misc[3];
      if (misc[1] && !misc[3]) atomicExch(a.status, misc[1]);
    }
    PHASE_CLK_END(602);  // emit
  }
}

#include "fold_cluster.inl"

}  // namespace

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
size_t wave_smem_bytes(int nmax, int esz) { return (size_t)WaveSmem::make(nmax, esz).total; }

// int16 tables hold values checked to lie in (-25000, 3000) or the stored infinity 30000; candidates are summed in
// 32-bit registers, but the partial minima of the interior items go back into int16 rows: the loop-size terms of the
// parameter set must keep table value + term inside int16, otherwise the int32 instances are used.
bool wave_int16_ok(const HostParams &hp) {
  const SmallParams &s = hp.dev.s;
  int lo = 0, hi = 0;
  for (int L = 2; L <= 30; ++L) {
    lo = std::min(lo, std::min(s.bulge[L], s.interior[L]));
    hi = std::max(hi, std::max(s.bulge[L], s.interior[L] + std::max(0, s.max_ninio)));
  }
  return hi < 2700 && lo > -7000 && s.ninio >= 0;
}

int wave_threads(int nmax) { return nmax <= kFullWinMax ? kSmallThreads : (nmax <= 224 ? 256 : 512); }

cudaError_t wave_upload_constants(const HostParams &hp, cudaStream_t st) {
  static WaveConst w;  // staging must outlive the async copy
  const SmallParams &s = hp.dev.s;
  for (int L = 0; L < 32; ++L) {
    const int lc = L <= 30 ? L : 30;
    auto asym = [&](int k) { const int v = (k < 0 ? -k : k) * s.ninio; return v < s.max_ninio ? v : s.max_ninio; };
    w.GB[L] = (L >= 2 && L <= 30) ? s.bulge[lc] : kBig;
    w.G1[L] = (L >= 4 && L <= 30) ? s.interior[lc] + asym(L - 2) : kBig;
    if (L <= 30)
      for (int n1 = 0; n1 < 32; ++n1)
        w.GI[L][n1] = (n1 <= L && shape_class(n1, L - n1) == kGen) ? s.interior[L] + asym(L - 2 * n1) : kBig;
  }
  memcpy(w.mmH, s.mmH, sizeof(w.mmH));
  memcpy(w.mmI, s.mmI, sizeof(w.mmI));
  memcpy(w.mm1nI, s.mm1nI, sizeof(w.mm1nI));
  memcpy(w.mmM, s.mmM, sizeof(w.mmM));
  memcpy(w.mmExt, s.mmExt, sizeof(w.mmExt));
  w.MLbase = s.MLbase;
  w.MLclosing = s.MLclosing;
  w.MLintern = s.MLintern;
  w.TerminalAU = s.TerminalAU;
  cudaError_t e = cudaMemcpyToSymbolAsync(cW, &w, sizeof(w), 0, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return e;
  static SmallParams sps;
  sps = s;
  e = cudaMemcpyToSymbolAsync(cSP, &sps, sizeof(sps), 0, cudaMemcpyHostToDevice, st);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(st);
}

// pitches with a specialised instance of the short-sequence int16 kernel: the launch classes whose longest
// sequence is 57..64, 73..80, 97..104 or 121..128 nt (the usual case when a class of a large batch is full)
#define WAVE_RS_LIST(X) X(64) X(80) X(104) X(128)
#define WAVE_RS_LIST_256(X) X(160) X(200)   // 256-thread instance: the 129..160 and 161..200 nt classes
#define WAVE_RS_LIST_512(X) X(104) X(256)   // 512-thread instance: 225..256 nt, and small batches of ~100 nt (RL steps)

template <typename ST, int MAXT, int RS>
static cudaError_t launch_wave_rs(const FoldParams *gp, const BatchArgs &a, int nmax, int grid, int *queue,
                                  cudaStream_t st) {
  const size_t smem = wave_smem_bytes(nmax, (int)sizeof(ST));
  cudaError_t err =
      cudaFuncSetAttribute(fold_wave_kernel<ST, MAXT, RS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (err != cudaSuccess) return err;
  fold_wave_kernel<ST, MAXT, RS><<<grid, MAXT, smem, st>>>(gp, a, nmax, queue);
  return cudaGetLastError();
}

template <typename ST, int MAXT>
static cudaError_t launch_wave_t(const FoldParams *gp, const BatchArgs &a, int nmax, int grid, int *queue,
                                 cudaStream_t st) {
  const int rs = (nmax + 7) / 8 * 8;
#define X(R) if (rs == R) return launch_wave_rs<ST, MAXT, R>(gp, a, nmax, grid, queue, st);
  if constexpr (sizeof(ST) == 2 && MAXT == kSmallThreads) { WAVE_RS_LIST(X) }
  if constexpr (sizeof(ST) == 2 && MAXT == 256) { WAVE_RS_LIST_256(X) }
  if constexpr (sizeof(ST) == 2 && MAXT == 512) { WAVE_RS_LIST_512(X) }
#undef X
  return launch_wave_rs<ST, MAXT, 0>(gp, a, nmax, grid, queue, st);
}

template <typename ST, int MAXT, int RS>
static int occupancy_rs(int nmax) {
  int nb = 0;
  const size_t smem = wave_smem_bytes(nmax, (int)sizeof(ST));
  if (cudaFuncSetAttribute(fold_wave_kernel<ST, MAXT, RS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) !=
      cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, fold_wave_kernel<ST, MAXT, RS>, MAXT, smem);
  return nb;
}

template <typename ST, int MAXT>
static int occupancy_t(int nmax) {
  const int rs = (nmax + 7) / 8 * 8;
#define X(R) if (rs == R) return occupancy_rs<ST, MAXT, R>(nmax);
  if constexpr (sizeof(ST) == 2 && MAXT == kSmallThreads) { WAVE_RS_LIST(X) }
  if constexpr (sizeof(ST) == 2 && MAXT == 256) { WAVE_RS_LIST_256(X) }
  if constexpr (sizeof(ST) == 2 && MAXT == 512) { WAVE_RS_LIST_512(X) }
#undef X
  return occupancy_rs<ST, MAXT, 0>(nmax);
}

#define WAVE_DISPATCH(FN, ...)                                                        \
  (esz == 2 ? (threads == kSmallThreads ? FN<int16_t, kSmallThreads>(__VA_ARGS__)      \
               : threads == 256 ? FN<int16_t, 256>(__VA_ARGS__)                        \
                                : FN<int16_t, 512>(__VA_ARGS__))                       \
            : (threads == kSmallThreads ? FN<int32_t, kSmallThreads>(__VA_ARGS__)      \
               : threads == 256 ? FN<int32_t, 256>(__VA_ARGS__)                        \
                                : FN<int32_t, 512>(__VA_ARGS__)))

cudaError_t launch_fold_wave(const FoldParams *gp, const BatchArgs &a, int nmax, int esz, int grid, int threads,
                             int *queue, cudaStream_t st) {
  return WAVE_DISPATCH(launch_wave_t, gp, a, nmax, grid, queue, st);
}

int wave_occupancy(int nmax, int esz, int threads) { return WAVE_DISPATCH(occupancy_t, nmax); }

#ifdef WAVE_TRACE
extern "C" int rnafold_debug_wave_trace(unsigned *out, int count) {
  if (count > 512 * 16 * 8) count = 512 * 16 * 8;
  return cudaMemcpyFromSymbol(out, g_wave_trace, sizeof(unsigned) * count) == cudaSuccess ? 0 : -1;
}
#endif
#ifdef CLUSTER_TRACE
extern "C" int rnafold_debug_cluster_trace(unsigned *out, int count) {
  if (count > 2 * 480 * 16 * 4) count = 2 * 480 * 16 * 4;
  return cudaMemcpyFromSymbol(out, g_cluster_trace, sizeof(unsigned) * count) == cudaSuccess ? 0 : -1;
}
extern "C" int rnafold_debug_cluster_kinds(unsigned *out, int count) {
  if (count > 480 * 16) count = 480 * 16;
  return cudaMemcpyFromSymbol(out, g_cluster_kind, sizeof(unsigned) * count) == cudaSuccess ? 0 : -1;
}
#endif
#ifdef WAVE_PHASE_CLK
extern "C" int rnafold_debug_phase_clocks(unsigned long long *out, int count, int reset) {
  if (count > 640) count = 640;
  cudaError_t e = cudaMemcpyFromSymbol(out, g_phase_clk, sizeof(unsigned long long) * count);
  if (e == cudaSuccess && reset) {
    static unsigned long long zeros[640];
    e = cudaMemcpyToSymbol(g_phase_clk, zeros, sizeof(zeros));
  }
  return e == cudaSuccess ? 0 : -1;
}
#endif

// ---- cluster path (long sequences)
size_t cluster_smem_bytes(int nmax, int esz) { return (size_t)ClusterSmem::make(nmax, esz).total; }
size_t cluster_table_stride(int nmax) { return (size_t)(table_cells(nmax) + 64) / 8 * 8; }  // elements per table per sequence

// one instance per cluster length class (capi.cu kClusterClasses: the class IS the row pitch) + the generic one
#define CLUSTER_RS_LIST(X) X(288) X(320) X(352) X(400) X(432) X(464)

template <int RS>
static int cluster_max_active_rs(size_t smem) {
  if (cudaFuncSetAttribute(fold_cluster_kernel<int16_t, RS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * 1024, 1, 1);
  cfg.blockDim = dim3(kClusterThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = 2;
  attr.val.clusterDim.y = 1;
  attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  int nc = 0;
  if (cudaOccupancyMaxActiveClusters(&nc, fold_cluster_kernel<int16_t, RS>, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return nc;
}

int cluster_max_active(int nmax) {
  // co-resident 2-CTA clusters on the device for this length class
  const size_t smem = cluster_smem_bytes(nmax, 2);
  const int rs = (nmax + 7) / 8 * 8;
#define X(R) if (rs == R) return cluster_max_active_rs<R>(smem);
  CLUSTER_RS_LIST(X)
#undef X
  return cluster_max_active_rs<0>(smem);
}

template <int RS>
static cudaError_t launch_fold_cluster_rs(const FoldParams *gp, const BatchArgs &a, int nmax, int num_clusters, int *queue,
                                          void *cbg, void *fg, cudaStream_t st) {
  const size_t smem = cluster_smem_bytes(nmax, 2);
  cudaError_t err = cudaFuncSetAttribute(fold_cluster_kernel<int16_t, RS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (err != cudaSuccess) return err;
  fold_cluster_kernel<int16_t, RS><<<2 * num_clusters, kClusterThreads, smem, st>>>(
      gp, a, nmax, queue, (int16_t *)cbg, (int16_t *)fg, (long long)cluster_table_stride(nmax));
  return cudaGetLastError();
}

cudaError_t launch_fold_cluster(const FoldParams *gp, const BatchArgs &a, int nmax, int num_clusters, int *queue,
                                void *cbg, void *fg, cudaStream_t st) {
  const int rs = (nmax + 7) / 8 * 8;
#define X(R) if (rs == R) return launch_fold_cluster_rs<R>(gp, a, nmax, num_clusters, queue, cbg, fg, st);
  CLUSTER_RS_LIST(X)
#undef X
  return launch_fold_cluster_rs<0>(gp, a, nmax, num_clusters, queue, cbg, fg, st);
}

cudaError_t launch_f5_backtrack(const FoldParams *gp, const BatchArgs &a, int nmax, int count, const void *cbg,
                                const void *fg, cudaStream_t st) {
  const size_t smem = (size_t)kBtWarps * bt_warp_bytes(nmax);
  cudaError_t err = cudaFuncSetAttribute(f5_backtrack_kernel<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (err != cudaSuccess) return err;
  const int grid = (count + kBtWarps - 1) / kBtWarps;  // count bounds the batch size when it lives on the device
  f5_backtrack_kernel<int16_t><<<grid, kBtWarps * 32, smem, st>>>(gp, a, nmax, (const int16_t *)cbg, (const int16_t *)fg,
                                                                 (long long)cluster_table_stride(nmax));
  return cudaGetLastError();
}

}  // namespace rnafold
