// fold_cluster.inl -- long sequences: one thread-block cluster (2 CTAs on 2 SMs) per sequence.
// Included by fold_wave.cu (same translation unit: shares cW, the table helpers and backtrack_warp).
//
// At n = 400 one sequence needs 157 KB for fML alone (int16), so the tables are split by ROLE across
// the distributed shared memory of a 2-CTA cluster, and only the wavefront rows cross between them:
//
//   rank 0 "pair CTA"       32-diagonal windows of cB, cI, c1; interior-loop / tabulated-shape work
//                           items over the list of pairable cells; finalises c of diagonal d-1 and
//                           PUSHES the row  c + E_MLstem  into rank 1's shared memory (DSMEM store).
//   rank 1 "multiloop CTA"  the full fML triangle; multibranch split work items; finalises fML of
//                           diagonal d-2 and PUSHES the row DML (best split) into rank 0's shared memory,
//                           where it closes multiloops two diagonals later.
//
// One cluster barrier per anti-diagonal orders the pushes.  Both CTAs also stream their finished
// rows (cB, fML) to a global workspace; the exterior recursion f5, the backtrack and the Hamming
// distance then run in a second kernel with one WARP per sequence (f5_backtrack_kernel), so the long
// serial tail of a fold never idles a cluster.
namespace cg = cooperative_groups;

constexpr int kClusterThreads = 512;

// Split cluster barrier (barrier.cluster.arrive / .wait): a CTA announces the end of its phase without blocking and
// only waits -- right before the items that consume the row its peer pushed -- for the peer's announcement.
__device__ __forceinline__ void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
// End of a phase, called by every thread right after a __syncthreads: the CTA barrier has ordered every thread's stores
// (local, DSMEM pushes, global table copies) before warp 0's release, which is cumulative -- so ONE warp pays the fence
// (MEMBAR.ALL.GPU in SASS) and the others announce without one (the pattern of a grid-wide sync: barrier, then one fence).
__device__ __forceinline__ void cluster_arrive_phase(int warp) {
#ifdef CLUSTER_ALL_FENCE   // the round-2 v3 behaviour, kept for A/B runs (tools/var_bench.sh): 64.2 k vs 72.4 k folds/s at n = 400
  (void)warp;
  cluster_arrive_release();
#else
  if (warp == 0) cluster_arrive_release();
  else cluster_arrive_relaxed();
#endif
}
__device__ __forceinline__ void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

#ifdef CLUSTER_TRACE   // debug builds only (tools/cluster_trace.py): per-warp time stamps of every phase of ONE fold
__device__ unsigned g_cluster_trace[2 * 480 * 16 * 4];
__device__ unsigned g_cluster_kind[480 * 16];   // [d][kind]: cycles (kinds 0..7) and item counts (8..15), pair CTA
#define TRACE_AT(slot_) do { if (tracing && lane == 0 && d < 480) g_cluster_trace[((rank * 480 + d) * 16 + warp) * 4 + (slot_)] = (unsigned)(clock64() - trace_t0); } while (0)
#define KIND_BEGIN() long long kind_t0 = clock64()
#define KIND_END(kind_) do { if (tracing && lane == 0 && d < 480) { const long long t1_ = clock64(); atomicAdd(&g_cluster_kind[d * 16 + (kind_)], (unsigned)(t1_ - kind_t0)); atomicAdd(&g_cluster_kind[d * 16 + 8 + (kind_)], 1u); kind_t0 = t1_; } } while (0)
#else
#define TRACE_AT(slot_) do { } while (0)
#define KIND_BEGIN() do { } while (0)
#define KIND_END(kind_) do { } while (0)
#endif

struct ClusterSmem {
  // common
  int misc_off, ptab_off, S_off;
  // rank 0
  int wB_off, wI_off, w1_off, part_off, dml_off, pl_off, plc_off, gd_off;
  // rank 1 (overlays rank 0's region)
  int F_off, mlp_off, cst_off;
  int total, rs, mlparts, np;
  __host__ __device__ static ClusterSmem make(int nmax, int esz) {
    ClusterSmem L;
    int o = 0;
    auto take = [&](int bytes) { int r = o; o += (bytes + 15) / 16 * 16; return r; };
    L.np = wave_np(nmax);
    L.rs = (nmax + 7) / 8 * 8;
    L.mlparts = 1;  // split chunks are merged by atomicMin into ONE row per diagonal (three diagonals in flight)
    L.misc_off = take(64);
    L.ptab_off = take(64);
    L.S_off = take(L.np);
    const int base = o;
    L.wB_off = take((kWin * L.rs + 64) * esz);
    L.wI_off = take((kWin * L.rs + 64) * esz);
    L.w1_off = take((kWin * L.rs + 64) * esz);
    L.part_off = take(2 * kWSlots * L.np * esz);  // slot 0 generic shapes, 1 tabulated, 2.. bulge/1xn parts
    L.dml_off = take(3 * L.np * 4);
    L.pl_off = take(2 * L.np * 2);
    L.plc_off = take(16);
    L.gd_off = take(kGenSlots * 2);
    const int end0 = o;
    o = base;
    L.F_off = take((fold_rows(L.rs) * L.rs + 64) * esz);  // folded layout, pitch rs
    L.mlp_off = take(3 * L.mlparts * L.np * 4);
    L.cst_off = take(2 * L.np * 4);
    L.total = o > end0 ? o : end0;
    return L;
  }
};

// RS > 0: compile-time row pitch (= the launch's length class), as in fold_wave_kernel
template <typename ST, int RS = 0>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kClusterThreads, 1)
fold_cluster_kernel(const FoldParams *__restrict__ gp, BatchArgs a, int nmax, int *__restrict__ queue,
                    ST *__restrict__ CBg, ST *__restrict__ Fg, long long gstride) {
  extern __shared__ __align__(16) unsigned char smem[];
  if (a.dyn && a.dyn[0] == 0) return;  // device-counted launch with nothing routed to it (uniform over the cluster)
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const ClusterSmem Ls = ClusterSmem::make(nmax, (int)sizeof(ST));
  int *misc = reinterpret_cast<int *>(smem + Ls.misc_off);
  unsigned char *ptab = smem + Ls.ptab_off;
  unsigned char *S = smem + Ls.S_off;
  ST *wB = reinterpret_cast<ST *>(smem + Ls.wB_off);
  ST *wI = reinterpret_cast<ST *>(smem + Ls.wI_off);
  ST *w1 = reinterpret_cast<ST *>(smem + Ls.w1_off);
  ST *part = reinterpret_cast<ST *>(smem + Ls.part_off);
  int *dml = reinterpret_cast<int *>(smem + Ls.dml_off);
  unsigned short *pl = reinterpret_cast<unsigned short *>(smem + Ls.pl_off);
  int *plc = reinterpret_cast<int *>(smem + Ls.plc_off);
  ST *F = reinterpret_cast<ST *>(smem + Ls.F_off);
  int *mlp = reinterpret_cast<int *>(smem + Ls.mlp_off);
  int *cst = reinterpret_cast<int *>(smem + Ls.cst_off);
  // the peer's copies of the rows this CTA pushes
  int *peer_cst = cluster.map_shared_rank(cst, 1);
  int *peer_dml = cluster.map_shared_rank(dml, 0);
  int *peer_misc = cluster.map_shared_rank(misc, 1);
  const int np = Ls.np, rs = RS > 0 ? RS : Ls.rs, mlparts = Ls.mlparts;
  const int tid = threadIdx.x, lane = tid & 31, NW = kClusterThreads >> 5, NT = kClusterThreads;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const SmallParams &sp = gp->s;
  constexpr int kInfS = Store<ST>::kInfS;
  const int tau = cW.TerminalAU;
  const unsigned wI_sa = (unsigned)__cvta_generic_to_shared(wI), w1_sa = (unsigned)__cvta_generic_to_shared(w1),
                 wB_sa = (unsigned)__cvta_generic_to_shared(wB);

  if (tid < 64) ptab[tid] = (unsigned char)(((tid >> 3) < 5 && (tid & 7) < 5) ? pair_type(tid >> 3, tid & 7) : 0);
  auto PT = [&](int x, int y) -> int { return ptab[x * 8 + y]; };
  auto craw = [&](int cb, int type) -> int { return cb - (type > 2 ? tau : 0); };
  // lane slots of the shape-parallel generic interior item (see fold_wave_kernel): slot -> L | n1 << 8
  unsigned short *gdesc = reinterpret_cast<unsigned short *>(smem + Ls.gd_off);
  if (rank == 0)
    for (int sidx = tid; sidx < kGenSlots; sidx += NT) {
      const int sc = min(sidx, kGenShapes - 1);
      int L = 6;
      while (gen_before(L + 1) <= sc) ++L;
      gdesc[sidx] = (unsigned short)(L | ((2 + sc - gen_before(L)) << 8));
    }

  for (;;) {
    cluster.sync();
    if (rank == 0 && tid == 0) {
      const int s = atomicAdd(queue, 1);
      misc[0] = s;
      peer_misc[0] = s;
    }
    cluster.sync();
    const int slot = misc[0];
    if (slot >= batch_count(a)) break;
    const int b = batch_entry(a, slot);
    const int beg = a.off[b];
    const int n = a.off[b + 1] - beg;
    ST *cbg = CBg + (long long)slot * gstride;
    ST *fg = Fg + (long long)slot * gstride;
    int flag = 0;

    for (int k = tid; k < n; k += NT) S[k + 1] = (unsigned char)enc_base(a.seq[beg + k]);
    if (rank == 1)
      for (int k = tid; k < 3 * mlparts * np; k += NT) mlp[k] = kInf;
    __syncthreads();
    if (tid == 0 && n > 0) {
      S[0] = S[n];
      S[n + 1] = S[1];
    }
    cluster.sync();

    // pair CTA: window offsets of the generic-shape slots at d = 4 (one row down per diagonal, wrapping)
    int gterm[kGenIters], goff[kGenIters];
    unsigned gpack[kGenIters / 4] = {};
    if (rank == 0) {
#pragma unroll
      for (int k = 0; k < kGenIters; ++k) {
        const int ds = gdesc[32 * k + lane];
        const int L = ds & 255, n1 = ds >> 8;
        gterm[k] = (32 * k + lane < kGenShapes) ? cW.GI[L][n1] : kBig;
        goff[k] = (int)wI_sa + (((kTurn + 1 - 2 - L) & (kWin - 1)) * rs + 1 + n1) * (int)sizeof(ST);  // shared-space addresses
        gpack[k >> 2] |= (unsigned)((1 + n1) * (int)sizeof(ST)) << (8 * (k & 3));
      }
    } else {
#pragma unroll
      for (int k = 0; k < kGenIters; ++k) gterm[k] = goff[k] = 0;
    }

    // Protocol per phase: [items that only read this CTA's tables] -> wait for the peer's previous phase -> [items that
    // read the row the peer pushed: Fp / Fm] -> __syncthreads -> arrive.  The rows are double / triple buffered, and a row
    // is only overwritten by an item that runs behind a wait, i.e. after its consumer's phase has been announced.
    cluster_arrive_release();
#ifdef CLUSTER_TRACE
    const bool tracing = blockIdx.x < 2 && slot < 74;   // cluster 0's first fold
    const long long trace_t0 = clock64();
#endif
    for (int d = kTurn + 1; d <= n + 1; ++d) {
#ifdef WAVE_PHASE_CLK
      const long long clk_p0 = clock64();
#endif
      TRACE_AT(0);
      const int par = d & 1, par1 = par ^ 1;
      const int cells_d = (d < n) ? n - d : 0;
      const int nb_d = (cells_d + 31) >> 5;
      const int lmax = min(kMaxLoop, d - 6);
      if (rank == 0) {
        // ================= pair CTA
        const int d1 = d - 1;
        const int cells_1 = (d1 >= kTurn + 1 && d1 < n) ? n - d1 : 0;
        const int nb_1 = (cells_1 + 31) >> 5;
        const int npair = (lmax >= 0 && d < n) ? plc[par] : 0;
        const unsigned short *pld = pl + par * np;
        const int nbI = (npair + 31) >> 5;
        const int nP = (d + 1 < n && d + 1 >= 6) ? 1 : 0;
        int neparts = 0;
#pragma unroll
        for (int r = 0; r < kEParts; ++r) neparts += (epart_beg(r) <= lmax) ? 1 : 0;
        const int nI = nbI * neparts, nT = (lmax >= 0) ? nbI : 0, nF = nb_1;
        const int total = nI + nT + nP + nF;
        auto finalise_pairs = [&](int kf) {
          // ---- Fp(b): finalise c of diagonal d1 (needs the DML row the multiloop CTA pushed in its previous phase),
          // push c + E_MLstem to the multiloop CTA; lookups up front and unconditional, masked afterwards
          {
            const int x = kf * 32 + lane;
            const bool live = x < cells_1;
            const int xs = live ? x : 0;
            const int i = xs + 1, j = i + d1;
            const int type = PT(S[i], S[j]);
            const int si1 = S[i + 1], sj1 = S[j - 1], sim1 = S[i - 1], sjp1 = S[j + 1];
            const int rt = rtype_of(type);
            const int v_mmh = __ldg(&sp.mmH[type][si1][sj1]);
            const int v_mlc = e_mlstem(sp, rt, sj1, si1);
            const int v_mmii = __ldg(&sp.mmI[rt][sjp1][sim1]);
            const int v_mm1i = __ldg(&sp.mm1nI[rt][sjp1][sim1]);
            const int v_mls = e_mlstem(sp, type, sim1, sjp1);
            const int hbase = (d1 - 1 <= 30) ? sp.hairpin[d1 - 1] : __ldg(&gp->hairpin_len[d1 - 1]);
            const int v_mmio = __ldg(&sp.mmI[type][si1][sj1]);
            const ST *src = part + (par1 * kWSlots) * np + xs;
            const int s_g = (int)src[0], s_t = (int)src[np];
            int s_e = kInfS;
#pragma unroll
            for (int r = 0; r < kEParts; ++r) s_e = (d1 - 6 >= epart_beg(r)) ? min(s_e, (int)src[(2 + r) * np]) : s_e;
            const int v_dml = dml[((d1 - 2) % 3) * np + xs + 1];
            int cij = hbase + v_mmh;
            if (d1 == 4 || d1 == 5 || d1 == 7) cij = type ? e_hairpin(sp, gp, S, i, j, type) : cij;
            cij = (d1 >= 12) ? min(cij, s_g + v_mmio) : cij;
            cij = (d1 >= 6) ? min(cij, s_t) : cij;
            cij = (d1 >= 8) ? min(cij, s_e) : cij;
            cij = (d1 >= 2 * kTurn + 4) ? min(cij, v_dml + v_mlc + cW.MLclosing) : cij;
            if (live) {
              const int wrow = (d1 & (kWin - 1)) * rs + x;
              int stem = kInf;
              ST cb = (ST)kInfS;
              if (type) {
                cb = put<ST>(cij + (type > 2 ? tau : 0), flag);
                wI[wrow] = put<ST>(cij + v_mmii, flag);
                w1[wrow] = put<ST>(cij + v_mm1i, flag);
                stem = cij + v_mls;
              } else {
                wI[wrow] = (ST)kInfS;
                w1[wrow] = (ST)kInfS;
              }
              wB[wrow] = cb;
              cbg[tri(n, d1) + x] = cb;
              peer_cst[(d1 & 1) * np + x] = stem;
            }
          }
        };
        bool waited = false;
        KIND_BEGIN();
        for (int k = warp; k < total; k += NW) {
          if (k >= nI + nT + nP) {
            if (!waited) {
              cluster_wait_acquire();   // the multiloop CTA's previous phase (its DML push) is complete
              waited = true;
            }
            finalise_pairs(k - nI - nT - nP);
            KIND_END(3);
            continue;
          }
          if (k < nI) {
            // ---- E(b,r): bulges and 1xn loops of the loop sizes in part r, lanes across pairable cells
            const int r = div_small(k, nbI, kEParts), blk = k - r * nbI;
            const int l0 = r == 0 ? epart_beg(0) : r == 1 ? epart_beg(1) : epart_beg(2);
            const int l1 = min((r == 0 ? epart_beg(1) : r == 1 ? epart_beg(2) : epart_beg(3)) - 1, lmax);
            const int idx = blk * 32 + lane;
            const int x = pld[min(idx, npair - 1)];
            const int oi = x + 1, oj = oi + d;
            const int type = PT(S[oi], S[oj]);
            const int v_mm1 = __ldg(&sp.mm1nI[type][S[oi + 1]][S[oj - 1]]);
            int a1 = kInfS, aB = kInfS;
            e_item_sa<ST, true>(w1_sa + (unsigned)((x + 1) * (int)sizeof(ST)), wB_sa + (unsigned)((x + 1) * (int)sizeof(ST)), d, n,
                                rs, l0, l1, a1, aB);
            if (idx < npair) {
              const int e = min(a1 + v_mm1, aB + (type > 2 ? tau : 0));
              part[(par * kWSlots + 2 + r) * np + x] = (ST)min(e, kInfS);
            }
            KIND_END(0);
          } else if (k < nI + nT) {
            // ---- T(b): tabulated shapes, inner pairs from the cB window
            const int idx = (k - nI) * 32 + lane;
            if (idx < npair) {
              const int x = pld[idx];
              const int i = x + 1, j = i + d;
              const int type = PT(S[i], S[j]);
              int best = kInfS;
              int si[5], sj[5];
#pragma unroll
              for (int u = 0; u < 5; ++u) {
                si[u] = S[i + u];
                sj[u] = S[j - u];
              }
#pragma unroll
              for (int s = 0; s < 9; ++s) {  // branch-free: the nine lookup chains overlap (see fold_wave_kernel)
                constexpr int kN1[9] = {0, 0, 1, 1, 1, 2, 2, 2, 3};
                constexpr int kN2[9] = {0, 1, 0, 1, 2, 1, 2, 3, 2};
                const int n1 = kN1[s], n2 = kN2[s];
                const int t2 = PT(si[n1 + 1], sj[n2 + 1]);
                const int cpq = craw((int)wB[((d - 2 - n1 - n2) & (kWin - 1)) * rs + x + 1 + n1], t2);
                const int e = cpq + e_intloop(sp, gp, n1, n2, type, rtype_of(t2), si[1], sj[1], si[n1], sj[n2]);
                best = (n1 + n2 <= d - 6 && t2 != 0) ? min(best, e) : best;
              }
              part[(par * kWSlots + 1) * np + x] = (ST)min(best, kInfS);
            }
            KIND_END(1);
          } else {
            // ---- P: list of the pairable cells of diagonal d+1 (used by the next phase)
            const int dn = d + 1, cells_n = n - dn;
            unsigned short *dst = pl + par1 * np;
            int pos = 0;
            for (int base = 0; base < cells_n; base += 32) {
              const int x = base + lane;
              const bool pr = x < cells_n && PT(S[x + 1], S[x + 1 + dn]) != 0;
              const unsigned m = __ballot_sync(0xffffffffu, pr);
              if (pr) dst[pos + __popc(m & ((1u << lane) - 1u))] = (unsigned short)x;
              pos += __popc(m);
            }
            if (lane == 0) plc[par1] = pos;
            KIND_END(2);
          }
        }
        // ---- G(cell): generic interior shapes, lanes across shapes, one pairable cell at a time
        const int nvalid = (d < n) ? gen_valid(d) : 0;
        int cg0 = warp - (total % NW);
        if (cg0 < 0) cg0 += NW;
        if (nvalid == kGenSlots) {
          int c = cg0;
#pragma unroll 1
          for (; c + NW < npair; c += 2 * NW) {
            const int xv0 = pld[c], xv1 = pld[c + NW];
            int r0, r1;
            gen_cell2_sa<ST>(__reduce_max_sync(0xffffffffu, (unsigned)xv0 * (unsigned)sizeof(ST)),
                             __reduce_max_sync(0xffffffffu, (unsigned)xv1 * (unsigned)sizeof(ST)), goff, gterm, r0, r1);
            if (lane == 0) {
              part[(par * kWSlots + 0) * np + xv0] = (ST)min(r0, kInfS);
              part[(par * kWSlots + 0) * np + xv1] = (ST)min(r1, kInfS);
            }
          }
#pragma unroll 1
          for (; c < npair; c += NW) {
            const int xv = pld[c];
            const int r = gen_cell_sa<ST, true>(__reduce_max_sync(0xffffffffu, (unsigned)xv * (unsigned)sizeof(ST)), goff, gterm, nvalid, lane);
            if (lane == 0) part[(par * kWSlots + 0) * np + xv] = (ST)min(r, kInfS);
          }
        } else if (nvalid > 0) {
#pragma unroll 1
          for (int c = cg0; c < npair; c += NW) {
            const int xv = pld[c];
            const int r = gen_cell_sa<ST, false>(__reduce_max_sync(0xffffffffu, (unsigned)xv * (unsigned)sizeof(ST)), goff, gterm, nvalid, lane);
            if (lane == 0) part[(par * kWSlots + 0) * np + xv] = (ST)min(r, kInfS);
          }
        }
#pragma unroll
        for (int k = 0; k < kGenIters; ++k) {
          const int bb = (int)((gpack[k >> 2] >> (8 * (k & 3))) & 0xffu);
          const int t = goff[k] + rs * (int)sizeof(ST), u = t - kWin * rs * (int)sizeof(ST);
          goff[k] = (u >= bb + (int)wI_sa) ? u : t;
        }
        KIND_END(4);
        TRACE_AT(1);
        if (!waited) cluster_wait_acquire();
        TRACE_AT(2);
      } else {
        // ================= multiloop CTA
        // split range t = 4 .. d-5 cut into enough chunks that every warp of this CTA gets work even on the
        // late diagonals (few cells, many split points); chunk minima are merged with shared-memory atomicMin
        const int nsplit = d - 2 * kTurn - 2;
        int nch = 0, chunk = 8;
        if (d < n && nsplit > 0) {
          const int want = (WAVE_MLWANT * NW + nb_d - 1) / nb_d;
          nch = max(1, min(want, (nsplit + 15) >> 4));
          chunk = ((nsplit + nch - 1) / nch + 7) & ~7;
          nch = (nsplit + chunk - 1) / chunk;
        }
        const int d2 = d - 2;
        const int cells_2 = (d2 >= kTurn + 1 && d2 < n) ? n - d2 : 0;
        const int nb_2 = (cells_2 + 31) >> 5;
        const int nM = nb_d * nch;
        bool waited = false;
        for (int k = warp; k < nM; k += NW) {
          // ---- M(b,m)
          const int m = k / nb_d, blk = k - m * nb_d;
          const int x = blk * 32 + lane, xc = min(x, cells_d - 1);
          const int t0 = kTurn + 1 + m * chunk, t1 = min(d - kTurn - 1, t0 + chunk);
          const int best = split_scan<ST>(F, rs, d, xc, t0, t1);
          if (x < cells_d) atomicMin(&mlp[(d % 3) * mlparts * np + x], best);
        }
        {
          int f0 = warp - (nM % NW);
          if (f0 < 0) f0 += NW;
          for (int kf = f0; kf < nb_2; kf += NW) {
            if (!waited) {
              cluster_wait_acquire();   // the pair CTA's previous phase (its stem-row push) is complete
              waited = true;
            }
            // ---- Fm(b): finalise fML of diagonal d2, push DML to the pair CTA
            const int x = kf * 32 + lane;
            if (x < cells_2) {
              int e = kInf;
              if (d2 > kTurn + 1) {
                const int pb = fold_idx(rs, d2 - 1, x);
                e = min((int)F[pb + 1], (int)F[pb]) + cW.MLbase;
              }
              e = min(e, cst[(d2 & 1) * np + x]);
              const int dm = mlp[(d2 % 3) * mlparts * np + x];
              mlp[(d2 % 3) * mlparts * np + x] = kInf;  // ready for diagonal d2 + 3
              peer_dml[(d2 % 3) * np + x] = dm;
              const ST f = put<ST>(min(e, dm), flag);
              F[fold_idx(rs, d2, x)] = f;
              fg[tri(n, d2) + x] = f;
            }
          }
        }
        TRACE_AT(1);
        if (!waited) cluster_wait_acquire();
        TRACE_AT(2);
      }
#ifdef WAVE_PHASE_CLK
      __syncthreads();
      if (tid == 0) atomicAdd(&g_phase_clk[(rank ? 64 : 0) + (d >> 4)], (unsigned long long)(clock64() - clk_p0));
#endif
      __syncthreads();            // this CTA's tables of the phase are complete ...
      TRACE_AT(3);
      cluster_arrive_phase(warp); // ... and so are its pushes: announce the phase, do not wait here
#ifdef WAVE_PHASE_CLK
      if (tid == 0 && rank == 0) atomicAdd(&g_phase_clk[128 + (d >> 4)], (unsigned long long)(clock64() - clk_p0));
#endif
    }
    cluster_wait_acquire();       // drain the last announcement (pairs with the arrive above)
    if (Store<ST>::kChecked && flag && a.redo) a.redo[b] = 1;
  }
}

// ---------------------------------------------------------------------------------------------
// f5 + backtrack + Hamming/reward from the global tables: one warp per sequence.
// ---------------------------------------------------------------------------------------------
constexpr int kBtWarps = 4;

__host__ __device__ inline int bt_warp_bytes(int nmax) {
  const int np = wave_np(nmax);
  return ((np * 4 + 15) / 16 * 16) * 2 + ((np + 15) / 16 * 16) * 2;  // f5, bt (int) + S, db (bytes)
}

template <typename ST>
__global__ void __launch_bounds__(kBtWarps * 32)
f5_backtrack_kernel(const FoldParams *__restrict__ gp, BatchArgs a, int nmax, const ST *__restrict__ CBg,
                    const ST *__restrict__ Fg, long long gstride) {
  extern __shared__ __align__(16) unsigned char smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = blockIdx.x * kBtWarps + warp;
  if (slot >= batch_count(a)) return;
  const int b = batch_entry(a, slot);
  if (a.redo && a.redo[b]) return;  // left to the int32 re-run
  const int np = wave_np(nmax);
  unsigned char *base = smem + (size_t)warp * bt_warp_bytes(nmax);
  int *f5 = reinterpret_cast<int *>(base);
  int *bt = f5 + ((np * 4 + 15) / 16 * 16) / 4;
  unsigned char *S = reinterpret_cast<unsigned char *>(bt) + (np * 4 + 15) / 16 * 16;
  unsigned char *db = S + (np + 15) / 16 * 16;
  const SmallParams &sp = gp->s;
  const int beg = a.off[b];
  const int n = a.off[b + 1] - beg;
  const ST *CB = CBg + (long long)slot * gstride;
  const ST *F = Fg + (long long)slot * gstride;

  for (int k = lane; k < n; k += 32) {
    S[k + 1] = (unsigned char)enc_base(a.seq[beg + k]);
    db[k] = '.';
  }
  for (int k = lane; k <= n; k += 32) f5[k] = 0;
  __syncwarp();
  if (lane == 0 && n > 0) {
    S[0] = S[n];
    S[n + 1] = S[1];
  }
  __syncwarp();
  for (int j = kTurn + 2; j <= n; ++j) {
    const int s3 = (j < n) ? (int)S[j + 1] : -1;
    int best = kInf;
    for (int i = j - kTurn - 1 - lane; i >= 1; i -= 32) {
      const int type = pair_type(S[i], S[j]);
      if (!type) continue;
      // cB = c + TerminalAU(pair): plus the mismatch / dangle part of E_ExtLoop it is c + E_ExtLoop
      best = min(best, f5[i - 1] + (int)CB[tri(n, j - i) + i - 1] + e_extloop_mm(sp, type, i > 1 ? (int)S[i - 1] : -1, s3));
    }
    for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) f5[j] = min(f5[j - 1], best);
    __syncwarp();
  }
  const int failc = backtrack_warp<ST, false, true>(n, S, f5, CB, F, 0, bt, db, gp, lane);
  __syncwarp();

  int hd = 0;
  for (int k = lane; k < n; k += 32) {
    const unsigned char ch = db[k];
    if (a.db) a.db[beg + k] = ch;
    if (a.target) {
      const int mm = a.target[beg + k] != ch;
      hd += mm;
      if (a.mismatch) a.mismatch[beg + k] = (unsigned char)mm;
    }
  }
  for (int o = 16; o; o >>= 1) hd += __shfl_xor_sync(0xffffffffu, hd, o);
  if (lane == 0) {
    if (a.mfe) a.mfe[b] = (n > 0) ? f5[n] : 0;
    if (a.target) {
      if (a.hd) a.hd[b] = hd;
      if (a.reward) {
        double r = (n > 0) ? pow(1.0 - (double)hd / (double)n, a.reward_exp) : 0.0;
        if (hd == 0) r += 0.25;
        a.reward[b] = r;
      }
    }
    if (failc) atomicExch(a.status, failc);
  }
}
