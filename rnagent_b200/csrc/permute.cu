// permute.cu -- batched local search of Solution.permute (reference: rlif/rna/solution.py:157-219) as it is
// reached from Solution.evaluate(permute=True) (rlif/rna/solution.py:133-141): fold, and for every sequence
// with 0 < HD <= permutation_threshold run up to permutation_budget rounds of
//     mutate the positions around the mismatches -> refold -> keep the candidate iff its HD improves.
// All candidates of a round are folded by ONE call of the library's batched fold (fold + Hamming + mismatch
// flags), so a round costs one fold latency for the whole batch instead of one RNA.fold per solution.
//
// What is kept from the reference, quirks included:
//   * get_surrounding (solution.py:166-181): the candidate set holds only the NEIGHBOURS (distance 1..radius)
//     of the mismatches; the probability table is written in the reference's order (mismatches ascending,
//     distance ascending, `probs[index] = p` first, then index-i with linspace(0.01, p, radius+1)[i-1], then
//     index+i with linspace(...)[:-1][-i]) and later writes override earlier ones;
//   * the mutation (solution.py:188-201): action in 0..3, or 0..5 on paired positions when
//     allow_gu_permutations; the partner of a paired position receives the complementary letter;
//   * acceptance (solution.py:206-212): strictly smaller HD only; the search stops at HD 0.
// Python's `random` stream cannot be reproduced, so the draws are a counter-based hash of
// (seed + sequence number, round, position, kind) -- the SAME function the golden-vector generator
// (tools/make_permute_golden.py) patches into the reference's `random`, which makes the whole search
// bit-comparable with the reference's own permute() code.  Candidate positions are visited in the order in which
// the reference's `for mismatch in best_mismatch_indices` walks its Python SET (solution.py:183, :189): the order
// matters when two mutated positions write the same nucleotide (a position and the partner of another one), the
// later write wins.  pm_set_order() replays CPython's set insertion for the insertion sequence of get_surrounding
// and lists the table's slots -- checked against real `set` objects in tests/test_host_side.py.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

#include "../../include/rnafold_b200.h"
#include "fold_kernels.h"

namespace {

using rnafold::set_error;

#define PM_TRY(expr)                                                                    \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      char buf_[256];                                                                   \
      std::snprintf(buf_, sizeof(buf_), "%s: %s", #expr, cudaGetErrorString(e_));       \
      return set_error(RNAFOLD_ECUDA, buf_);                                            \
    }                                                                                   \
  } while (0)

__host__ __device__ __forceinline__ unsigned long long pm_mix64(unsigned long long x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}
// kind 0: random.random(); 1: random.randint(0, 3); 2: random.randint(0, 5)
__host__ __device__ __forceinline__ unsigned long long pm_draw(unsigned long long key, int step, int pos, int kind) {
  return pm_mix64(key ^ (((unsigned long long)(unsigned)step << 32) | ((unsigned long long)(unsigned)pos << 4) | (unsigned long long)kind));
}

struct PermuteCfg {
  int budget, radius, threshold, allow_gu;
  double p;
  unsigned long long seed;
  const unsigned long long *keys;  // per sequence: what the draws are keyed on instead of the batch index (device; may be null)
  int order;  // candidate visiting order: 0 CPython >= 3.7 set, 1 CPython 3.6 set (the reference's pinned interpreter), 2 ascending
};
int g_pm_order = 0;

// Iteration order of the Python set built by `set(all_indices)` (solution.py:181) from the keys in the order
// get_surrounding appends them.  CPython's setobject.c for small non-negative ints (hash(i) = i): slot = key & mask;
// a run of up to 10 consecutive slots is probed (only when it does not cross the table's end), then
// i = (5 i + 1 + perturb) & mask with perturb >>= 5; when fill * 5 >= mask * 3 (3.6: fill * 3 >= mask * 2) the table
// is rebuilt with the smallest power of two > 4 * used slots, re-inserting the keys in slot order.
// tab / tmp: 8 n + 16 ints each.  -> number of keys written to cand[]
__host__ __device__ inline int pm_set_order(int n, const uint8_t *mm, int radius, int order, int32_t *tab, int32_t *tmp, int32_t *cand) {
  int ncand = 0;
  if (order == 2) {  // ascending positions
    for (int i = 0; i < n; ++i) tab[i] = 0;
    for (int index = 0; index < n; ++index) {
      if (!mm[index]) continue;
      for (int i = 1; i <= radius; ++i) {
        if (index - i >= 0) tab[index - i] = 1;
        if (index + i <= n - 1) tab[index + i] = 1;
      }
    }
    for (int i = 0; i < n; ++i)
      if (tab[i]) cand[ncand++] = i;
    return ncand;
  }
  const int ra = order == 1 ? 3 : 5, rb = order == 1 ? 2 : 3;
  int mask = 7, fill = 0;
  for (int i = 0; i < 8; ++i) tab[i] = -1;
  auto insert_clean = [](int32_t *t, int msk, int key) {
    unsigned perturb = (unsigned)key;
    int i = key & msk;
    for (;;) {
      if (t[i] == -1) { t[i] = key; return; }
      if (i + 9 <= msk)
        for (int j = 1; j <= 9; ++j)
          if (t[i + j] == -1) { t[i + j] = key; return; }
      perturb >>= 5;
      i = (int)(((unsigned)i * 5u + 1u + perturb) & (unsigned)msk);
    }
  };
  auto add = [&](int key) {
    unsigned perturb = (unsigned)key;
    int i = key & mask;
    for (;;) {
      const int probes = (i + 9 <= mask) ? 9 : 0;
      for (int j = 0; j <= probes; ++j) {
        const int e = tab[i + j];
        if (e == key) return;
        if (e == -1) {
          tab[i + j] = key;
          ++fill;
          if (fill * ra >= mask * rb) {
            int ns = 8;
            while (ns <= fill * 4) ns <<= 1;
            for (int q = 0; q < ns; ++q) tmp[q] = -1;
            for (int q = 0; q <= mask; ++q)
              if (tab[q] != -1) insert_clean(tmp, ns - 1, tab[q]);
            int32_t *sw = tab;
            tab = tmp;
            tmp = sw;
            mask = ns - 1;
          }
          return;
        }
      }
      perturb >>= 5;
      i = (int)(((unsigned)i * 5u + 1u + perturb) & (unsigned)mask);
    }
  };
  for (int index = 0; index < n; ++index) {
    if (!mm[index]) continue;
    for (int i = 1; i <= radius; ++i) {
      if (index - i >= 0) add(index - i);
      if (index + i <= n - 1) add(index + i);
    }
  }
  for (int q = 0; q <= mask; ++q)
    if (tab[q] != -1) cand[ncand++] = tab[q];
  return ncand;
}

// partner of every paired position, both directions (DotBracket.find_base_pairs, rlif/rna/dotbracket.py:94-111:
// unbalanced closers are skipped, unmatched openers have no partner); one thread per sequence, the stack lives
// in the partner array's own slots of a scratch buffer
__global__ void pm_pairs_kernel(int B, const int32_t *__restrict__ off, const uint8_t *__restrict__ tgt, int32_t *partner,
                                int32_t *stack) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int o = off[b], n = off[b + 1] - o;
  int top = 0;
  for (int i = 0; i < n; ++i) partner[o + i] = -1;
  for (int i = 0; i < n; ++i) {
    const uint8_t c = tgt[o + i];
    if (c == '(') stack[o + top++] = i;
    else if (c == ')' && top > 0) {
      const int a = stack[o + --top];
      partner[o + a] = i;
      partner[o + i] = a;
    }
  }
}

// get_surrounding (solution.py:166-181) for one sequence: the candidate positions in the reference's visiting
// order (cand[0..ncand)) and their mutation probabilities
__device__ int pm_surround(int n, const uint8_t *__restrict__ mm, int32_t *cand, int32_t *tab, double *prob, int radius, double p, int order) {
  const int ncand = pm_set_order(n, mm, radius, order, tab, tab + 8 * n + 16, cand);
  const double stepw = (p - 0.01) / (double)radius;  // numpy.linspace(0.01, p, radius + 1): start + k * step
  for (int index = 0; index < n; ++index) {
    if (!mm[index]) continue;
    for (int i = 1; i <= radius; ++i) {
      prob[index] = p;
      if (index - i >= 0) prob[index - i] = 0.01 + (double)(i - 1) * stepw;
      if (index + i <= n - 1) prob[index + i] = 0.01 + (double)(radius - i) * stepw;
    }
  }
  return ncand;
}

// after the first fold: which sequences enter the search (0 < HD <= threshold), their first candidate sets
__global__ void pm_init_kernel(int B, const int32_t *__restrict__ off, const int32_t *__restrict__ hd, const uint8_t *__restrict__ mm,
                               int32_t *cand, int32_t *ncand, int32_t *tab, double *prob, uint8_t *active, int32_t *rounds,
                               PermuteCfg cfg) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int o = off[b], n = off[b + 1] - o;
  const int act = (hd[b] > 0 && hd[b] <= cfg.threshold && cfg.budget > 0) ? 1 : 0;
  active[b] = (uint8_t)act;
  rounds[b] = 0;
  if (act) ncand[b] = pm_surround(n, mm + o, cand + o, tab + 16 * (size_t)o + 32 * (size_t)b, prob + o, cfg.radius, cfg.p, cfg.order);
}

// one round's candidates: gather the best sequence (and its target) of every searching solution into the
// compacted batch, then apply the mutations of this round in the reference's set-iteration order
__global__ void pm_mutate_kernel(int count, int step, const int32_t *__restrict__ list, const int32_t *__restrict__ boff,
                                 const int32_t *__restrict__ off, const uint8_t *__restrict__ best, const uint8_t *__restrict__ tgt,
                                 const int32_t *__restrict__ partner, const int32_t *__restrict__ cand,
                                 const int32_t *__restrict__ ncand, const double *__restrict__ prob, uint8_t *bseq,
                                 uint8_t *btgt, PermuteCfg cfg) {
  const int s = blockIdx.x;
  if (s >= count) return;
  const int b = list[s], o = off[b], n = off[b + 1] - o, bo = boff[s];
  for (int q = threadIdx.x; q < n; q += blockDim.x) {
    bseq[bo + q] = best[o + q];
    btgt[bo + q] = tgt[o + q];
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const unsigned long long key = pm_mix64(cfg.seed + (cfg.keys ? cfg.keys[b] : (unsigned long long)b));
  const int nc = ncand[b];
  for (int c = 0; c < nc; ++c) {
    const int m = cand[o + c];
    const double u = (double)(pm_draw(key, step, m, 0) >> 11) * (1.0 / 9007199254740992.0);
    if (!(u < prob[o + m])) continue;
    int action = (int)((pm_draw(key, step, m, 1) >> 33) % 4ull);
    const bool paired = tgt[o + m] != '.';
    if (cfg.allow_gu && paired) action = (int)((pm_draw(key, step, m, 2) >> 33) % 6ull);
    bseq[bo + m] = "ACGUGU"[action];          // mapping {0:'AU', 1:'CG', 2:'GC', 3:'UA', 4:'GU', 5:'UG'} (solution.py:20)
    if (paired) {
      const int pr = partner[o + m];
      if (pr >= 0) bseq[bo + pr] = "UGCAUG"[action];
    }
  }
}

// keep a candidate iff its HD is strictly smaller (solution.py:206-212); new candidate set from its mismatches
__global__ void pm_accept_kernel(int count, int step, const int32_t *__restrict__ list, const int32_t *__restrict__ boff,
                                 const int32_t *__restrict__ off, const uint8_t *__restrict__ bseq, const uint8_t *__restrict__ bdb,
                                 const uint8_t *__restrict__ bmm, const int32_t *__restrict__ bhd, const int32_t *__restrict__ bmfe,
                                 uint8_t *best, uint8_t *best_db, int32_t *min_hd, int32_t *best_mfe, int32_t *cand,
                                 int32_t *ncand, int32_t *tab, double *prob, uint8_t *active, int32_t *rounds, PermuteCfg cfg) {
  const int s = blockIdx.x;
  if (s >= count) return;
  const int b = list[s], o = off[b], n = off[b + 1] - o, bo = boff[s];
  const int hd = bhd[s];
  const bool better = hd < min_hd[b];  // uniform over the block (read before anyone writes)
  __syncthreads();
  if (better) {
    for (int q = threadIdx.x; q < n; q += blockDim.x) {
      best[o + q] = bseq[bo + q];
      best_db[o + q] = bdb[bo + q];
    }
  }
  if (threadIdx.x == 0) {
    rounds[b] = step + 1;
    if (better) {
      min_hd[b] = hd;
      best_mfe[b] = bmfe[s];
      ncand[b] = pm_surround(n, bmm + bo, cand + o, tab + 16 * (size_t)o + 32 * (size_t)b, prob + o, cfg.radius, cfg.p, cfg.order);
    }
    const int cur = better ? hd : min_hd[b];
    active[b] = (uint8_t)((cur != 0 && step + 1 < cfg.budget) ? 1 : 0);
  }
}

__global__ void pm_reward_kernel(int B, const int32_t *__restrict__ off, const int32_t *__restrict__ hd, double reward_exp,
                                 double *reward) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int n = off[b + 1] - off[b];
  double r = (n > 0) ? pow(1.0 - (double)hd[b] / (double)n, reward_exp) : 0.0;  // solution.py:145-148
  if (hd[b] == 0) r += 0.25;
  reward[b] = r;
}

bool pm_is_device_ptr(const void *p) {
  if (!p) return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

// grow-only device scratch shared by all calls (the environment calls this once per terminal step)
struct Scratch {
  std::vector<void *> ptrs;
  cudaStream_t st = nullptr;
  size_t cap_total = 0, cap_b = 0;
  void release() {
    for (void *p : ptrs) cudaFree(p);
    ptrs.clear();
    cap_total = cap_b = 0;
  }
  template <typename T>
  cudaError_t get(T **p, size_t n) {
    cudaError_t e = cudaMalloc((void **)p, sizeof(T) * (n ? n : 1));
    if (e == cudaSuccess) ptrs.push_back(*p);
    return e;
  }
  uint8_t *d_best, *d_tgt, *d_bestdb, *d_mm, *d_bseq, *d_btgt, *d_bdb, *d_bmm, *d_active;
  int32_t *d_off, *d_partner, *d_stack, *d_minhd, *d_mfe, *d_rounds, *d_list, *d_boff, *d_bhd, *d_bmfe;
  int32_t *d_cand, *d_ncand, *d_tab;
  double *d_prob, *d_reward;
  cudaError_t reserve(size_t total, size_t NB) {
    cudaError_t e;
    if (!st && (e = cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking)) != cudaSuccess) return e;
    if (total <= cap_total && NB <= cap_b) return cudaSuccess;
    release();
    total += total / 4 + 64;
    NB += NB / 4 + 16;
#define PM_GET(p, n) if ((e = get(&p, n)) != cudaSuccess) { release(); return e; }
    PM_GET(d_best, total) PM_GET(d_tgt, total) PM_GET(d_bestdb, total) PM_GET(d_mm, total) PM_GET(d_bseq, total)
    PM_GET(d_btgt, total) PM_GET(d_bdb, total) PM_GET(d_bmm, total) PM_GET(d_active, NB) PM_GET(d_off, NB + 1)
    PM_GET(d_partner, total) PM_GET(d_stack, total) PM_GET(d_minhd, NB) PM_GET(d_mfe, NB) PM_GET(d_rounds, NB)
    PM_GET(d_list, NB) PM_GET(d_boff, NB + 1) PM_GET(d_bhd, NB) PM_GET(d_bmfe, NB) PM_GET(d_cand, total) PM_GET(d_ncand, NB) PM_GET(d_tab, 16 * total + 32 * NB)
    PM_GET(d_prob, total) PM_GET(d_reward, NB)
#undef PM_GET
    cap_total = total;
    cap_b = NB;
    return cudaSuccess;
  }
};
Scratch g_scratch;

std::mutex g_pm_mu;

}  // namespace

extern "C" {

int rnafold_permute_set_order(int32_t mode) {
  if (mode < 0 || mode > 2) return set_error(RNAFOLD_EINVAL, "permute order: 0 CPython >= 3.7 set, 1 CPython 3.6 set, 2 ascending");
  std::lock_guard<std::mutex> lk(g_pm_mu);
  g_pm_order = mode;
  return RNAFOLD_OK;
}

// host-only: the visiting order itself, for the CPU test-suite (mismatch flags of one sequence -> positions)
int rnafold_permute_candidate_order(const uint8_t *mismatch, int32_t n, int32_t radius, int32_t mode, int32_t *out) {
  if (!mismatch || n < 0 || radius < 1 || mode < 0 || mode > 2 || !out) return set_error(RNAFOLD_EINVAL, "bad arguments");
  std::vector<int32_t> tab(2 * (8 * (size_t)n + 16));
  return pm_set_order(n, mismatch, radius, mode, tab.data(), tab.data() + 8 * n + 16, out);
}

uint64_t rnafold_permute_draw(uint64_t seed, int32_t index, int32_t step, int32_t pos, int32_t kind) {
  return pm_draw(pm_mix64(seed + (uint64_t)(uint32_t)index), step, pos, kind);
}

int rnafold_permute_batch(uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target, int32_t budget,
                          int32_t radius, int32_t threshold, double mutation_probability, int32_t allow_gu,
                          uint64_t seed, double reward_exp, uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out,
                          double *reward_out, int32_t *rounds_out, int32_t *folds_out) {
  return rnafold::permute_batch_keyed(seq, off, B, target, budget, radius, threshold, mutation_probability, allow_gu, seed,
                                      reward_exp, db_out, mfe_dcal_out, hd_out, reward_out, rounds_out, folds_out, nullptr);
}

}  // extern "C"

// The same with the draws of sequence b keyed on d_keys[b] (device array) instead of b: the batched environment passes
// (environment index, episode), so a search does not depend on which other environments finished in the same step.
int rnafold::permute_batch_keyed(uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target, int32_t budget,
                                 int32_t radius, int32_t threshold, double mutation_probability, int32_t allow_gu,
                                 uint64_t seed, double reward_exp, uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out,
                                 double *reward_out, int32_t *rounds_out, int32_t *folds_out, const unsigned long long *d_keys) {
  std::lock_guard<std::mutex> lk(g_pm_mu);
  if (B < 0 || !off || (B > 0 && (!seq || !target))) return set_error(RNAFOLD_EINVAL, "null seq/off/target or negative batch size");
  if (pm_is_device_ptr(off)) return set_error(RNAFOLD_EINVAL, "permute: the CSR offsets must be a host array");
  if (budget < 0 || radius < 1 || radius > 64 || threshold < 0 || !(mutation_probability >= 0.0 && mutation_probability <= 1.0))
    return set_error(RNAFOLD_EINVAL, "permute: budget >= 0, 1 <= radius <= 64, threshold >= 0, 0 <= mutation_probability <= 1");
  if (folds_out) *folds_out = 0;
  if (B == 0) return RNAFOLD_OK;
  const int dev = rnafold_current_device();
  if (dev < 0) return set_error(RNAFOLD_ENOINIT, "rnafold_init was not called");
  PM_TRY(cudaSetDevice(dev));
  if (off[0] != 0) return set_error(RNAFOLD_EINVAL, "off[0] must be 0");
  for (int b = 0; b < B; ++b)
    if (off[b + 1] < off[b] || off[b + 1] - off[b] > RNAFOLD_MAX_LEN) return set_error(RNAFOLD_EINVAL, "bad sequence length");
  const size_t total = (size_t)off[B], NB = (size_t)B;
  const PermuteCfg cfg{budget, radius, threshold, allow_gu ? 1 : 0, mutation_probability, (unsigned long long)seed, d_keys, g_pm_order};

  Scratch &S = g_scratch;
  PM_TRY(S.reserve(total, NB));
  cudaStream_t st = S.st;
  uint8_t *d_best = S.d_best, *d_tgt = S.d_tgt, *d_bestdb = S.d_bestdb, *d_mm = S.d_mm, *d_bseq = S.d_bseq, *d_btgt = S.d_btgt,
          *d_bdb = S.d_bdb, *d_bmm = S.d_bmm, *d_active = S.d_active;
  int32_t *d_off = S.d_off, *d_partner = S.d_partner, *d_stack = S.d_stack, *d_minhd = S.d_minhd, *d_mfe = S.d_mfe,
          *d_rounds = S.d_rounds, *d_list = S.d_list, *d_boff = S.d_boff, *d_bhd = S.d_bhd, *d_bmfe = S.d_bmfe;
  int32_t *d_cand = S.d_cand, *d_ncand = S.d_ncand, *d_tab = S.d_tab;
  double *d_prob = S.d_prob, *d_reward = S.d_reward;

  const cudaMemcpyKind k_seq = pm_is_device_ptr(seq) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  const cudaMemcpyKind k_tgt = pm_is_device_ptr(target) ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  PM_TRY(cudaMemcpyAsync(d_best, seq, total, k_seq, st));
  PM_TRY(cudaMemcpyAsync(d_tgt, target, total, k_tgt, st));
  PM_TRY(cudaMemcpyAsync(d_off, off, sizeof(int32_t) * (NB + 1), cudaMemcpyHostToDevice, st));
  PM_TRY(cudaStreamSynchronize(st));

  // the evaluation every solution gets (solution.py:133-134): fold + HD + mismatch flags
  int folds = 0;
  int rc = rnafold_fold_hamming_batch(d_best, off, B, d_tgt, d_bestdb, d_mfe, d_minhd, nullptr, d_mm, reward_exp, (void *)st);
  if (rc) return rc;
  folds += B;
  const int tpb = 128, nblk = (B + tpb - 1) / tpb;
  pm_pairs_kernel<<<nblk, tpb, 0, st>>>(B, d_off, d_tgt, d_partner, d_stack);
  pm_init_kernel<<<nblk, tpb, 0, st>>>(B, d_off, d_minhd, d_mm, d_cand, d_ncand, d_tab, d_prob, d_active, d_rounds, cfg);
  PM_TRY(cudaGetLastError());

  std::vector<uint8_t> h_active(NB);
  std::vector<int32_t> h_list(NB), h_boff(NB + 1);
  for (int step = 0; step < budget; ++step) {
    PM_TRY(cudaMemcpyAsync(h_active.data(), d_active, NB, cudaMemcpyDeviceToHost, st));
    PM_TRY(cudaStreamSynchronize(st));
    int count = 0;
    h_boff[0] = 0;
    for (int b = 0; b < B; ++b)
      if (h_active[b]) {
        h_list[count] = b;
        h_boff[count + 1] = h_boff[count] + (off[b + 1] - off[b]);
        ++count;
      }
    if (count == 0) break;
    PM_TRY(cudaMemcpyAsync(d_list, h_list.data(), sizeof(int32_t) * (size_t)count, cudaMemcpyHostToDevice, st));
    PM_TRY(cudaMemcpyAsync(d_boff, h_boff.data(), sizeof(int32_t) * ((size_t)count + 1), cudaMemcpyHostToDevice, st));
    pm_mutate_kernel<<<count, 64, 0, st>>>(count, step, d_list, d_boff, d_off, d_best, d_tgt, d_partner, d_cand, d_ncand,
                                           d_prob, d_bseq, d_btgt, cfg);
    PM_TRY(cudaGetLastError());
    // h_list / h_boff stay untouched until the next round's synchronize, so the async copies above are safe
    rc = rnafold_fold_hamming_batch(d_bseq, h_boff.data(), count, d_btgt, d_bdb, d_bmfe, d_bhd, nullptr, d_bmm, reward_exp, (void *)st);
    if (rc) return rc;
    folds += count;
    pm_accept_kernel<<<count, 64, 0, st>>>(count, step, d_list, d_boff, d_off, d_bseq, d_bdb, d_bmm, d_bhd, d_bmfe, d_best,
                                           d_bestdb, d_minhd, d_mfe, d_cand, d_ncand, d_tab, d_prob, d_active, d_rounds, cfg);
    PM_TRY(cudaGetLastError());
  }
  pm_reward_kernel<<<nblk, tpb, 0, st>>>(B, d_off, d_minhd, reward_exp, d_reward);
  PM_TRY(cudaGetLastError());

  auto out = [&](void *user, const void *devp, size_t bytes) -> cudaError_t {
    if (!user || !bytes) return cudaSuccess;
    return cudaMemcpyAsync(user, devp, bytes, pm_is_device_ptr(user) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st);
  };
  PM_TRY(out(seq, d_best, total));
  PM_TRY(out(db_out, d_bestdb, total));
  PM_TRY(out(mfe_dcal_out, d_mfe, 4 * NB));
  PM_TRY(out(hd_out, d_minhd, 4 * NB));
  PM_TRY(out(reward_out, d_reward, 8 * NB));
  PM_TRY(out(rounds_out, d_rounds, 4 * NB));
  PM_TRY(cudaStreamSynchronize(st));
  if (folds_out) *folds_out = folds;
  return RNAFOLD_OK;
}
