// env.cu -- batched RnaDesign environment on the device (reference: rlif/environments/RLIFenv.py:6-114,
// rlif/rna/solution.py:84-155, rlif/rna/dotbracket.py:94-211).
//
// N environments advance in lock-step: one kernel applies the N actions (insert_nucleotide +
// find_next_unfilled), the environments that reached their last position claim slots of a CSR
// batch and folded by the library's own fold kernels (fold + Hamming + reward), the results are
// scattered back, finished environments are optionally re-armed with their next target (the
// auto-reset of the reference's VecEnv), and one kernel emits the [N, 10, 2k] observation windows.
// Nothing here runs on the CPU except target parsing (once per data set) and launch bookkeeping.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <stdint.h>

#include <cstdio>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rnafold_b200.h"
#include "fold_kernels.h"

namespace {

using rnafold::set_error;

#define ENV_TRY(expr)                                                                   \
  do {                                                                                  \
    cudaError_t e_ = (expr);                                                            \
    if (e_ != cudaSuccess) {                                                            \
      char buf_[256];                                                                   \
      std::snprintf(buf_, sizeof(buf_), "%s: %s", #expr, cudaGetErrorString(e_));       \
      return set_error(RNAFOLD_ECUDA, buf_);                                            \
    }                                                                                   \
  } while (0)

constexpr int kMaxDelay = 64;
#ifndef ENV_STEP_THREADS
#define ENV_STEP_THREADS 64
#endif
constexpr int kStepThreads = ENV_STEP_THREADS;   // threads of the one-block-per-environment step kernel
constexpr int kMaxFinishes = 8;   // an environment may finish this often within one window of deferred rewards
constexpr int kRows = 10;  // 6 structure-motif rows (O C I H M E) + 4 nucleotide rows (A C G U)

struct Env {
  int n_envs = 0, k = 25, n_targets = 0, lmax = 0;
  bool auto_reset = true, randomize = false;
  bool gc_penalty = false;  // reward shaping of solution.py:150-151 (config keys gc_penalty / gc)
  uint64_t seed = 0;
  // Solution.permute at the terminal evaluation (RLIFenv.py:13,69-73)
  bool permute = false;
  int pm_budget = 50, pm_radius = 1, pm_threshold = 5, pm_allow_gu = 0;
  double pm_p = 0.5;
  cudaStream_t stream = nullptr;       // where the environments' work is enqueued: own_stream, or the caller's (rnafold_env_set_stream)
  cudaStream_t own_stream = nullptr;
  // data set (device, CSR over targets)
  uint8_t *d_tgt = nullptr, *d_mot = nullptr;
  int16_t *d_pair = nullptr;
  int32_t *d_toff = nullptr;
  // per-environment state
  int32_t *d_cur = nullptr, *d_idx = nullptr, *d_ep = nullptr, *d_act = nullptr;
  uint8_t *d_seq = nullptr, *d_mask = nullptr, *d_done = nullptr;
  double *d_reward = nullptr;
  // terminal batch
  int32_t *d_list = nullptr, *d_off = nullptr, *d_hd = nullptr, *d_mfe = nullptr, *d_slot = nullptr;
  unsigned long long *d_keys = nullptr;
  uint8_t *d_bseq = nullptr, *d_btgt = nullptr, *d_bdb = nullptr;
  double *d_brw = nullptr;
  int32_t *h_off = nullptr;  // pinned: [0] = total length, [1] = count, [2..] offsets (host-routed path: local search)
  int32_t *h_status = nullptr;  // pinned: the library's sticky status word, read with the end-of-step copies
  // deferred rewards (rnafold_env_config_delay): terminal sequences of up to `delay` consecutive steps are folded together
  int delay = 1;                // requested window (1 = every step folds its own terminals)
  int min_episode = 1;          // shortest episode of the data set, in steps: a window of w steps sees <= ceil(w / min_episode) finishes per environment
  int win = 0;                  // steps accumulated in the open window
  double win_exp = 0.0;         // their reward exponent
  double *win_r[kMaxDelay] = {};   // the callers' reward buffers of those steps (device pointers or null)
  int32_t *d_pend = nullptr, *d_pend_t = nullptr;   // per batch slot of the open window: the window step it finished in / its target
  std::vector<int32_t> lens;    // distinct target lengths (what the device-routed fold must be prepared for)
  void *plan = nullptr;         // launch list of the device-routed fold (capi.cu)
  void *plan_win = nullptr;     // ... of a window of deferred rewards (more entries)
  // statistics: episodes, solved, sum_hd, steps (int64) ; sum_reward, sum_mfe_kcal (double)
  unsigned long long *d_istat = nullptr;
  double *d_fstat = nullptr;
  uint8_t *d_obs = nullptr;
  // last folded structures of each env (for get_attr('prev_solution'))
  uint8_t *d_last_db = nullptr, *d_last_seq = nullptr;
  int32_t *d_last_hd = nullptr, *d_last_mfe = nullptr, *d_last_t = nullptr;
};

std::mutex g_env_mu;
std::vector<Env *> g_envs;

// ---- host: parse one dot-bracket target (rlif/rna/dotbracket.py:94-111 and :171-211)
void parse_target(const std::string &s, std::vector<int16_t> &pair, std::vector<uint8_t> &mot) {
  const int n = (int)s.size();
  pair.assign(n, -1);
  std::vector<int> stack;
  for (int i = 0; i < n; ++i) {
    if (s[i] == '(') stack.push_back(i);
    else if (s[i] == ')' && !stack.empty()) {  // unbalanced closers are ignored like the reference's try/except
      const int b = stack.back();
      stack.pop_back();
      pair[b] = (int16_t)i;  // base_pair_indices maps opener -> closer only
    }
  }
  // motif markers: O C I H M E = 0..5; mapping {'()':H, ')(':M, '((':I, '))':I, 'X(':E, ')X':E, 'XX':E}
  mot.assign(n, 5);
  char prev = 'X';
  int run = 0;
  auto motif = [](char a, char b) -> int {
    if (a == '(' && b == ')') return 3;
    if (a == ')' && b == '(') return 4;
    if ((a == '(' && b == '(') || (a == ')' && b == ')')) return 2;
    return 5;
  };
  for (int i = 0; i <= n; ++i) {
    const char c = i < n ? s[i] : 'X';
    if (c == '.') {
      ++run;
    } else if (c == 'X') {
      const int m = motif(prev, c);
      for (int q = i - run; q < i; ++q) mot[q] = (uint8_t)m;
    } else {
      mot[i] = c == '(' ? 0 : 1;
      if (run > 0) {
        const int m = motif(prev, c);
        for (int q = i - run; q < i; ++q) mot[q] = (uint8_t)m;
      }
      run = 0;
      prev = c;
    }
  }
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long x) {
  x += 0x9e3779b97f4a7c15ull;
  x = (x ^ (x >> 30)) * 0xbf58476d1ce4e5b9ull;
  x = (x ^ (x >> 27)) * 0x94d049bb133111ebull;
  return x ^ (x >> 31);
}

// (re)arm environments: next target (round-robin per env, or hashed random), empty sequence
__global__ void env_reset_kernel(int n_envs, int n_targets, int lmax, int randomize, unsigned long long seed,
                                 const int32_t *__restrict__ list, const int32_t *__restrict__ count_ptr, int32_t *cur,
                                 int32_t *idx, int32_t *ep, uint8_t *seq, uint8_t *mask) {
  // list: the finished environments of this step (their number lives on the device); nullptr: every environment
  const int count = list ? min(*count_ptr, n_envs) : n_envs;
  for (int s = blockIdx.x; s < count; s += gridDim.x) {
    const int e = list ? list[s] : s;
    if (threadIdx.x == 0) {
      const int epn = ep[e] + 1;
      int t;
      if (randomize) t = (int)(mix64(seed ^ ((unsigned long long)e << 32) ^ (unsigned long long)epn) % (unsigned long long)n_targets);
      else t = (cur[e] + 1) % n_targets;
      cur[e] = t;
      idx[e] = 0;
      ep[e] = epn;
    }
    for (int q = threadIdx.x; q < lmax; q += blockDim.x) {
      seq[(size_t)e * lmax + q] = '-';
      mask[(size_t)e * lmax + q] = 0;
    }
  }
}

// insert_nucleotide + find_next_unfilled (solution.py:84-114) and the terminal test of RnaDesign.step
// The observation window encoding[:, i:i+2k] of the zero-padded [10, len+2k] encoding: one COLUMN per thread -- the motif code
// and the one-hot nucleotide mask of position i - k + col are loaded once and give the column's ten bytes (rows 0..5: motif ==
// row, rows 6..9: mask bits); consecutive threads write consecutive bytes of a row.  (A thread per output byte paid an integer
// division and two dependent loads for each of the 500 bytes: 1,100 warp instructions per environment.)
__device__ __forceinline__ void write_observation(uint8_t *dst, const uint8_t *__restrict__ mot_row, const uint8_t *m, int n, int i,
                                                  int k) {
  const int w = 2 * k;
  for (int col = threadIdx.x; col < w; col += blockDim.x) {
    const int pos = i - k + col;
    const bool in = pos >= 0 && pos < n;
    const int mo = in ? (int)mot_row[pos] : 255;
    const unsigned mk = in ? (unsigned)m[pos] : 0u;
#pragma unroll
    for (int row = 0; row < 6; ++row) dst[row * w + col] = (uint8_t)(mo == row);
#pragma unroll
    for (int row = 6; row < kRows; ++row) dst[row * w + col] = (uint8_t)((mk >> (row - 6)) & 1u);
  }
}

// observation window encoding[:, i:i+2k] of the zero-padded [10, len+2k] encoding (solution.py:62-75,116-123)
__global__ void env_obs_kernel(int n_envs, int lmax, int k, const int32_t *__restrict__ cur, const int32_t *__restrict__ idx,
                               const int32_t *__restrict__ toff, const uint8_t *__restrict__ mot,
                               const uint8_t *__restrict__ mask, uint8_t *obs) {
  const int e = blockIdx.x;
  if (e >= n_envs) return;
  const int t = cur[e], base = toff[t], n = toff[t + 1] - base, i = idx[e];
  write_observation(obs + (size_t)e * kRows * 2 * k, mot + base, mask + (size_t)e * lmax, n, i, k);
}

// End of a step, one block per environment: a terminal environment takes its fold results (reward, last solution,
// rollout counters), is re-armed with its next target when auto_reset is on, and every environment emits its observation
// window -- the three former kernels in one launch (an environment's observation only depends on its own state).
__global__ void env_finish_kernel(int n_envs, int n_targets, int lmax, int k, int auto_reset, int randomize, unsigned long long seed,
                                  int gc_penalty, const uint8_t *__restrict__ done, const int32_t *__restrict__ slot_of,
                                  const int32_t *__restrict__ off, const double *__restrict__ brw, const int32_t *__restrict__ bhd,
                                  const int32_t *__restrict__ bmfe, const uint8_t *__restrict__ bdb, const uint8_t *__restrict__ bseq,
                                  const int32_t *__restrict__ toff, const uint8_t *__restrict__ mot, int32_t *cur, int32_t *idx,
                                  int32_t *ep, uint8_t *seq, uint8_t *mask, double *reward, uint8_t *last_db, uint8_t *last_seq,
                                  int32_t *last_hd, int32_t *last_mfe, int32_t *last_t, unsigned long long *istat, double *fstat,
                                  uint8_t *obs, unsigned long long *packed) {
  __shared__ int n_u;
  const int e = blockIdx.x;
  if (e >= n_envs) return;
  if (e == 0 && threadIdx.x == 0) *packed = 0ull;  // nobody reads the terminal counter any more: the next step starts empty
  if (done[e]) {
    const int s = slot_of[e];
    const int t0 = cur[e];
    const int o0 = off[s], len = toff[t0 + 1] - toff[t0];
    if (threadIdx.x == 0) n_u = 0;
    __syncthreads();
    int my_u = 0;
    for (int q = threadIdx.x; q < len; q += blockDim.x) {
      const uint8_t ch = bseq[o0 + q];
      last_db[(size_t)e * lmax + q] = bdb[o0 + q];
      last_seq[(size_t)e * lmax + q] = ch;
      my_u += ch == 'U';
    }
    for (int q = len + threadIdx.x; q < lmax; q += blockDim.x) {   // nothing of an earlier, longer solution stays behind
      last_db[(size_t)e * lmax + q] = '.';
      last_seq[(size_t)e * lmax + q] = '-';
    }
    if (gc_penalty && my_u) atomicAdd(&n_u, my_u);
    __syncthreads();
    if (threadIdx.x == 0) {
      double rw = brw[s];
      if (gc_penalty) {
        // r * 0.85 + clip(gcau['U'], 0, .15); gcau['U'] accumulates one increment per U like gcau_content (solution.py:257-268)
        const double inc = 1.0 / (double)len;
        double u = 0.0;
        for (int q = 0; q < n_u; ++q) u += inc;
        rw = rw * 0.85 + fmin(fmax(u, 0.0), 0.15);
      }
      reward[e] = rw;
      last_hd[e] = bhd[s];
      last_mfe[e] = bmfe[s];
      last_t[e] = t0;
      atomicAdd(&istat[0], 1ull);
      if (bhd[s] == 0) atomicAdd(&istat[1], 1ull);
      atomicAdd(&istat[2], (unsigned long long)bhd[s]);
      atomicAdd(&fstat[0], rw);
      atomicAdd(&fstat[1], (double)bmfe[s] / 100.0);
      if (auto_reset) {   // RnaDesign.reset: next target, empty solution
        const int epn = ep[e] + 1;
        int t;
        if (randomize) t = (int)(mix64(seed ^ ((unsigned long long)e << 32) ^ (unsigned long long)epn) % (unsigned long long)n_targets);
        else t = (t0 + 1) % n_targets;
        cur[e] = t;
        idx[e] = 0;
        ep[e] = epn;
      }
    }
    if (auto_reset)
      for (int q = threadIdx.x; q < lmax; q += blockDim.x) {
        seq[(size_t)e * lmax + q] = '-';
        mask[(size_t)e * lmax + q] = 0;
      }
    __syncthreads();
  }
  const int t = cur[e], base = toff[t], n = toff[t + 1] - base, i = idx[e];
  write_observation(obs + (size_t)e * kRows * 2 * k, mot + base, mask + (size_t)e * lmax, n, i, k);
}

// ---- deferred rewards (window of several steps, one fold)
// (a) RnaDesign.step for one environment per BLOCK: the cursor update is one thread's work, the copy of a finished sequence into
// the fold batch is spread over the block.  full != 0 (deferred rewards: ONE launch per step) adds the half of
// env_finish_kernel that does not need the fold -- auto-reset and observation window -- and marks the environment as pending
__global__ void env_step_kernel(int n_envs, int n_targets, int lmax, int k, int auto_reset, int randomize, unsigned long long seed,
                                   const int32_t *__restrict__ act, const int32_t *__restrict__ toff,
                                   const uint8_t *__restrict__ tgt, const int16_t *__restrict__ pair,
                                   const uint8_t *__restrict__ mot, int32_t *cur, int32_t *idx, int32_t *ep, uint8_t *seq,
                                   uint8_t *mask, uint8_t *done, double *reward, unsigned long long *istat,
                                   unsigned long long *packed, int32_t *list, int32_t *boff, uint8_t *bseq, uint8_t *btgt,
                                   int32_t *slot_of, unsigned long long *keys, int32_t *pend, int32_t *pend_t, int jstep,
                                   uint8_t *obs, int full) {
  __shared__ int sh[7];   // done, offset in the batch, target base, length; then what the observation needs: base, length, cursor
  const int e = blockIdx.x;
  if (e >= n_envs) return;
  uint8_t *s = seq + (size_t)e * lmax, *m = mask + (size_t)e * lmax;
  if (threadIdx.x == 0) {
    const int t = cur[e];
    const int base = toff[t], n = toff[t + 1] - base;
    const int i = idx[e];
    const int a = act[e] & 3;
    m[i] |= (uint8_t)(1u << a);
    s[i] = "ACGU"[a];
    if (tgt[base + i] == '(') {
      const int p = pair[base + i];
      if (p >= 0) {
        m[p] |= (uint8_t)(1u << (3 - a));  // reverse_action: the Watson-Crick partner
        s[p] = "UGCA"[a];
      }
    }
    reward[e] = 0.0;
    int fin = 0, o0 = 0, inext = i;
    if (i < n - 1) {
      int q = i;
      while (q < n - 1 && s[q] != '-') ++q;  // stops at n-1 even if that slot is filled (reference quirk)
      idx[e] = q;
      inext = q;
    } else {
      fin = 1;
      const unsigned long long old = atomicAdd(packed, (1ull << 32) | (unsigned long long)n);
      const int slot = (int)(old >> 32);
      o0 = (int)(old & 0xffffffffull);
      list[slot] = e;
      slot_of[e] = slot;
      keys[slot] = ((unsigned long long)(unsigned)e << 32) | (unsigned long long)(unsigned)ep[e];
      boff[slot] = o0;
      if (full) {   // deferred rewards: which step of the window this batch entry belongs to, and its target
        pend[slot] = jstep;
        pend_t[slot] = t;
      }
    }
    done[e] = (uint8_t)fin;
    if (e == 0) atomicAdd(&istat[3], (unsigned long long)n_envs);
    sh[0] = fin;
    sh[1] = o0;
    sh[2] = base;
    sh[3] = n;
    sh[4] = base;   // (replaced below if the environment is reset)
    sh[5] = n;
    sh[6] = inext;
  }
  __syncthreads();
  if (sh[0]) {
    const int o0 = sh[1], base = sh[2], n = sh[3];
    for (int q = threadIdx.x; q < n; q += blockDim.x) {
      bseq[o0 + q] = s[q];
      btgt[o0 + q] = tgt[base + q];
    }
    if (auto_reset && full) {
      __syncthreads();   // the copy reads the solution that is cleared next
      if (threadIdx.x == 0) {   // RnaDesign.reset: next target, empty solution
        const int t0 = cur[e], epn = ep[e] + 1;
        int t;
        if (randomize) t = (int)(mix64(seed ^ ((unsigned long long)e << 32) ^ (unsigned long long)epn) % (unsigned long long)n_targets);
        else t = (t0 + 1) % n_targets;
        cur[e] = t;
        idx[e] = 0;
        ep[e] = epn;
        sh[4] = toff[t];
        sh[5] = toff[t + 1] - sh[4];
        sh[6] = 0;
      }
      for (int q = threadIdx.x; q < lmax; q += blockDim.x) {
        s[q] = '-';
        m[q] = 0;
      }
    }
    __syncthreads();
  }
  if (!full) return;   // step-synchronous rewards: env_finish_kernel resets and writes the observation after the fold
  const int base = sh[4], n = sh[5], i = sh[6];   // handed over in shared memory: no second chain of global loads
  write_observation(obs + (size_t)e * kRows * 2 * k, mot + base, m, n, i, k);
}

// (b) at the end of a window, after the fold of everything that finished in it: rewards into the reward buffer of the step
// each environment finished in, last solution, rollout statistics
struct WindowRewards {
  double *r[kMaxDelay];
};
__global__ void env_rewards_kernel(int cap, int lmax, int gc_penalty, const int32_t *__restrict__ list,
                                   const int32_t *__restrict__ pend, const int32_t *__restrict__ pend_t,
                                   const int32_t *__restrict__ slot_of, const int32_t *__restrict__ off,
                                   const double *__restrict__ brw, const int32_t *__restrict__ bhd,
                                   const int32_t *__restrict__ bmfe, const uint8_t *__restrict__ bdb,
                                   const uint8_t *__restrict__ bseq, const int32_t *__restrict__ toff, WindowRewards out,
                                   uint8_t *last_db, uint8_t *last_seq, int32_t *last_hd, int32_t *last_mfe, int32_t *last_t,
                                   unsigned long long *istat, double *fstat, const unsigned long long *packed) {
  // one block per batch entry (slot) of the window; an environment may own several, the last one is its "last solution"
  __shared__ int n_u;
  const int s = blockIdx.x;
  if (s >= cap || s >= (int)(*packed >> 32)) return;
  const int e = list[s], t0 = pend_t[s];
  const int o0 = off[s], len = toff[t0 + 1] - toff[t0];
  const bool latest = slot_of[e] == s;
  if (threadIdx.x == 0) n_u = 0;
  __syncthreads();
  int my_u = 0;
  for (int q = threadIdx.x; q < len; q += blockDim.x) {
    const uint8_t ch = bseq[o0 + q];
    if (latest) {
      last_db[(size_t)e * lmax + q] = bdb[o0 + q];
      last_seq[(size_t)e * lmax + q] = ch;
    }
    my_u += ch == 'U';
  }
  if (latest)
    for (int q = len + threadIdx.x; q < lmax; q += blockDim.x) {   // nothing of an earlier, longer solution stays behind
      last_db[(size_t)e * lmax + q] = '.';
      last_seq[(size_t)e * lmax + q] = '-';
    }
  if (gc_penalty && my_u) atomicAdd(&n_u, my_u);
  __syncthreads();
  if (threadIdx.x == 0) {
    double rw = brw[s];
    if (gc_penalty) {
      const double inc = 1.0 / (double)len;
      double u = 0.0;
      for (int q = 0; q < n_u; ++q) u += inc;
      rw = rw * 0.85 + fmin(fmax(u, 0.0), 0.15);
    }
    double *dst = out.r[pend[s]];
    if (dst) dst[e] = rw;
    if (latest) {
      last_hd[e] = bhd[s];
      last_mfe[e] = bmfe[s];
      last_t[e] = t0;
    }
    atomicAdd(&istat[0], 1ull);
    if (bhd[s] == 0) atomicAdd(&istat[1], 1ull);
    atomicAdd(&istat[2], (unsigned long long)bhd[s]);
    atomicAdd(&fstat[0], rw);
    atomicAdd(&fstat[1], (double)bmfe[s] / 100.0);
  }
}

Env *get_env(int h) {
  if (h < 0 || h >= (int)g_envs.size()) return nullptr;
  return g_envs[h];
}

template <typename T>
cudaError_t dalloc(T **p, size_t n) {
  return cudaMalloc((void **)p, sizeof(T) * (n ? n : 1));
}

}  // namespace

static bool env_is_device_ptr(const void *p) {
  if (!p) return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

static cudaError_t copy_out(void *user, const void *dev, size_t bytes, cudaStream_t st) {
  if (!user || !bytes) return cudaSuccess;
  return cudaMemcpyAsync(user, dev, bytes, env_is_device_ptr(user) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, st);
}

static void env_free(Env *e) {
  void *ptrs[] = {e->d_tgt, e->d_mot, e->d_pair, e->d_toff, e->d_cur, e->d_idx, e->d_ep, e->d_act, e->d_seq, e->d_mask,
                  e->d_done, e->d_reward, e->d_list, e->d_off, e->d_hd, e->d_mfe, e->d_slot, e->d_keys, e->d_bseq, e->d_btgt, e->d_bdb,
                  e->d_brw, e->d_istat, e->d_fstat, e->d_obs, e->d_last_db, e->d_last_seq, e->d_last_hd,
                  e->d_last_mfe, e->d_last_t, e->d_pend, e->d_pend_t};
  for (void *p : ptrs)
    if (p) cudaFree(p);
  if (e->h_off) cudaFreeHost(e->h_off);
  if (e->h_status) cudaFreeHost(e->h_status);
  if (e->plan) rnafold::fold_dyn_plan_free(e->plan);
  if (e->plan_win) rnafold::fold_dyn_plan_free(e->plan_win);
  if (e->own_stream) cudaStreamDestroy(e->own_stream);
  delete e;
}

// Close the open window of deferred rewards: ONE device-routed fold of everything that finished in it, then the rewards
// into the reward buffers of the steps they belong to.  Called with g_env_mu held.
static int env_window_end(Env *e) {
  if (e->win == 0) return RNAFOLD_OK;
  cudaStream_t st = e->stream;
  const int N = e->n_envs;
  // entries a window of this many steps can hold at most; only wave-length targets can finish more than once in it
  // (a cluster-length episode takes more steps than the longest window), so the cluster workspace stays sized for N
  const int per_env = std::min(kMaxFinishes, (e->win + e->min_episode - 1) / e->min_episode);
  const int cap = N * std::max(1, per_env);
  int rc = rnafold::fold_hamming_dyn_enqueue(&e->plan_win, e->lens.data(), (int)e->lens.size(), cap, N, e->d_bseq, e->d_off + 2,
                                             e->d_off + 1, e->d_off, e->d_btgt, e->d_bdb, e->d_mfe, e->d_hd, e->d_brw, e->win_exp,
                                             (void *)st);
  WindowRewards out{};
  for (int j = 0; j < e->win; ++j) out.r[j] = e->win_r[j];
  e->win = 0;
  if (rc) {
    cudaMemsetAsync(e->d_off, 0, sizeof(unsigned long long), st);
    return rc;
  }
  env_rewards_kernel<<<cap, 128, 0, st>>>(cap, e->lmax, e->gc_penalty ? 1 : 0, e->d_list, e->d_pend, e->d_pend_t, e->d_slot,
                                          e->d_off + 2, e->d_brw, e->d_hd, e->d_mfe, e->d_bdb, e->d_bseq, e->d_toff, out,
                                          e->d_last_db, e->d_last_seq, e->d_last_hd, e->d_last_mfe, e->d_last_t, e->d_istat,
                                          e->d_fstat, reinterpret_cast<const unsigned long long *>(e->d_off));
  ENV_TRY(cudaGetLastError());
  ENV_TRY(cudaMemsetAsync(e->d_off, 0, sizeof(unsigned long long), st));   // the next window starts empty
  return RNAFOLD_OK;
}

// episode length (steps) of one target: the cursor walk of env_step_kernel on an empty solution
static int episode_steps(const std::string &s, const std::vector<int16_t> &pair) {
  const int n = (int)s.size();
  std::vector<char> filled((size_t)n, 0);
  int i = 0, steps = 0;
  for (;;) {
    ++steps;
    filled[i] = 1;
    if (s[i] == '(' && pair[i] >= 0) filled[pair[i]] = 1;
    if (i >= n - 1) return steps;
    int q = i;
    while (q < n - 1 && filled[q]) ++q;
    i = q;
  }
}

extern "C" {

// Deferred rewards for rollouts that keep one reward buffer per step: the terminal sequences of up to `steps` consecutive
// steps are folded by ONE call (a step is bound by the latency of a fold, not by its throughput), and every reward lands in
// the buffer that was passed to the step it belongs to -- at the latest when the window closes: after `steps` steps, on
// rnafold_env_flush, or before any call that reads or changes the environments' state.  Observations, dones and auto-resets
// are not deferred.  The window never exceeds the shortest episode of the data set (no environment finishes twice in it);
// steps with host output buffers or with the local search run undeferred.
int rnafold_env_config_delay(int32_t h, int32_t steps) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  if (steps < 1) return set_error(RNAFOLD_EINVAL, "reward delay must be at least 1 step");
  int rc = env_window_end(e);
  e->delay = steps > kMaxDelay ? kMaxDelay : steps;
  return rc;
}

// Enqueue the environments' work on the CALLER's stream from now on (e.g. torch.cuda.current_stream().cuda_stream, so that
// a policy network and the step are ordered without host synchronisation); use_own != 0 goes back to the environments' own
// non-blocking stream.  Work already enqueued on the previous stream is waited for on the new one.
int rnafold_env_set_stream(int32_t h, void *stream, int32_t use_own) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  // a null handle is the legacy default stream (torch's default stream reports 0): named explicitly, because the fold entry
  // points read a null stream as "the library's own"
  cudaStream_t next = use_own ? e->own_stream : (stream ? (cudaStream_t)stream : cudaStreamLegacy);
  if (next == e->stream) return RNAFOLD_OK;
  cudaEvent_t ev;
  ENV_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  cudaError_t err = cudaEventRecord(ev, e->stream);
  if (err == cudaSuccess) err = cudaStreamWaitEvent(next, ev, 0);
  cudaEventDestroy(ev);
  ENV_TRY(err);
  e->stream = next;
  return RNAFOLD_OK;
}

int rnafold_env_flush(int32_t h) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  return env_window_end(e);
}

int rnafold_env_create(int32_t n_envs, int32_t kernel_size, int32_t *handle) {
  if (n_envs <= 0 || kernel_size <= 0 || kernel_size > 512 || !handle) return set_error(RNAFOLD_EINVAL, "bad env arguments");
  const int dev = rnafold_current_device();
  if (dev < 0) return set_error(RNAFOLD_ENOINIT, "rnafold_init was not called");
  ENV_TRY(cudaSetDevice(dev));
  Env *e = new Env();
  e->n_envs = n_envs;
  e->k = kernel_size;
  const size_t N = (size_t)n_envs;
  ENV_TRY(cudaStreamCreateWithFlags(&e->own_stream, cudaStreamNonBlocking));
  e->stream = e->own_stream;
  ENV_TRY(dalloc(&e->d_cur, N));
  ENV_TRY(dalloc(&e->d_idx, N));
  ENV_TRY(dalloc(&e->d_ep, N));
  ENV_TRY(dalloc(&e->d_act, N));
  ENV_TRY(dalloc(&e->d_done, N));
  ENV_TRY(dalloc(&e->d_reward, N));
  const size_t NE = N * kMaxFinishes;   // batch entries of one window
  ENV_TRY(dalloc(&e->d_list, NE));
  ENV_TRY(dalloc(&e->d_slot, N));
  ENV_TRY(dalloc(&e->d_pend, NE));
  ENV_TRY(dalloc(&e->d_pend_t, NE));
  ENV_TRY(dalloc(&e->d_keys, NE));
  ENV_TRY(dalloc(&e->d_off, NE + 4));
  ENV_TRY(cudaMemset(e->d_off, 0, (N + 4) * sizeof(int32_t)));
  ENV_TRY(dalloc(&e->d_hd, NE));
  ENV_TRY(dalloc(&e->d_mfe, NE));
  ENV_TRY(dalloc(&e->d_brw, NE));
  ENV_TRY(dalloc(&e->d_last_hd, N));
  ENV_TRY(dalloc(&e->d_last_mfe, N));
  ENV_TRY(dalloc(&e->d_last_t, N));
  ENV_TRY(dalloc(&e->d_istat, 4));
  ENV_TRY(dalloc(&e->d_fstat, 2));
  ENV_TRY(dalloc(&e->d_obs, N * kRows * 2 * (size_t)kernel_size));
  ENV_TRY(cudaMallocHost((void **)&e->h_off, sizeof(int32_t) * (N + 4)));
  ENV_TRY(cudaMallocHost((void **)&e->h_status, sizeof(int32_t)));
  *e->h_status = 0;
  ENV_TRY(cudaMemset(e->d_istat, 0, 4 * sizeof(unsigned long long)));
  ENV_TRY(cudaMemset(e->d_fstat, 0, 2 * sizeof(double)));
  ENV_TRY(cudaMemset(e->d_cur, 0xff, N * sizeof(int32_t)));  // current_sequence = -1 (RLIFenv.py:25)
  ENV_TRY(cudaMemset(e->d_ep, 0, N * sizeof(int32_t)));
  ENV_TRY(cudaMemset(e->d_last_hd, 0, N * sizeof(int32_t)));
  ENV_TRY(cudaMemset(e->d_last_mfe, 0, N * sizeof(int32_t)));
  ENV_TRY(cudaMemset(e->d_last_t, 0xff, N * sizeof(int32_t)));
  std::lock_guard<std::mutex> lk(g_env_mu);
  g_envs.push_back(e);
  *handle = (int32_t)g_envs.size() - 1;
  return RNAFOLD_OK;
}

int rnafold_env_destroy(int32_t h) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  env_window_end(e);
  cudaStreamSynchronize(e->stream);
  env_free(e);
  g_envs[h] = nullptr;
  return RNAFOLD_OK;
}

int rnafold_env_config(int32_t h, int32_t auto_reset, int32_t randomize, uint64_t seed) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  e->auto_reset = auto_reset != 0;
  e->randomize = randomize != 0;
  e->seed = seed;
  return RNAFOLD_OK;
}

// env.set_attr('current_sequence', v) (rlif/learning/tester.py:63): every environment's round-robin position;
// the next reset moves on to (v + 1) % n_targets (RLIFenv.py:89-90)
int rnafold_env_set_current(int32_t h, int32_t value) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e || !e->n_targets) return set_error(RNAFOLD_EINVAL, "bad env handle or no targets set");
  if (value < -1 || value >= e->n_targets) return set_error(RNAFOLD_EINVAL, "current_sequence out of range");
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  std::vector<int32_t> v((size_t)e->n_envs, value);
  ENV_TRY(cudaMemcpyAsync(e->d_cur, v.data(), sizeof(int32_t) * v.size(), cudaMemcpyHostToDevice, e->stream));
  ENV_TRY(cudaStreamSynchronize(e->stream));
  return RNAFOLD_OK;
}

int rnafold_env_config_reward(int32_t h, int32_t gc_penalty) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  e->gc_penalty = gc_penalty != 0;
  return RNAFOLD_OK;
}

int rnafold_env_config_permute(int32_t h, int32_t enabled, int32_t budget, int32_t radius, int32_t threshold,
                               double mutation_probability, int32_t allow_gu) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e) return set_error(RNAFOLD_EINVAL, "bad env handle");
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  if (enabled && (budget < 0 || radius < 1 || radius > 64 || threshold < 0 || !(mutation_probability >= 0.0 && mutation_probability <= 1.0)))
    return set_error(RNAFOLD_EINVAL, "bad permutation settings");
  e->permute = enabled != 0;
  if (enabled) {
    e->pm_budget = budget;
    e->pm_radius = radius;
    e->pm_threshold = threshold;
    e->pm_p = mutation_probability;
    e->pm_allow_gu = allow_gu ? 1 : 0;
  }
  return RNAFOLD_OK;
}

// RnaDesign.set_data (RLIFenv.py:106-109): install a data set of dot-bracket targets; every environment
// starts over at current_sequence = -1 and takes its first target.
int rnafold_env_set_targets(int32_t h, const uint8_t *bytes, const int32_t *off, int32_t n_targets) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e || !bytes || !off || n_targets <= 0) return set_error(RNAFOLD_EINVAL, "bad env targets");
  {
    int rc = env_window_end(e);
    if (rc) return rc;
    ENV_TRY(cudaStreamSynchronize(e->stream));
  }
  std::vector<int16_t> pair_all;
  std::vector<uint8_t> mot_all;
  int lmax = 0, min_episode = 1 << 30;
  for (int t = 0; t < n_targets; ++t) {
    const int n = off[t + 1] - off[t];
    if (n <= 0 || n > RNAFOLD_MAX_LEN) return set_error(RNAFOLD_EINVAL, "target length out of range");
    std::string s((const char *)bytes + off[t], (size_t)n);
    for (char c : s)
      if (c != '(' && c != ')' && c != '.') return set_error(RNAFOLD_EINVAL, "targets must be dot-bracket strings");
    std::vector<int16_t> pr;
    std::vector<uint8_t> mo;
    parse_target(s, pr, mo);
    min_episode = std::min(min_episode, episode_steps(s, pr));
    pair_all.insert(pair_all.end(), pr.begin(), pr.end());
    mot_all.insert(mot_all.end(), mo.begin(), mo.end());
    lmax = n > lmax ? n : lmax;
  }
  const size_t total = (size_t)off[n_targets], N = (size_t)e->n_envs;
  for (void *p : {(void *)e->d_tgt, (void *)e->d_mot, (void *)e->d_pair, (void *)e->d_toff, (void *)e->d_seq, (void *)e->d_mask,
                  (void *)e->d_bseq, (void *)e->d_btgt, (void *)e->d_bdb, (void *)e->d_last_db, (void *)e->d_last_seq})
    if (p) cudaFree(p);
  ENV_TRY(dalloc(&e->d_tgt, total));
  ENV_TRY(dalloc(&e->d_mot, total));
  ENV_TRY(dalloc(&e->d_pair, total));
  ENV_TRY(dalloc(&e->d_toff, (size_t)n_targets + 1));
  ENV_TRY(dalloc(&e->d_seq, N * lmax));
  ENV_TRY(dalloc(&e->d_mask, N * lmax));
  // bases of a window's batch: the first episode an environment finishes in it may be as long as lmax, every further one ran
  // entirely inside the window, and an episode of L bases takes at least L / 2 steps
  const size_t NB = N * ((size_t)lmax + 2 * kMaxDelay) + 16;
  ENV_TRY(dalloc(&e->d_bseq, NB));
  ENV_TRY(dalloc(&e->d_btgt, NB));
  ENV_TRY(dalloc(&e->d_bdb, NB));
  ENV_TRY(dalloc(&e->d_last_db, N * lmax));
  ENV_TRY(dalloc(&e->d_last_seq, N * lmax));
  ENV_TRY(cudaMemcpy(e->d_tgt, bytes, total, cudaMemcpyHostToDevice));
  ENV_TRY(cudaMemcpy(e->d_mot, mot_all.data(), total, cudaMemcpyHostToDevice));
  ENV_TRY(cudaMemcpy(e->d_pair, pair_all.data(), total * sizeof(int16_t), cudaMemcpyHostToDevice));
  ENV_TRY(cudaMemcpy(e->d_toff, off, sizeof(int32_t) * ((size_t)n_targets + 1), cudaMemcpyHostToDevice));
  ENV_TRY(cudaMemset(e->d_last_db, '.', N * lmax));
  ENV_TRY(cudaMemset(e->d_last_seq, '-', N * lmax));
  e->n_targets = n_targets;
  e->lmax = lmax;
  e->min_episode = min_episode;
  ENV_TRY(cudaMemset(e->d_cur, 0xff, N * sizeof(int32_t)));
  ENV_TRY(cudaMemset(e->d_ep, 0xff, N * sizeof(int32_t)));  // the constructor's next_target_structure() is not an episode
  e->lens.clear();
  for (int t = 0; t < n_targets; ++t) {
    const int32_t n = off[t + 1] - off[t];
    if (std::find(e->lens.begin(), e->lens.end(), n) == e->lens.end()) e->lens.push_back(n);
  }
  if (e->plan) rnafold::fold_dyn_plan_free(e->plan);
  if (e->plan_win) rnafold::fold_dyn_plan_free(e->plan_win);
  e->plan = e->plan_win = nullptr;
  env_reset_kernel<<<e->n_envs, 64, 0, e->stream>>>(e->n_envs, e->n_targets, e->lmax, e->randomize ? 1 : 0, e->seed, nullptr,
                                                     nullptr, e->d_cur, e->d_idx, e->d_ep, e->d_seq, e->d_mask);
  ENV_TRY(cudaGetLastError());
  ENV_TRY(cudaStreamSynchronize(e->stream));
  return RNAFOLD_OK;
}

static int env_emit_obs(Env *e, uint8_t *obs_out) {
  // a device destination is written by the kernel itself (no 2 MB device-to-device copy per step)
  const bool direct = obs_out && env_is_device_ptr(obs_out);
  env_obs_kernel<<<e->n_envs, 128, 0, e->stream>>>(e->n_envs, e->lmax, e->k, e->d_cur, e->d_idx, e->d_toff, e->d_mot,
                                                    e->d_mask, direct ? obs_out : e->d_obs);
  ENV_TRY(cudaGetLastError());
  if (!direct) ENV_TRY(copy_out(obs_out, e->d_obs, (size_t)e->n_envs * kRows * 2 * e->k, e->stream));
  return RNAFOLD_OK;
}

// RnaDesign.reset for every environment (RLIFenv.py:34-57): next target, empty solution, first observation
int rnafold_env_reset(int32_t h, uint8_t *obs_out) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e || !e->n_targets) return set_error(RNAFOLD_EINVAL, "bad env handle or no targets set");
  ENV_TRY(cudaSetDevice(rnafold_current_device()));
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  env_reset_kernel<<<e->n_envs, 64, 0, e->stream>>>(e->n_envs, e->n_targets, e->lmax, e->randomize ? 1 : 0, e->seed, nullptr,
                                                     nullptr, e->d_cur, e->d_idx, e->d_ep, e->d_seq, e->d_mask);
  ENV_TRY(cudaGetLastError());
  ENV_TRY(cudaMemsetAsync(e->d_off, 0, sizeof(unsigned long long), e->stream));
  int rc = env_emit_obs(e, obs_out);
  if (rc) return rc;
  ENV_TRY(cudaStreamSynchronize(e->stream));
  return RNAFOLD_OK;
}

// RnaDesign.step for every environment (RLIFenv.py:59-81) + the VecEnv auto-reset when enabled.
//
// The whole step is enqueued without a host round trip: the terminal environments compact themselves into a CSR
// batch on the device, a routing kernel bins that batch by length class, the fold kernels read their counts from
// device memory (CTAs without work exit), and scatter / reset / observation follow on the same stream.  The host
// synchronises ONCE, at the end, and only when an output is a host buffer; with device outputs the call returns
// as soon as the work is enqueued.  (The local search, permute=True, sizes its rounds on the host and keeps the
// host-routed path.)
int rnafold_env_step(int32_t h, const int32_t *actions, double reward_exp, uint8_t *obs_out, double *reward_out,
                     uint8_t *done_out) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e || !e->n_targets || !actions) return set_error(RNAFOLD_EINVAL, "bad env handle, no targets set or null actions");
  ENV_TRY(cudaSetDevice(rnafold_current_device()));
  const int N = e->n_envs;
  cudaStream_t st = e->stream;
  const int32_t *d_act = actions;
  if (!env_is_device_ptr(actions)) {
    ENV_TRY(cudaMemcpyAsync(e->d_act, actions, sizeof(int32_t) * (size_t)N, cudaMemcpyHostToDevice, st));
    d_act = e->d_act;
  }
  {
    // ---- deferred rewards: device outputs only, no local search
    const int window = std::min(e->delay, kMaxFinishes * e->min_episode);
    const bool dev_out = (!obs_out || env_is_device_ptr(obs_out)) && (!reward_out || env_is_device_ptr(reward_out)) &&
                         (!done_out || env_is_device_ptr(done_out));
    if (window > 1 && dev_out && !e->permute && e->auto_reset) {   // (without auto-reset a finished environment finishes again every step)
      if (e->win > 0 && e->win_exp != reward_exp) {
        int rc = env_window_end(e);
        if (rc) return rc;
      }
      uint8_t *d_done = done_out ? done_out : e->d_done;
      double *d_rw = reward_out ? reward_out : e->d_reward;
      env_step_kernel<<<N, kStepThreads, 0, st>>>(N, e->n_targets, e->lmax, e->k, 1, e->randomize ? 1 : 0, e->seed, d_act, e->d_toff, e->d_tgt,
                                            e->d_pair, e->d_mot, e->d_cur, e->d_idx, e->d_ep, e->d_seq, e->d_mask, d_done, d_rw,
                                            e->d_istat, reinterpret_cast<unsigned long long *>(e->d_off), e->d_list, e->d_off + 2,
                                            e->d_bseq, e->d_btgt, e->d_slot, e->d_keys, e->d_pend, e->d_pend_t, e->win,
                                            obs_out ? obs_out : e->d_obs, 1);
      ENV_TRY(cudaGetLastError());
      e->win_r[e->win] = reward_out;
      e->win_exp = reward_exp;
      if (++e->win >= window) return env_window_end(e);
      return RNAFOLD_OK;
    }
    int rc = env_window_end(e);   // an undeferred step closes whatever is open first
    if (rc) return rc;
  }
  // d_off: [0] total length, [1] count (one 64-bit word advanced atomically by the terminal environments), [2..] offsets
  // (the terminal counter is zero here: cleared by the last kernel of the previous step, by reset(), and on error returns)
  env_step_kernel<<<N, kStepThreads, 0, st>>>(N, e->n_targets, e->lmax, e->k, 0, 0, e->seed, d_act, e->d_toff, e->d_tgt, e->d_pair, e->d_mot,
                                        e->d_cur, e->d_idx, e->d_ep, e->d_seq, e->d_mask, e->d_done, e->d_reward, e->d_istat,
                                        reinterpret_cast<unsigned long long *>(e->d_off), e->d_list, e->d_off + 2, e->d_bseq,
                                        e->d_btgt, e->d_slot, e->d_keys, nullptr, nullptr, 0, nullptr, 0);
  ENV_TRY(cudaGetLastError());
  int rc = RNAFOLD_OK;
  if (e->permute) {
    // evaluate(permute=True): fold, local search on the near misses, results of the best sequence (solution.py:133-141).
    // The count and the first offsets travel together (1 KB): a second round trip only for large terminal batches.
    const int prefetch = N + 3 < 259 ? N + 3 : 259;
    ENV_TRY(cudaMemcpyAsync(e->h_off, e->d_off, sizeof(int32_t) * (size_t)prefetch, cudaMemcpyDeviceToHost, st));
    ENV_TRY(cudaStreamSynchronize(st));
    const int count = e->h_off[1];
    if (count > 0) {
      if (count + 2 > prefetch) {
        ENV_TRY(cudaMemcpyAsync(e->h_off + prefetch, e->d_off + prefetch, sizeof(int32_t) * (size_t)(count + 2 - prefetch),
                                cudaMemcpyDeviceToHost, st));
        ENV_TRY(cudaStreamSynchronize(st));
      }
      int32_t *boff = e->h_off + 2;   // CSR offsets of the terminal batch (slot order = arrival order)
      boff[count] = e->h_off[0];
      // draws keyed on (environment, episode): reproducible from the seed whatever else finishes in the same step
      rc = rnafold::permute_batch_keyed(e->d_bseq, boff, count, e->d_btgt, e->pm_budget, e->pm_radius, e->pm_threshold, e->pm_p,
                                        e->pm_allow_gu, e->seed, reward_exp, e->d_bdb, e->d_mfe, e->d_hd, e->d_brw, nullptr,
                                        nullptr, e->d_keys);
      if (rc) {
        cudaMemsetAsync(e->d_off, 0, sizeof(unsigned long long), st);
        return rc;
      }
    }
  } else {
    rc = rnafold::fold_hamming_dyn_enqueue(&e->plan, e->lens.data(), (int)e->lens.size(), N, N, e->d_bseq, e->d_off + 2,
                                           e->d_off + 1, e->d_off, e->d_btgt, e->d_bdb, e->d_mfe, e->d_hd, e->d_brw,
                                           reward_exp, (void *)st);
    if (rc) {
      cudaMemsetAsync(e->d_off, 0, sizeof(unsigned long long), st);
      return rc;
    }
  }
  {
    // scatter + auto-reset + observation in one launch; a device destination is written by the kernel itself
    const bool direct = obs_out && env_is_device_ptr(obs_out);
    env_finish_kernel<<<N, 128, 0, st>>>(N, e->n_targets, e->lmax, e->k, e->auto_reset ? 1 : 0, e->randomize ? 1 : 0, e->seed,
                                         e->gc_penalty ? 1 : 0, e->d_done, e->d_slot, e->d_off + 2, e->d_brw, e->d_hd, e->d_mfe,
                                         e->d_bdb, e->d_bseq, e->d_toff, e->d_mot, e->d_cur, e->d_idx, e->d_ep, e->d_seq, e->d_mask,
                                         e->d_reward, e->d_last_db, e->d_last_seq, e->d_last_hd, e->d_last_mfe, e->d_last_t,
                                         e->d_istat, e->d_fstat, direct ? obs_out : e->d_obs,
                                         reinterpret_cast<unsigned long long *>(e->d_off));
    ENV_TRY(cudaGetLastError());
    if (!direct) ENV_TRY(copy_out(obs_out, e->d_obs, (size_t)N * kRows * 2 * e->k, st));
  }
  ENV_TRY(copy_out(reward_out, e->d_reward, sizeof(double) * (size_t)N, st));
  ENV_TRY(copy_out(done_out, e->d_done, (size_t)N, st));
  const bool host_out = (obs_out && !env_is_device_ptr(obs_out)) || (reward_out && !env_is_device_ptr(reward_out)) ||
                        (done_out && !env_is_device_ptr(done_out));
  if (host_out) {
    // the one synchronisation of the step; the library's sticky status word (a backtrack that found no
    // decomposition) is read with it
    ENV_TRY(cudaMemcpyAsync(e->h_status, rnafold::status_word(), sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    ENV_TRY(cudaStreamSynchronize(st));
    if (*e->h_status) {
      char buf[96];
      std::snprintf(buf, sizeof(buf), "backtrack found no decomposition (code %d)", *e->h_status);
      return set_error(RNAFOLD_ECUDA, buf);
    }
  }
  return RNAFOLD_OK;
}

// rollout statistics since the last clear: istat = {episodes, solved (HD == 0), sum HD, env steps},
// fstat = {sum reward, sum MFE (kcal/mol)}
int rnafold_env_stats(int32_t h, uint64_t istat[4], double fstat[2], int32_t clear) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e || !istat || !fstat) return set_error(RNAFOLD_EINVAL, "bad env handle");
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  ENV_TRY(cudaMemcpyAsync(istat, e->d_istat, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost, e->stream));
  ENV_TRY(cudaMemcpyAsync(fstat, e->d_fstat, 2 * sizeof(double), cudaMemcpyDeviceToHost, e->stream));
  if (clear) {
    ENV_TRY(cudaMemsetAsync(e->d_istat, 0, 4 * sizeof(uint64_t), e->stream));
    ENV_TRY(cudaMemsetAsync(e->d_fstat, 0, 2 * sizeof(double), e->stream));
  }
  ENV_TRY(cudaStreamSynchronize(e->stream));
  return RNAFOLD_OK;
}

// state snapshot for get_attr: current sequences ('-' = unfilled), cursors, target indices, episode counters
// and the last evaluated solution of every environment (sequence, folded structure, HD, MFE dcal, target)
int rnafold_env_snapshot(int32_t h, uint8_t *seq, int32_t *idx, int32_t *cur, int32_t *ep, uint8_t *last_seq,
                         uint8_t *last_db, int32_t *last_hd, int32_t *last_mfe, int32_t *last_t, int32_t *lmax_out) {
  std::lock_guard<std::mutex> lk(g_env_mu);
  Env *e = get_env(h);
  if (!e || !e->n_targets) return set_error(RNAFOLD_EINVAL, "bad env handle or no targets set");
  {
    int rc_ = env_window_end(e);   // pending deferred rewards first
    if (rc_) return rc_;
  }
  const size_t N = (size_t)e->n_envs, NL = N * (size_t)e->lmax;
  cudaStream_t st = e->stream;
  ENV_TRY(copy_out(seq, e->d_seq, NL, st));
  ENV_TRY(copy_out(idx, e->d_idx, 4 * N, st));
  ENV_TRY(copy_out(cur, e->d_cur, 4 * N, st));
  ENV_TRY(copy_out(ep, e->d_ep, 4 * N, st));
  ENV_TRY(copy_out(last_seq, e->d_last_seq, NL, st));
  ENV_TRY(copy_out(last_db, e->d_last_db, NL, st));
  ENV_TRY(copy_out(last_hd, e->d_last_hd, 4 * N, st));
  ENV_TRY(copy_out(last_mfe, e->d_last_mfe, 4 * N, st));
  ENV_TRY(copy_out(last_t, e->d_last_t, 4 * N, st));
  if (lmax_out) *lmax_out = e->lmax;
  ENV_TRY(cudaStreamSynchronize(st));
  return RNAFOLD_OK;
}

// host-only: the reference's target parsing (base pairs + motif markers) for the CPU test-suite
int rnafold_parse_target(const uint8_t *dotbracket, int32_t n, int16_t *pair_out, uint8_t *motif_out) {
  if (!dotbracket || n < 0 || !pair_out || !motif_out) return set_error(RNAFOLD_EINVAL, "bad arguments");
  std::vector<int16_t> pr;
  std::vector<uint8_t> mo;
  parse_target(std::string((const char *)dotbracket, (size_t)n), pr, mo);
  std::memcpy(pair_out, pr.data(), sizeof(int16_t) * (size_t)n);
  std::memcpy(motif_out, mo.data(), (size_t)n);
  return RNAFOLD_OK;
}

}  // extern "C"
