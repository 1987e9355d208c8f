// capi.cu -- C ABI (include/rnafold_b200.h) over the CUDA kernels: context, parameter upload,
// host/device buffer handling, length bucketing and kernel routing.
//
// Product code.  There is no CPU fold path in this library: every fold runs on the device, and
// every entry point fails with RNAFOLD_ECUDA when no usable device is present.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <numeric>
#include <string>
#include <vector>

#include "../../include/rnafold_b200.h"
#include "fold_kernels.h"
#include "fold_params.h"

extern "C" const char rnafold_builtin_turner2004[];  // builtin_params.cpp (generated at build time)
extern "C" const unsigned long rnafold_builtin_turner2004_len;

namespace {

using namespace rnafold;

thread_local std::string g_err;
int fail(int code, const char *fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  std::vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CUDA_TRY(expr)                                                                              \
  do {                                                                                              \
    cudaError_t e_ = (expr);                                                                        \
    if (e_ != cudaSuccess) return fail(RNAFOLD_ECUDA, "%s: %s", #expr, cudaGetErrorString(e_));     \
  } while (0)

// grow-only device / pinned buffers
struct DevBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    want += want / 4;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};
struct PinBuf {
  void *p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    size_t want = std::max(bytes, (size_t)4096);
    want += want / 4;
    cudaError_t e = cudaMallocHost(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() {
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
  }
};

struct Context {
  int device = -1;
  int num_sms = 0;
  int max_smem_optin = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  // ev_done: end of the last fold call's device work (the next call, on whatever stream, waits for it before it
  // touches the shared scratch); ev_staged: the last call's H2D copies out of the pinned staging blocks
  cudaEvent_t ev_done = nullptr, ev_staged = nullptr;
  // device-routed calls run their two kernel families side by side: the wavefront launches go to stream2 between a fork
  // and a join event (few sequences per family: they occupy disjoint SMs)
  cudaStream_t stream2 = nullptr, stream3 = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join3 = nullptr;
  bool have_done = false, have_staged = false;
  int32_t *d_status = nullptr;   // sticky: set by a kernel whose backtrack found no decomposition
  uint64_t params_epoch = 0;     // bumped by every parameter upload (device-routed plans are rebuilt)
  std::map<uint64_t, int> occ_cache;
  HostParams hp;
  FoldParams *d_params = nullptr;
  // model details (RNA.cvar)
  double md_temperature = 37.0, md_dangles = 2, md_noGU = 0, md_no_closingGU = 0, md_uniq_ML = 0;
  // device staging for host-side callers
  DevBuf d_seq, d_off, d_order, d_target, d_db, d_mfe, d_hd, d_reward, d_mismatch, d_ctl, d_ws, d_ws2, d_redo, d_dyn;
  PinBuf h_in, h_out, h_off;
  std::vector<int32_t> off_host, order_host;
  int32_t last_launches = 0, last_routed[3] = {0, 0, 0};
  bool timed = false;
};

Context *g_ctx = nullptr;
std::mutex g_mu;

// ---- scratch of rnafold_fold_hamming_batch_noalloc: lives in the library image, never cudaMalloc'd.
// Layout (int32 words): [0] queue counter of the fold kernel, [1] queue counter of the int32 re-run, [2] number of
// sequences the launches may fold (written by the validation kernel: B, or 0 when a sequence is longer than nmax),
// [3] first entry (always 0), [4..7] unused, [8 .. 8 + RNAFOLD_NOALLOC_MAX_BATCH) range-check flags, one per sequence.
// In HOST code the name below is only a stand-in for the symbol; its device address is resolved once, at set-up,
// with cudaGetSymbolAddress (g_noalloc_ptr) -- kernels are always handed that pointer.
constexpr int kNoallocHeader = 8;
constexpr size_t kNoallocWords = (size_t)kNoallocHeader + RNAFOLD_NOALLOC_MAX_BATCH;
static __device__ __align__(256) int32_t g_noalloc_scratch[kNoallocWords];
int32_t *g_noalloc_ptr = nullptr;        // device address of g_noalloc_scratch on the context's device
cudaError_t g_noalloc_err = cudaSuccess;  // outcome of the address lookup

// every sequence must fit the launch's length class (the fold kernels size their shared memory by nmax): otherwise
// nothing is folded and the sticky status word reports code 100
__global__ void noalloc_validate_kernel(const int32_t *__restrict__ off, int B, int nmax, int32_t *scratch, int32_t *status) {
  __shared__ int bad;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    const int n = off[b + 1] - off[b];
    if (n < 0 || n > nmax) bad = 1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    scratch[2] = bad ? 0 : B;
    scratch[3] = 0;
    if (bad) atomicExch(status, 100);
  }
}

bool is_device_ptr(const void *p) {
  if (!p) return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

bool is_pinned_host_ptr(const void *p) {
  // page-locked host memory (cudaMallocHost / cudaHostRegister / torch pin_memory): DMA-able in place
  if (!p) return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

int upload_params(Context *c) {
  CUDA_TRY(cudaDeviceSynchronize());  // folds in flight on any stream still read the live tables
  CUDA_TRY(cudaMemcpyAsync(c->d_params, &c->hp.dev, sizeof(FoldParams), cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(wave_upload_constants(c->hp, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  c->params_epoch++;
  return RNAFOLD_OK;
}

void init_host_params(HostParams *hp) {
  std::memset(hp, 0, sizeof(*hp));
  hp->lxc = 107.856;
  hp->dev.s.max_ninio = 300;
}

// Length classes: a launch is sized (shared memory, hence CTAs/SM) by the longest sequence in it,
// so mixed batches are split into classes.  Sequences that fit the fast wavefront kernel (int16
// tables in shared memory) go there; longer ones use the int32 CTA kernel with global tables.
const int kWaveClasses[] = {32, 48, 64, 80, 100, 128, 160, 200, 256};
const int kNumWaveClasses = (int)(sizeof(kWaveClasses) / sizeof(int));
const int kLegacySmemLimit = 232;

const int kClusterClasses[] = {288, 320, 352, 400, 432, 464};
const int kNumClusterClasses = (int)(sizeof(kClusterClasses) / sizeof(int));
const size_t kClusterWorkspaceCap = (size_t)24 << 30;  // bytes of global tables per cluster launch

enum { kKindWave = 0, kKindLegacySmem = 1, kKindLegacyGlobal = 2, kKindCluster = 3 };
struct Launch {
  int begin, end;  // slice of the length-sorted order (host-planned path); end - begin = upper bound otherwise
  int nmax;
  int kind;
  bool wide;       // wave kernel: latency-bound launch (fewer sequences than SMs) -> 512-thread CTAs
};

// what the live parameter set and the device allow
struct Limits {
  int wave_esz;       // int16 tables unless the parameter set's terms are too large
  int wave_limit;     // longest sequence of the single-CTA wavefront kernel
  int cluster_limit;  // longest sequence of the 2-CTA cluster kernel (= wave_limit when unavailable)
};
Limits make_limits(Context *c) {
  Limits m;
  m.wave_esz = wave_int16_ok(c->hp) ? 2 : 4;
  m.wave_limit = 0;
  for (int k = 0; k < kNumWaveClasses; ++k)
    if (wave_smem_bytes(kWaveClasses[k], m.wave_esz) + 1024 <= (size_t)c->max_smem_optin) m.wave_limit = kWaveClasses[k];
  m.cluster_limit = m.wave_limit;
  const int wave_keep = m.wave_limit;
  if (m.wave_esz == 2)
    for (int q = 0; q < kNumClusterClasses; ++q)
      if (cluster_smem_bytes(kClusterClasses[q], 2) + 1024 <= (size_t)c->max_smem_optin) m.cluster_limit = kClusterClasses[q];
  // experiments only: RNAFOLD_WAVE_LIMIT=<n> sends everything longer than n to the cluster kernel
  static const int forced_limit = getenv("RNAFOLD_WAVE_LIMIT") ? atoi(getenv("RNAFOLD_WAVE_LIMIT")) : 0;
  if (forced_limit > 0 && m.cluster_limit > wave_keep) m.wave_limit = std::min(m.wave_limit, forced_limit);
  return m;
}

// occupancy queries are CUDA API calls: cached per (kernel family, length class, element size, threads)
int cached_occ(Context *c, int family, int nmax, int esz, int threads) {
  const uint64_t key = ((uint64_t)family << 48) | ((uint64_t)nmax << 24) | ((uint64_t)esz << 16) | (uint64_t)threads;
  auto it = c->occ_cache.find(key);
  if (it != c->occ_cache.end()) return it->second;
  int v = 0;
  if (family == 0) v = wave_occupancy(nmax, esz, threads);
  else if (family == 1) v = cluster_max_active(nmax);
  else v = cta_kernel_occupancy(nmax, esz != 0);
  c->occ_cache[key] = v;
  return v;
}

// Issue the kernels of one launch record.  `count` is the number of sequences of the launch when the host knows
// it, otherwise (a.dyn != nullptr: the count lives on the device) an upper bound that sizes grids and workspaces.
// q: two queue counters (main kernel, range-check re-run).
int issue_launch(Context *c, const Limits &lim, const Launch &L, int count, BatchArgs a, int32_t *q, cudaStream_t st,
                 bool with_rerun = true, DevBuf *wsbuf = nullptr) {
  DevBuf &ws_buf = wsbuf ? *wsbuf : c->d_ws;   // global tables of this launch (launches that overlap need their own)
  auto legacy = [&](BatchArgs la, bool smem_tables, int32_t *queue) -> int {
    const int occ = std::max(1, cached_occ(c, 2, L.nmax, smem_tables ? 1 : 0, kCtaThreads));
    const int grid = std::min(count, c->num_sms * occ);
    int *ws = nullptr;
    long long stride = 0;
    if (!smem_tables) {
      stride = 2LL * table_cells(L.nmax);
      CUDA_TRY(ws_buf.reserve(sizeof(int32_t) * (size_t)stride * (size_t)grid));
      ws = (int *)ws_buf.p;
    }
    CUDA_TRY(launch_fold_cta(c->d_params, la, L.nmax, smem_tables, grid, ws, stride, queue, st));
    c->last_launches++;
    return RNAFOLD_OK;
  };
  a.redo_filter = 0;
  if (L.kind == kKindWave) {
    // A launch that cannot even fill the SMs is latency-bound per fold: give each sequence a wider CTA
    // (the per-diagonal items spread over more warps).  RNAFOLD_SMALL_THREADS overrides for experiments.
    int threads = wave_threads(L.nmax);
    if (L.wide) {
      static const int forced = getenv("RNAFOLD_SMALL_THREADS") ? atoi(getenv("RNAFOLD_SMALL_THREADS")) : 512;
      if (forced == 256 || forced == 512) threads = std::max(threads, forced);
    }
    const int occ = std::max(1, cached_occ(c, 0, L.nmax, lim.wave_esz, threads));
    const int grid = std::min(count, c->num_sms * occ);
    CUDA_TRY(launch_fold_wave(c->d_params, a, L.nmax, lim.wave_esz, grid, threads, q, st));
    c->last_launches++;
    c->last_routed[0] += count;
    if (lim.wave_esz == 2 && with_rerun) {
      // sequences whose values left the int16 range (flagged on the device) are re-run in int32
      a.redo_filter = 1;
      return legacy(a, L.nmax <= kLegacySmemLimit, q + 1);
    }
  } else if (L.kind == kKindCluster) {
    const size_t stride = cluster_table_stride(L.nmax);
    CUDA_TRY(ws_buf.reserve(2 * stride * sizeof(int16_t) * (size_t)count));
    int16_t *cbg = (int16_t *)ws_buf.p, *fg = cbg + stride * (size_t)count;
    const int ncl = std::max(1, std::min(count, cached_occ(c, 1, L.nmax, 2, 0)));
    CUDA_TRY(launch_fold_cluster(c->d_params, a, L.nmax, ncl, q, cbg, fg, st));
    CUDA_TRY(launch_f5_backtrack(c->d_params, a, L.nmax, count, cbg, fg, st));
    c->last_launches += 2;
    c->last_routed[1] += count;
    if (!with_rerun) return RNAFOLD_OK;
    a.redo_filter = 1;  // int16 range violations -> int32 kernel with global tables
    return legacy(a, false, q + 1);
  } else {
    int rc = legacy(a, L.kind == kKindLegacySmem, q);
    if (rc) return rc;
    c->last_routed[L.kind == kKindLegacySmem ? 0 : 2] += count;
  }
  return RNAFOLD_OK;
}

// Calls may arrive on different streams while sharing the context's scratch (order, queue counters, redo flags,
// workspace): every call first waits, on its stream, for the previous call's device work.
int enter_call(Context *c, cudaStream_t st) {
  if (c->have_done) CUDA_TRY(cudaStreamWaitEvent(st, c->ev_done, 0));
  return RNAFOLD_OK;
}
int leave_call(Context *c, cudaStream_t st) {
  CUDA_TRY(cudaEventRecord(c->ev_done, st));
  c->have_done = true;
  return RNAFOLD_OK;
}
// the pinned staging blocks are rewritten by the host: the previous call's H2D copies out of them must be over
int staging_ready(Context *c) {
  if (c->have_staged) CUDA_TRY(cudaEventSynchronize(c->ev_staged));
  c->have_staged = false;
  return RNAFOLD_OK;
}

// sticky device status word -> error (and clear); the stream must have been synchronised past the D2H of `status`
int status_result(Context *c, int32_t status, cudaStream_t st) {
  if (!status) return RNAFOLD_OK;
  cudaMemsetAsync(c->d_status, 0, sizeof(int32_t), st);
  return fail(RNAFOLD_ECUDA, "backtrack found no decomposition (code %d)", status);
}

int run_batch(Context *c, const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target,
              uint8_t *db_out, int32_t *mfe_out, int32_t *hd_out, double *reward_out,
              uint8_t *mismatch_out, double reward_exp, void *stream_v, bool defer_sync = false) {
  if (B < 0 || !off || (B > 0 && !seq)) return fail(RNAFOLD_EINVAL, "null seq/off or negative batch size");
  c->last_launches = 0;
  c->last_routed[0] = c->last_routed[1] = c->last_routed[2] = 0;
  if (B == 0) return RNAFOLD_OK;
  CUDA_TRY(cudaSetDevice(c->device));

  const bool off_dev = is_device_ptr(off);
  const bool seq_dev = is_device_ptr(seq);
  const bool tgt_dev = is_device_ptr(target);
  const bool any_host_out = (db_out && !is_device_ptr(db_out)) || (mfe_out && !is_device_ptr(mfe_out)) ||
                            (hd_out && !is_device_ptr(hd_out)) || (reward_out && !is_device_ptr(reward_out)) ||
                            (mismatch_out && !is_device_ptr(mismatch_out));
  cudaStream_t st = stream_v ? (cudaStream_t)stream_v : c->stream;
  {
    int rc = enter_call(c, st);
    if (rc) return rc;
  }

  // offsets on the host (needed for routing)
  c->off_host.resize((size_t)B + 1);
  if (off_dev) {
    CUDA_TRY(cudaMemcpyAsync(c->off_host.data(), off, sizeof(int32_t) * ((size_t)B + 1), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
  } else {
    std::memcpy(c->off_host.data(), off, sizeof(int32_t) * ((size_t)B + 1));
  }
  const int32_t *oh = c->off_host.data();
  if (oh[0] != 0) return fail(RNAFOLD_EINVAL, "off[0] must be 0");
  int nmax_all = 0;
  for (int b = 0; b < B; ++b) {
    const int n = oh[b + 1] - oh[b];
    if (n < 0 || n > RNAFOLD_MAX_LEN) return fail(RNAFOLD_EINVAL, "sequence %d has length %d (limit %d)", b, n, RNAFOLD_MAX_LEN);
    nmax_all = std::max(nmax_all, n);
  }
  const size_t total = (size_t)oh[B];

  // processing order: longest first (cubic cost), stable
  c->order_host.resize((size_t)B);
  std::iota(c->order_host.begin(), c->order_host.end(), 0);
  std::stable_sort(c->order_host.begin(), c->order_host.end(),
                   [&](int x, int y) { return oh[x + 1] - oh[x] > oh[y + 1] - oh[y]; });
  auto len_at = [&](int k) { return oh[c->order_host[k] + 1] - oh[c->order_host[k]]; };

  // split into launches
  const Limits lim = make_limits(c);
  std::vector<Launch> launches;
  {
    const int wave_limit = lim.wave_limit, cluster_limit = lim.cluster_limit;
    // longest first: beyond every shared-memory design -> int32 CTA kernel with global tables
    int k = 0;
    while (k < B && len_at(k) > cluster_limit) ++k;
    if (k > 0) launches.push_back(Launch{0, k, len_at(0), len_at(0) <= kLegacySmemLimit ? kKindLegacySmem : kKindLegacyGlobal, false});
    // 2-CTA cluster kernel (tables split over DSMEM) for lengths above the single-CTA kernel's reach: ONE launch of the
    // longest one's class for all of them (the kernel runs one CTA per SM whatever the class, so separate classes would
    // buy no occupancy, only one launch tail each; the queue is sorted longest first), cut only by the workspace cap.
    {
      int e_all = k;
      while (e_all < B && len_at(e_all) > wave_limit) ++e_all;
      if (e_all > k) {
        int q0 = 0;
        while (q0 + 1 < kNumClusterClasses && kClusterClasses[q0] < len_at(k)) ++q0;
        const size_t per_seq = 2 * cluster_table_stride(kClusterClasses[q0]) * sizeof(int16_t);
        const int chunk = (int)std::max<size_t>(1, kClusterWorkspaceCap / per_seq);
        for (int s0 = k; s0 < e_all; s0 += chunk)
          launches.push_back(Launch{s0, std::min(e_all, s0 + chunk), kClusterClasses[q0], kKindCluster, false});
        k = e_all;
      }
    }
    if (k < B && B - k <= c->num_sms && !getenv("RNAFOLD_NO_MERGE")) {
      // fewer sequences than SMs: one launch of the longest sequence's class (every CTA has an SM to itself,
      // so the step costs one fold latency instead of one per length class)
      launches.push_back(Launch{k, B, std::max(len_at(k), 8), kKindWave, true});
      k = B;
    }
    for (int cls = kNumWaveClasses - 1; cls >= 0 && k < B; --cls) {
      if (kWaveClasses[cls] > wave_limit) continue;
      const int lo = cls > 0 ? kWaveClasses[cls - 1] : -1;
      int e = k;
      while (e < B && len_at(e) > lo) ++e;
      if (e > k) launches.push_back(Launch{k, e, std::max(len_at(k), 8), kKindWave, e - k <= c->num_sms});
      k = e;
    }
  }

  // ---- device views of the inputs
  {
    int rc = staging_ready(c);
    if (rc) return rc;
  }
  const uint8_t *d_seq = seq;
  const int32_t *d_off = off;
  const uint8_t *d_target = target;
  if (!seq_dev) {
    CUDA_TRY(c->d_seq.reserve(total + 16));
    const void *src = seq;
    if (!is_pinned_host_ptr(seq)) {  // pageable memory is staged through the library's pinned block
      CUDA_TRY(c->h_in.reserve(2 * total + 32));
      std::memcpy(c->h_in.p, seq, total);
      src = c->h_in.p;
    }
    CUDA_TRY(cudaMemcpyAsync(c->d_seq.p, src, total, cudaMemcpyHostToDevice, st));
    d_seq = (const uint8_t *)c->d_seq.p;
  }
  if (target && !tgt_dev) {
    CUDA_TRY(c->d_target.reserve(total + 16));
    const void *src = target;
    if (!is_pinned_host_ptr(target)) {
      CUDA_TRY(c->h_in.reserve(2 * total + 32));
      uint8_t *stage = (uint8_t *)c->h_in.p + total + 16;
      std::memcpy(stage, target, total);
      src = stage;
    }
    CUDA_TRY(cudaMemcpyAsync(c->d_target.p, src, total, cudaMemcpyHostToDevice, st));
    d_target = (const uint8_t *)c->d_target.p;
  }
  {
    // offsets + order travel together from one pinned block
    CUDA_TRY(c->h_off.reserve(sizeof(int32_t) * (2 * (size_t)B + 2)));
    int32_t *ho = (int32_t *)c->h_off.p;
    std::memcpy(ho, oh, sizeof(int32_t) * ((size_t)B + 1));
    std::memcpy(ho + B + 1, c->order_host.data(), sizeof(int32_t) * (size_t)B);
    CUDA_TRY(c->d_order.reserve(sizeof(int32_t) * (2 * (size_t)B + 2)));
    CUDA_TRY(cudaMemcpyAsync(c->d_order.p, ho, sizeof(int32_t) * (2 * (size_t)B + 1), cudaMemcpyHostToDevice, st));
    if (!off_dev) d_off = (const int32_t *)c->d_order.p;
  }
  CUDA_TRY(cudaEventRecord(c->ev_staged, st));
  c->have_staged = true;
  const int32_t *d_order = (const int32_t *)c->d_order.p + B + 1;

  // ---- device views of the outputs
  auto out_view = [&](void *user, DevBuf &buf, size_t bytes, void **dev) -> cudaError_t {
    if (!user) {
      *dev = nullptr;
      return cudaSuccess;
    }
    if (is_device_ptr(user)) {
      *dev = user;
      return cudaSuccess;
    }
    cudaError_t e = buf.reserve(bytes);
    *dev = buf.p;
    return e;
  };
  void *v_db, *v_mfe, *v_hd, *v_rw, *v_mm;
  CUDA_TRY(out_view(db_out, c->d_db, total + 16, &v_db));
  CUDA_TRY(out_view(mfe_out, c->d_mfe, sizeof(int32_t) * (size_t)B, &v_mfe));
  CUDA_TRY(out_view(target ? hd_out : nullptr, c->d_hd, sizeof(int32_t) * (size_t)B, &v_hd));
  CUDA_TRY(out_view(target ? reward_out : nullptr, c->d_reward, sizeof(double) * (size_t)B, &v_rw));
  CUDA_TRY(out_view(target ? mismatch_out : nullptr, c->d_mismatch, total + 16, &v_mm));

  // ---- control words: two queue counters per launch (+ range-check re-run); redo flags
  const int nl = (int)launches.size();
  CUDA_TRY(c->d_ctl.reserve(sizeof(int32_t) * (size_t)(2 * nl)));
  CUDA_TRY(cudaMemsetAsync(c->d_ctl.p, 0, sizeof(int32_t) * (size_t)(2 * nl), st));
  int32_t *d_ctl = (int32_t *)c->d_ctl.p;
  CUDA_TRY(c->d_redo.reserve(sizeof(int32_t) * (size_t)B));
  CUDA_TRY(cudaMemsetAsync(c->d_redo.p, 0, sizeof(int32_t) * (size_t)B, st));

  if (c->timed) CUDA_TRY(cudaEventRecord(c->ev0, st));
  // Length classes overlap: the wavefront launches go to an auxiliary stream beside the cluster /
  // global-table launches on the caller's stream, so that the tail of one persistent launch is filled by the next
  // (RNAFOLD_OVERLAP=0 serialises them on one stream; 3: two auxiliary streams -- measured on C5: 237.8 k folds/s serial,
  // 252.8 k with one auxiliary stream (the default), 250.7 k with two)
  static const int overlap_env = getenv("RNAFOLD_OVERLAP") ? atoi(getenv("RNAFOLD_OVERLAP")) : 2;
  int n_w = 0, n_o = 0;
  for (const Launch &L : launches) (L.kind == kKindWave ? n_w : n_o)++;
  const bool two = overlap_env != 0 && n_w > 0 && (n_o > 0 || n_w > 1);
  const bool three = two && overlap_env >= 3 && n_w > 1;
  if (two) {
    CUDA_TRY(cudaEventRecord(c->ev_fork, st));
    CUDA_TRY(cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
    if (three) CUDA_TRY(cudaStreamWaitEvent(c->stream3, c->ev_fork, 0));
  }
  int wave_no = 0;
  for (int li = 0; li < nl; ++li) {
    const Launch &L = launches[li];
    BatchArgs a;
    a.seq = d_seq;
    a.off = d_off;
    a.order = d_order + L.begin;
    a.B = L.end - L.begin;
    a.target = d_target;
    a.db = (uint8_t *)v_db;
    a.mfe = (int32_t *)v_mfe;
    a.hd = (int32_t *)v_hd;
    a.reward = (double *)v_rw;
    a.mismatch = (uint8_t *)v_mm;
    a.reward_exp = reward_exp;
    a.status = c->d_status;
    a.redo = (int32_t *)c->d_redo.p;
    a.redo_filter = 0;
    a.dyn = nullptr;
    const bool on2 = two && L.kind == kKindWave;
    cudaStream_t ls = !on2 ? st : (three && (wave_no++ & 1)) ? c->stream3 : c->stream2;
    // (only the 256-nt class re-runs with global tables: one user of d_ws2 at a time)
    int rc = issue_launch(c, lim, L, a.B, a, d_ctl + 2 * li, ls, true, on2 ? &c->d_ws2 : nullptr);
    if (rc) return rc;
  }
  if (two) {
    CUDA_TRY(cudaEventRecord(c->ev_join, c->stream2));
    CUDA_TRY(cudaStreamWaitEvent(st, c->ev_join, 0));
    if (three) {
      CUDA_TRY(cudaEventRecord(c->ev_join3, c->stream3));
      CUDA_TRY(cudaStreamWaitEvent(st, c->ev_join3, 0));
    }
  }
  if (c->timed) CUDA_TRY(cudaEventRecord(c->ev1, st));

  // ---- copy back for host-side callers
  if (any_host_out) {
    size_t need = 0;
    auto slot = [&](void *user, size_t bytes) -> size_t {
      if (!user || is_device_ptr(user)) return 0;
      size_t o = need;
      need += (bytes + 15) / 16 * 16;
      return o;
    };
    const size_t o_db = slot(db_out, total), o_mfe = slot(mfe_out, 4 * (size_t)B),
                 o_hd = slot(target ? hd_out : nullptr, 4 * (size_t)B),
                 o_rw = slot(target ? reward_out : nullptr, 8 * (size_t)B),
                 o_mm = slot(target ? mismatch_out : nullptr, total);
    CUDA_TRY(c->h_out.reserve(need + 16));
    uint8_t *ho = (uint8_t *)c->h_out.p;
    auto back = [&](void *user, void *dev, size_t off_b, size_t bytes) -> cudaError_t {
      if (!user || is_device_ptr(user) || bytes == 0) return cudaSuccess;
      // pinned user buffers receive the DMA directly; pageable ones go through the pinned block
      return cudaMemcpyAsync(is_pinned_host_ptr(user) ? user : (void *)(ho + off_b), dev, bytes, cudaMemcpyDeviceToHost, st);
    };
    CUDA_TRY(back(db_out, v_db, o_db, total));
    CUDA_TRY(back(mfe_out, v_mfe, o_mfe, 4 * (size_t)B));
    CUDA_TRY(back(target ? hd_out : nullptr, v_hd, o_hd, 4 * (size_t)B));
    CUDA_TRY(back(target ? reward_out : nullptr, v_rw, o_rw, 8 * (size_t)B));
    CUDA_TRY(back(target ? mismatch_out : nullptr, v_mm, o_mm, total));
    int32_t status = 0;
    CUDA_TRY(cudaMemcpyAsync(&status, c->d_status, 4, cudaMemcpyDeviceToHost, st));
    {
      int rc = leave_call(c, st);
      if (rc) return rc;
    }
    CUDA_TRY(cudaStreamSynchronize(st));
    auto fin = [&](void *user, size_t off_b, size_t bytes) {
      if (user && !is_device_ptr(user) && !is_pinned_host_ptr(user) && bytes) std::memcpy(user, ho + off_b, bytes);
    };
    fin(db_out, o_db, total);
    fin(mfe_out, o_mfe, 4 * (size_t)B);
    fin(target ? hd_out : nullptr, o_hd, 4 * (size_t)B);
    fin(target ? reward_out : nullptr, o_rw, 8 * (size_t)B);
    fin(target ? mismatch_out : nullptr, o_mm, total);
    return status_result(c, status, st);
  }
  {
    int rc = leave_call(c, st);
    if (rc) return rc;
  }
  if ((!seq_dev || (target && !tgt_dev)) && !defer_sync) {
    // host inputs are consumed when the call returns (a caller may rewrite them at once)
    CUDA_TRY(cudaStreamSynchronize(st));
  }
  return RNAFOLD_OK;
}

// ---------------------------------------------------------------------------------------------
// Device-routed batches: the batch (CSR on the device) and its SIZE are produced by a kernel and never cross to
// the host.  The host only knows which lengths can occur (the environment's target lengths) and an upper bound
// of the count: it enqueues one routing kernel that bins the sequences by length class, then one launch per class
// that can occur; every kernel reads its own count from device memory and CTAs without work exit.
// ---------------------------------------------------------------------------------------------
constexpr int kMaxDynLaunches = 24;
struct RouteTable {
  int nl;                         // launches
  int merged;                     // index of the latency launch that takes a batch of <= num_sms sequences whole, or -1
  int num_sms;
  int first_class;                // first class launch (the classes follow in ASCENDING order of `upper`)
  int wfirst, wlast;              // wave class launches [wfirst, wlast] (-1: none)
  int cfirst, clast, cmax;        // cluster class launches [cfirst, clast]; cmax: clusters resident at once in clast's class
                                  // (a batch with <= cmax cluster-length sequences sends them all to launch clast)
  int ccap;                       // most cluster-length sequences the caller promised (the cluster workspace is sized for it)
  int upper[kMaxDynLaunches];     // class launches: longest sequence routed to launch k
};

__global__ void __launch_bounds__(1024) route_kernel(const int32_t *__restrict__ count_ptr, const int32_t *__restrict__ total_ptr,
                                                     int32_t *off, int cap, RouteTable rt, int32_t *__restrict__ order,
                                                     int32_t *__restrict__ dyn, int32_t *__restrict__ ctl, int nctl,
                                                     int32_t *__restrict__ redo, int32_t *status) {
  __shared__ int cnt[kMaxDynLaunches], beg[kMaxDynLaunches], fill[kMaxDynLaunches];
  const int tid = threadIdx.x;
  const int count = min(max(*count_ptr, 0), cap);
  const int total = total_ptr ? *total_ptr : 0;
  if (tid < kMaxDynLaunches) cnt[tid] = fill[tid] = 0;
  // queue counters of every launch and this batch's range-check flags start at zero (no separate memsets)
  for (int k = tid; k < nctl; k += blockDim.x) ctl[k] = 0;
  for (int s = tid; s < count; s += blockDim.x) redo[s] = 0;
  if (tid == 0) {   // entry rt.nl: the whole batch (the single int32 re-run launch walks all of it)
    dyn[2 * rt.nl] = count;
    dyn[2 * rt.nl + 1] = 0;
  }
  __syncthreads();
  auto cls_of = [&](int s) -> int {
    const int n = ((s + 1 == count && total_ptr) ? total : off[s + 1]) - off[s];
    int k = rt.first_class;
    while (k + 1 < rt.nl && n > rt.upper[k]) ++k;
    return k;
  };
  __shared__ int merge_w, merge_c;
  for (int s = tid; s < count; s += blockDim.x) atomicAdd(&cnt[cls_of(s)], 1);
  __syncthreads();
  if (tid == 0) {
    // Few sequences of a kernel family (no more than the device runs at once): ONE launch of the family's longest class,
    // so that the call costs one fold latency per family instead of one per length class.
    merge_w = merge_c = 0;
    if (rt.cfirst >= 0) {
      // more cluster-length sequences than the caller's bound (and the workspace) allows: fold NOTHING of this batch and say so
      int nc = 0;
      for (int k = rt.cfirst; k <= rt.clast; ++k) nc += cnt[k];
      if (nc > rt.ccap) {
        for (int k = 0; k < rt.nl; ++k) cnt[k] = 0;
        dyn[2 * rt.nl] = 0;
        atomicExch(status, 101);
      }
    }
    if (rt.merged >= 0 && rt.wfirst >= 0) {
      int nw = 0;
      for (int k = rt.wfirst; k <= rt.wlast; ++k) nw += cnt[k];
      if (nw <= rt.num_sms) {
        merge_w = 1;
        for (int k = rt.wfirst; k <= rt.wlast; ++k) cnt[k] = 0;
        cnt[rt.merged] = nw;
      }
    }
    if (rt.cfirst >= 0 && rt.clast > rt.cfirst) {
      int nc = 0;
      for (int k = rt.cfirst; k <= rt.clast; ++k) nc += cnt[k];
      if (nc <= rt.cmax) {
        merge_c = 1;
        for (int k = rt.cfirst; k < rt.clast; ++k) cnt[k] = 0;
        cnt[rt.clast] = nc;
      }
    }
    int o = 0;
    for (int k = 0; k < rt.nl; ++k) {
      beg[k] = o;
      o += cnt[k];
      dyn[2 * k] = cnt[k];
      dyn[2 * k + 1] = beg[k];
    }
  }
  __syncthreads();
  for (int s = tid; s < count; s += blockDim.x) {
    int k = cls_of(s);
    if (merge_w && k >= rt.wfirst && k <= rt.wlast) k = rt.merged;
    else if (merge_c && k >= rt.cfirst && k < rt.clast) k = rt.clast;
    order[beg[k] + atomicAdd(&fill[k], 1)] = s;
  }
  __syncthreads();
  if (tid == 0 && total_ptr) off[count] = total;  // closes the CSR (the producers wrote off[0..count-1])
}

struct DynPlanImpl {
  uint64_t epoch = ~0ull;
  int cap = 0, ccap = 0, lmax = 0;
  Limits lim;
  std::vector<Launch> launches;
  RouteTable rt;
};

int build_dyn_plan(Context *c, DynPlanImpl *p, const int32_t *lengths, int nlen, int cap, int ccap) {
  p->lim = make_limits(c);
  p->cap = cap;
  p->ccap = ccap;
  p->launches.clear();
  const Limits &lim = p->lim;
  int lmax = 0;
  for (int k = 0; k < nlen; ++k) lmax = std::max(lmax, lengths[k]);
  p->lmax = lmax;
  auto any_in = [&](int lo, int hi) {  // a possible length in (lo, hi]
    for (int k = 0; k < nlen; ++k)
      if (lengths[k] > lo && lengths[k] <= hi) return true;
    return false;
  };
  auto top_in = [&](int lo, int hi) {
    int top = 8;
    for (int k = 0; k < nlen; ++k)
      if (lengths[k] > lo && lengths[k] <= hi) top = std::max(top, lengths[k]);
    return top;
  };
  RouteTable &rt = p->rt;
  std::memset(&rt, 0, sizeof(rt));
  rt.merged = -1;
  rt.cfirst = rt.clast = rt.wfirst = rt.wlast = -1;
  rt.num_sms = c->num_sms;
  // the latency launch: the sequences of the single-CTA kernel's reach in one wide-CTA launch of their longest class
  {
    int wmax = 0;
    for (int k = 0; k < nlen; ++k)
      if (lengths[k] <= lim.wave_limit) wmax = std::max(wmax, lengths[k]);
    if (any_in(-1, lim.wave_limit)) {
      rt.merged = 0;
      p->launches.push_back(Launch{0, std::min(cap, c->num_sms), std::max(wmax, 8), kKindWave, true});
    }
  }
  rt.first_class = (int)p->launches.size();
  // class launches in ascending order of their upper bound (the routing kernel walks them upwards)
  int lo = -1;
  for (int cls = 0; cls < kNumWaveClasses; ++cls) {
    if (kWaveClasses[cls] > lim.wave_limit) break;
    if (any_in(lo, kWaveClasses[cls]) && (int)p->launches.size() < kMaxDynLaunches) {
      p->launches.push_back(Launch{0, cap, top_in(lo, kWaveClasses[cls]), kKindWave, false});
      rt.upper[p->launches.size() - 1] = kWaveClasses[cls];
      if (rt.wfirst < 0) rt.wfirst = (int)p->launches.size() - 1;
      rt.wlast = (int)p->launches.size() - 1;
    }
    lo = kWaveClasses[cls];
  }
  {
    // one cluster launch, of the class of the longest cluster-length target (see run_batch)
    int cmaxlen = 0;
    for (int k = 0; k < nlen; ++k)
      if (lengths[k] > lim.wave_limit && lengths[k] <= lim.cluster_limit) cmaxlen = std::max(cmaxlen, lengths[k]);
    if (cmaxlen > 0 && (int)p->launches.size() < kMaxDynLaunches) {
      int q0 = 0;
      while (q0 + 1 < kNumClusterClasses && kClusterClasses[q0] < cmaxlen) ++q0;
      const size_t per_seq = 2 * cluster_table_stride(kClusterClasses[q0]) * sizeof(int16_t);
      if (per_seq * (size_t)ccap > kClusterWorkspaceCap) return fail(RNAFOLD_EINVAL, "device-routed batch too large for the cluster workspace");
      p->launches.push_back(Launch{0, ccap, kClusterClasses[q0], kKindCluster, false});
      rt.ccap = ccap;
      rt.upper[p->launches.size() - 1] = lim.cluster_limit;
      rt.cfirst = rt.clast = (int)p->launches.size() - 1;
      rt.cmax = cached_occ(c, 1, kClusterClasses[q0], 2, 0);
    }
  }
  if (lmax > lim.cluster_limit && (int)p->launches.size() < kMaxDynLaunches) {
    p->launches.push_back(Launch{0, cap, lmax, lmax <= kLegacySmemLimit ? kKindLegacySmem : kKindLegacyGlobal, false});
    rt.upper[p->launches.size() - 1] = RNAFOLD_MAX_LEN;
  }
  rt.nl = (int)p->launches.size();
  if (rt.nl == 0 || rt.nl > kMaxDynLaunches) return fail(RNAFOLD_EINVAL, "bad length classes for a device-routed batch");
  p->epoch = c->params_epoch;
  return RNAFOLD_OK;
}

}  // namespace

namespace rnafold {
// env.cu: fold the compacted terminal batch (device buffers, host offsets) WITHOUT the trailing stream
// synchronisation -- the environment step enqueues its scatter / reset / observation kernels behind the fold and
// synchronises once at its end, before anything can reuse the staging buffers.
int fold_hamming_batch_enqueue(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target, uint8_t *db_out,
                               int32_t *mfe_out, int32_t *hd_out, double *reward_out, double reward_exp, void *stream) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  return run_batch(g_ctx, seq, off, B, target, db_out, mfe_out, hd_out, reward_out, nullptr, reward_exp, stream, true);
}

// env.cu: the same for a batch whose size lives on the device (no host round trip at all).  `lengths` lists the
// sequence lengths that can occur, `cap` bounds the count; off[0..count-1] are filled by the producer and
// off[count] = *total_ptr is written here.  *plan caches the launch list between calls.  cluster_cap (<= cap, 0 = cap) bounds
// the sequences of cluster length among them (their global tables are what the workspace is sized by); a batch that breaks
// the promise is not folded at all and sets status code 101.
int fold_hamming_dyn_enqueue(void **plan, const int32_t *lengths, int nlen, int cap, int cluster_cap, const uint8_t *seq, int32_t *off,
                             const int32_t *count_ptr, const int32_t *total_ptr, const uint8_t *target, uint8_t *db_out,
                             int32_t *mfe_out, int32_t *hd_out, double *reward_out, double reward_exp, void *stream) {
  std::lock_guard<std::mutex> lk(g_mu);
  Context *c = g_ctx;
  if (!c) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (!plan || !lengths || nlen <= 0 || cap <= 0 || !seq || !off || !count_ptr) return fail(RNAFOLD_EINVAL, "bad arguments");
  DynPlanImpl *p = (DynPlanImpl *)*plan;
  if (!p) *plan = p = new DynPlanImpl();
  const int ccap = (cluster_cap > 0 && cluster_cap < cap) ? cluster_cap : cap;
  if (p->epoch != c->params_epoch || p->cap != cap || p->ccap != ccap) {
    int rc = build_dyn_plan(c, p, lengths, nlen, cap, ccap);
    if (rc) return rc;
  }
  cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
  c->last_launches = 0;
  c->last_routed[0] = c->last_routed[1] = c->last_routed[2] = 0;
  int rc = enter_call(c, st);
  if (rc) return rc;
  const int nl = p->rt.nl;
  CUDA_TRY(c->d_order.reserve(sizeof(int32_t) * (2 * (size_t)cap + 2)));
  CUDA_TRY(c->d_dyn.reserve(sizeof(int32_t) * 2 * (kMaxDynLaunches + 1)));
  CUDA_TRY(c->d_ctl.reserve(sizeof(int32_t) * (size_t)(2 * nl + 2)));
  CUDA_TRY(c->d_redo.reserve(sizeof(int32_t) * (size_t)cap));
  route_kernel<<<1, 1024, 0, st>>>(count_ptr, total_ptr, off, cap, p->rt, (int32_t *)c->d_order.p, (int32_t *)c->d_dyn.p,
                                   (int32_t *)c->d_ctl.p, 2 * nl + 2, (int32_t *)c->d_redo.p, c->d_status);
  CUDA_TRY(cudaGetLastError());
  c->last_launches++;
  // both kernel families present: the wavefront launches run on stream2 beside the cluster / global-table launches
  const bool two = p->rt.wfirst >= 0 && (p->rt.cfirst >= 0 || p->launches.back().kind == kKindLegacySmem ||
                                         p->launches.back().kind == kKindLegacyGlobal);
  if (two) {
    CUDA_TRY(cudaEventRecord(c->ev_fork, st));
    CUDA_TRY(cudaStreamWaitEvent(c->stream2, c->ev_fork, 0));
  }
  bool joined = !two;
  for (int li = 0; li < nl; ++li) {
    const Launch &L = p->launches[li];
    BatchArgs a;
    a.seq = seq;
    a.off = off;
    a.order = (const int32_t *)c->d_order.p;
    a.B = 0;
    a.target = target;
    a.db = db_out;
    a.mfe = mfe_out;
    a.hd = target ? hd_out : nullptr;
    a.reward = target ? reward_out : nullptr;
    a.mismatch = nullptr;
    a.reward_exp = reward_exp;
    a.status = c->d_status;
    a.redo = (int32_t *)c->d_redo.p;
    a.redo_filter = 0;
    a.dyn = (const int32_t *)c->d_dyn.p + 2 * li;
    rc = issue_launch(c, p->lim, L, L.end - L.begin, a, (int32_t *)c->d_ctl.p + 2 * li,
                      (two && L.kind == kKindWave) ? c->stream2 : st, false);
    if (rc) return rc;
    if (li + 1 == nl && !joined) {
      CUDA_TRY(cudaEventRecord(c->ev_join, c->stream2));
      CUDA_TRY(cudaStreamWaitEvent(st, c->ev_join, 0));
      joined = true;
    }
    if (li + 1 == nl && p->lim.wave_esz == 2) {
      // ONE int32 re-run for the sequences any class flagged (values beyond the int16 tables' range): it walks the
      // whole batch in routed order with the longest possible length's layout
      a.redo_filter = 1;
      a.dyn = (const int32_t *)c->d_dyn.p + 2 * nl;
      const bool smem_tables = p->lmax <= kLegacySmemLimit;
      const int occ = std::max(1, cached_occ(c, 2, std::max(p->lmax, 8), smem_tables ? 1 : 0, kCtaThreads));
      const int grid = std::min(cap, c->num_sms * occ);
      int *wsp = nullptr;
      long long stride = 0;
      if (!smem_tables) {
        // the cluster launches' tables are dead by now (stream order); the re-run's own tables go behind them
        stride = 2LL * table_cells(p->lmax);
        CUDA_TRY(c->d_ws2.reserve(sizeof(int32_t) * (size_t)stride * (size_t)grid));
        wsp = (int *)c->d_ws2.p;
      }
      CUDA_TRY(launch_fold_cta(c->d_params, a, std::max(p->lmax, 8), smem_tables, grid, wsp, stride, (int32_t *)c->d_ctl.p + 2 * nl, st));
      c->last_launches++;
    }
  }
  return leave_call(c, st);
}
void fold_dyn_plan_free(void *plan) { delete (DynPlanImpl *)plan; }

// device address of the sticky status word (env.cu reads it with its end-of-step copies)
const int32_t *status_word() { return g_ctx ? g_ctx->d_status : nullptr; }

int ctx_view(CtxView *out) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return -1;
  out->device = g_ctx->device;
  out->num_sms = g_ctx->num_sms;
  out->stream = g_ctx->stream;
  out->hp = &g_ctx->hp;
  out->d_params = g_ctx->d_params;
  out->params_epoch = g_ctx->params_epoch;
  return 0;
}

// error channel shared with env.cu
int set_error(int code, const char *msg) {
  g_err = msg ? msg : "";
  return code;
}
}  // namespace rnafold

extern "C" {

const char *rnafold_last_error(void) { return g_err.c_str(); }
int rnafold_current_device(void) { return g_ctx ? g_ctx->device : -1; }
const char *rnafold_version(void) { return "rnafold_b200 0.1 (sm_100a)"; }

int rnafold_init(int device) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_ctx && g_ctx->device == device) return RNAFOLD_OK;
  if (g_ctx) return fail(RNAFOLD_EINVAL, "context already bound to device %d", g_ctx->device);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(RNAFOLD_ECUDA, "no CUDA device available (%s); this library has no CPU path",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(RNAFOLD_EINVAL, "device %d out of range (0..%d)", device, ndev - 1);
  CUDA_TRY(cudaSetDevice(device));
  Context *c = new Context();
  c->device = device;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    delete c;
    return fail(RNAFOLD_ECUDA, "device %d is sm_%d%d; this build targets sm_100a (B200)", device, prop.major, prop.minor);
  }
  c->num_sms = prop.multiProcessorCount;
  c->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
  CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreate(&c->ev0));
  CUDA_TRY(cudaEventCreate(&c->ev1));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_done, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_staged, cudaEventDisableTiming));
  CUDA_TRY(cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking));
  CUDA_TRY(cudaStreamCreateWithFlags(&c->stream3, cudaStreamNonBlocking));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_join3, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  CUDA_TRY(cudaMalloc((void **)&c->d_status, sizeof(int32_t)));
  CUDA_TRY(cudaMemset(c->d_status, 0, sizeof(int32_t)));
  CUDA_TRY(cudaMalloc((void **)&c->d_params, sizeof(FoldParams)));
  init_host_params(&c->hp);
  char err[256] = "";
  if (parse_parameter_text(rnafold_builtin_turner2004, (size_t)rnafold_builtin_turner2004_len, &c->hp, err, sizeof(err))) {
    delete c;
    return fail(RNAFOLD_EIO, "built-in Turner 2004 parameters: %s", err);
  }
  int rc = upload_params(c);
  if (rc) {
    delete c;
    return rc;
  }
  g_ctx = c;
  // set-up of the allocation-free entry point: device address of its static scratch, once per device, with that
  // device current (a failure only disables that entry point)
  g_noalloc_ptr = nullptr;
  g_noalloc_err = cudaGetSymbolAddress((void **)&g_noalloc_ptr, g_noalloc_scratch);
  if (g_noalloc_err != cudaSuccess) {
    g_noalloc_ptr = nullptr;
    cudaGetLastError();
  }
  return RNAFOLD_OK;
}

void rnafold_shutdown(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  Context *c = g_ctx;
  if (!c) return;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  for (DevBuf *b : {&c->d_seq, &c->d_off, &c->d_order, &c->d_target, &c->d_db, &c->d_mfe, &c->d_hd,
                    &c->d_reward, &c->d_mismatch, &c->d_ctl, &c->d_ws, &c->d_ws2, &c->d_redo, &c->d_dyn})
    b->release();
  for (PinBuf *b : {&c->h_in, &c->h_out, &c->h_off}) b->release();
  cudaFree(c->d_params);
  cudaFree(c->d_status);
  cudaEventDestroy(c->ev0);
  cudaEventDestroy(c->ev1);
  cudaEventDestroy(c->ev_done);
  cudaEventDestroy(c->ev_staged);
  cudaEventDestroy(c->ev_fork);
  cudaEventDestroy(c->ev_join);
  cudaStreamDestroy(c->stream2);
  cudaStreamDestroy(c->stream3);
  cudaEventDestroy(c->ev_join3);
  cudaStreamDestroy(c->stream);
  delete c;
  g_ctx = nullptr;
}

int rnafold_read_parameter_file(const char *path) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (!path) return fail(RNAFOLD_EINVAL, "null path");
  HostParams *hp = new HostParams(g_ctx->hp);  // absent sections keep the live values
  char err[256] = "";
  if (load_parameter_file(path, hp, err, sizeof(err))) {
    delete hp;
    return fail(RNAFOLD_EIO, "%s", err);
  }
  g_ctx->hp = *hp;
  delete hp;
  CUDA_TRY(cudaSetDevice(g_ctx->device));
  return upload_params(g_ctx);
}

int rnafold_export_params(int32_t *out, int cap) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  int n = export_params(&g_ctx->hp, out, cap);
  if (n < 0) return fail(RNAFOLD_EINVAL, "export buffer too small");
  return n;
}

int rnafold_parse_params(const char *path, int32_t *out, int cap) {
  // host-only: parse `path` (NULL = the built-in Turner 2004 text) and dump the tables; no device needed
  HostParams *hp = new HostParams();
  init_host_params(hp);
  char err[256] = "";
  int rc = parse_parameter_text(rnafold_builtin_turner2004, (size_t)rnafold_builtin_turner2004_len, hp, err, sizeof(err));
  if (!rc && path) rc = load_parameter_file(path, hp, err, sizeof(err));
  if (rc) {
    delete hp;
    return fail(RNAFOLD_EIO, "%s", err);
  }
  int n = export_params(hp, out, cap);
  delete hp;
  if (n < 0) return fail(RNAFOLD_EINVAL, "export buffer too small");
  return n;
}

static double *md_slot(Context *c, const char *name, double *dflt) {
  struct {
    const char *n;
    double *p;
    double d;
  } tab[] = {{"temperature", &c->md_temperature, 37.0}, {"dangles", &c->md_dangles, 2.0},
             {"noGU", &c->md_noGU, 0.0},                {"no_closingGU", &c->md_no_closingGU, 0.0},
             {"uniq_ML", &c->md_uniq_ML, 0.0}};
  for (auto &t : tab)
    if (!std::strcmp(t.n, name)) {
      *dflt = t.d;
      return t.p;
    }
  return nullptr;
}

int rnafold_set_md(const char *name, double value) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  double dflt = 0;
  double *p = name ? md_slot(g_ctx, name, &dflt) : nullptr;
  if (!p) return fail(RNAFOLD_EINVAL, "unknown model detail '%s'", name ? name : "(null)");
  if (value != dflt)
    return fail(RNAFOLD_ENOTSUP, "model detail %s=%g is not implemented (only the reference default %g)", name, value, dflt);
  *p = value;
  return RNAFOLD_OK;
}

int rnafold_get_md(const char *name, double *value) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  double dflt = 0;
  double *p = name ? md_slot(g_ctx, name, &dflt) : nullptr;
  if (!p || !value) return fail(RNAFOLD_EINVAL, "unknown model detail");
  *value = *p;
  return RNAFOLD_OK;
}

int rnafold_workspace_bytes(int32_t B, int32_t nmax, size_t *bytes) {
  if (!bytes || B < 0 || nmax < 0 || nmax > RNAFOLD_MAX_LEN) return fail(RNAFOLD_EINVAL, "bad arguments");
  // upper bound of the device scratch one call allocates: staged I/O (3 bytes per nucleotide + per-sequence
  // words), plus the global tables of the long-sequence paths
  size_t ws = 0;
  const int sms = g_ctx ? g_ctx->num_sms : 148;
  if (nmax > kWaveClasses[sizeof(kWaveClasses) / sizeof(int) - 1]) {
    if (nmax <= 464) {  // 2-CTA cluster kernel: int16 c and fML triangles per sequence, capped per launch
      const size_t per_seq = 2 * cluster_table_stride(nmax) * sizeof(int16_t);
      ws = std::min(kClusterWorkspaceCap, per_seq * (size_t)B);
    }
    // int32 kernel with tables in global memory (n > 464, and re-runs of range-check flags): 8 CTAs per SM
    ws += sizeof(int32_t) * 2 * (size_t)table_cells(nmax) * (size_t)std::min<long long>(B, 8LL * sms);
  }
  *bytes = ws + (size_t)B * (size_t)nmax * 3 + (size_t)B * 32;
  return RNAFOLD_OK;
}

int rnafold_fold_hamming_batch(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target,
                               uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out, double *reward_out,
                               uint8_t *mismatch_out, double reward_exp, void *stream) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  return run_batch(g_ctx, seq, off, B, target, db_out, mfe_dcal_out, hd_out, reward_out, mismatch_out,
                   reward_exp, stream);
}

int rnafold_fold_batch(const uint8_t *seq, const int32_t *off, int32_t B, uint8_t *db_out,
                       int32_t *mfe_dcal_out, void *stream) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  return run_batch(g_ctx, seq, off, B, nullptr, db_out, mfe_dcal_out, nullptr, nullptr, nullptr, 1.0, stream);
}

// Allocation-free variant of rnafold_fold_hamming_batch (see include/rnafold_b200.h): no cudaMalloc / cudaFree /
// growing buffer anywhere on this path; the kernels' scratch is the static device array above.
int rnafold_fold_hamming_batch_noalloc(const uint8_t *seq, const int32_t *off, int32_t B, int32_t nmax, const uint8_t *target,
                                       uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out, double *reward_out,
                                       double reward_exp, void *stream) {
  std::lock_guard<std::mutex> lk(g_mu);
  Context *c = g_ctx;
  if (!c) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (!g_noalloc_ptr)
    return fail(RNAFOLD_ECUDA, "static scratch unavailable: cudaGetSymbolAddress: %s", cudaGetErrorString(g_noalloc_err));
  if (B < 0 || !off || (B > 0 && !seq)) return fail(RNAFOLD_EINVAL, "null seq/off or negative batch size");
  if (B > RNAFOLD_NOALLOC_MAX_BATCH)
    return fail(RNAFOLD_EINVAL, "noalloc: batch of %d sequences exceeds the static scratch (RNAFOLD_NOALLOC_MAX_BATCH = %d)", B,
                RNAFOLD_NOALLOC_MAX_BATCH);
  if (nmax < 0 || nmax > RNAFOLD_NOALLOC_MAX_LEN)
    return fail(RNAFOLD_EINVAL, "noalloc: nmax %d exceeds RNAFOLD_NOALLOC_MAX_LEN = %d", nmax, RNAFOLD_NOALLOC_MAX_LEN);
  if (B == 0) return RNAFOLD_OK;
  if (!is_device_ptr(seq) || !is_device_ptr(off) || (target && !is_device_ptr(target)) || (db_out && !is_device_ptr(db_out)) ||
      (mfe_dcal_out && !is_device_ptr(mfe_dcal_out)) || (hd_out && !is_device_ptr(hd_out)) ||
      (reward_out && !is_device_ptr(reward_out)))
    return fail(RNAFOLD_EINVAL, "noalloc: every buffer must be a device pointer (host buffers would need staging memory)");
  CUDA_TRY(cudaSetDevice(c->device));
  const Limits lim = make_limits(c);
  const int cls = std::max(nmax, 8);
  if (cls > lim.wave_limit) return fail(RNAFOLD_EINVAL, "noalloc: nmax %d is beyond the shared-memory kernel's reach (%d)", nmax, lim.wave_limit);
  cudaStream_t st = stream ? (cudaStream_t)stream : c->stream;
  int32_t *scr = g_noalloc_ptr;
  // the array keeps the previous call's contents: queue counters, count and this batch's flags start at zero
  CUDA_TRY(cudaMemsetAsync(scr, 0, sizeof(int32_t) * ((size_t)kNoallocHeader + (size_t)B), st));
  noalloc_validate_kernel<<<1, 1024, 0, st>>>(off, B, nmax, scr, c->d_status);
  CUDA_TRY(cudaGetLastError());
  BatchArgs a;
  a.seq = seq;
  a.off = off;
  a.order = nullptr;  // sequences are folded in batch order (no sorted copy: that would need scratch per call)
  a.B = B;
  a.target = target;
  a.db = db_out;
  a.mfe = mfe_dcal_out;
  a.hd = target ? hd_out : nullptr;
  a.reward = target ? reward_out : nullptr;
  a.mismatch = nullptr;
  a.reward_exp = reward_exp;
  a.status = c->d_status;
  a.redo = scr + kNoallocHeader;
  a.redo_filter = 0;
  a.dyn = scr + 2;
  int threads = wave_threads(cls);
  if (B <= c->num_sms) threads = 512;
  const int occ = std::max(1, wave_occupancy(cls, lim.wave_esz, threads));
  CUDA_TRY(launch_fold_wave(c->d_params, a, cls, lim.wave_esz, std::min(B, c->num_sms * occ), threads, scr + 0, st));
  if (lim.wave_esz == 2) {
    // sequences whose values left the int16 range are re-folded with int32 tables in shared memory
    a.redo_filter = 1;
    const int occ2 = std::max(1, cta_kernel_occupancy(cls, true));
    CUDA_TRY(launch_fold_cta(c->d_params, a, cls, true, std::min(B, c->num_sms * occ2), nullptr, 0, scr + 1, st));
  }
  return RNAFOLD_OK;
}

int rnafold_last_status(int32_t *status, int32_t clear) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  Context *c = g_ctx;
  CUDA_TRY(cudaSetDevice(c->device));
  CUDA_TRY(cudaDeviceSynchronize());  // whatever stream the last (asynchronous) call was enqueued on
  int32_t v = 0;
  CUDA_TRY(cudaMemcpy(&v, c->d_status, sizeof(int32_t), cudaMemcpyDeviceToHost));
  if (v && clear) CUDA_TRY(cudaMemset(c->d_status, 0, sizeof(int32_t)));
  if (status) *status = v;
  if (v) return fail(RNAFOLD_ECUDA, "backtrack found no decomposition (code %d)", v);
  return RNAFOLD_OK;
}

int rnafold_last_launch_info(int32_t *launches, int32_t routed[3]) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (launches) *launches = g_ctx->last_launches;
  if (routed) std::memcpy(routed, g_ctx->last_routed, sizeof(g_ctx->last_routed));
  return RNAFOLD_OK;
}

int rnafold_set_timing(int enabled) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  g_ctx->timed = enabled != 0;
  return RNAFOLD_OK;
}

int rnafold_last_kernel_ms(float *ms) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (!ms || !g_ctx->timed) return fail(RNAFOLD_EINVAL, "timing not enabled (rnafold_set_timing)");
  CUDA_TRY(cudaEventSynchronize(g_ctx->ev1));
  CUDA_TRY(cudaEventElapsedTime(ms, g_ctx->ev0, g_ctx->ev1));
  return RNAFOLD_OK;
}

int rnafold_measure_onchip_peaks(double *lds_per_s, double *minplus_per_s) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (!lds_per_s || !minplus_per_s) return fail(RNAFOLD_EINVAL, "null output");
  CUDA_TRY(cudaSetDevice(g_ctx->device));
  CUDA_TRY(measure_onchip_peaks(g_ctx->num_sms, g_ctx->stream, lds_per_s, minplus_per_s));
  return RNAFOLD_OK;
}

int rnafold_wave_smem_bytes(int32_t nmax, int32_t elem_bytes) {
  // host-only planning helper: dynamic shared memory of the wavefront kernel for a length class
  if (nmax < 0 || nmax > RNAFOLD_MAX_LEN || (elem_bytes != 2 && elem_bytes != 4)) return fail(RNAFOLD_EINVAL, "bad arguments");
  return (int)wave_smem_bytes(nmax, elem_bytes);
}

int rnafold_device_info(int32_t *num_sms, int32_t *smem_optin_bytes) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_ctx) return fail(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (num_sms) *num_sms = g_ctx->num_sms;
  if (smem_optin_bytes) *smem_optin_bytes = g_ctx->max_smem_optin;
  return RNAFOLD_OK;
}

}  // extern "C"
