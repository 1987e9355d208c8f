// evaluate.cu -- rnafold_eval_structure_batch: fc.eval_structure for a batch of (sequence, structure) pairs
// (reference: rlif/rna/solution.py:232,234 evaluate the centroid and the MEA structure of a design).
// One CTA per pair: brackets -> pair table (one thread, the table doubles as the bracket stack), then the loop energies
// of evaluate.cuh spread over the threads.
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <vector>

#include "../../include/rnafold_b200.h"
#include "evaluate.cuh"
#include "fold_kernels.h"

namespace rnafold {
namespace {

constexpr int kEvalThreads = 128;

__global__ void __launch_bounds__(kEvalThreads)
eval_structure_kernel(const FoldParams *__restrict__ gp, const uint8_t *__restrict__ seq, const int32_t *__restrict__ off,
                      int B, const uint8_t *__restrict__ db, int32_t *e_out, int32_t *err_out) {
  extern __shared__ __align__(16) unsigned char smem[];
  __shared__ SmallParams sp;
  __shared__ int acc[4];
  const int tid = threadIdx.x;
  for (int k = tid; k < (int)(sizeof(SmallParams) / 4); k += kEvalThreads)
    reinterpret_cast<int32_t *>(&sp)[k] = reinterpret_cast<const int32_t *>(&gp->s)[k];
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const int beg = off[b], n = off[b + 1] - beg;
    int *pt = reinterpret_cast<int *>(smem);
    unsigned char *S = reinterpret_cast<unsigned char *>(pt + n + 2);
    __syncthreads();
    for (int k = tid; k < n; k += kEvalThreads) {
      int c = 0;
      switch (seq[beg + k]) {
        case 'A': case 'a': c = 1; break;
        case 'C': case 'c': c = 2; break;
        case 'G': case 'g': c = 3; break;
        case 'U': case 'u': case 'T': case 't': c = 4; break;
      }
      S[k + 1] = (unsigned char)c;
    }
    if (tid == 0) {
      // open brackets form a linked stack inside the table itself: pt[k] = previous open position until k is closed
      int top = 0, bad = 0;
      pt[0] = pt[n + 1] = 0;
      for (int k = 1; k <= n; ++k) {
        const uint8_t c = db[beg + k - 1];
        pt[k] = 0;
        if (c == '(') {
          pt[k] = top;
          top = k;
        } else if (c == ')') {
          if (!top) {
            bad = 3;
            break;
          }
          const int o = top;
          top = pt[o];
          pt[o] = k;
          pt[k] = o;
        }
      }
      acc[2] = (bad || top) ? 3 : 0;
    }
    __syncthreads();
    if (tid == 0 && n > 0) {
      S[0] = S[n];
      S[n + 1] = S[1];
    }
    const int bad = acc[2];
    int err = 0, e = 0;
    if (!bad && n > 0) e = eval_structure_cta(sp, gp, S, pt, n, acc, &err);
    if (tid == 0) {
      const int code = bad ? bad : err;
      e_out[b] = code ? 0 : e;
      if (err_out) err_out[b] = code;
    }
  }
}

}  // namespace
}  // namespace rnafold

using namespace rnafold;

extern "C" int rnafold_eval_structure_batch(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *structure,
                                            int32_t *e_dcal_out, int32_t *err_out) {
  CtxView cv;
  if (ctx_view(&cv)) return set_error(RNAFOLD_ENOINIT, "rnafold_init was not called");
  if (B < 0 || !off || (B > 0 && (!seq || !structure || !e_dcal_out)))
    return set_error(RNAFOLD_EINVAL, "eval_structure: null seq / off / structure / output or negative batch size");
  if (B == 0) return RNAFOLD_OK;
  if (off[0] != 0) return set_error(RNAFOLD_EINVAL, "off[0] must be 0");
  int nmax = 0;
  for (int b = 0; b < B; ++b) {
    const int n = off[b + 1] - off[b];
    if (n < 0 || n > kMaxLen) {
      char buf[128];
      std::snprintf(buf, sizeof(buf), "eval_structure: sequence %d has length %d (limit %d)", b, n, kMaxLen);
      return set_error(RNAFOLD_EINVAL, buf);
    }
    nmax = std::max(nmax, n);
  }
  std::vector<void *> tmp;
  auto fail = [&](const char *what, cudaError_t e) {
    char buf[256];
    std::snprintf(buf, sizeof(buf), "eval_structure: %s: %s", what, cudaGetErrorString(e));
    for (void *p : tmp) cudaFree(p);
    return set_error(RNAFOLD_ECUDA, buf);
  };
  cudaError_t e = cudaSetDevice(cv.device);
  if (e != cudaSuccess) return fail("cudaSetDevice", e);
  cudaStream_t st = cv.stream;
  const size_t total = (size_t)off[B];
  auto stage = [&](const void *src, size_t bytes, void **dev) -> cudaError_t {
    cudaError_t r = cudaMalloc(dev, bytes ? bytes : 1);
    if (r != cudaSuccess) return r;
    tmp.push_back(*dev);
    return src ? cudaMemcpyAsync(*dev, src, bytes, cudaMemcpyDefault, st) : cudaSuccess;
  };
  void *d_seq = nullptr, *d_off = nullptr, *d_db = nullptr, *d_e = nullptr, *d_err = nullptr;
  if ((e = stage(seq, total, &d_seq)) != cudaSuccess) return fail("stage seq", e);
  if ((e = stage(off, sizeof(int32_t) * ((size_t)B + 1), &d_off)) != cudaSuccess) return fail("stage off", e);
  if ((e = stage(structure, total, &d_db)) != cudaSuccess) return fail("stage structure", e);
  if ((e = stage(nullptr, sizeof(int32_t) * (size_t)B, &d_e)) != cudaSuccess) return fail("alloc energies", e);
  if ((e = stage(nullptr, sizeof(int32_t) * (size_t)B, &d_err)) != cudaSuccess) return fail("alloc codes", e);
  const size_t smem = (size_t)(nmax + 2) * sizeof(int) + (size_t)nmax + 8;
  const int grid = std::min(B, cv.num_sms * 8);
  eval_structure_kernel<<<grid, kEvalThreads, smem, st>>>(cv.d_params, (const uint8_t *)d_seq, (const int32_t *)d_off, B,
                                                          (const uint8_t *)d_db, (int32_t *)d_e, (int32_t *)d_err);
  if ((e = cudaGetLastError()) != cudaSuccess) return fail("launch", e);
  if ((e = cudaMemcpyAsync(e_dcal_out, d_e, sizeof(int32_t) * (size_t)B, cudaMemcpyDefault, st)) != cudaSuccess)
    return fail("copy energies", e);
  if (err_out && (e = cudaMemcpyAsync(err_out, d_err, sizeof(int32_t) * (size_t)B, cudaMemcpyDefault, st)) != cudaSuccess)
    return fail("copy codes", e);
  if ((e = cudaStreamSynchronize(st)) != cudaSuccess) return fail("synchronize", e);
  for (void *p : tmp) cudaFree(p);
  return RNAFOLD_OK;
}
