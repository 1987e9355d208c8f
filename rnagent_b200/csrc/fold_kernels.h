// fold_kernels.h -- launch interface between the C ABI (capi.cu) and the CUDA kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "fold_params.h"

namespace rnafold {

constexpr int kCtaThreads = 128;

// Cells of the diagonal-major triangle (diagonals 4..n-1) of one DP table.
__host__ __device__ inline int table_cells(int n) { return n > 4 ? ((n - 4) * (n - 3)) / 2 : 0; }

// Device pointers of one batch (CSR layout, see include/rnafold_b200.h).
struct BatchArgs {
  const uint8_t *seq;
  const int32_t *off;
  const int32_t *order;  // processing order (longest first) or nullptr
  int32_t B;
  const uint8_t *target;  // nullptr: fold only
  uint8_t *db;
  int32_t *mfe;
  int32_t *hd;
  double *reward;
  uint8_t *mismatch;
  double reward_exp;
  int32_t *status;  // set non-zero by a CTA whose backtrack found no decomposition
  int32_t *redo;    // per sequence: set by the int16 kernel when a value left its checked range
  int32_t redo_filter;  // non-zero: fold only sequences whose redo flag is set (int32 re-run)
  // Batch size known only on the device (environment step: the terminal batch is compacted by a kernel and never
  // crosses to the host): dyn[0] = number of sequences of this launch, dyn[1] = its first entry in `order`.
  // nullptr: B / order as given.
  const int32_t *dyn;
};
__device__ __forceinline__ int batch_count(const BatchArgs &a) { return a.dyn ? a.dyn[0] : a.B; }
__device__ __forceinline__ int batch_entry(const BatchArgs &a, int slot) {
  return a.order ? a.order[slot + (a.dyn ? a.dyn[1] : 0)] : slot;
}

size_t cta_kernel_smem_bytes(int nmax, bool tables_in_smem);
int cta_kernel_occupancy(int nmax, bool tables_in_smem);
cudaError_t launch_fold_cta(const FoldParams *gp, const BatchArgs &a, int nmax, bool tables_in_smem,
                            int grid, int *gws, long long ws_stride, int *queue, cudaStream_t st);

// fold_wave.cu: the fast wavefront kernel (tables in shared memory, int16 or int32 storage)
size_t wave_smem_bytes(int nmax, int esz);
int wave_threads(int nmax);
bool wave_int16_ok(const HostParams &hp);
int wave_occupancy(int nmax, int esz, int threads);
cudaError_t wave_upload_constants(const HostParams &hp, cudaStream_t st);
cudaError_t launch_fold_wave(const FoldParams *gp, const BatchArgs &a, int nmax, int esz, int grid, int threads,
                             int *queue, cudaStream_t st);

// cluster path for long sequences: 2-CTA cluster per sequence (DSMEM row exchange) + warp-per-sequence
// f5/backtrack kernel over the global tables (int16 tables; cluster_table_stride elements per table)
size_t cluster_smem_bytes(int nmax, int esz);
size_t cluster_table_stride(int nmax);
int cluster_max_active(int nmax);
cudaError_t launch_fold_cluster(const FoldParams *gp, const BatchArgs &a, int nmax, int num_clusters, int *queue,
                                void *cbg, void *fg, cudaStream_t st);
cudaError_t launch_f5_backtrack(const FoldParams *gp, const BatchArgs &a, int nmax, int count, const void *cbg,
                                const void *fg, cudaStream_t st);

// capi.cu: fold + Hamming of a device-resident batch with host offsets, enqueued on `stream` without a trailing
// synchronisation (the caller synchronises before the next fold call)
int fold_hamming_batch_enqueue(const uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target, uint8_t *db_out,
                               int32_t *mfe_out, int32_t *hd_out, double *reward_out, double reward_exp, void *stream);

// capi.cu: the same for a batch whose SIZE lives on the device (count_ptr; cap bounds it; `lengths` = the sequence
// lengths that can occur): a routing kernel bins the sequences by length class, every fold kernel reads its own
// count from device memory -- nothing crosses to the host.  *plan caches the launch list (fold_dyn_plan_free).
int fold_hamming_dyn_enqueue(void **plan, const int32_t *lengths, int nlen, int cap, int cluster_cap, const uint8_t *seq, int32_t *off,
                             const int32_t *count_ptr, const int32_t *total_ptr, const uint8_t *target, uint8_t *db_out,
                             int32_t *mfe_out, int32_t *hd_out, double *reward_out, double reward_exp, void *stream);
void fold_dyn_plan_free(void *plan);
const int32_t *status_word();

// permute.cu: rnafold_permute_batch with the draws of sequence b keyed on d_keys[b] (device array; nullptr: on b)
int permute_batch_keyed(uint8_t *seq, const int32_t *off, int32_t B, const uint8_t *target, int32_t budget, int32_t radius,
                        int32_t threshold, double mutation_probability, int32_t allow_gu, uint64_t seed, double reward_exp,
                        uint8_t *db_out, int32_t *mfe_dcal_out, int32_t *hd_out, double *reward_out, int32_t *rounds_out,
                        int32_t *folds_out, const unsigned long long *d_keys);

// capi.cu: what other translation units need from the per-process context (ensemble.cu)
struct CtxView {
  int device, num_sms;
  cudaStream_t stream;
  const HostParams *hp;        // live parameter set (host copy)
  const FoldParams *d_params;  // ... and its device copy
  uint64_t params_epoch;       // changes with every parameter upload
};
int ctx_view(CtxView *out);    // 0, or non-zero before rnafold_init

// capi.cu: records the message returned by rnafold_last_error() and passes `code` through
int set_error(int code, const char *msg);

// peaks.cu: on-box shared-memory-load and min-plus (ALU) peak rates, whole device, per second
cudaError_t measure_onchip_peaks(int num_sms, cudaStream_t st, double *lds_per_s, double *minplus_per_s);

}  // namespace rnafold
