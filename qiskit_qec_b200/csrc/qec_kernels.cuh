// qec_kernels.cuh -- device-side parameter blocks shared by qec_kernels.cu and qec_capi.cu
#pragma once
#include <stdint.h>
#include "qec_device.cuh"

#define QEC_KERNEL_MAX_ORDER 4
// k_int_peak: unrolled chain steps per loop trip x 8 chains = LOP3 executed per trip and thread
#define QEC_INT_PEAK_UNROLL 16
#define QEC_INT_PEAK_OPS_PER_TRIP (QEC_INT_PEAK_UNROLL * 8)

struct QecLaunchCfg {
  int grid, block;
  size_t smem;
};

// launchers (defined in qec_kernels.cu); all asynchronous on `stream`
void qec_launch_batch(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                      const QecOutDev &O, const long long *ev_off, const uint32_t *ev);
void qec_launch_enumerate(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                          const QecOutDev &O, const QecEnumDev &E, long long lo, long long hi,
                          uint32_t *scratch, const long long *word_list = nullptr,
                          const unsigned int *list_count = nullptr);
void qec_launch_sample(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                       const QecOutDev &O, const QecSamplerDev &S, long long word_lo,
                       long long n_shots, uint32_t *scratch, int cap, const QecSampleExport &X,
                       uint32_t *overflow_flag);
// linear-response sampler (qec_linear.cu)
void qec_launch_sample_lin(const QecLaunchCfg &cfg, cudaStream_t stream, const QecOutDev &O,
                           const QecLinDev &S, long long shot_lo, long long n_shots,
                           const QecSampleExport &X);
int qec_lin_set_smem_attr(size_t smem);
// half-word lane-stream sampler (qec_lanes16.cu)
void qec_launch_sample_lanes16(const QecLaunchCfg &cfg, cudaStream_t stream, const QecOutDev &O,
                               const QecLanes16Dev &S, long long shot_lo, long long n_shots,
                               const QecSampleExport &X);
int qec_lanes16_set_smem_attr(size_t smem);
// linear-response enumerator (qec_linear.cu); order 1..QEC_KERNEL_MAX_ORDER
void qec_launch_enum_lanes(const QecLaunchCfg &cfg, cudaStream_t stream, const QecOutDev &O,
                           const QecEnumDev &E, const QecEnumLinDev &EL, long long lo, long long hi);
int qec_enum_lanes_set_smem_attr(size_t smem);
void qec_launch_outcome_hist(cudaStream_t stream, const uint32_t *bits, int nc, long long n_paths,
                             long long n_words, unsigned long long *keys, unsigned long long *counts,
                             uint32_t *rows, long long table_cap, unsigned long long *status);
void qec_launch_hist_compact(cudaStream_t stream, const unsigned long long *keys,
                             const unsigned long long *counts, const uint32_t *rows,
                             long long table_cap, int rw, unsigned long long *out_keys,
                             unsigned long long *out_counts, uint32_t *out_rows,
                             unsigned long long *n_out);
// sampler fault export: fixed-capacity event rows -> dense pool (block_sums: one word per 1024 lane words)
void qec_launch_event_compact(cudaStream_t stream, const uint32_t *cnt, const uint32_t *ev, long long n_words,
                              int two_ended, int cap, uint32_t *block_sums, uint32_t *overflow,
                              unsigned long long *total, uint16_t *out_cnt, uint32_t *out_ev,
                              unsigned long long pool_cap, int ev24 = 0);
// 16-bit delta-coded form of the same (events in slot order per word; qec_kernels.cu)
void qec_launch_event_compact16(cudaStream_t stream, const uint32_t *cnt, const uint16_t *rows16, long long n_words,
                                int cap, uint32_t *block_sums, uint32_t *overflow, unsigned long long *total,
                                uint16_t *out_cnt, uint16_t *out_ev, unsigned long long pool_cap);
// multi-GPU: device 0 sums the peers' tallies through peer pointers
void qec_launch_tally_reduce_peers(cudaStream_t stream, const unsigned long long *const *ptrs, int n_dev,
                                   int len, unsigned long long *out);
void qec_launch_int_peak(cudaStream_t stream, int grid, int block, int iters, uint32_t *sink);
size_t qec_kernel_smem_bytes(int kind, int nq, int nc, int T, uint32_t B, uint32_t K1);
int qec_kernel_set_smem_attr(int kind, int order, size_t smem);
