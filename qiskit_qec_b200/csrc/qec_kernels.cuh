// qec_kernels.cuh -- device-side parameter blocks shared by qec_kernels.cu and qec_capi.cu
#pragma once
#include <stdint.h>
#include "qec_device.cuh"

#define QEC_KERNEL_MAX_ORDER 4

// FaultEnumerator tables (all device pointers).
//   E[m * (L + 2) + l], m = 0..order:  E[0][l] = 1;  E[m][l] = sum_{j >= l} npaul[j] * E[m-1][j+1]
//   = number of order-m paths whose first location is >= l.  Total = E[order][0].
struct QecEnumDev {
  int order, L;
  const unsigned long long *E;
  const uint32_t *npaul;  // [L]   Paulis offered by location l
  const uint32_t *poff;   // [L+1] prefix of npaul
  const uint8_t *pcode;   // [poff[L]] 4-bit Pauli codes in digit order
  const uint32_t *lop;    // [L]   op index of location l
};

// FaultSampler tables (all device pointers).  "sl" = location index after a stable sort of
// the locations by op index; a 32-shot word owns the slot space slot = sl * 32 + shot_bit.
struct QecSamplerDev {
  int L;
  long long n_slots;       // L * 32
  uint32_t B, K1;          // two-level geometric gap tables: gap = a * B + b
  const uint32_t *T1;      // [K1]  T1[i] = round(2^32 * Q^(i+1)), Q = (1-pmax)^B
  const uint32_t *T2;      // [B]   T2[j] = round(2^32 * P(b >= j | b < B)), T2[0] unused
  const uint32_t *slop;    // [L] op index of sorted location sl
  const uint32_t *slloc;   // [L] original location index of sl (for the fault export)
  const uint16_t *slcls;   // [L] class of sl
  const uint32_t *cls_acc; // [n_classes] accept threshold round(2^32 * p_class / pmax)
  const uint8_t *cls_always;  // [n_classes] 1 if p_class == pmax
  const uint32_t *cls_off;    // [n_classes+1]
  const uint32_t *cdf;        // [cls_off[n]] cumulative Pauli thresholds (digit d: r < cdf[d])
  const uint8_t *codes;       // [cls_off[n]] Pauli codes
  uint32_t k0, k1;            // Philox key = seed
};

struct QecSampleExport {
  uint32_t *cnt;  // [n_words] or nullptr
  uint32_t *ev;   // [n_words][cap]
  int cap;
};

struct QecLaunchCfg {
  int grid, block;
  size_t smem;
};

// launchers (defined in qec_kernels.cu); all asynchronous on `stream`
void qec_launch_batch(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                      const QecOutDev &O, const long long *ev_off, const uint32_t *ev);
void qec_launch_enumerate(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                          const QecOutDev &O, const QecEnumDev &E, long long lo, long long hi,
                          uint32_t *scratch);
void qec_launch_sample(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                       const QecOutDev &O, const QecSamplerDev &S, long long word_lo,
                       long long n_shots, uint32_t *scratch, int cap, const QecSampleExport &X,
                       uint32_t *overflow_flag);
void qec_launch_outcome_hist(cudaStream_t stream, const uint32_t *bits, int nc, long long n_paths,
                             unsigned long long *keys, unsigned long long *counts,
                             long long table_cap);
void qec_launch_int_peak(cudaStream_t stream, int grid, int block, int iters, uint32_t *sink);
size_t qec_kernel_smem_bytes(int kind, int nq, int nc, int T, uint32_t B, uint32_t K1);
int qec_kernel_set_smem_attr(int kind, int order, size_t smem);
