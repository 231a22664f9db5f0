// qec_device.cuh -- shared host/device definitions and the frame-propagation core.
//
// Data layout:
//   * gate table: one packed uint32 per op, padded with NOPs to a multiple of 4 so a warp
//     streams it with one uniform 128-bit load per 4 gates.
//       bits 0..3  internal opcode (QOP_*), bits 4..17 operand a, bits 18..31 operand b
//   * frame: bit-packed, 32 paths per uint32 "lane word".  One thread owns one lane word;
//     its 2*nq frame words live in shared memory at frame[row * T + tid] (row = qubit for X,
//     nq + qubit for Z), so consecutive threads hit consecutive banks.
//   * fault events of one lane word: uint32 = op << 9 | path_bit << 4 | pauli_code, sorted by
//     op (stable); inside one (op, path_bit) group the LAST event wins, which is the
//     reference's `fail[icomb[j]] = j` rule (errorpropagator.cpp:265-267).
//   * packed outcomes: out[c * n_words + w].
#pragma once
#ifdef __CUDACC_RTC__
// NVRTC has no <stdint.h>: this header is also compiled at run time for the
// circuit-specialised kernels (qec_jit.cpp embeds its text)
typedef unsigned char uint8_t;
typedef unsigned short uint16_t;
typedef unsigned int uint32_t;
typedef int int32_t;
typedef unsigned long long uint64_t;
typedef long long int64_t;
#else
#include <stdint.h>
#endif

#define QOP_NOP 0u    // x, y, z, id (b = 0: may host a fault); barrier, padding (b = 1)
#define QOP_H 1u
#define QOP_S 2u
#define QOP_CX 3u
#define QOP_RESET 4u
#define QOP_MEAS 5u   // measure whose clbit is overwritten later (value is dead)
#define QOP_MEASF 6u  // measure that is the final write to its clbit
#define QOP_ZEROC 7u  // pseudo-op appended by the lowering: clbit `a` is never measured

#define QEC_EV_SENTINEL 0xFFFFFFFFu
#define QEC_EV_OP_SHIFT 9

#if !defined(__CUDACC__) && !defined(__host__)
#define __host__
#define __device__
#endif
static inline __host__ __device__ uint32_t qec_pack_op(uint32_t opc, uint32_t a, uint32_t b) {
  return (opc & 15u) | ((a & 0x3FFFu) << 4) | ((b & 0x3FFFu) << 18);
}

struct QecCircuitDev {
  int nq, nc, n_ops, n_ops4;  // n_ops4 = padded op count / 4 (includes ZEROC pseudo-ops)
  const uint32_t *ops;        // device, n_ops4 * 4 words
};

struct QecOutDev {
  uint32_t *bits;              // [nc][n_words] or nullptr
  uint32_t *frame;             // [2*nq][n_words] or nullptr
  unsigned long long *tally;   // QEC_TALLY_LEN(nc) or nullptr (accumulated with atomics)
  long long n_words;
};

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
  // [L][4] (specialised kernels only) bit i of tile[4 l + c] = bit c of the Pauli code of
  // digit (i mod npaul[l]) of location l: the 32 consecutive paths of a lane word read their
  // last-location masks from it with one funnel shift (needs npaul[l] <= 32)
  const unsigned long long *tile;
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
  const uint32_t *slinf;   // [L] op | class << 20 | (op is a cx) << 30 | cls_always << 31
  int n_classes;           //     (specialised kernels; needs op < 2^20, n_classes <= 1024)
  uint32_t lgP;            // classes padded to P = 2^lgP Paulis in the kernels' shared tables
  int n_seg;               // gate-stream segments of the specialised sampler
  const uint32_t *seg_end; // [n_seg] first slot past segment s (slot = sorted location * 32 + shot)
  const uint32_t *slloc;   // [L] original location index of sl (for the fault export)
  const uint16_t *slcls;   // [L] class of sl
  const uint32_t *cls_acc; // [n_classes] accept threshold round(2^32 * p_class / pmax)
  const uint8_t *cls_always;  // [n_classes] 1 if p_class == pmax
  const uint32_t *cls_off;    // [n_classes+1]
  const uint32_t *cdf;        // [cls_off[n]] cumulative Pauli thresholds (digit d: r < cdf[d])
  const uint8_t *codes;       // [cls_off[n]] Pauli codes
  uint32_t k0, k1;            // Philox key = seed
};

// Linear-response sampler (qec_linear.cu).  Every gate of the path (cx, h, s, reset, measure,
// Pauli injection) is GF(2)-linear in the frame, so the outcome of a shot is the XOR of the
// outcome signatures of its faults.  The signature of fault f = (sorted location sl, digit d)
// is the list of clbits it flips, computed once per sampler by the frame-propagation kernel.
//   batch  = 2^lgb consecutive shots; its slot space is (sorted location, shot in batch),
//            location major: slot = sl << lgb | shot
//   stream = one of n_streams equal slot ranges of a batch; a warp walks a stream with 32
//            geometric gaps per step (Philox counter = batch, draw index, stream) and a warp
//            prefix sum, so every lane carries one fault per step
struct QecLinDev {
  int nc, L, lgb, n_streams;
  uint32_t slen;            // slots per stream = ceil(n_slots / n_streams)
  uint32_t n_slots;         // L << lgb  (< 2^30)
  uint32_t B, K1, gshift;   // gap tables as in QecSamplerDev; 2B = 2^(32 - gshift) guide buckets
  uint32_t n_classes, lgP;
  const uint32_t *T1, *T2;
  const uint32_t *info;     // [L] first fault index of sl | class << 21 | cls_always << 31
  const uint32_t *cls_acc, *cls_off, *cdf;
  const uint32_t *slloc;    // [L] original location index (fault export)
  const uint32_t *sigdesc;  // [n_faults] offset into siglist | length << 22
  const uint16_t *siglist;  // clbits flipped by each fault
  const uint4 *sigrec;      // [n_faults][1 << rec2] uint16 x 8: length, first 7 (15) clbits of the fault
  int rec2;
  int rec8, rec_in_smem;    // 8-byte records (length + 7 clbits as bytes), optionally in shared memory
  int uniform_digits;       // every class has equal Pauli weights: digit = floor(r * n / 2^32)
  uint32_t rk[20];          // Philox round keys (k0, k1 of round 0..9)
  // warp-autonomous variant: read-out program, [n_steps][32 lanes]: clbit row | chain start << 30
  // | valid << 31.  A lane XORs the rows of a chain cumulatively (time-difference basis).
  const uint32_t *rowprog;
  int n_steps;
  uint32_t n_faults, n_sig;
  int info_in_smem, sig_in_smem;
  int warp_mode;            // 1: k_sample_linw (lgb = 8, one warp per batch)
  uint32_t k0, k1;
  // draw scheme: 0 = two-level gap tables (64 bits per gap; any p), 1 / 2 = computed gap
  // (qec_gap_len: 32 bits per gap) with 1 / 2 candidate faults per Philox call (2 when no class
  // needs thinning).  Results depend on the scheme, never on the launch geometry.
  int draw;
  float gapC;               // -1 / log2(1 - pmax)
  const uint2 *info2;       // [L] {first fault of sl, number of Paulis of sl} (lean warp kernel)
  int fast;                 // lean warp kernel eligible: draw 2, uniform digits, 8-byte records and info2 in smem
  // lane-stream variant (k_sample_lanes): lgb = 10, lane = word column, no shared-memory atomics;
  // rowprog is then [n_steps passes][4 rows]: row | previous row of its chain << 8 | flags (qec_linear.cu)
  int lane_mode;
  uint32_t flush_every;     // lane-stream kernel: batches between flushes of the register tally counters (at most 448)
  int lanes_np;             // Philox calls (two candidates each) per loop trip of the lane-stream kernel: 1 or 2
  uint32_t *batch_ctr;      // lane-stream kernel: batches handed out so far in this launch (zeroed per launch)
  // CTA variant: read-out program lvl_prog[step * G + group], G = threads / words per batch row,
  // n_lvl steps; entry = row | 1 << 30 valid | 1 << 31 the row continues the chain (time-difference
  // basis) of the group's previous step.  (lvl_off: unused, kept for layout.)
  const uint32_t *lvl_prog, *lvl_off;
  int n_lvl;
};

// Linear-response enumerator (k_enum_lanes in qec_linear.cu): outcome(path) = XOR of the
// signatures of its faults (the locations of a path are distinct, so no op carries two faults).
// Same lane-owned accumulator and read-out as the lane-stream sampler: a warp owns 1024
// consecutive paths, lane l the 32 paths of word l.
struct QecEnumLinDev {
  int nc, n_steps;          // read-out passes (rowprog: [n_steps][4], see QecLinDev)
  const uint32_t *rowprog;
  const uint2 *sigrec;      // [n_faults] 8-byte records: fault = poff[location] + digit
  uint32_t n_faults;
  uint32_t *batch_ctr;      // batches handed out so far in this launch (zeroed per launch)
  int e32;                  // fewer than 2^32 paths: suffix-sum tables as 32-bit words in shared memory
  uint32_t flush_every;     // batches between flushes of the register tally counters (at most 448)
};
static inline __host__ __device__ size_t qec_enum_lanes_smem_words(int nc, int warps, uint32_t n_faults,
                                                                   uint32_t prog_words, uint32_t L, int order,
                                                                   int e32) {
  const size_t ncp = ((size_t)nc + 3u) & ~(size_t)3u;
  return (size_t)warps * (((size_t)nc + 2u) * 32u + ncp) + prog_words + L +
         (e32 ? (size_t)order * ((size_t)L + 2u) : 0u) + 2u + 2u * (size_t)n_faults;
}

// shared-memory words of the lane-stream variant (1024-shot batches, `warps` warps per CTA)
static inline __host__ __device__ size_t qec_lanes_smem_words(int nc, int warps, uint32_t L, uint32_t n_faults,
                                                              uint32_t prog_words) {
  const size_t ncp = ((size_t)nc + 3u) & ~(size_t)3u;
  const size_t per_warp = ((size_t)nc + 2u) * 32u + ncp;
  return (size_t)warps * per_warp + 64u + prog_words + (((size_t)L + 1u) & ~(size_t)1u) + 2u * (size_t)n_faults;
}

// Half-word lane-stream sampler (k_sample_lanes16 in qec_lanes16.cu): the lane-stream idea with
// 16 shots per lane, so that the accumulator of a warp is half the size and twice the warps fit
// an SM.  A warp owns a batch of 512 shots, lane l the shots 16 l .. 16 l + 15.  The accumulator
// is acc[word-row][lane] (uint32): a word-row packs TWO clbit rows, row A in the low and row B in
// the high half-word; the host pairs whole chains of the time-difference basis (A_k, B_k = k-th
// rows of two chains), so that "previous row of the chain" is the previous word-row for both
// halves.  Fault records are deduplicated 16-byte entries of seven 16-bit byte offsets
// (word-row * 128 + 2 * half; absent = the dummy word-row) that the lanes apply with 16-bit
// shared-memory loads and stores at their own column (bank = lane: no conflicts, no atomics).
struct QecLanes16Dev {
  int nc, L;
  uint32_t n_faults, n_rec, n_wrows;  // word-rows in use (the dummy and the zero word-row follow)
  int n_steps;                         // read-out passes of four word-rows (multiple of 4)
  // tables, each padded to a multiple of 16 bytes (bulk-copied into shared memory)
  const uint32_t *info;     // [L] first fault of sorted location | number of Paulis << 21
  const uint16_t *recidx;   // [n_faults] record of fault (sorted location, digit)
  const uint4 *rec;         // [n_rec] eight uint16: 7 offsets, spare
  const uint32_t *prog;     // [4 n_steps][2] {word-row | previous << 8 | row A has next << 17 | valid << 18 | row B has next << 19,
                            //                 clbit of row A | clbit of row B << 16 (0xFFFF: none)}
  const uint32_t *slloc;    // [L] original location index (fault export)
  uint32_t info_bytes, recidx_bytes, rec_bytes, prog_bytes;
  uint32_t rk[20];          // Philox round keys
  float gapC;
  int noclamp;              // no 32-bit draw reaches QEC_GAP_CAP slots: skip the clamp test
  uint32_t *batch_ctr;      // batches handed out so far in this launch (zeroed per launch)
};
static inline __host__ __device__ size_t qec_lanes16_smem_bytes(int nc, int warps, uint32_t n_wrows,
                                                                uint32_t info_bytes, uint32_t recidx_bytes,
                                                                uint32_t rec_bytes, uint32_t prog_bytes) {
  const size_t ncp = ((size_t)nc + 3u) & ~(size_t)3u;
  return (size_t)warps * ((size_t)n_wrows + 2u) * 128u + (ncp + 64u) * 4u + info_bytes + recidx_bytes +
         rec_bytes + prog_bytes;
}

// Computed geometric gap (draw schemes 1 and 2).  u = (r | 1) / 2^32 is uniform on (0, 1);
// the gap is floor(C * -log2(u)), C = -1 / log2(1 - pmax), so that P(gap >= k) = (1-pmax)^k.
// -log2(u) = (clz + 1) - log2(f), f in [1, 2) the top 24 bits of the normalised word; log2(f)
// by a degree-7 minimax polynomial in f - 1.5 (|error| < 2.8e-7) evaluated with IEEE fused
// multiply-adds only, so that the CPU restatement (oracle/sampler_port.c: fmaf) reproduces every
// gap bit for bit.  Gaps are clamped to QEC_GAP_CAP: a clamped draw advances the stream by
// QEC_GAP_CAP slots without a candidate (exact by memorylessness).
#define QEC_GAP_CAP 16777216u
#define QEC_GAP_COEFFS {0.584962785243988f, 0.9617953896522522f, -0.32063233852386475f, 0.14254532754421234f, \
                        -0.07058960944414139f, 0.037311289459466934f, -0.025123370811343193f, 0.015125359408557415f}

// shared-memory words of the warp-autonomous variant (256-shot batches, `warps` warps per CTA)
static inline __host__ __device__ size_t qec_linw_smem_words(int nc, int warps, uint32_t K1, uint32_t B,
                                                             uint32_t L, uint32_t ncls, uint32_t lgP,
                                                             int info_in_smem, uint32_t rec_words) {
  const size_t per_warp = (size_t)nc * 8u + 128u + (((size_t)nc + 3u) & ~(size_t)3u);
  size_t w = (size_t)warps * per_warp + 64u + K1 + 2u * B + 1u + 2u * ncls + ((size_t)ncls << lgP);
  if (info_in_smem) w += L;
  return w + 2u + rec_words;  // (records of the faults, when they sit in shared memory)
}

// shared-memory words of the linear sampler's CTA (T threads)
static inline __host__ __device__ size_t qec_lin_smem_words(int nc, int lgb, int T, uint32_t K1,
                                                            uint32_t B, uint32_t L, uint32_t ncls,
                                                            uint32_t lgP, uint32_t n_faults,
                                                            uint32_t n_sig, int info_in_smem,
                                                            int sig_in_smem) {
  const size_t BW = (size_t)1 << (lgb - 5);
  size_t w = ((size_t)nc + 1u) * BW + ((size_t)1 << (lgb - 1)) + BW;  // acc (+ a zero row), per-shot fault counts, any
  w += (size_t)nc + 64;                                         // the CTA's tally row, histogram
  (void)T;
  w += (size_t)K1 + 2u * B + 1u + 2u * ncls + ((size_t)ncls << lgP); // T1 T2 guide cacc cnp1 cdf
  if (info_in_smem) w += L;
  if (sig_in_smem) w += (size_t)n_faults + (n_sig + 1u) / 2u;
  return w;
}

struct QecSampleExport {
  uint32_t *cnt;  // [n_words] or nullptr
  uint32_t *ev;   // [n_words][cap]
  int cap;
  int fmt16;      // lane-stream kernel only: rows hold 2 * cap delta-coded 16-bit entries (qec_kernels.cu)
};

// words of shared memory the specialised sampler keeps its small tables in:
// T1[K1] T2[B] guide[2B x u16] info[L] cacc[n] cnp1[n] cdf[n][P] code[n][P x u8]
static inline __host__ __device__ uint32_t qec_samp_tab_words(uint32_t K1, uint32_t B, uint32_t L,
                                                               uint32_t ncls, uint32_t lgP) {
  const uint32_t P = 1u << lgP;
  return K1 + B + B + L + 2u * ncls + ncls * P + (ncls * P + 3u) / 4u;
}

#ifdef __CUDACC__

// One lane word through the whole gate table.  `fr` points at this thread's column of the
// shared-memory frame (stride T).  Events come from ev[k * ev_stride], k < n_ev.
// Gate semantics: reference errorpropagator.cpp:211-254 (cx/h/s/reset/measure) and
// :272-310 (fault after the gate, except measure where it comes first).
template <bool TALLY>
__device__ __forceinline__ void qec_propagate_word(const QecCircuitDev &C, uint32_t *fr,
                                                   const int T, const uint32_t *ev,
                                                   const long long ev_stride, const int n_ev,
                                                   const bool valid, const long long w,
                                                   const QecOutDev &O, uint32_t *s_tally,
                                                   uint32_t &any_out) {
  const int nq = C.nq;
  for (int r = 0; r < 2 * nq; ++r) fr[r * T] = 0u;
  int ei = 0;
  uint32_t cur = (n_ev > 0) ? ev[0] : QEC_EV_SENTINEL;
  uint32_t nxt = (n_ev > 1) ? ev[ev_stride] : QEC_EV_SENTINEL;
  uint32_t cur_op = cur >> QEC_EV_OP_SHIFT;
  const uint4 *ops4 = reinterpret_cast<const uint4 *>(C.ops);
  uint32_t *obits = (valid && O.bits) ? O.bits + w : nullptr;
  const long long ostride = O.n_words;
  const unsigned lane = threadIdx.x & 31u;

  uint4 g4 = __ldg(ops4);
  for (int i4 = 0; i4 < C.n_ops4; ++i4) {
    const uint4 gcur = g4;
    if (i4 + 1 < C.n_ops4) g4 = __ldg(ops4 + i4 + 1);  // prefetch next 4 gates
    const uint32_t gs[4] = {gcur.x, gcur.y, gcur.z, gcur.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const uint32_t g = gs[j];
      const uint32_t opc = g & 15u;
      const uint32_t a = (g >> 4) & 0x3FFFu;
      const uint32_t b = g >> 18;
      const uint32_t idx = (uint32_t)(i4 * 4 + j);
      uint32_t mxa = 0u, mza = 0u, mxb = 0u, mzb = 0u;
      const bool hit = (cur_op == idx);
      if (hit) {
        // gather this op's events into per-bit masks; a later event on the same path bit
        // overwrites an earlier one (last duplicate wins)
        do {
          const uint32_t bm = 1u << ((cur >> 4) & 31u);
          const uint32_t code = cur & 15u;
          mxa = (mxa & ~bm) | ((0u - (code & 1u)) & bm);
          mza = (mza & ~bm) | ((0u - ((code >> 1) & 1u)) & bm);
          mxb = (mxb & ~bm) | ((0u - ((code >> 2) & 1u)) & bm);
          mzb = (mzb & ~bm) | ((0u - ((code >> 3) & 1u)) & bm);
          ++ei;
          cur = nxt;
          nxt = (ei + 1 < n_ev) ? ev[(long long)(ei + 1) * ev_stride] : QEC_EV_SENTINEL;
          cur_op = cur >> QEC_EV_OP_SHIFT;
        } while (cur_op == idx);
      }
      uint32_t *xa = fr + a * T;
      uint32_t *za = fr + (nq + a) * T;
      if (opc == QOP_MEAS || opc == QOP_MEASF) {
        uint32_t v = *xa;
        if (hit) {  // fault BEFORE the measurement
          v ^= mxa;
          *xa = v;
          *za ^= mza;
        }
        if (opc == QOP_MEASF) {
          if (obits) obits[(long long)b * ostride] = v;
          if (TALLY) {
            any_out |= v;
            int cnt = __popc(v);
            cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
            if (lane == 0 && cnt) atomicAdd(&s_tally[b], (uint32_t)cnt);
          }
        }
      } else if (opc == QOP_CX) {
        uint32_t *xb = fr + b * T;
        uint32_t *zb = fr + (nq + b) * T;
        const uint32_t vxa = *xa, vzb = *zb;
        uint32_t vxb = *xb ^ vxa;
        uint32_t vza = *za ^ vzb;
        if (hit) {
          *xa = vxa ^ mxa;
          *zb = vzb ^ mzb;
          vxb ^= mxb;
          vza ^= mza;
        }
        *xb = vxb;
        *za = vza;
      } else if (opc == QOP_H) {
        const uint32_t vx = *xa, vz = *za;
        *xa = vz ^ mxa;
        *za = vx ^ mza;
      } else if (opc == QOP_RESET) {
        *xa = mxa;
        *za = mza;
      } else if (opc == QOP_S) {
        const uint32_t vx = *xa;
        *za = (*za ^ vx) ^ mza;
        if (hit) *xa = vx ^ mxa;
      } else if (opc == QOP_ZEROC) {
        if (obits) obits[(long long)a * ostride] = 0u;
      } else {  // NOP: only a fault, if any (b != 0 marks a barrier: faults ignored)
        if (hit && b == 0u) {
          *xa ^= mxa;
          *za ^= mza;
        }
      }
    }
  }
  if (valid && O.frame) {
    for (int r = 0; r < 2 * nq; ++r) O.frame[(long long)r * O.n_words + w] = fr[r * T];
  }
}

// ---- FaultEnumerator unranking ---------------------------------------------------------
// Rank -> (k-subset, Pauli digits) by closed-form unranking over the suffix tables E[m][l];
// successor steps for the following paths.  Order of paths: reference combinations.cpp:18-32
// (lexicographic subsets) and productiterator.cpp:17-32 (last digit fastest).
template <int K>
struct EnumState {
  uint32_t loc[K], dig[K], np[K], op[K], pb[K];
};

template <int K>
__device__ __forceinline__ void enum_load(const QecEnumDev &E, EnumState<K> &s, int j) {
  const uint32_t l = s.loc[j];
  s.np[j] = E.npaul[l];
  s.op[j] = E.lop[l];
  s.pb[j] = E.poff[l];
}

// (64-bit / 32-bit division costs ~100 instructions; ranks below 2^32 -- every BASELINE
// enumeration except order 3 -- take the 32-bit divide)
__device__ __forceinline__ unsigned long long qec_div_u64_u32(unsigned long long a, uint32_t b) {
  if ((a >> 32) == 0ull) return (unsigned long long)((uint32_t)a / b);
  return a / b;
}

template <int K>
__device__ __forceinline__ void enum_unrank(const QecEnumDev &E, unsigned long long r,
                                            EnumState<K> &s) {
  const int L2 = E.L + 2;
  uint32_t mult = 1;  // product of the Pauli counts of the locations chosen so far (<= 512^(K-1))
  uint32_t start = 0;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    const unsigned long long *Em = E.E + (size_t)(K - j) * L2;
    const unsigned long long u = qec_div_u64_u32(r, mult);
    const unsigned long long thr = Em[start] - u;  // >= 1
    // largest l in [start, L] with Em[l] >= thr (Em non-increasing, Em[L] = 0 < thr)
    uint32_t lo = start, hi = (uint32_t)E.L;
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (Em[mid] >= thr) lo = mid; else hi = mid;
    }
    r -= (unsigned long long)mult * (Em[start] - Em[lo]);
    s.loc[j] = lo;
    enum_load<K>(E, s, j);
    mult *= s.np[j];
    start = lo + 1;
  }
  // r is now the index inside the subset's product: below prod(np) < 2^32
  uint32_t rr = (uint32_t)r;
#pragma unroll
  for (int j = K - 1; j >= 0; --j) {
    const uint32_t q = rr / s.np[j];
    s.dig[j] = rr - q * s.np[j];
    rr = q;
  }
}

// advance to the next path; returns false past the last one
template <int K>
__device__ __forceinline__ bool enum_next(const QecEnumDev &E, EnumState<K> &s) {
#pragma unroll
  for (int j = K - 1; j >= 0; --j) {
    if (++s.dig[j] < s.np[j]) return true;
    s.dig[j] = 0;
  }
  int i = K - 1;
  while (i >= 0 && s.loc[i] >= (uint32_t)(E.L - K + i)) --i;
  if (i < 0) return false;
  const uint32_t base = s.loc[i] + 1;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    if (j >= i) {
      s.loc[j] = base + (uint32_t)(j - i);
      enum_load<K>(E, s, j);
    }
  }
  return true;
}

// ---- Philox4x32-10 (Salmon et al. 2011), counter-based ------------------------------
struct QecPhilox {
  uint32_t c[4];
};
__device__ __forceinline__ QecPhilox qec_philox(uint32_t c0, uint32_t c1, uint32_t c2,
                                                uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  QecPhilox o;
  o.c[0] = c0; o.c[1] = c1; o.c[2] = c2; o.c[3] = c3;
  return o;
}

// the same with the ten round keys precomputed (kernel parameters: constant-bank operands)
__device__ __forceinline__ QecPhilox qec_philox_rk(uint32_t c0, uint32_t c1, uint32_t c2,
                                                   uint32_t c3, const uint32_t (&rk)[20]) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    const uint32_t n0 = hi1 ^ c1 ^ rk[2 * r], n2 = hi0 ^ c3 ^ rk[2 * r + 1];
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
  }
  QecPhilox o;
  o.c[0] = c0; o.c[1] = c1; o.c[2] = c2; o.c[3] = c3;
  return o;
}

// gap of one 32-bit draw, clamped to QEC_GAP_CAP (see the comment at QecLinDev)
__device__ __forceinline__ uint32_t qec_gap_len(uint32_t r, float C) {
  const float k[8] = QEC_GAP_COEFFS;
  const uint32_t x = r | 1u;
  const int n = __clz((int)x);
  const float f = __uint_as_float(((x << n) >> 8) + 0x3F000000u);
  const float z = __fadd_rn(f, -1.5f);
  float p = k[7];
#pragma unroll
  for (int j = 6; j >= 0; --j) p = __fmaf_rn(p, z, k[j]);
  const float l = __fadd_rn((float)(n + 1), -p);
  const float t = fminf(__fmul_rn(C, l), (float)QEC_GAP_CAP);
  return (uint32_t)max(__float2int_rz(t), 0);
}

// number of entries of the non-increasing table t[0..n) that are > r
__device__ __forceinline__ uint32_t qec_count_greater(const uint32_t *t, uint32_t n, uint32_t r) {
  uint32_t lo = 0, hi = n;  // invariant: t[<lo] > r, t[>=hi] <= r
  while (lo < hi) {
    const uint32_t mid = (lo + hi) >> 1;
    if (t[mid] > r) lo = mid + 1; else hi = mid;
  }
  return lo;
}

#endif  // __CUDACC__
