// qec_jit_kernels.cuh -- kernel bodies of the circuit-specialised path.
//
// The table-driven kernels of qec_kernels.cu keep a lane word's frame in shared memory and
// decode every gate at run time.  For circuits whose frame fits the register file
// (2*nq <= ~200) the library instead emits, per circuit, straight-line CUDA C for the gate
// stream (qec_jit.cpp: every frame word is a named register, H is a renaming, a CX is two
// LOP3s), compiles it for sm_100a with NVRTC and launches it through the driver API.  This
// header is the hand-written part that is compiled together with the generated
// `qec_jit_gates()`: event cursors, fault generation (in-kernel unranking / Philox) and the
// persistent kernel loops.  The text of this file and of qec_device.cuh is embedded in
// libqecb200.so (qec_jit_embed.h, written by the build).
//
// Macros set by the generated prelude:
//   QEC_JIT_T      threads per CTA (lane words per chunk)
//   QEC_JIT_MINB   CTAs per SM the register budget is sized for
//   QEC_JIT_KIND   1 = FaultEnumerator, 2 = FaultSampler
//   QEC_JIT_ORDER  enumerator order K
//   QEC_JIT_TALLY / QEC_JIT_BITS   tallies wanted / packed outcomes wanted
//   QEC_JIT_NC     number of clbits
#pragma once

#define QEC_OP_END 0x7FFFFFFFu
#define QEC_EVW 5  // words per mask event: op, X/Z masks of operand a, X/Z masks of operand b

// ---- mask events (enumerator): one event per distinct op, 32 path bits per mask ----------
// word k of event e lives at p[(5 e + k) * es] (es = threads in the grid: coalesced).
// The list is sorted by op, ops are unique, and two END headers follow the last event.
struct QecEvList {
  uint32_t *p;
  long long es;
  int n;
  __device__ __forceinline__ uint32_t &at(int e, int k) { return p[(long long)(e * QEC_EVW + k) * es]; }
  __device__ __forceinline__ void insert(uint32_t op, uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3) {
    if ((a0 | a1 | a2 | a3) == 0u) return;  // identity Paulis only
    int q = n;
    while (q > 0 && at(q - 1, 0) > op) --q;
    if (q > 0 && at(q - 1, 0) == op) {  // same op reached again by a later subset: other bits
      at(q - 1, 1) |= a0; at(q - 1, 2) |= a1; at(q - 1, 3) |= a2; at(q - 1, 4) |= a3;
      return;
    }
    for (int i = n; i > q; --i) {
#pragma unroll
      for (int k = 0; k < QEC_EVW; ++k) at(i, k) = at(i - 1, k);
    }
    at(q, 0) = op; at(q, 1) = a0; at(q, 2) = a1; at(q, 3) = a2; at(q, 4) = a3;
    ++n;
  }
  __device__ __forceinline__ void finish() {
    at(n, 0) = QEC_OP_END;
    at(n + 1, 0) = QEC_OP_END;
  }
};

struct QecCurMask {
  const uint32_t *p;
  long long es;
  uint32_t nop, m0, m1, m2, m3;   // event under the cursor
  uint32_t n_op, n0, n1, n2, n3;  // the one after it, already loaded
  __device__ __forceinline__ void init(const uint32_t *base, long long stride) {
    p = base; es = stride;
    nop = p[0]; m0 = p[es]; m1 = p[2 * es]; m2 = p[3 * es]; m3 = p[4 * es];
    p += QEC_EVW * es;
    n_op = p[0]; n0 = p[es]; n1 = p[2 * es]; n2 = p[3 * es]; n3 = p[4 * es];
  }
  __device__ __forceinline__ void adv() {
    // only called on a hit, i.e. with a real event under the cursor: the event two
    // places on is at worst the second END header
    nop = n_op; m0 = n0; m1 = n1; m2 = n2; m3 = n3;
    p += QEC_EVW * es;
    n_op = p[0]; n0 = p[es]; n1 = p[2 * es]; n2 = p[3 * es]; n3 = p[4 * es];
  }
  __device__ __forceinline__ bool hit(uint32_t i) const { return nop == i; }
  __device__ __forceinline__ bool before(uint32_t end) const { return nop < end; }
  // (ops are unique in the list, so these loops run once; written as loops because ptxas
  // turns a plain `if` body into ~18 predicated instructions that every lane word then pays
  // for at every location, and a loop stays a branch)
  __device__ __forceinline__ void inj1(uint32_t i, uint32_t &xa, uint32_t &za) {
    do {
      xa ^= m0; za ^= m1;
      adv();
    } while (nop == i);
  }
  __device__ __forceinline__ void inj2(uint32_t i, uint32_t &xa, uint32_t &za, uint32_t &xb,
                                       uint32_t &zb) {
    do {
      xa ^= m0; za ^= m1; xb ^= m2; zb ^= m3;
      adv();
    } while (nop == i);
  }
};

// ---- bit events (sampler): (op << 4 | code, 1 << shot_bit), in shared memory --------------
// word k of event e at p[(2 e + k) * QEC_JIT_T]; sorted by op; two 0xFFFFFFFF headers follow.
struct QecCurBit {
  const uint32_t *p;
  uint32_t nop, hdr, msk, nh, nm;
  __device__ __forceinline__ void init(const uint32_t *base) {
    p = base;
    hdr = p[0]; msk = p[QEC_JIT_T]; nop = hdr >> 4;
    p += 2 * QEC_JIT_T;
    nh = p[0]; nm = p[QEC_JIT_T];
  }
  __device__ __forceinline__ void adv() {
    hdr = nh; msk = nm; nop = hdr >> 4;
    p += 2 * QEC_JIT_T;
    nh = p[0]; nm = p[QEC_JIT_T];
  }
  __device__ __forceinline__ bool hit(uint32_t i) const { return nop == i; }
  __device__ __forceinline__ bool before(uint32_t end) const { return nop < end; }
  __device__ __forceinline__ void inj1(uint32_t i, uint32_t &xa, uint32_t &za) {
    do {
      if (hdr & 1u) xa ^= msk;
      if (hdr & 2u) za ^= msk;
      adv();
    } while (nop == i);
  }
  __device__ __forceinline__ void inj2(uint32_t i, uint32_t &xa, uint32_t &za, uint32_t &xb,
                                       uint32_t &zb) {
    do {
      if (hdr & 1u) xa ^= msk;
      if (hdr & 2u) za ^= msk;
      if (hdr & 4u) xb ^= msk;
      if (hdr & 8u) zb ^= msk;
      adv();
    } while (nop == i);
  }
};

// ---- what a final measurement does with its value ------------------------------------------
template <bool TALLY, bool BITS>
__device__ __forceinline__ void qec_jit_out(const uint32_t c, const uint32_t v, const bool valid,
                                            uint32_t *__restrict__ ob, const long long os,
                                            uint32_t *s_tally, uint32_t &any_out) {
  if (BITS) {
    if (valid) ob[(long long)c * os] = v;
  }
  if (TALLY) {
    any_out |= v;
    const int cnt = __reduce_add_sync(0xFFFFFFFFu, __popc(v));
    if ((threadIdx.x & 31u) == 0u && cnt) atomicAdd(&s_tally[c], (uint32_t)cnt);
  }
}

// qec_jit_gates<CUR, TALLY, BITS>(cur, valid, ob, os, s_tally, any_out): the gate stream of
// the circuit at hand, generated by qec_jit.cpp and handed to NVRTC as a virtual header
#include "qec_jit_gates.inc"

__device__ __forceinline__ void qec_jit_flush_tally(const QecOutDev &O, uint32_t *s_tally,
                                                    uint32_t *s_hist, bool with_hist,
                                                    unsigned long long nz_local,
                                                    long long n_paths) {
  unsigned long long nz = nz_local;
  for (int off = 16; off > 0; off >>= 1) nz += __shfl_down_sync(0xFFFFFFFFu, nz, off);
  if ((threadIdx.x & 31) == 0 && nz) atomicAdd(&O.tally[QEC_JIT_NC], nz);
  __syncthreads();
  for (int c = threadIdx.x; c < QEC_JIT_NC; c += QEC_JIT_T) {
    const uint32_t v = s_tally[c];
    if (v) atomicAdd(&O.tally[c], (unsigned long long)v);
  }
  if (with_hist) {
    for (int c = threadIdx.x; c < 64; c += QEC_JIT_T) {
      const uint32_t v = s_hist[c];
      if (v) atomicAdd(&O.tally[QEC_JIT_NC + 2 + c], (unsigned long long)v);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    atomicAdd(&O.tally[QEC_JIT_NC + 1], (unsigned long long)n_paths);
}

#if QEC_JIT_KIND == 1
// ---- FaultEnumerator ---------------------------------------------------------------------------
// The 32 consecutive ranks of a lane word are walked segment by segment: a segment is a
// stretch in which only the LAST location's digit moves (productiterator.cpp:23-29: last digit
// fastest), so its masks come from the tile table with one 64-bit shift and every other
// location contributes a constant code over the segment's bit range.
template <int K>
__device__ __forceinline__ void qec_jit_enum_gen(const QecEnumDev &E, const unsigned long long r0,
                                                 const int nv, QecEvList &l) {
  EnumState<K> s;
  enum_unrank<K>(E, r0, s);
  uint32_t acc[K][4];
#pragma unroll
  for (int j = 0; j < K; ++j) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0u; }
  int b = 0;
  for (;;) {
    const int seg = min((int)(s.np[K - 1] - s.dig[K - 1]), nv - b);
    const uint32_t segmask = (seg >= 32 ? 0xFFFFFFFFu : ((1u << seg) - 1u)) << b;
    const unsigned long long *Wl = E.tile + 4 * (size_t)s.loc[K - 1];
#pragma unroll
    for (int c = 0; c < 4; ++c)
      acc[K - 1][c] |= (((uint32_t)(__ldg(Wl + c) >> s.dig[K - 1])) << b) & segmask;
#pragma unroll
    for (int j = 0; j < K - 1; ++j) {
      const uint32_t code = E.pcode[s.pb[j] + s.dig[j]];
#pragma unroll
      for (int c = 0; c < 4; ++c) acc[j][c] |= (0u - ((code >> c) & 1u)) & segmask;
    }
    b += seg;
    if (b >= nv) break;
    // odometer: the last digit wrapped; carry into the slower digits
    s.dig[K - 1] = 0;
    bool carried = true;
#pragma unroll
    for (int j = K - 2; j >= 0; --j) {
      if (carried) {
        if (++s.dig[j] < s.np[j]) carried = false;
        else s.dig[j] = 0;
      }
    }
    if (carried) {  // every digit wrapped: next k-subset (combinations.cpp:18-32)
      int i = K - 1;
      while (i >= 0 && s.loc[i] >= (uint32_t)(E.L - K + i)) --i;
      if (i < 0) break;  // past the last path
      const uint32_t base = s.loc[i] + 1;
#pragma unroll
      for (int j = 0; j < K; ++j) {
        if (j >= i) {
          l.insert(s.op[j], acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
          acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0u;
          s.loc[j] = base + (uint32_t)(j - i);
          enum_load<K>(E, s, j);
        }
      }
    }
  }
#pragma unroll
  for (int j = 0; j < K; ++j) l.insert(s.op[j], acc[j][0], acc[j][1], acc[j][2], acc[j][3]);
}

extern "C" __global__ void __launch_bounds__(QEC_JIT_T, QEC_JIT_MINB)
qec_jit_kernel(const QecOutDev O, const QecEnumDev E, const long long lo, const long long hi,
               uint32_t *__restrict__ scratch) {
  extern __shared__ __align__(16) uint32_t qec_smem[];
  uint32_t *s_tally = qec_smem;
  if (QEC_JIT_TALLY) {
    for (int c = threadIdx.x; c < QEC_JIT_NC; c += QEC_JIT_T) s_tally[c] = 0;
    __syncthreads();
  }
  const long long es = (long long)gridDim.x * QEC_JIT_T;
  uint32_t *ev = scratch + (long long)blockIdx.x * QEC_JIT_T + threadIdx.x;
  unsigned long long nz = 0;
  for (long long ch = blockIdx.x; ch * QEC_JIT_T < O.n_words; ch += gridDim.x) {
    const long long w = ch * QEC_JIT_T + threadIdx.x;
    const bool valid = w < O.n_words;
    QecEvList l;
    l.p = ev; l.es = es; l.n = 0;
    if (valid) {
      const long long r0 = lo + 32 * w;
      const int nv = (int)min((long long)32, hi - r0);
      qec_jit_enum_gen<QEC_JIT_ORDER>(E, (unsigned long long)r0, nv, l);
    }
    l.finish();
    QecCurMask cur;
    cur.init(ev, es);
    uint32_t any = 0;
    qec_jit_gates<QecCurMask, QEC_JIT_TALLY != 0, QEC_JIT_BITS != 0>(
        cur, valid, O.bits + w, O.n_words, s_tally, any);
    if (QEC_JIT_TALLY) nz += __popc(any);
  }
  if (QEC_JIT_TALLY) qec_jit_flush_tally(O, s_tally, nullptr, false, nz, hi - lo);
}
#endif  // enumerator

#if QEC_JIT_KIND == 2
// ---- FaultSampler ------------------------------------------------------------------------------
// Same random decisions, in the same order, as k_sample in qec_kernels.cu (so both kernels
// return identical bits for one (seed, shot id)); the events land in shared memory as
// (op << 4 | code, 1 << shot_bit) pairs.  `cap` = events a lane word can hold.
extern "C" __global__ void __launch_bounds__(QEC_JIT_T, QEC_JIT_MINB)
qec_jit_kernel(const QecOutDev O, const QecSamplerDev S, const long long word_lo,
               const long long n_shots, const int cap, uint32_t *__restrict__ overflow_flag) {
  extern __shared__ __align__(16) uint32_t qec_smem[];
  uint32_t *s_tally = qec_smem;
  uint32_t *s_hist = s_tally + QEC_JIT_NC;
  uint32_t *s_T1 = s_hist + 64;
  uint32_t *s_T2 = s_T1 + S.K1;
  uint32_t *s_ev = s_T2 + S.B;  // [2 * (cap + 2)][QEC_JIT_T]
  for (int c = threadIdx.x; c < QEC_JIT_NC + 64; c += QEC_JIT_T) s_tally[c] = 0;
  for (uint32_t i = threadIdx.x; i < S.K1; i += QEC_JIT_T) s_T1[i] = S.T1[i];
  for (uint32_t i = threadIdx.x; i < S.B; i += QEC_JIT_T) s_T2[i] = S.T2[i];
  __syncthreads();
  uint32_t *ev = s_ev + threadIdx.x;
  unsigned long long nz = 0;
  for (long long ch = blockIdx.x; ch * QEC_JIT_T < O.n_words; ch += gridDim.x) {
    const long long w = ch * QEC_JIT_T + threadIdx.x;
    const bool valid = w < O.n_words;
    int n_ev = 0;    // faults drawn (for the fault-count histogram and the overflow check)
    int n_st = 0;    // events stored
    if (valid) {
      const int nv = (int)min((long long)32, n_shots - 32 * w);
      const unsigned long long gw = (unsigned long long)(word_lo + w);
      const uint32_t wlo = (uint32_t)gw, whi = (uint32_t)(gw >> 32);
      uint32_t draw = 0;
      long long slot = -1;
      uint32_t cp[6] = {0, 0, 0, 0, 0, 0};  // bit-sliced per-shot fault counters
      for (;;) {
        QecPhilox r;
        unsigned long long skip = 0;
        for (;;) {
          r = qec_philox(wlo, whi, draw++, 0u, S.k0, S.k1);
          const uint32_t a = qec_count_greater(s_T1, S.K1, r.c[0]);
          skip += (unsigned long long)a * S.B;
          if (a < S.K1 || skip >= (unsigned long long)S.n_slots) break;
        }
        const uint32_t b = qec_count_greater(s_T2 + 1, S.B - 1, r.c[1]);
        slot += 1 + (long long)skip + b;
        if (slot >= S.n_slots) break;
        const uint32_t sl = (uint32_t)(slot >> 5), bit = (uint32_t)slot & 31u;
        const uint32_t cls = S.slcls[sl];
        if (!S.cls_always[cls] && r.c[2] >= S.cls_acc[cls]) continue;
        const uint32_t base = S.cls_off[cls], np = S.cls_off[cls + 1] - base;
        uint32_t d = 0;
        while (d + 1 < np && r.c[3] >= S.cdf[base + d]) ++d;
        if ((int)bit >= nv) continue;  // shot past the end of the range
        const uint32_t code = S.codes[base + d];
        if (code != 0u && n_st < cap) {
          ev[(2 * n_st) * QEC_JIT_T] = (S.slop[sl] << 4) | code;
          ev[(2 * n_st + 1) * QEC_JIT_T] = 1u << bit;
          ++n_st;
        } else if (code != 0u) {
          n_st = cap + 1;  // remembered as an overflow below
        }
        ++n_ev;
        if (QEC_JIT_TALLY) {
          uint32_t carry = 1u << bit;
#pragma unroll
          for (int p = 0; p < 6; ++p) {
            const uint32_t t = cp[p] & carry;
            cp[p] ^= carry;
            carry = t;
          }
          if (carry) {
#pragma unroll
            for (int p = 0; p < 6; ++p) cp[p] |= carry;  // saturate at 63
          }
        }
      }
      if (n_st > cap) {
        atomicMax(overflow_flag, (uint32_t)n_ev);
        n_st = cap;
      }
      if (QEC_JIT_TALLY) {
        for (int bit = 0; bit < nv; ++bit) {
          uint32_t k = 0;
#pragma unroll
          for (int p = 0; p < 6; ++p) k |= ((cp[p] >> bit) & 1u) << p;
          atomicAdd(&s_hist[k], 1u);
        }
      }
    }
    ev[(2 * n_st) * QEC_JIT_T] = 0xFFFFFFFFu;
    ev[(2 * n_st + 2) * QEC_JIT_T] = 0xFFFFFFFFu;
    QecCurBit cur;
    cur.init(ev);
    uint32_t any = 0;
    qec_jit_gates<QecCurBit, QEC_JIT_TALLY != 0, QEC_JIT_BITS != 0>(
        cur, valid, O.bits + w, O.n_words, s_tally, any);
    if (QEC_JIT_TALLY) nz += __popc(any);
  }
  if (QEC_JIT_TALLY) qec_jit_flush_tally(O, s_tally, s_hist, true, nz, n_shots);
}
#endif  // sampler
