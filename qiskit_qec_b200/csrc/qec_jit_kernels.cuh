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
#define QEC_NWARP (QEC_JIT_T / 32)
#define QEC_TALLYW (QEC_JIT_NC * QEC_NWARP)  // shared tally words: one row of clbit counts per warp
#define QEC_BMW (QEC_JIT_NSEG * QEC_JIT_SEGW)  // words of a lane word's "an event sits on op i" bitmap

// How a lane word meets its events in the generated gate stream
// ------------------------------------------------------------
// Every lane word (thread) owns a bitmap over the op indices, one bit per op that carries
// an event of this lane word, at s_bm[(i >> 5) * QEC_JIT_T + tid], and a queue of mask
// entries in op order.  The generated code loads bitmap word k once per 32 ops
// (`cur.word(k)`), tests bit j of it at a fault location (one LOP3 with a predicate result),
// votes (`__any_sync`) so that the warp skips the few instructions of the injection as a
// whole, and the lanes that hit pop their next queue entry and XOR it into the (statically
// named) frame registers.  No per-event op compare, no loop, no out-of-line code: the hit
// path is 4 (one-qubit op) or 7 (cx) predicated instructions placed right behind the gate.

// ---- mask events (enumerator): one event per distinct op, 32 path bits per mask ----------
// word k of event e lives at p[(5 e + k) * es] (es = threads in the grid: coalesced).
// The list is sorted by op and ops are unique.
struct QecEvList {
  uint32_t *p;
  long long es;
  uint32_t *bmw;   // this lane word's bitmap column: the bit of an op is set when its entry is made
  int n;
  uint32_t last;   // op of the last entry: an op beyond it is appended without reading the list
  __device__ __forceinline__ uint32_t &at(int e, int k) { return p[(long long)(e * QEC_EVW + k) * es]; }
  __device__ __forceinline__ void insert(uint32_t op, uint32_t a0, uint32_t a1, uint32_t a2,
                                         uint32_t a3) {
    if ((a0 | a1 | a2 | a3) == 0u) return;  // identity Paulis only
    int q = n;
    if (n > 0 && op <= last) {  // (rare: the slower locations of a later subset)
      while (q > 0 && at(q - 1, 0) > op) --q;
      if (q > 0 && at(q - 1, 0) == op) {  // same op reached again by a later subset: other bits
        at(q - 1, 1) |= a0; at(q - 1, 2) |= a1; at(q - 1, 3) |= a2; at(q - 1, 4) |= a3;
        return;
      }
      for (int i = n; i > q; --i) {
#pragma unroll
        for (int k = 0; k < QEC_EVW; ++k) at(i, k) = at(i - 1, k);
      }
    }
    if (q == n) last = op;
    at(q, 0) = op; at(q, 1) = a0; at(q, 2) = a1; at(q, 3) = a2; at(q, 4) = a3;
    bmw[(op >> 5) * QEC_JIT_T] |= 1u << (op & 31u);
    ++n;
  }
};

// cursor over a QecEvList (global memory, L2 resident): the entry under the cursor is held in
// registers and the next one is fetched when it is consumed, so a hit never waits for memory
struct QecCurMask {
  bool lane0;
  const uint32_t *bmp;  // this lane word's bitmap column in shared memory
  const uint32_t *p;
  long long es;
  uint32_t bm, m0, m1, m2, m3;
  __device__ __forceinline__ void init(const uint32_t *bitmap, const uint32_t *base, long long stride) {
    bmp = bitmap; p = base; es = stride; bm = 0u; lane0 = (threadIdx.x & 31u) == 0u;
    m0 = p[es]; m1 = p[2 * es]; m2 = p[3 * es]; m3 = p[4 * es];
  }
  __device__ __forceinline__ void word(int k) { bm = bmp[k * QEC_JIT_T]; }
  __device__ __forceinline__ bool hit(uint32_t bit) const { return (bm & bit) != 0u; }
  __device__ __forceinline__ bool any(uint32_t mask) const { return (bm & mask) != 0u; }
  __device__ __forceinline__ void adv() {  // (reads one entry past the list at the end: scratch has room)
    p += QEC_EVW * es;
    m0 = p[es]; m1 = p[2 * es]; m2 = p[3 * es]; m3 = p[4 * es];
  }
  __device__ __forceinline__ void inj1(uint32_t &xa, uint32_t &za) {
    xa ^= m0; za ^= m1;
    adv();
  }
  __device__ __forceinline__ void inj2(uint32_t &xa, uint32_t &za, uint32_t &xb, uint32_t &zb) {
    xa ^= m0; za ^= m1; xb ^= m2; zb ^= m3;
    adv();
  }
};

// ---- mask queue (sampler), in shared memory, one contiguous run per lane word --------------
// entry of a one-qubit op: {X mask, Z mask}; of a cx: {X a, Z a, X b, Z b}; entries in op order
struct QecCurQueue {
  bool lane0;
  const uint32_t *bmp;
  const uint32_t *pe;
  uint32_t bm;
  __device__ __forceinline__ void init(const uint32_t *bitmap, const uint32_t *queue) {
    bmp = bitmap; pe = queue; bm = 0u; lane0 = (threadIdx.x & 31u) == 0u;
  }
  __device__ __forceinline__ void word(int k) { bm = bmp[k * QEC_JIT_T]; }
  __device__ __forceinline__ bool hit(uint32_t bit) const { return (bm & bit) != 0u; }
  __device__ __forceinline__ bool any(uint32_t mask) const { return (bm & mask) != 0u; }
  __device__ __forceinline__ void inj1(uint32_t &xa, uint32_t &za) {
    const uint2 t = *reinterpret_cast<const uint2 *>(pe);
    pe += 2;
    xa ^= t.x; za ^= t.y;
  }
  __device__ __forceinline__ void inj2(uint32_t &xa, uint32_t &za, uint32_t &xb, uint32_t &zb) {
    const uint2 t = *reinterpret_cast<const uint2 *>(pe);
    const uint2 u = *reinterpret_cast<const uint2 *>(pe + 2);
    pe += 4;
    xa ^= t.x; za ^= t.y; xb ^= u.x; zb ^= u.y;
  }
};

// ---- what a final measurement does with its value ------------------------------------------
// (the marginal of clbit c: POPC, one REDUX warp sum, and lane 0 adds it to its WARP's row of
// the shared tally, s_tally[warp][c] -- a plain read-modify-write, no atomic: ptxas expands a
// shared-memory RED on a uniform address into a ~12-instruction warp aggregation)
template <bool TALLY, bool BITS>
__device__ __forceinline__ void qec_jit_out(const uint32_t c, const uint32_t v, const bool valid,
                                            const bool lane0, uint32_t *__restrict__ ob,
                                            const long long os, uint32_t *s_tally,
                                            uint32_t &any_out) {
  if (BITS) {
    if (valid) ob[(long long)c * os] = v;
  }
  if (TALLY) {
    any_out |= v;
    const uint32_t cnt = (uint32_t)__reduce_add_sync(0xFFFFFFFFu, __popc(v));
    if (lane0) s_tally[c] += cnt;
  }
}

// struct QecFrame (every frame word a member, i.e. a register), qec_jit_frame_zero(F) and
// qec_jit_seg<CUR, TALLY, BITS>(seg, F, cur, valid, ob, os, s_tally, any_out): segment `seg`
// (QEC_JIT_SEGW * 32 ops) of the gate stream of the circuit at hand, generated by qec_jit.cpp
// and handed to NVRTC as a virtual header
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
    uint32_t v = 0;
#pragma unroll
    for (int wp = 0; wp < QEC_NWARP; ++wp) v += s_tally[wp * QEC_JIT_NC + c];
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
// Returns false when the lane word would need more than `max_seg` segments (its 32 paths step
// through a run of one-Pauli locations): such words are rare but cost a warp -- and, through
// the CTA barriers, the whole CTA -- ten times the usual prologue, so the kernel leaves them
// to a second pass of the table-driven kernel over the list of deferred words.
template <int K>
__device__ __forceinline__ bool qec_jit_enum_gen(const QecEnumDev &E, const unsigned long long r0,
                                                 const int nv, QecEvList &l, const int max_seg) {
  EnumState<K> s;
  enum_unrank<K>(E, r0, s);
  uint32_t acc[K][4];
#pragma unroll
  for (int j = 0; j < K; ++j) { acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0u; }
  int b = 0, n_seg = 0;
  for (;;) {
    if (++n_seg > max_seg) return false;
    const int seg = min((int)(s.np[K - 1] - s.dig[K - 1]), nv - b);
    const uint32_t segmask = (seg >= 32 ? 0xFFFFFFFFu : ((1u << seg) - 1u)) << b;
    const unsigned long long *Wl = E.tile + 4 * (size_t)s.loc[K - 1];
#pragma unroll
    for (int c = 0; c < 4; ++c)
      acc[K - 1][c] |= (((uint32_t)(Wl[c] >> s.dig[K - 1])) << b) & segmask;
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
  return true;
}

// The enumerator's tables (suffix sums E, per-location Pauli counts / ops / code lists, tile
// masks) are copied to shared memory once per CTA when the host found room for them
// (`tab_in_smem`): unranking is a chain of dependent binary-search loads, 30 cycles each from
// shared memory against ~300+ from L2.
extern "C" __global__ void __launch_bounds__(QEC_JIT_T, QEC_JIT_MINB)
qec_jit_kernel(const QecOutDev O, const QecEnumDev Eg, const long long lo, const long long hi,
               uint32_t *__restrict__ scratch, const int tab_in_smem, const int max_seg,
               long long *__restrict__ redo_list, unsigned int *__restrict__ redo_count) {
  extern __shared__ __align__(16) uint32_t qec_smem[];
  uint32_t *s_tally = qec_smem;
  uint32_t *w_tally = s_tally + (threadIdx.x >> 5) * QEC_JIT_NC;  // this warp's row
  uint32_t *bmw = qec_smem + QEC_TALLYW + threadIdx.x;  // bitmap column of this lane word
  QecEnumDev E = Eg;
  if (tab_in_smem) {
    uint32_t *t = qec_smem + QEC_TALLYW + QEC_BMW * QEC_JIT_T;
    t += (uint32_t)(t - qec_smem) & 1u;  // 8-byte aligned
    const int L = Eg.L, nE = (QEC_JIT_ORDER + 1) * (L + 2), nP = (int)Eg.poff[L];
    unsigned long long *sE = reinterpret_cast<unsigned long long *>(t);
    unsigned long long *sTile = sE + nE;
    uint32_t *sNp = reinterpret_cast<uint32_t *>(sTile + 4 * L), *sPoff = sNp + L, *sLop = sPoff + L + 1;
    uint8_t *sCode = reinterpret_cast<uint8_t *>(sLop + L);
    for (int i = threadIdx.x; i < nE; i += QEC_JIT_T) sE[i] = Eg.E[i];
    for (int i = threadIdx.x; i < 4 * L; i += QEC_JIT_T) sTile[i] = Eg.tile[i];
    for (int i = threadIdx.x; i < L; i += QEC_JIT_T) { sNp[i] = Eg.npaul[i]; sLop[i] = Eg.lop[i]; }
    for (int i = threadIdx.x; i <= L; i += QEC_JIT_T) sPoff[i] = Eg.poff[i];
    for (int i = threadIdx.x; i < nP; i += QEC_JIT_T) sCode[i] = Eg.pcode[i];
    E.E = sE; E.tile = sTile; E.npaul = sNp; E.poff = sPoff; E.lop = sLop; E.pcode = sCode;
  }
  if (QEC_JIT_TALLY) {
    for (int c = threadIdx.x; c < QEC_TALLYW; c += QEC_JIT_T) s_tally[c] = 0;
  }
  __syncthreads();
  const long long es = (long long)gridDim.x * QEC_JIT_T;
  uint32_t *ev = scratch + (long long)blockIdx.x * QEC_JIT_T + threadIdx.x;
  unsigned long long nz = 0;
  for (long long ch = blockIdx.x; ch * QEC_JIT_T < O.n_words; ch += gridDim.x) {
    const long long w = ch * QEC_JIT_T + threadIdx.x;
    bool valid = w < O.n_words;
    QecEvList l;
    l.p = ev; l.es = es; l.n = 0; l.bmw = bmw; l.last = 0u;
#pragma unroll
    for (int k = 0; k < QEC_BMW; ++k) bmw[k * QEC_JIT_T] = 0u;
    if (valid) {
      const long long r0 = lo + 32 * w;
      const int nv = (int)min((long long)32, hi - r0);
      if (!qec_jit_enum_gen<QEC_JIT_ORDER>(E, (unsigned long long)r0, nv, l, max_seg)) {
        // deferred: this lane word runs empty here (zero frame, no stores, nothing tallied)
        redo_list[atomicAdd(redo_count, 1u)] = w;
        valid = false;
#pragma unroll
        for (int k = 0; k < QEC_BMW; ++k) bmw[k * QEC_JIT_T] = 0u;
      }
    }
    QecCurMask cur;
    cur.init(bmw, ev, es);
    uint32_t any = 0;
    QecFrame F;
    qec_jit_frame_zero(F);
#pragma unroll
    for (int seg = 0; seg < QEC_JIT_NSEG; ++seg) {
      cur.bmp = bmw + seg * QEC_JIT_SEGW * QEC_JIT_T;
      qec_jit_seg<QecCurMask, QEC_JIT_TALLY != 0, QEC_JIT_BITS != 0>(
          seg, F, cur, valid, O.bits + w, O.n_words, w_tally, any);
    }
    if (QEC_JIT_TALLY) nz += __popc(any);
  }
  if (QEC_JIT_TALLY) qec_jit_flush_tally(O, s_tally, nullptr, false, nz, hi - lo);
}
#endif  // enumerator

#if QEC_JIT_KIND == 2
// ---- FaultSampler ------------------------------------------------------------------------------
// Same random decisions, in the same order, as k_sample in qec_kernels.cu (so both kernels
// return identical bits for one (seed, shot id)); only the table lookups are quicker here: the
// sampler's small tables sit in shared memory at compile-time offsets (the table sizes are
// constants of the generated kernel), the second-level gap comes from a guide table plus
// ~1 step instead of a binary search, and the Pauli digit from an unrolled binary search over
// the class's cdf padded to P = 2^QEC_JIT_LGP entries.  All of them return exactly what
// qec_count_greater / the linear cdf scan return.
//
// The gate stream runs segment by segment: before a segment the lane word draws the faults
// that fall on its ops into its mask queue (`capw` words of shared memory) and bitmap, so the
// queue only ever holds one segment's faults.  `cap` = events per segment the host sized the
// queue for (reported back on overflow).
#define QEC_P (1u << QEC_JIT_LGP)
#define QEC_TAB_T1 0u
#define QEC_TAB_T2 (QEC_TAB_T1 + QEC_JIT_K1)
#define QEC_TAB_G (QEC_TAB_T2 + QEC_JIT_B)        // uint16[2 B]
#define QEC_TAB_INFO (QEC_TAB_G + QEC_JIT_B)
#define QEC_TAB_CACC (QEC_TAB_INFO + QEC_JIT_L)
#define QEC_TAB_CNP1 (QEC_TAB_CACC + QEC_JIT_NCLS)
#define QEC_TAB_CDF (QEC_TAB_CNP1 + QEC_JIT_NCLS)
#define QEC_TAB_CODE (QEC_TAB_CDF + QEC_JIT_NCLS * QEC_P)  // uint8[n][P]
#define QEC_TAB_WORDS (QEC_TAB_CODE + (QEC_JIT_NCLS * QEC_P + 3u) / 4u)

__device__ __forceinline__ uint32_t qec_jit_gshift() {  // 2 B = 2^(32 - gshift) guide buckets
  uint32_t g = 32;
  while ((1u << (32 - g)) < 2u * QEC_JIT_B) --g;
  return g;
}

__device__ __forceinline__ void qec_jit_tabs_fill(uint32_t *tab, const QecSamplerDev &S) {
  const uint32_t gshift = qec_jit_gshift();
  uint16_t *sG = reinterpret_cast<uint16_t *>(tab + QEC_TAB_G);
  uint8_t *scode = reinterpret_cast<uint8_t *>(tab + QEC_TAB_CODE);
  for (uint32_t i = threadIdx.x; i < QEC_JIT_K1; i += QEC_JIT_T) tab[QEC_TAB_T1 + i] = S.T1[i];
  for (uint32_t i = threadIdx.x; i < QEC_JIT_B; i += QEC_JIT_T) tab[QEC_TAB_T2 + i] = S.T2[i];
  for (uint32_t k = threadIdx.x; k < 2u * QEC_JIT_B; k += QEC_JIT_T) {
    // entries of T2[1..B) greater than the largest r of bucket k: where the scan for any r of
    // the bucket may start
    const uint32_t rmax = (k << gshift) | ((1u << gshift) - 1u);
    sG[k] = (uint16_t)qec_count_greater(S.T2 + 1, QEC_JIT_B - 1u, rmax);
  }
  for (uint32_t i = threadIdx.x; i < QEC_JIT_L; i += QEC_JIT_T) tab[QEC_TAB_INFO + i] = S.slinf[i];
  for (uint32_t c = threadIdx.x; c < QEC_JIT_NCLS; c += QEC_JIT_T) {
    tab[QEC_TAB_CACC + c] = S.cls_acc[c];
    tab[QEC_TAB_CNP1 + c] = S.cls_off[c + 1] - S.cls_off[c] - 1u;
  }
  for (uint32_t i = threadIdx.x; i < QEC_JIT_NCLS * QEC_P; i += QEC_JIT_T) {
    const uint32_t c = i >> QEC_JIT_LGP, d = i & (QEC_P - 1u), o = S.cls_off[c], n = S.cls_off[c + 1] - o;
    tab[QEC_TAB_CDF + i] = d < n ? S.cdf[o + d] : 0xFFFFFFFFu;
    scode[i] = d < n ? S.codes[o + d] : (uint8_t)0;
  }
}

// the draw stream of one 32-shot word: candidate slots in increasing order
struct QecRng {
  uint32_t wlo, whi, draw;
  int slot;          // candidate under the cursor (QEC_JIT_NSLOTS: none left)
  uint32_t r2, r3;   // its accept / digit words
};

__device__ __forceinline__ void qec_rng_advance(QecRng &g, const uint32_t *tab, const uint32_t k0,
                                                const uint32_t k1) {
  const uint16_t *sG = reinterpret_cast<const uint16_t *>(tab + QEC_TAB_G);
  QecPhilox r;
  uint32_t skip = 0;  // (saturates at QEC_JIT_NSLOTS < 2^24: 32 bits are plenty)
  for (;;) {
    r = qec_philox(g.wlo, g.whi, g.draw++, 0u, k0, k1);
    uint32_t a = 0;
    if (tab[QEC_TAB_T1] > r.c[0]) {  // first-level gap > 0: uncommon by the choice of B
      a = 1;
      while (a < QEC_JIT_K1 && tab[QEC_TAB_T1 + a] > r.c[0]) ++a;
      skip = min(skip + a * QEC_JIT_B, (uint32_t)QEC_JIT_NSLOTS);
    }
    if (a < QEC_JIT_K1 || skip >= (uint32_t)QEC_JIT_NSLOTS) break;  // a == K1: truncated table, draw on
  }
  uint32_t b = sG[r.c[1] >> qec_jit_gshift()];
  while (b < QEC_JIT_B - 1u && tab[QEC_TAB_T2 + 1u + b] > r.c[1]) ++b;
  g.slot = (int)min((uint32_t)(g.slot + 1) + skip + b, (uint32_t)QEC_JIT_NSLOTS);
  g.r2 = r.c[2];
  g.r3 = r.c[3];
}

// class thinning and Pauli digit of the candidate under the cursor; false = no fault there
__device__ __forceinline__ bool qec_rng_resolve(const QecRng &g, const uint32_t *tab, uint32_t &info,
                                                uint32_t &code) {
  info = tab[QEC_TAB_INFO + ((uint32_t)g.slot >> 5)];
  const uint32_t cls = (info >> 20) & 1023u;
  if (!(info >> 31) && g.r2 >= tab[QEC_TAB_CACC + cls]) return false;
  const uint32_t *c = tab + QEC_TAB_CDF + (cls << QEC_JIT_LGP);
  uint32_t d = 0;
#pragma unroll
  for (uint32_t step = QEC_P >> 1; step; step >>= 1)
    if (c[d + step - 1u] <= g.r3) d += step;
  d = min(d, tab[QEC_TAB_CNP1 + cls]);
  code = reinterpret_cast<const uint8_t *>(tab + QEC_TAB_CODE)[(cls << QEC_JIT_LGP) + d];
  return true;
}

extern "C" __global__ void __launch_bounds__(QEC_JIT_T, QEC_JIT_MINB)
qec_jit_kernel(const QecOutDev O, const QecSamplerDev S, const long long word_lo,
               const long long n_shots, const int cap, const int capw,
               uint32_t *__restrict__ overflow_flag) {
  extern __shared__ __align__(16) uint32_t qec_smem[];
  uint32_t *s_tally = qec_smem;
  uint32_t *w_tally = s_tally + (threadIdx.x >> 5) * QEC_JIT_NC;  // this warp's row
  uint32_t *s_hist = s_tally + QEC_TALLYW;
  uint32_t *s_tab = s_hist + 64;
  uint32_t *s_bm = s_tab + QEC_TAB_WORDS;                // [QEC_JIT_SEGW][QEC_JIT_T]
  uint32_t *s_q = s_bm + QEC_JIT_SEGW * QEC_JIT_T;       // [QEC_JIT_T][capw]
  s_q += (4u - ((uint32_t)(s_q - qec_smem) & 3u)) & 3u;  // 16-byte aligned
  qec_jit_tabs_fill(s_tab, S);
  for (int c = threadIdx.x; c < QEC_TALLYW + 64; c += QEC_JIT_T) s_tally[c] = 0;
  __syncthreads();
  const uint32_t k0 = S.k0, k1 = S.k1;
  uint32_t *bmw = s_bm + threadIdx.x;
  uint32_t *q = s_q + (size_t)threadIdx.x * capw;
  unsigned long long nz = 0;
  for (long long ch = blockIdx.x; ch * QEC_JIT_T < O.n_words; ch += gridDim.x) {
    const long long w = ch * QEC_JIT_T + threadIdx.x;
    const bool valid = w < O.n_words;
    const int nv = valid ? (int)min((long long)32, n_shots - 32 * w) : 0;
    QecRng g;
    {
      const unsigned long long gw = (unsigned long long)(word_lo + w);
      g.wlo = (uint32_t)gw; g.whi = (uint32_t)(gw >> 32); g.draw = 0u; g.slot = -1;
      g.r2 = g.r3 = 0u;
    }
    if (valid) qec_rng_advance(g, s_tab, k0, k1);
    else g.slot = QEC_JIT_NSLOTS;
    int n_ev = 0;  // faults drawn (fault-count histogram, overflow report)
    bool overflow = false;
    uint32_t cp[6] = {0, 0, 0, 0, 0, 0};  // bit-sliced per-shot fault counters
    uint32_t any = 0;
    QecFrame F;
    qec_jit_frame_zero(F);
#pragma unroll
    for (int seg = 0; seg < QEC_JIT_NSEG; ++seg) {
#pragma unroll
      for (int k = 0; k < QEC_JIT_SEGW; ++k) bmw[k * QEC_JIT_T] = 0u;
      const int slot_end = (int)S.seg_end[seg];
      const uint32_t op0 = (uint32_t)seg * (QEC_JIT_SEGW * 32u);
      int nw = 0;              // queue words in use
      uint32_t *ent = q;       // the entry being filled
      uint32_t last_op = 0xFFFFFFFFu;
      while (g.slot < slot_end) {
        uint32_t info, code;
        const uint32_t bit = (uint32_t)g.slot & 31u;
        if (qec_rng_resolve(g, s_tab, info, code) && (int)bit < nv) {
          ++n_ev;
          const uint32_t msk = 1u << bit;
          if (code != 0u) {
            const uint32_t op = info & 0xFFFFFu;
            const uint32_t m0 = (code & 1u) ? msk : 0u, m1 = (code & 2u) ? msk : 0u;
            const uint32_t m2 = (code & 4u) ? msk : 0u, m3 = (code & 8u) ? msk : 0u;
            const bool two = (info >> 30) & 1u;
            if (op != last_op) {  // a new entry (events arrive in op order)
              const int len = two ? 4 : 2;
              if (nw + len > capw) {
                overflow = true;
              } else {
                ent = q + nw;
                nw += len;
                last_op = op;
                *reinterpret_cast<uint2 *>(ent) = make_uint2(m0, m1);
                if (two) *reinterpret_cast<uint2 *>(ent + 2) = make_uint2(m2, m3);
                const uint32_t rel = op - op0;
                bmw[(rel >> 5) * QEC_JIT_T] |= 1u << (rel & 31u);
              }
            } else if (!overflow) {  // another shot of the same lane word on the same op
              ent[0] |= m0; ent[1] |= m1;
              if (two) { ent[2] |= m2; ent[3] |= m3; }
            }
          }
          if (QEC_JIT_TALLY) {
            uint32_t carry = msk;
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
        qec_rng_advance(g, s_tab, k0, k1);
      }
      QecCurQueue cur;
      cur.init(bmw, q);
      qec_jit_seg<QecCurQueue, QEC_JIT_TALLY != 0, QEC_JIT_BITS != 0>(
          seg, F, cur, valid, O.bits + w, O.n_words, w_tally, any);
    }
    if (overflow) atomicMax(overflow_flag, (uint32_t)max(n_ev, cap + 1));
    if (QEC_JIT_TALLY) {
      nz += __popc(any);
      // faults-per-shot histogram: shots with 0..3 faults are counted with bit logic on the
      // counter planes and one warp sum per bin; the few shots with more go one by one
      const uint32_t live = nv >= 32 ? 0xFFFFFFFFu : ((1u << nv) - 1u);
      const uint32_t hi4 = cp[2] | cp[3] | cp[4] | cp[5];
      const uint32_t lo4 = live & ~hi4;
      const int c0 = __reduce_add_sync(0xFFFFFFFFu, __popc(lo4 & ~cp[1] & ~cp[0]));
      const int c1 = __reduce_add_sync(0xFFFFFFFFu, __popc(lo4 & ~cp[1] & cp[0]));
      const int c2 = __reduce_add_sync(0xFFFFFFFFu, __popc(lo4 & cp[1] & ~cp[0]));
      const int c3 = __reduce_add_sync(0xFFFFFFFFu, __popc(lo4 & cp[1] & cp[0]));
      if ((threadIdx.x & 31u) == 0u) {
        atomicAdd(&s_hist[0], (uint32_t)c0); atomicAdd(&s_hist[1], (uint32_t)c1);
        atomicAdd(&s_hist[2], (uint32_t)c2); atomicAdd(&s_hist[3], (uint32_t)c3);
      }
      for (uint32_t m = live & hi4; m; m &= m - 1u) {
        const int bit = __ffs(m) - 1;
        uint32_t k = 0;
#pragma unroll
        for (int p = 0; p < 6; ++p) k |= ((cp[p] >> bit) & 1u) << p;
        atomicAdd(&s_hist[k], 1u);
      }
    }
  }
  if (QEC_JIT_TALLY) qec_jit_flush_tally(O, s_tally, s_hist, true, nz, n_shots);
}
#endif  // sampler
