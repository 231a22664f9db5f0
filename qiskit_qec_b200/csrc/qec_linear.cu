// qec_linear.cu -- linear-response fault sampler for sm_100a (FaultSampler::sample, bulk path).
//
// Frame propagation is linear over GF(2): cx / h / s are linear maps of the frame, reset is a
// projection, measure reads one frame bit, a fault XORs a Pauli in (reference
// errorpropagator.cpp:211-254, :272-310).  So
//     outcome(shot) = XOR over the shot's faults f of signature(f),
// signature(f) = the outcome of the circuit with the single fault f, which qec_capi.cu
// computes once per sampler with the frame-propagation kernel (k_propagate_batch).  A shot at
// p = 1e-3 carries ~0.8 faults, each flipping a handful of clbits, against ~760 gates per shot
// for explicit propagation: this kernel never touches the gate list.
//
// Kernels in this file (all share the signature tables built by qec_capi.cu):
//   k_sample_lanes  lane-stream sampler (default where it fits): a warp owns 1024 shots, lane l
//                   owns word column l and walks its own fault stream; plain load/xor/store on
//                   acc[row][l], no shared-memory atomics (the atomic unit bounded the two
//                   variants below at ~2 cycles per lane)
//   k_sample_linw   warp variant: one warp per 256-shot batch, one stream, warp prefix sum of
//                   the gaps, red.shared.xor into the warp's accumulator
//   k_sample_lin    CTA variant for circuits with many clbits (d=7, d=11): one CTA of 1024 threads per
//                   batch of 2^lgb shots, one stream per warp, CTA barrier, chain-walk read-out
//                   (lin_readout: running XOR in a register, row sums in registers)
//   k_enum_lanes    the same idea for FaultEnumerator: outcome(path) = XOR of the signatures of
//                   its faults; lane l owns the 32 paths of output word l
// Draw schemes (QecLinDev::draw): computed 32-bit geometric gaps (qec_gap_len), two candidates
// per Philox call when no class needs thinning; two-level gap tables for very small p.
#include <cuda_runtime.h>
#include "qec_kernels.cuh"

namespace {

struct LinSmem {
  uint32_t *acc, *cnt, *anyw, *tally, *hist, *T1, *T2, *cacc, *cnp1, *cdf;
  uint16_t *G;
  const uint32_t *info, *desc;
  const uint16_t *list;
};

__device__ __forceinline__ LinSmem lin_carve(uint32_t *sm, const QecLinDev &S, int T) {
  LinSmem m;
  const size_t BW = (size_t)1 << (S.lgb - 5);
  m.acc = sm;
  m.cnt = m.acc + ((size_t)S.nc + 1u) * BW;  // (row nc: always zero, the target of the read-out's pad entries)
  m.anyw = m.cnt + ((size_t)1 << (S.lgb - 1));
  m.tally = m.anyw + BW;
  m.hist = m.tally + (size_t)S.nc;  // one tally row per CTA (red.shared.add from any warp)
  m.T1 = m.hist + 64;
  m.T2 = m.T1 + S.K1;
  m.G = reinterpret_cast<uint16_t *>(m.T2 + S.B + 1u);
  m.cacc = m.T2 + 2u * S.B + 1u;
  m.cnp1 = m.cacc + S.n_classes;
  m.cdf = m.cnp1 + S.n_classes;
  uint32_t *p = m.cdf + ((size_t)S.n_classes << S.lgP);
  if (S.info_in_smem) { m.info = p; p += S.L; } else m.info = S.info;
  if (S.sig_in_smem) {
    m.desc = p;
    p += S.n_faults;
    m.list = reinterpret_cast<const uint16_t *>(p);
  } else {
    m.desc = S.sigdesc;
    m.list = S.siglist;
  }
  return m;
}

// shared-memory accesses by 32-bit shared-window address: the hot loops do their own address
// arithmetic (one IMAD per access) instead of 64-bit generic pointers
__device__ __forceinline__ void red_xor_s(uint32_t addr, uint32_t v) {
  asm volatile("red.shared.xor.b32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_s(uint32_t addr, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_if(bool cond, uint32_t addr, uint32_t v) {
  asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; @p red.shared.add.u32 [%1], %2; }" ::"r"((uint32_t)cond), "r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_s(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_s16(uint32_t addr) {
  uint32_t v;
  asm volatile("{ .reg .u16 t; ld.shared.u16 t, [%1]; cvt.u32.u16 %0, t; }" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint2 ld_s64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_s(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// Pauli digit: entries of the class's cdf (padded to 2^LGP with 0xFFFFFFFF) that are <= r
template <int LGP>
__device__ __forceinline__ uint32_t lin_digit(uint32_t cdf_s, uint32_t r) {
  uint32_t d = 0;
#pragma unroll
  for (int k = LGP - 1; k >= 0; --k)
    if (ld_s(cdf_s + ((d + (1u << k) - 1u) << 2)) <= r) d += 1u << k;
  return d;
}

// Read-out steps of one thread of the CTA variant (see k_sample_lin): program entries four steps
// ahead in registers, branch-free (pad entries name the always-zero row nc and only switch the
// stores off).  FULL: a warp spans 32 columns of one row (REDUX row sums).
#define QEC_CTA_REGCNT 4  // register tally counters of the CTA read-out: programs of up to 32 * 4 steps
template <bool TALLY, bool BITS, bool FULL, bool REGCNT>
__device__ __forceinline__ uint32_t lin_readout(const uint32_t *prog, const uint32_t G, const int n_steps,
                                                const uint32_t col_s, const int row_shift, uint32_t *const ocol,
                                                const long long n_words, const bool inr, const uint32_t tal_s,
                                                const bool head, const uint32_t grp, const int lane,
                                                uint32_t (&rc)[QEC_CTA_REGCNT]) {
  // (the program is padded to a multiple of four steps and followed by four pad steps: the
  // prefetch needs no bound check; n_steps counts the padded steps)
  uint32_t anyv = 0, pv = 0;
  uint32_t en[4];
  uint32_t off = 0;
#pragma unroll
  for (int u = 0; u < 4; ++u) { en[u] = __ldg(prog + off); off += G; }
  // REGCNT: the row sums (REDUX: every lane has them) stay in registers -- lane l keeps the count
  // of step 32 k + l in rc[k], the same row in every batch -- and reach the tally row once, at the
  // end of the kernel; the outer loop is unrolled so that k is static
#pragma unroll
  for (int k = 0; k < (REGCNT ? QEC_CTA_REGCNT : 1); ++k) {
    const int st_end = REGCNT ? min(n_steps, 32 * (k + 1)) : n_steps;
    for (int st = REGCNT ? 32 * k : 0; st < st_end; st += 4) {
      uint32_t e[4], v[4], a[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        e[u] = en[u];
        a[u] = col_s + ((e[u] & 0xFFFFu) << row_shift);
        v[u] = ld_s(a[u]);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) { en[u] = __ldg(prog + off); off += G; }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        st_s(a[u], 0u);
        v[u] ^= pv & (uint32_t)((int)e[u] >> 31);
        pv = v[u];
        const bool valid = (e[u] & 0x40000000u) != 0u;
        if (BITS && inr && valid) ocol[(long long)(e[u] & 0xFFFFu) * n_words] = v[u];
        if (TALLY) {
          anyv |= v[u];
          int cnt = __popc(v[u]);
          if (FULL) cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
          else for (uint32_t o = grp >> 1; o; o >>= 1) cnt += __shfl_xor_sync(0xFFFFFFFFu, cnt, o);
          if (REGCNT) rc[k] += (lane == st + u - 32 * k) ? (uint32_t)cnt : 0u;  // (pad rows count nothing: they are zero)
          else red_add_if(head && valid, tal_s + ((e[u] & 0xFFFFu) << 2), (uint32_t)cnt);
        }
      }
    }
  }
  return anyv;
}

// Constants of the draw loop that both kernels keep in registers.
struct LinDraw {
  uint32_t t1_0, t1_1, t1_2, kb, cdf_s, desc_s, list_s, rec_s, bw4, P;
  int lgb;
};

// One candidate fault at stream position p: range checks, class thinning (word rt), Pauli digit
// (word rd), then the fault's clbit list is XORed into the accumulator at shared address acc_s
// (row pitch dc.bw4 bytes) with shared-memory reductions.
template <bool TALLY>
__device__ __forceinline__ void lin_apply(const QecLinDev &S, const LinSmem &m, const LinDraw &dc,
                                          const int p, const bool has, const uint32_t rd,
                                          const uint32_t rt, const int end, const uint32_t slo,
                                          const uint32_t shi, const long long base_shot,
                                          const long long w_lo, const uint32_t acc_s,
                                          const uint32_t cnt_s, const QecSampleExport &X) {
  const int lgb = dc.lgb;
  const uint32_t BATCH = 1u << lgb, P = dc.P, bw4 = dc.bw4;
  const uint32_t cdf_s = dc.cdf_s, desc_s = dc.desc_s, list_s = dc.list_s;
  const uint32_t sl = (uint32_t)p >> lgb, shot = (uint32_t)p & (BATCH - 1u);
  bool ok = has && p < end && shot >= slo && shot < shi;  // inside the requested range
  uint32_t inf = 0, cls = 0;
  if (ok) {
    inf = m.info[sl];
    cls = (inf >> 21) & 1023u;
    ok = (inf >> 31) || rt < m.cacc[cls];  // class thinning
  }
  if (ok) {
    const uint32_t cd = cdf_s + (cls << (S.lgP + 2u));
    uint32_t d;
    if (S.uniform_digits) d = __umulhi(rd, m.cnp1[cls] + 1u);  // == the cdf count (host: ceil thresholds)
    else if (S.lgP == 4u) d = lin_digit<4>(cd, rd);
    else if (S.lgP == 2u) d = lin_digit<2>(cd, rd);
    else {
      d = 0;
      for (uint32_t step = P >> 1; step; step >>= 1)
        if (ld_s(cd + ((d + step - 1u) << 2)) <= rd) d += step;
    }
    d = min(d, m.cnp1[cls]);
    const uint32_t f = (inf & 0x1FFFFFu) + d;
    const uint32_t colb = acc_s + ((shot >> 5) << 2);
    const uint32_t msk = 1u << (shot & 31u);
    if (S.sig_in_smem) {
      const uint32_t desc = ld_s(desc_s + (f << 2));
      uint32_t li = list_s + ((desc & 0x3FFFFFu) << 1);
      for (uint32_t n = desc >> 22; n; --n, li += 2u) red_xor_s(colb + ld_s16(li) * bw4, msk);
    } else {
      if (S.rec8) {
        // 8-byte record: validity mask of the entries, then up to 7 clbits, one byte each
        // (circuits with <= 255 clbits)
        const uint2 q = S.rec_in_smem ? ld_s64(dc.rec_s + (f << 3)) : __ldg(reinterpret_cast<const uint2 *>(S.sigrec) + f);
#define QEC_RED8(K, E) if (q.x & (1u << (K))) red_xor_s(colb + (E) * bw4, msk)
        QEC_RED8(0, (q.x >> 8) & 0xFFu); QEC_RED8(1, (q.x >> 16) & 0xFFu); QEC_RED8(2, q.x >> 24);
        QEC_RED8(3, q.y & 0xFFu); QEC_RED8(4, (q.y >> 8) & 0xFFu);
        QEC_RED8(5, (q.y >> 16) & 0xFFu); QEC_RED8(6, q.y >> 24);
#undef QEC_RED8
      } else {
        // 16 / 32-byte record: length, then the first 7 / 15 clbits (one or two 128-bit loads
        // per fault instead of a descriptor and one 16-bit load per clbit)
#define QEC_RED(K, E) if (len > (K)) red_xor_s(colb + (E) * bw4, msk)
        const uint4 *rec = S.sigrec + (f << S.rec2);
        const uint4 q0 = __ldg(rec);
        const uint32_t len = q0.x & 0xFFFFu;
        QEC_RED(0u, q0.x >> 16);
        QEC_RED(1u, q0.y & 0xFFFFu); QEC_RED(2u, q0.y >> 16);
        QEC_RED(3u, q0.z & 0xFFFFu); QEC_RED(4u, q0.z >> 16);
        QEC_RED(5u, q0.w & 0xFFFFu); QEC_RED(6u, q0.w >> 16);
        if (len > 7u) {  // (only with 32-byte records)
          const uint4 q1 = __ldg(rec + 1);
          QEC_RED(7u, q1.x & 0xFFFFu); QEC_RED(8u, q1.x >> 16);
          QEC_RED(9u, q1.y & 0xFFFFu); QEC_RED(10u, q1.y >> 16);
          QEC_RED(11u, q1.z & 0xFFFFu); QEC_RED(12u, q1.z >> 16);
          QEC_RED(13u, q1.w & 0xFFFFu); QEC_RED(14u, q1.w >> 16);
          if (len > 15u) {
            const uint16_t *li = S.siglist + (__ldg(S.sigdesc + f) & 0x3FFFFFu) + 15u;
            for (uint32_t n = len - 15u; n; --n, ++li) red_xor_s(colb + (uint32_t)__ldg(li) * bw4, msk);
          }
        }
#undef QEC_RED
      }
    }
    if (TALLY) red_add_s(cnt_s + ((shot >> 1) << 2), 1u << ((shot & 1u) << 4));
    if (X.cnt) {
      const long long wr = ((base_shot + shot) >> 5) - w_lo;
      const uint32_t k = atomicAdd(X.cnt + wr, 1u);
      if (k < (uint32_t)X.cap)
        X.ev[wr * X.cap + k] = (S.slloc[sl] << 14) | ((shot & 31u) << 9) | d;
    }
  }
}

// The two candidates of a lane in one go, for the common configuration of the CTA variant (draw
// scheme 2, no class thinning, equal Pauli weights, 16-byte records in device memory): both
// records are requested before either is applied, so a step exposes one L2 latency, not two
// (ncu on d=11: the wait for the record was the largest single stall of the draw phase).
template <bool TALLY>
__device__ __forceinline__ void lin_apply_pair16(const QecLinDev &S, const LinSmem &m, const LinDraw &dc,
                                                 const int p0, const int p1, const bool has0, const bool has1,
                                                 const uint32_t rd0, const uint32_t rd1, const int end,
                                                 const uint32_t slo, const uint32_t shi, const long long base_shot,
                                                 const long long w_lo, const uint32_t acc_s, const uint32_t cnt_s,
                                                 const QecSampleExport &X) {
  const int lgb = dc.lgb;
  const uint32_t BATCH = 1u << lgb, bw4 = dc.bw4;
  const int pp[2] = {p0, p1};
  const bool hh[2] = {has0, has1};
  const uint32_t rr[2] = {rd0, rd1};
  uint32_t sl[2], shot[2], d[2];
  bool ok[2];
  uint4 q[2];
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    sl[j] = (uint32_t)pp[j] >> lgb;
    shot[j] = (uint32_t)pp[j] & (BATCH - 1u);
    ok[j] = hh[j] && pp[j] < end && shot[j] >= slo && shot[j] < shi;
    const uint32_t inf = ok[j] ? m.info[sl[j]] : 0u;
    const uint32_t np1 = m.cnp1[(inf >> 21) & 1023u];
    d[j] = min(__umulhi(rr[j], np1 + 1u), np1);  // == the cdf count (host: ceil thresholds)
    q[j] = make_uint4(0u, 0u, 0u, 0u);           // (length 0: nothing to apply)
    if (ok[j]) q[j] = __ldg(S.sigrec + ((inf & 0x1FFFFFu) + d[j]));
  }
#pragma unroll
  for (int j = 0; j < 2; ++j) {
    const uint32_t colb = acc_s + ((shot[j] >> 5) << 2);
    const uint32_t msk = 1u << (shot[j] & 31u);
    const uint32_t len = q[j].x & 0xFFFFu;
#define QEC_RED(K, E) if (len > (K)) red_xor_s(colb + (E) * bw4, msk)
    QEC_RED(0u, q[j].x >> 16);
    QEC_RED(1u, q[j].y & 0xFFFFu); QEC_RED(2u, q[j].y >> 16);
    QEC_RED(3u, q[j].z & 0xFFFFu); QEC_RED(4u, q[j].z >> 16);
    QEC_RED(5u, q[j].w & 0xFFFFu); QEC_RED(6u, q[j].w >> 16);
#undef QEC_RED
    if (ok[j]) {
      if (TALLY) red_add_s(cnt_s + ((shot[j] >> 1) << 2), 1u << ((shot[j] & 1u) << 4));
      if (X.cnt) {
        const long long wr = ((base_shot + shot[j]) >> 5) - w_lo;
        const uint32_t k = atomicAdd(X.cnt + wr, 1u);
        if (k < (uint32_t)X.cap)
          X.ev[wr * X.cap + k] = (S.slloc[sl[j]] << 14) | ((shot[j] & 31u) << 9) | d[j];
      }
    }
  }
}

// One fault stream of one batch, walked by one warp: 32 Philox calls per step (lane = call),
// each giving one candidate fault (draw schemes 0 and 1) or two (scheme 2); a warp prefix sum
// of the gaps places them, then every lane resolves its candidates (lin_apply).
template <bool TALLY>
__device__ __forceinline__ void lin_stream(const QecLinDev &S, const LinSmem &m, const LinDraw &dc,
                                           const int s, const int lane, const uint32_t blo,
                                           const uint32_t bhi, const uint32_t slo, const uint32_t shi,
                                           const long long base_shot, const long long w_lo,
                                           const uint32_t acc_s, const uint32_t cnt_s,
                                           const QecSampleExport &X) {
  const uint32_t kb = dc.kb;
  const uint32_t t1_0 = dc.t1_0, t1_1 = dc.t1_1, t1_2 = dc.t1_2;
  const int start = (int)((uint32_t)s * S.slen);
  const int end = (int)min((uint32_t)start + S.slen, S.n_slots);
  const int draw = S.draw;
  // (uniform_digits implies no thinning only together with all-always classes: draw == 2)
  const bool pair16 = draw == 2 && S.uniform_digits && !S.sig_in_smem && !S.rec8 && !S.rec2;
  int pos = start - 1;
  for (uint32_t it = 0; pos < end; ++it) {
    const QecPhilox r = qec_philox_rk(blo, bhi, it * 32u + (uint32_t)lane, 0x4C000000u | (uint32_t)s,
                                      S.rk);
    uint32_t len0, len1 = 0u, rd0, rd1 = 0u;
    bool has0, has1 = false;
    if (draw == 0) {
      // first-level gap a = entries of T1 above r0 (a == K1: the table ends, no fault in the
      // next K1 * B slots -- exact by memorylessness)
      uint32_t a = (t1_0 > r.c[0]) + (t1_1 > r.c[0]) + (t1_2 > r.c[0]);
      if (a == 3u) while (a < S.K1 && m.T1[a] > r.c[0]) ++a;
      // second-level gap: guide bucket, two unconditional steps, then the (rare) rest of the
      // scan; T2 carries a 0 sentinel at index B, so the steps need no bound check
      uint32_t g = m.G[r.c[1] >> S.gshift];
      g += m.T2[1u + g] > r.c[1];
      g += m.T2[1u + g] > r.c[1];
      while (m.T2[1u + g] > r.c[1]) ++g;
      has0 = a < S.K1;
      len0 = has0 ? a * S.B + g + 1u : kb;
      rd0 = r.c[3];
    } else {
      const uint32_t g0 = qec_gap_len(r.c[0], S.gapC);
      has0 = g0 < QEC_GAP_CAP;
      len0 = g0 + (has0 ? 1u : 0u);
      rd0 = r.c[1];
      if (draw == 2) {
        const uint32_t g1 = qec_gap_len(r.c[2], S.gapC);
        has1 = g1 < QEC_GAP_CAP;
        len1 = g1 + (has1 ? 1u : 0u);
        rd1 = r.c[3];
      }
    }
    uint32_t incl = len0 + len1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += t;
    }
    const int p1 = pos + (int)incl, p0 = p1 - (int)len1;
    pos = __shfl_sync(0xFFFFFFFFu, p1, 31);
    // (nested ifs, no `continue`: the warp reconverges before the next step's shuffles)
    if (pair16) {
      lin_apply_pair16<TALLY>(S, m, dc, p0, p1, has0, has1, rd0, rd1, end, slo, shi, base_shot, w_lo, acc_s, cnt_s, X);
    } else {
      lin_apply<TALLY>(S, m, dc, p0, has0, rd0, r.c[2], end, slo, shi, base_shot, w_lo, acc_s, cnt_s, X);
      if (draw == 2)
        lin_apply<TALLY>(S, m, dc, p1, has1, rd1, 0u, end, slo, shi, base_shot, w_lo, acc_s, cnt_s, X);
    }
  }
}

template <bool TALLY, bool BITS>
__global__ void __launch_bounds__(1024, 1)
k_sample_lin(const QecOutDev O, const QecLinDev S, const long long shot_lo, const long long n_shots,
             const QecSampleExport X) {
  extern __shared__ __align__(16) uint32_t qec_lin_smem[];
  const int T = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
  const LinSmem m = lin_carve(qec_lin_smem, S, T);
  const int lgb = S.lgb, lgw = lgb - 5;
  const uint32_t BATCH = 1u << lgb, BW = 1u << lgw;
  const uint32_t P = 1u << S.lgP;

  // ---- tables into shared memory, accumulators cleared
  {
    uint32_t *w = const_cast<uint32_t *>(m.info);
    if (S.info_in_smem) for (uint32_t i = tid; i < (uint32_t)S.L; i += T) w[i] = S.info[i];
    if (S.sig_in_smem) {
      uint32_t *d = const_cast<uint32_t *>(m.desc);
      for (uint32_t i = tid; i < S.n_faults; i += T) d[i] = S.sigdesc[i];
      uint32_t *l = reinterpret_cast<uint32_t *>(const_cast<uint16_t *>(m.list));
      const uint32_t *src = reinterpret_cast<const uint32_t *>(S.siglist);  // padded to a word
      for (uint32_t i = tid; i < (S.n_sig + 1u) / 2u; i += T) l[i] = src[i];
    }
  }
  for (uint32_t i = tid; i < S.K1; i += T) m.T1[i] = S.T1[i];
  for (uint32_t i = tid; i <= S.B; i += T) m.T2[i] = i < S.B ? S.T2[i] : 0u;
  for (uint32_t k = tid; k < 2u * S.B; k += T) {
    // entries of T2[1..B) greater than the largest r of guide bucket k: where the scan for
    // any r of the bucket may start
    const uint32_t rmax = (k << S.gshift) | ((1u << S.gshift) - 1u);
    m.G[k] = (uint16_t)qec_count_greater(S.T2 + 1, S.B - 1u, rmax);
  }
  for (uint32_t c = tid; c < S.n_classes; c += T) {
    m.cacc[c] = S.cls_acc[c];
    m.cnp1[c] = S.cls_off[c + 1] - S.cls_off[c] - 1u;
  }
  for (uint32_t i = tid; i < (S.n_classes << S.lgP); i += T) {
    const uint32_t c = i >> S.lgP, d = i & (P - 1u), o = S.cls_off[c], n = S.cls_off[c + 1] - o;
    m.cdf[i] = d < n ? S.cdf[o + d] : 0xFFFFFFFFu;
  }
  {
    const size_t nz = ((size_t)S.nc + 1u) * BW + (BATCH >> 1) + BW + (size_t)S.nc + 64;
    for (size_t i = tid; i < nz; i += T) m.acc[i] = 0u;  // acc, cnt, anyw, tally, hist are contiguous
  }
  __syncthreads();

  const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(m.acc);
  const uint32_t cnt_s = (uint32_t)__cvta_generic_to_shared(m.cnt);
  LinDraw dc;
  dc.t1_0 = m.T1[0]; dc.t1_1 = S.K1 > 1 ? m.T1[1] : 0u; dc.t1_2 = S.K1 > 2 ? m.T1[2] : 0u;
  dc.kb = S.K1 * S.B;
  dc.cdf_s = (uint32_t)__cvta_generic_to_shared(m.cdf);
  dc.desc_s = S.sig_in_smem ? (uint32_t)__cvta_generic_to_shared(m.desc) : 0u;
  dc.list_s = S.sig_in_smem ? (uint32_t)__cvta_generic_to_shared(m.list) : 0u;
  dc.rec_s = 0u;
  dc.bw4 = 4u << lgw;  // bytes per accumulator row
  dc.P = P;
  dc.lgb = lgb;
  const uint32_t tal_s = (uint32_t)__cvta_generic_to_shared(m.tally);
  // register tally counters of the read-out (lin_readout): a warp spans 32 columns of one row and the
  // program fits 32 * QEC_CTA_REGCNT steps
  const bool regcnt = TALLY && BW >= 32u && S.n_lvl <= 32 * QEC_CTA_REGCNT;
  uint32_t rc[QEC_CTA_REGCNT];
#pragma unroll
  for (int k = 0; k < QEC_CTA_REGCNT; ++k) rc[k] = 0u;
  unsigned long long nz_shots = 0;
  const long long b0 = shot_lo >> lgb, b1 = (shot_lo + n_shots - 1) >> lgb;
  const long long w_lo = shot_lo >> 5;

  for (long long b = b0 + blockIdx.x; b <= b1; b += gridDim.x) {
    const long long base_shot = b << lgb;
    // shots of this batch inside the requested range: [slo, shi)
    const uint32_t slo = shot_lo > base_shot ? (uint32_t)(shot_lo - base_shot) : 0u;
    const uint32_t shi = (uint32_t)min((long long)BATCH, shot_lo + n_shots - base_shot);
    const uint32_t blo = (uint32_t)b, bhi = (uint32_t)((unsigned long long)b >> 32);

    // ---- draw the batch's faults, XOR their signatures into acc
    for (int s = warp; s < S.n_streams; s += NW)
      lin_stream<TALLY>(S, m, dc, s, lane, blo, bhi, slo, shi, base_shot, w_lo, acc_s, cnt_s, X);
    __syncthreads();

    // ---- read-out: thread = (word column wb, row group g = tid / BW).  The host deals whole chains
    // of the time-difference basis to the T / BW groups (cta_readout_prog, qec_capi.cu), so a thread
    // meets the rows of a chain in consecutive steps and undoes the basis with a running XOR in a
    // register: one LDS, one zeroing STS, one STG per word, no second pass over the accumulator and
    // no barrier per chain depth.  Entry: row | 1 << 30 valid | 1 << 31 continues the chain of the
    // group's previous step.  Per-row ones: warp sum (REDUX when a warp spans one row) and one
    // red.shared.add into the CTA's tally row.
    {
      const uint32_t wb = (uint32_t)tid & (BW - 1u), G = (uint32_t)T >> lgw;
      const long long w = ((b << lgw) + wb) - w_lo;
      const bool inr = w >= 0 && w < O.n_words;
      const uint32_t grp = BW < 32u ? BW : 32u;  // lanes of a warp that share one clbit
      const bool head = (lane & (grp - 1u)) == 0u;
      const uint32_t *pp = S.lvl_prog + ((uint32_t)tid >> lgw);
      const uint32_t col_s = acc_s + (wb << 2);
      uint32_t *const ocol = O.bits + (inr ? w : 0);
      uint32_t anyv = 0;
      if (regcnt) anyv = lin_readout<TALLY, BITS, true, true>(pp, G, S.n_lvl, col_s, lgw + 2, ocol, O.n_words, inr, tal_s, head, grp, lane, rc);
      else if (grp == 32u) anyv = lin_readout<TALLY, BITS, true, false>(pp, G, S.n_lvl, col_s, lgw + 2, ocol, O.n_words, inr, tal_s, head, grp, lane, rc);
      else anyv = lin_readout<TALLY, BITS, false, false>(pp, G, S.n_lvl, col_s, lgw + 2, ocol, O.n_words, inr, tal_s, head, grp, lane, rc);
      if (TALLY) {
        if (anyv) atomicOr(m.anyw + wb, anyv);
        // faults-per-shot histogram from the 16-bit counters (two shots per word)
        uint32_t h0 = 0, h1 = 0, h2 = 0;
        for (uint32_t i = tid; i < (BATCH >> 1); i += T) {
          const uint32_t cw = m.cnt[i];
          m.cnt[i] = 0u;
#pragma unroll
          for (uint32_t h = 0; h < 2u; ++h) {
            const uint32_t shot = 2u * i + h, k = (cw >> (16u * h)) & 0xFFFFu;
            if (shot < slo || shot >= shi) continue;
            if (k == 0u) ++h0;
            else if (k == 1u) ++h1;
            else if (k == 2u) ++h2;
            else atomicAdd(m.hist + min(k, 63u), 1u);
          }
        }
        h0 = __reduce_add_sync(0xFFFFFFFFu, h0);
        h1 = __reduce_add_sync(0xFFFFFFFFu, h1);
        h2 = __reduce_add_sync(0xFFFFFFFFu, h2);
        if (lane == 0) {
          atomicAdd(m.hist + 0, h0); atomicAdd(m.hist + 1, h1); atomicAdd(m.hist + 2, h2);
        }
      }
    }
    __syncthreads();
    if (TALLY && (uint32_t)tid < BW) {
      nz_shots += __popc(m.anyw[tid]);
      m.anyw[tid] = 0u;
    }
  }

  // ---- tallies: per-warp rows -> global (layout: QEC_TALLY_LEN in qec_b200.h)
  if (TALLY) {
    if (regcnt) {
      // lane l of a warp kept the count of step 32 k + l of the warp's row group
      const uint32_t G = (uint32_t)T >> lgw;
#pragma unroll
      for (int k = 0; k < QEC_CTA_REGCNT; ++k) {
        const int st = 32 * k + lane;
        if (st < S.n_lvl) {
          const uint32_t e = __ldg(S.lvl_prog + (size_t)st * G + ((uint32_t)tid >> lgw));
          if ((e & 0x40000000u) && rc[k]) red_add_s(tal_s + ((e & 0xFFFFu) << 2), rc[k]);
        }
      }
    }
    __syncthreads();
    for (int off = 16; off > 0; off >>= 1) nz_shots += __shfl_down_sync(0xFFFFFFFFu, nz_shots, off);
    if (lane == 0 && nz_shots) atomicAdd(&O.tally[S.nc], nz_shots);
    for (int c = tid; c < S.nc; c += T)
      if (m.tally[c]) atomicAdd(&O.tally[c], (unsigned long long)m.tally[c]);
    for (int c = tid; c < 64; c += T)
      if (m.hist[c]) atomicAdd(&O.tally[S.nc + 2 + c], (unsigned long long)m.hist[c]);
    if (blockIdx.x == 0 && tid == 0) atomicAdd(&O.tally[S.nc + 1], (unsigned long long)n_shots);
  }
}

// ------------------------------------------------------------------------------------------
// Warp-autonomous variant for circuits with few clbits: batch = 256 shots, one stream, owned by
// ONE warp with a private accumulator acc[nc][8 words] -- no CTA barrier between drawing and
// read-out, no accumulator traffic between warps.  Read-out is row-parallel: lane = clbit, a
// lane moves its row's 32 bytes with two 128-bit loads and one 256-bit store, sums its popcounts
// in registers and updates its own tally entry without a warp reduction.
// Shared memory: [warp regions: acc nc*8 | cnt 128 | tally nc (padded to 4)] then the tables.
#define QEC_LINW_LGB 8
#define QEC_LINW_BW 8u
#define QEC_LINW_REGSTEPS 8  // read-out programs up to this many steps live in registers

__device__ __forceinline__ uint4 ld_s128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void st_s128z(uint32_t addr) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %1, %1, %1};" ::"r"(addr), "r"(0u) : "memory");
}
// predicated shared-memory reductions (no branch: the two candidates of a lane interleave)
__device__ __forceinline__ void red_xor_if(uint32_t cond, uint32_t addr, uint32_t v) {
  asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; @p red.shared.xor.b32 [%1], %2; }" ::"r"(cond), "r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void red_add_if(uint32_t cond, uint32_t addr, uint32_t v) {
  asm volatile("{ .reg .pred p; setp.ne.u32 p, %0, 0; @p red.shared.add.u32 [%1], %2; }" ::"r"(cond), "r"(addr), "r"(v) : "memory");
}

// Lean path of the warp variant (draw scheme 2, equal Pauli weights, 8-byte records and
// {first fault, number of Paulis} per location in shared memory): one candidate, branch-free.
// A lane without a fault reads location 0's tables and has every reduction predicated off.
template <bool TALLY>
__device__ __forceinline__ void linw_apply_fast(const QecLinDev &S, const int p, const bool has,
                                                const uint32_t rd, const int end, const uint32_t slo,
                                                const uint32_t nsh, const uint32_t info2_s,
                                                const uint32_t rec_s, const uint32_t acc_s,
                                                const uint32_t cnt_s, const long long base_shot,
                                                const long long w_lo, const QecSampleExport &X) {
  const uint32_t shot = (uint32_t)p & 255u;
  const bool ok = has && p < end && (shot - slo) < nsh;
  const uint32_t sl = ok ? (uint32_t)p >> QEC_LINW_LGB : 0u;
  const uint2 inf = ld_s64(info2_s + (sl << 3));
  const uint32_t d = __umulhi(rd, inf.y);
  const uint2 q = ld_s64(rec_s + ((inf.x + d) << 3));
  const uint32_t vm = ok ? q.x : 0u;  // bits 0..6: entry k of the record is present
  const uint32_t colb = acc_s + ((shot >> 3) & 0x1Cu);
  const uint32_t msk = 1u << (shot & 31u);
  red_xor_if(vm & 1u, colb + (__byte_perm(q.x, 0u, 0x4441u) << 5), msk);
  red_xor_if(vm & 2u, colb + (__byte_perm(q.x, 0u, 0x4442u) << 5), msk);
  red_xor_if(vm & 4u, colb + (__byte_perm(q.x, 0u, 0x4443u) << 5), msk);
  red_xor_if(vm & 8u, colb + (__byte_perm(q.y, 0u, 0x4440u) << 5), msk);
  red_xor_if(vm & 16u, colb + (__byte_perm(q.y, 0u, 0x4441u) << 5), msk);
  red_xor_if(vm & 32u, colb + (__byte_perm(q.y, 0u, 0x4442u) << 5), msk);
  red_xor_if(vm & 64u, colb + (__byte_perm(q.y, 0u, 0x4443u) << 5), msk);
  if (TALLY) red_add_if(ok ? 1u : 0u, cnt_s + ((shot << 1) & 0x1FCu), 1u << ((shot & 1u) << 4));
  if (X.cnt) {
    if (ok) {
      const long long wr = ((base_shot + shot) >> 5) - w_lo;
      const uint32_t k = atomicAdd(X.cnt + wr, 1u);
      if (k < (uint32_t)X.cap) X.ev[wr * X.cap + k] = (S.slloc[sl] << 14) | ((shot & 31u) << 9) | d;
    }
  }
}

template <bool TALLY>
__device__ __forceinline__ void linw_stream_fast(const QecLinDev &S, const int s, const int lane,
                                                 const uint32_t blo, const uint32_t bhi,
                                                 const uint32_t slo, const uint32_t shi,
                                                 const uint32_t info2_s, const uint32_t rec_s,
                                                 const uint32_t acc_s, const uint32_t cnt_s,
                                                 const long long base_shot, const long long w_lo,
                                                 const QecSampleExport &X) {
  const int start = (int)((uint32_t)s * S.slen);
  const int end = (int)min((uint32_t)start + S.slen, S.n_slots);
  const uint32_t nsh = shi - slo;
  const float C = S.gapC;
  int pos = start - 1;
  for (uint32_t it = 0; pos < end; ++it) {
    const QecPhilox r = qec_philox_rk(blo, bhi, it * 32u + (uint32_t)lane, 0x4C000000u | (uint32_t)s,
                                      S.rk);
    const uint32_t g0 = qec_gap_len(r.c[0], C), g1 = qec_gap_len(r.c[2], C);
    const bool has0 = g0 < QEC_GAP_CAP, has1 = g1 < QEC_GAP_CAP;
    const uint32_t len1 = g1 + (has1 ? 1u : 0u);
    uint32_t incl = g0 + (has0 ? 1u : 0u) + len1;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if (lane >= o) incl += t;
    }
    const int p1 = pos + (int)incl, p0 = p1 - (int)len1;
    pos = __shfl_sync(0xFFFFFFFFu, p1, 31);
    linw_apply_fast<TALLY>(S, p0, has0, r.c[1], end, slo, nsh, info2_s, rec_s, acc_s, cnt_s, base_shot, w_lo, X);
    linw_apply_fast<TALLY>(S, p1, has1, r.c[3], end, slo, nsh, info2_s, rec_s, acc_s, cnt_s, base_shot, w_lo, X);
  }
}

// One read-out step of a lane: row e of the accumulator is folded into the running XOR of the
// lane's chain (time-difference basis), cleared, stored (one 256-bit store = the row's whole
// 32-byte sector; two 128-bit stores are two partial-sector writes, and L2 then fetches the
// sector from DRAM first) and counted.
template <bool TALLY, bool BITS>
__device__ __forceinline__ void linw_row(const uint32_t e, const uint32_t acc_s, const uint32_t tal_s,
                                         uint4 &mlo, uint4 &mhi, uint4 &alo, uint4 &ahi,
                                         const QecOutDev &O, const long long w0, const bool fastst) {
  if (!(e >> 31)) return;
  const uint32_t c = e & 0xFFFFu;
  const uint32_t a = acc_s + c * 32u;
  const uint4 lo4 = ld_s128(a), hi4 = ld_s128(a + 16u);
  st_s128z(a);
  st_s128z(a + 16u);
  const uint32_t keep = (e & (1u << 30)) ? 0u : 0xFFFFFFFFu;  // a chain starts: drop the running XOR
  mlo.x = (mlo.x & keep) ^ lo4.x; mlo.y = (mlo.y & keep) ^ lo4.y;
  mlo.z = (mlo.z & keep) ^ lo4.z; mlo.w = (mlo.w & keep) ^ lo4.w;
  mhi.x = (mhi.x & keep) ^ hi4.x; mhi.y = (mhi.y & keep) ^ hi4.y;
  mhi.z = (mhi.z & keep) ^ hi4.z; mhi.w = (mhi.w & keep) ^ hi4.w;
  if (BITS) {
    uint32_t *po = O.bits + ((long long)c * O.n_words + w0);
    if (fastst) {
      asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(po), "r"(mlo.x),
                   "r"(mlo.y), "r"(mlo.z), "r"(mlo.w), "r"(mhi.x), "r"(mhi.y), "r"(mhi.z), "r"(mhi.w)
                   : "memory");
    } else {
      const uint32_t v[8] = {mlo.x, mlo.y, mlo.z, mlo.w, mhi.x, mhi.y, mhi.z, mhi.w};
#pragma unroll
      for (int j = 0; j < 8; ++j)
        if (w0 + j >= 0 && w0 + j < O.n_words) po[j] = v[j];
    }
  }
  if (TALLY) {
    alo.x |= mlo.x; alo.y |= mlo.y; alo.z |= mlo.z; alo.w |= mlo.w;
    ahi.x |= mhi.x; ahi.y |= mhi.y; ahi.z |= mhi.z; ahi.w |= mhi.w;
    const uint32_t cnt = __popc(mlo.x) + __popc(mlo.y) + __popc(mlo.z) + __popc(mlo.w) +
                         __popc(mhi.x) + __popc(mhi.y) + __popc(mhi.z) + __popc(mhi.w);
    const uint32_t ta = tal_s + (c << 2);
    st_s(ta, ld_s(ta) + cnt);
  }
}

template <bool TALLY, bool BITS, bool FAST>
__global__ void __launch_bounds__(1024, 1)
k_sample_linw(const QecOutDev O, const QecLinDev S, const long long shot_lo, const long long n_shots,
              const QecSampleExport X) {
  extern __shared__ __align__(16) uint32_t qec_lin_smem[];
  const int T = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
  const uint32_t ncp = ((uint32_t)S.nc + 3u) & ~3u;
  const uint32_t wwords = (uint32_t)S.nc * QEC_LINW_BW + 128u + ncp;  // words per warp region
  const uint32_t P = 1u << S.lgP;
  LinSmem m;
  uint32_t *rec_sm = nullptr, *info2_sm = nullptr;
  m.acc = qec_lin_smem + (size_t)warp * wwords;
  m.cnt = m.acc + (size_t)S.nc * QEC_LINW_BW;
  m.tally = m.cnt + 128;
  m.anyw = nullptr;
  m.hist = qec_lin_smem + (size_t)NW * wwords;
  m.T1 = m.hist + 64;
  m.T2 = m.T1 + S.K1;
  m.G = reinterpret_cast<uint16_t *>(m.T2 + S.B + 1u);
  m.cacc = m.T2 + 2u * S.B + 1u;
  m.cnp1 = m.cacc + S.n_classes;
  m.cdf = m.cnp1 + S.n_classes;
  {
    uint32_t *p = m.cdf + ((size_t)S.n_classes << S.lgP);
    if (S.info_in_smem) {
      m.info = p;
      for (uint32_t i = tid; i < (uint32_t)S.L; i += T) p[i] = S.info[i];
    } else {
      m.info = S.info;
    }
    m.desc = S.sigdesc;
    m.list = S.siglist;
    if (S.info_in_smem) p += S.L;
    p += (2u - ((uint32_t)(p - qec_lin_smem) & 1u)) & 1u;  // 8-byte aligned
    if (S.fast) {
      info2_sm = p;
      const uint32_t *src = reinterpret_cast<const uint32_t *>(S.info2);
      for (uint32_t i = tid; i < 2u * (uint32_t)S.L; i += T) p[i] = src[i];
      p += 2u * (uint32_t)S.L;
    }
    rec_sm = p;
    if (S.rec_in_smem) {
      const uint32_t *src = reinterpret_cast<const uint32_t *>(S.sigrec);
      for (uint32_t i = tid; i < 2u * S.n_faults; i += T) p[i] = src[i];
    }
  }
  for (uint32_t i = tid; i < S.K1; i += T) m.T1[i] = S.T1[i];
  if (S.draw == 0) {
    for (uint32_t i = tid; i <= S.B; i += T) m.T2[i] = i < S.B ? S.T2[i] : 0u;
    for (uint32_t k = tid; k < 2u * S.B; k += T) {
      const uint32_t rmax = (k << S.gshift) | ((1u << S.gshift) - 1u);
      m.G[k] = (uint16_t)qec_count_greater(S.T2 + 1, S.B - 1u, rmax);
    }
  }
  for (uint32_t c = tid; c < S.n_classes; c += T) {
    m.cacc[c] = S.cls_acc[c];
    m.cnp1[c] = S.cls_off[c + 1] - S.cls_off[c] - 1u;
  }
  for (uint32_t i = tid; i < (S.n_classes << S.lgP); i += T) {
    const uint32_t c = i >> S.lgP, d = i & (P - 1u), o = S.cls_off[c], n = S.cls_off[c + 1] - o;
    m.cdf[i] = d < n ? S.cdf[o + d] : 0xFFFFFFFFu;
  }
  for (uint32_t i = tid; i < (uint32_t)NW * wwords + 64u; i += T) qec_lin_smem[i] = 0u;
  __syncthreads();

  const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(m.acc);
  const uint32_t cnt_s = (uint32_t)__cvta_generic_to_shared(m.cnt);
  const uint32_t tal_s = (uint32_t)__cvta_generic_to_shared(m.tally);
  const uint32_t rec_s = (uint32_t)__cvta_generic_to_shared(rec_sm);
  const uint32_t info2_s = (uint32_t)__cvta_generic_to_shared(info2_sm);
  LinDraw dc;
  dc.t1_0 = S.K1 > 0 ? m.T1[0] : 0u; dc.t1_1 = S.K1 > 1 ? m.T1[1] : 0u; dc.t1_2 = S.K1 > 2 ? m.T1[2] : 0u;
  dc.kb = S.K1 * S.B;
  dc.cdf_s = (uint32_t)__cvta_generic_to_shared(m.cdf);
  dc.desc_s = 0u; dc.list_s = 0u;
  dc.rec_s = rec_s;
  dc.bw4 = 4u * QEC_LINW_BW;
  dc.P = P;
  dc.lgb = QEC_LINW_LGB;
  unsigned long long nz_shots = 0;
  uint32_t h1 = 0, h2 = 0, hx = 0;  // shots with exactly 1 / exactly 2 / at least 3 faults
  const long long b0 = shot_lo >> QEC_LINW_LGB, b1 = (shot_lo + n_shots - 1) >> QEC_LINW_LGB;
  const long long w_lo = shot_lo >> 5;
  const bool vec_ok = ((O.n_words | w_lo) & 7) == 0 && ((size_t)O.bits & 31) == 0;
  // the lane's read-out program: constant over batches, kept in registers when short
  const bool regprog = S.n_steps <= QEC_LINW_REGSTEPS;
  uint32_t prog[QEC_LINW_REGSTEPS];
#pragma unroll
  for (int k = 0; k < QEC_LINW_REGSTEPS; ++k)
    prog[k] = (regprog && k < S.n_steps) ? __ldg(S.rowprog + k * 32 + lane) : 0u;

  for (long long b = b0 + (long long)blockIdx.x * NW + warp; b <= b1; b += (long long)gridDim.x * NW) {
    const long long base_shot = b << QEC_LINW_LGB;
    const uint32_t slo = shot_lo > base_shot ? (uint32_t)(shot_lo - base_shot) : 0u;
    const uint32_t shi = (uint32_t)min(256ll, shot_lo + n_shots - base_shot);
    const uint32_t blo = (uint32_t)b, bhi = (uint32_t)((unsigned long long)b >> 32);
    for (int s = 0; s < S.n_streams; ++s) {
      if (FAST) linw_stream_fast<TALLY>(S, s, lane, blo, bhi, slo, shi, info2_s, rec_s, acc_s, cnt_s, base_shot, w_lo, X);
      else lin_stream<TALLY>(S, m, dc, s, lane, blo, bhi, slo, shi, base_shot, w_lo, acc_s, cnt_s, X);
    }
    __syncwarp();

    // ---- read-out: lane = chain of clbit rows
    const long long w0 = (b << 3) - w_lo;  // first output word of the batch (may be < 0 at the range start)
    const bool fastst = w0 >= 0 && w0 + 8 <= O.n_words && vec_ok;
    uint4 alo = make_uint4(0u, 0u, 0u, 0u), ahi = alo;  // OR of the lane's rows (non-zero shots)
    uint4 mlo = alo, mhi = alo;                         // running XOR along the lane's chain
    if (regprog) {
#pragma unroll
      for (int k = 0; k < QEC_LINW_REGSTEPS; ++k)
        if (k < S.n_steps) linw_row<TALLY, BITS>(prog[k], acc_s, tal_s, mlo, mhi, alo, ahi, O, w0, fastst);
    } else {
      uint32_t e_next = __ldg(S.rowprog + lane);
      for (int step = 0; step < S.n_steps; ++step) {
        const uint32_t e = e_next;
        if (step + 1 < S.n_steps) e_next = __ldg(S.rowprog + (step + 1) * 32 + lane);
        linw_row<TALLY, BITS>(e, acc_s, tal_s, mlo, mhi, alo, ahi, O, w0, fastst);
      }
    }
    if (TALLY) {
      const uint32_t nzb =
          __popc(__reduce_or_sync(0xFFFFFFFFu, alo.x)) + __popc(__reduce_or_sync(0xFFFFFFFFu, alo.y)) +
          __popc(__reduce_or_sync(0xFFFFFFFFu, alo.z)) + __popc(__reduce_or_sync(0xFFFFFFFFu, alo.w)) +
          __popc(__reduce_or_sync(0xFFFFFFFFu, ahi.x)) + __popc(__reduce_or_sync(0xFFFFFFFFu, ahi.y)) +
          __popc(__reduce_or_sync(0xFFFFFFFFu, ahi.z)) + __popc(__reduce_or_sync(0xFFFFFFFFu, ahi.w));
      if (lane == 0) nz_shots += nzb;
      // faults-per-shot histogram: 128 counter words (two 16-bit counters each), four per lane.
      // Shots outside the requested range never receive a fault, and the zero-fault count is
      // derived at the end, so no range test here.  pk: byte j counts the lane's shots with
      // min(faults, 3) == j.
      const uint4 cw = ld_s128(cnt_s + ((uint32_t)lane << 4));
      st_s128z(cnt_s + ((uint32_t)lane << 4));
      uint32_t pk = 0;
#define QEC_CNT(W) pk += (1u << (min((W) & 0xFFFFu, 3u) << 3)) + (1u << (min((W) >> 16, 3u) << 3))
      QEC_CNT(cw.x); QEC_CNT(cw.y); QEC_CNT(cw.z); QEC_CNT(cw.w);
#undef QEC_CNT
      h1 += (pk >> 8) & 0xFFu;
      h2 += (pk >> 16) & 0xFFu;
      if (pk >> 24) {  // rare: some shot of this lane carries 3 or more faults
        hx += pk >> 24;
        const uint32_t v[4] = {cw.x, cw.y, cw.z, cw.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t k0 = v[j] & 0xFFFFu, k1 = v[j] >> 16;
          if (k0 >= 3u) atomicAdd(m.hist + min(k0, 63u), 1u);
          if (k1 >= 3u) atomicAdd(m.hist + min(k1, 63u), 1u);
        }
      }
    }
    __syncwarp();
  }

  if (TALLY) {
    // faults-per-shot histogram: entries 1 and 2 from the registers, >= 3 from shared memory;
    // entry 0 = shots of the range minus the shots with a fault (64-bit wrap-around adds)
    unsigned long long t1 = h1, t2 = h2, tx = hx;
    for (int off = 16; off > 0; off >>= 1) {
      t1 += __shfl_down_sync(0xFFFFFFFFu, t1, off);
      t2 += __shfl_down_sync(0xFFFFFFFFu, t2, off);
      tx += __shfl_down_sync(0xFFFFFFFFu, tx, off);
    }
    if (lane == 0) {
      if (t1) atomicAdd(&O.tally[S.nc + 2 + 1], t1);
      if (t2) atomicAdd(&O.tally[S.nc + 2 + 2], t2);
      if (t1 + t2 + tx) atomicAdd(&O.tally[S.nc + 2], 0ull - (t1 + t2 + tx));
      if (nz_shots) atomicAdd(&O.tally[S.nc], nz_shots);
    }
    __syncthreads();
    for (int c = tid; c < S.nc; c += T) {
      unsigned long long v = 0;
      for (int k = 0; k < NW; ++k) v += qec_lin_smem[(size_t)k * wwords + (size_t)S.nc * QEC_LINW_BW + 128u + c];
      if (v) atomicAdd(&O.tally[c], v);
    }
    for (int c = tid; c < 64; c += T)
      if (m.hist[c]) atomicAdd(&O.tally[S.nc + 2 + c], (unsigned long long)m.hist[c]);
    if (blockIdx.x == 0 && tid == 0) {
      atomicAdd(&O.tally[S.nc + 1], (unsigned long long)n_shots);
      atomicAdd(&O.tally[S.nc + 2], (unsigned long long)n_shots);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Lane-stream variant: no shared-memory atomics at all.  ncu on k_sample_linw showed the SM's
// shared-memory atomic unit (about 2 cycles per lane) as the bound: ~700 lane-atomics per 256
// shots, whatever the instruction count or the occupancy.  Here a warp owns a batch of 1024
// shots and LANE l OWNS WORD COLUMN l (shots 32 l .. 32 l + 31 of the batch): every lane walks
// its own fault stream over slot = sorted location * 32 + shot bit (computed gaps, two
// candidates per Philox call, counter = (batch, call * 32 + lane, 0x4D000000)), and all flips of
// column l are plain load / xor / store on acc[row][l] -- bank = lane, so neither conflicts nor
// races.  The price is the imbalance between lanes (a step lasts as long as its busiest lane:
// ~70 % useful at d=5, p=1e-3).  Faults per shot are counted in six bit-sliced register planes.
// Read-out moves four rows per pass (8 lanes x 128 bits per row: conflict-free shared loads, one
// full 128-byte line per row stored); the time-difference basis is undone in place (row ^=
// previous row of the chain; the host schedules the passes so that a row and its previous row
// never share one).  Eligible: draw scheme 2, equal Pauli weights, 8-byte
// records.
// Shared memory: [warp regions: acc (nc+2)*32 (row nc = dummy target of absent record entries,
// row nc+1 = always zero, the "previous row" of rows without one) | tally ncp] then hist 64,
// prog 4 * n_steps (read-out passes), info L (first fault | number of Paulis << 21), rec 2F.
#define QEC_LANES_LGB 10
#define QEC_LANES_UNROLL 4  // read-out passes per trip (the host pads the schedule: qec_capi.cu)

// record table -> shared memory, 128 bits per thread and trip (device allocations are 256-byte
// aligned; the shared copy is 8-byte aligned: two 64-bit stores)
__device__ __forceinline__ void lin_copy_records(uint32_t *dst, const uint2 *src_rec, uint32_t n_faults, int tid, int T) {
  const uint4 *src = reinterpret_cast<const uint4 *>(src_rec);
  const uint32_t n4 = n_faults >> 1;  // two records per 128-bit load
  uint2 *d2 = reinterpret_cast<uint2 *>(dst);
  for (uint32_t i = tid; i < n4; i += T) {
    const uint4 v = __ldg(src + i);
    d2[2 * i] = make_uint2(v.x, v.y);
    d2[2 * i + 1] = make_uint2(v.z, v.w);
  }
  if ((n_faults & 1u) && tid == 0) d2[n_faults - 1] = src_rec[n_faults - 1];
}

// One candidate of a lane.  Absent record entries name the dummy row nc and a lane without a
// fault XORs a zero mask (into whatever rows location 0's record names), so the seven load /
// xor / store triples need no predicate.
__device__ __forceinline__ void lanes_apply(const uint32_t p, const bool ok, const uint2 q, const uint32_t lane_acc) {
  const uint32_t bit = p & 31u, msk = ok ? 1u << bit : 0u;
  const uint32_t a0 = lane_acc + (__byte_perm(q.x, 0u, 0x4441u) << 7);
  const uint32_t a1 = lane_acc + (__byte_perm(q.x, 0u, 0x4442u) << 7);
  const uint32_t a2 = lane_acc + (__byte_perm(q.x, 0u, 0x4443u) << 7);
  const uint32_t a3 = lane_acc + (__byte_perm(q.y, 0u, 0x4440u) << 7);
  const uint32_t a4 = lane_acc + (__byte_perm(q.y, 0u, 0x4441u) << 7);
  const uint32_t a5 = lane_acc + (__byte_perm(q.y, 0u, 0x4442u) << 7);
  const uint32_t a6 = lane_acc + (__byte_perm(q.y, 0u, 0x4443u) << 7);
  // the rows of one fault are distinct (or the dummy): all loads first, then all stores
  const uint32_t v0 = ld_s(a0), v1 = ld_s(a1), v2 = ld_s(a2), v3 = ld_s(a3), v4 = ld_s(a4),
                 v5 = ld_s(a5), v6 = ld_s(a6);
  st_s(a0, v0 ^ msk); st_s(a1, v1 ^ msk); st_s(a2, v2 ^ msk); st_s(a3, v3 ^ msk);
  st_s(a4, v4 ^ msk); st_s(a5, v5 ^ msk); st_s(a6, v6 ^ msk);
}

// Read-out of a lane-owned accumulator acc[row][32 word columns] (batch = 32 output words
// starting at word w0).  Four rows per pass: lane group g = lane >> 3 takes one row, its eight lanes
// 128 bits each (conflict-free shared loads, one full 128-byte line per row stored).  The host deals
// whole chains of the time-difference basis (c0 -> succ(c0) -> ...) to the four groups, so a group
// meets the rows of a chain in consecutive passes and undoes the basis with a running XOR held in
// registers: one LDS.128 and one zeroing STS.128 per row, no second load and no write-back.
// Program entry (qec_capi.cu, lanes_readout_prog): row | 1 << 9 valid | 1 << 31 the row continues
// the chain of the group's previous pass; pad entries name the always-zero row nc + 1.
// Per-clbit tallies: the program is the same for every batch, so pass k of group g always meets the
// same row -- on the fast path (whole, aligned batches) the trips are unrolled and every lane keeps
// its own partial counts of "its" rows in registers (LaneCounts: two rows per register, 16 bits
// each), with no shuffle and no shared-memory traffic per batch; they are reduced over the eight
// lanes of a group and added to the warp's tally row every QEC_LANES_FLUSH batches and at the end
// (ncu: the per-pass shuffle tree + red.shared.add cost 0.14 of 0.88 ms per 1e8 shots at d=5).
// Partial batches take the rolled loop with the per-pass reduction.
// Returns nothing; anyv collects the OR of all rows (non-zero outcomes).
#define QEC_LANES_MAXTRIPS 16  // 64 passes of four rows >= 254 clbits
#define QEC_LANES_FLUSH 448u   // most batches between flushes (QecLinDev::flush_every): 448 * 128 < 2^16 per lane and row

struct LaneCounts {
  uint32_t c[2 * QEC_LANES_MAXTRIPS];
  uint32_t nb;
};

struct LaneReadCtx {
  uint32_t col_s, pe_s, c4;
  char *colp;
  unsigned long long rowbytes;
  bool row32;
  long long w0;
  long long n_words;
};

// one trip: QEC_LANES_UNROLL passes whose entries e[] and rows v[] are already loaded.  ST: how a
// lane stores its four words of a row -- 2: one 128-bit store (rows 16-byte aligned), 1: four
// 32-bit stores (the batch's words all exist, rows not aligned: n_words not a multiple of 4),
// 0: each word checked against the range
template <bool TALLY, bool BITS, int ST>
__device__ __forceinline__ void lanes_trip(const LaneReadCtx &R, const uint32_t (&e)[QEC_LANES_UNROLL],
                                           uint4 (&v)[QEC_LANES_UNROLL], uint4 &pv, uint4 &anyv,
                                           uint32_t (&cpk)[QEC_LANES_UNROLL / 2]) {
#pragma unroll
  for (int u = 0; u < QEC_LANES_UNROLL; ++u) {
    st_s128z(R.col_s + ((e[u] & 0xFFu) << 7));
    const uint32_t m = (uint32_t)((int)e[u] >> 31);  // all ones: XOR the previous row of the chain in
    v[u].x ^= pv.x & m; v[u].y ^= pv.y & m; v[u].z ^= pv.z & m; v[u].w ^= pv.w & m;
    pv = v[u];
    const bool rowok = (e[u] & 0x200u) != 0u;
    if (BITS && rowok) {
      const uint32_t row = e[u] & 0xFFu;
      // (row * bytes per row in one 32 x 32 -> 64 multiply-add; rows of 2^32 bytes or more take the slow form)
      char *po = R.row32 ? R.colp + (unsigned long long)row * (uint32_t)R.rowbytes : R.colp + (unsigned long long)row * R.rowbytes;
      if (ST == 2) *reinterpret_cast<uint4 *>(po) = v[u];
      else if (ST == 1) {
        uint32_t *pw = reinterpret_cast<uint32_t *>(po);
        pw[0] = v[u].x; pw[1] = v[u].y; pw[2] = v[u].z; pw[3] = v[u].w;
      } else {
        uint32_t *pw = reinterpret_cast<uint32_t *>(po);
        const uint32_t vv[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (R.w0 + R.c4 + j >= 0 && R.w0 + R.c4 + j < R.n_words) pw[j] = vv[j];
      }
    }
    if (TALLY) {
      anyv.x |= v[u].x; anyv.y |= v[u].y; anyv.z |= v[u].z; anyv.w |= v[u].w;
      const uint32_t cnt = __popc(v[u].x) + __popc(v[u].y) + __popc(v[u].z) + __popc(v[u].w);
      if (u & 1) cpk[u >> 1] |= cnt << 16; else cpk[u >> 1] = cnt;  // two rows' counts per register
    }
  }
}

// 8-lane sums of a packed pair of row counts -> the warp's tally row (head lane of each group)
__device__ __forceinline__ void lanes_tally_pair(uint32_t lo, uint32_t hi, const uint32_t e0, const uint32_t e1,
                                                 const uint32_t tal_s, const bool head) {
  lo += __shfl_xor_sync(0xFFFFFFFFu, lo, 1); hi += __shfl_xor_sync(0xFFFFFFFFu, hi, 1);
  lo += __shfl_xor_sync(0xFFFFFFFFu, lo, 2); hi += __shfl_xor_sync(0xFFFFFFFFu, hi, 2);
  lo += __shfl_xor_sync(0xFFFFFFFFu, lo, 4); hi += __shfl_xor_sync(0xFFFFFFFFu, hi, 4);
  // (pad entries name the zero row, whose tally entry does not exist: skip them)
  red_add_if(head && (e0 & 0x200u), tal_s + ((e0 & 0xFFu) << 2), lo);
  red_add_if(head && (e1 & 0x200u), tal_s + ((e1 & 0xFFu) << 2), hi);
}

__device__ __forceinline__ void lanes_counts_flush(LaneCounts &LC, const uint32_t tal_s, const uint32_t prog_s,
                                                   const uint32_t n_steps, const int lane) {
  const uint32_t pe_s = prog_s + (((uint32_t)lane >> 3) << 2);
  const bool head = ((uint32_t)lane & 7u) == 0u;
#pragma unroll
  for (int t = 0; t < QEC_LANES_MAXTRIPS; ++t) {
    if ((uint32_t)t * QEC_LANES_UNROLL >= n_steps) break;
#pragma unroll
    for (int h = 0; h < QEC_LANES_UNROLL / 2; ++h) {
      const uint32_t c = LC.c[t * (QEC_LANES_UNROLL / 2) + h];
      LC.c[t * (QEC_LANES_UNROLL / 2) + h] = 0u;
      const uint32_t e0 = ld_s(pe_s + ((uint32_t)(t * QEC_LANES_UNROLL + 2 * h) << 4));
      const uint32_t e1 = ld_s(pe_s + ((uint32_t)(t * QEC_LANES_UNROLL + 2 * h + 1) << 4));
      lanes_tally_pair(c & 0xFFFFu, c >> 16, e0, e1, tal_s, head);
    }
  }
  LC.nb = 0u;
}

template <bool TALLY, bool BITS, int ST>
__device__ __forceinline__ void lanes_readout_unrolled(const LaneReadCtx &R, const uint32_t n_steps,
                                                       uint32_t (&en)[QEC_LANES_UNROLL], uint4 (&vn)[QEC_LANES_UNROLL],
                                                       uint4 &pv, uint4 &anyv, LaneCounts &LC) {
#pragma unroll
  for (int t = 0; t < QEC_LANES_MAXTRIPS; ++t) {
    const uint32_t pass = (uint32_t)t * QEC_LANES_UNROLL;
    if (pass >= n_steps) break;
    uint32_t e[QEC_LANES_UNROLL], cpk[QEC_LANES_UNROLL / 2];
    uint4 v[QEC_LANES_UNROLL];
#pragma unroll
    for (int u = 0; u < QEC_LANES_UNROLL; ++u) { e[u] = en[u]; v[u] = vn[u]; }
    if (t + 1 < QEC_LANES_MAXTRIPS && pass + QEC_LANES_UNROLL < n_steps) {
#pragma unroll
      for (int u = 0; u < QEC_LANES_UNROLL; ++u) en[u] = ld_s(R.pe_s + ((pass + QEC_LANES_UNROLL + u) << 4));
#pragma unroll
      for (int u = 0; u < QEC_LANES_UNROLL; ++u) vn[u] = ld_s128(R.col_s + ((en[u] & 0xFFu) << 7));
    }
    lanes_trip<TALLY, BITS, ST>(R, e, v, pv, anyv, cpk);
    if (TALLY) { LC.c[2 * t] += cpk[0]; LC.c[2 * t + 1] += cpk[1]; }
  }
}

template <bool TALLY, bool BITS>
__device__ __forceinline__ uint32_t lanes_readout(const QecOutDev &O, const uint32_t acc_s, const uint32_t tal_s,
                                                  const uint32_t prog_s, const uint32_t n_steps,
                                                  const long long w0, const bool vec_ok, const int lane,
                                                  LaneCounts &LC, const uint32_t flush_every) {
  static_assert(QEC_LANES_UNROLL == 4, "two packed count registers per trip");
  const bool whole = w0 >= 0 && w0 + 32 <= O.n_words;  // the batch's 32 words all exist
  LaneReadCtx R;
  R.c4 = ((uint32_t)lane & 7u) << 2;
  R.col_s = acc_s + (R.c4 << 2);
  R.pe_s = prog_s + (((uint32_t)lane >> 3) << 2);
  R.rowbytes = (unsigned long long)O.n_words << 2;
  R.row32 = R.rowbytes < (1ull << 32);
  R.colp = reinterpret_cast<char *>(O.bits) + ((w0 + (long long)R.c4) << 2);
  R.w0 = w0; R.n_words = O.n_words;
  uint4 anyv = make_uint4(0u, 0u, 0u, 0u);
  uint4 pv = make_uint4(0u, 0u, 0u, 0u);
  // software-pipelined: the next trip's program entries and rows are loaded before this trip's
  // stores and counts (n_steps is a multiple of QEC_LANES_UNROLL: lanes_readout_prog)
  uint32_t en[QEC_LANES_UNROLL];
  uint4 vn[QEC_LANES_UNROLL];
#pragma unroll
  for (int u = 0; u < QEC_LANES_UNROLL; ++u) en[u] = ld_s(R.pe_s + ((uint32_t)u << 4));
#pragma unroll
  for (int u = 0; u < QEC_LANES_UNROLL; ++u) vn[u] = ld_s128(R.col_s + ((en[u] & 0xFFu) << 7));
  // (a long chain dealt to one lane group can push the program past the unrolled trips -- e.g. a
  // repetition code with a hundred rounds: such programs take the rolled loop for every batch)
  if (whole && n_steps <= (uint32_t)(QEC_LANES_MAXTRIPS * QEC_LANES_UNROLL)) {
    // the trips unrolled: per-clbit counts in registers (LaneCounts).  One straight-line copy per
    // store mode (the kernel is sensitive to instruction fetch: the two must not interleave).
    if (vec_ok) lanes_readout_unrolled<TALLY, BITS, 2>(R, n_steps, en, vn, pv, anyv, LC);
    else lanes_readout_unrolled<TALLY, BITS, 1>(R, n_steps, en, vn, pv, anyv, LC);
    if (TALLY && ++LC.nb >= flush_every) lanes_counts_flush(LC, tal_s, prog_s, n_steps, lane);
  } else {
    const bool head = R.c4 == 0u;
    for (uint32_t pass = 0; pass < n_steps; pass += QEC_LANES_UNROLL) {
      uint32_t e[QEC_LANES_UNROLL], cpk[QEC_LANES_UNROLL / 2];
      uint4 v[QEC_LANES_UNROLL];
#pragma unroll
      for (int u = 0; u < QEC_LANES_UNROLL; ++u) { e[u] = en[u]; v[u] = vn[u]; }
      if (pass + QEC_LANES_UNROLL < n_steps) {
#pragma unroll
        for (int u = 0; u < QEC_LANES_UNROLL; ++u) en[u] = ld_s(R.pe_s + ((pass + QEC_LANES_UNROLL + u) << 4));
#pragma unroll
        for (int u = 0; u < QEC_LANES_UNROLL; ++u) vn[u] = ld_s128(R.col_s + ((en[u] & 0xFFu) << 7));
      }
      lanes_trip<TALLY, BITS, 0>(R, e, v, pv, anyv, cpk);
      if (TALLY) {
        lanes_tally_pair(cpk[0] & 0xFFFFu, cpk[0] >> 16, e[0], e[1], tal_s, head);
        lanes_tally_pair(cpk[1] & 0xFFFFu, cpk[1] >> 16, e[2], e[3], tal_s, head);
      }
    }
  }
  // every lane zeroed the rows it read; the next batch's faults may land on any of them
  __syncwarp();
  uint32_t nz = 0;
  if (TALLY) {
    // non-zero outcomes: OR over the four row groups, then lanes 0..7 hold the batch's 32 words
    anyv.x |= __shfl_xor_sync(0xFFFFFFFFu, anyv.x, 8); anyv.y |= __shfl_xor_sync(0xFFFFFFFFu, anyv.y, 8);
    anyv.z |= __shfl_xor_sync(0xFFFFFFFFu, anyv.z, 8); anyv.w |= __shfl_xor_sync(0xFFFFFFFFu, anyv.w, 8);
    anyv.x |= __shfl_xor_sync(0xFFFFFFFFu, anyv.x, 16); anyv.y |= __shfl_xor_sync(0xFFFFFFFFu, anyv.y, 16);
    anyv.z |= __shfl_xor_sync(0xFFFFFFFFu, anyv.z, 16); anyv.w |= __shfl_xor_sync(0xFFFFFFFFu, anyv.w, 16);
    if (lane < 8) nz = __popc(anyv.x) + __popc(anyv.y) + __popc(anyv.z) + __popc(anyv.w);
  }
  return nz;
}

// faults-per-shot counters of a lane's 32 shots: six bit planes, ripple-carry increment of the
// shots in mask c (the carry rarely leaves plane 1)
struct LanePlanes {
  uint32_t p0, p1, p2, p3, p4, p5;
};
__device__ __forceinline__ void planes_carry(LanePlanes &P, uint32_t c) {  // c: carries into plane 2
  uint32_t t = P.p2 & c; P.p2 ^= c; c = t;
  t = P.p3 & c; P.p3 ^= c; c = t;
  t = P.p4 & c; P.p4 ^= c; c = t;
  t = P.p5 & c; P.p5 ^= c; c = t;
  // 63 + 1: stay at 63 (the last histogram bin is ">= 63")
  P.p0 |= c; P.p1 |= c; P.p2 |= c; P.p3 |= c; P.p4 |= c; P.p5 |= c;
}
// up to four increments at once: a shot takes at most four, so it carries out of plane 1 at most
// once and the carries can be collected -- one test per group instead of one per candidate
__device__ __forceinline__ void planes_add_group(LanePlanes &P, const uint32_t *m, const int n) {  // n <= 4
  uint32_t cc = 0;
#pragma unroll
  for (int j = 0; j < n; ++j) {
    uint32_t c = m[j];
    uint32_t t = P.p0 & c; P.p0 ^= c; c = t;
    t = P.p1 & c; P.p1 ^= c;
    cc |= t;
  }
  if (cc) planes_carry(P, cc);
}

// The lane's fault stream of one batch.  NP Philox calls (2 NP candidates) per trip,
// software-pipelined: the next trip's Philox and gaps do not depend on the position, so they are
// computed while this trip's candidates wait on shared memory.  The position saturates at `end`
// (a finished lane's candidates all land beyond it).  FAST: every shot of the batch is inside the
// requested range and no 32-bit draw reaches the gap clamp.
template <bool TALLY, bool EXPORT, int NP, bool FAST>
__device__ __forceinline__ void lanes_stream(const QecLinDev &S, const QecSampleExport &X, const uint32_t blo,
                                             const uint32_t bhi, const int lane, const uint32_t end, const float C,
                                             const uint32_t lane_sh, const uint32_t nsh, const uint32_t info2_s,
                                             const uint32_t rec_s, const uint32_t lane_acc, LanePlanes &P,
                                             uint32_t *xrow, uint32_t &xk) {
  uint32_t pos = 0xFFFFFFFFu;  // -1
  uint32_t xprev = 0xFFFFFFFFu;  // (16-bit export: slot of the lane's previous fault)
  QecPhilox rr[NP];
  uint32_t gg[2 * NP];
#pragma unroll
  for (int c = 0; c < NP; ++c) {
    rr[c] = qec_philox_rk(blo, bhi, (uint32_t)c * 32u + (uint32_t)lane, 0x4D000000u, S.rk);
    gg[2 * c] = qec_gap_len(rr[c].c[0], C);
    gg[2 * c + 1] = qec_gap_len(rr[c].c[2], C);
  }
  // (pos + 1 <= end: the lane is still inside its stream; positions stay below 2^27 + end)
  for (uint32_t it = 0; __any_sync(0xFFFFFFFFu, pos + 1u <= end); it += NP) {
    QecPhilox nr[NP];
#pragma unroll
    for (int c = 0; c < NP; ++c)
      nr[c] = qec_philox_rk(blo, bhi, (it + NP + (uint32_t)c) * 32u + (uint32_t)lane, 0x4D000000u, S.rk);
    uint32_t pp[2 * NP], sl[2 * NP], dd[2 * NP];
    bool ok[2 * NP];
    uint2 qq[2 * NP];
    uint32_t run = pos;
    if (FAST) {
#pragma unroll
      for (int j = 0; j < 2 * NP; ++j) {
        run += gg[j] + 1u;
        pp[j] = run;
        ok[j] = run < end;
        sl[j] = ok[j] ? run >> 5 : 0u;
      }
      pos = min(run, end);
    } else {
      const bool live = pos + 1u <= end;
#pragma unroll
      for (int j = 0; j < 2 * NP; ++j) {
        const bool has = gg[j] < QEC_GAP_CAP;
        run += gg[j] + (has ? 1u : 0u);
        pp[j] = run;
        ok[j] = live && has && run < end && (lane_sh + (run & 31u)) < nsh;
        sl[j] = ok[j] ? run >> 5 : 0u;
      }
      pos = live ? min(run, end) : end;
    }
#pragma unroll
    for (int j = 0; j < 2 * NP; ++j) {
      const uint32_t inf = ld_s(info2_s + (sl[j] << 2));
      dd[j] = __umulhi(rr[j >> 1].c[(j & 1) * 2 + 1], inf >> 21);
      qq[j] = ld_s64(rec_s + (((inf & 0x1FFFFFu) + dd[j]) << 3));
    }
    uint32_t ng[2 * NP];
#pragma unroll
    for (int c = 0; c < NP; ++c) {
      ng[2 * c] = qec_gap_len(nr[c].c[0], C);
      ng[2 * c + 1] = qec_gap_len(nr[c].c[2], C);
    }
#pragma unroll
    for (int j = 0; j < 2 * NP; ++j) lanes_apply(pp[j], ok[j], qq[j], lane_acc);
    if (EXPORT) {
      // the lane owns its 32-shot word: events go to the word's row in stream order, the count
      // stays in a register (no atomics)
      if (X.fmt16) {
        // delta-coded 16-bit entries (qec_sampler_sample_events16): the position IS the slot
        uint16_t *xrow16 = reinterpret_cast<uint16_t *>(xrow);
        const uint32_t cap16 = 2u * (uint32_t)X.cap;
#pragma unroll
        for (int j = 0; j < 2 * NP; ++j)
          if (ok[j]) {
            uint32_t delta = pp[j] - xprev - 1u;  // (xprev starts at -1)
            for (; delta >= 4095u; delta -= 4095u) { if (xk < cap16) xrow16[xk] = 0xFFF0u; ++xk; }
            if (xk < cap16) xrow16[xk] = (uint16_t)((delta << 4) | dd[j]);
            ++xk;
            xprev = pp[j];
          }
      } else {
#pragma unroll
        for (int j = 0; j < 2 * NP; ++j)
          if (ok[j]) {
            if (xk < (uint32_t)X.cap) xrow[xk] = (S.slloc[sl[j]] << 14) | ((pp[j] & 31u) << 9) | dd[j];
            ++xk;
          }
      }
    }
    if (TALLY) {
      uint32_t mk[2 * NP];
#pragma unroll
      for (int j = 0; j < 2 * NP; ++j) mk[j] = ok[j] ? 1u << (pp[j] & 31u) : 0u;
#pragma unroll
      for (int j = 0; j < 2 * NP; j += 4) planes_add_group(P, mk + j, 2 * NP - j < 4 ? 2 * NP - j : 4);
    }
#pragma unroll
    for (int c = 0; c < NP; ++c) rr[c] = nr[c];
#pragma unroll
    for (int j = 0; j < 2 * NP; ++j) gg[j] = ng[j];
  }
}

template <bool TALLY, bool BITS, bool EXPORT, int NP>
__global__ void __launch_bounds__(384, 1)
k_sample_lanes(const QecOutDev O, const QecLinDev S, const long long shot_lo, const long long n_shots,
               const QecSampleExport X) {
  extern __shared__ __align__(16) uint32_t qec_lin_smem[];
  const int T = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
  const uint32_t nc = (uint32_t)S.nc, ncp = (nc + 3u) & ~3u;
  const uint32_t wwords = (nc + 2u) * 32u + ncp;  // words per warp region
  uint32_t *acc = qec_lin_smem + (size_t)warp * wwords;
  uint32_t *hist = qec_lin_smem + (size_t)NW * wwords;
  uint32_t *prog = hist + 64;
  uint32_t *info2 = prog + 4u * (uint32_t)S.n_steps;
  uint32_t *rec = info2 + (((uint32_t)S.L + 1u) & ~1u);
  {
    // one word per location: first fault | number of Paulis << 21
    for (uint32_t i = tid; i < (uint32_t)S.L; i += T) info2[i] = S.info2[i].x | (S.info2[i].y << 21);
    lin_copy_records(rec, reinterpret_cast<const uint2 *>(S.sigrec), S.n_faults, tid, T);
    for (uint32_t i = tid; i < 4u * (uint32_t)S.n_steps; i += T) prog[i] = S.rowprog[i];
    for (uint32_t i = tid; i < (uint32_t)NW * wwords + 64u; i += T) qec_lin_smem[i] = 0u;
  }
  __syncthreads();

  const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(acc);
  const uint32_t lane_acc = acc_s + ((uint32_t)lane << 2);
  const uint32_t tal_s = acc_s + (((nc + 2u) * 32u) << 2);
  const uint32_t prog_s = (uint32_t)__cvta_generic_to_shared(prog);
  const uint32_t info2_s = (uint32_t)__cvta_generic_to_shared(info2);
  const uint32_t rec_s = (uint32_t)__cvta_generic_to_shared(rec);
  const uint32_t end = (uint32_t)S.L << 5;
  const float C = S.gapC;
  // no 32-bit draw reaches the clamp when 32.5 C stays below it (-log2 u <= 32): skip the test
  const bool g_noclamp = C * 32.5f < (float)QEC_GAP_CAP;
  unsigned long long nz_shots = 0;
  uint32_t h1 = 0, h2 = 0, h3 = 0, hx = 0;  // shots with exactly 1 / 2 / 3 / at least 4 faults
  const long long b0 = shot_lo >> QEC_LANES_LGB, b1 = (shot_lo + n_shots - 1) >> QEC_LANES_LGB;
  const long long w_lo = shot_lo >> 5;
  const bool vec_ok = ((O.n_words | w_lo) & 3) == 0 && ((size_t)O.bits & 15) == 0;

  LaneCounts LC;
#pragma unroll
  for (int i = 0; i < 2 * QEC_LANES_MAXTRIPS; ++i) LC.c[i] = 0u;
  LC.nb = 0u;
  // batches are handed out dynamically (one atomic per 1024 shots): results depend only on the
  // batch index, and warps that share a scheduler with more neighbours simply take fewer
  for (;;) {
    uint32_t kb = 0;
    if (lane == 0) kb = atomicAdd(S.batch_ctr, 1u);
    const long long b = b0 + (long long)__shfl_sync(0xFFFFFFFFu, kb, 0);
    if (b > b1) break;
    const long long base_shot = b << QEC_LANES_LGB;
    const uint32_t slo = shot_lo > base_shot ? (uint32_t)(shot_lo - base_shot) : 0u;
    const uint32_t shi = (uint32_t)min(1024ll, shot_lo + n_shots - base_shot);
    const uint32_t nsh = shi - slo, lane_sh = ((uint32_t)lane << 5) - slo;
    const bool full = nsh == 1024u;  // every shot of the batch is inside the requested range
    const uint32_t blo = (uint32_t)b, bhi = (uint32_t)((unsigned long long)b >> 32);

    // ---- the lane's fault stream
    LanePlanes P = {0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t xk = 0;
    const long long xw = (b << 5) + lane - w_lo;  // the lane's word in the export arrays
    const bool xin = EXPORT && xw >= 0 && xw < O.n_words;
    uint32_t *xrow = EXPORT ? X.ev + (xin ? xw : 0) * X.cap : nullptr;
    // (FAST: every shot of the batch is inside the requested range and no 32-bit draw reaches the
    // gap clamp -- the common case: no range test, no clamp test per candidate)
    if (full && g_noclamp)
      lanes_stream<TALLY, EXPORT, NP, true>(S, X, blo, bhi, lane, end, C, lane_sh, nsh, info2_s, rec_s, lane_acc, P, xrow, xk);
    else
      lanes_stream<TALLY, EXPORT, NP, false>(S, X, blo, bhi, lane, end, C, lane_sh, nsh, info2_s, rec_s, lane_acc, P, xrow, xk);
    if (EXPORT && xin) X.cnt[xw] = xk;
    __syncwarp();

    // ---- read-out (lanes_readout): four rows per pass, lane = (row of the pass, 4-word piece)
    const long long w0 = (b << 5) - w_lo;  // first output word of the batch (may be < 0 at the range start)
    const uint32_t nzb = lanes_readout<TALLY, BITS>(O, acc_s, tal_s, prog_s, (uint32_t)S.n_steps, w0, vec_ok, lane, LC, S.flush_every);
    if (TALLY) {
      nz_shots += nzb;
      // faults-per-shot histogram from the planes
      const uint32_t hi = P.p2 | P.p3 | P.p4 | P.p5;
      h1 += __popc(P.p0 & ~P.p1 & ~hi);
      h2 += __popc(~P.p0 & P.p1 & ~hi);
      h3 += __popc(P.p0 & P.p1 & ~hi);
      hx += __popc(hi);
      for (uint32_t m = hi; m; m &= m - 1u) {  // rare: shots with four or more faults
        const uint32_t j = __ffs((int)m) - 1;
        const uint32_t k = ((P.p0 >> j) & 1u) | (((P.p1 >> j) & 1u) << 1) | (((P.p2 >> j) & 1u) << 2) |
                           (((P.p3 >> j) & 1u) << 3) | (((P.p4 >> j) & 1u) << 4) | (((P.p5 >> j) & 1u) << 5);
        atomicAdd(hist + k, 1u);
      }
    }
    __syncwarp();
  }

  if (TALLY) {
    lanes_counts_flush(LC, tal_s, prog_s, (uint32_t)S.n_steps, lane);
    unsigned long long t1 = h1, t2 = h2, t3 = h3, tx = hx;
    for (int off = 16; off > 0; off >>= 1) {
      t1 += __shfl_down_sync(0xFFFFFFFFu, t1, off);
      t2 += __shfl_down_sync(0xFFFFFFFFu, t2, off);
      t3 += __shfl_down_sync(0xFFFFFFFFu, t3, off);
      tx += __shfl_down_sync(0xFFFFFFFFu, tx, off);
      nz_shots += __shfl_down_sync(0xFFFFFFFFu, nz_shots, off);
    }
    if (lane == 0) {
      if (t1) atomicAdd(&O.tally[S.nc + 2 + 1], t1);
      if (t2) atomicAdd(&O.tally[S.nc + 2 + 2], t2);
      if (t3) atomicAdd(&O.tally[S.nc + 2 + 3], t3);
      if (t1 + t2 + t3 + tx) atomicAdd(&O.tally[S.nc + 2], 0ull - (t1 + t2 + t3 + tx));
      if (nz_shots) atomicAdd(&O.tally[S.nc], nz_shots);
    }
    __syncthreads();
    for (uint32_t c = tid; c < nc; c += T) {
      unsigned long long v = 0;
      for (int k = 0; k < NW; ++k) v += qec_lin_smem[(size_t)k * wwords + (nc + 2u) * 32u + c];
      if (v) atomicAdd(&O.tally[c], v);
    }
    for (int c = tid; c < 64; c += T)
      if (hist[c]) atomicAdd(&O.tally[S.nc + 2 + c], (unsigned long long)hist[c]);
    if (blockIdx.x == 0 && tid == 0) {
      atomicAdd(&O.tally[S.nc + 1], (unsigned long long)n_shots);
      atomicAdd(&O.tally[S.nc + 2], (unsigned long long)n_shots);
    }
  }
}

// ------------------------------------------------------------------------------------------
// Linear-response enumerator: the paths of FaultEnumerator::enumerate (reference
// faultenumerator.cpp:71-111) have distinct locations, so outcome(path) = XOR of the signatures
// of its K faults -- no gate stream.  A warp owns 1024 consecutive paths, lane l the 32 paths of
// output word l: it unranks its first path once (enum_unrank) and steps through the other 31
// (enum_next), XORing each fault's 8-byte record into its own accumulator column with plain
// load / xor / store (bank = lane), exactly as the lane-stream sampler does; all lanes have the
// same 32 paths of work, so there is no imbalance.  Read-out: lanes_readout.
// Shared memory: [warp regions: acc (nc+2)*32 | tally ncp] then prog 4 * n_steps, rec 2F.
// Path stepping of the linear enumerator with its tables in shared memory (the device-memory
// versions in qec_device.cuh put three dependent global loads into every change of location):
// linfo[l] = first fault of location l | number of Paulis << 21; Es = the suffix-sum tables
// E[1..K][0..L+1] as 32-bit words when the enumeration has fewer than 2^32 paths (E32), else
// the 64-bit tables stay in device memory.
template <int K>
struct LinEnumState {
  uint32_t loc[K], dig[K], np[K], pb[K];
};
template <int K>
__device__ __forceinline__ void lin_enum_load(const uint32_t linfo_s, LinEnumState<K> &s, int j) {
  const uint32_t w = ld_s(linfo_s + (s.loc[j] << 2));
  s.pb[j] = w & 0x1FFFFFu;
  s.np[j] = w >> 21;
}
template <int K>
__device__ __forceinline__ void lin_enum_next(const uint32_t linfo_s, const uint32_t L, LinEnumState<K> &s) {
#pragma unroll
  for (int j = K - 1; j >= 0; --j) {
    if (++s.dig[j] < s.np[j]) return;
    s.dig[j] = 0;
  }
  int i = K - 1;
  while (i >= 0 && s.loc[i] >= L - K + i) --i;
  if (i < 0) return;  // past the last path: the caller never uses the state again
  const uint32_t base = s.loc[i] + 1;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    if (j >= i) {
      s.loc[j] = base + (uint32_t)(j - i);
      lin_enum_load<K>(linfo_s, s, j);
    }
  }
}
template <int K, bool E32>
__device__ __forceinline__ void lin_enum_unrank(const QecEnumDev &E, const uint32_t es_s, const uint32_t linfo_s,
                                                unsigned long long r, LinEnumState<K> &s) {
  const uint32_t L2 = (uint32_t)E.L + 2u;
  uint32_t mult = 1, start = 0;
#pragma unroll
  for (int j = 0; j < K; ++j) {
    // level K - j of the tables; Es holds levels 1..K
    const uint32_t lvl_s = es_s + (((uint32_t)(K - j - 1) * L2) << 2);
    const unsigned long long *Em = E.E + (size_t)(K - j) * L2;
    auto at = [&](uint32_t l) -> unsigned long long {
      return E32 ? (unsigned long long)ld_s(lvl_s + (l << 2)) : Em[l];
    };
    const unsigned long long u = E32 ? (unsigned long long)((uint32_t)r / mult) : qec_div_u64_u32(r, mult);
    const unsigned long long e0 = at(start);
    const unsigned long long thr = e0 - u;  // >= 1
    uint32_t lo = start, hi = (uint32_t)E.L;  // largest l in [start, L] with Em[l] >= thr
    while (hi - lo > 1) {
      const uint32_t mid = (lo + hi) >> 1;
      if (at(mid) >= thr) lo = mid; else hi = mid;
    }
    r -= (unsigned long long)mult * (e0 - at(lo));
    s.loc[j] = lo;
    lin_enum_load<K>(linfo_s, s, j);
    mult *= s.np[j];
    start = lo + 1;
  }
  uint32_t rr = (uint32_t)r;  // index inside the subset's product
#pragma unroll
  for (int j = K - 1; j >= 0; --j) {
    const uint32_t q = rr / s.np[j];
    s.dig[j] = rr - q * s.np[j];
    rr = q;
  }
}

template <int K, bool TALLY, bool BITS, bool E32>
__global__ void __launch_bounds__(384, 1)
k_enum_lanes(const QecOutDev O, const QecEnumDev E, const QecEnumLinDev EL, const long long lo, const long long hi) {
  extern __shared__ __align__(16) uint32_t qec_lin_smem[];
  const int T = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
  const uint32_t nc = (uint32_t)EL.nc, ncp = (nc + 3u) & ~3u;
  const uint32_t wwords = (nc + 2u) * 32u + ncp;
  const uint32_t L = (uint32_t)E.L, L2 = L + 2u;
  uint32_t *acc = qec_lin_smem + (size_t)warp * wwords;
  uint32_t *prog = qec_lin_smem + (size_t)NW * wwords;
  uint32_t *linfo = prog + 4u * (uint32_t)EL.n_steps;
  uint32_t *es = linfo + L;
  uint32_t *rec = es + (E32 ? (uint32_t)K * L2 : 0u);
  rec += (uint32_t)(rec - qec_lin_smem) & 1u;  // 8-byte aligned
  {
    lin_copy_records(rec, EL.sigrec, EL.n_faults, tid, T);
    for (uint32_t i = tid; i < 4u * (uint32_t)EL.n_steps; i += T) prog[i] = EL.rowprog[i];
    for (uint32_t i = tid; i < L; i += T) linfo[i] = E.poff[i] | (E.npaul[i] << 21);
    if (E32)
      for (uint32_t i = tid; i < (uint32_t)K * L2; i += T) es[i] = (uint32_t)E.E[L2 + i];  // levels 1..K
    for (uint32_t i = tid; i < (uint32_t)NW * wwords; i += T) qec_lin_smem[i] = 0u;
  }
  __syncthreads();
  const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(acc);
  const uint32_t lane_acc = acc_s + ((uint32_t)lane << 2);
  const uint32_t tal_s = acc_s + (((nc + 2u) * 32u) << 2);
  const uint32_t prog_s = (uint32_t)__cvta_generic_to_shared(prog);
  const uint32_t linfo_s = (uint32_t)__cvta_generic_to_shared(linfo);
  const uint32_t es_s = (uint32_t)__cvta_generic_to_shared(es);
  const uint32_t rec_s = (uint32_t)__cvta_generic_to_shared(rec);
  const bool vec_ok = (O.n_words & 3) == 0 && ((size_t)O.bits & 15) == 0;
  const long long n_paths = hi - lo;
  const long long n_batches = (n_paths + 1023) >> 10;
  unsigned long long nz_paths = 0;
  LaneCounts LC;
#pragma unroll
  for (int i = 0; i < 2 * QEC_LANES_MAXTRIPS; ++i) LC.c[i] = 0u;
  LC.nb = 0u;

  for (;;) {
    uint32_t kb = 0;
    if (lane == 0) kb = atomicAdd(EL.batch_ctr, 1u);
    const long long b = (long long)__shfl_sync(0xFFFFFFFFu, kb, 0);
    if (b >= n_batches) break;
    const long long first = (b << 10) + ((long long)lane << 5);  // relative to lo
    const int nv = (int)max(0ll, min(32ll, n_paths - first));
    if (nv > 0) {
      LinEnumState<K> st;
      lin_enum_unrank<K, E32>(E, es_s, linfo_s, (unsigned long long)(lo + first), st);
      // (software-pipelined: the next path's records are loaded before this path's flips)
      uint2 qn[K];
#pragma unroll
      for (int j = 0; j < K; ++j) qn[j] = ld_s64(rec_s + ((st.pb[j] + st.dig[j]) << 3));
      for (int i = 0; i < nv; ++i) {
        const uint32_t msk = 1u << i;
        uint2 q[K];
#pragma unroll
        for (int j = 0; j < K; ++j) q[j] = qn[j];
        lin_enum_next<K>(linfo_s, L, st);
#pragma unroll
        for (int j = 0; j < K; ++j) qn[j] = ld_s64(rec_s + ((st.pb[j] + st.dig[j]) << 3));
#pragma unroll
        for (int j = 0; j < K; ++j) {
          // (two faults of a path may flip the same row: one record at a time, loads first)
          const uint32_t a0 = lane_acc + (__byte_perm(q[j].x, 0u, 0x4441u) << 7);
          const uint32_t a1 = lane_acc + (__byte_perm(q[j].x, 0u, 0x4442u) << 7);
          const uint32_t a2 = lane_acc + (__byte_perm(q[j].x, 0u, 0x4443u) << 7);
          const uint32_t a3 = lane_acc + (__byte_perm(q[j].y, 0u, 0x4440u) << 7);
          const uint32_t a4 = lane_acc + (__byte_perm(q[j].y, 0u, 0x4441u) << 7);
          const uint32_t a5 = lane_acc + (__byte_perm(q[j].y, 0u, 0x4442u) << 7);
          const uint32_t a6 = lane_acc + (__byte_perm(q[j].y, 0u, 0x4443u) << 7);
          const uint32_t v0 = ld_s(a0), v1 = ld_s(a1), v2 = ld_s(a2), v3 = ld_s(a3), v4 = ld_s(a4),
                         v5 = ld_s(a5), v6 = ld_s(a6);
          st_s(a0, v0 ^ msk); st_s(a1, v1 ^ msk); st_s(a2, v2 ^ msk); st_s(a3, v3 ^ msk);
          st_s(a4, v4 ^ msk); st_s(a5, v5 ^ msk); st_s(a6, v6 ^ msk);
        }
      }
    }
    __syncwarp();
    const uint32_t nzb = lanes_readout<TALLY, BITS>(O, acc_s, tal_s, prog_s, (uint32_t)EL.n_steps, b << 5, vec_ok, lane, LC, EL.flush_every);
    if (TALLY) nz_paths += nzb;
    __syncwarp();
  }

  if (TALLY) {
    lanes_counts_flush(LC, tal_s, prog_s, (uint32_t)EL.n_steps, lane);
    for (int off = 16; off > 0; off >>= 1) nz_paths += __shfl_down_sync(0xFFFFFFFFu, nz_paths, off);
    if (lane == 0 && nz_paths) atomicAdd(&O.tally[nc], nz_paths);
    __syncthreads();
    for (uint32_t c = tid; c < nc; c += T) {
      unsigned long long v = 0;
      for (int k = 0; k < NW; ++k) v += qec_lin_smem[(size_t)k * wwords + (nc + 2u) * 32u + c];
      if (v) atomicAdd(&O.tally[c], v);
    }
    if (blockIdx.x == 0 && tid == 0) atomicAdd(&O.tally[nc + 1], (unsigned long long)n_paths);
  }
}

}  // namespace

int qec_lin_set_smem_attr(size_t smem) {
  const int s = (int)smem;
  cudaError_t e;
#define QEC_SET(fn) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, s); \
  if (e != cudaSuccess) return (int)e;
  QEC_SET((k_sample_lin<true, true>)) QEC_SET((k_sample_lin<true, false>))
  QEC_SET((k_sample_lin<false, true>)) QEC_SET((k_sample_lin<false, false>))
  QEC_SET((k_sample_linw<true, true, false>)) QEC_SET((k_sample_linw<true, false, false>))
  QEC_SET((k_sample_linw<false, true, false>)) QEC_SET((k_sample_linw<false, false, false>))
#define QEC_SETL(NP) QEC_SET((k_sample_lanes<true, true, false, NP>)) QEC_SET((k_sample_lanes<true, false, false, NP>)) \
  QEC_SET((k_sample_lanes<false, true, false, NP>)) QEC_SET((k_sample_lanes<false, false, false, NP>)) \
  QEC_SET((k_sample_lanes<true, true, true, NP>)) QEC_SET((k_sample_lanes<true, false, true, NP>)) \
  QEC_SET((k_sample_lanes<false, true, true, NP>)) QEC_SET((k_sample_lanes<false, false, true, NP>))
  QEC_SETL(1) QEC_SETL(2)
#undef QEC_SETL
  QEC_SET((k_sample_linw<true, true, true>)) QEC_SET((k_sample_linw<true, false, true>))
  QEC_SET((k_sample_linw<false, true, true>)) QEC_SET((k_sample_linw<false, false, true>))
#undef QEC_SET
  return 0;
}

void qec_launch_sample_lin(const QecLaunchCfg &cfg, cudaStream_t stream, const QecOutDev &O,
                           const QecLinDev &S, long long shot_lo, long long n_shots,
                           const QecSampleExport &X) {
  const bool t = O.tally != nullptr, b = O.bits != nullptr;
#define QEC_GO(K) K<<<cfg.grid, cfg.block, cfg.smem, stream>>>(O, S, shot_lo, n_shots, X)
  if (S.lane_mode) {
    const bool x = X.cnt != nullptr;
#define QEC_GOL(NP)                                                          \
  if (x) {                                                                   \
    if (t && b) QEC_GO((k_sample_lanes<true, true, true, NP>));              \
    else if (t) QEC_GO((k_sample_lanes<true, false, true, NP>));             \
    else if (b) QEC_GO((k_sample_lanes<false, true, true, NP>));             \
    else QEC_GO((k_sample_lanes<false, false, true, NP>));                   \
  } else {                                                                   \
    if (t && b) QEC_GO((k_sample_lanes<true, true, false, NP>));             \
    else if (t) QEC_GO((k_sample_lanes<true, false, false, NP>));            \
    else if (b) QEC_GO((k_sample_lanes<false, true, false, NP>));            \
    else QEC_GO((k_sample_lanes<false, false, false, NP>));                  \
  }
    // (three and four calls per trip measured slower: the loop body outgrows the instruction cache)
    if (S.lanes_np == 1) { QEC_GOL(1) } else { QEC_GOL(2) }
#undef QEC_GOL
  } else if (S.warp_mode && S.fast) {
    if (t && b) QEC_GO((k_sample_linw<true, true, true>));
    else if (t) QEC_GO((k_sample_linw<true, false, true>));
    else if (b) QEC_GO((k_sample_linw<false, true, true>));
    else QEC_GO((k_sample_linw<false, false, true>));
  } else if (S.warp_mode) {
    if (t && b) QEC_GO((k_sample_linw<true, true, false>));
    else if (t) QEC_GO((k_sample_linw<true, false, false>));
    else if (b) QEC_GO((k_sample_linw<false, true, false>));
    else QEC_GO((k_sample_linw<false, false, false>));
  } else {
    if (t && b) QEC_GO((k_sample_lin<true, true>));
    else if (t) QEC_GO((k_sample_lin<true, false>));
    else if (b) QEC_GO((k_sample_lin<false, true>));
    else QEC_GO((k_sample_lin<false, false>));
  }
#undef QEC_GO
}

int qec_enum_lanes_set_smem_attr(size_t smem) {
  const int s = (int)smem;
  cudaError_t e;
#define QEC_SET(fn) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, s); \
  if (e != cudaSuccess) return (int)e;
#define QEC_SETK(K) QEC_SET((k_enum_lanes<K, true, true, true>)) QEC_SET((k_enum_lanes<K, true, false, true>)) \
  QEC_SET((k_enum_lanes<K, false, true, true>)) QEC_SET((k_enum_lanes<K, false, false, true>)) \
  QEC_SET((k_enum_lanes<K, true, true, false>)) QEC_SET((k_enum_lanes<K, true, false, false>)) \
  QEC_SET((k_enum_lanes<K, false, true, false>)) QEC_SET((k_enum_lanes<K, false, false, false>))
  QEC_SETK(1) QEC_SETK(2) QEC_SETK(3) QEC_SETK(4)
#undef QEC_SETK
#undef QEC_SET
  return 0;
}

void qec_launch_enum_lanes(const QecLaunchCfg &cfg, cudaStream_t stream, const QecOutDev &O,
                           const QecEnumDev &E, const QecEnumLinDev &EL, long long lo, long long hi) {
  const bool t = O.tally != nullptr, b = O.bits != nullptr;
#define QEC_GO2(K, E32) \
  if (t && b) k_enum_lanes<K, true, true, E32><<<cfg.grid, cfg.block, cfg.smem, stream>>>(O, E, EL, lo, hi); \
  else if (t) k_enum_lanes<K, true, false, E32><<<cfg.grid, cfg.block, cfg.smem, stream>>>(O, E, EL, lo, hi); \
  else if (b) k_enum_lanes<K, false, true, E32><<<cfg.grid, cfg.block, cfg.smem, stream>>>(O, E, EL, lo, hi); \
  else k_enum_lanes<K, false, false, E32><<<cfg.grid, cfg.block, cfg.smem, stream>>>(O, E, EL, lo, hi);
#define QEC_GO(K) if (EL.e32) { QEC_GO2(K, true) } else { QEC_GO2(K, false) }
  switch (E.order) {
    case 1: QEC_GO(1) break;
    case 2: QEC_GO(2) break;
    case 3: QEC_GO(3) break;
    default: QEC_GO(4) break;
  }
#undef QEC_GO
#undef QEC_GO2
}
