// qec_lanes16.cu -- half-word lane-stream sampler for sm_100a (FaultSampler::sample, bulk path).
//
// Same idea as k_sample_lanes (qec_linear.cu): outcome(shot) = XOR of the signatures of the
// shot's faults (reference errorpropagator.cpp:211-254, :272-310 are GF(2)-linear), every lane
// owns a column of shots and walks its own geometric-gap fault stream, all flips are plain
// load / xor / store on the lane's own shared-memory bank.  ncu on that kernel showed neither
// HBM nor a pipe as the bound but latency: 19 KB of accumulator per warp leave 9 warps per SM
// (14 % occupancy, issue slots 56 % busy).  Here a lane owns 16 shots instead of 32 and a 32-bit
// accumulator word packs TWO clbit rows (QecLanes16Dev in qec_device.cuh): 9.6 KB per warp at
// d=5, 16 warps per SM.  What changes with it:
//   * batch = 512 shots per warp; lane stream over slot = sorted location * 16 + shot bit,
//     Philox counter = (batch lo, batch hi, call * 32 + lane, 0x4E000000);
//   * records are 16-bit entries (word-row * 128 | half * 16) applied with 32-bit load / xor /
//     store, the shot mask shifted into the entry's half (both halves of a word belong to the
//     same lane: still bank = lane); the host pairs only rows that no fault flips together;
//   * the read-out un-interleaves the two rows of a word-row with byte permutes and stores 64
//     bytes per clbit row and batch (two full sectors);
//   * per-clbit tallies are one CTA-wide row updated with red.shared.add by one lane per row;
//   * the tables (records, record index, per-location info, read-out program: ~50 KB at d=5)
//     arrive in shared memory by four cp.async.bulk copies on an mbarrier instead of a copy loop;
//   * fault export without atomics: the two lanes of a 32-shot word fill its event row from
//     the two ends.
#include <cuda_runtime.h>
#include "qec_kernels.cuh"

namespace {

#define QEC_L16_LGB 9
#define QEC_L16_UNROLL 4

__device__ __forceinline__ uint32_t l16_ld(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint2 l16_ld64(uint32_t addr) {
  uint2 v;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ uint4 l16_ld128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void l16_st128(uint32_t addr, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(x), "r"(y), "r"(z), "r"(w) : "memory");
}
__device__ __forceinline__ void l16_st(uint32_t addr, uint32_t v) {
  asm volatile("st.shared.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t l16_ldh(uint32_t addr) {
  uint32_t v;
  asm volatile("{ .reg .u16 t; ld.shared.u16 t, [%1]; cvt.u32.u16 %0, t; }" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void l16_red_add(uint32_t addr, uint32_t v) {
  asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}

// One candidate of a lane: seven load / xor / store triples on the lane's own column.  A record
// entry is 16 bits: word-row * 128 | half * 16 -- bits 15..7 address the word-row, and the low five
// bits ARE the shift that moves the shot mask into the entry's half-word (0 or 16: a wrapping
// shift reads exactly those bits), so there is no half test.  32-bit accesses on purpose: with
// ld/st.shared.u16 the same kernel ran 3.6x slower (sub-word shared-memory stores; measured,
// profiles/r02).  Absent entries name the dummy word-row; a lane without a fault XORs a zero mask.
// All loads come first, so two entries of one fault must never share a word: the host pairs only
// rows that no single fault flips together (sampler_plan_lanes16).
__device__ __forceinline__ uint32_t l16_shl_wrap(uint32_t v, uint32_t n) {
  uint32_t r;
  asm("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(0u), "r"(v), "r"(n));
  return r;
}
__device__ __forceinline__ void l16_apply(const uint4 q, const uint32_t msk, const uint32_t lane_acc) {
  // fields: q.x lo, q.x hi, q.y lo, q.y hi, q.z lo, q.z hi, q.w lo (q.w hi: number of entries)
  const uint32_t f1 = q.x >> 16, f3 = q.y >> 16, f5 = q.z >> 16;
  const uint32_t a0 = lane_acc + (q.x & 0xFF80u), a1 = lane_acc + (f1 & 0xFF80u);
  const uint32_t a2 = lane_acc + (q.y & 0xFF80u), a3 = lane_acc + (f3 & 0xFF80u);
  const uint32_t a4 = lane_acc + (q.z & 0xFF80u), a5 = lane_acc + (f5 & 0xFF80u);
  const uint32_t a6 = lane_acc + (q.w & 0xFF80u);
  const uint32_t v0 = l16_ld(a0), v1 = l16_ld(a1), v2 = l16_ld(a2), v3 = l16_ld(a3),
                 v4 = l16_ld(a4), v5 = l16_ld(a5), v6 = l16_ld(a6);
  l16_st(a0, v0 ^ l16_shl_wrap(msk, q.x)); l16_st(a1, v1 ^ l16_shl_wrap(msk, f1));
  l16_st(a2, v2 ^ l16_shl_wrap(msk, q.y)); l16_st(a3, v3 ^ l16_shl_wrap(msk, f3));
  l16_st(a4, v4 ^ l16_shl_wrap(msk, q.z)); l16_st(a5, v5 ^ l16_shl_wrap(msk, f5));
  l16_st(a6, v6 ^ l16_shl_wrap(msk, q.w));
}

struct L16Planes {  // faults-per-shot counters of a lane's 16 shots, six bit planes
  uint32_t p0, p1, p2, p3, p4, p5;
};
__device__ __forceinline__ void l16_planes_add(L16Planes &P, uint32_t c) {
  uint32_t t = P.p0 & c; P.p0 ^= c; c = t;
  t = P.p1 & c; P.p1 ^= c; c = t;
  if (c) {
    t = P.p2 & c; P.p2 ^= c; c = t;
    t = P.p3 & c; P.p3 ^= c; c = t;
    t = P.p4 & c; P.p4 ^= c; c = t;
    t = P.p5 & c; P.p5 ^= c; c = t;
    P.p0 |= c; P.p1 |= c; P.p2 |= c; P.p3 |= c; P.p4 |= c; P.p5 |= c;  // 63 + 1 stays 63
  }
}

// The lane's fault stream of one batch.  Two Philox calls (four candidates) per trip,
// software-pipelined as in k_sample_lanes.  FULL: every shot of the batch is inside the
// requested range (no per-candidate range test).
template <bool TALLY, bool EXPORT, bool NOCLAMP, bool FULL>
__device__ __forceinline__ void l16_stream(const QecLanes16Dev &S, const uint32_t blo, const uint32_t bhi,
                                           const int lane, const uint32_t lane_acc, const uint32_t info_s,
                                           const uint32_t ridx_s, const uint32_t rec_s,
                                           const uint32_t lane_sh, const uint32_t nsh, L16Planes &P,
                                           const QecSampleExport &X, uint32_t *xrow, const int xdir,
                                           uint32_t &xk) {
  const uint32_t end = (uint32_t)S.L << 4;
  const float C = S.gapC;
  uint32_t pos = 0xFFFFFFFFu;  // -1; saturates at `end`
  QecPhilox ra = qec_philox_rk(blo, bhi, (uint32_t)lane, 0x4E000000u, S.rk);
  QecPhilox rb = qec_philox_rk(blo, bhi, 32u + (uint32_t)lane, 0x4E000000u, S.rk);
  uint32_t g0 = qec_gap_len(ra.c[0], C), g1 = qec_gap_len(ra.c[2], C);
  uint32_t g2 = qec_gap_len(rb.c[0], C), g3 = qec_gap_len(rb.c[2], C);
  for (uint32_t it = 0; __any_sync(0xFFFFFFFFu, pos + 1u <= end); it += 2u) {
    const QecPhilox na = qec_philox_rk(blo, bhi, (it + 2u) * 32u + (uint32_t)lane, 0x4E000000u, S.rk);
    const QecPhilox nb = qec_philox_rk(blo, bhi, (it + 3u) * 32u + (uint32_t)lane, 0x4E000000u, S.rk);
    const bool has0 = NOCLAMP || g0 < QEC_GAP_CAP, has1 = NOCLAMP || g1 < QEC_GAP_CAP;
    const bool has2 = NOCLAMP || g2 < QEC_GAP_CAP, has3 = NOCLAMP || g3 < QEC_GAP_CAP;
    const uint32_t p0 = pos + g0 + (has0 ? 1u : 0u), p1 = p0 + g1 + (has1 ? 1u : 0u);
    const uint32_t p2 = p1 + g2 + (has2 ? 1u : 0u), p3 = p2 + g3 + (has3 ? 1u : 0u);
    const bool live = pos + 1u <= end;
    bool ok0 = live && has0 && p0 < end, ok1 = live && has1 && p1 < end;
    bool ok2 = live && has2 && p2 < end, ok3 = live && has3 && p3 < end;
    if (!FULL) {
      ok0 = ok0 && (lane_sh + (p0 & 15u)) < nsh; ok1 = ok1 && (lane_sh + (p1 & 15u)) < nsh;
      ok2 = ok2 && (lane_sh + (p2 & 15u)) < nsh; ok3 = ok3 && (lane_sh + (p3 & 15u)) < nsh;
    }
    pos = live ? min(p3, end) : end;
    const uint32_t sl0 = ok0 ? p0 >> 4 : 0u, sl1 = ok1 ? p1 >> 4 : 0u;
    const uint32_t sl2 = ok2 ? p2 >> 4 : 0u, sl3 = ok3 ? p3 >> 4 : 0u;
    const uint32_t i0 = l16_ld(info_s + (sl0 << 2)), i1 = l16_ld(info_s + (sl1 << 2));
    const uint32_t i2 = l16_ld(info_s + (sl2 << 2)), i3 = l16_ld(info_s + (sl3 << 2));
    const uint32_t d0 = __umulhi(ra.c[1], i0 >> 21), d1 = __umulhi(ra.c[3], i1 >> 21);
    const uint32_t d2 = __umulhi(rb.c[1], i2 >> 21), d3 = __umulhi(rb.c[3], i3 >> 21);
    const uint32_t r0 = l16_ldh(ridx_s + (((i0 & 0x1FFFFFu) + d0) << 1)), r1 = l16_ldh(ridx_s + (((i1 & 0x1FFFFFu) + d1) << 1));
    const uint32_t r2 = l16_ldh(ridx_s + (((i2 & 0x1FFFFFu) + d2) << 1)), r3 = l16_ldh(ridx_s + (((i3 & 0x1FFFFFu) + d3) << 1));
    const uint4 q0 = l16_ld128(rec_s + (r0 << 4)), q1 = l16_ld128(rec_s + (r1 << 4));
    const uint4 q2 = l16_ld128(rec_s + (r2 << 4)), q3 = l16_ld128(rec_s + (r3 << 4));
    const uint32_t n0 = qec_gap_len(na.c[0], C), n1 = qec_gap_len(na.c[2], C);
    const uint32_t n2 = qec_gap_len(nb.c[0], C), n3 = qec_gap_len(nb.c[2], C);
    const uint32_t m0 = ok0 ? 1u << (p0 & 15u) : 0u, m1 = ok1 ? 1u << (p1 & 15u) : 0u;
    const uint32_t m2 = ok2 ? 1u << (p2 & 15u) : 0u, m3 = ok3 ? 1u << (p3 & 15u) : 0u;
    l16_apply(q0, m0, lane_acc);
    l16_apply(q1, m1, lane_acc);
    l16_apply(q2, m2, lane_acc);
    l16_apply(q3, m3, lane_acc);
    if (TALLY) {
      l16_planes_add(P, m0); l16_planes_add(P, m1); l16_planes_add(P, m2); l16_planes_add(P, m3);
    }
    if (EXPORT) {
      // event = original location << 14 | shot bit inside the 32-shot word << 9 | digit; the even
      // lane of a word fills the word's row from the front, the odd lane from the back
      const uint32_t hb = ((uint32_t)lane & 1u) << 4;
#define QEC_L16_X(OK, PP, SL, D)                                                                  \
  if (OK) {                                                                                       \
    if (xk < (uint32_t)X.cap) xrow[(int)xk * xdir] = (S.slloc[SL] << 14) | ((hb | ((PP) & 15u)) << 9) | (D); \
    ++xk;                                                                                         \
  }
      QEC_L16_X(ok0, p0, sl0, d0) QEC_L16_X(ok1, p1, sl1, d1) QEC_L16_X(ok2, p2, sl2, d2) QEC_L16_X(ok3, p3, sl3, d3)
#undef QEC_L16_X
    }
    ra = na; rb = nb; g0 = n0; g1 = n1; g2 = n2; g3 = n3;
  }
}

// Read-out of the warp's accumulator: four word-rows per pass (8 lanes x 128 bits per word-row:
// conflict-free), QEC_L16_UNROLL passes per trip with all loads first (the host schedules the
// passes so that no word-row shares a trip with the previous word-row of its chains); the
// time-difference basis is undone in place (word-row ^= previous word-row: both packed rows at
// once) and every word-row is left zero.  A lane ends up with 4 x 16 bits of row A and of row B:
// two byte permutes each give the two output words it stores (64 bits per row).
template <bool TALLY, bool BITS>
__device__ __forceinline__ uint32_t l16_readout(const QecOutDev &O, const uint32_t acc_s, const uint32_t tal_s,
                                                const uint32_t prog_s, const uint32_t n_steps,
                                                const long long w0, const bool vec_ok, const int lane) {
  const uint32_t rsub = (uint32_t)lane >> 3, k = (uint32_t)lane & 7u;
  const bool whole = w0 >= 0 && w0 + 16 <= O.n_words;
  const bool fastst = whole && vec_ok;
  uint32_t any01 = 0, any23 = 0;
  for (uint32_t pass = 0; pass < n_steps; pass += QEC_L16_UNROLL) {
    uint2 e[QEC_L16_UNROLL];
    uint32_t a[QEC_L16_UNROLL], ap[QEC_L16_UNROLL];
    uint4 v[QEC_L16_UNROLL], pv[QEC_L16_UNROLL];
#pragma unroll
    for (int u = 0; u < QEC_L16_UNROLL; ++u) e[u] = l16_ld64(prog_s + ((((pass + u) << 2) + rsub) << 3));
#pragma unroll
    for (int u = 0; u < QEC_L16_UNROLL; ++u) {
      a[u] = acc_s + ((e[u].x & 0xFFu) << 7) + (k << 4);
      ap[u] = acc_s + (((e[u].x >> 8) & 0xFFu) << 7) + (k << 4);
      v[u] = l16_ld128(a[u]);
      pv[u] = l16_ld128(ap[u]);
    }
    uint32_t cpk[QEC_L16_UNROLL];
#pragma unroll
    for (int u = 0; u < QEC_L16_UNROLL; ++u) {
      v[u].x ^= pv[u].x; v[u].y ^= pv[u].y; v[u].z ^= pv[u].z; v[u].w ^= pv[u].w;
      l16_st128(ap[u], 0u, 0u, 0u, 0u);
      // the next word-row of the chains needs this one: per half (chain B may end before chain A)
      const uint32_t keep = ((e[u].x & 0x20000u) ? 0x0000FFFFu : 0u) | ((e[u].x & 0x80000u) ? 0xFFFF0000u : 0u);
      l16_st128(a[u], v[u].x & keep, v[u].y & keep, v[u].z & keep, v[u].w & keep);
      const uint32_t A01 = __byte_perm(v[u].x, v[u].y, 0x5410u), A23 = __byte_perm(v[u].z, v[u].w, 0x5410u);
      const uint32_t B01 = __byte_perm(v[u].x, v[u].y, 0x7632u), B23 = __byte_perm(v[u].z, v[u].w, 0x7632u);
      const bool rowok = (e[u].x & 0x40000u) != 0u;
      const uint32_t rowA = e[u].y & 0xFFFFu, rowB = e[u].y >> 16;
      if (BITS && rowok) {
        uint32_t *pa = O.bits + ((long long)rowA * O.n_words + w0 + 2 * (long long)k);
        uint32_t *pb = O.bits + ((long long)rowB * O.n_words + w0 + 2 * (long long)k);
        const bool hasB = rowB != 0xFFFFu;
        if (fastst) {
          *reinterpret_cast<uint2 *>(pa) = make_uint2(A01, A23);
          if (hasB) *reinterpret_cast<uint2 *>(pb) = make_uint2(B01, B23);
        } else if (whole) {
          pa[0] = A01; pa[1] = A23;
          if (hasB) { pb[0] = B01; pb[1] = B23; }
        } else {
          const long long j = w0 + 2 * (long long)k;
          if (j >= 0 && j < O.n_words) { pa[0] = A01; if (hasB) pb[0] = B01; }
          if (j + 1 >= 0 && j + 1 < O.n_words) { pa[1] = A23; if (hasB) pb[1] = B23; }
        }
      }
      if (TALLY) {
        any01 |= A01 | B01; any23 |= A23 | B23;
        cpk[u] = (uint32_t)(__popc(A01) + __popc(A23)) | ((uint32_t)(__popc(B01) + __popc(B23)) << 16);
      }
    }
    if (TALLY) {
#pragma unroll
      for (int u = 0; u < QEC_L16_UNROLL; ++u) {
        uint32_t c = cpk[u];  // (8-lane sums stay below 2^16: at most 512 per row)
        c += __shfl_xor_sync(0xFFFFFFFFu, c, 1);
        c += __shfl_xor_sync(0xFFFFFFFFu, c, 2);
        c += __shfl_xor_sync(0xFFFFFFFFu, c, 4);
        if (k == 0u && (e[u].x & 0x40000u)) {
          const uint32_t rowA = e[u].y & 0xFFFFu, rowB = e[u].y >> 16;
          if (c & 0xFFFFu) l16_red_add(tal_s + (rowA << 2), c & 0xFFFFu);
          if ((c >> 16) && rowB != 0xFFFFu) l16_red_add(tal_s + (rowB << 2), c >> 16);
        }
      }
    }
    __syncwarp();  // kept word-rows are read by other lanes in a later trip
  }
  uint32_t nz = 0;
  if (TALLY) {
    any01 |= __shfl_xor_sync(0xFFFFFFFFu, any01, 8); any23 |= __shfl_xor_sync(0xFFFFFFFFu, any23, 8);
    any01 |= __shfl_xor_sync(0xFFFFFFFFu, any01, 16); any23 |= __shfl_xor_sync(0xFFFFFFFFu, any23, 16);
    if (lane < 8) nz = __popc(any01) + __popc(any23);
  }
  return nz;
}

template <bool TALLY, bool BITS, bool EXPORT, bool NOCLAMP>
__global__ void __launch_bounds__(512, 1)
k_sample_lanes16(const QecOutDev O, const QecLanes16Dev S, const long long shot_lo, const long long n_shots,
                 const QecSampleExport X) {
  extern __shared__ __align__(128) uint32_t qec_l16_smem[];
  __shared__ __align__(8) unsigned long long mbar;
  const int T = blockDim.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, NW = T >> 5;
  const uint32_t nc = (uint32_t)S.nc, ncp = (nc + 3u) & ~3u;
  const uint32_t wwords = (S.n_wrows + 2u) * 32u;  // words per warp region
  uint32_t *acc = qec_l16_smem + (size_t)warp * wwords;
  uint32_t *tally = qec_l16_smem + (size_t)NW * wwords;
  uint32_t *hist = tally + ncp;
  uint32_t *prog = hist + 64;
  uint32_t *info = prog + (S.prog_bytes >> 2);
  uint32_t *ridx = info + (S.info_bytes >> 2);
  uint32_t *rec = ridx + (S.recidx_bytes >> 2);
  const uint32_t mbar_s = (uint32_t)__cvta_generic_to_shared(&mbar);

  // ---- tables: four bulk copies (global -> shared) completing on one mbarrier, issued by one
  // thread while the CTA clears its accumulators
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar_s) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    const uint32_t total = S.prog_bytes + S.info_bytes + S.recidx_bytes + S.rec_bytes;
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar_s), "r"(total) : "memory");
#define QEC_BULK(DST, SRC, BYTES)                                                                              \
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"( \
                   (uint32_t)__cvta_generic_to_shared(DST)),                                                   \
               "l"(SRC), "r"(BYTES), "r"(mbar_s)                                                               \
               : "memory")
    QEC_BULK(prog, S.prog, S.prog_bytes);
    QEC_BULK(info, S.info, S.info_bytes);
    QEC_BULK(ridx, S.recidx, S.recidx_bytes);
    QEC_BULK(rec, S.rec, S.rec_bytes);
#undef QEC_BULK
  }
  for (uint32_t i = tid; i < (uint32_t)NW * wwords + ncp + 64u; i += T) qec_l16_smem[i] = 0u;
  {
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(mbar_s) : "memory");
  }
  __syncthreads();

  const uint32_t acc_s = (uint32_t)__cvta_generic_to_shared(acc);
  const uint32_t lane_acc = acc_s + ((uint32_t)lane << 2);
  const uint32_t tal_s = (uint32_t)__cvta_generic_to_shared(tally);
  const uint32_t prog_s = (uint32_t)__cvta_generic_to_shared(prog);
  const uint32_t info_s = (uint32_t)__cvta_generic_to_shared(info);
  const uint32_t ridx_s = (uint32_t)__cvta_generic_to_shared(ridx);
  const uint32_t rec_s = (uint32_t)__cvta_generic_to_shared(rec);
  unsigned long long nz_shots = 0;
  uint32_t h1 = 0, h2 = 0, h3 = 0, hx = 0;  // shots with exactly 1 / 2 / 3 / at least 4 faults
  const long long b0 = shot_lo >> QEC_L16_LGB, b1 = (shot_lo + n_shots - 1) >> QEC_L16_LGB;
  const long long w_lo = shot_lo >> 5;
  const bool vec_ok = ((O.n_words | w_lo) & 1) == 0 && ((size_t)O.bits & 7) == 0;

  // batches are handed out dynamically (one atomic per 512 shots): results depend only on the
  // batch index
  for (;;) {
    uint32_t kb = 0;
    if (lane == 0) kb = atomicAdd(S.batch_ctr, 1u);
    const long long b = b0 + (long long)__shfl_sync(0xFFFFFFFFu, kb, 0);
    if (b > b1) break;
    const long long base_shot = b << QEC_L16_LGB;
    const uint32_t slo = shot_lo > base_shot ? (uint32_t)(shot_lo - base_shot) : 0u;
    const uint32_t shi = (uint32_t)min(512ll, shot_lo + n_shots - base_shot);
    const uint32_t nsh = shi - slo, lane_sh = ((uint32_t)lane << 4) - slo;
    const uint32_t blo = (uint32_t)b, bhi = (uint32_t)((unsigned long long)b >> 32);
    const long long w0 = (b << 4) - w_lo;  // first output word of the batch (may be < 0 at the range start)

    L16Planes P = {0u, 0u, 0u, 0u, 0u, 0u};
    uint32_t xk = 0;
    uint32_t *xrow = nullptr;
    int xdir = 1;
    if (EXPORT) {
      const long long wr = w0 + (lane >> 1);
      const bool inw = wr >= 0 && wr < O.n_words;  // (outside: every shot of the word is outside the range)
      xrow = X.ev + (inw ? wr : 0) * X.cap + ((lane & 1) ? X.cap - 1 : 0);
      xdir = (lane & 1) ? -1 : 1;
    }
    if (nsh == 512u) l16_stream<TALLY, EXPORT, NOCLAMP, true>(S, blo, bhi, lane, lane_acc, info_s, ridx_s, rec_s, lane_sh, nsh, P, X, xrow, xdir, xk);
    else l16_stream<TALLY, EXPORT, NOCLAMP, false>(S, blo, bhi, lane, lane_acc, info_s, ridx_s, rec_s, lane_sh, nsh, P, X, xrow, xdir, xk);
    if (EXPORT) {
      // events of the word: even lane's count | odd lane's count << 16
      const uint32_t other = __shfl_down_sync(0xFFFFFFFFu, xk, 1);
      const long long wr = w0 + (lane >> 1);
      if (!(lane & 1) && wr >= 0 && wr < O.n_words) X.cnt[wr] = (xk & 0xFFFFu) | (other << 16);
    }
    __syncwarp();

    const uint32_t nzb = l16_readout<TALLY, BITS>(O, acc_s, tal_s, prog_s, (uint32_t)S.n_steps, w0, vec_ok, lane);
    if (TALLY) {
      nz_shots += nzb;
      const uint32_t hi = P.p2 | P.p3 | P.p4 | P.p5;
      h1 += __popc(P.p0 & ~P.p1 & ~hi);
      h2 += __popc(~P.p0 & P.p1 & ~hi & 0xFFFFu);
      h3 += __popc(P.p0 & P.p1 & ~hi);
      hx += __popc(hi);
      for (uint32_t m = hi; m; m &= m - 1u) {  // rare: shots with four or more faults
        const uint32_t j = __ffs((int)m) - 1;
        const uint32_t kk = ((P.p0 >> j) & 1u) | (((P.p1 >> j) & 1u) << 1) | (((P.p2 >> j) & 1u) << 2) |
                            (((P.p3 >> j) & 1u) << 3) | (((P.p4 >> j) & 1u) << 4) | (((P.p5 >> j) & 1u) << 5);
        atomicAdd(hist + kk, 1u);
      }
    }
    __syncwarp();
  }

  if (TALLY) {
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
    for (uint32_t c = tid; c < nc; c += T)
      if (tally[c]) atomicAdd(&O.tally[c], (unsigned long long)tally[c]);
    for (int c = tid; c < 64; c += T)
      if (hist[c]) atomicAdd(&O.tally[S.nc + 2 + c], (unsigned long long)hist[c]);
    if (blockIdx.x == 0 && tid == 0) {
      atomicAdd(&O.tally[S.nc + 1], (unsigned long long)n_shots);
      atomicAdd(&O.tally[S.nc + 2], (unsigned long long)n_shots);
    }
  }
}

}  // namespace

int qec_lanes16_set_smem_attr(size_t smem) {
  const int s = (int)smem;
  cudaError_t e;
#define QEC_SET(fn) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, s); \
  if (e != cudaSuccess) return (int)e;
#define QEC_SET4(X, C) QEC_SET((k_sample_lanes16<true, true, X, C>)) QEC_SET((k_sample_lanes16<true, false, X, C>)) \
  QEC_SET((k_sample_lanes16<false, true, X, C>)) QEC_SET((k_sample_lanes16<false, false, X, C>))
  QEC_SET4(false, false) QEC_SET4(false, true) QEC_SET4(true, false) QEC_SET4(true, true)
#undef QEC_SET4
#undef QEC_SET
  return 0;
}

void qec_launch_sample_lanes16(const QecLaunchCfg &cfg, cudaStream_t stream, const QecOutDev &O,
                               const QecLanes16Dev &S, long long shot_lo, long long n_shots,
                               const QecSampleExport &X) {
  const bool t = O.tally != nullptr, b = O.bits != nullptr, x = X.cnt != nullptr, c = S.noclamp != 0;
#define QEC_GO(TT, BB, XX, CC) k_sample_lanes16<TT, BB, XX, CC><<<cfg.grid, cfg.block, cfg.smem, stream>>>(O, S, shot_lo, n_shots, X)
#define QEC_GO2(XX, CC)                      \
  if (t && b) QEC_GO(true, true, XX, CC);    \
  else if (t) QEC_GO(true, false, XX, CC);   \
  else if (b) QEC_GO(false, true, XX, CC);   \
  else QEC_GO(false, false, XX, CC);
  if (x && c) { QEC_GO2(true, true) }
  else if (x) { QEC_GO2(true, false) }
  else if (c) { QEC_GO2(false, true) }
  else { QEC_GO2(false, false) }
#undef QEC_GO2
#undef QEC_GO
}
