/*
 * sampler_port.c -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of THIS repository's counter-based fault sampler (the kernel k_sample in
 * qiskit_qec_b200/csrc/qec_kernels.cu), so that tests can demand the GPU's sampled fault
 * lists bit for bit.  It has no reference counterpart: the reference draws from one
 * sequential std::mt19937 stream through libstdc++ distributions (reference
 * src/qiskit_qec/analysis/intern/faultsampler.cpp:49-77), which no parallel sampler can
 * reproduce; against the reference the sampler is checked statistically (binomial 5 sigma)
 * and by replaying these fault lists through the reference propagate().
 * What it keeps from the reference: per (shot, location) an independent Bernoulli(p_label)
 * failure (:55-60), then a Pauli drawn from the label's discrete weights (:64-71).
 *
 * Input: the fixed-point tables exactly as qec_sampler_tables() serialises them
 * (header[8] = B, K1, n_classes, L, n_paulis, seed_lo, seed_hi, active; then T1[K1], T2[B],
 * slop[L], slloc[L], slcls[L], cls_acc[n], cls_always[n], cls_off[n+1], cdf[np], codes[np]).
 */
#include <stdint.h>

static void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    c[0] = n0; c[1] = (uint32_t)p1; c[2] = n2; c[3] = (uint32_t)p0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

/* entries of the non-increasing table t[0..n) that are > r */
static uint32_t count_greater(const uint32_t *t, uint32_t n, uint32_t r) {
  uint32_t k = 0;
  while (k < n && t[k] > r) ++k;
  return k;
}

/* Philox4x32-10 known-answer hook for the tests (Random123 kat vectors) */
void orc_philox(uint32_t c[4], uint32_t k0, uint32_t k1) { philox4x32_10(c, k0, k1); }

/*
 * Fault events of the 32-shot word `gw` (global word index), of which the first `nv` shots
 * are inside the requested range.  ev_out[cap] receives location << 14 | shot_bit << 9 |
 * digit in generation order (sorted location major).  Returns the number of events (may
 * exceed cap; extra events are counted but not stored).
 */
int orc_sample_word(const uint32_t *tab, uint64_t gw, int nv, uint32_t *ev_out, int cap) {
  const uint32_t B = tab[0], K1 = tab[1], ncls = tab[2], L = tab[3], np = tab[4];
  const uint32_t k0 = tab[5], k1 = tab[6], active = tab[7];
  const uint32_t *T1 = tab + 8, *T2 = T1 + K1, *slop = T2 + B, *slloc = slop + L;
  const uint32_t *slcls = slloc + L, *cls_acc = slcls + L, *cls_always = cls_acc + ncls;
  const uint32_t *cls_off = cls_always + ncls, *cdf = cls_off + ncls + 1, *codes = cdf + np;
  (void)slop; (void)codes;
  const int64_t n_slots = active ? (int64_t)L * 32 : 0;
  int64_t slot = -1;
  uint32_t draw = 0;
  int n = 0;
  for (;;) {
    uint32_t c[4];
    uint64_t skip = 0;
    for (;;) {
      c[0] = (uint32_t)gw; c[1] = (uint32_t)(gw >> 32); c[2] = draw++; c[3] = 0;
      philox4x32_10(c, k0, k1);
      const uint32_t a = count_greater(T1, K1, c[0]);
      skip += (uint64_t)a * B;
      if (a < K1 || skip >= (uint64_t)n_slots) break;
    }
    const uint32_t b = count_greater(T2 + 1, B - 1, c[1]);
    slot += 1 + (int64_t)skip + b;
    if (slot >= n_slots) break;
    const uint32_t sl = (uint32_t)(slot >> 5), bit = (uint32_t)slot & 31u;
    const uint32_t cls = slcls[sl];
    if (!cls_always[cls] && c[2] >= cls_acc[cls]) continue;
    const uint32_t base = cls_off[cls], npc = cls_off[cls + 1] - base;
    uint32_t d = 0;
    while (d + 1 < npc && c[3] >= cdf[base + d]) ++d;
    if ((int)bit >= nv) continue;
    if (n < cap) ev_out[n] = (slloc[sl] << 14) | (bit << 9) | d;
    ++n;
  }
  return n;
}

/*
 * Computed geometric gap of draw schemes 1 and 2 (qec_gap_len in qec_device.cuh): u = (r | 1) /
 * 2^32, gap = floor(C * -log2(u)) clamped to 2^24, -log2(u) = (clz + 1) - log2(f) with f the top
 * 24 bits of the normalised word and log2(f) a degree-7 polynomial in f - 1.5 evaluated with
 * fused multiply-adds only (fmaf is exact IEEE: the same bits as the GPU's FFMA; this file is
 * compiled with -ffp-contract=off so that no other operation is fused).
 */
#include <math.h>
#include <string.h>
#define GAP_CAP 16777216u
static uint32_t gap_len(uint32_t r, float C) {
  static const float k[8] = {0.584962785243988f, 0.9617953896522522f, -0.32063233852386475f,
                             0.14254532754421234f, -0.07058960944414139f, 0.037311289459466934f,
                             -0.025123370811343193f, 0.015125359408557415f};
  const uint32_t x = r | 1u;
  const int n = __builtin_clz(x);
  const uint32_t fb = ((x << n) >> 8) + 0x3F000000u;
  float f;
  memcpy(&f, &fb, 4);
  const float z = f + -1.5f;
  float p = k[7];
  for (int j = 6; j >= 0; --j) p = fmaf(p, z, k[j]);
  const float l = (float)(n + 1) + -p;
  const float t = fminf(C * l, (float)GAP_CAP);
  const int32_t g = (int32_t)t; /* truncation, as __float2int_rz */
  return g > 0 ? (uint32_t)g : 0u;
}
uint32_t orc_gap_len(uint32_t r, float C) { return gap_len(r, C); }

/*
 * Draw order of the linear-response sampler (qec_linear.cu): batches of 2^lgb shots, slot =
 * sorted location << lgb | shot in batch, the slot space cut into n_streams equal ranges, one
 * Philox stream (counter = batch lo, batch hi, draw index, 0x4C000000 | stream) per range.
 * Scheme 0 (tab[7] bits 2-3): a draw advances by a * B + b + 1 slots to a candidate fault, or
 * -- when the first-level table is exhausted (a == K1) -- by K1 * B slots without one; words
 * c0, c1 = gap, c2 = class thinning, c3 = Pauli digit.  Schemes 1 / 2: computed gaps (gap_len),
 * one candidate per draw (c0 gap, c1 digit, c2 thinning) or two (c0 gap, c1 digit; c2 gap, c3
 * digit; only when no class needs thinning); a clamped gap advances GAP_CAP slots without a
 * candidate.  The stream's draws are taken 32 at a time (one warp step): a step starts only
 * while the position is before the stream's end, but then all of its 32 draws are placed.
 * Writes the accepted faults of batch `batch` as triples (sorted location, shot in batch,
 * digit) in stream order; returns their number (may exceed cap).  The same tables as above;
 * tab[7] carries lgb and n_streams, the gap constant C follows the tables.
 */
int orc_sample_batch(const uint32_t *tab, uint64_t batch, uint32_t *ev_out, int cap) {
  const uint32_t B = tab[0], K1 = tab[1], ncls = tab[2], L = tab[3], np = tab[4];
  const uint32_t k0 = tab[5], k1 = tab[6];
  const uint32_t lgb = (tab[7] >> 8) & 255u, ns = tab[7] >> 16, scheme = (tab[7] >> 2) & 3u;
  const uint32_t *T1 = tab + 8, *T2 = T1 + K1, *slop = T2 + B, *slloc = slop + L;
  const uint32_t *slcls = slloc + L, *cls_acc = slcls + L, *cls_always = cls_acc + ncls;
  const uint32_t *cls_off = cls_always + ncls, *cdf = cls_off + ncls + 1, *codes = cdf + np;
  float C;
  memcpy(&C, codes + np, 4);
  if (!(tab[7] & 2u)) return -1;
  const int64_t n_slots = (int64_t)L << lgb;
  const int64_t slen = (n_slots + ns - 1) / ns;
  const uint32_t nper = scheme == 2u ? 2u : 1u;
  int n = 0;
  if (tab[7] & 16u) {
    /* lane-stream order (k_sample_lanes: lgb = 10, k_sample_lanes16: lgb = 9): batch = 2^lgb
     * shots, lane l owns the 2^(lgb-5) shots spl * l .. spl * l + spl - 1 and walks its own stream
     * over slot = sorted location * spl + shot bit; Philox counter = (batch lo, batch hi,
     * call * 32 + lane, 0x4D000000 for 32 shots per lane / 0x4E000000 for 16), scheme 2 (two
     * candidates per call, no thinning).  Faults are written lane by lane, each lane's in stream
     * order. */
    const uint32_t lgs = lgb - 5u, spl = 1u << lgs;
    const uint32_t tag = lgb == 10u ? 0x4D000000u : 0x4E000000u;
    const int64_t lend = (int64_t)L << lgs;
    for (uint32_t lane = 0; lane < 32u; ++lane) {
      int64_t pos = -1;
      for (uint32_t call = 0; pos < lend; ++call) {
        uint32_t c[4] = {(uint32_t)batch, (uint32_t)(batch >> 32), call * 32u + lane, tag};
        philox4x32_10(c, k0, k1);
        for (uint32_t j = 0; j < 2u && pos < lend; ++j) {
          const uint32_t g = gap_len(c[2 * j], C);
          if (g >= GAP_CAP) { pos += GAP_CAP; continue; }
          pos += (int64_t)g + 1;
          if (pos >= lend) break;
          const uint32_t sl = (uint32_t)(pos >> lgs), shot = lane * spl + (uint32_t)(pos & (spl - 1u));
          const uint32_t cls = slcls[sl], base = cls_off[cls], npc = cls_off[cls + 1] - base;
          uint32_t d = 0;
          while (d + 1 < npc && c[2 * j + 1] >= cdf[base + d]) ++d;
          if (n < cap) { ev_out[3 * n] = sl; ev_out[3 * n + 1] = shot; ev_out[3 * n + 2] = d; }
          ++n;
        }
      }
    }
    return n;
  }
  for (uint32_t s = 0; s < ns; ++s) {
    const int64_t start = (int64_t)s * slen;
    const int64_t end = start + slen < n_slots ? start + slen : n_slots;
    int64_t pos = start - 1;
    for (uint32_t draw = 0;; ++draw) {
      if ((draw & 31u) == 0u && pos >= end) break; /* a warp step starts only before the end */
      uint32_t c[4] = {(uint32_t)batch, (uint32_t)(batch >> 32), draw, 0x4C000000u | s};
      philox4x32_10(c, k0, k1);
      for (uint32_t j = 0; j < nper; ++j) {
        uint32_t rd, rt = c[2];
        if (scheme == 0u) {
          const uint32_t a = count_greater(T1, K1, c[0]);
          if (a == K1) { pos += (int64_t)K1 * B; continue; }
          pos += (int64_t)a * B + count_greater(T2 + 1, B - 1, c[1]) + 1;
          rd = c[3];
        } else {
          const uint32_t g = gap_len(c[2 * j], C);
          if (g >= GAP_CAP) { pos += GAP_CAP; continue; }
          pos += (int64_t)g + 1;
          rd = c[2 * j + 1];
        }
        if (pos >= end) continue;
        const uint32_t sl = (uint32_t)(pos >> lgb), shot = (uint32_t)(pos & ((1u << lgb) - 1u));
        const uint32_t cls = slcls[sl];
        if (!cls_always[cls] && rt >= cls_acc[cls]) continue;
        const uint32_t base = cls_off[cls], npc = cls_off[cls + 1] - base;
        uint32_t d = 0;
        while (d + 1 < npc && rd >= cdf[base + d]) ++d;
        if (n < cap) { ev_out[3 * n] = sl; ev_out[3 * n + 1] = shot; ev_out[3 * n + 2] = d; }
        ++n;
      }
    }
  }
  return n;
}
