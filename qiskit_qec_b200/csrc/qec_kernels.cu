// qec_kernels.cu -- sm_100a kernels of the fault-analysis hot path.
//
//   k_propagate_batch  (K1) explicit fault lists            <- ErrorPropagator::propagate
//   k_enumerate<K>     (K2) in-kernel unranking of order-K  <- FaultEnumerator::enumerate
//   k_sample           (K3) Philox fault sampling           <- FaultSampler::sample
//   k_outcome_hist     (K4) hashed full-outcome histogram
//   k_int_peak              LOP3 throughput microbenchmark (roofline denominator)
//
// All three propagation kernels are persistent: grid = SMs x resident CTAs, each CTA walks
// chunks of T lane words (T*32 paths); one thread owns one lane word whose frame lives in
// shared memory (qec_device.cuh).  GF(2) bit logic: no tensor cores by design.
#include <cuda_runtime.h>
#include "qec_kernels.cuh"

namespace {

// shared memory carve-up: [frame 2*nq*T][tally nc][hist 64][misc 4][tables ...]
__device__ __forceinline__ uint32_t *smem_base() {
  extern __shared__ __align__(16) uint32_t qec_smem[];
  return qec_smem;
}

__device__ __forceinline__ void flush_tally(const QecCircuitDev &C, const QecOutDev &O,
                                            uint32_t *s_tally, uint32_t *s_hist, bool with_hist,
                                            unsigned long long nz_local, long long n_paths) {
  // nonzero-outcome count: warp reduce then one atomic per warp
  unsigned long long nz = nz_local;
  for (int off = 16; off > 0; off >>= 1) nz += __shfl_down_sync(0xFFFFFFFFu, nz, off);
  if ((threadIdx.x & 31) == 0 && nz) atomicAdd(&O.tally[C.nc], nz);
  __syncthreads();
  for (int c = threadIdx.x; c < C.nc; c += blockDim.x) {
    const uint32_t v = s_tally[c];
    if (v) atomicAdd(&O.tally[c], (unsigned long long)v);
  }
  if (with_hist) {
    for (int c = threadIdx.x; c < 64; c += blockDim.x) {
      const uint32_t v = s_hist[c];
      if (v) atomicAdd(&O.tally[C.nc + 2 + c], (unsigned long long)v);
    }
  }
  if (blockIdx.x == 0 && threadIdx.x == 0)
    atomicAdd(&O.tally[C.nc + 1], (unsigned long long)n_paths);
}

// ------------------------------------------------------------------------------------------
// K1: explicit fault lists, prepared per lane word (CSR) by the host side of the C ABI.
__global__ void __launch_bounds__(256)
k_propagate_batch(const QecCircuitDev C, const QecOutDev O, const long long *__restrict__ ev_off,
                  const uint32_t *__restrict__ ev) {
  uint32_t *sm = smem_base();
  const int T = blockDim.x;
  uint32_t *fr = sm + threadIdx.x;
  uint32_t any = 0;
  for (long long ch = blockIdx.x; ch * T < O.n_words; ch += gridDim.x) {
    const long long w = ch * T + threadIdx.x;
    const bool valid = w < O.n_words;
    long long base = 0;
    int n = 0;
    if (valid) {
      base = ev_off[w];
      n = (int)(ev_off[w + 1] - base);
    }
    qec_propagate_word<false>(C, fr, T, ev + base, 1, n, valid, w, O, nullptr, any);
  }
}

// ------------------------------------------------------------------------------------------
// K2: order-K enumeration.  Rank -> (k-subset, Pauli digits) by closed-form unranking over
// the suffix tables E[m][l]; the following 31 paths of the lane word by successor steps.
// Order of paths: reference combinations.cpp:18-32 (lexicographic subsets) and
// productiterator.cpp:17-32 (last digit fastest).
template <int K, bool TALLY>
__global__ void __launch_bounds__(256)
k_enumerate(const QecCircuitDev C, const QecOutDev O, const QecEnumDev E, const long long lo,
            const long long hi, uint32_t *__restrict__ scratch,
            const long long *__restrict__ word_list, const unsigned int *__restrict__ list_count,
            const long long count_paths) {
  uint32_t *sm = smem_base();
  const int T = blockDim.x;
  uint32_t *fr = sm + threadIdx.x;
  uint32_t *s_tally = sm + (size_t)2 * C.nq * T;
  if (TALLY) {
    for (int c = threadIdx.x; c < C.nc; c += T) s_tally[c] = 0;
    __syncthreads();
  }
  const long long stride = (long long)gridDim.x * T;
  uint32_t *ev = scratch + (long long)blockIdx.x * T + threadIdx.x;
  unsigned long long nz = 0;
  // items: every lane word of the range, or (word_list) the lane words the specialised
  // kernel deferred to this one
  const long long n_items = word_list ? (long long)*list_count : O.n_words;
  for (long long ch = blockIdx.x; ch * T < n_items; ch += gridDim.x) {
    const long long it = ch * T + threadIdx.x;
    const bool valid = it < n_items;
    const long long w = !valid ? 0 : (word_list ? word_list[it] : it);
    int n_ev = 0;
    if (valid) {
      const long long r0 = lo + 32 * w;
      const int nv = (int)min((long long)32, hi - r0);
      EnumState<K> s;
      enum_unrank<K>(E, (unsigned long long)r0, s);
      for (int bit = 0; bit < nv; ++bit) {
#pragma unroll
        for (int j = 0; j < K; ++j) {
          const uint32_t code = E.pcode[s.pb[j] + s.dig[j]];
          const uint32_t e = (s.op[j] << QEC_EV_OP_SHIFT) | ((uint32_t)bit << 4) | code;
          // stable insertion by op: emission is nearly sorted already
          int p = n_ev;
          while (p > 0) {
            const uint32_t q = ev[(long long)(p - 1) * stride];
            if ((q >> QEC_EV_OP_SHIFT) <= s.op[j]) break;
            ev[(long long)p * stride] = q;
            --p;
          }
          ev[(long long)p * stride] = e;
          ++n_ev;
        }
        if (bit + 1 < nv) enum_next<K>(E, s);
      }
    }
    uint32_t any = 0;
    qec_propagate_word<TALLY>(C, fr, T, ev, stride, n_ev, valid, w, O, s_tally, any);
    if (TALLY) nz += __popc(any);
  }
  if (TALLY) flush_tally(C, O, s_tally, nullptr, false, nz, count_paths);
}

// ------------------------------------------------------------------------------------------
// K3: fault sampling.  Geometric-gap skipping over the word's slot space (sorted location x
// 32 shots) at rate pmax, thinned to each class's probability; all 32 lanes of a warp draw
// in lockstep.  Every random decision is a 32-bit integer compare against a host-built
// table, so a CPU restatement reproduces the fault lists bit for bit.
template <bool TALLY>
__global__ void __launch_bounds__(256)
k_sample(const QecCircuitDev C, const QecOutDev O, const QecSamplerDev S, const long long word_lo,
         const long long n_shots, uint32_t *__restrict__ scratch, const int cap,
         const QecSampleExport X, uint32_t *__restrict__ overflow_flag) {
  uint32_t *sm = smem_base();
  const int T = blockDim.x;
  uint32_t *fr = sm + threadIdx.x;
  uint32_t *s_tally = sm + (size_t)2 * C.nq * T;
  uint32_t *s_hist = s_tally + C.nc;
  uint32_t *s_T1 = s_hist + 64;
  uint32_t *s_T2 = s_T1 + S.K1;
  for (int c = threadIdx.x; c < C.nc + 64; c += T) s_tally[c] = 0;
  for (uint32_t i = threadIdx.x; i < S.K1; i += T) s_T1[i] = S.T1[i];
  for (uint32_t i = threadIdx.x; i < S.B; i += T) s_T2[i] = S.T2[i];
  __syncthreads();
  const long long stride = (long long)gridDim.x * T;
  uint32_t *ev = scratch + (long long)blockIdx.x * T + threadIdx.x;
  unsigned long long nz = 0;
  for (long long ch = blockIdx.x; ch * T < O.n_words; ch += gridDim.x) {
    const long long w = ch * T + threadIdx.x;
    const bool valid = w < O.n_words;
    int n_ev = 0;
    if (valid) {
      const int nv = (int)min((long long)32, n_shots - 32 * w);
      const unsigned long long gw = (unsigned long long)(word_lo + w);
      const uint32_t wlo = (uint32_t)gw, whi = (uint32_t)(gw >> 32);
      uint32_t draw = 0;
      long long slot = -1;
      uint32_t cp[6] = {0, 0, 0, 0, 0, 0};  // bit-sliced per-shot fault counters
      uint32_t *xev = X.cnt ? X.ev + w * (long long)X.cap : nullptr;
      for (;;) {
        QecPhilox r;
        unsigned long long skip = 0;
        for (;;) {
          r = qec_philox(wlo, whi, draw++, 0u, S.k0, S.k1);
          const uint32_t a = qec_count_greater(s_T1, S.K1, r.c[0]);
          skip += (unsigned long long)a * S.B;
          if (a < S.K1 || skip >= (unsigned long long)S.n_slots) break;  // a == K1: truncated table, keep drawing
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
        if (n_ev < cap)
          ev[(long long)n_ev * stride] =
              (S.slop[sl] << QEC_EV_OP_SHIFT) | (bit << 4) | S.codes[base + d];
        if (xev && n_ev < X.cap) xev[n_ev] = (S.slloc[sl] << 14) | (bit << 9) | d;
        ++n_ev;
        if (TALLY) {
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
      if (X.cnt) X.cnt[w] = (uint32_t)n_ev;
      if (n_ev > cap) {
        atomicMax(overflow_flag, (uint32_t)n_ev);
        n_ev = cap;
      }
      if (TALLY) {
        for (int bit = 0; bit < nv; ++bit) {
          uint32_t k = 0;
#pragma unroll
          for (int p = 0; p < 6; ++p) k |= ((cp[p] >> bit) & 1u) << p;
          atomicAdd(&s_hist[k], 1u);
        }
      }
    }
    uint32_t any = 0;
    qec_propagate_word<TALLY>(C, fr, T, ev, stride, n_ev, valid, w, O, s_tally, any);
    if (TALLY) nz += __popc(any);
  }
  if (TALLY) flush_tally(C, O, s_tally, s_hist, true, nz, n_shots);
}

// ------------------------------------------------------------------------------------------
// K4: hashed full-outcome histogram over a bit-sliced outcome buffer.  One warp per lane
// word: lane j loads clbit row c0+j, 32 ballots transpose the 32x32 bit tile so lane b ends
// up holding clbits c0..c0+31 of path b; tiles are folded into a 64-bit hash.
__device__ __forceinline__ unsigned long long qec_mix64(unsigned long long h, uint32_t chunk) {
  h ^= (unsigned long long)chunk;
  h *= 0x9E3779B97F4A7C15ull;
  h ^= h >> 29;
  return h;
}

// rows (may be null): uint32[table_cap][ceil(nc/32)], the outcome itself (bit c % 32 of word c / 32
// = clbit c), written once by the thread that claims the slot; status[0] = slots claimed,
// status[1] = paths dropped because the table was full, status[2] = paths with the all-zero
// outcome (counted apart: at small p it is by far the hottest key, and it never enters the table).
__global__ void __launch_bounds__(256)
k_outcome_hist(const uint32_t *__restrict__ bits, const int nc, const long long n_paths,
               const long long n_words, unsigned long long *__restrict__ keys,
               unsigned long long *__restrict__ counts, uint32_t *__restrict__ rows,
               const long long table_cap, unsigned long long *__restrict__ status) {
  const unsigned lane = threadIdx.x & 31u;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long n_warps = ((long long)gridDim.x * blockDim.x) >> 5;
  const int rw = (nc + 31) / 32;
  const long long used_words = (n_paths + 31) / 32;
  unsigned long long zeros = 0, dropped = 0, claimed = 0;
  for (long long w = warp; w < used_words; w += n_warps) {
    unsigned long long h = 0x243F6A8885A308D3ull;
    uint32_t anyv = 0;
    for (int c0 = 0; c0 < nc; c0 += 32) {
      const int c = c0 + (int)lane;
      const uint32_t row = (c < nc) ? bits[(long long)c * n_words + w] : 0u;
      uint32_t mine = 0;
#pragma unroll
      for (int b = 0; b < 32; ++b) {
        const uint32_t col = __ballot_sync(0xFFFFFFFFu, (row >> b) & 1u);
        if ((int)lane == b) mine = col;
      }
      anyv |= mine;
      h = qec_mix64(h, mine);
    }
    h ^= h >> 32;
    if (h == 0) h = 1;  // 0 marks an empty slot
    const bool valid = 32 * w + lane < n_paths;
    long long won = -1;  // slot this lane claimed (its outcome row still has to be written)
    if (valid && anyv == 0u) ++zeros;
    else if (valid) {
      unsigned long long slot = h & (unsigned long long)(table_cap - 1);
      long long probe = 0;
      for (; probe < table_cap; ++probe) {
        const unsigned long long prev = atomicCAS(&keys[slot], 0ull, h);
        if (prev == 0ull || prev == h) {
          atomicAdd(&counts[slot], 1ull);
          if (prev == 0ull) { won = (long long)slot; ++claimed; }
          break;
        }
        slot = (slot + 1) & (unsigned long long)(table_cap - 1);
      }
      if (probe == table_cap) ++dropped;
    }
    if (rows && __any_sync(0xFFFFFFFFu, won >= 0)) {
      // second transpose, only when some lane of the word claimed a new slot (rare once the
      // table holds the common outcomes): winners store their outcome row
      for (int c0 = 0; c0 < nc; c0 += 32) {
        const int c = c0 + (int)lane;
        const uint32_t row = (c < nc) ? bits[(long long)c * n_words + w] : 0u;
        uint32_t mine = 0;
#pragma unroll
        for (int b = 0; b < 32; ++b) {
          const uint32_t col = __ballot_sync(0xFFFFFFFFu, (row >> b) & 1u);
          if ((int)lane == b) mine = col;
        }
        if (won >= 0) rows[won * rw + (c0 >> 5)] = mine;
      }
    }
  }
  if (status) {
    zeros = __reduce_add_sync(0xFFFFFFFFu, (unsigned)zeros);  // (< 2^32 per warp: 32 paths per trip)
    if (claimed) atomicAdd(&status[0], claimed);
    if (dropped) atomicAdd(&status[1], dropped);
    if (lane == 0 && zeros) atomicAdd(&status[2], zeros);
  }
}

// Occupied slots of the table -> dense arrays (order arbitrary): out_counts[i], out_rows[i][rw],
// out_keys[i]; *n_out counts them.
__global__ void __launch_bounds__(256)
k_hist_compact(const unsigned long long *__restrict__ keys, const unsigned long long *__restrict__ counts,
               const uint32_t *__restrict__ rows, const long long table_cap, const int rw,
               unsigned long long *__restrict__ out_keys, unsigned long long *__restrict__ out_counts,
               uint32_t *__restrict__ out_rows, unsigned long long *__restrict__ n_out) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < table_cap;
       i += (long long)gridDim.x * blockDim.x) {
    const unsigned long long k = keys[i];
    if (!k) continue;
    const unsigned long long o = atomicAdd(n_out, 1ull);
    out_keys[o] = k;
    out_counts[o] = counts[i];
    if (rows) for (int j = 0; j < rw; ++j) out_rows[o * rw + j] = rows[i * rw + j];
  }
}

// ------------------------------------------------------------------------------------------
// Compaction of the sampler's fault export: per 32-shot word a fixed-capacity event row
// ev[w][cap] and a count (two_ended: the half-word lane-stream kernel fills a row from both
// ends, count = front | back << 16) -> one dense pool, the words' events back to back, plus a
// 16-bit count per word.  Three small launches per chunk: block sums, scan of the block sums,
// copy.  A block covers 1024 words.
#define QEC_EVC_T 256
#define QEC_EVC_WPB 1024
__device__ __forceinline__ uint32_t evc_total(uint32_t c, int two_ended) {
  return two_ended ? (c & 0xFFFFu) + (c >> 16) : c;
}
__global__ void __launch_bounds__(QEC_EVC_T)
k_ev_block_sums(const uint32_t *__restrict__ cnt, const long long n_words, const int two_ended, const int cap,
                uint32_t *__restrict__ block_sums, uint32_t *__restrict__ overflow) {
  __shared__ uint32_t part[QEC_EVC_T / 32];
  const long long w0 = (long long)blockIdx.x * QEC_EVC_WPB;
  uint32_t sum = 0, mx = 0;
  for (int j = threadIdx.x; j < QEC_EVC_WPB; j += QEC_EVC_T) {
    const long long w = w0 + j;
    const uint32_t t = w < n_words ? evc_total(cnt[w], two_ended) : 0u;
    sum += t;
    mx = max(mx, t);
  }
  sum = __reduce_add_sync(0xFFFFFFFFu, sum);
  mx = __reduce_max_sync(0xFFFFFFFFu, mx);
  if ((threadIdx.x & 31) == 0) {
    part[threadIdx.x >> 5] = sum;
    if (mx > (uint32_t)cap) atomicMax(overflow, mx);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t t = 0;
    for (int k = 0; k < QEC_EVC_T / 32; ++k) t += part[k];
    block_sums[blockIdx.x] = t;
  }
}
// one block: exclusive scan of the block sums in place, grand total out
__global__ void __launch_bounds__(1024)
k_ev_scan_blocks(uint32_t *__restrict__ block_sums, const int n_blocks, unsigned long long *__restrict__ total) {
  __shared__ unsigned long long warp_tot[32];
  __shared__ unsigned long long carry;
  if (threadIdx.x == 0) carry = 0ull;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  for (int base = 0; base < n_blocks; base += 1024) {
    const int i = base + (int)threadIdx.x;
    const unsigned long long v = i < n_blocks ? block_sums[i] : 0u;
    unsigned long long incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned long long t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
      if ((int)lane >= o) incl += t;
    }
    if (lane == 31u) warp_tot[warp] = incl;
    __syncthreads();
    unsigned long long before = carry;
    for (unsigned k = 0; k < warp; ++k) before += warp_tot[k];
    if (i < n_blocks) block_sums[i] = (uint32_t)(before + incl - v);  // (chunk totals stay below 2^32)
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry;
}
__global__ void __launch_bounds__(QEC_EVC_T)
k_ev_compact(const uint32_t *__restrict__ cnt, const uint32_t *__restrict__ ev, const long long n_words,
             const int two_ended, const int cap, const uint32_t *__restrict__ block_off,
             uint16_t *__restrict__ out_cnt, uint32_t *__restrict__ out_ev, const unsigned long long pool_cap,
             const int ev24) {
  __shared__ uint32_t warp_tot[QEC_EVC_T / 32];
  const long long w0 = (long long)blockIdx.x * QEC_EVC_WPB;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t run = block_off[blockIdx.x];
  // thread t owns words w0 + 4 t .. w0 + 4 t + 3 (consecutive words, consecutive pool ranges)
  uint32_t c[4], tot[4], mine = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const long long w = w0 + 4 * (long long)threadIdx.x + j;
    c[j] = w < n_words ? cnt[w] : 0u;
    tot[j] = min(evc_total(c[j], two_ended), (uint32_t)cap);
    mine += tot[j];
  }
  uint32_t incl = mine;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  if (lane == 31u) warp_tot[warp] = incl;
  __syncthreads();
  for (unsigned k = 0; k < warp; ++k) run += warp_tot[k];
  uint32_t o = run + incl - mine;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const long long w = w0 + 4 * (long long)threadIdx.x + j;
    if (w >= n_words) break;
    out_cnt[w] = (uint16_t)tot[j];
    if ((unsigned long long)o + tot[j] > pool_cap) break;  // (the host sees total > pool and re-runs)
    const uint32_t *row = ev + w * cap;
    const uint32_t lo = two_ended ? min(c[j] & 0xFFFFu, tot[j]) : tot[j];
    if (ev24) {
      // three bytes per event: location << 9 | shot bit << 4 | digit (the host checked the ranges)
      // (bytes collected in a shift register and stored four at a time once the position is
      // 4-byte aligned: three single-byte stores per event made this kernel 3x slower)
      uint8_t *o8 = reinterpret_cast<uint8_t *>(out_ev);
      unsigned long long bpos = 3ull * o, acc = 0ull;
      int nb = 0;
      for (uint32_t k = 0; k < tot[j]; ++k) {
        const uint32_t e = k < lo ? row[k] : row[cap - 1 - (k - lo)];
        const uint32_t e24 = ((e >> 14) << 9) | (((e >> 9) & 31u) << 4) | (e & 15u);
        acc |= (unsigned long long)e24 << (8 * nb);
        nb += 3;
        while (nb > 0 && (bpos & 3ull)) { o8[bpos++] = (uint8_t)acc; acc >>= 8; --nb; }
        if (nb >= 4) { *reinterpret_cast<uint32_t *>(o8 + bpos) = (uint32_t)acc; acc >>= 32; nb -= 4; bpos += 4; }
      }
      while (nb > 0) { o8[bpos++] = (uint8_t)acc; acc >>= 8; --nb; }
    } else {
      for (uint32_t k = 0; k < lo; ++k) out_ev[o + k] = row[k];
      for (uint32_t k = lo; k < tot[j]; ++k) out_ev[o + k] = row[cap - 1 - (k - lo)];
    }
    o += tot[j];
  }
}

// ---- 16-bit delta-coded events (qec_sampler_sample_events16) --------------------------------------
// A word's events arrive in slot order (slot = sorted location * 32 + shot bit: the lane-stream
// kernel appends them as its stream meets them).  Entry = (delta << 4) | digit, delta = slot -
// previous slot - 1 (previous = -1 at the start of a word), delta < 4095; 0xFFF0 = "4095 slots
// without an event".  The lane-stream kernel writes the entries itself (QecSampleExport::fmt16:
// rows of 2 * cap entries, cnt[w] = their number); the block sums / scan above place them and
// k_ev16_compact copies them out.
__global__ void __launch_bounds__(QEC_EVC_T)
k_ev16_compact(const uint32_t *__restrict__ cnt16, const uint16_t *__restrict__ rows16, const long long n_words,
               const int cap, const uint32_t *__restrict__ block_off, uint16_t *__restrict__ out_cnt,
               uint16_t *__restrict__ out_ev, const unsigned long long pool_cap) {
  __shared__ uint32_t warp_tot[QEC_EVC_T / 32];
  const long long w0 = (long long)blockIdx.x * QEC_EVC_WPB;
  const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
  uint32_t run = block_off[blockIdx.x];
  uint32_t tot[4], mine = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const long long w = w0 + 4 * (long long)threadIdx.x + j;
    tot[j] = w < n_words ? min(cnt16[w], 2u * (uint32_t)cap) : 0u;
    mine += tot[j];
  }
  uint32_t incl = mine;
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xFFFFFFFFu, incl, o);
    if ((int)lane >= o) incl += t;
  }
  if (lane == 31u) warp_tot[warp] = incl;
  __syncthreads();
  for (unsigned k = 0; k < warp; ++k) run += warp_tot[k];
  uint32_t o = run + incl - mine;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const long long w = w0 + 4 * (long long)threadIdx.x + j;
    if (w >= n_words) break;
    out_cnt[w] = (uint16_t)tot[j];
    if ((unsigned long long)o + tot[j] > pool_cap) break;  // (the host sees total > pool and re-runs)
    const uint16_t *row = rows16 + w * (2 * (long long)cap);
    for (uint32_t k = 0; k < tot[j]; ++k) out_ev[o + k] = row[k];
    o += tot[j];
  }
}

// ------------------------------------------------------------------------------------------
// Multi-GPU tally reduce without a collective launch: device 0 sums the tallies of its peers,
// read straight from their memory over NVLink peer mappings (ptrs[k] = tally of device k; the
// caller's stream has waited on every device's kernels).  ~1.7 KB per device: latency-bound.
__global__ void __launch_bounds__(256)
k_tally_reduce_peers(const unsigned long long *const *__restrict__ ptrs, const int n_dev, const int len,
                     unsigned long long *__restrict__ out) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < len; i += gridDim.x * blockDim.x) {
    unsigned long long v = 0;
    for (int k = 0; k < n_dev; ++k) v += ptrs[k][i];
    out[i] = v;
  }
}

// ------------------------------------------------------------------------------------------
// Integer-pipe roofline denominator: 8 independent LOP3 chains per thread, no memory.  Every
// chain step is ONE `lop3.b32` written as inline PTX (asm volatile), so that the ops counted by
// qec_measure_int_peak are the LOP3 instructions the SM executes: ptxas neither merges nor
// splits them (round 1 wrote the chains in C and ptxas re-associated 128 counted ops into ~190
// executed ones).  QEC_INT_PEAK_OPS_PER_TRIP is the count per loop trip and per thread;
// profiles/r02/sass_k_int_peak.txt holds the SASS of the loop.
#define QEC_LOP3(d, a, b, c, lut) asm volatile("lop3.b32 %0, %1, %2, %3, " #lut ";" : "=r"(d) : "r"(a), "r"(b), "r"(c))
__global__ void __launch_bounds__(256) k_int_peak(const int iters, uint32_t *__restrict__ sink) {
  uint32_t a0 = threadIdx.x, a1 = a0 * 3u + 1u, a2 = a0 * 5u + 2u, a3 = a0 * 7u + 3u;
  uint32_t a4 = a0 ^ 0x55u, a5 = a0 ^ 0xAAu, a6 = a0 + 11u, a7 = a0 + 13u;
  const uint32_t k0 = blockIdx.x * 0x9E3779B9u + 1u, k1 = ~k0;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int u = 0; u < QEC_INT_PEAK_UNROLL; ++u) {
      // 0x6A = (a & b) ^ c, 0x1E = (a | b) ^ c: one three-input logic op per chain per step
      QEC_LOP3(a0, a0, k0, a1, 0x6A); QEC_LOP3(a1, a1, k1, a2, 0x1E);
      QEC_LOP3(a2, a2, k0, a3, 0x6A); QEC_LOP3(a3, a3, k1, a4, 0x1E);
      QEC_LOP3(a4, a4, k0, a5, 0x6A); QEC_LOP3(a5, a5, k1, a6, 0x1E);
      QEC_LOP3(a6, a6, k0, a7, 0x6A); QEC_LOP3(a7, a7, k1, a0, 0x1E);
    }
  }
  const uint32_t r = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
  if (r == 0x12345678u) sink[0] = r;  // practically never: keeps the chains live
}
#undef QEC_LOP3

}  // namespace

// ---- launchers -----------------------------------------------------------------------------
size_t qec_kernel_smem_bytes(int kind, int nq, int nc, int T, uint32_t B, uint32_t K1) {
  size_t words = (size_t)2 * nq * T;
  if (kind >= 1) words += (size_t)nc;            // tally
  if (kind == 2) words += 64 + (size_t)B + K1;   // hist + gap tables
  return words * sizeof(uint32_t) + 16;
}

void qec_launch_batch(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                      const QecOutDev &O, const long long *ev_off, const uint32_t *ev) {
  k_propagate_batch<<<cfg.grid, cfg.block, cfg.smem, stream>>>(C, O, ev_off, ev);
}

template <int K>
static void launch_enum_k(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                          const QecOutDev &O, const QecEnumDev &E, long long lo, long long hi,
                          uint32_t *scratch, const long long *word_list,
                          const unsigned int *list_count) {
  const long long count_paths = word_list ? 0 : hi - lo;  // (a deferred-word pass adds no path count)
  if (O.tally)
    k_enumerate<K, true><<<cfg.grid, cfg.block, cfg.smem, stream>>>(C, O, E, lo, hi, scratch,
                                                                    word_list, list_count, count_paths);
  else
    k_enumerate<K, false><<<cfg.grid, cfg.block, cfg.smem, stream>>>(C, O, E, lo, hi, scratch,
                                                                     word_list, list_count, count_paths);
}

void qec_launch_enumerate(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                          const QecOutDev &O, const QecEnumDev &E, long long lo, long long hi,
                          uint32_t *scratch, const long long *word_list,
                          const unsigned int *list_count) {
  switch (E.order) {
    case 1: launch_enum_k<1>(cfg, stream, C, O, E, lo, hi, scratch, word_list, list_count); break;
    case 2: launch_enum_k<2>(cfg, stream, C, O, E, lo, hi, scratch, word_list, list_count); break;
    case 3: launch_enum_k<3>(cfg, stream, C, O, E, lo, hi, scratch, word_list, list_count); break;
    default: launch_enum_k<4>(cfg, stream, C, O, E, lo, hi, scratch, word_list, list_count); break;
  }
}

void qec_launch_sample(const QecLaunchCfg &cfg, cudaStream_t stream, const QecCircuitDev &C,
                       const QecOutDev &O, const QecSamplerDev &S, long long word_lo,
                       long long n_shots, uint32_t *scratch, int cap, const QecSampleExport &X,
                       uint32_t *overflow_flag) {
  if (O.tally)
    k_sample<true><<<cfg.grid, cfg.block, cfg.smem, stream>>>(C, O, S, word_lo, n_shots, scratch,
                                                            cap, X, overflow_flag);
  else
    k_sample<false><<<cfg.grid, cfg.block, cfg.smem, stream>>>(C, O, S, word_lo, n_shots, scratch,
                                                             cap, X, overflow_flag);
}

void qec_launch_outcome_hist(cudaStream_t stream, const uint32_t *bits, int nc, long long n_paths,
                             long long n_words, unsigned long long *keys, unsigned long long *counts,
                             uint32_t *rows, long long table_cap, unsigned long long *status) {
  const long long used = (n_paths + 31) / 32;
  long long blocks = (used * 32 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  k_outcome_hist<<<(int)blocks, 256, 0, stream>>>(bits, nc, n_paths, n_words, keys, counts, rows,
                                                    table_cap, status);
}

void qec_launch_hist_compact(cudaStream_t stream, const unsigned long long *keys,
                             const unsigned long long *counts, const uint32_t *rows,
                             long long table_cap, int rw, unsigned long long *out_keys,
                             unsigned long long *out_counts, uint32_t *out_rows,
                             unsigned long long *n_out) {
  long long blocks = (table_cap + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  k_hist_compact<<<(int)blocks, 256, 0, stream>>>(keys, counts, rows, table_cap, rw, out_keys,
                                                    out_counts, out_rows, n_out);
}

void qec_launch_event_compact(cudaStream_t stream, const uint32_t *cnt, const uint32_t *ev, long long n_words,
                              int two_ended, int cap, uint32_t *block_sums, uint32_t *overflow,
                              unsigned long long *total, uint16_t *out_cnt, uint32_t *out_ev,
                              unsigned long long pool_cap, int ev24) {
  const int n_blocks = (int)((n_words + QEC_EVC_WPB - 1) / QEC_EVC_WPB);
  if (n_blocks < 1) return;
  k_ev_block_sums<<<n_blocks, QEC_EVC_T, 0, stream>>>(cnt, n_words, two_ended, cap, block_sums, overflow);
  k_ev_scan_blocks<<<1, 1024, 0, stream>>>(block_sums, n_blocks, total);
  k_ev_compact<<<n_blocks, QEC_EVC_T, 0, stream>>>(cnt, ev, n_words, two_ended, cap, block_sums, out_cnt, out_ev, pool_cap, ev24);
}

void qec_launch_event_compact16(cudaStream_t stream, const uint32_t *cnt, const uint16_t *rows16, long long n_words,
                                int cap, uint32_t *block_sums, uint32_t *overflow, unsigned long long *total,
                                uint16_t *out_cnt, uint16_t *out_ev, unsigned long long pool_cap) {
  const int n_blocks = (int)((n_words + QEC_EVC_WPB - 1) / QEC_EVC_WPB);
  if (n_blocks < 1) return;
  // (a row holds 2 * cap entries: longer words are reported through `overflow` and the host re-runs)
  k_ev_block_sums<<<n_blocks, QEC_EVC_T, 0, stream>>>(cnt, n_words, 0, 2 * cap, block_sums, overflow);
  k_ev_scan_blocks<<<1, 1024, 0, stream>>>(block_sums, n_blocks, total);
  k_ev16_compact<<<n_blocks, QEC_EVC_T, 0, stream>>>(cnt, rows16, n_words, cap, block_sums, out_cnt, out_ev, pool_cap);
}

void qec_launch_tally_reduce_peers(cudaStream_t stream, const unsigned long long *const *ptrs, int n_dev,
                                   int len, unsigned long long *out) {
  k_tally_reduce_peers<<<(len + 255) / 256, 256, 0, stream>>>(ptrs, n_dev, len, out);
}

void qec_launch_int_peak(cudaStream_t stream, int grid, int block, int iters, uint32_t *sink) {
  k_int_peak<<<grid, block, 0, stream>>>(iters, sink);
}

int qec_kernel_set_smem_attr(int kind, int order, size_t smem) {
  cudaError_t e = cudaSuccess;
  const int s = (int)smem;
#define QEC_SET(fn) e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, s); \
  if (e != cudaSuccess) return (int)e;
  if (kind == 0) { QEC_SET(k_propagate_batch) }
  else if (kind == 1) {
    switch (order) {
      case 1: QEC_SET((k_enumerate<1, true>)) QEC_SET((k_enumerate<1, false>)) break;
      case 2: QEC_SET((k_enumerate<2, true>)) QEC_SET((k_enumerate<2, false>)) break;
      case 3: QEC_SET((k_enumerate<3, true>)) QEC_SET((k_enumerate<3, false>)) break;
      default: QEC_SET((k_enumerate<4, true>)) QEC_SET((k_enumerate<4, false>)) break;
    }
  } else { QEC_SET(k_sample<true>) QEC_SET(k_sample<false>) }
#undef QEC_SET
  return 0;
}
