// qec_distance.cu -- brute-force minimum distance of stabilizer / subsystem codes on sm_100a
// (reference analysis/intern/distance.cpp:17-53 minimum_distance, :71-159
// distance_test, :172-202 minimum_distance_by_tests; linear algebra analysis/intern/linear.cpp).
//
// minimum_distance: the reference walks weight = 1, 2, ... and, per weight, every w-subset of the n
// qubits (Combinations) times every assignment of {X, Z, Y} to it (ProductIterator, 3^w), and asks
// of each error e: is e in W-perp (commutes with every stabilizer row) and is e in V (rank of the
// gauge rows + e unchanged)?  Both questions are GF(2)-LINEAR in e:
//   * the syndrome of e against the stabilizer rows is the XOR of single-qubit syndromes;
//   * with the gauge rows in REDUCED row echelon form (basis rows V_b, pivot columns p_b) the
//     residual  e ^ XOR_b e[p_b] V_b  is linear in e and zero exactly when e is in V.
// The host packs [stabilizer syndrome | residual on the non-pivot columns] of the 3n single-qubit
// Paulis into one table of 64-bit words; a GPU thread unranks its w-subset (combinatorial number
// system, binomials in constant memory), walks the 3^w assignments with an odometer over prefix
// XORs and raises a flag on the first hit.  The answer is only "does weight w have a hit", so no
// order has to be kept; the host walks the weights as the reference does.
//
// distance_test (meet in the middle, the reference's own algorithm): the 64-bit syndromes of all
// w1-subsets of the 3n single-qubit terms go into an open-addressed hash set on the device, the
// w2-subsets probe it for "same stabilizer syndrome, opposite commutation with the logical".
#include <cuda_runtime.h>
#include <stdint.h>

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "qec_distance.h"

namespace {

#define QD_MAX_WORDS 4    // 256 bits of [syndrome | residual]
#define QD_MAX_WEIGHT 16  // subsets of at most 16 elements
#define QD_MAX_ELEMS 256  // n (minimum_distance) or 3n (distance_test) at most

// binomials C(i, j), i <= QD_MAX_ELEMS, j <= QD_MAX_WEIGHT, saturated at 2^63 - 1
__constant__ unsigned long long qd_binom[(QD_MAX_ELEMS + 1) * (QD_MAX_WEIGHT + 1)];

__device__ __forceinline__ unsigned long long qd_C(int i, int j) {
  return qd_binom[i * (QD_MAX_WEIGHT + 1) + j];
}

// rank -> sorted w-subset of {0..n-1} in the order of Combinations::next_combination
// (reference combinations.cpp:18-32: lexicographic)
__device__ __forceinline__ void qd_unrank(unsigned long long r, const int n, const int w, int *sub) {
  int x = 0;
  for (int i = 0; i < w; ++i) {
    // subsets that start (at this position) with x: C(n - 1 - x, w - 1 - i)
    for (;; ++x) {
      const unsigned long long c = qd_C(n - 1 - x, w - 1 - i);
      if (r < c) break;
      r -= c;
    }
    sub[i] = x++;
  }
}

// table T[(3 q + p) * NW + k]: p = 0 X, 1 Z, 2 Y (the reference's pstate 0 / 1 / 2)
template <int NW>
__global__ void __launch_bounds__(256) k_min_distance(const unsigned long long *__restrict__ T, const int n, const int w,
                                                      const unsigned long long n_subsets,
                                                      const unsigned long long mask_s0, const unsigned long long mask_s1,
                                                      const unsigned long long mask_s2, const unsigned long long mask_s3,
                                                      const int n_stab_bits, const int k_is_zero, unsigned int *flag) {
  const unsigned long long mask_s[4] = {mask_s0, mask_s1, mask_s2, mask_s3};
  (void)n_stab_bits;
  for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_subsets;
       r += (unsigned long long)gridDim.x * blockDim.x) {
    if (*reinterpret_cast<volatile unsigned int *>(flag)) return;  // some thread already has a hit
    int sub[QD_MAX_WEIGHT];
    qd_unrank(r, n, w, sub);
    // prefix XORs: pre[i] = XOR of the terms of positions < i; dig[] = the odometer (last fastest)
    unsigned long long pre[QD_MAX_WEIGHT + 1][NW];
    int dig[QD_MAX_WEIGHT];
#pragma unroll
    for (int k = 0; k < NW; ++k) pre[0][k] = 0ull;
    for (int i = 0; i < w; ++i) {
      dig[i] = 0;
#pragma unroll
      for (int k = 0; k < NW; ++k) pre[i + 1][k] = pre[i][k] ^ T[(size_t)(3 * sub[i]) * NW + k];
    }
    for (;;) {
      bool in_wperp = true, in_v = true;
#pragma unroll
      for (int k = 0; k < NW; ++k) {
        const unsigned long long a = pre[w][k];
        in_wperp = in_wperp && (a & mask_s[k]) == 0ull;
        in_v = in_v && (a & ~mask_s[k]) == 0ull;
      }
      // reference distance.cpp:45-47: in W-perp and (k > 0 ? not in V : in V)
      if (in_wperp && (k_is_zero ? in_v : !in_v)) { atomicExch(flag, 1u); return; }
      int j = w - 1;
      for (; j >= 0; --j) { if (++dig[j] < 3) break; dig[j] = 0; }
      if (j < 0) break;
      for (int i = j; i < w; ++i) {
#pragma unroll
        for (int k = 0; k < NW; ++k) pre[i + 1][k] = pre[i][k] ^ T[(size_t)(3 * sub[i] + dig[i]) * NW + k];
      }
    }
  }
}

// ---- distance_test: hash set of 64-bit syndromes ---------------------------------------------
#define QD_EMPTY 0xFFFFFFFFFFFFFFFFull  // (a syndrome has at most 63 bits: m < 63)

__device__ __forceinline__ unsigned long long qd_hash(unsigned long long x) {
  x ^= x >> 33; x *= 0xFF51AFD7ED558CCDull; x ^= x >> 33; x *= 0xC4CEB9FE1A85EC53ull; x ^= x >> 33;
  return x;
}

__device__ __forceinline__ unsigned long long qd_subset_syndrome(const unsigned long long *__restrict__ sq,
                                                                 unsigned long long r, const int N, const int w) {
  int sub[QD_MAX_WEIGHT];
  qd_unrank(r, N, w, sub);
  unsigned long long s = 0ull;
  for (int i = 0; i < w; ++i) s ^= sq[sub[i]];
  return s;
}

// insert the syndromes of all w-subsets (bit m = anticommutes with the logical stays in the key)
__global__ void __launch_bounds__(256) k_dt_insert(const unsigned long long *__restrict__ sq, const int N, const int w,
                                                   const unsigned long long n_subsets, unsigned long long *table,
                                                   const unsigned long long cap_mask, unsigned int *overflow) {
  for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_subsets;
       r += (unsigned long long)gridDim.x * blockDim.x) {
    const unsigned long long key = qd_subset_syndrome(sq, r, N, w);
    unsigned long long h = qd_hash(key) & cap_mask;
    unsigned long long probes = 0;
    for (;; h = (h + 1) & cap_mask) {
      const unsigned long long old = atomicCAS(table + h, QD_EMPTY, key);
      if (old == QD_EMPTY || old == key) break;
      if (++probes > cap_mask) { atomicExch(overflow, 1u); break; }
    }
  }
}

// probe: a w-subset with syndrome s meets the set when s ^ top is in it (same stabilizer syndrome,
// one of the two commutes with the logical and the other does not: reference distance.cpp:150-157)
__global__ void __launch_bounds__(256) k_dt_probe(const unsigned long long *__restrict__ sq, const int N, const int w,
                                                  const unsigned long long n_subsets, const unsigned long long *table,
                                                  const unsigned long long cap_mask, const unsigned long long top,
                                                  unsigned int *flag) {
  for (unsigned long long r = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; r < n_subsets;
       r += (unsigned long long)gridDim.x * blockDim.x) {
    if (*reinterpret_cast<volatile unsigned int *>(flag)) return;
    const unsigned long long want = qd_subset_syndrome(sq, r, N, w) ^ top;
    for (unsigned long long h = qd_hash(want) & cap_mask;; h = (h + 1) & cap_mask) {
      const unsigned long long x = table[h];
      if (x == want) { atomicExch(flag, 1u); return; }
      if (x == QD_EMPTY) break;
    }
  }
}

// ---- host side ----------------------------------------------------------------------------------
struct Bits {  // rows packed 64 columns per word
  size_t cols = 0, words = 0;
  std::vector<std::vector<uint64_t>> r;
  Bits(const uint8_t *m, int rows, size_t cols_) : cols(cols_), words((cols_ + 63) / 64) {
    for (int i = 0; i < rows; ++i) {
      std::vector<uint64_t> w(std::max<size_t>(words, 1), 0);
      for (size_t k = 0; k < cols; ++k)
        if (m[(size_t)i * cols + k] & 1) w[k >> 6] |= 1ull << (k & 63);
      r.push_back(w);
    }
  }
  bool get(size_t i, size_t k) const { return (r[i][k >> 6] >> (k & 63)) & 1; }
};

// reduced row echelon form in place; returns the rank, pivot columns in `piv`
int rref(Bits &b, std::vector<size_t> *piv) {
  size_t i = 0;
  for (size_t k = 0; k < b.cols && i < b.r.size(); ++k) {
    size_t j = i;
    while (j < b.r.size() && !b.get(j, k)) ++j;
    if (j == b.r.size()) continue;
    std::swap(b.r[i], b.r[j]);
    for (size_t l = 0; l < b.r.size(); ++l)
      if (l != i && b.get(l, k))
        for (size_t w = 0; w < b.words; ++w) b.r[l][w] ^= b.r[i][w];
    if (piv) piv->push_back(k);
    ++i;
  }
  return (int)i;
}

int sip(const uint8_t *a, const uint8_t *b, int n) {  // symplectic inner product (linear.cpp:11-21)
  int s = 0;
  for (int i = 0; i < n; ++i) s ^= (a[i] & b[i + n]) ^ (a[i + n] & b[i]);
  return s & 1;
}

int set_err(char *err, size_t len, int code, const char *msg) {
  if (err && len) snprintf(err, len, "%s", msg);
  return code;
}

#define QD_CU(call)                                                                  \
  do {                                                                               \
    cudaError_t e__ = (call);                                                        \
    if (e__ != cudaSuccess) {                                                        \
      if (err && errlen) snprintf(err, errlen, "CUDA error %s at %s:%d (%s)",        \
                                  cudaGetErrorString(e__), __FILE__, __LINE__, #call); \
      cleanup();                                                                     \
      return QD_ERR_CUDA;                                                            \
    }                                                                                \
  } while (0)

std::vector<unsigned long long> binomials() {
  std::vector<unsigned long long> b((QD_MAX_ELEMS + 1) * (QD_MAX_WEIGHT + 1), 0ull);
  const unsigned long long cap = 0x7FFFFFFFFFFFFFFFull;
  for (int i = 0; i <= QD_MAX_ELEMS; ++i) {
    b[i * (QD_MAX_WEIGHT + 1)] = 1ull;
    for (int j = 1; j <= QD_MAX_WEIGHT && j <= i; ++j) {
      const unsigned long long x = b[(i - 1) * (QD_MAX_WEIGHT + 1) + j - 1];
      const unsigned long long y = j <= i - 1 ? b[(i - 1) * (QD_MAX_WEIGHT + 1) + j] : 0ull;
      b[i * (QD_MAX_WEIGHT + 1) + j] = (x > cap - y) ? cap : x + y;
    }
  }
  return b;
}

int upload_binomials(int device, char *err, size_t errlen) {
  static bool done[64] = {false};
  auto cleanup = []() {};
  if (device >= 0 && device < 64 && done[device]) return 0;
  const std::vector<unsigned long long> b = binomials();
  QD_CU(cudaMemcpyToSymbol(qd_binom, b.data(), b.size() * sizeof(unsigned long long)));
  if (device >= 0 && device < 64) done[device] = true;
  return 0;
}

int grid_for(unsigned long long items, int sms) {
  const unsigned long long blocks = (items + 255) / 256;
  return (int)std::min<unsigned long long>(std::max<unsigned long long>(blocks, 1), (unsigned long long)sms * 16);
}

}  // namespace

int qd_minimum_distance(int device, int sms, int n, int n_stab, const uint8_t *stab, int n_gauge,
                        const uint8_t *gauge, int max_weight, int *distance, char *err, size_t errlen) {
  unsigned long long *d_T = nullptr;
  unsigned int *d_flag = nullptr;
  auto cleanup = [&]() {
    if (d_T) cudaFree(d_T);
    if (d_flag) cudaFree(d_flag);
  };
  if (n < 1 || n_stab < 1 || n_gauge < 1 || !stab || !gauge || !distance)
    return set_err(err, errlen, QD_ERR_INVALID, "minimum_distance: empty input");
  // reference distance.cpp:20-21 (is_isotropic, linear.cpp:28-37)
  for (int i = 0; i < n_stab; ++i)
    for (int j = i + 1; j < n_stab; ++j)
      if (sip(stab + (size_t)i * 2 * n, stab + (size_t)j * 2 * n, n))
        return set_err(err, errlen, QD_ERR_LOGIC, "symplectic vectors must span isotropic space");
  Bits W(stab, n_stab, 2 * (size_t)n);
  const int rk = rref(W, nullptr);  // n - k - r
  Bits V(gauge, n_gauge, 2 * (size_t)n);
  std::vector<size_t> piv;
  const int rk2 = rref(V, &piv);    // n - k + r
  const int k = n - (rk + rk2) / 2;
  // bit layout of a packed word vector: [0, n_stab) syndrome against the stabilizer rows as given,
  // then one bit per NON-pivot column of the gauge basis (the residual vanishes on pivot columns)
  std::vector<int> col_bit(2 * (size_t)n, -1);
  {
    std::vector<uint8_t> is_piv(2 * (size_t)n, 0);
    for (size_t p : piv) is_piv[p] = 1;
    int nb = n_stab;
    for (size_t cidx = 0; cidx < 2 * (size_t)n; ++cidx)
      if (!is_piv[cidx]) col_bit[cidx] = nb++;
    if (nb > 64 * QD_MAX_WORDS || n > QD_MAX_ELEMS)
      return set_err(err, errlen, QD_ERR_LIMIT,
                     "minimum_distance: more than 256 syndrome + residual bits, or more than 256 qubits");
  }
  int total_bits = n_stab + 2 * n - rk2;
  const int NW = std::max(1, (total_bits + 63) / 64);
  // unit vector u_j (column j of e set): syndrome bit i = stab_i[j +- n]; residual = u_j ^ (j pivot of
  // basis row b ? V_b : 0), restricted to the non-pivot columns
  auto unit = [&](size_t j) {
    std::vector<uint64_t> v((size_t)NW, 0);
    const size_t partner = j < (size_t)n ? j + n : j - n;
    for (int i = 0; i < n_stab; ++i)
      if (stab[(size_t)i * 2 * n + partner] & 1) v[i >> 6] |= 1ull << (i & 63);
    if (col_bit[j] >= 0) v[col_bit[j] >> 6] |= 1ull << (col_bit[j] & 63);
    for (int b = 0; b < rk2; ++b)
      if (piv[b] == j)
        for (size_t cidx = 0; cidx < 2 * (size_t)n; ++cidx)
          if (col_bit[cidx] >= 0 && V.get(b, cidx)) v[col_bit[cidx] >> 6] ^= 1ull << (col_bit[cidx] & 63);
    return v;
  };
  std::vector<unsigned long long> T((size_t)3 * n * NW);
  for (int q = 0; q < n; ++q) {
    const std::vector<uint64_t> x = unit((size_t)q), z = unit((size_t)q + n);
    for (int kk = 0; kk < NW; ++kk) {
      T[((size_t)3 * q + 0) * NW + kk] = x[kk];           // pstate 0: X
      T[((size_t)3 * q + 1) * NW + kk] = z[kk];           // pstate 1: Z
      T[((size_t)3 * q + 2) * NW + kk] = x[kk] ^ z[kk];   // pstate 2: Y
    }
  }
  unsigned long long mask_s[4] = {0, 0, 0, 0};
  for (int i = 0; i < n_stab; ++i) mask_s[i >> 6] |= 1ull << (i & 63);

  QD_CU(cudaSetDevice(device));
  if (int rc = upload_binomials(device, err, errlen)) return rc;
  QD_CU(cudaMalloc(&d_T, T.size() * sizeof(unsigned long long)));
  QD_CU(cudaMalloc(&d_flag, sizeof(unsigned int)));
  QD_CU(cudaMemcpy(d_T, T.data(), T.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
  const std::vector<unsigned long long> binom = binomials();
  *distance = 0;
  const int wmax = std::min(max_weight, n);
  for (int w = 1; w <= wmax; ++w) {
    if (w > QD_MAX_WEIGHT) {
      cleanup();
      return set_err(err, errlen, QD_ERR_LIMIT, "minimum_distance: weights above 16 are not enumerated");
    }
    const unsigned long long n_subsets = binom[(size_t)n * (QD_MAX_WEIGHT + 1) + w];
    QD_CU(cudaMemset(d_flag, 0, sizeof(unsigned int)));
    const int grid = grid_for(n_subsets, sms);
#define QD_GO(NWV)                                                                                          \
  k_min_distance<NWV><<<grid, 256>>>(d_T, n, w, n_subsets, mask_s[0], mask_s[1], mask_s[2], mask_s[3], n_stab, \
                                     k == 0 ? 1 : 0, d_flag)
    if (NW == 1) QD_GO(1); else if (NW == 2) QD_GO(2); else if (NW == 3) QD_GO(3); else QD_GO(4);
#undef QD_GO
    QD_CU(cudaGetLastError());
    unsigned int hit = 0;
    QD_CU(cudaMemcpy(&hit, d_flag, sizeof hit, cudaMemcpyDeviceToHost));
    if (hit) { *distance = w; break; }
  }
  cleanup();
  return 0;
}

// one distance_test (reference distance.cpp:71-159); *found = 0 / 1
static int qd_distance_test(int sms, int n, int m, const uint8_t *stab, const uint8_t *logical, int weight,
                            unsigned long long *d_sq, unsigned long long *d_table, unsigned long long table_cap,
                            unsigned int *d_flags, const std::vector<unsigned long long> &binom, int *found,
                            char *err, size_t errlen) {
  auto cleanup = []() {};
  *found = 0;
  const int N = 3 * n;
  std::vector<unsigned long long> sq;
  for (int q = 0; q < n; ++q) {
    unsigned long long sx = 0, sz = 0;
    for (int i = 0; i < m; ++i) {
      if (stab[(size_t)i * 2 * n + n + q] == 1) sx |= 1ull << i;
      if (stab[(size_t)i * 2 * n + q] == 1) sz |= 1ull << i;
    }
    if (logical[n + q] == 1) sx |= 1ull << m;
    if (logical[q] == 1) sz |= 1ull << m;
    sq.push_back(sx); sq.push_back(sz); sq.push_back(sx ^ sz);
  }
  const int w1 = (weight + 1) / 2, w2 = weight / 2;
  if (w2 <= 0) return 0;  // (weight 1: the reference's second pair of sets is empty -- no meeting)
  if (w1 > N) return set_err(err, errlen, QD_ERR_INVALID, "0 <= k <= n");
  if (w1 > QD_MAX_WEIGHT) return set_err(err, errlen, QD_ERR_LIMIT, "distance_test: weights above 32 are not enumerated");
  const unsigned long long c1 = binom[(size_t)N * (QD_MAX_WEIGHT + 1) + w1];
  const unsigned long long c2 = binom[(size_t)N * (QD_MAX_WEIGHT + 1) + w2];
  if (c1 > table_cap / 2)
    return set_err(err, errlen, QD_ERR_LIMIT, "distance_test: the syndrome set does not fit the device hash table (2^27 entries)");
  // the smallest power-of-two table with load <= 1/2
  unsigned long long cap = 1024;
  while (cap < 2 * c1) cap <<= 1;
  QD_CU(cudaMemcpy(d_sq, sq.data(), sq.size() * sizeof(unsigned long long), cudaMemcpyHostToDevice));
  QD_CU(cudaMemset(d_table, 0xFF, cap * sizeof(unsigned long long)));
  QD_CU(cudaMemset(d_flags, 0, 2 * sizeof(unsigned int)));
  k_dt_insert<<<grid_for(c1, sms), 256>>>(d_sq, N, w1, c1, d_table, cap - 1, d_flags + 1);
  QD_CU(cudaGetLastError());
  // (w1 == w2: the reference copies the sets; the same subsets probe the table they filled)
  k_dt_probe<<<grid_for(c2, sms), 256>>>(d_sq, N, w2, c2, d_table, cap - 1, 1ull << m, d_flags);
  QD_CU(cudaGetLastError());
  unsigned int fl[2] = {0, 0};
  QD_CU(cudaMemcpy(fl, d_flags, sizeof fl, cudaMemcpyDeviceToHost));
  if (fl[1]) return set_err(err, errlen, QD_ERR_LIMIT, "distance_test: hash table overflow");
  *found = fl[0] ? 1 : 0;
  return 0;
}

int qd_minimum_distance_by_tests(int device, int sms, int n, int n_stab, const uint8_t *stab, int n_xl,
                                 const uint8_t *xl, int n_zl, const uint8_t *zl, int max_weight,
                                 int *distance, char *err, size_t errlen) {
  unsigned long long *d_sq = nullptr, *d_table = nullptr;
  unsigned int *d_flags = nullptr;
  auto cleanup = [&]() {
    if (d_sq) cudaFree(d_sq);
    if (d_table) cudaFree(d_table);
    if (d_flags) cudaFree(d_flags);
  };
  if (n < 1 || n_stab < 1 || !stab || !distance || n_xl < 0 || n_zl < 0)
    return set_err(err, errlen, QD_ERR_INVALID, "minimum_distance_by_tests: empty input");
  if (n_stab >= 63)  // (the reference asserts; a drop-in cannot abort the interpreter)
    return set_err(err, errlen, QD_ERR_LOGIC, "n-k must be < 63 for this version not to overflow.");
  if (3 * n > QD_MAX_ELEMS)
    return set_err(err, errlen, QD_ERR_LIMIT, "minimum_distance_by_tests: more than 85 qubits");
  const unsigned long long table_cap = 1ull << 28;  // 2 GiB of 64-bit keys at most
  QD_CU(cudaSetDevice(device));
  if (int rc = upload_binomials(device, err, errlen)) return rc;
  const std::vector<unsigned long long> binom = binomials();
  // size the table for the largest set any weight needs
  unsigned long long need = 1024;
  for (int w = 1; w <= max_weight; ++w) {
    const int w1 = (w + 1) / 2;
    if (w1 > QD_MAX_WEIGHT || w1 > 3 * n) break;
    const unsigned long long c1 = binom[(size_t)3 * n * (QD_MAX_WEIGHT + 1) + w1];
    if (c1 > table_cap / 2) break;  // (reported when that weight is actually reached)
    while (need < 2 * c1) need <<= 1;
  }
  QD_CU(cudaMalloc(&d_sq, (size_t)3 * n * sizeof(unsigned long long)));
  QD_CU(cudaMalloc(&d_table, need * sizeof(unsigned long long)));
  QD_CU(cudaMalloc(&d_flags, 2 * sizeof(unsigned int)));
  int best = max_weight + 1;
  for (int pass = 0; pass < 2; ++pass) {
    const uint8_t *ops = pass ? zl : xl;
    const int rows = pass ? n_zl : n_xl;
    for (int row = 0; row < rows; ++row)
      for (int w = 1; w <= max_weight; ++w) {
        int found = 0;
        const int rc = qd_distance_test(sms, n, n_stab, stab, ops + (size_t)row * 2 * n, w, d_sq, d_table, need,
                                        d_flags, binom, &found, err, errlen);
        if (rc) { cleanup(); return rc; }
        if (found) { best = std::min(best, w); break; }
      }
  }
  cleanup();
  *distance = best <= max_weight ? best : 0;
  return 0;
}
