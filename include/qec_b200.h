/*
 * qec_b200.h -- C ABI of libqecb200.so: B200-native (sm_100a) Pauli-frame fault analysis.
 *
 * This is the drop-in boundary for the hot path of qiskit-qec's `analysis` package:
 * what the reference's pybind11 module `qiskit_qec.analysis._c_analysis`
 * (reference src/qiskit_qec/analysis/bindings/analysis_bindings.cpp:12-54) binds as
 * C++ classes is exposed here as opaque handles + plain pointers and sizes.
 * No torch / pybind / STL types appear in any signature.
 *
 * Conventions
 *   - every function returns 0 on success or a negative qec_status; the message is
 *     available from qec_last_error() (thread local).  The messages for conditions the
 *     reference also checks are the reference's own strings (see each function).
 *   - "host" pointers are ordinary (ideally pinned) host memory; "_dev" entry points
 *     take device pointers on the handle's device and a cudaStream_t passed as void*
 *     (NULL = the legacy default stream) and are asynchronous.
 *   - the library never retains a caller pointer past the call.
 *   - there is NO CPU fallback: without a usable CUDA device every compute entry
 *     point fails with QEC_ERR_CUDA.
 *
 * Packed outcome layout ("bit-sliced"): uint32 out[nc][n_words], n_words = ceil(n/32);
 *   bit b of out[c*n_words + w] is outcome bit c (clbit c flipped?) of path/shot
 *   lo + 32*w + b.  Bits past the end of the range are 0.
 * Tally layout: uint64 tally[QEC_TALLY_LEN(nc)]:
 *   [0..nc)        number of paths whose outcome bit c is 1 (per-clbit marginals)
 *   [nc]           number of paths with any outcome bit set
 *   [nc+1]         number of paths processed
 *   [nc+2..nc+66)  histogram of faults per shot (sampler only; last bin = ">=63")
 * Pauli code (4 bits) for a fault on an op with qubits (a, b):
 *   bit0: X on a, bit1: Z on a, bit2: X on b, bit3: Z on b  (b only for cx: a=control,
 *   b=target; reference errorpropagator.cpp:179-206).  'y' = X|Z, 'i'/anything else = 0.
 */
#ifndef QEC_B200_H
#define QEC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define QEC_TALLY_HIST_BINS 64
#define QEC_TALLY_LEN(nc) ((size_t)(nc) + 2 + QEC_TALLY_HIST_BINS)
#define QEC_MAX_ORDER 8
#define QEC_MAX_QUBITS 16384
#define QEC_MAX_CLBITS 16384

typedef enum qec_status {
  QEC_OK = 0,
  QEC_ERR_LOGIC = -1,    /* reference std::logic_error   -> Python RuntimeError */
  QEC_ERR_INVALID = -2,  /* reference std::invalid_argument -> Python ValueError */
  QEC_ERR_CUDA = -3,     /* CUDA failure / no device / extension unusable -> RuntimeError */
  QEC_ERR_LIMIT = -4,    /* size beyond what this library encodes -> RuntimeError */
  QEC_ERR_OVERFLOW = -5  /* internal: sampler event buffer overflow (retried internally) */
} qec_status;

typedef struct qec_circuit qec_circuit;       /* compiled circuit IR, resident on one GPU */
typedef struct qec_enumerator qec_enumerator; /* FaultEnumerator state on one GPU */
typedef struct qec_sampler qec_sampler;       /* FaultSampler state on one GPU */

/* ---- library ------------------------------------------------------------------------- */
const char *qec_last_error(void);
const char *qec_version(void);
/* number of visible CUDA devices (0 if none / driver missing; never fails) */
int qec_device_count(void);
/* Dependent-free INT32 logic-op (LOP3) throughput of `device` in lane-ops/s: the
 * integer-pipe roofline denominator of SURVEY.md 8(d).  Runs ~`ms` milliseconds. */
int qec_measure_int_peak(int device, double ms, double *lane_ops_per_s, double *sm_clock_mhz);

/* ---- circuit IR: replaces ErrorPropagator::load_circuit ------------------------------
 * (reference src/qiskit_qec/analysis/intern/errorpropagator.cpp:103-149)
 * rows are the reference's `encoded_circ` in CSR form: row i = row_data[row_off[i]..row_off[i+1]).
 * Validation and messages as the reference: "bad 1-qubit instruction", "bad measure
 * instruction", "bad cx instruction", "bad barrier instruction", "bad opcode",
 * "qubit index out of range", "bit index out of range" (all QEC_ERR_LOGIC).
 * The IR is lowered to a flat packed gate table and uploaded to `device` once. */
int qec_circuit_create(int device, int nq, int nc, int n_ops, const int32_t *row_off,
                       const int32_t *row_data, qec_circuit **out);
void qec_circuit_destroy(qec_circuit *c);
int qec_circuit_sizes(const qec_circuit *c, int *nq, int *nc, int *n_ops);
/* W_bit of SURVEY.md 8(d) for this circuit (2 n_cx + 2 n_h + n_s + 2 n_reset + n_measure) */
int64_t qec_circuit_bit_ops(const qec_circuit *c);

/* ---- K1: batched ErrorPropagator::propagate -------------------------------------------
 * (reference errorpropagator.cpp:259-313)  path p injects faults
 * (fault_op[j], fault_pauli[j]) for j in [path_off[p], path_off[p+1]).  Duplicate op
 * indices inside one path: the LAST one wins, as in the reference (:265-267).
 * fault_op out of [0, n_ops) is rejected (the reference leaves it undefined).
 * out_bits: host uint32[nc][ceil(n_paths/32)]; out_frame (may be NULL): host
 * uint32[2*nq][ceil(n_paths/32)], final X frame rows then Z frame rows. */
int qec_propagate_batch(qec_circuit *c, int64_t n_paths, const int64_t *path_off,
                        const int32_t *fault_op, const uint8_t *fault_pauli,
                        uint32_t *out_bits, uint32_t *out_frame);

/* ---- K2: FaultEnumerator ---------------------------------------------------------------
 * (reference src/qiskit_qec/analysis/intern/faultenumerator.cpp:6-24,71-111)
 * L fault locations; location l sits on op loc_op[l] and offers the Pauli codes
 * pauli_codes[pauli_off[l]..pauli_off[l+1]) in product-digit order.
 * Path order = the reference's: k-subsets of locations in lexicographic order, inside a
 * subset the mixed-radix product with the LAST location fastest
 * (combinations.cpp:18-32, productiterator.cpp:17-32); path index = rank in that order.
 * Errors as the reference's Combinations/ProductIterator constructors:
 * "n must be positive", "0 <= k <= n", "each element must be positive" (QEC_ERR_INVALID). */
int qec_enumerator_create(qec_circuit *c, int order, int L, const int32_t *loc_op,
                          const int32_t *pauli_off, const uint8_t *pauli_codes,
                          qec_enumerator **out);
void qec_enumerator_destroy(qec_enumerator *e);
/* total number of fault paths (fails with QEC_ERR_LIMIT if >= 2^62) */
int qec_enumerator_num_paths(const qec_enumerator *e, int64_t *n);
/* rank -> (locations[order], digits[order]); host-side closed form of what the kernel does */
int qec_enumerator_unrank(const qec_enumerator *e, int64_t index, int32_t *locs, int32_t *digits);
/* first path index of the subset `locs` and its number of paths */
int qec_enumerator_subset_rank(const qec_enumerator *e, const int32_t *locs, int64_t *first,
                               int64_t *count);
/* propagate paths [lo, hi): out_bits host uint32[nc][ceil((hi-lo)/32)] (may be NULL),
 * tally host uint64[QEC_TALLY_LEN(nc)] (may be NULL; overwritten, not accumulated) */
int qec_enumerator_range(qec_enumerator *e, int64_t lo, int64_t hi, uint32_t *out_bits,
                         uint64_t *tally);
/* same with device pointers, asynchronous on `stream`; d_tally is ACCUMULATED into
 * (zero it first) so it can be the send buffer of a following all-reduce */
int qec_enumerator_range_dev(qec_enumerator *e, int64_t lo, int64_t hi, uint32_t *d_out_bits,
                             uint64_t *d_tally, void *stream);

/* ---- K3: FaultSampler ------------------------------------------------------------------
 * (reference src/qiskit_qec/analysis/intern/faultsampler.cpp:6-32,49-77)
 * Location l belongs to class (label) loc_class[l] in [0, n_classes); class k fails with
 * probability class_p[k] per shot and then draws digit d with probability proportional to
 * class_weight[class_off[k] + d]; digit d injects class_codes[class_off[k] + d].
 * RNG: Philox4x32-10 keyed by `seed`; the counter is built from the global shot position (a
 * 32-shot word or a batch of shots) and a draw index, so that results never depend on the grid
 * shape, on how a shot range is cut into calls or on the GPU count: 1 GPU == N GPUs bit for bit.
 * They DO depend on the sampler's kernel variant and draw scheme (frame / linear lane-streams /
 * warp / cta; gap tables or computed gaps), which the library picks from the circuit and the
 * noise model (clbit count, equal Pauli weights, pmax) and which the QEC_B200_SAMPLER /
 * QEC_B200_LIN_* environment variables can force: the same (seed, shot id) gives different --
 * equally distributed -- faults under a different variant.  qec_sampler_variant() names the
 * variant in use so that a caller can record or pin it.
 * Every probability is realised to within 2^-31 absolute (32-bit fixed point tables). */
int qec_sampler_create(qec_circuit *c, int L, const int32_t *loc_op, const int32_t *loc_class,
                       int n_classes, const double *class_p, const int32_t *class_off,
                       const uint8_t *class_codes, const double *class_weight, uint64_t seed,
                       qec_sampler **out);
void qec_sampler_destroy(qec_sampler *s);
/* kernel variant and draw scheme this sampler uses (see above), e.g.
 * "linear lane-streams draw=2"; the string lives as long as the sampler */
const char *qec_sampler_variant(const qec_sampler *s);
/* shots [shot_lo, shot_lo + n_shots); shot_lo must be a multiple of 32.
 * out_bits host uint32[nc][ceil(n/32)] (may be NULL); tally host (may be NULL; overwritten);
 * fault export (all three NULL or all non-NULL): fault_cnt host uint32[ceil(n/32)] = events
 * of each 32-shot word, fault_ev host uint32[ceil(n/32)][fault_cap], event =
 * location << 14 | shot_bit << 9 | digit, sorted by op order; *fault_cap_needed receives the
 * largest count seen (if > fault_cap the lists are truncated and QEC_ERR_LIMIT is returned). */
int qec_sampler_sample(qec_sampler *s, int64_t shot_lo, int64_t n_shots, uint32_t *out_bits,
                       uint64_t *tally, uint32_t *fault_cnt, uint32_t *fault_ev, int fault_cap,
                       int *fault_cap_needed);
int qec_sampler_sample_dev(qec_sampler *s, int64_t shot_lo, int64_t n_shots,
                           uint32_t *d_out_bits, uint64_t *d_tally, void *stream);
/* Compact form of qec_sampler_sample: the sampled faults themselves instead of the dense packed
 * outcomes (what the reference's sample() returns per shot is (labels, errors, outcome), reference
 * faultsampler.cpp:64-74; labels and errors ARE this list, the outcome is the XOR of the faults'
 * single-fault outcomes).  word_cnt host uint16[ceil(n/32)]: events of each 32-shot word; events
 * host uint32[events_cap]: all events, the words' back to back in word order, event = location
 * << 14 | shot_bit << 9 | digit; *n_events = their number (QEC_ERR_LIMIT if > events_cap: *n_events
 * then says how many were needed so far).  tally as above (may be NULL).  About 3.2 bytes per
 * shot cross PCIe at p = 1e-3 on the d=5 circuit, against 18.1 for the dense form.  The order of a
 * word's events is the kernel's (grouped by lane, stream order inside); sort by (location,
 * shot_bit) for a canonical order. */
int qec_sampler_sample_events(qec_sampler *s, int64_t shot_lo, int64_t n_shots, uint16_t *word_cnt,
                              uint32_t *events, int64_t events_cap, int64_t *n_events,
                              uint64_t *tally);
/* The same with three bytes per event (little endian): location << 9 | shot_bit << 4 | digit --
 * 2.4 bytes per shot on the d=5 circuit.  events24 host uint8[3 * events_cap].  QEC_ERR_LIMIT for
 * more than 32768 locations or more than 16 Paulis on a location. */
int qec_sampler_sample_events24(qec_sampler *s, int64_t shot_lo, int64_t n_shots, uint16_t *word_cnt,
                                uint8_t *events24, int64_t events_cap, int64_t *n_events,
                                uint64_t *tally);
/* The same, delta-coded in two bytes per entry: 1.6 bytes per shot on the d=5 circuit.  A word's
 * events come in SLOT order (slot = sorted location * 32 + shot_bit; qec_sampler_tables carries the
 * sorted -> original location map); entry = delta << 4 | digit with delta = slot - previous slot
 * - 1 (previous = -1 at the start of each word), and 0xFFF0 = "4095 slots without an event".
 * word_cnt counts ENTRIES (events + skips).  Served by the lane-stream kernel only (its lanes
 * export in slot order): QEC_ERR_LIMIT otherwise, and for more than 16 Paulis on a location. */
int qec_sampler_sample_events16(qec_sampler *s, int64_t shot_lo, int64_t n_shots, uint16_t *word_cnt,
                                uint16_t *entries16, int64_t entries_cap, int64_t *n_entries,
                                uint64_t *tally);
/* asynchronous sample_dev calls cannot report an event-buffer overflow (probability < 1e-20
 * per word with the default sizing); this checks for one: QEC_ERR_OVERFLOW means the last
 * batch of sample_dev results is invalid and must be re-run (the buffer has been grown). */
int qec_sampler_check(qec_sampler *s);
/* the fixed-point tables the sampler realises (for an independent CPU restatement to
 * consume): writes up to `cap` uint32 and returns the count in *n.  Layout documented in
 * DESIGN.md "sampler tables". */
int qec_sampler_tables(const qec_sampler *s, uint32_t *buf, int cap, int *n);

/* ---- K4: full-outcome histogram over a packed outcome buffer --------------------------
 * (new: the reference returns per-shot tuples only; this is the device-side form of the counts
 * dictionary its consumers build, reference utils/stim_tools.py:234-283,
 * decoders/decoding_graph.py:165-209.)
 * Counts distinct outcomes of n_paths paths held bit-sliced on the device of `c` (row stride
 * n_words >= ceil(n_paths/32)).  Open-addressed table of `table_cap` (power of two) slots:
 *   d_keys   uint64[table_cap]  64-bit hash of the outcome (0 = empty slot; zero the table first)
 *   d_counts uint64[table_cap]  multiplicity (accumulated: several calls may share one table)
 *   d_rows   uint32[table_cap][ceil(nc/32)] (may be NULL) the outcome itself, bit c%32 of word
 *            c/32 = clbit c, written by the thread that claims the slot
 *   d_status uint64[3] (may be NULL, accumulated): slots claimed, paths dropped because the
 *            table was full, paths with the all-zero outcome (counted here, NOT in the table)
 * Two distinct outcomes with equal 64-bit hashes would share a slot (probability about
 * n_distinct^2 / 2^65; 3e-6 for 1e7 distinct outcomes). */
int qec_outcome_histogram_rows_dev(qec_circuit *c, const uint32_t *d_bits, int64_t n_paths,
                                   int64_t n_words, uint64_t *d_keys, uint64_t *d_counts,
                                   uint32_t *d_rows, int64_t table_cap, uint64_t *d_status,
                                   void *stream);
/* the same without rows / status, row stride ceil(n_paths/32); the all-zero outcome is then
 * not reported anywhere */
int qec_outcome_histogram_dev(qec_circuit *c, const uint32_t *d_bits, int64_t n_paths,
                              uint64_t *d_keys, uint64_t *d_counts, int64_t table_cap,
                              void *stream);
/* FaultSampler -> {outcome: multiplicity} with host buffers: shots [shot_lo, shot_lo + n_shots)
 * are drawn and histogrammed on the device chunk by chunk; only the distinct outcomes cross
 * PCIe.  out_counts uint64[max_out], out_rows uint32[max_out][ceil(nc/32)] (order arbitrary);
 * *n_out = distinct non-zero outcomes, *n_zero = shots with the all-zero outcome; tally as in
 * qec_sampler_sample (may be NULL).  QEC_ERR_LIMIT if the table (table_cap slots, power of two)
 * or the output arrays are too small: *n_out then holds the number found so far / needed.
 * max_out = 0 only counts (n_out, n_zero). */
int qec_sampler_counts(qec_sampler *s, int64_t shot_lo, int64_t n_shots, int64_t table_cap,
                       int64_t max_out, uint64_t *out_counts, uint32_t *out_rows, int64_t *n_out,
                       uint64_t *n_zero, uint64_t *tally);

/* ---- several GPUs, one process ---------------------------------------------------------
 * (new: the reference is single-threaded and single-device, reference faultsampler.cpp:52,
 * faultenumerator.cpp:82)  A multi handle owns one circuit + sampler /
 * enumerator on each of `devices[n_dev]`.  A call cuts its shot / path range into n_dev
 * contiguous pieces of whole 32-unit lane words; there is no data-path exchange, only the
 * tallies are summed.  Results are those of ONE GPU, bit for bit (the Philox counters are
 * functions of the global shot position; path indices are global).
 *   *_sample / *_range: host buffers as in the single-GPU calls (out_bits host
 *     uint32[nc][ceil(n/32)] or NULL, tally host or NULL), one host thread per device.
 *   *_tally_dev: tallies only, all devices launched asynchronously, then device devices[0]
 *     sums its peers' tallies read over NVLink peer mappings (no collective launch; falls back
 *     to per-device copies where no peer mapping exists); *elapsed_ms (may be NULL) = device
 *     time from before the first kernel to after the reduce.
 * Creation arguments: those of qec_circuit_create + qec_sampler_create / qec_enumerator_create. */
typedef struct qec_multi_sampler qec_multi_sampler;
typedef struct qec_multi_enumerator qec_multi_enumerator;
int qec_multi_sampler_create(int n_dev, const int *devices, int nq, int nc, int n_ops,
                             const int32_t *row_off, const int32_t *row_data, int L,
                             const int32_t *loc_op, const int32_t *loc_class, int n_classes,
                             const double *class_p, const int32_t *class_off,
                             const uint8_t *class_codes, const double *class_weight, uint64_t seed,
                             qec_multi_sampler **out);
void qec_multi_sampler_destroy(qec_multi_sampler *ms);
int qec_multi_sampler_sample(qec_multi_sampler *ms, int64_t shot_lo, int64_t n_shots,
                             uint32_t *out_bits, uint64_t *tally);
int qec_multi_sampler_tally_dev(qec_multi_sampler *ms, int64_t shot_lo, int64_t n_shots,
                                uint64_t *tally, float *elapsed_ms);
int qec_multi_enumerator_create(int n_dev, const int *devices, int nq, int nc, int n_ops,
                                const int32_t *row_off, const int32_t *row_data, int order, int L,
                                const int32_t *loc_op, const int32_t *pauli_off,
                                const uint8_t *pauli_codes, qec_multi_enumerator **out);
void qec_multi_enumerator_destroy(qec_multi_enumerator *me);
int qec_multi_enumerator_num_paths(const qec_multi_enumerator *me, int64_t *n);
int qec_multi_enumerator_range(qec_multi_enumerator *me, int64_t lo, int64_t hi, uint32_t *out_bits,
                               uint64_t *tally);
int qec_multi_enumerator_tally_dev(qec_multi_enumerator *me, int64_t lo, int64_t hi, uint64_t *tally,
                                   float *elapsed_ms);

/* ---- K4 without any allocation ---------------------------------------------------------
 * (no counterpart in the reference: added for callers
 * that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it.)
 * The outcome histogram of qec_outcome_histogram_rows_dev with its hash table in a fixed device
 * array inside the library instead of caller memory: this call allocates and frees nothing.
 * Limits: circuits of at most 256 clbits; at most 65536 distinct non-zero outcomes per call
 * (fewer in practice: open addressing slows as the table fills; outcomes that find no slot
 * are counted in status[1], never silently lost); at most 2^40 paths.  Beyond the clbit or path
 * limit the call fails with QEC_ERR_LIMIT.  The array (3 MiB) is held by every GPU that loads
 * the library until the process exits.
 * d_bits: packed outcomes on the device of `c` (row stride n_words >= ceil(n_paths/32)).
 * Outputs (device pointers, overwritten): d_out_counts uint64[max_out], d_out_rows
 * uint32[max_out][ceil(nc/32)] -- the distinct non-zero outcomes, order arbitrary;
 * d_out_status uint64[4] = {distinct non-zero outcomes found, paths dropped because the table
 * was full, paths with the all-zero outcome, entries written = min(found, max_out)}.
 * Asynchronous on `stream`.  ALL calls on one GPU share the array, whatever the circuit or
 * stream: calls that overlap on the GPU race.  Serialise them (one stream, or events).  The
 * first call on each device looks up the array's address; make it before any graph capture. */
int qec_outcome_counts_noalloc_dev(qec_circuit *c, const uint32_t *d_bits, int64_t n_paths,
                                   int64_t n_words, uint64_t *d_out_counts, uint32_t *d_out_rows,
                                   int64_t max_out, uint64_t *d_out_status, void *stream);

/* ---- minimum distance by brute force ----------------------------------
 * Replace reference analysis/intern/distance.cpp:17-53 minimum_distance and :172-202
 * minimum_distance_by_tests (with distance_test :71-159), bound in the reference as
 * _c_minimum_distance / _c_minimum_distance_by_tests (bindings/analysis_bindings.cpp:56-61).
 * Matrices are row-major uint8 [rows][2 n] in the reference's [X part | Z part] order.
 * *distance = the minimum distance, or 0 when it exceeds max_weight (the reference's contract).
 * Errors: "symplectic vectors must span isotropic space" (QEC_ERR_LOGIC, reference :20-21);
 * "n-k must be < 63 for this version not to overflow." (QEC_ERR_LOGIC: the reference asserts);
 * QEC_ERR_LIMIT beyond 256 qubits / 256 syndrome + residual bits / weight 16 (minimum_distance),
 * beyond 85 qubits or 2^27 syndromes per half (by_tests); QEC_ERR_CUDA without a device. */
int qec_minimum_distance(int device, int n, int n_stab, const uint8_t *stab, int n_gauge,
                         const uint8_t *gauge, int max_weight, int *distance);
int qec_minimum_distance_by_tests(int device, int n, int n_stab, const uint8_t *stab, int n_xl,
                                  const uint8_t *xl, int n_zl, const uint8_t *zl, int max_weight,
                                  int *distance);

/* ---- device memory helpers (so a ctypes/cgo caller needs no CUDA binding) ------------- */
int qec_dev_alloc(int device, size_t bytes, void **out);
int qec_dev_free(int device, void *p);
int qec_dev_memset(int device, void *p, int value, size_t bytes, void *stream);
int qec_dev_to_host(int device, void *host, const void *dev, size_t bytes, void *stream);
int qec_host_to_dev(int device, void *dev, const void *host, size_t bytes, void *stream);
int qec_host_alloc_pinned(size_t bytes, void **out);
int qec_host_free_pinned(void *p);
int qec_stream_sync(int device, void *stream);
/* elapsed device time (ms) of the last *_dev / host call on handle, via CUDA events */
int qec_last_kernel_ms(int device, float *ms);


/* ---- circuit-specialised kernels ("JIT") ------------------------------------------------
 * For circuits whose frame fits the register file the enumerator and the sampler run a
 * kernel generated for that circuit (every frame word a register) and compiled for sm_100a
 * with NVRTC at first use; otherwise, or with QEC_B200_JIT=0, the table-driven kernels run.
 * Both produce identical bits.  qec_last_kernel() names the kernel that served this
 * thread's latest enumerator/sampler launch.  qec_jit_compile_to_file() needs no device: it
 * writes the generated gate-stream source and the sm_100a cubin of one variant
 * (kind 1 = enumerator, 2 = sampler) so that the SASS can be inspected. */
const char *qec_last_kernel(void);
int qec_jit_compile_to_file(int nq, int nc, int n_ops, const int32_t *row_off,
                            const int32_t *row_data, int L, const int32_t *loc_op, int kind,
                            int order, int tally, int bits, int threads, int min_blocks,
                            int group, const char *cubin_path, const char *source_path);

/* The host step of the linear-response planners on caller-given single-fault signatures, for
 * tests and inspection; needs no device.  bits[nc][ceil(n_faults / 32)]: bit f of row c = fault f
 * flips clbit c (what qec_propagate_batch returns for single-fault paths).  pred[nc] (or NULL):
 * the previous final measurement of the same qubit, -1 for none; diff_basis != 0 lists the rows
 * c ^ pred(c) (the time-difference basis the kernels accumulate in) instead of the raw ones.
 * Writes list_off[n_faults + 1] and the clbits of every fault back to back (ascending inside a
 * list) into entries[entries_cap]; raw_total / diff_total (may be NULL) receive the entry counts
 * of the two bases, whose comparison decides which one a sampler uses.  expect_short selects the
 * one-pass extraction the planners use for the time-difference lists (same result). */
int qec_planner_signature_lists(const uint32_t *bits, int nc, int64_t n_faults,
                                const int32_t *pred, int diff_basis, int expect_short,
                                int64_t *list_off, uint16_t *entries, int64_t entries_cap,
                                int64_t *raw_total, int64_t *diff_total);

#ifdef __cplusplus
}
#endif
#endif /* QEC_B200_H */
