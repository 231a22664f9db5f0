// bindings.cpp -- pybind11 module `_c_analysis`: the reference-facing names over the C ABI.
//
// Mirrors reference src/qiskit_qec/analysis/bindings/analysis_bindings.cpp:12-54 name for
// name (_CErrorPropagator, _CFaultEnumerator, _CFaultSampler, same constructor and method
// signatures, same return shapes, same exception types and messages), so the reference's
// Python drivers (analysis/extensions.py, cerrorpropagator.py, faultenumerator.py,
// faultsampler.py) work unmodified on top of it.  All propagation happens in libqecb200.so
// kernels; this file only marshals and keeps the small host-side state the reference API
// exposes (the interactive 1-shot frame, enumerator cursor).
#include <pybind11/numpy.h>
#include <pybind11/pybind11.h>
#include <pybind11/stl.h>

#include <algorithm>
#include <cstdint>
#include <map>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "qec_b200.h"

namespace py = pybind11;

namespace qecb200 {  // keeps these types distinct from the reference module's in one process

typedef std::tuple<long, std::vector<std::string>, std::vector<std::string>, std::vector<int>>
    FaultPath;  // reference faultenumerator.h:22-25

static void check(int rc) {
  if (rc == 0) return;
  const std::string msg = qec_last_error();
  if (rc == QEC_ERR_INVALID) throw std::invalid_argument(msg);  // -> ValueError
  throw std::logic_error(msg);                                  // -> RuntimeError
}

static uint8_t pauli_code(const std::string &p) {
  auto bits = [](char c) -> uint8_t { return c == 'x' ? 1 : c == 'z' ? 2 : c == 'y' ? 3 : 0; };
  uint8_t code = 0;
  if (p.size() > 0) code |= bits(p[0]);
  if (p.size() > 1) code |= (uint8_t)(bits(p[1]) << 2);
  return code;
}

struct CircuitHandle {
  qec_circuit *h = nullptr;
  CircuitHandle() = default;
  CircuitHandle(const CircuitHandle &) = delete;
  CircuitHandle &operator=(const CircuitHandle &) = delete;
  ~CircuitHandle() { reset(); }
  void reset() {
    if (h) qec_circuit_destroy(h);
    h = nullptr;
  }
  void create(int device, int nq, int nc, const std::vector<std::vector<int>> &rows) {
    reset();
    std::vector<int32_t> off(rows.size() + 1, 0), data;
    for (size_t i = 0; i < rows.size(); ++i) {
      off[i + 1] = off[i] + (int32_t)rows[i].size();
      data.insert(data.end(), rows[i].begin(), rows[i].end());
    }
    if (data.empty()) data.push_back(0);
    check(qec_circuit_create(device, nq, nc, (int)rows.size(), off.data(), data.data(), &h));
  }
};

static int default_device() {
  const char *e = std::getenv("QEC_B200_DEVICE");
  return e ? std::atoi(e) : 0;
}

// unpack row `p` of a bit-sliced buffer [rows][n_words] into ints
static std::vector<int> unpack_path(const std::vector<uint32_t> &bits, int rows, int64_t n_words,
                                    int64_t p) {
  std::vector<int> out(rows);
  const int64_t w = p >> 5;
  const int b = (int)(p & 31);
  for (int r = 0; r < rows; ++r) out[r] = (int)((bits[(size_t)r * n_words + w] >> b) & 1u);
  return out;
}

// ---------------------------------------------------------------------------------------------
// _CErrorPropagator  (reference errorpropagator.h:9-42, errorpropagator.cpp)
class ErrorPropagator {
 public:
  ErrorPropagator(int qreg_size, int creg_size)
      : nq_(qreg_size), nc_(creg_size), qa_(2 * (size_t)std::max(qreg_size, 0), 0),
        ca_((size_t)std::max(creg_size, 0), 0) {}

  void load_circuit(int qreg_size, int creg_size, std::vector<std::vector<int>> encoded_circ) {
    // same order as the reference (:106-111): sizes and circuit are stored before validation
    nq_ = qreg_size;
    nc_ = creg_size;
    qa_.assign(2 * (size_t)std::max(nq_, 0), 0);
    ca_.assign((size_t)std::max(nc_, 0), 0);
    circ_ = std::move(encoded_circ);
    loaded_ = false;
    py::gil_scoped_release nogil;
    handle_.create(default_device(), nq_, nc_, circ_);
    loaded_ = true;
  }

  void apply_error(std::vector<int> qubits, std::string paulis) {
    if (qubits.size() != paulis.length()) throw std::logic_error("qubits and pauli length disagree");
    for (size_t i = 0; i < qubits.size(); ++i) {
      const int q = rq(qubits[i]);
      if (paulis[i] == 'x' || paulis[i] == 'y') qa_[q] ^= 1;
      if (paulis[i] == 'y' || paulis[i] == 'z') qa_[q + nq_] ^= 1;
    }
  }
  // interactive gates on the 1-shot host frame (:211-254); indices are range checked here,
  // the reference leaves them undefined
  void cx(int qc, int qt) {
    rq(qc); rq(qt);
    qa_[qt] ^= qa_[qc];
    qa_[qc + nq_] ^= qa_[qt + nq_];
  }
  void h(int q) { rq(q); std::swap(qa_[q], qa_[q + nq_]); }
  void s(int q) { rq(q); if (qa_[q] == 1) qa_[q + nq_] ^= 1; }
  void reset(int q) { rq(q); qa_[q] = 0; qa_[q + nq_] = 0; }
  int measure(int q, int c) {
    rq(q);
    if (c < 0 || c >= nc_) throw std::logic_error("bit index out of range");
    ca_[c] = qa_[q];
    return ca_[c];
  }

  // one path through the K1 kernel; like the reference it leaves the final frame and clbits
  // in the object (:262-263, :312)
  std::vector<int> propagate(std::vector<int> icomb, std::vector<std::string> error) {
    std::vector<std::vector<int>> ic{std::move(icomb)};
    std::vector<std::vector<std::string>> er{std::move(error)};
    auto res = propagate_many(ic, er, true);
    return res[0];
  }

  std::vector<std::vector<int>> propagate_many(const std::vector<std::vector<int>> &icombs,
                                               const std::vector<std::vector<std::string>> &errors,
                                               bool keep_last_frame = false) {
    if (icombs.size() != errors.size()) throw std::logic_error("icomb and error lists disagree");
    ensure_handle();
    const int64_t n = (int64_t)icombs.size();
    std::vector<int64_t> off(n + 1, 0);
    std::vector<int32_t> ops;
    std::vector<uint8_t> codes;
    for (int64_t p = 0; p < n; ++p) {
      if (errors[p].size() < icombs[p].size())
        throw std::logic_error("fewer Pauli strings than failed operations");
      for (size_t j = 0; j < icombs[p].size(); ++j) {
        ops.push_back(icombs[p][j]);
        codes.push_back(pauli_code(errors[p][j]));
      }
      off[p + 1] = (int64_t)ops.size();
    }
    if (ops.empty()) { ops.push_back(0); codes.push_back(0); }
    const int64_t nw = (n + 31) / 32;
    std::vector<uint32_t> bits((size_t)std::max(nc_, 0) * nw + 1), frame((size_t)2 * nq_ * nw + 1);
    {
      py::gil_scoped_release nogil;
      check(qec_propagate_batch(handle_.h, n, off.data(), ops.data(), codes.data(), bits.data(),
                                keep_last_frame ? frame.data() : nullptr));
    }
    std::vector<std::vector<int>> out((size_t)n);
    for (int64_t p = 0; p < n; ++p) out[p] = unpack_path(bits, nc_, nw, p);
    if (keep_last_frame && n > 0) {
      qa_ = unpack_path(frame, 2 * nq_, nw, n - 1);
      ca_ = out[n - 1];
    }
    return out;
  }

  std::vector<int> get_cbits() const { return ca_; }
  std::vector<int> get_qubit_array() const { return qa_; }
  std::string get_qubits() const {
    std::string pauli;
    for (int i = 0; i < nq_; ++i) pauli.push_back("izxy"[qa_[i + nq_] + 2 * qa_[i]]);
    return pauli;
  }
  int get_qreg_size() const { return nq_; }
  int get_creg_size() const { return nc_; }
  size_t get_circuit_size() const { return circ_.size(); }

 private:
  int rq(int q) const {
    if (q < 0 || q >= nq_) throw std::logic_error("qubit index out of range");
    return q;
  }
  void ensure_handle() {
    if (!loaded_) {  // no circuit loaded: the reference propagates through an empty circuit
      handle_.create(default_device(), nq_, nc_, circ_);
      loaded_ = true;
    }
  }
  int nq_, nc_;
  std::vector<int> qa_, ca_;
  std::vector<std::vector<int>> circ_;
  CircuitHandle handle_;
  bool loaded_ = false;
};

// ---------------------------------------------------------------------------------------------
// _CFaultEnumerator  (reference faultenumerator.h:27-67, faultenumerator.cpp)
class FaultEnumerator {
 public:
  FaultEnumerator(int order, int qreg_size, int creg_size,
                  std::vector<std::vector<int>> &encoded_circ, std::vector<int> &faulty_ops_indices,
                  std::vector<std::string> &faulty_ops_labels,
                  std::vector<std::vector<std::string>> &faulty_ops_pauli_errors)
      : order_(order), nq_(qreg_size), nc_(creg_size), indices_(faulty_ops_indices),
        labels_(faulty_ops_labels), paulis_(faulty_ops_pauli_errors) {
    circuit_.create(default_device(), nq_, nc_, encoded_circ);  // validates like load_circuit
    const size_t L = indices_.size();
    if (labels_.size() < L || paulis_.size() < L)
      throw std::logic_error("faulty_ops_labels / faulty_ops_pauli_errors shorter than faulty_ops_indices");
    std::vector<int32_t> loc_op(indices_.begin(), indices_.end()), poff(L + 1, 0);
    std::vector<uint8_t> codes;
    for (size_t l = 0; l < L; ++l) {
      for (const auto &p : paulis_[l]) codes.push_back(pauli_code(p));
      poff[l + 1] = (int32_t)codes.size();
    }
    if (loc_op.empty()) loc_op.push_back(0);
    if (codes.empty()) codes.push_back(0);
    check(qec_enumerator_create(circuit_.h, order_, (int)L, loc_op.data(), poff.data(), codes.data(), &h_));
  }
  FaultEnumerator(const FaultEnumerator &) = delete;
  ~FaultEnumerator() { if (h_) qec_enumerator_destroy(h_); }

  void reset() { index_ = 0; state_.clear(); done_ = false; }
  long get_index() const { return index_; }
  std::vector<int> get_state() const { return state_; }
  bool done() const { return done_; }

  int64_t num_paths() const { int64_t n = 0; check(qec_enumerator_num_paths(h_, &n)); return n; }

  std::pair<std::vector<int>, std::vector<int>> unrank(int64_t index) const {
    std::vector<int32_t> l(std::max(order_, 1)), d(std::max(order_, 1));
    check(qec_enumerator_unrank(h_, index, l.data(), d.data()));
    return {std::vector<int>(l.begin(), l.begin() + order_), std::vector<int>(d.begin(), d.begin() + order_)};
  }

  // Blocks as the reference cuts them (faultenumerator.cpp:71-111): at least `blocksize`
  // paths, ending only after a whole subset's product; `done` is raised only by the call
  // that finds no further subset, so a stream ending exactly on a block boundary is followed
  // by one empty block.
  std::vector<FaultPath> enumerate(int blocksize) {
    std::vector<FaultPath> block;
    if (done_) return block;
    if (state_.empty()) { index_ = 0; done_ = false; }
    const int64_t total = num_paths();  // raises the Combinations / ProductIterator errors
    const int64_t start = index_;
    if (start >= total) { done_ = true; return block; }
    const int64_t target = start + std::max<int64_t>(blocksize, 1) - 1;
    int64_t end = total;
    std::vector<int32_t> locs(order_), digs(order_);
    if (target < total) {
      int64_t first = 0, count = 0;
      check(qec_enumerator_unrank(h_, target, locs.data(), digs.data()));
      check(qec_enumerator_subset_rank(h_, locs.data(), &first, &count));
      end = first + count;
    }
    const int64_t n = end - start, nw = (n + 31) / 32;
    std::vector<uint32_t> bits((size_t)nc_ * nw + 1);
    {
      py::gil_scoped_release nogil;
      check(qec_enumerator_range(h_, start, end, bits.data(), nullptr));
    }
    // labels / errors of each path: host successor walk from the unranked first path
    check(qec_enumerator_unrank(h_, start, locs.data(), digs.data()));
    const int L = (int)indices_.size();
    block.reserve((size_t)n);
    for (int64_t p = 0; p < n; ++p) {
      std::vector<std::string> lab(order_), err(order_);
      for (int j = 0; j < order_; ++j) {
        lab[j] = labels_[locs[j]];
        err[j] = paulis_[locs[j]][digs[j]];
      }
      block.emplace_back((long)(start + p), std::move(lab), std::move(err), unpack_path(bits, nc_, nw, p));
      if (p + 1 == n) break;
      int j = order_ - 1;
      for (; j >= 0; --j) {
        if (++digs[j] < (int)paulis_[locs[j]].size()) break;
        digs[j] = 0;
      }
      if (j < 0) {
        int i = order_ - 1;
        while (locs[i] >= L - order_ + i) --i;
        ++locs[i];
        for (int k = i + 1; k < order_; ++k) locs[k] = locs[i] + k - i;
      }
    }
    state_.assign(locs.begin(), locs.end());
    index_ = (long)end;
    if (target >= total) done_ = true;  // ran out of subsets inside this call
    return block;
  }

  // bulk API: packed outcomes uint32[nc][ceil((hi-lo)/32)] and tallies for paths [lo, hi)
  py::tuple enumerate_packed(int64_t lo, int64_t hi, bool want_tally) {
    const int64_t nw = (std::max<int64_t>(hi - lo, 0) + 31) / 32;
    py::array_t<uint32_t> bits({(py::ssize_t)nc_, (py::ssize_t)nw});
    py::array_t<uint64_t> tally((py::ssize_t)QEC_TALLY_LEN(nc_));
    {
      uint32_t *pb = bits.mutable_data();
      uint64_t *pt = want_tally ? tally.mutable_data() : nullptr;
      py::gil_scoped_release nogil;
      check(qec_enumerator_range(h_, lo, hi, pb, pt));
    }
    if (want_tally) return py::make_tuple(bits, tally);
    return py::make_tuple(bits, py::none());
  }

 private:
  int order_, nq_, nc_;
  std::vector<int> indices_;
  std::vector<std::string> labels_;
  std::vector<std::vector<std::string>> paulis_;
  CircuitHandle circuit_;
  qec_enumerator *h_ = nullptr;
  long index_ = 0;
  std::vector<int> state_;
  bool done_ = false;
};

// ---------------------------------------------------------------------------------------------
// _CFaultSampler  (reference faultsampler.h:26-66, faultsampler.cpp)
typedef std::vector<std::pair<std::string, double>> PauliWeight;
typedef std::map<std::string, PauliWeight> LabelToPauliWeightMap;

class FaultSampler {
 public:
  FaultSampler(int qreg_size, int creg_size, std::vector<std::vector<int>> &encoded_circ,
               std::vector<int> &faulty_ops_indices, std::vector<std::string> &faulty_ops_labels,
               LabelToPauliWeightMap &label_to_pauli_weight,
               std::map<std::string, double> &label_to_error_probability, uint64_t seed)
      : nq_(qreg_size), nc_(creg_size), indices_(faulty_ops_indices), labels_(faulty_ops_labels) {
    circuit_.create(default_device(), nq_, nc_, encoded_circ);
    const size_t L = indices_.size();
    if (labels_.size() < L) throw std::logic_error("faulty_ops_labels shorter than faulty_ops_indices");
    std::map<std::string, int> cls;
    std::vector<int32_t> loc_op(indices_.begin(), indices_.end()), loc_cls(L);
    for (size_t l = 0; l < L; ++l) {
      auto it = cls.find(labels_[l]);
      if (it == cls.end()) {
        it = cls.emplace(labels_[l], (int)class_names_.size()).first;
        class_names_.push_back(labels_[l]);
      }
      loc_cls[l] = it->second;
    }
    std::vector<double> p, w;
    std::vector<int32_t> off{0};
    std::vector<uint8_t> codes;
    for (const auto &name : class_names_) {
      // a label absent from the probability map never fails (the reference's
      // std::map::operator[] default-inserts 0.0, faultsampler.cpp:58)
      auto pit = label_to_error_probability.find(name);
      const double pc = pit == label_to_error_probability.end() ? 0.0 : pit->second;
      auto wit = label_to_pauli_weight.find(name);
      std::vector<std::string> strs;
      if (wit == label_to_pauli_weight.end() || wit->second.empty()) {
        if (pc > 0) throw std::invalid_argument("no Pauli weights for faulty label \"" + name + "\"");
        strs.push_back("i"); codes.push_back(0); w.push_back(1.0);
      } else {
        for (const auto &pw : wit->second) {
          strs.push_back(pw.first); codes.push_back(pauli_code(pw.first)); w.push_back(pw.second);
        }
      }
      class_paulis_.push_back(strs);
      p.push_back(pc);
      off.push_back((int32_t)codes.size());
    }
    if (loc_op.empty()) { loc_op.push_back(0); loc_cls.push_back(0); }
    if (p.empty()) p.push_back(0.0);
    if (codes.empty()) { codes.push_back(0); w.push_back(1.0); }
    check(qec_sampler_create(circuit_.h, (int)L, loc_op.data(), loc_cls.data(), (int)class_names_.size(),
                             p.data(), off.data(), codes.data(), w.data(), seed, &h_));
    loc_cls_.assign(loc_cls.begin(), loc_cls.begin() + L);
  }
  FaultSampler(const FaultSampler &) = delete;
  ~FaultSampler() { if (h_) qec_sampler_destroy(h_); }

  // blocksize fresh shots; successive calls continue with new global shot ids (the cursor
  // moves in whole 32-shot words), so blocks never repeat -- like the reference's RNG stream.
  // Returns the reference's list of (-1, labels, errors, outcome) tuples, built straight with the
  // CPython API: a 1e6-shot block is 1e6 tuples + 3e6 lists, and going through
  // std::vector<FaultPath> and pybind's casters built every one of them twice.
  py::list sample(int blocksize) {
    py::list block;
    if (blocksize <= 0) return block;
    const int64_t n = blocksize, nw = (n + 31) / 32;
    std::vector<uint32_t> bits((size_t)nc_ * nw + 1), cnt((size_t)nw), ev;
    int cap = cap_, need = 0;
    for (;;) {
      ev.assign((size_t)nw * cap, 0u);
      int rc;
      {
        py::gil_scoped_release nogil;
        rc = qec_sampler_sample(h_, cursor_, n, bits.data(), nullptr, cnt.data(), ev.data(), cap, &need);
      }
      if (rc == QEC_ERR_LIMIT && need > cap) { cap = cap_ = need + 8; continue; }  // same seed, same shots
      check(rc);
      break;
    }
    cursor_ += nw * 32;
    if (py_labels_.empty()) {  // Python strings of the labels / Paulis, made once
      for (const auto &l : labels_) py_labels_.push_back(py::str(l));
      for (const auto &cp : class_paulis_) {
        py_paulis_.emplace_back();
        for (const auto &ps : cp) py_paulis_.back().push_back(py::str(ps));
      }
    }
    PyObject *zero = PyLong_FromLong(0), *one = PyLong_FromLong(1), *minus = PyLong_FromLong(-1);
    PyObject *out = PyList_New((Py_ssize_t)n);
    // Host side of the drop-in call: one tuple of three lists per shot.  The packed outcomes are
    // bit-sliced [clbit][word]: 16 words at a time are gathered into word-major columns (one cache
    // line per clbit serves 16 words), and a shot's outcome list is filled with zeros first, then
    // ones only on the clbits that are set somewhere in its 32-shot word (about half of them at
    // p = 1e-3 on the d=5 circuit).
    constexpr int WB = 16;
    std::vector<uint32_t> col((size_t)WB * nc_), sorted;
    std::vector<int> live;
    for (int64_t wb0 = 0; wb0 < nw; wb0 += WB) {
      const int nwb = (int)std::min<int64_t>(WB, nw - wb0);
      for (int c = 0; c < nc_; ++c) {
        const uint32_t *src = bits.data() + (size_t)c * nw + wb0;
        for (int j = 0; j < nwb; ++j) col[(size_t)j * nc_ + c] = src[j];
      }
      for (int jw = 0; jw < nwb; ++jw) {
        const int64_t w = wb0 + jw;
        const uint32_t *cw = col.data() + (size_t)jw * nc_;
        live.clear();
        for (int c = 0; c < nc_; ++c) if (cw[c]) live.push_back(c);
        // the word's events by shot (stable: inside a shot they stay in location order)
        sorted.assign(ev.begin() + (size_t)w * cap, ev.begin() + (size_t)w * cap + cnt[w]);
        std::stable_sort(sorted.begin(), sorted.end(),
                         [](uint32_t a, uint32_t b) { return ((a >> 9) & 31u) < ((b >> 9) & 31u); });
        size_t k = 0;
        const int nv = (int)std::min<int64_t>(32, n - 32 * w);
        for (int bit = 0; bit < nv; ++bit) {
          size_t k1 = k;
          while (k1 < sorted.size() && ((sorted[k1] >> 9) & 31u) == (uint32_t)bit) ++k1;
          PyObject *labs = PyList_New((Py_ssize_t)(k1 - k)), *errs = PyList_New((Py_ssize_t)(k1 - k));
          for (size_t j = k; j < k1; ++j) {
            const uint32_t loc = sorted[j] >> 14, dig = sorted[j] & 511u;
            PyObject *l = py_labels_[loc].ptr(), *e = py_paulis_[loc_cls_[loc]][dig].ptr();
            Py_INCREF(l); Py_INCREF(e);
            PyList_SET_ITEM(labs, (Py_ssize_t)(j - k), l);
            PyList_SET_ITEM(errs, (Py_ssize_t)(j - k), e);
          }
          k = k1;
          PyObject *oc = PyList_New((Py_ssize_t)nc_);
          PyObject **items = reinterpret_cast<PyListObject *>(oc)->ob_item;
          for (int c = 0; c < nc_; ++c) items[c] = zero;
          // (one reference per slot, added in bulk; a no-op where small ints are immortal: 3.12 on)
          Py_SET_REFCNT(zero, Py_REFCNT(zero) + nc_);
          for (int c : live)
            if ((cw[c] >> bit) & 1u) { Py_DECREF(zero); Py_INCREF(one); items[c] = one; }
          PyObject *t = PyTuple_New(4);
          Py_INCREF(minus);
          PyTuple_SET_ITEM(t, 0, minus);
          PyTuple_SET_ITEM(t, 1, labs);
          PyTuple_SET_ITEM(t, 2, errs);
          PyTuple_SET_ITEM(t, 3, oc);
          PyList_SET_ITEM(out, (Py_ssize_t)(32 * w + bit), t);
        }
      }
    }
    Py_DECREF(zero); Py_DECREF(one); Py_DECREF(minus);
    return py::reinterpret_steal<py::list>(out);
  }

  // bulk API: packed outcomes + tallies of `n` fresh shots
  py::tuple sample_packed(int64_t n, bool want_bits, bool want_tally) {
    const int64_t nw = (std::max<int64_t>(n, 0) + 31) / 32;
    py::array_t<uint32_t> bits({(py::ssize_t)(want_bits ? nc_ : 0), (py::ssize_t)nw});
    py::array_t<uint64_t> tally((py::ssize_t)QEC_TALLY_LEN(nc_));
    {
      uint32_t *pb = want_bits ? bits.mutable_data() : nullptr;
      uint64_t *pt = want_tally ? tally.mutable_data() : nullptr;
      py::gil_scoped_release nogil;
      check(qec_sampler_sample(h_, cursor_, n, pb, pt, nullptr, nullptr, 0, nullptr));
    }
    cursor_ += nw * 32;
    return py::make_tuple(want_bits ? py::object(bits) : py::none(),
                          want_tally ? py::object(tally) : py::none());
  }
  // bulk API: {outcome: multiplicity} of `n` fresh shots, histogrammed on the device (K4);
  // returns (rows uint32[k][ceil(nc/32)], counts uint64[k], n_zero): the k distinct non-zero
  // outcomes, row-packed (bit c % 32 of word c / 32 = clbit c), and the all-zero multiplicity
  py::tuple sample_counts(int64_t n) {
    const int64_t nw = (std::max<int64_t>(n, 0) + 31) / 32;
    const int rw = (nc_ + 31) / 32;
    int64_t cap = 1 << 16;
    while (cap < 2 * std::min<int64_t>(std::max<int64_t>(n, 1), (int64_t)1 << 26)) cap <<= 1;
    int64_t room = std::min<int64_t>(std::max<int64_t>(n, 1), 1 << 16), n_out = 0;
    uint64_t n_zero = 0;
    std::vector<uint64_t> cnt;
    std::vector<uint32_t> rows;
    for (;;) {
      cnt.assign((size_t)room, 0);
      rows.assign((size_t)room * rw, 0);
      int rc;
      {
        py::gil_scoped_release nogil;
        rc = qec_sampler_counts(h_, cursor_, n, cap, room, cnt.data(), rows.data(), &n_out, &n_zero, nullptr);
      }
      if (rc == QEC_ERR_LIMIT && n_out > room) { room = std::max(n_out, 2 * room); continue; }  // same seed, same shots
      check(rc);
      break;
    }
    cursor_ += nw * 32;
    py::array_t<uint32_t> r({(py::ssize_t)n_out, (py::ssize_t)rw});
    py::array_t<uint64_t> c((py::ssize_t)n_out);
    if (n_out) {
      memcpy(r.mutable_data(), rows.data(), (size_t)n_out * rw * 4);
      memcpy(c.mutable_data(), cnt.data(), (size_t)n_out * 8);
    }
    return py::make_tuple(r, c, (uint64_t)n_zero);
  }
  // bulk API, compact form: (word_cnt uint16[ceil(n/32)], events uint32[n_events], tally or None)
  // of `n` fresh shots: the sampled faults (location << 14 | shot bit << 9 | digit), the words'
  // events back to back -- see qec_sampler_sample_events
  py::tuple sample_events(int64_t n, bool want_tally) {
    const int64_t nw = (std::max<int64_t>(n, 0) + 31) / 32;
    py::array_t<uint16_t> cnt((py::ssize_t)nw);
    py::array_t<uint64_t> tally((py::ssize_t)QEC_TALLY_LEN(nc_));
    std::vector<uint32_t> ev((size_t)std::max<int64_t>(1024, n + n / 4 + 4096));
    int64_t n_ev = 0;
    for (int attempt = 0;; ++attempt) {
      int rc;
      {
        uint16_t *pc = cnt.mutable_data();
        uint64_t *pt = want_tally ? tally.mutable_data() : nullptr;
        py::gil_scoped_release nogil;
        rc = qec_sampler_sample_events(h_, cursor_, n, pc, ev.data(), (int64_t)ev.size(), &n_ev, pt);
      }
      if (rc == QEC_ERR_OVERFLOW && attempt < 6) continue;  // same seed, same shots
      if (rc == QEC_ERR_LIMIT && n_ev > (int64_t)ev.size() && attempt < 6) {
        ev.resize((size_t)(2 * n_ev * std::max<int64_t>(1, nw * 32 / std::max<int64_t>(1, n)) + 4096));
        continue;
      }
      check(rc);
      break;
    }
    cursor_ += nw * 32;
    py::array_t<uint32_t> e((py::ssize_t)n_ev);
    if (n_ev) memcpy(e.mutable_data(), ev.data(), (size_t)n_ev * 4);
    return py::make_tuple(cnt, e, want_tally ? py::object(tally) : py::none());
  }
  std::string variant() const { return qec_sampler_variant(h_); }
  int64_t get_cursor() const { return cursor_; }
  void set_cursor(int64_t c) {
    if (c < 0 || (c & 31)) throw std::invalid_argument("cursor must be a non-negative multiple of 32");
    cursor_ = c;
  }

 private:
  int nq_, nc_;
  std::vector<int> indices_;
  std::vector<std::string> labels_, class_names_;
  std::vector<std::vector<std::string>> class_paulis_;
  std::vector<int> loc_cls_;
  CircuitHandle circuit_;
  qec_sampler *h_ = nullptr;
  int64_t cursor_ = 0;
  int cap_ = 64;
  std::vector<py::str> py_labels_;
  std::vector<std::vector<py::str>> py_paulis_;
};

// ---------------------------------------------------------------------------------------------
// Free functions the reference module also exports (analysis_bindings.cpp:56-61).  They are
// outside the accelerated path; these are small bit-packed host versions
// with the reference's results (reference intern/linear.cpp:11-83, intern/distance.cpp:17-202)
// so that analysis/extensions.py keeps every feature flag on.
typedef std::vector<std::vector<int>> IntMatrix;

struct BitRows {  // rows packed 64 columns per word
  size_t cols = 0, words = 0;
  std::vector<std::vector<uint64_t>> r;
  explicit BitRows(const IntMatrix &m) {
    cols = m.empty() ? 0 : m[0].size();
    words = (cols + 63) / 64;
    for (const auto &row : m) r.push_back(pack(row));
  }
  std::vector<uint64_t> pack(const std::vector<int> &row) const {
    std::vector<uint64_t> w(std::max<size_t>(words, 1), 0);
    for (size_t k = 0; k < row.size() && k < cols; ++k)
      if (row[k] & 1) w[k >> 6] |= 1ull << (k & 63);
    return w;
  }
};

// reduce to echelon form in place; returns (rank, pivot column of each basis row)
static int echelon(BitRows &b, std::vector<size_t> *pivots = nullptr) {
  size_t i = 0;
  for (size_t k = 0; k < b.cols && i < b.r.size(); ++k) {
    size_t j = i;
    while (j < b.r.size() && !((b.r[j][k >> 6] >> (k & 63)) & 1)) ++j;
    if (j == b.r.size()) continue;
    std::swap(b.r[i], b.r[j]);
    for (size_t l = 0; l < b.r.size(); ++l)
      if (l != i && ((b.r[l][k >> 6] >> (k & 63)) & 1))
        for (size_t w = 0; w < b.words; ++w) b.r[l][w] ^= b.r[i][w];
    if (pivots) pivots->push_back(k);
    ++i;
  }
  return (int)i;
}

static int gf2_rank(IntMatrix vectors) {
  if (vectors.empty()) throw std::logic_error("rank of an empty set of vectors");
  BitRows b(vectors);
  return echelon(b);
}

static inline int sip(const std::vector<int> &a, const std::vector<int> &b) {
  const size_t n = a.size() / 2;
  int s = 0;
  for (size_t i = 0; i < n; ++i) s ^= (a[i] & b[i + n]) ^ (a[i + n] & b[i]);
  return s & 1;
}

static bool gf2_isotropic(IntMatrix &v) {
  for (size_t i = 0; i < v.size(); ++i)
    for (size_t j = i + 1; j < v.size(); ++j)
      if (sip(v[i], v[j])) return false;
  return true;
}

// The brute-force distance functions run on the GPU (csrc/qec_distance.cu through the C ABI):
// reference analysis/intern/distance.cpp:17-53 and :172-202.  No host version: without a CUDA
// device they fail like every other compute entry point.
static std::vector<uint8_t> flatten(const IntMatrix &m, size_t cols, const char *what) {
  std::vector<uint8_t> out;
  out.reserve(m.size() * cols);
  for (const auto &row : m) {
    if (row.size() != cols) throw std::invalid_argument(std::string(what) + ": rows of unequal length");
    for (int v : row) out.push_back((uint8_t)(v & 1));
  }
  return out;
}

static int min_distance(IntMatrix &stab, IntMatrix &gauge, int max_weight) {
  if (stab.empty() || gauge.empty() || stab[0].empty() || (stab[0].size() & 1))
    throw std::invalid_argument("minimum_distance: empty or odd-length symplectic vectors");
  const size_t cols = stab[0].size();
  const std::vector<uint8_t> s = flatten(stab, cols, "symplectic_vectors"), g = flatten(gauge, cols, "symplectic_gauge_vectors");
  int d = 0;
  check(qec_minimum_distance(default_device(), (int)(cols / 2), (int)stab.size(), s.data(), (int)gauge.size(),
                             g.data(), max_weight, &d));
  return d;
}

static int min_distance_by_tests(IntMatrix &stab, IntMatrix &xl, IntMatrix &zl, int max_weight) {
  if (stab.empty() || stab[0].empty() || (stab[0].size() & 1))
    throw std::invalid_argument("minimum_distance_by_tests: empty or odd-length symplectic vectors");
  const size_t cols = stab[0].size();
  const std::vector<uint8_t> s = flatten(stab, cols, "symplectic_vectors"), x = flatten(xl, cols, "symplectic_xl"),
                             z = flatten(zl, cols, "symplectic_zl");
  int d = 0;
  check(qec_minimum_distance_by_tests(default_device(), (int)(cols / 2), (int)stab.size(), s.data(), (int)xl.size(),
                                      x.empty() ? nullptr : x.data(), (int)zl.size(), z.empty() ? nullptr : z.data(),
                                      max_weight, &d));
  return d;
}

}  // namespace qecb200
using namespace qecb200;

PYBIND11_MODULE(_c_analysis, module) {
  module.doc() = "qiskit-qec core analysis extensions -- B200-native (libqecb200.so) drop-in";
  module.attr("backend") = "qec_b200/sm_100a";

  py::class_<ErrorPropagator>(module, "_CErrorPropagator", py::module_local())
      .def(py::init<int, int>())
      .def("load_circuit", &ErrorPropagator::load_circuit)
      .def("apply_error", &ErrorPropagator::apply_error)
      .def("cx", &ErrorPropagator::cx)
      .def("h", &ErrorPropagator::h)
      .def("s", &ErrorPropagator::s)
      .def("reset", &ErrorPropagator::reset)
      .def("measure", &ErrorPropagator::measure)
      .def("propagate", &ErrorPropagator::propagate)
      .def("propagate_many", &ErrorPropagator::propagate_many, py::arg("icombs"), py::arg("errors"),
           py::arg("keep_last_frame") = false)
      .def("get_cbits", &ErrorPropagator::get_cbits)
      .def("get_qubits", &ErrorPropagator::get_qubits)
      .def("get_qubit_array", &ErrorPropagator::get_qubit_array)
      .def("get_qreg_size", &ErrorPropagator::get_qreg_size)
      .def("get_creg_size", &ErrorPropagator::get_creg_size)
      .def("get_circuit_size", &ErrorPropagator::get_circuit_size)
      .def("__repr__", [](const ErrorPropagator &a) {
        return std::string("<ErrorPropagator with ") + std::to_string(a.get_qreg_size()) +
               std::string(" qubits, ") + std::to_string(a.get_creg_size()) + std::string(" bits, ") +
               std::to_string(a.get_circuit_size()) + std::string(" instructions>");
      });

  py::class_<FaultEnumerator>(module, "_CFaultEnumerator", py::module_local())
      .def(py::init<int, int, int, std::vector<std::vector<int>> &, std::vector<int> &,
                    std::vector<std::string> &, std::vector<std::vector<std::string>> &>())
      .def("enumerate", &FaultEnumerator::enumerate)
      .def("reset", &FaultEnumerator::reset)
      .def("get_index", &FaultEnumerator::get_index)
      .def("get_state", &FaultEnumerator::get_state)
      .def("done", &FaultEnumerator::done)
      .def("num_paths", &FaultEnumerator::num_paths)
      .def("unrank", &FaultEnumerator::unrank)
      .def("enumerate_packed", &FaultEnumerator::enumerate_packed, py::arg("lo"), py::arg("hi"),
           py::arg("want_tally") = false);

  py::class_<FaultSampler>(module, "_CFaultSampler", py::module_local())
      .def(py::init<int, int, std::vector<std::vector<int>> &, std::vector<int> &,
                    std::vector<std::string> &, LabelToPauliWeightMap &,
                    std::map<std::string, double> &, uint64_t>())
      .def("sample", &FaultSampler::sample)
      .def("sample_packed", &FaultSampler::sample_packed, py::arg("n"), py::arg("want_bits") = true,
           py::arg("want_tally") = false)
      .def("sample_counts", &FaultSampler::sample_counts, py::arg("n"))
      .def("sample_events", &FaultSampler::sample_events, py::arg("n"), py::arg("want_tally") = false)
      .def("variant", &FaultSampler::variant)
      .def("get_cursor", &FaultSampler::get_cursor)
      .def("set_cursor", &FaultSampler::set_cursor);

  module.def("_c_minimum_distance", &min_distance,
             "compute minimum distance of subsystem stabilizer code", py::arg("symplectic_vectors"),
             py::arg("symplectic_gauge_vectors"), py::arg("max_weight") = 10);
  module.def("_c_minimum_distance_by_tests", &min_distance_by_tests,
             "compute minimum distance of subsystem stabilizer code", py::arg("symplectic_vectors"),
             py::arg("symplectic_xl"), py::arg("symplectic_zl"), py::arg("max_weight") = 10);
  module.def("_c_rank", &gf2_rank, "compute rank of set of vectors", py::arg("vectors"));
  module.def("_c_isotropic", &gf2_isotropic, "test if set of symplectic vectors is isotropic",
             py::arg("symplectic_vectors"));
}
