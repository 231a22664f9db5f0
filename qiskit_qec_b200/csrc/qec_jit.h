// qec_jit.h -- internal interface of the circuit-specialised kernel builder (qec_jit.cpp).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <string>
#include <vector>

struct QecJitSpec {
  int nq = 0, nc = 0;
  const std::vector<uint32_t> *packed = nullptr;  // lowered gate table (qec_pack_op words)
  std::vector<uint8_t> is_loc;                    // per table entry: may a fault sit on it?
  int kind = 1;                                   // 1 enumerator, 2 sampler
  int order = 1;
  bool tally = false, bits = true;
  int threads = 128, min_blocks = 3;
  int group = 8;  // ops per "any event in here?" block; 0 = test every location
  int segments = 1;     // the sampler draws the faults of one gate-stream segment at a time
  int sync_every = 96;
  // sampler only: table sizes become compile-time constants of the generated kernel
  unsigned K1 = 0, B = 0, L = 0, n_classes = 0, lgP = 1, n_slots = 0;  // a CTA barrier every so many ops (keeps its warps in one stretch of code)
};

struct QecJitKernel {
  void *module = nullptr;    // CUmodule
  void *function = nullptr;  // CUfunction
  int threads = 0, min_blocks = 0, regs = 0;
  size_t static_smem = 0;
  double compile_ms = 0;
};

// true when NVRTC and the driver API could be loaded (no GPU needed for this)
bool qec_jit_available(std::string *why);
// the CUDA C text of the gate stream (exposed for tests and for reading the generated code)
std::string qec_jit_gates_source(const QecJitSpec &spec);
// ops per segment (a multiple of 32) and number of segments the generator will emit
int qec_jit_segment_ops(const QecJitSpec &spec);
int qec_jit_segment_count(const QecJitSpec &spec);
// generate + compile to a cubin for sm_100a (works without a GPU)
int qec_jit_compile(const QecJitSpec &spec, std::vector<char> *cubin, std::string *log);
// load a cubin on the current device and look the kernel up
int qec_jit_load(const std::vector<char> &cubin, size_t dyn_smem, QecJitKernel *out,
                 std::string *err);
int qec_jit_launch(const QecJitKernel &k, int grid, size_t dyn_smem, cudaStream_t stream,
                   void **args, std::string *err);
void qec_jit_unload(QecJitKernel *k);
