// qec_jit.cpp -- circuit-specialised kernels: code generator + NVRTC + driver-API loader.
//
// The gate stream of one circuit becomes straight-line CUDA C in which every frame word is
// a named local (a register): `h` is a renaming, `cx` is two XORs, `reset` two constants and
// a final `measure` one store.  Fault injection is a compare of the next event's op index
// against the (constant) index of the location.  The hand-written surroundings (event
// cursors, unranking, Philox, kernel loops) are qec_jit_kernels.cuh.  NVRTC and libcuda are
// dlopen()ed, so libqecb200.so itself links against cudart only; when either is missing the
// callers stay on the table-driven kernels of qec_kernels.cu (still the GPU, never the CPU).
#include "qec_jit.h"

#include <cuda.h>
#include <dlfcn.h>
#include <nvrtc.h>

#include <sys/stat.h>
#include <unistd.h>

#include <chrono>
#include <cstring>
#include <map>
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <sstream>

#include "qec_device.cuh"
#include "qec_jit_embed.h"  // qec_embed_device_cuh[], qec_embed_jit_kernels_cuh[] (build step)

namespace {

struct Api {
  bool tried = false, ok = false;
  std::string why;
  // NVRTC
  nvrtcResult (*CreateProgram)(nvrtcProgram *, const char *, const char *, int,
                               const char *const *, const char *const *) = nullptr;
  nvrtcResult (*CompileProgram)(nvrtcProgram, int, const char *const *) = nullptr;
  nvrtcResult (*GetProgramLogSize)(nvrtcProgram, size_t *) = nullptr;
  nvrtcResult (*GetProgramLog)(nvrtcProgram, char *) = nullptr;
  nvrtcResult (*GetCUBINSize)(nvrtcProgram, size_t *) = nullptr;
  nvrtcResult (*GetCUBIN)(nvrtcProgram, char *) = nullptr;
  nvrtcResult (*DestroyProgram)(nvrtcProgram *) = nullptr;
  const char *(*GetErrorString)(nvrtcResult) = nullptr;
  // driver
  bool drv_tried = false, drv_ok = false;
  std::string drv_why;
  CUresult (*ModuleLoadData)(CUmodule *, const void *) = nullptr;
  CUresult (*ModuleUnload)(CUmodule) = nullptr;
  CUresult (*ModuleGetFunction)(CUfunction *, CUmodule, const char *) = nullptr;
  CUresult (*FuncSetAttribute)(CUfunction, CUfunction_attribute, int) = nullptr;
  CUresult (*FuncGetAttribute)(int *, CUfunction_attribute, CUfunction) = nullptr;
  CUresult (*LaunchKernel)(CUfunction, unsigned, unsigned, unsigned, unsigned, unsigned, unsigned,
                           unsigned, CUstream, void **, void **) = nullptr;
  CUresult (*GetErrorStringDrv)(CUresult, const char **) = nullptr;
};

Api g_api;
std::mutex g_mu;

template <class F>
bool sym(void *h, const char *name, F *out) {
  *out = reinterpret_cast<F>(dlsym(h, name));
  return *out != nullptr;
}

void load_nvrtc() {
  if (g_api.tried) return;
  g_api.tried = true;
  const char *names[] = {"libnvrtc.so.12", "/usr/local/cuda/lib64/libnvrtc.so.12",
                         "/usr/local/cuda/lib64/libnvrtc.so", "libnvrtc.so"};
  void *h = nullptr;
  for (const char *n : names) {
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) {
    g_api.why = "libnvrtc.so.12 not found";
    return;
  }
  bool ok = sym(h, "nvrtcCreateProgram", &g_api.CreateProgram) &&
            sym(h, "nvrtcCompileProgram", &g_api.CompileProgram) &&
            sym(h, "nvrtcGetProgramLogSize", &g_api.GetProgramLogSize) &&
            sym(h, "nvrtcGetProgramLog", &g_api.GetProgramLog) &&
            sym(h, "nvrtcGetCUBINSize", &g_api.GetCUBINSize) &&
            sym(h, "nvrtcGetCUBIN", &g_api.GetCUBIN) &&
            sym(h, "nvrtcDestroyProgram", &g_api.DestroyProgram) &&
            sym(h, "nvrtcGetErrorString", &g_api.GetErrorString);
  if (!ok) g_api.why = "libnvrtc lacks an expected symbol";
  g_api.ok = ok;
}

void load_driver() {
  if (g_api.drv_tried) return;
  g_api.drv_tried = true;
  void *h = dlopen("libcuda.so.1", RTLD_NOW | RTLD_GLOBAL);
  if (!h) {
    g_api.drv_why = "libcuda.so.1 not found";
    return;
  }
  bool ok = sym(h, "cuModuleLoadData", &g_api.ModuleLoadData) &&
            sym(h, "cuModuleUnload", &g_api.ModuleUnload) &&
            sym(h, "cuModuleGetFunction", &g_api.ModuleGetFunction) &&
            sym(h, "cuFuncSetAttribute", &g_api.FuncSetAttribute) &&
            sym(h, "cuFuncGetAttribute", &g_api.FuncGetAttribute) &&
            sym(h, "cuLaunchKernel", &g_api.LaunchKernel) &&
            sym(h, "cuGetErrorString", &g_api.GetErrorStringDrv);
  if (!ok) g_api.drv_why = "libcuda lacks an expected symbol";
  g_api.drv_ok = ok;
}

std::string drv_err(CUresult r) {
  const char *s = nullptr;
  if (g_api.GetErrorStringDrv) g_api.GetErrorStringDrv(r, &s);
  return s ? s : "unknown driver error";
}

}  // namespace

bool qec_jit_available(std::string *why) {
  std::lock_guard<std::mutex> lk(g_mu);
  load_nvrtc();
  if (!g_api.ok && why) *why = g_api.why;
  return g_api.ok;
}

// ---- code generator ----------------------------------------------------------------------------
namespace {

struct Namer {  // frame word -> register name; h() renames instead of moving
  std::vector<int> x, z;
  explicit Namer(int nq) : x(nq), z(nq) {
    for (int q = 0; q < nq; ++q) { x[q] = q; z[q] = nq + q; }
  }
};

inline void reg(std::ostringstream &o, int id) { o << "F.f" << id; }

// emit ops [s, e) of the table; with_faults = also the injection tests
void emit_range(std::ostringstream &o, const QecJitSpec &spec, Namer &nm, size_t s, size_t e,
                bool with_faults, const char *ind) {
  const std::vector<uint32_t> &tab = *spec.packed;
  for (size_t i = s; i < e; ++i) {
    const uint32_t g = tab[i], opc = g & 15u, a = (g >> 4) & 0x3FFFu, b = g >> 18;
    const bool loc = with_faults && i < spec.is_loc.size() && spec.is_loc[i];
    // bit (i & 31) of the bitmap word loaded by cur.word(i >> 5): the lanes that hit pop their
    // queue entry into the named registers (ptxas predicates these few instructions)
    auto inj1 = [&]() {
      if (!loc) return;
      o << ind << "if (cur.hit(0x" << std::hex << (1u << (i & 31)) << std::dec << "u)) cur.inj1(";
      reg(o, nm.x[a]); o << ", "; reg(o, nm.z[a]); o << ");\n";
    };
    switch (opc) {
      case QOP_H:
        std::swap(nm.x[a], nm.z[a]);
        inj1();
        break;
      case QOP_S:
        o << ind; reg(o, nm.z[a]); o << " ^= "; reg(o, nm.x[a]); o << ";\n";
        inj1();
        break;
      case QOP_CX:
        o << ind; reg(o, nm.x[b]); o << " ^= "; reg(o, nm.x[a]); o << "; ";
        reg(o, nm.z[a]); o << " ^= "; reg(o, nm.z[b]); o << ";\n";
        if (loc) {
          o << ind << "if (cur.hit(0x" << std::hex << (1u << (i & 31)) << std::dec << "u)) cur.inj2(";
          reg(o, nm.x[a]); o << ", "; reg(o, nm.z[a]); o << ", ";
          reg(o, nm.x[b]); o << ", "; reg(o, nm.z[b]); o << ");\n";
        }
        break;
      case QOP_RESET:
        o << ind; reg(o, nm.x[a]); o << " = 0u; "; reg(o, nm.z[a]); o << " = 0u;\n";
        inj1();
        break;
      case QOP_MEAS:
        inj1();  // fault BEFORE the measurement; the value itself is overwritten later
        break;
      case QOP_MEASF:
        inj1();
        o << ind << "qec_jit_out<TALLY, BITS>(" << b << "u, "; reg(o, nm.x[a]);
        o << ", valid, cur.lane0, ob, os, s_tally, any_out);\n";
        break;
      case QOP_ZEROC:
        o << ind << "if (BITS) { if (valid) ob[" << a << "ll * os] = 0u; }\n";
        break;
      default:  // x y z id (b == 0): only a fault; barrier / padding (b == 1): nothing
        if (b == 0u) inj1();
        break;
    }
  }
}

}  // namespace

// ops per gate-stream segment (a multiple of 32, so that segments own whole bitmap words)
static size_t seg_len(const QecJitSpec &spec) {
  const size_t n = spec.packed->size();
  const size_t S = spec.segments > 1 ? (size_t)spec.segments : 1;
  size_t len = ((n + S - 1) / S + 31) / 32 * 32;
  return len ? len : 32;
}

int qec_jit_segment_ops(const QecJitSpec &spec) { return (int)seg_len(spec); }
int qec_jit_segment_count(const QecJitSpec &spec) {
  const size_t n = spec.packed->size(), len = seg_len(spec);
  const size_t S = (n + len - 1) / len;
  return S ? (int)S : 1;
}

std::string qec_jit_gates_source(const QecJitSpec &spec) {
  const std::vector<uint32_t> &tab = *spec.packed;
  std::ostringstream o;
  const size_t n = tab.size(), SL = seg_len(spec);
  const int NSEG = qec_jit_segment_count(spec);
  o << "// generated: gate stream of a " << spec.nq << "-qubit, " << spec.nc << "-clbit circuit, "
    << tab.size() << " table entries, " << NSEG << " segment(s) of " << SL << " ops\n";
  o << "#define QEC_JIT_NSEG " << NSEG << "\n#define QEC_JIT_SEGW " << SL / 32 << "\n";
  o << "struct QecFrame {\n";
  for (int r = 0; r < 2 * spec.nq; ++r)
    o << (r % 8 == 0 ? "  uint32_t " : " ") << 'f' << r << ((r % 8 == 7 || r == 2 * spec.nq - 1) ? ";\n" : ",");
  o << "};\n__device__ __forceinline__ void qec_jit_frame_zero(QecFrame &F) {\n";
  for (int r = 0; r < 2 * spec.nq; ++r)
    o << (r % 8 == 0 ? "  " : " ") << "F.f" << r << " = 0u;" << ((r % 8 == 7 || r == 2 * spec.nq - 1) ? "\n" : "");
  o << "}\n";
  Namer nm(spec.nq);
  // a block never straddles a bitmap word: its size is a power of two <= 32
  size_t G = 0;
  if (spec.group > 0) {
    G = 1;
    while (2 * G <= (size_t)spec.group && 2 * G <= 32) G *= 2;
    if (G < 2) G = 0;
  }
  const size_t step = G ? G : 32;
  for (int seg = 0; seg < NSEG; ++seg) {
    const size_t s0 = (size_t)seg * SL, s1 = std::min(n, s0 + SL);
    o << "template <class CUR, bool TALLY, bool BITS>\n"
         "__device__ __forceinline__ void qec_jit_seg" << seg << "(QecFrame &F, CUR &cur, const bool valid,\n"
         "    uint32_t *__restrict__ ob, const long long os, uint32_t *s_tally, uint32_t &any_out) {\n";
    size_t since_sync = 0;
    std::vector<std::string> slow_blocks;
    for (size_t s = s0; s < s1; s += step) {
      const size_t e = std::min(s1, s + step);
      uint32_t mask = 0;
      for (size_t i = s; i < e; ++i)
        if (i < spec.is_loc.size() && spec.is_loc[i]) mask |= 1u << (i & 31);
      // keep the warps of a CTA within one stretch of the code (instruction-cache locality)
      if (spec.sync_every > 0 && since_sync >= (size_t)spec.sync_every) {
        o << "  __syncthreads();\n";
        since_sync = 0;
      }
      since_sync += e - s;
      if ((s & 31) == 0) {
        bool word_has_loc = false;
        for (size_t i = s; i < std::min(s1, s + 32); ++i)
          word_has_loc |= (i < spec.is_loc.size() && spec.is_loc[i]);
        if (word_has_loc) o << "  cur.word(" << ((s - s0) >> 5) << ");\n";
      }
      if (!G || !mask) {
        emit_range(o, spec, nm, s, e, mask != 0, "  ");
        continue;
      }
      // one warp vote per block: does any lane word have an event on ops [s, e)?  The copy
      // with the injection tests is placed behind the end of the function (goto), so that the
      // common path is one contiguous run of code that the instruction prefetcher can follow.
      const size_t id = slow_blocks.size();
      o << "  if (__any_sync(0xFFFFFFFFu, cur.any(0x" << std::hex << mask << std::dec
        << "u))) goto qslow_" << seg << "_" << id << ";\n  {\n";
      Namer fast = nm;
      emit_range(o, spec, fast, s, e, false, "    ");
      o << "  }\n  qback_" << seg << "_" << id << ":;\n";
      std::ostringstream sl;
      sl << "  qslow_" << seg << "_" << id << ": {\n";
      emit_range(sl, spec, nm, s, e, true, "    ");
      sl << "  }\n  goto qback_" << seg << "_" << id << ";\n";
      slow_blocks.push_back(sl.str());
    }
    o << "  return;\n";
    for (const std::string &b : slow_blocks) o << b;
    o << "}\n";
  }
  o << "template <class CUR, bool TALLY, bool BITS>\n"
       "__device__ __forceinline__ void qec_jit_seg(const int seg, QecFrame &F, CUR &cur, const bool valid,\n"
       "    uint32_t *__restrict__ ob, const long long os, uint32_t *s_tally, uint32_t &any_out) {\n"
       "  switch (seg) {\n";
  for (int seg = 0; seg < NSEG; ++seg)
    o << "    case " << seg << ": qec_jit_seg" << seg
      << "<CUR, TALLY, BITS>(F, cur, valid, ob, os, s_tally, any_out); break;\n";
  o << "  }\n}\n";
  return o.str();
}


// ---- NVRTC ---------------------------------------------------------------------------------------
namespace {
std::map<uint64_t, std::vector<char>> g_cache;

std::string cache_dir() {
  std::string dir;
  if (const char *e = std::getenv("QEC_B200_JIT_CACHE")) {
    if (!*e || !strcmp(e, "0")) return "";
    dir = e;
  } else {
    Dl_info info;
    if (dladdr((void *)&qec_jit_available, &info) && info.dli_fname) {
      dir = info.dli_fname;
      const size_t slash = dir.rfind('/');
      dir = (slash == std::string::npos ? std::string(".") : dir.substr(0, slash)) + "/jit_cache";
    } else {
      return "";
    }
  }
  mkdir(dir.c_str(), 0755);
  return access(dir.c_str(), W_OK) == 0 ? dir : "";
}
}  // namespace

int qec_jit_compile(const QecJitSpec &spec, std::vector<char> *cubin, std::string *log) {
  {
    std::lock_guard<std::mutex> lk(g_mu);
    load_nvrtc();
  }
  if (!g_api.ok) {
    if (log) *log = g_api.why;
    return -1;
  }
  const auto t0 = std::chrono::steady_clock::now();
  std::ostringstream main_src;
  main_src << "#define QEC_JIT_T " << spec.threads << "\n"
           << "#define QEC_JIT_MINB " << spec.min_blocks << "\n"
           << "#define QEC_JIT_KIND " << spec.kind << "\n"
           << "#define QEC_JIT_ORDER " << spec.order << "\n"
           << "#define QEC_JIT_TALLY " << (spec.tally ? 1 : 0) << "\n"
           << "#define QEC_JIT_BITS " << (spec.bits ? 1 : 0) << "\n"
           << "#define QEC_JIT_NC " << spec.nc << "\n"
           << "#define QEC_JIT_NOPS " << spec.packed->size() << "\n"
           << "#define QEC_JIT_K1 " << spec.K1 << "u\n#define QEC_JIT_B " << spec.B << "u\n"
           << "#define QEC_JIT_L " << spec.L << "u\n#define QEC_JIT_NCLS " << spec.n_classes << "u\n"
           << "#define QEC_JIT_LGP " << spec.lgP << "u\n#define QEC_JIT_NSLOTS " << spec.n_slots << "\n"
           << "#include \"qec_device.cuh\"\n"
           << "#include \"qec_jit_kernels.cuh\"\n";
  const std::string src = main_src.str();
  const std::string gates = qec_jit_gates_source(spec);
  const bool lineinfo = []() { const char *e = std::getenv("QEC_B200_JIT_LINEINFO"); return e && e[0] == '1'; }();
  // cache: this process first, then <library dir>/jit_cache (or QEC_B200_JIT_CACHE), keyed by
  // a hash of everything that goes into the compile
  uint64_t key = 0xcbf29ce484222325ull;
  auto mix = [&](const char *p, size_t n) {
    for (size_t i = 0; i < n; ++i) { key ^= (unsigned char)p[i]; key *= 0x100000001b3ull; }
  };
  mix(src.data(), src.size());
  mix(gates.data(), gates.size());
  mix(qec_embed_device_cuh, sizeof qec_embed_device_cuh);
  mix(qec_embed_jit_kernels_cuh, sizeof qec_embed_jit_kernels_cuh);
  mix(lineinfo ? "L" : "-", 1);
  mix("nvrtc-12.9-sm_100a", 18);
  char keyname[40];
  snprintf(keyname, sizeof keyname, "%016llx.cubin", (unsigned long long)key);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_cache.find(key);
    if (it != g_cache.end()) { *cubin = it->second; return 0; }
  }
  const std::string dir = cache_dir();
  if (!dir.empty()) {
    FILE *f = fopen((dir + "/" + keyname).c_str(), "rb");
    if (f) {
      fseek(f, 0, SEEK_END);
      const long n = ftell(f);
      fseek(f, 0, SEEK_SET);
      cubin->resize(n > 0 ? (size_t)n : 0);
      const bool okr = n > 0 && fread(cubin->data(), 1, (size_t)n, f) == (size_t)n;
      fclose(f);
      if (okr) {
        std::lock_guard<std::mutex> lk(g_mu);
        g_cache[key] = *cubin;
        return 0;
      }
    }
  }
  const char *hdr_src[] = {qec_embed_device_cuh, qec_embed_jit_kernels_cuh, gates.c_str()};
  const char *hdr_name[] = {"qec_device.cuh", "qec_jit_kernels.cuh", "qec_jit_gates.inc"};
  nvrtcProgram prog;
  nvrtcResult r = g_api.CreateProgram(&prog, src.c_str(), "qec_jit_program.cu", 3, hdr_src, hdr_name);
  if (r != NVRTC_SUCCESS) {
    if (log) *log = std::string("nvrtcCreateProgram: ") + g_api.GetErrorString(r);
    return -1;
  }
  std::vector<const char *> opts = {"--gpu-architecture=sm_100a", "-std=c++17", "-default-device"};
  if (lineinfo) opts.push_back("-lineinfo");
  r = g_api.CompileProgram(prog, (int)opts.size(), opts.data());
  size_t ls = 0;
  g_api.GetProgramLogSize(prog, &ls);
  std::string lg(ls, '\0');
  if (ls > 1) g_api.GetProgramLog(prog, &lg[0]);
  if (log) *log = lg;
  if (r != NVRTC_SUCCESS) {
    if (log) *log = std::string("nvrtcCompileProgram: ") + g_api.GetErrorString(r) + "\n" + lg;
    g_api.DestroyProgram(&prog);
    return -1;
  }
  size_t cs = 0;
  g_api.GetCUBINSize(prog, &cs);
  cubin->resize(cs);
  g_api.GetCUBIN(prog, cubin->data());
  g_api.DestroyProgram(&prog);
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_cache[key] = *cubin;
  }
  if (!dir.empty()) {  // write-then-rename so a concurrent rank never reads half a file
    const std::string tmp = dir + "/" + keyname + ".tmp" + std::to_string((long)getpid());
    FILE *f = fopen(tmp.c_str(), "wb");
    if (f) {
      const bool okw = fwrite(cubin->data(), 1, cubin->size(), f) == cubin->size();
      fclose(f);
      if (okw) rename(tmp.c_str(), (dir + "/" + keyname).c_str());
      else remove(tmp.c_str());
    }
  }
  if (std::getenv("QEC_B200_JIT_VERBOSE")) {
    const double ms =
        std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    fprintf(stderr, "[qec_b200 jit] kind %d order %d tally %d bits %d T %d: %zu B cubin, %.0f ms\n",
            spec.kind, spec.order, (int)spec.tally, (int)spec.bits, spec.threads, cs, ms);
  }
  return 0;
}

// ---- driver API -------------------------------------------------------------------------------
int qec_jit_load(const std::vector<char> &cubin, size_t dyn_smem, QecJitKernel *out,
                 std::string *err) {
  {
    std::lock_guard<std::mutex> lk(g_mu);
    load_driver();
  }
  if (!g_api.drv_ok) {
    if (err) *err = g_api.drv_why;
    return -1;
  }
  cudaFree(0);  // make sure the runtime's primary context is current on this thread
  CUmodule mod = nullptr;
  CUresult r = g_api.ModuleLoadData(&mod, cubin.data());
  if (r != CUDA_SUCCESS) {
    if (err) *err = "cuModuleLoadData: " + drv_err(r);
    return -1;
  }
  CUfunction fn = nullptr;
  r = g_api.ModuleGetFunction(&fn, mod, "qec_jit_kernel");
  if (r != CUDA_SUCCESS) {
    if (err) *err = "cuModuleGetFunction: " + drv_err(r);
    g_api.ModuleUnload(mod);
    return -1;
  }
  if (dyn_smem > 0) {
    r = g_api.FuncSetAttribute(fn, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, (int)dyn_smem);
    if (r != CUDA_SUCCESS) {
      if (err) *err = "cuFuncSetAttribute(max dynamic smem): " + drv_err(r);
      g_api.ModuleUnload(mod);
      return -1;
    }
  }
  int regs = 0, ssm = 0;
  g_api.FuncGetAttribute(&regs, CU_FUNC_ATTRIBUTE_NUM_REGS, fn);
  g_api.FuncGetAttribute(&ssm, CU_FUNC_ATTRIBUTE_SHARED_SIZE_BYTES, fn);
  out->module = mod;
  out->function = fn;
  out->regs = regs;
  out->static_smem = (size_t)ssm;
  return 0;
}

int qec_jit_launch(const QecJitKernel &k, int grid, size_t dyn_smem, cudaStream_t stream,
                   void **args, std::string *err) {
  CUresult r = g_api.LaunchKernel((CUfunction)k.function, (unsigned)grid, 1, 1,
                                  (unsigned)k.threads, 1, 1, (unsigned)dyn_smem, (CUstream)stream,
                                  args, nullptr);
  if (r != CUDA_SUCCESS) {
    if (err) *err = "cuLaunchKernel: " + drv_err(r);
    return -1;
  }
  return 0;
}

void qec_jit_unload(QecJitKernel *k) {
  if (k && k->module && g_api.ModuleUnload) g_api.ModuleUnload((CUmodule)k->module);
  if (k) { k->module = nullptr; k->function = nullptr; }
}

