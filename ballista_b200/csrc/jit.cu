// jit.cu — run-time compilation of fused scan -> filter -> expression -> aggregate kernels for plan shapes that are
// not in the static registry (agg_registry.cu).  compile_expr in the reference accepts ANY expression
// (rust/core/src/serde/physical_plan/from_proto.rs:348-364); the bandwidth kernel (agg_tma.cuh) is a template over the
// expression trees, so an unseen (predicate, keys, arguments) shape is turned into the same template instantiation the
// registry holds for Q1 / Q6 - by NVRTC, for sm_100a, from the very headers the library was built from (embedded at build
// time: build/jit_headers.inc) - loaded through the driver API and cached in memory and on disk.  First use of a shape
// costs a compilation (seconds); every later execution is a registry hit.
//
// libnvrtc and libcuda are dlopen'ed on first use: the library keeps no link-time dependency on either, and when one is
// missing the plan simply stays on the column-at-a-time path (HashAggregateExec::last_path says so).
#include <cuda.h>
#include <dlfcn.h>
#include <nvrtc.h>
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <mutex>
#include <sstream>

#include "agg.h"
#include "jit.h"

namespace b2 {

namespace {
#include "build/jit_headers.inc"  // static const char* const JIT_HEADER_NAMES[] / JIT_HEADER_TEXT[]; JIT_N_HEADERS

struct Api {
  void* nvrtc = nullptr;
  void* cuda = nullptr;
  decltype(&nvrtcCreateProgram) createProgram = nullptr;
  decltype(&nvrtcCompileProgram) compileProgram = nullptr;
  decltype(&nvrtcDestroyProgram) destroyProgram = nullptr;
  decltype(&nvrtcGetCUBINSize) getCUBINSize = nullptr;
  decltype(&nvrtcGetCUBIN) getCUBIN = nullptr;
  decltype(&nvrtcGetProgramLogSize) getLogSize = nullptr;
  decltype(&nvrtcGetProgramLog) getLog = nullptr;
  decltype(&nvrtcAddNameExpression) addName = nullptr;
  decltype(&nvrtcGetLoweredName) getLowered = nullptr;
  decltype(&cuModuleLoadData) moduleLoadData = nullptr;
  decltype(&cuModuleGetFunction) moduleGetFunction = nullptr;
  decltype(&cuFuncGetAttribute) funcGetAttribute = nullptr;
  decltype(&cuFuncSetAttribute) funcSetAttribute = nullptr;
  decltype(&cuLaunchKernel) launchKernel = nullptr;
  decltype(&cuOccupancyMaxActiveBlocksPerMultiprocessor) occupancy = nullptr;
  std::string error;
  bool ok = false;        // compiler and driver
  bool ok_nvrtc = false;  // compiler alone (enough to check that a shape compiles, without a GPU)
};
Api& api() {
  static Api a;
  static std::once_flag once;
  std::call_once(once, [] {
    for (const char* n : {"libnvrtc.so.12", "libnvrtc.so", "/usr/local/cuda/lib64/libnvrtc.so.12"})
      if (!a.nvrtc) a.nvrtc = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    for (const char* n : {"libcuda.so.1", "libcuda.so"})
      if (!a.cuda) a.cuda = dlopen(n, RTLD_NOW | RTLD_LOCAL);
    if (!a.nvrtc) {
      a.error = "libnvrtc.so.12 not found";
      return;
    }
#define B2_SYM(lib, field, name)                                   \
  a.field = reinterpret_cast<decltype(a.field)>(dlsym(lib, name)); \
  if (!a.field) {                                                  \
    a.error = std::string("missing symbol ") + name;               \
    return;                                                        \
  }
    B2_SYM(a.nvrtc, createProgram, "nvrtcCreateProgram")
    B2_SYM(a.nvrtc, compileProgram, "nvrtcCompileProgram")
    B2_SYM(a.nvrtc, destroyProgram, "nvrtcDestroyProgram")
    B2_SYM(a.nvrtc, getCUBINSize, "nvrtcGetCUBINSize")
    B2_SYM(a.nvrtc, getCUBIN, "nvrtcGetCUBIN")
    B2_SYM(a.nvrtc, getLogSize, "nvrtcGetProgramLogSize")
    B2_SYM(a.nvrtc, getLog, "nvrtcGetProgramLog")
    B2_SYM(a.nvrtc, addName, "nvrtcAddNameExpression")
    B2_SYM(a.nvrtc, getLowered, "nvrtcGetLoweredName")
    a.ok_nvrtc = true;
    if (!a.cuda) {
      a.error = "libcuda.so.1 not found";
      return;
    }
    B2_SYM(a.cuda, moduleLoadData, "cuModuleLoadData")
    B2_SYM(a.cuda, moduleGetFunction, "cuModuleGetFunction")
    B2_SYM(a.cuda, funcGetAttribute, "cuFuncGetAttribute")
    B2_SYM(a.cuda, funcSetAttribute, "cuFuncSetAttribute")
    B2_SYM(a.cuda, launchKernel, "cuLaunchKernel")
    B2_SYM(a.cuda, occupancy, "cuOccupancyMaxActiveBlocksPerMultiprocessor")
#undef B2_SYM
    a.ok = true;
  });
  return a;
}

// ---- signature text (expr_signature, expr.cpp) -> the C++ template type of fused.cuh ----
// value classes: 'd' Float64, 'i' Int32 / integer literal, 's' Utf8, 'b' Boolean; 0 = not expressible
struct Typed {
  std::string type;
  char cls = 0;
};
struct SigParser {
  const std::string& s;
  size_t pos = 0;
  explicit SigParser(const std::string& t) : s(t) {}
  bool done() const { return pos >= s.size(); }
  std::string ident() {
    size_t b = pos;
    while (pos < s.size() && (isalnum((unsigned char)s[pos]) || s[pos] == '_')) ++pos;
    return s.substr(b, pos - b);
  }
  Typed expr() {
    std::string id = ident();
    if (id.empty()) return {};
    if (pos < s.size() && s[pos] == '(') {
      ++pos;
      std::vector<Typed> args;
      while (true) {
        Typed a = expr();
        if (!a.cls) return {};
        args.push_back(a);
        if (pos < s.size() && s[pos] == ',') {
          ++pos;
          continue;
        }
        break;
      }
      if (pos >= s.size() || s[pos] != ')') return {};
      ++pos;
      return apply(id, args);
    }
    return leaf(id);
  }
  static Typed leaf(const std::string& id) {
    auto num = [&](size_t from) {
      if (from >= id.size()) return std::string();
      for (size_t k = from; k < id.size(); ++k)
        if (!isdigit((unsigned char)id[k])) return std::string();
      return id.substr(from);
    };
    if (id == "true") return {"fx::True", 'b'};
    if (id[0] == 'f' && !num(1).empty()) return {"fx::F<" + num(1) + ">", 'd'};
    if (id[0] == 'i' && !num(1).empty()) return {"fx::I<" + num(1) + ">", 'i'};
    if (id[0] == 's' && !num(1).empty()) return {"fx::S<" + num(1) + ">", 's'};
    if (id.compare(0, 2, "LF") == 0 && !num(2).empty()) return {"fx::LF<" + num(2) + ">", 'd'};
    if (id.compare(0, 2, "LI") == 0 && !num(2).empty()) return {"fx::LI<" + num(2) + ">", 'i'};
    if (id.compare(0, 2, "LS") == 0 && !num(2).empty()) return {"fx::LS<" + num(2) + ">", 's'};
    return {};
  }
  static Typed apply(const std::string& fn, const std::vector<Typed>& a) {
    auto bin = [&](const char* t, char cls) { return Typed{std::string("fx::") + t + "<" + a[0].type + "," + a[1].type + ">", cls}; };
    if (a.size() == 2) {
      const char x = a[0].cls, y = a[1].cls;
      if (fn == "add" && x == 'd' && y == 'd') return bin("Add", 'd');
      if (fn == "sub" && x == 'd' && y == 'd') return bin("Sub", 'd');
      if (fn == "mul" && x == 'd' && y == 'd') return bin("Mul", 'd');
      if (fn == "div" && x == 'd' && y == 'd') return bin("Div", 'd');
      const bool num = (x == 'd' && y == 'd') || (x == 'i' && y == 'i');
      if (fn == "lt" && num) return bin("Lt", 'b');
      if (fn == "le" && num) return bin("Le", 'b');
      if (fn == "gt" && num) return bin("Gt", 'b');
      if (fn == "ge" && num) return bin("Ge", 'b');
      if (fn == "eq" && num) return bin("Eq", 'b');
      if (fn == "ne" && num) return bin("Ne", 'b');
      if (fn == "eq" && x == 's' && y == 's') return bin("SEq", 'b');
      if (fn == "ne" && x == 's' && y == 's') return bin("SNe", 'b');
      if (fn == "and" && x == 'b' && y == 'b') return bin("And", 'b');
      if (fn == "or" && x == 'b' && y == 'b') return bin("Or", 'b');
      return {};
    }
    if (a.size() == 1) {
      if (fn == "not" && a[0].cls == 'b') return {"fx::Not<" + a[0].type + ">", 'b'};
      if (fn == "neg" && a[0].cls == 'd') return {"fx::Neg<" + a[0].type + ">", 'd'};
      if (fn == "f64" && a[0].cls == 'i') return {"fx::ToF64<" + a[0].type + ">", 'd'};
    }
    return {};
  }
};

uint64_t fnv(const std::string& s, uint64_t h = 1469598103934665603ull) {
  for (unsigned char c : s) h = (h ^ c) * 1099511628211ull;
  return h;
}
std::string cache_dir() {
  if (const char* e = getenv("B2_JIT_CACHE")) return e;
  const char* home = getenv("HOME");
  return std::string(home ? home : "/tmp") + "/.cache/ballista_b200/jit";
}
void mkdirs(const std::string& path) {
  std::string cur;
  for (size_t i = 0; i <= path.size(); ++i) {
    if (i == path.size() || path[i] == '/') {
      if (!cur.empty()) mkdir(cur.c_str(), 0755);
    }
    if (i < path.size()) cur += path[i];
  }
}

std::mutex g_mu;
std::map<std::string, std::shared_ptr<JitKernels>> g_cache;  // signature -> kernels (null = known not to compile)

}  // namespace

bool jit_signature_types(const std::string& sig, std::string* pred_t, std::string* keys_t, std::string* args_t,
                         int* n_int_keys, int* n_str_keys) {
  // "p:<expr>|k:<e>,<e>|a:<e>,<e>"
  size_t kp = sig.find("|k:"), ap = sig.find("|a:");
  if (sig.compare(0, 2, "p:") != 0 || kp == std::string::npos || ap == std::string::npos || ap < kp) return false;
  const std::string ps = sig.substr(2, kp - 2), ks = sig.substr(kp + 3, ap - kp - 3), as = sig.substr(ap + 3);
  {
    SigParser p(ps);
    Typed t = p.expr();
    if (t.cls != 'b' || !p.done()) return false;
    *pred_t = t.type;
  }
  auto list = [&](const std::string& text, bool keys, std::string* out) {
    *out = "fx::List<";
    SigParser p(text);
    bool first = true;
    while (!p.done()) {
      Typed t = p.expr();
      if (!t.cls) return false;
      if (keys) {  // group keys: plain Int32 / Utf8 column slots
        if (t.type.compare(0, 6, "fx::I<") == 0) ++*n_int_keys;
        else if (t.type.compare(0, 6, "fx::S<") == 0) ++*n_str_keys;
        else return false;
      } else if (t.cls != 'd') {
        return false;  // aggregate arguments of the fused kernels are Float64
      }
      *out += (first ? "" : ",") + t.type;
      first = false;
      if (!p.done()) {
        if (text[p.pos] != ',') return false;
        ++p.pos;
      }
    }
    *out += ">";
    return true;
  };
  *n_int_keys = *n_str_keys = 0;
  return list(ks, true, keys_t) && list(as, false, args_t);
}

// source + instantiated kernel names of a shape; false when the shape is outside the fused vocabulary
static bool jit_source(const std::string& sig, std::string* src, std::vector<std::string>* names, bool* narrow_ok,
                       std::string* why) {
  std::string pred_t, keys_t, args_t;
  int ki = 0, ks = 0;
  if (!jit_signature_types(sig, &pred_t, &keys_t, &args_t, &ki, &ks)) {
    *why = "expression outside the fused vocabulary";
    return false;
  }
  if (4 * ki > 15 || (ks && (15 - 4 * ki) / ks < 2)) {
    *why = "key wider than 15 packed bytes";
    return false;
  }
  *narrow_ok = 4 * ki <= 8 && (ks == 0 || (8 - 4 * ki) / ks >= 2);  // fx::narrow_key_ok
  *src = "#include \"agg_tma.cuh\"\n#include \"agg_kernels.cuh\"\nnamespace b2 {\nusing JitSrc = FusedSrc<" + pred_t + ", " +
         keys_t + ", " + args_t + ">;\n}\n";
  names->clear();
  if (*narrow_ok) names->push_back("b2::agg_tma_kernel<b2::JitSrc, true>");
  names->push_back("b2::agg_tma_kernel<b2::JitSrc, false>");
  names->push_back("b2::agg_small_kernel<b2::JitSrc>");
  names->push_back("b2::agg_global_kernel<b2::JitSrc>");
  return true;
}
static bool jit_compile(const std::string& src, const std::vector<std::string>& names, std::string* cubin,
                        std::vector<std::string>* lowered, std::string* why) {
  Api& a = api();
  if (!a.ok_nvrtc) {
    *why = "run-time compilation unavailable: " + a.error;
    return false;
  }
  nvrtcProgram prog;
  if (a.createProgram(&prog, src.c_str(), "b2_jit.cu", JIT_N_HEADERS, JIT_HEADER_TEXT, JIT_HEADER_NAMES) != NVRTC_SUCCESS) {
    *why = "nvrtcCreateProgram failed";
    return false;
  }
  for (auto& n : names) a.addName(prog, n.c_str());
  const char* cuda_home = getenv("CUDA_HOME");
  std::string inc = std::string("-I") + (cuda_home ? cuda_home : "/usr/local/cuda") + "/include";
  // -default-device: the headers' constexpr helpers carry no execution-space annotation
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo", "-default-device", inc.c_str()};
  nvrtcResult r = a.compileProgram(prog, 5, opts);
  if (r != NVRTC_SUCCESS) {
    size_t n = 0;
    a.getLogSize(prog, &n);
    std::string log(n, 0);
    if (n) a.getLog(prog, &log[0]);
    a.destroyProgram(&prog);
    *why = "NVRTC: " + log.substr(0, 2000);
    return false;
  }
  size_t n = 0;
  a.getCUBINSize(prog, &n);
  cubin->resize(n);
  a.getCUBIN(prog, &(*cubin)[0]);
  lowered->clear();
  for (auto& nm : names) {
    const char* low = nullptr;
    a.getLowered(prog, nm.c_str(), &low);
    lowered->push_back(low ? low : "");
  }
  a.destroyProgram(&prog);
  return true;
}
// compile only (no device needed): the size of the cubin, or 0 and *why
size_t jit_compile_check(const std::string& sig, std::string* why) {
  std::string src, cubin;
  std::vector<std::string> names, lowered;
  bool narrow = false;
  if (!jit_source(sig, &src, &names, &narrow, why)) return 0;
  if (!jit_compile(src, names, &cubin, &lowered, why)) return 0;
  return cubin.size();
}

std::shared_ptr<JitKernels> jit_fused_kernels(const std::string& sig, std::string* why) {
  std::lock_guard<std::mutex> g(g_mu);
  auto hit = g_cache.find(sig);
  if (hit != g_cache.end()) {
    if (!hit->second && why) *why = "shape known not to compile";
    return hit->second;
  }
  auto fail_with = [&](const std::string& m) {
    if (why) *why = m;
    g_cache[sig] = nullptr;
    return std::shared_ptr<JitKernels>();
  };
  Api& a = api();
  auto out = std::make_shared<JitKernels>();
  std::string src, err;
  std::vector<std::string> names;
  if (!jit_source(sig, &src, &names, &out->narrow_ok, &err)) return fail_with(err);
  if (!a.ok) return fail_with("run-time compilation unavailable: " + a.error);
  // ---- disk cache: cubin + lowered names, keyed on the source, the embedded headers and the compiler version ----
  uint64_t h = fnv(src);
  for (int k = 0; k < JIT_N_HEADERS; ++k) h = fnv(JIT_HEADER_TEXT[k], h);
  h = fnv("nvrtc-12.9-sm_100a", h);
  char hex[32];
  snprintf(hex, sizeof hex, "%016llx", (unsigned long long)h);
  const std::string dir = cache_dir(), base = dir + "/" + hex;
  std::string cubin;
  std::vector<std::string> lowered;
  if (!getenv("B2_JIT_NO_DISK_CACHE")) {
    std::ifstream fc(base + ".cubin", std::ios::binary), fn(base + ".names");
    if (fc && fn) {
      std::stringstream ss;
      ss << fc.rdbuf();
      cubin = ss.str();
      std::string line;
      while (std::getline(fn, line)) lowered.push_back(line);
      if (lowered.size() != names.size()) cubin.clear();
    }
  }
  if (cubin.empty()) {
    if (!jit_compile(src, names, &cubin, &lowered, &err)) return fail_with(err);
    if (!getenv("B2_JIT_NO_DISK_CACHE")) {
      mkdirs(dir);
      std::ofstream fc(base + ".cubin", std::ios::binary), fn(base + ".names");
      fc.write(cubin.data(), (std::streamsize)cubin.size());
      for (auto& l : lowered) fn << l << "\n";
    }
    out->compiled_now = true;
  }
  CUmodule mod;
  CUresult cr = a.moduleLoadData(&mod, cubin.data());
  if (cr != CUDA_SUCCESS) return fail_with("cuModuleLoadData failed (" + std::to_string((int)cr) + ")");
  size_t k = 0;
  auto get = [&](CUfunction* f) { return a.moduleGetFunction(f, mod, lowered[k++].c_str()) == CUDA_SUCCESS; };
  bool ok = true;
  if (out->narrow_ok) ok = ok && get(reinterpret_cast<CUfunction*>(&out->tma_narrow));
  ok = ok && get(reinterpret_cast<CUfunction*>(&out->tma_wide)) && get(reinterpret_cast<CUfunction*>(&out->small)) &&
       get(reinterpret_cast<CUfunction*>(&out->global));
  if (!ok) return fail_with("kernel not found in the compiled module");
  for (void* f : {out->tma_narrow, out->tma_wide}) {
    if (!f) continue;
    int st = 0;
    a.funcGetAttribute(&st, CU_FUNC_ATTRIBUTE_SHARED_SIZE_BYTES, (CUfunction)f);
    a.funcSetAttribute((CUfunction)f, CU_FUNC_ATTRIBUTE_MAX_DYNAMIC_SHARED_SIZE_BYTES, 232448 - st);
    (f == out->tma_narrow ? out->static_smem_narrow : out->static_smem_wide) = st;
  }
  a.occupancy(&out->per_sm_small, (CUfunction)out->small, 256, 0);
  a.occupancy(&out->per_sm_global, (CUfunction)out->global, 256, 0);
  if (out->per_sm_small < 1) out->per_sm_small = 1;
  if (out->per_sm_global < 1) out->per_sm_global = 1;
  g_cache[sig] = out;
  return out;
}

void jit_launch(void* fn, int grid, int block, size_t dyn_smem, cudaStream_t stream, void** args) {
  Api& a = api();
  CUresult r = a.launchKernel((CUfunction)fn, (unsigned)grid, 1, 1, (unsigned)block, 1, 1, (unsigned)dyn_smem, (CUstream)stream,
                              args, nullptr);
  if (r != CUDA_SUCCESS) fail(B2_ERR_CUDA, "cuLaunchKernel of a run-time compiled kernel failed (" + std::to_string((int)r) + ")");
}

}  // namespace b2
