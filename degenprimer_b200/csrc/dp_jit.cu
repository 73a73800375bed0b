// Scan kernel 3: the bit-sliced prefilter, compiled at run time for one pattern set.
//
// The generic kernel (dp_scan_bitslice.cuh) has to pick the one-hot plane of every
// anchor letter by a run-time index, which costs a shared-memory store per plane and
// a shared-memory load per term: 80 of its ~300 instructions per word at 4 patterns,
// plus the uniform loads of the offsets (profiles/README.md).  With the letters known
// at compile time the plane of (shift j, letter set S) is one LOP3 of the two shifted
// code words, lives in a register and is shared by every pattern that uses it: no
// shared memory, no offset loads, one launch for all patterns of a call, and a
// degenerate letter costs the same single LOP3.
//
// NVRTC is bound at run time (dlopen); the cubin is loaded with cudaLibraryLoadData.
// Kernels are cached per process by their source text.  A call uses a specialised
// kernel when its pattern set was prepared (dp_scan_prepare_kernels, which is what
// Context.prepare() calls) or when the caller asks for DP_ALGO_JIT; otherwise the
// generic kernel runs, so a one-off query never pays the ~0.3 s compile.
#include <dlfcn.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>

#include <map>
#include <mutex>
#include <string>

#include "dp_jit.cuh"

static const char DP_DEVICE_HEADER[] =
#include "../build/dp_jit_embed.inc"
    ;

// ---- NVRTC, bound with dlopen ------------------------------------------------------
typedef struct _nvrtcProgram* nvrtcProgram;
struct dp_nvrtc_api {
  void* lib;
  int (*CreateProgram)(nvrtcProgram*, const char*, const char*, int, const char* const*, const char* const*);
  int (*CompileProgram)(nvrtcProgram, int, const char* const*);
  int (*GetCUBINSize)(nvrtcProgram, size_t*);
  int (*GetCUBIN)(nvrtcProgram, char*);
  int (*GetProgramLogSize)(nvrtcProgram, size_t*);
  int (*GetProgramLog)(nvrtcProgram, char*);
  int (*DestroyProgram)(nvrtcProgram*);
  const char* (*GetErrorString)(int);
};

static dp_nvrtc_api* nvrtc_api(std::string* why) {
  static dp_nvrtc_api api;
  static bool tried = false, ok = false;
  static std::string reason;
  if (!tried) {
    tried = true;
    const char* env = getenv("DEGENPRIMER_B200_NVRTC");
    const char* names[] = {env, "/usr/local/cuda/lib64/libnvrtc.so.12", "libnvrtc.so.12", "libnvrtc.so"};
    for (const char* n : names) {
      if (!n || !*n) continue;
      api.lib = dlopen(n, RTLD_NOW | RTLD_LOCAL);
      if (api.lib) break;
    }
    if (!api.lib) {
      reason = "libnvrtc.so.12 could not be loaded (set DEGENPRIMER_B200_NVRTC to its path)";
    } else {
#define DP_SYM(field, name) *(void**)&api.field = dlsym(api.lib, name)
      DP_SYM(CreateProgram, "nvrtcCreateProgram");
      DP_SYM(CompileProgram, "nvrtcCompileProgram");
      DP_SYM(GetCUBINSize, "nvrtcGetCUBINSize");
      DP_SYM(GetCUBIN, "nvrtcGetCUBIN");
      DP_SYM(GetProgramLogSize, "nvrtcGetProgramLogSize");
      DP_SYM(GetProgramLog, "nvrtcGetProgramLog");
      DP_SYM(DestroyProgram, "nvrtcDestroyProgram");
      DP_SYM(GetErrorString, "nvrtcGetErrorString");
#undef DP_SYM
      ok = api.CreateProgram && api.CompileProgram && api.GetCUBINSize && api.GetCUBIN &&
           api.GetProgramLogSize && api.GetProgramLog && api.DestroyProgram && api.GetErrorString;
      if (!ok) reason = "libnvrtc is missing a symbol";
    }
  }
  if (!ok && why) *why = reason;
  return ok ? &api : nullptr;
}

// source -> cubin.  Empty string on failure (`log` then holds the reason).
static std::string nvrtc_compile(const std::string& source, std::string& log) {
  std::string why;
  dp_nvrtc_api* api = nvrtc_api(&why);
  if (!api) {
    log = why;
    return std::string();
  }
  nvrtcProgram prog = nullptr;
  const char* hdr_src[] = {DP_DEVICE_HEADER};
  const char* hdr_name[] = {"dp_device.cuh"};
  int rc = api->CreateProgram(&prog, source.c_str(), "dp_jit_scan.cu", 1, hdr_src, hdr_name);
  if (rc) {
    log = std::string("nvrtcCreateProgram: ") + api->GetErrorString(rc);
    return std::string();
  }
  const char* opts[] = {"--gpu-architecture=sm_100a", "-std=c++17", "-lineinfo"};
  rc = api->CompileProgram(prog, 3, opts);
  std::string cubin;
  if (rc) {
    size_t n = 0;
    api->GetProgramLogSize(prog, &n);
    std::string text(n ? n : 1, '\0');
    if (n) api->GetProgramLog(prog, &text[0]);
    log = std::string("nvrtcCompileProgram: ") + api->GetErrorString(rc) + "\n" + text;
  } else {
    size_t n = 0;
    if (api->GetCUBINSize(prog, &n) == 0 && n) {
      cubin.resize(n);
      if (api->GetCUBIN(prog, &cubin[0])) cubin.clear();
    }
    if (cubin.empty()) log = "nvrtcGetCUBIN failed";
  }
  api->DestroyProgram(&prog);
  return cubin;
}

// compile without loading (no GPU needed): used by the CPU-side build check
std::string dp_jit_compile_only(const std::string& source, std::string& log) { return nvrtc_compile(source, log); }

// ---- the kernel cache ------------------------------------------------------------------
static std::mutex g_jit_mutex;
static std::map<std::string, dp_jit_kernel*> g_jit_kernels;      // by source text; entries live for the process

dp_jit_kernel* dp_jit_lookup(const std::string& source) {
  std::lock_guard<std::mutex> lock(g_jit_mutex);
  auto it = g_jit_kernels.find(source);
  return it == g_jit_kernels.end() ? nullptr : it->second;
}

int dp_jit_compile(dp_ctx* ctx, const std::string& source, dp_jit_kernel** out) {
  *out = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_jit_mutex);
    auto it = g_jit_kernels.find(source);
    if (it != g_jit_kernels.end()) {
      *out = it->second;
      return (*out)->fn ? DP_OK : dp_set_error(ctx, DP_ERR_STATE, "run-time compile failed before: %s", (*out)->log.c_str());
    }
  }
  dp_jit_kernel* k = new dp_jit_kernel();
  k->lib = nullptr;
  k->fn = nullptr;
  k->per_sm = 0;
  k->regs = 0;
  std::string cubin = nvrtc_compile(source, k->log);
  if (!cubin.empty()) {
    cudaError_t e = cudaLibraryLoadData(&k->lib, cubin.data(), nullptr, nullptr, 0, nullptr, nullptr, 0);
    if (e == cudaSuccess) e = cudaLibraryGetKernel(&k->fn, k->lib, "dp_jit_scan");
    if (e != cudaSuccess) {
      k->log = std::string("loading the compiled kernel: ") + cudaGetErrorString(e);
      k->fn = nullptr;
      cudaGetLastError();
    } else {
      cudaFuncAttributes attr;
      if (cudaFuncGetAttributes(&attr, (const void*)k->fn) == cudaSuccess) k->regs = attr.numRegs;
      int per_sm = 0;
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)k->fn, DP_JIT_THREADS, 0) == cudaSuccess && per_sm > 0)
        k->per_sm = per_sm;
      else
        k->per_sm = k->regs ? std::max(1, std::min(8, 65536 / (k->regs * DP_JIT_THREADS))) : 4;
      cudaGetLastError();
    }
  }
  {
    std::lock_guard<std::mutex> lock(g_jit_mutex);
    auto it = g_jit_kernels.find(source);
    if (it != g_jit_kernels.end()) {      // another thread was faster; keep its entry
      delete k;
      k = it->second;
    } else {
      g_jit_kernels[source] = k;
    }
  }
  *out = k;
  return k->fn ? DP_OK : dp_set_error(ctx, DP_ERR_STATE, "run-time compile of the scan kernel failed: %s", k->log.c_str());
}

// ---- source generation -----------------------------------------------------------------
// All offsets are window positions: bit b of the thread's word w is the window that
// starts at base 32w+b, position j of it sits at bit b of the text shifted right by j.
//
// Two prefilters, chosen per pattern by instruction count (both are necessary
// conditions for a hit, K = hard mismatches a hit can have):
//   EXACT  K+1 disjoint sets of positions; a hit matches at least one set completely
//          (pigeonhole).  A set is an AND chain, one LOP3 per position:
//          acc = acc & f(h_j, l_j).  ~9 positions per set keep the verify rate low.
//   ADDER  the generic kernel's anchor: J = 8+2K positions, bit-sliced count >= J-K
//          (one LOP3 per plane + ~1.8 per adder input).
// The integer pipe (LOP3/SHF, 16 lanes per SM sub-partition and clock) is what bounds
// this kernel.  Moving the funnel shifts to the FMA pipe as multiply-adds
// (hi32(h0 * 2^(32-j)) + lo32(hn * 2^(32-j)), multiplier from a kernel parameter so that
// ptxas keeps the IMADs) was measured and is slower (IMAD.HI), so the shifts stay SHF.

// match plane of letter set `set` (bit c = plane code c: A C T G = (hi,lo) 00 01 10 11)
static std::string plane_expr(uint32_t set, const std::string& h, const std::string& l) {
  if ((set & 15u) == 15u) return "0xffffffffu";
  std::string s = "(";
  bool first = true;
  for (int c = 0; c < 4; ++c) {
    if (!(set & (1u << c))) continue;
    if (!first) s += " | ";
    first = false;
    s += (c & 2) ? "(" + h : "(~" + h;
    s += (c & 1) ? " & " + l + ")" : " & ~" + l + ")";
  }
  if (first) s += "0u";
  return s + ")";
}

static double set_fraction(uint32_t set) { return __builtin_popcount(set & 15u) / 4.0; }

// expected instructions per word spent in the out-of-line verify for a prefilter that
// lets a random window through with probability p
static double cold_cost(double p) {
  const double any = 1.0 - pow(1.0 - std::min(p, 0.5), 1024.0);   // some lane of the warp, some bit of its word
  return 250.0 * any;
}

// EXACT: K+1 disjoint position sets within the first `span` positions.  Returns the
// estimated instructions per word (sets + verify), or a negative number if infeasible.
static double choose_exact(const std::vector<uint8_t>& lt, int K, int span, std::vector<std::vector<int>>& sets) {
  const int L = std::min<int>((int)lt.size(), span);
  std::vector<int> order;
  for (int j = 0; j < L; ++j)
    if (set_fraction(dp_letter_set(lt[j])) < 1.0) order.push_back(j);
  std::stable_sort(order.begin(), order.end(), [&](int x, int y) {
    return set_fraction(dp_letter_set(lt[x])) < set_fraction(dp_letter_set(lt[y]));
  });
  double best = -1.0;
  std::vector<std::vector<int>> best_sets;
  // deal the most selective positions round-robin; try every common depth
  for (int depth = 6; depth * (K + 1) <= (int)order.size() && depth <= 12; ++depth) {
    std::vector<std::vector<int>> g(K + 1);
    for (int i = 0; i < depth * (K + 1); ++i) g[i % (K + 1)].push_back(order[i]);
    double p = 0.0;
    for (auto& v : g) {
      double q = 1.0;
      for (int j : v) q *= set_fraction(dp_letter_set(lt[j]));
      p += q;
    }
    const double cost = depth * (K + 1) + K + cold_cost(p);
    if (best < 0 || cost < best) {
      best = cost;
      best_sets = g;
    }
  }
  if (best >= 0) {
    for (auto& v : best_sets) std::sort(v.begin(), v.end());
    sets = best_sets;
  }
  return best;
}

static double adder_cost(const std::vector<uint8_t>& lt, const dp_jit_entry& e) {
  const int J = 8 + 2 * e.K;
  // alive probability as in dp_bitslice_make_plan
  std::vector<double> d(e.K + 2, 0.0);
  d[0] = 1.0;
  for (int j = 0; j < J; ++j) {
    const double pm = set_fraction(dp_letter_set(lt[e.anchor + j]));
    for (int m = e.K + 1; m >= 0; --m) {
      double stay = m == e.K + 1 ? d[m] : d[m] * pm;
      double from = m > 0 ? d[m - 1] * (1.0 - pm) : 0.0;
      d[m] = stay + from;
    }
  }
  double p = 0.0;
  for (int m = 0; m <= e.K; ++m) p += d[m];
  return 2.8 * J + 5.0 + cold_cost(p);
}

struct unit_pattern {
  dp_jit_entry e;
  bool exact;
  std::vector<std::vector<int>> sets;     // EXACT
};

static int jit_env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

static std::string unit_source(const std::vector<dp_pattern>& pats, const std::vector<std::vector<uint8_t>>& letters,
                               const std::vector<dp_jit_entry>& entries) {
  const int use_exact = jit_env_int("DEGENPRIMER_B200_JIT_EXACT", 1);
  std::vector<unit_pattern> ups;
  bool used[64] = {false};
  for (const dp_jit_entry& e : entries) {
    unit_pattern u;
    u.e = e;
    u.exact = false;
    const std::vector<uint8_t>& lt = letters[e.pattern];
    if (use_exact) {
      std::vector<std::vector<int>> sets;
      const double ce = choose_exact(lt, e.K, 24, sets);
      if (ce >= 0 && ce < adder_cost(lt, e)) {
        u.exact = true;
        u.sets = sets;
      }
    }
    if (u.exact) {
      for (auto& v : u.sets)
        for (int j : v) used[j] = true;
    } else {
      for (int j = 0; j < 8 + 2 * e.K; ++j) used[e.anchor + j] = true;
    }
    ups.push_back(u);
  }
  int max_off = 0;
  for (int j = 1; j < 64; ++j)
    if (used[j]) max_off = j;
  const bool three = max_off >= 32;        // positions 32..63 need the third word

  std::string s;
  s.reserve(4096 + 1024 * entries.size());
  s += "#include \"dp_device.cuh\"\n";
  s += "static_assert(sizeof(dp_pattern) == " + std::to_string(sizeof(dp_pattern)) + ", \"dp_pattern layout\");\n";
  s += "static_assert(sizeof(dp_scan_args) == " + std::to_string(sizeof(dp_scan_args)) + ", \"dp_scan_args layout\");\n";
  s += "static_assert(sizeof(dp_raw_hit) == " + std::to_string(sizeof(dp_raw_hit)) + ", \"dp_raw_hit layout\");\n";
  s += "#define NP " + std::to_string(entries.size()) + "\n";
  // cold path, kept out of line so that the word loop stays small
  s += "__device__ __noinline__ void dp_jit_alive(const dp_scan_args& A, const dp_pattern* pats, uint32_t slot,\n"
       "                                          uint32_t pattern, uint32_t alive, uint64_t w,\n"
       "                                          unsigned long long& n_cand) {\n"
       "  while (alive) {\n"
       "    const int b = __ffs(alive) - 1;\n"
       "    alive &= alive - 1;\n"
       "    ++n_cand;\n"
       "    const uint64_t g = w * 32 + b;\n"
       "    if (g < A.seg_begin || g >= A.seg_end) continue;\n"
       "    dp_verify_window(A, pats[slot], g, pattern, false);\n"
       "  }\n"
       "}\n";
  s += "extern \"C\" __global__ void __launch_bounds__(" + std::to_string(DP_JIT_THREADS) + ", 3)\n"
       "dp_jit_scan(const __grid_constant__ dp_scan_args A, const __grid_constant__ dp_pattern_set<NP> PS) {\n"
       "  const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;\n"
       "  const uint64_t w_end = A.word_end;\n"
       "  unsigned long long n_cand = 0;\n"
       "  uint64_t w = A.word_begin + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;\n"
       "  uint32_t h0 = 0, hn = 0, hm = 0, l0 = 0, ln = 0, lm = 0;\n"
       "  if (w < w_end) {\n"
       "    h0 = __ldg(A.hi + w); hn = __ldg(A.hi + w + 1);\n"
       "    l0 = __ldg(A.lo + w); ln = __ldg(A.lo + w + 1);\n";
  if (three) s += "    hm = __ldg(A.hi + w + 2); lm = __ldg(A.lo + w + 2);\n";
  s += "  }\n"
       "  while (w < w_end) {\n"
       "    const uint64_t wn = w + stride;\n"
       "    uint32_t ph0 = 0, phn = 0, phm = 0, pl0 = 0, pln = 0, plm = 0;\n"
       "    if (wn < w_end) {\n"
       "      ph0 = __ldg(A.hi + wn); phn = __ldg(A.hi + wn + 1);\n"
       "      pl0 = __ldg(A.lo + wn); pln = __ldg(A.lo + wn + 1);\n";
  if (three) s += "      phm = __ldg(A.hi + wn + 2); plm = __ldg(A.lo + wn + 2);\n";
  s += "    }\n";
  for (int j = 1; j < 64; ++j) {
    if (!used[j]) continue;
    const std::string n = std::to_string(j);
    if (j == 32) {
      s += "    const uint32_t h32 = hn, l32 = ln;\n";
      continue;
    }
    const std::string x0 = j < 32 ? "0" : "n", x1 = j < 32 ? "n" : "m";
    const int sh = j & 31;
      s += "    const uint32_t h" + n + " = __funnelshift_r(h" + x0 + ", h" + x1 + ", " + std::to_string(sh) + "), l" + n +
           " = __funnelshift_r(l" + x0 + ", l" + x1 + ", " + std::to_string(sh) + ");\n";
  }
  // patterns in chunks of 8: one branch per chunk on the hot path
  for (size_t c0 = 0; c0 < ups.size(); c0 += 8) {
    const size_t c1 = std::min(ups.size(), c0 + 8);
    s += "    {\n";
    for (size_t i = c0; i < c1; ++i) {
      const unit_pattern& u = ups[i];
      const std::vector<uint8_t>& lt = letters[u.e.pattern];
      const std::string v = "alive" + std::to_string(i);
      if (u.exact) {
        s += "      uint32_t " + v + ";\n      {\n";
        for (size_t g = 0; g < u.sets.size(); ++g) {
          const std::string acc = "e" + std::to_string(g);
          bool first = true;
          for (int j : u.sets[g]) {
            const std::string n = std::to_string(j);
            const std::string pe = plane_expr(dp_letter_set(lt[j]), "h" + n, "l" + n);
            s += first ? "        uint32_t " + acc + " = " + pe + ";\n" : "        " + acc + " &= " + pe + ";\n";
            first = false;
          }
        }
        s += "        " + v + " = e0";
        for (size_t g = 1; g < u.sets.size(); ++g) s += " | e" + std::to_string(g);
        s += ";\n      }\n";
      } else {
        const int J = 8 + 2 * u.e.K, THR = J - u.e.K;
        // fresh names for the shifted words behind an empty asm: keeps the compiler from
        // computing every plane of every pattern up front and holding them in registers
        s += "      uint32_t " + v + ";\n      {\n        uint32_t";
        for (int j = 0; j < J; ++j) {
          const std::string n = std::to_string(j), o = std::to_string(u.e.anchor + j);
          s += (j ? ", a" : " a") + n + " = h" + o + ", b" + n + " = l" + o;
        }
        s += ";\n";
        for (int j0 = 0; j0 < J; j0 += 8) {
          s += "        asm volatile(\"\" :";
          for (int j = j0; j < std::min(J, j0 + 8); ++j) {
            const std::string n = std::to_string(j);
            s += (j > j0 ? ", \"+r\"(a" : " \"+r\"(a") + n + "), \"+r\"(b" + n + ")";
          }
          s += ");\n";
        }
        s += "        const uint32_t x[" + std::to_string(J) + "] = {";
        for (int j = 0; j < J; ++j) {
          const std::string n = std::to_string(j);
          s += (j ? ", " : "") + plane_expr(dp_letter_set(lt[u.e.anchor + j]), "a" + n, "b" + n);
        }
        s += "};\n";
        s += "        " + v + " = dp_count_ge<" + std::to_string(J) + ", " + std::to_string(THR) + ">(x);\n      }\n";
      }
    }
    s += "      if (alive" + std::to_string(c0);
    for (size_t i = c0 + 1; i < c1; ++i) s += " | alive" + std::to_string(i);
    s += ") {\n";
    for (size_t i = c0; i < c1; ++i)
      s += "        if (alive" + std::to_string(i) + ") dp_jit_alive(A, PS.pat, " + std::to_string(i) + "u, " +
           std::to_string(ups[i].e.pattern) + "u, alive" + std::to_string(i) + ", w, n_cand);\n";
    s += "      }\n    }\n";
  }
  s += "    w = wn;\n"
       "    h0 = ph0; hn = phn; hm = phm; l0 = pl0; ln = pln; lm = plm;\n"
       "  }\n"
       "  if (n_cand) atomicAdd(&A.counters[DP_C_CAND], n_cand);\n"
       "}\n";
  return s;
}

void dp_jit_make_units(const std::vector<dp_pattern>& pats, const std::vector<std::vector<uint8_t>>& letters,
                       const std::vector<dp_jit_entry>& entries, std::vector<dp_jit_unit>& units) {
  units.clear();
  if (entries.empty()) return;
  const size_t n_units = (entries.size() + DP_JIT_MAX_PATTERNS - 1) / DP_JIT_MAX_PATTERNS;
  const size_t per = (entries.size() + n_units - 1) / n_units;
  for (size_t at = 0; at < entries.size(); at += per) {
    std::vector<dp_jit_entry> part(entries.begin() + at, entries.begin() + std::min(entries.size(), at + per));
    dp_jit_unit u;
    u.source = unit_source(pats, letters, part);
    for (const dp_jit_entry& e : part) u.patterns.push_back(e.pattern);
    u.kernel = nullptr;
    units.push_back(u);
  }
}

int dp_jit_launch(dp_ctx* ctx, const dp_scan_args& A, const std::vector<dp_pattern>& pats, const dp_jit_unit& unit) {
  if (!unit.kernel || !unit.kernel->fn) return dp_set_error(ctx, DP_ERR_STATE, "internal: scan kernel not compiled");
  static thread_local std::vector<dp_pattern> set;
  set.resize(unit.patterns.size());
  for (size_t i = 0; i < unit.patterns.size(); ++i) set[i] = pats[unit.patterns[i]];
  const uint64_t words = A.word_end - A.word_begin;
  const unsigned blocks = (unsigned)std::min<uint64_t>((words + DP_JIT_THREADS - 1) / DP_JIT_THREADS,
                                                       (uint64_t)ctx->sm_count * unit.kernel->per_sm);
  void* args[] = {(void*)&A, (void*)set.data()};
  DP_CUDA(ctx, cudaLaunchKernel((const void*)unit.kernel->fn, dim3(blocks), dim3(DP_JIT_THREADS), args, 0, ctx->stream));
  return DP_OK;
}
