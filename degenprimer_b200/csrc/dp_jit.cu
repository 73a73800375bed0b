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

int dp_jit_compile(dp_ctx* ctx, const std::string& source, int threads, dp_jit_kernel** out) {
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
  k->threads = threads;
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
      if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, (const void*)k->fn, threads, 0) == cudaSuccess && per_sm > 0)
        k->per_sm = per_sm;
      else
        k->per_sm = k->regs ? std::max(1, std::min(8, 65536 / (k->regs * threads))) : 4;
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
// The prefilter of a pattern (K = hard mismatches a hit can have) is a GROUP COUNT: some
// of its positions are dealt into G disjoint groups; a group matches iff every one of its
// positions matches.  K mismatches spoil at most K groups, so a hit has at least G-K
// matching groups -- a necessary condition, evaluated for 32 windows at once:
//   group i   = AND of the match planes of its positions
//   alive     = [number of matching groups >= G-K]   (carry-save adders + a small comparator)
// G = K+1 is the pigeonhole filter (alive = OR of the groups), groups of one position are the
// plain mismatch count; in between lies the cheapest form for K >= 2: e.g. L = 18, K = 3 ->
// 6 groups of 3 positions, alive = "at least 3 of 6" = 12 LOP3 per pattern and 32 windows
// against 14 planes + a 14-input adder (~42) for the anchor count.  The generator picks G and
// the positions per pattern by estimated instruction count (hot path + expected verify work).
//
// Match planes: the plane of (position j, letter set S) is one LOP3 of the two shifted code
// words (h_j, l_j).  With many patterns (a multiplex) every plane that is used is computed
// ONCE per word and kept in a register ("shared planes"): a group of three positions is then
// a single 3-input AND.  With few patterns the planes are folded into the AND chain
// (acc &= f(h_j, l_j): one LOP3 per position, nothing kept).
//
// Alive windows are not verified where they are found: they go to a per-warp queue and are
// verified 32 at a time (dp_queue_push / dp_verify_warp in dp_device.cuh).
//
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

static int jit_env_int(const char* name, int dflt) {
  const char* v = getenv(name);
  return v && *v ? atoi(v) : dflt;
}

// expected instructions per word and pattern spent on alive windows: the warp-wide vote and
// queue push when any of the warp's 1024 windows is alive, plus the batched verify per candidate
static double cold_cost(double p) {
  const double any = 1.0 - pow(1.0 - std::min(p, 0.5), 1024.0);
  return 30.0 * any + 1024.0 * p * 14.0;
}

// Emits the threshold function [x_0 + ... + x_{n-1} >= T] over one-bit planes as C statements.
// Full adders reduce three planes of one weight to a sum and a carry; with the sums W and the
// carries C of one level,  [sum >= T] = OR_j ([C-count >= j] & [W-count >= T - 2j]), which
// recurses on two shorter lists.  Unused sums are dead code for the compiler.
struct thr_emitter {
  std::string code, indent, prefix;
  int n_tmp = 0, ops = 0;
  std::map<std::string, std::string> memo;

  std::string tmp(const std::string& expr, int cost = 1) {
    auto it = memo.find(expr);
    if (it != memo.end()) return it->second;
    const std::string name = prefix + std::to_string(n_tmp++);
    code += indent + "const uint32_t " + name + " = " + expr + ";\n";
    ops += cost;
    memo[expr] = name;
    return name;
  }
  static std::string join(const std::vector<std::string>& x, const char* op) {
    std::string s = "(";
    for (size_t i = 0; i < x.size(); ++i) s += (i ? std::string(" ") + op + " " : std::string()) + x[i];
    return s + ")";
  }
  std::string ge(const std::vector<std::string>& x, int T) {
    const int n = (int)x.size();
    if (T <= 0) return "0xffffffffu";
    if (T > n) return "0u";
    if (n == 1) return x[0];
    if (T == 1) return tmp(join(x, "|"), n / 2);
    if (T == n) return tmp(join(x, "&"), n / 2);
    if (n == 3) return tmp("dp_maj(" + x[0] + ", " + x[1] + ", " + x[2] + ")");
    std::vector<std::string> W, C;
    for (int i = 0; i + 3 <= n; i += 3) {
      const std::string &a = x[i], &b = x[i + 1], &c = x[i + 2];
      W.push_back(tmp("dp_xor3(" + a + ", " + b + ", " + c + ")"));
      C.push_back(tmp("dp_maj(" + a + ", " + b + ", " + c + ")"));
    }
    for (int i = n - n % 3; i < n; ++i) W.push_back(x[i]);
    std::vector<std::string> terms;
    for (int j = 0; j <= (int)C.size(); ++j) {
      const int t1 = T - 2 * j;
      if (t1 > (int)W.size()) continue;
      const std::string cj = ge(C, j);
      if (cj == "0u") break;
      if (t1 <= 0) { terms.push_back(cj); break; }            // larger j are implied by this term
      const std::string wj = ge(W, t1);
      if (wj == "0u") continue;
      terms.push_back(cj == "0xffffffffu" ? wj : "(" + cj + " & " + wj + ")");
    }
    if (terms.empty()) return "0u";
    if (terms.size() == 1 && terms[0][0] != '(') return terms[0];
    return tmp(join(terms, "|"), (int)terms.size());
  }
};

// P(at least `need` of the independent events q[] happen)
static double tail_probability(const std::vector<double>& q, int need) {
  std::vector<double> d(q.size() + 1, 0.0);
  d[0] = 1.0;
  for (size_t i = 0; i < q.size(); ++i)
    for (size_t m = i + 1; m-- > 0;) {
      d[m + 1] += d[m] * q[i];
      d[m] *= 1.0 - q[i];
    }
  double s = 0.0;
  for (size_t m = (size_t)std::max(0, need); m < d.size(); ++m) s += d[m];
  return s;
}

struct unit_pattern {
  dp_jit_entry e;
  std::vector<std::vector<int>> groups;     // window positions of every group
  double p_alive;
};

// Chooses the groups of one pattern.  shared: planes come from registers (a group of d positions
// costs (d-1)/2 rounded up three-input ANDs), else one LOP3 per position.
static bool choose_groups(const std::vector<uint8_t>& lt, int K, bool shared, unit_pattern& out) {
  std::vector<int> order;
  for (int j = 0; j < (int)lt.size() && j < 64; ++j)
    if (set_fraction(dp_letter_set(lt[j])) < 1.0) order.push_back(j);
  std::stable_sort(order.begin(), order.end(), [&](int x, int y) {
    return set_fraction(dp_letter_set(lt[x])) < set_fraction(dp_letter_set(lt[y]));
  });
  const int n_avail = (int)order.size();
  double best = -1.0;
  for (int G = K + 1; G <= std::min(n_avail, 16); ++G) {
    thr_emitter dry;
    std::vector<std::string> names;
    for (int i = 0; i < G; ++i) names.push_back("g" + std::to_string(i));
    dry.ge(names, G - K);
    for (int n_use = G; n_use <= n_avail; ++n_use) {
      std::vector<std::vector<int>> g(G);
      for (int i = 0; i < n_use; ++i) g[i % G].push_back(order[i]);
      std::vector<double> q;
      double hot = dry.ops;
      for (auto& v : g) {
        double x = 1.0;
        for (int j : v) x *= set_fraction(dp_letter_set(lt[j]));
        q.push_back(x);
        hot += shared ? (double)(v.size() / 2) : (double)v.size();
      }
      const double p = tail_probability(q, G - K);
      const double cost = hot + cold_cost(p);
      if (best < 0 || cost < best - 1e-9) {
        best = cost;
        out.groups = g;
        out.p_alive = p;
      }
    }
  }
  if (best < 0) return false;
  for (auto& v : out.groups) std::sort(v.begin(), v.end());
  return out.p_alive <= 0.02;           // a filter that lets more through is not worth a launch
}

static std::string plane_name(int j, uint32_t set) { return "p" + std::to_string(j) + "_" + std::to_string(set & 15u); }

static int jit_threads() {
  const int t = jit_env_int("DEGENPRIMER_B200_JIT_THREADS", DP_JIT_THREADS);
  return (t == 64 || t == 128 || t == 256 || t == 512) ? t : DP_JIT_THREADS;
}

static std::string unit_source(const std::vector<dp_pattern>& pats, const std::vector<std::vector<uint8_t>>& letters,
                               const std::vector<dp_jit_entry>& entries, int unit_index) {
  (void)pats;
  const int shared_from = jit_env_int("DEGENPRIMER_B200_JIT_SHARED_FROM", 12);
  const bool shared = (int)entries.size() >= shared_from;
  std::vector<unit_pattern> ups;
  bool used[64] = {false};
  std::map<std::string, std::pair<int, uint32_t>> planes;       // shared planes by name
  for (const dp_jit_entry& e : entries) {
    unit_pattern u;
    u.e = e;
    u.p_alive = 1.0;
    const std::vector<uint8_t>& lt = letters[e.pattern];
    if (!choose_groups(lt, e.K, shared, u)) {
      // cannot happen for patterns the plan accepted (they have a selective anchor); keep the
      // kernel correct anyway: one group per position of the anchor
      u.groups.clear();
      for (int j = 0; j < 8 + 2 * e.K; ++j) u.groups.push_back(std::vector<int>(1, (int)e.anchor + j));
    }
    for (auto& v : u.groups)
      for (int j : v) {
        used[j] = true;
        if (shared) planes[plane_name(j, dp_letter_set(lt[j]))] = std::make_pair(j, dp_letter_set(lt[j]));
      }
    ups.push_back(u);
  }
  int max_off = 0;
  for (int j = 1; j < 64; ++j)
    if (used[j]) max_off = j;
  const bool three = max_off >= 32;        // positions 32..63 need the third word
  const int min_blocks = jit_env_int("DEGENPRIMER_B200_JIT_MIN_BLOCKS", shared ? 1 : 3);

  std::string s;
  s.reserve(8192 + 1024 * entries.size());
  // the plan, readable by dp_jit_describe and by whoever looks at the source
  for (const unit_pattern& u : ups) {
    char head[96];
    snprintf(head, sizeof head, "//PLAN pattern %u K %d need %d p_alive %.3g groups", u.e.pattern, u.e.K,
             (int)u.groups.size() - u.e.K, u.p_alive);
    s += head;
    for (auto& v : u.groups) {
      s += " ";
      for (size_t t = 0; t < v.size(); ++t) s += (t ? "," : "") + std::to_string(v[t]);
    }
    s += "\n";
  }
  s += std::string("//PLAN mode ") + (shared ? "shared-planes" : "folded-planes") + "\n";
  s += "#include \"dp_device.cuh\"\n";
  // The three-input functions of the group / threshold network as explicit LOP3: left to itself the compiler
  // re-associates the and/or/xor trees of a 40-pattern kernel into more instructions than the source has statements
  // (713 -> 524 LOP3 in the word loop of config 4).  DEGENPRIMER_B200_JIT_LOP3=0: the same functions in plain C.
  if (jit_env_int("DEGENPRIMER_B200_JIT_LOP3", 1) != 0)
    s += "#define DP_LOP3(name, lut) __device__ __forceinline__ uint32_t name(uint32_t a, uint32_t b, uint32_t c) { \\\n"
         "  uint32_t r; asm(\"lop3.b32 %0, %1, %2, %3, \" #lut \";\" : \"=r\"(r) : \"r\"(a), \"r\"(b), \"r\"(c)); return r; }\n"
         "DP_LOP3(dp_and3, 0x80)\nDP_LOP3(dp_xor3, 0x96)\nDP_LOP3(dp_maj, 0xE8)\n";
  else
    s += "__device__ __forceinline__ uint32_t dp_and3(uint32_t a, uint32_t b, uint32_t c) { return a & b & c; }\n"
         "__device__ __forceinline__ uint32_t dp_xor3(uint32_t a, uint32_t b, uint32_t c) { return a ^ b ^ c; }\n"
         "__device__ __forceinline__ uint32_t dp_maj(uint32_t a, uint32_t b, uint32_t c) { return (a & b) | (c & (a | b)); }\n";
  s += "static_assert(sizeof(dp_pattern) == " + std::to_string(sizeof(dp_pattern)) + ", \"dp_pattern layout\");\n";
  s += "static_assert(sizeof(dp_scan_args) == " + std::to_string(sizeof(dp_scan_args)) + ", \"dp_scan_args layout\");\n";
  s += "static_assert(sizeof(dp_raw_hit) == " + std::to_string(sizeof(dp_raw_hit)) + ", \"dp_raw_hit layout\");\n";
  s += "#define NP " + std::to_string(entries.size()) + "\n";
  s += "#define NT " + std::to_string(jit_threads()) + "\n";
  s += "__device__ const uint8_t dp_jit_pattern_of[NP] = {";
  for (size_t i = 0; i < ups.size(); ++i) s += (i ? ", " : "") + std::to_string(ups[i].e.pattern);
  s += "};\n";
  s += "extern \"C\" __global__ void __launch_bounds__(NT, " + std::to_string(min_blocks) + ")\n"
       "dp_jit_scan(const __grid_constant__ dp_scan_args A, const __grid_constant__ dp_pattern_set<NP> PS) {\n"
       "  // patterns and the candidate queues of the block's warps\n"
       "  __shared__ dp_pattern s_pat[NP];\n"
       "  __shared__ unsigned long long s_queue[NT / 32][DP_Q_CAP];\n"
       "  for (uint32_t i = threadIdx.x; i < NP * (uint32_t)(sizeof(dp_pattern) / 4); i += NT)\n"
       "    reinterpret_cast<uint32_t*>(s_pat)[i] = reinterpret_cast<const uint32_t*>(PS.pat)[i];\n"
       "  __syncthreads();\n"
       "  unsigned long long* const q = s_queue[threadIdx.x >> 5];\n"
       "  uint32_t qn = 0;\n"
       "  const uint64_t stride = (uint64_t)gridDim.x * NT;\n"
       "  unsigned long long n_cand = 0;\n"
       "  // Work is handed out per WARP in chunks of consecutive words from a counter in global memory\n"
       "  // (zeroed by the previous call's publish kernel), not by a fixed stride: a block that starts\n"
       "  // late -- its SM was still held by a small kernel of another call in flight -- then simply\n"
       "  // takes fewer chunks instead of stretching the whole launch by its full share.  No block barrier:\n"
       "  // with one resident block per SM a barrier per chunk would idle the integer pipe.  The first\n"
       "  // 15/16 of the words go out in chunks of DP_JIT_CHUNK x 32, the rest in chunks of\n"
       "  // DP_JIT_CHUNK_TAIL x 32, so that the warps finish within a few microseconds of each other;\n"
       "  // the index of the next chunk is asked for before the current one is worked on.\n"
       "  const uint32_t lane = threadIdx.x & 31u;\n"
       "  const uint64_t n_words = A.word_end > A.word_begin ? A.word_end - A.word_begin : 0;\n"
       "  const uint64_t big_words = (uint64_t)DP_JIT_CHUNK * 32u, small_words = (uint64_t)DP_JIT_CHUNK_TAIL * 32u;\n"
       "  const uint64_t n_big = (n_words - n_words / 16) / big_words;\n"
       "  const uint64_t big_total = n_big * big_words;\n"
       "  const uint64_t n_chunks = n_big + (n_words - big_total + small_words - 1) / small_words;\n"
       "  unsigned long long* const next_chunk = &A.counters[DP_C_CHUNK0 + " + std::to_string(unit_index) + "];\n"
       "  unsigned long long chunk = 0;\n"
       "  if (lane == 0) chunk = atomicAdd(next_chunk, 1ull);\n"
       "  chunk = __shfl_sync(0xffffffffu, chunk, 0);\n"
       "  while (chunk < n_chunks) {\n"
       "  unsigned long long after = 0;\n"
       "  if (lane == 0) after = atomicAdd(next_chunk, 1ull);        // in flight while this chunk is scanned\n"
       "  const uint64_t c_begin = chunk < n_big ? chunk * big_words : big_total + (chunk - n_big) * small_words;\n"
       "  const uint64_t w_end = min(A.word_end, A.word_begin + c_begin + (chunk < n_big ? big_words : small_words));\n"
       "  uint64_t w = A.word_begin + c_begin + lane;\n"
       "  uint32_t h0 = 0, hn = 0, hm = 0, l0 = 0, ln = 0, lm = 0;\n"
       "  if (w < w_end) {\n"
       "    h0 = __ldg(A.hi + w); hn = __ldg(A.hi + w + 1);\n"
       "    l0 = __ldg(A.lo + w); ln = __ldg(A.lo + w + 1);\n";
  if (three) s += "    hm = __ldg(A.hi + w + 2); lm = __ldg(A.lo + w + 2);\n";
  s += "  }\n"
       "  // the loop is uniform over the warp (its first lane decides): the queue is warp-collective\n"
       "  while (w - (threadIdx.x & 31u) < w_end) {\n"
       "    const uint32_t live = w < w_end ? 0xffffffffu : 0u;\n"
       "    const uint64_t wn = w + 32u;\n"
       "    uint32_t ph0 = 0, phn = 0, phm = 0, pl0 = 0, pln = 0, plm = 0;\n"
       "    if (wn < w_end) {\n"
       "      ph0 = __ldg(A.hi + wn); phn = __ldg(A.hi + wn + 1);\n"
       "      pl0 = __ldg(A.lo + wn); pln = __ldg(A.lo + wn + 1);\n";
  if (three) s += "      phm = __ldg(A.hi + wn + 2); plm = __ldg(A.lo + wn + 2);\n";
  s += "    }\n";
  // Shared planes, "letter planes first" (default): the plane of letter set S is made ONCE per template word
  // (one LOP3 on the unshifted code words of this word and of the next) and the plane of (position j, S) is one
  // funnel shift of those -- instead of two funnel shifts per position (both code words) plus one LOP3 per plane.
  const bool shift_planes = shared && jit_env_int("DEGENPRIMER_B200_JIT_SHIFT_PLANES", 1) != 0;
  if (!shift_planes)
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
  if (shared && !shift_planes)
    for (auto& kv : planes) {
      const std::string n = std::to_string(kv.second.first);
      s += "    const uint32_t " + kv.first + " = " + plane_expr(kv.second.second, "h" + n, "l" + n) + ";\n";
    }
  if (shift_planes) {
    std::map<uint32_t, int> sets;           // letter set -> highest word it is needed in (0, 1 = next, 2 = third)
    for (auto& kv : planes) {
      const int j = kv.second.first;
      int& top = sets[kv.second.second];
      top = std::max(top, j == 0 ? 0 : j <= 32 ? 1 : 2);
    }
    for (auto& kv : sets) {
      const std::string n = "s" + std::to_string(kv.first);
      s += "    const uint32_t " + n + "_0 = " + plane_expr(kv.first, "h0", "l0");
      if (kv.second >= 1) s += ", " + n + "_n = " + plane_expr(kv.first, "hn", "ln");
      if (kv.second >= 2) s += ", " + n + "_m = " + plane_expr(kv.first, "hm", "lm");
      s += ";\n";
    }
    for (auto& kv : planes) {
      const int j = kv.second.first;
      const std::string n = "s" + std::to_string(kv.second.second);
      std::string e;
      // (an unshifted plane takes the live-lane mask along: as a two-input function of the code words ptxas would
      // otherwise re-derive it inside every group that uses it -- one more LOP3 per pattern)
      if (j == 0) e = n + "_0 & live";
      else if (j < 32) e = "__funnelshift_r(" + n + "_0, " + n + "_n, " + std::to_string(j) + ")";
      else if (j == 32) e = n + "_n & live";
      else e = "__funnelshift_r(" + n + "_n, " + n + "_m, " + std::to_string(j - 32) + ")";
      s += "    const uint32_t " + kv.first + " = " + e + ";\n";
    }
  }
  // patterns in chunks of 8: one vote + branch per chunk on the hot path
  for (size_t c0 = 0; c0 < ups.size(); c0 += 8) {
    const size_t c1 = std::min(ups.size(), c0 + 8);
    s += "    {\n";
    for (size_t i = c0; i < c1; ++i) {
      const unit_pattern& u = ups[i];
      const std::vector<uint8_t>& lt = letters[u.e.pattern];
      const std::string v = "alive" + std::to_string(i);
      s += "      uint32_t " + v + ";\n      {\n";
      std::vector<std::string> gnames;
      for (size_t g = 0; g < u.groups.size(); ++g) {
        const std::string acc = "g" + std::to_string(g);
        gnames.push_back(acc);
        if (shared) {
          std::vector<std::string> in;
          for (int j : u.groups[g]) in.push_back(plane_name(j, dp_letter_set(lt[j])));
          std::string e;
          if (shift_planes) {
            // chains of explicit three-input ANDs; an even number of planes leaves one input free and the
            // live-lane mask fills it
            if (in.size() % 2 == 0) in.push_back("live");
            e = in[0];
            for (size_t t = 1; t + 1 < in.size(); t += 2) e = "dp_and3(" + e + ", " + in[t] + ", " + in[t + 1] + ")";
          } else {
            for (size_t t = 0; t < in.size(); ++t) e += (t ? " & " : "") + in[t];
          }
          s += "        const uint32_t " + acc + " = " + e + ";\n";
        } else {
          bool first = true;
          for (int j : u.groups[g]) {
            const std::string n = std::to_string(j);
            const std::string pe = plane_expr(dp_letter_set(lt[j]), "h" + n, "l" + n);
            s += first ? "        uint32_t " + acc + " = " + pe + ";\n" : "        " + acc + " &= " + pe + ";\n";
            first = false;
          }
        }
      }
      thr_emitter em;
      em.indent = "        ";
      em.prefix = "t";
      const std::string res = em.ge(gnames, (int)u.groups.size() - u.e.K);
      s += em.code;
      s += "        " + v + " = " + res + ";\n      }\n";
    }
    s += "      if (__any_sync(0xffffffffu, ((alive" + std::to_string(c0);
    for (size_t i = c0 + 1; i < c1; ++i) s += " | alive" + std::to_string(i);
    // lanes past the end of the chunk hold zeros, not template words: their bits are dropped here, once per
    // group of patterns, and again where a pattern's word is handed to the queue
    s += ") & live) != 0u)) {\n";
    for (size_t i = c0; i < c1; ++i)
      s += "        if (__any_sync(0xffffffffu, (alive" + std::to_string(i) + " & live) != 0u)) qn = dp_queue_push(A, s_pat, dp_jit_pattern_of, q, qn, alive" +
           std::to_string(i) + " & live, w, " + std::to_string(i) + "u, n_cand);\n";
    s += "      }\n    }\n";
  }
  s += "    w = wn;\n"
       "    h0 = ph0; hn = phn; hm = phm; l0 = pl0; ln = pln; lm = plm;\n"
       "  }\n"
       "  chunk = __shfl_sync(0xffffffffu, after, 0);\n"
       "  }\n"
       "  if (qn) dp_verify_warp(A, s_pat, dp_jit_pattern_of, q[threadIdx.x & 31u], (threadIdx.x & 31u) < qn);\n"
       "  if (n_cand) atomicAdd(&A.counters[DP_C_CAND], n_cand);\n"
       "  // windows that touch a non-ACGT base (the prefilter skips them): enumerated from the exception-run table\n"
       "  // for this unit's patterns, as dp_scan_runs_kernel does for the generic kernel -- no extra launch\n"
       "  dp_scan_runs_device(A, s_pat, dp_jit_pattern_of, NP, (uint64_t)blockIdx.x * NT + threadIdx.x, stride);\n"
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
    u.source = unit_source(pats, letters, part, (int)units.size());
    u.threads = jit_threads();
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
  const uint64_t nt = (uint64_t)unit.kernel->threads;
  const unsigned blocks = (unsigned)std::min<uint64_t>((words + nt - 1) / nt, (uint64_t)ctx->sm_count * unit.kernel->per_sm);
  void* args[] = {(void*)&A, (void*)set.data()};
  DP_CUDA(ctx, cudaLaunchKernel((const void*)unit.kernel->fn, dim3(blocks), dim3((unsigned)nt), args, 0, ctx->stream));
  return DP_OK;
}
