// Run-time compiled scan kernels (dp_jit.cu).
#pragma once
#include <string>
#include <vector>

#include "dp_scan.cuh"

#define DP_JIT_THREADS 256         // default block size; DEGENPRIMER_B200_JIT_THREADS overrides (tuning)
#define DP_JIT_MAX_PATTERNS 48     // per launch: bounds code size and the parameter block

struct dp_jit_kernel {
  cudaLibrary_t lib;
  cudaKernel_t fn;       // null: the compile failed, `log` says why
  int per_sm;            // resident CTAs per SM
  int regs;
  int threads;           // block size the source was generated for
  std::string log;
};

// one prefiltered pattern of a call: which pattern, where its anchor starts, how many
// hard mismatches the anchor may have (J = 8 + 2K letters)
struct dp_jit_entry {
  uint32_t pattern;
  uint32_t anchor;
  int K;
};

// one launch: source text + the patterns it was generated for (kernel parameter order)
struct dp_jit_unit {
  std::string source;
  std::vector<uint32_t> patterns;
  int threads;               // block size (NT of the source)
  dp_jit_kernel* kernel;     // null until compiled / found in the cache
};

// Generates the launch units for the prefiltered patterns of a call.
void dp_jit_make_units(const std::vector<dp_pattern>& pats, const std::vector<std::vector<uint8_t>>& letters,
                       const std::vector<dp_jit_entry>& entries, std::vector<dp_jit_unit>& units);
std::string dp_jit_compile_only(const std::string& source, std::string& log);
dp_jit_kernel* dp_jit_lookup(const std::string& source);
int dp_jit_compile(dp_ctx* ctx, const std::string& source, int threads, dp_jit_kernel** out);
int dp_jit_launch(dp_ctx* ctx, const dp_scan_args& A, const std::vector<dp_pattern>& pats, const dp_jit_unit& unit);
