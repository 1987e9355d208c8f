// fold_params.h -- energy parameter tables as the kernels see them (plain data, host + device).
//
// Shapes follow ViennaRNA's vrna_param_t so that the loop-energy functions can index them the way
// E_Hairpin / E_IntLoop / E_MLstem / E_ExtLoop do (SURVEY.md section 8a' rows F2-F5): pair types
// 1..7 = CG GC GU UG AU UA NN (0 = no pair), bases 0..4 = N A C G U.  All energies are integer
// dcal/mol, already rescaled for 37 C with dangles=2 (dangles and multiloop/exterior mismatches
// clamped to <= 0).
#pragma once
#include <stddef.h>
#include <stdint.h>

namespace rnafold {

constexpr int kInf = 10000000;
constexpr int kTurn = 3;
constexpr int kMaxLoop = 30;
constexpr int kMaxSpecial = 200;  // ViennaRNA keeps up to 200 special hairpins per class
constexpr int kMaxLen = 4000;

// Small, hot tables: copied into shared memory by every CTA.
struct SmallParams {
  int32_t stack[8][8];
  int32_t hairpin[32];
  int32_t bulge[32];
  int32_t interior[32];
  int32_t mmH[8][5][5];
  int32_t mmI[8][5][5];
  int32_t mm1nI[8][5][5];
  int32_t mm23I[8][5][5];
  int32_t mmM[8][5][5];
  int32_t mmExt[8][5][5];
  int32_t dangle5[8][5];
  int32_t dangle3[8][5];
  int32_t MLbase, MLclosing, MLintern, ninio, max_ninio, TerminalAU, DuplexInit;
  int32_t n_tetra, n_tri, n_hexa;
  int32_t lxc_milli;  // lxc * 1000, informational (the log term is pre-evaluated on the host)
  int32_t pad_;
};

// Large tables: stay in global memory (read through the read-only path, L1/L2 resident).
struct FoldParams {
  SmallParams s;
  int32_t int11[8][8][5][5];
  int32_t int21[8][8][5][5][5];
  int32_t int22[8][8][5][5][5][5];
  // Special hairpins: key = packed base codes (3 bits each, first base in the low bits), in file order.
  uint32_t tetra_key[kMaxSpecial];
  uint32_t tri_key[kMaxSpecial];
  uint32_t hexa_key[kMaxSpecial];
  int32_t tetra_E[kMaxSpecial];
  int32_t tri_E[kMaxSpecial];
  int32_t hexa_E[kMaxSpecial];
  // hairpin_len[u] = unpaired-size term of E_Hairpin for every loop size 0..kMaxLen:
  // hairpin[u] for u <= 30, hairpin[30] + (int)(lxc * log(u / 30.)) above (host libm, truncation).
  int32_t hairpin_len[kMaxLen + 1];
};

// Host-side extras that never travel to the device.
struct HostParams {
  FoldParams dev;
  double lxc;
  char tetra[kMaxSpecial][8];
  char tri[kMaxSpecial][8];
  char hexa[kMaxSpecial][12];
  // dangle / multiloop- and exterior-mismatch tables as read from the file, before the <= 0 clamp of the MFE model
  // (the partition function passes them through ViennaRNA's SMOOTH() instead, csrc/ensemble.cu)
  int32_t raw_mmM[8][5][5], raw_mmExt[8][5][5], raw_dangle5[8][5], raw_dangle3[8][5];
  int32_t raw_valid;
};

// params_io.cpp
// Parse "RNAfold parameter file v2.0" text (comments, INF/DEF/NST tokens, sections in any order;
// sections that are absent keep the values already in *hp).  Returns 0 or a negative code and
// fills `err`.
int parse_parameter_text(const char *text, size_t len, HostParams *hp, char *err, size_t errlen);
int load_parameter_file(const char *path, HostParams *hp, char *err, size_t errlen);
// Recompute derived members (clamps, hairpin_len, keys) after the raw tables changed.
void finalize_params(HostParams *hp);
// Oracle-compatible flat dump (see oracle/rnafold_oracle.c orc_params_export).
int export_params(const HostParams *hp, int32_t *out, int cap);

}  // namespace rnafold
