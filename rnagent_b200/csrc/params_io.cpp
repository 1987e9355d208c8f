// params_io.cpp -- loader for "RNAfold parameter file v2.0" energy parameter files.
//
// Replaces RNA.read_parameter_file (reference call sites rlif/rna/vienna.py:18, rlif/CLI.py:21,
// rlif/interface/widgets.py:31) and the 37 C rescaling ViennaRNA performs on every fold
// (SURVEY.md section 8a' row F2, Appendix A).  Product code: does not depend on oracle/.
#include "fold_params.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <unordered_map>
#include <vector>

namespace rnafold {
namespace {

struct Section {
  std::string name;
  std::vector<std::string> tok;
};

int token_value(const std::string &t) {
  if (t == "INF") return kInf;
  if (t == "DEF") return -50;
  if (t == "NST") return 0;
  return (int)std::strtol(t.c_str(), nullptr, 10);
}

// Split the file into sections; C comments are removed first (they may sit on their own line or
// trail a row of numbers and may contain anything but "*/").
std::vector<Section> split_sections(const char *text, size_t len) {
  std::string s(text, len);
  for (size_t p = s.find("/*"); p != std::string::npos; p = s.find("/*", p)) {
    size_t e = s.find("*/", p + 2);
    size_t stop = (e == std::string::npos) ? s.size() : e + 2;
    for (size_t k = p; k < stop; ++k)
      if (s[k] != '\n') s[k] = ' ';
    p = stop;
  }
  std::vector<Section> out;
  size_t pos = 0;
  bool ended = false;
  while (pos < s.size() && !ended) {
    size_t eol = s.find('\n', pos);
    if (eol == std::string::npos) eol = s.size();
    std::string line = s.substr(pos, eol - pos);
    pos = eol + 1;
    if (!line.empty() && line[0] == '#') {
      if (line.size() > 1 && line[1] == '#') continue;  // "## RNAfold parameter file v2.0"
      size_t b = line.find_first_not_of("# \t");
      std::string name = (b == std::string::npos) ? "" : line.substr(b);
      size_t e = name.find_first_of(" \t\r");
      if (e != std::string::npos) name.resize(e);
      if (name == "END") {
        ended = true;
        break;
      }
      out.push_back(Section{name, {}});
      continue;
    }
    if (out.empty()) continue;
    size_t k = 0;
    while (k < line.size()) {
      while (k < line.size() && std::isspace((unsigned char)line[k])) ++k;
      size_t b = k;
      while (k < line.size() && !std::isspace((unsigned char)line[k])) ++k;
      if (k > b) out.back().tok.emplace_back(line, b, k - b);
    }
  }
  return out;
}

bool fill(const Section &sec, int32_t *dst, const std::vector<int> &index, char *err, size_t errlen) {
  // index[k] = flat destination of the k-th value
  if (sec.tok.size() < index.size()) {
    std::snprintf(err, errlen, "section '%s' holds %zu values, expected %zu", sec.name.c_str(),
                  sec.tok.size(), index.size());
    return false;
  }
  for (size_t k = 0; k < index.size(); ++k) dst[index[k]] = token_value(sec.tok[k]);
  return true;
}

// Destination index lists for ViennaRNA's table shapes (pair types 1..7, bases 0..4).
std::vector<int> idx_pairs_bases(int nb) {  // [8][5]^nb with pairs 1..7
  std::vector<int> v;
  int inner = 1;
  for (int k = 0; k < nb; ++k) inner *= 5;
  for (int p = 1; p <= 7; ++p)
    for (int r = 0; r < inner; ++r) v.push_back(p * inner + r);
  return v;
}

std::vector<int> idx_pairpair_bases(int nb) {  // [8][8][5]^nb with pairs 1..7 x 1..7
  std::vector<int> v;
  int inner = 1;
  for (int k = 0; k < nb; ++k) inner *= 5;
  for (int p = 1; p <= 7; ++p)
    for (int q = 1; q <= 7; ++q)
      for (int r = 0; r < inner; ++r) v.push_back((p * 8 + q) * inner + r);
  return v;
}

uint32_t pack_key(const char *s, int n) {
  uint32_t key = 0;
  for (int k = 0; k < n; ++k) {
    uint32_t c;
    switch (s[k]) {
      case 'A': c = 1; break;
      case 'C': c = 2; break;
      case 'G': c = 3; break;
      case 'U': c = 4; break;
      default: c = 6; break;  // never produced by a sequence window (those use 7 for non-ACGU)
    }
    key |= c << (3 * k);
  }
  return key;
}

// int22 entries that touch base N: maximum over all A/C/G/U completions (ViennaRNA fills these
// after reading the 6x6x4x4x4x4 block; only reachable by sequences with non-ACGU letters).
void complete_int22(int32_t a[8][8][5][5][5][5]) {
  for (int i = 1; i <= 6; ++i)
    for (int j = 1; j <= 6; ++j)
      for (int code = 0; code < 625; ++code) {
        int d[4] = {code / 125, (code / 25) % 5, (code / 5) % 5, code % 5};
        if (d[0] && d[1] && d[2] && d[3]) continue;
        int best = -kInf;
        for (int sub = 0; sub < 256; ++sub) {
          int e[4] = {(sub >> 6) & 3, (sub >> 4) & 3, (sub >> 2) & 3, sub & 3};
          bool ok = true;
          int x[4];
          for (int k = 0; k < 4; ++k) {
            x[k] = d[k] ? d[k] : e[k] + 1;
            if (d[k] && e[k] != 0) ok = false;  // fixed digits enumerate once
          }
          if (!ok) continue;
          int v = a[i][j][x[0]][x[1]][x[2]][x[3]];
          if (v > best) best = v;
        }
        a[i][j][d[0]][d[1]][d[2]][d[3]] = best;
      }
}

}  // namespace

void finalize_params(HostParams *hp) {
  SmallParams &s = hp->dev.s;
  std::memcpy(hp->raw_mmM, s.mmM, sizeof(s.mmM));
  std::memcpy(hp->raw_mmExt, s.mmExt, sizeof(s.mmExt));
  std::memcpy(hp->raw_dangle5, s.dangle5, sizeof(s.dangle5));
  std::memcpy(hp->raw_dangle3, s.dangle3, sizeof(s.dangle3));
  hp->raw_valid = 1;
  for (int p = 0; p < 8; ++p)
    for (int a = 0; a < 5; ++a) {
      if (s.dangle5[p][a] > 0) s.dangle5[p][a] = 0;
      if (s.dangle3[p][a] > 0) s.dangle3[p][a] = 0;
      for (int b = 0; b < 5; ++b) {
        if (s.mmM[p][a][b] > 0) s.mmM[p][a][b] = 0;
        if (s.mmExt[p][a][b] > 0) s.mmExt[p][a][b] = 0;
      }
    }
  for (int u = 0; u <= kMaxLen; ++u) {
    if (u <= 30) hp->dev.hairpin_len[u] = s.hairpin[u];
    else hp->dev.hairpin_len[u] = s.hairpin[30] + (int)(hp->lxc * std::log(u / 30.));
  }
  for (int k = 0; k < kMaxSpecial; ++k) {
    hp->dev.tetra_key[k] = k < s.n_tetra ? pack_key(hp->tetra[k], 6) : 0xffffffffu;
    hp->dev.tri_key[k] = k < s.n_tri ? pack_key(hp->tri[k], 5) : 0xffffffffu;
    hp->dev.hexa_key[k] = k < s.n_hexa ? pack_key(hp->hexa[k], 8) : 0xffffffffu;
  }
  s.lxc_milli = (int32_t)std::lround(hp->lxc * 1000.0);
}

int parse_parameter_text(const char *text, size_t len, HostParams *hp, char *err, size_t errlen) {
  static const char kMagic[] = "## RNAfold parameter file v2.0";
  if (len < sizeof(kMagic) - 1 || std::strncmp(text, kMagic, sizeof(kMagic) - 1) != 0) {
    std::snprintf(err, errlen, "not an 'RNAfold parameter file v2.0'");
    return -1;
  }
  FoldParams &d = hp->dev;
  SmallParams &s = d.s;
  if (hp->raw_valid) {  // sections absent from this file keep the live values: un-clamp them first
    std::memcpy(s.mmM, hp->raw_mmM, sizeof(s.mmM));
    std::memcpy(s.mmExt, hp->raw_mmExt, sizeof(s.mmExt));
    std::memcpy(s.dangle5, hp->raw_dangle5, sizeof(s.dangle5));
    std::memcpy(s.dangle3, hp->raw_dangle3, sizeof(s.dangle3));
  }
  const std::vector<int> i_mm = idx_pairs_bases(2), i_dg = idx_pairs_bases(1);
  std::vector<int> i_stack;
  for (int p = 1; p <= 7; ++p)
    for (int q = 1; q <= 7; ++q) i_stack.push_back(p * 8 + q);
  std::vector<int> i_lin(31);
  for (int k = 0; k < 31; ++k) i_lin[k] = k;
  std::vector<int> i_22;
  for (int p = 1; p <= 6; ++p)
    for (int q = 1; q <= 6; ++q)
      for (int a = 1; a <= 4; ++a)
        for (int b = 1; b <= 4; ++b)
          for (int c = 1; c <= 4; ++c)
            for (int e = 1; e <= 4; ++e) i_22.push_back(((((p * 8 + q) * 5 + a) * 5 + b) * 5 + c) * 5 + e);

  std::unordered_map<std::string, int32_t *> mm = {
      {"mismatch_hairpin", &s.mmH[0][0][0]},      {"mismatch_interior", &s.mmI[0][0][0]},
      {"mismatch_interior_1n", &s.mm1nI[0][0][0]}, {"mismatch_interior_23", &s.mm23I[0][0][0]},
      {"mismatch_multi", &s.mmM[0][0][0]},        {"mismatch_exterior", &s.mmExt[0][0][0]}};

  for (const Section &sec : split_sections(text, len)) {
    const std::string &n = sec.name;
    bool ok = true;
    if (n == "stack") ok = fill(sec, &s.stack[0][0], i_stack, err, errlen);
    else if (mm.count(n)) ok = fill(sec, mm[n], i_mm, err, errlen);
    else if (n == "dangle5") ok = fill(sec, &s.dangle5[0][0], i_dg, err, errlen);
    else if (n == "dangle3") ok = fill(sec, &s.dangle3[0][0], i_dg, err, errlen);
    else if (n == "int11") ok = fill(sec, &d.int11[0][0][0][0], idx_pairpair_bases(2), err, errlen);
    else if (n == "int21") ok = fill(sec, &d.int21[0][0][0][0][0], idx_pairpair_bases(3), err, errlen);
    else if (n == "int22") {
      ok = fill(sec, &d.int22[0][0][0][0][0][0], i_22, err, errlen);
      if (ok) complete_int22(d.int22);
    } else if (n == "hairpin") ok = fill(sec, s.hairpin, i_lin, err, errlen);
    else if (n == "bulge") ok = fill(sec, s.bulge, i_lin, err, errlen);
    else if (n == "interior") ok = fill(sec, s.interior, i_lin, err, errlen);
    else if (n == "ML_params") {
      if (sec.tok.size() < 6) ok = false;
      else {
        s.MLbase = token_value(sec.tok[0]);
        s.MLclosing = token_value(sec.tok[2]);
        s.MLintern = token_value(sec.tok[4]);
      }
    } else if (n == "NINIO") {
      if (sec.tok.size() < 3) ok = false;
      else {
        s.ninio = token_value(sec.tok[0]);
        s.max_ninio = token_value(sec.tok[2]);
      }
    } else if (n == "Misc") {
      if (sec.tok.size() < 4) ok = false;
      else {
        s.DuplexInit = token_value(sec.tok[0]);
        s.TerminalAU = token_value(sec.tok[2]);
        if (sec.tok.size() >= 5) hp->lxc = std::strtod(sec.tok[4].c_str(), nullptr);
      }
    } else if (n == "Tetraloops" || n == "Triloops" || n == "Hexaloops") {
      int width = n[0] == 'H' ? 8 : (n[1] == 'e' ? 6 : 5);
      int cnt = (int)(sec.tok.size() / 3);
      if (cnt > kMaxSpecial) cnt = kMaxSpecial;
      for (int k = 0; k < cnt; ++k) {
        char *dst = width == 8 ? hp->hexa[k] : width == 6 ? hp->tetra[k] : hp->tri[k];
        std::memset(dst, 0, width == 8 ? 12 : 8);
        std::strncpy(dst, sec.tok[3 * k].c_str(), (size_t)width);
        int e = token_value(sec.tok[3 * k + 1]);
        (width == 8 ? d.hexa_E : width == 6 ? d.tetra_E : d.tri_E)[k] = e;
      }
      (width == 8 ? s.n_hexa : width == 6 ? s.n_tetra : s.n_tri) = cnt;
    }
    // *_enthalpies and unknown sections are ignored: unused at 37 C.
    if (!ok) {
      if (!err[0]) std::snprintf(err, errlen, "section '%s' is too short", n.c_str());
      return -1;
    }
  }
  finalize_params(hp);
  return 0;
}

int load_parameter_file(const char *path, HostParams *hp, char *err, size_t errlen) {
  std::FILE *f = std::fopen(path, "rb");
  if (!f) {
    std::snprintf(err, errlen, "cannot open parameter file '%s'", path);
    return -1;
  }
  std::string buf;
  char chunk[65536];
  size_t got;
  while ((got = std::fread(chunk, 1, sizeof(chunk), f)) > 0) buf.append(chunk, got);
  std::fclose(f);
  err[0] = 0;
  return parse_parameter_text(buf.data(), buf.size(), hp, err, errlen);
}

int export_params(const HostParams *hp, int32_t *out, int cap) {
  const FoldParams &d = hp->dev;
  const SmallParams &s = d.s;
  std::vector<int32_t> v;
  auto dump = [&](const int32_t *p, int n) { v.insert(v.end(), p, p + n); };
  dump(&s.stack[0][0], 64);
  dump(s.hairpin, 31);
  dump(s.bulge, 31);
  dump(s.interior, 31);
  dump(&s.mmH[0][0][0], 200);
  dump(&s.mmI[0][0][0], 200);
  dump(&s.mm1nI[0][0][0], 200);
  dump(&s.mm23I[0][0][0], 200);
  dump(&s.mmM[0][0][0], 200);
  dump(&s.mmExt[0][0][0], 200);
  dump(&s.dangle5[0][0], 40);
  dump(&s.dangle3[0][0], 40);
  dump(&d.int11[0][0][0][0], 8 * 8 * 25);
  dump(&d.int21[0][0][0][0][0], 8 * 8 * 125);
  dump(&d.int22[0][0][0][0][0][0], 8 * 8 * 625);
  v.push_back(s.MLbase);
  v.push_back(s.MLclosing);
  v.push_back(s.MLintern);
  v.push_back(s.ninio);
  v.push_back(s.max_ninio);
  v.push_back(s.DuplexInit);
  v.push_back(s.TerminalAU);
  v.push_back(s.lxc_milli);
  for (int cls = 0; cls < 3; ++cls) {
    int n = cls == 0 ? s.n_tetra : cls == 1 ? s.n_tri : s.n_hexa;
    int w = cls == 0 ? 6 : cls == 1 ? 5 : 8;
    v.push_back(n);
    for (int k = 0; k < kMaxSpecial; ++k) {
      const char *sq = cls == 0 ? hp->tetra[k] : cls == 1 ? hp->tri[k] : hp->hexa[k];
      for (int c = 0; c < 8; ++c) v.push_back((k < n && c < w) ? sq[c] : 0);
      v.push_back(k < n ? (cls == 0 ? d.tetra_E[k] : cls == 1 ? d.tri_E[k] : d.hexa_E[k]) : 0);
    }
  }
  if ((int)v.size() > cap) return -1;
  std::memcpy(out, v.data(), v.size() * sizeof(int32_t));
  return (int)v.size();
}

}  // namespace rnafold
