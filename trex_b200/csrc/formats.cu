// File formats — SURVEY 8(f) rank 4.  The reference has NO on-disk or wire format (inputs exist only as in-memory
// arrays), so there is nothing to be parity-checked against: what is fixed here is the OUTPUT contract, which is the
// input contract of the hot path:
//   alignment (FASTA, PHYLIP)  ->  u8 states [taxon][site], 0..A-1 = index in the alphabet (case-insensitive),
//                                  255 = "no state" (gap, ambiguity code, unknown character): the leaf row that stays
//                                  all-1e5 in the reference (src/trex/sankoff.py:50, dropped out-of-bounds scatter);
//   tree (Newick)              ->  parent[] of a rooted binary tree in the ordering run_dp requires
//                                  (src/trex/sankoff.py:45-46, 87-92): leaves are 0..n-1 (in the order given by the
//                                  caller's taxon names, or of appearance), internal nodes n..2n-2 in post-order, so
//                                  every child index < its parent's and the root is the last node; adjacency[i][parent[i]] = 1
//                                  (src/trex/utils/types.py:33), no root self-loop.
// Pure host code (all pointers are HOST pointers: the `_host` suffix of include/trex_b200.h); no CUDA call is made.
#include <cctype>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"

namespace trex {
namespace {

struct Alphabet {
  uint8_t code[256];
  explicit Alphabet(const char* letters) {
    std::memset(code, 255, sizeof(code));
    for (int i = 0; letters && letters[i] && i < 255; ++i) {
      code[(unsigned char)std::toupper((unsigned char)letters[i])] = (uint8_t)i;
      code[(unsigned char)std::tolower((unsigned char)letters[i])] = (uint8_t)i;
    }
  }
};

struct Alignment {
  std::vector<std::string> names;
  std::vector<std::string> seqs;
};

bool parse_fasta(const char* text, size_t len, Alignment& out, std::string& err) {
  size_t i = 0;
  while (i < len) {
    size_t e = i;
    while (e < len && text[e] != '\n') ++e;
    size_t le = e;
    while (le > i && (text[le - 1] == '\r' || text[le - 1] == ' ' || text[le - 1] == '\t')) --le;
    if (le > i) {
      if (text[i] == '>') {
        size_t a = i + 1;
        while (a < le && (text[a] == ' ' || text[a] == '\t')) ++a;
        size_t b = a;
        while (b < le && !std::isspace((unsigned char)text[b])) ++b;  // the name is the first word of the header
        out.names.emplace_back(text + a, b - a);
        out.seqs.emplace_back();
      } else if (text[i] != ';') {  // ';' comment lines of old FASTA files
        if (out.seqs.empty()) { err = "FASTA: sequence data before the first '>' header"; return false; }
        for (size_t j = i; j < le; ++j)
          if (!std::isspace((unsigned char)text[j])) out.seqs.back().push_back(text[j]);
      }
    }
    i = e + 1;
  }
  if (out.names.empty()) { err = "FASTA: no records"; return false; }
  return true;
}

// PHYLIP, sequential or interleaved, relaxed names (first whitespace-delimited word) — strict 10-column names are
// accepted when the name contains no blank.
bool parse_phylip(const char* text, size_t len, Alignment& out, std::string& err) {
  std::vector<std::string> lines;
  size_t i = 0;
  while (i < len) {
    size_t e = i;
    while (e < len && text[e] != '\n') ++e;
    std::string l(text + i, e - i);
    while (!l.empty() && (l.back() == '\r' || l.back() == ' ' || l.back() == '\t')) l.pop_back();
    if (!l.empty() && l.find_first_not_of(" \t") != std::string::npos) lines.push_back(l);
    i = e + 1;
  }
  if (lines.empty()) { err = "PHYLIP: empty input"; return false; }
  long n_taxa = 0, n_sites = 0;
  if (std::sscanf(lines[0].c_str(), "%ld %ld", &n_taxa, &n_sites) != 2 || n_taxa < 1 || n_sites < 1) {
    err = "PHYLIP: the first line must hold the number of taxa and of sites";
    return false;
  }
  if ((long)lines.size() < 1 + n_taxa) { err = "PHYLIP: fewer lines than taxa"; return false; }
  out.names.assign(n_taxa, "");
  out.seqs.assign(n_taxa, "");
  auto append = [](std::string& dst, const std::string& src, size_t from) {
    for (size_t j = from; j < src.size(); ++j)
      if (!std::isspace((unsigned char)src[j])) dst.push_back(src[j]);
  };
  auto name_and_data = [&](long t, const std::string& l) {
    size_t a = l.find_first_not_of(" \t"), b = a;
    while (b < l.size() && !std::isspace((unsigned char)l[b])) ++b;
    out.names[t] = l.substr(a, b - a);
    append(out.seqs[t], l, b);
  };
  const long body = (long)lines.size() - 1;
  if (body % n_taxa == 0) {
    // blocks of one line per taxon: the first block carries the names, the others data only (interleaved; a
    // sequential file with one line per taxon is the one-block case)
    for (long t = 0; t < n_taxa; ++t) name_and_data(t, lines[1 + t]);
    for (long li = 1 + n_taxa; li < (long)lines.size(); ++li) append(out.seqs[(li - 1) % n_taxa], lines[li], 0);
  } else {
    // sequential with wrapped sequences: a taxon continues on the following lines until it has n_sites characters
    size_t li = 1;
    for (long t = 0; t < n_taxa; ++t) {
      if (li >= lines.size()) { err = "PHYLIP: file ends before taxon " + std::to_string(t + 1); return false; }
      name_and_data(t, lines[li++]);
      while ((long)out.seqs[t].size() < n_sites && li < lines.size()) append(out.seqs[t], lines[li++], 0);
    }
  }
  for (long t = 0; t < n_taxa; ++t)
    if ((long)out.seqs[t].size() != n_sites) {
      err = "PHYLIP: taxon '" + out.names[t] + "' has " + std::to_string(out.seqs[t].size()) + " sites, header says " +
            std::to_string(n_sites);
      return false;
    }
  return true;
}

int emit_alignment(const Alignment& al, const char* alphabet, int* n_taxa, int64_t* n_sites, uint8_t* states,
                   int64_t site_stride, char* names, size_t names_capacity) {
  const size_t L = al.seqs[0].size();
  for (size_t t = 0; t < al.seqs.size(); ++t)
    TREX_CHECK_ARG(al.seqs[t].size() == L, "alignment: taxon '%s' has %zu sites, the first one %zu (sequences must be aligned)",
                   al.names[t].c_str(), al.seqs[t].size(), L);
  TREX_CHECK_ARG(L >= 1, "alignment: empty sequences");
  if (n_taxa) *n_taxa = (int)al.seqs.size();
  if (n_sites) *n_sites = (int64_t)L;
  if (states) {
    TREX_CHECK_ARG(site_stride >= (int64_t)L, "alignment: site_stride %lld < %zu sites", (long long)site_stride, L);
    const Alphabet ab(alphabet);
    for (size_t t = 0; t < al.seqs.size(); ++t) {
      uint8_t* row = states + t * (size_t)site_stride;
      for (size_t j = 0; j < L; ++j) row[j] = ab.code[(unsigned char)al.seqs[t][j]];
      for (size_t j = L; j < (size_t)site_stride; ++j) row[j] = 255;
    }
  }
  if (names) {  // '\n'-separated, NUL-terminated
    size_t need = 1;
    for (const auto& s : al.names) need += s.size() + 1;
    TREX_CHECK_ARG(names_capacity >= need, "alignment: names need %zu bytes, %zu given", need, names_capacity);
    char* p = names;
    for (const auto& s : al.names) {
      std::memcpy(p, s.data(), s.size());
      p += s.size();
      *p++ = '\n';
    }
    *p = 0;
  }
  return TREX_OK;
}

// ---- Newick -------------------------------------------------------------------------------------------------------------
struct NwkNode {
  std::string name;
  std::vector<int> kids;
};

struct NewickParser {
  const char* s;
  size_t n, i = 0;
  std::vector<NwkNode> nodes;
  std::string err;
  void skip() {
    for (;;) {
      while (i < n && std::isspace((unsigned char)s[i])) ++i;
      if (i < n && s[i] == '[') {  // comment
        while (i < n && s[i] != ']') ++i;
        if (i < n) ++i;
      } else break;
    }
  }
  std::string label() {
    skip();
    std::string out;
    if (i < n && s[i] == '\'') {  // quoted label, '' = a quote
      ++i;
      while (i < n) {
        if (s[i] == '\'') {
          if (i + 1 < n && s[i + 1] == '\'') { out.push_back('\''); i += 2; continue; }
          ++i;
          break;
        }
        out.push_back(s[i++]);
      }
    } else {
      while (i < n && !std::strchr("(),:;[", s[i]) && !std::isspace((unsigned char)s[i])) out.push_back(s[i++]);
    }
    return out;
  }
  void branch_length() {
    skip();
    if (i < n && s[i] == ':') {
      ++i;
      skip();
      while (i < n && (std::isdigit((unsigned char)s[i]) || std::strchr("+-.eE", s[i]))) ++i;
    }
  }
  int subtree(int depth) {
    if (depth > 10000) { err = "Newick: nesting deeper than 10000 levels"; return -1; }  // (recursive descent: bounded stack)
    skip();
    const int id = (int)nodes.size();
    nodes.emplace_back();
    if (i < n && s[i] == '(') {
      ++i;
      for (;;) {
        const int c = subtree(depth + 1);
        if (c < 0) return -1;
        nodes[id].kids.push_back(c);
        skip();
        if (i < n && s[i] == ',') { ++i; continue; }
        if (i < n && s[i] == ')') { ++i; break; }
        err = "Newick: expected ',' or ')' at offset " + std::to_string(i);
        return -1;
      }
    }
    nodes[id].name = label();
    branch_length();
    return id;
  }
};

}  // namespace
}  // namespace trex

extern "C" {

int trex_parse_alignment_host(const char* text, size_t text_len, int format, const char* alphabet, int* n_taxa,
                              int64_t* n_sites, uint8_t* states_host, int64_t site_stride, char* names_host,
                              size_t names_capacity) {
  using namespace trex;
  TREX_CHECK_ARG(text && alphabet && (format == TREX_FORMAT_FASTA || format == TREX_FORMAT_PHYLIP),
                 "trex_parse_alignment_host: bad arguments");
  Alignment al;
  std::string err;
  const bool ok = format == TREX_FORMAT_FASTA ? parse_fasta(text, text_len, al, err) : parse_phylip(text, text_len, al, err);
  if (!ok) {
    set_error("trex_parse_alignment_host: %s", err.c_str());
    return TREX_ERR_INVALID_ARGUMENT;
  }
  return emit_alignment(al, alphabet, n_taxa, n_sites, states_host, site_stride, names_host, names_capacity);
}

int trex_parse_newick_host(const char* text, size_t text_len, const char* taxon_names, int* n_leaves,
                           int32_t* parent_host, float* adjacency_host, char* names_host, size_t names_capacity) {
  using namespace trex;
  TREX_CHECK_ARG(text, "trex_parse_newick_host: null text");
  NewickParser ps{text, text_len};
  const int root = ps.subtree(0);
  if (root < 0) {
    set_error("trex_parse_newick_host: %s", ps.err.c_str());
    return TREX_ERR_INVALID_ARGUMENT;
  }
  // unary nodes (a pair of parentheses around one subtree) are transparent; anything but 2 children is refused
  std::vector<NwkNode>& nd = ps.nodes;
  std::vector<std::string> leaf_names;
  std::vector<int> leaf_of(nd.size(), -1);
  for (size_t v = 0; v < nd.size(); ++v) {
    if (nd[v].kids.empty()) {
      leaf_of[v] = (int)leaf_names.size();
      leaf_names.push_back(nd[v].name);
    } else {
      TREX_CHECK_ARG(nd[v].kids.size() <= 2,
                     "trex_parse_newick_host: node with %zu children — the Sankoff DP of the reference reads exactly two "
                     "(src/trex/sankoff.py:60): resolve multifurcations first", nd[v].kids.size());
    }
  }
  const int n = (int)leaf_names.size();
  TREX_CHECK_ARG(n >= 2, "trex_parse_newick_host: a tree needs at least two leaves");
  // leaf index = position in the caller's taxon list (alignment row order) when given, else order of appearance
  std::vector<int> leaf_index(n);
  if (taxon_names) {
    std::vector<std::string> want;
    const char* p = taxon_names;
    while (*p) {
      const char* e = std::strchr(p, '\n');
      if (!e) e = p + std::strlen(p);
      if (e > p) want.emplace_back(p, e - p);
      p = *e ? e + 1 : e;
    }
    TREX_CHECK_ARG((int)want.size() == n, "trex_parse_newick_host: %zu taxon names given, the tree has %d leaves", want.size(), n);
    std::vector<char> used(n, 0);
    for (int l = 0; l < n; ++l) {
      int hit = -1;
      for (int w = 0; w < n; ++w)
        if (!used[w] && want[w] == leaf_names[l]) { hit = w; break; }
      TREX_CHECK_ARG(hit >= 0, "trex_parse_newick_host: leaf '%s' of the tree is not among the taxon names (or appears twice)",
                     leaf_names[l].c_str());
      used[hit] = 1;
      leaf_index[l] = hit;
    }
  } else {
    for (int l = 0; l < n; ++l) leaf_index[l] = l;
  }
  // post-order numbering of the binary nodes: children before parents, root last (iterative: deep caterpillars)
  std::vector<int> id(nd.size(), -1);
  int next = n;
  std::vector<std::pair<int, size_t>> stack{{root, 0}};
  std::vector<int> parent(2 * (size_t)n - 1, -1);
  auto resolve = [&](int v) {  // skip unary chains
    while (nd[v].kids.size() == 1) v = nd[v].kids[0];
    return v;
  };
  stack[0].first = resolve(root);
  while (!stack.empty()) {
    auto& top = stack.back();
    const int v = top.first;
    if (nd[v].kids.empty()) {
      id[v] = leaf_index[leaf_of[v]];
      stack.pop_back();
    } else if (top.second < 2) {
      const int c = resolve(nd[v].kids[top.second++]);
      stack.push_back({c, 0});
    } else {
      TREX_CHECK_ARG(next < 2 * n - 1 + 1, "trex_parse_newick_host: internal error (node count)");
      id[v] = next++;
      for (int k = 0; k < 2; ++k) parent[id[resolve(nd[v].kids[k])]] = id[v];
      stack.pop_back();
    }
  }
  TREX_CHECK_ARG(next == 2 * n - 1, "trex_parse_newick_host: %d internal nodes for %d leaves — not a binary tree", next - n, n);
  if (n_leaves) *n_leaves = n;
  const int n_all = 2 * n - 1;
  if (parent_host) std::memcpy(parent_host, parent.data(), sizeof(int32_t) * (size_t)n_all);
  if (adjacency_host) {
    std::memset(adjacency_host, 0, sizeof(float) * (size_t)n_all * n_all);
    for (int v = 0; v < n_all - 1; ++v) adjacency_host[(size_t)v * n_all + parent[v]] = 1.f;
  }
  if (names_host) {  // leaf names in LEAF-INDEX order, '\n'-separated
    std::vector<std::string> by_index(n);
    for (int l = 0; l < n; ++l) by_index[leaf_index[l]] = leaf_names[l];
    size_t need = 1;
    for (const auto& s : by_index) need += s.size() + 1;
    TREX_CHECK_ARG(names_capacity >= need, "trex_parse_newick_host: names need %zu bytes, %zu given", need, names_capacity);
    char* p = names_host;
    for (const auto& s : by_index) {
      std::memcpy(p, s.data(), s.size());
      p += s.size();
      *p++ = '\n';
    }
    *p = 0;
  }
  return TREX_OK;
}

}  // extern "C"
