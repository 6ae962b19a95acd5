// Matrix ingest: Harwell-Boeing and MatrixMarket files -> row-sorted CSR with global column ids,
// ready for b200_op_create_csr.  Host code, no CUDA.
//
// Reference path replaced: Belos::Util::createEpetraProblem (belos/epetra/src/BelosEpetraUtils.cpp:
// 81-200) = Trilinos_Util_read_hb (triutils/src/Trilinos_Util_read_hb.cpp:49-230: header, CSC
// arrays, optional right-hand side / exact solution, symmetric -> full expansion, CSC -> CSR ->
// MSR) + one InsertGlobalValues per row for the off-diagonals and one for the diagonal + FillComplete.
// The matrix Epetra ends up with -- and that this reader reproduces bit for bit -- therefore has
//   * every stored entry of the file (explicit zeros included), symmetric files mirrored,
//   * an explicit diagonal entry in EVERY row (0.0 when the file has none: MSR always stores it),
//   * column indices sorted within each row (FillComplete).
// MatrixMarket: Trilinos_Util_ReadMatrixMarket2Epetra (triutils/src/
// Trilinos_Util_ReadMatrixMarket2Epetra.cpp): "coordinate real general|symmetric", 1-based
// triples, repeated entries summed by Epetra's InsertGlobalValues/FillComplete.
#include <algorithm>
#include <cctype>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "b200_belos.h"
#include "internal.h"

namespace {

using b200::fail;
using b200::failf;

struct Triple {
  int64_t r, c;
  double v;
};

// Fortran edit descriptor of an HB header, e.g. "(10I8)", "(1P5E16.8)", "(1P,4D20.12)", "(3E25.16)":
// values per line and field width
bool parse_fmt(const std::string& f, int* per_line, int* width) {
  std::string s;
  for (char ch : f)
    if (!isspace((unsigned char)ch)) s.push_back((char)toupper((unsigned char)ch));
  size_t a = s.find('('), b = s.rfind(')');
  if (a == std::string::npos || b == std::string::npos || b <= a) return false;
  s = s.substr(a + 1, b - a - 1);
  // drop a scale factor "nP" and an optional comma
  size_t p = s.find('P');
  if (p != std::string::npos) {
    s = s.substr(p + 1);
    if (!s.empty() && s[0] == ',') s = s.substr(1);
  }
  size_t t = s.find_first_of("IEDFG");
  if (t == std::string::npos) return false;
  *per_line = t == 0 ? 1 : atoi(s.substr(0, t).c_str());
  *width = atoi(s.c_str() + t + 1);
  return *per_line > 0 && *width > 0;
}

bool read_line(FILE* f, std::string* out) {
  out->clear();
  int ch;
  bool any = false;
  while ((ch = fgetc(f)) != EOF) {
    any = true;
    if (ch == '\n') break;
    if (ch != '\r') out->push_back((char)ch);
  }
  return any;
}

// `count` fixed-width fields laid out `per_line` to a line (fields may touch each other)
template <class T, class Conv>
bool read_fields(FILE* f, int64_t count, int per_line, int width, std::vector<T>* out, Conv conv) {
  out->resize((size_t)count);
  std::string line;
  int64_t i = 0;
  while (i < count) {
    if (!read_line(f, &line)) return false;
    for (int k = 0; k < per_line && i < count; ++k) {
      const size_t off = (size_t)k * width;
      if (off >= line.size()) break;
      std::string fld = line.substr(off, (size_t)width);
      for (char& ch : fld)
        if (ch == 'D' || ch == 'd') ch = 'E';
      (*out)[(size_t)i++] = conv(fld.c_str());
    }
  }
  return true;
}

// rows sorted by column, repeated (row, col) summed, explicit diagonal in every row if asked
void triples_to_csr(int64_t n, std::vector<Triple>& t, bool force_diag, std::vector<int32_t>* rp,
                    std::vector<int64_t>* ci, std::vector<double>* va) {
  if (force_diag) {
    std::vector<char> has((size_t)n, 0);
    for (const Triple& e : t)
      if (e.r == e.c) has[(size_t)e.r] = 1;
    for (int64_t i = 0; i < n; ++i)
      if (!has[(size_t)i]) t.push_back(Triple{i, i, 0.0});
  }
  std::stable_sort(t.begin(), t.end(), [](const Triple& a, const Triple& b) {
    return a.r != b.r ? a.r < b.r : a.c < b.c;
  });
  rp->assign((size_t)n + 1, 0);
  ci->clear();
  va->clear();
  for (size_t i = 0; i < t.size(); ++i) {
    if (i > 0 && t[i].r == t[i - 1].r && t[i].c == t[i - 1].c) {
      va->back() += t[i].v;
      continue;
    }
    ci->push_back(t[i].c);
    va->push_back(t[i].v);
    (*rp)[(size_t)t[i].r + 1] += 1;
  }
  for (int64_t i = 0; i < n; ++i) (*rp)[(size_t)i + 1] += (*rp)[(size_t)i];
}

struct Parsed {
  std::string path;
  int64_t n = 0;
  std::vector<int32_t> rp;
  std::vector<int64_t> ci;
  std::vector<double> va, rhs, xexact;
  bool has_rhs = false;
};
thread_local Parsed g_last;  // the two-call protocol parses once

int parse_hb(const char* path, Parsed* out) {
  FILE* f = fopen(path, "r");
  if (!f) return failf(B200_ERR_ARG, "read_hb: cannot open file: %s", path);
  std::string l1, l2, l3, l4, l5;
  if (!read_line(f, &l1) || !read_line(f, &l2) || !read_line(f, &l3) || !read_line(f, &l4)) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_hb: truncated header");
  }
  long totcrd = 0, ptrcrd = 0, indcrd = 0, valcrd = 0, rhscrd = 0;
  sscanf(l2.c_str(), "%ld %ld %ld %ld %ld", &totcrd, &ptrcrd, &indcrd, &valcrd, &rhscrd);
  char type[8] = "XXX";
  long nrow = 0, ncol = 0, nnzero = 0, neltvl = 0;
  if (sscanf(l3.c_str(), "%3s %ld %ld %ld %ld", type, &nrow, &ncol, &nnzero, &neltvl) < 4) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_hb: bad header line 3");
  }
  for (char& ch : type) ch = (char)toupper((unsigned char)ch);
  // the reference: "Can only handle real valued / assembled / square matrices"
  if (type[0] != 'R' || type[2] != 'A' || nrow != ncol) {
    fclose(f);
    return failf(B200_ERR_ARG, "read_hb: matrix type %s (%ld x %ld) is not a real assembled square matrix",
                 type, nrow, ncol);
  }
  l4.resize(std::max<size_t>(l4.size(), 72), ' ');
  int pp, pw, ip, iw, vp, vw, rpl = 0, rw = 0;
  if (!parse_fmt(l4.substr(0, 16), &pp, &pw) || !parse_fmt(l4.substr(16, 16), &ip, &iw) ||
      !parse_fmt(l4.substr(32, 20), &vp, &vw)) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_hb: cannot parse the Fortran formats of header line 4");
  }
  char rhstype[4] = "   ";
  long nrhs = 0;
  if (rhscrd > 0) {
    if (!read_line(f, &l5)) {
      fclose(f);
      return fail(B200_ERR_ARG, "read_hb: missing right-hand-side header line");
    }
    l5.resize(std::max<size_t>(l5.size(), 3), ' ');
    memcpy(rhstype, l5.c_str(), 3);
    for (char& ch : rhstype) ch = (char)toupper((unsigned char)ch);
    sscanf(l5.c_str() + 3, "%ld", &nrhs);
    parse_fmt(l4.substr(52, 20), &rpl, &rw);
  }
  std::vector<int64_t> ptr, ind;
  std::vector<double> val;
  auto toi = [](const char* s) { return (int64_t)atol(s); };
  auto tod = [](const char* s) { return strtod(s, nullptr); };
  bool ok = read_fields(f, ncol + 1, pp, pw, &ptr, toi) && read_fields(f, nnzero, ip, iw, &ind, toi) &&
            read_fields(f, nnzero, vp, vw, &val, tod);
  if (!ok) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_hb: truncated matrix data");
  }
  // right-hand side 'F' (full vectors) followed by guess (if 'G') and exact solution (if 'X'):
  // the reference reads them only when an exact solution is there (read_hb.cpp:134-143)
  out->has_rhs = false;
  out->rhs.clear();
  out->xexact.clear();
  if (nrhs > 0 && rhstype[0] == 'F' && rhstype[2] == 'X' && rpl > 0) {
    std::vector<double> all;
    const int64_t nvec = nrhs * (1 + (rhstype[1] == 'G' ? 1 : 0) + 1);
    if (read_fields(f, nvec * nrow, rpl, rw, &all, tod)) {
      out->rhs.assign(all.begin(), all.begin() + nrow);
      const int64_t xoff = nrhs * (1 + (rhstype[1] == 'G' ? 1 : 0)) * nrow;
      out->xexact.assign(all.begin() + xoff, all.begin() + xoff + nrow);
      out->has_rhs = true;
    }
  }
  fclose(f);
  std::vector<Triple> t;
  t.reserve((size_t)nnzero * (type[1] == 'S' ? 2 : 1));
  for (int64_t j = 0; j < ncol; ++j)
    for (int64_t e = ptr[(size_t)j] - 1; e < ptr[(size_t)j + 1] - 1; ++e) {
      const int64_t i = ind[(size_t)e] - 1;
      if (i < 0 || i >= nrow) return fail(B200_ERR_ARG, "read_hb: row index out of range");
      t.push_back(Triple{i, j, val[(size_t)e]});
      if (type[1] == 'S' && i != j) t.push_back(Triple{j, i, val[(size_t)e]});
    }
  out->n = nrow;
  triples_to_csr(nrow, t, /*force_diag=*/true, &out->rp, &out->ci, &out->va);
  out->path = path;
  return 0;
}

int parse_mm(const char* path, Parsed* out) {
  FILE* f = fopen(path, "r");
  if (!f) return failf(B200_ERR_ARG, "read_matrix_market: cannot open file: %s", path);
  std::string line;
  if (!read_line(f, &line)) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_matrix_market: empty file");
  }
  std::string banner;
  for (char ch : line) banner.push_back((char)tolower((unsigned char)ch));
  const bool sym = banner.find("symmetric") != std::string::npos;
  // like the reference reader, any '%' header line is accepted and symmetry is read off the word
  // "symmetric"; a proper MatrixMarket banner must describe a real coordinate matrix
  const bool proper = banner.find("%%matrixmarket") == 0;
  if (banner.empty() || banner[0] != '%' ||
      (proper && (banner.find("coordinate") == std::string::npos || banner.find("real") == std::string::npos))) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_matrix_market: only 'matrix coordinate real general|symmetric' is supported");
  }
  do {
    if (!read_line(f, &line)) {
      fclose(f);
      return fail(B200_ERR_ARG, "read_matrix_market: missing size line");
    }
  } while (line.empty() || line[0] == '%');
  long nr = 0, nc = 0, nz = 0;
  if (sscanf(line.c_str(), "%ld %ld %ld", &nr, &nc, &nz) != 3 || nr != nc) {
    fclose(f);
    return fail(B200_ERR_ARG, "read_matrix_market: bad size line (square matrices only)");
  }
  std::vector<Triple> t;
  t.reserve((size_t)nz * (sym ? 2 : 1));
  for (long e = 0; e < nz; ++e) {
    long i, j;
    double v;
    if (!read_line(f, &line) || sscanf(line.c_str(), "%ld %ld %lf", &i, &j, &v) != 3 || i < 1 || j < 1 ||
        i > nr || j > nc) {
      fclose(f);
      return failf(B200_ERR_ARG, "read_matrix_market: bad entry %ld", e + 1);
    }
    t.push_back(Triple{i - 1, j - 1, v});
    if (sym && i != j) t.push_back(Triple{j - 1, i - 1, v});
  }
  fclose(f);
  out->n = nr;
  out->has_rhs = false;
  out->rhs.clear();
  out->xexact.clear();
  triples_to_csr(nr, t, /*force_diag=*/false, &out->rp, &out->ci, &out->va);
  out->path = path;
  return 0;
}

int deliver(const Parsed& p, int64_t row_begin, int64_t row_end, int64_t* n, int64_t* nnz, int32_t* rowptr,
            int64_t* colind, double* values, double* rhs, double* xexact, int* has_rhs) {
  if (row_end < 0) row_end = p.n;
  if (row_begin < 0 || row_begin > row_end || row_end > p.n)
    return fail(B200_ERR_ARG, "matrix reader: bad row range");
  const int32_t b = p.rp[(size_t)row_begin], e = p.rp[(size_t)row_end];
  if (n) *n = p.n;
  if (nnz) *nnz = e - b;
  if (has_rhs) *has_rhs = p.has_rhs ? 1 : 0;
  if (!rowptr) return 0;  // size query: the parse stays cached for the delivering call
  for (int64_t i = row_begin; i <= row_end; ++i) rowptr[i - row_begin] = p.rp[(size_t)i] - b;
  if (colind) std::copy(p.ci.begin() + b, p.ci.begin() + e, colind);
  if (values) std::copy(p.va.begin() + b, p.va.begin() + e, values);
  if (p.has_rhs && rhs) std::copy(p.rhs.begin() + row_begin, p.rhs.begin() + row_end, rhs);
  if (p.has_rhs && xexact) std::copy(p.xexact.begin() + row_begin, p.xexact.begin() + row_end, xexact);
  return 0;
}

}  // namespace

extern "C" int b200_read_hb(const char* path, int64_t row_begin, int64_t row_end, int64_t* n, int64_t* nnz,
                            int32_t* rowptr, int64_t* colind_global, double* values, double* rhs,
                            double* xexact, int* has_rhs) {
  if (!path) return fail(B200_ERR_ARG, "read_hb: null path");
  if (g_last.path != std::string("hb:") + path) {
    Parsed p;
    const int rc = parse_hb(path, &p);
    if (rc) return rc;
    p.path = std::string("hb:") + path;
    g_last = std::move(p);
  }
  const int rc = deliver(g_last, row_begin, row_end, n, nnz, rowptr, colind_global, values, rhs, xexact, has_rhs);
  if (rowptr) g_last = Parsed();  // delivered: a later call re-reads the file
  return rc;
}

extern "C" int b200_read_matrix_market(const char* path, int64_t row_begin, int64_t row_end, int64_t* n,
                                       int64_t* nnz, int32_t* rowptr, int64_t* colind_global,
                                       double* values) {
  if (!path) return fail(B200_ERR_ARG, "read_matrix_market: null path");
  if (g_last.path != std::string("mm:") + path) {
    Parsed p;
    const int rc = parse_mm(path, &p);
    if (rc) return rc;
    p.path = std::string("mm:") + path;
    g_last = std::move(p);
  }
  const int rc = deliver(g_last, row_begin, row_end, n, nnz, rowptr, colind_global, values, nullptr, nullptr, nullptr);
  if (rowptr) g_last = Parsed();
  return rc;
}
