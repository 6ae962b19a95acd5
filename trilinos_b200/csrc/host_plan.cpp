// Host-side planning for the B200 Block GMRES path: linear row-block partition,
// column-map construction (ghost ordering), and Galeri-style stencil generators.
// Pure host C++ (no CUDA) so the multi-rank logic is testable on CPU.
//
// Reference behaviour restated (not copied):
//   partition : Epetra_Map(n,0,comm)            epetra/src/Epetra_BlockMap.cpp:83-90
//   colmap    : Epetra_CrsGraph::MakeColMap     epetra/src/Epetra_CrsGraph.cpp:1348-1524
//   stencils  : galeri/src-epetra/CrsMatrices/Galeri_Cross2D.h:57-117, Galeri_Cross3D.h:53-131,
//               Galeri_Recirc2D.h:55-124, Galeri_Utils.cpp:330-398,
//               galeri/src-xpetra/Galeri_XpetraMatrixTypes.hpp:531-633,736-761 (Brick3D)
#include <algorithm>
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "b200_belos.h"
#include "internal.h"

extern "C" int b200_linear_partition(int64_t n_global, int nranks, int rank, int64_t* row_begin,
                                     int64_t* row_end) {
  if (n_global < 0 || nranks <= 0 || rank < 0 || rank >= nranks || !row_begin || !row_end)
    return b200::fail(B200_ERR_ARG, "b200_linear_partition: bad arguments");
  // first (n mod P) ranks own ceil(n/P) rows, the rest floor(n/P)
  const int64_t q = n_global / nranks, r = n_global % nranks;
  const int64_t lo = rank * q + std::min<int64_t>(rank, r);
  *row_begin = lo;
  *row_end = lo + q + (rank < r ? 1 : 0);
  return B200_OK;
}

namespace b200 {
int owner_of(int64_t gid, int64_t n_global, int nranks) {
  const int64_t q = n_global / nranks, r = n_global % nranks;
  const int64_t split = r * (q + 1);
  if (gid < split) return (int)(gid / (q + 1));
  if (q == 0) return nranks - 1;
  return (int)(r + (gid - split) / q);
}
}  // namespace b200

extern "C" int64_t b200_build_colmap(int64_t n_global, int nranks, int rank, int64_t nnz,
                                     const int64_t* colind_global, int32_t* colind_local,
                                     int64_t* ghost_gids, int32_t* ghost_owner) {
  int64_t rb, re;
  if (b200_linear_partition(n_global, nranks, rank, &rb, &re) != 0) return -1;
  const int64_t nloc = re - rb;
  // Ghost columns: distinct remote GIDs, grouped by owning rank (ascending rank) and,
  // inside one owner, ascending GID.  With a linear map "sort by (owner, gid)" equals
  // "sort by gid"; MakeColMap sorts remote GIDs by PID and then by GID inside each PID.
  std::vector<int64_t> remote;
  remote.reserve(1024);
  for (int64_t k = 0; k < nnz; ++k) {
    const int64_t g = colind_global[k];
    if (g < 0 || g >= n_global) {
      b200::fail(B200_ERR_ARG, "b200_build_colmap: column id out of range");
      return -1;
    }
    if (g < rb || g >= re) remote.push_back(g);
  }
  std::sort(remote.begin(), remote.end());
  remote.erase(std::unique(remote.begin(), remote.end()), remote.end());
  const int64_t nghost = (int64_t)remote.size();
  if (nloc + nghost > 0x7fffffffLL) {
    b200::fail(B200_ERR_ARG, "b200_build_colmap: local column count exceeds int32");
    return -1;
  }
  for (int64_t i = 0; i < nghost; ++i) {
    if (ghost_gids) ghost_gids[i] = remote[i];
    if (ghost_owner) ghost_owner[i] = b200::owner_of(remote[i], n_global, nranks);
  }
  if (colind_local) {
    for (int64_t k = 0; k < nnz; ++k) {
      const int64_t g = colind_global[k];
      if (g >= rb && g < re) {
        colind_local[k] = (int32_t)(g - rb);
      } else {
        const int64_t pos = std::lower_bound(remote.begin(), remote.end(), g) - remote.begin();
        colind_local[k] = (int32_t)(nloc + pos);
      }
    }
  }
  return nghost;
}

// ------------------------------------------------------------------ halo plan (host part)
// Receive side of the Epetra_Import plan (epetra/src/Epetra_Import.cpp:319-460): ghosts are in
// column-map order, i.e. grouped by owner, so rank r's block is [recv_off[r], recv_off[r]+need[r]).
extern "C" int b200_halo_recv_plan(int nranks, int64_t n_ghost, const int32_t* ghost_owner,
                                   int64_t* need_cnt, int64_t* recv_off) {
  if (nranks <= 0 || n_ghost < 0 || (n_ghost > 0 && !ghost_owner) || !need_cnt || !recv_off)
    return b200::fail(B200_ERR_ARG, "b200_halo_recv_plan: bad arguments");
  for (int r = 0; r < nranks; ++r) need_cnt[r] = 0;
  int prev = -1;
  for (int64_t g = 0; g < n_ghost; ++g) {
    const int o = ghost_owner[g];
    if (o < 0 || o >= nranks || o < prev)
      return b200::fail(B200_ERR_ARG, "b200_halo_recv_plan: ghosts are not grouped by ascending owner");
    prev = o;
    need_cnt[o]++;
  }
  recv_off[0] = 0;
  for (int r = 1; r < nranks; ++r) recv_off[r] = recv_off[r - 1] + need_cnt[r - 1];
  return B200_OK;
}

// Send side: the global ids my peers asked for (concatenated by peer rank) become local row
// ids to pack (Epetra_Import's ExportLIDs).  Every id must be owned by this rank.
extern "C" int b200_halo_send_index(int64_t n_global, int nranks, int rank, int64_t n_asked,
                                    const int64_t* asked_gids, int32_t* send_idx) {
  int64_t rb, re;
  if (b200_linear_partition(n_global, nranks, rank, &rb, &re) != 0) return B200_ERR_ARG;
  if (n_asked < 0 || (n_asked > 0 && (!asked_gids || !send_idx)))
    return b200::fail(B200_ERR_ARG, "b200_halo_send_index: bad arguments");
  for (int64_t i = 0; i < n_asked; ++i) {
    const int64_t g = asked_gids[i];
    if (g < rb || g >= re)
      return b200::fail(B200_ERR_ARG, "halo setup: a peer asked for a row this rank does not own");
    send_idx[i] = (int32_t)(g - rb);
  }
  return B200_OK;
}

// ------------------------------------------------------------------ Galeri
namespace {

struct Row {
  int n = 0;
  int64_t ind[27];
  double val[27];
  void push(int64_t i, double v) {
    ind[n] = i;
    val[n] = v;
    ++n;
  }
  // Epetra/Tpetra fillComplete sorts each row by column id and sums duplicates
  // (stencils here never produce duplicates).
  void sort() {
    for (int a = 1; a < n; ++a) {
      int64_t i = ind[a];
      double v = val[a];
      int b = a - 1;
      while (b >= 0 && ind[b] > i) {
        ind[b + 1] = ind[b];
        val[b + 1] = val[b];
        --b;
      }
      ind[b + 1] = i;
      val[b + 1] = v;
    }
  }
};

inline void nbr2d(int64_t i, int64_t nx, int64_t ny, int64_t& left, int64_t& right,
                  int64_t& lower, int64_t& upper) {
  const int64_t ix = i % nx, iy = (i - ix) / nx;
  left = (ix == 0) ? -1 : i - 1;
  right = (ix == nx - 1) ? -1 : i + 1;
  lower = (iy == 0) ? -1 : i - nx;
  upper = (iy == ny - 1) ? -1 : i + nx;
}

inline void nbr3d(int64_t i, int64_t nx, int64_t ny, int64_t nz, int64_t& left, int64_t& right,
                  int64_t& lower, int64_t& upper, int64_t& below, int64_t& above) {
  const int64_t nxy = nx * ny;
  const int64_t ixy = i % nxy, iz = (i - ixy) / nxy;
  below = (iz == 0) ? -1 : i - nxy;
  above = (iz == nz - 1) ? -1 : i + nxy;
  nbr2d(ixy, nx, ny, left, right, lower, upper);
  if (left != -1) left += iz * nxy;
  if (right != -1) right += iz * nxy;
  if (lower != -1) lower += iz * nxy;
  if (upper != -1) upper += iz * nxy;
}

inline void cross2d_row(Row& r, int64_t g, int64_t nx, int64_t ny, double a, double b, double c,
                        double d, double e) {
  int64_t left, right, lower, upper;
  nbr2d(g, nx, ny, left, right, lower, upper);
  if (left != -1) r.push(left, b);
  if (right != -1) r.push(right, c);
  if (lower != -1) r.push(lower, d);
  if (upper != -1) r.push(upper, e);
  r.push(g, a);
}

inline void nine_point(Row& r, int64_t centre, int64_t nx, int64_t ny, int64_t nz, double b,
                       double c, double d, double e, double z1, double z2, double z3, double z4) {
  int64_t left, right, lower, upper, below, above;
  nbr3d(centre, nx, ny, nz, left, right, lower, upper, below, above);
  if (left != -1) r.push(left, b);
  if (right != -1) r.push(right, c);
  if (lower != -1) r.push(lower, d);
  if (upper != -1) r.push(upper, e);
  if (left != -1 && lower != -1) r.push(lower - 1, z1);
  if (right != -1 && lower != -1) r.push(lower + 1, z2);
  if (left != -1 && upper != -1) r.push(upper - 1, z3);
  if (right != -1 && upper != -1) r.push(upper + 1, z4);
}

void make_row(const std::string& name, int64_t g, int64_t nx, int64_t ny, int64_t nz,
              double conv, double diff, Row& r) {
  r.n = 0;
  if (name == "Laplace2D") {
    cross2d_row(r, g, nx, ny, 4.0, -1.0, -1.0, -1.0, -1.0);
  } else if (name == "Laplace3D") {
    int64_t left, right, lower, upper, below, above;
    nbr3d(g, nx, ny, nz, left, right, lower, upper, below, above);
    if (left != -1) r.push(left, -1.0);
    if (right != -1) r.push(right, -1.0);
    if (lower != -1) r.push(lower, -1.0);
    if (upper != -1) r.push(upper, -1.0);
    if (below != -1) r.push(below, -1.0);
    if (above != -1) r.push(above, -1.0);
    r.push(g, 6.0);
  } else if (name == "Recirc2D") {
    // upwinded convection (recirculating field) + diffusion on the unit square
    const double lx = 1.0, ly = 1.0;
    const double hx = lx / (nx + 1), hy = ly / (ny + 1);
    const int64_t ix = g % nx, iy = (g - ix) / nx;
    const double x = hx * (ix + 1), y = hy * (iy + 1);
    const double cx = conv * 4 * x * (x - 1.) * (1. - 2 * y) / hx;
    const double cy = -conv * 4 * y * (y - 1.) * (1. - 2 * x) / hy;
    double A = 0, B = 0, C = 0, D = 0, E = 0;
    if (cx < 0) { C += cx; A -= cx; } else { B -= cx; A += cx; }
    if (cy < 0) { E += cy; A -= cy; } else { D -= cy; A += cy; }
    A += diff * 2. / (hx * hx) + diff * 2. / (hy * hy);
    B -= diff / (hx * hx);
    C -= diff / (hx * hx);
    D -= diff / (hy * hy);
    E -= diff / (hy * hy);
    cross2d_row(r, g, nx, ny, A, B, C, D, E);
  } else if (name == "Brick3D") {
    // 27-point stencil, centre 26, every neighbour -1; default boundary flags are all
    // "Dirichlet" but only the x/y faces are tested, so nodes lying only on a z face get the
    // "Neumann" diagonal (minus the sum of the off-diagonals).
    const double a = 26.0, b = -1.0, c = -1.0, d = -1.0, e = -1.0;
    int64_t left, right, front, back, below, above;
    nbr3d(g, nx, ny, nz, left, right, front, back, below, above);
    const bool is_dirichlet = (left == -1) || (right == -1) || (front == -1) || (back == -1);
    nine_point(r, g, nx, ny, nz, b, b, b, b, c, c, c, c);
    if (below != -1) {
      r.push(below, b);
      nine_point(r, below, nx, ny, nz, d, d, d, d, e, e, e, e);
    }
    if (above != -1) {
      r.push(above, b);
      nine_point(r, above, nx, ny, nz, d, d, d, d, e, e, e, e);
    }
    const int64_t nxy = nx * ny;
    const int64_t ix = g % nx, ixy = g % nxy, iy = (ixy - ix) / nx, iz = (g - ixy) / nxy;
    const bool boundary =
        ix == 0 || ix == nx - 1 || iy == 0 || iy == ny - 1 || iz == 0 || iz == nz - 1;
    double z = a;
    if (boundary && !is_dirichlet) {
      z = 0.0;
      for (int j = 0; j < r.n; ++j) z -= r.val[j];
    }
    r.push(g, z);
  } else if (name == "Tridiag") {
    // Laplace1D: tridiag(-1, 2, -1) on n = nx unknowns
    if (g > 0) r.push(g - 1, -1.0);
    if (g < nx - 1) r.push(g + 1, -1.0);
    r.push(g, 2.0);
  } else if (name == "Diag") {
    // diagonal operator with entries (g+1) * conv  (known-answer tests, cf.
    // belos/epetra/test/BlockGmres/test_bl_gmres_diag.cpp)
    r.push(g, (double)(g + 1) * (conv == 0.0 ? 1.0 : conv));
  }
  r.sort();
}

}  // namespace

extern "C" int b200_galeri_generate(const char* name_c, int nx, int ny, int nz, double conv,
                                    double diff, int64_t row_begin, int64_t row_end,
                                    int64_t* nnz_out, int32_t* rowptr, int64_t* colind,
                                    double* values) {
  if (!name_c || !nnz_out) return b200::fail(B200_ERR_ARG, "b200_galeri_generate: null argument");
  const std::string name(name_c);
  int64_t n_global;
  if (name == "Laplace2D" || name == "Recirc2D") {
    if (nx <= 0 || ny <= 0) return b200::fail(B200_ERR_ARG, "galeri: nx, ny must be positive");
    n_global = (int64_t)nx * ny;
  } else if (name == "Laplace3D" || name == "Brick3D") {
    if (nx <= 0 || ny <= 0 || nz <= 0)
      return b200::fail(B200_ERR_ARG, "galeri: nx, ny, nz must be positive");
    n_global = (int64_t)nx * ny * nz;
  } else if (name == "Tridiag" || name == "Diag") {
    if (nx <= 0) return b200::fail(B200_ERR_ARG, "galeri: nx must be positive");
    n_global = nx;
  } else {
    return b200::fail(B200_ERR_ARG, "galeri: unknown matrix type");
  }
  if (row_begin < 0 || row_end < row_begin || row_end > n_global)
    return b200::fail(B200_ERR_ARG, "galeri: row range out of bounds");
  Row r;
  int64_t nnz = 0;
  if (!rowptr) {
    for (int64_t g = row_begin; g < row_end; ++g) {
      make_row(name, g, nx, ny, nz, conv, diff, r);
      nnz += r.n;
    }
    *nnz_out = nnz;
    return B200_OK;
  }
  if (!colind || !values) return b200::fail(B200_ERR_ARG, "galeri: null output arrays");
  rowptr[0] = 0;
  for (int64_t g = row_begin; g < row_end; ++g) {
    make_row(name, g, nx, ny, nz, conv, diff, r);
    for (int j = 0; j < r.n; ++j) {
      colind[nnz + j] = r.ind[j];
      values[nnz + j] = r.val[j];
    }
    nnz += r.n;
    if (nnz > 0x7fffffffLL) return b200::fail(B200_ERR_ARG, "galeri: local nnz exceeds int32");
    rowptr[g - row_begin + 1] = (int32_t)nnz;
  }
  *nnz_out = nnz;
  return B200_OK;
}
