// b200slam — block-sparse Cholesky for the large-graph solve, HOST side: fill-reducing ordering, symbolic
// factorisation and the static schedule ("plan") of the numeric kernel (gn_sparse.cuh).
//
// What it replaces: the analysis half of Eigen's SimplicialLLT (AMD ordering + symbolic factorisation), which the
// reference runs on the host inside SparseBlock::solve on every iteration (gn_kernels.cu:131-152).  Here it runs
// once per graph TOPOLOGY (the plan is cached by a hash of the edge list) and only the numeric factorisation runs
// per iteration, on the device.  Unknowns are the 7x7 blocks of the free poses; the fixed pose (row 0) is dropped
// and duplicate edges are summed, like SparseBlock::update_lhs/update_rhs (:70-112).
//
// Ordering: greedy minimum degree with MULTIPLE elimination — every round eliminates an independent set of
// minimum-degree poses — so that chain-like graphs (odometry + a few loop closures: the graphs on which conjugate
// gradients need ~N iterations) get a bushy elimination tree: few levels, each with many independent columns.
// The plan is left-looking and fully static: per destination column the list of block products to subtract,
// per block the list of edge contributions to gather, columns grouped by tree level.  Everything is summed in a
// fixed order => deterministic and bit-identical across ranks.
#pragma once
#include <stdint.h>

#include <math.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <vector>

namespace b200sparse {

struct Plan {
  int M = 0;                        // free poses (unknown blocks)
  int nL = 0;                       // off-diagonal blocks of L
  long long work = 0;               // block products of the factorisation (left-looking update triples)
  int levels = 0;
  std::vector<int> perm;            // perm[k] = free pose (0-based) eliminated k-th; column k of L belongs to it
  std::vector<int> colptr, rowidx;  // CSC of the strictly lower block structure of L (positions, ascending)
  std::vector<int> level_ptr, level_cols;  // columns by elimination-tree level (leaves first)
  std::vector<int> upd_ptr;         // M+1: update triples of destination column j
  std::vector<int> upd;             // 3 ints per triple: slot a (L_ik), slot b (L_jk), dst (slot in column j, or -1 = diagonal)
  std::vector<int> row_ptr;         // M+1: forward-substitution entries of row j
  std::vector<int> row_ent;         // 2 ints per entry: column k, slot of L_jk
  std::vector<int> asm_ptr;         // M + nL + 1: edge contributions of [diagonal block k | off-diagonal slot s]
  std::vector<int> asm_ent;         // (edge << 1) | 1 if subtracted (off-diagonal), | 0 if added (diagonal)
  std::vector<int> rhs_ptr;         // M+1: edge contributions of the right-hand side of column k
  std::vector<int> rhs_ent;         // (edge << 1) | 1: + g_e (pose is j of the edge), | 0: - g_e (pose is i)
};

// ii_row / jj_row: rows of Twc (0 = the fixed pose) of the E directed edges.  Returns false for an empty problem
// or as soon as the factorisation would need more than max_work block products (the graph fills in: conjugate
// gradients are the better solver there, and the elimination is abandoned before it gets expensive).
inline bool analyze(const int64_t *ii_row, const int64_t *jj_row, int E, int N, Plan &P, long long max_work) {
  const int M = N - 1;
  P = Plan();
  P.M = M;
  if (M < 1) return false;
  const int W = (M + 63) / 64;
  std::vector<uint64_t> adj((size_t)M * W, 0);
  auto set = [&](int a, int b) { adj[(size_t)a * W + (b >> 6)] |= 1ull << (b & 63); };
  auto test = [&](const uint64_t *row, int b) { return (row[b >> 6] >> (b & 63)) & 1ull; };
  for (int e = 0; e < E; e++) {
    const int i = (int)ii_row[e] - 1, j = (int)jj_row[e] - 1;
    if (i >= 0 && j >= 0 && i != j && i < M && j < M) { set(i, j); set(j, i); }
  }
  std::vector<int> deg(M), pos(M, -1);
  std::vector<char> alive(M, 1), blocked(M);
  for (int v = 0; v < M; v++) {
    int d = 0;
    for (int w = 0; w < W; w++) d += __builtin_popcountll(adj[(size_t)v * W + w]);
    deg[v] = d;
  }
  P.perm.reserve(M);
  std::vector<std::vector<int>> colv(M);  // structure of each column in ORIGINAL ids (remaining neighbours at elimination)
  std::vector<uint64_t> rowv(W);
  int remaining = M;
  long long work = 0;
  while (remaining > 0) {
    int dmin = 1 << 30;
    for (int v = 0; v < M; v++)
      if (alive[v] && deg[v] < dmin) dmin = deg[v];
    std::fill(blocked.begin(), blocked.end(), 0);
    // independent set of poses of (near-)minimum degree, in index order
    std::vector<int> round;
    for (int v = 0; v < M; v++) {
      // degree <= 2 always qualifies: on a chain (the fixed pose cuts the odometry ring open) only the two end
      // poses have minimum degree 1, and eliminating just those would serialise the whole factorisation
      if (!alive[v] || blocked[v] || deg[v] > (dmin < 2 ? 2 : dmin)) continue;
      round.push_back(v);
      const uint64_t *row = &adj[(size_t)v * W];
      for (int w = 0; w < W; w++) {
        uint64_t m = row[w];
        while (m) {
          const int b = w * 64 + __builtin_ctzll(m);
          m &= m - 1;
          blocked[b] = 1;
        }
      }
      blocked[v] = 1;
    }
    for (int v : round) {
      // neighbours still alive = structure of this column; connect them pairwise (fill), drop v
      uint64_t *row = &adj[(size_t)v * W];
      std::vector<int> &nb = colv[v];
      for (int w = 0; w < W; w++) {
        uint64_t m = row[w];
        while (m) {
          nb.push_back(w * 64 + __builtin_ctzll(m));
          m &= m - 1;
        }
      }
      work += (long long)nb.size() * ((long long)nb.size() + 1) / 2;
      if (work > max_work) return false;
      for (int w = 0; w < W; w++) rowv[w] = row[w];
      for (int a : nb) {
        uint64_t *ra = &adj[(size_t)a * W];
        int d = 0;
        for (int w = 0; w < W; w++) {
          ra[w] |= rowv[w];
          if (w == (a >> 6)) ra[w] &= ~(1ull << (a & 63));
          if (w == (v >> 6)) ra[w] &= ~(1ull << (v & 63));
          d += __builtin_popcountll(ra[w]);
        }
        deg[a] = d;
      }
      for (int w = 0; w < W; w++) row[w] = 0;
      alive[v] = 0;
      pos[v] = (int)P.perm.size();
      P.perm.push_back(v);
      remaining--;
    }
  }
  (void)test;
  // CSC structure in positions
  P.colptr.assign(M + 1, 0);
  for (int k = 0; k < M; k++) P.colptr[k + 1] = P.colptr[k] + (int)colv[P.perm[k]].size();
  P.nL = P.colptr[M];
  P.rowidx.resize(P.nL);
  for (int k = 0; k < M; k++) {
    int o = P.colptr[k];
    for (int a : colv[P.perm[k]]) P.rowidx[o++] = pos[a];
    std::sort(P.rowidx.begin() + P.colptr[k], P.rowidx.begin() + P.colptr[k + 1]);
  }
  // elimination-tree levels: level(k) = 1 + max level of the columns that update k (its descendants)
  std::vector<int> level(M, 0);
  for (int k = 0; k < M; k++)
    for (int s = P.colptr[k]; s < P.colptr[k + 1]; s++) level[P.rowidx[s]] = std::max(level[P.rowidx[s]], level[k] + 1);
  P.levels = 0;
  for (int k = 0; k < M; k++) P.levels = std::max(P.levels, level[k] + 1);
  P.level_ptr.assign(P.levels + 1, 0);
  for (int k = 0; k < M; k++) P.level_ptr[level[k] + 1]++;
  for (int l = 0; l < P.levels; l++) P.level_ptr[l + 1] += P.level_ptr[l];
  P.level_cols.resize(M);
  {
    std::vector<int> fill(P.level_ptr.begin(), P.level_ptr.end() - 1);
    for (int k = 0; k < M; k++) P.level_cols[fill[level[k]]++] = k;
  }
  // row structure (forward substitution) and left-looking update triples, grouped by destination column
  std::vector<std::vector<int>> rows(M);  // per row j: (k, slot) pairs flattened
  for (int k = 0; k < M; k++)
    for (int s = P.colptr[k]; s < P.colptr[k + 1]; s++) {
      rows[P.rowidx[s]].push_back(k);
      rows[P.rowidx[s]].push_back(s);
    }
  P.row_ptr.assign(M + 1, 0);
  for (int j = 0; j < M; j++) P.row_ptr[j + 1] = P.row_ptr[j] + (int)rows[j].size() / 2;
  P.row_ent.reserve((size_t)2 * P.nL);
  for (int j = 0; j < M; j++) P.row_ent.insert(P.row_ent.end(), rows[j].begin(), rows[j].end());
  P.upd_ptr.assign(M + 1, 0);
  std::vector<int> where(M, -1);  // position -> slot within the destination column
  for (int j = 0; j < M; j++) {
    for (int s = P.colptr[j]; s < P.colptr[j + 1]; s++) where[P.rowidx[s]] = s;
    std::vector<std::array<int, 3>> trip;
    for (size_t t = 0; t < rows[j].size(); t += 2) {
      const int k = rows[j][t], sb = rows[j][t + 1];  // L_jk lives in slot sb of column k
      for (int sa = sb; sa < P.colptr[k + 1]; sa++) {  // rows i >= j of column k (ascending: sb is row j itself)
        const int i = P.rowidx[sa];
        trip.push_back({sa, sb, i == j ? -1 : where[i]});
      }
    }
    // grouped by DESTINATION block (stable: the sources of one block keep their column order), so that the numeric
    // kernel accumulates all updates of a block in registers and touches the block once
    std::stable_sort(trip.begin(), trip.end(), [](const std::array<int, 3> &a, const std::array<int, 3> &b) { return a[2] < b[2]; });
    for (const auto &tr : trip) { P.upd.push_back(tr[0]); P.upd.push_back(tr[1]); P.upd.push_back(tr[2]); }
    P.upd_ptr[j + 1] = (int)P.upd.size() / 3;
    for (int s = P.colptr[j]; s < P.colptr[j + 1]; s++) where[P.rowidx[s]] = -1;
  }
  P.work = P.upd_ptr[M];
  // assembly gather lists, in edge order
  std::vector<std::vector<int>> contrib((size_t)M + P.nL), rhs(M);
  for (int e = 0; e < E; e++) {
    const int i = (int)ii_row[e] - 1, j = (int)jj_row[e] - 1;
    if (i == j) continue;  // a self-edge contributes +B +B -B -B = 0 and g - g = 0
    const int pi = i >= 0 ? pos[i] : -1, pj = j >= 0 ? pos[j] : -1;
    if (pi >= 0) { contrib[pi].push_back(e << 1); rhs[pi].push_back(e << 1); }
    if (pj >= 0) { contrib[pj].push_back(e << 1); rhs[pj].push_back((e << 1) | 1); }
    if (pi >= 0 && pj >= 0) {
      const int c = std::min(pi, pj), r = std::max(pi, pj);
      const int *b = &P.rowidx[P.colptr[c]], *en = &P.rowidx[P.colptr[c + 1]];
      const int *it = std::lower_bound(b, en, r);
      const int slot = (int)(it - &P.rowidx[0]);
      // H_ij = -B and H_ji = -B are the same lower block: the two directed edges of a pair each contribute once,
      // exactly like the four update_lhs blocks of the reference (ii/ij/ji/jj) restricted to the lower triangle
      contrib[(size_t)M + slot].push_back((e << 1) | 1);
    }
  }
  P.asm_ptr.assign((size_t)M + P.nL + 1, 0);
  for (size_t b = 0; b < contrib.size(); b++) P.asm_ptr[b + 1] = P.asm_ptr[b] + (int)contrib[b].size();
  P.asm_ent.reserve(P.asm_ptr.back());
  for (auto &c : contrib) P.asm_ent.insert(P.asm_ent.end(), c.begin(), c.end());
  P.rhs_ptr.assign(M + 1, 0);
  for (int k = 0; k < M; k++) P.rhs_ptr[k + 1] = P.rhs_ptr[k] + (int)rhs[k].size();
  for (auto &c : rhs) P.rhs_ent.insert(P.rhs_ent.end(), c.begin(), c.end());
  return true;
}

}  // namespace b200sparse

// ---- host execution of the plan (reference for the device kernel; used by the CPU tests through
// b200slam_gn_sparse_host_solve) ---------------------------------------------------------------------------------
namespace b200sparse {

inline bool chol7(double *A /* in: SPD 7x7 row-major; out: W = C^-1 (lower), C C^T = A */) {
  double C[49] = {0}, Wm[49] = {0};
  for (int j = 0; j < 7; j++) {
    double s = A[j * 7 + j];
    for (int k = 0; k < j; k++) s -= C[j * 7 + k] * C[j * 7 + k];
    if (!(s > 0.0)) return false;
    const double d = std::sqrt(s);
    C[j * 7 + j] = d;
    for (int i = j + 1; i < 7; i++) {
      double t = A[i * 7 + j];
      for (int k = 0; k < j; k++) t -= C[i * 7 + k] * C[j * 7 + k];
      C[i * 7 + j] = t / d;
    }
  }
  for (int j = 0; j < 7; j++) {
    Wm[j * 7 + j] = 1.0 / C[j * 7 + j];
    for (int i = j + 1; i < 7; i++) {
      double t = 0.0;
      for (int k = j; k < i; k++) t -= C[i * 7 + k] * Wm[k * 7 + j];
      Wm[i * 7 + j] = t / C[i * 7 + i];
    }
  }
  for (int t = 0; t < 49; t++) A[t] = Wm[t];
  return true;
}

// eb: (E,56) [B | g].  x: 7M doubles in ORIGINAL pose order (solution of H x = b).  false: not positive definite.
inline bool solve_host(const Plan &P, const double *eb, double *x) {
  const int M = P.M;
  std::vector<double> Dg((size_t)M * 49, 0.0), L((size_t)P.nL * 49, 0.0), y((size_t)M * 7, 0.0), xs((size_t)M * 7, 0.0);
  for (int b = 0; b < M + P.nL; b++) {
    double *dst = b < M ? &Dg[(size_t)b * 49] : &L[(size_t)(b - M) * 49];
    for (int t = P.asm_ptr[b]; t < P.asm_ptr[b + 1]; t++) {
      const int code = P.asm_ent[t];
      const double *B = eb + (size_t)(code >> 1) * 56;
      for (int q = 0; q < 49; q++) dst[q] += (code & 1) ? -B[q] : B[q];
    }
  }
  for (int k = 0; k < M; k++)
    for (int t = P.rhs_ptr[k]; t < P.rhs_ptr[k + 1]; t++) {
      const int code = P.rhs_ent[t];
      const double *g = eb + (size_t)(code >> 1) * 56 + 49;
      for (int q = 0; q < 7; q++) y[(size_t)k * 7 + q] += (code & 1) ? g[q] : -g[q];
    }
  for (int l = 0; l < P.levels; l++)
    for (int c = P.level_ptr[l]; c < P.level_ptr[l + 1]; c++) {
      const int j = P.level_cols[c];
      double *D = &Dg[(size_t)j * 49];
      for (int t = P.upd_ptr[j]; t < P.upd_ptr[j + 1]; t++) {
        const double *A = &L[(size_t)P.upd[3 * t] * 49], *Bk = &L[(size_t)P.upd[3 * t + 1] * 49];
        double *dst = P.upd[3 * t + 2] < 0 ? D : &L[(size_t)P.upd[3 * t + 2] * 49];
        for (int r = 0; r < 7; r++)
          for (int cc = 0; cc < 7; cc++) {
            double s = 0.0;
            for (int q = 0; q < 7; q++) s += A[r * 7 + q] * Bk[cc * 7 + q];  // L_ik L_jk^T
            dst[r * 7 + cc] -= s;
          }
      }
      if (!chol7(D)) return false;  // D now holds W = C^-1
      for (int s = P.colptr[j]; s < P.colptr[j + 1]; s++) {  // L_ij = A_ij C^-T = A_ij W^T
        double *A = &L[(size_t)s * 49], T[49];
        for (int r = 0; r < 7; r++)
          for (int cc = 0; cc < 7; cc++) {
            double v = 0.0;
            for (int q = 0; q <= cc; q++) v += A[r * 7 + q] * D[cc * 7 + q];
            T[r * 7 + cc] = v;
          }
        for (int q = 0; q < 49; q++) A[q] = T[q];
      }
      double rhs[7];
      for (int q = 0; q < 7; q++) rhs[q] = y[(size_t)j * 7 + q];
      for (int t = P.row_ptr[j]; t < P.row_ptr[j + 1]; t++) {
        const int k = P.row_ent[2 * t];
        const double *Lk = &L[(size_t)P.row_ent[2 * t + 1] * 49];
        for (int r = 0; r < 7; r++)
          for (int q = 0; q < 7; q++) rhs[r] -= Lk[r * 7 + q] * y[(size_t)k * 7 + q];
      }
      for (int r = 0; r < 7; r++) {
        double v = 0.0;
        for (int q = 0; q <= r; q++) v += D[r * 7 + q] * rhs[q];
        y[(size_t)j * 7 + r] = v;
      }
    }
  for (int l = P.levels - 1; l >= 0; l--)
    for (int c = P.level_ptr[l]; c < P.level_ptr[l + 1]; c++) {
      const int j = P.level_cols[c];
      const double *D = &Dg[(size_t)j * 49];
      double rhs[7];
      for (int q = 0; q < 7; q++) rhs[q] = y[(size_t)j * 7 + q];
      for (int s = P.colptr[j]; s < P.colptr[j + 1]; s++) {
        const double *Ls = &L[(size_t)s * 49], *xi = &xs[(size_t)P.rowidx[s] * 7];
        for (int q = 0; q < 7; q++)
          for (int r = 0; r < 7; r++) rhs[q] -= Ls[r * 7 + q] * xi[r];  // L_ij^T x_i
      }
      for (int r = 0; r < 7; r++) {
        double v = 0.0;
        for (int q = r; q < 7; q++) v += D[q * 7 + r] * rhs[q];  // W^T
        xs[(size_t)j * 7 + r] = v;
      }
    }
  for (int k = 0; k < M; k++)
    for (int q = 0; q < 7; q++) x[(size_t)P.perm[k] * 7 + q] = xs[(size_t)k * 7 + q];
  return true;
}

}  // namespace b200sparse
