// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.cpp).
//
// Block-sparse storage of the reduced camera system and a sparse Cholesky for it, so that the oracle can run the
// refinement-sized graphs (BASELINE.json configs 4 and 5: reduced systems of 12 000 and 60 000 unknowns) that a dense
// n x n matrix cannot hold. Ceres' SPARSE_SCHUR does the same job with SuiteSparse / Eigen's SimplicialLLT behind
// `SparseSchurComplementSolver` (schur_complement_solver.cc; third-party, not vendored, version un-pinned): a block
// sparse S over the f-blocks, a fill-reducing ordering, a sparse LL^T. Restated here with textbook algorithms:
//   * pattern of S = cliques of f-blocks per eliminated landmark and per residual block without e-block
//     (what BlockRandomAccessSparseMatrix is given in SparseSchurComplementSolver::InitStorage);
//   * ordering: greedy minimum degree on the block graph (the role AMD plays in SuiteSparse);
//   * numeric factorisation: up-looking sparse Cholesky with an elimination tree (Davis, "Direct Methods for Sparse
//     Linear Systems", ch. 4), scalar entries.
// The result of a linear solve does not depend on the ordering beyond rounding, so the small cases solved with both the
// dense and the sparse path check each other (tests/test_oracle_golden.py).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <queue>
#include <vector>

namespace orc {

// Full symmetric block pattern over f-blocks, CSR by block row, values stored per block row-major (ls_i x ls_j).
struct BlockPattern {
  int nfb = 0, n = 0;                       // f-blocks, scalar dimension
  std::vector<int> off, ls;                 // scalar offset / size per f-block
  std::vector<int> rowptr, col;             // block CSR
  std::vector<int64_t> voff;                // value offset per block
  int64_t nvals = 0;
  std::vector<int> blk_of;                  // scalar -> f-block
  int Slot(int bi, int bj) const {          // -1 if absent
    const int* b0 = col.data() + rowptr[bi]; const int* b1 = col.data() + rowptr[bi + 1];
    const int* p = std::lower_bound(b0, b1, bj);
    return (p != b1 && *p == bj) ? (int)(p - col.data()) : -1;
  }
};

// adjacency lists with lazy de-duplication: cliques of a few dozen blocks are pushed millions of times
struct LazyAdj {
  std::vector<std::vector<int>> rows; std::vector<size_t> clean;
  explicit LazyAdj(int n) : rows(n), clean(n, 0) {}
  void Push(int r, int c) {
    std::vector<int>& v = rows[r]; v.push_back(c);
    if (v.size() > 64 && v.size() > 4 * clean[r]) Compact(r);
  }
  void Compact(int r) { std::vector<int>& v = rows[r]; std::sort(v.begin(), v.end()); v.erase(std::unique(v.begin(), v.end()), v.end()); clean[r] = v.size(); }
  void AddClique(const std::vector<int>& c) { for (int a : c) for (int b : c) Push(a, b); }
};

inline void FinishPattern(LazyAdj* A, const std::vector<int>& ls, BlockPattern* P) {
  const int nfb = (int)ls.size();
  P->nfb = nfb; P->ls = ls; P->off.assign(nfb + 1, 0);
  for (int i = 0; i < nfb; ++i) P->off[i + 1] = P->off[i] + ls[i];
  P->n = P->off[nfb];
  P->blk_of.resize(P->n);
  for (int i = 0; i < nfb; ++i) for (int k = 0; k < ls[i]; ++k) P->blk_of[P->off[i] + k] = i;
  P->rowptr.assign(nfb + 1, 0); P->col.clear(); P->voff.clear(); P->nvals = 0;
  for (int i = 0; i < nfb; ++i) {
    A->Push(i, i); A->Compact(i);
    for (int c : A->rows[i]) { P->col.push_back(c); P->voff.push_back(P->nvals); P->nvals += (int64_t)ls[i] * ls[c]; }
    P->rowptr[i + 1] = (int)P->col.size();
    std::vector<int>().swap(A->rows[i]);
  }
}

// Greedy minimum-degree ordering on the block graph (explicit elimination graph, lazy priority queue).
inline std::vector<int> MinimumDegreeOrder(const BlockPattern& P) {
  const int n = P.nfb;
  std::vector<std::vector<int>> adj(n);
  for (int i = 0; i < n; ++i) for (int k = P.rowptr[i]; k < P.rowptr[i + 1]; ++k) if (P.col[k] != i) adj[i].push_back(P.col[k]);
  typedef std::pair<int, int> DN;   // (degree, node); ties broken by node index: deterministic
  std::priority_queue<DN, std::vector<DN>, std::greater<DN>> pq;
  for (int i = 0; i < n; ++i) pq.push({(int)adj[i].size(), i});
  std::vector<char> done(n, 0);
  std::vector<int> order; order.reserve(n);
  std::vector<int> merged;
  while (!pq.empty()) {
    const DN top = pq.top(); pq.pop();
    const int v = top.second;
    if (done[v] || top.first != (int)adj[v].size()) continue;
    done[v] = 1; order.push_back(v);
    const std::vector<int> nb = adj[v];
    for (int u : nb) {
      merged.clear(); merged.reserve(adj[u].size() + nb.size());
      std::set_union(adj[u].begin(), adj[u].end(), nb.begin(), nb.end(), std::back_inserter(merged));
      merged.erase(std::remove_if(merged.begin(), merged.end(), [&](int w) { return w == u || w == v; }), merged.end());
      adj[u].swap(merged);
      pq.push({(int)adj[u].size(), u});
    }
    std::vector<int>().swap(adj[v]);
  }
  return order;
}

// Up-looking sparse Cholesky of C = P A P^T. Symbolic structure is kept between factorisations of the same pattern.
struct SparseCholesky {
  int n = 0;
  std::vector<int> perm, iperm;             // scalar permutation: new -> old, old -> new
  std::vector<int> Cp, Ci; std::vector<int64_t> Csrc;   // upper triangle of C, CSC; Csrc: index into the block values
  std::vector<int> parent, Lp, Li; std::vector<double> Lx;
  bool analysed = false;

  void Analyse(const BlockPattern& P) {
    n = P.n;
    const std::vector<int> border = MinimumDegreeOrder(P);
    perm.clear(); perm.reserve(n);
    for (int b : border) for (int k = 0; k < P.ls[b]; ++k) perm.push_back(P.off[b] + k);
    iperm.assign(n, 0); for (int k = 0; k < n; ++k) iperm[perm[k]] = k;
    // upper triangle of C in CSC: column j holds rows i <= j
    std::vector<std::vector<std::pair<int, int64_t>>> cols(n);
    for (int bi = 0; bi < P.nfb; ++bi)
      for (int s = P.rowptr[bi]; s < P.rowptr[bi + 1]; ++s) {
        const int bj = P.col[s];
        for (int a = 0; a < P.ls[bi]; ++a) for (int b = 0; b < P.ls[bj]; ++b) {
          const int i = iperm[P.off[bi] + a], j = iperm[P.off[bj] + b];
          if (i <= j) cols[j].push_back({i, P.voff[s] + (int64_t)a * P.ls[bj] + b});
        }
      }
    Cp.assign(n + 1, 0); Ci.clear(); Csrc.clear();
    for (int j = 0; j < n; ++j) {
      std::sort(cols[j].begin(), cols[j].end());
      for (auto& e : cols[j]) { Ci.push_back(e.first); Csrc.push_back(e.second); }
      Cp[j + 1] = (int)Ci.size();
    }
    // elimination tree
    parent.assign(n, -1);
    std::vector<int> anc(n, -1);
    for (int k = 0; k < n; ++k)
      for (int p = Cp[k]; p < Cp[k + 1]; ++p)
        for (int i = Ci[p]; i != -1 && i < k;) { const int nx = anc[i]; anc[i] = k; if (nx == -1) parent[i] = k; i = nx; }
    // column counts by walking the row subtrees once
    std::vector<int> cnt(n, 1), mark(n, -1);
    for (int k = 0; k < n; ++k) {
      mark[k] = k;
      for (int p = Cp[k]; p < Cp[k + 1]; ++p)
        for (int i = Ci[p]; i < k && mark[i] != k; i = parent[i]) { mark[i] = k; cnt[i]++; }
    }
    Lp.assign(n + 1, 0);
    for (int k = 0; k < n; ++k) Lp[k + 1] = Lp[k] + cnt[k];
    Li.assign(Lp[n], 0); Lx.assign(Lp[n], 0.0);
    analysed = true;
  }

  bool Factor(const std::vector<double>& vals) {
    std::vector<int> c(Lp.begin(), Lp.end() - 1), s(n), mark(n, -1);
    std::vector<double> x(n, 0.0);
    for (int k = 0; k < n; ++k) {
      // pattern of row k of L in topological order (ereach)
      int top = n; mark[k] = k;
      for (int p = Cp[k]; p < Cp[k + 1]; ++p) {
        int i = Ci[p]; if (i > k) continue;
        x[i] = vals[Csrc[p]];
        int len = 0;
        for (; mark[i] != k; i = parent[i]) { s[len++] = i; mark[i] = k; }
        while (len > 0) s[--top] = s[--len];
      }
      double d = x[k]; x[k] = 0.0;
      for (; top < n; ++top) {
        const int i = s[top];
        const double lki = x[i] / Lx[Lp[i]]; x[i] = 0.0;
        for (int p = Lp[i] + 1; p < c[i]; ++p) x[Li[p]] -= Lx[p] * lki;
        d -= lki * lki;
        const int p = c[i]++; Li[p] = k; Lx[p] = lki;
      }
      if (!(d > 0.0) || !std::isfinite(d)) return false;
      const int p = c[k]++; Li[p] = k; Lx[p] = std::sqrt(d);
    }
    return true;
  }
  // b (original order) <- A^-1 b
  void Solve(double* b) const {
    std::vector<double> y(n);
    for (int k = 0; k < n; ++k) y[k] = b[perm[k]];
    for (int j = 0; j < n; ++j) { y[j] /= Lx[Lp[j]]; for (int p = Lp[j] + 1; p < Lp[j + 1]; ++p) y[Li[p]] -= Lx[p] * y[j]; }
    for (int j = n - 1; j >= 0; --j) { for (int p = Lp[j] + 1; p < Lp[j + 1]; ++p) y[j] -= Lx[p] * y[Li[p]]; y[j] /= Lx[Lp[j]]; }
    for (int k = 0; k < n; ++k) b[perm[k]] = y[k];
  }
};

}  // namespace orc
