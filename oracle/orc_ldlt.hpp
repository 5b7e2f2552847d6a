// oracle/orc_ldlt.hpp - CPU ORACLE (test infrastructure, see orc.h): sparse simplicial LDL^T.
// Stands in for Eigen::SimplicialCholesky<SparseMatrix<float>, Lower> (default mode LDL^T with a
// fill-reducing ordering) used at impl/random_walker.hpp:194-203.  Eigen is not in this image; this is the
// textbook up-looking algorithm (elimination tree + sparse triangular solves, Davis, "Direct Methods for
// Sparse Linear Systems", ch. 4) with an exact minimum-degree ordering on the quotient graph.  The pivot
// order differs from Eigen's AMD, so float results agree with the reference to rounding, not bit-wise.
#ifndef TCS_ORACLE_LDLT_HPP
#define TCS_ORACLE_LDLT_HPP

#include <algorithm>
#include <set>
#include <vector>

namespace orc {

// Symmetric sparse matrix given by its strictly-lower entries (row > col) and its diagonal.
struct SymPattern {
  int n = 0;
  std::vector<int> row, col;  // strictly lower triplets
};

// Exact minimum (external) degree ordering with element absorption (George & Liu quotient graph).
inline std::vector<int> minimum_degree_order(int n, const std::vector<int>& ti, const std::vector<int>& tj) {
  std::vector<std::vector<int>> adjv(n), adje(n), elem(n);
  for (size_t k = 0; k < ti.size(); ++k) {
    if (ti[k] == tj[k]) continue;
    adjv[ti[k]].push_back(tj[k]);
    adjv[tj[k]].push_back(ti[k]);
  }
  std::vector<int> degree(n), status(n, 0), mark(n, -1), mark2(n, -1), order;
  order.reserve(n);
  for (int v = 0; v < n; ++v) {
    std::sort(adjv[v].begin(), adjv[v].end());
    adjv[v].erase(std::unique(adjv[v].begin(), adjv[v].end()), adjv[v].end());
    degree[v] = (int)adjv[v].size();
  }
  std::set<std::pair<int, int>> queue;
  for (int v = 0; v < n; ++v) queue.insert(std::make_pair(degree[v], v));
  int stamp = 0, stamp2 = 0;
  std::vector<int> lp;
  while (!queue.empty()) {
    int p = queue.begin()->second;
    queue.erase(queue.begin());
    order.push_back(p);
    ++stamp;
    lp.clear();
    mark[p] = stamp;
    for (int u : adjv[p])
      if (status[u] == 0 && mark[u] != stamp) { mark[u] = stamp; lp.push_back(u); }
    for (int e : adje[p]) {
      if (status[e] != 1) continue;
      for (int u : elem[e])
        if (status[u] == 0 && mark[u] != stamp) { mark[u] = stamp; lp.push_back(u); }
      status[e] = 2;  // absorbed into the new element p
      std::vector<int>().swap(elem[e]);
    }
    status[p] = 1;
    std::vector<int>().swap(adjv[p]);
    std::vector<int>().swap(adje[p]);
    elem[p] = lp;
    for (int u : lp) {
      size_t w = 0;
      for (int x : adjv[u])
        if (status[x] == 0 && mark[x] != stamp) adjv[u][w++] = x;
      adjv[u].resize(w);
      w = 0;
      for (int e : adje[u])
        if (status[e] == 1 && e != p) adje[u][w++] = e;
      adje[u].resize(w);
      adje[u].push_back(p);
    }
    for (int u : lp) {
      ++stamp2;
      mark2[u] = stamp2;
      int d = 0;
      for (int x : adjv[u])
        if (mark2[x] != stamp2) { mark2[x] = stamp2; ++d; }
      for (int e : adje[u])
        for (int x : elem[e])
          if (status[x] == 0 && mark2[x] != stamp2) { mark2[x] = stamp2; ++d; }
      if (d != degree[u]) {
        queue.erase(std::make_pair(degree[u], u));
        degree[u] = d;
        queue.insert(std::make_pair(d, u));
      }
    }
  }
  return order;  // order[k] = original index of the k-th pivot
}

template <typename T>
class SimplicialLDLT {
 public:
  // diag[n]; strictly-lower triplets (ti > tj) with values tv (duplicates are summed)
  bool compute(int n, const std::vector<T>& diag, const std::vector<int>& ti, const std::vector<int>& tj,
               const std::vector<T>& tv) {
    n_ = n;
    perm_ = minimum_degree_order(n, ti, tj);
    iperm_.assign(n, 0);
    for (int k = 0; k < n; ++k) iperm_[perm_[k]] = k;
    // permuted matrix, upper triangle by columns (row <= col)
    std::vector<int> cnt(n + 1, 0);
    for (size_t k = 0; k < ti.size(); ++k) {
      int a = iperm_[ti[k]], b = iperm_[tj[k]];
      cnt[std::max(a, b) + 1]++;
    }
    for (int k = 0; k < n; ++k) cnt[k + 1] += cnt[k];
    std::vector<int> ap(cnt), ai(ti.size());
    std::vector<T> ax(ti.size());
    {
      std::vector<int> fill(cnt.begin(), cnt.end() - 1);
      for (size_t k = 0; k < ti.size(); ++k) {
        int a = iperm_[ti[k]], b = iperm_[tj[k]];
        int c = std::max(a, b), r = std::min(a, b);
        ai[fill[c]] = r;
        ax[fill[c]++] = tv[k];
      }
    }
    // symbolic: elimination tree and column counts
    parent_.assign(n, -1);
    std::vector<int> flag(n), lnz(n, 0);
    for (int k = 0; k < n; ++k) {
      flag[k] = k;
      for (int p = ap[k]; p < ap[k + 1]; ++p)
        for (int i = ai[p]; flag[i] != k; i = parent_[i]) {
          if (parent_[i] == -1) parent_[i] = k;
          lnz[i]++;
          flag[i] = k;
        }
    }
    lp_.assign(n + 1, 0);
    for (int k = 0; k < n; ++k) lp_[k + 1] = lp_[k] + lnz[k];
    li_.assign(lp_[n], 0);
    lx_.assign(lp_[n], T(0));
    d_.assign(n, T(0));
    // numeric: row k of L from the sparse triangular solve L(0:k-1,0:k-1) y = A(0:k-1,k)
    std::vector<T> y(n, T(0));
    std::vector<int> pattern(n);
    std::fill(lnz.begin(), lnz.end(), 0);
    ok_ = true;
    for (int k = 0; k < n; ++k) {
      int top = n;
      flag[k] = k;
      y[k] = diag[perm_[k]];
      for (int p = ap[k]; p < ap[k + 1]; ++p) {
        int i = ai[p];
        y[i] += ax[p];
        int len = 0;
        for (; flag[i] != k; i = parent_[i]) { pattern[len++] = i; flag[i] = k; }
        while (len > 0) pattern[--top] = pattern[--len];
      }
      T dk = y[k];
      y[k] = T(0);
      for (; top < n; ++top) {
        int i = pattern[top];
        T yi = y[i];
        y[i] = T(0);
        int p2 = lp_[i] + lnz[i];
        for (int p = lp_[i]; p < p2; ++p) y[li_[p]] -= lx_[p] * yi;
        T lki = yi / d_[i];
        dk -= lki * yi;
        li_[p2] = k;
        lx_[p2] = lki;
        lnz[i]++;
      }
      d_[k] = dk;
      if (dk == T(0) || dk != dk) { ok_ = false; return false; }
    }
    return true;
  }
  // solve A x = b (b, x: n)
  void solve(const T* b, T* x) const {
    std::vector<T> w(n_);
    for (int k = 0; k < n_; ++k) w[k] = b[perm_[k]];
    for (int j = 0; j < n_; ++j) {
      T wj = w[j];
      for (int p = lp_[j]; p < lp_[j + 1]; ++p) w[li_[p]] -= lx_[p] * wj;
    }
    for (int j = 0; j < n_; ++j) w[j] /= d_[j];
    for (int j = n_ - 1; j >= 0; --j) {
      T s = w[j];
      for (int p = lp_[j]; p < lp_[j + 1]; ++p) s -= lx_[p] * w[li_[p]];
      w[j] = s;
    }
    for (int k = 0; k < n_; ++k) x[perm_[k]] = w[k];
  }
  size_t nnz_l() const { return li_.size(); }
  bool ok() const { return ok_; }

 private:
  int n_ = 0;
  bool ok_ = false;
  std::vector<int> perm_, iperm_, parent_, lp_, li_;
  std::vector<T> lx_, d_;
};

}  // namespace orc
#endif
