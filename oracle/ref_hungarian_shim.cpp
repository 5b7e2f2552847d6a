// oracle/ref_hungarian_shim.cpp - TEST INFRASTRUCTURE.  C entry point around the reference's OWN vendored assignment
// solver (third-party/hungarian-v2.0, compiled from where it lies under /root/reference by `make _ref`; no reference
// source is copied into this repo).  It drives the library exactly like the reference's wrapper does
// (src/hungarian/hungarian.cpp): ref_find_assignment follows findAssignment(const Eigen::MatrixXi&) (:57-84,
// transpose when rows > cols), ref_find_assignment_triples follows findAssignment(triples) (:8-54, the overload
// src/segmentation_evaluation.cpp:148-157 calls: labels numbered in first-encounter order, no transpose); both fill
// one Matrix cell per weight, run BipartiteGraph -> Hungarian::HungarianAlgo and read the matching from Hungarian::M.
// Used only by tests/ to pin the evaluation scorer (include/tcs/evaluation.h, oracle orc_score) against real
// reference code.
#include <cstdint>
#include <cstdio>
#include <map>
#include <vector>
#include <unistd.h>

#include "Assignment.h"
#include "BipartiteGraph.h"
#include "CmdParser.h"  // defines the library's globals _Verbose / _Plot (zero, as in the reference, which never parses)
#include "Hungarian.h"

// weights: rows x cols row-major, non-negative.  out_col[r] = column matched to row r or -1.
// Returns the number of matched pairs.
extern "C" int ref_find_assignment(const int32_t* weights, int rows_in, int cols_in, int32_t* out_col) {
  const bool transpose = rows_in > cols_in;
  const int rows = !transpose ? rows_in : cols_in;
  const int cols = !transpose ? cols_in : rows_in;
  Matrix m(rows);
  for (int i = 0; i < rows; ++i) {
    m[i].resize(cols);
    for (int j = 0; j < cols; ++j)
      m[i][j].SetWeight(!transpose ? weights[(size_t)i * cols_in + j] : weights[(size_t)j * cols_in + i]);
  }
  for (int r = 0; r < rows_in; ++r) out_col[r] = -1;
  // the library prints its progress on stdout; keep the test log clean
  fflush(stdout);
  int saved = dup(1);
  FILE* nul = freopen("/dev/null", "w", stdout);
  (void)nul;
  // BipartiteGraph(Matrix&) leaves num_matched uninitialised (the reference's wrapper relies on fresh stack memory
  // being zero); the (agents, tasks) constructor zeroes it, ConstructBG then does what BipartiteGraph(Matrix&) does
  BipartiteGraph bg(m.size(), m[0].size());
  bg.ConstructBG(m);
  Hungarian h(bg);
  h.HungarianAlgo();
  fflush(stdout);
  dup2(saved, 1);
  close(saved);
  int matched = 0;
  for (const auto& pr : h.M) {
    int r = !transpose ? (int)pr.first : (int)pr.second;
    int c = !transpose ? (int)pr.second : (int)pr.first;
    if (r >= 0 && r < rows_in && c >= 0 && c < cols_in) { out_col[r] = c; ++matched; }
  }
  return matched;
}

// (l1[i], l2[i], count[i]) triples as segmentation_evaluation.cpp:148-154 emits them.  Writes the matched
// (l1, l2) pairs to out_l1 / out_l2 (capacity n) and returns how many.
extern "C" int ref_find_assignment_triples(const uint32_t* l1, const uint32_t* l2, const uint64_t* count, int n,
                                           uint32_t* out_l1, uint32_t* out_l2) {
  std::map<uint32_t, size_t> m1, m2;
  std::vector<uint32_t> inv1, inv2;
  for (int i = 0; i < n; ++i) {
    if (!m1.count(l1[i])) { m1[l1[i]] = inv1.size(); inv1.push_back(l1[i]); }
    if (!m2.count(l2[i])) { m2[l2[i]] = inv2.size(); inv2.push_back(l2[i]); }
  }
  Matrix m(inv1.size());
  for (size_t i = 0; i < inv1.size(); ++i) m[i].resize(inv2.size());
  for (int i = 0; i < n; ++i) m[m1[l1[i]]][m2[l2[i]]].SetWeight((double)count[i]);
  fflush(stdout);
  int saved = dup(1);
  FILE* nul = freopen("/dev/null", "w", stdout);
  (void)nul;
  // BipartiteGraph(Matrix&) leaves num_matched uninitialised (the reference's wrapper relies on fresh stack memory
  // being zero); the (agents, tasks) constructor zeroes it, ConstructBG then does what BipartiteGraph(Matrix&) does
  BipartiteGraph bg(m.size(), m[0].size());
  bg.ConstructBG(m);
  Hungarian h(bg);
  h.HungarianAlgo();
  fflush(stdout);
  dup2(saved, 1);
  close(saved);
  int k = 0;
  for (const auto& pr : h.M) {
    if (pr.first >= inv1.size() || pr.second >= inv2.size() || k >= n) continue;
    out_l1[k] = inv1[pr.first];
    out_l2[k] = inv2[pr.second];
    ++k;
  }
  return k;
}
