// ORACLE / TEST INFRASTRUCTURE -- CPU restatement of Tree::buildByUPGMA (reference tree.cpp:362-454) and
// Tree::buildByNeighborJoining (tree.cpp:240-352) on plain arrays: same loops over an ordered set of active nodes, same
// arithmetic.  Output as hist_build_tree: child [2n-1][2], branch_length [2n-1] (before the reference's 6-digit
// serialise / parse round trip).  Pinned: the Newick text built from it equals the reference's on tests/golden/distmat_*.npz.
// Only tests/ may use this file.
#include <algorithm>
#include <cstdint>
#include <set>
#include <vector>

extern "C" int oracle_build_tree(int32_t n, const double* dm, int32_t method, double minLen, int32_t* child, double* blen) {
  const int M = 2 * n - 1;
  std::vector<std::vector<double> > dist(n, std::vector<double>(n));
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) dist[i][j] = dm[(size_t)i * n + j];
  std::set<int> active;
  for (int k = 0; k < n; ++k) active.insert(k);
  for (int k = 0; k < M; ++k) { child[2 * k] = child[2 * k + 1] = -1; blen[k] = 0; }
  std::vector<double> height(n, 0), avg;
  int nodes = n;
  while (active.size() > 2) {
    const int nAct = (int)active.size();
    if (method == 1) {
      avg.assign(nodes, 0.0);
      for (int ni : active) {
        double a = 0;
        for (int nj : active)
          if (nj != ni) a += dist[ni][nj];
        avg[ni] = a / (double)(nAct - 2);
      }
    }
    bool first = true;
    double minDist = 0;
    int mi = -1, mj = -1;
    for (auto pi = active.begin(); pi != active.end(); ++pi)
      for (auto pj = std::next(pi); pj != active.end(); ++pj) {
        const double d = method == 1 ? dist[*pi][*pj] - avg[*pi] - avg[*pj] : dist[*pi][*pj];
        if (first || d < minDist) { mi = *pi; mj = *pj; minDist = d; first = false; }
      }
    const int k = nodes++;
    dist.push_back(std::vector<double>(k + 1));
    dist[k][k] = 0;
    const double dij = dist[mi][mj];
    double dik, djk;
    if (method == 1) {
      for (int m = 0; m < k; ++m) dist[m].push_back(dist[k][m] = 0.5 * (dist[mi][m] + dist[mj][m] - dij));
      dik = 0.5 * (dij + avg[mi] - avg[mj]);
      djk = dij - dik;
      if (dik < minLen) { djk -= dik - minLen; dik = minLen; }
      if (djk < 0) { dik -= djk - minLen; djk = minLen; }
    } else {
      height.push_back(std::max(height[mi] + minLen, std::max(height[mj] + minLen, (height[mi] + height[mj] + dij) / 2)));
      dik = height[k] - height[mi];
      djk = height[k] - height[mj];
      for (int m = 0; m < k; ++m) dist[m].push_back(dist[k][m] = (dist[mi][m] + dist[mj][m]) / 2);
    }
    dist[mi][k] = dist[k][mi] = dik;
    dist[mj][k] = dist[k][mj] = djk;
    child[2 * k] = mi; child[2 * k + 1] = mj;
    blen[mi] = std::max(0., dik);
    blen[mj] = std::max(0., djk);
    active.erase(mi);
    active.erase(mj);
    active.insert(k);
  }
  auto it = active.begin();
  const int i = *it, j = *++it, k = nodes;
  child[2 * k] = i; child[2 * k + 1] = j;
  if (method == 1) {
    const double d = std::max(dist[i][j], 0.);
    blen[i] = std::max(0., d / 2);
    blen[j] = std::max(0., d / 2);
  } else {
    const double hk = std::max(height[i] + minLen, std::max(height[j] + minLen, (height[i] + height[j] + dist[i][j]) / 2));
    blen[i] = std::max(0., hk - height[i]);
    blen[j] = std::max(0., hk - height[j]);
  }
  return 0;
}
