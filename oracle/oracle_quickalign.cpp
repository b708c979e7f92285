// ORACLE / TEST INFRASTRUCTURE -- CPU restatement of the guide-alignment stage's pairwise step, QuickAlignMatrix
// (reference quickalign.h:8-76, quickalign.cpp:7-214) over a DiagonalEnvelope given by its diagonal list
// (diagenv.h:38-155).  Row-major loops and a dense (xLen+1) x (yLen+1) x 3 cell array, like the reference (which stores
// only the envelope's storage diagonals; cells outside the envelope read as -inf either way, quickalign.h:30-33).
// Pinned: bit-equal scores, end cells and paths on tests/golden/quickalign_*.npz (tests/test_quickalign.py).
// Only tests/ may use this file.
#include <algorithm>
#include <cstdint>
#include <limits>
#include <vector>

#include "historian_b200.h"

namespace {
const double NEG = -std::numeric_limits<double>::infinity();

struct QA {
  const hist_quickalign_scores& S;
  unsigned xLen, yLen;
  double startGap(unsigned i, unsigned j) const {   // quickalign.h:60-63 (SeqIdx is a 32-bit unsigned)
    return (i == 1 ? S.no_gap : (S.gap_open + (i - 2) * S.gap_extend)) + (j == 1 ? S.no_gap : (S.gap_open + (j - 2) * S.gap_extend));
  }
  double endGap(unsigned i, unsigned j) const {     // quickalign.h:65-68, unsigned wrap-around included
    return (i == xLen ? S.no_gap : (S.gap_open + (xLen - i - 2) * S.gap_extend)) + (j == yLen ? S.no_gap : (S.gap_open + (yLen - j - 2) * S.gap_extend));
  }
};
}  // namespace

extern "C" int oracle_quickalign(const hist_quickalign_pair* p, const hist_quickalign_scores* S, double* result, int32_t* x_end,
                                 int32_t* y_end, uint8_t* path, int32_t* path_len) {
  const unsigned xLen = p->x_len, yLen = p->y_len;
  const int A = S->alph_size;
  const QA q{*S, xLen, yLen};
  std::vector<double> cell((size_t)(xLen + 1) * (yLen + 1) * 3, NEG);
  auto at = [&](unsigned i, unsigned j, int s) -> double& { return cell[((size_t)i * (yLen + 1) + j) * 3 + s]; };
  auto emit = [&](unsigned i, unsigned j) {   // matchEmitScore, quickalign.h:41-45
    const int xt = p->x_tok[i - 1], yt = p->y_tok[j - 1];
    return (xt < 0 || yt < 0 || xt >= A || yt >= A) ? 0.0 : S->submat[xt * A + yt];
  };
  const std::vector<int> diags(p->diagonals, p->diagonals + p->n_diagonals);
  const double start = 0;
  double end = NEG;
  unsigned xEnd = 0, yEnd = 0;
  for (unsigned j = 1; j <= yLen; ++j) {   // quickalign.cpp:60-94
    const auto b = std::upper_bound(diags.begin(), diags.end(), -(int)j), e = std::upper_bound(diags.begin(), diags.end(), (int)xLen - (int)j);
    for (auto it = b; it != e; ++it) {
      const unsigned i = (unsigned)(*it + (int)j);
      double m = std::max(std::max(at(i - 1, j - 1, 0) + S->m2m, at(i - 1, j - 1, 2) + S->d2m), at(i - 1, j - 1, 1) + S->i2m);
      m = std::max(m, start + q.startGap(i, j));
      m += emit(i, j);
      at(i, j, 0) = m;
      at(i, j, 1) = std::max(at(i, j - 1, 1) + S->i2i, at(i, j - 1, 0) + S->m2i);
      at(i, j, 2) = std::max(std::max(at(i - 1, j, 1) + S->i2d, at(i - 1, j, 2) + S->d2d), at(i - 1, j, 0) + S->m2d);
      const double ijEnd = m + q.endGap(i, j);
      if (ijEnd > end) { xEnd = i; yEnd = j; end = ijEnd; }
    }
  }
  *result = end;
  *x_end = (int32_t)xEnd;
  *y_end = (int32_t)yEnd;
  *path_len = 0;
  if (!(end > NEG) || !path) return 0;
  // alignPath, quickalign.cpp:144-207
  std::vector<uint8_t> rev;
  unsigned i = xEnd, j = yEnd;
  int state = 1;   // Start 0, Match 1, Insert 2, Delete 3
  auto upd = [](double& cur, int& st, double cand, int cs) { if (cand > cur) { cur = cand; st = cs; } };
  while (state != 0) {
    double sc = NEG;
    if (state == 1) {
      const double em = emit(i, j);
      --i; --j;
      rev.push_back(3);
      upd(sc, state, at(i, j, 0) + S->m2m + em, 1);
      upd(sc, state, at(i, j, 1) + S->i2m + em, 2);
      upd(sc, state, at(i, j, 2) + S->d2m + em, 3);
      upd(sc, state, start + q.startGap(i + 1, j + 1) + em, 0);
      if (sc != at(i + 1, j + 1, 0)) return -1;
    } else if (state == 2) {
      --j;
      rev.push_back(2);
      upd(sc, state, at(i, j, 0) + S->m2i, 1);
      upd(sc, state, at(i, j, 1) + S->i2i, 2);
      if (sc != at(i, j + 1, 1)) return -1;
    } else {
      --i;
      rev.push_back(1);
      upd(sc, state, at(i, j, 0) + S->m2d, 1);
      upd(sc, state, at(i, j, 1) + S->i2d, 2);
      upd(sc, state, at(i, j, 2) + S->d2d, 3);
      if (sc != at(i + 1, j, 2)) return -1;
    }
  }
  int c = 0;
  for (unsigned k = 0; k < j; ++k) path[c++] = 2;
  for (unsigned k = 0; k < i; ++k) path[c++] = 1;
  for (size_t k = rev.size(); k-- > 0;) path[c++] = rev[k];
  for (unsigned k = 0; k < xLen - xEnd; ++k) path[c++] = 1;
  for (unsigned k = 0; k < yLen - yEnd; ++k) path[c++] = 2;
  *path_len = c;
  return 0;
}
