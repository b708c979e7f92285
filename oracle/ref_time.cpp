// ORACLE INFRASTRUCTURE: CPU baseline timer.  Compiled against the UNMODIFIED reference
// (headers and objects where they lie, see oracle/Makefile), it times the reference's own
// ForwardMatrix and BackwardMatrix constructors (forward.cpp:68-223, 975-1097) — exactly the
// calls Reconstructor::reconstruct makes per sibling merge (recon.cpp:946-984) — on the
// leaf-by-leaf merges of a gapped FASTA: rows (2k, 2k+1) for k in [first, first+count).
//
//   ref_time <preset> <gapped.fa> <band|-1> <tl> <tr> <first> <count> [fwdonly]
// prints one JSON line: {"pairs":..,"cells":..,"fwd_s":..,"bwd_s":..,"lp_sum":..}
// cells = in-envelope (i,j) pairs per matrix, counted as forward.cpp:78-86 visits them.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>

#include "forward.h"
#include "presets.h"
#include "fastseq.h"
#include "alignpath.h"
#include "logger.h"

using namespace std;

int main(int argc, char** argv) {
  if (argc < 8) {
    fprintf(stderr, "usage: ref_time <preset> <gapped.fa> <band|-1> <tl> <tr> <first> <count> [fwdonly]\n");
    return 2;
  }
  const RateModel model = namedModel(argv[1]);
  const vguard<FastSeq> gapped = readFastSeqs(argv[2]);
  const int band = atoi(argv[3]);
  const double tl = atof(argv[4]), tr = atof(argv[5]);
  const size_t first = (size_t)atol(argv[6]), count = (size_t)atol(argv[7]);
  const bool fwdOnly = argc > 8 && !strcmp(argv[8], "fwdonly");
  ProbModel lProbs(model, tl), rProbs(model, tr);
  PairHMM hmm(lProbs, rProbs, model.insProb);
  double fwdS = 0, bwdS = 0, lpSum = 0;
  long long cells = 0;
  size_t pairs = 0;
  for (size_t k = first; k < first + count && 2 * k + 1 < gapped.size(); ++k, ++pairs) {
    vguard<FastSeq> two;
    two.push_back(gapped[2 * k]);
    two.push_back(gapped[2 * k + 1]);
    const Alignment aln(two);
    const Profile x(model.components(), model.alphabet, aln.ungapped[0], 0);
    const Profile y(model.components(), model.alphabet, aln.ungapped[1], 1);
    const GuideAlignmentEnvelope env = band >= 0 ? GuideAlignmentEnvelope(aln.path, 0, 1, band) : GuideAlignmentEnvelope();
    const auto t0 = chrono::steady_clock::now();
    ForwardMatrix fwd(x, y, hmm, 2, env);
    const auto t1 = chrono::steady_clock::now();
    fwdS += chrono::duration<double>(t1 - t0).count();
    lpSum += fwd.lpEnd;
    for (ProfileStateIndex i = 0; i + 1 < fwd.xSize; ++i)
      for (ProfileStateIndex j = 0; j + 1 < fwd.ySize; ++j)
        if (fwd.inEnvelope(i, j)) ++cells;
    if (!fwdOnly) {
      const auto t2 = chrono::steady_clock::now();
      BackwardMatrix bwd(fwd);
      const auto t3 = chrono::steady_clock::now();
      bwdS += chrono::duration<double>(t3 - t2).count();
    }
  }
  printf("{\"pairs\": %zu, \"cells\": %lld, \"fwd_s\": %.6f, \"bwd_s\": %.6f, \"lp_sum\": %.6f}\n", pairs, cells, fwdS, bwdS, lpSum);
  return 0;
}
