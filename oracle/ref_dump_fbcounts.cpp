// ORACLE INFRASTRUCTURE: golden-vector generator for BackwardMatrix::getCounts (forward.cpp:1183-1214).
// Compiled against the UNMODIFIED reference (oracle/Makefile).  For a two-leaf tree it builds the leaf
// profiles, the PairHMM and the SumProduct exactly as Reconstructor::reconstruct does (recon.cpp:932-1013),
// runs the reference's Forward and Backward fills and dumps getCounts() in the model's natural basis
// (EigenCounts::transform, model.cpp:924-931) so the fixture does not depend on an eigenvector convention.
// tools/make_golden_fbcounts.py turns the dump into tests/golden/fbcounts_*.npz.
//
//   ref_dump_fbcounts <out.bin> <preset:NAME[:bins:shape]> <gapped2.fa> <tree.nh> <band | -1>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iostream>
#include <string>

#include "alignpath.h"
#include "fastseq.h"
#include "forward.h"
#include "gamma.h"
#include "jsonutil.h"
#include "logger.h"
#include "model.h"
#include "presets.h"
#include "sumprod.h"
#include "tree.h"
#include "util.h"

using namespace std;

static FILE* gOut = NULL;

static void putRaw(const string& name, char dtype, const vector<uint64_t>& shape, const void* data, size_t bytes) {
  const uint32_t nl = (uint32_t)name.size();
  fwrite(&nl, 4, 1, gOut);
  fwrite(name.data(), 1, nl, gOut);
  fwrite(&dtype, 1, 1, gOut);
  const uint32_t nd = (uint32_t)shape.size();
  fwrite(&nd, 4, 1, gOut);
  for (auto d : shape) fwrite(&d, 8, 1, gOut);
  const uint64_t nb = bytes;
  fwrite(&nb, 8, 1, gOut);
  if (bytes) fwrite(data, 1, bytes, gOut);
}
static void putD(const string& n, const vector<double>& v, vector<uint64_t> shape) { putRaw(n, 'd', shape, v.data(), v.size() * 8); }
static void putS(const string& n, const string& s) { putRaw(n, 's', {(uint64_t)s.size()}, s.data(), s.size()); }

int main(int argc, char** argv) {
  if (argc != 6) {
    cerr << "Usage: " << argv[0] << " <out.bin> <preset:NAME[:bins:shape]> <gapped2.fa> <tree.nh> <band|-1>\n";
    return 1;
  }
  gOut = fopen(argv[1], "wb");
  const string spec = argv[2];
  vector<string> f;
  for (size_t p = 7;;) {
    const size_t q = spec.find(':', p);
    f.push_back(spec.substr(p, q == string::npos ? string::npos : q - p));
    if (q == string::npos) break;
    p = q + 1;
  }
  RateModel model = namedModel(f[0]);
  if (f.size() == 3) model = makeDiscretizedGammaModel(model, atoi(f[1].c_str()), atof(f[2].c_str()));
  vguard<FastSeq> gapped = readFastSeqs(argv[3]);
  ifstream treeStream(argv[4]);
  Tree tree(JsonUtil::readStringFromStream(treeStream));
  const int band = atoi(argv[5]);
  if (tree.nodes() != 3 || gapped.size() != 2 || !tree.isLeaf(0) || !tree.isLeaf(1)) {
    cerr << "two-leaf trees only" << endl;
    return 1;
  }
  if (gapped[0].name != tree.nodeName(0)) swap(gapped[0], gapped[1]);  // alignment row == tree node index
  if (gapped[0].name != tree.nodeName(0) || gapped[1].name != tree.nodeName(1)) {
    cerr << "sequence names do not match the tree's leaves" << endl;
    return 1;
  }
  const Alignment align(gapped);
  const int C = model.components(), A = model.alphabetSize();
  const TreeNodeIndex root = tree.root(), l = tree.getChild(root, 0), r = tree.getChild(root, 1);

  const Profile lProf(C, model.alphabet, align.ungapped[l], l), rProf(C, model.alphabet, align.ungapped[r], r);
  ProbModel lProbs(model, tree.branchLength(l)), rProbs(model, tree.branchLength(r));
  PairHMM hmm(lProbs, rProbs, model.insProb);
  SumProduct sumProd(model, tree);
  const GuideAlignmentEnvelope env = band < 0 ? GuideAlignmentEnvelope() : GuideAlignmentEnvelope(align.path, l, r, band);
  ForwardMatrix forward(lProf, rProf, hmm, root, env, &sumProd);
  BackwardMatrix backward(forward);
  const EigenCounts eigenCounts = backward.getCounts();
  const EventCounts ev = eigenCounts.transform(model);

  putS("preset", spec.substr(7));
  putS("x_seq", align.ungapped[l].seq);
  putS("y_seq", align.ungapped[r].seq);
  putS("x_row", gapped[l].seq);
  putS("y_row", gapped[r].seq);
  putD("branch", {tree.branchLength(l), tree.branchLength(r)}, {2});
  putD("band", {(double)band}, {1});
  putD("lp_end", {forward.lpEnd}, {1});
  putD("waits", {lProbs.t, rProbs.t, lProbs.insWait, lProbs.delWait, rProbs.insWait, rProbs.delWait}, {6});
  const IndelCounts& ic = ev.indelCounts;
  putD("indel", {ic.ins, ic.del, ic.insExt, ic.delExt, ic.insTime, ic.delTime, ic.lp}, {7});
  vector<double> rootCount((size_t)C * A), subCount((size_t)C * A * A);
  for (int c = 0; c < C; ++c)
    for (int i = 0; i < A; ++i) {
      rootCount[(size_t)c * A + i] = ev.rootCount[c][i];
      for (int j = 0; j < A; ++j) subCount[((size_t)c * A + i) * A + j] = ev.subCount[c][i][j];
    }
  putD("root_count", rootCount, {(uint64_t)C, (uint64_t)A});
  putD("sub_count", subCount, {(uint64_t)C, (uint64_t)A, (uint64_t)A});
  fclose(gOut);
  return 0;
}
