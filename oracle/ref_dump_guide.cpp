// ORACLE INFRASTRUCTURE: golden-vector generator for the guide-alignment stage as a whole, AlignGraph (span.cpp:35-158)
// over QuickAlignMatrix, DiagonalEnvelope and alignPathMerge.  Compiled against the UNMODIFIED reference (oracle/Makefile).
// Builds the graph the way `historian recon -seqs` does (recon.cpp:806-821: time 1, the generator re-seeded before the
// random graph), takes the spanning tree and prints the merged guide alignment as gapped FASTA on stdout; the edges with
// their scores and the tree's joins go to stderr ("edge a b lp" / "join a b").
//
//   ref_dump_guide <preset> <seqs.fa> <time> rnd|all <seed> full | sparse:<band>:<kmerLen>:<kmerThreshold>:<maxSize>
//
// tools/make_golden_guide.py turns the output into tests/golden/guide_*.npz.
#include <chrono>
#include <cstdio>
#include <iostream>
#include <string>

#include "diagenv.h"
#include "fastseq.h"
#include "logger.h"
#include "model.h"
#include "presets.h"
#include "quickalign.h"
#include "span.h"
#include "util.h"

using namespace std;

int main(int argc, char** argv) {
  if (argc != 7) {
    cerr << "Usage: " << argv[0] << " <preset> <seqs.fa> <time> rnd|all <seed> full|sparse:<band>:<kmerLen>:<kmerThreshold>:<maxSize>\n";
    return 1;
  }
  const RateModel model = namedModel(argv[1]);
  const vguard<FastSeq> seqs = readFastSeqs(argv[2]);
  const double time = atof(argv[3]);
  const bool all = string(argv[4]) == "all";
  const unsigned seed = (unsigned)atol(argv[5]);
  const string mode = argv[6];
  DiagEnvParams dep;
  if (mode == "full")
    dep.sparse = false;
  else {
    int band = 0, kmerLen = 0, kmerThreshold = 0;
    long long maxSize = 0;
    if (sscanf(mode.c_str(), "sparse:%d:%d:%d:%lld", &band, &kmerLen, &kmerThreshold, &maxSize) != 4) {
      cerr << "bad mode " << mode << endl;
      return 1;
    }
    dep.sparse = true;
    dep.bandSize = band;
    dep.kmerLen = kmerLen;
    dep.kmerThreshold = kmerThreshold;
    dep.autoMemSize = false;
    dep.maxSize = (size_t)maxSize;
  }
  {   // the score table every pairwise alignment uses (quickalign.cpp:22-50)
    DiagonalEnvelope env(seqs[0], seqs[1]);
    env.initFull();
    const QuickAlignMatrix mx(env, model, time);
    fprintf(stderr, "alphabet %s\n", model.alphabet.c_str());
    fprintf(stderr, "scores %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", mx.noGap, mx.gapOpen, mx.gapExtend, mx.m2m, mx.m2i, mx.m2d,
            mx.i2i, mx.i2m, mx.i2d, mx.d2d, mx.d2m);
    fprintf(stderr, "submat");
    for (size_t i = 0; i < model.alphabet.size(); ++i)
      for (size_t j = 0; j < model.alphabet.size(); ++j) fprintf(stderr, " %.17g", mx.submat[i][j]);
    fprintf(stderr, "\n");
  }
  const auto t0 = chrono::steady_clock::now();
  AlignGraph* ag = NULL;
  if (all)
    ag = new AlignGraph(seqs, model, time, dep);
  else {
    ForwardMatrix::random_engine generator = ForwardMatrix::newRNG();
    generator.seed(seed);
    ag = new AlignGraph(seqs, model, time, dep, generator);
  }
  const double tGraph = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
  for (size_t src = 0; src < seqs.size(); ++src)
    for (const auto& dest_path : ag->edgePath[src]) {
      // the score sits in the priority queues: find the edge in the source's queue
      auto q = ag->edges[src];
      double lp = 0;
      while (!q.empty()) {
        if (q.top().row1 == src && q.top().row2 == dest_path.first) lp = q.top().lp;
        q.pop();
      }
      fprintf(stderr, "edge %zu %zu %.17g\n", src, (size_t)dest_path.first, lp);
    }
  const auto t1 = chrono::steady_clock::now();
  const vguard<FastSeq> gapped = ag->mstGapped();
  const double tMerge = chrono::duration<double>(chrono::steady_clock::now() - t1).count();
  writeFastaSeqs(cout, gapped);
  fprintf(stderr, "timing graph %.6f merge %.6f\n", tGraph, tMerge);
  delete ag;
  return 0;
}
