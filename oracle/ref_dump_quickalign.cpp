// ORACLE INFRASTRUCTURE: golden-vector generator for the guide-alignment stage, QuickAlignMatrix (quickalign.cpp:7-214)
// inside a DiagonalEnvelope (diagenv.cpp:93-222).  Compiled against the UNMODIFIED reference (oracle/Makefile).
// For every pair (a, b), a < b, of the sequences in a FASTA file it builds the envelope as AlignGraph does
// (span.cpp:99-107), runs the reference's Viterbi fill and traceback and dumps the inputs the device needs (tokens,
// envelope diagonals, the score table the constructor derives) and the results (score, end cell, alignment path).
// tools/make_golden_quickalign.py turns the dump into tests/golden/quickalign_*.npz.
//
//   ref_dump_quickalign <out.bin> <preset> <seqs.fa> <time> full | sparse:<band>:<kmerLen>:<kmerThreshold>[:<maxSize>]
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <string>

#include "diagenv.h"
#include "fastseq.h"
#include "logger.h"
#include "model.h"
#include "presets.h"
#include "quickalign.h"
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
static void putD(const string& n, const vector<double>& v) { putRaw(n, 'd', {(uint64_t)v.size()}, v.data(), v.size() * 8); }
static void putI(const string& n, const vector<int32_t>& v) { putRaw(n, 'i', {(uint64_t)v.size()}, v.data(), v.size() * 4); }
static void putB(const string& n, const vector<uint8_t>& v) { putRaw(n, 'b', {(uint64_t)v.size()}, v.data(), v.size()); }

int main(int argc, char** argv) {
  if (argc != 6) {
    cerr << "Usage: " << argv[0] << " <out.bin> <preset> <seqs.fa> <time> full|sparse:<band>:<kmerLen>:<kmerThreshold>\n";
    return 1;
  }
  gOut = fopen(argv[1], "wb");
  const RateModel model = namedModel(argv[2]);
  const vguard<FastSeq> seqs = readFastSeqs(argv[3]);
  const double time = atof(argv[4]);
  const string mode = argv[5];
  int band = 0, kmerLen = 0, kmerThreshold = 0;
  long long maxSize = 0;   // optional fourth field: the memory limit of the kmerThreshold < 0 branch (-kmatchmb), cell size 40
  const bool sparse = mode != "full";
  if (sparse && sscanf(mode.c_str(), "sparse:%d:%d:%d:%lld", &band, &kmerLen, &kmerThreshold, &maxSize) < 3) {
    cerr << "bad mode " << mode << endl;
    return 1;
  }
  const int A = model.alphabetSize();
  bool scoresDone = false;
  int k = 0;
  double fillSeconds = 0, traceSeconds = 0, cells = 0;   // time inside the reference's own constructor / alignPath()
  for (size_t a = 0; a < seqs.size(); ++a)
    for (size_t b = a + 1; b < seqs.size(); ++b, ++k) {
      DiagonalEnvelope env(seqs[a], seqs[b]);
      if (sparse) {
        KmerIndex yKmerIndex(seqs[b], model.alphabet, kmerLen);
        env.initSparse(yKmerIndex, band, kmerThreshold, maxSize > 0 ? 40 : 3 * sizeof(double), (size_t)maxSize);
      } else
        env.initFull();
      const auto t0 = chrono::steady_clock::now();
      QuickAlignMatrix mx(env, model, time);
      fillSeconds += chrono::duration<double>(chrono::steady_clock::now() - t0).count();
      for (auto d : env.diagonals) {
        const long lo = max(1L, 1L - d), hi = min((long)env.yLen, (long)env.xLen - d);
        if (hi >= lo) cells += (double)(hi - lo + 1);
      }
      if (!scoresDone) {
        vector<double> sub;
        for (int i = 0; i < A; ++i)
          for (int j = 0; j < A; ++j) sub.push_back(mx.submat[i][j]);
        putD("submat", sub);
        putD("scores", {mx.noGap, mx.gapOpen, mx.gapExtend, mx.m2m, mx.m2i, mx.m2d, mx.i2i, mx.i2m, mx.i2d, mx.d2d, mx.d2m});
        putRaw("alphabet", 's', {(uint64_t)model.alphabet.size()}, model.alphabet.data(), model.alphabet.size());
        scoresDone = true;
      }
      const string p = "p" + to_string(k) + ".";
      putI(p + "x_tok", vector<int32_t>(mx.xTok.begin(), mx.xTok.end()));
      putI(p + "y_tok", vector<int32_t>(mx.yTok.begin(), mx.yTok.end()));
      putI(p + "diagonals", vector<int32_t>(env.diagonals.begin(), env.diagonals.end()));
      putD(p + "result", {mx.result});
      putI(p + "end", {(int32_t)mx.xEnd, (int32_t)mx.yEnd});
      if (mx.resultIsFinite()) {
        const auto t1 = chrono::steady_clock::now();
        const AlignPath path = mx.alignPath();
        traceSeconds += chrono::duration<double>(chrono::steady_clock::now() - t1).count();
        const vector<bool>& r0 = path.at(0);
        const vector<bool>& r1 = path.at(1);
        vector<uint8_t> ops(r0.size());
        for (size_t c = 0; c < r0.size(); ++c) ops[c] = (r0[c] ? 1 : 0) | (r1[c] ? 2 : 0);   // 3: both, 1: x only, 2: y only
        putB(p + "path", ops);
      }
    }
  putI("n_pairs", {k});
  putD("timing", {fillSeconds, traceSeconds, cells});
  printf("pairs %d cells %.0f fill_s %.6f trace_s %.6f\n", k, cells, fillSeconds, traceSeconds);
  fclose(gOut);
  return 0;
}
