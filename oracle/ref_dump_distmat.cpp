// ORACLE INFRASTRUCTURE: golden-vector generator for the distance-matrix stage:
// RateModel::distanceMatrix / mlDistance / DistanceMatrixParams::tML (model.cpp:506-655) on a gapped FASTA alignment,
// and the neighbour-joining and UPGMA trees the reference builds from it (tree.cpp:240-461).
// Compiled against the UNMODIFIED reference (oracle/Makefile).
//
//   ref_dump_distmat <out.bin> <preset[:bins:shape]> <gapped.fa> <maxIterations>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <iostream>
#include <string>

#include "fastseq.h"
#include "gamma.h"
#include "logger.h"
#include "model.h"
#include "presets.h"
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
static void putS(const string& n, const string& s) { putRaw(n, 's', {(uint64_t)s.size()}, s.data(), s.size()); }

int main(int argc, char** argv) {
  if (argc != 5) {
    cerr << "Usage: " << argv[0] << " <out.bin> <preset[:bins:shape]> <gapped.fa> <maxIterations>\n";
    return 1;
  }
  gOut = fopen(argv[1], "wb");
  const string spec = argv[2];
  vector<string> f;
  for (size_t p = 0;;) {
    const size_t q = spec.find(':', p);
    f.push_back(spec.substr(p, q == string::npos ? string::npos : q - p));
    if (q == string::npos) break;
    p = q + 1;
  }
  RateModel model = namedModel(f[0]);
  if (f.size() == 3) model = makeDiscretizedGammaModel(model, atoi(f[1].c_str()), atof(f[2].c_str()));
  const vguard<FastSeq> gapped = readFastSeqs(argv[3]);
  const int maxIter = atoi(argv[4]);
  const size_t n = gapped.size();
  const auto t0 = chrono::steady_clock::now();
  const auto dist = model.distanceMatrix(gapped, maxIter);
  const double distSeconds = chrono::duration<double>(chrono::steady_clock::now() - t0).count();
  vector<double> flat;
  for (const auto& row : dist) flat.insert(flat.end(), row.begin(), row.end());
  putRaw("dist", 'd', {(uint64_t)n, (uint64_t)n}, flat.data(), flat.size() * 8);
  const double rate = model.expectedSubstitutionRate();
  putRaw("expected_sub_rate", 'd', {1}, &rate, 8);
  string rows;
  for (const auto& s : gapped) rows += s.seq;
  putRaw("rows", 'b', {(uint64_t)n, (uint64_t)gapped[0].seq.size()}, rows.data(), rows.size());
  string names;
  for (const auto& s : gapped) names += s.name + "\n";
  putS("names", names);
  putS("alphabet", model.alphabet);
  {
    Tree nj, up;
    const auto t1 = chrono::steady_clock::now();
    nj.buildByNeighborJoining(gapped, dist);
    const auto t2 = chrono::steady_clock::now();
    up.buildByUPGMA(gapped, dist);
    const auto t3 = chrono::steady_clock::now();
    const double timing[3] = {distSeconds, chrono::duration<double>(t2 - t1).count(), chrono::duration<double>(t3 - t2).count()};
    putRaw("timing", 'd', {3}, timing, 24);
    printf("rows %zu dist_s %.6f nj_s %.6f upgma_s %.6f\n", n, timing[0], timing[1], timing[2]);
    putS("nj", nj.toString());
    putS("upgma", up.toString());
  }
  fclose(gOut);
  return 0;
}
