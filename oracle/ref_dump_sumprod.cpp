// ORACLE INFRASTRUCTURE: golden-vector generator for the sum-product / EM-count path.
// Compiled against the UNMODIFIED reference (oracle/Makefile), it runs the reference's own
// AlignColSumProduct (sumprod.cpp) column loop exactly as EigenCounts::accumulateSubstitutionCounts
// does (model.cpp:900-915) and dumps the model's eigen-system, the per-branch tables, every column's
// E/F/G messages with their log scale factors, column log-likelihoods, node posteriors, MAP states and
// the accumulated root / eigen / substitution counts.  tools/make_golden.py turns the dump into a
// fixture under tests/golden/.
//
//   ref_dump_sumprod <out.bin> <model.json | preset:NAME[:bins:shape]> <gapped.fa> <tree.nh>
#include <cstdint>
#include <cstdio>
#include <fstream>
#include <iostream>

#include <gsl/gsl_matrix.h>
#include <gsl/gsl_matrix_complex_double.h>
#include <gsl/gsl_vector.h>
#include <gsl/gsl_vector_complex_double.h>
#include <string>

#include "alignpath.h"
#include "fastseq.h"
#include "jsonutil.h"
#include "logsumexp.h"
#include "tree.h"
// The E/F/G message tables of SumProduct and the eigen-system of EigenModel are private members.
// Every header those two include is already included above, so only model.h / sumprod.h are affected.
#define private public
#include "model.h"
#include "sumprod.h"
#undef private
#include "gamma.h"
#include "jsonutil.h"
#include "logger.h"
#include "presets.h"
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
static void putI(const string& n, const vector<int32_t>& v) { putRaw(n, 'i', {(uint64_t)v.size()}, v.data(), v.size() * 4); }
static void putB(const string& n, const vector<uint8_t>& v, vector<uint64_t> shape) { putRaw(n, 'b', shape, v.data(), v.size()); }
static void putS(const string& n, const string& s) { putRaw(n, 's', {(uint64_t)s.size()}, s.data(), s.size()); }

int main(int argc, char** argv) {
  if (argc != 5) {
    cerr << "Usage: " << argv[0] << " <out.bin> <model.json|preset:NAME[:bins:shape]> <gapped.fa> <tree.nh>\n";
    return 1;
  }
  gOut = fopen(argv[1], "wb");
  RateModel rates;
  const string spec = argv[2];
  if (spec.compare(0, 7, "preset:") == 0) {
    vector<string> f;
    size_t p = 7;
    while (true) {
      const size_t q = spec.find(':', p);
      f.push_back(spec.substr(p, q == string::npos ? string::npos : q - p));
      if (q == string::npos) break;
      p = q + 1;
    }
    rates = namedModel(f[0]);
    if (f.size() == 3) rates = makeDiscretizedGammaModel(rates, atoi(f[1].c_str()), atof(f[2].c_str()));
  } else {
    ifstream in(argv[2]);
    ParsedJson pj(in);
    rates.read(pj.value);
  }
  vguard<FastSeq> gapped = readFastSeqs(argv[3]);
  ifstream treeStream(argv[4]);
  Tree tree(JsonUtil::readStringFromStream(treeStream));
  tree.reorderSeqs(gapped);

  const int C = rates.components(), A = rates.alphabetSize(), N = tree.nodes();
  const uint64_t uC = C, uA = A, uN = N;
  putS("alphabet", rates.alphabet);
  putI("dims", {C, A, N});
  vector<double> cw(C), ins(C * A), rate(C * A * A);
  for (int c = 0; c < C; ++c) {
    cw[c] = rates.cptWeight[c];
    for (int i = 0; i < A; ++i) {
      ins[c * A + i] = gsl_vector_get(rates.insProb[c], i);
      for (int j = 0; j < A; ++j) rate[(c * A + i) * A + j] = gsl_matrix_get(rates.subRate[c], i, j);
    }
  }
  putD("cpt_weight", cw, {uC});
  putD("ins_prob", ins, {uC, uA});
  putD("sub_rate", rate, {uC, uA, uA});
  putD("indel", {rates.insRate, rates.delRate, rates.insExtProb, rates.delExtProb}, {4});

  vector<int32_t> parent(N);
  vector<double> blen(N);
  for (int r = 0; r < N; ++r) {
    parent[r] = tree.parentNode(r);
    blen[r] = r == tree.root() ? 0 : tree.branchLength(r);
  }
  putI("parent", parent);
  putD("branch_length", blen, {uN});

  AlignColSumProduct sp(rates, tree, gapped);
  {
    vector<double> er(C * A), ei(C * A), vr(C * A * A), vi(C * A * A), wr(C * A * A), wi(C * A * A);
    for (int c = 0; c < C; ++c)
      for (int i = 0; i < A; ++i) {
        er[c * A + i] = GSL_REAL(sp.eigen.ev[c][i]);
        ei[c * A + i] = GSL_IMAG(sp.eigen.ev[c][i]);
        for (int j = 0; j < A; ++j) {
          const gsl_complex v = gsl_matrix_complex_get(sp.eigen.evec[c], i, j), w = gsl_matrix_complex_get(sp.eigen.evecInv[c], i, j);
          vr[(c * A + i) * A + j] = GSL_REAL(v); vi[(c * A + i) * A + j] = GSL_IMAG(v);
          wr[(c * A + i) * A + j] = GSL_REAL(w); wi[(c * A + i) * A + j] = GSL_IMAG(w);
        }
      }
    putD("eval_re", er, {uC, uA}); putD("eval_im", ei, {uC, uA});
    putD("evec_re", vr, {uC, uA, uA}); putD("evec_im", vi, {uC, uA, uA});
    putD("evec_inv_re", wr, {uC, uA, uA}); putD("evec_inv_im", wi, {uC, uA, uA});
  }
  {
    vector<double> bs((size_t)C * N * A * A, 0), br((size_t)C * N * A * A, 0), bi((size_t)C * N * A * A, 0);
    for (int c = 0; c < C; ++c)
      for (int r = 0; r < N - 1; ++r)
        for (int i = 0; i < A; ++i)
          for (int j = 0; j < A; ++j) {
            const size_t k = (((size_t)c * N + r) * A + i) * A + j;
            bs[k] = sp.branchSubProb[c][r][i][j];
            const gsl_complex e = gsl_matrix_complex_get(sp.branchEigenSubCount[c][r], i, j);
            br[k] = GSL_REAL(e); bi[k] = GSL_IMAG(e);
          }
    putD("branch_sub", bs, {uC, uN, uA, uA});
    putD("branch_esc_re", br, {uC, uN, uA, uA});
    putD("branch_esc_im", bi, {uC, uN, uA, uA});
  }

  const size_t ncol = gapped.front().length();
  const uint64_t uL = ncol;
  vector<uint8_t> cols(ncol * N), mapst(ncol * N);
  vector<double> E(ncol * C * N * A), F(E.size()), G(E.size()), lE(ncol * C * N), lF(lE.size()), lG(lE.size()), cll(ncol),
      npost(ncol * N * A, -numeric_limits<double>::infinity());
  vguard<vguard<vguard<gsl_complex> > > eigenCount(C, vguard<vguard<gsl_complex> >(A, vguard<gsl_complex>(A, gsl_complex_rect(0, 0))));
  vguard<vguard<vguard<double> > > subCount(C, vguard<vguard<double> >(A, vguard<double>(A, 0)));
  vguard<vguard<double> > root(C, vguard<double>(A, 0)), root2(C, vguard<double>(A, 0));
  vector<double> cumRe(ncol * C * A * A), cumIm(cumRe.size()), cumRoot(ncol * C * A);
  double lpTotal = 0;
  while (!sp.alignmentDone()) {
    const size_t col = sp.col;
    sp.fillUp();
    sp.fillDown();
    const bool single = sp.roots.size() == 1;
    if (single) {
      sp.accumulateEigenCounts(root, eigenCount);
      sp.accumulateSubCounts(root2, subCount);
    }
    lpTotal += sp.columnLogLikelihood();
    cll[col] = sp.columnLogLikelihood();
    for (int r = 0; r < N; ++r) {
      cols[col * N + r] = (uint8_t)gapped[r].seq[col];
      const char g = gapped[r].seq[col];
      mapst[col * N + r] = (uint8_t)((single && Alignment::isWildcard(g)) ? rates.alphabet[sp.maxPostState(r)] : g);
      if (single && !sp.isGap(r)) {
        const auto lpp = sp.logNodePostProb(r);
        for (int i = 0; i < A; ++i) npost[(col * N + r) * A + i] = lpp[i];
      }
    }
    for (int c = 0; c < C; ++c) {
      for (int r = 0; r < N; ++r) {
        lE[(col * C + c) * N + r] = sp.logE[c][r];
        lF[(col * C + c) * N + r] = sp.logF[c][r];
        lG[(col * C + c) * N + r] = sp.logG[c][r];
        for (int i = 0; i < A; ++i) {
          const size_t k = (((size_t)col * C + c) * N + r) * A + i;
          E[k] = sp.E[c][r][i]; F[k] = sp.F[c][r][i]; G[k] = sp.G[c][r][i];
        }
      }
      for (int i = 0; i < A; ++i) {
        cumRoot[(col * C + c) * A + i] = root[c][i];
        for (int j = 0; j < A; ++j) {
          cumRe[((col * C + c) * A + i) * A + j] = GSL_REAL(eigenCount[c][i][j]);
          cumIm[((col * C + c) * A + i) * A + j] = GSL_IMAG(eigenCount[c][i][j]);
        }
      }
    }
    sp.nextColumn();
  }
  putB("columns", cols, {uL, uN});
  putB("map_state", mapst, {uL, uN});
  putD("E", E, {uL, uC, uN, uA}); putD("F", F, {uL, uC, uN, uA}); putD("G", G, {uL, uC, uN, uA});
  putD("logE", lE, {uL, uC, uN}); putD("logF", lF, {uL, uC, uN}); putD("logG", lG, {uL, uC, uN});
  putD("col_loglike", cll, {uL});
  putD("node_post", npost, {uL, uN, uA});
  putD("cum_root_count", cumRoot, {uL, uC, uA});
  putD("cum_eigen_re", cumRe, {uL, uC, uA, uA});
  putD("cum_eigen_im", cumIm, {uL, uC, uA, uA});
  putD("lp_total", {lpTotal}, {1});
  const auto sc = sp.eigen.getSubCounts(eigenCount);
  vector<double> s1(C * A * A), s2(C * A * A);
  for (int c = 0; c < C; ++c)
    for (int i = 0; i < A; ++i)
      for (int j = 0; j < A; ++j) {
        s1[(c * A + i) * A + j] = sc[c][i][j];
        s2[(c * A + i) * A + j] = subCount[c][i][j];
      }
  putD("sub_count_eigen", s1, {uC, uA, uA});
  putD("sub_count_direct", s2, {uC, uA, uA});
  fclose(gOut);
  return 0;
}
