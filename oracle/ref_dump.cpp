// ORACLE INFRASTRUCTURE: golden-vector generator.  Compiled against the UNMODIFIED
// reference (headers and objects where they lie, see oracle/Makefile target
// `refdump`), it replays Reconstructor::reconstruct (recon.cpp:917-1052) with the
// reference's own classes and dumps, for every internal node asked for, the
// flattened inputs of the hot path (profiles, PairHMM, envelope) together with
// the reference's outputs (constructor tables, Forward/Backward cells, traces,
// posterior cells, the parent profile).  tools/make_golden.py turns the dumps
// into the fixtures under tests/golden/.
//
//   ref_dump <out.bin> <node[,node...]|all> <mode> [historian recon options...]
//     mode: sample | best | post   (profile construction used at non-root nodes)
#include <cstdint>
#include <cstdio>
#include <deque>
#include <fstream>
#include <iostream>
#include <sstream>

#include "recon.h"
#include "logger.h"
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
static void putD(const string& n, const vector<double>& v, vector<uint64_t> shape = {}) {
  if (shape.empty()) shape = {(uint64_t)v.size()};
  putRaw(n, 'd', shape, v.data(), v.size() * 8);
}
static void putI(const string& n, const vector<int32_t>& v, vector<uint64_t> shape = {}) {
  if (shape.empty()) shape = {(uint64_t)v.size()};
  putRaw(n, 'i', shape, v.data(), v.size() * 4);
}
static void putB(const string& n, const vector<uint8_t>& v, vector<uint64_t> shape = {}) {
  if (shape.empty()) shape = {(uint64_t)v.size()};
  putRaw(n, 'b', shape, v.data(), v.size());
}
static void putU(const string& n, const vector<uint32_t>& v) { putRaw(n, 'u', {(uint64_t)v.size()}, v.data(), v.size() * 4); }
static void putS(const string& n, const string& s) { putRaw(n, 's', {(uint64_t)s.size()}, s.data(), s.size()); }

static void dumpProfile(const string& pfx, const Profile& p, int closestRow, bool haveEnv) {
  const int n = (int)p.size(), C = (int)p.components, A = (int)p.alphSize;
  vector<uint8_t> emit(n);
  vector<double> lpAbs((size_t)n * C * A, -numeric_limits<double>::infinity());
  vector<int32_t> leafPos(n, 0);
  for (int i = 0; i < n; ++i) {
    emit[i] = !p.state[i].isNull();
    if (emit[i])
      for (int c = 0; c < C; ++c)
        for (int a = 0; a < A; ++a) lpAbs[((size_t)i * C + c) * A + a] = p.state[i].lpAbsorb[c][a];
    if (haveEnv && i > 0) leafPos[i] = (int32_t)p.state[i].seqCoords.at(closestRow);
  }
  vector<int32_t> src, dest;
  vector<double> lp;
  for (const auto& t : p.trans) {
    src.push_back((int32_t)t.src);
    dest.push_back((int32_t)t.dest);
    lp.push_back(t.lpTrans);
  }
  // per-state list orders, to verify "ascending transition index" (see historian_b200.h)
  vector<int32_t> inFlat, inOff(1, 0), nullFlat, nullOff(1, 0), absFlat, absOff(1, 0);
  for (int i = 0; i < n; ++i) {
    for (auto t : p.state[i].in) inFlat.push_back((int32_t)t);
    inOff.push_back((int32_t)inFlat.size());
    for (auto t : p.state[i].nullOut) nullFlat.push_back((int32_t)t);
    nullOff.push_back((int32_t)nullFlat.size());
    for (auto t : p.state[i].absorbOut) absFlat.push_back((int32_t)t);
    absOff.push_back((int32_t)absFlat.size());
  }
  putI(pfx + "dims", {n, (int32_t)p.trans.size(), A, C});
  putB(pfx + "is_emit", emit);
  putI(pfx + "trans_src", src);
  putI(pfx + "trans_dest", dest);
  putD(pfx + "trans_lp", lp);
  putD(pfx + "lp_absorb", lpAbs, {(uint64_t)n, (uint64_t)C, (uint64_t)A});
  putI(pfx + "closest_leaf_pos", leafPos);
  putI(pfx + "in_flat", inFlat);
  putI(pfx + "in_off", inOff);
  putI(pfx + "null_flat", nullFlat);
  putI(pfx + "null_off", nullOff);
  putI(pfx + "abs_flat", absFlat);
  putI(pfx + "abs_off", absOff);
  putS(pfx + "name", p.name);
}

static void dumpHmm(const string& pfx, const PairHMM& h) {
  const int C = h.components(), A = (int)h.alphabetSize();
  vector<double> w = {h.imm_imm, h.imm_imd, h.imm_idm, h.imm_imi, h.imm_iiw, h.imm_eee, h.imd_imm, h.imd_imd,
                      h.imd_idm, h.imd_eee, h.idm_imm, h.idm_imd, h.idm_idm, h.idm_eee, h.imi_imm, h.imi_imd,
                      h.imi_imi, h.imi_iiw, h.imi_eee, h.iiw_imm, h.iiw_idm, h.iiw_iiw, h.iiw_eee};
  putD(pfx + "lp_trans", w);
  vector<double> root, lsub, rsub, lins, rins;
  for (int c = 0; c < C; ++c)
    for (int a = 0; a < A; ++a) {
      root.push_back(h.logRoot[c][a]);
      lins.push_back(h.logl.logInsProb[c][a]);
      rins.push_back(h.logr.logInsProb[c][a]);
      for (int d = 0; d < A; ++d) {
        lsub.push_back(gsl_matrix_get(h.l.subMat[c], a, d));
        rsub.push_back(gsl_matrix_get(h.r.subMat[c], a, d));
      }
    }
  putD(pfx + "log_root", root, {(uint64_t)C, (uint64_t)A});
  putD(pfx + "l_sub", lsub, {(uint64_t)C, (uint64_t)A, (uint64_t)A});
  putD(pfx + "r_sub", rsub, {(uint64_t)C, (uint64_t)A, (uint64_t)A});
  putD(pfx + "l_log_ins", lins, {(uint64_t)C, (uint64_t)A});
  putD(pfx + "r_log_ins", rins, {(uint64_t)C, (uint64_t)A});
  putD(pfx + "l_log_cpt_weight", h.logl.logCptWeight);
  putD(pfx + "r_log_cpt_weight", h.logr.logCptWeight);
  putD(pfx + "branch", {h.l.t, h.r.t, h.l.insWait, h.l.delWait, h.r.insWait, h.r.delWait});
}

static void dumpEnv(const string& pfx, const GuideAlignmentEnvelope& e) {
  putI(pfx + "max_distance", {(int32_t)e.maxDistance});
  vector<int32_t> cm(e.cumulativeMatches.begin(), e.cumulativeMatches.end());
  vector<int32_t> r1(e.row1PosToCol.begin(), e.row1PosToCol.end()), r2(e.row2PosToCol.begin(), e.row2PosToCol.end());
  putI(pfx + "cumulative_matches", cm);
  putI(pfx + "row1_pos_to_col", r1);
  putI(pfx + "row2_pos_to_col", r2);
}

struct FwdPeek : ForwardMatrix {
  FwdPeek(const Profile& x, const Profile& y, const PairHMM& hmm, AlignRowIndex row, const GuideAlignmentEnvelope& env, SumProduct* sp = NULL)
      : ForwardMatrix(x, y, hmm, row, env, sp) {}
  void dumpTables(const string& pfx) {
    const int C = components(), A = (int)alphSize;
    vector<double> sx((size_t)xSize * C * A, -numeric_limits<double>::infinity()), sy((size_t)ySize * C * A, -numeric_limits<double>::infinity());
    for (size_t i = 0; i < xSize; ++i)
      if (!subx.state[i].isNull())
        for (int c = 0; c < C; ++c)
          for (int a = 0; a < A; ++a) sx[(i * C + c) * A + a] = subx.state[i].lpAbsorb[c][a];
    for (size_t j = 0; j < ySize; ++j)
      if (!suby.state[j].isNull())
        for (int c = 0; c < C; ++c)
          for (int a = 0; a < A; ++a) sy[(j * C + c) * A + a] = suby.state[j].lpAbsorb[c][a];
    putD(pfx + "subx", sx, {(uint64_t)xSize, (uint64_t)C, (uint64_t)A});
    putD(pfx + "suby", sy, {(uint64_t)ySize, (uint64_t)C, (uint64_t)A});
    putD(pfx + "insx", vector<double>(insx.begin(), insx.end()));
    putD(pfx + "insy", vector<double>(insy.begin(), insy.end()));
    putD(pfx + "rootsubx", vector<double>(rootsubx.begin(), rootsubx.end()));
    putD(pfx + "rootsuby", vector<double>(rootsuby.begin(), rootsuby.end()));
    vector<uint8_t> ns(xSize), ne(ySize);
    for (size_t i = 0; i < xSize; ++i) ns[i] = xNearStart[i];
    for (size_t j = 0; j < ySize; ++j) ne[j] = yNearEnd[j];
    putB(pfx + "x_near_start", ns);
    putB(pfx + "y_near_end", ne);
    if (!getenv("REF_DUMP_SPARSE")) {
      vector<uint8_t> env((xSize - 1) * (ySize - 1));
      for (size_t i = 0; i + 1 < xSize; ++i)
        for (size_t j = 0; j + 1 < ySize; ++j) env[i * (ySize - 1) + j] = inEnvelope(i, j);
      putB(pfx + "in_envelope", env, {(uint64_t)xSize - 1, (uint64_t)ySize - 1});
    }
  }
};

template <class M>
static void dumpCells(const string& name, const M& m) {
  vector<double> c((m.xSize - 1) * (m.ySize - 1) * 5);
  for (size_t i = 0; i + 1 < m.xSize; ++i)
    for (size_t j = 0; j + 1 < m.ySize; ++j)
      for (int s = 0; s < 5; ++s)
        c[(i * (m.ySize - 1) + j) * 5 + s] = m.cell(i, j, (PairHMM::State)s);
  putD(name, c, {(uint64_t)m.xSize - 1, (uint64_t)m.ySize - 1, 5});
}

// REF_DUMP_SPARSE=<stride>: large matrices.  Instead of the dense cell array, every stride-th in-envelope cell in
// row-major order (coordinates + five values), plus the per-row in-envelope counts.
template <class M>
static void dumpCellsSparse(const string& name, const M& m, int stride) {
  vector<int32_t> xy, rowCount(m.xSize - 1, 0);
  vector<double> v;
  size_t k = 0;
  for (size_t i = 0; i + 1 < m.xSize; ++i)
    for (size_t j = 0; j + 1 < m.ySize; ++j)
      if (m.inEnvelope(i, j)) {
        ++rowCount[i];
        if (k++ % stride == 0) {
          xy.push_back((int32_t)i);
          xy.push_back((int32_t)j);
          for (int s = 0; s < 5; ++s) v.push_back(m.cell(i, j, (PairHMM::State)s));
        }
      }
  putI(name + "_xy", xy, {(uint64_t)xy.size() / 2, 2});
  putD(name + "_vals", v, {(uint64_t)xy.size() / 2, 5});
  putI(name + "_row_count", rowCount);
}

static void dumpPath(const string& name, const ForwardMatrix::Path& p) {
  vector<int32_t> v;
  for (const auto& c : p) {
    v.push_back((int32_t)c.xpos);
    v.push_back((int32_t)c.ypos);
    v.push_back((int32_t)c.state);
  }
  putI(name, v, {(uint64_t)p.size(), 3});
}

static vector<uint32_t> rngState(const ForwardMatrix::random_engine& g) {
  stringstream ss;
  ss << g;
  vector<uint32_t> v(625);
  for (auto& x : v) {
    unsigned long t;
    ss >> t;
    x = (uint32_t)t;
  }
  return v;
}

int main(int argc, char** argv) {
  if (argc < 4) {
    cerr << "Usage: " << argv[0] << " <out.bin> <node|all> <sample|best|post> [recon options]\n";
    return 1;
  }
  gOut = fopen(argv[1], "wb");
  const string which(argv[2]), mode(argv[3]);
  deque<string> argvec;
  for (int i = 4; i < argc; ++i) argvec.push_back(argv[i]);

  Reconstructor recon;
  recon.reconstructRoot = true;
  while (argvec.size()) {
    if (logger.parseLogArgs(argvec) || recon.parseReconArgs(argvec) || recon.parseModelArgs(argvec) ||
        recon.parseProfileArgs(argvec, false) || recon.parseSamplerArgs(argvec) || recon.parseAncSeqArgs(argvec))
      continue;
    cerr << "unknown option " << argvec.front() << endl;
    return 1;
  }
  recon.loadModel();
  recon.loadSeqs();
  Reconstructor::Dataset& dataset = recon.datasets.front();
  RateModel& model = recon.model;

  recon.generator = ForwardMatrix::newRNG();
  recon.generator.seed(recon.rndSeed);

  vguard<gsl_vector*> rootProb = model.insProb;
  // REF_DUMP_COUNTS=2: the profiling strategy of `historian fit` on unaligned sequences (recon.cpp:925-930 with
  // accumulateSubstCounts and accumulateIndelCounts): every profile transition carries the EigenCounts of the paths it sums
  // over (forward.cpp:753-790), and getCounts at the root reads them (forward.cpp:1183-1214).
  const bool propagate = getenv("REF_DUMP_COUNTS") != NULL && string(getenv("REF_DUMP_COUNTS")) == "2";
  const int strategy = ForwardMatrix::CollapseChains | (recon.keepGapsOpen ? ForwardMatrix::KeepGapsOpen : 0) |
                       (recon.includeBestTraceInProfile ? ForwardMatrix::IncludeBestTrace : 0) |
                       (propagate ? (ForwardMatrix::CountSubstEvents | ForwardMatrix::CountIndelEvents) : 0);
  putI("strategy", {strategy});
  putS("alphabet", model.alphabet);
  putI("profile_samples", {(int32_t)recon.profileSamples});
  putI("max_profile_states", {(int32_t)recon.maxProfileStates()});
  putD("min_post_prob", {recon.minPostProb});
  putI("root_node", {(int32_t)dataset.tree.root()});

  // REF_DUMP_COUNTS=1: also run BackwardMatrix::getCounts (forward.cpp:1183-1214) at the dumped nodes, with the
  // SumProduct attached as recon.cpp:932-934 does.  The profiling strategy above carries no Count* flags, so
  // profile transitions hold no counts of their own.
  const bool wantCounts = getenv("REF_DUMP_COUNTS") != NULL;
  SumProduct* sumProd = wantCounts ? new SumProduct(model, dataset.tree) : NULL;
  if (wantCounts) {
    vector<int32_t> parent;
    vector<double> blen;
    for (TreeNodeIndex n = 0; n < dataset.tree.nodes(); ++n) {
      parent.push_back((int32_t)dataset.tree.parentNode(n));
      blen.push_back(n == dataset.tree.root() ? 0. : dataset.tree.branchLength(n));
    }
    putI("tree.parent", parent);
    putD("tree.branch_length", blen);
  }

  vector<int32_t> dumped;
  map<int, Profile> prof;
  for (TreeNodeIndex node = 0; node < dataset.tree.nodes(); ++node) {
    if (dataset.tree.isLeaf(node)) {
      prof[node] = Profile(model.components(), model.alphabet, dataset.seqs[dataset.nodeToSeqIndex[node]], node);
      continue;
    }
    const int l = dataset.tree.getChild(node, 0), r = dataset.tree.getChild(node, 1);
    const Profile& lProf = prof[l];
    const Profile& rProf = prof[r];
    ProbModel lProbs(model, dataset.tree.branchLength(l));
    ProbModel rProbs(model, dataset.tree.branchLength(r));
    PairHMM hmm(lProbs, rProbs, rootProb);
    const bool haveGuide = !dataset.guide.empty();
    int maxDist = recon.maxDistanceFromGuide;
    FwdPeek* forward = NULL;
    GuideAlignmentEnvelope env;
    while (true) {  // recon.cpp:954-975
      env = haveGuide ? GuideAlignmentEnvelope(dataset.guide, dataset.closestLeaf[l], dataset.closestLeaf[r], maxDist)
                      : GuideAlignmentEnvelope();
      forward = new FwdPeek(lProf, rProf, hmm, node, env, sumProd);
      if (forward->lpEnd > -numeric_limits<double>::infinity()) break;
      if (maxDist < 0) { cerr << "zero likelihood" << endl; return 1; }
      if (maxDist * 2 > (int)alignPathColumns(dataset.guide)) maxDist = -1;
      else if (maxDist == 0) maxDist = 1;
      else maxDist *= 2;
      delete forward;
    }
    const bool isRoot = node == dataset.tree.root();
    bool want = which == "all";
    if (which == "auto3") {  // the root and those of its children that are internal nodes
      const TreeNodeIndex rt = dataset.tree.root();
      want = isRoot || ((node == dataset.tree.getChild(rt, 0) || node == dataset.tree.getChild(rt, 1)) && !dataset.tree.isLeaf(node));
    } else if (!want) {  // comma-separated node list
      stringstream ws(which);
      string tok;
      while (getline(ws, tok, ',')) want = want || atoi(tok.c_str()) == node;
    }
    const string pfx = "n" + to_string(node) + ".";
    BackwardMatrix* backward = NULL;
    if (want || (mode == "post" && !isRoot)) backward = new BackwardMatrix(*forward);
    if (want) {
      dumped.push_back(node);
      putI(pfx + "children", {l, r});
      putI(pfx + "closest_rows", {haveGuide ? (int32_t)dataset.closestLeaf[l] : -1, haveGuide ? (int32_t)dataset.closestLeaf[r] : -1});
      dumpProfile(pfx + "x.", lProf, haveGuide ? dataset.closestLeaf[l] : 0, env.initialized());
      dumpProfile(pfx + "y.", rProf, haveGuide ? dataset.closestLeaf[r] : 0, env.initialized());
      dumpHmm(pfx + "hmm.", hmm);
      dumpEnv(pfx + "env.", env);
      forward->dumpTables(pfx);
      const int sparse = getenv("REF_DUMP_SPARSE") ? max(1, atoi(getenv("REF_DUMP_SPARSE"))) : 0;
      if (sparse) dumpCellsSparse(pfx + "fwd", *forward, sparse); else dumpCells(pfx + "fwd", *forward);
      putD(pfx + "lp_end", {forward->lpEnd});
      if (sparse) dumpCellsSparse(pfx + "bwd", *backward, sparse); else dumpCells(pfx + "bwd", *backward);
      putD(pfx + "lp_start", {backward->lpStart()});
      dumpPath(pfx + "best_trace", forward->bestTrace());
      // posterior cells in pop order
      {
        auto pq = backward->cellsAbovePostProbThreshold(recon.minPostProb > 0 ? recon.minPostProb : .01);
        vector<int32_t> pc;
        vector<double> pl;
        while (!pq.empty()) {
          const auto& t = pq.top();
          pc.push_back((int32_t)t.xpos); pc.push_back((int32_t)t.ypos); pc.push_back((int32_t)t.state);
          pl.push_back(t.logPostProb);
          pq.pop();
        }
        putI(pfx + "post_cells", pc, {(uint64_t)pl.size(), 3});
        putD(pfx + "post_lp", pl);
        putD(pfx + "post_threshold", {recon.minPostProb > 0 ? recon.minPostProb : .01});
      }
      if (wantCounts) {
        const int N = (int)dataset.tree.nodes();
        for (int side = 0; side < 2; ++side) {
          const Profile& p = side ? rProf : lProf;
          vector<uint8_t> cols(p.size() * (size_t)N, (uint8_t)'-');
          for (size_t s = 0; s < p.size(); ++s)
            for (const auto& rc : p.alignColumn(s)) cols[s * N + rc.first] = (uint8_t)rc.second;
          putB(pfx + (side ? "y." : "x.") + "align_col", cols, {(uint64_t)p.size(), (uint64_t)N});
          if (propagate) {   // the counts every transition of the child profile carries: indel slots, then the transformed event counts
            const int C = model.components(), A = model.alphabetSize();
            vector<double> tIndel, tRoot, tSub;
            for (const auto& t : p.trans) {
              EigenCounts ec = t.counts;
              if (ec.rootCount.empty()) ec = EigenCounts(C, A) + ec;   // a transition without substitution counts
              const IndelCounts& ic = ec.indelCounts;
              for (double v : {ic.ins, ic.del, ic.insExt, ic.delExt, ic.insTime, ic.delTime, ic.lp}) tIndel.push_back(v);
              const EventCounts ev = ec.transform(model);
              for (int c = 0; c < C; ++c)
                for (int i = 0; i < A; ++i) {
                  tRoot.push_back(ev.rootCount[c][i]);
                  for (int j = 0; j < A; ++j) tSub.push_back(ev.subCount[c][i][j]);
                }
            }
            putD(pfx + (side ? "y." : "x.") + "trans_indel", tIndel, {(uint64_t)p.trans.size(), 7});
            putD(pfx + (side ? "y." : "x.") + "trans_root_count", tRoot, {(uint64_t)p.trans.size(), (uint64_t)C, (uint64_t)A});
            putD(pfx + (side ? "y." : "x.") + "trans_sub_count", tSub, {(uint64_t)p.trans.size(), (uint64_t)C, (uint64_t)A, (uint64_t)A});
          }
        }
        const EigenCounts ec = backward->getCounts();
        const EventCounts ev = ec.transform(model);
        const IndelCounts& ic = ev.indelCounts;
        putD(pfx + "fbc.indel", {ic.ins, ic.del, ic.insExt, ic.delExt, ic.insTime, ic.delTime, ic.lp});
        const int C = model.components(), A = model.alphabetSize();
        vector<double> rootCount, subCount;
        for (int c = 0; c < C; ++c)
          for (int i = 0; i < A; ++i) {
            rootCount.push_back(ev.rootCount[c][i]);
            for (int j = 0; j < A; ++j) subCount.push_back(ev.subCount[c][i][j]);
          }
        putD(pfx + "fbc.root_count", rootCount, {(uint64_t)C, (uint64_t)A});
        putD(pfx + "fbc.sub_count", subCount, {(uint64_t)C, (uint64_t)A, (uint64_t)A});
        putD(pfx + "fbc.waits", {lProbs.t, rProbs.t, lProbs.insWait, lProbs.delWait, rProbs.insWait, rProbs.delWait});
      }
      // three sampled traces from a COPY of the generator (does not disturb the run)
      {
        ForwardMatrix::random_engine g = recon.generator;
        putU(pfx + "rng_before", rngState(g));
        for (int k = 0; k < 3; ++k) dumpPath(pfx + "sample_trace" + to_string(k), forward->sampleTrace(g));
        putU(pfx + "rng_after", rngState(g));
      }
    }
    Profile& nodeProf = prof[node];
    if (isRoot) nodeProf = forward->bestProfile();
    else if (mode == "post") nodeProf = backward->postProbProfile(recon.minPostProb, recon.maxProfileStates(), (ForwardMatrix::ProfilingStrategy)strategy);
    else if (mode == "best") nodeProf = forward->bestProfile((ForwardMatrix::ProfilingStrategy)strategy);
    else nodeProf = forward->sampleProfile(recon.generator, recon.profileSamples, recon.maxProfileStates(), (ForwardMatrix::ProfilingStrategy)strategy);
    if (want) {
      dumpProfile(pfx + "parent.", nodeProf, haveGuide ? dataset.closestLeaf[node] : 0, haveGuide);
      vector<double> cum;
      nodeProf.calcSumPathAbsorbProbs(log_vector(model.cptWeight), log_vector_gsl_vector(rootProb), "cumLogProb");
      for (size_t s = 0; s < nodeProf.size(); ++s)
        cum.push_back(s == 0 ? 0. : atof(nodeProf.state[s].meta.count("cumLogProb") ? nodeProf.state[s].meta.at("cumLogProb").c_str() : "0"));
      putD(pfx + "parent.cum_lp6", cum);
      if (propagate) {   // the indel counts every transition of the new profile carries (forward.cpp:764-765, 779-790, 825-826)
        vector<double> tIndel;
        for (const auto& t : nodeProf.trans) {
          const IndelCounts& ic = t.counts.indelCounts;
          for (double v : {ic.ins, ic.del, ic.insExt, ic.delExt, ic.insTime, ic.delTime, ic.lp}) tIndel.push_back(v);
        }
        putD(pfx + "parent.trans_indel", tIndel, {(uint64_t)nodeProf.trans.size(), 7});
      }
    }
    if (backward) delete backward;
    delete forward;
  }
  putI("dumped_nodes", dumped);
  fclose(gOut);
  return 0;
}
