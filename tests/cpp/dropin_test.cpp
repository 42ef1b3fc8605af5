// dropin_test.cpp — drives the C++ drop-ins the way Cbc's callers do, on top
// of the COIN test double.  Usage:
//   cbp_dropin_test <instance.bin> <out.bin> <level 1|2|3> <maxRounds> <logLevel> <timeLimit>
//                   [dive | strong <cands.bin> | probe <maxProbes> | divebatch <bounds.bin> | bench <reps>]
// Prints one "RESULT {json}" line; writes the solver's final column bounds to
// out.bin (n doubles lb, n doubles ub).  With "dive" it instead performs the
// single-round CoinBoundPropagation step of CbcHeuristicDive
// (/root/reference/src/CbcHeuristicDive.cpp:967-1001).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>

#include <thread>

#include "CbcBoundPropagation.hpp"
#include "CbcDiveBatch.hpp"
#include "CbcProbePropagation.hpp"
#include "CbcStrongPropagation.hpp"
#include "CoinBoundPropagation.hpp"
#include "OsiSolverInterface.hpp"
#include "cbp_capi.h"
#include "../../cbc_b200/host/CbpDeviceCache.hpp"

double CoinGetTimeOfDay()
{
  return std::chrono::duration<double>(std::chrono::system_clock::now().time_since_epoch()).count();
}
double CoinCpuTime() { return static_cast<double>(clock()) / CLOCKS_PER_SEC; }

struct TestMatrix : CoinPackedMatrix {
  std::vector<double> el;
  std::vector<int> idx, start, len;
  const double *getElements() const override { return el.data(); }
  const int *getIndices() const override { return idx.data(); }
  const CoinBigIndex *getVectorStarts() const override { return start.data(); }
  const int *getVectorLengths() const override { return len.data(); }
  int getNumRows() const override { return static_cast<int>(len.size()); }
  CoinBigIndex getNumElements() const override { return static_cast<int>(el.size()); }
};

struct TestSolver : OsiSolverInterface {
  TestMatrix mat;
  std::vector<char> sense, ctype;
  std::vector<double> rhs, range, lb, ub;
  double inf = 1e30, tol = 1e-7;
  int nSetLower = 0, nSetUpper = 0;

  CbpProblemView view() const
  {
    CbpProblemView v;
    v.m = mat.getNumRows();
    v.n = static_cast<int>(lb.size());
    v.rowStart32 = mat.start.data();
    v.rowLen = mat.len.data();
    v.colIdx = mat.idx.data();
    v.val = mat.el.data();
    v.rowSense = sense.data();
    v.rhs = rhs.data();
    v.range = range.data();
    v.colType = ctype.data();
    v.infinity = inf;
    return v;
  }
  int getNumCols() const override { return static_cast<int>(lb.size()); }
  int getNumRows() const override { return mat.getNumRows(); }
  // what an Osi implementation would bind: phase 1 on the device (cbp_singletons)
  int tightenBoundsFromSingletonRows(int &nFixed) override
  {
    nFixed = 0;
    cbp_matrix *m = cbpAcquireMatrix(view());
    if (!m)
      return 0;
    int nT = 0;
    static thread_local std::vector<double> l, u;
    l = lb;
    u = ub;
    if (cbp_singletons(m, l.data(), u.data(), tol, &nT, &nFixed, 0) != CBP_OK)
      return 0;
    if (nT > 0)
      for (size_t j = 0; j < lb.size(); ++j) {
        if (l[j] != lb[j])
          setColLower(static_cast<int>(j), l[j]);
        if (u[j] != ub[j])
          setColUpper(static_cast<int>(j), u[j]);
      }
    return nT;
  }
  const char *getColType(bool) const override { return ctype.data(); }
  const CoinPackedMatrix *getMatrixByRow() const override { return &mat; }
  const char *getRowSense() const override { return sense.data(); }
  const double *getRightHandSide() const override { return rhs.data(); }
  const double *getRowRange() const override { return range.data(); }
  bool getDblParam(OsiDblParam key, double &value) const override
  {
    if (key == OsiPrimalTolerance) {
      value = tol;
      return true;
    }
    return false;
  }
  double getInfinity() const override { return inf; }
  const double *getColLower() const override { return lb.data(); }
  const double *getColUpper() const override { return ub.data(); }
  void setColLower(int col, double value) override
  {
    lb[col] = value;
    ++nSetLower;
  }
  void setColUpper(int col, double value) override
  {
    ub[col] = value;
    ++nSetUpper;
  }
  std::string getRowName(int row, unsigned) const override { return "R" + std::to_string(row); }
  std::string getColName(int col, unsigned) const override { return "C" + std::to_string(col); }
};

template <class T>
static bool rd(FILE *f, std::vector<T> &v, size_t n)
{
  v.resize(n);
  return n == 0 || fread(v.data(), sizeof(T), n, f) == n;
}

static bool load(const char *path, TestSolver &s)
{
  FILE *f = fopen(path, "rb");
  if (!f)
    return false;
  int m = 0, n = 0;
  long long nnz = 0;
  bool ok = fread(&m, 4, 1, f) == 1 && fread(&n, 4, 1, f) == 1 && fread(&nnz, 8, 1, f) == 1;
  std::vector<int> start;
  ok = ok && rd(f, start, static_cast<size_t>(m) + 1) && rd(f, s.mat.idx, nnz) && rd(f, s.mat.el, nnz)
    && rd(f, s.sense, m) && rd(f, s.rhs, m) && rd(f, s.range, m) && rd(f, s.ctype, n) && rd(f, s.lb, n)
    && rd(f, s.ub, n) && fread(&s.inf, 8, 1, f) == 1;
  fclose(f);
  if (!ok)
    return false;
  s.mat.start.assign(start.begin(), start.begin() + m);
  s.mat.len.resize(m);
  for (int i = 0; i < m; ++i)
    s.mat.len[i] = start[i + 1] - start[i];
  return true;
}

int main(int argc, char **argv)
{
  if (argc < 7) {
    fprintf(stderr, "usage: %s inst.bin out.bin level maxRounds logLevel timeLimit [dive]\n", argv[0]);
    return 2;
  }
  TestSolver solver;
  if (!load(argv[1], solver)) {
    fprintf(stderr, "cannot read %s\n", argv[1]);
    return 2;
  }
  const int level = atoi(argv[3]), maxRounds = atoi(argv[4]), logLevel = atoi(argv[5]);
  const double timeLimit = atof(argv[6]);
  const bool dive = argc > 7 && !strcmp(argv[7], "dive");
  const bool strong = argc > 8 && !strcmp(argv[7], "strong");
  if (strong) {
    // what CbcNode::chooseDynamicBranch would do before its LP hot starts (src/CbcNode.cpp:3508-3626):
    // cands.bin = int32 count, int32 columns[count], double values[count]
    FILE *c = fopen(argv[8], "rb");
    int cnt = 0;
    if (!c || fread(&cnt, 4, 1, c) != 1 || cnt <= 0) {
      fprintf(stderr, "cannot read %s\n", argv[8]);
      return 2;
    }
    std::vector<int> cols(cnt);
    std::vector<double> vals(cnt);
    if (fread(cols.data(), 4, cnt, c) != static_cast<size_t>(cnt) || fread(vals.data(), 8, cnt, c) != static_cast<size_t>(cnt))
      return 2;
    fclose(c);
    CbcStrongPropagation sp;
    const bool ran = sp.run(&solver, cnt, cols.data(), vals.data(), maxRounds);
    printf("RESULT {\"strong\": 1, \"ran\": %d, \"nInfeasibleArms\": %d, \"nodeInfeasible\": %d}\n", ran ? 1 : 0,
      sp.numberInfeasibleArms(), sp.nodeInfeasible() ? 1 : 0);
    // per arm: infeasible, row, col, rounds, number of implied bounds, and the implied bounds themselves
    FILE *o = fopen(argv[2], "wb");
    for (int k = 0; ran && k < cnt; ++k)
      for (int w = 0; w < 2; ++w) {
        const CbcStrongPropagation::Way way = static_cast<CbcStrongPropagation::Way>(w);
        const CbcStrongPropagation::BoundList bl = sp.impliedBounds(k, way);
        bool up = false;
        double fb = 0.0;
        const bool forced = sp.forcedBound(k, up, fb);
        double head[8] = { sp.infeasible(k, way) ? 1.0 : 0.0, static_cast<double>(sp.infeasibleRow(k, way)),
          static_cast<double>(sp.infeasibleCol(k, way)), static_cast<double>(sp.rounds(k, way)),
          static_cast<double>(bl.size()), forced ? 1.0 : 0.0, up ? 1.0 : 0.0, fb };
        fwrite(head, 8, 8, o);
        for (const auto &e : bl) {
          double rec[3] = { static_cast<double>(e.first), e.second.first, e.second.second };
          fwrite(rec, 8, 3, o);
        }
      }
    if (o)
      fclose(o);
    cbpReleaseThreadCache();
    return 0;
  }

  if (argc > 8 && !strcmp(argv[7], "probe")) {
    // the probing sweep of fixVubs (src/CbcSolverHeuristics.cpp:265-490) as one batch, consequences applied to the solver
    CbcProbePropagation pp;
    const bool ran = pp.run(&solver, 0, nullptr, atoi(argv[8]), maxRounds);
    FILE *o = fopen(argv[2], "wb");
    int nb = static_cast<int>(pp.nodeBounds().size()), np = pp.numberProbed();
    fwrite(&np, 4, 1, o);
    for (int k = 0; k < np; ++k) {
      int rec[3] = { pp.column(k), pp.infeasible(k, CbcProbePropagation::Down) ? 1 : 0,
        pp.infeasible(k, CbcProbePropagation::Up) ? 1 : 0 };
      fwrite(rec, 4, 3, o);
    }
    fwrite(&nb, 4, 1, o);
    for (const auto &e : pp.nodeBounds()) {
      double rec[3] = { static_cast<double>(e.first), e.second.first, e.second.second };
      fwrite(rec, 8, 3, o);
    }
    const int moved = pp.apply(&solver);
    fwrite(solver.lb.data(), 8, solver.lb.size(), o);
    fwrite(solver.ub.data(), 8, solver.ub.size(), o);
    fclose(o);
    printf("RESULT {\"probe\": 1, \"ran\": %d, \"numberProbed\": %d, \"failedProbes\": %d, \"nodeInfeasible\": %d, "
           "\"nodeBounds\": %d, \"moved\": %d}\n",
      ran ? 1 : 0, np, pp.numberFailedProbes(), pp.nodeInfeasible() ? 1 : 0, nb, moved);
    cbpReleaseThreadCache();
    return 0;
  }
  if (argc > 8 && !strcmp(argv[7], "divebatch")) {
    // T dive threads as CbcDiveScheduler::runParallel starts them (src/CbcDiveScheduler.cpp:158-214); every thread walks
    // through its own sequence of bound vectors (one per dive iteration) and stops at the first infeasible one.
    // bounds.bin: int32 T, then per dive: int32 count, count x (n doubles lb, n doubles ub)
    FILE *c = fopen(argv[8], "rb");
    int T = 0;
    if (!c || fread(&T, 4, 1, c) != 1 || T <= 0)
      return 2;
    const size_t n = solver.lb.size();
    std::vector<std::vector<double> > seq(T);
    std::vector<int> count(T);
    for (int t = 0; t < T; ++t) {
      if (fread(&count[t], 4, 1, c) != 1)
        return 2;
      seq[t].resize(static_cast<size_t>(count[t]) * 2 * n);
      if (fread(seq[t].data(), 8, seq[t].size(), c) != seq[t].size())
        return 2;
    }
    fclose(c);
    CbcDiveBatch batch(&solver, T, maxRounds);
    struct Rec {
      int feasible, row, col;
      CbcDiveBatch::BoundList upd;
    };
    std::vector<std::vector<Rec> > out(T);
    std::vector<std::thread> threads;
    for (int t = 0; t < T; ++t)
      threads.emplace_back([&, t]() {
        for (int w = 0; w < count[t]; ++w) {
          Rec r;
          const double *lb = seq[t].data() + static_cast<size_t>(w) * 2 * n, *ub = lb + n;
          r.feasible = batch.propagate(t, lb, ub, r.upd, r.row, r.col) ? 1 : 0;
          out[t].push_back(r);
          if (!r.feasible)
            break;  // reasonToStop = 1 in CbcHeuristicDive::solution
        }
        batch.finish(t);
      });
    for (auto &th : threads)
      th.join();
    FILE *o = fopen(argv[2], "wb");
    long long total = 0;
    for (int t = 0; t < T; ++t) {
      const int steps = static_cast<int>(out[t].size());
      fwrite(&steps, 4, 1, o);
      for (const Rec &r : out[t]) {
        int head[4] = { r.feasible, r.row, r.col, static_cast<int>(r.upd.size()) };
        fwrite(head, 4, 4, o);
        for (const auto &e : r.upd) {
          double rec[3] = { static_cast<double>(e.first), e.second.first, e.second.second };
          fwrite(rec, 8, 3, o);
        }
        ++total;
      }
    }
    fclose(o);
    printf("RESULT {\"divebatch\": 1, \"usable\": %d, \"dives\": %d, \"waves\": %d, \"vectors\": %lld, \"steps\": %lld}\n",
      batch.usable() ? 1 : 0, T, batch.wavesRun(), batch.vectorsPropagated(), total);
    return 0;
  }
  if (argc > 8 && !strcmp(argv[7], "bench")) {
    // node-mode cost of the drop-in: `reps` calls of run() on the same solver (bounds reset before each), beside the
    // same number of bare C-ABI calls with the same host buffers; the difference is what the C++ layer adds per call
    // (finding the matrix again, copying bounds, writing changes back through setColLower / setColUpper)
    const int reps = atoi(argv[8]);
    const std::vector<double> lb0 = solver.lb, ub0 = solver.ub;
    auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    CbcBoundPropagation warm;
    warm.run(&solver, NULL, 0, static_cast<CbcBoundPropagation::Level>(level), maxRounds, true, 1e100, 0.0);
    double tRun = 0.0, tAcq = 0.0, tCapi = 0.0;
    int rounds = 0;
    for (int r = 0; r < reps; ++r) {
      solver.lb = lb0;
      solver.ub = ub0;
      const double t0 = now();
      CbcBoundPropagation bp;
      bp.run(&solver, NULL, 0, static_cast<CbcBoundPropagation::Level>(level), maxRounds, true, 1e100, 0.0);
      tRun += now() - t0;
      rounds = bp.nRoundsRun();
    }
    cbp_matrix *m = cbpAcquireMatrix(solver.view());
    for (int r = 0; r < reps; ++r) {
      const double t0 = now();
      m = cbpAcquireMatrix(solver.view());
      tAcq += now() - t0;
    }
    std::vector<double> l, u;
    double tCapiSing = 0.0, tSolverSing = 0.0;
    for (int r = 0; r < reps && m; ++r) {
      l = lb0;
      u = ub0;
      const double t0 = now();
      cbp_result res;
      int nT = 0, nF = 0;
      cbp_singletons(m, l.data(), u.data(), solver.tol, &nT, &nF, 0);
      const double t1 = now();
      cbp_propagate(m, l.data(), u.data(), solver.tol, level == 3 ? 0 : maxRounds, 1e100, &res, nullptr, 0, 0);
      tCapi += now() - t0;
      tCapiSing += t1 - t0;
    }
    for (int r = 0; r < reps; ++r) {  // phase 1 alone, as the (test double's) solver does it
      solver.lb = lb0;
      solver.ub = ub0;
      const double t0 = now();
      int nF = 0;
      solver.tightenBoundsFromSingletonRows(nF);
      tSolverSing += now() - t0;
    }
    {  // the steps of run()'s phase 2, one by one
      double tCopy = 0, tProp = 0, tBack = 0;
      std::vector<double> lb2, ub2;
      std::vector<int> perRound(1024), touched;
      for (int r = 0; r < reps && m; ++r) {
        solver.lb = lb0;
        solver.ub = ub0;
        int nF = 0;
        solver.tightenBoundsFromSingletonRows(nF);
        double t0 = now();
        lb2.assign(solver.getColLower(), solver.getColLower() + solver.getNumCols());
        ub2.assign(solver.getColUpper(), solver.getColUpper() + solver.getNumCols());
        double t1 = now();
        cbp_result res;
        cbp_propagate(m, lb2.data(), ub2.data(), solver.tol, level == 3 ? 0 : maxRounds, 1e100, &res, perRound.data(), 1024, 0);
        double t2 = now();
        touched.clear();
        for (int j = 0; j < solver.getNumCols(); ++j)
          if (lb2[j] != solver.lb[j] || ub2[j] != solver.ub[j])
            touched.push_back(j);
        for (int j : touched) {
          solver.setColLower(j, lb2[j]);
          solver.setColUpper(j, ub2[j]);
        }
        double t3 = now();
        tCopy += t1 - t0;
        tProp += t2 - t1;
        tBack += t3 - t2;
      }
      printf("STEPS {\"copy_us\": %.2f, \"propagate_us\": %.2f, \"write_back_us\": %.2f}\n", 1e6 * tCopy / reps,
        1e6 * tProp / reps, 1e6 * tBack / reps);
    }
    printf("PHASES {\"capi_singletons_us\": %.2f, \"solver_singletons_us\": %.2f}\n", 1e6 * tCapiSing / reps,
      1e6 * tSolverSing / reps);
    const CbpCacheStats cs = cbpCacheStats();
    printf("RESULT {\"bench\": 1, \"reps\": %d, \"rounds\": %d, \"run_us\": %.2f, \"capi_us\": %.2f, \"acquire_us\": %.2f, "
           "\"host_share\": %.4f, \"acquire_share\": %.4f, \"cache\": {\"calls\": %lld, \"pointerHits\": %lld, "
           "\"contentHits\": %lld, \"fullHashes\": %lld, \"uploads\": %lld}}\n",
      reps, rounds, 1e6 * tRun / reps, 1e6 * tCapi / reps, 1e6 * tAcq / reps, tRun > 0 ? (tRun - tCapi) / tRun : 0.0,
      tRun > 0 ? tAcq / tRun : 0.0, cs.calls, cs.pointerHits, cs.contentHits, cs.fullHashes, cs.uploads);
    cbpReleaseThreadCache();
    return 0;
  }
  if (dive) {
    // the block of CbcHeuristicDive::solution, verbatim in shape
    const CoinPackedMatrix *matByRow = solver.getMatrixByRow();
    int totalBoundFixed = 0, reasonToStop = 0;
    CoinBoundPropagation bt(solver.getNumCols(), solver.getColType(false), solver.getColLower(),
      solver.getColUpper(), matByRow, solver.getRowSense(), solver.getRightHandSide(), solver.getRowRange(),
      solver.tol, solver.getInfinity());
    if (bt.isInfeasible()) {
      reasonToStop = 1;
    } else {
      const auto &bounds = bt.updatedBounds();
      for (const auto &p : bounds) {
        int col = static_cast<int>(p.first);
        solver.setColLower(col, p.second.first);
        solver.setColUpper(col, p.second.second);
        totalBoundFixed++;
      }
    }
    printf("RESULT {\"dive\": 1, \"ran\": %d, \"infeasible\": %d, \"infeasibleRow\": %d, \"infeasibleCol\": %d, "
           "\"totalBoundFixed\": %d, \"reasonToStop\": %d}\n",
      bt.ran() ? 1 : 0, bt.isInfeasible() ? 1 : 0, bt.infeasibleRow(), bt.infeasibleCol(), totalBoundFixed,
      reasonToStop);
  } else {
    // the calling sequence of CbcSolver::applyLpMethod (src/CbcSolver.cpp:1092-1105)
    CbcBoundPropagation bp;
    const double startTime = CoinGetTimeOfDay();
    const bool feasible = bp.run(&solver, NULL, logLevel, static_cast<CbcBoundPropagation::Level>(level), maxRounds,
      true, timeLimit, startTime);
    printf("RESULT {\"feasible\": %d, \"nSingletonTightened\": %d, \"nSingletonFixed\": %d, \"nBoundPropFixed\": %d, "
           "\"nFixed\": %d, \"nRoundsRun\": %d, \"stopReason\": %d, \"infeasibleRow\": %d, \"infeasibleCol\": %d, "
           "\"nSetLower\": %d, \"nSetUpper\": %d, \"timeUsed\": %.6f}\n",
      feasible ? 1 : 0, bp.nSingletonTightened(), bp.nSingletonFixed(), bp.nBoundPropFixed(), bp.nFixed(),
      bp.nRoundsRun(), static_cast<int>(bp.stopReason()), bp.infeasibleRow(), bp.infeasibleCol(), solver.nSetLower,
      solver.nSetUpper, bp.timeUsed());
    // a second run on the now-propagated solver: same matrix, must hit the cache and be a fixpoint
    if (feasible && bp.stopReason() == CbcBoundPropagation::ReachedFixpoint && level >= 2) {
      CbcBoundPropagation again;
      const bool f2 = again.run(&solver, NULL, 0, CbcBoundPropagation::Fixpoint, 0, true, 1e100, 0.0);
      printf("AGAIN {\"feasible\": %d, \"nRoundsRun\": %d, \"nBoundPropFixed\": %d, \"stopReason\": %d}\n", f2 ? 1 : 0,
        again.nRoundsRun(), again.nBoundPropFixed(), static_cast<int>(again.stopReason()));
    }
  }
  FILE *o = fopen(argv[2], "wb");
  if (o) {
    fwrite(solver.lb.data(), 8, solver.lb.size(), o);
    fwrite(solver.ub.data(), 8, solver.ub.size(), o);
    fclose(o);
  }
  cbpReleaseThreadCache();
  return 0;
}
