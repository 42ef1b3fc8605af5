// CbcBoundPropagation.cpp — run() on top of the C ABI (cbp_capi.h).
// Control flow, counters, stop reasons and the printf lines follow
// /root/reference/src/CbcBoundPropagation.cpp:30-210; the difference is that the
// whole round loop (:107-195) is one call into the device engine.
#include "CbcBoundPropagation.hpp"

#include <cassert>
#include <climits>
#include <cstdio>
#include <string>
#include <vector>

#include "CbpDeviceCache.hpp"
#include "OsiSolverInterface.hpp"

CbcBoundPropagation::CbcBoundPropagation()
  : nSingletonTightened_(0)
  , nSingletonFixed_(0)
  , nBoundPropFixed_(0)
  , nRoundsRun_(0)
  , stopReason_(NotRun)
  , timeUsed_(0.0)
  , infeasibleRow_(-1)
  , infeasibleCol_(-1)
{
}

namespace {
inline double now(bool elapsed) { return elapsed ? CoinGetTimeOfDay() : CoinCpuTime(); }

std::string nameOrUnknown(bool inRange, const std::string &name) { return inRange ? name : std::string("(unknown)"); }
}  // namespace

bool CbcBoundPropagation::run(OsiSolverInterface *solver, CoinMessageHandler * /*handler*/, int logLevel,
  Level level, int maxRounds, bool useElapsed, double timeLimit, double startTime)
{
  assert(level != Off);
  if (!solver || solver->getNumCols() == 0)
    return true;  // reference :41-42
  const double t0 = now(useElapsed);

  // ---- phase 1: the solver's own singleton-row tightening (reference :49-72)
  int nFixed = 0;
  const int nTightened = solver->tightenBoundsFromSingletonRows(nFixed);
  if (nTightened < 0) {
    stopReason_ = InfeasibleDetected;
    timeUsed_ = now(useElapsed) - t0;
    if (logLevel >= 1)
      printf("  Bound propagation: INFEASIBLE (singleton tightening), %.3f s.\n", timeUsed_);
    return false;
  }
  nSingletonFixed_ = nFixed;
  nSingletonTightened_ = nTightened - nFixed;
  if (logLevel >= 2 && nTightened > 0)
    printf("  Bound propagation (singletons): %d tightened, %d fixed.\n", nSingletonTightened_, nSingletonFixed_);
  if (level == Singletons) {  // reference :74-83
    stopReason_ = ReachedFixpoint;
    timeUsed_ = now(useElapsed) - t0;
    if (logLevel >= 1)
      printf("  Bound propagation fixed %d vars in %.3f s.\n", nSingletonFixed_, timeUsed_);
    return true;
  }

  // ---- phase 2: rounds on the device
  const int nCols = solver->getNumCols();
  const CoinPackedMatrix *byRow = solver->getMatrixByRow();
  double primalTol = 1e-7;
  solver->getDblParam(OsiPrimalTolerance, primalTol);
  // (per-thread scratch, kept between calls: a node-level caller comes back thousands of times, and fresh
  // megabyte-sized vectors per call cost more in page faults than the device run)
  static thread_local std::vector<double> lb, ub;
  static thread_local std::vector<int> perRound, touched;
  lb.assign(solver->getColLower(), solver->getColLower() + nCols);
  ub.assign(solver->getColUpper(), solver->getColUpper() + nCols);

  bool ranOnDevice = false;
  cbp_result res;
  const int roundLimit = (level == Fixpoint) ? INT_MAX : maxRounds;  // reference :88
  if (roundLimit > 0) {
    // the reference tests the clock before every round, the first included (:109-121)
    const double tNow = now(useElapsed);
    if (tNow - startTime >= timeLimit) {
      stopReason_ = HitTimeLimit;
      timeUsed_ = tNow - t0;
      if (logLevel >= 1)
        printf("  Bound propagation: fixed %d (%d singleton + %d propagation, %d round(s), TIME LIMIT), %.3f s.\n",
          nSingletonFixed_ + nBoundPropFixed_, nSingletonFixed_, nBoundPropFixed_, nRoundsRun_, timeUsed_);
      return true;
    }
    CbpProblemView view;
    view.m = byRow ? byRow->getNumRows() : 0;
    view.n = nCols;
    view.rowStart32 = byRow ? byRow->getVectorStarts() : nullptr;
    view.rowLen = byRow ? byRow->getVectorLengths() : nullptr;
    view.colIdx = byRow ? byRow->getIndices() : nullptr;
    view.val = byRow ? byRow->getElements() : nullptr;
    view.rowSense = solver->getRowSense();
    view.rhs = solver->getRightHandSide();
    view.range = solver->getRowRange();
    view.colType = solver->getColType(false);
    view.infinity = solver->getInfinity();
    cbp_matrix *mat = byRow ? cbpAcquireMatrix(view) : nullptr;
    if (mat) {
      perRound.assign(1024, 0);
      const int rc = cbp_propagate(mat, lb.data(), ub.data(), primalTol, level == Fixpoint ? 0 : maxRounds,
        timeLimit - (tNow - startTime), &res, perRound.data(), static_cast<int>(perRound.size()), 0);
      if (rc == CBP_OK)
        ranOnDevice = true;
      else
        fprintf(stderr, "CbcBoundPropagation: device run failed (%d); no bounds were tightened\n", rc);
    }
    if (!ranOnDevice) {  // documented: tighten nothing, claim nothing
      timeUsed_ = now(useElapsed) - t0;
      return true;
    }
  }

  if (ranOnDevice) {
    nRoundsRun_ = res.rounds;
    nBoundPropFixed_ = res.nChanged;
    // write back what completed rounds changed (reference :181-187 sets both sides of a listed column)
    const double *curLb = solver->getColLower();
    const double *curUb = solver->getColUpper();
    touched.clear();
    for (int j = 0; j < nCols; ++j)
      if (lb[j] != curLb[j] || ub[j] != curUb[j])
        touched.push_back(j);
    for (int j : touched) {
      solver->setColLower(j, lb[j]);
      solver->setColUpper(j, ub[j]);
    }
    if (logLevel >= 2) {
      int total = 0;
      const int done = res.rounds - (res.feasible ? 0 : 1);
      for (int r = 0; r < done && r < static_cast<int>(perRound.size()); ++r) {
        if (perRound[r] == 0)
          continue;
        total += perRound[r];
        printf("  Bound propagation: round %d fixed %d variables (total %d).\n", r + 1, perRound[r], total);
      }
    }
    if (!res.feasible) {  // reference :131-163
      infeasibleRow_ = res.infeasibleRow;
      infeasibleCol_ = res.infeasibleCol;
      stopReason_ = InfeasibleDetected;
      timeUsed_ = now(useElapsed) - t0;
      if (logLevel >= 1) {
        if (infeasibleRow_ >= 0 && infeasibleCol_ >= 0) {
          printf("  Bound propagation: INFEASIBLE in round %d — row %d (%s), col %d (%s), %.3f s.\n", nRoundsRun_,
            infeasibleRow_,
            nameOrUnknown(infeasibleRow_ < solver->getNumRows(), solver->getRowName(infeasibleRow_)).c_str(),
            infeasibleCol_,
            nameOrUnknown(infeasibleCol_ < solver->getNumCols(), solver->getColName(infeasibleCol_)).c_str(),
            timeUsed_);
        } else if (infeasibleRow_ >= 0) {
          printf("  Bound propagation: INFEASIBLE in round %d — row %d (%s), %.3f s.\n", nRoundsRun_, infeasibleRow_,
            nameOrUnknown(infeasibleRow_ < solver->getNumRows(), solver->getRowName(infeasibleRow_)).c_str(),
            timeUsed_);
        } else {
          printf("  Bound propagation: INFEASIBLE in round %d, %.3f s.\n", nRoundsRun_, timeUsed_);
        }
      }
      return false;
    }
    if (res.stopReason == CBP_STOP_TIME_LIMIT) {  // reference :109-121
      stopReason_ = HitTimeLimit;
      timeUsed_ = now(useElapsed) - t0;
      if (logLevel >= 1)
        printf("  Bound propagation: fixed %d (%d singleton + %d propagation, %d round(s), TIME LIMIT), %.3f s.\n",
          nSingletonFixed_ + nBoundPropFixed_, nSingletonFixed_, nBoundPropFixed_, nRoundsRun_, timeUsed_);
      return true;
    }
    if (res.stopReason == CBP_STOP_FIXPOINT) {
      stopReason_ = ReachedFixpoint;
      if (logLevel >= 2)
        printf("  Bound propagation: fixpoint reached after %d round(s).\n", nRoundsRun_);
    }
  }
  if (stopReason_ == NotRun)
    stopReason_ = HitMaxRounds;  // reference :197-200
  timeUsed_ = now(useElapsed) - t0;
  if (logLevel >= 1)
    printf("  Bound propagation fixed %d vars in %.3f s.\n", nSingletonFixed_ + nBoundPropFixed_, timeUsed_);
  return true;
}
