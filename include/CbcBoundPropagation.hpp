// CbcBoundPropagation.hpp — the class Cbc's callers already use, backed by the
// B200 engine.  Same name, enums, run() signature, accessors and data-member
// order as /root/reference/src/CbcBoundPropagation.hpp:42-132 (the accessors
// are inline there, so the member layout is ABI); what differs is what run()
// does behind it: the round loop executes on the GPU through the C ABI in
// cbp_capi.h instead of constructing a CoinBoundPropagation per round
// (/root/reference/src/CbcBoundPropagation.cpp:107-195).
//
// Callers that keep working unchanged: CbcSolver::applyLpMethod
// (src/CbcSolver.cpp:1092-1100), the -doBoundProp action (:13471-13476),
// CbcModel::doOneNode (src/CbcModel.cpp:18268-18271),
// CbcHeuristic::smallBranchAndBound (src/CbcHeuristic.cpp:795-811) and
// Cbc_resolve (src/Cbc_C_Interface.cpp:2136).
#ifndef CbcBoundPropagation_hpp
#define CbcBoundPropagation_hpp

#if defined(__has_include)
#if __has_include("CbcConfig.h")
#include "CbcConfig.h"
#endif
#endif
#ifndef CBCLIB_EXPORT
#define CBCLIB_EXPORT __attribute__((visibility("default")))
#endif

class OsiSolverInterface;
class CoinMessageHandler;

class CBCLIB_EXPORT CbcBoundPropagation {
public:
  // how far run() goes (reference :52-57)
  enum Level { Off = 0, Singletons, MILPbt, Fixpoint };
  // why the round loop ended (reference :60-66)
  enum StopReason { NotRun = 0, ReachedFixpoint, HitMaxRounds, HitTimeLimit, InfeasibleDetected };

  CbcBoundPropagation();

  // Phase 1 = solver->tightenBoundsFromSingletonRows(); phase 2 = propagation
  // rounds on the device until fixpoint (Fixpoint), maxRounds (MILPbt), the
  // time limit, or proven infeasibility.  Tightened bounds are written to the
  // solver with setColLower / setColUpper.  Returns false iff infeasibility
  // was proved.  `handler` is ignored, output goes to stdout gated on logLevel,
  // exactly as the reference does (:31).  If the device cannot be used the
  // call prints one line to stderr, tightens nothing and returns true with
  // stopReason() == NotRun — it never claims infeasibility it did not prove.
  bool run(OsiSolverInterface *solver, CoinMessageHandler *handler, int logLevel, Level level,
    int maxRounds, bool useElapsed, double timeLimit, double startTime);

  int nSingletonTightened() const { return nSingletonTightened_; }
  int nSingletonFixed() const { return nSingletonFixed_; }
  int nBoundPropFixed() const { return nBoundPropFixed_; }
  int nFixed() const { return nSingletonFixed_ + nBoundPropFixed_; }
  int nRoundsRun() const { return nRoundsRun_; }
  StopReason stopReason() const { return stopReason_; }
  double timeUsed() const { return timeUsed_; }
  int infeasibleRow() const { return infeasibleRow_; }
  int infeasibleCol() const { return infeasibleCol_; }

private:
  // order and types as in the reference (:123-131)
  int nSingletonTightened_;
  int nSingletonFixed_;
  int nBoundPropFixed_;
  int nRoundsRun_;
  StopReason stopReason_;
  double timeUsed_;
  int infeasibleRow_;
  int infeasibleCol_;
};

#endif
