// CbcDiveBatch.hpp — the propagation steps of several concurrent dives served by ONE device call per wave.
//
// Where it plugs in: CbcDiveScheduler::runParallel runs its dive heuristics as std::threads
// (/root/reference/src/CbcDiveScheduler.cpp:158-214); every dive, after each fixing, constructs a
// CoinBoundPropagation over its own bounds — one propagation round — and stops without an LP when that
// proves infeasibility (src/CbcHeuristicDive.cpp:967-1001).  All dives of a schedule work on clones of one
// model: same matrix, bound vectors that differ from the node's only by their own fixings.  With this object
// the T threads deposit their bound vectors instead; when the last running dive has arrived the wave goes to
// the device as one batch of node bound diffs (cbp_propagate_batch_delta, include/cbp_capi.h) and every
// thread leaves with its own verdict and its own list of changed columns — the four things the dive reads
// from CoinBoundPropagation (isInfeasible / infeasibleRow / infeasibleCol / updatedBounds).
//
// No counterpart in the reference (each dive thread talks to CoinUtils on its own there); it is the host
// side of SURVEY.md 8(f) N2.  The object owns one device context and one device copy of the matrix, uploaded
// once when it is constructed (on the thread that sets the schedule up), whatever thread ends up issuing a wave.
#ifndef CbcDiveBatch_hpp
#define CbcDiveBatch_hpp

#include <condition_variable>
#include <cstddef>
#include <mutex>
#include <utility>
#include <vector>

class OsiSolverInterface;
struct cbp_ctx;
struct cbp_matrix;

class CbcDiveBatch {
public:
  typedef std::vector<std::pair<std::size_t, std::pair<double, double> > > BoundList;

  /** node: the solver every dive was cloned from (matrix, row data, column types, tolerance and the node's column
      bounds are taken from it now).  numberDives: dives that take part, numbered 0 .. numberDives-1.
      maxRounds: rounds per step (1 = what CbcHeuristicDive does; <= 0 = to the fixpoint). */
  CbcDiveBatch(OsiSolverInterface *node, int numberDives, int maxRounds = 1);
  ~CbcDiveBatch();

  /// false if no device could be used (then propagate() changes nothing and reports every vector feasible)
  bool usable() const { return mat_ != nullptr; }

  /** Called by the thread of dive d with the dive's current column bounds (n doubles each).  Blocks until every dive
      that is still running has called it for this wave and the wave's device call has returned.  Returns false if
      this dive's vector was proved infeasible (infeasibleRow / infeasibleCol as CoinBoundPropagation reports them);
      otherwise `updated` lists the columns whose bounds changed, with their new bounds. */
  bool propagate(int d, const double *lb, const double *ub, BoundList &updated, int &infeasibleRow, int &infeasibleCol);

  /// dive d is over (solution found, infeasible, aborted): later waves do not wait for it
  void finish(int d);

  int wavesRun() const { return waves_; }
  long long vectorsPropagated() const { return vectors_; }

private:
  void runWaveLocked();

  struct Slot {
    const double *lb, *ub;
    bool active, deposited, feasible;
    int infRow, infCol;
    BoundList updated;
  };
  cbp_ctx *ctx_;
  cbp_matrix *mat_;
  int n_, maxRounds_;
  double tol_;
  std::vector<double> baseLb_, baseUb_;
  std::vector<Slot> slot_;
  int nActive_, nDeposited_;
  unsigned long generation_;
  int waves_;
  long long vectors_;
  std::mutex mtx_;
  std::condition_variable cv_;
};

#endif
