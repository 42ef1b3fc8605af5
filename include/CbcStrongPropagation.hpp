// CbcStrongPropagation.hpp — all arms of a strong-branching step propagated in one device call.
//
// Where it plugs in: CbcNode::chooseDynamicBranch builds, for each of its numberToDo candidates, a down
// arm and an up arm and hot-starts an LP for each (/root/reference/src/CbcNode.cpp:3508-3626); an arm whose
// LP is infeasible lets it fix the variable the other way (choice.fix, :3557).  An integer candidate's arms
// are one bound change each: down sets ub = floor(value), up sets lb = ceil(value)
// (src/CbcSimpleInteger.cpp:450-454,499-501).  run() propagates all 2*numberToDo of them to their fixpoints
// as ONE batch over the node's bounds, before any LP is touched: arms that propagation proves infeasible need
// no LP at all, and every surviving arm comes with the bounds it implies.  The same call serves the root
// sweep of CbcNode::analyze (src/CbcNode.cpp:5354,5958-5997).
//
// The class has no counterpart in the reference (there the arms are only ever seen by the LP); it is the
// host side of SURVEY.md §8(f) N1 over cbp_propagate_batch_delta (include/cbp_capi.h).
#ifndef CbcStrongPropagation_hpp
#define CbcStrongPropagation_hpp

#include <cstddef>
#include <utility>
#include <vector>

class OsiSolverInterface;

class CbcStrongPropagation {
public:
  typedef std::vector<std::pair<std::size_t, std::pair<double, double> > > BoundList;
  enum Way { Down = 0, Up = 1 };

  CbcStrongPropagation();

  /** Propagate the 2*numberToDo arms of the candidates (columns[k] with LP value values[k]) over the solver's
      current column bounds.  maxRounds <= 0: to the fixpoint.  Returns false if the device could not be used
      (then no arm is reported infeasible and no bound implied). */
  bool run(OsiSolverInterface *solver, int numberToDo, const int *columns, const double *values,
    int maxRounds = 0, double timeLimit = 1.0e100);

  int numberCandidates() const { return static_cast<int>(arm_.size() / 2); }
  /// the arm cannot hold: propagation ran into an infeasible row or crossing column bounds
  bool infeasible(int k, Way way) const { return arm_[2 * k + way].infeasible; }
  int infeasibleRow(int k, Way way) const { return arm_[2 * k + way].row; }
  int infeasibleCol(int k, Way way) const { return arm_[2 * k + way].col; }
  int rounds(int k, Way way) const { return arm_[2 * k + way].rounds; }
  /// both arms infeasible: the node itself is infeasible
  bool nodeInfeasible() const { return nodeInfeasible_; }
  int numberInfeasibleArms() const { return nInfeasible_; }
  /// columns whose bounds an arm's propagation changed, with the new bounds (the branched column's own
  /// change is part of the arm, not of this list unless propagation tightened it further)
  BoundList impliedBounds(int k, Way way) const;
  /// For a candidate with exactly one infeasible arm: the bound change that fixes it the other way
  /// (what choice.fix does after an infeasible LP); false if neither or both arms are infeasible.
  bool forcedBound(int k, bool &isUpper, double &bound) const;

private:
  struct Arm {
    bool infeasible;
    int row, col, rounds;
    long long updBegin, updEnd;
  };
  std::vector<Arm> arm_;
  std::vector<int> updCol_;
  std::vector<double> updLb_, updUb_, down_, up_;
  bool nodeInfeasible_;
  int nInfeasible_;
};

#endif
