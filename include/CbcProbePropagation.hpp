// CbcProbePropagation.hpp — probing on all candidate columns as one device batch, with the consequences
// Cbc's probing loop draws from it.
//
// Where it plugs in: fixVubs probes every column with lower bound 0 and a positive range by faking its
// upper bound to 0 and propagating the rows it touches (/root/reference/src/CbcSolverHeuristics.cpp:265-490):
// columns that the probe forces to 0 are collected (:394-437), and a probe that turns out infeasible on an
// integer column makes that column's lower bound 1 (:483-486).  CglProbing's bound changes reach the solver the
// same way (src/CbcCutGenerator.cpp:495-624), and CbcNode::analyze sweeps all integer columns both ways at the
// root (src/CbcNode.cpp:5354,5912-5991).  run() does the sweep as ONE batch of node bound diffs: for every
// probed column a "down" vector (ub = lb) and an "up" vector (lb = lb + 1 for integer columns), each run to its
// propagation fixpoint, and then:
//   - a failed down probe on an integer column  => lb = lb + 1   (fixVubs :483-486)
//   - a failed up probe on an integer column    => ub = lb       (the mirror image)
//   - both fail                                 => the node is infeasible
//   - a column that BOTH probes of one candidate leave with bounds tighter than the node's gets the weaker of
//     the two (the classic probing implication; CglProbing's column cuts)
// The per-probe implied bounds are kept as well (impliedBounds): the list fixVubs builds per column.
//
// No counterpart class in the reference; it is the host side of SURVEY.md 8(f) N4 over the same batch entry
// point as CbcStrongPropagation.
#ifndef CbcProbePropagation_hpp
#define CbcProbePropagation_hpp

#include <cstddef>
#include <utility>
#include <vector>

#include "CbcStrongPropagation.hpp"

class OsiSolverInterface;

class CbcProbePropagation {
public:
  typedef CbcStrongPropagation::BoundList BoundList;
  enum Way { Down = 0, Up = 1 };

  CbcProbePropagation();

  /** Probe `columns` (numberColumns of them), or — columns == nullptr — every integer column with a finite range of at
      least 1, at most maxProbes of them (<= 0: all).  maxRounds <= 0: every probe runs to its fixpoint.
      Returns false if the device could not be used (then nothing is concluded). */
  bool run(OsiSolverInterface *solver, int numberColumns = 0, const int *columns = nullptr, int maxProbes = 0,
    int maxRounds = 0);

  int numberProbed() const { return static_cast<int>(cols_.size()); }
  int column(int k) const { return cols_[k]; }
  bool infeasible(int k, Way way) const { return sp_.infeasible(k, static_cast<CbcStrongPropagation::Way>(way)); }
  BoundList impliedBounds(int k, Way way) const { return sp_.impliedBounds(k, static_cast<CbcStrongPropagation::Way>(way)); }
  bool nodeInfeasible() const { return nodeInfeasible_; }
  int numberFailedProbes() const { return sp_.numberInfeasibleArms(); }
  /// what the sweep proves about the node: (column, (new lb, new ub)), every entry at least as tight as the node's bounds
  const BoundList &nodeBounds() const { return nodeBounds_; }
  /// writes nodeBounds() into the solver (setColLower / setColUpper for columns whose bound really moves); returns how many columns moved
  int apply(OsiSolverInterface *solver) const;

private:
  CbcStrongPropagation sp_;
  std::vector<int> cols_;
  BoundList nodeBounds_;
  bool nodeInfeasible_;
};

#endif
