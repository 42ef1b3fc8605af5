// CoinBoundPropagation.hpp — GPU-backed stand-in for the CoinUtils class that
// CbcHeuristicDive constructs directly, one round per dive step
// (/root/reference/src/CbcHeuristicDive.cpp:969-999).  Constructing it runs ONE
// propagation round on the device; the results are read with the same four
// calls the reference makes (src/CbcBoundPropagation.cpp:131-133,166).
#ifndef CoinBoundPropagation_hpp
#define CoinBoundPropagation_hpp

#include <cstddef>
#include <utility>
#include <vector>

class CoinPackedMatrix;

class CoinBoundPropagation {
public:
  typedef std::vector<std::pair<std::size_t, std::pair<double, double> > > BoundList;

  CoinBoundPropagation(int nCols, const char *colType, const double *colLB, const double *colUB,
    const CoinPackedMatrix *matByRow, const char *rowSense, const double *rowRHS, const double *rowRange,
    double primalTol, double infinity);

  bool isInfeasible() const { return infeasible_; }
  int infeasibleRow() const { return infeasibleRow_; }
  int infeasibleCol() const { return infeasibleCol_; }
  // columns whose bounds changed in this round: (col, (newLB, newUB)); empty = fixpoint
  const BoundList &updatedBounds() const { return updated_; }
  // false if the device could not be used (then nothing is reported as changed)
  bool ran() const { return ran_; }

private:
  bool infeasible_, ran_;
  int infeasibleRow_, infeasibleCol_;
  BoundList updated_;
};

#endif
