// One propagation round on the device behind the CoinUtils-style interface
// that CbcHeuristicDive uses (/root/reference/src/CbcHeuristicDive.cpp:969-999).
#include "CoinBoundPropagation.hpp"

#include <cstdio>
#include <vector>

#include "CbpDeviceCache.hpp"
#include "OsiSolverInterface.hpp"

CoinBoundPropagation::CoinBoundPropagation(int nCols, const char *colType, const double *colLB,
  const double *colUB, const CoinPackedMatrix *matByRow, const char *rowSense, const double *rowRHS,
  const double *rowRange, double primalTol, double infinity)
  : infeasible_(false)
  , ran_(false)
  , infeasibleRow_(-1)
  , infeasibleCol_(-1)
{
  if (!matByRow || nCols <= 0)
    return;
  CbpProblemView view;
  view.m = matByRow->getNumRows();
  view.n = nCols;
  view.rowStart32 = matByRow->getVectorStarts();
  view.rowLen = matByRow->getVectorLengths();
  view.colIdx = matByRow->getIndices();
  view.val = matByRow->getElements();
  view.rowSense = rowSense;
  view.rhs = rowRHS;
  view.range = rowRange;
  view.colType = colType;
  view.infinity = infinity;
  cbp_matrix *mat = cbpAcquireMatrix(view);
  if (!mat)
    return;
  std::vector<double> lb(colLB, colLB + nCols), ub(colUB, colUB + nCols);
  cbp_result res;
  if (cbp_propagate(mat, lb.data(), ub.data(), primalTol, 1, 1.0e100, &res, nullptr, 0, 0) != CBP_OK) {
    fprintf(stderr, "CoinBoundPropagation: device run failed; nothing tightened\n");
    return;
  }
  ran_ = true;
  if (!res.feasible) {
    infeasible_ = true;
    infeasibleRow_ = res.infeasibleRow;
    infeasibleCol_ = res.infeasibleCol;
    return;
  }
  for (int j = 0; j < nCols; ++j)
    if (lb[j] != colLB[j] || ub[j] != colUB[j])
      updated_.push_back(std::make_pair(static_cast<std::size_t>(j), std::make_pair(lb[j], ub[j])));
}
