// Probing sweep as one batch (/root/reference/src/CbcSolverHeuristics.cpp:265-490; consequences :483-486).
#include "CbcProbePropagation.hpp"

#include <algorithm>
#include <cmath>
#include <map>

#include "OsiSolverInterface.hpp"

CbcProbePropagation::CbcProbePropagation()
  : nodeInfeasible_(false)
{
}

bool CbcProbePropagation::run(OsiSolverInterface *solver, int numberColumns, const int *columns, int maxProbes,
  int maxRounds)
{
  cols_.clear();
  nodeBounds_.clear();
  nodeInfeasible_ = false;
  if (!solver)
    return true;
  const int n = solver->getNumCols();
  const double *lb = solver->getColLower(), *ub = solver->getColUpper();
  const char *type = solver->getColType(false);
  const double inf = solver->getInfinity();
  if (columns) {
    cols_.assign(columns, columns + numberColumns);
  } else {
    for (int j = 0; j < n; ++j)
      if (type && type[j] != 0 && std::fabs(lb[j]) < inf && std::fabs(ub[j]) < inf && ub[j] - lb[j] >= 1.0)
        cols_.push_back(j);
  }
  if (maxProbes > 0 && static_cast<int>(cols_.size()) > maxProbes)
    cols_.resize(maxProbes);
  if (cols_.empty())
    return true;
  // a value strictly between lb and lb + 1: the down arm is ub = lb (fixVubs' "fake upper bound", :270-276, for
  // lb = 0), the up arm lb = lb + 1
  std::vector<double> value(cols_.size());
  for (size_t k = 0; k < cols_.size(); ++k)
    value[k] = lb[cols_[k]] + 0.5;
  if (!sp_.run(solver, static_cast<int>(cols_.size()), cols_.data(), value.data(), maxRounds))
    return false;
  nodeInfeasible_ = sp_.nodeInfeasible();
  if (nodeInfeasible_)
    return true;
  // consequences for the node: tightest bound per column over everything the sweep proves
  std::map<std::size_t, std::pair<double, double> > best;
  auto tighten = [&](std::size_t j, double l, double u) {
    auto it = best.find(j);
    if (it == best.end())
      it = best.insert(std::make_pair(j, std::make_pair(lb[j], ub[j]))).first;
    it->second.first = std::max(it->second.first, l);
    it->second.second = std::min(it->second.second, u);
  };
  for (size_t k = 0; k < cols_.size(); ++k) {
    const int j = cols_[k];
    const bool dBad = sp_.infeasible(static_cast<int>(k), CbcStrongPropagation::Down);
    const bool uBad = sp_.infeasible(static_cast<int>(k), CbcStrongPropagation::Up);
    bool isUpper = false;
    double b = 0.0;
    if (sp_.forcedBound(static_cast<int>(k), isUpper, b)) {  // one probe fails: the column goes the other way
      if (isUpper)
        tighten(j, lb[j], b);
      else
        tighten(j, b, ub[j]);
    }
    if (dBad || uBad)
      continue;
    // both probes hold: what both of them imply holds at the node (the weaker of the two bounds per column)
    const BoundList d = sp_.impliedBounds(static_cast<int>(k), CbcStrongPropagation::Down);
    const BoundList u = sp_.impliedBounds(static_cast<int>(k), CbcStrongPropagation::Up);
    size_t a = 0, c = 0;  // both lists are in ascending column order
    while (a < d.size() && c < u.size()) {
      if (d[a].first < u[c].first)
        ++a;
      else if (u[c].first < d[a].first)
        ++c;
      else {
        const std::size_t col = d[a].first;
        if (static_cast<int>(col) != j)
          tighten(col, std::min(d[a].second.first, u[c].second.first), std::max(d[a].second.second, u[c].second.second));
        ++a;
        ++c;
      }
    }
  }
  for (const auto &e : best)
    if (e.second.first > lb[e.first] || e.second.second < ub[e.first]) {
      if (e.second.first > e.second.second) {  // two conclusions contradict each other: no feasible point at this node
        nodeInfeasible_ = true;
        nodeBounds_.clear();
        return true;
      }
      nodeBounds_.push_back(std::make_pair(e.first, e.second));
    }
  return true;
}

int CbcProbePropagation::apply(OsiSolverInterface *solver) const
{
  if (!solver)
    return 0;
  int moved = 0;
  for (const auto &e : nodeBounds_) {
    const int j = static_cast<int>(e.first);
    const double l = solver->getColLower()[j], u = solver->getColUpper()[j];
    if (e.second.first > l)
      solver->setColLower(j, e.second.first);
    if (e.second.second < u)
      solver->setColUpper(j, e.second.second);
    moved += (e.second.first > l || e.second.second < u) ? 1 : 0;
  }
  return moved;
}
