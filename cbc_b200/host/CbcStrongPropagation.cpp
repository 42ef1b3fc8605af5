// All arms of a strong-branching step as one batch of node bound diffs
// (/root/reference/src/CbcNode.cpp:3508-3626; diff encoding src/CbcNode.cpp:265-273).
#include "CbcStrongPropagation.hpp"

#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <algorithm>

#include "CbpDeviceCache.hpp"
#include "OsiSolverInterface.hpp"

CbcStrongPropagation::CbcStrongPropagation()
  : nodeInfeasible_(false)
  , nInfeasible_(0)
{
}

bool CbcStrongPropagation::run(OsiSolverInterface *solver, int numberToDo, const int *columns, const double *values,
  int maxRounds, double timeLimit)
{
  arm_.clear();
  updCol_.clear();
  updLb_.clear();
  updUb_.clear();
  down_.clear();
  up_.clear();
  nodeInfeasible_ = false;
  nInfeasible_ = 0;
  if (!solver || numberToDo <= 0 || !columns || !values)
    return true;
  const CoinPackedMatrix *matByRow = solver->getMatrixByRow();
  const int nCols = solver->getNumCols();
  if (!matByRow || nCols <= 0)
    return true;
  CbpProblemView view;
  view.m = matByRow->getNumRows();
  view.n = nCols;
  view.rowStart32 = matByRow->getVectorStarts();
  view.rowLen = matByRow->getVectorLengths();
  view.colIdx = matByRow->getIndices();
  view.val = matByRow->getElements();
  view.rowSense = solver->getRowSense();
  view.rhs = solver->getRightHandSide();
  view.range = solver->getRowRange();
  view.colType = solver->getColType(false);
  view.infinity = solver->getInfinity();
  cbp_matrix *mat = cbpAcquireMatrix(view);
  if (!mat)
    return false;
  double primalTol = 1.0e-7;
  solver->getDblParam(OsiPrimalTolerance, primalTol);

  // arm 2k: down (ub = floor(value)), arm 2k+1: up (lb = ceil(value))
  const int B = 2 * numberToDo;
  std::vector<int64_t> start(B + 1);
  std::vector<uint32_t> var(B);
  std::vector<double> bnd(B);
  down_.resize(numberToDo);
  up_.resize(numberToDo);
  for (int k = 0; k < numberToDo; ++k) {
    down_[k] = std::floor(values[k]);
    up_[k] = std::ceil(values[k]);
    if (up_[k] == down_[k])  // integral value: the arms of CbcSimpleInteger still differ by one
      up_[k] = down_[k] + 1.0;
    var[2 * k] = static_cast<uint32_t>(columns[k]) | 0x80000000u;
    bnd[2 * k] = down_[k];
    var[2 * k + 1] = static_cast<uint32_t>(columns[k]);
    bnd[2 * k + 1] = up_[k];
  }
  for (int b = 0; b <= B; ++b)
    start[b] = b;

  // The device keeps a batch's bounds as [column][vector] cells (two buffers of 16 bytes): a sweep over every integer
  // column of a large model (CbcNode::analyze, fixVubs) does not fit as one batch, so the arms go in pieces of at most
  // CBP_BATCH_BYTES (default 16 GB) of cells — whole candidates, a multiple of 32 arms.
  static const double batchBytes = getenv("CBP_BATCH_BYTES") ? atof(getenv("CBP_BATCH_BYTES")) : 16.0e9;
  long long perCall = static_cast<long long>(batchBytes / (33.0 * static_cast<double>(nCols)));
  perCall = std::max<long long>(64, perCall / 32 * 32);
  std::vector<cbp_result> res(B);
  std::vector<int64_t> updStart(B + 1, 0);
  std::vector<int64_t> pieceStart;
  std::vector<int> pieceCol;
  std::vector<double> pieceLb, pieceUb;
  for (long long o = 0; o < B; o += perCall) {
    const int Bc = static_cast<int>(std::min<long long>(perCall, B - o));
    pieceStart.assign(static_cast<size_t>(Bc) + 1, 0);
    int64_t cap = 64 * static_cast<int64_t>(Bc) + 1024;
    for (int attempt = 0; attempt < 2; ++attempt) {
      pieceCol.resize(cap);
      pieceLb.resize(cap);
      pieceUb.resize(cap);
      const int rc = attempt == 0
        ? cbp_propagate_batch_delta(mat, solver->getColLower(), solver->getColUpper(), Bc, start.data(), var.data() + o,
            bnd.data() + o, primalTol, maxRounds, timeLimit, res.data() + o, cap, pieceStart.data(), pieceCol.data(),
            pieceLb.data(), pieceUb.data(), 0)
        : cbp_batch_fetch(mat, res.data() + o, cap, pieceStart.data(), pieceCol.data(), pieceLb.data(), pieceUb.data(), 0);
      if (rc == CBP_OK)
        break;
      if (rc == CBP_ERR_CAPACITY && attempt == 0) {  // the offsets are valid: ask again with room for everything
        cap = pieceStart[Bc];
        continue;
      }
      fprintf(stderr, "CbcStrongPropagation: device run failed (%d); no arm filtered\n", rc);
      updCol_.clear();
      updLb_.clear();
      updUb_.clear();
      return false;
    }
    const int64_t base = updStart[o];
    for (int i = 0; i < Bc; ++i)
      updStart[o + i + 1] = base + pieceStart[i + 1];
    updCol_.insert(updCol_.end(), pieceCol.begin(), pieceCol.begin() + pieceStart[Bc]);
    updLb_.insert(updLb_.end(), pieceLb.begin(), pieceLb.begin() + pieceStart[Bc]);
    updUb_.insert(updUb_.end(), pieceUb.begin(), pieceUb.begin() + pieceStart[Bc]);
  }
  arm_.resize(B);
  for (int b = 0; b < B; ++b) {
    Arm a;
    a.infeasible = !res[b].feasible;
    a.row = res[b].infeasibleRow;
    a.col = res[b].infeasibleCol;
    a.rounds = res[b].rounds;
    a.updBegin = updStart[b];
    a.updEnd = updStart[b + 1];
    arm_[b] = a;
    nInfeasible_ += a.infeasible ? 1 : 0;
  }
  for (int k = 0; k < numberToDo; ++k)
    nodeInfeasible_ = nodeInfeasible_ || (arm_[2 * k].infeasible && arm_[2 * k + 1].infeasible);
  return true;
}

CbcStrongPropagation::BoundList CbcStrongPropagation::impliedBounds(int k, Way way) const
{
  BoundList out;
  const Arm &a = arm_[2 * k + way];
  for (long long i = a.updBegin; i < a.updEnd; ++i)
    out.push_back(std::make_pair(static_cast<std::size_t>(updCol_[i]), std::make_pair(updLb_[i], updUb_[i])));
  return out;
}

bool CbcStrongPropagation::forcedBound(int k, bool &isUpper, double &bound) const
{
  const bool d = arm_[2 * k].infeasible, u = arm_[2 * k + 1].infeasible;
  if (d == u)
    return false;
  if (d) {  // down cannot hold: the variable is at least ceil(value)
    isUpper = false;
    bound = up_[k];
  } else {
    isUpper = true;
    bound = down_[k];
  }
  return true;
}
