// The propagation steps of T concurrent dives as one batch of node bound diffs per wave
// (/root/reference/src/CbcDiveScheduler.cpp:158-214, src/CbcHeuristicDive.cpp:967-1001).
#include "CbcDiveBatch.hpp"

#include <cstdint>
#include <cstdio>

#include "OsiSolverInterface.hpp"
#include "cbp_capi.h"

CbcDiveBatch::CbcDiveBatch(OsiSolverInterface *node, int numberDives, int maxRounds)
  : ctx_(nullptr)
  , mat_(nullptr)
  , n_(0)
  , maxRounds_(maxRounds)
  , tol_(1.0e-7)
  , nActive_(numberDives > 0 ? numberDives : 0)
  , nDeposited_(0)
  , generation_(0)
  , waves_(0)
  , vectors_(0)
{
  slot_.resize(nActive_);
  for (Slot &s : slot_) {
    s.lb = s.ub = nullptr;
    s.active = true;
    s.deposited = false;
    s.feasible = true;
    s.infRow = s.infCol = -1;
  }
  if (!node)
    return;
  const CoinPackedMatrix *m = node->getMatrixByRow();
  n_ = node->getNumCols();
  if (!m || n_ <= 0)
    return;
  node->getDblParam(OsiPrimalTolerance, tol_);
  baseLb_.assign(node->getColLower(), node->getColLower() + n_);
  baseUb_.assign(node->getColUpper(), node->getColUpper() + n_);
  ctx_ = cbp_create(-1);
  if (!ctx_) {
    fprintf(stderr, "CbcDiveBatch: no usable CUDA device (%s); dives are not propagated\n", cbp_last_error(nullptr));
    return;
  }
  const int rows = m->getNumRows();
  const int *start = m->getVectorStarts(), *len = m->getVectorLengths();
  std::vector<int64_t> starts(static_cast<size_t>(rows) + 1, 0);
  for (int i = 0; i < rows; ++i)
    starts[i] = start[i];
  starts[rows] = rows > 0 ? static_cast<int64_t>(start[rows - 1]) + len[rows - 1] : 0;
  mat_ = cbp_matrix_upload(ctx_, rows, n_, starts.data(), len, m->getIndices(), m->getElements(), node->getRowSense(),
    node->getRightHandSide(), node->getRowRange(), node->getColType(false), node->getInfinity());
  if (!mat_)
    fprintf(stderr, "CbcDiveBatch: matrix upload failed (%s); dives are not propagated\n", cbp_last_error(ctx_));
}

CbcDiveBatch::~CbcDiveBatch()
{
  if (mat_)
    cbp_matrix_free(mat_);
  if (ctx_)
    cbp_destroy(ctx_);
}

// every running dive has deposited: build the wave's diffs (each vector against the node's bounds), one device call,
// hand the answers out.  Called with the mutex held.
void CbcDiveBatch::runWaveLocked()
{
  std::vector<int> who;
  for (size_t d = 0; d < slot_.size(); ++d)
    if (slot_[d].deposited)
      who.push_back(static_cast<int>(d));
  const int B = static_cast<int>(who.size());
  std::vector<int64_t> start(B + 1, 0);
  std::vector<uint32_t> var;
  std::vector<double> bnd;
  for (int i = 0; i < B; ++i) {
    const Slot &s = slot_[who[i]];
    for (int j = 0; j < n_; ++j) {
      if (s.lb[j] != baseLb_[j]) {
        var.push_back(static_cast<uint32_t>(j));
        bnd.push_back(s.lb[j]);
      }
      if (s.ub[j] != baseUb_[j]) {
        var.push_back(static_cast<uint32_t>(j) | 0x80000000u);
        bnd.push_back(s.ub[j]);
      }
    }
    start[i + 1] = static_cast<int64_t>(var.size());
  }
  std::vector<cbp_result> res(B);
  std::vector<int64_t> updStart(B + 1, 0);
  std::vector<int> updCol;
  std::vector<double> updLb, updUb;
  bool ok = mat_ != nullptr && B > 0;
  if (ok) {
    int64_t cap = 256 * static_cast<int64_t>(B) + 1024;
    for (int attempt = 0; attempt < 2 && ok; ++attempt) {
      updCol.resize(cap);
      updLb.resize(cap);
      updUb.resize(cap);
      const int rc = attempt == 0
        ? cbp_propagate_batch_delta(mat_, baseLb_.data(), baseUb_.data(), B, start.data(), var.data(), bnd.data(), tol_,
            maxRounds_, 1.0e100, res.data(), cap, updStart.data(), updCol.data(), updLb.data(), updUb.data(), 0)
        : cbp_batch_fetch(mat_, res.data(), cap, updStart.data(), updCol.data(), updLb.data(), updUb.data(), 0);
      if (rc == CBP_OK)
        break;
      if (rc == CBP_ERR_CAPACITY && attempt == 0) {
        cap = updStart[B];
        continue;
      }
      fprintf(stderr, "CbcDiveBatch: device run failed (%d: %s); this wave is not propagated\n", rc, cbp_last_error(ctx_));
      ok = false;
    }
  }
  for (int i = 0; i < B; ++i) {
    Slot &s = slot_[who[i]];
    s.updated.clear();
    s.feasible = true;
    s.infRow = s.infCol = -1;
    if (ok) {
      s.feasible = res[i].feasible != 0;
      s.infRow = res[i].infeasibleRow;
      s.infCol = res[i].infeasibleCol;
      for (int64_t k = updStart[i]; s.feasible && k < updStart[i + 1]; ++k)
        s.updated.push_back(std::make_pair(static_cast<std::size_t>(updCol[k]), std::make_pair(updLb[k], updUb[k])));
    }
    s.deposited = false;
  }
  nDeposited_ = 0;
  ++waves_;
  vectors_ += B;
  ++generation_;
  cv_.notify_all();
}

bool CbcDiveBatch::propagate(int d, const double *lb, const double *ub, BoundList &updated, int &infeasibleRow,
  int &infeasibleCol)
{
  updated.clear();
  infeasibleRow = infeasibleCol = -1;
  if (d < 0 || d >= static_cast<int>(slot_.size()) || !lb || !ub)
    return true;
  std::unique_lock<std::mutex> lock(mtx_);
  Slot &s = slot_[d];
  if (!s.active)
    return true;
  s.lb = lb;
  s.ub = ub;
  s.deposited = true;
  ++nDeposited_;
  if (nDeposited_ == nActive_) {
    runWaveLocked();
  } else {
    const unsigned long gen = generation_;
    cv_.wait(lock, [&] { return generation_ != gen; });
  }
  updated.swap(s.updated);
  infeasibleRow = s.infRow;
  infeasibleCol = s.infCol;
  return s.feasible;
}

void CbcDiveBatch::finish(int d)
{
  if (d < 0 || d >= static_cast<int>(slot_.size()))
    return;
  std::lock_guard<std::mutex> lock(mtx_);
  Slot &s = slot_[d];
  if (!s.active)
    return;
  s.active = false;
  --nActive_;
  if (nActive_ > 0 && nDeposited_ == nActive_)
    runWaveLocked();
}
