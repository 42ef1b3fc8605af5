/*
 * cbp_oracle.h — CPU oracle for the bound-propagation hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under cbc_b200/ or include/ may link,
 * import or call this; only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs use it, and only as the checker or as
 * the timed CPU baseline.
 *
 * PARITY UNPINNED: the arithmetic of the reference path lives in CoinUtils
 * (class CoinBoundPropagation, github.com/coin-or/CoinUtils @ master, un-pinned:
 * /root/reference/.coin-or/Dependencies:6) and in Osi
 * (OsiSolverInterface::tightenBoundsFromSingletonRows, Osi @ master,
 * .coin-or/Dependencies:7).  Neither source tree is on this machine and the
 * reference's tests hold no vectors for this boundary, so this file restates
 * the published algorithm (min/max row activity, per-nonzero candidate
 * bounds, integer rounding, infeasibility detection, iterated to a fixpoint)
 * and anchors on:
 *   - loop / stop / counter semantics: src/CbcBoundPropagation.cpp:30-210
 *   - activity arithmetic with separate infinite-bound counters, redundancy
 *     and infeasibility tests: src/CbcSolverHeuristics.cpp:282-332
 *   - per-nonzero candidate bounds for +/- coefficients:
 *     src/CbcSolverHeuristics.cpp:335-477
 *   - row-sense meaning: src/Cbc_C_Interface.cpp:2219-2235
 *   - column type codes (0 continuous, 1 binary, 2 general integer):
 *     src/CbcHeuristicFPump.cpp:962,991
 * The hand-derivable micro-instances of test/CInterfaceTest.c
 * (:32-59, :328-358, :401-436) are the only known answers it is pinned on
 * (tests/golden/micro_instances.json).
 */
#ifndef CBP_ORACLE_H
#define CBP_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Problem data exactly as CbcBoundPropagation::run extracts it from the
 * solver (src/CbcBoundPropagation.cpp:90-105). */
typedef struct orc_problem {
  int m;                   /* rows    */
  int n;                   /* columns */
  const int64_t *rowStart; /* m+1 entries (CoinBigIndex vectorStarts) */
  const int *rowLen;       /* m entries or NULL (then rowStart[i+1]-rowStart[i]) */
  const int *col;          /* column index per nonzero */
  const double *val;       /* coefficient per nonzero  */
  const char *sense;       /* 'L','G','E','R','N' per row */
  const double *rhs;       /* right-hand side per row */
  const double *range;     /* range per row ('R' rows) */
  const char *colType;     /* 0 continuous, 1 binary, 2 general integer */
  double inf;              /* solver infinity (getInfinity()) */
} orc_problem;

/* stop reasons, same numbering as CbcBoundPropagation::StopReason
 * (src/CbcBoundPropagation.hpp:60-66) */
enum { ORC_NOT_RUN = 0, ORC_FIXPOINT = 1, ORC_MAX_ROUNDS = 2, ORC_TIME_LIMIT = 3, ORC_INFEASIBLE = 4 };

typedef struct orc_result {
  int feasible;      /* 1 unless infeasibility was proved */
  int rounds;        /* rounds executed (the empty fixpoint round counts) */
  int nChanged;      /* sum over rounds of the number of changed columns */
  int infeasibleRow; /* smallest violating row of the failing round, else -1 */
  int infeasibleCol; /* smallest crossing column of the failing round, else -1 */
  int stopReason;
  double seconds;
  double minMargin;  /* smallest relative distance of any accept / round / infeasible
                        decision to its threshold (knife-edge detector) */
} orc_result;

/* Row activities with exact infinite-bound counters (one pass, no tightening). */
void orc_row_activity(const orc_problem *p, const double *lb, const double *ub,
  double *minAct, double *maxAct, int *infMin, int *infMax);

/* Phase 1: rows with exactly one nonzero.  Returns the number of columns whose
 * bounds changed, or -1 if a column's bounds cross (infeasible).  *nFixed = how
 * many of the changed columns ended with lb == ub. */
int orc_singletons(const orc_problem *p, double *lb, double *ub, double tol, int *nFixed);

/* One Jacobi round.  newLb/newUb receive the tightened bounds (equal to lb/ub
 * where nothing changed); changed[j] is set to 1 for changed columns (callers
 * may pre-seed it).  Returns 0 feasible / 1 infeasible. */
int orc_round(const orc_problem *p, const double *lb, const double *ub, double tol,
  double *newLb, double *newUb, unsigned char *changed, int *nChanged,
  int *infRow, int *infCol, double *minMargin);

/* The row phase of one round only (rules R3-R8): the unsettled Jacobi proposals.  Returns the
 * smallest violated row or -1.  (What one rank computes in the row-sharded multi-GPU mode.) */
int orc_round_rows(const orc_problem *p, const double *lb, const double *ub, double tol,
  double *newLb, double *newUb, unsigned char *changed, double *minMargin);

/* Phase 2: rounds to a fixpoint / maxRounds (<=0: no limit) / time limit.
 * lb/ub are updated in place with every completed feasible round.
 * roundChanged (nullable, roundCap entries) receives per-round counts. */
int orc_propagate(const orc_problem *p, double *lb, double *ub, double tol,
  int maxRounds, double timeLimitSec, orc_result *out, int *roundChanged, int roundCap);

/* Same rules applied Gauss-Seidel style (each row sees the bounds already
 * tightened by earlier rows of the same sweep).  Self-check only. */
int orc_propagate_gs(const orc_problem *p, double *lb, double *ub, double tol,
  int maxRounds, orc_result *out);

/* B independent runs of orc_propagate; lb/ub are B x n, row-major per vector.
 * nThreads > 1 spreads the vectors over OpenMP threads. */
int orc_propagate_batch(const orc_problem *p, int B, double *lb, double *ub, double tol,
  int maxRounds, double timeLimitSec, orc_result *out, int nThreads);

/* Delta form: vector b = base bounds + the changes deltaVar/deltaBound[deltaStart[b] ..
 * deltaStart[b+1]); deltaVar carries the column, bit 31 set = upper bound
 * (the encoding of src/CbcNode.cpp:265-273).  lbOut/ubOut (nullable) are B x n. */
int orc_propagate_batch_delta(const orc_problem *p, const double *baseLb, const double *baseUb,
  int B, const int64_t *deltaStart, const unsigned int *deltaVar, const double *deltaBound,
  double tol, int maxRounds, double timeLimitSec, orc_result *out,
  double *lbOut, double *ubOut, int nThreads);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
