/*
 * cbp_oracle.c — CPU oracle (plain C, single pass per round, sequential sums).
 *
 * TEST INFRASTRUCTURE ONLY; see cbp_oracle.h for what it restates, what it is
 * pinned on (parity unpinned at the CoinBoundPropagation boundary) and who may
 * call it.
 *
 * THE RULES (frozen; DESIGN.md §3 repeats them; the CUDA path implements the
 * same ones):
 *   R1  row bounds from (sense, rhs, range): L (-INF, rhs], G [rhs, INF),
 *       E [rhs, rhs], R [rhs-|range|, rhs], N row skipped.
 *       (src/Cbc_C_Interface.cpp:2219-2235)
 *   R2  a bound x is infinite iff !(|x| < INF)  (INF = solver infinity).
 *   R3  per row, over its nonzeros a != 0 in CSR order:
 *         lo = a>0 ? lb : ub,  hi = a>0 ? ub : lb
 *         lo infinite ? ++infMin : minAct = fma(a, lo, minAct)
 *         hi infinite ? ++infMax : maxAct = fma(a, hi, maxAct)
 *       (src/CbcSolverHeuristics.cpp:282-318: separate integer counters)
 *   R4  row infeasible iff (rowUB finite, infMin==0, minAct > rowUB + 100 tol)
 *       or (rowLB finite, infMax==0, maxAct < rowLB - 100 tol)     (:328-332)
 *   R5  the rowUB side is used iff rowUB finite, infMin <= 1 and NOT
 *       (infMax==0 and maxAct <= rowUB + tol); the rowLB side symmetrically
 *       (:324-326 redundancy test, applied per side).
 *   R6  candidates, slackU = rowUB - minAct, slackL = maxAct - rowLB:
 *         infMin==0: c = lo + slackU/a, evaluated only when the column's
 *                    range is infinite or slackU < |a|((ub-lb) - g), where
 *                    g = 0 for integer columns and 1e-3*max(1, ub-lb) for
 *                    continuous ones (the smallest step R8 would accept)
 *         infMin==1: c = slackU/a, only for the nonzero whose lo is infinite
 *         infMax==0: c = hi - slackL/a (same pre-check with slackL)
 *         infMax==1: c = -slackL/a, only for the nonzero whose hi is infinite
 *       rowUB side: a>0 gives a new upper bound, a<0 a new lower bound;
 *       rowLB side the other way round.                           (:340-477)
 *   R7  integer columns (colType != 0): upper c = floor(c + tol), lower
 *       c = ceil(c - tol); then c += 0.0 (no negative zero); a candidate with
 *       !(|c| < INF) is dropped.
 *   R8  a candidate is accepted iff it improves the ROUND-START bound by more
 *       than thr: 0.5 for integer columns; for continuous ones 1e-3*max(1,
 *       ub-lb) when both round-start bounds are finite, else 1e-3*max(1,|old
 *       bound|); an infinite old bound accepts any finite candidate.
 *       new lb = max of accepted candidates, new ub = min (Jacobi round).
 *   R9  after all rows, for every changed column: lb' > ub' + 100 tol means
 *       infeasible (column); a smaller crossing is snapped: if ub moved,
 *       lb' = ub' = min(lb', old ub), else lb' = ub'.              (:386-392)
 *   R10 reporting: the smallest violating row index of the round, else the
 *       smallest crossing column.  A failing round leaves lb/ub untouched.
 *   R11 before round 1, integer columns get lb = ceil(lb - tol),
 *       ub = floor(ub + tol) and every column is checked as in R9; columns
 *       changed here count among round 1's changed columns.
 *   R12 loop: src/CbcBoundPropagation.cpp:107-195 — time check before a round,
 *       ++rounds, infeasible -> stop, zero changed columns -> fixpoint (that
 *       round counts), else apply and continue; maxRounds exhausted ->
 *       HitMaxRounds.
 */
#include "cbp_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define FEAS_MULT 100.0
#define CONT_REL_THR 1.0e-3
#define INT_THR 0.5

static double now_sec(void)
{
  struct timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return (double)ts.tv_sec + 1.0e-9 * (double)ts.tv_nsec;
}

static inline int is_inf(double x, double INF) { return !(fabs(x) < INF); }

static inline void note_margin(double *mm, double dist, double scale)
{
  if (!mm)
    return;
  double s = fabs(scale) > 1.0 ? fabs(scale) : 1.0;
  double r = fabs(dist) / s;
  if (r < *mm)
    *mm = r;
}

/* R1 */
static inline int row_bounds(const orc_problem *p, int i, double *rl, double *ru)
{
  const double INF = p->inf;
  const double r = p->rhs[i];
  switch (p->sense[i]) {
  case 'L': *rl = -INF; *ru = r; return 1;
  case 'G': *rl = r; *ru = INF; return 1;
  case 'E': *rl = r; *ru = r; return 1;
  case 'R': *rl = r - fabs(p->range ? p->range[i] : 0.0); *ru = r; return 1;
  default: return 0; /* 'N' */
  }
}

static inline int64_t row_end(const orc_problem *p, int i)
{
  return p->rowLen ? p->rowStart[i] + p->rowLen[i] : p->rowStart[i + 1];
}

typedef struct round_ctx {
  const orc_problem *p;
  const double *lb, *ub; /* round-start bounds */
  double *nlb, *nub;     /* tightened bounds   */
  unsigned char *changed;
  double tol;
  double *mm;
} round_ctx;

/* R7 + R8 */
static inline void propose_ub(round_ctx *c, int j, double v)
{
  const double INF = c->p->inf;
  const int isInt = c->p->colType && c->p->colType[j] != 0;
  double rmargin = 1.0;
  if (isInt) {
    double t = v + c->tol;
    rmargin = fabs(t - nearbyint(t));
    v = floor(t);
  }
  v += 0.0;
  if (!(fabs(v) < INF))
    return;
  const double old = c->ub[j], other = c->lb[j];
  if (fabs(old) < INF) {
    const double thr = isInt ? INT_THR
                             : CONT_REL_THR * fmax(1.0, fabs(other) < INF ? old - other : fabs(old));
    note_margin(c->mm, v - (old - thr), v);
    if (isInt && v - 1.0 < old - thr) /* a flipped rounding would matter */
      note_margin(c->mm, rmargin, v);
    if (!(v < old - thr))
      return;
  } else if (!(v < old)) {
    return;
  }
  if (v < c->nub[j])
    c->nub[j] = v;
  c->changed[j] = 1;
}

static inline void propose_lb(round_ctx *c, int j, double v)
{
  const double INF = c->p->inf;
  const int isInt = c->p->colType && c->p->colType[j] != 0;
  double rmargin = 1.0;
  if (isInt) {
    double t = v - c->tol;
    rmargin = fabs(t - nearbyint(t));
    v = ceil(t);
  }
  v += 0.0;
  if (!(fabs(v) < INF))
    return;
  const double old = c->lb[j], other = c->ub[j];
  if (fabs(old) < INF) {
    const double thr = isInt ? INT_THR
                             : CONT_REL_THR * fmax(1.0, fabs(other) < INF ? other - old : fabs(old));
    note_margin(c->mm, v - (old + thr), v);
    if (isInt && v + 1.0 > old + thr)
      note_margin(c->mm, rmargin, v);
    if (!(v > old + thr))
      return;
  } else if (!(v > old)) {
    return;
  }
  if (v > c->nlb[j])
    c->nlb[j] = v;
  c->changed[j] = 1;
}

/* R3 */
static inline void row_sums(const orc_problem *p, int i, const double *lb, const double *ub,
  double *minAct, double *maxAct, int *infMin, int *infMax)
{
  const double INF = p->inf;
  double mn = 0.0, mx = 0.0;
  int im = 0, ix = 0;
  const int64_t e = row_end(p, i);
  for (int64_t k = p->rowStart[i]; k < e; ++k) {
    const double a = p->val[k];
    if (a == 0.0)
      continue;
    const int j = p->col[k];
    const double lo = a > 0.0 ? lb[j] : ub[j];
    const double hi = a > 0.0 ? ub[j] : lb[j];
    if (is_inf(lo, INF))
      ++im;
    else
      mn = fma(a, lo, mn);
    if (is_inf(hi, INF))
      ++ix;
    else
      mx = fma(a, hi, mx);
  }
  *minAct = mn;
  *maxAct = mx;
  *infMin = im;
  *infMax = ix;
}

void orc_row_activity(const orc_problem *p, const double *lb, const double *ub,
  double *minAct, double *maxAct, int *infMin, int *infMax)
{
  for (int i = 0; i < p->m; ++i)
    row_sums(p, i, lb, ub, &minAct[i], &maxAct[i], &infMin[i], &infMax[i]);
}

/* R4-R6 for one row.  Returns 1 if the row is infeasible. */
static inline int process_row(round_ctx *c, int i, const double *lb, const double *ub)
{
  const orc_problem *p = c->p;
  const double INF = p->inf, tol = c->tol;
  double rl, ru;
  if (!row_bounds(p, i, &rl, &ru))
    return 0;
  /* an empty (or all-zero) row is the constraint rl <= 0 <= ru and is tested like any other */
  const int64_t s = p->rowStart[i], e = row_end(p, i);
  double minAct, maxAct;
  int infMin, infMax;
  row_sums(p, i, lb, ub, &minAct, &maxAct, &infMin, &infMax);
  const int ruFin = ru < INF, rlFin = rl > -INF;
  if (ruFin && infMin == 0) {
    note_margin(c->mm, minAct - (ru + FEAS_MULT * tol), ru);
    if (minAct > ru + FEAS_MULT * tol)
      return 1;
  }
  if (rlFin && infMax == 0) {
    note_margin(c->mm, maxAct - (rl - FEAS_MULT * tol), rl);
    if (maxAct < rl - FEAS_MULT * tol)
      return 1;
  }
  const int sideU = ruFin && infMin <= 1 && !(infMax == 0 && maxAct <= ru + tol);
  const int sideL = rlFin && infMax <= 1 && !(infMin == 0 && minAct >= rl - tol);
  if (!sideU && !sideL)
    return 0;
  const double slackU = ru - minAct, slackL = maxAct - rl;
  for (int64_t k = s; k < e; ++k) {
    const double a = p->val[k];
    if (a == 0.0)
      continue;
    const int j = p->col[k];
    const double l = lb[j], u = ub[j];
    const double lo = a > 0.0 ? l : u;
    const double hi = a > 0.0 ? u : l;
    const int loInf = is_inf(lo, INF), hiInf = is_inf(hi, INF);
    const int finite = !loInf && !hiInf;
    const int isInt = p->colType && p->colType[j] != 0;
    const double width = u - l;
    const double w = finite ? fabs(a) * (isInt ? width : width - CONT_REL_THR * fmax(1.0, width)) : 0.0;
    if (sideU) {
      if (infMin == 0) {
        if (!finite || slackU < w) {
          const double v = lo + slackU / a;
          if (a > 0.0)
            propose_ub(c, j, v);
          else
            propose_lb(c, j, v);
        }
      } else if (loInf) {
        const double v = slackU / a;
        if (a > 0.0)
          propose_ub(c, j, v);
        else
          propose_lb(c, j, v);
      }
    }
    if (sideL) {
      if (infMax == 0) {
        if (!finite || slackL < w) {
          const double v = hi - slackL / a;
          if (a > 0.0)
            propose_lb(c, j, v);
          else
            propose_ub(c, j, v);
        }
      } else if (hiInf) {
        const double v = -slackL / a;
        if (a > 0.0)
          propose_lb(c, j, v);
        else
          propose_ub(c, j, v);
      }
    }
  }
  return 0;
}

/* R9 for one changed column; returns 1 if its bounds cross beyond tolerance. */
static inline int settle_column(double *L, double *U, double oldUb, double tol, double *mm)
{
  if (*L > *U) {
    note_margin(mm, *L - (*U + FEAS_MULT * tol), *U);
    if (*L > *U + FEAS_MULT * tol)
      return 1;
    if (*U < oldUb) {
      const double x = *L < oldUb ? *L : oldUb;
      *U = x;
      *L = x;
    } else {
      *L = *U;
    }
  }
  return 0;
}

int orc_round_rows(const orc_problem *p, const double *lb, const double *ub, double tol,
  double *newLb, double *newUb, unsigned char *changed, double *minMargin)
{
  round_ctx c = { p, lb, ub, newLb, newUb, changed, tol, minMargin };
  if (newLb != lb)
    memcpy(newLb, lb, sizeof(double) * (size_t)p->n);
  if (newUb != ub)
    memcpy(newUb, ub, sizeof(double) * (size_t)p->n);
  int badRow = -1;
  for (int i = 0; i < p->m; ++i) {
    if (process_row(&c, i, lb, ub) && badRow < 0)
      badRow = i;
  }
  return badRow;
}

int orc_round(const orc_problem *p, const double *lb, const double *ub, double tol,
  double *newLb, double *newUb, unsigned char *changed, int *nChanged,
  int *infRow, int *infCol, double *minMargin)
{
  const int badRow = orc_round_rows(p, lb, ub, tol, newLb, newUb, changed, minMargin);
  int badCol = -1;
  int cnt = 0;
  for (int j = 0; j < p->n; ++j) {
    if (!changed[j])
      continue;
    ++cnt;
    if (settle_column(&newLb[j], &newUb[j], ub[j], tol, minMargin) && badCol < 0)
      badCol = j;
  }
  *nChanged = cnt;
  if (badRow >= 0) {
    *infRow = badRow;
    *infCol = -1;
    return 1;
  }
  if (badCol >= 0) {
    *infRow = -1;
    *infCol = badCol;
    return 1;
  }
  *infRow = -1;
  *infCol = -1;
  return 0;
}

/* R11.  Returns the smallest crossing column or -1. */
static int prepass(const orc_problem *p, double *lb, double *ub, double tol,
  unsigned char *changed, double *mm)
{
  const double INF = p->inf;
  int bad = -1;
  for (int j = 0; j < p->n; ++j) {
    double l = lb[j], u = ub[j];
    int touched = 0;
    if (p->colType && p->colType[j] != 0) {
      if (fabs(l) < INF) {
        const double t = ceil(l - tol) + 0.0;
        if (t != l) {
          l = t;
          touched = 1;
        }
      }
      if (fabs(u) < INF) {
        const double t = floor(u + tol) + 0.0;
        if (t != u) {
          u = t;
          touched = 1;
        }
      }
    }
    if (l > u) {
      /* the upper bound counts as "moved" only if rounding moved it */
      if (settle_column(&l, &u, ub[j], tol, mm)) {
        if (bad < 0)
          bad = j;
        continue;
      }
      touched = 1;
    }
    if (touched) {
      lb[j] = l;
      ub[j] = u;
      changed[j] = 1;
    }
  }
  return bad;
}

int orc_propagate(const orc_problem *p, double *lb, double *ub, double tol,
  int maxRounds, double timeLimitSec, orc_result *out, int *roundChanged, int roundCap)
{
  const double t0 = now_sec();
  const int n = p->n;
  orc_result r;
  memset(&r, 0, sizeof(r));
  r.feasible = 1;
  r.infeasibleRow = -1;
  r.infeasibleCol = -1;
  r.minMargin = 1.0e300;
  double *nlb = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  double *nub = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  unsigned char *changed = (unsigned char *)calloc((size_t)(n > 0 ? n : 1), 1);
  if (!nlb || !nub || !changed) {
    free(nlb);
    free(nub);
    free(changed);
    return -1;
  }
  const int limit = maxRounds > 0 ? maxRounds : 0x7fffffff;
  const int bad0 = prepass(p, lb, ub, tol, changed, &r.minMargin);
  if (bad0 >= 0) {
    r.rounds = 1;
    r.feasible = 0;
    r.infeasibleCol = bad0;
    r.stopReason = ORC_INFEASIBLE;
  }
  for (int round = 0; r.stopReason == ORC_NOT_RUN && round < limit; ++round) {
    if (now_sec() - t0 >= timeLimitSec) {
      r.stopReason = ORC_TIME_LIMIT;
      break;
    }
    int cnt = 0, ir = -1, ic = -1;
    const int bad = orc_round(p, lb, ub, tol, nlb, nub, changed, &cnt, &ir, &ic, &r.minMargin);
    ++r.rounds;
    if (bad) {
      r.feasible = 0;
      r.infeasibleRow = ir;
      r.infeasibleCol = ic;
      r.stopReason = ORC_INFEASIBLE;
      break;
    }
    if (roundChanged && r.rounds <= roundCap)
      roundChanged[r.rounds - 1] = cnt;
    if (cnt == 0) {
      r.stopReason = ORC_FIXPOINT;
      break;
    }
    memcpy(lb, nlb, sizeof(double) * (size_t)n);
    memcpy(ub, nub, sizeof(double) * (size_t)n);
    memset(changed, 0, (size_t)n);
    r.nChanged += cnt;
  }
  if (r.stopReason == ORC_NOT_RUN)
    r.stopReason = ORC_MAX_ROUNDS;
  r.seconds = now_sec() - t0;
  free(nlb);
  free(nub);
  free(changed);
  if (out)
    *out = r;
  return 0;
}

int orc_propagate_gs(const orc_problem *p, double *lb, double *ub, double tol,
  int maxRounds, orc_result *out)
{
  const double t0 = now_sec();
  const int n = p->n;
  orc_result r;
  memset(&r, 0, sizeof(r));
  r.feasible = 1;
  r.infeasibleRow = -1;
  r.infeasibleCol = -1;
  r.minMargin = 1.0e300;
  unsigned char *changed = (unsigned char *)calloc((size_t)(n > 0 ? n : 1), 1);
  double *sl = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  double *su = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  if (!changed || !sl || !su) {
    free(changed);
    free(sl);
    free(su);
    return -1;
  }
  const int limit = maxRounds > 0 ? maxRounds : 0x7fffffff;
  const int bad0 = prepass(p, lb, ub, tol, changed, NULL);
  if (bad0 >= 0) {
    r.rounds = 1;
    r.feasible = 0;
    r.infeasibleCol = bad0;
    r.stopReason = ORC_INFEASIBLE;
  }
  for (int round = 0; r.stopReason == ORC_NOT_RUN && round < limit; ++round) {
    /* in place: proposals are compared with, and written to, the live bounds */
    round_ctx c = { p, lb, ub, lb, ub, changed, tol, NULL };
    memcpy(sl, lb, sizeof(double) * (size_t)n);
    memcpy(su, ub, sizeof(double) * (size_t)n);
    int badRow = -1, badCol = -1;
    for (int i = 0; i < p->m && badRow < 0; ++i) {
      if (process_row(&c, i, lb, ub))
        badRow = i;
    }
    int cnt = 0;
    for (int j = 0; j < n; ++j) {
      if (!changed[j])
        continue;
      ++cnt;
      if (settle_column(&lb[j], &ub[j], su[j], tol, NULL) && badCol < 0)
        badCol = j;
      changed[j] = 0;
    }
    ++r.rounds;
    if (badRow >= 0 || badCol >= 0) {
      r.feasible = 0;
      r.infeasibleRow = badRow;
      r.infeasibleCol = badRow >= 0 ? -1 : badCol;
      r.stopReason = ORC_INFEASIBLE;
      break;
    }
    if (cnt == 0) {
      r.stopReason = ORC_FIXPOINT;
      break;
    }
    r.nChanged += cnt;
  }
  if (r.stopReason == ORC_NOT_RUN)
    r.stopReason = ORC_MAX_ROUNDS;
  r.seconds = now_sec() - t0;
  free(changed);
  free(sl);
  free(su);
  if (out)
    *out = r;
  return 0;
}

int orc_singletons(const orc_problem *p, double *lb, double *ub, double tol, int *nFixed)
{
  const int n = p->n;
  const double INF = p->inf;
  double *nlb = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  double *nub = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
  unsigned char *changed = (unsigned char *)calloc((size_t)(n > 0 ? n : 1), 1);
  if (!nlb || !nub || !changed) {
    free(nlb);
    free(nub);
    free(changed);
    return -2;
  }
  memcpy(nlb, lb, sizeof(double) * (size_t)n);
  memcpy(nub, ub, sizeof(double) * (size_t)n);
  round_ctx c = { p, lb, ub, nlb, nub, changed, tol, NULL };
  for (int i = 0; i < p->m; ++i) {
    double rl, ru;
    if (!row_bounds(p, i, &rl, &ru))
      continue;
    const int64_t e = row_end(p, i);
    int cnt = 0;
    int64_t at = -1;
    for (int64_t k = p->rowStart[i]; k < e; ++k) {
      if (p->val[k] != 0.0) {
        ++cnt;
        at = k;
      }
    }
    if (cnt != 1)
      continue;
    const double a = p->val[at];
    const int j = p->col[at];
    if (ru < INF) {
      if (a > 0.0)
        propose_ub(&c, j, ru / a);
      else
        propose_lb(&c, j, ru / a);
    }
    if (rl > -INF) {
      if (a > 0.0)
        propose_lb(&c, j, rl / a);
      else
        propose_ub(&c, j, rl / a);
    }
  }
  int nT = 0, nF = 0, bad = 0;
  for (int j = 0; j < n; ++j) {
    if (!changed[j])
      continue;
    if (settle_column(&nlb[j], &nub[j], ub[j], tol, NULL)) {
      bad = 1;
      break;
    }
    ++nT;
    if (nlb[j] == nub[j])
      ++nF;
  }
  if (!bad) {
    memcpy(lb, nlb, sizeof(double) * (size_t)n);
    memcpy(ub, nub, sizeof(double) * (size_t)n);
  }
  free(nlb);
  free(nub);
  free(changed);
  if (nFixed)
    *nFixed = bad ? 0 : nF;
  return bad ? -1 : nT;
}

int orc_max_threads(void)
{
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

int orc_propagate_batch(const orc_problem *p, int B, double *lb, double *ub, double tol,
  int maxRounds, double timeLimitSec, orc_result *out, int nThreads)
{
  int rc = 0;
  const size_t n = (size_t)p->n;
#ifdef _OPENMP
  if (nThreads < 1)
    nThreads = 1;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nThreads)
#endif
  for (int b = 0; b < B; ++b) {
    if (orc_propagate(p, lb + n * (size_t)b, ub + n * (size_t)b, tol, maxRounds, timeLimitSec,
          &out[b], NULL, 0)
      != 0)
      rc = -1;
  }
  (void)nThreads;
  return rc;
}

int orc_propagate_batch_delta(const orc_problem *p, const double *baseLb, const double *baseUb,
  int B, const int64_t *deltaStart, const unsigned int *deltaVar, const double *deltaBound,
  double tol, int maxRounds, double timeLimitSec, orc_result *out,
  double *lbOut, double *ubOut, int nThreads)
{
  int rc = 0;
  const size_t n = (size_t)p->n;
#ifdef _OPENMP
  if (nThreads < 1)
    nThreads = 1;
#pragma omp parallel num_threads(nThreads)
#endif
  {
    double *l = (double *)malloc(sizeof(double) * (n ? n : 1));
    double *u = (double *)malloc(sizeof(double) * (n ? n : 1));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 1)
#endif
    for (int b = 0; b < B; ++b) {
      if (!l || !u) {
        rc = -1;
        continue;
      }
      memcpy(l, baseLb, sizeof(double) * n);
      memcpy(u, baseUb, sizeof(double) * n);
      for (int64_t k = deltaStart[b]; k < deltaStart[b + 1]; ++k) {
        /* CbcPartialNodeInfo's encoding (src/CbcPartialNodeInfo.cpp:139,196,208): bit 31 = upper
         * bound, bit 30 = "can go odd way" flag, bits 0-29 = column */
        const unsigned int v = deltaVar[k];
        const unsigned int j = v & 0x3fffffffu;
        if (j >= (unsigned int)p->n) {
          rc = -2;
          continue;
        }
        if (v & 0x80000000u)
          u[j] = deltaBound[k];
        else
          l[j] = deltaBound[k];
      }
      if (orc_propagate(p, l, u, tol, maxRounds, timeLimitSec, &out[b], NULL, 0) != 0)
        rc = -1;
      if (lbOut)
        memcpy(lbOut + n * (size_t)b, l, sizeof(double) * n);
      if (ubOut)
        memcpy(ubOut + n * (size_t)b, u, sizeof(double) * n);
    }
    free(l);
    free(u);
  }
  (void)nThreads;
  return rc;
}
