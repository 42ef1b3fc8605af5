// TEST DOUBLE.  The real OsiSolverInterface / CoinPackedMatrix / CoinTime
// headers (COIN-OR Osi and CoinUtils) are not on this machine; this file
// declares exactly the members /root/reference/src/CbcBoundPropagation.cpp
// touches (:41,51,91-105,139-151,185-186) so that the drop-in and the caller
// snippets compile and run here.  A real Cbc build puts the COIN-OR include
// directories first and never sees this file.
#ifndef OsiSolverInterface_H
#define OsiSolverInterface_H

#include <string>

typedef int CoinBigIndex;

class CoinPackedMatrix {
public:
  virtual ~CoinPackedMatrix() {}
  virtual const double *getElements() const = 0;
  virtual const int *getIndices() const = 0;
  virtual const CoinBigIndex *getVectorStarts() const = 0;
  virtual const int *getVectorLengths() const = 0;
  virtual int getNumRows() const = 0; // major dimension of a row-ordered copy
  virtual CoinBigIndex getNumElements() const = 0;
};

class CoinMessageHandler;

enum OsiDblParam { OsiDualObjectiveLimit = 0, OsiPrimalObjectiveLimit, OsiDualTolerance, OsiPrimalTolerance,
  OsiObjOffset, OsiLastDblParam };

class OsiSolverInterface {
public:
  virtual ~OsiSolverInterface() {}
  virtual int getNumCols() const = 0;
  virtual int getNumRows() const = 0;
  virtual int tightenBoundsFromSingletonRows(int &nFixed) = 0;
  virtual const char *getColType(bool refresh = false) const = 0;
  virtual const CoinPackedMatrix *getMatrixByRow() const = 0;
  virtual const char *getRowSense() const = 0;
  virtual const double *getRightHandSide() const = 0;
  virtual const double *getRowRange() const = 0;
  virtual bool getDblParam(OsiDblParam key, double &value) const = 0;
  virtual double getInfinity() const = 0;
  virtual const double *getColLower() const = 0;
  virtual const double *getColUpper() const = 0;
  virtual void setColLower(int col, double value) = 0;
  virtual void setColUpper(int col, double value) = 0;
  virtual std::string getRowName(int row, unsigned maxLen = static_cast<unsigned>(std::string::npos)) const = 0;
  virtual std::string getColName(int col, unsigned maxLen = static_cast<unsigned>(std::string::npos)) const = 0;
};

double CoinGetTimeOfDay();
double CoinCpuTime();

#endif
