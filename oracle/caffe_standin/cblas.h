/* TEST INFRASTRUCTURE (oracle). Stand-in for <cblas.h>: the enum types caffe/util/math_functions.hpp names in its
 * declarations, and the two routines mkl_alternate.hpp's inline helpers mention (never called by the oracle). */
#pragma once
enum CBLAS_ORDER { CblasRowMajor = 101, CblasColMajor = 102 };
enum CBLAS_TRANSPOSE { CblasNoTrans = 111, CblasTrans = 112, CblasConjTrans = 113 };
void cblas_saxpy(int n, float alpha, const float* x, int incx, float* y, int incy);
void cblas_daxpy(int n, double alpha, const double* x, int incx, double* y, int incy);
void cblas_sscal(int n, float alpha, float* x, int incx);
void cblas_dscal(int n, double alpha, double* x, int incx);
