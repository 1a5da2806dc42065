/* linpack_port.c -- CPU restatement of the two LINPACK routines the reference
 * reaches through R (src/math/linear.algebra.c:197-204): dqrdc2 (Householder
 * QR with R's limited column pivoting, tolerance sqrt(eps)) and dqrsl (apply
 * the factorisation). They belong to R (src/appl/dqrdc2.f, dqrsl.f), not to the
 * reference tree, and are only reached for Gaussian nodes with >= 3 parents
 * (src/math/least.squares.c:339-352). TEST INFRASTRUCTURE ONLY.
 * PARITY UNPINNED: written from the published LINPACK algorithm; checked in
 * tests against numpy.linalg.lstsq. Arrays are column-major, 1-based in the
 * comments, 0-based in the code. */
#include "shim/minir.h"

#define X(i, j) x[(size_t)(i) + (size_t)(j) * (size_t)ld]

static double nrm2(int n, const double *v) {
  /* scaled 2-norm, as the BLAS does, to avoid overflow. */
  double scale = 0.0, ssq = 1.0;
  for (int i = 0; i < n; i++) {
    if (v[i] != 0.0) {
      double a = fabs(v[i]);
      if (scale < a) { ssq = 1.0 + ssq * (scale / a) * (scale / a); scale = a; }
      else ssq += (a / scale) * (a / scale);
    }
  }
  return scale * sqrt(ssq);
}

static double dot(int n, const double *a, const double *b) {
  double s = 0.0;
  for (int i = 0; i < n; i++) s += a[i] * b[i];
  return s;
}

static void axpy(int n, double t, const double *a, double *b) {
  for (int i = 0; i < n; i++) b[i] += t * a[i];
}

void minir_f77_dqrdc2(double *x, int *ldx, int *pn, int *pp, double *tol,
    int *rank, double *qraux, int *jpvt, double *work) {
  int ld = *ldx, n = *pn, p = *pp;
  double *work1 = work, *work2 = work + p;

  for (int j = 0; j < p; j++) {
    qraux[j] = nrm2(n, &X(0, j));
    work1[j] = qraux[j];
    work2[j] = qraux[j] == 0.0 ? 1.0 : qraux[j];
  }

  int lup = n < p ? n : p;
  int k = p + 1;                       /* 1-based: columns >= k are "dropped". */

  for (int l = 1; l <= lup; l++) {
    /* move negligible columns to the far right, keeping the others in order. */
    while (l < k && qraux[l - 1] < work2[l - 1] * (*tol)) {
      for (int i = 0; i < n; i++) {
        double t = X(i, l - 1);
        for (int j = l; j < p; j++) X(i, j - 1) = X(i, j);
        X(i, p - 1) = t;
      }
      int ip = jpvt[l - 1]; double tq = qraux[l - 1], t1 = work1[l - 1], t2 = work2[l - 1];
      for (int j = l; j < p; j++) {
        jpvt[j - 1] = jpvt[j]; qraux[j - 1] = qraux[j];
        work1[j - 1] = work1[j]; work2[j - 1] = work2[j];
      }
      jpvt[p - 1] = ip; qraux[p - 1] = tq; work1[p - 1] = t1; work2[p - 1] = t2;
      k--;
    }
    if (l == n) break;

    /* Householder reflection for column l. */
    int len = n - l + 1;
    double *col = &X(l - 1, l - 1);
    double nrmxl = nrm2(len, col);
    if (nrmxl == 0.0) continue;
    if (col[0] != 0.0) nrmxl = copysign(nrmxl, col[0]);
    for (int i = 0; i < len; i++) col[i] /= nrmxl;
    col[0] += 1.0;

    for (int j = l + 1; j <= p; j++) {
      double *cj = &X(l - 1, j - 1);
      double t = -dot(len, col, cj) / col[0];
      axpy(len, t, col, cj);
      if (qraux[j - 1] == 0.0) continue;
      double r = fabs(cj[0]) / qraux[j - 1];
      double tt = 1.0 - r * r;
      if (tt < 0.0) tt = 0.0;
      if (fabs(tt) < 1e-6) {
        qraux[j - 1] = nrm2(n - l, &X(l, j - 1));
        work1[j - 1] = qraux[j - 1];
      } else {
        qraux[j - 1] *= sqrt(tt);
      }
    }
    qraux[l - 1] = col[0];
    col[0] = -nrmxl;
  }
  k = k - 1;
  *rank = k < n ? k : n;
}

void minir_f77_dqrsl(double *x, int *ldx, int *pn, int *pk, double *qraux,
    double *y, double *qy, double *qty, double *b, double *rsd, double *xb,
    int *job, int *info) {
  int ld = *ldx, n = *pn, k = *pk;
  *info = 0;
  int cqy = (*job / 10000) != 0;
  int cqty = (*job % 10000) != 0;
  int cb = ((*job % 1000) / 100) != 0;
  int cr = ((*job % 100) / 10) != 0;
  int cxb = (*job % 10) != 0;
  int ju = k < n - 1 ? k : n - 1;

  if (ju == 0) {
    if (cqy) qy[0] = y[0];
    if (cqty) qty[0] = y[0];
    if (cxb) xb[0] = y[0];
    if (cb) { if (X(0, 0) != 0.0) b[0] = y[0] / X(0, 0); else *info = 1; }
    if (cr) rsd[0] = 0.0;
    return;
  }

  if (cqy && qy != y) memmove(qy, y, (size_t)n * sizeof(double));
  if (cqty && qty != y) memmove(qty, y, (size_t)n * sizeof(double));

  if (cqy) {
    for (int jj = 1; jj <= ju; jj++) {
      int j = ju - jj + 1;
      if (qraux[j - 1] == 0.0) continue;
      double temp = X(j - 1, j - 1); X(j - 1, j - 1) = qraux[j - 1];
      double t = -dot(n - j + 1, &X(j - 1, j - 1), &qy[j - 1]) / X(j - 1, j - 1);
      axpy(n - j + 1, t, &X(j - 1, j - 1), &qy[j - 1]);
      X(j - 1, j - 1) = temp;
    }
  }
  if (cqty) {
    for (int j = 1; j <= ju; j++) {
      if (qraux[j - 1] == 0.0) continue;
      double temp = X(j - 1, j - 1); X(j - 1, j - 1) = qraux[j - 1];
      double t = -dot(n - j + 1, &X(j - 1, j - 1), &qty[j - 1]) / X(j - 1, j - 1);
      axpy(n - j + 1, t, &X(j - 1, j - 1), &qty[j - 1]);
      X(j - 1, j - 1) = temp;
    }
  }

  /* set up b, rsd and xb from trans(q)*y. */
  if (cb && b != qty) memmove(b, qty, (size_t)k * sizeof(double));
  if (cxb && xb != qty) memmove(xb, qty, (size_t)k * sizeof(double));
  if (cr && k < n && rsd != qty)
    memmove(&rsd[k], &qty[k], (size_t)(n - k) * sizeof(double));
  if (cxb) for (int i = k; i < n; i++) xb[i] = 0.0;
  if (cr) for (int i = 0; i < k; i++) rsd[i] = 0.0;

  if (cb) {
    for (int jj = 1; jj <= k; jj++) {
      int j = k - jj + 1;
      if (X(j - 1, j - 1) == 0.0) { *info = j; break; }
      b[j - 1] /= X(j - 1, j - 1);
      if (j != 1) axpy(j - 1, -b[j - 1], &X(0, j - 1), b);
    }
  }

  if (cr || cxb) {
    for (int jj = 1; jj <= ju; jj++) {
      int j = ju - jj + 1;
      if (qraux[j - 1] == 0.0) continue;
      double temp = X(j - 1, j - 1); X(j - 1, j - 1) = qraux[j - 1];
      if (cr) {
        double t = -dot(n - j + 1, &X(j - 1, j - 1), &rsd[j - 1]) / X(j - 1, j - 1);
        axpy(n - j + 1, t, &X(j - 1, j - 1), &rsd[j - 1]);
      }
      if (cxb) {
        double t = -dot(n - j + 1, &X(j - 1, j - 1), &xb[j - 1]) / X(j - 1, j - 1);
        axpy(n - j + 1, t, &X(j - 1, j - 1), &xb[j - 1]);
      }
      X(j - 1, j - 1) = temp;
    }
  }
}
