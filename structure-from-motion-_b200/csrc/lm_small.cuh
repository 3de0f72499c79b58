// lm_small.cuh -- dense Levenberg-Marquardt for tiny problems (N = 3 or 6 parameters): a
// restatement of MINPACK's lmder/lmpar control flow on the NORMAL EQUATIONS.
//
// The reference refines points and poses with scipy.optimize.least_squares(method='lm')
// (NonLinearTriangulation.py:116-122, NonlinearPnP.py:97-104), i.e. MINPACK lmder with
// mode=1 scaling (scipy >= 1.16 'jac' x_scale), factor=100, ftol=xtol=gtol=1e-8 and a
// forward-difference Jacobian. With its loose 1e-8 tolerances the reference stops measurably
// short of the minimum when the residual is large (PnP with gross outliers: ~1e-5 relative in
// the pose), so matching it to 1e-6 needs the same trust-region schedule and stopping rules,
// not just the same minimum. Every quantity lmder/lmpar derive from the pivoted QR of J is a
// function of A = J^T J and g = J^T r alone:
//     R^T R = A,  qtf = R^-T g,  |J p|^2 = p^T A p,
//     qrsolv(par) solves (A + par D^2) x = g,  |S^-T w|^2 = w^T (A + par D^2)^-1 w,
// so the same iteration runs on 3x3 / 6x6 Cholesky factors in registers (the Jacobian here is
// analytic; the reference's differs from it by its ~1e-8 differencing error).
// Passing tol << 1e-8 gives the "converged" solution instead of the reference's stopping point.
#pragma once
#include "sfm_common.cuh"

namespace sfm {

enum LMStatus : int {
    LM_RUNNING = 0,
    LM_FTOL = 1,      // MINPACK info values
    LM_XTOL = 2,
    LM_FXTOL = 3,
    LM_GTOL = 4,
    LM_MAXITER = 5,   // nfev reached maxfev
    LM_FTOL_EPS = 6,
    LM_XTOL_EPS = 7,
    LM_GTOL_EPS = 8,
    LM_FAILED = -1,   // non-finite residuals at the start (scipy raises -> reference keeps its input)
};

SFM_HD bool is_finite(double v) { return fabs(v) <= 1.7976931348623157e308; }  // false for NaN/inf

// Packed symmetric index (lower triangle, row-major).
SFM_HD constexpr int sym_idx(int r, int c) { return r >= c ? r * (r + 1) / 2 + c : c * (c + 1) / 2 + r; }

template <int N>
struct NormalEq {
    double A[N * (N + 1) / 2];   // J^T J, packed lower
    double g[N];                 // J^T r
    double cost;                 // 0.5 |r|^2
};

// Cholesky factor of (A + par D^2) into L (packed lower). false if not positive definite.
template <int N>
SFM_HD bool chol_factor(const double *A, const double (&D)[N], double par, double *L) {
#pragma unroll
    for (int i = 0; i < N * (N + 1) / 2; ++i) L[i] = A[i];
#pragma unroll
    for (int i = 0; i < N; ++i) L[sym_idx(i, i)] += par * D[i] * D[i];
#pragma unroll
    for (int j = 0; j < N; ++j) {
        double d = L[sym_idx(j, j)];
#pragma unroll
        for (int k = 0; k < j; ++k) d -= L[sym_idx(j, k)] * L[sym_idx(j, k)];
        if (!(d > 0.0) || !is_finite(d)) return false;
        d = sqrt(d);
        L[sym_idx(j, j)] = d;
        double inv = 1.0 / d;
#pragma unroll
        for (int i = j + 1; i < N; ++i) {
            double s = L[sym_idx(i, j)];
#pragma unroll
            for (int k = 0; k < j; ++k) s -= L[sym_idx(i, k)] * L[sym_idx(j, k)];
            L[sym_idx(i, j)] = s * inv;
        }
    }
    return true;
}

// y = L^-1 b
template <int N>
SFM_HD void chol_forward(const double *L, const double (&b)[N], double (&y)[N]) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
        double s = b[i];
#pragma unroll
        for (int k = 0; k < i; ++k) s -= L[sym_idx(i, k)] * y[k];
        y[i] = s / L[sym_idx(i, i)];
    }
}

// x = L^-T y
template <int N>
SFM_HD void chol_backward(const double *L, const double (&y)[N], double (&x)[N]) {
#pragma unroll
    for (int i = N - 1; i >= 0; --i) {
        double s = y[i];
#pragma unroll
        for (int k = i + 1; k < N; ++k) s -= L[sym_idx(k, i)] * x[k];
        x[i] = s / L[sym_idx(i, i)];
    }
}

template <int N>
struct LMState {
    double x[N];
    double D[N];          // MINPACK diag
    double delta, par, xnorm, pnorm;
    int iter;             // MINPACK outer iteration counter (starts at 1)
    int nfev;             // residual evaluations (1 = the initial one)
    int need_outer;       // the next propose must run the per-linearisation prologue first
    int status;
};

template <int N>
SFM_HD void lm_init(LMState<N> &s, const NormalEq<N> &ne) {
    s.par = 0.0;
    s.iter = 1;
    s.nfev = 1;
    s.need_outer = 1;
    s.delta = 0.0;
    s.xnorm = 0.0;
    s.pnorm = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) s.D[j] = 1.0;
    s.status = is_finite(ne.cost) ? LM_RUNNING : LM_FAILED;
}

// lmder, top of the outer loop: scaling on the first iteration, gtol test, diag update.
template <int N>
SFM_HD void lm_begin_outer(LMState<N> &s, const NormalEq<N> &ne, double gtol) {
    double acnorm[N];
#pragma unroll
    for (int j = 0; j < N; ++j) acnorm[j] = sqrt(ne.A[sym_idx(j, j)]);
    if (s.iter == 1) {
        double xn = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) {
            s.D[j] = acnorm[j] == 0.0 ? 1.0 : acnorm[j];
            double t = s.D[j] * s.x[j];
            xn += t * t;
        }
        s.xnorm = sqrt(xn);
        s.delta = 100.0 * s.xnorm;
        if (s.delta == 0.0) s.delta = 100.0;
    }
    double fnorm = sqrt(2.0 * ne.cost);
    double gnorm = 0.0;
    if (fnorm != 0.0) {
#pragma unroll
        for (int j = 0; j < N; ++j)
            if (acnorm[j] != 0.0) gnorm = fmax(gnorm, fabs(ne.g[j] / fnorm) / acnorm[j]);
    }
    if (gnorm <= gtol) s.status = LM_GTOL;
#pragma unroll
    for (int j = 0; j < N; ++j) s.D[j] = fmax(s.D[j], acnorm[j]);
    s.need_outer = 0;
}

// lmpar on the normal equations: x solves (A + par D^2) x = g with | |D x| - delta | <= 0.1 delta
// (or par = 0 when the Gauss-Newton step is inside the region).
template <int N>
SFM_HD void lm_par(const NormalEq<N> &ne, const double (&D)[N], double delta, double &par,
                   double (&x)[N]) {
    const double dwarf = 2.2250738585072014e-308;
    double L[N * (N + 1) / 2], y[N], w[N];
    bool full_rank = chol_factor<N>(ne.A, D, 0.0, L);
    double dxnorm = 0.0, fp = 1.7976931348623157e308;
    if (full_rank) {
        chol_forward<N>(L, ne.g, y);
        chol_backward<N>(L, y, x);
#pragma unroll
        for (int j = 0; j < N; ++j) dxnorm += (D[j] * x[j]) * (D[j] * x[j]);
        dxnorm = sqrt(dxnorm);
        fp = dxnorm - delta;
        if (fp <= 0.1 * delta) { par = 0.0; return; }
    }
    double parl = 0.0;
    if (full_rank) {
#pragma unroll
        for (int j = 0; j < N; ++j) w[j] = D[j] * (D[j] * x[j] / dxnorm);
        chol_forward<N>(L, w, y);
        double t2 = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) t2 += y[j] * y[j];
        parl = fp / delta / t2;
    }
    double gnorm = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) gnorm += (ne.g[j] / D[j]) * (ne.g[j] / D[j]);
    gnorm = sqrt(gnorm);
    double paru = gnorm / delta;
    if (paru == 0.0) paru = dwarf / fmin(delta, 0.1);
    par = fmax(par, parl);
    par = fmin(par, paru);
    if (par == 0.0) par = full_rank ? gnorm / dxnorm : paru * 0.001;
    for (int it = 1;; ++it) {
        if (par == 0.0) par = fmax(dwarf, 0.001 * paru);
        if (!chol_factor<N>(ne.A, D, par, L)) {
            // cannot happen for par > 0 in exact arithmetic; keep going with more damping
            par = fmax(par * 10.0, 1e-300);
            if (it >= 40) {
#pragma unroll
                for (int j = 0; j < N; ++j) x[j] = 0.0;
                return;
            }
            continue;
        }
        chol_forward<N>(L, ne.g, y);
        chol_backward<N>(L, y, x);
        dxnorm = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) dxnorm += (D[j] * x[j]) * (D[j] * x[j]);
        dxnorm = sqrt(dxnorm);
        double temp = fp;
        fp = dxnorm - delta;
        if (fabs(fp) <= 0.1 * delta || (parl == 0.0 && fp <= temp && temp < 0.0) || it >= 10) break;
#pragma unroll
        for (int j = 0; j < N; ++j) w[j] = D[j] * (D[j] * x[j] / dxnorm);
        chol_forward<N>(L, w, y);
        double t2 = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) t2 += y[j] * y[j];
        double parc = fp / delta / t2;
        if (fp > 0.0) parl = fmax(parl, par);
        if (fp < 0.0) paru = fmin(paru, par);
        par = fmax(parl, par + parc);
    }
}

// Propose the next trial point (lmder: lmpar, p = -x, first-iteration delta clamp).
template <int N>
SFM_HD void lm_propose(LMState<N> &s, const NormalEq<N> &ne, double (&p)[N], double (&xtrial)[N]) {
    double step[N];
    lm_par<N>(ne, s.D, s.delta, s.par, step);
    double pn = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) {
        p[j] = -step[j];
        xtrial[j] = s.x[j] + p[j];
        pn += (s.D[j] * p[j]) * (s.D[j] * p[j]);
    }
    s.pnorm = sqrt(pn);
    if (s.iter == 1) s.delta = fmin(s.delta, s.pnorm);
}

// Judge the trial (lmder inner loop after the function evaluation). `net` = normal equations at
// xtrial. Returns true if the step was accepted (ne/x updated, next propose starts an outer pass).
template <int N>
SFM_HD bool lm_judge(LMState<N> &s, NormalEq<N> &ne, const NormalEq<N> &net, const double (&p)[N],
                     const double (&xtrial)[N], double ftol, double xtol, int maxfev) {
    const double epsmch = 2.220446049250313e-16;
    s.nfev += 1;
    double fnorm = sqrt(2.0 * ne.cost);
    double fnorm1 = is_finite(net.cost) ? sqrt(2.0 * net.cost) : 1.7976931348623157e308;
    double actred = -1.0;
    if (0.1 * fnorm1 < fnorm) { double t = fnorm1 / fnorm; actred = 1.0 - t * t; }
    // |J p|^2 = p^T A p
    double pAp = 0.0;
#pragma unroll
    for (int r = 0; r < N; ++r) {
        double row = 0.0;
#pragma unroll
        for (int c = 0; c < N; ++c) row += ne.A[sym_idx(r, c)] * p[c];
        pAp += p[r] * row;
    }
    double temp1 = sqrt(fmax(pAp, 0.0)) / fnorm;
    double temp2 = sqrt(s.par) * s.pnorm / fnorm;
    double prered = temp1 * temp1 + temp2 * temp2 / 0.5;
    double dirder = -(temp1 * temp1 + temp2 * temp2);
    double ratio = prered != 0.0 ? actred / prered : 0.0;
    if (!(fnorm > 0.0)) { prered = 0.0; dirder = 0.0; ratio = 0.0; }
    if (ratio <= 0.25) {
        double temp = actred >= 0.0 ? 0.5 : 0.5 * dirder / (dirder + 0.5 * actred);
        if (0.1 * fnorm1 >= fnorm || temp < 0.1) temp = 0.1;
        s.delta = temp * fmin(s.delta, s.pnorm / 0.1);
        s.par = s.par / temp;
    } else if (s.par == 0.0 || ratio >= 0.75) {
        s.delta = s.pnorm / 0.5;
        s.par = 0.5 * s.par;
    }
    bool accept = ratio >= 1e-4;
    if (accept) {
        ne = net;
        double xn = 0.0;
#pragma unroll
        for (int j = 0; j < N; ++j) {
            s.x[j] = xtrial[j];
            double t = s.D[j] * s.x[j];
            xn += t * t;
        }
        s.xnorm = sqrt(xn);
        s.iter += 1;
        s.need_outer = 1;
    }
    int info = 0;
    if (fabs(actred) <= ftol && prered <= ftol && 0.5 * ratio <= 1.0) info = 1;
    if (s.delta <= xtol * s.xnorm) info = 2;
    if (fabs(actred) <= ftol && prered <= ftol && 0.5 * ratio <= 1.0 && info == 2) info = 3;
    if (info == 0) {
        if (s.nfev >= maxfev) info = 5;
        if (fabs(actred) <= epsmch && prered <= epsmch && 0.5 * ratio <= 1.0) info = 6;
        if (s.delta <= epsmch * s.xnorm) info = 7;
    }
    if (info != 0) s.status = info;
    return accept;
}

}  // namespace sfm
