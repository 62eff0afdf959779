"""TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

numpy fp64 restatement of what ``LogisticRegression(solver="lbfgs").fit`` does
for one (candidate, fold) fit -- the arithmetic the reference worker lands in at
aws-prod/worker/worker.py:315 (hold-out fit) and :326 (cross_val_score).

Third-party algorithm restated here (source is NOT in /root/reference):
  * scikit-learn 1.9.0  sklearn/linear_model/_logistic.py:438-461,580-598
    (objective, w0 = 0, l2 = 1/(C*n_train), L-BFGS-B options),
    sklearn/linear_model/_linear_loss.py:291-379 (loss + gradient),
    sklearn/_loss/_loss.pyx.tp:686-751 (half binomial), :1300-1335 (half
    multinomial).
  * scipy 1.18.1 L-BFGS-B (Byrd, Lu, Nocedal, Zhu 1995/1997/2011), unbounded
    case, driven by scipy/optimize/_lbfgsb_py.py:393-461; line search is
    MINPACK-2 dcsrch/dcstep (More' & Thuente 1994), readable twin at
    scipy/optimize/_dcsrch.py.

Rule set: m = 10 pairs, H0 = I/theta with
theta = y'y / s'y, first trial step min(1/||d||_2, 1e10) then 1.0, dcsrch
(ftol 1e-3, gtol 0.9, xtol 0.1, stpmin 0, stpmax 1e10), stop when
||g||_inf <= tol or (f_old - f) <= 64*eps*max(|f_old|,|f|,1), iteration cap
checked before the convergence tests, pair skipped when
s'y <= eps * (-g_old'd * stp), line-search failure => memory reset.
"""
from __future__ import annotations

import numpy as np

EPS = float(np.finfo(np.float64).eps)
FTOL_LS, GTOL_LS, XTOL_LS = 1e-3, 0.9, 0.1
STPMAX = 1e10
MAXCOR = 10
FACTR = 64.0  # sklearn passes ftol = 64*eps  ->  factr = 64


# --------------------------------------------------------------------------
# MINPACK-2 dcstep / dcsrch (More'-Thuente), as used by L-BFGS-B's lnsrlb
# --------------------------------------------------------------------------
def _dcstep(stx, fx, dx, sty, fy, dy, stp, fp, dp, brackt, stpmin, stpmax):
    sgnd = dp * (dx / abs(dx))
    if fp > fx:  # case 1: higher function value -> minimum bracketed
        theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp
        s = max(abs(theta), abs(dx), abs(dp))
        gamma = s * np.sqrt((theta / s) ** 2 - (dx / s) * (dp / s))
        if stp < stx:
            gamma = -gamma
        p = (gamma - dx) + theta
        q = ((gamma - dx) + gamma) + dp
        r = p / q
        stpc = stx + r * (stp - stx)
        stpq = stx + ((dx / ((fx - fp) / (stp - stx) + dx)) / 2.0) * (stp - stx)
        if abs(stpc - stx) < abs(stpq - stx):
            stpf = stpc
        else:
            stpf = stpc + (stpq - stpc) / 2.0
        brackt = True
    elif sgnd < 0.0:  # case 2: derivatives of opposite sign
        theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp
        s = max(abs(theta), abs(dx), abs(dp))
        gamma = s * np.sqrt((theta / s) ** 2 - (dx / s) * (dp / s))
        if stp > stx:
            gamma = -gamma
        p = (gamma - dp) + theta
        q = ((gamma - dp) + gamma) + dx
        r = p / q
        stpc = stp + r * (stx - stp)
        stpq = stp + (dp / (dp - dx)) * (stx - stp)
        stpf = stpc if abs(stpc - stp) > abs(stpq - stp) else stpq
        brackt = True
    elif abs(dp) < abs(dx):  # case 3: derivative magnitude decreases
        theta = 3.0 * (fx - fp) / (stp - stx) + dx + dp
        s = max(abs(theta), abs(dx), abs(dp))
        gamma = s * np.sqrt(max(0.0, (theta / s) ** 2 - (dx / s) * (dp / s)))
        if stp > stx:
            gamma = -gamma
        p = (gamma - dp) + theta
        q = (gamma + (dx - dp)) + gamma
        r = p / q
        if r < 0.0 and gamma != 0.0:
            stpc = stp + r * (stx - stp)
        elif stp > stx:
            stpc = stpmax
        else:
            stpc = stpmin
        stpq = stp + (dp / (dp - dx)) * (stx - stp)
        if brackt:
            stpf = stpc if abs(stpc - stp) < abs(stpq - stp) else stpq
            if stp > stx:
                stpf = min(stp + 0.66 * (sty - stp), stpf)
            else:
                stpf = max(stp + 0.66 * (sty - stp), stpf)
        else:
            stpf = stpc if abs(stpc - stp) > abs(stpq - stp) else stpq
            stpf = min(stpmax, stpf)
            stpf = max(stpmin, stpf)
    else:  # case 4
        if brackt:
            theta = 3.0 * (fp - fy) / (sty - stp) + dy + dp
            s = max(abs(theta), abs(dy), abs(dp))
            gamma = s * np.sqrt((theta / s) ** 2 - (dy / s) * (dp / s))
            if stp > sty:
                gamma = -gamma
            p = (gamma - dp) + theta
            q = ((gamma - dp) + gamma) + dy
            r = p / q
            stpf = stp + r * (sty - stp)
        elif stp > stx:
            stpf = stpmax
        else:
            stpf = stpmin
    if fp > fx:
        sty, fy, dy = stp, fp, dp
    else:
        if sgnd < 0.0:
            sty, fy, dy = stx, fx, dx
        stx, fx, dx = stp, fp, dp
    return stx, fx, dx, sty, fy, dy, stpf, brackt


class Dcsrch:
    """Reverse-communication More'-Thuente search; ``start`` then ``step``.

    Return value of both is (stp, task) with task in {"FG", "CONV", "WARN"}.
    """

    def __init__(self, ftol=FTOL_LS, gtol=GTOL_LS, xtol=XTOL_LS, stpmin=0.0, stpmax=STPMAX):
        self.ftol, self.gtol, self.xtol = ftol, gtol, xtol
        self.stpmin, self.stpmax = stpmin, stpmax

    def start(self, stp, f, g):
        assert g < 0.0 and self.stpmin <= stp <= self.stpmax
        self.brackt = False
        self.stage = 1
        self.finit, self.ginit = f, g
        self.gtest = self.ftol * g
        self.width = self.stpmax - self.stpmin
        self.width1 = self.width / 0.5
        self.stx, self.fx, self.gx = 0.0, f, g
        self.sty, self.fy, self.gy = 0.0, f, g
        self.stmin = 0.0
        self.stmax = stp + 4.0 * stp
        return stp, "FG"

    def step(self, stp, f, g):
        ftest = self.finit + stp * self.gtest
        if self.stage == 1 and f <= ftest and g >= 0.0:
            self.stage = 2
        task = "FG"
        if self.brackt and (stp <= self.stmin or stp >= self.stmax):
            task = "WARN"  # rounding errors prevent progress
        if self.brackt and self.stmax - self.stmin <= self.xtol * self.stmax:
            task = "WARN"  # xtol test satisfied
        if stp == self.stpmax and f <= ftest and g <= self.gtest:
            task = "WARN"
        if stp == self.stpmin and (f > ftest or g >= self.gtest):
            task = "WARN"
        if f <= ftest and abs(g) <= self.gtol * (-self.ginit):
            task = "CONV"
        if task != "FG":
            return stp, task
        if self.stage == 1 and f <= self.fx and f > ftest:
            fm = f - stp * self.gtest
            fxm = self.fx - self.stx * self.gtest
            fym = self.fy - self.sty * self.gtest
            gm = g - self.gtest
            gxm = self.gx - self.gtest
            gym = self.gy - self.gtest
            (self.stx, fxm, gxm, self.sty, fym, gym, stp, self.brackt) = _dcstep(
                self.stx, fxm, gxm, self.sty, fym, gym, stp, fm, gm,
                self.brackt, self.stmin, self.stmax)
            self.fx = fxm + self.stx * self.gtest
            self.fy = fym + self.sty * self.gtest
            self.gx = gxm + self.gtest
            self.gy = gym + self.gtest
        else:
            (self.stx, self.fx, self.gx, self.sty, self.fy, self.gy, stp, self.brackt) = _dcstep(
                self.stx, self.fx, self.gx, self.sty, self.fy, self.gy, stp, f, g,
                self.brackt, self.stmin, self.stmax)
        if self.brackt:
            if abs(self.sty - self.stx) >= 0.66 * self.width1:
                stp = self.stx + 0.5 * (self.sty - self.stx)
            self.width1 = self.width
            self.width = abs(self.sty - self.stx)
        if self.brackt:
            self.stmin = min(self.stx, self.sty)
            self.stmax = max(self.stx, self.sty)
        else:
            self.stmin = stp + 1.1 * (stp - self.stx)
            self.stmax = stp + 4.0 * (stp - self.stx)
        stp = min(max(stp, self.stpmin), self.stpmax)
        if (self.brackt and (stp <= self.stmin or stp >= self.stmax)) or (
                self.brackt and self.stmax - self.stmin <= self.xtol * self.stmax):
            stp = self.stx
        return stp, "FG"


# --------------------------------------------------------------------------
# Objective: mean pointwise loss + 0.5 * l2 * ||w||^2, intercept unpenalised
# --------------------------------------------------------------------------
def logistic_loss_grad(coef, X, y, n_classes, fit_intercept, l2, dtype=np.float64):
    """Loss and gradient for coef laid out (T, d[+1]); T = 1 (binary) or n_classes.

    ``dtype`` is the dtype of the two matrix products (sklearn uses X.dtype,
    _linear_loss.py:216-222,355-368); everything else is fp64.
    """
    n, d = X.shape
    T = 1 if n_classes == 2 else n_classes
    W = coef.reshape(T, -1)
    w, b = (W[:, :d], W[:, d]) if fit_intercept else (W, np.zeros(T))
    Xc = X.astype(dtype, copy=False)
    raw = (Xc @ w.T.astype(dtype)).astype(np.float64) + b
    if T == 1:
        z = raw[:, 0]
        # log(1+e^z) - y z, stable (sklearn _loss.pyx.tp:686-751)
        loss = np.where(z > 0, z + np.log1p(np.exp(-np.abs(z))), np.log1p(np.exp(-np.abs(z)))) - y * z
        p = np.where(z >= 0, 1.0 / (1.0 + np.exp(-np.abs(z))), np.exp(-np.abs(z)) / (1.0 + np.exp(-np.abs(z))))
        gp = (p - y)[:, None]
    else:
        m = raw.max(axis=1, keepdims=True)
        e = np.exp(raw - m)
        s = e.sum(axis=1, keepdims=True)
        yi = y.astype(np.int64)
        loss = (np.log(s) + m)[:, 0] - raw[np.arange(n), yi]
        gp = e / s
        gp[np.arange(n), yi] -= 1.0
    f = float(loss.sum() / n) + 0.5 * l2 * float((w * w).sum())
    gp = gp / n
    G = np.empty_like(W)
    G[:, :d] = (gp.T.astype(dtype) @ Xc).astype(np.float64) + l2 * w
    if fit_intercept:
        G[:, d] = gp.sum(axis=0)
    return f, G.reshape(coef.shape)


def _two_loop(g, S, Y, theta):
    q = g.copy()
    al = []
    for s, yv in zip(reversed(S), reversed(Y)):
        a = (s @ q) / (yv @ s)
        al.append(a)
        q -= a * yv
    q /= theta
    for (s, yv), a in zip(zip(S, Y), reversed(al)):
        q += (a - (yv @ q) / (yv @ s)) * s
    return -q


def lbfgs_minimize(fg, x0, tol=1e-4, max_iter=100, maxls=50, record=None):
    """Unbounded L-BFGS-B.  Returns dict(x, f, nit, nfev, status, g).

    status 0 = converged, 1 = iteration cap (scipy warnflag 1), 2 = abnormal
    line-search termination.
    """
    x = np.asarray(x0, dtype=np.float64).copy()
    f, g = fg(x)
    nfev, nit = 1, 0
    S, Y, theta = [], [], 1.0
    if np.abs(g).max() <= tol:
        return dict(x=x, f=f, g=g, nit=0, nfev=nfev, status=0)
    status = 0
    while True:
        d = _two_loop(g, S, Y, theta) if S else -g
        gd = float(g @ d)
        x_old, f_old, g_old, gd_old = x, f, g, gd
        failed = gd >= 0.0  # lnsrlb info = -4: not a descent direction
        if not failed:
            stp = min(1.0 / float(np.sqrt(d @ d)), STPMAX) if nit == 0 else 1.0
            ls = Dcsrch()
            stp, task = ls.start(stp, f, gd)
            ifun = 0
            while task == "FG":
                if ifun >= maxls:  # mainlb: iback >= maxls
                    failed = True
                    break
                ifun += 1
                x = x_old + stp * d
                f, g = fg(x)
                nfev += 1
                gd = float(g @ d)
                stp, task = ls.step(stp, f, gd)
        if failed:
            x, f, g = x_old, f_old, g_old
            if not S:  # col == 0: ABNORMAL_TERMINATION_IN_LNSRCH
                status = 2
                break
            S, Y, theta = [], [], 1.0  # RESTART_FROM_LNSRCH: drop the memory
            continue
        nit += 1
        if record is not None:
            record.append(x.copy())
        if nit >= max_iter:  # _lbfgsb_py.py:429-431 checks the cap first
            status = 1
            break
        if np.abs(g).max() <= tol:
            break
        if (f_old - f) <= EPS * FACTR * max(abs(f_old), abs(f), 1.0):
            break
        yv = g - g_old
        dr = (gd - gd_old) * stp  # s'y as L-BFGS-B forms it
        ddum = -gd_old * stp
        if dr > EPS * ddum:
            s = stp * d
            S.append(s)
            Y.append(yv)
            theta = float(yv @ yv) / dr
            if len(S) > MAXCOR:
                S.pop(0)
                Y.pop(0)
    return dict(x=x, f=f, g=g, nit=nit, nfev=nfev, status=status)


def fit_logreg(X, y_idx, n_classes, C=1.0, tol=1e-4, fit_intercept=True, max_iter=100,
               matmul_dtype=np.float64):
    """One LogisticRegression(lbfgs) fit on (X, y_idx in 0..n_classes-1).

    Returns dict(coef (T,d), intercept (T,), n_iter, nfev, status).
    """
    n, d = X.shape
    T = 1 if n_classes == 2 else n_classes
    l2 = 1.0 / (C * n)
    yv = y_idx.astype(np.float64)
    x0 = np.zeros(T * (d + int(fit_intercept)))
    res = lbfgs_minimize(
        lambda w: logistic_loss_grad(w, X, yv, n_classes, fit_intercept, l2, matmul_dtype),
        x0, tol=tol, max_iter=max_iter)
    W = res["x"].reshape(T, -1)
    return dict(coef=W[:, :d].copy(),
                intercept=(W[:, d].copy() if fit_intercept else np.zeros(T)),
                n_iter=min(res["nit"], max_iter), nfev=res["nfev"], status=res["status"])
