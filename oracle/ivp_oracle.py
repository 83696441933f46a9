"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the initial-value integration dynamo runs through its vector
field.  Nothing in the product imports this file.

The reference (prediction/utils.py:62-259, integrate_vf_ivp) calls scipy.integrate.solve_ivp once per initial state
with the default explicit Runge-Kutta 5(4) pair of Dormand and Prince, rtol = 1e-3, atol = 1e-6,
max_step = |t[-1] - t[0]| / interpolation_num (:96), dense_output = True and ONE terminal event
`np.all(abs(f(x)) < 1e-5) - 1` (:107), i.e. an event function that takes the two values -1 and 0.  scipy is a
third-party dependency of the reference (requirements.txt gives `scipy>=1.4.1`, no pin; 1.18.1 is installed here and
on the GPU box); the algorithm restated below is the published one of scipy/integrate/_ivp/{ivp,rk,base,common}.py:
  * initial step: Hairer, Norsett, Wanner, Solving ODEs I, sec. II.4 (select_initial_step),
  * step control: error norm = RMS of err / (atol + rtol max(|y|, |y_new|)), factor = 0.9 err^-1/5 clipped to
    [0.2, 10], no growth right after a rejection, steps clipped to [10 ulp(t), max_step], last step cut to t_bound,
  * dense output: the 4th-order interpolant  y(t) = y_old + h Q [x, x^2, x^3, x^4],  Q = K^T P,  x = (t - t_old) / h,
  * event: with a two-valued event function the bracketing root finder (brentq) returns an end point at once: the
    step's end when the previous value was -1, its start when it was already 0; the state stored for that time is
    the interpolant's value there (ivp.py handle_events / `y = sol(t)`).
Pinned by tests/test_ivp_cpu.py against scipy.integrate.solve_ivp itself (same t, y, sol(t), status, bit for bit)
and against golden vectors from the executed reference integrate_vf_ivp (tests/golden/ivp/, oracle/gen_golden_ivp.py).
"""
from __future__ import annotations

import numpy as np

# Dormand-Prince tableau (scipy/integrate/_ivp/rk.py, class RK45)
C = np.array([0, 1 / 5, 3 / 10, 4 / 5, 8 / 9, 1])
A = np.array([
    [0, 0, 0, 0, 0],
    [1 / 5, 0, 0, 0, 0],
    [3 / 40, 9 / 40, 0, 0, 0],
    [44 / 45, -56 / 15, 32 / 9, 0, 0],
    [19372 / 6561, -25360 / 2187, 64448 / 6561, -212 / 729, 0],
    [9017 / 3168, -355 / 33, 46732 / 5247, 49 / 176, -5103 / 18656],
])
B = np.array([35 / 384, 0, 500 / 1113, 125 / 192, -2187 / 6784, 11 / 84])
E = np.array([-71 / 57600, 0, 71 / 16695, -71 / 1920, 17253 / 339200, -22 / 525, 1 / 40])
P = np.array([
    [1, -8048581381 / 2820520608, 8663915743 / 2820520608, -12715105075 / 11282082432],
    [0, 0, 0, 0],
    [0, 131558114200 / 32700410799, -68118460800 / 10900136933, 87487479700 / 32700410799],
    [0, -1754552775 / 470086768, 14199869525 / 1410260304, -10690763975 / 1880347072],
    [0, 127303824393 / 49829197408, -318862633887 / 49829197408, 701980252875 / 199316789632],
    [0, -282668133 / 205662961, 2019193451 / 616988883, -1453857185 / 822651844],
    [0, 40617522 / 29380423, -110615467 / 29380423, 69997945 / 29380423],
])
SAFETY, MIN_FACTOR, MAX_FACTOR = 0.9, 0.2, 10.0
EVENT_TOL = 1e-5   # prediction/utils.py:107

FINISHED, TERMINATED, FAILED = 0, 1, -1


def _rms(x):
    return np.linalg.norm(x) / x.size ** 0.5


def _at_rest(f):
    return bool(np.all(np.abs(f) < EVENT_TOL))


def first_step(fun, t0, y0, t_bound, max_step, f0, direction, rtol, atol):
    span = abs(t_bound - t0)
    if span == 0.0:
        return 0.0
    scale = atol + np.abs(y0) * rtol
    d0, d1 = _rms(y0 / scale), _rms(f0 / scale)
    h0 = 1e-6 if (d0 < 1e-5 or d1 < 1e-5) else 0.01 * d0 / d1
    h0 = min(h0, span)
    f1 = fun(y0 + h0 * direction * f0)
    d2 = _rms((f1 - f0) / scale) / h0
    h1 = max(1e-6, h0 * 1e-3) if (d1 <= 1e-15 and d2 <= 1e-15) else (0.01 / max(d1, d2)) ** (1 / 5)
    return min(100 * h0, h1, span, max_step)


def rk45_solve(fun, t0, t_bound, y0, max_step=np.inf, rtol=1e-3, atol=1e-6):
    """Integrate y' = fun(y) from t0 to t_bound the way the reference's solve_ivp call does.
    Returns dict(t (n+1,), y (D, n+1), Q (n, D, 4), h (n,), status): status 0 reached t_bound, 1 stopped by the event,
    -1 step size fell below 10 ulp.  h[s] is the step the interpolant of segment s was built on: t[s+1] - t[s]
    except when the start is already at rest and so is the end of the first step -- then the event's root is the
    step's START, t = [t0, t0], and the one interpolant keeps the step that was taken."""
    y = np.array(y0, dtype=np.float64)
    n = y.size
    t = float(t0)
    direction = float(np.sign(t_bound - t0)) if t_bound != t0 else 1.0
    f = fun(y)
    h_abs = first_step(fun, t, y, t_bound, max_step, f, direction, rtol, atol)
    K = np.empty((7, n))
    ts, ys, Qs, hs = [t], [y], [], []
    rest_old = _at_rest(f)
    status = None
    while status is None:
        if t == t_bound:     # OdeSolver.step corner case (zero span)
            status = FINISHED
            break
        min_step = 10 * np.abs(np.nextafter(t, direction * np.inf) - t)
        h_abs = max_step if h_abs > max_step else (min_step if h_abs < min_step else h_abs)
        rejected = False
        while True:
            if h_abs < min_step:
                status = FAILED
                break
            h = h_abs * direction
            t_new = t + h
            if direction * (t_new - t_bound) > 0:
                t_new = t_bound
            h = t_new - t
            h_abs = np.abs(h)
            K[0] = f
            for s in range(1, 6):
                K[s] = fun(y + np.dot(K[:s].T, A[s, :s]) * h)
            y_new = y + h * np.dot(K[:-1].T, B)
            f_new = fun(y_new)
            K[6] = f_new
            scale = atol + np.maximum(np.abs(y), np.abs(y_new)) * rtol
            err = _rms(np.dot(K.T, E) * h / scale)
            if err < 1:
                factor = MAX_FACTOR if err == 0 else min(MAX_FACTOR, SAFETY * err ** -0.2)
                if rejected:
                    factor = min(1, factor)
                h_abs *= factor
                break
            h_abs *= max(MIN_FACTOR, SAFETY * err ** -0.2)
            rejected = True
        if status == FAILED:
            break
        Q = K.T.dot(P)
        t_old, y_old = t, y
        t, y, f = t_new, y_new, f_new
        Qs.append(Q)
        hs.append(h)
        if direction * (t - t_bound) >= 0:
            status = FINISHED
        rest_new = _at_rest(f)
        if rest_new or rest_old:
            # find_active_events: the event function (0 at rest, -1 otherwise) fires on -1 -> 0 (brentq returns t),
            # 0 -> 0 and 0 -> -1 (a start at rest; brentq returns t_old, so t = [t0, t0])
            root = t_old if rest_old else t
            x = (root - t_old) / h
            p = np.cumprod(np.tile(x, 4))
            y = h * np.dot(Q, p) + y_old
            t = root
            status = TERMINATED
        rest_old = rest_new
        ts.append(t)
        ys.append(y)
    return {"t": np.array(ts), "y": np.array(ys).T, "Q": np.array(Qs).reshape(len(Qs), n, 4), "h": np.array(hs),
            "status": status}


def dense_eval(sol, t):
    """scipy's OdeSolution.__call__ on the pieces of rk45_solve (segment choice of common.py, RkDenseOutput of
    rk.py); t 1-d -> (D, len(t))."""
    ts, ys, Q = sol["t"], sol["y"], sol["Q"]
    t = np.asarray(t, dtype=np.float64)
    nseg = len(Q)
    if ts[-1] >= ts[0]:
        seg = np.searchsorted(ts, t, side="left") - 1
        seg = np.clip(seg, 0, nseg - 1)
    else:
        seg = np.searchsorted(ts[::-1], t, side="right") - 1
        seg = nseg - 1 - np.clip(seg, 0, nseg - 1)
    out = np.empty((ys.shape[0], t.size))
    for i, (s, ti) in enumerate(zip(seg, t)):
        h = sol["h"][s]
        x = (ti - ts[s]) / h
        out[:, i] = h * np.dot(Q[s], np.cumprod(np.tile(x, 4))) + ys[:, s]
    return out
