"""CPU tests of the trajectory-integration row: the oracle is pinned to scipy.integrate.solve_ivp
(the third-party routine the reference calls, prediction/utils.py:115-135) and to golden vectors of the executed
reference; the host-side sampling logic of dynamo_release_b200.integrate is checked against the same golden vectors
with the oracle standing in for the device solver."""
import numpy as np
import pytest
from scipy.integrate import solve_ivp

import ivp_util
from golden_util import rel
from oracle import ivp_oracle as io
from oracle import vfc_oracle as vo


def _scipy_call(f, a, b, y0, max_step):
    ev = lambda t, x: np.all(abs(f(x)) < 1e-5) - 1  # noqa: E731  (prediction/utils.py:107)
    ev.terminal = True
    return solve_ivp(lambda t, x: f(x), [a, b], y0, events=ev, max_step=max_step, dense_output=True)


@pytest.mark.parametrize("D,M,T,interp", [(2, 20, 50.0, 250), (5, 40, 10.0, 100), (30, 100, 3.0, 250),
                                          (3, 25, 400.0, 4), (4, 30, 30.0, 1)])
def test_oracle_equals_scipy_solve_ivp_bit_for_bit(D, M, T, interp):
    rng = np.random.default_rng(D * 1000 + M)
    vf = {"X_ctrl": rng.normal(size=(M, D)), "C": rng.normal(size=(M, D)) * 0.3, "beta": 0.5 / D}
    f = lambda x: vo.field_eval(x, vf)  # noqa: E731
    statuses, steps = set(), set()
    for k in range(3):
        y0 = rng.normal(size=D)
        for a, b in ((0.0, T), (-0.0, -T)):
            r = _scipy_call(f, a, b, y0, T / interp)
            g = io.rk45_solve(f, a, b, y0, max_step=T / interp)
            assert r.status == g["status"]
            assert np.array_equal(r.t, g["t"]) and np.array_equal(r.y, g["y"])
            tt = np.linspace(a, r.t[-1], 41)
            assert rel(io.dense_eval(g, tt), r.sol(tt)) < 1e-14    # scipy groups times per segment (GEMM vs dot)
            statuses.add(r.status)
            steps.add(len(r.t))
    print("statuses", statuses, "steps", sorted(steps))


def test_oracle_matches_reference_raw_solver_output():
    z = ivp_util.load()
    fr = z["sol_fractions"]
    seen = set()
    for name in ivp_util.FIELDS:
        vf = ivp_util.field(z, name)
        f = lambda x: vo.field_eval(x, vf)  # noqa: E731
        for te in z[f"{name}/t_end"]:
            for sign, tag in ((1.0, "fw"), (-1.0, "bw")):
                for i, y0 in enumerate(z[f"{name}/init"]):
                    key = f"{name}/T{te:g}/raw_{tag}/{i}"
                    g = io.rk45_solve(f, sign * 0.0, sign * te, y0, max_step=te / int(z["interp"]))
                    assert g["status"] == int(z[key + "/status"])
                    assert g["t"].shape == z[key + "/t"].shape
                    assert rel(g["t"], z[key + "/t"]) < 1e-12 and rel(g["y"], z[key + "/y"]) < 1e-12
                    tt = g["t"][0] + fr * (g["t"][-1] - g["t"][0])
                    assert rel(io.dense_eval(g, tt), z[key + "/sol"]) < 1e-12
                    seen.add(g["status"])
    assert seen == {0, 1}      # both "reached the end time" and "came to rest" are covered


def test_host_sampling_logic_matches_reference(monkeypatch):
    from dynamo_release_b200 import integrate as it

    monkeypatch.setattr(it, "solve_ivp_batch", ivp_util.oracle_batch(it.DenseSolution, it.IvpResult))
    z = ivp_util.load()
    interp = int(z["interp"])
    worst = 0.0
    for name, te, direction, sampling, average in ivp_util.combos(z):
        f = it.RkhsField(ivp_util.field(z, name))
        t, Y = it.integrate_vf_ivp(z[f"{name}/init"], np.linspace(0, te, 7), direction, f, interpolation_num=interp,
                                   average=average, sampling=sampling, disable=True)
        key = f"{name}/T{te:g}/{direction}/{sampling}/avg{int(average)}"
        assert len(Y) == int(z[key + "/n"])
        for i in range(len(Y)):
            assert Y[i].shape == z[f"{key}/Y{i}"].shape and np.shape(t[i]) == z[f"{key}/t{i}"].shape
            e = max(rel(t[i], z[f"{key}/t{i}"]), rel(Y[i], z[f"{key}/Y{i}"]))
            worst = max(worst, e)
            assert e < 1e-10, (key, i, e)
    print("worst relative difference to the reference:", worst)


def test_fate_wrapper_matches_reference(monkeypatch):
    from dynamo_release_b200 import integrate as it

    monkeypatch.setattr(it, "solve_ivp_batch", ivp_util.oracle_batch(it.DenseSolution, it.IvpResult))
    print("_fate vs reference: worst", ivp_util.check_fate(ivp_util.load(), it, 1e-10))


def test_dense_solution_equals_scipy_ode_solution():
    from dynamo_release_b200.integrate import DenseSolution

    rng = np.random.default_rng(3)
    vf = {"X_ctrl": rng.normal(size=(30, 4)), "C": rng.normal(size=(30, 4)) * 0.4, "beta": 0.2}
    f = lambda x: vo.field_eval(x, vf)  # noqa: E731
    for a, b in ((0.0, 5.0), (-0.0, -5.0)):
        r = _scipy_call(f, a, b, rng.normal(size=4), 0.1)
        g = io.rk45_solve(f, a, b, r.y[:, 0], max_step=0.1)
        s = DenseSolution(g["t"], g["y"], g["Q"], g["h"])
        tt = np.concatenate([rng.uniform(min(a, b) - 0.5, max(a, b) + 0.5, 200), r.t])   # outside, inside, on the knots
        assert rel(s(tt), r.sol(tt)) < 1e-14
        assert rel(s(tt[7]), r.sol(tt[7])) < 1e-14 and s(tt[7]).shape == (4,)
    with pytest.raises(ValueError):
        DenseSolution(np.array([0.0, 0.0, 1.0]), np.zeros((4, 3)), np.zeros((2, 4, 4)))


def test_start_at_rest_like_scipy():
    """f == 0: the event's root is the start of the first step; scipy returns t = [0, 0], status 1."""
    f = lambda x: np.zeros(2)  # noqa: E731
    y0 = np.array([1.0, 2.0])
    r = _scipy_call(f, 0.0, 1.0, y0, 0.1)
    g = io.rk45_solve(f, 0.0, 1.0, y0, max_step=0.1)
    assert r.status == g["status"] == 1 and np.array_equal(r.t, g["t"]) and np.array_equal(r.y, g["y"])
    assert np.array_equal(r.sol(np.array([0.0, 5e-7])), io.dense_eval(g, np.array([0.0, 5e-7])))


def test_start_at_rest_and_leave_it_like_scipy():
    """Round-1 advisor finding: a start just under the at-rest threshold on a growing field.  The event function goes
    0 -> -1 over the first step; scipy's find_active_events fires on that too and brentq returns t_old: t = [t0, t0],
    status 1."""
    f = lambda x: 9e-6 + 1e5 * x  # noqa: E731
    y0 = np.array([0.0])
    r = _scipy_call(f, 0.0, 1.0, y0, 0.1)
    g = io.rk45_solve(f, 0.0, 1.0, y0, max_step=0.1)
    assert r.status == g["status"] == 1
    assert np.array_equal(r.t, g["t"]) and np.array_equal(r.t, [0.0, 0.0]) and np.array_equal(r.y, g["y"])


def test_large_cell_sets_are_sliced_without_changing_results(monkeypatch):
    """More cells than the HBM budget for resident dense solutions: solve -> resample -> evaluate runs per slice."""
    from dynamo_release_b200 import integrate as it

    calls = []
    base = ivp_util.oracle_batch(it.DenseSolution, it.IvpResult)

    def counting(init_states, *a, **k):
        calls.append(len(init_states))
        return base(init_states, *a, **k)

    monkeypatch.setattr(it, "solve_ivp_batch", counting)
    z = ivp_util.load()
    f = it.RkhsField(ivp_util.field(z, "sink"))
    init, t = z["sink/init"], np.linspace(0, 3.0, 7)
    whole = it.integrate_vf_ivp(init, t, "both", f, interpolation_num=40, sampling="arc_length")
    assert calls == [5, 5]
    calls.clear()
    monkeypatch.setattr(it, "RESIDENT_BYTES", 2 * 2 * it._resident_bytes_per_cell(t, 40, 3))     # two cells per slice
    sliced = it.integrate_vf_ivp(init, t, "both", f, interpolation_num=40, sampling="arc_length")
    assert calls == [2, 2, 2, 2, 1, 1]
    for a, b in zip(whole[0] + whole[1], sliced[0] + sliced[1]):
        assert np.array_equal(a, b)
    calls.clear()
    it.integrate_vf_ivp(init, t, "forward", f, interpolation_num=40, sampling="uniform_indices")
    assert calls == [5]          # one shared time grid: never sliced


def test_no_initial_states_gives_empty_lists():
    from dynamo_release_b200.integrate import RkhsField, integrate_vf_ivp

    f = RkhsField({"X_ctrl": np.zeros((3, 2)), "C": np.zeros((3, 2)), "beta": 1.0})
    assert integrate_vf_ivp(np.zeros((0, 2)), np.linspace(0, 1, 3), "forward", f) == ([], [])


def test_generic_callable_is_refused():
    from dynamo_release_b200.integrate import integrate_vf_ivp

    with pytest.raises(NotImplementedError):
        integrate_vf_ivp(np.zeros((1, 2)), np.linspace(0, 1, 3), "forward", lambda x: x)


def _traced_vff(x, vf_dict, dim=None, X_ctrl_ind=None):
    """Stand-in for the package's vector_field_function on a box without CUDA: oracle arithmetic, same trace hook."""
    from dynamo_release_b200 import vf as vfm

    r = vo.field_eval(x, vf_dict)
    if dim is not None:
        r = r[..., :dim] if np.isscalar(dim) else r[..., dim]
    vfm._record_trace(x, vf_dict, dim, X_ctrl_ind, r)
    return r


def test_lambda_wrapped_fields_are_recognised(monkeypatch):
    """dynamo hands integrate_vf_ivp `lambda x: scale * vector_field_function(x=x, vf_dict=VecFld, dim=dims)`
    (prediction/fate.py:133-135): the trace recovers the field and the exact scale, and refuses everything else."""
    from dynamo_release_b200 import integrate as it

    rng = np.random.default_rng(8)
    vf = {"X_ctrl": rng.normal(size=(20, 3)), "C": rng.normal(size=(20, 3)), "beta": 0.3}
    other = {"X_ctrl": vf["X_ctrl"], "C": 2 * vf["C"], "beta": 0.3}
    probes = rng.normal(size=(5, 3))
    for scale in (1, 2.5, 0.1, 1 / 3, 1e-3, -7.0):
        for dims in (None, 3, [0, 1, 2]):
            d, s = it._field_of(lambda x: scale * _traced_vff(x=x, vf_dict=vf, dim=dims), probes)
            assert d is vf and s == float(scale)
    refused = [lambda x: x,                                                        # never touches the field
               lambda x: _traced_vff(x, vf) * x[0],                                # not a constant multiple
               lambda x: _traced_vff(x, vf) + 1.0,
               lambda x: _traced_vff(x, vf, dim=2),                                # a subset of the dimensions
               lambda x: _traced_vff(x, vf, X_ctrl_ind=[0, 1]),
               lambda x: _traced_vff(x, vf) + _traced_vff(x, other),              # two evaluations
               lambda x: _traced_vff(x, vf if x[0] > probes[1, 0] else other),     # different fields per call
               lambda x: _traced_vff(x * 2 + 1, vf),                               # the field somewhere else than at x
               lambda x: _traced_vff((x - 0.25) / 1.5, vf)]                        # (a normalising wrapper)
    for g in refused:
        with pytest.raises(NotImplementedError):
            it._field_of(g, probes)
    # a single initial state: perturbed copies are probed too, so a state-dependent factor is still seen to vary
    for g in (lambda x: _traced_vff(x, vf) * x[0], lambda x: _traced_vff(x * 2 + 1, vf)):
        with pytest.raises(NotImplementedError):
            it._field_of(g, probes[:1])
    d, s = it._field_of(lambda x: 0.7 * _traced_vff(x, vf), probes[:1])
    assert d is vf and s == 0.7
    # and end to end (oracle standing in for the device solver): same lists as with the field object
    monkeypatch.setattr(it, "solve_ivp_batch", ivp_util.oracle_batch(it.DenseSolution, it.IvpResult))
    t = np.linspace(0, 2.0, 5)
    a = it.integrate_vf_ivp(probes, t, "forward", lambda x: 2.5 * _traced_vff(x=x, vf_dict=vf, dim=None),
                            interpolation_num=20, sampling="logspace")
    b = it.integrate_vf_ivp(probes, t, "forward", it.ScaledRkhsField(vf, 2.5), interpolation_num=20, sampling="logspace")
    for u, v in zip(a[0] + a[1], b[0] + b[1]):
        assert np.array_equal(u, v)


def test_repeat_cut_equals_the_reference_on_periodic_trajectories():
    """The batched restatement of dup_osc_idx_iter (vectorfield/topography.py:332-397) against the executed reference,
    on limit-cycle-like curves where the cut really fires (several lengths, several periods, plus noise and a drift that
    must NOT be cut).  Needs /root/reference (build container only)."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("reference tree not present")
    topo = ref_loader.load("dynamo.vectorfield.topography")
    from dynamo_release_b200.integrate import _repeat_cuts

    rng = np.random.default_rng(9)
    xs = []
    for n in (40, 251, 251, 252, 300, 3, 1):
        for periods in (0.4, 3.0, 9.0, 27.0):
            s = np.linspace(0, 2 * np.pi * periods, n)
            x = np.stack([np.cos(s), np.sin(s), 0.3 * np.cos(2 * s)], 1) * rng.uniform(0.5, 2.0)
            xs.append(x)
            xs.append(x + 1e-3 * rng.normal(size=x.shape))
            xs.append(x + np.linspace(0, 5, n)[:, None])
    ours = _repeat_cuts(xs, max_iter=100)
    fired = 0
    for x, got in zip(xs, ours):
        ref = topo.dup_osc_idx_iter(x, max_iter=100, tol=(x.max(0) - x.min(0)).mean() / 1000)[0]
        assert got == ref
        fired += ref != len(x)
    assert fired >= 5


def test_threaded_scipy_fft_equals_numpy_fft_bitwise():
    """The repeat cut uses scipy.fft.rfft(workers=...) where the reference calls numpy.fft.rfft per column."""
    import scipy.fft

    rng = np.random.default_rng(2)
    a = rng.normal(size=(37, 251, 7))
    for lo, hi in ((83, 166), (0, 1), (5, 7), (10, 13), (100, 200)):
        got = scipy.fft.rfft(a[:, lo:hi], axis=1, workers=4)
        ref = np.stack([[np.fft.rfft(a[i, lo:hi, k]) for k in range(7)] for i in range(37)]).transpose(0, 2, 1)
        assert np.array_equal(got, ref)


def test_time_sequence_helpers_equal_the_reference():
    """svc._t_end / _t_seq against the executed getTend / getTseq (tools/utils.py:3175-3209); build container only."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("reference tree not present")
    tu = ref_loader.load("dynamo.tools.utils")
    from dynamo_release_b200.svc import _t_end, _t_seq

    rng = np.random.default_rng(4)
    X, V = rng.normal(size=(500, 7)) * 3, rng.normal(size=(500, 7)) * 0.1
    V[::17] = 0.0
    assert _t_end(X, V) == tu.getTend(X, V)
    for init_cols, t_end, step in ((7, 250.0, None), (7, 12345.6, None), (400, 1e9, None), (7, 30.0, 0.5), (7, 3.0, 1)):
        init = np.zeros((3, init_cols))
        assert np.array_equal(_t_seq(init, t_end, step), tu.getTseq(init, t_end, step))


def test_svc_integrate_method_equals_the_reference_method(monkeypatch):
    """SvcVectorField.integrate (scVectorField.py:721-784) incl. the default t_end of getTend, against the executed
    reference class on its own fit; the oracle stands in for the device solver.  Build container only."""
    from oracle import ref_loader

    if not ref_loader.available():
        pytest.skip("reference tree not present")
    sc = ref_loader.load("dynamo.vectorfield.scVectorField")
    from dynamo_release_b200 import integrate as it
    from dynamo_release_b200.svc import SvcVectorField

    rng = np.random.default_rng(12)
    X = rng.normal(size=(300, 3))
    V = -(X - 0.2) + 0.05 * rng.normal(size=X.shape)
    ref = sc.SvcVectorField(X, V, None, M=30, MaxIter=20, normalize=False)
    ref.train()
    want = ref.integrate(X[:4], interpolation_num=20, sampling="logspace", integration_direction="both", disable=True)
    monkeypatch.setattr(it, "solve_ivp_batch", ivp_util.oracle_batch(it.DenseSolution, it.IvpResult))
    vf = {k: ref.vf_dict[k] for k in ("X_ctrl", "C", "beta")}
    ours = SvcVectorField(vf_dict=vf, func=it.RkhsField(vf))
    ours.data["X"], ours.data["V"] = ref.get_X(), ref.get_V()
    got = ours.integrate(X[:4], interpolation_num=20, sampling="logspace", integration_direction="both", disable=True)
    assert len(got[1]) == len(want[1]) == 4
    for a, b in zip(got[0] + got[1], want[0] + want[1]):
        assert np.shape(a) == np.shape(b) and rel(a, b) < 1e-10
