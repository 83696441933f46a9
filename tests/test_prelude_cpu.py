"""CPU: the torch-based prelude (stable lexicographic row sort, first-occurrence unique, sampling, bandwidth)
against the reference semantics restated in the oracle (np.unique / Generator.choice / sklearn kNN)."""
import numpy as np
import pytest
import torch

from dynamo_release_b200 import prelude
from golden_util import CASES, Golden
from oracle import vfc_oracle as o


def test_unique_rows_matches_numpy_with_ties_and_duplicates():
    rng = np.random.default_rng(0)
    X = rng.normal(size=(500, 4))
    X[:, 0] = np.round(X[:, 0], 1)          # many ties in the first column
    X[100:140] = X[20:60]                   # exact duplicate rows
    X[300, :2] = X[301, :2]                 # tie in two columns
    u, first = np.unique(X, axis=0, return_index=True)
    tu, tf = prelude.unique_rows(torch.from_numpy(X))
    assert np.array_equal(tu.numpy(), u)
    assert np.array_equal(tf.numpy(), first)


def test_unique_rows_one_column_and_empty():
    X = np.array([[3.0], [1.0], [3.0], [2.0]])
    u, first = np.unique(X, axis=0, return_index=True)
    tu, tf = prelude.unique_rows(torch.from_numpy(X))
    assert np.array_equal(tu.numpy(), u) and np.array_equal(tf.numpy(), first)


@pytest.mark.parametrize("name", [c for c in CASES if not c.startswith("lean_n8000")])
def test_prelude_matches_golden_ctrl_points(name):
    g = Golden(name)
    valid, Xv, Yv, ctrl, idx, beta = prelude.run(torch.from_numpy(g.X.copy()), torch.from_numpy(g.V.copy()),
                                                 g.params["M"], seed=0)
    if not g.lean:
        assert np.array_equal(valid.numpy(), g["final_valid_ind"])
    assert np.array_equal(idx, g["final_ctrl_idx"])
    assert np.array_equal(ctrl, g["final_X_ctrl"])
    assert abs(beta - float(g["final_beta"])) <= 1e-13 * abs(beta)


def test_bandwidth_matches_sklearn_tree():
    rng = np.random.default_rng(3)
    for shape in [(100, 30), (64, 2), (7, 3), (2, 5)]:
        c = rng.normal(size=shape)
        assert abs(prelude.bandwidth(c) - o.knn_bandwidth(c)) <= 1e-14 * o.knn_bandwidth(c)


def test_velocity_weighted_choice_equals_numpy_generator_choice():
    """The restated replace=False algorithm (C cumsum helper) picks exactly the cells Generator.choice picks."""
    for trial, (n, M) in enumerate([(200000, 500), (1000, 100), (300, 300), (50, 7), (5000, 100), (200, 199)]):
        rng = np.random.default_rng(trial)
        speed = np.abs(rng.normal(size=n)) + (0.0 if trial % 2 else 1e-3)
        for seed in (0, 1, 19491001):
            ref = np.random.default_rng(seed).choice(np.arange(n), size=M, p=speed / np.sum(speed), replace=False)
            assert np.array_equal(ref, prelude.velocity_weighted_choice(speed, M, seed)), (n, M, seed)


def test_choice_input_validation_like_numpy():
    """Round-1 advisor finding: Generator.choice raises for all-zero speeds / fewer non-zero entries than M; the restated
    sampler must do the same instead of spinning (tools/sampling.py:242-245 -> numpy/random/_generator.pyx)."""
    import pytest

    from dynamo_release_b200 import prelude

    speed = np.zeros(50)
    with pytest.raises(ValueError):
        np.random.default_rng(0).choice(50, size=5, p=speed / speed.sum() if speed.sum() else np.full(50, np.nan), replace=False)
    with pytest.raises(ValueError, match="NaN"):
        prelude.velocity_weighted_choice(speed, 5, 0)
    speed[[3, 7, 11]] = 1.0
    with pytest.raises(ValueError, match="Fewer non-zero"):
        np.random.default_rng(0).choice(50, size=5, p=speed / speed.sum(), replace=False)
    with pytest.raises(ValueError, match="Fewer non-zero"):
        prelude.velocity_weighted_choice(speed, 5, 0)
    assert sorted(prelude.velocity_weighted_choice(speed, 3, 0)) == [3, 7, 11]


def _host_cumsum(v, total=1.0, skip=()):
    import ctypes as C

    from dynamo_release_b200 import _lib
    v = np.ascontiguousarray(v, dtype=np.float64)
    raw = np.empty_like(v)
    sk = np.ascontiguousarray(sorted(skip), dtype=np.int64)
    with np.errstate(all="ignore"):
        _lib.check(_lib.load().dvf_host_choice_cumsum(v.ctypes.data_as(C.c_void_p), v.shape[0], float(total),
                                                      sk.ctypes.data_as(C.c_void_p) if sk.size else None, sk.size,
                                                      raw.ctypes.data_as(C.c_void_p)))
    return raw


def test_threaded_cumsum_is_numpy_cumsum_bit_for_bit():
    """host_cumsum.cpp forms cumsum(speed / total) on several threads (exact block sums inside one binade + a serial
    scan); every partial sum must carry the rounding of numpy's sequential recurrence, whatever the input looks like:
    half-way cases, binade crossings, zeros at the head, an outlier, tiny / subnormal values, negative or non-finite
    entries (the serial path), and the drawn cells zeroed through `skip`."""
    rng = np.random.default_rng(0)
    n = 700_003                          # 43 blocks of 16384 and a ragged tail
    cases = {
        "uniform": rng.random(n),
        "norms": np.linalg.norm(rng.normal(size=(n, 4)), axis=1),
        "lognormal": rng.lognormal(0, 5, n),
        "integers": rng.integers(0, 1 << 20, n).astype(np.float64),                       # every sum exact, ties
        "dyadic": rng.integers(0, 1 << 12, n).astype(np.float64) * 2.0 ** -40 + (rng.random(n) < 0.001) * 3.0,
        "leading zeros": np.concatenate([np.zeros(100_000), rng.random(n - 100_000)]),
        "mostly zeros": rng.random(n) * (rng.random(n) < 0.1),
        "outlier": np.concatenate([rng.random(n // 2), [1e12], rng.random(n // 2)]),
        "tiny": rng.random(n) * 1e-300,
        "subnormal": rng.random(n) * 1e-315,
        "negative": rng.normal(size=n),
        "inf": np.concatenate([rng.random(n - 5), [np.inf], rng.random(4)]),
        "halves": np.full(n, 0.5),
        "half ulp": np.concatenate([[1.0], np.full(n - 1, 2.0 ** -53)]),                   # every step is a tie
        "ulp mix": np.concatenate([[1.0], rng.choice([2.0 ** -53, 3 * 2.0 ** -53, 2.0 ** -52], n - 1)]),
        "short": rng.random(1000),
        "one": np.array([0.25]),
    }
    for name, v in cases.items():
        with np.errstate(all="ignore"):
            ref = np.cumsum(v)
        got = _host_cumsum(v)
        assert np.array_equal(ref.view(np.int64), got.view(np.int64)), name
    # quotients and zeroed cells, as Generator.choice's second round sees them
    v = cases["norms"]
    total = float(np.sum(v))
    skip = rng.choice(n, 300, replace=False)
    p = v / total
    p[skip] = 0.0
    assert np.array_equal(np.cumsum(p), _host_cumsum(v, total, skip))
    assert np.array_equal(np.cumsum(v / total), _host_cumsum(v, total))


def test_row_norms_leaves_no_cyclic_garbage():
    """row_norms' temporaries (two copies of the velocity matrix) must die by reference count when it returns.  A
    nested recursive helper once tied them into a reference cycle; on the GPU they then came back to the caching
    allocator only when the cyclic collector happened to run, and the benchmark loop hit cudaMalloc at random."""
    import gc
    import weakref

    Y = torch.from_numpy(np.random.default_rng(0).normal(size=(500, 30)))
    seen = []
    orig = torch.Tensor.__mul__

    def spy(self, *a, **k):
        out = orig(self, *a, **k)
        if out.shape == (30, 500):            # the squared, transposed velocity matrix
            seen.append(weakref.ref(out))
        return out

    gc.collect()
    gc.disable()
    try:
        torch.Tensor.__mul__ = spy
        ref = np.linalg.norm(Y.numpy(), axis=1)
        assert np.array_equal(prelude.row_norms(Y).numpy(), ref)
    finally:
        torch.Tensor.__mul__ = orig
        gc.enable()
    assert seen and all(r() is None for r in seen)
