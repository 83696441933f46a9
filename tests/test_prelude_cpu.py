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
