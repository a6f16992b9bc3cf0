"""Data layer vs the golden fingerprints of SURVEY.md §6.3 and vs sklearn itself."""
import hashlib
import os

import numpy as np
import pytest

from fedb200.data import income as inc
from fedb200.data import partition as part


def test_table_shape_and_balance():
    t = inc.load_table()
    assert t.n_rows == 10000 and len(t.names) == 15
    assert (t.columns["income"] == ">50K").sum() == 5000


def test_label_encoding_is_sorted_rank():
    t = inc.load_table()
    enc, vocab = inc.label_encode(t)
    assert list(vocab["income"]) == ["<=50K", ">50K"]
    assert len(vocab["native.country"]) == 40 and len(vocab["education"]) == 16
    X = np.stack([enc[n] for n in t.names if n != "income"], 1).astype(np.float64)
    np.testing.assert_allclose(X.mean(0)[:4], [40.188, 2.215, 189072.749, 10.532], atol=2e-3)
    np.testing.assert_allclose(X.std(0)[:4], [12.549, 0.996, 105448.645, 3.503], atol=2e-3)


def test_split_matches_reference_permutation(income):
    tr, te = inc.split_indices(10000)
    assert list(te[:5]) == [6252, 4684, 1731, 4742, 4521]
    assert len(tr) == 8000 and len(te) == 2000
    assert income.y_train.sum() == 3987 and income.y_test.sum() == 1013


def test_scaled_sums(income, income_nomean):
    assert income.X_train.shape == (8000, 14) and income.X_train.dtype == np.float64
    assert abs(income.X_train.sum() - 66.019061) < 1e-4
    assert abs(income_nomean.X_train.sum() - 277813.390045) < 1e-3
    assert hashlib.sha256(income.X_train.tobytes()).hexdigest()[:16] == "b8b6ed9e2a11752f"


def test_matches_sklearn_pipeline_bitwise(income_nomean):
    from sklearn.model_selection import train_test_split
    from sklearn.preprocessing import LabelEncoder, StandardScaler
    t = inc.load_table()
    cols = []
    for n in t.names:
        c = t.columns[n]
        cols.append(LabelEncoder().fit_transform(c) if c.dtype.kind in "US" else c)
    data = dict(zip(t.names, cols))
    X = np.stack([data[n] for n in t.names if n != "income"], 1)
    Xs = StandardScaler(with_mean=False).fit_transform(X)
    Xtr, Xte, ytr, yte = train_test_split(Xs, data["income"], test_size=0.2, random_state=42)
    assert np.array_equal(Xtr, income_nomean.X_train) and np.array_equal(yte, income_nomean.y_test)


@pytest.mark.skipif(not os.path.exists("/root/reference/balanced_income_data.csv"), reason="reference CSV not mounted")
def test_packed_table_equals_csv():
    a, b = inc.load_table(), inc.load_table("/root/reference/balanced_income_data.csv")
    assert a.names == b.names
    for n in a.names:
        assert (a.columns[n] == b.columns[n]).all()


def test_missing_label_raises():
    with pytest.raises(KeyError):
        inc.prepare(label="Outcome")


def test_contiguous_partition_bounds():
    s = part.contiguous(8000, 3)
    assert [(int(x[0]), int(x[-1]) + 1) for x in s] == [(0, 2666), (2666, 5332), (5332, 8000)]
    assert [len(x) for x in part.contiguous(8000, 8)] == [1000] * 8


def test_per_client_shuffle_overlaps_and_is_seeded():
    a, b = part.per_client_shuffle(8000, 3, seed=5), part.per_client_shuffle(8000, 3, seed=5)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert [len(x) for x in a] == [2666, 2666, 2668]
    assert len(np.intersect1d(a[0], a[1])) > 0  # random subsets, not a partition (C:48-61)


def test_dirichlet_is_skewed(income):
    s = part.dirichlet(income.y_train, 16, alpha=0.1, seed=0)
    fr = np.array([income.y_train[i].mean() for i in s])
    assert fr.std() > 0.2 and min(len(i) for i in s) >= 8


def test_synthetic_shape():
    d = inc.synthetic_income_like(5000, seed=1)
    assert d.X_train.shape == (4000, 14) and set(np.unique(d.y_train)) == {0, 1}
    assert abs(d.y_train.mean() - 0.5) < 0.05
