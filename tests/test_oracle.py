"""CPU suite: the oracle against the hand-derived golden vectors (tests/golden/, minted by
make_golden.py because the reference has none) and against its independently written numpy
twin.  No GPU, no product code."""
import math

import numpy as np
import pytest
from hypothesis import given, settings
from hypothesis import strategies as st

from conftest import golden
from oracle import oracle_np as onp


def rel(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1e-300))) if a.size else 0.0


# ---------------------------------------------------------------- Utils.round
def test_round_kats(oracle):
    for x, want in golden("round_kat.json")["cases"]:
        for impl in (oracle.round_half_up, onp.round_half_up):
            got = impl(x)
            assert got == want and math.copysign(1, got) == math.copysign(1, want), (x, got, want)


@settings(max_examples=300, deadline=None)
@given(st.floats(allow_nan=False, allow_infinity=False, width=64))
def test_round_c_equals_decimal(oracle, x):
    assert oracle.round_half_up(x) == onp.round_half_up(x)


@settings(max_examples=200, deadline=None)
@given(st.integers(-10**6, 10**6))
def test_round_ties_away_from_zero(oracle, k):
    x = k + 0.5
    assert oracle.round_half_up(x) == (k + 1 if x > 0 else k)


# ---------------------------------------------------------------- hashing / featuriser
def test_hash_kats(oracle):
    for c in golden("featurizer_kat.json")["hash_kats"]:
        assert onp.java_hash(c["units"]) == c["hash"]
        assert oracle.hash_term(c["units"], 1000) == c["bucket"]
        assert onp.hash_term(c["units"], 1000) == c["bucket"]


def _pack(tweets):
    units = [u for t in tweets for u in t["units"]]
    off = np.cumsum([0] + [len(t["units"]) for t in tweets]).astype(np.int64)
    col = lambda k, dt: np.array([t[k] for t in tweets], dt)
    return (np.array(units, np.uint16), off, col("retweet_count", np.int32), col("followers", np.int32),
            col("favourites", np.int32), col("friends", np.int32), col("created_ms", np.int64),
            col("is_retweet", np.uint8))


def test_featurizer_kat(oracle):
    g = golden("featurizer_kat.json")
    rowptr, colidx, vals, labels = oracle.featurize(*_pack(g["tweets"]), g["now_ms"], g["rt_begin"],
                                                    g["rt_end"], g["num_text_features"])
    exp = g["expected_rows"]
    assert len(labels) == len(exp) == len(rowptr) - 1          # row count is bit-exact
    for r, e in enumerate(exp):
        sl = slice(rowptr[r], rowptr[r + 1])
        assert colidx[sl].tolist() == e["idx"]
        assert vals[sl].tolist() == e["val"]                   # exact doubles
        assert labels[r] == e["label"]


def test_featurizer_edge_kats(oracle):
    """Lone surrogates, empty and one-unit texts, F = 1 and F = 7 (tests/golden/featurizer_edge_kat.json)."""
    g = golden("featurizer_edge_kat.json")
    for grp in g["groups"]:
        F = grp["num_text_features"]
        rowptr, colidx, vals, labels = oracle.featurize(*_pack(grp["tweets"]), g["now_ms"], g["rt_begin"], g["rt_end"], F)
        assert len(labels) == len(grp["expected_rows"])
        for r, e in enumerate(grp["expected_rows"]):
            sl = slice(rowptr[r], rowptr[r + 1])
            assert colidx[sl].tolist() == e["idx"] and vals[sl].tolist() == e["val"] and labels[r] == e["label"]
        py = onp.featurize([dict(t, text=_units_to_str(t["units"])) for t in grp["tweets"]], g["now_ms"], g["rt_begin"], g["rt_end"], F)
        assert py[1].tolist() == colidx.tolist() and py[2].tolist() == vals.tolist()


def _units_to_str(units):
    return np.array(units, np.uint16).tobytes().decode("utf-16-le", "surrogatepass")


@settings(max_examples=60, deadline=None)
@given(st.lists(st.tuples(st.text(max_size=40), st.integers(0, 2000), st.integers(0, 2**31 - 1),
                          st.integers(0, 10**5), st.integers(0, 10**5), st.integers(0, 10**9),
                          st.booleans()), max_size=8))
def test_featurizer_c_equals_python(oracle, rows):
    now = 1464739200000
    tweets = [dict(text=t.lower(), units=onp.utf16_units(t.lower()), retweet_count=rc, followers=fo,
                   favourites=fa, friends=fr, created_ms=now - age, is_retweet=int(rt))
              for t, rc, fo, fa, fr, age, rt in rows]
    a = oracle.featurize(*_pack(tweets), now) if tweets else None
    b = onp.featurize(tweets, now)
    if a is None:
        assert len(b[3]) == 0
        return
    for x, y in zip(a, b):
        assert x.tolist() == y.tolist()
    # CSR invariants: strictly increasing indices per row, numeric slots always present
    rowptr, colidx = a[0], a[1]
    for r in range(len(rowptr) - 1):
        idx = colidx[rowptr[r]:rowptr[r + 1]]
        assert np.all(np.diff(idx) > 0) and idx[-4:].tolist() == [1000, 1001, 1002, 1003]


# ---------------------------------------------------------------- SGD
@pytest.mark.parametrize("case", golden("sgd_kat.json")["cases"], ids=lambda c: c["name"])
def test_sgd_kats(oracle, case):
    X = np.array(case["X"], float).reshape(len(case["y"]), len(case["w0"]))
    for parts in (1, 2):
        w, hist = oracle.sgd(case["y"], len(case["w0"]), case["w0"], X=X, step_size=case["step"],
                             num_iterations=case["iters"], convergence_tol=case["tol"], partitions=parts)
        wn, hn = onp.sgd(case["y"], len(case["w0"]), case["w0"], X=X, step_size=case["step"],
                         num_iterations=case["iters"], convergence_tol=case["tol"])
        assert len(hist) == len(case["loss_history"]) == len(hn)   # iterations actually run
        if case["weights_per_iter"]:
            assert rel(w, case["weights_per_iter"][-1]) < 1e-13
            assert rel(hist, case["loss_history"]) < 1e-13
            assert rel(wn, case["weights_per_iter"][-1]) < 1e-13
        else:
            assert w.tolist() == case["w0"]


@pytest.mark.parametrize("n,d,parts", [(257, 4, 1), (1000, 128, 3), (513, 1004, 8)])
def test_sgd_c_vs_numpy_dense(oracle, n, d, parts):
    X, y = oracle.synth_dense(n, d, seed=7)
    X2, y2 = onp.synth_dense(n, d, seed=7)
    assert np.array_equal(X, X2) and np.array_equal(y, y2)         # generator is bit-defined
    w0 = np.linspace(-0.1, 0.1, d)
    for upd, reg in ((0, 0.0), (1, 0.01), (2, 0.001)):
        w, h = oracle.sgd(y, d, w0, X=X, step_size=0.1, num_iterations=12, convergence_tol=0.0,
                          reg_param=reg, updater=upd, partitions=parts, threads=2)
        wn, hn = onp.sgd(y, d, w0, X=X, step_size=0.1, num_iterations=12, convergence_tol=0.0,
                         reg_param=reg, updater=upd)
        assert len(h) == 12 and rel(h, hn) < 1e-12
        assert np.max(np.abs(w - wn)) < 1e-12 * max(1.0, np.max(np.abs(wn)))


def test_sgd_c_vs_numpy_csr_and_fraction(oracle):
    n, dt = 700, 2000
    rowptr, colidx, vals, labels = oracle.synth_csr(n, dt, seed=3)
    d = dt + 4
    w0 = np.zeros(d)
    for frac in (1.0, 0.5):
        w, h = oracle.sgd(labels, d, w0, csr=(rowptr, colidx, vals), step_size=0.005, num_iterations=10,
                          mini_batch_fraction=frac, convergence_tol=0.0, partitions=4, row0=1000)
        wn, hn = onp.sgd(labels, d, w0, csr=(rowptr, colidx, vals), step_size=0.005, num_iterations=10,
                         mini_batch_fraction=frac, convergence_tol=0.0, row0=1000)
        assert len(h) == len(hn) == 10 and rel(h, hn) < 1e-12
        assert np.max(np.abs(w - wn)) < 1e-12 * max(1.0, np.max(np.abs(wn)))
    # csr and dense forms of the same rows agree
    Xd = onp._densify((rowptr, colidx, vals), n, d)
    wd, hd = oracle.sgd(labels, d, w0, X=Xd, step_size=0.005, num_iterations=10, convergence_tol=0.0)
    wc, hc = oracle.sgd(labels, d, w0, csr=(rowptr, colidx, vals), step_size=0.005, num_iterations=10,
                        convergence_tol=0.0)
    assert rel(hd, hc) < 1e-13 and np.max(np.abs(wd - wc)) < 1e-13


def test_sgd_default_tolerance_stops_early(oracle):
    X, y = oracle.synth_dense(400, 16, seed=1)
    w, h = oracle.sgd(y, 16, np.zeros(16), X=X, step_size=1e-5, num_iterations=50)   # tol 0.001 default
    assert len(h) == 2
    wn, hn = onp.sgd(y, 16, np.zeros(16), X=X, step_size=1e-5, num_iterations=50)
    assert len(hn) == 2 and rel(w, wn) < 1e-12


def first_sample_empty_case(onp_mod):
    """miniBatchFraction under which the one-row batch at global row 0 draws NO row at iteration 1 but its row at
    iterations 2 and 3 (the counter-based sampler gives u = 0.844, 0.610, 0.019 there)."""
    frac = 0.7
    assert [bool(onp_mod.sample_mask(i, 0, 1, frac)[0]) for i in (1, 2, 3)] == [False, True, True]
    return frac


def test_convergence_check_waits_for_a_previous_nonempty_iteration(oracle):
    """GradientDescent [U] tests convergence only when previousWeights != None, i.e. from the SECOND iteration that drew
    a non-empty sample -- not from i = 2.  With an empty first sample and a step so small that any two consecutive
    weight vectors are within tolerance, the loop must run iterations 2 and 3 (two loss entries) and stop there; gating on
    i >= 2 would stop after iteration 2 with one entry."""
    frac = first_sample_empty_case(onp)
    X, y = np.array([[1.0, 1.0]]), np.array([1.0])
    kw = dict(step_size=1e-4, num_iterations=50, mini_batch_fraction=frac, convergence_tol=0.001, row0=0)
    w, h = oracle.sgd(y, 2, np.zeros(2), X=X, **kw)
    wn, hn = onp.sgd(y, 2, np.zeros(2), X=X, **kw)
    assert len(h) == 2 and len(hn) == 2
    assert rel(w, wn) < 1e-13 and rel(h, hn) < 1e-13
    # by hand: iteration 2 moves w by eta_2 * (0 - 1) * (1, 1) = 1e-4/sqrt(2) per coordinate, iteration 3 by 1e-4/sqrt(3) * (1 - 2 w)
    e2 = 1e-4 / math.sqrt(2)
    assert abs(w[0] - (e2 + 1e-4 / math.sqrt(3) * (1 - 2 * e2))) < 1e-18 and h[0] == 0.5


def test_sampler_is_partition_invariant(oracle):
    X, y = oracle.synth_dense(300, 8, seed=5)
    a = oracle.sgd(y, 8, np.zeros(8), X=X, num_iterations=5, mini_batch_fraction=0.3, convergence_tol=0.0,
                   partitions=1)
    b = oracle.sgd(y, 8, np.zeros(8), X=X, num_iterations=5, mini_batch_fraction=0.3, convergence_tol=0.0,
                   partitions=7)
    assert rel(a[0], b[0]) < 1e-12 and rel(a[1], b[1]) < 1e-12
    mask = onp.sample_mask(3, 0, 300, 0.3)
    assert [oracle.sampled(3, r, 0.3) for r in range(300)] == mask.tolist()
    assert 50 < mask.sum() < 130


# ---------------------------------------------------------------- predict + stats
def test_stats_kat(oracle):
    g = golden("stats_kat.json")
    X = np.array(g["X"], float)
    for parts in (1, 2, 4):
        s, raw, rnd = oracle.predict_stats(g["y"], 2, g["w"], X=X, partitions=parts)
        assert raw.tolist() == g["pred_raw"] and rnd.tolist() == g["pred_rounded"]
        assert s["batch"] == g["batch"] and s["mse"] == g["mse"]
        assert s["real_stdev"] == g["real_stdev"] and s["pred_stdev"] == g["pred_stdev"]
        assert abs(s["mse_raw"] - g["mse_raw"]) < 1e-15
        assert abs(s["real_stdev_raw"] - g["real_stdev_raw"]) < 1e-15
        assert abs(s["pred_stdev_raw"] - g["pred_stdev_raw"]) < 1e-15
    s2, raw2, rnd2 = onp.predict_stats(g["y"], 2, g["w"], X=X)
    assert rnd2.tolist() == g["pred_rounded"] and s2["mse"] == g["mse"]


def test_stats_c_vs_numpy(oracle):
    rowptr, colidx, vals, labels = oracle.synth_csr(5000, 1000, seed=11)
    d = 1004
    w = np.random.default_rng(0).normal(0, 30, d)
    s, raw, rnd = oracle.predict_stats(labels, d, w, csr=(rowptr, colidx, vals), partitions=5)
    s2, raw2, rnd2 = onp.predict_stats(labels, d, w, csr=(rowptr, colidx, vals))
    assert rel(raw, raw2) < 1e-12 and np.array_equal(rnd, rnd2)
    for k in ("mse_raw", "real_stdev_raw", "pred_stdev_raw", "real_mean", "pred_mean"):
        assert abs(s[k] - s2[k]) <= 1e-12 * abs(s2[k]), k
    for k in ("batch", "mse", "real_stdev", "pred_stdev"):
        assert s[k] == s2[k]
