"""CPU suite for the k-means app's oracle (oracle/twtml_oracle_kmeans.c): the hand-derived interval of
tests/golden/kmeans_kat.json and agreement with the independently written numpy twin.  No GPU, no product code."""
import numpy as np

from conftest import golden
from oracle import oracle_np as onp


def test_kmeans_kat(oracle):
    g = golden("kmeans_kat.json")
    X = np.array(g["points"], float)
    assert oracle.kmeans_predict(np.array(g["centers"], float), X).tolist() == g["closest_before"]
    for unit, ck, wk in ((0, "centers_after", "weights_after"), (1, "centers_after_points", "weights_after_points")):
        for impl in (oracle, onp):
            c, w = impl.kmeans_update(g["centers"], g["weights"], X, g["decay"], unit)
            assert np.allclose(c, g[ck], rtol=1e-14, atol=0) and w.tolist() == g[wk]
    c, w = oracle.kmeans_update(g["centers"], g["weights"], X, g["decay"], 0)
    assert oracle.kmeans_predict(c, X).tolist() == g["closest_after"]
    assert np.allclose(oracle.column_std(X), g["column_std"], rtol=1e-14, atol=0)
    d = g["dying"]                                             # the largest cluster is split when another one dies
    for impl in (oracle, onp):
        c, w = impl.kmeans_update(d["centers"], d["weights"], X, g["decay"], 0)
        assert w.tolist() == d["weights_after"]
        mid = np.array(d["center_mid"])
        assert np.max(np.abs(c[0] - mid)) < 1e-13 and np.max(np.abs(c[1] - mid)) < 1e-13
        assert c[0][0] > c[1][0] and c[0][1] > c[1][1]         # x + p and x - p, p = 1e-14 max(|x|, 1)


def test_kmeans_c_vs_numpy(oracle):
    rng = np.random.default_rng(3)
    for n, k, dim in [(1, 3, 2), (2, 3, 2), (400, 3, 2), (3000, 5, 7)]:
        X = rng.normal(0, 1, (n, dim)) * rng.uniform(1, 1e6, dim) + rng.uniform(-5, 5, dim)
        X[rng.random((n, dim)) < 0.1] = 0.0                     # zeros: the summarizer treats them apart [U]
        sd, sd2 = oracle.column_std(X), onp.column_std(X)
        assert np.max(np.abs(sd - sd2)) <= 1e-12 * np.max(np.abs(sd2)) or n == 1
        Xs = oracle.scale(X, sd)
        assert np.max(np.abs(Xs - onp.scale(X, sd))) < 1e-15 * max(1.0, np.max(np.abs(Xs)))
        c, w = rng.normal(0, 1, (k, dim)), rng.uniform(0, 5, k)
        for t in range(3):
            c1, w1 = oracle.kmeans_update(c, w, Xs, 0.87, t % 2)
            c2, w2 = onp.kmeans_update(c, w, Xs, 0.87, t % 2)
            assert np.max(np.abs(c1 - c2)) < 1e-12 and np.max(np.abs(w1 - w2)) < 1e-12
            c, w = c1, w1
        assert np.array_equal(oracle.kmeans_predict(c, Xs), onp.kmeans_predict(c, Xs))
    assert oracle.column_std(np.array([[3.0, 4.0]])).tolist() == [0.0, 0.0]   # one point: variance 0, scaled to 0.0
