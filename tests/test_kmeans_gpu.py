"""GPU parity tests of the k-means app (KMeans.scala:69-113 -> twtml_kmeans_*), through the C ABI: the hand-derived
interval, random intervals against the oracle (centres / weights / deviations to 1e-9 relative -- observed ~1e-15 --,
assignments exact), and the reference's own loop (StandardScaler.fit.transform, latestModel.update, predict) replayed
through the Python mirror over several intervals."""
import numpy as np
import pytest

from conftest import golden

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tw():
    import twtml_b200
    return twtml_b200


@pytest.fixture(scope="module")
def ctx(tw):
    c = tw.Context(0)
    yield c
    c.close()


def test_kmeans_golden_interval(tw, ctx):
    g = golden("kmeans_kat.json")
    X = np.array(g["points"], float)
    for unit, ck, wk in (("batches", "centers_after", "weights_after"), ("points", "centers_after_points", "weights_after_points")):
        km = tw.StreamingKMeans(ctx=ctx).setK(2).setInitialCenters(g["centers"], g["weights"])
        assert km.latestModel().predict(X).tolist() == g["closest_before"]
        m = km.latestModel().update(X, g["decay"], unit)
        assert np.allclose(m.clusterCenters, g[ck], rtol=1e-14, atol=0) and m.clusterWeights.tolist() == g[wk]
        assert m.predict(X).tolist() == g["closest_after"]
        km.free()
    d = g["dying"]
    km = tw.StreamingKMeans(ctx=ctx).setK(2).setInitialCenters(d["centers"], d["weights"])
    m = km.latestModel().update(X, g["decay"], "batches")
    c = m.clusterCenters
    assert m.clusterWeights.tolist() == d["weights_after"] and np.max(np.abs(c - np.array(d["center_mid"]))) < 1e-13
    assert c[0][0] > c[1][0]
    scaled = tw.StandardScaler(False, True).fit(X).transform(X)
    km2 = tw.StreamingKMeans(ctx=ctx).setK(2).setInitialCenters(g["centers"], g["weights"])
    km2.latestModel().update(scaled, 1.0, "batches")
    assert np.allclose(scaled.std, g["column_std"], rtol=1e-14, atol=0)
    assert np.allclose(scaled.toArray(), X * (1.0 / np.array(g["column_std"])), rtol=1e-14, atol=0)


@pytest.mark.parametrize("n,k,dim", [(1, 3, 2), (2, 3, 2), (37, 3, 2), (5000, 3, 2), (200000, 3, 2), (3000, 7, 16), (1000, 16, 5)])
def test_kmeans_intervals_match_oracle(tw, ctx, oracle, n, k, dim):
    rng = np.random.default_rng(n + k)
    c, w = rng.normal(0, 1, (k, dim)), np.zeros(k)                 # setRandomCenters(dim, 0.0): weights start at 0
    km = tw.StreamingKMeans(ctx=ctx).setK(k).setHalfLife(5, "batches").setInitialCenters(c, w)
    for t in range(4):
        X = np.floor(np.exp(rng.uniform(0, 12, (n, dim))))          # retweet / follower counts: log-uniform integers
        if t == 2:
            X[:, 0] = 7.0                                           # a constant column: deviation 0, scaled to 0.0
        scaled = tw.StandardScaler(False, True).fit(X).transform(X)
        # every interval starts from the oracle's state, bit for bit: once a dying cluster has been split, two centres differ
        # by 1e-14 (MLlib's own perturbation) and a last-bit difference between the two implementations' centres would hand
        # the points between the twins to different sides
        km.setInitialCenters(c, w)
        m = km.latestModel().update(scaled, km.decayFactor, km.timeUnit)
        sd = oracle.column_std(X)
        Xs = oracle.scale(X, sd)
        assert np.max(np.abs(scaled.std - sd)) <= 1e-9 * max(1e-300, np.max(np.abs(sd)))
        assert np.max(np.abs(scaled.toArray() - Xs)) <= 1e-9 * max(1.0, np.max(np.abs(Xs)))
        # the update is checked on the points the GPU scaled: a last-bit difference in a deviation may move a point that sits
        # on a cluster boundary (the counts are integers with many duplicates) to the other side, and with it the centres
        c, w = oracle.kmeans_update(c, w, scaled.toArray(), km.decayFactor, 0)
        assert np.max(np.abs(m.clusterCenters - c)) <= 1e-9 * max(1.0, np.max(np.abs(c)))
        assert np.max(np.abs(m.clusterWeights - w)) <= 1e-9 * max(1.0, np.max(np.abs(w)))
        pred = m.predict(scaled)
        ref = oracle.kmeans_predict(m.clusterCenters, scaled.toArray())     # same centres, same points: the same labels
        assert np.array_equal(pred, ref)
    km.free()


def test_kmeans_api_errors_and_streams(tw, ctx):
    km = tw.StreamingKMeans(ctx=ctx).setK(3)
    with pytest.raises(ValueError):
        km.latestModel()                                           # assertInitialized [U]
    km.setRandomCenters(2, 0.0, seed=5)
    c0 = km.latestModel().clusterCenters
    assert c0.shape == (3, 2) and np.all(np.isfinite(c0)) and np.all(km.latestModel().clusterWeights == 0.0)
    km.latestModel().update(np.zeros((0, 2)), 0.5, "batches")      # `if (rdd.count > 0)`: an empty interval changes nothing
    assert np.array_equal(km.latestModel().clusterCenters, c0)
    stream = tw.DStream()
    km.trainOn(stream)
    labels = []
    km.predictOn(stream).foreachRDD(lambda p: labels.append(p))
    rng = np.random.default_rng(0)
    for _ in range(3):
        stream.push(rng.normal(0, 1, (500, 2)) + rng.choice([-4, 0, 4], (500, 1)))
    assert len(labels) == 3 and all(len(l) == 500 and l.min() >= 0 and l.max() <= 2 for l in labels)
    assert km.latestModel().clusterWeights.sum() > 0
    with pytest.raises(tw.TwtmlError):
        tw.StreamingKMeans(ctx=ctx).setK(40).setRandomCenters(2, 0.0)   # k beyond the library's limit is refused
    km.free()
