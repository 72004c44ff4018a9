"""MLlib 1.6.1's own acceptance test for StreamingLinearRegressionWithSGD (restated in
tests/upstream_cases.py), run against the oracle (CPU) and against the CUDA path (GPU)."""
import numpy as np
import pytest

from upstream_cases import JavaRandom, parameter_accuracy_batches


def test_java_random_known_answers():
    # java.util.Random(42).nextInt() == -1170105035, nextDouble of a fresh Random(42) == 0.7275636800328681 (well-known values)
    r = JavaRandom(42)
    assert r._next(32) == -1170105035
    assert JavaRandom(42).next_double() == 0.7275636800328681
    assert abs(JavaRandom(42).next_gaussian() - 1.1419053154730547) < 1e-15


def _check(w, batches, predict):
    assert abs(w[0] - 10.0) < 0.1 and abs(w[1] - 10.0) < 0.1          # upstream: ~== 10.0 absTol 0.1
    X, y = batches[0]
    assert np.mean(np.abs(predict(X) - y)) < 0.1                      # upstream: error < 0.1


def test_parameter_accuracy_oracle(oracle):
    batches = parameter_accuracy_batches()
    w = np.zeros(2)
    for X, y in batches:
        w, _ = oracle.sgd(y, 2, w, X=X, step_size=0.2, num_iterations=25, convergence_tol=1e-4)
    _check(w, batches, lambda X: X @ w)


@pytest.mark.gpu
def test_parameter_accuracy_gpu(oracle):
    import twtml_b200 as tw
    batches = parameter_accuracy_batches()
    with tw.Context(0) as ctx:
        model = (tw.StreamingLinearRegressionWithSGD(ctx=ctx).setStepSize(0.2).setNumIterations(25).setConvergenceTol(1e-4)
                 .setInitialWeights(tw.Vectors.dense(0.0, 0.0)))
        stream = tw.DStream()
        model.trainOn(stream)
        w_ref = np.zeros(2)
        for X, y in batches:
            stream.push([tw.LabeledPoint(t, tw.Vectors.dense(x)) for x, t in zip(X, y)])
            w_ref, _ = oracle.sgd(y, 2, w_ref, X=X, step_size=0.2, num_iterations=25, convergence_tol=1e-4)
        w = model.latestModel().weights.toArray()
        _check(w, batches, lambda X: model.latestModel().predict(ctx.batch_from_dense(X, np.zeros(len(X)))))
        assert np.max(np.abs(w - w_ref)) < 1e-6 * np.max(np.abs(w_ref))
