"""The training restatement (oracle/train_oracle.py) against first principles: finite differences,
the reference's label / learning-rate rules.  No GPU."""
import numpy as np
import torch

from oracle import train_oracle as to


def _batch(rng, B, D, p=9):
    return (rng.standard_normal((B, p, p, 3)).astype(np.float32),
            rng.standard_normal((B, p, p + D - 1, 3)).astype(np.float32), to.soft_labels(B, D))


def test_layout_counts_match_keras():
    layers, n_p, n_s = to.param_layout("win9_dep9")
    assert n_p == 113088 and n_s == 8 * 64                      # 113 088 fp32 gradients
    _, n_p37, _ = to.param_layout("win37_dep9")
    assert n_p37 == 617280 + 1920 // 2                          # conv params + gamma/beta


def test_soft_labels_and_schedule():
    t = to.soft_labels(3, 201)
    assert t.shape == (3, 201) and np.allclose(t.sum(axis=1), 1.0)
    assert list(np.nonzero(t[0])[0]) == [98, 99, 100, 101, 102] and t[0, 100] == 0.5       # data_handler.py:78-85
    assert to.learning_rate(1) == 1e-3 and to.learning_rate(23999) == 1e-3
    assert np.isclose(to.learning_rate(24000), 2e-4) and np.isclose(to.learning_rate(31999), 2e-4)
    assert np.isclose(to.learning_rate(32000), 4e-5) and np.isclose(to.learning_rate(40000), 8e-6)   # train.py:125-130


def test_gradient_matches_finite_differences():
    rng = np.random.default_rng(0)
    B, D = 4, 7
    params, stats = to.init_params("win9_dep9", seed=1)
    params = params.astype(np.float64)
    params += 0.05 * rng.standard_normal(params.size)           # biases / beta away from zero
    left, right, targets = _batch(rng, B, D)
    loss, correct, grads, logits = to.forward_backward(params, stats.copy(), left, right, targets)
    assert logits.shape == (B, D) and 0 <= correct <= B and np.isfinite(loss)
    layers, _, _ = to.param_layout("win9_dep9")
    probe = [layers[0]["kernel"].start + 5, layers[1]["kernel"].start + 1234, layers[2]["bias"].start + 3,
             layers[3]["gamma"].start + 7, layers[0]["beta"].start + 11, layers[3]["kernel"].start + 999]
    h = 1e-5
    for i in probe:
        pp, pm = params.copy(), params.copy()
        pp[i] += h
        pm[i] -= h
        lp = to.forward_backward(pp, stats.copy(), left, right, targets)[0]
        lm = to.forward_backward(pm, stats.copy(), left, right, targets)[0]
        fd = (lp - lm) / (2 * h)
        assert abs(fd - grads[i]) <= 1e-6 + 1e-4 * abs(fd), (i, fd, grads[i])


def test_moving_statistics_follow_keras_rule():
    rng = np.random.default_rng(2)
    params, stats = to.init_params("win9_dep9", seed=3)
    left, right, targets = _batch(rng, 3, 5)
    before = stats.copy()
    to.forward_backward(params, stats, left, right, targets)
    layers, _, _ = to.param_layout("win9_dep9")
    # two tower passes: mean moves towards the batch means by (1 - 0.99) twice; variance stays positive
    assert not np.array_equal(before, stats) and np.all(stats[layers[0]["var"]] > 0)
    x = torch.tensor(left, dtype=torch.float64).permute(0, 3, 1, 2)
    w = torch.tensor(params[layers[0]["kernel"]].reshape(3, 3, 3, 64), dtype=torch.float64).permute(3, 2, 0, 1)
    y1 = torch.relu(torch.nn.functional.conv2d(x, w, torch.tensor(params[layers[0]["bias"]], dtype=torch.float64)))
    xr = torch.tensor(right, dtype=torch.float64).permute(0, 3, 1, 2)
    y2 = torch.relu(torch.nn.functional.conv2d(xr, w, torch.tensor(params[layers[0]["bias"]], dtype=torch.float64)))
    m = before[layers[0]["mean"]].astype(np.float64)
    for y in (y1, y2):
        m = m * 0.99 + y.mean(dim=(0, 2, 3)).numpy() * 0.01
    assert np.allclose(stats[layers[0]["mean"]], m, rtol=1e-5, atol=1e-7)


def test_adam_first_step_is_lr_sized():
    g = np.array([0.3, -2.0, 5.0])
    p, m, v = to.adam_step(np.zeros(3), g, np.zeros(3), np.zeros(3), 1, 1e-3)
    assert np.allclose(p, -1e-3 * np.sign(g), rtol=1e-3)        # |update| = lr on the first step
