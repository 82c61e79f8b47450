"""Optimizers follow the Keras 2.2 update rules, checked against independent NumPy formulas."""
import numpy as np
import pytest

from cori_intml_examples_b200.nn import optimizers as O

EPS = 1e-7


def run(opt, steps=4, seed=0):
    rng = np.random.RandomState(seed)
    p = [rng.randn(5, 3).astype(np.float32), rng.randn(7).astype(np.float32)]
    slots = opt.init_slots(p)
    grads = [[rng.randn(*a.shape).astype(np.float32) for a in p] for _ in range(steps)]
    p0 = [a.copy() for a in p]
    for g in grads:
        opt.apply(p, g, slots)
    return p0, grads, p


def test_adam_matches_formula():
    p0, grads, p = run(O.Adam(lr=1e-3))
    for i in range(2):
        w = p0[i].astype(np.float64); m = np.zeros_like(w); v = np.zeros_like(w)
        for t, gs in enumerate(grads, 1):
            g = gs[i].astype(np.float64)
            lr_t = 1e-3 * np.sqrt(1 - 0.999 ** t) / (1 - 0.9 ** t)
            m = 0.9 * m + 0.1 * g; v = 0.999 * v + 0.001 * g * g
            w = w - lr_t * m / (np.sqrt(v) + EPS)
        np.testing.assert_allclose(p[i], w, rtol=2e-5, atol=1e-7)


def test_nadam_matches_formula():
    p0, grads, p = run(O.Nadam())
    for i in range(2):
        w = p0[i].astype(np.float64); m = np.zeros_like(w); v = np.zeros_like(w); sched = 1.0
        for t, gs in enumerate(grads, 1):
            g = gs[i].astype(np.float64)
            mu_t = 0.9 * (1 - 0.5 * 0.96 ** (t * 0.004)); mu_t1 = 0.9 * (1 - 0.5 * 0.96 ** ((t + 1) * 0.004))
            sched_new = sched * mu_t; sched_next = sched_new * mu_t1; sched = sched_new
            gp = g / (1 - sched_new)
            m = 0.9 * m + 0.1 * g; mp = m / (1 - sched_next)
            v = 0.999 * v + 0.001 * g * g; vp = v / (1 - 0.999 ** t)
            w = w - 0.002 * ((1 - mu_t) * gp + mu_t1 * mp) / (np.sqrt(vp) + EPS)
        np.testing.assert_allclose(p[i], w, rtol=2e-5, atol=1e-7)


def test_adadelta_matches_formula():
    p0, grads, p = run(O.Adadelta())
    for i in range(2):
        w = p0[i].astype(np.float64); a = np.zeros_like(w); d = np.zeros_like(w)
        for gs in grads:
            g = gs[i].astype(np.float64)
            a = 0.95 * a + 0.05 * g * g
            u = g * np.sqrt(d + EPS) / np.sqrt(a + EPS)
            w = w - 1.0 * u
            d = 0.95 * d + 0.05 * u * u
        np.testing.assert_allclose(p[i], w, rtol=1e-4, atol=1e-6)


def test_sgd_momentum():
    p0, grads, p = run(O.SGD(lr=0.1, momentum=0.9))
    w = p0[0].astype(np.float64); vel = np.zeros_like(w)
    for gs in grads:
        vel = 0.9 * vel - 0.1 * gs[0]; w = w + vel
    np.testing.assert_allclose(p[0], w, rtol=1e-5, atol=1e-6)


def test_lookup_by_name_and_defaults():
    assert isinstance(O.get("Adadelta"), O.Adadelta) and O.get("Adadelta").lr == 1.0
    assert O.get("Nadam").lr == 0.002 and O.get("adam").epsilon == EPS
    assert O.get("Nadam", lr=1e-3).lr == 1e-3            # rpv.py:62: getattr(optimizers, name)(lr=lr)
    inst = O.Adadelta(1.0 * 8)                          # DistTrain_mnist.ipynb:307
    assert O.get(inst) is inst and inst.lr == 8.0
    with pytest.raises(ValueError):
        O.get("RMSpropX")


def test_padding_is_a_fixed_point():
    """Zero parameters with zero gradients stay zero under every optimizer (the device layout pads channels)."""
    for opt in (O.SGD(momentum=0.9), O.Adam(), O.Nadam(), O.Adadelta()):
        p = [np.zeros(8, np.float32)]
        s = opt.init_slots(p)
        for _ in range(3):
            opt.apply(p, [np.zeros(8, np.float32)], s)
        assert not p[0].any()


def test_nadam_state_accepts_the_keras_layout_without_m_schedule():
    """Keras 2.2's Nadam saves [iterations] + ms + vs -- no m_schedule.  Loading such a list must rebuild the schedule from the step
    count (not pop a moment tensor as the schedule), and a malformed list must leave the optimizer untouched."""
    import numpy as np
    from cori_intml_examples_b200.nn import optimizers
    opt = optimizers.Nadam(lr=0.002)
    params = [np.ones((3, 2), np.float32), np.ones((2,), np.float32)]
    slots = [[np.zeros_like(p) for p in params] for _ in range(2)]
    for _ in range(5):
        opt.apply(params, [0.1 * np.ones_like(p) for p in params], slots)
    ours = opt.state_arrays(slots)
    keras_like = [ours[0]] + ours[2:]
    fresh = optimizers.Nadam(lr=0.002)
    fslots = [[np.zeros_like(p) for p in params] for _ in range(2)]
    fresh.load_state_arrays(keras_like, fslots)
    assert fresh.iterations == 5 and abs(fresh.m_schedule - opt.m_schedule) < 1e-9 * max(1.0, abs(opt.m_schedule))
    assert all(np.array_equal(a, b) for sa, sb in zip(fslots, slots) for a, b in zip(sa, sb))
    again = optimizers.Nadam(lr=0.002)
    aslots = [[np.zeros_like(p) for p in params] for _ in range(2)]
    again.load_state_arrays(ours, aslots)
    assert again.iterations == 5 and abs(again.m_schedule - opt.m_schedule) < 1e-6
    broken = optimizers.Nadam(lr=0.002)
    try:
        broken.load_state_arrays(ours[:-1][:3], aslots)
        raise AssertionError("a truncated state list must be rejected")
    except ValueError:
        pass
    assert broken.iterations == 0 and broken.m_schedule == 1.0
