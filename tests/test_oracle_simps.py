"""CPU: the scipy-1.2.1 simps restatement -- closed forms and weight-vector equivalence."""
import numpy as np

from oracle import simps121 as S


def test_odd_matches_textbook_simpson_exact_on_cubics():
    x = np.linspace(0, 2, 201)
    assert abs(S.simps(x ** 3, dx=x[1] - x[0]) - 4.0) < 1e-13
    assert abs(S.simps(x ** 3, x) - 4.0) < 1e-13


def test_even_avg_rule_known_answer():
    # y = x^2 on 4 points 0..3: first-panels 0..2 Simpson (8/3) + trapz(2..3)=6.5 -> 9.1666..; last: trapz(0..1)=.5 + Simpson 1..3 (26/3)
    y = np.arange(4.0) ** 2
    expect = 0.5 * ((8 / 3 + 6.5) + (0.5 + 26 / 3))
    assert abs(S.simps(y, dx=1.0) - expect) < 1e-14
    assert abs(S.simps(y, np.arange(4.0)) - expect) < 1e-14


def test_weights_reproduce_rule():
    rng = np.random.default_rng(0)
    for n in (5, 200, 201, 400, 2000, 9999, 10000):
        y = rng.random(n)
        x = np.cumsum(rng.random(n) + 0.5)
        assert abs(np.dot(S.simps_weights(n, 0.37), y) / S.simps(y, dx=0.37) - 1) < 1e-13
        assert abs(np.dot(S.simps_weights_x(x), y) / S.simps(y, x) - 1) < 1e-13
        assert abs(np.dot(S.trapz_weights_x(x), y) / np.trapezoid(y, x) - 1) < 1e-13


def test_axis_handling():
    rng = np.random.default_rng(1)
    y = rng.random((6, 7, 3))
    a = S.simps(y, dx=0.1, axis=0)
    b = np.tensordot(S.simps_weights(6, 0.1), y, axes=(0, 0))
    assert np.allclose(a, b, rtol=1e-13)
    a = S.simps(y, np.linspace(0, 1, 7), axis=1)
    b = np.einsum('j,ijk->ik', S.simps_weights_x(np.linspace(0, 1, 7)), y)
    assert np.allclose(a, b, rtol=1e-13)
