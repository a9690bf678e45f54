"""Scalar problem definition: membership, chord lengths, coefficients (W1 of SURVEY.md section 8)."""
import math
import random

import pytest

from poisson_ellipse_openmp_mpi_cuda_b200.models import EllipseProblem


def ref_chord(fixed, lo, hi, vertical):
    if vertical:
        if abs(fixed) >= 1.0:
            return 0.0
        root = math.sqrt(max(0.0, (1.0 - fixed * fixed) / 4.0))
    else:
        if abs(2.0 * fixed) >= 1.0:
            return 0.0
        root = math.sqrt(max(0.0, 1.0 - 4.0 * fixed * fixed))
    return max(0.0, min(hi, root) - max(lo, -root))


def test_chord_random_and_edge_cases():
    rnd = random.Random(3)
    for _ in range(2000):
        f, a, b = rnd.uniform(-1.2, 1.2), rnd.uniform(-1, 1), rnd.uniform(-1, 1)
        lo, hi = min(a, b), max(a, b)
        for v in (True, False):
            assert EllipseProblem.chord(f, lo, hi, v) == ref_chord(f, lo, hi, v)
    # |x0| = 1 and |2 y0| = 1: outside by the >= test; tangent faces
    assert EllipseProblem.chord(1.0, -1, 1, True) == 0.0
    assert EllipseProblem.chord(-1.0, -1, 1, True) == 0.0
    assert EllipseProblem.chord(0.5, -1, 1, False) == 0.0
    assert EllipseProblem.chord(0.0, -1, 1, True) == 1.0       # full vertical chord through the centre: 2 * 0.5
    assert EllipseProblem.chord(0.0, -2, 2, False) == 2.0


def test_coefficients_three_kinds():
    p = EllipseProblem(40, 40)
    inv_eps = 1.0 / p.eps
    assert p.coef_a(20, 20) == 1.0 and p.coef_b(20, 20) == 1.0           # centre: faces fully inside
    assert p.coef_a(1, 1) == inv_eps and p.coef_b(1, 1) == inv_eps       # corner: fully outside
    blended = [p.coef_a(i, j) for i in range(1, 40) for j in range(1, 40)]
    mixed = [v for v in blended if v not in (1.0, inv_eps)]
    assert mixed and all(1.0 < v < inv_eps for v in mixed)
    assert p.rhs(20, 20) == 1.0 and p.rhs(1, 1) == 0.0
    assert p.diag(20, 20) == pytest.approx(2.0 / p.h1 ** 2 + 2.0 / p.h2 ** 2, rel=1e-15)


def test_membership_is_strict():
    assert EllipseProblem.in_domain(0.0, 0.0)
    assert not EllipseProblem.in_domain(1.0, 0.0)
    assert not EllipseProblem.in_domain(0.0, 0.5)
    assert EllipseProblem.in_domain(0.6, 0.39)


def test_exact_solution_field():
    p = EllipseProblem(20, 20)
    u = p.exact()
    assert u.shape == (21, 21)
    assert float(u.max()) == pytest.approx(0.1, abs=1e-12)
    assert float(u[0].abs().max()) == 0.0
