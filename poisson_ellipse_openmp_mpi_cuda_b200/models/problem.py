"""The model problem: -div(k grad u) = F on [-1,1] x [-0.6,0.6], k = 1 in the ellipse D, 1/eps outside.

Python view of csrc/core/problem.h (constants S4:11-13, membership S4:32-34, chord length S4:46-76,
coefficients S4:162-177, right-hand side S4:181-191, diagonal S4:558-561).  The scalar functions call the
native library, so they are the same code the CPU oracle and the device assembly kernel run.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from ..ops import _native


@dataclass(frozen=True)
class EllipseProblem:
    M: int
    N: int
    A1: float = -1.0
    B1: float = 1.0
    A2: float = -0.6
    B2: float = 0.6
    F: float = 1.0

    @property
    def h1(self) -> float:
        return (self.B1 - self.A1) / self.M

    @property
    def h2(self) -> float:
        return (self.B2 - self.A2) / self.N

    @property
    def eps(self) -> float:
        h = max(self.h1, self.h2)
        return h * h

    @property
    def default_max_iter(self) -> int:
        return (self.M - 1) * (self.N - 1)

    def x(self, i: int) -> float:
        return self.A1 + i * self.h1

    def y(self, j: int) -> float:
        return self.A2 + j * self.h2

    @staticmethod
    def in_domain(x: float, y: float) -> bool:
        return x * x + 4.0 * y * y < 1.0

    def coef_a(self, i: int, j: int) -> float:
        return _native.lib().pb_coef_a(self.M, self.N, i, j)

    def coef_b(self, i: int, j: int) -> float:
        return _native.lib().pb_coef_b(self.M, self.N, i, j)

    def rhs(self, i: int, j: int) -> float:
        return _native.lib().pb_rhs(self.M, self.N, i, j)

    def diag(self, i: int, j: int) -> float:
        return _native.lib().pb_diag(self.M, self.N, i, j)

    @staticmethod
    def chord(fixed: float, lo: float, hi: float, vertical: bool) -> float:
        return _native.lib().pb_chord(fixed, lo, hi, int(vertical))

    def exact(self) -> torch.Tensor:
        """Analytic solution u = (1 - x^2 - 4y^2)/10 in D, 0 outside, on the (M+1) x (N+1) nodes."""
        x = self.A1 + torch.arange(self.M + 1, dtype=torch.float64) * self.h1
        y = self.A2 + torch.arange(self.N + 1, dtype=torch.float64) * self.h2
        s = x[:, None] ** 2 + 4.0 * y[None, :] ** 2
        return torch.where(s < 1.0, (1.0 - s) / 10.0, torch.zeros_like(s))
