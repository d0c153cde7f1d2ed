"""Problem data and synthetic workloads for the configurations BASELINE.json names.

Numbers are fixtures taken from the reference's examples (paths relative to /root/reference/):

* cartpole   — examples/cartpole_example_one_solve.jl:11-18 (A, B, Q, R, rho = 1)
* quadrotor  — tinympc/TinyMPC/examples/problem_data/quadrotor_20hz_params.hpp:5-37
               (row-major arrays there; stored row-major here too and used as numpy 2-D arrays)
* rocket     — tinympc/TinyMPC/examples/problem_data/rocket_landing_params_20hz.hpp:5-29,
               bounds/cones from examples/rocket_landing_constraints.jl:38-60

The workload generators follow SURVEY.md §8d exactly (seeds, distributions) so that the CUDA
path, the oracle and the reference all see the same inputs.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional

import numpy as np


@dataclass
class System:
    name: str
    A: np.ndarray
    B: np.ndarray
    f: np.ndarray
    Q: np.ndarray          # full nx x nx (diagonal)
    R: np.ndarray
    rho: float
    x_min: Optional[np.ndarray] = None   # per-state bound vectors (replicated over the horizon)
    x_max: Optional[np.ndarray] = None
    u_min: Optional[np.ndarray] = None
    u_max: Optional[np.ndarray] = None
    extra: dict = field(default_factory=dict)

    @property
    def nx(self) -> int:
        return self.A.shape[0]

    @property
    def nu(self) -> int:
        return self.B.shape[1]

    def bounds(self, N: int):
        """(x_min, x_max, u_min, u_max) as nx x N / nu x (N-1) Fortran-ordered matrices."""
        rep = lambda v, k: np.asfortranarray(np.repeat(np.asarray(v, dtype=np.float64).reshape(-1, 1), k, axis=1))
        return rep(self.x_min, N), rep(self.x_max, N), rep(self.u_min, N - 1), rep(self.u_max, N - 1)


def cartpole() -> System:
    A = np.array([[1.0, 0.01, 0.0, 0.0],
                  [0.0, 1.0, 0.039, 0.0],
                  [0.0, 0.0, 1.002, 0.01],
                  [0.0, 0.0, 0.458, 1.002]])
    B = np.array([[0.0], [0.02], [0.0], [0.067]])
    # bounds: examples/cartpole_example_reference_constrained.jl:16-18
    return System("cartpole", A, B, np.zeros(4), np.diag([10.0, 1.0, 10.0, 1.0]), np.diag([1.0]), 1.0,
                  x_min=np.array([-2.0, -np.inf, -np.inf, -np.inf]), x_max=np.array([2.0, np.inf, np.inf, np.inf]),
                  u_min=np.array([-5.0]), u_max=np.array([5.0]))


_QUAD_A = [
    1.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0245250, 0.0000000, 0.0500000, 0.0000000, 0.0000000, 0.0000000, 0.0002044, 0.0000000,
    0.0000000, 1.0000000, 0.0000000, -0.0245250, 0.0000000, 0.0000000, 0.0000000, 0.0500000, 0.0000000, -0.0002044, 0.0000000, 0.0000000,
    0.0000000, 0.0000000, 1.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0500000, 0.0000000, 0.0000000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0250000, 0.0000000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0250000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0250000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.9810000, 0.0000000, 1.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0122625, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, -0.9810000, 0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000, -0.0122625, 0.0000000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000, 0.0000000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 1.0000000, 0.0000000,
    0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 0.0000000, 1.0000000]
_QUAD_B = [
    -0.0007069, 0.0007773, 0.0007091, -0.0007795,
    0.0007034, 0.0007747, -0.0007042, -0.0007739,
    0.0052554, 0.0052554, 0.0052554, 0.0052554,
    -0.1720966, -0.1895213, 0.1722891, 0.1893288,
    -0.1729419, 0.1901740, 0.1734809, -0.1907131,
    0.0123423, -0.0045148, -0.0174024, 0.0095748,
    -0.0565520, 0.0621869, 0.0567283, -0.0623632,
    0.0562756, 0.0619735, -0.0563386, -0.0619105,
    0.2102143, 0.2102143, 0.2102143, 0.2102143,
    -13.7677303, -15.1617018, 13.7831318, 15.1463003,
    -13.8353509, 15.2139209, 13.8784751, -15.2570451,
    0.9873856, -0.3611820, -1.3921880, 0.7659845]
_QUAD_Q = [100.0, 100.0, 100.0, 4.0, 4.0, 400.0, 4.0, 4.0, 4.0, 2.0408163, 2.0408163, 4.0]
_QUAD_R = [4.0, 4.0, 4.0, 4.0]


def quadrotor() -> System:
    """Quadrotor hover at 20 Hz; bounds u in +-0.5, x in +-5 (quadrotor_hovering.cpp:41-44)."""
    A = np.array(_QUAD_A).reshape(12, 12)
    B = np.array(_QUAD_B).reshape(12, 4)
    return System("quadrotor", A, B, np.zeros(12), np.diag(_QUAD_Q), np.diag(_QUAD_R), 5.0,
                  x_min=np.full(12, -5.0), x_max=np.full(12, 5.0), u_min=np.full(4, -0.5), u_max=np.full(4, 0.5))


def rocket() -> System:
    A = np.eye(6)
    A[0, 3] = A[1, 4] = A[2, 5] = 0.05
    B = np.zeros((6, 3))
    B[0, 0] = B[1, 1] = B[2, 2] = 0.000125
    B[3, 0] = B[4, 1] = B[5, 2] = 0.005
    f = np.array([0.0, 0.0, -0.0122625, 0.0, 0.0, -0.4905])
    s = System("rocket", A, B, f, np.diag([101.0] * 6), np.diag([2.0] * 3), 1.0,
               x_min=np.array([-5.0, -5.0, -0.5, -10.0, -10.0, -20.0]), x_max=np.array([5.0, 5.0, 100.0, 10.0, 10.0, 20.0]),
               u_min=np.full(3, -10.0), u_max=np.full(3, 105.0))
    # cone triples as the Julia example passes them: (Acu, qcu, cu) first, (Acx, qcx, cx) second
    s.extra.update(Acu=[0], qcu=[3], cu=[0.25], Acx=[0], qcx=[3], cx=[0.5],
                   xinit=np.array([4.0, 2.0, 20.0, -3.0, 2.0, -4.5]))
    return s


# ----------------------------------------------------------------------------- synthetic workloads

def cartpole_batch_x0(B: int, seed: int = 2) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return np.asfortranarray(rng.uniform(-1.0, 1.0, size=(4, B)) * np.array([1.0, 0.5, 0.2, 0.5]).reshape(4, 1))


def quadrotor_batch_x0(B: int, hard: bool = False, seed: int = 3) -> np.ndarray:
    """easy: x0 ~ U(+-0.1)^12 (5-9 iterations, all solved); hard: U(+-0.5)^12 (about half hit max_iter)."""
    rng = np.random.default_rng(seed)
    amp = 0.5 if hard else 0.1
    return np.asfortranarray(rng.uniform(-amp, amp, size=(12, B)))


def rocket_batch(B: int, N: int = 50, seed: int = 4):
    """x0 = 1.1 xinit (1 + U(+-0.1)^6); Xref = line xinit -> 0 over 99 steps (first N columns); Uref[2,:] = 10."""
    rng = np.random.default_rng(seed)
    xinit = rocket().extra["xinit"]
    x0 = np.asfortranarray((1.1 * xinit).reshape(6, 1) * (1.0 + rng.uniform(-0.1, 0.1, size=(6, B))))
    Xref = np.zeros((6, N), order="F")
    for i in range(N):
        Xref[:, i] = xinit + (0.0 - xinit) * i / 99.0
    Uref = np.zeros((3, N - 1), order="F")
    Uref[2, :] = 10.0
    return x0, Xref, Uref


def rocket_linear_constraints():
    """Synthetic half-spaces of config 4: -z <= 0 on the state, sum(u) <= 150 on the input."""
    Ax = np.zeros((1, 6), order="F"); Ax[0, 2] = -1.0
    Au = np.ones((1, 3), order="F")
    return Ax, np.array([0.0]), Au, np.array([150.0])


def quadrotor_line_refs(B: int, N: int, nsteps: int, seed: int = 5):
    """Config 5: x0 ~ U(+-0.1)^12 and, per instance, a straight line in y at speed v_y ~ U(0, 0.5) m/s
    (z = 1, ydot = v_y, 20 Hz), laid out like trajectory_data/quadrotor_20hz_y_axis_line.hpp.
    Returns x0 (12, B) and Xref_long (12, N + nsteps - 1, B)."""
    rng = np.random.default_rng(seed)
    x0 = np.asfortranarray(rng.uniform(-0.1, 0.1, size=(12, B)))
    vy = rng.uniform(0.0, 0.5, size=B)
    T = N + nsteps - 1
    t = np.arange(T) * 0.05
    X = np.zeros((12, T, B), order="F")
    X[1, :, :] = t.reshape(T, 1) * vy.reshape(1, B)
    X[2, :, :] = 1.0
    X[7, :, :] = vy.reshape(1, B)
    return x0, X


# FLOP model for the roofline: per ADMM iteration per instance, box constraints.
def flops_per_iteration(nx: int, nu: int, N: int) -> dict:
    M = (N - 1) * (2 * nx * nx + 4 * nx * nu + nu * nu) + nx * nx
    V = (N - 1) * (5 * nx + 3 * nu)
    E = 13 * (nx * N + nu * (N - 1))
    return {"M": M, "V": V, "E": E, "F": 2 * M + V + E}
