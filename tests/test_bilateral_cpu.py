"""CPU checks of the bilateral-row restatement (oracle/np_oracle.py): MINRES converges to the solution of an SPD system
and the constrained velocities satisfy the rows."""
import numpy as np

from oracle import np_oracle as no


def test_minres_solves_spd_system():
    rng = np.random.default_rng(0)
    A = rng.normal(size=(12, 12))
    N = A @ A.T + 0.5 * np.eye(12)
    b = rng.normal(size=12)
    x, it, res = no.minres(N, b, 200, 1e-13)
    assert res < 1e-13 and it < 40
    assert np.allclose(N @ x, b, rtol=0, atol=1e-10)


def test_rows_hold_after_the_solve():
    rng = np.random.default_rng(1)
    nb, n = 7, 5
    mass, inv_in = rng.uniform(1, 2, nb), rng.uniform(5, 9, (nb, 3))
    active = np.ones(nb, dtype=bool)
    active[0] = False
    ids = np.array([[0, 1], [1, 2], [2, 3], [0, 4], [5, 6]])
    cq_a, cq_b, rhs = rng.normal(size=(n, 6)), rng.normal(size=(n, 6)), rng.normal(size=n)
    vel, hf = rng.normal(size=(nb, 6)), rng.normal(size=(nb, 6)) * 0.1
    vel[0] = 0
    v, gamma, it, res = no.bilateral_solve(ids, cq_a, cq_b, rhs, vel, hf, mass, inv_in, active, 200, 1e-14)
    viol = np.array([cq_a[i] @ v[ids[i, 0]] + cq_b[i] @ v[ids[i, 1]] + rhs[i] for i in range(n)])
    assert np.abs(viol).max() < 1e-10
    assert np.array_equal(v[0], vel[0])  # an inactive body has no rows in M^-1
