"""Exact finite-size results for the 2D Ising model on an m x n torus (Kaufman 1949).
TEST INFRASTRUCTURE ONLY -- reference values for the physics tests.

Convention of examples/ising2D.py: E = J * sum_bonds s_i s_j with J = -1 (ferromagnet), K = -J/kT.
"""
import numpy as np


def log_z(m, n, K):
    """ln Z of the m x n periodic lattice at coupling K = |J|/kT."""
    l = np.arange(2 * n)
    c = np.cosh(2 * K) / np.tanh(2 * K) - np.cos(l * np.pi / n)
    gamma = np.arccosh(np.maximum(c, 1.0))
    gamma[0] = 2 * K + np.log(np.tanh(K))
    odd, even = gamma[1::2], gamma[0::2]
    # work with logarithms of the four products; Z4 can be negative (gamma_0 < 0 above T_c)
    lz1 = np.sum(np.log(2 * np.cosh(m * odd / 2)))
    lz2 = np.sum(np.log(2 * np.sinh(m * odd / 2)))
    lz3 = np.sum(np.log(2 * np.cosh(m * even / 2)))
    s4 = np.prod(np.sign(np.sinh(m * even / 2)))
    lz4 = np.sum(np.log(np.abs(2 * np.sinh(m * even / 2))))
    top = max(lz1, lz2, lz3, lz4)
    total = np.exp(lz1 - top) + np.exp(lz2 - top) + np.exp(lz3 - top) + s4 * np.exp(lz4 - top)
    return -np.log(2.0) + 0.5 * m * n * np.log(2 * np.sinh(2 * K)) + top + np.log(total)


def energy_and_cv_per_site(m, n, kT, J=-1.0, h=1e-4):
    """(<E>/N, C_v/N) with C_v = (<E^2> - <E>^2)/kT^2, by central differences of ln Z in K."""
    K = abs(J) / kT
    f0, fp, fm = log_z(m, n, K), log_z(m, n, K + h), log_z(m, n, K - h)
    d1 = (fp - fm) / (2 * h)          # d lnZ / dK = -<E>/|J|
    d2 = (fp - 2 * f0 + fm) / (h * h)  # d^2 lnZ / dK^2 = var(E)/J^2
    N = m * n
    return -abs(J) * d1 / N, d2 * K * K / N


def onsager_tc(J=-1.0):
    return 2.0 * abs(J) / np.log(1.0 + np.sqrt(2.0))
