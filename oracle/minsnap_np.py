"""minsnap_np — TEST INFRASTRUCTURE.  numpy restatement of the reference's linear min-snap solve for the
vertex pattern poly_traj::generateTrajectory builds (external/poly_traj/src/trajectory_generator.cpp:12-58):
start vertex = position + velocity + acceleration given, jerk and snap zero (makeStartOrEnd, vertex.cpp:146-170),
interior vertices = position only, end vertex = position given, all derivatives zero; N = 10 coefficients,
derivative_to_optimize = SNAP.

Follows external/poly_traj/include/mav_trajectory_generation/impl/polynomial_optimization_linear_impl.h in the
reference's OWN (unscaled) coordinates, so that it checks the scaled formulation of the CUDA kernel independently:
  setupMappingMatrix (:111-121), invertMappingMatrix (:142-179, Schur complement), computeQuadraticCostJacobian
  (:567-583), constructR (:307-336), setupConstraintReorderingMatrix (:181-260: fixed constraints first, then
  free), solveLinear (:338-380: dp = -Rpp^-1 Rpf df), updateSegmentsFromCompactConstraints (:262-283).
The reference solves with Eigen's SparseQR; here np.linalg.solve.  Parity of this row is by tolerance, not bits:
the solver itself (Eigen) cannot be built in this image — "parity unpinned" for §8(f) N3, stated in DESIGN.md.
"""
from math import factorial

import numpy as np

N, HALF, SNAP = 10, 5, 4


def base_coefficient(r, j):
    """Polynomial::base_coefficients_(r, j) (polynomial.cpp:145-160): j! / (j - r)!."""
    return float(factorial(j) // factorial(j - r)) if j >= r else 0.0


def mapping_matrix(T):
    A = np.zeros((N, N))
    for r in range(HALF):
        A[r, r] = base_coefficient(r, r)                       # baseCoeffsWithTime(N, r, 0.0)
        for j in range(r, N):
            A[HALF + r, j] = base_coefficient(r, j) * T ** (j - r)
    return A


def inverse_mapping_matrix(A):
    Ad = np.diag(A)[:HALF]
    A_inv = np.diag(1.0 / Ad)
    C = A[HALF:, :HALF]
    D_inv = np.linalg.inv(A[HALF:, HALF:])
    out = np.zeros((N, N))
    out[:HALF, :HALF] = A_inv
    out[HALF:, :HALF] = -D_inv @ C @ A_inv
    out[HALF:, HALF:] = D_inv
    return out


def cost_jacobian(T, derivative=SNAP):
    Q = np.zeros((N, N))
    for col in range(N - derivative):
        for row in range(N - derivative):
            exponent = (N - 1 - derivative) * 2 + 1 - row - col
            Q[N - 1 - row, N - 1 - col] = (base_coefficient(derivative, N - 1 - row) *
                                           base_coefficient(derivative, N - 1 - col) * T ** exponent * 2.0 / exponent)
    return Q


def solve(waypoints, durations, start_vel=None, start_acc=None):
    """waypoints (K+1, 3), durations (K,) -> coeffs (K, 3, 10), increasing powers."""
    wp = np.asarray(waypoints, dtype=np.float64)
    T = np.asarray(durations, dtype=np.float64)
    K = T.shape[0]
    D = wp.shape[1]
    v0 = np.zeros(D) if start_vel is None else np.asarray(start_vel, dtype=np.float64)
    a0 = np.zeros(D) if start_acc is None else np.asarray(start_acc, dtype=np.float64)

    # variables = (vertex, derivative); fixed ones first, then free ones (std::set order: vertex, then derivative)
    fixed, free = {}, []
    for v in range(K + 1):
        for r in range(HALF):
            if r == 0:
                fixed[(v, r)] = wp[v]
            elif v == 0:
                fixed[(v, r)] = v0 if r == 1 else (a0 if r == 2 else np.zeros(D))
            elif v == K:
                fixed[(v, r)] = np.zeros(D)
            else:
                free.append((v, r))
    fixed_keys = sorted(fixed)
    index = {k: i for i, k in enumerate(fixed_keys)}
    nf = len(fixed_keys)
    index.update({k: nf + i for i, k in enumerate(free)})
    n_all = nf + len(free)

    A_inv = [inverse_mapping_matrix(mapping_matrix(t)) for t in T]
    M = np.zeros((N * K, n_all))                                 # constraint_reordering_
    for k in range(K):
        for row in range(N):
            M[k * N + row, index[(k + row // HALF, row % HALF)]] = 1.0
    Hbig = np.zeros((N * K, N * K))
    for k in range(K):
        Hbig[k * N:(k + 1) * N, k * N:(k + 1) * N] = A_inv[k].T @ cost_jacobian(T[k]) @ A_inv[k]
    R = M.T @ Hbig @ M
    Rpf, Rpp = R[nf:, :nf], R[nf:, nf:]
    df = np.stack([fixed[k] for k in fixed_keys])                # (nf, D)
    dp = np.linalg.solve(Rpp, -Rpf @ df) if free else np.zeros((0, D))
    d_all = np.vstack([df, dp])
    coeffs = np.zeros((K, D, N))
    for k in range(K):
        new_d = M[k * N:(k + 1) * N] @ d_all                     # (N, D)
        coeffs[k] = (A_inv[k] @ new_d).T
    return coeffs


def evaluate(coeffs_kd, t, derivative=0):
    """Value of one polynomial (increasing powers) or its derivative at t."""
    c = np.asarray(coeffs_kd, dtype=np.float64)
    acc = 0.0
    for j in range(N - 1, derivative - 1, -1):
        acc = acc * t + base_coefficient(derivative, j) * c[j]
    return acc
