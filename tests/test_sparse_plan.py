"""CPU tier: the block-sparse Cholesky's PLAN (minimum-degree ordering with multiple elimination, symbolic
factorisation, left-looking schedule, assembly gather lists: splatt3r_slam_b200/csrc/gn_sparse_plan.h) executed on
the host through the C ABI (b200slam_gn_sparse_host_solve) against a dense fp64 solve of the system the reference
assembles (SparseBlock::update_lhs/update_rhs, gn_kernels.cu:70-112: fixed pose dropped, duplicates summed).  The
device kernel (gn_sparse.cuh) executes the same plan; tests/test_gpu_gn.py checks it on the GPU."""
import ctypes

import numpy as np
import pytest

import splatt3r_slam_b200.mast3r_slam_backends as be


def make(N, Eu, seed, ring=True):
    rng = np.random.default_rng(seed)
    und = [(k, (k + 1) % N) for k in range(min(N, Eu))] if ring else []
    while len(und) < Eu:
        a, b = rng.integers(0, N, 2)
        if a != b:
            und.append((int(a), int(b)))
    ii = np.array([a for a, _ in und] + [b for _, b in und], dtype=np.int64)
    jj = np.array([b for _, b in und] + [a for a, _ in und], dtype=np.int64)
    E = len(ii)
    eb = np.zeros((E, 56))
    for e in range(E):
        J = rng.standard_normal((30, 7)) * np.array([10, 10, 10, 3, 3, 3, 1.0])
        eb[e, :49] = (J.T @ J).ravel()
        eb[e, 49:] = J.T @ rng.standard_normal(30)
    return ii, jj, eb


def dense(ii, jj, eb, N):
    D = 7 * (N - 1)
    H, b = np.zeros((D, D)), np.zeros(D)
    for e in range(len(ii)):
        B, g, i, j = eb[e, :49].reshape(7, 7), eb[e, 49:], ii[e] - 1, jj[e] - 1
        for (r, c, s) in ((i, i, 1), (i, j, -1), (j, i, -1), (j, j, 1)):
            if r >= 0 and c >= 0:
                H[7 * r:7 * r + 7, 7 * c:7 * c + 7] += s * B
        if i >= 0:
            b[7 * i:7 * i + 7] -= g
        if j >= 0:
            b[7 * j:7 * j + 7] += g
    return H, b


def host_solve(ii, jj, eb, N):
    x = np.zeros(7 * (N - 1))
    st = (ctypes.c_longlong * 4)()
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    rc = be._lib.b200slam_gn_sparse_host_solve(p(ii), p(jj), p(eb), N, len(ii), p(x), st)
    return rc, x, list(st)


@pytest.mark.parametrize("N,Eu,ring", [(2, 1, True), (8, 8, True), (40, 60, True), (130, 130, True), (130, 150, True),
                                       (300, 330, True), (64, 200, False), (1024, 1100, True)])
def test_plan_solves_the_reference_system(N, Eu, ring):
    ii, jj, eb = make(N, Eu, seed=N + Eu, ring=ring)
    rc, x, st = host_solve(ii, jj, eb, N)
    assert rc == 0
    H, b = dense(ii, jj, eb, N)
    xe = np.linalg.solve(H, b)
    assert np.abs(x - xe).max() <= 1e-10 * np.abs(xe).max()


def test_chain_gets_a_bushy_tree_and_random_chords_are_left_to_the_pcg():
    ii, jj, eb = make(1024, 1024, seed=1)  # bare odometry ring: the fixed pose cuts it into a chain of 1023 poses
    rc, _, st = host_solve(ii, jj, eb, 1024)
    assert rc == 0 and st[2] <= 16 and st[3] == 1, st  # ~log2(N) levels, tiny fill
    ii, jj, eb = make(300, 1200, seed=2)  # random chords: the factor fills in
    rc, _, st = host_solve(ii, jj, eb, 300)
    assert rc == 0 and st[3] == 0, st


def test_duplicates_self_edges_and_the_fixed_pose():
    ii = np.array([0, 1, 1, 2, 2, 1, 3, 3], dtype=np.int64)  # duplicates (1,2) twice, a self-edge (3,3), edges to pose 0
    jj = np.array([1, 0, 2, 1, 1, 2, 3, 0], dtype=np.int64)
    rng = np.random.default_rng(3)
    eb = np.zeros((8, 56))
    for e in range(8):
        J = rng.standard_normal((20, 7))
        eb[e, :49] = (J.T @ J).ravel()
        eb[e, 49:] = J.T @ rng.standard_normal(20)
    rc, x, _ = host_solve(ii, jj, eb, 4)
    H, b = dense(ii, jj, eb, 4)
    assert rc == 0 and np.abs(x - np.linalg.solve(H, b)).max() < 1e-10


def test_not_positive_definite_is_reported():
    ii, jj, eb = make(8, 8, seed=5)
    eb[:, :49] = 0.0
    rc, _, _ = host_solve(ii, jj, eb, 8)
    assert rc == 1
