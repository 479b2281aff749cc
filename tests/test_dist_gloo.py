"""CPU tier, world_size 2 over gloo: the edge-sharded Gauss-Newton protocol of
splatt3r_slam_b200.dist_gn (shard ranges, ONE all-gather of the edge blocks per iteration,
replicated solve in lock-step).  The compute hooks are backed by the CPU oracle here; on the GPU
the same driver runs the CUDA hooks (tests/test_gpu_dist.py)."""
import os

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from splatt3r_slam_b200 import dist_gn, synthetic

GN = dict(sigma_ray=0.003, sigma_dist=10.0, C_thresh=0.0, Q_thresh=1.5)


class OracleHooks:
    """edge_blocks / solve_retract on CPU tensors through the oracle (test infrastructure)."""

    def __init__(self, g):
        self.g, self.done_flag, self.iters = g, False, 0

    def edge_blocks(self, Twc, Xs, Cs, K, ii_row, jj_row, idx_s, valid_s, Q_s, e0, Ec, out):
        import oracle
        if self.done_flag or Ec == 0:
            return
        ii = ii_row[e0:e0 + Ec].contiguous()
        jj = jj_row[e0:e0 + Ec].contiguous()
        # rows are already dense ranks of the FULL graph: bypass the oracle's own remap
        Hs = torch.zeros(4, Ec, 7, 7)
        gs = torch.zeros(2, Ec, 7)
        import ctypes
        f = ctypes.c_float
        oracle.lib().oracle_gn_build(0, oracle._p(Twc), oracle._p(Xs), oracle._p(Cs), oracle._p(None),
                                     oracle._p(ii), oracle._p(jj), oracle._p(idx_s.contiguous()),
                                     oracle._p(valid_s.view(torch.uint8).contiguous()), oracle._p(Q_s.contiguous()),
                                     Ec, Xs.shape[1], f(GN["sigma_ray"]), f(GN["sigma_dist"]), f(GN["C_thresh"]),
                                     f(GN["Q_thresh"]), 0, 0, 0, f(0.0), oracle._p(Hs), oracle._p(gs))
        out[e0:e0 + Ec, :49] = Hs[3].reshape(Ec, 49).double()
        out[e0:e0 + Ec, 49:] = gs[1].double()

    def solve_retract(self, Twc, ii_row, jj_row, edge_blocks, delta_thresh, dx):
        import oracle
        if self.done_flag:
            return
        E = ii_row.shape[0]
        H = edge_blocks[:, :49].reshape(E, 7, 7).float()
        g = edge_blocks[:, 49:].float()
        Hs = torch.stack([H, -H, -H, H]).contiguous()
        gs = torch.stack([-g, g]).contiguous()
        d, ok = oracle.gn_solve(Hs, gs, (ii_row - 1).contiguous(), (jj_row - 1).contiguous(), Twc.shape[0] - 1)
        dx.copy_(d)
        oracle.pose_retr(Twc, dx)
        self.iters += 1
        if dx.norm().item() < delta_thresh:
            self.done_flag = True


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = synthetic.make_graph(seed=9, N=4, E_undirected=5, h=24, w=32)
        E = g["ii"].numel()
        e0, Ec = dist_gn.shard_range(E, rank, world)
        Twc = g["Twc"].clone()
        hooks = OracleHooks(g)
        rows = torch.searchsorted(torch.unique(torch.cat([g["ii"], g["jj"]])), g["ii"]), \
            torch.searchsorted(torch.unique(torch.cat([g["ii"], g["jj"]])), g["jj"])
        dx = dist_gn.gauss_newton_sharded(None, Twc, g["Xs"], g["Cs"], None, g["ii"], g["jj"],
                                          g["idx_ii2jj"][e0:e0 + Ec], g["valid_match"][e0:e0 + Ec],
                                          g["Q"][e0:e0 + Ec], 4, 0.0, hooks=hooks, rows=rows)
        ret[rank] = (Twc, dx, hooks.iters)
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_edges():
    for E in (0, 1, 7, 128, 8192):
        for world in (1, 2, 3, 8):
            spans = [dist_gn.shard_range(E, r, world) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == E
            for (a, ca), (b, _) in zip(spans, spans[1:]):
                assert a + ca == b
            assert max(c for _, c in spans) - min(c for _, c in spans) <= 1


@pytest.mark.parametrize("world", [2, 3])  # E = 10: equal shards (in-place all-gather) / ragged shards (padded)
def test_sharded_gn_matches_single_process(world):
    import oracle
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, 29531 + os.getpid() % 200, ret), nprocs=world, join=True)
    T0, dx0, it0 = ret[0]
    for r in range(1, world):  # ranks in lock-step, bit-identical
        Tr, dxr, itr = ret[r]
        assert torch.equal(T0, Tr) and torch.equal(dx0, dxr) and it0 == itr == 4
    g = synthetic.make_graph(seed=9, N=4, E_undirected=5, h=24, w=32)
    Tref = g["Twc"].clone()
    oracle.gauss_newton_rays(Tref, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"],
                             GN["sigma_ray"], GN["sigma_dist"], GN["C_thresh"], GN["Q_thresh"], 4, 0.0)
    assert (T0 - Tref).abs().max().item() < 1e-5
    assert (T0 - g["Twc"]).abs().max().item() > 1e-4  # it actually moved
