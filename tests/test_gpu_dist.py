"""GPU tier: the sharded GN driver with the CUDA hooks.  Two shards are emulated on ONE device by
running both ranks' stage 1+2 into the shared block buffer (what the allreduce produces), then the
replicated solve; the result must equal the unsharded single-call driver bit for bit."""
import pytest
import torch

from splatt3r_slam_b200 import dist_gn, synthetic

pytestmark = pytest.mark.gpu


def test_two_shards_equal_unsharded(backend):
    dev = torch.device("cuda:0")
    g = synthetic.make_graph(seed=4, N=12, E_undirected=14, h=48, w=64)  # D = 77 > fused small path
    G = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    N, n, E = 12, 48 * 64, 28
    prm = backend.make_params(0, 0.003, 10.0, 0.0, 1.5)
    Tref = G["Twc"].clone()
    backend.gauss_newton_rays(Tref, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"],
                              0.003, 10.0, 0.0, 1.5, 4, 0.0)
    rows = backend.edge_rows(G["ii"], G["jj"])
    hooks = dist_gn.CudaHooks(prm, N, E, n, dev)
    T = G["Twc"].clone()
    eb = torch.zeros(E, 56, dtype=torch.float64, device=dev)
    dx = torch.zeros(N - 1, 7, device=dev)
    for _ in range(4):
        eb.zero_()
        for rank in range(2):
            e0, Ec = dist_gn.shard_range(E, rank, 2)
            sl = slice(e0, e0 + Ec)
            hooks.edge_blocks(T, G["Xs"], G["Cs"], None, rows[0], rows[1], G["idx_ii2jj"][sl].contiguous(),
                              G["valid_match"][sl].contiguous(), G["Q"][sl].contiguous(), e0, Ec, eb)
        hooks.solve_retract(T, rows[0], rows[1], eb, 0.0, dx)
    torch.cuda.synchronize()
    assert torch.equal(T, Tref)
    # world_size 1 through the public driver
    T1 = G["Twc"].clone()
    dist_gn.gauss_newton_sharded(prm, T1, G["Xs"], G["Cs"], None, G["ii"], G["jj"], G["idx_ii2jj"],
                                 G["valid_match"], G["Q"], 4, 0.0)
    assert torch.equal(T1, Tref)
