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


def test_fused_push_two_emulated_ranks(backend):
    """b200slam_gn_edge_blocks_push / b200slam_gn_wait_peers with two ranks emulated on ONE device: each "rank" owns
    a pair of ping-pong (E,56) arrays and a flag array in ordinary device memory, and the peer pointer tables point
    at both (on NVLink boxes the same tables hold symmetric-memory peer addresses).  After each rank's push both
    ranks' arrays hold all rows, the flags carry the sequence number, and the solve matches the unsharded driver."""
    import ctypes
    from types import SimpleNamespace
    dev = torch.device("cuda:0")
    g = synthetic.make_graph(seed=4, N=12, E_undirected=14, h=48, w=64)
    G = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    N, n, E, world = 12, 48 * 64, 28, 2
    prm = backend.make_params(0, 0.003, 10.0, 0.0, 1.5)
    Tref = G["Twc"].clone()
    backend.gauss_newton_rays(Tref, G["Xs"], G["Cs"], G["ii"], G["jj"], G["idx_ii2jj"], G["valid_match"], G["Q"],
                              0.003, 10.0, 0.0, 1.5, 4, 0.0)
    rows = backend.edge_rows(G["ii"], G["jj"])
    hooks = dist_gn.CudaHooks(prm, N, E, n, dev)
    blocks = [[torch.zeros(E, 56, dtype=torch.float64, device=dev) for _ in range(2)] for _ in range(world)]  # [rank][parity]
    flags = [torch.zeros(64, dtype=torch.int32, device=dev) for _ in range(world)]
    arr = ctypes.c_void_p * world
    pushes = [SimpleNamespace(world=world, rank=r, blocks=blocks[r], flags=flags[r],
                              block_ptrs=[arr(*[blocks[p][par].data_ptr() for p in range(world)]) for par in (0, 1)],
                              flag_ptrs=arr(*[flags[p].data_ptr() for p in range(world)])) for r in range(world)]
    T = [G["Twc"].clone() for _ in range(world)]
    dx = torch.zeros(N - 1, 7, device=dev)
    for it in range(4):
        seq = it + 1
        for r in range(world):  # both ranks build + push (one stream: the waits come after both pushes)
            e0, Ec = dist_gn.shard_range(E, r, world)
            sl = slice(e0, e0 + Ec)
            hooks.edge_blocks_push(T[r], G["Xs"], G["Cs"], None, rows[0], rows[1], G["idx_ii2jj"][sl].contiguous(),
                                   G["valid_match"][sl].contiguous(), G["Q"][sl].contiguous(), e0, Ec, pushes[r], seq,
                                   wait=False)
        for r in range(world):
            hooks.wait_peers(pushes[r], seq)
            hooks.solve_retract(T[r], rows[0], rows[1], blocks[r][seq & 1], 0.0, dx)
        torch.cuda.synchronize()
        assert torch.equal(blocks[0][seq & 1], blocks[1][seq & 1])
        assert flags[0][:2].tolist() == [seq, seq] and flags[1][:2].tolist() == [seq, seq]
    assert torch.equal(T[0], Tref) and torch.equal(T[1], Tref)
