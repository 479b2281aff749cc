"""GPU tier: the two-way GN entry and the device-resident factor-graph glue against the
reference's own sequence — torch.stack / torch.cat (global_opt.py:104-119) followed by gauss_newton_rays — on the
same graphs.  Same kernels, same summation order: poses must be bit-identical."""
from types import SimpleNamespace

import pytest
import torch

from splatt3r_slam_b200 import global_opt, synthetic

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def one_way_graph(N, Eu, h, w, seed):
    """Synthetic two-way graph (E = 2*Eu rows, second half = reversed edges) split back into the FactorGraph's
    one-way fields."""
    g = synthetic.make_graph(seed=seed, N=N, E_undirected=Eu, h=h, w=w)
    G = {k: (v.to(DEV) if torch.is_tensor(v) else v) for k, v in g.items()}
    assert torch.equal(G["ii"][:Eu], G["jj"][Eu:]) and torch.equal(G["jj"][:Eu], G["ii"][Eu:])
    c = lambda t: t.contiguous()
    graph = SimpleNamespace(ii=c(G["ii"][:Eu]), jj=c(G["jj"][:Eu]), idx_ii2jj=c(G["idx_ii2jj"][:Eu]),
                            idx_jj2ii=c(G["idx_ii2jj"][Eu:]), valid_match_j=c(G["valid_match"][:Eu]),
                            valid_match_i=c(G["valid_match"][Eu:]), Q_ii2jj=c(G["Q"][:Eu]), Q_jj2ii=c(G["Q"][Eu:]))
    return G, graph


@pytest.mark.parametrize("N,Eu,h,w", [(2, 1, 96, 128), (6, 9, 46, 61), (12, 60, 24, 32), (40, 60, 24, 32)])
def test_twoway_equals_concatenated(backend, N, Eu, h, w):
    G, graph = one_way_graph(N, Eu, h, w, seed=N + Eu)
    cfg = dict(global_opt.DEFAULT_CFG, max_iters=4, delta_norm=0.0)
    T_ref = G["Twc"].clone()
    ii, jj, idx, valid, Q = global_opt.prep_two_way_edges(graph)
    backend.gauss_newton_rays(T_ref, G["Xs"], G["Cs"], ii, jj, idx, valid, Q, cfg["sigma_ray"], cfg["sigma_dist"],
                              cfg["C_conf"], cfg["Q_conf"], cfg["max_iters"], cfg["delta_norm"])
    # keyframe buffers with spare capacity, ids 0..N-1: zero-copy views, poses updated in place
    cap = N + 3
    X_buf = torch.zeros((cap,) + tuple(G["Xs"].shape[1:]), device=DEV); X_buf[:N] = G["Xs"]
    C_buf = torch.zeros((cap,) + tuple(G["Cs"].shape[1:]), device=DEV); C_buf[:N] = G["Cs"]
    T_buf = torch.zeros((cap, 8), device=DEV); T_buf[:N] = G["Twc"]
    dx = global_opt.solve_GN_rays(graph, X_buf, C_buf, T_buf, cfg=cfg)
    assert dx.shape == (N - 1, 7)
    assert torch.equal(T_buf[:N], T_ref) and torch.equal(T_buf[0], G["Twc"][0])
    assert not torch.equal(T_ref[1:], G["Twc"][1:])


def test_non_contiguous_keyframe_ids_read_the_buffers_in_place(backend):
    """Any id set: the pointmap / confidence buffers go to the kernels as they are (kf_slots map), only the pose rows
    are gathered and scattered; poses bit-identical to the reference's stack + cat sequence."""
    N, Eu, h, w = 5, 6, 24, 32
    G, graph = one_way_graph(N, Eu, h, w, seed=4)
    ids = torch.tensor([1, 4, 5, 8, 9], device=DEV)  # keyframe ids of the 5 graph nodes in a 12-slot buffer
    graph.ii, graph.jj = ids[graph.ii].contiguous(), ids[graph.jj].contiguous()
    cfg = dict(global_opt.DEFAULT_CFG, max_iters=3, delta_norm=0.0)
    T_ref = G["Twc"].clone()
    ii, jj, idx, valid, Q = global_opt.prep_two_way_edges(graph)
    backend.gauss_newton_rays(T_ref, G["Xs"], G["Cs"], ii, jj, idx, valid, Q, cfg["sigma_ray"], cfg["sigma_dist"],
                              cfg["C_conf"], cfg["Q_conf"], cfg["max_iters"], cfg["delta_norm"])
    X_buf = torch.randn((12,) + tuple(G["Xs"].shape[1:]), device=DEV); X_buf[ids] = G["Xs"]
    C_buf = torch.rand((12,) + tuple(G["Cs"].shape[1:]), device=DEV); C_buf[ids] = G["Cs"]
    T_buf = torch.zeros((12, 8), device=DEV); T_buf[:, 6:] = 1.0; T_buf[ids] = G["Twc"]
    untouched = T_buf.clone()
    Xv, Cv, _, slots = global_opt.keyframe_views(X_buf, C_buf, T_buf, ids)
    assert Xv.data_ptr() == X_buf.data_ptr() and Cv.data_ptr() == C_buf.data_ptr() and torch.equal(slots, ids)
    global_opt.solve_GN_rays(graph, X_buf, C_buf, T_buf, cfg=cfg)
    assert torch.equal(T_buf[ids], T_ref)
    mask = torch.ones(12, dtype=torch.bool, device=DEV); mask[ids] = False
    assert torch.equal(T_buf[mask], untouched[mask])


def test_calib_twoway(backend):
    N, Eu, h, w = 4, 4, 24, 32
    G, graph = one_way_graph(N, Eu, h, w, seed=6)
    cfg = dict(global_opt.DEFAULT_CFG, max_iters=3, delta_norm=0.0)
    T_ref = G["Twc"].clone()
    ii, jj, idx, valid, Q = global_opt.prep_two_way_edges(graph)
    backend.gauss_newton_calib(T_ref, G["Xs"], G["Cs"], G["K"], ii, jj, idx, valid, Q, h, w, cfg["pixel_border"],
                               cfg["depth_eps"], cfg["sigma_pixel"], cfg["sigma_depth"], cfg["C_conf"], cfg["Q_conf"],
                               cfg["max_iters"], cfg["delta_norm"])
    T_buf = G["Twc"].clone()
    global_opt.solve_GN_calib(graph, G["Xs"], G["Cs"], T_buf, G["K"], (h, w), cfg=cfg)
    assert torch.equal(T_buf, T_ref)


@pytest.mark.parametrize("N,Eu,h,w,name", [(2, 1, 96, 128, "rays"), (12, 60, 24, 32, "calib"), (40, 60, 24, 32, "rays")])
def test_kf_slots_on_every_gn_path(backend, N, Eu, h, w, name):
    """The slot map through the persistent kernel (N = 2), the fused per-iteration path and the large-graph path."""
    G, graph = one_way_graph(N, Eu, h, w, seed=3 * N + Eu)
    ids = (torch.arange(N, device=DEV) * 2 + 3)  # keyframe ids 3, 5, 7, ... in a (2N + 4)-slot buffer
    cap = 2 * N + 4
    gslots = SimpleNamespace(**vars(graph))
    gslots.ii, gslots.jj = ids[graph.ii].contiguous(), ids[graph.jj].contiguous()
    cfg = dict(global_opt.DEFAULT_CFG, max_iters=3, delta_norm=0.0)
    X_buf = torch.randn((cap,) + tuple(G["Xs"].shape[1:]), device=DEV); X_buf[ids] = G["Xs"]
    C_buf = torch.rand((cap,) + tuple(G["Cs"].shape[1:]), device=DEV); C_buf[ids] = G["Cs"]
    T_buf = torch.zeros((cap, 8), device=DEV); T_buf[:, 6:] = 1.0; T_buf[ids] = G["Twc"]
    T_ref = G["Twc"].clone()
    if name == "rays":
        global_opt.solve_GN_rays(graph, G["Xs"], G["Cs"], T_ref, cfg=cfg)
        global_opt.solve_GN_rays(gslots, X_buf, C_buf, T_buf, cfg=cfg)
    else:
        global_opt.solve_GN_calib(graph, G["Xs"], G["Cs"], T_ref, G["K"], (h, w), cfg=cfg)
        global_opt.solve_GN_calib(gslots, X_buf, C_buf, T_buf, G["K"], (h, w), cfg=cfg)
    assert torch.equal(T_buf[ids], T_ref) and not torch.equal(T_ref[1:], G["Twc"][1:])
