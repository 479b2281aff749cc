"""Edge-sharded multi-GPU Gauss-Newton — new functionality on top of the
reference's single-GPU `gauss_newton_*` (gn_kernels.cu:1139-1227): one process per GPU, the E
directed edges split into contiguous ranges, poses/pointmaps replicated.

Per iteration, on every rank:
  1. stage 1+2 on the rank's own edges -> rows [e0, e0+Ec) of `edge_blocks` (E,56) fp64
     (= [H_jj | g_j] per edge, all other rows zero)                     [b200slam_gn_edge_blocks]
  2. the exchange.  On NVLink-connected GPUs it is FUSED into step 1: the finalizing kernel stores each block into
     every rank's array through peer pointers (symmetric memory) and releases a per-rank flag; the stream then
     waits for all flags (EdgeBlockPush: no collective call at all).  Otherwise (gloo / CPU tests, no symmetric
     memory) ONE collective: all-gather of the ranks' row ranges of `edge_blocks` over NCCL/NVLink (rows are
     disjoint, so nothing is summed: every rank ends with bit-identical blocks and the replicated
     solve below stays in lock-step).  Equal shards use all_gather_into_tensor in place; ragged
     shards (E % world != 0) a padded one.
  3. replicated solve + retraction + ||dx|| test                         [b200slam_gn_solve_retract]
     (large graphs: pipelined PCG on the edge blocks, well under a millisecond at N = 1024, so
     replicating it costs less than one extra collective per CG iteration would)
No other exchange: the matcher shards by pair with no collective at all.

The compute hooks are injectable so the protocol can be exercised on CPU tensors with the gloo
backend (tests/test_dist_gloo.py passes oracle-backed hooks); the default hooks call the CUDA
library and fail loudly without it.
"""
import ctypes

import torch
import torch.distributed as dist


def shard_range(E, rank, world):
    """Contiguous, balanced edge range of `rank` (all edges carry the same n points)."""
    base, rem = divmod(E, world)
    e0 = rank * base + min(rank, rem)
    return e0, base + (1 if rank < rem else 0)


class CudaHooks:
    """Default compute hooks: the sm_100a library through its C ABI."""

    def __init__(self, prm, N, E, n, device):
        from . import mast3r_slam_backends as be
        self.be, self.lib, self.prm = be, be._lib, prm
        self.N, self.E, self.n, self.device = N, E, n, device
        self.nbytes = self.lib.b200slam_gn_workspace_bytes(N, E, n)
        self.ws = torch.empty(int(self.nbytes), dtype=torch.uint8, device=device)
        self._stream = lambda: torch.cuda.current_stream(device).cuda_stream
        be._check(self.lib.b200slam_gn_reset(self.ws.data_ptr(), self.nbytes, self._stream()), "gn_reset")

    def edge_blocks(self, Twc, Xs, Cs, K, ii_row, jj_row, idx_s, valid_s, Q_s, e0, Ec, out):
        # shard tensors hold rows [e0, e0+Ec) only: hand the library virtual full-E base pointers
        n = self.n
        vp = lambda t, esz: ctypes.c_void_p(t.data_ptr() - e0 * n * esz)
        self.be._check(self.lib.b200slam_gn_edge_blocks(
            ctypes.byref(self.prm), Twc.data_ptr(), Xs.data_ptr(), Cs.data_ptr(),
            K.data_ptr() if K is not None else None, ii_row.data_ptr(), jj_row.data_ptr(), vp(idx_s, 8),
            vp(valid_s, 1), vp(Q_s, 4), self.N, self.E, n, e0, Ec, out.data_ptr(), self.ws.data_ptr(),
            self.nbytes, self._stream()), "gn_edge_blocks")

    def edge_blocks_push(self, Twc, Xs, Cs, K, ii_row, jj_row, idx_s, valid_s, Q_s, e0, Ec, push, seq, wait=True):
        """Stage 1+2 with the exchange fused into the finalizing kernel (`push`: an EdgeBlockPush)."""
        n = self.n
        vp = lambda t, esz: ctypes.c_void_p(t.data_ptr() - e0 * n * esz)
        par = seq & 1
        self.be._check(self.lib.b200slam_gn_edge_blocks_push(
            ctypes.byref(self.prm), Twc.data_ptr(), Xs.data_ptr(), Cs.data_ptr(),
            K.data_ptr() if K is not None else None, ii_row.data_ptr(), jj_row.data_ptr(), vp(idx_s, 8),
            vp(valid_s, 1), vp(Q_s, 4), self.N, self.E, n, e0, Ec, push.block_ptrs[par], push.flag_ptrs, push.world,
            push.rank, seq, self.ws.data_ptr(), self.nbytes, self._stream()), "gn_edge_blocks_push")
        if wait:
            self.wait_peers(push, seq)
        return push.blocks[par]

    def wait_peers(self, push, seq):
        self.be._check(self.lib.b200slam_gn_wait_peers(push.flags.data_ptr(), push.world, seq, self._stream()),
                       "gn_wait_peers")

    def analyze(self, ii_row, jj_row):
        """Large graphs: sparse direct solve or PCG for this call (one host sync; cached per topology)."""
        self.be._check(self.lib.b200slam_gn_analyze(ii_row.data_ptr(), jj_row.data_ptr(), self.N, self.E, self.n,
                                                    self.ws.data_ptr(), self.nbytes, self._stream()), "gn_analyze")

    def solve_retract(self, Twc, ii_row, jj_row, edge_blocks, delta_thresh, dx):
        self.be._check(self.lib.b200slam_gn_solve_retract(
            Twc.data_ptr(), ii_row.data_ptr(), jj_row.data_ptr(), edge_blocks.data_ptr(), self.N, self.E,
            self.n, float(delta_thresh), dx.data_ptr(), self.ws.data_ptr(), self.nbytes, self._stream()),
            "gn_solve_retract")

    def done(self):
        """Device-side convergence flag (costs a sync; the loop below never calls it)."""
        return self.be.gn_status_ws(self.ws)


class EdgeBlockGather:
    """The one collective of an iteration: every rank contributes rows [e0, e0+Ec) of the (E,56) fp64
    edge-block array and ends with all of them (disjoint rows: nothing is summed)."""

    def __init__(self, E, rank, world, device, group=None):
        self.E, self.rank, self.world, self.group = E, rank, world, group
        self.e0, self.Ec = shard_range(E, rank, world)
        self.even = E % world == 0
        if world > 1 and not self.even:  # ragged shards: gather through a padded (world, Emax, 56) buffer
            self.Emax = (E + world - 1) // world
            self.pad = torch.zeros((world, self.Emax, 56), dtype=torch.float64, device=device)
            self.mine = torch.zeros((self.Emax, 56), dtype=torch.float64, device=device)
            self.spans = [shard_range(E, r, world) for r in range(world)]

    def run(self, edge_blocks):
        if self.world == 1:
            return
        e0, Ec = self.e0, self.Ec
        if self.even:  # in place: this rank's rows are already at their final position
            dist.all_gather_into_tensor(edge_blocks, edge_blocks[e0:e0 + Ec], group=self.group)
            return
        self.mine[:Ec].copy_(edge_blocks[e0:e0 + Ec])
        dist.all_gather_into_tensor(self.pad.view(self.world * self.Emax, 56), self.mine, group=self.group)
        for r, (s0, sc) in enumerate(self.spans):
            if r != self.rank:
                edge_blocks[s0:s0 + sc].copy_(self.pad[r, :sc])


class EdgeBlockPush:
    """The exchange FUSED into the kernel that produces the blocks: two (E,56) fp64 edge-block arrays (ping-pong by
    iteration parity) and a flag array per rank live in NVLink-mapped symmetric memory; `CudaHooks.edge_blocks_push`
    has the finalizing kernel store every block of its shard into all ranks' arrays and release a per-rank sequence
    flag, and makes the stream wait for all ranks' flags.  No NCCL call, no staging copy on the data path.
    `available()` is False where symmetric memory cannot be set up (CPU / gloo tests, a single rank)."""
    _seq = 0  # process-wide exchange counter: flags only ever increase, identical on all ranks (lock-step)

    @staticmethod
    def available(device):
        if not (dist.is_initialized() and dist.get_world_size() > 1 and torch.device(device).type == "cuda"):
            return False
        try:
            import importlib
            importlib.import_module("torch.distributed._symmetric_memory")
            return dist.get_backend() == "nccl"
        except Exception:  # noqa: BLE001
            return False

    def __init__(self, E, rank, world, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        self.E, self.rank, self.world = E, rank, world
        name = (group or dist.group.WORLD).group_name
        # one symmetric allocation: [blocks parity 0 | blocks parity 1 | flags (64 x u32, zero)]
        nd = 2 * E * 56
        self.buf = symm_mem.empty(nd + 32, dtype=torch.float64, device=device)
        self.buf.zero_()
        self.hdl = symm_mem.rendezvous(self.buf, name)
        self.blocks = [self.buf[:E * 56].view(E, 56), self.buf[E * 56:nd].view(E, 56)]
        self.flags = self.buf[nd:].view(torch.int32)
        ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        arr = ctypes.c_void_p * world
        self.block_ptrs = [arr(*[p + par * E * 56 * 8 for p in ptrs]) for par in (0, 1)]
        self.flag_ptrs = arr(*[p + nd * 8 for p in ptrs])
        self.hdl.barrier()  # every rank's zeroed flags are in place before anyone pushes

    def next_seq(self):
        EdgeBlockPush._seq += 1
        return EdgeBlockPush._seq


def gauss_newton_sharded(prm, Twc, Xs, Cs, K, ii, jj, idx_shard, valid_shard, Q_shard, max_iter, delta_thresh,
                         group=None, hooks=None, rows=None, push=None):
    """Edge-sharded GN.  `push`: an EdgeBlockPush to reuse (False: force the NCCL all-gather path).  Twc (N,8) is updated in place on every rank (identically).
    ii/jj: ALL E keyframe ids (replicated); idx/valid/Q: this rank's rows [e0, e0+Ec) of the
    (E,n,...) arrays, with (e0, Ec) = shard_range(E, rank, world).  Returns dx (N-1,7)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    N, n, E = Xs.shape[0], Xs.shape[1], ii.shape[0]
    e0, Ec = shard_range(E, rank, world)
    assert idx_shard.shape[0] == Ec and valid_shard.shape[0] == Ec and Q_shard.shape[0] == Ec
    if rows is None:
        from .mast3r_slam_backends import edge_rows
        rows = edge_rows(ii, jj)
    ii_row, jj_row = rows
    if hooks is None:
        hooks = CudaHooks(prm, N, E, n, Twc.device)
    edge_blocks = torch.zeros((E, 56), dtype=torch.float64, device=Twc.device)
    dx = torch.zeros((N - 1, 7), dtype=torch.float32, device=Twc.device)
    if hasattr(hooks, "analyze"):
        hooks.analyze(ii_row, jj_row)
    if push is None and hasattr(hooks, "edge_blocks_push") and EdgeBlockPush.available(Twc.device):
        push = EdgeBlockPush(E, rank, world, Twc.device, group)
    if push:  # fused build + exchange over NVLink peer memory
        for _ in range(max_iter):
            eb = hooks.edge_blocks_push(Twc, Xs, Cs, K, ii_row, jj_row, idx_shard, valid_shard, Q_shard, e0, Ec, push,
                                        push.next_seq())
            hooks.solve_retract(Twc, ii_row, jj_row, eb, delta_thresh, dx)
        return dx
    gather = EdgeBlockGather(E, rank, world, Twc.device, group)
    for _ in range(max_iter):
        hooks.edge_blocks(Twc, Xs, Cs, K, ii_row, jj_row, idx_shard, valid_shard, Q_shard, e0, Ec, edge_blocks)
        gather.run(edge_blocks)
        hooks.solve_retract(Twc, ii_row, jj_row, edge_blocks, delta_thresh, dx)
    return dx
