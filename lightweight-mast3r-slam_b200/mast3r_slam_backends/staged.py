"""Staged and multi-GPU Gauss-Newton on top of the C ABI (m3r_gn_begin / m3r_gn_linearize /
m3r_gn_solve_retract, include/m3r_backends.h).

Multi-GPU: one process per GPU.  Every rank holds a replica of
`Twc`, `Xs`, `Cs` and the full `ii`/`jj` lists, and the rows of `idx_ii2jj`, `valid_match`,
`Q` for its contiguous shard of the directed edges.  Per iteration each rank linearises its
shard into the dense fp64 system, the systems are summed with one all-reduce (NCCL over
NVLink), and every rank runs the same Cholesky solve and retraction on its replica: the
all-reduce result is identical on all ranks, so the replicas stay bitwise identical with no
broadcast.  The reference has no multi-GPU path; this is new.
"""
import ctypes

import torch

from . import _lib, _workspace, _ptr, _stream
from ._lib import GnProblem

KINDS = {"rays": _lib.GN_RAYS, "calib": _lib.GN_CALIB, "points": _lib.GN_POINTS}


def shard_range(num_edges, rank, world_size):
    """Contiguous, near-equal split of the directed edges (all edges cost n points)."""
    base, rem = divmod(num_edges, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


class StagedGN:
    """One Gauss-Newton problem, driven stage by stage.

    `idx_ii2jj`, `valid_match`, `Q` hold ONLY the rows [edge_begin, edge_end) of the graph's
    directed edges; `ii`, `jj` are the whole lists."""

    def __init__(self, kind, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, *, edge_begin=0,
                 edge_end=None, K=None, height=0, width=0, pixel_border=0, z_eps=0.0,
                 sigma_ray=1.0, sigma_dist=1.0, sigma_pixel=1.0, sigma_depth=1.0,
                 sigma_point=1.0, C_thresh=0.0, Q_thresh=0.0):
        for name, t in dict(Twc=Twc, Xs=Xs, Cs=Cs, ii=ii, jj=jj, idx_ii2jj=idx_ii2jj,
                            valid_match=valid_match, Q=Q).items():
            if not t.is_cuda:
                raise RuntimeError(f"{name} must be a CUDA tensor")
            if not t.is_contiguous():
                raise RuntimeError(f"{name} must be contiguous")
        self.dev = Twc.device
        N, n, E = Xs.shape[0], Xs.shape[1], ii.shape[0]
        if edge_end is None:
            edge_end = E
        rows = edge_end - edge_begin
        if idx_ii2jj.numel() != rows * n or valid_match.numel() != rows * n or Q.numel() != rows * n:
            raise RuntimeError("edge data must hold exactly the rows of this shard")
        self.N, self.n, self.E = N, n, E
        self.keep = (Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, K)  # keep storage alive
        p = GnProblem()
        p.kind = KINDS[kind]
        p.num_poses, p.num_points, p.num_edges = N, n, E
        p.edge_begin, p.edge_end = edge_begin, edge_end
        p.Twc, p.Xs, p.Cs = Twc.data_ptr(), Xs.data_ptr(), Cs.data_ptr()
        p.ii, p.jj = ii.data_ptr(), jj.data_ptr()
        p.idx_ii2jj, p.valid_match, p.Q = idx_ii2jj.data_ptr(), valid_match.data_ptr(), Q.data_ptr()
        p.sigma_ray, p.sigma_dist = sigma_ray, sigma_dist
        p.K = K.data_ptr() if K is not None else None
        p.height, p.width, p.pixel_border = int(height), int(width), int(pixel_border)
        p.z_eps, p.sigma_pixel, p.sigma_depth = z_eps, sigma_pixel, sigma_depth
        p.sigma_point = sigma_point
        p.C_thresh, p.Q_thresh = C_thresh, Q_thresh
        self.prob = p
        self.ws = _workspace(self.dev)
        self.D = 7 * (N - 1)
        self.system = torch.empty(int(_lib.gn_system_size(N)), dtype=torch.float64, device=self.dev)
        self.dx = torch.zeros((N - 1, 7), dtype=torch.float32, device=self.dev)

    def begin(self):
        with torch.cuda.device(self.dev):
            _lib.check(_lib.gn_begin(self.ws.handle, ctypes.byref(self.prob), _stream()), "gn_begin")

    def linearize(self, into=None):
        """Accumulate this shard and assemble the dense system into `self.system`, or into the
        device buffer at address `into` (an exchange buffer of PeerExchange)."""
        dst = _ptr(self.system) if into is None else ctypes.c_void_p(int(into))
        with torch.cuda.device(self.dev):
            _lib.check(_lib.gn_linearize(self.ws.handle, ctypes.byref(self.prob), dst, _stream()),
                       "gn_linearize")

    def solve_retract(self, delta_thresh=0.0):
        with torch.cuda.device(self.dev):
            _lib.check(_lib.gn_solve_retract(self.ws.handle, ctypes.byref(self.prob),
                                             _ptr(self.system), _ptr(self.dx), float(delta_thresh),
                                             _stream()), "gn_solve_retract")

    def H(self):
        """Dense [D,D] view of the assembled system (lower triangle meaningful)."""
        return self.system[: self.D * self.D].view(self.D, self.D)

    def b(self):
        return self.system[self.D * self.D:]

    def edge_blocks(self):
        """(Hjj [E_shard,7,7], gj [E_shard,7]) fp64 copies of the last linearisation."""
        hp, gp = ctypes.c_void_p(), ctypes.c_void_p()
        _lib.check(_lib.gn_edge_blocks(self.ws.handle, ctypes.byref(hp), ctypes.byref(gp)),
                   "gn_edge_blocks")
        rows = self.prob.edge_end - self.prob.edge_begin
        H = _from_ptr(hp.value, (rows, 7, 7), self.dev)
        g = _from_ptr(gp.value, (rows, 7), self.dev)
        return H.clone(), g.clone()

    def iterations_done(self):
        it = ctypes.c_int(0)
        with torch.cuda.device(self.dev):
            _lib.check(_lib.gn_iterations_done(self.ws.handle, _stream(), ctypes.byref(it)),
                       "gn_iterations_done")
        return it.value


class _CudaArray:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {
            "shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 3,
            "strides": None,
        }


def _from_ptr(ptr, shape, device):
    with torch.cuda.device(device):
        return torch.as_tensor(_CudaArray(ptr, shape, "<f8"), device=device)


class TorchAllReduce:
    """Exchange through torch.distributed (NCCL on GPUs, gloo in the CPU tests): one
    all_reduce(sum) of the assembled system per iteration."""

    def __init__(self, group=None):
        self.group = group

    def round_buffer(self, stage):
        return None  # linearise into stage.system

    def allreduce(self, stage):
        import torch.distributed as dist

        if dist.is_initialized() and dist.get_world_size(self.group) > 1:
            dist.all_reduce(stage.system, op=dist.ReduceOp.SUM, group=self.group)


class PeerExchange:
    """Exchange through this library's own NVLink peer-memory kernels (m3r_gn_comm_*,
    include/m3r_backends.h): each rank linearises straight into its exchange buffer, flags it
    complete in every peer, and sums all ranks' buffers in rank order into its local system.
    One process per GPU; the exchange areas are mapped across processes with CUDA IPC, the
    handles travel once through torch.distributed at construction."""

    def __init__(self, num_poses, group=None, device=None):
        import torch.distributed as dist

        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.size = int(_lib.gn_system_size(num_poses))
        self.handle = ctypes.c_void_p()
        with torch.cuda.device(self.dev):
            _lib.check(_lib.gn_comm_create(ctypes.byref(self.handle), self.rank, self.world, self.size),
                       "gn_comm_create")
            if self.world > 1:
                mine = (ctypes.c_ubyte * _lib.IPC_HANDLE_BYTES)()
                _lib.check(_lib.gn_comm_handle(self.handle, mine), "gn_comm_handle")
                gathered = [None] * self.world
                dist.all_gather_object(gathered, bytes(mine), group=group)
                blob = b"".join(gathered)
                buf = (ctypes.c_ubyte * len(blob)).from_buffer_copy(blob)
                _lib.check(_lib.gn_comm_connect(self.handle, buf), "gn_comm_connect")
                dist.barrier(group=group)  # nobody signals into an area that is not mapped yet

    def __del__(self):
        try:
            if self.handle:
                _lib.gn_comm_destroy(self.handle)
        except Exception:
            pass

    def round_buffer(self, stage):
        ptr = ctypes.c_void_p()
        _lib.check(_lib.gn_comm_next_round(self.handle, ctypes.byref(ptr)), "gn_comm_next_round")
        return ptr.value

    def allreduce(self, stage):
        with torch.cuda.device(self.dev):
            _lib.check(_lib.gn_comm_signal(self.handle, stage.ws.handle, _stream()), "gn_comm_signal")
            _lib.check(_lib.gn_comm_reduce(self.handle, stage.ws.handle, _ptr(stage.system),
                                           stage.N, _stream()), "gn_comm_reduce")


def run_spmd(stage, exchange, max_iter, delta_thresh):
    """The SPMD Gauss-Newton loop every rank runs: linearise the own shard, sum the systems over
    the ranks, solve and retract the replica.  `stage` offers begin / linearize(into) /
    solve_retract / system (StagedGN here; the CPU tests plug in an oracle-backed stage)."""
    stage.begin()
    for _ in range(int(max_iter)):
        stage.linearize(into=exchange.round_buffer(stage))
        exchange.allreduce(stage)
        stage.solve_retract(delta_thresh)


def gauss_newton_sharded(kind, Twc, Xs, Cs, ii, jj, idx_shard, valid_shard, Q_shard, *,
                         max_iter, delta_thresh, group=None, exchange=None, **params):
    """SPMD Gauss-Newton: call on every rank of `group` with the same replicated tensors and
    this rank's shard of the edge data (rows `shard_range(E, rank, world)`).

    `exchange`: a PeerExchange (NVLink peer memory, default when world > 1 is not given: NCCL)
    or TorchAllReduce.  Returns ([dx], stage); Twc is updated in place on every rank (bitwise
    identical replicas)."""
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    begin, end = shard_range(ii.shape[0], rank, world)
    gn = StagedGN(kind, Twc, Xs, Cs, ii, jj, idx_shard, valid_shard, Q_shard,
                  edge_begin=begin, edge_end=end, **params)
    run_spmd(gn, exchange if exchange is not None else TorchAllReduce(group), max_iter, delta_thresh)
    # raises if a peer never arrived (the device then stopped the solve before touching Twc again)
    # or if ii/jj name more keyframes than Twc holds; synchronises this rank's stream
    gn.iterations_done()
    return [gn.dx], gn
