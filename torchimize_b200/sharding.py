"""Multi-GPU sharding of the batch dimension.

Problems are independent, so B is split into contiguous shards, one process per GPU
(torch.distributed, NCCL over NVLink/NVSwitch).  The data path has NO per-iteration collective:

  mode='local'  each rank iterates until ITS shard is done; one all-gather of the finals.
  mode='exact'  additionally all-reduces the 4 batch-global fp64 sums of the control block
                (32 bytes) once per iteration, so the global break (lma:118, gna:88, gda:84), the
                "any rho > 0" gate (lma:110) and GD's batch-global fcon (gda:76) see the whole
                batch and iteration counts match a single-device run of the reference.
  mode='exact-peer'  the same exchange done by our own kernel (nlls_iter_end_peer): every rank stores its
                32 bytes + a stamp straight into the peers' mailboxes over NVLink (symmetric memory) and
                sums what it received — no collective library call inside the iteration.
"""
import os

import torch
import torch.distributed as dist


def shard_bounds(B, world_size, rank):
    """Contiguous [lo, hi) slice of rank; the first B % world_size ranks hold one extra problem."""
    base, extra = divmod(B, world_size)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(t, world_size, rank, dim=0):
    lo, hi = shard_bounds(t.shape[dim], world_size, rank)
    return t.narrow(dim, lo, hi - lo)


def make_reduce_hook(group=None):
    """Hook for _engine.solve(reduce_hook=...): sums the control-block reductions over ranks."""
    def hook(red):
        dist.all_reduce(red, op=dist.ReduceOp.SUM, group=group)
    return hook


class PeerExchange:
    """Peer-mapped mailboxes for nlls_iter_end_peer: the 32-byte exchange of exact mode done by our own kernel with NVLink
    P2P stores instead of an NCCL all-reduce per iteration (one launch instead of two, no collective launch latency).
    Needs every rank of `group` on one NVLink/NVSwitch box (torch symmetric memory over CUDA IPC)."""

    def __init__(self, group=None):
        import torch.distributed._symmetric_memory as symm_mem
        group = dist.group.WORLD if group is None else group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        if self.world > 32:
            raise ValueError('PeerExchange: at most 32 ranks')
        dev = torch.device('cuda', torch.cuda.current_device())
        self.buf = symm_mem.empty(4 * self.world * 8, dtype=torch.float64, device=dev)   # [2 epoch parities][2 iteration parities][world][8]
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, group=group.group_name)
        self.ptrs = torch.tensor([int(q) for q in self.handle.buffer_ptrs], dtype=torch.int64, device=dev)
        self.epoch = 0
        self.group = group
        self.timeout_s = float(os.environ.get('NLLS_PEER_TIMEOUT_S', '60'))   # a silent peer sets ctrl->error after this long
        torch.cuda.synchronize()
        dist.barrier(group)                                  # every mailbox is zeroed before anybody writes into a peer


def make_peer_hook(group=None):
    """Hook for _engine.solve(reduce_hook=...) that selects the peer-memory iteration end (see PeerExchange)."""
    def hook(red):                                           # never called: the engine launches nlls_iter_end_peer instead
        raise RuntimeError('peer hook is driven by the engine')
    hook.peer = PeerExchange(group)
    return hook


def gather_results(p_local, B_total, group=None):
    """All-gather the per-rank [B_r, P] finals into the full [B_total, P] tensor on every rank
    (one collective; shards may differ by one row, so they are padded to the largest)."""
    world = dist.get_world_size(group)
    P = p_local.shape[1]
    rows = (B_total + world - 1) // world
    pad = torch.zeros(rows, P, dtype=p_local.dtype, device=p_local.device)
    pad[:p_local.shape[0]] = p_local
    out = torch.empty(world * rows, P, dtype=p_local.dtype, device=p_local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    parts = []
    for r in range(world):
        lo, hi = shard_bounds(B_total, world, r)
        parts.append(out[r * rows: r * rows + (hi - lo)])
    return torch.cat(parts, dim=0)


_peer_hooks = {}


def solve_sharded(solver_fn, p_local, function, jac_function=None, args=(), wvec=None, mode='local',
                  group=None, B_total=None, gather=True, **kw):
    """Run one of the drop-in solvers on this rank's shard and (optionally) gather the finals.

    solver_fn: lsq_lma_parallel | lsq_gna_parallel | gradient_descent_parallel
    p_local / args: this rank's shard (use shard()/shard_bounds() to cut them)
    mode: how the SHARDS agree on stopping (module docstring); the solvers' own mode='retire' goes through **kw as
          solver_mode='retire' and needs mode='local' here (a retired problem has nothing batch-global left to exchange)
    returns (coeffs_list_local, p_full or None)"""
    from .functions.parallel import _engine
    if mode not in ('local', 'exact', 'exact-peer'):
        raise ValueError("mode must be 'local', 'exact' or 'exact-peer'")
    if 'solver_mode' in kw:
        kw['mode'] = kw.pop('solver_mode')
    multi = dist.is_initialized() and dist.get_world_size(group) > 1
    if multi and mode != 'local' and (p_local.shape[0] == 0 or (B_total is not None and B_total < dist.get_world_size(group))):
        # a rank without problems returns before the first iteration and would leave the others waiting in the exchange
        raise ValueError(f"mode='{mode}' needs at least one problem on every rank")
    hook = None
    if mode == 'exact' and multi:
        hook = make_reduce_hook(group)
    elif mode == 'exact-peer' and multi:                     # exact semantics, exchange by nlls_iter_end_peer (NVLink P2P)
        if group not in _peer_hooks:
            _peer_hooks[group] = make_peer_hook(group)
        hook = _peer_hooks[group]
    if hook is None:
        coeffs = solver_fn(p_local, function, jac_function, args=args, wvec=wvec, **kw)
    else:
        if getattr(hook, 'peer', None) is not None:
            # align the ranks before the first stamp: whatever a rank did before this call (data loading, a plugin
            # compile, a GC pause) must not count against the peers' timeout
            dist.barrier(group)
        token = _engine.set_default_reduce_hook(hook)
        try:
            coeffs = solver_fn(p_local, function, jac_function, args=args, wvec=wvec, **kw)
        finally:
            _engine.set_default_reduce_hook(token)
    full = None
    if gather and dist.is_initialized() and dist.get_world_size(group) > 1:
        if B_total is None:
            n = torch.tensor([p_local.shape[0]], device=p_local.device)
            dist.all_reduce(n, group=group)
            B_total = int(n.item())
        full = gather_results(p_local, B_total, group)
    elif gather:
        full = p_local
    return coeffs, full
