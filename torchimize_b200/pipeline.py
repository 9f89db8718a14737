"""Host-resident batches through the device solvers with the PCIe copies hidden behind the solves.

The drop-in solvers take device tensors (like the reference on CUDA).  A caller whose data lives in
host memory pays, per batch, a host->device copy that is as long as the solve itself at the headline
shape (2.7 GB at ~58 GB/s = 46 ms against a 38 ms Gauss-Newton solve).  ``solve_host_batches`` keeps
``depth`` device slots and a copy stream: the copy of batch k+1 is enqueued BEFORE the solve of batch k
starts, so in steady state a batch costs max(copy, solve) instead of copy + solve.  Every batch is
still solved by one ordinary call of the drop-in solver on its own device tensors — semantics, iteration
counts and results are those of calling the solver directly (tests/test_gpu_pipeline.py).

    for p_final, n_iter in solve_host_batches(lsq_gna_parallel, batches, emg.cost, emg.jac, wvec=w, max_iter=80):
        ...

``batches`` is any iterable of ``(p0_host, args_host)`` with pinned (or pageable) CPU tensors;
results come back as pinned CPU tensors (one per slot, reused — copy them if you keep them).
"""
import torch

from . import _lib


def slot_schedule(n_batches, depth):
    """Order of operations for n_batches and `depth` device slots, as a list of
    ('copy', batch, slot) / ('solve', batch, slot) tuples: batch k lives in slot k % depth, and the copy
    of batch k+depth-1 is issued right before the solve of batch k (the order solve_host_batches enqueues; CPU-testable)."""
    if depth < 1:
        raise ValueError('depth must be >= 1')
    ops, copied = [], 0
    for k in range(n_batches):
        while copied < min(n_batches, k + depth):
            ops.append(('copy', copied, copied % depth))
            copied += 1
        ops.append(('solve', k, k % depth))
    return ops


class _Slot:
    def __init__(self):
        self.p = self.args = self.out = None
        self.ready = self.free = None

    def fit(self, p_host, args_host, device):
        def like(h, d):
            return d if d is not None and d.shape == h.shape and d.dtype == h.dtype else torch.empty(h.shape, dtype=h.dtype, device=device)
        self.p = like(p_host, self.p)
        old = self.args if self.args is not None and len(self.args) == len(args_host) else [None] * len(args_host)
        self.args = [like(h, d) if torch.is_tensor(h) else h for h, d in zip(args_host, old)]
        if self.out is None or self.out.shape != p_host.shape or self.out.dtype != p_host.dtype:
            self.out = torch.empty(p_host.shape, dtype=p_host.dtype).pin_memory()


def solve_host_batches(solver, batches, function, jac_function=None, wvec=None, device=None, depth=2, **solver_kw):
    """Generator: yields ``(p_final_host, n_iterations)`` per batch, in order.

    solver       lsq_lma_parallel | lsq_gna_parallel | gradient_descent_parallel (this package's drop-ins)
    batches      iterable of (p0_host [B,P], args_host tuple) — CPU tensors, ideally pinned
    wvec         device or host tensor, copied once
    depth        device slots (2 = double buffering)
    solver_kw    forwarded to the solver (ftol, ptol, gtol, max_iter, meth, l, fused, ...)
    """
    if not torch.cuda.is_available():
        raise _lib.NllsError('solve_host_batches needs a CUDA device (no CPU fallback by design)')
    device = torch.device('cuda', torch.cuda.current_device()) if device is None else torch.device(device)
    compute = torch.cuda.current_stream(device)
    copy_stream = torch.cuda.Stream(device)
    if wvec is not None:
        wvec = wvec.to(device)
    slots = [_Slot() for _ in range(depth)]
    it = iter(batches)
    pending = []                                   # batches copied (or being copied) but not solved yet: slot indices
    n_copied = 0
    exhausted = False

    def enqueue_copy():
        nonlocal n_copied, exhausted
        try:
            p_host, args_host = next(it)
        except StopIteration:
            exhausted = True
            return
        s = slots[n_copied % depth]
        with torch.cuda.device(device):
            s.fit(p_host, tuple(args_host), device)
            with torch.cuda.stream(copy_stream):
                if s.free is not None:
                    copy_stream.wait_event(s.free)   # the previous solve in this slot (and its result read) is over
                s.p.copy_(p_host, non_blocking=True)
                for d, h in zip(s.args, args_host):
                    if torch.is_tensor(h):
                        d.copy_(h, non_blocking=True)
                s.ready = torch.cuda.Event()
                s.ready.record(copy_stream)
        pending.append(n_copied % depth)
        n_copied += 1

    while True:
        while not exhausted and len(pending) < depth:
            enqueue_copy()
        if not pending:
            return
        s = slots[pending.pop(0)]
        compute.wait_event(s.ready)
        with torch.cuda.device(device):              # the solver launches on the current device's stream
            coeffs = solver(s.p, function, jac_function, args=tuple(s.args), wvec=wvec, **solver_kw)
            s.out.copy_(s.p, non_blocking=True)
            s.free = torch.cuda.Event()
            s.free.record(compute)
        # top the pipe up BEFORE blocking on this batch's result, so the next copy is already running
        while not exhausted and len(pending) < depth - 1:
            enqueue_copy()
        s.free.synchronize()
        yield s.out, len(coeffs)


# ------------------------------------------------------------------------------------ host memory placement
def numa_node_of_gpu(index):
    """NUMA node the GPU's PCIe root hangs off (sysfs), or None when it cannot be determined."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(':', 1)
        with open(f'/sys/bus/pci/devices/{dom[-4:].lower()}:{rest.lower()}/numa_node') as f:
            node = int(f.read())
        return node if node >= 0 else None
    except Exception:                                             # noqa: BLE001
        return None


class host_memory_near:
    """Context manager: pinned buffers allocated inside it are placed on the NUMA node of GPU ``index``
    (set_mempolicy(MPOL_PREFERRED) on the calling thread, which the pinned-memory allocation inherits), so that
    with one process per GPU every rank's host->device copies read node-local memory instead of all ranks
    pulling from the node the launcher happened to start on.  Best effort: a no-op when the node is unknown or
    the call is not permitted; ``.node`` / ``.ok`` say what happened."""
    _SYS_SET_MEMPOLICY = {'x86_64': 238, 'aarch64': 237}

    def __init__(self, index):
        self.node, self.ok = numa_node_of_gpu(index), False

    def _policy(self, mode, node):
        import ctypes
        import platform
        nr = self._SYS_SET_MEMPOLICY.get(platform.machine())
        if nr is None:
            return False
        libc = ctypes.CDLL(None, use_errno=True)
        if node is None:
            return libc.syscall(nr, mode, None, 0) == 0
        mask = (ctypes.c_ulong * 4)()
        mask[node // 64] = 1 << (node % 64)
        return libc.syscall(nr, mode, mask, 4 * 64 + 1) == 0

    def __enter__(self):
        if self.node is not None:
            try:
                self.ok = self._policy(1, self.node)              # MPOL_PREFERRED
            except Exception:                                     # noqa: BLE001
                self.ok = False
        return self

    def __exit__(self, *exc):
        if self.ok:
            try:
                self._policy(0, None)                             # MPOL_DEFAULT
            except Exception:                                     # noqa: BLE001
                pass
        return False
