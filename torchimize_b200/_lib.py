"""ctypes binding of libnlls_b200.so (C ABI: include/nlls_b200.h).  Plumbing only: torch owns
device memory and streams, the library owns every kernel."""
import ctypes
import functools
import os
import subprocess
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('NLLS_LIB') or os.path.join(_HERE, 'libnlls_b200.so')   # NLLS_LIB: A/B an older build
CSRC = os.path.join(_HERE, 'csrc')

F32, F64 = 0, 1
SOLVER_GN, SOLVER_LM, SOLVER_GD = 0, 1, 2
LM_LEVENBERG, LM_MARQUARDT = 0, 1
MODEL_EMG, MODEL_SKEWGAUSS, MODEL_MULTI_EMG, MODEL_USER = 0, 1, 2, 100
CTRL_BYTES = 48
CTRL_E_PEER_TIMEOUT = 1

_vp, _i, _ll, _d = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_double


class Track(ctypes.Structure):
    """nlls_track (include/nlls_b200.h): per-problem tracking / retirement; device pointers in a host struct"""
    _fields_ = [('pstat', _vp), ('active', _vp), ('count', _vp), ('retire', _i), ('reserved', _i), ('B', _ll)]


_tp = ctypes.POINTER(Track)
# entry points that take `const nlls_track* track` as their last argument before `stream`; call() inserts it (track=None: none)
TRACKED = ('nlls_lm_solve', 'nlls_lm_update', 'nlls_gn_update', 'nlls_gd_update', 'nlls_iter_end', 'nlls_model_normal_eq',
           'nlls_model_normal_eq_gda', 'nlls_model_lm_trial', 'nlls_lm_update_spec', 'nlls_model_normal_eq_static')

# name -> argtypes; mirrors include/nlls_b200.h one to one (tests check the exported symbols)
SIGNATURES = {
    'nlls_abi_version': [],
    'nlls_device_check': [],
    'nlls_set_force_direct': [_i],
    'nlls_normal_eq': [_i, _vp, _vp, _vp, _vp, _i, _ll, _i, _i, _i, _vp, _vp, _vp, _vp, _vp],
    'nlls_normal_eq_gda': [_i, _vp, _vp, _vp, _vp, _i, _ll, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp],
    'nlls_cost_stat': [_i, _vp, _ll, _i, _i, _vp, _vp, _vp],
    'nlls_lm_solve': [_i, _vp, _vp, _vp, _vp, _vp, _i, _ll, _i, _vp, _vp, _vp, _tp, _vp],
    'nlls_lm_update': [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _d, _d, _d, _d, _d, _d, _d, _ll, _i, _vp, _vp, _vp, _tp, _vp],
    'nlls_gn_update': [_i, _vp, _vp, _vp, _vp, _d, _d, _d, _d, _i, _ll, _i, _vp, _vp, _vp, _tp, _vp],
    'nlls_gd_update': [_i, _vp, _vp, _vp, _vp, _d, _d, _d, _d, _i, _ll, _i, _vp, _vp, _vp, _tp, _vp],
    'nlls_iter_end': [_vp, _i, _d, _tp, _vp],
    'nlls_iter_end_peer': [_vp, _i, _d, _vp, _i, _i, _ll, _d, _vp],
    'nlls_sumsq_partials': [],
    'nlls_sumsq': [_i, _vp, _vp, _ll, _vp, _vp],
    'nlls_model_eval': [_i, _i, _vp, _i, _vp, _vp, _ll, _ll, _ll, _i, _i, _vp, _vp, _vp, _vp],
    'nlls_model_normal_eq': [_i, _i, _vp, _i, _vp, _vp, _ll, _ll, _vp, _i, _ll, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _tp, _vp],
    'nlls_model_normal_eq_static': [_i, _i, _vp, _i, _vp, _vp, _ll, _ll, _vp, _i, _ll, _i, _i, _vp, _vp, _vp, _vp, _tp, _vp],
    'nlls_model_normal_eq_gda': [_i, _i, _vp, _i, _vp, _vp, _ll, _ll, _vp, _i, _ll, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _tp, _vp],
    'nlls_model_cost_stat': [_i, _i, _vp, _i, _vp, _vp, _ll, _ll, _ll, _i, _i, _vp, _vp, _vp],
    'nlls_model_lm_trial': [_i, _i, _vp, _i, _vp, _vp, _ll, _ll, _vp, _i, _ll, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _tp, _vp],
    'nlls_lm_update_spec': [_i, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _d, _d, _d, _d, _d, _d, _d, _ll, _i,
                            _vp, _vp, _vp, _tp, _vp],
}

_lock = threading.Lock()
_lib = None
launches = 0          # kernels enqueued through this binding (bench.py reports it as gpu_launches)
_LAUNCHES_PER_CALL = {'nlls_gd_update': 2}


class NllsError(RuntimeError):
    pass


def build(verbose=False):
    """Compile csrc/*.cu for sm_100a into libnlls_b200.so (in-tree, next to this file)."""
    r = subprocess.run(['make', '-C', CSRC, '-j8'], capture_output=True, text=True)
    if r.returncode != 0:
        raise NllsError('building libnlls_b200.so failed:\n' + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


def load():
    """Load the library; raises loudly when it is missing (no fallback path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise NllsError(f'{LIB_PATH} not found: run `python -c "import __graft_entry__ as g; g.build()"` '
                                f'or `make -C {CSRC}`; torchimize_b200 has no CPU or torch fallback')
            lib = ctypes.CDLL(LIB_PATH)
            for name, argtypes in SIGNATURES.items():
                fn = getattr(lib, name)
                fn.argtypes = argtypes
                fn.restype = _i
            lib.nlls_last_error.restype = ctypes.c_char_p
            lib.nlls_last_error.argtypes = []
            if os.environ.get('NLLS_FORCE_DIRECT') == '1':     # debug: keep K1 off the bulk-copy staged kernel
                lib.nlls_set_force_direct(1)
            _lib = lib
    return _lib


def source_hash():
    """sha1 (16 hex digits) of the kernel sources and the C-ABI header: identifies the build an ncu capture belongs to"""
    import hashlib
    h = hashlib.sha1()
    for fn in sorted(os.listdir(CSRC)):
        if fn.endswith(('.cu', '.cuh')) or fn == 'Makefile':
            with open(os.path.join(CSRC, fn), 'rb') as f:
                h.update(f.read())
    with open(os.path.join(_HERE, '..', 'include', 'nlls_b200.h'), 'rb') as f:
        h.update(f.read())
    return h.hexdigest()[:16]


JIT_DIR = os.path.join(_HERE, '_jit')
PLUGIN_ENTRY_POINTS = ('nlls_model_eval', 'nlls_model_normal_eq', 'nlls_model_normal_eq_gda', 'nlls_model_cost_stat',
                       'nlls_model_lm_trial')
_plugins = {}


def build_plugin(source: str, name: str = 'user', verbose=False):
    """Compile the fused-model kernels (csrc/fused_models.cu) for ONE user-supplied device functor
    ``nlls::UserModel<T>`` into an in-tree plugin library exporting the nlls_model_* entry points of
    include/nlls_b200.h.  Cached by a hash of the functor source and of the kernel sources."""
    import hashlib
    h = hashlib.sha1(source.encode())
    for fn in ('fused_models.cu', 'models.cuh', 'common.cuh', 'rowacc.cuh', 'api.cu'):
        with open(os.path.join(CSRC, fn), 'rb') as f:
            h.update(f.read())
    with open(os.path.join(_HERE, '..', 'include', 'nlls_b200.h'), 'rb') as f:
        h.update(f.read())
    safe = ''.join(ch if ch.isalnum() else '_' for ch in name)[:40]
    out_dir = os.path.join(JIT_DIR, f'{safe}_{h.hexdigest()[:16]}')
    so = os.path.join(out_dir, 'libnlls_user.so')
    if os.path.exists(so):
        return so
    os.makedirs(out_dir, exist_ok=True)
    # Several ranks may compile the same new model at once (torchrun): one compiles under an exclusive lock on the build
    # directory, the others wait and find the library.  Header and output are per-process temp names; only finished files
    # are renamed into place; a failure removes this process's temp files and nothing else.
    import fcntl
    with open(os.path.join(out_dir, '.lock'), 'w') as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if os.path.exists(so):
                return so
            hdr = os.path.join(out_dir, f'user_model.{os.getpid()}.cuh')
            tmp = so + f'.tmp{os.getpid()}'
            with open(hdr, 'w') as f:
                f.write(source)
            cmd = ['nvcc', '-O3', '-std=c++17', '-gencode', 'arch=compute_100a,code=sm_100a', '-lineinfo', '-Xcompiler', '-fPIC',
                   f'-DNLLS_USER_MODEL_HEADER="{hdr}"', '-I', CSRC, '-shared', '-o', tmp,
                   os.path.join(CSRC, 'fused_models.cu'), os.path.join(CSRC, 'api.cu'), '-lcudart']
            try:
                r = subprocess.run(cmd, capture_output=True, text=True)
            except FileNotFoundError as e:
                raise NllsError('compile_model needs nvcc on PATH (CUDA 12.8+ with sm_100a support): ' + str(e)) from e
            finally:
                if os.path.exists(hdr) and not os.path.exists(tmp):
                    os.unlink(hdr)
            if r.returncode != 0:
                for leftover in (hdr, tmp):
                    if os.path.exists(leftover):
                        os.unlink(leftover)
                raise NllsError(f'compiling user model {name!r} failed:\n' + r.stdout[-3000:] + r.stderr[-6000:])
            os.replace(hdr, os.path.join(out_dir, 'user_model.cuh'))     # kept beside the library for reference
            os.replace(tmp, so)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    if verbose:
        print('built', so)
    return so


def load_plugin(path):
    """ctypes handle of a plugin library (RTLD_LOCAL: its nlls_* symbols do not clash with libnlls_b200.so)."""
    with _lock:
        if path not in _plugins:
            lib = ctypes.CDLL(path, mode=ctypes.RTLD_LOCAL)
            for name in PLUGIN_ENTRY_POINTS:
                fn = getattr(lib, name)
                fn.argtypes = SIGNATURES[name]
                fn.restype = _i
            lib.nlls_last_error.restype = ctypes.c_char_p
            lib.nlls_last_error.argtypes = []
            lib.nlls_user_model_params.restype = _i
            lib.nlls_set_force_direct.argtypes = [_i]
            _plugins[path] = lib
    return _plugins[path]


def dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float64:
        return F64
    if t.dtype == torch.float32:
        return F32
    raise NllsError(f'unsupported dtype {t.dtype}: the B200 core computes in float32 or float64')


def ptr(t):
    return None if t is None else t.data_ptr()


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NllsError('torchimize_b200 runs on CUDA tensors only (no CPU fallback by design); got a '
                            f'{t.device} tensor')


def stream_ptr():
    return torch.cuda.current_stream().cuda_stream


def on_device_of(argpos, name='p'):
    """Decorator: run the function with the CUDA device of its tensor argument (position `argpos` or keyword `name`)
    made current.  The library launches on the CURRENT device's stream and sizes its grids from the current device;
    the reference accepts tensors on any device, so a solve of a cuda:1 tensor while cuda:0 is current has to switch."""
    def deco(fn):
        @functools.wraps(fn)
        def wrapped(*a, **kw):
            t = a[argpos] if len(a) > argpos else kw.get(name)
            if torch.is_tensor(t) and t.is_cuda and t.device.index != torch.cuda.current_device():
                with torch.cuda.device(t.device):
                    return fn(*a, **kw)
            return fn(*a, **kw)
        return wrapped
    return deco


_timing = None        # when a list: (name, start_event, end_event) per call (bench.py's per-kernel breakdown)


def enable_timing(on=True):
    global _timing
    _timing = [] if on else None


def timing_records():
    """[(name, milliseconds)] in call order; synchronises."""
    torch.cuda.synchronize()
    return [(n, a.elapsed_time(b)) for n, a, b in (_timing or [])]


def call(name, *args, lib=None, track=None):
    """Invoke an entry point on the current stream of the current device; raise on error.
    lib: a plugin library handle (load_plugin) for the nlls_model_* entry points of a user model.
    track: a Track (nlls_track) for the entry points in TRACKED — inserted before their `stream` argument (None = none)."""
    global launches
    lib = lib or load()
    if name in TRACKED:
        args = args[:-1] + (ctypes.byref(track) if track is not None else None, args[-1])
    if _NVTX:                                          # NLLS_NVTX=1: one NVTX range per entry point (K1..K4 in nsys / ncu timelines)
        torch.cuda.nvtx.range_push(name)
        try:
            return _call(lib, name, args)
        finally:
            torch.cuda.nvtx.range_pop()
    return _call(lib, name, args)


_NVTX = os.environ.get('NLLS_NVTX') == '1'


def _call(lib, name, args):
    global launches
    if _timing is not None:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        rc = getattr(lib, name)(*args)
        b.record()
        _timing.append((name, a, b))
    else:
        rc = getattr(lib, name)(*args)
    if rc != 0:
        raise NllsError(f'{name} failed (code {rc}): {lib.nlls_last_error().decode()}')
    launches += _LAUNCHES_PER_CALL.get(name, 1)
    return rc


def new_ctrl(device):
    """Zeroed device-resident nlls_ctrl block (48 bytes) as an int32 view + fp64 view."""
    raw = torch.zeros(CTRL_BYTES // 8, dtype=torch.float64, device=device)
    return raw
