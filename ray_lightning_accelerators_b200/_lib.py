"""ctypes binding of the b2c C-ABI (include/b2c.h).

The CUDA library is the product; this module only loads it.  There is NO
fallback: if ``libb2c.so`` is missing or a symbol is absent the import of the
GPU path fails loudly (``B2cLibraryError``).
"""
import ctypes
import os
import subprocess
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libb2c.so")
CSRC_DIR = os.path.join(_HERE, "csrc")
HEADER_PATH = os.path.join(os.path.dirname(_HERE), "include", "b2c.h")

# mirrors of the enums in include/b2c.h
F32, BF16, U8 = 0, 1, 2
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS = 0, 1, 2, 3
PARAM_CROSSOVER_BYTES = 0
PARAM_MAX_BLOCKS_ONESHOT = 1
PARAM_MAX_BLOCKS_TWOSHOT = 2
PARAM_THREADS = 3
PARAM_TIMEOUT_MS = 4
PARAM_TWOSHOT_VARIANT = 5
PARAM_MAX_BLOCKS_BCAST = 6
PARAM_ZEROCOPY = 7
PARAM_ZEROCOPY_MIN_BYTES = 8
PARAM_NVLS_MIN_BYTES = 9
PARAM_MAX_BLOCKS_NVLS = 10
PARAM_UNROLL_NVLS = 11
PARAM_UNROLL_ZC = 12
PARAM_PDL = 13
PARAM_THREADS_NVLS = 14
PARAM_NVLS_STAGE_BYTES = 15
PARAM_NVLS_LEAN = 16
PARAM_BCAST_MC = 17
PARAM_MC_ACQ_POLL = 18
PARAM_BCAST_INPLACE_MIN_BYTES = 19
PHASE_RING = 32
REG_BLOB_BYTES = 128
NVLS_BLOB_BYTES = 64
NOALLOC_MAX_BYTES = 256 * 1024
MAX_WORLD = 8
MAX_BLOCKS = 256
EXPORT_BLOB_BYTES = 256
E_INVAL, E_STATE, E_TOOBIG, E_TIMEOUT, E_NOPEER, E_UNSUPPORTED = -1, -2, -3, -4, -5, -6


class B2cLibraryError(RuntimeError):
    """libb2c.so is missing / unloadable / incomplete."""


class B2cError(RuntimeError):
    """A b2c_* call returned non-zero."""

    def __init__(self, code, message):
        super().__init__(f"b2c error {code}: {message}")
        self.code = code


OPT_SGD, OPT_ADAM, OPT_ADAMW = 1, 2, 3


class Opt(ctypes.Structure):
    """b2c_opt of include/b2c.h."""
    _fields_ = [("kind", ctypes.c_int), ("nesterov", ctypes.c_int), ("first_step", ctypes.c_int),
                ("lr", ctypes.c_double), ("momentum", ctypes.c_double), ("dampening", ctypes.c_double),
                ("weight_decay", ctypes.c_double), ("beta1", ctypes.c_double), ("beta2", ctypes.c_double),
                ("eps", ctypes.c_double), ("bias_correction1", ctypes.c_double), ("bias_correction2", ctypes.c_double)]


class Stats(ctypes.Structure):
    _fields_ = [
        ("calls_allreduce", ctypes.c_uint64),
        ("calls_broadcast", ctypes.c_uint64),
        ("calls_barrier", ctypes.c_uint64),
        ("launches", ctypes.c_uint64),
        ("bytes_allreduce", ctypes.c_uint64),
        ("bytes_broadcast", ctypes.c_uint64),
        ("oneshot_launches", ctypes.c_uint64),
        ("twoshot_launches", ctypes.c_uint64),
        ("zerocopy_launches", ctypes.c_uint64),
        ("nvls_launches", ctypes.c_uint64),
        ("last_algo", ctypes.c_uint32),
        ("last_blocks", ctypes.c_uint32),
        ("last_threads", ctypes.c_uint32),
        ("device_status", ctypes.c_uint32),
        ("device_status_detail", ctypes.c_uint32 * 3),
        ("reserved", ctypes.c_uint32),
        ("bcast_mc_launches", ctypes.c_uint64),
        ("noop_calls", ctypes.c_uint64),
        ("fused_steps", ctypes.c_uint64),
    ]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in ("device_status_detail", "reserved")}
        d["device_status_detail"] = list(self.device_status_detail)
        return d


_c = ctypes
_vp, _sz, _int = _c.c_void_p, _c.c_size_t, _c.c_int
_pp = _c.POINTER(_c.c_void_p)

# name -> (restype, argtypes); kept in the order of include/b2c.h
SIGNATURES = {
    "b2c_version": (_int, []),
    "b2c_last_error": (_c.c_char_p, []),
    "b2c_comm_create": (_int, [_int, _int, _int, _sz, _pp]),
    "b2c_comm_export": (_int, [_vp, _vp, _c.POINTER(_sz)]),
    "b2c_comm_import": (_int, [_vp, _vp, _sz]),
    "b2c_comm_destroy": (_int, [_vp]),
    "b2c_group_create_local": (_int, [_int, _int, _sz, _pp]),
    "b2c_group_destroy_local": (_int, [_pp, _int]),
    "b2c_allreduce": (_int, [_vp, _vp, _sz, _int, _int, _c.c_float, _int, _vp]),
    "b2c_broadcast": (_int, [_vp, _vp, _sz, _int, _vp]),
    "b2c_barrier": (_int, [_vp, _vp]),
    "b2c_group_allreduce": (_int, [_pp, _pp, _sz, _int, _int, _c.c_float, _int, _vp]),
    "b2c_group_broadcast": (_int, [_pp, _pp, _sz, _int, _vp]),
    "b2c_group_barrier": (_int, [_pp, _vp]),
    "b2c_register_export": (_int, [_vp, _vp, _sz, _vp, _c.POINTER(_sz)]),
    "b2c_register_import": (_int, [_vp, _vp, _sz, _vp, _sz]),
    "b2c_deregister": (_int, [_vp, _vp]),
    "b2c_nvls_export": (_int, [_vp, _sz, _vp, _c.POINTER(_sz)]),
    "b2c_nvls_import": (_int, [_vp, _vp, _sz]),
    "b2c_nvls_bind": (_int, [_vp]),
    "b2c_nvls_arena": (_int, [_vp, _pp, _c.POINTER(_sz)]),
    "b2c_nvls_alloc": (_int, [_vp, _sz, _pp]),
    "b2c_nvls_free": (_int, [_vp, _vp]),
    "b2c_nvls_pool_bind": (_int, [_vp]),
    "b2c_pool_alloc": (_vp, [_c.c_ssize_t, _int, _vp]),
    "b2c_pool_free": (None, [_vp, _c.c_ssize_t, _int, _vp]),
    "b2c_fused_shard": (_int, [_sz, _int, _int, _c.POINTER(_sz), _c.POINTER(_sz)]),
    "b2c_fused_step": (_int, [_vp, _vp, _vp, _sz, _int, _c.c_float, _c.POINTER(Opt), _vp, _vp, _vp, _sz, _int, _vp]),
    "b2c_group_fused_step": (_int, [_pp, _pp, _pp, _sz, _int, _c.c_float, _c.POINTER(Opt), _pp, _pp, _vp]),
    "b2c_allreduce_local_noalloc": (_int, [_pp, _int, _sz, _int, _c.c_float, _vp]),
    "b2c_noalloc_status": (_int, [_c.POINTER(_c.c_uint32 * 4), _int]),
    "b2c_set_param": (_int, [_vp, _int, _c.c_long]),
    "b2c_get_param": (_int, [_vp, _int, _c.POINTER(_c.c_long)]),
    "b2c_get_stats": (_int, [_vp, _c.POINTER(Stats)]),
    "b2c_comm_arena": (_int, [_vp, _pp, _c.POINTER(_sz)]),
    "b2c_get_phase_ns": (_int, [_vp, _c.POINTER(_c.c_uint64 * 4)]),
    "b2c_get_phase_ring": (_int, [_vp, _c.POINTER(_c.c_uint64), _int, _c.POINTER(_c.c_uint32)]),
    "b2c_comm_info": (_int, [_vp, _c.POINTER(_int), _c.POINTER(_int), _c.POINTER(_int)]),
}

_lib = None
_lock = threading.Lock()


def build(force=False, verbose=False):
    """Compile libb2c.so in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    if force and os.path.exists(LIB_PATH):
        os.remove(LIB_PATH)
    proc = subprocess.run(["make", "-C", CSRC_DIR], capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout)
        print(proc.stderr)
    if proc.returncode != 0 or not os.path.exists(LIB_PATH):
        raise B2cLibraryError(f"building libb2c.so failed (make -C {CSRC_DIR}):\n{proc.stdout}\n{proc.stderr}")
    return LIB_PATH


def load():
    """Load libb2c.so and bind every symbol of include/b2c.h.  Raises loudly."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise B2cLibraryError(
                f"{LIB_PATH} not found. The b2c CUDA library is required for use_gpu=True and has no "
                f"fallback; build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                f"or `make -C {CSRC_DIR}`.")
        try:
            lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        except OSError as e:
            raise B2cLibraryError(f"cannot load {LIB_PATH}: {e}") from e
        for name, (res, args) in SIGNATURES.items():
            try:
                fn = getattr(lib, name)
            except AttributeError as e:
                raise B2cLibraryError(f"{LIB_PATH} does not export {name}") from e
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib


def last_error():
    return load().b2c_last_error().decode("utf-8", "replace")


def check(rc):
    if rc != 0:
        raise B2cError(rc, last_error())


# ---------------------------------------------------------------- optional C++ DDP hook (torch extension)
HOOK_SRC = os.path.join(CSRC_DIR, "b2c_torch_hook.cpp")
HOOK_PATH = os.path.join(_HERE, "_b2c_torch_hook.so")
_hook_mod = None


def build_torch_hook(verbose=False):
    """Compile csrc/b2c_torch_hook.cpp in-tree (needs torch headers; no GPU).  The only torch-aware
    binary of the package; libb2c.so itself stays torch-free."""
    import shutil
    from torch.utils.cpp_extension import load
    build_dir = os.path.join(_HERE, "_hook_build")
    os.makedirs(build_dir, exist_ok=True)
    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0a")
    mod = load(name="_b2c_torch_hook", sources=[HOOK_SRC], build_directory=build_dir, with_cuda=True,
               extra_cflags=["-O2"], verbose=verbose)
    shutil.copyfile(mod.__file__, HOOK_PATH)
    return HOOK_PATH


def load_torch_hook():
    """The built extension module, or None when it has not been built (the Python hook is used then)."""
    global _hook_mod
    if _hook_mod is None and os.path.exists(HOOK_PATH):
        import importlib.util
        import torch  # noqa: F401  (its symbols must be loaded first)
        spec = importlib.util.spec_from_file_location("_b2c_torch_hook", HOOK_PATH)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        _hook_mod = mod
    return _hook_mod


def fn_address(name):
    return ctypes.cast(getattr(load(), name), ctypes.c_void_p).value
