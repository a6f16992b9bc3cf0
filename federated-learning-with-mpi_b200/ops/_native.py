"""Build and bind the native library (``libfedb200.so``, sm_100a, C ABI in include/fedb200.h).

The library is plain CUDA runtime code with ``extern "C"`` entry points — no torch headers,
no pybind — so it builds in seconds with ``nvcc -shared`` and binds with ctypes.  It is
built IN-TREE (``federated-learning-with-mpi_b200/_lib/``) so the ``.so`` travels with the
repository snapshot to the GPU box.

``lib()`` is strict: on a machine with a CUDA device a missing or unloadable library raises
(after one build attempt) instead of silently falling back to PyTorch; on a CPU-only
machine it returns ``None`` and the ops run their fp32 PyTorch restatement (ops/reference.py),
which is what the CPU test-suite exercises.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
import sys
import threading
from concurrent.futures import ThreadPoolExecutor
from typing import List, Optional

import numpy as np

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPO = os.path.dirname(_PKG)
CSRC = os.path.join(REPO, "csrc")
INCLUDE = os.path.join(REPO, "include")
LIB_DIR = os.path.join(_PKG, "_lib")
LIB_PATH = os.path.join(LIB_DIR, "libfedb200.so")
OBJ_DIR = os.path.join(REPO, "build", "obj")

NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-diag-suppress", "177",
]

_lock = threading.Lock()
_lib: Optional[ctypes.CDLL] = None


def _nvcc() -> Optional[str]:
    cand = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    return cand if os.path.exists(cand) else None


def sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))


def _deps_mtime() -> float:
    files = sources() + [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    files += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)]
    return max(os.path.getmtime(f) for f in files)


def needs_build() -> bool:
    return not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < _deps_mtime()


def build(force: bool = False, verbose: bool = False, extra_flags: Optional[List[str]] = None) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a and link ``libfedb200.so``.  Cross-compiles on a
    machine without a GPU.  Returns the library path."""
    with _lock:
        if not force and not needs_build():
            return LIB_PATH
        nvcc = _nvcc()
        if nvcc is None:
            raise RuntimeError("nvcc not found; cannot build libfedb200.so")
        os.makedirs(OBJ_DIR, exist_ok=True)
        os.makedirs(LIB_DIR, exist_ok=True)
        flags = NVCC_FLAGS + (extra_flags or [])
        hdr_mtime = max([os.path.getmtime(os.path.join(CSRC, f)) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
                        + [os.path.getmtime(os.path.join(INCLUDE, f)) for f in os.listdir(INCLUDE)])

        def compile_one(src: str) -> str:
            obj = os.path.join(OBJ_DIR, os.path.basename(src) + ".o")
            if not force and os.path.exists(obj) and os.path.getmtime(obj) >= max(os.path.getmtime(src), hdr_mtime):
                return obj
            cmd = [nvcc, *flags, "-I", INCLUDE, "-c", src, "-o", obj]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            r = subprocess.run(cmd, capture_output=True, text=True)
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
            return obj

        with ThreadPoolExecutor(max_workers=min(8, max(1, len(sources())))) as ex:
            objs = list(ex.map(compile_one, sources()))
        tmp = LIB_PATH + ".tmp"
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs, "-lcuda"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
        os.replace(tmp, LIB_PATH)
        global _lib
        _lib = None
        return LIB_PATH


class NativeError(RuntimeError):
    pass


def _cuda_present() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def lib(required: Optional[bool] = None) -> Optional[ctypes.CDLL]:
    """The loaded library, or ``None`` on a CPU-only machine (unless ``required``)."""
    global _lib
    if _lib is not None:
        return _lib
    required = _cuda_present() if required is None else required
    if not required and not os.path.exists(LIB_PATH):
        return None
    if needs_build():
        try:
            build()
        except Exception as e:  # stale or missing and cannot rebuild
            if not os.path.exists(LIB_PATH):
                if required:
                    raise NativeError(f"libfedb200.so is missing and could not be built: {e}") from e
                return None
    try:
        l = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    except OSError as e:
        if required:
            raise NativeError(f"cannot load {LIB_PATH}: {e}") from e
        return None
    _declare(l)
    sizes = (ctypes.c_int32 * 5)()
    l.fedb_struct_sizes(sizes)
    expect = [GROUP_DTYPE.itemsize, FED_DTYPE.itemsize, FEDSTATE_DTYPE.itemsize, EVALJOB_DTYPE.itemsize, TILE_DTYPE.itemsize]
    if list(sizes) != expect:
        raise NativeError(f"struct layout mismatch between include/fedb200.h {list(sizes)} and the ctypes binding {expect}")
    _lib = l
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        msg = _lib.fedb_last_error().decode() if _lib is not None else "?"
        raise NativeError(f"fedb200 native call failed (code {rc}): {msg}")


# ---- struct layouts (must mirror include/fedb200.h) ------------------------------------
MAX_LAYERS = 6
MAX_CLASSES = 8
TILE_ROWS = 32
LOSS_SLOTS = 32

GROUP_DTYPE = np.dtype([
    ("n_layers", "i4"), ("dims", "i4", (MAX_LAYERS + 1,)), ("head", "i4"), ("act", "i4"), ("slab_size", "i4"),
    ("fed", "i4"), ("slab_off", "i8"), ("idx_off", "i8"), ("n_rows", "i4"), ("batch_size", "i4"), ("opt", "i4"),
    ("nesterov", "i4"), ("lr", "f4"), ("beta1", "f4"), ("beta2", "f4"), ("eps", "f4"), ("momentum", "f4"),
    ("alpha", "f4"), ("lr_step_size", "i4"), ("lr_gamma", "f4"), ("agg_num", "f4"), ("agg_den", "f4"),
    ("scratch_off", "i8"), ("max_tiles", "i4"), ("client", "i4"),
], align=True)
FED_DTYPE = np.dtype([
    ("first_group", "i4"), ("n_clients", "i4"), ("slab_size", "i4"), ("n_classes", "i4"), ("global_off", "i8"),
    ("metric_mode", "i4"), ("patience", "i4"), ("atol", "f4"), ("rtol", "f4"), ("max_rounds", "i4"), ("pad_", "i4"),
], align=True)
FEDSTATE_DTYPE = np.dtype([
    ("stop", "i4"), ("stop_pending", "i4"), ("round", "i4"), ("count", "i4"), ("has_anchor", "i4"), ("done", "i4"),
    ("anchor", "f8", (4,)), ("last", "f8", (4,)),
], align=True)
EVALJOB_DTYPE = np.dtype([
    ("group", "i4"), ("n_rows", "i4"), ("slab_off", "i8"), ("idx_off", "i8"), ("conf_slot", "i4"), ("from_global", "i4"),
], align=True)
TILE_DTYPE = np.dtype([("owner", "i4"), ("tile", "i4")], align=True)

HEAD_IDS = {"softmax_ce": 0, "sigmoid_logloss": 1}
ACT_IDS = {"relu": 0, "logistic": 1, "tanh": 2, "identity": 3}
OPT_IDS = {"sgd": 0, "adam_torch": 1, "adam_sklearn": 2}
METRIC_IDS = {"mean_of_local": 0, "pooled": 1}
INIT_IDS = {"zeros": 0, "torch_default": 1, "sklearn_glorot": 2}

_P = ctypes.c_void_p
_I32, _I64, _U64, _F64 = ctypes.c_int32, ctypes.c_int64, ctypes.c_uint64, ctypes.c_double


def _declare(l: ctypes.CDLL) -> None:
    l.fedb_last_error.restype = ctypes.c_char_p
    l.fedb_version.restype = _I32
    l.fedb_struct_sizes.argtypes = [ctypes.POINTER(_I32)]
    l.fedb_struct_sizes.restype = None
    l.fedb_simt_smem_bytes.argtypes = [ctypes.POINTER(_I32), _I32]
    l.fedb_simt_smem_bytes.restype = _I64
    l.fedb_batch_stats_stride.argtypes = [_I32]
    l.fedb_batch_stats_stride.restype = _I32
    sigs = {
        "fedb_init_params": [_P, _I32, _I32, _P, _P, _I32, _U64, _P],
        "fedb_round_begin": [_P, _P, _I32, _P, _I32, _P],
        "fedb_shuffle": [_P, _I32, _P, _P, _P, _P, _U64, _P],
        "fedb_fwd_bwd": [_P, _P, _I32, _P, _P, _I32, _P, _P, _I32, _P, _P, _P, _P, _I64, _P],
        "fedb_reduce_opt": [_P, _I32, _I32, _P, _P, _P, _P, _I32, _P, _P, _P, _P, _P, _I32, _P],
        "fedb_epoch_end": [_P, _I32, _P, _I32, _I32, _F64, _I32, _I32, _P, _P, _P, _P, _P, _P],
        "fedb_eval": [_P, _P, _P, _I32, _P, _P, _P, _I32, _P, _P, _P, _P, _I64, _P],
        "fedb_fedavg": [_P, _P, _P, _I32, _I32, _I32, _P, _P, _P, _P, _P, _I32, _I32, _I32, _P, _P],
    }
    for name, args in sigs.items():
        fn = getattr(l, name)
        fn.argtypes = args
        fn.restype = _I32
    # optional entry points (present once their translation unit is part of the build)
    for name, (args, res) in OPTIONAL_SIGS.items():
        if hasattr(l, name):
            fn = getattr(l, name)
            fn.argtypes = args
            fn.restype = res


# entry points that exist only once their translation unit is part of the build
OPTIONAL_SIGS = {
    "fedb_tc_probe": ([_P, _P, ctypes.POINTER(_I32), _P, _P], _I32),
    "fedb_tc_probe_args_size": ([], _I32),
    "fedb_local_round_tc": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _P], _I32),
    "fedb_tc_smem_bytes": ([], _I64),
    "fedb_tc_max_slab": ([], _I32),
    "fedb_nvls_supported": ([_I32], _I32),
    "fedb_nvls_create": ([_I32, _I64, _P, _P, _P], _I32),
    "fedb_nvls_import": ([_I32, _P], _I32),
    "fedb_nvls_add_device": ([_U64, _I32], _I32),
    "fedb_nvls_bind_map": ([_U64, _I32, _I64, _P, _P, _P], _I32),
    "fedb_nvls_release": ([_U64, _U64, _P, _P, _I64], _I32),
    "fedb_fit_cluster": ([_P, _I32, _I32, _I32, _P, _P, _P, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _F64, _I32,
                          _I32, _I32, _I32, _U64, _I64, _P], _I32),
    "fedb_round_tc_p2p_hostin": ([_P, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _P, _P], _I32),
    "fedb_round_tc_hostin": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _I32, _P, _P, _P, _I32, _P, _P, _P, _P], _I32),
    "fedb_round_tc_p2p": ([_P, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _P, _P], _I32),
    "fedb_round_tc_prof": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _I32, _P, _P, _P, _I32, _P, _P, _P, _P], _I32),
    "fedb_round_tc": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _I32, _P, _P, _P, _I32, _P, _P, _P], _I32),
    "fedb_local_round_tc_prof": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _P, _P], _I32),
    "fedb_tcx_smem_bytes": ([], _I64),
    "fedb_tcx_fits": ([_I32, _I32, _I32, _I32, _I32], _I32),
    "fedb_tcx_rows": ([], _I32),
    "fedb_tcx_max_cluster": ([], _I32),
    "fedb_local_round_tcx": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _I32, _I32, _P, _P], _I32),
    "fedb_round_tcx": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _I32, _P, _P, _P, _I32, _P, _P, _P, _P], _I32),
    "fedb_round_tcx_p2p": ([_P, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _P, _P], _I32),
    "fedb_round_tcx_p2p_many": ([_P, _I32, _I32, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I32, _P, _P, _P, _P], _I32),
    "fedb_fitx_smem_bytes": ([], _I64),
    "fedb_fitx_fits": ([_I32, _I32, _I32, _I32], _I32),
    "fedb_fit_tcx": ([_P, _I32, _P, _P, _P, _P, _I32, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _F64, _I32, _I32, _I32, _I32, _U64, _P, _P], _I32),
    "fedb_noalloc_scratch_floats": ([], _I64),
    "fedb_wide_gather_x": ([_P, _I32, _P, _I32, _P, _P, _I32, _I32, _P], _I32),
    "fedb_wide_shadow": ([_P, _P, _I64, _P], _I32),
    "fedb_wide_head_fwd": ([_P, _I32, _P, _P, _I32, _I32, _P, _P, _P, _P, _P, _P, _P, _I32, _P], _I32),
    "fedb_wide_head_bwd": ([_P, _I32, _P, _P, _P, _P, _P, _I32, _I32, _P, _P, _P, _P, _P, _P, _P], _I32),
    "fedb_wide_bias_opt": ([_P, _I32, _P, _P, _P, _P, _I32, _I32, _I32, _P, _P, _P], _I32),
    "fedb_gemm_args_size": ([], _I32),
    "fedb_gemm_tc": ([_P, _I32, _P], _I32),
    "fedb_gemm_tma": ([_P, _I32, _P], _I32),
    "fedb_gemm_tma2": ([_P, _I32, _P, _P], _I32),
    "fedb_gemm_tma2p": ([_P, _I32, _P, _P], _I32),
    "fedb_gemm_tma2_prof": ([_P, _I32, _P, _P, _P], _I32),
    "fedb_p2p_rank_size": ([], _I32),
    "fedb_fedavg_p2p": ([_P, _I32, _I32, _P], _I32),
    "fedb_symm_alloc": ([_I64, ctypes.POINTER(_P)], _I32),
    "fedb_symm_free": ([_P], _I32),
    "fedb_ipc_get_handle": ([_P, _P], _I32),
    "fedb_ipc_open_handle": ([_P, ctypes.POINTER(_P)], _I32),
    "fedb_ipc_close_handle": ([_P], _I32),
}

P2PRANK_DTYPE = np.dtype([
    ("rank", "i4"), ("world", "i4"), ("n_feds", "i4"), ("n_groups_local", "i4"), ("n_groups_total", "i4"),
    ("conf_slots", "i4"), ("global_total", "i8"), ("groups", "u8"), ("feds", "u8"), ("feds_global", "u8"),
    ("fed_state", "u8"), ("params", "u8"), ("global", "u8"), ("conf", "u8"), ("conf_all", "u8"), ("hist_local", "u8"),
    ("hist_global", "u8"), ("conf_src_rank", "u8"), ("conf_src_slot", "u8"), ("sym", "u8", (8,)),
    ("sym_partial_off", "i8"), ("sym_conf_off", "i8"), ("sym_flag_off", "i8"), ("seq", "u8"), ("bar", "u8"), ("err", "u8"),
    ("timeout_cycles", "i8"), ("write_back", "i4"), ("do_metrics", "i4"), ("begin_next", "i4"), ("pad_", "i4"), ("shadow", "u8"),
    ("nvls_uc_partial", "u8"), ("nvls_mc_partial", "u8"), ("nvls_uc_result", "u8"), ("nvls_mc_result", "u8"),
], align=True)


def ptr(t) -> Optional[int]:
    """Device (or host) address of a torch tensor / None."""
    if t is None:
        return None
    return t.data_ptr()


def cuda_stream() -> int:
    import torch
    return torch.cuda.current_stream().cuda_stream
