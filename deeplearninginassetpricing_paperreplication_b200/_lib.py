"""ctypes binding of ``libdlap_b200.so`` (C ABI declared in ``include/dlap.h``).

The library is built in-tree by ``build()`` (``nvcc -gencode arch=compute_100a,code=sm_100a``)
and loaded with ``ctypes.CDLL``.  There is no fallback of any kind: if the shared object is
missing, or a call returns a non-zero status, a ``RuntimeError`` is raised.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
import threading

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libdlap_b200.so")
OBJ_DIR = os.path.join(PKG_DIR, "build")          # objects: git- and gpurun-ignored
SOURCES = ("capi.cu", "cell_fwd.cu", "cell_bwd.cu", "reduce.cu", "step_reduce.cu", "macro_lstm.cu", "adam.cu", "ingest.cu", "panel_image.cu", "tc_selftest.cu",
           "cell_fwd_tc.cu", "cell_bwd_tc.cu", "panel_index.cu")
NVCC_FLAGS = ("-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC")

DLAP_PREC_FP32 = 0
DLAP_PREC_BF16X3 = 1
DLAP_REDUCE_NO_DEMEAN = 1
DLAP_REDUCE_Q_DENSE = 2
DLAP_REDUCE_STOCK_SHARD = 4


class DlapDims(C.Structure):
    _fields_ = [("T", C.c_int32), ("N", C.c_int32), ("F", C.c_int32), ("M", C.c_int32),
                ("H", C.c_int32)]


class DlapPlan(C.Structure):
    _fields_ = [("dims", DlapDims), ("tpm", C.c_int32), ("sms", C.c_int32), ("grid_cell", C.c_int32),
                ("cooperative", C.c_int32), ("col_splits", C.c_int32), ("ldw1", C.c_int32), ("ldwq", C.c_int32),
                ("col_part_bytes", C.c_int64), ("loss_part_bytes", C.c_int64), ("row_part_bytes", C.c_int64),
                ("bar_bytes", C.c_int64),
                ("wpart_bytes", C.c_int64), ("dadc_part_bytes", C.c_int64),
                ("lstm_save_bytes", C.c_int64), ("dpre_bytes", C.c_int64), ("index_bytes", C.c_int64),
                ("image_bytes", C.c_int64)]


class DlapTensors(C.Structure):
    """Twelve device pointers in the reference's state_dict order."""
    FIELDS = ("w_ih", "w_hh", "b_ih", "b_hh", "w1", "b1", "w2", "b2", "w3", "b3", "wq", "bq")
    _fields_ = [(n, C.c_void_p) for n in FIELDS]


class DlapMacroIO(C.Structure):
    """dlap_macro_io: the macro side of a per-cell call (include/dlap.h)."""
    _fields_ = [("macro", C.c_void_p), ("series_T", C.c_int32), ("month0", C.c_int32), ("month_stride", C.c_int32),
                ("tables_ready", C.c_int32), ("h0", C.c_void_p), ("c0", C.c_void_p), ("lstm_save", C.c_void_p), ("a_t", C.c_void_p),
                ("c_t", C.c_void_p), ("hn", C.c_void_p), ("cn", C.c_void_p), ("sync", C.c_void_p)]


ADAM_MAX_TENSORS = 16


class DlapAdamTensors(C.Structure):
    """dlap_adam_tensors: parameter / gradient / moment pointers of one optimiser step."""
    _fields_ = [("count", C.c_int32), ("param", C.c_void_p * ADAM_MAX_TENSORS), ("grad", C.c_void_p * ADAM_MAX_TENSORS),
                ("exp_avg", C.c_void_p * ADAM_MAX_TENSORS), ("exp_avg_sq", C.c_void_p * ADAM_MAX_TENSORS),
                ("step", C.c_void_p * ADAM_MAX_TENSORS), ("numel", C.c_int64 * ADAM_MAX_TENSORS)]


class DlapPeers(C.Structure):
    """dlap_peers: every rank's exchange buffer as this device addresses it (stock shards over NVLink peer memory)."""
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("capacity", C.c_int32), ("buf", C.c_void_p * 8),
                ("epoch", C.c_void_p)]


_P = C.c_void_p
_MIO = C.POINTER(DlapMacroIO)
_PLAN = C.POINTER(DlapPlan)
_TENS = C.POINTER(DlapTensors)
_I = C.c_int

# name -> argtypes; every function returns int status (except the two introspection calls)
SIGNATURES = {
    "dlap_make_plan": (C.POINTER(DlapDims), _I, _PLAN),
    "dlap_macro_fwd": (_PLAN, _P, _TENS, _P, _P, _P, _P, _P, _P, _P, _P),
    "dlap_panel_index": (_PLAN, _P, _P, _P),
    "dlap_panel_pull": (_PLAN, _P, _P, _P, _I, _I, _P),
    "dlap_panel_ingest": (_PLAN, _P, _I, C.c_float, _P, _P, _P, _I, _P),
    "dlap_macro_zscore": (_P, _I, _I, _P, _P, _I, _I, _P, _P),
    "dlap_panel_image": (_PLAN, _P, _P, _P, _P),
    "dlap_cell_fwd": (_PLAN, _P, _P, _P, _P, _TENS, _MIO, _P, _P, _I, C.c_float, C.c_uint64, _P, _P),
    "dlap_step_reduce": (_PLAN, _P, _P, _P, _P, _P, _I, _I, _P, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _MIO, _P, _P, _I,
                         C.POINTER(DlapPeers), _P),
    "dlap_loss_finish": (_PLAN, _P, _I, _P, _P, _P, _P, _P, _P),
    "dlap_loss_finish_noalloc": (_PLAN, _P, _I, _P, _P, _P, _P),
    "dlap_cell_bwd": (_PLAN, _P, _P, _P, _P, _P, _TENS, _MIO, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _P, _P, _I,
                      C.c_float, C.c_uint64, _P, _P, _P, _P),
    "dlap_grad_finish": (_PLAN, _P, _P, _P, _I, _I, _I, _TENS, _P, _I, _P),
    "dlap_macro_bwd": (_PLAN, _P, _TENS, _P, _P, _P, _P, _P, _I, _I, _TENS, _P, _P, _P, _I, _P, C.POINTER(DlapPeers), _P),
    "dlap_weights": (_PLAN, _P, _P, _P, _I, _I, _P, _P),
    "dlap_clip_adam": (C.POINTER(DlapAdamTensors), C.c_float, C.c_float, C.c_float, C.c_float, C.c_float, _P, _P),
    "dlap_debug_timeline": (_P,),
    "dlap_umma_selftest": (_P, _I, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P),
    "dlap_umma_ts_selftest": (_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P),
    "dlap_umma_cycles": (_P, _I, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P, _P, _P),
}
EXPORTS = ("dlap_abi_version", "dlap_error_string") + tuple(SIGNATURES)

_lock = threading.Lock()
_lib = None


def _needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "dlap.h")]
    return any(os.path.getmtime(d) > t for d in deps if os.path.isfile(d))


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every CUDA source for sm_100a into the in-tree shared library: one object per source
    (compiled side by side, and only when the source or a header is newer than the object), then one link."""
    if not force and not _needs_build():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libdlap_b200.so")
    from concurrent.futures import ThreadPoolExecutor
    os.makedirs(OBJ_DIR, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers.append(os.path.join(ROOT, "include", "dlap.h"))
    hdr_time = max(os.path.getmtime(h) for h in headers)
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + os.environ.get("DLAP_NVCC_EXTRA", "").split()

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, src[:-3] + ".o")
        path = os.path.join(CSRC, src)
        if not force and os.path.exists(obj) and os.path.getmtime(obj) > max(hdr_time, os.path.getmtime(path)):
            return obj, ""
        cmd = [nvcc, *flags, "-c", "-o", obj, path]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + res.stdout + res.stderr)
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 1)) as pool:
        done = list(pool.map(compile_one, SOURCES))
    if verbose:
        print("".join(log for _, log in done))
    res = subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB_PATH,
                          *[o for o, _ in done]], capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    return LIB_PATH


def load():
    """The loaded library, with argtypes set.  Raises if it has not been built."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: the CUDA library has not been built "
                "(run `python -c 'import __graft_entry__ as g; g.build()'`). There is no CPU fallback.")
        lib = C.CDLL(LIB_PATH)
        lib.dlap_abi_version.restype = C.c_int
        lib.dlap_error_string.restype = C.c_char_p
        lib.dlap_error_string.argtypes = (C.c_int,)
        for name, argt in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.argtypes = argt
            fn.restype = C.c_int
        _lib = lib
        return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().dlap_error_string(rc).decode()
        raise RuntimeError(f"{what} failed with status {rc}: {msg}")


def call(name: str, *args) -> None:
    check(getattr(load(), name)(*args), name)
