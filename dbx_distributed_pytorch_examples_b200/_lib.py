"""ctypes binding of libb200dp.so (the C ABI declared in include/b200dp.h).

There is deliberately no fallback: if the shared library is missing or a call
returns an error, the caller gets an exception.  Build the library with
``python -c "import __graft_entry__ as g; g.build()"`` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
import threading
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_uint32, c_void_p

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC_DIR = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libb200dp.so")
SOURCES = ["augment.cu", "resample.cu", "bn.cu", "pool.cu", "comm.cu", "update.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
]

F32, BF16 = 0, 1
OPT_NONE, OPT_SGD, OPT_ADAM = 0, 1, 2
MAX_WORLD = 8
COMM_HANDLE_BYTES = 128


class B200DPError(RuntimeError):
    """A libb200dp entry point returned a non-zero code."""


class OptArgs(Structure):
    """Mirror of ``b200dp_opt_args`` (include/b200dp.h)."""

    _fields_ = [
        ("kind", c_int), ("lr", c_float), ("momentum", c_float), ("dampening", c_float),
        ("nesterov", c_int), ("first_step", c_int), ("weight_decay", c_float),
        ("beta1", c_float), ("beta2", c_float), ("eps", c_float),
        ("step_size", c_float), ("bias_correction2_sqrt", c_float),
    ]


def _sources():
    hdrs = [os.path.join(CSRC_DIR, "common.cuh"), os.path.join(PKG_DIR, "..", "include", "b200dp.h")]
    return [os.path.join(CSRC_DIR, s) for s in SOURCES], hdrs


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu into libb200dp.so with nvcc for sm_100a (no GPU needed)."""
    srcs, hdrs = _sources()
    if not force and os.path.exists(LIB_PATH):
        newest = max(os.path.getmtime(p) for p in srcs + hdrs if os.path.exists(p))
        if os.path.getmtime(LIB_PATH) >= newest:
            build_fastcall()
            build_autograd()
            return LIB_PATH
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB_PATH] + srcs
    proc = subprocess.run(cmd, capture_output=True, text=True, stdin=subprocess.DEVNULL, timeout=1200)
    if proc.returncode != 0:
        raise B200DPError("nvcc failed:\n" + proc.stdout + proc.stderr)
    if verbose:
        print(proc.stderr)
    check_sass(LIB_PATH)
    build_fastcall(force=True)
    build_autograd(force=True)
    return LIB_PATH


def check_sass(path: str) -> None:
    """Refuse a build that contains the one miscompilation met so far: ptxas 12.9 turned the cp.async ring of
    `bn_bwd_apply_kernel<float, *>` into `LDGSTS [R+UR0], desc[UR1]` with the uniform operands never
    written, which B200 answers with "illegal instruction" (csrc/bn.cu, RowRing::init has the workaround)."""
    import re
    import shutil

    tool = shutil.which("cuobjdump") or os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "bin", "cuobjdump")
    if not os.path.exists(tool):
        return
    proc = subprocess.run([tool, "-sass", path], capture_output=True, text=True, stdin=subprocess.DEVNULL, timeout=600)
    bad, fn = {}, "?"
    for line in proc.stdout.splitlines():
        if "Function :" in line:
            fn = line.split()[-1]
        elif "LDGSTS" in line and re.search(r"\[R\d+\+UR|desc\[UR[01]\]", line):
            bad[fn] = bad.get(fn, 0) + 1
    if bad:
        raise B200DPError("the build contains LDGSTS instructions with unset uniform operands in: " + ", ".join(sorted(bad)))


def _fast_path() -> str:
    import sysconfig

    return os.path.join(PKG_DIR, "_b200dp_fast" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_fastcall(force: bool = False) -> str:
    """gcc-compile csrc/fastcall.c (CPython METH_FASTCALL shim for the per-layer SyncBN calls)."""
    import sysconfig

    src, out = os.path.join(CSRC_DIR, "fastcall.c"), _fast_path()
    if not force and os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(src):
        return out
    cmd = [os.environ.get("B200DP_CC", "gcc"), "-O2", "-shared", "-fPIC", "-I" + sysconfig.get_paths()["include"], "-o", out, src]
    proc = subprocess.run(cmd, capture_output=True, text=True, stdin=subprocess.DEVNULL, timeout=300)
    if proc.returncode != 0:
        raise B200DPError("gcc failed on fastcall.c:\n" + proc.stdout + proc.stderr)
    return out


def _autograd_path() -> str:
    import sysconfig

    return os.path.join(PKG_DIR, "_b200dp_autograd" + sysconfig.get_config_var("EXT_SUFFIX"))


def build_autograd(force: bool = False) -> str:
    """g++-compile csrc/autograd_bn.cpp (the SyncBN autograd node as a C++ torch::autograd::Function,
    host plumbing over the same C ABI) against the installed torch headers, in-tree."""
    import sysconfig

    import torch
    from torch.utils import cpp_extension

    src, out = os.path.join(CSRC_DIR, "autograd_bn.cpp"), _autograd_path()
    if not force and os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(src):
        return out
    inc = cpp_extension.include_paths() + [sysconfig.get_paths()["include"],
                                           os.path.join(os.environ.get("CUDA_HOME", "/usr/local/cuda"), "include")]
    libdir = os.path.join(os.path.dirname(torch.__file__), "lib")
    # the distribution's g++ (the one whose libstdc++ the torch wheel runs against here), not $CXX: an
    # extension built by a toolchain with its own libstdc++ crashed on its first formatted exception
    cmd = ([os.environ.get("B200DP_CXX", "g++"), "-O2", "-std=c++17", "-shared", "-fPIC", "-DTORCH_EXTENSION_NAME=_b200dp_autograd",
            f"-D_GLIBCXX_USE_CXX11_ABI={int(torch._C._GLIBCXX_USE_CXX11_ABI)}"] + [f"-I{i}" for i in inc] +
           [src, "-o", out, f"-L{libdir}", "-lc10", "-lc10_cuda", "-ltorch_cpu", "-ltorch_cuda", "-ltorch",
            "-ltorch_python", f"-Wl,-rpath,{libdir}", "-ldl"])
    proc = subprocess.run(cmd, capture_output=True, text=True, stdin=subprocess.DEVNULL, timeout=1800)
    if proc.returncode != 0:
        raise B200DPError("g++ failed on autograd_bn.cpp:\n" + proc.stdout + proc.stderr)
    return out


_lock = threading.Lock()
_lib = None

# ---- launch accounting -----------------------------------------------------------
# CALLS counts every entry-point call (bench.py reports the kernel launches of a timed
# region from it).  PROFILER, when set, brackets each call with CUDA events on the
# stream the call was given (bench.py's per-kernel roofline pass).
CALLS = {}
PROFILER = None
_COMPUTE = ("b200dp_augment_u8_bf16", "b200dp_resized_crop_u8_bf16", "b200dp_maxpool_fwd", "b200dp_maxpool_bwd", "b200dp_bn_fwd_stats", "b200dp_bn_fwd_stats_noalloc", "b200dp_bn_fwd_apply", "b200dp_bn_inference",
            "b200dp_bn_bwd_reduce", "b200dp_bn_bwd_apply", "b200dp_bucket_pack", "b200dp_allreduce_update",
            "b200dp_comm_barrier", "b200dp_comm_all_gather32", "b200dp_bn_forward", "b200dp_bn_backward")


class _Entry:
    __slots__ = ("name", "fn")

    def __init__(self, name, fn):
        self.name, self.fn = name, fn

    def __call__(self, *args):
        CALLS[self.name] = CALLS.get(self.name, 0) + 1
        if PROFILER is None:
            return self.fn(*args)
        return PROFILER(self.name, self.fn, args)


class _Handle:
    """Attribute access like a CDLL; compute entry points go through _Entry."""

    def __init__(self, cdll):
        for name in _SIGNATURES:
            fn = getattr(cdll, name)
            setattr(self, name, _Entry(name, fn) if name in _COMPUTE else fn)
        self._cdll = cdll

_VP = c_void_p
_SIGNATURES = {
    # name: (restype, argtypes)
    "b200dp_version": (c_int, []),
    "b200dp_kernel_launches": (ctypes.c_ulonglong, []),
    "b200dp_error_string": (c_char_p, [c_int]),
    "b200dp_augment_u8_bf16": (c_int, [_VP, c_int, c_int, c_int, _VP, c_int, c_int, c_int,
                                       POINTER(c_float), POINTER(c_float), _VP, _VP]),
    "b200dp_resized_crop_u8_bf16": (c_int, [_VP, c_int, c_int, c_int, _VP, c_int, c_int,
                                            POINTER(c_float), POINTER(c_float), _VP, _VP]),
    "b200dp_maxpool_fwd": (c_int, [_VP, _VP, _VP, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _VP]),
    "b200dp_maxpool_bwd": (c_int, [_VP, _VP, _VP, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_int, _VP]),
    "b200dp_bn_scratch_bytes": (c_size_t, [c_int]),
    "b200dp_bn_fwd_stats": (c_int, [_VP, c_int, c_int64, c_int, _VP, _VP, _VP, c_uint32, _VP]),
    "b200dp_bn_fwd_stats_noalloc": (c_int, [_VP, c_int, c_int64, c_int, _VP, _VP]),
    "b200dp_bn_fwd_apply": (c_int, [_VP, _VP, _VP, c_int, c_int64, c_int, _VP, _VP, _VP, c_float,
                                    c_float, _VP, _VP, _VP, _VP, _VP, c_int, _VP, c_uint32, _VP, _VP]),
    "b200dp_bn_inference": (c_int, [_VP, _VP, _VP, c_int, c_int64, c_int, _VP, _VP, _VP, _VP,
                                    c_float, c_int, _VP]),
    "b200dp_bn_bwd_reduce": (c_int, [_VP, _VP, _VP, c_int, c_int64, c_int, _VP, _VP, _VP, _VP,
                                     c_int, _VP, _VP, _VP, _VP, _VP, c_uint32, _VP]),
    "b200dp_bn_bwd_apply": (c_int, [_VP, _VP, _VP, _VP, _VP, c_int, c_int64, c_int, _VP, _VP, _VP,
                                    _VP, _VP, _VP, c_int, _VP, c_uint32, _VP]),
    "b200dp_bn_forward": (c_int, [_VP, _VP, _VP, c_int, c_int64, c_int, _VP, _VP, _VP, _VP, c_float, c_float,
                                  _VP, _VP, _VP, c_int, _VP, c_uint32, _VP, _VP, _VP]),
    "b200dp_bn_backward": (c_int, [_VP, _VP, _VP, _VP, _VP, c_int, c_int64, c_int, _VP, _VP, _VP, c_int, _VP, _VP,
                                   _VP, _VP, _VP, c_uint32, _VP]),
    "b200dp_comm_create": (c_int, [c_int, c_int, c_size_t, POINTER(_VP)]),
    "b200dp_comm_create_external": (c_int, [c_int, c_int, c_size_t, POINTER(_VP), _VP, POINTER(_VP)]),
    "b200dp_comm_has_multicast": (c_int, [_VP]),
    "b200dp_comm_export": (c_int, [_VP, _VP]),
    "b200dp_comm_open_peers": (c_int, [_VP, _VP]),
    "b200dp_comm_destroy": (c_int, [_VP]),
    "b200dp_comm_rank": (c_int, [_VP]),
    "b200dp_comm_world": (c_int, [_VP]),
    "b200dp_comm_heap": (_VP, [_VP, c_int]),
    "b200dp_comm_heap_bytes": (c_size_t, [_VP]),
    "b200dp_comm_next_seq": (c_uint32, [_VP]),
    "b200dp_comm_barrier": (c_int, [_VP, _VP]),
    "b200dp_comm_all_gather32": (c_int, [_VP, _VP, c_int, _VP, _VP]),
    "b200dp_comm_error": (c_int, [_VP]),
    "b200dp_comm_error_async": (c_int, [_VP, _VP, _VP]),
    "b200dp_bn_one_kernel": (c_int, [c_int, c_int64, c_int, c_int]),
    "b200dp_bn_set_one_kernel_limits": (c_int, [c_int, c_int, c_int, c_int]),
    "b200dp_bucket_pack": (c_int, [POINTER(_VP), POINTER(c_int64), POINTER(c_int64), c_int, _VP,
                                   c_int, c_float, _VP]),
    "b200dp_allreduce_update": (c_int, [_VP, c_size_t, c_int, c_size_t, c_int, c_int64, c_int64,
                                        _VP, _VP, _VP, POINTER(OptArgs), c_int, c_int, c_int,
                                        c_uint32, _VP]),
}


def exported_symbols():
    """Names every build of the library must export (checked by the CPU tests)."""
    return sorted(_SIGNATURES)


def lib():
    """Load libb200dp.so once; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise B200DPError(
                    f"{LIB_PATH} is missing: build it with __graft_entry__.build() (nvcc, sm_100a). "
                    "This package has no CPU or eager fallback.")
            handle = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in _SIGNATURES.items():
                fn = getattr(handle, name)  # AttributeError if the symbol is not exported
                fn.restype = res
                fn.argtypes = args
            _lib = _Handle(handle)
    return _lib


_fast = None


def fast():
    """The fast-call shim bound to the loaded library, or None when it has not been built
    (callers then use ctypes: same C ABI, ~10 us more per call)."""
    global _fast
    if _fast is None:
        path = _fast_path()
        if not os.path.exists(path):
            _fast = False
        else:
            import importlib.util

            spec = importlib.util.spec_from_file_location("_b200dp_fast", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            cdll = lib()._cdll
            mod.bind(ctypes.cast(cdll.b200dp_bn_forward, c_void_p).value,
                     ctypes.cast(cdll.b200dp_bn_backward, c_void_p).value)
            _fast = mod
    return _fast or None


_native = None


def native():
    """The C++ autograd node module bound to the loaded library, or None when it has not been built
    (the python autograd.Function in syncbn.py then drives the same two C-ABI calls).
    B200DP_PY_AUTOGRAD=1 forces the python node."""
    global _native
    if _native is None:
        path = _autograd_path()
        if not os.path.exists(path) or os.environ.get("B200DP_PY_AUTOGRAD", "0") == "1":
            _native = False
        else:
            import importlib.util

            import torch  # noqa: F401  (the extension links against libtorch)

            lib()  # the ctypes handle first: both must refer to the same loaded library
            spec = importlib.util.spec_from_file_location("_b200dp_autograd", path)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            mod.bind(LIB_PATH)
            _native = mod
    return _native or None


def total_calls() -> dict:
    """CALLS plus the calls the C++ autograd node made (it does not go through _Entry)."""
    out = dict(CALLS)
    if _native:
        f, b = _native.counters()
        out["b200dp_bn_forward"] = out.get("b200dp_bn_forward", 0) + f - _native_base[0]
        out["b200dp_bn_backward"] = out.get("b200dp_bn_backward", 0) + b - _native_base[1]
    return out


_native_base = [0, 0]


def reset_calls() -> None:
    CALLS.clear()
    if _native:
        _native_base[0], _native_base[1] = _native.counters()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200dp_error_string(rc).decode()
        raise B200DPError(f"{what or 'libb200dp'} failed with code {rc}: {msg}")


def dtype_code(dtype) -> int:
    import torch

    if dtype == torch.float32:
        return F32
    if dtype == torch.bfloat16:
        return BF16
    raise B200DPError(f"libb200dp handles float32 and bfloat16 tensors, not {dtype}")


def ptr(t) -> int:
    """Device (or pinned host) address of a tensor, 0 for None."""
    return 0 if t is None else t.data_ptr()


def current_stream() -> int:
    import torch

    return torch.cuda.current_stream().cuda_stream


def raw_stream(device_index: int) -> int:
    """cudaStream_t of torch's current stream on that device, without building a Stream object
    (torch.cuda.current_stream() costs ~10 us per call; this is the call TorchInductor uses)."""
    import torch

    return torch._C._cuda_getCurrentRawStream(device_index)


# launches of bn_forward / bn_backward are two kernels each: bench.py counts kernels, not calls
LAUNCHES_PER_CALL = {"b200dp_bn_forward": 2, "b200dp_bn_backward": 2}
