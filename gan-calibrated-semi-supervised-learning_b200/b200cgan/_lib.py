"""ctypes binding of libb2c.so (C-ABI declared in include/b2c.h).

There is NO fallback: if the library is missing or a call fails, an exception is raised.  Signatures are read
from the header so the binding cannot drift from the ABI.
"""
from __future__ import annotations

import ctypes
import os
import re
from pathlib import Path

import torch

_PKG = Path(__file__).resolve().parent
_ROOT = _PKG.parent.parent
HEADER = _ROOT / "include" / "b2c.h"
LIBPATH = _PKG / "libb2c.so"

_CTYPES = {"int": ctypes.c_int, "long long": ctypes.c_longlong, "float": ctypes.c_float,
           "unsigned long long": ctypes.c_ulonglong}


def parse_header(path: Path = HEADER):
    """-> {name: (restype, [(ctype, is_pointer)])} for every b2c_* prototype."""
    text = re.sub(r"/\*.*?\*/", "", path.read_text(), flags=re.S)
    protos = {}
    for m in re.finditer(r"(const char\*|long long|int)\s+(b2c_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.group(1), m.group(2), m.group(3).strip()
        sig = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    sig.append((ctypes.c_void_p, True))
                else:
                    ty = re.sub(r"\s+\w+$", "", a).replace("const ", "").strip()
                    sig.append((_CTYPES[ty], False))
        protos[name] = ({"const char*": ctypes.c_char_p, "long long": ctypes.c_longlong}.get(ret, ctypes.c_int), sig)
    return protos


class B2CError(RuntimeError):
    pass


class _Native:
    def __init__(self):
        if not LIBPATH.exists():
            raise B2CError(f"{LIBPATH} not found: build it with __graft_entry__.build() "
                           f"(csrc/build.sh); there is no CPU or library fallback")
        self.lib = ctypes.CDLL(str(LIBPATH))
        self.protos = parse_header()
        self.fns = {}
        for name, (ret, sig) in self.protos.items():
            fn = getattr(self.lib, name)
            fn.restype = ret
            fn.argtypes = [t for t, _ in sig]
            self.fns[name] = fn
        self.launches = 0
        self.timed = None  # bench hook: {kernel name: [(start_event, end_event, argument pattern)]} (see time_kernels)

    def call(self, name, *args):
        fn = self.fns[name]
        conv = []
        for a in args:
            if isinstance(a, torch.Tensor):
                if not a.is_cuda:
                    raise B2CError(f"{name}: tensor argument is not on a CUDA device")
                conv.append(a.data_ptr())
            elif a is None:
                conv.append(None)
            else:
                conv.append(a)
        stream = torch.cuda.current_stream()
        rec = self.timed.get(name) if self.timed is not None else None
        if rec is not None:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
        rc = fn(*conv, stream.cuda_stream)
        if rec is not None:
            ev1.record(stream)
            # scalars as they are, every other argument as "was it given": enough to size the launch (bench.py, kprof.py)
            rec.append((ev0, ev1, tuple(a if isinstance(a, (int, float)) else (a is not None) for a in args)))
        self.launches += 1
        if rc != 0:
            raise B2CError(f"{name} failed ({rc}): {self.lib.b2c_last_error().decode()}")


_backend = None


def backend():
    global _backend
    if _backend is None:
        _backend = _Native()
    return _backend


def call(name, *args):
    backend().call(name, *args)


def query(name, *args):
    """Host-side helper entry points that return a value and launch nothing (e.g. workspace sizes)."""
    return backend().fns[name](*args)


def launch_count() -> int:
    return backend().launches


def time_kernels(names):
    """Bench hook: bracket every later launch of the named kernels with CUDA events on the launching stream.
    Returns the dict the (start, end, scalar-args) records are appended to; pass None to switch it off."""
    b = backend()
    b.timed = None if names is None else {n: [] for n in names}
    return b.timed


_sm_count = None


def sm_count() -> int:
    global _sm_count
    if _sm_count is None:
        _sm_count = torch.cuda.get_device_properties(torch.cuda.current_device()).multi_processor_count \
            if torch.cuda.is_available() else 148
    return _sm_count
