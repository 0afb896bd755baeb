"""ctypes binding of libnlosb200.so (include/nlosb200.h).  No torch types cross this boundary:
device pointers go in as integers (tensor.data_ptr()), the stream as torch's current cudaStream_t.

There is NO fallback: if the library is missing or a call fails, a RuntimeError is raised."""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "lib", "libnlosb200.so")
_HEADER = os.path.join(os.path.dirname(_HERE), "include", "nlosb200.h")
_lib = None

_CT = {"int": C.c_int, "float": C.c_float, "long long": C.c_longlong, "void*": C.c_void_p}


def _parse_header():
    """[(name, restype, [argtypes])] for every function the header declares."""
    text = open(_HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = []
    for m in re.finditer(r"(const char\*|long long|int)\s+(nlosb200_\w+)\s*\(([^)]*)\)\s*;", text):
        ret, name, args = m.groups()
        argt = []
        for a in [x.strip() for x in args.split(",")]:
            if a in ("void", ""):
                continue
            if "*" in a:
                argt.append(C.c_void_p)
            elif a.startswith("float"):
                argt.append(C.c_float)
            elif a.startswith("long long"):
                argt.append(C.c_longlong)
            else:
                argt.append(C.c_int)
        res = C.c_char_p if ret.startswith("const char") else (C.c_longlong if ret == "long long" else C.c_int)
        out.append((name, res, argt))
    return out


def declared_symbols():
    return [n for n, _, _ in _parse_header()]


def lib_path():
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            from . import build as _b
            _b.build()
        L = C.CDLL(_LIB_PATH)
        for name, res, argt in _parse_header():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = argt
        _lib = L
    return _lib


def call(name, *args):
    """Invoke an int-returning entry point; raise RuntimeError(last_error) on failure."""
    L = lib()
    rc = getattr(L, name)(*args)
    if rc != 0:
        raise RuntimeError(f"{name} failed ({rc}): {L.nlosb200_last_error().decode()}")


def ptr(t):
    """device pointer of a torch tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream():
    import torch
    return torch.cuda.current_stream().cuda_stream
