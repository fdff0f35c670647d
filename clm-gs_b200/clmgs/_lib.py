"""ctypes binding of libclmgs.so, generated from include/clmgs.h.

The header is the single source of truth: every ``clmgs_*`` prototype in it is
parsed into ctypes argtypes, so python and C cannot drift apart.  There is NO
fallback: if the library is missing or a symbol cannot be resolved, importing
or calling fails loudly (the product never routes through the CPU oracle or
plain torch ops).
"""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_PKG = os.path.dirname(_HERE)
_ROOT = os.path.dirname(_PKG)
HEADER = os.path.join(_ROOT, "include", "clmgs.h")
LIB_PATH = os.environ.get("CLMGS_LIB_PATH") or os.path.join(_PKG, "libclmgs.so")   # override: tuning variants only

_SCALARS = {"int": C.c_int, "int64_t": C.c_int64, "float": C.c_float, "size_t": C.c_size_t,
            "int32_t": C.c_int32}
_NO_STATUS = ("clmgs_abi_version", "clmgs_isect_count_static_max")  # int-like returns that are values


def parse_header(path=HEADER):
    """-> {name: (restype, [(ctype, argname, is_pointer)])}"""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    src = re.sub(r"//[^\n]*", " ", src)
    protos = {}
    for m in re.finditer(r"([\w\s\*]+?)\b(clmgs_\w+)\s*\(([^)]*)\)\s*;", src):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if "*" in ret:
            restype = C.c_char_p
        else:
            restype = _SCALARS[ret.split()[-1]]
        parsed = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                if "*" in a:
                    nm = a.split("*")[-1].strip()
                    parsed.append((C.c_void_p, nm, True))
                else:
                    toks = a.split()
                    parsed.append((_SCALARS[toks[-2]], toks[-1], False))
        protos[name] = (restype, parsed)
    return protos


class ClmgsError(RuntimeError):
    pass


class _Lib:
    def __init__(self):
        if not os.path.exists(LIB_PATH):
            raise ClmgsError(
                "libclmgs.so not found at %s -- build it with `python clm-gs_b200/build.py` "
                "(there is no CPU or torch fallback)" % LIB_PATH)
        self.cdll = C.CDLL(LIB_PATH)
        self.protos = parse_header()
        self.fn = {}
        self.cdll.clmgs_last_error.restype = C.c_char_p

    def missing_symbols(self):
        return [n for n in self.protos if not hasattr(self.cdll, n)]

    def _resolve(self, name):
        f = self.fn.get(name)
        if f is None:
            restype, args = self.protos[name]
            try:
                f = getattr(self.cdll, name)
            except AttributeError:
                raise ClmgsError("libclmgs.so does not export %s (declared in include/clmgs.h)" % name)
            f.restype = restype
            f.argtypes = [t for t, _, _ in args]
            self.fn[name] = f
        return f

    @staticmethod
    def _conv(v, ctype, is_ptr, fname, aname):
        if is_ptr:
            if v is None:
                return None
            if hasattr(v, "data_ptr"):  # torch.Tensor
                if not v.is_contiguous():
                    raise ClmgsError("%s: argument %s must be contiguous" % (fname, aname))
                return v.data_ptr()
            if isinstance(v, np.ndarray):
                if not v.flags["C_CONTIGUOUS"]:
                    raise ClmgsError("%s: argument %s must be contiguous" % (fname, aname))
                return v.ctypes.data
            if isinstance(v, int):
                return v
            raise ClmgsError("%s: cannot pass %r as pointer %s" % (fname, type(v), aname))
        return v

    def call(self, name, *args):
        restype, spec = self.protos[name]
        if len(args) != len(spec):
            raise ClmgsError("%s expects %d arguments, got %d" % (name, len(spec), len(args)))
        cargs = [self._conv(v, t, p, name, an) for v, (t, an, p) in zip(args, spec)]
        rc = self._resolve(name)(*cargs)
        if restype is C.c_int and name not in _NO_STATUS and rc != 0:
            raise ClmgsError("%s failed (%d): %s" % (name, rc, self.cdll.clmgs_last_error().decode()))
        return rc


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _Lib()
    return _lib


def call(name, *args):
    return lib().call(name, *args)


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream
