"""ctypes binding of libunetb200.so.

The prototypes are read from include/unetb200.h, so the header is the single source of truth for the C ABI.
There is no fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes
import os
import re

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_REPO_DIR = os.path.dirname(_PKG_DIR)
LIB_PATH = os.path.join(_PKG_DIR, "libunetb200.so")
HEADER_PATH = os.path.join(_REPO_DIR, "include", "unetb200.h")


class UnetB200Error(RuntimeError):
    pass


# element-type enum of include/unetb200.h
BF16, TF32, F32, F64 = 0, 1, 2, 3


_SCALARS = {
    "int": ctypes.c_int,
    "float": ctypes.c_float,
    "long long": ctypes.c_longlong,
    "size_t": ctypes.c_size_t,
    "uint64_t": ctypes.c_uint64,
    "uint32_t": ctypes.c_uint32,
    "void": None,
}


def _ctype(decl: str, is_return=False):
    decl = decl.strip()
    if "*" in decl:
        if is_return and decl.replace(" ", "") == "constchar*":
            return ctypes.c_char_p
        return ctypes.c_void_p
    base = decl.replace("const", "").strip()
    return _SCALARS[base]


def parse_header(path=HEADER_PATH):
    """-> {name: (restype, [argtypes])} for every unetb200_* function the header declares."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"^\s*#.*$", " ", text, flags=re.M)
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(unetb200_\w+)\s*\(([^;{}]*?)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        if ret.startswith("typedef"):
            continue
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                # drop the parameter name (last identifier) unless the declaration is a bare type
                mm = re.match(r"^(.*?)(\b[A-Za-z_]\w*)$", a)
                typ = mm.group(1).strip() if mm and mm.group(1).strip() else a
                argtypes.append(_ctype(typ))
        protos[name] = (_ctype(ret, is_return=True), argtypes)
    return protos


_lib = None


def lib():
    """Load (once) and return the bound library; raises UnetB200Error when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise UnetB200Error(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C unet_distributed_b200/csrc`). There is no CPU fallback.")
    cdll = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in parse_header().items():
        try:
            fn = getattr(cdll, name)
        except AttributeError as e:
            raise UnetB200Error(f"{LIB_PATH} does not export {name} declared in {HEADER_PATH}") from e
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = cdll
    return _lib


def last_error() -> str:
    s = lib().unetb200_last_error_string()
    return s.decode() if s else ""


def check(rc: int, what: str = "libunetb200 call"):
    if rc != 0:
        raise UnetB200Error(f"{what} failed with code {rc}: {last_error()}")


def call(name: str, *args):
    """Call an int-returning entry point and raise on a non-zero status."""
    check(getattr(lib(), name)(*args), name)
