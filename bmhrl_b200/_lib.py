"""ctypes binding of libbmhrl_b200.so.

The prototypes are parsed from include/bmhrl_b200.h, so the header is the single source of truth for the C ABI.
No torch types cross the boundary: tensors are passed as raw device pointers (`Tensor.data_ptr()`), the stream as the
raw `cudaStream_t` of torch's current stream.  There is no fallback: a missing library or a failing call raises.
"""
import ctypes
import os
import re

_PKG = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG)
HEADER = os.path.join(_ROOT, "include", "bmhrl_b200.h")
LIB_PATH = os.path.join(_PKG, "lib", "libbmhrl_b200.so")

_CT = {"int": ctypes.c_int, "unsigned int": ctypes.c_uint, "long long": ctypes.c_longlong, "float": ctypes.c_float,
       "void": None, "double": ctypes.c_double}


def _ctype(decl):
    decl = decl.strip()
    if "*" in decl:
        return ctypes.c_char_p if decl.replace(" ", "") == "constchar*" else ctypes.c_void_p
    decl = re.sub(r"\bconst\b", "", decl).strip()
    return _CT[decl]


def parse_header(path=HEADER):
    """Returns {name: (restype, [argtypes])} for every prototype in the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = "\n".join(l for l in text.splitlines() if not l.lstrip().startswith("#") and 'extern "C"' not in l)
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\b(bmhrl_\w+)\s*\(([^;{}]*?)\)\s*;", text):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        argtypes = []
        if args.strip() not in ("", "void"):
            for a in args.split(","):
                a = a.strip()
                # drop the parameter name (last identifier) keeping pointer stars with the type
                mm = re.match(r"(.*?)(\w+)$", a)
                argtypes.append(_ctype(mm.group(1)))
        protos[name] = (_ctype(ret), argtypes)
    return protos


class _Lib:
    def __init__(self):
        self._dll = None
        self.protos = None

    def load(self):
        if self._dll is not None:
            return self._dll
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"bmhrl_b200: CUDA library not built ({LIB_PATH} missing). Run `python -c 'import __graft_entry__ as g; "
                f"g.build()'` (or `make -C bmhrl_b200/csrc`). There is no CPU fallback.")
        dll = ctypes.CDLL(LIB_PATH)
        self.protos = parse_header()
        for name, (res, args) in self.protos.items():
            fn = getattr(dll, name)  # AttributeError here = header/library mismatch
            fn.restype = res
            fn.argtypes = args
        if dll.bmhrl_abi_version() != 1:
            raise RuntimeError("bmhrl_b200: ABI version mismatch between header and library")
        self._dll = dll
        return dll

    def call(self, name, *args):
        dll = self.load()
        rc = getattr(dll, name)(*args)
        if rc != 0:
            raise RuntimeError(f"{name} failed ({rc}): {dll.bmhrl_last_error().decode()}")

    def raw(self, name):
        return getattr(self.load(), name)


LIB = _Lib()


def launch_count():
    return int(LIB.raw("bmhrl_launch_count")())


def reset_launch_count():
    LIB.raw("bmhrl_reset_launch_count")()
