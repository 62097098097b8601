"""ctypes binding of include/rbx.h (the C ABI of librbx.so).

Prototypes are read from the header itself so that the Python side can never drift from it.
There is no fallback: if librbx.so is missing or cannot be loaded, importing this module raises.
"""
import ctypes
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(_HERE), "include", "rbx.h")
# RBX_LIBRARY: another build of the same library (kernel experiments); the default is the in-tree one
LIBPATH = os.environ.get("RBX_LIBRARY") or os.path.join(_HERE, "librbx.so")

_CT = {
    "void": None, "char": ctypes.c_char, "int": ctypes.c_int, "int64_t": ctypes.c_int64, "uint64_t": ctypes.c_uint64,
    "float": ctypes.c_float, "double": ctypes.c_double, "uint8_t": ctypes.c_uint8,
}


def _ctype(decl):
    decl = decl.strip()
    is_ptr = "*" in decl or "[" in decl
    base = decl.replace("const", " ").replace("*", " ")
    base = re.sub(r"\[\d*\]", " ", base).split()
    tname = base[0]
    if is_ptr:
        return ctypes.c_char_p if tname == "char" else ctypes.c_void_p
    return _CT[tname]


def parse_header(path=HEADER):
    """-> {name: (restype, [argtypes], [argnames])} for every function declared in rbx.h"""
    src = open(path).read()
    src = re.sub(r"/\*.*?\*/", " ", src, flags=re.S)
    protos = {}
    for m in re.finditer(r"(?:^|\n)\s*((?:const\s+)?\w+\s*\*?)\s*(rbx_\w+)\s*\(([^;{]*?)\)\s*;", src, flags=re.S):
        ret, name, args = m.group(1), m.group(2), m.group(3)
        args = " ".join(args.split())
        argt, argn = [], []
        if args and args != "void":
            for a in args.split(","):
                argt.append(_ctype(a))
                argn.append(re.sub(r"\[\d*\]", "", a.strip().split()[-1].replace("*", "")))
        protos[name] = (_ctype(ret) if "*" not in ret else ctypes.c_char_p, argt, argn)
    return protos


PROTOS = parse_header()

if not os.path.exists(LIBPATH):
    raise ImportError(f"{LIBPATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                      "(randblas_b200 has no CPU fallback)")
lib = ctypes.CDLL(LIBPATH)
for _name, (_ret, _argt, _argn) in PROTOS.items():
    _fn = getattr(lib, _name)          # AttributeError here == header/library mismatch
    _fn.restype = _ret
    _fn.argtypes = _argt


class RbxError(RuntimeError):
    def __init__(self, code, where):
        self.code = code
        super().__init__(f"{where}: {lib.rbx_strerror(code).decode()} (code {code})")


def check(code, where):
    if code != 0:
        raise RbxError(code, where)
