"""ctypes binding of libjtsm_b200.so (the C-ABI of include/jtsm_b200.h).

There is no CPU fallback: if the library is missing or a call returns a
non-zero status, a RuntimeError is raised.
"""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libjtsm_b200.so")
HEADER = os.path.join(HERE, "..", "include", "jtsm_b200.h")

F32, BF16, F16 = 0, 1, 2
ARGMAX_I32, ARGMAX_U16 = 0, 1

_lib = None

_vp = ctypes.c_void_p
_i = ctypes.c_int
_f = ctypes.c_float
_ll = ctypes.c_longlong
_sz = ctypes.c_size_t

_CTYPE = {
    "void*": _vp, "float*": _vp, "int*": _vp, "char*": ctypes.c_char_p, "int": _i, "float": _f,
    "long long": _ll, "size_t": _sz, "unsigned long long": ctypes.c_ulonglong,
    "int64_t*": _vp, "int32_t*": _vp, "uint8_t*": _vp, "unsigned char*": _vp, "signed char*": _vp,
}


def declared_functions(header=HEADER):
    """Parse include/jtsm_b200.h -> {name: (restype_str, [argtype_str, ...])}."""
    src = open(header).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    out = {}
    for m in re.finditer(r"JTSM_API\s+([\w\s\*]+?)\s*\b(jtsm_\w+)\s*\(([^)]*)\)\s*;", src):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip().replace("const ", "")
                ptr = "*" in a
                a = a.replace("*", " ")
                toks = a.split()
                base = " ".join(toks[:-1]) if len(toks) > 1 else toks[0]
                argtypes.append(base + ("*" if ptr else ""))
        out[name] = (ret.replace("const ", ""), argtypes)
    return out


def _ctype_of(s):
    s = s.strip()
    if s in _CTYPE:
        return _CTYPE[s]
    if s.endswith("*"):
        return _vp
    raise KeyError("no ctypes mapping for %r" % s)


def lib():
    """Load (once) and return the ctypes handle with argtypes/restypes set from the header."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "jtsm_b200: %s not found. Build it with `python -m jtsm_b200.build` "
            "(there is no CPU / eager fallback)." % LIB_PATH
        )
    h = ctypes.CDLL(LIB_PATH)
    for name, (ret, args) in declared_functions().items():
        fn = getattr(h, name)
        fn.restype = ctypes.c_char_p if ret == "char*" else _ctype_of(ret)
        fn.argtypes = [_ctype_of(a) for a in args]
    if h.jtsm_abi_version() != 1:
        raise RuntimeError("jtsm_b200: ABI version mismatch")
    _lib = h
    return h


def check(status, what):
    if status != 0:
        msg = lib().jtsm_error_string(status)
        raise RuntimeError("jtsm_b200.%s failed: %s (status %d)" % (what, msg.decode() if msg else "?", status))


def launch_count():
    return int(lib().jtsm_launch_count())


def set_tuning(key, value):
    check(lib().jtsm_set_tuning(key.encode(), int(value)), "set_tuning(%s)" % key)
