"""ctypes binding of libmmvae_b200.so (the C ABI declared in include/mmvae_b200.h).

There is NO fallback: if the shared library is missing or a symbol the header declares is not exported, importing
the product path raises.  argtypes/restype are derived from the header so the binding cannot drift from the ABI.
"""
import ctypes
import os
import re

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "csrc", "libmmvae_b200.so")
HEADER_PATH = os.path.join(os.path.dirname(HERE), "include", "mmvae_b200.h")

_CTYPES = {
    "int": ctypes.c_int, "long": ctypes.c_long, "float": ctypes.c_float,
    "const float*": ctypes.c_void_p, "float*": ctypes.c_void_p, "void*": ctypes.c_void_p,
    "const int*": ctypes.c_void_p, "const void*": ctypes.c_void_p,
    "const char*": ctypes.c_char_p, "void": None,
}
_PROTO = re.compile(r"^\s*(int|long|const char\*)\s+(mmvae_\w+)\s*\(([^;{]*?)\)\s*;", re.M | re.S)


def parse_header(path=HEADER_PATH):
    """-> {name: (restype, [argtypes])} and the #define constants of the header."""
    src = open(path).read()
    src_nc = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    protos = {}
    for ret, name, args in _PROTO.findall(src_nc):
        args = " ".join(args.split())
        argtypes = []
        if args and args != "void":
            for a in args.split(","):
                a = a.strip()
                ty = a.rsplit(" ", 1)[0].strip() if not a.endswith("*") else a
                ty = ty.replace(" *", "*")
                if ty not in _CTYPES:
                    raise RuntimeError(f"mmvae_b200.h: unknown type {ty!r} in {name}")
                argtypes.append(_CTYPES[ty])
        protos[name] = (_CTYPES[ret], argtypes)
    consts = {k: int(v) for k, v in re.findall(r"^#define\s+(MMVAE_\w+)\s+(\d+)\s", src_nc, flags=re.M)}
    return protos, consts


PROTOS, CONSTS = parse_header()
_lib = None


def load():
    """Loads the library once; raises (never falls back) if it is absent or incomplete."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m mmvae_b200.build` (or __graft_entry__.build()). "
            "mmvae_b200 has no CPU or eager-PyTorch fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (restype, argtypes) in PROTOS.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise RuntimeError(f"{LIB_PATH} does not export {name} declared in mmvae_b200.h") from e
        fn.restype = restype
        fn.argtypes = argtypes
    if lib.mmvae_abi_version() != CONSTS["MMVAE_ABI_VERSION"]:
        raise RuntimeError("libmmvae_b200.so ABI version does not match include/mmvae_b200.h; rebuild")
    _lib = lib
    return lib


def check(rc, what=""):
    if rc != 0:
        msg = load().mmvae_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"libmmvae_b200 {what} failed (code {rc}): {msg}")
