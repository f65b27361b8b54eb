"""ctypes binding of libssb200.so (the C ABI declared in include/ssb200.h).

The product path has no CPU fallback: if the CUDA library is missing this raises.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libssb200.so")

_lib = None


class SSBError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SSBError("libssb200.so not built (%s); run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "or `make -C speech-separation_b200/csrc` -- there is no CPU fallback" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.ss_last_error.restype = ctypes.c_char_p
    return _lib


def _conv(a):
    import torch
    if a is None:
        return ctypes.c_void_p(0)
    if isinstance(a, torch.Tensor):
        return ctypes.c_void_p(a.data_ptr())
    if isinstance(a, float):
        return ctypes.c_float(a)
    if isinstance(a, bool):
        return ctypes.c_int(int(a))
    if isinstance(a, int):
        return ctypes.c_longlong(a) if abs(a) >= 2 ** 31 else ctypes.c_int(a)
    return a


class LL(int):
    """Marks an integer argument that the C signature declares as `long long`."""


def call(name, *args):
    """Call ss_<name>; tensors -> device pointers, LL -> long long, int -> int, float -> float."""
    fn = getattr(lib(), "ss_" + name)
    cargs = []
    for a in args:
        if isinstance(a, LL):
            cargs.append(ctypes.c_longlong(int(a)))
        else:
            cargs.append(_conv(a))
    rc = fn(*cargs)
    if rc < 0:
        raise SSBError("ss_%s failed (%d): %s" % (name, rc, lib().ss_last_error().decode()))
    return rc


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
