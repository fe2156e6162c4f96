"""ctypes binding of libfcaf3d_b200.so.

The prototypes are parsed from include/fcaf3d_b200.h, so the header is the single source of truth for the
C ABI.  There is no CPU fallback: if the library is missing or a call fails, an exception is raised.
"""
import ctypes
import os
import re

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(os.path.dirname(HERE), "include", "fcaf3d_b200.h")
LIB_PATH = os.path.join(HERE, "libfcaf3d_b200.so")

_SCALARS = {
    "int": ctypes.c_int,
    "long long": ctypes.c_longlong,
    "float": ctypes.c_float,
    "size_t": ctypes.c_size_t,
    "unsigned int": ctypes.c_uint,
    "double": ctypes.c_double,
}


class B2Error(RuntimeError):
    pass


def parse_header(path=HEADER):
    """Returns {name: (restype, [argtypes], [argnames])} for every prototype of the header."""
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    text = re.sub(r"//[^\n]*", " ", text)
    text = re.sub(r"#[^\n]*", " ", text)
    protos = {}
    for m in re.finditer(r"([A-Za-z_][\w\s\*]*?)\s*\b(b2_\w+)\s*\(([^;{}]*)\)\s*;", text):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        ret = ret.replace('extern "C"', "").strip()
        if "*" in ret:
            restype = ctypes.c_char_p if "char" in ret else ctypes.c_void_p
        else:
            restype = _SCALARS[ret]
        argtypes, argnames = [], []
        if args and args != "void":
            for a in args.split(","):
                a = " ".join(a.split())
                if "*" in a:
                    argtypes.append(ctypes.c_void_p)
                    argnames.append(a.split("*")[-1].strip())
                else:
                    toks = a.replace("const ", "").split(" ")
                    argnames.append(toks[-1])
                    argtypes.append(_SCALARS[" ".join(toks[:-1])])
        protos[name] = (restype, argtypes, argnames)
    return protos


_lib = None
_protos = None


def lib():
    global _lib, _protos
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B2Error(
                "libfcaf3d_b200.so is missing (%s). Build it with `python -m fcaf3d_b200.build`; "
                "there is no CPU or PyTorch fallback for this path." % LIB_PATH)
        l = ctypes.CDLL(LIB_PATH)
        _protos = parse_header()
        for name, (restype, argtypes, _) in _protos.items():
            fn = getattr(l, name)  # AttributeError if the library lacks a declared symbol
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = l
    return _lib


def last_error():
    return lib().b2_last_error().decode()


def check(rc, what=""):
    if rc != 0:
        raise B2Error("%s failed (%d): %s" % (what, rc, last_error()))


def ptr(t):
    """Device (or host) pointer of a tensor, None -> NULL."""
    if t is None:
        return None
    return t.data_ptr()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def stream():
    """cudaStream_t of torch's current stream (the raw-handle query is ~10x cheaper than building a Stream object, and
    every entry point asks for it: ~200 calls per inference step)."""
    if _raw_stream is not None:
        return _raw_stream(torch.cuda.current_device())
    return torch.cuda.current_stream().cuda_stream


_fns = {}


def call(name, *args):
    """Calls an int-returning entry point on the current torch stream (appended as last argument)."""
    fn = _fns.get(name)
    if fn is None:
        fn = _fns[name] = getattr(lib(), name)
    rc = fn(*args, stream())
    if rc != 0:
        raise B2Error("%s failed (%d): %s" % (name, rc, last_error()))


def require_cuda(t, name="tensor"):
    if not t.is_cuda:
        raise B2Error("%s must be a CUDA tensor: fcaf3d_b200 has no CPU path" % name)
    if not t.is_contiguous():
        raise B2Error("%s must be contiguous" % name)


DT = {torch.float32: 0, torch.bfloat16: 1}

_ws = {}


def workspace(nbytes, device, slot=0):
    """Cached, grow-only scratch buffer per (device, slot, current stream): kernels enqueued on different streams (the
    trainer builds the next batch's maps on a side stream) must not share scratch."""
    key = (device.index if device.index is not None else torch.cuda.current_device(), slot, stream())
    buf = _ws.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes), 1 << 16), dtype=torch.uint8, device=device)
        _ws[key] = buf
    return buf
