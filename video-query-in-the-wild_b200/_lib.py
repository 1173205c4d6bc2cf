"""ctypes binding of libvqwild.so (include/vqwild.h).  No fallback: if the CUDA extension is
missing or the device is not sm_100 every compute call raises."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
SO_PATH = os.path.join(_HERE, "libvqwild.so")

VQ_F32, VQ_BF16 = 0, 1
PREC_AUTO, PREC_BF16, PREC_SPLIT3, PREC_EXACT = 0, 1, 2, 3

_lib = None


class VqError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("vqwild error %d: %s" % (code, msg))
        self.code = code


def build(extra_flags=()):
    """Compile the CUDA extension in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import subprocess
    subprocess.run(["bash", os.path.join(_HERE, "csrc", "build.sh"), *extra_flags], check=True)
    return SO_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                "libvqwild.so is not built (%s). Run `python -c 'import __graft_entry__ as g; g.build()'`; "
                "there is no CPU fallback." % SO_PATH)
        L = C.CDLL(SO_PATH)
        L.vq_last_error.restype = C.c_char_p
        L.vq_index_ntotal.restype = C.c_int64
        L.vq_index_ntotal.argtypes = [C.c_void_p]
        L.vq_index_dim.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def debug_set(hook, value):
    """Test / A-B hooks (FORCE_FALLBACK, NO_PAIR, NO_CHAIN, DENSE_ROWS, NO_TC_TRIPLET, TRIPLET_CLOCKS, NO_TC_EVAL, NO_INSERT,
    KPRIME, TRIGGER, NO_VMAJOR, NO_NMS_MASKS, NO_SEG_UNIQUE, WALKERS, NO_L2SQ_WIDE); see include/vqwild.h."""
    check(lib().vq_debug_set(hook.encode(), C.c_int(int(value))))


def check(rc):
    if rc != 0:
        raise VqError(rc, lib().vq_last_error().decode())


def _is_torch(x):
    return type(x).__module__.startswith("torch")


def ptr(x):
    """c_void_p of a numpy array (host) or torch tensor (host or CUDA).  None -> NULL."""
    if x is None:
        return C.c_void_p(0)
    if _is_torch(x):
        assert x.is_contiguous(), "tensor must be contiguous"
        return C.c_void_p(x.data_ptr())
    assert x.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return C.c_void_p(x.ctypes.data)


def stream_of(*xs):
    for x in xs:
        if x is not None and _is_torch(x) and x.is_cuda:
            import torch
            return C.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
    return C.c_void_p(0)


def emb_dtype(x):
    """VQ_F32 / VQ_BF16 code of an embedding array; converts other numpy dtypes to float32."""
    if _is_torch(x):
        import torch
        if x.dtype == torch.bfloat16:
            return x.contiguous(), VQ_BF16
        return x.float().contiguous(), VQ_F32
    return np.ascontiguousarray(x, dtype=np.float32), VQ_F32


def as_i32(x):
    if x is None:
        return None
    if _is_torch(x):
        import torch
        return x.to(torch.int32).contiguous()
    return np.ascontiguousarray(x, dtype=np.int32)


def as_f32(x):
    if x is None:
        return None
    if _is_torch(x):
        import torch
        return x.to(torch.float32).contiguous()
    return np.ascontiguousarray(x, dtype=np.float32)


def empty_like_kind(ref, shape, dtype):
    """Output buffer living where `ref` lives (torch CUDA tensor -> CUDA tensor, else numpy)."""
    if _is_torch(ref):
        import torch
        td = {np.float32: torch.float32, np.int64: torch.int64, np.float64: torch.float64,
              np.int32: torch.int32, np.uint8: torch.uint8}[dtype]
        return torch.empty(shape, dtype=td, device=ref.device)
    return np.empty(shape, dtype=dtype)
