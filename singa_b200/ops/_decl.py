"""argtypes for the non-GEMM entry points (kept next to include/singa_b200.h, one block per section)."""
import ctypes as C

vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_longlong, C.c_float


def declare_all(lib):
    def d(name, *args, res=C.c_int):
        if hasattr(lib, name):
            fn = getattr(lib, name)
            fn.argtypes = list(args)
            fn.restype = res
