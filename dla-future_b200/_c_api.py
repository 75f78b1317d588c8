"""ctypes signatures of the reference-facing C API (include/dlaf_c/*.h) and of the MPI shim."""
import ctypes

c_int = ctypes.c_int
c_char = ctypes.c_char
c_vp = ctypes.c_void_p
P_int = ctypes.POINTER(ctypes.c_int)


class DLAFDescriptor(ctypes.Structure):
    """struct DLAF_descriptor (include/dlaf_c/desc.h)."""
    _fields_ = [(k, c_int) for k in ("m", "n", "mb", "nb", "isrc", "jsrc", "i", "j", "ld")]


DESC9 = c_int * 9

C_API = {
    "dlaf_initialize": (None, [c_int, c_vp, c_int, c_vp]),
    "dlaf_finalize": (None, []),
    "dlaf_create_grid": (c_int, [c_vp, c_int, c_int, c_char]),
    "dlaf_free_grid": (None, [c_int]),
    "grid_ordering": (c_char, [c_vp, c_int, c_int, c_int, c_int]),
    "dlaf_create_grid_from_blacs_info": (None, [c_int, c_vp, c_int, c_int, c_char]),
    "make_dlaf_descriptor": (DLAFDescriptor, [c_int, c_int, c_int, c_int, DESC9]),
    "dlafb_mpi_comm_world": (c_vp, []),
    "dlafb_mpi_init": (c_int, [c_int, c_int, c_vp]),
    "dlafb_mpi_finalize": (c_int, []),
    "dlafb_grid_create_from_comm": (c_int, [ctypes.POINTER(c_vp), c_vp, c_int, c_int, c_char]),
    "MPI_Comm_rank": (c_int, [c_vp, P_int]),
    "MPI_Comm_size": (c_int, [c_vp, P_int]),
}
for _x in "sdcz":
    C_API[f"dlaf_cholesky_factorization_{_x}"] = (c_int, [c_int, c_char, c_vp, DLAFDescriptor])
    C_API[f"dlaf_p{_x}potrf"] = (None, [c_char, c_int, c_vp, c_int, c_int, DESC9, P_int])


def symbols():
    return list(C_API)


def bind(L):
    for k, (res, args) in C_API.items():
        f = getattr(L, k)
        f.restype, f.argtypes = res, args
