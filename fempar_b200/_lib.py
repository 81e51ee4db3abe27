"""ctypes binding of libfempar_b200.so (the C-ABI declared in include/fempar_b200.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is visible,
every entry point raises.
"""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libfempar_b200.so")


class ILSParams(C.Structure):
    _fields_ = [("rtol", C.c_double), ("atol", C.c_double), ("max_num_iterations", C.c_int),
                ("stopping_criterium", C.c_int), ("track_convergence_history", C.c_int), ("check_every", C.c_int)]


class ILSExtParams(C.Structure):
    _fields_ = [("dkrymax", C.c_int), ("orthonorm_strat", C.c_int), ("relaxation", C.c_double)]


class Fb2Error(RuntimeError):
    pass


_lib = None
_inited = False

_VP = C.c_void_p
_SIGS = {
    "fb2_init": [C.c_int],
    "fb2_device_synchronize": [],
    "fb2_last_phase_ms": [C.c_char_p, C.POINTER(C.c_double)],
    "fb2_comm_unique_id": [_VP],
    "fb2_comm_init": [C.c_int, C.c_int, _VP],
    "fb2_comm_finalize": [],
    "fb2_fe_space_create": [C.POINTER(_VP), _VP, _VP, C.c_int, C.c_int, C.c_int, C.c_uint32],
    "fb2_fe_space_destroy": [_VP],
    "fb2_fe_space_get_sizes": [_VP, _VP],
    "fb2_fe_space_get_lst_dofs_gids": [_VP, _VP],
    "fb2_fe_space_get_cells_gids": [_VP, _VP, _VP],
    "fb2_fe_space_interpolate_dirichlet": [_VP, C.c_int, _VP],
    "fb2_fe_space_interpolate": [_VP, C.c_int, _VP, _VP],
    "fb2_fe_space_get_fixed_values": [_VP, _VP],
    "fb2_fe_space_set_fixed_values": [_VP, _VP],
    "fb2_dof_import_get_sizes": [_VP, _VP],
    "fb2_dof_import_get": [_VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP],
    "fb2_csr_create": [C.POINTER(_VP), _VP, C.c_int],
    "fb2_csr_destroy": [_VP],
    "fb2_fe_space_get_quadrature_points": [_VP, _VP],
    "fb2_fe_space_set_source_values": [_VP, _VP],
    "fb2_fe_space_create_bare": [C.POINTER(_VP), C.c_int64],
    "fb2_fe_space_create_from_host": [C.POINTER(_VP), C.c_int32, C.c_int32, C.c_int, C.c_int, _VP, _VP, C.c_int64, C.c_int64],
    "fb2_dof_import_set": [_VP, C.c_int32, _VP, _VP, _VP, C.c_int32, _VP, _VP, _VP, _VP],
    "fb2_csr_create_from_host": [C.POINTER(_VP), _VP, C.c_int64, _VP, _VP, _VP, C.c_int],
    "fb2_csr_initialize_values": [_VP, C.c_double],
    "fb2_assemble_batch": [_VP, _VP, C.c_int32, C.c_int32, _VP, _VP, _VP, _VP, C.c_int64],
    "fb2_csr_get_sizes": [_VP, C.POINTER(C.c_int64), C.POINTER(C.c_int64)],
    "fb2_csr_get_pattern": [_VP, _VP, _VP],
    "fb2_csr_get_values": [_VP, _VP],
    "fb2_csr_set_values": [_VP, _VP],
    "fb2_integrate_assemble": [_VP, _VP, C.c_int, C.c_int, _VP, _VP, C.c_int],
    "fb2_integrate_residual": [_VP, _VP, C.c_int, _VP, C.c_double, _VP],
    "fb2_integrate_tangent": [_VP, _VP, C.c_int, _VP, C.c_double],
    "fb2_spmv": [_VP, _VP, _VP],
    "fb2_extract_diagonal": [_VP, _VP],
    "fb2_set_spmv_variant": [C.c_int],
    "fb2_csr_set_initial_solution": [_VP, _VP],
    "fb2_csr_free_solver_workspace": [_VP],
    "fb2_csr_spmv_reset": [_VP],
    "fb2_peer_allreduce_selftest": [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double)],
    "fb2_csr_get_convergence_data": [_VP, _VP, _VP],
    "fb2_set_assembly_variant": [C.c_int],
    "fb2_vec_create": [C.POINTER(_VP), _VP],
    "fb2_vec_create_n": [C.POINTER(_VP), C.c_int64],
    "fb2_vec_destroy": [_VP],
    "fb2_vec_size": [_VP, C.POINTER(C.c_int64)],
    "fb2_vec_set": [_VP, _VP],
    "fb2_vec_get": [_VP, _VP],
    "fb2_vec_init": [_VP, C.c_double],
    "fb2_vec_copy": [_VP, _VP],
    "fb2_vec_scal": [_VP, C.c_double, _VP],
    "fb2_vec_axpby": [_VP, C.c_double, _VP, C.c_double],
    "fb2_vec_dot": [_VP, _VP, C.POINTER(C.c_double)],
    "fb2_vec_nrm2": [_VP, C.POINTER(C.c_double)],
    "fb2_vec_local_dot": [_VP, _VP, C.POINTER(C.c_double)],
    "fb2_vec_dot_no_alloc": [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(C.c_double)],
    "fb2_vec_nullify_non_owned_dofs": [_VP],
    "fb2_vec_extract_subvector": [_VP, C.c_int32, _VP, _VP],
    "fb2_vec_insert_subvector": [_VP, C.c_int32, _VP, _VP],
    "fb2_vec_entrywise_product": [_VP, _VP, _VP],
    "fb2_vec_entrywise_invert": [_VP],
    "fb2_vec_comm": [_VP],
    "fb2_jacobi_setup": [_VP, _VP],
    "fb2_pcg_solve": [_VP, _VP, _VP, _VP, C.POINTER(ILSParams), C.POINTER(C.c_int), C.POINTER(C.c_int),
                      C.POINTER(C.c_double), _VP],
    "fb2_pcg_solve_host": [_VP, _VP, _VP, _VP, C.POINTER(ILSParams), C.POINTER(C.c_int), C.POINTER(C.c_int),
                           C.POINTER(C.c_double), _VP],
    "fb2_ils_solve": [_VP, _VP, _VP, _VP, C.c_int, C.POINTER(ILSParams), C.POINTER(ILSExtParams), C.POINTER(C.c_int),
                      C.POINTER(C.c_int), C.POINTER(C.c_double), _VP],
    "fb2_comm_set_parts_per_rank": [C.c_int],
    "fb2_vec_comm_multi": [C.c_int, _VP],
    "fb2_vec_dot_multi": [C.c_int, _VP, _VP, C.POINTER(C.c_double)],
    "fb2_spmv_multi": [C.c_int, _VP, _VP, _VP],
    "fb2_jacobi_setup_multi": [C.c_int, _VP, _VP],
    "fb2_pcg_solve_multi": [C.c_int, _VP, _VP, _VP, _VP, C.POINTER(ILSParams), C.POINTER(C.c_int), C.POINTER(C.c_int),
                            C.POINTER(C.c_double), _VP],
    "fb2_pcg_iterate_multi": [C.c_int, _VP, _VP, _VP, _VP, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)],
    "fb2_pcg_iterate": [_VP, _VP, _VP, _VP, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)],
    "fb2_error_norms": [_VP, _VP, C.c_int, _VP, C.POINTER(C.c_double), C.POINTER(C.c_double)],
    "fb2_error_norms_all": [_VP, _VP, C.c_int, _VP, C.c_double, _VP],
}
EXPORTED_SYMBOLS = sorted(list(_SIGS) + ["fb2_last_error_string", "fb2_version", "fb2_vec_device_ptr"])


def load():
    """dlopen the library and declare the signatures (no CUDA call is made)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Fb2Error("%s is missing: run `python -m fempar_b200.build` (there is no CPU fallback)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        for name, args in _SIGS.items():
            f = getattr(L, name)
            f.argtypes = args
            f.restype = C.c_int
        L.fb2_last_error_string.restype = C.c_char_p
        L.fb2_version.restype = C.c_int
        L.fb2_vec_device_ptr.restype = C.c_void_p
        L.fb2_vec_device_ptr.argtypes = [_VP]
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise Fb2Error(load().fb2_last_error_string().decode())


def init(device=None):
    """FEMPAR_INIT analogue: selects the CUDA device of this rank."""
    global _inited
    L = load()
    if device is None:
        device = int(os.environ.get("LOCAL_RANK", "0"))
    check(L.fb2_init(int(device)))
    _inited = True
    return L


def lib():
    if not _inited:
        init()
    return _lib
