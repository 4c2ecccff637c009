"""ctypes binding of liblsc_b200.so (include/lsc_b200.h).

The shared library is built in-tree by `make -C lsc_internship_b200/csrc` (or
`__graft_entry__.build()`).  There is deliberately NO fallback: if the library is
missing, or no CUDA device is present when a solver is created, the caller gets
an exception -- never a CPU path.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# LSC_B200_LIB selects an alternative build of the SAME library (kernel tuning experiments)
LIB_PATH = os.environ.get("LSC_B200_LIB") or os.path.join(_HERE, "_lib", "liblsc_b200.so")

LSC_OK = 0
BC_PERIODIC, BC_REFLECTIVE, BC_TRANSMISSIVE, BC_TRANSMISSIVE_E = 0, 1, 2, 3
MHD_IC_KELVIN_HELMHOLTZ, MHD_IC_ORSZAG_TANG = 0, 1

BC_NAMES = {"periodic": BC_PERIODIC, "reflective": BC_REFLECTIVE, "transmissive": BC_TRANSMISSIVE,
            "transmissive_e": BC_TRANSMISSIVE_E}
MHD_PROBLEMS = {"kh": MHD_IC_KELVIN_HELMHOLTZ, "kelvin-helmholtz": MHD_IC_KELVIN_HELMHOLTZ,
                "ot": MHD_IC_ORSZAG_TANG, "orszag-tang": MHD_IC_ORSZAG_TANG}


class LscError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"liblsc_b200 error {code}: {msg}")
        self.code = code


class MhdConfig(ctypes.Structure):
    """struct lsc_mhd_config -- the constants of MUSCL_NEW.cpp:659-671."""
    _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int),
                ("x0", ctypes.c_double), ("x1", ctypes.c_double),
                ("y0", ctypes.c_double), ("y1", ctypes.c_double),
                ("gamma", ctypes.c_double), ("cfl", ctypes.c_double),
                ("bc_x", ctypes.c_int), ("bc_y", ctypes.c_int),
                ("device", ctypes.c_int), ("rank", ctypes.c_int), ("nranks", ctypes.c_int),
                ("slab_i0", ctypes.c_int), ("slab_nx", ctypes.c_int)]


class Hllc2dConfig(ctypes.Structure):
    """struct lsc_hllc2d_config -- the constants of HLLC_MHD_2D.cpp:616-627."""
    _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int),
                ("x0", ctypes.c_double), ("x1", ctypes.c_double),
                ("y0", ctypes.c_double), ("y1", ctypes.c_double),
                ("gamma", ctypes.c_double), ("cfl", ctypes.c_double), ("device", ctypes.c_int)]


class EulerConfig(ctypes.Structure):
    """struct lsc_euler_config -- the constants of SLIC_euler_2d.cpp:529-540."""
    _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int),
                ("x0", ctypes.c_double), ("x1", ctypes.c_double),
                ("y0", ctypes.c_double), ("y1", ctypes.c_double),
                ("gamma", ctypes.c_double), ("cfl", ctypes.c_double), ("omega", ctypes.c_double),
                ("bc", ctypes.c_int), ("device", ctypes.c_int),
                ("rank", ctypes.c_int), ("nranks", ctypes.c_int)]


EULER_PROBLEMS = {"ot": 0, "orszag-tang": 0, "quadrants": 1, "cylinder": 2, "tabeos": 3}
EULER_LOG_CB = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_long, ctypes.c_double, ctypes.c_double, ctypes.c_void_p)

DUMP_CB = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_double, ctypes.c_double,
                           ctypes.c_void_p)
LOG_CB = ctypes.CFUNCTYPE(ctypes.c_int, ctypes.c_long, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_void_p)

# every symbol include/lsc_b200.h declares
EXPORTS = [
    "lsc_last_error", "lsc_device_count", "lsc_host_alloc", "lsc_host_free",
    "lsc_mhd_create", "lsc_mhd_destroy", "lsc_mhd_local_nx", "lsc_mhd_local_i0",
    "lsc_mhd_initial_condition", "lsc_mhd_upload", "lsc_mhd_init_problem", "lsc_mhd_step",
    "lsc_mhd_run_until", "lsc_mhd_get_scalars", "lsc_mhd_download", "lsc_mhd_write_dat",
    "lsc_comm_unique_id", "lsc_mhd_comm_init", "lsc_mhd_kat_hllc", "lsc_mhd_kat_fluxdef", "lsc_mhd_kat_cfl_bound", "lsc_mhd_set_arithmetic",
    "lsc_kernel_launch_count", "lsc_mhd_step_timed", "lsc_mhd_set_profiling", "lsc_mhd_kernel_times", "lsc_mhd_comm_times", "lsc_mhd_set_overlap", "lsc_kat_division", "lsc_mhd_ieee_fallback_count", "lsc_mhd_group_step", "lsc_mhd_fill_ghost_ring", "lsc_mhd_state_hash", "lsc_mhd_state_hash_noalloc", "lsc_euler_state_hash",
    "lsc_mhd_group_step_timed", "lsc_mhd_group_run_until", "lsc_mhd_loop_stats", "lsc_mhd_group_download",
    "lsc_guard_check", "lsc_euler_group_run_until",
    "lsc_hllc2d_create", "lsc_hllc2d_destroy", "lsc_hllc2d_initial_condition", "lsc_hllc2d_upload",
    "lsc_hllc2d_init_problem", "lsc_hllc2d_step", "lsc_hllc2d_run_until", "lsc_hllc2d_get_scalars",
    "lsc_hllc2d_download", "lsc_hllc2d_write_dat", "lsc_euler_group_download", "lsc_euler_loop_stats",
    "lsc_euler_create", "lsc_euler_destroy", "lsc_euler_initial_condition", "lsc_euler_upload",
    "lsc_euler_init_problem", "lsc_euler_step", "lsc_euler_step_timed", "lsc_euler_run_until",
    "lsc_euler_get_scalars", "lsc_euler_download", "lsc_euler_write_dat", "lsc_euler_ieee_fallback_count",
    "lsc_euler_set_profiling", "lsc_euler_set_overlap", "lsc_euler_kernel_times", "lsc_euler_local_nx", "lsc_euler_local_i0",
    "lsc_euler_comm_init", "lsc_euler_group_step", "lsc_eos_table_read", "lsc_euler_set_eos_table",
    "lsc_euler_tabeos_initial_condition", "lsc_euler_tabeos_write_dat", "lsc_euler_kat_cfl_bound", "lsc_euler_set_eos_search", "lsc_euler_eos_search_is_warm",
]

_lib = None


def lib():
    """Load liblsc_b200.so (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LscError(-2, f"{LIB_PATH} not built: run `make -C lsc_internship_b200/csrc` "
                           "(no CPU fallback exists)")
    L = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    dp = ctypes.POINTER(ctypes.c_double)
    vp = ctypes.c_void_p
    L.lsc_last_error.restype = ctypes.c_char_p
    L.lsc_kernel_launch_count.restype = ctypes.c_ulonglong
    L.lsc_host_alloc.argtypes = [ctypes.POINTER(vp), ctypes.c_ulonglong]
    L.lsc_host_free.argtypes = [vp]
    L.lsc_mhd_create.argtypes = [ctypes.POINTER(MhdConfig), ctypes.POINTER(vp)]
    L.lsc_mhd_destroy.argtypes = [vp]
    L.lsc_mhd_local_nx.argtypes = [vp]
    L.lsc_mhd_local_i0.argtypes = [vp]
    L.lsc_mhd_initial_condition.argtypes = [vp, ctypes.c_int, dp]
    L.lsc_mhd_upload.argtypes = [vp, dp]
    L.lsc_mhd_init_problem.argtypes = [vp, ctypes.c_int]
    L.lsc_mhd_step.argtypes = [vp, ctypes.c_long]
    L.lsc_mhd_step_timed.argtypes = [vp, ctypes.c_long, ctypes.POINTER(ctypes.c_float)]
    L.lsc_mhd_set_profiling.argtypes = [vp, ctypes.c_int]
    L.lsc_mhd_set_overlap.argtypes = [vp, ctypes.c_int]
    L.lsc_mhd_kernel_times.argtypes = [vp, dp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_mhd_comm_times.argtypes = [vp, dp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_mhd_run_until.argtypes = [vp, ctypes.c_double, DUMP_CB, LOG_CB, vp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_mhd_get_scalars.argtypes = [vp, dp, dp, dp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_mhd_download.argtypes = [vp, dp]
    L.lsc_mhd_write_dat.argtypes = [ctypes.POINTER(MhdConfig), dp, ctypes.c_char_p]
    L.lsc_comm_unique_id.argtypes = [vp]
    L.lsc_mhd_comm_init.argtypes = [vp, vp]
    L.lsc_mhd_kat_hllc.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, dp, dp, ctypes.c_double,
                                   ctypes.c_double, dp]
    L.lsc_mhd_kat_fluxdef.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_long, dp, ctypes.c_double,
                                      ctypes.c_double, dp]
    L.lsc_mhd_set_arithmetic.argtypes = [vp, ctypes.c_int]
    L.lsc_mhd_kat_cfl_bound.argtypes = [ctypes.c_int, ctypes.c_long, dp, ctypes.c_double, ctypes.c_double, dp]
    L.lsc_kat_division.argtypes = [ctypes.c_int, ctypes.c_ulonglong, ctypes.c_int, ctypes.c_long, ctypes.c_int,
                                   ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong)]
    L.lsc_mhd_group_step.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_long]
    L.lsc_mhd_group_step_timed.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_long,
                                           ctypes.POINTER(ctypes.c_float)]
    L.lsc_mhd_group_run_until.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_double, DUMP_CB, LOG_CB, vp,
                                          ctypes.POINTER(ctypes.c_long)]
    L.lsc_mhd_group_download.argtypes = [ctypes.POINTER(vp), ctypes.c_int, dp]
    L.lsc_mhd_loop_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_long), ctypes.POINTER(ctypes.c_long)]
    ecfg = ctypes.POINTER(EulerConfig)
    L.lsc_euler_create.argtypes = [ecfg, ctypes.POINTER(vp)]
    L.lsc_euler_destroy.argtypes = [vp]
    L.lsc_eos_table_read.argtypes = [ctypes.c_char_p, dp, dp, dp, dp]
    L.lsc_euler_set_eos_table.argtypes = [vp, dp, dp, dp, dp]
    L.lsc_euler_set_eos_search.argtypes = [vp, ctypes.c_int]
    L.lsc_euler_eos_search_is_warm.argtypes = [vp]
    L.lsc_euler_tabeos_initial_condition.argtypes = [ecfg, dp, dp, dp, dp]
    L.lsc_euler_tabeos_write_dat.argtypes = [ecfg, dp, ctypes.c_char_p, dp, dp, dp]
    L.lsc_euler_kat_cfl_bound.argtypes = [ctypes.c_int, ctypes.c_long, dp, ctypes.c_double, ctypes.c_double, dp]
    L.lsc_euler_local_nx.argtypes = [vp]
    L.lsc_euler_local_i0.argtypes = [vp]
    L.lsc_euler_comm_init.argtypes = [vp, vp]
    L.lsc_euler_group_step.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_long]
    L.lsc_euler_group_run_until.argtypes = [ctypes.POINTER(vp), ctypes.c_int, ctypes.c_double, EULER_LOG_CB, vp,
                                            ctypes.POINTER(ctypes.c_long)]
    hcfg = ctypes.POINTER(Hllc2dConfig)
    L.lsc_hllc2d_create.argtypes = [hcfg, ctypes.POINTER(vp)]
    L.lsc_hllc2d_destroy.argtypes = [vp]
    L.lsc_hllc2d_initial_condition.argtypes = [hcfg, dp]
    L.lsc_hllc2d_upload.argtypes = [vp, dp]
    L.lsc_hllc2d_init_problem.argtypes = [vp]
    L.lsc_hllc2d_step.argtypes = [vp, ctypes.c_long]
    L.lsc_hllc2d_run_until.argtypes = [vp, ctypes.c_double, LOG_CB, vp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_hllc2d_get_scalars.argtypes = [vp, dp, dp, dp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_hllc2d_download.argtypes = [vp, dp]
    L.lsc_hllc2d_write_dat.argtypes = [hcfg, dp, ctypes.c_char_p]
    L.lsc_guard_check.argtypes = [ctypes.POINTER(ctypes.c_ulonglong), ctypes.POINTER(ctypes.c_ulonglong)]
    L.lsc_euler_loop_stats.argtypes = [vp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_euler_group_download.argtypes = [ctypes.POINTER(vp), ctypes.c_int, dp]
    L.lsc_euler_initial_condition.argtypes = [ecfg, ctypes.c_int, dp]
    L.lsc_euler_upload.argtypes = [vp, dp]
    L.lsc_euler_init_problem.argtypes = [vp, ctypes.c_int]
    L.lsc_euler_step.argtypes = [vp, ctypes.c_long]
    L.lsc_euler_step_timed.argtypes = [vp, ctypes.c_long, ctypes.POINTER(ctypes.c_float)]
    L.lsc_euler_run_until.argtypes = [vp, ctypes.c_double, EULER_LOG_CB, vp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_euler_get_scalars.argtypes = [vp, dp, dp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_euler_download.argtypes = [vp, dp]
    L.lsc_euler_write_dat.argtypes = [ecfg, dp, ctypes.c_char_p, ctypes.c_int]
    L.lsc_euler_ieee_fallback_count.argtypes = [vp]
    L.lsc_euler_ieee_fallback_count.restype = ctypes.c_long
    L.lsc_euler_set_profiling.argtypes = [vp, ctypes.c_int]
    L.lsc_euler_set_overlap.argtypes = [vp, ctypes.c_int]
    L.lsc_euler_kernel_times.argtypes = [vp, dp, ctypes.POINTER(ctypes.c_long)]
    L.lsc_mhd_fill_ghost_ring.argtypes = [ctypes.POINTER(MhdConfig), dp]
    L.lsc_mhd_state_hash.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong)]
    L.lsc_mhd_state_hash_noalloc.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong)]
    L.lsc_euler_state_hash.argtypes = [vp, ctypes.POINTER(ctypes.c_ulonglong)]
    L.lsc_mhd_ieee_fallback_count.argtypes = [vp]
    L.lsc_mhd_ieee_fallback_count.restype = ctypes.c_long
    _lib = L
    return L


_M64 = (1 << 64) - 1


def state_hash_host(u_interior, i0_global=0, ny_total=None):
    """lsc_mhd_state_hash / lsc_euler_state_hash evaluated with numpy on a host array [nxl][ny][nvar] of interior
    values (include/lsc_b200.h): the same splitmix64 sum modulo 2^64.  For checking the device fingerprint."""
    import numpy as np
    a = np.ascontiguousarray(u_interior, dtype=np.float64)
    nxl, ny, nv = a.shape
    ny_total = ny if ny_total is None else ny_total
    with np.errstate(over="ignore"):
        cell = (np.arange(nxl, dtype=np.uint64)[:, None] + np.uint64(i0_global)) * np.uint64(ny_total) + \
            np.arange(ny, dtype=np.uint64)[None, :]
        idx = cell[:, :, None] * np.uint64(nv) + (np.arange(nv, dtype=np.uint64) + np.uint64(1))[None, None, :]
        z = a.view(np.uint64) + np.uint64(0x9E3779B97F4A7C15) * idx
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
        return int(z.sum(dtype=np.uint64)) & _M64


def guard_check():
    """(corrupted guard bytes, watched buffers) -- lsc_guard_check; meaningful when LSC_GUARD=1 was set at create"""
    a, b = ctypes.c_ulonglong(), ctypes.c_ulonglong()
    check(lib().lsc_guard_check(ctypes.byref(a), ctypes.byref(b)))
    return a.value, b.value


def check(rc):
    if rc != LSC_OK:
        raise LscError(rc, lib().lsc_last_error().decode(errors="replace"))
