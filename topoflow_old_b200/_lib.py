"""ctypes binding of libtopoflow_b200.so (include/topoflow_b200.h).

There is no CPU fallback: if the shared library is missing this module raises
at import, and every compute call needs a CUDA device.  Build the library with
``python -c "import __graft_entry__ as g; g.build()"`` or
``make -C topoflow_old_b200/csrc``.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('TF_B200_LIB') or os.path.join(_HERE, 'libtopoflow_b200.so')

if not os.path.exists(LIB_PATH):
    raise ImportError(
        'topoflow_old_b200: %s not found. The routing and D8 components run only '
        'on the CUDA library (no CPU fallback); build it with '
        '`make -C topoflow_old_b200/csrc`.' % LIB_PATH)

lib = C.CDLL(LIB_PATH)

handle_t = C.c_void_p
_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_u8p = C.POINTER(C.c_uint8)

# name -> (restype, argtypes); mirrors include/topoflow_b200.h one to one
SIGNATURES = {
    'tf_last_error': (C.c_char_p, []),
    'tf_version': (C.c_char_p, []),
    'tf_device_count': (C.c_int, []),
    'tf_create': (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_uint, C.POINTER(handle_t)]),
    'tf_set_strip': (C.c_int, [handle_t, C.c_int, C.c_int, C.c_int, C.c_int]),
    'tf_destroy': (C.c_int, [handle_t]),
    'tf_sync': (C.c_int, [handle_t]),
    'tf_d8_build': (C.c_int, [handle_t, _fp, _fp, _fp, _fp, C.c_float, C.c_int, C.c_int,
                              C.c_int, _i64p]),
    'tf_d8_set_row0_sizes': (C.c_int, [handle_t, C.c_float, C.c_float, C.c_float]),
    'tf_d8_set_codes': (C.c_int, [handle_t, _u8p, _fp, _fp, _fp, C.c_int, C.c_int]),
    'tf_d8_area': (C.c_int, [handle_t, _dp, C.c_int, _i64p]),
    'tf_d8_area_rounds': (C.c_longlong, [handle_t]),
    'tf_d8_get': (C.c_int, [handle_t, C.c_int, C.c_void_p]),
    'tf_d8_set_halo_dem': (C.c_int, [handle_t, _fp, _fp]),
    'tf_chan_set_param': (C.c_int, [handle_t, C.c_int, _dp, C.c_double, C.c_int]),
    'tf_chan_set_angle': (C.c_int, [handle_t, _dp, C.c_double, C.c_int, _dp, C.POINTER(C.c_int)]),
    'tf_chan_set_angle_table': (C.c_int, [handle_t, _dp, _dp, C.c_int]),
    'tf_chan_init': (C.c_int, [handle_t, C.c_int, C.c_int, C.c_double, _dp, _i32p, C.c_int,
                               C.c_int]),
    'tf_chan_set_rho': (C.c_int, [handle_t, C.c_double]),
    'tf_chan_set_state': (C.c_int, [handle_t, C.c_int, _dp]),
    'tf_chan_set_R': (C.c_int, [handle_t, _dp, C.c_double, _dp]),
    'tf_chan_set_forcing': (C.c_int, [handle_t, C.c_int, _dp, C.c_double]),
    'tf_chan_assemble_R': (C.c_int, [handle_t, _dp]),
    'tf_chan_set_points': (C.c_int, [handle_t, C.c_int, _i32p, _i32p, _dp, C.c_int]),
    'tf_chan_step': (C.c_int, [handle_t, C.c_int, _dp, _dp, _i64p]),
    'tf_chan_step_async': (C.c_int, [handle_t, C.c_int]),
    'tf_chan_get': (C.c_int, [handle_t, C.c_int, _dp]),
    'tf_chan_minmax': (C.c_int, [handle_t, _dp]),
    'tf_chan_timing': (C.c_int, [handle_t, _dp]),
    'tf_chan_frame': (C.c_int, [handle_t, C.c_int, C.c_int, _fp, C.c_int]),
    'tf_chan_frame_wait': (C.c_int, [handle_t]),
    'tf_host_alloc': (C.c_int, [C.c_size_t, C.POINTER(C.c_void_p)]),
    'tf_host_free': (C.c_int, [C.c_void_p]),
    'tf_sum_R_noalloc': (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_void_p, C.c_double,
                                   C.c_double, C.c_void_p, C.c_void_p]),
    'tf_grid_new': (C.c_int, [handle_t, C.POINTER(C.c_int)]),
    'tf_grid_set': (C.c_int, [handle_t, C.c_int, _dp, C.c_double]),
    'tf_grid_get': (C.c_int, [handle_t, C.c_int, _dp]),
    'tf_grid_copy': (C.c_int, [handle_t, C.c_int, C.c_int]),
    'tf_grid_axpb': (C.c_int, [handle_t, C.c_int, C.c_double, C.c_int, C.c_double, C.c_double, C.c_int]),
    'tf_interp_linear': (C.c_int, [handle_t, C.c_int, C.c_int, C.c_double, C.c_double, C.c_int, C.c_int,
                                   C.POINTER(C.c_int)]),
    'tf_prov_met_rain': (C.c_int, [handle_t, C.c_int, C.c_double, C.c_int, C.c_double, C.c_int, C.c_int]),
    'tf_prov_snow_degree_day': (C.c_int, [handle_t] + [C.c_int, C.c_double] * 4 + [C.c_double, C.c_int, C.c_double,
                                                                                  C.c_double, C.c_int, C.c_int,
                                                                                  C.c_int, _dp]),
    'tf_prov_evap_priestley_taylor': (C.c_int, [handle_t] + [C.c_int, C.c_double] * 8 + [C.c_double, C.c_int, _dp]),
    'tf_prov_infil_green_ampt': (C.c_int, [handle_t] + [C.c_int, C.c_double] * 7 + [C.c_int, C.c_int, C.c_double,
                                                                                   C.c_int, C.c_int, _dp]),
    'tf_chan_set_forcing_dev': (C.c_int, [handle_t, C.c_int, C.c_int, C.c_double, C.c_int, C.c_double]),
    'tf_chan_assemble_R_dev': (C.c_int, [handle_t, _dp, _dp]),
    'tf_fill_pits': (C.c_int, [handle_t, _fp, C.c_float, _fp, _i64p, _i64p]),
    'tf_satzone_dzw': (C.c_int, [handle_t, _dp, _dp, C.c_double, _dp, C.c_double, _dp]),
    'tf_halo_plan': (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_int)]),
    'tf_comm_unique_id': (C.c_int, [_u8p]),
    'tf_comm_init': (C.c_int, [handle_t, C.c_int, C.c_int, _u8p]),
}

for _name, (_res, _args) in SIGNATURES.items():
    _fn = getattr(lib, _name)       # AttributeError here = header / library mismatch
    _fn.restype = _res
    _fn.argtypes = _args

# enums of the header
KINEMATIC, DIFFUSIVE, DYNAMIC = 0, 1, 2
MANNING, LAW_OF_WALL = 0, 1
P_WIDTH, P_TAN_ANGLE, P_COS_ANGLE, P_SLOPE, P_ROUGH, P_DS = range(6)
(F_VOL, F_D, F_U, F_Q, F_QNEXT, F_F, F_RH, F_A_WET, F_P_WET, F_TAU, F_U_STAR,
 F_FROUDE, F_S_FREE, F_WIDTH, F_MUTATED, F_R) = range(16)
R_P_RAIN, R_SM, R_GW, R_MR, R_ET, R_IN = range(6)
D8_CODE, D8_PARENT_ID, D8_DS, D8_DW, D8_AREA, D8_INMASK, D8_RAW = range(7)
PT_SOURCES, PT_SINKS, PT_CANALS_IN, PT_CANALS_OUT = range(4)


class TopoflowB200Error(RuntimeError):
    """A C-ABI call returned non-zero; the message is tf_last_error()."""


def check(rc):
    if rc != 0:
        raise TopoflowB200Error(lib.tf_last_error().decode('utf-8', 'replace'))


def ptr(a, ctype):
    """Pointer to a C-contiguous NumPy array (or None)."""
    if a is None:
        return None
    return a.ctypes.data_as(C.POINTER(ctype))
