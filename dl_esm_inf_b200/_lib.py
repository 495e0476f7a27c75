"""ctypes loader for libdlesm_b200.so (the product).  Fails loudly if the library is missing:
there is no Python / CPU fallback for any device operation."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdlesm_b200.so")
MAXCOMM = 16


class Region(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("xstart", C.c_int), ("xstop", C.c_int),
                ("ystart", C.c_int), ("ystop", C.c_int)]

    def astuple(self):
        return (self.nx, self.ny, self.xstart, self.xstop, self.ystart, self.ystop)

    def box(self):
        return (self.xstart, self.xstop, self.ystart, self.ystop)

    def __repr__(self):
        return "region(nx=%d, ny=%d, x=%d:%d, y=%d:%d)" % self.astuple()


class Subdomain(C.Structure):
    _fields_ = [("glob", Region), ("internal", Region)]  # `global` is a Python keyword


class Halo(C.Structure):
    _fields_ = [("needs_update", C.c_int), ("source", Region), ("dest", Region)]


_ia = C.c_int * MAXCOMM


class Comms(C.Structure):
    _fields_ = [("nsend", C.c_int), ("nrecv", C.c_int)] + [
        (n, _ia) for n in ("dirsend", "destination", "isrcsend", "jsrcsend", "idessend",
                           "jdessend", "nxsend", "nysend", "nsendp", "dirrecv", "source",
                           "isrcrecv", "jsrcrecv", "idesrecv", "jdesrecv", "nxrecv", "nyrecv",
                           "nrecvp")]


_ll = C.c_longlong * MAXCOMM


class CommPlan(C.Structure):
    """dlesm_comm_plan: the per-peer message plan of the device exchange"""
    _fields_ = [("npeers", C.c_int), ("peer", _ia), ("nsend", _ia), ("nrecv", _ia),
                ("send_slot", _ia * MAXCOMM), ("recv_slot", _ia * MAXCOMM),
                ("send_off", _ll), ("recv_off", _ll), ("send_n", _ll), ("recv_n", _ll),
                ("send_total", C.c_longlong), ("recv_total", C.c_longlong)]


class GridInfo(C.Structure):
    _fields_ = [("name", C.c_int), ("offset", C.c_int), ("global_nx", C.c_int),
                ("global_ny", C.c_int), ("nx", C.c_int), ("ny", C.c_int), ("dx", C.c_double),
                ("dy", C.c_double), ("boundary_conditions", C.c_int * 3),
                ("subdomain", Subdomain), ("alignment", C.c_int), ("padding", C.c_int),
                ("initialised", C.c_int)]


class FieldInfo(C.Structure):
    _fields_ = [("defined_on", C.c_int), ("internal", Region), ("whole", Region),
                ("num_halos", C.c_int), ("halo", Halo * 4), ("nx", C.c_int), ("ny", C.c_int),
                ("ntiles", C.c_int), ("data_on_device", C.c_int), ("device_ptr", C.c_void_p)]


class GoArg(C.Structure):
    """dlesm_go_arg: go_arg of src/argument_mod.f90:57-61 plus the element family"""
    _fields_ = [("arg_intent", C.c_int), ("kind", C.c_int), ("element", C.c_int), ("stencil", C.c_int * 3)]


# every symbol include/dlesm_b200.h declares; tests/test_abi.py checks header == this == .so
_V = C.c_void_p
_I = C.c_int
_D = C.c_double
_B = C.c_bool
_PD = C.POINTER(C.c_double)
_PI = C.POINTER(C.c_int)
SIGNATURES = {
    "dlesm_last_error": (C.c_char_p, []),
    "dlesm_version": (C.c_char_p, []),
    "dlesm_source_hash": (C.c_char_p, []),
    "dlesm_parallel_init": (_I, [_I, _I]),
    "dlesm_device_init": (_I, [_I]),
    "dlesm_comm_unique_id": (_I, [_V]),
    "dlesm_comm_init": (_I, [_V]),
    "dlesm_parallel_finalise": (_I, []),
    "dlesm_parallel_abort": (None, [C.c_char_p]),
    "dlesm_get_rank": (_I, []),
    "dlesm_get_num_ranks": (_I, []),
    "dlesm_on_master": (_I, []),
    "dlesm_global_sum": (_I, [_PD]),
    "dlesm_gather_host": (_I, [_PD, C.c_longlong, _PD]),
    "dlesm_reset_saved_state": (None, []),
    "dlesm_set_extensions": (_I, [_I]),
    "dlesm_go_decompose": (_V, [_I] * 6),
    "dlesm_last_status": (_I, []),
    "dlesm_decomp_info": (_I, [_V, _PI]),
    "dlesm_decomp_subdomain": (_I, [_V, _I, C.POINTER(Subdomain)]),
    "dlesm_decomp_proc_subdomain": (_I, [_V, _I, _I]),
    "dlesm_decomp_free": (None, [_V]),
    "dlesm_map_comms": (_I, [_V, _I, _I, _I, _I, _I, _I, C.POINTER(Comms)]),
    "dlesm_iprocmap": (_I, [_V, _I, _I, _I]),
    "dlesm_comm_plan_build": (_I, [C.POINTER(Comms), _I, C.POINTER(CommPlan)]),
    "dlesm_grid_create": (_V, [_I, _PI, _I]),
    "dlesm_grid_decompose": (_I, [_V] + [_I] * 6),
    "dlesm_grid_init": (_I, [_V, _D, _D, _PI, _I, _I]),
    "dlesm_grid_get_info": (_I, [_V, C.POINTER(GridInfo)]),
    "dlesm_grid_decomp": (_V, [_V]),
    "dlesm_grid_comms": (_I, [_V, C.POINTER(Comms)]),
    "dlesm_grid_host_array": (_I, [_V, _I, _V]),
    "dlesm_grid_device_array": (_I, [_V, _I, C.POINTER(_V)]),
    "dlesm_grid_free": (None, [_V]),
    "dlesm_r2d_field_create": (_V, [_V, _I, _I, _PD, _I, _I]),
    "dlesm_field_get_info": (_I, [_V, C.POINTER(FieldInfo)]),
    "dlesm_field_grid": (_V, [_V]),
    "dlesm_field_tile": (_I, [_V, _I, C.POINTER(Region), C.POINTER(Region)]),
    "dlesm_field_halo_is_dirty": (_I, [_V, _I]),
    "dlesm_field_set_halo_dirty": (_I, [_V]),
    "dlesm_grid_field_bounds": (_I, [_V, _I, C.POINTER(FieldInfo)]),
    "dlesm_field_set_data": (_I, [_V, _PD]),
    "dlesm_field_get_data": (_I, [_V, _PD]),
    "dlesm_field_write_to_device": (_I, [_V, _PD, _I, _I, _I, _I, _B]),
    "dlesm_field_read_from_device": (_I, [_V, _PD, _I, _I, _I, _I, _B]),
    "dlesm_field_halo_exchange": (_I, [_V, _I]),
    "dlesm_fields_halo_exchange": (_I, [C.POINTER(_V), _I, _I]),
    "dlesm_field_apply_periodic_halos": (_I, [_V]),
    "dlesm_field_gather_inner_data": (_I, [_V, _PD]),
    "dlesm_copy_field": (_I, [_V, _V]),
    "dlesm_copy_field_patch": (_I, [_V, C.POINTER(Region), C.POINTER(Region)]),
    "dlesm_set_field": (_I, [_V, _D]),
    "dlesm_field_checksum": (_I, [_V, _PD]),
    "dlesm_free_field": (_I, [_V]),
    "dlesm_dev_field_alloc": (_I, [_I, _I, C.POINTER(_V)]),
    "dlesm_dev_field_free": (_I, [_V]),
    "dlesm_read_from_device_c": (None, [_V, _V, _I, _I, _I, _I, _B]),
    "dlesm_write_to_device_c": (None, [_V, _V, _I, _I, _I, _I, _B]),
    "dlesm_dev_sync": (_I, []),
    "dlesm_host_alloc": (_I, [C.c_size_t, C.POINTER(_V)]),
    "dlesm_host_free": (_I, [_V]),
    "dlesm_host_register": (_I, [_V, C.c_size_t]),
    "dlesm_host_unregister": (_I, [_V]),
    "dlesm_event_record": (_I, [_I]),
    "dlesm_event_elapsed_ms": (_I, [_I, _I, _PD]),
    "dlesm_dev_copy_field": (_I, [_V, _V]),
    "dlesm_dev_copy_patch": (_I, [_V] + [_I] * 8),
    "dlesm_dev_copy_patch_static": (_I, [_V] + [_I] * 10 + [_V, _I]),
    "dlesm_dev_set_field": (_I, [_V, _D]),
    "dlesm_dev_checksum": (_I, [_V, _I, _I, _I, _I, _PD]),
    "dlesm_dev_periodic_halos": (_I, [_V, _I, _I, _I, _I, _I, _I]),
    "dlesm_dev_set_comms": (_I, [C.POINTER(Comms)]),
    "dlesm_dev_halo_exchange": (_I, [C.POINTER(_V), _I]),
    "dlesm_host_halo_exchange": (_I, [_PD, _I, _I]),
    "dlesm_dev_set_exchange_mode": (_I, [_I]),
    "dlesm_dev_peer_register": (_I, [C.POINTER(_V), _I]),
    "dlesm_dev_peer_unregister": (_I, [C.POINTER(_V), _I]),
    "dlesm_dev_peer_unregister_all": (_I, []),
    "dlesm_set_arith": (_I, [_I]),
    "dlesm_get_arith": (_I, []),
    "dlesm_dev_set_sweep_tiles": (_I, [_I, _I]),
    "dlesm_dev_grid_upload": (_I, [_I, _I, _PI, C.POINTER(_PD), C.POINTER(_V)]),
    "dlesm_dev_grid_release": (_I, [C.POINTER(_V)]),
    "dlesm_compute_cu": (_I, [_V, _V, _V, _I, _I, _I, _I, _V]),
    "dlesm_compute_cv": (_I, [_V, _V, _V, _I, _I, _I, _I, _V]),
    "dlesm_compute_z": (_I, [_V, _V, _V, _V, _D, _D, _I, _I, _I, _I, _V]),
    "dlesm_compute_h": (_I, [_V, _V, _V, _V, _I, _I, _I, _I, _V]),
    "dlesm_compute_unew": (_I, [_V] * 5 + [_D, _D, _I, _I, _I, _I, _V, _V]),
    "dlesm_compute_vnew": (_I, [_V] * 5 + [_D, _D, _I, _I, _I, _I, _V, _V]),
    "dlesm_compute_pnew": (_I, [_V] * 4 + [_D, _D, _D, _I, _I, _I, _I, _V, _V]),
    "dlesm_time_smooth": (_I, [_V, _V, _V, _D, _I, _I, _I, _I, _V, _I]),
    "dlesm_sw_step_fused": (_I, [_V] * 9 + [_D, _D, _D, _D, _I, _I, _I, _I, _I, _V, _I]),
    "dlesm_sw_create": (_V, [_V, _D, _D, _I, _I]),
    "dlesm_sw_init_device": (_I, [_V]),
    "dlesm_sw_set_state": (_I, [_V, _PD, _PD, _PD]),
    "dlesm_sw_step": (_I, [_V, _I]),
    "dlesm_sw_step_host": (_I, [_V, C.POINTER(_PD), C.POINTER(_PD), _I]),
    "dlesm_sw_step_host_n": (_I, [_V, C.POINTER(_PD), C.POINTER(_PD), _I, _I]),
    "dlesm_sw_field": (_V, [_V, _I]),
    "dlesm_sw_set_variant": (_I, [_V, _I]),
    "dlesm_sw_set_halo_mode": (_I, [_V, _I]),
    "dlesm_sw_set_two_step": (_I, [_V, _I, _I]),
    "dlesm_sw_step2_fused": (_I, [_V] * 12 + [_D, _D, _D, _D, _I, _I, _I, _I, _V, _I]),
    "dlesm_dev_periodic_halos_deep": (_I, [_V, _I, _I, _I, _I, _I]),
    "dlesm_kernel_launch_count": (C.c_longlong, []),
    "dlesm_profile_enable": (_I, [_I]),
    "dlesm_profile_read": (_I, [_PD, C.POINTER(C.c_longlong), _I]),
    "dlesm_debug_cta_trace_begin": (_I, [_I]),
    "dlesm_debug_cta_trace_end": (_I, [C.POINTER(C.c_longlong), _I]),
    "dlesm_sw_free": (None, [_V]),
    "dlesm_sw_free_local": (None, [_V]),
    "dlesm_sw_launch_grid": (_I, [_V, _I, _PI]),
    "dlesm_tileset_create": (_V, [_I, _PI, _I, _I, _I, _I, _I, _I, _I]),
    "dlesm_tileset_ntiles": (_I, [_V]),
    "dlesm_tileset_grid": (_V, [_V, _I]),
    "dlesm_tileset_tile_id": (_I, [_V, _I]),
    "dlesm_tileset_decomp": (_V, [_V]),
    "dlesm_tileset_free": (None, [_V]),
    "dlesm_tileset_plan": (_I, [_V, _I, _I, _I, _I, _I, C.POINTER(C.c_longlong), _I, _PI, C.POINTER(C.c_longlong)]),
    "dlesm_tileset_halo_exchange": (_I, [_V, C.POINTER(_V), _I, _I]),
    "dlesm_tileset_gather_inner_data": (_I, [_V, C.POINTER(_V), _PD]),
    "dlesm_tileset_sw_step": (_I, [_V, C.POINTER(_V), _I]),
    "dlesm_tileset_sw_sync_state": (_I, [_V, C.POINTER(_V)]),
    "dlesm_kernel_create": (_V, [C.c_char_p, C.c_char_p, C.POINTER(GoArg), _I, _I, _I]),
    "dlesm_kernel_generated_source": (C.c_char_p, [_V]),
    "dlesm_kernel_launch": (_I, [_V, C.POINTER(_V), _I, _PD, _I, _PI, _I]),
    "dlesm_kernel_launch_dev": (_I, [_V, C.POINTER(_V), _I, _PD, _I, _PI, _I, _I, _I, _I, _I]),
    "dlesm_kernel_free": (None, [_V]),
}

_lib = None


def _preload_nccl():
    """libdlesm_b200.so needs `libnccl.so.2`.  If the launcher (torch.distributed) is imported
    later in the same process it insists on its own, newer wheel copy of that soname; loading
    the system copy first would leave torch with missing symbols.  So prefer the wheel's copy
    when it exists (same soname: the dynamic loader then reuses it for both)."""
    import importlib.util
    try:
        spec = importlib.util.find_spec("nvidia")
    except (ImportError, ValueError):
        spec = None
    for base in (spec.submodule_search_locations if spec and spec.submodule_search_locations else []):
        cand = os.path.join(base, "nccl", "lib", "libnccl.so.2")
        if os.path.exists(cand):
            C.CDLL(cand, mode=C.RTLD_GLOBAL)
            return


def lib():
    """The loaded product library.  Raises if it has not been built (python -m
    dl_esm_inf_b200.build, or __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            "libdlesm_b200.so is missing (%s). Build it with `python -m dl_esm_inf_b200.build`; "
            "there is no CPU fallback." % LIB_PATH)
    _preload_nccl()
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L
