"""ctypes binding of ``libkolmo_b200.so`` (C ABI declared in ``include/kolmo_b200.h``).

There is no CPU fallback: if the shared library is missing, loading raises, and
if no B200 is present ``kf_create`` fails -- both surface as exceptions.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("KF_LIB") or os.path.join(_HERE, "libkolmo_b200.so")  # KF_LIB: experiments only

KF_F32, KF_F64 = 0, 1

# every symbol include/kolmo_b200.h declares
SYMBOLS = (
    "kf_last_error", "kf_create", "kf_destroy", "kf_set_constants", "kf_rollout", "kf_rk_steps",
    "kf_reward_obs", "kf_step", "kf_step_host", "kf_profile_kernels", "kf_launch_count", "kf_workspace_bytes",
    "kf_chunk",
    "kf_lanes",
    "kf_slab_create", "kf_slab_destroy", "kf_slab_set_constants", "kf_slab_set_action", "kf_slab_import",
    "kf_slab_export", "kf_slab_exchange_bytes", "kf_slab_col", "kf_slab_row", "kf_slab_obs",
    "kf_slab_launch_count", "kf_slab_workspace_bytes", "kf_energy_spectrum",
    "kf_slab_p2p_init", "kf_slab_p2p_attach", "kf_slab_p2p_attach_local", "kf_physical_fields",
    "kf_slab_barrier", "kf_slab_rk_steps", "kf_reward_obs_noalloc",
    "kf_set_control_field", "kf_nonlinear_terms", "kf_linear_terms", "kf_implicit_timestep", "kf_strain_diagnostics",
    # include/channel_b200.h
    "kc_create", "kc_destroy", "kc_set_operators", "kc_set_state", "kc_get_state", "kc_step", "kc_to_spectral",
    "kc_rk4_steps", "kc_rhs", "kc_get_spectral", "kc_to_physical", "kc_launch_count", "kc_workspace_bytes",
)


class KfConfig(ctypes.Structure):
    _fields_ = [
        ("N", ctypes.c_int), ("M", ctypes.c_int), ("B", ctypes.c_int), ("dtype", ctypes.c_int),
        ("nu", ctypes.c_double), ("dt", ctypes.c_double),
        ("n_inner", ctypes.c_int), ("n_outer", ctypes.c_int), ("dealias", ctypes.c_int),
        ("drag", ctypes.c_double), ("device", ctypes.c_int), ("obs_size", ctypes.c_int),
        ("min_action", ctypes.c_double), ("max_action", ctypes.c_double),
        ("chunk", ctypes.c_int), ("use_graph", ctypes.c_int),
    ]


class KcConfig(ctypes.Structure):
    """``kc_config`` of include/channel_b200.h."""
    _fields_ = [
        ("B", ctypes.c_int), ("dtype", ctypes.c_int), ("device", ctypes.c_int),
        ("nu", ctypes.c_double), ("lx", ctypes.c_double), ("ly", ctypes.c_double), ("fx", ctypes.c_double),
        ("target_bulk_u", ctypes.c_double), ("k_det", ctypes.c_int), ("wss_plane", ctypes.c_int),
        ("jets_on", ctypes.c_int),
    ]


_lib = None


def load():
    """Load the shared library once; raise if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(hydrogym_b200 has no CPU fallback)")
    lib = ctypes.CDLL(LIB_PATH)
    vp, ip = ctypes.c_void_p, ctypes.POINTER(ctypes.c_int)
    lib.kf_last_error.restype = ctypes.c_char_p
    lib.kf_last_error.argtypes = []
    lib.kf_create.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(KfConfig)]
    lib.kf_destroy.argtypes = [vp]
    lib.kf_set_constants.argtypes = [vp, vp, vp]
    lib.kf_rollout.argtypes = [vp, vp, vp, vp, vp]
    lib.kf_rk_steps.argtypes = [vp, vp, vp, ctypes.c_int, vp]
    lib.kf_reward_obs.argtypes = [vp, vp, ctypes.c_int, vp, vp, vp, vp]
    lib.kf_reward_obs_noalloc.argtypes = [vp, vp, ctypes.c_int, vp, vp, vp, vp]
    lib.kf_reward_obs_noalloc.restype = ctypes.c_int
    lib.kf_step.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
    lib.kf_step_host.argtypes = [vp, vp, vp, vp, vp, vp]
    fp = ctypes.POINTER(ctypes.c_float)
    lib.kf_profile_kernels.argtypes = [vp, ctypes.c_int, fp, fp, vp]
    for name in ("kf_create", "kf_destroy", "kf_set_constants", "kf_rollout", "kf_rk_steps", "kf_reward_obs",
                 "kf_step", "kf_step_host", "kf_profile_kernels", "kf_chunk", "kf_lanes"):
        getattr(lib, name).restype = ctypes.c_int
    lib.kf_launch_count.argtypes = [vp]
    lib.kf_launch_count.restype = ctypes.c_longlong
    lib.kf_workspace_bytes.argtypes = [vp]
    lib.kf_workspace_bytes.restype = ctypes.c_size_t
    lib.kf_chunk.argtypes = [vp]
    lib.kf_lanes.argtypes = [vp]
    # slab-decomposed single grid
    lib.kf_slab_create.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(KfConfig), ctypes.c_int, ctypes.c_int]
    lib.kf_slab_destroy.argtypes = [vp]
    lib.kf_slab_set_constants.argtypes = [vp, vp, vp]
    lib.kf_slab_set_action.argtypes = [vp, vp, vp]
    lib.kf_slab_import.argtypes = [vp, vp, vp]
    lib.kf_slab_export.argtypes = [vp, vp, vp]
    lib.kf_slab_exchange_bytes.argtypes = [vp, ctypes.c_int]
    lib.kf_slab_exchange_bytes.restype = ctypes.c_size_t
    lib.kf_slab_col.argtypes = [vp, ctypes.c_int, vp, vp, vp]
    lib.kf_slab_row.argtypes = [vp, vp, vp, vp]
    lib.kf_slab_obs.argtypes = [vp, vp, vp, vp, ctypes.c_double, vp]
    for name in ("kf_slab_create", "kf_slab_destroy", "kf_slab_set_constants", "kf_slab_set_action",
                 "kf_slab_import", "kf_slab_export", "kf_slab_col", "kf_slab_row", "kf_slab_obs"):
        getattr(lib, name).restype = ctypes.c_int
    lib.kf_slab_launch_count.argtypes = [vp]
    lib.kf_slab_launch_count.restype = ctypes.c_longlong
    lib.kf_slab_workspace_bytes.argtypes = [vp]
    lib.kf_slab_workspace_bytes.restype = ctypes.c_size_t
    lib.kf_slab_p2p_init.argtypes = [vp, vp]
    lib.kf_slab_p2p_attach.argtypes = [vp, vp]
    lib.kf_slab_p2p_attach_local.argtypes = [vp, ctypes.POINTER(vp)]
    lib.kf_slab_barrier.argtypes = [vp, vp]
    lib.kf_slab_rk_steps.argtypes = [vp, ctypes.c_int, vp]
    for name in ("kf_slab_p2p_init", "kf_slab_p2p_attach", "kf_slab_p2p_attach_local", "kf_slab_barrier",
                 "kf_slab_rk_steps"):
        getattr(lib, name).restype = ctypes.c_int
    lib.kf_physical_fields.argtypes = [vp, vp, vp, ctypes.c_int, vp]
    lib.kf_physical_fields.restype = ctypes.c_int
    lib.kf_set_control_field.argtypes = [vp, vp]
    lib.kf_nonlinear_terms.argtypes = [vp, vp, vp, vp, vp]
    cd, ci = ctypes.c_double, ctypes.c_int
    lib.kf_linear_terms.argtypes = [ci, ci, ci, cd, vp, vp, vp]
    lib.kf_implicit_timestep.argtypes = [ci, ci, ci, cd, cd, vp, vp, vp]
    lib.kf_strain_diagnostics.argtypes = [ci, ci, ci, cd, vp, vp, vp, vp, vp]
    for name in ("kf_set_control_field", "kf_nonlinear_terms", "kf_linear_terms", "kf_implicit_timestep",
                 "kf_strain_diagnostics"):
        getattr(lib, name).restype = ctypes.c_int
    lib.kf_energy_spectrum.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, vp, vp, ctypes.c_int, vp]
    lib.kf_energy_spectrum.restype = ctypes.c_int
    # 3-D channel flow (include/channel_b200.h)
    dp = ctypes.POINTER(ctypes.c_double)
    lib.kc_create.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(KcConfig)]
    lib.kc_destroy.argtypes = [vp]
    lib.kc_set_operators.argtypes = [vp] + [dp] * 8
    lib.kc_set_state.argtypes = [vp, vp, vp, vp, vp]
    lib.kc_get_state.argtypes = [vp, vp, vp, vp, vp]
    lib.kc_step.argtypes = [vp, vp, dp, ci, cd, cd, cd, vp, vp, vp]
    lib.kc_to_spectral.argtypes = [vp, vp]
    lib.kc_to_physical.argtypes = [vp, vp]
    lib.kc_rk4_steps.argtypes = [vp, vp, dp, ci, cd, cd, cd, vp]
    lib.kc_rhs.argtypes = [vp, vp, dp, cd, cd, vp, vp]
    lib.kc_get_spectral.argtypes = [vp, vp, vp]
    for name in ("kc_create", "kc_destroy", "kc_set_operators", "kc_set_state", "kc_get_state", "kc_step",
                 "kc_to_spectral", "kc_to_physical", "kc_rk4_steps", "kc_rhs", "kc_get_spectral"):
        getattr(lib, name).restype = ctypes.c_int
    lib.kc_launch_count.argtypes = [vp]
    lib.kc_launch_count.restype = ctypes.c_longlong
    lib.kc_workspace_bytes.argtypes = [vp]
    lib.kc_workspace_bytes.restype = ctypes.c_ulonglong
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().kf_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed: {msg}")
