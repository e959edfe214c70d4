"""ctypes binding of libshipsim.so (include/shipsim.h).  There is no fallback: if the library is
missing or no CUDA device is present, every compute call raises."""
from __future__ import annotations

import ctypes as C
import os

MAX_WP, MAX_MODES, MAX_EDGES = 16, 4, 128
N_OBS, N_TRAJ, N_LOG, N_STATS = 10, 8, 27, 16
ABI_VERSION = 3
F64, F32 = 0, 1
KIND_FLEET, KIND_ENV = 1, 2
SG = {"MOTOR": 0, "GEN": 1, "OFF": 2}
DT_F64, DT_F32, DT_I32 = 0, 1, 2

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SIT_LIBSHIPSIM", os.path.join(_HERE, "libshipsim.so"))   # override: kernel experiments

D, I32, I64 = C.c_double, C.c_int32, C.c_int64


class SitShipParams(C.Structure):
    _fields_ = [
        ("dt", D),
        ("dead_weight_tonnage", D), ("coefficient_of_deadweight_to_displacement", D), ("bunkers", D), ("ballast", D),
        ("length_of_ship", D), ("width_of_ship", D),
        ("added_mass_coefficient_in_surge", D), ("added_mass_coefficient_in_sway", D),
        ("added_mass_coefficient_in_yaw", D),
        ("mass_over_linear_friction_coefficient_in_surge", D), ("mass_over_linear_friction_coefficient_in_sway", D),
        ("mass_over_linear_friction_coefficient_in_yaw", D),
        ("nonlinear_friction_coefficient__in_surge", D), ("nonlinear_friction_coefficient__in_sway", D),
        ("nonlinear_friction_coefficient__in_yaw", D),
        ("n_modes", I32), ("machinery_kind", I32),
        ("mode_main_engine_capacity", D * MAX_MODES), ("mode_electrical_capacity", D * MAX_MODES),
        ("mode_shaft_generator_state", I32 * MAX_MODES),
        ("hotel_load", D),
        ("linear_friction_main_engine", D), ("linear_friction_hybrid_shaft_generator", D),
        ("gear_ratio_between_main_engine_and_propeller", D),
        ("gear_ratio_between_hybrid_shaft_generator_and_propeller", D),
        ("propeller_inertia", D), ("propeller_speed_to_torque_coefficient", D), ("propeller_diameter", D),
        ("propeller_speed_to_thrust_force_coefficient", D),
        ("rudder_angle_to_sway_force_coefficient", D), ("rudder_angle_to_yaw_force_coefficient", D),
        ("max_rudder_angle_degrees", D),
        ("sfc_me", D * 3), ("sfc_dg", D * 3),
        ("kp_ship_speed", D), ("ki_ship_speed", D), ("kp_shaft_speed", D), ("ki_shaft_speed", D),
        ("kp_heading", D), ("kd_heading", D), ("ki_heading", D),
        ("radius_of_acceptance", D), ("lookahead_distance", D), ("los_integral_gain", D),
        ("los_integrator_windup_limit", D),
        ("thrust_force_dynamic_time_constant", D), ("kp_simplified", D), ("ki_simplified", D),
    ]


class SitMap(C.Structure):
    _fields_ = [("n_edges", I32), ("n_poly", I32),
                ("x1", D * MAX_EDGES), ("y1", D * MAX_EDGES), ("x2", D * MAX_EDGES), ("y2", D * MAX_EDGES),
                ("poly", I32 * MAX_EDGES)]


class SitConfig(C.Structure):
    _fields_ = [("abi_version", I32), ("kind", I32), ("precision", I32), ("max_wp", I32),
                ("ship", SitShipParams * 2), ("map", SitMap),
                ("sampling_frequency", D), ("theta", D)]


class SitField(C.Structure):
    _fields_ = [("name", C.c_char * 32), ("offset", I64), ("dtype", I32), ("rows", I32), ("count", I64)]


class SitFleetIO(C.Structure):
    _fields_ = [("dist", C.c_void_p), ("dist_steps", I32), ("avoid", I32), ("store", I32), ("traj_steps", I32),
                ("traj", C.c_void_p), ("log27", C.c_void_p), ("cmd", C.c_void_p)]


class SitEnvIO(C.Structure):
    _fields_ = [("action", C.c_void_p), ("sac_update", C.c_void_p), ("act_steps", I32), ("init", I32),
                ("auto_reset", I32), ("rec_steps", I32),
                ("obs", C.c_void_p), ("reward", C.c_void_p), ("done", C.c_void_p), ("status", C.c_void_p),
                ("reward_sum", C.c_void_p), ("done_any", C.c_void_p), ("stats", C.c_void_p),
                ("reward_parts", C.c_void_p), ("snap", C.c_void_p)]


class SitPolicy(C.Structure):
    _fields_ = [("mode", I32), ("hidden", I32), ("w0", C.c_void_p), ("b0", C.c_void_p), ("w1", C.c_void_p),
                ("b1", C.c_void_p), ("w2", C.c_void_p), ("b2", C.c_void_p), ("obs_scale", C.c_void_p)]


class SitRollout(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("env_id_offset", I64), ("max_episode_steps", I32),
                ("max_route_resampling", I32), ("replay", C.c_void_p), ("capacity", I64), ("counters", C.c_void_p),
                ("truncate_episodes", I32), ("_pad", I32)]


class SitSacConfig(C.Structure):
    _fields_ = [("hidden", I32), ("batch_size", I32), ("gamma", C.c_float), ("tau", C.c_float), ("lr", C.c_float),
                ("alpha", C.c_float), ("automatic_entropy_tuning", I32), ("target_update_interval", I32),
                ("seed", C.c_uint64), ("obs_scale", C.c_float * 10), ("tensor_cores", I32), ("_pad", I32)]


N_REWARD_PARTS = 8
TR_FLOATS = 24
MACH_FULL, MACH_SIMPLIFIED = 0, 1
MAX_CIRCLES = 64
NOALLOC_MAX_ENVS = 16384
POLICY_UNIFORM, POLICY_GAUSSIAN, POLICY_MEAN = 0, 1, 2
POLICY_TC_BF16 = 0x100
RC_TRANSITIONS, RC_DECISIONS, RC_RESAMPLES, RC_EPISODES, RC_CLOSED, RC_N = 0, 1, 2, 3, 4, 8

EXPORTS = ("sit_create", "sit_destroy", "sit_state_bytes", "sit_layout", "sit_n_ships", "sit_fleet_step_k", "sit_fleet_step_host", "sit_fleet_step_host_sel",
           "sit_capture_init", "sit_reset", "sit_env_init_step", "sit_env_step_k", "sit_env_step_host", "sit_env_step_host_noalloc", "sit_map_query", "sit_rollout_begin", "sit_rollout_group_count", "sit_rollout_policy", "sit_rollout_step_k", "sit_rollout_collect",
           "sit_sac_create", "sit_sac_destroy", "sit_sac_workspace_bytes", "sit_sac_field", "sit_sac_init", "sit_sac_policy",
           "sit_sac_update", "sit_replay_pack", "sit_replay_append", "sit_circle_query", "sit_ship_outline",
           "sit_last_error", "sit_status_string", "sit_abi_version", "sit_measure_fma_peak")

_lib = None


class ShipSimError(RuntimeError):
    pass


def lib():
    """Load libshipsim.so (built by sac_maritime_ast_b200/csrc/build.py).  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ShipSimError(f"{LIB_PATH} not found: build it with `python -m sac_maritime_ast_b200.csrc.build` "
                           "(there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, P = C.c_void_p, C.POINTER
    L.sit_create.argtypes = [P(SitConfig), I64, C.c_int, P(vp)]
    L.sit_destroy.argtypes = [vp]
    L.sit_state_bytes.argtypes = [vp, P(C.c_size_t)]
    L.sit_layout.argtypes = [vp, P(SitField), I32, P(I32)]
    L.sit_n_ships.argtypes = [vp]
    L.sit_n_ships.restype = I64
    L.sit_fleet_step_k.argtypes = [vp, vp, I32, P(SitFleetIO), vp]
    L.sit_fleet_step_host.argtypes = [vp, vp, I32, vp, I32, vp, vp]
    L.sit_fleet_step_host_sel.argtypes = [vp, vp, I32, vp, I32, vp, I32, vp, vp]
    L.sit_capture_init.argtypes = [vp, vp, vp]
    L.sit_reset.argtypes = [vp, vp, vp, I32, vp]
    L.sit_env_init_step.argtypes = [vp, vp, vp, vp]
    L.sit_env_step_k.argtypes = [vp, vp, I32, P(SitEnvIO), vp]
    L.sit_env_step_host.argtypes = [vp, vp, I32, vp, vp, I32, vp, vp, vp, vp, vp]
    L.sit_env_step_host_noalloc.argtypes = [vp, vp, I32, vp, vp, I32, vp, vp, vp, vp, vp]
    L.sit_rollout_begin.argtypes = [vp, vp, vp]
    L.sit_rollout_group_count.argtypes = [vp, C.POINTER(C.c_int32)]
    L.sit_rollout_policy.argtypes = [vp, vp, P(SitPolicy), P(SitRollout), vp]
    L.sit_rollout_step_k.argtypes = [vp, vp, I32, P(SitRollout), P(SitEnvIO), vp]
    L.sit_rollout_collect.argtypes = [vp, vp, I32, I32, P(SitPolicy), P(SitRollout), P(SitEnvIO), I32, vp]
    L.sit_map_query.argtypes = [vp, I64, vp, vp, vp, vp, vp]
    L.sit_sac_create.argtypes = [P(SitSacConfig), C.c_int, P(vp)]
    L.sit_sac_destroy.argtypes = [vp]
    L.sit_sac_workspace_bytes.argtypes = [vp, P(C.c_size_t)]
    L.sit_sac_field.argtypes = [vp, C.c_char_p, P(I64), P(I64)]
    L.sit_sac_init.argtypes = [vp, vp, vp]
    L.sit_sac_policy.argtypes = [vp, vp, P(SitPolicy)]
    L.sit_sac_update.argtypes = [vp, vp, vp, I64, vp, I32, vp, I32, vp]
    L.sit_replay_pack.argtypes = [P(SitRollout), vp, vp, vp, I64, vp]
    L.sit_replay_append.argtypes = [vp, I64, vp, vp, I32, I64, vp]
    L.sit_circle_query.argtypes = [I32, I32, vp, I64, vp, vp, vp, vp, vp]
    L.sit_ship_outline.argtypes = [I32, I64, vp, vp, vp, vp, vp]
    L.sit_last_error.restype = C.c_char_p
    L.sit_status_string.argtypes = [I32]
    L.sit_status_string.restype = C.c_char_p
    L.sit_measure_fma_peak.argtypes = [C.c_int, I32, P(D), P(D)]
    if L.sit_abi_version() != ABI_VERSION:
        raise ShipSimError("libshipsim ABI version mismatch")
    _lib = L
    return L


def check(rc: int):
    if rc != 0:
        raise ShipSimError(f"libshipsim error {rc}: {lib().sit_last_error().decode()}")


def status_strings():
    L = lib()
    return tuple(L.sit_status_string(i).decode() for i in range(12))
