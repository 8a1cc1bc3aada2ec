"""ctypes mirror of include/tfode_batch.h (the C ABI of the batched rover stepping engine).

Field order and types must match the header exactly; tests/test_abi.py checks the sizes
against the compiled library.
"""
import ctypes as C

ABI_VERSION = 2
NBODY_ROVER = 10
NBODY = 12
NOBS = 17
NACT = 2
MLP_HIDDEN = 512
TRANSITION_WIDTH = 40
ROLLOUT_WIDTH = 24
NOALLOC_MAX_ENVS = 128   # TFODE_NOALLOC_MAX_ENVS
POLICY_NFLOATS = 17 * 512 + 512 + 512 * 512 + 512 + 512 * 2 + 2  # 272 898, scripts/rover_training_1.py:26-35

SCENE_PLANE, SCENE_STEP_BOXES, SCENE_HEIGHTFIELD = 0, 1, 2
CTRL_EXTERNAL, CTRL_RANDOM, CTRL_MLP, CTRL_CRAWLER, CTRL_MODEL_TREE = 0, 1, 2, 3, 4
TREE_INPUTS, TREE_MAX_FEATURES = 14, 119
POLICY_FP32, POLICY_TC = 0, 1
SOLVER_QUICKSTEP, SOLVER_DANTZIG = 0, 1
TERM_NONE, TERM_TIMEOUT, TERM_Y_MAX, TERM_X_GOAL, TERM_UPSIDE_DOWN = 0, 1, 2, 3, 4
PRESET_SCENE_1, PRESET_MT_RANDOM, PRESET_TRAINING, PRESET_CRAWLING = 0, 1, 2, 3


class BatchConfig(C.Structure):
    _fields_ = [
        ("abi_version", C.c_int32), ("n_envs", C.c_int32), ("device", C.c_int32), ("scene", C.c_int32),
        ("controller", C.c_int32), ("policy_precision", C.c_int32), ("env_id_offset", C.c_int32),
        ("auto_reset", C.c_int32),
        ("mu", C.c_float), ("dt", C.c_float), ("erp", C.c_float), ("cfm", C.c_float),
        ("pgs_iters", C.c_int32), ("sor_w", C.c_float),
        ("crawling_mode", C.c_int32), ("ic_period", C.c_float), ("ic_initially_active", C.c_int32),
        ("exploration", C.c_int32), ("fork_callback", C.c_int32),
        ("speedf", C.c_float), ("term", C.c_float), ("ic_start", C.c_float), ("ic_start_jitter", C.c_float),
        ("timeout", C.c_float), ("x_goal", C.c_float), ("y_max", C.c_float), ("x_goal_abs", C.c_int32),
        ("check_upside_down", C.c_int32), ("terminate", C.c_int32),
        ("step_yaw_deg", C.c_float), ("step_yaw_jitter_deg", C.c_float),
        ("crawler_amp", C.c_float), ("crawler_dt_torque", C.c_float), ("crawler_rate", C.c_float),
        ("crawler_amp_jitter", C.c_float), ("crawler_rate_jitter", C.c_float),
        ("hf_heights", C.POINTER(C.c_float)), ("hf_ncol", C.c_int32), ("hf_nrow", C.c_int32),
        ("hf_length", C.c_float), ("hf_width", C.c_float), ("hf_pos", C.c_float * 3),
        ("seed", C.c_uint64), ("max_contacts", C.c_int32), ("transition_capacity", C.c_int32), ("solver", C.c_int32), ("reserved_", C.c_int32),
    ]


class TreeNode(C.Structure):
    _fields_ = [("id", C.c_int32), ("terminal", C.c_int32), ("child", C.c_int32 * 2), ("param_off", C.c_int32), ("n_params", C.c_int32)]


class EnvState(C.Structure):
    _fields_ = [
        ("pos", (C.c_double * 3) * NBODY), ("quat", (C.c_double * 4) * NBODY),
        ("lvel", (C.c_double * 3) * NBODY), ("avel", (C.c_double * 3) * NBODY),
        ("facc", (C.c_double * 3) * NBODY), ("tacc", (C.c_double * 3) * NBODY),
        ("wheel_vel", C.c_double * 4), ("wheel_fmax", C.c_double * 4), ("servo_vel", C.c_double),
        ("robot_speed", C.c_double), ("steering_rate", C.c_double), ("boggie_torque", C.c_double),
        ("W", C.c_double * 4), ("ic_clock", C.c_double), ("ic_period", C.c_double),
        ("filt", (C.c_double * 4) * 12), ("ft", C.c_double * 12),
        ("last_pos", C.c_double * 3), ("total_reward", C.c_double), ("last_state", C.c_double * NOBS),
        ("last_reward", C.c_double), ("step_yaw_deg", C.c_double),
        ("crawler_amp", C.c_double), ("crawler_rate", C.c_double), ("crawler_dtorque_dt", C.c_double),
        ("speed", C.c_float), ("ic_start", C.c_float), ("time_steps", C.c_int64),
        ("lcg_seed", C.c_uint32), ("episode", C.c_uint32), ("tick", C.c_uint32),
        ("ic_activated", C.c_int32), ("ic_tick", C.c_int32), ("servo_mode_vel", C.c_int32),
        ("has_last_state", C.c_int32), ("explore", C.c_int32), ("fork_collision", C.c_int32),
        ("crawler_phase", C.c_int32), ("done", C.c_int32), ("term_reason", C.c_int32),
        ("n_contacts", C.c_int32), ("episodes_done", C.c_int32),
    ]
