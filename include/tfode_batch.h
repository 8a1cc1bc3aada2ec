/*
 * tfode_batch.h -- C ABI of the B200-native batched rover stepping engine.
 *
 * The reference (arthur-bouton/TF_ODE) has no FFI for this path: its seams are
 * C++ virtuals and one Boost.Python module.  Each entry point
 * below names the reference seam it replaces (paths relative to the reference
 * root).  One handle = one GPU = one host thread.  All calls return 0 on
 * success or a negative TFODE_E* code; tfode_last_error() gives the message.
 * No exceptions, no torch/STL types cross this boundary.
 *
 * Build: tf_ode_b200/csrc/build.sh -> tf_ode_b200/csrc/libtfode_b200.so
 */
#ifndef TFODE_BATCH_H_
#define TFODE_BATCH_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TFODE_ABI_VERSION 2

/* ---- model sizes (src/rover_1.cc:116-217; src/rover_training_1.cc:90-98) -- */
#define TFODE_NBODY_ROVER 10 /* main, battery, rear, boggie, front fork, rear fork, wheel0..3 */
#define TFODE_NBODY 12       /* + step, step_c (dynamic 1 kg bodies held by a slider, ode/object.cc:155-179) */
#define TFODE_NOBS 17        /* src/rover_1_tf.cc:60-77 */
#define TFODE_NACT 2         /* steering rate [deg/s], boggie torque [N.m] */
#define TFODE_MLP_HIDDEN 512 /* scripts/rover_training_1.py:26-35 */
#define TFODE_TRANSITION_WIDTH 40 /* env_id, s[17], a[2], r, terminal, s'[17], pad */
#define TFODE_ROLLOUT_WIDTH 24    /* obs17, reward, done, last episode return / steps, episodes done, global env id, pad */
#define TFODE_TREE_INPUTS 14       /* Rover_1_mt::GetState(flip=false, full=false), src/rover_1_mt.cc:30-52 */
#define TFODE_TREE_MAX_FEATURES 119 /* 14 + 14*15/2: Polynomial_model_tree of degree 2 (model_tree.cc:91-123) */

/* ---- error codes ---------------------------------------------------------- */
enum {
  TFODE_OK = 0,
  TFODE_EINVAL = -1,   /* bad argument (std::runtime_error in the reference) */
  TFODE_ECUDA = -2,    /* CUDA runtime failure; message has cudaGetErrorString */
  TFODE_ENOMEM = -3,
  TFODE_ESTATE = -4    /* call not valid in this state (e.g. MLP controller without a policy) */
};

/* ---- scene kinds (what the step function's terrain is) -------------------- */
enum {
  TFODE_SCENE_PLANE = 0,      /* src/crawling.cc:19 (plane only) */
  TFODE_SCENE_STEP_BOXES = 1, /* src/rover_training_1.cc:84-98, src/scene_1.cc:45-53 */
  TFODE_SCENE_HEIGHTFIELD = 2 /* commented lines src/scene_1.cc:60-62, scene_1_mt.cc:65-67 */
};

/* ---- controller kinds (Rover_1::_InternalControl overrides) --------------- */
enum {
  TFODE_CTRL_EXTERNAL = 0, /* src/scene_1.cc: IC off; commands come from tfode_batch_set_actions (keyboard in the reference) */
  TFODE_CTRL_RANDOM = 1,   /* BASELINE config 2/5: U(-15,15) deg/s, U(-25,25) N.m every IC tick */
  TFODE_CTRL_MLP = 2,      /* src/rover_1_tf.cc:111-199 (17-512-512-2 actor, e-greedy optional) */
  TFODE_CTRL_CRAWLER = 3,  /* src/rover_1_crawlers.cc:40-77 */
  TFODE_CTRL_MODEL_TREE = 4 /* src/rover_1_mt.cc:72-83 (two linear / polynomial model trees loaded from YAML) */
};

/* ---- world step (ode/environment.hh:132-133) ------------------------------- */
enum {
  TFODE_SOLVER_QUICKSTEP = 0, /* dWorldQuickStep: 20 SOR-LCP sweeps with random reordering (the north-star path, the fast one) */
  TFODE_SOLVER_DANTZIG = 1    /* dWorldStep: dense A = J M^-1 J^T per island, Dantzig pivoting LCP (include/tfode_lcp.h) -- what the shipped
                                 binaries call; a correctness path (one lane per island), an order of magnitude slower than QuickStep */
};

/* ---- policy arithmetic ----------------------------------------------------- */
enum {
  TFODE_POLICY_FP32 = 0, /* fp32, one fma per weight in input order (bit-identical to the oracle): inside tfode_batch_step the
                            actor runs fused in step_kernel at each environment's own IC tick and writes the servo / boggie
                            setpoints directly; tfode_policy_infer uses the batched CUDA-core kernel */
  TFODE_POLICY_TC = 1    /* tcgen05 tensor-core path (bf16 operands, fp32 TMEM accumulators): environments park at their IC
                            tick and a batched GEMM kernel serves the parked list between step_kernel launches */
};

/* ---- termination reasons (src/rover_training_1.cc:142) -------------------- */
enum {
  TFODE_TERM_NONE = 0,
  TFODE_TERM_TIMEOUT = 1,
  TFODE_TERM_Y_MAX = 2,
  TFODE_TERM_X_GOAL = 3,
  TFODE_TERM_UPSIDE_DOWN = 4
};

/*
 * Batch configuration (POD).  Defaults follow the reference; fill with
 * tfode_batch_config_default() and override.
 */
typedef struct TfodeBatchConfig {
  int32_t abi_version;      /* must be TFODE_ABI_VERSION */
  int32_t n_envs;           /* environments on this GPU */
  int32_t device;           /* CUDA device ordinal */
  int32_t scene;            /* TFODE_SCENE_* */
  int32_t controller;       /* TFODE_CTRL_* */
  int32_t policy_precision; /* TFODE_POLICY_* */
  int32_t env_id_offset;    /* global id of local env 0 (multi-GPU sharding; RNG is keyed on the global id) */
  int32_t auto_reset;       /* 1: an env that terminates is re-initialised in the same step (new episode) */

  /* world (ode/environment.cc:29-55; libode defaults) */
  float mu;                 /* Environment(mu): 0.5 scene_1/_mt/training, 0.6 crawling */
  float dt;                 /* Sim_loop timestep, 0.001f (kept as float, Q5) */
  float erp;                /* 0.2 */
  float cfm;                /* 1e-10 */
  int32_t pgs_iters;        /* 20 */
  float sor_w;              /* 1.3 */

  /* rover flags (src/rover_training_1.cc:58-65) */
  int32_t crawling_mode;    /* SetCrawlingMode */
  float ic_period;          /* SetCmdPeriod, seconds; 0 = every step */
  int32_t ic_initially_active; /* Rover_1 ctor: true; training/mt scenes call DeactivateIC */
  int32_t exploration;      /* Rover_1_tf::SetExploration */
  int32_t fork_callback;    /* Rover_1_tf: fork geoms carry the "ground" AABB callback (Q10) */

  /* scene rules (src/rover_training_1.cc:100-125) */
  float speedf;             /* cruise speed 0.04 */
  float term;               /* time to reach cruise speed 0.5; <=0 disables the ramp (crawling.cc) */
  float ic_start;           /* 1.0; <0 = IC never auto-activated */
  float ic_start_jitter;    /* trial: 0.25 (IC_start += jitter*U(-1,1)) */
  float timeout;            /* 60 */
  float x_goal;             /* 1.5 training, 1.0 mt, 10 crawling */
  float y_max;              /* 0.6; <=0 disables (crawling.cc has none) */
  int32_t x_goal_abs;       /* 1: |x|>=x_goal (training/mt); 0: x>=x_goal (crawling.cc:62) */
  int32_t check_upside_down;/* 1 except crawling.cc */
  int32_t terminate;        /* 0: the step function never ends the episode (scene_1.cc:96) */

  /* step boxes (TFODE_SCENE_STEP_BOXES) */
  float step_yaw_deg;       /* orientation argument; */
  float step_yaw_jitter_deg;/* trial: 5 (orientation = 5*U(-1,1)) */

  /* crawler (src/crawling.cc:27) */
  float crawler_amp, crawler_dt_torque, crawler_rate;
  float crawler_amp_jitter, crawler_rate_jitter; /* BASELINE config 4: amp in U(20,30), rate in U(5,15) => 5,5 around 25,10 */

  /* heightfield (ode/heightfield.hh:42-59,91-105) */
  const float* hf_heights;  /* HOST pointer, [hf_nrow*hf_ncol], h[iz*ncol+ix] metres; iz=0 is the image's top row */
  int32_t hf_ncol, hf_nrow;
  float hf_length, hf_width;/* extent along HF-local x (world x) and z (world -y) */
  float hf_pos[3];          /* geom position (world) */

  uint64_t seed;            /* Philox key (20260411) */
  int32_t max_contacts;     /* per-env contact capacity (1..24); extra contacts are dropped in generation order */
  int32_t transition_capacity; /* rows in the transition ring (0 = 4*n_envs) */
  int32_t solver;           /* TFODE_SOLVER_*: which world step Environment::next_step runs (ode/environment.hh:132-133) */
  int32_t reserved_;        /* keeps the struct a multiple of 8 bytes; must be 0 */
} TfodeBatchConfig;

/*
 * Full per-env state in exchange format (doubles), used by the parity tests to
 * copy a state between the engine and the CPU oracle.  Field names follow the
 * reference members they mirror.
 */
typedef struct TfodeEnvState {
  double pos[TFODE_NBODY][3];
  double quat[TFODE_NBODY][4];   /* w,x,y,z (ODE order) */
  double lvel[TFODE_NBODY][3];
  double avel[TFODE_NBODY][3];
  double facc[TFODE_NBODY][3];   /* force accumulators written by robot.next_step, consumed by the next world step (Q6) */
  double tacc[TFODE_NBODY][3];
  double wheel_vel[4];           /* dParamVel of the wheel hinges (rover_1.cc:398) */
  double wheel_fmax[4];          /* dParamFMax (rover_1.cc:397) */
  double servo_vel;              /* Servo::_vel == dParamVel of the centre hinge */
  double robot_speed, steering_rate, boggie_torque; /* Rover_1 members */
  double W[4];
  double ic_clock, ic_period;
  double filt[12][4];            /* LP_second_order: y_k1, y_k2, x_k1, x_k2 (Filters/cpp/filters.cc:137-144) */
  double ft[12];                 /* FT_sensor _F/_T after in-place filtering (Q9) */
  double last_pos[3];            /* Rover_1_tf::_last_pos */
  double total_reward;
  double last_state[TFODE_NOBS];
  double last_reward;            /* reward computed at the last IC tick */
  double step_yaw_deg;           /* this episode's step orientation */
  double crawler_amp, crawler_rate, crawler_dtorque_dt;
  float speed;                   /* the scene lambda's float ramp variable */
  float ic_start;
  int64_t time_steps;            /* Sim_loop::_time */
  uint32_t lcg_seed;             /* dRand() state of this env's world */
  uint32_t episode;
  uint32_t tick;                 /* IC ticks this episode */
  int32_t ic_activated, ic_tick, servo_mode_vel;
  int32_t has_last_state, explore, fork_collision;
  int32_t crawler_phase;
  int32_t done, term_reason;     /* set by the step in which the episode ended */
  int32_t n_contacts;            /* contacts created by the last world step */
  int32_t episodes_done;
} TfodeEnvState;

typedef struct TfodeBatch TfodeBatch;

/* Fill *cfg with the reference defaults of scene `preset`:
 * 0 scene_1 (external commands), 1 scene_1_mt-style random actions,
 * 2 rover_training_1 (MLP), 3 crawling.  */
int tfode_batch_config_default(TfodeBatchConfig* cfg, int preset);

/* ode::Environment + robot::Rover_1* constructors for n_envs worlds
 * (ode/environment.cc:29-55, src/rover_1.cc:18-245, src/rover_training_1.cc:41-125). */
int tfode_batch_create(const TfodeBatchConfig* cfg, TfodeBatch** out);
void tfode_batch_destroy(TfodeBatch* b);
const char* tfode_last_error(const TfodeBatch* b); /* b may be NULL: error of the failed create */

/* All device work of this handle is enqueued on `stream` (a cudaStream_t). */
int tfode_batch_set_stream(TfodeBatch* b, void* stream);

/* New episode = new objects in the reference (src/rover_training_1.cc:41-125).
 * env_ids: HOST int32 array of local ids, or NULL = all envs. */
int tfode_batch_reset(TfodeBatch* b, const int32_t* env_ids, int n);

/* Rover_1::SetSteeringRate / SetBoggieTorque (src/rover_1.cc:326-341) for every env.
 * actions: DEVICE float [n_envs,2] = (steering rate deg/s, boggie torque N.m). */
int tfode_batch_set_actions(TfodeBatch* b, const float* actions_dev);
/* Rover_1::SetRobotSpeed (src/rover_1.cc:308-311). speed_dev: DEVICE float[n_envs], NULL = built-in ramp. */
int tfode_batch_set_speed(TfodeBatch* b, const float* speed_dev);

/* TF_model<float> ctor (tf_cpp_binding.cc:14-66): install the actor.
 * weights: HOST fp32, Keras layout [in,out] row-major, concatenated
 * W1[17,512] b1[512] W2[512,512] b2[512] W3[512,2] b3[2]  (272 898 floats). */
int tfode_batch_load_policy(TfodeBatch* b, const float* weights, size_t n_floats);

/* TF_model<float>( actor_dir, s_dim, a_dim ) (tf_cpp_binding.cc:14-66; called with the directory written by
 * `td3.actor.save( session_dir + '/actor' )`, scripts/rover_training_1.py:90,121, from src/rover_1_tf.cc:24):
 * read the three Dense layers of scripts/rover_training_1.py:26-35 out of
 * <saved_model_dir>/variables/variables.index + variables.data-XXXXX-of-YYYYY (TensorFlow tensor-bundle checkpoint:
 * keys layer_with_weights-{0,1,2}/{kernel,bias}/.ATTRIBUTES/VARIABLE_VALUE, DT_FLOAT, little-endian) into the blob
 * layout of tfode_batch_load_policy.  *weights_out is malloc'ed (tfode_free).  Anything else -- missing file or
 * variable, wrong dtype or shape, compressed index blocks, checksum mismatch -- is TFODE_EINVAL with the reason in
 * tfode_last_error(NULL).  The graph (saved_model.pb) is not read: the architecture is the reference's fixed actor.
 * FORMAT UNPINNED (no SavedModel ships with the reference, TensorFlow absent here). */
int tfode_savedmodel_read_actor(const char* saved_model_dir, float** weights_out, size_t* n_floats_out);
/* The two calls above in one: what Rover_1_tf's constructor does with its `actor_path`. */
int tfode_batch_load_policy_saved_model(TfodeBatch* b, const char* saved_model_dir);

/* n x { step_function(dt,time) } of src/rover_training_1.cc:127-146 for every env:
 * speed ramp, IC activation, Environment::next_step (ode/environment.hh:127-136 with
 * dWorldQuickStep), Rover_1::next_step (src/rover_1.cc:478-503), termination test. */
int tfode_batch_step(TfodeBatch* b, int n_substeps);

/* Rover_1_tf::GetState etc.  All pointers DEVICE, any may be NULL.
 * obs17 [n,17] float (unscaled), reward [n] (last IC-tick reward), done [n] (episode ended
 * during the last tfode_batch_step call), contact_count [n] (last world step). */
int tfode_batch_observe(TfodeBatch* b, float* obs17, float* reward, uint8_t* done, int32_t* contact_count);
/* What one policy tick hands the trainer, one packed row per env so that a single collective (or copy) moves it
 * (the experience list / result line of src/rover_training_1.cc:203-224,243-246): DEVICE float [n, TFODE_ROLLOUT_WIDTH] =
 * obs17 | reward of the last IC tick | done during the last call | return, steps of the last finished episode |
 * episodes finished | global env id | 0. */
int tfode_batch_pack_rollout(TfodeBatch* b, float* rows_dev);
/* pos(3) quat(4) linvel(3) angvel(3) of the 12 bodies: DEVICE float [n,12,13]. */
int tfode_batch_get_bodies(TfodeBatch* b, float* out_dev);

/* Experience tuples (src/rover_1_tf.cc:120-122, rover_training_1.cc:212-224) appended since the last drain, in
 * completion order: DEVICE float [max_rows, TFODE_TRANSITION_WIDTH].
 *   *n_rows_host    rows written to out_dev (HOST int);
 *   *n_dropped_host rows produced beyond the ring's transition_capacity since the last drain and therefore lost
 *                   (HOST int, may be NULL; 0 means the experience list is complete -- size transition_capacity for
 *                   episodes_per_call x ticks_per_episode rows, the default 4*n_envs covers 2 s of simulated time per call).
 * If the pending rows do not fit max_rows (or out_dev is NULL) NOTHING is consumed: the call returns TFODE_EINVAL with
 * *n_rows_host = rows pending, so the caller can retry with a buffer of that many rows. */
int tfode_batch_drain_transitions(TfodeBatch* b, float* out_dev, int max_rows, int* n_rows_host, int* n_dropped_host);

/* Contact-capacity audit.  The reference creates every contact dCollide returns (ode/environment.cc:101-128: up to 10 per
 * geom pair, no per-world limit); this engine keeps max_contacts (<= 24) per environment and drops the excess in generation
 * order.  *over_capacity_steps = environment sub-steps since creation (or the last reset of the counters) in which more
 * contacts were generated than kept; *max_generated = the largest number one environment generated in one sub-step.
 * A non-zero first value means trajectories deviate from the reference's for that reason.  HOST pointers, either may be NULL. */
int tfode_batch_contact_stats(TfodeBatch* b, int64_t* over_capacity_steps, int64_t* max_generated, int reset);

/* The reference's debug print formats that scripts/bin/monitor-policies and the tree-fitting scripts read from stdout, built
 * from one experience row (HOST float[TFODE_TRANSITION_WIDTH] as drained above):
 *   TFODE_PRINT_TRANSITIONS        src/rover_1_tf.cc:125-136   "s0 .. s16 steer torque s'0 .. s'16\n"   (all "%f")
 *   TFODE_PRINT_STATE_AND_ACTIONS  src/rover_1_tf.cc:190-195   "s0 .. s16 steer torque\n" -- the state of a tick and the action chosen
 *                                  at it are the (state, action) of the tuple the NEXT tick appends, so the same rows serve.
 * Writes a NUL-terminated line into out[cap]; returns its length, or TFODE_EINVAL (bad kind, cap too small: 1024 is enough). */
enum { TFODE_PRINT_TRANSITIONS = 0, TFODE_PRINT_STATE_AND_ACTIONS = 1 };
int tfode_format_transition(const float* row_host, int kind, char* out, size_t cap);

/* TFODE_SOLVER_DANTZIG: islands (summed over environments and sub-steps since creation) whose pivoting drive stalled and was abandoned with
 * the remaining multipliers at zero -- lcp.cpp's "s <= 0" exit.  0 in every tested scene. */
int tfode_batch_lcp_failures(TfodeBatch* b, int64_t* stalled_islands);

/* A stepping call that allocates nothing, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  NO COUNTERPART IN THE REFERENCE.
 * It is tfode_batch_step with the dense Dantzig solver (step_kernel<TFODE_SOLVER_DANTZIG>), whose per-island LCP work arrays --
 * cudaMalloc'd at creation for a TFODE_SOLVER_DANTZIG handle, 211 KB per environment -- live here in a file-scope
 * `static __device__` array of the library instead.  So a handle created with TFODE_SOLVER_QUICKSTEP (no dense-solver scratch)
 * can take dWorldStep-equivalent steps, and the call itself makes no cudaMalloc / cudaFree / allocating library call:
 *   actions_dev     DEVICE float [n,2] as tfode_batch_set_actions takes, or NULL to keep the current commands       (input)
 *   n_substeps      sub-steps to run, >= 0
 *   bodies_out_dev  DEVICE float [n,12,13] as tfode_batch_get_bodies writes, or NULL                                (output)
 * all enqueued on the handle's stream, three launches at most.  Results are bit-identical to tfode_batch_set_actions +
 * tfode_batch_step + tfode_batch_get_bodies on a TFODE_SOLVER_DANTZIG handle in the same state.
 * Limits: n_envs <= TFODE_NOALLOC_MAX_ENVS (else TFODE_EINVAL with a message in tfode_last_error); controllers that park
 * environments between launches (model trees, TFODE_POLICY_TC) are refused with TFODE_ESTATE.  The array (26.2 MiB) is held by every
 * GPU that loads the library until the process exits.  All calls on one device share it -- from any handle or stream -- and it keeps
 * the last call's contents (nothing in it needs clearing: every island writes its work arrays before reading them): two such
 * calls that overlap on the GPU (different streams, no ordering between them) RACE and corrupt each other's solves. */
#define TFODE_NOALLOC_MAX_ENVS 128
int tfode_batch_step_dantzig_static(TfodeBatch* b, const float* actions_dev, int n_substeps, float* bodies_out_dev);

/* Host-buffer round trip used by the end-to-end measurement: H2D actions (may be NULL),
 * n_substeps, D2H obs17/reward/done (HOST pointers, pinned or pageable; any may be NULL). */
int tfode_batch_rollout_host(TfodeBatch* b, const float* actions_host, int n_substeps,
                             float* obs17_host, float* reward_host, uint8_t* done_host);

/* Exchange-format state of one env (HOST pointers). */
int tfode_batch_get_state(TfodeBatch* b, int env, TfodeEnvState* out);
int tfode_batch_set_state(TfodeBatch* b, int env, const TfodeEnvState* in);

/* Counters: kernels launched by this handle since creation. */
int64_t tfode_batch_launch_count(const TfodeBatch* b);
int tfode_batch_n_envs(const TfodeBatch* b);
int tfode_batch_sync(TfodeBatch* b);

/* Batched policy op on its own (TF_model::infer, tf_cpp_binding.cc:70-113):
 * obs_scaled DEVICE [m,17] -> act DEVICE [m,2] (tanh outputs, unscaled). */
int tfode_policy_infer(TfodeBatch* b, const float* obs_scaled_dev, int m, float* act_dev);

/* ---- model-tree policy (src/rover_mt.hh, scripts/ModelTree/cpp/model_tree.{hh,cc}) -------------- */

/* One node of a flattened tree, in the order Linear_model_tree::_build_tree_recursively visits them
 * (model_tree.cc:21-45: root id 1, ids handed out depth-first, child 0 before child 1). */
typedef struct TfodeTreeNode {
  int32_t id;          /* Node::id */
  int32_t terminal;    /* Node::terminal */
  int32_t child[2];    /* indices into the node array, -1 for a terminal node */
  int32_t param_off;   /* Node::params = params[param_off .. param_off + n_params) */
  int32_t n_params;
} TfodeTreeNode;

/* Linear_model_tree<double>(yaml_file_path, oblique) (model_tree.cc:11-45) on YAML text already in memory.
 * Outputs are malloc'ed (release with tfode_free); errors -> TFODE_EINVAL, message in tfode_last_error(NULL)
 * (std::runtime_error in the reference, e.g. "Read 3 split parameters for the node 2 instead of the 2 required ..."). */
int tfode_model_tree_parse(const char* yaml_text, size_t len, int oblique,
                           TfodeTreeNode** nodes_out, int32_t* n_nodes_out, double** params_out, int32_t* n_params_out);

/* Rover_1_mt ctor (src/rover_1_mt.cc:13-27): install tree 1 (steering rate) and tree 2 (boggie torque).
 * degree 1 = Linear_model_tree, 2 = Polynomial_model_tree(degree 2, interaction_only); higher degrees -> TFODE_EINVAL.
 * Every terminal node (and every split of an oblique tree) must hold n_features + 1 parameters, every straight
 * split a feature index inside the feature vector: what the reference only detects at predict() time
 * (model_tree.cc:50-51) is rejected here at load time. */
int tfode_batch_load_model_trees(TfodeBatch* b, const char* yaml_1, size_t len_1, const char* yaml_2, size_t len_2,
                                 int oblique, int degree, int interaction_only);

/* Rover_1_mt::InferAction (src/rover_1_mt.cc:56-69) on its own, batched:
 * state DEVICE double [m,14] -> act DEVICE double [m,2] (steering rate, boggie torque, unclamped),
 * node DEVICE int32 [m,2] (terminal node ids node_1, node_2; may be NULL). */
int tfode_model_tree_infer(TfodeBatch* b, const double* state_dev, int m, double* act_dev, int32_t* node_dev);

/* Rover_1_mt::node_1 / node_2 of every env after its last IC tick: DEVICE int32 [n,2]. */
int tfode_batch_get_tree_nodes(TfodeBatch* b, int32_t* out_dev);

/* ---- heightmap images (ode/heightfield.hh:42-59) -------------------------------------------------- */

/* HeightField(env, pos, heightimage_path, z_scale, ...): decode a PNG and build the height array the reference hands to
 * dGeomHeightfieldDataBuildDouble: heights[row*ncol + col] = first byte of pixel (col,row) * z_scale / 255, row 0 = top
 * row of the picture.  NOTE the reference ignores its own `int z_scale` argument and hard-codes 0.3
 * (heightfield.hh:54); pass 0.3 to reproduce it.  Supported: non-interlaced, 8-bit grey / grey+alpha / RGB / RGBA,
 * 1/2/4/8-bit grey or palette.  *heights_out is malloc'ed (tfode_free).  Unreadable file -> TFODE_EINVAL with the
 * reference's message "Can't open <path>". */
int tfode_heightmap_load_png(const char* path, double z_scale, float** heights_out, int32_t* nrow_out, int32_t* ncol_out);
int tfode_heightmap_decode_png(const void* png_bytes, size_t len, double z_scale, float** heights_out, int32_t* nrow_out, int32_t* ncol_out);
void tfode_free(void* p);

#ifdef __cplusplus
}
#endif
#endif /* TFODE_BATCH_H_ */
