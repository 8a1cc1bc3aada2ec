// Device-side data layout of the batched rover stepping engine (sm_100a).
//
// One environment = one EnvRec (env-major records, 16-byte aligned) in HBM.  A warp owns
// one environment: lane b (< 12) owns rigid body b, so inside a record every per-body
// scalar is stored body-contiguous (struct-of-arrays over bodies: body[field][b]) and the
// warp's loads/stores of a record are fully coalesced float4 accesses.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/tfode_batch.h"

namespace tfode {

constexpr int NB = TFODE_NBODY;       // 12 bodies (10 rover + 2 step boxes)
constexpr int NBR = TFODE_NBODY_ROVER;
constexpr int NJ = 9;                 // battery slider, centre hinge, boggie hinge, 4 wheel hinges, 2 step fixes
constexpr int MAXC = 24;              // contact capacity per env
constexpr int MAXJROWS = 6 * NJ;      // 54
constexpr int MAXROWS = MAXJROWS + 3 * MAXC;  // 126
constexpr int NSPHERE = 200;          // 4 wheels x 50 tire spheres (ode/wheel.hh:63-73)
constexpr int NNS = 11;               // non-sphere rover geoms: 3 chassis boxes, 4 motor cylinders, 4 rims
constexpr int LCP_ROW_F = 3 * MAXROWS + 40, LCP_ROW_I = 8;   // per-row scratch of the Dantzig path: an island of m rows needs 3 m^2 + 37 m floats, 7 m ints
constexpr int NISL = 5;               // base islands: {main,battery,rear,boggie} {ffork,w0,w1} {rfork,w2,w3} {step} {step_c}

// body[field][b] rows
enum { F_POS = 0, F_QUAT = 3, F_LVEL = 7, F_AVEL = 10, F_FACC = 13, F_TACC = 16, F_NFIELD = 19 };

// Controller / episode block of a record (Rover_1, Servo, Rover_1_tf, Crawler_1 and scene-lambda members).
struct Ctrl {
  double filt[12][4];      // LP_second_order state y1,y2,x1,x2 (fp64: double pole at z~0.9875)
  double ic_clock;         // Rover_1::_ic_clock (fp64 so that 500 x 0.001f crosses 0.5 on the same step as the reference)
  double robot_speed, steering_rate, boggie_torque;
  double total_reward, last_reward;
  double last_pos[3];
  double cr_dtorque_dt;
  double pending_reward;   // reward of the held-back tuple (fp64 until the terminal penalty is known)
  long long time_steps;    // Sim_loop::_time
  float wheel_vel[4], wheel_fmax[4], W[4];
  float servo_vel;
  float ft[12];            // filtered F/T (what GetState sees)
  float last_state[17];
  float pending[38];       // held-back experience tuple (terminal flag is patched retroactively, rover_training_1.cc:212-224)
  float speed, ic_start, ic_period;
  float step_yaw_deg, cr_amp, cr_rate;
  float fix_qrel[2][4], fix_offset[2][3];
  float jq_angle[NJ], jq_rate[NJ];  // hinge angle / rate of the current pose (servo, boggie, wheels)
  float last_ep_return, last_ep_x, last_ep_y;
  float policy_act[2];     // tanh outputs written by the policy kernel while the env is parked
  uint32_t lcg_seed, episode, tick;
  int32_t ic_activated, ic_tick, servo_mode_vel, has_last_state, explore, fork_collision, pending_valid;
  int32_t cr_phase, done_call, done_step, term_reason, n_contacts, episodes_done, frozen, parked;
  int32_t last_ep_steps, n_rows;
  int32_t tree_nodes;      // Rover_1_mt::node_1 | node_2 << 16 (terminal nodes reached at the last IC tick)
};

struct __align__(16) EnvRec {
  float body[F_NFIELD][NB];  // 228 floats
  Ctrl c;
};

// Per-GPU constants of the rover model (built on the host in double, src/rover_1.cc:49-233 restated as tables).
struct Model {
  double robot_max_speed;
  double filt_c[3];                   // Cy1, Cy2, Cx0 of LP_second_order_bilinear(0.001f, 4*pi, 0.5)
  float invMass[NB], mass[NB];
  float Ib[NB][3], invIb[NB][3];
  float pos0[NB][3], quat0[NB][4];    // construction pose
  int jb1[NJ], jb2[NJ], jslider[NJ];
  float anchor1[NJ][3], anchor2[NJ][3], axis1[NJ][3], axis2[NJ][3], qrel[NJ][4], offset[NJ][3];
  float lostop[NJ], histop[NJ];
  float ring[50][3];                   // tire sphere centres in the wheel frame (z = 0)
  float sphere_r;
  int ns_type[NNS], ns_body[NNS];      // 1 = box, 2 = cylinder
  float ns_off[NNS][3];                // geom offset position (body frame)
  int ns_rot[NNS];                     // 1: has an offset rotation
  float ns_offR[NNS][9];               // dRFromEulerAngles(pi/2,0,0) as built in double by libm, cast to float
  float ns_dim[NNS][3];                // box sides | (radius, length, -)
  int ns_cb[NNS];                      // carries the Rover_1_tf fork callback
  float ft_ocA[2][3], ft_ocB[2][3];
  float getpos_rel[3];
  float step_side[2][3];
};

struct HfParams {
  const float* h;          // [nrow*ncol]
  const float* blockmax;   // 8x8 vertex block maxima
  int ncol, nrow, nbx, nbz;
  float sx, sz, inv_sx, inv_sz, length, width, hmax;
  float pos[3];
};

// Flattened model trees (tfode_formats.hpp), shared by all environments of a GPU.
struct TreeDev { const TfodeTreeNode* nodes; const double* params; int n_nodes; };
struct TreeParams { TreeDev tree[2]; int oblique, degree, interaction_only, n_features; };

struct StepParams {
  EnvRec* rec;
  int* budget;             // [n_envs] remaining substeps of the current tfode_batch_step call
  int n_envs;
  int set_budget;          // >=0: budget[e] = set_budget and done_call cleared (first launch of a call); -1: keep
  TfodeBatchConfig cfg;    // hf_heights unused on the device
  HfParams hf;
  // policy hand-off (TFODE_CTRL_MLP)
  float* policy_in;        // [n_envs,17] scaled observations of parked envs
  int* parked_list;        // env ids parked in this launch
  int* n_parked;
  const float* policy_w;   // fp32 actor blob W1 b1 W2 b2 W3 b3 (TFODE_POLICY_FP32: the actor runs inside step_kernel)
  // transitions
  float* trans; int trans_cap; int* n_trans;
  // contact-capacity audit: [0] sub-steps in which an environment generated more contacts than max_contacts (the excess is
  // dropped), [1] largest number of contacts one environment generated in one sub-step
  unsigned long long* cstats;   // [2]: islands whose Dantzig drive stalled (TFODE_SOLVER_DANTZIG)
  // TFODE_SOLVER_DANTZIG scratch: LCP_ROW_F floats and LCP_ROW_I ints per row slot, MAXROWS row slots per environment (global memory)
  float* lcp_f; int* lcp_i;
  // persistent launch (slice_len > 0): item counter, per-group count of finished slices, item geometry
  int* sched; int* slice_done; int slice_len, n_items, n_groups;
  // external commands
  const float* ext_speed;  // NULL or [n_envs]
  // tire-sphere centres in the wheel frame (Model::ring, padded to float4) in global memory: the lanes of a warp read 32
  // different entries at once, which the constant cache would serialise
  const float4* ring4;
  // the model tables once more in global memory: per-lane indexed reads (body / joint / geom tables) gather through L1
  // instead of being serialised address by address by the constant cache
  const Model* model;
};

}  // namespace tfode
