// C ABI of the batched rover stepping engine (include/tfode_batch.h) and the host-side
// model tables.  No CPU stepping path exists here: every compute entry point launches the
// sm_100a kernels and fails with TFODE_ECUDA when no device is usable.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "tfode_formats.hpp"
#include "tfode_kernels.cuh"
#include "tfode_model_tree.cuh"
#include "tfode_policy_tc.cuh"

using namespace tfode;

static thread_local std::string g_create_error;

struct TfodeBatch {
  TfodeBatchConfig cfg;
  StepParams P;
  cudaStream_t stream = nullptr;
  std::string err;
  int64_t launches = 0;
  // device buffers
  EnvRec* rec = nullptr; int* budget = nullptr;
  float* hf_h = nullptr; float* hf_bmax = nullptr;
  float* policy_in = nullptr; int* parked_list = nullptr; int* n_parked = nullptr;
  float* trans = nullptr; int* n_trans = nullptr; int trans_cap = 0;
  int* sched = nullptr;   // [0] item counter, [1 .. 1 + n_groups) finished slices per group (persistent launches)
  int resident_ctas = 0;  // CTAs of step_kernel that are co-resident on this device
  int persistent = 1;     // TFODE_PERSISTENT=0 turns the persistent launch off (A/B)
  float* lcp_f = nullptr; int* lcp_i = nullptr;   // TFODE_SOLVER_DANTZIG scratch
  unsigned long long* cstats = nullptr;   // contact-capacity audit counters (StepParams::cstats)
  float* ext_speed = nullptr;
  float4* ring4 = nullptr; Model* model_g = nullptr;    // Model::ring as float4 in global memory
  float* weights = nullptr;   // fp32 W1 b1 W2 b2 W3 b3
  PolicyTC tc;                // bf16 operand copies for the tcgen05 path
  bool has_policy = false;
  // model trees (TFODE_CTRL_MODEL_TREE): flattened nodes + parameter pools of tree 1 / tree 2
  TfodeTreeNode* tree_nodes[2] = {nullptr, nullptr}; double* tree_params[2] = {nullptr, nullptr};
  TreeParams TP{};
  bool has_trees = false;
  // staging for the host round trip
  float* d_obs = nullptr; float* d_rew = nullptr; uint8_t* d_done = nullptr; float* d_act = nullptr;
  float* h_pin = nullptr; size_t h_pin_bytes = 0;
  int* d_ids = nullptr; int ids_cap = 0;
  size_t smem_bytes = 0, smem_step = 0;
  int lsu = 2;   // lock-step unroll factor of the step_kernel instantiation this batch launches (4 on the bare plane)
};

// Every entry point that takes a handle runs on the handle's own device, whatever device the calling thread had current
// (a thread may own handles on several GPUs, or another library may have switched devices); the caller's device is restored.
struct DeviceGuard {
  int prev = -1; bool ok = true;
  explicit DeviceGuard(const TfodeBatch* b) {
    if (!b) return;
    if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; ok = false; return; }
    if (prev != b->cfg.device) { if (cudaSetDevice(b->cfg.device) != cudaSuccess) ok = false; } else prev = -1;
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};
#define GUARD(b) DeviceGuard guard_((b)); if ((b) && !guard_.ok) { (b)->err = "cudaSetDevice failed for this handle's device"; return TFODE_ECUDA; }

#define CK(call)                                                                            \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      b->err = std::string(#call) + ": " + cudaGetErrorString(e_);                          \
      return TFODE_ECUDA;                                                                   \
    }                                                                                       \
  } while (0)

// ------------------------------------------------------------------------------------------------
// The rover as tables (restates the constructor of src/rover_1.cc:49-233, ode/wheel.hh:56-74,
// ode/box.hh:54-62, ode/ft_sensor.cc:50-55 and the step boxes of src/rover_training_1.cc:88-98).
// ------------------------------------------------------------------------------------------------
namespace {
// A miniature of the construction calls the reference makes (dBodySetPosition/Rotation, dMassSet*Total,
// dJointSetHingeAnchor/Axis, dJointSetSliderAxis, dGeomSetOffsetWorldPosition), evaluated in double like libode
// and rounded to float once, when the tables are written.
struct DBody { double pos[3], q[4], R[9], mass, I[3]; };

void d_RFromEuler(double* R, double phi, double theta, double psi) {  // dRFromEulerAngles
  double sphi = std::sin(phi), cphi = std::cos(phi), sth = std::sin(theta), cth = std::cos(theta), spsi = std::sin(psi), cpsi = std::cos(psi);
  R[0] = cpsi * cth; R[1] = spsi * cth; R[2] = -sth;
  R[3] = cpsi * sth * sphi - spsi * cphi; R[4] = spsi * sth * sphi + cpsi * cphi; R[5] = cth * sphi;
  R[6] = cpsi * sth * cphi + spsi * sphi; R[7] = spsi * sth * cphi - cpsi * sphi; R[8] = cth * cphi;
}
void d_QfromR(double* q, const double* R) {  // dRtoQ, tr >= 0 branch is the only one the rover's construction reaches
  double tr = R[0] + R[4] + R[8];
  double s = std::sqrt(tr + 1); q[0] = 0.5 * s; s = 0.5 / s;
  q[1] = (R[7] - R[5]) * s; q[2] = (R[2] - R[6]) * s; q[3] = (R[3] - R[1]) * s;
}
void d_RfromQ(double* R, const double* q) {
  double qq1 = 2 * q[1] * q[1], qq2 = 2 * q[2] * q[2], qq3 = 2 * q[3] * q[3];
  R[0] = 1 - qq2 - qq3; R[1] = 2 * (q[1] * q[2] - q[0] * q[3]); R[2] = 2 * (q[1] * q[3] + q[0] * q[2]);
  R[3] = 2 * (q[1] * q[2] + q[0] * q[3]); R[4] = 1 - qq1 - qq3; R[5] = 2 * (q[2] * q[3] - q[0] * q[1]);
  R[6] = 2 * (q[1] * q[3] - q[0] * q[2]); R[7] = 2 * (q[2] * q[3] + q[0] * q[1]); R[8] = 1 - qq1 - qq2;
}
void d_setRotation(DBody& b, const double* Rin) {  // dBodySetRotation: R -> q -> normalise -> R
  d_QfromR(b.q, Rin);
  double l = b.q[0] * b.q[0] + b.q[1] * b.q[1] + b.q[2] * b.q[2] + b.q[3] * b.q[3];
  l = 1.0 / std::sqrt(l);
  for (int k = 0; k < 4; k++) b.q[k] *= l;
  d_RfromQ(b.R, b.q);
}
DBody d_body(double x, double y, double z) {
  DBody b; b.pos[0] = x; b.pos[1] = y; b.pos[2] = z; b.q[0] = 1; b.q[1] = b.q[2] = b.q[3] = 0;
  for (int k = 0; k < 9; k++) b.R[k] = (k % 4 == 0) ? 1.0 : 0.0;
  b.mass = 1; b.I[0] = b.I[1] = b.I[2] = 1; return b;
}
void d_massBox(DBody& b, double m, double lx, double ly, double lz) {  // dMassSetBoxTotal
  b.mass = m; b.I[0] = m / 12.0 * (ly * ly + lz * lz); b.I[1] = m / 12.0 * (lx * lx + lz * lz); b.I[2] = m / 12.0 * (lx * lx + ly * ly);
}
void d_massCylZ(DBody& b, double m, double r, double l) {  // dMassSetCylinderTotal(direction 3)
  double r2 = r * r;
  b.mass = m; b.I[0] = b.I[1] = m * (0.25 * r2 + (1.0 / 12.0) * l * l); b.I[2] = m * 0.5 * r2;
}
void d_posRelPoint(const DBody& b, const double* p, double* out) {  // dBodyGetPosRelPoint
  double d[3] = {p[0] - b.pos[0], p[1] - b.pos[1], p[2] - b.pos[2]};
  tfn::mulRtv(out, b.R, d, 3);
}
void d_qmulInvA(double* r, const double* b, const double* c) {  // b^-1 * c
  r[0] = b[0] * c[0] + b[1] * c[1] + b[2] * c[2] + b[3] * c[3];
  r[1] = b[0] * c[1] - b[1] * c[0] - b[2] * c[3] + b[3] * c[2];
  r[2] = b[0] * c[2] - b[2] * c[0] - b[3] * c[1] + b[1] * c[3];
  r[3] = b[0] * c[3] - b[3] * c[0] - b[1] * c[2] + b[2] * c[1];
}

void build_model(Model& M) {
  std::memset(&M, 0, sizeof(M));
  // parameters of src/rover_1.cc:49-103
  const double wheelbase = 0.58, wheeltrack = 0.61, wheel_r = 0.105, wheel_w = 0.1, belly = 0.33;
  const double front_len = 0.345, front_h = 0.225, front_w = 0.200;
  const double fx = front_len / 2 + -0.063, fy = 0.023, fz = belly + front_h / 2;
  const double rear_len = 0.312, rear_h = 0.177, rear_w = 0.126;
  const double rx = -rear_len / 2 + 0.046, ry = -0.040, rz = belly + rear_h / 2;
  const double sea_elev = belly + 0.12;
  const double boggie_h = sea_elev - belly;
  const double fork_elev = belly - 0.1 / 2;
  const double rim_r = wheel_r - wheel_w / 2;
  const float step_h = (float)(0.105 * 2);  // `float step_height` in src/rover_training_1.cc:84
  DBody B[NB];
  B[0] = d_body(fx, fy, fz); d_massBox(B[0], 5.25, front_len, front_w, front_h);
  B[1] = d_body(0.285, 0, belly + 0.155); d_massBox(B[1], 3, 0.0975, 0.151, 0.065);
  B[2] = d_body(rx, ry, rz); d_massBox(B[2], 3.67, rear_len, rear_w, rear_h);
  B[3] = d_body(-wheelbase / 2 + 0.035, 0, belly + boggie_h / 2); d_massBox(B[3], 1, 0.1, 0.1, boggie_h);
  B[4] = d_body(wheelbase / 2, 0, fork_elev); d_massBox(B[4], 1, 0.05, 0.3, 0.1);
  B[5] = d_body(-wheelbase / 2, 0, fork_elev); d_massBox(B[5], 1, 0.05, 0.3, 0.1);
  double Rw[9]; d_RFromEuler(Rw, M_PI / 2, 0.0, 0.0);  // src/rover_1.cc:214
  for (int i = 0; i < 4; i++) {
    B[6 + i] = d_body((i / 2 ? -1 : 1) * wheelbase / 2, (i % 2 ? -1 : 1) * wheeltrack / 2, wheel_r);
    d_massCylZ(B[6 + i], 1.4 + 1, rim_r, wheel_w);
    d_setRotation(B[6 + i], Rw);
  }
  B[10] = d_body(1, 0, step_h / 2); d_massBox(B[10], 1, 1, 3, step_h);   // yaw is per episode (reset_env)
  B[11] = d_body(2, 0, step_h / 2); d_massBox(B[11], 1, 2, 3, step_h);
  for (int b = 0; b < NB; b++) {
    M.mass[b] = (float)B[b].mass; M.invMass[b] = (float)(1.0 / B[b].mass);
    for (int k = 0; k < 3; k++) { M.Ib[b][k] = (float)B[b].I[k]; M.invIb[b][k] = (float)(1.0 / B[b].I[k]); M.pos0[b][k] = (float)B[b].pos[k]; }
    for (int k = 0; k < 4; k++) M.quat0[b][k] = (float)B[b].q[k];
  }
  M.step_side[0][0] = 1; M.step_side[0][1] = 3; M.step_side[0][2] = step_h;
  M.step_side[1][0] = 2; M.step_side[1][1] = 3; M.step_side[1][2] = step_h;

  auto set_axes = [&](int j, int b1, int b2, double x, double y, double z) {  // setAxes (joint.cpp)
    double a[3] = {x, y, z}; double l = tfn::dot3(a, a); l = 1.0 / std::sqrt(l); a[0] *= l; a[1] *= l; a[2] *= l;
    double o[3];
    tfn::mulRtv(o, B[b1].R, a, 3); for (int k = 0; k < 3; k++) M.axis1[j][k] = (float)o[k];
    if (b2 < NB) { tfn::mulRtv(o, B[b2].R, a, 3); for (int k = 0; k < 3; k++) M.axis2[j][k] = (float)o[k]; }
    else for (int k = 0; k < 3; k++) M.axis2[j][k] = (float)a[k];
  };
  auto set_qrel = [&](int j, int b1, int b2) {  // computeInitialRelativeRotation
    double q[4];
    if (b2 < NB) d_qmulInvA(q, B[b1].q, B[b2].q); else { q[0] = B[b1].q[0]; for (int k = 1; k < 4; k++) q[k] = -B[b1].q[k]; }
    for (int k = 0; k < 4; k++) M.qrel[j][k] = (float)q[k];
  };
  auto hinge = [&](int j, int b1, int b2, const double* anchor, double ax, double ay, double az) {
    M.jb1[j] = b1; M.jb2[j] = b2; M.jslider[j] = 0;
    double o[3];
    d_posRelPoint(B[b1], anchor, o); for (int k = 0; k < 3; k++) M.anchor1[j][k] = (float)o[k];
    d_posRelPoint(B[b2], anchor, o); for (int k = 0; k < 3; k++) M.anchor2[j][k] = (float)o[k];
    set_axes(j, b1, b2, ax, ay, az);
    set_qrel(j, b1, b2);
    M.lostop[j] = -INFINITY; M.histop[j] = INFINITY;
  };
  auto slider = [&](int j, int b1, int b2, double ax, double ay, double az) {
    M.jb1[j] = b1; M.jb2[j] = b2; M.jslider[j] = 1;
    set_axes(j, b1, b2, ax, ay, az);
    if (b2 < NB) {  // computeOffset
      double c[3] = {B[b1].pos[0] - B[b2].pos[0], B[b1].pos[1] - B[b2].pos[1], B[b1].pos[2] - B[b2].pos[2]}, o[3];
      tfn::mulRtv(o, B[b2].R, c, 3);
      for (int k = 0; k < 3; k++) M.offset[j][k] = (float)o[k];
    } else for (int k = 0; k < 3; k++) M.offset[j][k] = (float)B[b1].pos[k];
    set_qrel(j, b1, b2);
    M.lostop[j] = 0; M.histop[j] = 0;
  };
  slider(0, 1, 0, 0, 1, 0);                                                    // battery clamp, src/rover_1.cc:128-132
  { const double a[3] = {0, 0, (fz + rz) / 2}; hinge(1, 2, 0, a, 0, 0, 1); }   // centre hinge servo, :181-189
  { const double a[3] = {-wheelbase / 2, 0, sea_elev}; hinge(2, 3, 2, a, 1, 0, 0);  // boggie hinge, :194-200
    M.lostop[2] = (float)(-45 * 0.017453292519943295); M.histop[2] = (float)(45 * 0.017453292519943295); }
  for (int i = 0; i < 4; i++) hinge(3 + i, B[6 + i].pos[0] > 0 ? 4 : 5, 6 + i, B[6 + i].pos, 0, -1, 0);  // wheel hinges, :224-229
  slider(7, 10, 15, 0, 0, 1); slider(8, 11, 15, 0, 0, 1);                      // Object::fix, ode/object.cc:155-179

  // geoms (ode/wheel.hh:63-73, src/rover_1.cc:116-176)
  for (int k = 0; k < 50; k++) { M.ring[k][0] = (float)(rim_r * std::cos(k * 2 * M_PI / 50)); M.ring[k][1] = (float)(rim_r * std::sin(k * 2 * M_PI / 50)); M.ring[k][2] = 0; }
  M.sphere_r = (float)(wheel_w / 2);
  const double chassis[3][3] = {{front_len, front_w, front_h}, {0.0975, 0.151, 0.065}, {rear_len, rear_w, rear_h}};
  for (int g = 0; g < 3; g++) { M.ns_type[g] = 1; M.ns_body[g] = g; for (int k = 0; k < 3; k++) M.ns_dim[g][k] = (float)chassis[g][k]; }
  for (int g = 0; g < 4; g++) {  // motor bulks on the forks
    const int body = 4 + g / 2; const double sx = g / 2 == 0 ? 1 : -1;
    M.ns_type[3 + g] = 2; M.ns_body[3 + g] = body; M.ns_rot[3 + g] = 1; M.ns_cb[3 + g] = 1;
    M.ns_dim[3 + g][0] = (float)0.025; M.ns_dim[3 + g][1] = (float)0.1;
    const double ap[3] = {sx * wheelbase / 2, (g % 2 == 0 ? 1 : -1) * (wheeltrack - wheel_w - 0.1) / 2, wheel_r};
    double o[3]; d_posRelPoint(B[body], ap, o);
    for (int k = 0; k < 3; k++) M.ns_off[3 + g][k] = (float)o[k];
    for (int k = 0; k < 9; k++) M.ns_offR[3 + g][k] = (float)Rw[k];
  }
  for (int g = 0; g < 4; g++) { M.ns_type[7 + g] = 2; M.ns_body[7 + g] = 6 + g; M.ns_dim[7 + g][0] = (float)rim_r; M.ns_dim[7 + g][1] = (float)wheel_w; }
  // FT sensors (src/rover_1.cc:205-206, ode/ft_sensor.cc:50-55): A = main / boggie, B = forks
  for (int s = 0; s < 2; s++) {
    const double c[3] = {(s == 0 ? 1 : -1) * wheelbase / 2, 0, belly};
    const int A = s == 0 ? 0 : 3, Bb = s == 0 ? 4 : 5;
    double o[3];
    d_posRelPoint(B[A], c, o); for (int k = 0; k < 3; k++) M.ft_ocA[s][k] = (float)o[k];
    d_posRelPoint(B[Bb], c, o); for (int k = 0; k < 3; k++) M.ft_ocB[s][k] = (float)o[k];
  }
  M.getpos_rel[0] = (float)(-fx); M.getpos_rel[1] = (float)(-fy); M.getpos_rel[2] = (float)(-fz + sea_elev);
  M.robot_max_speed = 7.4351 / wheel_r;  // Q4: divide for wheel 0, multiply for 1..3
  for (int i = 1; i < 4; i++) M.robot_max_speed = std::fmin(M.robot_max_speed, 7.4351 * wheel_r);
  // LP_second_order_bilinear(Te = 0.001f, w0 = 4 pi, Q = 0.5f), Filters/cpp/filters.cc:148-158
  { const float Te = 0.001f; const double w0 = 4 * M_PI; const float Q = 0.5f;
    double w2T2 = std::pow(w0 * Te, 2), den = 2 * w0 * Te + 4 * Q + Q * w2T2;
    M.filt_c[0] = 2 * Q * (4 - w2T2) / den; M.filt_c[1] = (2 * w0 * Te - 4 * Q - Q * w2T2) / den; M.filt_c[2] = Q * w2T2 / den; }
}

int launch_blocks(int n) { return (n + WARPS_PER_CTA - 1) / WARPS_PER_CTA; }
}  // namespace

extern "C" {

int tfode_batch_config_default(TfodeBatchConfig* c, int preset) {
  if (!c) return TFODE_EINVAL;
  std::memset(c, 0, sizeof(*c));
  c->abi_version = TFODE_ABI_VERSION; c->n_envs = 1; c->scene = TFODE_SCENE_STEP_BOXES; c->controller = TFODE_CTRL_EXTERNAL;
  c->policy_precision = TFODE_POLICY_FP32; c->auto_reset = 1;
  c->mu = 0.5f; c->dt = 0.001f; c->erp = 0.2f; c->cfm = 1e-10f; c->pgs_iters = 20; c->sor_w = 1.3f;
  c->crawling_mode = 1; c->ic_period = 0; c->ic_initially_active = 0;
  c->speedf = 0.04f; c->term = 0.5f; c->ic_start = -1.0f; c->timeout = 60; c->x_goal = 2.0f; c->y_max = 0.5f;
  c->x_goal_abs = 1; c->check_upside_down = 1; c->terminate = 0;
  c->crawler_amp = 30; c->crawler_dt_torque = 2; c->crawler_rate = 10;
  c->hf_length = 3; c->hf_width = 3;
  c->seed = 20260411ull; c->max_contacts = 24; c->solver = TFODE_SOLVER_QUICKSTEP;
  switch (preset) {
    case 0: break;
    case 1: c->controller = TFODE_CTRL_RANDOM; c->ic_period = 0.5f; c->ic_start = 1.0f; c->x_goal = 1.0f; c->y_max = 0.6f; c->terminate = 1; break;
    case 2: c->controller = TFODE_CTRL_MLP; c->ic_period = 0.5f; c->ic_start = 1.0f; c->ic_start_jitter = 0.25f; c->exploration = 1; c->fork_callback = 1;
            c->x_goal = 1.5f; c->y_max = 0.6f; c->terminate = 1; c->step_yaw_jitter_deg = 5.0f; break;
    case 3: c->scene = TFODE_SCENE_PLANE; c->controller = TFODE_CTRL_CRAWLER; c->mu = 0.6f; c->ic_initially_active = 1; c->term = 0;
            c->x_goal = 10; c->y_max = 0; c->x_goal_abs = 0; c->check_upside_down = 0; c->terminate = 1; break;
    default: return TFODE_EINVAL;
  }
  return TFODE_OK;
}

const char* tfode_last_error(const TfodeBatch* b) { return b ? b->err.c_str() : g_create_error.c_str(); }
int tfode_sizeof_config() { return (int)sizeof(TfodeBatchConfig); }
int tfode_sizeof_state() { return (int)sizeof(TfodeEnvState); }
int tfode_sizeof_record() { return (int)sizeof(EnvRec); }

constexpr int TFODE_SLICE_LEN = 25;   // sub-steps per work item of a persistent launch

static int launch_reset(TfodeBatch* b, const int* d_ids, int n, int first) {
  reset_kernel<<<launch_blocks(n), WARPS_PER_CTA * 32, b->smem_bytes, b->stream>>>(b->P, d_ids, n, first);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_create(const TfodeBatchConfig* cfg, TfodeBatch** out) {
  if (!cfg || !out) { g_create_error = "null argument"; return TFODE_EINVAL; }
  *out = nullptr;
  if (cfg->abi_version != TFODE_ABI_VERSION) { g_create_error = "abi_version mismatch"; return TFODE_EINVAL; }
  if (cfg->n_envs <= 0) { g_create_error = "n_envs must be positive"; return TFODE_EINVAL; }
  if (cfg->scene < 0 || cfg->scene > 2 || cfg->controller < 0 || cfg->controller > TFODE_CTRL_MODEL_TREE) { g_create_error = "bad scene/controller"; return TFODE_EINVAL; }
  if (cfg->scene == TFODE_SCENE_HEIGHTFIELD && (!cfg->hf_heights || cfg->hf_ncol < 2 || cfg->hf_nrow < 2)) { g_create_error = "heightfield scene without heights"; return TFODE_EINVAL; }
  if (cfg->max_contacts < 1 || cfg->max_contacts > MAXC) { g_create_error = "max_contacts must be in 1..24"; return TFODE_EINVAL; }
  if (cfg->solver != TFODE_SOLVER_QUICKSTEP && cfg->solver != TFODE_SOLVER_DANTZIG) { g_create_error = "bad solver"; return TFODE_EINVAL; }
  if (!(cfg->dt > 0) || cfg->pgs_iters < 1 || cfg->pgs_iters > 1000) { g_create_error = "bad dt / pgs_iters"; return TFODE_EINVAL; }
  if (std::fabs(cfg->step_yaw_deg) + std::fabs(cfg->step_yaw_jitter_deg) >= 119.0f) { g_create_error = "step yaw must stay below 119 degrees"; return TFODE_EINVAL; }
  TfodeBatch* b = new TfodeBatch();
  b->cfg = *cfg;
  auto fail = [&](int code) { g_create_error = b->err; tfode_batch_destroy(b); return code; };
#define CKC(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { b->err = std::string(#call) + ": " + cudaGetErrorString(e_); return fail(TFODE_ECUDA); } } while (0)
  CKC(cudaSetDevice(cfg->device));
  cudaDeviceProp prop; CKC(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major < 10) { b->err = "this engine is built for sm_100a (B200) only"; return fail(TFODE_ECUDA); }
  const int n = cfg->n_envs;
  Model M; build_model(M);
  CKC(cudaMemcpyToSymbol(g_model, &M, sizeof(M)));
  {
    float4 r4[50];
    for (int k = 0; k < 50; k++) r4[k] = make_float4(M.ring[k][0], M.ring[k][1], M.ring[k][2], 0.0f);
    CKC(cudaMalloc(&b->model_g, sizeof(Model)));
    CKC(cudaMemcpy(b->model_g, &M, sizeof(Model), cudaMemcpyHostToDevice));
    CKC(cudaMalloc(&b->ring4, sizeof(r4)));
    CKC(cudaMemcpy(b->ring4, r4, sizeof(r4), cudaMemcpyHostToDevice));
  }
  CKC(cudaMalloc(&b->rec, sizeof(EnvRec) * (size_t)n));
  CKC(cudaMalloc(&b->budget, sizeof(int) * (size_t)n));
  CKC(cudaMemset(b->budget, 0, sizeof(int) * (size_t)n));
  CKC(cudaMalloc(&b->policy_in, sizeof(float) * 17 * (size_t)n));
  CKC(cudaMalloc(&b->parked_list, sizeof(int) * (size_t)n));
  CKC(cudaMalloc(&b->n_parked, sizeof(int) * 4));
  CKC(cudaMemset(b->n_parked, 0, sizeof(int) * 4));
  b->n_trans = b->n_parked + 1;
  CKC(cudaMalloc(&b->cstats, sizeof(unsigned long long) * 4));
  CKC(cudaMemset(b->cstats, 0, sizeof(unsigned long long) * 4));
  if (cfg->solver == TFODE_SOLVER_DANTZIG) {   // dense-LCP scratch: 211 KB per environment
    const size_t nf = (size_t)n * MAXROWS * LCP_ROW_F, ni = (size_t)n * MAXROWS * LCP_ROW_I;
    if (cudaMalloc(&b->lcp_f, nf * sizeof(float)) != cudaSuccess || cudaMalloc(&b->lcp_i, ni * sizeof(int)) != cudaSuccess) {
      cudaGetLastError(); b->err = "TFODE_SOLVER_DANTZIG needs " + std::to_string((nf * 4 + ni * 4) >> 20) + " MiB of device scratch for " + std::to_string(n) + " environments"; return fail(TFODE_ENOMEM);
    }
  }
  b->trans_cap = cfg->transition_capacity > 0 ? cfg->transition_capacity : 4 * n;
  CKC(cudaMalloc(&b->trans, sizeof(float) * TFODE_TRANSITION_WIDTH * (size_t)b->trans_cap));
  CKC(cudaMalloc(&b->ext_speed, sizeof(float) * (size_t)n));
  CKC(cudaMalloc(&b->d_obs, sizeof(float) * 17 * (size_t)n));
  CKC(cudaMalloc(&b->d_rew, sizeof(float) * (size_t)n));
  CKC(cudaMalloc(&b->d_done, (size_t)n));
  CKC(cudaMalloc(&b->d_act, sizeof(float) * 2 * (size_t)n));
  b->h_pin_bytes = sizeof(float) * (17 + 1 + 2) * (size_t)n + (size_t)n + 64;
  CKC(cudaMallocHost(&b->h_pin, b->h_pin_bytes));
  StepParams& P = b->P;
  std::memset(&P, 0, sizeof(P));
  P.rec = b->rec; P.budget = b->budget; P.n_envs = n; P.set_budget = -1; P.cfg = *cfg; P.cfg.hf_heights = nullptr;
  P.policy_in = b->policy_in; P.parked_list = b->parked_list; P.n_parked = b->n_parked;
  P.trans = b->trans; P.trans_cap = b->trans_cap; P.n_trans = b->n_trans; P.cstats = b->cstats; P.policy_w = nullptr; P.lcp_f = b->lcp_f; P.lcp_i = b->lcp_i; P.ext_speed = nullptr; P.ring4 = b->ring4; P.model = b->model_g;
  if (cfg->scene == TFODE_SCENE_HEIGHTFIELD) {
    const int nc = cfg->hf_ncol, nr = cfg->hf_nrow;
    const int nbx = (nc + 7) / 8, nbz = (nr + 7) / 8;
    std::vector<float> bm((size_t)nbx * nbz, -1e30f);
    float hmax = -1e30f;
    for (int iz = 0; iz < nr; iz++) for (int ix = 0; ix < nc; ix++) {
      float v = cfg->hf_heights[(size_t)iz * nc + ix];
      if (!(v == v)) { b->err = "NaN in heightfield"; return fail(TFODE_EINVAL); }
      hmax = std::fmax(hmax, v);
      float& m = bm[(size_t)(iz / 8) * nbx + ix / 8]; m = std::fmax(m, v);
    }
    CKC(cudaMalloc(&b->hf_h, sizeof(float) * (size_t)nc * nr));
    CKC(cudaMemcpy(b->hf_h, cfg->hf_heights, sizeof(float) * (size_t)nc * nr, cudaMemcpyHostToDevice));
    CKC(cudaMalloc(&b->hf_bmax, sizeof(float) * bm.size()));
    CKC(cudaMemcpy(b->hf_bmax, bm.data(), sizeof(float) * bm.size(), cudaMemcpyHostToDevice));
    HfParams& H = P.hf;
    H.h = b->hf_h; H.blockmax = b->hf_bmax; H.ncol = nc; H.nrow = nr; H.nbx = nbx; H.nbz = nbz;
    H.sx = (float)((double)cfg->hf_length / (nc - 1)); H.sz = (float)((double)cfg->hf_width / (nr - 1));
    H.inv_sx = (float)((nc - 1) / (double)cfg->hf_length); H.inv_sz = (float)((nr - 1) / (double)cfg->hf_width);
    H.length = cfg->hf_length; H.width = cfg->hf_width; H.hmax = hmax;
    for (int k = 0; k < 3; k++) H.pos[k] = cfg->hf_pos[k];
  }
  b->cfg.hf_heights = nullptr;
  b->smem_bytes = sizeof(WarpSm) * WARPS_PER_CTA;
  b->smem_step = sizeof(WarpSm) * ENVS_PER_CTA;
  {
    const int groups = (n + ENVS_PER_CTA - 1) / ENVS_PER_CTA;
    CKC(cudaMalloc(&b->sched, sizeof(int) * (size_t)(groups + 1)));
    if (const char* o = std::getenv("TFODE_PERSISTENT")) b->persistent = std::atoi(o);
  }
  b->lsu = cfg->scene == TFODE_SCENE_PLANE ? 4 : 2;
  if (const char* o = std::getenv("TFODE_LSU")) b->lsu = std::atoi(o) == 4 ? 4 : 2;   // developer override for A/B runs
  CKC(cudaFuncSetAttribute(step_kernel<TFODE_SOLVER_QUICKSTEP, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->smem_step));
  CKC(cudaFuncSetAttribute(step_kernel<TFODE_SOLVER_QUICKSTEP, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->smem_step));
  CKC(cudaFuncSetAttribute(step_kernel<TFODE_SOLVER_DANTZIG, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->smem_step));
  CKC(cudaFuncSetAttribute(reset_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)b->smem_bytes));
  {
    int per_sm = 0;
    if (b->lsu == 4) CKC(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, step_kernel<TFODE_SOLVER_QUICKSTEP, 4>, ENVS_PER_CTA * 32, b->smem_step));
    else CKC(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, step_kernel<TFODE_SOLVER_QUICKSTEP, 2>, ENVS_PER_CTA * 32, b->smem_step));
    b->resident_ctas = cfg->solver == TFODE_SOLVER_QUICKSTEP ? per_sm * prop.multiProcessorCount : 0;   // the dense-solver path keeps the plain launch
  }
  if (launch_reset(b, nullptr, n, 1) != TFODE_OK) return fail(TFODE_ECUDA);
  CKC(cudaStreamSynchronize(b->stream));
  *out = b;
  return TFODE_OK;
#undef CKC
}

void tfode_batch_destroy(TfodeBatch* b) {
  if (!b) return;
  DeviceGuard guard_(b);
  for (int t = 0; t < 2; t++) { cudaFree(b->tree_nodes[t]); cudaFree(b->tree_params[t]); }
  cudaFree(b->rec); cudaFree(b->budget); cudaFree(b->hf_h); cudaFree(b->hf_bmax); cudaFree(b->policy_in); cudaFree(b->parked_list);
  cudaFree(b->sched); cudaFree(b->n_parked); cudaFree(b->cstats); cudaFree(b->lcp_f); cudaFree(b->lcp_i); cudaFree(b->trans); cudaFree(b->ext_speed); cudaFree(b->ring4); cudaFree(b->model_g); cudaFree(b->weights); cudaFree(b->d_obs); cudaFree(b->d_rew);
  cudaFree(b->d_done); cudaFree(b->d_act); cudaFree(b->d_ids);
  policy_tc_free(b->tc);
  if (b->h_pin) cudaFreeHost(b->h_pin);
  delete b;
}

int tfode_batch_set_stream(TfodeBatch* b, void* stream) { if (!b) return TFODE_EINVAL; b->stream = (cudaStream_t)stream; return TFODE_OK; }
int tfode_batch_n_envs(const TfodeBatch* b) { return b ? b->cfg.n_envs : 0; }
int64_t tfode_batch_launch_count(const TfodeBatch* b) { return b ? b->launches : 0; }
int tfode_batch_sync(TfodeBatch* b) { if (!b) return TFODE_EINVAL; GUARD(b); CK(cudaStreamSynchronize(b->stream)); return TFODE_OK; }

int tfode_batch_reset(TfodeBatch* b, const int32_t* env_ids, int n) {
  if (!b) return TFODE_EINVAL;
  GUARD(b);
  if (!env_ids) return launch_reset(b, nullptr, b->cfg.n_envs, 0);
  if (n <= 0) return TFODE_OK;
  for (int i = 0; i < n; i++) if (env_ids[i] < 0 || env_ids[i] >= b->cfg.n_envs) { b->err = "env id out of range"; return TFODE_EINVAL; }
  if (n > b->ids_cap) { cudaFree(b->d_ids); b->d_ids = nullptr; CK(cudaMalloc(&b->d_ids, sizeof(int) * (size_t)n)); b->ids_cap = n; }
  CK(cudaMemcpyAsync(b->d_ids, env_ids, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return launch_reset(b, b->d_ids, n, 0);
}

int tfode_batch_set_actions(TfodeBatch* b, const float* actions_dev) {
  if (!b || !actions_dev) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs;
  set_actions_kernel<<<(n + 127) / 128, 128, 0, b->stream>>>(b->rec, n, actions_dev, b->cfg.controller == TFODE_CTRL_CRAWLER ? 1 : 0);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_set_speed(TfodeBatch* b, const float* speed_dev) {
  if (!b) return TFODE_EINVAL;
  GUARD(b);
  if (!speed_dev) { b->P.ext_speed = nullptr; return TFODE_OK; }
  CK(cudaMemcpyAsync(b->ext_speed, speed_dev, sizeof(float) * (size_t)b->cfg.n_envs, cudaMemcpyDeviceToDevice, b->stream));
  b->P.ext_speed = b->ext_speed;
  return TFODE_OK;
}

int tfode_batch_load_policy(TfodeBatch* b, const float* w, size_t n_floats) {
  if (!b || !w) return TFODE_EINVAL;
  GUARD(b);
  const size_t H = TFODE_MLP_HIDDEN, need = 17 * H + H + H * H + H + H * 2 + 2;
  if (n_floats != need) { b->err = "policy blob must hold 272898 floats (W1 b1 W2 b2 W3 b3)"; return TFODE_EINVAL; }
  if (!b->weights) CK(cudaMalloc(&b->weights, sizeof(float) * need));
  CK(cudaMemcpyAsync(b->weights, w, sizeof(float) * need, cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  int rc = policy_tc_upload(b->tc, w, b->stream);
  if (rc != 0) { b->err = "policy_tc_upload failed"; return TFODE_ECUDA; }
  b->P.policy_w = b->weights;
  b->has_policy = true;
  return TFODE_OK;
}

static int run_policy(TfodeBatch* b, const float* in, const int* list, const int* n_list, int m_direct, int max_rows, float* act_direct) {
  const size_t H = TFODE_MLP_HIDDEN;
  const float* W1 = b->weights; const float* b1 = W1 + 17 * H; const float* W2 = b1 + H; const float* b2 = W2 + H * H; const float* W3 = b2 + H; const float* b3 = W3 + H * 2;
  if (b->cfg.policy_precision == TFODE_POLICY_TC) {
    const cudaError_t e = policy_tc_launch(b->tc, in, list, n_list, m_direct, max_rows, b1, b2, W3, b3, b->rec, act_direct, b->stream);
    b->launches += 1;
    if (e != cudaSuccess) { b->err = std::string("policy_tc_launch: ") + cudaGetErrorString(e); return TFODE_ECUDA; }
    return TFODE_OK;
  }
  const int blocks = (max_rows + POL_ROWS - 1) / POL_ROWS;
  policy_fp32_kernel<<<blocks, 256, 0, b->stream>>>(in, list, n_list, m_direct, W1, b1, W2, b2, W3, b3, b->rec, act_direct);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_step(TfodeBatch* b, int n_substeps) {
  if (!b || n_substeps < 0) return TFODE_EINVAL;
  GUARD(b);
  if (n_substeps == 0) return TFODE_OK;
  const int n = b->cfg.n_envs;
  const bool mlp = b->cfg.controller == TFODE_CTRL_MLP, mtree = b->cfg.controller == TFODE_CTRL_MODEL_TREE;
  if (mlp && !b->has_policy) { b->err = "TFODE_CTRL_MLP needs tfode_batch_load_policy first"; return TFODE_ESTATE; }
  if (mtree && !b->has_trees) { b->err = "TFODE_CTRL_MODEL_TREE needs tfode_batch_load_model_trees first"; return TFODE_ESTATE; }
  // TFODE_POLICY_FP32: the actor runs inside step_kernel at each environment's own IC tick (actor_fused), so a call is ONE launch.
  // TFODE_POLICY_TC / model trees: environments park at their IC tick, the batched policy kernel serves the parked list, the next
  // launch resumes them.
  const bool fused_actor = mlp && b->cfg.policy_precision == TFODE_POLICY_FP32;
  int rounds = 1;
  if ((mlp && !fused_actor) || mtree) {
    int period_steps = b->cfg.ic_period > 0 ? (int)std::floor(b->cfg.ic_period / b->cfg.dt) : 1;
    if (period_steps < 1) period_steps = 1;
    rounds = n_substeps / period_steps + 2;
  }
  for (int r = 0; r < rounds; r++) {
    StepParams P = b->P;
    P.set_budget = (r == 0) ? n_substeps : -1;
    const int groups = (n + ENVS_PER_CTA - 1) / ENVS_PER_CTA;
    dim3 grid(groups), block(ENVS_PER_CTA * 32);
    // Persistent launch: when the groups do not fit one wave of resident CTAs and nothing parks, one wave of CTAs pulls (group, 25-sub-step
    // slice) items instead, so the last wave of a call is as full as the others (config 2: 512 groups on 296 slots = 1.73 waves, not 2).
    P.slice_len = 0;
    if (b->persistent && rounds == 1 && b->resident_ctas > 0 && groups > b->resident_ctas && groups < 16 * b->resident_ctas && n_substeps > TFODE_SLICE_LEN) {   // beyond 16 waves the tail is < 3 % and not worth the slices' record traffic
      P.slice_len = TFODE_SLICE_LEN; P.n_groups = groups; P.n_items = groups * ((n_substeps + TFODE_SLICE_LEN - 1) / TFODE_SLICE_LEN);
      P.sched = b->sched; P.slice_done = b->sched + 1;
      CK(cudaMemsetAsync(b->sched, 0, sizeof(int) * (size_t)(groups + 1), b->stream));
      grid = dim3(b->resident_ctas);
    }
    // Instantiation by measurement (profiles/r02_ab_table.md): the lock-step loop unrolled by 4 where sweeps are short (bare plane: ~4 contacts),
    // by 2 where they are long (step boxes, heightfields) -- the loop's instruction footprint against its per-iteration overhead.
    if (b->cfg.solver == TFODE_SOLVER_DANTZIG) step_kernel<TFODE_SOLVER_DANTZIG, 2><<<grid, block, b->smem_step, b->stream>>>(P);
    else if (b->lsu == 4) step_kernel<TFODE_SOLVER_QUICKSTEP, 4><<<grid, block, b->smem_step, b->stream>>>(P);
    else step_kernel<TFODE_SOLVER_QUICKSTEP, 2><<<grid, block, b->smem_step, b->stream>>>(P);
    b->launches++;
    CK(cudaGetLastError());
    if (mlp && !fused_actor) {
      int rc = run_policy(b, b->policy_in, b->parked_list, b->n_parked, 0, n, nullptr);
      if (rc != TFODE_OK) return rc;
      zero_int_kernel<<<1, 1, 0, b->stream>>>(b->n_parked);
      b->launches++;
    }
    if (mtree) {
      const int blocks = std::min((n + 127) / 128, 148 * 8);
      model_tree_kernel<<<blocks, 128, 0, b->stream>>>(b->TP, b->policy_in, b->parked_list, b->n_parked, b->rec, nullptr, 0, nullptr, nullptr);
      CK(cudaGetLastError());
      zero_int_kernel<<<1, 1, 0, b->stream>>>(b->n_parked);
      b->launches += 2;
    }
  }
  return TFODE_OK;
}

int tfode_policy_infer(TfodeBatch* b, const float* obs_scaled_dev, int m, float* act_dev) {
  if (!b || !obs_scaled_dev || !act_dev || m < 0) return TFODE_EINVAL;
  GUARD(b);
  if (!b->has_policy) { b->err = "no policy loaded"; return TFODE_ESTATE; }
  if (m == 0) return TFODE_OK;
  return run_policy(b, obs_scaled_dev, nullptr, nullptr, m, m, act_dev);
}

int tfode_batch_observe(TfodeBatch* b, float* obs17, float* reward, uint8_t* done, int32_t* contact_count) {
  if (!b) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs;
  observe_kernel<<<(n + 127) / 128, 128, 0, b->stream>>>(b->rec, n, obs17, reward, done, contact_count);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_pack_rollout(TfodeBatch* b, float* rows_dev) {
  if (!b || !rows_dev) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs;
  pack_rollout_kernel<<<(n + 127) / 128, 128, 0, b->stream>>>(b->rec, n, b->cfg.env_id_offset, rows_dev);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_get_bodies(TfodeBatch* b, float* out_dev) {
  if (!b || !out_dev) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs * NB;
  bodies_kernel<<<(n + 127) / 128, 128, 0, b->stream>>>(b->rec, b->cfg.n_envs, out_dev);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_drain_transitions(TfodeBatch* b, float* out_dev, int max_rows, int* n_rows_host, int* n_dropped_host) {
  if (!b || !n_rows_host || max_rows < 0) return TFODE_EINVAL;
  GUARD(b);
  int cnt = 0;
  CK(cudaMemcpyAsync(&cnt, b->n_trans, sizeof(int), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  const int avail = cnt < b->trans_cap ? cnt : b->trans_cap;   // rows held by the ring
  if (n_dropped_host) *n_dropped_host = cnt - avail;             // produced beyond transition_capacity since the last drain: lost
  if (avail > max_rows || (avail > 0 && !out_dev)) {             // nothing is consumed: the caller retries with room for *n_rows_host rows
    *n_rows_host = avail;
    b->err = "tfode_batch_drain_transitions: " + std::to_string(avail) + " rows pending, buffer holds " + std::to_string(max_rows);
    return TFODE_EINVAL;
  }
  if (avail > 0) CK(cudaMemcpyAsync(out_dev, b->trans, sizeof(float) * TFODE_TRANSITION_WIDTH * (size_t)avail, cudaMemcpyDeviceToDevice, b->stream));
  CK(cudaMemsetAsync(b->n_trans, 0, sizeof(int), b->stream));
  *n_rows_host = avail;   // rows written to out_dev
  return TFODE_OK;
}

int tfode_batch_lcp_failures(TfodeBatch* b, int64_t* stalled_islands) {
  if (!b || !stalled_islands) return TFODE_EINVAL;
  GUARD(b);
  unsigned long long h = 0;
  CK(cudaMemcpyAsync(&h, b->cstats + 2, sizeof(h), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  *stalled_islands = (int64_t)h;
  return TFODE_OK;
}

int tfode_batch_contact_stats(TfodeBatch* b, int64_t* over_capacity_steps, int64_t* max_generated, int reset) {
  if (!b) return TFODE_EINVAL;
  GUARD(b);
  unsigned long long h[2] = {0, 0};
  CK(cudaMemcpyAsync(h, b->cstats, sizeof(h), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  if (over_capacity_steps) *over_capacity_steps = (int64_t)h[0];
  if (max_generated) *max_generated = (int64_t)h[1];
  if (reset) CK(cudaMemsetAsync(b->cstats, 0, sizeof(h), b->stream));
  return TFODE_OK;
}

// ------------------------------------------------------------------------------------------------
// tfode_batch_step_dantzig_static: the dense-solver step with its LCP scratch in a file-scope device array instead of
// cudaMalloc'd memory, for callers that may not allocate (include/tfode_batch.h).  No counterpart in the reference.
// Every GPU that loads this library holds the array (26.2 MiB) until the process exits; every call on a device shares it.
// ------------------------------------------------------------------------------------------------
namespace {
constexpr size_t NOALLOC_F_BYTES = (size_t)TFODE_NOALLOC_MAX_ENVS * MAXROWS * LCP_ROW_F * sizeof(float);
constexpr size_t NOALLOC_I_BYTES = (size_t)TFODE_NOALLOC_MAX_ENVS * MAXROWS * LCP_ROW_I * sizeof(int);
static_assert(NOALLOC_F_BYTES % 256 == 0, "the int part of the static scratch must stay 256-byte aligned");
constexpr int NOALLOC_MAX_DEVICES = 64;
}
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_F_BYTES + NOALLOC_I_BYTES];
// In host code `g_noalloc_scratch` is only a handle for the runtime, not a device address: the address a kernel can use is looked
// up once per device (each device has its own copy of the array) with that device current.
static std::mutex g_noalloc_mu;
static void* g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

static int noalloc_scratch(TfodeBatch* b, unsigned char** out) {   // the caller holds GUARD(b): b's device is current
  const int dev = b->cfg.device;
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) { b->err = "tfode_batch_step_dantzig_static: device index beyond " + std::to_string(NOALLOC_MAX_DEVICES - 1); return TFODE_EINVAL; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    void* p = nullptr;
    CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    cudaPointerAttributes at{};
    CK(cudaPointerGetAttributes(&at, p));
    if (!p || at.type != cudaMemoryTypeDevice || at.device != dev) { b->err = "tfode_batch_step_dantzig_static: the static scratch did not resolve to memory of this handle's device"; return TFODE_ECUDA; }
    g_noalloc_ptr[dev] = p;
  }
  *out = (unsigned char*)g_noalloc_ptr[dev];
  return TFODE_OK;
}

int tfode_batch_step_dantzig_static(TfodeBatch* b, const float* actions_dev, int n_substeps, float* bodies_out_dev) {
  if (!b || n_substeps < 0) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs;
  if (n > TFODE_NOALLOC_MAX_ENVS) {
    b->err = "tfode_batch_step_dantzig_static: " + std::to_string(n) + " environments, the static scratch holds " + std::to_string(TFODE_NOALLOC_MAX_ENVS) +
             " (use a TFODE_SOLVER_DANTZIG handle and tfode_batch_step beyond that)";
    return TFODE_EINVAL;
  }
  const bool mlp = b->cfg.controller == TFODE_CTRL_MLP;
  if (b->cfg.controller == TFODE_CTRL_MODEL_TREE || (mlp && b->cfg.policy_precision != TFODE_POLICY_FP32)) {
    b->err = "tfode_batch_step_dantzig_static: controllers that park environments between launches (model trees, TFODE_POLICY_TC) are not served here";
    return TFODE_ESTATE;
  }
  if (mlp && !b->has_policy) { b->err = "TFODE_CTRL_MLP needs tfode_batch_load_policy first"; return TFODE_ESTATE; }
  unsigned char* scratch = nullptr;
  if (int rc = noalloc_scratch(b, &scratch)) return rc;
  if (actions_dev) {
    set_actions_kernel<<<(n + 127) / 128, 128, 0, b->stream>>>(b->rec, n, actions_dev, b->cfg.controller == TFODE_CTRL_CRAWLER ? 1 : 0);
    b->launches++;
    CK(cudaGetLastError());
  }
  if (n_substeps > 0) {
    // Nothing in the scratch has to start at zero: dantzig_env writes an island's rows and lcp_dantzig initialises its work arrays before
    // reading them (the cudaMalloc'd scratch of the existing path starts uninitialised too), so the last call's contents are harmless.
    StepParams P = b->P;
    P.set_budget = n_substeps; P.slice_len = 0;
    P.lcp_f = (float*)scratch; P.lcp_i = (int*)(scratch + NOALLOC_F_BYTES);
    const int groups = (n + ENVS_PER_CTA - 1) / ENVS_PER_CTA;
    step_kernel<TFODE_SOLVER_DANTZIG, 2><<<groups, ENVS_PER_CTA * 32, b->smem_step, b->stream>>>(P);
    b->launches++;
    CK(cudaGetLastError());
  }
  if (bodies_out_dev) {
    bodies_kernel<<<(n * NB + 127) / 128, 128, 0, b->stream>>>(b->rec, n, bodies_out_dev);
    b->launches++;
    CK(cudaGetLastError());
  }
  return TFODE_OK;
}

int tfode_batch_rollout_host(TfodeBatch* b, const float* actions_host, int n_substeps, float* obs17_host, float* reward_host, uint8_t* done_host) {
  if (!b) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs;
  float* pin_act = b->h_pin; float* pin_obs = pin_act + 2 * (size_t)n; float* pin_rew = pin_obs + 17 * (size_t)n; uint8_t* pin_done = (uint8_t*)(pin_rew + n);
  if (actions_host) {
    std::memcpy(pin_act, actions_host, sizeof(float) * 2 * (size_t)n);
    CK(cudaMemcpyAsync(b->d_act, pin_act, sizeof(float) * 2 * (size_t)n, cudaMemcpyHostToDevice, b->stream));
    int rc = tfode_batch_set_actions(b, b->d_act); if (rc) return rc;
  }
  int rc = tfode_batch_step(b, n_substeps); if (rc) return rc;
  rc = tfode_batch_observe(b, obs17_host ? b->d_obs : nullptr, reward_host ? b->d_rew : nullptr, done_host ? b->d_done : nullptr, nullptr); if (rc) return rc;
  if (obs17_host) CK(cudaMemcpyAsync(pin_obs, b->d_obs, sizeof(float) * 17 * (size_t)n, cudaMemcpyDeviceToHost, b->stream));
  if (reward_host) CK(cudaMemcpyAsync(pin_rew, b->d_rew, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, b->stream));
  if (done_host) CK(cudaMemcpyAsync(pin_done, b->d_done, (size_t)n, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  if (obs17_host) std::memcpy(obs17_host, pin_obs, sizeof(float) * 17 * (size_t)n);
  if (reward_host) std::memcpy(reward_host, pin_rew, sizeof(float) * (size_t)n);
  if (done_host) std::memcpy(done_host, pin_done, (size_t)n);
  return TFODE_OK;
}

int tfode_batch_get_state(TfodeBatch* b, int env, TfodeEnvState* s) {
  if (!b || !s || env < 0 || env >= b->cfg.n_envs) return TFODE_EINVAL;
  GUARD(b);
  EnvRec r;
  CK(cudaMemcpyAsync(&r, b->rec + env, sizeof(EnvRec), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  std::memset(s, 0, sizeof(*s));
  const Ctrl& c = r.c;
  for (int bd = 0; bd < NB; bd++) {
    for (int k = 0; k < 3; k++) { s->pos[bd][k] = r.body[F_POS + k][bd]; s->lvel[bd][k] = r.body[F_LVEL + k][bd]; s->avel[bd][k] = r.body[F_AVEL + k][bd]; s->facc[bd][k] = r.body[F_FACC + k][bd]; s->tacc[bd][k] = r.body[F_TACC + k][bd]; }
    for (int k = 0; k < 4; k++) s->quat[bd][k] = r.body[F_QUAT + k][bd];
  }
  if (b->cfg.scene != TFODE_SCENE_STEP_BOXES) for (int bd = NBR; bd < NB; bd++) { for (int k = 0; k < 3; k++) s->pos[bd][k] = 0; s->quat[bd][0] = 1; s->quat[bd][1] = s->quat[bd][2] = s->quat[bd][3] = 0; }
  for (int i = 0; i < 4; i++) { s->wheel_vel[i] = c.wheel_vel[i]; s->wheel_fmax[i] = c.wheel_fmax[i]; s->W[i] = c.W[i]; }
  s->servo_vel = c.servo_vel; s->robot_speed = c.robot_speed; s->steering_rate = c.steering_rate; s->boggie_torque = c.boggie_torque;
  s->ic_clock = c.ic_clock; s->ic_period = c.ic_period;
  for (int i = 0; i < 12; i++) { for (int k = 0; k < 4; k++) s->filt[i][k] = c.filt[i][k]; s->ft[i] = c.ft[i]; }
  for (int k = 0; k < 3; k++) s->last_pos[k] = c.last_pos[k];
  s->total_reward = c.total_reward; s->last_reward = c.last_reward;
  for (int i = 0; i < 17; i++) s->last_state[i] = c.last_state[i];
  s->step_yaw_deg = c.step_yaw_deg; s->crawler_amp = c.cr_amp; s->crawler_rate = c.cr_rate; s->crawler_dtorque_dt = c.cr_dtorque_dt;
  s->speed = c.speed; s->ic_start = c.ic_start; s->time_steps = c.time_steps; s->lcg_seed = c.lcg_seed; s->episode = c.episode; s->tick = c.tick;
  s->ic_activated = c.ic_activated; s->ic_tick = c.ic_tick; s->servo_mode_vel = c.servo_mode_vel; s->has_last_state = c.has_last_state;
  s->explore = c.explore; s->fork_collision = c.fork_collision; s->crawler_phase = c.cr_phase; s->done = c.done_step; s->term_reason = c.term_reason;
  s->n_contacts = c.n_contacts; s->episodes_done = c.episodes_done;
  return TFODE_OK;
}

int tfode_batch_set_state(TfodeBatch* b, int env, const TfodeEnvState* s) {
  if (!b || !s || env < 0 || env >= b->cfg.n_envs) return TFODE_EINVAL;
  GUARD(b);
  EnvRec r;
  CK(cudaMemcpyAsync(&r, b->rec + env, sizeof(EnvRec), cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  if (s->episode != r.c.episode) { b->err = "set_state: episode mismatch (per-episode scene constants differ)"; return TFODE_EINVAL; }
  Ctrl& c = r.c;
  const int nb = b->cfg.scene == TFODE_SCENE_STEP_BOXES ? NB : NBR;
  for (int bd = 0; bd < nb; bd++) {
    for (int k = 0; k < 3; k++) { r.body[F_POS + k][bd] = (float)s->pos[bd][k]; r.body[F_LVEL + k][bd] = (float)s->lvel[bd][k]; r.body[F_AVEL + k][bd] = (float)s->avel[bd][k]; r.body[F_FACC + k][bd] = (float)s->facc[bd][k]; r.body[F_TACC + k][bd] = (float)s->tacc[bd][k]; }
    for (int k = 0; k < 4; k++) r.body[F_QUAT + k][bd] = (float)s->quat[bd][k];
  }
  for (int i = 0; i < 4; i++) { c.wheel_vel[i] = (float)s->wheel_vel[i]; c.wheel_fmax[i] = (float)s->wheel_fmax[i]; c.W[i] = (float)s->W[i]; }
  c.servo_vel = (float)s->servo_vel; c.robot_speed = s->robot_speed; c.steering_rate = s->steering_rate; c.boggie_torque = s->boggie_torque;
  c.ic_clock = s->ic_clock; c.ic_period = (float)s->ic_period;
  for (int i = 0; i < 12; i++) { for (int k = 0; k < 4; k++) c.filt[i][k] = s->filt[i][k]; c.ft[i] = (float)s->ft[i]; }
  for (int k = 0; k < 3; k++) c.last_pos[k] = s->last_pos[k];
  c.total_reward = s->total_reward; c.last_reward = s->last_reward;
  for (int i = 0; i < 17; i++) c.last_state[i] = (float)s->last_state[i];
  c.speed = s->speed; c.ic_start = s->ic_start; c.time_steps = s->time_steps; c.lcg_seed = s->lcg_seed; c.tick = s->tick;
  c.ic_activated = s->ic_activated; c.ic_tick = s->ic_tick; c.servo_mode_vel = s->servo_mode_vel; c.has_last_state = s->has_last_state;
  c.explore = s->explore; c.fork_collision = s->fork_collision; c.cr_phase = s->crawler_phase; c.done_call = s->done; c.done_step = s->done; c.term_reason = s->term_reason;
  c.n_contacts = s->n_contacts; c.episodes_done = s->episodes_done; c.parked = 0; c.frozen = 0;
  CK(cudaMemcpyAsync(b->rec + env, &r, sizeof(EnvRec), cudaMemcpyHostToDevice, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return TFODE_OK;
}

// ------------------------------------------------------------------------------------------------ model trees / heightmap images
void tfode_free(void* p) { std::free(p); }

int tfode_model_tree_parse(const char* yaml_text, size_t len, int oblique, TfodeTreeNode** nodes_out, int32_t* n_nodes_out, double** params_out, int32_t* n_params_out) {
  if (!yaml_text || !nodes_out || !n_nodes_out || !params_out || !n_params_out) { g_create_error = "null argument"; return TFODE_EINVAL; }
  try {
    const tfode_fmt::FlatTree t = tfode_fmt::parse_model_tree(yaml_text, len, oblique != 0);
    static_assert(sizeof(tfode_fmt::TreeNode) == sizeof(TfodeTreeNode), "TreeNode mirrors TfodeTreeNode");
    TfodeTreeNode* nd = (TfodeTreeNode*)std::malloc(sizeof(TfodeTreeNode) * std::max<size_t>(1, t.nodes.size()));
    double* pr = (double*)std::malloc(sizeof(double) * std::max<size_t>(1, t.params.size()));
    if (!nd || !pr) { std::free(nd); std::free(pr); g_create_error = "out of memory"; return TFODE_ENOMEM; }
    std::memcpy(nd, t.nodes.data(), sizeof(TfodeTreeNode) * t.nodes.size());
    std::memcpy(pr, t.params.data(), sizeof(double) * t.params.size());
    *nodes_out = nd; *n_nodes_out = (int32_t)t.nodes.size(); *params_out = pr; *n_params_out = (int32_t)t.params.size();
  } catch (const std::exception& e) { g_create_error = e.what(); return TFODE_EINVAL; }
  return TFODE_OK;
}

int tfode_batch_load_model_trees(TfodeBatch* b, const char* yaml_1, size_t len_1, const char* yaml_2, size_t len_2, int oblique, int degree, int interaction_only) {
  if (!b) return TFODE_EINVAL;
  GUARD(b);
  if (!yaml_1 || !yaml_2) { b->err = "null argument"; return TFODE_EINVAL; }
  if (degree < 1 || degree > 2) { b->err = "model tree degree must be 1 (Linear_model_tree) or 2 (Polynomial_model_tree)"; return TFODE_EINVAL; }
  const int nf = tfode_fmt::poly_feature_count(TFODE_TREE_INPUTS, degree, interaction_only != 0);
  if (nf < TFODE_TREE_INPUTS || nf > TFODE_TREE_MAX_FEATURES) { b->err = "model tree feature count out of range"; return TFODE_EINVAL; }
  tfode_fmt::FlatTree t[2];
  try {
    t[0] = tfode_fmt::parse_model_tree(yaml_1, len_1, oblique != 0);
    t[1] = tfode_fmt::parse_model_tree(yaml_2, len_2, oblique != 0);
    for (int k = 0; k < 2; k++)
      for (const auto& nd : t[k].nodes) {
        if ((nd.terminal || oblique) && nd.n_params != nf + 1)   // model_tree.cc:50-51, detected at load time here
          throw std::runtime_error("The dimension of the input vector does not fit the number of parameters of the node " + std::to_string(nd.id));
        if (!nd.terminal && !oblique) {
          const double idx = t[k].params[nd.param_off];
          if (!(idx >= 0) || idx >= nf || idx != std::floor(idx)) throw std::runtime_error("split feature index of node " + std::to_string(nd.id) + " is outside the feature vector");
        }
      }
  } catch (const std::exception& e) { b->err = e.what(); return TFODE_EINVAL; }
  CK(cudaStreamSynchronize(b->stream));
  for (int k = 0; k < 2; k++) {
    cudaFree(b->tree_nodes[k]); cudaFree(b->tree_params[k]); b->tree_nodes[k] = nullptr; b->tree_params[k] = nullptr;
    CK(cudaMalloc(&b->tree_nodes[k], sizeof(TfodeTreeNode) * t[k].nodes.size()));
    CK(cudaMalloc(&b->tree_params[k], sizeof(double) * t[k].params.size()));
    CK(cudaMemcpy(b->tree_nodes[k], t[k].nodes.data(), sizeof(TfodeTreeNode) * t[k].nodes.size(), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(b->tree_params[k], t[k].params.data(), sizeof(double) * t[k].params.size(), cudaMemcpyHostToDevice));
    b->TP.tree[k].nodes = b->tree_nodes[k]; b->TP.tree[k].params = b->tree_params[k]; b->TP.tree[k].n_nodes = (int)t[k].nodes.size();
  }
  b->TP.oblique = oblique != 0; b->TP.degree = degree; b->TP.interaction_only = interaction_only != 0; b->TP.n_features = nf;
  b->has_trees = true;
  return TFODE_OK;
}

int tfode_model_tree_infer(TfodeBatch* b, const double* state_dev, int m, double* act_dev, int32_t* node_dev) {
  if (!b || !state_dev || !act_dev || m < 0) return TFODE_EINVAL;
  GUARD(b);
  if (!b->has_trees) { b->err = "no model trees loaded"; return TFODE_ESTATE; }
  if (m == 0) return TFODE_OK;
  const int blocks = std::min((m + 127) / 128, 148 * 8);
  model_tree_kernel<<<blocks, 128, 0, b->stream>>>(b->TP, nullptr, nullptr, nullptr, nullptr, state_dev, m, act_dev, node_dev);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_batch_get_tree_nodes(TfodeBatch* b, int32_t* out_dev) {
  if (!b || !out_dev) return TFODE_EINVAL;
  GUARD(b);
  const int n = b->cfg.n_envs;
  tree_nodes_kernel<<<(n + 255) / 256, 256, 0, b->stream>>>(b->rec, n, out_dev);
  b->launches++;
  CK(cudaGetLastError());
  return TFODE_OK;
}

int tfode_heightmap_decode_png(const void* png_bytes, size_t len, double z_scale, float** heights_out, int32_t* nrow_out, int32_t* ncol_out) {
  if (!png_bytes || !heights_out || !nrow_out || !ncol_out) { g_create_error = "null argument"; return TFODE_EINVAL; }
  try {
    std::vector<unsigned char> lv; int nrow = 0, ncol = 0;
    tfode_fmt::decode_png((const unsigned char*)png_bytes, len, lv, nrow, ncol);
    float* h = (float*)std::malloc(sizeof(float) * lv.size());
    if (!h) { g_create_error = "out of memory"; return TFODE_ENOMEM; }
    for (size_t i = 0; i < lv.size(); i++) h[i] = (float)((double)lv[i] * z_scale / 255);   // ode/heightfield.hh:54 (double), then the engine's fp32
    *heights_out = h; *nrow_out = nrow; *ncol_out = ncol;
  } catch (const std::exception& e) { g_create_error = e.what(); return TFODE_EINVAL; }
  return TFODE_OK;
}

int tfode_heightmap_load_png(const char* path, double z_scale, float** heights_out, int32_t* nrow_out, int32_t* ncol_out) {
  if (!path) { g_create_error = "null argument"; return TFODE_EINVAL; }
  std::vector<unsigned char> buf;
  try { buf = tfode_fmt::read_file(path); } catch (const std::exception& e) { g_create_error = e.what(); return TFODE_EINVAL; }
  return tfode_heightmap_decode_png(buf.data(), buf.size(), z_scale, heights_out, nrow_out, ncol_out);
}

int tfode_format_transition(const float* row, int kind, char* out, size_t cap) {
  if (!row || !out || cap == 0 || (kind != TFODE_PRINT_TRANSITIONS && kind != TFODE_PRINT_STATE_AND_ACTIONS)) return TFODE_EINVAL;
  // src/rover_1_tf.cc:125-136 / :190-195: every value through "%f", blanks exactly where the reference's printf calls put them
  size_t n = 0;
  auto put = [&](const char* fmt, double v) { if (n < cap) { const int k = snprintf(out + n, cap - n, fmt, v); n += k > 0 ? (size_t)k : 0; } };
  for (int i = 0; i < 17; i++) put("%f ", (double)row[1 + i]);
  if (n < cap) { const int k = snprintf(out + n, cap - n, "%f %f", (double)row[18], (double)row[19]); n += k > 0 ? (size_t)k : 0; }
  if (kind == TFODE_PRINT_TRANSITIONS) for (int i = 0; i < 17; i++) put(" %f", (double)row[22 + i]);
  if (n + 1 >= cap) return TFODE_EINVAL;
  out[n++] = '\n'; out[n] = 0;
  return (int)n;
}

int tfode_savedmodel_read_actor(const char* saved_model_dir, float** weights_out, size_t* n_floats_out) {
  if (!saved_model_dir || !weights_out || !n_floats_out) { g_create_error = "null argument"; return TFODE_EINVAL; }
  *weights_out = nullptr; *n_floats_out = 0;
  try {
    const int dims[4] = {TFODE_NOBS, TFODE_MLP_HIDDEN, TFODE_MLP_HIDDEN, TFODE_NACT};
    const std::vector<float> w = tfode_fmt::read_saved_model_actor(saved_model_dir, dims, 3);
    float* out = (float*)std::malloc(w.size() * sizeof(float));
    if (!out) { g_create_error = "out of memory"; return TFODE_EINVAL; }
    std::memcpy(out, w.data(), w.size() * sizeof(float));
    *weights_out = out; *n_floats_out = w.size();
  } catch (const std::exception& e) { g_create_error = e.what(); return TFODE_EINVAL; }
  return TFODE_OK;
}

int tfode_batch_load_policy_saved_model(TfodeBatch* b, const char* saved_model_dir) {
  if (!b) return TFODE_EINVAL;
  float* w = nullptr; size_t n = 0;
  const int rc = tfode_savedmodel_read_actor(saved_model_dir, &w, &n);
  if (rc != TFODE_OK) { b->err = g_create_error; return rc; }
  const int rc2 = tfode_batch_load_policy(b, w, n);
  std::free(w);
  return rc2;
}

#ifdef TFODE_PHASE_TIMING
// developer build only: cumulative cycles per phase (collide, build, wait, pgs, wait, post), summed over CTAs (thread 0's clock)
int tfode_debug_phase_cycles(unsigned long long* out8, int reset) {
  if (cudaMemcpyFromSymbol(out8, g_phase_cycles, sizeof(unsigned long long) * 16) != cudaSuccess) return TFODE_ECUDA;
  if (reset) { unsigned long long z[16] = {0}; cudaMemcpyToSymbol(g_phase_cycles, z, sizeof(z)); }
  return TFODE_OK;
}
#endif

}  // extern "C"
