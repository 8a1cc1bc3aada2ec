// C++ host side of the batched rover stepping engine: the reference's class API
// (ode::Environment, ode::Object/Box/HeightField, ode::Servo, robot::Robot, robot::Rover_1,
// robot::Rover_1_tf, robot::Crawler_1, Sim_loop) re-expressed over N environments that live on
// one B200 and are stepped by the C ABI of include/tfode_batch.h.  Header-only; link with
// -ltfode_b200.  There is no CPU stepping path behind these classes: every next_step() is a
// kernel launch, and construction throws std::runtime_error (the reference's error convention,
// ode/heightfield.hh:48, tf_cpp_binding.cc:32) when the library cannot reach a B200.
//
// Mapping of the reference's scene code (src/scene_1_mt.cc, src/rover_training_1.cc, src/crawling.cc):
//
//   ode::Environment env( 0.5 );                       ode::Environment env( 0.5, n_envs, device );
//   robot::Rover_1_tf robot( env, pose, actor_dir );   robot::Rover_1_tf robot( env, pose, weights, n );
//   ode::Box step( env, pos, 1, 1, 3, h, false );      same; only the reference's two step boxes are accepted
//   step.set_rotation( 0, 0, yaw ); step.fix();        same (yaw becomes cfg.step_yaw_deg)
//   ode::HeightField field( env, pos, data, ... );     same (heights as float array; PNG decode is the caller's)
//   env.next_step( dt ); robot.next_step( dt );        same two calls; the fused kernel runs at robot.next_step()
//   robot.GetPosition()                                robot.GetPosition( env_index )
//
// The objects are recorded while the scene is being described and the batch is created lazily at the
// first step (or by env.finalize()), which is when the reference's constructors would have run.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include <cuda_runtime_api.h>

#include "../../include/tfode_batch.h"

namespace tfode_host {
struct Vec3 { double v[3]; Vec3(double x = 0, double y = 0, double z = 0) : v{x, y, z} {} double x() const { return v[0]; } double y() const { return v[1]; } double z() const { return v[2]; } double operator[](int i) const { return v[i]; } };
inline void check(int rc, const TfodeBatch* b, const char* what) {
  if (rc != TFODE_OK) throw std::runtime_error(std::string(what) + ": " + tfode_last_error(b));
}
}  // namespace tfode_host

namespace ode {
using Vector3d = tfode_host::Vec3;  // stands in for Eigen::Vector3d (Eigen is not a dependency of this path)

class Object;

// ode/environment.hh:54-165 for a batch of worlds.
class Environment {
 public:
  static constexpr double time_step = 0.05;  // ode/environment.hh:57 (unused default, kept for signature parity)
  explicit Environment(double mu = 0.7, int n_envs = 1, int device = 0) {
    tfode_batch_config_default(&_cfg, 0);
    _cfg.mu = (float)mu; _cfg.n_envs = n_envs; _cfg.device = device; _cfg.scene = TFODE_SCENE_PLANE;
  }
  ~Environment() { if (_d_bodies) cudaFree(_d_bodies); if (_batch) tfode_batch_destroy(_batch); }
  Environment(const Environment&) = delete;
  Environment& operator=(const Environment&) = delete;

  // The world step.  The device kernel fuses Environment::next_step and Rover_1::next_step (state stays on chip
  // between the two), so this call records the request and the launch happens in Robot::next_step, which the
  // reference's step functions always call right after (src/scene_1.cc:88-89, Q6).
  void next_step(double dt = time_step) {
    finalize();
    if (std::fabs(dt - (double)_cfg.dt) > 1e-12) throw std::runtime_error("Environment::next_step: dt differs from the batch's time step");
    _pending_steps++;
  }
  // n fused steps at once (what Sim_loop::run uses): n x { env.next_step; robot.next_step }
  void step_fused(int n) { finalize(); tfode_host::check(tfode_batch_step(_batch, n), _batch, "tfode_batch_step"); _cache_valid = false; _time_steps += n; }

  void finalize() {
    if (_batch) return;
    if (_hf.size()) { _cfg.scene = TFODE_SCENE_HEIGHTFIELD; _cfg.hf_heights = _hf.data(); }
    else if (_n_steps_boxes == 2) _cfg.scene = TFODE_SCENE_STEP_BOXES;
    else if (_n_steps_boxes != 0) throw std::runtime_error("Environment: the step scene needs both boxes of src/rover_training_1.cc:90-98");
    int rc = tfode_batch_create(&_cfg, &_batch);
    if (rc != TFODE_OK) throw std::runtime_error(std::string("tfode_batch_create: ") + tfode_last_error(nullptr));
    if (_policy.size()) tfode_host::check(tfode_batch_load_policy(_batch, _policy.data(), _policy.size()), _batch, "tfode_batch_load_policy");
    if (_has_trees) _load_trees();
    size_t n = (size_t)_cfg.n_envs;
    _h_bodies.resize(n * TFODE_NBODY * 13); _h_obs.resize(n * TFODE_NOBS); _h_rew.resize(n); _h_done.resize(n);
    if (cudaMalloc((void**)&_d_bodies, _h_bodies.size() * sizeof(float)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
  }
  TfodeBatch* batch() { finalize(); return _batch; }
  bool finalized() const { return _batch != nullptr; }
  TfodeBatchConfig& config() { if (_batch) throw std::runtime_error("Environment: configuration is frozen once the batch exists"); return _cfg; }
  const TfodeBatchConfig& cconfig() const { return _cfg; }
  int n_envs() const { return _cfg.n_envs; }
  double get_mu() const { return _cfg.mu; }
  long time_steps() const { return _time_steps; }

  // body poses/velocities and observations of every env, downloaded once per step on demand
  const float* bodies() {
    refresh();
    return _h_bodies.data();
  }
  const float* observations() { refresh(); return _h_obs.data(); }
  const float* rewards() { refresh(); return _h_rew.data(); }
  const uint8_t* dones() { refresh(); return _h_done.data(); }

  // used by the scene objects while the scene is described
  void _register_step_box(int which, double yaw_deg) { if (which == 0) _cfg.step_yaw_deg = (float)yaw_deg; _n_steps_boxes |= 0; _box_seen[which] = true; _n_steps_boxes = (int)_box_seen[0] + (int)_box_seen[1]; }
  void _register_heightfield(const float* h, int ncol, int nrow, double length, double width, const Vector3d& pos) {
    _hf.assign(h, h + (size_t)ncol * nrow); _cfg.hf_ncol = ncol; _cfg.hf_nrow = nrow; _cfg.hf_length = (float)length; _cfg.hf_width = (float)width;
    for (int k = 0; k < 3; k++) _cfg.hf_pos[k] = (float)pos[k];
  }
  void _set_policy(const float* w, size_t n) { _policy.assign(w, w + n); if (_batch) tfode_host::check(tfode_batch_load_policy(_batch, w, n), _batch, "tfode_batch_load_policy"); }
  void _set_model_trees(const std::string& y1, const std::string& y2, bool oblique, unsigned degree, bool interaction_only) {
    _tree_yaml[0] = y1; _tree_yaml[1] = y2; _tree_oblique = oblique; _tree_degree = (int)degree; _tree_io = interaction_only; _has_trees = true;
    if (_batch) _load_trees();
  }
  void _load_trees() {
    tfode_host::check(tfode_batch_load_model_trees(_batch, _tree_yaml[0].data(), _tree_yaml[0].size(), _tree_yaml[1].data(), _tree_yaml[1].size(),
                                                   _tree_oblique, _tree_degree, _tree_io), _batch, "tfode_batch_load_model_trees");
  }
  int _take_pending() { int p = _pending_steps; _pending_steps = 0; return p; }

 private:
  std::string _tree_yaml[2]; bool _tree_oblique = false, _tree_io = false, _has_trees = false; int _tree_degree = 1;
  void refresh() {
    finalize();
    if (_cache_valid) return;
    tfode_host::check(tfode_batch_get_bodies(_batch, _d_bodies), _batch, "tfode_batch_get_bodies");
    tfode_host::check(tfode_batch_sync(_batch), _batch, "tfode_batch_sync");
    if (cudaMemcpy(_h_bodies.data(), _d_bodies, _h_bodies.size() * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) throw std::runtime_error("cudaMemcpy failed");
    tfode_host::check(tfode_batch_rollout_host(_batch, nullptr, 0, _h_obs.data(), _h_rew.data(), _h_done.data()), _batch, "tfode_batch_rollout_host");
    _cache_valid = true;
  }
  TfodeBatchConfig _cfg;
  TfodeBatch* _batch = nullptr;
  std::vector<float> _hf, _policy, _h_bodies, _h_obs, _h_rew;
  std::vector<uint8_t> _h_done;
  float* _d_bodies = nullptr;
  bool _cache_valid = false, _box_seen[2] = {false, false};
  int _n_steps_boxes = 0, _pending_steps = 0;
  long _time_steps = 0;
};

// ode/object.hh:42-157 -- a rigid body of every environment; getters take the environment index.
class Object {
 public:
  Object(Environment& env, int body_index) : _env(&env), _body(body_index) {}
  virtual ~Object() {}
  Vector3d get_pos(int e = 0) const { const float* b = rec(e); return Vector3d(b[0], b[1], b[2]); }
  void get_quaternion(double q[4], int e = 0) const { const float* b = rec(e); for (int k = 0; k < 4; k++) q[k] = b[3 + k]; }
  Vector3d get_vel(int e = 0) const { const float* b = rec(e); return Vector3d(b[7], b[8], b[9]); }
  Vector3d get_angvel(int e = 0) const { const float* b = rec(e); return Vector3d(b[10], b[11], b[12]); }
  // 3x3 rotation (row-major) of the body, dQtoR
  void get_rotation(double R[9], int e = 0) const {
    double q[4]; get_quaternion(q, e);
    double qq1 = 2 * q[1] * q[1], qq2 = 2 * q[2] * q[2], qq3 = 2 * q[3] * q[3];
    R[0] = 1 - qq2 - qq3; R[1] = 2 * (q[1] * q[2] - q[0] * q[3]); R[2] = 2 * (q[1] * q[3] + q[0] * q[2]);
    R[3] = 2 * (q[1] * q[2] + q[0] * q[3]); R[4] = 1 - qq1 - qq3; R[5] = 2 * (q[2] * q[3] - q[0] * q[1]);
    R[6] = 2 * (q[1] * q[3] - q[0] * q[2]); R[7] = 2 * (q[2] * q[3] + q[0] * q[1]); R[8] = 1 - qq1 - qq2;
  }
  virtual void fix() { _fixed = true; }
  bool get_fix() const { return _fixed; }
  virtual void set_collision_group(const char* group, int = -1) { _group = group ? group : ""; }
  const char* get_collision_group(int = 0) const { return _group.c_str(); }
  Environment& get_env() const { return *_env; }
  int body_index() const { return _body; }
 protected:
  const float* rec(int e) const {
    if (e < 0 || e >= _env->n_envs() || _body < 0) throw std::runtime_error("Object: environment/body index out of range");
    return _env->bodies() + ((size_t)e * TFODE_NBODY + _body) * 13;
  }
  Environment* _env; int _body; bool _fixed = false; std::string _group;
};

// ode/box.hh:33-62.  Free-standing boxes are the terrain of the step scene: only the two boxes every shipped scene
// builds (1 x 3 x 0.21 at (1,0,h/2) with a yaw, 2 x 3 x 0.21 at (2,0,h/2); src/rover_training_1.cc:88-98) exist on the device.
class Box : public Object {
 public:
  Box(Environment& env, const Vector3d& pos, double mass, double l, double w, double h, bool /*casts_shadow*/ = true)
      : Object(env, -1), _l(l), _w(w), _h(h), _pos(pos) {
    const double sh = (double)(float)(0.105 * 2);
    if (mass == 1 && std::fabs(w - 3) < 1e-9 && std::fabs(h - sh) < 1e-6 && std::fabs(pos.y()) < 1e-12 && std::fabs(pos.z() - sh / 2) < 1e-6) {
      if (std::fabs(l - 1) < 1e-9 && std::fabs(pos.x() - 1) < 1e-9) _which = 0;
      else if (std::fabs(l - 2) < 1e-9 && std::fabs(pos.x() - 2) < 1e-9) _which = 1;
    }
    if (_which < 0) throw std::runtime_error("ode::Box: only the reference's two step boxes are supported as free-standing bodies");
    _body = TFODE_NBODY_ROVER + _which;
    env._register_step_box(_which, 0.0);
  }
  void set_rotation(double phi, double theta, double psi) {  // ode/object.cc:116-121; the step scenes only yaw the first box
    if (phi != 0 || theta != 0 || (_which != 0 && psi != 0)) throw std::runtime_error("ode::Box::set_rotation: only a yaw of the first step box is supported");
    _env->_register_step_box(_which, psi * 180.0 / M_PI);
  }
 private:
  double _l, _w, _h; Vector3d _pos; int _which = -1;
};

// ode/heightfield.hh:33-59,91-105 (the float-array constructor; an image is decoded by the caller: h[iz*ncol+ix], iz = 0 is the top row)
class HeightField : public Object {
 public:
  HeightField(Environment& env, const Vector3d& pos, const float* heights, int ncol, int nrow, double length, double width)
      : Object(env, -1) { if (!heights || ncol < 2 || nrow < 2) throw std::runtime_error("HeightField: empty height data"); env._register_heightfield(heights, ncol, nrow, length, width, pos); }
  // the image constructor (ode/heightfield.hh:42-59): PNG decoded by the engine's own reader; like the reference, the
  // `int z_scale` argument is ignored and the heights are level * 0.3 / 255; throws "Can't open <path>" like :48
  HeightField(Environment& env, const Vector3d& pos, const char* heightimage_path, int /*z_scale*/, double length, double width) : Object(env, -1) {
    float* h = nullptr; int32_t nrow = 0, ncol = 0;
    if (tfode_heightmap_load_png(heightimage_path, 0.3, &h, &nrow, &ncol) != TFODE_OK) throw std::runtime_error(tfode_last_error(nullptr));
    env._register_heightfield(h, ncol, nrow, length, width, pos);
    tfode_free(h);
  }
};

// ode/servo.hh:13-88 -- the centre steering hinge; the software limit stop and the P/velocity logic run on the device.
class Servo {
 public:
  explicit Servo(Environment& env) : _env(&env) {}
  double get_true_angle(int e = 0) const { return (double)_env->observations()[(size_t)e * TFODE_NOBS + 1] * M_PI / 180.0; }  // obs[1] = steering angle [deg]
 private:
  Environment* _env;
};
}  // namespace ode

namespace robot {
using ode::Vector3d;

// ode/robot.hh:18-74
class Robot {
 public:
  explicit Robot(ode::Environment& env) : _env(&env), _servo(env) {}
  virtual ~Robot() {}
  const std::vector<std::shared_ptr<ode::Object>>& bodies() const { return _bodies; }
  const ode::Servo& servo() const { return _servo; }
  // Runs the fused device step for every env.next_step() recorded since the last call (normally one).
  virtual void next_step(double /*dt*/ = ode::Environment::time_step) {
    int n = _env->_take_pending();
    if (n > 0) _env->step_fused(n);
  }
 protected:
  ode::Environment* _env;
  std::vector<std::shared_ptr<ode::Object>> _bodies;
  ode::Servo _servo;
};

// src/rover.hh:18-156.  Per-env commands are sent with the array setters; the scalar setters broadcast.
class Rover_1 : public Robot {
 public:
  Rover_1(ode::Environment& env, const Vector3d& pose) : Robot(env) {
    if (pose.x() != 0 || pose.y() != 0 || pose.z() != 0) throw std::runtime_error("Rover_1: only pose (0,0,0) is supported (every shipped scene uses it)");
    for (int b = 0; b < TFODE_NBODY_ROVER; b++) _bodies.push_back(std::make_shared<ode::Object>(env, b));
    TfodeBatchConfig& c = env.config();
    c.controller = TFODE_CTRL_EXTERNAL; c.ic_initially_active = 1; c.crawling_mode = 0; c.ic_period = 0;  // Rover_1 ctor defaults (src/rover_1.cc:18-28)
  }
  ~Rover_1() { if (_d_act) cudaFree(_d_act); if (_d_speed) cudaFree(_d_speed); }
  void SetCrawlingMode(bool crawl) { _env->config().crawling_mode = crawl; }
  void SetCmdPeriod(double period) { _env->config().ic_period = (float)period; }
  void SetCmdFreq(double freq) { _env->config().ic_period = (float)(1.0 / freq); }
  // After the batch exists the device owns the IC clock: activation happens at cfg.ic_start (SetICStart), as the step
  // lambdas do with `if (!robot.IsICActivated() && time >= IC_start) robot.ActivateIC()` (src/scene_1_mt.cc:102-103).
  void ActivateIC() {
    if (_env->finalized()) { if (_env->cconfig().ic_start < 0 && !_env->cconfig().ic_initially_active) throw std::runtime_error("ActivateIC after the first step: call SetICStart() before stepping"); return; }
    if (!_env->config().ic_initially_active) _env->config().ic_start = 0.0f;
  }
  void DeactivateIC() { _env->config().ic_initially_active = 0; }
  // the scene lambda's ramp (src/rover_training_1.cc:129-133) and IC start run on the device:
  void UseBuiltInSpeedRamp(float speedf, float term) { _env->config().speedf = speedf; _env->config().term = term; }
  void SetICStart(float t, float jitter = 0) { _env->config().ic_start = t; _env->config().ic_start_jitter = jitter; }

  void SetRobotSpeed(double speed) { std::vector<float> v((size_t)_env->n_envs(), (float)speed); SetRobotSpeed(v.data()); }
  void SetRobotSpeed(const float* per_env) {
    size_t n = (size_t)_env->n_envs();
    if (!_d_speed && cudaMalloc((void**)&_d_speed, n * sizeof(float)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
    cudaMemcpy(_d_speed, per_env, n * sizeof(float), cudaMemcpyHostToDevice);
    tfode_host::check(tfode_batch_set_speed(_env->batch(), _d_speed), _env->batch(), "tfode_batch_set_speed");
  }
  void SetCommands(const float* steer_rate_boggie_torque /*[n,2]*/) {
    size_t n = (size_t)_env->n_envs();
    if (!_d_act && cudaMalloc((void**)&_d_act, n * 2 * sizeof(float)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
    cudaMemcpy(_d_act, steer_rate_boggie_torque, n * 2 * sizeof(float), cudaMemcpyHostToDevice);
    tfode_host::check(tfode_batch_set_actions(_env->batch(), _d_act), _env->batch(), "tfode_batch_set_actions");
  }
  void SetSteeringRateAndBoggieTorque(double rate, double torque) {
    std::vector<float> v((size_t)_env->n_envs() * 2);
    for (size_t i = 0; i < v.size(); i += 2) { v[i] = (float)rate; v[i + 1] = (float)torque; }
    SetCommands(v.data());
  }

  Vector3d GetPosition(int e = 0) const {  // src/rover_1.cc:248-253: main-body point (-0.1095,-0.023,0.0075)
    double R[9]; _bodies[0]->get_rotation(R, e); Vector3d p = _bodies[0]->get_pos(e);
    const double r[3] = {-(0.345 / 2 - 0.063), -0.023, -(0.33 + 0.225 / 2) + 0.45};
    return Vector3d(p.x() + R[0] * r[0] + R[1] * r[1] + R[2] * r[2], p.y() + R[3] * r[0] + R[4] * r[1] + R[5] * r[2], p.z() + R[6] * r[0] + R[7] * r[1] + R[8] * r[2]);
  }
  bool IsUpsideDown(int e = 0) const { double R[9]; _bodies[0]->get_rotation(R, e); return R[8] < 0; }  // :268-273
  double GetDirection(int e = 0) const { return obs(e, 0); }
  double GetSteeringTrueAngle(int e = 0) const { return obs(e, 1); }
  double GetRollAngle(int e = 0) const { return obs(e, 2); }
  double GetPitchAngle(int e = 0) const { return obs(e, 3); }
  double GetBoggieAngle(int e = 0) const { return obs(e, 4); }
  const float* GetState(int e = 0) const { return _env->observations() + (size_t)e * TFODE_NOBS; }  // Rover_1_tf::GetState, 17 values
  bool EpisodeEnded(int e = 0) const { return _env->dones()[e] != 0; }
 protected:
  double obs(int e, int k) const { if (e < 0 || e >= _env->n_envs()) throw std::runtime_error("Rover_1: environment index out of range"); return _env->observations()[(size_t)e * TFODE_NOBS + k]; }
  float* _d_act = nullptr; float* _d_speed = nullptr;
};

// src/rover_tf.hh / src/rover_1_tf.cc: the actor is either the reference's own `actor_path` (the SavedModel directory written by
// td3.actor.save, read without TensorFlow by tfode_savedmodel_read_actor) or a raw fp32 blob (W1 b1 W2 b2 W3 b3, Keras [in,out] layout).
class Rover_1_tf : public Rover_1 {
 public:
  Rover_1_tf(ode::Environment& env, const Vector3d& pose, const char* actor_path, bool tensor_cores = true) : Rover_1(env, pose) {
    float* w = nullptr; size_t n = 0;
    if (tfode_savedmodel_read_actor(actor_path, &w, &n) != TFODE_OK) throw std::runtime_error(tfode_last_error(nullptr));
    TfodeBatchConfig& c = env.config();
    c.controller = TFODE_CTRL_MLP; c.fork_callback = 1; c.policy_precision = tensor_cores ? TFODE_POLICY_TC : TFODE_POLICY_FP32;
    try { env._set_policy(w, n); } catch (...) { tfode_free(w); throw; }
    tfode_free(w);
  }
  Rover_1_tf(ode::Environment& env, const Vector3d& pose, const float* actor_weights, size_t n_floats, bool tensor_cores = true) : Rover_1(env, pose) {
    TfodeBatchConfig& c = env.config();
    c.controller = TFODE_CTRL_MLP; c.fork_callback = 1; c.policy_precision = tensor_cores ? TFODE_POLICY_TC : TFODE_POLICY_FP32;
    env._set_policy(actor_weights, n_floats);
  }
  void SetExploration(bool expl) { _env->config().exploration = expl; }
  double GetLastReward(int e = 0) const { return _env->rewards()[e]; }
  // experience tuples appended since the last call: rows of TFODE_TRANSITION_WIDTH floats (env id, s[17], a[2], r, terminal, s'[17])
  std::vector<float> GetExperience() {
    TfodeBatch* b = _env->batch();
    int cap = 4 * _env->n_envs(), n = 0;
    float* d = nullptr;
    if (cudaMalloc((void**)&d, (size_t)cap * TFODE_TRANSITION_WIDTH * sizeof(float)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
    int dropped = 0;
    int rc = tfode_batch_drain_transitions(b, d, cap, &n, &dropped);   // cap = the ring's default capacity, so pending rows always fit
    if (rc != TFODE_OK) { cudaFree(d); tfode_host::check(rc, b, "tfode_batch_drain_transitions"); }
    if (dropped) { cudaFree(d); throw std::runtime_error(std::to_string(dropped) + " experience tuples were lost: raise transition_capacity"); }
    std::vector<float> out((size_t)n * TFODE_TRANSITION_WIDTH);
    tfode_batch_sync(b);
    cudaMemcpy(out.data(), d, out.size() * sizeof(float), cudaMemcpyDeviceToHost);
    cudaFree(d);
    return out;
  }
};

// src/rover_mt.hh / src/rover_1_mt.cc:13-27: two model trees loaded from YAML files (tree 1 -> steering rate, tree 2 -> boggie torque)
class Rover_1_mt : public Rover_1 {
 public:
  Rover_1_mt(ode::Environment& env, const Vector3d& pose, const std::string yaml_file_path_1, const std::string yaml_file_path_2,
             bool oblique_trees = false, unsigned int degree = 1, bool interaction_only = false) : Rover_1(env, pose) {
    env.config().controller = TFODE_CTRL_MODEL_TREE;
    env._set_model_trees(_slurp(yaml_file_path_1), _slurp(yaml_file_path_2), oblique_trees, degree, interaction_only);
  }
  // node_1 / node_2 (terminal nodes reached at the last IC tick) of environment e
  std::pair<int, int> GetNodes(int e = 0) {
    TfodeBatch* b = _env->batch();
    const int n = _env->n_envs();
    int32_t* d = nullptr;
    if (cudaMalloc((void**)&d, (size_t)n * 2 * sizeof(int32_t)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
    int rc = tfode_batch_get_tree_nodes(b, d);
    if (rc != TFODE_OK) { cudaFree(d); tfode_host::check(rc, b, "tfode_batch_get_tree_nodes"); }
    tfode_batch_sync(b);
    int32_t v[2] = {0, 0};
    cudaMemcpy(v, d + (size_t)e * 2, sizeof(v), cudaMemcpyDeviceToHost);
    cudaFree(d);
    return {v[0], v[1]};
  }

 private:
  static std::string _slurp(const std::string& path) {
    FILE* f = std::fopen(path.c_str(), "rb");
    if (!f) throw std::runtime_error("bad file: " + path);   // YAML::BadFile from YAML::LoadFile (model_tree.cc:15)
    std::string out; char buf[65536]; size_t k;
    while ((k = std::fread(buf, 1, sizeof buf, f)) > 0) out.append(buf, k);
    std::fclose(f);
    return out;
  }
};

// src/rover.hh:245-262, src/rover_1_crawlers.cc:13-27
class Crawler_1 : public Rover_1 {
 public:
  Crawler_1(ode::Environment& env, const Vector3d& pose, float torque_amplitude, float dt_torque, float steering_rate) : Rover_1(env, pose) {
    TfodeBatchConfig& c = env.config();
    c.controller = TFODE_CTRL_CRAWLER; c.crawler_amp = torque_amplitude; c.crawler_dt_torque = dt_torque; c.crawler_rate = steering_rate; c.crawling_mode = 1;
  }
};
}  // namespace robot

// renderer/sim_loop.hh:39 + sim_loop.cc:204-216, headless branch: calls step_function(dt, time) until it returns true.
class Sim_loop {
 public:
  explicit Sim_loop(float timestep) : _timestep(timestep) {}
  void loop(const std::function<bool(float, double)>& step_function) {
    while (true) {
      const double time = (double)((float)_time * _timestep);  // long * float -> float (Q5)
      if (step_function(_timestep, time)) break;
      _time++;
    }
  }
  // fast path: n fused steps with the scene rules (ramp, IC start, termination, auto-reset) evaluated on the device
  void run(ode::Environment& env, int n_steps) { env.step_fused(n_steps); _time += n_steps; }
  double get_time() const { return (double)((float)_time * _timestep); }
 private:
  float _timestep; long _time = 0;
};
