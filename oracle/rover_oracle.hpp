// =============================================================================
// CPU ORACLE -- TEST INFRASTRUCTURE ONLY (see ode_restated.hpp header).
//
// Line-by-line restatement of the reference's first-party hot path for one
// environment:
//   Rover_1 ctor                 src/rover_1.cc:18-245
//   Rover_1::next_step           src/rover_1.cc:478-503  (+ :346-416, :248-305)
//   FT_sensor::Update            ode/ft_sensor.cc:58-88
//   LP_second_order_bilinear     Filters/cpp/filters.cc:137-158   (pinned against oracle/_ref, tests/golden)
//   Servo                        ode/servo.cc:12-25,123-167
//   Environment::_collision      ode/environment.cc:56-150
//   Environment::next_step       ode/environment.hh:127-136 (dWorldQuickStep variant)
//   Rover_1_tf                   src/rover_1_tf.cc:60-199
//   Crawler_1                    src/rover_1_crawlers.cc:13-77
//   simulation() step lambda     src/rover_training_1.cc:84-146,212-224
//   Sim_loop headless loop       renderer/sim_loop.cc:204-216
// Randomness (std::mt19937 + std::random_device in the reference) is replaced by
// counter-based Philox4x32-10 keyed (seed; env, episode, tick, stream).
// PARITY UNPINNED for everything that bottoms out in libode (see ode_restated.hpp).
// =============================================================================
#pragma once
#include <map>
#include <memory>
#include <stdexcept>
#include <string>

#include "../include/tfode_batch.h"
#include "ode_restated.hpp"

namespace rover_oracle {
using namespace ode_restated;

static const double RAD_TO_DEG = 57.29577951308232;
static const double DEG_TO_RAD = 0.017453292519943295;

// ------------------------------------------------------------------ Philox
struct Philox {
  static inline void round(uint32_t* c, const uint32_t* k) {
    uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k[0], n1 = (uint32_t)p1, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k[1], n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
  }
  static inline void gen(uint64_t seed, uint32_t env, uint32_t episode, uint32_t tick, uint32_t stream, uint32_t out[4]) {
    uint32_t c[4] = {env, episode, tick, stream};
    uint32_t k[2] = {(uint32_t)seed, (uint32_t)(seed >> 32)};
    for (int i = 0; i < 10; i++) { round(c, k); k[0] += 0x9E3779B9u; k[1] += 0xBB67AE85u; }
    for (int i = 0; i < 4; i++) out[i] = c[i];
  }
  static inline float u11(uint32_t x) { return 2.0f * ((float)(x >> 8) * (1.0f / 16777216.0f)) - 1.0f; }  // U(-1,1), exact in fp32
};

// ------------------------------------------------------------------ biquad (always double, as the reference)
struct Biquad {
  double Cy1, Cy2, Cx0, Cx1, Cx2, y1 = 0, y2 = 0, x1 = 0, x2 = 0;
  void init(float Te, double w0, float Q) {  // LP_second_order_bilinear ctor, Filters/cpp/filters.cc:148-158
    double w2T2 = std::pow(w0 * Te, 2);
    double den = 2 * w0 * Te + 4 * Q + Q * w2T2;
    Cy1 = 2 * Q * (4 - w2T2) / den;
    Cy2 = (2 * w0 * Te - 4 * Q - Q * w2T2) / den;
    Cx0 = Q * w2T2 / den; Cx1 = 2 * Cx0; Cx2 = Cx0;
  }
  double update(double x) {  // Filters/cpp/filters.cc:137-144
    double y = Cy1 * y1 + Cy2 * y2 + Cx0 * x + Cx1 * x1 + Cx2 * x2;
    y2 = y1; y1 = y; x2 = x1; x1 = x;
    return y;
  }
};

// ------------------------------------------------------------------ geoms
enum { G_SPHERE = 0, G_BOX = 1, G_CYL = 2 };
enum { FEAT_NULL = 0, FEAT_EMPTY_HARD = 1, FEAT_EMPTY_SOFT = 2 };  // dGeomGetData == NULL / feature{group "", HARD} / {group "", SOFT}
template <class T> struct Geom {
  int type = G_SPHERE, body = 0, feat = FEAT_NULL;
  bool callback = false, has_rot = false;
  T off[3] = {0, 0, 0}, offR[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  T a = 0, b = 0, c = 0;  // sphere: a=radius; cylinder: a=radius,b=length; box: sides
};

// MLP actor (scripts/rover_training_1.py:26-35), fp32 as TF_model<float>
struct Policy {
  std::vector<float> W1, b1, W2, b2, W3, b3;
  bool loaded = false;
  void load(const float* w, size_t n) {
    const size_t H = TFODE_MLP_HIDDEN, need = 17 * H + H + H * H + H + H * 2 + 2;
    if (n != need) { loaded = false; return; }
    W1.assign(w, w + 17 * H); w += 17 * H; b1.assign(w, w + H); w += H;
    W2.assign(w, w + H * H); w += H * H; b2.assign(w, w + H); w += H;
    W3.assign(w, w + H * 2); w += H * 2; b3.assign(w, w + 2);
    loaded = true;
  }
  void infer(const float* x, float* y) const {
    const int H = TFODE_MLP_HIDDEN;
    std::vector<float> h1(H), h2(H);
    // fp32 with one fma per weight, accumulated in input order starting from the bias (numerics contract)
    for (int j = 0; j < H; j++) { float s = b1[j]; for (int i = 0; i < 17; i++) s = tfn::fma(x[i], W1[i * H + j], s); h1[j] = s > 0 ? s : 0; }
    for (int j = 0; j < H; j++) { float s = b2[j]; for (int i = 0; i < H; i++) s = tfn::fma(h1[i], W2[i * H + j], s); h2[j] = s > 0 ? s : 0; }
    for (int j = 0; j < 2; j++) { float s = b3[j]; for (int i = 0; i < H; i++) s = tfn::fma(h2[i], W3[i * 2 + j], s); y[j] = tfn::tanh(s); }
  }
};

// Model-tree policy (scripts/ModelTree/cpp/model_tree.{hh,cc}), restated with the reference's own node structure.
// yaml-cpp is absent here, so the constructor's YAML::Node is replaced by a map  node id -> (terminal, params)  that the
// tests fill from PyYAML (the library that wrote the files, scripts/ModelTree/model_tree.py:514-520).
struct ModelTree {
  struct Node { int id; bool terminal = true; std::vector<double> params; std::unique_ptr<Node> children[2]; explicit Node(int i) : id(i) {} };
  struct Entry { bool terminal; std::vector<double> params; };
  bool oblique = false; unsigned degree = 1; bool interaction_only = false;
  std::unique_ptr<Node> root; int node_count = 1;
  void build(Node& node, const std::map<int, Entry>& y) {  // _build_tree_recursively, model_tree.cc:21-45
    const Entry& e = y.at(node.id);
    node.params = e.params;
    if (e.terminal) return;
    node.terminal = false;
    if (!oblique && node.params.size() != 2) throw std::runtime_error("Read " + std::to_string(node.params.size()) + " split parameters for the node " + std::to_string(node.id) + " instead of the 2 required for a straight tree");
    for (int child : {0, 1}) { node.children[child].reset(new Node(++node_count)); build(*node.children[child], y); }
  }
  void load(const std::map<int, Entry>& y, bool oblique_, unsigned degree_, bool io) {
    oblique = oblique_; degree = degree_; interaction_only = io; node_count = 1; root.reset(new Node(1)); build(*root, y);
  }
  static double inner(const std::vector<double>& in, const std::vector<double>& p) {  // std::inner_product(in.begin(), in.end(), p.begin(), p.back())
    double acc = p.back();
    for (size_t i = 0; i < in.size(); i++) acc = acc + in[i] * p[i];
    return acc;
  }
  double traverse(const Node& node, const std::vector<double>& input, int& terminal_node_id) const {  // model_tree.cc:47-65
    if ((node.terminal || oblique) && input.size() != node.params.size() - 1)
      throw std::runtime_error("The dimension of the input vector does not fit the number of parameters of the node " + std::to_string(node.id));
    if (node.terminal) { terminal_node_id = node.id; return inner(input, node.params); }
    bool side;
    if (oblique) side = inner(input, node.params) >= 0;
    else side = input[(size_t)node.params[0]] >= node.params[1];
    return traverse(*node.children[side ? 0 : 1], input, terminal_node_id);
  }
  double predict(const std::vector<double>& input, int& terminal_node_id) const {
    if (degree <= 1) return traverse(*root, input, terminal_node_id);
    // Polynomial_model_tree::predict, model_tree.cc:91-123
    std::vector<double> features = input, prev_chunk = input;
    std::vector<size_t> indices(input.size());
    for (size_t i = 0; i < indices.size(); i++) indices[i] = i;
    for (unsigned d = 1; d < degree; ++d) {
      std::vector<double> new_chunk;
      for (size_t i = 0; i < input.size() - (interaction_only ? d : 0); ++i) {
        const size_t next_index = new_chunk.size();
        for (size_t c = indices[i + (interaction_only ? 1 : 0)]; c < prev_chunk.size(); ++c) new_chunk.push_back(input[i] * prev_chunk[c]);
        indices[i] = next_index;
      }
      features.insert(features.end(), new_chunk.begin(), new_chunk.end());
      prev_chunk = new_chunk;
    }
    return traverse(*root, features, terminal_node_id);
  }
};

struct SharedScene {  // shared by all envs of a batch
  TfodeBatchConfig cfg;
  std::vector<float> hf_data;
  Policy policy;
  ModelTree tree[2];   // Rover_1_mt::_lmt_ptr_1/_2 (src/rover_1_mt.cc:13-27)
};

template <class T> struct RoverSim {
  const SharedScene* sc = nullptr;
  uint32_t env_id = 0;
  World<T> w;
  std::vector<Geom<T>> geoms;
  HeightField<T> hf;
  // step boxes
  T step_side[2][3];
  // joints
  int j_battery = 0, j_servo = 1, j_boggie = 2, j_wheel[4] = {3, 4, 5, 6}, j_fix[2] = {7, 8};
  // Rover_1 members (src/rover.hh)
  double robot_speed = 0, steering_rate = 0, boggie_torque = 0;
  double steering_max_vel = 15, boggie_max_torque = 25, robot_max_speed = 0;
  double wheelbase = 0.58, wheeltrack = 0.61, wheel_radius[4] = {0.105, 0.105, 0.105, 0.105};
  double front_pos[3], sea_elev = 0;
  T W[4] = {0, 0, 0, 0};
  bool ic_activated = true, ic_tick = false, crawling_mode = false;
  double ic_period = 0, ic_clock = 0;
  // Servo members (ode/servo.hh)
  T servo_min = 0, servo_max = 0, servo_vel_max = 0, servo_vel = 0; bool servo_mode_vel = false;
  // FT sensors
  struct FT { int A, B; T ocA[3], ocB[3]; T F[3] = {0, 0, 0}, Tq[3] = {0, 0, 0}; } ft[2];
  T Klin[3] = {T(2e4), T(2e4), T(1e5)}, Kang[3] = {T(1e3), T(1e3), T(1e3)}, Clin[3] = {T(5e2), T(5e2), T(5e2)}, Cang[3] = {1, 1, 1};
  Biquad filt[12];
  // Rover_1_tf
  double last_pos[3] = {0, 0, 0}, total_reward = 0, last_state[17], last_reward = 0;
  bool has_last_state = false, explore = false, fork_collision = false, exploration = false;
  double pending[38]; bool pending_valid = false;
  std::vector<std::vector<float>> transitions;  // emitted rows of TFODE_TRANSITION_WIDTH
  // Crawler_1
  float cr_amp = 30, cr_rate = 10, cr_span = 45; double cr_dtorque_dt = 30; int cr_phase = 1;
  int node_1 = 0, node_2 = 0;   // Rover_1_mt::node_1/node_2
  // scene lambda state
  float speed = 0, ic_start = 1; int64_t time_steps = 0; uint32_t episode = 0, tick = 0;
  double step_yaw_deg = 0;
  bool done = false, frozen = false; int term_reason = 0, n_contacts = 0, episodes_done = 0;
  double last_ep_return = 0, last_ep_x = 0, last_ep_y = 0; int64_t last_ep_steps = 0;
  int max_pair_contacts = 0;
  uint32_t contact_hist[129] = {0};   // steps by number of generated contacts (before the engine's capacity), last bin = 128+

  // -------------------------------------------------------------- construction
  void init(const SharedScene* scene, uint32_t env_global_id) {
    sc = scene; env_id = env_global_id;
    if (sc->cfg.scene == TFODE_SCENE_HEIGHTFIELD) {
      double p[3] = {sc->cfg.hf_pos[0], sc->cfg.hf_pos[1], sc->cfg.hf_pos[2]};
      hf.init(sc->hf_data.data(), sc->cfg.hf_ncol, sc->cfg.hf_nrow, sc->cfg.hf_length, sc->cfg.hf_width, p);
    }
    explore = false;  // function-static in the reference (src/rover_1_tf.cc:168): survives episodes
    reset(0);
  }

  // Everything `simulation()` constructs, in double, then cast to T (the reference constructs in double).
  void reset(uint32_t episode_) {
    const TfodeBatchConfig& cfg = sc->cfg;
    episode = episode_; tick = 0;
    uint32_t rnd[4]; Philox::gen(cfg.seed, env_id, episode, 0, 0, rnd);
    step_yaw_deg = (double)cfg.step_yaw_deg + (double)cfg.step_yaw_jitter_deg * (double)Philox::u11(rnd[0]);
    ic_start = (float)((double)cfg.ic_start + (double)cfg.ic_start_jitter * (double)Philox::u11(rnd[1]));
    cr_amp = (float)((double)cfg.crawler_amp + (double)cfg.crawler_amp_jitter * (double)Philox::u11(rnd[2]));
    cr_rate = (float)((double)cfg.crawler_rate + (double)cfg.crawler_rate_jitter * (double)Philox::u11(rnd[3]));

    World<double> wd; std::vector<Geom<double>> gd;
    wd.gravity[2] = -9.80665;  // cst::g, ode/misc.hh:26
    wd.erp = (double)cfg.erp; wd.cfm = (double)cfg.cfm; wd.iters = cfg.pgs_iters; wd.sor_w = (double)cfg.sor_w; wd.solver = cfg.solver;
    build_rover(wd, gd);
    if (cfg.scene == TFODE_SCENE_STEP_BOXES) build_steps(wd);
    // FT sensors (src/rover_1.cc:205-206; ode/ft_sensor.cc:50-55)
    double ocd[2][2][3];
    { double c0[3] = {wheelbase / 2, 0, 0.33}, c1[3] = {-wheelbase / 2, 0, 0.33};
      wd.body[0].posRelPoint(c0, ocd[0][0]); wd.body[4].posRelPoint(c0, ocd[0][1]);
      wd.body[3].posRelPoint(c1, ocd[1][0]); wd.body[5].posRelPoint(c1, ocd[1][1]); }
    ft[0].A = 0; ft[0].B = 4; ft[1].A = 3; ft[1].B = 5;
    for (int s = 0; s < 2; s++) for (int k = 0; k < 3; k++) { ft[s].ocA[k] = (T)ocd[s][0][k]; ft[s].ocB[k] = (T)ocd[s][1][k]; ft[s].F[k] = 0; ft[s].Tq[k] = 0; }
    convert_world(wd, gd);
    for (int i = 0; i < 12; i++) { filt[i] = Biquad(); filt[i].init(0.001, 4 * M_PI, 0.5); }  // src/rover_1.cc:244 (Te is float in the ctor)
    // Rover_1 ctor defaults + scene overrides (src/rover_training_1.cc:58-65)
    robot_speed = steering_rate = boggie_torque = 0;
    for (int i = 0; i < 4; i++) W[i] = 0;
    ic_clock = 0; ic_tick = false;
    ic_period = (double)cfg.ic_period; ic_activated = cfg.ic_initially_active != 0; crawling_mode = cfg.crawling_mode != 0;
    boggie_max_torque = 25; steering_max_vel = 15;
    servo_min = (T)(-45 * DEG_TO_RAD); servo_max = (T)(45 * DEG_TO_RAD); servo_vel_max = (T)(15 * DEG_TO_RAD); servo_vel = 0; servo_mode_vel = false;
    exploration = cfg.exploration != 0;
    if (cfg.controller == TFODE_CTRL_CRAWLER) {  // Crawler_1 ctor, src/rover_1_crawlers.cc:13-27
      cr_span = 45; boggie_max_torque = cr_amp;
      cr_dtorque_dt = (double)(2 * cr_amp / cfg.crawler_dt_torque);
      cr_phase = 1; crawling_mode = true;
    }
    total_reward = 0; last_reward = 0; has_last_state = false; fork_collision = false; pending_valid = false;
    for (int i = 0; i < 17; i++) last_state[i] = 0;
    for (int i = 0; i < 38; i++) pending[i] = 0;
    { T p[3]; getPosition(p); for (int k = 0; k < 3; k++) last_pos[k] = (double)p[k]; }  // Rover_1_tf ctor :19
    speed = 0; time_steps = 0; done = false; frozen = false; term_reason = 0; n_contacts = 0;
    w.rand_seed = 0;  // a fresh process' dRand() state
  }

  void build_rover(World<double>& wd, std::vector<Geom<double>>& gd) {
    // parameters, src/rover_1.cc:49-103
    const double wheel_mass = 1.4 + 1, wheel_width = 0.1; const int wheel_def = 50;
    const double belly_elev = 0.33;
    const double front_mass = 5.25, front_length = 0.345, front_height = 0.225, front_width = 0.200, front_x_offset = -0.063, front_y_offset = 0.023;
    front_pos[0] = front_length / 2 + front_x_offset; front_pos[1] = front_y_offset; front_pos[2] = belly_elev + front_height / 2;
    const double rear_mass = 3.67, rear_length = 0.312, rear_height = 0.177, rear_width = 0.126, rear_x_offset = 0.046, rear_y_offset = -0.040;
    const double rear_pos[3] = {-rear_length / 2 + rear_x_offset, rear_y_offset, belly_elev + rear_height / 2};
    const double hinge_pos[3] = {0, 0, (front_pos[2] + rear_pos[2]) / 2};
    sea_elev = belly_elev + 0.12;
    const double sea_pos[3] = {-wheelbase / 2, 0, sea_elev};
    const double boggie_mass = 1, boggie_length = 0.1, boggie_height = sea_elev - belly_elev, boggie_width = 0.1;
    const double boggie_pos[3] = {-wheelbase / 2 + 0.035, 0, belly_elev + boggie_height / 2};
    const double fork_mass = 1, fork_length = 0.05, fork_height = 0.1, fork_width = 0.3, fork_elev = belly_elev - fork_height / 2;
    robot_max_speed = 7.4351 / wheel_radius[0];
    for (int i = 1; i < 4; i++) robot_max_speed = std::min(robot_max_speed, 7.4351 * wheel_radius[i]);  // Q4

    auto add_box_body = [&](const double* p, double m, double l, double wd_, double h, bool geom) {
      Body<double> b; copy3(b.pos, p); b.setMassBoxTotal(m, l, wd_, h); wd.body.push_back(b);
      if (geom) { Geom<double> g; g.type = G_BOX; g.body = (int)wd.body.size() - 1; g.a = l; g.b = wd_; g.c = h; gd.push_back(g); }
      return (int)wd.body.size() - 1;
    };
    // bodies 0..5 (:116-176)
    add_box_body(front_pos, front_mass, front_length, front_width, front_height, true);
    { double p[3] = {0.285, 0, belly_elev + 0.155}; add_box_body(p, 3, 0.0975, 0.151, 0.065, true); }
    add_box_body(rear_pos, rear_mass, rear_length, rear_width, rear_height, true);
    add_box_body(boggie_pos, boggie_mass, boggie_length, boggie_width, boggie_height, false);
    const double motor_radius = 0.025, motor_length = 0.1;
    for (int f = 0; f < 2; f++) {
      double sx = f == 0 ? 1 : -1;
      double p[3] = {sx * wheelbase / 2, 0, fork_elev};
      int bi = add_box_body(p, fork_mass, fork_length, fork_width, fork_height, false);
      for (int s = 0; s < 2; s++) {
        Geom<double> g; g.type = G_CYL; g.body = bi; g.a = motor_radius; g.b = motor_length;
        RFromEulerAngles(g.offR, M_PI / 2, 0.0, 0.0); g.has_rot = true;
        double ap[3] = {sx * wheelbase / 2, (s == 0 ? 1 : -1) * (wheeltrack - wheel_width - motor_length) / 2, wheel_radius[f * 2 + s]};
        wd.body[bi].posRelPoint(ap, g.off);  // dGeomSetOffsetWorldPosition
        if (sc->cfg.fork_callback) { g.feat = FEAT_EMPTY_HARD; g.callback = true; }  // src/rover_1_tf.cc:50-56 (Q2: type uninitialised -> HARD)
        gd.push_back(g);
      }
    }
    // battery clamp (:128-132)
    j_battery = wd.createSlider(1, 0); wd.setSliderAxis(j_battery, 0, 1, 0);
    wd.joint[j_battery].limot.lostop = 0; wd.joint[j_battery].limot.histop = 0;
    // centre hinge servo (:181-189, ode/servo.cc:12-25)
    j_servo = wd.createHinge(2, 0);
    wd.setHingeAnchor(j_servo, hinge_pos[0], hinge_pos[1], hinge_pos[2]);
    wd.setHingeAxis(j_servo, 0, 0, 1);
    wd.joint[j_servo].limot.setFMax(180);  // dInfinity then set_torque_max(STEERING_MAX_TORQUE)
    // boggie hinge (:194-200)
    j_boggie = wd.createHinge(3, 2);
    wd.setHingeAxis(j_boggie, 1, 0, 0);
    wd.setHingeAnchor(j_boggie, sea_pos[0], sea_pos[1], sea_pos[2]);
    wd.joint[j_boggie].limot.lostop = -45 * DEG_TO_RAD; wd.joint[j_boggie].limot.histop = 45 * DEG_TO_RAD;
    // wheels (:209-233, ode/wheel.hh:33-74)
    for (int i = 0; i < 4; i++) {
      double wp[3] = {(i / 2 ? -1 : 1) * wheelbase / 2, (i % 2 ? -1 : 1) * wheeltrack / 2, wheel_radius[i]};
      Body<double> b; copy3(b.pos, wp);
      double rr = wheel_radius[i] - wheel_width / 2;
      b.setMassCylinderTotalZ(wheel_mass, rr, wheel_width);
      double Rw[9]; RFromEulerAngles(Rw, M_PI / 2, 0.0, 0.0); b.setRotation(Rw);
      wd.body.push_back(b);
      int bi = (int)wd.body.size() - 1;
      { Geom<double> g; g.type = G_CYL; g.body = bi; g.a = rr; g.b = wheel_width; gd.push_back(g); }
      for (int k = 0; k < wheel_def; k++) {
        Geom<double> g; g.type = G_SPHERE; g.body = bi; g.a = wheel_width / 2;
        g.off[0] = rr * std::cos(k * 2 * M_PI / wheel_def); g.off[1] = rr * std::sin(k * 2 * M_PI / wheel_def); g.off[2] = 0;
        if (k == wheel_def - 1) g.feat = FEAT_EMPTY_SOFT;  // Q1: set_contact_type(SOFT) touches only the last geom
        gd.push_back(g);
      }
      j_wheel[i] = wd.createHinge(wp[0] > 0 ? 4 : 5, bi);
      wd.setHingeAnchor(j_wheel[i], wp[0], wp[1], wp[2]);
      wd.setHingeAxis(j_wheel[i], 0, -1, 0);
      wd.joint[j_wheel[i]].limot.setFMax(25.4);
    }
  }

  void build_steps(World<double>& wd) {  // src/rover_training_1.cc:84-98
    const float step_height = (float)(0.105 * 2);
    for (int s = 0; s < 2; s++) {
      Body<double> b;
      b.pos[0] = s == 0 ? 1 : 2; b.pos[1] = 0; b.pos[2] = step_height / 2;
      double l = s == 0 ? 1 : 2;
      b.setMassBoxTotal(1, l, 3, step_height);
      if (s == 0) {  // dRFromEulerAngles(0,0,psi) with the contract's sin/cos (the yaw is re-drawn on the device at every auto-reset)
        double sp, cp; tfn::d_sincos(step_yaw_deg * 3.14159265358979323846 / 180.0, &sp, &cp);
        double Rs[9] = {cp, sp, 0, -sp, cp, 0, 0, 0, 1};
        b.setRotation(Rs);
      }
      wd.body.push_back(b);
      step_side[s][0] = (T)l; step_side[s][1] = 3; step_side[s][2] = (T)step_height;
      int bi = (int)wd.body.size() - 1;
      j_fix[s] = wd.createSlider(bi, -1);  // Object::fix, ode/object.cc:155-179
      wd.setSliderAxis(j_fix[s], 0, 0, 1);
      wd.joint[j_fix[s]].limot.lostop = 0; wd.joint[j_fix[s]].limot.histop = 0;
    }
  }

  void convert_world(const World<double>& wd, const std::vector<Geom<double>>& gd) {
    w = World<T>();
    for (int k = 0; k < 3; k++) w.gravity[k] = (T)wd.gravity[k];
    w.erp = (T)wd.erp; w.cfm = (T)wd.cfm; w.iters = wd.iters; w.sor_w = (T)wd.sor_w; w.solver = wd.solver;
    for (auto& s : wd.body) {
      Body<T> b;
      for (int k = 0; k < 3; k++) { b.pos[k] = (T)s.pos[k]; b.lvel[k] = 0; b.avel[k] = 0; b.facc[k] = 0; b.tacc[k] = 0; b.I[k] = (T)s.I[k]; b.invI[k] = (T)s.invI[k]; }
      for (int k = 0; k < 4; k++) b.q[k] = (T)s.q[k];
      RfromQ(b.R, b.q);
      b.mass = (T)s.mass; b.invMass = (T)s.invMass;
      w.body.push_back(b);
    }
    for (auto& s : wd.joint) {
      Joint<T> j; j.type = s.type; j.b1 = s.b1; j.b2 = s.b2; j.reverse = s.reverse;
      for (int k = 0; k < 3; k++) { j.anchor1[k] = (T)s.anchor1[k]; j.anchor2[k] = (T)s.anchor2[k]; j.axis1[k] = (T)s.axis1[k]; j.axis2[k] = (T)s.axis2[k]; j.offset[k] = (T)s.offset[k]; }
      for (int k = 0; k < 4; k++) j.qrel[k] = (T)s.qrel[k];
      j.limot.vel = (T)s.limot.vel; j.limot.fmax = (T)s.limot.fmax; j.limot.lostop = (T)s.limot.lostop; j.limot.histop = (T)s.limot.histop;
      j.limot.fudge_factor = (T)s.limot.fudge_factor; j.limot.normal_cfm = (T)s.limot.normal_cfm; j.limot.stop_erp = (T)s.limot.stop_erp; j.limot.stop_cfm = (T)s.limot.stop_cfm;
      w.joint.push_back(j);
    }
    geoms.clear();
    for (auto& s : gd) {
      Geom<T> g; g.type = s.type; g.body = s.body; g.feat = s.feat; g.callback = s.callback; g.has_rot = s.has_rot;
      for (int k = 0; k < 3; k++) g.off[k] = (T)s.off[k];
      for (int k = 0; k < 9; k++) g.offR[k] = (T)s.offR[k];
      g.a = (T)s.a; g.b = (T)s.b; g.c = (T)s.c;
      geoms.push_back(g);
    }
  }

  // -------------------------------------------------------------- observation getters (src/rover_1.cc:248-305)
  void getPosition(T* out) const {
    T p[3] = {(T)(-front_pos[0]), (T)(-front_pos[1]), (T)(-front_pos[2] + sea_elev)};
    w.body[0].relPointPos(p, out);
  }
  T getDirection() const {
    T v[3], e[3] = {1, 0, 0}; w.body[0].vectorToWorld(e, v);
    T direction = tfn::asin(v[1]) * (T)RAD_TO_DEG;
    if (v[0] < 0) direction = (v[1] > 0 ? 1 : -1) * T(180) - direction;
    return direction;
  }
  bool isUpsideDown() const { T v[3], e[3] = {0, 0, 1}; w.body[0].vectorToWorld(e, v); return v[2] < 0; }
  T getRollAngle() const { T v[3], e[3] = {0, 1, 0}; w.body[0].vectorToWorld(e, v); return tfn::asin(v[2]) * (T)RAD_TO_DEG; }
  T getPitchAngle() const { T v[3], e[3] = {1, 0, 0}; w.body[0].vectorToWorld(e, v); return tfn::asin(-v[2]) * (T)RAD_TO_DEG; }
  T getBoggieAngle() const { return w.hingeAngle(w.joint[j_boggie]) * (T)RAD_TO_DEG; }
  T getSteeringTrueAngle() const { return w.hingeAngle(w.joint[j_servo]) * (T)RAD_TO_DEG; }
  void getState(double* s) const {  // Rover_1_tf::GetState, src/rover_1_tf.cc:60-77
    s[0] = getDirection(); s[1] = getSteeringTrueAngle(); s[2] = getRollAngle(); s[3] = getPitchAngle(); s[4] = getBoggieAngle();
    for (int k = 0; k < 3; k++) { s[5 + k] = ft[0].F[k]; s[8 + k] = ft[0].Tq[k]; s[11 + k] = ft[1].F[k]; s[14 + k] = ft[1].Tq[k]; }
  }

  // -------------------------------------------------------------- setters (src/rover_1.cc:308-341)
  void setRobotSpeed(double s) { robot_speed = std::min(std::max(-robot_max_speed, s), robot_max_speed); }
  void setSteeringRate(double r) { steering_rate = std::min(std::max(-steering_max_vel, r), steering_max_vel); }
  void setBoggieTorque(double t) { boggie_torque = std::min(std::max(-boggie_max_torque, t), boggie_max_torque); }

  // -------------------------------------------------------------- Environment::_collision (ode/environment.cc:56-150)
  void geomPose(const Geom<T>& g, T* p, T* R) const {
    const Body<T>& b = w.body[g.body];
    b.relPointPos(g.off, p);
    if (g.has_rot) { for (int r = 0; r < 3; r++) for (int c = 0; c < 3; c++) R[r * 3 + c] = tfn::dot3(b.R[r * 3], b.R[r * 3 + 1], b.R[r * 3 + 2], g.offR[c], g.offR[3 + c], g.offR[6 + c]); }
    else for (int k = 0; k < 9; k++) R[k] = b.R[k];
  }
  void cylAABB(const T* p, const T* R, T radius, T lz, T* lo, T* hi) const {  // dxCylinder::computeAABB
    const int ax[3] = {2, 5, 8};
    for (int k = 0; k < 3; k++) {
      T rr = R[ax[k]];
      T range = std::fabs(rr * lz * T(0.5)) + radius * std::sqrt(std::max(T(0), T(1) - rr * rr));
      lo[k] = p[k] - range; hi[k] = p[k] + range;
    }
  }
  void addContact(const ContactGeom<T>& cg, int b1, int b2, bool soft) {
    if ((int)w.contact.size() >= sc->cfg.max_contacts) return;  // engine capacity (DESIGN.md D-CAP)
    Joint<T> j; j.type = J_CONTACT; World<T>::attach(j, b1, b2);
    copy3(j.cpos, cg.pos); copy3(j.cnormal, cg.normal); j.cdepth = cg.depth;
    j.mu = (T)sc->cfg.mu;
    if (soft) { j.soft = true; j.soft_cfm = T(0.005); j.soft_erp = T(0.4); }
    w.contact.push_back(j);
  }

  // Order (our canonical definition, D-ORDER): terrain geoms in creation order (plane; step, step_c |
  // heightfield); against each, first the 11 non-sphere rover geoms in creation order (chassis boxes, motor
  // cylinders, rims), then the 200 tire spheres in creation order.  ODE's own order is the hash-space
  // traversal order, which cannot be reproduced.
  void collide() {
    const TfodeBatchConfig& cfg = sc->cfg;
    const T up[3] = {0, 0, 1};
    ContactGeom<T> cg[10];
    int total = 0;
    const int n_terrain = cfg.scene == TFODE_SCENE_STEP_BOXES ? 3 : (cfg.scene == TFODE_SCENE_HEIGHTFIELD ? 2 : 1);
    for (int t = 0; t < n_terrain; t++) {
      for (int pass = 0; pass < 2; pass++) {
        for (const Geom<T>& g : geoms) {
          if ((g.type == G_SPHERE) != (pass == 1)) continue;
          T p[3], R[9]; geomPose(g, p, R);
          const bool soft = g.feat == FEAT_EMPTY_SOFT;
          int n = 0, b2 = -1;
          T lo[3], hi[3];
          if (g.callback) cylAABB(p, R, g.a, g.b, lo, hi);
          if (t == 0) {  // ground plane z = 0 (feature "ground")
            if (g.type == G_SPHERE) n = collideSpherePlane(p, g.a, up, T(0), cg);
            else if (g.type == G_CYL) n = collideCylinderPlane(p, R, g.a, g.b, up, T(0), cg, 10);
            else { T side[3] = {g.a, g.b, g.c}; n = collideBoxPlane(p, R, side, up, T(0), cg, 10); }
            if (g.callback && lo[2] <= 0) fork_collision = true;  // Q10: the half-space AABB of the z=0 plane is z <= 0
          } else if (cfg.scene == TFODE_SCENE_STEP_BOXES) {
            const int s = t - 1;
            const Body<T>& sb = w.body[TFODE_NBODY_ROVER + s];
            b2 = TFODE_NBODY_ROVER + s;
            if (g.callback) {  // AABB overlap test of the near callback (dxBox::computeAABB)
              bool ov = true;
              for (int k = 0; k < 3; k++) {
                T range = T(0.5) * (std::fabs(sb.R[k * 3 + 0] * step_side[s][0]) + std::fabs(sb.R[k * 3 + 1] * step_side[s][1]) + std::fabs(sb.R[k * 3 + 2] * step_side[s][2]));
                if (lo[k] > sb.pos[k] + range || hi[k] < sb.pos[k] - range) ov = false;
              }
              if (ov) fork_collision = true;
            }
            auto test = [&](const T* q, T* nrm, T* dep) { return pointInBox(q, sb.pos, sb.R, step_side[s], nrm, dep); };
            if (g.type == G_SPHERE) n = collideSphereBox(p, g.a, sb.pos, sb.R, step_side[s], cg);
            else if (g.type == G_CYL) n = cylSamples(p, R, g.a, g.b, test, cg);
            else n = boxSamples(p, R, g.a, g.b, g.c, test, cg);
          } else {
            if (g.callback) {  // heightfield AABB: footprint x [pos.z - 1, pos.z + 1] (bounds -1..1, ode/heightfield.hh:95)
              bool ov = !(lo[0] > hf.pos[0] + hf.length * T(0.5) || hi[0] < hf.pos[0] - hf.length * T(0.5) ||
                          lo[1] > hf.pos[1] + hf.width * T(0.5) || hi[1] < hf.pos[1] - hf.width * T(0.5) ||
                          lo[2] > hf.pos[2] + T(1) || hi[2] < hf.pos[2] - T(1));
              if (ov) fork_collision = true;
            }
            auto test = [&](const T* q, T* nrm, T* dep) { return pointVsHF(q, hf, nrm, dep); };
            if (g.type == G_SPHERE) n = collideSphereHF(p, g.a, hf, cg, 4);
            else if (g.type == G_CYL) n = cylSamples(p, R, g.a, g.b, test, cg);
            else n = boxSamples(p, R, g.a, g.b, g.c, test, cg);
          }
          for (int i = 0; i < n; i++) addContact(cg[i], g.body, b2, soft);
          total += n; max_pair_contacts = std::max(max_pair_contacts, n);
        }
      }
    }
    n_contacts = total;
    contact_hist[total < 128 ? total : 128]++;
  }

  // cylinder vs box/heightfield (D-CYLBOX, D-HF): 8 rim points per cap, the deepest penetrating
  // sample of each cap gives one contact (ties: lower index).
  template <class F> int cylSamples(const T* p, const T* R, T radius, T length, F test, ContactGeom<T>* out) const {
    static const double cs[8][2] = {{1, 0}, {0.70710678118654752, 0.70710678118654752}, {0, 1}, {-0.70710678118654752, 0.70710678118654752},
                                    {-1, 0}, {-0.70710678118654752, -0.70710678118654752}, {0, -1}, {0.70710678118654752, -0.70710678118654752}};
    int cnt = 0;
    for (int cap = 0; cap < 2; cap++) {
      T s = (cap == 0 ? T(-0.5) : T(0.5)) * length;
      T best = 0; bool any = false; ContactGeom<T> bc;
      for (int k = 0; k < 8; k++) {
        T lx = radius * (T)cs[k][0], ly = radius * (T)cs[k][1];
        T q[3];
        for (int i = 0; i < 3; i++) q[i] = p[i] + R[i * 3 + 0] * lx + R[i * 3 + 1] * ly + R[i * 3 + 2] * s;
        T nrm[3], dep;
        if (test(q, nrm, &dep) && (!any || dep > best)) { any = true; best = dep; copy3(bc.pos, q); copy3(bc.normal, nrm); bc.depth = dep; }
      }
      if (any) out[cnt++] = bc;
    }
    return cnt;
  }
  // box vs box/heightfield: the 8 corners, first 4 penetrating in index order (bit0=x, bit1=y, bit2=z)
  template <class F> int boxSamples(const T* p, const T* R, T lx, T ly, T lz, F test, ContactGeom<T>* out) const {
    int cnt = 0;
    for (int k = 0; k < 8 && cnt < 4; k++) {
      T hx = ((k & 1) ? T(0.5) : T(-0.5)) * lx, hy = ((k & 2) ? T(0.5) : T(-0.5)) * ly, hz = ((k & 4) ? T(0.5) : T(-0.5)) * lz;
      T q[3];
      for (int i = 0; i < 3; i++) q[i] = p[i] + R[i * 3 + 0] * hx + R[i * 3 + 1] * hy + R[i * 3 + 2] * hz;
      T nrm[3], dep;
      if (test(q, nrm, &dep)) { copy3(out[cnt].pos, q); copy3(out[cnt].normal, nrm); out[cnt].depth = dep; cnt++; }
    }
    return cnt;
  }

  // -------------------------------------------------------------- FT_sensor::Update (ode/ft_sensor.cc:58-88)
  void ftUpdate(FT& s) {
    Body<T>& A = w.body[s.A]; Body<T>& B = w.body[s.B];
    T pA[3], pB[3], vA[3], vB[3];
    A.relPointPos(s.ocA, pA); B.relPointPos(s.ocB, pB);
    A.relPointVel(s.ocA, vA); B.relPointVel(s.ocB, vB);
    for (int k = 0; k < 3; k++) s.F[k] = Klin[k] * (pB[k] - pA[k]) + Clin[k] * (vB[k] - vA[k]);
    // AngleAxisd( qB * qA.inverse() )  (Eigen 3.3 semantics)
    T n2 = A.q[0] * A.q[0] + A.q[1] * A.q[1] + A.q[2] * A.q[2] + A.q[3] * A.q[3];
    T qi[4] = {A.q[0] / n2, -A.q[1] / n2, -A.q[2] / n2, -A.q[3] / n2};
    T qr[4]; QMul0(qr, B.q, qi);
    T n = std::sqrt(qr[1] * qr[1] + qr[2] * qr[2] + qr[3] * qr[3]);
    T angle, axis[3];
    if (n != T(0)) {
      angle = T(2) * tfn::atan2(n, std::fabs(qr[0]));
      if (qr[0] < T(0)) n = -n;
      axis[0] = qr[1] / n; axis[1] = qr[2] / n; axis[2] = qr[3] / n;
    } else { angle = 0; axis[0] = 1; axis[1] = 0; axis[2] = 0; }
    for (int k = 0; k < 3; k++) s.Tq[k] = angle * Kang[k] * axis[k] + Cang[k] * (B.avel[k] - A.avel[k]);
    T nF[3] = {-s.F[0], -s.F[1], -s.F[2]}, nT[3] = {-s.Tq[0], -s.Tq[1], -s.Tq[2]};
    A.addForceAtRelPos(s.F, s.ocA); B.addForceAtRelPos(nF, s.ocB);
    A.addTorque(s.Tq); B.addTorque(nT);
  }

  // -------------------------------------------------------------- Servo (ode/servo.cc:123-167)
  void servoSetDesiredVel(T vel) {
    vel = std::max(-servo_vel_max, vel); vel = std::min(vel, servo_vel_max);
    T angle = w.hingeAngle(w.joint[j_servo]);
    if ((angle <= servo_min && vel < 0) || (angle >= servo_max && vel > 0)) vel = 0;
    servo_vel = vel; servo_mode_vel = true;
    w.joint[j_servo].limot.vel = servo_vel;
  }
  void servoNextStep() {
    T angle = w.hingeAngle(w.joint[j_servo]);
    if (!servo_mode_vel) {  // POS mode, desired angle 0 (never reached after the first robot step)
      T vel = T(0.8) / T((double)sc->cfg.dt) * (T(0) - angle);
      vel = std::max(-servo_vel_max, vel); vel = std::min(vel, servo_vel_max);
      w.joint[j_servo].limot.vel = vel;
    } else if ((angle <= servo_min && servo_vel < 0) || (angle >= servo_max && servo_vel > 0)) {
      servo_vel = 0; w.joint[j_servo].limot.vel = 0;
    }
  }

  // -------------------------------------------------------------- wheel control (src/rover_1.cc:346-400)
  void updateWheelControl() {
    T gamma = w.hingeAngle(w.joint[j_servo]);
    T dgamma_dt = w.hingeAngleRate(w.joint[j_servo]);
    T diff[4], trans[4], min_speed = 0, adjusted_speed;
    const T wb = (T)wheelbase, wt = (T)wheeltrack, rs = (T)robot_speed;
    for (int i = 0; i < 4; i++) {
      diff[i] = (i % 2 ? -1 : 1) * wt / wb * tfn::tan(gamma / 2);
      trans[i] = (i / 2 ? -1 : 1) * (-wb * tfn::tan(gamma / 2) + (i % 2 ? -1 : 1) * wt) * dgamma_dt / 4;
      if (crawling_mode) {
        if (rs >= 0) min_speed = std::max(min_speed, -trans[i] / (1 + diff[i]));
        else min_speed = std::min(min_speed, -trans[i] / (1 + diff[i]));
      }
    }
    if (rs >= 0) adjusted_speed = std::max(rs, min_speed); else adjusted_speed = std::min(rs, min_speed);
    const T WMS = T(7.4351);
    for (int i = 0; i < 4; i++) {
      T r = (T)wheel_radius[i];
      if (rs >= 0) adjusted_speed = std::min(adjusted_speed, (WMS * r - trans[i]) / (1 + diff[i]));
      else adjusted_speed = std::max(adjusted_speed, (-WMS * r - trans[i]) / (1 + diff[i]));
    }
    for (int i = 0; i < 4; i++) W[i] = (adjusted_speed * (1 + diff[i]) + trans[i]) / (T)wheel_radius[i];
  }
  void applyWheelControl() {
    const T WMS = T(7.4351);
    for (int i = 0; i < 4; i++) {
      W[i] = std::min(std::max(-WMS, W[i]), WMS);
      T max_torque = T(25.4) - T(3.416228430014391) * std::fabs(w.hingeAngleRate(w.joint[j_wheel[i]]));
      w.joint[j_wheel[i]].limot.setFMax(max_torque);
      w.joint[j_wheel[i]].limot.vel = W[i];
    }
  }

  // -------------------------------------------------------------- _InternalControl variants
  void emitTransition(bool terminal, double penalty) {
    std::vector<float> row(TFODE_TRANSITION_WIDTH, 0.f);
    row[0] = (float)env_id;
    for (int i = 0; i < 17; i++) row[1 + i] = (float)pending[i];
    row[18] = (float)pending[17]; row[19] = (float)pending[18];
    row[20] = (float)(pending[19] - penalty); row[21] = terminal ? 1.f : 0.f;
    for (int i = 0; i < 17; i++) row[22 + i] = (float)pending[20 + i];
    transitions.push_back(row);
  }
  void internalControl(double delta_t) {
    const TfodeBatchConfig& cfg = sc->cfg;
    tick++;
    switch (cfg.controller) {
      case TFODE_CTRL_RANDOM: {
        uint32_t r[4]; Philox::gen(cfg.seed, env_id, episode, tick, 1, r);
        setSteeringRate((double)Philox::u11(r[0]) * steering_max_vel);
        setBoggieTorque((double)Philox::u11(r[1]) * boggie_max_torque);
      } break;
      case TFODE_CTRL_MLP: {  // src/rover_1_tf.cc:80-199
        T np[3]; getPosition(np);
        double reward = ((double)np[0] - last_pos[0]) * (25. / delta_t);
        reward -= std::fabs(boggie_torque) / boggie_max_torque * 0.5;
        if (fork_collision) { reward -= 1; fork_collision = false; }
        for (int k = 0; k < 3; k++) last_pos[k] = (double)np[k];
        total_reward += reward; last_reward = reward;
        double cur[17]; getState(cur);
        if (has_last_state) {
          if (pending_valid) emitTransition(false, 0);
          for (int i = 0; i < 17; i++) pending[i] = last_state[i];
          pending[17] = steering_rate; pending[18] = boggie_torque; pending[19] = reward;
          for (int i = 0; i < 17; i++) pending[20 + i] = cur[i];
          pending_valid = true;
        }
        static const float scaling[17] = {90, 45, 25, 25, 45, 100, 100, 100, 30, 30, 30, 100, 100, 100, 30, 30, 30};
        float in[17]; for (int i = 0; i < 17; i++) in[i] = (float)cur[i] / scaling[i];
        uint32_t r[4]; Philox::gen(cfg.seed, env_id, episode, tick, 1, r);
        double draw = ((double)Philox::u11(r[0]) + 1) / 2;
        if (exploration && ((!explore && draw > 0.8) || (explore && draw > 0.7))) {
          explore = !explore;
          if (explore) { steering_rate = (double)Philox::u11(r[1]) * steering_max_vel; boggie_torque = (double)Philox::u11(r[2]) * boggie_max_torque; }
        }
        if (!exploration || !explore) {
          float out[2] = {0, 0};
          if (sc->policy.loaded) sc->policy.infer(in, out);
          steering_rate = (double)out[0] * steering_max_vel; boggie_torque = (double)out[1] * boggie_max_torque;
        }
        for (int i = 0; i < 17; i++) last_state[i] = cur[i];
        has_last_state = true;
      } break;
      case TFODE_CTRL_MODEL_TREE: {  // Rover_1_mt::_InternalControl / GetState(false, false) / InferAction, src/rover_1_mt.cc:30-83
        double cur[17]; getState(cur);
        std::vector<double> state(cur, cur + 11);          // direction, steering, roll, pitch, boggie, front F, front T
        state.insert(state.end(), cur + 14, cur + 17);     // rear T (i == 2, the rear forces, is skipped unless full)
        steering_rate = sc->tree[0].predict(state, node_1);
        boggie_torque = sc->tree[1].predict(state, node_2);
      } break;
      case TFODE_CTRL_CRAWLER: {  // src/rover_1_crawlers.cc:40-77
        switch (cr_phase) {
          case 1: boggie_torque += cr_dtorque_dt * delta_t; if (boggie_torque >= cr_amp) { boggie_torque = cr_amp; ++cr_phase; } break;
          case 2: steering_rate = cr_rate; if ((double)getSteeringTrueAngle() >= cr_span) { steering_rate = 0; ++cr_phase; } break;
          case 3: boggie_torque -= cr_dtorque_dt * delta_t; if (boggie_torque <= -cr_amp) { boggie_torque = -cr_amp; ++cr_phase; } break;
          case 4: steering_rate = -cr_rate; if ((double)getSteeringTrueAngle() <= -cr_span) { steering_rate = 0; cr_phase = 1; } break;
        }
      } break;
      default: break;  // Rover_1::_InternalControl only prints (src/rover_1.cc:506-519)
    }
  }

  // -------------------------------------------------------------- Rover_1::next_step (src/rover_1.cc:478-503)
  void robotNextStep(double dt) {
    ftUpdate(ft[0]); ftUpdate(ft[1]);
    for (int s = 0; s < 2; s++) for (int k = 0; k < 3; k++) {  // Q9: in place
      ft[s].F[k] = (T)filt[s * 6 + k].update((double)ft[s].F[k]);
      ft[s].Tq[k] = (T)filt[s * 6 + 3 + k].update((double)ft[s].Tq[k]);
    }
    ic_clock += dt;
    if (ic_activated && ic_clock >= ic_period) { internalControl(ic_clock); ic_clock = 0; ic_tick = true; }
    else ic_tick = false;
    updateWheelControl();
    applyWheelControl();
    servoSetDesiredVel((T)(steering_rate * DEG_TO_RAD));
    w.addHingeTorque(w.joint[j_boggie], (T)std::min(std::max(-boggie_max_torque, boggie_torque), boggie_max_torque));
    servoNextStep();
  }

  // -------------------------------------------------------------- one Sim_loop iteration (src/rover_training_1.cc:127-146)
  // Returns true when the episode ended in this step.
  bool step() {
    const TfodeBatchConfig& cfg = sc->cfg;
    done = false;
    if (frozen) return false;
    const float timestep = cfg.dt;
    const float timef = (float)time_steps * timestep;  // renderer/sim_loop.cc:211 (long*float -> float)
    const double time = (double)timef;
    if (cfg.term > 0 && std::fabs(speed) <= std::fabs(cfg.speedf)) {
      speed += cfg.speedf / cfg.term * timestep;
      setRobotSpeed((double)speed);
    }
    if (!ic_activated && cfg.ic_start >= 0 && time >= (double)ic_start) ic_activated = true;
    collide();
    w.quickstep((T)(double)timestep);
    robotNextStep((double)timestep);
    bool end = false; int reason = 0;
    if (cfg.terminate) {
      T p[3]; getPosition(p);
      if (time >= (double)cfg.timeout) { end = true; reason = TFODE_TERM_TIMEOUT; }
      else if (cfg.y_max > 0 && std::fabs((double)p[1]) >= (double)cfg.y_max) { end = true; reason = TFODE_TERM_Y_MAX; }
      else if ((cfg.x_goal_abs ? std::fabs((double)p[0]) : (double)p[0]) >= (double)cfg.x_goal) { end = true; reason = TFODE_TERM_X_GOAL; }
      else if (cfg.check_upside_down && isUpsideDown()) { end = true; reason = TFODE_TERM_UPSIDE_DOWN; }
      if (end) {
        // epilogue of simulation(), src/rover_training_1.cc:212-224 (penalty precedence: y first, then upside-down)
        if (pending_valid) {
          double pen = 0;
          if (cfg.y_max > 0 && std::fabs((double)p[1]) >= (double)cfg.y_max) pen = 2; else if (cfg.check_upside_down && isUpsideDown()) pen = 5;
          emitTransition(true, pen); pending_valid = false;
        }
        last_ep_return = total_reward; last_ep_x = p[0]; last_ep_y = p[1]; last_ep_steps = time_steps;
      }
    }
    if (end) {
      done = true; term_reason = reason; episodes_done++;
      if (cfg.auto_reset) { int ed = episodes_done; int tr = term_reason; reset(episode + 1); episodes_done = ed; term_reason = tr; done = true; }
      else frozen = true;
      return true;
    }
    time_steps++;
    return false;
  }

  // -------------------------------------------------------------- exchange format
  void getStateX(TfodeEnvState* s) const {
    std::memset(s, 0, sizeof(*s));
    for (size_t b = 0; b < w.body.size(); b++) {
      const Body<T>& B = w.body[b];
      for (int k = 0; k < 3; k++) { s->pos[b][k] = B.pos[k]; s->lvel[b][k] = B.lvel[k]; s->avel[b][k] = B.avel[k]; s->facc[b][k] = B.facc[k]; s->tacc[b][k] = B.tacc[k]; }
      for (int k = 0; k < 4; k++) s->quat[b][k] = B.q[k];
    }
    for (size_t b = w.body.size(); b < TFODE_NBODY; b++) s->quat[b][0] = 1;
    for (int i = 0; i < 4; i++) { s->wheel_vel[i] = w.joint[j_wheel[i]].limot.vel; s->wheel_fmax[i] = w.joint[j_wheel[i]].limot.fmax; s->W[i] = W[i]; }
    s->servo_vel = w.joint[j_servo].limot.vel;
    s->robot_speed = robot_speed; s->steering_rate = steering_rate; s->boggie_torque = boggie_torque;
    s->ic_clock = ic_clock; s->ic_period = ic_period;
    for (int i = 0; i < 12; i++) { s->filt[i][0] = filt[i].y1; s->filt[i][1] = filt[i].y2; s->filt[i][2] = filt[i].x1; s->filt[i][3] = filt[i].x2; }
    for (int k = 0; k < 3; k++) { s->ft[k] = ft[0].F[k]; s->ft[3 + k] = ft[0].Tq[k]; s->ft[6 + k] = ft[1].F[k]; s->ft[9 + k] = ft[1].Tq[k]; s->last_pos[k] = last_pos[k]; }
    s->total_reward = total_reward; s->last_reward = last_reward;
    for (int i = 0; i < 17; i++) s->last_state[i] = last_state[i];
    s->step_yaw_deg = step_yaw_deg; s->crawler_amp = cr_amp; s->crawler_rate = cr_rate; s->crawler_dtorque_dt = cr_dtorque_dt;
    s->speed = speed; s->ic_start = ic_start; s->time_steps = time_steps; s->lcg_seed = w.rand_seed; s->episode = episode; s->tick = tick;
    s->ic_activated = ic_activated; s->ic_tick = ic_tick; s->servo_mode_vel = servo_mode_vel;
    s->has_last_state = has_last_state; s->explore = explore; s->fork_collision = fork_collision; s->crawler_phase = cr_phase;
    s->done = done; s->term_reason = term_reason; s->n_contacts = n_contacts; s->episodes_done = episodes_done;
  }
  void setStateX(const TfodeEnvState* s) {
    if (s->episode != episode) { int ed = episodes_done; reset(s->episode); episodes_done = ed; }  // per-episode scene constants (step yaw, crawler jitter)
    for (size_t b = 0; b < w.body.size(); b++) {
      Body<T>& B = w.body[b];
      for (int k = 0; k < 3; k++) { B.pos[k] = (T)s->pos[b][k]; B.lvel[k] = (T)s->lvel[b][k]; B.avel[k] = (T)s->avel[b][k]; B.facc[k] = (T)s->facc[b][k]; B.tacc[k] = (T)s->tacc[b][k]; }
      for (int k = 0; k < 4; k++) B.q[k] = (T)s->quat[b][k];
      RfromQ(B.R, B.q);
    }
    for (int i = 0; i < 4; i++) { w.joint[j_wheel[i]].limot.vel = (T)s->wheel_vel[i]; w.joint[j_wheel[i]].limot.fmax = (T)s->wheel_fmax[i]; W[i] = (T)s->W[i]; }
    w.joint[j_servo].limot.vel = (T)s->servo_vel; servo_vel = (T)s->servo_vel;
    robot_speed = s->robot_speed; steering_rate = s->steering_rate; boggie_torque = s->boggie_torque;
    ic_clock = s->ic_clock; ic_period = s->ic_period;
    for (int i = 0; i < 12; i++) { filt[i].y1 = s->filt[i][0]; filt[i].y2 = s->filt[i][1]; filt[i].x1 = s->filt[i][2]; filt[i].x2 = s->filt[i][3]; }
    for (int k = 0; k < 3; k++) { ft[0].F[k] = (T)s->ft[k]; ft[0].Tq[k] = (T)s->ft[3 + k]; ft[1].F[k] = (T)s->ft[6 + k]; ft[1].Tq[k] = (T)s->ft[9 + k]; last_pos[k] = s->last_pos[k]; }
    total_reward = s->total_reward; last_reward = s->last_reward;
    for (int i = 0; i < 17; i++) last_state[i] = s->last_state[i];
    speed = s->speed; ic_start = s->ic_start; time_steps = s->time_steps; w.rand_seed = s->lcg_seed; tick = s->tick;
    ic_activated = s->ic_activated != 0; ic_tick = s->ic_tick != 0; servo_mode_vel = s->servo_mode_vel != 0;
    has_last_state = s->has_last_state != 0; explore = s->explore != 0; fork_collision = s->fork_collision != 0; cr_phase = s->crawler_phase;
    done = s->done != 0; term_reason = s->term_reason; n_contacts = s->n_contacts; episodes_done = s->episodes_done;
  }
};

}  // namespace rover_oracle
