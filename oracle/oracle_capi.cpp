// =============================================================================
// CPU ORACLE -- TEST INFRASTRUCTURE ONLY (see ode_restated.hpp header).
// C entry points for tests/ (ctypes), __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs.  N environments are partitioned
// statically over T std::threads (the reference itself is single-threaded, one
// world per process: BASELINE.md section 4).
// =============================================================================
#include <atomic>
#include <cstdio>
#include <memory>
#include <thread>

#include "rover_oracle.hpp"

using namespace rover_oracle;

namespace {

struct BatchBase {
  virtual ~BatchBase() {}
  virtual void step(int n) = 0;
  virtual void get_state(int env, TfodeEnvState* s) = 0;
  virtual void set_state(int env, const TfodeEnvState* s) = 0;
  virtual void set_actions(const float* a) = 0;
  virtual void set_speed(const float* v) = 0;
  virtual void observe(float* obs, float* rew, uint8_t* done, int32_t* nc) = 0;
  virtual int drain(float* out, int max_rows) = 0;
  virtual void reset(const int32_t* ids, int n) = 0;
  virtual int last_rows(int env) = 0;
  virtual int max_pair_contacts() = 0;
  virtual int lcp_failures() = 0;
  virtual void contact_hist(uint64_t* out129) = 0;
  SharedScene scene;
  int n_threads = 1;
};

template <class T> struct Batch : BatchBase {
  std::vector<RoverSim<T>> envs;
  std::vector<uint8_t> done_call;
  void create() {
    envs.resize(scene.cfg.n_envs);
    done_call.assign(scene.cfg.n_envs, 0);
    for (int i = 0; i < scene.cfg.n_envs; i++) envs[i].init(&scene, (uint32_t)(scene.cfg.env_id_offset + i));
  }
  void step(int n) override {
    const int N = (int)envs.size();
    std::fill(done_call.begin(), done_call.end(), 0);
    auto work = [&](int lo, int hi) {
      for (int e = lo; e < hi; e++) for (int s = 0; s < n; s++) if (envs[e].step()) done_call[e] = 1;
    };
    int Tn = std::max(1, std::min(n_threads, N));
    if (Tn == 1) { work(0, N); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < Tn; t++) { int lo = (int)((int64_t)N * t / Tn), hi = (int)((int64_t)N * (t + 1) / Tn); th.emplace_back(work, lo, hi); }
    for (auto& t : th) t.join();
  }
  void get_state(int env, TfodeEnvState* s) override { envs[env].getStateX(s); }
  void set_state(int env, const TfodeEnvState* s) override { envs[env].setStateX(s); }
  void set_actions(const float* a) override {
    for (size_t e = 0; e < envs.size(); e++) { envs[e].setSteeringRate((double)a[e * 2]); envs[e].setBoggieTorque((double)a[e * 2 + 1]); }
  }
  void set_speed(const float* v) override { for (size_t e = 0; e < envs.size(); e++) envs[e].setRobotSpeed((double)v[e]); }
  void observe(float* obs, float* rew, uint8_t* done, int32_t* nc) override {
    for (size_t e = 0; e < envs.size(); e++) {
      if (obs) { double s[17]; envs[e].getState(s); for (int i = 0; i < 17; i++) obs[e * 17 + i] = (float)s[i]; }
      if (rew) rew[e] = (float)envs[e].last_reward;
      if (done) done[e] = done_call[e];
      if (nc) nc[e] = envs[e].n_contacts;
    }
  }
  int drain(float* out, int max_rows) override {
    int n = 0;
    for (auto& e : envs) {
      for (auto& r : e.transitions) { if (n < max_rows) { std::memcpy(out + (size_t)n * TFODE_TRANSITION_WIDTH, r.data(), sizeof(float) * TFODE_TRANSITION_WIDTH); } n++; }
      e.transitions.clear();
    }
    return n;
  }
  void reset(const int32_t* ids, int n) override {
    if (!ids) { for (auto& e : envs) { e.reset(e.episode + 1); } return; }
    for (int i = 0; i < n; i++) envs[ids[i]].reset(envs[ids[i]].episode + 1);
  }
  int last_rows(int env) override { return envs[env].w.last_rows; }
  void contact_hist(uint64_t* out) override { for (int i = 0; i < 129; i++) out[i] = 0; for (auto& e : envs) for (int i = 0; i < 129; i++) out[i] += e.contact_hist[i]; }
  int lcp_failures() override { int m = 0; for (auto& e : envs) m += e.w.lcp_failures; return m; }
  int max_pair_contacts() override { int m = 0; for (auto& e : envs) m = std::max(m, e.max_pair_contacts); return m; }
};

}  // namespace

extern "C" {

// precision: 0 = float (GPU-equivalent arithmetic), 1 = double (reference-equivalent)
void* orc_create(const TfodeBatchConfig* cfg, int precision, int n_threads) {
  BatchBase* b = nullptr;
  if (precision == 0) b = new Batch<float>(); else b = new Batch<double>();
  b->scene.cfg = *cfg; b->n_threads = n_threads;
  if (cfg->scene == TFODE_SCENE_HEIGHTFIELD) {
    if (!cfg->hf_heights || cfg->hf_ncol < 2 || cfg->hf_nrow < 2) { delete b; return nullptr; }
    b->scene.hf_data.assign(cfg->hf_heights, cfg->hf_heights + (size_t)cfg->hf_ncol * cfg->hf_nrow);
    b->scene.cfg.hf_heights = nullptr;
  }
  if (precision == 0) static_cast<Batch<float>*>(b)->create(); else static_cast<Batch<double>*>(b)->create();
  return b;
}
void orc_destroy(void* h) { delete static_cast<BatchBase*>(h); }
int orc_load_policy(void* h, const float* w, size_t n) { auto* b = static_cast<BatchBase*>(h); b->scene.policy.load(w, n); return b->scene.policy.loaded ? 0 : -1; }
// Model trees from per-id entries (what YAML::LoadFile gives the reference): ids[k], terminal[k], params[off[k] .. off[k]+cnt[k]).
int orc_load_model_tree(void* h, int which, int n_entries, const int32_t* ids, const int32_t* terminal, const int32_t* off, const int32_t* cnt,
                        const double* params, int oblique, int degree, int interaction_only) {
  auto* b = static_cast<BatchBase*>(h);
  std::map<int, ModelTree::Entry> y;
  for (int k = 0; k < n_entries; k++) y[ids[k]] = ModelTree::Entry{terminal[k] != 0, std::vector<double>(params + off[k], params + off[k] + cnt[k])};
  try { b->scene.tree[which & 1].load(y, oblique != 0, (unsigned)degree, interaction_only != 0); } catch (const std::exception&) { return -1; }
  return 0;
}
int orc_model_tree_predict(void* h, int which, const double* x, int n_in, int m, double* y, int32_t* node) {
  auto* b = static_cast<BatchBase*>(h);
  try {
    for (int i = 0; i < m; i++) { int nd = 0; y[i] = b->scene.tree[which & 1].predict(std::vector<double>(x + (size_t)i * n_in, x + (size_t)(i + 1) * n_in), nd); if (node) node[i] = nd; }
  } catch (const std::exception&) { return -1; }
  return 0;
}
void orc_step(void* h, int n) { static_cast<BatchBase*>(h)->step(n); }
void orc_get_state(void* h, int env, TfodeEnvState* s) { static_cast<BatchBase*>(h)->get_state(env, s); }
void orc_set_state(void* h, int env, const TfodeEnvState* s) { static_cast<BatchBase*>(h)->set_state(env, s); }
void orc_set_actions(void* h, const float* a) { static_cast<BatchBase*>(h)->set_actions(a); }
void orc_set_speed(void* h, const float* v) { static_cast<BatchBase*>(h)->set_speed(v); }
void orc_observe(void* h, float* obs, float* rew, uint8_t* done, int32_t* nc) { static_cast<BatchBase*>(h)->observe(obs, rew, done, nc); }
int orc_drain_transitions(void* h, float* out, int max_rows) { return static_cast<BatchBase*>(h)->drain(out, max_rows); }
void orc_reset(void* h, const int32_t* ids, int n) { static_cast<BatchBase*>(h)->reset(ids, n); }
int orc_last_rows(void* h, int env) { return static_cast<BatchBase*>(h)->last_rows(env); }
int orc_max_pair_contacts(void* h) { return static_cast<BatchBase*>(h)->max_pair_contacts(); }
void orc_contact_hist(void* h, uint64_t* out129) { static_cast<BatchBase*>(h)->contact_hist(out129); }
// Dantzig LCP on its own (include/tfode_lcp.h), double: returns the solver's status; x = lambda
int orc_lcp_solve(int n, const double* A, const double* b, const double* lo, const double* hi, const int32_t* findex, double* x) {
  std::vector<double> fw((size_t)tfl::lcp_fwork(n)); std::vector<int> iw((size_t)tfl::lcp_iwork(n));
  return tfl::lcp_dantzig<double>(n, A, x, b, lo, hi, findex, fw.data(), iw.data());
}
int orc_lcp_failures(void* h) { return static_cast<BatchBase*>(h)->lcp_failures(); }
int orc_sizeof_state() { return (int)sizeof(TfodeEnvState); }
int orc_sizeof_config() { return (int)sizeof(TfodeBatchConfig); }

// ---- unit hooks for known-answer tests --------------------------------------
void orc_rand_ints(uint32_t seed, const int32_t* n, int count, int32_t* out, uint32_t* seed_out) {
  for (int i = 0; i < count; i++) out[i] = dRandInt(seed, n[i]);
  if (seed_out) *seed_out = seed;
}
void orc_plane_space(const double* n, double* p, double* q) { planeSpace(n, p, q); }
void orc_philox(uint64_t seed, uint32_t env, uint32_t episode, uint32_t tick, uint32_t stream, uint32_t* out4, float* u4) {
  Philox::gen(seed, env, episode, tick, stream, out4);
  if (u4) for (int i = 0; i < 4; i++) u4[i] = Philox::u11(out4[i]);
}
void orc_biquad_coeffs(double* c5) { Biquad b; b.init(0.001, 4 * M_PI, 0.5); c5[0] = b.Cy1; c5[1] = b.Cy2; c5[2] = b.Cx0; c5[3] = b.Cx1; c5[4] = b.Cx2; }
void orc_biquad_run(const double* x, int n, double* y) { Biquad b; b.init(0.001, 4 * M_PI, 0.5); for (int i = 0; i < n; i++) y[i] = b.update(x[i]); }
void orc_mlp_infer(void* h, const float* x, int m, float* y) { auto* b = static_cast<BatchBase*>(h); for (int i = 0; i < m; i++) b->scene.policy.infer(x + i * 17, y + i * 2); }
// colliders (double)
int orc_sphere_plane(const double* c, double r, const double* n, double d, double* out7) {
  ContactGeom<double> g; int k = collideSpherePlane(c, r, n, d, &g);
  if (k) { for (int i = 0; i < 3; i++) { out7[i] = g.pos[i]; out7[3 + i] = g.normal[i]; } out7[6] = g.depth; }
  return k;
}
int orc_sphere_box(const double* c, double r, const double* bpos, const double* bR, const double* side, double* out7) {
  ContactGeom<double> g; int k = collideSphereBox(c, r, bpos, bR, side, &g);
  if (k) { for (int i = 0; i < 3; i++) { out7[i] = g.pos[i]; out7[3 + i] = g.normal[i]; } out7[6] = g.depth; }
  return k;
}
int orc_cylinder_plane(const double* p, const double* R, double radius, double length, const double* n, double d, double* out /*[4][7]*/) {
  ContactGeom<double> g[4]; int k = collideCylinderPlane(p, R, radius, length, n, d, g, 4);
  for (int j = 0; j < k; j++) { for (int i = 0; i < 3; i++) { out[j * 7 + i] = g[j].pos[i]; out[j * 7 + 3 + i] = g[j].normal[i]; } out[j * 7 + 6] = g[j].depth; }
  return k;
}
int orc_box_plane(const double* p, const double* R, const double* side, const double* n, double d, double* out /*[3][7]*/) {
  ContactGeom<double> g[3]; int k = collideBoxPlane(p, R, side, n, d, g, 3);
  for (int j = 0; j < k; j++) { for (int i = 0; i < 3; i++) { out[j * 7 + i] = g[j].pos[i]; out[j * 7 + 3 + i] = g[j].normal[i]; } out[j * 7 + 6] = g[j].depth; }
  return k;
}
int orc_sphere_hf(const double* c, double r, const float* h, int ncol, int nrow, double length, double width, const double* pos, double* out /*[4][7]*/) {
  HeightField<double> hf; hf.init(h, ncol, nrow, length, width, pos);
  ContactGeom<double> g[4]; int k = collideSphereHF(c, r, hf, g, 4);
  for (int j = 0; j < k; j++) { for (int i = 0; i < 3; i++) { out[j * 7 + i] = g[j].pos[i]; out[j * 7 + 3 + i] = g[j].normal[i]; } out[j * 7 + 6] = g[j].depth; }
  return k;
}
void orc_R_from_euler(double phi, double theta, double psi, double* R9) { RFromEulerAngles(R9, phi, theta, psi); }
void orc_mass_box(double m, double lx, double ly, double lz, double* I3) { Body<double> b; b.setMassBoxTotal(m, lx, ly, lz); for (int i = 0; i < 3; i++) I3[i] = b.I[i]; }
void orc_mass_cyl_z(double m, double r, double l, double* I3) { Body<double> b; b.setMassCylinderTotalZ(m, r, l); for (int i = 0; i < 3; i++) I3[i] = b.I[i]; }

}  // extern "C"
