// rover_1_mt-style rollout driver over the batched engine (the drop-in for src/scene_1_mt.cc's main()).
//
//   rover_batch_mt nodisplay [orientation_deg] [policy] [ic_start_offset] [--envs N] [--gpus G] [--steps S]
//                  [--scene steps|groove|plane|<heightmap.png>] [--degree D] [--interaction-only]
//
// argv[1..4] keep the meaning they have in src/scene_1_mt.cc:25-48,78-84 (display option, step orientation,
// policy parameter file, IC start offset).  `policy`:
//   a path prefix P such that P1.yaml and P2.yaml exist -> Rover_1_mt with those two model trees, exactly the
//       reference's argv[3] (scene_1_mt.cc:25-30; default there "../scripts/tree_params2_");
//   "random" (default here; BASELINE config 2) -> U(-15,15) deg/s and U(-25,25) N.m at every IC tick;
//   a directory holding variables/variables.index -> the actor SavedModel of scripts/rover_training_1.py:90,121
//       (Rover_1_tf's `actor_path`, src/rover_1_tf.cc:24), read without TensorFlow;
//   any other path -> a raw fp32 actor blob (272 898 floats); both select the Rover_1_tf MLP controller.
// `--scene file.png` builds the HeightField of scene_1_mt.cc:66 from that image (3 x 3 m at (1,0,-0.3)).
// Environments are sharded over G GPUs by contiguous ranges, one host thread and one handle per GPU; there is no
// data-path collective (environments are independent).  Prints the reference's result line for the first
// environments of rank 0 and the aggregate statistics of all of them.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <thread>

#include "tfode_host.hh"

namespace {
// rock-groove stand-in (the PNGs stay in the reference tree): plateau at 0.3 m with a groove, 200 x 200, 8-bit levels
std::vector<float> groove_map(int n) {
  std::vector<float> h((size_t)n * n);
  for (int iz = 0; iz < n; iz++) for (int ix = 0; ix < n; ix++) {
    double x = ix / double(n - 1), y = iz / double(n - 1);
    double centre = 0.45 + 0.12 * y + 0.03 * std::sin(y * 9.0), w = 0.035 + 0.01 * std::sin(y * 5.0);
    double d = std::fabs(x - centre) / w, dip = std::fmin(std::fmax(1.6 - d, 0.0), 1.0);
    double lvl = std::round(255 - dip * 165);
    if (y < 0.05 || y > 0.97) lvl = 255;
    h[(size_t)iz * n + ix] = (float)(lvl * (0.3 / 255.0));
  }
  return h;
}
struct Result { long n = 0, success = 0, episodes = 0; double sum_x = 0, sum_y = 0, seconds = 0; std::string lines, error; };
}  // namespace

int main(int argc, char* argv[]) {
  double orientation = 0, ic_offset = 0; std::string policy = "random", scene = "steps";
  int n_envs = 4096, n_gpus = 1, n_steps = 10000, degree = 1; bool interaction_only = false;
  std::vector<std::string> pos;
  for (int i = 1; i < argc; i++) {
    std::string a = argv[i];
    auto need = [&](const char* what) { if (i + 1 >= argc) throw std::runtime_error(std::string("missing value for ") + what); return argv[++i]; };
    if (a == "--envs") n_envs = atoi(need("--envs"));
    else if (a == "--gpus") n_gpus = atoi(need("--gpus"));
    else if (a == "--steps") n_steps = atoi(need("--steps"));
    else if (a == "--scene") scene = need("--scene");
    else if (a == "--degree") degree = atoi(need("--degree"));
    else if (a == "--interaction-only") interaction_only = true;
    else pos.push_back(a);
  }
  if (pos.size() > 0 && pos[0] != "nodisplay") { fprintf(stderr, "display/capture: unavailable on this box (no graphics stack); running headless\n"); }
  if (pos.size() > 1) { char* e; orientation = strtod(pos[1].c_str(), &e); if (*e) throw std::runtime_error("Invalide orientation: " + pos[1]); }
  if (pos.size() > 2) policy = pos[2];
  if (pos.size() > 3) { char* e; ic_offset = strtod(pos[3].c_str(), &e); if (*e) throw std::runtime_error("Invalide starting offset: " + pos[3]); }
  if (n_envs < 1 || n_gpus < 1 || n_steps < 1) throw std::runtime_error("--envs/--gpus/--steps must be positive");
  std::vector<float> weights;
  bool trees = false;
  if (policy != "random") { std::ifstream t1(policy + "1.yaml"), t2(policy + "2.yaml"); trees = (bool)t1 && (bool)t2; }
  if (policy != "random" && !trees && (bool)std::ifstream(policy + "/variables/variables.index")) {
    float* w = nullptr; size_t n = 0;
    if (tfode_savedmodel_read_actor(policy.c_str(), &w, &n) != TFODE_OK) throw std::runtime_error(tfode_last_error(nullptr));
    weights.assign(w, w + n); tfode_free(w);
  } else if (policy != "random" && !trees) {
    std::ifstream f(policy, std::ios::binary);
    if (!f) throw std::runtime_error("cannot open actor blob " + policy);
    f.seekg(0, std::ios::end); size_t bytes = (size_t)f.tellg(); f.seekg(0);
    weights.resize(bytes / 4); f.read((char*)weights.data(), (std::streamsize)(weights.size() * 4));
  }
  const float x_goal = 1.0f, y_max = 0.6f;  // src/scene_1_mt.cc:88-90
  std::vector<Result> res((size_t)n_gpus);
  std::vector<std::thread> th;
  const int per = (n_envs + n_gpus - 1) / n_gpus;
  for (int g = 0; g < n_gpus; g++) th.emplace_back([&, g]() {
    Result& R = res[(size_t)g];
    try {
      const int first = g * per, n = std::min(per, n_envs - first);
      if (n <= 0) return;
      ode::Environment env(0.5, n, g);
      std::unique_ptr<robot::Rover_1> robot;
      if (trees) robot.reset(new robot::Rover_1_mt(env, ode::Vector3d(0, 0, 0), policy + "1.yaml", policy + "2.yaml", false, (unsigned)degree, interaction_only));
      else if (weights.empty()) { robot.reset(new robot::Rover_1(env, ode::Vector3d(0, 0, 0))); env.config().controller = TFODE_CTRL_RANDOM; }
      else { auto* r = new robot::Rover_1_tf(env, ode::Vector3d(0, 0, 0), weights.data(), weights.size()); r->SetExploration(false); robot.reset(r); }
      robot->SetCrawlingMode(true); robot->SetCmdPeriod(0.5); robot->DeactivateIC();        // scene_1_mt.cc:31-34
      robot->UseBuiltInSpeedRamp(0.04f, 0.5f); robot->SetICStart((float)(1.0 + ic_offset));  // :72-84
      TfodeBatchConfig& c = env.config();
      c.env_id_offset = first; c.timeout = 60; c.x_goal = x_goal; c.y_max = y_max; c.x_goal_abs = 1; c.check_upside_down = 1; c.terminate = 1; c.auto_reset = 0;
      std::unique_ptr<ode::Box> step, step_c; std::unique_ptr<ode::HeightField> field; std::vector<float> hm;
      if (scene == "steps") {
        const float sh = (float)(0.105 * 2);
        step.reset(new ode::Box(env, ode::Vector3d(1, 0, sh / 2), 1, 1, 3, sh, false)); step->set_rotation(0, 0, orientation * M_PI / 180); step->fix(); step->set_collision_group("ground");
        step_c.reset(new ode::Box(env, ode::Vector3d(2, 0, sh / 2), 1, 2, 3, sh, false)); step_c->fix(); step_c->set_collision_group("ground");
      } else if (scene == "groove") {
        hm = groove_map(200);
        field.reset(new ode::HeightField(env, ode::Vector3d(1, 0, -0.3), hm.data(), 200, 200, 3, 3)); field->set_collision_group("ground");
      } else if (scene.size() > 4 && scene.substr(scene.size() - 4) == ".png") {   // scene_1_mt.cc:66
        field.reset(new ode::HeightField(env, ode::Vector3d(1, 0, -0.3), scene.c_str(), 0.3, 3, 3)); field->set_collision_group("ground");
      } else if (scene != "plane") throw std::runtime_error("unknown --scene " + scene);
      Sim_loop sim(0.001f);
      auto t0 = std::chrono::steady_clock::now();
      sim.run(env, n_steps);
      tfode_host::check(tfode_batch_sync(env.batch()), env.batch(), "sync");
      R.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
      // one download for the statistics of every environment (the packed rollout rows), full state only for the four printed lines
      std::vector<float> rows((size_t)n * TFODE_ROLLOUT_WIDTH);
      {
        float* d = nullptr;
        if (cudaMalloc((void**)&d, rows.size() * sizeof(float)) != cudaSuccess) throw std::runtime_error("cudaMalloc failed");
        int rc = tfode_batch_pack_rollout(env.batch(), d);
        if (rc == TFODE_OK) rc = tfode_batch_sync(env.batch());
        if (rc == TFODE_OK && cudaMemcpy(rows.data(), d, rows.size() * sizeof(float), cudaMemcpyDeviceToHost) != cudaSuccess) { cudaFree(d); throw std::runtime_error("cudaMemcpy failed"); }
        cudaFree(d);
        tfode_host::check(rc, env.batch(), "tfode_batch_pack_rollout");
      }
      for (int e = 0; e < n; e++) {
        ode::Vector3d p = robot->GetPosition(e);
        const bool ok = std::fabs(p.x()) >= x_goal;
        R.n++; R.success += ok; R.episodes += (long)rows[(size_t)e * TFODE_ROLLOUT_WIDTH + 21]; R.sum_x += p.x(); R.sum_y += p.y();
        if (g == 0 && e < 4) {
          TfodeEnvState s; tfode_host::check(tfode_batch_get_state(env.batch(), e, &s), env.batch(), "get_state");
          char buf[200];
          snprintf(buf, sizeof buf, "%s t %6.3f | x %5.3f | y %+6.3f\n", ok ? "\033[1;32m[Success]\033[0;39m" : "\033[1;31m[Failure]\033[0;39m",
                   (double)((float)s.time_steps * 0.001f), p.x(), p.y());  // the result line of src/scene_1_mt.cc:191-194
          R.lines += buf;
        }
      }
    } catch (const std::exception& ex) { R.error = ex.what(); }
  });
  for (auto& t : th) t.join();
  Result T; double tmax = 0;
  for (auto& r : res) { if (!r.error.empty()) { fprintf(stderr, "error: %s\n", r.error.c_str()); return 1; } T.n += r.n; T.success += r.success; T.episodes += r.episodes; T.sum_x += r.sum_x; T.sum_y += r.sum_y; tmax = std::max(tmax, r.seconds); }
  fputs(res[0].lines.c_str(), stdout);
  printf("envs %ld | gpus %d | steps %d | finished episodes %ld | success %ld | mean x %.3f | mean y %+.3f | %.3e env-steps/s (wall, max over GPUs)\n",
         T.n, n_gpus, n_steps, T.episodes, T.success, T.sum_x / T.n, T.sum_y / T.n, (double)T.n * n_steps / tmax);
  return 0;
}
