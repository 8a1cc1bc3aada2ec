// Trace of the CPU oracle (oracle/rover_oracle.hpp, double precision) on the harness scene -- the file trace_libode.cpp's output is
// compared with.  TEST INFRASTRUCTURE ONLY.
//   trace_oracle <plane|steps> <n_steps> <out.trace>
// Scene = src/scene_1.cc as shipped: Environment(0.5) (+ the two fixed step boxes of :45-53 with "steps"), rover_1 at the origin,
// crawling mode on, internal control off, speed ramp 0.04 m/s over 0.5 s (:68-86), steering rate and boggie torque held at the
// harness's scripted commands (harness_commands.h).
#include <cstring>
#include "../rover_oracle.hpp"
#include "harness_commands.h"
#include "trace_format.h"
using namespace rover_oracle;
int main(int argc, char** argv) {
  if (argc != 4) { fprintf(stderr, "usage: trace_oracle <plane|steps> <n_steps> <out.trace>\n"); return 2; }
  SharedScene sc; std::memset(&sc.cfg, 0, sizeof sc.cfg);
  TfodeBatchConfig& c = sc.cfg;
  c.abi_version = TFODE_ABI_VERSION; c.n_envs = 1; c.scene = std::strcmp(argv[1], "steps") == 0 ? TFODE_SCENE_STEP_BOXES : TFODE_SCENE_PLANE;
  c.controller = TFODE_CTRL_EXTERNAL; c.auto_reset = 0; c.mu = 0.5f; c.dt = 0.001f; c.erp = 0.2f; c.cfm = 1e-10f; c.pgs_iters = 20; c.sor_w = 1.3f;
  c.crawling_mode = 1; c.speedf = 0.04f; c.term = 0.5f; c.ic_start = -1.0f; c.timeout = 1e9f; c.x_goal = 1e9f; c.seed = 20260411ull; c.max_contacts = 24;
  const int n_steps = atoi(argv[2]);
  RoverSim<double> sim; sim.init(&sc, 0);
  FILE* f = fopen(argv[3], "wb"); if (!f) { perror(argv[3]); return 1; }
  const int nb = c.scene == TFODE_SCENE_STEP_BOXES ? 12 : 10;
  trace_write_header(f, nb, n_steps);
  for (int k = 1; k <= n_steps; k++) {
    double steer, torque; harness_commands(k - 1, &steer, &torque);
    sim.setSteeringRate(steer); sim.setBoggieTorque(torque);
    sim.step();
    TfodeEnvState s; sim.getStateX(&s);
    double b[12][13];
    for (int i = 0; i < nb; i++) { for (int j = 0; j < 3; j++) { b[i][j] = s.pos[i][j]; b[i][7 + j] = s.lvel[i][j]; b[i][10 + j] = s.avel[i][j]; } for (int j = 0; j < 4; j++) b[i][3 + j] = s.quat[i][j]; }
    trace_write_step(f, k, s.n_contacts, &b[0][0], nb);
  }
  fclose(f);
  return 0;
}
