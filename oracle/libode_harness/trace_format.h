// Binary trace shared by trace_oracle (the CPU restatement) and trace_libode (the same scene on real libode).
// TEST INFRASTRUCTURE -- nothing under tf_ode_b200/ includes this.
//   header : char magic[4] = "TFTR"; int32 version = 1; int32 n_bodies; int32 n_steps;
//   n_steps x { int32 step; int32 n_contacts; double body[n_bodies][13] }   body = pos(3) quat wxyz(4) linvel(3) angvel(3)
// taken after every { Environment::next_step; Rover_1::next_step } pair (src/scene_1.cc:80-97), step = 1, 2, ...
#pragma once
#include <cstdint>
#include <cstdio>
struct TraceHeader { char magic[4]; int32_t version, n_bodies, n_steps; };
inline void trace_write_header(FILE* f, int n_bodies, int n_steps) { TraceHeader h = {{'T', 'F', 'T', 'R'}, 1, n_bodies, n_steps}; fwrite(&h, sizeof h, 1, f); }
inline void trace_write_step(FILE* f, int step, int n_contacts, const double* bodies13, int n_bodies) {
  int32_t hd[2] = {step, n_contacts}; fwrite(hd, sizeof hd, 1, f); fwrite(bodies13, sizeof(double) * 13, (size_t)n_bodies, f);
}
