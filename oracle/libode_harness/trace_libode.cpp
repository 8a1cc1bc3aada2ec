// WITH_LIBODE cross-check harness: the oracle's scene on REAL libode, dumping the trace format of trace_oracle.cpp.
// TEST INFRASTRUCTURE ONLY -- never linked into the product.  Needs libode (double precision, 0.13 / 0.14):
//     g++ -O2 -std=c++11 -DdDOUBLE trace_libode.cpp $(pkg-config --cflags --libs ode) -o trace_libode
//     ./trace_libode <plane|steps> <n_steps> <out.trace> [worldstep]
// libode is not installed in the build container ("unavailable on this box"): this program exists so that anyone who HAS libode
// can pin oracle/ode_restated.hpp against it (compare_traces.py).  The solver is dWorldQuickStep, i.e. ode/environment.hh:132-133
// with the two lines swapped; pass `worldstep` as 4th argument to run dWorldStep (what the shipped binaries call) instead.
//
// The scene restates, call for call, what the reference executes for src/scene_1.cc as shipped:
//   Environment(0.5)                      ode/environment.cc:29-55   world, gravity, hash space, ground plane "ground", contact group
//   Rover_1(env, (0,0,0))                 src/rover_1.cc:18-245      10 bodies, 7 joints, geoms, 2 FT spring-dampers
//   step / step_c boxes + fix()           src/scene_1.cc:45-53, ode/object.cc:155-179  ("steps" only)
//   per step                              src/scene_1.cc:80-97: speed ramp, env.next_step() (ode/environment.hh:127-136,
//                                         ode/environment.cc:56-150), robot.next_step() (src/rover_1.cc:478-503) with the internal
//                                         control off and the commands of harness_commands.h.
// The 12 F/T biquads only feed observations, not forces, and are left out.
#include <ode/ode.h>

#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "harness_commands.h"
#include "trace_format.h"

struct Feature { const char* group; bool soft; };   // ode/object.hh collision_feature: group + contact type (callbacks unused here)
static Feature F_GROUND = {"ground", false}, F_SOFT = {"", true};
static dWorldID world; static dSpaceID space; static dJointGroupID contacts; static int n_contacts_step;
static const double MU = 0.5;

static void near_cb(void*, dGeomID o1, dGeomID o2) {   // Environment::_collision, ode/environment.cc:56-150
  Feature* f1 = (Feature*)dGeomGetData(o1); Feature* f2 = (Feature*)dGeomGetData(o2);
  bool soft = false;
  if (!f1 && !f2) return;
  if (f1 && f2) { if (std::strcmp(f1->group, f2->group) == 0) return; soft = f1->soft || f2->soft; }
  else { Feature* f = f1 ? f1 : f2; if (*f->group == '\0') return; soft = f->soft; }
  dContact c[10];
  const int n = dCollide(o1, o2, 10, &c[0].geom, sizeof(dContact));
  for (int i = 0; i < n; i++) {
    c[i].surface.mode = dContactApprox1; c[i].surface.mu = MU;
    if (soft) { c[i].surface.mode |= dContactSoftCFM | dContactSoftERP; c[i].surface.soft_cfm = 0.005; c[i].surface.soft_erp = 0.4; }
    dJointID j = dJointCreateContact(world, contacts, &c[i]);
    dJointAttach(j, dGeomGetBody(c[i].geom.g1), dGeomGetBody(c[i].geom.g2));
  }
  n_contacts_step += n;
}

static dBodyID make_box(double x, double y, double z, double m, double lx, double ly, double lz, bool geom, void* data) {   // ode/box.hh:33-62
  dBodyID b = dBodyCreate(world); dBodySetPosition(b, x, y, z);
  dMass mm; dMassSetBoxTotal(&mm, m, lx, ly, lz); dBodySetMass(b, &mm);
  if (geom) { dGeomID g = dCreateBox(space, lx, ly, lz); dGeomSetBody(g, b); dGeomSetData(g, data); }
  return b;
}
static dJointID hinge(dBodyID a, dBodyID b, double ax, double ay, double az, double x, double y, double z) {
  dJointID j = dJointCreateHinge(world, 0); dJointAttach(j, a, b); dJointSetHingeAnchor(j, ax, ay, az); dJointSetHingeAxis(j, x, y, z); return j;
}

struct FT { dBodyID A, B; dVector3 ocA, ocB; };
static void ft_init(FT& s, dBodyID A, dBodyID B, double cx, double cy, double cz) { s.A = A; s.B = B; dBodyGetPosRelPoint(A, cx, cy, cz, s.ocA); dBodyGetPosRelPoint(B, cx, cy, cz, s.ocB); }
static void ft_update(FT& s) {   // ode/ft_sensor.cc:58-88, K_lin (2e4,2e4,1e5), K_ang 1e3, C_lin 5e2, C_ang 1 (src/rover_1.cc:44-47,205-206)
  dVector3 pA, pB, vA, vB;
  dBodyGetRelPointPos(s.A, s.ocA[0], s.ocA[1], s.ocA[2], pA); dBodyGetRelPointPos(s.B, s.ocB[0], s.ocB[1], s.ocB[2], pB);
  dBodyGetRelPointVel(s.A, s.ocA[0], s.ocA[1], s.ocA[2], vA); dBodyGetRelPointVel(s.B, s.ocB[0], s.ocB[1], s.ocB[2], vB);
  const double Kl[3] = {2e4, 2e4, 1e5};
  double F[3], T[3];
  for (int k = 0; k < 3; k++) F[k] = Kl[k] * (pB[k] - pA[k]) + 5e2 * (vB[k] - vA[k]);
  const dReal* qA = dBodyGetQuaternion(s.A); const dReal* qB = dBodyGetQuaternion(s.B);
  // Eigen: AngleAxisd( qB * qA.inverse() ); inverse = conjugate / squaredNorm
  const double n2 = qA[0] * qA[0] + qA[1] * qA[1] + qA[2] * qA[2] + qA[3] * qA[3];
  const double iw = qA[0] / n2, ix = -qA[1] / n2, iy = -qA[2] / n2, iz = -qA[3] / n2;
  const double w = qB[0] * iw - qB[1] * ix - qB[2] * iy - qB[3] * iz, x = qB[0] * ix + qB[1] * iw + qB[2] * iz - qB[3] * iy;
  const double y = qB[0] * iy + qB[2] * iw + qB[3] * ix - qB[1] * iz, z = qB[0] * iz + qB[3] * iw + qB[1] * iy - qB[2] * ix;
  double n = std::sqrt(x * x + y * y + z * z), angle = 0, axis[3] = {1, 0, 0};
  if (n != 0) { angle = 2 * std::atan2(n, std::fabs(w)); if (w < 0) n = -n; axis[0] = x / n; axis[1] = y / n; axis[2] = z / n; }
  const dReal* wA = dBodyGetAngularVel(s.A); const dReal* wB = dBodyGetAngularVel(s.B);
  for (int k = 0; k < 3; k++) T[k] = angle * 1e3 * axis[k] + 1.0 * (wB[k] - wA[k]);
  dBodyAddForceAtRelPos(s.A, F[0], F[1], F[2], s.ocA[0], s.ocA[1], s.ocA[2]);
  dBodyAddForceAtRelPos(s.B, -F[0], -F[1], -F[2], s.ocB[0], s.ocB[1], s.ocB[2]);
  dBodyAddTorque(s.A, T[0], T[1], T[2]); dBodyAddTorque(s.B, -T[0], -T[1], -T[2]);
}

int main(int argc, char** argv) {
  if (argc < 4) { fprintf(stderr, "usage: trace_libode <plane|steps> <n_steps> <out.trace> [worldstep]\n"); return 2; }
  const bool steps_scene = std::strcmp(argv[1], "steps") == 0, world_step = argc > 4 && std::strcmp(argv[4], "worldstep") == 0;
  const int n_steps = atoi(argv[2]);
  dInitODE2(0);
  world = dWorldCreate(); dWorldSetGravity(world, 0, 0, -9.80665);
  space = dHashSpaceCreate(0); contacts = dJointGroupCreate(0);
  dGeomSetData(dCreatePlane(space, 0, 0, 1, 0), &F_GROUND);

  // ---- Rover_1 (src/rover_1.cc:49-233): dimensions as in the constructor
  const double wheelbase = 0.58, wheeltrack = 0.61, wheel_r = 0.105, wheel_w = 0.1, belly = 0.33;
  const double front_len = 0.345, front_h = 0.225, front_w = 0.200, fx = front_len / 2 - 0.063, fy = 0.023, fz = belly + front_h / 2;
  const double rear_len = 0.312, rear_h = 0.177, rear_w = 0.126, rx = -rear_len / 2 + 0.046, ry = -0.040, rz = belly + rear_h / 2;
  const double sea_elev = belly + 0.12, boggie_h = sea_elev - belly, fork_elev = belly - 0.1 / 2, rim_r = wheel_r - wheel_w / 2;
  std::vector<dBodyID> B;
  B.push_back(make_box(fx, fy, fz, 5.25, front_len, front_w, front_h, true, 0));                          // 0 main
  B.push_back(make_box(0.285, 0, belly + 0.155, 3, 0.0975, 0.151, 0.065, true, 0));                       // 1 battery
  B.push_back(make_box(rx, ry, rz, 3.67, rear_len, rear_w, rear_h, true, 0));                             // 2 rear
  B.push_back(make_box(-wheelbase / 2 + 0.035, 0, belly + boggie_h / 2, 1, 0.1, 0.1, boggie_h, false, 0)); // 3 boggie (no geom)
  dMatrix3 Rw; dRFromEulerAngles(Rw, M_PI / 2, 0, 0);
  for (int s = 0; s < 2; s++) {                                                                            // 4 front fork, 5 rear fork
    const double sx = s == 0 ? 1 : -1;
    dBodyID b = make_box(sx * wheelbase / 2, 0, fork_elev, 1, 0.05, 0.3, 0.1, false, 0);
    for (int k = 0; k < 2; k++) {  // motor bulks: cylinders r .025 l .1 with axis along y (:153-163)
      dGeomID g = dCreateCylinder(space, 0.025, 0.1); dGeomSetBody(g, b);
      dGeomSetOffsetWorldPosition(g, sx * wheelbase / 2, (k == 0 ? 1 : -1) * (wheeltrack - wheel_w - 0.1) / 2, wheel_r);
      dGeomSetOffsetRotation(g, Rw);
    }
    B.push_back(b);
  }
  dJointID j_clamp = dJointCreateSlider(world, 0); dJointAttach(j_clamp, B[1], B[0]); dJointSetSliderAxis(j_clamp, 0, 1, 0);   // :128-132
  dJointSetSliderParam(j_clamp, dParamLoStop, 0); dJointSetSliderParam(j_clamp, dParamHiStop, 0);
  dJointID j_servo = hinge(B[2], B[0], 0, 0, (fz + rz) / 2, 0, 0, 1);                                     // :181-189, ode/servo.cc:12-25
  dJointSetHingeParam(j_servo, dParamFMax, 180);
  dJointID j_boggie = hinge(B[3], B[2], -wheelbase / 2, 0, sea_elev, 1, 0, 0);                             // :194-200 (axis set before the anchor there)
  dJointSetHingeParam(j_boggie, dParamLoStop, -45 * M_PI / 180); dJointSetHingeParam(j_boggie, dParamHiStop, 45 * M_PI / 180);
  dJointID j_wheel[4];
  for (int i = 0; i < 4; i++) {                                                                            // :209-233, ode/wheel.hh:33-74
    const double wx = (i / 2 ? -1 : 1) * wheelbase / 2, wy = (i % 2 ? -1 : 1) * wheeltrack / 2;
    dBodyID b = dBodyCreate(world); dBodySetPosition(b, wx, wy, wheel_r);
    dMass mm; dMassSetCylinderTotal(&mm, 1.4 + 1, 3, rim_r, wheel_w); dBodySetMass(b, &mm);
    dGeomID rim = dCreateCylinder(space, rim_r, wheel_w); dGeomSetBody(rim, b);
    for (int k = 0; k < 50; k++) {
      dGeomID g = dCreateSphere(space, wheel_w / 2); dGeomSetBody(g, b);
      dGeomSetOffsetPosition(g, rim_r * std::cos(k * 2 * M_PI / 50), rim_r * std::sin(k * 2 * M_PI / 50), 0);
      if (k == 49) dGeomSetData(g, &F_SOFT);   // Q1: set_contact_type(SOFT) reaches only the last geom
    }
    dBodySetRotation(b, Rw);
    B.push_back(b);
    j_wheel[i] = hinge(wx > 0 ? B[4] : B[5], b, wx, wy, wheel_r, 0, -1, 0);
    dJointSetHingeParam(j_wheel[i], dParamFMax, 25.4);
  }
  FT ft[2]; ft_init(ft[0], B[0], B[4], wheelbase / 2, 0, belly); ft_init(ft[1], B[3], B[5], -wheelbase / 2, 0, belly);
  if (steps_scene) {                                                                                       // src/scene_1.cc:45-53
    const float sh = (float)(0.105 * 2);
    for (int s = 0; s < 2; s++) {
      dBodyID b = make_box(s == 0 ? 1 : 2, 0, sh / 2, 1, s == 0 ? 1 : 2, 3, sh, true, &F_GROUND);
      dJointID f = dJointCreateSlider(world, 0); dJointAttach(f, b, 0); dJointSetSliderAxis(f, 0, 0, 1);   // Object::fix
      dJointSetSliderParam(f, dParamLoStop, 0); dJointSetSliderParam(f, dParamHiStop, 0);
      B.push_back(b);
    }
  }

  FILE* out = fopen(argv[3], "wb"); if (!out) { perror(argv[3]); return 1; }
  const int nb = (int)B.size();
  trace_write_header(out, nb, n_steps);
  const float timestep = 0.001f; float speed = 0, speedf = 0.04f, term = 0.5f;
  const double robot_max_speed = 7.4351 * wheel_r;   // Q4
  double robot_speed = 0, servo_vel = 0;
  for (int k = 1; k <= n_steps; k++) {
    if (std::fabs(speed) <= std::fabs(speedf)) { speed += speedf / term * timestep; robot_speed = std::fmin(std::fmax(-robot_max_speed, (double)speed), robot_max_speed); }
    // env.next_step (ode/environment.hh:127-136)
    n_contacts_step = 0;
    dSpaceCollide(space, 0, &near_cb);
    if (world_step) dWorldStep(world, timestep); else dWorldQuickStep(world, timestep);
    dJointGroupEmpty(contacts);
    // robot.next_step (src/rover_1.cc:478-503), internal control off
    ft_update(ft[0]); ft_update(ft[1]);
    double steering_rate, boggie_torque; harness_commands(k - 1, &steering_rate, &boggie_torque);
    steering_rate = std::fmin(std::fmax(-15.0, steering_rate), 15.0); boggie_torque = std::fmin(std::fmax(-25.0, boggie_torque), 25.0);
    {  // _UpdateWheelControl + _ApplyWheelControl (:346-400), crawling mode on
      const double gamma = dJointGetHingeAngle(j_servo), dgamma = dJointGetHingeAngleRate(j_servo), WMS = 7.4351;
      double diff[4], trans[4], min_speed = 0, adj;
      for (int i = 0; i < 4; i++) {
        diff[i] = (i % 2 ? -1 : 1) * wheeltrack / wheelbase * std::tan(gamma / 2);
        trans[i] = (i / 2 ? -1 : 1) * (-wheelbase * std::tan(gamma / 2) + (i % 2 ? -1 : 1) * wheeltrack) * dgamma / 4;
        if (robot_speed >= 0) min_speed = std::fmax(min_speed, -trans[i] / (1 + diff[i])); else min_speed = std::fmin(min_speed, -trans[i] / (1 + diff[i]));
      }
      adj = robot_speed >= 0 ? std::fmax(robot_speed, min_speed) : std::fmin(robot_speed, min_speed);
      for (int i = 0; i < 4; i++) adj = robot_speed >= 0 ? std::fmin(adj, (WMS * wheel_r - trans[i]) / (1 + diff[i])) : std::fmax(adj, (-WMS * wheel_r - trans[i]) / (1 + diff[i]));
      for (int i = 0; i < 4; i++) {
        double W = (adj * (1 + diff[i]) + trans[i]) / wheel_r; W = std::fmin(std::fmax(-WMS, W), WMS);
        dJointSetHingeParam(j_wheel[i], dParamFMax, 25.4 - 3.416228430014391 * std::fabs(dJointGetHingeAngleRate(j_wheel[i])));
        dJointSetHingeParam(j_wheel[i], dParamVel, W);
      }
      // Servo::set_desired_vel + next_step (ode/servo.cc:123-167), VEL mode
      const double vmax = 15 * M_PI / 180, smin = -45 * M_PI / 180, smax = 45 * M_PI / 180;
      double vel = steering_rate * M_PI / 180; vel = std::fmax(-vmax, vel); vel = std::fmin(vel, vmax);
      if ((gamma <= smin && vel < 0) || (gamma >= smax && vel > 0)) vel = 0;
      servo_vel = vel; dJointSetHingeParam(j_servo, dParamVel, servo_vel);
      dJointAddHingeTorque(j_boggie, boggie_torque);   // _ApplyBoggieControl (:411-416)
    }
    std::vector<double> row((size_t)nb * 13);
    for (int i = 0; i < nb; i++) {
      const dReal *p = dBodyGetPosition(B[i]), *q = dBodyGetQuaternion(B[i]), *v = dBodyGetLinearVel(B[i]), *w = dBodyGetAngularVel(B[i]);
      for (int j = 0; j < 3; j++) { row[i * 13 + j] = p[j]; row[i * 13 + 7 + j] = v[j]; row[i * 13 + 10 + j] = w[j]; }
      for (int j = 0; j < 4; j++) row[i * 13 + 3 + j] = q[j];
    }
    trace_write_step(out, k, n_contacts_step, row.data(), nb);
  }
  fclose(out);
  dJointGroupDestroy(contacts); dSpaceDestroy(space); dWorldDestroy(world); dCloseODE();
  return 0;
}
