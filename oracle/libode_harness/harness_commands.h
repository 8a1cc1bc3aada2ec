// Scripted commands of the cross-check scene (BASELINE config 1: "scripted servo commands"): piecewise-constant steering rate
// [deg/s] and boggie torque [N.m], changed every 500 steps, from a fixed table so that both programs need no RNG.
#pragma once
inline void harness_commands(int step_index, double* steering_rate_deg_s, double* boggie_torque) {
  static const double tab[8][2] = {{0, 0}, {10, 5}, {-15, -20}, {5, 25}, {0, -10}, {-10, 15}, {15, 0}, {-5, -25}};
  const int k = (step_index / 500) % 8;
  *steering_rate_deg_s = tab[k][0]; *boggie_torque = tab[k][1];
}
