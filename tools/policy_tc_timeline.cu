// Developer tool: phase timeline of policy_tc_kernel (clock64 at the role hand-overs), built from the library's own header.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -DTFODE_TC_TIMELINE -I tf_ode_b200/csrc -I include -o tools/policy_tc_timeline tools/policy_tc_timeline.cu
//   tools/policy_tc_timeline [rows]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "tfode_numerics.h"
#include "tfode_policy_tc.cuh"

using namespace tfode;

int main(int argc, char** argv) {
  const int m = argc > 1 ? atoi(argv[1]) : 16384;
  const size_t H = TC_H, nw = 17 * H + H + H * H + H + H * 2 + 2;
  std::vector<float> w(nw), x((size_t)m * 17);
  srand(1);
  for (auto& v : w) v = (rand() / (float)RAND_MAX - 0.5f) * 0.1f;
  for (auto& v : x) v = rand() / (float)RAND_MAX * 2 - 1;
  float *dw, *dx, *dy;
  cudaMalloc(&dw, nw * 4); cudaMalloc(&dx, x.size() * 4); cudaMalloc(&dy, (size_t)m * 2 * 4);
  cudaMemcpy(dw, w.data(), nw * 4, cudaMemcpyHostToDevice); cudaMemcpy(dx, x.data(), x.size() * 4, cudaMemcpyHostToDevice);
  PolicyTC tc;
  if (policy_tc_upload(tc, w.data(), 0)) { printf("upload failed\n"); return 1; }
  const float* b1 = dw + 17 * H; const float* b2 = b1 + H + H * H; const float* W3 = b2 + H; const float* b3 = W3 + 2 * H;
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int i = 0; i < 5; i++) policy_tc_launch(tc, dx, nullptr, nullptr, m, m, b1, b2, W3, b3, nullptr, dy, 0);
  cudaEventRecord(e0);
  const int reps = 200;
  for (int i = 0; i < reps; i++) policy_tc_launch(tc, dx, nullptr, nullptr, m, m, b1, b2, W3, b3, nullptr, dy, 0);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  printf("rows %d: %s, %.2f us per launch\n", m, cudaGetErrorString(e), ms * 1000 / reps);
#ifdef TFODE_TC_TIMELINE
  static long long tl[2048 * 16];
  cudaMemcpyFromSymbol(tl, g_tc_timeline, sizeof(tl));
  const int blocks = (m + TC_ROWS - 1) / TC_ROWS;
  const char* names[16] = {"entry", "setup done", "W1+A1 ready", "A2 first half ready (L2 start)", "last slab arrived", "L1 done", "epi1 done", "H0 done", "epi2 done", "H1 done", "end", "producer done", "slab 15 arrived", "slab 0 arrived", "", ""};
  const int order[13] = {0, 1, 2, 5, 3, 6, 12, 7, 11, 4, 9, 8, 10};
  long long t0min = tl[0];
  for (int b = 0; b < blocks && b < 2048; b++) if (tl[b * 16] < t0min) t0min = tl[b * 16];
  for (int b : {0, 1, blocks / 2, blocks - 1}) {
    if (b < 0 || b >= 2048) continue;
    printf("CTA %d (entry at +%lld clk after the first CTA's entry; clock64 is per SM, compare within a CTA):\n", b, tl[b * 16] - t0min);
    for (int k = 0; k < 13; k++) printf("  %-22s %7lld clk\n", names[order[k]], tl[b * 16 + order[k]] - tl[b * 16]);
  }
#ifdef TFODE_TC_LOOPMARKS
  {
    static unsigned lp[32 * 8];
    cudaMemcpyFromSymbol(lp, g_tc_loop, sizeof(lp));
    printf("CTA 0, MMA issue loop, per W2 slab (clk since the previous slab's loop top): loop top | wait done | - | MMA 1 issued | MMA 2 issued | commit issued\n");
    for (int k = 1; k < TC_NSLAB; k++) {
      printf("  slab %2d  +%4u |", k, lp[k * 8] - lp[(k - 1) * 8]);
      for (int j = 1; j < 6; j++) printf(" %5d", (int)(lp[k * 8 + j] - lp[k * 8]));
      printf("\n");
    }
  }
#endif
  double avg[16] = {0};
  const int nb = blocks < 2048 ? blocks : 2048;
  for (int b = 0; b < nb; b++) for (int k = 0; k < 16; k++) avg[k] += (double)(tl[b * 16 + k] - tl[b * 16]) / nb;
  printf("mean over %d CTAs:\n", nb);
  for (int k = 0; k < 13; k++) printf("  %-22s %9.0f clk\n", names[order[k]], avg[order[k]]);
#endif
  return 0;
}
