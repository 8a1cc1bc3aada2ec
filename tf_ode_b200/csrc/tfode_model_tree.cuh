// Batched model-tree controller: Rover_1_mt::_InternalControl (src/rover_1_mt.cc:72-83) for many environments.
//
// The reference holds two Linear_model_tree<double> (or Polynomial_model_tree<double>) objects per rover
// (src/rover_1_mt.cc:13-27): tree 1 predicts the steering rate, tree 2 the boggie torque, both from the 14-value
// Rover_1_mt::GetState(flip = false, full = false) vector (:30-52).  Here the two trees live flattened in HBM
// (tfode_formats.hpp: nodes in the loader's own depth-first order + one parameter pool), shared by all environments,
// and one thread walks both trees for one environment:
//   Polynomial_model_tree::predict   scripts/ModelTree/cpp/model_tree.cc:91-123   feature expansion (degree <= 2 here)
//   _traverse_and_predict            scripts/ModelTree/cpp/model_tree.cc:47-65    straight split input[p0] >= p1 -> child 0,
//                                                                                 oblique split <input,p> + p_last >= 0 -> child 0,
//                                                                                 terminal: std::inner_product(input, p, p_last)
// All arithmetic is fp64 with separate multiply and add in std::inner_product's order (acc = acc + x*p), so the result is
// bit-identical to the loader's when its inputs are.
#pragma once
#include "tfode_device.cuh"

namespace tfode {

constexpr int TREE_NIN = TFODE_TREE_INPUTS;          // 14
constexpr int TREE_MAXF = TFODE_TREE_MAX_FEATURES;   // 14 + 14*15/2 = 119 (degree 2)

__device__ __forceinline__ double tree_inner(const double* x, const double* p, int n) {   // std::inner_product(x, x+n, p, p[n])
  double acc = p[n];
  for (int i = 0; i < n; i++) acc = __dadd_rn(acc, __dmul_rn(x[i], p[i]));
  return acc;
}

__device__ double tree_predict(const TreeDev& T, const double* feat, int nfeat, int oblique, int& node_id) {
  int k = 0;
  for (int guard = 0; guard < T.n_nodes; guard++) {
    const TfodeTreeNode nd = T.nodes[k];
    const double* p = T.params + nd.param_off;
    if (nd.terminal) { node_id = nd.id; return tree_inner(feat, p, nfeat); }
    bool side;
    if (oblique) side = tree_inner(feat, p, nfeat) >= 0.0;
    else side = feat[(int)p[0]] >= p[1];
    k = nd.child[side ? 0 : 1];
  }
  node_id = -1;
  return 0.0;
}

// list != NULL: environments parked by step_kernel at an IC tick (state in policy_in[env*17 + 0..13] as float), results go
//               into the record (Rover_1_mt writes _steering_rate/_boggie_torque directly, src/rover_1_mt.cc:58-59).
// list == NULL: stand-alone op on m rows of fp64 states -> act[m,2], node[m,2].
__global__ void __launch_bounds__(128) model_tree_kernel(TreeParams TP, const float* __restrict__ policy_in, const int* __restrict__ list,
                                                         const int* __restrict__ n_list, EnvRec* rec, const double* __restrict__ in_direct, int m_direct,
                                                         double* __restrict__ act_out, int32_t* __restrict__ node_out) {
  const int m = list ? *n_list : m_direct;
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < m; r += gridDim.x * blockDim.x) {
    double feat[TREE_MAXF];
    const int env = list ? list[r] : r;
    if (list) { for (int i = 0; i < TREE_NIN; i++) feat[i] = (double)policy_in[(size_t)env * 17 + i]; }
    else { for (int i = 0; i < TREE_NIN; i++) feat[i] = in_direct[(size_t)r * TREE_NIN + i]; }
    int nf = TREE_NIN;
    if (TP.degree == 2) {   // chunk of degree-2 products, model_tree.cc:101-121 with d = 1
      const int io = TP.interaction_only ? 1 : 0;
      for (int i = 0; i < TREE_NIN - io; i++)
        for (int j = i + io; j < TREE_NIN; j++) feat[nf++] = __dmul_rn(feat[i], feat[j]);
    }
    int n1 = 0, n2 = 0;
    const double a1 = tree_predict(TP.tree[0], feat, nf, TP.oblique, n1);
    const double a2 = tree_predict(TP.tree[1], feat, nf, TP.oblique, n2);
    if (list) {
      Ctrl& c = rec[env].c;
      c.steering_rate = a1; c.boggie_torque = a2;
      c.tree_nodes = (n1 & 0xffff) | (n2 << 16);
      c.parked = 2;
    } else {
      act_out[(size_t)r * 2] = a1; act_out[(size_t)r * 2 + 1] = a2;
      if (node_out) { node_out[(size_t)r * 2] = n1; node_out[(size_t)r * 2 + 1] = n2; }
    }
  }
}

__global__ void tree_nodes_kernel(const EnvRec* rec, int n, int32_t* out) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e < n) { const int v = rec[e].c.tree_nodes; out[e * 2] = v & 0xffff; out[e * 2 + 1] = (v >> 16) & 0xffff; }
}

}  // namespace tfode
