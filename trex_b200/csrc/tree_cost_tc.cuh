// Interface between tree_cost.cu (entry point, exact-f32 path) and tree_cost_tc.cu (tcgen05 path).
#pragma once
#include "common.cuh"

namespace trex {
namespace tc {

constexpr int kMaxNodes = 256;   // one CTA holds all of G = S S^T in TMEM (2 x 128 lanes x 256 columns)
constexpr int kMaxSlabs = 64;    // split-K slices of the Gram contraction

struct TcBuffers {
  int npad;            // 128 or 256
  float* d;            // [N] (rowsum + colsum) / 2
  double* cost;        // [1]
  unsigned* ticket;    // [1]
  float* m_hi;         // [npad][npad]  M + M^T, tf32 head
  float* m_lo;         //               and exact remainder
  float* g;            // [npad][npad] G = S S^T (f32, reduced with red.global.add)
};

size_t workspace_bytes(int n_nodes, long long k);
// identity cost, N <= kMaxNodes, and S addressable by TMA (16-byte aligned base and row pitch)
bool supported(const float* seqs, int n_nodes, long long k, int cost_kind);
// value + gradients for cost_kind == 0; d_seqs / d_adj may be null
int run(const float* seqs, const float* adj, int n_nodes, long long k, float* out_cost, float* d_seqs,
        float* d_adj, void* workspace, cudaStream_t stream);

}  // namespace tc
}  // namespace trex
