// Interface between tree_cost.cu (entry point, exact-f32 path) and tree_cost_tc.cu (tcgen05 path).
#pragma once
#include "common.cuh"

namespace trex {
namespace tc {

constexpr int kMaxNodes = 256;   // one CTA holds all of G = S S^T in TMEM (2 x 128 lanes x 256 columns)
constexpr int kMaxSlabs = 64;    // split-K slices of the Gram contraction
constexpr int kOneMaxGram = 32;  // one-kernel mode: split-K slices, each with its own partial-G slot
constexpr int kOneMaxCtas = 256; // one-kernel mode: flag / partial-cost slots (the grid never exceeds the SM count)

struct TcBuffers {
  int npad;            // 128 or 256
  float* d;            // [N] (rowsum + colsum) / 2
  double* cost;        // [1]
  unsigned* ticket;    // [1]
  float* m_hi;         // [npad][npad]  M + M^T, tf32 head
  float* m_lo;         //               and exact remainder
  float* g;            // [npad][npad] G = S S^T (f32, reduced with red.global.add)
  // one-kernel mode (N <= 128, grid <= SM count): no prep / final launches, no atomics, no zeroing
  unsigned* epoch;             // [1] launch counter kept ON THE DEVICE (so a replayed CUDA graph advances it too)
  unsigned long long* flags;   // [3][kOneMaxCtas] tags: Msym rows ready / partial G stored / partial cost stored
  float* slots;                // [kOneMaxGram][128][128] partial Gram tiles of the split-K slices
  float* diag_part;            // [kOneMaxGram][128] their diagonals
  double* cost_part;           // [kOneMaxCtas]
  unsigned* err;               // [1] set when a flag wait gave up (never on a healthy device)
};

size_t workspace_bytes(int n_nodes, long long k);
// identity cost, N <= kMaxNodes, and S addressable by TMA (16-byte aligned base and row pitch)
bool supported(const float* seqs, int n_nodes, long long k, int cost_kind);
// value + gradients for cost_kind == 0; d_seqs / d_adj may be null
int run(const float* seqs, const float* adj, int n_nodes, long long k, float* out_cost, float* d_seqs,
        float* d_adj, void* workspace, cudaStream_t stream);

}  // namespace tc
}  // namespace trex
