// Internal interfaces between knn.cu, knn_tc.cu and select.cu.
#pragma once
#include "common.cuh"

// Candidate layout shared by both candidate variants: for query qi and reference chunk c
// (references [c*chunk, min(nr,(c+1)*chunk))) LC list slots at cand_s[(qi*S + c)*LC + t] (approximate
// scores, smaller is better) with their GLOBAL reference index in cand_idx (unused: -1 / +inf).
// SIMT variant: LC = KC, the KC best in ascending order.

// tcgen05 candidate pass (knn_tc.cu), dp in {16,32} (rows of width d <= dp are zero-filled by its packing
// kernel), KC in {8,16}, LC = 6 KC: every reference of the chunk whose score is at or below the chunk's
// threshold (the KC-th best of the column-group minima of the chunk's first half, so at least KC
// references qualify), in no particular order; bound[qi*S + c] = that threshold (no dropped reference
// scores below it), -inf when more than LC qualified (the re-rank then sends the row to the exact
// fallback).  *rmax_bits (zeroed by the caller) receives the largest |r|^2 as fp32 bits.
// cand_idx holds columns of the (row-scrambled) score matrix; *unscramble_out [nr padded to 256] maps a
// column to its reference index (the re-rank applies it).
int knn_cand_tc_launch(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int dp, int KC, int LC,
                       int chunk, int S, int maxdot, int32_t* cand_idx, float* bound, unsigned int* rmax_bits,
                       const int32_t** unscramble_out, cudaStream_t st);
int knn_cand_tc_list_slots(int KC);
// tile passes (MMA + drain of one 128 x 256 accumulator) one CTA of a `tiles`-tile chunk runs
int knn_cand_tc_passes(int tiles);
// relative error bound of the tcgen05 split-bf16 inner product: |err| <= gamma * |q| |r|
double knn_cand_tc_gamma(int d);

// Engine E2 (select.cu): exact float64 rows + per-row radix select.
//   metric: 0 squared L2 (unfused, sequential: bit-identical with the oracle), 1 Euclidean
//   (sqrt of 0), 2 negative inner product, 3 cosine distance.
//   Output rows sorted ascending by (value, index).  idx64/idx32/dist64/dist32 may each be NULL.
//   sort_by_index != 0 returns the k selected indices in ascending INDEX order instead.
int cb_rows_topk(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int k, int metric,
                 int exclude_self, int self_offset, int sort_by_index, int64_t* idx64, int32_t* idx32,
                 double* dist64, float* dist32, int dist32_mode /*0 raw,1 sqrt(v+1e-7),2 -v*/,
                 const int32_t* q_gather /*optional [nq] row ids into q*/, cudaStream_t st);
size_t cb_rows_topk_workspace(int nq, int nr, int k);
