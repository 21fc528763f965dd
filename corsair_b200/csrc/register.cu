// Composite entry point: every registration of one query-CAD pair in ONE call.
// Replaces the body of sym_pose (utils/symmetry.py:262-358 of the reference) after the two
// symmetry segmentations: the vanilla registration (:267-283: full-cloud feature kNN ->
// correspondence RANSAC -> one-direction Chamfer of the float32-rounded pose) and the P symmetry
// pairings of its sweeps (:303-356: per-segment kNN between paired segments (split_corr), the
// concatenated correspondences -> RANSAC -> Chamfer).  The host passes the segment permutations;
// everything else — ~25 launches per registration — is enqueued from C on one stream, so a host
// thread of the drop-in spends its time outside the Python interpreter (symmetry.sym_pose_many).
#include "common.cuh"

namespace {

// correspondences of one kNN problem: query row t of the (segment-sorted) query block pairs with its
// k neighbours; src = a_perm[q0 + t] (or q0 + t), tgt = b_perm[r0 + idx] (or r0 + idx)
__global__ void reg_corr_kernel(const int64_t* __restrict__ idx, int nq, int k, int q0, int r0,
                                const int64_t* __restrict__ a_perm, const int64_t* __restrict__ b_perm,
                                int64_t* __restrict__ src_idx, int64_t* __restrict__ tgt_idx) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nq * k) return;
    const int t = i / k;
    const size_t o = (size_t)q0 * k + i;
    src_idx[o] = a_perm ? a_perm[q0 + t] : (int64_t)(q0 + t);
    const int64_t j = (int64_t)r0 + idx[(size_t)q0 * k + i];
    tgt_idx[o] = b_perm ? b_perm[j] : j;
}

__global__ void reg_gather_rows_kernel(const float* __restrict__ x, const int64_t* __restrict__ perm, int n, int d,
                                       float* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)n * d) return;
    const int row = (int)(i / d), c = (int)(i % d);
    out[i] = x[(size_t)perm[row] * d + c];
}

// the reference evaluates Chamfer with (and returns) the pose rounded to float32 (:274,:280,:318)
__global__ void reg_round_pose_kernel(const double* __restrict__ T, double* __restrict__ out) {
    if (threadIdx.x < 16) out[threadIdx.x] = (double)(float)T[threadIdx.x];
}

}  // namespace

extern "C" int cb_register_pairings(cb_ctx* ctx, const float* base_f, const float* base_xyz, int n0,
                                    const float* pos_f, const float* pos_xyz, int n1, int d, int k_nn,
                                    const int64_t* a_perm, const int32_t* h_q_off, const int64_t* b_perm,
                                    const int32_t* h_r_off, int n_seg, const int32_t* h_pair_map, int n_pairings,
                                    double max_dist, uint64_t seed, int h_total, int adaptive, double confidence,
                                    double* T_out, double* chamfer_out, double* stats_out, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (!base_f || !base_xyz || !pos_f || !pos_xyz || !T_out || !chamfer_out || n0 <= 0 || n1 <= 0 || d <= 0 ||
        k_nn <= 0 || n_pairings < 0 || n_seg < 0)
        CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: bad arguments");
    if (n_pairings > 0 && (!a_perm || !b_perm || !h_q_off || !h_r_off || !h_pair_map || n_seg <= 0))
        CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: pairings need the segment permutations");
    if (k_nn > n1) CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: k_nn=%d exceeds the %d target points", k_nn, n1);
    if (n_pairings > 0 && (h_q_off[n_seg] != n0 || h_r_off[n_seg] != n1))
        CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: segment offsets do not cover the clouds");
    const int m = n0 * k_nn;
    // persistent buffers + the largest transient step (each step releases its scratch: the stream
    // orders the kernels, so the next step may overwrite it)
    size_t persistent = cb_align((size_t)m * 8) * 3 + cb_align((size_t)n0 * d * 4) + cb_align((size_t)n1 * d * 4) + 4096;
    size_t transient = cb_knn_workspace(n0, n1, d, k_nn);
    transient = transient > cb_ransac_workspace(ctx, m, h_total) ? transient : cb_ransac_workspace(ctx, m, h_total);
    const size_t ch = (size_t)(n0 + n1) * 40 + (2 << 16);
    transient = transient > ch ? transient : ch;
    if ((rc = cb_reserve(ctx, persistent + transient + (1 << 20)))) return rc;
    int64_t* idx = cb_alloc_t<int64_t>(ctx, m);
    int64_t* src_idx = cb_alloc_t<int64_t>(ctx, m);
    int64_t* tgt_idx = cb_alloc_t<int64_t>(ctx, m);
    float* f0s = cb_alloc_t<float>(ctx, (size_t)n0 * d);
    float* f1s = cb_alloc_t<float>(ctx, (size_t)n1 * d);
    double* T64 = cb_alloc_t<double>(ctx, 16);
    double* stats = cb_alloc_t<double>(ctx, 4);
    if (!idx || !src_idx || !tgt_idx || !f0s || !f1s || !T64 || !stats) return CB_ENOMEM;
    const size_t mark = ctx->arena_off;

    auto finish = [&](int slot) -> int {
        // RANSAC on the correspondence list, pose rounded to float32, one-direction Chamfer with it
        int r2 = cb_ransac_impl(ctx, base_xyz, pos_xyz, src_idx, tgt_idx, m, nullptr, seed, h_total, 10, max_dist,
                                adaptive, confidence, T64, stats_out ? stats_out + 4 * slot : stats, nullptr, nullptr,
                                st);
        ctx->arena_off = mark;
        if (r2) return r2;
        reg_round_pose_kernel<<<1, 32, 0, st>>>(T64, T_out + 16 * slot);
        CB_LAUNCH_CHECK(ctx);
        r2 = cb_chamfer_impl(ctx, base_xyz, n0, pos_xyz, n1, T_out + 16 * slot, chamfer_out + slot, st);
        ctx->arena_off = mark;
        return r2;
    };

    // vanilla registration
    if ((rc = cb_knn_impl(ctx, base_f, n0, pos_f, n1, d, k_nn, CB_KNN_SQL2, idx, nullptr, nullptr, st))) return rc;
    ctx->arena_off = mark;
    reg_corr_kernel<<<cb_div_up(m, 256), 256, 0, st>>>(idx, n0, k_nn, 0, 0, nullptr, nullptr, src_idx, tgt_idx);
    CB_LAUNCH_CHECK(ctx);
    if ((rc = finish(0))) return rc;
    if (n_pairings == 0) return CB_OK;

    // segment-sorted feature rows, gathered once; pairing p matches base segment s with pos segment
    // pair_map[p][s]
    reg_gather_rows_kernel<<<(unsigned)(((size_t)n0 * d + 255) / 256), 256, 0, st>>>(base_f, a_perm, n0, d, f0s);
    CB_LAUNCH_CHECK(ctx);
    reg_gather_rows_kernel<<<(unsigned)(((size_t)n1 * d + 255) / 256), 256, 0, st>>>(pos_f, b_perm, n1, d, f1s);
    CB_LAUNCH_CHECK(ctx);
    for (int p = 0; p < n_pairings; ++p) {
        for (int s = 0; s < n_seg; ++s) {
            const int q0 = h_q_off[s], nq = h_q_off[s + 1] - q0;
            const int ps = h_pair_map[p * n_seg + s];
            if (ps < 0 || ps >= n_seg) CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: pair_map entry out of range");
            const int r0 = h_r_off[ps], nr = h_r_off[ps + 1] - r0;
            if (nq < 0 || nr < 0) CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: offsets must not decrease");
            if (nq == 0) continue;
            if (k_nn > nr)
                CB_FAIL(ctx, CB_EINVAL, "cb_register_pairings: pairing %d, segment %d has %d target points, k_nn=%d", p, s,
                        nr, k_nn);
            rc = cb_knn_impl(ctx, f0s + (size_t)q0 * d, nq, f1s + (size_t)r0 * d, nr, d, k_nn, CB_KNN_SQL2,
                             idx + (size_t)q0 * k_nn, nullptr, nullptr, st);
            ctx->arena_off = mark;
            if (rc) return rc;
            reg_corr_kernel<<<cb_div_up(nq * k_nn, 256), 256, 0, st>>>(idx, nq, k_nn, q0, r0, a_perm, b_perm, src_idx,
                                                                       tgt_idx);
            CB_LAUNCH_CHECK(ctx);
        }
        if ((rc = finish(1 + p))) return rc;
    }
    return CB_OK;
}
