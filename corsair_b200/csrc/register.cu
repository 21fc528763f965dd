// Composite entry points: every registration of one query-CAD pair in ONE call (cb_register_pairings),
// and the whole of sym_pose including its two symmetry segmentations (cb_sym_pose).
// Replaces the body of sym_pose (utils/symmetry.py:262-358 of the reference) after the two
// symmetry segmentations: the vanilla registration (:267-283: full-cloud feature kNN ->
// correspondence RANSAC -> one-direction Chamfer of the float32-rounded pose) and the P symmetry
// pairings of its sweeps (:303-356: per-segment kNN between paired segments (split_corr), the
// concatenated correspondences -> RANSAC -> Chamfer).  The host passes the segment permutations;
// everything else — ~25 launches per registration — is enqueued from C on one stream, so a host
// thread of the drop-in spends its time outside the Python interpreter (symmetry.sym_pose_many).
#include "common.cuh"

#include <stdlib.h>

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

struct RegArgs {
    const float *base_f, *base_xyz, *pos_f, *pos_xyz;
    int n0, n1, d, k_nn;
    double max_dist;
    uint64_t seed;
    int h_total, adaptive;
    double confidence;
};

// segment-sorted feature rows + the permutations / offsets behind them
struct SegView {
    const float *f0s, *f1s;
    const int64_t *a_perm, *b_perm;
    const int32_t *h_q_off, *h_r_off;
    int n_seg;
};

size_t reg_one_workspace(const cb_ctx* ctx, const RegArgs& a) {
    const int m = a.n0 * a.k_nn;
    size_t transient = cb_knn_workspace(a.n0, a.n1, a.d, a.k_nn);
    const size_t rw = cb_ransac_workspace(ctx, m, a.h_total);
    transient = transient > rw ? transient : rw;
    const size_t ch = (size_t)(a.n0 + a.n1) * 40 + (2 << 16);
    transient = transient > ch ? transient : ch;
    return cb_align((size_t)m * 8) * 3 + 4096 + transient;
}

// ONE registration on stream st out of ctx's arena (from its current offset, restored on return):
// correspondences (pair_map == nullptr: kNN over the whole clouds, utils/symmetry.py:267-271; else the
// per-segment kNN of split_corr for base segment s against pos segment pair_map[s], :305-316) ->
// RANSAC -> pose rounded to float32 -> one-direction Chamfer.
int register_one(cb_ctx* ctx, cudaStream_t st, const RegArgs& a, const SegView* sv, const int32_t* pair_map,
                 double* T_out, double* chamfer_out, double* stats_out) {
    const size_t entry = ctx->arena_off;
    const int m = a.n0 * a.k_nn;
    int64_t* idx = cb_alloc_t<int64_t>(ctx, m);
    int64_t* src_idx = cb_alloc_t<int64_t>(ctx, m);
    int64_t* tgt_idx = cb_alloc_t<int64_t>(ctx, m);
    double* T64 = cb_alloc_t<double>(ctx, 16);
    double* stats = cb_alloc_t<double>(ctx, 4);
    if (!idx || !src_idx || !tgt_idx || !T64 || !stats) return CB_ENOMEM;
    const size_t mark = ctx->arena_off;
    int rc;
    if (!pair_map) {
        rc = cb_knn_impl(ctx, a.base_f, a.n0, a.pos_f, a.n1, a.d, a.k_nn, CB_KNN_SQL2, idx, nullptr, nullptr, st);
        ctx->arena_off = mark;
        if (rc) return rc;
        reg_corr_kernel<<<cb_div_up(m, 256), 256, 0, st>>>(idx, a.n0, a.k_nn, 0, 0, nullptr, nullptr, src_idx, tgt_idx);
        CB_LAUNCH_CHECK(ctx);
    } else {
        for (int s = 0; s < sv->n_seg; ++s) {
            const int q0 = sv->h_q_off[s], nq = sv->h_q_off[s + 1] - q0;
            const int ps = pair_map[s];
            if (ps < 0 || ps >= sv->n_seg) CB_FAIL(ctx, CB_EINVAL, "registration: pair_map entry out of range");
            const int r0 = sv->h_r_off[ps], nr = sv->h_r_off[ps + 1] - r0;
            if (nq < 0 || nr < 0) CB_FAIL(ctx, CB_EINVAL, "registration: segment offsets must not decrease");
            if (nq == 0) continue;
            if (a.k_nn > nr)
                CB_FAIL(ctx, CB_EINVAL, "registration: segment %d is matched with %d target points, k_nn=%d", s, nr,
                        a.k_nn);
            rc = cb_knn_impl(ctx, sv->f0s + (size_t)q0 * a.d, nq, sv->f1s + (size_t)r0 * a.d, nr, a.d, a.k_nn,
                             CB_KNN_SQL2, idx + (size_t)q0 * a.k_nn, nullptr, nullptr, st);
            ctx->arena_off = mark;
            if (rc) return rc;
            reg_corr_kernel<<<cb_div_up(nq * a.k_nn, 256), 256, 0, st>>>(idx, nq, a.k_nn, q0, r0, sv->a_perm, sv->b_perm,
                                                                         src_idx, tgt_idx);
            CB_LAUNCH_CHECK(ctx);
        }
    }
    rc = cb_ransac_impl(ctx, a.base_xyz, a.pos_xyz, src_idx, tgt_idx, m, nullptr, a.seed, a.h_total, 10, a.max_dist,
                        a.adaptive, a.confidence, T64, stats_out ? stats_out : stats, nullptr, nullptr, st);
    ctx->arena_off = mark;
    if (rc) return rc;
    reg_round_pose_kernel<<<1, 32, 0, st>>>(T64, T_out);
    CB_LAUNCH_CHECK(ctx);
    rc = cb_chamfer_impl(ctx, a.base_xyz, a.n0, a.pos_xyz, a.n1, T_out, chamfer_out, st);
    ctx->arena_off = entry;
    return rc;
}

int gather_segment_rows(cb_ctx* ctx, cudaStream_t st, const RegArgs& a, const int64_t* a_perm, const int64_t* b_perm,
                        float* f0s, float* f1s) {
    reg_gather_rows_kernel<<<(unsigned)(((size_t)a.n0 * a.d + 255) / 256), 256, 0, st>>>(a.base_f, a_perm, a.n0, a.d, f0s);
    CB_LAUNCH_CHECK(ctx);
    reg_gather_rows_kernel<<<(unsigned)(((size_t)a.n1 * a.d + 255) / 256), 256, 0, st>>>(a.pos_f, b_perm, a.n1, a.d, f1s);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
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
    const RegArgs a{base_f, base_xyz, pos_f, pos_xyz, n0, n1, d, k_nn, max_dist, seed, h_total, adaptive, confidence};
    // each step releases its scratch: the stream orders the kernels, so the next step may overwrite it
    const size_t feat_rows = cb_align((size_t)n0 * d * 4) + cb_align((size_t)n1 * d * 4);
    if ((rc = cb_reserve(ctx, feat_rows + reg_one_workspace(ctx, a) + (1 << 20)))) return rc;
    if ((rc = register_one(ctx, st, a, nullptr, nullptr, T_out, chamfer_out, stats_out))) return rc;
    if (n_pairings == 0) return CB_OK;
    float* f0s = cb_alloc_t<float>(ctx, (size_t)n0 * d);
    float* f1s = cb_alloc_t<float>(ctx, (size_t)n1 * d);
    if (!f0s || !f1s) return CB_ENOMEM;
    if ((rc = gather_segment_rows(ctx, st, a, a_perm, b_perm, f0s, f1s))) return rc;
    const SegView sv{f0s, f1s, a_perm, b_perm, h_q_off, h_r_off, n_seg};
    for (int p = 0; p < n_pairings; ++p)
        if ((rc = register_one(ctx, st, a, &sv, h_pair_map + (size_t)p * n_seg, T_out + 16 * (1 + p),
                               chamfer_out + 1 + p, stats_out ? stats_out + 4 * (1 + p) : nullptr)))
            return rc;
    return CB_OK;
}

// ---- sym_pose in one call ---------------------------------------------------------------------
namespace {

constexpr size_t kPinnedBytes = 64 << 10;

int sym_default_streams() {
    static const int n = [] {
        const char* e = getenv("CB_SYM_STREAMS");
        int v = e ? atoi(e) : 3;
        return v < 0 ? 0 : (v > 8 ? 8 : v);
    }();
    return n;
}

// sub-contexts: the 1 + P registrations of a pair are independent, each is a chain of ~25 mostly
// small kernels, so they run on a few streams of their own (own arena each) and overlap on the device
// Wait for the stream.  Measured: the default spinning cudaStreamSynchronize beats sleeping on a
// blocking-sync event both at 1 GPU / 16 cores (782 vs 771 pairs/s) and at 8 GPUs / 32 cores with six
// host threads per rank (5 243 vs 4 804); CB_SYNC_BLOCK=1 selects the sleeping wait for hosts that
// need their cores back.
cudaError_t sym_wait(cb_ctx* ctx, cudaStream_t st) {
    static const bool block = getenv("CB_SYNC_BLOCK") && atoi(getenv("CB_SYNC_BLOCK"));
    if (!block) return cudaStreamSynchronize(st);
    cudaError_t e = cudaEventRecord(ctx->sync_event, st);
    return e != cudaSuccess ? e : cudaEventSynchronize(ctx->sync_event);
}

int sym_prepare(cb_ctx* ctx, int n_sub) {
    if (!ctx->pinned) CB_CUDA(ctx, cudaHostAlloc((void**)&ctx->pinned, kPinnedBytes, cudaHostAllocDefault));
    if (!ctx->fork_event) CB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->fork_event, cudaEventDisableTiming));
    if (!ctx->sync_event)
        CB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->sync_event, cudaEventDisableTiming | cudaEventBlockingSync));
    while ((int)ctx->sub.size() < n_sub) {
        cb_ctx* c = nullptr;
        if (cb_create(&c, ctx->device, size_t(64) << 20) != CB_OK)
            CB_FAIL(ctx, CB_ENOMEM, "cb_sym_pose: sub-context could not be created");
        cudaStream_t s = nullptr;
        cudaEvent_t e = nullptr;
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
            cb_destroy(c);
            CB_FAIL(ctx, CB_ECUDA, "cb_sym_pose: stream / event creation failed");
        }
        ctx->sub.push_back(c);
        ctx->sub_stream.push_back(s);
        ctx->sub_event.push_back(e);
    }
    return CB_OK;
}

}  // namespace

extern "C" int cb_sym_pose(cb_ctx* ctx, const float* base_f, const float* base_xyz, int n0, const float* pos_f,
                           const float* pos_xyz, int n1, int d, int pos_sym, int k_nn, const int32_t* h_samples_base,
                           const int32_t* h_samples_pos, int n_samples, int64_t np_seed, int n_nn,
                           const int32_t* h_km_first,
                           const double* h_km_unif, int n_init, int n_trials, double min_center, double max_dist,
                           uint64_t seed, int h_total, int adaptive, double confidence, double* h_out, void* stream) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    if (!base_f || !base_xyz || !pos_f || !pos_xyz || !h_out || n0 <= 0 || n1 <= 0 || d <= 0 || k_nn <= 0 ||
        n_samples < 0)
        CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose: bad arguments");
    if (k_nn > n1) CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose: k_nn=%d exceeds the %d target points", k_nn, n1);
    const int K = pos_sym >= 2 ? 4 : 2;
    int32_t drawn[2 * 1024];
    if (n_samples > 0 && !h_samples_base && !h_samples_pos && np_seed >= 0 && n_samples <= 1024) {
        // the draws of np.random.seed(np_seed); choice(n0, ..); choice(n1, ..) (utils/symmetry.py:193);
        // a cloud with fewer points than samples makes the reference raise and fall back (:299-301)
        const int32_t ns[2] = {n0, n1};
        if (cb_np_seeded_choices((uint32_t)np_seed, ns, 2, n_samples, drawn) == CB_OK) {
            h_samples_base = drawn;
            h_samples_pos = drawn + n_samples;
        } else {
            n_samples = 0;
        }
    }
    const int P = n_samples > 0 ? (pos_sym >= 2 ? 2 * K : K) : 0;
    if (n_samples > 0) {
        if (!h_samples_base || !h_samples_pos || !h_km_first || !h_km_unif || n_init < 1 || n_init > 16 ||
            n_trials < 1 || n_trials > 4 || n_nn < 1)
            CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose: the segmentation needs its sample rows and k-means++ draws");
        if (n_samples > 1024) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_sym_pose: at most 1024 seed samples (got %d)", n_samples);
    }
    const RegArgs a{base_f, base_xyz, pos_f, pos_xyz, n0, n1, d, k_nn, max_dist, seed, h_total, adaptive, confidence};
    const int n_sub = ctx->timing ? 0 : sym_default_streams();  // per-class event timing needs one stream
    if ((rc = sym_prepare(ctx, n_sub))) return rc;

    // host -> pinned -> device: sample rows of both clouds and the k-means++ draws, in one copy
    const size_t n_unif = (size_t)n_init * (K - 1) * n_trials;
    const size_t in_bytes = n_samples > 0 ? cb_align((size_t)2 * n_samples * 4 + (size_t)n_init * 4) + n_unif * 8 : 0;
    const size_t off_bytes = 256, res_bytes = (size_t)(1 + 8) * 17 * 8;
    if (in_bytes + off_bytes + res_bytes > kPinnedBytes) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_sym_pose: staging too small");
    char* pin_in = ctx->pinned;
    int32_t* pin_off = reinterpret_cast<int32_t*>(ctx->pinned + cb_align(in_bytes));
    double* pin_res = reinterpret_cast<double*>(ctx->pinned + cb_align(in_bytes) + off_bytes);

    const int nmax = n0 > n1 ? n0 : n1;
    size_t need = cb_align(in_bytes) + 2 * cb_align((size_t)nmax * 8) + 512 + cb_align((size_t)(1 + P) * 17 * 8) +
                  cb_align((size_t)n0 * d * 4) + cb_align((size_t)n1 * d * 4) + 8192;
    size_t scratch = n_samples > 0 ? cb_segment_workspace(nmax, n_samples, n_nn, K) : 0;
    if (n_sub == 0 && reg_one_workspace(ctx, a) > scratch) scratch = reg_one_workspace(ctx, a);
    if ((rc = cb_reserve(ctx, need + scratch + (1 << 20)))) return rc;
    for (int j = 0; j < n_sub; ++j) {
        cb_ctx* c = ctx->sub[j];
        c->knn_variant = ctx->knn_variant;
        c->ransac_variant = ctx->ransac_variant;
        size_t sub_need = reg_one_workspace(c, a);
        if (j == 1 && scratch > sub_need) sub_need = scratch;  // sub-context 1 also runs the pos segmentation
        if ((rc = cb_begin(c)) || (rc = cb_reserve(c, sub_need + (1 << 20)))) {
            snprintf(ctx->err, sizeof(ctx->err), "%s", c->err);
            return rc;
        }
    }
    char* d_in = in_bytes ? reinterpret_cast<char*>(cb_alloc(ctx, in_bytes)) : nullptr;
    int64_t* a_perm = cb_alloc_t<int64_t>(ctx, n0);
    int64_t* b_perm = cb_alloc_t<int64_t>(ctx, n1);
    int32_t* d_off = cb_alloc_t<int32_t>(ctx, 64);
    double* T_all = cb_alloc_t<double>(ctx, (size_t)(1 + P) * 17);
    double* chamf = T_all + (size_t)(1 + P) * 16;
    float* f0s = cb_alloc_t<float>(ctx, (size_t)n0 * d);
    float* f1s = cb_alloc_t<float>(ctx, (size_t)n1 * d);
    if ((in_bytes && !d_in) || !a_perm || !b_perm || !d_off || !T_all || !f0s || !f1s) return CB_ENOMEM;

    bool forked = false;
    unsigned used = 0;
    auto bail = [&](int code) {  // nothing of this call may still run when the caller sees the error
        if (forked)
            for (int j = 0; j < n_sub; ++j) cudaStreamSynchronize(ctx->sub_stream[j]);
        cudaStreamSynchronize(st);
        return code;
    };
    // registration `slot` (0: vanilla) on sub-stream slot % n_sub, ordered after the work on st so far
    auto run = [&](int slot, const SegView* sv, const int32_t* pair_map) -> int {
        cb_ctx* c = ctx;
        cudaStream_t s = st;
        if (n_sub > 0) {
            const int j = slot % n_sub;
            c = ctx->sub[j];
            s = ctx->sub_stream[j];
            if (cudaEventRecord(ctx->fork_event, st) != cudaSuccess ||
                cudaStreamWaitEvent(s, ctx->fork_event, 0) != cudaSuccess)
                CB_FAIL(ctx, CB_ECUDA, "cb_sym_pose: fork failed: %s", cudaGetErrorString(cudaGetLastError()));
            forked = true;
            used |= 1u << j;
        }
        const int r = register_one(c, s, a, sv, pair_map, T_all + 16 * slot, chamf + slot, nullptr);
        if (r && c != ctx) snprintf(ctx->err, sizeof(ctx->err), "%s", c->err);
        return r;
    };

    if (in_bytes) {
        int32_t* p32 = reinterpret_cast<int32_t*>(pin_in);
        memcpy(p32, h_samples_base, (size_t)n_samples * 4);
        memcpy(p32 + n_samples, h_samples_pos, (size_t)n_samples * 4);
        memcpy(p32 + 2 * n_samples, h_km_first, (size_t)n_init * 4);
        memcpy(pin_in + cb_align((size_t)2 * n_samples * 4 + (size_t)n_init * 4), h_km_unif, n_unif * 8);
        CB_CUDA(ctx, cudaMemcpyAsync(d_in, pin_in, in_bytes, cudaMemcpyHostToDevice, st));
    }
    // vanilla registration (utils/symmetry.py:267-283); on its own stream it overlaps the segmentation
    if ((rc = run(0, nullptr, nullptr))) return bail(rc);

    int status = 3;  // 0 both segmentations accepted a model, 1 base failed, 2 pos failed, 3 not attempted
    if (n_samples > 0) {
        const int32_t* d_s = reinterpret_cast<const int32_t*>(d_in);
        const int32_t* d_first = d_s + 2 * n_samples;
        const double* d_unif = reinterpret_cast<const double*>(d_in + cb_align((size_t)2 * n_samples * 4 + (size_t)n_init * 4));
        const size_t mark = ctx->arena_off;
        // the pos cloud's segmentation is independent of the base cloud's: second internal stream
        if (n_sub >= 2) {
            cb_ctx* c = ctx->sub[1];
            cudaStream_t s = ctx->sub_stream[1];
            if (cudaEventRecord(ctx->fork_event, st) != cudaSuccess ||
                cudaStreamWaitEvent(s, ctx->fork_event, 0) != cudaSuccess)
                return bail(CB_ECUDA);
            forked = true;
            used |= 2u;
            const size_t m1 = c->arena_off;
            rc = cb_segment_impl(c, pos_f, pos_xyz, n1, d, K, d_s + n_samples, n_samples, n_nn < n1 ? n_nn : n1, d_first,
                                 d_unif, n_init, n_trials, min_center, b_perm, d_off + 8, s);
            c->arena_off = m1;
            if (rc) {
                snprintf(ctx->err, sizeof(ctx->err), "%s", c->err);
                return bail(rc);
            }
            rc = cb_segment_impl(ctx, base_f, base_xyz, n0, d, K, d_s, n_samples, n_nn < n0 ? n_nn : n0, d_first, d_unif,
                                 n_init, n_trials, min_center, a_perm, d_off, st);
            ctx->arena_off = mark;
            if (rc) return bail(rc);
            if (cudaEventRecord(ctx->sub_event[1], s) != cudaSuccess ||
                cudaStreamWaitEvent(st, ctx->sub_event[1], 0) != cudaSuccess)
                return bail(CB_ECUDA);
        } else {
            rc = cb_segment_impl(ctx, base_f, base_xyz, n0, d, K, d_s, n_samples, n_nn < n0 ? n_nn : n0, d_first, d_unif,
                                 n_init, n_trials, min_center, a_perm, d_off, st);
            ctx->arena_off = mark;
            if (rc) return bail(rc);
            rc = cb_segment_impl(ctx, pos_f, pos_xyz, n1, d, K, d_s + n_samples, n_samples, n_nn < n1 ? n_nn : n1,
                                 d_first, d_unif, n_init, n_trials, min_center, b_perm, d_off + 8, st);
            ctx->arena_off = mark;
            if (rc) return bail(rc);
        }
        if (cudaMemcpyAsync(pin_off, d_off, 16 * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
            sym_wait(ctx, st) != cudaSuccess) {
            snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: segment offsets: %s", cudaGetErrorString(cudaGetLastError()));
            return bail(CB_ECUDA);
        }
        status = pin_off[K + 1] < 0 ? 1 : (pin_off[8 + K + 1] < 0 ? 2 : 0);
    }
    int n_reg = 1;
    if (status == 0) {
        if ((rc = gather_segment_rows(ctx, st, a, a_perm, b_perm, f0s, f1s))) return bail(rc);
        const SegView sv{f0s, f1s, a_perm, b_perm, pin_off, pin_off + 8, K};
        // pairings of utils/symmetry.py:303-356: the K cyclic rotations of the pos segments, then
        // (pos_sym >= 2) those of [B0, B3, B2, B1]
        static const int order2[4] = {0, 3, 2, 1};
        for (int p = 0; p < P; ++p) {
            int32_t map[4];
            for (int s = 0; s < K; ++s) map[s] = p < K ? (s + p) % K : order2[(s + (p - K)) % K];
            if ((rc = run(1 + p, &sv, map))) return bail(rc);
        }
        n_reg = 1 + P;
    }
    // join, read the poses back, keep the strictly lowest Chamfer (:321,:353)
    for (int j = 0; j < n_sub; ++j)
        if (used >> j & 1) {
            if (cudaEventRecord(ctx->sub_event[j], ctx->sub_stream[j]) != cudaSuccess ||
                cudaStreamWaitEvent(st, ctx->sub_event[j], 0) != cudaSuccess) {
                snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: join failed: %s", cudaGetErrorString(cudaGetLastError()));
                return bail(CB_ECUDA);
            }
            ctx->launches += ctx->sub[j]->launches;
            ctx->sub[j]->launches = 0;
        }
    if (cudaMemcpyAsync(pin_res, T_all, (size_t)(1 + P) * 17 * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
        sym_wait(ctx, st) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: result read failed: %s", cudaGetErrorString(cudaGetLastError()));
        return bail(CB_ECUDA);
    }
    const double* ch = pin_res + (size_t)(1 + P) * 16;
    int best = 0;
    for (int i = 1; i < n_reg; ++i)
        if (ch[best] > ch[i]) best = i;
    memcpy(h_out, pin_res + 16 * best, 16 * 8);
    h_out[16] = ch[best];
    memcpy(h_out + 17, pin_res, 16 * 8);
    h_out[33] = ch[0];
    h_out[34] = (double)status;
    h_out[35] = (double)best;
    h_out[36] = (double)(n_reg - 1);
    return CB_OK;
}
