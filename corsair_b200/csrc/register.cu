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

// cb_sym_pose as a resumable job: stage A enqueues everything up to the read of the segment offsets,
// stage B (once those are on the host) the pairings and the read of the poses, finish picks the winner.
// cb_sym_pose runs the three stages with two waits in between; cb_sym_pose_many keeps several jobs in
// flight from ONE host thread and polls their events.
namespace {

struct SymJob {
    cb_sym_job in;
    cudaStream_t st = nullptr;
    int K = 2, P = 0, n_sub = 0, status = 3, n_reg = 1;
    RegArgs a{};
    int32_t drawn[2 * 1024];
    const int32_t *samples_base = nullptr, *samples_pos = nullptr;
    size_t in_bytes = 0;
    char* d_in = nullptr;
    int64_t *a_perm = nullptr, *b_perm = nullptr;
    int32_t* d_off = nullptr;
    double *T_all = nullptr, *chamf = nullptr;
    float *f0s = nullptr, *f1s = nullptr;
    char* pin_in = nullptr;
    int32_t* pin_off = nullptr;
    double* pin_res = nullptr;
    bool forked = false;
    unsigned used = 0;
};

// nothing of this job may still run when the caller sees the error
int sym_bail(cb_ctx* ctx, SymJob& j, int code) {
    if (j.forked)
        for (int s = 0; s < j.n_sub; ++s) cudaStreamSynchronize(ctx->sub_stream[s]);
    cudaStreamSynchronize(j.st);
    return code;
}

// registration `slot` (0: vanilla) on sub-stream slot % n_sub, ordered after the work on st so far
int sym_run(cb_ctx* ctx, SymJob& j, int slot, const SegView* sv, const int32_t* pair_map) {
    cb_ctx* c = ctx;
    cudaStream_t s = j.st;
    if (j.n_sub > 0) {
        const int q = slot % j.n_sub;
        c = ctx->sub[q];
        s = ctx->sub_stream[q];
        if (cudaEventRecord(ctx->fork_event, j.st) != cudaSuccess || cudaStreamWaitEvent(s, ctx->fork_event, 0) != cudaSuccess)
            CB_FAIL(ctx, CB_ECUDA, "cb_sym_pose: fork failed: %s", cudaGetErrorString(cudaGetLastError()));
        j.forked = true;
        j.used |= 1u << q;
    }
    const int r = register_one(c, s, j.a, sv, pair_map, j.T_all + 16 * slot, j.chamf + slot, nullptr);
    if (r && c != ctx) snprintf(ctx->err, sizeof(ctx->err), "%s", c->err);
    return r;
}

int sym_stage_a(cb_ctx* ctx, cudaStream_t st, SymJob& j) {
    int rc = cb_begin(ctx);
    if (rc) return rc;
    const cb_sym_job& in = j.in;
    j.st = st;
    if (!in.base_f || !in.base_xyz || !in.pos_f || !in.pos_xyz || !in.h_out || in.n0 <= 0 || in.n1 <= 0 || in.d <= 0 ||
        in.k_nn <= 0 || in.n_samples < 0)
        CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose: bad arguments");
    if (in.k_nn > in.n1) CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose: k_nn=%d exceeds the %d target points", in.k_nn, in.n1);
    const int n0 = in.n0, n1 = in.n1, d = in.d;
    const int K = j.K = in.pos_sym >= 2 ? 4 : 2;
    int n_samples = in.n_samples;
    j.samples_base = in.h_samples_base;
    j.samples_pos = in.h_samples_pos;
    if (n_samples > 0 && !j.samples_base && !j.samples_pos && in.np_seed >= 0 && n_samples <= 1024) {
        // the draws of np.random.seed(np_seed); choice(n0, ..); choice(n1, ..) (utils/symmetry.py:193);
        // a cloud with fewer points than samples makes the reference raise and fall back (:299-301)
        const int32_t ns[2] = {n0, n1};
        if (cb_np_seeded_choices((uint32_t)in.np_seed, ns, 2, n_samples, j.drawn) == CB_OK) {
            j.samples_base = j.drawn;
            j.samples_pos = j.drawn + n_samples;
        } else {
            n_samples = 0;
        }
    }
    const int P = j.P = n_samples > 0 ? (in.pos_sym >= 2 ? 2 * K : K) : 0;
    if (n_samples > 0) {
        if (!j.samples_base || !j.samples_pos || !in.h_km_first || !in.h_km_unif || in.n_init < 1 || in.n_init > 16 ||
            in.n_trials < 1 || in.n_trials > 4 || in.n_nn < 1)
            CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose: the segmentation needs its sample rows and k-means++ draws");
        if (n_samples > 1024) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_sym_pose: at most 1024 seed samples (got %d)", n_samples);
    }
    j.a = RegArgs{in.base_f, in.base_xyz, in.pos_f, in.pos_xyz, n0, n1, d, in.k_nn, in.max_dist, in.seed, in.h_total,
                  in.adaptive, in.confidence};
    const RegArgs& a = j.a;
    const int n_sub = j.n_sub = ctx->timing ? 0 : sym_default_streams();  // per-class event timing needs one stream
    if ((rc = sym_prepare(ctx, n_sub))) return rc;

    // host -> pinned -> device: sample rows of both clouds and the k-means++ draws, in one copy
    const size_t n_unif = (size_t)in.n_init * (K - 1) * in.n_trials;
    const size_t in_bytes = j.in_bytes = n_samples > 0 ? cb_align((size_t)2 * n_samples * 4 + (size_t)in.n_init * 4) + n_unif * 8 : 0;
    const size_t off_bytes = 256, res_bytes = (size_t)(1 + 8) * 17 * 8;
    if (in_bytes + off_bytes + res_bytes > kPinnedBytes) CB_FAIL(ctx, CB_EUNSUPPORTED, "cb_sym_pose: staging too small");
    j.pin_in = ctx->pinned;
    j.pin_off = reinterpret_cast<int32_t*>(ctx->pinned + cb_align(in_bytes));
    j.pin_res = reinterpret_cast<double*>(ctx->pinned + cb_align(in_bytes) + off_bytes);

    const int nmax = n0 > n1 ? n0 : n1;
    size_t need = cb_align(in_bytes) + 2 * cb_align((size_t)nmax * 8) + 512 + cb_align((size_t)(1 + P) * 17 * 8) +
                  cb_align((size_t)n0 * d * 4) + cb_align((size_t)n1 * d * 4) + 8192;
    size_t scratch = n_samples > 0 ? cb_segment_workspace(nmax, n_samples, in.n_nn, K) : 0;
    if (n_sub == 0 && reg_one_workspace(ctx, a) > scratch) scratch = reg_one_workspace(ctx, a);
    if ((rc = cb_reserve(ctx, need + scratch + (1 << 20)))) return rc;
    for (int q = 0; q < n_sub; ++q) {
        cb_ctx* c = ctx->sub[q];
        c->knn_variant = ctx->knn_variant;
        c->ransac_variant = ctx->ransac_variant;
        size_t sub_need = reg_one_workspace(c, a);
        if (q == 1 && scratch > sub_need) sub_need = scratch;  // sub-context 1 also runs the pos segmentation
        if ((rc = cb_begin(c)) || (rc = cb_reserve(c, sub_need + (1 << 20)))) {
            snprintf(ctx->err, sizeof(ctx->err), "%s", c->err);
            return rc;
        }
    }
    j.d_in = in_bytes ? reinterpret_cast<char*>(cb_alloc(ctx, in_bytes)) : nullptr;
    j.a_perm = cb_alloc_t<int64_t>(ctx, n0);
    j.b_perm = cb_alloc_t<int64_t>(ctx, n1);
    j.d_off = cb_alloc_t<int32_t>(ctx, 64);
    j.T_all = cb_alloc_t<double>(ctx, (size_t)(1 + P) * 17);
    j.chamf = j.T_all + (size_t)(1 + P) * 16;
    j.f0s = cb_alloc_t<float>(ctx, (size_t)n0 * d);
    j.f1s = cb_alloc_t<float>(ctx, (size_t)n1 * d);
    if ((in_bytes && !j.d_in) || !j.a_perm || !j.b_perm || !j.d_off || !j.T_all || !j.f0s || !j.f1s) return CB_ENOMEM;

    if (in_bytes) {
        int32_t* p32 = reinterpret_cast<int32_t*>(j.pin_in);
        memcpy(p32, j.samples_base, (size_t)n_samples * 4);
        memcpy(p32 + n_samples, j.samples_pos, (size_t)n_samples * 4);
        memcpy(p32 + 2 * n_samples, in.h_km_first, (size_t)in.n_init * 4);
        memcpy(j.pin_in + cb_align((size_t)2 * n_samples * 4 + (size_t)in.n_init * 4), in.h_km_unif, n_unif * 8);
        CB_CUDA(ctx, cudaMemcpyAsync(j.d_in, j.pin_in, in_bytes, cudaMemcpyHostToDevice, st));
        ctx->h2d_bytes += (int64_t)in_bytes;
    }
    // vanilla registration (utils/symmetry.py:267-283); on its own stream it overlaps the segmentation
    if ((rc = sym_run(ctx, j, 0, nullptr, nullptr))) return sym_bail(ctx, j, rc);

    j.status = 3;  // 0 both segmentations accepted a model, 1 base failed, 2 pos failed, 3 not attempted
    if (n_samples > 0) {
        const int32_t* d_s = reinterpret_cast<const int32_t*>(j.d_in);
        const int32_t* d_first = d_s + 2 * n_samples;
        const double* d_unif = reinterpret_cast<const double*>(j.d_in + cb_align((size_t)2 * n_samples * 4 + (size_t)in.n_init * 4));
        const size_t mark = ctx->arena_off;
        const int nn0 = in.n_nn < n0 ? in.n_nn : n0, nn1 = in.n_nn < n1 ? in.n_nn : n1;
        // the pos cloud's segmentation is independent of the base cloud's: second internal stream
        if (n_sub >= 2) {
            cb_ctx* c = ctx->sub[1];
            cudaStream_t s = ctx->sub_stream[1];
            if (cudaEventRecord(ctx->fork_event, st) != cudaSuccess || cudaStreamWaitEvent(s, ctx->fork_event, 0) != cudaSuccess)
                return sym_bail(ctx, j, CB_ECUDA);
            j.forked = true;
            j.used |= 2u;
            const size_t m1 = c->arena_off;
            rc = cb_segment_impl(c, in.pos_f, in.pos_xyz, n1, d, K, d_s + n_samples, n_samples, nn1, d_first, d_unif, in.n_init,
                                 in.n_trials, in.min_center, j.b_perm, j.d_off + 8, s);
            c->arena_off = m1;
            if (rc) {
                snprintf(ctx->err, sizeof(ctx->err), "%s", c->err);
                return sym_bail(ctx, j, rc);
            }
            rc = cb_segment_impl(ctx, in.base_f, in.base_xyz, n0, d, K, d_s, n_samples, nn0, d_first, d_unif, in.n_init,
                                 in.n_trials, in.min_center, j.a_perm, j.d_off, st);
            ctx->arena_off = mark;
            if (rc) return sym_bail(ctx, j, rc);
            if (cudaEventRecord(ctx->sub_event[1], s) != cudaSuccess || cudaStreamWaitEvent(st, ctx->sub_event[1], 0) != cudaSuccess)
                return sym_bail(ctx, j, CB_ECUDA);
        } else {
            rc = cb_segment_impl(ctx, in.base_f, in.base_xyz, n0, d, K, d_s, n_samples, nn0, d_first, d_unif, in.n_init,
                                 in.n_trials, in.min_center, j.a_perm, j.d_off, st);
            ctx->arena_off = mark;
            if (rc) return sym_bail(ctx, j, rc);
            rc = cb_segment_impl(ctx, in.pos_f, in.pos_xyz, n1, d, K, d_s + n_samples, n_samples, nn1, d_first, d_unif,
                                 in.n_init, in.n_trials, in.min_center, j.b_perm, j.d_off + 8, st);
            ctx->arena_off = mark;
            if (rc) return sym_bail(ctx, j, rc);
        }
        if (cudaMemcpyAsync(j.pin_off, j.d_off, 16 * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess) {
            snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: segment offsets: %s", cudaGetErrorString(cudaGetLastError()));
            return sym_bail(ctx, j, CB_ECUDA);
        }
        ctx->d2h_bytes += 16 * 4;
        j.status = -1;  // decided in stage B from the offsets
    }
    return CB_OK;
}

// the caller has waited for the work of stage A on j.st
int sym_stage_b(cb_ctx* ctx, SymJob& j) {
    int rc;
    const int K = j.K, P = j.P;
    cudaStream_t st = j.st;
    if (j.status == -1) j.status = j.pin_off[K + 1] < 0 ? 1 : (j.pin_off[8 + K + 1] < 0 ? 2 : 0);
    j.n_reg = 1;
    if (j.status == 0) {
        if ((rc = gather_segment_rows(ctx, st, j.a, j.a_perm, j.b_perm, j.f0s, j.f1s))) return sym_bail(ctx, j, rc);
        const SegView sv{j.f0s, j.f1s, j.a_perm, j.b_perm, j.pin_off, j.pin_off + 8, K};
        // pairings of utils/symmetry.py:303-356: the K cyclic rotations of the pos segments, then
        // (pos_sym >= 2) those of [B0, B3, B2, B1]
        static const int order2[4] = {0, 3, 2, 1};
        for (int p = 0; p < P; ++p) {
            int32_t map[4];
            for (int s = 0; s < K; ++s) map[s] = p < K ? (s + p) % K : order2[(s + (p - K)) % K];
            if ((rc = sym_run(ctx, j, 1 + p, &sv, map))) return sym_bail(ctx, j, rc);
        }
        j.n_reg = 1 + P;
    }
    // join, read the poses back
    for (int q = 0; q < j.n_sub; ++q)
        if (j.used >> q & 1) {
            if (cudaEventRecord(ctx->sub_event[q], ctx->sub_stream[q]) != cudaSuccess ||
                cudaStreamWaitEvent(st, ctx->sub_event[q], 0) != cudaSuccess) {
                snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: join failed: %s", cudaGetErrorString(cudaGetLastError()));
                return sym_bail(ctx, j, CB_ECUDA);
            }
            ctx->launches += ctx->sub[q]->launches;
            ctx->sub[q]->launches = 0;
        }
    if (cudaMemcpyAsync(j.pin_res, j.T_all, (size_t)(1 + P) * 17 * 8, cudaMemcpyDeviceToHost, st) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: result read failed: %s", cudaGetErrorString(cudaGetLastError()));
        return sym_bail(ctx, j, CB_ECUDA);
    }
    ctx->d2h_bytes += (int64_t)(1 + P) * 17 * 8;
    return CB_OK;
}

// the caller has waited for the work of stage B on j.st: keep the strictly lowest Chamfer (:321,:353)
void sym_finish(SymJob& j) {
    const int P = j.P;
    const double* ch = j.pin_res + (size_t)(1 + P) * 16;
    int best = 0;
    for (int i = 1; i < j.n_reg; ++i)
        if (ch[best] > ch[i]) best = i;
    double* h_out = j.in.h_out;
    memcpy(h_out, j.pin_res + 16 * best, 16 * 8);
    h_out[16] = ch[best];
    memcpy(h_out + 17, j.pin_res, 16 * 8);
    h_out[33] = ch[0];
    h_out[34] = (double)j.status;
    h_out[35] = (double)best;
    h_out[36] = (double)(j.n_reg - 1);
}

int sym_run_blocking(cb_ctx* ctx, cudaStream_t st, SymJob& j) {
    int rc = sym_stage_a(ctx, st, j);
    if (rc) return rc;
    if (j.status == -1 && sym_wait(ctx, st) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: segment offsets: %s", cudaGetErrorString(cudaGetLastError()));
        return sym_bail(ctx, j, CB_ECUDA);
    }
    if ((rc = sym_stage_b(ctx, j))) return rc;
    if (sym_wait(ctx, st) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "cb_sym_pose: result read failed: %s", cudaGetErrorString(cudaGetLastError()));
        return sym_bail(ctx, j, CB_ECUDA);
    }
    sym_finish(j);
    return CB_OK;
}

}  // namespace

extern "C" int cb_sym_pose(cb_ctx* ctx, const float* base_f, const float* base_xyz, int n0, const float* pos_f,
                           const float* pos_xyz, int n1, int d, int pos_sym, int k_nn, const int32_t* h_samples_base,
                           const int32_t* h_samples_pos, int n_samples, int64_t np_seed, int n_nn,
                           const int32_t* h_km_first,
                           const double* h_km_unif, int n_init, int n_trials, double min_center, double max_dist,
                           uint64_t seed, int h_total, int adaptive, double confidence, double* h_out, void* stream) {
    if (!ctx) return CB_EINVAL;
    SymJob j;
    j.in = cb_sym_job{base_f, base_xyz, pos_f, pos_xyz, n0, n1, d, pos_sym, k_nn, n_samples, n_nn, n_init, n_trials, h_total,
                      adaptive, 0, np_seed, h_samples_base, h_samples_pos, h_km_first, h_km_unif, min_center, max_dist,
                      confidence, seed, h_out};
    return sym_run_blocking(ctx, (cudaStream_t)stream, j);
}

// Several independent pairs (evaluation.py:297-313 is such a loop) kept in flight by ONE host thread:
// `concurrency` slots, each a context with its own stream, arena, pinned staging and internal streams.
// The thread enqueues a job's stage, records an event and moves on; it comes back to the slot when the
// event has completed (cudaEventQuery), so no thread ever blocks inside the driver while other pairs
// have work to enqueue.  Round 1 ran this loop on five Python threads spinning in
// cudaStreamSynchronize (six host threads per GPU on four cores per rank at 8 GPUs).
extern "C" int cb_sym_pose_many(cb_ctx* ctx, const cb_sym_job* jobs, int n_jobs, int concurrency, int32_t* h_status,
                                void* stream) {
    if (!ctx) return CB_EINVAL;
    ctx->err[0] = 0;
    if (n_jobs < 0 || (n_jobs > 0 && !jobs)) CB_FAIL(ctx, CB_EINVAL, "cb_sym_pose_many: bad arguments");
    CB_CUDA(ctx, cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    // a pair that never starts (an earlier one failed) keeps CB_ENOTRUN and gets a recognisable h_out
    for (int i = 0; i < n_jobs && h_status; ++i) h_status[i] = CB_ENOTRUN;
    auto mark_not_run = [&](int from) {
        for (int i = from; i < n_jobs; ++i)
            if (jobs[i].h_out) {
                for (int t = 0; t < 40; ++t) jobs[i].h_out[t] = 0.0;
                jobs[i].h_out[34] = -1.0;
            }
    };
    if (n_jobs == 0) return CB_OK;
    if (ctx->timing || concurrency <= 1) {
        // per-class event timing needs one stream: one pair after the other on the caller's
        for (int i = 0; i < n_jobs; ++i) {
            SymJob j;
            j.in = jobs[i];
            const int rc = sym_run_blocking(ctx, st, j);
            if (h_status) h_status[i] = rc;
            if (rc) {
                mark_not_run(i + 1);
                return rc;
            }
        }
        return CB_OK;
    }
    const int W = std::min(std::min(concurrency, 32), n_jobs);  // every slot owns an arena, staging and streams
    while ((int)ctx->slots.size() < W) {
        cb_ctx* c = nullptr;
        cudaStream_t s = nullptr;
        cudaEvent_t e = nullptr;
        if (cb_create(&c, ctx->device, size_t(64) << 20) != CB_OK) CB_FAIL(ctx, CB_ENOMEM, "cb_sym_pose_many: slot context");
        if (cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking) != cudaSuccess ||
            cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) {
            cb_destroy(c);
            CB_FAIL(ctx, CB_ECUDA, "cb_sym_pose_many: stream / event creation failed");
        }
        ctx->slots.push_back(c);
        ctx->slot_stream.push_back(s);
        ctx->slot_event.push_back(e);
    }
    if (!ctx->fork_event) CB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->fork_event, cudaEventDisableTiming));
    // the inputs were produced on the caller's stream
    CB_CUDA(ctx, cudaEventRecord(ctx->fork_event, st));
    for (int w = 0; w < W; ++w) {
        CB_CUDA(ctx, cudaStreamWaitEvent(ctx->slot_stream[w], ctx->fork_event, 0));
        ctx->slots[w]->knn_variant = ctx->knn_variant;
        ctx->slots[w]->ransac_variant = ctx->ransac_variant;
    }
    std::vector<SymJob> job(W);
    std::vector<int> state(W, 0), which(W, -1);  // 0 idle, 1 stage A in flight, 2 stage B in flight
    int next = 0, done = 0, first_rc = CB_OK;
    auto fail = [&](int w, int rc) {
        if (h_status) h_status[which[w]] = rc;
        if (first_rc == CB_OK) {
            first_rc = rc;
            snprintf(ctx->err, sizeof(ctx->err), "pair %d: %s", which[w], ctx->slots[w]->err);
        }
        state[w] = 0;
        ++done;
    };
    while (done < n_jobs) {
        bool progress = false;
        for (int w = 0; w < W; ++w) {
            cb_ctx* c = ctx->slots[w];
            cudaStream_t s = ctx->slot_stream[w];
            if (state[w] == 0) {
                if (next >= n_jobs || first_rc != CB_OK) {
                    if (first_rc != CB_OK && next < n_jobs) {
                        mark_not_run(next);
                        done += n_jobs - next;
                        next = n_jobs;
                    }
                    continue;
                }
                which[w] = next++;
                job[w] = SymJob();
                job[w].in = jobs[which[w]];
                progress = true;
                int rc = sym_stage_a(c, s, job[w]);
                if (rc == CB_OK && job[w].status != -1) rc = sym_stage_b(c, job[w]);  // no segmentation: nothing to wait for
                if (rc != CB_OK || cudaEventRecord(ctx->slot_event[w], s) != cudaSuccess) { fail(w, rc ? rc : CB_ECUDA); continue; }
                state[w] = job[w].status == -1 ? 1 : 2;
            } else {
                const cudaError_t q = cudaEventQuery(ctx->slot_event[w]);
                if (q == cudaErrorNotReady) continue;
                progress = true;
                if (q != cudaSuccess) {
                    snprintf(c->err, sizeof(c->err), "device error: %s", cudaGetErrorString(q));
                    sym_bail(c, job[w], CB_ECUDA);
                    fail(w, CB_ECUDA);
                    continue;
                }
                if (state[w] == 1) {
                    const int rc = sym_stage_b(c, job[w]);
                    if (rc != CB_OK || cudaEventRecord(ctx->slot_event[w], s) != cudaSuccess) { fail(w, rc ? rc : CB_ECUDA); continue; }
                    state[w] = 2;
                } else {
                    sym_finish(job[w]);
                    if (h_status) h_status[which[w]] = CB_OK;
                    state[w] = 0;
                    ++done;
                }
            }
        }
        if (!progress) {
#if defined(__x86_64__)
            __builtin_ia32_pause();
#endif
        }
    }
    for (int w = 0; w < W; ++w) {
        cb_ctx* c = ctx->slots[w];
        ctx->launches += c->launches;
        ctx->h2d_bytes += c->h2d_bytes;
        ctx->d2h_bytes += c->d2h_bytes;
        c->launches = 0;
        c->h2d_bytes = c->d2h_bytes = 0;
    }
    return first_rc;
}
