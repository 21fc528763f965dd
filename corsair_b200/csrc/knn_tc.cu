// tcgen05 candidate pass of the kNN engine (K1): the distance GEMM on the 5th-generation tensor
// cores in split precision, fused with the running top-KC of every query row.
//
// Split-bf16 ("bf16x3"): q = qh + ql, r = rh + rl (bf16 each);  q.r ~= qh.rh + qh.rl + ql.rh.
// The three products are ONE GEMM over a widened K axis:
//     A row (query)     [ qh (D) | qh (D) | ql (D) | 1 1 1 0 ... (16) ]
//     B row (reference) [ rh (D) | rl (D) | rh (D) | c0 c1 c2 0 ... (16) ],  c0+c1+c2 = -|r|^2/2
// so the accumulator is dot' = q.r - |r|^2/2 and the squared distance ordering is simply
// "largest dot' first" (score s = -2 dot' = |r|^2 - 2 q.r, the candidate-list convention of
// knn.cu).  Error: |dot' - exact| <= 3*2^-18 |q||r| + fp32 accumulation, covered by
// knn_cand_tc_gamma() in the re-rank's proof obligation.
//
// Pipeline per CTA (one 128-query tile x one reference chunk), 320 threads:
//   warp 8  producer: cp.async.bulk (TMA engine) of pre-packed operand tiles -> shared memory
//   warp 9  one elected thread issues tcgen05.mma (M=128, N=256, K=16 per instruction) into one
//           of two 256-column TMEM accumulators and commits to mbarriers
//   warps 0-7 epilogue: warp w drains lane quadrant w%4 (one query row per thread), column half w/4
//           of every accumulator, tcgen05.ld 32 columns at a time — two warps per SM sub-partition,
//           so one warp's tensor-memory load overlaps the other's arithmetic.  The two threads of a
//           row merge their pass-1 lists through shared memory (same threshold as one thread over
//           all columns would find) and share the row's candidate list through a shared counter.
// The chunk's GEMM runs TWICE (the tensor pipe is nearly idle, the operands sit in L2):
//   pass 1  per row, the maxima of column groups (16 or 8 columns) go through a branch-free
//           sorted insertion into the row's KC best group maxima; their KC-th is the row's
//           threshold tau: at least KC references of the chunk score at or above it
//   pass 2  every reference with score >= tau is appended to the row's candidate list: a branch-free
//           32-bit qualification mask per tcgen05.ld, then one short loop over its set bits.
// A per-lane sorted top-k over all columns would make the warp take the insertion path whenever
// ANY of its 32 rows inserts, i.e. for nearly every column of a cold list (measured: 1 400
// instructions per 32 columns); the threshold form costs ~0.6 + ~4 instructions per element.
#include "common.cuh"
#include "knn_internal.cuh"
#include "tc_common.cuh"

#include <math.h>

namespace {

constexpr int kTileM = 128;   // queries per CTA (TMEM lanes)
constexpr int kTileN = 256;   // references per MMA / accumulator stage
constexpr int kExtraK = 16;   // the norm block of the widened K axis
constexpr int kEpiWarps = 8;
constexpr int kThreads = (kEpiWarps + 2) * 32;

__host__ __device__ constexpr int kprime(int D) { return 3 * D + kExtraK; }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// fp32 rows [n, D] -> packed bf16 operand tiles.  One thread per (row, 8-column group).
// is_b: 0 = A layout (hi|hi|lo|ones), 1 = B layout (hi|lo|hi|-norm/2 split in three).
// B rows are stored at the scrambled position (row * perm_a) mod n_pad (perm_a coprime to n_pad;
// 1 for A): the column groups of the threshold pass then hold pseudo-random subsets of the
// references, whatever spatial order (and feature similarity between neighbours) the cloud has.
__global__ void knn_tc_pack_kernel(const float* __restrict__ x, const float* __restrict__ norms, int n, int n_pad,
                                   int D, int TR, int is_b, int maxdot, unsigned perm_a,
                                   __nv_bfloat16* __restrict__ out) {
    const int KP = kprime(D);
    const int groups = KP / 8;
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= (size_t)n_pad * groups) return;
    const int row = (int)(gid / groups), kc = (int)(gid % groups);
    __nv_bfloat16 v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) v[e] = __float2bfloat16_rn(0.f);
    const int k0 = kc * 8;
    if (k0 < 3 * D) {
        if (row < n) {
            const int part = k0 / D;          // 0,1,2
            const int c0 = k0 - part * D;     // column inside the part
            const bool want_lo = is_b ? (part == 1) : (part == 2);
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                __nv_bfloat16 hi, lo;
                tc::split2(x[(size_t)row * D + c0 + e], hi, lo);
                v[e] = want_lo ? lo : hi;
            }
        }
    } else if (k0 == 3 * D) {
        if (!is_b) {
            if (row < n) { v[0] = v[1] = v[2] = __float2bfloat16_rn(1.f); }
        } else if (row < n) {
            if (!maxdot) tc::split3(-0.5f * norms[row], v[0], v[1], v[2]);
        } else {
            v[0] = __float2bfloat16_rn(-1e30f);  // padding references can never enter a list
        }
    }
    const int prow = (int)(((unsigned long long)row * perm_a) % (unsigned)n_pad);
    const int tile = prow / TR, r = prow % TR;
    char* dst = reinterpret_cast<char*>(out) + (size_t)tile * TR * KP * 2 + tc::tile_offset(TR, r, kc);
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
}

struct TcSmemCtl {
    uint64_t a_full, full[2], empty[2], tmem_full[2], tmem_empty[2];
    uint32_t tmem_base;
};

template <int D, int KC>
__global__ void __launch_bounds__(kThreads, 1)
knn_tc_kernel(const __nv_bfloat16* __restrict__ Apack, const __nv_bfloat16* __restrict__ Bpack, int nq, int nr,
              int chunk, int S, int maxdot, unsigned perm_inv, int nr_pad, int32_t* __restrict__ cand_idx,
              float* __restrict__ bound) {
    constexpr int KP = kprime(D);
    constexpr int LC = 3 * KC;            // list slots per (query, chunk)
    constexpr int GROUP = 128 / KC;       // columns per group maximum: 16 (KC = 8) or 8 (KC = 16)
    constexpr uint32_t A_BYTES = kTileM * KP * 2, B_BYTES = kTileN * KP * 2;
    constexpr uint32_t LBO_A = (kTileM / 8) * 128, LBO_B = (kTileN / 8) * 128;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;
    unsigned char* sB = smem + A_BYTES;
    TcSmemCtl* ctl = reinterpret_cast<TcSmemCtl*>(smem + A_BYTES + 2 * B_BYTES);
    float* s_top = reinterpret_cast<float*>(smem + A_BYTES + 2 * B_BYTES + 128);  // [2][kTileM][KC]
    int* s_cnt = reinterpret_cast<int*>(s_top + 2 * kTileM * KC);                // [kTileM]
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int c = blockIdx.y;
    const int t0 = (c * chunk) / kTileN;
    const int t1 = min((nr + kTileN - 1) / kTileN, ((c + 1) * chunk) / kTileN);
    const int ntiles = t1 - t0;
    const int nsteps = 2 * ntiles;  // the chunk's tiles, twice

    if (threadIdx.x == 0) {
        tc::mbar_init(&ctl->a_full, 1);
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(&ctl->full[i], 1);
            tc::mbar_init(&ctl->empty[i], 1);
            tc::mbar_init(&ctl->tmem_full[i], 1);
            tc::mbar_init(&ctl->tmem_empty[i], kEpiWarps);
        }
        tc::fence_barrier_init();
    }
    if (threadIdx.x < kTileM) s_cnt[threadIdx.x] = 0;
    if (warp == kEpiWarps) tc::tmem_alloc(&ctl->tmem_base, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tbase = ctl->tmem_base;

    if (warp == kEpiWarps) {
        // ===== producer: bulk copies of pre-packed tiles =====
        if (lane == 0) {
            tc::mbar_arrive_expect_tx(&ctl->a_full, A_BYTES);
            tc::bulk_g2s(sA, reinterpret_cast<const char*>(Apack) + (size_t)blockIdx.x * A_BYTES, A_BYTES, &ctl->a_full);
            for (int it = 0; it < nsteps; ++it) {
                const int st = it & 1, use = it >> 1;
                const int tile = t0 + (it >= ntiles ? it - ntiles : it);
                tc::mbar_wait(&ctl->empty[st], (use & 1) ^ 1);
                tc::mbar_arrive_expect_tx(&ctl->full[st], B_BYTES);
                tc::bulk_g2s(sB + st * B_BYTES, reinterpret_cast<const char*>(Bpack) + (size_t)tile * B_BYTES, B_BYTES,
                             &ctl->full[st]);
            }
        }
    } else if (warp == kEpiWarps + 1) {
        // ===== MMA issuer: one thread =====
        if (lane == 0) {
            constexpr uint32_t idesc = tc::idesc_bf16(kTileM, kTileN);
            const uint32_t a_addr = tc::smem_u32(sA);
            tc::mbar_wait(&ctl->a_full, 0);
            for (int it = 0; it < nsteps; ++it) {
                const int st = it & 1, use = it >> 1;
                tc::mbar_wait(&ctl->full[st], use & 1);
                tc::mbar_wait(&ctl->tmem_empty[st], (use & 1) ^ 1);
                tc::fence_after_sync();
                const uint32_t b_addr = tc::smem_u32(sB + st * B_BYTES);
#pragma unroll
                for (int ks = 0; ks < KP / 16; ++ks) {
                    const uint64_t ad = tc::smem_desc(a_addr + ks * 2 * LBO_A, LBO_A, 128);
                    const uint64_t bd = tc::smem_desc(b_addr + ks * 2 * LBO_B, LBO_B, 128);
                    tc::mma_bf16(tbase + st * kTileN, ad, bd, idesc, ks > 0);
                }
                tc::mma_commit(&ctl->empty[st]);      // smem stage reusable once these MMAs finish
                tc::mma_commit(&ctl->tmem_full[st]);  // accumulator ready for the epilogue
            }
        }
    } else {
        // ===== epilogue: 256 threads = 128 query rows (TMEM lanes) x 2 column halves =====
        const int half = warp >> 2;
        const int row = (warp & 3) * 32 + lane;
        const int qi = blockIdx.x * kTileM + row;
        const size_t o = ((size_t)min(qi, nq - 1) * S + c) * LC;
        const bool live = qi < nq;
        constexpr int kHalfN = kTileN / 2;
        const uint32_t lane_addr = tbase + ((uint32_t)((warp & 3) * 32) << 16) + half * kHalfN;
        float top[KC];  // the KC best group maxima of this thread's columns, descending
#pragma unroll
        for (int t = 0; t < KC; ++t) top[t] = -INFINITY;
        float tau = -INFINITY;
        for (int it = 0; it < nsteps; ++it) {
            const int st = it & 1, use = it >> 1;
            const bool second = it >= ntiles;
            if (it == ntiles) {
                // the row's threshold: KC-th best of the union of both halves' group maxima
#pragma unroll
                for (int t = 0; t < KC; ++t) s_top[(half * kTileM + row) * KC + t] = top[t];
                named_bar_sync(1, kEpiWarps * 32);
#pragma unroll
                for (int u = 0; u < KC; ++u) {
                    float x = s_top[((half ^ 1) * kTileM + row) * KC + u];
#pragma unroll
                    for (int t = 0; t < KC; ++t) {
                        const float hi = fmaxf(top[t], x);
                        x = fminf(top[t], x);
                        top[t] = hi;
                    }
                }
                tau = fmaxf(top[KC - 1], -1e29f);  // padding (-1e30) never qualifies
            }
            tc::mbar_wait(&ctl->tmem_full[st], use & 1);
            tc::fence_after_sync();
            const int col0 = (t0 + (second ? it - ntiles : it)) * kTileN + half * kHalfN;
            // one 32-column chunk of this thread's half of the stage
            auto process = [&](const float (&v)[32], int cb) {
                if (!second) {
#pragma unroll
                    for (int g = 0; g < 32 / GROUP; ++g) {
                        float x = v[g * GROUP];
#pragma unroll
                        for (int e = 1; e < GROUP; ++e) x = fmaxf(x, v[g * GROUP + e]);
                        // branch-free insertion into the descending list (x sinks to its place)
#pragma unroll
                        for (int t = 0; t < KC; ++t) {
                            const float hi = fmaxf(top[t], x);
                            x = fminf(top[t], x);
                            top[t] = hi;
                        }
                    }
                } else {
                    // which of the 32 columns qualify: a bit mask, two instructions per column on two
                    // pipes — the sign bit of v - tau (set: below the threshold) is shifted into the word
                    // by one funnel shift; column e lands at bit 31 - e, hence the bit reversal
                    unsigned below = 0;
#pragma unroll
                    for (int e = 0; e < 32; ++e) below = __funnelshift_l(__float_as_uint(v[e] - tau), below, 1);
                    unsigned mask = __brev(~below);
                    // padding rows (all-equal scores) and rows whose list already overflowed (they go
                    // to the exact fallback whatever else qualifies) must not spin in the loop below
                    if (!live || (mask && *(volatile int*)&s_cnt[row] > LC)) mask = 0;
                    // append their reference indices (the re-rank recomputes every candidate exactly,
                    // the approximate scores are not needed); ~1 % of the columns qualify
                    while (mask) {
                        const int e = __ffs(mask) - 1;
                        mask &= mask - 1;
                        const int pos = atomicAdd(&s_cnt[row], 1);
                        if (pos < LC)
                            cand_idx[o + pos] = (int)(((unsigned long long)(col0 + cb * 32 + e) * perm_inv) % (unsigned)nr_pad);
                    }
                }
            };
            // software-pipelined drain: the next 32 columns are in flight while these are processed;
            // the stage goes back to the MMA warp as soon as its last load has landed
            static_assert(kHalfN / 32 == 4, "four chunks per stage and thread");
            float va[32], vb[32];
            tc::tmem_ld32(lane_addr + st * kTileN, va);
            tc::tmem_ld_wait32(va);
            tc::tmem_ld32(lane_addr + st * kTileN + 32, vb);
            process(va, 0);
            tc::tmem_ld_wait32(vb);
            tc::tmem_ld32(lane_addr + st * kTileN + 64, va);
            process(vb, 1);
            tc::tmem_ld_wait32(va);
            tc::tmem_ld32(lane_addr + st * kTileN + 96, vb);
            process(va, 2);
            tc::tmem_ld_wait32(vb);
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&ctl->tmem_empty[st]);
            process(vb, 3);
        }
        named_bar_sync(1, kEpiWarps * 32);  // both halves of every row have appended
        if (live && half == 0) {
            const int cnt = s_cnt[row];
            for (int t = min(cnt, LC); t < LC; ++t) cand_idx[o + t] = -1;
            // no dropped reference scores below the threshold; an overflowing list proves nothing
            bound[(size_t)qi * S + c] = cnt > LC ? -INFINITY : (maxdot ? -tau : -2.f * tau);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == kEpiWarps) tc::tmem_dealloc(tbase, 512);
}

template <int D, int KC>
cudaError_t launch_tc(const __nv_bfloat16* A, const __nv_bfloat16* B, int nq, int nr, int chunk, int S, int maxdot,
                      unsigned perm_inv, int nr_pad, int32_t* ci, float* bound, cudaStream_t st) {
    constexpr int KP = kprime(D);
    const size_t smem = (size_t)kTileM * KP * 2 + 2 * (size_t)kTileN * KP * 2 + 128 + 2 * kTileM * KC * 4 + kTileM * 4 + 64;
    cudaError_t e = cudaFuncSetAttribute(knn_tc_kernel<D, KC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    dim3 grid((nq + kTileM - 1) / kTileM, S);
    knn_tc_kernel<D, KC><<<grid, kThreads, smem, st>>>(A, B, nq, nr, chunk, S, maxdot, perm_inv, nr_pad, ci, bound);
    return cudaGetLastError();
}

}  // namespace

double knn_cand_tc_gamma(int) { return 1.5e-5; }

extern "C" const char* cb_knn_variant(const cb_ctx* ctx) {
    return (ctx && ctx->knn_variant == 1) ? "simt" : "tcgen05";
}

int knn_cand_tc_launch(cb_ctx* ctx, const float* q, int nq, const float* r, const float* rn, int nr, int d, int KC,
                       int LC, int chunk, int S, int maxdot, int32_t* cand_idx, float* cand_s, float* bound,
                       cudaStream_t st) {
    if ((d != 16 && d != 32) || (KC != 8 && KC != 16) || LC != 3 * KC || chunk % kTileN != 0)
        CB_FAIL(ctx, CB_EUNSUPPORTED, "knn_cand_tc_launch: unsupported shape (d=%d KC=%d LC=%d chunk=%d)", d, KC, LC, chunk);
    const int KP = kprime(d);
    const int nq_pad = cb_div_up(nq, kTileM) * kTileM, nr_pad = cb_div_up(nr, kTileN) * kTileN;
    __nv_bfloat16* A = cb_alloc_t<__nv_bfloat16>(ctx, (size_t)nq_pad * KP);
    __nv_bfloat16* B = cb_alloc_t<__nv_bfloat16>(ctx, (size_t)nr_pad * KP);
    if (!A || !B) return CB_ENOMEM;
    const int groups = KP / 8;
    (void)cand_s;
    // multiplicative scramble of the reference rows: perm_a ~ 0.618 n_pad, coprime to n_pad
    unsigned perm_a = ((unsigned)(0.6180339887 * nr_pad)) | 1u, perm_inv = 1;
    for (;; perm_a += 2) {
        long long g0 = nr_pad, g1 = perm_a % (unsigned)nr_pad, x0 = 0, x1 = 1;  // extended Euclid
        while (g1) { long long qq = g0 / g1, t = g0 - qq * g1; g0 = g1; g1 = t; t = x0 - qq * x1; x0 = x1; x1 = t; }
        if (g0 == 1) { perm_inv = (unsigned)((x0 % nr_pad + nr_pad) % nr_pad); perm_a %= (unsigned)nr_pad; break; }
    }
    knn_tc_pack_kernel<<<(unsigned)(((size_t)nq_pad * groups + 255) / 256), 256, 0, st>>>(q, nullptr, nq, nq_pad, d,
                                                                                          kTileM, 0, maxdot, 1u, A);
    CB_LAUNCH_CHECK(ctx);
    knn_tc_pack_kernel<<<(unsigned)(((size_t)nr_pad * groups + 255) / 256), 256, 0, st>>>(r, rn, nr, nr_pad, d, kTileN, 1,
                                                                                          maxdot, perm_a, B);
    CB_LAUNCH_CHECK(ctx);
    cudaError_t e;
    if (d == 16 && KC == 8) e = launch_tc<16, 8>(A, B, nq, nr, chunk, S, maxdot, perm_inv, nr_pad, cand_idx, bound, st);
    else if (d == 16) e = launch_tc<16, 16>(A, B, nq, nr, chunk, S, maxdot, perm_inv, nr_pad, cand_idx, bound, st);
    else if (KC == 8) e = launch_tc<32, 8>(A, B, nq, nr, chunk, S, maxdot, perm_inv, nr_pad, cand_idx, bound, st);
    else e = launch_tc<32, 16>(A, B, nq, nr, chunk, S, maxdot, perm_inv, nr_pad, cand_idx, bound, st);
    ctx->launches++;
    if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "knn_tc_kernel launch failed: %s", cudaGetErrorString(e));
    return CB_OK;
}
