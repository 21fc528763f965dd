// tcgen05 candidate pass of the kNN engine (K1): the distance GEMM on the 5th-generation tensor
// cores in split precision, fused with a threshold top-k of every query row.
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
// Pipeline per CTA (one 128-query tile x one reference chunk of T column tiles), 576 threads:
//   warp 16 producer: cp.async.bulk (TMA engine) of pre-packed operand tiles -> shared memory
//   warp 17 one elected thread issues tcgen05.mma (M=128, N=256, K=16 per instruction) into one
//           of two 256-column TMEM accumulators and commits to mbarriers
//   warps 0-15 epilogue: warp w drains lane quadrant w%4 (one query row per thread), column
//           quarter w/4 of every accumulator (two 32-column tcgen05.ld per stage) — four warps per
//           SM sub-partition.  Round 1 ran 8 epilogue warps and ncu showed the kernel bound by the
//           epilogue's dependent-issue latency (issue 40 %, tensor pipe 44 %, stall `wait` 1.9 per
//           issue with two warps per scheduler), not by tensor memory.
// Schedule of a chunk (T tiles): T0 = ceil(T/2) "threshold" steps, then T "collect" steps:
//   threshold  the maxima of 16- (8-) column groups of the first T0 tiles go through a branch-free
//              sorted insertion into the thread's KC best group maxima; the four threads of a row
//              merge their lists through shared memory: the KC-th best is the row's threshold tau
//              (at least KC references of the chunk reach it)
//   collect    every reference of the chunk with score >= tau is appended to the row's candidate
//              list: a packed subtract + funnel shift per column builds a 32-bit qualification
//              mask per tcgen05.ld, then one short loop over its set bits.
// Round 1 ran the threshold pass over ALL tiles (every score left tensor memory twice); a threshold
// from half of the chunk lets ~2 KC references qualify instead of ~KC (list slots: 6 KC) and cuts
// MMA + drain work from 2 T to 1.5 T tile passes.
#include "common.cuh"
#include "knn_internal.cuh"
#include "tc_common.cuh"

#include <math.h>

namespace {

constexpr int kTileM = 128;   // queries per CTA (TMEM lanes)
constexpr int kTileN = 256;   // references per MMA / accumulator stage
constexpr int kExtraK = 16;   // the norm block of the widened K axis
constexpr int kEpiWarps = 16;
constexpr int kColParts = kEpiWarps / 4;      // threads per query row
constexpr int kPartN = kTileN / kColParts;    // 64 columns per thread and stage
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads = (kEpiWarps + 2) * 32;

__host__ __device__ constexpr int kprime(int D) { return 3 * D + kExtraK; }

__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float lo, float hi) {
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void unpk2(f32x2 v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// fp32 rows -> packed bf16 operand tiles, ONE launch for both operands (round 1: a norms kernel,
// two pad kernels and two pack kernels).  One thread per padded row: rows [0, nq_pad) are queries (A
// layout hi|hi|lo|ones, 128-row tiles), rows [nq_pad, nq_pad + nr_pad) references (B layout
// hi|lo|hi|-norm/2 in three exact bf16 pieces, 256-row tiles).  Rows of width d < D are zero-filled.
// B rows are stored at the scrambled position (row * perm_a) mod nr_pad (perm_a coprime to nr_pad):
// the column groups of the threshold pass then hold pseudo-random subsets of the references,
// whatever spatial order (and feature similarity between neighbours) the cloud has.  The largest
// reference norm (fp32, the re-rank's error scale) goes to rmax_bits.
template <int D>
__global__ void __launch_bounds__(128)
knn_tc_pack_kernel(const float* __restrict__ q, int nq, int nq_pad, const float* __restrict__ r, int nr, int nr_pad,
                   int d, int vec4, int maxdot, unsigned perm_a, __nv_bfloat16* __restrict__ Apack,
                   __nv_bfloat16* __restrict__ Bpack, int32_t* __restrict__ unscramble,
                   unsigned int* __restrict__ rmax_bits) {
    constexpr int KP = kprime(D);
    const int gid = blockIdx.x * blockDim.x + threadIdx.x;
    const bool is_b = gid >= nq_pad;
    const int row = is_b ? gid - nq_pad : gid;
    float norm = 0.f;
    if (row < (is_b ? nr_pad : nq_pad)) {
        const int n = is_b ? nr : nq;
        const float* x = (is_b ? r : q) + (size_t)row * d;
        uint32_t hp[D / 2], lp[D / 2];  // bf16 pairs: hi and lo pieces of columns (2c, 2c + 1)
        auto put = [&](int c2, float a, float b) {
            __nv_bfloat16 ah, al, bh, bl;
            tc::split2(a, ah, al);
            tc::split2(b, bh, bl);
            hp[c2] = (uint32_t)__bfloat16_as_ushort(ah) | ((uint32_t)__bfloat16_as_ushort(bh) << 16);
            lp[c2] = (uint32_t)__bfloat16_as_ushort(al) | ((uint32_t)__bfloat16_as_ushort(bl) << 16);
            norm = fmaf(a, a, norm);
            norm = fmaf(b, b, norm);
        };
        if (row < n) {
            if (vec4) {  // d == D and both arrays 16-byte aligned (the caller's pointers carry no such promise)
#pragma unroll
                for (int c = 0; c < D; c += 4) {
                    const float4 v = *reinterpret_cast<const float4*>(x + c);
                    put(c / 2, v.x, v.y);
                    put(c / 2 + 1, v.z, v.w);
                }
            } else {
#pragma unroll
                for (int c = 0; c < D; c += 2) put(c / 2, c < d ? x[c] : 0.f, c + 1 < d ? x[c + 1] : 0.f);
            }
        } else {
#pragma unroll
            for (int c = 0; c < D / 2; ++c) hp[c] = lp[c] = 0u;
        }
        // the norm block: A = (1, 1, 1, 0...), B = -|r|^2/2 in three exact bf16 pieces
        uint32_t ex0 = 0u, ex1 = 0u;
        if (!is_b) {
            if (row < n) { ex0 = 0x3F803F80u; ex1 = 0x3F80u; }
        } else if (row < n) {
            if (!maxdot) {
                __nv_bfloat16 c0, c1, c2;
                tc::split3(-0.5f * norm, c0, c1, c2);
                ex0 = (uint32_t)__bfloat16_as_ushort(c0) | ((uint32_t)__bfloat16_as_ushort(c1) << 16);
                ex1 = (uint32_t)__bfloat16_as_ushort(c2);
            }
        } else {
            ex0 = (uint32_t)__bfloat16_as_ushort(__float2bfloat16_rn(-1e30f));  // padding references can never enter a list
        }
        const int TR = is_b ? kTileN : kTileM;
        const int prow = is_b ? (int)(((unsigned long long)row * perm_a) % (unsigned)nr_pad) : row;
        if (is_b) unscramble[prow] = row < n ? row : -1;  // column of the score matrix -> reference index
        const int tile = prow / TR, rr = prow % TR;
        char* base = reinterpret_cast<char*>(is_b ? Bpack : Apack) + (size_t)tile * TR * KP * 2 +
                     (size_t)(rr >> 3) * 128 + (size_t)(rr & 7) * 16;
        const size_t kstride = (size_t)(TR / 8) * 128;  // tc::tile_offset: [kc][r/8][r%8][8 x bf16]
#pragma unroll
        for (int kc = 0; kc < KP / 8; ++kc) {
            constexpr int kGroupsPerPart = D / 8;
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (kc < 3 * kGroupsPerPart) {
                const int part = kc / kGroupsPerPart, w0 = (kc % kGroupsPerPart) * 4;
                const bool want_lo = is_b ? (part == 1) : (part == 2);
                v.x = want_lo ? lp[w0] : hp[w0];
                v.y = want_lo ? lp[w0 + 1] : hp[w0 + 1];
                v.z = want_lo ? lp[w0 + 2] : hp[w0 + 2];
                v.w = want_lo ? lp[w0 + 3] : hp[w0 + 3];
            } else if (kc == 3 * kGroupsPerPart) {
                v.x = ex0;
                v.y = ex1;
            }
            *reinterpret_cast<uint4*>(base + kc * kstride) = v;
        }
        if (!(is_b && row < nr)) norm = 0.f;
    }
    // block max of the reference norms -> one atomic per warp (non-negative floats order like their bits)
    for (int o = 16; o > 0; o >>= 1) norm = fmaxf(norm, __shfl_xor_sync(0xffffffffu, norm, o));
    if ((threadIdx.x & 31) == 0 && norm > 0.f) atomicMax(rmax_bits, __float_as_uint(norm));
}

struct TcSmemCtl {
    uint64_t a_full, full[4], empty[4], tmem_full[2], tmem_empty[2];
    uint32_t tmem_base;
};

// NB = shared-memory stages of B tiles.  The tile of step it + NB - 1 is requested when the MMAs of step
// it - 1 have completed; a 56 KB bulk copy from L2 takes longer than the 7 MMAs of one step, so two
// stages (round 1) left the tensor pipe waiting for operands.
template <int D, int KC, int NB>
__global__ void __launch_bounds__(kThreads, 1)
knn_tc_kernel(const __nv_bfloat16* __restrict__ Apack, const __nv_bfloat16* __restrict__ Bpack, int nq, int nr,
              int chunk, int S, int maxdot, int nr_pad, int32_t* __restrict__ cand_idx, float* __restrict__ bound) {
    constexpr int KP = kprime(D);
    constexpr int LC = 6 * KC;            // list slots per (query, chunk): P(overflow) ~ 3e-7
    constexpr int GROUP = 8;              // columns per group maximum (16 let 4 % of the rows of clustered
                                          // features overflow their list in two-tile chunks, 8: < 0.1 %)
    constexpr uint32_t A_BYTES = kTileM * KP * 2, B_BYTES = kTileN * KP * 2;
    constexpr uint32_t LBO_A = (kTileM / 8) * 128, LBO_B = (kTileN / 8) * 128;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;
    unsigned char* sB = smem + A_BYTES;
    TcSmemCtl* ctl = reinterpret_cast<TcSmemCtl*>(smem + A_BYTES + NB * B_BYTES);
    float* s_top = reinterpret_cast<float*>(smem + A_BYTES + NB * B_BYTES + 128);  // [kColParts][kTileM][KC]
    int* s_cnt = reinterpret_cast<int*>(s_top + kColParts * kTileM * KC);        // [kTileM]
    const int warp = tc::uniform_warp_idx(), lane = threadIdx.x & 31;
    const int c = blockIdx.y;
    const int t0 = (c * chunk) / kTileN;
    const int t1 = min((nr + kTileN - 1) / kTileN, ((c + 1) * chunk) / kTileN);
    const int ntiles = t1 - t0;
    const int nthr = (ntiles + 1) >> 1;     // threshold steps: the first half of the chunk's tiles
    const int nsteps = nthr + ntiles;       // then every tile once more (collect)

    if (threadIdx.x == 0) {
        tc::mbar_init(&ctl->a_full, 1);
        for (int i = 0; i < NB; ++i) {
            tc::mbar_init(&ctl->full[i], 1);
            tc::mbar_init(&ctl->empty[i], 1);
        }
        for (int i = 0; i < 2; ++i) {
            tc::mbar_init(&ctl->tmem_full[i], 1);
            tc::mbar_init(&ctl->tmem_empty[i], kEpiWarps);
        }
        tc::fence_barrier_init();
    }
    CB_DEV_CHECK(ntiles >= 1 && t1 * kTileN <= nr_pad);
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
                const int st = it % NB, use = it / NB;
                const int tile = t0 + (it >= nthr ? it - nthr : it);
                tc::mbar_wait(&ctl->empty[st], (use & 1) ^ 1);
                tc::mbar_arrive_expect_tx(&ctl->full[st], B_BYTES);
                tc::bulk_g2s(sB + st * B_BYTES, reinterpret_cast<const char*>(Bpack) + (size_t)tile * B_BYTES, B_BYTES,
                             &ctl->full[st]);
            }
        }
    } else if (warp == kEpiWarps + 1) {
        // ===== MMA issuer: the whole warp walks the loop (uniform control flow), one elected lane issues =====
        constexpr uint32_t idesc = tc::idesc_bf16(kTileM, kTileN);
        const uint32_t a_addr = tc::smem_u32(sA);
        const bool leader = tc::elect_one();
        tc::mbar_wait(&ctl->a_full, 0);
        for (int it = 0; it < nsteps; ++it) {
            const int st = it % NB, use = it / NB;   // shared-memory stage
            const int ts = it & 1, tuse = it >> 1;   // tensor-memory stage
            tc::mbar_wait(&ctl->full[st], use & 1);
            tc::mbar_wait(&ctl->tmem_empty[ts], (tuse & 1) ^ 1);
            tc::fence_after_sync();
            if (leader) {
                const uint32_t b_addr = tc::smem_u32(sB + st * B_BYTES);
#pragma unroll
                for (int ks = 0; ks < KP / 16; ++ks) {
                    const uint64_t ad = tc::smem_desc(a_addr + ks * 2 * LBO_A, LBO_A, 128);
                    const uint64_t bd = tc::smem_desc(b_addr + ks * 2 * LBO_B, LBO_B, 128);
                    tc::mma_bf16(tbase + ts * kTileN, ad, bd, idesc, ks > 0);
                }
                tc::mma_commit(&ctl->empty[st]);      // smem stage reusable once these MMAs finish
                tc::mma_commit(&ctl->tmem_full[ts]);  // accumulator ready for the epilogue
            }
            __syncwarp();
        }
    } else {
        // ===== epilogue: 512 threads = 128 query rows (TMEM lanes) x 4 column quarters =====
        const int part = warp >> 2;
        const int row = (warp & 3) * 32 + lane;
        const int qi = blockIdx.x * kTileM + row;
        const size_t o = ((size_t)min(qi, nq - 1) * S + c) * LC;
        const bool live = qi < nq;
        const uint32_t lane_addr = tbase + ((uint32_t)((warp & 3) * 32) << 16) + part * kPartN;
        // the KT = KC / 2 best group maxima of this thread's columns, descending.  The row's threshold is the
        // KC-th best of the four threads' lists: a thread that holds more than KT of the row's KC best
        // group maxima (11 % of the rows) lowers it slightly, which only lets a few more references qualify
        constexpr int KT = KC / 2;
        float top[KC];
#pragma unroll
        for (int t = 0; t < KC; ++t) top[t] = -INFINITY;
        float tau = -INFINITY;
        f32x2 ntau2 = pk2(0.f, 0.f);
        for (int it = 0; it < nsteps; ++it) {
            const int st = it & 1, use = it >> 1;
            const bool collect = it >= nthr;
            if (it == nthr) {
                // the row's threshold: KC-th best of the union of its four threads' group maxima
#pragma unroll
                for (int t = 0; t < KT; ++t) s_top[(part * kTileM + row) * KC + t] = top[t];
                named_bar_sync(1, kEpiThreads);
#pragma unroll
                for (int p = 1; p < kColParts; ++p) {
                    const float* other = s_top + (((part + p) % kColParts) * kTileM + row) * KC;
#pragma unroll
                    for (int u = 0; u < KT; ++u) {
                        float x = other[u];
#pragma unroll
                        for (int t = 0; t < KC; ++t) {
                            const float hi = fmaxf(top[t], x);
                            x = fminf(top[t], x);
                            top[t] = hi;
                        }
                    }
                }
                tau = fmaxf(top[KC - 1], -1e29f);  // padding (-1e30) never qualifies
                ntau2 = pk2(-tau, -tau);
            }
            tc::mbar_wait(&ctl->tmem_full[st], use & 1);
            tc::fence_after_sync();
            const int col0 = (t0 + (collect ? it - nthr : it)) * kTileN + part * kPartN;
            // both 32-column loads of this thread's quarter, then the stage goes straight back to the
            // MMA warp: the arithmetic below overlaps the other warps' loads and the next MMAs
            float v[kPartN];
            tc::tmem_ld32(lane_addr + st * kTileN, v);
            tc::tmem_ld32(lane_addr + st * kTileN + 32, v + 32);
            tc::tmem_ld_wait64(v);
            tc::fence_before_sync();
            __syncwarp();
            if (lane == 0) tc::mbar_arrive(&ctl->tmem_empty[st]);
            if (!collect) {
#pragma unroll
                for (int g = 0; g < kPartN / GROUP; ++g) {
                    // group maximum by three-input maxima (FMNMX3), then a branch-free insertion into the
                    // descending list (x sinks to its place)
                    const float* m = v + g * GROUP;
                    static_assert(GROUP == 8, "the maximum tree below is written for groups of 8");
                    const float a0 = fmaxf(fmaxf(m[0], m[1]), m[2]);
                    const float a1 = fmaxf(fmaxf(m[3], m[4]), m[5]);
                    float x = fmaxf(fmaxf(a0, a1), fmaxf(m[6], m[7]));
#pragma unroll
                    for (int t = 0; t < KT; ++t) {
                        const float hi = fmaxf(top[t], x);
                        x = fminf(top[t], x);
                        top[t] = hi;
                    }
                }
            } else {
#pragma unroll
                for (int h = 0; h < kPartN / 32; ++h) {
                    // which of the 32 columns qualify: packed v - tau (two columns per instruction), the
                    // sign bit (set: below the threshold) shifted into the word by one funnel shift;
                    // column e lands at bit 31 - e, hence the bit reversal
                    // (four independent 8-column chains: one 32-long chain of dependent shifts left the
                    // ALU pipe waiting on its own latency)
                    unsigned b4[4] = {0u, 0u, 0u, 0u};
#pragma unroll
                    for (int e = 0; e < 32; e += 2) {
                        float d0, d1;
                        unpk2(add2(pk2(v[h * 32 + e], v[h * 32 + e + 1]), ntau2), d0, d1);
                        b4[e >> 3] = __funnelshift_l(__float_as_uint(d0), b4[e >> 3], 1);
                        b4[e >> 3] = __funnelshift_l(__float_as_uint(d1), b4[e >> 3], 1);
                    }
                    const unsigned below = (b4[0] << 24) | (b4[1] << 16) | (b4[2] << 8) | b4[3];
                    unsigned mask = __brev(~below);
                    // padding rows (all-equal scores) and rows whose list already overflowed (they go
                    // to the exact fallback whatever else qualifies) must not spin in the loop below
                    if (!live || (mask && *(volatile int*)&s_cnt[row] > LC)) mask = 0;
                    // append their reference indices (the re-rank recomputes every candidate exactly,
                    // the approximate scores are not needed)
                    while (mask) {
                        const int e = __ffs(mask) - 1;
                        mask &= mask - 1;
                        const int pos = atomicAdd(&s_cnt[row], 1);
                        CB_DEV_CHECK(pos >= 0 && col0 + h * 32 + e < nr_pad && e >= 0 && e < 32);
                        if (pos < LC) cand_idx[o + pos] = col0 + h * 32 + e;  // scrambled column: the re-rank maps it back
                    }
                }
            }
        }
        named_bar_sync(1, kEpiThreads);  // every thread of every row has appended
        if (live && part == 0) {
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
constexpr int b_stages() {
    // as many 256-row B stages as fit beside the A tile and the threshold lists (227 KB per CTA)
    return (227 * 1024 - kTileM * kprime(D) * 2 - 128 - kColParts * kTileM * KC * 4 - kTileM * 4 - 64) / (kTileN * kprime(D) * 2) >= 4
               ? 4
               : (227 * 1024 - kTileM * kprime(D) * 2 - 128 - kColParts * kTileM * KC * 4 - kTileM * 4 - 64) / (kTileN * kprime(D) * 2);
}

template <int D, int KC>
cudaError_t launch_tc(const __nv_bfloat16* A, const __nv_bfloat16* B, int nq, int nr, int chunk, int S, int maxdot,
                      int nr_pad, int32_t* ci, float* bound, cudaStream_t st) {
    constexpr int KP = kprime(D);
    constexpr int NB = b_stages<D, KC>();
    static_assert(NB >= 2, "at least two B stages");
    const size_t smem = (size_t)kTileM * KP * 2 + NB * (size_t)kTileN * KP * 2 + 128 + kColParts * kTileM * KC * 4 + kTileM * 4 + 64;
    cudaError_t e = cudaFuncSetAttribute(knn_tc_kernel<D, KC, NB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    dim3 grid((nq + kTileM - 1) / kTileM, S);
    knn_tc_kernel<D, KC, NB><<<grid, kThreads, smem, st>>>(A, B, nq, nr, chunk, S, maxdot, nr_pad, ci, bound);
    return cudaGetLastError();
}

}  // namespace

double knn_cand_tc_gamma(int) { return 1.5e-5; }

extern "C" const char* cb_knn_variant(const cb_ctx* ctx) {
    return (ctx && ctx->knn_variant == 1) ? "simt" : "tcgen05";
}

int knn_cand_tc_list_slots(int KC) { return 6 * KC; }

// tile passes (MMA + drain of one 128 x 256 accumulator) one CTA of a `tiles`-tile chunk runs
int knn_cand_tc_passes(int tiles) { return tiles + (tiles + 1) / 2; }

// q [nq, d], r [nr, d] raw rows (d <= dp, dp in {16, 32}); writes *rmax_bits (zeroed by the caller).
// cand_idx receives COLUMNS of the scrambled score matrix; *unscramble_out [nr_pad] maps them to
// reference indices (the re-rank applies it).
int knn_cand_tc_launch(cb_ctx* ctx, const float* q, int nq, const float* r, int nr, int d, int dp, int KC, int LC,
                       int chunk, int S, int maxdot, int32_t* cand_idx, float* bound, unsigned int* rmax_bits,
                       const int32_t** unscramble_out, cudaStream_t st) {
    if ((dp != 16 && dp != 32) || d > dp || (KC != 8 && KC != 16) || LC != 6 * KC || chunk % kTileN != 0)
        CB_FAIL(ctx, CB_EUNSUPPORTED, "knn_cand_tc_launch: unsupported shape (d=%d KC=%d LC=%d chunk=%d)", d, KC, LC, chunk);
    const int KP = kprime(dp);
    const int nq_pad = cb_div_up(nq, kTileM) * kTileM, nr_pad = cb_div_up(nr, kTileN) * kTileN;
    __nv_bfloat16* A = cb_alloc_t<__nv_bfloat16>(ctx, (size_t)nq_pad * KP);
    __nv_bfloat16* B = cb_alloc_t<__nv_bfloat16>(ctx, (size_t)nr_pad * KP);
    int32_t* unscramble = cb_alloc_t<int32_t>(ctx, nr_pad);
    if (!A || !B || !unscramble) return CB_ENOMEM;
    *unscramble_out = unscramble;
    // multiplicative scramble of the reference rows: perm_a ~ 0.618 n_pad, coprime to n_pad
    unsigned perm_a = ((unsigned)(0.6180339887 * nr_pad)) | 1u;
    for (;; perm_a += 2) {
        unsigned g0 = (unsigned)nr_pad, g1 = perm_a % (unsigned)nr_pad;
        while (g1) { const unsigned t = g0 % g1; g0 = g1; g1 = t; }
        if (g0 == 1) { perm_a %= (unsigned)nr_pad; break; }
    }
    const unsigned pack_blocks = (unsigned)cb_div_up(nq_pad + nr_pad, 128);
    const int vec4 = d == dp && (((size_t)q | (size_t)r) & 15) == 0;
    if (dp == 16)
        knn_tc_pack_kernel<16><<<pack_blocks, 128, 0, st>>>(q, nq, nq_pad, r, nr, nr_pad, d, vec4, maxdot, perm_a, A, B, unscramble, rmax_bits);
    else
        knn_tc_pack_kernel<32><<<pack_blocks, 128, 0, st>>>(q, nq, nq_pad, r, nr, nr_pad, d, vec4, maxdot, perm_a, A, B, unscramble, rmax_bits);
    CB_LAUNCH_CHECK(ctx);
    cudaError_t e;
    if (dp == 16 && KC == 8) e = launch_tc<16, 8>(A, B, nq, nr, chunk, S, maxdot, nr_pad, cand_idx, bound, st);
    else if (dp == 16) e = launch_tc<16, 16>(A, B, nq, nr, chunk, S, maxdot, nr_pad, cand_idx, bound, st);
    else if (KC == 8) e = launch_tc<32, 8>(A, B, nq, nr, chunk, S, maxdot, nr_pad, cand_idx, bound, st);
    else e = launch_tc<32, 16>(A, B, nq, nr, chunk, S, maxdot, nr_pad, cand_idx, bound, st);
    ctx->launches++;
    if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "knn_tc_kernel launch failed: %s", cudaGetErrorString(e));
    return CB_OK;
}
