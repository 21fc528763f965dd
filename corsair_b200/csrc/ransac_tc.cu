// K3 scoring on the 5th-generation tensor cores.
//
// The inlier test of hypothesis h on correspondence c, u = |R_h s_c + t_h - q_c|^2 - thr^2 < 0, is
// one inner product between a per-hypothesis weight row and a per-correspondence feature row
// (ransac_internal.cuh), so a whole wave is the GEMM  U[h, c] = A[h, :] . B[c, :]  in split bf16
// (three bf16 pieces per factor, six products per bilinear term: fp32-grade accuracy), fp32 accumulate.
// The epilogue never stores U: each thread owns one hypothesis (one TMEM lane) and folds its row
// into an inlier count and a sum as it streams out of tensor memory.
//
// Exactness: |U - u_float64| <= band[h] (a rigorous per-hypothesis rounding bound).  Evaluations
// with |U| <= band are queued in shared memory and re-decided by the oracle's unfused float64
// expression on the raw float32 points, so the COUNT is exactly the float64 count; everything
// outside the band (> 99.9 %) is decided by the tensor-core value alone.
//
// CTA = 256 hypotheses (two A tiles) x one split of the correspondences, 320 threads:
//   warp 8  producer: cp.async.bulk of the pre-packed A tiles and of a 3-stage ring of B tiles
//   warp 9  one elected thread issues tcgen05.mma (M=128, N=128, K=16; 6 per tile and A tile) into
//           a double-buffered TMEM accumulator (2 stages x 2 A tiles x 128 columns = 512)
//   warps 0-7 epilogue: warp w reads lane quadrant w%4 of A tile w/4 in 64-column tcgen05.ld.
//           Measured (profiles/r01_tmem_ld_microbench.txt): with the arithmetic compiled out the kernel
//           still needs 70 % of its time — the drain runs at ~88 B/clk/SM beside the MMAs' accumulator
//           traffic (267+ B/clk/SM on an idle SM) — so the epilogue is kept to 1.5 instructions per value
#include "ransac_internal.cuh"

#include <stdlib.h>

#include <math.h>

namespace {

constexpr int kHT = kRtHT;        // A tiles (of 128 hypotheses) per CTA
// "phased" epilogue (needs kHT == 1): one CTA of 128 hypotheses per SM, FOUR 128-column accumulator
// stages; epilogue warp group g = warp / 4 drains the whole tiles t = g (mod 4), so the MMA warp can run
// up to four tiles ahead and no warp waits for the slowest of the sixteen
#ifndef CB_RT_PHASED
#define CB_RT_PHASED 0
#endif
constexpr bool kPhased = CB_RT_PHASED != 0;
static_assert(!kPhased || kHT == 1, "the phased epilogue drains one A tile per CTA");
constexpr int kTS = kPhased ? 4 : 2;  // tensor-memory accumulator stages
constexpr int kStages = (kHT == 2 || kPhased) ? 6 : 3;  // shared-memory ring of 24 KB B tiles
// epilogue warps: 4 lane quadrants x kHT A tiles x (EW / (4 kHT)) column parts.  16 (four per SM sub-partition)
// is the default: with 8 the kernel was bound by the epilogue's dependent-issue latency (ncu, round 2:
// issue active 48 %, tensor pipe 49 %, warps active 15.6 %, stall `wait` 1.0 per issue)
constexpr int kMaxEpiWarps = kPhased ? 16 : 8 * kHT;
constexpr int kListCap = (kHT == 2 || kPhased) ? 2048 : 1024;  // queued in-band evaluations per CTA (overflow is settled inline)
constexpr int kListDrain = kListCap / 2;          // drain the queue at the next tile boundary once it holds this many
constexpr uint32_t kATileBytes = kRtTileM * kRtKP * 2;  // 24 KB
constexpr uint32_t kBTileBytes = kRtTileN * kRtKP * 2;  // 24 KB

// ------------------------------------------------------------------------------ B pack
// One thread per (correspondence row, 8-column group).  Rows >= m are padding: n = +big so the
// evaluation is never an inlier and never inside a band.
__global__ void __launch_bounds__(256)
ransac_tc_pack_corr_kernel(const float4* __restrict__ src4, const float4* __restrict__ tgt4, int m, int m_pad,
                           RansacState* __restrict__ state, __nv_bfloat16* __restrict__ Bpack) {
    const size_t gid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    constexpr int kGroups = kRtKP / 8;
    const int row = (int)(gid / kGroups), kc = (int)(gid % kGroups);
    float ns = 0.f, nq = 0.f;
    if (row < m_pad) {
        __nv_bfloat16 v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) v[e] = __float2bfloat16_rn(0.f);
        if (row < m) {
            const double inv = 1.0 / (double)m;
            const float4 s4 = src4[row], q4 = tgt4[row];
            const double s[3] = {(double)s4.x - state->csum[0] * inv, (double)s4.y - state->csum[1] * inv,
                                 (double)s4.z - state->csum[2] * inv};
            const double q[3] = {(double)q4.x - state->csum[3] * inv, (double)q4.y - state->csum[4] * inv,
                                 (double)q4.z - state->csum[5] * inv};
            const double n = s[0] * s[0] + s[1] * s[1] + s[2] * s[2] + q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
            if (kc == 0) {
                // maxima of the centred norms, rounded up, for the fit kernel's band
                ns = (float)sqrt(s[0] * s[0] + s[1] * s[1] + s[2] * s[2]) * 1.000001f;
                nq = (float)sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2]) * 1.000001f;
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                const int col = kc * 8 + e;
                if (col < 90) {
                    const int k = col % kRtTerms, part = col / kRtTerms;  // part: pa pb pa pc pa pb
                    const double phi = k < 9 ? q[k / 3] * s[k % 3] : (k < 12 ? s[k - 9] : q[k - 12]);
                    __nv_bfloat16 pa, pb, pc;
                    rt_split3(phi, pa, pb, pc);
                    v[e] = (part == 1 || part == 5) ? pb : (part == 3 ? pc : pa);
                } else if (col < 93) {
                    __nv_bfloat16 a, b, c;
                    rt_split3(n, a, b, c);
                    v[e] = col == 90 ? a : (col == 91 ? b : c);
                } else {
                    v[e] = __float2bfloat16_rn(1.f);
                }
            }
        } else if (kc == 11) {
            v[2] = __float2bfloat16_rn(1e30f);  // column 90
        }
        const int tile = row / kRtTileN, r = row % kRtTileN;
        char* dst = reinterpret_cast<char*>(Bpack) + (size_t)tile * kBTileBytes + tc::tile_offset(kRtTileN, r, kc);
        *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
    }
    for (int o = 16; o > 0; o >>= 1) {
        ns = fmaxf(ns, __shfl_xor_sync(0xffffffffu, ns, o));
        nq = fmaxf(nq, __shfl_xor_sync(0xffffffffu, nq, o));
    }
    if ((threadIdx.x & 31) == 0) {
        if (ns > 0.f) atomicMax(&state->smaxc_bits, __float_as_uint(ns));
        if (nq > 0.f) atomicMax(&state->tmaxc_bits, __float_as_uint(nq));
    }
}

// ------------------------------------------------------------------------------ scoring
struct RtCtl {
    uint64_t a_full, full[kStages], empty[kStages], tmem_full[4], tmem_empty[4];
    uint32_t tmem_base;
    int list_n;
};

// queue entry (int2): x = CTA-local hypothesis row (8 bits) << 24 | correspondence index (24 bits),
// y = the tensor-core value of u (float bits)

__device__ __forceinline__ float fma_sat(float a, float b, float c) {
    float d;
    asm("fma.rn.sat.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
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
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) {
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 add2(f32x2 a, f32x2 b) {
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}

// what the streaming pass added for a value u: its sign bit to the count (one integer instruction,
// LEA.HI in SASS; a saturating-FMA gate measured at a quarter of the FMA pipe's rate), min(u, 0) to
// the sum
__device__ __forceinline__ unsigned rt_sign(float u) { return __float_as_uint(u) >> 31; }
// ... for half of the columns; the other half is counted on the FMA pipe (fma.sat gate + packed add)
// so that both pipes share the epilogue: column c takes the float path when bit 1 of c is clear
constexpr float kNegBig = -1.2676506e30f;  // -2^100: fma.sat(u, kNegBig, 0) = 1 for u < 0, else 0
__device__ __forceinline__ float rt_gate(float u) { return fma_sat(u, kNegBig, 0.f); }
#ifndef CB_RT_FLOAT_SHARE
#define CB_RT_FLOAT_SHARE 2  // quarters of the columns counted on the FMA pipe (0..4)
#endif
__device__ __forceinline__ bool rt_float_path(int col) { return ((col >> 1) & 3) < CB_RT_FLOAT_SHARE; }
__device__ __forceinline__ double rt_streamed_count(int col, float u) {
    return rt_float_path(col) ? (double)rt_gate(u) : (double)rt_sign(u);
}

// exact decision for one evaluation; returns the correction (exact - streamed) of count and sum
__device__ __forceinline__ void rt_exact_delta(const double* __restrict__ Tj, const float4* __restrict__ src4,
                                               const float4* __restrict__ tgt4, int col, float u, double thr2d,
                                               double& dc, double& ds) {
    const double dd = ransac_exact_d2_inl(Tj, src4[col], tgt4[col]);
    const bool in = dd < thr2d;
    dc = (in ? 1.0 : 0.0) - rt_streamed_count(col, u);
    ds = (in ? dd - thr2d : 0.0) - (double)fminf(u, 0.f);
}

// the CTA's queue is full: settle one in-band evaluation on the spot
__device__ __noinline__ void rt_overflow(int col, float u, const double* __restrict__ Tj, const float4* __restrict__ src4,
                                         const float4* __restrict__ tgt4, double thr2d, double& dcnt, double& dsum) {
    double dc, ds;
    rt_exact_delta(Tj, src4, tgt4, col, u, thr2d, dc, ds);
    dcnt += dc;
    dsum += ds;
}

// all epilogue threads: re-decide the queued evaluations exactly, corrections into s_dcnt/s_dsum
template <int kEpiThreads>
__device__ __forceinline__ void rt_drain(RtCtl* ctl, const int2* list, double* s_dcnt, double* s_dsum, int jcta, int w,
                                         const double* __restrict__ Tw, const float4* __restrict__ src4,
                                         const float4* __restrict__ tgt4, double thr2d) {
    const int nlist = min(*(volatile int*)&ctl->list_n, kListCap);
    for (int i = threadIdx.x; i < nlist; i += kEpiThreads) {
        const int2 en = list[i];
        const int row = (en.x >> 24) & 0xff, col = en.x & 0xffffff;
        CB_DEV_CHECK(col < (1 << 24) && row < kHT * kRtTileM);
        double dc, ds;
        rt_exact_delta(Tw + (size_t)min(jcta + row, w - 1) * 12, src4, tgt4, col, __int_as_float(en.y), thr2d, dc, ds);
        atomicAdd(&s_dcnt[row], dc);
        atomicAdd(&s_dsum[row], ds);
    }
}

template <bool DUMP, int LDW, bool WSUM, int kEpiWarps>
__global__ void __launch_bounds__(kEpiWarps * 32 + 64, (kHT == 2 || kPhased) ? 1 : 2)
ransac_tc_score_kernel(const __nv_bfloat16* __restrict__ Apack, const __nv_bfloat16* __restrict__ Bpack,
                       const float* __restrict__ band_g, const double* __restrict__ Tw,
                       const float4* __restrict__ src4, const float4* __restrict__ tgt4, int m, int ntiles_total,
                       int tiles_per_split, int h0, int w, double thr2d, const RansacState* __restrict__ state,
                       int32_t* __restrict__ pc, double* __restrict__ ps, float* __restrict__ dump, int dump_ld) {
    constexpr int kColSplit = kPhased ? 1 : kEpiWarps / (4 * kHT);
    constexpr int kColsPerThread = kRtTileN / kColSplit;
    constexpr int kChunks = kColsPerThread / 32;  // 32-column tcgen05.ld per thread and tile (even)
    constexpr int kEpiThreads = kEpiWarps * 32;
    // hypotheses at or beyond the (device-side, shrinking) budget are never looked at by the fold
    if (h0 + (int)blockIdx.x * (kHT * kRtTileM) >= state->budget) return;
    extern __shared__ __align__(1024) unsigned char smem[];
    unsigned char* sA = smem;                               // kHT A tiles
    unsigned char* sB = smem + kHT * kATileBytes;           // kStages B tiles
    int2* list = reinterpret_cast<int2*>(sB + kStages * kBTileBytes);
    double* s_dsum = reinterpret_cast<double*>(list + kListCap);
    double* s_dcnt = s_dsum + kHT * kRtTileM;
    RtCtl* ctl = reinterpret_cast<RtCtl*>(s_dcnt + kHT * kRtTileM);
    const int warp = tc::uniform_warp_idx(), lane = threadIdx.x & 31;
    const int t0 = blockIdx.y * tiles_per_split;
    const int t1 = min(ntiles_total, t0 + tiles_per_split);
    const int ntiles = t1 - t0;
    const int jcta = blockIdx.x * (kHT * kRtTileM);  // first hypothesis (wave-relative) of this CTA

    if (threadIdx.x == 0) {
        tc::mbar_init(&ctl->a_full, 1);
        for (int i = 0; i < kStages; ++i) {
            tc::mbar_init(&ctl->full[i], 1);
            tc::mbar_init(&ctl->empty[i], 1);
        }
        for (int i = 0; i < kTS; ++i) {
            tc::mbar_init(&ctl->tmem_full[i], 1);
            tc::mbar_init(&ctl->tmem_empty[i], kPhased ? 4 : kEpiWarps);
        }
        ctl->list_n = 0;
        tc::fence_barrier_init();
    }
    if (threadIdx.x < kHT * kRtTileM) { s_dsum[threadIdx.x] = 0.0; s_dcnt[threadIdx.x] = 0.0; }
    if (warp == kEpiWarps) tc::tmem_alloc(&ctl->tmem_base, kTS * kHT * kRtTileN);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tbase = ctl->tmem_base;

    if (warp == kEpiWarps) {
        // ===== producer =====
        if (lane == 0) {
            tc::mbar_arrive_expect_tx(&ctl->a_full, kHT * kATileBytes);
            tc::bulk_g2s(sA, reinterpret_cast<const char*>(Apack) + (size_t)blockIdx.x * kHT * kATileBytes,
                         kHT * kATileBytes, &ctl->a_full);
            for (int it = 0; it < ntiles; ++it) {
                const int st = it % kStages, use = it / kStages;
                tc::mbar_wait(&ctl->empty[st], (use & 1) ^ 1);
                tc::mbar_arrive_expect_tx(&ctl->full[st], kBTileBytes);
                tc::bulk_g2s(sB + st * kBTileBytes, reinterpret_cast<const char*>(Bpack) + (size_t)(t0 + it) * kBTileBytes,
                             kBTileBytes, &ctl->full[st]);
            }
        }
    } else if (warp == kEpiWarps + 1) {
        // ===== MMA issuer: the whole warp walks the loop (uniform control flow), one elected lane issues =====
        constexpr uint32_t idesc = tc::idesc_bf16(kRtTileM, kRtTileN);
        constexpr uint32_t LBO = (kRtTileM / 8) * 128;  // both operands: 128-row tiles
        const uint32_t a_addr = tc::smem_u32(sA);
        const bool leader = tc::elect_one();
        tc::mbar_wait(&ctl->a_full, 0);
        for (int it = 0; it < ntiles; ++it) {
            const int st = it % kStages, use = it / kStages;
            const int ts = it % kTS, tuse = it / kTS;
            tc::mbar_wait(&ctl->full[st], use & 1);
            tc::mbar_wait(&ctl->tmem_empty[ts], (tuse & 1) ^ 1);
            tc::fence_after_sync();
            if (leader) {
                const uint32_t b_addr = tc::smem_u32(sB + st * kBTileBytes);
#pragma unroll
                for (int a = 0; a < kHT; ++a) {
#pragma unroll
                    for (int ks = 0; ks < kRtKP / 16; ++ks) {
                        const uint64_t ad = tc::smem_desc(a_addr + a * kATileBytes + ks * 2 * LBO, LBO, 128);
                        const uint64_t bd = tc::smem_desc(b_addr + ks * 2 * LBO, LBO, 128);
                        tc::mma_bf16(tbase + (ts * kHT + a) * kRtTileN, ad, bd, idesc, ks > 0);
                    }
                }
                tc::mma_commit(&ctl->empty[st]);
                tc::mma_commit(&ctl->tmem_full[ts]);
            }
            __syncwarp();
        }
    } else {
        // ===== epilogue: one thread per hypothesis row and column part =====
        const int a = kPhased ? 0 : (warp >> 2) % kHT;     // A tile of this warp
        const int half = kPhased ? 0 : warp / (4 * kHT);   // column part of every tile this warp drains
        const int group = warp >> 2;                       // phased: this warp drains the tiles t = group (mod 4)
        const bool first = kPhased ? group == 0 : half == 0;  // the thread of a row that adds the queue's corrections
        const int rloc = a * kRtTileM + (warp & 3) * 32 + lane;  // row inside the CTA
        const int j = jcta + rloc;                         // wave-relative hypothesis
        const bool valid = j < w;
        const float band = valid ? band_g[j] : -1.f;
        const double* Tj = Tw + (size_t)min(j, w - 1) * 12;
        const uint32_t lane_addr = tbase + ((uint32_t)((warp & 3) * 32) << 16) + a * kRtTileN + half * kColsPerThread;
        double sum_d = 0.0, dcnt = 0.0, dsum = 0.0;
        unsigned cnt_i = 0;
        f32x2 s2, c2;
        // one 32-column chunk: gate, count, sum, and the in-band test
        auto process = [&](const float (&v)[32], int colc) {
            float bm0 = INFINITY, bm1 = INFINITY;
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
                if (!rt_float_path(e)) cnt_i += rt_sign(v[e]) + rt_sign(v[e + 1]);
                else c2 = add2(c2, pk2(rt_gate(v[e]), rt_gate(v[e + 1])));
                // the streamed sum only feeds cb_ransac's optional `sums` output: the fold recomputes the
                // sum of every hypothesis tied at the top count exactly
                if (WSUM) s2 = add2(s2, pk2(fminf(v[e], 0.f), fminf(v[e + 1], 0.f)));
                if (e & 2) bm1 = fminf(bm1, fminf(fabsf(v[e]), fabsf(v[e + 1])));
                else bm0 = fminf(bm0, fminf(fabsf(v[e]), fabsf(v[e + 1])));
            }
            if (DUMP) {
                if (valid)
                    for (int e = 0; e < 32; ++e)
                        if (colc + e < m) dump[(size_t)j * dump_ld + colc + e] = v[e];
            }
            // rare (about one lane-chunk in 500): some evaluation sits inside the rounding band.  Only
            // the lanes concerned run this: a 32-bit mask of their in-band columns, ONE reservation of
            // queue slots, then the entries.
            if (fminf(bm0, bm1) <= band) {
                // the opaque copy keeps the 32 comparisons inside the branch (the compiler otherwise
                // hoists them into the streaming loop: 64 more instructions per 32 columns)
                float band_in;
                asm volatile("mov.f32 %0, %1;" : "=f"(band_in) : "f"(band));
                unsigned mask = 0;
#pragma unroll
                for (int e = 0; e < 32; ++e) mask |= (fabsf(v[e]) <= band_in ? 1u : 0u) << e;
                int pos = atomicAdd(&ctl->list_n, __popc(mask));
                while (mask) {
                    const int e = __ffs(mask) - 1;
                    mask &= mask - 1;
                    float u = 0.f;  // v[e] by static selects: a runtime index would put v in local memory
#pragma unroll
                    for (int t = 0; t < 32; ++t) u = t == e ? v[t] : u;
                    CB_DEV_CHECK(pos >= 0 && rloc >= 0 && rloc < kHT * kRtTileM && colc + e < m && colc + e < (1 << 24));
                    if (pos < kListCap) list[pos] = make_int2((rloc << 24) | (colc + e), __float_as_int(u));
                    else rt_overflow(colc + e, u, Tj, src4, tgt4, thr2d, dcnt, dsum);
                    ++pos;
                }
            }
        };
        constexpr bool kPipelined = LDW == 64 && kChunks == 4;
        constexpr bool kBoth = LDW == 64 && kChunks == 2;  // 16 warps: the thread's 64 columns in two loads, one wait
        float va[kPipelined || kBoth || kPhased ? 64 : 1], vb[kPipelined ? 64 : 1];
        if constexpr (kPhased) {
            const int n_iter = (ntiles + 3) >> 2;
            for (int k = 0; k < n_iter; ++k) {
                const int it = 4 * k + group;
                if (it < ntiles) {
                    tc::mbar_wait(&ctl->tmem_full[group], k & 1);
                    tc::fence_after_sync();
                    const int col0 = (t0 + it) * kRtTileN;
                    const uint32_t taddr = lane_addr + group * kRtTileN;
                    s2 = pk2(0.f, 0.f);
                    c2 = pk2(0.f, 0.f);
                    tc::tmem_ld32(taddr, va);
                    tc::tmem_ld32(taddr + 32, va + 32);
                    tc::tmem_ld_wait64(va);
                    process(*reinterpret_cast<const float(*)[32]>(va), col0);
                    process(*reinterpret_cast<const float(*)[32]>(va + 32), col0 + 32);
                    tc::tmem_ld32(taddr + 64, va);
                    tc::tmem_ld32(taddr + 96, va + 32);
                    tc::tmem_ld_wait64(va);
                    tc::fence_before_sync();
                    __syncwarp();
                    if (lane == 0) tc::mbar_arrive(&ctl->tmem_empty[group]);
                    process(*reinterpret_cast<const float(*)[32]>(va), col0 + 64);
                    process(*reinterpret_cast<const float(*)[32]>(va + 32), col0 + 96);
                    float x, y;
                    unpk2(c2, x, y);  // at most 64 per tile: exact in fp32
                    cnt_i += (unsigned)(x + y);
                    if (WSUM) {
                        unpk2(s2, x, y);
                        sum_d += (double)x + (double)y;
                    }
                }
                // every 16 tiles: if the queue is half full (a CTA-uniform decision) drain it
                if ((k & 3) != 3) continue;
                named_bar_sync(1, kEpiThreads);
                if (*(volatile int*)&ctl->list_n >= kListDrain) {
                    rt_drain<kEpiThreads>(ctl, list, s_dcnt, s_dsum, jcta, w, Tw, src4, tgt4, thr2d);
                    named_bar_sync(1, kEpiThreads);
                    if (threadIdx.x == 0) ctl->list_n = 0;
                    named_bar_sync(1, kEpiThreads);
                }
            }
        } else
        for (int it = 0; it < ntiles; ++it) {
            const int ts = it & 1, tuse = it >> 1;
            tc::mbar_wait(&ctl->tmem_full[ts], tuse & 1);
            tc::fence_after_sync();
            const int col0 = (t0 + it) * kRtTileN + half * kColsPerThread;
            const uint32_t taddr = lane_addr + ts * (kHT * kRtTileN);
            s2 = pk2(0.f, 0.f);
            c2 = pk2(0.f, 0.f);
            // drain this thread's columns, 64 (or 32) at a time (measured: neither double-buffering the
            // loads nor 16 epilogue warps changes the time)
            if constexpr (kPipelined) {
                // two 64-column loads per tile: the second is in flight while the first 64 columns are
                // processed, and the accumulator stage goes back to the MMA warp as soon as it has
                // landed (562 -> 524 us per registration; also prefetching the next tile's first load
                // during the second half measured slower, 553 us)
                tc::tmem_ld64(taddr, va);
                tc::tmem_ld_wait64(va);
                tc::tmem_ld64(taddr + 64, vb);
                process(*reinterpret_cast<const float(*)[32]>(va), col0);
                process(*reinterpret_cast<const float(*)[32]>(va + 32), col0 + 32);
                tc::tmem_ld_wait64(vb);
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&ctl->tmem_empty[ts]);
                process(*reinterpret_cast<const float(*)[32]>(vb), col0 + 64);
                process(*reinterpret_cast<const float(*)[32]>(vb + 32), col0 + 96);
            } else if constexpr (kBoth) {
                // the stage goes back to the MMA warp as soon as both loads have landed; the arithmetic
                // overlaps the other three warps of this sub-partition and the next tile's MMAs
                tc::tmem_ld32(taddr, va);
                tc::tmem_ld32(taddr + 32, va + 32);
                tc::tmem_ld_wait64(va);
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&ctl->tmem_empty[ts]);
                process(*reinterpret_cast<const float(*)[32]>(va), col0);
                process(*reinterpret_cast<const float(*)[32]>(va + 32), col0 + 32);
            } else if constexpr (LDW == 64) {
#pragma unroll 1
                for (int k = 0; k < kChunks; k += 2) {
                    float v[64];
                    tc::tmem_ld64(taddr + k * 32, v);
                    tc::tmem_ld_wait();
                    process(*reinterpret_cast<const float(*)[32]>(v), col0 + k * 32);
                    process(*reinterpret_cast<const float(*)[32]>(v + 32), col0 + k * 32 + 32);
                }
            } else {
#pragma unroll 1
                for (int k = 0; k < kChunks; ++k) {
                    float v[32];
                    tc::tmem_ld32(taddr + k * 32, v);
                    tc::tmem_ld_wait();
                    process(v, col0 + k * 32);
                }
            }
            // accumulator stage drained: hand it back to the MMA warp
            if constexpr (!kPipelined && !kBoth) {
                tc::fence_before_sync();
                __syncwarp();
                if (lane == 0) tc::mbar_arrive(&ctl->tmem_empty[ts]);
            }
            {
                float x, y;
                unpk2(c2, x, y);  // at most 64 per tile: exact in fp32
                cnt_i += (unsigned)(x + y);
                if (WSUM) {
                    unpk2(s2, x, y);
                    sum_d += (double)x + (double)y;
                }
            }
            // every 16 tiles: if the queue is half full (a CTA-uniform decision) drain it
            if ((it & 15) != 15) continue;
            named_bar_sync(1, kEpiThreads);
            if (*(volatile int*)&ctl->list_n >= kListDrain) {
                rt_drain<kEpiThreads>(ctl, list, s_dcnt, s_dsum, jcta, w, Tw, src4, tgt4, thr2d);
                named_bar_sync(1, kEpiThreads);
                if (threadIdx.x == 0) ctl->list_n = 0;
                named_bar_sync(1, kEpiThreads);
            }
        }
        // settle the queued in-band evaluations of the whole CTA
        named_bar_sync(1, kEpiThreads);
        rt_drain<kEpiThreads>(ctl, list, s_dcnt, s_dsum, jcta, w, Tw, src4, tgt4, thr2d);
        named_bar_sync(1, kEpiThreads);
        if (valid) {
            // this thread's share of the row; the queue's corrections ride with the first half
            const double c = (double)cnt_i + dcnt + (first ? s_dcnt[rloc] : 0.0);
            const long long ci = llrint(c);
            // totals over column halves and correspondence splits (zeroed by the fit kernel)
            atomicAdd(&pc[j], (int)ci);
            // sum of squared inlier distances = sum (d^2 - thr^2) + count * thr^2
            if (WSUM) atomicAdd(&ps[j], sum_d + dsum + (first ? s_dsum[rloc] : 0.0) + (double)ci * thr2d);
        }
    }
    tc::fence_before_sync();
    __syncthreads();
    if (warp == kEpiWarps) tc::tmem_dealloc(tbase, kTS * kHT * kRtTileN);
}

constexpr size_t kScoreSmem = kHT * kATileBytes + kStages * kBTileBytes + kListCap * sizeof(int2) +
                              2 * kHT * kRtTileM * sizeof(double) + sizeof(RtCtl) + 64;

// Splits of the correspondence axis: one CTA per SM is resident (TMEM: 512 columns each), so the
// kernel runs ceil(groups * P / SMs) rounds of CTAs that each stream ~1/P of the tiles; pick the P
// with the shortest makespan (a small per-CTA constant stands for the prologue).
void rt_split_plan(const cb_ctx* ctx, int m, int w, int* P, int* tiles_per_split) {
    const int ntiles = cb_div_up(m, kRtTileN);
    const int groups = cb_div_up(w, kHT * kRtTileM);
    int best_p = 1;
    double best_cost = 1e300;
    for (int p = 1; p <= min(ntiles, 32); ++p) {
        const int tps = cb_div_up(ntiles, p);
        const int pp = cb_div_up(ntiles, tps);
        const int rounds = cb_div_up(groups * pp, ctx->sm_count * ((kHT == 2 || kPhased) ? 1 : 2));  // resident CTAs per SM
        const double cost = rounds * (tps + 2.0);
        if (cost < best_cost - 1e-9) { best_cost = cost; best_p = p; }
    }
    const int tps = cb_div_up(ntiles, best_p);
    *tiles_per_split = tps;
    *P = cb_div_up(ntiles, tps);
}

}  // namespace

int ransac_tc_pack_corr_launch(cb_ctx* ctx, const float4* src4, const float4* tgt4, int m, RansacState* state,
                               __nv_bfloat16* Bpack, cudaStream_t st) {
    const int m_pad = cb_div_up(m, kRtTileN) * kRtTileN;
    const size_t threads = (size_t)m_pad * (kRtKP / 8);
    ransac_tc_pack_corr_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(src4, tgt4, m, m_pad, state, Bpack);
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}

int ransac_tc_score_launch(cb_ctx* ctx, const __nv_bfloat16* Apack, const __nv_bfloat16* Bpack, const float* band,
                           const double* Tw, const float4* src4, const float4* tgt4, int m, int h0, int w,
                           double thr2, const RansacState* state, int32_t* pc, double* ps, bool want_sum, cudaStream_t st) {
    if (m >= (1 << 24)) CB_FAIL(ctx, CB_EUNSUPPORTED, "tensor-core RANSAC scoring holds at most 2^24 correspondences");
    int P, tps;
    rt_split_plan(ctx, m, w, &P, &tps);
    float* dump = ctx->ransac_dump;
    const int dump_ld = ctx->ransac_dump_ld;
    cudaError_t e;
    // the sum is streamed only when the caller asked for cb_ransac's per-hypothesis `sums`
    static const int epi_warps = [] {
        const char* e = getenv("CB_RT_EPI_WARPS");  // tuning knob: 8 or 16 epilogue warps (one CTA per SM only)
        return kHT == 2 && e && atoi(e) == 8 ? 8 : kMaxEpiWarps;
    }();
    using kern_t = void (*)(const __nv_bfloat16*, const __nv_bfloat16*, const float*, const double*, const float4*,
                            const float4*, int, int, int, int, int, double, const RansacState*, int32_t*, double*, float*, int);
    kern_t kern;
    if (epi_warps == 8)
        kern = dump ? ransac_tc_score_kernel<true, 32, true, 8>
                    : (want_sum ? ransac_tc_score_kernel<false, 64, true, 8> : ransac_tc_score_kernel<false, 64, false, 8>);
    else
        kern = dump ? ransac_tc_score_kernel<true, 32, true, kMaxEpiWarps>
                    : (want_sum ? ransac_tc_score_kernel<false, 64, true, kMaxEpiWarps> : ransac_tc_score_kernel<false, 64, false, kMaxEpiWarps>);
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kScoreSmem);
    if (e == cudaSuccess)
        kern<<<dim3(cb_div_up(w, kHT * kRtTileM), P), epi_warps * 32 + 64, kScoreSmem, st>>>(
            Apack, Bpack, band, Tw, src4, tgt4, m, cb_div_up(m, kRtTileN), tps, h0, w, thr2, state, pc, ps,
            dump ? dump + (size_t)h0 * dump_ld : nullptr, dump_ld);
    if (e != cudaSuccess) CB_FAIL(ctx, CB_ECUDA, "ransac_tc_score_kernel: %s", cudaGetErrorString(e));
    CB_LAUNCH_CHECK(ctx);
    return CB_OK;
}
