// jt_resident_kernel: the junction-tree pipeline of SMALL trees with the whole tile state in shared memory (resident.h).
//   One CTA per SM, persistent over tiles of 64 test cases (lane = 2 cases, 128-bit rows as everywhere, plan.h). The tile's
//   live message rows (up messages, down ratios, unnormalised marginals) never leave shared memory; HBM sees the
//   evidence coming in and the posteriors going out and nothing else. The constants of a phase (static per-warp work lists,
//   per-entry records, pattern-folded potentials) arrive as ONE block per phase through cp.async.bulk + mbarrier (TMA) into a
//   double buffer, requested one phase ahead, so the inner loop is: one broadcast LDS.128 (record), per-case LDS.64
//   (potential under the case's evidence pattern), one LDS.128 per gathered message row, multiply-add.
//   Same SumOps and the same arithmetic order per target as jt_flat_kernel (flat_kernel.cu): collect (reference
//   src/JunctionTree.cpp:829-941, :1056-1148), distribute (:700-816, :1150-1238; division done algebraically), marginals
//   (:1392-1454), Normalize (PotentialTableBase.cpp:433-445) once, on the final marginals (resident.h), ArgMax (Inference.cpp:92-102).
#include <cstdint>
#include <type_traits>

#include <cuda_runtime.h>

#include "plan.h"
#include "resident.h"

namespace fbn {

#ifndef FBN_RES_CLOCK_BUILD
#define FBN_RES_CLOCK_BUILD 0   // developer builds (tools/build_variant.sh ... -DFBN_RES_CLOCK_BUILD=1): FBN_RES_CLOCK=1 records cycles per block
#endif
constexpr bool kResClock = FBN_RES_CLOCK_BUILD != 0;
#if FBN_RES_CLOCK_BUILD == 1
// developer build: cycles of thread 0 of CTA 0 per segment of a phase (summed over all phases and tiles)
__shared__ long long s_clk_t, s_clk_sum[16], s_clk_cnt[16];
#define RES_STAMP(k)                                                     \
    do {                                                                 \
        if (blockIdx.x == 0 && threadIdx.x == 0) {                       \
            const long long n__ = clock64();                             \
            s_clk_sum[k] += n__ - s_clk_t; s_clk_cnt[k] += 1; s_clk_t = n__; \
        }                                                                \
    } while (0)
#else
#define RES_STAMP(k) do { } while (0)
#endif

namespace {

struct V2 {
    double x, y;
};

__device__ __forceinline__ uint32_t lds32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint32_t lds16(uint32_t a) {
    uint16_t v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds128u(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ double lds64(uint32_t a) {
    double v;
    asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ V2 lds128(uint32_t a) {
    V2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t a, V2 v) { asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(a), "d"(v.x), "d"(v.y) : "memory"); }

// start the bulk copy of one constant block into a buffer; completion is counted in bytes on the buffer's mbarrier
__device__ __forceinline__ void block_issue(uint32_t dst_s, const unsigned char *src, uint32_t bytes, uint32_t bar_s) {
    // (no proxy fence: the buffer was only READ through the generic proxy, and those reads are ordered before this point by the
    // CTA barrier; measured: the fence cost ~270 cycles per block on the issuing warp's critical path)
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s), "l"(src), "r"(bytes),
                 "r"(bar_s)
                 : "memory");
}
__device__ __forceinline__ void block_wait(uint32_t bar_s, uint32_t parity) {
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar_s), "r"(parity)
                     : "memory");
    }
}

// evidence patterns of one tile (ops.cuh: load_patterns), by ONE warp: a lane walks the pairs of its two cases (they are
// adjacent in the CSR arrays) four at a time; pat_s must be zero. dims_s: one byte per variable.
__device__ __forceinline__ void res_load_patterns(const ResDev &R, uint32_t pat_s, uint32_t dims_s, int64_t tile, int64_t n_cases,
                                                  const int64_t *__restrict__ ev_off, int64_t ev_base, const int32_t *__restrict__ ev_vars,
                                                  const int32_t *__restrict__ ev_vals, int lane) {
    const int64_t c0 = tile * kTileCases + lane * 2;
    if (c0 >= n_cases) return;
    const int64_t k0 = ev_off[c0] - ev_base, kmid = ev_off[c0 + 1] - ev_base;
    const int64_t k1 = c0 + 1 < n_cases ? ev_off[c0 + 2] - ev_base : kmid;
    constexpr int kEvBatch = 8;   // (variable, value) pairs requested together: the item's time is global-memory round trips
    for (int64_t k = k0; k < k1; k += kEvBatch) {
        int vv[kEvBatch], ww[kEvBatch];
#pragma unroll
        for (int u = 0; u < kEvBatch; ++u) {
            const int64_t kk = min(k + u, k1 - 1);
            vv[u] = __ldg(ev_vars + kk);
            ww[u] = __ldg(ev_vals + kk);
        }
#pragma unroll
        for (int u = 0; u < kEvBatch; ++u) {
            if (k + u >= k1) break;
            const int v = vv[u], val = ww[u];
            if (v < 0 || v >= R.n_vars) continue;   // src/JunctionTree.cpp:326-331
            uint32_t dim;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(dim) : "r"(dims_s + (uint32_t)v));
            const uint32_t np = (val < 0 || val >= (int)dim) ? dim + 1 : 1 + (uint32_t)val;
            const uint32_t a = pat_s + (uint32_t)v * kTileCases + lane * 2 + (k + u >= kmid ? 1 : 0);
            uint32_t old;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(old) : "r"(a));
            const uint32_t nw = (old == 0 || old == np) ? np : dim + 1;
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(a), "r"(nw) : "memory");
        }
    }
}

// lambda rows (evidence indicators 1.0 / 0.0 per state) of the variables some view gathers, from the pattern bytes `patl` of a
// tile; the variables are dealt to the CTA's warps
__device__ __forceinline__ void res_write_lams(uint32_t lams_s, int n_lams, uint32_t Sb, uint32_t patl, int warp, int T) {
    for (int k = warp; k < n_lams; k += T) {
        const uint4 l = lds128u(lams_s + 16 * k);   // variable, dim, state offset
        const uint32_t pp = lds16(patl + l.x * kTileCases), q0 = pp & 255u, q1 = pp >> 8;
        for (uint32_t d = 0; d < l.y; ++d) {
            V2 x;
            x.x = (q0 == 0 || q0 == 1 + d) ? 1.0 : 0.0;
            x.y = (q1 == 0 || q1 == 1 + d) ? 1.0 : 0.0;
            sts128(Sb + l.z + d * kResRowBytes, x);
        }
    }
}

// records per batch (all loads of a batch are issued before the first use): 16 warps x 128 registers take 4 / 2, 32 warps x 64
// registers 2 / 1
template <int NG, int MAXT>
struct ResBatch {
    static constexpr int value = MAXT <= 16 ? (NG <= 2 ? 4 : 2) : (MAXT <= 24 ? 2 : (NG <= 2 ? 2 : 1));
};

// this lane's pattern slots of an op's potentials (one per case): block base + sum over the folded variables of pattern x stride
__device__ __noinline__ uint2 res_fold(uint32_t blk, uint32_t desc, uint32_t patl) {
    const uint32_t nfold = lds32(desc);
    uint32_t pot0 = blk, pot1 = blk;
#pragma unroll 1
    for (uint32_t k = 0; k < nfold; ++k) {
        const uint32_t v = lds32(desc + 16 + 8 * k), st = lds32(desc + 20 + 8 * k);
        const uint32_t pp = lds16(patl + v * kTileCases);
        pot0 += (pp & 255u) * st;
        pot1 += (pp >> 8) * st;
    }
    return make_uint2(pot0, pot1);
}

// one SUM item: the `n_targets` x `rows` records of a piece as ONE stream, target-major, in batches of BATCH records whose loads
// are all issued before the first use; a batch may span targets (their sums overlap in time: a warp that finished one target
// at a time spent most of its cycles waiting for a chain of dependent shared-memory loads), the end of a target is a
// warp-uniform counter. One compact body per NG: the code of all item kinds must stay in the instruction cache, the warps of
// a CTA run different kinds at the same time (a first version with one fully unrolled body per (gathers, remainder, checks)
// was 380 KB of SASS and lost 13 % of the warp time to instruction fetches).
template <int NG, int BATCH>
__device__ __noinline__ void res_sum(uint32_t blk, const uint4 a, const uint4 b4, const uint4 c4, const uint4 d4, uint32_t Sb, uint32_t patl) {
    constexpr uint32_t kRec = NG > 3 ? 32u : 16u;
    const uint32_t rows = a.w;
    uint32_t n = a.z * rows, dst = Sb + b4.x;
    RES_STAMP(3);
    // set-up from the item alone (resident.h): the pattern bytes of the first two folded variables and of the (at most
    // kResMaxChecks) checked target variables are requested together; unused slots name a valid variable and are masked
    const uint32_t nfold = c4.x & 255u, ncm = c4.x >> 8;
    const uint32_t f0 = lds16(patl + c4.y * kTileCases), f1 = lds16(patl + c4.w * kTileCases);
    const uint32_t k0 = lds16(patl + d4.y * kTileCases), k1 = lds16(patl + d4.z * kTileCases), k2 = lds16(patl + d4.w * kTileCases);
    // this lane's pattern slots of the op's potentials (one per case): block base + sum over the folded variables of pattern x stride
    uint32_t pot0 = blk + (f0 & 255u) * c4.z + (f1 & 255u) * d4.x, pot1 = blk + (f0 >> 8) * c4.z + (f1 >> 8) * d4.x;
    if (nfold > 2) {
        const uint2 pots = res_fold(blk, blk + b4.y, patl);
        pot0 = pots.x; pot1 = pots.y;
    }
    RES_STAMP(4);
    // (measured and rejected: one potential load for both cases when the view has no folded variable, 4 of 5 ALARM entries:
    // the warp-uniform branch costs more than the load, 256 against 268 M cases/s)
    uint32_t rec = blk + a.y, r = 0;
    // checked target variables: this lane's two pattern bytes of each, read once per item; per finished target the byte the
    // target wants (a word per target and variable); cw = 0: no checks
    uint32_t cw = ncm ? blk + b4.w : 0u;
    const uint32_t c0 = ncm > 0 ? k0 : 0u, c1 = ncm > 1 ? k1 : 0u, c2 = ncm > 2 ? k2 : 0u;
    RES_STAMP(5);
    auto check = [&](uint32_t c, uint32_t want, V2 &x) {
        const uint32_t q0 = c & 255u, q1 = c >> 8;
        if (q0 != 0 && q0 != want) x.x = 0.0;
        if (q1 != 0 && q1 != want) x.y = 0.0;
    };
    V2 acc = {0.0, 0.0};
    // a batch = N records: their data loads are all issued before the first use; the RECORDS of the next full batch are requested
    // before this batch's arithmetic (they do not depend on it), which takes one shared-memory round trip out of every batch
    auto batch = [&](auto nb, uint4 (&A)[decltype(nb)::value], uint4 (&B)[decltype(nb)::value], uint32_t next_rec, auto prefetch) {
        constexpr int N = decltype(nb)::value;
        constexpr bool kPrefetch = decltype(prefetch)::value;
        double p0[N], p1[N];
        V2 m[N][NG > 0 ? NG : 1];
#pragma unroll
        for (int b = 0; b < N; ++b) {
            p0[b] = lds64(pot0 + A[b].x);
            p1[b] = lds64(pot1 + A[b].x);
            if (NG > 0) m[b][0] = lds128(Sb + A[b].y);
            if (NG > 1) m[b][1] = lds128(Sb + A[b].z);
            if (NG > 2) m[b][2] = lds128(Sb + A[b].w);
            if (NG > 3) m[b][3] = lds128(Sb + B[b].x);
        }
        if (kPrefetch) {
#pragma unroll
            for (int b = 0; b < N; ++b) {
                A[b] = lds128u(next_rec + b * kRec);
                if (NG > 3) B[b] = lds128u(next_rec + b * kRec + 16);
            }
        }
#pragma unroll
        for (int b = 0; b < N; ++b) {
            double v0 = p0[b], v1 = p1[b];
#pragma unroll
            for (int g = 0; g < NG; ++g) { v0 *= m[b][g].x; v1 *= m[b][g].y; }
            acc.x += v0; acc.y += v1;
            if (++r == rows) {   // end of target j
                if (cw) {
                    // all three slots, no branches: an unused slot holds pattern 0 (nothing observed), the words read past the
                    // target's ncm are the next target's (or the bytes after the table: still this CTA's shared memory)
                    const uint32_t w0 = lds32(cw), w1 = lds32(cw + 4), w2 = lds32(cw + 8);
                    check(c0, w0, acc);
                    check(c1, w1, acc);
                    check(c2, w2, acc);
                    cw += 4 * ncm;
                }
                sts128(dst, acc);
                dst += kResRowBytes;
                r = 0;
                acc.x = 0.0; acc.y = 0.0;
            }
        }
    };
    auto load_records = [&](auto nb, uint4 (&A)[decltype(nb)::value], uint4 (&B)[decltype(nb)::value]) {
#pragma unroll
        for (int b = 0; b < decltype(nb)::value; ++b) {
            A[b] = lds128u(rec + b * kRec);
            if (NG > 3) B[b] = lds128u(rec + b * kRec + 16);
        }
    };
    if (n >= BATCH) {
        uint4 A[BATCH], B[BATCH];
        load_records(std::integral_constant<int, BATCH>(), A, B);
#pragma unroll 1
        for (; n >= BATCH; n -= BATCH) {
            const uint32_t cur = rec;
            rec += BATCH * kRec;
            // the last full batch re-reads its own records (valid addresses, values unused) instead of branching
            batch(std::integral_constant<int, BATCH>(), A, B, n >= 2 * BATCH ? rec : cur, std::true_type());
        }
    }
    if constexpr (BATCH > 1) {
#pragma unroll 1
        for (; n > 0; --n) {
            uint4 A[1], B[1];
            load_records(std::integral_constant<int, 1>(), A, B);
            rec += kRec;
            batch(std::integral_constant<int, 1>(), A, B, rec, std::false_type());
        }
    }
    RES_STAMP(6);
}

// ops with five to seven gathers per entry (cliques with more than four neighbours): plain loop
__device__ __noinline__ void res_sum_generic(uint32_t blk, const uint4 a, const uint4 b4, uint32_t Sb, uint32_t patl, uint32_t ng, bool cm) {
    const uint32_t n_targets = a.z, rows = a.w, desc = blk + b4.y;
    uint32_t dst = Sb + b4.x;
    const uint2 pots = res_fold(blk, desc, patl);
    uint32_t rec = blk + a.y, j = b4.z;
    for (uint32_t t = 0; t < n_targets; ++t) {
        V2 acc = {0.0, 0.0};
        for (uint32_t r = 0; r < rows; ++r, rec += 32u) {
            const uint32_t po = lds32(rec);
            double v0 = lds64(pots.x + po), v1 = lds64(pots.y + po);
            for (uint32_t g = 0; g < ng; ++g) {
                const V2 m = lds128(Sb + lds32(rec + 4 + 4 * g));
                v0 *= m.x; v1 *= m.y;
            }
            acc.x += v0; acc.y += v1;
        }
        if (cm) {
            const uint4 h = lds128u(desc);   // nfold, ncm, want_off
            for (uint32_t k = 0; k < h.y; ++k) {
                const uint32_t c = lds16(patl + lds32(desc + 16 + 8 * h.x + 4 * k) * kTileCases), want = lds32(blk + h.z + (j * h.y + k) * 4);
                if ((c & 255u) != 0 && (c & 255u) != want) acc.x = 0.0;
                if ((c >> 8) != 0 && (c >> 8) != want) acc.y = 0.0;
            }
            ++j;
        }
        sts128(dst, acc);
        dst += kResRowBytes;
    }
}

// OUT item: posteriors + label of one variable (GetProbabilitiesOneNode src/JunctionTree.cpp:1392-1454, Inference::ArgMax
// src/Inference.cpp:92-102): normalise the unnormalised marginal rows, zeros for an observed variable
__device__ __noinline__ void res_out(const ResDev &R, const uint4 a, const uint4 b4, uint32_t Sb, uint32_t patl, int64_t tile, int64_t n_cases,
                                        int lane, int32_t *__restrict__ labels, double *__restrict__ post) {
    const uint32_t v = a.y, dim = a.z, off = a.w, mg = Sb + b4.x;   // ResItem: variable, dim, posterior column, marginal rows
    RES_STAMP(9);
    const uint32_t pp = lds16(patl + v * kTileCases);
    const bool obs0 = (pp & 255u) != 0, obs1 = (pp >> 8) != 0;
    const int64_t c0 = tile * kTileCases + lane * 2;
    const bool ok0 = c0 < n_cases, ok1 = c0 + 1 < n_cases;
    double s0 = 0.0, s1 = 0.0;
    for (uint32_t d = 0; d < dim; ++d) {
        const V2 x = lds128(mg + d * kResRowBytes);
        s0 += x.x; s1 += x.y;
    }
    // one division per case, then products (a division per state costs ~35 instructions; x * (1 / s) is within one ulp of x / s).
    // 0 / 0 (evidence with probability zero) must stay NaN as in the reference: 1 / 0 = inf, 0 * inf = NaN
    RES_STAMP(10);
    const double r0 = 1.0 / s0, r1 = 1.0 / s1;
    double best0 = 0.0, best1 = 0.0;
    int arg0 = -1, arg1 = -1;
    double *q0 = post ? post + c0 * R.sum_dims + off : nullptr;
    double *q1 = q0 ? q0 + R.sum_dims : nullptr;
    for (uint32_t d = 0; d < dim; ++d) {
        const V2 x = lds128(mg + d * kResRowBytes);
        const double pr0 = obs0 ? 0.0 : x.x * r0, pr1 = obs1 ? 0.0 : x.y * r1;
        if (q0 && ok0) q0[d] = pr0;
        if (q1 && ok1) q1[d] = pr1;
        if (pr0 > best0) { best0 = pr0; arg0 = (int)d; }
        if (pr1 > best1) { best1 = pr1; arg1 = (int)d; }
    }
    RES_STAMP(11);
    if ((int)v == R.query_var && labels) {
        if (ok0) labels[c0] = arg0;
        if (ok1) labels[c0 + 1] = arg1;
    }
}

}  // namespace

template <int MAXT>
__global__ void __launch_bounds__(MAXT * 32, 1) jt_resident_kernel(const ResDev R, int64_t n_cases, const int64_t *__restrict__ ev_off, int64_t ev_base,
                                                                   const int32_t *__restrict__ ev_vars, const int32_t *__restrict__ ev_vals,
                                                                   int32_t *__restrict__ labels, double *__restrict__ post, int64_t n_tiles) {
    extern __shared__ __align__(128) unsigned char sm[];
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sm);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, T = blockDim.x >> 5;
    const ResSmem L = R.L;
    for (int v = threadIdx.x; v < R.n_vars; v += blockDim.x) sm[L.dims + v] = (unsigned char)R.vars[v].dim;
    for (int b = threadIdx.x; b <= R.n_blocks; b += blockDim.x) reinterpret_cast<uint32_t *>(sm + L.boff)[b] = R.block_off[b];
    for (int i = threadIdx.x; i < R.n_lams * 4; i += blockDim.x) reinterpret_cast<uint32_t *>(sm + L.lams)[i] = reinterpret_cast<const uint32_t *>(R.lams)[i];
    const uint32_t pat_words = (L.pat1 - L.pat0) / 4;
    for (uint32_t i = threadIdx.x; i < 2 * pat_words; i += blockDim.x) reinterpret_cast<uint32_t *>(sm + L.pat0)[i] = 0u;
    const uint32_t bars = sbase + L.bars;
#if FBN_RES_CLOCK_BUILD == 1
    if (threadIdx.x == 0) {
        for (int k = 0; k < 16; ++k) { s_clk_sum[k] = 0; s_clk_cnt[k] = 0; }
        s_clk_t = clock64();
    }
#endif
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if ((int64_t)blockIdx.x >= n_tiles) return;
    const uint32_t dims_s = sbase + L.dims;
    const uint32_t *boff = reinterpret_cast<const uint32_t *>(sm + L.boff);
    if (warp == 0) res_load_patterns(R, sbase + L.pat0, dims_s, blockIdx.x, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane);
    if (threadIdx.x == 0) block_issue(sbase + L.buf0, R.blob, boff[1] - boff[0], bars);
    const uint32_t Sb = sbase + L.state + lane * 16;   // this lane's two cases of state row 0
    if (R.n_lams) {
        __syncthreads();
        res_write_lams(sbase + L.lams, R.n_lams, Sb, sbase + L.pat0 + lane * 2, warp, T);
    }
    __syncthreads();

    uint32_t g = 0;   // blocks consumed so far: buffer g & 1, mbarrier parity (g >> 1) & 1
    int it = 0;
    for (int64_t tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
        const uint32_t pat_cur = sbase + ((it & 1) ? L.pat1 : L.pat0), pat_nxt = sbase + ((it & 1) ? L.pat0 : L.pat1);
        const uint32_t patl = pat_cur + lane * 2;
        const int64_t next_tile = tile + gridDim.x;
        for (int b = 0; b < R.n_blocks; ++b, ++g) {
            if (kResClock && R.clk && blockIdx.x == 0 && threadIdx.x == 0 && g < 2048) R.clk[4 * g] = clock64();
            if (threadIdx.x == blockDim.x - 32) {   // (the last warp: longest-processing-time-first gives it the lightest load) request the next block: its buffer was last read before the CTA barrier that ended block g - 1
                const int nb = b + 1 < R.n_blocks ? b + 1 : (next_tile < n_tiles ? 0 : -1);
                if (nb >= 0) block_issue(sbase + (((g + 1) & 1) ? L.buf1 : L.buf0), R.blob + boff[nb], boff[nb + 1] - boff[nb], bars + 8 * ((g + 1) & 1));
            }
            RES_STAMP(0);
            block_wait(bars + 8 * (g & 1), (g >> 1) & 1);
            RES_STAMP(1);
            if (kResClock && R.clk && blockIdx.x == 0 && threadIdx.x == 0 && g < 2048) R.clk[4 * g + 1] = clock64();
#if FBN_RES_CLOCK_BUILD == 2   // -DFBN_RES_CLOCK_BUILD=2: cycles every warp of CTA 0 needs for its items, per block (summed over tiles)
            const long long w_t0 = clock64();
#endif
            const uint32_t blk = sbase + ((g & 1) ? L.buf1 : L.buf0);
            const uint4 hdr = lds128u(blk);   // n_items, -, items_off, bytes
            // the last block sums nothing: the next tile's lambda rows can be written (its patterns came in an earlier block)
            if (R.n_lams && b == R.n_blocks - 1 && next_tile < n_tiles) res_write_lams(sbase + L.lams, R.n_lams, Sb, pat_nxt + lane * 2, warp, T);
            const uint32_t i0 = lds32(blk + 16 + 4 * warp), i1 = lds32(blk + 20 + 4 * warp);
            // (measured and rejected: the warps taking the items from a queue, an atomic counter in shared memory, instead of the
            // planner's static lists: 228-238 M ALARM cases/s against 258-268 M, the counter serialises the warps at every block start)
            for (uint32_t i = i0; i < i1; ++i) {
                const uint32_t ia = blk + hdr.z + i * (uint32_t)sizeof(ResItem);
                RES_STAMP(2);
                const uint4 a = lds128u(ia), b4 = lds128u(ia + 16), c4 = lds128u(ia + 32), d4 = lds128u(ia + 48);
                const uint32_t kind = a.x & 255u, ng = (a.x >> 8) & 255u;
                if (kind == RES_SUM || kind == RES_SUMC) {
                    // bit 16 + min(ng, 5) of the item's first word names the body: a chain of warp-uniform bit tests, most
                    // frequent first (a switch on ng becomes a jump table: a constant load and an indirect branch per item)
                    if (a.x & (1u << 17)) res_sum<1, ResBatch<1, MAXT>::value>(blk, a, b4, c4, d4, Sb, patl);
                    else if (a.x & (1u << 18)) res_sum<2, ResBatch<2, MAXT>::value>(blk, a, b4, c4, d4, Sb, patl);
                    else if (a.x & (1u << 16)) res_sum<0, ResBatch<0, MAXT>::value>(blk, a, b4, c4, d4, Sb, patl);
                    else if (a.x & (1u << 19)) res_sum<3, ResBatch<3, MAXT>::value>(blk, a, b4, c4, d4, Sb, patl);
                    else if (a.x & (1u << 20)) res_sum<4, ResBatch<4, MAXT>::value>(blk, a, b4, c4, d4, Sb, patl);
                    else res_sum_generic(blk, a, b4, Sb, patl, ng, kind == RES_SUMC);
                } else if (kind == RES_OUT) {
                    res_out(R, a, b4, Sb, patl, tile, n_cases, lane, labels, post);
                } else if (next_tile < n_tiles) {   // RES_EVID: the next tile's patterns into the other pattern buffer
                    for (uint32_t w = lane; w < pat_words; w += 32) asm volatile("st.shared.u32 [%0], %1;" ::"r"(pat_nxt + 4 * w), "r"(0u) : "memory");
                    __syncwarp();
                    res_load_patterns(R, pat_nxt, dims_s, next_tile, n_cases, ev_off, ev_base, ev_vars, ev_vals, lane);
                }
            }
            if (kResClock && R.clk && blockIdx.x == 0 && threadIdx.x == 0 && g < 2048) R.clk[4 * g + 2] = clock64();
#if FBN_RES_CLOCK_BUILD == 2
            if (R.clk && blockIdx.x == 0 && lane == 0 && b < 32) atomicAdd(reinterpret_cast<unsigned long long *>(R.clk) + 8300 + b * 32 + warp, (unsigned long long)(clock64() - w_t0));
            if (R.clk && blockIdx.x == 0 && threadIdx.x == 0 && b == 0) atomicAdd(reinterpret_cast<unsigned long long *>(R.clk) + 8299, 1ull);
#endif
            RES_STAMP(7);
            __syncthreads();
            RES_STAMP(8);
        }
    }
#if FBN_RES_CLOCK_BUILD == 1
    if (R.clk && blockIdx.x == 0 && threadIdx.x == 0)
        for (int k = 0; k < 16; ++k) { R.clk[8100 + k] = s_clk_sum[k]; R.clk[8120 + k] = s_clk_cnt[k]; }
#endif
}

// ---- launch helpers (called from capi.cu) ------------------------------------------------------------------------
static int res_maxt(int warps) { return warps <= 8 ? 8 : (warps <= 16 ? 16 : (warps <= 24 ? 24 : 32)); }

template <int MAXT>
static cudaError_t res_prep(int smem_optin) {
    cudaFuncAttributes fa;
    cudaError_t e = cudaFuncGetAttributes(&fa, jt_resident_kernel<MAXT>);
    if (e != cudaSuccess) return e;
    // per function and per context: the device's opt-in maximum minus the kernel's static shared memory (none in product builds)
    return cudaFuncSetAttribute(jt_resident_kernel<MAXT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_optin - (int)fa.sharedSizeBytes);
}

cudaError_t resident_prepare(int warps, int smem_optin) {
    switch (res_maxt(warps)) {
        case 8: return res_prep<8>(smem_optin);
        case 16: return res_prep<16>(smem_optin);
        case 24: return res_prep<24>(smem_optin);
        default: return res_prep<32>(smem_optin);
    }
}

cudaError_t resident_launch(const ResDev &R, int warps, int grid, int64_t n_cases, const int64_t *ev_off, int64_t ev_base, const int32_t *ev_vars,
                            const int32_t *ev_vals, int32_t *labels, double *post, cudaStream_t st) {
    const int64_t n_tiles = (n_cases + kTileCases - 1) / kTileCases;
    if (n_tiles == 0) return cudaSuccess;
    if (grid > n_tiles) grid = (int)n_tiles;
    const size_t smem = R.L.total;
    switch (res_maxt(warps)) {
        case 8: jt_resident_kernel<8><<<grid, warps * 32, smem, st>>>(R, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles); break;
        case 16: jt_resident_kernel<16><<<grid, warps * 32, smem, st>>>(R, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles); break;
        case 24: jt_resident_kernel<24><<<grid, warps * 32, smem, st>>>(R, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles); break;
        default: jt_resident_kernel<32><<<grid, warps * 32, smem, st>>>(R, n_cases, ev_off, ev_base, ev_vars, ev_vals, labels, post, n_tiles); break;
    }
    return cudaGetLastError();
}

}  // namespace fbn
