// Second-order IIR (biquad) over whole clips as a SINGLE-PASS blocked scan with a deterministic look-back.
// Replaces torchaudio.functional.lowpass_biquad / highpass_biquad -> biquad -> lfilter(clamp=True) behind
// features/processing.py:141-150 (low_pass / mid_pass / high_pass):
//     v[n] = b0 x[n] + b1 x[n-1] + b2 x[n-2]          (coefficients already divided by a0)
//     y[n] = v[n] - (a2 y[n-2] + a1 y[n-1]),  y[-1] = y[-2] = 0,  output clamped to [-1, 1] at the very end.
//
// One CTA owns a SPAN of 256 chunks x 32 samples, staged once in shared memory (the input is read from HBM once):
//   sweep A: every chunk (= thread) runs the recurrence from a ZERO state and keeps its end state z_j;
//   scan:    the CTA scans its z_j in fp64 (the state follows the affine recurrence S_j = A^32 S_{j-1} + z_j, A = companion
//            matrix; shuffles inside a warp, one barrier across warps) and publishes the zero-state end state Z of the span;
//   look-back: the true state entering the span is  P^g S_{base-1} + sum_{i<g} P^(g-1-i) Z_{base+i}  (P = A^8192, spans in
//            groups of 32, g = position in the group, base = first span of the group); the last span of every group
//            also publishes its true end state, so a clip's chain has n_spans / 32 sequential hops.  Which records a CTA
//            reads and the order it sums them in are FIXED (one lane per record, xor-shuffle tree), so the result does
//            not depend on timing: every run gives the same bits.  (A two-level variant without the chain — group
//            super-aggregates read by every span — measured slower: 345 vs 272 us per 8-clip chain launch.)
//   sweep C: every chunk re-runs the SAME fp32 recurrence from its true entering state and writes clamp(y), so within a
//            chunk the operation order is the sequential reference's.
// Up to three filters per launch: filters 0 and 1 read the same input; filter 2 (CHAIN) reads filter 1's clamped output
// straight from shared memory — mid_pass(x) = low_pass(high_pass(x, 4000), 200) next to low_pass(x) and high_pass(x)
// (features/processing.py:141-150) costs one read of x.  Filter 2's sweep A rides in filters 0/1's sweep C.
// CTAs take their span from a ticket counter, so a CTA only ever waits on CTAs that started before it.
#include "common.cuh"

namespace {

constexpr int L = 32;        // samples per chunk
constexpr int LOG_L = 5;
constexpr int CPB = 256;     // chunks (= threads) per CTA
constexpr int LOG_CPB = 8;
constexpr int LDT = L + 1;   // padded chunk stride in shared memory
constexpr int G = 32;        // spans per look-back group
// powers AL^(2^s) of AL = A^L: s < LOG_CPB inside a span, then P^(2^b) (P = one span, b < 5) across the spans of a group
constexpr int NPOW = LOG_CPB + 5;
static_assert((1 << LOG_L) == L && (1 << LOG_CPB) == CPB, "chunk geometry");

struct Coef {
    float b0, b1, b2, a1, a2;
};

__device__ __forceinline__ float step(const Coef& c, float x0, float x1, float x2, float y1, float y2) {
    // explicit roundings: the contraction of these sums must not depend on what else the compiler sees in the loop (every
    // instantiation has to give the same bits per filter)
    const float v = __fmaf_rn(c.b0, x0, __fmaf_rn(c.b1, x1, __fmul_rn(c.b2, x2)));
    return __fsub_rn(v, __fmaf_rn(c.a1, y1, __fmul_rn(c.a2, y2)));
}

struct M2 {
    double a, b, c, d;  // [[a, b], [c, d]]
};
inline M2 mm_host(const M2& x, const M2& y) {
    return {x.a * y.a + x.b * y.c, x.a * y.b + x.b * y.d, x.c * y.a + x.d * y.c, x.c * y.b + x.d * y.d};
}
struct V2 {
    double x, y;
};
__device__ __forceinline__ V2 mv(const M2& m, const V2& v) { return {m.a * v.x + m.b * v.y, m.c * v.x + m.d * v.y}; }

struct Filter {
    Coef c;
    float* y;          // output signal or NULL
    float* block_ss;   // per-1024-sample sums of squares of the output (+ y[0], y[1024]), or NULL
    M2 p[NPOW];        // AL^(2^s) in fp64 (host-computed once per launch)
};
struct Filters {
    Filter f[3];
};

// look-back records: [filter][clip][span], 16 bytes each.  A record IS its own flag: the launch fills the arrays with
// all-ones words (a NaN pattern no arithmetic produces); a reader polls the record until neither half is that pattern, so
// there is no separate flag, no fence, and a torn 16-byte read only costs another poll.
struct Lookback {
    unsigned* counter;  // span tickets
    V2* agg;            // zero-state end state of the span
    V2* inc;            // true end state of the span (written for the last span of every group)
};
constexpr unsigned long long EMPTY = 0xffffffffffffffffull;

__device__ __forceinline__ void publish(V2* p, V2 v) {
    unsigned long long bx = (unsigned long long)__double_as_longlong(v.x), by = (unsigned long long)__double_as_longlong(v.y);
    if (bx == EMPTY) bx = 0x7ff8000000000000ull;  // (a NaN state stays a NaN, never the marker)
    if (by == EMPTY) by = 0x7ff8000000000000ull;
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(bx), "l"(by) : "memory");
}
__device__ __forceinline__ V2 wait_record(const V2* p) {
    const long long t0 = clock64();
    unsigned long long bx, by;
    while (true) {
        asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(bx), "=l"(by) : "l"(p) : "memory");
        if (bx != EMPTY && by != EMPTY) break;
        if (clock64() - t0 > (1LL << 32)) __trap();  // ~2 s: a predecessor never published (fail loudly, never hang)
    }
    return {__longlong_as_double((long long)bx), __longlong_as_double((long long)by)};
}

// Scan of the CTA's chunk end states z_j (zero entering state per chunk) under S_j = AL S_{j-1} + z_j, S_{-1} = 0, in fp64:
// shuffles inside a warp, the warp totals through shared memory (ONE barrier).  -> incl = S_j, excl = S_{j-1}.
__device__ __forceinline__ void scan_states(V2 z, const M2* pw, V2* wt, int lane, int warp, V2& incl, V2& excl) {
    V2 q = z;
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const double ox = __shfl_up_sync(0xffffffffu, q.x, 1 << s), oy = __shfl_up_sync(0xffffffffu, q.y, 1 << s);
        if (lane >= (1 << s)) {
            const V2 t = mv(pw[s], V2{ox, oy});
            q.x += t.x;
            q.y += t.y;
        }
    }
    if (lane == 31) wt[warp] = q;
    __syncthreads();
    V2 acc = {0.0, 0.0};  // state after the last chunk of the previous warp
    for (int w = 0; w < warp; ++w) {
        acc = mv(pw[5], acc);
        const V2 t = wt[w];
        acc.x += t.x;
        acc.y += t.y;
    }
    V2 c = acc;
    const int e = lane + 1;
#pragma unroll
    for (int b = 0; b < 6; ++b)
        if (e & (1 << b)) c = mv(pw[b], c);
    incl = {q.x + c.x, q.y + c.y};
    const double ux = __shfl_up_sync(0xffffffffu, incl.x, 1), uy = __shfl_up_sync(0xffffffffu, incl.y, 1);
    excl = lane == 0 ? acc : V2{ux, uy};
}

// state entering chunk `tid` of the span: its zero-entering-state value + AL^tid S_in
__device__ __forceinline__ V2 chunk_state_in(V2 excl, const M2* pw, V2 sin, int tid) {
    V2 prop = sin;
#pragma unroll
    for (int b = 0; b < LOG_CPB; ++b)
        if (tid & (1 << b)) prop = mv(pw[b], prop);
    return {excl.x + prop.x, excl.y + prop.y};
}

// True state entering span `span` of (filter, clip): called by ONE warp; Z = this span's own aggregate.
__device__ __forceinline__ V2 span_state_in(const Lookback& lb, const M2* pw, long long rec0, int span, int n_spans, V2 Z,
                                            int lane) {
    const int g = span % G, base = span - g;
    V2 term = {0.0, 0.0};
    if (lane < g) {  // aggregate of span base + lane, carried over the g - 1 - lane spans after it
        term = wait_record(lb.agg + rec0 + base + lane);
        const int e = g - 1 - lane;
#pragma unroll
        for (int b = 0; b < 5; ++b)
            if (e & (1 << b)) term = mv(pw[LOG_CPB + b], term);
    } else if (lane == 31 && base > 0) {  // true end state of the previous group, carried over g spans (g <= 31 < 32)
        term = wait_record(lb.inc + rec0 + base - 1);
#pragma unroll
        for (int b = 0; b < 5; ++b)
            if (g & (1 << b)) term = mv(pw[LOG_CPB + b], term);
    }
    __syncwarp();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {  // fixed tree; a + b == b + a, so every lane holds the same bits
        term.x += __shfl_xor_sync(0xffffffffu, term.x, o);
        term.y += __shfl_xor_sync(0xffffffffu, term.y, o);
    }
    if (lane == 0 && g == G - 1 && span + 1 < n_spans) {
        V2 S = mv(pw[LOG_CPB], term);
        S.x += Z.x;
        S.y += Z.y;
        publish(lb.inc + rec0 + span, S);
    }
    return term;
}

// NF = 1: one filter; 2: two filters of the same input; 3: + filter 2 over filter 1's (clamped) output
template <int NF>
__global__ void __launch_bounds__(CPB) biquad_span_kernel(const float* __restrict__ x, long long n, long long stride_x,
                                                          const Filters fl, long long stride_y, long long n_chunks,
                                                          int n_spans, int n_clips, int clamp, long long n_ss_blocks,
                                                          const Lookback lb) {
    constexpr int NI = NF >= 2 ? 2 : 1;  // filters reading x
    constexpr bool CHAIN = NF == 3;
    constexpr int NW = CPB / 32;
    __shared__ float tile[CPB * LDT];
    extern __shared__ float otile[];  // NF > 1: filter 1's output tile (then filter 2's), CPB * LDT floats
    __shared__ V2 wt[3][NW];
    __shared__ V2 s_z[3], s_in[3];
    __shared__ unsigned s_ticket;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_ticket = atomicAdd(lb.counter, 1u);
    __syncthreads();
    const int span = (int)(s_ticket / (unsigned)n_clips), clip = (int)(s_ticket % (unsigned)n_clips);
    const float* xc = x + (long long)clip * stride_x;
    const long long tile0 = (long long)span * CPB * L;
    const bool full = tile0 + (long long)CPB * L <= n;
    if (full && ((reinterpret_cast<uintptr_t>(xc + tile0) & 15) == 0)) {  // 128-bit loads, 8 in flight per thread
        const float4* __restrict__ x4 = reinterpret_cast<const float4*>(xc + tile0);
#pragma unroll 8
        for (int q = tid; q < CPB * L / 4; q += CPB) {
            const float4 v = __ldg(x4 + q);
            float* d = tile + ((4 * q) / L) * LDT + ((4 * q) % L);
            d[0] = v.x;
            d[1] = v.y;
            d[2] = v.z;
            d[3] = v.w;
        }
    } else {
#pragma unroll 4
        for (int i = tid; i < CPB * L; i += CPB) {
            const long long g = tile0 + i;
            tile[(i / L) * LDT + (i % L)] = g < n ? __ldg(xc + g) : 0.f;
        }
    }
    __syncthreads();
    const long long chunk = (long long)span * CPB + tid;
    const bool live = chunk < n_chunks;
    const long long n0 = chunk * L;
    const bool inside = n0 + L <= n;  // every sample of the chunk is a sample of the clip
    const float xp1 = (live && n0 >= 1) ? __ldg(xc + n0 - 1) : 0.f;
    const float xp2 = (live && n0 >= 2) ? __ldg(xc + n0 - 2) : 0.f;
    float* row = tile + tid * LDT;
    float* orow = otile + (NF > 1 ? tid * LDT : 0);

    // ---- sweep A: zero-state end states of filters 0 (and 1)
    float y1[NI], y2[NI];
#pragma unroll
    for (int f = 0; f < NI; ++f) y1[f] = y2[f] = 0.f;
    if (live) {
        float x1 = xp1, x2 = xp2;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            const float x0 = row[i];
#pragma unroll
            for (int f = 0; f < NI; ++f) {
                const float yy = step(fl.f[f].c, x0, x1, x2, y1[f], y2[f]);
                y2[f] = y1[f];
                y1[f] = yy;
            }
            x2 = x1;
            x1 = x0;
        }
    }
    // ---- scan + look-back: true state entering every chunk
    V2 excl[NI];
#pragma unroll
    for (int f = 0; f < NI; ++f) {
        V2 z = {0.0, 0.0}, incl;
        if (live) z = {(double)y1[f], (double)y2[f]};
        scan_states(z, fl.f[f].p, wt[f], lane, warp, incl, excl[f]);
        if (tid == CPB - 1) {
            s_z[f] = incl;
            publish(lb.agg + ((long long)f * n_clips + clip) * n_spans + span, incl);
        }
    }
    __syncthreads();  // s_z visible
    if (warp < NI) {  // warp f resolves filter f
        const V2 sin = span_state_in(lb, fl.f[warp].p, ((long long)warp * n_clips + clip) * n_spans, span, n_spans, s_z[warp], lane);
        if (lane == 0) s_in[warp] = sin;
    }
    __syncthreads();
    float hy1 = 0.f, hy2 = 0.f;  // filter 1's true entering state (filter 2's input history)
#pragma unroll
    for (int f = 0; f < NI; ++f) {
        const V2 init = chunk_state_in(excl[f], fl.f[f].p, s_in[f], tid);
        y1[f] = (float)init.x;
        y2[f] = (float)init.y;
        if (f == 1) {
            hy1 = y1[f];
            hy2 = y2[f];
        }
    }
    if (clamp) {
        hy1 = fminf(fmaxf(hy1, -1.f), 1.f);
        hy2 = fminf(fmaxf(hy2, -1.f), 1.f);
    }
    // ---- sweep C: outputs of filters 0 (and 1) from the true states; filter 2's sweep A rides along
    float ssq[3] = {0.f, 0.f, 0.f};
    float first[3] = {0.f, 0.f, 0.f};  // output at the chunk's first sample (y[0] / y[1024] of the rms edge frame)
    float m1 = 0.f, m2 = 0.f;
    if (live) {
        float x1 = xp1, x2 = xp2, h1 = hy1, h2 = hy2;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            const float x0 = row[i];
            float yo1 = 0.f;
#pragma unroll
            for (int f = 0; f < NI; ++f) {
                const float yy = step(fl.f[f].c, x0, x1, x2, y1[f], y2[f]);
                const float yo = clamp ? fminf(fmaxf(yy, -1.f), 1.f) : yy;
                if (f == 0) row[i] = yo;
                else orow[i] = yo;
                if (f == 1) yo1 = yo;
                if (inside || n0 + i < n) ssq[f] = fmaf(yo, yo, ssq[f]);
                if (i == 0) first[f] = yo;
                y2[f] = y1[f];
                y1[f] = yy;
            }
            if (CHAIN) {
                const float mmv = step(fl.f[2].c, yo1, h1, h2, m1, m2);
                m2 = m1;
                m1 = mmv;
                h2 = h1;
                h1 = yo1;
            }
            x2 = x1;
            x1 = x0;
        }
    }
    // sum of squares of the (clamped) output per 1024-sample block = 1024 / L consecutive chunks (lanes of one warp)
    constexpr int CPS = 1024 / L;  // chunks per block sum
    static_assert(CPS <= 32 && (CPS & (CPS - 1)) == 0, "block_ss: a block's chunks must sit in one warp");
    auto store_block_ss = [&](int f) {
        if (!fl.f[f].block_ss) return;
        float* bs = fl.f[f].block_ss + (long long)clip * (n_ss_blocks + 2);
        float v = ssq[f];
#pragma unroll
        for (int o = CPS / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        const long long blk = (long long)span * (CPB / CPS) + tid / CPS;
        if (tid % CPS == 0 && blk < n_ss_blocks) bs[blk] = v;
        if (live && (n0 == 0 || n0 == 1024)) bs[n_ss_blocks + (n0 == 0 ? 0 : 1)] = first[f];  // reflected edge of rms frame 0
    };
    auto store_tile = [&](int f, const float* src) {
        if (!fl.f[f].y) return;
        float* yc = fl.f[f].y + (long long)clip * stride_y;
        if (full && ((reinterpret_cast<uintptr_t>(yc + tile0) & 15) == 0)) {
            float4* __restrict__ y4 = reinterpret_cast<float4*>(yc + tile0);
#pragma unroll 4
            for (int qq = tid; qq < CPB * L / 4; qq += CPB) {
                const float* d = src + ((4 * qq) / L) * LDT + ((4 * qq) % L);
                y4[qq] = make_float4(d[0], d[1], d[2], d[3]);
            }
        } else {
            for (int i = tid; i < CPB * L; i += CPB) {
                const long long g = tile0 + i;
                if (g < n) yc[g] = src[(i / L) * LDT + (i % L)];
            }
        }
    };
#pragma unroll
    for (int f = 0; f < NI; ++f) store_block_ss(f);
    V2 exm = {0.0, 0.0};
    if (CHAIN) {  // publish filter 2's aggregate before the stores: its successors' waits overlap them
        V2 z = {0.0, 0.0}, incl;
        if (live) z = {(double)m1, (double)m2};
        scan_states(z, fl.f[2].p, wt[2], lane, warp, incl, exm);  // (its barrier also orders the tile writes before the stores)
        if (tid == CPB - 1) {
            s_z[2] = incl;
            publish(lb.agg + ((long long)2 * n_clips + clip) * n_spans + span, incl);
        }
    } else {
        __syncthreads();
    }
    if (CHAIN && warp == NW - 1) {  // the last warp (it wrote s_z[2]) resolves filter 2 while the others store the tiles
        __syncwarp();
        const V2 sin = span_state_in(lb, fl.f[2].p, ((long long)2 * n_clips + clip) * n_spans, span, n_spans, s_z[2], lane);
        if (lane == 0) s_in[2] = sin;
    }
    store_tile(0, tile);
    if (NF > 1) store_tile(1, otile);
    if (!CHAIN) return;

    // ---- filter 2: its own sweep C over filter 1's tile (in place)
    __syncthreads();  // s_in[2] visible; every thread has read the tiles it stores
    {
        const V2 init = chunk_state_in(exm, fl.f[2].p, s_in[2], tid);
        m1 = (float)init.x;
        m2 = (float)init.y;
    }
    if (live) {
        float h1 = hy1, h2 = hy2;
#pragma unroll
        for (int i = 0; i < L; ++i) {
            const float h0 = orow[i];
            const float yy = step(fl.f[2].c, h0, h1, h2, m1, m2);
            const float yo = clamp ? fminf(fmaxf(yy, -1.f), 1.f) : yy;
            orow[i] = yo;
            if (inside || n0 + i < n) ssq[2] = fmaf(yo, yo, ssq[2]);
            if (i == 0) first[2] = yo;
            m2 = m1;
            m1 = yy;
            h2 = h1;
            h1 = h0;
        }
    }
    store_block_ss(2);
    __syncthreads();
    store_tile(2, otile);
}

// rms(y, frame_length 2048, hop 1024, centre, reflect) (features/audio.py:31-37) from the per-1024-sample sums of squares
// the biquad's output sweep leaves behind: frame t >= 1 covers blocks t - 1 and t; frame 0 covers the reflection
// y[1024..1] followed by y[0..1023].  y is only read for that one edge frame (two samples per clip).
__global__ void rms_from_block_ss_kernel(const float* __restrict__ ss, long long n_blocks, int n_frames,
                                         float* __restrict__ out) {
    const int clip = blockIdx.y;
    const float* s = ss + (long long)clip * (n_blocks + 2);  // [n_blocks sums | y[0] | y[1024]]
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n_frames; t += gridDim.x * blockDim.x) {
        float acc;
        if (t == 0) {
            const float a = s[n_blocks], b = s[n_blocks + 1];
            acc = s[0] + ((s[0] - a * a) + b * b);
        } else {
            acc = s[t - 1] + s[t];
        }
        out[(long long)clip * n_frames + t] = sqrtf(acc / 2048.f);
    }
}

}  // namespace

static long long span_count(long long n_samples) {
    const long long n_chunks = (n_samples + L - 1) / L;
    return (n_chunks + CPB - 1) / CPB;
}

// scratch: [ticket counter (16 bytes, zeroed before every launch) | agg | inc (filled with the EMPTY marker)], room for 3 filters
extern "C" long long ssar_biquad_scratch_bytes(int n_clips, long long n_samples) {
    const long long rec = 3LL * n_clips * span_count(n_samples);
    return 16 + 2 * rec * (long long)sizeof(V2);
}

struct FilterSpec {
    float b0, b1, b2, a1, a2;
    float* y;
    float* block_ss;
};

// n_filters = 1, 2 (same input) or 3 (the third filter reads the second one's output)
static int biquad_impl(const float* x, long long n_samples, long long stride_x, int n_clips, const FilterSpec* specs,
                       int n_filters, int clamp, long long stride_y, void* scratch, void* stream) {
    if (n_clips == 0 || n_samples == 0) return SSAR_OK;
    SSAR_REQUIRE(x && scratch, "biquad: null pointer");
    SSAR_REQUIRE(n_filters >= 1 && n_filters <= 3, "biquad: 1 to 3 filters per launch");
    cudaStream_t st = (cudaStream_t)stream;
    const long long n_chunks = (n_samples + L - 1) / L;
    const long long n_spans = span_count(n_samples);
    const long long n_ss = (n_samples + 1023) / 1024;
    SSAR_REQUIRE(n_spans * n_clips < (1LL << 31), "biquad: too many spans per launch");
    Filters fl;
    for (int f = 0; f < 3; ++f) {
        const FilterSpec& sp = specs[f < n_filters ? f : 0];
        SSAR_REQUIRE(f >= n_filters || sp.y || sp.block_ss || (n_filters == 3 && f == 1),
                     "biquad: a filter needs an output (signal or block sums)");
        SSAR_REQUIRE(f >= n_filters || !sp.y || sp.y != x, "biquad: y must not alias x");
        Filter& F = fl.f[f];
        F.c = {sp.b0, sp.b1, sp.b2, sp.a1, sp.a2};
        F.y = f < n_filters ? sp.y : nullptr;
        F.block_ss = f < n_filters ? sp.block_ss : nullptr;
        // A = companion matrix on (y[n-1], y[n-2]); AL = A^L by repeated squaring, then AL^(2^s)
        M2 AL = {-(double)sp.a1, -(double)sp.a2, 1.0, 0.0};
        for (int i = 0; i < LOG_L; ++i) AL = mm_host(AL, AL);
        F.p[0] = AL;
        for (int i = 1; i < NPOW; ++i) F.p[i] = mm_host(F.p[i - 1], F.p[i - 1]);
    }
    const long long rec = 3LL * n_clips * n_spans;
    char* base = reinterpret_cast<char*>(scratch);
    SSAR_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "biquad: scratch must be 16-byte aligned");
    Lookback lb;
    lb.counter = reinterpret_cast<unsigned*>(base);
    lb.agg = reinterpret_cast<V2*>(base + 16);
    lb.inc = lb.agg + rec;
    cudaError_t e = cudaMemsetAsync(base, 0, 16, st);
    if (e == cudaSuccess) e = cudaMemsetAsync(base + 16, 0xff, (size_t)(2 * rec) * sizeof(V2), st);
    if (e != cudaSuccess) {
        ssar_set_error("biquad: cudaMemsetAsync: %s", cudaGetErrorString(e));
        return SSAR_ERR_CUDA;
    }
    const unsigned grid = (unsigned)(n_spans * n_clips);
    const size_t osm = (size_t)CPB * LDT * sizeof(float);  // + 34 KB static: above the 48 KB default
    if (n_filters > 1) {
        e = n_filters == 2
                ? cudaFuncSetAttribute(biquad_span_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)osm)
                : cudaFuncSetAttribute(biquad_span_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)osm);
        if (e != cudaSuccess) {
            ssar_set_error("biquad: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
            return SSAR_ERR_CUDA;
        }
    }
    if (n_filters == 1)
        biquad_span_kernel<1><<<grid, CPB, 0, st>>>(x, n_samples, stride_x, fl, stride_y, n_chunks, (int)n_spans, n_clips, clamp, n_ss, lb);
    else if (n_filters == 2)
        biquad_span_kernel<2><<<grid, CPB, osm, st>>>(x, n_samples, stride_x, fl, stride_y, n_chunks, (int)n_spans, n_clips, clamp, n_ss, lb);
    else
        biquad_span_kernel<3><<<grid, CPB, osm, st>>>(x, n_samples, stride_x, fl, stride_y, n_chunks, (int)n_spans, n_clips, clamp, n_ss, lb);
    return ssar_launch_status("biquad");
}

extern "C" int ssar_biquad(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                           float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* scratch,
                           void* stream) {
    SSAR_REQUIRE(y != nullptr, "biquad: null pointer");
    const FilterSpec sp = {b0, b1, b2, a1, a2, y, nullptr};
    return biquad_impl(x, n_samples, stride_x, n_clips, &sp, 1, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_biquad_ss(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                              float b2, float a1, float a2, int clamp, float* y, long long stride_y, float* block_ss,
                              void* scratch, void* stream) {
    SSAR_REQUIRE(block_ss != nullptr, "biquad_ss: null block_ss");
    const FilterSpec sp = {b0, b1, b2, a1, a2, y, block_ss};
    return biquad_impl(x, n_samples, stride_x, n_clips, &sp, 1, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_biquad_pair(const float* x, long long n_samples, long long stride_x, int n_clips, const float* coeffs,
                                int clamp, float* y0, float* y1, long long stride_y, float* block_ss0, float* block_ss1,
                                void* scratch, void* stream) {
    SSAR_REQUIRE(coeffs != nullptr, "biquad_pair: null coefficients");
    const FilterSpec sp[2] = {{coeffs[0], coeffs[1], coeffs[2], coeffs[3], coeffs[4], y0, block_ss0},
                              {coeffs[5], coeffs[6], coeffs[7], coeffs[8], coeffs[9], y1, block_ss1}};
    return biquad_impl(x, n_samples, stride_x, n_clips, sp, 2, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_biquad_chain(const float* x, long long n_samples, long long stride_x, int n_clips, const float* coeffs,
                                 int clamp, float* y0, float* y1, float* y2, long long stride_y, float* block_ss0,
                                 float* block_ss1, float* block_ss2, void* scratch, void* stream) {
    SSAR_REQUIRE(coeffs != nullptr, "biquad_chain: null coefficients");
    const FilterSpec sp[3] = {{coeffs[0], coeffs[1], coeffs[2], coeffs[3], coeffs[4], y0, block_ss0},
                              {coeffs[5], coeffs[6], coeffs[7], coeffs[8], coeffs[9], y1, block_ss1},
                              {coeffs[10], coeffs[11], coeffs[12], coeffs[13], coeffs[14], y2, block_ss2}};
    return biquad_impl(x, n_samples, stride_x, n_clips, sp, 3, clamp, stride_y, scratch, stream);
}

extern "C" int ssar_rms_from_block_ss(const float* block_ss, long long n_blocks, int n_clips, int n_frames, float* out,
                                      void* stream) {
    if (n_clips == 0 || n_frames == 0) return SSAR_OK;
    SSAR_REQUIRE(block_ss && out, "rms_from_block_ss: null pointer");
    SSAR_REQUIRE(n_frames <= n_blocks && n_blocks >= 2, "rms_from_block_ss: frame t needs blocks t - 1 and t");
    rms_from_block_ss_kernel<<<dim3(ssar_div_up(n_frames, 256), n_clips), 256, 0, (cudaStream_t)stream>>>(
        block_ss, n_blocks, n_frames, out);
    return ssar_launch_status("rms_from_block_ss");
}

// ---------------------------------------------------------------------------------------------------------------
// ssar_biquad_static: ssar_biquad for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc).  No counterpart in the reference.  The scan scratch
// lives in a file-scope device array of SSAR_BIQUAD_STATIC_SCRATCH_BYTES (2 MiB; every GPU that loads the library holds
// it until exit).  Every call shares the array and it keeps the last call's contents: what has to start from a known state
// (the span ticket counter = 0, the look-back records = the EMPTY marker) is reset on the stream before the kernel; calls
// that overlap on one GPU (two streams, two threads) race on it.
// ---------------------------------------------------------------------------------------------------------------
#include <mutex>

namespace {
constexpr long long STATIC_SCRATCH_BYTES = SSAR_BIQUAD_STATIC_SCRATCH_BYTES;
static __device__ __align__(256) unsigned char g_biquad_static_scratch[STATIC_SCRATCH_BYTES];

constexpr int MAX_DEVICES = 64;
std::mutex g_static_mu;
void* g_static_ptr[MAX_DEVICES] = {};  // device address of the array on every device, resolved once per device

// In host code the array's name is a host-side stand-in, not a device address: resolve it with the device current.
int static_scratch_ptr(void** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= MAX_DEVICES) {
        ssar_set_error("biquad_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return SSAR_ERR_CUDA;
    }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_biquad_static_scratch);
        if (e != cudaSuccess || !p) {
            ssar_set_error("biquad_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return SSAR_ERR_CUDA;
        }
        g_static_ptr[dev] = p;
    }
    *out = g_static_ptr[dev];
    return SSAR_OK;
}
}  // namespace

extern "C" int ssar_biquad_static(const float* x, long long n_samples, long long stride_x, int n_clips, float b0, float b1,
                                  float b2, float a1, float a2, int clamp, float* y, long long stride_y, void* stream) {
    if (n_clips == 0 || n_samples == 0) return SSAR_OK;
    SSAR_REQUIRE(x != nullptr && y != nullptr, "biquad_static: null pointer");
    SSAR_REQUIRE(n_clips > 0 && n_samples > 0, "biquad_static: negative size");
    const long long need = ssar_biquad_scratch_bytes(n_clips, n_samples);
    SSAR_REQUIRE(need <= STATIC_SCRATCH_BYTES,
                 "biquad_static: %d clips x %lld samples need %lld bytes of scratch, the static array holds %lld "
                 "(use ssar_biquad with caller-provided scratch)", n_clips, n_samples, need, STATIC_SCRATCH_BYTES);
    void* scratch = nullptr;
    const int rc = static_scratch_ptr(&scratch);
    if (rc != SSAR_OK) return rc;
    const FilterSpec sp = {b0, b1, b2, a1, a2, y, nullptr};
    return biquad_impl(x, n_samples, stride_x, n_clips, &sp, 1, clamp, stride_y, scratch, stream);
}
