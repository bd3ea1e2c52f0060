// Persistent, weight-stationary training kernel — one cooperative launch runs `nsteps` optimisation steps
// of the MNIST MLP at a small per-rank batch (the launch/latency-bound regime).
//
// Why: at batch 64 a step moves 2.5 MB and does 32 MFLOP (0.4 us of HBM time); separate kernels spend their
// time in launch gaps and dependent global round trips (profiles/r01_v1_*).  Here the grid stays resident and
// the 401 KB fc1 weight matrix never leaves the chip: CTA c < 112 OWNS the W1 tile [16 hidden units x 56
// pixels], keeps it in shared memory across steps, computes that tile's contribution to z1 (phase A) and that
// tile's gradient (phase C) from the same decoded pixel tile, and applies the SGD update in place.
//
//   phase A  (CTAs 0..111)      part[nt][b][16 units] = X[b][56 px of tile nt] . W1tile^T
//   phase B  (CTA r -> rows r, r+148, ..)  z1 = sum of 14 partials + b1 -> ReLU -> dropout -> fc2 -> ReLU -> fc3
//            -> log-softmax/NLL -> dlogits -> dz2 -> dz1   (W2 in registers, W3 / biases in shared memory)
//   phase C  CTAs 0..111: dW1 tile (+ SGD in smem);  112..143: dW2 tiles, db2;  144..147: dW3 tiles, batch loss
//
// Round 2: NO grid barrier between the phases.  Everything one phase hands to the next travels as 8-byte
// (fp32 value, 32-bit step tag) words in L2 - the "flag in the data" protocol the cross-GPU exchange already used:
// the producer stores the word, the consumer polls the very word it needs until the tag is this step's.  A grid
// barrier cost 3,800 cycles between A and B and 2,600 between B and C (release fence behind the CTA's stores +
// atomic + poll, profiles/r02_*timeline*); a tagged word is visible one L2 hop after it was stored.  Buffers are
// double-buffered by step parity: a producer can reach step t+2 only after every consumer of step t has finished
// with it (each phase needs the output of ALL CTAs of the phase before).  The updated W2 / W3 / b2 are published
// through one monotonic counter (release by their 36 owner CTAs, sampled by every thread of the row CTAs); the
// updated b1 travels as tagged words.
//
// Every sum has a fixed order: results are bit-identical run to run and independent of timing.
#include <cuda_pipeline.h>
#include <stdlib.h>

#include "common.cuh"

#ifndef DDPM_POLL_BACKOFF_NS
#define DDPM_POLL_BACKOFF_NS 0           // sleep between polls of a word that has not arrived (0 = spin).  Measured with
                                         // %globaltimer stamps (scripts/persist_timeline_abs.py): with __nanosleep(64) a CTA
                                         // that had been waiting for a while saw a word 1.3 - 2.3 us after it was stored,
                                         // a CTA that had just started polling after 0.58 us; spinning pollers do not slow
                                         // the L2 down measurably (scripts/micro/bcast_bench.cu)
#endif

namespace ddpm {

constexpr int P_THREADS = 256;
constexpr int MT = 16, NT = 56;                       // W1 tile: 16 hidden units x 56 input pixels
constexpr int N_MT = HID / MT, N_NT = IN / NT;        // 8 x 14
constexpr int N_W1 = N_MT * N_NT;                     // 112 owner CTAs
constexpr int N_W2T = 32;                             // dW2 tiles: 8 row tiles (16 rows) x 4 column tiles (32 cols)
constexpr int N_W3T = 4;                              // dW3 column tiles (32 cols); the same CTAs do db1, db2, loss
constexpr int CTA_W2 = N_W1, CTA_W3 = N_W1 + N_W2T;
constexpr int P_GRID = CTA_W3 + N_W3T;                // 148 = one CTA per SM
constexpr int N_SMALL = N_W2T + N_W3T;                // 36 CTAs update W2, W3, b2
constexpr int N_READY = N_SMALL;                      // bumps of the 'ready' counter per step (b1 travels as tagged words)
constexpr int CR = 64;                                // batch rows per smem chunk
constexpr int XS_LD = 60;                             // padded row of the decoded pixel tile (16B aligned)
constexpr int W1_ACTIVE = 4 * 4 * (NT / 4);           // 224 threads hold the 16 x 56 gradient tile (4 values each)
#define DDPM_MAX_WORLD 8
constexpr int ACT_DZ1 = 0, ACT_DZ2 = 128, ACT_H1D = 256, ACT_H2 = 384, ACT_DZ3 = 512;   // words of a row in ll_act
constexpr int XACT_ROW_WORDS = HID;                   // cross-GPU dz1 exchange: 128 words per (parity, source rank, row)
constexpr long long POLL_TIMEOUT_CYCLES = 6000000000LL;   // ~3 s

struct PersistArgs {
    float* params; float* grads;
    const uint8_t* images; const uint8_t* labels; int64_t n_samples;
    const int64_t* idx; const uint8_t* masks;
    uint64_t seed, step0;
    int B, nsteps; float gscale, lr; int train;
    float* loss2; float* loss_steps;
    unsigned long long* llpart;                       // [2][14][B][128] tagged fc1 partial products
    unsigned long long* llact;                        // [2][B][LL_ACT_WORDS] tagged per-row activations / gradients
    unsigned long long* llloss;                       // [2][B] tagged per-row losses
    unsigned long long* llb1;                         // [2][128] tagged fc1 bias of the NEXT step (its producers get their
                                                      // input last, so it does not travel with the 'ready' counter)
    unsigned long long* ctl;                          // the workspace's control block as u64 words (CTRL_*64)
    int* lerr;                                        // dataflow timeout flag (ctrl word CTRL_ERR32)
    long long* dbg;                                   // optional [grid][8 steps][16] clock64 stamps (profiling aid)
    // ---- fused data-parallel mode (world > 1): gradient tiles are exchanged over NVLink peer memory
    int world, rank, two_stage;
    int two_stage_small;                              // ... for the 71 KB behind fc1's weights (b1, W2, b2, W3)
    unsigned long long comm_step;                     // number of fused steps this communicator ran before this launch
    unsigned long long* xrecv[DDPM_MAX_WORLD];        // every rank's receive buffer [2 parities][world sources][NP] of
                                                      // 8-byte (value, step) words, peer-mapped
    int* xerr;                                        // local error word (exchange timeout)
    // dz1 exchange (batch <= 64): instead of fc1's 401 KB gradient, ranks exchange their dz1 rows (32 KB) as they are
    // produced in phase B, and every rank forms the GLOBAL-batch dW1 tiles itself, summing sources in rank order
    // (all ranks hold the whole dataset and know every shard)
    unsigned long long* xrecv_dz[DDPM_MAX_WORLD];     // every rank's buffer [2 parities][world sources][CR rows][XACT_ROW_WORDS]
    const int64_t* idx_all; size_t idx_stride;        // i64 [world][idx_stride]: every rank's batches of this launch
    int act_exchange;
    float* xout;                                      // exchange-only kernel: reduced values of the last step (or NULL)
    // ---- host-fed mode: the batches of this launch arrive in a device ring while the kernel runs; step s may read
    // its slot once feed_flags[(feed_g0 + s) % feed_slots] == feed_g0 + s + 1
    const unsigned long long* feed_flags; unsigned long long feed_g0; int feed_slots; int* feed_err;
};

struct __align__(16) PersistSmem {
    float w2[HID * HID / 2];                          // 32 KB scratch: partial products (phase B), staging / partial gradients (phase C)
    float w3[NC * HID];
    float b1[HID], b2[HID];
    float w1t[NT * MT];                               // [n][m]  owned W1 tile (persistent across steps)
    float xs[2][CR * XS_LD];                          // decoded pixels of the tile, [row][60]; double buffered
    float xsx[CR * XS_LD];                            // pixel tile of another rank's batch (dz1 exchange)
    float dzt[CR * MT];                               // dz1 tile [row][16]
    float hs[HID], h2s[HID], ds[HID];                 // row vectors
    float red[2 * HID];
    float zs[NCP];
    int ylab;
    unsigned long long epoch0, ready_base;
    float lut[256];
};

__device__ __forceinline__ unsigned long long ld_acquire64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release64(unsigned long long* p) {
    asm volatile("red.release.gpu.global.add.u64 [%0], 1;" :: "l"(p) : "memory");
}

// Called when a polled word has not arrived yet.  Backs off a little (a whole grid polling L2 flat out slows the
// producers down), and every 16th miss looks at the clock and at the error word: a wait gives up after ~3 s, and
// once ANY wait has given up every later one gives up within microseconds (the error word is sticky), so a dead
// peer or a lost producer can never hold the non-preemptible grid for more than a few seconds in total.
__device__ __forceinline__ bool poll_miss(int* err, long long t0, unsigned& spins) {
    ++spins;
#if DDPM_POLL_BACKOFF_NS > 0
    if (spins > 2) __nanosleep(spins > 6 ? DDPM_POLL_BACKOFF_NS : DDPM_POLL_BACKOFF_NS / 2);
#endif
    if ((spins & 15u) == 0u) {
        if (*(volatile int*)err != 0) return true;
        if (clock64() - t0 > POLL_TIMEOUT_CYCLES) { *err = 1; return true; }
    }
    return false;
}

// host-fed mode: slot of step `step` filled?  The copy stream carries pixels, labels and then a stream memory
// operation that writes the sequence word behind a system-wide fence (cuStreamWriteValue64); the data loads that
// follow go to L2 (.cg), never to a stale L1 line.
__device__ __forceinline__ unsigned long long feed_word(const PersistArgs& a, int step) {
    const unsigned long long g = a.feed_g0 + (unsigned long long)step;
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(a.feed_flags + (g & (unsigned long long)(a.feed_slots - 1))) : "memory");
    return v;
}
__device__ __forceinline__ bool feed_ready(const PersistArgs& a, int step) {
    return feed_word(a, step) == a.feed_g0 + (unsigned long long)step + 1;
}
__device__ __forceinline__ void feed_wait(const PersistArgs& a, int step) {
    if (feed_ready(a, step)) return;
    if (*(volatile int*)a.feed_err) return;                                // already timed out once: do not wait again
    const long long t0 = clock64();
    while (!feed_ready(a, step))
        if (clock64() - t0 > POLL_TIMEOUT_CYCLES) { *a.feed_err = 1; break; }     // ~3 s: the host stopped feeding
}

__device__ __forceinline__ int w1_idx(int m, int n) { return (((n >> 2) * MT + m) << 2) + (n & 3); }   // [n/4][m][n%4]

// Pixel tile [rows c0..c0+CR) x [56 pixels of tile nt]: loads first (all index loads, then all
// pixel loads: two dependent round trips in total), LUT decode + store to smem later, so the
// latency can sit behind the wait for the next phase's operands.
constexpr int PX_PER_ROW = NT / 4;                                         // 14 words of 4 pixels
constexpr int PX_ITERS = (CR * PX_PER_ROW + P_THREADS - 1) / P_THREADS;    // 4
struct PixelRegs { int64_t sidx[PX_ITERS]; uint32_t v[PX_ITERS]; unsigned ok; };

__device__ __forceinline__ void tile_load_idx(PixelRegs& px, const PersistArgs& a, const int64_t* idx, int c0) {
#pragma unroll
    for (int i = 0; i < PX_ITERS; ++i) {
        const int t = threadIdx.x + i * P_THREADS;
        const int row = c0 + t / PX_PER_ROW;
        px.sidx[i] = (t < CR * PX_PER_ROW && row < a.B) ? __ldg(idx + row) : -1;
    }
}
template <bool FEED>
__device__ __forceinline__ void tile_load_px(PixelRegs& px, const PersistArgs& a, int nt) {
    px.ok = 0;
#pragma unroll
    for (int i = 0; i < PX_ITERS; ++i) {
        const int t = threadIdx.x + i * P_THREADS;
        const int q = t % PX_PER_ROW;
        const bool ok = (px.sidx[i] >= 0 && px.sidx[i] < a.n_samples);
        px.ok |= (ok ? 1u : 0u) << i;
        if (FEED) px.v[i] = ok ? __ldcg(reinterpret_cast<const uint32_t*>(a.images + px.sidx[i] * IN + nt * NT) + q) : 0u;
        else      px.v[i] = ok ? __ldg(reinterpret_cast<const uint32_t*>(a.images + px.sidx[i] * IN + nt * NT) + q) : 0u;
    }
}
template <bool FEED>
__device__ __forceinline__ void tile_load(PixelRegs& px, const PersistArgs& a, const int64_t* idx, int c0, int nt) {
    tile_load_idx(px, a, idx, c0);
    tile_load_px<FEED>(px, a, nt);
}

__device__ __forceinline__ void tile_store(const PixelRegs& px, PersistSmem& s, float* xs) {
#pragma unroll
    for (int i = 0; i < PX_ITERS; ++i) {
        const int t = threadIdx.x + i * P_THREADS;
        if (t < CR * PX_PER_ROW) {
            const int r = t / PX_PER_ROW, q = t - r * PX_PER_ROW;
            const bool ok = (px.ok >> i) & 1u;
            const uint32_t v = px.v[i];
            float4 o;
            o.x = ok ? s.lut[v & 0xff] : 0.f;          o.y = ok ? s.lut[(v >> 8) & 0xff] : 0.f;
            o.z = ok ? s.lut[(v >> 16) & 0xff] : 0.f;  o.w = ok ? s.lut[v >> 24] : 0.f;
            *reinterpret_cast<float4*>(&xs[r * XS_LD + q * 4]) = o;
        }
    }
}

__device__ __forceinline__ void load_w2_block(float (&blk)[8][8], const float* W2, int jb, int kb) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
        const float4 v0 = __ldcg(reinterpret_cast<const float4*>(W2 + (size_t)(jb * 8 + r) * HID + kb * 8));
        const float4 v1 = __ldcg(reinterpret_cast<const float4*>(W2 + (size_t)(jb * 8 + r) * HID + kb * 8) + 1);
        blk[r][0] = v0.x; blk[r][1] = v0.y; blk[r][2] = v0.z; blk[r][3] = v0.w;
        blk[r][4] = v1.x; blk[r][5] = v1.y; blk[r][6] = v1.z; blk[r][7] = v1.w;
    }
}

// gout/lr are the per-step destinations: the caller's d_grads and a.lr normally; in fused
// data-parallel mode the local exchange bucket and 0 (the update happens after the exchange).
__device__ __forceinline__ void grad_out(float* gout, float lr, float* params, int off, float g, float pold) {
    if (gout) gout[off] = g;
    if (lr != 0.f) params[off] = fmaf(-lr, g, pold);
}

// ---- tagged words ("flag in the data"): every value travels as one 8-byte word (fp32 value, 32-bit step tag).
// 8-byte stores are single transactions, so a word whose tag equals T carries step T's value: no fences, no
// separate flags.  Used inside the GPU (phase A -> B -> C, through L2) and between GPUs (gradient exchange, straight
// into the peer's receive buffer over NVLink).
template <int N>
__device__ __forceinline__ void ll_store(unsigned long long* dst, unsigned t32, const float (&v)[N]) {
    if (N == 1) {
        asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(dst), "r"(__float_as_uint(v[0])), "r"(t32) : "memory");
    } else {
#pragma unroll
        for (int i = 0; i < N; i += 2)
            asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" :: "l"(dst + i), "r"(__float_as_uint(v[i])),
                         "r"(t32), "r"(__float_as_uint(v[i + (N > 1 ? 1 : 0)])), "r"(t32) : "memory");
    }
}
__device__ __forceinline__ void ll_store1(unsigned long long* dst, unsigned t32, float v) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(dst), "r"(__float_as_uint(v)), "r"(t32) : "memory");
}
// N tagged words at src (N == 1, or even N read as 16-byte pairs): all loads in flight at once, then only what has
// not arrived is re-polled (a dependent round trip per word made a 4-word poll cost 2 L2 round trips even when
// everything had arrived).
template <int N>
__device__ __forceinline__ void ll_load(const unsigned long long* src, unsigned (&x)[N], unsigned (&f)[N]) {
    if (N == 1) {
        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(x[0]), "=r"(f[0]) : "l"(src) : "memory");
    } else {
#pragma unroll
        for (int i = 0; i < N; i += 2)
            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x[i]), "=r"(f[i]), "=r"(x[i + (N > 1 ? 1 : 0)]),
                         "=r"(f[i + (N > 1 ? 1 : 0)]) : "l"(src + i) : "memory");
    }
}
template <int N>
__device__ __forceinline__ bool ll_missing(const unsigned (&f)[N], unsigned t32) {
    bool m = false;
#pragma unroll
    for (int i = 0; i < N; ++i) m |= f[i] != t32;
    return m;
}
template <int N>
__device__ __forceinline__ void ll_poll(int* err, const unsigned long long* src, unsigned t32, long long t0,
                                        float (&got)[N]) {
    unsigned x[N], f[N], spins = 0;
    ll_load<N>(src, x, f);
    while (ll_missing<N>(f, t32)) {
        if (poll_miss(err, t0, spins)) break;
        ll_load<N>(src, x, f);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) got[i] = __uint_as_float(x[i]);
}
// Cross-GPU all-reduce of this thread's N gradient values.  v: this rank's contribution on entry, the sum over all
// ranks (rank order) on exit.
//   one-shot  (a.two_stage == 0): everyone pushes to everyone, everyone sums: 1 NVLink hop, (ws-1) x bytes
//   two-stage (a.two_stage == 1): the values of this THREAD have a home rank `home`; the others push their
//             contribution to the home, the home sums and pushes the result back: 2 hops, 2(ws-1)/ws x bytes.
//             Homes are spread over the threads of every CTA (thread slices of a tile), so on every rank every SM
//             sends and receives the same share (round 1 made whole CTAs home: 14 SMs per rank did all the summing).
// All pushes of a thread come before any of its waits, so diverged lanes of a warp never wait for each other.
// The summing thread has the words of ALL sources in flight at once and re-polls only the missing ones: the wait is
// one L2 round trip after the last source arrived, not one per source.
// Split in two so that a thread with several exchanges (a tile and a bias, two classes of W3) has all of them on the
// wire before it waits for any: ll_ar_begin = this thread's first-hop pushes, ll_ar_end = everything else.
template <int N, bool SMALL = false>
__device__ __forceinline__ void ll_ar_begin(const PersistArgs& a, unsigned long long T, int home, int off, const float (&v)[N]) {
    const unsigned t32 = (unsigned)T;
    const size_t par = (size_t)(T & 1) * a.world;
    const int two_stage = SMALL ? a.two_stage_small : a.two_stage;
    if (two_stage) {
        if (home != a.rank) ll_store<N>(a.xrecv[home] + (par + a.rank) * NP + off, t32, v);
    } else {
        for (int r = 0; r < a.world; ++r)
            if (r != a.rank) ll_store<N>(a.xrecv[r] + (par + a.rank) * NP + off, t32, v);
    }
}
template <int N, bool SMALL = false>
__device__ __forceinline__ void ll_ar_end(const PersistArgs& a, unsigned long long T, int home, int off, float (&v)[N]) {
    const unsigned t32 = (unsigned)T;
    const size_t par = (size_t)(T & 1) * a.world;
    const long long t0 = clock64();
    const int two_stage = SMALL ? a.two_stage_small : a.two_stage;
    const bool is_home = !two_stage || home == a.rank;
    if (!is_home) {
        ll_poll<N>(a.xerr, a.xrecv[a.rank] + (par + home) * NP + off, t32, t0, v);
        return;
    }
    const unsigned long long* mine = a.xrecv[a.rank] + par * NP + off;
    unsigned x[DDPM_MAX_WORLD][N], f[DDPM_MAX_WORLD][N], spins = 0;
#pragma unroll
    for (int r = 0; r < DDPM_MAX_WORLD; ++r) {
        if (r < a.world && r != a.rank) ll_load<N>(mine + (size_t)r * NP, x[r], f[r]);
        else {
#pragma unroll
            for (int i = 0; i < N; ++i) { x[r][i] = __float_as_uint(v[i]); f[r][i] = t32; }
        }
    }
    for (;;) {
        bool missing = false;
#pragma unroll
        for (int r = 0; r < DDPM_MAX_WORLD; ++r) missing |= ll_missing<N>(f[r], t32);
        if (!missing || poll_miss(a.xerr, t0, spins)) break;
#pragma unroll
        for (int r = 0; r < DDPM_MAX_WORLD; ++r)
            if (ll_missing<N>(f[r], t32)) ll_load<N>(mine + (size_t)r * NP, x[r], f[r]);
    }
    float sum[N];
#pragma unroll
    for (int i = 0; i < N; ++i) sum[i] = __uint_as_float(x[0][i]);
#pragma unroll
    for (int r = 1; r < DDPM_MAX_WORLD; ++r)
        if (r < a.world) {
#pragma unroll
            for (int i = 0; i < N; ++i) sum[i] += __uint_as_float(x[r][i]);
        }
    if (two_stage) {
        for (int r = 0; r < a.world; ++r)
            if (r != a.rank) ll_store<N>(a.xrecv[r] + (par + a.rank) * NP + off, t32, sum);
    }
#pragma unroll
    for (int i = 0; i < N; ++i) v[i] = sum[i];
}
template <int N, bool SMALL = false>
__device__ __forceinline__ void ll_allreduce(const PersistArgs& a, unsigned long long T, int home, int off, float (&v)[N]) {
    ll_ar_begin<N, SMALL>(a, T, home, off, v);
    ll_ar_end<N, SMALL>(a, T, home, off, v);
}

// dst[r][0..NCOLS) for rows r < CR from tagged words at base[r * row_stride + c] (polled until they carry tag
// t32); rows >= nrows are zero.  First round: every poll of this thread in flight at once; then re-poll only
// what has not arrived.
template <int NCOLS>
__device__ __forceinline__ void ll_stage(const unsigned long long* base, size_t row_stride, int nrows, unsigned t32,
                                         float* dst, int dst_ld, int* err) {
    constexpr int PAIRS = NCOLS / 2, PER_THREAD = CR * PAIRS / P_THREADS;      // 2 (16 columns) or 4 (32 columns)
    static_assert(CR * PAIRS % P_THREADS == 0, "tile must divide evenly");
    unsigned x0[PER_THREAD], f0[PER_THREAD], x1[PER_THREAD], f1[PER_THREAD];
#pragma unroll
    for (int i = 0; i < PER_THREAD; ++i) {
        const int e = threadIdx.x + i * P_THREADS, r = e / PAIRS, c2 = (e - r * PAIRS) * 2;
        x0[i] = x1[i] = 0u; f0[i] = f1[i] = t32;
        if (r < nrows)
            asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x0[i]), "=r"(f0[i]), "=r"(x1[i]), "=r"(f1[i])
                         : "l"(base + (size_t)r * row_stride + c2) : "memory");
    }
    const long long t0 = clock64();
    unsigned spins = 0;
    for (;;) {                                    // every round re-issues ALL loads that are still missing, together
        bool missing = false;
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) missing |= (f0[i] != t32 || f1[i] != t32);
        if (!missing || poll_miss(err, t0, spins)) break;
#pragma unroll
        for (int i = 0; i < PER_THREAD; ++i) {
            const int e = threadIdx.x + i * P_THREADS, r = e / PAIRS, c2 = (e - r * PAIRS) * 2;
            if (f0[i] != t32 || f1[i] != t32)
                asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(x0[i]), "=r"(f0[i]), "=r"(x1[i]), "=r"(f1[i])
                             : "l"(base + (size_t)r * row_stride + c2) : "memory");
        }
    }
#pragma unroll
    for (int i = 0; i < PER_THREAD; ++i) {
        const int e = threadIdx.x + i * P_THREADS, r = e / PAIRS, c2 = (e - r * PAIRS) * 2;
        *reinterpret_cast<float2*>(&dst[r * dst_ld + c2]) = make_float2(__uint_as_float(x0[i]), __uint_as_float(x1[i]));
    }
}

// Two tiles of the same rows at once (the dW2 / dW3 CTAs need an early and a late quantity of every row): the polls of
// both are in flight together, so the wait ends one L2 round trip after the LATER one arrived instead of two.
template <int NA, int NB>
__device__ __forceinline__ void ll_stage_pair(const unsigned long long* base_a, float* dst_a, int ld_a,
                                              const unsigned long long* base_b, float* dst_b, int ld_b,
                                              size_t row_stride, int nrows, unsigned t32, int* err) {
    constexpr int PA = NA / 2, PB = NB / 2, TA = CR * PA / P_THREADS, TB = CR * PB / P_THREADS;
    static_assert(CR * PA % P_THREADS == 0 && CR * PB % P_THREADS == 0, "tiles must divide evenly");
    unsigned xa0[TA], fa0[TA], xa1[TA], fa1[TA], xb0[TB], fb0[TB], xb1[TB], fb1[TB];
    auto load_a = [&](int i) {
        const int e = threadIdx.x + i * P_THREADS, r = e / PA, c2 = (e - r * PA) * 2;
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(xa0[i]), "=r"(fa0[i]), "=r"(xa1[i]), "=r"(fa1[i])
                     : "l"(base_a + (size_t)r * row_stride + c2) : "memory");
    };
    auto load_b = [&](int i) {
        const int e = threadIdx.x + i * P_THREADS, r = e / PB, c2 = (e - r * PB) * 2;
        asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(xb0[i]), "=r"(fb0[i]), "=r"(xb1[i]), "=r"(fb1[i])
                     : "l"(base_b + (size_t)r * row_stride + c2) : "memory");
    };
#pragma unroll
    for (int i = 0; i < TA; ++i) {
        xa0[i] = xa1[i] = 0u; fa0[i] = fa1[i] = t32;
        if ((threadIdx.x + i * P_THREADS) / PA < nrows) load_a(i);
    }
#pragma unroll
    for (int i = 0; i < TB; ++i) {
        xb0[i] = xb1[i] = 0u; fb0[i] = fb1[i] = t32;
        if ((threadIdx.x + i * P_THREADS) / PB < nrows) load_b(i);
    }
    const long long t0 = clock64();
    unsigned spins = 0;
    for (;;) {
        bool missing = false;
#pragma unroll
        for (int i = 0; i < TA; ++i) missing |= (fa0[i] != t32 || fa1[i] != t32);
#pragma unroll
        for (int i = 0; i < TB; ++i) missing |= (fb0[i] != t32 || fb1[i] != t32);
        if (!missing || poll_miss(err, t0, spins)) break;
#pragma unroll
        for (int i = 0; i < TA; ++i) if (fa0[i] != t32 || fa1[i] != t32) load_a(i);
#pragma unroll
        for (int i = 0; i < TB; ++i) if (fb0[i] != t32 || fb1[i] != t32) load_b(i);
    }
#pragma unroll
    for (int i = 0; i < TA; ++i) {
        const int e = threadIdx.x + i * P_THREADS, r = e / PA, c2 = (e - r * PA) * 2;
        *reinterpret_cast<float2*>(&dst_a[r * ld_a + c2]) = make_float2(__uint_as_float(xa0[i]), __uint_as_float(xa1[i]));
    }
#pragma unroll
    for (int i = 0; i < TB; ++i) {
        const int e = threadIdx.x + i * P_THREADS, r = e / PB, c2 = (e - r * PB) * 2;
        *reinterpret_cast<float2*>(&dst_b[r * ld_b + c2]) = make_float2(__uint_as_float(xb0[i]), __uint_as_float(xb1[i]));
    }
}

// FEED: host-fed mode (batches arrive in a device ring while the kernel runs); DDP: world > 1 (in-kernel gradient
// exchange) - a template parameter so that the single-GPU kernel carries none of the exchange code or its registers
template <bool FEED, bool DDP>
__global__ void __launch_bounds__(P_THREADS, 1)
train_persist_kernel(PersistArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    PersistSmem& s = *reinterpret_cast<PersistSmem*>(smem_raw);
    const int tid = threadIdx.x, cta = blockIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int B = a.B;
    const float invB = 1.0f / (float)B;
    const float scale = a.train ? 1.25f : 1.0f;
    unsigned long long* const ready = a.ctl + CTRL_READY64;
    if (tid == 0) {
        s.epoch0 = ld_acquire64(a.ctl + CTRL_EPOCH64);      // steps this workspace has seen: the base of the step tags
        s.ready_base = ld_acquire64(ready);                 // nobody bumps it before every row CTA has produced step 0's rows
    }
    unsigned long long ready_seen = 0;
    s.lut[tid] = decode_px(tid);

    const bool own_w1 = cta < N_W1;
    const int mt = cta / N_NT, nt = cta - mt * N_NT;
    const int nchunks = (B + CR - 1) / CR;
    const bool single = nchunks == 1;                 // batch <= 64: the pixel tile is prefetched a step ahead
    // home rank of this thread's gradient values in the two-stage exchange (slices of every tile, see ll_allreduce)
    const int home_w1 = DDP ? (tid * a.world) / W1_ACTIVE : 0;
    const int home_sm = DDP ? (tid * a.world) >> 8 : 0;
    PixelRegs px;
    if (own_w1) {
        for (int t = tid; t < MT * (NT / 4); t += P_THREADS) {
            const int m = t / (NT / 4), q = t - m * (NT / 4);
            const float4 v = *reinterpret_cast<const float4*>(a.params + DDPM_OFF_W1 + (size_t)(mt * MT + m) * IN + nt * NT + q * 4);
            *reinterpret_cast<float4*>(&s.w1t[w1_idx(m, q * 4)]) = v;
        }
        if (single) {
            if (FEED) feed_wait(a, 0);
            tile_load<FEED>(px, a, a.idx, 0, nt);
        }
    }
    __syncthreads();                                   // lut, epoch0 ready
    const unsigned long long epoch0 = s.epoch0;
    if (own_w1 && single) tile_store(px, s, s.xs[0]);
    __syncthreads();

// clock stamp that cannot be taken before a preceding bar.sync has actually released (BAR is
// deferred-blocking: a bare clock read right after it reports the issue time): read a shared word first
#ifdef DDPM_DBG_GLOBALTIMER      /* nanoseconds on one clock for the whole GPU: lines CTAs up against each other */
#define DBG_CLOCK_ASM(t, dep) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "r"(dep) : "memory")
#else
#define DBG_CLOCK_ASM(t, dep) asm volatile("mov.u64 %0, %%clock64;" : "=l"(t) : "r"(dep) : "memory")
#endif
#define DBG_STAMP(slot) do { if (a.dbg && tid == 0 && step < 8) { \
        unsigned v__; long long t__; \
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v__) : "r"((unsigned)__cvta_generic_to_shared(&s.zs[0])) : "memory"); \
        DBG_CLOCK_ASM(t__, v__); \
        a.dbg[((size_t)cta * 8 + step) * 16 + (slot)] = t__; } } while (0)
    for (int step = 0; step < a.nsteps; ++step) {
        DBG_STAMP(0);
        const int64_t* idx = a.idx + (size_t)step * B;
        const uint8_t* mask = a.masks ? a.masks + (size_t)step * B * HID : nullptr;
        float* xs_cur = s.xs[step & 1];
        const bool prefetch = own_w1 && single && (step + 1 < a.nsteps);
        constexpr bool ddp = DDP;
        const bool actx = ddp && a.act_exchange && single;           // W1's gradient through the dz1 exchange
        const unsigned long long T = a.comm_step + step + 1;         // step number of this communicator (cross-GPU tags)
        const unsigned long long TL = epoch0 + step + 1;             // step number of this workspace (in-GPU tags)
        const unsigned tl = (unsigned)TL;
        const size_t parl = (size_t)(TL & 1);
        unsigned long long* const my_part = a.llpart + parl * N_NT * B * HID;
        unsigned long long* const my_act = a.llact + parl * B * LL_ACT_WORDS;
        unsigned long long* const my_loss = a.llloss + parl * B;
        float host_loss = 0.f;
        unsigned long long feed_probe = 0;
        // Row CTAs fetch this step's W2 (into registers: thread (jb, kb) keeps the 8x8 block W2[jb*8.., kb*8..], used by
        // both the forward product, reduced over kb, and the backward product, reduced over jb) and W3 / b2 (into smem)
        // once the 'ready' counter shows that the 36 owner CTAs finished last step's updates.  EVERY thread samples the
        // counter itself (no broadcast, no barrier): at the top of the step and again together with the partial-sum
        // polls of stage 1; the loads are issued when the thread's polls are home (see there), else after stage 1.
        // (Tried, %globaltimer-stamped: looking at the sample INSIDE phase A and issuing the loads
        // there - the 64 KB arrive earlier, but the loop around them takes 2,250 ns instead of 1,310 and the step time is
        // the same; a TMA bulk copy into smem issued by one thread - same result.  The chain that binds is
        // dz2 -> dW2 tiles -> release -> this fetch -> z2, profiles/r02_phase_timeline_1gpu_globaltimer.txt.)
        // (No fence: the owners released their updates before bumping the counter; the loads are L2-coherent .cg /
        // cp.async loads issued after the sample showed the count.)
        float blk[8][8];
        bool params_loaded = false;
        const unsigned long long ready_need = s.ready_base + (unsigned long long)N_READY * step;
        if (cta < B && step > 0) ready_seen = ld_relaxed64(ready);
#define PARAM_LOADS() do { \
            load_w2_block(blk, a.params + DDPM_OFF_W2, tid >> 4, tid & 15); \
            for (int t__ = tid; t__ < NC * HID / 4; t__ += P_THREADS) \
                __pipeline_memcpy_async(&s.w3[t__ * 4], a.params + DDPM_OFF_W3 + t__ * 4, 16); \
            if (tid < HID / 4) __pipeline_memcpy_async(&s.b2[tid * 4], a.params + DDPM_OFF_B2 + tid * 4, 16); \
            else if (tid < HID / 2 && step == 0) __pipeline_memcpy_async(&s.b1[(tid - HID / 4) * 4], a.params + DDPM_OFF_B1 + (tid - HID / 4) * 4, 16); \
            __pipeline_commit(); params_loaded = true; DBG_STAMP(13); } while (0)

        // ================================ phase A ================================
        if (own_w1) {
            const int m = tid & 15, rg = tid >> 4;            // 16 row groups x 4 rows
            // (tried: 4 rows x 4 units per thread over four K slices combined through smem - half the LDS per FMA, but
            //  bank conflicts between row groups and the extra pass made phase A 13 % slower; not kept)
            if (prefetch) tile_load_idx(px, a, idx + B, 0);   // next step's sample indices: in flight during phase A
            // early probe of the next step's sequence word by every thread: the load is in flight during phase A and
            // is looked at only before the pixel loads (a probe through shared memory would put an L2 round trip in
            // front of the bar.sync below: + 0.7 us per step, measured)
            if (FEED && prefetch) feed_probe = feed_word(a, step + 1);
            if (FEED && !single && tid == 64) feed_wait(a, step);
            __syncthreads();                                  // last step's in-place W1 tile update is complete
            for (int c = 0; c < nchunks; ++c) {
                if (!single) {
                    __syncthreads();
                    tile_load<FEED>(px, a, idx, c * CR, nt);
                    tile_store(px, s, xs_cur);
                    __syncthreads();
                }
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int n4 = 0; n4 < NT / 4; ++n4) {
                    const float4 w = *reinterpret_cast<const float4*>(&s.w1t[w1_idx(m, n4 * 4)]);
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const float4 x = *reinterpret_cast<const float4*>(&xs_cur[(rg * 4 + r) * XS_LD + n4 * 4]);
                        acc[r] = fmaf(x.x, w.x, acc[r]); acc[r] = fmaf(x.y, w.y, acc[r]);
                        acc[r] = fmaf(x.z, w.z, acc[r]); acc[r] = fmaf(x.w, w.w, acc[r]);
                    }
                }
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int row = c * CR + rg * 4 + r;
                    if (row < B) ll_store1(my_part + ((size_t)nt * B + row) * HID + mt * MT + m, tl, acc[r]);
                }
            }
            // next step's pixels do not depend on the weights: start fetching them now, decode
            // them after phase B; the two round trips hide behind the waits for the rows
            if (prefetch) {
                if (FEED && feed_probe != a.feed_g0 + (unsigned long long)step + 2) feed_wait(a, step + 1);
                tile_load_px<FEED>(px, a, nt);
            }
        }
        DBG_STAMP(1);

        // ================================ phase B ================================
        if (cta < B) {
            const int j = tid & 127, h = tid >> 7;
            const int jb = tid >> 4, kb = tid & 15;
            float* red2 = s.w2;                               // [16][RLD] partial products (padded: conflict-free float4 stores)
            constexpr int RLD = HID + 4;
            bool staged = false;
            for (int row = cta; row < B; row += gridDim.x) {
                if (tid == 32) {                              // label of this row: two dependent loads, off warp 0
                    const int64_t sidx = __ldg(idx + row);
                    if (FEED) feed_wait(a, step);
                    const bool ok = sidx >= 0 && sidx < a.n_samples;
                    s.ylab = !ok ? 0 : (FEED ? __ldcg(a.labels + sidx) : __ldg(a.labels + sidx));   // read after three more barriers
                }
                // ---- 1. z1 = sum of the 14 tile partials (two fixed halves) + b1 ; ReLU ; dropout
                float zp = 0.f, b1v = 0.f;
                {
                    const unsigned long long* p = my_part + ((size_t)(h * 7) * B + row) * HID + j;
                    unsigned vx[7], vf[7];
#pragma unroll
                    for (int i = 0; i < 7; ++i)
                        asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(vx[i]), "=r"(vf[i])
                                     : "l"(p + (size_t)i * B * HID) : "memory");
                    // b1 as updated in the last step (its tagged word): in flight together with the partial sums
                    const bool poll_b1 = step > 0 && h == 0;
                    const unsigned long long* pb = a.llb1 + parl * HID + j;
                    unsigned bx = 0, bf = tl;
                    if (poll_b1) asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(bx), "=r"(bf) : "l"(pb) : "memory");
                    // the launch's first step takes b1 from the parameters too (no tagged word yet), and nothing has to be
                    // waited for: request W2 / W3 / b2 / b1 right away.  Later steps: a fresh sample of the 'ready' counter
                    // travels with the polls (the one from the top of the step is ~1.2 us old by now)
                    if (!params_loaded) {
                        if (step == 0) PARAM_LOADS();
                        else if (ready_seen < ready_need) ready_seen = ld_relaxed64(ready);
                    }
                    const long long t0 = clock64();
                    unsigned spins = 0;
                    for (;;) {                                // re-issue everything still missing, together
                        bool missing = bf != tl;
#pragma unroll
                        for (int i = 0; i < 7; ++i) missing |= vf[i] != tl;
                        if (!missing || poll_miss(a.lerr, t0, spins)) break;
                        if (bf != tl) asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(bx), "=r"(bf) : "l"(pb) : "memory");
#pragma unroll
                        for (int i = 0; i < 7; ++i)
                            if (vf[i] != tl)
                                asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(vx[i]), "=r"(vf[i])
                                             : "l"(p + (size_t)i * B * HID) : "memory");
                        if (!params_loaded && ready_seen < ready_need) ready_seen = ld_relaxed64(ready);
                    }
                    // this thread's polls are home: if last step's W2 / W3 / b2 updates have been published by now (the sample
                    // travelled with the polls), request its share of the 72 KB here, BEHIND its polls (a poll queued
                    // behind the 64 KB came back ~0.5 us later), so that they arrive while stage 1 runs.  (Requesting
                    // CTA-wide after the barrier below instead: the sample is stale by then, 6.7 instead of 6.4 us/step.)
                    if (!params_loaded && ready_seen >= ready_need) { PARAM_LOADS(); DBG_STAMP(14); }
                    b1v = __uint_as_float(bx);
                    if (row == cta) DBG_STAMP(15);
#pragma unroll
                    for (int i = 0; i < 7; ++i) zp += __uint_as_float(vx[i]);
                }
                bool keep = true;
                if (a.train && h == 0) {
                    if (mask) keep = mask[(size_t)row * HID + j] != 0;
                    else keep = philox_keep(a.seed, a.step0 + step, (uint32_t)(row * HID + j));
                }
                s.red[h * HID + j] = zp;
                if (row == cta) DBG_STAMP(12);
                if (!staged && step == 0) __pipeline_wait_prior(0);          // b1 of the launch's first step comes with W3 / b2
                __syncthreads();
                if (row == cta) DBG_STAMP(2);
                DBG_STAMP(6);
                unsigned long long* const arow = my_act + (size_t)row * LL_ACT_WORDS;
                float h1 = 0.f;
                if (h == 0) {
                    const float z = (s.red[j] + s.red[HID + j]) + (step > 0 ? b1v : s.b1[j]);
                    h1 = (z > 0.f && keep) ? z * scale : 0.f;
                    ll_store1(arow + ACT_H1D + j, tl, h1);
                    s.hs[j] = h1;
                }
                // W2 / W3 / b2 are first needed now.  Stage 1 above ran on the partial sums and b1 alone (they arrive
                // ~0.4 us before the 'ready' count of last step's W2 / W3 / b2 updates shows); every thread now waits for
                // the count itself and requests its share of the 72 KB (a request issued before the polls of stage 1 made
                // them queue behind 64 KB: + 0.5 us).
                if (!params_loaded) {
                    const long long tr = clock64();
                    unsigned rspins = 0;
                    while (ready_seen < ready_need) {
                        if (poll_miss(a.lerr, tr, rspins)) break;
                        ready_seen = ld_relaxed64(ready);
                    }
                    PARAM_LOADS();
                    DBG_STAMP(14);
                }
                __syncthreads();
                // ---- 2. z2 = h1d . W2^T + b2 : 8x8 block products, reduced over the 16 column blocks
                {
                    const float4 x0 = *reinterpret_cast<const float4*>(&s.hs[kb * 8]);
                    const float4 x1 = *reinterpret_cast<const float4*>(&s.hs[kb * 8 + 4]);
                    const float x[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
                    float zr[8];
#pragma unroll
                    for (int r = 0; r < 8; ++r) {
                        float v = 0.f;
#pragma unroll
                        for (int c = 0; c < 8; ++c) v = fmaf(blk[r][c], x[c], v);
                        zr[r] = v;
                    }
                    *reinterpret_cast<float4*>(&red2[kb * RLD + jb * 8]) = make_float4(zr[0], zr[1], zr[2], zr[3]);
                    *reinterpret_cast<float4*>(&red2[kb * RLD + jb * 8 + 4]) = make_float4(zr[4], zr[5], zr[6], zr[7]);
                }
                if (!staged) { __pipeline_wait_prior(0); staged = true; }    // W3 / b2 are in smem behind the next barrier
                __syncthreads();
                bool gate2 = false;
                if (h == 0) {
                    float z2 = 0.f;
#pragma unroll
                    for (int q = 0; q < 16; ++q) z2 += red2[q * RLD + j];
                    z2 += s.b2[j];
                    gate2 = z2 > 0.f;
                    const float hv = gate2 ? z2 : 0.f;
                    ll_store1(arow + ACT_H2 + j, tl, hv);
                    s.h2s[j] = hv;
                }
                __syncthreads();
                DBG_STAMP(7);
                // ---- 3. logits: warp w < 5 computes classes 2w and 2w+1 together (one pass, interleaved)
                if (warp < NC / 2) {
                    float v0 = 0.f, v1 = 0.f;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float hv = s.h2s[lane + 32 * i];
                        v0 = fmaf(hv, s.w3[(2 * warp) * HID + lane + 32 * i], v0);
                        v1 = fmaf(hv, s.w3[(2 * warp + 1) * HID + lane + 32 * i], v1);
                    }
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) {
                        v0 += __shfl_xor_sync(0xffffffffu, v0, o);
                        v1 += __shfl_xor_sync(0xffffffffu, v1, o);
                    }
                    if (lane == 0) { s.zs[2 * warp] = v0; s.zs[2 * warp + 1] = v1; }
                }
                __syncthreads();
                // ---- 4+5. log-softmax / NLL / dlogits, evaluated by every thread of the first half from the ten logits
                //      in smem (no warp-0 section, no extra barrier), then dz2 = (dz3 . W3) * (z2 > 0)
                if (h == 0) {
                    float z[NC], e[NC];
                    {
                        const float4 z0 = *reinterpret_cast<const float4*>(&s.zs[0]);
                        const float4 z1 = *reinterpret_cast<const float4*>(&s.zs[4]);
                        const float2 z2v = *reinterpret_cast<const float2*>(&s.zs[8]);
                        z[0] = z0.x; z[1] = z0.y; z[2] = z0.z; z[3] = z0.w; z[4] = z1.x; z[5] = z1.y; z[6] = z1.z; z[7] = z1.w;
                        z[8] = z2v.x; z[9] = z2v.y;
                    }
                    float mx = z[0];
#pragma unroll
                    for (int c = 1; c < NC; ++c) mx = fmaxf(mx, z[c]);
                    float sum = 0.f;
#pragma unroll
                    for (int c = 0; c < NC; ++c) { e[c] = expf(z[c] - mx); sum += e[c]; }
                    const int yv = s.ylab;
                    float d = 0.f, dmine = 0.f, zy = 0.f;
#pragma unroll
                    for (int c = 0; c < NC; ++c) {
                        const float dc = (e[c] / sum - (c == yv ? 1.f : 0.f)) * invB;
                        d = fmaf(dc, s.w3[c * HID + j], d);
                        if (c == j) dmine = dc;
                        if (c == yv) zy = z[c];
                    }
                    if (j < NCP) ll_store1(arow + ACT_DZ3 + j, tl, dmine);          // classes 10..15: zero padding
                    if (j == 0) ll_store1(my_loss + row, tl, -((zy - mx) - logf(sum)));
                    d = gate2 ? d : 0.f;
                    ll_store1(arow + ACT_DZ2 + j, tl, d);
                    s.ds[j] = d;
                }
                __syncthreads();
                DBG_STAMP(8);
                // ---- 6. dz1 = (dz2 . W2) * scale * (kept && z1 > 0): same blocks, reduced over the 16 row blocks
                {
                    const float4 d0 = *reinterpret_cast<const float4*>(&s.ds[jb * 8]);
                    const float4 d1 = *reinterpret_cast<const float4*>(&s.ds[jb * 8 + 4]);
                    const float d[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
                    float dc[8];
#pragma unroll
                    for (int c = 0; c < 8; ++c) {
                        float v = 0.f;
#pragma unroll
                        for (int r = 0; r < 8; ++r) v = fmaf(d[r], blk[r][c], v);
                        dc[c] = v;
                    }
                    *reinterpret_cast<float4*>(&red2[jb * RLD + kb * 8]) = make_float4(dc[0], dc[1], dc[2], dc[3]);
                    *reinterpret_cast<float4*>(&red2[jb * RLD + kb * 8 + 4]) = make_float4(dc[4], dc[5], dc[6], dc[7]);
                }
                __syncthreads();
                if (h == 0) {
                    float v = 0.f;
#pragma unroll
                    for (int q = 0; q < 16; ++q) v += red2[q * RLD + j];
                    const float dzv = (h1 > 0.f) ? v * scale : 0.f;
                    ll_store1(arow + ACT_DZ1 + j, tl, dzv);
                    if (actx) {                               // the same word to every peer (dz1 exchange)
                        const size_t slot = (((size_t)(T & 1) * a.world + a.rank) * CR + row) * XACT_ROW_WORDS + j;
                        for (int r = 0; r < a.world; ++r)
                            if (r != a.rank) ll_store1(a.xrecv_dz[r] + slot, (unsigned)T, dzv);
                    }
                }
                __syncthreads();
            }
        }
        if (prefetch) tile_store(px, s, s.xs[(step + 1) & 1]);    // the other buffer: phase C below still reads xs_cur
        DBG_STAMP(3);

        // ================================ phase C ================================
        if (own_w1) {
            // thread -> (batch quarter qd, 4 units mq*4.., 4 pixels n4*4..): 16 FMAs per two LDS.128
            const bool act = tid < W1_ACTIVE;                   // 224 threads
            const int qd = tid / 56, u = tid - qd * 56, mq = u / (NT / 4), n4 = u - mq * (NT / 4);
            // db1 of this tile's 16 units: by 16 threads of the otherwise idle warp 7 of the pixel-tile-0 CTAs, from the
            // dz1 tile every W1 CTA stages anyway.  The updated bias goes to the next step's rows as tagged words (no
            // release fence in this CTA: its next phase A is on everybody's critical path).
            const bool b1_thread = nt == 0 && tid >= W1_ACTIVE && tid < W1_ACTIVE + MT;
            const int offb1 = DDPM_OFF_B1 + mt * MT + (tid - W1_ACTIVE);
            const float pold_b1 = b1_thread ? __ldcg(a.params + offb1) : 0.f;
            float db_acc = 0.f;
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int jx = 0; jx < 4; ++jx) acc[i][jx] = 0.f;
            // dz1-exchange mode: one pass per SOURCE RANK, in rank order on every rank (identical sums);
            // the next remote source's pixel tile is fetched while the current one is being multiplied
            const int npass = actx ? a.world : nchunks;
            int nxt = -1;
            if (actx) {
                nxt = (a.rank == 0) ? 1 : 0;
                tile_load<FEED>(px, a, a.idx_all + (size_t)nxt * a.idx_stride + (size_t)step * B, 0, nt);
            }
            for (int c = 0; c < npass; ++c) {
                __syncthreads();
                const float* xs_src = xs_cur;
                if (!actx) {
                    if (!single) { tile_load<FEED>(px, a, idx, c * CR, nt); tile_store(px, s, xs_cur); }
                    const int nrows = B - c * CR < CR ? B - c * CR : CR;
                    ll_stage<MT>(my_act + (size_t)(c * CR) * LL_ACT_WORDS + ACT_DZ1 + mt * MT, LL_ACT_WORDS, nrows, tl,
                                 s.dzt, MT, a.lerr);
                } else if (c == a.rank) {
                    ll_stage<MT>(my_act + ACT_DZ1 + mt * MT, LL_ACT_WORDS, B, tl, s.dzt, MT, a.lerr);
                } else {
                    tile_store(px, s, s.xsx);
                    xs_src = s.xsx;
                    nxt = c + 1 + ((c + 1 == a.rank) ? 1 : 0);
                    if (nxt < a.world) tile_load<FEED>(px, a, a.idx_all + (size_t)nxt * a.idx_stride + (size_t)step * B, 0, nt);
                    // rank c's dz1 rows, this tile's columns, from the words it pushed into this rank's buffer
                    ll_stage<MT>(a.xrecv_dz[a.rank] + (((size_t)(T & 1) * a.world + c) * CR) * XACT_ROW_WORDS + mt * MT,
                                 XACT_ROW_WORDS, B, (unsigned)T, s.dzt, MT, a.xerr);
                }
                __syncthreads();
                if (c == 0) DBG_STAMP(4);
                DBG_STAMP(9);
                if (b1_thread) {
                    const int m = tid - W1_ACTIVE;
                    float p[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int r = 0; r < CR; r += 4) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) p[i] += s.dzt[(r + i) * MT + m];
                    }
                    db_acc += (p[0] + p[1]) + (p[2] + p[3]);
                }
                if (act) {
#pragma unroll
                    for (int rr = 0; rr < CR / 4; ++rr) {
                        const int r = qd * (CR / 4) + rr;
                        const float4 d = *reinterpret_cast<const float4*>(&s.dzt[r * MT + mq * 4]);
                        const float4 x = *reinterpret_cast<const float4*>(&xs_src[r * XS_LD + n4 * 4]);
                        const float dv[4] = {d.x, d.y, d.z, d.w}, xv[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i)
#pragma unroll
                            for (int jx = 0; jx < 4; ++jx) acc[i][jx] = fmaf(dv[i], xv[jx], acc[i][jx]);
                    }
                }
            }
            float* pbuf = s.w2;                                  // [4 quarters][16 units][56 px] partial gradients
            if (act) {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    *reinterpret_cast<float4*>(&pbuf[(qd * MT + mq * 4 + i) * NT + n4 * 4]) =
                        make_float4(acc[i][0], acc[i][1], acc[i][2], acc[i][3]);
            }
            __syncthreads();
            DBG_STAMP(10);
            if (act) {
                const int m = tid / (NT / 4), n4b = tid - m * (NT / 4);       // (unit, 4 pixels)
                float4 p[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) p[q] = *reinterpret_cast<const float4*>(&pbuf[(q * MT + m) * NT + n4b * 4]);
                float g[4];
                g[0] = ((p[0].x + p[1].x) + (p[2].x + p[3].x)) * a.gscale;
                g[1] = ((p[0].y + p[1].y) + (p[2].y + p[3].y)) * a.gscale;
                g[2] = ((p[0].z + p[1].z) + (p[2].z + p[3].z)) * a.gscale;
                g[3] = ((p[0].w + p[1].w) + (p[2].w + p[3].w)) * a.gscale;
                const int off = DDPM_OFF_W1 + (mt * MT + m) * IN + nt * NT + n4b * 4;
                if (ddp && !actx) {                             // exchange with every peer, then everyone holds the sum
                    DBG_STAMP(11);
                    ll_allreduce<4>(a, T, home_w1, off, g);
                    DBG_STAMP(12);
                }
                if (a.grads) *reinterpret_cast<float4*>(a.grads + off) = make_float4(g[0], g[1], g[2], g[3]);
                if (a.lr != 0.f) {
                    float4 w = *reinterpret_cast<float4*>(&s.w1t[w1_idx(m, n4b * 4)]);
                    w.x = fmaf(-a.lr, g[0], w.x); w.y = fmaf(-a.lr, g[1], w.y);
                    w.z = fmaf(-a.lr, g[2], w.z); w.w = fmaf(-a.lr, g[3], w.w);
                    *reinterpret_cast<float4*>(&s.w1t[w1_idx(m, n4b * 4)]) = w;
                }
            } else if (b1_thread) {                              // alongside the tile update above
                float gb[1] = {db_acc * a.gscale};
                if (ddp && !actx) ll_allreduce<1, true>(a, T, (tid - W1_ACTIVE) % a.world, offb1, gb);
                grad_out(a.grads, a.lr, a.params, offb1, gb[0], pold_b1);
                if (a.lr != 0.f)
                    ll_store1(a.llb1 + (size_t)((TL + 1) & 1) * HID + mt * MT + (tid - W1_ACTIVE), (unsigned)(TL + 1),
                              fmaf(-a.lr, gb[0], pold_b1));
            }
        } else if (cta < CTA_W3) {
            // ---- dW2 tile [16 rows jt][32 cols kt].  thread -> (batch quarter bq, 4 rows jq*4.., 2 cols k2*2..):
            //      8 FMAs per LDS.128 + LDS.64; the four batch quarters are combined in fixed order.
            const int t2 = cta - CTA_W2, jt = t2 >> 2, kt = t2 & 3;
            float* hst = s.w2;                                   // [CR][32] staging of h1d
            float* pb = s.w2 + CR * 32;                          // [4][16][32] partial gradients
            const int bq = tid >> 6, jq = (tid >> 4) & 3, k2 = tid & 15;
            const int jj = tid >> 4;                             // final pass: (row jj, cols k2*2..)
            const int off = DDPM_OFF_W2 + (jt * 16 + jj) * HID + kt * 32 + k2 * 2;
            const float2 pold = __ldcg(reinterpret_cast<const float2*>(a.params + off));
            const bool do_db = (kt == 0 && tid < 16);
            const float pbias = do_db ? __ldcg(a.params + DDPM_OFF_B2 + jt * 16 + tid) : 0.f;
            float acc[4][2] = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}};
            float db_acc = 0.f;
            for (int c = 0; c < nchunks; ++c) {
                __syncthreads();
                const int nrows = B - c * CR < CR ? B - c * CR : CR;
                const unsigned long long* rows0 = my_act + (size_t)(c * CR) * LL_ACT_WORDS;
                ll_stage_pair<32, MT>(rows0 + ACT_H1D + kt * 32, hst, 32, rows0 + ACT_DZ2 + jt * 16, s.dzt, MT,
                                      LL_ACT_WORDS, nrows, tl, a.lerr);
                __syncthreads();
                if (c == 0) DBG_STAMP(4);
#pragma unroll
                for (int rr = 0; rr < CR / 4; ++rr) {
                    const int r = bq * (CR / 4) + rr;
                    const float4 d = *reinterpret_cast<const float4*>(&s.dzt[r * MT + jq * 4]);
                    const float2 x = *reinterpret_cast<const float2*>(&hst[r * 32 + k2 * 2]);
                    acc[0][0] = fmaf(d.x, x.x, acc[0][0]); acc[0][1] = fmaf(d.x, x.y, acc[0][1]);
                    acc[1][0] = fmaf(d.y, x.x, acc[1][0]); acc[1][1] = fmaf(d.y, x.y, acc[1][1]);
                    acc[2][0] = fmaf(d.z, x.x, acc[2][0]); acc[2][1] = fmaf(d.z, x.y, acc[2][1]);
                    acc[3][0] = fmaf(d.w, x.x, acc[3][0]); acc[3][1] = fmaf(d.w, x.y, acc[3][1]);
                }
                if (do_db) {                                       // db2 for this tile's 16 units
                    float p[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int r = 0; r < CR; r += 4) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) p[i] += s.dzt[(r + i) * MT + tid];
                    }
                    db_acc += (p[0] + p[1]) + (p[2] + p[3]);
                }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
                *reinterpret_cast<float2*>(&pb[(bq * 16 + jq * 4 + i) * 32 + k2 * 2]) = make_float2(acc[i][0], acc[i][1]);
            __syncthreads();
            {
                float2 p[4];
#pragma unroll
                for (int q = 0; q < 4; ++q) p[q] = *reinterpret_cast<const float2*>(&pb[(q * 16 + jj) * 32 + k2 * 2]);
                float g[2] = {((p[0].x + p[1].x) + (p[2].x + p[3].x)) * a.gscale,
                              ((p[0].y + p[1].y) + (p[2].y + p[3].y)) * a.gscale};
                const int offb = DDPM_OFF_B2 + jt * 16 + tid;
                float gb[1] = {db_acc * a.gscale};
                if (ddp) {                                       // tile and bias on the wire before either wait
                    ll_ar_begin<2, true>(a, T, home_sm, off, g);
                    if (do_db) ll_ar_begin<1, true>(a, T, tid % a.world, offb, gb);
                    ll_ar_end<2, true>(a, T, home_sm, off, g);
                    if (do_db) ll_ar_end<1, true>(a, T, tid % a.world, offb, gb);
                }
                grad_out(a.grads, a.lr, a.params, off, g[0], pold.x);
                grad_out(a.grads, a.lr, a.params, off + 1, g[1], pold.y);
                if (do_db) grad_out(a.grads, a.lr, a.params, offb, gb[0], pbias);
            }
        } else {
            // ---- dW3 column tile [10 classes][32 cols].  thread -> (batch eighth bq, column k), all 10 classes:
            //      10 FMAs per LDS + 3 broadcast LDS.128.  CTA kt == 2 also reduces the batch loss.
            const int kt = cta - CTA_W3;
            float* hst = s.w2;                                   // [CR][32] staging of h2
            float* pb = s.w2 + CR * 32;                          // [8][10][32] partial gradients
            const int k = tid & 31, bq = tid >> 5;
            const int c0 = tid >> 5, c1 = c0 + 8;                // final pass: classes c0 and (threads < 64) c1
            const float p0 = __ldcg(a.params + DDPM_OFF_W3 + c0 * HID + kt * 32 + k);
            const float p1 = c1 < NC ? __ldcg(a.params + DDPM_OFF_W3 + c1 * HID + kt * 32 + k) : 0.f;
            float acc[NC];
#pragma unroll
            for (int i = 0; i < NC; ++i) acc[i] = 0.f;
            for (int c = 0; c < nchunks; ++c) {
                __syncthreads();
                const int nrows = B - c * CR < CR ? B - c * CR : CR;
                const unsigned long long* rows0 = my_act + (size_t)(c * CR) * LL_ACT_WORDS;
                ll_stage_pair<32, NCP>(rows0 + ACT_H2 + kt * 32, hst, 32, rows0 + ACT_DZ3, s.dzt, MT,
                                       LL_ACT_WORDS, nrows, tl, a.lerr);
                __syncthreads();
                if (c == 0) DBG_STAMP(4);
#pragma unroll
                for (int rr = 0; rr < CR / 8; ++rr) {
                    const int r = bq * (CR / 8) + rr;
                    const float hv = hst[r * 32 + k];
                    const float4 d0 = *reinterpret_cast<const float4*>(&s.dzt[r * MT]);
                    const float4 d1 = *reinterpret_cast<const float4*>(&s.dzt[r * MT + 4]);
                    const float2 d2 = *reinterpret_cast<const float2*>(&s.dzt[r * MT + 8]);
                    acc[0] = fmaf(d0.x, hv, acc[0]); acc[1] = fmaf(d0.y, hv, acc[1]); acc[2] = fmaf(d0.z, hv, acc[2]);
                    acc[3] = fmaf(d0.w, hv, acc[3]); acc[4] = fmaf(d1.x, hv, acc[4]); acc[5] = fmaf(d1.y, hv, acc[5]);
                    acc[6] = fmaf(d1.z, hv, acc[6]); acc[7] = fmaf(d1.w, hv, acc[7]); acc[8] = fmaf(d2.x, hv, acc[8]);
                    acc[9] = fmaf(d2.y, hv, acc[9]);
                }
            }
#pragma unroll
            for (int i = 0; i < NC; ++i) pb[(bq * NC + i) * 32 + k] = acc[i];
            __syncthreads();
            {
                float g0 = 0.f, g1 = 0.f;
#pragma unroll
                for (int q = 0; q < 8; ++q) g0 += pb[(q * NC + c0) * 32 + k];
                const int o0 = DDPM_OFF_W3 + c0 * HID + kt * 32 + k, o1 = DDPM_OFF_W3 + c1 * HID + kt * 32 + k;
                if (c1 < NC) {
#pragma unroll
                    for (int q = 0; q < 8; ++q) g1 += pb[(q * NC + c1) * 32 + k];
                }
                float ga[1] = {g0 * a.gscale}, gc[1] = {g1 * a.gscale};
                if (ddp) {                                       // both classes on the wire before either wait
                    ll_ar_begin<1, true>(a, T, home_sm, o0, ga);
                    if (c1 < NC) ll_ar_begin<1, true>(a, T, home_sm, o1, gc);
                    ll_ar_end<1, true>(a, T, home_sm, o0, ga);
                    if (c1 < NC) ll_ar_end<1, true>(a, T, home_sm, o1, gc);
                }
                grad_out(a.grads, a.lr, a.params, o0, ga[0], p0);
                if (c1 < NC) grad_out(a.grads, a.lr, a.params, o1, gc[0], p1);
            }
            if (a.lr != 0.f) {                                   // W3 of the next step is in place: publish it BEFORE the loss
                __syncthreads();                                 // bookkeeping below (the rows of the next step wait for
                if (tid == 0) red_release64(ready);              // the slowest of the 36 publishers, and this was it)
            }
            if (kt == 2) {                                       // per-row losses (stored before that row's dz3)
                float lsum = 0.f;
                const long long t0 = clock64();
                for (int row = tid; row < B; row += P_THREADS) {
                    float lv[1];
                    ll_poll<1>(a.lerr, my_loss + row, tl, t0, lv);
                    lsum += lv[0];
                }
                s.red[tid] = lsum;
                __syncthreads();
            }
            if (kt == 2 && warp == 0) {
                // batch loss: 256 strided partial sums -> 32 lanes (fixed order) -> shuffle tree
                float v = 0.f;
#pragma unroll
                for (int q = 0; q < P_THREADS / 32; ++q) v += s.red[q * 32 + lane];
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) {
                    const float mean = v * invB;
                    // host-fed mode: loss_steps is HOST memory; a store over PCIe issued here would sit in front of this
                    // thread's release of the 'ready' counter below (+ ~1 us on the next step's critical path)
                    if (FEED) host_loss = mean;
                    else if (a.loss_steps) a.loss_steps[step] = mean;
                    if (a.loss2) { a.loss2[0] = mean; a.loss2[1] += mean * invB; }
                }
            }
        }
        if (!own_w1 && cta < CTA_W3 && a.lr != 0.f) {         // W2 / b2 of the next step are in place
            __syncthreads();
            if (tid == 0) red_release64(ready);
        }
        if (FEED && a.loss_steps && tid == 0 && cta == CTA_W3 + 2) a.loss_steps[step] = host_loss;
        DBG_STAMP(5);
        // nothing else orders step t+1 after step t: W1 tiles are private, everything else is gated by tags / 'ready'
    }

    // write the owned W1 tile back
    if (own_w1 && a.lr != 0.f) {
        __syncthreads();
        for (int t = tid; t < MT * (NT / 4); t += P_THREADS) {
            const int m = t / (NT / 4), q = t - m * (NT / 4);
            const float4 v = *reinterpret_cast<const float4*>(&s.w1t[w1_idx(m, q * 4)]);
            *reinterpret_cast<float4*>(a.params + DDPM_OFF_W1 + (size_t)(mt * MT + m) * IN + nt * NT + q * 4) = v;
        }
    }
    // the last CTA to leave advances the workspace's step count (the tag base of the next launch).  Every CTA read it at
    // the top of the kernel, and a CTA can only finish after all CTAs took part in the last step.
    if (tid == 0) {
        const unsigned long long old = atomicAdd(a.ctl + CTRL_FIN64, 1ULL);
        if (old % gridDim.x == gridDim.x - 1) a.ctl[CTRL_EPOCH64] = epoch0 + (unsigned long long)a.nsteps;
    }
}

// The gradient exchange alone (measurement aid + test of the exchange): the same threads exchange the same elements
// of the 473,088-byte bucket as in train_persist_kernel<*, true>, `nsteps` times, with no forward/backward work.
// Contribution of rank r to element e at step t: (r + 1) * (1 + (e & 1023)) / 1024 + t (exactly representable sums).
__global__ void __launch_bounds__(P_THREADS, 1)
exchange_only_kernel(PersistArgs a) {
    const int tid = threadIdx.x, cta = blockIdx.x;
    const int home_w1 = (tid * a.world) / W1_ACTIVE, home_sm = (tid * a.world) >> 8;
    const float rk = (float)(a.rank + 1);
    auto contrib = [&](int e, int step) { return rk * (float)(1 + (e & 1023)) * (1.0f / 1024.0f) + (float)(step & 255); };
    for (int step = 0; step < a.nsteps; ++step) {
        const unsigned long long T = a.comm_step + step + 1;
        const bool last = a.xout != nullptr && step == a.nsteps - 1;
        if (cta < N_W1) {
            const int mt = cta / N_NT, nt = cta - mt * N_NT;
            if (tid < W1_ACTIVE) {
                const int m = tid / (NT / 4), n4b = tid - m * (NT / 4);
                const int off = DDPM_OFF_W1 + (mt * MT + m) * IN + nt * NT + n4b * 4;
                float g[4] = {contrib(off, step), contrib(off + 1, step), contrib(off + 2, step), contrib(off + 3, step)};
                ll_allreduce<4>(a, T, home_w1, off, g);
                if (last) *reinterpret_cast<float4*>(a.xout + off) = make_float4(g[0], g[1], g[2], g[3]);
            } else if (nt == 0 && tid < W1_ACTIVE + MT) {
                const int offb = DDPM_OFF_B1 + mt * MT + (tid - W1_ACTIVE);
                float gb[1] = {contrib(offb, step)};
                ll_allreduce<1, true>(a, T, (tid - W1_ACTIVE) % a.world, offb, gb);
                if (last) a.xout[offb] = gb[0];
            }
        } else if (cta < CTA_W3) {
            const int t2 = cta - CTA_W2, jt = t2 >> 2, kt = t2 & 3, jj = tid >> 4, k2 = tid & 15;
            const int off = DDPM_OFF_W2 + (jt * 16 + jj) * HID + kt * 32 + k2 * 2;
            float g[2] = {contrib(off, step), contrib(off + 1, step)};
            const bool do_db = kt == 0 && tid < 16;
            const int offb = DDPM_OFF_B2 + jt * 16 + tid;
            float gb[1] = {contrib(offb, step)};
            ll_ar_begin<2, true>(a, T, home_sm, off, g);
            if (do_db) ll_ar_begin<1, true>(a, T, tid % a.world, offb, gb);
            ll_ar_end<2, true>(a, T, home_sm, off, g);
            if (do_db) ll_ar_end<1, true>(a, T, tid % a.world, offb, gb);
            if (last) { a.xout[off] = g[0]; a.xout[off + 1] = g[1]; }
            if (last && do_db) a.xout[offb] = gb[0];
        } else {
            const int kt = cta - CTA_W3, k = tid & 31, c0 = tid >> 5, c1 = c0 + 8;
            const int o0 = DDPM_OFF_W3 + c0 * HID + kt * 32 + k, o1 = DDPM_OFF_W3 + c1 * HID + kt * 32 + k;
            float ga[1] = {contrib(o0, step)}, gc[1] = {contrib(o1, step)};
            ll_ar_begin<1, true>(a, T, home_sm, o0, ga);
            if (c1 < NC) ll_ar_begin<1, true>(a, T, home_sm, o1, gc);
            ll_ar_end<1, true>(a, T, home_sm, o0, ga);
            if (c1 < NC) ll_ar_end<1, true>(a, T, home_sm, o1, gc);
            if (last) a.xout[o0] = ga[0];
            if (last && c1 < NC) a.xout[o1] = gc[0];
        }
    }
}

}  // namespace ddpm

using namespace ddpm;

static long long* g_persist_dbg = nullptr;
// profiling aid (not part of the public header): clock64 stamps of the first 8 steps of each CTA
extern "C" void ddpm_debug_set_persist_timing(long long* d_buf) { g_persist_dbg = d_buf; }

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int persist_configure() {
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    const int smem = (int)sizeof(PersistSmem);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        int sms = 0;
        DDPM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        const void* fns[4] = {(const void*)train_persist_kernel<false, false>, (const void*)train_persist_kernel<true, false>,
                              (const void*)train_persist_kernel<false, true>, (const void*)train_persist_kernel<true, true>};
        for (const void* fn : fns) {
            int per_sm = 0;
            DDPM_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
            DDPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, P_THREADS, smem));
            DDPM_CHECK_ARG(per_sm * sms >= P_GRID, "persistent kernel: device cannot co-schedule %d blocks", P_GRID);
        }
        configured[dev] = true;
    }
    return 0;
}

static int persist_fill_and_launch(PersistArgs& a, float* d_params, float* d_grads,
                                   const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                                   const int64_t* d_idx, const uint8_t* d_masks, uint64_t philox_seed,
                                   uint64_t philox_step0, int batch, int nsteps, unsigned flags, float grad_scale,
                                   float lr, float* d_loss, float* d_loss_steps, void* d_workspace, void* stream,
                                   const char* who) {
    DDPM_CHECK_ARG(batch > 0 && batch <= PERSIST_MAX_BATCH,
                   "%s: batch %d outside [1, %d] (use ddpm_step_forward/backward for larger batches)", who, batch,
                   PERSIST_MAX_BATCH);
    DDPM_CHECK_ARG(nsteps > 0, "%s: nsteps must be positive (got %d)", who, nsteps);
    DDPM_CHECK_ARG(d_params && d_images && d_labels && d_idx && d_workspace && n_samples > 0,
                   "%s: null pointer or empty dataset", who);
    DDPM_CHECK_ARG((flags & ~DDPM_FWD_PHILOX_MASK) == 0, "%s: unknown flags %#x", who, flags);
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0 && ((uintptr_t)d_params & 15) == 0 &&
                   ((uintptr_t)d_grads & 15) == 0 && ((uintptr_t)d_images & 3) == 0, "%s: misaligned pointer", who);
    if (persist_configure()) return 1;
    char* ws = (char*)d_workspace;
    const WsLayout L = ws_layout(batch);
    a.params = d_params; a.grads = d_grads;
    a.images = d_images; a.labels = d_labels; a.n_samples = n_samples;
    a.idx = d_idx;
    a.masks = (flags & DDPM_FWD_PHILOX_MASK) ? nullptr : d_masks;
    a.seed = philox_seed; a.step0 = philox_step0;
    a.B = batch; a.nsteps = nsteps; a.gscale = grad_scale; a.lr = lr;
    a.train = ((flags & DDPM_FWD_PHILOX_MASK) || d_masks != nullptr) ? 1 : 0;
    a.loss2 = d_loss; a.loss_steps = d_loss_steps;
    a.llpart = (unsigned long long*)(ws + L.ll_part);
    a.llact = (unsigned long long*)(ws + L.ll_act);
    a.llloss = (unsigned long long*)(ws + L.ll_loss);
    a.llb1 = (unsigned long long*)(ws + L.ll_b1);
    a.ctl = (unsigned long long*)(ws + L.ctrl);
    a.lerr = (int*)(ws + L.ctrl) + CTRL_ERR32;
    a.dbg = g_persist_dbg;
    void* kargs[] = {&a};
    const bool feed = a.feed_flags != nullptr, ddp = a.world > 1;
    const void* fn = feed ? (ddp ? (const void*)train_persist_kernel<true, true> : (const void*)train_persist_kernel<true, false>)
                          : (ddp ? (const void*)train_persist_kernel<false, true> : (const void*)train_persist_kernel<false, false>);
    DDPM_CUDA(cudaLaunchCooperativeKernel(fn, dim3(P_GRID), dim3(P_THREADS), kargs, sizeof(PersistSmem), (cudaStream_t)stream));
    return 0;
}

// Launch floor of the step kernel's launch shape: an empty kernel, same grid / block / dynamic shared memory,
// launched cooperatively.  What a step can never be faster than per LAUNCH (not per step: a launch runs many steps).
__global__ void __launch_bounds__(P_THREADS, 1) persist_floor_kernel(int* sink) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    if (sink != nullptr && threadIdx.x == 0 && blockIdx.x == 0xffffffu) *sink = (int)smem_raw[0];
}

extern "C" int ddpm_persist_launch_floor(int nlaunches, void* stream) {
    DDPM_CHECK_ARG(nlaunches > 0 && nlaunches <= 100000, "ddpm_persist_launch_floor: nlaunches %d out of range", nlaunches);
    if (persist_configure()) return 1;
    static bool attr_set[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !attr_set[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute((const void*)persist_floor_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       (int)sizeof(PersistSmem)));
        attr_set[dev] = true;
    }
    int* sink = nullptr;
    void* kargs[] = {&sink};
    for (int i = 0; i < nlaunches; ++i)
        DDPM_CUDA(cudaLaunchCooperativeKernel((const void*)persist_floor_kernel, dim3(P_GRID), dim3(P_THREADS), kargs,
                                              sizeof(PersistSmem), (cudaStream_t)stream));
    return 0;
}

extern "C" int ddpm_train_steps_persistent(
        float* d_params, float* d_grads,
        const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
        const int64_t* d_idx, const uint8_t* d_masks, uint64_t philox_seed, uint64_t philox_step0,
        int batch, int nsteps, unsigned flags, float grad_scale, float lr,
        float* d_loss, float* d_loss_steps, void* d_workspace, void* stream) {
    DDPM_CHECK_ARG(d_grads == nullptr || nsteps == 1 || lr != 0.f,
                   "ddpm_train_steps_persistent: with lr == 0 only one step per launch makes sense");
    DDPM_CHECK_ARG(d_grads != nullptr || lr != 0.f, "ddpm_train_steps_persistent: nothing to do (no d_grads and lr == 0)");
    PersistArgs a;
    memset(&a, 0, sizeof(a));
    a.world = 1;
    return persist_fill_and_launch(a, d_params, d_grads, d_images, d_labels, n_samples, d_idx, d_masks, philox_seed,
                                   philox_step0, batch, nsteps, flags, grad_scale, lr, d_loss, d_loss_steps,
                                   d_workspace, stream, "ddpm_train_steps_persistent");
}

// ------------------------------------------------------------------------------------------
// fused data-parallel communicator: one symmetric buffer per GPU, mapped into every peer process
// ------------------------------------------------------------------------------------------
struct ddpm_comm {
    int rank = 0, world = 1, device = 0;
    void* local = nullptr;                       // cudaMalloc'd: receive buffer [2][world][NP] x 8 B | i32 error
    size_t recv_bytes = 0, dz_bytes = 0;         // gradient words | dz1 words (activation exchange)
    void* peer[DDPM_MAX_WORLD] = {};             // peer[r]: rank r's buffer in this process's address space
    bool opened[DDPM_MAX_WORLD] = {};
    unsigned long long steps = 0;                // fused steps run so far (same on every rank)
    int two_stage = 0;                           // exchange algorithm (see ll_allreduce)
};

extern "C" int ddpm_comm_create(int rank, int world, ddpm_comm** out) {
    DDPM_CHECK_ARG(out != nullptr, "ddpm_comm_create: null output");
    *out = nullptr;
    DDPM_CHECK_ARG(world >= 1 && world <= DDPM_MAX_WORLD && rank >= 0 && rank < world,
                   "ddpm_comm_create: rank %d / world %d unsupported (world <= %d: one NVLink domain)", rank, world,
                   DDPM_MAX_WORLD);
    ddpm_comm* c = new ddpm_comm();
    c->rank = rank; c->world = world;
    {   // default: one hop while the (ws-1)x traffic is small, two hops with 2(ws-1)/ws x traffic beyond that
        const char* e = getenv("DDPM_EXCHANGE");
        c->two_stage = e ? (strcmp(e, "two_stage") == 0) : (world > 2);
    }
    DDPM_CUDA(cudaGetDevice(&c->device));
    c->recv_bytes = 2 * (size_t)world * NP * sizeof(unsigned long long);     // 15 MB at world 8
    c->dz_bytes = 2 * (size_t)world * CR * XACT_ROW_WORDS * sizeof(unsigned long long);     // 1 MB at world 8
    cudaError_t e = cudaMalloc(&c->local, c->recv_bytes + c->dz_bytes + 256);
    if (e != cudaSuccess) { delete c; return set_error("ddpm_comm_create: cudaMalloc failed: %s", cudaGetErrorString(e)); }
    e = cudaMemset(c->local, 0, c->recv_bytes + c->dz_bytes + 256);   // step field 0 never matches a real step (steps start at 1)
    if (e != cudaSuccess) { cudaFree(c->local); delete c; return set_error("ddpm_comm_create: cudaMemset failed: %s", cudaGetErrorString(e)); }
    c->peer[rank] = c->local;
    *out = c;
    return 0;
}

extern "C" int ddpm_comm_ipc_handle(const ddpm_comm* c, void* h_out64) {
    DDPM_CHECK_ARG(c && h_out64, "ddpm_comm_ipc_handle: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    DDPM_CUDA(cudaIpcGetMemHandle((cudaIpcMemHandle_t*)h_out64, c->local));
    return 0;
}

extern "C" int ddpm_comm_connect(ddpm_comm* c, const void* h_handles) {
    DDPM_CHECK_ARG(c && h_handles, "ddpm_comm_connect: null argument");
    const cudaIpcMemHandle_t* hs = (const cudaIpcMemHandle_t*)h_handles;
    for (int r = 0; r < c->world; ++r) {
        if (r == c->rank || c->opened[r]) continue;
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, hs[r], cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess)
            return set_error("ddpm_comm_connect: cannot map rank %d's buffer (%s); the fused data-parallel step needs "
                             "peer access between all GPUs of the job", r, cudaGetErrorString(e));
        c->peer[r] = p; c->opened[r] = true;
    }
    return 0;
}

extern "C" int ddpm_comm_error(const ddpm_comm* c, int* h_err) {
    DDPM_CHECK_ARG(c && h_err, "ddpm_comm_error: null argument");
    DDPM_CUDA(cudaMemcpy(h_err, (char*)c->local + c->recv_bytes + c->dz_bytes, sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" void ddpm_comm_destroy(ddpm_comm* c) {
    if (!c) return;
    for (int r = 0; r < c->world; ++r)
        if (c->opened[r]) cudaIpcCloseMemHandle(c->peer[r]);
    if (c->local) cudaFree(c->local);
    delete c;
}

static void comm_fill(PersistArgs& a, const ddpm_comm* comm) {
    a.world = comm->world; a.rank = comm->rank; a.comm_step = comm->steps;
    a.two_stage = comm->two_stage;
    // the 71 KB behind fc1's weights: one-shot (one NVLink hop instead of two in front of the W2 / W3 update that the
    // next step's rows wait for) unless DDPM_SMALL_EXCHANGE=two_stage
    {
        const char* e = getenv("DDPM_SMALL_EXCHANGE");
        a.two_stage_small = e ? (strcmp(e, "two_stage") == 0) : comm->two_stage;
    }
    for (int r = 0; r < comm->world; ++r) {
        a.xrecv[r] = (unsigned long long*)comm->peer[r];
        a.xrecv_dz[r] = (unsigned long long*)((char*)comm->peer[r] + comm->recv_bytes);
    }
    a.xerr = (int*)((char*)comm->local + comm->recv_bytes + comm->dz_bytes);
}

extern "C" int ddpm_train_steps_fused_ddp(
        ddpm_comm* comm, float* d_params,
        const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
        const int64_t* d_idx, const int64_t* d_idx_all, int64_t idx_stride,
        const uint8_t* d_masks, uint64_t philox_seed, uint64_t philox_step0,
        int batch, int nsteps, unsigned flags, float lr,
        float* d_loss, float* d_loss_steps, void* d_workspace, void* stream) {
    DDPM_CHECK_ARG(comm != nullptr, "ddpm_train_steps_fused_ddp: null communicator");
    DDPM_CHECK_ARG(d_idx_all == nullptr || idx_stride >= (int64_t)nsteps * batch,
                   "ddpm_train_steps_fused_ddp: idx_stride %lld shorter than nsteps*batch", (long long)idx_stride);
    DDPM_CHECK_ARG(lr != 0.f, "ddpm_train_steps_fused_ddp: lr must be non-zero (the update is part of the exchange)");
    for (int r = 0; r < comm->world; ++r)
        DDPM_CHECK_ARG(comm->peer[r] != nullptr, "ddpm_train_steps_fused_ddp: rank %d is not connected (ddpm_comm_connect)", r);
    PersistArgs a;
    memset(&a, 0, sizeof(a));
    comm_fill(a, comm);
    // fc1's gradient through the dz1 exchange when every rank's indices are known and the batch is one chunk
    // (measured, batch 64: ws 2: 9.96 -> 11.0 M img/s, ws 4: 14.8 -> 15.7 M, ws 8: 23.9 -> 19.9 M - every rank then
    //  multiplies ws x 64 rows of pixels per tile, which outweighs the saved traffic beyond 4 ranks; DDPM_ACT_EXCHANGE=1 forces it)
    const bool act_default = comm->world <= 4 || getenv("DDPM_ACT_EXCHANGE") != nullptr;
    a.act_exchange = (d_idx_all != nullptr && batch <= CR && comm->world > 1 && act_default && getenv("DDPM_NO_ACT_EXCHANGE") == nullptr) ? 1 : 0;
    // with the dz1 exchange only 71 KB of the bucket are left to exchange: everyone pushes them to everyone (one NVLink
    // hop, (ws-1) x 142 KB out per GPU) instead of the two-stage push-sum-push that the full 946 KB need beyond 2 ranks
    if (a.act_exchange && getenv("DDPM_EXCHANGE") == nullptr) a.two_stage = a.two_stage_small = 0;
    a.idx_all = d_idx_all; a.idx_stride = (size_t)idx_stride;
    if (d_idx == nullptr && d_idx_all != nullptr) d_idx = d_idx_all + (size_t)comm->rank * idx_stride;
    const int rc = persist_fill_and_launch(a, d_params, nullptr, d_images, d_labels, n_samples, d_idx, d_masks,
                                           philox_seed, philox_step0, batch, nsteps, flags, 1.0f / (float)comm->world, lr,
                                           d_loss, d_loss_steps, d_workspace, stream, "ddpm_train_steps_fused_ddp");
    if (rc == 0) comm->steps += (unsigned long long)nsteps;
    return rc;
}

extern "C" int ddpm_comm_exchange_only(ddpm_comm* comm, int nsteps, float* d_out, void* stream) {
    DDPM_CHECK_ARG(comm != nullptr && comm->world > 1, "ddpm_comm_exchange_only: needs a communicator of world > 1");
    DDPM_CHECK_ARG(nsteps > 0, "ddpm_comm_exchange_only: nsteps must be positive (got %d)", nsteps);
    DDPM_CHECK_ARG(((uintptr_t)d_out & 15) == 0, "ddpm_comm_exchange_only: d_out must be 16-byte aligned");
    for (int r = 0; r < comm->world; ++r)
        DDPM_CHECK_ARG(comm->peer[r] != nullptr, "ddpm_comm_exchange_only: rank %d is not connected (ddpm_comm_connect)", r);
    PersistArgs a;
    memset(&a, 0, sizeof(a));
    comm_fill(a, comm);
    a.nsteps = nsteps; a.xout = d_out;
    void* kargs[] = {&a};
    DDPM_CUDA(cudaLaunchCooperativeKernel((const void*)exchange_only_kernel, dim3(P_GRID), dim3(P_THREADS), kargs, 0,
                                          (cudaStream_t)stream));
    comm->steps += (unsigned long long)nsteps;
    return 0;
}

extern "C" int ddpm_workspace_error(const void* d_workspace, int* h_err) {
    DDPM_CHECK_ARG(d_workspace && h_err, "ddpm_workspace_error: null argument");
    DDPM_CUDA(cudaMemcpy(h_err, (const int*)d_workspace + CTRL_ERR32, sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}

// ------------------------------------------------------------------------------------------
// host-fed training loop: batches stream from pinned host memory into a device ring while the
// persistent kernel consumes them (include/ddpmnist_b200.h, "host-fed training loop")
// ------------------------------------------------------------------------------------------
struct ddpm_feed {
    int device = 0, batch = 0, slots = 0;
    uint8_t* ring_x = nullptr;                   // [slots][batch][784]
    uint8_t* ring_y = nullptr;                   // [slots][batch]
    unsigned long long* flags = nullptr;         // [slots] sequence words, then one i32 error word
    int64_t* ident = nullptr;                    // [slots*batch] = 0, 1, 2, ...: the ring viewed as a dataset
    // step t is fed on copy stream t % nstreams: the three stream-ordered operations of a step (pixels, labels,
    // sequence word) cost ~4.8 us of copy-queue latency on one stream (measured: 14.3 us/step at batch 64 against
    // 8.5 us for the kernel); slots are independent, so consecutive steps overlap on different streams
    static constexpr int MAX_STREAMS = 8;
    int nstreams = 4;
    cudaStream_t copy_stream[MAX_STREAMS] = {};
    cudaEvent_t done[2] = {nullptr, nullptr};    // kernel that last read ring half h has finished
    bool done_valid[2] = {false, false};
    unsigned long long chunks = 0;               // chunk c uses ring half c & 1 and sequence numbers c * slots/2 + i
};

extern "C" void ddpm_feed_destroy(ddpm_feed* f) {
    if (!f) return;
    for (int i = 0; i < ddpm_feed::MAX_STREAMS; ++i)
        if (f->copy_stream[i]) { cudaStreamSynchronize(f->copy_stream[i]); cudaStreamDestroy(f->copy_stream[i]); }
    for (int h = 0; h < 2; ++h) if (f->done[h]) cudaEventDestroy(f->done[h]);
    if (f->ring_x) cudaFree(f->ring_x);
    if (f->ring_y) cudaFree(f->ring_y);
    if (f->flags) cudaFree(f->flags);
    if (f->ident) cudaFree(f->ident);
    delete f;
}

// cuStreamWriteValue64 through the runtime's driver entry point lookup (no link-time dependency on libcuda)
typedef int (*write_value64_fn)(cudaStream_t, unsigned long long, unsigned long long, unsigned);
static write_value64_fn write_value64 = nullptr;
static int resolve_write_value64() {
    if (write_value64) return 0;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
    DDPM_CUDA(cudaGetDriverEntryPoint("cuStreamWriteValue64", &p, cudaEnableDefault, &q));
    DDPM_CHECK_ARG(q == cudaDriverEntryPointSuccess && p != nullptr, "driver does not provide cuStreamWriteValue64 (status %d)", (int)q);
    write_value64 = (write_value64_fn)p;
    return 0;
}

extern "C" int ddpm_feed_create(int batch, int slots, ddpm_feed** out) {
    DDPM_CHECK_ARG(out != nullptr, "ddpm_feed_create: null output");
    *out = nullptr;
    DDPM_CHECK_ARG(batch > 0 && batch <= PERSIST_MAX_BATCH, "ddpm_feed_create: batch %d outside [1, %d]", batch, PERSIST_MAX_BATCH);
    DDPM_CHECK_ARG(slots >= 4 && slots <= 4096 && (slots & (slots - 1)) == 0,
                   "ddpm_feed_create: slots must be a power of two in [4, 4096] (got %d)", slots);
    if (resolve_write_value64()) return 1;
    ddpm_feed* f = new ddpm_feed();
    f->batch = batch; f->slots = slots;
    const size_t nrow = (size_t)slots * batch;
    cudaError_t e = cudaGetDevice(&f->device);
    if (e == cudaSuccess) e = cudaMalloc((void**)&f->ring_x, nrow * IN);
    if (e == cudaSuccess) e = cudaMalloc((void**)&f->ring_y, nrow);
    if (e == cudaSuccess) e = cudaMalloc((void**)&f->flags, (size_t)slots * 8 + 64);
    if (e == cudaSuccess) e = cudaMalloc((void**)&f->ident, nrow * sizeof(int64_t));
    if (e == cudaSuccess) e = cudaMemset(f->flags, 0, (size_t)slots * 8 + 64);       // 0 never matches a sequence word (>= 1)
    if (e == cudaSuccess) {
        int64_t* h = (int64_t*)malloc(nrow * sizeof(int64_t));
        if (!h) e = cudaErrorMemoryAllocation;
        else {
            for (size_t i = 0; i < nrow; ++i) h[i] = (int64_t)i;
            e = cudaMemcpy(f->ident, h, nrow * sizeof(int64_t), cudaMemcpyHostToDevice);
            free(h);
        }
    }
    if (const char* env = getenv("DDPM_FEED_STREAMS")) {
        const int k = atoi(env);
        if (k >= 1 && k <= ddpm_feed::MAX_STREAMS) f->nstreams = k;
    }
    for (int i = 0; i < f->nstreams && e == cudaSuccess; ++i) e = cudaStreamCreateWithFlags(&f->copy_stream[i], cudaStreamNonBlocking);
    for (int h = 0; h < 2 && e == cudaSuccess; ++h) e = cudaEventCreateWithFlags(&f->done[h], cudaEventDisableTiming);
    if (e != cudaSuccess) {
        ddpm_feed_destroy(f);
        return set_error("ddpm_feed_create: %s", cudaGetErrorString(e));
    }
    *out = f;
    return 0;
}

extern "C" int ddpm_feed_error(const ddpm_feed* f, int* h_err) {
    DDPM_CHECK_ARG(f && h_err, "ddpm_feed_error: null argument");
    DDPM_CUDA(cudaMemcpy(h_err, f->flags + f->slots, sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int ddpm_train_steps_host_fed(
        ddpm_feed* f, ddpm_comm* comm, float* d_params,
        const uint8_t* h_pixels, const uint8_t* h_labels,
        const uint8_t* d_masks, uint64_t philox_seed, uint64_t philox_step0,
        int nsteps, unsigned flags, float lr,
        float* d_loss, float* h_loss_steps, void* d_workspace, void* stream) {
    const char* who = "ddpm_train_steps_host_fed";
    DDPM_CHECK_ARG(f != nullptr && h_pixels != nullptr && h_labels != nullptr, "%s: null feed or host buffer", who);
    DDPM_CHECK_ARG(nsteps > 0, "%s: nsteps must be positive (got %d)", who, nsteps);
    DDPM_CHECK_ARG(lr != 0.f, "%s: lr must be non-zero", who);
    int dev = -1;
    DDPM_CUDA(cudaGetDevice(&dev));
    DDPM_CHECK_ARG(dev == f->device, "%s: feed was created on device %d, current device is %d", who, f->device, dev);
    {   // both host buffers must be page-locked: a pageable source would make every cudaMemcpyAsync a staged, blocking copy
        cudaPointerAttributes pa;
        DDPM_CUDA(cudaPointerGetAttributes(&pa, h_pixels));
        DDPM_CHECK_ARG(pa.type == cudaMemoryTypeHost, "%s: h_pixels is not pinned host memory", who);
        DDPM_CUDA(cudaPointerGetAttributes(&pa, h_labels));
        DDPM_CHECK_ARG(pa.type == cudaMemoryTypeHost, "%s: h_labels is not pinned host memory", who);
    }
    float* d_loss_steps = nullptr;
    if (h_loss_steps) {
        cudaError_t e = cudaHostGetDevicePointer((void**)&d_loss_steps, h_loss_steps, 0);
        if (e != cudaSuccess) return set_error("%s: h_loss_steps is not mapped pinned host memory (%s)", who, cudaGetErrorString(e));
    }
    if (comm)
        for (int r = 0; r < comm->world; ++r)
            DDPM_CHECK_ARG(comm->peer[r] != nullptr, "%s: rank %d is not connected (ddpm_comm_connect)", who, r);
    const int B = f->batch, H = f->slots / 2;
    cudaStream_t st = (cudaStream_t)stream;
    for (int done = 0; done < nsteps;) {
        const int n = nsteps - done < H ? nsteps - done : H;
        const int half = (int)(f->chunks & 1);
        const unsigned long long g0 = f->chunks * (unsigned long long)H;
        const size_t slot0 = (size_t)half * H;
        // the copies of this chunk overwrite the half that the kernel of chunk c-2 read
        if (f->done_valid[half])
            for (int k = 0; k < f->nstreams; ++k) DDPM_CUDA(cudaStreamWaitEvent(f->copy_stream[k], f->done[half], 0));
        PersistArgs a;
        memset(&a, 0, sizeof(a));
        a.world = 1;
        if (comm) comm_fill(a, comm);
        a.feed_flags = f->flags; a.feed_g0 = g0; a.feed_slots = f->slots; a.feed_err = (int*)(f->flags + f->slots);
        // the kernel goes first: it starts consuming while the copies below are still being enqueued
        const int rc = persist_fill_and_launch(
            a, d_params, nullptr, f->ring_x, f->ring_y, (int64_t)f->slots * B, f->ident + slot0 * B,
            d_masks ? d_masks + (size_t)done * B * HID : nullptr, philox_seed, philox_step0 + (uint64_t)done, B, n, flags,
            comm ? 1.0f / (float)comm->world : 1.0f, lr, d_loss, d_loss_steps ? d_loss_steps + done : nullptr,
            d_workspace, stream, who);
        if (rc) return rc;
        if (comm) comm->steps += (unsigned long long)n;
        DDPM_CUDA(cudaEventRecord(f->done[half], st));
        f->done_valid[half] = true;
        f->chunks += 1;
        for (int i = 0; i < n; ++i) {
            const size_t slot = slot0 + i, t = (size_t)done + i;
            cudaStream_t cs = f->copy_stream[i % f->nstreams];
            DDPM_CUDA(cudaMemcpyAsync(f->ring_x + slot * B * IN, h_pixels + t * B * IN, (size_t)B * IN, cudaMemcpyHostToDevice, cs));
            DDPM_CUDA(cudaMemcpyAsync(f->ring_y + slot * B, h_labels + t * B, (size_t)B, cudaMemcpyHostToDevice, cs));
            // sequence word: a stream memory operation, i.e. ordered after both copies AND preceded by a system-wide
            // fence - two back-to-back DMA copies alone give a polling kernel no such guarantee
            const int wr = write_value64(cs, (unsigned long long)(uintptr_t)(f->flags + slot), g0 + i + 1, 0u);
            DDPM_CHECK_ARG(wr == 0, "%s: cuStreamWriteValue64 failed (CUresult %d)", who, wr);
        }
        done += n;
    }
    return 0;
}
