// Persistent, weight-stationary training kernel (v2) — one cooperative launch runs `nsteps`
// optimisation steps of the MNIST MLP at a small per-rank batch (the launch/latency-bound regime).
//
// Why: at batch 64 a step moves 2.5 MB and does 32 MFLOP (0.4 us of HBM time); three separate
// kernels spend their time in launch gaps and dependent global round trips (profiles/r01_v1_*).
// Here the grid stays resident, the phases of a step are separated by two grid barriers, and
// the 401 KB fc1 weight matrix never leaves the chip: CTA c < 112 OWNS the W1 tile
// [16 hidden units x 56 pixels], keeps it in shared memory across steps, computes that tile's
// contribution to z1 (phase A) and that tile's gradient (phase C) from the same decoded pixel
// tile, and applies the SGD update in place.  No other CTA ever needs W1, so nothing orders
// phase C of step t against phase A of step t+1.
//
//   phase A  (CTAs 0..111)      part[nt][b][16 units] = X[b][56 px of tile nt] . W1tile^T
//   -- grid barrier --
//   phase B  (CTA r -> rows r, r+G, ..)  z1 = sum of 14 partials + b1 -> ReLU -> dropout -> fc2
//            -> ReLU -> fc3 -> log-softmax/NLL -> dlogits -> dz2 -> dz1   (W2 re-staged per step
//            with cp.async into an XOR-swizzled layout: conflict-free by row and by column)
//   -- grid barrier --
//   phase C  CTAs 0..111: dW1 tile (+SGD in smem, + bucket);  112..119: dW2 tiles;  120: dW3;
//            121: db1, db2;  122: batch loss (fixed-order sum)
//
// Every sum has a fixed order: results are bit-identical run to run and independent of timing.
#include <cooperative_groups.h>
#include <cuda_pipeline.h>

#include "common.cuh"

namespace ddpm {

constexpr int P_THREADS = 256;
constexpr int MT = 16, NT = 56;                       // W1 tile: 16 hidden units x 56 input pixels
constexpr int N_MT = HID / MT, N_NT = IN / NT;        // 8 x 14
constexpr int N_W1 = N_MT * N_NT;                     // 112 owner CTAs
constexpr int N_W2T = 32;                             // dW2 tiles: 8 row tiles (16 rows) x 4 column tiles (32 cols)
constexpr int N_W3T = 4;                              // dW3 column tiles (32 cols); the same CTAs do db1, db2, loss
constexpr int CTA_W2 = N_W1, CTA_W3 = N_W1 + N_W2T;
constexpr int P_GRID = CTA_W3 + N_W3T;                // 148 = one CTA per SM
constexpr int CR = 64;                                // batch rows per smem chunk
constexpr int XS_LD = 60;                             // padded row of the decoded pixel tile (16B aligned)
constexpr int PERSIST_MAX_BATCH = 512;

struct PersistArgs {
    float* params; float* grads;
    const uint8_t* images; const uint8_t* labels; int64_t n_samples;
    const int64_t* idx; const uint8_t* masks;
    uint64_t seed, step0;
    int B, nsteps; float gscale, lr; int train;
    float* loss2; float* loss_steps;
    float* part; float* h1d; float* h2; float* dz2; float* dz1; float* dz3; float* rowloss;
    unsigned* bar;                                    // [0] arrival count, [1] generation
    long long* dbg;                                   // optional [grid][8 steps][8] clock64 stamps (profiling aid)
};

struct __align__(16) PersistSmem {
    float w2[HID * HID];                              // swizzled at float4 granularity; reused as staging in phase C
    float w3[NC * HID];
    float b1[HID], b2[HID];
    float w1t[NT * MT];                               // [n][m]  owned W1 tile (persistent across steps)
    float xs[2][CR * XS_LD];                          // decoded pixels of the tile, [row][60]; double buffered
    float dzt[CR * MT];                               // dz1 tile [row][16]
    float hs[HID], h2s[HID], ds[HID];                 // row vectors
    float red[2 * HID];
    float zs[NCP], dzs[NCP];
    float lut[256];
};

__device__ __forceinline__ unsigned ld_acquire(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Sense-reversing grid barrier; all CTAs are co-resident (cooperative launch).  Arrival is one
// acq_rel atomic (it publishes this CTA's writes: the bar.sync before it makes them visible to
// thread 0, release is cumulative); the last arriver resets the count and bumps the generation
// with a release store; waiters spin on an acquire load.
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& gen) {
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned arrived;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(arrived) : "l"(bar) : "memory");
        if (arrived == gridDim.x - 1) {
            asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(bar), "r"(0u) : "memory");
            asm volatile("st.release.gpu.global.u32 [%0], %1;" :: "l"(bar + 1), "r"(gen + 1) : "memory");
        } else {
            while (ld_acquire(bar + 1) == gen) { }
        }
    }
    gen += 1;
    __syncthreads();
}

__device__ __forceinline__ int w2_slot(int j, int q) { return j * 32 + (q ^ (j & 31)); }   // float4 index

// Pixel tile [rows c0..c0+CR) x [56 pixels of tile nt]: loads first (all index loads, then all
// pixel loads: two dependent round trips in total), LUT decode + store to smem later, so the
// latency can sit behind a barrier wait.
constexpr int PX_PER_ROW = NT / 4;                                         // 14 words of 4 pixels
constexpr int PX_ITERS = (CR * PX_PER_ROW + P_THREADS - 1) / P_THREADS;    // 4
struct PixelRegs { uint32_t v[PX_ITERS]; unsigned ok; };

__device__ __forceinline__ void tile_load(PixelRegs& px, const PersistArgs& a, const int64_t* idx, int c0, int nt) {
    int64_t sidx[PX_ITERS];
#pragma unroll
    for (int i = 0; i < PX_ITERS; ++i) {
        const int t = threadIdx.x + i * P_THREADS;
        const int row = c0 + t / PX_PER_ROW;
        sidx[i] = (t < CR * PX_PER_ROW && row < a.B) ? __ldg(idx + row) : -1;
    }
    px.ok = 0;
#pragma unroll
    for (int i = 0; i < PX_ITERS; ++i) {
        const int t = threadIdx.x + i * P_THREADS;
        const int q = t % PX_PER_ROW;
        const bool ok = (sidx[i] >= 0 && sidx[i] < a.n_samples);
        px.ok |= (ok ? 1u : 0u) << i;
        px.v[i] = ok ? __ldg(reinterpret_cast<const uint32_t*>(a.images + sidx[i] * IN + nt * NT) + q) : 0u;
    }
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

__device__ __forceinline__ void grad_out(const PersistArgs& a, int off, float g, float pold) {
    if (a.grads) a.grads[off] = g;
    if (a.lr != 0.f) a.params[off] = fmaf(-a.lr, g, pold);
}

__global__ void __launch_bounds__(P_THREADS, 1)
train_persist_kernel(PersistArgs a) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    PersistSmem& s = *reinterpret_cast<PersistSmem*>(smem_raw);
    const int tid = threadIdx.x, cta = blockIdx.x;
    const int lane = tid & 31, warp = tid >> 5;
    const int B = a.B;
    const float invB = 1.0f / (float)B;
    const float scale = a.train ? 1.25f : 1.0f;
    unsigned gen = 0;
    if (tid == 0) gen = ld_acquire(&a.bar[1]);        // only thread 0's copy is used by the barrier
    s.lut[tid] = decode_px(tid);

    const bool own_w1 = cta < N_W1;
    const int mt = cta / N_NT, nt = cta - mt * N_NT;
    const int nchunks = (B + CR - 1) / CR;
    const bool single = nchunks == 1;                 // batch <= 64: the pixel tile is prefetched a step ahead
    PixelRegs px;
    if (own_w1) {
        for (int t = tid; t < MT * (NT / 4); t += P_THREADS) {
            const int m = t / (NT / 4), q = t - m * (NT / 4);
            const float4 v = *reinterpret_cast<const float4*>(a.params + DDPM_OFF_W1 + (size_t)(mt * MT + m) * IN + nt * NT + q * 4);
            s.w1t[(q * 4 + 0) * MT + m] = v.x; s.w1t[(q * 4 + 1) * MT + m] = v.y;
            s.w1t[(q * 4 + 2) * MT + m] = v.z; s.w1t[(q * 4 + 3) * MT + m] = v.w;
        }
        if (single) tile_load(px, a, a.idx, 0, nt);
    }
    __syncthreads();                                   // lut ready
    if (own_w1 && single) tile_store(px, s, s.xs[0]);
    __syncthreads();

#define DBG_STAMP(slot) do { if (a.dbg && tid == 0 && step < 8) a.dbg[((size_t)cta * 8 + step) * 8 + (slot)] = clock64(); } while (0)
    for (int step = 0; step < a.nsteps; ++step) {
        DBG_STAMP(0);
        const int64_t* idx = a.idx + (size_t)step * B;
        const uint8_t* mask = a.masks ? a.masks + (size_t)step * B * HID : nullptr;
        float* xs_cur = s.xs[step & 1];
        const bool prefetch = own_w1 && single && (step + 1 < a.nsteps);

        // ================================ phase A ================================
        if (own_w1) {
            const int m = tid & 15, rg = tid >> 4;            // 16 row groups x 4 rows
            __syncthreads();                                  // last step's in-place W1 tile update is complete
            for (int c = 0; c < nchunks; ++c) {
                if (!single) {
                    __syncthreads();
                    tile_load(px, a, idx, c * CR, nt);
                    tile_store(px, s, xs_cur);
                    __syncthreads();
                }
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int n4 = 0; n4 < NT / 4; ++n4) {
                    float w[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) w[i] = s.w1t[(n4 * 4 + i) * MT + m];
#pragma unroll
                    for (int r = 0; r < 4; ++r) {
                        const float4 x = *reinterpret_cast<const float4*>(&xs_cur[(rg * 4 + r) * XS_LD + n4 * 4]);
                        acc[r] = fmaf(x.x, w[0], acc[r]); acc[r] = fmaf(x.y, w[1], acc[r]);
                        acc[r] = fmaf(x.z, w[2], acc[r]); acc[r] = fmaf(x.w, w[3], acc[r]);
                    }
                }
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const int row = c * CR + rg * 4 + r;
                    if (row < B) a.part[((size_t)nt * B + row) * HID + mt * MT + m] = acc[r];
                }
            }
            // next step's pixels do not depend on the weights: start fetching them now, decode
            // them after phase B; the two round trips hide behind the barrier waits
            if (prefetch) tile_load(px, a, idx + B, 0, nt);
        }
        DBG_STAMP(1);
        grid_barrier(a.bar, gen);
        DBG_STAMP(2);

        // ================================ phase B ================================
        if (cta < B) {
            // stage W2 (swizzled), W3, b1, b2: they were updated by their owners in the last phase C
            const float* W2 = a.params + DDPM_OFF_W2;
            for (int t = tid; t < HID * HID / 4; t += P_THREADS) {
                const int j = t >> 5, q = t & 31;
                __pipeline_memcpy_async(&s.w2[w2_slot(j, q) * 4], W2 + (size_t)t * 4, 16);
            }
            for (int t = tid; t < NC * HID / 4; t += P_THREADS)
                __pipeline_memcpy_async(&s.w3[t * 4], a.params + DDPM_OFF_W3 + t * 4, 16);
            if (tid < HID / 4) __pipeline_memcpy_async(&s.b1[tid * 4], a.params + DDPM_OFF_B1 + tid * 4, 16);
            else if (tid < HID / 2) __pipeline_memcpy_async(&s.b2[(tid - HID / 4) * 4], a.params + DDPM_OFF_B2 + (tid - HID / 4) * 4, 16);
            __pipeline_commit();
            const int j = tid & 127, h = tid >> 7;
            bool staged = false;
            for (int row = cta; row < B; row += gridDim.x) {
                // label: two dependent loads, issued now, consumed in step 4
                int yv = 0;
                if (tid == 0) {
                    const int64_t sidx = __ldg(idx + row);
                    yv = (sidx >= 0 && sidx < a.n_samples) ? __ldg(a.labels + sidx) : 0;
                }
                // ---- 1. z1 = sum of the 14 tile partials (two fixed halves) + b1 ; ReLU ; dropout
                float zp = 0.f;
                {
                    const float* p = a.part + ((size_t)(h * 7) * B + row) * HID + j;
                    float v[7];
#pragma unroll
                    for (int i = 0; i < 7; ++i) v[i] = __ldcg(p + (size_t)i * B * HID);
#pragma unroll
                    for (int i = 0; i < 7; ++i) zp += v[i];
                }
                bool keep = true;
                if (a.train && h == 0) {
                    if (mask) keep = mask[(size_t)row * HID + j] != 0;
                    else keep = philox_keep(a.seed, a.step0 + step, (uint32_t)(row * HID + j));
                }
                s.red[h * HID + j] = zp;
                if (!staged) { __pipeline_wait_prior(0); staged = true; }
                __syncthreads();
                float h1 = 0.f;
                if (h == 0) {
                    const float z = (s.red[j] + s.red[HID + j]) + s.b1[j];
                    h1 = (z > 0.f && keep) ? z * scale : 0.f;
                    a.h1d[(size_t)row * HID + j] = h1;
                    s.hs[j] = h1;
                }
                __syncthreads();
                // ---- 2. z2 = h1d . W2^T + b2 (k range split over the two thread halves) ; ReLU
                {
                    float a0 = 0.f, a1 = 0.f;
#pragma unroll
                    for (int qq = 0; qq < 16; qq += 2) {
                        const int q0 = h * 16 + qq, q1 = q0 + 1;
                        const float4 w0 = *reinterpret_cast<const float4*>(&s.w2[w2_slot(j, q0) * 4]);
                        const float4 w1 = *reinterpret_cast<const float4*>(&s.w2[w2_slot(j, q1) * 4]);
                        const float4 x0 = *reinterpret_cast<const float4*>(&s.hs[q0 * 4]);
                        const float4 x1 = *reinterpret_cast<const float4*>(&s.hs[q1 * 4]);
                        a0 = fmaf(x0.x, w0.x, a0); a0 = fmaf(x0.y, w0.y, a0); a0 = fmaf(x0.z, w0.z, a0); a0 = fmaf(x0.w, w0.w, a0);
                        a1 = fmaf(x1.x, w1.x, a1); a1 = fmaf(x1.y, w1.y, a1); a1 = fmaf(x1.z, w1.z, a1); a1 = fmaf(x1.w, w1.w, a1);
                    }
                    s.red[h * HID + j] = a0 + a1;
                }
                __syncthreads();
                bool gate2 = false;
                if (h == 0) {
                    const float z2 = (s.red[j] + s.red[HID + j]) + s.b2[j];
                    gate2 = z2 > 0.f;
                    const float hv = gate2 ? z2 : 0.f;
                    a.h2[(size_t)row * HID + j] = hv;
                    s.h2s[j] = hv;
                }
                __syncthreads();
                // ---- 3. logits: class c handled by warp c (warps 0..7) and c+8 (warps 0..1)
                for (int c = warp; c < NC; c += 8) {
                    float v = 0.f;
#pragma unroll
                    for (int i = 0; i < 4; ++i) v = fmaf(s.h2s[lane + 32 * i], s.w3[c * HID + lane + 32 * i], v);
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                    if (lane == 0) s.zs[c] = v;
                }
                __syncthreads();
                // ---- 4. log-softmax / NLL / dlogits by warp 0 (lane c < 10 holds class c)
                if (warp == 0) {
                    const float z = lane < NC ? s.zs[lane] : -INFINITY;
                    float m = z;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
                    const float e = lane < NC ? expf(z - m) : 0.f;
                    float sum = e;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
                    yv = __shfl_sync(0xffffffffu, yv, 0);
                    const float zy = __shfl_sync(0xffffffffu, z, yv);
                    if (lane == 0) a.rowloss[row] = -((zy - m) - logf(sum));
                    if (lane < NCP) {
                        const float d = lane < NC ? (e / sum - (lane == yv ? 1.f : 0.f)) * invB : 0.f;
                        s.dzs[lane] = d;
                        a.dz3[(size_t)row * NCP + lane] = d;
                    }
                }
                __syncthreads();
                // ---- 5. dz2 = (dz3 . W3) * (z2 > 0)
                if (h == 0) {
                    float d = 0.f;
#pragma unroll
                    for (int c = 0; c < NC; ++c) d = fmaf(s.dzs[c], s.w3[c * HID + j], d);
                    d = gate2 ? d : 0.f;
                    a.dz2[(size_t)row * HID + j] = d;
                    s.ds[j] = d;
                }
                __syncthreads();
                // ---- 6. dz1 = (dz2 . W2) * scale * (kept && z1 > 0); j range split over the halves
                {
                    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
                    const int kq = j >> 2, kr = j & 3;
#pragma unroll
                    for (int jj = 0; jj < 64; jj += 4) {
                        const int j0 = h * 64 + jj;
                        const float4 d4 = *reinterpret_cast<const float4*>(&s.ds[j0]);
                        a0 = fmaf(d4.x, s.w2[w2_slot(j0 + 0, kq) * 4 + kr], a0);
                        a1 = fmaf(d4.y, s.w2[w2_slot(j0 + 1, kq) * 4 + kr], a1);
                        a2 = fmaf(d4.z, s.w2[w2_slot(j0 + 2, kq) * 4 + kr], a2);
                        a3 = fmaf(d4.w, s.w2[w2_slot(j0 + 3, kq) * 4 + kr], a3);
                    }
                    s.red[h * HID + j] = (a0 + a1) + (a2 + a3);
                }
                __syncthreads();
                if (h == 0) {
                    const float v = s.red[j] + s.red[HID + j];
                    a.dz1[(size_t)row * HID + j] = (h1 > 0.f) ? v * scale : 0.f;
                }
                __syncthreads();
            }
            if (!staged) __pipeline_wait_prior(0);
        }
        if (prefetch) tile_store(px, s, s.xs[(step + 1) & 1]);    // the other buffer: phase C below still reads xs_cur
        DBG_STAMP(3);
        grid_barrier(a.bar, gen);
        DBG_STAMP(4);

        // ================================ phase C ================================
        if (own_w1) {
            const bool act = tid < MT * (NT / 4);               // 224 threads: (m, 4 pixels)
            const int m = tid / (NT / 4), n4 = tid - m * (NT / 4);
            float g[4] = {0.f, 0.f, 0.f, 0.f};
            for (int c = 0; c < nchunks; ++c) {
                __syncthreads();
                if (!single) { tile_load(px, a, idx, c * CR, nt); tile_store(px, s, xs_cur); }
                {
                    const int r = tid >> 2, q = tid & 3, row = c * CR + r;      // 64 rows x 4 float4
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (row < B) v = __ldcg(reinterpret_cast<const float4*>(a.dz1 + (size_t)row * HID + mt * MT) + q);
                    *reinterpret_cast<float4*>(&s.dzt[r * MT + q * 4]) = v;
                }
                __syncthreads();
                if (act) {
#pragma unroll 16
                    for (int r = 0; r < CR; ++r) {
                        const float d = s.dzt[r * MT + m];
                        const float4 x = *reinterpret_cast<const float4*>(&xs_cur[r * XS_LD + n4 * 4]);
                        g[0] = fmaf(d, x.x, g[0]); g[1] = fmaf(d, x.y, g[1]);
                        g[2] = fmaf(d, x.z, g[2]); g[3] = fmaf(d, x.w, g[3]);
                    }
                }
            }
            if (act) {
#pragma unroll
                for (int i = 0; i < 4; ++i) g[i] *= a.gscale;
                const int off = DDPM_OFF_W1 + (mt * MT + m) * IN + nt * NT + n4 * 4;
                if (a.grads) *reinterpret_cast<float4*>(a.grads + off) = make_float4(g[0], g[1], g[2], g[3]);
                if (a.lr != 0.f) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        float& w = s.w1t[(n4 * 4 + i) * MT + m];
                        w = fmaf(-a.lr, g[i], w);
                    }
                }
            }
        } else if (cta < CTA_W3) {
            // ---- dW2 tile [16 rows jt][32 cols kt]; thread -> (row jj, columns 2*k2, 2*k2+1)
            const int t2 = cta - CTA_W2, jt = t2 >> 2, kt = t2 & 3;
            float* hst = s.w2;                                   // [CR][32] staging
            const int jj = tid >> 4, k2 = tid & 15;
            const int off = DDPM_OFF_W2 + (jt * 16 + jj) * HID + kt * 32 + k2 * 2;
            const float2 pold = __ldcg(reinterpret_cast<const float2*>(a.params + off));
            float g0 = 0.f, g1 = 0.f;
            for (int c = 0; c < nchunks; ++c) {
                __syncthreads();
                {
                    const int r = tid >> 2, q = tid & 3, row = c * CR + r;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f), u0 = v, u1 = v;
                    if (row < B) {
                        v = __ldcg(reinterpret_cast<const float4*>(a.dz2 + (size_t)row * HID + jt * 16) + q);
                        u0 = __ldcg(reinterpret_cast<const float4*>(a.h1d + (size_t)row * HID + kt * 32) + q);
                        u1 = __ldcg(reinterpret_cast<const float4*>(a.h1d + (size_t)row * HID + kt * 32) + q + 4);
                    }
                    *reinterpret_cast<float4*>(&s.dzt[r * MT + q * 4]) = v;
                    *reinterpret_cast<float4*>(&hst[r * 32 + q * 4]) = u0;
                    *reinterpret_cast<float4*>(&hst[r * 32 + q * 4 + 16]) = u1;
                }
                __syncthreads();
#pragma unroll 16
                for (int r = 0; r < CR; ++r) {
                    const float d = s.dzt[r * MT + jj];
                    const float2 x = *reinterpret_cast<const float2*>(&hst[r * 32 + k2 * 2]);
                    g0 = fmaf(d, x.x, g0); g1 = fmaf(d, x.y, g1);
                }
            }
            grad_out(a, off, g0 * a.gscale, pold.x);
            grad_out(a, off + 1, g1 * a.gscale, pold.y);
        } else {
            // ---- dW3 column tile [10 classes][32 cols]; then db1 / db2 / loss on three of these CTAs
            const int kt = cta - CTA_W3;
            float* hst = s.w2;                                   // [CR][32] staging of h2
            const int k = tid & 31, cg = tid >> 5;               // classes cg and cg + 8
            const bool two = cg + 8 < NC;
            const int off0 = DDPM_OFF_W3 + cg * HID + kt * 32 + k, off1 = off0 + 8 * HID;
            const float p0 = __ldcg(a.params + off0), p1 = two ? __ldcg(a.params + off1) : 0.f;
            float g0 = 0.f, g1 = 0.f;
            for (int c = 0; c < nchunks; ++c) {
                __syncthreads();
                {
                    const int r = tid >> 2, q = tid & 3, row = c * CR + r;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f), u0 = v, u1 = v;
                    if (row < B) {
                        v = __ldcg(reinterpret_cast<const float4*>(a.dz3 + (size_t)row * NCP) + q);
                        u0 = __ldcg(reinterpret_cast<const float4*>(a.h2 + (size_t)row * HID + kt * 32) + q);
                        u1 = __ldcg(reinterpret_cast<const float4*>(a.h2 + (size_t)row * HID + kt * 32) + q + 4);
                    }
                    *reinterpret_cast<float4*>(&s.dzt[r * MT + q * 4]) = v;
                    *reinterpret_cast<float4*>(&hst[r * 32 + q * 4]) = u0;
                    *reinterpret_cast<float4*>(&hst[r * 32 + q * 4 + 16]) = u1;
                }
                __syncthreads();
#pragma unroll 16
                for (int r = 0; r < CR; ++r) {
                    const float hv = hst[r * 32 + k];
                    g0 = fmaf(s.dzt[r * MT + cg], hv, g0);
                    if (two) g1 = fmaf(s.dzt[r * MT + cg + 8], hv, g1);
                }
            }
            grad_out(a, off0, g0 * a.gscale, p0);
            if (two) grad_out(a, off1, g1 * a.gscale, p1);
            if (kt < 2) {
                // ---- db1 (kt 0) / db2 (kt 1): stage the [rows][128] block, column sums in fixed order
                const float* D = kt == 0 ? a.dz1 : a.dz2;
                const int col = tid & 127, half = tid >> 7;
                const int offb = (kt == 0 ? DDPM_OFF_B1 : DDPM_OFF_B2) + col;
                const float pb = __ldcg(a.params + offb);
                float acc = 0.f;
                float* st = s.w2;                                // [CR][128]
                for (int c = 0; c < nchunks; ++c) {
                    __syncthreads();
                    float4 v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int t = tid + i * P_THREADS, r = t >> 5, q = t & 31, row = c * CR + r;
                        v[i] = row < B ? __ldcg(reinterpret_cast<const float4*>(D + (size_t)row * HID) + q)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int t = tid + i * P_THREADS;
                        *reinterpret_cast<float4*>(&st[t * 4]) = v[i];
                    }
                    __syncthreads();
                    float p[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                    for (int r = 0; r < CR / 2; r += 4) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) p[i] += st[(half * (CR / 2) + r + i) * HID + col];
                    }
                    acc += (p[0] + p[1]) + (p[2] + p[3]);
                }
                __syncthreads();
                s.red[half * HID + col] = acc;
                __syncthreads();
                if (half == 0) grad_out(a, offb, (s.red[col] + s.red[HID + col]) * a.gscale, pb);
            } else if (kt == 2) {
                // ---- batch loss: fixed-order sum of the per-row losses
                float v = 0.f;
                for (int row = tid; row < B; row += P_THREADS) v += __ldcg(a.rowloss + row);
                __syncthreads();
                s.red[tid] = v;
                __syncthreads();
                for (int o = P_THREADS / 2; o > 0; o >>= 1) {
                    if (tid < o) s.red[tid] += s.red[tid + o];
                    __syncthreads();
                }
                if (tid == 0) {
                    const float mean = s.red[0] * invB;
                    if (a.loss_steps) a.loss_steps[step] = mean;
                    if (a.loss2) { a.loss2[0] = mean; a.loss2[1] += mean * invB; }
                }
            }
        }
        DBG_STAMP(5);
        // no barrier here: W1 tiles are private; everything else is ordered by the next step's
        // first barrier (phase B of step t+1 starts only after every CTA finished phase C of t)
    }

    // write the owned W1 tile back
    if (own_w1 && a.lr != 0.f) {
        __syncthreads();
        for (int t = tid; t < MT * (NT / 4); t += P_THREADS) {
            const int m = t / (NT / 4), q = t - m * (NT / 4);
            float4 v;
            v.x = s.w1t[(q * 4 + 0) * MT + m]; v.y = s.w1t[(q * 4 + 1) * MT + m];
            v.z = s.w1t[(q * 4 + 2) * MT + m]; v.w = s.w1t[(q * 4 + 3) * MT + m];
            *reinterpret_cast<float4*>(a.params + DDPM_OFF_W1 + (size_t)(mt * MT + m) * IN + nt * NT + q * 4) = v;
        }
    }
}

}  // namespace ddpm

using namespace ddpm;

static long long* g_persist_dbg = nullptr;
// profiling aid (not part of the public header): clock64 stamps of the first 8 steps of each CTA
extern "C" void ddpm_debug_set_persist_timing(long long* d_buf) { g_persist_dbg = d_buf; }

extern "C" int ddpm_train_steps_persistent(
        float* d_params, float* d_grads,
        const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
        const int64_t* d_idx, const uint8_t* d_masks, uint64_t philox_seed, uint64_t philox_step0,
        int batch, int nsteps, unsigned flags, float grad_scale, float lr,
        float* d_loss, float* d_loss_steps, void* d_workspace, void* stream) {
    DDPM_CHECK_ARG(batch > 0 && batch <= PERSIST_MAX_BATCH,
                   "ddpm_train_steps_persistent: batch %d outside [1, %d] (use ddpm_step_forward/backward for larger batches)",
                   batch, PERSIST_MAX_BATCH);
    DDPM_CHECK_ARG(nsteps > 0, "ddpm_train_steps_persistent: nsteps must be positive (got %d)", nsteps);
    DDPM_CHECK_ARG(d_params && d_images && d_labels && d_idx && d_workspace && n_samples > 0,
                   "ddpm_train_steps_persistent: null pointer or empty dataset");
    DDPM_CHECK_ARG((flags & ~DDPM_FWD_PHILOX_MASK) == 0, "ddpm_train_steps_persistent: unknown flags %#x", flags);
    DDPM_CHECK_ARG(d_grads == nullptr || nsteps == 1 || lr != 0.f,
                   "ddpm_train_steps_persistent: with lr == 0 only one step per launch makes sense");
    DDPM_CHECK_ARG(d_grads != nullptr || lr != 0.f, "ddpm_train_steps_persistent: nothing to do (no d_grads and lr == 0)");
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0 && ((uintptr_t)d_params & 15) == 0 &&
                   ((uintptr_t)d_grads & 15) == 0 && ((uintptr_t)d_images & 3) == 0,
                   "ddpm_train_steps_persistent: misaligned pointer");
    cudaStream_t st = (cudaStream_t)stream;
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    const int smem = (int)sizeof(PersistSmem);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(train_persist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        int per_sm = 0, sms = 0;
        DDPM_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, train_persist_kernel, P_THREADS, smem));
        DDPM_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        DDPM_CHECK_ARG(per_sm * sms >= P_GRID, "ddpm_train_steps_persistent: device cannot co-schedule %d blocks", P_GRID);
        configured[dev] = true;
    }
    char* ws = (char*)d_workspace;
    const WsLayout L = ws_layout(batch);
    PersistArgs a;
    memset(&a, 0, sizeof(a));
    a.params = d_params; a.grads = d_grads;
    a.images = d_images; a.labels = d_labels; a.n_samples = n_samples;
    a.idx = d_idx;
    a.masks = (flags & DDPM_FWD_PHILOX_MASK) ? nullptr : d_masks;
    a.seed = philox_seed; a.step0 = philox_step0;
    a.B = batch; a.nsteps = nsteps; a.gscale = grad_scale; a.lr = lr;
    a.train = ((flags & DDPM_FWD_PHILOX_MASK) || d_masks != nullptr) ? 1 : 0;
    a.loss2 = d_loss; a.loss_steps = d_loss_steps;
    a.part = (float*)(ws + L.part1);
    a.h1d = (float*)(ws + L.h1d); a.h2 = (float*)(ws + L.h2);
    a.dz2 = (float*)(ws + L.dz2); a.dz1 = (float*)(ws + L.dz1);
    a.dz3 = (float*)(ws + L.dz3); a.rowloss = (float*)(ws + L.rowloss);
    a.bar = (unsigned*)(ws + L.ctrl) + 2;
    a.dbg = g_persist_dbg;
    void* kargs[] = {&a};
    DDPM_CUDA(cudaLaunchCooperativeKernel((const void*)train_persist_kernel, dim3(P_GRID), dim3(P_THREADS), kargs,
                                          (size_t)smem, st));
    return 0;
}
