// libddpmnist_b200 — CUDA kernels for the MNIST data-parallel training step (sm_100a).
//
// One training step is three dependent phases:
//   G1    fc1 forward as a split-K GEMM whose A operand is decoded on the fly from the resident
//         uint8 dataset (gather + ToTensor + Normalize fused into the operand load)
//   ROWS  everything that is local to one batch row once z1 is known: bias+ReLU+dropout, fc2,
//         ReLU, fc3, log-softmax/NLL, dlogits, and the data-gradient chain back to dz1
//   GRADS the five weight/bias gradients (reductions over the batch), scaled for DDP and with the
//         SGD update optionally fused into the epilogue
// All arithmetic is fp32 FFMA with fixed summation orders (run-to-run deterministic).
#include <stdarg.h>

#include "common.cuh"

namespace ddpm {

thread_local char g_err[512] = "";

int set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return 1;
}

// ------------------------------------------------------------------------------------------
// workspace layout
// ------------------------------------------------------------------------------------------
static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

constexpr int G1_BM_SMALL = 32, G1_BN_SMALL = 64;
constexpr int G1_BM_LARGE = 64, G1_BN_LARGE = 128;
constexpr int G1_LARGE_BATCH = 1024;
constexpr int TARGET_ITEMS = 296;                 // 2 x 148 SMs

// grads_kernel cuts large batches into slices so that more than 67 CTAs have work
int grads_bsplit(int batch) { return batch >= 1024 ? (batch / 128 < 32 ? batch / 128 : 32) : 1; }

int fc1_ksplit(int batch) {
    const bool large = batch >= G1_LARGE_BATCH;
    const int bm = large ? G1_BM_LARGE : G1_BM_SMALL, bn = large ? G1_BN_LARGE : G1_BN_SMALL;
    const int items_mn = ((batch + bm - 1) / bm) * (HID / bn);
    if (items_mn >= 148) return 1;
    int chunks_per_split = (NCHUNKS * items_mn) / TARGET_ITEMS;
    if (chunks_per_split < 1) chunks_per_split = 1;
    return (NCHUNKS + chunks_per_split - 1) / chunks_per_split;
}

size_t part1_rows_capacity(int max_batch) {
    size_t cap = 0;
    for (int b = 1; b <= max_batch; ++b) {
        size_t need = (size_t)fc1_ksplit(b) * (size_t)b;
        if (need > cap) cap = need;
    }
    return cap;
}

static WsLayout ws_layout_compute(int max_batch);

// ws_layout is called on every step; part1_rows_capacity loops over the batch, so memoise.
WsLayout ws_layout(int max_batch) {
    thread_local int cached_b[4] = {0, 0, 0, 0};
    thread_local WsLayout cached_l[4];
    thread_local int next = 0;
    for (int i = 0; i < 4; ++i)
        if (cached_b[i] == max_batch) return cached_l[i];
    const WsLayout L = ws_layout_compute(max_batch);
    cached_b[next] = max_batch; cached_l[next] = L; next = (next + 1) & 3;
    return L;
}

// plain arrays of a `max_batch`-row layout, starting at byte `off`; returns their end
static size_t lay_plain(WsLayout& L, size_t off, int max_batch) {
    const size_t B = (size_t)max_batch;
    size_t part_rows = part1_rows_capacity(max_batch);
    if (part_rows < (size_t)14 * B) part_rows = (size_t)14 * B;      // (v2 persistent kernel: 14 pixel-tile partials)
    L.part1 = off;   off = al256(off + part_rows * HID * sizeof(float));
    L.h1d = off;     off = al256(off + B * HID * sizeof(float));
    L.h2 = off;      off = al256(off + B * HID * sizeof(float));
    L.dz2 = off;     off = al256(off + B * HID * sizeof(float));
    L.dz1 = off;     off = al256(off + B * HID * sizeof(float));
    L.z3 = off;      off = al256(off + B * NCP * sizeof(float));
    L.dz3 = off;     off = al256(off + B * NCP * sizeof(float));
    L.rowloss = off; off = al256(off + B * sizeof(float));
    L.yb = off;      off = al256(off + B);
    // gpart holds the batch-slice partial gradients of grads_kernel (grads_bsplit slices, capped at 32) AND, on the
    // tcgen05 path, dw1_tc_kernel's slices (one per 224 rows: 37 at batch 8192, 74 at 16384): size it for the larger
    // (both counts grow monotonically with the batch, so a workspace sized for max_batch holds every smaller batch)
    int gslices = grads_bsplit(max_batch) > 1 ? grads_bsplit(max_batch) : 0;
    if (max_batch >= 2048 && dw1_tc_slices(max_batch) > gslices) gslices = dw1_tc_slices(max_batch);
    L.gpart = off;   off = al256(off + (size_t)gslices * NP * sizeof(float));
    return off;
}

static size_t lay_ll(WsLayout& L, size_t off, int batch) {
    const size_t B = (size_t)batch;
    L.ll = off;
    L.ll_part = off; off = al256(off + 2 * 14 * B * HID * sizeof(unsigned long long));
    L.ll_act = off;  off = al256(off + 2 * B * LL_ACT_WORDS * sizeof(unsigned long long));
    L.ll_loss = off; off = al256(off + 2 * B * sizeof(unsigned long long));
    L.ll_b1 = off;   off = al256(off + 2 * HID * sizeof(unsigned long long));
    L.ll_bytes = off - L.ll;
    return off;
}

static WsLayout ws_layout_compute(int max_batch) {
    WsLayout L;
    memset(&L, 0, sizeof(L));
    L.ctrl = 0;
    const size_t first = al256(64 * sizeof(uint32_t));
    WsLayout cap;
    const size_t plain_end_512 = lay_plain(cap, first, PERSIST_MAX_BATCH);
    if (max_batch <= PERSIST_MAX_BATCH) {
        L.fend = lay_plain(L, first, max_batch);               // <= plain_end_512: every size grows with the batch
        L.total = lay_ll(L, plain_end_512, max_batch);
    } else {                                                   // the persistent kernel never runs at this batch, but a
        const size_t ll_end = lay_ll(L, plain_end_512, PERSIST_MAX_BATCH);   // smaller one may on the same workspace
        L.fend = lay_plain(L, ll_end, max_batch);
        L.total = L.fend;
    }
    return L;
}

// ------------------------------------------------------------------------------------------
// K1: gather + decode (standalone loader kernel)
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
gather_decode_kernel(const uint8_t* __restrict__ images, const uint8_t* __restrict__ labels,
                     int64_t n_samples, const int64_t* __restrict__ idx, int batch,
                     float* __restrict__ x, uint8_t* __restrict__ y, int* status) {
    __shared__ float lut[256];
    lut[threadIdx.x] = decode_px(threadIdx.x);
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * 8 + warp;
    if (row >= batch) return;
    const int64_t s = idx[row];
    const bool ok = (s >= 0 && s < n_samples);
    if (!ok && lane == 0 && status) atomicExch(status, 1);
    const uint4* src = reinterpret_cast<const uint4*>(images + (ok ? s : 0) * IN);
    float4* dst = reinterpret_cast<float4*>(x + (size_t)row * IN);
    const float qnan = __int_as_float(0x7fc00000);
#pragma unroll
    for (int c = lane; c < NCHUNKS; c += 32) {
        uint4 v = ok ? __ldg(src + c) : make_uint4(0, 0, 0, 0);
        const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float4 o;
            o.x = ok ? lut[w[q] & 0xff] : qnan;
            o.y = ok ? lut[(w[q] >> 8) & 0xff] : qnan;
            o.z = ok ? lut[(w[q] >> 16) & 0xff] : qnan;
            o.w = ok ? lut[w[q] >> 24] : qnan;
            dst[c * 4 + q] = o;
        }
    }
    if (lane == 0 && y) y[row] = ok ? labels[s] : (uint8_t)255;
}

// ------------------------------------------------------------------------------------------
// K7: Philox dropout mask
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
dropout_mask_kernel(uint8_t* __restrict__ mask, int n_blocks4, uint64_t seed, uint64_t step) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_blocks4) return;
    uint32_t r[4];
    Philox::gen(seed, step, (uint32_t)i, r);
    uchar4 o;
    o.x = r[0] < KEEP_THRESHOLD; o.y = r[1] < KEEP_THRESHOLD;
    o.z = r[2] < KEEP_THRESHOLD; o.w = r[3] < KEEP_THRESHOLD;
    reinterpret_cast<uchar4*>(mask)[i] = o;
}

// ------------------------------------------------------------------------------------------
// G1: fc1 forward, split-K partial sums.   part[s][b][n] = sum_{k in split s} X[b][k] * W1[n][k]
// X comes from the resident u8 dataset through idx (U8) or from an fp32 batch.
// ------------------------------------------------------------------------------------------
template <int BM, int BN, int TM, int TN, bool U8>
__global__ void __launch_bounds__(256)
fc1_fwd_kernel(const float* __restrict__ W1, const uint8_t* __restrict__ images, int64_t n_samples,
               const int64_t* __restrict__ idx, const float* __restrict__ x, int B, int ksplit,
               float* __restrict__ part, int* status) {
    constexpr int TX = BN / TN, TY = BM / TM;
    static_assert(TX * TY == 256, "thread layout must cover the tile with 256 threads");
    static_assert(TM == 2 || TM == 4, "TM");
    static_assert(TN == 4 || TN == 8, "TN");
    constexpr int AS = BM + 4, BS = BN + 4;
    __shared__ __align__(16) float As[KCHUNK][AS];
    __shared__ __align__(16) float Bs[KCHUNK][BS];
    __shared__ float lut[256];
    const int tid = threadIdx.x;
    if (U8) lut[tid] = decode_px(tid);
    const int tx = tid % TX, ty = tid / TX;
    const int mt = (B + BM - 1) / BM, nt = HID / BN;
    const int items = mt * nt * ksplit;
    for (int item = blockIdx.x; item < items; item += gridDim.x) {
        const int s = item / (mt * nt);
        const int rem = item - s * (mt * nt);
        const int m0 = (rem / nt) * BM, n0 = (rem % nt) * BN;
        const int c0 = (s * NCHUNKS) / ksplit, c1 = ((s + 1) * NCHUNKS) / ksplit;
        float acc[TM][TN];
#pragma unroll
        for (int i = 0; i < TM; ++i)
#pragma unroll
            for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

        for (int c = c0; c < c1; ++c) {
            __syncthreads();                      // previous tile fully consumed (and lut ready)
            if (U8) {
                if (tid < BM) {
                    const int row = m0 + tid;
                    uint4 v = make_uint4(0, 0, 0, 0);
                    bool ok = false;
                    if (row < B) {
                        const int64_t sidx = idx[row];
                        ok = (sidx >= 0 && sidx < n_samples);
                        if (ok) v = __ldg(reinterpret_cast<const uint4*>(images + sidx * IN) + c);
                        else if (status) atomicExch(status, 1);
                    }
                    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        As[q * 4 + 0][tid] = ok ? lut[w[q] & 0xff] : 0.f;
                        As[q * 4 + 1][tid] = ok ? lut[(w[q] >> 8) & 0xff] : 0.f;
                        As[q * 4 + 2][tid] = ok ? lut[(w[q] >> 16) & 0xff] : 0.f;
                        As[q * 4 + 3][tid] = ok ? lut[w[q] >> 24] : 0.f;
                    }
                }
            } else {
                for (int t = tid; t < BM * 4; t += 256) {
                    const int r = t >> 2, q = t & 3, row = m0 + r;
                    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (row < B) v = __ldg(reinterpret_cast<const float4*>(x + (size_t)row * IN + c * KCHUNK) + q);
                    As[q * 4 + 0][r] = v.x; As[q * 4 + 1][r] = v.y;
                    As[q * 4 + 2][r] = v.z; As[q * 4 + 3][r] = v.w;
                }
            }
            for (int t = tid; t < BN * 4; t += 256) {
                const int n = t >> 2, q = t & 3;
                const float4 v = __ldg(reinterpret_cast<const float4*>(W1 + (size_t)(n0 + n) * IN + c * KCHUNK) + q);
                Bs[q * 4 + 0][n] = v.x; Bs[q * 4 + 1][n] = v.y;
                Bs[q * 4 + 2][n] = v.z; Bs[q * 4 + 3][n] = v.w;
            }
            __syncthreads();
#pragma unroll
            for (int k = 0; k < KCHUNK; ++k) {
                float a[TM], b[TN];
                if (TM == 2) {
                    const float2 t2 = *reinterpret_cast<const float2*>(&As[k][ty * 2]);
                    a[0] = t2.x; a[1] = t2.y;
                } else {
                    const float4 t4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
                    a[0] = t4.x; a[1] = t4.y; a[TM - 2] = t4.z; a[TM - 1] = t4.w;
                }
#pragma unroll
                for (int j4 = 0; j4 < TN / 4; ++j4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(&Bs[k][tx * TN + j4 * 4]);
                    b[j4 * 4 + 0] = t4.x; b[j4 * 4 + 1] = t4.y; b[j4 * 4 + 2] = t4.z; b[j4 * 4 + 3] = t4.w;
                }
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
            }
        }
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            const int row = m0 + ty * TM + i;
            if (row < B) {
                float4* dst = reinterpret_cast<float4*>(part + ((size_t)s * B + row) * HID + n0 + tx * TN);
#pragma unroll
                for (int j4 = 0; j4 < TN / 4; ++j4)
                    dst[j4] = make_float4(acc[i][j4 * 4], acc[i][j4 * 4 + 1], acc[i][j4 * 4 + 2], acc[i][j4 * 4 + 3]);
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// ROWS: per-row chain.  128 threads, thread j owns hidden unit j; R rows per pass.
// ------------------------------------------------------------------------------------------
constexpr int W2S = HID + 1;                       // padded row stride of W2 in shared memory
constexpr unsigned ROWS_BWD_ONLY = 0x100u;         // internal: backward half from external dlogits

struct RowsArgs {
    const float* params;
    const float* part; int ksplit;
    const uint8_t* mask; uint64_t seed, step;
    const uint8_t* y; const uint8_t* labels; const int64_t* idx; int64_t n_samples;
    const float* dlogits;                          // BWD_ONLY input [B][10]
    int B; unsigned flags;
    float* logits; float* loss;
    float* h1d; float* h2; float* dz2; float* dz1; float* z3; float* dz3; float* rowloss;
    uint32_t* ctrl;
};

template <int R> struct RowVec;
template <> struct RowVec<1> { typedef float T; };
template <> struct RowVec<2> { typedef float2 T; };
template <> struct RowVec<4> { typedef float4 T; };

template <int R>
__device__ __forceinline__ void ld_rows(const float* p, float (&v)[R]) {
    typename RowVec<R>::T t = *reinterpret_cast<const typename RowVec<R>::T*>(p);
    const float* f = reinterpret_cast<const float*>(&t);
#pragma unroll
    for (int r = 0; r < R; ++r) v[r] = f[r];
}

template <int R>
__global__ void __launch_bounds__(128)
rows_kernel(RowsArgs a) {
    extern __shared__ __align__(16) float smem[];
    float* w2s = smem;                              // [128][129]
    float* w3s = w2s + HID * W2S;                   // [10][128]
    float* sA = w3s + NC * HID;                     // [128][R]  h1d (k-major)
    float* sB = sA + HID * R;                       // [128][R]  h2
    float* sC = sB + HID * R;                       // [128][R]  dz2
    float* sZ = sC + HID * R;                       // [R][16]   logits
    float* sD = sZ + R * NCP;                       // [R][16]   dlogits
    __shared__ float red[128];
    __shared__ int s_last;

    const int j = threadIdx.x, lane = j & 31, warp = j >> 5;
    const float* W2 = a.params + DDPM_OFF_W2;
    const float* W3 = a.params + DDPM_OFF_W3;
    for (int t = j; t < HID * HID / 4; t += 128) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(W2) + t);
        const int row = t >> 5, col = (t & 31) * 4;
        float* d = w2s + row * W2S + col;
        d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
    }
    for (int t = j; t < NC * HID; t += 128) w3s[t] = __ldg(W3 + t);
    const float b1 = __ldg(a.params + DDPM_OFF_B1 + j), b2 = __ldg(a.params + DDPM_OFF_B2 + j);
    const bool bwd_only = (a.flags & ROWS_BWD_ONLY) != 0;
    const bool prep_bwd = bwd_only || (a.flags & DDPM_FWD_PREP_BACKWARD);
    const bool philox = (a.flags & DDPM_FWD_PHILOX_MASK) != 0;
    // the backward half must use the dropout scale of the forward that filled the workspace
    const bool train = bwd_only ? (__ldcg(a.ctrl + 1) != 0u) : (philox || a.mask != nullptr);
    if (!bwd_only && blockIdx.x == 0 && j == 0) a.ctrl[1] = train ? 1u : 0u;
    const bool have_y = a.y != nullptr || (a.labels != nullptr && a.idx != nullptr);
    const float scale = train ? 1.25f : 1.0f;       // 1/(1-0.2), exact
    const int B = a.B;
    const float invB = 1.0f / (float)B;
    __syncthreads();

    for (int row0 = blockIdx.x * R; row0 < B; row0 += gridDim.x * R) {
        float h1[R];
        bool gate2[R];
        if (!bwd_only) {
            // ---- 1. z1 = sum of split-K partials + b1 ; ReLU ; dropout
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int row = row0 + r;
                float h = 0.f;
                if (row < B) {
                    float z = 0.f;
                    const float* p = a.part + (size_t)row * HID + j;
                    const size_t sstride = (size_t)B * HID;
#pragma unroll 7
                    for (int s = 0; s < a.ksplit; ++s) z += __ldcg(p + s * sstride);
                    z += b1;
                    bool keep = true;
                    if (philox) keep = philox_keep(a.seed, a.step, (uint32_t)(row * HID + j));
                    else if (a.mask) keep = a.mask[(size_t)row * HID + j] != 0;
                    h = (z > 0.f && keep) ? z * scale : 0.f;
                    if (prep_bwd) a.h1d[(size_t)row * HID + j] = h;
                }
                h1[r] = h;
                sA[j * R + r] = h;
            }
            __syncthreads();
            // ---- 2. z2 = h1d . W2^T + b2 ; ReLU
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = b2;
#pragma unroll 8
            for (int k = 0; k < HID; ++k) {
                const float w = w2s[j * W2S + k];
                float hv[R];
                ld_rows<R>(sA + k * R, hv);
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fmaf(hv[r], w, acc[r]);
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int row = row0 + r;
                gate2[r] = acc[r] > 0.f;
                const float h = gate2[r] ? acc[r] : 0.f;
                sB[j * R + r] = h;
                if (prep_bwd && row < B) a.h2[(size_t)row * HID + j] = h;
            }
            __syncthreads();
            // ---- 3. logits z3 = h2 . W3^T (10 x R dot products, one warp each, shuffle reduce)
            for (int p = warp; p < NC * R; p += 4) {
                const int r = p / NC, c = p - r * NC;
                float v = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) v = fmaf(sB[(lane + 32 * i) * R + r], w3s[c * HID + lane + 32 * i], v);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
                if (lane == 0) sZ[r * NCP + c] = v;
            }
            __syncthreads();
            // ---- 4. log-softmax, NLL, dlogits (one thread per row)
            if (j < R) {
                const int row = row0 + j;
                if (row < B) {
                    float z[NC];
                    float m = -INFINITY;
#pragma unroll
                    for (int c = 0; c < NC; ++c) { z[c] = sZ[j * NCP + c]; m = fmaxf(m, z[c]); }
                    if (a.logits) {
#pragma unroll
                        for (int c = 0; c < NC; ++c) a.logits[(size_t)row * NC + c] = z[c];
                    }
                    if (have_y) {
                        int yv;
                        if (a.y) yv = a.y[row];
                        else {
                            const int64_t sidx = a.idx[row];
                            yv = (sidx >= 0 && sidx < a.n_samples) ? a.labels[sidx] : 0;
                        }
                        float e[NC], sum = 0.f;
#pragma unroll
                        for (int c = 0; c < NC; ++c) { e[c] = expf(z[c] - m); sum += e[c]; }
                        const float lse = logf(sum);
                        float zy = 0.f;
#pragma unroll
                        for (int c = 0; c < NC; ++c) zy = (c == yv) ? z[c] : zy;
                        a.rowloss[row] = -((zy - m) - lse);
                        const float inv = 1.0f / sum;
#pragma unroll
                        for (int c = 0; c < NCP; ++c) {
                            float d = 0.f;
                            if (c < NC) d = (e[c < NC ? c : 0] * inv - ((c == yv) ? 1.f : 0.f)) * invB;
                            sD[j * NCP + c] = d;
                            if (prep_bwd) a.dz3[(size_t)row * NCP + c] = d;
                        }
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < NCP; ++c) sD[j * NCP + c] = 0.f;
                }
            }
            __syncthreads();
        } else {
            // backward half only: gates and dlogits come from memory
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int row = row0 + r;
                h1[r] = (row < B) ? a.h1d[(size_t)row * HID + j] : 0.f;
                gate2[r] = (row < B) ? (a.h2[(size_t)row * HID + j] > 0.f) : false;
            }
            if (j < R * NCP) {
                const int r = j / NCP, c = j - r * NCP, row = row0 + r;
                float d = 0.f;
                if (row < B && c < NC) d = a.dlogits[(size_t)row * NC + c];
                sD[r * NCP + c] = d;
                if (row < B) a.dz3[(size_t)row * NCP + c] = d;
            }
            __syncthreads();
        }
        if (prep_bwd && (bwd_only || have_y)) {
            // ---- 5. dz2 = (dz3 . W3) * (z2 > 0)
#pragma unroll
            for (int r = 0; r < R; ++r) {
                float d = 0.f;
#pragma unroll
                for (int c = 0; c < NC; ++c) d = fmaf(sD[r * NCP + c], w3s[c * HID + j], d);
                d = gate2[r] ? d : 0.f;
                sC[j * R + r] = d;
                const int row = row0 + r;
                if (row < B) a.dz2[(size_t)row * HID + j] = d;
            }
            __syncthreads();
            // ---- 6. dz1 = (dz2 . W2) * dropout scale * (z1 > 0 && kept)
            float acc[R];
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = 0.f;
#pragma unroll 8
            for (int i = 0; i < HID; ++i) {
                const float w = w2s[i * W2S + j];
                float dv[R];
                ld_rows<R>(sC + i * R, dv);
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fmaf(dv[r], w, acc[r]);
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const int row = row0 + r;
                if (row < B) a.dz1[(size_t)row * HID + j] = (h1[r] > 0.f) ? acc[r] * scale : 0.f;
            }
        }
        __syncthreads();
    }

    // ---- batch loss: the last block to finish sums the per-row losses in a fixed order
    if (!bwd_only && have_y && a.loss) {
        __threadfence();
        __syncthreads();
        if (j == 0) {
            const unsigned ticket = atomicInc(a.ctrl, gridDim.x - 1);   // wraps back to 0
            s_last = (ticket == gridDim.x - 1);
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            float v = 0.f;
            for (int row = j; row < B; row += 128) v += __ldcg(a.rowloss + row);
            red[j] = v;
            __syncthreads();
            for (int o = 64; o > 0; o >>= 1) {
                if (j < o) red[j] += red[j + o];
                __syncthreads();
            }
            if (j == 0) {
                const float mean = red[0] * invB;
                a.loss[0] = mean;
                a.loss[1] += mean * invB;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------
// ROWS for large batches: the same per-row chain as rows_kernel, organised as small GEMMs.
// 256 threads handle 32 batch rows per pass; thread (rg, cg) owns the 4x4 block rows 4rg.., units 4cg..
// of every [32 x 128] intermediate.  W2 sits in shared memory twice (row-major for dz2.W2, transposed
// for h1d.W2^T) so that both products read one broadcast LDS.128 (4 rows of the k-major activation
// tile) and one conflict-free LDS.128 (4 weights) per 16 FMAs.  Exact fp32, fixed summation order.
// ------------------------------------------------------------------------------------------
constexpr int RG_ROWS = 32;                       // batch rows per pass
constexpr int RG_LD = RG_ROWS + 4;                // k-major activation tiles [128][36]: float4 rows, conflict-free stores

struct RowsGemmSmem {
    float w2[HID * HID];                          // W2[j][k]
    float w2t[HID * HID];                         // W2^T[k][j]
    float w3[NC * HID];
    float b1[HID], b2[HID];
    float aT[HID * RG_LD];                        // h1d^T  [k][row]
    float hT[HID * RG_LD];                        // h2^T   [j][row]
    float dT[HID * RG_LD];                        // dz2^T  [j][row]
    float z3[RG_ROWS * NCP];                      // logits
    float d3[RG_ROWS * NCP];                      // dlogits (zero padded)
    float red[256];
    int last;
};

// W2^T is stored with its 16-byte column chunks XOR-swizzled by (k / 4): the transposing stores of the staging
// loop (32 lanes, same j, k = 4*lane + i) then spread over the banks instead of hitting one (32-way conflict)
__device__ __forceinline__ int w2t_idx(int k, int j) { return k * HID + ((((j >> 2) ^ (k >> 2)) & 31) << 2) + (j & 3); }

template <bool SWZ>
__device__ __forceinline__ void rg_mma(const float* __restrict__ actT, const float* __restrict__ w, int rg, int cg,
                                       float (&acc)[4][4]) {
    // acc[r][c] += sum_k actT[k][4rg + r] * w[k][4cg + c]
#pragma unroll 8
    for (int k = 0; k < HID; ++k) {
        const float4 a4 = *reinterpret_cast<const float4*>(&actT[k * RG_LD + rg * 4]);
        const float4 w4 = *reinterpret_cast<const float4*>(&w[SWZ ? w2t_idx(k, cg * 4) : k * HID + cg * 4]);
        const float av[4] = {a4.x, a4.y, a4.z, a4.w}, wv[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 4; ++c) acc[r][c] = fmaf(av[r], wv[c], acc[r][c]);
    }
}

__global__ void __launch_bounds__(256, 1)
rows_gemm_kernel(RowsArgs a) {
    extern __shared__ __align__(16) unsigned char rg_smem_raw[];
    RowsGemmSmem& s = *reinterpret_cast<RowsGemmSmem*>(rg_smem_raw);
    const int tid = threadIdx.x, rg = tid >> 5, cg = tid & 31;
    const float* W2 = a.params + DDPM_OFF_W2;
    for (int t = tid; t < HID * HID / 4; t += 256) {
        const float4 v = __ldg(reinterpret_cast<const float4*>(W2) + t);
        *reinterpret_cast<float4*>(&s.w2[t * 4]) = v;
        const int j = t >> 5, k = (t & 31) * 4;
        s.w2t[w2t_idx(k + 0, j)] = v.x; s.w2t[w2t_idx(k + 1, j)] = v.y;
        s.w2t[w2t_idx(k + 2, j)] = v.z; s.w2t[w2t_idx(k + 3, j)] = v.w;
    }
    for (int t = tid; t < NC * HID; t += 256) s.w3[t] = __ldg(a.params + DDPM_OFF_W3 + t);
    if (tid < HID) { s.b1[tid] = __ldg(a.params + DDPM_OFF_B1 + tid); s.b2[tid] = __ldg(a.params + DDPM_OFF_B2 + tid); }
    const bool prep_bwd = (a.flags & DDPM_FWD_PREP_BACKWARD) != 0;
    const bool philox = (a.flags & DDPM_FWD_PHILOX_MASK) != 0;
    const bool train = philox || a.mask != nullptr;
    if (blockIdx.x == 0 && tid == 0) a.ctrl[1] = train ? 1u : 0u;
    const bool have_y = a.y != nullptr || (a.labels != nullptr && a.idx != nullptr);
    const float scale = train ? 1.25f : 1.0f;
    const int B = a.B;
    const float invB = 1.0f / (float)B;
    __syncthreads();

    for (int row0 = blockIdx.x * RG_ROWS; row0 < B; row0 += gridDim.x * RG_ROWS) {
        // ---- 1. z1 = sum of split-K partials + b1 ; ReLU ; dropout  -> h1 (registers), aT (smem), h1d (global)
        float h1[4][4];
        {
            const float4 bb = *reinterpret_cast<const float4*>(&s.b1[cg * 4]);
            const float bv[4] = {bb.x, bb.y, bb.z, bb.w};
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int row = row0 + rg * 4 + r;
                float z[4] = {0.f, 0.f, 0.f, 0.f};
                if (row < B) {
                    for (int sp = 0; sp < a.ksplit; ++sp) {
                        const float4 p = __ldcg(reinterpret_cast<const float4*>(a.part + ((size_t)sp * B + row) * HID) + cg);
                        z[0] += p.x; z[1] += p.y; z[2] += p.z; z[3] += p.w;
                    }
                }
                uchar4 mk = make_uchar4(1, 1, 1, 1);
                if (row < B && a.mask && !philox) mk = *reinterpret_cast<const uchar4*>(a.mask + (size_t)row * HID + cg * 4);
                const unsigned char mv[4] = {mk.x, mk.y, mk.z, mk.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const float zz = z[c] + bv[c];
                    bool keep = mv[c] != 0;
                    if (philox) keep = philox_keep(a.seed, a.step, (uint32_t)(row * HID + cg * 4 + c));
                    h1[r][c] = (row < B && zz > 0.f && keep) ? zz * scale : 0.f;
                }
                if (prep_bwd && row < B)
                    *reinterpret_cast<float4*>(a.h1d + (size_t)row * HID + cg * 4) = make_float4(h1[r][0], h1[r][1], h1[r][2], h1[r][3]);
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<float4*>(&s.aT[(cg * 4 + c) * RG_LD + rg * 4]) = make_float4(h1[0][c], h1[1][c], h1[2][c], h1[3][c]);
        }
        __syncthreads();
        // ---- 2. z2 = h1d . W2^T + b2 ; ReLU -> gate2 (registers), hT (smem), h2 (global)
        bool gate2[4][4];
        {
            float acc[4][4];
            const float4 bb = *reinterpret_cast<const float4*>(&s.b2[cg * 4]);
#pragma unroll
            for (int r = 0; r < 4; ++r) { acc[r][0] = bb.x; acc[r][1] = bb.y; acc[r][2] = bb.z; acc[r][3] = bb.w; }
            rg_mma<true>(s.aT, s.w2t, rg, cg, acc);
            float hv[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r) {
#pragma unroll
                for (int c = 0; c < 4; ++c) { gate2[r][c] = acc[r][c] > 0.f; hv[r][c] = gate2[r][c] ? acc[r][c] : 0.f; }
                const int row = row0 + rg * 4 + r;
                if (prep_bwd && row < B)
                    *reinterpret_cast<float4*>(a.h2 + (size_t)row * HID + cg * 4) = make_float4(hv[r][0], hv[r][1], hv[r][2], hv[r][3]);
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<float4*>(&s.hT[(cg * 4 + c) * RG_LD + rg * 4]) = make_float4(hv[0][c], hv[1][c], hv[2][c], hv[3][c]);
        }
        __syncthreads();
        // ---- 3. logits: thread (row = tid & 31, part = tid >> 5) computes classes part and part + 8
        {
            const int r = tid & 31, c0 = tid >> 5, c1 = c0 + 8;
            float v0 = 0.f, v1 = 0.f;
#pragma unroll 8
            for (int k = 0; k < HID; ++k) {
                const float hv = s.hT[k * RG_LD + r];
                v0 = fmaf(hv, s.w3[c0 * HID + k], v0);
                if (c1 < NC) v1 = fmaf(hv, s.w3[c1 * HID + k], v1);
            }
            s.z3[r * NCP + c0] = v0;
            if (c1 < NC) s.z3[r * NCP + c1] = v1;
        }
        __syncthreads();
        // ---- 4. log-softmax, NLL, dlogits: one thread per row
        if (tid < RG_ROWS) {
            const int row = row0 + tid;
            float z[NC], m = -INFINITY;
#pragma unroll
            for (int c = 0; c < NC; ++c) { z[c] = s.z3[tid * NCP + c]; m = fmaxf(m, z[c]); }
            if (row < B && a.logits) {
#pragma unroll
                for (int c = 0; c < NC; ++c) a.logits[(size_t)row * NC + c] = z[c];
            }
            float d[NCP];
#pragma unroll
            for (int c = 0; c < NCP; ++c) d[c] = 0.f;
            if (row < B && have_y) {
                int yv;
                if (a.y) yv = a.y[row];
                else {
                    const int64_t sidx = a.idx[row];
                    yv = (sidx >= 0 && sidx < a.n_samples) ? a.labels[sidx] : 0;
                }
                float e[NC], sum = 0.f;
#pragma unroll
                for (int c = 0; c < NC; ++c) { e[c] = expf(z[c] - m); sum += e[c]; }
                float zy = 0.f;
#pragma unroll
                for (int c = 0; c < NC; ++c) zy = (c == yv) ? z[c] : zy;
                a.rowloss[row] = -((zy - m) - logf(sum));
                const float inv = 1.0f / sum;
#pragma unroll
                for (int c = 0; c < NC; ++c) d[c] = (e[c] * inv - ((c == yv) ? 1.f : 0.f)) * invB;
                if (prep_bwd) {
#pragma unroll
                    for (int c = 0; c < NCP; c += 4)
                        *reinterpret_cast<float4*>(a.dz3 + (size_t)row * NCP + c) = make_float4(d[c], d[c + 1], d[c + 2], d[c + 3]);
                }
            }
#pragma unroll
            for (int c = 0; c < NCP; ++c) s.d3[tid * NCP + c] = d[c];
        }
        __syncthreads();
        if (prep_bwd && have_y) {
            // ---- 5. dz2 = (dz3 . W3) * (z2 > 0) -> dT (smem), dz2 (global)
            float dv[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) dv[r][c] = 0.f;
#pragma unroll
            for (int k = 0; k < NC; ++k) {
                const float4 w4 = *reinterpret_cast<const float4*>(&s.w3[k * HID + cg * 4]);
#pragma unroll
                for (int r = 0; r < 4; ++r) {
                    const float dd = s.d3[(rg * 4 + r) * NCP + k];
                    dv[r][0] = fmaf(dd, w4.x, dv[r][0]); dv[r][1] = fmaf(dd, w4.y, dv[r][1]);
                    dv[r][2] = fmaf(dd, w4.z, dv[r][2]); dv[r][3] = fmaf(dd, w4.w, dv[r][3]);
                }
            }
#pragma unroll
            for (int r = 0; r < 4; ++r) {
#pragma unroll
                for (int c = 0; c < 4; ++c) dv[r][c] = gate2[r][c] ? dv[r][c] : 0.f;
                const int row = row0 + rg * 4 + r;
                if (row < B)
                    *reinterpret_cast<float4*>(a.dz2 + (size_t)row * HID + cg * 4) = make_float4(dv[r][0], dv[r][1], dv[r][2], dv[r][3]);
            }
#pragma unroll
            for (int c = 0; c < 4; ++c)
                *reinterpret_cast<float4*>(&s.dT[(cg * 4 + c) * RG_LD + rg * 4]) = make_float4(dv[0][c], dv[1][c], dv[2][c], dv[3][c]);
            __syncthreads();
            // ---- 6. dz1 = (dz2 . W2) * dropout scale * (z1 > 0 && kept)
            float acc[4][4];
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) acc[r][c] = 0.f;
            rg_mma<false>(s.dT, s.w2, rg, cg, acc);
#pragma unroll
            for (int r = 0; r < 4; ++r) {
                const int row = row0 + rg * 4 + r;
                if (row < B)
                    *reinterpret_cast<float4*>(a.dz1 + (size_t)row * HID + cg * 4) =
                        make_float4(h1[r][0] > 0.f ? acc[r][0] * scale : 0.f, h1[r][1] > 0.f ? acc[r][1] * scale : 0.f,
                                    h1[r][2] > 0.f ? acc[r][2] * scale : 0.f, h1[r][3] > 0.f ? acc[r][3] * scale : 0.f);
            }
        }
        __syncthreads();
    }

    // ---- batch loss: the last block to finish sums the per-row losses in a fixed order
    if (have_y && a.loss) {
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const unsigned ticket = atomicInc(a.ctrl, gridDim.x - 1);
            s.last = (ticket == gridDim.x - 1);
        }
        __syncthreads();
        if (s.last) {
            __threadfence();
            float v = 0.f;
            for (int row = tid; row < B; row += 256) v += __ldcg(a.rowloss + row);
            s.red[tid] = v;
            __syncthreads();
            for (int o = 128; o > 0; o >>= 1) {
                if (tid < o) s.red[tid] += s.red[tid + o];
                __syncthreads();
            }
            if (tid == 0) {
                const float mean = s.red[0] * invB;
                a.loss[0] = mean;
                a.loss[1] += mean * invB;
            }
        }
    }
}

static int launch_rows_gemm(const RowsArgs& a, cudaStream_t st) {
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    const size_t smem = sizeof(RowsGemmSmem);
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(rows_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev] = true;
    }
    int grid = (a.B + RG_ROWS - 1) / RG_ROWS;
    if (grid > 148) grid = 148;
    rows_gemm_kernel<<<grid, 256, smem, st>>>(a);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------------------------------
// GRADS: C[m][n] = scale * sum_b A[b][m] * Bm[b][n]   (reduction over the batch)
// ------------------------------------------------------------------------------------------
struct GradArgs {
    float* params; float* grads;
    const uint8_t* images; int64_t n_samples; const int64_t* idx; const float* x;
    const float* h1d; const float* h2; const float* dz1; const float* dz2; const float* dz3;
    int B; float scale, lr;
    int bsplit_w1, skip_w1;            // dW1 comes from the tensor-core kernel (its own slice count); grads_kernel skips it
    int skip_small;                    // only dW1 (DDPM_BWD_PART_W1): the other items were done by an earlier call
    int bsplit; float* gpart;          // large batches: the batch is cut in bsplit slices, raw partial gradients [bsplit][NP]
};

__device__ __forceinline__ void grad_store4(const GradArgs& a, int off, float4 g) {
    if (a.grads) *reinterpret_cast<float4*>(a.grads + off) = g;
    if (a.lr != 0.f) {
        float4 p = *reinterpret_cast<float4*>(a.params + off);
        p.x = fmaf(-a.lr, g.x, p.x); p.y = fmaf(-a.lr, g.y, p.y);
        p.z = fmaf(-a.lr, g.z, p.z); p.w = fmaf(-a.lr, g.w, p.w);
        *reinterpret_cast<float4*>(a.params + off) = p;
    }
}
__device__ __forceinline__ void grad_store1(const GradArgs& a, int off, float g) {
    if (a.grads) a.grads[off] = g;
    if (a.lr != 0.f) a.params[off] = fmaf(-a.lr, g, a.params[off]);
}

// One output tile.  A: f32 [B][lda] (m contiguous).  Bm: f32 [B][ldb] or u8 rows through idx.
template <int BM, int BN, int TM, int TN, bool U8>
__device__ void grad_tile(const GradArgs& a, const float* __restrict__ A, int lda, int m0, int M,
                          const float* __restrict__ Bf, int ldb, int n0, int c_off, int ldc,
                          float* As, float* Bs, const float* lut, int b_begin, int b_end, int bs) {
    constexpr int TX = BN / TN, TY = BM / TM;
    static_assert(TX * TY <= 256, "tile too large for 256 threads");
    static_assert(TN == 4 || TN == 8, "TN");
    constexpr int AS = BM + 4, BS = BN + 4;
    const int tid = threadIdx.x;
    const bool active = tid < TX * TY;
    const int tx = tid % TX, ty = tid / TX;
    const int B = b_end;                 // rows at or past the end of this batch slice read as zero
    float acc[TM][TN];
#pragma unroll
    for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    for (int k0 = b_begin; k0 < B; k0 += KCHUNK) {
        __syncthreads();
        for (int t = tid; t < KCHUNK * (BM / 4); t += 256) {
            const int k = t / (BM / 4), q = t - k * (BM / 4), b = k0 + k;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (b < B) v = __ldcg(reinterpret_cast<const float4*>(A + (size_t)b * lda + m0) + q);
            *reinterpret_cast<float4*>(As + k * AS + q * 4) = v;
        }
        if (U8) {
            for (int t = tid; t < KCHUNK * (BN / 8); t += 256) {
                const int k = t / (BN / 8), q = t - k * (BN / 8), b = k0 + k;
                uint2 v = make_uint2(0, 0);
                bool ok = false;
                if (b < B) {
                    const int64_t sidx = a.idx[b];
                    ok = (sidx >= 0 && sidx < a.n_samples);
                    if (ok) v = __ldg(reinterpret_cast<const uint2*>(a.images + sidx * IN + n0) + q);
                }
                float4 lo, hi;
                lo.x = ok ? lut[v.x & 0xff] : 0.f;          lo.y = ok ? lut[(v.x >> 8) & 0xff] : 0.f;
                lo.z = ok ? lut[(v.x >> 16) & 0xff] : 0.f;  lo.w = ok ? lut[v.x >> 24] : 0.f;
                hi.x = ok ? lut[v.y & 0xff] : 0.f;          hi.y = ok ? lut[(v.y >> 8) & 0xff] : 0.f;
                hi.z = ok ? lut[(v.y >> 16) & 0xff] : 0.f;  hi.w = ok ? lut[v.y >> 24] : 0.f;
                *reinterpret_cast<float4*>(Bs + k * BS + q * 8) = lo;
                *reinterpret_cast<float4*>(Bs + k * BS + q * 8 + 4) = hi;
            }
        } else {
            for (int t = tid; t < KCHUNK * (BN / 4); t += 256) {
                const int k = t / (BN / 4), q = t - k * (BN / 4), b = k0 + k;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (b < B) v = __ldcg(reinterpret_cast<const float4*>(Bf + (size_t)b * ldb + n0) + q);
                *reinterpret_cast<float4*>(Bs + k * BS + q * 4) = v;
            }
        }
        __syncthreads();
        if (active) {
#pragma unroll
            for (int k = 0; k < KCHUNK; ++k) {
                float av[TM], bv[TN];
                if (TM == 4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(As + k * AS + ty * 4);
                    av[0] = t4.x; av[1] = t4.y; av[TM - 2] = t4.z; av[TM - 1] = t4.w;
                } else {
#pragma unroll
                    for (int i = 0; i < TM; ++i) av[i] = As[k * AS + ty * TM + i];
                }
#pragma unroll
                for (int j4 = 0; j4 < TN / 4; ++j4) {
                    const float4 t4 = *reinterpret_cast<const float4*>(Bs + k * BS + tx * TN + j4 * 4);
                    bv[j4 * 4] = t4.x; bv[j4 * 4 + 1] = t4.y; bv[j4 * 4 + 2] = t4.z; bv[j4 * 4 + 3] = t4.w;
                }
#pragma unroll
                for (int i = 0; i < TM; ++i)
#pragma unroll
                    for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
            }
        }
    }
    if (active) {
#pragma unroll
        for (int i = 0; i < TM; ++i) {
            const int m = m0 + ty * TM + i;
            if (m < M) {
#pragma unroll
                for (int j4 = 0; j4 < TN / 4; ++j4) {
                    const int off = c_off + m * ldc + n0 + tx * TN + j4 * 4;
                    if (a.bsplit > 1) {       // raw partial sum of this batch slice; grads_finalize_kernel combines them
                        *reinterpret_cast<float4*>(a.gpart + (size_t)bs * NP + off) =
                            make_float4(acc[i][j4 * 4], acc[i][j4 * 4 + 1], acc[i][j4 * 4 + 2], acc[i][j4 * 4 + 3]);
                    } else {
                        float4 g = make_float4(acc[i][j4 * 4] * a.scale, acc[i][j4 * 4 + 1] * a.scale,
                                               acc[i][j4 * 4 + 2] * a.scale, acc[i][j4 * 4 + 3] * a.scale);
                        grad_store4(a, off, g);
                    }
                }
            }
        }
    }
}

constexpr int GW1_BM = 32, GW1_BN = 56;           // dW1 tiles: 4 x 14 = 56
constexpr int GW2_BM = 32, GW2_BN = 64;           // dW2 tiles: 4 x 2  = 8
constexpr int N_GW1 = (HID / GW1_BM) * (IN / GW1_BN);
constexpr int N_GW2 = (HID / GW2_BM) * (HID / GW2_BN);
constexpr int GRAD_ITEMS = N_GW1 + N_GW2 + 3;     // + dW3, db1, db2

template <bool U8>
__global__ void __launch_bounds__(256)
grads_kernel(GradArgs a) {
    __shared__ __align__(16) float As[KCHUNK * (64 + 4)];
    __shared__ __align__(16) float Bs[KCHUNK * (HID + 4)];
    __shared__ float lut[256];
    if (U8) lut[threadIdx.x] = decode_px(threadIdx.x);
    const int slice = (a.B + a.bsplit - 1) / a.bsplit;
    for (int work = blockIdx.x; work < GRAD_ITEMS * a.bsplit; work += gridDim.x) {
        const int bs = work / GRAD_ITEMS, item = work - bs * GRAD_ITEMS;
        const int b0 = bs * slice, b1 = min(a.B, b0 + slice);
        if (item >= N_GW1 && a.skip_small) continue;
        if (item < N_GW1) {
            if (a.skip_w1) continue;
            if (a.bsplit > 1) {
                // sliced (large) batches: 14 tiles of 64 x 112 with 4 x 8 register blocks (32 FMAs per 3 LDS.128);
                // items 14..55 of this slice have nothing to do
                if (item < (HID / 64) * (IN / 112)) {
                    const int mt = item / (IN / 112), nt = item - mt * (IN / 112);
                    grad_tile<64, 112, 4, 8, U8>(a, a.dz1, HID, mt * 64, HID, a.x, IN, nt * 112, DDPM_OFF_W1, IN, As, Bs, lut,
                                                 b0, b1, bs);
                }
                continue;
            }
            const int mt = item / (IN / GW1_BN), nt = item - mt * (IN / GW1_BN);
            grad_tile<GW1_BM, GW1_BN, 2, 4, U8>(a, a.dz1, HID, mt * GW1_BM, HID, a.x, IN, nt * GW1_BN,
                                                DDPM_OFF_W1, IN, As, Bs, lut, b0, b1, bs);
        } else if (item < N_GW1 + N_GW2) {
            const int t = item - N_GW1;
            const int mt = t / (HID / GW2_BN), nt = t - mt * (HID / GW2_BN);
            grad_tile<GW2_BM, GW2_BN, 2, 4, false>(a, a.dz2, HID, mt * GW2_BM, HID, a.h1d, HID, nt * GW2_BN,
                                                   DDPM_OFF_W2, HID, As, Bs, lut, b0, b1, bs);
        } else if (item == N_GW1 + N_GW2) {
            grad_tile<16, HID, 1, 8, false>(a, a.dz3, NCP, 0, NC, a.h2, HID, 0, DDPM_OFF_W3, HID, As, Bs, lut, b0, b1, bs);
        } else {
            // bias gradients: column sums over the batch slice, two interleaved halves per column
            const bool first = (item == N_GW1 + N_GW2 + 1);
            const float* D = first ? a.dz1 : a.dz2;
            const int off = first ? DDPM_OFF_B1 : DDPM_OFF_B2;
            const int col = threadIdx.x & 127, half = threadIdx.x >> 7;
            float v = 0.f;
            // eight rows in flight per round, added in row order (one dependent L2 round trip per row made this item the
            // long pole of the kernel: 64 rows x ~0.3 us per batch slice)
            int b = b0 + half;
            for (; b + 14 < b1; b += 16) {
                float t[8];
#pragma unroll
                for (int k = 0; k < 8; ++k) t[k] = __ldcg(D + (size_t)(b + 2 * k) * HID + col);
#pragma unroll
                for (int k = 0; k < 8; ++k) v += t[k];
            }
            for (; b < b1; b += 2) v += __ldcg(D + (size_t)b * HID + col);
            __syncthreads();
            if (half) As[col] = v;
            __syncthreads();
            if (!half) {
                if (a.bsplit > 1) a.gpart[(size_t)bs * NP + off + col] = v + As[col];
                else grad_store1(a, off + col, (v + As[col]) * a.scale);
            }
        }
    }
}

// Combine the batch slices' partial gradients in fixed order, scale, write the bucket and/or apply SGD.
__global__ void __launch_bounds__(256)
grads_finalize_kernel(GradArgs a) {
    // range of the bucket this launch finalises: all of it, only fc1's weights, or everything behind them
    const int lo4 = a.skip_w1 == 2 ? DDPM_OFF_B1 / 4 : 0, n4 = a.skip_small ? DDPM_OFF_B1 / 4 : NP / 4;
    for (int i = lo4 + blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += gridDim.x * blockDim.x) {
        float4 g = __ldcg(reinterpret_cast<const float4*>(a.gpart) + i);
        const int nsl = (i * 4 < DDPM_OFF_B1) ? a.bsplit_w1 : a.bsplit;
        // eight slices in flight per round, added in slice order (one dependent L2 round trip per slice made this
        // kernel 10 us at batch 4096: 32-37 slices x ~0.3 us)
        int s = 1;
        for (; s + 8 <= nsl; s += 8) {
            float4 v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = __ldcg(reinterpret_cast<const float4*>(a.gpart + (size_t)(s + k) * NP) + i);
#pragma unroll
            for (int k = 0; k < 8; ++k) { g.x += v[k].x; g.y += v[k].y; g.z += v[k].z; g.w += v[k].w; }
        }
        for (; s < nsl; ++s) {
            const float4 v = __ldcg(reinterpret_cast<const float4*>(a.gpart + (size_t)s * NP) + i);
            g.x += v.x; g.y += v.y; g.z += v.z; g.w += v.w;
        }
        g.x *= a.scale; g.y *= a.scale; g.z *= a.scale; g.w *= a.scale;
        grad_store4(a, i * 4, g);
    }
}

// ------------------------------------------------------------------------------------------
// K4 alone: cross entropy forward + gradient, one block
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
cross_entropy_kernel(const float* __restrict__ logits, const uint8_t* __restrict__ y, int B,
                     float* loss, float* __restrict__ dlogits) {
    __shared__ float red[256];
    const float invB = 1.0f / (float)B;
    float part = 0.f;
    for (int row = threadIdx.x; row < B; row += 256) {
        float z[NC], m = -INFINITY;
#pragma unroll
        for (int c = 0; c < NC; ++c) { z[c] = logits[(size_t)row * NC + c]; m = fmaxf(m, z[c]); }
        float e[NC], sum = 0.f;
#pragma unroll
        for (int c = 0; c < NC; ++c) { e[c] = expf(z[c] - m); sum += e[c]; }
        const int yv = y[row];
        float zy = 0.f;
#pragma unroll
        for (int c = 0; c < NC; ++c) zy = (c == yv) ? z[c] : zy;
        part += -((zy - m) - logf(sum));
        if (dlogits) {
            const float inv = 1.0f / sum;
#pragma unroll
            for (int c = 0; c < NC; ++c)
                dlogits[(size_t)row * NC + c] = (e[c] * inv - ((c == yv) ? 1.f : 0.f)) * invB;
        }
    }
    red[threadIdx.x] = part;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0 && loss) loss[0] = red[0] * invB;
}

// ------------------------------------------------------------------------------------------
// K6: fused flat SGD
// ------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, int64_t n, float lr, float gscale) {
    const int64_t n4 = n >> 2;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        float4 pv = reinterpret_cast<float4*>(p)[i];
        const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
        pv.x = fmaf(-lr, gv.x * gscale, pv.x); pv.y = fmaf(-lr, gv.y * gscale, pv.y);
        pv.z = fmaf(-lr, gv.z * gscale, pv.z); pv.w = fmaf(-lr, gv.w * gscale, pv.w);
        reinterpret_cast<float4*>(p)[i] = pv;
    }
    for (int64_t i = (n4 << 2) + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
        p[i] = fmaf(-lr, g[i] * gscale, p[i]);
}

// ------------------------------------------------------------------------------------------
// host-side launch helpers
// ------------------------------------------------------------------------------------------
static size_t rows_smem_bytes(int R) {
    return sizeof(float) * (size_t)(HID * W2S + NC * HID + 3 * HID * R + 2 * R * NCP);
}

template <int R>
static int launch_rows(const RowsArgs& a, cudaStream_t st) {
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    const size_t smem = rows_smem_bytes(R);
    if (dev < 64 && !configured[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(rows_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev] = true;
    }
    int grid = (a.B + R - 1) / R;
    if (grid > 296) grid = 296;
    rows_kernel<R><<<grid, 128, smem, st>>>(a);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

static int launch_rows_any(const RowsArgs& a, cudaStream_t st) {
    if (a.B >= 512 && !(a.flags & ROWS_BWD_ONLY)) return launch_rows_gemm(a, st);   // GEMM-style chain, 32 rows per pass
    if (a.B <= 148) return launch_rows<1>(a, st);
    if (a.B <= 592) return launch_rows<2>(a, st);
    return launch_rows<4>(a, st);
}

static int launch_fc1(const float* params, const uint8_t* images, int64_t n_samples, const int64_t* idx,
                      const float* x, int B, int ksplit, float* part, int* status, cudaStream_t st) {
    const float* W1 = params + DDPM_OFF_W1;
    const bool large = B >= G1_LARGE_BATCH;
    const int bm = large ? G1_BM_LARGE : G1_BM_SMALL, bn = large ? G1_BN_LARGE : G1_BN_SMALL;
    int items = ((B + bm - 1) / bm) * (HID / bn) * ksplit;
    int grid = items < 148 * 4 ? items : 148 * 4;
    if (x == nullptr) {
        if (large) fc1_fwd_kernel<G1_BM_LARGE, G1_BN_LARGE, 4, 8, true><<<grid, 256, 0, st>>>(W1, images, n_samples, idx, x, B, ksplit, part, status);
        else       fc1_fwd_kernel<G1_BM_SMALL, G1_BN_SMALL, 2, 4, true><<<grid, 256, 0, st>>>(W1, images, n_samples, idx, x, B, ksplit, part, status);
    } else {
        if (large) fc1_fwd_kernel<G1_BM_LARGE, G1_BN_LARGE, 4, 8, false><<<grid, 256, 0, st>>>(W1, images, n_samples, idx, x, B, ksplit, part, status);
        else       fc1_fwd_kernel<G1_BM_SMALL, G1_BN_SMALL, 2, 4, false><<<grid, 256, 0, st>>>(W1, images, n_samples, idx, x, B, ksplit, part, status);
    }
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace ddpm

// ==========================================================================================
// C ABI
// ==========================================================================================
using namespace ddpm;

extern "C" {

const char* ddpm_last_error(void) { return g_err; }

int ddpm_abi_version(void) { return 6; }

int ddpm_decode_lut_host(float* h_lut256) {
    DDPM_CHECK_ARG(h_lut256 != nullptr, "ddpm_decode_lut_host: null output");
    for (unsigned v = 0; v < 256; ++v) h_lut256[v] = decode_px(v);
    return 0;
}

int ddpm_gather_decode(const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                       const int64_t* d_idx, int batch, float* d_x, uint8_t* d_y,
                       int* d_status, void* stream) {
    DDPM_CHECK_ARG(batch >= 0, "ddpm_gather_decode: negative batch %d", batch);
    DDPM_CHECK_ARG(n_samples > 0, "ddpm_gather_decode: n_samples must be positive");
    if (batch == 0) return 0;
    DDPM_CHECK_ARG(d_images && d_idx && d_x, "ddpm_gather_decode: null pointer");
    DDPM_CHECK_ARG(d_y == nullptr || d_labels != nullptr, "ddpm_gather_decode: d_y given without d_labels");
    DDPM_CHECK_ARG(((uintptr_t)d_images & 15) == 0 && ((uintptr_t)d_x & 15) == 0,
                   "ddpm_gather_decode: d_images and d_x must be 16-byte aligned");
    gather_decode_kernel<<<(batch + 7) / 8, 256, 0, (cudaStream_t)stream>>>(
        d_images, d_labels, n_samples, d_idx, batch, d_x, d_y, d_status);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

int ddpm_dropout_mask_philox(uint8_t* d_mask, int batch, uint64_t seed, uint64_t step, void* stream) {
    DDPM_CHECK_ARG(batch >= 0, "ddpm_dropout_mask_philox: negative batch");
    if (batch == 0) return 0;
    DDPM_CHECK_ARG(d_mask != nullptr, "ddpm_dropout_mask_philox: null mask");
    const int n4 = batch * HID / 4;
    dropout_mask_kernel<<<(n4 + 255) / 256, 256, 0, (cudaStream_t)stream>>>(d_mask, n4, seed, step);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

size_t ddpm_workspace_bytes(int max_batch) {
    if (max_batch <= 0) return 0;
    return ws_layout(max_batch).total;
}

int ddpm_workspace_peek(const void* d_workspace, int batch, int which, float* d_out, void* stream) {
    DDPM_CHECK_ARG(d_workspace && d_out && batch > 0, "ddpm_workspace_peek: bad arguments");
    const WsLayout L = ws_layout(batch);
    const size_t off[5] = {L.h1d, L.h2, L.dz2, L.dz1, L.dz3};
    DDPM_CHECK_ARG(which >= 0 && which < 5, "ddpm_workspace_peek: which must be 0..4 (h1d, h2, dz2, dz1, dz3)");
    const size_t n = (size_t)batch * (which < 4 ? HID : NCP) * sizeof(float);
    DDPM_CUDA(cudaMemcpyAsync(d_out, (const char*)d_workspace + off[which], n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}

size_t ddpm_workspace_forward_bytes(int max_batch) {
    if (max_batch <= 0) return 0;
    return ws_layout(max_batch).fend;
}

int ddpm_workspace_init(void* d_workspace, int max_batch, void* stream) {
    DDPM_CHECK_ARG(d_workspace != nullptr && max_batch > 0, "ddpm_workspace_init: bad arguments");
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0, "ddpm_workspace_init: workspace must be 256-byte aligned");
    DDPM_CUDA(cudaMemsetAsync(d_workspace, 0, 64 * sizeof(uint32_t), (cudaStream_t)stream));
    // the persistent kernel's dataflow words carry step tags that only ever grow on one workspace; memory handed
    // out by a caching allocator may hold the tagged words of an earlier workspace: start from tag 0 everywhere
    const WsLayout L = ws_layout(max_batch);
    DDPM_CUDA(cudaMemsetAsync((char*)d_workspace + L.ll, 0, L.ll_bytes, (cudaStream_t)stream));
    return 0;
}

int ddpm_step_forward(const float* d_params,
                      const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                      const int64_t* d_idx, const float* d_x, const uint8_t* d_y,
                      const uint8_t* d_mask, uint64_t philox_seed, uint64_t philox_step,
                      int batch, unsigned flags,
                      float* d_logits, float* d_loss, void* d_workspace, void* stream) {
    DDPM_CHECK_ARG(batch > 0, "ddpm_step_forward: batch must be positive (got %d)", batch);
    DDPM_CHECK_ARG(d_params && d_workspace, "ddpm_step_forward: null params/workspace");
    DDPM_CHECK_ARG(d_x != nullptr || (d_images != nullptr && d_idx != nullptr && n_samples > 0),
                   "ddpm_step_forward: need either d_x or (d_images, d_idx, n_samples)");
    DDPM_CHECK_ARG((flags & ~(DDPM_FWD_PREP_BACKWARD | DDPM_FWD_PHILOX_MASK | DDPM_FWD_TF32)) == 0, "ddpm_step_forward: unknown flags %#x", flags);
    DDPM_CHECK_ARG(!(flags & DDPM_FWD_TF32) || d_x == nullptr,
                   "ddpm_step_forward: DDPM_FWD_TF32 needs the resident uint8 dataset path (d_x must be NULL)");
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0 && ((uintptr_t)d_params & 15) == 0,
                   "ddpm_step_forward: workspace must be 256-byte and params 16-byte aligned");
    DDPM_CHECK_ARG(d_x == nullptr || ((uintptr_t)d_x & 15) == 0, "ddpm_step_forward: d_x must be 16-byte aligned");
    DDPM_CHECK_ARG(d_x != nullptr || ((uintptr_t)d_images & 15) == 0, "ddpm_step_forward: d_images must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_workspace;
    const WsLayout L = ws_layout(batch);       // offsets depend only on the batch of THIS call
    const bool tf32 = (flags & DDPM_FWD_TF32) != 0;
    const int ks = tf32 ? fc1_tc_ksplit() : fc1_ksplit(batch);
    float* part = (float*)(ws + L.part1);
    if (tf32) {
        if (launch_fc1_tc(d_params + DDPM_OFF_W1, d_images, n_samples, d_idx, batch, part, st)) return 1;
    } else if (launch_fc1(d_params, d_images, n_samples, d_idx, d_x, batch, ks, part, nullptr, st)) return 1;
    RowsArgs a;
    memset(&a, 0, sizeof(a));
    a.params = d_params; a.part = part; a.ksplit = ks;
    a.mask = (flags & DDPM_FWD_PHILOX_MASK) ? nullptr : d_mask;
    a.seed = philox_seed; a.step = philox_step;
    a.y = d_y;
    a.labels = (d_x == nullptr) ? d_labels : nullptr;
    a.idx = (d_x == nullptr) ? d_idx : nullptr;
    a.n_samples = n_samples;
    a.B = batch; a.flags = flags & ~DDPM_FWD_TF32;
    a.logits = d_logits; a.loss = d_loss;
    a.h1d = (float*)(ws + L.h1d); a.h2 = (float*)(ws + L.h2);
    a.dz2 = (float*)(ws + L.dz2); a.dz1 = (float*)(ws + L.dz1);
    a.z3 = (float*)(ws + L.z3); a.dz3 = (float*)(ws + L.dz3);
    a.rowloss = (float*)(ws + L.rowloss);
    a.ctrl = (uint32_t*)(ws + L.ctrl);
    return launch_rows_any(a, st);
}

int ddpm_step_backward(float* d_params,
                       const uint8_t* d_images, int64_t n_samples,
                       const int64_t* d_idx, const float* d_x,
                       const float* d_dlogits, int batch, float grad_scale, float lr, unsigned flags,
                       float* d_grads, void* d_workspace, void* stream) {
    DDPM_CHECK_ARG(batch > 0, "ddpm_step_backward: batch must be positive (got %d)", batch);
    DDPM_CHECK_ARG((flags & ~(DDPM_BWD_TF32 | DDPM_BWD_PART_SMALL | DDPM_BWD_PART_W1)) == 0, "ddpm_step_backward: unknown flags %#x", flags);
    const bool part_small = (flags & DDPM_BWD_PART_SMALL) != 0, part_w1 = (flags & DDPM_BWD_PART_W1) != 0;
    DDPM_CHECK_ARG(!(part_small && part_w1), "ddpm_step_backward: DDPM_BWD_PART_SMALL and DDPM_BWD_PART_W1 are two separate calls");
    DDPM_CHECK_ARG(!(part_small || part_w1) || (d_dlogits == nullptr && grads_bsplit(batch) > 1),
                   "ddpm_step_backward: the two-part backward is for batches >= 1024 of the fused-loss path");
    DDPM_CHECK_ARG(d_params && d_workspace, "ddpm_step_backward: null params/workspace");
    DDPM_CHECK_ARG(d_grads != nullptr || lr != 0.f, "ddpm_step_backward: nothing to do (no d_grads and lr == 0)");
    DDPM_CHECK_ARG(d_x != nullptr || (d_images != nullptr && d_idx != nullptr && n_samples > 0),
                   "ddpm_step_backward: need either d_x or (d_images, d_idx, n_samples)");
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0 && ((uintptr_t)d_params & 15) == 0 &&
                   ((uintptr_t)d_grads & 15) == 0, "ddpm_step_backward: misaligned pointer");
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_workspace;
    const WsLayout L = ws_layout(batch);
    if (d_dlogits != nullptr) {
        RowsArgs r;
        memset(&r, 0, sizeof(r));
        r.params = d_params; r.B = batch; r.flags = ROWS_BWD_ONLY; r.dlogits = d_dlogits;
        r.h1d = (float*)(ws + L.h1d); r.h2 = (float*)(ws + L.h2);
        r.dz2 = (float*)(ws + L.dz2); r.dz1 = (float*)(ws + L.dz1);
        r.dz3 = (float*)(ws + L.dz3); r.ctrl = (uint32_t*)(ws + L.ctrl);
        if (launch_rows_any(r, st)) return 1;
    }
    GradArgs g;
    memset(&g, 0, sizeof(g));
    g.params = d_params; g.grads = d_grads;
    g.images = d_images; g.n_samples = n_samples; g.idx = d_idx; g.x = d_x;
    g.h1d = (const float*)(ws + L.h1d); g.h2 = (const float*)(ws + L.h2);
    g.dz1 = (const float*)(ws + L.dz1); g.dz2 = (const float*)(ws + L.dz2);
    g.dz3 = (const float*)(ws + L.dz3);
    g.B = batch; g.scale = grad_scale; g.lr = lr;
    g.bsplit = grads_bsplit(batch); g.gpart = (float*)(ws + L.gpart);
    g.bsplit_w1 = g.bsplit; g.skip_w1 = 0;
    g.skip_small = part_w1 ? 1 : 0;
    if (part_small) g.skip_w1 = 2;                       // 2: skipped AND outside the range this call finalises
    else if ((flags & DDPM_BWD_TF32) && d_x == nullptr && batch >= 2048) {
        // fc1's weight gradient (85 % of the flops of the backward pass) on the tensor cores.  Below 2048 rows the
        // 7 x slices grid leaves most SMs idle and the FFMA tiles are faster (measured: 46 vs 37 us at 1024).
        g.bsplit_w1 = dw1_tc_slices(batch); g.skip_w1 = 1;
        DDPM_CHECK_ARG(L.gpart + (size_t)g.bsplit_w1 * NP * sizeof(float) <= L.total,
                       "ddpm_step_backward: batch %d needs %d gradient slices, the workspace layout holds fewer", batch, g.bsplit_w1);
        if (launch_dw1_tc(g.dz1, d_images, n_samples, d_idx, batch, g.gpart, st)) return 1;
    }
    if (!(g.skip_small && g.skip_w1)) {                  // (dW1 alone from the tensor-core kernel: nothing left for this one)
        if (d_x == nullptr) grads_kernel<true><<<GRAD_ITEMS * g.bsplit, 256, 0, st>>>(g);
        else                grads_kernel<false><<<GRAD_ITEMS * g.bsplit, 256, 0, st>>>(g);
        DDPM_CUDA(cudaGetLastError());
    }
    if (g.bsplit > 1) {
        grads_finalize_kernel<<<116, 256, 0, st>>>(g);
        DDPM_CUDA(cudaGetLastError());
    }
    return 0;
}

int ddpm_cross_entropy(const float* d_logits, const uint8_t* d_y, int batch,
                       float* d_loss, float* d_dlogits, void* stream) {
    DDPM_CHECK_ARG(batch > 0, "ddpm_cross_entropy: batch must be positive (got %d)", batch);
    DDPM_CHECK_ARG(d_logits && d_y, "ddpm_cross_entropy: null pointer");
    cross_entropy_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(d_logits, d_y, batch, d_loss, d_dlogits);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

int ddpm_sgd_step(float* d_params, const float* d_grads, int64_t n, float lr, float grad_scale,
                  void* stream) {
    DDPM_CHECK_ARG(n >= 0, "ddpm_sgd_step: negative n");
    if (n == 0) return 0;
    DDPM_CHECK_ARG(d_params && d_grads, "ddpm_sgd_step: null pointer");
    DDPM_CHECK_ARG(((uintptr_t)d_params & 15) == 0 && ((uintptr_t)d_grads & 15) == 0,
                   "ddpm_sgd_step: pointers must be 16-byte aligned");
    int64_t blocks = ((n >> 2) + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 8) blocks = 148 * 8;
    sgd_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(d_params, d_grads, n, lr, grad_scale);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

}  // extern "C"

// ==========================================================================================
// Forward/eval entry point that owns its scratch (no allocation anywhere on the call path)
// ==========================================================================================
// Some hosts hand a plug-in its input and output buffers and nothing else: a validation hook
// inside another framework's data pipeline whose memory pool is closed to plug-ins, a C harness
// that may not call cudaMalloc after start-up because cudaMalloc synchronises the whole device,
// a process running under a fixed device-memory budget.  Such a caller cannot provide the
// workspace ddpm_step_forward needs.  For them the scratch lives here, in a fixed-size device
// array that every GPU loading this library carries until the process exits.
//
// g_static_scratch is a HOST-side handle in host code, not a device address.  The address is
// looked up once per device with cudaGetSymbolAddress while that device is current, and that
// pointer is what the kernels receive.
namespace ddpm {

constexpr size_t STATIC_SCRATCH_BYTES = 3490816;   // the forward's arrays of a DDPM_STATIC_SCRATCH_MAX_BATCH-row workspace layout
static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

static void* g_static_scratch_ptr[64] = {};        // per device, filled on first use

static int static_scratch_for_current_device(void** out) {
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    DDPM_CHECK_ARG(dev >= 0 && dev < 64, "static scratch: device ordinal %d not supported", dev);
    void* p = __atomic_load_n(&g_static_scratch_ptr[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        DDPM_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        DDPM_CHECK_ARG(p != nullptr && ((uintptr_t)p & 255) == 0, "static scratch: bad symbol address");
        __atomic_store_n(&g_static_scratch_ptr[dev], p, __ATOMIC_RELEASE);
    }
    *out = p;
    return 0;
}

}  // namespace ddpm

extern "C" int ddpm_step_forward_static_scratch(
        const float* d_params,
        const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
        const int64_t* d_idx, const float* d_x, const uint8_t* d_y,
        const uint8_t* d_mask, uint64_t philox_seed, uint64_t philox_step,
        int batch, unsigned flags, float* d_logits, float* d_loss, void* stream) {
    DDPM_CHECK_ARG(batch > 0, "ddpm_step_forward_static_scratch: batch must be positive (got %d)", batch);
    DDPM_CHECK_ARG(batch <= DDPM_STATIC_SCRATCH_MAX_BATCH,
                   "ddpm_step_forward_static_scratch: batch %d exceeds the built-in scratch limit of %d rows; "
                   "use ddpm_step_forward with a caller-provided workspace", batch, DDPM_STATIC_SCRATCH_MAX_BATCH);
    DDPM_CHECK_ARG((flags & DDPM_FWD_PREP_BACKWARD) == 0,
                   "ddpm_step_forward_static_scratch: forward/eval only (DDPM_FWD_PREP_BACKWARD needs a "
                   "caller-owned workspace that ddpm_step_backward can read)");
    DDPM_CHECK_ARG(ws_layout(batch).fend <= STATIC_SCRATCH_BYTES,
                   "ddpm_step_forward_static_scratch: internal scratch too small for batch %d", batch);
    void* scratch = nullptr;
    if (static_scratch_for_current_device(&scratch)) return 1;
    // The array keeps whatever the previous call left.  Only the control words must start at
    // zero (the ticket counter that elects the block summing the per-row losses); everything
    // else is written before it is read.
    DDPM_CUDA(cudaMemsetAsync(scratch, 0, 64 * sizeof(uint32_t), (cudaStream_t)stream));
    return ddpm_step_forward(d_params, d_images, d_labels, n_samples, d_idx, d_x, d_y, d_mask,
                             philox_seed, philox_step, batch, flags, d_logits, d_loss, scratch, stream);
}
