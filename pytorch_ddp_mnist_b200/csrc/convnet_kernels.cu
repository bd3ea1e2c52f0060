// BASELINE.json's "ConvNet"  behind the same boundary as the reference's MLP:
//   conv 1->32 3x3 + ReLU | conv 32->64 3x3 + ReLU + maxpool 2 + dropout .25 + flatten |
//   fc 9216->128 + ReLU + dropout .5 | fc 128->10 | log_softmax + nll_loss | backward | Adadelta
// The model is NOT in the reference: the checker is oracle/convnet_oracle.py, a restatement of
// BASELINE.json's layer list - "parity unpinned".  Same input path as the MLP: resident u8 dataset, int64
// sample indices, decode through the 256-entry table.
//
// fp32 FFMA throughout, fixed summation order (no atomics on data).  One CTA works on one image (x one half of
// the channels) with the image's activations in shared memory:
//   cn_conv1_kernel      x (gathered, decoded) -> a1 [B][32][26x26]                               (J1)
//   cn_conv2_pool_kernel a1 -> conv2 + bias + ReLU + 2x2 max + dropout -> p2d [B][9216], flags    (J2)
//   cn_fc1_fwd_kernel    split-K partials of p2d . Wfc1^T                                         (J3)
//   cn_head_kernel       partial sums + bias + ReLU + dropout, fc2, log_softmax, nll, dlogits, dz(fc1)
//   cn_fc1_bwd_kernel    dWfc1 tile + gradient of the pooled activations                          (J4)
//   cn_conv2_dgrad_kernel / cn_conv2_wgrad_kernel / cn_conv1_bwd_kernel, cn_reduce_kernel
//   adadelta_kernel      flat fused Adadelta                                                      (J5)
#include "convnet.cuh"

namespace ddpm {

// ------------------------------------------------------------------------------------------ J1: conv1 + ReLU
// grid = (B, 3), 256 threads.  The 28x28 image is gathered through idx and decoded through the table into shared
// memory; a thread owns an output pixel: its 3x3 input window sits in registers and the 32 channels' weights are warp-
// broadcast from shared memory.  Output: planar [B][32][676] and, for the tensor-core conv2 (TC), also pixel-major
// [B][676][32] with every value rounded to TF32 (the planar copy carries the same rounded values, so the backward
// pass differentiates the function the forward pass computed).
template <bool TC>
__global__ void __launch_bounds__(256)
cn_conv1_kernel(const float* __restrict__ params, const uint8_t* __restrict__ images, int64_t n_samples,
                const int64_t* __restrict__ idx, float* __restrict__ a1, float* __restrict__ a1n,
                float* __restrict__ w2t, float* __restrict__ w2d) {
    if (TC && blockIdx.y == 3) {
        // fourth layer of CTAs (tensor-core mode): conv2.weight [co][ci][tap] -> TF32 copies w2t [tap][co][ci] (forward,
        // K = ci) and w2d [tap][ci][co] (data gradient, K = co) - same launch, one dependent kernel less
        for (int t = blockIdx.x * 256 + threadIdx.x; t < C2 * C1 * 9; t += gridDim.x * 256) {
            const int co = t / (C1 * 9), r = t - co * (C1 * 9), ci = r / 9, tap = r - ci * 9;
            const float v = tf32_round(params[OFF_C2W + t]);
            w2t[(tap * C2 + co) * C1 + ci] = v;
            w2d[(tap * C1 + ci) * C2 + co] = v;
        }
        return;
    }
    __shared__ float xs[IN];
    __shared__ __align__(16) float ws[9 * C1 + C1];          // [tap][32 channels], then bias[32]
    __shared__ float lut[256];
    const int tid = threadIdx.x, b = blockIdx.x;
    lut[tid] = decode_px(tid);
    for (int t = tid; t < C1 * 9; t += 256) ws[(t % 9) * C1 + t / 9] = params[OFF_C1W + t];     // conv1.weight is [c][tap]
    if (tid < C1) ws[9 * C1 + tid] = params[OFF_C1B + tid];
    __syncthreads();
    const int64_t s = idx[b];
    const bool ok = s >= 0 && s < n_samples;
    for (int t = tid; t < IN / 4; t += 256) {
        const uint32_t v = ok ? __ldg(reinterpret_cast<const uint32_t*>(images + s * IN) + t) : 0u;
        xs[t * 4 + 0] = ok ? lut[v & 0xff] : 0.f;          xs[t * 4 + 1] = ok ? lut[(v >> 8) & 0xff] : 0.f;
        xs[t * 4 + 2] = ok ? lut[(v >> 16) & 0xff] : 0.f;  xs[t * 4 + 3] = ok ? lut[v >> 24] : 0.f;
    }
    __syncthreads();
    float* outp = a1 + (size_t)b * C1 * A1;
    const int pos = blockIdx.y * 256 + tid;                  // grid.y = 3 covers the 676 pixels
    if (pos < A1) {
        const int y = pos / H1, x = pos - y * H1;
        float win[9];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky)
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) win[ky * 3 + kx] = xs[(y + ky) * 28 + x + kx];
#pragma unroll 2
        for (int c4 = 0; c4 < C1; c4 += 4) {
            float4 acc = *reinterpret_cast<const float4*>(&ws[9 * C1 + c4]);
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const float4 w = *reinterpret_cast<const float4*>(&ws[t * C1 + c4]);
                acc.x = fmaf(win[t], w.x, acc.x); acc.y = fmaf(win[t], w.y, acc.y);
                acc.z = fmaf(win[t], w.z, acc.z); acc.w = fmaf(win[t], w.w, acc.w);
            }
            acc.x = fmaxf(acc.x, 0.f); acc.y = fmaxf(acc.y, 0.f); acc.z = fmaxf(acc.z, 0.f); acc.w = fmaxf(acc.w, 0.f);
            if (TC) {
                acc.x = tf32_round(acc.x); acc.y = tf32_round(acc.y); acc.z = tf32_round(acc.z); acc.w = tf32_round(acc.w);
                *reinterpret_cast<float4*>(a1n + ((size_t)b * A1 + pos) * C1 + c4) = acc;
            }
            outp[(c4 + 0) * A1 + pos] = acc.x; outp[(c4 + 1) * A1 + pos] = acc.y;
            outp[(c4 + 2) * A1 + pos] = acc.z; outp[(c4 + 3) * A1 + pos] = acc.w;
        }
    }
}

// ------------------------------------------------------------------------------------------ J2: conv2 + ReLU + pool + dropout
// grid = (B, 2): CTA (b, h) computes output channels [32h, 32h+32) of image b.  288 threads: thread -> (pooled
// position p, group parity); each round it owns one 2x2 pooling window x 8 output channels = 32 accumulators, fed
// from a 4x4 input patch per input channel (8 LDS.64) and 9 x 8 weights (18 LDS.128, warp-broadcast).
constexpr int C2_THREADS = 288;
constexpr size_t C2_SMEM = (size_t)(C1 * A1S + 288 * 32 + 32) * 4;         // a1 image + weights [ci*9+tap][32 co] + bias
__global__ void __launch_bounds__(C2_THREADS, 1)
cn_conv2_pool_kernel(const float* __restrict__ params, const float* __restrict__ a1, const uint8_t* __restrict__ mask1,
                     uint64_t seed, uint64_t step, int use_philox, int train,
                     float* __restrict__ p2d, uint8_t* __restrict__ pflag) {
    extern __shared__ __align__(16) float sm[];
    float* a1s = sm;                          // [32][677]
    float* w2s = a1s + C1 * A1S;              // [288][32]
    float* b2s = w2s + 288 * 32;
    const int tid = threadIdx.x, b = blockIdx.x, half = blockIdx.y;
    const float* src = a1 + (size_t)b * C1 * A1;
#pragma unroll 8
    for (int t = tid; t < C1 * A1; t += C2_THREADS) {
        const int c = t / A1, pos = t - c * A1;
        a1s[c * A1S + pos] = src[t];
    }
    // conv2.weight is [co][ci][3][3]: element (co, r) with r = ci*9+tap sits at co*288 + r; transposed here to [r][co]
    // (groups of 4 channels XOR-swizzled by r so that this transposing store is 4-way, not 32-way, bank-conflicted)
#pragma unroll 8
    for (int t = tid; t < 288 * 32; t += C2_THREADS) {
        const int co = t / 288, r = t - co * 288;
        w2s[r * 32 + (((co >> 2) ^ (r & 7)) << 2) + (co & 3)] = params[OFF_C2W + (size_t)(half * 32 + co) * 288 + r];
    }
    if (tid < 32) b2s[tid] = params[OFF_C2B + half * 32 + tid];
    __syncthreads();
    const int p = tid % NPOOL, gh = tid / NPOOL;            // gh in {0, 1}
    const int py = p / HP, px = p - py * HP;
    for (int round = 0; round < 2; ++round) {
        const int co0 = (gh + 2 * round) * 8;               // 8 output channels of this half
        float acc[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[q][j] = 0.f;
        for (int ci = 0; ci < C1; ++ci) {
            float patch[4][4];
            const float* base = a1s + ci * A1S + (2 * py) * H1 + 2 * px;
#pragma unroll
            for (int r = 0; r < 4; ++r)
#pragma unroll
                for (int c = 0; c < 4; ++c) patch[r][c] = base[r * H1 + c];
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const int r = ci * 9 + ky * 3 + kx;
                    const float4 w0 = *reinterpret_cast<const float4*>(&w2s[r * 32 + ((((co0 >> 2)) ^ (r & 7)) << 2)]);
                    const float4 w1 = *reinterpret_cast<const float4*>(&w2s[r * 32 + ((((co0 >> 2) + 1) ^ (r & 7)) << 2)]);
                    const float w[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                        for (int dx = 0; dx < 2; ++dx)
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                acc[dy * 2 + dx][j] = fmaf(patch[dy + ky][dx + kx], w[j], acc[dy * 2 + dx][j]);
                }
        }
        const unsigned keep8 = (train && use_philox) ? philox_keep8(seed, step, b, p, half * 32 + co0, KEEP1_THR) : 0xffu;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int co = half * 32 + co0 + j;
            const float bias = b2s[co0 + j];
            float m = acc[0][j] + bias;
            unsigned arg = 0;
#pragma unroll
            for (int q = 1; q < 4; ++q) {
                const float v = acc[q][j] + bias;
                if (v > m) { m = v; arg = q; }              // first maximum wins, like max_pool2d's window scan
            }
            const size_t e = (size_t)b * FEAT + (size_t)co * NPOOL + p;
            bool keep = true;
            if (train) keep = use_philox ? ((keep8 >> j) & 1u) != 0 : (mask1[e] != 0);
            const bool pos = m > 0.f;
            p2d[e] = (pos && keep) ? (train ? m * SCALE1 : m) : 0.f;
            pflag[e] = (uint8_t)(arg | ((pos && keep) ? PF_PASS : 0u));
        }
    }
}

// ------------------------------------------------------------------------------------------ J3: fc1 forward, split-K
// grid = 72 (slices of 128 features), 256 threads.  part[ks][b][n] = sum_{k in slice} p2d[b][k] . W[n][k].
// W slice transposed into Ws[k][n] (stride 129), rows in chunks of 64: thread tile 4 rows x 8 units.
constexpr int F1_LD = 129;
constexpr size_t F1_SMEM = (size_t)(128 * F1_LD + 64 * F1_LD) * 4;
__global__ void __launch_bounds__(256, 1)
cn_fc1_fwd_kernel(const float* __restrict__ params, const float* __restrict__ p2d, int B, float* __restrict__ part) {
    extern __shared__ __align__(16) float sm[];
    float* Ws = sm;                           // [k][n] stride 129
    float* Xs = Ws + 128 * F1_LD;             // [row][k] stride 129
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, ks = blockIdx.x, k0 = ks * 128;
#pragma unroll 16
    for (int seg = warp; seg < 128 * 4; seg += 8) {          // (unit n, 32 consecutive k): coalesced read, conflict-free write
        const int n = seg >> 2, kk = (seg & 3) * 32 + lane;
        Ws[kk * F1_LD + n] = params[OFF_F1W + (size_t)n * FEAT + k0 + kk];
    }
    const int tn = tid & 15, rg = tid >> 4;
    for (int r0 = 0; r0 < B; r0 += 64) {
        __syncthreads();
#pragma unroll 16
        for (int seg = warp; seg < 64 * 4; seg += 8) {
            const int r = seg >> 2, kk = (seg & 3) * 32 + lane;
            Xs[r * F1_LD + kk] = (r0 + r < B) ? p2d[(size_t)(r0 + r) * FEAT + k0 + kk] : 0.f;
        }
        __syncthreads();
        float acc[4][8];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
#pragma unroll 4
        for (int k = 0; k < 128; ++k) {
            float x[4], w[8];
#pragma unroll
            for (int i = 0; i < 4; ++i) x[i] = Xs[(rg * 4 + i) * F1_LD + k];
#pragma unroll
            for (int j = 0; j < 8; ++j) w[j] = Ws[k * F1_LD + tn + 16 * j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(x[i], w[j], acc[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = r0 + rg * 4 + i;
            if (row < B)
#pragma unroll
                for (int j = 0; j < 8; ++j) part[((size_t)ks * B + row) * FC + tn + 16 * j] = acc[i][j];
        }
    }
}

// ------------------------------------------------------------------------------------------ head: fc1 epilogue .. dz(fc1)
// grid = B (one batch row per CTA), 128 threads (thread = fc1 unit).  Sum of the 72 split-K partials in slice order
// + bias, ReLU, dropout(.5); fc2 + bias; log_softmax; nll (mean over the batch); dlogits = (softmax - onehot)/B;
// gradient of fc1's pre-activation.  The last CTA to finish sums the per-row losses in row order.
__global__ void __launch_bounds__(FC)
cn_head_kernel(const float* __restrict__ params, const float* __restrict__ part, const uint8_t* __restrict__ labels,
               int64_t n_samples, const int64_t* __restrict__ idx, const uint8_t* __restrict__ mask2,
               uint64_t seed, uint64_t step, int use_philox, int train, int prep_backward, int B,
               float* __restrict__ hd, float* __restrict__ dzf, float* __restrict__ dlog, float* __restrict__ logprobs,
               float* __restrict__ rowloss, uint8_t* __restrict__ yb, float* __restrict__ loss2, uint32_t* __restrict__ ctrl) {
    __shared__ float hs[FC];
    __shared__ float zs[NCP], dzs[NCP];
    __shared__ int last;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, b = blockIdx.x;
    float z = 0.f;
#pragma unroll 24
    for (int ks = 0; ks < KSPLIT; ++ks) z += part[((size_t)ks * B + b) * FC + tid];
    z += params[OFF_F1B + tid];
    const bool pos = z > 0.f;
    bool keep = true;
    if (train) keep = use_philox ? philox_keep_thr(seed, step + (1ull << 62), (uint32_t)(b * FC + tid), KEEP2_THR)
                                 : (mask2[(size_t)b * FC + tid] != 0);
    const float h = (pos && keep) ? (train ? z * SCALE2 : z) : 0.f;
    hs[tid] = h;
    if (prep_backward) hd[(size_t)b * FC + tid] = h;
    __syncthreads();
    for (int c = warp; c < NC; c += FC / 32) {             // fc2: one warp per class
        float v = 0.f;
#pragma unroll
        for (int k = 0; k < FC / 32; ++k) v = fmaf(hs[lane + 32 * k], params[OFF_F2W + c * FC + lane + 32 * k], v);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) zs[c] = v + params[OFF_F2B + c];
    }
    __syncthreads();
    if (warp == 0) {
        const int64_t s = idx[b];
        const int y = (s >= 0 && s < n_samples) ? labels[s] : 0;
        const float v = lane < NC ? zs[lane] : -INFINITY;
        float m = v;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
        const float e = lane < NC ? expf(v - m) : 0.f;
        float sum = e;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        const float lse = m + logf(sum);
        const float lp = v - lse;
        if (lane < NC && logprobs) logprobs[(size_t)b * NC + lane] = lp;
        const float ly = __shfl_sync(0xffffffffu, lp, y);
        if (lane == 0) { rowloss[b] = -ly; yb[b] = (uint8_t)y; }
        if (lane < NCP) {
            const float d = lane < NC ? (e / sum - (lane == y ? 1.f : 0.f)) / (float)B : 0.f;
            dzs[lane] = d;
            if (prep_backward) dlog[(size_t)b * NCP + lane] = d;
        }
    }
    __syncthreads();
    if (prep_backward) {
        float d = 0.f;
#pragma unroll
        for (int c = 0; c < NC; ++c) d = fmaf(dzs[c], params[OFF_F2W + c * FC + tid], d);
        dzf[(size_t)b * FC + tid] = (pos && keep) ? (train ? d * SCALE2 : d) : 0.f;
    }
    // batch loss: the last CTA sums the rows in order
    __threadfence();
    if (tid == 0) last = (atomicAdd(&ctrl[0], 1u) == (unsigned)(B - 1));
    __syncthreads();
    if (last && warp == 0) {
        __threadfence();
        float v = 0.f;
        for (int r0 = 0; r0 < B; r0 += 32) {
            float x = (r0 + lane < B) ? __ldcg(rowloss + r0 + lane) : 0.f;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
            v += x;
        }
        if (lane == 0) {
            const float mean = v / (float)B;
            if (loss2) { loss2[0] = mean; loss2[1] += mean / (float)B; }
            ctrl[0] = 0;
        }
    }
}

// ------------------------------------------------------------------------------------------ J4: fc1 backward
// grid = 144 feature tiles of 64 (one wave on 148 SMs), 256 threads.  CTA kt owns columns [64kt, 64kt+64) of fc1.weight:
//   dW[n][k] = scale . sum_b dzf[b][n] . p2d[b][k]      (complete: every row of the batch passes through this CTA)
//   g[b][k]  = sum_n dzf[b][n] . W[n][k], passed through the pooled element's flag (ReLU > 0 and kept) x 1/0.75
// Rows in chunks of 64; dW accumulates in registers (8 x 4 per thread) across chunks.
constexpr int F1B_KT = 64, F1B_LD = F1B_KT + 1;
constexpr int F1B_TILES = FEAT / F1B_KT;                   // 144
constexpr size_t F1B_SMEM = (size_t)(128 * F1B_LD + 64 * 128 + 64 * F1B_KT) * 4;
__global__ void __launch_bounds__(256, 1)
cn_fc1_bwd_kernel(const float* __restrict__ params, const float* __restrict__ p2d, const uint8_t* __restrict__ pflag,
                  const float* __restrict__ dzf, int B, int train, float scale, float* __restrict__ grads,
                  float* __restrict__ dp2) {
    extern __shared__ __align__(16) float sm[];
    float* Ws = sm;                           // [n][k] stride 65
    float* Ds = Ws + 128 * F1B_LD;            // [row][n]
    float* Xs = Ds + 64 * 128;                // [row][k]
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, k0 = blockIdx.x * F1B_KT;
#pragma unroll 16
    for (int seg = warp; seg < 128 * 2; seg += 8) {          // (unit n, 32 consecutive k)
        const int n = seg >> 1, kk = (seg & 1) * 32 + lane;
        Ws[n * F1B_LD + kk] = params[OFF_F1W + (size_t)n * FEAT + k0 + kk];
    }
    const int tk = tid & 15, tn = tid >> 4;                 // dW tile: units tn + 16 i, columns tk + 16 j
    float dw[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dw[i][j] = 0.f;
    for (int r0 = 0; r0 < B; r0 += 64) {
        __syncthreads();
#pragma unroll 8
        for (int t = tid; t < 64 * 128; t += 256) {
            const int r = t >> 7, c = t & 127;
            Ds[t] = (r0 + r < B) ? dzf[(size_t)(r0 + r) * FC + c] : 0.f;
        }
#pragma unroll 8
        for (int t = tid; t < 64 * F1B_KT; t += 256) {
            const int r = t >> 6, c = t & 63;
            Xs[t] = (r0 + r < B) ? p2d[(size_t)(r0 + r) * FEAT + k0 + c] : 0.f;
        }
        __syncthreads();
#pragma unroll 4
        for (int r = 0; r < 64; ++r) {
            float d[8], x[4];
#pragma unroll
            for (int i = 0; i < 8; ++i) d[i] = Ds[r * 128 + tn + 16 * i];
#pragma unroll
            for (int j = 0; j < 4; ++j) x[j] = Xs[r * F1B_KT + tk + 16 * j];
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) dw[i][j] = fmaf(d[i], x[j], dw[i][j]);
        }
        // gradient of the pooled activations of these 64 rows: thread -> rows rg*4.., columns tk + 16 j
        const int rg = tid >> 4;
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 4
        for (int n = 0; n < FC; ++n) {
            float d[4], w[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) d[i] = Ds[(rg * 4 + i) * 128 + n];
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = Ws[n * F1B_LD + tk + 16 * j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(d[i], w[j], acc[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int row = r0 + rg * 4 + i;
            if (row < B)
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const size_t e = (size_t)row * FEAT + k0 + tk + 16 * j;
                    const bool pass = (pflag[e] & PF_PASS) != 0;
                    dp2[e] = pass ? (train ? acc[i][j] * SCALE1 : acc[i][j]) : 0.f;
                }
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
            grads[OFF_F1W + (size_t)(tn + 16 * i) * FEAT + k0 + tk + 16 * j] = dw[i][j] * scale;
}

// ------------------------------------------------------------------------------------------ J4: conv2 backward
// The gradient reaching conv2's pre-activation is sparse: of each 2x2 window only the argmax element gets the pooled
// gradient g (and only if its value was > 0 and it was kept by the dropout).  Both kernels rebuild a dense, zero-
// filled dz2 plane set in shared memory from (g, flags) and then run a dense register-tiled product over it.

// data gradient: da1[ci][y][x] = sum_{co,ky,kx} dz2[co][y-ky][x-kx] . W[co][ci][ky][kx]; dz1 = da1 . (a1 > 0).
// grid = (B, 2): CTA (b, h) produces input channels [16h, 16h+16).  dz2 is held with a 2-wide zero border
// ([32 co][28x28], two passes over the output-channel halves).  352 threads, 338 active: thread -> (2x2 window of
// the 26x26 plane, 8 input channels).
constexpr int DG_THREADS = 352;
constexpr size_t DG_SMEM = (size_t)(32 * 784 + 32 * 9 * 16) * 4;
__global__ void __launch_bounds__(DG_THREADS, 1)
cn_conv2_dgrad_kernel(const float* __restrict__ params, const float* __restrict__ a1, const float* __restrict__ dp2,
                      const uint8_t* __restrict__ pflag, float* __restrict__ dz1) {
    extern __shared__ __align__(16) float sm[];
    float* dzs = sm;                          // [32][28*28] zero-bordered
    float* wsm = dzs + 32 * 784;              // [co 32][tap 9][ci 16]
    const int tid = threadIdx.x, b = blockIdx.x, half = blockIdx.y;
    const bool active = tid < 338;
    const int w = tid % 169, cg = (tid / 169) & 1;
    const int wy = w / 13, wx = w - wy * 13;
    float acc[4][8];
#pragma unroll
    for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[q][j] = 0.f;
    for (int hh = 0; hh < 2; ++hh) {
        __syncthreads();
        for (int t = tid; t < 32 * 784; t += DG_THREADS) dzs[t] = 0.f;
#pragma unroll 8
        for (int t = tid; t < 32 * 9 * 16; t += DG_THREADS) {
            const int ci = t & 15, tap = (t >> 4) % 9, co = t / 144;
            wsm[t] = params[OFF_C2W + ((size_t)(hh * 32 + co) * C1 + half * 16 + ci) * 9 + tap];
        }
        __syncthreads();
#pragma unroll 4
        for (int t = tid; t < 32 * NPOOL; t += DG_THREADS) {
            const int co = t / NPOOL, p = t - co * NPOOL, py = p / HP, px = p - py * HP;
            const size_t e = (size_t)b * FEAT + (size_t)(hh * 32 + co) * NPOOL + p;
            const unsigned f = pflag[e];
            if (f & PF_PASS) {
                const int yo = 2 * py + ((f >> 1) & 1), xo = 2 * px + (f & 1);
                dzs[co * 784 + (yo + 2) * 28 + xo + 2] = dp2[e];
            }
        }
        __syncthreads();
        if (active) {
            for (int co = 0; co < 32; ++co) {
                float patch[4][4];
                const float* base = dzs + co * 784 + (2 * wy) * 28 + 2 * wx;
#pragma unroll
                for (int r = 0; r < 4; ++r)
#pragma unroll
                    for (int c = 0; c < 4; ++c) patch[r][c] = base[r * 28 + c];
#pragma unroll
                for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) {
                        const float4 w0 = *reinterpret_cast<const float4*>(&wsm[((co * 9) + ky * 3 + kx) * 16 + cg * 8]);
                        const float4 w1 = *reinterpret_cast<const float4*>(&wsm[((co * 9) + ky * 3 + kx) * 16 + cg * 8 + 4]);
                        const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
                        for (int r = 0; r < 2; ++r)
#pragma unroll
                            for (int c = 0; c < 2; ++c)
#pragma unroll
                                for (int j = 0; j < 8; ++j)
                                    acc[r * 2 + c][j] = fmaf(patch[r - ky + 2][c - kx + 2], wv[j], acc[r * 2 + c][j]);
                    }
            }
        }
    }
    if (active) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int ci = half * 16 + cg * 8 + j;
#pragma unroll
            for (int r = 0; r < 2; ++r)
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    const size_t e = ((size_t)b * C1 + ci) * A1 + (2 * wy + r) * H1 + 2 * wx + c;
                    dz1[e] = a1[e] > 0.f ? acc[r * 2 + c][j] : 0.f;
                }
        }
    }
}

// weight gradient of one image: dW2[co][ci][tap] = sum_pos dz2[co][pos] . a1[ci][pos + tap], db2[co] = sum_pos dz2.
// grid = (B, 2): CTA (b, h) produces output channels [32h, 32h+32) into this image's partial-gradient row.
// 256 threads: thread -> (input channel = lane, 8 output channels = warp & 3, half of the rows = warp >> 2):
// 72 accumulators; walking a row it keeps the 3x3 input window in registers and loads 3 new values per step.
constexpr size_t WG_SMEM = (size_t)(C1 * A1S + 32 * 576) * 4;
__global__ void __launch_bounds__(256, 1)
cn_conv2_wgrad_kernel(const float* __restrict__ a1, const float* __restrict__ dp2, const uint8_t* __restrict__ pflag,
                      float* __restrict__ gpart) {
    extern __shared__ __align__(16) float sm[];
    float* a1s = sm;                          // [32][677]
    float* dzd = a1s + C1 * A1S;              // [32 co][24*24] dense
    const int tid = threadIdx.x, b = blockIdx.x, half = blockIdx.y;
    const float* src = a1 + (size_t)b * C1 * A1;
#pragma unroll 8
    for (int t = tid; t < C1 * A1; t += 256) {
        const int c = t / A1, pos = t - c * A1;
        a1s[c * A1S + pos] = src[t];
    }
    for (int t = tid; t < 32 * 576; t += 256) dzd[t] = 0.f;
    __syncthreads();
#pragma unroll 6
    for (int t = tid; t < 32 * NPOOL; t += 256) {
        const int co = t / NPOOL, p = t - co * NPOOL, py = p / HP, px = p - py * HP;
        const size_t e = (size_t)b * FEAT + (size_t)(half * 32 + co) * NPOOL + p;
        const unsigned f = pflag[e];
        if (f & PF_PASS) dzd[co * 576 + (2 * py + ((f >> 1) & 1)) * H2 + 2 * px + (f & 1)] = dp2[e];
    }
    __syncthreads();
    float* grow = gpart + (size_t)b * CONV_GRADS;
    {                                                        // db2: 8 threads per channel, 72 positions each, fixed-order combine
        const int co = tid >> 3, part = tid & 7;
        float v = 0.f;
#pragma unroll 8
        for (int i = 0; i < 72; ++i) v += dzd[co * 576 + part * 72 + i];
        v += __shfl_xor_sync(0xffffffffu, v, 1);
        v += __shfl_xor_sync(0xffffffffu, v, 2);
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        if (part == 0) grow[OFF_C2B + half * 32 + co] = v;
    }
    const int ci = tid & 31, oct = (tid >> 5) & 3, ph = tid >> 7;
    float acc[8][9];
#pragma unroll
    for (int j = 0; j < 8; ++j)
#pragma unroll
        for (int t = 0; t < 9; ++t) acc[j][t] = 0.f;
    const float* ap = a1s + ci * A1S;
    for (int yo = ph * 12; yo < ph * 12 + 12; ++yo) {
        float win[3][3];
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) { win[ky][1] = ap[(yo + ky) * H1]; win[ky][2] = ap[(yo + ky) * H1 + 1]; }
        for (int xo = 0; xo < H2; ++xo) {
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                win[ky][0] = win[ky][1]; win[ky][1] = win[ky][2]; win[ky][2] = ap[(yo + ky) * H1 + xo + 2];
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = dzd[(oct * 8 + j) * 576 + yo * H2 + xo];
#pragma unroll
                for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) acc[j][ky * 3 + kx] = fmaf(d, win[ky][kx], acc[j][ky * 3 + kx]);
            }
        }
    }
    __syncthreads();                                         // everyone is done with dzd: reuse it to combine the halves
    if (ph == 1) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
#pragma unroll
            for (int t = 0; t < 9; ++t) dzd[(j * 9 + t) * 128 + (tid & 127)] = acc[j][t];
    }
    __syncthreads();
    if (ph == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int co = half * 32 + oct * 8 + j;
#pragma unroll
            for (int t = 0; t < 9; ++t)
                grow[OFF_C2W + ((size_t)co * C1 + ci) * 9 + t] = acc[j][t] + dzd[(j * 9 + t) * 128 + tid];
        }
    }
}

// conv1's weight gradient of one image: dW1[c][tap] = sum_pos dz1[c][pos] . x[pos + tap], db1[c] = sum_pos dz1[c][pos]
// (conv1's input needs no gradient).  grid = B, 960 threads: (288 (channel, tap) pairs + 32 bias sums) x 3 row
// groups (rows 0-8, 9-17, 18-25), combined in fixed order through shared memory.
constexpr int C1B_THREADS = 960;
constexpr size_t C1B_SMEM = (size_t)(C1 * A1S + IN + 256 + 3 * 320) * 4;
__global__ void __launch_bounds__(C1B_THREADS, 2)
cn_conv1_bwd_kernel(const uint8_t* __restrict__ images, int64_t n_samples, const int64_t* __restrict__ idx,
                    const float* __restrict__ dz1, float* __restrict__ gpart) {
    extern __shared__ __align__(16) float sm[];
    float* dzs = sm;                          // [32][677]
    float* xs = dzs + C1 * A1S;               // [28*28]
    float* lut = xs + IN;
    float* red = lut + 256;                   // [3][320]
    const int tid = threadIdx.x, b = blockIdx.x;
    if (tid < 256) lut[tid] = decode_px(tid);
    const float* src = dz1 + (size_t)b * C1 * A1;
#pragma unroll 8
    for (int t = tid; t < C1 * A1; t += C1B_THREADS) {
        const int c = t / A1, pos = t - c * A1;
        dzs[c * A1S + pos] = src[t];
    }
    __syncthreads();
    const int64_t s = idx[b];
    const bool ok = s >= 0 && s < n_samples;
    for (int t = tid; t < IN; t += C1B_THREADS) xs[t] = ok ? lut[images[s * IN + t]] : 0.f;
    __syncthreads();
    const int part = tid / 320, item = tid - part * 320;       // item < 288: (channel, tap); else bias of channel item - 288
    const int y0 = part * 9, y1 = min(H1, y0 + 9);
    float v = 0.f;
    if (item < 288) {
        const int c = item / 9, tap = item - c * 9, ky = tap / 3, kx = tap - ky * 3;
        for (int y = y0; y < y1; ++y)
#pragma unroll 13
            for (int x = 0; x < H1; ++x) v = fmaf(dzs[c * A1S + y * H1 + x], xs[(y + ky) * 28 + x + kx], v);
    } else {
        const int c = item - 288;
        for (int pos = y0 * H1; pos < y1 * H1; ++pos) v += dzs[c * A1S + pos];
    }
    red[part * 320 + item] = v;
    __syncthreads();
    if (tid < 320) {
        const float sum = (red[tid] + red[320 + tid]) + red[640 + tid];
        float* grow = gpart + (size_t)b * CONV_GRADS;
        if (tid < 288) grow[OFF_C1W + tid] = sum;
        else grow[OFF_C1B + tid - 288] = sum;
    }
}

// db2[co] = scale . sum of the channel's pooled gradients over the batch (fixed order)
__device__ __forceinline__ void db2_block(const float* __restrict__ dp2, int B, float scale, float* __restrict__ grads, int co) {
    __shared__ float red[256];
    const int tid = threadIdx.x;
    float v = 0.f;
#pragma unroll 8
    for (int i = tid; i < B * NPOOL; i += 256) {
        const int b = i / NPOOL, p = i - b * NPOOL;
        v += dp2[(size_t)b * FEAT + (size_t)co * NPOOL + p];
    }
    red[tid] = v;
    __syncthreads();
    for (int s = 128; s > 0; s >>= 1) {
        if (tid < s) red[tid] += red[tid + s];
        __syncthreads();
    }
    if (tid == 0) grads[OFF_C2B + co] = red[0] * scale;
}


// sum of the tensor-core weight-gradient kernel's per-CTA partials: four quarter-sums (partials in CTA order) per
// element, combined in fixed order; 64 elements per block
__device__ __forceinline__ void wpart_reduce_block(const float* __restrict__ wpart, int nparts, float scale,
                                                   float* __restrict__ grads, int blk) {
    __shared__ float red4[4][64];
    const int jl = threadIdx.x & 63, grp = threadIdx.x >> 6;
    const int j = blk * 64 + jl;
    const int per = (nparts + 3) / 4, c0 = grp * per, c1 = min(nparts, c0 + per);
    float v = 0.f;
#pragma unroll 8
    for (int c = c0; c < c1; ++c) v += wpart[(size_t)c * (C2 * C1 * 9) + j];
    red4[grp][jl] = v;
    __syncthreads();
    if (grp == 0) grads[OFF_C2W + j] = ((red4[0][jl] + red4[1][jl]) + (red4[2][jl] + red4[3][jl])) * scale;
}

// sum of the per-image partial gradients and the three small gradients of the dense layers (fc1.bias, fc2.weight,
// fc2.bias: sums over the batch of dzf, dlogits x hd, dlogits).  One warp per output element: lane l takes images
// l, l+32, ... (in order), then the 32 lane sums are combined by a fixed shuffle tree - the same order every run.
constexpr int SMALL_N = FC + NC * FC + NC;                 // 1418
__global__ void __launch_bounds__(256)
cn_reduce_kernel(const float* __restrict__ gpart, const float* __restrict__ dzf, const float* __restrict__ dlog,
                 const float* __restrict__ hd, int B, float scale, int nconv, float* __restrict__ grads,
                 const float* __restrict__ wpart, int nparts, const float* __restrict__ dp2, int first_tc_block) {
    // tensor-core backward: the blocks from first_tc_block on finish conv2's gradients (288 blocks sum the weight-
    // gradient partials, 64 form the bias gradient) - same launch as the other reductions
    if ((int)blockIdx.x >= first_tc_block) {
        const int blk = blockIdx.x - first_tc_block;
        if (blk < C2 * C1 * 9 / 64) wpart_reduce_block(wpart, nparts, scale, grads, blk);
        else db2_block(dp2, B, scale, grads, blk - C2 * C1 * 9 / 64);
        return;
    }
    const int lane = threadIdx.x & 31;
    const int e = blockIdx.x * 8 + (threadIdx.x >> 5);     // 8 elements per block
    if (e >= nconv + SMALL_N) return;
    float v = 0.f;
    int out;
    if (e < nconv) {                                       // conv1 (+ conv2 in the FFMA path): per-image partials
        for (int b = lane; b < B; b += 32) v += gpart[(size_t)b * CONV_GRADS + e];
        out = e;
    } else if (e < nconv + FC) {
        const int n = e - nconv;
        for (int b = lane; b < B; b += 32) v += dzf[(size_t)b * FC + n];
        out = OFF_F1B + n;
    } else if (e < nconv + FC + NC * FC) {
        const int c = (e - nconv - FC) / FC, n = (e - nconv - FC) - c * FC;
        for (int b = lane; b < B; b += 32) v = fmaf(dlog[(size_t)b * NCP + c], hd[(size_t)b * FC + n], v);
        out = OFF_F2W + c * FC + n;
    } else {
        const int c = e - nconv - FC - NC * FC;
        for (int b = lane; b < B; b += 32) v += dlog[(size_t)b * NCP + c];
        out = OFF_F2B + c;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) grads[out] = v * scale;
}

// ------------------------------------------------------------------------------------------ J5: Adadelta
// torch.optim.Adadelta's single-tensor rule (torch/optim/adadelta.py:292-302), one pass over the flat arrays:
//   sq <- rho sq + (1-rho) g g ; std = sqrt(sq + eps) ; delta = sqrt(acc + eps) / std . g ;
//   acc <- rho acc + (1-rho) delta delta ; p <- p - lr delta
// Every operation individually rounded (no contraction), as the CPU's vectorised elementwise kernels do.
__device__ __forceinline__ void adadelta_elem(float& p, float g, float& sq, float& acc, float lr, float rho,
                                              float one_minus_rho, float eps, float gscale) {
    const float gv = __fmul_rn(g, gscale);
    const float s = __fadd_rn(__fmul_rn(sq, rho), __fmul_rn(__fmul_rn(one_minus_rho, gv), gv));
    const float std = __fsqrt_rn(__fadd_rn(s, eps));
    const float delta = __fmul_rn(__fdiv_rn(__fsqrt_rn(__fadd_rn(acc, eps)), std), gv);
    acc = __fadd_rn(__fmul_rn(acc, rho), __fmul_rn(__fmul_rn(one_minus_rho, delta), delta));
    sq = s;
    p = fmaf(-lr, delta, p);
}
__global__ void __launch_bounds__(256)
adadelta_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ sq, float* __restrict__ acc,
                int64_t n, float lr, float rho, float one_minus_rho, float eps, float gscale, int vec) {
    const int64_t n4 = vec ? n / 4 : 0;                       // 16-byte path when all four arrays are 16-byte aligned
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (int64_t)gridDim.x * 256) {
        float4 pv = reinterpret_cast<float4*>(p)[i], sv = reinterpret_cast<float4*>(sq)[i], av = reinterpret_cast<float4*>(acc)[i];
        const float4 gv = reinterpret_cast<const float4*>(g)[i];
        adadelta_elem(pv.x, gv.x, sv.x, av.x, lr, rho, one_minus_rho, eps, gscale);
        adadelta_elem(pv.y, gv.y, sv.y, av.y, lr, rho, one_minus_rho, eps, gscale);
        adadelta_elem(pv.z, gv.z, sv.z, av.z, lr, rho, one_minus_rho, eps, gscale);
        adadelta_elem(pv.w, gv.w, sv.w, av.w, lr, rho, one_minus_rho, eps, gscale);
        reinterpret_cast<float4*>(p)[i] = pv; reinterpret_cast<float4*>(sq)[i] = sv; reinterpret_cast<float4*>(acc)[i] = av;
    }
    for (int64_t i = n4 * 4 + (int64_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (int64_t)gridDim.x * 256)
        adadelta_elem(p[i], g[i], sq[i], acc[i], lr, rho, one_minus_rho, eps, gscale);
}

}  // namespace ddpm

// ------------------------------------------------------------------------------------------ C ABI
using namespace ddpm;

static int cn_configure() {
    static bool done[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !done[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(cn_conv2_pool_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C2_SMEM));
        DDPM_CUDA(cudaFuncSetAttribute(cn_fc1_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F1_SMEM));
        DDPM_CUDA(cudaFuncSetAttribute(cn_fc1_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F1B_SMEM));
        DDPM_CUDA(cudaFuncSetAttribute(cn_conv2_dgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)DG_SMEM));
        DDPM_CUDA(cudaFuncSetAttribute(cn_conv2_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)WG_SMEM));
        DDPM_CUDA(cudaFuncSetAttribute(cn_conv1_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C1B_SMEM));
        done[dev] = true;
    }
    return 0;
}

extern "C" size_t ddpm_cn_workspace_bytes(int max_batch) {
    if (max_batch <= 0 || max_batch > DDPM_CN_MAX_BATCH) return 0;
    return cn_layout(max_batch).total;
}

extern "C" int ddpm_cn_workspace_init(void* d_workspace, int max_batch, void* stream) {
    DDPM_CHECK_ARG(d_workspace != nullptr && max_batch > 0 && max_batch <= DDPM_CN_MAX_BATCH,
                   "ddpm_cn_workspace_init: bad arguments (max_batch %d, limit %d)", max_batch, DDPM_CN_MAX_BATCH);
    DDPM_CUDA(cudaMemsetAsync(d_workspace, 0, 256, (cudaStream_t)stream));
    return 0;
}

extern "C" int ddpm_cn_step_forward(const float* d_params, const uint8_t* d_images, const uint8_t* d_labels,
                                    int64_t n_samples, const int64_t* d_idx, const uint8_t* d_mask1,
                                    const uint8_t* d_mask2, uint64_t philox_seed, uint64_t philox_step, int batch,
                                    unsigned flags, float* d_logprobs, float* d_loss, void* d_workspace, void* stream) {
    const char* who = "ddpm_cn_step_forward";
    DDPM_CHECK_ARG(batch > 0 && batch <= DDPM_CN_MAX_BATCH, "%s: batch %d outside [1, %d]", who, batch, DDPM_CN_MAX_BATCH);
    DDPM_CHECK_ARG(d_params && d_images && d_labels && d_idx && d_workspace && n_samples > 0, "%s: null pointer or empty dataset", who);
    DDPM_CHECK_ARG((flags & ~(DDPM_FWD_PREP_BACKWARD | DDPM_FWD_PHILOX_MASK | DDPM_FWD_TF32)) == 0, "%s: unknown flags %#x", who, flags);
    DDPM_CHECK_ARG((d_mask1 == nullptr) == (d_mask2 == nullptr), "%s: give both dropout masks or neither", who);
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0 && ((uintptr_t)d_images & 3) == 0, "%s: misaligned pointer", who);
    if (cn_configure()) return 1;
    const bool philox = (flags & DDPM_FWD_PHILOX_MASK) != 0;
    const int train = (philox || d_mask1 != nullptr) ? 1 : 0;
    const int prep = (flags & DDPM_FWD_PREP_BACKWARD) ? 1 : 0;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_workspace;
    const CnLayout L = cn_layout(batch);
    float* a1 = (float*)(ws + L.a1);
    float* p2d = (float*)(ws + L.p2d);
    uint8_t* pflag = (uint8_t*)(ws + L.pflag);
    float* part = (float*)(ws + L.part);
    uint32_t* ctrl = (uint32_t*)(ws + L.ctrl);
    if (flags & DDPM_FWD_TF32) {
        // conv2 (89 % of the forward flops) as an implicit GEMM on the tensor cores (convnet_tc.cu)
        float* a1n = (float*)(ws + L.a1n);
        float* w2t = (float*)(ws + L.w2t);
        cn_conv1_kernel<true><<<dim3(batch, 4), 256, 0, st>>>(d_params, d_images, n_samples, d_idx, a1, a1n, w2t, (float*)(ws + L.w2d));
        if (launch_cn_conv2_tc(w2t, d_params, a1n, philox ? nullptr : d_mask1, philox_seed, philox_step, philox ? 1 : 0,
                               train, batch, p2d, pflag, st)) return 1;
    } else {
        cn_conv1_kernel<false><<<dim3(batch, 3), 256, 0, st>>>(d_params, d_images, n_samples, d_idx, a1, nullptr, nullptr, nullptr);
        cn_conv2_pool_kernel<<<dim3(batch, 2), C2_THREADS, C2_SMEM, st>>>(d_params, a1, philox ? nullptr : d_mask1, philox_seed,
                                                                        philox_step, philox ? 1 : 0, train, p2d, pflag);
    }
    if (flags & DDPM_FWD_TF32) {
        if (launch_cn_fc1_fwd_tc(d_params, p2d, batch, part, st)) return 1;
    } else {
        cn_fc1_fwd_kernel<<<KSPLIT, 256, F1_SMEM, st>>>(d_params, p2d, batch, part);
    }
    cn_head_kernel<<<batch, FC, 0, st>>>(d_params, part, d_labels, n_samples, d_idx, philox ? nullptr : d_mask2, philox_seed,
                                         philox_step, philox ? 1 : 0, train, prep, batch, (float*)(ws + L.hd),
                                         (float*)(ws + L.dzf), (float*)(ws + L.dlog), d_logprobs, (float*)(ws + L.rowloss),
                                         (uint8_t*)(ws + L.yb), d_loss, ctrl);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int ddpm_cn_step_backward(const float* d_params, const uint8_t* d_images, int64_t n_samples,
                                     const int64_t* d_idx, int batch, unsigned flags, float grad_scale, float* d_grads,
                                     void* d_workspace, void* stream) {
    const char* who = "ddpm_cn_step_backward";
    DDPM_CHECK_ARG(batch > 0 && batch <= DDPM_CN_MAX_BATCH, "%s: batch %d outside [1, %d]", who, batch, DDPM_CN_MAX_BATCH);
    DDPM_CHECK_ARG(d_params && d_images && d_idx && d_grads && d_workspace && n_samples > 0, "%s: null pointer or empty dataset", who);
    DDPM_CHECK_ARG(((uintptr_t)d_workspace & 255) == 0, "%s: misaligned workspace", who);
    DDPM_CHECK_ARG((flags & ~(DDPM_CN_BWD_TRAIN | DDPM_CN_BWD_TF32 | DDPM_CN_BWD_FWD_WAS_TF32)) == 0, "%s: unknown flags %#x", who, flags);
    if (cn_configure()) return 1;
    const int train = (flags & DDPM_CN_BWD_TRAIN) ? 1 : 0;
    cudaStream_t st = (cudaStream_t)stream;
    char* ws = (char*)d_workspace;
    const CnLayout L = cn_layout(batch);
    float* a1 = (float*)(ws + L.a1);
    float* p2d = (float*)(ws + L.p2d);
    uint8_t* pflag = (uint8_t*)(ws + L.pflag);
    float* dzf = (float*)(ws + L.dzf);
    float* dp2 = (float*)(ws + L.dp2);
    float* dz1 = (float*)(ws + L.dz1);
    float* gpart = (float*)(ws + L.gpart);
    const bool tc = (flags & DDPM_CN_BWD_TF32) != 0;
    if (tc) {
        if (launch_cn_fc1_bwd_tc(d_params, p2d, pflag, dzf, batch, train ? 1 : 0, grad_scale, d_grads, dp2, st)) return 1;
    } else {
        cn_fc1_bwd_kernel<<<F1B_TILES, 256, F1B_SMEM, st>>>(d_params, p2d, pflag, dzf, batch, train ? 1 : 0, grad_scale, d_grads, dp2);
    }
    if (tc) {
        // conv2's data and weight gradients (2/3 of the backward flops) as implicit GEMMs on the tensor cores
        float* w2d = (float*)(ws + L.w2d);
        float* dz2n = (float*)(ws + L.dz2n);
        // DDPM_CN_BWD_FWD_WAS_TF32: the forward left the TF32 weight copies and the pixel-major a1 in this workspace
        const int ready = (flags & DDPM_CN_BWD_FWD_WAS_TF32) ? 1 : 0;
        if (!ready && launch_cn_w2_prep(d_params, (float*)(ws + L.w2t), w2d, st)) return 1;
        if (launch_cn_conv2_dgrad_tc(w2d, dp2, pflag, a1, batch, dz2n, dz1, st)) return 1;
        if (launch_cn_conv2_wgrad_tc(a1, dz2n, batch, (float*)(ws + L.a1n), ready, (float*)(ws + L.wpart), st)) return 1;
    } else {
        cn_conv2_wgrad_kernel<<<dim3(batch, 2), 256, WG_SMEM, st>>>(a1, dp2, pflag, gpart);
        cn_conv2_dgrad_kernel<<<dim3(batch, 2), DG_THREADS, DG_SMEM, st>>>(d_params, a1, dp2, pflag, dz1);
    }
    cn_conv1_bwd_kernel<<<batch, C1B_THREADS, C1B_SMEM, st>>>(d_images, n_samples, d_idx, dz1, gpart);
    const int nconv = tc ? OFF_C2W : CONV_GRADS;            // conv2's gradients came from the tensor-core kernels
    const int red_blocks = (nconv + SMALL_N + 7) / 8;
    const int tc_blocks = tc ? C2 * C1 * 9 / 64 + C2 : 0;
    cn_reduce_kernel<<<red_blocks + tc_blocks, 256, 0, st>>>(gpart, dzf, (const float*)(ws + L.dlog), (const float*)(ws + L.hd),
                                                            batch, grad_scale, nconv, d_grads, (const float*)(ws + L.wpart),
                                                            tc ? cn_wgrad_tc_parts(batch) : 0, dp2, red_blocks);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int ddpm_adadelta_step(float* d_params, const float* d_grads, float* d_square_avg, float* d_acc_delta,
                                  int64_t n, float lr, double rho, double eps, float grad_scale, void* stream) {
    DDPM_CHECK_ARG(d_params && d_grads && d_square_avg && d_acc_delta && n > 0, "ddpm_adadelta_step: null pointer or n <= 0");
    const float one_minus_rho = (float)(1.0 - rho);          // torch forms 1 - rho in double, then rounds once
    const int vec = ((((uintptr_t)d_params | (uintptr_t)d_grads | (uintptr_t)d_square_avg | (uintptr_t)d_acc_delta) & 15) == 0) ? 1 : 0;
    const int64_t work = vec ? (n + 3) / 4 : n;
    const int blocks = (int)((work + 255) / 256 < 148 * 8 ? (work + 255) / 256 : 148 * 8);
    adadelta_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(d_params, d_grads, d_square_avg, d_acc_delta, n, lr, (float)rho,
                                                              one_minus_rho, (float)eps, grad_scale, vec);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}
