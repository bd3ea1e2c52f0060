// conv2 of the BASELINE.json ConvNet (row J2) as an implicit GEMM on the 5th-generation tensor cores.
//
//   out[pos][co] = sum_{tap, ci} a1[pos + tap][ci] . W[co][ci][tap]        M = B.576 positions, N = 64, K = 288
//
// Opt-in (DDPM_FWD_TF32): both operands are rounded to TF32, accumulation is FP32 in TMEM.
//
// conv1 leaves a1 pixel-major ([B][26x26][32 channels], TF32-rounded), so the 32 input channels of one pixel are one
// 128-byte row: exactly one row of a K-major SWIZZLE_128B operand tile.  The im2col matrix never exists: for each of
// the 9 taps the A tile [128 positions x 32 ci] is gathered with 16-byte cp.async copies (source = the pixel shifted
// by the tap) straight into the swizzled layout; all 9 tap tiles of a position tile (144 KB) and the whole weight
// matrix as 9 tiles [64 co x 32 ci] (72 KB) are resident in shared memory, then one thread issues 36 tcgen05.mma
// (M128 N64 K8) into a 64-column TMEM accumulator.
//
// Rows are ordered so that the four positions of a 2x2 pooling window are four consecutive TMEM lanes
// (row = 4 . window + dy.2 + dx): the epilogue reads the accumulator with tcgen05.ld, adds the bias, reduces max and
// argmax over the four lanes with two shuffle rounds, applies ReLU and dropout and writes the pooled activations in
// fc1's (c, y, x) order plus the flag byte the backward pass needs - conv2's 24x24 output never reaches memory.
#include "convnet.cuh"
#include "tc_common.cuh"

namespace ddpm {

constexpr int CT_THREADS = 256;
constexpr int CT_A_TILE = 128 * 128;                    // one tap: 128 positions x 128 B
constexpr int CT_W_TILE = 64 * 128;                     // one tap: 64 output channels x 128 B
constexpr int CT_A_BYTES = 9 * CT_A_TILE;               // 147,456
constexpr int CT_W_BYTES = 9 * CT_W_TILE;               // 73,728
constexpr size_t CT_SMEM = (size_t)CT_A_BYTES + CT_W_BYTES + 256 /*bias*/ + 64 /*mbarrier, tmem ptr*/ + 1024 /*align*/;
constexpr uint32_t CT_IDESC = umma_idesc_tf32(128, 64);

// conv2.weight [co][ci][tap] -> w2t [tap][co][ci] (forward: K = ci) and w2d [tap][ci][co] (data gradient: K = co),
// TF32-rounded.  18,432 elements; runs once per step (the weights change every step).
__global__ void __launch_bounds__(256)
cn_w2_prep_kernel(const float* __restrict__ params, float* __restrict__ w2t, float* __restrict__ w2d) {
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= C2 * C1 * 9) return;
    const int co = t / (C1 * 9), r = t - co * (C1 * 9), ci = r / 9, tap = r - ci * 9;
    const float v = to_tf32(params[OFF_C2W + t]);
    w2t[(tap * C2 + co) * C1 + ci] = v;
    w2d[(tap * C1 + ci) * C2 + co] = v;
}

__global__ void __launch_bounds__(CT_THREADS, 1)
cn_conv2_tc_kernel(const float* __restrict__ w2t, const float* __restrict__ params, const float* __restrict__ a1n,
                   const uint8_t* __restrict__ mask1, uint64_t seed, uint64_t step, int use_philox, int train, int B,
                   float* __restrict__ p2d, uint8_t* __restrict__ pflag) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t* As = smem;
    uint8_t* Ws = smem + CT_A_BYTES;
    float* bs = reinterpret_cast<float*>(Ws + CT_W_BYTES);
    uint64_t* mbar = reinterpret_cast<uint64_t*>(bs + 64);
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(mbar + 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    if (tid == 0) mbar_init(mbar, 1);
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" :: "r"(smem_u32(tmem_slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (tid < C2) bs[tid] = params[OFF_C2B + tid];
    // weights: 9 x 64 rows of 128 B, 8 chunks each
    for (int t = tid; t < 9 * C2 * 8; t += CT_THREADS) {
        const int row = t >> 3, q = t & 7;
        cp_async16(smem_u32(Ws) + (uint32_t)(row >> 6) * CT_W_TILE + sw128(row & 63, q), w2t + (size_t)row * C1 + q * 4, 16);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;

    const int nwin = B * NPOOL;
    const int ntiles = (nwin + 31) / 32;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        // ---- gather the 9 tap tiles: 8 consecutive threads copy one 128-byte pixel row
        {
            const int q = tid & 7;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = (tid >> 3) + 32 * i;
                const int w = tile * 32 + (r >> 2), j = r & 3;
                const bool valid = w < nwin;
                const int wv = valid ? w : 0;
                const int b = wv / NPOOL, p = wv - b * NPOOL, py = p / HP, px = p - py * HP;
                const int yo = 2 * py + (j >> 1), xo = 2 * px + (j & 1);
                const float* base = a1n + ((size_t)(b * H1 + yo) * H1 + xo) * C1 + q * 4;
                const uint32_t dst = smem_u32(As) + sw128(r, q);
#pragma unroll
                for (int tap = 0; tap < 9; ++tap)
                    cp_async16_ca(dst + (uint32_t)tap * CT_A_TILE, base + ((tap / 3) * H1 + (tap % 3)) * C1, valid ? 16 : 0);
            }
        }
        cp_async_wait_all();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // generic-proxy writes -> async proxy (tcgen05.mma)
        __syncthreads();
        if (tid == 0) {
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t a0 = smem_u32(As), b0 = smem_u32(Ws);
#pragma unroll
            for (int tap = 0; tap < 9; ++tap)
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    umma_tf32(tmem_base, umma_desc_sw128(a0 + tap * CT_A_TILE + k * 32), umma_desc_sw128(b0 + tap * CT_W_TILE + k * 32),
                              CT_IDESC, (tap | k) ? 1u : 0u);
            umma_commit(mbar);
        }
        mbar_wait(mbar, phase);
        phase ^= 1;
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

        // ---- epilogue: warp -> TMEM lanes 32.(warp & 3).., columns 32.(warp >> 2)..; lane = position within the tile
        {
            uint32_t v[32];
            const int lrow = 32 * (warp & 3) + lane, c0 = 32 * (warp >> 2);
            tmem_ld_32x32(tmem_base + ((uint32_t)(32 * (warp & 3)) << 16) + (uint32_t)c0, v);
            const int j = lane & 3;
            const int w = tile * 32 + (lrow >> 2);
            const bool valid = w < nwin;
            const int wv = valid ? w : 0;
            const int b = wv / NPOOL, p = wv - b * NPOOL;
            // max / argmax over the window's 4 lanes for all 32 channels; lane j keeps channels [8j, 8j+8)
            float mx[8];
            int ar[8];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                float m = __uint_as_float(v[i]) + bs[c0 + i];
                int arg = j;
#pragma unroll
                for (int off = 1; off <= 2; off <<= 1) {
                    const float om = __shfl_xor_sync(0xffffffffu, m, off);
                    const int oa = __shfl_xor_sync(0xffffffffu, arg, off);
                    if (om > m || (om == m && oa < arg)) { m = om; arg = oa; }     // first maximum of the window scan wins
                }
                if ((i >> 3) == j) { mx[i & 7] = m; ar[i & 7] = arg; }
            }
            // (kept out of the loop above: there the dropout draw would run, divergently, once per channel per warp)
            if (valid) {
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    const size_t e = (size_t)b * FEAT + (size_t)(c0 + 8 * j + k) * NPOOL + p;
                    bool keep = true;
                    if (train) keep = use_philox ? philox_keep_thr(seed, step, (uint32_t)e, KEEP1_THR) : (mask1[e] != 0);
                    const bool pos = mx[k] > 0.f;
                    p2d[e] = (pos && keep) ? (train ? mx[k] * SCALE1 : mx[k]) : 0.f;
                    pflag[e] = (uint8_t)((unsigned)ar[k] | ((pos && keep) ? PF_PASS : 0u));
                }
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();                                   // accumulator and tap tiles are free for the next tile
    }
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" :: "r"(tmem_base) : "memory");
}

int launch_cn_w2_prep(const float* params, float* w2t, float* w2d, cudaStream_t st) {
    cn_w2_prep_kernel<<<(C2 * C1 * 9 + 255) / 256, 256, 0, st>>>(params, w2t, w2d);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

int launch_cn_conv2_tc(const float* w2t, const float* params, const float* a1n, const uint8_t* mask1, uint64_t seed,
                       uint64_t step, int use_philox, int train, int B, float* p2d, uint8_t* pflag, cudaStream_t st) {
    static bool configured[64] = {};
    int dev = 0;
    DDPM_CUDA(cudaGetDevice(&dev));
    if (dev >= 0 && dev < 64 && !configured[dev]) {
        DDPM_CUDA(cudaFuncSetAttribute(cn_conv2_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)CT_SMEM));
        configured[dev] = true;
    }
    const int ntiles = (B * NPOOL + 31) / 32;
    const int grid = ntiles < 148 ? ntiles : 148;
    cn_conv2_tc_kernel<<<grid, CT_THREADS, CT_SMEM, st>>>(w2t, params, a1n, mask1, seed, step, use_philox, train, B, p2d, pflag);
    DDPM_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace ddpm
