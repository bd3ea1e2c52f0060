// Constants, workspace layout and launchers shared by the ConvNet kernels (convnet_kernels.cu, convnet_tc.cu).
#pragma once
#include "common.cuh"

namespace ddpm {

constexpr int C1 = 32, C2 = 64, H1 = 26, H2 = 24, HP = 12;
constexpr int A1 = H1 * H1;                    // 676 conv1 outputs per channel
constexpr int A1S = A1 + 1;                    // padded channel stride in shared memory (677: odd -> lanes = channels hit 32 banks)
constexpr int NPOOL = HP * HP;                 // 144
constexpr int FEAT = C2 * NPOOL;               // 9216
constexpr int FC = 128;
constexpr int OFF_C1W = DDPM_CN_OFF_CONV1_W, OFF_C1B = DDPM_CN_OFF_CONV1_B, OFF_C2W = DDPM_CN_OFF_CONV2_W,
              OFF_C2B = DDPM_CN_OFF_CONV2_B, OFF_F1W = DDPM_CN_OFF_FC1_W, OFF_F1B = DDPM_CN_OFF_FC1_B,
              OFF_F2W = DDPM_CN_OFF_FC2_W, OFF_F2B = DDPM_CN_OFF_FC2_B;
constexpr int CONV_GRADS = OFF_F1W;            // 18,816 conv1 + conv2 gradient elements come as per-image partials
constexpr int KSPLIT = FEAT / 128;             // 72 split-K slices of fc1
constexpr uint32_t KEEP1_THR = 3221225472u;    // 0.75 * 2^32
constexpr uint32_t KEEP2_THR = 2147483648u;    // 0.5  * 2^32
constexpr float SCALE1 = 1.0f / 0.75f, SCALE2 = 2.0f;
// pooled-activation flags (one byte per pooled element)
constexpr unsigned PF_PASS = 4u;               // bits 0-1: argmax inside the 2x2 window (dy*2+dx); bit 2: gradient passes (value > 0 and kept)

struct CnLayout {
    size_t ctrl, a1, p2d, pflag, part, hd, dzf, dlog, rowloss, yb, dp2, dz1, gpart;
    size_t a1n, w2t, w2d, dz2n, wpart;               // tensor-core path: a1 as NHWC TF32, conv2.weight as [tap][co][ci] / [tap][half][ci][co], dense padded dz2
    size_t total;
};
static inline CnLayout cn_layout(int B) {
    CnLayout L;
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o += (bytes + 255) & ~(size_t)255; return at; };
    L.ctrl = take(256);
    L.a1 = take((size_t)B * C1 * A1 * 4);
    L.p2d = take((size_t)B * FEAT * 4);
    L.pflag = take((size_t)B * FEAT);
    L.part = take((size_t)KSPLIT * B * FC * 4);
    L.hd = take((size_t)B * FC * 4);
    L.dzf = take((size_t)B * FC * 4);
    L.dlog = take((size_t)B * NCP * 4);
    L.rowloss = take((size_t)B * 4);
    L.yb = take((size_t)B);
    L.dp2 = take((size_t)B * FEAT * 4);
    L.dz1 = take((size_t)B * C1 * A1 * 4);
    L.gpart = take((size_t)B * CONV_GRADS * 4);
    L.a1n = take((size_t)B * A1 * C1 * 4);
    L.w2t = take((size_t)9 * C2 * C1 * 4);
    L.w2d = take((size_t)9 * C2 * C1 * 4);
    L.dz2n = take((size_t)B * 28 * 28 * C2 * 4);
    L.wpart = take((size_t)148 * C2 * C1 * 9 * 4);        // per-CTA partial dW2 of the tensor-core weight gradient
    L.total = o;
    return L;
}

__host__ __device__ inline bool philox_keep_thr(uint64_t seed, uint64_t step, uint32_t elem, uint32_t thr) {
    uint32_t r[4];
    Philox::gen(seed, step, elem >> 2, r);
    return r[elem & 3] < thr;
}

// Dropout(.25) draws for 8 consecutive channels [co0, co0+8) (co0 % 8 == 0) of pooled position p of image b: the
// random stream is indexed channel-fastest, (b.144 + p).64 + co, so the 8 draws are two Philox blocks of four
// 32-bit words - a draw per element indexed in fc1's (c, y, x) order would cost a whole block per element.
__host__ __device__ inline unsigned philox_keep8(uint64_t seed, uint64_t step, int b, int p, int co0, uint32_t thr) {
    const uint32_t blk = ((uint32_t)(b * NPOOL + p) * C2 + (uint32_t)co0) >> 2;
    uint32_t r0[4], r1[4];
    Philox::gen(seed, step, blk, r0);
    Philox::gen(seed, step, blk + 1, r1);
    unsigned m = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) m |= (r0[i] < thr ? 1u : 0u) << i | (r1[i] < thr ? 1u : 0u) << (4 + i);
    return m;
}

// the same stream for 4 consecutive channels [co0, co0+4), co0 % 4 == 0: one block
__host__ __device__ inline unsigned philox_keep4(uint64_t seed, uint64_t step, int b, int p, int co0, uint32_t thr) {
    uint32_t r0[4];
    Philox::gen(seed, step, ((uint32_t)(b * NPOOL + p) * C2 + (uint32_t)co0) >> 2, r0);
    unsigned m = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) m |= (r0[i] < thr ? 1u : 0u) << i;
    return m;
}

__device__ __forceinline__ float tf32_round(float x) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
    return __uint_as_float(r);
}

// tensor-core kernels (convnet_tc.cu)
int launch_cn_w2_prep(const float* params, float* w2t, float* w2d, cudaStream_t st);
int launch_cn_conv2_tc(const float* w2t, const float* params, const float* a1n, const uint8_t* mask1, uint64_t seed,
                       uint64_t step, int use_philox, int train, int B, float* p2d, uint8_t* pflag, cudaStream_t st);

int launch_cn_conv2_dgrad_tc(const float* w2d, const float* dp2, const uint8_t* pflag, const float* a1, int B, float* dz2n,
                             float* dz1, cudaStream_t st);

int cn_wgrad_tc_parts(int B);
int launch_cn_conv2_wgrad_tc(const float* a1, const float* dz2n, int B, float* a1n, int a1n_ready, float* wpart,
                             cudaStream_t st);

int launch_cn_fc1_fwd_tc(const float* params, const float* p2d, int B, float* part, cudaStream_t st);

int launch_cn_fc1_bwd_tc(const float* params, const float* p2d, const uint8_t* pflag, const float* dzf, int B, int train,
                         float scale, float* grads, float* dp2, cudaStream_t st);

}  // namespace ddpm
