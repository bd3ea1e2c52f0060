// Shared definitions for libddpmnist_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/ddpmnist_b200.h"

namespace ddpm {

constexpr int IN = DDPM_IN;          // 784 = 49 * 16
constexpr int HID = DDPM_HID;        // 128
constexpr int NC = DDPM_CLASSES;     // 10
constexpr int NCP = 16;              // logits / dlogits rows padded to 16 floats in the workspace
constexpr int KCHUNK = 16;           // fc1 K is walked in 16-element chunks (one uint4 of pixels)
constexpr int NCHUNKS = IN / KCHUNK; // 49
constexpr int NP = DDPM_N_PARAMS;

// ---- error plumbing -----------------------------------------------------------------------
extern thread_local char g_err[512];
int set_error(const char* fmt, ...);

#define DDPM_CHECK_ARG(cond, ...)                       \
    do {                                                \
        if (!(cond)) return ::ddpm::set_error(__VA_ARGS__); \
    } while (0)

#define DDPM_CUDA(call)                                                                     \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess)                                                             \
            return ::ddpm::set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), \
                                     __FILE__, __LINE__);                                   \
    } while (0)

// ---- decode table -------------------------------------------------------------------------
// ToTensor (v/255) then Normalize ((t - 0.1307)/0.3081), mp:374-377.  Plain `/` on floats is the
// IEEE-rounded division in CUDA device code unless -use_fast_math is given (it is not), and on
// the host; __fdiv_rn/__fsub_rn additionally forbid contraction into an FMA.
__host__ __device__ __forceinline__ float decode_px(unsigned v) {
#ifdef __CUDA_ARCH__
    return __fdiv_rn(__fsub_rn(__fdiv_rn((float)v, 255.0f), 0.1307f), 0.3081f);
#else
    volatile float t = (float)v / 255.0f;
    volatile float u = t - 0.1307f;
    return u / 0.3081f;
#endif
}

// ---- Philox4x32-10 ------------------------------------------------------------------------
struct Philox {
    static constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    __host__ __device__ static inline void round(uint32_t c[4], uint32_t k0, uint32_t k1) {
        uint64_t p0 = (uint64_t)M0 * c[0], p1 = (uint64_t)M1 * c[2];
        uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
        uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
        uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
        c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
    }
    __host__ __device__ static inline void gen(uint64_t seed, uint64_t step, uint32_t blk, uint32_t out[4]) {
        uint32_t c[4] = {blk, (uint32_t)step, (uint32_t)(step >> 32), 0x44726F70u /* "Drop" */};
        uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
        for (int i = 0; i < 10; ++i) { round(c, k0, k1); k0 += W0; k1 += W1; }
        out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; out[3] = c[3];
    }
};
// keep iff u32 < 0.8 * 2^32 (P = 0.8 to within 2^-32)
constexpr uint32_t KEEP_THRESHOLD = 3435973837u;
__host__ __device__ inline bool philox_keep(uint64_t seed, uint64_t step, uint32_t elem) {
    uint32_t r[4];
    Philox::gen(seed, step, elem >> 2, r);
    return r[elem & 3] < KEEP_THRESHOLD;
}

// ---- workspace layout (all offsets in bytes, 256-aligned) ---------------------------------
struct WsLayout {
    size_t ctrl;      // 64 x u32 control words
    size_t part1;     // f32 [ksplit][B][128] split-K partial sums of z1
    size_t h1d;       // f32 [B][128]  dropout(relu(z1))
    size_t h2;        // f32 [B][128]  relu(z2)
    size_t dz2;       // f32 [B][128]
    size_t dz1;       // f32 [B][128]
    size_t z3;        // f32 [B][16]   logits (10 used)
    size_t dz3;       // f32 [B][16]   dL/dlogits, zero padded
    size_t rowloss;   // f32 [B]
    size_t yb;        // u8  [B]
    size_t gpart;     // f32 [bsplit][NP] partial gradients of the batch slices (batch >= 1024 only)
    size_t fend;      // end of the plain (float / byte) arrays above
    // persistent kernel only (batch <= PERSIST_MAX_BATCH): dataflow buffers of 8-byte (fp32 value, 32-bit step tag)
    // words, double-buffered by step parity.  The region starts at the SAME offset for every batch <= 512 (right
    // behind the plain arrays of a 512-row layout) and is zeroed by ddpm_workspace_init, so that no plain array of
    // any layout ever aliases a tagged word.
    size_t ll;        // start of the region
    size_t ll_part;   // u64 [2][14 pixel tiles][B][128]  fc1 partial products
    size_t ll_act;    // u64 [2][B][LL_ACT_WORDS]         per-row dz1 | dz2 | h1d | h2 | dz3
    size_t ll_loss;   // u64 [2][B]                       per-row loss
    size_t ll_b1;     // u64 [2][128]                     fc1 bias after the update (read by the next step's rows)
    size_t ll_bytes;
    size_t total;
};
constexpr int PERSIST_MAX_BATCH = 512;
constexpr int LL_ACT_WORDS = 528;       // dz1 128 | dz2 128 | h1d 128 | h2 128 | dz3 16
// control words (u32 index into the 256-byte ctrl block at the start of the workspace)
//   0 ticket (rows kernels)   1 train flag   4,5 finished-CTA counter (u64)   6,7 'parameters ready' counter (u64)
//   8,9 steps run so far by the persistent kernel on this workspace (u64: the tag base)   12 dataflow timeout flag
constexpr int CTRL_FIN64 = 2, CTRL_READY64 = 3, CTRL_EPOCH64 = 4, CTRL_ERR32 = 12;
int fc1_ksplit(int batch);              // how many K splits G1 uses for this batch
int grads_bsplit(int batch);            // how many batch slices grads_kernel uses
size_t part1_rows_capacity(int max_batch);
WsLayout ws_layout(int max_batch);

// fc1 forward on tcgen05 tensor cores (csrc/fc1_tc.cu)
int fc1_tc_ksplit();
int dw1_tc_slices(int batch);
int launch_dw1_tc(const float* dz1, const uint8_t* images, int64_t n_samples, const int64_t* idx, int B, float* gpart,
                  cudaStream_t st);
int launch_fc1_tc(const float* W1, const uint8_t* images, int64_t n_samples, const int64_t* idx, int B, float* part,
                  cudaStream_t st);

}  // namespace ddpm
