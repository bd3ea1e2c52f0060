/* ddpmnist_b200.h — C ABI of libddpmnist_b200.so
 *
 * B200 (sm_100a) implementation of ONE hot path: the data-parallel MNIST training step of
 * Jonathanlyj/pytorch_ddp_mnist, `mnist_pnetcdf_cpu_mp.py` ("mp:" below).  The reference has no
 * FFI/plugin seam (it is pure Python on top of torch); each entry point therefore cites the
 * reference statement(s) whose work it performs, and INTEGRATION.md shows the ctypes binding a
 * maintainer of the reference would add.
 *
 * Conventions
 *   - plain C types only; every pointer named d_* is a DEVICE pointer owned by the caller
 *     (PyTorch tensors in practice), h_* is a HOST pointer.  The library allocates no device
 *     memory and keeps no pointer after a call returns (two exceptions, both explicit objects the
 *     caller creates and destroys: the ddpm_comm_* communicator of the fused data-parallel step owns
 *     one small symmetric buffer, a ddpm_feed_* host feed owns its ring of batch slots).
 *   - every GPU entry point is asynchronous and ordered on `stream` (a cudaStream_t passed as
 *     void*; NULL = legacy default stream); none of them synchronises or allocates (the host-fed loop
 *     additionally uses the feed's own copy streams).
 *   - return value: 0 = ok, non-zero = error; ddpm_last_error() returns a thread-local message.
 *   - flat parameter layout (118,272 fp32 = 473,088 B, the reference's single DDP bucket), in
 *     `model.parameters()` order, each matrix row-major [out,in] (mp:412-422):
 *         W1 [128,784] @ 0 | b1 [128] @ 100352 | W2 [128,128] @ 100480 | b2 [128] @ 116864 |
 *         W3 [10,128] @ 116992
 *     gradients use the same layout.
 *   - indices are int64 (what torch's samplers produce); labels are uint8 (what the reference's
 *     loader yields, mp:45-49).
 */
#ifndef DDPMNIST_B200_H
#define DDPMNIST_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DDPM_IN        784
#define DDPM_HID       128
#define DDPM_CLASSES   10
#define DDPM_N_PARAMS  118272
#define DDPM_OFF_W1    0
#define DDPM_OFF_B1    100352
#define DDPM_OFF_W2    100480
#define DDPM_OFF_B2    116864
#define DDPM_OFF_W3    116992

/* -------------------------------------------------------------------------------- misc */
const char* ddpm_last_error(void);
/* ABI version of this header; bumped on any signature change. */
int ddpm_abi_version(void);
/* Fills the 256-entry decode table on the host with the same arithmetic the kernels use:
 * fp32 IEEE  ((v / 255) - 0.1307f) / 0.3081f   — ToTensor + Normalize, mp:374-377. */
int ddpm_decode_lut_host(float* h_lut256);

/* -------------------------------------------------------------------------------- loader
 * K1. Replaces B calls of MNISTNetCDF.__getitem__ + default_collate + x.view(B,-1)
 * (mp:39-49, mp:454, mp:458) for a dataset that is resident in HBM.
 *   d_images  u8 [n_samples,784] (row i = the 784 bytes at begin_images + 784*i of the CDF-5 file)
 *   d_labels  u8 [n_samples]
 *   d_idx     i64[batch]  sample indices (a slice of the DistributedSampler shard, mp:390-394)
 *   d_x       f32[batch,784] out, bit-exact with the reference's transform stack
 *   d_y       u8 [batch] out
 * Fails (returns non-zero) on batch < 0 / n_samples <= 0; indices are range-checked on the
 * device: an out-of-range index writes NaN rows and sets *d_status (i32, may be NULL) to 1. */
int ddpm_gather_decode(const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                       const int64_t* d_idx, int batch, float* d_x, uint8_t* d_y,
                       int* d_status, void* stream);

/* K7. Dropout keep-mask for nn.Dropout(0.2) (mp:417), perf mode: Philox4x32-10 keyed by `seed`,
 * counter = (step, element/4).  d_mask u8[batch,128] in {0,1}, P(1) = 0.8.  Parity mode does not
 * call this: the caller supplies the oracle's mask. */
int ddpm_dropout_mask_philox(uint8_t* d_mask, int batch, uint64_t seed, uint64_t step, void* stream);

/* -------------------------------------------------------------------------------- fused step
 * Workspace: one caller-owned device buffer of ddpm_workspace_bytes(max_batch) bytes, 256-byte
 * aligned, reused by every step (control block, activations, gates, split-K partials, gradient slices, and the
 * persistent kernel's tagged dataflow words: 8.3 MB at batch 64, 24.9 MB at 512, 78 MB at 4096).
 * Call ddpm_workspace_init once after allocating it (it zeroes the control block and the tagged-word region - memory
 * from a caching allocator may hold the tagged words of an earlier workspace; stream-ordered).
 * Contents between calls are meaningful only to ddpm_step_backward, which reads what the last
 * ddpm_step_forward on the same workspace left.  One workspace serves one stream at a time. */
size_t ddpm_workspace_bytes(int max_batch);
int    ddpm_workspace_init(void* d_workspace, int max_batch, void* stream);
/* The leading part of a workspace that ddpm_step_forward / ddpm_step_backward use (control block + plain arrays); the
 * rest of ddpm_workspace_bytes() holds the persistent kernel's tagged dataflow words.  ddpm_step_forward_static_scratch's
 * built-in array is this many bytes for its row limit. */
size_t ddpm_workspace_forward_bytes(int max_batch);
/* Test aid (no counterpart in the reference): copies one of the per-layer arrays that ddpm_step_forward(PREP_BACKWARD)
 * / ddpm_step_backward left in the workspace to d_out - which = 0 h1d (dropout(relu(fc1)), mp:415-417), 1 h2
 * (relu(fc2), mp:418-419), 2 dz2, 3 dz1 (f32 [batch][128]), 4 dlogits (f32 [batch][16], 10 used; the logits
 * themselves leave through ddpm_step_forward's d_logits).  The parity
 * tests compare them with the reference's own per-layer activations (tests/golden/traj_ws1.npz acts0.*). */
int ddpm_workspace_peek(const void* d_workspace, int batch, int which, float* d_out, void* stream);

#define DDPM_FWD_PREP_BACKWARD 1u  /* also compute dL/dz2, dL/dz1 from the CE gradient (training) */
#define DDPM_FWD_PHILOX_MASK   2u  /* draw the dropout mask in-kernel (seed, step); d_mask unused */
#define DDPM_FWD_TF32          4u  /* fc1 on the tcgen05 tensor cores: TF32 operands, FP32 accumulate in TMEM
                                      (csrc/fc1_tc.cu).  Results agree with the fp32 path to ~1e-3 relative
                                      instead of 2e-5; worthwhile for large batches only.  Needs d_x == NULL. */

/* Forward of create_model() + CrossEntropyLoss (mp:412-422, mp:460-461; eval form mp:476-477).
 * Input rows come either from the resident u8 dataset through d_idx (d_x == NULL; the decode is
 * fused into the fc1 operand load) or from an fp32 batch d_x[batch,784] (d_images/d_idx unused).
 *   d_params   f32[118272]
 *   d_mask     u8[batch,128] dropout keep mask; NULL (and no PHILOX flag) = eval mode
 *              (model.eval(), mp:470: dropout is the identity)
 *   d_y        u8[batch] labels; if NULL and d_idx is given, labels are gathered from d_labels;
 *              if no labels are available the loss outputs are not written
 *   d_logits   f32[batch,10] out (may be NULL)
 *   d_loss     f32[2] (may be NULL): [0] = mean loss of this batch (nn.CrossEntropyLoss default);
 *              [1] += loss/batch — the reference's epoch bookkeeping `epoch_loss += batch_loss /
 *              x.shape[0]` (mp:465, 481) accumulated on the device, so no per-step host sync is
 *              needed; the caller zeroes d_loss[1] at the start of an epoch. */
int ddpm_step_forward(const float* d_params,
                      const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                      const int64_t* d_idx, const float* d_x, const uint8_t* d_y,
                      const uint8_t* d_mask, uint64_t philox_seed, uint64_t philox_step,
                      int batch, unsigned flags,
                      float* d_logits, float* d_loss, void* d_workspace, void* stream);

/* Backward of the same step (batch_loss.backward(), mp:462) for the batch last given to
 * ddpm_step_forward with this workspace (same d_images/d_idx/d_x must be passed again: the input
 * is re-decoded for the fc1 weight gradient instead of being stored).
 * Writes dL/dparams into d_grads f32[118272] (may be NULL when lr != 0) scaled by grad_scale
 * (1/world_size folds DDP's pre-division, c10d reducer semantics behind mp:439; 1.0 otherwise).
 * If d_dlogits != NULL it is used as dL/dlogits f32[batch,10] (autograd path: any upstream loss);
 * otherwise the forward must have been called with DDPM_FWD_PREP_BACKWARD.
 * If lr != 0 the SGD update p <- fma(-lr, g, p) (optimizer.step(), mp:463) is fused into the
 * gradient epilogue (single-GPU fast path; d_params is written).
 * flags: DDPM_BWD_TF32 computes fc1's weight gradient dW1 = dz1^T . X with tcgen05 TF32 tensor-core MMAs
 * (batch >= 2048 and the resident uint8 dataset path only; ignored otherwise); ~1e-3 relative. */
#define DDPM_BWD_TF32 1u
/* Two-part backward (batch >= 1024, data-parallel runs): DDPM_BWD_PART_SMALL computes everything except fc1's weight
 * gradient (bucket elements 100352..118271: b1, W2, b2, W3), DDPM_BWD_PART_W1 computes only fc1's weight gradient
 * (elements 0..100351).  Called in that order they equal one plain call, and the all-reduce of the small part (71,680 B)
 * can run while the second call computes - the overlap of the bucket all-reduce with backward that
 * DistributedDataParallel's hooks give the reference (mp:439, 462). */
#define DDPM_BWD_PART_SMALL 2u
#define DDPM_BWD_PART_W1 4u
int ddpm_step_backward(float* d_params,
                       const uint8_t* d_images, int64_t n_samples,
                       const int64_t* d_idx, const float* d_x,
                       const float* d_dlogits, int batch, float grad_scale, float lr, unsigned flags,
                       float* d_grads, void* d_workspace, void* stream);

/* Persistent step kernel (small-batch fast path): ONE cooperative launch runs `nsteps` complete
 * optimisation steps — the loop body mp:455-463 repeated — with the fc1 weights held in shared
 * memory across steps (see csrc/persist_kernel.cu).  Same arithmetic contract as
 * ddpm_step_forward(PREP_BACKWARD) + ddpm_step_backward, other summation order.
 *   d_idx        i64[nsteps*batch]   consecutive batches of the shard (the DataLoader's chunks, mp:396-401)
 *   d_masks      u8[nsteps][batch][128] injected dropout masks, or NULL with DDPM_FWD_PHILOX_MASK
 *                (step s uses Philox counter philox_step0 + s), or NULL without the flag = no dropout
 *   d_grads      f32[118272] or NULL: gradient bucket scaled by grad_scale (DDP: 1/world_size)
 *   lr           != 0: SGD fused, d_params updated after every step; == 0: gradients only
 *   d_loss       f32[2] as in ddpm_step_forward (may be NULL); d_loss_steps f32[nsteps] per-step
 *                mean loss (may be NULL)
 * batch must be <= 512 (larger batches are compute-bound: use the step_forward/backward pair).
 * The device must be able to co-schedule the kernel's 148 blocks, one per SM (any B200 does). */
int ddpm_train_steps_persistent(float* d_params, float* d_grads,
                                const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                                const int64_t* d_idx, const uint8_t* d_masks,
                                uint64_t philox_seed, uint64_t philox_step0,
                                int batch, int nsteps, unsigned flags, float grad_scale, float lr,
                                float* d_loss, float* d_loss_steps, void* d_workspace, void* stream);

/* Measurement aid (no counterpart in the reference): `nlaunches` cooperative launches of an EMPTY kernel with the
 * step kernel's launch shape (148 blocks x 256 threads, same dynamic shared memory) on `stream`.  bench.py times
 * it as the launch floor that SURVEY.md §8(d) asks to be reported beside the step time. */
int ddpm_persist_launch_floor(int nlaunches, void* stream);
/* Dataflow time-out flag of a workspace (no counterpart in the reference): the persistent kernel passes data between
 * its phases as tagged words and waits for them with a ~3 s bound; a wait that gave up sets a sticky flag in the
 * workspace's control block (and every later wait gives up at once, so a lost producer cannot hold the GPU).
 * *h_err != 0 means the results of the launches since ddpm_workspace_init are not to be trusted.  Blocking (D2H). */
int ddpm_workspace_error(const void* d_workspace, int* h_err);

/* -------------------------------------------------------------------------------- fused data-parallel step
 * DistributedDataParallel's gradient averaging (mp:439: one 473,088-byte bucket, pre-divided by
 * world_size, summed over ranks) + optimizer.step() (mp:463), done INSIDE the persistent step
 * kernel over NVLink peer memory with a low-latency "flag in the data" protocol: the thread that
 * produced a gradient element stores it as an 8-byte (value, step) word straight into every peer's
 * receive buffer, then polls the same element of every source in its own buffer, sums in rank order
 * (bit-identical replicas) and applies the SGD update in place.  No fences, no collective library
 * call and no kernel launch between steps.
 *
 * A communicator owns one symmetric device buffer (2.0 MB per rank of the job; besides a host feed's ring, below,
 * the only device memory this library allocates) that every peer process maps with CUDA IPC.  One process per GPU, all GPUs in one
 * NVLink/P2P domain, world <= 8.  Set-up (host side, blocking):
 *   ddpm_comm_create(rank, world, &c)            on every rank (its GPU current)
 *   ddpm_comm_ipc_handle(c, handle64)            64-byte handle of the local buffer
 *   -- exchange the handles (e.g. torch.distributed.all_gather_object) --
 *   ddpm_comm_connect(c, handles[world][64])     maps every peer's buffer
 * Every rank must then call ddpm_train_steps_fused_ddp with the same batch-independent arguments
 * (nsteps, lr) in the same order.  A peer that never shows up makes the waiting kernels time out
 * after ~3 s and set an error word readable with ddpm_comm_error (no GPU hang). */
typedef struct ddpm_comm ddpm_comm;
int  ddpm_comm_create(int rank, int world, ddpm_comm** out);
int  ddpm_comm_ipc_handle(const ddpm_comm* c, void* h_out64);
int  ddpm_comm_connect(ddpm_comm* c, const void* h_handles);
int  ddpm_comm_error(const ddpm_comm* c, int* h_err);
void ddpm_comm_destroy(ddpm_comm* c);
/* Same arguments as ddpm_train_steps_persistent minus d_grads/grad_scale (the bucket is the
 * communicator's; gradients are scaled by 1/world).  d_loss reports THIS rank's batch loss, as the
 * reference's per-rank `batch_loss` does (mp:461).
 *   d_idx      i64[nsteps*batch]: this rank's batches (may be NULL when d_idx_all is given)
 *   d_idx_all  i64[world][idx_stride], row r = rank r's batches of this launch (DistributedSampler shards are a
 *              pure function of (seed, epoch, rank), so every rank can compute all of them), or NULL.
 *              When given, batch <= 64 and world <= 4 (beyond that the extra multiply work outweighs the saved
 *              traffic), fc1's gradient (85 % of the bucket) is not exchanged at all: ranks
 *              exchange their dz1 rows (32 KB) and each forms the global-batch dW1 tile itself from the
 *              resident dataset — same sums in the same (rank) order on every rank. */
/* Measurement / test aid (no counterpart in the reference): the gradient exchange of the fused step ALONE - the same
 * threads exchange the same elements of the 473,088-byte bucket as in ddpm_train_steps_fused_ddp, `nsteps` times, with
 * no forward/backward work.  Rank r contributes (r + 1) * (1 + (e & 1023)) / 1024 + (step & 255) to element e; with
 * d_out != NULL (f32[118272], 16-byte aligned) the reduced values of the last step are stored there.  Collective:
 * every rank calls it with the same nsteps. */
int ddpm_comm_exchange_only(ddpm_comm* comm, int nsteps, float* d_out, void* stream);
int ddpm_train_steps_fused_ddp(ddpm_comm* comm, float* d_params,
                               const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                               const int64_t* d_idx, const int64_t* d_idx_all, int64_t idx_stride,
                               const uint8_t* d_masks,
                               uint64_t philox_seed, uint64_t philox_step0,
                               int batch, int nsteps, unsigned flags, float lr,
                               float* d_loss, float* d_loss_steps, void* d_workspace, void* stream);

/* -------------------------------------------------------------------------------- host-fed training loop
 * The reference's inner loop with a HOST-side loader (mp:452-466): every step the batch arrives in
 * host memory, is copied to the device, trained on, and the loss is read back:
 *     for x, y in train_loader:  x, y = x.to(device), y.to(device) ... batch_loss.item()
 * Here the whole loop is one call.  A feed owns a ring of `slots` batch slots in device memory, a
 * few copy streams and a sequence word per slot.  ddpm_train_steps_host_fed launches the persistent
 * step kernel for up to slots/2 steps at a time and, while it runs, streams every step's u8 pixels
 * and labels from the caller's PINNED host buffers into the ring (two cudaMemcpyAsync per step;
 * step t on copy stream t % 4, since one stream's per-operation latency would cap the rate)
 * followed by the slot's 8-byte sequence word (cuStreamWriteValue64: ordered after the copies and
 * preceded by a system-wide fence); the kernel waits for the sequence word of a step before it
 * reads that step's pixels (bounded ~3 s, then it sets the feed's error word and carries on: no
 * GPU hang).  Ring halves are recycled with events (the copy streams wait for the kernel that last
 * read the half).  Each step's mean loss is stored by the kernel straight into the caller's
 * pinned host array (h_loss_steps; mapped host memory, one 4-byte PCIe write per step), so the
 * host->device copy of the inputs and the device->host delivery of the result happen every step
 * without a launch or a synchronisation per step.
 *   batch <= 512; slots: power of two, 4 <= slots <= 4096
 *   h_pixels  u8[nsteps][batch][784], h_labels u8[nsteps][batch]: pinned (cudaHostAlloc /
 *             cudaHostRegister) host memory; must stay valid until `stream` has passed this call
 *   comm      NULL (single GPU) or a connected communicator: the gradient bucket is then averaged
 *             inside the kernel exactly as in ddpm_train_steps_fused_ddp with d_idx_all == NULL
 *   h_loss_steps f32[nsteps] pinned host memory or NULL; d_loss as in ddpm_step_forward
 * Arithmetic and results are those of ddpm_train_steps_persistent on the same batches (bit-exact). */
typedef struct ddpm_feed ddpm_feed;
int  ddpm_feed_create(int batch, int slots, ddpm_feed** out);
int  ddpm_feed_error(const ddpm_feed* f, int* h_err);
void ddpm_feed_destroy(ddpm_feed* f);
int  ddpm_train_steps_host_fed(ddpm_feed* feed, ddpm_comm* comm, float* d_params,
                               const uint8_t* h_pixels, const uint8_t* h_labels,
                               const uint8_t* d_masks, uint64_t philox_seed, uint64_t philox_step0,
                               int nsteps, unsigned flags, float lr,
                               float* d_loss, float* h_loss_steps, void* d_workspace, void* stream);

/* -------------------------------------------------------------------------------- BASELINE.json "ConvNet"
 * SURVEY.md §8(a) rows J1-J5.  NOT IN THE REFERENCE (the reference defines only the MLP
 * above); these entry points implement the layer list BASELINE.json names behind the same boundary
 * (resident u8 dataset + int64 sample indices in, loss out) and are checked against
 * oracle/convnet_oracle.py, a restatement - parity unpinned.
 *   conv1 1->32 3x3 + ReLU; conv2 32->64 3x3 + ReLU; max_pool2d(2); dropout(.25); flatten (c, y, x);
 *   fc1 9216->128 + ReLU; dropout(.5); fc2 128->10; log_softmax; nll_loss (mean)        -- J1-J3
 *   backward of all of it into one flat gradient bucket (4,799,528 B)                   -- J4
 *   Adadelta (lr 1.0, rho 0.9, eps 1e-6), one fused pass over the flat arrays           -- J5
 * Flat parameter layout = state_dict order: conv1.weight [32,1,3,3], conv1.bias, conv2.weight
 * [64,32,3,3], conv2.bias, fc1.weight [128,9216], fc1.bias, fc2.weight [10,128], fc2.bias. */
#define DDPM_CN_N_PARAMS     1199882
#define DDPM_CN_OFF_CONV1_W  0
#define DDPM_CN_OFF_CONV1_B  288
#define DDPM_CN_OFF_CONV2_W  320
#define DDPM_CN_OFF_CONV2_B  18752
#define DDPM_CN_OFF_FC1_W    18816
#define DDPM_CN_OFF_FC1_B    1198464
#define DDPM_CN_OFF_FC2_W    1198592
#define DDPM_CN_OFF_FC2_B    1199872
#define DDPM_CN_MAX_BATCH    8192
size_t ddpm_cn_workspace_bytes(int max_batch);               /* 0 if max_batch is out of range; ~0.66 MB per image + 11 MB */
int ddpm_cn_workspace_init(void* d_workspace, int max_batch, void* stream);
/* Forward (+ loss).  d_mask1 u8[batch][9216] / d_mask2 u8[batch][128]: injected dropout keep masks
 * (both or neither); DDPM_FWD_PHILOX_MASK: masks from Philox4x32-10 at (seed, step); neither: eval
 * mode.  DDPM_FWD_PREP_BACKWARD keeps what ddpm_cn_step_backward needs.  DDPM_FWD_TF32: conv2 (89 % of
 * the forward flops) as an implicit GEMM on the tcgen05 tensor cores - TF32 operands, fp32 accumulate,
 * pooling fused into the TMEM epilogue; ~1e-3 relative.  d_logprobs f32[batch][10] or NULL; d_loss
 * f32[2] as in ddpm_step_forward. */
int ddpm_cn_step_forward(const float* d_params,
                         const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                         const int64_t* d_idx, const uint8_t* d_mask1, const uint8_t* d_mask2,
                         uint64_t philox_seed, uint64_t philox_step, int batch, unsigned flags,
                         float* d_logprobs, float* d_loss, void* d_workspace, void* stream);
/* Backward of the step whose forward just ran with DDPM_FWD_PREP_BACKWARD on the same workspace:
 * d_grads f32[1199882] = grad_scale x gradient of the mean loss (grad_scale = 1/world_size under
 * data parallelism, then allreduce(SUM) the bucket).
 * flags: DDPM_CN_BWD_TRAIN = the forward ran in training mode (with dropout); DDPM_CN_BWD_TF32 =
 * conv2's data and weight gradients as implicit GEMMs on the tcgen05 tensor cores (TF32 operands, fp32
 * accumulate), like DDPM_FWD_TF32 for the forward. */
#define DDPM_CN_BWD_TRAIN 1u
#define DDPM_CN_BWD_TF32  2u
/* with DDPM_CN_BWD_TF32: this step's forward ran with DDPM_FWD_TF32 on the same workspace, so its TF32 copies of
 * conv2.weight and of conv1's output are reused instead of rebuilt (saves two small kernels) */
#define DDPM_CN_BWD_FWD_WAS_TF32 4u
int ddpm_cn_step_backward(const float* d_params, const uint8_t* d_images, int64_t n_samples,
                          const int64_t* d_idx, int batch, unsigned flags, float grad_scale,
                          float* d_grads, void* d_workspace, void* stream);
/* torch.optim.Adadelta's update on flat arrays (square_avg and acc_delta start at zero); the gradient
 * is multiplied by grad_scale first.  Every operation individually rounded as on the CPU. */
int ddpm_adadelta_step(float* d_params, const float* d_grads, float* d_square_avg, float* d_acc_delta,
                       int64_t n, float lr, double rho, double eps, float grad_scale, void* stream);

/* Forward/eval step for callers that cannot allocate.  No counterpart in the reference (added for
 * embedding; the reference is a script that allocates freely through torch).
 *
 * Who needs it: a host that hands a plug-in its input and output buffers and nothing else — a
 * validation hook inside another framework's pipeline whose memory pool is closed to plug-ins, a C
 * harness that must not call cudaMalloc once running because cudaMalloc synchronises the device,
 * a process held to a fixed device-memory budget.  Such a caller has no workspace to give
 * ddpm_step_forward, so this variant keeps the scratch inside the library: a fixed
 * 3,490,816-byte device array (sized for DDPM_STATIC_SCRATCH_MAX_BATCH rows) that every GPU which
 * loads the library holds until the process exits.  The call path performs no cudaMalloc/cudaFree
 * and uses no allocating library routine; inputs and outputs are still the caller's.
 *
 * Same arguments and results as ddpm_step_forward, minus d_workspace, with these limits:
 *   - batch > DDPM_STATIC_SCRATCH_MAX_BATCH is rejected (non-zero return, message says so);
 *   - forward/eval only: DDPM_FWD_PREP_BACKWARD is rejected, because ddpm_step_backward reads a
 *     caller-owned workspace;
 *   - the first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress, with that device current); make that first call before starting any
 *     stream capture;
 *   - all calls on a device share the one array, which still holds the previous call's contents;
 *     the entry point re-zeroes the control words it depends on at every call, but two calls
 *     that overlap on the GPU (different streams, or different host threads) race on it —
 *     serialise them. */
#define DDPM_STATIC_SCRATCH_MAX_BATCH 128
int ddpm_step_forward_static_scratch(const float* d_params,
                      const uint8_t* d_images, const uint8_t* d_labels, int64_t n_samples,
                      const int64_t* d_idx, const float* d_x, const uint8_t* d_y,
                      const uint8_t* d_mask, uint64_t philox_seed, uint64_t philox_step,
                      int batch, unsigned flags,
                      float* d_logits, float* d_loss, void* stream);

/* K4 alone. nn.CrossEntropyLoss()(logits, y) forward + gradient (mp:444, 461): d_loss[0] = mean
 * NLL of log_softmax; d_dlogits = (softmax - onehot)/batch (may be NULL).  y is uint8. */
int ddpm_cross_entropy(const float* d_logits, const uint8_t* d_y, int batch,
                       float* d_loss, float* d_dlogits, void* stream);

/* -------------------------------------------------------------------------------- optimizer
 * K6. optim.SGD(lr).step() over the flat buffer (mp:443, 463; torch/optim/sgd.py
 * `param.add_(grad, alpha=-lr)`): p[i] = fma(-lr, grad_scale * g[i], p[i]).  With grad_scale == 1
 * this is bit-exact with the reference's update. */
int ddpm_sgd_step(float* d_params, const float* d_grads, int64_t n, float lr, float grad_scale,
                  void* stream);

/* -------------------------------------------------------------------------------- file reader
 * Host-side CDF-5 ("64BIT_DATA") reader: the subset of PnetCDF the path uses
 * (pncpy.File(path,'r') mp:31, variables[...].shape mp:37, per-sample gets mp:43/46 — served here
 * as row-range reads so a rank can pull a contiguous slab straight into pinned memory). */
typedef struct ddpm_cdf5 ddpm_cdf5;
int  ddpm_cdf5_open(const char* path, ddpm_cdf5** out);
void ddpm_cdf5_close(ddpm_cdf5* f);
/* ndims <= 8; dims[] receives the shape; *elem_size the element size in bytes; *begin the
 * variable's file offset. Returns non-zero if the variable does not exist. */
int  ddpm_cdf5_inq_var(const ddpm_cdf5* f, const char* name, int* ndims, int64_t* dims,
                       int* nc_type, int* elem_size, int64_t* begin);
/* Reads rows [row0, row0+nrows) of a variable (row = index along the first dimension) into
 * h_dst with positioned reads; thread-safe for concurrent calls on one handle. */
int  ddpm_cdf5_read_rows(const ddpm_cdf5* f, const char* name, int64_t row0, int64_t nrows,
                         void* h_dst);

#ifdef __cplusplus
}
#endif
#endif /* DDPMNIST_B200_H */
