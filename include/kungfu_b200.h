/* kungfu_b200 device library — C ABI (libkungfu_cuda.so).
 *
 * Every entry point returns 0 on success or a negative KFB_ERR_* code; the
 * message for the last failure on the calling thread is kfb_last_error().
 * All launches are asynchronous on the given stream unless stated otherwise.
 *
 * Reference parity notes cite files under /root/reference/.
 */
#ifndef KUNGFU_B200_H
#define KUNGFU_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define KFB_MAX_RANKS 16  /* one NVSwitch domain; B200 HGX has 8 */
#define KFB_MAX_BLOCKS 128 /* persistent comm CTAs per launch (<= 148 SMs) */
/* Layout of a rank's signal pad in symmetric memory (uint32 words):
 *   [0, FLAG_WORDS)                 barrier flags   [block][src_rank]
 *   [FLAG_WORDS, 2*FLAG_WORDS)      exchanged f32   [block][src_rank]
 *   [2*FLAG_WORDS, +64)             seqlock / misc words
 */
#define KFB_PAD_FLAG_WORDS (KFB_MAX_BLOCKS * KFB_MAX_RANKS)
#define KFB_PAD_MISC_WORD (2 * KFB_PAD_FLAG_WORDS)
#define KFB_PAD_BYTES 32768
/* local (non-peer) per-rank state: epoch counters + reduction scratch */
#define KFB_STATE_EPOCH_WORDS KFB_MAX_BLOCKS
#define KFB_SCRATCH_FLOATS (KFB_MAX_BLOCKS + 8)
#define KFB_FUSE_SCRATCH_FLOATS 1024
#define KFB_VAR_WORK_WORDS (4 * KFB_MAX_BLOCKS + 8)

enum kfb_error {
    KFB_OK          = 0,
    KFB_ERR_ARG     = -1,
    KFB_ERR_CUDA    = -2,
    KFB_ERR_UNSUPP  = -3,
    KFB_ERR_DRIVER  = -4,
};

enum kfb_dtype {
    KFB_F32  = 0,
    KFB_BF16 = 1,
    KFB_F16  = 2,
    KFB_I32  = 3,
    KFB_F64  = 4,
    KFB_I64  = 5,
    KFB_U8   = 6,
};

/* same numeric values as KungFu_Op (srcs/cpp/include/kungfu/op.h:8-13) */
enum kfb_op { KFB_OP_SUM = 0, KFB_OP_MIN = 1, KFB_OP_MAX = 2, KFB_OP_PROD = 3 };

enum kfb_algo {
    KFB_ALGO_ONESHOT = 0, /* every rank pulls all peers, reduces locally     */
    KFB_ALGO_TWOSHOT = 1, /* pull-reduce own shard, push result to all peers */
    KFB_ALGO_NVLS    = 2, /* multimem.ld_reduce + multimem.st (switch math)  */
};

/* A communicator as the kernels see it. rank < 0 selects single-GPU
 * emulation: one cooperative launch whose blockIdx.y plays the rank, used by
 * tests when there are fewer GPUs than ranks (all "peer" pointers are then
 * local). */
typedef struct kfb_comm {
    int32_t rank;
    int32_t np;
    void *pads[KFB_MAX_RANKS]; /* peer-mapped signal pads, KFB_PAD_BYTES each */
    void *local_epoch; /* uint32[KFB_STATE_EPOCH_WORDS] (emulation: x np)     */
    void *scratch;     /* float[KFB_SCRATCH_FLOATS]      (emulation: x np)    */
    void *mc_pad;      /* multicast alias of the pads (NVLS) or NULL          */
} kfb_comm;

const char *kfb_last_error(void);
int kfb_device_info(int device, int *sm_count, int *cc_major, int *cc_minor,
                    int *multicast_supported);

/* ---- collectives over peer memory (replace K1-K3, K8-K12 of SURVEY §2.7;
 *      reference: srcs/cpp/src/nccl/gpu_collective.cpp:96-127,
 *      srcs/go/kungfu/session/session.go:98-286) ---------------------------- */

/* out = scale * reduce_op(in[0..np)) elementwise. in[r]/out[r] are the
 * peer-mapped addresses of rank r's buffers (same offset in every rank's
 * symmetric arena). ONESHOT writes out[rank] only (may be any local buffer);
 * TWOSHOT/NVLS write every out[r] and allow out == in. count is in elements;
 * buffers must be 16-byte aligned and padded to a 16-byte multiple.
 * sumsq_out (device float*, may be NULL, SUM only): receives ||out||^2 summed
 * in a fixed order (bit-identical on every rank and run to run).
 * mc_in/mc_out: multicast addresses (NVLS only). */
int kfb_allreduce(const kfb_comm *comm, int algo, void *const *in,
                  void *const *out, void *mc_in, void *mc_out, int64_t count,
                  int dtype, int op, float scale, float *sumsq_out, int blocks,
                  int threads, void *stream);

/* Rank `root` stores its buffer into every peer's out (push fan-out, or one
 * multimem.st when mc_out != NULL). Byte-wise. (session.go:158-161) */
int kfb_broadcast(const kfb_comm *comm, int root, const void *in,
                  void *const *out, void *mc_out, int64_t bytes, int blocks,
                  int threads, void *stream);

/* Every rank stores its `bytes` block into slot `rank` of every peer's out.
 * (srcs/go/kungfu/session/allgather.go:17-45) */
int kfb_allgather(const kfb_comm *comm, const void *in, void *const *out,
                  int64_t bytes, int blocks, int threads, void *stream);

/* Root pulls and reduces (reduce) or pulls and concatenates (gather).
 * (session.go:153-207) */
int kfb_reduce(const kfb_comm *comm, int root, void *const *in, void *out,
               int64_t count, int dtype, int op, int blocks, int threads,
               void *stream);
int kfb_gather(const kfb_comm *comm, int root, void *const *in, void *out,
               int64_t bytes, int blocks, int threads, void *stream);

/* Device-side barrier (K11) and byte consensus (K12): *agree_out (device
 * int32) = 1 iff every rank's `bytes` of `in` are identical. */
int kfb_barrier(const kfb_comm *comm, void *stream);
int kfb_consensus(const kfb_comm *comm, void *const *in, int64_t bytes,
                  int32_t *agree_out, int blocks, int threads, void *stream);

/* One edge of a strategy graph (K8): dst = op(dst, src) or dst = src where
 * src is a peer-mapped pointer. Used by the stream-ordered graph executor
 * for ring/tree/star/binary-tree/set_tree strategies. No in-kernel waiting:
 * ordering comes from events + kfb_barrier on the executor's stream. */
int kfb_edge_reduce(void *dst, const void *src, int64_t count, int dtype,
                    int op, int blocks, int threads, void *stream);
int kfb_edge_copy(void *dst, const void *src, int64_t bytes, int blocks,
                  int threads, void *stream);

/* A whole strategy walk (reduce graph then broadcast graph, several graph pairs
 * over chunks of the buffer) in ONE launch: in round r this rank pulls from
 * edge_src[e] for e in [round_begin[r], round_begin[r+1]) into chunk
 * edge_plan[e] of its own buffer, reducing onto it while r < n_reduce_rounds
 * and overwriting it afterwards; peer barriers separate the rounds. bufs[r]:
 * rank r's buffer as mapped here; chunk i covers chunk_bytes[i] bytes at
 * chunk_lo_bytes[i] (16-byte aligned). The tables are this rank's view,
 * compiled by the host runtime (kf_device_plan, include/kungfu_host.h) from the
 * strategy's graphs.
 * (srcs/go/kungfu/session/session.go:218-313) */
int kfb_graph_allreduce(const kfb_comm *comm, void *const *bufs, int dtype,
                        int op, int n_plans, const int64_t *chunk_lo_bytes,
                        const int64_t *chunk_bytes, int n_rounds,
                        int n_reduce_rounds, const int32_t *round_begin,
                        const int32_t *edge_plan, const int32_t *edge_src,
                        int blocks, int threads, void *stream);

/* ---- multi-tensor helpers (replace K13 fuse/defuse + K17 local norm) ----- */

/* Copy+cast n tensors into one flat buffer (the symmetric bucket) and
 * accumulate sum(x^2) over everything copied into *sumsq_out (deterministic).
 * srcs[i] has counts[i] elements of src_dtype, lands at dst + dst_offs[i]
 * (elements of dst_dtype). Host arrays; any n (chunked launches of 64).
 * scratch: device float[KFB_FUSE_SCRATCH_FLOATS], needed iff sumsq_out.
 * (srcs/python/kungfu/tensorflow/ops/__init__.py:29-46) */
int kfb_fuse(void *const *srcs, const int64_t *counts, const int64_t *dst_offs,
             int n, int src_dtype, void *dst, int dst_dtype, float *sumsq_out,
             float *scratch, void *stream);
/* kfb_fuse without a caller-supplied scratch buffer, for callers that cannot
 * allocate device memory where they call us (code running inside another
 * framework's memory-planned step, latency-critical hooks where cudaMalloc's
 * device-wide synchronisation is not acceptable, threads barred from the
 * allocator). This entry point has no counterpart in the reference. The block
 * partials of the norm live in a static device array inside the library
 * (64 KiB per GPU that loads it, held until exit), so:
 *  - n <= KFB_FUSE_NOALLOC_MAX_TENSORS per call, larger n is rejected;
 *  - it never calls cudaMalloc/cudaFree or any allocating routine;
 *  - the first call on a device looks the array's address up
 *    (cudaGetSymbolAddress) - issue that call before any stream capture;
 *  - all calls on a device share the array: two calls whose kernels can
 *    overlap on the GPU (different streams, no ordering) race on it; calls
 *    ordered on one stream are fine. Nothing in it needs clearing: every slot
 *    that is read was written by the same call.
 * Inputs, outputs and semantics are those of kfb_fuse; the norm is summed in
 * one fixed order (equal to kfb_fuse within fp32 rounding). Unlike kfb_fuse it
 * does not cut a very large tensor into pieces first (that needs a host-side
 * list whose length depends on the sizes): every tensor is streamed by 16 CTAs,
 * so a single gradient of tens of millions of elements is copied well below
 * HBM speed; hand such tensors to kfb_fuse, or split them in the caller. */
#define KFB_FUSE_NOALLOC_MAX_TENSORS 1024
int kfb_fuse_noalloc(void *const *srcs, const int64_t *counts,
                     const int64_t *dst_offs, int n, int src_dtype, void *dst,
                     int dst_dtype, float *sumsq_out, void *stream);
/* out[0] = slots[0] + ... + slots[n-1], added in index order by one block:
 * how per-bucket norm partials become one reproducible number (K17). */
int kfb_sum_slots(const float *slots, int n, float *out, void *stream);
/* Inverse: scatter flat buffer back into n tensors (cast dst_dtype<-src). */
int kfb_defuse(const void *src, int src_dtype, void *const *dsts,
               const int64_t *counts, const int64_t *src_offs, int n,
               int dst_dtype, void *stream);

/* ---- fused optimizer steps ------------------------------------------------ */

typedef struct kfb_sgd_hyper {
    float lr;
    float momentum;
    float dampening;
    float weight_decay;
    int32_t nesterov;
    int32_t first_step; /* momentum buffer := grad (torch.optim.SGD rule) */
    /* Update rule. 0: torch.optim.SGD (fields above). 1: torch.optim.AdamW
     * (decoupled weight decay), 2: torch.optim.Adam (L2 decay). For 1 and 2
     * the `mom` argument of the step functions is the first moment and
     * exp_avg_sq the second (same layout); the step number t of the bias
     * corrections 1 - beta^t is *step + 1 when `step` (a DEVICE int64 the
     * caller advances once per training step; keeps a captured CUDA graph
     * valid across replays) is given, otherwise bias_corr1/2 are used. */
    int32_t kind;
    float beta1, beta2, eps;
    float bias_corr1, bias_corr2;
    const void *step;
    float *exp_avg_sq;
} kfb_sgd_hyper;

/* S-SGD fused step (K1+K14+optimizer): reduce-scatter the gradient shard over
 * peer memory, scale by grad_scale, apply SGD(momentum, wd) to the fp32 master
 * shard this rank owns, and push the updated parameters (param_dtype) into
 * every peer's flat parameter buffer. np==1 degenerates to a local fused SGD.
 * grads[r]/params[r]: peer-mapped flat buffers. master/mom: local fp32 flat
 * buffers (only this rank's shard is live). sumsq_out: ||avg grad||^2.
 * (srcs/python/kungfu/tensorflow/optimizers/sync_sgd.py:78-109) */
int kfb_ssgd_fused_step(const kfb_comm *comm, void *const *grads,
                        int grad_dtype, void *const *params, int param_dtype,
                        float *master, float *mom, int64_t count,
                        float grad_scale, const kfb_sgd_hyper *h,
                        float *sumsq_out, int blocks, int threads,
                        void *stream);

/* Split-phase entry barrier for steps that run UNDER other work (bucketed,
 * overlapped S-SGD). kfb_peer_signal (one tiny CTA) sends this rank's arrival
 * for `rows` CTA rows of the communicator ("my gradients of the next bucket
 * are final"); the step launched later with KFB_STEP_PRESIGNALED only checks
 * that every peer's arrival is there instead of arriving and spinning itself.
 * KFB_STEP_NO_EXIT_BARRIER drops the barrier on the way out; the caller must
 * end the sequence with a launch that keeps it (the last bucket) before
 * anything reads the parameters. kfb_set_step_flags applies to the NEXT
 * kfb_ssgd_fused_step[_nvls] call of the calling thread only. With
 * NO_EXIT_BARRIER sumsq_out receives this rank's shard partial (sum them
 * over the ranks afterwards). `rows` must equal the step's `blocks`. */
#define KFB_STEP_PRESIGNALED 1
#define KFB_STEP_NO_EXIT_BARRIER 2
#define KFB_STEP_PARTIAL_NORM 4 /* sumsq_out := this rank's shard partial even when the exit barrier is kept */
int kfb_peer_signal(const kfb_comm *comm, int rows, void *stream);
void kfb_set_step_flags(int flags);

/* NVLS variant: the switch sums the gradient shard (multimem.ld_reduce) and
 * multicasts the updated parameters (multimem.st). mc_grads / mc_params are
 * the multicast aliases of the flat gradient / parameter buffers. np >= 2. */
int kfb_ssgd_fused_step_nvls(const kfb_comm *comm, void *mc_grads,
                             int grad_dtype, void *mc_params, int param_dtype,
                             float *master, float *mom, int64_t count,
                             float grad_scale, const kfb_sgd_hyper *h,
                             float *sumsq_out, int blocks, int threads,
                             void *stream);

/* Local flat SGD on an already-reduced gradient buffer (any-optimizer path's
 * fast sibling; also the np==1 path). */
int kfb_sgd_flat(const void *grad, int grad_dtype, void *param,
                 int param_dtype, float *master, float *mom, int64_t count,
                 float grad_scale, const kfb_sgd_hyper *h, int blocks,
                 int threads, void *stream);

/* SMA fused step (K15 + optimizer): avg = mean_r(params_r) over peer memory,
 * v <- (1-alpha) v + alpha avg, then SGD with the LOCAL gradient, one kernel.
 * avg_stage[r]: peer-mapped staging buffers (param_dtype, same size).
 * (srcs/python/kungfu/tensorflow/optimizers/sma_sgd.py:50-74) */
int kfb_sma_fused_step(const kfb_comm *comm, void *const *params,
                       int param_dtype, void *const *avg_stage,
                       const void *grad, int grad_dtype, float *master,
                       float *mom, int64_t count, float alpha,
                       float grad_scale, const kfb_sgd_hyper *h, int blocks,
                       int threads, void *stream);

/* The same step with two refinements (what the optimizer wrappers call):
 *  - mc_params / mc_stage: multicast aliases of the parameter and staging
 *    buffers; when both are given the switch does the averaging
 *    (multimem.ld_reduce + multimem.st), moving ~1x the model per GPU;
 *  - phase 1 runs only the averaging (it only READS parameters, so it may be
 *    launched at the start of a step on a side stream and overlap forward and
 *    backward), phase 2 only the blend + update (purely local); 0 = both. */
int kfb_sma_fused_step_ex(const kfb_comm *comm, void *const *params,
                          int param_dtype, void *const *avg_stage,
                          void *mc_params, void *mc_stage, const void *grad,
                          int grad_dtype, float *master, float *mom,
                          int64_t count, float alpha, float grad_scale,
                          const kfb_sgd_hyper *h, int phase, int blocks,
                          int threads, void *stream);

/* PairAveraging fused step (K16+K21): one-sided pull of the target peer's
 * published snapshot (seqlock + double buffer), v <- 0.5 (v + v_t), SGD with
 * the local gradient, publish own new snapshot. No barrier, no NCCL.
 * snap_peer: peer-mapped base of the target's 2-slot snapshot area
 * (2 * slot_bytes); snap_self: own area; seq words live in the pads.
 * torn_out (device int32, may be NULL) counts snapshots whose slot the owner
 * had begun to overwrite before the read finished (the owner announces every
 * write in a second pad word before touching the slot, so nothing is missed;
 * such a blend is kept, as in the reference, and counted). (srcs/python/kungfu/tensorflow/optimizers/async_sgd.py:110-142) */
int kfb_pair_avg_step(const kfb_comm *comm, int target, const void *snap_peer,
                      void *snap_self, int64_t slot_bytes, void *param,
                      int param_dtype, const void *grad, int grad_dtype,
                      float *master, float *mom, int64_t count,
                      float grad_scale, const kfb_sgd_hyper *h,
                      int32_t *torn_out, int blocks, int threads,
                      void *stream);
/* The same step with the target drawn ON THE DEVICE from (seed, *step, rank),
 * uniformly over the other ranks: no launch argument depends on the host's
 * choice, so the whole training step can be captured in one CUDA graph.
 * snaps[r] = rank r's snapshot area as mapped in this process (np entries);
 * `step` is a device scalar the caller advances once per step. */
int kfb_pair_avg_step_dev(const kfb_comm *comm, void *const *snaps,
                          int64_t slot_bytes, void *param, int param_dtype,
                          const void *grad, int grad_dtype, float *master,
                          float *mom, int64_t count, float grad_scale,
                          const kfb_sgd_hyper *h, int32_t *torn_out,
                          const long long *step, unsigned long long seed,
                          int blocks, int threads, void *stream);
/* Pull + blend only: param <- 0.5 (param + target's snapshot), no update and
 * no publication. Device path of PairAveraging around any torch optimizer:
 * kfb_pair_blend, the wrapped optimizer's own step, kfb_pair_publish.
 * (async_sgd.py:130-133) */
int kfb_pair_blend(const kfb_comm *comm, int target, const void *snap_peer,
                   int64_t slot_bytes, void *param, int param_dtype,
                   int64_t count, int32_t *torn_out, int blocks, int threads,
                   void *stream);
/* Copy the target's current snapshot (`bytes`, a multiple of 16) into the
 * local buffer dst; *got_out (device int32, may be NULL) := whether the target
 * had published anything yet. For models of several dtypes: one pull, then a
 * blend per dtype. */
int kfb_pair_pull(const kfb_comm *comm, int target, const void *snap_peer,
                  int64_t slot_bytes, void *dst, int64_t bytes,
                  int32_t *torn_out, int32_t *got_out, int blocks, int threads,
                  void *stream);
/* Publish the current parameters as snapshot version (seq+1) (init_store). */
int kfb_pair_publish(const kfb_comm *comm, const void *param, void *snap_self,
                     int64_t slot_bytes, int64_t bytes, int blocks,
                     int threads, void *stream);

/* ---- monitors -------------------------------------------------------------- */

/* Gradient variance (K18): one pass over every rank's gradients with two
 * accumulators. For each segment (tensor) s this rank adds, over its shard,
 * sum((E[g^2] - E[g]^2)^2) into seg_out[s] (device float[n_seg], zeroed by the
 * caller); all-reducing seg_out and summing sqrt() over s gives the
 * reference's "sum of per-tensor norms". seg_ends_dev: device int64[n_seg]
 * cumulative end offsets (elements). work: device scratch of
 * KFB_VAR_WORK_WORDS 32-bit words (x np under emulation), zeroed once by the
 * caller; sums are formed without atomics, in one fixed order.
 * (srcs/python/kungfu/tensorflow/optimizers/grad_variance.py:46-59) */
int kfb_grad_variance(const kfb_comm *comm, void *const *grads, int dtype,
                      int64_t count, const int64_t *seg_ends_dev, int n_seg,
                      float *seg_out, float *work, int blocks, int threads,
                      void *stream);

/* Gradient noise scale finalize (K17+K19) on device scalars: state[0..3] =
 * {g_ema, s_ema, initialised, last_noise_scale}.
 * (srcs/cpp/src/tensorflow/ops/cpu/collective.cpp:223-258) */
int kfb_gns_finalize(const float *sumsq_local, const float *sumsq_global,
                     float batch_small, float batch_big, float alpha,
                     float *state, void *stream);

/* ---- symmetric memory (replaces C5-C8) ------------------------------------- */

/* VMM allocation that can be exported as a POSIX fd and mapped by peers. */
typedef struct kfb_symm_block kfb_symm_block;
int kfb_symm_alloc(int device, size_t bytes, kfb_symm_block **out);
int kfb_symm_export_fd(kfb_symm_block *b, int *fd_out);
int kfb_symm_import_fd(int device, int fd, size_t bytes,
                       kfb_symm_block **out);
void *kfb_symm_ptr(kfb_symm_block *b);
size_t kfb_symm_size(kfb_symm_block *b);
int kfb_symm_free(kfb_symm_block *b);
/* Legacy CUDA-IPC fallback (64-byte handles through any byte transport). */
int kfb_ipc_alloc(int device, size_t bytes, void **ptr_out,
                  unsigned char handle_out[64]);
int kfb_ipc_open(int device, const unsigned char handle[64], void **ptr_out);
int kfb_ipc_close(void *ptr);
int kfb_ipc_free(void *ptr);
/* Multicast object over `np` devices of the node (NVLS). Rank 0 creates and
 * exports; others import; every rank adds its device, then binds its block. */
typedef struct kfb_mc_object kfb_mc_object;
int kfb_mc_create(int device, int np, size_t bytes, kfb_mc_object **out,
                  int *fd_out);
int kfb_mc_import(int device, int np, int fd, size_t bytes,
                  kfb_mc_object **out);
int kfb_mc_add_device(kfb_mc_object *mc);
int kfb_mc_bind(kfb_mc_object *mc, kfb_symm_block *mem, void **mc_ptr_out);
int kfb_mc_free(kfb_mc_object *mc);
size_t kfb_symm_granularity(int device, int multicast);
/* 1 if the device can take part in multicast objects (NVLS), else 0 */
int kfb_multicast_supported(int device);
/* close a descriptor obtained from kfb_symm_export_fd / kfb_mc_create or
 * received from a peer, once it has been imported */
int kfb_close_fd(int fd);

/* ---- model kernels (channels-last bf16 training blocks) ------------------- */

/* y = relu?(bn(x) [+ residual]) forward/backward for NHWC bf16/f16 activations
 * (dtype) with fp32 statistics. rows = N*H*W; channels % 8 == 0, <= 2048.
 * scratch: device float[kfb_bn_scratch_floats(rows, channels)], private to the
 * call. training == 0 normalises with the running statistics. relu_mask:
 * uint8[rows * channels / 8], one bit per element (y > 0), written by the
 * forward when relu != 0 and read by the backward instead of re-reading y.
 * dresidual may be NULL. */
int kfb_bn_scratch_floats(long long rows, int channels);
int kfb_bn_act_fwd(const void *x, const void *residual, void *y,
                   const float *weight, const float *bias, float *running_mean,
                   float *running_var, float *save_mean, float *save_invstd,
                   int64_t rows, int channels, float eps, float momentum,
                   int relu, int training, int dtype, float *scratch,
                   void *relu_mask, void *stream);
int kfb_bn_act_bwd(const void *dy, const void *x, const void *relu_mask,
                   const float *weight, const float *save_mean,
                   const float *save_invstd, void *dx, void *dresidual,
                   float *dweight, float *dbias, int64_t rows, int channels,
                   int relu, int dtype, float *scratch, void *stream);

/* 3x3 / stride 2 / pad 1 max pooling, NHWC bf16/f16. idx: uint8[N*OH*OW*C]
 * window slot (0..8) of the maximum, written by fwd, read by bwd (a gather:
 * deterministic, no atomics). OH = (H-1)/2+1, OW = (W-1)/2+1. */
int kfb_maxpool3x3s2_fwd(const void *x, void *y, void *idx, int N, int H, int W,
                         int C, int dtype, void *stream);
int kfb_maxpool3x3s2_bwd(const void *dy, const void *idx, void *dx, int N, int H,
                         int W, int C, int dtype, void *stream);

/* Global average pooling (N,H,W,C)->(N,C), NHWC bf16/f16, and its backward
 * (broadcast of dy/HW) written directly in channels-last layout. */
int kfb_gap_fwd(const void *x, void *y, int N, int HW, int C, int dtype,
                void *stream);
int kfb_gap_bwd(const void *dy, void *dx, int N, int HW, int C, int dtype,
                void *stream);

/* ---- tcgen05 linear layer (csrc/kernels/linear_tc.cu) ------------------------
 * y[M,N] = act(x[M,K] . w[N,K]^T + bias[N]); x, w, y (and z) bf16 row-major,
 * fp32 accumulation in tensor memory, operands staged by TMA. bias may be NULL,
 * KFB_F32 or KFB_BF16. With z != NULL the pre-activation is stored as well
 * (what kfb_act_bwd needs). K and N must be multiples of 8, pointers 16-byte
 * aligned. No counterpart in the reference (its models are TensorFlow's). */
enum kfb_act {
    KFB_ACT_NONE = 0,
    KFB_ACT_GELU = 1, /* erf form */
    KFB_ACT_RELU = 2,
};
int kfb_linear_act(const void *x, const void *w, const void *bias, int bias_dtype,
                   void *y, void *z, int M, int N, int K, int act, void *stream);
/* dz = dy * act'(z), n bf16 elements (n % 8 == 0) */
int kfb_act_bwd(const void *dy, const void *z, void *dz, size_t n, int act,
                void *stream);
/* -1: choose per shape, 0: one CTA per tile (cta_group::1), 1: CTA pairs (cta_group::2) */
void kfb_linear_act_set_mode(int mode);

/* ---- fused "add & norm" (csrc/kernels/ln_fused.cu) ---------------------------
 * y = LayerNorm(x + dropout(h)) * gamma + beta over rows of H bf16 elements,
 * H in {256, 512, 768, 1024}; all tensors bf16, mean / rstd fp32 [M]. x may be
 * NULL (no residual), s_out / mean / rstd may be NULL in inference. The dropout
 * mask is never stored: forward and backward regenerate it from
 * (seed, stream_id, *step, element index); `step` is a device scalar the caller
 * advances once per training step (NULL = 0). Backward writes dx (= ds), dh
 * (= ds * mask / (1-p); either may be NULL) and dgamma / dbeta through
 * `partial`, a scratch of kfb_ln_partial_floats(M, H) floats. No counterpart in
 * the reference (its models are TensorFlow's). */
int kfb_ln_partial_floats(long long M, int H);
int kfb_add_dropout_ln_fwd(const void *h, const void *x, const void *gamma,
                           const void *beta, void *y, void *s_out, float *mean,
                           float *rstd, long long M, int H, float eps, float p,
                           unsigned long long seed, const long long *step,
                           unsigned int stream_id, void *stream);
/* out[N] = column sums of a[M, N] (bf16; a bias gradient), N % 64 == 0. `scratch`
 * holds kfb_colsum_scratch_floats(N) floats, ZEROED once by the caller (the
 * kernel leaves its ticket counters at zero again); calls sharing a scratch
 * must be stream-ordered. Deterministic. */
int kfb_colsum_scratch_floats(int N);
int kfb_colsum(const void *a, long long M, int N, void *out, float *scratch,
               void *stream);
/* dqkv[B, S, 3, h, d] = gather of dq, dk, dv (each logically [B, h, S, d], d
 * contiguous; strides9 = {stride_b, stride_h, stride_s} x 3 in elements). */
int kfb_pack_qkv_grad(const void *dq, const void *dk, const void *dv,
                      const long long *strides9, int B, int h, int S, int d,
                      void *out, void *stream);
int kfb_add_dropout_ln_bwd(const void *dy, const void *s, const float *mean,
                           const float *rstd, const void *gamma, void *dx,
                           void *dh, void *dgamma, void *dbeta, float *partial,
                           long long M, int H, float p, unsigned long long seed,
                           const long long *step, unsigned int stream_id,
                           void *stream);

#ifdef __cplusplus
}
#endif
#endif /* KUNGFU_B200_H */
