/* libb200dp — C ABI of the sm_100a kernels and the peer-memory collective runtime.
 *
 * Every function returns 0 on success; on failure b200_last_error() describes why.
 * Pointers are device pointers unless stated; `stream` is a cudaStream_t passed as void*.
 * Activations are NHWC bf16, weights use TensorFlow layouts (conv HWIO, dense [in,out]).
 * Reference citations (path:line) are relative to the upstream repository.
 */
#ifndef B200DP_H
#define B200DP_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_MAX_RANKS 8
#define B200_MAX_CHANNELS 16
#define B200_MAX_COMM_BLOCKS 296
#define B200_MAX_RANGES 4
#define B200_MAX_SEGMENTS 32
#define B200_ADASUM_STATIC_MAX_BLOCKS 64

typedef void* b200_comm_t;

const char* b200_last_error(void);

/* ---- collective runtime (replaces Horovod core + MPI/NCCL ops; horovod/Dockerfile:64-65) ---- */
size_t b200_comm_pad_bytes(void);
int b200_comm_create(int rank, int nranks, int emulated, b200_comm_t* out);
int b200_comm_rank(b200_comm_t h);
int b200_comm_nranks(b200_comm_t h);
int b200_comm_is_emulated(b200_comm_t h);
int b200_comm_set_timeout_ms(b200_comm_t h, double ms);
/* The fused kernel launched on `channel` (the last kernel of a training step) also copies that step's slot of the
 * loss / correct-count rings into pinned host memory: the per-step result the reference prints
 * (horovod/tensorflow_mnist.py:148-149) reaches the host without a copy operation between two steps. */
/* Push inbox for the latency-bound one-shot buckets: launches of b200_fused_allreduce_adam on `channel` (blocks of at most
 * 256 threads, one-shot ranges) store each rank's gradient slice into every peer's inbox BEFORE raising the barrier flag,
 * so the peers' data arrives with the flag (one one-way NVLink latency) instead of being fetched after it (two round
 * trips).  seg: symmetric segment of 2 * bytes_per_parity bytes. */
int b200_comm_set_inbox(b200_comm_t h, int channel, int seg, size_t bytes_per_parity);
/* Fused kernels read the step's Adam coefficients from (ctrl_seg, off) (written once per step by the first kernel of the
 * step, bump_step = 2 of b200_conv1_pool_fwd / b200_conv1_valid_fwd) instead of computing two fp64 pow() each. */
int b200_comm_set_coef_cache(b200_comm_t h, int ctrl_seg, size_t off);
int b200_comm_set_step_mirror(b200_comm_t h, int channel, int ctrl_seg, size_t loss_off, size_t correct_off, int ring,
                              float* host_loss, int* host_correct);
int b200_comm_set_debug_skew_ns(b200_comm_t h, unsigned max_ns);   /* race hunting: random delay before each barrier arrival */
int b200_comm_set_debug_poison(b200_comm_t h, int on);             /* race hunting: consumed gradient buffers get NaN, not zeros */
int b200_comm_seg_alloc(b200_comm_t h, size_t bytes, int* seg_id, void* ipc_handle_out /* 64 B or NULL */);
int b200_comm_seg_open(b200_comm_t h, int seg_id, const void* all_handles /* nranks x 64 B */);
void* b200_comm_seg_ptr(b200_comm_t h, int seg_id, int peer);
int b200_comm_destroy(b200_comm_t h);

/* DistributedOptimizer's hot path (horovod/tensorflow_mnist.py:133-136): gradient allreduce fused with the
 * 1/size scale, the Adam update (horovod/tensorflow_mnist.py:130) and the bf16 weight refresh. */
int b200_fused_allreduce_adam(b200_comm_t h, int channel, int nblocks, int nthreads, int reduce,
                              int grad_seg, size_t grad_cur_off, size_t grad_other_off,
                              int state_seg, size_t master_off, size_t m_off, size_t v_off,
                              int w16_seg, size_t w16_off, int ctrl_seg, size_t step_off, size_t hyper_off,
                              float scale, int nranges, const long long* range_start, const long long* range_numel,
                              const int* range_algo, const int* range_zero_other, void* stream);
/* hvd.allreduce (Average/Sum) on a staged fp32 buffer */
int b200_allreduce(b200_comm_t h, int channel, int nblocks, int nthreads, int algo, int src_seg, size_t src_off,
                   int dst_seg, size_t dst_off, long long numel, float scale, void* stream);
/* hvd.broadcast / BroadcastGlobalVariablesHook (horovod/tensorflow_mnist.py:143) */
int b200_broadcast(b200_comm_t h, int channel, int nblocks, int nthreads, int root, int src_seg, size_t src_off,
                   int dst_seg, size_t dst_off, long long numel, void* stream);
int b200_allgather_slices(b200_comm_t h, int channel, int nblocks, int nthreads, int seg, size_t off, long long numel,
                          void* stream);
/* one-sided copy of a peer's [off, off+bytes) into the same range of the local segment (rank-0 checkpoint gather,
 * horovod/tensorflow_mnist.py:157-159: only worker 0 saves) */
int b200_comm_pull_from_peer(b200_comm_t h, int seg, size_t off, size_t bytes, int peer, void* stream);
/* The dense layer's gradient "allreduce" turned inside out (horovod/tensorflow_mnist.py:133-136 for layers.dense, :65-66):
 * every rank gathers all ranks' dz [B,D] and its own column slice of all ranks' activations a2 [B,F] (2.5 MB over NVLink at
 * N=8 instead of the 12.8 MB gradient), then b200_gemm_tc_wgrad_adam contracts over the global batch. */
int b200_allgather_acts(b200_comm_t h, int channel, int nblocks, int act_seg, size_t act_off, int dz_seg, size_t dz_off,
                        int out_seg, size_t act_out_off, size_t dz_out_off, int B, int F, int D, void* stream);
/* NVLS: segments bound to an NVSwitch multicast object (in-switch reduction). Three steps with the
 * caller's control plane in between: alloc (-> POSIX fds to ship to the peers), open (import + add device),
 * [barrier], bind.  A multicast segment is also peer-mapped, so every P2P entry point accepts it. */
int b200_nvls_supported(void);
int b200_comm_mcseg_alloc(b200_comm_t h, size_t bytes, int* seg_id, int* mem_fd_out, int* mc_fd_out, size_t* actual_bytes);
int b200_comm_mcseg_open(b200_comm_t h, int seg_id, const int* mem_fds, int mc_fd);
int b200_comm_mcseg_bind(b200_comm_t h, int seg_id);
void* b200_comm_seg_mc_ptr(b200_comm_t h, int seg_id);
int b200_allreduce_nvls(b200_comm_t h, int channel, int nblocks, int nthreads, int src_seg, size_t src_off, int dst_seg,
                        size_t dst_off, long long numel, float scale, void* stream);
int b200_broadcast_nvls(b200_comm_t h, int channel, int nblocks, int nthreads, int root, int src_seg, size_t src_off,
                        int dst_seg, size_t dst_off, long long numel, void* stream);
int b200_fused_allreduce_adam_nvls(b200_comm_t h, int channel, int nblocks, int nthreads, int grad_seg, size_t grad_cur_off,
                                   size_t grad_other_off, int state_seg, size_t master_off, size_t m_off, size_t v_off,
                                   int w16_seg, size_t w16_off, int ctrl_seg, size_t step_off, size_t hyper_off, float scale,
                                   long long start, long long numel, int zero_other, void* stream);
/* hvd.Adasum (horovod/tensorflow_mnist.py:133), one tree level per call */
size_t b200_adasum_scratch_bytes(int nblocks, int nseg);
int b200_adasum_level(b200_comm_t h, int channel, int nblocks, int level, int cur_seg, size_t cur_off, int next_seg,
                      size_t next_off, int scratch_seg, size_t scratch_off, int nseg, const long long* seg_start,
                      const long long* seg_numel, void* stream);

/* No counterpart in the reference (added for callers of this library, not taken from upstream).
 * Pairwise Adasum of two fp32 vectors, out = adasum(a, b) per segment — the same arithmetic as one
 * b200_adasum_level — for callers that may not allocate device memory: it needs no communicator and calls no
 * cudaMalloc/cudaFree; its scratch is a 48 KiB array built into the library (resident on every GPU that loads
 * it, until exit).  Limits: nblocks <= B200_ADASUM_STATIC_MAX_BLOCKS, nseg <= B200_MAX_SEGMENTS; larger requests
 * fail with an error.  All calls on a device share that scratch: run them on one stream, never overlapping (several
 * host threads may call into that one stream: the two kernels of a call are enqueued together under a lock).
 * The first call on each device resolves the scratch address (cudaGetSymbolAddress) — make it before any stream
 * capture starts. */
int b200_adasum_pair_static(const float* a, const float* b, float* out, int nblocks, int nseg,
                            const long long* seg_start, const long long* seg_numel, void* stream);

/* ---- model kernels, CUDA-core family (first conv, head, fallbacks) ---- */
/* conv 5x5 SAME Cin=1 + bias + ReLU + maxpool2 (horovod/tensorflow_mnist.py:49-53). step_ptr (may be NULL)
 * is incremented by one when bump_step != 0: the step counter Adam and dropout read; with ring > 0 the
 * new step's slot of the loss/correct rings is cleared at the same time.  bump_step == 2: step_ptr is the start of the
 * engine's control block (hyper-parameters at +16 B, beta powers at +96 B) and the step's Adam coefficients are written
 * to +128 B (see b200_comm_set_coef_cache). */
int b200_conv1_pool_fwd(const float* x, const void* w_bf16, const void* bias_bf16, void* out_bf16, uint8_t* argmax,
                        int B, int C1, int* step_ptr, int bump_step, float* loss_ring, int* correct_ring, int ring,
                        void* stream);
int b200_conv1_wgrad(const float* x, const void* da1_bf16, const void* a1_bf16, const uint8_t* argmax,
                     float* dw, float* dbias, int B, int C1, void* stream);
/* dense1 epilogue + dropout + logits + softmax-xent fwd/bwd + accuracy (horovod/tensorflow_mnist.py:65-73).
 * ring == 0: loss_sum/correct are running sums; ring > 0: they are rings indexed by *step_ptr % ring and
 * receive the batch-mean loss / correct count of that step. */
int b200_head_fwd_bwd(float* z, const void* b1_bf16, const void* w2_bf16, const void* b2_bf16, const int* labels,
                      void* h_bf16, void* dz_bf16, float* dlogits, float* logits_out, float* loss_sum, int* correct,
                      int B, int D, int NC, float drop_rate, int training, uint64_t seed, const int* step_ptr,
                      int layer_id, int zero_z, int ring, void* stream);
int b200_head_wgrad(const void* h_bf16, const void* dz_bf16, const float* dlogits, float* dw2, float* db2, float* db1,
                    int B, int D, int NC, void* stream);
/* max-pool 2x2/2 backward fused with the ReLU mask: dy[b,2h+i,2w+j,c] */
/* dbias (may be NULL) accumulates sum_{b,h,w} dy[b,h,w,c]: the bias gradient of the convolution that produced `pooled` */
int b200_unpool_relu_bwd(const void* dpool_bf16, const void* pooled_bf16, const uint8_t* argmax, void* dy_bf16,
                         float* dbias, int B, int HP, int WP, int C, void* stream);
int b200_maxpool_fwd(const void* in_bf16, void* out_bf16, uint8_t* argmax, int B, int H, int W, int C, void* stream);
/* same, when `pooled` is the tensor after an inverted dropout of keep-scale gscale = 1/(1-rate) (Keras: MaxPooling2D ->
 * Dropout(0.25)): zero entries (ReLU-clamped or dropped) pass nothing, the rest pass dpool * gscale */
int b200_unpool_relu_bwd_scaled(const void* dpool_bf16, const void* pooled_bf16, const uint8_t* argmax, void* dy_bf16,
                                float* dbias, int B, int HP, int WP, int C, float gscale, void* stream);
/* keras128 first layer: k x k VALID conv (k = 3 or 5), Cin=1 -> C1, + bias + ReLU, out [B,28-k+1,28-k+1,C1] bf16.
 * step_ptr / ring: as b200_conv1_pool_fwd. */
int b200_conv1_valid_fwd(const float* x, const void* w_bf16, const void* bias_bf16, void* out_bf16, int B, int C1, int k,
                         int* step_ptr, int bump_step, float* loss_ring, int* correct_ring, int ring, void* stream);
/* its weight/bias gradient (accumulates): dW[tap,c] += sum x[pix+tap] * da1[pix,c] * (a1[pix,c] > 0) */
int b200_conv1_valid_wgrad(const float* x, const void* da1_bf16, const void* a1_bf16, float* dw, float* dbias, int B,
                           int C1, int k, void* stream);
/* inverted dropout on a bf16 tensor of n (multiple of 4) elements, in place allowed; Philox4x32-10 keyed by
 * (seed, *step_ptr, layer_id), the stream b200dp/utils/philox.py reproduces on the host */
int b200_dropout_bf16(const void* in_bf16, void* out_bf16, long long n, float rate, uint64_t seed, const int* step_ptr,
                      int layer_id, void* stream);
/* direct convolution fallbacks (any k, stride 1), NHWC bf16, fp32 accumulate */
int b200_conv_direct_fwd(const void* in, const void* w, const void* bias, void* out, int B, int H, int W, int Ci,
                         int Co, int k, int pad, int relu, void* stream);
int b200_conv_direct_dgrad(const void* dout, const void* w, void* din, int B, int H, int W, int Ci, int Co, int k,
                           int pad, void* stream);
int b200_conv_direct_wgrad(const void* in, const void* dout, float* dw, float* dbias, int B, int H, int W, int Ci,
                           int Co, int k, int pad, void* stream);
/* C[M,N] (+)= op(A) op(B); bf16 in, fp32 accumulate; CUDA-core fallback and on-device oracle.
 * transA: A is [K,M]; transB: B is [N,K]. out_bf16: C stored as bf16 (else fp32). */
int b200_gemm_simple(const void* A, const void* B, void* C, int M, int N, int K, int transA, int transB,
                     int accumulate, int out_bf16, void* stream);
int b200_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream);
int b200_adam_local(float* master, float* m, float* v, const float* grad, void* w16, long long n, int step, double lr,
                    double b1, double b2, double eps, int eps_mode, float grad_scale, void* stream);
int b200_bias_relu_bf16(const float* z, const void* bias_bf16, void* out_bf16, int rows, int cols, int relu, void* stream);

/* ---- model kernels, tcgen05 family (UMMA + TMEM + TMA) ---- */
/* D[M,N] = A.B with bf16 operands, fp32 accumulation in TMEM.
 *   a_mn_major=0: A stored [M,K] (K contiguous);  =1: A stored [K,M] (M contiguous)
 *   b_mn_major=0: B stored [N,K] (K contiguous);  =1: B stored [K,N] (N contiguous)
 *   out_mode 0: fp32 store, 1: bf16 store, 2: fp32 atomic accumulate (required when split_k > 1)
 *   lda/ldb/ldc are row pitches in elements of the stored matrices. */
int b200_gemm_tc(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                 int a_mn_major, int b_mn_major, int out_mode, int split_k, int block_n, void* stream);
/* GEMM fused with the reduce-scatter leg of the gradient allreduce (dense wgrad -> owner ranks, tile by tile over
 * NVLink with red.add at system scope).  The [M,N] fp32 buffer at (c_seg, c_off) of every rank must be zero when
 * the step starts; rank r ends up holding the cross-rank SUM of rows [r*M/n, (r+1)*M/n).  Pair it with range
 * algo 3 of b200_fused_allreduce_adam.  (horovod/tensorflow_mnist.py:133-136: gradient of layers.dense -> allreduce) */
int b200_gemm_tc_reduce_scatter(b200_comm_t comm, const void* A, const void* B, int c_seg, size_t c_off, int nranks,
                                int M, int N, int K, int lda, int ldb, int ldc, int a_mn_major, int b_mn_major,
                                int block_n, void* stream);
/* Dense wgrad GEMM over the global batch with the optimizer in its epilogue: D = A^T.B (A stored [K,M], B stored [K,N],
 * K = nranks*batch rows gathered by b200_allgather_acts) is the cross-rank gradient sum of the M rows this rank owns; the
 * epilogue scales it, applies Adam (horovod/tensorflow_mnist.py:130) to the fp32 master weights/moments in state_seg and
 * stores the bf16 weights into EVERY rank's w16_seg over NVLink.  Element (row, col) is parameter
 * param_start + (row0+row)*N + col.  as_rank -1: the communicator's own rank (emulated communicators name one). */
int b200_gemm_tc_wgrad_adam(b200_comm_t comm, const void* A, const void* B, int M, int N, int K, int lda, int ldb,
                            int state_seg, size_t master_off, size_t m_off, size_t v_off, int w16_seg, size_t w16_off,
                            int ctrl_seg, size_t step_off, size_t hyper_off, long long param_start, int row0, int as_rank,
                            float gscale, int block_n, void* stream);
/* Implicit-GEMM 5x5/3x3 SAME convolution on tensor cores, taps applied by shifted TMA boxes (no im2col buffer).
 *   mode 0 forward: out = relu(conv(in, w) + bias), optionally followed by maxpool2 (+argmax)
 *   mode 1 dgrad  : din = conv_transpose(dout, w)  */
int b200_conv_tc_fwd(const void* in, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H, int W,
                     int Ci, int Co, int k, int relu, int pool, void* stream);
int b200_conv_tc_dgrad(const void* dout, const void* w, void* din, const void* mask_src, int B, int H, int W, int Ci,
                       int Co, int k, void* stream);
int b200_conv_tc_wgrad(const void* in, const void* dout, float* dw, float* dbias, int B, int H, int W, int Ci, int Co,
                       int k, void* stream);
/* The same three kernels with an explicit padding 0 <= pad < k (pad = 0: VALID) and any spatial size (the pixel tiles
 * walk W as well as H).  H x W is always the forward INPUT size; the forward output / dout is
 * (H+2*pad-k+1) x (W+2*pad-k+1). */
int b200_conv_tc_fwd_pad(const void* in, const void* w, const void* bias, void* out, uint8_t* argmax, int B, int H, int W,
                         int Ci, int Co, int k, int pad, int relu, int pool, void* stream);
int b200_conv_tc_dgrad_pad(const void* dout, const void* w, void* din, int B, int H, int W, int Ci, int Co, int k,
                           int pad, void* stream);
int b200_conv_tc_wgrad_pad(const void* in, const void* dout, float* dw, float* dbias, int B, int H, int W, int Ci,
                           int Co, int k, int pad, void* stream);

/* ---- native minibatch loader (host threads; no CUDA calls) ----
 * Infinite shuffled-minibatch stream with the reference generator's semantics (horovod/tensorflow_mnist.py:76-85): a fresh
 * permutation per pass, consecutive full batches, tail dropped.  One producer thread copies rows into the caller's ring of
 * slots (slot_x[i]: batch*row_bytes bytes, slot_y[i]: batch int32 labels; pinned memory on a GPU box).  Batches come out in
 * order; a slot must be released before the producer may overwrite it. */
typedef void* b200_loader_t;
int b200_loader_create(const void* x, const int* y, long long n, size_t row_bytes, int batch, int nslots, int shuffle,
                       uint64_t seed, void* const* slot_x, int* const* slot_y, b200_loader_t* out);
int b200_loader_next(b200_loader_t loader, int* slot, long long* batch_index);
int b200_loader_release(b200_loader_t loader, int slot);
long long b200_loader_batches_per_epoch(b200_loader_t loader);
int b200_loader_destroy(b200_loader_t loader);

/* ---- host runtime ---- */
/* JSON description of the visible GPUs: names, SM counts, memory, peer-access matrix. Returns bytes written. */
int b200_topology_json(char* buf, int buflen);
int b200_l2_flush(void* scratch, size_t bytes, void* stream);
/* L2 access-policy window on a stream (persisting hits, streaming misses); bytes == 0 clears it.  Opt-in experiment of the
 * engine (B200DP_L2_PERSIST=1): keep the Adam state resident between steps. */
int b200_stream_l2_persist(void* base, size_t bytes, float hit_ratio, void* stream);
/* Programmatic dependent launch for the step kernels (default on; also B200DP_PDL=0). */
int b200_set_pdl(int on);
/* Launches made while this is set promise that the weights they read were not written by the kernel launched just before
 * them (true inside a training step): such kernels fetch their weights ahead of the programmatic dependency wait. */
int b200_set_stable_weights(int on);

#ifdef __cplusplus
}
#endif
#endif
