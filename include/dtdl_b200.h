/* C ABI of libdtdl_b200.so -- the sm_100a kernels and peer-memory runtime of the
 * B200-native data-parallel training framework.
 *
 * Conventions
 *   - Every function returns 0 on success; on failure it returns non-zero and
 *     dtdl_last_error() describes why (thread local).
 *   - Activations are NHWC, bf16 or fp32 (the `act` argument), with the channel count padded
 *     to a multiple of 8 ("Cp"); pad channels hold zeros.  `P` is the number of pixels N*H*W.
 *   - `stream` is a cudaStream_t passed as void* (0 = default stream).  All work
 *     is asynchronous on that stream; nothing here synchronises the device.
 *   - Functions are safe to capture in a CUDA graph unless stated otherwise.
 *
 * Each block names the reference behaviour it provides (paths are relative to
 * MyXiaoPao/distributed-training-dl).
 */
#ifndef DTDL_B200_H_
#define DTDL_B200_H_

#ifdef __cplusplus
extern "C" {
#endif

const char* dtdl_last_error(void);
int dtdl_version(void);

/* ---- tensor-core ops (tcgen05 + TMEM + TMA) ------------------------------------------
 * `act` selects the activation / operand storage type of a call: 0 = bf16 (tcgen05 kind::f16),
 * 1 = fp32 read by the tensor core as TF32 (kind::tf32) -- the reference's own precision (fp32
 * tensors, cuDNN's default TF32 convolutions: pytorch/single_gpu.py:65-72).  Accumulation is fp32
 * in both.
 *
 * Convolution (stride 1 or 2) / linear forward as implicit GEMM (nn.Conv2d / nn.Linear in
 * pytorch/model.py:28-32,62-63,69,71,77; Keras Conv2D/Dense in tensorflow2/mnist_single.py:17-26;
 * Chainer L.Linear in chainer/train_mnist.py:19-21).  `wpack` comes from dtdl_pack_weights.
 * The data gradient of a stride-1 layer is the same call on dy with the dgrad pack and
 * pad' = K-1-pad.  flags: 1 = ReLU epilogue, 2 = fp32 output (act 0 only; act 1 always fp32). */
int dtdl_conv_fprop(const void* x, const void* wpack, void* y, const float* bias, int N, int H,
                    int W, int Cin_p, int Ho, int Wo, int Cout_p, int KH, int KW, int pad_h,
                    int pad_w, int flags, int stride, int act, void* stream);

/* Same, and additionally accumulates the per-channel sum / sum of squares of the stored output
 * into stats[2*Cout_p] (pre-zeroed): the batch statistics of the BatchNorm that follows the
 * convolution (pytorch/model.py:41-46) come out of the GEMM epilogue instead of a separate pass. */
int dtdl_conv_fprop_stats(const void* x, const void* wpack, void* y, const float* bias, float* stats,
                          int N, int H, int W, int Cin_p, int Ho, int Wo, int Cout_p, int KH, int KW,
                          int pad_h, int pad_w, int flags, int stride, int act, void* stream);

/* dtdl_conv_fprop for a replica's LAST layer in the single-process engine: the first y2_cols output
 * channels are also stored as fp32 rows (stride y2_ld floats) at y2, which may be peer-mapped memory of
 * GPU 0 -- nn.DataParallel's gather (pytorch/data_parallel.py:104) becomes a store of the GEMM epilogue. */
int dtdl_conv_fprop_gather(const void* x, const void* wpack, void* y, const float* bias, float* y2,
                           int y2_ld, int y2_cols, int N, int H, int W, int Cin_p, int Ho, int Wo,
                           int Cout_p, int KH, int KW, int pad_h, int pad_w, int flags, int act,
                           void* stream);

/* Data gradient of a stride-2 convolution (pytorch/model.py:28-29 with stride=2 from :69,71) in
 * gather form: four launches, one per output phase (h%2, w%2), each a stride-1 convolution over
 * dy [N,Hy,Wy,Cout_p] with the taps that reach that phase, stored to every other pixel of
 * dx [N,H,W,Cin_p].  `wdpack` is the ordinary dgrad pack.  Filters of 2x2 and larger. */
int dtdl_conv_dgrad_s2(const void* dy, const void* wdpack, void* dx, int N, int Hy, int Wy,
                       int Cout_p, int H, int W, int Cin_p, int KH, int KW, int pad_h, int pad_w,
                       int act, void* stream);

/* Weight gradient (autograd of the layers above; pytorch/single_gpu.py:102).  Writes (or
 * accumulates into) the fp32 OIHW gradient `dw`.  Needs scratch of the size returned by
 * dtdl_conv_wgrad_workspace_bytes_ex for the same shape, stride and act
 * (dtdl_conv_wgrad_workspace_bytes = the stride-1 bf16 case). */
long long dtdl_conv_wgrad_workspace_bytes(int N, int Ho, int Wo, int Cin_p, int Cout_p, int KH,
                                          int KW);
long long dtdl_conv_wgrad_workspace_bytes_ex(int N, int Ho, int Wo, int Cin_p, int Cout_p, int KH,
                                             int KW, int stride, int act);
int dtdl_conv_wgrad(const void* x, const void* dy, float* dw, void* workspace,
                    long long workspace_bytes, int N, int H, int W, int Cin, int Cin_p, int Ho,
                    int Wo, int Cout, int Cout_p, int KH, int KW, int pad_h, int pad_w,
                    int accumulate, int stride, int act, void* stream);

/* dtdl_conv_wgrad for callers that cannot allocate device memory (no counterpart in the
 * reference, which never manages scratch itself).  Such callers exist: programs whose own pool
 * owns all GPU memory, plug-ins that receive tensors but no allocator, code entered during
 * stream capture where cudaMalloc is not allowed.  The split-K scratch is a 32 MiB array
 * inside the library (held by every GPU that loads it until exit); shapes needing more are
 * rejected.  Same arguments and result as the stride-1, bf16 (act 0) dtdl_conv_wgrad minus the
 * workspace pair.  The first
 * call on a device resolves the array's device address -- make it before any stream capture.
 * Calls share the array: calls that can overlap on the GPU race, so keep them stream-ordered.
 * The call allocates nothing. */
long long dtdl_conv_wgrad_static_scratch_bytes(void);
int dtdl_conv_wgrad_static(const void* x, const void* dy, float* dw, int N, int H, int W, int Cin,
                           int Cin_p, int Ho, int Wo, int Cout, int Cout_p, int KH, int KW,
                           int pad_h, int pad_w, int accumulate, void* stream);

/* fp32 master weights -> operand packs (bf16, or fp32 for act 1) for a whole model in one launch. */
int dtdl_pack_entry_size(void);
long long dtdl_pack_entry_fill(void* entry, const float* w, void* wf, void* wd, int Cout, int Cin,
                               int taps, int Cin_p, int Cout_p, long long first_block, int act);
int dtdl_pack_weights(const void* device_table, int n_entries, long long total_blocks,
                      void* stream);

/* ---- bandwidth-bound layer ops --------------------------------------------------------
 * BatchNorm2d train/eval (pytorch/model.py:27,30,33,64,74), ReLU (:34,75), residual add with
 * zero-padded shortcut (:18,49), pooling (:11,76; tensorflow2/mnist_single.py:19,21),
 * softmax cross entropy + accuracy count (pytorch/single_gpu.py:99,105-108).
 * Activations are NHWC, bf16 (act 0) or fp32 (act 1), channels padded to 8. */
int dtdl_nchw_to_nhwc(const float* x, void* y, int N, int C, int H, int W, int Cp, int act,
                      void* stream);
int dtdl_nchw_to_s2d(const float* x, void* y, int N, int C, int H, int W, int Cp, int act,
                     void* stream);
int dtdl_bn_stats(const void* x, float* sums /* [2*Cp], zeroed */, long long P, int Cp, int act,
                  void* stream);
int dtdl_bn_apply(const void* x, void* y, const float* sums, const float* gamma, const float* beta,
                  float* running_mean, float* running_var, long long* num_batches_tracked,
                  float* save_mean, float* save_rstd, const void* residual, int Cres_p, long long P,
                  int C, int Cp, float eps, float momentum, int relu, int training, float* zero_buf,
                  int zero_n, int act, void* stream);
/* Same; additionally accumulates sum / sum of squares of the stored y into out_sums[2*Cp]
 * (pre-zeroed): the statistics of the BatchNorm that opens the next block (pytorch/model.py:27). */
int dtdl_bn_apply_stats(const void* x, void* y, const float* sums, const float* gamma, const float* beta,
                  float* running_mean, float* running_var, long long* num_batches_tracked,
                  float* save_mean, float* save_rstd, const void* residual, int Cres_p, long long P,
                  int C, int Cp, float eps, float momentum, int relu, int training, float* zero_buf,
                  int zero_n, float* out_sums, int act, void* stream);
int dtdl_bn_bwd(const void* dy, const void* x, const void* y, const float* save_mean,
                const float* save_rstd, const float* gamma, float* sums2 /* [2*Cp], zeroed */,
                void* dx, void* dres, int Cres_p, float* dgamma, float* dbeta, int accumulate,
                long long P, int C, int Cp, int relu, float* zero_buf, int zero_n, int act, void* stream);
/* Same; dx additionally receives add_in (shape of x): the gradient that reaches the same
 * tensor through the identity shortcut (pytorch/model.py:37,49), so autograd needs no add pass. */
int dtdl_bn_bwd_add(const void* dy, const void* x, const void* y, const float* save_mean,
                const float* save_rstd, const float* gamma, float* sums2 /* [2*Cp], zeroed */,
                void* dx, void* dres, int Cres_p, float* dgamma, float* dbeta, int accumulate,
                long long P, int C, int Cp, int relu, float* zero_buf, int zero_n, const void* add_in,
                int act, void* stream);
/* dtdl_bn_bwd_add as ONE cooperative kernel: each CTA keeps its slice of the masked gradient and of x in
 * shared memory across a grid barrier, so dy / x / y are read once instead of twice.  Only for layers
 * whose dy + x fit the GPU's aggregate shared memory (dtdl_bn_bwd_fused_fits: 1 / 0).  `barrier` = 8
 * zeroed bytes of device memory owned by the caller, not shared with a concurrently running call. */
int dtdl_bn_bwd_fused_fits(long long P, int Cp, int act);
int dtdl_bn_bwd_fused(const void* dy, const void* x, const void* y, const float* save_mean,
                const float* save_rstd, const float* gamma, float* sums2 /* [2*Cp], zeroed */,
                void* dx, void* dres, int Cres_p, float* dgamma, float* dbeta, int accumulate,
                long long P, int C, int Cp, int relu, float* zero_buf, int zero_n, const void* add_in,
                void* barrier, int act, void* stream);
int dtdl_relu_bwd(const void* dy, const void* y, void* dx, long long n_elems, int act, void* stream);
int dtdl_add(const void* a, const void* b, void* out, long long n_elems, int act, void* stream);
int dtdl_add_prefix(void* a, const void* b, long long P, int Ca_p, int Cb_p, int act, void* stream);
int dtdl_pool_fwd(const void* x, void* y, int N, int H, int W, int Ho, int Wo, int Cp, int k,
                  int stride, int pad, int mode /* 0 avg, 1 max */, int act, void* stream);
int dtdl_pool_bwd(const void* dy, const void* x, const void* y, void* dx, int N, int H, int W,
                  int Ho, int Wo, int Cp, int k, int stride, int pad, int mode, int act, void* stream);
int dtdl_subsample2(const void* x, void* y, int N, int H, int W, int Cp, int adjoint, int act,
                    void* stream);
int dtdl_gap_fwd(const void* x, void* y, int N, int HW, int Cp, int act, void* stream);
int dtdl_gap_bwd(const void* dy, void* dx, int N, int HW, int Cp, int act, void* stream);
int dtdl_softmax_ce(const float* logits, const long long* targets, float* out2 /* loss, correct */,
                    void* dlogits_bf16, float* dlogits_f32, int N, int C, int Cp, float inv_n,
                    float grad_scale, void* stream);
int dtdl_colsum(const void* x, float* out /* zeroed */, long long P, int C, int Cp, int act,
                void* stream);
/* fp32 [rows][cols] <-> activation type [rows][cols_p] (zero padded) */
int dtdl_pad_cast(const float* x, void* y, long long rows, int cols, int cols_p, int act,
                  void* stream);
int dtdl_unpad_cast(const void* x, float* y, long long rows, int cols, int cols_p, int act,
                    void* stream);
/* Full-fp32 Linear on CUDA cores for the fp32 activation mode: torch keeps TF32 off for matmuls,
 * so the reference's nn.Linear (pytorch/model.py:77,111) is plain fp32.  x [N][Kp], y / dy [N][Op]
 * fp32 padded rows; w [O][K], dw [O][K]; dx may be null (first layer). */
int dtdl_linear_f32_fwd(const float* x, const float* w, const float* bias, float* y,
                        float* y2 /* optional second copy, rows of y2_ld floats (may be peer memory) */,
                        int y2_ld, int N, int K, int Kp, int O, int Op, int relu, void* stream);
int dtdl_linear_f32_bwd(const float* dy, const float* x, const float* w, float* dx, float* dw, int N,
                        int K, int Kp, int O, int Op, int accumulate, void* stream);

/* ---- peer-memory runtime and collectives ----------------------------------------------
 * Replaces the NCCL calls made by DistributedDataParallel / nn.DataParallel / ChainerMN /
 * MirroredStrategy (pytorch/distributed_data_parallel.py:61-62,74,132; pytorch/data_parallel.py:71;
 * chainer/train_mnist_multi.py:55,81-82; tensorflow2/mnist_mirror_strategy.py:12).
 * "Symmetric" buffers have the same size on every rank; `bufs[r]` is rank r's base pointer as
 * seen from the calling process (cudaIpcOpenMemHandle across processes, peer access inside
 * one).  All ranks must issue the same sequence of collective calls. */
int dtdl_comm_create(int rank, int world, void** handle, void** flags_ptr);
int dtdl_comm_set_peer_flags(void* handle, int peer, void* flags_ptr);
int dtdl_comm_destroy(void* handle);
int dtdl_symm_alloc(long long bytes, void** ptr);
int dtdl_symm_free(void* ptr);
int dtdl_ipc_get_handle(void* ptr, void* out64);
int dtdl_ipc_open_handle(const void* in64, void** ptr);
int dtdl_ipc_close_handle(void* ptr);
int dtdl_enable_peer_access(int dev, int peer);
/* 1 / 0: device can / cannot join an NVLS multicast object (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED),
 * -1: query failed.  The shipped collectives are the P2P variants; multicast ones are gated on this. */
int dtdl_multicast_supported(int device);
int dtdl_comm_barrier(void* handle, void* stream);
/* Each collective = START exchange (one-CTA kernel) + data kernel + DONE exchange (one-CTA kernel).
 * defer = 1 drops the DONE exchange of the collectives that follow: a training step then closes all
 * its buckets with ONE dtdl_comm_barrier on the same stream (buckets touch disjoint memory).  The
 * reference's DDP overlaps all but the last bucket the same way
 * (pytorch/distributed_data_parallel.py:128-133). */
int dtdl_comm_defer_done(void* handle, int defer);
/* Device-side timeline of the collectives (works inside a replayed CUDA graph): `base` = zeroed device
 * buffer of `capacity` slots of 4 x u64 {start ns, end ns, kind, bytes}; NULL = off.  Kinds: 1 START
 * wait, 2 DONE wait, 3 reduce, 4 NVLS reduce, 5 broadcast, 6 all-gather, 7 mark. */
int dtdl_comm_trace(void* handle, void* base, int capacity);
int dtdl_comm_trace_mark(void* handle, int kind, void* stream);
int dtdl_comm_trace_count(void* handle);
/* dst[r] = scale * sum over ranks of grads[r].  two_shot 0 = one-shot (every rank reads all peers;
 * needs dst != grads), 1 = two-shot; an in-place call (dst == grads) always runs two-shot. */
int dtdl_comm_allreduce(void* handle, void* const* grads, void* const* dst, long long offset_bytes,
                        long long n, float scale, int two_shot, int max_blocks, void* stream);
/* opt: 1 = SGD momentum, 2 = Adam (torch), 3 = Adam (Keras epsilon placement).
 * hyper: device float[16] = lr, momentum, weight_decay, 1/world, beta1, beta2, eps,
 *        1-beta1^t, sqrt(1-beta2^t), dampening. */
int dtdl_comm_reduce_update(void* handle, void* const* grads, void* const* params, float* mom,
                            float* var, const float* hyper, long long offset_bytes, long long n,
                            int opt, int two_shot, int max_blocks, void* stream);
/* bf16 gradient transport (option; the north star's "cast fused with the collective and the update"):
 * dtdl_cast_f32_bf16 fills the bf16 gradient arena, the *16 variants reduce it over NVLink in half the
 * bytes (fp32 accumulation), apply 1/W + weight decay + SGD/Adam in fp32 and store fp32 parameters to
 * every rank.  offset_bytes / n address the fp32 arenas (n multiple of 8); the bf16 arena is addressed
 * at offset_bytes / 2.  The reference's DDP moves fp32 gradients
 * (pytorch/distributed_data_parallel.py:132), so this is off by default. */
int dtdl_cast_f32_bf16(const float* src, void* dst, long long n, void* stream);
int dtdl_comm_reduce_update16(void* handle, void* const* grads16, void* const* params, float* mom,
                              float* var, const float* hyper, long long offset_bytes, long long n,
                              int opt, int max_blocks, void* stream);
int dtdl_comm_nvls_reduce_update16(void* handle, void* mc_grads16, void* mc_params, float* param, float* mom,
                                   float* var, const float* hyper, long long offset_bytes, long long n,
                                   int opt, int max_blocks, void* stream);
int dtdl_comm_broadcast(void* handle, void* const* bufs, long long offset_bytes, long long bytes,
                        int root, int max_blocks, void* stream);
int dtdl_comm_allgather(void* handle, void* const* bufs, long long offset_bytes, long long bytes,
                        int max_blocks, void* stream);
/* Single-rank optimizer step (torch.optim.SGD / Adam: pytorch/single_gpu.py:71-72,103;
 * chainer/train_mnist.py:69). */
int dtdl_local_update(float* param, const float* grad, float* mom, float* var, const float* hyper,
                      long long n, int opt, void* stream);

/* ---- NVLS: symmetric memory through cuMem* + multicast objects (csrc/comm/vmm.cpp) ----------
 * The reference reaches NVSwitch reduction only inside NCCL (pytorch/distributed_data_parallel.py:61-62);
 * here the buffers are allocated so that our own kernels can issue multimem.ld_reduce / multimem.st.
 * Handles are CUmemGenericAllocationHandle values carried as long long. */
long long dtdl_vmm_granularity(int dev, int world);
int dtdl_vmm_create(int dev, long long bytes, int shareable, long long* handle);
int dtdl_vmm_export_fd(long long handle, int* fd);
int dtdl_vmm_import_fd(int fd, long long* handle);
int dtdl_vmm_map(long long handle, long long bytes, long long align, const int* devs, int ndevs, void** ptr);
int dtdl_vmm_unmap(void* ptr, long long bytes);
int dtdl_vmm_release(long long handle);
int dtdl_vmm_close_fd(int fd);
int dtdl_mc_create(int world, long long bytes, int shareable, long long* handle);
int dtdl_mc_add_device(long long mc, int dev);
int dtdl_mc_bind(long long mc, long long mem, long long bytes);

/* Collectives through the multicast address (two-shot shape: rank r reduces slice r in the
 * switch and stores the result -- or the updated parameters -- to every rank at once). */
int dtdl_comm_nvls_allreduce(void* handle, void* mc_src, void* mc_dst, long long offset_bytes, long long n,
                             float scale, int max_blocks, void* stream);
int dtdl_comm_nvls_reduce_update(void* handle, void* mc_grads, void* mc_params, float* param, float* mom,
                                 float* var, const float* hyper, long long offset_bytes, long long n, int opt,
                                 int max_blocks, void* stream);
int dtdl_comm_nvls_broadcast(void* handle, void* local, void* mc, long long offset_bytes, long long bytes,
                             int root, int max_blocks, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DTDL_B200_H_ */
