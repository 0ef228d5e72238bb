/* b200dp C ABI - everything the two native libraries export.
 *
 *   libb200dp_cuda.so   sm_100a kernels + symmetric-memory runtime   (b200_*)
 *   libb200dp_host.so   TCP store, shared-memory collectives, DDP bucket planner (b200h_*)
 *
 * Conventions: every b200_* entry point returns 0 on success, non-zero on failure, and
 * b200_last_error() returns the message of the calling thread's last failure.  Pointers are
 * device pointers unless the name says host; `stream` is a cudaStream_t.  Activations are bf16
 * NHWC, parameters / gradients / statistics fp32.  No entry point allocates device memory
 * unless it says so; callers own every buffer, including scratch ("partials").
 *
 * Citations "ref:" point at the reference repo (AditMeh/torch-distributed-sandbox) call site
 * whose behaviour the entry point provides; "K#/C#/D#" are the SURVEY.md inventory rows.
 */
#ifndef B200DP_H
#define B200DP_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ common --- */
const char* b200_last_error(void);
int b200_sm_count(void);
int b200_l2_flush(void* buf, size_t bytes, void* stream);            /* timing hygiene helper */

/* ------------------------------------------------- input pipeline (K1/K2) --- */
/* ref: transforms.Resize(IMAGE_SHAPE) + ToTensor, mnist_onegpu.py:53 / mnist_distributed.py:71.
 * uint8 [B,sh,sw] -> fp32 [B,H,W]; bit-exact with Pillow's 8-bit bilinear resample + /255. */
int b200_resize_pil_bilinear(const void* src_u8, int B, int sh, int sw, const void* xtab_i32, const void* ytab_i32,
                             void* dst_f32, int H, int W, void* stream);

/* ------------------------------------------------ layer 1 (K3/K4/K5/K13) --- */
/* ref: nn.Conv2d(1,16,5,1,2) forward, mnist_distributed.py:29.  partials: [b200_conv1_fwd_grid()][32]. */
int b200_conv1_fwd_grid(void);
int b200_conv1_fwd(const void* in_f32, const void* w1, const void* b1, void* x1_bf16, void* partials, int B, int H,
                   int W, void* stream);
/* ref: nn.BatchNorm2d train-mode statistics + running-stat update, mnist_distributed.py:30,35.
 * out = [scale | shift | mean | rstd] (4*C floats). */
int b200_bn_finalize(const void* partials, int nparts, int C, double count, const void* gamma, const void* beta,
                     void* running_mean, void* running_var, void* num_batches_tracked_i64, float momentum, float eps,
                     void* out, void* stream);
/* ref: BatchNorm2d apply + nn.ReLU + nn.MaxPool2d(2,2), mnist_distributed.py:30-32,35-37. C in {16,32}. */
int b200_bn_relu_pool(const void* x_bf16, const void* bn, void* p_bf16, int B, int H, int W, int C, void* stream);
/* ref: loss.backward() through MaxPool2d/ReLU/BatchNorm2d, mnist_distributed.py:96.
 * mode 0: partials [b200_pool_bn_bwd_grid()][2C];  mode 1: dx (may alias x). */
int b200_pool_bn_bwd_grid(void);
int b200_pool_bn_bwd(const void* x_bf16, const void* dp_bf16, const void* bn, const void* red, void* partials,
                     void* dx_bf16, int B, int H, int W, int C, int mode, void* stream);
int b200_bn_bwd_finalize(const void* partials, int nparts, int C, double count, void* red, void* dgamma, void* dbeta,
                         float gscale, void* stream);
/* ref: conv1 weight gradient inside loss.backward().  partials: [b200_conv1_wgrad_grid()][400] scratch. */
int b200_conv1_wgrad_grid(void);
int b200_conv1_wgrad(const void* in_f32, const void* x1_bf16, const void* dp1_bf16, const void* bn, const void* red,
                     void* partials, void* dw1, int B, int H, int W, float gscale, void* stream);

/* Allocation-free variant of b200_conv1_wgrad.  No counterpart in the reference: it exists for callers
 * that cannot obtain device scratch themselves (inside a stream capture or a no-new-allocations memory
 * policy, a fixed up-front memory budget, a thin driver that owns only its inputs and outputs).  The
 * per-block partial sums go to a 256,000-byte device array built into the library (one per GPU, held
 * until exit) instead of a caller buffer.  nblocks = thread blocks / partial rows (0: the SM count);
 * more than b200_conv1_wgrad_noalloc_max_blocks() (160) is rejected.  The first call on each device
 * resolves the array's address - make it outside stream capture.  All calls on a device share the
 * array: serialise calls that could overlap on the GPU.  Never calls cudaMalloc/cudaFree.
 * It launches the CUDA-core (FFMA) kernel of b200_conv1_wgrad - about 2 ms at batch 5, 3000x3000 - not the
 * 0.7 ms tcgen05 kernel of b200_conv1_wgrad_tc, and takes bf16 x1 / dp1 only. */
int b200_conv1_wgrad_noalloc_max_blocks(void);
int b200_conv1_wgrad_noalloc(const void* in_f32, const void* x1_bf16, const void* dp1_bf16, const void* bn,
                             const void* red, void* dw1, int B, int H, int W, float gscale, int nblocks,
                             void* stream);

/* conv1 on tcgen05 (same contracts as b200_conv1_fwd / b200_conv1_wgrad; partial rows = grid, returned). */
int b200_conv1_fwd_tc(const void* in_f32, const void* w1, const void* b1, void* x1_bf16, void* partials, int B, int H,
                      int W, int* nparts_out, void* stream);
int b200_conv1_wgrad_tc(const void* in_f32, const void* x1_bf16, const void* dp1_bf16, const void* bn, const void* red,
                        void* partials, void* dw1, int B, int H, int W, float gscale, void* stream);

/* ------------------------------------------- layer 2, tcgen05 (K6/K12) --- */
/* ref: nn.Conv2d(16,32,5,1,2) forward / input gradient / weight gradient (cuDNN in the reference),
 * mnist_distributed.py:34,96.  partials (fwd): [b200_conv_tc_max_grid()][64];
 * partials (wgrad): [b200_conv_tc_max_grid()][12800] scratch. */
int b200_conv_tc_max_grid(void);
int b200_conv_tc_set_version(int v);   /* 1 tap-stationary+TMA, 2 row-stationary+TMA, 3 row-stationary+cp.async,
                                          4 = 3 + dx-replicated M=128 weight-gradient kernel,
                                          5 = 4 + one TMA request per input row into swizzled pixel rows (default) */
int b200_conv_tc_get_version(void);
int b200_set_reserved_sms(int n);      /* SM partition for an overlapped collective (DDP); 0 = none, at most SM count - 1 */
int b200_get_reserved_sms(void);
int b200_conv2_fwd_tc(const void* p1_bf16, const void* w2, const void* b2, void* x2_bf16, void* partials, int B, int H,
                      int W, int* nparts_out, void* stream);
int b200_conv2_dgrad_tc(const void* dx2_bf16, const void* w2, void* dp1_bf16, int B, int H, int W, void* stream);
int b200_conv2_dgrad_stats_tc(const void* dx2_bf16, const void* w2, const void* pooled_p1, const void* bn1,
                              void* dp1_bf16, void* partials, int B, int H, int W, int* nparts_out, void* stream);
int b200_conv2_wgrad_tc(const void* p1_bf16, const void* dx2_bf16, void* partials, void* dw2, int B, int H, int W,
                        float gscale, void* stream);

/* --------------------------------------------- head + loss (K8/K9/K10) --- */
/* ref: nn.LazyLinear(10) forward/backward and nn.CrossEntropyLoss, mnist_distributed.py:38,44,64,91,96.
 * W is [10][F] in HWC-flatten order; batch chunks of <= 8 rows.  partials: [b200_fc_grid()][80]. */
int b200_fc_grid(void);
int b200_fc_fwd(const void* p2_bf16, const void* W, const void* bias, void* partials, void* logits, int bt, size_t F,
                void* stream);
int b200_cross_entropy(const void* logits, const void* labels_i64, int B, void* loss, void* dlogits, void* dbias,
                       float gscale, void* stream);
/* accumulate: 0 = dW (fp32) overwritten, 1 = added to (batch chunks of the <= 8-row kernel), 2 = dW is a BF16 bucket
 * [10][F], overwritten (DDP gradient exchange in bf16, b200_ar_nvls_bf16_sgd). */
int b200_fc_bwd(const void* dlogits, const void* p2_bf16, void* W, void* dW_or_null, void* dp2_bf16, int bt, size_t F,
                float lr_fused_sgd, int accumulate, float gscale, const void* bn2_or_null, void* bwd_partials_or_null,
                void* stream);

/* ------------------------------- fp32 storage / TF32 math ("tf32" mode) --- */
/* ref: the reference's own numerics - fp32 tensors everywhere, cuDNN convolutions with allow_tf32 (the default),
 * fp32 cuBLAS fc (mnist_distributed.py:28-38, no autocast).  Same contracts as the entry points above with every
 * activation / gradient tensor in fp32 (NHWC); conv2 runs tcgen05.mma.kind::tf32, conv1 forward is kind::tf32 too (row-operand
 * kernel, below) or keeps its bf16 hi/lo split (fp32-level products), conv1 weight gradient is kind::tf32 or exact fp32 FFMA. */
int b200_conv1_fwd_tc_f32(const void* in_f32, const void* w1, const void* b1, void* x1_f32, void* partials, int B,
                          int H, int W, int* nparts_out, void* stream);
int b200_bn_relu_pool_f32(const void* x, const void* bn, void* p, int B, int H, int W, int C, void* stream);
int b200_pool_bn_bwd_f32(const void* x, const void* dp, const void* bn, const void* red, void* partials, void* dx,
                         int B, int H, int W, int C, int mode, void* stream);
int b200_conv1_wgrad_f32(const void* in_f32, const void* x1, const void* dp1, const void* bn, const void* red,
                         void* partials, void* dw1, int B, int H, int W, float gscale, void* stream);
int b200_conv1_wgrad_tc_f32(const void* in_f32, const void* x1, const void* dp1, const void* bn, const void* red,
                            void* partials, void* dw1, int B, int H, int W, float gscale, void* stream);
int b200_conv2_fwd_tc_f32(const void* p1, const void* w2, const void* b2, void* x2, void* partials, int B, int H,
                          int W, int* nparts_out, void* stream);
int b200_conv2_dgrad_tc_f32(const void* dx2, const void* w2, void* dp1, int B, int H, int W, void* stream);
int b200_conv2_dgrad_stats_tc_f32(const void* dx2, const void* w2, const void* pooled_p1, const void* bn1, void* dp1,
                                  void* partials, int B, int H, int W, int* nparts_out, void* stream);
int b200_conv2_wgrad_tc_f32(const void* p1, const void* dx2, void* partials, void* dw2, int B, int H, int W,
                            float gscale, void* stream);
int b200_fc_fwd_f32(const void* p2, const void* W, const void* bias, void* partials, void* logits, int bt, size_t F,
                    void* stream);
int b200_fc_bwd_f32(const void* dlogits, const void* p2, void* W, void* dW_or_null, void* dp2, int bt, size_t F,
                    float lr_fused_sgd, int accumulate, float gscale, const void* bn2_or_null,
                    void* bwd_partials_or_null, void* stream);

/* Layer 1 of the tf32 mode without its 2.88 GB conv output (mnist_distributed.py:28-31 forward, its backward inside
 * loss.backward()).  Two forward passes over the image instead of one over x1:
 *   b200_conv1_stats_*            BatchNorm batch statistics of conv1(in)+bias, nothing stored
 *                                 (partials: [b200_conv1_fwd_grid()][32], *nparts_out rows written);
 *   b200_conv1_fwd_pool_*         conv again (same MMAs in the same order: bit-identical values), BN (bn = [scale 16 |
 *                                 shift 16 | mean | rstd]) + ReLU + 2x2 max-pool -> p1 [B][H/2][W/2][16]; the _mask_
 *                                 forms also write one byte per pooled element (arg-max position | ReLU-active << 2)
 *                                 and take x1 = NULL.
 * *_tc_f32: bf16 hi/lo vertical-Toeplitz kernel (fp32-level products; may also store x1).
 * *_tf32:   kind::tf32 kernel whose MMA operand is the image row as TMA delivered it (csrc/conv1_hz.cu; W % 4 == 0,
 *           input truncated to tf32 by the tensor core as in cuDNN's TF32 convolution; stores no x1).
 * Use the same family for both passes.  b200_conv1_wgrad_nox1_f32 is the matching weight gradient: dp1 is routed with
 * the mask bytes on the tensor cores, the dense BatchNorm term comes from the input's 25x25 Gram matrix
 * (partials: [b200_conv1_wgrad_grid()][b200_conv1_wgrad_nox1_partial()]; H % 8 == 0, W % 4 == 0). */
int b200_conv1_stats_tc_f32(const void* in_f32, const void* w1, const void* b1, void* partials, int B, int H, int W,
                            int* nparts_out, void* stream);
int b200_conv1_fwd_pool_tc_f32(const void* in_f32, const void* w1, const void* b1, const void* bn, void* x1_f32,
                               void* p1_f32, void* partials, int B, int H, int W, void* stream);
int b200_conv1_fwd_pool_mask_tc_f32(const void* in_f32, const void* w1, const void* b1, const void* bn,
                                    void* x1_f32_or_null, void* p1_f32, void* amask_u8, void* partials, int B, int H,
                                    int W, void* stream);
int b200_conv1_stats_tf32(const void* in_f32, const void* w1, const void* b1, void* partials, int B, int H, int W,
                          int* nparts_out, void* stream);
int b200_conv1_fwd_pool_mask_tf32(const void* in_f32, const void* w1, const void* b1, const void* bn, void* p1_f32,
                                  void* amask_u8, int B, int H, int W, void* stream);
int b200_conv1_wgrad_nox1_partial(void);
int b200_conv1_wgrad_nox1_f32(const void* in_f32, const void* dp1, const void* amask_u8, const void* w1, const void* b1,
                              const void* bn, const void* red, void* partials, void* dw1, int B, int H, int W,
                              float gscale, void* stream);
/* fc backward that publishes its weight gradient slab by slab (ready[slab] counts finished sweeps, release at system
 * scope) for b200_ar_tile_f32 to all-reduce while later slabs are still being computed.  dtype: 0 bf16, 1 fp32. */
size_t b200_fc_bwd_tile_slab_floats(int sweeps_per_slab);
int b200_fc_bwd_tiles(const void* dlogits, const void* p2, void* W, void* dW, void* dp2, int bt, size_t F, float gscale,
                      const void* bn2, void* bwd_partials, void* ready, int sweeps_per_slab, int dtype, void* stream);

/* ------------------------------------------------------ optimizer (K14) --- */
/* ref: torch.optim.SGD(model.parameters(), 1e-4).step(), mnist_distributed.py:65,97. */
int b200_sgd(void* p, const void* g, size_t n, float lr, void* stream);

/* ----------------------------------------------------- layout helpers --- */
int b200_reduce_partials(const void* partials, int nparts, int len, void* out, float scale, void* stream);
int b200_fcw_permute(const void* src, void* dst, int C, int Hd, int Wd, int to_hwc, void* stream);
int b200_nhwc_bf16_to_nchw_f32(const void* src, void* dst, int B, int C, int H, int W, void* stream);
int b200_nchw_f32_to_nhwc_bf16(const void* src, void* dst, int B, int C, int H, int W, void* stream);

/* ------------------------------------- symmetric memory (W3) --- */
/* ref: what ProcessGroupNCCL sets up under dist.init_process_group('nccl'), mnist_distributed.py:50.
 * All of these are COLLECTIVE over the group and allocate (cold path). */
void* b200_symm_comm_create(const char* tag, int rank, int world, int device, int timeout_ms);
void b200_symm_comm_destroy(void* comm);
int b200_symm_multicast_supported(int device);
void* b200_symm_alloc(void* comm, size_t bytes, int want_multicast, void** ptrs_out, void** mc_out,
                      size_t* padded_out, void** dev_table_out, int timeout_ms);
void b200_symm_free(void* comm, void* alloc);
int b200_symm_topology(int device, int* p2p_out, int* native_atomics_out, int cap);

/* -------------------------------------------- collectives (K16, C1-C6) --- */
/* ref: dist.all_reduce / dist.barrier / DDP's bucket all-reduce and broadcasts,
 * allreduce_toy.py:31,33; mnist_distributed.py:67,96.  ptrs/flags: host arrays of `world` device VAs
 * (this process's mappings of every rank's buffer / flag page).  Each call consumes up to 4 epochs. */
size_t b200_ar_flag_bytes(void);
int b200_ar_max_blocks(void);
int b200_ar_barrier(void* const* flags, int rank, int world, uint32_t epoch, void* stream);
int b200_ar_oneshot(void* const* ptrs, void* const* flags, int rank, int world, void* out, size_t n, int dtype, int op,
                    float scale, int blocks, uint32_t epoch, void* stream);
/* in-kernel waits for a peer give up (diagnostic + trap) after this long; set from init_process_group(timeout=) */
int b200_ar_set_timeout_ms(long long ms);
int b200_ar_oneshot_f32x4(void* const* ptrs, void* const* flags, int rank, int world, void* out, void* param, size_t n,
                          float scale, float lr, int blocks, uint32_t epoch, void* stream);
int b200_ar_large_f32(void* const* ptrs, void* mc, void* const* flags, void* const* wptrs, void* wmc, int rank,
                      int world, size_t n, float scale, float lr, int algo, int blocks, uint32_t epoch, void* stream);
int b200_ar_nvls_bf16(void* mc, void* const* flags, int rank, int world, size_t n, float scale, int blocks,
                      uint32_t epoch, void* stream);
int b200_ar_broadcast(void* const* ptrs, void* const* flags, int rank, int world, void* out, size_t bytes, int root,
                      int blocks, uint32_t epoch, void* stream);
int b200_ar_allgather(void* const* ptrs, void* const* flags, int rank, int world, void* out, size_t bytes, int blocks,
                      uint32_t epoch, void* stream);
/* DDP bucket exchanged in bf16 (torch's bf16_compress_hook for the reference's DistributedDataParallel,
 * mnist_distributed.py:67): `mc` = multicast address of the symmetric bf16 gradient bucket (b200_fc_bwd* writes it with
 * accumulate = 2), `wptrs` / `wmc` = the fp32 master weights.  One pass: switch-side fp32 accumulation of the ranks'
 * bf16 values, weights -= lr * scale * sum, new weights multicast to every rank.  n = gradient elements (multiple of 8). */
int b200_ar_nvls_bf16_sgd(void* mc, void* const* flags, void* const* wptrs, void* wmc, int rank, int world, size_t n,
                          float scale, float lr, int blocks, uint32_t epoch, void* stream);
/* Small all-reduce through per-peer push rings (each rank writes its data + flag into every peer's slot; two ring
 * halves, no trailing barrier): the <= 64 KB path of SymmBackend.all_reduce. */
int b200_ar_push(void* const* ptrs, void* const* flags, int rank, int world, const void* in, void* out, size_t n,
                 int dtype, int op, float scale, size_t slot_bytes, int buf, int blocks, uint32_t epoch, void* stream);
/* One-shot fp32 all-reduce + SGD with a broadcast of bc_floats floats from bc_root folded into the same launch
 * (the BatchNorm-buffer broadcast DDP does before each forward). */
int b200_ar_oneshot_f32x4_bcast(void* const* ptrs, void* const* flags, int rank, int world, void* out, void* param,
                                size_t n, float scale, float lr, int blocks, uint32_t epoch, void* stream,
                                void* const* bc_ptrs, size_t bc_floats, int bc_root);
/* Tile-driven one-pass all-reduce + SGD of the fc weight gradient: consumes the slab counters b200_fc_bwd_tiles
 * publishes (rdy[r] = rank r's counters), reduces each slab as soon as every rank has produced it. */
int b200_ar_tile_f32(void* const* ptrs, void* mc, void* const* flags, void* const* wptrs, void* wmc, void* const* rdy,
                     int rank, int world, int rows, size_t F, size_t slab_floats, uint32_t target, float scale, float lr,
                     int blocks, uint32_t epoch, void* stream);
/* NVLink topology of `device` from NVML (dlopen): active links, links that end at an NVSwitch, per-peer direct links
 * (peer_links[cap]), per-link GB/s, multicast capability. */
int b200_nvlink_report(int device, int* links_active, int* links_to_switch, int* peer_links, int cap, int* link_gbs,
                       int* multicast);

/* =========================================================== host library === */
/* ref: the TCPStore behind env:// rendezvous, test_init.py:76-91 (D2). */
const char* b200h_last_error(void);
void* b200h_store_server_start(const char* host, int port, int* actual_port);
void b200h_store_server_stop(void* server);
void* b200h_store_connect(const char* host, int port, int timeout_ms);
void b200h_store_close(void* client);
int b200h_store_set(void* client, const char* key, const void* val, int len);
int b200h_store_get(void* client, const char* key, void* buf, int cap, int* len_out, int timeout_ms);
int b200h_store_add(void* client, const char* key, long long delta, long long* result);
int b200h_store_check(void* client, const char* key);
int b200h_store_delete(void* client, const char* key);
long long b200h_store_num_keys(void* client);
/* ref: the gloo backend of test_init.py:55 (CPU collectives; D3-D6). dtype: 0 f32,1 f64,2 i32,3 i64,4 u8,5 bf16;
 * op: 0 SUM,1 AVG,2 MAX,3 MIN,4 PROD. */
const char* b200h_coll_last_error(void);
void* b200h_coll_open(const char* name, int rank, int world, int is_creator, size_t slot_bytes, int timeout_ms);
void b200h_coll_close(void* group, int unlink_segment);
int b200h_coll_barrier(void* group);
int b200h_coll_allreduce(void* group, void* data, size_t count, int dtype, int op);
int b200h_coll_broadcast(void* group, void* data, size_t bytes, int root);
int b200h_coll_allgather(void* group, const void* in, void* out, size_t bytes);
/* ref: torch Reducer's compute_bucket_assignment_by_size behind DistributedDataParallel, mnist_distributed.py:67 (D7). */
int b200h_plan_buckets(const int64_t* sizes, const int32_t* keys, int n, const int64_t* limits, int nlimits,
                       int32_t* bucket_of);
int64_t b200h_bucket_offsets(const int64_t* numels, int n, int64_t align_elems, int64_t* offsets);

#ifdef __cplusplus
}
#endif
#endif /* B200DP_H */
