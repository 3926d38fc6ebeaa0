/* dpt_b200 — C ABI of the B200-native UNet training kernels (libdpt_b200.so).
 *
 * All tensors are raw device pointers.  Activations are NHWC bf16, parameters / gradients / optimizer state
 * are fp32 in PyTorch's native layouts (so reference checkpoints load unchanged), bf16 kernel-side weight
 * copies are produced by dpt_prep_* .  Every launcher takes a cudaStream_t (as void*) and returns 0 on
 * success, -1 on error (text from dpt_last_error()).  Nothing here synchronises the stream.
 *
 * Reference citations (file:line) are relative to the upstream repository notnitsuj/DistributedPyTorch.
 */
#ifndef DPT_B200_H_
#define DPT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* dpt_last_error(void);
int dpt_version(void);
/* Number of kernels this library has launched in the calling process (all threads, all devices). */
long long dpt_launch_count(void);
/* Leave `n` SMs free: every plan / launch created AFTERWARDS sizes its grid to the remaining SMs (pipeline mode: NCCL's
 * point-to-point kernels run beside the persistent one-CTA-per-SM compute kernels instead of stalling them).  Also read
 * from DPT_SM_RESERVE at load time.  dpt_get_sm_count returns the SM count the launchers use. */
int dpt_set_sm_reserve(int n);
int dpt_get_sm_count(void);

/* ------------------------------------------------------------------------------------------------------
 * Implicit-GEMM convolution family on tcgen05/TMEM fed by TMA  (replaces cuDNN fprop/dgrad reached from
 * model/unet_parts.py:10-12,51-54 and the autograd conv backward of utils/train_utils.py:69).
 *
 *   out[pixel, n] = epilogue( sum_{tap, k} A_tap[pixel, k] * W[tap][n][k] )
 *
 * The A operand is read straight from one or two NHWC bf16 activation tensors (two = channel concat fused
 * into the K loop, replacing torch.cat at model/unet_parts.py:59,64,69,74).
 * ------------------------------------------------------------------------------------------------------ */
enum {
  DPT_CONV_3X3 = 0,     /* 3x3, pad 1, stride 1: one halo tile in smem, nine shifted UMMA views of it      */
  DPT_CONV_3X3_TAPBOX = 1, /* same math, one TMA box per tap (no smem reuse); debugging / fallback        */
  DPT_CONV_2X2S2 = 2,   /* 2x2 stride-2 gather (ConvTranspose2d backward-data): src is [B,2H,2W,C]         */
  DPT_CONV_1X1 = 3      /* pointwise (ConvTranspose2d forward together with DPT_OUT_SHUFFLE2)              */
};
enum {
  DPT_OUT_NHWC = 0,     /* dst[b, y, x, n]                                                                */
  DPT_OUT_SHUFFLE2 = 1  /* n = (dy*2+dx)*C + c  ->  dst[b, 2y+dy, 2x+dx, c]   (pixel shuffle, C = cout/4)  */
};

typedef struct dpt_conv_desc {
  int mode;             /* DPT_CONV_*                                                                     */
  int out_mode;         /* DPT_OUT_*                                                                      */
  int batch, h, w;      /* pixel space of the GEMM M dimension (= output pixels before any shuffle)        */
  int num_src;          /* 1 or 2                                                                         */
  const void* src[2];   /* bf16 NHWC; for DPT_CONV_2X2S2 the tensor is [batch, 2h, 2w, src_c]              */
  int src_c[2];         /* channels of each source (multiples of 16)                                       */
  const void* weight;   /* bf16 [taps][cout][src_c[0]+src_c[1]]                                            */
  int cout;             /* GEMM N (multiple of 32)                                                         */
  const float* bias;    /* fp32 [cout] or NULL                                                             */
  int relu;             /* apply max(x,0) in the epilogue                                                  */
  int num_dst;          /* 1 or 2: output channels [0,dst_c[0]) go to dst[0], the rest to dst[1]           */
  void* dst[2];         /* bf16                                                                            */
  int dst_c[2];
  const void* mask[2];  /* optional bf16 tensor shaped like dst[i]: out *= (mask > 0)  (fused ReLU backward,
                           replaces threshold_backward of autograd)                                        */
  float* stats;         /* optional fp32 [grid][2][cout] (grid <= SM count; dpt_conv_plan_info): per-CTA partial sums /
                           sums of squares of the fp32 result per output channel (BatchNorm batch statistics).  No
                           atomics, fixed summation order.  A CTA writes only the tile_n channels of its N tile:
                           channel ch is found in rows ch/tile_n + k*(cout/tile_n), k < grid/(cout/tile_n), which is what
                           dpt_bn_finalize(rows_per_tile = grid/(cout/tile_n), nt = tile_n) sums.  Excludes bias, relu,
                           mask and split / shuffled stores.                                                          */
  void* pool_dst;       /* optional bf16 [batch, h/2, w/2, cout]: nn.MaxPool2d(2,2) of the bias+ReLU result, written by
                           the same epilogue (model/unet_parts.py:26,30-39 without a pooling kernel); needs bias, relu,
                           one NHWC destination, even h and w                                                 */
  int tile_s;           /* 0 = auto; else sub-tiles per CTA tile (tile = 16 rows x 8*tile_s pixels)        */
  int tile_n;           /* 0 = auto; else N tile (multiple of 32, <= 256)                                  */
  int src_h[2], src_w[2]; /* 0 = the source is exactly [batch, h, w, src_c]; else its real height / width: the GEMM reads  */
  int src_y0[2], src_x0[2]; /* the [h, w] window whose top-left corner is (src_y0, src_x0) and sees zeros around it - the
                           CenterCrop of a skip tensor (model/unet_parts.py:58-74) as a tensor-map view, no copy
                           (DPT_CONV_3X3 / DPT_CONV_1X1 sources)                                              */
  int cluster;          /* 0 = auto (DPT_CONV_CLUSTER, default 1 = off), 1 = off, 2 / 4: thread-block cluster whose CTAs
                           share every streamed filter tile through TMA multicast (3x3, N tile 128, 64-channel K
                           chunks).  Only taken when `weight` is plain device memory of the current GPU.       */
} dpt_conv_desc;

typedef struct dpt_conv_plan dpt_conv_plan;  /* tensor maps + launch geometry, built once per (shapes, pointers) */
int dpt_conv_plan_create(const dpt_conv_desc* desc, dpt_conv_plan** out);
int dpt_conv_plan_launch(const dpt_conv_plan* plan, void* stream);
void dpt_conv_plan_destroy(dpt_conv_plan* plan);
/* Introspection for tests / profiles: writes {grid, smem_bytes, tile_s, tile_n, a_stages, b_stages, b_resident,
 * acc_bufs, chunk_ch} */
int dpt_conv_plan_info(const dpt_conv_plan* plan, int info[9]);
int dpt_conv_plan_cluster(const dpt_conv_plan* plan); /* CTAs per cluster the plan launches with (1 = no cluster) */

/* ------------------------------------------------------------------------------------------------------
 * Weight gradients (cuDNN backward-filter in the reference's autograd pass, utils/train_utils.py:69):
 *   dw[m*stride_m + n*stride_n + t*stride_t] = scale * sum_pixels z[pixel][m] * x[pixel + tap t][n]
 * 3x3 conv : z = dZ [B,h,w,C_out], x = layer input (1 or 2 concatenated sources) [B,h,w,C_in],
 *            strides (C_in*9, 9, 1) give PyTorch's [C_out][C_in][3][3].
 * convT 2x2: z = layer input [B,h,w,C_in], x = dY [B,2h,2w,C_out] (mode DPT_CONV_2X2S2),
 *            strides (C_out*4, 4, 1) give PyTorch's [C_in][C_out][2][2].
 * Split-K partials go to caller-provided scratch (dpt_wgrad_scratch_bytes); a second kernel reduces them
 * in a fixed order, so results are bit-reproducible.
 * ------------------------------------------------------------------------------------------------------ */
typedef struct dpt_wgrad_desc {
  int mode;            /* DPT_CONV_3X3, DPT_CONV_3X3_TAPBOX or DPT_CONV_2X2S2                              */
  int batch, h, w;     /* pixel space of z                                                                 */
  const void* z;       /* bf16 NHWC                                                                        */
  int z_c;             /* multiple of 32                                                                   */
  int num_x;
  const void* x[2];
  int x_c[2];          /* multiples of 16                                                                  */
  float* dw;
  long long dw_stride_m, dw_stride_n, dw_stride_t;
  float scale;         /* 0 is read as 1                                                                   */
  void* scratch;
  size_t scratch_bytes;
  int ksplit;          /* 0 = auto                                                                         */
  int n_valid;         /* 0 = all; else only x channels [0, n_valid) are written (zero-padded first layer) */
  int accumulate;      /* 1: dw += result (gradient accumulation over micro-batches), 0: dw = result       */
  float* dbias;        /* may be NULL: fp32 bias gradient = per-channel sum of the output gradient (3x3: z, [z_c];   */
                       /* ConvTranspose: x, [x_c]), produced by the same two kernels (same scale / accumulate)       */
} dpt_wgrad_desc;
typedef struct dpt_wgrad_plan dpt_wgrad_plan;
size_t dpt_wgrad_scratch_bytes(const dpt_wgrad_desc* desc);
int dpt_wgrad_plan_create(const dpt_wgrad_desc* desc, dpt_wgrad_plan** out);
int dpt_wgrad_plan_launch(const dpt_wgrad_plan* plan, void* stream);
void dpt_wgrad_plan_destroy(dpt_wgrad_plan* plan);

/* ------------------------------------------------------------------------------------------------------
 * Symmetric memory on the CUDA VMM API + NVLS multicast objects (csrc/comm/vmm.cu).  File descriptors are
 * exchanged by the caller (AF_UNIX / SCM_RIGHTS).  Sizes must come from dpt_vmm_round_size.
 * ------------------------------------------------------------------------------------------------------ */
int dpt_vmm_query(int device, int* vmm_supported, int* multicast_supported);
int dpt_vmm_round_size(int device, size_t bytes, int multicast_world, size_t* rounded);
int dpt_vmm_alloc(int device, size_t bytes, void** ptr, unsigned long long* handle, int* fd);
int dpt_vmm_import(int device, int fd, size_t bytes, void** ptr, unsigned long long* handle);
int dpt_vmm_map_local(int device, unsigned long long handle, size_t bytes, void** ptr);
int dpt_vmm_grant(void* ptr, size_t bytes, int device); /* one process, several devices: `device` may use the mapping too */
int dpt_vmm_unmap(void* ptr, size_t bytes);
int dpt_vmm_release(unsigned long long handle);
int dpt_close_fd(int fd);
int dpt_mc_create(int world, size_t bytes, unsigned long long* mc, int* fd);
int dpt_mc_import(int fd, unsigned long long* mc);
int dpt_mc_add_device(unsigned long long mc, int device);
int dpt_mc_bind(unsigned long long mc, int device, unsigned long long mem_handle, size_t bytes);
int dpt_mc_map(unsigned long long mc, int device, size_t bytes, void** ptr);

/* ------------------------------------------------------------------------------------------------------
 * Bandwidth-bound kernels (ATen elementwise / pooling / cat kernels in the reference).
 * ------------------------------------------------------------------------------------------------------ */
/* fp32 NCHW [B,C<=16,H,W] -> bf16 NHWC padded to 16 channels (utils/train_utils.py:60). */
int dpt_pack_input(const float* src, void* dst, int batch, int channels, int h, int w, void* stream);
/* Same with the source type chosen at run time: src_dtype 0 = float32 (scale usually 1), 1 = uint8 raw pixels (scale 1/255:
 * the normalisation of utils/dataloading.py:40 done on the device, a quarter of the host->device bytes). */
int dpt_pack_input_ex(const void* src, int src_dtype, float scale, void* dst, int batch, int channels, int h, int w,
                      void* stream);
/* nn.MaxPool2d(2,2) on NHWC bf16 (model/unet_parts.py:26). */
int dpt_maxpool2x2(const void* x, void* y, int batch, int h, int w, int c, void* stream);
/* dz = (d_skip + route(d_pool -> first arg-max)) * (act > 0); d_skip may be NULL and may alias dz. */
int dpt_pool_bwd_fused(const void* act, const void* d_pool, const void* d_skip, void* dz, int batch, int h, int w,
                       int c, int relu_mask, void* stream);
/* out[c] = scale * sum_pixels x[pixel][c]  (bias gradients), deterministic; scratch from the _bytes call. */
size_t dpt_channel_sum_scratch_bytes(int c);
int dpt_channel_sum(const void* x, float* out, long long pixels, int c, float scale, void* scratch,
                    size_t scratch_bytes, void* stream);
/* As above; accumulate = 1 adds to out instead of overwriting it (micro-batch gradient accumulation). */
int dpt_channel_sum_ex(const void* x, float* out, long long pixels, int c, float scale, int accumulate, void* scratch,
                       size_t scratch_bytes, void* stream);
/* Same result as dpt_channel_sum, but the partial sums go to an array built into the library instead of
 * caller-provided scratch, for hosts that may not allocate device memory after start-up (fixed memory
 * budgets, plug-ins without access to the host framework's allocator, loops that cannot afford the
 * device-wide synchronisation of cudaMalloc).  This one has no counterpart in the reference; it exists for
 * embedders of this library.
 *  - limit: c <= DPT_CHANNEL_SUM_STATIC_MAX_C (even); wider tensors are rejected with an error.
 *  - cost: 296 * DPT_CHANNEL_SUM_STATIC_MAX_C floats (592 KiB) of device memory on every GPU that loads the
 *    library, held until the process exits.
 *  - the first call on each device looks up the array's device address (cudaGetSymbolAddress); make that call
 *    during set-up, before any stream capture on that device.  The array used is the CURRENT device's: the stream
 *    must belong to the current device (checked; a foreign stream is refused).
 *  - one array per GPU, shared by all calls: calls that overlap on the same GPU race; order them on one
 *    stream.  No zeroing is needed between calls (every slot read is first rewritten).
 *  - allocates nothing (no cudaMalloc / cudaFree / allocating library routine). */
#define DPT_CHANNEL_SUM_STATIC_MAX_C 512
int dpt_channel_sum_static(const void* x, float* out, long long pixels, int c, float scale, void* stream);

/* Kernel-side bf16 weight layouts from the fp32 master parameters; `table` lives in DEVICE memory. */
typedef struct dpt_prep_entry {
  const float* w;     /* kind 0: [cout][cin][3][3]; kind 1: [cin][cout][2][2]                               */
  const float* bias;  /* kind 1 only (may be NULL)                                                          */
  void* wf;           /* kind 0: bf16 [9][cout][cin_pad]; kind 1: bf16 [1][4*cout][cin]                     */
  void* wd;           /* kind 0: bf16 [9][cin][cout] (taps flipped) or NULL; kind 1: bf16 [4][cin][cout]    */
  float* bias4;       /* kind 1: fp32 [4*cout] bias replicated per output quadrant                          */
  int cout, cin, cin_pad, kind;
  const float* oscale; /* kind 0, may be NULL: fp32 [cout] factor folded into wf (evaluation-mode BatchNorm:
                          gamma * rsqrt(running_var + eps); the shift goes in as the conv's epilogue bias)         */
} dpt_prep_entry;
int dpt_prep_weights(const dpt_prep_entry* device_table, int n_entries, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Output head + loss (model/unet_model.py:10-11,59-60 and utils/utils.py:9-25), fp32.
 * state (8 floats): bce_sum, I, sum p, sum [t==1], loss, pixel count, bce mean, dice term.
 * target_dtype: 0 = float32, 1 = uint8.
 * ------------------------------------------------------------------------------------------------------ */
int dpt_head_blocks(void); /* partial arrays hold dpt_head_blocks() rows */
int dpt_head_sigmoid_fwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                         float* prob, float* partial /* [blocks][4] */, long long pixels, int c, void* stream);
int dpt_head_finalize(const float* const* partials, const int* n_blocks, int n_parts, double total_pixels,
                      float dice_weight, float* state, float* loss_log, int log_slot, void* stream);
int dpt_head_sigmoid_bwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                         const float* state, void* dx, float* partial_w /* [blocks][c+1] */, float* dw, float* db,
                         long long pixels, int c, float gscale, float dice_weight, int relu_mask, void* stream);
/* As above; accumulate = 1 adds to dw / db instead of overwriting them. */
int dpt_head_sigmoid_bwd_ex(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                            const float* state, void* dx, float* partial_w, float* dw, float* db, long long pixels, int c,
                            float gscale, float dice_weight, int relu_mask, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * North-star model family (the upstream DoubleConv / Up / OutConv blocks the reference documents in
 * model/modelsummary.txt:153-247; cuDNN batch-norm / ATen upsample_bilinear2d / softmax + NLL kernels there).
 * bn = fp32 [4][c]: mean, invstd, scale = gamma*invstd, shift = beta - mean*scale.
 * ------------------------------------------------------------------------------------------------------ */
size_t dpt_bn_scratch_bytes(int c); /* forward needs this many bytes, backward this + 2*c floats */
/* Training-mode BatchNorm2d (+ReLU) of a raw conv output: batch statistics (two-stage, deterministic), running
 * statistics update (momentum, unbiased variance), y = relu(y_raw*scale + shift).  y may alias y_raw only if the
 * backward pass is not needed (it re-reads y_raw). */
int dpt_bn_forward_train(const void* y_raw, void* y, const float* gamma, const float* beta, float* running_mean,
                         float* running_var, float* bn, float* sums_out, long long pixels, int c, float eps, float momentum,
                         int relu, void* scratch, size_t scratch_bytes, void* stream);
int dpt_bn_forward_eval(const void* y_raw, void* y, const float* gamma, const float* beta, const float* running_mean,
                        const float* running_var, float* bn, long long pixels, int c, float eps, int relu, void* stream);
/* The pieces of dpt_bn_forward_train for a conv that already produced the statistics in its epilogue
 * (dpt_conv_desc.stats): finalize sums the per-CTA rows (see there; a plain [rows][2][c] array is nt = c) into bn and
 * updates the running statistics; apply writes y = relu(y_raw*scale + shift) and, if pooled != NULL, nn.MaxPool2d(2,2)
 * of y in the same pass ([batch, h/2, w/2, c]; the Down blocks' pooling, model/unet_parts.py:26 in the reference's
 * own model).  c is 16 times a power of two; h and w must be even when a pooled output is asked for. */
int dpt_bn_finalize(const float* partial, int rows_per_tile, int nt, const float* gamma, const float* beta,
                    float* running_mean, float* running_var, float* bn, long long pixels, int c, float eps, float momentum,
                    void* stream);
int dpt_bn_apply(const void* y_raw, void* y, void* pooled, const float* bn, int batch, int h, int w, int c, int relu,
                 void* stream);
/* bn <- running statistics (evaluation mode): bn[2] = gamma*rsqrt(var+eps) folds into the conv weights through
 * dpt_prep_entry.oscale, bn[3] = beta - mean*bn[2] is the conv's epilogue bias. */
int dpt_bn_eval_prepare(const float* gamma, const float* beta, const float* running_mean, const float* running_var,
                        float* bn, int c, float eps, void* stream);
/* dz = d(loss)/d(y_raw) given dy = d(loss)/d(y) (ReLU mask recomputed from y_raw and bn); dgamma, dbeta (+= if accumulate) */
int dpt_bn_backward(const void* y_raw, const void* dy, void* dz, const float* bn, float* dgamma, float* dbeta,
                    long long pixels, int c, int accumulate, void* scratch, size_t scratch_bytes, void* stream);
/* nn.Upsample(scale_factor=2, mode='bilinear', align_corners=True): [B,h,w,c] -> [B,2h,2w,c]; backward is a gather */
int dpt_bilinear_up2x_fwd(const void* in, void* out, int batch, int h, int w, int c, void* stream);
int dpt_bilinear_up2x_bwd(const void* dout, void* din, int batch, int h, int w, int c, void* stream);
/* OutConv (1x1, c -> classes) + softmax + cross-entropy + Dice loss; logits_out fp32 planar [B][classes][hw] or NULL.
 * state (8 floats): ce_sum, sum p_target, -, -, loss, pixel count, ce mean, dice loss.  target: class ids. */
int dpt_head_softmax_fwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                         float* logits_out, float* partial, float* state, long long pixels, long long hw, int c, int classes,
                         void* stream);
/* Loss over several forward calls that were given state = NULL (micro-batches of a pipeline stage; replicas of
 * DataParallel after an all-reduce of the partial arrays): sums their partials, total_pixels = pixels of the whole
 * batch.  The backward kernel reads only state[5] (that pixel count), so it may run before or after this. */
int dpt_head_softmax_finalize(const float* const* partials, int n_parts, double total_pixels, float* state, void* stream);
size_t dpt_head_softmax_scratch_bytes(long long pixels, int c, int classes);
int dpt_head_softmax_bwd(const void* x, const float* w, const float* b, const void* target, int target_dtype,
                         const float* state, void* dx, float* dw, float* db, long long pixels, int c, int classes, float gscale,
                         int accumulate, void* scratch, size_t scratch_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Optimizers over a flat fp32 arena (torch.optim.Adam at utils/train_utils.py:45; RMSprop for the
 * north-star recipe).  opt state = 8 floats in device memory: lr, step, 1-b1^t, 1-b2^t, b1, b2, -, -.
 * ------------------------------------------------------------------------------------------------------ */
enum { DPT_OPT_ADAM = 0, DPT_OPT_RMSPROP = 1 };
int dpt_opt_init(float* state, float lr, int step, float beta1, float beta2, void* stream);
int dpt_opt_set_lr(float* state, float lr, void* stream);
int dpt_opt_tick(float* state, void* stream); /* step += 1 and refresh the bias corrections */
int dpt_adam_step(float* p, float* g, float* m, float* v, long long n, const float* state, float eps,
                  float weight_decay, float grad_scale, int zero_grad, void* stream);
int dpt_rmsprop_step(float* p, float* g, float* sq, float* buf, long long n, const float* state, float alpha, float eps,
                     float weight_decay, float momentum, float grad_scale, int zero_grad, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Fused collectives over NVLink peer memory (DDP Reducer + NCCL all-reduce + optimizer of
 * utils/train_utils.py:196,224-225; DataParallel reduce/broadcast of utils/train_utils.py:98,138,144).
 * A team is `world` ranks (processes or devices of one process) that mapped each other's buffers.
 * ------------------------------------------------------------------------------------------------------ */
#define DPT_MAX_PEERS 8
#define DPT_COMM_MAX_BLOCKS 128
/* bytes of a signal pad (uint32 [DPT_COMM_MAX_BLOCKS][DPT_MAX_PEERS]) and of the local epoch buffer; both
 * must start zeroed */
#define DPT_PAD_BYTES (DPT_COMM_MAX_BLOCKS * DPT_MAX_PEERS * 4)
#define DPT_EPOCH_BYTES (DPT_COMM_MAX_BLOCKS * 4)
typedef struct dpt_team {
  int rank, world;
  void* pads[DPT_MAX_PEERS]; /* pads[r]: rank r's signal pad as mapped into this rank's address space      */
  void* epoch;               /* this rank's private epoch buffer                                            */
} dpt_team;
enum { DPT_AR_TWO_SHOT = 0, DPT_AR_ONE_SHOT = 1, DPT_AR_MULTIMEM = 2 };
typedef struct dpt_fused_ar_desc {
  void* grads[DPT_MAX_PEERS];  /* fp32 gradient arena of every rank (this rank's own included)              */
  void* params[DPT_MAX_PEERS]; /* fp32 parameter arena of every rank                                         */
  void* mc_grads;              /* NVLS multicast addresses of the same arenas (DPT_AR_MULTIMEM only)         */
  void* mc_params;
  float* state0;               /* Adam exp_avg / RMSprop square_avg  (local, full length)                    */
  float* state1;               /* Adam exp_avg_sq / RMSprop momentum buffer (may be NULL for RMSprop)        */
  const float* opt_state;
  long long n;                 /* elements, multiple of 4                                                    */
  int algo, optimizer, blocks;
  float eps, weight_decay, grad_scale, alpha, momentum;
} dpt_fused_ar_desc;
int dpt_fused_allreduce_opt(const dpt_team* team, const dpt_fused_ar_desc* desc, void* stream);
int dpt_team_bcast(const dpt_team* team, void* const* bufs, long long n, int root, int blocks, void* stream);
/* Broadcast through the NVLS multicast mapping `mc` of the ranks' buffers: the root streams its copy (`root_src`, its
 * unicast address) once with multimem.st and the NVSwitch writes every rank's copy (DDP's constructor broadcast and the
 * per-step BatchNorm-buffer sync, utils/train_utils.py:196). */
int dpt_team_bcast_multimem(const dpt_team* team, const void* root_src, void* mc, long long n, int root, int blocks,
                            void* stream);
int dpt_team_allreduce_sum(const dpt_team* team, void* const* bufs, long long n, float scale, int blocks,
                           void* stream);
int dpt_enable_peer_access(int device, int peer);
/* Symmetric buffers for multi-process teams (what c10d/NCCL buffer registration gives the reference's DDP,
 * train.py:58): zeroed cudaMalloc + CUDA IPC handle (64 opaque bytes) export / import. */
int dpt_symm_alloc(size_t bytes, void** out);
int dpt_symm_free(void* p);
int dpt_ipc_export(void* p, unsigned char handle[64]);
int dpt_ipc_import(const unsigned char handle[64], void** out);
int dpt_ipc_close(void* p);

/* ------------------------------------------------------------------------------------------------------
 * Launch programs: a training / evaluation step as a fixed array of op records, executed natively
 * (replaces the Python interpreter + autograd engine + DataParallel's parallel_apply threads of
 * utils/train_utils.py:59-70,138).  Argument slots per kind are documented in csrc/runtime/program.cu and
 * mirrored by distributedpytorch_b200/engine/program.py.
 * ------------------------------------------------------------------------------------------------------ */
enum {
  DPT_OP_CONV = 0, DPT_OP_WGRAD, DPT_OP_PACK_INPUT, DPT_OP_MAXPOOL, DPT_OP_POOL_BWD, DPT_OP_CHANNEL_SUM,
  DPT_OP_PREP_WEIGHTS, DPT_OP_HEAD_FWD, DPT_OP_HEAD_FINALIZE, DPT_OP_HEAD_BWD, DPT_OP_OPT_TICK, DPT_OP_ADAM,
  DPT_OP_RMSPROP, DPT_OP_FUSED_AR, DPT_OP_BCAST, DPT_OP_ALLREDUCE_SUM, DPT_OP_MEMSET, DPT_OP_MEMCPY,
  DPT_OP_EVENT_RECORD, DPT_OP_STREAM_WAIT, DPT_OP_BN_FWD_TRAIN, DPT_OP_BN_FWD_EVAL, DPT_OP_BN_BWD, DPT_OP_BILINEAR_FWD,
  DPT_OP_BILINEAR_BWD, DPT_OP_HEAD_SOFTMAX_FWD, DPT_OP_HEAD_SOFTMAX_BWD, DPT_OP_BN_FINALIZE, DPT_OP_BN_APPLY,
  DPT_OP_BN_EVAL_PREPARE, DPT_OP_HEAD_SOFTMAX_FINALIZE, DPT_OP_BCAST_MULTIMEM, DPT_OP__COUNT
};
typedef struct dpt_op {
  int kind;      /* DPT_OP_*                                   */
  int stream;    /* 0 = main (compute) stream, 1 = aux (communication) stream */
  void* p[12];
  long long i[8];
  float f[6];
} dpt_op;
typedef struct dpt_program dpt_program;
int dpt_program_create(int device, int n_events, dpt_program** out);
void dpt_program_destroy(dpt_program* program);
int dpt_program_add(dpt_program* program, const dpt_op* op); /* returns the op index, or -1 */
int dpt_program_size(const dpt_program* program);
/* `stream` waits for the program's event `event` as last recorded by a DPT_OP_EVENT_RECORD op (input staging). */
int dpt_program_stream_wait(const dpt_program* program, int event, void* stream);
/* Launch ops [first, last) (last < 0: to the end).  Enqueues only; never synchronises. */
int dpt_program_run(const dpt_program* program, int first, int last, void* stream_main, void* stream_aux);
/* Run one program per device concurrently from persistent native worker threads (one per device). */
int dpt_run_parallel(int n, dpt_program* const* programs, const int* first, const int* last,
                     void* const* streams_main, void* const* streams_aux);

#ifdef __cplusplus
}
#endif
#endif /* DPT_B200_H_ */
