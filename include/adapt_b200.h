/*
 * adapt_b200.h -- C ABI of the B200 stage executor behind ADAPT/DEFER's pipeline-partitioned
 * ResNet-50 inference path.
 *
 * The reference (pure Python, /root/reference) has no FFI of its own: its operator boundary is
 * the Python surface DEFER / Node, and the arithmetic is delegated to TensorFlow through
 * `model.predict` (reference src/node.py:177).  This header is the boundary a maintainer would
 * bind instead of that call (ctypes stub in INTEGRATION.md).  Every entry point names the
 * reference interface it stands in for.
 *
 * Conventions
 *   - plain C: pointers, sizes, ints.  No torch / C++ types cross this boundary.
 *   - every function returns 0 on success and a negative ADAPT_E* code on failure;
 *     adapt_last_error() returns a thread-local human-readable message for the last failure.
 *   - device pointers are ordinary CUDA device pointers (e.g. torch.Tensor.data_ptr()).
 *   - `stream` is a cudaStream_t passed as void* (0 = legacy default stream).
 *   - a stage handle is owned by one host thread at a time (not internally locked).
 *   - no CPU fallback exists: without a CUDA device every compute entry point fails.
 */
#ifndef ADAPT_B200_H_
#define ADAPT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADAPT_OK 0
#define ADAPT_EINVAL (-1)   /* bad argument / inconsistent program            */
#define ADAPT_ECUDA (-2)    /* a CUDA runtime or driver call failed           */
#define ADAPT_ENODEV (-3)   /* no usable sm_100 device                        */
#define ADAPT_ENOMEM (-4)   /* device allocation failed                       */
#define ADAPT_ESTATE (-5)   /* call not valid in the handle's current state   */

/* numeric mode of a stage (north_star: bf16 activations + fp32 accumulate; fp32 check mode) */
#define ADAPT_MODE_BF16 0
#define ADAPT_MODE_FP32_CHECK 1

/* op kinds of a stage program.  One op = one fused group of Keras layers. */
#define ADAPT_OP_INGEST 0   /* fp32 NHWC input image -> working dtype, channel-padded (input_1)          */
#define ADAPT_OP_CONV 1     /* [ZeroPadding2D+]Conv2D+bias+BatchNormalization[+Add][+ReLU]; Dense = 1x1  */
#define ADAPT_OP_MAXPOOL 2  /* pool1_pad + pool1_pool                                                    */
#define ADAPT_OP_GAP 3      /* avg_pool (GlobalAveragePooling2D)                                         */
#define ADAPT_OP_SOFTMAX 4  /* softmax activation of `predictions`                                       */
#define ADAPT_OP_AFFINE 5   /* stand-alone BatchNormalization and/or ReLU (a cut separated it from its  */
                            /* convolution): y = relu?(x * scale[c] + bias[c]); weight = -1: ReLU only  */
#define ADAPT_OP_ZEROPAD 6  /* stand-alone ZeroPadding2D (a cut at conv1_pad / pool1_pad)               */
#define ADAPT_OP_CONV_CHAIN 7 /* bf16 mode: convN_blockB_2_conv+bn+relu (3x3 / stride 1 / pad 1, C -> C) chained  */
                            /* with _3_conv+bn+add+out (1x1, C -> 4C, shortcut, ReLU) in one kernel; the         */
                            /* _2_relu tensor never reaches HBM.  C in {64, 128}, map width >= 28.               */
                            /* weight = the 3x3's entry, weight2 = the 1x1's; Cin = C, Cout = 4C                  */

#define ADAPT_OP_STEM_POOL 8 /* bf16 mode: conv1_conv+bn+relu in its space-to-depth form (4x4 / stride 1 / VALID over   */
                            /* 16-channel pixels stored 128 wide, 64 output channels) fused with pool1_pad +        */
                            /* pool1_pool (3x3 / stride 2 / zero pad 1): the 112x112x64 map never reaches HBM.     */
                            /* H, W, Cin = stored input; Ho, Wo = the CONVOLUTION's output size (the op's output    */
                            /* tensor is (Ho/2, Wo/2, 64)); kh = kw = 4, stride 1, pad 0, relu 1                    */

/* tensor roles inside a stage program */
#define ADAPT_T_INTERNAL 0  /* library-owned scratch activation                       */
#define ADAPT_T_INPUT 1     /* stage input, caller-owned, bound per slot (bind_io)    */
#define ADAPT_T_OUTPUT 2    /* stage output, caller-owned, bound per slot (bind_io)   */

/* tensor element types */
#define ADAPT_DT_WORK 0     /* bf16 in ADAPT_MODE_BF16, fp32 in ADAPT_MODE_FP32_CHECK */
#define ADAPT_DT_F32 1
#define ADAPT_DT_U8 2       /* stage INPUT only: raw uint8 image (see ADAPT_IN_U8_RGB_CAFFE)  */

/* One Conv2D/Dense layer's parameters in Keras layout (what `get_weights()` returns,
 * reference src/dispatcher.py:243) with BatchNormalization already folded on the host into
 * a per-output-channel fp32 scale and bias:  y = conv(x) * scale + bias. */
typedef struct adapt_weight_desc {
    const float* kernel; /* host, HWIO: [kh][kw][cin][cout] fp32 (Dense: [in][out]);  */
                         /* NULL with kh = kw = 0: scale/bias only (ADAPT_OP_AFFINE)  */
    const float* scale;  /* host, [cout]                                             */
    const float* bias;   /* host, [cout]                                             */
    int32_t kh, kw, cin, cout;
} adapt_weight_desc;

typedef struct adapt_tensor_desc {
    int64_t elems_per_image; /* stored elements per image (channel padding included) */
    int32_t dtype;           /* ADAPT_DT_*                                           */
    int32_t role;            /* ADAPT_T_*                                            */
} adapt_tensor_desc;

typedef struct adapt_op_desc {
    int32_t kind;                    /* ADAPT_OP_*                                             */
    int32_t in_tensor;               /* index into the tensor table                            */
    int32_t res_tensor;              /* residual (shortcut) input of a fused Add, or -1        */
    int32_t out_tensor;
    int32_t H, W, Cin;               /* input spatial size and STORED channels per pixel       */
    int32_t Ho, Wo, Cout;            /* output spatial size and channels                       */
    int32_t kh, kw, stride, pad;     /* filter / pool window, stride, symmetric zero padding   */
    int32_t relu;                    /* apply ReLU in the epilogue                             */
    int32_t weight;                  /* index into the weight table (CONV) or -1               */
    int32_t out_f32;                 /* CONV: write fp32 instead of the working dtype (logits) */
    int32_t in_format;               /* INGEST: ADAPT_IN_* (0 elsewhere)                       */
    int32_t weight2;                 /* CONV_CHAIN: weight-table index of the 1x1 (-1 elsewhere) */
} adapt_op_desc;

/* INGEST input formats.  ADAPT_IN_F32 is the reference's queue contract: an fp32 NHWC image that already went
 * through `preprocess_input` (reference test/test.py:20-23).  ADAPT_IN_U8_RGB_CAFFE (no counterpart in the
 * reference) takes the raw uint8 RGB image and applies Keras' caffe-mode preprocessing -- RGB -> BGR,
 * minus the ImageNet channel means (103.939, 116.779, 123.68), no scaling -- inside the ingest kernel, so an input
 * costs 150,528 bytes of PCIe instead of 602,112. */
#define ADAPT_IN_F32 0
#define ADAPT_IN_U8_RGB_CAFFE 1

typedef struct adapt_stage adapt_stage_t;

/* Library / device probes.  adapt_device_count() is 0 on a machine without a GPU. */
const char* adapt_version(void);
const char* adapt_last_error(void);
int adapt_device_count(void);

/* Replaces: `model_from_json` + `set_weights` on the worker (reference src/node.py:40,45,77-78)
 * and the weight shipping of `_send_full_configuration` (reference src/dispatcher.py:223-264).
 * Uploads ALL layers' weights of the model to `device` once (so a later re-slice is metadata
 * only) and installs the stage program `ops` over `tensors` (the layer range (start, end] that
 * `DEFER._partition` / `construct_model` would cut, reference src/dispatcher.py:39-53,
 * src/dag_util.py:50-62). */
int adapt_stage_create(int device, int dtype_mode, int max_batch, const adapt_weight_desc* weights, int n_weights,
                       const adapt_tensor_desc* tensors, int n_tensors, const adapt_op_desc* ops, int n_ops,
                       adapt_stage_t** out);

/* Replaces: re-configuring a worker with another partition (`_send_full_configuration`,
 * reference src/dispatcher.py:223-264) after a failure; weights are already resident. */
int adapt_stage_reslice(adapt_stage_t* stage, const adapt_tensor_desc* tensors, int n_tensors,
                        const adapt_op_desc* ops, int n_ops);

/* Replaces: the socket payload hand-off into / out of `model.predict` (reference
 * src/node.py:137-179).  Binds the caller-owned input and output buffers of double-buffer
 * `slot` (0 or 1).  `out_ptr` may be a peer-mapped pointer on the next stage's GPU: the last
 * layer's epilogue then stores the cut tensor straight over NVLink. */
int adapt_stage_bind_io(adapt_stage_t* stage, int slot, void* in_ptr, void* out_ptr);

/* Optional warm-up, part of (re)configuration: build the launch records of (`slot`, `batch`) and
 * capture them into a CUDA graph.  May synchronise the device; call it before streaming starts.
 * A batch size that was never prepared still works through enqueue(), as plain launches.
 * (Reference counterpart: Keras builds its predict function on the first `predict` call.) */
int adapt_stage_prepare(adapt_stage_t* stage, int slot, int batch);

/* Replaces: `model.predict(inpt)` (reference src/node.py:177).  Asynchronously runs the stage
 * program on `batch` images from slot's input buffer into slot's output buffer on `stream`. */
int adapt_stage_enqueue(adapt_stage_t* stage, int slot, int batch, void* stream);

/* Debug/test tap: copy an INTERNAL tensor (first `bytes` bytes) to host memory after
 * synchronising `stream`.  No reference counterpart (the reference never inspects activations). */
int adapt_stage_read_tensor(adapt_stage_t* stage, int tensor, void* host_dst, size_t bytes, void* stream);

/* Measurement aid: run the ops of (`slot`, `batch`) one by one `reps` times on `stream`, each bracketed
 * by CUDA events, and return the mean duration of every op (ms) in ms_per_op[n_ops].  Synchronises. */
int adapt_stage_profile(adapt_stage_t* stage, int slot, int batch, void* stream, int reps, float* ms_per_op,
                        int n_ops);

/* Measurement aid: the ops whose kind is in `kind_mask` (bit ADAPT_OP_* set) of (`slot`, `batch`) captured into their own CUDA graph -- same
 * launches and launch attributes as inside the stage's graph, so consecutive convolutions keep their programmatic
 * dependent launch overlap -- and timed over `reps` graph launches with one pair of CUDA events on `stream`.
 * *ms_per_rep = device time of that class per step; *n_launches = launches of that kind.  Overwrites the stage's
 * internal activations with meaningless values (inputs of the skipped ops are stale); synchronises.
 * No reference counterpart (bench.py's roofline accounting). */
int adapt_stage_time_kind(adapt_stage_t* stage, int slot, int batch, void* stream, int kind_mask, int reps,
                          float* ms_per_rep, int* n_launches);

/* Number of kernel launches one enqueue() of `batch` images performs (for bench accounting). */
int adapt_stage_num_launches(adapt_stage_t* stage, int batch);

void adapt_stage_destroy(adapt_stage_t* stage);

/* Single-layer entry point (kernel-level parity tests and micro-benchmarks):
 *   out = act((conv(in, w) * scale + bias) (+ residual)),  all pointers DEVICE pointers.
 *   in   : bf16 NHWC [batch][H][W][Cin]      (Cin a multiple of 8)
 *   w    : bf16      [Cout][kh][kw][Cin]
 *   out  : bf16 NHWC [batch][Ho][Wo][Cout], or fp32 when out_is_f32 != 0
 * a_mode: -1 auto, 0 tiled TMA, 1 im2col TMA, 2 software gather, 3 row reuse (3x3 / stride 1 / pad 1 without
 * residual only; other shapes fall back to im2col), 4 stem by column phase (4x4 / stride 1 / VALID, 16 -> 64
 * channels, W a multiple of 4; other shapes fall back).  block_n: 0 auto, 64/128/256.
 * The optional cluster variant (ADAPT_CLUSTER=2/4) multicasts the weight tile; it is taken only when `w` is plain
 * device memory allocated on `device` itself (TMA multicast loads here read only this GPU's memory, since on B200 a
 * TMA multicast load from another GPU's memory hung the GPU).  Any other `w` runs without multicast.
 * Stands in for one Conv2D+BatchNormalization(+Add)(+ReLU) group inside `model.predict`
 * (reference src/node.py:177). */
int adapt_conv2d_bf16(int device, const void* in, const void* w, const float* scale, const float* bias,
                      const void* residual, void* out, int out_is_f32, int batch, int H, int W, int Cin, int Cout,
                      int kh, int kw, int stride, int pad, int relu, int a_mode, int block_n, void* stream);

/* Single-block entry point of ADAPT_OP_CONV_CHAIN (kernel-level parity tests and micro-benchmarks), all pointers
 * DEVICE pointers on `device`:
 *   mid = relu(conv3x3(in, w2) * scale1 + bias1)  rounded to bf16 (never stored),
 *   out = relu(conv1x1(mid, w3) * scale2 + bias2 + residual)
 *   in : bf16 NHWC [batch][H][W][C]      w2 : bf16 [C][3][3][C]      w3 : bf16 [4C][C]
 *   residual, out : bf16 NHWC [batch][H][W][4C];  C in {64, 128}, W >= 28.
 * Stands in for the layers convN_blockB_2_conv ... convN_blockB_out inside `model.predict` (reference src/node.py:177). */
int adapt_conv_chain_bf16(int device, const void* in, const void* w2, const float* scale1, const float* bias1,
                          const void* w3, const float* scale2, const float* bias2, const void* residual, void* out,
                          int batch, int H, int W, int C, void* stream);

/* Single-op entry point of ADAPT_OP_STEM_POOL (kernel-level parity tests and micro-benchmarks), DEVICE pointers:
 *   y   = relu(conv4x4_valid(in, w) * scale + bias)     (never stored)
 *   out = maxpool3x3_stride2(zero_pad1(y))
 *   in : bf16 [batch][H][W = 128][16] (the ingest kernel's space-to-depth image)   w : bf16 [64][4][4][16]
 *   out: bf16 NHWC [batch][Ho/2][Wo/2][64];  Ho, Wo = size of y (multiples of 4, Wo <= 128, Ho <= H - 3).
 * Stands in for the layers conv1_conv ... pool1_pool inside `model.predict` (reference src/node.py:177). */
int adapt_stem_pool_bf16(int device, const void* in, const void* w, const float* scale, const float* bias, void* out, int batch,
                         int H, int W, int Ho, int Wo, void* stream);

/* Classifier head (Dense + softmax) for callers that must not allocate device memory.
 * This entry point has NO counterpart in the reference; it exists for users of this library whose
 * process cannot call cudaMalloc when it needs a prediction: a serving process that handed all of
 * its HBM to its own arenas at start-up (any later allocation would fail), a latency-critical loop
 * that cannot afford the device-wide synchronisation and unbounded duration of an allocation, or a
 * liveness probe of the fail-over logic that has to keep working on a GPU that is out of memory.
 *
 *   probs[b, :] = softmax(in[b, :] @ w^T * scale + bias)      in: bf16 [batch][cin], w: bf16 [cout][cin],
 *                                                              scale/bias: fp32 [cout], probs: fp32 [batch][cout]
 * All pointers are device pointers owned by the caller and must be 16-byte aligned (the kernels use 128-bit
 * loads and TMA; a caller sub-allocating from its own arena has to keep that alignment -- anything else is
 * refused with ADAPT_EINVAL); cin and cout must be multiples of 8.
 * The only intermediate, the fp32 logits, lives in a fixed array inside the library
 * (ADAPT_NOALLOC_MAX_LOGITS floats = 1 MiB per GPU), so the call performs no cudaMalloc / cudaFree and calls no
 * allocating library routine.  The array is part of the library's device code: the CUDA runtime reserves it
 * when it loads that code on a device (with lazy loading, at the first call that touches it there) and keeps it
 * until the process exits -- so make one warm-up call per device while memory is still available.
 *   - LIMIT: batch * cout <= ADAPT_NOALLOC_MAX_LOGITS, otherwise ADAPT_EINVAL with a message.
 *   - The first call on each device looks up the array's device address and sets kernel attributes; make that
 *     first call before starting any stream capture.  First calls racing from several threads are serialised
 *     inside the library.
 *   - Every call uses the same array and leaves its logits there.  Nothing needs clearing (each call
 *     overwrites all batch * cout entries it reads), but two calls whose GPU work overlaps -- different
 *     streams, or different threads -- race on it: serialise them on one stream. */
#define ADAPT_NOALLOC_MAX_LOGITS (256 * 1024)
int adapt_dense_softmax_noalloc(int device, const void* in, const void* w, const float* scale, const float* bias,
                                float* probs, int batch, int cin, int cout, void* stream);

/* ---- stage-to-stage hand-off plumbing (replaces the TCP framing of reference
 *      src/node_state.py:39-161 and the routing of src/dispatcher.py:176-220) ---- */

/* cudaMalloc / cudaFree on `device` (cut buffers and flags must be raw allocations to be
 * exportable over CUDA IPC). */
int adapt_device_alloc(int device, size_t bytes, void** out);
int adapt_device_free(int device, void* ptr);
/* Enable peer access device a <-> b (same process). */
int adapt_peer_enable(int a, int b);
/* CUDA IPC: export a raw allocation as a 64-byte handle / map a peer process's handle. */
int adapt_ipc_export(int device, void* ptr, void* handle64_out);
int adapt_ipc_import(int device, const void* handle64, void** out);
int adapt_ipc_close(int device, void* ptr);
/* Stream-ordered 32-bit flag write / wait-until->= on (possibly peer-mapped) device memory. */
int adapt_stream_write_flag(int device, void* stream, void* flag_ptr, uint32_t value);
int adapt_stream_wait_flag_geq(int device, void* stream, void* flag_ptr, uint32_t value);
/* Async copies on `stream` (host memory should be pinned). */
int adapt_memcpy_h2d(int device, void* dst, const void* src, size_t bytes, void* stream);
int adapt_memcpy_d2h(int device, void* dst, const void* src, size_t bytes, void* stream);
int adapt_memcpy_d2d(int device, void* dst, const void* src, size_t bytes, void* stream);
int adapt_memset(int device, void* dst, int value, size_t bytes, void* stream);
int adapt_stream_sync(int device, void* stream);
/* Non-blocking: *busy_out = 0 when all work queued on `stream` has completed, 1 otherwise. */
int adapt_stream_query(int device, void* stream, int* busy_out);
/* Liveness beacon: a native thread stores CLOCK_MONOTONIC nanoseconds into *word every
 * `interval_us` (immune to the Python GIL).  Replaces the worker's etcd lease/heartbeat
 * (reference src/node_state.py:18-23).  Needs no GPU. */
int adapt_heartbeat_start(volatile int64_t* word, int interval_us, void** handle_out);
int adapt_heartbeat_stop(void* handle);
/* Pin + map an existing host range (e.g. a shared-memory segment) so that async copies and
 * stream-ordered flag writes can target it; *dev_ptr_out is the device-side alias. */
int adapt_host_register(int device, void* host_ptr, size_t bytes, void** dev_ptr_out);
int adapt_host_unregister(int device, void* host_ptr);
/* Streams and timing events owned by the library (a worker uses one compute and one copy stream). */
int adapt_stream_create(int device, void** stream_out);
int adapt_stream_destroy(int device, void* stream);
int adapt_event_create(int device, void** event_out);
int adapt_event_record(int device, void* event, void* stream);
int adapt_event_sync(int device, void* event);
int adapt_event_elapsed_ms(int device, void* start, void* stop, float* ms_out);
/* Non-blocking: *done_out = 1 when `event` has completed, 0 otherwise (per-micro-batch busy times of a stage,
 * folded into the health table for the straggler watchdog -- the reference's `start_time` idea, src/dispatcher.py:193). */
int adapt_event_query(int device, void* event, int* done_out);
int adapt_event_destroy(int device, void* event);
/* Fault injection (tests): occupy `stream` for `microseconds` without touching memory -- a stage that is alive but
 * slow (clock-capped GPU, noisy neighbour).  No reference counterpart. */
int adapt_debug_spin(int device, void* stream, int microseconds);

#ifdef __cplusplus
}
#endif
#endif /* ADAPT_B200_H_ */
