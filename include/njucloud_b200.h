/*
 * njucloud_b200.h -- C ABI of the B200-native hot path behind NJUCloud (view-sonic/Cloud-Server).
 *
 * The reference is pure Python; its arithmetic sits in third-party wheels (OpenCV, Pillow,
 * TensorFlow 1.x).  This library replaces those calls for ONE path only: the batched image
 * preprocess chain of apps/preprocess + construct_inference.py, and forward / backward / update of
 * the CNN graphs that construct_distribute.py builds.  Each entry point cites the reference
 * call it stands in for (paths relative to the reference root).
 *
 * Conventions
 *  - plain C, no exceptions cross the boundary; every function returns NC_OK (0) or a negative
 *    NC_E* code; nc_last_error(ctx) returns the message of the last failure on that ctx
 *    (ctx == NULL: the last failure of a call that had no ctx, thread-local).
 *  - the caller allocates and owns every data buffer.  Pointers are DEVICE pointers unless the
 *    parameter name starts with host_.  The library owns only its handles, small coefficient
 *    tables, workspaces and the flat parameter / gradient / optimizer-state buffers of a net.
 *  - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream).  Calls are
 *    asynchronous on that stream unless stated otherwise.
 *  - a ctx is bound to one device; it is not thread-safe: one ctx per thread / process.
 *    Data parallel training = one process per GPU; gradients are summed either by the host
 *    binding (torch.distributed / NCCL) on the flat gradient buffer between nc_net_backward and
 *    nc_net_apply, or inside the library once a net is attached to a communicator
 *    (nc_net_attach_nccl), which is also what batch-norm layers need.
 */
#ifndef NJUCLOUD_B200_H
#define NJUCLOUD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NC_OK            0
#define NC_EINVAL       -1   /* bad argument (the Python shim raises ValueError)            */
#define NC_ECUDA        -2   /* a CUDA runtime call or kernel launch failed                  */
#define NC_ENOMEM       -3   /* device or host allocation failed                            */
#define NC_EUNSUPPORTED -4   /* valid in the reference but not implemented by this library  */
#define NC_ECONFIG      -5   /* malformed --config JSON / unknown layer (reference: assert) */
#define NC_ECV          -6   /* the reference's cv2 call would raise cv2.error here          */

typedef struct nc_ctx nc_ctx;
typedef struct nc_net nc_net;

/* ---- context -------------------------------------------------------------------------- */
int nc_version(void);
int nc_init(int device, nc_ctx** out);
int nc_destroy(nc_ctx* ctx);
const char* nc_last_error(nc_ctx* ctx);
/* number of kernels this ctx has launched so far (bench.py's gpu_launches) */
int64_t nc_launch_count(nc_ctx* ctx);

/* ---- P1: JPEG decode ---------------------------------------------------------------------
 * Replaces cv2.imread(dir) (apps/preprocess/preprocess.py:28, :40 -- BGR) and PIL.Image.open
 * (apps/construction/util/construct_distribute.py:446, construct_inference.py:312 -- RGB / L) for baseline
 * (sequential Huffman, 8-bit, 1 or 3 components, chroma 4:4:4 / 4:2:2 / 4:2:0) files, bit for bit with the
 * libjpeg-turbo both wheels embed: "islow" integer IDCT, "fancy" triangle chroma upsampling, fixed-point
 * YCbCr -> RGB.  Progressive / arithmetic files: NC_EUNSUPPORTED (decode those on the host).
 *   nc_jpeg_info    header only (host): width, height, components, chroma subsampling factors (1 or 2).
 *   nc_jpeg_decode  count pictures of the SAME width x height: host_jpegs[i] / sizes[i] are the files' bytes in host
 *                   memory.  Huffman decoding runs on the host (host_threads threads, <= 0: all cores); the
 *                   nonzero coefficients cross PCIe (*h2d_bytes, nullable, reports how many bytes did) and
 *                   dequantisation, IDCT, upsampling and colour conversion run on the device.
 *                   out[i] (device) = uint8 [height][width][out_channels]; out_channels 3: channel_order 0 = RGB
 *                   (PIL), 1 = BGR (cv2; a gray file is replicated as cv2.imread(dir) does); out_channels 1: the
 *                   luma plane (what both libraries return for a single-component file).
 *                   Synchronises `stream` once (the pinned staging buffer of the previous call). */
int nc_jpeg_info(const uint8_t* host_jpeg, size_t size, int* width, int* height, int* components, int* hsamp, int* vsamp);
/* Host only: the Huffman stage of nc_jpeg_decode alone, expanded to dense quantised coefficients coefs[block][64]
 * (natural order; component planes back to back, blocks row-major, MCU padded).  coefs NULL: *nblocks only. */
int nc_jpeg_entropy_decode(const uint8_t* host_jpeg, size_t size, int16_t* coefs, int64_t capacity_blocks, int64_t* nblocks);
int nc_jpeg_decode(nc_ctx* ctx, const uint8_t* const* host_jpegs, const size_t* sizes, int count, int width, int height,
                   int out_channels, int channel_order, uint8_t* out, int64_t img_stride, int host_threads,
                   int64_t* h2d_bytes, void* stream);

/* ---- P2+P3+P4: train-side preprocess chain ---------------------------------------------
 * Replaces, fused and batched, for n decoded uploads src[n][h][w][c] (uint8, c == 3 BGR as
 * cv2.imread returns it, or c == 1):
 *   cv2.resize(im1,(28,28),interpolation=cv2.INTER_CUBIC)     apps/preprocess/preprocess.py:29
 *   the BGR->gray of cv2.imread(dir, 0) on the resized picture apps/preprocess/preprocess.py:40
 *   images.astype(float32) * (1/255)    apps/construction/util/construct_distribute.py:503-504
 * img_stride = bytes between consecutive images (>= h*w*c).  mode 0 = OpenCV's own arithmetic
 * (cv2 4.13 with IPP disabled: int32 horizontal pass, float32 SIMD-body / integer-tail
 * vertical pass) -- bit exact.  mode 1 = the arithmetic the opencv-python wheel uses by DEFAULT
 * (Intel IPP, closed source; requirment.txt:14 does not pin the wheel): unquantised float32
 * coefficients, float32 products and sums tap by tap, horizontal pass first, round half to even
 * -- a float model fitted to the live call, 100 % equal on 94 080 probed outputs, gated at 99.9 %.
 * The two modes differ by +-1 on 3-5 % of the outputs.
 * out_u8[n][28][28] and/or out_f32[n][784] (either may be NULL). */
int nc_pp_resize_gray_norm(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c,
                           int64_t img_stride, int mode, uint8_t* out_u8, float* out_f32,
                           void* stream);

/* The resize alone: out[n][28][28][c] = cv2.resize(src[i], (28,28), interpolation=cv2.INTER_CUBIC)
 * (apps/preprocess/preprocess.py:29), the picture preprocess.resize writes back to disk (:30). */
int nc_pp_resize_cubic(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c, int64_t img_stride,
                       int mode, uint8_t* out, void* stream);

/* Same chain with the source in HOST memory (what the REST view has after decoding).  If
 * host_src is pinned (cudaHostAlloc / cudaHostRegister) the kernel gathers the 4x4 taps it
 * needs straight over PCIe; pageable memory is staged through the ctx workspace.  Outputs are
 * device pointers.  Synchronous with respect to host_src on return only if the function says
 * so via *staged (1 = a staging copy was made and host_src may be reused immediately). */
int nc_pp_resize_gray_norm_host(nc_ctx* ctx, const uint8_t* host_src, int n, int h, int w, int c,
                                int64_t img_stride, int mode, uint8_t* out_u8, float* out_f32,
                                int* staged, void* stream);

/* ---- P5: predict-side preprocess chain -------------------------------------------------
 * Replaces apps/construction/util/construct_inference.py:312-330 for n decoded pictures:
 *   Image.open(f).convert('L')        (:312)  rgb2l fixed point; c == 1 is taken as 'L'
 *   im.resize((20,20))                (:315)  Pillow BICUBIC, antialiased, 8bpc two-pass
 *   > 150 -> 254 on a 28x28 zero canvas at offset 4, /255.0   (:317-330)
 * channel_order: 0 = RGB (PIL), 1 = BGR (cv2 buffers).  compat != 0 reproduces the reference's
 * `arr[bizar+i][bizar+j] = 220` statement (:325-326) including its IndexError: status[i] is set
 * to 1 where the reference raises (out_f32 row i is then all zero), else 0.  status may be NULL
 * when compat == 0.  out_small20 (nullable) receives the 20x20 'L' picture as the resize returns it
 * (before the compat statement patches 220s into its working copy). */
int nc_pp_predict_chain(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c,
                        int64_t img_stride, int channel_order, int compat, float* out_f32,
                        uint8_t* out_small20, int* status, void* stream);

/* ---- P2-P5 in one pass: both chains over the same uploads ------------------------------
 * The platform runs both chains over every upload: the train-side one when a dataset is
 * preprocessed (apps/preprocess/preprocess.py:28-30,40 + construct_distribute.py:503-504) and the
 * predict-side one when a picture is classified (construct_inference.py:312-330).  For a batch
 * resident in HBM this entry point produces exactly the bytes of nc_pp_resize_gray_norm(mode 0)
 * (out_train_u8 / out_train_f32, either nullable) and of nc_pp_predict_chain (out_predict_f32,
 * out_small20, status); its single-pass kernel reads each upload from HBM once: the cubic taps of the
 * train-side chain are taken from the rows the predict-side chain streams through shared memory.  The train
 * side reads the buffer as BGR (cv2), channel_order applies to the predict side as above.
 * Two launches by default; with NC_FUSED_CHAINS set in the environment, one single-pass kernel
 * where the shape is inside its plan (3 channels, 16-byte aligned rows, >= ~160 rows, width
 * <= ~2000).  On B200 the two take the same time (the single pass trades the 13 % of HBM traffic
 * it saves for issue slots), which is why it is opt-in.  *fused (nullable) reports which ran;
 * the outputs are identical either way. */
int nc_pp_both_chains(nc_ctx* ctx, const uint8_t* src, int n, int h, int w, int c, int64_t img_stride,
                      int channel_order, int compat, uint8_t* out_train_u8, float* out_train_f32,
                      float* out_predict_f32, uint8_t* out_small20, int* status, int* fused, void* stream);

/* ---- P6-P19: the 14 per-picture ops on 28x28 8-bit gray --------------------------------
 * in/out: [n][28][28] uint8 device buffers (in == out allowed).  One call = one op applied
 * to every picture, as PreprocessView.execute's loop does (apps/preprocess/views.py:118-126).
 * v1 / v2 are the op's value1 / value2 (views.py:123-126).  v_is_int: bit0 = v1 was a Python
 * int, bit1 = v2 was (only adjust_brightness_contrast cares).  host_draws: the uniform [0,1)
 * doubles the reference draws on the host for NC_OP_SALT_PEPPER (4*m per picture, m =
 * int(784*percent), in call order: preprocess.py:163-165, :174-176); NULL otherwise. */
enum nc_op {
    NC_OP_FLIP_UP_DOWN = 0,        /* preprocess.py:33-45   cv2.flip(img, 0)                  */
    NC_OP_FLIP_LEFT_RIGHT = 1,     /* preprocess.py:48-60   cv2.flip(img, 1)                  */
    NC_OP_TRANSPOSE_IMAGE = 2,     /* preprocess.py:63-75   cv2.flip(img, -1)                 */
    NC_OP_BRIGHTNESS_CONTRAST = 3, /* preprocess.py:78-99   p*alpha+beta stored into uint8    */
    NC_OP_MEAN_FILTER = 4,         /* preprocess.py:123-129 cv2.blur(img,(k,k))               */
    NC_OP_GAUSSIAN_BLUR = 5,       /* preprocess.py:132-138 cv2.GaussianBlur(img,(k,k),0)     */
    NC_OP_MEDIAN_FILTER = 6,       /* preprocess.py:141-148 cv2.medianBlur(img,k)             */
    NC_OP_NL_DENOISE = 7,          /* preprocess.py:151-157 cv2.fastNlMeansDenoising(,h,7,21) */
    NC_OP_SALT_PEPPER = 8,         /* preprocess.py:160-188                                   */
    NC_OP_EQUALIZE_HIST = 9,       /* preprocess.py:191-197 cv2.equalizeHist                  */
    NC_OP_CLAHE = 10,              /* preprocess.py:200-207 cv2.createCLAHE().apply           */
    NC_OP_ERODE = 11,              /* preprocess.py:210-217 cv2.erode(img, ones(k,k))         */
    NC_OP_DILATE = 12              /* preprocess.py:220-227 cv2.dilate(img, ones(k,k))        */
    /* random_brightness_contrast (preprocess.py:102-120) draws alpha/beta on the host with
     * Python's `random` and is then NC_OP_BRIGHTNESS_CONTRAST.                               */
};
int nc_pp_op(nc_ctx* ctx, int op_id, const uint8_t* in, uint8_t* out, int n, double v1, double v2,
             int v_is_int, const double* host_draws, int64_t draws_per_image, void* stream);

/* Host-only introspection of the per-shape coefficient tables the kernels consume (no GPU
 * needed; used by the CPU tests to check the host logic against the oracle).
 *  nc_pp_cubic_table: OpenCV INTER_CUBIC taps for one axis src -> 28 (preprocess.py:29):
 *                     ofs[28*4] clamped source indices, coef[28*4] = round(c*2048).
 *  nc_pp_cubic_table_f32: the same taps in the IPP arithmetic of mode 1: ofs[28*4], float32 coef[28*4].
 *  nc_pp_pil_table:   Pillow BICUBIC antialias taps for one axis in_size -> out_size
 *                     (construct_inference.py:315): xmin[out], cnt[out], kk[out*ksize] 22-bit
 *                     fixed point; returns ksize (or a negative error); kk may be NULL to query.
 *  nc_pp_gaussian_q8: cv2.GaussianBlur(k,k,0) 8-bit fixed-point kernel (preprocess.py:134). */
int nc_pp_cubic_table(int src, int32_t* ofs, int16_t* coef);
int nc_pp_cubic_table_f32(int src, int32_t* ofs, float* coef);
int nc_pp_pil_table(int in_size, int out_size, int32_t* xmin, int32_t* cnt, int32_t* kk, int kk_capacity);
int nc_pp_gaussian_q8(int k, int32_t* q);

/* CRC-32C (Castagnoli) of n host bytes, continuing from `crc` (0 to start): the checksum TensorFlow's
 * checkpoint bundles keep per tensor and per index block -- tf.train.Saver() / saver.restore
 * (construct_distribute.py:375, :385-392; construct_inference.py:336-339).  Host-only. */
uint32_t nc_crc32c(const void* host_data, size_t n, uint32_t crc);

/* [n][28][28] uint8 -> [n][784] float32, x * float32(1/255)
 * (DataSet.__init__, construct_distribute.py:503-504). */
int nc_pp_normalise(nc_ctx* ctx, const uint8_t* in, int64_t count, float* out, void* stream);

/* Batch assembly on the device: out[r][:] = src[idx[r]][:] for float32 rows of `width` floats --
 * the `self._images[perm]` / slice of DataSet.next_batch (construct_distribute.py:526-569) with the
 * dataset resident in HBM.  idx is a DEVICE array of int64 row numbers in [0, n_src). */
int nc_gather_rows(nc_ctx* ctx, const float* src, int64_t n_src, int width, const int64_t* idx, int n,
                   float* out, void* stream);
/* numpy raises IndexError for an index outside [0, n_src) (construct_distribute.py:541, :556); the kernel
 * zero-fills such a row and sets a sticky flag.  This call synchronises `stream`, returns NC_EINVAL if any
 * nc_gather_rows on ctx since the previous status call saw such an index, and clears the flag. */
int nc_gather_rows_status(nc_ctx* ctx, void* stream);

/* ---- CNN graph: construct / forward / backward / update --------------------------------
 * nc_net_create replaces cnn() + get_net() (construct_distribute.py:208-281, identical copy in
 * construct_inference.py:53-279): it interprets net_config_json = the value of
 * config["net_config"] ({"middle_layer":[...]}) with the defaults of :222-250 and appends the
 * implicit dense->classes output layer (:262-264).  The reference fixes the input at
 * 28x28x1 (:216) and classes at 10 (:261-264); they are parameters here.
 * Parameters live in ONE flat float32 buffer in variable-creation order (conv: HWIO filter then
 * bias if any; norm: scale then offset; connect: W[in][out] then b; output W then b).
 * Initial values follow weight_variable/bias_variable (:57-87) from `seed`; parity tests
 * overwrite them with nc_net_set_params. */
int nc_net_create(nc_ctx* ctx, const char* net_config_json, int in_h, int in_w, int in_c,
                  int classes, uint64_t seed, nc_net** out);
int nc_net_destroy(nc_net* net);
int64_t nc_net_param_count(nc_net* net);
/* number of parameter tensors and, for tensor i, its element offset / size in the flat buffer */
int nc_net_tensor_count(nc_net* net);
int nc_net_tensor_info(nc_net* net, int i, int64_t* offset, int64_t* size, int* ndim, int64_t dims[4]);
/* device pointers to the library-owned flat buffers (float32[param_count]) */
float* nc_net_params(nc_net* net);
float* nc_net_grads(nc_net* net);
int nc_net_set_params(nc_net* net, const float* host_params);   /* synchronous H2D */
int nc_net_get_params(nc_net* net, float* host_params);         /* synchronous D2H */
/* math: 0 = fp32 SIMT everywhere; 1 = tcgen05 tensor-core tiles in 3xTF32 (fp32-class accuracy);
 * 2 = tcgen05 tiles with one round-to-nearest TF32 product (logits within ~2e-3 relative);
 * 3 = 3xTF32 for the forward pass (logits / loss as in mode 1), one rounded TF32 product for the
 *     two gradient contractions (gradients ~1e-3 relative, post-step weights far inside 1e-3). */
int nc_net_set_math(nc_net* net, int math);

/* logits[n][classes] = net(x[n][in_h*in_w*in_c])   -- sess.run(y_conv / argmax)
 * (construct_distribute.py:408, construct_inference.py:340-342) */
int nc_net_forward(nc_net* net, const float* x, int n, float* logits, void* stream);
/* labels[n] = argmax(logits, 1)  (construct_inference.py:335) */
int nc_net_predict(nc_net* net, const float* x, int n, int32_t* labels, void* stream);
/* forward + loss (loss_function :285-298; loss_name 0 = entropy, 1 = mse) + backward into the
 * flat gradient buffer; loss_out (device float, nullable) receives the batch-mean loss and
 * correct_out (device int32, nullable) the number of argmax hits (accuracy :381-382).
 * grad_scale multiplies dLoss (1/world for data parallel ranks holding n of world*n rows). */
int nc_net_backward(nc_net* net, const float* x, const float* y_onehot, int n, int loss_name,
                    float grad_scale, float* loss_out, int32_t* correct_out, void* stream);
/* optimizer update on the flat buffers (ApplyAdagrad / ApplyGradientDescent / ApplyAdam):
 * opt 0 = Adagrad (what runs: AdagradOptimizer(1e-4), construct_distribute.py:372-373;
 * accumulator starts at 0.1, no epsilon), 1 = SGD (optimizer_function :309), 2 = Adam (TF1
 * defaults b1=.9 b2=.999 eps=1e-8; offered by options.py:28).  Increments global_step. */
int nc_net_apply(nc_net* net, int opt, float lr, void* stream);
/* nc_net_backward + nc_net_apply: one sess.run(train_op) (construct_distribute.py:413) */
int nc_net_train_step(nc_net* net, const float* x, const float* y_onehot, int n, int loss_name,
                      int opt, float lr, float* loss_out, int32_t* correct_out, void* stream);
int64_t nc_net_global_step(nc_net* net);
/* global_step += k without running anything.  For hosts that capture nc_net_train_step into a CUDA
 * graph (every call of this library on a non-NULL stream is capturable once a first, un-captured call
 * with the same batch size has sized the workspaces): replaying the graph repeats the device work
 * of sess.run(train_op) (construct_distribute.py:413) but not the host-side counter the reference
 * keeps in tf.train.get_or_create_global_step() (:371).  Adam must not be captured: its step size
 * depends on global_step and is computed on the host. */
int nc_net_advance_step(nc_net* net, int64_t k);
/* Per-stage timing of a step, for benchmarks: after nc_net_profile(net, 1) every forward / backward call on this net
 * records a CUDA event on its stream in front of each layer's launches (and of the loss head and the optimizer);
 * nc_net_profile_read closes the last stage, synchronises `stream`, returns the stage durations in milliseconds (ms[0..max))
 * and their names (newline separated) and clears the record.  The duration of the reference's sess.run(train_op)
 * (construct_distribute.py:412-414) is the sum; the stages are this library's own kernels. */
int nc_net_profile(nc_net* net, int on);
int nc_net_profile_read(nc_net* net, void* stream, float* ms, int max, char* names, size_t cap, int* count);
/* The library-owned activation arena, split-K workspace and im2col buffer grow on demand (a forward,
 * predict or backward with a larger batch than any before frees and re-allocates them).  A CUDA graph
 * captured earlier holds the OLD device addresses: replaying it would read and write freed memory.
 * nc_net_arena_generation returns a counter that changes whenever any of those buffers moved; a host
 * that replays captured steps records it at capture time and re-captures when it differs.
 * nc_net_reserve sizes every batch-proportional buffer for max_batch rows up front (the split-K and
 * im2col workspaces are sized by the first call that needs them), so later calls up to that
 * batch move nothing. */
int64_t nc_net_arena_generation(nc_net* net);
int nc_net_reserve(nc_net* net, int max_batch);

/* ---- data parallel inside the library (NCCL bound with dlopen, one process per GPU) ----------
 * Replaces the 1 ps + 1 worker gRPC exchange of construct_distribute.py:344-357 (cmd.py:63-69) by
 * a synchronous all-reduce; parity target = the single-device step at the GLOBAL batch.
 * nc_nccl_unique_id: rank 0 creates the 128-byte id and hands it to the other ranks through
 *   whatever channel the host has (torch.distributed broadcast, a file, the Django job row).
 *   libnccl_path NULL / "" = "libnccl.so.2" on the loader path.
 * nc_net_attach_nccl: collective over all ranks (ncclCommInitRank).  Afterwards, on this net:
 *   - nc_net_backward / nc_net_train_step compute normalize() (:155-165) from the moments of the
 *     global batch (per-channel partial sums all-reduced: two small exchanges forward, one
 *     backward), assuming every rank holds the same number of rows; nc_net_forward / _predict
 *     keep using the statistics of the batch they are given (they are per-rank replica calls);
 *   - nc_net_train_step scales dLoss by 1/world, sums the flat gradient buffer over the ranks and
 *     applies the update, so every rank ends the step with the weights of the global-batch step;
 *     the tail of the buffer (the dense layers: most of the bytes) is exchanged on a side stream
 *     while the convolution gradients are still being computed (NC_DP_NO_OVERLAP in the
 *     environment at attach time turns that off);
 *   - nc_net_allreduce_grads does that sum alone (between nc_net_backward(grad_scale = 1/world)
 *     and nc_net_apply).
 *   All ranks must make the same calls in the same order (they contain collectives).
 * nc_net_detach_nccl: synchronises the device, destroys the communicator, back to world = 1. */
int nc_nccl_unique_id(nc_ctx* ctx, const char* libnccl_path, unsigned char out_id[128]);
int nc_net_attach_nccl(nc_net* net, const char* libnccl_path, const unsigned char id[128], int rank, int world);
int nc_net_allreduce_grads(nc_net* net, void* stream);
/* Same as nc_net_attach_nccl for a host that already owns a communicator: nccl_comm is its ncclComm_t (one per
 * process, spanning `world` ranks); the library uses it on its own side stream and never destroys it.  While the net is
 * attached the host must not issue collectives on that communicator itself. */
int nc_net_attach_comm(nc_net* net, const char* libnccl_path, void* nccl_comm, int rank, int world);
int nc_net_detach_nccl(nc_net* net);
/* Gradient all-reduce over NVLink peer memory instead of ncclAllReduce (one node, 2..8 ranks, one process per GPU).  At the
 * bucket sizes of this path NCCL is latency-bound (8 B200: 30 us for the 0.2 MB convolution bucket, 76-90 us for the 12.9 MB
 * dense bucket); the library's own two-shot kernel reads and writes the peers' gradient buffers directly, sums in rank order
 * (bit-identical on every rank) and can be captured in a CUDA graph.
 *   nc_net_p2p_handles  after nc_net_attach_nccl: 128 bytes = the CUDA IPC handles of this rank's gradient buffer and flag words
 *   nc_net_p2p_attach   all_handles = world x 128 bytes in rank order (gathered by the host through any channel); maps the
 *                       peers (NC_EUNSUPPORTED if two GPUs have no peer access: the net then keeps using NCCL)
 *   nc_net_p2p_status   *timed_out = 1 if a wait inside the kernel gave up (~20 s: a peer never launched its step)
 * Every rank must call nc_net_train_step / nc_net_allreduce_grads in the same order, as with NCCL. */
int nc_net_p2p_handles(nc_net* net, unsigned char out[128]);
int nc_net_p2p_attach(nc_net* net, const unsigned char* all_handles, int rank, int world);
int nc_net_p2p_status(nc_net* net, int* timed_out);
/* optimizer state (Adagrad accumulator / Adam m,v) for checkpoints: float32[slots*param_count] */
int nc_net_state_slots(nc_net* net, int opt);
int nc_net_get_state(nc_net* net, int opt, float* host_state);
int nc_net_set_state(nc_net* net, int opt, const float* host_state, int64_t global_step);

/* tf.matmul(a, b) (+ bias) as used by connect_layer() and the output layer
 * (construct_distribute.py:182, :264) on the TF32 tensor-core path (tcgen05.mma, fp32 accumulate):
 * out[m][n] = op(a) . op(b) + bias[n];  a is [m][k] ([k][m] if trans_a), b is [k][n] ([n][k] if
 * trans_b), all row-major float32 device buffers, 16-byte aligned with row pitches that are
 * multiples of 4 floats (NC_EUNSUPPORTED otherwise).  bias may be NULL.
 * precision 0 = 3xTF32 (operands split into hi + lo TF32 parts, three MMAs per k step: fp32-class
 * accuracy, ~1e-6 relative), 1 = one round-to-nearest TF32 product (~1e-3 relative). */
int nc_matmul(nc_ctx* ctx, const float* a, const float* b, const float* bias, int m, int k, int n,
              int trans_a, int trans_b, int precision, float* out, void* stream);

/* ---- dense layer without any allocation -------------------------------------------------
 * No counterpart in the reference: this entry point exists for hosts of the library that must not
 * allocate device memory while serving -- a process whose own pool already owns all of HBM, a
 * latency-critical request thread (cudaMalloc locks and may synchronise the whole device), or a
 * tenant whose memory budget is audited.  It computes y[m][n] = x[m][k] . w[k][n] (+ bias[n]) --
 * the arithmetic of one `connect` / output layer -- with the same deterministic split-K scheme as
 * nc_net_forward, but keeps the partial sums in a fixed 4 MiB array inside the library image
 * instead of a cudaMalloc'ed workspace.  It calls no cudaMalloc / cudaFree / allocating routine.
 *   limit:   m * n <= 1 048 576 outputs (NC_EINVAL beyond that; split the batch or use nc_net_forward)
 *   cost:    every GPU that loads the library holds those 4 MiB until the process exits
 *   sharing: all calls on a device use the same array; it keeps the previous call's partial sums
 *            (never read: every partial is written before it is summed, nothing needs zeroing).
 *            Calls whose kernels overlap on the GPU (other streams / threads) race: serialise them.
 *   set-up:  the first call on a device looks up the array's device address (no allocation);
 *            make that first call before any stream capture on that device.
 * x, w, bias (nullable), y are device pointers, row-major float32. */
int nc_dense_forward_noalloc(nc_ctx* ctx, const float* x, const float* w, const float* bias,
                             int m, int k, int n, float* y, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NJUCLOUD_B200_H */
