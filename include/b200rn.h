/* b200rn.h -- C ABI of the B200-native ResNet-50 v2 backend (libb200rn.so).
 *
 * This is the drop-in boundary for the one hot path of
 * AbhiramDodda/distributed-inference-engine-cpp: the batched forward behind `InferenceEngine`
 * (reference: include/inference_engine.h:10-21, src/inference_engine.cpp).  The reference has no
 * C ABI of its own -- it calls ONNX Runtime's C++ API in-process -- so each entry point below names
 * the reference code it replaces.  Plain pointers and sizes only; no exceptions cross this
 * boundary: every function returns 0 on success and a non-zero code on failure, with a message
 * available from b200rn_last_error() (per handle) or b200rn_global_error() (no handle).
 *
 * Threading: a handle serialises its own calls internally (the reference holds mutex_ for the
 * whole of predict/batchPredict, src/inference_engine.cpp:90,136).  Different handles are
 * independent (one per GPU for the replica-per-device deployment).
 */
#ifndef B200RN_H
#define B200RN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200rn b200rn_t;

/* Arithmetic modes.
 *  EXACT: fp32 storage; convolutions as 3 x TF32 split products on the tensor cores with two-level fp32
 *         accumulation (chunks in tensor memory, chunk sums in registers), FC on CUDA cores; logits within 1e-3
 *         max-abs of the fp32 CPU oracle (the north-star's "fp32 accumulate" bar).  B200RN_EXACT_SIMT=1 in the
 *         environment keeps every convolution on the fp32 CUDA-core kernel instead.
 *  BF16 / FP16: 16-bit NHWC activations and weights, tcgen05 tensor cores, fp32 accumulation in
 *         TMEM; judged by top-1 agreement. */
enum { B200RN_MODE_EXACT = 0, B200RN_MODE_BF16 = 1, B200RN_MODE_FP16 = 2 };

enum {
  B200RN_OK = 0,
  B200RN_ERR_ARG = 1,      /* bad argument (null pointer, over-long sample, ...) */
  B200RN_ERR_MODEL = 2,    /* unreadable / malformed / unsupported ONNX file */
  B200RN_ERR_CUDA = 3,     /* CUDA runtime or driver error, or no usable device */
  B200RN_ERR_STATE = 4     /* call sequence error (e.g. wait without submit) */
};

/* Replaces InferenceEngine::InferenceEngine + initializeSession (src/inference_engine.cpp:6-87):
 * parses the ONNX file, folds BatchNorm, re-lays the weights and uploads them ONCE to `device`
 * (the reference hard-codes CUDA device 0, :23; its shard_id is stored and never used, :7).
 * max_batch bounds the activation arena; larger batches are processed in slices of max_batch.
 * max_batch <= 0 selects 256. */
int b200rn_create(const char* onnx_path, int device, int mode, int max_batch, b200rn_t** out);

/* Replaces InferenceEngine::~InferenceEngine (src/inference_engine.cpp:12-14). NULL is a no-op. */
void b200rn_destroy(b200rn_t* h);

/* Replaces getInputShape / getOutputShape (src/inference_engine.cpp:211-217).  Dynamic dims are
 * reported as 1, as the reference does (:47-51,64-68).  Writes up to `cap` dims; returns the rank
 * (or -1 on error). */
int b200rn_input_shape(const b200rn_t* h, int64_t* dims, int cap);
int b200rn_output_shape(const b200rn_t* h, int64_t* dims, int cap);

/* Elements per sample on the way in (product of input dims after dim 0) and out. */
int64_t b200rn_input_elems(const b200rn_t* h);
int64_t b200rn_output_elems(const b200rn_t* h);

/* Replaces InferenceEngine::batchPredict (src/inference_engine.cpp:134-209) for the caller's ragged
 * batch: samples[i] points at sample_len[i] host floats.  Short samples are zero-padded (:157-159).
 * A sample LONGER than the model input is rejected with B200RN_ERR_ARG -- the reference appends it
 * whole (:155) and mis-sizes the tensor; see DESIGN.md "documented divergence".  batch == 0 is OK and
 * writes nothing (:137-139).  out receives batch x b200rn_output_elems() floats, caller-owned. */
int b200rn_infer(b200rn_t* h, const float* const* samples, const size_t* sample_len, size_t batch, float* out);

/* Same forward for an already packed host batch [batch, C, H, W] (what batchPredict builds at
 * src/inference_engine.cpp:152-160).  If `nchw` / `out` lie in pinned memory (b200rn_host_alloc, or
 * cudaHostRegister'ed by the caller) the transfers are true async DMA with no staging copy. */
int b200rn_infer_packed(b200rn_t* h, const float* nchw, size_t batch, float* out);

/* Same forward with input and output already resident in device memory on the handle's device
 * (kernel-only path; no host<->device traffic). */
int b200rn_infer_device(b200rn_t* h, const float* d_nchw, size_t batch, float* d_out);

/* Double-buffered asynchronous form of b200rn_infer_packed: two slots (0, 1).  submit() queues the
 * host->device copy on a side stream, then the forward, then the device->host copy, and returns at
 * once; wait() blocks until that slot's results are in `out`.  While slot 0 computes, slot 1's input
 * is already crossing PCIe.  batch must be <= max_batch.  Buffers must stay valid until wait(). */
int b200rn_submit_packed(b200rn_t* h, int slot, const float* nchw, size_t batch, float* out);
int b200rn_wait(b200rn_t* h, int slot);

/* Replaces InferenceEngine::predict (src/inference_engine.cpp:89-132): one sample, padded AND
 * truncated to the model input size (:100-103).  out receives b200rn_output_elems() floats. */
int b200rn_predict(b200rn_t* h, const float* sample, size_t len, float* out);

/* Pinned host memory for zero-staging transfers. */
void* b200rn_host_alloc(size_t bytes);
void b200rn_host_free(void* p);

const char* b200rn_last_error(const b200rn_t* h);
const char* b200rn_global_error(void);

/* ---- introspection / measurement (no reference counterpart; used by bench.py and tests) ---- */

/* Human-readable fused layer plan built from an ONNX file, no GPU needed.  Returns the number of
 * bytes the full text needs (excluding NUL); writes at most cap-1 bytes + NUL. */
int64_t b200rn_describe_plan(const char* onnx_path, int mode, char* buf, size_t cap);

/* Number of device launches one forward of `batch` samples performs (gpu_launches accounting). */
int b200rn_launch_count(const b200rn_t* h);
int b200rn_layer_count(const b200rn_t* h);

/* Runs `iters` forwards of `batch` (input taken from d_nchw, device memory) with a CUDA event
 * between every launch and writes the mean milliseconds per layer into ms[0..layer_count).
 * kinds[i] receives the layer kind (0 InputXform, 1 Conv, 2 MaxPool, 3 Affine, 4 Add, 5 Pool,
 * 6 Gemm, 7 Softmax), flops[i] the algorithmic FLOPs of layer i for this batch (2*MACs; 0 for
 * non-GEMM layers) and bytes[i] its algorithmic bytes (inputs + weights + outputs once). */
int b200rn_profile_layers(b200rn_t* h, const float* d_nchw, size_t batch, int iters, float* ms, int* kinds,
                          double* flops, double* bytes);
/* Copies layer i's name (fused ONNX node names) into buf. */
int b200rn_layer_name(const b200rn_t* h, int i, char* buf, size_t cap);

/* Times `iters` back-to-back forwards of `batch` on the device (CUDA events on the compute stream,
 * inputs resident in HBM, taken round-robin from the n_inputs device batches so that consecutive
 * steps do not re-read a cached input); returns total milliseconds in *ms_total. */
int b200rn_time_forward(b200rn_t* h, const float* const* d_inputs, int n_inputs, size_t batch, int iters,
                        float* ms_total);

/* ---- single-operator entry point (tests of the conv kernels against a framework reference) ----
 * NHWC convolution with the fused epilogue, device pointers, on the current device.
 * dtype: 0 fp32 (exact mode: 3 x TF32 on tcgen05 where Cin % 32 == 0 and Cout % 64 == 0, else -- or with
 * B200RN_EXACT_SIMT=1 -- the CUDA-core kernel), 1 bf16, 2 fp16 (tcgen05 path).  w: fp32 host weights
 * [Cout][kh][kw][Cin]; bias/act_scale/act_shift: fp32 host arrays [Cout] (act_* may be NULL).
 * residual / out / out_act are device NHWC tensors of dtype (any of them may be NULL except that
 * one output must exist).  pre_scale / pre_shift (fp32 host arrays [Cin], may be NULL; 1x1 pad-0 convs
 * only) make the conv read relu?(pre_scale[c]*x + pre_shift[c]) instead of x. */
int b200rn_conv2d(int dtype, const void* d_in, int B, int H, int W, int Cin, const float* w, const float* bias,
                  int Cout, int kh, int kw, int stride, int pad, int relu, const void* d_residual, void* d_out,
                  void* d_out_act, const float* act_scale, const float* act_shift, int act_relu,
                  const float* pre_scale, const float* pre_shift, int pre_relu);
/* The fused pair "1x1 conv (Cin -> Cout = 256) + residual, then the next block's 1x1 conv (256 -> tail_cout)"
 * as one launch (tensor-core dtypes 1 / 2 only; the graph planner's tail fusion, csrc/plan.h).  Replaces two
 * ORT Conv nodes + Add + BatchNormalization + Relu of the reference's session (src/inference_engine.cpp:134).
 *   d_out      = conv1x1(d_in, w, bias) (ReLU'd if relu) + d_residual                       [B,H,W,256]
 *   d_tail_out = relu?( conv1x1( relu?(act_scale*d_out + act_shift), tail_w ) + tail_bias )   [B,H,W,tail_cout]
 * w: [256][Cin], tail_w: [tail_cout][256] fp32 host arrays; tail_cout is 64 or 128; Cin is 64. */
int b200rn_conv1x1_tail(int dtype, const void* d_in, int B, int H, int W, int Cin, const float* w, const float* bias, int relu,
                        const void* d_residual, void* d_out, const float* act_scale, const float* act_shift, int act_relu,
                        const float* tail_w, const float* tail_bias, int tail_cout, int tail_relu, void* d_tail_out);
/* Two-input 1x1 conv (tensor-core dtypes 1 / 2 only; the planner's fusion of a projection block's "main conv +
 * shortcut conv -> Add", csrc/plan.h): replaces two ORT Conv nodes and the Add between them.
 *   d_out     = conv1x1(d_in, w[:, :Cin]) + conv1x1_stride_b(d_in_b, w[:, Cin:]) + bias      [B,H,W,Cout]
 *   d_out_act = relu?(act_scale * d_out + act_shift)                                          (optional)
 * d_in is [B,H,W,Cin], d_in_b is [B,Hb,Wb,Cin_b] read at (y*stride_b, x*stride_b) with (Hb-1)/stride_b+1 == H (same
 * for W); w: fp32 host [Cout][Cin + Cin_b]; Cin, Cin_b multiples of 64, Cout a multiple of 256. */
int b200rn_conv1x1_two_input(int dtype, const void* d_in, int B, int H, int W, int Cin, const void* d_in_b, int Hb, int Wb,
                             int Cin_b, int stride_b, const float* w, const float* bias, int Cout, void* d_out,
                             void* d_out_act, const float* act_scale, const float* act_shift, int act_relu);
/* Same, for the 7x7/2 stem through the tensor-core path: d_in_nchw is fp32 NCHW [B,3,H,W]; the
 * input re-layout (with per-channel scale/shift) and the conv run back to back.  dtype 0 (fp32 output, even H) is the
 * exact-mode form: zero-framed fp32 input, 3 x TF32 split products with two-level accumulation. */
int b200rn_stem_conv(int dtype, const float* d_in_nchw, int B, int H, int W, const float* in_scale,
                     const float* in_shift, const float* w, const float* bias, int Cout, int relu, void* d_out);

/* No counterpart in the reference.  The same stem convolution as b200rn_stem_conv for callers that
 * cannot allocate device memory when they call in (a worker whose device pool was carved up at start-up,
 * a request thread that cannot afford cudaMalloc's device-wide synchronisation, a plugin host that
 * forbids plugins to own device memory).  All scratch -- the zero-framed re-laid input, the packed
 * filters, bias and input scale/shift -- lives in ONE fixed array inside the library
 * (B200RN_STATIC_SCRATCH_BYTES of device memory on every GPU that loads libb200rn.so, held until the
 * process exits), so the call performs no cudaMalloc/cudaFree and no heap allocation.
 * Limits: B * (H+6) * round_up(W+8, 8) * 8 bytes must fit in B200RN_STATIC_FRAME_BYTES (nine 224x224
 * images), Cout must be 64 or 128, H even; anything larger is refused with a message (use
 * b200rn_stem_conv).  The first call on each device looks up the array's device address; make that first
 * call before any stream capture begins.  Every call shares the array and starts by clearing the part it
 * uses; calls whose GPU work overlaps would race on it (calls through this entry point are serialised and
 * synchronous, so that can only happen if the caller overlaps it with a replayed capture of itself).
 * err (optional, err_cap bytes) receives the failure text.  Returns 0 on success. */
#define B200RN_STATIC_FRAME_BYTES (4u << 20)
#define B200RN_STATIC_SCRATCH_BYTES ((4u << 20) + 57344u + 1024u)
int b200rn_stem_conv_static(int dtype, const float* d_in_nchw, int B, int H, int W, const float* in_scale,
                            const float* in_shift, const float* w, const float* bias, int Cout, int relu,
                            void* d_out, char* err, size_t err_cap);

/* ---- request routing and cache keys (bit-exact restatements used by the serving shell) ----
 * Replaces ConsistentHash (include/consistent_hash.h, src/consistent_hash.cpp:6-45) and
 * VectorHash (include/lru_cache.h:84-96). */
uint32_t b200rn_fnv1a(const char* key, size_t len);
typedef struct b200rn_ring b200rn_ring_t;
b200rn_ring_t* b200rn_ring_create(int virtual_nodes);
void b200rn_ring_destroy(b200rn_ring_t* r);
void b200rn_ring_add(b200rn_ring_t* r, const char* node);
void b200rn_ring_remove(b200rn_ring_t* r, const char* node);
/* Writes the owning node's name (NUL-terminated, "" for an empty ring); returns its length. */
int b200rn_ring_get(const b200rn_ring_t* r, const char* key, size_t len, char* out, size_t cap);
/* Distinct nodes in ring order, '\n'-separated. */
int b200rn_ring_nodes(const b200rn_ring_t* r, char* out, size_t cap);
uint64_t b200rn_hash_float(float f);
uint64_t b200rn_vector_hash(const float* v, size_t n);

/* LRU cache keyed by the raw request vector (include/lru_cache.h:14-82). */
typedef struct b200rn_lru b200rn_lru_t;
b200rn_lru_t* b200rn_lru_create(size_t capacity);
void b200rn_lru_destroy(b200rn_lru_t* c);
/* Returns the value length on a hit (copying up to cap floats into out), -1 on a miss. */
long b200rn_lru_get(b200rn_lru_t* c, const float* key, size_t n, float* out, size_t cap);
void b200rn_lru_put(b200rn_lru_t* c, const float* key, size_t n, const float* val, size_t m);
size_t b200rn_lru_size(const b200rn_lru_t* c);
size_t b200rn_lru_hits(const b200rn_lru_t* c);
size_t b200rn_lru_misses(const b200rn_lru_t* c);

/* BatchProcessor self-test harness (include/batch_processor.h): `submitters` threads each submit
 * `per_thread` integers x to a BatchProcessor<int,int> whose callback sleeps callback_ms and returns
 * x+1.  out[0..4] = total_requests, total_batches, timeout_batches, full_batches, wrong answers. */
void b200rn_batch_selftest(int max_batch, int timeout_ms, int callback_ms, int submitters, int per_thread,
                           long long* out, double* avg_batch);
/* The same with the two optional policies of this repository's BatchProcessor (host/batch_processor.h; neither
 * exists in the reference): fill_wait = wait up to timeout_ms for a batch to fill (first half),
 * in_flight = consumer threads (second half: two batches in flight).  out[5] = most callbacks seen running at once. */
void b200rn_batch_selftest2(int max_batch, int timeout_ms, int callback_ms, int submitters, int per_thread, int fill_wait,
                            int in_flight, long long* out, double* avg_batch);

#ifdef __cplusplus
}
#endif
#endif /* B200RN_H */
