/*
 * serann_b200.h - C ABI of the B200-native SeRANN population-evaluation engine.
 *
 * One engine handle = one GPU = one shard of a generation's population.  The
 * reference has no FFI for this path: its worker is pure Python on TF 2.6
 * (reference evolutionary_experiment/logic/experiment_worker.py).  Each entry
 * point below cites the reference code whose work it replaces; the Python side
 * that binds them with ctypes is self-replicating-artificial-neural-networks_b200/engine.py
 * and the stub a reference maintainer would add is in INTEGRATION.md.
 *
 * Conventions: every function returns SRN_OK (0) or a negative error code and
 * never throws; srn_last_error() gives the message.  All pointers are HOST
 * pointers unless a parameter is named d_* (device).  The caller owns every
 * buffer; sizes are explicit.  A handle is not thread-safe; calls block until
 * their results are on the host.  fp16 values cross the ABI as uint16_t bit
 * patterns (numpy float16).
 */
#ifndef SERANN_B200_H
#define SERANN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SRN_OK              0
#define SRN_ERR_ARG        -1   /* bad argument */
#define SRN_ERR_CUDA       -2   /* CUDA runtime error (message has the detail) */
#define SRN_ERR_STATE      -3   /* call order (e.g. fit before build) */
#define SRN_ERR_UNSUPPORTED -4  /* layer graph the engine cannot run */
#define SRN_ERR_LIMIT      -5   /* size beyond a documented maximum */

/* layer kinds: must match frontend.py */
#define SRN_KIND_DENSE   0   /* Dense on the last axis = 1x1 convolution on (rows, 1, C) */
#define SRN_KIND_CONV    1   /* Conv2D, or Conv1D given as a 1 x k convolution; 'valid' padding */
#define SRN_KIND_MAXPOOL 2
#define SRN_KIND_BN      3   /* reference common/BatchNormalizationF16.py */
#define SRN_KIND_CONCAT  4   /* last axis */
#define SRN_KIND_RESHAPE 5   /* view */

#define SRN_ACT_NONE 0
#define SRN_ACT_RELU 1
#define SRN_ACT_SIGMOID 2

#define SRN_MAX_CONCAT_INPUTS 4

typedef struct srn_engine srn_engine;
typedef struct srn_decoder srn_decoder;

typedef struct {
    int32_t device;          /* CUDA device ordinal */
    int32_t image_h, image_w;/* classification_image_dimensions (example.json:8) */
    int32_t genotype_size;   /* example.json:5 */
    int32_t n_classes;       /* example.json:9 */
    int32_t max_batch;       /* largest number of rows one forward pass will carry (>= training_batch_size) */
} srn_config;

/* One layer of one individual.  Tensor ids: 0 = image input, 1 = genotype input,
 * 2 + i = output of the individual's layer i.  Shapes are per sample, channel last,
 * flattened to (H, W, C): Dense/BN/concat/reshape use H = product of the leading axes, W = 1. */
typedef struct {
    int32_t kind;
    int32_t act;
    int32_t n_in;
    int32_t in[SRN_MAX_CONCAT_INPUTS];
    int32_t units;           /* Dense units / conv filters */
    int32_t kh, kw, sh, sw;  /* conv kernel+stride, pool window+stride */
    int32_t in_h, in_w, in_c;
    int32_t out_h, out_w, out_c;
} srn_layer_desc;

const char* srn_last_error(void);
/* Number of CUDA kernels this library has launched in the calling process (bench.py's gpu_launches). */
uint64_t srn_launch_count(void);

/* Number of cudaMalloc calls this library has made in the calling process.  Device memory of a handle is grow-only:
 * a steady-state generation (same population size, similar architectures) allocates nothing, which this counter lets
 * a caller verify (bench.py reports the allocations per generation). */
uint64_t srn_alloc_count(void);

/* Replaces EvolutionaryExperimentWorker.__init__ (experiment_worker.py:24-43). */
int srn_create(const srn_config* cfg, srn_engine** out);
void srn_destroy(srn_engine* h);

/* Replaces _build_dataset (experiment_worker.py:197-220) / get_serann_training_data (common/logic.py:38-54).
 * x_*: float32 [n][image_h*image_w] already divided by 255; y_*: class index; g_*: 0/1 per locus
 * [n][genotype_size].  Copied to the device once and kept (fp16 images). */
int srn_set_dataset(srn_engine* h,
                    const float* x_train, const uint8_t* y_train, const uint8_t* g_train, int32_t n_train,
                    const float* x_test, const uint8_t* y_test, const uint8_t* g_test, int32_t n_test);

/* Replaces _build_keras_models + Model(...) + compile (experiment_worker.py:63-80,222-246) for the
 * n trainable individuals of a shard.  layer_offsets has n + 1 entries into layers[].  Allocates the
 * packed parameter / Adam / gradient arenas and the activation arena; parameters get Keras' default
 * initialisation (glorot_uniform kernels, zero biases, BN gamma 1, beta 0, mean 0, variance 1) from
 * `seed`.  A second build discards the first. */
int srn_build(srn_engine* h, const srn_layer_desc* layers, const int32_t* layer_offsets,
              const float* loss_balance, int32_t n, uint64_t seed);
/* The same with one initialisation seed per individual, so that an individual's initial weights do not depend on
 * which shard (GPU) it lands in or on its position there: a generation sharded over 1, 2, 4 or 8 GPUs starts every
 * individual from the same weights (the reference draws them from TF's unseeded global generator,
 * experiment_worker.py:63-70, so any assignment is equally faithful). */
int srn_build_seeded(srn_engine* h, const srn_layer_desc* layers, const int32_t* layer_offsets,
                     const float* loss_balance, int32_t n, const uint64_t* individual_seeds);

/* Weights in Keras layouts (Dense (in, out); Conv2D HWIO; Conv1D (k, in, out); BN gamma, beta,
 * moving_mean, moving_variance), fp16.  `which` indexes layer.get_weights().  *count receives the
 * element count; buf may be NULL to query it. */
int srn_get_weights(srn_engine* h, int32_t individual, int32_t layer, int32_t which, uint16_t* buf, int64_t* count);
int srn_set_weights(srn_engine* h, int32_t individual, int32_t layer, int32_t which, const uint16_t* buf, int64_t count);
int64_t srn_parameters_count(srn_engine* h, int32_t individual);   /* Keras count_params() */
int64_t srn_device_bytes(srn_engine* h);                          /* arena bytes held by the build */

/* Replaces joined_model.fit (experiment_worker.py:101-128): `epochs` passes over the first
 * n_train - n_val rows in mini-batches of `batch` (last batch of an epoch may be short), then an
 * inference-mode pass over the last n_val rows after every epoch.  perms: [epochs][n_train - n_val]
 * row order per epoch, or NULL to draw them from shuffle_seed.  max_steps > 0 stops after that many
 * steps (benchmarks).  Outputs (any may be NULL), S = steps per epoch, V = ceil(n_val / batch):
 *   train_correct [epochs][S][n] int32    correctly classified rows of each step
 *   train_loss    [epochs][S][n][2] float mean CE and mean MSE of each step
 *   val_correct   [epochs][V][n] int32,  val_mse_sum [epochs][V][n] float (sum over the chunk of per-row MSE)
 *   device_ms     device time of the training steps (CUDA events), excluding validation */
int srn_fit(srn_engine* h, int32_t epochs, int32_t batch, int32_t n_val, const int32_t* perms,
            uint64_t shuffle_seed, int32_t max_steps,
            int32_t* train_correct, float* train_loss, int32_t* val_correct, float* val_mse_sum,
            float* device_ms);

/* Replaces joined_model.evaluate (experiment_worker.py:87) and the validation pass: inference-mode
 * forward over rows [start, start + rows) of split 0 (train) or 1 (test), metrics tallied per
 * `chunk` rows (Keras' batch of 32 for evaluate): correct [ceil(rows/chunk)][n], mse_sum likewise. */
int srn_evaluate(srn_engine* h, int32_t split, int32_t start, int32_t rows, int32_t chunk,
                 int32_t* correct, float* mse_sum);

/* Replaces _replicate (experiment_worker.py:130-167) plus the threshold and mutation count the
 * driver applies afterwards (experiment.py:210,229).  parent_bits [n][genotype_size] 0/1;
 * image_index [n*R] rows of the test split (NULL = the reference's rule, row (i*R + r) / ceil(n*R / n_test)).
 * probs [n][R][genotype_size] fp16 sigmoid outputs (may be NULL); packed_bits [n][R][ceil(G/32)]
 * uint32, bit j of word j/32 = (p > 0.5); mutations [n][R] = Hamming distance to the parent. */
int srn_replicate(srn_engine* h, const uint8_t* parent_bits, int32_t R, const int32_t* image_index,
                  uint16_t* probs, uint32_t* packed_bits, int32_t* mutations);

/* Device-time accounting for bench.py's roofline: with profiling on, every kernel launch of this engine is
 * bracketed by CUDA events on the engine's stream and tagged with one of the categories below;
 * srn_profile_read sums the elapsed times and counts per category since profiling was switched on.
 * (No reference counterpart: the reference times whole phases with datetime.now(), experiment_worker.py:114-119.) */
#define SRN_CAT_GATHER      0   /* batch gather from the resident dataset */
#define SRN_CAT_GEMM_FWD    1   /* Dense / Conv forward */
#define SRN_CAT_GEMM_DGRAD  2   /* input gradients */
#define SRN_CAT_GEMM_WGRAD  3   /* weight gradients */
#define SRN_CAT_ACT_BWD     4   /* activation backward + bias gradient */
#define SRN_CAT_BN          5
#define SRN_CAT_POOL        6
#define SRN_CAT_CONCAT      7
#define SRN_CAT_LOSS        8
#define SRN_CAT_ADAM        9
#define SRN_CAT_READOUT    10
#define SRN_CAT_OTHER      11
#define SRN_CAT_COUNT      12
int srn_profile(srn_engine* h, int32_t on);
int srn_profile_read(srn_engine* h, float* ms /* [SRN_CAT_COUNT] */, int32_t* launches /* [SRN_CAT_COUNT] */);

/* Debug/parity: inference- or training-mode forward of rows [start, start+rows) of a split,
 * returning the fused head logits [n][rows][n_classes + genotype_size] as fp16. */
int srn_forward_logits(srn_engine* h, int32_t split, int32_t start, int32_t rows, int32_t training,
                       uint16_t* logits);

/* Ribosomal autoencoder decode: replaces RiboAeProcess.decode (ribosomal_autoencoder.py:119-126) over the
 * decoder of ribosomal_autoencoder/model.py:35-48.  Weights fp32 in Keras layouts: conv kernel (5, 2, 32),
 * conv bias 32, BN1 gamma/beta/mean/var 32 each, dense kernel (96*32, max_tokens*vocab), dense bias,
 * BN2 gamma/beta/mean/var `vocab` each.  Creation folds the binary-input convolution and the first BN into the
 * Dense kernel (a table over the five-bit windows, accumulated in double) and splits it into two fp16 halves; decode
 * is then one tcgen05 GEMM with the second BN and the per-token argmax in its epilogue, plus an exact (double)
 * recomputation of the tokens whose top-two margin is too small for the split to be trusted.  Shapes the tensor-core
 * kernel does not serve (vocab not a multiple of 8, genotype_size > 104) run the exact kernel for every token. */
int srn_decoder_create(int32_t device, int32_t genotype_size, int32_t max_tokens, int32_t vocab,
                       const float* conv_kernel, const float* conv_bias, const float* bn1,
                       const float* dense_kernel, const float* dense_bias, const float* bn2,
                       srn_decoder** out);
void srn_decoder_destroy(srn_decoder* d);
/* bits [n][genotype_size] 0/1 -> tokens [n][max_tokens] (argmax over the vocabulary, first maximum wins).
 * device_ms (may be NULL) receives the device time of the decode kernels. */
int srn_decoder_decode(srn_decoder* d, const uint8_t* bits, int32_t n, int32_t* tokens, float* device_ms);
/* Of the last srn_decoder_decode call: tokens that went through the exact recomputation, and whether the tensor-core
 * kernel served the call (no reference counterpart: bench.py and the tests report them). */
int srn_decoder_stats(srn_decoder* d, int64_t* rechecked, int32_t* tensor_path);

/* Stand-alone readout (experiment.py:210,229; serann_evaluation/logic.py:199-203,223):
 * probs [rows][G] fp16 -> packed bits [rows][ceil(G/32)], mutations [rows] vs parent_bits [rows/rows_per_parent][G]. */
int srn_readout(int32_t device, const uint16_t* probs, const uint8_t* parent_bits, int32_t rows,
                int32_t rows_per_parent, int32_t G, uint32_t* packed_bits, int32_t* mutations);

/* Levenshtein distance of two byte strings (host code; replaces edlib.align(a, b)['editDistance'],
 * evolutionary_experiment/logic/experiment.py:16-17).  Returns -1 on a bad argument. */
int32_t srn_levenshtein(const char* a, int32_t la, const char* b, int32_t lb);

/* Stand-alone BatchNormalizationF16 training-mode forward on a channel-last fp16 tensor [rows][C]
 * (common/BatchNormalizationF16.py:130-153).  params = gamma, beta, moving_mean, moving_variance
 * (each [C], fp16); moving statistics are updated in place; y [rows][C]. */
int srn_bn_train_forward(int32_t device, const uint16_t* x, int64_t rows, int32_t C,
                         const uint16_t* gamma, const uint16_t* beta,
                         uint16_t* moving_mean, uint16_t* moving_variance, uint16_t* y);

/* The same BatchNormalizationF16 forward for callers that must not allocate device memory.  It has no
 * counterpart in the reference.  Who needs it: code running inside a framework whose own allocator has
 * already claimed the GPU's memory (a raw cudaMalloc from a plug-in can fail there, or break the
 * framework's accounting), and latency-sensitive loops, because cudaMalloc/cudaFree synchronise the
 * whole device.  This entry point calls no cudaMalloc, cudaFree or allocating library routine: its working
 * storage (per-channel statistics, per-block partial sums, the launch descriptor) is a fixed array of
 * SRN_NOALLOC_SCRATCH_BYTES that is part of the library image - every GPU that loads the library holds
 * those bytes until the process exits.
 *   - All d_* pointers are DEVICE pointers on `device` that the caller allocated; d_moving_variance must be
 *     d_moving_mean + C (one [2][C] fp16 array).  When C is a multiple of 8 the kernels move 16-byte groups of
 *     8 channels, so d_x and d_y must then be 16-byte aligned (a slice of a larger tensor that starts at an
 *     odd element offset is refused with SRN_ERR_ARG instead of faulting inside the caller's CUDA context).
 *     `stream` is a cudaStream_t created on `device` (NULL = that device's default stream); the call returns
 *     after the stream has finished the work and leaves the calling thread's current device as it found it.
 *   - Limit: C * (4 + 2 * ceil(rows / R)) floats must fit in SRN_NOALLOC_SCRATCH_BYTES - 512, where R is the
 *     number of rows one thread block reduces: R = 16 * max(1, 256 / n) with n = C / 8 when C is a multiple
 *     of 8, else n = C (e.g. 750*784 rows x 128 channels fits; 16M rows x 64 channels does not).  Larger
 *     problems are refused with SRN_ERR_LIMIT; use srn_bn_train_forward for those.
 *   - The first call on each device looks up the array's device address (cudaGetSymbolAddress); make that
 *     first call before starting any stream capture.
 *   - Every call uses the same array and leaves its last contents there.  Nothing in it has to be zero on
 *     entry (each value is written before it is read), but two calls whose GPU work overlaps - different
 *     streams, or different host threads - race on it: serialise them. */
#define SRN_NOALLOC_SCRATCH_BYTES (4u << 20)
int srn_bn_train_forward_noalloc(int32_t device, void* stream, const uint16_t* d_x, int64_t rows, int32_t C,
                                 const uint16_t* d_gamma, const uint16_t* d_beta,
                                 uint16_t* d_moving_mean, uint16_t* d_moving_variance, uint16_t* d_y);

#ifdef __cplusplus
}
#endif
#endif /* SERANN_B200_H */
