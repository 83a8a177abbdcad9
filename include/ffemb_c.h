/* ffemb_c.h -- C-ABI of the B200-native DLRM embedding-bag path.
 *
 * Drop-in boundary for ONE hot path of TravisDai/DLRM-FlexFlow: the embedding
 * operator (forward gather-and-pool, backward, SGD row update, table-wise
 * placement).  Written in the idiom of the reference's python/flexflow_c.h:
 * extern "C", opaque `struct { void *impl; }` handles, plain pointers and
 * sizes, no C++ and no torch types, parseable by `gcc -E | cffi.cdef`.
 *
 * Two layers:
 *
 *   (K) kernel-level entry points -- stateless, device pointers in, device
 *       pointers out, explicit stream.  These are what the reference's Legion
 *       task bodies would call in place of their own kernel launches
 *       (Embedding::forward_task / backward_task, SGDOptimizer::update_task,
 *       UniformInitializer::init_task).  INTEGRATION.md shows the patch.
 *
 *   (M) model-level entry points -- the flexflow_c.h surface for this path
 *       (config / model / tensor / op / parameter / optimizer / initializer
 *       handles; add_embedding, compile, forward, zero_gradients, backward,
 *       update, get/set_weights), each citing the flexflow_c.h declaration it
 *       stands in for.  Names carry the prefix ffemb_ instead of flexflow_ so
 *       both libraries can live in one process.
 *
 * Conventions (differences from the reference are deliberate):
 *   - every call that can fail returns an int status (FFEMB_OK = 0); the
 *     message is available from ffemb_last_error().  Nothing aborts.
 *   - a model handle is not thread-safe; one host thread per GPU.
 *   - no hidden device synchronisation: calls enqueue on the given stream (or
 *     the model's stream) and return; ffemb_model_sync / the caller syncs.
 *   - handles are owned by the caller (destroy them); data buffers passed in
 *     are borrowed for the call; tables are owned by the library.
 *   - all tables / outputs / gradients fp32, all indices int64.
 *   - tables are row-major [num_entries][out_dim]; index streams are
 *     flattened [bag][position]; bag order is summation order.
 *   - There is no CPU fallback: without a CUDA device every compute call
 *     returns FFEMB_ERR_CUDA.
 */
#ifndef FFEMB_C_H
#define FFEMB_C_H

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#include "ffemb_const.h"

#ifdef __cplusplus
extern "C" {
#endif

#define FFEMB_MAX_PARTS 16 /* sample-blocks (ranks) one matrix may be split over */

/* ------------------------------------------------------------------------
 * Library
 * --------------------------------------------------------------------- */

/* Message of the last failing call on this host thread ("" if none). */
const char *ffemb_last_error(void);

/* "major.minor.patch"; sm architecture the kernels were compiled for. */
const char *ffemb_version(void);
int ffemb_compiled_sm(void);

/* Number of this library's own kernels launched so far by this process
 * (all streams).  Used by bench.py for its gpu_launches figure. */
int64_t ffemb_kernel_launch_count(void);

/* ------------------------------------------------------------------------
 * (K) Kernel-level entry points
 * --------------------------------------------------------------------- */

/* One table's share of a batched launch.
 * Stands in for the three regions of Embedding::forward_task
 * (src/ops/embedding.cu:226-263: input, output, kernel). */
typedef struct ffemb_lookup_t {
  float *weight;          /* device, [num_entries][out_dim], 16-byte aligned  */
  const int64_t *indices; /* device, flattened [bag][position]               */
  const int64_t *offsets; /* device, batch+1 CSR bag offsets; NULL = fixed   */
  int64_t num_entries;    /* rows of the table                               */
  int64_t num_indices;    /* indices in this batch: exactly batch*bag_size
                             for fixed bags (checked, both directions);
                             exactly offsets[batch] for CSR bags (device
                             data, NOT checked: positions past offsets[batch]
                             would be sorted and applied uninitialised)      */
  int64_t col;            /* offset (floats) of this table's block from the
                             start of a row of the output / gradient matrix:
                             its first column in a [batch][T*D] concat
                             matrix; any float offset from base[0] when the
                             tables' blocks are separate allocations that
                             share one row stride                            */
  int32_t bag_size;       /* indices per bag when offsets == NULL            */
  int32_t reserved;       /* set to 0                                        */
} ffemb_lookup_t;

/* A sample-major fp32 matrix, optionally split into equal blocks of
 * consecutive samples ("parts"), each with its own base pointer: part r holds
 * samples [r*rows_per_part, (r+1)*rows_per_part).  With num_parts == 1 it is
 * an ordinary [batch][ld] matrix.  With num_parts == world size and base[r]
 * pointing into rank r's memory (peer mapping, ffemb_comm_*) it expresses the
 * reference's table-wise -> sample-wise repartition done at Concat
 * (src/ops/concat.cu:85-93) without a separate exchange step. */
typedef struct ffemb_matrix_t {
  float *base[FFEMB_MAX_PARTS];
  int32_t num_parts;
  int32_t rows_per_part;
  int64_t ld; /* row stride in floats */
} ffemb_matrix_t;

/* Forward gather-and-pool over `num_tables` tables in one batched launch.
 *   out[b, col_t + d] = sum_{j in bag b of table t} W_t[idx, d]   (+0 start,
 *   bag order), times 1.0f/len for AGGR_MODE_AVG (CPU semantics).
 * Replaces embed_forward<<<>>> (src/ops/embedding.cu:173-197, launch :254)
 * and EmbeddingLookup_int64_t_float_float__avx2_fma
 * (src/ops/embedding_avx2.cc:5-296).  `lookups` is a HOST array.
 * An out-of-range index contributes nothing and sets *err_flag (device int,
 * may be NULL) to FFEMB_ERR_INDEX_OUT_OF_RANGE. */
int ffemb_embed_forward(const ffemb_lookup_t *lookups, int num_tables,
                        int batch, int out_dim, enum AggrMode aggr,
                        const ffemb_matrix_t *out, int *err_flag,
                        void *stream);

/* Scratch bytes ffemb_embed_backward_sgd needs for a batch with
 * `total_indices` lookups over tables with `total_entries` rows in all. */
size_t ffemb_embed_backward_workspace_bytes(int64_t total_indices,
                                            int64_t total_entries,
                                            int out_dim);

/* Deterministic backward fused with the sparse SGD row update:
 *   sort (table,row) keys, segment-reduce duplicate rows in original
 *   (bag, position) order, W[row,:] = fmaf(-lr, gsum, W[row,:]) on touched
 *   rows only.
 * Replaces embed_backward<<<>>> (src/ops/embedding.cu:199-224, launch :318),
 * the dense zero of the gradient (src/runtime/initializer_kernel.cu:232-262)
 * and dense sgd_update<<<>>> (src/runtime/optimizer_kernel.cu:23-41,104-106)
 * for weight_decay == 0 and momentum == 0, where the three are equivalent.
 * `grad` is d(loss)/d(out) in the layout of ffemb_embed_forward's `out`.
 * `workspace` is device memory of at least
 * ffemb_embed_backward_workspace_bytes() bytes, 256-byte aligned. */
int ffemb_embed_backward_sgd(const ffemb_lookup_t *lookups, int num_tables,
                             int batch, int out_dim, enum AggrMode aggr,
                             const ffemb_matrix_t *grad, float lr,
                             void *workspace, size_t workspace_bytes,
                             int *err_flag, void *stream);

/* Which key sort the backward uses (process-wide).
 *   2 (default): this library's one-launch on-chip sort (csrc/ffemb_csort.cuh):
 *      one thread-block cluster per table keeps the table's whole segment in
 *      shared memory across its radix passes.  Used when every table of the
 *      launch has at most 65536 lookups and batch <= 65536; otherwise 0 is
 *      used (measured faster than 1 on half-million-lookup segments).
 *   1: this library's segmented sort through global memory
 *      (csrc/ffemb_sort.cuh), every table's segment sorted by row number with
 *      only as many passes as that table's size needs.
 *   0: the CUDA toolkit's DeviceRadixSort over the global (table,row) keys.
 * All are stable, so they produce the same order and bit-identical results.
 * Launch groups whose tables hold 2^32-1 rows or more always use 0. */
int ffemb_set_sort_impl(int impl);
int ffemb_get_sort_impl(void);

/* The two halves of ffemb_embed_backward_sgd, for callers that keep the
 * reference's backward() / update() split: `backward_sort` leaves the sorted
 * sparse gradient description in `workspace`; `sgd_update` consumes it.
 * (Embedding::backward, src/ops/embedding.cu:329-355; SGDOptimizer::update,
 * src/runtime/optimizer.cc:75-102.) */
int ffemb_embed_backward_sort(const ffemb_lookup_t *lookups, int num_tables,
                              int batch, int out_dim, void *workspace,
                              size_t workspace_bytes, int *err_flag,
                              void *stream);
int ffemb_embed_sgd_update(const ffemb_lookup_t *lookups, int num_tables,
                           int batch, int out_dim, enum AggrMode aggr,
                           const ffemb_matrix_t *grad, float lr,
                           void *workspace, size_t workspace_bytes,
                           void *stream);

/* ffemb_embed_backward_sgd for callers that cannot allocate device memory.
 * (No counterpart in the reference, which allocates through Legion.)
 * Some hosts of this library must not call cudaMalloc: a plugin inside a
 * framework that owns the whole memory pool and lends no scratch, or a
 * latency-critical loop that may not take the allocator's device-wide lock.
 * This variant therefore takes no workspace: it uses a fixed array inside the
 * library image -- ffemb_static_scratch_bytes() = 8 MiB of device memory that
 * every GPU which loads the library keeps until the process exits.
 *   - A batch needing more scratch than that (about 250k lookups at out_dim
 *     128) is refused with FFEMB_ERR_WORKSPACE_TOO_SMALL.  The need is the
 *     exact layout size, ffemb_embed_backward_workspace_bytes() minus the
 *     1 MiB of slack that function adds for callers with several launch
 *     groups; at most 64 tables per call.
 *   - The first call on each device resolves the array's device address
 *     (two host threads may make it at the same time); make that call before
 *     any stream capture begins.
 *   - All calls on a device share the array and find the previous call's
 *     data in it; the reduce kernel's per-block status words, the only part
 *     that must start at zero, are cleared by each call.  Calls that overlap on the GPU race on the
 *     array: issue them one after the other on one stream.
 * Inputs, outputs and results are those of ffemb_embed_backward_sgd. */
size_t ffemb_static_scratch_bytes(void);
int ffemb_embed_backward_sgd_static(const ffemb_lookup_t *lookups,
                                    int num_tables, int batch, int out_dim,
                                    enum AggrMode aggr,
                                    const ffemb_matrix_t *grad, float lr,
                                    int *err_flag, void *stream);

/* W[i] = (max-min) * u_i + min, u_i in (0,1] from a counter-based generator
 * keyed by (seed, i).  Stands in for UniformInitializer::init_task
 * (src/runtime/initializer_kernel.cu:24-85) + scale_kernel
 * (src/runtime/cuda_helper.cu:3-11); same distribution, reproducible stream. */
int ffemb_uniform_init(float *weight, int64_t count, int seed, float min_val,
                       float max_val, void *stream);

/* Dense SGD step W[i] = fmaf(-lr, G[i], W[i]) followed by G[i] = 0: the
 * reference's sgd_update (src/runtime/optimizer_kernel.cu:23-41, wd = m = 0)
 * plus its zero_gradients (src/runtime/model.cc:256-280).  Used for small
 * tables that are replicated on every rank instead of sharded -- the
 * reference's data-parallel embedding (replicated table, summed gradient
 * replicas: src/runtime/model.cc:697-724, optimizer_kernel.cu:96-101): each
 * rank builds the dense gradient of its own samples with
 * ffemb_embed_backward_sgd(lr = -1) into a zeroed G, the replicas are summed
 * by an all-reduce, and every rank applies the same dense step. */
int ffemb_dense_sgd(float *weight, float *grad, int64_t count, float lr,
                    void *stream);

/* Peer-memory plumbing of the table-wise sharded path (one process per GPU).
 * Stands in for what Legion DMA does for the reference: embedding outputs and
 * their gradients cross GPUs through zero-copy host memory
 * (src/ops/embedding.cu:280-284,343-347; src/ops/concat.cu:85-93).  A rank
 * allocates its [batch/N][T*D] output and gradient matrices with
 * ffemb_device_malloc (plain cudaMalloc: IPC-exportable), publishes the
 * handles, opens its peers', and passes the mapped pointers as the parts of
 * an ffemb_matrix_t: the forward kernel then stores pooled rows straight into
 * the consuming rank's matrix over NVLink and the update kernel loads
 * gradient rows straight from the producing rank's. */
int ffemb_device_malloc(void **ptr, size_t bytes);
int ffemb_device_free(void *ptr);
int ffemb_ipc_get_handle(void *device_ptr, uint8_t handle[64]);
int ffemb_ipc_open_handle(const uint8_t handle[64], void **device_ptr);
int ffemb_ipc_close_handle(void *device_ptr);
/* Stream-ordered barrier among the `world` ranks of one box through peer
 * memory: peer_flags[q] (HOST array of device pointers) is the mapped base of
 * rank q's 32 zero-initialised uint32 words (word r < world: flag written by
 * rank r; word 16: rank q's own barrier count, kept on the device so that a
 * captured launch can be replayed).  All ranks must call it the same number
 * of times.  After it, everything every peer enqueued before ITS barrier has
 * completed.  A peer that does not arrive within 5 s sets
 * *err_flag = FFEMB_ERR_COMM instead of hanging. */
int ffemb_peer_barrier(uint32_t *const *peer_flags, int rank, int world,
                       int *err_flag, void *stream);
/* Backward exchange as a push.  A rank holds the upstream gradient of its own
 * samples for all tables, src = [rows][src_ld]; item i copies the column block
 * [src_col, src_col + width) of every row to dst + row*dst_ld, where dst is
 * (a peer mapping of) the owning rank's staging matrix, already advanced to
 * this rank's first sample and the table's first column.  Stands in for the
 * gradient half of the reference's repartition at Concat
 * (src/ops/concat.cu:263-320, src/ops/embedding.cu:343-347).  `items` is a
 * HOST array. */
typedef struct ffemb_push_t {
  float *dst;
  int64_t dst_ld;
  int64_t src_col;
} ffemb_push_t;
int ffemb_push_columns(const ffemb_push_t *items, int num_items, int rows,
                       int width, const float *src, int64_t src_ld,
                       void *stream);
/* Replicated small tables: sum of the ranks' dense gradient replicas fused
 * with the SGD step, over peer memory.  weights[q] / grads[q] (HOST arrays of
 * device pointers, 16-byte aligned) are the mapped bases of rank q's flat
 * [count] copy of the replicated tables and of its dense gradient.  Rank r
 * handles the r-th slice: loads that slice of every rank's gradient, adds
 * them in rank order (replica 0 += replica q, src/runtime/
 * optimizer_kernel.cu:96-101), applies W = fmaf(-lr, sum, W) (sgd_update,
 * :23-41) and stores the result into every rank's copy.  Bracket it with two
 * ffemb_peer_barrier calls: gradients complete before, copies written after. */
int ffemb_replica_reduce_update(float *const *weights, const float *const *grads,
                                int64_t count, int rank, int world, float lr,
                                void *stream);

/* ------------------------------------------------------------------------
 * (M) Model-level entry points (flexflow_c.h surface for this path)
 * --------------------------------------------------------------------- */

#define FFEMB_NEW_OPAQUE_TYPE(T) typedef struct T { void *impl; } T

FFEMB_NEW_OPAQUE_TYPE(ffemb_config_t);            /* flexflow_c.h:27 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_model_t);             /* :28 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_tensor_t);            /* :29 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_sgd_optimizer_t);     /* :30 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_initializer_t);       /* :32 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_zero_initializer_t);  /* :34 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_uniform_initializer_t); /* :35 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_op_t);                /* :37 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_parameter_t);         /* :38 */
FFEMB_NEW_OPAQUE_TYPE(ffemb_strategy_t);          /* FFProtoBuf::Strategy */

/* --- FFConfig (flexflow_c.h:49-79; flags of src/runtime/model.cc:1313-1381) */
ffemb_config_t ffemb_config_create(void);
void ffemb_config_destroy(ffemb_config_t handle);
/* "-b" batch size (global batch: every embedding op sees the whole batch,
 * reference strategy dims [1,1]); "-ll:gpu" workers per node; "--nodes". */
int ffemb_config_set_batch_size(ffemb_config_t handle, int batch_size);
int ffemb_config_get_batch_size(ffemb_config_t handle);
int ffemb_config_set_workers_per_node(ffemb_config_t handle, int n);
int ffemb_config_get_workers_per_node(ffemb_config_t handle);
int ffemb_config_set_num_nodes(ffemb_config_t handle, int n);
int ffemb_config_get_num_nodes(ffemb_config_t handle);
/* This process's rank in [0, workers_per_node*num_nodes) and CUDA device.
 * (The reference is one process for all GPUs under Legion; here it is one
 * process per GPU.) */
int ffemb_config_set_rank(ffemb_config_t handle, int rank, int cuda_device);
int ffemb_config_get_rank(ffemb_config_t handle);
/* "--import[-strategy] <file>" (src/mapper/mapper.cc:282-286). */
int ffemb_config_import_strategy_file(ffemb_config_t handle, const char *path);
/* With more than one rank: tables with fewer than `rows` rows are kept on
 * EVERY rank instead of sharded -- the reference's data-parallel embedding
 * (replicated table, gradient replicas summed: src/runtime/model.cc:697-724,
 * optimizer_kernel.cu:96-101) -- because for them one sum of the dense
 * gradient replicas is cheaper than two exchanges of pooled rows.  0 (the
 * default): shard everything, as the reference's strategy files do; -1: the
 * break-even batch / ranks.  Without a strategy the remaining sharded tables
 * are dealt round-robin, in table order, over the ranks. */
int ffemb_config_set_replicate_below(ffemb_config_t handle, int64_t rows);
/* How the upstream gradients reach the tables' owners.  1 push: one kernel
 * stores every table's block of a rank's gradient into its owner's staging
 * matrix and the update reads local memory; 2 pull: the update kernel loads
 * gradient rows straight from the producing ranks' matrices (no staging
 * pass, but remote loads reach roughly half the NVLink rate of remote
 * stores); 0 (default): pull with two ranks and one-hot bags, push otherwise --
 * where each measured faster (with bags of L lookups a pull would fetch every
 * remote gradient row L times). */
int ffemb_config_set_gradient_exchange(ffemb_config_t handle, int mode);
/* Run the sharded engine (exchange buffer, barriers, staging) even with one
 * rank.  For tests on a single GPU. */
int ffemb_config_set_exchange_engine(ffemb_config_t handle, int force);

/* --- FFModel (flexflow_c.h:86-135) */
ffemb_model_t ffemb_model_create(ffemb_config_t config);
void ffemb_model_destroy(ffemb_model_t handle);

/* flexflow_tensor_create (flexflow_c.h:371-378): a model input.  On this path
 * that is the sparse input of one table: dims = {batch, bag_size}, DT_INT64
 * (examples/cpp/DLRM/dlrm.cc:100-108). */
ffemb_tensor_t ffemb_tensor_create(ffemb_model_t model, int num_dims,
                                   const int *dims, const char *name,
                                   enum DataType data_type, bool create_grad);
void ffemb_tensor_destroy(ffemb_tensor_t handle);
int ffemb_tensor_get_num_dims(ffemb_tensor_t handle);
int ffemb_tensor_get_dim(ffemb_tensor_t handle, int axis);
int ffemb_tensor_get_data_type(ffemb_tensor_t handle);

/* flexflow_model_add_embedding (flexflow_c.h:193-200; flexflow_c.cc:376-393;
 * FFModel::embedding, src/ops/embedding.cu:19-33).  Same argument order and
 * meaning, including the reference's spelling `num_entires`.  `shared_op`
 * only shares the op name, not storage (src/runtime/model.cc:164-169).
 * Returns the output tensor [batch, out_dim] DT_FLOAT; impl == NULL on error. */
ffemb_tensor_t ffemb_model_add_embedding(ffemb_model_t handle,
                                         const ffemb_tensor_t input,
                                         int num_entires, int out_dim,
                                         enum AggrMode aggr,
                                         ffemb_op_t shared_op,
                                         ffemb_initializer_t kernel_initializer);

/* flexflow_model_set_sgd_optimizer (flexflow_c.h:338-341). */
int ffemb_model_set_sgd_optimizer(ffemb_model_t handle,
                                  ffemb_sgd_optimizer_t optimizer);

/* flexflow_model_compile (flexflow_c.h:122-127): allocates the tables this
 * rank owns under the placement, runs the initialisers, sizes the backward
 * workspace.  Loss/metrics arguments of the reference do not exist on this
 * path. */
int ffemb_model_compile(ffemb_model_t handle);
/* --- Sharded mode (workers_per_node * num_nodes > 1; one process per GPU).
 * The same program runs on every rank: all tables are added on all ranks,
 * compile() places them (strategy -> owner, FFMapper::slice_task) and
 * allocates, per rank, the tables it owns plus ONE exchange buffer holding
 * its [batch/ranks][sum out_dim] output and gradient matrices (every table's
 * block side by side: the reference's Concat layout), the barrier words, the
 * replicated tables and the gradient staging.  The ranks then swap the 64-byte
 * IPC handles of those buffers out of band (torch.distributed, MPI, a file):
 * every rank calls get_exchange_handle, gathers all ranks' handles in rank
 * order and passes them to connect_peers.  After that
 *   forward()   gathers the whole batch for the tables this rank owns and
 *               stores each pooled row straight into the consuming rank's
 *               matrix over NVLink; replicated tables are gathered locally;
 *               a peer-memory barrier ends it;
 *   backward()  pushes every sharded table's block of this rank's gradient
 *               matrix into its owner's staging and builds the replicas' dense
 *               gradients; barrier;
 *   update()    applies the sparse SGD step to the owned tables and, for the
 *               replicas, sums the ranks' gradients and writes the stepped
 *               rows into every rank's copy in one kernel over peer memory.
 * Stands in for the reference's implicit repartition at Concat
 * (src/ops/concat.cu:85-93; embedding.cu:280-284,343-347) and the mapper's
 * placement (src/mapper/mapper.cc:33-97).  Output tensors then describe this
 * rank's samples: get_float / set_grad_float take (batch/ranks)*out_dim
 * values, get_device_ptr points into the exchange matrices.  Work that
 * consumes the outputs or produces the gradients must be enqueued on the
 * model's stream; every rank must make the same calls in the same order. */
int ffemb_model_is_sharded(ffemb_model_t handle);
int ffemb_model_get_exchange_handle(ffemb_model_t handle, uint8_t ipc_handle[64]);
/* `handles`: count * 64 bytes, rank q's handle at handles + 64*q. */
int ffemb_model_connect_peers(ffemb_model_t handle, const uint8_t *handles, int count);
/* This rank's [batch/ranks][ld] output (grad = false) or gradient matrix. */
int ffemb_model_get_exchange_matrix(ffemb_model_t handle, bool grad, float **ptr,
                                    int64_t *ld, int *rows);

/* Development aid, sharded mode: record timed events at the phase boundaries
 * of the following steps (not under stream capture); the report lists, per
 * boundary, the mean time since the step's forward began, and clears the
 * record.  Returns the size the text needs. */
int ffemb_model_trace(ffemb_model_t handle, int enable);
size_t ffemb_model_trace_report(ffemb_model_t handle, char *buf, size_t cap);

/* flexflow_model_init_layers (:98-100) -- EMBED_INIT_TASK is a no-op in the
 * reference too (src/ops/embedding.cu:137-144). */
int ffemb_model_init_layers(ffemb_model_t handle);
/* flexflow_model_forward / zero_gradients / backward / update (:106-135).
 * zero_gradients is a no-op: no dense gradient exists.  backward builds the
 * sorted sparse gradient; update applies SGD to the touched rows.  The
 * reference's backward accumulates into the dense gradient (embedding.cu:
 * 350-352); here a second backward() before update() / zero_gradients()
 * returns FFEMB_ERR_UNSUPPORTED instead of replacing the first gradient. */
int ffemb_model_forward(ffemb_model_t handle);
int ffemb_model_zero_gradients(ffemb_model_t handle);
int ffemb_model_backward(ffemb_model_t handle);
int ffemb_model_update(ffemb_model_t handle);
/* Blocks until the model's stream is idle, then reports (and clears) a
 * pending out-of-range-index error. */
int ffemb_model_sync(ffemb_model_t handle);
/* forward() of a model with an optimizer starts the coming backward's key
 * sort on a side stream (it needs only the indices) so that it runs under the
 * gather; `enable = false` leaves the sort to backward() (inference-style
 * forward calls, measurements of the kernels in isolation).  Default: true. */
int ffemb_model_set_presort(ffemb_model_t handle, bool enable);
/* Use `stream` (a cudaStream_t) for all model work; NULL = default stream. */
int ffemb_model_set_stream(ffemb_model_t handle, void *stream);

int ffemb_model_get_num_layers(ffemb_model_t handle);
ffemb_op_t ffemb_model_get_layer_by_id(ffemb_model_t handle, int layer_id);

/* --- Op (flexflow_c.h:720-743) */
ffemb_parameter_t ffemb_op_get_parameter_by_id(ffemb_op_t handle, int id);
ffemb_tensor_t ffemb_op_get_input_by_id(ffemb_op_t handle, int id);
ffemb_tensor_t ffemb_op_get_output_by_id(ffemb_op_t handle, int id);
/* "Embed_<N>x<D>_<guid>" (src/runtime/model.cc:157-171) */
const char *ffemb_op_get_name(ffemb_op_t handle);
/* Rank that owns this table under the compiled placement. */
int ffemb_op_get_device(ffemb_op_t handle);

/* --- Tensor data movement.  The reference moves data with dataloader tasks
 * (examples/cpp/DLRM/dlrm.cu:19-56) and inline mappings
 * (flexflow_tensor_inline_map, flexflow_c.h:395-430); here it is explicit. */
/* Host -> device copy into an int64 input tensor (enqueued on the model
 * stream; `data` should be pinned for the copy to be asynchronous). */
int ffemb_tensor_set_int64(ffemb_tensor_t handle, ffemb_model_t model,
                           const int64_t *data, int64_t count);
/* Host -> device copy into a float tensor's gradient (d loss / d out). */
int ffemb_tensor_set_grad_float(ffemb_tensor_t handle, ffemb_model_t model,
                                const float *data, int64_t count);
/* Device -> host copy of a float tensor (an embedding output). */
int ffemb_tensor_get_float(ffemb_tensor_t handle, ffemb_model_t model,
                           float *data, int64_t count);
/* --- Double-buffered host ingest / egress.  The reference stages every
 * table's indices per step through a freshly allocated pinned buffer and a
 * synchronous cudaMemcpy (examples/cpp/DLRM/dlrm.cu:19-56, from
 * DataLoader::next_batch, dlrm.cc:486-524) and keeps embedding outputs and
 * their gradients in zero-copy host memory (src/ops/embedding.cu:280-284,
 * 343-347).  Here two slots of device buffers and two library-owned copy
 * streams let the H2D of batch i+1 and the D2H of batch i-1 run under the
 * kernels of batch i.  Per step:
 *     ffemb_model_begin_batch(m);             the loaded slot becomes current
 *     ffemb_model_forward(m);
 *     ffemb_model_fetch_outputs_async(m, o);  pooled outputs -> host
 *     ffemb_model_zero_gradients(m); ffemb_model_backward(m); ffemb_model_update(m);
 *     ffemb_model_load_batch_async(m, idx, g);   the NEXT batch, host -> idle slot
 *     ffemb_model_wait_outputs(m, false);     the previous step's outputs are on the host
 * (one ffemb_model_load_batch_async before the first step).  Host buffers
 * should be pinned.  `indices[i]`: host [batch][bag] int64 of layer i (NULL for
 * a table this rank does not read; of a replicated table only this rank's own
 * samples are copied).  `out_grad` (may be NULL) and `out_host`: host
 * [rows][sum of out_dim] fp32, every table's block side by side in layer
 * order, rows = batch (one rank) or batch / ranks (sharded: this rank's
 * samples).  While the pipeline is in use it owns the tensors' buffers:
 * output / gradient blocks attached with ffemb_tensor_attach_device_ptr are
 * replaced. */
int ffemb_model_load_batch_async(ffemb_model_t handle, const int64_t *const *indices,
                                 const float *out_grad);
int ffemb_model_begin_batch(ffemb_model_t handle);
int ffemb_model_fetch_outputs_async(ffemb_model_t handle, float *out_host);
/* Blocks the host until every fetch but the most recent one has landed (all
 * of them when `all` is true). */
int ffemb_model_wait_outputs(ffemb_model_t handle, bool all);

/* Borrow caller-owned device memory instead of the library's own buffer
 * (torch tensors: data_ptr()).  `ld` is the row stride in elements; for an
 * output tensor, attaching column blocks of one [batch][T*D] matrix yields
 * the Concat layout directly.  grad = true attaches the gradient buffer. */
int ffemb_tensor_attach_device_ptr(ffemb_tensor_t handle, void *device_ptr,
                                   int64_t ld, bool grad);
void *ffemb_tensor_get_device_ptr(ffemb_tensor_t handle, bool grad);
int64_t ffemb_tensor_get_ld(ffemb_tensor_t handle, bool grad);

/* --- Parameter (flexflow_c.h:447-459; src/runtime/model.cu:260-334).
 * Layout is row-major [num_entries][out_dim] (what the reference's kernels
 * index, src/ops/embedding.cu:189,221).  On a rank that does not own the
 * table both calls fail with FFEMB_ERR_INVALID_ARGUMENT. */
bool ffemb_parameter_set_weights_float(ffemb_parameter_t handle,
                                       ffemb_model_t model, int num_dim,
                                       const int *dims, const float *data);
bool ffemb_parameter_get_weights_float(ffemb_parameter_t handle,
                                       ffemb_model_t model, float *data);
/* Device pointer of the table (NULL when not owned), for zero-copy wrapping. */
float *ffemb_parameter_get_device_ptr(ffemb_parameter_t handle);
int64_t ffemb_parameter_get_num_entries(ffemb_parameter_t handle);
int ffemb_parameter_get_out_dim(ffemb_parameter_t handle);

/* --- SGDOptimizer (flexflow_c.h:465-480; include/optimizer.h:36-54).
 * The sparse fused update is exact only for momentum == 0 and
 * weight_decay == 0 (DLRM: examples/cpp/DLRM/dlrm.cc:134); other settings
 * make ffemb_model_update return FFEMB_ERR_UNSUPPORTED. */
ffemb_sgd_optimizer_t ffemb_sgd_optimizer_create(ffemb_model_t model, double lr,
                                                 double momentum, bool nesterov,
                                                 double weight_decay);
void ffemb_sgd_optimizer_destroy(ffemb_sgd_optimizer_t handle);
void ffemb_sgd_optimizer_set_lr(ffemb_sgd_optimizer_t handle, double lr);

/* --- Initializers (flexflow_c.h:526-545) */
ffemb_uniform_initializer_t ffemb_uniform_initializer_create(int seed, float min,
                                                             float max);
void ffemb_uniform_initializer_destroy(ffemb_uniform_initializer_t handle);
ffemb_zero_initializer_t ffemb_zero_initializer_create(void);
void ffemb_zero_initializer_destroy(ffemb_zero_initializer_t handle);

/* --- Placement: strategy files and the mapper hook.
 * FFProtoBuf::Strategy wire format, src/runtime/strategy.proto:5-23;
 * load/save, src/runtime/strategy.cc:96-172; FFMapper::slice_task,
 * src/mapper/mapper.cc:33-97 (point -> devices[device_ids[i] % ndevices]). */
ffemb_strategy_t ffemb_strategy_create(void);
void ffemb_strategy_destroy(ffemb_strategy_t handle);
int ffemb_strategy_load_file(ffemb_strategy_t handle, const char *path);
int ffemb_strategy_save_file(ffemb_strategy_t handle, const char *path);
int ffemb_strategy_parse(ffemb_strategy_t handle, const uint8_t *bytes, size_t n);
/* Serialises into `bytes` (capacity `cap`); returns the size needed. */
size_t ffemb_strategy_serialize(ffemb_strategy_t handle, uint8_t *bytes, size_t cap);
int ffemb_strategy_num_ops(ffemb_strategy_t handle);
const char *ffemb_strategy_op_name(ffemb_strategy_t handle, int i);
int ffemb_strategy_op_device_type(ffemb_strategy_t handle, int i);
int ffemb_strategy_op_num_dims(ffemb_strategy_t handle, int i);
int ffemb_strategy_op_dim(ffemb_strategy_t handle, int i, int k);
int ffemb_strategy_op_num_device_ids(ffemb_strategy_t handle, int i);
int ffemb_strategy_op_device_id(ffemb_strategy_t handle, int i, int k);
int ffemb_strategy_op_num_memory_types(ffemb_strategy_t handle, int i);
int ffemb_strategy_op_memory_type(ffemb_strategy_t handle, int i, int k);
int ffemb_strategy_add_op(ffemb_strategy_t handle, const char *name,
                          int device_type, const int *dims, int num_dims,
                          const int *device_ids, int num_device_ids,
                          const int *memory_types, int num_memory_types);
/* Index of the op called `name`, or -1 (the reference hashes the name,
 * src/runtime/strategy.cc:23-26). */
int ffemb_strategy_find(ffemb_strategy_t handle, const char *name);
/* The generator of src/runtime/dlrm_strategy.cc:242-296: embedding<i> ->
 * device i % (gpus_per_node*num_nodes); concat / linear / mse_loss entries as
 * the reference writes them. */
int ffemb_strategy_generate_dlrm(ffemb_strategy_t handle, int gpus_per_node,
                                 int embs_per_node, int num_nodes);
/* The generator of src/runtime/dlrm_strategy.py:26-46, which produced the
 * reference's checked-in dlrm_strategy_<E>embs_<G>gpus.pb fixtures:
 * embedding<i> -> dims [1,1], device i % num_gpu; linear, mse_loss, concat ->
 * dims [1,num_gpu], devices 0..num_gpu-1.  Byte-identical output. */
int ffemb_strategy_generate_dlrm_simple(ffemb_strategy_t handle, int num_gpu,
                                        int num_emb);
/* Mapper hook: device of table `table_index` (also looked up under the names
 * "embedding<i>" and `op_name`), folded by `% num_devices` as slice_task
 * does; round-robin table_index % num_devices when the strategy has no entry
 * (or handle.impl == NULL). */
int ffemb_strategy_device_for_table(ffemb_strategy_t handle, int table_index,
                                    const char *op_name, int num_devices);
/* Cost-aware placement (the reference cannot cost embeddings,
 * src/ops/embedding.cu:357-364): greedy longest-processing-time assignment of
 * tables to devices by lookups per step, subject to bytes_capacity per
 * device.  Writes device ids into owner[num_tables].  */
int ffemb_placement_balance(const int64_t *num_entries, const int64_t *lookups,
                            int num_tables, int out_dim, int num_devices,
                            int64_t bytes_capacity, int *owner);
/* Cost-aware placement WITH the replicate option (the reference's
 * simulator cannot cost an embedding: src/ops/embedding.cu:357-364,
 * src/runtime/simulator.cc:200-275).  Models a step of the sharded engine per
 * rank -- NVLink bytes of the two exchanges and of the replicas' sum, HBM
 * bytes of the gathers and updates -- and picks, among "replicate the k
 * smallest tables, balance the rest by lookups", the k with the shortest
 * modelled step.  owner[t] = rank, or -1 = keep table t on every rank.
 * `lookups[t]`: lookups of table t per step over the whole batch.  cost ==
 * NULL: this pool's measured B200 rates.  modelled_step_us may be NULL. */
typedef struct ffemb_cost_model_t {
  double hbm_gbs;            /* rate the gather / reduce kernels sustain        */
  double nvlink_gbs;         /* rate of peer stores issued from inside a kernel */
  double hbm_passes;         /* row-sized HBM accesses per lookup (5: row read,
                                output write, gradient read, row read + write)  */
  double replica_latency_us; /* fixed cost of having any replicated table       */
} ffemb_cost_model_t;
int ffemb_placement_plan(const int64_t *num_entries, const int64_t *lookups,
                         int num_tables, int out_dim, int batch, int num_devices,
                         int64_t bytes_capacity, const ffemb_cost_model_t *cost,
                         int *owner, double *modelled_step_us);
/* Model-level: use this strategy at compile (overrides the config's file). */
int ffemb_model_set_strategy(ffemb_model_t model, ffemb_strategy_t strategy);

#ifdef __cplusplus
}
#endif

#endif /* FFEMB_C_H */
