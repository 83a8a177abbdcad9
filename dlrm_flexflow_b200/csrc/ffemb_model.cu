// ffemb_model.cu -- model-level C-ABI (the flexflow_c.h surface for this path).
//
// Mirrors, for the embedding path only, the handle objects of the reference's
// python/flexflow_c.cc: FFConfig, FFModel, Tensor, Op, Parameter,
// SGDOptimizer, Uniform/Zero initialisers.  Everything Legion did for this
// path (regions, partitions, mapping) collapses to: which rank owns which
// table (the placement hook), device buffers, one stream.
#include <string.h>

#include <memory>
#include <vector>

#include "ffemb_common.cuh"

namespace ffemb {

struct Model;
struct Op;

struct Config {
  int batch_size = 64;        // FFConfig default, src/runtime/model.cc:1273-1289
  int workers_per_node = 1;
  int num_nodes = 1;
  int rank = 0;
  int device = -1;            // -1: whatever device is current at model_create
  std::string strategy_file;
};

struct Initializer {
  enum Kind { kUniform, kZero } kind;
  int seed;
  float lo, hi;
};

struct Buffer {  // a device matrix view, owned or borrowed
  void *ptr = nullptr;
  int64_t ld = 0;
  bool owned = false;
};

struct Tensor {
  Model *model = nullptr;
  Op *producer = nullptr;  // embedding op whose output this is (else an input)
  std::string name;
  int num_dims = 0;
  int dims[4] = {0, 0, 0, 0};
  int data_type = DT_FLOAT;
  bool create_grad = false;
  Buffer data, grad;
};

struct Parameter {
  Op *op;
};

struct Op {
  Model *model;
  int index;
  std::string name;
  Tensor *input, *output;
  int64_t N;
  int D;
  int aggr;
  Initializer init;
  float *W = nullptr;
  int owner = 0;
  Parameter param;
};

struct Sgd {
  double lr, momentum;
  bool nesterov;
  double weight_decay;
};

struct Strategy;  // ffemb_strategy.cc

struct Model {
  Config cfg;
  int device = 0;
  int world = 1;
  std::vector<std::unique_ptr<Tensor>> tensors;
  std::vector<std::unique_ptr<Op>> ops;
  Sgd *opt = nullptr;
  ffemb_strategy_t strategy{nullptr};
  bool own_strategy = false;
  cudaStream_t stream = nullptr;
  bool compiled = false;
  bool sorted = false;  // backward() ran since the last update()
  // forward() starts the backward's key sort on a side stream (it needs only
  // the indices), so that it runs under the bandwidth-bound gather
  bool presort = true;
  bool presorted = false;
  std::vector<int64_t> presort_sig;
  cudaStream_t side = nullptr;
  cudaEvent_t ev_inputs = nullptr, ev_sorted = nullptr;
  void *workspace = nullptr;
  size_t workspace_bytes = 0;
  int *err_dev = nullptr;
  float *out_mat = nullptr, *grad_mat = nullptr;  // library-owned [B][sum D] concat matrices
  int64_t mat_ld = 0;
  int op_guid = 100;  // Op guids start at 100 in the reference (model.cc:1165)
};

static Config *C(ffemb_config_t h) { return (Config *)h.impl; }
static Model *M(ffemb_model_t h) { return (Model *)h.impl; }
static Tensor *T(ffemb_tensor_t h) { return (Tensor *)h.impl; }
static Op *O(ffemb_op_t h) { return (Op *)h.impl; }
static Parameter *P(ffemb_parameter_t h) { return (Parameter *)h.impl; }

static bool owned(const Op *op) { return op->owner == op->model->cfg.rank; }

// One batched launch: owned ops that share out_dim, aggr and row strides.
struct LaunchGroup {
  int D, aggr;
  int64_t ld_out, ld_grad;
  std::vector<Op *> ops;
};

static std::vector<LaunchGroup> groups_of(Model *m, bool need_grad) {
  std::vector<LaunchGroup> gs;
  for (auto &up : m->ops) {
    Op *op = up.get();
    if (!owned(op)) continue;
    const int64_t ldo = op->output->data.ld, ldg = need_grad ? op->output->grad.ld : 0;
    LaunchGroup *g = nullptr;
    for (auto &c : gs)
      if (c.D == op->D && c.aggr == op->aggr && c.ld_out == ldo && c.ld_grad == ldg &&
          (int)c.ops.size() < kMaxTables)
        g = &c;
    if (!g) {
      gs.push_back(LaunchGroup{op->D, op->aggr, ldo, ldg, {}});
      g = &gs.back();
    }
    g->ops.push_back(op);
  }
  return gs;
}

// lookups + matrix for one group; `grad` selects output-gradient buffers.
static int describe(const LaunchGroup &g, bool grad, std::vector<ffemb_lookup_t> &lk,
                    ffemb_matrix_t &mat, bool indices_only = false) {
  float *base = nullptr;
  for (Op *op : g.ops) {
    if (indices_only) break;
    float *p = (float *)(grad ? op->output->grad.ptr : op->output->data.ptr);
    if (!p)
      return fail(FFEMB_ERR_INVALID_ARGUMENT,
                  op->name + (grad ? ": no output gradient set" : ": no output buffer"));
    if (!base || p < base) base = p;
  }
  lk.clear();
  for (Op *op : g.ops) {
    if (!op->input->data.ptr)
      return fail(FFEMB_ERR_INVALID_ARGUMENT, op->name + ": input tensor has no data");
    float *p = (float *)(grad ? op->output->grad.ptr : op->output->data.ptr);
    ffemb_lookup_t l;
    memset(&l, 0, sizeof l);
    l.weight = op->W;
    l.indices = (const int64_t *)op->input->data.ptr;
    l.offsets = nullptr;
    l.num_entries = op->N;
    l.bag_size = op->input->dims[1];
    l.num_indices = (int64_t)op->input->dims[0] * l.bag_size;
    l.col = indices_only ? 0 : p - base;
    lk.push_back(l);
  }
  memset(&mat, 0, sizeof mat);
  if (indices_only) return FFEMB_OK;
  mat.base[0] = base;
  mat.num_parts = 1;
  mat.rows_per_part = g.ops[0]->input->dims[0];
  mat.ld = grad ? g.ld_grad : g.ld_out;
  return FFEMB_OK;
}

static void release(Buffer &b) {
  if (b.owned && b.ptr) cudaFree(b.ptr);
  b = Buffer();
}

}  // namespace ffemb

using namespace ffemb;

// ---- FFConfig ----------------------------------------------------------------
extern "C" ffemb_config_t ffemb_config_create(void) { return ffemb_config_t{new Config}; }
extern "C" void ffemb_config_destroy(ffemb_config_t h) { delete C(h); }
extern "C" int ffemb_config_set_batch_size(ffemb_config_t h, int b) {
  if (!C(h) || b <= 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_batch_size: bad argument");
  C(h)->batch_size = b;
  return FFEMB_OK;
}
extern "C" int ffemb_config_get_batch_size(ffemb_config_t h) { return C(h) ? C(h)->batch_size : -1; }
extern "C" int ffemb_config_set_workers_per_node(ffemb_config_t h, int n) {
  if (!C(h) || n <= 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_workers_per_node: bad argument");
  C(h)->workers_per_node = n;
  return FFEMB_OK;
}
extern "C" int ffemb_config_get_workers_per_node(ffemb_config_t h) { return C(h) ? C(h)->workers_per_node : -1; }
extern "C" int ffemb_config_set_num_nodes(ffemb_config_t h, int n) {
  if (!C(h) || n <= 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_num_nodes: bad argument");
  C(h)->num_nodes = n;
  return FFEMB_OK;
}
extern "C" int ffemb_config_get_num_nodes(ffemb_config_t h) { return C(h) ? C(h)->num_nodes : -1; }
extern "C" int ffemb_config_set_rank(ffemb_config_t h, int rank, int cuda_device) {
  if (!C(h) || rank < 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_rank: bad argument");
  C(h)->rank = rank;
  C(h)->device = cuda_device;
  return FFEMB_OK;
}
extern "C" int ffemb_config_get_rank(ffemb_config_t h) { return C(h) ? C(h)->rank : -1; }
extern "C" int ffemb_config_import_strategy_file(ffemb_config_t h, const char *path) {
  if (!C(h) || !path) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_import_strategy_file: NULL argument");
  C(h)->strategy_file = path;
  return FFEMB_OK;
}

// ---- FFModel -----------------------------------------------------------------
extern "C" ffemb_model_t ffemb_model_create(ffemb_config_t config) {
  if (!C(config)) {
    set_error("model_create: NULL config");
    return ffemb_model_t{nullptr};
  }
  Model *m = new Model;
  m->cfg = *C(config);
  m->world = m->cfg.workers_per_node * m->cfg.num_nodes;
  if (m->cfg.rank >= m->world) {
    set_error("model_create: rank >= workers_per_node * num_nodes");
    delete m;
    return ffemb_model_t{nullptr};
  }
  return ffemb_model_t{m};
}

extern "C" void ffemb_model_destroy(ffemb_model_t h) {
  Model *m = M(h);
  if (!m) return;
  if (m->compiled) cudaSetDevice(m->device);
  for (auto &t : m->tensors) {
    release(t->data);
    release(t->grad);
  }
  for (auto &op : m->ops)
    if (op->W) cudaFree(op->W);
  if (m->workspace) cudaFree(m->workspace);
  if (m->err_dev) cudaFree(m->err_dev);
  if (m->side) cudaStreamDestroy(m->side);
  if (m->ev_inputs) cudaEventDestroy(m->ev_inputs);
  if (m->ev_sorted) cudaEventDestroy(m->ev_sorted);
  if (m->out_mat) cudaFree(m->out_mat);
  if (m->grad_mat) cudaFree(m->grad_mat);
  if (m->own_strategy) ffemb_strategy_destroy(m->strategy);
  delete m;
}

extern "C" ffemb_tensor_t ffemb_tensor_create(ffemb_model_t model, int num_dims, const int *dims,
                                              const char *name, enum DataType data_type,
                                              bool create_grad) {
  Model *m = M(model);
  if (!m || !dims || num_dims < 1 || num_dims > 4) {
    set_error("tensor_create: bad arguments");
    return ffemb_tensor_t{nullptr};
  }
  if (m->compiled) {
    set_error("tensor_create: model is already compiled");
    return ffemb_tensor_t{nullptr};
  }
  auto t = std::make_unique<Tensor>();
  t->model = m;
  t->name = name ? name : "";
  t->num_dims = num_dims;
  for (int i = 0; i < num_dims; ++i) {
    if (dims[i] <= 0) {
      set_error("tensor_create: non-positive dimension");
      return ffemb_tensor_t{nullptr};
    }
    t->dims[i] = dims[i];
  }
  t->data_type = data_type;
  t->create_grad = create_grad;
  m->tensors.push_back(std::move(t));
  return ffemb_tensor_t{m->tensors.back().get()};
}
extern "C" void ffemb_tensor_destroy(ffemb_tensor_t) { /* storage belongs to the model */ }
extern "C" int ffemb_tensor_get_num_dims(ffemb_tensor_t h) { return T(h) ? T(h)->num_dims : -1; }
extern "C" int ffemb_tensor_get_dim(ffemb_tensor_t h, int axis) {
  return T(h) && axis >= 0 && axis < T(h)->num_dims ? T(h)->dims[axis] : -1;
}
extern "C" int ffemb_tensor_get_data_type(ffemb_tensor_t h) { return T(h) ? T(h)->data_type : -1; }

extern "C" ffemb_tensor_t ffemb_model_add_embedding(ffemb_model_t handle, const ffemb_tensor_t input,
                                                    int num_entires, int out_dim, enum AggrMode aggr,
                                                    ffemb_op_t shared_op,
                                                    ffemb_initializer_t kernel_initializer) {
  Model *m = M(handle);
  Tensor *in = T(input);
  const ffemb_tensor_t none{nullptr};
  if (!m || !in || in->model != m) {
    set_error("add_embedding: NULL model/input or input of another model");
    return none;
  }
  if (m->compiled) {
    set_error("add_embedding: model is already compiled");
    return none;
  }
  // src/ops/embedding.cu:60 assert(_input.numDim == 2); dlrm.cc:105-107 int64
  if (in->num_dims != 2 || in->data_type != DT_INT64) {
    set_error("add_embedding: input must be a 2-D [batch, bag] DT_INT64 tensor");
    return none;
  }
  if (in->dims[0] != m->cfg.batch_size) {
    set_error("add_embedding: input dim 0 must equal the config's batch size");
    return none;
  }
  if (num_entires <= 0 || out_dim <= 0) {
    set_error("add_embedding: num_entries and out_dim must be positive");
    return none;
  }
  if (aggr != AGGR_MODE_SUM && aggr != AGGR_MODE_AVG) {
    set_error("add_embedding: aggr must be AGGR_MODE_SUM or AGGR_MODE_AVG");
    return none;
  }
  auto op = std::make_unique<Op>();
  op->model = m;
  op->index = (int)m->ops.size();
  op->input = in;
  op->N = num_entires;
  op->D = out_dim;
  op->aggr = aggr;
  const int guid = m->op_guid++;
  // model.cc:157-171: "<type>_<guid>", or the shared op's name
  op->name = O(shared_op) ? O(shared_op)->name
                          : "Embed_" + std::to_string(num_entires) + "x" + std::to_string(out_dim) +
                                "_" + std::to_string(guid);
  if (kernel_initializer.impl)
    op->init = *(Initializer *)kernel_initializer.impl;
  else  // reference default: GlorotUniform(seed = rand()); keep its range on [N][D]
    op->init = Initializer{Initializer::kUniform, guid,
                           -sqrtf(6.0f / (float)(num_entires + out_dim)),
                           sqrtf(6.0f / (float)(num_entires + out_dim))};
  op->param.op = op.get();
  auto out = std::make_unique<Tensor>();
  out->model = m;
  out->producer = op.get();
  out->name = op->name + "_out";
  out->num_dims = 2;
  out->dims[0] = in->dims[0];
  out->dims[1] = out_dim;
  out->data_type = DT_FLOAT;
  out->create_grad = true;
  op->output = out.get();
  m->tensors.push_back(std::move(out));
  m->ops.push_back(std::move(op));
  return ffemb_tensor_t{m->ops.back()->output};
}

extern "C" int ffemb_model_set_sgd_optimizer(ffemb_model_t h, ffemb_sgd_optimizer_t opt) {
  if (!M(h) || !opt.impl) return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_sgd_optimizer: NULL argument");
  M(h)->opt = (Sgd *)opt.impl;
  return FFEMB_OK;
}

extern "C" int ffemb_model_set_strategy(ffemb_model_t h, ffemb_strategy_t s) {
  Model *m = M(h);
  if (!m) return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_strategy: NULL model");
  if (m->compiled) return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_strategy: model is already compiled");
  if (m->own_strategy) ffemb_strategy_destroy(m->strategy);
  m->strategy = s;
  m->own_strategy = false;
  return FFEMB_OK;
}

extern "C" int ffemb_model_set_stream(ffemb_model_t h, void *stream) {
  if (!M(h)) return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_stream: NULL model");
  M(h)->stream = (cudaStream_t)stream;
  return FFEMB_OK;
}

extern "C" int ffemb_model_compile(ffemb_model_t h) {
  Model *m = M(h);
  if (!m) return fail(FFEMB_ERR_INVALID_ARGUMENT, "compile: NULL model");
  if (m->compiled) return fail(FFEMB_ERR_INVALID_ARGUMENT, "compile: model is already compiled");
  if (!m->strategy.impl && !m->cfg.strategy_file.empty()) {
    m->strategy = ffemb_strategy_create();
    m->own_strategy = true;
    if (int rc = ffemb_strategy_load_file(m->strategy, m->cfg.strategy_file.c_str())) return rc;
  }
  // A strategy may place an embedding on a CPU (the reference's "hetero"
  // generator, src/runtime/dlrm_strategy_hetero.cc:27-36, whose CPU tasks are
  // not even registered in the reference build).  There is no CPU
  // path here: refuse it rather than run it somewhere else silently.
  for (auto &op : m->ops) {
    int i = ffemb_strategy_find(m->strategy, ("embedding" + std::to_string(op->index)).c_str());
    if (i < 0) i = ffemb_strategy_find(m->strategy, op->name.c_str());
    if (i >= 0 && ffemb_strategy_op_device_type(m->strategy, i) == FFEMB_DEVICE_CPU)
      return fail(FFEMB_ERR_UNSUPPORTED, "compile: the strategy places " + op->name +
                                             " on a CPU; this library has no CPU path");
  }
  if (m->cfg.device >= 0) {
    FFEMB_CUDA_OK(cudaSetDevice(m->cfg.device));
    m->device = m->cfg.device;
  } else {
    FFEMB_CUDA_OK(cudaGetDevice(&m->device));
  }
  // placement: table i -> rank (FFMapper::slice_task, mapper.cc:33-97)
  int64_t lookups = 0, entries = 0, sum_d = 0;
  int max_d = 1, n_owned = 0;
  for (auto &op : m->ops) {
    op->owner = ffemb_strategy_device_for_table(m->strategy, op->index, op->name.c_str(), m->world);
    if (!owned(op.get())) continue;
    ++n_owned;
    lookups += (int64_t)op->input->dims[0] * op->input->dims[1];
    entries += op->N;
    sum_d += op->D;
    if (op->D > max_d) max_d = op->D;
  }
  const int B = m->cfg.batch_size;
  // tables + initialisers (create_weights, embedding.cu:95-106; model.cc:692-696)
  for (auto &op : m->ops) {
    if (!owned(op.get())) continue;
    FFEMB_CUDA_OK(cudaMalloc(&op->W, (size_t)op->N * op->D * sizeof(float)));
    if (op->init.kind == Initializer::kZero) {
      FFEMB_CUDA_OK(cudaMemsetAsync(op->W, 0, (size_t)op->N * op->D * sizeof(float), m->stream));
    } else if (int rc = ffemb_uniform_init(op->W, op->N * op->D, op->init.seed, op->init.lo,
                                           op->init.hi, m->stream)) {
      return rc;
    }
  }
  // outputs and their gradients: one [B][sum D] matrix each, so the owned
  // tables' blocks already sit side by side as Concat would lay them out
  // (src/ops/concat.cu:159-215).
  if (n_owned) {
    m->mat_ld = (sum_d + 3) / 4 * 4;
    FFEMB_CUDA_OK(cudaMalloc(&m->out_mat, (size_t)B * m->mat_ld * sizeof(float)));
    FFEMB_CUDA_OK(cudaMalloc(&m->grad_mat, (size_t)B * m->mat_ld * sizeof(float)));
    FFEMB_CUDA_OK(cudaMemsetAsync(m->grad_mat, 0, (size_t)B * m->mat_ld * sizeof(float), m->stream));
    int64_t col = 0;
    for (auto &op : m->ops) {
      if (!owned(op.get())) continue;
      if (!op->output->data.ptr) op->output->data = Buffer{m->out_mat + col, m->mat_ld, false};
      if (!op->output->grad.ptr) op->output->grad = Buffer{m->grad_mat + col, m->mat_ld, false};
      col += op->D;
    }
    m->workspace_bytes =
        ffemb_embed_backward_workspace_bytes(lookups, entries, max_d) + ((size_t)n_owned << 20);
    FFEMB_CUDA_OK(cudaMalloc(&m->workspace, m->workspace_bytes));
  }
  FFEMB_CUDA_OK(cudaMalloc(&m->err_dev, sizeof(int)));
  FFEMB_CUDA_OK(cudaMemsetAsync(m->err_dev, 0, sizeof(int), m->stream));
  {
    int lo = 0, hi = 0;  // numerically lowest value = highest priority
    FFEMB_CUDA_OK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    FFEMB_CUDA_OK(cudaStreamCreateWithPriority(&m->side, cudaStreamNonBlocking, hi));
  }
  FFEMB_CUDA_OK(cudaEventCreateWithFlags(&m->ev_inputs, cudaEventDisableTiming));
  FFEMB_CUDA_OK(cudaEventCreateWithFlags(&m->ev_sorted, cudaEventDisableTiming));
  m->compiled = true;
  return FFEMB_OK;
}

extern "C" int ffemb_model_init_layers(ffemb_model_t h) {
  if (!M(h) || !M(h)->compiled) return fail(FFEMB_ERR_INVALID_ARGUMENT, "init_layers: model is not compiled");
  return FFEMB_OK;  // EMBED_INIT_TASK returns NULL in the reference (embedding.cu:137-144)
}

static int ready(Model *m, const char *who) {
  if (!m || !m->compiled) return fail(FFEMB_ERR_INVALID_ARGUMENT, std::string(who) + ": model is not compiled");
  FFEMB_CUDA_OK(cudaSetDevice(m->device));
  return FFEMB_OK;
}

// Workspace region of each group, in groups_of() order.
static int backward_phase(Model *m, bool sort, bool update, cudaStream_t stream) {
  std::vector<ffemb_lookup_t> lk;
  ffemb_matrix_t mat;
  size_t off = 0;
  for (const LaunchGroup &g : groups_of(m, true)) {
    if (int rc = describe(g, true, lk, mat, /*indices_only=*/!update)) return rc;
    int64_t n = 0, e = 0;
    for (const ffemb_lookup_t &l : lk) {
      n += l.num_indices;
      e += l.num_entries;
    }
    const size_t need = (ffemb_embed_backward_workspace_bytes(n, e, g.D) + 255) & ~(size_t)255;
    if (off + need > m->workspace_bytes)
      return fail(FFEMB_ERR_WORKSPACE_TOO_SMALL, "backward: model workspace too small");
    char *ws = (char *)m->workspace + off;
    off += need;
    if (sort)
      if (int rc = ffemb_embed_backward_sort(lk.data(), (int)lk.size(), m->cfg.batch_size, g.D, ws,
                                             need, m->err_dev, stream))
        return rc;
    if (update)
      if (int rc = ffemb_embed_sgd_update(lk.data(), (int)lk.size(), m->cfg.batch_size, g.D,
                                          (enum AggrMode)g.aggr, &mat, (float)m->opt->lr, ws, need,
                                          stream))
        return rc;
  }
  return FFEMB_OK;
}

// What the sorted keys in the workspace depend on: grouping and index buffers.
static std::vector<int64_t> sort_signature(Model *m) {
  std::vector<int64_t> sig;
  for (const LaunchGroup &g : groups_of(m, true)) {
    sig.push_back(-1);
    sig.push_back(g.D);
    sig.push_back(g.ld_out);
    sig.push_back(g.ld_grad);
    for (Op *op : g.ops) {
      sig.push_back(op->index);
      sig.push_back((int64_t)(uintptr_t)op->input->data.ptr);
    }
  }
  return sig;
}

extern "C" int ffemb_model_forward(ffemb_model_t h) {
  Model *m = M(h);
  if (int rc = ready(m, "forward")) return rc;
  // Training: fork the key sort of the coming backward onto the side stream.
  // It reads only the index tensors, which are complete at this point of the
  // model stream, and it is small and latency-bound, so it hides under the
  // gather.  backward() joins it.
  m->presorted = false;
  // The sort goes first, on a high-priority stream: its small grids then get
  // SM slots as they free up instead of queueing behind the gather's
  // thousands of blocks.
  const bool presort = m->presort && m->opt;
  if (presort) {
    FFEMB_CUDA_OK(cudaEventRecord(m->ev_inputs, m->stream));
    FFEMB_CUDA_OK(cudaStreamWaitEvent(m->side, m->ev_inputs, 0));
    if (int rc = backward_phase(m, true, false, m->side)) return rc;
    FFEMB_CUDA_OK(cudaEventRecord(m->ev_sorted, m->side));
  }
  std::vector<ffemb_lookup_t> lk;
  ffemb_matrix_t mat;
  for (const LaunchGroup &g : groups_of(m, false)) {
    if (int rc = describe(g, false, lk, mat)) return rc;
    if (int rc = ffemb_embed_forward(lk.data(), (int)lk.size(), m->cfg.batch_size, g.D,
                                     (enum AggrMode)g.aggr, &mat, m->err_dev, m->stream))
      return rc;
  }
  if (presort) {
    // join right away in stream order: later work on the model stream (and a
    // stream capture) sees one timeline; the kernels still overlap the gather
    FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, m->ev_sorted, 0));
    m->presort_sig = sort_signature(m);
    m->presorted = true;
  }
  return FFEMB_OK;
}

extern "C" int ffemb_model_zero_gradients(ffemb_model_t h) {
  Model *m = M(h);
  if (int rc = ready(m, "zero_gradients")) return rc;
  m->sorted = false;  // the (sparse) table gradient is now empty
  return FFEMB_OK;
}

extern "C" int ffemb_model_backward(ffemb_model_t h) {
  Model *m = M(h);
  if (int rc = ready(m, "backward")) return rc;
  if (m->presorted && m->presort_sig == sort_signature(m)) {
    m->presorted = false;  // forward() already sorted these indices
  } else {
    if (int rc = backward_phase(m, true, false, m->stream)) return rc;
  }
  m->sorted = true;
  return FFEMB_OK;
}

extern "C" int ffemb_model_update(ffemb_model_t h) {
  Model *m = M(h);
  if (int rc = ready(m, "update")) return rc;
  if (!m->opt) return fail(FFEMB_ERR_INVALID_ARGUMENT, "update: no optimizer set");
  if (m->opt->momentum != 0.0 || m->opt->weight_decay != 0.0)
    return fail(FFEMB_ERR_UNSUPPORTED,
                "update: the sparse fused SGD equals the reference's dense sgd_update only for "
                "momentum == 0 and weight_decay == 0");
  if (!m->sorted) return FFEMB_OK;  // zero gradient: W -= lr*0
  if (int rc = backward_phase(m, false, true, m->stream)) return rc;
  m->sorted = false;
  return FFEMB_OK;
}

extern "C" int ffemb_model_sync(ffemb_model_t h) {
  Model *m = M(h);
  if (int rc = ready(m, "sync")) return rc;
  FFEMB_CUDA_OK(cudaStreamSynchronize(m->stream));
  int e = 0;
  FFEMB_CUDA_OK(cudaMemcpy(&e, m->err_dev, sizeof e, cudaMemcpyDeviceToHost));
  if (e) {
    FFEMB_CUDA_OK(cudaMemset(m->err_dev, 0, sizeof e));
    return fail(e, "an embedding index was outside [0, num_entries)");
  }
  return FFEMB_OK;
}

extern "C" int ffemb_model_get_num_layers(ffemb_model_t h) { return M(h) ? (int)M(h)->ops.size() : -1; }
extern "C" ffemb_op_t ffemb_model_get_layer_by_id(ffemb_model_t h, int id) {
  if (!M(h) || id < 0 || id >= (int)M(h)->ops.size()) return ffemb_op_t{nullptr};
  return ffemb_op_t{M(h)->ops[id].get()};
}

// ---- Op ----------------------------------------------------------------------
extern "C" ffemb_parameter_t ffemb_op_get_parameter_by_id(ffemb_op_t h, int id) {
  if (!O(h) || id != 0) return ffemb_parameter_t{nullptr};  // numWeights = 1 (embedding.cu:66)
  return ffemb_parameter_t{&O(h)->param};
}
extern "C" ffemb_tensor_t ffemb_op_get_input_by_id(ffemb_op_t h, int id) {
  return ffemb_tensor_t{O(h) && id == 0 ? O(h)->input : nullptr};
}
extern "C" ffemb_tensor_t ffemb_op_get_output_by_id(ffemb_op_t h, int id) {
  return ffemb_tensor_t{O(h) && id == 0 ? O(h)->output : nullptr};
}
extern "C" const char *ffemb_op_get_name(ffemb_op_t h) { return O(h) ? O(h)->name.c_str() : nullptr; }
extern "C" int ffemb_op_get_device(ffemb_op_t h) { return O(h) ? O(h)->owner : -1; }

// ---- Tensor data -------------------------------------------------------------
static Op *consumer_of(Tensor *t) {
  for (auto &op : t->model->ops)
    if (op->input == t) return op.get();
  return nullptr;
}

extern "C" int ffemb_tensor_set_int64(ffemb_tensor_t h, ffemb_model_t model, const int64_t *data,
                                      int64_t count) {
  Tensor *t = T(h);
  Model *m = M(model);
  if (!t || !m || t->model != m || !data)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_set_int64: bad arguments");
  if (int rc = ready(m, "tensor_set_int64")) return rc;
  if (t->data_type != DT_INT64 || t->producer)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_set_int64: not an int64 input tensor");
  const int64_t n = (int64_t)t->dims[0] * t->dims[1];
  if (count != n)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_set_int64: count != batch * bag_size");
  // indices are only needed where the consuming table lives (the reference's
  // loader also loads table t's indices on table t's device, dlrm.cc:491-524)
  bool needed = false;
  for (auto &op : m->ops) needed = needed || (op->input == t && owned(op.get()));
  if (!needed) return FFEMB_OK;
  if (!t->data.ptr) {
    void *p = nullptr;
    FFEMB_CUDA_OK(cudaMalloc(&p, n * sizeof(int64_t)));
    t->data = Buffer{p, t->dims[1], true};
  }
  FFEMB_CUDA_OK(cudaMemcpyAsync(t->data.ptr, data, n * sizeof(int64_t), cudaMemcpyHostToDevice,
                                m->stream));
  m->presorted = false;  // keys sorted during forward() belong to the old contents
  return FFEMB_OK;
}

extern "C" int ffemb_tensor_set_grad_float(ffemb_tensor_t h, ffemb_model_t model, const float *data,
                                           int64_t count) {
  Tensor *t = T(h);
  Model *m = M(model);
  if (!t || !m || t->model != m || !data)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_set_grad_float: bad arguments");
  if (int rc = ready(m, "tensor_set_grad_float")) return rc;
  if (!t->producer) return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_set_grad_float: not an embedding output");
  if (count != (int64_t)t->dims[0] * t->dims[1])
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_set_grad_float: count != batch * out_dim");
  if (!owned(t->producer)) return FFEMB_OK;
  FFEMB_CUDA_OK(cudaMemcpy2DAsync(t->grad.ptr, t->grad.ld * sizeof(float), data,
                                  t->dims[1] * sizeof(float), t->dims[1] * sizeof(float), t->dims[0],
                                  cudaMemcpyHostToDevice, m->stream));
  return FFEMB_OK;
}

extern "C" int ffemb_tensor_get_float(ffemb_tensor_t h, ffemb_model_t model, float *data,
                                      int64_t count) {
  Tensor *t = T(h);
  Model *m = M(model);
  if (!t || !m || t->model != m || !data)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_get_float: bad arguments");
  if (int rc = ready(m, "tensor_get_float")) return rc;
  if (!t->producer || !owned(t->producer))
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_get_float: not an output of a table this rank owns");
  if (count != (int64_t)t->dims[0] * t->dims[1])
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_get_float: count != batch * out_dim");
  FFEMB_CUDA_OK(cudaMemcpy2DAsync(data, t->dims[1] * sizeof(float), t->data.ptr,
                                  t->data.ld * sizeof(float), t->dims[1] * sizeof(float), t->dims[0],
                                  cudaMemcpyDeviceToHost, m->stream));
  FFEMB_CUDA_OK(cudaStreamSynchronize(m->stream));
  return FFEMB_OK;
}

extern "C" int ffemb_tensor_attach_device_ptr(ffemb_tensor_t h, void *device_ptr, int64_t ld,
                                              bool grad) {
  Tensor *t = T(h);
  if (!t || !device_ptr) return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_attach_device_ptr: NULL argument");
  if (ld < t->dims[t->num_dims - 1])
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_attach_device_ptr: ld smaller than the row length");
  if (grad && !t->producer)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_attach_device_ptr: index tensors have no gradient");
  if (!t->producer && ld != t->dims[t->num_dims - 1])
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_attach_device_ptr: index tensors must be contiguous (ld == bag size)");
  Buffer &b = grad ? t->grad : t->data;
  if (b.owned && b.ptr) {
    if (t->model->compiled) cudaSetDevice(t->model->device);
    cudaFree(b.ptr);
  }
  b = Buffer{device_ptr, ld, false};
  if (!t->producer) t->model->presorted = false;  // an index tensor changed under a finished sort
  return FFEMB_OK;
}
extern "C" void *ffemb_tensor_get_device_ptr(ffemb_tensor_t h, bool grad) {
  return T(h) ? (grad ? T(h)->grad.ptr : T(h)->data.ptr) : nullptr;
}
extern "C" int64_t ffemb_tensor_get_ld(ffemb_tensor_t h, bool grad) {
  return T(h) ? (grad ? T(h)->grad.ld : T(h)->data.ld) : -1;
}

// ---- Parameter ---------------------------------------------------------------
extern "C" bool ffemb_parameter_set_weights_float(ffemb_parameter_t h, ffemb_model_t model,
                                                  int num_dim, const int *dims, const float *data) {
  Parameter *p = P(h);
  Model *m = M(model);
  if (!p || !m || p->op->model != m || !data || !dims) {
    set_error("parameter_set_weights_float: bad arguments");
    return false;
  }
  if (ready(m, "parameter_set_weights_float")) return false;
  Op *op = p->op;
  if (!owned(op)) {
    set_error("parameter_set_weights_float: table " + op->name + " is owned by rank " +
              std::to_string(op->owner));
    return false;
  }
  // model.cu:277-290 checks the volume against the region; so do we
  int64_t vol = 1;
  for (int i = 0; i < num_dim; ++i) vol *= dims[i];
  if (vol != op->N * op->D) {
    set_error("parameter_set_weights_float: dims do not match num_entries * out_dim");
    return false;
  }
  if (cudaMemcpyAsync(op->W, data, vol * sizeof(float), cudaMemcpyHostToDevice, m->stream) != cudaSuccess ||
      cudaStreamSynchronize(m->stream) != cudaSuccess) {
    set_error("parameter_set_weights_float: copy failed");
    return false;
  }
  return true;
}

extern "C" bool ffemb_parameter_get_weights_float(ffemb_parameter_t h, ffemb_model_t model,
                                                  float *data) {
  Parameter *p = P(h);
  Model *m = M(model);
  if (!p || !m || p->op->model != m || !data) {
    set_error("parameter_get_weights_float: bad arguments");
    return false;
  }
  if (ready(m, "parameter_get_weights_float")) return false;
  Op *op = p->op;
  if (!owned(op)) {
    set_error("parameter_get_weights_float: table " + op->name + " is owned by rank " +
              std::to_string(op->owner));
    return false;
  }
  if (cudaMemcpyAsync(data, op->W, (size_t)op->N * op->D * sizeof(float), cudaMemcpyDeviceToHost,
                      m->stream) != cudaSuccess ||
      cudaStreamSynchronize(m->stream) != cudaSuccess) {
    set_error("parameter_get_weights_float: copy failed");
    return false;
  }
  return true;
}
extern "C" float *ffemb_parameter_get_device_ptr(ffemb_parameter_t h) { return P(h) ? P(h)->op->W : nullptr; }
extern "C" int64_t ffemb_parameter_get_num_entries(ffemb_parameter_t h) { return P(h) ? P(h)->op->N : -1; }
extern "C" int ffemb_parameter_get_out_dim(ffemb_parameter_t h) { return P(h) ? P(h)->op->D : -1; }

// ---- SGDOptimizer / initialisers ---------------------------------------------
extern "C" ffemb_sgd_optimizer_t ffemb_sgd_optimizer_create(ffemb_model_t, double lr, double momentum,
                                                            bool nesterov, double weight_decay) {
  return ffemb_sgd_optimizer_t{new Sgd{lr, momentum, nesterov, weight_decay}};
}
extern "C" void ffemb_sgd_optimizer_destroy(ffemb_sgd_optimizer_t h) { delete (Sgd *)h.impl; }
extern "C" void ffemb_sgd_optimizer_set_lr(ffemb_sgd_optimizer_t h, double lr) {
  if (h.impl) ((Sgd *)h.impl)->lr = lr;
}
extern "C" ffemb_uniform_initializer_t ffemb_uniform_initializer_create(int seed, float min, float max) {
  return ffemb_uniform_initializer_t{new Initializer{Initializer::kUniform, seed, min, max}};
}
extern "C" void ffemb_uniform_initializer_destroy(ffemb_uniform_initializer_t h) {
  delete (Initializer *)h.impl;
}
extern "C" ffemb_zero_initializer_t ffemb_zero_initializer_create(void) {
  return ffemb_zero_initializer_t{new Initializer{Initializer::kZero, 0, 0.f, 0.f}};
}
extern "C" void ffemb_zero_initializer_destroy(ffemb_zero_initializer_t h) {
  delete (Initializer *)h.impl;
}
