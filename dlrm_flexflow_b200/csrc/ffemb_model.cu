// ffemb_model.cu -- model-level C-ABI (the flexflow_c.h surface for this path).
//
// Mirrors, for the embedding path only, the handle objects of the reference's
// python/flexflow_c.cc: FFConfig, FFModel, Tensor, Op, Parameter,
// SGDOptimizer, Uniform/Zero initialisers.  Everything Legion did for this
// path (regions, partitions, mapping) collapses to: which rank owns which
// table (the placement hook), device buffers, one stream.
#include <stdlib.h>
#include <string.h>

#include <algorithm>
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
  int64_t replicate_below = 0;  // world > 1: tables with fewer rows live on every rank; -1 = batch / world
  int force_exchange = 0;       // run the sharded engine even with one rank (tests)
  int grad_exchange = 0;        // 0 auto (pull for 2 ranks, push from 3 up), 1 push, 2 pull
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
  int owner = 0;            // rank that holds the table; -1: replicated on every rank
  bool replicated = false;
  float *G = nullptr;       // replicated: this rank's dense gradient replica
  int64_t gcol = 0;         // sharded mode: first column in the [B/N][sum D] matrices
  int64_t scol = 0;         // sharded mode: first column in the owner's staging matrix
  Parameter param;
};

struct Sgd {
  double lr, momentum;
  bool nesterov;
  double weight_decay;
};

struct Strategy;  // ffemb_strategy.cc

// Table-wise sharded mode (world > 1, one process per GPU): this rank's side
// of the two exchanges.  One IPC-exportable allocation per rank,
//   [ out B/N x ldx | 64 barrier words | grad B/N x ldx | replica W | replica G | staging B x stage_ld ]
// whose first five parts have the same offsets on every rank.  Stands in for
// what Legion DMA does for the reference at Concat (src/ops/concat.cu:85-93,
// outputs in zero-copy memory: src/ops/embedding.cu:280-284,343-347).
struct Exchange {
  bool on = false, connected = false;
  int Bs = 0;                 // samples per rank
  int64_t ldx = 0;            // row stride of the out / grad matrices (all tables side by side)
  std::vector<int64_t> stage_ld;  // per rank: row stride of its staging matrix
  char *buf = nullptr;
  size_t bytes = 0;
  size_t off_out = 0, off_flags = 0, off_grad = 0, off_repW = 0, off_repG = 0, off_stage = 0;
  char *peer[kMaxParts] = {};
  int64_t rep_count = 0;      // floats in the flat replica arrays (multiple of 4)
  int n_rep = 0, n_owned = 0;
  void *rep_ws = nullptr;
  size_t rep_ws_bytes = 0;
  cudaStream_t rep_stream = nullptr;
  cudaEvent_t ev_rep_fwd = nullptr, ev_rep_sorted = nullptr, ev_grads = nullptr, ev_rep_done = nullptr;
  bool rep_pending = false;   // backward() left a replica gradient for update() to reduce
  bool pull = false;          // gradient exchange: the update kernel loads rows from the peers' matrices
  bool sort_in_forward = false;  // sharded tables' key sort: under the gather (true) or under the push
  cudaStream_t rep_side = nullptr;  // high priority: the replicas' key sort
  // development aid: timed events at the phase boundaries of a step (all streams)
  bool trace = false;
  std::vector<std::pair<const char *, cudaEvent_t>> marks;
};

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
  Exchange x;
  struct HostPipe *pipe = nullptr;  // double-buffered host ingest / egress (below)
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

static void x_destroy(Model *m);     // sharded engine, below
static void pipe_destroy(Model *m);  // host pipeline, below

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
extern "C" int ffemb_config_set_replicate_below(ffemb_config_t h, int64_t rows) {
  if (!C(h) || rows < -1) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_replicate_below: bad argument");
  C(h)->replicate_below = rows;
  return FFEMB_OK;
}
extern "C" int ffemb_config_set_gradient_exchange(ffemb_config_t h, int mode) {
  if (!C(h) || mode < 0 || mode > 2) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_gradient_exchange: 0 auto, 1 push, 2 pull");
  C(h)->grad_exchange = mode;
  return FFEMB_OK;
}
extern "C" int ffemb_config_set_exchange_engine(ffemb_config_t h, int force) {
  if (!C(h)) return fail(FFEMB_ERR_INVALID_ARGUMENT, "config_set_exchange_engine: NULL config");
  C(h)->force_exchange = force != 0;
  return FFEMB_OK;
}
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
  pipe_destroy(m);
  x_destroy(m);
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

extern "C" int ffemb_model_set_presort(ffemb_model_t h, bool enable) {
  if (!M(h)) return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_presort: NULL model");
  M(h)->presort = enable;
  return FFEMB_OK;
}

extern "C" int ffemb_model_set_stream(ffemb_model_t h, void *stream) {
  if (!M(h)) return fail(FFEMB_ERR_INVALID_ARGUMENT, "set_stream: NULL model");
  M(h)->stream = (cudaStream_t)stream;
  return FFEMB_OK;
}

// ===========================================================================
// Sharded engine (Exchange): forward / backward / update with world > 1.
// ===========================================================================
namespace ffemb {

static size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

static void x_mark(Model *m, const char *name, cudaStream_t s) {
  if (!m->x.trace || m->x.marks.size() >= 20000) return;
  cudaEvent_t e;
  if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, s);
  m->x.marks.emplace_back(name, e);
}

// Launch groups of the sharded mode: this rank's sharded tables, or the
// replicated ones, by (out_dim, aggr).
struct XGroup {
  int D, aggr;
  std::vector<Op *> ops;
};

static std::vector<XGroup> x_groups(Model *m, bool rep) {
  std::vector<XGroup> gs;
  for (auto &up : m->ops) {
    Op *op = up.get();
    if (rep ? !op->replicated : (op->replicated || !owned(op))) continue;
    XGroup *g = nullptr;
    for (auto &c : gs)
      if (c.D == op->D && c.aggr == op->aggr && (int)c.ops.size() < kMaxTables) g = &c;
    if (!g) {
      gs.push_back(XGroup{op->D, op->aggr, {}});
      g = &gs.back();
    }
    g->ops.push_back(op);
  }
  return gs;
}

enum XMode { kXFwd, kXBwd, kXRepFwd, kXRepBwd };

// lookups of one group.  Sharded tables see the whole batch; replicated ones
// this rank's own samples, rows [rank*Bs, (rank+1)*Bs) of the index tensor.
static int x_lookups(Model *m, const XGroup &g, XMode mode, std::vector<ffemb_lookup_t> &lk) {
  const Exchange &x = m->x;
  lk.clear();
  for (Op *op : g.ops) {
    if (!op->input->data.ptr)
      return fail(FFEMB_ERR_INVALID_ARGUMENT, op->name + ": input tensor has no data");
    const int L = op->input->dims[1];
    ffemb_lookup_t l;
    memset(&l, 0, sizeof l);
    l.num_entries = op->N;
    l.bag_size = L;
    l.offsets = nullptr;
    const int64_t *idx = (const int64_t *)op->input->data.ptr;
    if (mode == kXFwd || mode == kXBwd) {
      l.weight = op->W;
      l.indices = idx;
      l.num_indices = (int64_t)m->cfg.batch_size * L;
      l.col = (mode == kXFwd || x.pull) ? op->gcol : op->scol;
    } else {
      l.weight = mode == kXRepFwd ? op->W : op->G;
      l.indices = idx + (int64_t)m->cfg.rank * x.Bs * L;
      l.num_indices = (int64_t)x.Bs * L;
      l.col = op->gcol;
    }
    lk.push_back(l);
  }
  return FFEMB_OK;
}

static ffemb_matrix_t x_matrix_out_peers(const Model *m) {  // forward: every rank's [Bs][ldx] out
  ffemb_matrix_t mat;
  memset(&mat, 0, sizeof mat);
  for (int q = 0; q < m->world; ++q) mat.base[q] = (float *)(m->x.peer[q] + m->x.off_out);
  mat.num_parts = m->world;
  mat.rows_per_part = m->x.Bs;
  mat.ld = m->x.ldx;
  return mat;
}

static ffemb_matrix_t x_matrix_grad_peers(const Model *m) {  // pull: every rank's [Bs][ldx] gradient
  ffemb_matrix_t mat;
  memset(&mat, 0, sizeof mat);
  for (int q = 0; q < m->world; ++q) mat.base[q] = (float *)(m->x.peer[q] + m->x.off_grad);
  mat.num_parts = m->world;
  mat.rows_per_part = m->x.Bs;
  mat.ld = m->x.ldx;
  return mat;
}

static ffemb_matrix_t x_matrix_local(const Model *m, size_t off, int rows, int64_t ld) {
  ffemb_matrix_t mat;
  memset(&mat, 0, sizeof mat);
  mat.base[0] = (float *)(m->x.buf + off);
  mat.num_parts = 1;
  mat.rows_per_part = rows;
  mat.ld = ld;
  return mat;
}

static int x_barrier(Model *m, int channel, cudaStream_t s) {
  if (m->world <= 1) return FFEMB_OK;
  uint32_t *flags[kMaxParts];
  for (int q = 0; q < m->world; ++q) flags[q] = (uint32_t *)(m->x.peer[q] + m->x.off_flags) + 32 * channel;
  return ffemb_peer_barrier(flags, m->cfg.rank, m->world, m->err_dev, s);
}

// sort and/or update over the groups of one kind; every group has its own
// region of the workspace (sort results must survive until the update).
static int x_backward_groups(Model *m, bool rep, bool sort, bool update, float lr, cudaStream_t s) {
  const Exchange &x = m->x;
  char *ws = (char *)(rep ? x.rep_ws : m->workspace);
  const size_t ws_bytes = rep ? x.rep_ws_bytes : m->workspace_bytes;
  const int batch = rep ? x.Bs : m->cfg.batch_size;
  std::vector<ffemb_lookup_t> lk;
  size_t off = 0;
  for (const XGroup &g : x_groups(m, rep)) {
    if (int rc = x_lookups(m, g, rep ? kXRepBwd : kXBwd, lk)) return rc;
    int64_t n = 0, e = 0;
    for (const ffemb_lookup_t &l : lk) {
      n += l.num_indices;
      e += l.num_entries;
    }
    const size_t need = up256(ffemb_embed_backward_workspace_bytes(n, e, g.D));
    if (off + need > ws_bytes) return fail(FFEMB_ERR_WORKSPACE_TOO_SMALL, "backward: model workspace too small");
    if (sort)
      if (int rc = ffemb_embed_backward_sort(lk.data(), (int)lk.size(), batch, g.D, ws + off, need, m->err_dev, s))
        return rc;
    if (update) {
      // replicas: dense gradient of this rank's own samples, from its [Bs][ldx]
      // gradient matrix; sharded: the whole batch's gradient in the staging matrix
      ffemb_matrix_t mat = rep      ? x_matrix_local(m, x.off_grad, x.Bs, x.ldx)
                           : x.pull ? x_matrix_grad_peers(m)
                                    : x_matrix_local(m, x.off_stage, batch, x.stage_ld[m->cfg.rank]);
      if (int rc = ffemb_embed_sgd_update(lk.data(), (int)lk.size(), batch, g.D, (enum AggrMode)g.aggr, &mat, lr,
                                          ws + off, need, s))
        return rc;
    }
    off += need;
  }
  return FFEMB_OK;
}

static std::vector<int64_t> x_signature(Model *m) {
  std::vector<int64_t> sig;
  for (auto &op : m->ops)
    if (op->replicated || owned(op.get())) {
      sig.push_back(op->index);
      sig.push_back((int64_t)(uintptr_t)op->input->data.ptr);
    }
  return sig;
}

static int x_require_connected(Model *m, const char *who) {
  if (!m->x.connected)
    return fail(FFEMB_ERR_INVALID_ARGUMENT,
                std::string(who) + ": the ranks' exchange buffers are not connected yet "
                                   "(ffemb_model_get_exchange_handle / ffemb_model_connect_peers)");
  return FFEMB_OK;
}

static int x_forward(Model *m) {
  Exchange &x = m->x;
  if (int rc = x_require_connected(m, "forward")) return rc;
  m->presorted = false;
  const bool presort = m->presort && m->opt;
  std::vector<ffemb_lookup_t> lk;
  x_mark(m, "fwd_begin", m->stream);
  FFEMB_CUDA_OK(cudaEventRecord(m->ev_inputs, m->stream));
  // The key sorts of the coming backward need only the indices: they go first,
  // on their own streams, and run under the NVLink-bound gather.
  // (Push mode sorts the sharded tables' keys in backward() instead, under the
  // NVLink-bound gradient push: beside the gather the latency-bound sort ran
  // 2-3x slower than alone and ended up as the forward's critical path.)
  const bool presort_owned = presort && x.n_owned && x.sort_in_forward;
  if (presort_owned) {
    FFEMB_CUDA_OK(cudaStreamWaitEvent(m->side, m->ev_inputs, 0));
    if (int rc = x_backward_groups(m, false, true, false, 0.f, m->side)) return rc;
    FFEMB_CUDA_OK(cudaEventRecord(m->ev_sorted, m->side));
    x_mark(m, "  (side) sharded tables' keys sorted", m->side);
  }
  if (x.n_rep) {
    if (presort) {
      FFEMB_CUDA_OK(cudaStreamWaitEvent(x.rep_side, m->ev_inputs, 0));
      if (int rc = x_backward_groups(m, true, true, false, 0.f, x.rep_side)) return rc;
      FFEMB_CUDA_OK(cudaEventRecord(x.ev_rep_sorted, x.rep_side));
      x_mark(m, "  (rep side) replicas' keys sorted", x.rep_side);
    }
    // replicated tables: this rank's own samples, straight into its own slice;
    // local traffic only, so it runs beside the sharded tables' gather
    FFEMB_CUDA_OK(cudaStreamWaitEvent(x.rep_stream, m->ev_inputs, 0));
    ffemb_matrix_t own = x_matrix_local(m, x.off_out, x.Bs, x.ldx);
    for (const XGroup &g : x_groups(m, true)) {
      if (int rc = x_lookups(m, g, kXRepFwd, lk)) return rc;
      if (int rc = ffemb_embed_forward(lk.data(), (int)lk.size(), x.Bs, g.D, (enum AggrMode)g.aggr, &own,
                                       m->err_dev, x.rep_stream))
        return rc;
    }
    FFEMB_CUDA_OK(cudaEventRecord(x.ev_rep_fwd, x.rep_stream));
    x_mark(m, "  (rep) replicas gathered", x.rep_stream);
  }
  // sharded tables: the whole batch, every pooled row stored straight into the
  // consuming rank's [Bs][ldx] matrix over NVLink
  if (x.n_owned) {
    ffemb_matrix_t out = x_matrix_out_peers(m);
    for (const XGroup &g : x_groups(m, false)) {
      if (int rc = x_lookups(m, g, kXFwd, lk)) return rc;
      if (int rc = ffemb_embed_forward(lk.data(), (int)lk.size(), m->cfg.batch_size, g.D, (enum AggrMode)g.aggr,
                                       &out, m->err_dev, m->stream))
        return rc;
    }
  }
  x_mark(m, "sharded gather + remote stores done", m->stream);
  if (int rc = x_barrier(m, 0, m->stream)) return rc;  // every rank's rows have landed
  x_mark(m, "forward barrier passed", m->stream);
  // join the forks: one timeline for the caller (and for a stream capture)
  if (x.n_rep) {
    FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, x.ev_rep_fwd, 0));
    if (presort) FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, x.ev_rep_sorted, 0));
  }
  if (presort_owned) FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, m->ev_sorted, 0));
  if (presort) {
    m->presort_sig = x_signature(m);
    m->presorted = true;
  }
  x_mark(m, "forward end (forks joined)", m->stream);
  return FFEMB_OK;
}

static int x_backward(Model *m) {
  Exchange &x = m->x;
  if (int rc = x_require_connected(m, "backward")) return rc;
  const bool presorted = m->presorted && m->presort_sig == x_signature(m);
  m->presorted = false;
  x_mark(m, "bwd_begin", m->stream);
  FFEMB_CUDA_OK(cudaEventRecord(x.ev_grads, m->stream));  // this rank's gradient matrix is complete
  if (x.n_rep) {
    // replicas (the reference's data-parallel embedding): dense gradient of
    // this rank's own samples -- the segment-reduce kernel with lr = -1 into a
    // zeroed G; update() sums the replicas and applies the step
    FFEMB_CUDA_OK(cudaStreamWaitEvent(x.rep_stream, x.ev_grads, 0));
    // (a sort started by forward() was joined into the model stream there, so
    // ev_grads already orders this branch after it)
    FFEMB_CUDA_OK(cudaMemsetAsync(x.buf + x.off_repG, 0, (size_t)x.rep_count * 4, x.rep_stream));
    if (int rc = x_backward_groups(m, true, !presorted, true, -1.0f, x.rep_stream)) return rc;
    x_mark(m, "  (rep) own samples reduced into the dense replica gradient", x.rep_stream);
    x.rep_pending = true;
  }
  const bool fork_sort = x.n_owned && !x.sort_in_forward && m->presort;
  if (fork_sort) {  // the sharded tables' key sort runs under the gradient push
    FFEMB_CUDA_OK(cudaStreamWaitEvent(m->side, x.ev_grads, 0));
    if (int rc = x_backward_groups(m, false, true, false, 0.f, m->side)) return rc;
    FFEMB_CUDA_OK(cudaEventRecord(m->ev_sorted, m->side));
    x_mark(m, "  (side) sharded tables' keys sorted", m->side);
  } else if (x.n_owned && !(presorted && x.sort_in_forward)) {
    if (int rc = x_backward_groups(m, false, true, false, 0.f, m->stream)) return rc;
  }
  // gradient exchange.  Push: every sharded table's column block of this
  // rank's gradient goes to its owner's staging matrix, rows [rank*Bs, ..), and
  // the update then reads local memory.  Pull: nothing moves here; after the
  // barrier the update kernel loads gradient rows from the peers' matrices.
  if (!x.pull) {
    std::vector<int> dims;
    for (auto &op : m->ops)
      if (!op->replicated && std::find(dims.begin(), dims.end(), op->D) == dims.end()) dims.push_back(op->D);
    for (int D : dims) {
      std::vector<ffemb_push_t> items;
      for (auto &op : m->ops) {
        if (op->replicated || op->D != D) continue;
        const int q = op->owner;
        ffemb_push_t it;
        it.dst = (float *)(x.peer[q] + x.off_stage) + (int64_t)m->cfg.rank * x.Bs * x.stage_ld[q] + op->scol;
        it.dst_ld = x.stage_ld[q];
        it.src_col = op->gcol;
        items.push_back(it);
      }
      if (int rc = ffemb_push_columns(items.data(), (int)items.size(), x.Bs, D,
                                      (const float *)(x.buf + x.off_grad), x.ldx, m->stream))
        return rc;
    }
  }
  x_mark(m, "gradient push done", m->stream);
  if (int rc = x_barrier(m, 0, m->stream)) return rc;  // the blocks have landed in the owners' staging
  x_mark(m, "backward barrier passed", m->stream);
  if (fork_sort) FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, m->ev_sorted, 0));  // join: one timeline
  return FFEMB_OK;
}

static int x_update(Model *m) {
  Exchange &x = m->x;
  const float lr = (float)m->opt->lr;
  if (x.n_owned)
    if (int rc = x_backward_groups(m, false, false, true, lr, m->stream)) return rc;
  x_mark(m, "sharded tables updated", m->stream);
  // pull: the peers have read this rank's gradient matrix inside their update;
  // nobody may overwrite it (the next batch's gradients) before they are done
  if (x.pull)
    if (int rc = x_barrier(m, 0, m->stream)) return rc;
  if (x.rep_pending) {
    float *W[kMaxParts];
    const float *G[kMaxParts];
    for (int q = 0; q < m->world; ++q) {
      W[q] = (float *)(x.peer[q] + x.off_repW);
      G[q] = (const float *)(x.peer[q] + x.off_repG);
    }
    if (int rc = x_barrier(m, 1, x.rep_stream)) return rc;  // every rank's replica gradient is complete
    x_mark(m, "  (rep) replica barrier 1 passed", x.rep_stream);
    if (int rc = ffemb_replica_reduce_update(W, G, x.rep_count, m->cfg.rank, m->world, lr, x.rep_stream)) return rc;
    x_mark(m, "  (rep) replicas summed, stepped, written to every rank", x.rep_stream);
    if (int rc = x_barrier(m, 1, x.rep_stream)) return rc;  // every copy of the tables is written
    x_mark(m, "  (rep) replica barrier 2 passed", x.rep_stream);
    FFEMB_CUDA_OK(cudaEventRecord(x.ev_rep_done, x.rep_stream));
    FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, x.ev_rep_done, 0));
    x.rep_pending = false;
  }
  x_mark(m, "step end", m->stream);
  return FFEMB_OK;
}

// compile() for the sharded mode: placement, the exchange buffer, tables.
static int x_compile(Model *m) {
  Exchange &x = m->x;
  const int B = m->cfg.batch_size, N = m->world, rank = m->cfg.rank;
  if (N > kMaxParts) return fail(FFEMB_ERR_UNSUPPORTED, "compile: at most " + std::to_string(kMaxParts) + " ranks");
  if (B % N) return fail(FFEMB_ERR_INVALID_ARGUMENT, "compile: the batch size must be divisible by the number of ranks");
  x.on = true;
  x.Bs = B / N;
  int64_t below = m->cfg.replicate_below;
  if (below < 0) below = x.Bs;  // a replica's all-reduce is cheaper than two exchanges iff N_t < B/N
  if (N == 1 && !m->cfg.force_exchange) below = 0;
  // placement: a strategy names the owners (FFMapper::slice_task); without one
  // the sharded tables go round-robin, in table order, over the ranks
  bool any_rep = false;
  for (auto &op : m->ops) {
    op->replicated = op->N < below;
    any_rep = any_rep || op->replicated;
  }
  int k = 0;
  for (auto &op : m->ops) {
    if (op->replicated) {
      op->owner = -1;
      continue;
    }
    if (m->strategy.impl || !any_rep)
      op->owner = ffemb_strategy_device_for_table(m->strategy, op->index, op->name.c_str(), N);
    else
      op->owner = k % N;
    ++k;
  }
  // column layouts
  x.stage_ld.assign(N, 0);
  int64_t col = 0, rep_floats = 0;
  std::vector<int64_t> rep_off(m->ops.size(), 0);
  for (auto &op : m->ops) {
    op->gcol = col;
    col += op->D;
    if (op->replicated) {
      rep_off[op->index] = rep_floats;
      rep_floats += (op->N * op->D + 3) / 4 * 4;
      ++x.n_rep;
    } else {
      op->scol = x.stage_ld[op->owner];
      x.stage_ld[op->owner] += op->D;
      if (op->owner == rank) ++x.n_owned;
    }
  }
  x.ldx = (col + 3) / 4 * 4;
  for (int q = 0; q < N; ++q) x.stage_ld[q] = (std::max<int64_t>(x.stage_ld[q], 4) + 3) / 4 * 4;
  x.rep_count = rep_floats;
  // the one IPC-exportable allocation
  size_t off = 0;
  auto take = [&](size_t bytes) {
    const size_t o = off;
    off += up256(bytes);
    return o;
  };
  x.off_out = take((size_t)x.Bs * x.ldx * 4);
  x.off_flags = take(256);
  x.off_grad = take((size_t)x.Bs * x.ldx * 4);
  x.off_repW = take((size_t)x.rep_count * 4);
  x.off_repG = take((size_t)x.rep_count * 4);
  // auto: pull only with two ranks AND one-hot bags.  With bags of L lookups the
  // update would load each remote gradient row L times over NVLink (C4, L = 32,
  // N = 2: update 0.90 ms pulling), whereas a push moves it once.
  int max_bag = 1;
  for (auto &op : m->ops)
    if (!op->replicated) max_bag = std::max(max_bag, op->input->dims[1]);
  x.pull = m->cfg.grad_exchange == 2 || (m->cfg.grad_exchange == 0 && N == 2 && max_bag == 1);
  x.sort_in_forward = x.pull;  // pull has no push to hide the sort under
  if (const char *e = getenv("FFEMB_SHARDED_SORT"))  // development switch: "fwd" / "bwd"
    x.sort_in_forward = !strcmp(e, "fwd") || x.pull;
  x.off_stage = take(x.n_owned && !x.pull ? (size_t)B * x.stage_ld[rank] * 4 : 0);
  x.bytes = off;
  FFEMB_CUDA_OK(cudaMalloc((void **)&x.buf, x.bytes));
  FFEMB_CUDA_OK(cudaMemsetAsync(x.buf, 0, x.bytes, m->stream));
  x.peer[rank] = x.buf;
  // tables (create_weights, embedding.cu:95-106); replicas live in the exchange
  // buffer so that the owner of a slice can write every rank's copy
  int64_t lookups = 0, entries = 0, rlookups = 0, rentries = 0;
  int max_d = 1, rmax_d = 1;
  for (auto &op : m->ops) {
    const int L = op->input->dims[1];
    if (op->replicated) {
      op->W = (float *)(x.buf + x.off_repW) + rep_off[op->index];
      op->G = (float *)(x.buf + x.off_repG) + rep_off[op->index];
      rlookups += (int64_t)x.Bs * L;
      rentries += op->N;
      rmax_d = std::max(rmax_d, op->D);
    } else if (op->owner == rank) {
      FFEMB_CUDA_OK(cudaMalloc(&op->W, (size_t)op->N * op->D * sizeof(float)));
      lookups += (int64_t)B * L;
      entries += op->N;
      max_d = std::max(max_d, op->D);
    } else {
      continue;
    }
    if (op->init.kind == Initializer::kZero) {
      FFEMB_CUDA_OK(cudaMemsetAsync(op->W, 0, (size_t)op->N * op->D * sizeof(float), m->stream));
    } else if (int rc = ffemb_uniform_init(op->W, op->N * op->D, op->init.seed, op->init.lo, op->init.hi,
                                           m->stream)) {
      return rc;
    }
  }
  // every table's output / gradient block: a column block of this rank's matrices
  for (auto &op : m->ops) {
    op->output->data = Buffer{(float *)(x.buf + x.off_out) + op->gcol, x.ldx, false};
    op->output->grad = Buffer{(float *)(x.buf + x.off_grad) + op->gcol, x.ldx, false};
  }
  if (x.n_owned) {
    m->workspace_bytes = ffemb_embed_backward_workspace_bytes(lookups, entries, max_d) + ((size_t)(x.n_owned + 1) << 20);
    FFEMB_CUDA_OK(cudaMalloc(&m->workspace, m->workspace_bytes));
  }
  if (x.n_rep) {
    x.rep_ws_bytes = ffemb_embed_backward_workspace_bytes(rlookups, rentries, rmax_d) + ((size_t)(x.n_rep + 1) << 20);
    FFEMB_CUDA_OK(cudaMalloc(&x.rep_ws, x.rep_ws_bytes));
    // The replicas' branch is a chain of small kernels, barriers and one small
    // collective: at high priority its blocks are dispatched as slots free up
    // instead of queueing behind the bulk kernels' grids.
    int plo = 0, phi = 0;
    FFEMB_CUDA_OK(cudaDeviceGetStreamPriorityRange(&plo, &phi));
    FFEMB_CUDA_OK(cudaStreamCreateWithPriority(&x.rep_stream, cudaStreamNonBlocking, phi));
    FFEMB_CUDA_OK(cudaStreamCreateWithPriority(&x.rep_side, cudaStreamNonBlocking, phi));
    FFEMB_CUDA_OK(cudaEventCreateWithFlags(&x.ev_rep_fwd, cudaEventDisableTiming));
    FFEMB_CUDA_OK(cudaEventCreateWithFlags(&x.ev_rep_sorted, cudaEventDisableTiming));
    FFEMB_CUDA_OK(cudaEventCreateWithFlags(&x.ev_rep_done, cudaEventDisableTiming));
  }
  FFEMB_CUDA_OK(cudaEventCreateWithFlags(&x.ev_grads, cudaEventDisableTiming));
  x.connected = N == 1;
  return FFEMB_OK;
}

static void x_destroy(Model *m) {
  Exchange &x = m->x;
  if (!x.on) return;
  for (int q = 0; q < m->world; ++q)
    if (q != m->cfg.rank && x.peer[q]) cudaIpcCloseMemHandle(x.peer[q]);
  for (auto &op : m->ops)
    if (op->replicated) op->W = nullptr;  // inside the exchange buffer
  if (x.buf) cudaFree(x.buf);
  if (x.rep_ws) cudaFree(x.rep_ws);
  if (x.rep_stream) cudaStreamDestroy(x.rep_stream);
  if (x.rep_side) cudaStreamDestroy(x.rep_side);
  for (cudaEvent_t e : {x.ev_rep_fwd, x.ev_rep_sorted, x.ev_grads, x.ev_rep_done})
    if (e) cudaEventDestroy(e);
}

}  // namespace ffemb

using namespace ffemb;

// ---- Sharded mode: connecting the ranks --------------------------------------
extern "C" int ffemb_model_is_sharded(ffemb_model_t h) { return M(h) && M(h)->x.on ? 1 : 0; }

// Development aid (sharded mode): timed events at the phase boundaries of the
// steps that follow; the report gives, per mark, the mean time since the
// step's forward began.  Not for use under stream capture.
extern "C" int ffemb_model_trace(ffemb_model_t h, int enable) {
  Model *m = M(h);
  if (!m || !m->x.on) return fail(FFEMB_ERR_INVALID_ARGUMENT, "trace: not a sharded model");
  m->x.trace = enable != 0;
  return FFEMB_OK;
}

extern "C" size_t ffemb_model_trace_report(ffemb_model_t h, char *buf, size_t cap) {
  Model *m = M(h);
  if (!m || !m->x.on) return 0;
  cudaSetDevice(m->device);
  cudaDeviceSynchronize();
  std::vector<std::string> names;
  std::vector<double> sum;
  std::vector<int> cnt;
  cudaEvent_t begin = nullptr;
  size_t k = 0;
  int steps = 0;
  for (auto &mk : m->x.marks) {
    if (!strcmp(mk.first, "fwd_begin")) {
      begin = mk.second;
      k = 0;
      ++steps;
    }
    if (begin) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, begin, mk.second);
      if (k >= names.size()) {
        names.push_back(mk.first);
        sum.push_back(0.0);
        cnt.push_back(0);
      }
      if (names[k] == mk.first) {
        sum[k] += ms * 1e3;
        ++cnt[k];
      }
      ++k;
    }
  }
  std::string out = "rank " + std::to_string(m->cfg.rank) + " (" + std::to_string(m->x.n_owned) + " sharded, " +
                    std::to_string(m->x.n_rep) + " replicated tables), " + std::to_string(steps) +
                    " steps, mean us since the step's forward began:\n";
  char line[256];
  for (size_t i = 0; i < names.size(); ++i) {
    snprintf(line, sizeof line, "  %8.1f  %s\n", cnt[i] ? sum[i] / cnt[i] : 0.0, names[i].c_str());
    out += line;
  }
  if (buf && cap) {  // a sizing call (buf == NULL) keeps the record
    const size_t n = out.size() < cap - 1 ? out.size() : cap - 1;
    memcpy(buf, out.data(), n);
    buf[n] = 0;
    for (auto &mk : m->x.marks) cudaEventDestroy(mk.second);
    m->x.marks.clear();
  }
  return out.size() + 1;
}

extern "C" int ffemb_model_get_exchange_handle(ffemb_model_t h, uint8_t handle[64]) {
  Model *m = M(h);
  if (!m || !m->compiled || !m->x.on || !handle)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "get_exchange_handle: not a compiled sharded model");
  FFEMB_CUDA_OK(cudaSetDevice(m->device));
  FFEMB_CUDA_OK(cudaStreamSynchronize(m->stream));  // the buffer is zeroed before any peer can see it
  return ffemb_ipc_get_handle(m->x.buf, handle);
}

extern "C" int ffemb_model_connect_peers(ffemb_model_t h, const uint8_t *handles, int count) {
  Model *m = M(h);
  if (!m || !m->compiled || !m->x.on || !handles)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "connect_peers: not a compiled sharded model");
  if (count != m->world) return fail(FFEMB_ERR_INVALID_ARGUMENT, "connect_peers: one handle per rank is needed");
  if (m->x.connected) return fail(FFEMB_ERR_INVALID_ARGUMENT, "connect_peers: already connected");
  FFEMB_CUDA_OK(cudaSetDevice(m->device));
  for (int q = 0; q < m->world; ++q) {
    if (q == m->cfg.rank) continue;
    void *p = nullptr;
    if (int rc = ffemb_ipc_open_handle(handles + (size_t)q * 64, &p)) return rc;
    m->x.peer[q] = (char *)p;
  }
  m->x.connected = true;
  return FFEMB_OK;
}

extern "C" int ffemb_model_get_exchange_matrix(ffemb_model_t h, bool grad, float **ptr, int64_t *ld, int *rows) {
  Model *m = M(h);
  if (!m || !m->compiled || !m->x.on || !ptr || !ld || !rows)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "get_exchange_matrix: not a compiled sharded model");
  *ptr = (float *)(m->x.buf + (grad ? m->x.off_grad : m->x.off_out));
  *ld = m->x.ldx;
  *rows = m->x.Bs;
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
  const bool sharded = m->world > 1 || m->cfg.force_exchange;
  if (sharded) {
    if (int rc = x_compile(m)) return rc;
  }
  // placement: table i -> rank (FFMapper::slice_task, mapper.cc:33-97)
  int64_t lookups = 0, entries = 0, sum_d = 0;
  int max_d = 1, n_owned = 0;
  for (auto &op : m->ops) {
    if (sharded) break;
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
    if (sharded || !owned(op.get())) continue;
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
  if (n_owned && !sharded) {
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
  if (m->x.on) return x_forward(m);
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
  if (m->x.on && m->sorted)
    return fail(FFEMB_ERR_UNSUPPORTED,
                "zero_gradients: in sharded mode a backward() must be followed by update() on every "
                "rank (the ranks' barriers are counted)");
  m->sorted = false;  // the (sparse) table gradient is now empty
  return FFEMB_OK;
}

extern "C" int ffemb_model_backward(ffemb_model_t h) {
  Model *m = M(h);
  if (int rc = ready(m, "backward")) return rc;
  // The reference's backward ACCUMULATES into the table gradient
  // (embedding.cu:350-352, READ_WRITE region); the sparse gradient kept here
  // describes one batch only, so a second backward() before update() /
  // zero_gradients() is refused instead of silently replacing the first.
  if (m->sorted)
    return fail(FFEMB_ERR_UNSUPPORTED,
                "backward: a gradient is already pending; gradient accumulation over several "
                "backward() calls is not supported -- call update() or zero_gradients() first");
  if (m->x.on) {
    if (int rc = x_backward(m)) return rc;
    m->sorted = true;
    return FFEMB_OK;
  }
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
  if (m->x.on) {
    if (int rc = x_update(m)) return rc;
    m->sorted = false;
    return FFEMB_OK;
  }
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

// ===========================================================================
// Host ingest / egress, double-buffered.
//
// The reference stages every table's indices per step through a freshly
// allocated pinned buffer and a synchronous cudaMemcpy (examples/cpp/DLRM/
// dlrm.cu:19-56, called from DataLoader::next_batch, dlrm.cc:486-524), and
// keeps embedding outputs and their gradients in zero-copy host memory
// (src/ops/embedding.cu:280-284,343-347).  Here the host side of a step is
// three asynchronous library calls over two slots of device buffers and two
// copy streams, so the H2D of batch i+1 and the D2H of batch i-1 run under
// the kernels of batch i (PCIe is full duplex):
//   load_batch_async      indices (+ upstream gradients) of the NEXT batch,
//                         host -> the idle slot, on the copy-in stream
//   begin_batch           the loaded slot becomes current (stream-ordered)
//   fetch_outputs_async   pooled outputs of the current batch -> host, on the
//                         copy-out stream, after forward()
// ===========================================================================
namespace ffemb {

struct HostPipe {
  cudaStream_t s_in = nullptr, s_out = nullptr;
  cudaEvent_t ev_in[2] = {}, ev_done[2] = {}, ev_fwd[2] = {}, ev_out[2] = {};
  bool done_rec[2] = {false, false}, out_rec[2] = {false, false}, loaded[2] = {false, false};
  std::vector<int64_t *> idx[2];  // per op: device index buffer of the slot (nullptr: not read by this rank)
  float *out[2] = {}, *grad[2] = {};  // [rows][ld] per slot (sharded: staging next to the exchange matrices)
  int64_t ld = 0;
  int rows = 0;
  int64_t width = 0;  // sum of out_dim over all tables = row length of the host matrices
  int next = 0, cur = -1;
  int last_fetch = -1, prev_fetch = -1;
};

static void pipe_destroy(Model *m) {
  HostPipe *p = m->pipe;
  if (!p) return;
  for (int s = 0; s < 2; ++s) {
    for (int64_t *b : p->idx[s])
      if (b) cudaFree(b);
    // slot 0 of an unsharded model borrows the model's own matrices
    if (p->out[s] && p->out[s] != m->out_mat) cudaFree(p->out[s]);
    if (p->grad[s] && p->grad[s] != m->grad_mat) cudaFree(p->grad[s]);
    for (cudaEvent_t e : {p->ev_in[s], p->ev_done[s], p->ev_fwd[s], p->ev_out[s]})
      if (e) cudaEventDestroy(e);
  }
  if (p->s_in) cudaStreamDestroy(p->s_in);
  if (p->s_out) cudaStreamDestroy(p->s_out);
  delete p;
  m->pipe = nullptr;
}

static bool reads_indices(const Model *m, const Op *op) { return op->replicated || owned(op); }

static int pipe_create(Model *m) {
  if (m->pipe) return FFEMB_OK;
  auto p = std::make_unique<HostPipe>();
  const int B = m->cfg.batch_size;
  FFEMB_CUDA_OK(cudaStreamCreateWithFlags(&p->s_in, cudaStreamNonBlocking));
  FFEMB_CUDA_OK(cudaStreamCreateWithFlags(&p->s_out, cudaStreamNonBlocking));
  for (auto &op : m->ops) p->width += op->D;
  if (m->x.on) {
    p->rows = m->x.Bs;
    p->ld = m->x.ldx;
  } else {
    p->rows = B;
    p->ld = (p->width + 3) / 4 * 4;
    for (auto &op : m->ops)
      if (!owned(op.get()))
        return fail(FFEMB_ERR_UNSUPPORTED, "host pipeline: every table must be held by this rank");
  }
  m->pipe = p.release();  // from here on pipe_destroy cleans up after a failure
  HostPipe *q = m->pipe;
  for (int s = 0; s < 2; ++s) {
    for (cudaEvent_t *e : {&q->ev_in[s], &q->ev_done[s], &q->ev_fwd[s], &q->ev_out[s]})
      FFEMB_CUDA_OK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    q->idx[s].assign(m->ops.size(), nullptr);
    for (auto &op : m->ops)
      if (reads_indices(m, op.get()))
        FFEMB_CUDA_OK(cudaMalloc(&q->idx[s][op->index], (size_t)B * op->input->dims[1] * sizeof(int64_t)));
    const size_t bytes = (size_t)q->rows * q->ld * sizeof(float);
    if (!m->x.on && s == 0 && m->out_mat && m->mat_ld == q->ld) {
      q->out[0] = m->out_mat;
      q->grad[0] = m->grad_mat;
    } else {
      FFEMB_CUDA_OK(cudaMalloc(&q->out[s], bytes));
      FFEMB_CUDA_OK(cudaMalloc(&q->grad[s], bytes));
      FFEMB_CUDA_OK(cudaMemsetAsync(q->grad[s], 0, bytes, m->stream));
    }
  }
  return FFEMB_OK;
}

}  // namespace ffemb

using namespace ffemb;

extern "C" int ffemb_model_load_batch_async(ffemb_model_t h, const int64_t *const *indices, const float *out_grad) {
  Model *m = M(h);
  if (!m || !m->compiled) return fail(FFEMB_ERR_INVALID_ARGUMENT, "load_batch_async: model is not compiled");
  if (!indices) return fail(FFEMB_ERR_INVALID_ARGUMENT, "load_batch_async: indices is NULL");
  FFEMB_CUDA_OK(cudaSetDevice(m->device));
  if (int rc = pipe_create(m)) return rc;
  HostPipe *p = m->pipe;
  const int s = p->next;
  if (p->loaded[s]) return fail(FFEMB_ERR_INVALID_ARGUMENT, "load_batch_async: both slots hold a batch; call begin_batch");
  // the slot's previous batch has been consumed (kernels) and its outputs fetched
  if (p->done_rec[s]) FFEMB_CUDA_OK(cudaStreamWaitEvent(p->s_in, p->ev_done[s], 0));
  const int B = m->cfg.batch_size;
  for (auto &op : m->ops) {
    int64_t *dst = p->idx[s][op->index];
    if (!dst) continue;
    const int64_t *src = indices[op->index];
    if (!src)
      return fail(FFEMB_ERR_INVALID_ARGUMENT, "load_batch_async: no indices for " + op->name + ", which this rank reads");
    const int L = op->input->dims[1];
    int64_t first = 0, cnt = (int64_t)B * L;
    if (op->replicated) {  // only this rank's own samples of a replicated table are read
      cnt = (int64_t)m->x.Bs * L;
      first = (int64_t)m->cfg.rank * cnt;
    }
    FFEMB_CUDA_OK(cudaMemcpyAsync(dst + first, src + first, cnt * sizeof(int64_t), cudaMemcpyHostToDevice, p->s_in));
  }
  if (out_grad) {
    if (p->ld == p->width)  // contiguous on both sides: one linear copy runs at the full PCIe rate
      FFEMB_CUDA_OK(cudaMemcpyAsync(p->grad[s], out_grad, (size_t)p->rows * p->width * sizeof(float),
                                    cudaMemcpyHostToDevice, p->s_in));
    else
      FFEMB_CUDA_OK(cudaMemcpy2DAsync(p->grad[s], p->ld * sizeof(float), out_grad, p->width * sizeof(float),
                                      p->width * sizeof(float), p->rows, cudaMemcpyHostToDevice, p->s_in));
  }
  FFEMB_CUDA_OK(cudaEventRecord(p->ev_in[s], p->s_in));
  p->loaded[s] = true;
  p->next = s ^ 1;
  return FFEMB_OK;
}

extern "C" int ffemb_model_begin_batch(ffemb_model_t h) {
  Model *m = M(h);
  if (!m || !m->compiled || !m->pipe) return fail(FFEMB_ERR_INVALID_ARGUMENT, "begin_batch: no batch was loaded");
  FFEMB_CUDA_OK(cudaSetDevice(m->device));
  HostPipe *p = m->pipe;
  const int s = p->cur < 0 ? 0 : p->cur ^ 1;
  if (!p->loaded[s]) return fail(FFEMB_ERR_INVALID_ARGUMENT, "begin_batch: the next slot holds no batch; call load_batch_async");
  if (p->cur >= 0) {  // everything that used the old slot has been enqueued by now
    FFEMB_CUDA_OK(cudaEventRecord(p->ev_done[p->cur], m->stream));
    p->done_rec[p->cur] = true;
  }
  FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, p->ev_in[s], 0));
  if (p->out_rec[s]) FFEMB_CUDA_OK(cudaStreamWaitEvent(m->stream, p->ev_out[s], 0));  // its outputs have left
  for (auto &op : m->ops) {
    if (p->idx[s][op->index]) {
      release(op->input->data);
      op->input->data = Buffer{p->idx[s][op->index], op->input->dims[1], false};
    }
    if (!m->x.on) {  // the slot's matrices ARE the outputs / gradients of this batch
      int64_t col = 0;
      for (auto &o2 : m->ops) {
        if (o2.get() == op.get()) break;
        col += o2->D;
      }
      release(op->output->data);
      release(op->output->grad);
      op->output->data = Buffer{p->out[s] + col, p->ld, false};
      op->output->grad = Buffer{p->grad[s] + col, p->ld, false};
    }
  }
  if (m->x.on)  // the exchange matrices are fixed (peers hold their addresses): copy in
    FFEMB_CUDA_OK(cudaMemcpyAsync(m->x.buf + m->x.off_grad, p->grad[s], (size_t)p->rows * p->ld * sizeof(float),
                                  cudaMemcpyDeviceToDevice, m->stream));
  m->presorted = false;
  p->loaded[s] = false;
  p->cur = s;
  return FFEMB_OK;
}

extern "C" int ffemb_model_fetch_outputs_async(ffemb_model_t h, float *out_host) {
  Model *m = M(h);
  if (!m || !m->compiled || !m->pipe || m->pipe->cur < 0 || !out_host)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "fetch_outputs_async: no current batch (begin_batch) or NULL buffer");
  FFEMB_CUDA_OK(cudaSetDevice(m->device));
  HostPipe *p = m->pipe;
  const int s = p->cur;
  if (m->x.on)
    FFEMB_CUDA_OK(cudaMemcpyAsync(p->out[s], m->x.buf + m->x.off_out, (size_t)p->rows * p->ld * sizeof(float),
                                  cudaMemcpyDeviceToDevice, m->stream));
  FFEMB_CUDA_OK(cudaEventRecord(p->ev_fwd[s], m->stream));
  FFEMB_CUDA_OK(cudaStreamWaitEvent(p->s_out, p->ev_fwd[s], 0));
  if (p->ld == p->width)
    FFEMB_CUDA_OK(cudaMemcpyAsync(out_host, p->out[s], (size_t)p->rows * p->width * sizeof(float),
                                  cudaMemcpyDeviceToHost, p->s_out));
  else
    FFEMB_CUDA_OK(cudaMemcpy2DAsync(out_host, p->width * sizeof(float), p->out[s], p->ld * sizeof(float),
                                    p->width * sizeof(float), p->rows, cudaMemcpyDeviceToHost, p->s_out));
  FFEMB_CUDA_OK(cudaEventRecord(p->ev_out[s], p->s_out));
  p->out_rec[s] = true;
  p->prev_fetch = p->last_fetch;
  p->last_fetch = s;
  return FFEMB_OK;
}

extern "C" int ffemb_model_wait_outputs(ffemb_model_t h, bool all) {
  Model *m = M(h);
  if (!m || !m->pipe) return FFEMB_OK;
  HostPipe *p = m->pipe;
  if (p->prev_fetch >= 0 && p->prev_fetch != p->last_fetch) FFEMB_CUDA_OK(cudaEventSynchronize(p->ev_out[p->prev_fetch]));
  if (all && p->last_fetch >= 0) FFEMB_CUDA_OK(cudaEventSynchronize(p->ev_out[p->last_fetch]));
  return FFEMB_OK;
}

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
  bool needed = false, slice = false;
  for (auto &op : m->ops) {
    needed = needed || (op->input == t && owned(op.get()));
    slice = slice || (op->input == t && op->replicated);  // a replica reads this rank's own samples only
  }
  if (!needed && !slice) return FFEMB_OK;
  if (!t->data.ptr) {
    void *p = nullptr;
    FFEMB_CUDA_OK(cudaMalloc(&p, n * sizeof(int64_t)));
    t->data = Buffer{p, t->dims[1], true};
  }
  int64_t first = 0, cnt = n;
  if (!needed) {
    cnt = (int64_t)m->x.Bs * t->dims[1];
    first = (int64_t)m->cfg.rank * cnt;
  }
  FFEMB_CUDA_OK(cudaMemcpyAsync((int64_t *)t->data.ptr + first, data + first, cnt * sizeof(int64_t),
                                cudaMemcpyHostToDevice, m->stream));
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
  // sharded mode: every rank holds the gradient of ITS samples for every table
  const int rows = m->x.on ? m->x.Bs : t->dims[0];
  if (count != (int64_t)rows * t->dims[1])
    return fail(FFEMB_ERR_INVALID_ARGUMENT,
                m->x.on ? "tensor_set_grad_float: count != (batch / ranks) * out_dim"
                        : "tensor_set_grad_float: count != batch * out_dim");
  if (!m->x.on && !owned(t->producer)) return FFEMB_OK;
  FFEMB_CUDA_OK(cudaMemcpy2DAsync(t->grad.ptr, t->grad.ld * sizeof(float), data,
                                  t->dims[1] * sizeof(float), t->dims[1] * sizeof(float), rows,
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
  if (!t->producer || (!m->x.on && !owned(t->producer)))
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "tensor_get_float: not an output of a table this rank owns");
  // sharded mode: this rank's samples of every table's output
  const int rows = m->x.on ? m->x.Bs : t->dims[0];
  if (count != (int64_t)rows * t->dims[1])
    return fail(FFEMB_ERR_INVALID_ARGUMENT, m->x.on ? "tensor_get_float: count != (batch / ranks) * out_dim"
                                                    : "tensor_get_float: count != batch * out_dim");
  FFEMB_CUDA_OK(cudaMemcpy2DAsync(data, t->dims[1] * sizeof(float), t->data.ptr,
                                  t->data.ld * sizeof(float), t->dims[1] * sizeof(float), rows,
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
  if (t->producer && t->model->x.on)
    return fail(FFEMB_ERR_UNSUPPORTED,
                "tensor_attach_device_ptr: in sharded mode outputs and gradients live in the library's "
                "peer-mapped exchange buffer (ffemb_model_get_exchange_matrix)");
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
  if (!owned(op) && !op->replicated) {
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
  if (!owned(op) && !op->replicated) {
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
