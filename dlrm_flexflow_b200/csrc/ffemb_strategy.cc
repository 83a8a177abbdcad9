// ffemb_strategy.cc -- placement hook: strategy files and table -> device.
//
// Reads and writes the reference's FFProtoBuf::Strategy wire format
// (src/runtime/strategy.proto:5-23, proto2) without protobuf:
//   Strategy { repeated Op ops = 1; }
//   Op { required string name = 1; required DeviceType device_type = 2;
//        repeated int32 dims = 3; repeated int32 device_ids = 4;
//        repeated MemoryType memory_types = 5; }
// Fields are written in field-number order with unpacked repeated scalars,
// which is what protobuf's C++ serialiser emits for this schema, so files
// written here are byte-identical to the reference's checked-in
// dlrm_strategy_*.pb fixtures.  Packed encodings are accepted on input.
//
// The table -> device rule is FFMapper::slice_task's
// (src/mapper/mapper.cc:33-97): devices[device_ids[point] % devices.size()].
#include <stdint.h>
#include <stdio.h>

#include <algorithm>
#include <numeric>
#include <string>
#include <vector>

#include "ffemb_c.h"

namespace ffemb {
int fail(int status, const std::string &msg);

struct StrategyOp {
  std::string name;
  int device_type = FFEMB_DEVICE_GPU;
  std::vector<int> dims, device_ids, memory_types;
};

struct Strategy {
  std::vector<StrategyOp> ops;
};

// ---- varint wire helpers -----------------------------------------------------
static bool get_varint(const uint8_t *&p, const uint8_t *end, uint64_t &v) {
  v = 0;
  for (int shift = 0; shift < 64 && p < end; shift += 7) {
    const uint8_t b = *p++;
    v |= (uint64_t)(b & 0x7F) << shift;
    if (!(b & 0x80)) return true;
  }
  return false;
}

static void put_varint(std::string &out, uint64_t v) {
  while (v >= 0x80) {
    out.push_back((char)(v | 0x80));
    v >>= 7;
  }
  out.push_back((char)v);
}

static void put_int32_field(std::string &out, int field, int value) {
  put_varint(out, (uint64_t)field << 3);  // wire type 0
  put_varint(out, (uint64_t)(int64_t)value);  // negative int32 -> 10 bytes, as protobuf
}

static bool skip_field(const uint8_t *&p, const uint8_t *end, int wire) {
  uint64_t v;
  switch (wire) {
    case 0: return get_varint(p, end, v);
    case 1: if (end - p < 8) return false; p += 8; return true;
    case 2: if (!get_varint(p, end, v) || (uint64_t)(end - p) < v) return false; p += v; return true;
    case 5: if (end - p < 4) return false; p += 4; return true;
    default: return false;
  }
}

static bool get_repeated(const uint8_t *&p, const uint8_t *end, int wire, std::vector<int> &dst) {
  uint64_t v;
  if (wire == 0) {
    if (!get_varint(p, end, v)) return false;
    dst.push_back((int)(int64_t)v);
    return true;
  }
  if (wire != 2) return false;  // packed
  if (!get_varint(p, end, v) || (uint64_t)(end - p) < v) return false;
  const uint8_t *e = p + v;
  while (p < e) {
    uint64_t x;
    if (!get_varint(p, e, x)) return false;
    dst.push_back((int)(int64_t)x);
  }
  return true;
}

static bool parse_op(const uint8_t *p, const uint8_t *end, StrategyOp &op) {
  bool has_name = false;
  while (p < end) {
    uint64_t tag, v;
    if (!get_varint(p, end, tag)) return false;
    const int field = (int)(tag >> 3), wire = (int)(tag & 7);
    switch (field) {
      case 1:
        if (wire != 2 || !get_varint(p, end, v) || (uint64_t)(end - p) < v) return false;
        op.name.assign((const char *)p, (size_t)v);
        p += v;
        has_name = true;
        break;
      case 2:
        if (wire != 0 || !get_varint(p, end, v)) return false;
        op.device_type = (int)v;
        break;
      case 3: if (!get_repeated(p, end, wire, op.dims)) return false; break;
      case 4: if (!get_repeated(p, end, wire, op.device_ids)) return false; break;
      case 5: if (!get_repeated(p, end, wire, op.memory_types)) return false; break;
      default: if (!skip_field(p, end, wire)) return false;
    }
  }
  return has_name;  // `name` is a required field
}

static bool parse_strategy(const uint8_t *p, size_t n, Strategy &s) {
  const uint8_t *end = p + n;
  while (p < end) {
    uint64_t tag, len;
    if (!get_varint(p, end, tag)) return false;
    if ((tag >> 3) == 1 && (tag & 7) == 2) {
      if (!get_varint(p, end, len) || (uint64_t)(end - p) < len) return false;
      StrategyOp op;
      if (!parse_op(p, p + len, op)) return false;
      s.ops.push_back(std::move(op));
      p += len;
    } else if (!skip_field(p, end, (int)(tag & 7))) {
      return false;
    }
  }
  return true;
}

static std::string serialize(const Strategy &s) {
  std::string out;
  for (const StrategyOp &op : s.ops) {
    std::string body;
    put_varint(body, (1u << 3) | 2);
    put_varint(body, op.name.size());
    body += op.name;
    put_int32_field(body, 2, op.device_type);
    for (int d : op.dims) put_int32_field(body, 3, d);
    for (int d : op.device_ids) put_int32_field(body, 4, d);
    for (int m : op.memory_types) put_int32_field(body, 5, m);
    put_varint(out, (1u << 3) | 2);
    put_varint(out, body.size());
    out += body;
  }
  return out;
}

static Strategy *S(ffemb_strategy_t h) { return (Strategy *)h.impl; }

static void add(Strategy *s, const std::string &name, int dev_type, std::vector<int> dims,
                std::vector<int> ids, std::vector<int> mem) {
  StrategyOp op;
  op.name = name;
  op.device_type = dev_type;
  op.dims = std::move(dims);
  op.device_ids = std::move(ids);
  op.memory_types = std::move(mem);
  s->ops.push_back(std::move(op));
}

}  // namespace ffemb

using namespace ffemb;

extern "C" ffemb_strategy_t ffemb_strategy_create(void) { return ffemb_strategy_t{new Strategy}; }
extern "C" void ffemb_strategy_destroy(ffemb_strategy_t h) { delete S(h); }

extern "C" int ffemb_strategy_parse(ffemb_strategy_t h, const uint8_t *bytes, size_t n) {
  if (!S(h) || (!bytes && n)) return fail(FFEMB_ERR_INVALID_ARGUMENT, "strategy_parse: NULL argument");
  Strategy tmp;
  if (!parse_strategy(bytes, n, tmp))
    return fail(FFEMB_ERR_IO, "Failed to parse strategy file");  // strategy.cc:101-104
  // the reference asserts that no name occurs twice (strategy.cc:129)
  for (size_t i = 0; i < tmp.ops.size(); ++i)
    for (size_t j = 0; j < i; ++j)
      if (tmp.ops[i].name == tmp.ops[j].name)
        return fail(FFEMB_ERR_IO, "strategy has two entries named " + tmp.ops[i].name);
  for (const StrategyOp &op : tmp.ops) {
    // strategy.cc:124: product(dims) == #device_ids, or no device ids at all
    int64_t n_parts = 1;
    for (int d : op.dims) n_parts *= d;
    if (!op.device_ids.empty() && (int64_t)op.device_ids.size() != n_parts)
      return fail(FFEMB_ERR_IO, "strategy entry " + op.name + ": product(dims) != #device_ids");
    if (op.device_type != FFEMB_DEVICE_GPU && op.device_type != FFEMB_DEVICE_CPU)
      return fail(FFEMB_ERR_IO, "Unsupported Device Type");  // strategy.cc:114-116
  }
  *S(h) = std::move(tmp);
  return FFEMB_OK;
}

extern "C" size_t ffemb_strategy_serialize(ffemb_strategy_t h, uint8_t *bytes, size_t cap) {
  if (!S(h)) return 0;
  const std::string out = serialize(*S(h));
  if (bytes && cap >= out.size()) std::copy(out.begin(), out.end(), (char *)bytes);
  return out.size();
}

extern "C" int ffemb_strategy_load_file(ffemb_strategy_t h, const char *path) {
  if (!S(h) || !path) return fail(FFEMB_ERR_INVALID_ARGUMENT, "strategy_load_file: NULL argument");
  FILE *f = fopen(path, "rb");
  if (!f) return fail(FFEMB_ERR_IO, std::string("cannot open strategy file ") + path);
  std::vector<uint8_t> buf;
  uint8_t chunk[4096];
  size_t got;
  while ((got = fread(chunk, 1, sizeof chunk, f)) > 0) buf.insert(buf.end(), chunk, chunk + got);
  fclose(f);
  return ffemb_strategy_parse(h, buf.data(), buf.size());
}

extern "C" int ffemb_strategy_save_file(ffemb_strategy_t h, const char *path) {
  if (!S(h) || !path) return fail(FFEMB_ERR_INVALID_ARGUMENT, "strategy_save_file: NULL argument");
  const std::string out = serialize(*S(h));
  FILE *f = fopen(path, "wb");
  if (!f) return fail(FFEMB_ERR_IO, std::string("cannot write strategy file ") + path);
  const bool ok = fwrite(out.data(), 1, out.size(), f) == out.size();
  fclose(f);
  return ok ? FFEMB_OK : fail(FFEMB_ERR_IO, "Failed to save to strategy file");
}

extern "C" int ffemb_strategy_num_ops(ffemb_strategy_t h) { return S(h) ? (int)S(h)->ops.size() : 0; }

#define FFEMB_OP_OR(h, i, dflt)                                 \
  if (!S(h) || (i) < 0 || (i) >= (int)S(h)->ops.size()) return dflt; \
  const StrategyOp &op = S(h)->ops[i]

extern "C" const char *ffemb_strategy_op_name(ffemb_strategy_t h, int i) {
  FFEMB_OP_OR(h, i, nullptr);
  return op.name.c_str();
}
extern "C" int ffemb_strategy_op_device_type(ffemb_strategy_t h, int i) {
  FFEMB_OP_OR(h, i, -1);
  return op.device_type;
}
extern "C" int ffemb_strategy_op_num_dims(ffemb_strategy_t h, int i) {
  FFEMB_OP_OR(h, i, -1);
  return (int)op.dims.size();
}
extern "C" int ffemb_strategy_op_dim(ffemb_strategy_t h, int i, int k) {
  FFEMB_OP_OR(h, i, -1);
  return k >= 0 && k < (int)op.dims.size() ? op.dims[k] : -1;
}
extern "C" int ffemb_strategy_op_num_device_ids(ffemb_strategy_t h, int i) {
  FFEMB_OP_OR(h, i, -1);
  return (int)op.device_ids.size();
}
extern "C" int ffemb_strategy_op_device_id(ffemb_strategy_t h, int i, int k) {
  FFEMB_OP_OR(h, i, -1);
  return k >= 0 && k < (int)op.device_ids.size() ? op.device_ids[k] : -1;
}
extern "C" int ffemb_strategy_op_num_memory_types(ffemb_strategy_t h, int i) {
  FFEMB_OP_OR(h, i, -1);
  return (int)op.memory_types.size();
}
extern "C" int ffemb_strategy_op_memory_type(ffemb_strategy_t h, int i, int k) {
  FFEMB_OP_OR(h, i, -1);
  return k >= 0 && k < (int)op.memory_types.size() ? op.memory_types[k] : -1;
}

extern "C" int ffemb_strategy_add_op(ffemb_strategy_t h, const char *name, int device_type,
                                     const int *dims, int num_dims, const int *device_ids,
                                     int num_device_ids, const int *memory_types,
                                     int num_memory_types) {
  if (!S(h) || !name || num_dims < 0 || num_device_ids < 0 || num_memory_types < 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "strategy_add_op: bad arguments");
  if (ffemb_strategy_find(h, name) >= 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, std::string("strategy already has an entry named ") + name);
  add(S(h), name, device_type, std::vector<int>(dims, dims + num_dims),
      std::vector<int>(device_ids, device_ids + num_device_ids),
      std::vector<int>(memory_types, memory_types + num_memory_types));
  return FFEMB_OK;
}

extern "C" int ffemb_strategy_find(ffemb_strategy_t h, const char *name) {
  if (!S(h) || !name) return -1;
  for (size_t i = 0; i < S(h)->ops.size(); ++i)
    if (S(h)->ops[i].name == name) return (int)i;
  return -1;
}

// src/runtime/dlrm_strategy.cc:242-296 (entries and their order).
extern "C" int ffemb_strategy_generate_dlrm(ffemb_strategy_t h, int gpus_per_node,
                                            int embs_per_node, int num_nodes) {
  if (!S(h)) return fail(FFEMB_ERR_INVALID_ARGUMENT, "strategy_generate_dlrm: NULL handle");
  if (gpus_per_node <= 0 || embs_per_node <= 0 || num_nodes <= 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "Invalide input configurations...");  // dlrm_strategy.cc:91-94
  Strategy *s = S(h);
  s->ops.clear();
  const int G = FFEMB_DEVICE_GPU, FBM = FFEMB_MEMORY_FBM;
  const int ngpu = gpus_per_node * num_nodes;
  for (int i = 0; i < num_nodes * embs_per_node; ++i)
    add(s, "embedding" + std::to_string(i), G, {1, 1}, {i % ngpu}, {FBM, FBM, FBM});
  std::vector<int> node_heads, all(ngpu);
  for (int i = 0; i < num_nodes; ++i) node_heads.push_back(i * gpus_per_node);
  std::iota(all.begin(), all.end(), 0);
  add(s, "concat", G, {1, num_nodes}, node_heads, {FBM, FBM});
  add(s, "batch_matmul", G, {1, 1, ngpu}, all, {FBM, FBM, FBM});
  add(s, "transpose", G, {1, 1, ngpu}, all, {FBM, FBM});
  add(s, "linear", G, {1, ngpu}, all, {FBM, FBM, FBM});
  add(s, "mse_loss", G, {1, ngpu}, all, {FBM});
  return FFEMB_OK;
}

// src/runtime/dlrm_strategy.py:26-46 -- the generator the checked-in
// dlrm_strategy_<E>embs_<G>gpus.pb fixtures came from.
extern "C" int ffemb_strategy_generate_dlrm_simple(ffemb_strategy_t h, int num_gpu, int num_emb) {
  if (!S(h) || num_gpu <= 0 || num_emb < 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "strategy_generate_dlrm_simple: bad arguments");
  Strategy *s = S(h);
  s->ops.clear();
  for (int i = 0; i < num_emb; ++i)
    add(s, "embedding" + std::to_string(i), FFEMB_DEVICE_GPU, {1, 1}, {i % num_gpu}, {});
  std::vector<int> all(num_gpu);
  std::iota(all.begin(), all.end(), 0);
  for (const char *name : {"linear", "mse_loss", "concat"})
    add(s, name, FFEMB_DEVICE_GPU, {1, num_gpu}, all, {});
  return FFEMB_OK;
}

extern "C" int ffemb_strategy_device_for_table(ffemb_strategy_t h, int table_index,
                                               const char *op_name, int num_devices) {
  if (num_devices <= 0 || table_index < 0) return -1;
  if (S(h)) {
    int i = ffemb_strategy_find(h, ("embedding" + std::to_string(table_index)).c_str());
    if (i < 0 && op_name) i = ffemb_strategy_find(h, op_name);
    if (i >= 0 && !S(h)->ops[i].device_ids.empty()) {
      const int d = S(h)->ops[i].device_ids[0];
      return ((d % num_devices) + num_devices) % num_devices;  // mapper.cc:77-81
    }
  }
  return table_index % num_devices;  // dlrm_strategy.cc:252-256
}

extern "C" int ffemb_placement_balance(const int64_t *num_entries, const int64_t *lookups,
                                       int num_tables, int out_dim, int num_devices,
                                       int64_t bytes_capacity, int *owner) {
  if (!num_entries || !lookups || !owner || num_tables < 0 || num_devices <= 0 || out_dim <= 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "placement_balance: bad arguments");
  // Longest-processing-time first: heaviest table (lookups per step, ties by
  // bytes) to the device with the least work that still has room.
  std::vector<int> order(num_tables);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    if (lookups[a] != lookups[b]) return lookups[a] > lookups[b];
    return num_entries[a] > num_entries[b];
  });
  std::vector<int64_t> work(num_devices, 0), bytes(num_devices, 0);
  for (int t : order) {
    const int64_t tb = num_entries[t] * (int64_t)out_dim * 4;
    int best = -1;
    for (int d = 0; d < num_devices; ++d) {
      if (bytes_capacity > 0 && bytes[d] + tb > bytes_capacity) continue;
      if (best < 0 || work[d] < work[best] || (work[d] == work[best] && bytes[d] < bytes[best]))
        best = d;
    }
    if (best < 0)
      return fail(FFEMB_ERR_INVALID_ARGUMENT,
                  "placement_balance: table " + std::to_string(t) + " does not fit on any device");
    owner[t] = best;
    work[best] += lookups[t];
    bytes[best] += tb;
  }
  return FFEMB_OK;
}

// ---------------------------------------------------------------------------
// Cost-aware placement with the replicate option.
//
// The reference cannot cost an embedding at all: Embedding::measure_compute_time
// returns false (src/ops/embedding.cu:357-364), so its strategy search
// (src/runtime/simulator.cc:200-275, link constants simulator.cu:27-29) never
// moves the tables off the generator's round-robin.  This hook models one
// step of the sharded engine per rank r:
//   NVLink  2 * B*D*4*(N-1)/N bytes per sharded table r owns (pooled rows out,
//           gradients in) + 2 * replica_bytes*(N-1)/N (sum + broadcast of the
//           replicated tables' dense gradient / stepped rows)
//   HBM     hbm_passes * D*4 bytes per lookup r processes (whole batch for its
//           sharded tables, batch/N for every replicated one: row read, output
//           write, gradient read, row read + write) + 3 * replica_bytes
//   step    max(NVLink time, HBM time) [+ replica_latency_us if anything is
//           replicated: the chain of small kernels, two barriers and the
//           collective that the replicas add]
// and picks, among "replicate the k smallest tables, balance the rest by
// lookups" for k = 0..T, the k with the shortest modelled step.  Defaults are
// this pool's measured rates (650 GB/s peer stores inside a kernel, 5.5 TB/s
// gather/reduce kernels); with them the model lands within 15 % of the
// measured C3 steps at 2 and 8 ranks, sharded and replicated.
// ---------------------------------------------------------------------------
extern "C" int ffemb_placement_plan(const int64_t *num_entries, const int64_t *lookups, int num_tables,
                                    int out_dim, int batch, int num_devices, int64_t bytes_capacity,
                                    const ffemb_cost_model_t *cost, int *owner, double *modelled_step_us) {
  if (!num_entries || !lookups || !owner || num_tables < 0 || num_devices <= 0 || out_dim <= 0 || batch <= 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "placement_plan: bad arguments");
  ffemb_cost_model_t c = {5500.0, 650.0, 5.0, 60.0};
  if (cost) c = *cost;
  if (c.hbm_gbs <= 0 || c.nvlink_gbs <= 0 || c.hbm_passes <= 0 || c.replica_latency_us < 0)
    return fail(FFEMB_ERR_INVALID_ARGUMENT, "placement_plan: bad cost model");
  const int T = num_tables, N = num_devices;
  std::vector<int> by_rows(T);
  std::iota(by_rows.begin(), by_rows.end(), 0);
  std::stable_sort(by_rows.begin(), by_rows.end(), [&](int a, int b) { return num_entries[a] < num_entries[b]; });
  const double row = (double)out_dim * 4.0;
  const double frac = N > 1 ? (double)(N - 1) / N : 0.0;
  double best = -1.0;
  std::vector<int> best_owner(T, 0), cur(T, 0);
  for (int k = 0; k <= (N > 1 ? T : 0); ++k) {
    // the k smallest tables replicated, the rest balanced by lookups under the capacity
    std::vector<int64_t> ne, lk;
    std::vector<int> idx;
    double rep_bytes = 0, rep_lookups = 0;
    int64_t rep_cap = 0;
    for (int i = 0; i < T; ++i) {
      const int t = by_rows[i];
      if (i < k) {
        cur[t] = -1;
        rep_bytes += (double)num_entries[t] * row;
        rep_lookups += (double)lookups[t] / N;
        rep_cap += num_entries[t] * (int64_t)out_dim * 8;  // table + dense gradient on every rank
      } else {
        idx.push_back(t);
        ne.push_back(num_entries[t]);
        lk.push_back(lookups[t]);
      }
    }
    std::vector<int> own(idx.size(), 0);
    if (bytes_capacity > 0 && rep_cap > bytes_capacity) continue;  // the replicas alone do not fit
    if (!idx.empty()) {
      const int64_t cap = bytes_capacity > 0 ? bytes_capacity - rep_cap : 0;
      if (bytes_capacity > 0 && cap <= 0) continue;
      if (ffemb_placement_balance(ne.data(), lk.data(), (int)idx.size(), out_dim, N, cap, own.data())) continue;
    }
    std::vector<double> nv(N, 0.0), hb(N, 0.0);
    for (size_t i = 0; i < idx.size(); ++i) {
      cur[idx[i]] = own[i];
      nv[own[i]] += 2.0 * (double)batch * row * frac;
      hb[own[i]] += c.hbm_passes * (double)lk[i] * row;
    }
    double step = 0.0;
    for (int r = 0; r < N; ++r) {
      const double t_nv = (nv[r] + 2.0 * rep_bytes * frac) / (c.nvlink_gbs * 1e3);          // us
      const double t_hb = (hb[r] + c.hbm_passes * rep_lookups * row + 3.0 * rep_bytes) / (c.hbm_gbs * 1e3);
      step = std::max(step, std::max(t_nv, t_hb));
    }
    if (k > 0) step += c.replica_latency_us;
    if (best < 0 || step < best) {
      best = step;
      best_owner = cur;
    }
  }
  if (best < 0) return fail(FFEMB_ERR_INVALID_ARGUMENT, "placement_plan: the tables do not fit the devices");
  for (int t = 0; t < T; ++t) owner[t] = best_owner[t];
  if (modelled_step_us) *modelled_step_us = best;
  return FFEMB_OK;
}
