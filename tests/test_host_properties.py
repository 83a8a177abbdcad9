"""CPU property tests of the host-side logic: the strategy wire codec against
google.protobuf on random strategies, the cffi binding flow the reference uses
(gcc -E | cffi.cdef, python/flexflow/core/legion_cffi_build.py idiom), and the
sharded plan's layout algebra on random placements."""
import os
import subprocess
import sys

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))

import dlrm_flexflow_b200 as ff  # noqa: E402
from dlrm_flexflow_b200.sharded import ShardPlan  # noqa: E402


def _pb_class():
    """FFProtoBuf.Strategy built from the schema of src/runtime/strategy.proto:5-23."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name, fdp.package, fdp.syntax = "strategy_prop.proto", "FFProtoBufProp", "proto2"
    op = fdp.message_type.add()
    op.name = "Op"
    for ename, vals in (("DeviceType", ("GPU", "CPU")), ("MemoryType", ("FBM", "ZCM"))):
        e = op.enum_type.add()
        e.name = ename
        for i, v in enumerate(vals):
            ev = e.value.add()
            ev.name, ev.number = v, i
    F = descriptor_pb2.FieldDescriptorProto
    for name, num, label, typ, tn in (
            ("name", 1, F.LABEL_REQUIRED, F.TYPE_STRING, None),
            ("device_type", 2, F.LABEL_REQUIRED, F.TYPE_ENUM, ".FFProtoBufProp.Op.DeviceType"),
            ("dims", 3, F.LABEL_REPEATED, F.TYPE_INT32, None),
            ("device_ids", 4, F.LABEL_REPEATED, F.TYPE_INT32, None),
            ("memory_types", 5, F.LABEL_REPEATED, F.TYPE_ENUM, ".FFProtoBufProp.Op.MemoryType")):
        f = op.field.add()
        f.name, f.number, f.label, f.type = name, num, label, typ
        if tn:
            f.type_name = tn
    s = fdp.message_type.add()
    s.name = "Strategy"
    f = s.field.add()
    f.name, f.number, f.label, f.type, f.type_name = "ops", 1, F.LABEL_REPEATED, F.TYPE_MESSAGE, ".FFProtoBufProp.Op"
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    return message_factory.GetMessageClass(pool.FindMessageTypeByName("FFProtoBufProp.Strategy"))


PB = _pb_class()

op_strategy = st.tuples(
    st.integers(0, 1),                                                # device_type
    st.lists(st.integers(1, 4), min_size=0, max_size=3),               # dims
    st.lists(st.integers(0, 1), min_size=0, max_size=3),               # memory types
    st.integers(-3, 300),                                              # first device id
)


@settings(max_examples=60, deadline=None)
@given(st.lists(op_strategy, min_size=0, max_size=12))
def test_strategy_codec_agrees_with_protobuf(ops):
    s = ff.Strategy()
    msg = PB()
    for i, (dev, dims, mem, first) in enumerate(ops):
        n = int(np.prod(dims)) if dims else 1
        ids = [first + k for k in range(n)] if (i % 3) else []        # device_ids may be absent (strategy.cc:124)
        name = "op%d_%s" % (i, "x" * (i * 17 % 140))                   # names past 127 bytes: 2-byte length varint
        s.add_op(name, dev, dims, ids, mem)
        o = msg.ops.add()
        o.name, o.device_type = name, dev
        o.dims.extend(dims)
        o.device_ids.extend(ids)
        o.memory_types.extend(mem)
    ours = s.serialize()
    assert ours == msg.SerializeToString()                             # byte-identical to protobuf's serialiser
    back = PB()
    back.ParseFromString(ours)
    parsed = ff.Strategy(data=msg.SerializeToString()).ops()
    assert [(o["name"], o["device_type"], o["dims"], o["device_ids"], o["memory_types"]) for o in parsed] == \
           [(o.name, int(o.device_type), list(o.dims), list(o.device_ids), [int(m) for m in o.memory_types])
            for o in back.ops]


def test_header_binds_with_cffi_like_the_reference_does():
    cffi = pytest.importorskip("cffi")
    src = subprocess.check_output(["/usr/bin/gcc", "-E", "-P", "-I", os.path.join(ROOT, "include"),
                                   os.path.join(ROOT, "include", "ffemb_c.h")]).decode()
    ffi = cffi.FFI()
    ffi.cdef(src[src.index("enum AggrMode"):])
    lib = ffi.dlopen(ff.library_path())
    assert ffi.string(lib.ffemb_version()) == b"0.2.0" and lib.AGGR_MODE_SUM == 21 and lib.DT_INT64 == 43
    s = lib.ffemb_strategy_create()
    assert lib.ffemb_strategy_generate_dlrm_simple(s, 8, 16) == 0 and lib.ffemb_strategy_num_ops(s) == 19
    assert lib.ffemb_strategy_device_for_table(s, 11, ffi.NULL, 8) == 3
    lib.ffemb_strategy_destroy(s)
    cfg = lib.ffemb_config_create()
    lib.ffemb_config_set_batch_size(cfg, 32)
    m = lib.ffemb_model_create(cfg)
    t = lib.ffemb_tensor_create(m, 2, [32, 4], b"sparse0", lib.DT_INT64, False)
    out = lib.ffemb_model_add_embedding(m, t, 1000, 16, lib.AGGR_MODE_AVG, [ffi.NULL], [ffi.NULL])
    assert out.impl != ffi.NULL and lib.ffemb_tensor_get_dim(out, 1) == 16
    assert ffi.string(lib.ffemb_op_get_name(lib.ffemb_model_get_layer_by_id(m, 0))) == b"Embed_1000x16_100"
    lib.ffemb_model_destroy(m)
    lib.ffemb_config_destroy(cfg)


@settings(max_examples=40, deadline=None)
@given(st.integers(1, 6), st.integers(1, 9), st.integers(1, 4), st.integers(0, 40), st.data())
def test_shard_plan_layout_roundtrips(world, T, bs, rep_below, data):
    D = 4
    rows = data.draw(st.lists(st.integers(1, 60), min_size=T, max_size=T))
    owners = data.draw(st.lists(st.integers(0, world - 1), min_size=T, max_size=T))
    plan = ShardPlan(rows, D, bs * world, world, owners=owners, replicate_below=rep_below)
    rep = set(plan.replicated)
    assert rep == ({t for t in range(T) if rows[t] < rep_below} if world > 1 else set())
    assert sorted(sum((plan.tables_of(r) for r in range(world)), []) + sorted(rep)) == list(range(T))
    for r in range(world):
        assert plan.forward_send_splits(r) == [plan.Bs * plan.width(r)] * world
        assert plan.backward_recv_splits(r) == plan.forward_send_splits(r)
        assert plan.forward_recv_splits(r) == plan.backward_send_splits(r)
        assert sum(plan.forward_recv_splits(r)) == plan.Bs * D * (T - len(rep))
        for t in plan.tables_of(r):
            assert plan.local_col(t) == plan.tables_of(r).index(t) * D
    # pack (concat -> per-owner blocks) then unpack (blocks -> concat) restores the sharded columns
    rng = np.random.default_rng(T * 100 + world)
    concat = rng.normal(size=(plan.Bs, T * D)).astype(np.float32)
    flat = np.zeros(max(plan.Bs * D * (T - len(rep)), 1), np.float32)
    plan.pack_backward(concat, flat)
    back = np.full_like(concat, np.nan)
    plan.unpack_forward(flat, back)
    for t in range(T):
        blk = back[:, t * D:(t + 1) * D]
        if t in rep:
            assert np.isnan(blk).all()                                 # replicated columns are produced locally
        else:
            assert blk.tobytes() == concat[:, t * D:(t + 1) * D].tobytes()


def test_placement_plan_shard_or_replicate():
    """ffemb_placement_plan: the cost model replicates only small
    tables, never with one rank, never where the exchange is cheap relative to
    the HBM work (multi-hot), keeps the sharded tables balanced and respects
    the capacity."""
    from dlrm_flexflow_b200.workloads import TERABYTE_ROWS as R, WORKLOADS
    B = 65536
    own, us1 = ff.plan_placement(R, [B] * 26, 128, B, 1)
    assert own == [0] * 26
    last = us1
    for N in (2, 4, 8):
        own, us = ff.plan_placement(R, [B] * 26, 128, B, N, 150 << 30)
        rep = [t for t in range(26) if own[t] < 0]
        assert rep and all(0 <= o < N for o in own if o >= 0)
        assert max(R[t] for t in rep) < min(R[t] for t in range(26) if own[t] >= 0)      # the smallest ones
        assert max(R[t] for t in rep) < B                                                 # never a big table
        per = [sum(1 for o in own if o == r) for r in range(N)]
        assert max(per) - min(per) <= 1                                                   # balanced by lookups
        for r in range(N):      # capacity: sharded tables + every replica (table and dense gradient)
            assert sum(R[t] for t in range(26) if own[t] == r) * 512 + sum(R[t] for t in rep) * 1024 <= 150 << 30
        assert us < last        # more ranks, shorter modelled step
        last = us
    w = WORKLOADS["C4"]         # pooling 32: HBM work dwarfs the exchange, nothing to gain from replicas
    own, _ = ff.plan_placement(w["rows"], [w["batch"] * w["bag"]] * 8, 128, w["batch"], 8)
    assert sorted(own) == list(range(8))
    # a cost model with a free network shards everything; with a very slow one it replicates more
    own_fast, _ = ff.plan_placement(R, [B] * 26, 128, B, 8, cost=(5500.0, 1e9, 5.0, 60.0))
    own_slow, _ = ff.plan_placement(R, [B] * 26, 128, B, 8, cost=(5500.0, 50.0, 5.0, 60.0))
    assert sum(o < 0 for o in own_fast) == 0 < sum(o < 0 for o in own_slow)
    with pytest.raises(ff.FFEmbError):
        ff.plan_placement([10 ** 9] * 4, [B] * 4, 128, B, 2, bytes_capacity=1 << 30)       # nothing fits
