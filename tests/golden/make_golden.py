"""Generates the committed golden vectors from the REFERENCE itself.

Run in the authoring container (needs /root/reference):

    python tests/golden/make_golden.py

* embedding_avx2_golden.npz -- inputs and outputs of the reference's own CPU
  kernel EmbeddingLookup_int64_t_float_float__avx2_fma
  (src/ops/embedding_avx2.cc:5-296), compiled in place into oracle/_ref by
  oracle/Makefile and called through oracle.oracle.forward_ref.  Covers the
  D=128/64/32 specialisations and the generic path (D=16, 20), bag sizes 1, 4
  and 32, ragged bags with empty ones, SUM and normalize_by_lengths (AVG).
* embedding_bwd_golden.npz -- inputs and outputs of the reference's own CPU
  backward embed_backward_generic (src/ops/embedding.cc:54-72), compiled in
  place into oracle/_ref (only that Legion-free function) and called through
  oracle.oracle.backward_ref: bag size 1, SUM, heavy duplication, D = 128, 64,
  16 and 20, a fresh (zeroed) gradient and a second call accumulating into the
  first one's result (the reference opens the gradient READ_WRITE,
  src/ops/embedding.cu:350-352).
* strategy_fixtures.json -- size, sha256 and decoded content of the three
  checked-in strategy files src/runtime/dlrm_strategy_*.pb (decoded with the
  protobuf runtime from the schema in src/runtime/strategy.proto, not with our
  codec).

Nothing under /root/reference is copied; only values computed from it.
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402

REF = "/root/reference"


def embedding_cases():
    rng = np.random.default_rng(20261106)
    out = {}
    names = []
    for D in (128, 64, 32, 16, 20):
        for kind in ("L1", "L4", "L32", "ragged"):
            N, B = 61, 19
            W = rng.uniform(-1.0, 1.0, size=(N, D)).astype(np.float32)
            W[3, :] = -0.0  # the CPU path turns -0 rows into +0 (0 + -0)
            if kind == "ragged":
                lengths = rng.integers(0, 7, size=B).astype(np.int32)
                lengths[0] = 0
                lengths[B - 1] = 0
            else:
                lengths = np.full(B, int(kind[1:]), dtype=np.int32)
            idx = rng.integers(0, N, size=int(lengths.sum())).astype(np.int64)
            if idx.size:
                idx[0] = 3
            name = "D%d_%s" % (D, kind)
            names.append(name)
            out[name + "_W"] = W
            out[name + "_idx"] = idx
            out[name + "_len"] = lengths
            out[name + "_sum"] = O.forward_ref(W, idx, lengths, O.AGGR_MODE_SUM)
            out[name + "_avg"] = O.forward_ref(W, idx, lengths, O.AGGR_MODE_AVG)
    out["names"] = np.array(names)
    return out


def backward_cases():
    rng = np.random.default_rng(20261107)
    out, names = {}, []
    for D in (128, 64, 16, 20):
        for N, B in ((7, 130), (61, 97), (500, 64)):
            idx = rng.integers(0, N, size=B).astype(np.int64)
            g1 = rng.normal(size=(B, D)).astype(np.float32)
            g2 = rng.normal(size=(B, D)).astype(np.float32)
            G1 = O.backward_ref(N, D, idx, g1)
            G2 = O.backward_ref(N, D, idx[::-1].copy(), g2, gradW=G1.copy())
            name = "D%d_N%d_B%d" % (D, N, B)
            names.append(name)
            out[name + "_idx"] = idx
            out[name + "_g1"] = g1
            out[name + "_g2"] = g2
            out[name + "_G1"] = G1
            out[name + "_G2"] = G2      # second call (reversed indices) accumulated on top of G1
    out["names"] = np.array(names)
    return out


def decode_pb(path):
    """Decode with google.protobuf from a descriptor built from strategy.proto's
    published schema (src/runtime/strategy.proto:5-23)."""
    from google.protobuf import descriptor_pb2, descriptor_pool, message_factory
    fdp = descriptor_pb2.FileDescriptorProto()
    fdp.name = "strategy.proto"
    fdp.package = "FFProtoBuf"
    fdp.syntax = "proto2"
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
            ("device_type", 2, F.LABEL_REQUIRED, F.TYPE_ENUM, ".FFProtoBuf.Op.DeviceType"),
            ("dims", 3, F.LABEL_REPEATED, F.TYPE_INT32, None),
            ("device_ids", 4, F.LABEL_REPEATED, F.TYPE_INT32, None),
            ("memory_types", 5, F.LABEL_REPEATED, F.TYPE_ENUM, ".FFProtoBuf.Op.MemoryType")):
        f = op.field.add()
        f.name, f.number, f.label, f.type = name, num, label, typ
        if tn:
            f.type_name = tn
    st = fdp.message_type.add()
    st.name = "Strategy"
    f = st.field.add()
    f.name, f.number, f.label, f.type, f.type_name = "ops", 1, F.LABEL_REPEATED, F.TYPE_MESSAGE, ".FFProtoBuf.Op"
    pool = descriptor_pool.DescriptorPool()
    pool.Add(fdp)
    cls = message_factory.GetMessageClass(pool.FindMessageTypeByName("FFProtoBuf.Strategy"))
    msg = cls()
    msg.ParseFromString(open(path, "rb").read())
    return [{"name": o.name, "device_type": int(o.device_type), "dims": list(o.dims),
             "device_ids": list(o.device_ids), "memory_types": [int(m) for m in o.memory_types]}
            for o in msg.ops]


def strategy_fixtures():
    out = {}
    for fn in ("dlrm_strategy_8embs_8gpus.pb", "dlrm_strategy_16embs_8gpus.pb",
               "dlrm_strategy_16embs_16gpus.pb"):
        path = os.path.join(REF, "src", "runtime", fn)
        data = open(path, "rb").read()
        out[fn] = {"size": len(data), "sha256": hashlib.sha256(data).hexdigest(),
                   "ops": decode_pb(path)}
    return out


if __name__ == "__main__":
    assert O.have_ref(), "oracle/_ref missing: make -C oracle in a container with /root/reference"
    np.savez_compressed(os.path.join(HERE, "embedding_avx2_golden.npz"), **embedding_cases())
    assert O.have_ref_bwd(), "oracle/_ref/libembed_backward_generic.so missing: make -C oracle"
    np.savez_compressed(os.path.join(HERE, "embedding_bwd_golden.npz"), **backward_cases())
    with open(os.path.join(HERE, "strategy_fixtures.json"), "w") as f:
        json.dump(strategy_fixtures(), f, indent=1, sort_keys=True)
    print("wrote golden vectors")
