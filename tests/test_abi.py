"""CPU tests of the boundary: the C-ABI library loads, exports every symbol
include/ffemb_c.h declares, the header is plain C, and the host-side logic
(handles, argument checking, strategy codec, placement) behaves like the
reference.  No compute call is made: there is no GPU here."""
import ctypes as C
import hashlib
import json
import os
import re
import subprocess

import pytest

import dlrm_flexflow_b200 as ff
from dlrm_flexflow_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "ffemb_c.h")
FIXTURES = json.load(open(os.path.join(ROOT, "tests", "golden", "strategy_fixtures.json")))


def _declared_functions():
    src = subprocess.check_output(["/usr/bin/gcc", "-E", "-P", "-I", os.path.join(ROOT, "include"), HEADER],
                                  text=True)
    src = src[src.index("ffemb_last_error") - 20:]
    return sorted(set(re.findall(r"\b(ffemb_[a-z0-9_]+)\s*\(", src)))


def test_header_is_plain_c():
    # compiles as C (not C++): what `gcc -E | cffi.cdef` in the reference's
    # legion_cffi_build.py flow needs
    subprocess.check_call(["/usr/bin/gcc", "-std=c99", "-fsyntax-only", "-x", "c", "-I",
                           os.path.join(ROOT, "include"), HEADER])


def test_library_exports_every_declared_symbol():
    names = _declared_functions()
    assert len(names) > 70
    L = C.CDLL(_lib.library_path()) if os.path.exists(_lib.library_path()) else ff.lib()
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, missing


def test_python_binding_covers_every_declared_symbol():
    assert sorted(_lib.SIGNATURES) == _declared_functions()


def test_enum_values_are_the_references():
    # include/ffconst.h:11-28
    assert [m.value for m in ff.AggrMode] == [20, 21, 22]
    assert [m.value for m in ff.DataType] == [40, 41, 42, 43, 44]
    assert ff.lib().ffemb_compiled_sm() == 100


def test_struct_layout_matches_header():
    assert C.sizeof(_lib.Lookup) == 56
    assert C.sizeof(_lib.Matrix) == 16 * 8 + 16
    assert _lib.Lookup.col.offset == 40 and _lib.Lookup.bag_size.offset == 48


# ---- placement hook ----------------------------------------------------------
@pytest.mark.parametrize("fn,ngpu,nemb", [("dlrm_strategy_8embs_8gpus.pb", 8, 8),
                                          ("dlrm_strategy_16embs_8gpus.pb", 8, 16),
                                          ("dlrm_strategy_16embs_16gpus.pb", 16, 16)])
def test_generator_reproduces_reference_fixture_bytes(fn, ngpu, nemb):
    """The three checked-in .pb files (src/runtime/) are the reference's only
    golden artefacts on this path: our generator + serialiser must reproduce
    them byte for byte."""
    data = ff.Strategy.dlrm_simple(ngpu, nemb).serialize()
    assert len(data) == FIXTURES[fn]["size"]
    assert hashlib.sha256(data).hexdigest() == FIXTURES[fn]["sha256"]
    ref_path = os.path.join("/root/reference/src/runtime", fn)
    if os.path.exists(ref_path):
        assert data == open(ref_path, "rb").read()


@pytest.mark.parametrize("fn", sorted(FIXTURES))
def test_parser_decodes_fixture_like_protobuf(fn, tmp_path):
    ngpu = int(re.search(r"_(\d+)gpus", fn).group(1))
    nemb = int(re.search(r"_(\d+)embs", fn).group(1))
    data = ff.Strategy.dlrm_simple(ngpu, nemb).serialize()
    s = ff.Strategy(data=data)
    assert s.ops() == FIXTURES[fn]["ops"]          # decoded by google.protobuf
    p = tmp_path / fn
    s.save(str(p))
    s2 = ff.Strategy(path=str(p))
    assert s2.serialize() == data
    for i in range(nemb):                          # embedding i -> GPU i % ngpu
        assert s2.device_for_table(i, ngpu) == i % ngpu
        assert s2.device_for_table(i, 4) == (i % ngpu) % 4   # mapper.cc: % devices.size()


def test_parser_accepts_packed_and_unknown_fields():
    # Op{name="e", device_type=1, dims packed [1,2], unknown field 9 varint, device_ids [3,4]}
    op = b"\x0a\x01e\x10\x01\x1a\x02\x01\x02\x48\x05\x20\x03\x20\x04"
    data = b"\x0a" + bytes([len(op)]) + op
    s = ff.Strategy(data=data)
    assert s.ops() == [{"name": "e", "device_type": 1, "dims": [1, 2], "device_ids": [3, 4],
                        "memory_types": []}]


def test_parser_rejects_garbage_and_reference_asserts():
    with pytest.raises(ff.FFEmbError) as e:
        ff.Strategy(data=b"\x0a\x05abc")
    assert "Failed to parse strategy file" in str(e.value)       # strategy.cc:101-104
    op = b"\x0a\x01e\x10\x00\x18\x01\x18\x02\x20\x00"            # dims [1,2] but 1 device id
    with pytest.raises(ff.FFEmbError):                           # strategy.cc:124
        ff.Strategy(data=b"\x0a" + bytes([len(op)]) + op)
    s = ff.Strategy()
    s.add_op("embedding0", 0, [1, 1], [0])
    with pytest.raises(ff.FFEmbError):                           # strategy.cc:129 duplicate name
        s.add_op("embedding0", 0, [1, 1], [1])
    with pytest.raises(ff.FFEmbError):
        ff.Strategy(path="/nonexistent/strategy.pb")


def test_negative_int32_roundtrip():
    s = ff.Strategy()
    s.add_op("x", 0, [1], [-1])
    assert ff.Strategy(data=s.serialize()).ops()[0]["device_ids"] == [-1]


def test_full_dlrm_generator_layout():
    ops = ff.Strategy.dlrm(gpus_per_node=4, embs_per_node=6, num_nodes=2).ops()
    names = [o["name"] for o in ops]
    assert names == ["embedding%d" % i for i in range(12)] + ["concat", "batch_matmul", "transpose",
                                                              "linear", "mse_loss"]
    assert [o["device_ids"][0] for o in ops[:12]] == [i % 8 for i in range(12)]  # dlrm_strategy.cc:252-256
    assert ops[0]["memory_types"] == [0, 0, 0] and ops[0]["dims"] == [1, 1]
    assert ops[12]["device_ids"] == [0, 4] and ops[12]["dims"] == [1, 2]         # one per node
    assert ops[15]["device_ids"] == list(range(8))


def test_default_placement_is_round_robin_and_names_are_accepted():
    assert [ff.round_robin_device(i, 8) for i in range(26)] == [i % 8 for i in range(26)]
    s = ff.Strategy()
    s.add_op("Embed_100x16_100", 0, [1, 1], [5])      # reference op name (F9)
    assert s.device_for_table(0, 8, op_name="Embed_100x16_100") == 5
    assert s.device_for_table(0, 8) == 0               # not found -> round robin
    s.add_op("embedding0", 0, [1, 1], [3])             # strategy-file name wins
    assert s.device_for_table(0, 8, op_name="Embed_100x16_100") == 3


def test_balanced_placement():
    # 26 Terabyte-shaped tables, equal lookups: at most ceil(26/8) tables per GPU
    rows = [39884406, 39043, 17289, 7420, 20263, 3, 7120, 1543, 63, 38532951, 2953546, 403346, 10,
            2208, 11938, 155, 4, 976, 14, 39979771, 25641295, 39664984, 585935, 12972, 108, 36]
    owner = ff.balance_placement(rows, [65536] * 26, 128, 8)
    counts = [owner.count(d) for d in range(8)]
    assert max(counts) == 4 and min(counts) == 3
    big = [owner[i] for i, r in enumerate(rows) if r > 20_000_000]
    assert len(set(big)) == len(big)                   # the six 20+ GB tables land on six GPUs
    with pytest.raises(ff.FFEmbError):                 # capacity too small for a 20 GB table
        ff.balance_placement(rows, [65536] * 26, 128, 8, bytes_capacity=1 << 30)
    # skewed lookups: heavy table alone
    owner = ff.balance_placement([10, 10, 10, 10], [100, 1, 1, 1], 4, 2)
    assert owner[0] != owner[1] == owner[2] == owner[3]


# ---- handles and argument checking -------------------------------------------
def _model(batch=8):
    cfg = ff.FFConfig()
    cfg.set_batch_size(batch)
    return cfg, ff.FFModel(cfg)


def test_config_roundtrip_and_parse_args():
    cfg = ff.FFConfig()
    assert cfg.get_batch_size() == 64 and cfg.get_workers_per_node() == 1 and cfg.get_num_nodes() == 1
    cfg.parse_args(["-b", "256", "-ll:gpu", "8", "--nodes", "2", "-e", "5"])
    assert (cfg.get_batch_size(), cfg.get_workers_per_node(), cfg.get_num_nodes()) == (256, 8, 2)
    cfg.set_rank(3, -1)
    assert cfg.get_rank() == 3
    with pytest.raises(ff.FFEmbError):
        cfg.set_batch_size(0)


def test_add_embedding_shape_inference_and_names():
    cfg, m = _model(8)
    t = m.create_tensor([8, 3], ff.DataType.DT_INT64)
    o = m.embedding(t, 1000, 16, ff.AggrMode.AGGR_MODE_AVG, None, ff.UniformInitializer(1, -0.1, 0.1))
    assert o.dims == [8, 16] and o.data_type == ff.DataType.DT_FLOAT
    op0 = m.get_layer_by_id(0)
    assert op0.name == "Embed_1000x16_100"             # model.cc:157-171, guids from 100
    o2 = m.embedding(t, 50, 8, ff.AggrMode.AGGR_MODE_SUM, None, ff.ZeroInitializer())
    assert m.get_layer_by_id(1).name == "Embed_50x8_101" and o2.dims == [8, 8]
    o3 = m.embedding(t, 50, 8, ff.AggrMode.AGGR_MODE_SUM, shared_op=op0)
    assert m.get_layer_by_id(2).name == op0.name       # shared op shares the NAME only
    w = op0.get_weight_tensor()
    assert w.dims == [1000, 16]
    assert op0.get_input_tensor().dims == [8, 3] and op0.get_output_tensor().dims == [8, 16]


def test_add_embedding_rejects_what_the_reference_asserts():
    cfg, m = _model(8)
    t3 = m.create_tensor([8, 2, 2], ff.DataType.DT_INT64)
    with pytest.raises(ff.FFEmbError):                 # embedding.cu:60 numDim == 2
        m.embedding(t3, 10, 4, ff.AggrMode.AGGR_MODE_SUM)
    tf = m.create_tensor([8, 2], ff.DataType.DT_FLOAT)
    with pytest.raises(ff.FFEmbError):
        m.embedding(tf, 10, 4, ff.AggrMode.AGGR_MODE_SUM)
    tb = m.create_tensor([4, 2], ff.DataType.DT_INT64)
    with pytest.raises(ff.FFEmbError):                 # batch mismatch
        m.embedding(tb, 10, 4, ff.AggrMode.AGGR_MODE_SUM)
    t = m.create_tensor([8, 2], ff.DataType.DT_INT64)
    with pytest.raises(ff.FFEmbError):
        m.embedding(t, 0, 4, ff.AggrMode.AGGR_MODE_SUM)
    with pytest.raises(ff.FFEmbError):
        m.embedding(t, 10, 4, ff.AggrMode.AGGR_MODE_NONE)


def test_no_cpu_fallback_without_gpu():
    """On a box without a GPU every compute entry point fails loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    cfg, m = _model(8)
    t = m.create_tensor([8, 1], ff.DataType.DT_INT64)
    m.embedding(t, 10, 4, ff.AggrMode.AGGR_MODE_SUM)
    with pytest.raises(ff.FFEmbError) as e:
        m.compile(ff.SGDOptimizer(m, 0.01))
    assert e.value.status == _lib.FFEMB_ERR_CUDA
    with pytest.raises(ff.FFEmbError):
        m.forward()                                    # not compiled
    from dlrm_flexflow_b200 import kernels
    with pytest.raises(ValueError):
        kernels.make_lookups([torch.zeros(4, 4)], [torch.zeros(2, 1, dtype=torch.int64)])


def test_cpu_placement_is_refused_not_emulated():
    """The reference's hetero strategy (dlrm_strategy_hetero.cc:27-36) puts embeddings on CPUs; this
    library has no CPU path and says so at compile, before touching any device."""
    cfg, m = _model(8)
    t = m.create_tensor([8, 1], ff.DataType.DT_INT64)
    m.embedding(t, 10, 4, ff.AggrMode.AGGR_MODE_SUM)
    s = ff.Strategy()
    s.add_op("embedding0", 1, [1, 1], [0])        # device_type CPU
    m.set_strategy(s)
    with pytest.raises(ff.FFEmbError) as e:
        m.compile(ff.SGDOptimizer(m, 0.01))
    assert e.value.status == _lib.FFEMB_ERR_UNSUPPORTED and "CPU" in str(e.value)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "dlrm_flexflow_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src, f
                assert "libffemb_oracle" not in src and "oracle/" not in src.replace(
                    "oracle/ffemb_oracle.c:ffo_uniform_init", ""), f
