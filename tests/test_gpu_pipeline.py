"""The library's double-buffered host pipeline (ffemb_model_load_batch_async /
begin_batch / fetch_outputs_async / wait_outputs -- what the reference's
DataLoader::next_batch + zero-copy outputs do, dlrm.cc:486-524,
embedding.cu:280-284): three steps on three different batches fed from pinned
host memory must give the same pooled outputs (bit for bit) and the same tables
(bit for bit) as the same three steps fed through the plain tensor calls, and
both must match the oracle."""
import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

ROWS = [50, 3000, 700, 120000]
D, B, L, LR = 32, 384, 3, 0.02


def _model(ff, engine):
    cfg = ff.FFConfig()
    cfg.set_batch_size(B)
    if engine:
        cfg.set_exchange_engine(True)
        cfg.set_replicate_below(100)
    m = ff.FFModel(cfg)
    ins, outs = [], []
    for t, n in enumerate(ROWS):
        x = m.create_tensor([B, L], ff.DataType.DT_INT64)
        outs.append(m.embedding(x, n, D, ff.AggrMode.AGGR_MODE_SUM, None, ff.UniformInitializer(7 + t, -0.1, 0.1)))
        ins.append(x)
    m.compile(optimizer=ff.SGDOptimizer(m, LR))
    return m, ins, outs


@pytest.mark.parametrize("engine", [False, True])
def test_pipeline_equals_plain_calls_and_oracle(oracle, engine):
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    import dlrm_flexflow_b200 as ff
    T = len(ROWS)
    rng = np.random.default_rng(11)
    batches = [([rng.integers(0, n, size=(B, L)) for n in ROWS], rng.normal(size=(B, T * D)).astype(np.float32))
               for _ in range(3)]
    # plain path
    m, ins, outs = _model(ff, engine)
    W0 = [m.get_layer_by_id(t).get_weight_tensor().get_weights(m) for t in range(T)]
    plain_out = []
    for idx, g in batches:
        for t in range(T):
            ins[t].set_array(m, idx[t])
        m.forward()
        m.zero_gradients()
        for t in range(T):
            outs[t].set_grad(m, np.ascontiguousarray(g[:, t * D:(t + 1) * D]))
        m.backward()
        m.update()
        m.sync()
        plain_out.append(np.concatenate([outs[t].get_array(m) for t in range(T)], axis=1))
    plain_W = [m.get_layer_by_id(t).get_weight_tensor().get_weights(m) for t in range(T)]
    m.close()
    # pipelined path: batch i+1 is loaded while batch i computes; outputs are read one step late
    m, ins, outs = _model(ff, engine)
    hidx = [[torch.from_numpy(i).pin_memory() for i in idx] for idx, _ in batches]
    hg = [torch.from_numpy(g).pin_memory() for _, g in batches]
    hout = [torch.empty((B, T * D), pin_memory=True) for _ in batches]
    m.load_batch_async(hidx[0], hg[0])
    for i in range(3):
        m.begin_batch()
        m.forward()
        m.fetch_outputs_async(hout[i])
        m.zero_gradients()
        m.backward()
        m.update()
        if i + 1 < 3:
            m.load_batch_async(hidx[i + 1], hg[i + 1])
        m.wait_outputs()
        if i > 0:
            assert hout[i - 1].numpy().tobytes() == plain_out[i - 1].tobytes()
    m.wait_outputs(all=True)
    assert hout[2].numpy().tobytes() == plain_out[2].tobytes()
    m.sync()
    for t in range(T):
        got = m.get_layer_by_id(t).get_weight_tensor().get_weights(m)
        assert got.tobytes() == plain_W[t].tobytes()
    m.close()
    # and the oracle
    W = [w.copy() for w in W0]
    for i, (idx, g) in enumerate(batches):
        want = []
        for t in range(T):
            o, W[t] = oracle.step(W[t], idx[t], np.ascontiguousarray(g[:, t * D:(t + 1) * D]), LR)
            want.append(o)
        if i == 0:
            assert plain_out[0].tobytes() == np.concatenate(want, axis=1).tobytes()
    for t in range(T):
        np.testing.assert_allclose(plain_W[t], W[t], rtol=1e-5, atol=1e-5)
