"""End-to-end GPU step: gradients vs PyTorch autograd, graph vs eager, learning, checkpoint variables."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _cfg(**kw):
    from tensorflow_train_distributed_b200.config import TrainConfig
    base = dict(worker_hosts="127.0.0.1:2222", batch_size=128, save_checkpoint_secs=None, seed=11)
    base.update(kw)
    return TrainConfig(**base)


def _torch_reference(vars_, x, y, sizes, act="elu"):
    """fp32 PyTorch model with the SAME weights (TF layout [in,out])."""
    ws, bs = [], []
    names = [f"hidden{i + 1}" for i in range(len(sizes) - 1)] + ["output"]
    h = x
    for i, n in enumerate(names):
        w = torch.tensor(vars_[n + "/kernel"], device="cuda", requires_grad=True)
        b = torch.tensor(vars_[n + "/bias"], device="cuda", requires_grad=True)
        ws.append(w), bs.append(b)
        h = h @ w + b
        if i < len(names) - 1:
            h = torch.nn.functional.elu(h) if act == "elu" else torch.relu(h)
    loss = torch.nn.functional.cross_entropy(h, y)
    loss.backward()
    return loss.item(), (h.argmax(1) == y).float().mean().item(), ws, bs


@pytest.mark.parametrize("act,tail", [("elu", "chain"), ("relu", "chain"), ("elu", "cluster"), ("relu", "none"),
                                      ("elu", "column"), ("relu", "column")])
def test_step_gradients_match_autograd(act, tail):
    """All back ends of the narrow layers: the column kernel (activations resident in shared memory), the single-CTA
    tcgen05 chain kernel, the fp32 cluster kernel and one tcgen05 GEMM launch per layer."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    cfg = _cfg(dropout_rate=0.0, cuda_graph=False, activation=act, fused_tail=tail not in ("none", "column"),
               tail_impl=tail if tail not in ("none", "column") else "chain",
               step_impl="column" if tail == "column" else "layers")
    tr = GpuTrainer(cfg)
    tr.initialize()
    assert tr.model.use_column == (tail == "column")
    ds = synthetic_mnist(1024, seed=5)
    xs, ys = ds.next_batch(128)
    vars_ = tr.export_variables()
    with torch.cuda.stream(tr.stream):
        tr.load_batch(xs, ys)
        tr.wait_batch()
        tr.store.state.view(torch.float32)[4:6] = 0
        tr.model.launch_forward_backward()
    tr.synchronize()
    x, y = torch.tensor(xs, device="cuda"), torch.tensor(ys, device="cuda")
    loss, acc, ws, bs = _torch_reference(vars_, x, y, cfg.layer_sizes, act)
    st = tr.store.read_state()
    assert abs(st.metrics_acc[0] - loss) < 2e-2 * max(1.0, loss)
    assert abs(st.metrics_acc[1] - acc) <= 2.0 / 128
    for i, (w, b) in enumerate(zip(ws, bs)):
        gm = tr.store.grad_matrix(i)
        scale = w.grad.abs().max().item()
        assert torch.allclose(gm[:-1], w.grad, atol=5e-2 * scale), f"kernel grad of layer {i}"
        assert torch.allclose(gm[-1], b.grad, atol=5e-2 * max(b.grad.abs().max().item(), 1e-6)), f"bias grad {i}"
        cos = torch.nn.functional.cosine_similarity(gm[:-1].flatten(), w.grad.flatten(), dim=0).item()
        assert cos > 0.995, (i, cos)
    tr.close()


def test_graph_replay_equals_eager_and_learns():
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    ds1, ds2 = synthetic_mnist(4096, seed=5), synthetic_mnist(4096, seed=5)
    a = GpuTrainer(_cfg(cuda_graph=True))
    b = GpuTrainer(_cfg(cuda_graph=False))
    a.initialize(), b.initialize()
    assert a.launches_per_step == 2   # column kernel (forward + loss + data gradients) + weight gradients with the update
    c = GpuTrainer(_cfg(cuda_graph=False, fused_tail=True))
    c.initialize()
    assert c.launches_per_step == 5   # prep + first-layer fwd + chain kernel + first-layer wgrad + update
    c.close()
    c = GpuTrainer(_cfg(cuda_graph=False, step_impl="layers"))
    c.initialize()
    assert c.launches_per_step == 12  # prep + 5 fwd + 5 bwd + update: every layer is its own GEMM launch
    c.close()
    la = lb = None
    first = None
    for _ in range(60):
        xa, ya = ds1.next_batch(128)
        xb, yb = ds2.next_batch(128)
        la, lb = a.train_step(xa, ya), b.train_step(xb, yb)
        first = first or la.loss
    assert la.global_step == lb.global_step == 60 and la.local_step == 60
    assert abs(la.loss - lb.loss) < 1e-5           # same kernels, same RNG stream: bitwise-equal up to atomics order
    assert la.loss < 0.7 * first and la.accuracy > 0.5
    va, vb = a.export_variables(), b.export_variables()
    assert set(va) == set(vb) and len(va) == 21    # global_step + 10 variables + 10 EMA shadows
    for k in va:
        assert np.allclose(va[k], vb[k], atol=1e-5), k
    assert va["hidden1/kernel"].shape == (784, 200) and va["global_step"] == 60
    a.close(), b.close()


def test_restore_roundtrip_continues_identically():
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    ds = synthetic_mnist(2048, seed=9)
    a = GpuTrainer(_cfg(dropout_rate=0.0))
    a.initialize()
    batches = [ds.next_batch(128) for _ in range(12)]
    for xs, ys in batches[:6]:
        a.train_step(xs, ys)
    snap = a.export_variables()
    b = GpuTrainer(_cfg(dropout_rate=0.0, seed=999))
    b.initialize(restore=snap)
    assert b.global_step() == 6
    for xs, ys in batches[6:]:
        va, vb = a.train_step(xs, ys), b.train_step(xs, ys)
    assert va.global_step == vb.global_step == 12
    assert abs(va.loss - vb.loss) < 1e-5
    a.close(), b.close()


def test_device_resident_dataset_matches_fed_batches():
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    ds = synthetic_mnist(128 * 4, seed=2)
    xs = torch.tensor(ds.images).view(4, 128, 784)
    ys = torch.tensor(ds.labels).view(4, 128)
    a, b = GpuTrainer(_cfg()), GpuTrainer(_cfg())
    a.initialize(), b.initialize()
    a.use_device_dataset(xs.cuda(), ys.cuda())
    for i in range(10):
        with torch.cuda.stream(a.stream):
            a.run_step()
        vb = b.train_step(xs[i % 4], ys[i % 4])
    va = a.read_values()
    assert va.global_step == vb.global_step == 10 and abs(va.loss - vb.loss) < 1e-5
    a.close(), b.close()


def test_wide_mlp_step_runs_and_learns():
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    tr = GpuTrainer(_cfg(model="wide_mlp", activation="relu", regularizer="none"))
    tr.initialize()
    ds = synthetic_mnist(2048, seed=4)
    first = last = None
    for _ in range(30):
        xs, ys = ds.next_batch(128)
        last = tr.train_step(xs, ys)
        first = first or last.loss
    assert last.loss < first and last.global_step == 30
    tr.close()


def _lenet_reference(vars_, x, y, act="elu"):
    """fp32 PyTorch LeNet with the same weights (HWIO kernels -> OIHW)."""
    f = torch.nn.functional
    p = {k: torch.tensor(v, device="cuda", requires_grad=True) for k, v in vars_.items()
         if k != "global_step" and not k.endswith("ExponentialMovingAverage")}
    a = f.elu if act == "elu" else torch.relu
    h = x.view(-1, 1, 28, 28)
    h = f.max_pool2d(a(f.conv2d(h, p["conv1/kernel"].permute(3, 2, 0, 1), p["conv1/bias"], padding=2)), 2)
    h = f.max_pool2d(a(f.conv2d(h, p["conv2/kernel"].permute(3, 2, 0, 1), p["conv2/bias"])), 2)
    h = h.permute(0, 2, 3, 1).reshape(x.shape[0], 400)  # TF flatten of NHWC
    h = a(h @ p["fc1/kernel"] + p["fc1/bias"])
    h = a(h @ p["fc2/kernel"] + p["fc2/bias"])
    logits = h @ p["output/kernel"] + p["output/bias"]
    loss = f.cross_entropy(logits, y)
    loss.backward()
    return loss.item(), p


def test_lenet_gradients_match_autograd_and_it_learns():
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    cfg = _cfg(model="lenet", dropout_rate=0.0, cuda_graph=False)
    tr = GpuTrainer(cfg)
    tr.initialize()
    vars_ = tr.export_variables()
    assert vars_["conv1/kernel"].shape == (5, 5, 1, 6) and vars_["conv2/kernel"].shape == (5, 5, 6, 16)
    assert sum(v.size for k, v in vars_.items() if k != "global_step" and "Exponential" not in k) == 61706
    ds = synthetic_mnist(2048, seed=8)
    xs, ys = ds.next_batch(128)
    with torch.cuda.stream(tr.stream):
        tr.load_batch(xs, ys)
        tr.wait_batch()
        tr.store.state.view(torch.float32)[4:6] = 0
        tr.model.launch_forward_backward()
    tr.synchronize()
    loss, p = _lenet_reference(vars_, torch.tensor(xs, device="cuda"), torch.tensor(ys, device="cuda"))
    st = tr.store.read_state()
    assert abs(st.metrics_acc[0] - loss) < 3e-2 * max(1.0, loss)
    names = ["conv1", "conv2", "fc1", "fc2", "output"]
    for i, n in enumerate(names):
        gm = tr.store.grad_matrix(i)
        wg = p[n + "/kernel"].grad.reshape(gm.shape[0] - 1, gm.shape[1])
        cos = torch.nn.functional.cosine_similarity(gm[:-1].flatten(), wg.flatten(), dim=0).item()
        assert cos > 0.99, (n, cos)
        assert torch.allclose(gm[:-1], wg, atol=6e-2 * wg.abs().max().item()), n
        assert torch.allclose(gm[-1], p[n + "/bias"].grad, atol=6e-2 * max(p[n + "/bias"].grad.abs().max().item(), 1e-6)), n
    tr.close()
    tr = GpuTrainer(_cfg(model="lenet"))
    tr.initialize()
    first = last = None
    for _ in range(80):
        xs, ys = ds.next_batch(128)
        last = tr.train_step(xs, ys)
        first = first or last.loss
    assert last.loss < 0.6 * first and last.global_step == 80
    tr.close()


def test_fused_tail_follows_the_per_layer_path_with_dropout():
    """Same seeds, dropout on: the cluster kernel and the per-layer GEMM path draw identical masks, so the two
    trajectories stay together (fp32 tail vs bf16 GEMMs: small numeric drift only)."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    ds1, ds2 = synthetic_mnist(4096, seed=6), synthetic_mnist(4096, seed=6)
    a = GpuTrainer(_cfg(dropout_rate=0.2, fused_tail=True, tail_impl="cluster"))
    b = GpuTrainer(_cfg(dropout_rate=0.2, fused_tail=False, step_impl="layers"))
    a.initialize(), b.initialize()
    for i in range(40):
        xa, ya = ds1.next_batch(128)
        xb, yb = ds2.next_batch(128)
        va, vb = a.train_step(xa, ya), b.train_step(xb, yb)
        if i == 0:
            assert abs(va.loss - vb.loss) < 2e-2 * max(1.0, vb.loss)
    assert va.global_step == vb.global_step == 40
    assert abs(va.loss - vb.loss) < 0.15 * max(vb.loss, 0.2)
    wa, wb = a.export_variables(), b.export_variables()
    for k in ("hidden1/kernel", "hidden3/kernel", "output/bias"):
        cos = float(np.dot(wa[k].ravel(), wb[k].ravel()) / (np.linalg.norm(wa[k]) * np.linalg.norm(wb[k]) + 1e-12))
        assert cos > 0.999, (k, cos)
    a.close(), b.close()


@pytest.mark.parametrize("batch,hidden", [(128, "200,100,50,25,10"), (100, "203,97,31,10"), (256, "64,10")])
def test_column_kernel_follows_the_per_layer_path_with_dropout(batch, hidden):
    """The column kernel draws the same Philox dropout masks and rounds to bf16 at the same places as one GEMM launch
    per layer; what differs is the summation order inside the softmax and the operand roles of the MMAs. With equal
    seeds the two trajectories therefore agree to rounding noise — on the reference shape, on a ragged one (batch 100,
    widths that are no multiples of 8) and on two clusters' worth of batch."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    hu = [int(x) for x in hidden.split(",")]
    ds1, ds2 = synthetic_mnist(4096, seed=6), synthetic_mnist(4096, seed=6)
    a = GpuTrainer(_cfg(dropout_rate=0.1, batch_size=batch, hidden_units=hu, step_impl="column"))
    b = GpuTrainer(_cfg(dropout_rate=0.1, batch_size=batch, hidden_units=hu, step_impl="layers"))
    a.initialize(), b.initialize()
    assert a.model.use_column and not b.model.use_column and a.launches_per_step == 2
    for i in range(30):
        va, vb = a.train_step(*ds1.next_batch(batch)), b.train_step(*ds2.next_batch(batch))
        if i == 0:
            assert abs(va.loss - vb.loss) < 2e-3 * max(1.0, vb.loss), (va.loss, vb.loss)
            assert abs(va.accuracy - vb.accuracy) <= 1.5 / batch
    assert va.global_step == vb.global_step == 30 and va.local_step == 30
    assert abs(va.loss - vb.loss) < 2e-2 * max(1.0, vb.loss), (va.loss, vb.loss)
    assert abs(va.learning_rate - vb.learning_rate) < 1e-9
    wa, wb = a.export_variables(), b.export_variables()
    assert set(wa) == set(wb)
    for k in wa:
        if k == "global_step":
            continue
        scale = max(float(np.abs(wb[k]).max()), 1e-6)
        assert float(np.abs(wa[k] - wb[k]).max()) < 2e-2 * scale, (k, float(np.abs(wa[k] - wb[k]).max()), scale)
    # the compute copies are what the next forward reads: they must be the bf16 image of the fp32 master
    st = a.store
    for i, g in enumerate(st.layout.groups):
        m = st.master_matrix(i)
        assert torch.equal(st.w_t[i][:, :g.rows], m[:g.rows].t().to(torch.bfloat16))
        if st.w_rm[i] is not None:
            assert torch.equal(st.w_rm[i][:, :g.cols], m[:g.rows].to(torch.bfloat16))
        assert torch.equal(st.bias_c[i][:g.cols], m[g.rows])
    a.close(), b.close()


def test_chain_kernel_is_bitwise_the_per_layer_path():
    """The chain kernel runs the very same tcgen05 pipeline and epilogues as one launch per layer would: with
    identical seeds the two trainers must stay together to the last bit of the loss (atomics aside)."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    ds1, ds2 = synthetic_mnist(4096, seed=6), synthetic_mnist(4096, seed=6)
    a = GpuTrainer(_cfg(dropout_rate=0.1, fused_tail=True, tail_impl="chain"))
    b = GpuTrainer(_cfg(dropout_rate=0.1, fused_tail=False, step_impl="layers"))
    a.initialize(), b.initialize()
    assert a.launches_per_step == 5 and b.launches_per_step == 12
    for _ in range(30):
        xa, ya = ds1.next_batch(128)
        xb, yb = ds2.next_batch(128)
        va, vb = a.train_step(xa, ya), b.train_step(xb, yb)
    assert abs(va.loss - vb.loss) < 1e-5 and va.global_step == vb.global_step == 30
    wa, wb = a.export_variables(), b.export_variables()
    for k in wa:
        assert np.allclose(wa[k], wb[k], atol=1e-6), k
    a.close(), b.close()


def test_net_kernel_is_bitwise_the_per_layer_path():
    """``--fused_tail --tail_impl net``: every GEMM of the step in ONE launch of co-resident CTAs that hand tiles
    over through arrival counters. Same pipeline, same epilogues, same Philox streams: identical to the bit."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    ds1, ds2 = synthetic_mnist(4096, seed=6), synthetic_mnist(4096, seed=6)
    a = GpuTrainer(_cfg(dropout_rate=0.1, fused_tail=True, tail_impl="net"))
    b = GpuTrainer(_cfg(dropout_rate=0.1, fused_tail=False, step_impl="layers"))
    a.initialize(), b.initialize()
    assert a.launches_per_step == 3 and b.launches_per_step == 12  # prep + net + update
    assert a.model._net.ctas == 114  # one CTA per tile for the reference model
    for _ in range(30):
        va, vb = a.train_step(*ds1.next_batch(128)), b.train_step(*ds2.next_batch(128))
    assert abs(va.loss - vb.loss) < 1e-5 and va.global_step == vb.global_step == 30
    wa, wb = a.export_variables(), b.export_variables()
    for k in wa:
        assert np.allclose(wa[k], wb[k], atol=1e-6), k
    a.close(), b.close()


def test_loss_curve_tracks_fp32_torch_sgd():
    """Loss-curve parity (north-star): 150 steps of our bf16 tensor-core pipeline against a plain fp32 PyTorch
    SGD + L2 + staircase-decay loop started from the same weights and fed the same batches."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    cfg = _cfg(dropout_rate=0.0, batch_size=128)
    tr = GpuTrainer(cfg)
    tr.initialize()
    v0 = tr.export_variables()
    names = [f"hidden{i + 1}" for i in range(len(cfg.layer_sizes) - 1)] + ["output"]
    ws = [torch.tensor(v0[n + "/kernel"], device="cuda", requires_grad=True) for n in names]
    bs = [torch.tensor(v0[n + "/bias"], device="cuda", requires_grad=True) for n in names]
    ds1, ds2 = synthetic_mnist(4096, seed=21), synthetic_mnist(4096, seed=21)
    ours, ref = [], []
    for step in range(150):
        xs, ys = ds1.next_batch(128)
        ours.append(tr.train_step(xs, ys).loss)
        xr, yr = ds2.next_batch(128)
        h = torch.tensor(xr, device="cuda")
        for i in range(len(names)):
            h = h @ ws[i] + bs[i]
            if i < len(names) - 1:
                h = torch.nn.functional.elu(h)
        loss = torch.nn.functional.cross_entropy(h, torch.tensor(yr, device="cuda"))
        ref.append(loss.item())
        grads = torch.autograd.grad(loss, ws + bs)
        lr = cfg.initial_learning_rate * cfg.learning_rate_decay_factor ** (step // cfg.decay_steps)
        with torch.no_grad():
            for i, w in enumerate(ws):  # L2 on the hidden kernels only (:55 vs :61), none on biases
                reg = cfg.regularizer_scale * w if i < len(names) - 1 else 0.0
                w -= lr * (grads[i] + reg)
            for i, b in enumerate(bs):
                b -= lr * grads[len(ws) + i]
    ours, ref = np.array(ours), np.array(ref)
    assert ref[-1] < 0.5 * ref[0]  # the reference run itself learned
    assert np.max(np.abs(ours - ref)) < 0.05 * max(1.0, ref[0]), (ours[-5:], ref[-5:])
    assert abs(ours[-10:].mean() - ref[-10:].mean()) < 0.03
    vf = tr.export_variables()
    for i, n in enumerate(names):  # weights stay together too (bf16 operands, fp32 master)
        d = np.abs(vf[n + "/kernel"] - ws[i].detach().cpu().numpy()).max()
        assert d < 0.02, (n, d)
    tr.close()


def test_fetch_lag_pipelines_the_readback():
    """--fetch_lag L: run() hands back step s-L from the pinned ring the update kernel writes; same trajectory."""
    from tensorflow_train_distributed_b200.data import synthetic_mnist
    from tensorflow_train_distributed_b200.parallel.gpu_trainer import GpuTrainer
    L = 4
    a, b = GpuTrainer(_cfg(fetch_lag=0)), GpuTrainer(_cfg(fetch_lag=L))
    a.initialize(), b.initialize()
    ds1, ds2 = synthetic_mnist(4096, seed=9), synthetic_mnist(4096, seed=9)
    exact, lagged = [], []
    for _ in range(40):
        exact.append(a.train_step(*ds1.next_batch(128)))
        lagged.append(b.train_step(*ds2.next_batch(128)))
    assert [v.local_step for v in exact] == list(range(1, 41))
    assert all(v.local_step == 0 for v in lagged[:L])          # nothing has landed yet: the initial state
    for i in range(L, 40):
        assert lagged[i].local_step == i - L + 1 and lagged[i].global_step == i - L + 1
        assert abs(lagged[i].loss - exact[i - L].loss) < 1e-6
    last = b.drain()
    assert last.local_step == 40 and abs(last.loss - exact[-1].loss) < 1e-6
    assert b.drain() is None
    va, vb = a.export_variables(), b.export_variables()
    for k in va:
        assert np.allclose(va[k], vb[k], atol=1e-6), k
    a.close(), b.close()
