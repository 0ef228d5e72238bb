"""Full-step parity: N training steps of the fused sm_100a engine against the plain PyTorch
fp32 model (same architecture/initialisation as the reference, mnist_onegpu.py:11-49) -
loss curve, gradients, updated parameters, BatchNorm buffers and checkpoint round trip."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


@pytest.fixture(scope="module", autouse=True)
def _no_tf32():
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False


class _Q(torch.autograd.Function):
    """bf16 storage point: value rounded in forward; gradient rounded in backward iff the
    engine materialises that gradient in bf16 (dp2, dx2, dp1 - not the conv1 output grad)."""

    @staticmethod
    def forward(ctx, x, round_grad):
        ctx.round_grad = round_grad
        return x.to(torch.bfloat16).float()

    @staticmethod
    def backward(ctx, g):
        return (g.to(torch.bfloat16).float() if ctx.round_grad else g), None


def _emulated_forward(m, x):
    """The reference architecture with the engine's bf16 storage points made explicit."""
    l1, l2 = m.layer1, m.layer2
    x1 = _Q.apply(l1[0](x), False)
    p1 = _Q.apply(F.max_pool2d(F.relu(l1[1](x1)), 2), True)
    w2 = _Q.apply(l2[0].weight, False)
    x2 = _Q.apply(F.conv2d(p1, w2, l2[0].bias, padding=2), True)
    p2 = _Q.apply(F.max_pool2d(F.relu(l2[1](x2)), 2), True)
    return m.fc(p2.flatten(1))


def _run_pair(shape, B, steps, lr, backend, fuse, emulate=True, precision="bf16"):
    from torch_distributed_sandbox_b200.models import reference_init
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    ref = reference_init(shape, seed=0).to(DEV)
    eng = FusedConvNet(shape, batch_size=B, device=DEV, lr=lr, init="empty", conv_backend=backend,
                       fuse_fc_sgd=fuse, precision=precision)
    eng.load_state_dict(ref.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr)
    # MNIST-shaped synthetic digits with class-dependent content (so gradients carry signal instead of
    # being zero-mean noise), resized on the GPU by the engine's own PIL-exact kernel
    from torch_distributed_sandbox_b200.data.mnist import synthetic_digits
    imgs, labs = synthetic_digits(B * steps, seed=5)
    imgs = torch.from_numpy(imgs).view(steps, B, 28, 28).to(DEV)
    labs = torch.from_numpy(labs.astype("int64")).view(steps, B).to(DEV)
    out = []
    for s in range(steps):
        eng.set_input(imgs[s])
        x = eng.x[:, None].clone()
        y = labs[s]
        loss_ref = F.cross_entropy(_emulated_forward(ref, x) if emulate else ref(x), y)
        opt.zero_grad()
        loss_ref.backward()
        grads_ref = {k: p.grad.clone() for k, p in ref.named_parameters()}
        opt.step()
        eng.forward()
        loss = eng.loss_backward(y).clone()
        grads = eng.named_gradients()
        eng.step()
        out.append((float(loss), float(loss_ref), grads, grads_ref))
    return ref, eng, out


@pytest.mark.parametrize("backend", ["torch", "tc"])
def test_engine_matches_reference_model(backend):
    """Against the fp32 model with the engine's bf16 storage points emulated: tight bands."""
    shape, B = (64, 96), 3
    ref, eng, out = _run_pair(shape, B, steps=3, lr=1e-3, backend=backend, fuse=False)   # stable regime
    for s, (loss, loss_ref, grads, grads_ref) in enumerate(out):
        print(f"[engine/{backend}] step {s}: loss {loss:.6f} ref {loss_ref:.6f}")
        assert abs(loss - loss_ref) < 2e-3 * max(1.0, abs(loss_ref))
        if s > 0:
            continue   # later steps: trajectories of two bf16 pipelines drift (pool arg-max flips feed back
                       # through BN and SGD), so only the loss is compared; step 0 checks every kernel exactly
        for k, gref in grads_ref.items():
            if k in ("layer1.0.bias", "layer2.0.bias"):
                continue    # analytically zero (bias before BatchNorm); torch returns round-off
            e = rel(grads[k], gref)
            print(f"[engine/{backend}]   grad {k}: rel-L2 {e:.3e}")
            tol = 2e-2 if k == "layer1.0.weight" else 8e-3
            assert e < tol, f"step {s} grad {k}: {e}"
    sd, sd_ref = eng.state_dict(), ref.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    for k in sd_ref:
        assert sd[k].shape == sd_ref[k].shape and sd[k].dtype == sd_ref[k].dtype, k
        if k.endswith("num_batches_tracked"):
            assert int(sd[k]) == int(sd_ref[k])
        else:
            e = rel(sd[k], sd_ref[k])
            print(f"[engine/{backend}]   state {k}: rel-L2 {e:.3e}")
            # BN biases start at 0, so their state IS the accumulated gradient of drifting trajectories
            assert e < (1e-1 if k.endswith("1.bias") else 2e-2), f"state {k}: {e}"


def test_engine_tracks_pure_fp32_model():
    """Against the un-emulated fp32 model (what the reference computes, TF32 off): the loss
    curve agrees to bf16 precision and the weights stay close after a few SGD steps.  Per-step
    gradients of the conv layers are NOT compared here: on random data they are noise-like
    sums in which a handful of pool arg-max flips (bf16 vs fp32 activations) dominate."""
    shape, B = (64, 96), 4
    ref, eng, out = _run_pair(shape, B, steps=4, lr=1e-3, backend="tc", fuse=True, emulate=False)
    for s, (loss, loss_ref, grads, grads_ref) in enumerate(out):
        print(f"[engine/fp32] step {s}: loss {loss:.6f} ref {loss_ref:.6f}")
        assert abs(loss - loss_ref) < 1e-2 * max(1.0, abs(loss_ref))
    sd, sd_ref = eng.state_dict(), ref.state_dict()
    for k in ("fc.weight", "fc.bias", "layer2.0.weight", "layer1.0.weight", "layer2.1.running_var"):
        e = rel(sd[k], sd_ref[k])
        print(f"[engine/fp32]   state {k}: rel-L2 {e:.3e}")
        assert e < 2e-2, f"state {k}: {e}"


@pytest.mark.parametrize("conv1_fwd", ["vt", "hz"])
def test_tf32_engine_matches_pure_fp32_model(conv1_fwd, monkeypatch):
    """precision="tf32" (fp32 storage, TF32 conv2 math: the reference's numerics) against the UN-emulated fp32 model
    with TF32 off: every gradient of step 0, the loss curve and the final state, with nothing in the oracle rounded
    to match.  The bands are TF32's 10-bit operand mantissa, an order of magnitude tighter than the bf16 mode's."""
    shape, B = (64, 96), 4
    monkeypatch.setenv("B200DP_CONV1_FWD", conv1_fwd)     # "hz": layer-1 forward on kind::tf32 as well (opt-in)
    ref, eng, out = _run_pair(shape, B, steps=4, lr=1e-3, backend="tc", fuse=True, emulate=False, precision="tf32")
    assert eng.p1.dtype == torch.float32 and eng.p2.dtype == torch.float32
    assert eng.conv1_row_operand == (conv1_fwd == "hz")
    for s, (loss, loss_ref, grads, grads_ref) in enumerate(out):
        print(f"[engine/tf32] step {s}: loss {loss:.6f} ref {loss_ref:.6f}")
        # step 0 compares the same weights; afterwards two slightly different gradient streams drive the weights apart
        assert abs(loss - loss_ref) < (5e-4 if s == 0 else 5e-3) * max(1.0, abs(loss_ref))
        if s > 0:
            continue
        errs = {}
        for k, gref in grads_ref.items():
            if k in ("layer1.0.bias", "layer2.0.bias", "fc.weight"):
                continue    # zero by construction / fused into the SGD update in this configuration
            errs[k] = rel(grads[k], gref)
            print(f"[engine/tf32]   grad {k}: rel-L2 {errs[k]:.3e}")
        for k, e in errs.items():
            # Parameters whose gradient is a BatchNorm-backward sum with heavy cancellation (conv weights, BN-1, BN-2
            # beta) amplify the 2^-11 TF32 operand rounding to ~1e-2: torch's own cuDNN TF32 path shows 1e-2 .. 6e-2
            # on the same quantities at this shape (profiles/precision_error_vs_cudnn_tf32.txt); the rest is tight.
            # "hz": layer 1 runs kind::tf32 as well (csrc/conv1_hz.cu, input truncated to tf32 as in cuDNN's TF32 conv);
            # cuDNN's TF32 path shows 2.4e-2 .. 5.3e-2 on the layer-1 gradients at this shape.
            loose = k.startswith("layer1") or k in ("layer2.0.weight", "layer2.1.bias")
            band = 6e-2 if (conv1_fwd == "hz" and k.startswith("layer1")) else 3e-2
            assert e < (band if loose else 2e-3), f"grad {k}: {e}"
    sd, sd_ref = eng.state_dict(), ref.state_dict()
    for k in sd_ref:
        if k.endswith("num_batches_tracked"):
            assert int(sd[k]) == int(sd_ref[k])
            continue
        e = rel(sd[k], sd_ref[k])
        print(f"[engine/tf32]   state {k}: rel-L2 {e:.3e}")
        # BN biases start at 0: their state IS the sum of four gradients from slightly different trajectories
        assert e < (2e-1 if k.endswith("1.bias") else 5e-3), f"state {k}: {e}"


def test_autograd_front_end_matches_train_step():
    """model(images) -> criterion -> backward -> optimizer.step (the reference's loop idiom)
    produces the same update as the fused train_step."""
    from torch_distributed_sandbox_b200.models.module import FusedModule, SGD
    shape, B = (32, 64), 3
    g = torch.Generator(device=DEV).manual_seed(5)
    x = torch.rand(B, 1, *shape, device=DEV, generator=g)
    y = torch.randint(0, 10, (B,), device=DEV, generator=g)
    m1 = FusedModule(shape, batch_size=B, device=DEV, lr=1e-2)
    m2 = FusedModule(shape, batch_size=B, device=DEV, lr=1e-2)
    opt = SGD(m1, lr=1e-2)
    crit = torch.nn.CrossEntropyLoss()
    for it in range(2):
        loss1 = crit(m1(x), y)
        opt.zero_grad()
        loss1.backward()
        opt.step()
        loss2 = m2.train_step(x, y)
        # iteration 0 runs identical kernels on identical weights; afterwards torch's CE backward vs our
        # fused one differ in the last bit and bf16 rounding of activations amplifies that a little
        assert abs(float(loss1.detach()) - float(loss2)) < (1e-6 if it == 0 else 1e-3)
    assert rel(m1.engine.fcw, m2.engine.fcw) < 1e-3 and rel(m1.engine.small, m2.engine.small) < 1e-3
    assert list(m1.state_dict().keys())[0] == "layer1.0.weight"


def test_fused_fc_sgd_equals_unfused():
    shape, B = (32, 64), 5
    _, e1, _ = _run_pair(shape, B, steps=2, lr=1e-2, backend="tc", fuse=False)
    _, e2, _ = _run_pair(shape, B, steps=2, lr=1e-2, backend="tc", fuse=True)
    assert rel(e2.fcw, e1.fcw) < 1e-6
    assert rel(e2.small, e1.small) < 1e-6


def test_checkpoint_roundtrip_into_torch_model():
    from torch_distributed_sandbox_b200.models import ConvNet
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    shape = (32, 48)
    eng = FusedConvNet(shape, batch_size=2, device=DEV, init="reference", seed=0)
    model = ConvNet(shape)
    model.load_state_dict({k: v.cpu() for k, v in eng.state_dict().items()})   # strict
    eng2 = FusedConvNet(shape, batch_size=2, device=DEV, init="empty")
    eng2.load_state_dict({"module." + k: v for k, v in model.state_dict().items()})
    assert torch.equal(eng2.fcw, eng.fcw) and torch.equal(eng2.small, eng.small)


def test_uint8_input_path_and_batch_gt_8():
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    shape, B = (64, 64), 11
    eng = FusedConvNet(shape, batch_size=B, device=DEV, lr=1e-3, init="reference")
    g = torch.Generator().manual_seed(0)
    imgs = torch.randint(0, 256, (B, 28, 28), dtype=torch.uint8, generator=g)
    labels = torch.randint(0, 10, (B,), generator=g)
    l0 = float(eng.train_step(imgs.to(DEV), labels.to(DEV)))
    for _ in range(20):
        l1 = float(eng.train_step(imgs.to(DEV), labels.to(DEV)))
    print(f"[engine] overfit one batch: {l0:.4f} -> {l1:.4f}")
    assert l1 < l0


@pytest.mark.parametrize("precision", ["bf16", "tf32"])
def test_short_last_batch_equals_an_engine_of_that_size(precision):
    """drop_last=False (mnist_onegpu.py:55-59): a batch smaller than the engine's capacity runs on leading-dimension
    views of the same buffers and must give exactly what an engine built for that batch size gives."""
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    shape = (64, 96)
    big = FusedConvNet(shape, batch_size=5, device=DEV, lr=1e-2, precision=precision)
    small = FusedConvNet(shape, batch_size=3, device=DEV, lr=1e-2, precision=precision)
    g = torch.Generator().manual_seed(3)
    full = torch.randint(0, 256, (5, 28, 28), dtype=torch.uint8, generator=g).to(DEV)
    lab = torch.randint(0, 10, (5,), generator=g).to(DEV)
    big.train_step(full, lab)                      # leaves stale rows 3..4 in every activation buffer
    small.load_state_dict(big.state_dict())
    l_big = float(big.train_step(full[:3], lab[:3]))
    l_small = float(small.train_step(full[:3], lab[:3]))
    assert l_big == l_small
    assert torch.equal(big.fcw, small.fcw) and torch.equal(big.small, small.small)
    assert torch.equal(big.bn_flat, small.bn_flat)


def test_checkpoint_loads_into_the_installed_reference_convnet(tmp_path):
    """a checkpoint written by the engine must load (strict) into the UNMODIFIED reference's own
    ``ConvNet`` (baseline/_ref/mnist_onegpu.py:11-31) after its lazy-init forward (:36-39), and that model must then
    compute the same logits as the engine.  The class is imported from the installed reference, not re-stated."""
    import importlib.util
    import os
    ref_py = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref", "mnist_onegpu.py")
    if not os.path.exists(ref_py):
        pytest.skip("baseline/_ref is not installed (run baseline/install_ref.sh)")
    spec = importlib.util.spec_from_file_location("ref_mnist_onegpu", ref_py)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    shape, B = (32, 48), 2
    eng = FusedConvNet(shape, batch_size=B, device=DEV, lr=1e-2, precision="tf32")
    g = torch.Generator().manual_seed(9)
    x = torch.rand(B, 1, *shape, generator=g)
    y = torch.randint(0, 10, (B,), generator=g)
    for _ in range(2):
        eng.train_step(x.to(DEV), y.to(DEV))
    path = str(tmp_path / "ck.pt")
    torch.save(eng.state_dict(), path)
    torch.manual_seed(0)
    model = ref.ConvNet()
    model(torch.zeros([1, 1, *shape]))                       # the reference's lazy init (mnist_onegpu.py:39)
    model.load_state_dict(torch.load(path, map_location="cpu"))      # strict
    model = model.to(DEV).train()
    want = model(x.to(DEV))                                  # train mode: batch statistics, like the engine
    got = eng.forward(x.to(DEV))
    assert rel(got, want) < 2e-3, rel(got, want)


def test_graphed_step_equals_eager_step(monkeypatch):
    """The single-GPU training step replayed as one CUDA graph (engine._train_step_graphed) is bit-identical to the same
    steps launched one kernel at a time, counts its kernels, and survives a short batch in between."""
    from torch_distributed_sandbox_b200 import ops
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    from torch_distributed_sandbox_b200.data.mnist import synthetic_digits
    shape, B, steps = (64, 96), 4, 7
    imgs, labs = synthetic_digits(B * steps, seed=9)
    imgs = torch.from_numpy(imgs).view(steps, B, 28, 28).to(DEV)
    labs = torch.from_numpy(labs.astype("int64")).view(steps, B).to(DEV)
    out = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("B200DP_CUDA_GRAPH", mode)
        eng = FusedConvNet(shape, batch_size=B, device=DEV, lr=1e-3, seed=0, precision="tf32")
        losses, l0 = [], ops.LAUNCHES
        for s in range(steps):
            if s == 4:      # a short batch in the middle: eager (its own key has not warmed up), same buffers
                losses.append(float(eng.train_step(imgs[s, :3], labs[s, :3])))
            else:
                losses.append(float(eng.train_step(imgs[s], labs[s])))
        out[mode] = (losses, {k: v.clone() for k, v in eng.state_dict().items()}, ops.LAUNCHES - l0,
                     sum(1 for g in eng._graphs.values() if g[0] is not None))
    assert out["1"][3] == 1 and out["0"][3] == 0          # one captured graph (the full batch), none when disabled
    assert out["0"][0] == out["1"][0], (out["0"][0], out["1"][0])
    for k, v in out["0"][1].items():
        assert torch.equal(v, out["1"][1][k]), k
    assert out["0"][2] == out["1"][2] > 0                 # replayed kernels are counted
