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


def _run_pair(shape, B, steps, lr, backend, fuse):
    from torch_distributed_sandbox_b200.models import reference_init
    from torch_distributed_sandbox_b200.models.engine import FusedConvNet
    ref = reference_init(shape, seed=0).to(DEV)
    eng = FusedConvNet(shape, batch_size=B, device=DEV, lr=lr, init="empty", conv_backend=backend,
                       fuse_fc_sgd=fuse)
    eng.load_state_dict(ref.state_dict())
    opt = torch.optim.SGD(ref.parameters(), lr)
    g = torch.Generator(device=DEV).manual_seed(123)
    out = []
    for s in range(steps):
        x = torch.rand(B, 1, *shape, device=DEV, generator=g)
        y = torch.randint(0, 10, (B,), device=DEV, generator=g)
        loss_ref = F.cross_entropy(ref(x), y)
        opt.zero_grad()
        loss_ref.backward()
        grads_ref = {k: p.grad.clone() for k, p in ref.named_parameters()}
        opt.step()
        eng.forward(x)
        loss = eng.loss_backward(y).clone()
        grads = eng.named_gradients()
        eng.step()
        out.append((float(loss), float(loss_ref), grads, grads_ref))
    return ref, eng, out


@pytest.mark.parametrize("backend", ["torch", "tc"])
def test_engine_matches_reference_model(backend):
    shape, B = (64, 96), 3
    ref, eng, out = _run_pair(shape, B, steps=3, lr=1e-2, backend=backend, fuse=False)
    for s, (loss, loss_ref, grads, grads_ref) in enumerate(out):
        print(f"[engine/{backend}] step {s}: loss {loss:.6f} ref {loss_ref:.6f}")
        assert abs(loss - loss_ref) < 2e-2 * max(1.0, abs(loss_ref))
        for k, gref in grads_ref.items():
            if k in ("layer1.0.bias", "layer2.0.bias"):
                continue    # analytically zero (bias before BatchNorm); torch returns round-off
            e = rel(grads[k], gref)
            print(f"[engine/{backend}]   grad {k}: rel-L2 {e:.3e}")
            assert e < 6e-2, f"step {s} grad {k}: {e}"
    sd, sd_ref = eng.state_dict(), ref.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys())
    for k in sd_ref:
        assert sd[k].shape == sd_ref[k].shape and sd[k].dtype == sd_ref[k].dtype, k
        if k.endswith("num_batches_tracked"):
            assert int(sd[k]) == int(sd_ref[k])
        else:
            e = rel(sd[k], sd_ref[k])
            print(f"[engine/{backend}]   state {k}: rel-L2 {e:.3e}")
            assert e < 2e-2, f"state {k}: {e}"


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
