"""CPU oracles for the non-kernel components: sampler index parity with torch, bucket plan
parity with torch's Reducer planner, model/state_dict parity with the reference's lazy
initialisation, PIL-exact resize tables, idx dataset round trip."""
import numpy as np
import pytest
import torch
import torch.distributed as tdist
from torch import nn

from torch_distributed_sandbox_b200.models import ConvNet, STATE_KEYS, feature_size, reference_init
from torch_distributed_sandbox_b200.parallel import DistributedSampler, plan_buckets
from torch_distributed_sandbox_b200.parallel.buckets import FIRST_BUCKET_BYTES


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("n,ws", [(60000, 1), (60000, 2), (60000, 4), (60000, 8), (1003, 4), (7, 8)])
@pytest.mark.parametrize("shuffle", [True, False])
def test_sampler_indices_bit_exact_vs_torch(n, ws, shuffle):
    from torch.utils.data.distributed import DistributedSampler as TorchSampler
    for rank in range(ws):
        for epoch in (0, 3):
            ours = DistributedSampler(_DS(n), num_replicas=ws, rank=rank, shuffle=shuffle)
            ref = TorchSampler(_DS(n), num_replicas=ws, rank=rank, shuffle=shuffle)
            ours.set_epoch(epoch)
            ref.set_epoch(epoch)
            assert len(ours) == len(ref)
            assert list(ours) == list(ref)


def test_sampler_survey_oracle_and_drop_last():
    s = DistributedSampler(_DS(60000), num_replicas=2, rank=0)
    assert list(s)[:4] == [36044, 57327, 21567, 9481]
    assert [len(DistributedSampler(_DS(60000), num_replicas=w, rank=0)) for w in (1, 2, 4, 8)] == \
        [60000, 30000, 15000, 7500]
    from torch.utils.data.distributed import DistributedSampler as TorchSampler
    a = DistributedSampler(_DS(1003), num_replicas=4, rank=3, drop_last=True)
    b = TorchSampler(_DS(1003), num_replicas=4, rank=3, drop_last=True)
    assert list(a) == list(b)
    with pytest.raises(ValueError):
        DistributedSampler(_DS(10), num_replicas=2, rank=2)


def test_bucket_plan_matches_torch_reducer():
    shape = (3000, 3000)
    with torch.device("meta"):
        model = ConvNet(shape)
    params = list(model.parameters())
    cap = 25 * 1024 * 1024
    # iteration 1: one bucket (mnist_distributed.py:67 defaults -> [sys.maxsize])
    first = plan_buckets(list(reversed(params)), [1 << 62])
    assert len(first) == 1 and sum(params[i].numel() for i in range(len(params))) * 4 == 720_053_416
    # steady state: gradient-arrival order (reverse of registration), caps [1 MiB, 25 MiB]
    arrival = [params[i] for i in (9, 8, 7, 6, 5, 4, 3, 2, 1, 0)]
    # fc.bias is accumulated before fc.weight by autograd (40 B + 720 MB)
    arrival = [params[9], params[8]] + [params[i] for i in (7, 6, 5, 4, 3, 2, 1, 0)]
    ours = plan_buckets(arrival, [FIRST_BUCKET_BYTES, cap])
    sizes = [sum(arrival[i].numel() * 4 for i in b) for b in ours]
    assert sizes == [720_000_040, 53_376]
    real = [torch.empty(0) for _ in arrival]  # torch's planner needs real tensors: use small proxies
    small = [nn.Parameter(torch.empty(min(p.numel(), 300_000))) for p in arrival]
    ref_idx, _ = tdist._compute_bucket_assignment_by_size(small, [FIRST_BUCKET_BYTES, cap])
    ours_small = plan_buckets(small, [FIRST_BUCKET_BYTES, cap])
    assert [list(b) for b in ref_idx] == ours_small
    # random mixed sizes / dtypes
    g = torch.Generator().manual_seed(0)
    ts = []
    for i in range(40):
        n = int(torch.randint(1, 400_000, (1,), generator=g))
        ts.append(torch.empty(n, dtype=torch.float32 if i % 5 else torch.float64))
    ref_idx, _ = tdist._compute_bucket_assignment_by_size(ts, [FIRST_BUCKET_BYTES, 2 * 1024 * 1024])
    assert sorted(map(tuple, ref_idx)) == sorted(map(tuple, plan_buckets(ts, [FIRST_BUCKET_BYTES, 2 * 1024 * 1024])))


class _LazyRef(nn.Module):
    """The reference's construction order (conv stages, then a LazyLinear head materialised by
    a train-mode forward on zeros) restated for the parity test."""

    def __init__(self):
        super().__init__()
        mk = lambda a, b: nn.Sequential(nn.Conv2d(a, b, 5, 1, 2), nn.BatchNorm2d(b), nn.ReLU(), nn.MaxPool2d(2, 2))
        self.layer1, self.layer2, self.fc = mk(1, 16), mk(16, 32), nn.LazyLinear(10)

    def forward(self, x):
        o = self.layer2(self.layer1(x))
        return self.fc(o.reshape(o.size(0), -1))


@pytest.mark.parametrize("shape", [(32, 32), (64, 48)])
def test_reference_init_reproduces_lazy_init_and_bn_side_effect(shape):
    torch.manual_seed(0)
    lazy = _LazyRef()
    lazy(torch.zeros([1, 1] + list(shape)))
    ours = reference_init(shape, seed=0)
    sd_ref, sd = lazy.state_dict(), ours.state_dict()
    assert list(sd.keys()) == list(sd_ref.keys()) == list(STATE_KEYS)
    for k in sd_ref:
        assert sd[k].dtype == sd_ref[k].dtype and sd[k].shape == sd_ref[k].shape, k
        if k.endswith("running_mean"):
            assert torch.allclose(sd[k], sd_ref[k], atol=1e-7), k
        else:
            assert torch.equal(sd[k], sd_ref[k]), k
    assert int(sd["layer1.1.num_batches_tracked"]) == 1 and float(sd["layer2.1.running_var"][0]) == pytest.approx(0.9)
    lazy.load_state_dict(sd)                      # strict load into the reference-style module
    assert feature_size((3000, 3000)) == 18_000_000


def test_resize_tables_match_pillow_and_torchvision():
    from PIL import Image
    import torchvision.transforms as T
    from torch_distributed_sandbox_b200.data.resize import resize_u8_reference
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (28, 28), dtype=np.uint8)
    for hw in [(3000, 3000), (500, 500), (64, 96), (28, 28)]:
        ref = np.asarray(Image.fromarray(a, mode="L").resize((hw[1], hw[0]), Image.BILINEAR))
        assert np.array_equal(ref, resize_u8_reference(a, hw)), hw
    tv = T.Compose([T.Resize([300, 300]), T.ToTensor()])(Image.fromarray(a, mode="L"))[0]
    assert torch.equal(tv, torch.from_numpy(resize_u8_reference(a, (300, 300))).float() / 255.0)


def test_trace_step_timer_and_nvtx_noop():
    """step timer (host clock without a GPU) and NVTX ranges that cost nothing when disabled."""
    import time
    from torch_distributed_sandbox_b200.utils import trace
    t = trace.StepTimer(skip=1)
    for _ in range(4):
        with t.step():
            time.sleep(0.002)
    s = t.summary()
    assert s["steps"] == 3 and s["min_ms"] >= 1.5 and s["max_ms"] >= s["mean_ms"] >= s["min_ms"]
    assert not trace.nvtx_enabled()
    with trace.nvtx_range("x"):
        pass
    trace.enable_nvtx(True)
    try:
        with trace.nvtx_range("y"):      # no CUDA here: still a no-op, must not raise
            pass
    finally:
        trace.enable_nvtx(False)
