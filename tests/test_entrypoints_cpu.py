"""Entrypoint / CLI parity on the CPU plumbing configuration: flag names and
defaults of the four reference scripts, the toy's printed line format, DDP on the host
backend against torch's own DDP on gloo (the oracle of SURVEY.md section 4, item 2)."""
import os
import re
import subprocess
import sys

import pytest
import torch

import torch_distributed_sandbox_b200 as b2
from torch_distributed_sandbox_b200 import distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "examples"))


def test_cli_flag_surface_matches_reference():
    import mnist_distributed as md
    import mnist_onegpu as mo
    a = md.parse([])
    assert (a.nodes, a.gpus, a.nr, a.epochs, a.world_size) == (1, 1, 0, 2, 1)
    a = md.parse(["-n", "2", "-g", "4", "-nr", "1", "--epochs", "3"])
    assert (a.nodes, a.gpus, a.nr, a.epochs, a.world_size) == (2, 4, 1, 3, 8)
    assert mo.parse([]).epochs == 2 and mo.parse(["--epochs", "5"]).epochs == 5
    assert md.IMAGE_SHAPE == [3000, 3000] and mo.IMAGE_SHAPE == [3000, 3000]


def test_allreduce_toy_cpu_world2_prints_reference_lines():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "allreduce_toy.py"), "-s", "2", "--backend",
                          "gloo", "--steps", "4"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = re.findall(r"rank: (\d), step: (\d+), value: (\d+), reduced sum: (\d+)\.", out.stdout)
    assert len(lines) == 8
    by_step = {}
    for r, s, v, tot in lines:
        by_step.setdefault(int(s), []).append((int(v), int(tot)))
    for s, pairs in by_step.items():
        assert len(pairs) == 2 and pairs[0][1] == pairs[1][1] == pairs[0][0] + pairs[1][0]


def _ddp_vs_torch(rank, world, port, port2, out_dir):
    import torch.distributed as tdist
    from torch import nn
    from torch_distributed_sandbox_b200.models import reference_init
    from torch_distributed_sandbox_b200.parallel import DistributedDataParallel
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", port
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tdist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port2}", rank=rank, world_size=world)
    shape, B, steps = (32, 32), 3, 3
    torch.manual_seed(100 + rank)                               # deliberately different initial weights
    ours = DistributedDataParallel(reference_init(shape, seed=rank))
    theirs = nn.parallel.DistributedDataParallel(reference_init(shape, seed=rank))
    assert list(ours.state_dict().keys()) == list(theirs.state_dict().keys())
    for k, v in theirs.state_dict().items():                    # both constructors broadcast rank 0's state
        assert torch.equal(ours.state_dict()[k], v), k
    o1 = torch.optim.SGD(ours.parameters(), 0.05)
    o2 = torch.optim.SGD(theirs.parameters(), 0.05)
    g = torch.Generator().manual_seed(7)
    xs = torch.rand(steps, world, B, 1, *shape, generator=g)
    ys = torch.randint(0, 10, (steps, world, B), generator=g)
    crit = nn.CrossEntropyLoss()
    sizes = []
    for s in range(steps):
        for m, o in ((ours, o1), (theirs, o2)):
            loss = crit(m(xs[s, rank]), ys[s, rank])
            o.zero_grad()
            loss.backward()
            o.step()
        sizes.append(ours.bucket_sizes())
    for (k, a), (_, b) in zip(ours.state_dict().items(), theirs.state_dict().items()):
        assert torch.allclose(a.float(), b.float(), rtol=1e-5, atol=1e-6), k
    # iteration 1: one bucket with everything; afterwards rebuilt in arrival order [1 MiB, 25 MiB]
    total = sum(p.numel() * 4 for p in ours.module.parameters())
    assert sizes[0] == [total] and sum(sizes[-1]) == total
    with ours.no_sync():
        crit(ours(xs[0, rank]), ys[0, rank]).backward()
    tdist.destroy_process_group()
    dist.destroy_process_group()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("1")


def test_ddp_host_backend_matches_torch_ddp_gloo(tmp_path):
    b2.spawn(_ddp_vs_torch, args=(2, b2.find_free_port(), b2.find_free_port(), str(tmp_path)), nprocs=2)
    assert len(list(tmp_path.iterdir())) == 2


def test_mnist_distributed_cpu_eager_two_ranks(tmp_path):
    ck = str(tmp_path / "ck.pt")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "examples", "mnist_distributed.py"), "-g", "2", "--backend",
                          "gloo", "--model", "eager", "--image-shape", "32", "32", "--synthetic", "64", "--epochs", "1",
                          "--log-every", "2", "--save", ck], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-3000:]
    assert re.search(r"Rank \[0\], Epoch \[1/1\], Step \[2/7\], Loss: \d+\.\d{4}", out.stdout), out.stdout
    assert "Training complete in: " in out.stdout
    sd = torch.load(ck)
    assert list(sd.keys())[0] == "module.layer1.0.weight" and sd["module.fc.weight"].shape == (10, 32 * 8 * 8)
