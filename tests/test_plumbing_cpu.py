"""CPU-only plumbing tests: launcher, TCP store, env://
rendezvous + its sanity errors, host collectives (world 2 and 4), new_group caching.
Multi-process paths run on the shared-memory 'gloo' backend, as the reference's own
bring-up test does when no GPU is present (test_init.py:55,84-88)."""
import os
import time

import pytest
import torch

import torch_distributed_sandbox_b200 as b2
from torch_distributed_sandbox_b200 import distributed as dist
from torch_distributed_sandbox_b200.launch import ProcessExitedError, ProcessRaisedError
from torch_distributed_sandbox_b200.store import StoreTimeout, TCPStore


# ------------------------------------------------------------------ store ----
def test_store_set_get_add_timeout():
    m = TCPStore("127.0.0.1", 0, True, timeout=5)
    c = TCPStore("127.0.0.1", m.port, False, timeout=5)
    c.set("k", b"v1")
    assert m.get("k") == b"v1"
    assert c.add("n", 2) == 2 and m.add("n", 5) == 7
    assert c.check("k") and not c.check("missing")
    assert c.delete_key("k") and not c.check("k")
    t0 = time.time()
    with pytest.raises(StoreTimeout):
        c.get("never", 0.2)
    assert 0.15 < time.time() - t0 < 2.0
    big = os.urandom(300_000)
    m.set("big", big)
    assert c.get("big") == big
    c.close()
    m.close()


def test_find_free_port_is_str_and_bindable():
    p = b2.find_free_port()
    assert isinstance(p, str) and 1024 < int(p) < 65536
    TCPStore("127.0.0.1", int(p), True, timeout=2).close()


# --------------------------------------------------------------- launcher ----
def _ok(i, base):
    assert os.environ["B2_TEST_ENV"] == "inherited"
    open(f"{base}.{i}", "w").write(str(i))


def _boom(i):
    if i == 1:
        raise ValueError("rank one exploded")
    time.sleep(30)


def _die(i):
    if i == 0:
        os._exit(7)
    time.sleep(30)


def test_spawn_runs_every_rank_and_inherits_env(tmp_path):
    os.environ["B2_TEST_ENV"] = "inherited"
    b2.spawn(_ok, args=(str(tmp_path / "f"),), nprocs=3)
    assert sorted(p.name for p in tmp_path.iterdir()) == ["f.0", "f.1", "f.2"]


def test_spawn_propagates_traceback_and_kills_siblings():
    t0 = time.time()
    with pytest.raises(ProcessRaisedError) as ei:
        b2.spawn(_boom, nprocs=3)
    assert "rank one exploded" in str(ei.value) and ei.value.error_index == 1
    assert time.time() - t0 < 20
    with pytest.raises(ProcessExitedError) as ei:
        b2.spawn(_die, nprocs=2)
    assert ei.value.exit_code == 7


# ------------------------------------------------------------- rendezvous ----
def test_rendezvous_sanity_errors(monkeypatch):
    for k in ("MASTER_ADDR", "MASTER_PORT", "RANK", "WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    with pytest.raises(ValueError, match="MASTER_ADDR"):
        dist.init_process_group("gloo", rank=0, world_size=1)
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", b2.find_free_port())
    with pytest.raises(ValueError, match="RANK"):
        dist.init_process_group("gloo")
    with pytest.raises(ValueError, match="outside"):
        dist.init_process_group("gloo", rank=4, world_size=4)
    with pytest.raises(ValueError, match="unknown backend"):
        dist.init_process_group("mpi", rank=0, world_size=1)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="needs CUDA"):
            dist.init_process_group("nccl", rank=0, world_size=1)
    assert not dist.is_initialized()


def test_single_process_group_and_double_init(monkeypatch):
    monkeypatch.setenv("MASTER_ADDR", "127.0.0.1")
    monkeypatch.setenv("MASTER_PORT", b2.find_free_port())
    dist.init_process_group("gloo", rank=0, world_size=1)
    try:
        with pytest.raises(RuntimeError, match="twice"):
            dist.init_process_group("gloo", rank=0, world_size=1)
        t = torch.tensor(5, dtype=torch.int32)
        dist.all_reduce(t)
        dist.barrier()
        assert int(t) == 5 and dist.get_rank() == 0 and dist.get_world_size() == 1
    finally:
        dist.destroy_process_group()
    assert not dist.is_initialized()


def _timeout_rank(i, port):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", port
    dist.init_process_group("gloo", rank=0, world_size=2, timeout=1.0)   # rank 1 never shows up


def test_rendezvous_timeout_names_missing_ranks():
    with pytest.raises(ProcessRaisedError, match="only 1 of 2 ranks"):
        b2.spawn(_timeout_rank, args=(b2.find_free_port(),), nprocs=1)


def _dup_rank(i, port):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", port
    dist.init_process_group("gloo", rank=0 if i < 2 else 1, world_size=2, timeout=3.0)


def test_duplicate_rank_is_reported():
    with pytest.raises((ProcessRaisedError, ProcessExitedError)):
        b2.spawn(_dup_rank, args=(b2.find_free_port(),), nprocs=2)


# ------------------------------------------------------------ collectives ----
def _collectives(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", port
    dist.init_process_group("gloo", rank=rank, world_size=world)
    assert dist.get_rank() == rank and dist.get_world_size() == world
    # 0-dim int32 SUM: the toy's tensor (allreduce_toy.py:30-31)
    t = torch.tensor(rank + 1, dtype=torch.int32)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    assert int(t) == world * (world + 1) // 2
    # fp32 larger than one shared-memory slot (chunking), SUM / AVG / MAX
    n = (4 << 20) // 4 * 2 + 13
    base = torch.arange(n, dtype=torch.float32) % 97
    x = base * (rank + 1)
    dist.all_reduce(x)
    assert torch.equal(x, base * (world * (world + 1) // 2))
    y = torch.full((1000,), float(rank))
    dist.all_reduce(y, op=dist.ReduceOp.AVG)
    assert torch.allclose(y, torch.full((1000,), (world - 1) / 2))
    z = torch.tensor([rank, -rank], dtype=torch.int64)
    dist.all_reduce(z, op=dist.ReduceOp.MAX)
    assert z.tolist() == [world - 1, 0]
    bf = torch.full((64,), rank + 0.5, dtype=torch.bfloat16)
    dist.all_reduce(bf)
    assert torch.equal(bf.float(), torch.full((64,), sum(r + 0.5 for r in range(world))))
    # broadcast from a non-zero global rank, all_gather
    b = torch.full((5,), float(rank))
    dist.broadcast(b, src=world - 1)
    assert torch.equal(b, torch.full((5,), float(world - 1)))
    outs = [torch.zeros(3) for _ in range(world)]
    dist.all_gather(outs, torch.full((3,), float(rank)))
    assert [int(o[0]) for o in outs] == list(range(world))
    # new_group: cached by rank tuple, non-members get the sentinel, sub-group reduce
    g_all = dist.new_group(ranks=list(range(world)))
    assert g_all is dist.new_group(ranks=list(range(world)))
    sub = dist.new_group(ranks=[0, world - 1])
    s = torch.ones(2)
    if rank in (0, world - 1):
        dist.all_reduce(s, group=sub)
        assert s.tolist() == ([2.0, 2.0] if world > 1 else [1.0, 1.0])
        assert dist.get_rank(sub) == (0 if rank == 0 else 1)
    else:
        assert sub is dist.GroupMember.NON_MEMBER and dist.all_reduce(s, group=sub) is None
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, f"ok{rank}"), "w").write("1")


@pytest.mark.parametrize("world", [2, 4])
def test_host_collectives(world, tmp_path):
    b2.spawn(_collectives, args=(world, b2.find_free_port(), str(tmp_path)), nprocs=world)
    assert len(list(tmp_path.iterdir())) == world
