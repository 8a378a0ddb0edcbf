"""CPU, world_size 2 over gloo: the N>1 path of the hot path is data parallel -- independent prompts sharded across
replicas with no data-path collective; the only communication is bench.py's barrier + MAX-reduce of the timed
region.  These tests run the host-side sharding and aggregation logic in two real processes."""
import os
import socket
import sys
from pathlib import Path

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int, out_dir: str):
    sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from b200sd.serving import shard_range
    from oracle import sd21 as O

    n_prompts = 5
    mine = list(shard_range(n_prompts, rank, world))
    # every rank builds the same seeded inputs and takes its slice: no scatter, no collective on the data path
    lat, ctx = O.seeded_inputs(n_prompts, 8, 16, ctx_len=4)
    my_lat = lat[mine]
    # a stand-in "denoise": any deterministic per-prompt function; results must equal the unsharded run
    my_out = torch.stack([x.flip(-1) * (i + 1) for i, x in zip(mine, my_lat)]) if mine else lat[:0]
    torch.save({"idx": mine, "out": my_out}, os.path.join(out_dir, f"r{rank}.pt"))
    # bench.py's aggregation: barrier, then MAX over ranks of the per-rank elapsed time; value = all images / max time
    dist.barrier()
    t = torch.tensor([0.010 * (rank + 1), float(len(mine))], dtype=torch.float64)
    tmax = t.clone()
    dist.all_reduce(tmax[:1], op=dist.ReduceOp.MAX)
    tot = t[1:].clone()
    dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    if rank == 0:
        torch.save({"max_s": float(tmax[0]), "images": float(tot[0])}, os.path.join(out_dir, "agg.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_two_replicas_shard_prompts_and_aggregate(tmp_path, lib_built):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    parts = [torch.load(tmp_path / f"r{r}.pt") for r in range(world)]
    idx = sum((p["idx"] for p in parts), [])
    assert idx == list(range(5)) and [len(p["idx"]) for p in parts] == [3, 2]
    from oracle import sd21 as O
    lat, _ = O.seeded_inputs(5, 8, 16, ctx_len=4)
    full = torch.stack([x.flip(-1) * (i + 1) for i, x in enumerate(lat)])
    assert torch.equal(torch.cat([p["out"] for p in parts]), full)  # sharded == unsharded, bit for bit
    agg = torch.load(tmp_path / "agg.pt")
    assert agg["max_s"] == pytest.approx(0.020) and agg["images"] == 5.0
    assert agg["images"] / agg["max_s"] == pytest.approx(250.0)  # whole-job throughput = all images / slowest rank


def test_bench_reference_arm_only_rank0_prints(tmp_path):
    """under torchrun every rank starts bench.py --impl reference; only rank 0 may work and print"""
    import subprocess
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0"], env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0 and r.stdout.strip() == ""
