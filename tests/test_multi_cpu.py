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


# ------------------------------------------------------------------------------------------------ patch-parallel bands
def _band_worker(rank: int, world: int, port: int, out_dir: str):
    """BandComm (b200sd/patch.py) on gloo: the partition arithmetic and the three exchanges of the patch-parallel mode,
    with torch CPU convolutions standing in for the kernels -- band results must equal the bands of the unsplit result."""
    sys.path.insert(0, str(ROOT / "scalable-hw-agnostic-inference_b200"))
    sys.path.insert(0, str(ROOT))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch.nn.functional as F
    from b200sd.patch import BandComm

    comm = BandComm()
    g = torch.Generator("cpu").manual_seed(0)
    n, H, W, C = 2, 8 * world, 6, 5
    full = torch.randn(n, H, W, C, generator=g)
    wt = torch.randn(7, C, 3, 3, generator=g)
    r0, rows = comm.band(H)
    assert (r0, rows) == (rank * 8, 8)
    band = full[:, r0:r0 + rows].contiguous()
    res = {}

    # halo rows: neighbours' rows, zeros at the image edges
    ext = comm.halo_pad(band)
    want = F.pad(full, (0, 0, 0, 0, 1, 1))[:, r0:r0 + rows + 2]
    res["halo"] = torch.equal(ext, want)

    # stride-1 conv: extended band, the kernel's own zero padding, outer rows dropped == band of the full conv
    conv = lambda x, **kw: F.conv2d(x.permute(0, 3, 1, 2), wt, **kw).permute(0, 2, 3, 1)
    ref = conv(full, padding=1)[:, r0:r0 + rows]
    got = conv(ext, padding=1)[:, 1:rows + 1]
    res["conv"] = torch.allclose(got, ref, atol=1e-5)

    # stride-2 conv: halo + one zero column in front, then pad (0,1,0,1) + no padding (im2col's pad_lo = 0), the extra
    # last row dropped == band of the full stride-2, padding-1 conv
    ext2 = comm.halo_pad(band, left_pad=1)
    ref2 = conv(full, padding=1, stride=2)[:, r0 // 2:(r0 + rows) // 2]
    x2 = F.pad(ext2.permute(0, 3, 1, 2), (0, 1, 0, 1))
    got2 = F.conv2d(x2, wt, stride=2).permute(0, 2, 3, 1)[:, :rows // 2]
    res["conv_s2"] = got2.shape == ref2.shape and torch.allclose(got2, ref2, atol=1e-5)

    # GroupNorm statistics: local sums all-reduced == sums over the whole image
    groups = 1
    loc = torch.stack([band.sum(dim=(1, 2, 3)), (band * band).sum(dim=(1, 2, 3))], dim=-1).view(n, groups, 2).contiguous()
    comm.allreduce_sums(loc)
    tot = torch.stack([full.sum(dim=(1, 2, 3)), (full * full).sum(dim=(1, 2, 3))], dim=-1).view(n, groups, 2)
    res["sums"] = torch.allclose(loc, tot, rtol=1e-5)

    # K|V rows: per-image gathers land in row order, batch-major
    kv = comm.all_gather_rows(band.reshape(n, rows * W, C))
    res["gather"] = torch.equal(kv, full.reshape(n, H * W, C))
    res["bands"] = torch.equal(comm.all_gather_bands(band), full)
    res["bytes"] = (comm.bytes_gathered, comm.bytes_halo, comm.n_allreduce)
    torch.save(res, os.path.join(out_dir, f"band{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_patch_parallel_band_exchanges(tmp_path, lib_built, world):
    mp.spawn(_band_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        res = torch.load(tmp_path / f"band{r}.pt")
        for k in ("halo", "conv", "conv_s2", "sums", "gather", "bands"):
            assert res[k], (r, k)
        gathered, halo, n_ar = res["bytes"]
        n, rows, W, C = 2, 8, 6, 5
        assert gathered == 2 * (world - 1) * n * rows * W * C * 4 and n_ar == 1
        inner = 0 < r < world - 1
        assert halo == 2 * (2 if inner else 1) * n * W * C * 4   # two halo_pad calls
    with pytest.raises(ValueError):
        from b200sd.patch import BandComm  # noqa: F401  (partition check needs no process group state beyond rank / world)
        class _C(BandComm):
            def __init__(self):
                self.rank, self.world = 0, 3
        _C().band(16)
