#!/usr/bin/env python3
"""Link microbenchmarks of the box: the measured denominators of every "fraction of the NVLink roofline".

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 bench/link_bench.py [--mb 256] [--out FILE]

Measured with this library's own kernels on the peer arena the parameter server uses, CUDA events, best of 5:
  p2p_store   rank 0 stores --mb MiB into rank 1's memory (16-byte st.global, `pdnn_p2p_copy`)           GB/s
  p2p_load    rank 0 loads --mb MiB from rank 1's memory                                                   GB/s
  p2p_all_in  every rank r > 0 stores into rank 0 at once (the unicast gradient push pattern)              GB/s at rank 0
  mc_store    rank 0 stores --mb MiB once through the multicast address (`multimem.st`): every GPU gets it GB/s out of rank 0
  mc_reduce   rank 0 reads --mb MiB of fp32 through the multicast address (`multimem.ld_reduce.add`)       GB/s into rank 0
  flag_rtt    release-store / acquire-load ping-pong between rank 0 and rank 1                             ns per round trip
  local_copy  device-local copy with the same kernel (HBM read + write)                                    GB/s
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from pytorch_distributed_nn_b200.ops import _lib  # noqa: E402
from pytorch_distributed_nn_b200.parallel.peer import make_arena  # noqa: E402


def best_ms(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(reps):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        best = min(best, s.elapsed_time(e))
    return best


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=256)
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    assert world >= 2, "needs at least two GPUs"
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=dev)
    nbytes = a.mb << 20
    arena = make_arena(2 * nbytes + 8192, rank, world, dev)
    off_flag = arena.carve(4096)
    off_a = arena.carve(nbytes)
    off_b = arena.carve(nbytes)
    local_src = torch.randn(nbytes // 4, device=dev)
    local_dst = torch.empty(nbytes // 4, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    res = {"n_gpus": world, "mb": a.mb, "multicast": bool(getattr(arena, "mc_ptr", 0))}

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()

    barrier()
    if rank == 0:
        res["local_copy_GBps"] = 2 * nbytes / best_ms(lambda: _lib.call("pdnn_p2p_copy", local_src, local_dst, nbytes, 0, st)) / 1e6
        res["p2p_store_GBps"] = nbytes / best_ms(lambda: _lib.call("pdnn_p2p_copy", local_src, arena.ptr(1, off_a), nbytes, 0, st)) / 1e6
        res["p2p_load_GBps"] = nbytes / best_ms(lambda: _lib.call("pdnn_p2p_copy", arena.ptr(1, off_a), local_dst, nbytes, 0, st)) / 1e6
    barrier()
    # every other rank pushes into rank 0 at the same time: per-rank slice so that the ingress total is nbytes
    share = nbytes // max(1, world - 1) // 16 * 16
    t = torch.zeros(1, device=dev, dtype=torch.float64)
    if rank > 0:
        fn = lambda: _lib.call("pdnn_p2p_copy", local_src, arena.ptr(0, off_a) + (rank - 1) * share, share, 0, st)  # noqa: E731
        fn()
        barrier()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(5):
            fn()
        e.record()
        torch.cuda.synchronize()
        t[0] = s.elapsed_time(e) / 5
    else:
        barrier()
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    res["p2p_all_in_GBps_at_rank0"] = share * (world - 1) / float(t) / 1e6
    barrier()
    if res["multicast"]:
        if rank == 0:
            res["mc_store_GBps_out_of_rank0"] = nbytes / best_ms(lambda: _lib.call("pdnn_mc_store", local_src, arena.mc(off_b), nbytes, 0, st)) / 1e6
            res["mc_store_GBps_delivered"] = res["mc_store_GBps_out_of_rank0"] * world
        barrier()
        if rank == 0:
            ms = best_ms(lambda: _lib.call("pdnn_mc_ld_reduce_f32", arena.mc(off_a), local_dst, nbytes // 4, 0, st))
            res["mc_reduce_GBps_into_rank0"] = nbytes / ms / 1e6
            res["mc_reduce_GBps_summed"] = nbytes * world / ms / 1e6
        barrier()
    # flag round trip rank 0 <-> rank 1
    out_ns = torch.zeros(1, device=dev, dtype=torch.int64)
    iters = 2000
    if rank in (0, 1):
        mine = arena.ptr(rank, off_flag)
        peer = arena.ptr(1 - rank, off_flag)
    for rep in range(3):
        barrier()
        if rank in (0, 1):
            _lib.call("pdnn_flag_pingpong", mine, peer, iters, int(rank == 0), rep * iters, out_ns, st)
        barrier()
    if rank == 0:
        res["flag_round_trip_ns"] = float(out_ns) / iters
        print("LINK " + json.dumps(res), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(res, f, indent=1)
    barrier()
    arena.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
