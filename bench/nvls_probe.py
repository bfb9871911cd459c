#!/usr/bin/env python3
"""Probe torch symmetric memory / NVLS multicast on this box (run under torchrun with >= 2 ranks)."""
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    import torch.distributed._symmetric_memory as symm
    try:
        if hasattr(symm, "set_backend"):
            pass
        t = symm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
        hdl = symm.rendezvous(t, dist.group.WORLD.group_name if hasattr(dist.group.WORLD, "group_name") else dist.group.WORLD)
        info = {"rank": rank, "has_multicast": getattr(hdl, "has_multicast_support", None), "multicast_ptr": getattr(hdl, "multicast_ptr", None),
                "buffer_ptrs": list(getattr(hdl, "buffer_ptrs", []))[:world], "signal_pad_ptrs": list(getattr(hdl, "signal_pad_ptrs", []))[:2]}
        print("NVLS_PROBE", info, flush=True)
        # data check: rank 0 writes through unicast peer pointers using our p2p kernel
        sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        from pytorch_distributed_nn_b200.ops import _lib
        t.fill_(float(rank))
        torch.cuda.synchronize(); dist.barrier()
        if rank == 0:
            src = torch.full((1 << 20,), 7.0, device="cuda")
            for r in range(1, world):
                _lib.call("pdnn_p2p_copy", src, hdl.buffer_ptrs[r], (1 << 20) * 4, 0, _lib.stream_ptr())
            torch.cuda.synchronize()
        dist.barrier(); torch.cuda.synchronize()
        print("NVLS_PROBE rank", rank, "value after unicast write", float(t[12345]), flush=True)
    except Exception as e:  # noqa: BLE001
        import traceback; traceback.print_exc()
        print("NVLS_PROBE_FAILED", rank, repr(e), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
