#!/usr/bin/env python3
"""Flagship benchmark: synchronous-SGD parameter-server training of ResNet-18 on synthetic CIFAR-10-shape data.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--network ResNet18]
                    [--batch-size 1024] [--num-aggregate k] [--inject-delay SPEC]

Metric (BASELINE.json): images/sec (whole job) and p50/p99 step time. One rank per GPU; for N > 1 launch with
``python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N``. Rank 0 hosts the master AND a
worker (``colocate``), so N GPUs = N workers (weak scaling: per-worker batch fixed, reference run line
``--batch-size=1024``, pytorch_code/run_pytorch.sh:1).

Two timed regions, both device-timed with CUDA events and reduced with MAX over ranks:
  * ``value``     : K steps with inputs already resident on the device (kernel-side throughput);
  * ``e2e.value`` : K steps through the public API including, every step, the host->device copy of that step's
                    images+labels from pinned memory and the device->host read of the step's loss.
Every step runs the complete protocol: weight wait, forward, loss, backward with bucketed gradient push, master
k-of-n aggregation + momentum SGD + weight broadcast. Nothing is skipped or cached.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--network", default="ResNet18")
    ap.add_argument("--batch-size", type=int, default=1024, help="per-worker batch (reference run_pytorch.sh: 1024)")
    ap.add_argument("--num-aggregate", type=int, default=0, help="k of n (0 = all workers)")
    ap.add_argument("--lr", type=float, default=0.01)
    ap.add_argument("--momentum", type=float, default=0.5)
    ap.add_argument("--grad-dtype", default="bf16", choices=["bf16", "fp32"])
    ap.add_argument("--inject-delay", default="", help="straggler spec, e.g. 'const:ranks=1,2,3:ms=5' (see parallel/delays.py)")
    ap.add_argument("--n-buckets", type=int, default=8)
    ap.add_argument("--no-cuda-graph", dest="cuda_graph", action="store_false", default=True,
                    help="enqueue every kernel of every step from the host instead of replaying the step from a CUDA graph")
    ap.add_argument("--master-only", action="store_true", help="rank 0 is master-only as in the reference (n = N-1 workers)")
    ap.add_argument("--interval-ms", type=float, default=0.0, help="interval policy: the master closes each step after this deadline")
    ap.add_argument("--worker-times-cdf", action="store_true", help="report the per-(worker, step) time distribution (TF study)")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of this rank's GPU while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append(line.strip())

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            f = [t.strip() for t in s.split(",")]
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
                for n, v in zip(names, f[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        # under load = upper half of the samples (idle samples before/after the region read low)
        med = sm[len(sm) * 3 // 4] if sm else 0.0
        return {"sm_mhz": med, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


def _ref_unavailable(why):
    if int(os.environ.get("RANK", "0")) == 0:      # under torchrun every rank runs this: one line, from rank 0
        print(json.dumps({"impl": "reference", "unavailable": why}))
    return 0


def run_reference(args):
    """``--impl reference``: the UNMODIFIED reference tree (byte-identical copy under baseline/_ref) runs its own
    ``pytorch_code/distributed_nn.py`` (master = its SyncReplicasMaster_NN, workers = its DistributedWorker on its
    ResNetSplit18; reference run line pytorch_code/run_pytorch.sh:1) -- none of this repo's models, kernels or engine.
    Each torchrun rank starts one reference worker on its GPU (MPI rank r+1); rank 0 also starts the reference's
    master process (MPI rank 0, host memory, as written). ``baseline/run_reference.py`` documents the environment
    shims (mpi4py over local sockets, import stubs, torch compatibility patches) and how a step is timed."""
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    ref_dir = os.path.join(ROOT, "baseline", "_ref", "pytorch_code")
    if not os.path.isdir(ref_dir):
        src = "/root/reference"
        if os.path.isdir(os.path.join(src, "pytorch_code")) and rank == 0:
            import shutil
            shutil.copytree(src, os.path.join(ROOT, "baseline", "_ref"), symlinks=True, dirs_exist_ok=True,
                            ignore=shutil.ignore_patterns(".git"))
        if not os.path.isdir(ref_dir) and world == 1:
            return _ref_unavailable("baseline/_ref (byte-identical copy of the reference tree) is missing and /root/reference is not mounted")
    if args.network not in ("ResNet18", "ResNet34", "LeNet"):
        return _ref_unavailable(f"the reference's distributed entry point builds only LeNet / ResNet18 / ResNet34 "
                                f"(sync_replicas_master_nn.py:103-109), not {args.network}")
    import torch
    have_gpu = torch.cuda.is_available()
    mode = os.environ.get("PDNN_REF_DEVICE", "cuda" if have_gpu else "cpu")
    autocast = os.environ.get("PDNN_REF_AUTOCAST", "bf16" if mode == "cuda" else "none")
    device = f"cuda:{local_rank}" if mode == "cuda" else "cpu"
    tag = f"{os.getppid()}_{os.environ.get('MASTER_PORT', '0')}" if world > 1 else f"{os.getpid()}_solo"
    rdv = os.path.join(os.environ.get("PDNN_REF_TMP", "/tmp"), f"pdnn_ref_{tag}")
    data_dir = os.path.join(os.environ.get("PDNN_REF_TMP", "/tmp"), "pdnn_ref_data")
    os.makedirs(rdv, exist_ok=True)
    result = os.path.join(rdv, "result.json")
    launcher = os.path.join(ROOT, "baseline", "run_reference.py")
    dataset = "MNIST" if args.network == "LeNet" else "Cifar10"
    k = args.num_aggregate or world
    ref_argv = [f"--network={args.network}", f"--dataset={dataset}", f"--batch-size={args.batch_size}", "--comm-type=Bcast",
                f"--num-aggregate={k}", f"--lr={args.lr}", f"--momentum={args.momentum}"]
    ncpu = os.cpu_count() or 8
    threads = max(1, ncpu // (world + 1))
    common = [sys.executable, launcher, "--mpi-size", str(world + 1), "--rdv", rdv, "--data-dir", data_dir,
              "--bench", f"{args.warmup},{args.steps}", "--result", result, "--threads", str(threads)]
    procs = []
    if rank == 0:
        procs.append(subprocess.Popen(common + ["--mpi-rank", "0", "--device", "cpu", "--log", os.path.join(rdv, "master.log"), "--"] + ref_argv))
    procs.append(subprocess.Popen(common + ["--mpi-rank", str(rank + 1), "--device", device, "--autocast", autocast,
                                            "--log", os.path.join(rdv, f"worker{rank + 1}.log"), "--"] + ref_argv))
    sampler = ClockSampler(local_rank) if (rank == 0 and have_gpu) else None
    if sampler:
        sampler.start()
    rcs = [p.wait() for p in procs]
    clocks = sampler.stop() if sampler else None
    if rank != 0:
        return 0
    if not os.path.exists(result):
        return _ref_unavailable(f"reference run ended without a result (exit codes {rcs}); logs in {rdv}")
    with open(result) as f:
        res = json.load(f)
    if res.get("status") != "bench_done":
        return _ref_unavailable(f"reference run stopped early: {res.get('status')} after {res.get('stamps')} step announcements; logs in {rdv}")
    B, K = args.batch_size, args.steps
    gb = B * world                      # the reference divides by n and waits for every worker when k = n
    total_s = res["total_s"]
    st = sorted(res["step_s"])
    n_param = {"ResNet18": 11173962, "ResNet34": 21282122, "LeNet": 431080}[args.network]
    chw = (1, 28, 28) if args.network == "LeNet" else (3, 32, 32)
    batch_bytes = B * chw[0] * chw[1] * chw[2] * 4 + B * 8
    value = gb * K / total_s
    out = {
        "metric": "images/sec (sync-SGD parameter server, whole job) + p50/p99 step ms",
        "value": round(value, 1), "unit": "images/sec", "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": round(total_s / K * 1e3, 4), "p50_step_ms": round(st[len(st) // 2] * 1e3, 4),
        "p99_step_ms": round(st[min(len(st) - 1, int(len(st) * 0.99))] * 1e3, 4),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16 autocast (fp32 parameters, fp64 wire as written)" if autocast == "bf16" else "fp32 (fp64 wire as written)",
        "data": "synthetic", "impl": "reference",
        "config": {"model": args.network, "dataset_shape": "CIFAR-10 (3x32x32, 10 classes)" if chw[0] == 3 else "MNIST (1x28x28)",
                   "global_batch": gb, "per_worker_batch": B, "seq_len": None,
                   "parallelism": f"reference PS: 1 master process (host) + {world} worker processes, k={k} of n={world}",
                   "entry_point": "baseline/_ref/pytorch_code/distributed_nn.py (unmodified) " + " ".join(ref_argv),
                   "device": device if mode == "cpu" else "one B200 per worker (modules placed by torch.set_default_device; see baseline/ref_shims/pdnn_ref_compat.py)",
                   "transport": "mpi4py-compatible shim over AF_UNIX sockets (baseline/ref_shims/mpi4py)",
                   "timing": "host clock between the master's step announcements (tag 10); every step ends with a blocking D2H of all gradients, so the GPU is idle at both edges",
                   "l2_policy": "inputs come from a 50000-image host array, a new batch every step"},
        # the reference has one path only: every step copies its batch H2D and reads all gradients + loss back
        "e2e": {"value": round(value, 1), "unit": "images/sec", "ms_per_step": round(total_s / K * 1e3, 4),
                "h2d_bytes_per_step": (batch_bytes + 4 * n_param) if mode == "cuda" else 0,
                "d2h_bytes_per_step": (4 * n_param + 4) if mode == "cuda" else 0},
        "gpu_launches": None,
        "mpi_bytes": res.get("traffic"),
        "clocks": clocks,
    }
    print(json.dumps(out))
    import shutil
    shutil.rmtree(rdv, ignore_errors=True)
    return 0


def main():
    args = parse()
    if args.impl == "reference":
        return run_reference(args)
    import torch
    import torch.distributed as dist
    from pytorch_distributed_nn_b200.ops import _lib
    from pytorch_distributed_nn_b200.parallel.b200_ps import B200ParameterServer
    from pytorch_distributed_nn_b200.parallel.delays import DeviceDelay

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    B = args.batch_size
    # straggler injection: sampled on the device (Philox keyed by seed/worker/step/layer), so it replays inside the CUDA graph
    device_delay = DeviceDelay(args.inject_delay, seed=1) if args.inject_delay else None
    ps = B200ParameterServer(args.network, B, rank, world, lr=args.lr, momentum=args.momentum,
                             num_aggregate=args.num_aggregate or None, grad_dtype=args.grad_dtype, colocate=not (args.master_only and world > 1),
                             n_buckets=args.n_buckets, device_delay=device_delay, group=group, device=dev, deadline_ms=args.interval_ms)
    chw = ps.net.in_chw
    gen = torch.Generator().manual_seed(1234 + rank)
    # synthetic CIFAR-10-shape data: raw 0..255 pixels like the reference reader (cifar10.py:167-199), 4 batches in rotation
    host_x = [torch.empty(B, *chw).uniform_(0, 255, generator=gen).pin_memory() for _ in range(4)]
    host_y = [torch.randint(0, 10, (B,), generator=gen).pin_memory() for _ in range(4)]
    dev_x = [t.to(dev) for t in host_x]
    dev_y = [t.to(dev) for t in host_y]

    class HostBatches:
        """What a data set hands to the training loop: host tensors, one batch per call (utils/data.py next_batch contract)."""
        i = 0

        def next_batch(self, batch_size):
            j = self.i % 4
            self.i += 1
            return host_x[j], host_y[j]

    loader = HostBatches()

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_resident(n):
        for i in range(n):
            ps.train_step(dev_x[i % 4], dev_y[i % 4])

    def run_e2e(n):
        """The public training loop (B200ParameterServer.fit, the one SyncReplicasMaster_NN / DistributedWorker run):
        every step's batch goes host -> pinned -> device on a side stream, the loss of every step is read back."""
        return ps.fit(loader if ps.is_worker else None, ps.step + n - 1, log_interval=1 << 30, cuda_graph=args.cuda_graph,
                      graph_after=0, read_loss_every_step=True)

    # ---- warm-up (untimed) -------------------------------------------------------------------
    run_resident(2)
    if args.cuda_graph:
        # the whole step (wait, weight fetch, forward, loss, backward + bucket pushes, master kernels) becomes one graph launch
        ps.enable_cuda_graph(dev_x[0], dev_y[0])
    run_resident(args.warmup)
    run_e2e(max(1, min(args.warmup, 3)))
    sync_all()

    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.3)

    # ---- timed region 1: device-resident inputs ----------------------------------------------
    sync_all()
    sets_before_resident = 0
    if rank == 0:
        ps.last_result()
        sets_before_resident = ps.total_sets_applied
    _lib.reset_launch_count()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    per_step = []
    s.record()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record()
    for i in range(args.steps):
        ps.train_step(dev_x[i % 4], dev_y[i % 4])
        evs[i + 1].record()
    ps.join()                  # the master's stream: its last release belongs to the timed region
    e.record()
    sync_all()
    launches = _lib.launch_count()
    total_ms = s.elapsed_time(e)
    sets_before_e2e = None
    if rank == 0:
        ps.last_result()
        sets_after_resident = ps.total_sets_applied
    per_step_raw = [evs[i].elapsed_time(evs[i + 1]) for i in range(args.steps)]
    per_step = sorted(per_step_raw)
    cdf = None
    if args.worker_times_cdf:
        # per-(worker, iteration) times as in the reference's compute-time study (distributed_TF/tools/benchmark.py:60-111)
        mine = torch.tensor(per_step_raw, device=dev, dtype=torch.float64)
        allt = [torch.empty_like(mine) for _ in range(world)] if world > 1 else [mine]
        if world > 1:
            dist.all_gather(allt, mine)
        if rank == 0:
            from pytorch_distributed_nn_b200.parallel.delays import summarize_ms
            mat = torch.stack(allt).cpu()                        # [worker][iteration]
            per_iter = mat.max(dim=0).values.tolist()
            cdf = {"all": summarize_ms(mat.flatten().tolist()),
                   "per_iteration_slowest_worker": summarize_ms(per_iter),
                   "per_worker_mean_ms": [float(v) for v in mat.mean(dim=1)]}

    # ---- timed region 2: end to end (H2D inputs from pinned memory + D2H loss every step) ----
    sync_all()
    s2, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s2.record()
    losses = [v for _, v in run_e2e(args.steps)]
    ps.join()
    e2.record()
    sync_all()
    e2e_ms = s2.elapsed_time(e2)
    clocks = sampler.stop()

    t = torch.tensor([total_ms, e2e_ms, per_step[len(per_step) // 2], per_step[min(len(per_step) - 1, int(len(per_step) * 0.99))]],
                     device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, e2e_ms, p50, p99 = [float(v) for v in t.tolist()]
    res = ps.last_result() if rank == 0 else None
    if rank == 0:
        gb = B * ps.k            # images whose gradients are applied per step (k of n sets; k = n without stragglers)
        ps.last_result()
        # gradient sets the master actually applied (under the interval policy fewer than k may make a deadline)
        sets_e2e = ps.total_sets_applied - sets_after_resident
        sets_resident = sets_after_resident - sets_before_resident
        value = B * sets_resident / (total_ms / 1e3)
        e2e_value = B * sets_e2e / (e2e_ms / 1e3)
        h2d = B * chw[0] * chw[1] * chw[2] * 4 + B * 8
        out = {
            "metric": "images/sec (sync-SGD parameter server, whole job) + p50/p99 step ms",
            "value": round(value, 1), "unit": "images/sec", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(total_ms / args.steps, 4), "p50_step_ms": round(p50, 4), "p99_step_ms": round(p99, 4),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "impl": "ours",
            "config": {"model": args.network, "dataset_shape": "CIFAR-10 (3x32x32, 10 classes)" if chw[0] == 3 else "MNIST (1x28x28)",
                       "global_batch": gb, "per_worker_batch": B, "seq_len": None,
                       "parallelism": (f"ps-dp{ps.n_workers} (master+worker colocated on GPU0, k={ps.k} of n={ps.n_workers})" if ps.colocate else
                                       f"ps-dp{ps.n_workers} (GPU0 master-only as in the reference, k={ps.k} of n={ps.n_workers})"),
                       "optimizer": f"momentum-SGD lr={args.lr} mu={args.momentum} (fp32 master weights)",
                       "grad_wire_dtype": args.grad_dtype, "cuda_graph": bool(args.cuda_graph), "inject_delay": args.inject_delay or "none",
                       "l2_policy": "working set (activations >> 126 MB L2 at batch 1024; 4 input batches rotate) exceeds L2"},
            "e2e": {"value": round(e2e_value, 1), "unit": "images/sec", "ms_per_step": round(e2e_ms / args.steps, 4),
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4, "final_loss": losses[-1] if losses else None,
                    "path": "B200ParameterServer.fit (the loop of the master / worker roles): every step's batch pinned host -> device on a side stream, every step's loss device -> pinned host and read by the host one step later"},
            "gpu_launches": int(launches),
            "gpu_launches_per_step": round(launches / args.steps, 1),
            "gradient_sets_applied": {"resident": sets_resident, "e2e": sets_e2e},
            "last_master_result": {"selected_mask": res[0], "kill_mask": res[1], "n_received": res[2], "step": res[3],
                                   "bucket_masks": ps.last_bucket_masks, "device_timeline_us": ps.last_timing_us},
            "clocks": clocks,
        }
        if cdf is not None:
            out["worker_times_cdf"] = cdf
        print(json.dumps(out))
    ps.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
