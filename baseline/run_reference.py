#!/usr/bin/env python3
"""Run ONE rank of the unmodified reference (``baseline/_ref/pytorch_code/distributed_nn.py``) under shims.

    python baseline/run_reference.py --mpi-rank R --mpi-size S --rdv DIR [--device cpu|cuda:N] [--autocast none|bf16]
                                     [--bench W,K --result FILE] -- --network=ResNet18 --dataset=Cifar10 ...

What ``mpirun -n S python distributed_nn.py ...`` (reference pytorch_code/run_pytorch.sh:1) would start S times,
this starts once; ``bench.py --impl reference`` and ``tests/test_cpu_reference_arm.py`` start the S copies. The
reference script is executed with ``runpy`` from a byte-identical copy of the tree; the pieces of its 2017
environment that this image lacks are answered by ``baseline/ref_shims`` (mpi4py over local sockets, import
stubs for the TensorFlow / skimage names its data readers import, Python-2 builtins, a handful of torch
compatibility patches listed in ``ref_shims/pdnn_ref_compat.py``). Data: files of the exact on-disk formats its
readers parse (CIFAR-10 python pickles, MNIST idx.gz), filled with random pixels and labels.

Timing (``--bench W,K`` on rank 0): the master announces every step with ``comm.isend(step, dest, tag=10)``
(sync_replicas_master_nn.py:211-217). The launcher's send hook time-stamps those announcements; one step is the
interval between two of them (weights out, all workers' forward/backward, every gradient in, update applied).
After W untimed and K timed steps the master's launcher writes the result file and ends the job.
"""
import argparse
import gzip
import json
import os
import pickle
import runpy
import struct
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")
SHIMS = os.path.join(HERE, "ref_shims")


class BenchDone(BaseException):
    pass


def stage_synthetic_data(data_dir, seed=1234):
    """Random data in the formats the reference readers parse (cifar10/cifar10.py:124-199, mnist/mnist.py:133-156)."""
    import numpy as np
    marker = os.path.join(data_dir, ".complete")
    if os.path.exists(marker):
        return
    rng = np.random.RandomState(seed)
    cdir = os.path.join(data_dir, "cifar10_data", "cifar-10-batches-py")
    os.makedirs(cdir, exist_ok=True)
    for name in ["data_batch_%d" % i for i in range(1, 6)] + ["test_batch"]:
        d = {"data": rng.randint(0, 256, size=(10000, 3072)).astype(np.uint8),
             "labels": [int(v) for v in rng.randint(0, 10, size=10000)]}
        with open(os.path.join(cdir, name), "wb") as f:
            pickle.dump(d, f, protocol=2)
    # the reader only checks that the archive exists before skipping the download (cifar10.py:133-144)
    open(os.path.join(data_dir, "cifar10_data", "cifar-10-python.tar.gz"), "wb").close()
    mdir = os.path.join(data_dir, "mnist_data")
    os.makedirs(mdir, exist_ok=True)
    for prefix, n in (("train", 60000), ("t10k", 10000)):
        with gzip.open(os.path.join(mdir, "%s-images-idx3-ubyte.gz" % prefix), "wb", compresslevel=1) as f:
            f.write(struct.pack(">IIII", 2051, n, 28, 28))
            f.write(rng.randint(0, 256, size=n * 784).astype(np.uint8).tobytes())
        with gzip.open(os.path.join(mdir, "%s-labels-idx1-ubyte.gz" % prefix), "wb", compresslevel=1) as f:
            f.write(struct.pack(">II", 2049, n))
            f.write(rng.randint(0, 10, size=n).astype(np.uint8).tobytes())
    open(marker, "w").close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mpi-rank", type=int, required=True)
    ap.add_argument("--mpi-size", type=int, required=True)
    ap.add_argument("--rdv", required=True, help="rendezvous directory shared by all ranks of the job")
    ap.add_argument("--data-dir", default="", help="directory with cifar10_data/ and mnist_data/ (staged when missing)")
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--autocast", default="none", choices=["none", "bf16"])
    ap.add_argument("--script", default="distributed_nn.py")
    ap.add_argument("--bench", default="", help="W,K: warm-up and timed steps (rank 0 stops the job afterwards)")
    ap.add_argument("--result", default="")
    ap.add_argument("--log", default="", help="file for the reference's own stdout (it prints several lines per message)")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("ref_args", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    ref_args = [x for x in a.ref_args if x != "--"]

    os.makedirs(a.rdv, exist_ok=True)
    os.environ["PDNN_MPI_RANK"], os.environ["PDNN_MPI_SIZE"], os.environ["PDNN_MPI_DIR"] = str(a.mpi_rank), str(a.mpi_size), a.rdv
    data_dir = a.data_dir or os.path.join(a.rdv, "data")
    if a.mpi_rank == 0 or not a.data_dir:
        os.makedirs(data_dir, exist_ok=True)
    if a.mpi_rank == 0:
        stage_synthetic_data(data_dir)
    else:
        t_end = time.time() + 600
        while not os.path.exists(os.path.join(data_dir, ".complete")):
            if time.time() > t_end:
                raise RuntimeError("synthetic data was never staged in " + data_dir)
            time.sleep(0.1)

    code_dir = os.path.join(REF_ROOT, "pytorch_code")
    if not os.path.isdir(code_dir):
        raise RuntimeError("baseline/_ref is missing: copy the reference tree there (see DESIGN.md section 3)")
    # sys.path[0] = the script's directory as with ``python distributed_nn.py``; shims answer the missing imports
    sys.path[:0] = [code_dir, SHIMS]
    os.chdir(data_dir)                       # the readers use relative paths ('./mnist_data', 'cifar10_data/...')

    import torch
    if a.threads > 0:
        torch.set_num_threads(a.threads)
    import pdnn_ref_compat
    ctx = pdnn_ref_compat.apply(a.device, a.autocast)
    from mpi4py import MPI
    comm = MPI.COMM_WORLD

    stamps = []
    if a.bench and a.mpi_rank == 0:
        W, K = [int(v) for v in a.bench.split(",")]

        def hook(kind, payload, dest, tag):
            if kind == "obj" and tag == 10 and dest == 1:
                stamps.append(time.perf_counter())
                # stamps[0] is the announcement before the loop (sync_replicas_master_nn.py:116), stamps[j] opens step j
                if len(stamps) == W + K + 2:
                    raise BenchDone()

        comm.add_send_hook(hook)

    out = open(a.log, "w") if a.log else None
    old_stdout = sys.stdout
    if out is not None:
        sys.stdout = out
    sys.argv = [os.path.join(code_dir, a.script)] + ref_args
    status = "ended"
    t0 = time.perf_counter()
    try:
        if ctx is not None:
            with ctx:
                runpy.run_path(sys.argv[0], run_name="__main__")
        else:
            runpy.run_path(sys.argv[0], run_name="__main__")
    except BenchDone:
        status = "bench_done"
    except MPI.PeerFinished:
        status = "peer_finished"
    except BaseException:
        # a rank that dies must not leave the others blocked in a receive for ever
        import traceback
        traceback.print_exc(file=sys.stderr)
        sys.stderr.flush()
        status = "failed"
        try:
            comm.finish()
        except Exception:
            pass
    finally:
        sys.stdout = old_stdout
        if out is not None:
            out.close()
    if a.bench and a.mpi_rank == 0 and a.result:
        W, K = [int(v) for v in a.bench.split(",")]
        res = {"status": status, "stamps": len(stamps), "traffic": comm.traffic(), "wall_s": time.perf_counter() - t0}
        if status == "bench_done":
            steps = [stamps[j + 1] - stamps[j] for j in range(1, len(stamps) - 1)]
            timed = steps[W:]
            res.update({"warmup": W, "steps": K, "total_s": stamps[W + K + 1] - stamps[W + 1],
                        "step_s": timed, "warmup_step_s": steps[:W]})
        tmp = a.result + ".tmp"
        with open(tmp, "w") as f:
            json.dump(res, f)
        os.replace(tmp, a.result)
    if a.mpi_rank == 0 and status != "failed":
        comm.finish()
    sys.stdout.flush()
    # daemon threads of the shim may still sit in socket calls; leave without waiting for them
    os._exit(1 if status == "failed" else 0)


if __name__ == "__main__":
    main()
