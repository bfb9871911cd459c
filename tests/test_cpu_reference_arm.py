"""The reference arm of bench.py: the mpi4py stand-in's matching semantics, and one short run of the UNMODIFIED
reference (baseline/_ref/pytorch_code/distributed_nn.py, ResNet18, 1 master + 1 worker) on the CPU."""
import json
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHIMS = os.path.join(ROOT, "baseline", "ref_shims")

RANK_PROGRAM = textwrap.dedent("""
    import sys, numpy as np
    sys.path.insert(0, %r)
    from mpi4py import MPI
    comm = MPI.COMM_WORLD
    rank, size = comm.Get_rank(), comm.Get_size()
    if rank == 0:
        # any-source receives are matched in post order per tag; Waitany reports each exactly once
        bufs = [[np.zeros(4) for _ in range(size - 1)] for _ in range(2)]
        reqs = [comm.Irecv([bufs[t][k], MPI.DOUBLE], source=MPI.ANY_SOURCE, tag=88 + t) for t in range(2) for k in range(size - 1)]
        extra = comm.Irecv([np.zeros(4), MPI.DOUBLE], source=MPI.ANY_SOURCE, tag=99)      # never matched -> cancelled
        seen = []
        for _ in range(2 * (size - 1)):
            st = MPI.Status()
            i = MPI.Request.Waitany(reqs, st)
            seen.append((i, st.source, st.tag))
            t, k = divmod(i, size - 1)
            assert bufs[t][k][0] == st.source and bufs[t][k][1] == st.tag, (bufs[t][k], st.source, st.tag)
        assert sorted(i for i, _, _ in seen) == list(range(2 * (size - 1)))
        assert MPI.Request.Waitany(reqs, MPI.Status()) == MPI.UNDEFINED
        extra.Cancel()
        # pickled objects + Bcast + probe
        for d in range(1, size):
            comm.isend({"step": 7}, dest=d, tag=10).wait()
        w = np.arange(1000, dtype=np.float64)
        comm.Bcast([w, MPI.DOUBLE], root=0)
        st = MPI.Status()
        assert comm.Iprobe(1, 77, st) is False and st.Get_source() != 1
        comm.Barrier()
        print("RANK0 OK")
    else:
        for t in (1, 0):                       # tag order differs from the master's post order on purpose
            comm.Isend([np.array([rank, 88 + t, 0, 0], dtype=np.float64), MPI.DOUBLE], dest=0, tag=88 + t).wait()
        assert comm.irecv(source=0, tag=10).wait() == {"step": 7}
        w = np.zeros(1000)
        comm.Bcast([w, MPI.DOUBLE], root=0)
        assert w[999] == 999.0
        comm.Barrier()
        print("RANK%%d OK" %% rank)
""") % SHIMS


def test_mpi_shim_matching(tmp_path):
    size = 3
    procs = []
    for r in range(size):
        env = dict(os.environ, PDNN_MPI_RANK=str(r), PDNN_MPI_SIZE=str(size), PDNN_MPI_DIR=str(tmp_path))
        procs.append(subprocess.Popen([sys.executable, "-c", RANK_PROGRAM], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=120)[0] for p in procs]
    for r, (p, o) in enumerate(zip(procs, outs)):
        assert p.returncode == 0 and f"RANK{r} OK" in o, o


@pytest.mark.skipif(not os.path.isdir(os.path.join(ROOT, "baseline", "_ref", "pytorch_code")) and not os.path.isdir("/root/reference/pytorch_code"),
                    reason="no copy of the reference tree available")
def test_reference_arm_runs_the_unmodified_reference(tmp_path):
    env = dict(os.environ, PDNN_REF_DEVICE="cpu", PDNN_REF_TMP=str(tmp_path))
    for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE"):
        env.pop(k, None)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--batch-size", "8", "--steps", "2", "--warmup", "1"],
                       env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = [l for l in r.stdout.splitlines() if l.startswith("{")][-1]
    out = json.loads(line)
    assert out["impl"] == "reference" and "unavailable" not in out, out
    assert out["steps"] == 2 and out["warmup"] == 1 and out["value"] > 0 and out["config"]["model"] == "ResNet18"
    # weights out (fp64, one message per tensor) and gradients back, for 1 + 2 (+ the opening) steps
    assert out["mpi_bytes"]["bytes_recv"] >= 3 * 8 * 11173962
    # the copy that ran is the reference, byte for byte
    if os.path.isdir("/root/reference/pytorch_code"):
        for name in ("distributed_nn.py", "sync_replicas_master_nn.py", "distributed_worker.py", "model_ops/resnet_split.py"):
            with open(os.path.join("/root/reference/pytorch_code", name), "rb") as a, open(os.path.join(ROOT, "baseline", "_ref", "pytorch_code", name), "rb") as b:
                assert a.read() == b.read(), name
