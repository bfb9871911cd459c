#!/usr/bin/env python3
"""Step-level roofline from an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv`, one row per launch).

Groups the launches of one training step by kernel family, adds the work each family does for the named network (conv
FLOPs from the layer shapes, BatchNorm bytes from the activation sizes) and relates the achieved rates to the measured
peaks in MEASURED_PEAKS.json. This is how the "Step-level roofline" table in profiles/README.md is produced.

    python tools/roofline_from_launches.py profiles/launches_resnet18_b1024_v10.csv --network ResNet18 --batch 1024
"""
import argparse
import collections
import csv
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

GROUPS = [
    ("wgrad", ("conv_halo_wgrad", "gemm_tcgen05_kernel<1, ")),
    ("fprop+dgrad", ("conv_halo_kernel", "gemm_tcgen05_kernel<2, ", "conv_stem")),
    ("fc", ("gemm_tcgen05_kernel<0", "gemm_simt")),
    ("batchnorm", ("bn_",)),
    ("parameter server", ("ps_",)),
]


def launch_times(path):
    out = collections.defaultdict(float)
    n = 0
    for r in csv.reader(open(path)):
        if len(r) < 10 or not r[0].isdigit():
            continue
        us = float(r[-1].replace(",", ""))
        us = us / 1000 if r[-2] == "ns" else us * 1000 if r[-2] == "ms" else us
        group = next((g for g, pats in GROUPS if any(p in r[4] for p in pats)), "other")
        out[group] += us
        n += 1
    return out, n


def network_work(name, batch):
    from pytorch_distributed_nn_b200.engine import layers as L
    from pytorch_distributed_nn_b200.engine.network import build_network
    net = build_network(name, batch, device="meta")
    fprop = dgrad = 0
    bn_act_bytes = 0
    for l in net._all_layers():
        if isinstance(l, L.Conv2d):
            fl = 2 * l.out_shape[0] * l.oh * l.ow * l.cout * l.k * l.k * l.cin
            fprop += fl
            dgrad += 0 if l.first else fl
        elif isinstance(l, L.BatchNorm):
            bn_act_bytes += math.prod(l.out_shape) * 2
    return {"fprop": fprop, "dgrad": dgrad, "wgrad": fprop, "bn_act_bytes": bn_act_bytes, "params": net.total}


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("csv")
    ap.add_argument("--network", default="ResNet18")
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--peaks", default=os.path.join(ROOT, "MEASURED_PEAKS.json"))
    a = ap.parse_args(argv)
    peaks = {"hbm_gbs": 6553.9, "bf16_tflops": 1642.2, "bf16_tflops_sustained": 1345.5}
    if os.path.exists(a.peaks):
        peaks.update({k: v for k, v in json.load(open(a.peaks)).items() if k in peaks})
    t, n = launch_times(a.csv)
    w = network_work(a.network, a.batch)
    total = sum(t.values())
    print(f"{n} launches, {total / 1e3:.2f} ms of kernels; {a.network} batch {a.batch}: {w['params'] / 1e6:.1f} M parameters")
    rows = []
    if t.get("fprop+dgrad"):
        fl = w["fprop"] + w["dgrad"]
        rows.append(("conv fprop + dgrad", f"{fl / 1e12:.2f} TFLOP", t["fprop+dgrad"], fl / t["fprop+dgrad"] / 1e6, "tflops"))
    if t.get("wgrad"):
        rows.append(("conv wgrad", f"{w['wgrad'] / 1e12:.2f} TFLOP", t["wgrad"], w["wgrad"] / t["wgrad"] / 1e6, "tflops"))
    if t.get("batchnorm"):
        # statistics (1 read) + apply (read, write) + backward sums (2 reads) + backward apply (2 reads, 1 write) = 7 passes,
        # plus the residual streams of the block outputs (about a quarter more on the ResNets)
        by = w["bn_act_bytes"] * 7 * 1.3
        rows.append(("batchnorm passes", f"~{by / 1e9:.1f} GB", t["batchnorm"], by / t["batchnorm"] / 1e3, "gbs"))
    for k in ("parameter server", "fc", "other"):
        if t.get(k):
            rows.append((k, "", t[k], None, None))
    for name, work, us, rate, kind in rows:
        if kind == "tflops":
            tail = f"{rate:7.0f} TFLOP/s = {rate / peaks['bf16_tflops']:.2f} of burst cuBLAS, {rate / peaks['bf16_tflops_sustained']:.2f} of sustained"
        elif kind == "gbs":
            tail = f"{rate / 1e3:7.2f} TB/s   = {rate / peaks['hbm_gbs']:.2f} of the HBM copy rate"
        else:
            tail = ""
        print(f"  {name:20s} {work:12s} {us / 1e3:6.2f} ms  {tail}")
    return t, w


if __name__ == "__main__":
    main()
