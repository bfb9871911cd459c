#!/usr/bin/env python3
"""Chrome-trace (chrome://tracing, Perfetto) view of a run -- what ``--timeline_logging`` gave the TF study
(distributed_TF/src/distributed_train.py:284-286,306-310: one ``worker=<id>_timeline_iter=<step>.json`` per step), built from the
files every run already leaves in its ``--train-dir``:

* ``worker_rank<r>.jsonl``: per step, the phases fetch-weights / forward / backward (durations, record written at the end of
  the step) -> three spans per step on the worker's track, killed steps marked;
* ``master_rank0.jsonl``: per step, the step span, the gather span (message backends) or the device timeline of the master's
  kernels (device backend: microseconds from the first bucket's wait to each bucket's k-th arrival, the wait for the last one,
  the reduce + update + broadcast tail after it) -> spans and per-bucket instants on the master's track;
* ``timeline_master`` (``time_ms step event who`` rows): step starts and worker completions as instants.

    python tools/chrome_trace.py /path/to/train_dir [-o trace.json] [--per-step]

``--per-step`` additionally writes one file per step (``timeline_iter=<step>.json``) like the reference. Times are
microseconds since the first record of the run.
"""
import argparse
import glob
import json
import os


def _jsonl(path):
    if not os.path.exists(path):
        return []
    with open(path) as f:
        return [json.loads(l) for l in f if l.strip()]


def build_events(train_dir):
    master = [r for r in _jsonl(os.path.join(train_dir, "master_rank0.jsonl")) if "step" in r and "event" not in r]
    workers = {}
    for p in sorted(glob.glob(os.path.join(train_dir, "worker_rank*.jsonl"))):
        workers[int(os.path.basename(p)[len("worker_rank"):-len(".jsonl")])] = [r for r in _jsonl(p) if "step" in r and "event" not in r]
    stamps = [r["ts"] - r.get("step_ms", 0.0) / 1e3 for r in master] + [r["ts"] - r.get("total_s", 0.0) for rs in workers.values() for r in rs]
    if not stamps:
        return []
    t0 = min(stamps)
    us = lambda t: (t - t0) * 1e6                                      # noqa: E731
    ev = [{"ph": "M", "pid": 0, "tid": 0, "name": "thread_name", "args": {"name": "master (rank 0)"}}]

    def span(tid, name, start_us, dur_us, step, **args):
        ev.append({"ph": "X", "pid": 0, "tid": tid, "name": name, "ts": round(start_us, 3), "dur": round(max(dur_us, 0.0), 3),
                   "args": dict(step=step, **args)})

    def instant(tid, name, at_us, step, **args):
        ev.append({"ph": "i", "s": "t", "pid": 0, "tid": tid, "name": name, "ts": round(at_us, 3), "args": dict(step=step, **args)})

    for r in master:
        end = us(r["ts"])
        dur = r.get("step_ms", 0.0) * 1e3
        start = end - dur
        span(0, "step", start, dur, r["step"], selected=r.get("selected"), killed=r.get("killed"))
        if "gather_s" in r:                                           # message backends: first arrival -> threshold
            span(0, "gradient gather", end - r["gather_s"] * 1e6, r["gather_s"] * 1e6, r["step"])
        dt = r.get("device_timeline_us") or {}
        if dt:                                                        # device backend: the master kernels' own clock
            wait, tail = float(dt.get("wait_for_kth_arrival", 0)), float(dt.get("reduce_update_broadcast", 0))
            t_first = end - tail - wait                               # first bucket's wait kernel started here
            span(0, "wait for k-th arrival (last bucket)", t_first, wait, r["step"])
            span(0, "reduce + update + broadcast tail", end - tail, tail, r["step"])
            for b, t in enumerate(dt.get("bucket_ready_us", [])):
                instant(0, f"bucket {b} has k arrivals", t_first + float(t), r["step"], bucket=b)
    for rank, rs in sorted(workers.items()):
        ev.append({"ph": "M", "pid": 0, "tid": rank, "name": "thread_name", "args": {"name": f"worker (rank {rank})"}})
        for r in rs:
            end = us(r["ts"])
            tot = r.get("total_s", 0.0) * 1e6
            start = end - tot
            f, fw, bw = (r.get(k, 0.0) * 1e6 for k in ("fetch_s", "forward_s", "backward_s"))
            if f or fw or bw:
                span(rank, "fetch weights", start, f, r["step"])
                span(rank, "forward", start + f, fw, r["step"], loss=r.get("loss"))
                span(rank, "backward + push" + (" (killed)" if r.get("killed") else ""), start + f + fw, bw, r["step"], killed=bool(r.get("killed")))
            else:
                span(rank, "step", start, tot, r["step"], loss=r.get("loss"))
    tl = os.path.join(train_dir, "timeline_master")
    if os.path.exists(tl):
        rows = [l.split() for l in open(tl) if l.strip()]
        # the timeline file has its own zero (monotonic clock of the master); align its first step_start with the first step span
        starts = [float(t) for t, s, e, w in rows if e == "step_start"]
        first_span = min((e["ts"] for e in ev if e.get("name") == "step" and e["tid"] == 0), default=0.0)
        off = first_span - (starts[0] * 1e3 if starts else 0.0)
        for t, s, e, w in rows:
            instant(0, e + (f" worker {w}" if e != "step_start" else ""), float(t) * 1e3 + off, int(s), who=int(w))
    return ev


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("train_dir")
    ap.add_argument("-o", "--out", default="")
    ap.add_argument("--per-step", action="store_true")
    a = ap.parse_args(argv)
    ev = build_events(a.train_dir)
    out = a.out or os.path.join(a.train_dir, "trace.json")
    with open(out, "w") as f:
        json.dump({"traceEvents": ev, "displayTimeUnit": "ms"}, f)
    files = [out]
    if a.per_step:
        steps = sorted({e["args"]["step"] for e in ev if e["ph"] != "M"})
        meta = [e for e in ev if e["ph"] == "M"]
        for s in steps:
            p = os.path.join(os.path.dirname(out) or ".", f"timeline_iter={s}.json")
            with open(p, "w") as f:
                json.dump({"traceEvents": meta + [e for e in ev if e["ph"] != "M" and e["args"]["step"] == s], "displayTimeUnit": "ms"}, f)
            files.append(p)
    print(f"{len(ev)} events -> {out}" + (f" (+{len(files) - 1} per-step files)" if a.per_step else ""))
    return files


if __name__ == "__main__":
    main()
