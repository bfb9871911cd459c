#!/usr/bin/env python3
"""Cluster tooling (counterpart of tools/pytorch_ec2.py + local_script.sh / remote_script.sh and distributed_TF/tools/tf_ec2.py).

The reference's tooling rents EC2 instances with boto3, mounts NFS, writes ``hosts`` / ``hosts_alias`` /
``hosts_address`` files and starts ``mpirun --hostfile``. The B200 target is one NVSwitch box (or a few boxes reached
by ssh), so the same verbs act on a *host list* instead of an AWS account:

    python tools/cluster.py get_hosts  --cfg configs/cluster_local.cfg      # write hosts, hosts_alias, hosts_address
    python tools/cluster.py launch     --cfg ... -- --network ResNet18 ...  # torchrun on every host (ssh for remote ones)
    python tools/cluster.py run_command --cfg ... "nvidia-smi -L"
    python tools/cluster.py kill_all_python --cfg ...                       # kills only the PIDs this tool started
    python tools/cluster.py list_running --cfg ...
    python tools/cluster.py shutdown   --cfg ...                            # kill + remove the state file
    python tools/cluster.py download_file   --cfg ... <remote path> <local dir>   # one file from every host
    python tools/cluster.py download_outdir --cfg ... <remote dir>  <local dir>   # a log / checkpoint directory

The config is a python-literal dict with ``%(key)s`` self-interpolation, like the reference's ``Cfg``
(tools/pytorch_ec2.py:12-20). Nothing is ever killed by pattern: PIDs (process groups) are recorded in a state file.
"""
import argparse
import ast
import json
import os
import shlex
import signal
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class Cfg(dict):
    """dict whose string values may refer to other keys with %(key)s (resolved at lookup)."""

    def __getitem__(self, item):
        v = dict.__getitem__(self, item)
        if isinstance(v, str):
            for _ in range(8):
                if "%(" not in v:
                    break
                v = v % self
        return v


def load_cfg(path):
    with open(path) as f:
        raw = ast.literal_eval(f.read())          # literal_eval, not eval: a config cannot run code
    cfg = Cfg(raw)
    cfg.setdefault("name", "pdnn")
    cfg.setdefault("hosts", ["127.0.0.1"])
    cfg.setdefault("gpus_per_host", 8)
    cfg.setdefault("master_port", 29500)
    cfg.setdefault("ssh_user", os.environ.get("USER", "root"))
    cfg.setdefault("repo_dir", ROOT)
    cfg.setdefault("state_file", os.path.join("/tmp", f"{cfg['name']}_cluster_state.json"))
    cfg.setdefault("python", sys.executable)
    cfg.setdefault("entry", "distributed_nn.py")
    return cfg


def is_local(host):
    return host in ("127.0.0.1", "localhost")


def get_hosts(cfg):
    hosts = cfg["hosts"]
    with open("hosts", "w") as f:
        for i, h in enumerate(hosts):
            f.write(f"{h} deeplearning-worker{i}\n")
    with open("hosts_alias", "w") as f:
        for i in range(len(hosts)):
            f.write(f"deeplearning-worker{i}\n")
    with open("hosts_address", "w") as f:
        for h in hosts:
            f.write(f"{h} slots={cfg['gpus_per_host']}\n")
    print(f"wrote hosts, hosts_alias, hosts_address for {len(hosts)} host(s)")
    return hosts


def _spawn(cfg, host, cmd, log):
    if is_local(host):
        full = cmd
    else:
        full = f"ssh -o StrictHostKeyChecking=no {cfg['ssh_user']}@{host} {shlex.quote('cd ' + cfg['repo_dir'] + ' && ' + cmd)}"
    with open(log, "w") as lf:
        p = subprocess.Popen(full, shell=True, stdout=lf, stderr=subprocess.STDOUT, start_new_session=True, cwd=cfg["repo_dir"] if is_local(host) else None)
    return p.pid


def launch(cfg, extra):
    hosts = cfg["hosts"]
    nnodes = len(hosts)
    state = {"procs": []}
    for node, host in enumerate(hosts):
        cmd = (f"{cfg['python']} -m torch.distributed.run --nnodes={nnodes} --node-rank={node} --nproc-per-node={cfg['gpus_per_host']} "
               f"--master-addr {hosts[0]} --master-port {cfg['master_port']} {cfg['entry']} {' '.join(shlex.quote(a) for a in extra)}")
        log = f"out_node{node}"
        pid = _spawn(cfg, host, cmd, log)
        state["procs"].append({"host": host, "pid": pid, "log": log, "cmd": cmd})
        print(f"[{host}] pid {pid} -> {log}")
    with open(cfg["state_file"], "w") as f:
        json.dump(state, f, indent=1)


def _load_state(cfg):
    if not os.path.exists(cfg["state_file"]):
        return {"procs": []}
    with open(cfg["state_file"]) as f:
        return json.load(f)


def list_running(cfg):
    for p in _load_state(cfg)["procs"]:
        alive = False
        if is_local(p["host"]):
            try:
                os.kill(p["pid"], 0)
                alive = True
            except OSError:
                alive = False
        print(f"{p['host']} pid={p['pid']} alive={alive if is_local(p['host']) else 'unknown (remote)'} log={p['log']}")


def kill_all_python(cfg):
    """Stops exactly the process groups this tool started (recorded PIDs), never a name pattern."""
    for p in _load_state(cfg)["procs"]:
        if is_local(p["host"]):
            try:
                os.killpg(p["pid"], signal.SIGTERM)
                print(f"sent SIGTERM to process group {p['pid']}")
            except OSError as e:
                print(f"pid {p['pid']}: {e}")
        else:
            subprocess.call(f"ssh {cfg['ssh_user']}@{p['host']} kill -TERM -- -{p['pid']}", shell=True)


def run_command(cfg, command):
    for host in cfg["hosts"]:
        print(f"== {host}")
        if is_local(host):
            subprocess.call(command, shell=True)
        else:
            subprocess.call(f"ssh -o StrictHostKeyChecking=no {cfg['ssh_user']}@{host} {shlex.quote(command)}", shell=True)


def download(cfg, remote, local_dir, recursive):
    """Collect a file / directory from every host into local_dir/<host>/ (reference: tf_ec2.py download_file /
    download_outdir through the vendored scp module; here the system scp, or a plain copy for local hosts)."""
    import shutil
    done = []
    for host in cfg["hosts"]:
        dst = os.path.join(local_dir, host.replace(":", "_"))
        os.makedirs(dst, exist_ok=True)
        if is_local(host):
            src = remote if os.path.isabs(remote) else os.path.join(cfg["repo_dir"], remote)
            if not os.path.exists(src):
                print(f"{host}: {src} does not exist")
                continue
            target = os.path.join(dst, os.path.basename(src.rstrip("/")))
            if os.path.isdir(src):
                if not recursive:
                    print(f"{host}: {src} is a directory (use download_outdir)")
                    continue
                shutil.copytree(src, target, dirs_exist_ok=True)
            else:
                shutil.copy2(src, target)
            done.append(target)
        else:
            flag = "-r " if recursive else ""
            rc = subprocess.call(f"scp -o StrictHostKeyChecking=no {flag}{cfg['ssh_user']}@{host}:{shlex.quote(remote)} {shlex.quote(dst)}/", shell=True)
            if rc == 0:
                done.append(os.path.join(dst, os.path.basename(remote.rstrip("/"))))
    for d in done:
        print("downloaded", d)
    return done


def wait(cfg, timeout):
    t0 = time.time()
    while time.time() - t0 < timeout:
        alive = 0
        for p in _load_state(cfg)["procs"]:
            if is_local(p["host"]):
                try:
                    os.kill(p["pid"], 0)
                    alive += 1
                except OSError:
                    pass
        if alive == 0:
            return True
        time.sleep(1)
    return False


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("command", choices=["launch", "get_hosts", "shutdown", "kill_all_python", "kill_python", "run_command", "list_running", "wait",
                                        "download_file", "download_outdir"])
    ap.add_argument("--cfg", default=os.path.join(ROOT, "configs", "cluster_local.cfg"))
    ap.add_argument("--timeout", type=float, default=600)
    ap.add_argument("rest", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    cfg = load_cfg(a.cfg)
    rest = [r for r in a.rest if r != "--"]
    if a.command == "get_hosts":
        get_hosts(cfg)
    elif a.command == "launch":
        get_hosts(cfg)
        launch(cfg, rest)
    elif a.command in ("kill_all_python", "kill_python"):
        kill_all_python(cfg)
    elif a.command == "shutdown":
        kill_all_python(cfg)
        if os.path.exists(cfg["state_file"]):
            os.remove(cfg["state_file"])
    elif a.command == "run_command":
        run_command(cfg, " ".join(rest))
    elif a.command == "list_running":
        list_running(cfg)
    elif a.command == "wait":
        sys.exit(0 if wait(cfg, a.timeout) else 1)
    elif a.command in ("download_file", "download_outdir"):
        if len(rest) != 2:
            ap.error(f"{a.command} <remote path> <local dir>")
        download(cfg, rest[0], rest[1], recursive=(a.command == "download_outdir"))


if __name__ == "__main__":
    main()
