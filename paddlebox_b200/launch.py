"""``python -m paddlebox_b200.launch [--gpus 0,1,2,3 | --nproc_per_node N] [--ips a,b] train.py args...``

The job launcher: one trainer process per GPU of this node, with the environment both families of job scripts read.

Reference: ``python/paddle/distributed/fleet/launch.py`` (arguments :101-260, ``launch_collective`` / ``launch_ps``)
and ``launch_utils.py`` (``start_local_trainers``, ``watch_local_trainers``, ``terminate_local_procs``).  Each child
gets the Paddle launcher's variables — ``PADDLE_TRAINER_ID``, ``PADDLE_TRAINERS_NUM``, ``PADDLE_TRAINER_ENDPOINTS``,
``PADDLE_CURRENT_ENDPOINT``, ``TRAINING_ROLE=TRAINER``, ``FLAGS_selected_gpus`` (what ``role_maker.*`` read) — and the
``torch.distributed`` ones — ``RANK``, ``LOCAL_RANK``, ``WORLD_SIZE``, ``LOCAL_WORLD_SIZE``, ``MASTER_ADDR``,
``MASTER_PORT`` (what plain ``fleet.init()`` reads).  Output goes to ``<log_dir>/workerlog.<local rank>`` (rank 0 also
to the terminal).  The launcher watches its children: when one exits with an error the others are terminated (the
exact processes it started) and its exit code is returned; Ctrl-C / SIGTERM are passed on.  There are no
parameter-server processes to start (the table lives in the trainers' HBM): ``--server_num`` / ``--servers`` are
accepted and ignored with a note, ``--worker_num`` counts trainers.  Several nodes: run the same command on every
node with ``--ips`` listing all nodes in the same order (the first is the rendezvous) and ``--node_rank`` (or the
node's own address among ``--ips``).
"""
from __future__ import annotations

import argparse
import os
import signal
import socket
import subprocess
import sys
import time


def _parse(argv=None):
    ap = argparse.ArgumentParser(prog="python -m paddlebox_b200.launch", description=__doc__.split("\n\n")[1])
    ap.add_argument("--log_dir", default="log", help="directory of the per-process logs (workerlog.N)")
    ap.add_argument("--backend", default=os.environ.get("PADDLE_DISTRI_BACKEND", "auto"), help="nccl | gloo | auto")
    ap.add_argument("--nproc_per_node", type=int, default=None, help="processes on this node (default: one per visible GPU)")
    ap.add_argument("--run_mode", default=None, help="collective | ps (both start trainers only)")
    ap.add_argument("--gpus", "--selected_gpus", dest="gpus", default=None, help='e.g. "0,1,2,3": one process per listed GPU')
    ap.add_argument("--ips", default="127.0.0.1", help="addresses of all nodes, comma separated, same order everywhere")
    ap.add_argument("--node_rank", type=int, default=None, help="index of this node in --ips")
    ap.add_argument("--start_port", type=int, default=int(os.environ.get("FLAGS_START_PORT", "6170")))
    ap.add_argument("--master_port", type=int, default=None, help="rendezvous port (default: start_port - 1)")
    ap.add_argument("--worker_num", "--trainer_num", dest="worker_num", type=int, default=None)
    ap.add_argument("--workers", default="", help="accepted for compatibility (endpoints are derived from --ips)")
    ap.add_argument("--server_num", type=int, default=None)
    ap.add_argument("--servers", default="")
    ap.add_argument("--heter_worker_num", default=None)
    ap.add_argument("--http_port", type=int, default=None)
    ap.add_argument("training_script")
    ap.add_argument("training_script_args", nargs=argparse.REMAINDER)
    return ap.parse_args(argv)


def _visible_gpus() -> list:
    env = os.environ.get("CUDA_VISIBLE_DEVICES")
    if env is not None:
        return [g for g in env.split(",") if g.strip() != ""]
    try:
        import torch
        return [str(i) for i in range(torch.cuda.device_count())]
    except Exception:
        return []


def plan(args) -> dict:
    """Who runs where: the per-process environments of THIS node."""
    ips = [ip for ip in args.ips.split(",") if ip]
    gpus = [g for g in args.gpus.split(",") if g != ""] if args.gpus else _visible_gpus()
    nproc = args.nproc_per_node or args.worker_num or len(gpus) or 1
    if args.gpus and args.nproc_per_node and args.nproc_per_node > len(gpus):
        raise SystemExit(f"--nproc_per_node {args.nproc_per_node} exceeds the {len(gpus)} GPUs of --gpus")
    node_rank = args.node_rank
    if node_rank is None:
        node_rank = 0
        if len(ips) > 1:
            mine = {socket.gethostname(), "127.0.0.1"}
            try:
                mine.update(socket.gethostbyname_ex(socket.gethostname())[2])
            except OSError:
                pass
            found = [i for i, ip in enumerate(ips) if ip in mine]
            if not found:
                raise SystemExit("cannot tell which of --ips this node is: pass --node_rank")
            node_rank = found[0]
    endpoints = [f"{ip}:{args.start_port + i}" for ip in ips for i in range(nproc)]
    world = len(endpoints)
    master_port = args.master_port or (args.start_port - 1)
    envs = []
    for local in range(nproc):
        rank = node_rank * nproc + local
        e = {"PADDLE_TRAINER_ID": str(rank), "PADDLE_TRAINERS_NUM": str(world), "PADDLE_TRAINER_ENDPOINTS": ",".join(endpoints),
             "PADDLE_CURRENT_ENDPOINT": endpoints[rank], "TRAINING_ROLE": "TRAINER", "PADDLE_TRAINING_ROLE": "TRAINER",
             "PADDLE_RANK_IN_NODE": str(local), "PADDLE_LOCAL_DEVICE_IDS": gpus[local] if local < len(gpus) else "",
             "PADDLE_WORLD_DEVICE_IDS": ",".join(gpus), "POD_IP": ips[node_rank],
             "RANK": str(rank), "LOCAL_RANK": str(local), "WORLD_SIZE": str(world), "LOCAL_WORLD_SIZE": str(nproc),
             "MASTER_ADDR": ips[0], "MASTER_PORT": str(master_port)}
        if local < len(gpus):
            e["FLAGS_selected_gpus"] = gpus[local]
        if args.backend != "auto":
            e["PADDLE_DISTRI_BACKEND"] = args.backend
        envs.append(e)
    return {"envs": envs, "world": world, "node_rank": node_rank, "endpoints": endpoints}


def _terminate(procs) -> None:
    for p in procs:
        if p.poll() is None:
            p.terminate()
    deadline = time.time() + 10
    for p in procs:
        while p.poll() is None and time.time() < deadline:
            time.sleep(0.05)
        if p.poll() is None:
            p.kill()


def launch(argv=None) -> int:
    args = _parse(argv)
    if args.server_num or args.servers:
        print("[launch] no parameter-server processes are started: the sparse table lives in the trainers' HBM",
              file=sys.stderr)
    pl = plan(args)
    os.makedirs(args.log_dir, exist_ok=True)
    procs, logs = [], []
    for local, extra in enumerate(pl["envs"]):
        env = {**os.environ, **extra}
        cmd = [sys.executable, "-u", args.training_script] + list(args.training_script_args)
        log = open(os.path.join(args.log_dir, f"workerlog.{local}"), "w")
        logs.append(log)
        if local == 0:                       # rank 0 of the node is shown on the terminal as well
            p = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        else:
            p = subprocess.Popen(cmd, env=env, stdout=log, stderr=subprocess.STDOUT)
        procs.append(p)

    def on_signal(signum, frame):
        _terminate(procs)
        raise SystemExit(128 + signum)
    old = {s: signal.signal(s, on_signal) for s in (signal.SIGINT, signal.SIGTERM)}
    rc = 0
    try:
        import threading

        def tee():
            for line in procs[0].stdout:
                sys.stdout.write(line)
                logs[0].write(line)
            logs[0].flush()
        t = threading.Thread(target=tee, daemon=True)
        t.start()
        alive = set(range(len(procs)))
        while alive:
            for i in sorted(alive):
                r = procs[i].poll()
                if r is None:
                    continue
                alive.discard(i)
                if r != 0 and rc == 0:
                    rc = r
                    print(f"[launch] trainer {pl['envs'][i]['PADDLE_TRAINER_ID']} exited with code {r}: stopping the others "
                          f"(see {args.log_dir}/workerlog.{i})", file=sys.stderr)
                    _terminate([procs[j] for j in alive])
            time.sleep(0.05)
        t.join(timeout=5)
    finally:
        for s, h in old.items():
            signal.signal(s, h)
        for log in logs:
            log.close()
    return rc if rc >= 0 else 128 - rc       # a child killed by signal n reports -n


if __name__ == "__main__":
    sys.exit(launch())
