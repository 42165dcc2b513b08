"""``b200run`` — single-box replacement for ``mpirun`` + the MPI Operator's launcher pod.

The reference launches ranks with
``mpirun -np 2 --allow-run-as-root -bind-to none -map-by slot -x LD_LIBRARY_PATH -x PATH
-mca pml ob1 -mca btl ^openib python /examples/tensorflow_mnist.py``
(``horovod/tensorflow-mnist.yaml:17-38``), Open MPI built at ``horovod/Dockerfile:51-61``
and sshd set up at ``:67-78``.  ``b200run`` accepts that exact argv (so the MPIJob
manifest only swaps the command name) and provides what mpirun provided:

* N ranks with rank/size/local-rank environment (``RANK``/``WORLD_SIZE``/``LOCAL_RANK``
  and the ``OMPI_COMM_WORLD_*`` names) and a TCP rendezvous address;
* ``-x VAR`` / ``-x VAR=value`` environment export; ``-mca k v`` accepted and ignored;
* ``-bind-to none|core|numa`` CPU affinity; ``-map-by slot`` rank order over a hostfile
  (``--hostfile`` lines ``host slots=N``; non-local hosts are reached over ssh);
* every rank's stdout/stderr forwarded line-atomically to the launcher's stdout/stderr
  (what Promtail tails), ``--tag-output`` for ``[1,<rank>]<stdout>:`` prefixes;
* fail-fast: first non-zero exit => SIGTERM, then SIGKILL, to every other rank, and the
  launcher exits with that code;
* ``--max-restarts N``: restart the whole job from the latest checkpoint (the recovery
  model of the reference stack: operator restarts launcher -> rank 0 restores -> broadcast).
"""
from __future__ import annotations

import argparse
import os
import shlex
import signal
import socket
import subprocess
import sys
import threading
import time
from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple


@dataclass
class Slot:
    host: str
    local_rank: int
    local_size: int


def parse_hostfile(text: str) -> List[Tuple[str, int]]:
    out: List[Tuple[str, int]] = []
    for raw in text.splitlines():
        line = raw.split("#", 1)[0].strip()
        if not line:
            continue
        parts = line.split()
        host, slots = parts[0], 1
        for p in parts[1:]:
            if p.startswith("slots="):
                slots = int(p.split("=", 1)[1])
        if ":" in host and not parts[1:]:
            host, s = host.rsplit(":", 1)
            slots = int(s)
        out.append((host, slots))
    return out


def parse_hosts_arg(arg: str) -> List[Tuple[str, int]]:
    out = []
    for item in arg.split(","):
        item = item.strip()
        if not item:
            continue
        if ":" in item:
            h, s = item.rsplit(":", 1)
            out.append((h, int(s)))
        else:
            out.append((item, 1))
    return out


def map_by_slot(hosts: Sequence[Tuple[str, int]], np_: int) -> List[Slot]:
    """Fill each host's slots before moving to the next (``-map-by slot``)."""
    total = sum(s for _, s in hosts)
    if np_ > total:
        raise ValueError(f"-np {np_} exceeds the {total} slots in the host list")
    assign: List[Tuple[str, int]] = []
    for h, s in hosts:
        for i in range(s):
            if len(assign) < np_:
                assign.append((h, i))
    counts: Dict[str, int] = {}
    for h, _ in assign:
        counts[h] = counts.get(h, 0) + 1
    return [Slot(h, i, counts[h]) for h, i in assign]


def build_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(prog="b200run", add_help=True, allow_abbrev=False,
                                description="mpirun-compatible launcher for b200dp jobs")
    p.add_argument("-np", "-n", "--np", dest="np", type=int, default=None)
    p.add_argument("--allow-run-as-root", action="store_true")
    p.add_argument("-bind-to", "--bind-to", dest="bind_to", default="none", choices=["none", "core", "numa", "socket"])
    p.add_argument("-map-by", "--map-by", dest="map_by", default="slot")
    p.add_argument("-x", dest="exports", action="append", default=[])
    p.add_argument("-mca", "--mca", dest="mca", nargs=2, action="append", default=[], metavar=("KEY", "VALUE"))
    p.add_argument("-H", "-host", "--host", dest="hosts", default=None)
    p.add_argument("-hostfile", "--hostfile", "-machinefile", dest="hostfile", default=None)
    p.add_argument("--tag-output", "-tag-output", dest="tag_output", action="store_true")
    p.add_argument("--master-addr", default=None)
    p.add_argument("--master-port", type=int, default=None)
    p.add_argument("--max-restarts", type=int, default=0)
    p.add_argument("--grace-secs", type=float, default=5.0)
    p.add_argument("--resize-file", default=None,
                   help="elastic resize: ranks see this path as B200DP_RESIZE_FILE; when it exists (content: the new -np) the next "
                        "hvd.elastic commit exits with code 75 and the job is relaunched with that many ranks")
    p.add_argument("--cpu", action="store_true", help="force the CPU/gloo backend in every rank")
    p.add_argument("--log-dir", "--output-filename", "-output-filename", dest="log_dir", default=None,
                   help="also write each rank's output to <dir>/rank-<r>.log (mpirun: --output-filename)")
    p.add_argument("-wdir", "--wdir", "-wd", dest="wdir", default=None, help="working directory of every rank (mpirun: -wdir)")
    p.add_argument("--timeline", default=None, help="sets B200DP_TIMELINE for every rank")
    p.add_argument("--check-build", "-cb", dest="check_build", action="store_true",
                   help="print what this installation was built with (horovodrun --check-build) and exit")
    p.add_argument("command", nargs=argparse.REMAINDER)
    return p


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _is_local(host: str) -> bool:
    return host in ("localhost", "127.0.0.1", "::1", socket.gethostname(), socket.getfqdn())


def _cpu_sets(n_local: int, mode: str) -> List[Optional[set]]:
    if mode == "none":
        return [None] * n_local
    try:
        avail = sorted(os.sched_getaffinity(0))
    except AttributeError:
        return [None] * n_local
    if mode in ("numa", "socket"):
        nodes = []
        base = "/sys/devices/system/node"
        if os.path.isdir(base):
            for d in sorted(os.listdir(base)):
                if d.startswith("node") and d[4:].isdigit():
                    try:
                        txt = open(os.path.join(base, d, "cpulist")).read().strip()
                        cpus = set()
                        for part in txt.split(","):
                            if "-" in part:
                                a, b = part.split("-")
                                cpus.update(range(int(a), int(b) + 1))
                            elif part:
                                cpus.add(int(part))
                        cpus &= set(avail)
                        if cpus:
                            nodes.append(cpus)
                    except OSError:
                        pass
        if nodes:
            per = max(1, -(-n_local // len(nodes)))
            return [nodes[min(i // per, len(nodes) - 1)] for i in range(n_local)]
    per = max(1, len(avail) // max(n_local, 1))
    return [set(avail[i * per:(i + 1) * per]) or None for i in range(n_local)]


class _Pump(threading.Thread):
    def __init__(self, pipe, sink, prefix: bytes, lock: threading.Lock, tee=None):
        super().__init__(daemon=True)
        self.pipe, self.sink, self.prefix, self.lock, self.tee = pipe, sink, prefix, lock, tee

    def run(self):
        for line in iter(self.pipe.readline, b""):
            with self.lock:
                self.sink.write(self.prefix + line if self.prefix else line)
                self.sink.flush()
                if self.tee:
                    self.tee.write(line)
                    self.tee.flush()
        self.pipe.close()


def launch_once(args, slots: List[Slot], restart_count: int) -> int:
    np_ = len(slots)
    master_addr = args.master_addr or ("127.0.0.1" if all(_is_local(s.host) for s in slots) else slots[0].host)
    if args.master_port:
        master_port = args.master_port
    elif _is_local(slots[0].host):
        master_port = _free_port()          # rank 0 binds it on this host: probing here is meaningful
    else:
        # rank 0 lives on another host, where a port probed here says nothing: a fixed, documented default (override
        # with --master-port), spread by the restart count so a relaunch does not meet the old socket in TIME_WAIT
        master_port = 29512 + (restart_count % 16)
    base_env = dict(os.environ)
    exports: Dict[str, str] = {}
    for e in args.exports:
        if "=" in e:
            k, v = e.split("=", 1)
            exports[k] = v
        elif e in os.environ:
            exports[e] = os.environ[e]
    n_local = sum(1 for s in slots if _is_local(s.host))
    cpu_sets = _cpu_sets(n_local, args.bind_to)
    procs: List[subprocess.Popen] = []
    pumps: List[_Pump] = []
    lock = threading.Lock()
    tees = []
    out_b, err_b = sys.stdout.buffer, sys.stderr.buffer
    if args.log_dir:
        os.makedirs(args.log_dir, exist_ok=True)
    local_i = 0
    for r, s in enumerate(slots):
        env_rank = {
            "RANK": str(r), "WORLD_SIZE": str(np_), "LOCAL_RANK": str(s.local_rank),
            "LOCAL_WORLD_SIZE": str(s.local_size), "MASTER_ADDR": master_addr, "MASTER_PORT": str(master_port),
            "OMPI_COMM_WORLD_RANK": str(r), "OMPI_COMM_WORLD_SIZE": str(np_),
            "OMPI_COMM_WORLD_LOCAL_RANK": str(s.local_rank), "OMPI_COMM_WORLD_LOCAL_SIZE": str(s.local_size),
            "B200DP_RESTART_COUNT": str(restart_count), "PYTHONUNBUFFERED": "1",
        }
        if args.cpu:
            env_rank["B200DP_FORCE_CPU"] = "1"
        if args.timeline:
            env_rank["B200DP_TIMELINE"] = args.timeline
        if args.resize_file:
            env_rank["B200DP_RESIZE_FILE"] = os.path.abspath(args.resize_file)
        if _is_local(s.host):
            env = dict(base_env)
            env.update(exports)
            env.update(env_rank)
            cpus = cpu_sets[local_i] if local_i < len(cpu_sets) else None
            local_i += 1

            def _pre(cpus=cpus):
                os.setsid()
                if cpus:
                    try:
                        os.sched_setaffinity(0, cpus)
                    except OSError:
                        pass
            p = subprocess.Popen(args.command, env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                 preexec_fn=_pre, cwd=args.wdir or None)
        else:
            envs = " ".join(f"{k}={shlex.quote(v)}" for k, v in {**exports, **env_rank}.items())
            remote = f"cd {shlex.quote(args.wdir or os.getcwd())} && env {envs} {' '.join(shlex.quote(c) for c in args.command)}"
            p = subprocess.Popen(["ssh", "-o", "StrictHostKeyChecking=no", "-o", "UserKnownHostsFile=/dev/null",
                                  s.host, remote], stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                 preexec_fn=os.setsid)
        procs.append(p)
        tee = open(os.path.join(args.log_dir, f"rank-{r}.log"), "ab") if args.log_dir else None
        tees.append(tee)
        po = f"[1,{r}]<stdout>:".encode() if args.tag_output else b""
        pe = f"[1,{r}]<stderr>:".encode() if args.tag_output else b""
        pumps += [_Pump(p.stdout, out_b, po, lock, tee), _Pump(p.stderr, err_b, pe, lock, tee)]
    for t in pumps:
        t.start()

    def kill_all(sig):
        for p in procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, sig)
                except (ProcessLookupError, PermissionError):
                    try:
                        p.send_signal(sig)
                    except ProcessLookupError:
                        pass

    stop = {"sig": None}

    def on_signal(signum, _frame):
        stop["sig"] = signum
    old_int = signal.signal(signal.SIGINT, on_signal)
    old_term = signal.signal(signal.SIGTERM, on_signal)
    exit_code = 0
    try:
        alive = set(range(np_))
        while alive:
            if stop["sig"] is not None:
                exit_code = 128 + stop["sig"]
                args._launcher_signalled = True      # the operator (or the pod) stopped the JOB: never restart after this
                break
            for r in list(alive):
                rc = procs[r].poll()
                if rc is not None:
                    alive.discard(r)
                    if rc != 0 and exit_code == 0:
                        exit_code = rc if rc > 0 else 128 - rc
                        sys.stderr.write(f"b200run: rank {r} exited with code {rc}; terminating {len(alive)} remaining rank(s)\n")
                        sys.stderr.flush()
            if exit_code != 0:
                break
            time.sleep(0.02)
        if exit_code != 0:
            kill_all(signal.SIGTERM)
            deadline = time.time() + args.grace_secs
            while time.time() < deadline and any(p.poll() is None for p in procs):
                time.sleep(0.05)
            kill_all(signal.SIGKILL)
        for p in procs:
            try:
                p.wait(timeout=10)
            except subprocess.TimeoutExpired:
                pass
    finally:
        signal.signal(signal.SIGINT, old_int)
        signal.signal(signal.SIGTERM, old_term)
        for t in pumps:
            t.join(timeout=5)
        for tee in tees:
            if tee:
                tee.close()
    return exit_code


def check_build_report() -> str:
    """What ``horovodrun --check-build`` prints, for this framework: data planes, kernels library, devices."""
    import torch
    from ..ops import _lib
    from ..runtime import world as _w
    box = lambda b: "[X]" if b else "[ ]"
    lib_ok = _lib.available()
    lines = ["b200dp build:", "", "Available frameworks:", f"    {box(True)} PyTorch {torch.__version__}",
             f"    {box(False)} TensorFlow (the reference scripts' names are provided by `import b200dp as hvd`)", "",
             "Available controllers:", f"    {box(False)} MPI (b200run accepts mpirun's argv)", f"    {box(_w.gloo_built())} Gloo", "",
             "Available tensor operations:",
             f"    {box(lib_ok)} NVLink peer-memory kernels (libb200dp.so, sm_100a){'' if lib_ok else ' - run tools/build.py'}",
             f"    {box(False)} NCCL (baseline harness only)", f"    {box(_w.gloo_built())} Gloo (CPU tensors, control plane)", "",
             f"library: {_lib.lib_path()}"]
    n = torch.cuda.device_count() if torch.cuda.is_available() else 0
    lines.append(f"CUDA devices visible: {n}")
    for d in range(n):
        p = torch.cuda.get_device_properties(d)
        lines.append(f"    {d}: {p.name}, sm_{p.major}{p.minor}, {p.multi_processor_count} SMs, {p.total_memory / 2**30:.0f} GiB")
    return "\n".join(lines) + "\n"


def main(argv: Optional[Sequence[str]] = None) -> int:
    args = build_parser().parse_args(list(argv) if argv is not None else None)
    cmd = list(args.command)
    if cmd and cmd[0] == "--":
        cmd = cmd[1:]
    if args.check_build:
        sys.stdout.write(check_build_report())
        return 0
    if not cmd:
        sys.stderr.write("b200run: no command given\n")
        return 2
    args.command = cmd
    if args.hostfile:
        hosts = parse_hostfile(open(args.hostfile).read())
    elif args.hosts:
        hosts = parse_hosts_arg(args.hosts)
    else:
        hosts = [("localhost", args.np or 1)]
    np_ = args.np or sum(s for _, s in hosts)
    if args.map_by not in ("slot", "core", "node", "numa", "socket"):
        sys.stderr.write(f"b200run: unsupported -map-by {args.map_by}\n")
        return 2
    slots = map_by_slot(hosts, np_)
    restarts = 0
    incarnation = 0
    while True:
        rc = launch_once(args, slots, incarnation)
        if rc == 75 and args.resize_file and os.path.exists(args.resize_file):
            # elastic resize requested at a commit point: relaunch with the new world size (not a failure: no restart is spent)
            try:
                new_np = int(open(args.resize_file).read().strip() or np_)
            except ValueError:
                new_np = np_
            os.remove(args.resize_file)
            if not args.hostfile and not args.hosts:
                hosts = [("localhost", new_np)]
            try:
                slots = map_by_slot(hosts, new_np)
            except Exception as e:
                sys.stderr.write(f"b200run: cannot resize to {new_np} ranks: {e}\n")
                return 2
            sys.stderr.write(f"b200run: resizing {np_} -> {new_np} ranks at the last commit\n")
            sys.stderr.flush()
            np_ = new_np
            incarnation += 1
            continue
        # A rank that died of a signal (OOM-killer SIGKILL -> 137, SIGSEGV -> 139, SIGABRT -> 134) is a failure like any
        # other and is restarted from the latest checkpoint; only a signal delivered to the launcher itself ends the job.
        if rc == 0 or restarts >= args.max_restarts or getattr(args, "_launcher_signalled", False):
            return rc
        incarnation += 1
        restarts += 1
        sys.stderr.write(f"b200run: job failed with code {rc}; restart {restarts}/{args.max_restarts} from the latest checkpoint\n")
        sys.stderr.flush()
        time.sleep(0.5)


if __name__ == "__main__":
    sys.exit(main())
