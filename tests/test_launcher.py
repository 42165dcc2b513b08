import os
import subprocess
import sys
import textwrap

import pytest

from b200dp.launch import b200run as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RUN = [sys.executable, os.path.join(ROOT, "bin", "b200run")]


def test_parses_the_reference_mpirun_argv():
    # argv of horovod/tensorflow-mnist.yaml:19-38
    argv = ["-np", "2", "--allow-run-as-root", "-bind-to", "none", "-map-by", "slot", "-x", "LD_LIBRARY_PATH",
            "-x", "PATH", "-mca", "pml", "ob1", "-mca", "btl", "^openib", "python", "/examples/tensorflow_mnist.py"]
    a = L.build_parser().parse_args(argv)
    assert a.np == 2 and a.allow_run_as_root and a.bind_to == "none" and a.map_by == "slot"
    assert a.exports == ["LD_LIBRARY_PATH", "PATH"] and a.mca == [["pml", "ob1"], ["btl", "^openib"]]
    assert a.command == ["python", "/examples/tensorflow_mnist.py"]


def test_hostfile_and_slot_mapping():
    hosts = L.parse_hostfile("# workers\nw0 slots=2\nw1 slots=1\n\nw2:3\n")
    assert hosts == [("w0", 2), ("w1", 1), ("w2", 3)]
    slots = L.map_by_slot(hosts, 4)
    assert [(s.host, s.local_rank, s.local_size) for s in slots] == [("w0", 0, 2), ("w0", 1, 2), ("w1", 0, 1), ("w2", 0, 1)]
    assert L.parse_hosts_arg("a:2,b") == [("a", 2), ("b", 1)]
    with pytest.raises(ValueError):
        L.map_by_slot([("a", 1)], 2)


def _run(args, code, timeout=120, env=None):
    e = dict(os.environ, **(env or {}))
    return subprocess.run(RUN + args + [sys.executable, "-c", textwrap.dedent(code)], capture_output=True,
                          text=True, timeout=timeout, env=e)


def test_env_export_and_rank_env():
    r = _run(["-np", "3", "-x", "FOO=bar", "-x", "PATH", "-mca", "pml", "ob1", "--tag-output"], """
        import os
        print(os.environ["RANK"], os.environ["WORLD_SIZE"], os.environ["LOCAL_RANK"], os.environ["OMPI_COMM_WORLD_RANK"],
              os.environ["FOO"], os.environ["MASTER_ADDR"])
    """)
    assert r.returncode == 0, r.stderr
    lines = sorted(r.stdout.splitlines())
    assert lines == [f"[1,{i}]<stdout>:{i} 3 {i} {i} bar 127.0.0.1" for i in range(3)]


def test_fail_fast_kills_survivors_and_propagates_code():
    r = _run(["-np", "3", "--grace-secs", "1"], """
        import os, sys, time
        if os.environ["RANK"] == "1":
            time.sleep(0.3); sys.exit(7)
        time.sleep(60)
    """, timeout=30)
    assert r.returncode == 7
    assert "rank 1 exited with code 7" in r.stderr


def test_max_restarts_reruns_job(tmp_path):
    marker = tmp_path / "m"
    r = _run(["-np", "2", "--max-restarts", "2"], f"""
        import os, sys
        if os.environ["RANK"] == "0" and os.environ["B200DP_RESTART_COUNT"] == "0":
            sys.exit(3)
        if os.environ["RANK"] == "0":
            open({str(marker)!r}, "w").write(os.environ["B200DP_RESTART_COUNT"])
    """, timeout=60)
    assert r.returncode == 0, r.stderr
    assert marker.read_text() == "1"
    assert "restart 1/2" in r.stderr


def test_rank_killed_by_signal_is_restarted(tmp_path):
    """A rank that dies of SIGKILL (what the OOM killer sends) or a native crash is a failure like any other: the job is
    torn down and restarted from the latest checkpoint.  Only a signal sent to the launcher itself ends the job."""
    marker = tmp_path / "m"
    r = _run(["-np", "2", "--max-restarts", "1", "--grace-secs", "1"], f"""
        import os, signal, time
        if os.environ["RANK"] == "1" and os.environ["B200DP_RESTART_COUNT"] == "0":
            os.kill(os.getpid(), signal.SIGKILL)
        if os.environ["B200DP_RESTART_COUNT"] == "0":
            time.sleep(30)
        if os.environ["RANK"] == "0":
            open({str(marker)!r}, "w").write(os.environ["B200DP_RESTART_COUNT"])
    """, timeout=60)
    assert r.returncode == 0, r.stderr
    assert marker.read_text() == "1"
    assert "restart 1/1" in r.stderr


def test_signal_to_the_launcher_is_not_restarted(tmp_path):
    import signal, time
    up = tmp_path / "up"
    code = f"import os, time; open({str(up)!r} + os.environ['RANK'], 'w').close(); time.sleep(60)"
    p = subprocess.Popen(RUN + ["-np", "2", "--max-restarts", "3", "--grace-secs", "1", sys.executable, "-c", code],
                         stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    deadline = time.time() + 60
    while time.time() < deadline and not (os.path.exists(str(up) + "0") and os.path.exists(str(up) + "1")):
        time.sleep(0.1)                     # the ranks run, so the launcher's own signal handlers are installed
    time.sleep(0.3)
    p.send_signal(signal.SIGTERM)
    out, err = p.communicate(timeout=30)
    assert p.returncode == 128 + signal.SIGTERM
    assert "restart" not in err


def test_no_command_is_an_error():
    r = subprocess.run(RUN + ["-np", "1"], capture_output=True, text=True)
    assert r.returncode == 2


def test_check_build_reports_data_planes():
    r = subprocess.run(RUN + ["--check-build"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert "[X] Gloo" in r.stdout and "libb200dp.so" in r.stdout and "PyTorch" in r.stdout


def test_mpirun_wdir_and_output_filename(tmp_path):
    wd = tmp_path / "wd"
    wd.mkdir()
    logs = tmp_path / "logs"
    r = _run(["-np", "2", "-wdir", str(wd), "--output-filename", str(logs)], """
        import os
        print("cwd", os.getcwd(), os.environ["RANK"])
    """, timeout=60)
    assert r.returncode == 0, r.stderr
    assert r.stdout.count(f"cwd {wd}") == 2
    assert sorted(os.listdir(logs)) == ["rank-0.log", "rank-1.log"] and f"cwd {wd} 1" in (logs / "rank-1.log").read_text()


def test_master_port_for_a_remote_rank0_is_not_probed_locally(monkeypatch):
    """The rendezvous port is bound by rank 0 on ITS host.  With rank 0 on this host a free port is probed; with rank 0 on
    another host a local probe would be meaningless, so a fixed default (or --master-port) is used."""
    seen = {}

    class FakeProc:
        pid = 1
        stdout = stderr = None

        def poll(self):
            return 0

        def wait(self, timeout=None):
            return 0

    def fake_popen(cmd, **kw):
        seen.setdefault("envs", []).append(kw.get("env") or {})
        seen.setdefault("cmds", []).append(cmd)
        import io
        p = FakeProc()
        p.stdout, p.stderr = io.BytesIO(b""), io.BytesIO(b"")
        return p

    monkeypatch.setattr(L.subprocess, "Popen", fake_popen)
    args = L.build_parser().parse_args(["-np", "2", "-H", "otherhost:2", "python", "x.py"])
    args.command = ["python", "x.py"]
    slots = L.map_by_slot(L.parse_hosts_arg("otherhost:2"), 2)
    assert L.launch_once(args, slots, 0) == 0
    joined = " ".join(" ".join(c) if isinstance(c, (list, tuple)) else str(c) for c in seen["cmds"])
    assert "MASTER_PORT=29512" in joined and "MASTER_ADDR=otherhost" in joined, joined[:400]
