"""utils/guard.py: a multi-rank job run as watched children with fallback configurations (CPU, gloo).

Two launcher processes (ranks 0 and 1, environment as torchrun sets it, including TORCHELASTIC_USE_AGENT_STORE) each
run the same list of attempts: in the first one rank 1 never reaches the rendezvous (both children stop making
progress and are ended), the second raises on rank 0 only, the third completes an all-reduce.  Both launchers must end
up on the third attempt together."""

import json
import os
import socket
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = textwrap.dedent("""
    import importlib.util, json, os, sys, time
    # guard.py has no imports of its own package: load it by path so the first mark does not wait for `import torch`
    _spec = importlib.util.spec_from_file_location("guard", os.path.join(%(root)r, "cs336_a2_systems_b200", "utils", "guard.py"))
    _guard = importlib.util.module_from_spec(_spec); _spec.loader.exec_module(_guard)
    heartbeat = _guard.heartbeat
    mode = sys.argv[1]
    rank, ws = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    assert "TORCHELASTIC_USE_AGENT_STORE" not in os.environ
    heartbeat("started " + mode, echo=False)
    if mode == "hang" and rank == 1:
        time.sleep(3600)
    if mode == "raise" and rank == 0:
        sys.exit(7)
    import torch, torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    heartbeat("process group ready", echo=False)
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t)
    if rank == 0:
        print(json.dumps({"sum": float(t.item()), "mode": mode, "port": os.environ["MASTER_PORT"]}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
""")

LAUNCHER = textwrap.dedent("""
    import importlib.util, json, os, sys
    _spec = importlib.util.spec_from_file_location("guard", os.path.join(%(root)r, "cs336_a2_systems_b200", "utils", "guard.py"))
    _guard = importlib.util.module_from_spec(_spec); _spec.loader.exec_module(_guard)
    run_attempts = _guard.run_attempts
    rc, text, hist = run_attempts([sys.executable, %(child)r],
                                  [("a", ["hang"]), ("b", ["raise"]), ("c", ["ok"])], stall_s=25.0, total_s=150.0)
    print(json.dumps({"rc": rc, "text": text, "history": hist}), flush=True)
""")


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(600)
def test_two_launchers_fall_back_together(tmp_path):
    child = tmp_path / "child.py"
    child.write_text(CHILD % {"root": ROOT})
    launcher = tmp_path / "launcher.py"
    launcher.write_text(LAUNCHER % {"root": ROOT, "child": str(child)})
    port = 20000 + _free_port() % 20000
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), TORCHELASTIC_USE_AGENT_STORE="True")
        procs.append(subprocess.Popen([sys.executable, str(launcher)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=560) for p in procs]
    res = []
    for (so, se), p in zip(outs, procs):
        assert p.returncode == 0, se
        res.append(json.loads([l for l in so.splitlines() if l.startswith("{")][-1]))
    for r in res:
        assert r["rc"] == 0
        assert [h["config"] for h in r["history"]] == ["a", "b", "c"]
        assert r["history"][0]["outcome"].startswith("no progress for")
        assert r["history"][2]["outcome"] == "ok"
    assert res[0]["history"][1]["outcome"] == "exit code 7"
    line = json.loads(res[0]["text"].strip().splitlines()[-1])
    assert line["sum"] == 3.0 and line["mode"] == "ok" and int(line["port"]) != port
    assert res[1]["text"].strip() == ""


def test_all_attempts_fail(tmp_path):
    from cs336_a2_systems_b200.utils.guard import run_attempts

    script = tmp_path / "bad.py"
    script.write_text("import sys; sys.exit(int(sys.argv[1]))\n")
    rc, text, hist = run_attempts([sys.executable, str(script)], [("x", ["3"]), ("y", ["5"])], stall_s=0.5, total_s=20,
                                  log=lambda m: None)
    assert rc == 1 and text == ""
    assert [h["outcome"] for h in hist] == ["exit code 3", "exit code 5"]


def test_child_that_hangs_after_its_result_counts_as_done(tmp_path):
    from cs336_a2_systems_b200.utils.guard import run_attempts

    script = tmp_path / "late.py"
    script.write_text(textwrap.dedent("""
        import importlib.util, os, sys, time
        spec = importlib.util.spec_from_file_location("guard", os.path.join(%r, "cs336_a2_systems_b200", "utils", "guard.py"))
        guard = importlib.util.module_from_spec(spec); spec.loader.exec_module(guard)
        heartbeat, DONE_MARK = guard.heartbeat, guard.DONE_MARK
        print('{"value": 1}', flush=True)
        heartbeat(DONE_MARK, echo=False)
        time.sleep(3600)
    """ % ROOT))
    rc, text, hist = run_attempts([sys.executable, str(script)], [("only", [])], stall_s=5, total_s=60, log=lambda m: None)
    assert rc == 0 and json.loads(text.strip())["value"] == 1 and hist[0]["outcome"] == "ok"
