"""Run a multi-rank job as a child process per rank, with a stall detector and a list of fallback configurations.

Why: a collective job that stops making progress (a rank that never arrives at a barrier, a kernel that waits for a
peer) cannot be recovered inside the process — the CUDA context is busy until the process exits.  A launcher that is
itself one rank of a ``torchrun`` job can, however, run the real work in a child of its own, watch the child's progress
marks, and when they stop end exactly that child (its own process group id) and start the next configuration on a new
rendezvous port.  Every rank's launcher applies the same rule to the same marks, so all ranks move on together.

    parent (rank r of torchrun)                      child (rank r of its own process group)
    ---------------------------                      ---------------------------------------
    for each configuration:
        spawn child, MASTER_PORT = base + k   --->   init_process_group, build, warm up, time, print JSON (rank 0)
        poll: exit code / heartbeat file      <---   heartbeat(msg) after every phase
        no new mark for `stall_s`  -> kill child's process group, next configuration
        exit code 0                -> relay the child's stdout, done

The reference has no counterpart (its benchmark scripts run one configuration and rely on the NCCL watchdog,
cs336_systems/ddp_bucketed_overlapped_sharded.py:84-127).
"""

from __future__ import annotations

import os
import signal
import subprocess
import sys
import tempfile
import time

HEARTBEAT_ENV = "B200SYS_HEARTBEAT_FILE"
#: a child whose last mark is this one has delivered its result: ending it afterwards counts as success
DONE_MARK = "done"
_t0 = time.monotonic()


def heartbeat(msg: str, echo: bool = True) -> None:
    """Child side: record a progress mark (appends one line to the launcher's heartbeat file when there is one) and,
    on rank 0, echo it to stderr with the seconds since the process started."""
    dt = time.monotonic() - _t0
    path = os.environ.get(HEARTBEAT_ENV)
    if path:
        try:
            with open(path, "a") as f:
                f.write(f"{dt:.1f} {msg}\n")
        except OSError:
            pass
    if echo and int(os.environ.get("RANK", "0")) == 0:
        print(f"[bench +{dt:.1f}s] {msg}", file=sys.stderr, flush=True)


def _kill_group(proc: subprocess.Popen) -> None:
    """End the child and everything it started: the child was made the leader of its own session/process group."""
    try:
        os.killpg(proc.pid, signal.SIGKILL)
    except (ProcessLookupError, PermissionError):
        pass
    try:
        proc.wait(timeout=10)
    except subprocess.TimeoutExpired:  # pragma: no cover
        pass


def child_env(attempt: int, base_port: int | None = None) -> dict:
    """Environment of attempt ``attempt``: its own rendezvous port (the same on every rank), and rank 0 of the child
    group hosts the store itself (under torchrun the agent's store is only reachable on the agent's port)."""
    env = dict(os.environ)
    if base_port is None:
        base_port = int(env.get("MASTER_PORT", "29541"))
    port = 20000 + (base_port - 20000 + 101 + 13 * attempt) % 40000
    env["MASTER_PORT"] = str(port)
    env.setdefault("MASTER_ADDR", "127.0.0.1")
    env.pop("TORCHELASTIC_USE_AGENT_STORE", None)
    env["B200SYS_GUARD_ATTEMPT"] = str(attempt)
    return env


def run_attempts(base_cmd: list[str], attempts: list[tuple], *, stall_s: float = 75.0,
                 total_s: float = 300.0, poll_s: float = 0.25, log=None) -> tuple[int, str, list[dict]]:
    """Run ``base_cmd + extra`` for each ``(label, extra)`` or ``(label, extra, extra_env)`` in ``attempts`` until one
    exits 0.

    Returns ``(exit code, stdout of the successful child or "", history)``; ``history`` has one entry per attempt:
    ``{"config": label, "outcome": "ok" | "exit code N" | "no progress for S s after '<last mark>'" | ...}``.
    A child is ended when its heartbeat file has not grown for ``stall_s`` seconds or it has run ``total_s`` seconds.
    """
    log = log or (lambda m: print(m, file=sys.stderr, flush=True))
    history: list[dict] = []
    current: list[subprocess.Popen] = []

    def on_term(signum, _frame):  # torchrun ends its workers with SIGTERM: take the child down with us
        for p in current:
            _kill_group(p)
        os._exit(128 + signum)

    old = {s: signal.signal(s, on_term) for s in (signal.SIGTERM, signal.SIGINT)}
    try:
        for k, att in enumerate(attempts):
            label, extra = att[0], list(att[1])
            hb = tempfile.NamedTemporaryFile("w", prefix="b200_hb_", suffix=".log", delete=False)
            hb.close()
            env = child_env(k)
            if len(att) > 2 and att[2]:
                env.update(att[2])
            env[HEARTBEAT_ENV] = hb.name
            out = tempfile.TemporaryFile("w+")
            t_start = time.monotonic()
            proc = subprocess.Popen(base_cmd + extra, stdout=out, stderr=None, env=env, start_new_session=True)
            current[:] = [proc]
            last_size, t_mark, outcome = 0, t_start, None
            while True:
                rc = proc.poll()
                if rc is not None:
                    outcome = "ok" if rc == 0 else f"exit code {rc}"
                    break
                now = time.monotonic()
                try:
                    size = os.path.getsize(hb.name)
                except OSError:
                    size = last_size
                if size != last_size:
                    last_size, t_mark = size, now
                if now - t_mark > stall_s or now - t_start > total_s:
                    _kill_group(proc)
                    try:
                        marks = open(hb.name).read().strip().splitlines()
                    except OSError:
                        marks = []
                    where = marks[-1].split(" ", 1)[-1] if marks else "start"
                    if where == DONE_MARK:  # the result is out; only the tear-down of the process group did not return
                        outcome = "ok"
                        break
                    outcome = (f"no progress for {now - t_mark:.0f} s after '{where}'" if now - t_mark > stall_s
                               else f"still running after {now - t_start:.0f} s (last mark '{where}')")
                    break
                time.sleep(poll_s)
            current[:] = []
            out.seek(0)
            text = out.read()
            out.close()
            try:
                os.unlink(hb.name)
            except OSError:
                pass
            history.append({"config": label, "outcome": outcome, "seconds": round(time.monotonic() - t_start, 1)})
            if outcome == "ok":
                return 0, text, history
            log(f"[guard rank {os.environ.get('RANK', '0')}] '{label}': {outcome}"
                + ("; trying the next configuration" if k + 1 < len(attempts) else ""))
            if k + 1 < len(attempts):
                # A rank whose child failed at once must not run ahead of the ranks whose children are still being
                # watched: everybody starts the next attempt `stall_s` after the last mark (marks are written after
                # collective phases, so they carry nearly the same time on every rank).
                wait = t_mark + stall_s - time.monotonic()
                if wait > 0:
                    time.sleep(wait)
        return 1, "", history
    finally:
        for s, h in old.items():
            signal.signal(s, h)
