"""Cooperative pre-emption client.

The reference polls ``hfai.client.receive_suspend_command()`` every step and, when it
fires, checkpoints and calls ``hfai.client.go_suspend()`` (ref:restnet_ddp.py:36-47).
``hfai`` is a closed cluster service; here a suspend request is a signal (SIGUSR1 /
SIGTERM) or a sentinel file, and ``go_suspend`` is a clean exit with a distinctive code
so a launcher / scheduler can tell "parked" from "crashed" (SURVEY.md B22, §5.3).
"""
from __future__ import annotations

import os
import signal
import sys
import threading

SUSPEND_EXIT_CODE = 75          # EX_TEMPFAIL: "try again later"
SENTINEL_ENV = "PDB200_SUSPEND_FILE"

_flag = threading.Event()
_installed = False


def _on_signal(signum, frame):  # pragma: no cover - exercised through os.kill in tests
    _flag.set()


def install_signal_handlers(signals=(signal.SIGUSR1, signal.SIGTERM)) -> None:
    """Idempotent; only the main thread of a process may install handlers."""
    global _installed
    if _installed or threading.current_thread() is not threading.main_thread():
        return
    for s in signals:
        try:
            signal.signal(s, _on_signal)
        except (ValueError, OSError):
            pass
    _installed = True


def request_suspend() -> None:
    """Programmatic trigger (tests, embedding applications)."""
    _flag.set()


def clear() -> None:
    _flag.clear()


def receive_suspend_command() -> bool:
    if _flag.is_set():
        return True
    path = os.environ.get(SENTINEL_ENV)
    return bool(path) and os.path.exists(path)


def go_suspend(sleep_forever: bool = False) -> None:
    """Yield the job.  The reference blocks here until the scheduler kills the task;
    off-cluster the useful behaviour is to exit so that the launcher tears the group
    down (quirk Q10) and the same command line can be re-run to resume."""
    path = os.environ.get(SENTINEL_ENV)
    if path and os.path.exists(path):
        try:
            os.remove(path)
        except OSError:
            pass
    sys.stdout.flush()
    sys.stderr.flush()
    raise SystemExit(SUSPEND_EXIT_CODE)
