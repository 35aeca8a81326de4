"""Shell helpers (reference: ``core.run_cmd`` = ``shell_get_command_output`` with timeout and retry
interval, framework/io/shell.cc; used by user scripts to drive hadoop / afs commands)."""
from __future__ import annotations

import subprocess
import time


def run_cmd(cmd: str, time_out: int = -1, sleep_inter: int = -1) -> str:
    """Run ``cmd`` through ``sh -c`` and return its stdout.  ``time_out`` (ms, <= 0: none): keep retrying a
    failing command until that much time has passed, sleeping ``sleep_inter`` ms between attempts
    (<= 0: no retry) — the reference retries pipe failures the same way (shell.cc
    ``shell_get_command_output``)."""
    deadline = time.time() + time_out / 1000.0 if time_out and time_out > 0 else None
    while True:
        r = subprocess.run(["sh", "-c", cmd], capture_output=True, text=True)
        if r.returncode == 0 or deadline is None or sleep_inter is None or sleep_inter <= 0 or time.time() >= deadline:
            return r.stdout
        time.sleep(sleep_inter / 1000.0)
