"""subprocess.run with a hard timeout that kills the WHOLE process group (torchrun workers included), so a hung
multi-GPU test can never leave orphans spinning on the GPUs."""
import os
import signal
import subprocess


class Result:
    def __init__(self, returncode, stdout, stderr):
        self.returncode, self.stdout, self.stderr = returncode, stdout, stderr


def run_group(cmd, timeout, env=None, cwd=None) -> Result:
    p = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, env=env, cwd=cwd,
                         start_new_session=True)
    try:
        out, err = p.communicate(timeout=timeout)
        return Result(p.returncode, out, err)
    except subprocess.TimeoutExpired:
        try:
            os.killpg(p.pid, signal.SIGKILL)
        except ProcessLookupError:
            pass
        out, err = p.communicate()
        return Result(-9, out or "", (err or "") + f"\n[proc_util] killed process group after {timeout}s timeout")
