#!/usr/bin/env python
"""Start N copies of a command with RANK / LOCAL_RANK / WORLD_SIZE set and nothing else -- the environment an `mpirun` / `srun`
launch gives the reference's back-end tests, which pick their own MASTER_ADDR / MASTER_PORT (`tests/test_mpi_backend.py:10-11`
there).  (`torchrun` does not fit those tests: its workers are clients of the agent's store at the port torchrun chose.)

    python tools/run_ranks.py -n 2 -- python -m pytest /root/reference/tests/test_mpi_backend.py -q

Exit code = the largest exit code of the ranks; a rank that runs longer than --timeout seconds is killed (exit 124)."""
import argparse
import os
import subprocess
import sys
import tempfile
import time


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("-n", type=int, default=2)
    ap.add_argument("--timeout", type=float, default=600.0)
    ap.add_argument("cmd", nargs=argparse.REMAINDER)
    a = ap.parse_args()
    cmd = a.cmd[1:] if a.cmd and a.cmd[0] == "--" else a.cmd
    assert cmd, "no command"
    procs, logs = [], []
    for r in range(a.n):
        env = dict(os.environ, RANK=str(r), LOCAL_RANK=str(r), WORLD_SIZE=str(a.n), OMP_NUM_THREADS=os.environ.get("OMP_NUM_THREADS", "1"))
        # rank 0 reports live; the others would interleave the same lines, so theirs are kept and shown only when they fail
        log = None if r == 0 else tempfile.TemporaryFile(mode="w+")
        logs.append(log)
        procs.append(subprocess.Popen(cmd, env=env, stdout=log, stderr=None if r == 0 else subprocess.STDOUT))
    t0, codes = time.time(), [None] * a.n
    while any(c is None for c in codes):
        for i, p in enumerate(procs):
            if codes[i] is None:
                codes[i] = p.poll()
        if time.time() - t0 > a.timeout:
            for i, p in enumerate(procs):
                if codes[i] is None:
                    p.kill()
                    codes[i] = 124
        time.sleep(0.05)
    for r, (code, log) in enumerate(zip(codes, logs)):
        if log is not None and code != 0:
            log.seek(0)
            print(f"---- rank {r} (exit {code}), last lines ----\n" + "".join(log.readlines()[-40:]), file=sys.stderr)
    print(f"rank exit codes: {codes}")
    return max(codes)


if __name__ == "__main__":
    sys.exit(main())
