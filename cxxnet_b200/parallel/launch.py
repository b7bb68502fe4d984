"""One-worker-per-GPU launcher.

``dev = gpu:0-3`` in a conf asks for four devices.  The reference starts one pthread per device
inside a single process (reference ``src/nnet/nnet_impl-inl.hpp:356-359``); here the CLI re-executes
itself as ``len(devs)`` ranks of a ``torch.distributed`` job on this node (equivalent to
``torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1``), one process per GPU.

  python -m cxxnet_b200.parallel.launch <conf> [k=v ...]
"""
from __future__ import annotations

import os
import socket
import subprocess
import sys
from typing import List


def _free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def relaunch_per_device(argv: List[str], devs: List[int]) -> int:
    port = _free_port()
    procs = []
    n = len(devs)
    for r in range(n):
        env = dict(os.environ)
        env.update({"RANK": str(r), "LOCAL_RANK": str(r), "WORLD_SIZE": str(n), "LOCAL_WORLD_SIZE": str(n),
                    "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
        procs.append(subprocess.Popen([sys.executable, "-m", "cxxnet_b200"] + list(argv), env=env))
    return wait_all(procs)


def wait_all(procs, poll_seconds: float = 0.2, grace_seconds: float = 5.0) -> int:
    """Wait for the ranks.  When one of them fails, the others - which would otherwise sit in a collective or in
    the exchange kernel's bounded wait until a timeout - are stopped (the processes this launcher started, by PID):
    the job ends promptly with the failing rank's exit code.  The reference has no failure handling at all
    (errors call exit(-1) in the failing thread's process)."""
    import time
    failed = 0
    pending = list(procs)
    while pending and not failed:
        for p in list(pending):
            rc = p.poll()
            if rc is None:
                continue
            pending.remove(p)
            if rc != 0:
                failed = rc
                break
        if pending and not failed:
            time.sleep(poll_seconds)
    if failed:
        for p in pending:
            p.terminate()
        deadline = time.time() + grace_seconds
        for p in pending:
            try:
                p.wait(timeout=max(0.1, deadline - time.time()))
            except subprocess.TimeoutExpired:
                p.kill()
                p.wait()
    return failed


def main() -> int:
    from ..config import parse_cli_overrides, parse_config_file
    from ..trainer import parse_devices
    argv = sys.argv[1:]
    if not argv:
        print("Usage: python -m cxxnet_b200.parallel.launch <config> [name=value ...]")
        return 0
    dev = "cpu"
    for k, v in parse_config_file(argv[0]) + parse_cli_overrides(argv[1:]):
        if k == "dev":
            dev = v
    kind, devs = parse_devices(dev)
    if len(devs) <= 1:
        from ..cli import main as cli_main
        return cli_main(argv)
    return relaunch_per_device(argv, devs)


if __name__ == "__main__":
    sys.exit(main())
