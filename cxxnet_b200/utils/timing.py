"""Per-layer timers (`layer_timing = 1`).  The reference has no profiling facility at all (SURVEY.md
section 5.1: wall-clock progress lines only); this one brackets every layer's forward and backward with
CUDA events on the launching stream (perf_counter on the CPU backend) and reports mean microseconds
per step.  It serialises nothing: events are resolved once, when the report is asked for.  Captured
steps (`cuda_graph = 1`) cannot be bracketed, so the trainer runs eagerly while timing is on."""
from __future__ import annotations

import time
from typing import List

import torch


class LayerTimer:
    def __init__(self, names, cuda: bool):
        """names: list of layer names, or a callable returning it (resolved when a report is made)"""
        self._names = names
        self.cuda = cuda
        self.total = {"f": [], "b": []}
        self.count = {"f": [], "b": []}
        self._pending = []

    @property
    def names(self) -> List[str]:
        n = self._names() if callable(self._names) else self._names
        width = max(len(n), len(self.total["f"]), len(self.total["b"]))
        return list(n) + ["layer %d" % i for i in range(len(n), width)]

    def _add(self, kind: str, index: int, us: float) -> None:
        t, c = self.total[kind], self.count[kind]
        while len(t) <= index:
            t.append(0.0)
            c.append(0)
        t[index] += us
        c[index] += 1

    # -- bracket one layer call
    def begin(self):
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            return e
        return time.perf_counter()

    def end(self, kind: str, index: int, start) -> None:
        if self.cuda:
            e = torch.cuda.Event(enable_timing=True)
            e.record()
            self._pending.append((kind, index, start, e))
            if len(self._pending) > 4096:
                self.flush()
        else:
            self._add(kind, index, (time.perf_counter() - start) * 1e6)

    def flush(self) -> None:
        if self._pending:
            torch.cuda.synchronize()
            for kind, index, e0, e1 in self._pending:
                self._add(kind, index, e0.elapsed_time(e1) * 1e3)
            self._pending = []

    def reset(self) -> None:
        self.flush()
        for k in ("f", "b"):
            self.total[k], self.count[k] = [], []

    def report(self) -> str:
        """table of mean microseconds per call, forward and backward, with shares of the total"""
        self.flush()
        names = self.names
        mean = {}
        for k in ("f", "b"):
            m = [t / c if c else 0.0 for t, c in zip(self.total[k], self.count[k])]
            mean[k] = m + [0.0] * (len(names) - len(m))
        whole = sum(mean["f"]) + sum(mean["b"])
        lines = ["layer timing (mean us per step; stream time between launches of consecutive layers is attributed "
                 "to the earlier one)", "%4s %-28s %10s %10s %7s" % ("#", "layer", "forward", "backward", "share")]
        for i, n in enumerate(names):
            t = mean["f"][i] + mean["b"][i]
            lines.append("%4d %-28s %10.1f %10.1f %6.1f%%" % (i, n[:28], mean["f"][i], mean["b"][i],
                                                                100.0 * t / whole if whole else 0.0))
        lines.append("%4s %-28s %10.1f %10.1f" % ("", "total", sum(mean["f"]), sum(mean["b"])))
        return "\n".join(lines)
