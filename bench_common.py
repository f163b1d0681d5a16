"""Helpers shared by the bench_*.py configurations."""
import numpy as np
import torch


class LossReader:
    """Reads every step's loss on the host one step behind: the D2H copy of step i is queued behind step i and waited for only after
    step i + 1 has been queued, so the device never idles for the read (a training loop that logs its loss does the same)."""

    def __init__(self):
        self.slots = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.events = [torch.cuda.Event(), torch.cuda.Event()]
        self.i, self.read = 0, 0

    def _wait(self, k):
        self.events[k].synchronize()
        self.read += int(np.isfinite(float(self.slots[k][0])))

    def push(self, loss):
        k = self.i % 2
        self.slots[k].copy_(loss.detach().reshape(1), non_blocking=True)
        self.events[k].record()
        if self.i > 0:
            self._wait(1 - k)
        self.i += 1

    def finish(self):
        if self.i > 0:
            self._wait((self.i - 1) % 2)
        assert self.read == self.i, 'a step produced a non-finite loss'
