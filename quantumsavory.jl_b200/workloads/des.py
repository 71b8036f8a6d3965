"""A minimal discrete-event scheduler for the workload drivers (NOT part of the backend).

The reference runs its protocols as ResumableFunctions on ConcurrentSim; that control plane stays in
Julia and is out of scope.  The benchmark configs still need *a* host that
produces the op trace of examples/firstgenrepeater_lowlevel, so this file provides the three
primitives those protocols use -- timeout, request/unlock of a slot lock, and `a & b` -- with
ConcurrentSim's ordering rule (events fire in (time, insertion order)).
"""
from __future__ import annotations

import heapq
import itertools


class Event:
    __slots__ = ("sim", "triggered", "callbacks")

    def __init__(self, sim):
        self.sim, self.triggered, self.callbacks = sim, False, []

    def succeed(self):
        if not self.triggered:
            self.triggered = True
            for cb in self.callbacks:
                self.sim._schedule(0.0, cb)
            self.callbacks = []

    def wait(self, cb):
        if self.triggered:
            self.sim._schedule(0.0, cb)
        else:
            self.callbacks.append(cb)

    def __and__(self, other):
        return AllOf(self.sim, [self, other])


class AllOf(Event):
    def __init__(self, sim, events):
        super().__init__(sim)
        self.pending = 0
        evs = []
        for e in events:
            evs.extend(e.events if isinstance(e, AllOf) and not e.triggered else [e])
        self.events = evs
        for e in evs:
            if not e.triggered:
                self.pending += 1
                e.callbacks.append(self._one)
        if self.pending == 0:
            self.succeed()

    def _one(self):
        self.pending -= 1
        if self.pending == 0:
            self.succeed()


class Resource:
    """Capacity-1 lock with a FIFO queue (ConcurrentSim.Resource as used for register slots)."""

    def __init__(self, sim):
        self.sim, self.locked, self.queue = sim, False, []

    def request(self) -> Event:
        ev = Event(self.sim)
        if not self.locked:
            self.locked = True
            ev.succeed()
        else:
            self.queue.append(ev)
        return ev

    def release(self):
        if self.queue:
            self.queue.pop(0).succeed()
        else:
            self.locked = False

    def islocked(self):
        return self.locked


class Simulation:
    def __init__(self):
        self.now = 0.0
        self._heap = []
        self._seq = itertools.count()

    def _schedule(self, delay, cb):
        heapq.heappush(self._heap, (self.now + delay, next(self._seq), cb))

    def timeout(self, delay) -> Event:
        ev = Event(self)
        self._schedule(delay, ev.succeed)
        return ev

    def process(self, gen):
        def step():
            try:
                ev = next(gen)
            except StopIteration:
                return
            ev.wait(step)
        self._schedule(0.0, step)

    def run(self, until):
        """Process every event with time < until, then set now = until (ConcurrentSim.run(sim, t))."""
        while self._heap and self._heap[0][0] < until:
            t, _, cb = heapq.heappop(self._heap)
            self.now = t
            cb()
        self.now = until
