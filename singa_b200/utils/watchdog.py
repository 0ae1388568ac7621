"""Failure detection for a one-box job (the reference has none — ``WaitUpdate`` spins forever when a
peer dies, src/utils/param_manager.cc:232-233, and the Mesos scheduler ignores lost slaves).

Two small host-side threads:

* ``Watchdog``  — local stall detector: if the training step counter does not advance for ``timeout_s`` the process
  is terminated with exit code 86 and a diagnostic line (instead of hanging in a collective or spinning on a device
  flag until the kernels' own 30-60 s bounded waits trap).
* ``Heartbeat`` — peer liveness over the job's rendezvous store (the TCPStore ``torch.distributed`` already runs, so
  no extra port): every rank publishes ``hb/<rank> = <beat>:<step>`` every ``interval_s`` and watches the others; a
  peer whose beat has not changed for ``dead_after_s`` is declared dead and the survivor leaves with exit code 87
  (synchronous mode cannot make progress without it).  ``on_dead`` can be replaced, e.g. by the asynchronous
  exchange, which keeps training with the remaining groups.

Both are opt-in from the library (``Worker.run`` starts them when ``SINGA_B200_WATCHDOG`` / the ``--watchdog_s`` flag
is set) and cost one sleeping thread each.
"""
from __future__ import annotations

import logging
import os
import sys
import threading
import time
from typing import Callable, Dict, Optional

log = logging.getLogger("singa")

EXIT_STALL = 86
EXIT_PEER_DEAD = 87


def _die(code: int, msg: str) -> None:
    log.error(msg)
    sys.stderr.write(msg + "\n")
    sys.stderr.flush()
    os._exit(code)


class Watchdog:
    def __init__(self, progress: Callable[[], int], timeout_s: float,
                 on_stall: Optional[Callable[[int, float], None]] = None, poll_s: float = 0.0):
        self.progress, self.timeout_s = progress, float(timeout_s)
        self.on_stall = on_stall or (lambda step, idle: _die(
            EXIT_STALL, f"watchdog: no training progress for {idle:.0f} s at step {step}; terminating "
                        f"(a peer may have died or a device wait is stuck)"))
        self.poll_s = poll_s or max(0.05, min(1.0, self.timeout_s / 10))
        self._stop = threading.Event()
        self._paused = False
        self._t: Optional[threading.Thread] = None
        self.fired = False

    def _loop(self) -> None:
        last, since = self.progress(), time.monotonic()
        while not self._stop.wait(self.poll_s):
            cur = self.progress()
            now = time.monotonic()
            if cur != last or self._paused:
                last, since = cur, now
                continue
            if now - since > self.timeout_s:
                self.fired = True
                self.on_stall(cur, now - since)
                last, since = cur, now      # a custom handler may return: re-arm

    def start(self) -> "Watchdog":
        self._t = threading.Thread(target=self._loop, name="singa-watchdog", daemon=True)
        self._t.start()
        return self

    def pause(self, flag: bool = True) -> None:
        """Suspend detection around phases with no step progress (checkpoint gather, test passes)."""
        self._paused = flag

    def stop(self) -> None:
        self._stop.set()
        if self._t is not None:
            self._t.join(timeout=2)


class Heartbeat:
    def __init__(self, store, rank: int, world: int, step: Callable[[], int], interval_s: float = 1.0,
                 dead_after_s: float = 10.0, on_dead: Optional[Callable[[int, float], None]] = None):
        self.store, self.rank, self.world, self.step = store, rank, world, step
        self.interval_s, self.dead_after_s = float(interval_s), float(dead_after_s)
        self.on_dead = on_dead or (lambda peer, idle: _die(
            EXIT_PEER_DEAD, f"heartbeat: rank {peer} has been silent for {idle:.0f} s; rank {self.rank} is leaving"))
        self._stop = threading.Event()
        self._t: Optional[threading.Thread] = None
        self.dead: Dict[int, float] = {}
        self._beat = 0

    def _publish(self) -> None:
        self._beat += 1
        self.store.set(f"hb/{self.rank}", f"{self._beat}:{self.step()}")

    def _loop(self) -> None:
        seen: Dict[int, str] = {}
        since: Dict[int, float] = {}
        while not self._stop.wait(self.interval_s):
            try:
                self._publish()
                now = time.monotonic()
                for r in range(self.world):
                    if r == self.rank or r in self.dead:
                        continue
                    key = f"hb/{r}"
                    # check() avoids blocking in get() on a key a peer never wrote
                    val = self.store.get(key).decode() if self.store.check([key]) else ""
                    if val == "bye":
                        self.dead[r] = 0.0            # clean exit: not a failure
                        continue
                    if seen.get(r) != val:
                        seen[r], since[r] = val, now
                    elif now - since.get(r, now) > self.dead_after_s:
                        self.dead[r] = now - since[r]
                        self.on_dead(r, now - since[r])
            except Exception as e:  # noqa: BLE001 - the store lives in rank 0: losing it means rank 0 is gone
                if self._stop.is_set():
                    return
                self.dead[0] = self.dead_after_s
                self.on_dead(0, self.dead_after_s)
                log.error("heartbeat: store unreachable (%s)", e)
                return

    def start(self) -> "Heartbeat":
        self._publish()
        self._t = threading.Thread(target=self._loop, name="singa-heartbeat", daemon=True)
        self._t.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        try:
            self.store.set(f"hb/{self.rank}", "bye")
        except Exception:  # noqa: BLE001
            pass
        if self._t is not None:
            self._t.join(timeout=2)


def start_for(executor, timeout_s: float):
    """Watchdog (+ heartbeat when a process group exists) for ``executor``; returns the started objects."""
    out = []
    if timeout_s <= 0:
        return out
    progress = lambda: executor.progress_counter  # noqa: E731
    out.append(Watchdog(progress, timeout_s).start())
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            from torch.distributed.distributed_c10d import _get_default_store
            out.append(Heartbeat(_get_default_store(), dist.get_rank(), dist.get_world_size(), progress,
                                 interval_s=max(0.2, min(2.0, timeout_s / 10)), dead_after_s=timeout_s).start())
    except Exception as e:  # noqa: BLE001
        log.error("heartbeat not started: %s", e)
    return out
