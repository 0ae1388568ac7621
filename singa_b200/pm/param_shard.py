"""ParamShard: the server-side parameter store of Stack B (reference: include/utils/param_shard.h:26-47,
src/utils/param_shard.cc:12-64).  ``sync_now(id)`` becomes true every ``sync_interval`` updates of a parameter
(and resets its counter) — the trigger for the server-group neighbour exchange."""
from __future__ import annotations

import threading
from typing import Dict, Optional

import numpy as np

from ..param import Param, param_factory


class ParamShard:
    def __init__(self, shard_id: int, sync_interval: int, sync_rule: str = "replace"):
        self.id = shard_id
        self.sync_interval = max(1, int(sync_interval))
        self.sync_rule = sync_rule           # "replace" (reference placeholder) or "average"
        self.params: Dict[int, Param] = {}
        self.update_counters: Dict[int, int] = {}
        self._lock = threading.Lock()

    def is_local(self, param_id: int) -> bool:
        return param_id in self.params

    def put(self, param_id: int, payload: bytes) -> bool:
        with self._lock:
            p = self.params.get(param_id)
            if p is None:
                p = param_factory.create("Plain")
                p.id = param_id
                self.params[param_id] = p
                self.update_counters[param_id] = 0
            p.handle_put_msg(payload)
        return True

    def get(self, param_id: int) -> Optional[bytes]:
        with self._lock:
            p = self.params.get(param_id)
            return None if p is None else p.handle_get_msg()

    def update(self, param_id: int, payload: bytes) -> Optional[bytes]:
        """Apply a worker's update (placeholder semantics of the reference: replace and echo,
        src/utils/param.cc:57-61) and count it."""
        with self._lock:
            p = self.params.get(param_id)
            if p is None:
                return None
            out = p.handle_update_msg(payload)
            self.update_counters[param_id] += 1
            return out

    def sync_update(self, param_id: int, payload: bytes) -> Optional[bytes]:
        """A neighbour group's server sends its fresh value; fold it in and answer with ours."""
        with self._lock:
            p = self.params.get(param_id)
            if p is None:
                return None
            if self.sync_rule == "average":
                mine = np.frombuffer(p.to_bytes(), dtype=np.float32)
                theirs = np.frombuffer(payload, dtype=np.float32)
                p.from_bytes(((mine + theirs) * 0.5).astype(np.float32).tobytes())
            else:
                p.from_bytes(payload)
            return p.to_bytes()

    def sync_now(self, param_id: int) -> bool:
        with self._lock:
            if self.update_counters.get(param_id, 0) >= self.sync_interval:
                self.update_counters[param_id] = 0
                return True
            return False
