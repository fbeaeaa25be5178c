"""Expiring key-value storage with dictionary values (every record has an expiration time, newer
expiration wins; a key may hold a dictionary of sub-keys, each with its own expiration -- this is how
`{prefix}_metrics[pubkey]` and `{prefix}_progress[pubkey]` work)."""
from __future__ import annotations

import heapq
from dataclasses import dataclass
from typing import Any, Dict, Iterator, List, Optional, Tuple

from .routing import get_dht_time


@dataclass(frozen=True)
class ValueWithExpiration:
    value: Any
    expiration_time: float

    def __iter__(self):
        return iter((self.value, self.expiration_time))


class DictionaryDHTValue:
    """subkey -> ValueWithExpiration, with its own latest-expiration bookkeeping."""

    def __init__(self):
        self.data: Dict[Any, ValueWithExpiration] = {}

    @property
    def latest_expiration_time(self) -> float:
        return max((v.expiration_time for v in self.data.values()), default=-float("inf"))

    def store(self, subkey, value, expiration_time: float) -> bool:
        prev = self.data.get(subkey)
        if prev is not None and prev.expiration_time >= expiration_time:
            return False
        self.data[subkey] = ValueWithExpiration(value, expiration_time)
        return True

    def prune(self, now: float) -> None:
        for k in [k for k, v in self.data.items() if v.expiration_time <= now]:
            del self.data[k]

    def items(self):
        return self.data.items()

    def __len__(self):
        return len(self.data)


class DHTLocalStorage:
    def __init__(self, maxsize: Optional[int] = None):
        self.maxsize = maxsize
        self.data: Dict[bytes, ValueWithExpiration] = {}
        self._heap: List[Tuple[float, bytes]] = []

    def _prune(self) -> None:
        now = get_dht_time()
        while self._heap and (self._heap[0][0] <= now or (self.maxsize and len(self.data) > self.maxsize)):
            exp, key = heapq.heappop(self._heap)
            cur = self.data.get(key)
            if cur is None:
                continue
            if isinstance(cur.value, DictionaryDHTValue):
                cur.value.prune(now)
                if len(cur.value):
                    latest = cur.value.latest_expiration_time
                    self.data[key] = ValueWithExpiration(cur.value, latest)
                    heapq.heappush(self._heap, (min(v.expiration_time for v in cur.value.data.values()), key))
                    continue
                del self.data[key]
            elif cur.expiration_time <= now or exp == cur.expiration_time:
                del self.data[key]

    def store(self, key: bytes, value: Any, expiration_time: float, subkey: Any = None) -> bool:
        if expiration_time <= get_dht_time():
            return False
        self._prune()
        cur = self.data.get(key)
        if subkey is not None:
            if cur is None or not isinstance(cur.value, DictionaryDHTValue):
                if cur is not None and cur.expiration_time >= expiration_time:
                    return False  # a plain value that is at least as fresh shadows this dictionary entry (ties lose)
                d = DictionaryDHTValue()
            else:
                d = cur.value
            ok = d.store(subkey, value, expiration_time)
            if ok:
                self.data[key] = ValueWithExpiration(d, d.latest_expiration_time)
                heapq.heappush(self._heap, (expiration_time, key))
            return ok
        if cur is not None and cur.expiration_time >= expiration_time:
            return False
        self.data[key] = ValueWithExpiration(value, expiration_time)
        heapq.heappush(self._heap, (expiration_time, key))
        return True

    def get(self, key: bytes) -> Optional[ValueWithExpiration]:
        self._prune()
        cur = self.data.get(key)
        if cur is None:
            return None
        now = get_dht_time()
        if isinstance(cur.value, DictionaryDHTValue):
            cur.value.prune(now)
            if not len(cur.value):
                del self.data[key]
                return None
            return ValueWithExpiration(cur.value, cur.value.latest_expiration_time)
        if cur.expiration_time <= now:
            del self.data[key]
            return None
        return cur

    def items(self) -> Iterator[Tuple[bytes, ValueWithExpiration]]:
        self._prune()
        return iter(list(self.data.items()))

    def __len__(self):
        self._prune()
        return len(self.data)
