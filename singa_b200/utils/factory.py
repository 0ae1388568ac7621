"""String-keyed factories and lazy singletons (reference: include/utils/factory.h:11-55,
include/utils/singleton.h:4-39).  Users extend the framework by registering new Layer / Param classes."""
from __future__ import annotations

from typing import Callable, Dict, Generic, TypeVar

T = TypeVar("T")


class Factory(Generic[T]):
    def __init__(self, kind: str = "object"):
        self._kind = kind
        self._creators: Dict[str, Callable[[], T]] = {}

    def register(self, key: str, creator: Callable[[], T]) -> None:
        self._creators[key] = creator

    def registered(self, key: str) -> bool:
        return key in self._creators

    def keys(self):
        return sorted(self._creators)

    def create(self, key: str) -> T:
        if key not in self._creators:
            raise KeyError(f"no {self._kind} registered under id {key!r}; known: {self.keys()}")
        return self._creators[key]()


class Singleton:
    """``Singleton.get(cls)`` builds ``cls()`` once; ``get(cls, arg)`` (the reference's ASingleton) passes a
    constructor argument on first use."""
    _instances: dict = {}

    @classmethod
    def get(cls, klass, *args):
        if klass not in cls._instances:
            cls._instances[klass] = klass(*args)
        return cls._instances[klass]

    @classmethod
    def reset(cls, klass=None):
        if klass is None:
            cls._instances.clear()
        else:
            cls._instances.pop(klass, None)
