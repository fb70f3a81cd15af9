"""Side channel for closure objects (comparers, range keys, seeds) that a plan refers to by index —
reference: LinqToDryad/DryadLinqObjectStore.cs:43-173 (client `Put(obj) -> index`, serialised into
a resource file shipped with the job, vertex-side `Get(index)`)."""
from __future__ import annotations

import pickle
from typing import Any, List


class ObjectStore:
    def __init__(self):
        self._objects: List[Any] = []
        self._sealed = False

    def put(self, obj: Any) -> int:
        if self._sealed:
            raise RuntimeError("the object store was already shipped with a job")
        for i, o in enumerate(self._objects):
            if o is obj:
                return i
        self._objects.append(obj)
        return len(self._objects) - 1

    def get(self, index: int) -> Any:
        return self._objects[index]

    def __len__(self):
        return len(self._objects)

    def save(self, path: str) -> None:
        """Serialise to the job's resource file.  Objects that cannot be pickled (lambdas) are the
        reason single-process execution keeps closures in memory instead."""
        with open(path, "wb") as f:
            pickle.dump(self._objects, f)
        self._sealed = True

    @staticmethod
    def load(path: str) -> "ObjectStore":
        s = ObjectStore()
        with open(path, "rb") as f:
            s._objects = pickle.load(f)
        s._sealed = True
        return s
