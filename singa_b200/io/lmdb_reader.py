"""LMDB input (``kLMDBData``, src/worker/layer.cc:237-328).  liblmdb / the ``lmdb`` Python module are not
installed on the box, so an LMDB path can only be served when that module exists; otherwise the layer explains
the two supported alternatives instead of failing obscurely."""
from __future__ import annotations

import numpy as np

from ..proto import schema as S
from ..layers.data import BatchSource


class LMDBSource(BatchSource):
    def __init__(self, path: str):
        try:
            import lmdb  # type: ignore
        except ImportError as e:
            raise RuntimeError(
                f"kLMDBData layer needs the `lmdb` module to read {path!r}, which is not installed here. "
                "Convert the data to a Shard with `python -m singa_b200.tools.loader`, set "
                "`data_param { source: \"synthetic\" synthetic_shape: ... }`, or export SINGA_B200_SYNTHETIC=1 "
                "to substitute synthetic records for missing datasets.") from e
        self.env = lmdb.open(path, readonly=True, lock=False)
        self.txn = self.env.begin()
        self.cursor = self.txn.cursor()
        if not self.cursor.first():
            raise RuntimeError(f"LMDB {path} is empty")
        d = S.Datum()
        d.ParseFromString(self.cursor.value())
        self.sample_shape = [d.channels, d.height, d.width] if d.channels > 1 else [d.height, d.width]

    def next_batch(self, pixels, labels):
        for i in range(pixels.shape[0]):
            d = S.Datum()
            d.ParseFromString(self.cursor.value())
            pixels[i, :] = np.frombuffer(d.data, dtype=np.uint8)[:pixels.shape[1]]
            labels[i] = d.label
            if not self.cursor.next():
                self.cursor.first()

    def skip(self, n):
        for _ in range(n):
            if not self.cursor.next():
                self.cursor.first()

    def count(self):
        return self.env.stat()["entries"]
