"""``loader`` — build / split Shard record files (reference: tools/data_loader/data_loader.cc:20-146,
tools/data_loader/data_source.cc:25-196, script/load_data.py).

  python -m singa_b200.tools.loader --datasource mnist --imagefile I --labelfile L --shard_folder OUT
  python -m singa_b200.tools.loader --datasource imagenet --imagefolder DIR --rid_file rid.txt [--meanfile m.binaryproto]
                                    --width 256 --height 256 --shard_folder OUT
  python -m singa_b200.tools.loader --datasource synthetic --shape 28,28 --n 10000 --shard_folder OUT
  python -m singa_b200.tools.loader --input SHARD --prefix OUT_ --n K       split a shard into K shards (K=2 -> train/val)
  python -m singa_b200.tools.loader --partition rid.txt --ngroups G [--replicate R]   record lists per worker group

The loader is restartable: shards are opened in append mode, which drops a torn tail and skips keys already there.
"""
from __future__ import annotations

import argparse
import os
import struct
import sys
from typing import Iterator, List, Tuple

import numpy as np

from ..io.shard import Shard
from ..proto import schema as S


def _record(shape, label: int, pixels: np.ndarray = None, floats: np.ndarray = None) -> "S.Record":
    r = S.Record()
    r.image.shape.extend(int(s) for s in shape)
    r.image.label = int(label)
    if pixels is not None:
        r.image.pixel = np.ascontiguousarray(pixels, dtype=np.uint8).tobytes()
    if floats is not None:
        r.image.data.extend(np.asarray(floats, dtype=np.float32).ravel().tolist())
    return r


class MnistSource:
    """idx3 images + idx1 labels, big-endian headers (data_source.cc:25-77)."""

    def __init__(self, imagefile: str, labelfile: str):
        with open(imagefile, "rb") as f:
            magic, n, rows, cols = struct.unpack(">IIII", f.read(16))
            assert magic == 2051, f"{imagefile}: bad idx3 magic {magic}"
            self.images = np.frombuffer(f.read(n * rows * cols), dtype=np.uint8).reshape(n, rows, cols)
        with open(labelfile, "rb") as f:
            magic, n2 = struct.unpack(">II", f.read(8))
            assert magic == 2049 and n2 == n, f"{labelfile}: bad idx1 header"
            self.labels = np.frombuffer(f.read(n), dtype=np.uint8)

    def __iter__(self) -> Iterator[Tuple[str, "S.Record"]]:
        for i in range(len(self.labels)):
            yield f"{i:08d}", _record(self.images[i].shape, int(self.labels[i]), self.images[i])


class ImageNetSource:
    """``rid.txt`` lines ``<image file> <label>`` under ``imagefolder``; images are resized to width x height and
    the optional Caffe ``mean.binaryproto`` is subtracted, which makes the record float-valued
    (data_source.cc:97-196)."""

    def __init__(self, imagefolder: str, rid_file: str, width: int, height: int, meanfile: str = ""):
        self.folder, self.w, self.h = imagefolder, width, height
        self.lines = [l.split() for l in open(rid_file) if l.strip()]
        self.mean = None
        if meanfile:
            bp = S.BlobProto()
            bp.ParseFromString(open(meanfile, "rb").read())
            self.mean = np.asarray(bp.data, dtype=np.float32).reshape(bp.channels, bp.height, bp.width)

    def _load(self, path: str) -> np.ndarray:
        try:
            import cv2
            img = cv2.imread(path, cv2.IMREAD_COLOR)
            if img is None:
                raise IOError(f"cannot read image {path}")
            img = cv2.resize(img, (self.w, self.h))
            return img.transpose(2, 0, 1)                  # BGR, C,H,W like the reference's OpenCV path
        except ImportError:
            from PIL import Image
            img = Image.open(path).convert("RGB").resize((self.w, self.h))
            return np.asarray(img)[:, :, ::-1].transpose(2, 0, 1)

    def __iter__(self):
        for name, label in self.lines:
            img = self._load(os.path.join(self.folder, name))
            if self.mean is not None:
                yield name, _record(img.shape, int(label), floats=img.astype(np.float32) - self.mean)
            else:
                yield name, _record(img.shape, int(label), img)


class SyntheticRecords:
    def __init__(self, shape: List[int], n: int, classes: int = 10, seed: int = 0):
        from ..layers.data import SyntheticSource
        self.src = SyntheticSource(shape, classes, pool=n, seed=seed)
        self.shape = shape

    def __iter__(self):
        for i in range(self.src.pixels.shape[0]):
            yield f"{i:08d}", _record(self.shape, int(self.src.labels[i]), self.src.pixels[i])


def build(source, shard_folder: str, flush_every: int = 1000) -> int:
    n = 0
    with Shard(shard_folder, Shard.kAppend) as sh:
        for key, rec in source:
            if sh.insert(key, rec):
                n += 1
            if n % flush_every == 0:
                sh.flush()
        sh.flush()
    return n


def build_lmdb(source, folder: str) -> int:
    """Same records as ``build`` but as an LMDB environment of Caffe ``Datum`` values (what the reference's
    kLMDBData layer reads, src/worker/layer.cc:237-328), written by the in-tree writer (no liblmdb needed)."""
    from ..io.lmdb_reader import write_lmdb
    items = []
    for key, rec in source:
        r = rec
        if isinstance(rec, (bytes, bytearray)):
            r = S.Record()
            r.ParseFromString(bytes(rec))
        d = S.Datum()
        shp = list(r.image.shape)
        shp = [1] * (3 - len(shp)) + shp
        d.channels, d.height, d.width = shp[-3:]
        if r.image.pixel:
            d.data = r.image.pixel
        else:
            d.float_data.extend(r.image.data)
        d.label = r.image.label
        items.append((key if isinstance(key, bytes) else str(key).encode(), d.SerializeToString()))
    return write_lmdb(folder, items)


def split(input_folder: str, prefix: str, n: int) -> List[int]:
    """Round-robin split of a shard into ``n`` shards ``<prefix><i>`` (data_loader.cc:43-94)."""
    outs = [Shard(f"{prefix}{i}", Shard.kAppend) for i in range(n)]
    counts = [0] * n
    with Shard(input_folder, Shard.kRead) as src:
        i = 0
        while True:
            kv = src.next()
            if kv is None:
                break
            if outs[i % n].insert(kv[0], kv[1]):
                counts[i % n] += 1
            i += 1
    for o in outs:
        o.flush()
        o.close()
    return counts


def partition_record_list(rid_file: str, ngroups: int, replicate: int = 1) -> List[List[str]]:
    """Assign the records of ``rid_file`` to worker groups, optionally replicating each record on ``replicate``
    consecutive groups (script/load_data.py)."""
    lines = [l.rstrip("\n") for l in open(rid_file) if l.strip()]
    parts: List[List[str]] = [[] for _ in range(ngroups)]
    per = (len(lines) + ngroups - 1) // ngroups
    for g in range(ngroups):
        chunk = lines[g * per:(g + 1) * per]
        for r in range(replicate):
            parts[(g + r) % ngroups].extend(chunk)
    return parts


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(prog="loader")
    ap.add_argument("--datasource", default="")
    ap.add_argument("--imagefile", default="")
    ap.add_argument("--labelfile", default="")
    ap.add_argument("--imagefolder", default="")
    ap.add_argument("--rid_file", default="")
    ap.add_argument("--meanfile", default="")
    ap.add_argument("--width", type=int, default=256)
    ap.add_argument("--height", type=int, default=256)
    ap.add_argument("--shape", default="28,28")
    ap.add_argument("--classes", type=int, default=10)
    ap.add_argument("--shard_folder", default="")
    ap.add_argument("--format", default="shard", choices=["shard", "lmdb"],
                    help="output container: the Shard file (default) or an LMDB environment of Datum records")
    ap.add_argument("--input", default="")
    ap.add_argument("--prefix", default="")
    ap.add_argument("--n", type=int, default=2)
    ap.add_argument("--partition", default="")
    ap.add_argument("--ngroups", type=int, default=1)
    ap.add_argument("--replicate", type=int, default=1)
    a = ap.parse_args(argv)
    if a.partition:
        for g, part in enumerate(partition_record_list(a.partition, a.ngroups, a.replicate)):
            out = f"{a.partition}.group{g}"
            open(out, "w").write("\n".join(part) + "\n")
            print(f"{out}: {len(part)} records")
        return 0
    if a.input:
        print("records per shard:", split(a.input, a.prefix or (a.input.rstrip("/") + "_"), a.n))
        return 0
    if a.datasource == "mnist":
        src = MnistSource(a.imagefile, a.labelfile)
    elif a.datasource == "imagenet":
        src = ImageNetSource(a.imagefolder, a.rid_file, a.width, a.height, a.meanfile)
    elif a.datasource == "synthetic":
        src = SyntheticRecords([int(x) for x in a.shape.split(",")], a.n if a.n > 2 else 1000, a.classes)
    else:
        ap.error("--datasource must be mnist, imagenet or synthetic (or use --input / --partition)")
    n = build_lmdb(src, a.shard_folder) if a.format == "lmdb" else build(src, a.shard_folder)
    print(f"inserted {n} records into {a.shard_folder}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
