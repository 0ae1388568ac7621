"""Proto IO and small helpers (reference: src/utils/common.cc:18-87, include/utils/common.h:29-50)."""
from __future__ import annotations

import os
import random
import time

from google.protobuf import text_format


def read_proto_from_text_file(path: str, msg):
    """Parse a text-format protobuf file (``#`` comments allowed) into ``msg`` and return it."""
    with open(path, "r") as f:
        text_format.Parse(f.read(), msg)
    return msg


def write_proto_to_text_file(msg, path: str) -> None:
    with open(path, "w") as f:
        f.write(text_format.MessageToString(msg))


def read_proto_from_binary_file(path: str, msg):
    with open(path, "rb") as f:
        msg.ParseFromString(f.read())
    return msg


def write_proto_to_binary_file(msg, path: str) -> None:
    tmp = f"{path}.tmp.{os.getpid()}"
    with open(tmp, "wb") as f:
        f.write(msg.SerializeToString())
        f.flush()
        os.fsync(f.fileno())
    os.replace(tmp, path)  # atomic: a crash never leaves a torn checkpoint


def string_printf(fmt: str, *args) -> str:
    return fmt % args if args else fmt


def int_vec_to_string(vec) -> str:
    """``"(a, b, c, )"`` — the format the reference uses in shape logs (common.cc:18-26)."""
    return "(" + "".join(f"{int(v)}, " for v in vec) + ")"


def check_exists(path: str) -> bool:
    return os.path.exists(path)


def sleep_ms(ms: float) -> None:
    time.sleep(ms / 1000.0)


def rand_real() -> float:
    return random.random()


def debug_attach_spin() -> None:
    """``Debug()`` of the reference (common.cc:45-53): print the pid and spin until a debugger (or a file
    ``/tmp/singa_continue_<pid>``) releases the process."""
    pid = os.getpid()
    print(f"PID {pid} waiting for debugger; touch /tmp/singa_continue_{pid} to continue", flush=True)
    while not os.path.exists(f"/tmp/singa_continue_{pid}"):
        time.sleep(1.0)
