#!/usr/bin/env python
"""Place the UNMODIFIED reference node program under oracle/_ref/ -- TEST INFRASTRUCTURE ONLY.

The reference is pure Python: there is nothing to compile, and /root/reference does
not exist on the GPU box.  So that ``bench.py --impl reference`` and the
``cpu_baseline`` leg can time the real thing there (cpu_baseline.kind = "reference")
instead of the port, this recipe copies the four files the two-node CIFAR pipeline
consists of -- byte for byte, from where they lie -- into ``oracle/_ref/``:

    node.py  cifar_model_parts.py  node_service_pb2.py  node_service_pb2_grpc.py

``oracle/_ref/`` is git-ignored (the reference's sources never enter this
repository's history) but not gpurun-ignored, so it travels to the GPU box with the
snapshot exactly like the built .so.  ``__graft_entry__.build()`` runs this whenever
/root/reference is present; without it everything falls back to the port.

    python oracle/build_ref.py
"""
from __future__ import annotations

import hashlib
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("B2P_REFERENCE", "/root/reference")
OUT = os.path.join(HERE, "_ref")
FILES = ("node.py", "cifar_model_parts.py", "node_service_pb2.py", "node_service_pb2_grpc.py")


def available() -> bool:
    return all(os.path.exists(os.path.join(OUT, f)) for f in FILES)


def build() -> bool:
    """Returns True when oracle/_ref/ is complete afterwards."""
    if not all(os.path.exists(os.path.join(REF_SRC, f)) for f in FILES):
        return available()
    os.makedirs(OUT, exist_ok=True)
    lines = []
    for f in FILES:
        src, dst = os.path.join(REF_SRC, f), os.path.join(OUT, f)
        shutil.copyfile(src, dst)
        os.chmod(dst, 0o644)
        lines.append(f"{hashlib.sha256(open(dst, 'rb').read()).hexdigest()}  {f}")
    with open(os.path.join(OUT, "SHA256SUMS"), "w") as fh:
        fh.write("\n".join(lines) + "\n")
    return True


if __name__ == "__main__":
    ok = build()
    print("oracle/_ref:", "ready" if ok else "unavailable (no /root/reference here)")
    sys.exit(0 if ok else 1)
