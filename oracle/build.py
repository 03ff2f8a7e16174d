"""Build recipe for the oracle (test infrastructure).

``build_port()``  compiles swscale_port.c with gcc into oracle/_build/libtvoracle.so.
``fetch_ref()``   copies the genuine libswscale/libavutil binaries (and the two bundled libraries
                  libavutil needs) out of the opencv-python-headless wheel in this image into
                  oracle/_ref/, so the real reference travels to the GPU box with the snapshot even
                  if that box's site-packages differ.  Binaries only — libswscale's sources are not
                  in this container and nothing is copied from /root/reference.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(HERE, "_build", "libtvoracle.so")
REF_DIR = os.path.join(HERE, "_ref")
_WANTED = ("libdrm-", "libcrypto-", "libavutil-", "libswscale-")


def build_port(force: bool = False) -> str:
    src = os.path.join(HERE, "swscale_port.c")
    if not force and os.path.exists(PORT_SO) and os.path.getmtime(PORT_SO) >= os.path.getmtime(src):
        return PORT_SO
    os.makedirs(os.path.dirname(PORT_SO), exist_ok=True)
    tmp = PORT_SO + ".tmp.%d" % os.getpid()
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", tmp, src, "-lm"])
    os.replace(tmp, PORT_SO)
    return PORT_SO


def wheel_libs_dir() -> str | None:
    for base in {sysconfig.get_paths()["purelib"], sysconfig.get_paths()["platlib"]}:
        d = os.path.join(base, "opencv_python_headless.libs")
        if os.path.isdir(d):
            return d
    return None


def fetch_ref() -> str | None:
    src = wheel_libs_dir()
    if src is None:
        return REF_DIR if glob.glob(os.path.join(REF_DIR, "libswscale-*")) else None
    os.makedirs(REF_DIR, exist_ok=True)
    for prefix in _WANTED:
        for path in glob.glob(os.path.join(src, prefix + "*.so*")):
            dst = os.path.join(REF_DIR, os.path.basename(path))
            if not os.path.exists(dst) or os.path.getsize(dst) != os.path.getsize(path):
                shutil.copy2(path, dst)
    return REF_DIR


if __name__ == "__main__":
    print(build_port(force=True))
    print(fetch_ref())
