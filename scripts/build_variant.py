"""Tuning aid: build a second copy of the library with extra -D flags (lightzero_b200/liblzb200_<name>.so) for A/B runs
through the LZB200_LIB hook.  Usage: python scripts/build_variant.py NAME -DFLAG [-DFLAG ...]"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lightzero_b200 import _lib  # noqa: E402

name, flags = sys.argv[1], sys.argv[2:]
out = os.path.join(ROOT, "lightzero_b200", "liblzb200_%s.so" % name)
subprocess.run(["nvcc"] + _lib.NVCC_FLAGS + flags + [os.path.join(_lib.CSRC, "lzb_api.cu"), "-o", out], check=True)
print(out)
