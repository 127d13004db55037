"""TEST INFRASTRUCTURE ONLY: build recipe for the checker libraries.

* ``oracle/_ref/libref_{mz,ez,gmz,sez,smz}.so`` - the UNMODIFIED reference ctree
  (/root/reference/lzero/mcts/ctree/ctree_*/lib/cnode.cpp + common_lib/cminimax.cpp) compiled from
  where it lies, wrapped by oracle/ref_harness*.cpp and linked with oracle/rng_pin.c.  Only built when
  /root/reference exists (this container); the GPU box uses the prebuilt files that travel with the
  snapshot.  Flags mirror the reference's own build (setup.py:53-68 -> distutils: -O2 -DNDEBUG).
* ``oracle/_build/liboracle.so`` - the plain-C restatement (oracle/mcts_oracle.c).

Nothing under lightzero_b200/ imports this module.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.environ.get("LZB_REFERENCE_ROOT", "/root/reference")
CTREE = os.path.join(REF_ROOT, "lzero", "mcts", "ctree")
REF_OUT = os.path.join(HERE, "_ref")
ORC_OUT = os.path.join(HERE, "_build")

VARIANTS = {
    "mz": ("ctree_muzero", "REF_MZ", "ref_harness.cpp"),
    "ez": ("ctree_efficientzero", "REF_EZ", "ref_harness.cpp"),
    "gmz": ("ctree_gumbel_muzero", "REF_GMZ", "ref_harness.cpp"),
    "sez": ("ctree_sampled_efficientzero", "REF_SEZ", "ref_harness_sampled.cpp"),
    "smz": ("ctree_stochastic_muzero", "REF_SMZ", "ref_harness_stochastic.cpp"),
}


def _newer(target, sources):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(s) <= t for s in sources if os.path.exists(s))


def _run(cmd):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("build failed: %s\n%s\n%s" % (" ".join(cmd), res.stdout, res.stderr))


def reference_available():
    return os.path.isdir(CTREE)


def build_ref(force=False):
    """Compile the reference trees into oracle/_ref/. Returns dict variant -> path (existing files only)."""
    os.makedirs(REF_OUT, exist_ok=True)
    out = {}
    for name, (subdir, macro, harness) in VARIANTS.items():
        so = os.path.join(REF_OUT, "libref_%s.so" % name)
        hsrc = os.path.join(HERE, harness)
        pin = os.path.join(HERE, "rng_pin.c")
        if reference_available() and os.path.exists(hsrc):
            srcs = [hsrc, pin, os.path.join(CTREE, subdir, "lib", "cnode.cpp"),
                    os.path.join(CTREE, subdir, "lib", "cnode.h"), __file__]
            if force or not _newer(so, srcs):
                obj = os.path.join(REF_OUT, "rng_pin.o")
                _run(["gcc", "-O2", "-fPIC", "-c", pin, "-o", obj])
                _run(["g++", "-O2", "-DNDEBUG", "-g0", "-w", "-fPIC", "-shared", "-D" + macro,
                      "-I" + os.path.join(CTREE, subdir), "-I" + os.path.join(CTREE, subdir, "lib"),
                      hsrc, obj, "-Wl,-Bsymbolic", "-o", so])
        if os.path.exists(so):
            out[name] = so
    return out


def build_oracle(force=False):
    """Compile the plain-C restatement into oracle/_build/liboracle.so."""
    os.makedirs(ORC_OUT, exist_ok=True)
    so = os.path.join(ORC_OUT, "liboracle.so")
    srcs = [os.path.join(HERE, f) for f in ("mcts_oracle.c", "glibc_flt32.c", "libstdcxx_rng.c", "oracle.h")]
    srcs = [s for s in srcs if os.path.exists(s)]
    if force or not _newer(so, srcs + [__file__]):
        # -ffp-contract=off: the restatement states every fma() explicitly; nothing else may fuse.
        _run(["gcc", "-O2", "-std=c11", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-Wextra",
              "-Wno-unused-parameter"] + [s for s in srcs if s.endswith(".c")] + ["-lm", "-o", so])
    return so


if __name__ == "__main__":
    force = "--force" in sys.argv
    print(build_ref(force))
    print(build_oracle(force))
