"""Loader (and in-tree build recipe) for liblzb200.so, the C-ABI CUDA library of include/lzb200.h.

There is no CPU fallback: if the library cannot be loaded the import of any tree module fails loudly.
"""
import ctypes as C
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "liblzb200.so")
ABI_VERSION = 6  # LZB_ABI_VERSION of include/lzb200.h
SOURCES = ["lzb_api.cu", "lzb_scaling.cuh", "lzb_select.cuh", "lzb_expand.cuh", "lzb_step.cuh", "lzb_group.cuh", "lzb_sample.cuh", "lzb_collect.cuh", "lzb_peer.cuh", "lzb_pool.cuh", "lzb_math.cuh"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",  # B200 only
    "-fmad=false",  # the reference's x86-64 -O2 build never fuses float mul+add; fp64 fma() calls stay explicit
    "-lineinfo", "-O3", "-std=c++17", "--shared", "-Xcompiler", "-fPIC",
]


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + [os.path.join(ROOT, "include", "lzb200.h")]
    return any(os.path.exists(d) and os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile lightzero_b200/csrc/lzb_api.cu for sm_100a into lightzero_b200/liblzb200.so.

    One process per GPU may get here at the same time (torchrun, bench.py --gpus 8) with a stale or missing library:
    the build runs under an exclusive file lock, compiles to a temporary file in the same directory and renames it
    into place, so no rank ever maps a half-written library and only the first rank compiles."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        if os.path.exists(LIB_PATH):
            import warnings
            warnings.warn("lightzero_b200: sources are newer than %s but nvcc is not available; loading the existing "
                          "library (the ABI version is still checked)" % LIB_PATH)
            return LIB_PATH
        raise RuntimeError("nvcc not found and %s is not built" % LIB_PATH)
    import fcntl
    with open(LIB_PATH + ".lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not force and not _stale():  # another rank built it while this one waited for the lock
                return LIB_PATH
            tmp = "%s.tmp.%d" % (LIB_PATH, os.getpid())
            cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [os.path.join(CSRC, "lzb_api.cu"), "-o", tmp]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                if os.path.exists(tmp):
                    os.unlink(tmp)
                raise RuntimeError("nvcc failed:\n%s\n%s" % (res.stdout, res.stderr))
            os.replace(tmp, LIB_PATH)
            if verbose:
                print(res.stderr)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    pylist()
    return LIB_PATH


PYLIST_PATH = os.path.join(HERE, "_lzb_pylist.so")
_pylist = False  # False: not tried yet; None: unavailable


def pylist():
    """lzb_pylist.c behind ctypes.PyDLL: list-of-lists <-> flat buffers at C speed (host marshalling of the reference's
    list arguments, no arithmetic).  None when it cannot be built or loaded; the callers then walk the lists with numpy."""
    global _pylist
    if _pylist is False:
        _pylist = None
        try:
            src = os.path.join(CSRC, "lzb_pylist.c")
            if not os.path.exists(PYLIST_PATH) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(PYLIST_PATH)):
                import sysconfig
                cc = shutil.which("gcc") or shutil.which("cc")
                if cc:
                    subprocess.run([cc, "-O2", "-shared", "-fPIC", "-pthread", "-I" + sysconfig.get_paths()["include"], src, "-o", PYLIST_PATH],
                                   check=True, capture_output=True)
            P = C.PyDLL(PYLIST_PATH)
            P.lzb_rows_to_f32.restype = C.c_int
            P.lzb_rows_to_f32.argtypes = [C.py_object, C.c_void_p, C.c_long, C.c_long, C.c_void_p]
            P.lzb_seq_to_f32.restype = C.c_int
            P.lzb_seq_to_f32.argtypes = [C.py_object, C.c_void_p, C.c_long]
            P.lzb_rows_to_i32.restype = C.c_int
            P.lzb_rows_to_i32.argtypes = [C.py_object, C.c_void_p, C.c_long, C.c_long, C.c_void_p]
            P.lzb_seq_to_i32.restype = C.c_int
            P.lzb_seq_to_i32.argtypes = [C.py_object, C.c_void_p, C.c_long]
            P.lzb_rows_from_i32.restype = C.py_object
            P.lzb_rows_from_i32.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_long]
            _pylist = P
        except Exception:
            _pylist = None
    return _pylist


class SearchCfg(C.Structure):
    """lzb_search_cfg_t"""
    _fields_ = [
        ("pb_c_base", C.c_int32), ("pb_c_init", C.c_float), ("discount", C.c_float), ("players", C.c_int32),
        ("num_simulations", C.c_int32), ("max_num_considered_actions", C.c_int32), ("rng_mode", C.c_int32),
        ("rng_seed", C.c_uint64), ("rng_seed_dev", C.c_void_p),
    ]


class Gather(C.Structure):
    """lzb_gather_t"""
    _fields_ = [("src_dev", C.c_void_p), ("dst_dev", C.c_void_p), ("row_bytes", C.c_int64), ("zero_if_reset", C.c_int32)]


class Slots(C.Structure):
    """lzb_slots_t"""
    _fields_ = [("ptr", C.c_void_p * 3), ("n", C.c_int32)]


class Reuse(C.Structure):
    """lzb_reuse_t"""
    _fields_ = [("true_action_dev", C.c_void_p), ("reuse_value_dev", C.c_void_p), ("infer_count_dev", C.c_void_p),
                ("flags_dev", C.c_void_p), ("batch_index_dev", C.c_void_p)]


class Chance(C.Structure):
    """lzb_chance_t"""
    _fields_ = [("forced_outcome_dev", C.c_void_p), ("forced_stride", C.c_int32), ("leaf_is_chance_dev", C.c_void_p),
                ("seed_dev", C.c_void_p)]


class SelectAction(C.Structure):
    """lzb_select_action_t"""
    _fields_ = [("visits_dev", C.c_void_p), ("legal_count_dev", C.c_void_p), ("legal_actions_dev", C.c_void_p),
                ("uniforms_dev", C.c_void_p), ("action_pos_dev", C.c_void_p), ("action_dev", C.c_void_p),
                ("entropy_dev", C.c_void_p), ("temperature", C.c_double), ("num_roots", C.c_int32),
                ("action_space", C.c_int32), ("deterministic", C.c_int32), ("root_offset", C.c_int32),
                ("seed", C.c_uint64), ("counter", C.c_uint32)]


class SupportHead(C.Structure):
    """lzb_support_head_t"""
    _fields_ = [("logits_dev", C.c_void_p), ("row_stride", C.c_int64), ("scalar_dev", C.c_void_p)]


class ValueTargets(C.Structure):
    """lzb_value_targets_t"""
    _fields_ = [("root_values_dev", C.c_void_p), ("value_mask_dev", C.c_void_p), ("rewards_dev", C.c_void_p),
                ("td_steps_dev", C.c_void_p), ("reward_len_dev", C.c_void_p), ("segment_len_dev", C.c_void_p),
                ("pos_dev", C.c_void_p), ("to_play_dev", C.c_void_p), ("pow_table_dev", C.c_void_p),
                ("target_values_dev", C.c_void_p), ("target_rewards_dev", C.c_void_p), ("num_segments", C.c_int32),
                ("num_unroll_steps", C.c_int32), ("reward_stride", C.c_int32), ("to_play_stride", C.c_int32),
                ("pow_len", C.c_int32), ("board_games", C.c_int32), ("lstm_horizon_len", C.c_int32)]


class GumbelHelper(C.Structure):
    """lzb_gumbel_helper_t"""
    _fields_ = [("op", C.c_int32), ("n", C.c_int32), ("a_dev", C.c_void_p), ("b_dev", C.c_void_p), ("c_dev", C.c_void_p),
                ("v_dev", C.c_void_p), ("s0", C.c_float), ("s1", C.c_float), ("s2", C.c_float), ("s3", C.c_float),
                ("i0", C.c_int32), ("i1", C.c_int32), ("out_dev", C.c_void_p), ("out_i_dev", C.c_void_p), ("tmp_dev", C.c_void_p)]


class TraverseIO(C.Structure):
    """lzb_traverse_io_t"""
    _fields_ = [("to_play_dev", C.c_void_p), ("ix_dev", C.c_void_p), ("iy_dev", C.c_void_p),
                ("last_action_dev", C.c_void_p), ("last_action_f_dev", C.c_void_p), ("search_len_dev", C.c_void_p),
                ("vtp_dev", C.c_void_p), ("gathers", C.POINTER(Gather)), ("num_gathers", C.c_int32),
                ("reuse", C.POINTER(Reuse)), ("chance", C.POINTER(Chance))]


class ChanceHead(C.Structure):
    """lzb_chance_head_t"""
    _fields_ = [("reward_dev", C.c_void_p), ("value_dev", C.c_void_p), ("logits_dev", C.c_void_p),
                ("logits_stride", C.c_int32), ("decision_logits_stride", C.c_int32), ("latent_dev", C.c_void_p)]


class BackpropIO(C.Structure):
    """lzb_backprop_io_t"""
    _fields_ = [("latent_index", C.c_int32), ("discount", C.c_float), ("reward_dev", C.c_void_p),
                ("value_dev", C.c_void_p), ("logits_dev", C.c_void_p), ("to_play_dev", C.c_void_p),
                ("is_reset_dev", C.c_void_p), ("lstm_horizon_len", C.c_int32), ("is_reset_out_dev", C.c_void_p),
                ("sampled_actions_dev", C.c_void_p), ("new_slots", Slots), ("reuse", C.POINTER(Reuse)),
                ("active_dev", C.c_void_p), ("chance_head", C.POINTER(ChanceHead))]


_P = C.c_void_p
_PROTOS = {
    "lzb_last_error": (C.c_char_p, []),
    "lzb_abi_version": (C.c_int, []),
    "lzb_tiebreak_hash": (C.c_uint32, [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]),
    "lzb_chance_bits": (C.c_uint64, [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]),
    "lzb_forest_bytes": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "lzb_forest_create": (C.c_int, [C.POINTER(_P), _P, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    _P, _P, _P]),
    "lzb_forest_destroy": (None, [_P]),
    "lzb_forest_reset": (C.c_int, [_P, _P, _P, _P]),
    "lzb_set_delta": (C.c_int, [_P, C.c_float, _P]),
    "lzb_prepare": (C.c_int, [_P, _P, _P, C.c_float, _P, _P, _P, _P, _P]),
    "lzb_set_slots": (C.c_int, [_P, C.c_int, C.POINTER(Slots), _P]),
    "lzb_traverse": (C.c_int, [_P, C.POINTER(SearchCfg), C.POINTER(TraverseIO), _P]),
    "lzb_backprop": (C.c_int, [_P, C.POINTER(BackpropIO), _P]),
    "lzb_step": (C.c_int, [_P, C.POINTER(SearchCfg), C.POINTER(BackpropIO), C.POINTER(TraverseIO), _P]),
    "lzb_store_hidden": (C.c_int, [_P, _P, _P, C.c_int, C.c_int64, _P]),
    "lzb_root_outputs": (C.c_int, [_P, _P, _P, _P, C.c_float, _P, _P, _P]),
    "lzb_trajectories": (C.c_int, [_P, _P, C.c_int, _P, _P]),
    "lzb_dump_tree": (C.c_int, [_P, C.c_int, _P, _P, _P, _P, _P, _P, _P, _P]),
    "lzb_dump_paths": (C.c_int, [_P, _P, C.c_int, _P]),
    "lzb_minmax": (C.c_int, [_P, _P, _P, _P]),
    "lzb_root_stats": (C.c_int, [_P, _P, _P, _P, _P]),
    "lzb_sample_actions": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint32, _P]),
    "lzb_sample_actions_seeded": (C.c_int, [_P, _P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, _P,
                                            C.c_uint32, _P]),
    "lzb_sample_actions_libstdcxx": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64, _P]),
    "lzb_legal_from_mask": (C.c_int, [_P, C.c_int, C.c_int, _P, _P, _P]),
    "lzb_dirichlet_noise": (C.c_int, [C.c_float, _P, _P, C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint32, _P]),
    "lzb_select_action": (C.c_int, [C.POINTER(SelectAction), _P]),
    "lzb_target_policies": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_int, _P, _P]),
    "lzb_scalar_workspace_bytes": (C.c_int, [C.c_int, C.c_int, C.POINTER(C.c_size_t)]),
    "lzb_inverse_scalar_transform": (C.c_int, [C.POINTER(SupportHead), C.c_int, C.c_int, C.c_int, C.c_float, _P, _P]),
    "lzb_inverse_scalar_transform_static": (C.c_int, [C.POINTER(SupportHead), C.c_int, C.c_int, C.c_int, C.c_float, _P]),
    "lzb_value_targets": (C.c_int, [C.POINTER(ValueTargets), _P]),
    "lzb_peer_bytes": (C.c_int, [C.c_int, C.c_int64, C.c_int, C.POINTER(C.c_size_t)]),
    "lzb_peer_alloc": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "lzb_peer_free": (C.c_int, [_P]),
    "lzb_peer_export": (C.c_int, [_P, _P]),
    "lzb_peer_import": (C.c_int, [_P, C.POINTER(_P)]),
    "lzb_peer_unimport": (C.c_int, [_P]),
    "lzb_peer_allgather": (C.c_int, [C.POINTER(_P), C.c_int, C.c_int, C.c_int64, C.POINTER(_P), C.POINTER(C.c_int32), C.c_int,
                                     C.c_uint32, C.c_int64, C.POINTER(_P), _P]),
    "lzb_peer_error": (C.c_int, [_P, C.c_int, C.c_int64, C.c_int, C.POINTER(C.c_int), _P]),
    "lzb_gumbel_table": (C.c_int, [C.c_int, C.c_int, _P]),
    "lzb_gumbel_noise": (C.c_int, [C.c_float, C.c_int, _P]),
    "lzb_gumbel_helper": (C.c_int, [_P, _P]),
    "lzb_check_word": (C.c_int, [_P, _P, _P, _P]),
    "lzb_test_expf": (C.c_int, [_P, _P, C.c_int64, _P]),
    "lzb_test_logf": (C.c_int, [_P, _P, C.c_int64, _P]),
}

_lib = None


def lib():
    """The loaded library with prototypes declared; builds it first when the sources are newer."""
    global _lib
    if _lib is None:
        path = os.environ.get("LZB200_LIB")  # tuning hook (A/B and timeline builds): load another build of the library
        if path:
            import sys
            print("lightzero_b200: LZB200_LIB is set - loading %s instead of %s" % (path, LIB_PATH), file=sys.stderr)
        else:
            path = build()
        L = C.CDLL(path)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)  # AttributeError here == header and library disagree
            fn.restype = res
            fn.argtypes = args
        if L.lzb_abi_version() != ABI_VERSION:
            raise RuntimeError("liblzb200.so ABI version mismatch")
        _lib = L
    return _lib


def exported_symbols():
    return sorted(_PROTOS)


class LzbError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        raise LzbError("liblzb200 error %d: %s" % (rc, lib().lzb_last_error().decode()))
