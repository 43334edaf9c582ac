"""ctypes binding of the in-tree CUDA library ``_C/libfedmse_b200.so`` (C ABI in ``include/fedmse_b200.h``).

The library is built in-tree by ``make`` / ``build()`` with
``nvcc -gencode arch=compute_100a,code=sm_100a`` (cross-compiles without a GPU) and
loaded with ``ctypes.CDLL``.  There is no silent fallback: on a box with a CUDA
device, asking for the CUDA backend without the library raises.
"""
from __future__ import annotations

import ctypes as C
import os
import shutil
import subprocess
from typing import Optional

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPO_ROOT = os.path.dirname(_PKG)
LIB_PATH = os.environ.get("FEDMSE_B200_LIB") or os.path.join(_PKG, "_C", "libfedmse_b200.so")   # override: experiments
_lib: Optional[C.CDLL] = None

c_f32p = C.POINTER(C.c_float)
c_i32p = C.POINTER(C.c_int32)
c_f64p = C.POINTER(C.c_double)
c_vp = C.c_void_p


class FmTrainArgs(C.Structure):
    _fields_ = [("params", c_vp), ("adam_m", c_vp), ("adam_v", c_vp), ("adam_t", c_vp), ("anchor", c_vp),
                ("best", c_vp), ("loss_log", c_vp), ("epochs_run", c_vp), ("trained_rounds", c_vp),
                ("x_train", c_vp), ("train_off", c_vp), ("x_valid", c_vp), ("valid_off", c_vp),
                ("sel_local", c_vp), ("n_sel", C.c_int32), ("first_client", C.c_int32),
                ("epochs", C.c_int32), ("batch", C.c_int32), ("patience", C.c_int32),
                ("log_stride", C.c_int32), ("lr", C.c_float), ("shrink_lambda", C.c_float),
                ("fedprox_mu", C.c_float), ("phase_cycles", c_vp)]


class FmTrainGenericArgs(C.Structure):
    _fields_ = [("params", c_vp), ("adam_m", c_vp), ("adam_v", c_vp), ("adam_t", c_vp), ("anchor", c_vp),
                ("best", c_vp), ("loss_log", c_vp), ("epochs_run", c_vp), ("trained_rounds", c_vp),
                ("x_train", c_vp), ("train_off", c_vp), ("x_valid", c_vp), ("valid_off", c_vp),
                ("sel_local", c_vp), ("n_sel", C.c_int32), ("first_client", C.c_int32),
                ("epochs", C.c_int32), ("batch", C.c_int32), ("patience", C.c_int32),
                ("log_stride", C.c_int32), ("lr", C.c_float), ("shrink_lambda", C.c_float),
                ("fedprox_mu", C.c_float), ("dim", C.c_int32), ("hidden", C.c_int32), ("latent", C.c_int32),
                ("scratch", c_vp), ("scratch_stride", C.c_int64)]


class FmScoreGenericArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("slot_client", c_vp), ("first_client", C.c_int32), ("n_local", C.c_int32),
                ("X", c_vp), ("tiles", c_vp), ("n_tiles", C.c_int32), ("out_sse", c_vp), ("out_lat", c_vp),
                ("lat_pitch", C.c_int32), ("dim", C.c_int32), ("hidden", C.c_int32), ("latent", C.c_int32)]


class FmKdeJsArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("slot_client", c_vp), ("first_client", C.c_int32), ("n_local", C.c_int32),
                ("n_slots", C.c_int32), ("X", c_vp), ("n_rows", C.c_int32), ("rec_scratch", c_vp), ("dens", c_vp),
                ("dist_out", c_vp)]


class FmScoreArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("slot_client", c_vp), ("first_client", C.c_int32), ("n_local", C.c_int32),
                ("X", c_vp), ("tiles", c_vp), ("n_tiles", C.c_int32), ("out_sse", c_vp), ("out_lat", c_vp)]


class FmMultiScoreArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("slot_client", c_vp), ("n_slots", C.c_int32), ("first_client", C.c_int32),
                ("n_local", C.c_int32), ("X", c_vp), ("X_lo", c_vp), ("x_rows", C.c_int64), ("precision", C.c_int32),
                ("images", c_vp), ("owned_slots", c_vp), ("n_owned", c_vp), ("out_sse", c_vp),
                ("dbg_cycles", c_vp)]


class FmFusedScoreArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("slot_client", c_vp), ("n_slots", C.c_int32), ("first_client", C.c_int32),
                ("n_local", C.c_int32), ("X", c_vp), ("x_rows", C.c_int64), ("images", c_vp), ("owned_slots", c_vp),
                ("n_owned", c_vp), ("partial", c_vp), ("out_mse", c_vp), ("grid_barrier", c_vp), ("dbg_cycles", c_vp)]


TC_IMAGE_BYTES = 65536


class FmVoteArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("sel", c_vp), ("n_sel", C.c_int32), ("first_client", C.c_int32),
                ("n_local", C.c_int32), ("V", c_vp), ("rows", C.c_int32), ("chunk", C.c_int32),
                ("jitter", c_vp), ("voter_pos", C.c_int32), ("tiles", c_vp), ("n_tiles", C.c_int32),
                ("scratch_norm", c_vp), ("scratch_sse", c_vp), ("scores", c_vp)]


class FmVoteStaticArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("sel", c_vp), ("n_sel", C.c_int32), ("first_client", C.c_int32),
                ("n_local", C.c_int32), ("V", c_vp), ("rows", C.c_int32), ("chunk", C.c_int32),
                ("jitter", c_vp), ("voter_pos", C.c_int32), ("tiles", c_vp), ("n_tiles", C.c_int32),
                ("scores", c_vp)]


MAX_SEL = 4096
CTL_FLOATS = 2 * MAX_SEL + 16
STATIC_VOTE_MAX_ROWS = 1024
STATIC_VOTE_MAX_SEL = 64


class FmMergeArgs(C.Structure):
    _fields_ = [("peer_slabs", c_vp), ("world", C.c_int32), ("rank", C.c_int32), ("n_global", C.c_int32),
                ("sel", c_vp), ("n_sel", C.c_int32), ("mse", c_vp), ("mode", C.c_int32), ("compat", C.c_int32),
                ("agg_pos", c_vp), ("agg_count", c_vp), ("theta", c_vp), ("weights_out", c_vp),
                ("arrive", c_vp), ("expect", C.c_uint32), ("done_ctl", c_vp), ("done_counter", c_vp)]


class FmMergeNvlsArgs(C.Structure):
    _fields_ = [("slab", c_vp), ("world", C.c_int32), ("rank", C.c_int32), ("n_global", C.c_int32),
                ("sel", c_vp), ("n_sel", C.c_int32), ("mse", c_vp), ("mode", C.c_int32), ("compat", C.c_int32),
                ("agg_pos", c_vp), ("agg_count", c_vp), ("weights_out", c_vp), ("arrive_mse", c_vp),
                ("expect_mse", C.c_uint32), ("partial_local", c_vp), ("mc_partial", c_vp), ("mc_theta", c_vp),
                ("peer_ctl", c_vp), ("arrive_partial", c_vp), ("expect_partial", C.c_uint32), ("grid_counter", c_vp)]


ARRIVE_DONE, ARRIVE_GOSSIP, ARRIVE_PARTIAL = 2, 3, 4


class FmVerifyArgs(C.Structure):
    _fields_ = [("params", c_vp), ("anchor", c_vp), ("hist", c_vp), ("hist_perf", c_vp), ("hist_valid", c_vp),
                ("rejected", c_vp), ("n_local", C.c_int32), ("first_client", C.c_int32), ("theta", c_vp),
                ("agg_pos", c_vp), ("sse", c_vp), ("sse_off", c_vp), ("shared_segment", C.c_int32),
                ("thr_change", C.c_float), ("thr_perf", C.c_float), ("out_change", c_vp), ("out_dperf", c_vp),
                ("arrive", c_vp), ("expect", C.c_uint32)]


class FmAucArgs(C.Structure):
    _fields_ = [("mode", C.c_int32), ("n_local", C.c_int32), ("test_sse", c_vp), ("test_lat", c_vp),
                ("labels", c_vp), ("test_off", c_vp), ("train_lat", c_vp), ("train_off", c_vp),
                ("scores_out", c_vp), ("auc", c_vp), ("max_rows", C.c_int32), ("keys_scratch", c_vp)]


class FmGossipArgs(C.Structure):
    _fields_ = [("peer_src", c_vp), ("world", C.c_int32), ("rank", C.c_int32), ("n_global", C.c_int32),
                ("params", c_vp), ("anchor", c_vp), ("hist", c_vp), ("hist_perf", c_vp), ("hist_valid", c_vp),
                ("rejected", c_vp), ("n_local", C.c_int32), ("first_client", C.c_int32),
                ("sel_local", c_vp), ("n_sel", C.c_int32), ("nbr_off", c_vp), ("nbr_idx", c_vp),
                ("x_valid", c_vp), ("valid_off", c_vp), ("mode", C.c_int32), ("thr_change", C.c_float),
                ("thr_perf", C.c_float), ("topo_kind", C.c_int32), ("topo_k", C.c_int32), ("round", C.c_int32),
                ("topo_seed", C.c_uint64), ("snap_parity", C.c_int32), ("snap_stride", C.c_int64),
                ("arrive", c_vp), ("expect", C.c_uint32)]


TOPO_KIND = {"csr": 0, "full": 1, "ring": 2, "random-k": 3, "random_k": 3, "random": 3}


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every ``csrc/*.cu`` for sm_100a into the in-tree shared library."""
    srcs = sorted(os.path.join(REPO_ROOT, "csrc", f) for f in os.listdir(os.path.join(REPO_ROOT, "csrc"))
                  if f.endswith(".cu"))
    deps = srcs + [os.path.join(REPO_ROOT, "csrc", f) for f in os.listdir(os.path.join(REPO_ROOT, "csrc"))
                   if f.endswith(".cuh")] + [os.path.join(REPO_ROOT, "include", "fedmse_b200.h")]
    if not force and os.path.exists(LIB_PATH) and \
            all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(d) for d in deps):
        return LIB_PATH
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build libfedmse_b200.so")
    bdir = os.path.join(REPO_ROOT, "build")
    os.makedirs(bdir, exist_ok=True)
    os.makedirs(os.path.dirname(LIB_PATH), exist_ok=True)
    arch = ["-gencode", "arch=compute_100a,code=sm_100a"]
    flags = arch + ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
    procs = []
    objs = []
    for s in srcs:
        o = os.path.join(bdir, os.path.basename(s)[:-3] + ".o")
        objs.append(o)
        log = open(o[:-2] + ".ptxas.log", "w")
        procs.append((s, subprocess.Popen([nvcc] + flags + ["-c", s, "-o", o], stdout=log, stderr=log), log))
    for s, p, log in procs:
        rc = p.wait()
        log.close()
        if rc != 0:
            raise RuntimeError(f"nvcc failed on {s}:\n" + open(log.name).read())
    subprocess.check_call([nvcc] + arch + ["-shared", "-o", LIB_PATH] + objs + ["-lcudart"])
    if verbose:
        print("built", LIB_PATH)
    return LIB_PATH


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing. Build it with `make` (or python -c 'import __graft_entry__ as g; "
            "g.build()') -- the CUDA backend has no eager fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, argt in (
        ("fm_train_clients", [C.POINTER(FmTrainArgs), c_vp]),
        ("fm_train_generic", [C.POINTER(FmTrainGenericArgs), c_vp]),
        ("fm_score_generic", [C.POINTER(FmScoreGenericArgs), c_vp]),
        ("fm_kde_js", [C.POINTER(FmKdeJsArgs), c_vp]),
        ("fm_score_rows", [C.POINTER(FmScoreArgs), c_vp]),
        ("fm_score_rows_tc", [C.POINTER(FmScoreArgs), C.c_int64, C.c_int32, c_vp, c_vp]),
        ("fm_split_residual", [c_vp, c_vp, C.c_int64, c_vp]),
        ("fm_score_multi_tc", [C.POINTER(FmMultiScoreArgs), c_vp]),
        ("fm_score_fused_f16", [C.POINTER(FmFusedScoreArgs), c_vp]),
        ("fm_vote_normalise", [c_vp, C.c_int32, c_vp, c_vp]),
        ("fm_chunk_mean", [c_vp, C.c_int32, C.c_int32, C.c_int32, c_vp, c_vp, C.c_int32, C.c_int32,
                           C.c_int32, c_vp, c_vp]),
        ("fm_vote_scores", [C.POINTER(FmVoteArgs), c_vp]),
        ("fm_vote_scores_static", [C.POINTER(FmVoteStaticArgs), c_vp]),
        ("fm_elect", [c_vp, c_vp, C.c_int32, C.c_int32, c_vp, C.c_int32, c_vp, c_vp, C.c_uint32, c_vp]),
        ("fm_publish", [c_vp, c_vp, C.c_int32, C.c_int32, C.c_int32, c_vp, C.c_int32, C.c_int32, c_vp]),
        ("fm_merge", [C.POINTER(FmMergeArgs), c_vp]),
        ("fm_verify_update", [C.POINTER(FmVerifyArgs), c_vp]),
        ("fm_merge_nvls", [C.POINTER(FmMergeNvlsArgs), c_vp]),
        ("fm_flag_wait", [c_vp, C.c_uint32, c_vp]),
        ("fm_nvls_allreduce_slice", [c_vp, C.c_int32, C.c_int32, C.c_int64, c_vp, C.c_uint32, c_vp, C.c_int32, c_vp, c_vp]),
        ("fm_auc", [C.POINTER(FmAucArgs), c_vp]),
        ("fm_gossip", [C.POINTER(FmGossipArgs), c_vp]),
        ("fm_gossip_publish", [c_vp, c_vp, C.c_int64, c_vp, C.c_int32, c_vp, c_vp]),
        ("fm_random_neighbour", [C.c_uint64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32]),
        ("fm_weighted_rows", [c_vp, c_vp, C.c_int32, c_vp, C.c_int64, c_vp, C.c_uint32, c_vp, C.c_int32, C.c_int32,
                              c_vp, c_vp]),
        ("fm_slice_reduce", [c_vp, C.c_int32, C.c_int32, c_vp, C.c_int64, c_vp, C.c_uint32, c_vp, C.c_int32, c_vp, c_vp]),
        ("fm_slice_gather", [c_vp, C.c_int32, c_vp, C.c_int64, c_vp, C.c_uint32, c_vp]),
        ("fm_slab_alloc", [C.c_uint64, C.POINTER(c_vp)]),
        ("fm_slab_free", [c_vp]),
        ("fm_ipc_get_handle", [c_vp, C.c_char_p]),
        ("fm_ipc_open_handle", [C.c_char_p, C.POINTER(c_vp)]),
        ("fm_ipc_close_handle", [c_vp]),
        ("fm_enable_peer_access", [C.c_int32]),
        ("fm_device_info", [c_i32p, c_i32p, c_i32p, c_i32p]),
    ):
        fn = getattr(lib, name)
        fn.argtypes = argt
        fn.restype = C.c_int
    lib.fm_score_fused_scratch_floats.argtypes = [C.c_int64, C.c_int32]
    lib.fm_score_fused_scratch_floats.restype = C.c_int64
    lib.fm_train_generic_scratch_floats.argtypes = [C.c_int32] * 4
    lib.fm_train_generic_scratch_floats.restype = C.c_int64
    lib.fm_kde_js_scratch_floats.argtypes = [C.c_int32, C.c_int32]
    lib.fm_kde_js_scratch_floats.restype = C.c_int64
    lib.fm_auc_scratch_keys.argtypes = [C.c_int32]
    lib.fm_auc_scratch_keys.restype = C.c_int64
    lib.fm_vote_scores_static_error.argtypes = []
    lib.fm_vote_scores_static_error.restype = C.c_char_p
    lib.fm_error_string.argtypes = [C.c_int]
    lib.fm_error_string.restype = C.c_char_p
    _lib = lib
    return lib


def check(code: int, what: str = "") -> None:
    if code != 0:
        msg = load().fm_error_string(code).decode()
        raise RuntimeError(f"fedmse_b200 {what} failed: CUDA error {code}: {msg}")


def is_available() -> bool:
    return os.path.exists(LIB_PATH)
