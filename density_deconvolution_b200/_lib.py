"""ctypes binding of ``libdgmm_b200.so`` (C ABI in ``include/dgmm_b200.h``).

The library is built in-tree (``python __graft_entry__.py`` or ``make -C
density_deconvolution_b200/csrc``).  There is no fallback: if the shared object is
missing or a call fails, an exception is raised.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# DGMM_LIB: A/B runs of an experimental build of the same library (tools/gpu_ab_lib.sh); never a fallback
LIB_PATH = os.environ.get("DGMM_LIB") or os.path.join(_HERE, "libdgmm_b200.so")
CSRC_DIR = os.path.join(_HERE, "csrc")

NOISE_FULL, NOISE_DIAG, NOISE_NONE, NOISE_SHARED, NOISE_PACKED = 0, 1, 2, 3, 4
NOISE_LAYOUTS = {"full": NOISE_FULL, "diag": NOISE_DIAG, "none": NOISE_NONE, "shared": NOISE_SHARED,
                 "packed": NOISE_PACKED}
MAX_D, MAX_K = 32, 1024

_ERRORS = {
    -1: "bad argument", -2: "unsupported dimension D (1..32)", -3: "unsupported component count K (1..1024) for this call",
    -4: "workspace too small", -5: "unknown noise layout", -6: "no sm_100 CUDA device",
    -7: "out_ll is required for this shape (dgmm_stats_needs_ll)",
    -8: "shape needs more scratch than the library's static array holds (use the entry point that takes a workspace)",
}

# every symbol include/dgmm_b200.h declares (tests check the list against the header)
SYMBOLS = (
    "dgmm_abi_version", "dgmm_supported_dims", "dgmm_workspace_bytes", "dgmm_stats_needs_ll", "dgmm_loglik_f32",
    "dgmm_estep_stats_f32", "dgmm_estep_full_f32", "dgmm_mstep_update_f32", "dgmm_online_mstep_f32",
    "dgmm_fp32_probe",
    "dgmm_ctx_create", "dgmm_ctx_destroy", "dgmm_em_iteration_host", "dgmm_estep_stats_host", "dgmm_score_host",
    "dgmm_ctx_upload", "dgmm_em_iteration_resident", "dgmm_em_fit_resident", "dgmm_score_resident",
    "dgmm_pack_noise_f32", "dgmm_pack_noise_host",
    "dgmm_kmeans_workspace_bytes", "dgmm_kmeans_assign_f32", "dgmm_kmeans_update_f32",
    "dgmm_kmeans_static_scratch_bytes", "dgmm_kmeans_assign_static_f32",
    "dgmm_gather_points_f32", "dgmm_posterior_sample_f32", "dgmm_posterior_logprob_f32", "dgmm_sgd_params_f32", "dgmm_sgd_adam_step_f32", "dgmm_sgd_fused_step_f32", "dgmm_sgd_adam_step_peers_f32",
    "dgmm_debug_set_tuning", "dgmm_debug_get_plan", "dgmm_debug_pair_timing", "dgmm_debug_pair_times",
)


class DgmmError(RuntimeError):
    pass


def build(verbose=False, jobs=None):
    """Compile the CUDA sources for sm_100a into ``libdgmm_b200.so`` (nvcc cross-compiles
    without a GPU)."""
    import subprocess
    jobs = jobs or os.cpu_count() or 4
    cmd = ["make", "-C", CSRC_DIR, "-j%d" % jobs]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise DgmmError("building libdgmm_b200.so failed (exit %d)" % res.returncode)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DgmmError(
            "%s is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, i32, i64, f32, f64, sz = (ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float,
                                  ctypes.c_double, ctypes.c_size_t)
    L.dgmm_abi_version.restype = i32
    L.dgmm_abi_version.argtypes = []
    L.dgmm_supported_dims.restype = i32
    L.dgmm_supported_dims.argtypes = [ctypes.POINTER(i32), ctypes.POINTER(i32)]
    L.dgmm_workspace_bytes.restype = sz
    L.dgmm_workspace_bytes.argtypes = [i32, i32]
    L.dgmm_stats_needs_ll.restype = i32
    L.dgmm_stats_needs_ll.argtypes = [i32, i32, i32]
    L.dgmm_loglik_f32.restype = i32
    L.dgmm_loglik_f32.argtypes = [vp, vp, i32, i64, i32, i32, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    L.dgmm_estep_stats_f32.restype = i32
    L.dgmm_estep_stats_f32.argtypes = [vp, vp, i32, i64, i32, i32, vp, vp, vp, vp, f32, vp, vp, vp, sz, vp]
    L.dgmm_estep_full_f32.restype = i32
    L.dgmm_estep_full_f32.argtypes = [vp, vp, i32, i64, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, sz, vp]
    L.dgmm_mstep_update_f32.restype = i32
    L.dgmm_mstep_update_f32.argtypes = [vp, f64, f32, i32, i32, vp, vp, vp, vp, vp]
    L.dgmm_online_mstep_f32.restype = i32
    L.dgmm_online_mstep_f32.argtypes = [vp, f32, f32, f32, f64, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp]
    L.dgmm_fp32_probe.restype = i32
    L.dgmm_fp32_probe.argtypes = [vp, i32, ctypes.POINTER(f64), vp]
    L.dgmm_ctx_create.restype = i32
    L.dgmm_ctx_create.argtypes = [ctypes.POINTER(vp), i32, i32, i32, i64]
    L.dgmm_ctx_destroy.restype = None
    L.dgmm_ctx_destroy.argtypes = [vp]
    L.dgmm_em_iteration_host.restype = i32
    L.dgmm_em_iteration_host.argtypes = [vp, vp, vp, i32, i64, f32, vp, vp, vp, ctypes.POINTER(f64),
                                         ctypes.POINTER(i64)]
    L.dgmm_estep_stats_host.restype = i32
    L.dgmm_estep_stats_host.argtypes = [vp, vp, vp, i32, i64, vp, vp, vp, f32, vp]
    L.dgmm_score_host.restype = i32
    L.dgmm_score_host.argtypes = [vp, vp, vp, i32, i64, vp, vp, vp, vp, ctypes.POINTER(f64), ctypes.POINTER(i64)]
    L.dgmm_ctx_upload.restype = i32
    L.dgmm_ctx_upload.argtypes = [vp, vp, vp, i32, i64, i32]
    L.dgmm_em_iteration_resident.restype = i32
    L.dgmm_em_iteration_resident.argtypes = [vp, f32, vp, vp, vp, ctypes.POINTER(f64), ctypes.POINTER(i64)]
    L.dgmm_em_fit_resident.restype = i32
    L.dgmm_em_fit_resident.argtypes = [vp, i32, f32, f64, vp, vp, vp, ctypes.POINTER(i32), ctypes.POINTER(f64),
                                       ctypes.POINTER(i64)]
    L.dgmm_score_resident.restype = i32
    L.dgmm_score_resident.argtypes = [vp, vp, vp, vp, vp, ctypes.POINTER(f64), ctypes.POINTER(i64)]
    L.dgmm_pack_noise_f32.restype = i32
    L.dgmm_pack_noise_f32.argtypes = [vp, i64, i32, vp, vp]
    L.dgmm_pack_noise_host.restype = i32
    L.dgmm_pack_noise_host.argtypes = [vp, i64, i32, vp]
    L.dgmm_kmeans_workspace_bytes.restype = sz
    L.dgmm_kmeans_workspace_bytes.argtypes = [i32, i32]
    L.dgmm_kmeans_assign_f32.restype = i32
    L.dgmm_kmeans_assign_f32.argtypes = [vp, i64, i32, i32, vp, vp, vp, vp, sz, vp]
    L.dgmm_kmeans_static_scratch_bytes.restype = sz
    L.dgmm_kmeans_static_scratch_bytes.argtypes = []
    L.dgmm_kmeans_assign_static_f32.restype = i32
    L.dgmm_kmeans_assign_static_f32.argtypes = [vp, i64, i32, i32, vp, vp, vp, vp]
    L.dgmm_kmeans_update_f32.restype = i32
    L.dgmm_kmeans_update_f32.argtypes = [i32, vp, i32, i32, vp, vp, vp]
    L.dgmm_gather_points_f32.restype = i32
    L.dgmm_gather_points_f32.argtypes = [vp, vp, i32, vp, i64, i32, vp, vp, vp]
    L.dgmm_posterior_sample_f32.restype = i32
    L.dgmm_posterior_sample_f32.argtypes = [vp, vp, i32, i64, i32, i32, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, vp,
                                            vp, sz, vp]
    L.dgmm_posterior_logprob_f32.restype = i32
    L.dgmm_posterior_logprob_f32.argtypes = [vp, vp, i32, i64, i32, i32, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, sz,
                                             vp]
    L.dgmm_sgd_params_f32.restype = i32
    L.dgmm_sgd_params_f32.argtypes = [vp, vp, vp, i32, i32, vp, vp, vp]
    L.dgmm_sgd_adam_step_f32.restype = i32
    L.dgmm_sgd_adam_step_f32.argtypes = [vp, f64, f64, f32, f32, f32, f32, f32, i32, i32, i32, vp, vp, vp, vp,
                                         ctypes.POINTER(vp), ctypes.POINTER(vp), vp, vp, vp, vp]
    L.dgmm_sgd_fused_step_f32.restype = i32
    L.dgmm_sgd_fused_step_f32.argtypes = [vp, vp, i32, i32, i32, vp, i64, i64, vp, vp, vp, i32, vp, f64, f64, f32, f32,
                                          f32, f32, f32, i32, vp, vp, vp, vp, ctypes.POINTER(vp), ctypes.POINTER(vp),
                                          vp, vp, vp, vp, sz, vp]
    L.dgmm_sgd_adam_step_peers_f32.restype = i32
    L.dgmm_sgd_adam_step_peers_f32.argtypes = [ctypes.POINTER(vp), i32, i32, i64, i64, ctypes.c_uint32, vp, f64, f64,
                                               f32, f32, f32, f32, f32, i32, i32, i32, vp, vp, vp, vp,
                                               ctypes.POINTER(vp), ctypes.POINTER(vp), vp, vp, vp, vp]
    L.dgmm_debug_set_tuning.restype = i32
    L.dgmm_debug_set_tuning.argtypes = [i32] * 8
    L.dgmm_debug_get_plan.restype = i32
    L.dgmm_debug_get_plan.argtypes = [i32, i32, i32, i64, i32, i32, ctypes.POINTER(i32)]
    L.dgmm_debug_pair_timing.restype = i32
    L.dgmm_debug_pair_timing.argtypes = [i32]
    L.dgmm_debug_pair_times.restype = i32
    L.dgmm_debug_pair_times.argtypes = [ctypes.POINTER(f32), ctypes.POINTER(i32)]
    if L.dgmm_abi_version() != 2:
        raise DgmmError("libdgmm_b200.so ABI version mismatch")
    _lib = L
    return L


def check(rc, what):
    if rc == 0:
        return
    if rc < 0:
        raise DgmmError("%s: %s (code %d)" % (what, _ERRORS.get(rc, "error"), rc))
    raise DgmmError("%s: CUDA runtime error %d" % (what, rc))
