"""ctypes binding of the C ABI in include/nlbac.h (no torch types cross the boundary).

The shared library is built in-tree by ``build()`` (nvcc, sm_100a) and loaded lazily.  There is no
CPU fallback: if the library is missing or a call fails, a RuntimeError is raised.
"""
import ctypes
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB_PATH = os.path.join(HERE, "libnlbac_b200.so")
SOURCES = ["node_abi.cu", "mlp_abi.cu", "tail_abi.cu", "tc_abi.cu", "env_abi.cu", "comm_abi.cu"]
NVCC_FLAGS = ["-std=c++20", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC"]

_lib = None


class MlpDesc(ctypes.Structure):
    _fields_ = [("n_in", ctypes.c_int), ("hidden", ctypes.c_int), ("n_out", ctypes.c_int)]


MAX_HAZARDS = 15
MAX_TENSORS = 32


class ConstraintDesc(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int), ("m", ctypes.c_int), ("n_hazards", ctypes.c_int), ("dt", ctypes.c_float),
                ("gamma_b", ctypes.c_float), ("gamma_l", ctypes.c_float), ("lya_inv_dt", ctypes.c_float),
                ("batch_size", ctypes.c_float), ("collision_radius", ctypes.c_float),
                ("hazards", ctypes.c_float * (2 * MAX_HAZARDS)),
                ("operator_dist", ctypes.c_float), ("y_max", ctypes.c_float), ("y_min", ctypes.c_float),
                ("delta_y", ctypes.c_float)]


class NodeDesc(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int), ("n_s", ctypes.c_int), ("n_u", ctypes.c_int), ("n_c", ctypes.c_int),
                ("hidden", ctypes.c_int), ("n_f_layers", ctypes.c_int), ("n_g_layers", ctypes.c_int)]


def raw_stream():
    """cudaStream_t of torch's current stream on the current device, as an int.  ``torch.cuda.current_stream()``
    costs ~25 us of Python per call, which matters for the launch-bound B = 128 / 256 updates."""
    import torch
    return torch._C._cuda_getCurrentRawStream(torch._C._cuda_getDevice())


def _newer(src_files, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(f) > t for f in src_files)


def build(force=False, verbose=False):
    """Compile every CUDA source for sm_100a into libnlbac_b200.so (in-tree)."""
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(ROOT, "include", "nlbac.h")]
    if not force and not _newer(deps, LIB_PATH):
        return LIB_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    for s in srcs:
        o = s[:-3] + ".o"
        objs.append(o)
        cmd = [nvcc] + NVCC_FLAGS + os.environ.get("NLBAC_NVCC_EXTRA", "").split() + ["-c", s, "-o", o]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            print(out.decode())
        if p.returncode != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    cmd = [nvcc, "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ldl"]
    subprocess.check_call(cmd)
    return LIB_PATH


def _declare(lib):
    c = ctypes
    vp, ip, fp = c.c_void_p, c.POINTER(c.c_int), c.POINTER(c.c_float)
    dp = c.POINTER(NodeDesc)
    lib.nlbac_last_error.restype = c.c_char_p
    lib.nlbac_abi_version.restype = c.c_int
    lib.nlbac_node_supported.argtypes = [dp, c.c_int]
    lib.nlbac_node_param_count.argtypes = [dp]
    lib.nlbac_node_param_count.restype = c.c_int64
    lib.nlbac_node_packed_floats.argtypes = [dp]
    lib.nlbac_node_packed_floats.restype = c.c_int64
    lib.nlbac_node_pack.argtypes = [dp, c.POINTER(vp), vp, vp]
    lib.nlbac_node_tc_packed_floats.argtypes = [dp]
    lib.nlbac_node_tc_packed_floats.restype = c.c_int64
    lib.nlbac_node_tc_pack.argtypes = [dp, c.POINTER(vp), vp, vp]
    lib.nlbac_node_tc_bwd_workspace_bytes.argtypes = [dp, c.c_int64, c.c_int, c.c_int, c.c_int, c.c_int, c.c_int,
                                                      c.POINTER(c.c_size_t)]
    lib.nlbac_node_workspace_bytes.argtypes = [dp, c.c_int64, c.c_int, c.c_int, c.c_int,
                                               c.POINTER(c.c_size_t), c.POINTER(c.c_size_t), c.POINTER(c.c_size_t)]
    lib.nlbac_node_solve_fwd.argtypes = [dp, vp, vp, c.c_int64, fp, c.c_int, fp, c.c_int, c.c_int, c.c_int,
                                         vp, vp, vp, vp]
    lib.nlbac_node_solve_bwd.argtypes = [dp, vp, vp, c.c_int64, fp, c.c_int, fp, c.c_int, c.c_int, c.c_int,
                                         vp, vp, vp, vp, vp, vp]
    lib.nlbac_node_solve_bwd_noalloc.argtypes = [dp, vp, vp, c.c_int64, fp, c.c_int, fp, c.c_int, c.c_int,
                                                 vp, vp, vp, vp, vp]
    szp = c.POINTER(c.c_size_t)
    lib.nlbac_node_dopri5_workspace_bytes.argtypes = [dp, c.c_int64, c.c_int, c.c_int, szp, szp, szp]
    lib.nlbac_node_dopri5_fwd.argtypes = [dp, vp, vp, c.c_int64, fp, c.c_int, c.c_float, c.c_float, c.c_int, c.c_int,
                                          vp, vp, vp, ip, ip, fp, ip, fp, vp]
    lib.nlbac_node_dopri5_bwd.argtypes = [dp, vp, vp, c.c_int64, c.c_int, c.c_int, fp, ip, fp, c.c_int,
                                          vp, vp, vp, vp, vp, vp]
    lib.nlbac_node_dopri5_state_bytes.argtypes = [c.c_int, c.c_int, szp]
    lib.nlbac_node_dopri5_fwd_dev.argtypes = [dp, vp, vp, c.c_int64, fp, c.c_int, c.c_float, c.c_float, c.c_int,
                                              c.c_int, c.c_int, c.c_int, vp, vp, vp, vp, vp, c.c_int64, vp]
    lib.nlbac_node_dopri5_state_read.argtypes = [vp, c.c_int, c.c_int, ip, ip, ip, ip, fp, vp]
    lib.nlbac_node_dopri5_bwd_dev.argtypes = [dp, vp, vp, c.c_int64, c.c_int, c.c_int, vp, c.c_int, vp, vp, vp, vp,
                                              vp, vp]
    mp, cp, i64 = c.POINTER(MlpDesc), c.POINTER(ConstraintDesc), c.c_int64
    lib.nlbac_mlp_workspace_bytes.argtypes = [mp, i64, szp, szp]
    lib.nlbac_mlp_fwd.argtypes = [mp, c.POINTER(vp), vp, i64, vp, vp, vp]
    lib.nlbac_mlp_bwd.argtypes = [mp, c.POINTER(vp), vp, i64, vp, vp, vp, c.POINTER(vp), vp, vp]
    lib.nlbac_policy_head_fwd.argtypes = [vp, vp, vp, vp, i64, c.c_int, vp, vp, vp, vp]
    lib.nlbac_policy_head_bwd.argtypes = [vp, vp, vp, i64, c.c_int, vp, vp, vp, vp, vp]
    lib.nlbac_constraints_fwd.argtypes = [cp, vp, vp, vp, vp, vp, i64, vp, vp]
    lib.nlbac_constraints_bwd.argtypes = [cp, vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp]
    lib.nlbac_constraints4_fwd.argtypes = [cp, vp, vp, vp, vp, vp, vp, i64, vp, vp]
    lib.nlbac_constraints4_bwd.argtypes = [cp, vp, vp, vp, vp, vp, vp, i64, vp, vp, vp, vp, vp, vp]
    lib.nlbac_lookahead_fwd.argtypes = [vp, c.c_int, c.c_float, i64, vp, vp]
    lib.nlbac_lookahead_bwd.argtypes = [vp, c.c_int, c.c_float, i64, vp, vp, vp]
    lib.nlbac_get_state.argtypes = [c.c_int, vp, i64, vp, vp, vp]
    lib.nlbac_get_obs_fwd.argtypes = [c.c_int, vp, i64, vp, vp]
    lib.nlbac_get_obs_bwd.argtypes = [c.c_int, vp, i64, vp, vp, vp]
    lib.nlbac_adam_step.argtypes = [vp, vp, vp, vp, i64, c.c_int, c.c_float, c.c_float, c.c_float, c.c_float, vp]
    lib.nlbac_soft_update.argtypes = [vp, vp, i64, c.c_float, vp]
    pvp, pi64 = c.POINTER(vp), c.POINTER(c.c_int64)
    lib.nlbac_adam_step_multi.argtypes = [c.c_int, pvp, pvp, pvp, pvp, pi64, c.c_int, c.c_float, c.c_float, c.c_float,
                                          c.c_float, vp]
    lib.nlbac_adam_step_multi_dev.argtypes = [c.c_int, pvp, pvp, pvp, pvp, pi64, vp, c.c_int, c.c_float, c.c_float,
                                              c.c_float, c.c_float, vp]
    lib.nlbac_soft_update_multi.argtypes = [c.c_int, pvp, pvp, pi64, c.c_float, vp]
    lib.nlbac_replay_gather.argtypes = [vp, c.c_int, vp, i64, vp, vp]
    lib.nlbac_env_unicycle_reset.argtypes = [vp, vp, vp, vp, i64, vp, vp]
    lib.nlbac_env_unicycle_step.argtypes = [vp, vp, vp, vp, i64] + [vp] * 10
    lib.nlbac_env_cars_reset.argtypes = [vp, vp, vp, vp, vp, i64, vp, vp]
    lib.nlbac_env_cars_step.argtypes = [vp, vp, vp, vp, i64] + [vp] * 10
    lib.nlbac_env_pvtol_reset.argtypes = [vp, vp, vp, i64, vp, vp]
    lib.nlbac_env_pvtol_step.argtypes = [vp, vp, vp, i64] + [vp] * 9
    lib.nlbac_comm_unique_id.argtypes = [vp]
    lib.nlbac_comm_init.argtypes = [c.c_int, c.c_int, vp, c.POINTER(vp)]
    lib.nlbac_replay_gather_split.argtypes = [vp, c.c_int, vp, c.c_int64, c.c_int, ip, ip, pvp, vp]
    lib.nlbac_td_target.argtypes = [vp, vp, vp, vp, vp, vp, c.c_float, c.c_int64, vp, vp]
    lib.nlbac_mse_mean_fwd.argtypes = [vp, vp, c.c_int64, vp, vp]
    lib.nlbac_mse_mean_bwd.argtypes = [vp, vp, c.c_int64, vp, vp, vp]
    lib.nlbac_sac_actor_term_fwd.argtypes = [vp, vp, vp, vp, c.c_int64, vp, vp]
    lib.nlbac_sac_actor_term_bwd.argtypes = [vp, vp, vp, c.c_int64, vp, vp, vp, vp, vp]
    lib.nlbac_lagrangian_fwd.argtypes = [vp, c.c_int, c.c_int, vp, vp, vp, vp, c.c_int, c.c_double, c.c_double, c.c_double, c.c_int,
                                         c.c_int, c.c_double, vp, vp, vp]
    lib.nlbac_lagrangian_bwd.argtypes = [vp, vp, c.c_int, c.c_double, vp, vp, vp]
    lib.nlbac_allreduce_grads.argtypes = [vp, vp, c.c_size_t, vp]
    lib.nlbac_allreduce_sum_f64.argtypes = [vp, vp, c.c_size_t, vp]
    lib.nlbac_comm_destroy.argtypes = [vp]
    lib.nlbac_umma_selftest.argtypes = [c.c_int, vp, vp, vp, vp]
    lib.nlbac_kernel_timing.argtypes = [c.c_int]
    lib.nlbac_kernel_timing_read.argtypes = [c.POINTER(c.c_double), c.POINTER(c.c_int)]
    lib.nlbac_umma_probe.argtypes = [c.POINTER(c.c_int), vp, vp, vp, vp]
    lib.nlbac_umma_peak_probe.argtypes = [c.c_int, c.c_int, c.POINTER(c.c_double), vp]
    lib.nlbac_fp32_peak_probe.argtypes = [c.c_int, vp, c.POINTER(c.c_double), vp]
    lib.nlbac_fma_shape_probe.argtypes = [c.c_int, c.c_int, vp, c.POINTER(c.c_double), vp]
    return lib


def lib():
    """Load the CUDA library; fail loudly if it was never built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "nlbac_b200: %s is missing — run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        _lib = _declare(ctypes.CDLL(LIB_PATH))
    return _lib


def check(rc, what):
    if rc != 0:
        raise RuntimeError("nlbac_b200 %s failed (rc=%d): %s" % (what, rc, lib().nlbac_last_error().decode()))


def declared_symbols():
    """Every function name declared in include/nlbac.h (used by the CPU export test)."""
    import re
    text = open(os.path.join(ROOT, "include", "nlbac.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nlbac_[a-z0-9_]+)\s*\(", text)))
